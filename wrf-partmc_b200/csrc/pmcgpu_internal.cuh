// pmcgpu_internal.cuh -- shared declarations of the B200 aerosol-step library.
// Product code: never includes anything under oracle/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <map>
#include <string>
#include <utility>
#include <vector>
#include "../../include/pmcgpu.h"

#define PMC_MAX_SPEC 64
#define PMC_WARPS_PER_CTA 8

// draw-address phases; must agree with the documented stream layout
enum : uint32_t {
  PH_COAG = 1, PH_COAG_ACCEL = 2, PH_REWEIGHT = 3, PH_HALVE = 4, PH_DOUBLE = 5, PH_SCALE = 6, PH_TRANS = 7, PH_INIT = 8, PH_DEP = 9, PH_EMIT = 10, PH_BND = 11, PH_COMP = 12
};

struct PmcConst {  // partmc/src/constants.F90:30-65
  static constexpr double pi = 3.14159265358979323846;
  static constexpr double boltzmann = 1.3806505e-23;
  static constexpr double avagadro = 6.02214179e23;
  static constexpr double univ_gas_const = 8.314472;
  static constexpr double accom_coeff = 1.0;
  static constexpr double water_eq_vap_press = 611.0;
  static constexpr double water_freeze_temp = 273.15;
  static constexpr double water_surf_eng = 0.073;
  static constexpr double water_latent_heat = 2.272e6;
  static constexpr double air_spec_heat = 1005.0;
  static constexpr double water_molec_weight = 18e-3;
  static constexpr double water_density = 1e3;
  static constexpr double air_molec_weight = 2.89644e-2;
  static constexpr double air_std_press = 101325.0;
  static constexpr double air_dyn_visc = 1.78e-5;
};

// Cold part of the particle record (aero_particle.F90:53-61), carried only after pmcgpu_track_components:
// component(:) = up to PMC_MAX_COMP {source_id, create_time} pairs (MAX_AERO_COMPONENT_SIZE, aero_particle.F90:22)
// and n_opq opaque doubles per particle (the optical fields absorb / scatter_cross_sect, asymmetry, refract_*,
// core_vol, which only the host's optics module interprets).  Rows have the stride of the slab they belong to.
#define PMC_MAX_COMP 20
struct ColdSlab {
  int* n;        // [stride] number of components; nullptr = the cold part is not carried
  int* src;      // [PMC_MAX_COMP][stride]
  double* time;  // [PMC_MAX_COMP][stride]
  double* opq;   // [n_opq][stride]
};

// Device view of one patch: structure-of-arrays particle slabs, one fixed-size
// slab of `cap` slots per grid cell, so no phase needs a global compaction.
struct DevView {
  int n_cells, cap, S, G, C, nbcap;
  int i_water;  // 1-based, 0 none
  unsigned seed;
  long long n_cells_global;
  // particle slabs: live buffer (cur) and scratch buffer (alt) of a ping-pong pair
  double* vol;        // [S][n_cells*cap]
  long long* id;      // [n_cells*cap]
  int* wgc;           // packed (class-1)*G + (group-1)
  int* nprim;
  int* hyst;
  double* lct;
  double* gct;
  double* vol_alt; long long* id_alt; int* wgc_alt; int* nprim_alt; int* hyst_alt; double* lct_alt; double* gct_alt;
  ColdSlab cold, cold_alt;  // stride n_cells*cap; cold.n == nullptr unless pmcgpu_track_components was called
  int n_opq;
  int* count;         // [n_cells]
  long long* next_id; // [n_cells]
  long long* gcell;   // [n_cells] global cell index
  int* status;        // [n_cells] first error
  // sorted state
  int* pbin;          // [n_cells*cap] bin (1-based) of each particle, sorted layout
  int* pbin_alt;
  double* lnr;        // [n_cells*cap] scratch: ln(radius)
  int* bin_count;     // [n_cells][C][nbcap]   entries per (class, bin)
  int* grid_nb;       // [n_cells]
  double* grid_min;   // [n_cells]
  double* grid_max;
  int* bounds_valid;  // [n_cells] coag_kernel_bounds_valid
  double* kmax;       // [n_cells][nbcap*nbcap] (b1,b2) at [(b2-1)*nb + b1-1], nb = grid_nb
  // weights / env
  int* wtype;         // [n_cells][C][G]
  double* wmag;
  double* wexp;
  double* temp; double* pres; double* rh; double* rrho; double* cellvol;
  // species table
  const double* density; const double* kappa; const double* molw; const int* nions;
  double dens_min, dens_max;
  double ideal;       // n_part_ideal per (group, class) when split evenly (pmcgpu_config)
  double* ideal_gc;   // [C][G] n_part_ideal(group, class) (aero_state.F90:241, pmcgpu_set_n_part_ideal)
  // coagulation work space
  int* L;             // [n_cells][3*cap] list arena
  int* fidx;          // [n_cells*cap] position of particle in its (bin, class) list
  int* lptr;          // [n_cells][C*nbcap]
  int* llen;
  int* lcap;
  int* coag_samp;     // [n_cells]
  int* coag_ncoag;
  // condensation work space
  double* cond_T1;    // [n_cells] final temperature / pressure of the step
  double* cond_p1;
  double* cond_h;     // [n_cells] step size carried from one call to the next
  int* cond_stats;    // [n_cells][3] steps, rejected steps, rate evaluations per particle
  double* cond_wre;   // [n_cells] water balance (condense.F90:301-319)
  double* cond_T0;    // [n_cells] env_state_initial of the step (old_env_states, wrf_pmc_driver.F90:120)
  double* cond_p0;
  double* stage;      // [5][n_cells] H2D staging of wrf_to_partmc outputs
  double* w2p;        // [6][n_cells] wrf_to_partmc on the device: T, p, RH, 1/rho, cell volume, height
  float* numconc;     // [n_cells][C] partmc_to_wrf output
  // work counters (pmcgpu_get_work): [0] particle rate evaluations of condense_particles (rate evaluations x
  // particles of the cell, summed over cells and calls), [1] particles condense_particles integrated,
  // [2] particles at the start of partmc_timestep (particle-timesteps), [3] spare
  unsigned long long* work;
  // open (non-periodic) domain: kind of every cell, nullptr for a doubly periodic domain.  0 interior, 1 west,
  // 2 east, 3 south, 4 north edge of the GLOBAL domain, 5 corner.  The outermost ring holds the boundary condition:
  // the aerosol step and the multinomial transport sampling cover interior cells only
  // (wrf_pmc_driver.F90:157-167, wrf_pmc_trans.F90:116-136, wrf_pmc_trans_aero.F90:497-506)
  const unsigned char* kind;
  // aero_data%fractal (fractal.F90:55-66, pmcgpu_set_fractal): fr_on = 0 for spherical particles (the default and
  // the WRF setting); read by the sort, the majorant tables and the fractal build of mc_coag only
  double fr_df, fr_r0, fr_fill;
  int fr_on;
};

// Transport work space of one patch (trans.cu).  "ext" arrays carry a one-cell ghost ring in x and y:
// e = (i + 1) + (nx + 2) * (k + nz * (j + 1)).  Sides: 0 y-, 1 y+, 2 x-, 3 x+.
struct TransView {
  int nx, nz, ny, C;
  float dx, dy, dt;
  int vertmix;
  // WRF fields of the patch (float, WRF memory order, class slowest)
  float *u_flux, *v_flux, *w_flux;  // [(nx+1) nz ny C], [nx nz (ny+1) C], [nx (nz+1) ny C]
  float *xkhh, *alt;                // ext arrays
  float *zc, *z_at_w, *exch_h;      // [nx nz ny], [nx (nz+1) ny] x 2
  int* ccount;                      // [n_cells][C] particles of (group 1, class)
  double* num_old;                  // [n_cells][C] count * magnitude
  double* scale_ratio;              // [n_cells][C] weight_ratio of the preweighting, 0 = none
  double *padv, *pdiff, *pvert;     // [n_cells][C][6 | 4 | nz]
  double* cellvol_ext;              // [n_ext]
  double* wmag_ext;                 // [n_ext][C] magnitudes after the preweighting
  int* out_dir;                     // [n_cells][C][4 + nz][2] offset, count of the out list per direction
  int* out_slot;                    // [n_cells][cap] original slot of every transported particle
  long long* out_newid;             // [n_cells][cap] new id when a copy stays behind, else 0
  int* n_surv;                      // [n_cells]
  int* n_out;                       // [n_cells] entries of the cell's out list (all classes and directions)
  int* n_moved;                     // [n_cells] arrivals of the last call
  double* sbuf[4];                  // edge exchange buffers
  double* rbuf[4];
  int* scnt[4];
  int* soff[4];                     // [L nz + 1] offsets of the ghost cells in the send slab
  int* roff[4];
};
struct TrSlab {  // SoA particle records crossing one side
  int cap;
  double* vol;  // [S][cap]
  long long* id;
  int* wgc; int* nprim; int* hyst;
  double* lct; double* gct;
  ColdSlab cold;  // stride cap; cold.n == nullptr when the cold part is not carried
};

// one record's cold part from slot `src` of slab `from` (row stride sstride) to slot `dst` of slab `to`
__device__ __forceinline__ void cold_copy(const ColdSlab& to, size_t dstride, size_t dst, const ColdSlab& from,
                                          size_t sstride, size_t src, int n_opq) {
  const int n = from.n[src];
  to.n[dst] = n;
  for (int q = 0; q < n; q++) {
    to.src[(size_t)q * dstride + dst] = from.src[(size_t)q * sstride + src];
    to.time[(size_t)q * dstride + dst] = from.time[(size_t)q * sstride + src];
  }
  for (int q = 0; q < n_opq; q++) to.opq[(size_t)q * dstride + dst] = from.opq[(size_t)q * sstride + src];
}
// the same by the lanes of a warp (lane q takes component q; PMC_MAX_COMP <= 32)
__device__ __forceinline__ void cold_copy_warp(const ColdSlab& to, size_t dstride, size_t dst, const ColdSlab& from,
                                               size_t sstride, size_t src, int n_opq, int lane) {
  const int n = from.n[src];
  if (lane < n) {
    to.src[(size_t)lane * dstride + dst] = from.src[(size_t)lane * sstride + src];
    to.time[(size_t)lane * dstride + dst] = from.time[(size_t)lane * sstride + src];
  }
  for (int q = lane; q < n_opq; q += 32) to.opq[(size_t)q * dstride + dst] = from.opq[(size_t)q * sstride + src];
  if (lane == 0) to.n[dst] = n;
}
// a fresh particle (aero_particle_set_component, aero_particle.F90:169-183; optics zero, :114-124)
__device__ __forceinline__ void cold_new(const ColdSlab& to, size_t stride, size_t dst, int source, double create_time,
                                         int n_opq) {
  to.n[dst] = 1;
  to.src[dst] = source;
  to.time[dst] = create_time;
  for (int q = 0; q < n_opq; q++) to.opq[(size_t)q * stride + dst] = 0.0;
}

struct ProfEntry {
  const char* name;
  cudaEvent_t a, b;
};
struct pmcgpu_ctx {
  bool prof_on = false;                 // per-kernel CUDA-event timing (pmcgpu_profile_enable)
  std::vector<ProfEntry> prof_pending;  // recorded, not yet read
  std::vector<cudaEvent_t> ev_pool;
  std::map<std::string, std::pair<double, long long>> prof_acc;  // name -> (ms, launches)
  pmcgpu_config cfg;
  DevView v;
  cudaStream_t stream;
  std::string err;
  long long launches;
  std::vector<void*> allocs;
  std::vector<double> h_density;
  int cur;  // unused: pointers are swapped in v
  int* d_flag;  // first non-zero cell status of the last check
  int* h_flag;  // pinned mirror
  // transport
  TransView t;
  bool trans_ready = false;
  // C4 ghost particles: ONE contiguous message per side = int header (offsets of the side's ghost cells in the
  // slab, last entry = record count) followed by the SoA record arrays with row pitch `pitch`
  // ([S] vol rows, id, lct, gct: 8 bytes; wgc, nprim, hyst: 4 bytes).  The pitch of a step is derived from the
  // counts of the PREVIOUS step (both ends of a link know them: the sender its own count, the receiver the
  // count in the header it received), so no host read of device counts sits inside a transport step.
  char* smsg[4] = {nullptr, nullptr, nullptr, nullptr};
  char* rmsg[4] = {nullptr, nullptr, nullptr, nullptr};
  int slab_cap[4] = {0, 0, 0, 0};   // largest pitch the buffers hold
  int hdr_ints[4] = {0, 0, 0, 0};   // header length (multiple of 4 ints)
  int* h_tot = nullptr;             // pinned: records sent [0..3] / received [4..7] per side in the last step
  cudaEvent_t ev_tot = nullptr;     // completion of the copy into h_tot
  bool tot_valid = false;           // h_tot holds the counts of the previous step
  bool cfg_trans_smem = false, cfg_cond = false, cfg_cond_grid = false, cfg_sort = false;  // per-context (= per-device) opt-ins
  int cond_n_sm = 0, cond_max_blk4 = 0, cond_max_blk8 = 0;
  // side streams of condense_particles: the population-size variants of k_condense run side by side, so that the
  // tail of one launch (SMs idling behind its last cells) is filled by the next
  cudaStream_t cond_stream[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t cond_fork = nullptr, cond_join[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  size_t cfg_msamp_smem = 0, cfg_kmax_smem = 0, cfg_bnd_smem = 0, cfg_sort_smem = 0, cfg_sort_staged_smem = 0;
  // boundary conditions of an open domain (pmcgpu_set_boundary): device copies
  int bnd_n_mode = 0;
  char* bnd_buf = nullptr;
  int *bnd_type = nullptr, *bnd_wclass = nullptr, *bnd_outflow = nullptr, *bnd_throw = nullptr;
  double *bnd_cr = nullptr, *bnd_ls = nullptr, *bnd_vf = nullptr, *bnd_vs = nullptr, *bnd_nc = nullptr;
  bool w2p_ready = false;           // pmcgpu_wrf_to_partmc has staged an env_state for the next partmc_timestep
  int kmax_ktype = 0;               // kernel type the per-cell majorant tables were built for
  // patch neighbours (rank of the patch on side s; this rank itself on a periodic single patch)
  int rank = 0, n_ranks = 1, px = 1, py = 1;
  int nb[4] = {0, 0, 0, 0};
  void* nccl_comm = nullptr;
  int cond_method = 4;  // 4: RODAS4, 5: BDF (CVODE restatement) (pmcgpu_set_condense_method)
  char* dep_buf = nullptr;          // device copies of the 2-D surface fields of the deposition driver
  double* cond_grid_buf = nullptr;  // partial sums + barrier counter of the grid-wide condensation kernel
  unsigned rng_step = 0;     // step number addressing the draws of the stand-alone condensation calls (pmcgpu_set_rng_step)
  bool any_nonflat = false;  // some weight is size dependent (set by pmcgpu_set_weights): aero_state_reweight is not the identity
};
int pmc_exchange_sized(pmcgpu_ctx* ctx, void* const send[4], void* const recv[4], const size_t send_bytes[4],
                       const size_t recv_bytes[4]);
// bytes of one record in a C4 message and the SoA view of a message buffer with row pitch `pitch`
static inline size_t tr_rec_bytes(int S, bool cold = false, int n_opq = 0) {
  return (size_t)8 * S + 8 * 3 + 4 * 3 + (cold ? (size_t)8 * (PMC_MAX_COMP + n_opq) + 4 * (1 + PMC_MAX_COMP) + 4 /* pad */ : 0);
}
static inline TrSlab tr_slab_view(char* msg, int hdr_ints, int pitch, int S, bool cold = false, int n_opq = 0) {
  TrSlab s;
  s.cap = pitch;
  char* p = msg + (size_t)hdr_ints * sizeof(int);
  s.vol = (double*)p;       p += (size_t)8 * S * pitch;
  s.id = (long long*)p;     p += (size_t)8 * pitch;
  s.lct = (double*)p;       p += (size_t)8 * pitch;
  s.gct = (double*)p;       p += (size_t)8 * pitch;
  s.wgc = (int*)p;          p += (size_t)4 * pitch;
  s.nprim = (int*)p;        p += (size_t)4 * pitch;
  s.hyst = (int*)p;         p += (size_t)4 * pitch;
  s.cold.n = nullptr; s.cold.src = nullptr; s.cold.time = nullptr; s.cold.opq = nullptr;
  if (cold) {  // pitch is a multiple of 256: the 8-byte rows stay aligned behind the three 4-byte rows
    p += (size_t)4 * pitch;  // pad to 8 bytes
    s.cold.time = (double*)p; p += (size_t)8 * PMC_MAX_COMP * pitch;
    s.cold.opq = (double*)p;  p += (size_t)8 * n_opq * pitch;
    s.cold.n = (int*)p;       p += (size_t)4 * pitch;
    s.cold.src = (int*)p;
  }
  return s;
}

#define CUDA_OK(ctx, call)                                                              \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                  \
      return PMCGPU_ERR_CUDA;                                                           \
    }                                                                                   \
  } while (0)

#define LAUNCH_CHECK(ctx)                                                               \
  do {                                                                                  \
    (ctx)->launches++;                                                                  \
    cudaError_t e_ = cudaGetLastError();                                                \
    if (e_ != cudaSuccess) {                                                            \
      (ctx)->err = std::string("kernel launch: ") + cudaGetErrorString(e_);             \
      return PMCGPU_ERR_CUDA;                                                           \
    }                                                                                   \
  } while (0)

// a device buffer that lives for one call
template <typename T>
struct TmpDev {
  T* p = nullptr;
  ~TmpDev() {
    if (p) cudaFree(p);
  }
  cudaError_t upload(const T* h, size_t n, cudaStream_t s) {
    if (!h) return cudaSuccess;
    cudaError_t e = cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, s);
  }
  cudaError_t alloc(size_t n) { return cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(T)); }
};

int pmc_first_status(pmcgpu_ctx* ctx);
// brackets the launches of one scope with a pair of events on the context's stream
// entry points whose formulas use the radius of a SPHERICAL particle (or the mobility radius, fractal.F90:239-440,
// which is not built): refused while pmcgpu_set_fractal has a non-spherical setting
#define PMC_SPHERICAL_ONLY(ctx, what)                                                                   \
  do {                                                                                                  \
    if ((ctx) && (ctx)->v.fr_on) {                                                                      \
      (ctx)->err = what ": not built for fractal particles (pmcgpu_set_fractal has a non-spherical setting)"; \
      return PMCGPU_ERR_ARG;                                                                            \
    }                                                                                                   \
  } while (0)

struct ProfScope {
  pmcgpu_ctx* c;
  int idx;
  ProfScope(pmcgpu_ctx* ctx, const char* name) : c(ctx), idx(-1) {
    if (!c->prof_on) return;
    ProfEntry e;
    e.name = name;
    for (cudaEvent_t* ev : {&e.a, &e.b}) {
      if (!c->ev_pool.empty()) {
        *ev = c->ev_pool.back();
        c->ev_pool.pop_back();
      } else {
        cudaEventCreate(ev);
      }
    }
    cudaEventRecord(e.a, c->stream);
    idx = (int)c->prof_pending.size();
    c->prof_pending.push_back(e);
  }
  ~ProfScope() {
    if (idx >= 0) cudaEventRecord(c->prof_pending[idx].b, c->stream);
  }
};  // reduce per-cell status to one code (syncs)

// ---------------------------------------------------------------------------
// device math
// ---------------------------------------------------------------------------
#ifdef __CUDACC__
// cell of an open domain that takes no part in the aerosol step (a boundary-condition cell)
__device__ __forceinline__ bool cell_is_boundary(const DevView& v, int cell) { return v.kind != nullptr && v.kind[cell] != 0; }

// Philox4x32-10 (Salmon et al. SC'11).  Address = (cell, step, phase, a, b, j):
// counter = (j>>1, b, a, cell), key = (seed, step*16 + phase); one block yields
// two 53-bit uniforms (j even / odd).
__device__ __forceinline__ void philox_block(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                             uint32_t k1, uint32_t o[4]) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0;
    const uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  o[0] = c0; o[1] = c1; o[2] = c2; o[3] = c3;
}
struct RKey {
  uint32_t seed, cell, step, phase, a, b;
};
// PMC_RNG_NOINLINE (defined by a translation unit before this header): one out-of-line copy of the
// generator instead of one per call site; kernels with many draw sites (mc_coag) otherwise outgrow the
// instruction cache
#ifdef PMC_RNG_NOINLINE
static __device__ __noinline__ double rng_uniform(const RKey& k, uint32_t j) {
#else
__device__ __forceinline__ double rng_uniform(const RKey& k, uint32_t j) {
#endif
  uint32_t o[4];
  philox_block(j >> 1, k.b, k.a, k.cell, k.seed, k.step * 16u + k.phase, o);
  const uint32_t hi = (j & 1u) ? o[2] : o[0];
  const uint32_t lo = (j & 1u) ? o[3] : o[1];
  const unsigned long long m = ((unsigned long long)(hi >> 5) << 26) | (unsigned long long)(lo >> 6);
  return (double)m * (1.0 / 9007199254740992.0);
}
// rand.F90:172-205 (in-tree branch): mod(int(u*n), n) + 1
__device__ __forceinline__ int pmc_rand_int(double u, int n) {
  const int v = (int)(u * (double)n);
  return (v % n) + 1;
}
// rand.F90:214-226
__device__ __forceinline__ int prob_round(double val, double u) {
  return (u < ceil(val) - val) ? (int)floor(val) : (int)ceil(val);
}
struct RStream {
  RKey k;
  uint32_t j;
  __device__ __forceinline__ double next() { return rng_uniform(k, j++); }
};
// rand.F90:410-474 Box-Muller branch
__device__ inline double rand_normal(double mean, double stddev, RStream& s) {
  const double u1 = s.next();
  const double u2 = s.next();
  const double r = sqrt(-2.0 * log(u1));
  const double theta = 2.0 * PmcConst::pi * u2;
  return stddev * (r * cos(theta)) + mean;
}
// rand.F90:252-304 in-tree branch
#ifdef PMC_RNG_NOINLINE
static __device__ __noinline__ int rand_poisson(double mean, RStream& s) {
#else
__device__ inline int rand_poisson(double mean, RStream& s) {
#endif
  if (mean <= 10.0) {
    const double L = exp(-mean);
    int k = 0;
    double p = 1.0;
    do {
      k++;
      p = p * s.next();
    } while (!(p < L));
    return k - 1;
  }
  const int k = (int)llround(rand_normal(mean, sqrt(mean), s));
  return max(k, 0);
}
// rand.F90:319-404 in-tree branch
__device__ inline int rand_binomial(int n, double p, RStream& s) {
  const double np = (double)n * p;
  const double nomp = (double)n * (1.0 - p);
  if (np >= 10.0 && nomp >= 10.0) {
    const int k = (int)llround(rand_normal(np, sqrt(np * (1.0 - p)), s));
    return min(max(k, 0), n);
  } else if (p < 1e-15) {
    return 0;
  } else if (np < 1e-200) {
    return 0;
  } else if (nomp < 1e-200) {
    return n;
  }
  const double q = (p <= 0.5) ? p : 1.0 - p;
  int k = 0, sum = 0;
  for (;;) {
    const double G_real = log(s.next()) / log(1.0 - q);
    if (G_real > (double)(n - sum)) break;
    const int Gi = (int)ceil(G_real);
    sum += Gi;
    if (sum > n) break;
    k++;
  }
  return (p <= 0.5) ? k : n - k;
}

// util.F90:240 / :265 (spherical particles, fractal.F90:113)
#ifdef PMC_RNG_NOINLINE
static __device__ __noinline__ double vol2rad(double v) { return pow(3.0 * v / 4.0 / PmcConst::pi, 1.0 / 3.0); }
#else
__device__ __forceinline__ double vol2rad(double v) { return pow(3.0 * v / 4.0 / PmcConst::pi, 1.0 / 3.0); }
#endif
__device__ __forceinline__ double rad2vol(double r) { return 4.0 * PmcConst::pi * (r * r * r) / 3.0; }
// fractal.F90:90-100, 106-121, 142-158: number of monomers and geometric radius <-> mass-equivalent volume
__device__ __forceinline__ double fr_monomers(const DevView& v, double vol) {
  const double x = pow(3.0 * vol / 4.0 / PmcConst::pi, 1.0 / 3.0) / v.fr_r0;
  return x * x * x;
}
__device__ __forceinline__ double fr_vol2rad(const DevView& v, double vol) {
  if (!v.fr_on) return pow(3.0 * vol / 4.0 / PmcConst::pi, 1.0 / 3.0);
  return v.fr_r0 * pow(fr_monomers(v, vol) * v.fr_fill, 1.0 / v.fr_df);
}
__device__ __forceinline__ double fr_rad2vol(const DevView& v, double r) {
  if (!v.fr_on) return rad2vol(r);
  return rad2vol(v.fr_r0) * pow(r / v.fr_r0, v.fr_df) / v.fr_fill;
}
// kernel_brown_free_helper (coag_kernel_brown_free.F90:118-147) and kernel_brown_cont_helper
// (coag_kernel_brown_cont.F90:117-144): Eq. 4 and Eq. 6 of Vemury and Pratsinis [1995]
__device__ __forceinline__ double kernel_brown_free(const DevView& v, double v1, double d1, double v2, double d2,
                                                    double temp) {
  const double rho = (v1 / (v1 + v2)) * d1 + (v2 / (v1 + v2)) * d2;
  const double N_i = fr_monomers(v, v1), N_j = fr_monomers(v, v2);
  const double N_i_inv_df = pow(N_i, 1.0 / v.fr_df), N_j_inv_df = pow(N_j, 1.0 / v.fr_df);
  const double t = N_i_inv_df + N_j_inv_df;
  return sqrt(6.0 * PmcConst::boltzmann * temp * v.fr_r0 / rho * (1.0 / N_i + 1.0 / N_j)) * (t * t);
}
__device__ __forceinline__ double kernel_brown_cont(const DevView& v, double v1, double v2, double temp) {
  const double N_i = fr_monomers(v, v1), N_j = fr_monomers(v, v2);
  const double N_i_inv_df = pow(N_i, 1.0 / v.fr_df), N_j_inv_df = pow(N_j, 1.0 / v.fr_df);
  return 2.0 * PmcConst::boltzmann * temp / 3.0 / PmcConst::air_dyn_visc * (1.0 / N_i_inv_df + 1.0 / N_j_inv_df) *
         (N_i_inv_df + N_j_inv_df);
}
// util.F90:663
__device__ __forceinline__ double interp_linear_disc(double x_1, double x_n, int n, int i) {
  if (x_1 == x_n) return x_1;
  const double alpha = (double)(i - 1) / (double)(n - 1);
  return (1.0 - alpha) * x_1 + alpha * x_n;
}
// util.F90:532 on logs (util.F90:566): bin of ln(r) on a log grid with n_bin bins
__device__ __forceinline__ int log_grid_find(double ln_min, double ln_max, int n_bin, double ln_r) {
  const int n = n_bin + 1;
  if (ln_r == ln_max) return n - 1;
  const double f = floor((ln_r - ln_min) / (ln_max - ln_min) * (double)(n - 1));
  long long v = (f > 2.0e9) ? 2000000000LL : (f < -2.0e9 ? -2000000000LL : (long long)f);
  v = v + 1;
  v = v < (long long)n ? v : (long long)n;
  v = v > 0 ? v : 0;
  return (int)v;
}
// bin_grid.F90:83-134 edges = exp(linspace(ln min, ln max, n_bin+1)) with exact ends
#ifdef PMC_RNG_NOINLINE
static __device__ __noinline__ double grid_edge(double gmin, double gmax, int n_bin, int i /*1..n_bin+1*/) {
#else
__device__ __forceinline__ double grid_edge(double gmin, double gmax, int n_bin, int i /*1..n_bin+1*/) {
#endif
  if (i == 1) return gmin;
  if (i == n_bin + 1) return gmax;
  const double a = (double)(i - 1) / (double)n_bin;
  return exp((1.0 - a) * log(gmin) + a * log(gmax));
}
// bin_grid.F90:113-118 centers = logspace(c1, c2, n_bin)
#ifdef PMC_RNG_NOINLINE
static __device__ __noinline__ double grid_center(double gmin, double gmax, int n_bin, int i /*1..n_bin*/) {
#else
__device__ __forceinline__ double grid_center(double gmin, double gmax, int n_bin, int i /*1..n_bin*/) {
#endif
  const double c1 = exp(interp_linear_disc(log(gmin), log(gmax), 2 * n_bin + 1, 2));
  const double c2 = exp(interp_linear_disc(log(gmin), log(gmax), 2 * n_bin + 1, 2 * n_bin));
  if (i == 1) return c1;
  if (i == n_bin) return c2;
  const double a = (double)(i - 1) / (double)(n_bin - 1);
  return exp((1.0 - a) * log(c1) + a * log(c2));
}

// aero_weight.F90:127-146
__device__ __forceinline__ double weight_nc(int type, double mag, double ex, double radius) {
  if (type == PMCGPU_WEIGHT_NONE) return mag;
  return mag * pow(radius, ex);
}
// aero_weight_array.F90:232-251: harmonic combination over groups of class c (1-based)
#ifdef PMC_RNG_NOINLINE
static __device__ __noinline__ double awa_nc_at_radius(const DevView& v, int cell, int c, double radius) {
#else
__device__ __forceinline__ double awa_nc_at_radius(const DevView& v, int cell, int c, double radius) {
#endif
  const size_t base = ((size_t)cell * v.C + (c - 1)) * v.G;
  double s = 0.0;
  for (int g = 0; g < v.G; g++) s += 1.0 / weight_nc(v.wtype[base + g], v.wmag[base + g], v.wexp[base + g], radius);
  return 1.0 / s;
}
__device__ __forceinline__ bool class_is_flat(const DevView& v, int cell, int c) {
  const size_t base = ((size_t)cell * v.C + (c - 1)) * v.G;
  for (int g = 0; g < v.G; g++)
    if (v.wtype[base + g] != PMCGPU_WEIGHT_NONE) return false;
  return true;
}

// coag_kernel_brown.F90:111-197, split into the single-particle part and the
// pair combination (identical operation order, so identical values).
struct BrownEnv {
  double viscosd, gasfreepath, kT;
};
__device__ __forceinline__ BrownEnv brown_env(double temp, double pressure) {
  BrownEnv e;
  const double rhoair = (pressure * PmcConst::air_molec_weight) / (PmcConst::univ_gas_const * temp);
  e.viscosd = 1.8325e-05 * (416.16 / (temp + 120.0)) * pow(temp / 296.16, 1.5);
  const double viscosk = e.viscosd / rhoair;
  const double gasspeed =
      sqrt((8.0 * PmcConst::boltzmann * temp * PmcConst::avagadro) / (PmcConst::pi * PmcConst::air_molec_weight));
  e.gasfreepath = 2.0 * viscosk / gasspeed;
  e.kT = temp;
  return e;
}
struct BrownPart {
  double rad, diffus, speedsq, deltasq;
};
__device__ __forceinline__ BrownPart brown_part(const BrownEnv& e, double vol, double den) {
  BrownPart b;
  const double temp = e.kT;
  b.rad = vol2rad(vol);
  const double Rme = b.rad;
  const double knud = e.gasfreepath / Rme;
  const double cunning = 1.0 + knud * (1.249 + 0.42 * exp(-0.87 / knud));
  b.diffus = (PmcConst::boltzmann * temp * cunning) / (6.0 * PmcConst::pi * Rme * e.viscosd);
  b.speedsq = 8.0 * PmcConst::boltzmann * temp / (PmcConst::pi * den * vol);
  const double freepath = 8.0 * b.diffus / (PmcConst::pi * sqrt(b.speedsq));
  const double t = 2.0 * Rme + freepath;
  const double tmp1 = t * t * t;
  const double tmp2 = pow(4.0 * Rme * Rme + freepath * freepath, 1.5);
  const double dd = (tmp1 - tmp2) / (6.0 * Rme * freepath) - 2.0 * Rme;
  b.deltasq = dd * dd;
  return b;
}
__device__ __forceinline__ double brown_pair(const BrownPart& a, const BrownPart& b) {
  const double rad_sum = a.rad + b.rad;
  const double diffus_sum = a.diffus + b.diffus;
  const double tmp1 = rad_sum / (rad_sum + sqrt(a.deltasq + b.deltasq));
  const double tmp2 = 4.0 * diffus_sum / (rad_sum * sqrt(a.speedsq + b.speedsq));
  return 4.0 * PmcConst::pi * rad_sum * diffus_sum / (tmp1 + tmp2);
}

#endif  // __CUDACC__
