// process.cu -- the per-cell diagnostics of partmc_process (interface/wrf_pmc_driver.F90:671-1288, the part every
// history step needs, :826-944) as device reductions, so that a WRF history alarm does not have to pull the
// particles to the host:
//   tot_num_conc       sum(aero_state_num_concs)                                   :840-841
//   tot_wet_num_conc   aero_state_total_num_conc_wet (aero_state.F90:1287-1308)    :843-844
//   tot_mass_conc      sum(aero_state_masses * num_concs)  (all species, water in) :846-848
//   num_dist, mass_dist  bin_grid_histogram_1d (bin_grid.F90:225-248) of the DRY diameters on a logarithmic
//                      grid (bin_grid_make(LOG, 100, 1d-9, 1d-3), :782), weights num_conc and num_conc * mass
//   num_conc_by_class  aero_state_num_concs_by_source: one weight class per source in WRF (:858-861)
//   species_mass_conc  sum(aero_state_masses(include = species) * num_concs) for every species (:865-930 picks
//                      its named species out of this table on the host)
// One CTA per cell; cell totals by a fixed-order block reduction (deterministic), histograms by shared-memory
// atomics (the order of the additions inside a bin is not fixed: 1e-16 relative).
#include "pmcgpu_internal.cuh"

#define PROC_THREADS 128

struct ProcOut {
  double* tot_num_conc;      // [n_cells]
  double* tot_wet_num_conc;  // [n_cells]
  double* tot_mass_conc;     // [n_cells]
  double* num_dist;          // [n_cells][n_bin]
  double* mass_dist;         // [n_cells][n_bin]
  double* num_conc_by_class; // [n_cells][C]
  double* species_mass_conc; // [n_cells][S]
};

__global__ void __launch_bounds__(PROC_THREADS) k_partmc_process(DevView v, int n_bin, double d_min, double d_max,
                                                                 ProcOut o) {
  const int cell = blockIdx.x, tid = threadIdx.x;
  const int n = v.count[cell];
  const size_t base = (size_t)cell * v.cap, NS = (size_t)v.n_cells * v.cap;
  extern __shared__ double s_p[];
  double* s_num = s_p;                 // [n_bin]
  double* s_mass = s_num + n_bin;      // [n_bin]
  double* s_cls = s_mass + n_bin;      // [C]
  double* s_spec = s_cls + v.C;        // [S]
  double* s_red = s_spec + v.S;        // [3][PROC_THREADS / 32]
  for (int q = tid; q < 2 * n_bin + v.C + v.S; q += PROC_THREADS) s_p[q] = 0.0;
  __syncthreads();
  const int iw = v.i_water - 1;
  // bin_grid_make(LOG): edges = logspace(min, max, n_bin + 1), widths = (ln max - ln min) / n_bin (bin_grid.F90:83-134)
  const double lnmin = log(d_min), lnmax = log(d_max);
  const double width = (lnmax - lnmin) / (double)n_bin;
  double t_num = 0.0, t_wet = 0.0, t_mass = 0.0;
  for (int p = tid; p < n; p += PROC_THREADS) {
    double vdry = 0.0, mass = 0.0, water = 0.0;
    const int gc = v.wgc[base + p];
    double vt = 0.0;
    for (int s = 0; s < v.S; s++) {
      const double x = v.vol[(size_t)s * NS + base + p];
      vt += x;
      mass += x * v.density[s];  // aero_particle_mass (aero_particle.F90:252)
      if (s == iw) water = x; else vdry += x;
    }
    // aero_weight_array_num_conc (aero_weight_array.F90:257-271): harmonic over the groups of the particle's class
    const double nc = awa_nc_at_radius(v, cell, gc / v.G + 1, vol2rad(vt));
    t_num += nc;
    if (water > 0.0) t_wet += nc;
    t_mass += mass * nc;
    atomicAdd(&s_cls[gc / v.G], nc);
    for (int s = 0; s < v.S; s++) atomicAdd(&s_spec[s], v.vol[(size_t)s * NS + base + p] * v.density[s] * nc);
    // bin_grid_find on the dry diameter (aero_particle_dry_diameter, aero_particle.F90:474-536)
    const double dd = 2.0 * vol2rad(vdry);
    const int b = log_grid_find(lnmin, lnmax, n_bin, log(dd));
    if (b >= 1 && b <= n_bin) {
      atomicAdd(&s_num[b - 1], nc / width);
      atomicAdd(&s_mass[b - 1], nc * mass / width);
    }
  }
  // fixed-order block reduction of the three totals
  double r[3] = {t_num, t_wet, t_mass};
  for (int q = 0; q < 3; q++) {
    double t = r[q];
    for (int off = 16; off > 0; off >>= 1) t += __shfl_down_sync(0xffffffffu, t, off);
    if ((tid & 31) == 0) s_red[q * (PROC_THREADS / 32) + (tid >> 5)] = t;
  }
  __syncthreads();
  if (tid == 0) {
    double tot[3];
    for (int q = 0; q < 3; q++) {
      double t = 0.0;
      for (int w = 0; w < PROC_THREADS / 32; w++) t += s_red[q * (PROC_THREADS / 32) + w];
      tot[q] = t;
    }
    o.tot_num_conc[cell] = tot[0];
    o.tot_wet_num_conc[cell] = tot[1];
    o.tot_mass_conc[cell] = tot[2];
  }
  for (int q = tid; q < n_bin; q += PROC_THREADS) {
    o.num_dist[(size_t)cell * n_bin + q] = s_num[q];
    o.mass_dist[(size_t)cell * n_bin + q] = s_mass[q];
  }
  for (int q = tid; q < v.C; q += PROC_THREADS) o.num_conc_by_class[(size_t)cell * v.C + q] = s_cls[q];
  for (int q = tid; q < v.S; q += PROC_THREADS) o.species_mass_conc[(size_t)cell * v.S + q] = s_spec[q];
}

extern "C" int pmcgpu_partmc_process(pmcgpu_ctx* ctx, int n_bin, double d_min, double d_max, double* tot_num_conc,
                                     double* tot_wet_num_conc, double* tot_mass_conc, double* num_dist,
                                     double* mass_dist, double* num_conc_by_class, double* species_mass_conc) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_partmc_process");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n_bin < 1 || n_bin > 2000 || !(d_min > 0.0) || !(d_max > d_min)) return PMCGPU_ERR_ARG;
  CUDA_OK(ctx, cudaSetDevice(ctx->cfg.device));
  DevView& v = ctx->v;
  const size_t nc = (size_t)v.n_cells;
  const size_t n_out = nc * (3 + 2 * (size_t)n_bin + v.C + v.S);
  TmpDev<double> buf;
  CUDA_OK(ctx, buf.alloc(n_out));
  ProcOut o;
  o.tot_num_conc = buf.p;
  o.tot_wet_num_conc = o.tot_num_conc + nc;
  o.tot_mass_conc = o.tot_wet_num_conc + nc;
  o.num_dist = o.tot_mass_conc + nc;
  o.mass_dist = o.num_dist + nc * n_bin;
  o.num_conc_by_class = o.mass_dist + nc * n_bin;
  o.species_mass_conc = o.num_conc_by_class + nc * v.C;
  const size_t smem = (size_t)(2 * n_bin + v.C + v.S + 3 * (PROC_THREADS / 32)) * sizeof(double);
  {
    ProfScope ps_(ctx, "k_partmc_process");
    k_partmc_process<<<v.n_cells, PROC_THREADS, smem, ctx->stream>>>(v, n_bin, d_min, d_max, o);
    LAUNCH_CHECK(ctx);
  }
  cudaStream_t st = ctx->stream;
#define OUT(dst, src, n) \
  if (dst) CUDA_OK(ctx, cudaMemcpyAsync(dst, src, (size_t)(n) * sizeof(double), cudaMemcpyDefault, st));
  OUT(tot_num_conc, o.tot_num_conc, nc) OUT(tot_wet_num_conc, o.tot_wet_num_conc, nc)
  OUT(tot_mass_conc, o.tot_mass_conc, nc) OUT(num_dist, o.num_dist, nc * n_bin) OUT(mass_dist, o.mass_dist, nc * n_bin)
  OUT(num_conc_by_class, o.num_conc_by_class, nc * v.C) OUT(species_mass_conc, o.species_mass_conc, nc * v.S)
#undef OUT
  CUDA_OK(ctx, cudaStreamSynchronize(st));
  return PMCGPU_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// pmcgpu_partmc_process_static: the same diagnostics for callers that may not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scratch that pmcgpu_partmc_process takes from cudaMalloc (the per-cell result
// table the kernel writes before it is copied to the caller's host or device arrays) lives in a file-scope device
// array of PMCGPU_PROCESS_STATIC_MAX_DOUBLES doubles (8 MiB), held by every GPU that loads the library until exit.
//   * n_cells (3 + 2 n_bin + n_class + n_spec) beyond that maximum is rejected with PMCGPU_ERR_CAPACITY.
//   * The array is shared by every call and every context of the process on that GPU and keeps the last call's
//     contents: the kernel writes EVERY entry of the table it is given (totals, all bins, all classes, all species of
//     every cell, zeros for empty cells), so nothing has to be cleared; calls that overlap on the GPU (two contexts
//     = two streams on one device, or two host threads) race on it and must be serialised by the caller.
//   * In host code the name of the array is a host-side stand-in, not a device address: its device address is taken
//     once per device, with that device current, by cudaGetSymbolAddress and handed to the kernel.
// No cudaMalloc / cudaFree / allocating library call on this path (ProfScope is left out: it may create events).
#define PMC_PROC_MAX_DEVICES 64
static __device__ __align__(256) double g_proc_scratch[PMCGPU_PROCESS_STATIC_MAX_DOUBLES];
static double* g_proc_scratch_addr[PMC_PROC_MAX_DEVICES];  // device address per device ordinal, nullptr = not taken yet

extern "C" int pmcgpu_partmc_process_static(pmcgpu_ctx* ctx, int n_bin, double d_min, double d_max,
                                            double* tot_num_conc, double* tot_wet_num_conc, double* tot_mass_conc,
                                            double* num_dist, double* mass_dist, double* num_conc_by_class,
                                            double* species_mass_conc) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_partmc_process_static");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n_bin < 1 || n_bin > 2000 || !(d_min > 0.0) || !(d_max > d_min)) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  const size_t nc = (size_t)v.n_cells;
  const size_t n_out = nc * (3 + 2 * (size_t)n_bin + v.C + v.S);
  if (n_out > (size_t)PMCGPU_PROCESS_STATIC_MAX_DOUBLES) {
    ctx->err = "pmcgpu_partmc_process_static: n_cells (3 + 2 n_bin + n_class + n_spec) = " + std::to_string(n_out) +
               " doubles exceed the static scratch of " + std::to_string((size_t)PMCGPU_PROCESS_STATIC_MAX_DOUBLES) +
               " doubles; use pmcgpu_partmc_process or fewer bins";
    return PMCGPU_ERR_CAPACITY;
  }
  const int dev = ctx->cfg.device;
  if (dev < 0 || dev >= PMC_PROC_MAX_DEVICES) return PMCGPU_ERR_ARG;
  if (g_proc_scratch_addr[dev] == nullptr) {  // set-up, once per device, with that device current
    double* p = nullptr;
    CUDA_OK(ctx, cudaGetSymbolAddress((void**)&p, g_proc_scratch));
    g_proc_scratch_addr[dev] = p;
  }
  double* buf = g_proc_scratch_addr[dev];
  ProcOut o;
  o.tot_num_conc = buf;
  o.tot_wet_num_conc = o.tot_num_conc + nc;
  o.tot_mass_conc = o.tot_wet_num_conc + nc;
  o.num_dist = o.tot_mass_conc + nc;
  o.mass_dist = o.num_dist + nc * n_bin;
  o.num_conc_by_class = o.mass_dist + nc * n_bin;
  o.species_mass_conc = o.num_conc_by_class + nc * v.C;
  const size_t smem = (size_t)(2 * n_bin + v.C + v.S + 3 * (PROC_THREADS / 32)) * sizeof(double);  // < 48 KB
  cudaStream_t st = ctx->stream;
  k_partmc_process<<<v.n_cells, PROC_THREADS, smem, st>>>(v, n_bin, d_min, d_max, o);
  LAUNCH_CHECK(ctx);
#define OUT(dst, src, n) \
  if (dst) CUDA_OK(ctx, cudaMemcpyAsync(dst, src, (size_t)(n) * sizeof(double), cudaMemcpyDefault, st));
  OUT(tot_num_conc, o.tot_num_conc, nc) OUT(tot_wet_num_conc, o.tot_wet_num_conc, nc)
  OUT(tot_mass_conc, o.tot_mass_conc, nc) OUT(num_dist, o.num_dist, nc * n_bin) OUT(mass_dist, o.mass_dist, nc * n_bin)
  OUT(num_conc_by_class, o.num_conc_by_class, nc * v.C) OUT(species_mass_conc, o.species_mass_conc, nc * v.S)
#undef OUT
  CUDA_OK(ctx, cudaStreamSynchronize(st));
  return PMCGPU_OK;
}
