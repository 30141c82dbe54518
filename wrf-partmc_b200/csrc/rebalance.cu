// rebalance.cu -- aero_state_rebalance (partmc/src/aero_state.F90:1765-1825),
// aero_state_double (:1695-1718) and aero_state_halve (:1726-1758) for every
// cell of the patch; one warp per cell.
//
// Doubling appends one copy (new id) of every (group, class) particle in
// particle-index order and halves the weight magnitude.  Halving walks the
// particle array backwards, removing with probability 1/2 by swap-with-last and
// doubles the magnitude; the draw of particle i is addressed
// (PH_HALVE, pass*4096 + (group,class), i), so all draws are evaluated in
// parallel, the swap-with-last permutation is then replayed serially on
// indices only, and the records move in parallel.
#include "pmcgpu_internal.cuh"
#include <cstring>
#include <vector>

#define REB_WARPS 4

__device__ __forceinline__ int warp_sum(int x) {
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

__global__ void __launch_bounds__(REB_WARPS * 32) k_rebalance(DevView v, int allow_doubling, int allow_halving,
                                                               unsigned step) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * REB_WARPS + warp;
  if (cell >= v.n_cells) return;
  if (v.status[cell] != 0 || cell_is_boundary(v, cell)) return;
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  int n_part = v.count[cell];
  int err = 0;
  int* map = v.L + (size_t)cell * 3 * v.cap;  // scratch: current slot -> original slot

  if (allow_doubling) {
    for (int g = 1; g <= v.G && !err; g++)
      for (int c = 1; c <= v.C && !err; c++) {
        const int gc = (c - 1) * v.G + (g - 1);
        const double ideal = v.ideal_gc[gc];
        for (;;) {
          int m = 0;
          for (int p = lane; p < n_part; p += 32) m += (v.wgc[base + p] == gc);
          m = warp_sum(m);
          if (!(((double)m < ideal / 2.0) && (m > 0))) break;
          if (n_part + m > v.cap) {
            err = PMCGPU_ERR_CAPACITY;
            break;
          }
          // aero_state_double: copies appended in particle order, ids in the same order
          const long long nx = v.next_id[cell];
          int run = 0;
          for (int p0 = 0; p0 < n_part; p0 += 32) {
            const int p = p0 + lane;
            const bool match = (p < n_part) && (v.wgc[base + p] == gc);
            const unsigned bm = __ballot_sync(0xffffffffu, match);
            if (match) {
              const int r = run + __popc(bm & ((1u << lane) - 1u));
              const int dst = n_part + r;
              for (int s = 0; s < v.S; s++) v.vol[(size_t)s * NS + base + dst] = v.vol[(size_t)s * NS + base + p];
              v.id[base + dst] = (nx + r - 1) * v.n_cells_global + v.gcell[cell] + 1;
              v.wgc[base + dst] = gc;
              v.nprim[base + dst] = v.nprim[base + p];
              v.hyst[base + dst] = v.hyst[base + p];
              v.lct[base + dst] = v.lct[base + p];
              v.gct[base + dst] = v.gct[base + p];
              if (v.cold.n) cold_copy(v.cold, NS, base + dst, v.cold, NS, base + p, v.n_opq);
            }
            run += __popc(bm);
          }
          __syncwarp();
          n_part += m;
          if (lane == 0) {
            v.next_id[cell] = nx + m;
            v.wmag[(size_t)cell * v.C * v.G + gc] *= 0.5;
          }
          __syncwarp();
        }
      }
  }
  if (allow_halving && !err) {
    for (int g = 1; g <= v.G; g++)
      for (int c = 1; c <= v.C; c++) {
        const int gc = (c - 1) * v.G + (g - 1);
        const double ideal = v.ideal_gc[gc];
        unsigned pass = 0;
        for (;;) {
          int m = 0;
          for (int p = lane; p < n_part; p += 32) m += (v.wgc[base + p] == gc);
          m = warp_sum(m);
          if (!((double)m > ideal * 2.0)) break;
          // removal decisions, all draws in parallel; map[p] = -1 marks a removal
          for (int p = lane; p < n_part; p += 32) {
            int keep = p;
            if (v.wgc[base + p] == gc) {
              const RKey key{v.seed, (unsigned)v.gcell[cell], step, PH_HALVE, pass * 4096u + (unsigned)gc, (unsigned)(p + 1)};
              if (rng_uniform(key, 0) < 0.5) keep = -1;
            }
            map[p] = keep;
          }
          __syncwarp();
          // replay the backward swap-with-last loop on indices only
          int n_cur = n_part;
          if (lane == 0) {
            for (int i = n_part - 1; i >= 0; i--) {
              if (map[i] < 0) {
                if (i != n_cur - 1) map[i] = map[n_cur - 1];
                n_cur--;
              }
            }
          }
          n_cur = __shfl_sync(0xffffffffu, n_cur, 0);
          __syncwarp();
          // sources are original slots >= n_cur, destinations < n_cur: disjoint
          for (int p = lane; p < n_cur; p += 32) {
            const int src = map[p];
            if (src != p) {
              for (int s = 0; s < v.S; s++) v.vol[(size_t)s * NS + base + p] = v.vol[(size_t)s * NS + base + src];
              v.id[base + p] = v.id[base + src];
              v.wgc[base + p] = v.wgc[base + src];
              v.nprim[base + p] = v.nprim[base + src];
              v.hyst[base + p] = v.hyst[base + src];
              v.lct[base + p] = v.lct[base + src];
              v.gct[base + p] = v.gct[base + src];
              if (v.cold.n) cold_copy(v.cold, NS, base + p, v.cold, NS, base + src, v.n_opq);
            }
          }
          __syncwarp();
          n_part = n_cur;
          if (lane == 0) v.wmag[(size_t)cell * v.C * v.G + gc] *= 2.0;
          __syncwarp();
          pass++;
        }
      }
  }
  if (lane == 0) {
    v.count[cell] = n_part;
    if (err && v.status[cell] == 0) v.status[cell] = err;
  }
}

// aero_state_reweight (aero_state.F90:597-653) after a pass that changed particle sizes
// (condense_particles :297, condense_equilib_particles :767).  The pass left, per particle, the
// number concentration before the change in v.lnr (aero_state_num_conc_for_reweight :564-587) and the
// new radius in v.gct_alt.  Per (group, class) the magnitude is rescaled to keep the computational
// count in expectation; then, backwards through the particle array, particle i keeps
// prob_round(old_nc / new_nc) copies (aero_state_dup_particle :488-539): 0 copies = swap-with-last
// removal, extra copies are appended with new ids.  The draw of particle i is addressed
// (PH_REWEIGHT, pass, i), so a chunk of 32 particles is decided in parallel and only the (rare)
// particles whose count is not 1 are replayed in order.  Flat weights: ratio 1 exactly, nothing to do.
__global__ void __launch_bounds__(REB_WARPS * 32) k_reweight(DevView v, unsigned step, unsigned pass) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * REB_WARPS + warp;
  if (cell >= v.n_cells) return;
  if (v.status[cell] != 0 || cell_is_boundary(v, cell)) return;
  const int ngc = v.C * v.G;
  const size_t wb = (size_t)cell * ngc;
  bool nonflat = false;
  for (int q = 0; q < ngc; q++)
    if (v.wtype[wb + q] != PMCGPU_WEIGHT_NONE) nonflat = true;
  if (!nonflat) return;
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  const int n0 = v.count[cell];
  int n_part = n0;
  int err = 0;
  // ---- magnitudes: n_part_new / n_part_old per (group, class)
  for (int gc = 0; gc < ngc; gc++) {
    double sn = 0.0;
    int cnt = 0;
    const int wt = v.wtype[wb + gc];
    const double wm = v.wmag[wb + gc], we = v.wexp[wb + gc];
    for (int p = lane; p < n0; p += 32)
      if (v.wgc[base + p] == gc) {
        sn += v.lnr[base + p] / weight_nc(wt, wm, we, v.gct_alt[base + p]);
        cnt++;
      }
    for (int o = 16; o > 0; o >>= 1) sn += __shfl_xor_sync(0xffffffffu, sn, o);
    cnt = warp_sum(cnt);
    __syncwarp();
    if (lane == 0 && cnt > 0) v.wmag[wb + gc] = wm * (sn / (double)cnt);
  }
  __syncwarp();
  // ---- particles, backwards
  for (int hi = n0 - 1; hi >= 0 && !err; hi -= 32) {
    const int i = hi - lane;
    int n_copies = 1;
    if (i >= 0) {
      const int gc = v.wgc[base + i];
      const double new_nc = weight_nc(v.wtype[wb + gc], v.wmag[wb + gc], v.wexp[wb + gc], v.gct_alt[base + i]);
      const RKey key{v.seed, (unsigned)v.gcell[cell], step, PH_REWEIGHT, pass, (unsigned)(i + 1)};
      n_copies = prob_round(v.lnr[base + i] / new_nc, rng_uniform(key, 0));
    }
    unsigned todo = __ballot_sync(0xffffffffu, n_copies != 1);
    while (todo && !err) {
      const int l = __ffs(todo) - 1;
      todo &= todo - 1;
      const int ii = hi - l;
      const int nc = __shfl_sync(0xffffffffu, n_copies, l);
      if (nc == 0) {
        const int last = n_part - 1;
        if (ii != last) {
          for (int sp = lane; sp < v.S; sp += 32) v.vol[(size_t)sp * NS + base + ii] = v.vol[(size_t)sp * NS + base + last];
          if (lane == 0) {
            v.id[base + ii] = v.id[base + last];
            v.wgc[base + ii] = v.wgc[base + last];
            v.nprim[base + ii] = v.nprim[base + last];
            v.hyst[base + ii] = v.hyst[base + last];
            v.lct[base + ii] = v.lct[base + last];
            v.gct[base + ii] = v.gct[base + last];
          }
          if (v.cold.n) cold_copy_warp(v.cold, NS, base + ii, v.cold, NS, base + last, v.n_opq, lane);
        }
        n_part = last;
      } else {
        if (n_part + nc - 1 > v.cap) {
          err = PMCGPU_ERR_CAPACITY;
          break;
        }
        const long long nx = v.next_id[cell];
        for (int d = 0; d < nc - 1; d++) {
          const int dst = n_part + d;
          for (int sp = lane; sp < v.S; sp += 32) v.vol[(size_t)sp * NS + base + dst] = v.vol[(size_t)sp * NS + base + ii];
          if (lane == 0) {
            v.id[base + dst] = (nx + d - 1) * v.n_cells_global + v.gcell[cell] + 1;
            v.wgc[base + dst] = v.wgc[base + ii];
            v.nprim[base + dst] = v.nprim[base + ii];
            v.hyst[base + dst] = v.hyst[base + ii];
            v.lct[base + dst] = v.lct[base + ii];
            v.gct[base + dst] = v.gct[base + ii];
          }
          if (v.cold.n) cold_copy_warp(v.cold, NS, base + dst, v.cold, NS, base + ii, v.n_opq, lane);
        }
        __syncwarp();
        if (lane == 0) v.next_id[cell] = nx + (nc - 1);
        n_part += nc - 1;
      }
      __syncwarp();
    }
  }
  if (lane == 0) {
    v.count[cell] = n_part;
    if (err && v.status[cell] == 0) v.status[cell] = err;
  }
}

int pmc_reweight_async(pmcgpu_ctx* ctx, unsigned step, unsigned pass) {
  DevView& v = ctx->v;
  const int blocks = (v.n_cells + REB_WARPS - 1) / REB_WARPS;
  ProfScope ps_(ctx, "k_reweight");
  k_reweight<<<blocks, REB_WARPS * 32, 0, ctx->stream>>>(v, step, pass);
  LAUNCH_CHECK(ctx);
  return PMCGPU_OK;
}

// aero_state_scale_weight (aero_state.F90:1832-1895) for (group g, class c) of every cell, ratio[cell]
// (<= 0 or 1: leave the cell alone).  ratio > 1: magnitude x ratio and prob_round(n (1 - 1/ratio)) uniformly
// chosen particles of the (group, class) list removed one after the other (swap-with-last in the particle
// array AND in the list, integer_rmap2_remove :190-243) -- the draws are addressed (PH_SCALE, gc, removal
// number), the index surgery is replayed by one lane on three integer arrays, the records then move in
// parallel.  ratio < 1: magnitude x ratio and every particle of the list, backwards, keeps
// prob_round(1 / ratio) copies (aero_state_dup_particle :488-539): decided 32 at a time, copies appended in
// the reference's order with consecutive new ids.
// salt: distinguishes the draw addresses of several calls for the same (group, class) inside one step
__device__ void cell_scale_weight(const DevView& v, int cell, int lane, int g, int c, double weight_ratio,
                                  int allow_doubling, int allow_halving, unsigned step, unsigned salt) {
  if (!(weight_ratio > 0.0) || weight_ratio == 1.0) return;
  const int gc = (c - 1) * v.G + (g - 1);
  const unsigned gca = (unsigned)gc + 4096u * salt;  // draw address
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  const int n0 = v.count[cell];
  int* list = v.L + (size_t)cell * 3 * v.cap;  // slots of the (group, class) particles, ascending
  int* listpos = list + v.cap;                  // slot -> position in list, -1 if not in it
  int* orig = listpos + v.cap;                  // current slot -> original slot
  int n_part = 0;
  for (int p0 = 0; p0 < n0; p0 += 32) {
    const int p = p0 + lane;
    const bool is = (p < n0) && (v.wgc[base + p] == gc);
    const unsigned b = __ballot_sync(0xffffffffu, is);
    if (p < n0) {
      orig[p] = p;
      listpos[p] = -1;
    }
    if (is) {
      const int idx = n_part + __popc(b & ((1u << lane) - 1u));
      list[idx] = p;
      listpos[p] = idx;
    }
    n_part += __popc(b);
  }
  __syncwarp();
  const unsigned gcell = (unsigned)v.gcell[cell];
  const size_t wi = (size_t)cell * v.C * v.G + gc;
  int err = 0;
  if ((weight_ratio > 1.0) && (allow_halving || (n_part == 0))) {
    const int n_remove = prob_round((double)n_part * (1.0 - 1.0 / weight_ratio),
                                    rng_uniform(RKey{v.seed, gcell, step, PH_SCALE, gca, 0u}, 0));
    int n_cur = n0;
    if (lane == 0) {
      v.wmag[wi] *= weight_ratio;
      int len = n_part;
      for (int r = 1; r <= n_remove; r++) {
        const int ie = pmc_rand_int(rng_uniform(RKey{v.seed, gcell, step, PH_SCALE, gca, (unsigned)r}, 0), len);
        const int i = list[ie - 1];
        const int old_pos = listpos[i];
        const int last = n_cur - 1;
        if (last != i) {
          if (listpos[last] >= 0) list[listpos[last]] = i;
          orig[i] = orig[last];
          listpos[i] = listpos[last];
        }
        n_cur = last;
        const int idx_fix = len - 1;
        const int dom_fix = list[idx_fix];
        if (old_pos < idx_fix) list[old_pos] = list[idx_fix];
        len = idx_fix;
        if (idx_fix != old_pos) listpos[dom_fix] = old_pos;
      }
    }
    n_cur = __shfl_sync(0xffffffffu, n_cur, 0);
    __syncwarp();
    // sources are original slots >= n_cur, destinations < n_cur: disjoint
    for (int p = lane; p < n_cur; p += 32) {
      const int src = orig[p];
      if (src != p) {
        for (int s = 0; s < v.S; s++) v.vol[(size_t)s * NS + base + p] = v.vol[(size_t)s * NS + base + src];
        v.id[base + p] = v.id[base + src];
        v.wgc[base + p] = v.wgc[base + src];
        v.nprim[base + p] = v.nprim[base + src];
        v.hyst[base + p] = v.hyst[base + src];
        v.lct[base + p] = v.lct[base + src];
        v.gct[base + p] = v.gct[base + src];
        if (v.cold.n) cold_copy(v.cold, NS, base + p, v.cold, NS, base + src, v.n_opq);
      }
    }
    if (lane == 0) v.count[cell] = n_cur;
  } else if ((weight_ratio < 1.0) && (allow_doubling || (n_part == 0))) {
    if (lane == 0) v.wmag[wi] *= weight_ratio;
    int n_cur = n0;
    long long nx = v.next_id[cell];
    for (int hi = n_part; hi >= 1 && !err; hi -= 32) {
      const int ie = hi - lane;  // 1-based list entry
      int extra = 0, src = 0;
      if (ie >= 1) {
        src = list[ie - 1];
        const RKey key{v.seed, gcell, step, PH_SCALE, gca, 0x40000000u + (unsigned)ie};
        extra = prob_round(1.0 / weight_ratio, rng_uniform(key, 0)) - 1;
      }
      // exclusive prefix over the lanes (lane 0 = first processed entry)
      int incl = extra;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      const int tot = __shfl_sync(0xffffffffu, incl, 31);
      if (n_cur + tot > v.cap) {
        err = PMCGPU_ERR_CAPACITY;
        break;
      }
      const int first = n_cur + incl - extra;
      for (int d = 0; d < extra; d++) {
        const int dst = first + d;
        for (int s = 0; s < v.S; s++) v.vol[(size_t)s * NS + base + dst] = v.vol[(size_t)s * NS + base + src];
        v.id[base + dst] = (nx + (incl - extra) + d - 1) * v.n_cells_global + v.gcell[cell] + 1;
        v.wgc[base + dst] = gc;
        v.nprim[base + dst] = v.nprim[base + src];
        v.hyst[base + dst] = v.hyst[base + src];
        v.lct[base + dst] = v.lct[base + src];
        v.gct[base + dst] = v.gct[base + src];
        if (v.cold.n) cold_copy(v.cold, NS, base + dst, v.cold, NS, base + src, v.n_opq);
      }
      n_cur += tot;
      nx += tot;
    }
    __syncwarp();
    if (lane == 0) {
      v.count[cell] = n_cur;
      v.next_id[cell] = nx;
      if (err && v.status[cell] == 0) v.status[cell] = err;
    }
  }
}

__global__ void __launch_bounds__(REB_WARPS * 32) k_scale_weight(DevView v, int g, int c, const double* __restrict__ ratio,
                                                                  int allow_doubling, int allow_halving, unsigned step) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * REB_WARPS + warp;
  if (cell >= v.n_cells) return;
  if (v.status[cell] != 0) return;
  cell_scale_weight(v, cell, lane, g, c, ratio[cell], allow_doubling, allow_halving, step, 0u);
}

extern "C" int pmcgpu_aero_state_scale_weight(pmcgpu_ctx* ctx, int i_group, int i_class, const double* weight_ratio,
                                              int allow_doubling, int allow_halving, unsigned step) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !weight_ratio) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  if (i_group < 1 || i_group > v.G || i_class < 1 || i_class > v.C) return PMCGPU_ERR_ARG;
  // v.stage: [n_cells] staging doubles (also used by partmc_timestep for the new inverse density)
  CUDA_OK(ctx, cudaMemcpyAsync(v.stage, weight_ratio, v.n_cells * sizeof(double), cudaMemcpyDefault, ctx->stream));
  {
    ProfScope ps_(ctx, "k_scale_weight");
    const int blocks = (v.n_cells + REB_WARPS - 1) / REB_WARPS;
    k_scale_weight<<<blocks, REB_WARPS * 32, 0, ctx->stream>>>(v, i_group, i_class, v.stage, allow_doubling, allow_halving, step);
    LAUNCH_CHECK(ctx);
  }
  return pmc_first_status(ctx);
}

// ---------------------------------------------------------------------------------------------------
// aero_state_add_aero_dist_sample (aero_state.F90:758-831): emission / inflow sampling of new particles
// (SURVEY.md section 8f, N1).  Per cell, per (group, mode) in the reference's order:
// aero_state_prepare_weight_for_add (:703-757, PMC_USE_WRF branch) -> cell_scale_weight when the new count
// would leave [ideal / 2, 2 ideal]; Poisson number of new particles (draw (PH_EMIT, (group, mode), 0));
// every new particle i has its own draw stream (PH_EMIT, (group, mode), i), so 32 particles are sampled at a
// time: radius from the mode (aero_mode.F90:549-628, log-normal with the tail threshold of
// rand.F90:410-474, exponential, mono-disperse), volume fractions by the accept-reject loop of :633-668.
struct EmitDev {
  int n_mode;
  const int* type;             // [n_mode] 1 log-normal, 2 exponential, 3 mono-disperse
  const int* wclass;           // [n_mode] weight class of the mode's source (1-based)
  const double* char_radius;   // [n_mode]
  const double* log10_std;     // [n_mode]
  const double* vol_frac;      // [n_mode][S]
  const double* vol_frac_std;  // [n_mode][S]
  const double* num_conc;      // [n_cells][n_mode]
  const double* sample_prop;   // [n_cells]
  double char_factor, create_time;
  int allow_doubling, allow_halving;
  unsigned step;
  unsigned stream;             // added to the draw address a = (group, mode): keeps several callers of one step apart
  int* n_add;                  // [n_cells]
  // sampled (tabulated) modes, type 4 (aero_mode.F90:37): per mode n_sbin + 1 bin-edge radii and n_sbin number
  // concentrations; num_conc[cell][mode] is then a dimensionless factor on the table
  int n_sbin;
  const double* sample_radius;    // [n_mode][n_sbin + 1]
  const double* sample_num_conc;  // [n_mode][n_sbin]
};
struct SampTab {
  int nbin;
  const double* rad;
  const double* nc;
};
// aero_mode_weighted_sampled_num_conc (aero_mode.F90:456-491), entry k (0-based)
__device__ __forceinline__ double sampled_weighted_nc(const SampTab& t, int k, int wtype, double wexp) {
  if (wtype == PMCGPU_WEIGHT_NONE) return t.nc[k];
  const double x0 = log(t.rad[k]), x1 = log(t.rad[k + 1]);
  return t.nc[k] / wexp * (exp(-wexp * x0) - exp(-wexp * x1)) / (x1 - x0);
}

// aero_mode_number (aero_mode.F90:496-544)
__device__ __forceinline__ double aero_mode_number(int type, double num_conc, double char_radius, double lsd, int wtype,
                                                   double wmag, double wexp, const SampTab& tab) {
  if (type == 4) {  // :533-538: sum(weighted_num_conc) / magnitude, scaled by the cell's factor
    double sum = 0.0;
    for (int k = 0; k < tab.nbin; k++) sum = sum + sampled_weighted_nc(tab, k, wtype, wexp);
    return num_conc * (sum / wmag);
  }
  if (type == 1) {
    if (wtype == PMCGPU_WEIGHT_NONE) return num_conc / wmag;
    const double l10 = log10(char_radius);
    const double x_mean_prime = l10 - wexp * (lsd * lsd) * log(10.0);
    return num_conc / wmag * exp((x_mean_prime * x_mean_prime - l10 * l10) / (2.0 * (lsd * lsd)));
  } else if (type == 2) {
    return num_conc / wmag;
  }
  return num_conc / weight_nc(wtype, wmag, wexp, char_radius);
}

// aero_mode_sample_radius (aero_mode.F90:549-628: log-normal with the tail threshold of rand.F90:410-474,
// exponential, mono-disperse) + aero_mode_sample_vols (:633-668) for one new particle with its own draw stream;
// vols[S] receives the species volumes, the radius is returned
__device__ double sample_mode_particle(int S, int type, double cr, double lsd, int wt, double we, double size_factor,
                                       const double* __restrict__ vf, const double* __restrict__ vs, RStream& s,
                                       double* vols, const SampTab& tab = SampTab{0, nullptr, nullptr},
                                       double wm = 1.0, double fr_df = 0.0, double fr_r0 = 0.0, double fr_fill = 0.0) {
  // fr_df > 0: fractal particles, aero_data_rad2vol / aero_data_vol2rad (aero_mode.F90:611, aero_state.F90:817)
  // aero_mode_sample_radius
  double radius = cr;
  if (type == 4) {
    // sampled mode (:581-608): bin by sample_cts_pdf (rand.F90:550-576) on the weighted table, then a radius
    // log-uniform in the bin (flat weight) or uniform in 1 / num_conc (power weight)
    double pdf_max = 0.0;
    for (int k = 0; k < tab.nbin; k++) pdf_max = fmax(pdf_max, sampled_weighted_nc(tab, k, wt, we));
    int ks;
    for (;;) {
      ks = pmc_rand_int(s.next(), tab.nbin) - 1;
      if (s.next() < sampled_weighted_nc(tab, ks, wt, we) / pdf_max) break;
    }
    const double r = s.next();
    if (wt == PMCGPU_WEIGHT_NONE || we == 0.0) {
      const double x0 = log(tab.rad[ks]), x1 = log(tab.rad[ks + 1]);
      radius = exp((1.0 - r) * x0 + r * x1);
    } else {
      const double inv_nc0 = 1.0 / weight_nc(wt, wm, we, tab.rad[ks]);
      const double inv_nc1 = 1.0 / weight_nc(wt, wm, we, tab.rad[ks + 1]);
      const double inv_nc = (1.0 - r) * inv_nc0 + r * inv_nc1;
      radius = exp(log((1.0 / inv_nc) / wm) / we);  // aero_weight_radius_at_num_conc (aero_weight.F90:170-191)
    }
  } else if (type == 1) {
    double x_mean_prime = log10(cr);
    if (wt != PMCGPU_WEIGHT_NONE) x_mean_prime = log10(cr) - we * (lsd * lsd) * log(10.0);
    const double sqrt2 = (double)sqrtf(2.0f);  // 2.0**.5 is a default-real expression in rand.F90:463
    double x;
    for (;;) {
      const double u1 = s.next();
      const double u2 = s.next();
      const double r = sqrt(-2.0 * log(u1));
      const double theta = 2.0 * PmcConst::pi * u2;
      const double z0 = r * cos(theta);
      if (1.0 - fabs(erf(z0 / sqrt2)) >= size_factor) {
        x = lsd * z0 + x_mean_prime;
        break;
      }
    }
    radius = pow(10.0, x);
  } else if (type == 2) {
    if (fr_df > 0.0) {
      const double v_exp = -(rad2vol(fr_r0) * pow(cr / fr_r0, fr_df) / fr_fill) * log(s.next());
      const double x = pow(3.0 * v_exp / 4.0 / PmcConst::pi, 1.0 / 3.0) / fr_r0;
      radius = fr_r0 * pow((x * x * x) * fr_fill, 1.0 / fr_df);
    } else {
      radius = vol2rad(-rad2vol(cr) * log(s.next()));
    }
  }
  const double total_vol = (fr_df > 0.0) ? rad2vol(fr_r0) * pow(radius / fr_r0, fr_df) / fr_fill : rad2vol(radius);
  // aero_mode_sample_vols
    for (int it = 0; it < 1000000; it++) {
    double sum = 0.0;
    for (int k = 0; k < S; k++) {
      vols[k] = rand_normal(vf[k], vs[k], s);
      sum += vols[k];
    }
    const double offset = 1.0 - sum;
    double mn = 1e300;
    for (int k = 0; k < S; k++) {
      vols[k] = vols[k] + offset * vf[k];
      mn = fmin(mn, vols[k]);
    }
    if (mn >= 0.0) break;
  }
  double sum = 0.0;
  for (int k = 0; k < S; k++) sum += vols[k];
  for (int k = 0; k < S; k++) vols[k] = vols[k] / sum * total_vol;
  return radius;
}

// aero_state_add_aero_dist_sample for ONE cell (warp-cooperative); mode_nc[n_mode] = number concentration of every
// mode in this cell.  Returns the number of particles added; err receives a status.
__device__ int emit_cell(DevView& v, const EmitDev& e, int cell, int lane, double sample_prop,
                         const double* __restrict__ mode_nc_row, int& err) {
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  const unsigned gcell = (unsigned)v.gcell[cell];
  int n_add = 0;
  for (int g = 1; g <= v.G && !err; g++)
    for (int im = 1; im <= e.n_mode && !err; im++) {
      const int type = e.type[im - 1], c = e.wclass[im - 1];
      const double cr = e.char_radius[im - 1], lsd = e.log10_std[im - 1];
      const double mode_nc = mode_nc_row[im - 1];
      const SampTab tab{type == 4 ? e.n_sbin : 0,
                        type == 4 ? e.sample_radius + (size_t)(im - 1) * (e.n_sbin + 1) : nullptr,
                        type == 4 ? e.sample_num_conc + (size_t)(im - 1) * e.n_sbin : nullptr};
      const unsigned a = e.stream + (unsigned)((g - 1) * e.n_mode + (im - 1));
      const int gc = (c - 1) * v.G + (g - 1);
      const double ideal = v.ideal_gc[gc];
      const size_t wi = (size_t)cell * v.C * v.G + gc;
      if (type == 2 && v.wtype[wi] != PMCGPU_WEIGHT_NONE) {
        err = PMCGPU_ERR_ARG;  // "cannot use exponential modes with weighting"
        break;
      }
      __syncwarp();
      double n_samp_avg = sample_prop * aero_mode_number(type, mode_nc, cr, lsd, v.wtype[wi], v.wmag[wi], v.wexp[wi], tab);
      {  // aero_state_prepare_weight_for_add
        const int n_now = v.count[cell];
        int m = 0;
        for (int p = lane; p < n_now; p += 32) m += (v.wgc[base + p] == gc);
        m = warp_sum(m);
        const double n_part_new = (double)m + n_samp_avg;
        if (n_part_new != 0.0 && ((n_part_new < ideal / 2.0) || (n_part_new > ideal * 2.0))) {
          __syncwarp();
          cell_scale_weight(v, cell, lane, g, c, n_part_new / ideal, e.allow_doubling, e.allow_halving, e.step, 1u + a);
          __syncwarp();
          if (v.status[cell] != 0) return n_add;
        }
      }
      if (n_samp_avg == 0.0) continue;
      const int wt = v.wtype[wi];
      const double wm = v.wmag[wi], we = v.wexp[wi];
      n_samp_avg = sample_prop * aero_mode_number(type, mode_nc, cr, lsd, wt, wm, we, tab);
      RStream s0{RKey{v.seed, gcell, e.step, PH_EMIT, a, 0u}, 0u};
      const int n_samp = rand_poisson(n_samp_avg, s0);
      const double size_factor = fmin(1.0 / (e.char_factor * n_samp_avg), 0.32);
      const int n_cur = v.count[cell];
      if (n_cur + n_samp > v.cap) {
        err = PMCGPU_ERR_CAPACITY;
        break;
      }
      const long long nx = v.next_id[cell];
      const double* vf = e.vol_frac + (size_t)(im - 1) * v.S;
      const double* vs = e.vol_frac_std + (size_t)(im - 1) * v.S;
      for (int i_samp = 1 + lane; i_samp <= n_samp; i_samp += 32) {
        RStream s{RKey{v.seed, gcell, e.step, PH_EMIT, a, (unsigned)i_samp}, 0u};
        double vols[PMC_MAX_SPEC];
        const double radius = sample_mode_particle(v.S, type, cr, lsd, wt, we, size_factor, vf, vs, s, vols, tab, wm,
                                                   v.fr_on ? v.fr_df : 0.0, v.fr_r0, v.fr_fill);
        (void)radius;
        const int slot = n_cur + i_samp - 1;
        for (int k = 0; k < v.S; k++) v.vol[(size_t)k * NS + base + slot] = vols[k];
        v.id[base + slot] = (nx + (i_samp - 1) - 1) * v.n_cells_global + v.gcell[cell] + 1;
        v.wgc[base + slot] = gc;
        v.nprim[base + slot] = 1;
        v.hyst[base + slot] = 0;
        v.lct[base + slot] = e.create_time;
        v.gct[base + slot] = e.create_time;
        // aero_particle_set_component (aero_state.F90:823); the source of a mode is its weight class (one class
        // per source in WRF, aero_state_weight_class_for_source)
        if (v.cold.n) cold_new(v.cold, NS, base + slot, c, e.create_time, v.n_opq);
      }
      __syncwarp();
      if (lane == 0) {
        v.count[cell] = n_cur + n_samp;
        v.next_id[cell] = nx + n_samp;
      }
      __syncwarp();
      n_add += n_samp;
    }
  return n_add;
}

__global__ void __launch_bounds__(REB_WARPS * 32) k_emit(DevView v, EmitDev e) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * REB_WARPS + warp;
  if (cell >= v.n_cells) return;
  if (v.status[cell] != 0) return;
  int err = 0;
  const int n_add = emit_cell(v, e, cell, lane, e.sample_prop[cell], e.num_conc + (size_t)cell * e.n_mode, err);
  if (lane == 0) {
    if (e.n_add) e.n_add[cell] = n_add;
    if (err && v.status[cell] == 0) v.status[cell] = err;
  }
}

static int emit_run(pmcgpu_ctx* ctx, const pmcgpu_emit_args* a, int* n_part_add);
extern "C" int pmcgpu_aero_state_add_aero_dist_sample(pmcgpu_ctx* ctx, const pmcgpu_emit_args* a, int* n_part_add) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  return emit_run(ctx, a, n_part_add);
}
static int emit_run(pmcgpu_ctx* ctx, const pmcgpu_emit_args* a, int* n_part_add) {
  if (!ctx || !a || a->n_mode < 1 || !a->mode_type || !a->weight_class || !a->char_radius || !a->log10_std_dev ||
      !a->vol_frac || !a->num_conc || !a->sample_prop || !(a->characteristic_factor > 0.0))
    return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  const size_t nm = (size_t)a->n_mode, S = (size_t)v.S, nc = (size_t)v.n_cells;
  bool any_sampled = false;
  for (size_t m = 0; m < nm; m++) {
    if (a->weight_class[m] < 1 || a->weight_class[m] > v.C || a->mode_type[m] < 1 || a->mode_type[m] > 4) return PMCGPU_ERR_ARG;
    any_sampled = any_sampled || a->mode_type[m] == 4;
  }
  if (any_sampled && (a->n_sample_bins < 1 || a->n_sample_bins > 4096 || !a->sample_radius || !a->sample_num_conc)) {
    ctx->err = "sampled mode (type 4) without its table (n_sample_bins, sample_radius, sample_num_conc)";
    return PMCGPU_ERR_ARG;
  }
  const size_t nsb = any_sampled ? (size_t)a->n_sample_bins : 0;
  // one device buffer per call (mode tables are tiny; num_conc is [n_cells][n_mode])
  const size_t bytes = (2 * nm) * sizeof(int) + (2 * nm + 2 * nm * S + nc * nm + nc + nm * (2 * nsb + 1)) * sizeof(double) +
                       nc * sizeof(int) + 64;
  char* buf = nullptr;
  CUDA_OK(ctx, cudaMalloc((void**)&buf, bytes));
  double* d_cr = (double*)buf;
  double* d_ls = d_cr + nm;
  double* d_vf = d_ls + nm;
  double* d_vs = d_vf + nm * S;
  double* d_nc = d_vs + nm * S;
  double* d_sp = d_nc + nc * nm;
  double* d_sr = d_sp + nc;                // [n_mode][nsb + 1]
  double* d_sn = d_sr + nm * (nsb + 1);    // [n_mode][nsb]
  int* d_ty = (int*)(d_sn + nm * nsb);
  int* d_wc = d_ty + nm;
  int* d_na = d_wc + nm;
  cudaStream_t st = ctx->stream;
  CUDA_OK(ctx, cudaMemcpyAsync(d_cr, a->char_radius, nm * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_ls, a->log10_std_dev, nm * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_vf, a->vol_frac, nm * S * sizeof(double), cudaMemcpyDefault, st));
  if (a->vol_frac_std)
    CUDA_OK(ctx, cudaMemcpyAsync(d_vs, a->vol_frac_std, nm * S * sizeof(double), cudaMemcpyDefault, st));
  else
    CUDA_OK(ctx, cudaMemsetAsync(d_vs, 0, nm * S * sizeof(double), st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_nc, a->num_conc, nc * nm * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_sp, a->sample_prop, nc * sizeof(double), cudaMemcpyDefault, st));
  if (any_sampled) {
    CUDA_OK(ctx, cudaMemcpyAsync(d_sr, a->sample_radius, nm * (nsb + 1) * sizeof(double), cudaMemcpyDefault, st));
    CUDA_OK(ctx, cudaMemcpyAsync(d_sn, a->sample_num_conc, nm * nsb * sizeof(double), cudaMemcpyDefault, st));
  }
  CUDA_OK(ctx, cudaMemcpyAsync(d_ty, a->mode_type, nm * sizeof(int), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_wc, a->weight_class, nm * sizeof(int), cudaMemcpyDefault, st));
  EmitDev e;
  e.n_mode = a->n_mode;
  e.type = d_ty; e.wclass = d_wc; e.char_radius = d_cr; e.log10_std = d_ls; e.vol_frac = d_vf; e.vol_frac_std = d_vs;
  e.num_conc = d_nc; e.sample_prop = d_sp;
  e.char_factor = a->characteristic_factor; e.create_time = a->create_time;
  e.allow_doubling = a->allow_doubling; e.allow_halving = a->allow_halving; e.step = a->step;
  e.stream = a->stream;
  e.n_sbin = (int)nsb; e.sample_radius = d_sr; e.sample_num_conc = d_sn;
  e.n_add = d_na;
  {
    ProfScope ps_(ctx, "k_emit");
    const int blocks = (v.n_cells + REB_WARPS - 1) / REB_WARPS;
    k_emit<<<blocks, REB_WARPS * 32, 0, st>>>(v, e);
    LAUNCH_CHECK(ctx);
  }
  if (n_part_add) CUDA_OK(ctx, cudaMemcpyAsync(n_part_add, d_na, nc * sizeof(int), cudaMemcpyDefault, st));
  const int rc = pmc_first_status(ctx);  // synchronises the stream
  cudaFree(buf);
  return rc;
}

// ---- sea salt (interface/wrf_pmc_driver.F90:212-223, :1767-1924): the per-cell sampling proportion and mode
// fluxes on the device, then the emission kernel above.  Tabulated numbers are DEFAULT-REAL literals in the Fortran
// (single precision widened to double), like the exponent of wind_speed**3.41.
__global__ void k_seasalt_prop(int nx, int nz, int ny, const float* __restrict__ u10, const float* __restrict__ v10,
                               const float* __restrict__ xland, const double* __restrict__ height, int opt, double dt,
                               int n_mode, const double* __restrict__ flux, double* __restrict__ sample_prop,
                               double* __restrict__ num_conc) {
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= nx * nz * ny) return;
  const int i = cell % nx, k = (cell / nx) % nz, j = cell / (nx * nz);
  double p = 0.0;
  if (k == 0 && xland[j * nx + i] > 1.5f) {
    const float u = u10[j * nx + i], w = v10[j * nx + i];
    const float w2 = __fadd_rn(__fmul_rn(u, u), __fmul_rn(w, w));  // no contraction: the products round first
    const double wind_speed = (double)__fsqrt_rn(w2);
    const double scale = (opt == 1) ? pow(wind_speed, (double)3.41f) : 1.0;
    p = scale * dt / height[j * nx + i];
  }
  sample_prop[cell] = p;
  for (int m = 0; m < n_mode; m++) num_conc[(size_t)cell * n_mode + m] = flux[m];
}

extern "C" int pmcgpu_seasalt_emissions(pmcgpu_ctx* ctx, const pmcgpu_seasalt_args* a, int* n_part_add) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_seasalt_emissions");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !a || !a->u10 || !a->v10 || !a->xland || !a->height || !a->weight_class || !(a->dt > 0.0)) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  const int n_mode = a->sea_salt_opt == 1 ? 2 : (a->sea_salt_opt == 2 ? 5 : 0);
  if (n_mode == 0) {
    ctx->err = "invalid sea salt option";  // wrf_pmc_driver.F90:1911
    return PMCGPU_ERR_ARG;
  }
  if (a->i_na < 1 || a->i_na > v.S || a->i_cl < 1 || a->i_cl > v.S || a->i_na == a->i_cl) return PMCGPU_ERR_ARG;
  const int nx = ctx->cfg.nx, nz = ctx->cfg.nz, ny = ctx->cfg.ny;
  const size_t n2 = (size_t)nx * ny, nc = (size_t)v.n_cells;
  // mode tables (host)
  int type[5], wclass[5];
  double cr[5], ls[5], flux[5];
  std::vector<double> vf((size_t)n_mode * v.S, 0.0);
  const double mw_na = 22.990, mw_cl = 35.450;
  for (int m = 0; m < n_mode; m++) {
    type[m] = 1;
    wclass[m] = a->weight_class[m];
    vf[(size_t)m * v.S + a->i_na - 1] = mw_na / (mw_na + mw_cl);
    vf[(size_t)m * v.S + a->i_cl - 1] = 1.0 - (mw_na / (mw_na + mw_cl));
    if (a->sea_salt_opt == 1) {
      static const float f[2] = {73.7095947f, 10.1706810f};
      static const float diam[2] = {2.03347408e-07f, 1.09840811e-06f};
      static const float sd[2] = {2.0f, 2.2f};
      cr[m] = (double)diam[m] / 2;
      ls[m] = log10((double)sd[m]);
      flux[m] = (double)f[m];
    } else {
      static const float sigma[5] = {1.37f, 1.5f, 1.42f, 1.53f, 1.85f};
      static const float CMD[5] = {.018f, .041f, .09f, .23f, .83f};
      static const float coef[5] = {104.5f, .0442f, 149.6f, 2.96f, .51f};
      static const double offset[5] = {1e5, 1e5, 1e5, 1e5, 2e5};
      static const float expon[5] = {.556f, 1.08f, .545f, .79f, .87f};
      cr[m] = .5 * (double)CMD[m] * 1e-6;
      ls[m] = log10((double)sigma[m]);
      flux[m] = (double)coef[m] * pow(3.1e6 - offset[m], (double)expon[m]);
    }
  }
  TmpDev<float> f32;
  TmpDev<double> f64;
  CUDA_OK(ctx, f32.alloc(3 * n2));
  CUDA_OK(ctx, f64.alloc(n2 + 5 + nc + nc * n_mode));
  float* d_u = f32.p;
  float* d_v = d_u + n2;
  float* d_x = d_v + n2;
  double* d_h = f64.p;
  double* d_flux = d_h + n2;
  double* d_sp = d_flux + 5;
  double* d_nc = d_sp + nc;
  cudaStream_t st = ctx->stream;
  CUDA_OK(ctx, cudaMemcpyAsync(d_u, a->u10, n2 * sizeof(float), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_v, a->v10, n2 * sizeof(float), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_x, a->xland, n2 * sizeof(float), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_h, a->height, n2 * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_flux, flux, n_mode * sizeof(double), cudaMemcpyHostToDevice, st));
  {
    ProfScope ps_(ctx, "k_seasalt_prop");
    k_seasalt_prop<<<(unsigned)((nc + 255) / 256), 256, 0, st>>>(nx, nz, ny, d_u, d_v, d_x, d_h, a->sea_salt_opt, a->dt,
                                                                 n_mode, d_flux, d_sp, d_nc);
    LAUNCH_CHECK(ctx);
  }
  CUDA_OK(ctx, cudaStreamSynchronize(st));  // `flux` is a stack array
  pmcgpu_emit_args e;
  memset(&e, 0, sizeof(e));
  e.n_mode = n_mode;
  e.mode_type = type; e.weight_class = wclass; e.char_radius = cr; e.log10_std_dev = ls;
  e.vol_frac = vf.data(); e.vol_frac_std = nullptr;
  e.num_conc = d_nc; e.sample_prop = d_sp;
  const double sample_timescale = 3600.0;
  e.characteristic_factor = fmax(1.0, fmin(sample_timescale, a->elapsed_time)) / a->dt;
  e.create_time = a->elapsed_time;
  e.allow_doubling = a->allow_doubling; e.allow_halving = a->allow_halving;
  e.step = a->step;
  e.stream = 1024u;
  return emit_run(ctx, &e, n_part_add);
}

int pmc_rebalance_async(pmcgpu_ctx* ctx, int allow_doubling, int allow_halving, unsigned step) {
  DevView& v = ctx->v;
  const int blocks = (v.n_cells + REB_WARPS - 1) / REB_WARPS;
  {
    ProfScope ps_(ctx, "k_rebalance");
    k_rebalance<<<blocks, REB_WARPS * 32, 0, ctx->stream>>>(v, allow_doubling, allow_halving, step);
    LAUNCH_CHECK(ctx);
  }
  return PMCGPU_OK;
}

extern "C" int pmcgpu_aero_state_rebalance(pmcgpu_ctx* ctx, int allow_doubling, int allow_halving,
                                           unsigned int step) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  int rc = pmc_rebalance_async(ctx, allow_doubling, allow_halving, step);
  if (rc) return rc;
  return pmc_first_status(ctx);
}

// ---------------------------------------------------------------------------------------------------
// Open domains: boundary conditions (include/pmcgpu.h, pmcgpu_set_boundary) and the refresh of the boundary
// cells at the end of trans_aero (interface/wrf_pmc_trans_aero.F90:811-955).
// ---------------------------------------------------------------------------------------------------
extern "C" int pmcgpu_set_boundary(pmcgpu_ctx* ctx, const pmcgpu_boundary_args* a) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !a || a->n_mode < 1 || !a->mode_type || !a->weight_class || !a->char_radius || !a->log10_std_dev ||
      !a->vol_frac || !a->num_conc || !a->outflow || !a->throw_away)
    return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  if (!v.kind) {
    ctx->err = "pmcgpu_set_boundary: the domain is doubly periodic";
    return PMCGPU_ERR_ARG;
  }
  const size_t nm = (size_t)a->n_mode, S = (size_t)v.S, nc = (size_t)v.n_cells;
  if (ctx->bnd_buf && ctx->bnd_n_mode != a->n_mode) {
    cudaFree(ctx->bnd_buf);
    ctx->bnd_buf = nullptr;
  }
  if (!ctx->bnd_buf) {
    const size_t bytes = (2 * nm + 2 * nm * S + nc * nm) * sizeof(double) + (2 * nm + 2 * nc) * sizeof(int) + 64;
    CUDA_OK(ctx, cudaMalloc((void**)&ctx->bnd_buf, bytes));
    ctx->bnd_n_mode = a->n_mode;
    ctx->bnd_cr = (double*)ctx->bnd_buf;
    ctx->bnd_ls = ctx->bnd_cr + nm;
    ctx->bnd_vf = ctx->bnd_ls + nm;
    ctx->bnd_vs = ctx->bnd_vf + nm * S;
    ctx->bnd_nc = ctx->bnd_vs + nm * S;
    ctx->bnd_type = (int*)(ctx->bnd_nc + nc * nm);
    ctx->bnd_wclass = ctx->bnd_type + nm;
    ctx->bnd_outflow = ctx->bnd_wclass + nm;
    ctx->bnd_throw = ctx->bnd_outflow + nc;
  }
  cudaStream_t st = ctx->stream;
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_cr, a->char_radius, nm * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_ls, a->log10_std_dev, nm * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_vf, a->vol_frac, nm * S * sizeof(double), cudaMemcpyDefault, st));
  if (a->vol_frac_std)
    CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_vs, a->vol_frac_std, nm * S * sizeof(double), cudaMemcpyDefault, st));
  else
    CUDA_OK(ctx, cudaMemsetAsync(ctx->bnd_vs, 0, nm * S * sizeof(double), st));
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_nc, a->num_conc, nc * nm * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_type, a->mode_type, nm * sizeof(int), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_wclass, a->weight_class, nm * sizeof(int), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_outflow, a->outflow, nc * sizeof(int), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->bnd_throw, a->throw_away, nc * sizeof(int), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaStreamSynchronize(st));
  return PMCGPU_OK;
}

struct BndDev {
  EmitDev e;             // background modes (num_conc / sample_prop per cell are taken from here and from 1 / rrho)
  const int* outflow;    // [n_cells]
  const int* throw_away; // [n_cells]
  int nx, nz;
};

// record `src` of cell `from_base` -> slot `dst` of cell `to_base` (warp-cooperative, the id is set by the caller)
__device__ __forceinline__ void bnd_copy_record(DevView& v, size_t NS, size_t to, size_t from, int lane) {
  for (int s = lane; s < v.S; s += 32) v.vol[(size_t)s * NS + to] = v.vol[(size_t)s * NS + from];
  if (lane == 0) v.wgc[to] = v.wgc[from];
  if (lane == 1) v.nprim[to] = v.nprim[from];
  if (lane == 2) v.hyst[to] = v.hyst[from];
  if (lane == 3) v.lct[to] = v.lct[from];
  if (lane == 4) v.gct[to] = v.gct[from];
  if (lane == 5) v.pbin[to] = v.pbin[from];
  if (v.cold.n) cold_copy_warp(v.cold, NS, to, v.cold, NS, from, v.n_opq, lane);
}

// One warp (= one CTA) per boundary-edge cell.  Shared memory: LST[cap] class lists (positions), FXC[cap]
// position -> (class << 24 | index in its list) or -1 for particles appended by this call, len[C], off[C].
__global__ void __launch_bounds__(32) k_bnd_refresh(DevView v, BndDev b, unsigned step) {
  const int cell = blockIdx.x, lane = threadIdx.x;
  const int kind = v.kind[cell];
  if (kind < 1 || kind > 4) return;
  if (v.status[cell] != 0) return;
  extern __shared__ int s_b[];
  int* LST = s_b;
  int* FXC = LST + v.cap;
  int* len = FXC + v.cap;
  int* off = len + v.C;
  const size_t NS = (size_t)v.n_cells * v.cap;
  const size_t base = (size_t)cell * v.cap;
  const int nbr = cell + (kind == 1 ? 1 : (kind == 2 ? -1 : (kind == 3 ? b.nx * b.nz : -b.nx * b.nz)));
  const unsigned gcell = (unsigned)v.gcell[cell];
  int err = 0;
  if (b.outflow[cell]) {
    // ---- aero_state_dup_all_particles (aero_state.F90:3529-3551): the state of the interior neighbour, new ids
    const int n = v.count[nbr];
    if (n > 0) {
      const size_t nbase = (size_t)nbr * v.cap;
      const long long nxid = v.next_id[cell];
      for (int p = 0; p < n; p++) bnd_copy_record(v, NS, base + p, nbase + p, lane);
      for (int p = lane; p < n; p += 32) v.id[base + p] = (nxid + p - 1) * v.n_cells_global + v.gcell[cell] + 1;
      const int ncg = v.C * v.G;
      for (int q = lane; q < ncg; q += 32) {
        v.wtype[(size_t)cell * ncg + q] = v.wtype[(size_t)nbr * ncg + q];
        v.wmag[(size_t)cell * ncg + q] = v.wmag[(size_t)nbr * ncg + q];
        v.wexp[(size_t)cell * ncg + q] = v.wexp[(size_t)nbr * ncg + q];
      }
      for (int q = lane; q < v.C * v.nbcap; q += 32)
        v.bin_count[(size_t)cell * v.C * v.nbcap + q] = v.bin_count[(size_t)nbr * v.C * v.nbcap + q];
      if (lane == 0) {
        v.next_id[cell] = nxid + n;
        v.grid_nb[cell] = v.grid_nb[nbr];
        v.grid_min[cell] = v.grid_min[nbr];
        v.grid_max[cell] = v.grid_max[nbr];
        v.bounds_valid[cell] = 0;
      }
    }
    if (lane == 0) v.count[cell] = n;  // an empty neighbour: aero_state_zero
    return;
  }
  const double density = 1.0 / v.rrho[cell];
  const double* mode_nc = b.e.num_conc + (size_t)cell * b.e.n_mode;
  if (b.throw_away[cell]) {
    // ---- aero_state_resample, throw_away_previous (:1028-1032): aero_state_zero + aero_state_add_aero_dist_sample
    // (density, sample_timescale = 1, create_time, allow_doubling, allow_halving)
    if (lane == 0) v.count[cell] = 0;
    __syncwarp();
    emit_cell(v, b.e, cell, lane, density, mode_nc, err);
    if (lane == 0 && err && v.status[cell] == 0) v.status[cell] = err;
    return;
  }
  // ---- aero_state_resample (:1033-1095) from fresh class lists (ascending positions, a fresh aero_state_sort)
  int n_cur = v.count[cell];
  {
    int run = 0;
    for (int c = 0; c < v.C; c++) {
      if (lane == 0) off[c] = run;
      int m = 0;
      for (int p0 = 0; p0 < n_cur; p0 += 32) {
        const int p = p0 + lane;
        const bool is = (p < n_cur) && (v.wgc[base + p] == c * v.G);
        const unsigned bm = __ballot_sync(0xffffffffu, is);
        if (is) {
          const int idx = m + __popc(bm & ((1u << lane) - 1u));
          LST[run + idx] = p;
          FXC[p] = (c << 24) | idx;
        }
        m += __popc(bm);
      }
      if (lane == 0) len[c] = m;
      run += m;
    }
    if (run != n_cur) err = PMCGPU_ERR_ARG;  // one weight group in WRF
  }
  __syncwarp();
  const double sample_timescale = 1.0, factor = 0.1;
  for (int c = 0; c < v.C && !err; c++) {
    const unsigned gc = (unsigned)(c * v.G);  // group 1
    const size_t wi = (size_t)cell * v.C * v.G + gc;
    const int wt = v.wtype[wi];
    const double wm = v.wmag[wi], we = v.wexp[wi];
    const int n_part_before = len[c];
    double total_num_conc = 0.0;
    for (int im = 0; im < b.e.n_mode; im++)
      if (b.e.wclass[im] == c + 1) total_num_conc = total_num_conc + mode_nc[im] * density;
    const double n_parts_new = total_num_conc / wm;
    RStream s0{RKey{v.seed, gcell, step, PH_BND, gc, 0u}, 0u};
    const int n_samp = rand_poisson(n_parts_new, s0);
    const int n_parts_to_add = n_samp - n_part_before;
    if (n_parts_to_add > 0) {
      for (int im = 0; im < b.e.n_mode && !err; im++) {
        if (b.e.wclass[im] != c + 1) continue;
        const unsigned am = 1000u + gc * (unsigned)b.e.n_mode + (unsigned)im;
        const int type = b.e.type[im];
        const double cr = b.e.char_radius[im], lsd = b.e.log10_std[im];
        const double sample_proportion = (density * mode_nc[im]) / total_num_conc;
        RStream sm{RKey{v.seed, gcell, step, PH_BND, am, 0u}, 0u};
        const int n_samp_mode = rand_poisson(sample_proportion * (double)n_parts_to_add, sm);
        const double n_samp_avg = sample_proportion * (double)n_parts_to_add;
        const double num_conc_threshold = (density * mode_nc[im] / n_samp_avg) * factor;
        const double size_factor = fmin(1.0 / (sample_timescale * n_samp_avg), 0.32);
        const double* vf = b.e.vol_frac + (size_t)im * v.S;
        const double* vs = b.e.vol_frac_std + (size_t)im * v.S;
        const long long nxid = v.next_id[cell];
        for (int i0 = 0; i0 < n_samp_mode && !err; i0 += 32) {
          const int i_samp = i0 + lane + 1;
          const bool active = i_samp <= n_samp_mode;
          double vols[PMC_MAX_SPEC];
          bool keep = false;
          if (active) {
            RStream s{RKey{v.seed, gcell, step, PH_BND, am, (unsigned)i_samp}, 0u};
            const double radius = sample_mode_particle(v.S, type, cr, lsd, wt, we, size_factor, vf, vs, s, vols);
            keep = weight_nc(wt, wm, we, radius) > num_conc_threshold;  // :1069-1073
          }
          const unsigned km = __ballot_sync(0xffffffffu, keep);
          const int slot = n_cur + __popc(km & ((1u << lane) - 1u));
          if (n_cur + __popc(km) > v.cap) { err = PMCGPU_ERR_CAPACITY; break; }
          if (keep) {
            for (int k = 0; k < v.S; k++) v.vol[(size_t)k * NS + base + slot] = vols[k];
            v.id[base + slot] = (nxid + (i_samp - 1) - 1) * v.n_cells_global + v.gcell[cell] + 1;  // an id per SAMPLE (:1067)
            v.wgc[base + slot] = (int)gc;
            v.nprim[base + slot] = 1;
            v.hyst[base + slot] = 0;
            v.lct[base + slot] = b.e.create_time;
            v.gct[base + slot] = b.e.create_time;
            if (v.cold.n) cold_new(v.cold, NS, base + slot, (int)gc / v.G + 1, b.e.create_time, v.n_opq);
            FXC[slot] = -1;  // not in any list of this call
          }
          n_cur += __popc(km);
        }
        __syncwarp();
        if (lane == 0) v.next_id[cell] = nxid + n_samp_mode;
        __syncwarp();
      }
    } else {
      // removal of min(-n_parts_to_add, n_part) random entries of the (group, class) list (:1080-1091):
      // aero_particle_array swap-with-last + integer_rmap2_remove (integer_rmap2.F90:190-243)
      const int n_rm = min(-n_parts_to_add, len[c]);
      for (int r = 1; r <= n_rm; r++) {
        const int ne = len[c];
        const int ie = pmc_rand_int(rng_uniform(RKey{v.seed, gcell, step, PH_BND, gc, (unsigned)r}, 0), ne);
        const int p = LST[off[c] + ie - 1];
        const int last = n_cur - 1;
        __syncwarp();
        if (lane == 0) {
          const int pos_old = FXC[p] & 0xffffff;
          if (last != p) {
            const int ff = FXC[last];
            if (ff >= 0) LST[off[ff >> 24] + (ff & 0xffffff)] = p;
            FXC[p] = ff;
          }
          const int lp = off[c];
          const int pos_last = len[c] - 1;
          const int dom_fix = LST[lp + pos_last];
          if (pos_old < pos_last) LST[lp + pos_old] = dom_fix;
          len[c] = pos_last;
          if (pos_last != pos_old) FXC[dom_fix] = (c << 24) | pos_old;
        }
        __syncwarp();
        if (last != p) {
          bnd_copy_record(v, NS, base + p, base + last, lane);
          if (lane == 6) v.id[base + p] = v.id[base + last];
        }
        n_cur = last;
        __syncwarp();
      }
    }
  }
  if (lane == 0) {
    v.count[cell] = n_cur;
    if (err && v.status[cell] == 0) v.status[cell] = err;
  }
}

int pmc_boundary_refresh_async(pmcgpu_ctx* ctx, unsigned step) {
  DevView& v = ctx->v;
  if (!v.kind || !ctx->bnd_buf) return PMCGPU_ERR_ARG;
  BndDev b;
  b.e.n_mode = ctx->bnd_n_mode;
  b.e.type = ctx->bnd_type; b.e.wclass = ctx->bnd_wclass; b.e.char_radius = ctx->bnd_cr; b.e.log10_std = ctx->bnd_ls;
  b.e.vol_frac = ctx->bnd_vf; b.e.vol_frac_std = ctx->bnd_vs; b.e.num_conc = ctx->bnd_nc; b.e.sample_prop = nullptr;
  b.e.char_factor = 1.0; b.e.create_time = 0.0;   // sample_timescale, create_time of the call sites (:839-842)
  b.e.allow_doubling = 1; b.e.allow_halving = 1; b.e.step = step; b.e.stream = 0u; b.e.n_add = nullptr;
  b.e.n_sbin = 0; b.e.sample_radius = nullptr; b.e.sample_num_conc = nullptr;
  b.outflow = ctx->bnd_outflow; b.throw_away = ctx->bnd_throw;
  b.nx = ctx->cfg.nx; b.nz = ctx->cfg.nz;
  const size_t smem = ((size_t)2 * v.cap + 2 * v.C) * sizeof(int);
  if (smem > 200 * 1024) {
    ctx->err = "cell_capacity too large for the class lists of the boundary refresh";
    return PMCGPU_ERR_ARG;
  }
  if (smem > 48 * 1024 && smem > ctx->cfg_bnd_smem) {
    CUDA_OK(ctx, cudaFuncSetAttribute(k_bnd_refresh, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctx->cfg_bnd_smem = smem;
  }
  {
    ProfScope ps_(ctx, "k_bnd_refresh");
    k_bnd_refresh<<<v.n_cells, 32, smem, ctx->stream>>>(v, b, step);
    LAUNCH_CHECK(ctx);
  }
  // the bin-grid decision of the refreshed cells (a fresh aero_state_sort of every boundary cell)
  int pmc_sort_grid_only_async(pmcgpu_ctx*, int);
  return pmc_sort_grid_only_async(ctx, 2);
}
