// sort.cu -- aero_state_sort for every cell of the patch in one launch.
//
// Replaces, per cell: aero_sorted_remake_if_needed (partmc/src/aero_sorted.F90:
// 214-373: r_min/r_max scan, safety-factor test, adaptive log grid),
// aero_sorted_sort_particles (:185-209) and the integer_rmap2 multimap
// (integer_rmap2.F90) with a stable segmented counting sort keyed by
// (weight class, size bin).  The particle records themselves are permuted into
// (class, bin, original index) order, so every size_class%inverse(bin, class)
// list of the reference is a contiguous run whose order (ascending original
// particle index) equals the reference's append order.
//
// One CTA per grid cell.  HBM traffic per particle: read vol[S] (binning) +
// read record + write record (permutation) = 8S + 2P bytes, P = 8S + 44.
#include "pmcgpu_internal.cuh"

#define SORT_THREADS 256
#define SORT_WARPS (SORT_THREADS / 32)

static const double BINS_PER_DECADE = 10.0;   // aero_sorted.F90:69
static const double BIN_OVER_FACTOR = 10.0;   // :71
static const double BIN_SAFETY_FACTOR = 3.0;  // :73

__global__ void __launch_bounds__(SORT_THREADS) k_sort_cells(DevView v, int edge_recheck) {
  const int cell = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = v.count[cell];
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  extern __shared__ int s_hist[];  // [SORT_WARPS][nk] then [nk] totals
  __shared__ double s_min[SORT_WARPS], s_max[SORT_WARPS];
  __shared__ double s_lnmin, s_lnmax;
  __shared__ int s_nb, s_fail, s_redo;

  // ---- phase 1: radius of every particle, r_min / r_max (aero_sorted.F90:277-290)
  double rmin = 1e300, rmax = 0.0;
  for (int p = tid; p < n; p += SORT_THREADS) {
    double vt = 0.0;
    for (int s = 0; s < v.S; s++) vt += v.vol[(size_t)s * NS + base + p];  // aero_particle_volume
    const double r = vol2rad(vt);
    v.lnr[base + p] = log(r);
    rmin = fmin(rmin, r);
    rmax = fmax(rmax, r);
  }
  for (int o = 16; o > 0; o >>= 1) {
    rmin = fmin(rmin, __shfl_xor_sync(0xffffffffu, rmin, o));
    rmax = fmax(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
  }
  if (lane == 0) {
    s_min[warp] = rmin;
    s_max[warp] = rmax;
  }
  __syncthreads();
  // ---- grid decision (thread 0), aero_sorted.F90:313-372 with valid_sort = .false.
  if (tid == 0) {
    double r_min = s_min[0], r_max = s_max[0];
    for (int w = 1; w < SORT_WARPS; w++) {
      r_min = fmin(r_min, s_min[w]);
      r_max = fmax(r_max, s_max[w]);
    }
    if (n == 0) {
      r_min = 0.0;
      r_max = 0.0;
    }
    int nb = v.grid_nb[cell];
    double gmin = v.grid_min[cell], gmax = v.grid_max[cell];
    s_fail = 0;
    if (r_max == 0.0) {
      // no particles: bin_grid_make(n_bin = 0) + aero_sorted_set_bin_grid (:322-327)
      nb = 0;
      gmin = 0.0;
      gmax = 0.0;
      v.bounds_valid[cell] = 0;
    } else {
      bool need_new = false;
      if (nb < 1) {
        need_new = true;
      } else if ((r_min / gmin < BIN_SAFETY_FACTOR) || (gmax / r_max < BIN_SAFETY_FACTOR)) {
        need_new = true;
      }
      if (need_new) {
        gmin = r_min / BIN_OVER_FACTOR;
        gmax = r_max * BIN_OVER_FACTOR;
        nb = (int)ceil((log10(gmax) - log10(gmin)) * BINS_PER_DECADE);
        v.bounds_valid[cell] = 0;  // aero_sorted_set_bin_grid, :132
        if (nb > v.nbcap) {
          s_fail = 1;
          if (v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_NBINS;
          nb = 0;
        }
      }
    }
    v.grid_nb[cell] = nb;
    v.grid_min[cell] = gmin;
    v.grid_max[cell] = gmax;
    s_nb = nb;
    s_lnmin = (nb > 0) ? log(gmin) : 0.0;
    s_lnmax = (nb > 0) ? log(gmax) : 0.0;
  }
  __syncthreads();
  // zero the published counts, (class, bin) -> bin_count[cell][c][b]
  int* bc = v.bin_count + (size_t)cell * v.C * v.nbcap;
  for (int q = tid; q < v.C * v.nbcap; q += SORT_THREADS) bc[q] = 0;
  if (s_nb == 0 || s_fail) return;
  if (edge_recheck == 2) return;  // grid decision only (end of trans_aero): the records stay where they are
  // ---- phase 2: bin of every particle (bin_grid.F90:143) + per-warp histograms.
  // Warp w owns the contiguous index range [w*seg, (w+1)*seg) so that ranks
  // inside a key follow the original particle order (stable).
  // With edge_recheck (the aero_state_sort call at the top of mc_coag, which sees
  // valid_sort = .true.) the safety-factor test is repeated on the edges of the
  // lowest / highest occupied bins (aero_sorted.F90:258-276); if that asks for a
  // new grid the bins are recomputed once.
  const int seg = (n + SORT_WARPS - 1) / SORT_WARPS;
  const int lo = warp * seg, hi = min(n, lo + seg);
  int nb = s_nb;
  int nk = v.C * nb;
  int* myhist = s_hist + warp * nk;
  int* tot = s_hist + SORT_WARPS * nk;
  for (int attempt = 0; attempt < 2; attempt++) {
    nb = s_nb;
    nk = v.C * nb;
    myhist = s_hist + warp * nk;
    tot = s_hist + SORT_WARPS * nk;
    const double lnmin = s_lnmin, lnmax = s_lnmax;
    for (int q = tid; q < (SORT_WARPS + 1) * nk; q += SORT_THREADS) s_hist[q] = 0;
    __syncthreads();
    for (int p0 = lo; p0 < hi; p0 += 32) {
      const int p = p0 + lane;
      int key = -1;
      if (p < hi) {
        int b = log_grid_find(lnmin, lnmax, nb, v.lnr[base + p]);
        b = min(max(b, 1), nb);  // unreachable clamp: the grid was just checked against r_min / r_max
        v.pbin[base + p] = b;
        key = (v.wgc[base + p] / v.G) * nb + (b - 1);
      }
      const unsigned m = __match_any_sync(0xffffffffu, key);
      if (key >= 0 && (m & ((1u << lane) - 1u)) == 0) myhist[key] += __popc(m);
      __syncwarp();
    }
    __syncthreads();
    for (int k = tid; k < nk; k += SORT_THREADS) {
      int t = 0;
      for (int w = 0; w < SORT_WARPS; w++) t += s_hist[w * nk + k];
      tot[k] = t;
    }
    __syncthreads();
    if (tid == 0) {
      s_redo = 0;
      if (attempt == 0 && edge_recheck) {
        int i_min = 0, i_max = 0;
        for (int b = 1; b <= nb; b++) {
          int t = 0;
          for (int c = 0; c < v.C; c++) t += tot[c * nb + (b - 1)];
          if (t > 0) {
            if (i_min == 0) i_min = b;
            i_max = b;
          }
        }
        const double gmin = v.grid_min[cell], gmax = v.grid_max[cell];
        const double r_min = grid_edge(gmin, gmax, nb, i_min);
        const double r_max = grid_edge(gmin, gmax, nb, i_max + 1);
        if ((r_min / gmin < BIN_SAFETY_FACTOR) || (gmax / r_max < BIN_SAFETY_FACTOR)) {
          const double ngmin = r_min / BIN_OVER_FACTOR;
          const double ngmax = r_max * BIN_OVER_FACTOR;
          int nnb = (int)ceil((log10(ngmax) - log10(ngmin)) * BINS_PER_DECADE);
          v.bounds_valid[cell] = 0;
          if (nnb > v.nbcap) {
            s_fail = 1;
            if (v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_NBINS;
            nnb = 0;
          }
          v.grid_nb[cell] = nnb;
          v.grid_min[cell] = ngmin;
          v.grid_max[cell] = ngmax;
          s_nb = nnb;
          s_lnmin = (nnb > 0) ? log(ngmin) : 0.0;
          s_lnmax = (nnb > 0) ? log(ngmax) : 0.0;
          s_redo = 1;
        }
      }
    }
    __syncthreads();
    if (s_fail) return;
    if (!s_redo) break;
  }
  // ---- exclusive scan: key-major, warp-minor
  for (int k = tid; k < nk; k += SORT_THREADS) bc[(k / nb) * v.nbcap + (k % nb)] = tot[k];
  __syncthreads();
  if (warp == 0) {
    int carry = 0;
    for (int k0 = 0; k0 < nk; k0 += 32) {
      const int k = k0 + lane;
      const int x = (k < nk) ? tot[k] : 0;
      int incl = x;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (k < nk) tot[k] = carry + incl - x;
      carry += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
  __syncthreads();
  for (int k = tid; k < nk; k += SORT_THREADS) {
    int run = tot[k];
    for (int w = 0; w < SORT_WARPS; w++) {
      const int c = s_hist[w * nk + k];
      s_hist[w * nk + k] = run;
      run += c;
    }
  }
  __syncthreads();
  // ---- phase 3: permute the records into the alternate slab
  for (int p0 = lo; p0 < hi; p0 += 32) {
    const int p = p0 + lane;
    int key = -1, b = 0, gc = 0;
    if (p < hi) {
      b = v.pbin[base + p];
      gc = v.wgc[base + p];
      key = (gc / v.G) * nb + (b - 1);
    }
    const unsigned m = __match_any_sync(0xffffffffu, key);
    int dst = -1;
    if (key >= 0) {
      const int rank = __popc(m & ((1u << lane) - 1u));
      dst = myhist[key] + rank;
    }
    __syncwarp();
    if (key >= 0 && (m & ((1u << lane) - 1u)) == 0) myhist[key] += __popc(m);
    __syncwarp();
    if (dst >= 0) {
      for (int s = 0; s < v.S; s++) v.vol_alt[(size_t)s * NS + base + dst] = v.vol[(size_t)s * NS + base + p];
      v.id_alt[base + dst] = v.id[base + p];
      v.wgc_alt[base + dst] = gc;
      v.nprim_alt[base + dst] = v.nprim[base + p];
      v.hyst_alt[base + dst] = v.hyst[base + p];
      v.lct_alt[base + dst] = v.lct[base + p];
      v.gct_alt[base + dst] = v.gct[base + p];
      v.pbin_alt[base + dst] = b;
    }
  }
}

static void swap_slabs(DevView& v);
void pmc_swap_slabs(pmcgpu_ctx* ctx) { swap_slabs(ctx->v); }
// aero_state_sort at the end of trans_aero (wrf_pmc_trans_aero.F90:803): the bin-grid decision of
// aero_sorted_remake_if_needed only; the next mc_coag permutes the records anyway
int pmc_sort_grid_only_async(pmcgpu_ctx* ctx) {
  DevView& v = ctx->v;
  const size_t smem = (size_t)(SORT_WARPS + 1) * v.C * v.nbcap * sizeof(int);
  ProfScope ps_(ctx, "k_sort_cells_grid");
  k_sort_cells<<<v.n_cells, SORT_THREADS, smem, ctx->stream>>>(v, 2);
  LAUNCH_CHECK(ctx);
  return PMCGPU_OK;
}
static void swap_slabs(DevView& v) {
  std::swap(v.vol, v.vol_alt);
  std::swap(v.id, v.id_alt);
  std::swap(v.wgc, v.wgc_alt);
  std::swap(v.nprim, v.nprim_alt);
  std::swap(v.hyst, v.hyst_alt);
  std::swap(v.lct, v.lct_alt);
  std::swap(v.gct, v.gct_alt);
  std::swap(v.pbin, v.pbin_alt);
}

int pmc_sort_async(pmcgpu_ctx* ctx, int edge_recheck) {
  DevView& v = ctx->v;
  const size_t smem = (size_t)(SORT_WARPS + 1) * v.C * v.nbcap * sizeof(int);
  if (smem > 200 * 1024) {
    ctx->err = "sort: n_class * max_bins too large for the shared-memory histogram";
    return PMCGPU_ERR_ARG;
  }
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    CUDA_OK(ctx, cudaFuncSetAttribute(k_sort_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  {
    ProfScope ps_(ctx, "k_sort_cells");
    k_sort_cells<<<v.n_cells, SORT_THREADS, smem, ctx->stream>>>(v, edge_recheck);
    LAUNCH_CHECK(ctx);
  }
  // cells with an empty grid keep their (empty) slab: nothing to copy
  swap_slabs(v);
  return PMCGPU_OK;
}

extern "C" int pmcgpu_aero_state_sort(pmcgpu_ctx* ctx) {
  if (!ctx) return PMCGPU_ERR_ARG;
  int rc = pmc_sort_async(ctx, 0);
  if (rc) return rc;
  return pmc_first_status(ctx);
}
