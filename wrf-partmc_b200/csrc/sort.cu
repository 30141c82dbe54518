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

#define SORT_THREADS 256          // legacy (unstaged) variant
#define SORT_THREADS_STAGED 512   // staged variant: one CTA per SM, so more warps per CTA

static const double BINS_PER_DECADE = 10.0;   // aero_sorted.F90:69
static const double BIN_OVER_FACTOR = 10.0;   // :71
static const double BIN_SAFETY_FACTOR = 3.0;  // :73

// mbarrier + 1-D bulk-copy (TMA) primitives of the staged variant
__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
  } while (!done);
}

// STAGED: the cell's S species rows are pulled into shared memory once by bulk copies (TMA); radii, bins and
// the permutation are computed from there and the records leave in destination order with coalesced stores,
// so HBM sees each record once in and once out.  Handles the cells with n <= n_split; the legacy variant
// (rows read twice, scattered stores) takes the cells with n > n_split (n_split = -1: every cell).
// s_st: how many of the S species rows the staged variant keeps in shared memory (the first s_st); the others are
// streamed from HBM for the radius and gathered from the cell's own, just-read rows when the records leave.
// geometric radius of fractal particles (pmcgpu_set_fractal, fractal.F90:106-121); out of line, so that the spherical
// path keeps its code, and with scalar arguments: a reference to the by-value kernel parameter would put a copy of
// the whole DevView on every thread's stack
static __device__ __noinline__ double vol2rad_fractal(double vol, double r0, double fill, double df) {
  const double x = pow(3.0 * vol / 4.0 / PmcConst::pi, 1.0 / 3.0) / r0;
  return r0 * pow((x * x * x) * fill, 1.0 / df);
}

template <bool STAGED, int NT>
__global__ void __launch_bounds__(NT) k_sort_cells(DevView v, int edge_recheck, int n_split, int s_st, int n_lo) {
  constexpr int SORT_WARPS = NT / 32;
  constexpr int SORT_THREADS_ = NT;
  const int cell = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = v.count[cell];
  if (STAGED ? (n > n_split || n <= n_lo) : (n <= n_split)) return;  // staged: cells with n_lo < n <= n_split
  if (edge_recheck == 3) {  // grid decision only, interior cells of an open domain only (wrf_pmc_trans_aero.F90:777-804)
    if (cell_is_boundary(v, cell)) return;
    edge_recheck = 2;
  } else if (edge_recheck == 4) {  // grid decision only, boundary-edge cells only (after their refresh)
    if (!v.kind || v.kind[cell] < 1 || v.kind[cell] > 4) return;
    edge_recheck = 2;
  }
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  extern __shared__ __align__(16) int s_hist[];  // [SORT_WARPS][nk] then [nk] totals (then the staged rows)
  __shared__ double s_min[SORT_WARPS], s_max[SORT_WARPS];
  __shared__ double s_lnmin, s_lnmax;
  __shared__ int s_nb, s_fail, s_redo;
  __shared__ __align__(8) unsigned long long s_bar;
  // staged layout after the histograms: vol rows [S][n2], ln r [n2] (later the inverse permutation), key [n2]
  const int n2 = (n + 1) & ~1;  // rows are copied in 16-byte units
  double* svol = (double*)(s_hist + (((SORT_WARPS + 1) * v.C * v.nbcap + 3) & ~3));
  double* s_lnr = svol + (size_t)s_st * n2;
  int* s_src = (int*)s_lnr;
  int* s_key = (int*)(s_lnr + n2);
  if (STAGED) {
    if (tid == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    if (n > 0 && s_st > 0) {
      if (tid == 0) {
        mbar_expect_tx(&s_bar, (unsigned)((size_t)s_st * n2 * sizeof(double)));
        for (int s = 0; s < s_st; s++)
          bulk_g2s(svol + (size_t)s * n2, v.vol + (size_t)s * NS + base, (unsigned)(n2 * sizeof(double)), &s_bar);
      }
      mbar_wait(&s_bar, 0);
    }
  }

  // ---- phase 1: radius of every particle, r_min / r_max (aero_sorted.F90:277-290)
  double rmin = 1e300, rmax = 0.0;
  for (int p = tid; p < n; p += SORT_THREADS_) {
    double vt = 0.0;
    if (STAGED) {
      // rows below s_st from shared memory, the others straight from HBM (same order of the additions)
      for (int s = 0; s < s_st; s++) vt += svol[(size_t)s * n2 + p];
      for (int s = s_st; s < v.S; s++) vt += v.vol[(size_t)s * NS + base + p];
    } else {
      for (int s = 0; s < v.S; s++) vt += v.vol[(size_t)s * NS + base + p];  // aero_particle_volume
    }
    const double r = v.fr_on ? vol2rad_fractal(vt, v.fr_r0, v.fr_fill, v.fr_df) : vol2rad(vt);  // aero_particle_radius (aero_particle.F90:388-399)
    if (STAGED)
      s_lnr[p] = log(r);
    else
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
  // ---- grid decision, aero_sorted.F90:313-372 with valid_sort = .false.  Warp 0, every lane on the same
  // values; the two logarithms of a pair are taken by lanes 0 and 1 side by side (half the serial latency).
  if (warp == 0) {
    double r_min = (lane < SORT_WARPS) ? s_min[lane] : 1e300;
    double r_max = (lane < SORT_WARPS) ? s_max[lane] : 0.0;
    for (int o = 16; o > 0; o >>= 1) {
      r_min = fmin(r_min, __shfl_xor_sync(0xffffffffu, r_min, o));
      r_max = fmax(r_max, __shfl_xor_sync(0xffffffffu, r_max, o));
    }
    if (n == 0) {
      r_min = 0.0;
      r_max = 0.0;
    }
    int nb = v.grid_nb[cell];
    double gmin = v.grid_min[cell], gmax = v.grid_max[cell];
    int fail = 0;
    bool new_grid = false;
    if (r_max == 0.0) {
      // no particles: bin_grid_make(n_bin = 0) + aero_sorted_set_bin_grid (:322-327)
      nb = 0;
      gmin = 0.0;
      gmax = 0.0;
      new_grid = true;
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
        const double l10 = log10((lane & 1) ? gmin : gmax);
        const double l10max = __shfl_sync(0xffffffffu, l10, 0), l10min = __shfl_sync(0xffffffffu, l10, 1);
        nb = (int)ceil((l10max - l10min) * BINS_PER_DECADE);
        new_grid = true;  // aero_sorted_set_bin_grid, :132
        if (nb > v.nbcap) {
          fail = 1;
          nb = 0;
        }
      }
    }
    const double ln = (nb > 0) ? log((lane & 1) ? gmax : gmin) : 0.0;
    if (lane == 0) {
      if (new_grid) v.bounds_valid[cell] = 0;
      if (fail && v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_NBINS;
      v.grid_nb[cell] = nb;
      v.grid_min[cell] = gmin;
      v.grid_max[cell] = gmax;
      s_nb = nb;
      s_fail = fail;
      s_lnmin = ln;
    }
    if (lane == 1) s_lnmax = ln;
  }
  __syncthreads();
  // zero the published counts, (class, bin) -> bin_count[cell][c][b]
  int* bc = v.bin_count + (size_t)cell * v.C * v.nbcap;
  for (int q = tid; q < v.C * v.nbcap; q += SORT_THREADS_) bc[q] = 0;
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
    for (int q = tid; q < (SORT_WARPS + 1) * nk; q += SORT_THREADS_) s_hist[q] = 0;
    __syncthreads();
    for (int p0 = lo; p0 < hi; p0 += 32) {
      const int p = p0 + lane;
      int key = -1;
      if (p < hi) {
        int b = log_grid_find(lnmin, lnmax, nb, STAGED ? s_lnr[p] : v.lnr[base + p]);
        b = min(max(b, 1), nb);  // unreachable clamp: the grid was just checked against r_min / r_max
        key = (v.wgc[base + p] / v.G) * nb + (b - 1);
        if (STAGED)
          s_key[p] = key;
        else
          v.pbin[base + p] = b;
      }
      const unsigned m = __match_any_sync(0xffffffffu, key);
      if (key >= 0 && (m & ((1u << lane) - 1u)) == 0) myhist[key] += __popc(m);
      __syncwarp();
    }
    __syncthreads();
    for (int k = tid; k < nk; k += SORT_THREADS_) {
      int t = 0;
      for (int w = 0; w < SORT_WARPS; w++) t += s_hist[w * nk + k];
      tot[k] = t;
    }
    __syncthreads();
    if (warp == 0) {
      // edges of the lowest / highest occupied bins: occupancy by ballot over the bins, the two edges by
      // lanes 0 and 1 side by side
      int redo = 0;
      if (attempt == 0 && edge_recheck) {
        int i_min = 0, i_max = 0;
        for (int b0 = 0; b0 < nb; b0 += 32) {
          const int b = b0 + lane;
          int t = 0;
          if (b < nb)
            for (int c = 0; c < v.C; c++) t += tot[c * nb + b];
          const unsigned occ = __ballot_sync(0xffffffffu, t > 0);
          if (occ) {
            if (i_min == 0) i_min = b0 + __ffs(occ);
            i_max = b0 + 32 - __clz(occ);
          }
        }
        const double gmin = v.grid_min[cell], gmax = v.grid_max[cell];
        const double e = grid_edge(gmin, gmax, nb, (lane & 1) ? i_max + 1 : i_min);
        const double r_min = __shfl_sync(0xffffffffu, e, 0), r_max = __shfl_sync(0xffffffffu, e, 1);
        if ((r_min / gmin < BIN_SAFETY_FACTOR) || (gmax / r_max < BIN_SAFETY_FACTOR)) {
          const double ngmin = r_min / BIN_OVER_FACTOR;
          const double ngmax = r_max * BIN_OVER_FACTOR;
          const double l10 = log10((lane & 1) ? ngmin : ngmax);
          const double l10max = __shfl_sync(0xffffffffu, l10, 0), l10min = __shfl_sync(0xffffffffu, l10, 1);
          int nnb = (int)ceil((l10max - l10min) * BINS_PER_DECADE);
          int fail = 0;
          if (nnb > v.nbcap) {
            fail = 1;
            nnb = 0;
          }
          const double ln = (nnb > 0) ? log((lane & 1) ? ngmax : ngmin) : 0.0;
          if (lane == 0) {
            v.bounds_valid[cell] = 0;
            if (fail) {
              s_fail = 1;
              if (v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_NBINS;
            }
            v.grid_nb[cell] = nnb;
            v.grid_min[cell] = ngmin;
            v.grid_max[cell] = ngmax;
            s_nb = nnb;
            s_lnmin = ln;
          }
          if (lane == 1) s_lnmax = ln;
          redo = 1;
        }
      }
      if (lane == 0) s_redo = redo;
    }
    __syncthreads();
    if (s_fail) return;
    if (!s_redo) break;
  }
  // ---- exclusive scan: key-major, warp-minor
  for (int k = tid; k < nk; k += SORT_THREADS_) bc[(k / nb) * v.nbcap + (k % nb)] = tot[k];
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
  for (int k = tid; k < nk; k += SORT_THREADS_) {
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
      if (STAGED) {
        key = s_key[p];
      } else {
        b = v.pbin[base + p];
        gc = v.wgc[base + p];
        key = (gc / v.G) * nb + (b - 1);
      }
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
    if (STAGED) {
      if (dst >= 0) s_src[dst] = p;  // ln r is dead: its space holds the inverse permutation
    } else if (dst >= 0) {
      for (int s = 0; s < v.S; s++) v.vol_alt[(size_t)s * NS + base + dst] = v.vol[(size_t)s * NS + base + p];
      v.id_alt[base + dst] = v.id[base + p];
      v.wgc_alt[base + dst] = gc;
      v.nprim_alt[base + dst] = v.nprim[base + p];
      v.hyst_alt[base + dst] = v.hyst[base + p];
      v.lct_alt[base + dst] = v.lct[base + p];
      v.gct_alt[base + dst] = v.gct[base + p];
      v.pbin_alt[base + dst] = b;
      if (v.cold.n) cold_copy(v.cold_alt, NS, base + dst, v.cold, NS, base + p, v.n_opq);
    }
  }
  if (STAGED) {
    // records leave in destination order: shared-memory gather, coalesced stores
    __syncthreads();
    for (int q = tid; q < n; q += SORT_THREADS_) {
      const int src = s_src[q];
      // the gathered fields first (one round trip for all six), the staged rows while they are in flight
      const long long r_id = v.id[base + src];
      const int r_wgc = v.wgc[base + src], r_nprim = v.nprim[base + src], r_hyst = v.hyst[base + src];
      const double r_lct = v.lct[base + src], r_gct = v.gct[base + src];
      const int r_bin = s_key[src] % nb + 1;
#pragma unroll 4
      for (int s = 0; s < s_st; s++) v.vol_alt[(size_t)s * NS + base + q] = svol[(size_t)s * n2 + src];
      // the unstaged rows: gathered from the cell's own rows, which this CTA has just streamed (cache hits)
      for (int s = s_st; s < v.S; s += 4) {
        double a[4];
#pragma unroll
        for (int u = 0; u < 4; u++) a[u] = (s + u < v.S) ? v.vol[(size_t)(s + u) * NS + base + src] : 0.0;
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (s + u < v.S) v.vol_alt[(size_t)(s + u) * NS + base + q] = a[u];
      }
      v.id_alt[base + q] = r_id;
      v.wgc_alt[base + q] = r_wgc;
      v.nprim_alt[base + q] = r_nprim;
      v.hyst_alt[base + q] = r_hyst;
      v.lct_alt[base + q] = r_lct;
      v.gct_alt[base + q] = r_gct;
      v.pbin_alt[base + q] = r_bin;
      if (v.cold.n) cold_copy(v.cold_alt, NS, base + q, v.cold, NS, base + src, v.n_opq);
    }
  }
}

static void swap_slabs(DevView& v);
void pmc_swap_slabs(pmcgpu_ctx* ctx) { swap_slabs(ctx->v); }
// aero_state_sort at the end of trans_aero (wrf_pmc_trans_aero.F90:803): the bin-grid decision of
// aero_sorted_remake_if_needed only; the next mc_coag permutes the records anyway
int pmc_sort_grid_only_async(pmcgpu_ctx* ctx, int interior_only) {
  DevView& v = ctx->v;
  const size_t smem = (size_t)(SORT_THREADS / 32 + 1) * v.C * v.nbcap * sizeof(int);
  ProfScope ps_(ctx, "k_sort_cells_grid");
  k_sort_cells<false, SORT_THREADS><<<v.n_cells, SORT_THREADS, smem, ctx->stream>>>(v, interior_only == 1 ? 3 : (interior_only == 2 ? 4 : 2), -1, 0, -1);
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
  std::swap(v.cold, v.cold_alt);
}

// Largest cell the staged variant takes: histograms + n (8 S + 12) bytes must fit the 227 KB of one CTA.
// PMCGPU_SORT_STAGED_MAX (environment) lowers it (0: legacy variant only) -- used by the tests to reach both.
static int sort_staged_limit(const DevView& v, size_t* smem_out, int s_st, size_t budget) {
  const size_t hist = ((size_t)(SORT_THREADS_STAGED / 32 + 1) * v.C * v.nbcap + 3) / 4 * 16;
  const size_t per = 8 * (size_t)s_st + 12;
  *smem_out = 0;
  if ((v.cap & 1) || hist + 64 * per > budget) return -1;  // rows must start on 16-byte boundaries
  long long nmax = (long long)((budget - hist) / per) & ~1LL;
  if (nmax > v.cap) nmax = v.cap;
  if (const char* e = getenv("PMCGPU_SORT_STAGED_MAX")) {
    const long long lim = atoll(e) & ~1LL;
    if (lim < nmax) nmax = lim;
  }
  if (nmax < 2) return -1;
  *smem_out = hist + (size_t)nmax * per + 16;
  return (int)nmax;
}

int pmc_sort_async(pmcgpu_ctx* ctx, int edge_recheck) {
  DevView& v = ctx->v;
  const size_t smem = (size_t)(SORT_THREADS / 32 + 1) * v.C * v.nbcap * sizeof(int);
  if (smem > 200 * 1024) {
    ctx->err = "sort: n_class * max_bins too large for the shared-memory histogram";
    return PMCGPU_ERR_ARG;
  }
  if (smem > 48 * 1024 && smem > ctx->cfg_sort_smem) {
    CUDA_OK(ctx, cudaFuncSetAttribute(k_sort_cells<false, SORT_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem));
    ctx->cfg_sort_smem = smem;
  }
  // Staged rows and shared memory per CTA: 4 rows in <= 56 KB, i.e. 4 CTAs of 512 threads per SM whose load, compute
  // and store phases overlap (measured on K4, S = 20: 3.08 ms; all 20 rows in 224 KB = one CTA per SM: 3.51 ms;
  // 10 rows / 112 KB 3.21; 2 rows / 50 KB 3.31; none 3.20 -- profiles/r02_sort_staging_sweep.txt).
  // PMCGPU_SORT_STAGED_ROWS / PMCGPU_SORT_SMEM_KB (environment) override for measurements.
  size_t smem_small = 0, smem_big = 0;
  int s_st = std::min(v.S, 4);
  size_t budget = 56 * 1024;
  if (const char* e = getenv("PMCGPU_SORT_STAGED_ROWS")) s_st = std::max(0, std::min(v.S, atoi(e)));
  if (const char* e = getenv("PMCGPU_SORT_SMEM_KB")) budget = (size_t)std::max(16, std::min(224, atoi(e))) * 1024;
  // two staged launches: cells that fit the small budget (several CTAs per SM), then the larger ones with up to
  // 224 KB (one CTA per SM); what is larger still takes the legacy variant
  const int n_small = sort_staged_limit(v, &smem_small, s_st, budget);
  const int n_big = (n_small > 0 && n_small < v.cap) ? sort_staged_limit(v, &smem_big, s_st, 224 * 1024) : n_small;
  const int n_split = std::max(n_small, n_big);
  const size_t smem_max = std::max(smem_small, smem_big);
  if (n_split > 0 && smem_max > 48 * 1024 && smem_max > ctx->cfg_sort_staged_smem) {
    CUDA_OK(ctx, cudaFuncSetAttribute(k_sort_cells<true, SORT_THREADS_STAGED>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));
    ctx->cfg_sort_staged_smem = smem_max;
  }
  {
    ProfScope ps_(ctx, "k_sort_cells");
    if (n_small > 0) {
      k_sort_cells<true, SORT_THREADS_STAGED>
          <<<v.n_cells, SORT_THREADS_STAGED, smem_small, ctx->stream>>>(v, edge_recheck, n_small, s_st, -1);
      LAUNCH_CHECK(ctx);
    }
    if (n_big > n_small) {
      k_sort_cells<true, SORT_THREADS_STAGED>
          <<<v.n_cells, SORT_THREADS_STAGED, smem_big, ctx->stream>>>(v, edge_recheck, n_big, s_st, n_small);
      LAUNCH_CHECK(ctx);
    }
    if (n_split < v.cap) {
      k_sort_cells<false, SORT_THREADS><<<v.n_cells, SORT_THREADS, smem, ctx->stream>>>(v, edge_recheck, n_split, 0, -1);
      LAUNCH_CHECK(ctx);
    }
  }
  // cells with an empty grid keep their (empty) slab: nothing to copy
  swap_slabs(v);
  return PMCGPU_OK;
}

extern "C" int pmcgpu_aero_state_sort(pmcgpu_ctx* ctx) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  int rc = pmc_sort_async(ctx, 0);
  if (rc) return rc;
  return pmc_first_status(ctx);
}
