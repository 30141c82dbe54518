// ctx.cu -- context lifetime, device memory layout, host<->device transfer of
// aero_state / aero_weight_array / env_state for all cells of one patch.
// Reference types mirrored: aero_state_t (partmc/src/aero_state.F90:69-86),
// aero_particle_t (aero_particle.F90:30-61), env_state_t (env_state.F90:29-76).
#include "pmcgpu_internal.cuh"
#include <cstdio>
#include <cstring>
#include <algorithm>

template <typename T>
static int dev_alloc(pmcgpu_ctx* ctx, T** p, size_t n, bool zero = true) {
  void* q = nullptr;
  CUDA_OK(ctx, cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
  if (zero) CUDA_OK(ctx, cudaMemsetAsync(q, 0, std::max<size_t>(n, 1) * sizeof(T), ctx->stream));
  ctx->allocs.push_back(q);
  *p = (T*)q;
  return 0;
}

__global__ void k_init_cells(DevView v, long long i_off, long long j_off, int nx, int nz, int ny, long long gnx) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= v.n_cells) return;
  const int i = c % nx, k = (c / nx) % nz, j = c / (nx * nz);
  // global cell index in the same i,k,j order over the global domain
  v.gcell[c] = (i_off + i) + gnx * ((long long)k + (long long)nz * (j_off + j));
  v.next_id[c] = 1;
  v.temp[c] = 288.0;
  v.pres[c] = 1e5;
  v.rh[c] = 0.5;
  v.rrho[c] = 1.0;
  v.cellvol[c] = 1.0;
  for (int q = 0; q < v.C * v.G; q++) {
    v.wtype[(size_t)c * v.C * v.G + q] = PMCGPU_WEIGHT_NONE;
    v.wmag[(size_t)c * v.C * v.G + q] = 1.0;
    v.wexp[(size_t)c * v.C * v.G + q] = 0.0;
  }
}

extern "C" int pmcgpu_create(const pmcgpu_config* cfg, pmcgpu_ctx** out) {
  if (!cfg || !out) return PMCGPU_ERR_ARG;
  if (cfg->n_spec < 1 || cfg->n_spec > PMC_MAX_SPEC || cfg->n_group < 1 || cfg->n_class < 1 || cfg->nx < 1 ||
      cfg->ny < 1 || cfg->nz < 1 || cfg->cell_capacity < 1 || cfg->max_bins < 1 || cfg->max_bins > 255)
    return PMCGPU_ERR_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
    // no CPU fallback: the product path needs a CUDA device
    fprintf(stderr, "pmcgpu_create: no usable CUDA device (no CPU fallback exists)\n");
    return PMCGPU_ERR_CUDA;
  }
  pmcgpu_ctx* ctx = new pmcgpu_ctx();
  ctx->cfg = *cfg;
  ctx->launches = 0;
  ctx->cur = 0;
  if (cudaSetDevice(cfg->device) != cudaSuccess) {
    delete ctx;
    return PMCGPU_ERR_CUDA;
  }
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx;
    return PMCGPU_ERR_CUDA;
  }
  DevView& v = ctx->v;
  memset(&v, 0, sizeof(v));
  ctx->d_flag = nullptr;
  ctx->h_flag = nullptr;
  ctx->rank = 0;
  ctx->n_ranks = 1;
  for (int s = 0; s < 4; s++) ctx->nb[s] = 0;
  if (cudaMallocHost((void**)&ctx->h_flag, sizeof(int)) != cudaSuccess) {
    delete ctx;
    return PMCGPU_ERR_CUDA;
  }
  v.n_cells = cfg->nx * cfg->ny * cfg->nz;
  v.cap = cfg->cell_capacity;
  v.S = cfg->n_spec;
  v.G = cfg->n_group;
  v.C = cfg->n_class;
  v.nbcap = cfg->max_bins;
  v.i_water = cfg->i_water;
  v.seed = cfg->seed;
  v.n_cells_global = (long long)cfg->global_nx * cfg->global_ny * cfg->nz;
  v.ideal = cfg->n_part_ideal / (double)(cfg->n_group * cfg->n_class);
  const size_t NS = (size_t)v.n_cells * v.cap;
  int rc = 0;
#define A(p, n) if ((rc = dev_alloc(ctx, &(p), (n)))) { *out = ctx; return rc; }
  A(v.vol, NS * v.S) A(v.id, NS) A(v.wgc, NS) A(v.nprim, NS) A(v.hyst, NS) A(v.lct, NS) A(v.gct, NS)
  A(v.vol_alt, NS * v.S) A(v.id_alt, NS) A(v.wgc_alt, NS) A(v.nprim_alt, NS) A(v.hyst_alt, NS) A(v.lct_alt, NS)
  A(v.gct_alt, NS)
  A(v.count, v.n_cells) A(v.next_id, v.n_cells) A(v.gcell, v.n_cells) A(v.status, v.n_cells)
  A(v.pbin, NS) A(v.pbin_alt, NS) A(v.lnr, NS)
  A(v.bin_count, (size_t)v.n_cells * v.C * v.nbcap)
  A(v.grid_nb, v.n_cells) A(v.grid_min, v.n_cells) A(v.grid_max, v.n_cells) A(v.bounds_valid, v.n_cells)
  A(v.kmax, (size_t)v.n_cells * v.nbcap * v.nbcap)
  A(v.wtype, (size_t)v.n_cells * v.C * v.G) A(v.wmag, (size_t)v.n_cells * v.C * v.G)
  A(v.wexp, (size_t)v.n_cells * v.C * v.G)
  A(v.temp, v.n_cells) A(v.pres, v.n_cells) A(v.rh, v.n_cells) A(v.rrho, v.n_cells) A(v.cellvol, v.n_cells)
  A(v.L, (size_t)v.n_cells * 3 * v.cap) A(v.fidx, NS)
  A(v.lptr, (size_t)v.n_cells * v.C * v.nbcap) A(v.llen, (size_t)v.n_cells * v.C * v.nbcap)
  A(v.lcap, (size_t)v.n_cells * v.C * v.nbcap)
  A(v.coag_samp, v.n_cells) A(v.coag_ncoag, v.n_cells)
  A(v.cond_T1, v.n_cells) A(v.cond_p1, v.n_cells) A(v.cond_h, v.n_cells) A(v.cond_stats, (size_t)3 * v.n_cells)
  A(v.cond_wre, v.n_cells) A(v.cond_T0, v.n_cells) A(v.cond_p0, v.n_cells) A(v.stage, (size_t)5 * v.n_cells)
  A(v.numconc, (size_t)v.n_cells * v.C) A(ctx->d_flag, 1) A(v.work, 4) A(v.ideal_gc, (size_t)v.C * v.G) A(v.w2p, (size_t)6 * v.n_cells)
  double *d_density, *d_kappa, *d_molw;
  int* d_nions;
  A(d_density, v.S) A(d_kappa, v.S) A(d_molw, v.S) A(d_nions, v.S)
  CUDA_OK(ctx, cudaMemcpyAsync(d_density, cfg->density, v.S * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  std::vector<double> zero((size_t)v.S, 0.0), mw((size_t)v.S, 18e-3);
  std::vector<int> zi((size_t)v.S, 0);
  CUDA_OK(ctx, cudaMemcpyAsync(d_kappa, cfg->kappa ? cfg->kappa : zero.data(), v.S * sizeof(double),
                               cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(ctx, cudaMemcpyAsync(d_molw, cfg->molec_weight ? cfg->molec_weight : mw.data(), v.S * sizeof(double),
                               cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(ctx, cudaMemcpyAsync(d_nions, cfg->num_ions ? cfg->num_ions : zi.data(), v.S * sizeof(int),
                               cudaMemcpyHostToDevice, ctx->stream));
  {
    std::vector<double> ideal((size_t)v.C * v.G, v.ideal);
    CUDA_OK(ctx, cudaMemcpyAsync(v.ideal_gc, ideal.data(), ideal.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  }
  v.density = d_density;
  v.kappa = d_kappa;
  v.molw = d_molw;
  v.nions = d_nions;
  ctx->h_density.assign(cfg->density, cfg->density + v.S);
  v.dens_min = *std::min_element(ctx->h_density.begin(), ctx->h_density.end());
  v.dens_max = *std::max_element(ctx->h_density.begin(), ctx->h_density.end());
  v.fr_df = 3.0;  // fractal_set_spherical (fractal.F90:73-82)
  v.fr_r0 = 1e-8;
  v.fr_fill = 1.0;
  v.fr_on = 0;
  if (!cfg->periodic_x || !cfg->periodic_y) {
    // The reference treats every domain that is not doubly periodic as open in both directions
    // (get_array_extents, wrf_pmc_trans_aero.F90:1665-1705)
    if (cfg->periodic_x || cfg->periodic_y) {
      ctx->err = "a domain periodic in one direction only is not supported (the reference treats it as open in both)";
      *out = ctx;
      return PMCGPU_ERR_ARG;
    }
    if (cfg->global_nx < 3 || cfg->global_ny < 3 || (cfg->i_offset == 0 && cfg->nx < 2) || (cfg->j_offset == 0 && cfg->ny < 2) ||
        (cfg->i_offset + cfg->nx == cfg->global_nx && cfg->nx < 2) || (cfg->j_offset + cfg->ny == cfg->global_ny && cfg->ny < 2)) {
      ctx->err = "open domain: a patch on the domain edge must hold the boundary cells and their interior neighbours";
      *out = ctx;
      return PMCGPU_ERR_ARG;
    }
    std::vector<unsigned char> kind((size_t)v.n_cells, 0);
    for (int c = 0; c < v.n_cells; c++) {
      const int gi = cfg->i_offset + c % cfg->nx, gj = cfg->j_offset + c / (cfg->nx * cfg->nz);
      const bool bi = (gi == 0 || gi == cfg->global_nx - 1), bj = (gj == 0 || gj == cfg->global_ny - 1);
      kind[c] = (bi && bj) ? 5 : (gi == 0 ? 1 : (gi == cfg->global_nx - 1 ? 2 : (gj == 0 ? 3 : (gj == cfg->global_ny - 1 ? 4 : 0))));
    }
    unsigned char* dk = nullptr;
    A(dk, (size_t)v.n_cells)
    CUDA_OK(ctx, cudaMemcpyAsync(dk, kind.data(), kind.size(), cudaMemcpyHostToDevice, ctx->stream));
    CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
    v.kind = dk;
    for (int s = 0; s < 4; s++) ctx->nb[s] = -1;  // a single open patch has no neighbours
  }
#undef A
  k_init_cells<<<(v.n_cells + 255) / 256, 256, 0, ctx->stream>>>(v, cfg->i_offset, cfg->j_offset, cfg->nx, cfg->nz,
                                                                  cfg->ny, cfg->global_nx);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  // keep pointers to private copies of nothing host-side: cfg arrays may go away
  ctx->cfg.density = nullptr;
  ctx->cfg.kappa = nullptr;
  ctx->cfg.molec_weight = nullptr;
  ctx->cfg.num_ions = nullptr;
  *out = ctx;
  return PMCGPU_OK;
}

extern "C" int pmcgpu_destroy(pmcgpu_ctx* ctx) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  cudaStreamSynchronize(ctx->stream);
  for (void* p : ctx->allocs) cudaFree(p);
  if (ctx->h_flag) cudaFreeHost(ctx->h_flag);
  if (ctx->h_tot) cudaFreeHost(ctx->h_tot);
  if (ctx->bnd_buf) cudaFree(ctx->bnd_buf);
  pmcgpu_comm_destroy(ctx);
  for (int q = 0; q < 8; q++) {
    if (ctx->cond_stream[q]) cudaStreamDestroy(ctx->cond_stream[q]);
    if (ctx->cond_join[q]) cudaEventDestroy(ctx->cond_join[q]);
  }
  if (ctx->cond_fork) cudaEventDestroy(ctx->cond_fork);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
  return PMCGPU_OK;
}
extern "C" const char* pmcgpu_last_error(pmcgpu_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
extern "C" long long pmcgpu_launch_count(pmcgpu_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" void* pmcgpu_stream(pmcgpu_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" int pmcgpu_synchronize(pmcgpu_ctx* ctx) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

// ---------------------------------------------------------------------------
// packed host arrays <-> per-cell slabs
// ---------------------------------------------------------------------------
__global__ void k_scatter_particles(DevView v, const long long* __restrict__ off, long long N,
                                    const double* __restrict__ vol, const int* __restrict__ wgroup,
                                    const int* __restrict__ wclass, const long long* __restrict__ id,
                                    const int* __restrict__ nprim, const int* __restrict__ hyst,
                                    const double* __restrict__ lct, const double* __restrict__ gct) {
  const int cell = blockIdx.x;
  const long long o = off[cell];
  const int n = (int)(off[cell + 1] - o);
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  for (int p = threadIdx.x; p < n; p += blockDim.x) {
    for (int s = 0; s < v.S; s++) v.vol[(size_t)s * NS + base + p] = vol[(size_t)s * N + o + p];
    const int g = wgroup ? wgroup[o + p] : 1;
    const int c = wclass ? wclass[o + p] : 1;
    v.wgc[base + p] = (c - 1) * v.G + (g - 1);
    v.id[base + p] = id ? id[o + p] : 0;
    v.nprim[base + p] = nprim ? nprim[o + p] : 1;
    v.hyst[base + p] = hyst ? hyst[o + p] : 0;
    v.lct[base + p] = lct ? lct[o + p] : 0.0;
    v.gct[base + p] = gct ? gct[o + p] : 0.0;
    // default cold part: one component {class as source, least_create_time}; pmcgpu_set_components overrides
    if (v.cold.n) cold_new(v.cold, NS, base + p, c, lct ? lct[o + p] : 0.0, v.n_opq);
  }
  if (threadIdx.x == 0) {
    v.count[cell] = n;
    v.status[cell] = 0;
  }
}
__global__ void k_gather_particles(DevView v, const long long* __restrict__ off, long long N, double* vol,
                                   int* wgroup, int* wclass, long long* id, int* nprim, int* hyst, double* lct,
                                   double* gct, int* pbin) {
  const int cell = blockIdx.x;
  const long long o = off[cell];
  const int n = (int)(off[cell + 1] - o);
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  for (int p = threadIdx.x; p < n; p += blockDim.x) {
    if (vol)
      for (int s = 0; s < v.S; s++) vol[(size_t)s * N + o + p] = v.vol[(size_t)s * NS + base + p];
    const int gc = v.wgc[base + p];
    if (wgroup) wgroup[o + p] = gc % v.G + 1;
    if (wclass) wclass[o + p] = gc / v.G + 1;
    if (id) id[o + p] = v.id[base + p];
    if (nprim) nprim[o + p] = v.nprim[base + p];
    if (hyst) hyst[o + p] = v.hyst[base + p];
    if (lct) lct[o + p] = v.lct[base + p];
    if (gct) gct[o + p] = v.gct[base + p];
    if (pbin) pbin[o + p] = v.pbin[base + p];
  }
}


extern "C" int pmcgpu_set_particles(pmcgpu_ctx* ctx, const int* cell_count, const double* vol, const int* wgroup,
                                    const int* wclass, const int64_t* id, const int* n_primary, const int* hyst,
                                    const double* lct, const double* gct) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !cell_count) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  std::vector<long long> off((size_t)v.n_cells + 1, 0);
  for (int c = 0; c < v.n_cells; c++) {
    if (cell_count[c] < 0) return PMCGPU_ERR_ARG;
    if (cell_count[c] > v.cap) {
      ctx->err = "pmcgpu_set_particles: cell exceeds cell_capacity";
      return PMCGPU_ERR_CAPACITY;
    }
    off[c + 1] = off[c] + cell_count[c];
  }
  const long long N = off[v.n_cells];
  if (N > 0 && !vol) return PMCGPU_ERR_ARG;
  TmpDev<long long> d_off, d_id;
  TmpDev<double> d_vol, d_lct, d_gct;
  TmpDev<int> d_wg, d_wc, d_np, d_hy;
  CUDA_OK(ctx, d_off.upload(off.data(), off.size(), ctx->stream));
  CUDA_OK(ctx, d_vol.upload(vol, (size_t)N * v.S, ctx->stream));
  CUDA_OK(ctx, d_wg.upload(wgroup, N, ctx->stream));
  CUDA_OK(ctx, d_wc.upload(wclass, N, ctx->stream));
  CUDA_OK(ctx, d_id.upload((const long long*)id, N, ctx->stream));
  CUDA_OK(ctx, d_np.upload(n_primary, N, ctx->stream));
  CUDA_OK(ctx, d_hy.upload(hyst, N, ctx->stream));
  CUDA_OK(ctx, d_lct.upload(lct, N, ctx->stream));
  CUDA_OK(ctx, d_gct.upload(gct, N, ctx->stream));
  k_scatter_particles<<<v.n_cells, 128, 0, ctx->stream>>>(v, d_off.p, N, d_vol.p, d_wg.p, d_wc.p, d_id.p, d_np.p,
                                                          d_hy.p, d_lct.p, d_gct.p);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

extern "C" int pmcgpu_get_counts(pmcgpu_ctx* ctx, int* cell_count) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !cell_count) return PMCGPU_ERR_ARG;
  CUDA_OK(ctx, cudaMemcpyAsync(cell_count, ctx->v.count, ctx->v.n_cells * sizeof(int), cudaMemcpyDeviceToHost,
                               ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

static int gather_impl(pmcgpu_ctx* ctx, long long n_total, double* vol, int* wgroup, int* wclass, int64_t* id,
                       int* n_primary, int* hyst, double* lct, double* gct, int* pbin) {
  DevView& v = ctx->v;
  std::vector<int> cnt((size_t)v.n_cells);
  int rc = pmcgpu_get_counts(ctx, cnt.data());
  if (rc) return rc;
  std::vector<long long> off((size_t)v.n_cells + 1, 0);
  for (int c = 0; c < v.n_cells; c++) off[c + 1] = off[c] + cnt[c];
  const long long N = off[v.n_cells];
  if (N != n_total) {
    ctx->err = "pmcgpu_get_particles: n_total does not match the device population";
    return PMCGPU_ERR_ARG;
  }
  TmpDev<long long> d_off, d_id;
  TmpDev<double> d_vol, d_lct, d_gct;
  TmpDev<int> d_wg, d_wc, d_np, d_hy, d_pb;
  CUDA_OK(ctx, d_off.upload(off.data(), off.size(), ctx->stream));
  if (vol) CUDA_OK(ctx, d_vol.alloc((size_t)N * v.S));
  if (wgroup) CUDA_OK(ctx, d_wg.alloc(N));
  if (wclass) CUDA_OK(ctx, d_wc.alloc(N));
  if (id) CUDA_OK(ctx, d_id.alloc(N));
  if (n_primary) CUDA_OK(ctx, d_np.alloc(N));
  if (hyst) CUDA_OK(ctx, d_hy.alloc(N));
  if (lct) CUDA_OK(ctx, d_lct.alloc(N));
  if (gct) CUDA_OK(ctx, d_gct.alloc(N));
  if (pbin) CUDA_OK(ctx, d_pb.alloc(N));
  k_gather_particles<<<v.n_cells, 128, 0, ctx->stream>>>(v, d_off.p, N, d_vol.p, d_wg.p, d_wc.p, d_id.p, d_np.p,
                                                         d_hy.p, d_lct.p, d_gct.p, d_pb.p);
  LAUNCH_CHECK(ctx);
#define D2H(dst, src, n, T) \
  if (dst) CUDA_OK(ctx, cudaMemcpyAsync(dst, src, (size_t)(n) * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  D2H(vol, d_vol.p, N * v.S, double)
  D2H(wgroup, d_wg.p, N, int) D2H(wclass, d_wc.p, N, int) D2H(id, d_id.p, N, long long)
  D2H(n_primary, d_np.p, N, int) D2H(hyst, d_hy.p, N, int) D2H(lct, d_lct.p, N, double)
  D2H(gct, d_gct.p, N, double) D2H(pbin, d_pb.p, N, int)
#undef D2H
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_particles(pmcgpu_ctx* ctx, long long n_total, double* vol, int* wgroup, int* wclass,
                                    int64_t* id, int* n_primary, int* hyst, double* lct, double* gct) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  return gather_impl(ctx, n_total, vol, wgroup, wclass, id, n_primary, hyst, lct, gct, nullptr);
}
extern "C" int pmcgpu_get_bins(pmcgpu_ctx* ctx, long long n_total, int* bin_of_particle) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !bin_of_particle) return PMCGPU_ERR_ARG;
  return gather_impl(ctx, n_total, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                     bin_of_particle);
}

#define H2D_OPT(dst, src, n, T) \
  if (src) CUDA_OK(ctx, cudaMemcpyAsync(dst, src, (size_t)(n) * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
#define D2H_OPT(dst, src, n, T) \
  if (dst) CUDA_OK(ctx, cudaMemcpyAsync(dst, src, (size_t)(n) * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));

extern "C" int pmcgpu_set_next_id(pmcgpu_ctx* ctx, const int64_t* next_id) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !next_id) return PMCGPU_ERR_ARG;
  H2D_OPT(ctx->v.next_id, next_id, ctx->v.n_cells, long long)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_next_id(pmcgpu_ctx* ctx, int64_t* next_id) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !next_id) return PMCGPU_ERR_ARG;
  D2H_OPT(next_id, ctx->v.next_id, ctx->v.n_cells, long long)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_set_weights(pmcgpu_ctx* ctx, const int* type, const double* magnitude, const double* exponent) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  const size_t n = (size_t)ctx->v.n_cells * ctx->v.C * ctx->v.G;
  H2D_OPT(ctx->v.wtype, type, n, int) H2D_OPT(ctx->v.wmag, magnitude, n, double) H2D_OPT(ctx->v.wexp, exponent, n, double)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  if (type) {
    ctx->any_nonflat = false;
    for (size_t i = 0; i < n; i++)
      if (type[i] != PMCGPU_WEIGHT_NONE) ctx->any_nonflat = true;
  }
  return PMCGPU_OK;
}
extern "C" int pmcgpu_set_rng_step(pmcgpu_ctx* ctx, unsigned step) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  ctx->rng_step = step;
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_weights(pmcgpu_ctx* ctx, int* type, double* magnitude, double* exponent) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  const size_t n = (size_t)ctx->v.n_cells * ctx->v.C * ctx->v.G;
  D2H_OPT(type, ctx->v.wtype, n, int) D2H_OPT(magnitude, ctx->v.wmag, n, double) D2H_OPT(exponent, ctx->v.wexp, n, double)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_set_env(pmcgpu_ctx* ctx, const double* temp, const double* pressure, const double* rel_humid,
                              const double* inverse_density, const double* cell_volume) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  const size_t n = ctx->v.n_cells;
  H2D_OPT(ctx->v.temp, temp, n, double) H2D_OPT(ctx->v.pres, pressure, n, double) H2D_OPT(ctx->v.rh, rel_humid, n, double)
  H2D_OPT(ctx->v.rrho, inverse_density, n, double) H2D_OPT(ctx->v.cellvol, cell_volume, n, double)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_rel_humid(pmcgpu_ctx* ctx, double* rel_humid) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !rel_humid) return PMCGPU_ERR_ARG;
  D2H_OPT(rel_humid, ctx->v.rh, ctx->v.n_cells, double)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_grid(pmcgpu_ctx* ctx, int* n_bin, double* grid_min, double* grid_max) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  const size_t n = ctx->v.n_cells;
  D2H_OPT(n_bin, ctx->v.grid_nb, n, int) D2H_OPT(grid_min, ctx->v.grid_min, n, double)
  D2H_OPT(grid_max, ctx->v.grid_max, n, double)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_bin_counts(pmcgpu_ctx* ctx, int* counts) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !counts) return PMCGPU_ERR_ARG;
  D2H_OPT(counts, ctx->v.bin_count, (size_t)ctx->v.n_cells * ctx->v.C * ctx->v.nbcap, int)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_kernel_max(pmcgpu_ctx* ctx, int cell, double* kmax) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !kmax || cell < 0 || cell >= ctx->v.n_cells) return PMCGPU_ERR_ARG;
  int nb = 0;
  CUDA_OK(ctx, cudaMemcpyAsync(&nb, ctx->v.grid_nb + cell, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  D2H_OPT(kmax, ctx->v.kmax + (size_t)cell * ctx->v.nbcap * ctx->v.nbcap, (size_t)nb * nb, double)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_condense_stats(pmcgpu_ctx* ctx, int* stats) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !stats) return PMCGPU_ERR_ARG;
  D2H_OPT(stats, ctx->v.cond_stats, (size_t)3 * ctx->v.n_cells, int)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
// n_part_ideal(n_group, n_class) (aero_state.F90:241 sets every entry to n_part / (n_group n_class); a host
// may give each (group, class) its own value): n_part_ideal[(class - 1) * n_group + (group - 1)]
extern "C" int pmcgpu_set_n_part_ideal(pmcgpu_ctx* ctx, const double* n_part_ideal) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !n_part_ideal) return PMCGPU_ERR_ARG;
  const size_t n = (size_t)ctx->v.C * ctx->v.G;
  for (size_t q = 0; q < n; q++)
    if (!(n_part_ideal[q] > 0.0)) return PMCGPU_ERR_ARG;
  CUDA_OK(ctx, cudaMemcpyAsync(ctx->v.ideal_gc, n_part_ideal, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_work(pmcgpu_ctx* ctx, unsigned long long* work, int reset) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !work) return PMCGPU_ERR_ARG;
  CUDA_OK(ctx, cudaMemcpyAsync(work, ctx->v.work, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
  if (reset) CUDA_OK(ctx, cudaMemsetAsync(ctx->v.work, 0, 4 * sizeof(unsigned long long), ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
// particles at the start of a partmc_timestep: the unit of the benchmark's metric (particle-timesteps)
__global__ void k_count_particles(DevView v) {
  unsigned long long t = 0;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < v.n_cells; c += gridDim.x * blockDim.x) t += (unsigned)v.count[c];
  for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0 && t) atomicAdd(&v.work[2], t);
}
int pmc_count_particles_async(pmcgpu_ctx* ctx) {
  k_count_particles<<<std::min((ctx->v.n_cells + 255) / 256, 296), 256, 0, ctx->stream>>>(ctx->v);
  LAUNCH_CHECK(ctx);
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_coag_stats(pmcgpu_ctx* ctx, int* tot_n_samp, int* tot_n_coag) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  D2H_OPT(tot_n_samp, ctx->v.coag_samp, ctx->v.n_cells, int)
  D2H_OPT(tot_n_coag, ctx->v.coag_ncoag, ctx->v.n_cells, int)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_cell_status(pmcgpu_ctx* ctx, int* status) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !status) return PMCGPU_ERR_ARG;
  D2H_OPT(status, ctx->v.status, ctx->v.n_cells, int)
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

__global__ void k_first_status(const int* __restrict__ status, int n, int* out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    if (status[i] != 0) atomicCAS(out, 0, status[i]);
}
int pmc_first_status(pmcgpu_ctx* ctx) {
  int* d = ctx->d_flag;
  cudaMemsetAsync(d, 0, sizeof(int), ctx->stream);
  k_first_status<<<std::min((ctx->v.n_cells + 255) / 256, 1024), 256, 0, ctx->stream>>>(ctx->v.status, ctx->v.n_cells, d);
  ctx->launches++;
  cudaMemcpyAsync(ctx->h_flag, d, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
  cudaError_t e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    ctx->err = cudaGetErrorString(e);
    return PMCGPU_ERR_CUDA;
  }
  const int h = *ctx->h_flag;
  if (h != 0) ctx->err = "device reported cell status " + std::to_string(h);
  return h;
}

// scenario.F90:331-334 (WRF branch): awa *= old_rrho * (1 / rrho)
__global__ void k_update_density(DevView v, const double* __restrict__ new_rrho) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= v.n_cells) return;
  if (cell_is_boundary(v, c)) {  // scenario_update_aero_state runs for interior cells only; env_state is set everywhere
    v.rrho[c] = new_rrho[c];
    return;
  }
  const double f = v.rrho[c] * (1.0 / new_rrho[c]);
  for (int q = 0; q < v.C * v.G; q++) v.wmag[(size_t)c * v.C * v.G + q] *= f;
  v.rrho[c] = new_rrho[c];
}
int pmc_update_density_async(pmcgpu_ctx* ctx, const double* d_new_rrho) {
  k_update_density<<<(ctx->v.n_cells + 255) / 256, 256, 0, ctx->stream>>>(ctx->v, d_new_rrho);
  LAUNCH_CHECK(ctx);
  return PMCGPU_OK;
}
extern "C" int pmcgpu_update_density(pmcgpu_ctx* ctx, const double* new_inverse_density) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !new_inverse_density) return PMCGPU_ERR_ARG;
  TmpDev<double> d;
  CUDA_OK(ctx, d.upload(new_inverse_density, ctx->v.n_cells, ctx->stream));
  k_update_density<<<(ctx->v.n_cells + 255) / 256, 256, 0, ctx->stream>>>(ctx->v, d.p);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

// wrf_pmc_driver.F90:1704-1719
__global__ void k_num_conc_by_class(DevView v, float* __restrict__ out) {
  const int cell = blockIdx.x;
  const size_t base = (size_t)cell * v.cap;
  extern __shared__ int s_cnt[];
  for (int c = threadIdx.x; c < v.C; c += blockDim.x) s_cnt[c] = 0;
  __syncthreads();
  const int n = v.count[cell];
  for (int p = threadIdx.x; p < n; p += blockDim.x) {
    const int gc = v.wgc[base + p];
    if (gc % v.G == 0) atomicAdd(&s_cnt[gc / v.G], 1);  // i_group = 1
  }
  __syncthreads();
  for (int c = threadIdx.x; c < v.C; c += blockDim.x) {
    const double nc = (double)s_cnt[c] * v.wmag[((size_t)cell * v.C + c) * v.G + 0];
    out[(size_t)cell * v.C + c] = fmaxf((float)nc * (float)v.rrho[cell], 1e-16f);
  }
}
int pmc_num_conc_async(pmcgpu_ctx* ctx, float* d_out) {
  ProfScope ps_(ctx, "k_num_conc_by_class");
  k_num_conc_by_class<<<ctx->v.n_cells, 128, ctx->v.C * sizeof(int), ctx->stream>>>(ctx->v, d_out);
  LAUNCH_CHECK(ctx);
  return PMCGPU_OK;
}
extern "C" int pmcgpu_num_conc_by_class(pmcgpu_ctx* ctx, float* out) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !out) return PMCGPU_ERR_ARG;
  TmpDev<float> d;
  const size_t n = (size_t)ctx->v.n_cells * ctx->v.C;
  CUDA_OK(ctx, d.alloc(n));
  k_num_conc_by_class<<<ctx->v.n_cells, 128, ctx->v.C * sizeof(int), ctx->stream>>>(ctx->v, d.p);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaMemcpyAsync(out, d.p, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

// ---- stand-alone evaluations for known-answer tests
__global__ void k_eval_brown(int n, const double* v1, const double* d1, const double* v2, const double* d2,
                             const double* T, const double* P, double* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const BrownEnv e = brown_env(T[i], P[i]);
  out[i] = brown_pair(brown_part(e, v1[i], d1[i]), brown_part(e, v2[i], d2[i]));
}
extern "C" int pmcgpu_eval_kernel_brown(pmcgpu_ctx* ctx, int n, const double* v1, const double* d1, const double* v2,
                                        const double* d2, const double* temp, const double* pressure, double* out) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n < 0) return PMCGPU_ERR_ARG;
  TmpDev<double> a, b, c, d, t, p, o;
  CUDA_OK(ctx, a.upload(v1, n, ctx->stream));
  CUDA_OK(ctx, b.upload(d1, n, ctx->stream));
  CUDA_OK(ctx, c.upload(v2, n, ctx->stream));
  CUDA_OK(ctx, d.upload(d2, n, ctx->stream));
  CUDA_OK(ctx, t.upload(temp, n, ctx->stream));
  CUDA_OK(ctx, p.upload(pressure, n, ctx->stream));
  CUDA_OK(ctx, o.alloc(n));
  k_eval_brown<<<(n + 127) / 128, 128, 0, ctx->stream>>>(n, a.p, b.p, c.p, d.p, t.p, p.p, o.p);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaMemcpyAsync(out, o.p, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
__global__ void k_eval_philox(unsigned seed, int n, const unsigned* cell, const unsigned* step, const unsigned* phase,
                              const unsigned* a, const unsigned* b, const unsigned* j, double* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  RKey k{seed, cell[i], step[i], phase[i], a[i], b[i]};
  out[i] = rng_uniform(k, j[i]);
}
extern "C" int pmcgpu_eval_philox(pmcgpu_ctx* ctx, int n, const unsigned int* cell, const unsigned int* step,
                                  const unsigned int* phase, const unsigned int* a, const unsigned int* b,
                                  const unsigned int* j, double* out) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n < 0) return PMCGPU_ERR_ARG;
  TmpDev<unsigned> c, s, p, aa, bb, jj;
  TmpDev<double> o;
  CUDA_OK(ctx, c.upload(cell, n, ctx->stream));
  CUDA_OK(ctx, s.upload(step, n, ctx->stream));
  CUDA_OK(ctx, p.upload(phase, n, ctx->stream));
  CUDA_OK(ctx, aa.upload(a, n, ctx->stream));
  CUDA_OK(ctx, bb.upload(b, n, ctx->stream));
  CUDA_OK(ctx, jj.upload(j, n, ctx->stream));
  CUDA_OK(ctx, o.alloc(n));
  k_eval_philox<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->v.seed, n, c.p, s.p, p.p, aa.p, bb.p, jj.p, o.p);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaMemcpyAsync(out, o.p, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

// ---- per-kernel timing with CUDA events on the launching stream (bench.py's roofline)
extern "C" int pmcgpu_profile_enable(pmcgpu_ctx* ctx, int on) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  ctx->prof_on = on != 0;
  return PMCGPU_OK;
}
static int prof_drain(pmcgpu_ctx* ctx) {
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  for (ProfEntry& e : ctx->prof_pending) {
    float ms = 0.f;
    CUDA_OK(ctx, cudaEventElapsedTime(&ms, e.a, e.b));
    auto& acc = ctx->prof_acc[e.name];
    acc.first += ms;
    acc.second += 1;
    ctx->ev_pool.push_back(e.a);
    ctx->ev_pool.push_back(e.b);
  }
  ctx->prof_pending.clear();
  return PMCGPU_OK;
}
// writes "name ms launches\n" lines into buf; reset != 0 clears the accumulators afterwards
extern "C" int pmcgpu_profile_read(pmcgpu_ctx* ctx, char* buf, int buf_len, int reset) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !buf || buf_len < 1) return PMCGPU_ERR_ARG;
  int rc = prof_drain(ctx);
  if (rc) return rc;
  std::string out;
  for (auto& kv : ctx->prof_acc) {
    char line[256];
    snprintf(line, sizeof line, "%s %.6f %lld\n", kv.first.c_str(), kv.second.first, kv.second.second);
    out += line;
  }
  if ((int)out.size() + 1 > buf_len) return PMCGPU_ERR_ARG;
  memcpy(buf, out.c_str(), out.size() + 1);
  if (reset) ctx->prof_acc.clear();
  return PMCGPU_OK;
}
