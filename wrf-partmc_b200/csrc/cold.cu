// cold.cu -- the cold part of the particle record: aero_particle_t%component(:)
// (partmc/src/aero_particle.F90:53-54, aero_component.F90: {source_id, create_time}, at most
// MAX_AERO_COMPONENT_SIZE = 20, aero_particle.F90:22) and the optical fields (:38-49) as an opaque blob of
// n_opaque doubles per particle.  Opt-in (pmcgpu_track_components): a 204-byte record at S = 20 grows by
// 4 + 20 * 12 + 8 n_opaque bytes in both ping-pong slabs and in the halo messages, and every record move carries
// the components of that particle.  Semantics carried through the path:
//   new particle (emission, inflow)   one component {source of the mode, create time}   aero_state.F90:823
//   copy (doubling, dup_particle, resample, transport copy)  identical list               aero_particle copy
//   coagulation                        concatenation, or a sample of 20 split in proportion to n_primary_parts
//                                      (prob_round + two sample_without_replacement)      aero_particle.F90:955-972
//   accelerated coagulation            the source accumulates by the same rule, is then resampled to
//                                      min(n_primary_parts, 20) components                coagulation.F90:472-508
//   optics                             zero on new and coagulated particles (:920-929), copied otherwise
#include "pmcgpu_internal.cuh"

template <typename T>
static int cold_alloc(pmcgpu_ctx* ctx, T** p, size_t n) {
  void* q = nullptr;
  CUDA_OK(ctx, cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
  CUDA_OK(ctx, cudaMemsetAsync(q, 0, std::max<size_t>(n, 1) * sizeof(T), ctx->stream));
  ctx->allocs.push_back(q);
  *p = (T*)q;
  return 0;
}

// every particle already on the device: one component {weight class as source, least_create_time}
__global__ void k_cold_init(DevView v) {
  const int cell = blockIdx.x;
  const size_t base = (size_t)cell * v.cap, NS = (size_t)v.n_cells * v.cap;
  const int n = v.count[cell];
  for (int p = threadIdx.x; p < n; p += blockDim.x)
    cold_new(v.cold, NS, base + p, v.wgc[base + p] / v.G + 1, v.lct[base + p], v.n_opq);
}

extern "C" int pmcgpu_track_components(pmcgpu_ctx* ctx, int n_opaque) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n_opaque < 0 || n_opaque > 64) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  if (v.cold.n) {
    if (n_opaque == v.n_opq) return PMCGPU_OK;
    ctx->err = "pmcgpu_track_components: already enabled with another n_opaque";
    return PMCGPU_ERR_ARG;
  }
  if (ctx->trans_ready || ctx->nccl_comm) {
    ctx->err = "pmcgpu_track_components: call before pmcgpu_comm_init and the first transport step (message layout)";
    return PMCGPU_ERR_ARG;
  }
  const size_t NS = (size_t)v.n_cells * v.cap;
  int rc;
  for (ColdSlab* c : {&v.cold, &v.cold_alt}) {
    if ((rc = cold_alloc(ctx, &c->n, NS))) return rc;
    if ((rc = cold_alloc(ctx, &c->src, NS * PMC_MAX_COMP))) return rc;
    if ((rc = cold_alloc(ctx, &c->time, NS * PMC_MAX_COMP))) return rc;
    if ((rc = cold_alloc(ctx, &c->opq, NS * (size_t)n_opaque))) return rc;
  }
  v.n_opq = n_opaque;
  k_cold_init<<<v.n_cells, 128, 0, ctx->stream>>>(v);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

// packed [N][PMC_MAX_COMP] host layout <-> slabs; dir 0 = to the device, 1 = to the host
__global__ void k_cold_xfer(DevView v, const long long* __restrict__ off, int dir, int* n_comp, int* source,
                            double* create_time, double* opaque) {
  const int cell = blockIdx.x;
  const long long o = off[cell];
  const int n = (int)(off[cell + 1] - o);
  const size_t base = (size_t)cell * v.cap, NS = (size_t)v.n_cells * v.cap;
  for (int p = threadIdx.x; p < n; p += blockDim.x) {
    const size_t h = (size_t)(o + p);
    if (dir == 0) {
      int nc = n_comp[h];
      nc = nc < 0 ? 0 : (nc > PMC_MAX_COMP ? PMC_MAX_COMP : nc);
      v.cold.n[base + p] = nc;
      for (int q = 0; q < nc; q++) {
        v.cold.src[(size_t)q * NS + base + p] = source[h * PMC_MAX_COMP + q];
        v.cold.time[(size_t)q * NS + base + p] = create_time[h * PMC_MAX_COMP + q];
      }
      if (opaque)
        for (int q = 0; q < v.n_opq; q++) v.cold.opq[(size_t)q * NS + base + p] = opaque[h * v.n_opq + q];
    } else {
      const int nc = v.cold.n[base + p];
      if (n_comp) n_comp[h] = nc;
      for (int q = 0; q < PMC_MAX_COMP; q++) {
        if (source) source[h * PMC_MAX_COMP + q] = q < nc ? v.cold.src[(size_t)q * NS + base + p] : 0;
        if (create_time) create_time[h * PMC_MAX_COMP + q] = q < nc ? v.cold.time[(size_t)q * NS + base + p] : 0.0;
      }
      if (opaque)
        for (int q = 0; q < v.n_opq; q++) opaque[h * v.n_opq + q] = v.cold.opq[(size_t)q * NS + base + p];
    }
  }
}

static int cold_xfer(pmcgpu_ctx* ctx, long long n_total, int dir, int* n_comp, int* source, double* create_time,
                     double* opaque) {
  DevView& v = ctx->v;
  if (!v.cold.n) {
    ctx->err = "components are not tracked: call pmcgpu_track_components first";
    return PMCGPU_ERR_ARG;
  }
  std::vector<int> cnt((size_t)v.n_cells);
  CUDA_OK(ctx, cudaMemcpyAsync(cnt.data(), v.count, v.n_cells * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  std::vector<long long> off((size_t)v.n_cells + 1, 0);
  for (int c = 0; c < v.n_cells; c++) off[c + 1] = off[c] + cnt[c];
  const size_t N = (size_t)off[v.n_cells];
  if ((long long)N != n_total) {
    ctx->err = "pmcgpu_set/get_components: n_total does not match the device population";
    return PMCGPU_ERR_ARG;
  }
  TmpDev<long long> d_off;
  TmpDev<int> d_n, d_src;
  TmpDev<double> d_t, d_o;
  CUDA_OK(ctx, d_off.upload(off.data(), off.size(), ctx->stream));
  const size_t nq = N * (size_t)v.n_opq;
  if (dir == 0) {
    CUDA_OK(ctx, d_n.upload(n_comp, N, ctx->stream));
    CUDA_OK(ctx, d_src.upload(source, N * PMC_MAX_COMP, ctx->stream));
    CUDA_OK(ctx, d_t.upload(create_time, N * PMC_MAX_COMP, ctx->stream));
    if (opaque && nq) CUDA_OK(ctx, d_o.upload(opaque, nq, ctx->stream));
  } else {
    CUDA_OK(ctx, d_n.alloc(N));
    CUDA_OK(ctx, d_src.alloc(N * PMC_MAX_COMP));
    CUDA_OK(ctx, d_t.alloc(N * PMC_MAX_COMP));
    if (opaque && nq) CUDA_OK(ctx, d_o.alloc(nq));
  }
  k_cold_xfer<<<v.n_cells, 128, 0, ctx->stream>>>(v, d_off.p, dir, d_n.p, d_src.p, d_t.p, (opaque && nq) ? d_o.p : nullptr);
  LAUNCH_CHECK(ctx);
  if (dir == 1) {
    cudaStream_t st = ctx->stream;
    if (n_comp) CUDA_OK(ctx, cudaMemcpyAsync(n_comp, d_n.p, N * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (source) CUDA_OK(ctx, cudaMemcpyAsync(source, d_src.p, N * PMC_MAX_COMP * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (create_time)
      CUDA_OK(ctx, cudaMemcpyAsync(create_time, d_t.p, N * PMC_MAX_COMP * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (opaque && nq) CUDA_OK(ctx, cudaMemcpyAsync(opaque, d_o.p, nq * sizeof(double), cudaMemcpyDeviceToHost, st));
  }
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

extern "C" int pmcgpu_set_components(pmcgpu_ctx* ctx, long long n_total, const int* n_comp, const int* source,
                                     const double* create_time, const double* opaque) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !n_comp || !source || !create_time) return PMCGPU_ERR_ARG;
  return cold_xfer(ctx, n_total, 0, (int*)n_comp, (int*)source, (double*)create_time, (double*)opaque);
}
extern "C" int pmcgpu_get_components(pmcgpu_ctx* ctx, long long n_total, int* n_comp, int* source, double* create_time,
                                     double* opaque) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  return cold_xfer(ctx, n_total, 1, n_comp, source, create_time, opaque);
}

// aero_state_num_concs_by_source (aero_state.F90:1310-1347): every component of a particle credits its source with
// num_conc * (n_primary_parts / n_components) -- an INTEGER quotient in the Fortran, kept.  One thread per
// (cell, source) walks the particles and their components in order, so the sums are the reference's, term by term.
__global__ void k_num_concs_by_source(DevView v, int n_source, double* __restrict__ out) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= v.n_cells * n_source) return;
  const int cell = q / n_source, source = q % n_source + 1;
  const size_t base = (size_t)cell * v.cap, NS = (size_t)v.n_cells * v.cap;
  const int n = v.count[cell];
  double acc = 0.0;
  for (int p = 0; p < n; p++) {
    const int nc = v.cold.n[base + p];
    if (nc <= 0) continue;
    double vt = 0.0;
    bool hit = false;
    for (int k = 0; k < nc; k++) hit = hit || (v.cold.src[(size_t)k * NS + base + p] == source);
    if (!hit) continue;
    for (int s = 0; s < v.S; s++) vt += v.vol[(size_t)s * NS + base + p];
    const double num_conc = awa_nc_at_radius(v, cell, v.wgc[base + p] / v.G + 1, vol2rad(vt));
    const double term = num_conc * (double)(v.nprim[base + p] / nc);
    for (int k = 0; k < nc; k++)
      if (v.cold.src[(size_t)k * NS + base + p] == source) acc = acc + term;
  }
  out[q] = acc;
}

extern "C" int pmcgpu_num_concs_by_source(pmcgpu_ctx* ctx, int n_source, double* out) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_num_concs_by_source");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n_source < 1 || !out) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  if (!v.cold.n) {
    ctx->err = "components are not tracked: call pmcgpu_track_components first";
    return PMCGPU_ERR_ARG;
  }
  const size_t n = (size_t)v.n_cells * n_source;
  TmpDev<double> d;
  CUDA_OK(ctx, d.alloc(n));
  k_num_concs_by_source<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(v, n_source, d.p);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaMemcpyAsync(out, d.p, n * sizeof(double), cudaMemcpyDefault, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
