// checkpoint.cu -- save / restore of the whole device state of a context, the hand-over point for WRF's restart
// alarms (the reference writes one netCDF group per column, partmc/src/output.F90:827-967 and
// interface/wrf_pmc_init.F90:1084; netCDF is outside this library, so the state travels as ONE flat binary file
// per rank that a host-side writer can wrap).  What is saved is exactly what the next step depends on: particle
// records in cell order, weights, env_state, per-cell id counters, bin grids, and the majorant tables with their
// validity flags (they are only recomputed when a grid is remade, coagulation.F90:64-70, so they carry the (T, p)
// of that moment).  Random numbers are counter-based: there is no generator state.  A run restored from a
// checkpoint continues bit for bit (tests/test_gpu_round2.py).
#include "pmcgpu_internal.cuh"
#include <cstdio>
#include <cstring>
#include <vector>

namespace {
const char CK_MAGIC[8] = {'P', 'M', 'C', 'G', 'P', 'U', '0', '3'};
struct CkHeader {
  char magic[8];
  int n_cells, cap, S, G, C, nbcap;
  int nx, nz, ny, global_nx, global_ny, i_offset, j_offset, periodic_x, periodic_y;
  int kmax_ktype, any_nonflat;
  int cold, n_opq;  // component lists / opaque optics carried (pmcgpu_track_components)
  long long n_particles;
};
template <typename T>
int put(pmcgpu_ctx* ctx, FILE* f, const T* dev, size_t n) {
  std::vector<T> h(n ? n : 1);
  if (n) CUDA_OK(ctx, cudaMemcpy(h.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost));
  if (n && fwrite(h.data(), sizeof(T), n, f) != n) return PMCGPU_ERR_ARG;
  return PMCGPU_OK;
}
template <typename T>
int get(pmcgpu_ctx* ctx, FILE* f, T* dev, size_t n) {
  std::vector<T> h(n ? n : 1);
  if (n && fread(h.data(), sizeof(T), n, f) != n) return PMCGPU_ERR_ARG;
  if (n) CUDA_OK(ctx, cudaMemcpy(dev, h.data(), n * sizeof(T), cudaMemcpyHostToDevice));
  return PMCGPU_OK;
}
}  // namespace

#define CK(call)            \
  do {                      \
    int rc_ = (call);       \
    if (rc_) {              \
      fclose(f);            \
      return rc_;           \
    }                       \
  } while (0)

extern "C" int pmcgpu_checkpoint_write(pmcgpu_ctx* ctx, const char* path) {
  if (!ctx || !path) return PMCGPU_ERR_ARG;
  if (cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  DevView& v = ctx->v;
  const size_t nc = (size_t)v.n_cells;
  std::vector<int> counts(nc);
  int rc = pmcgpu_get_counts(ctx, counts.data());
  if (rc) return rc;
  long long N = 0;
  for (size_t c = 0; c < nc; c++) N += counts[c];
  FILE* f = fopen(path, "wb");
  if (!f) {
    ctx->err = std::string("cannot open ") + path;
    return PMCGPU_ERR_ARG;
  }
  CkHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, CK_MAGIC, 8);
  h.n_cells = v.n_cells; h.cap = v.cap; h.S = v.S; h.G = v.G; h.C = v.C; h.nbcap = v.nbcap;
  h.nx = ctx->cfg.nx; h.nz = ctx->cfg.nz; h.ny = ctx->cfg.ny; h.global_nx = ctx->cfg.global_nx;
  h.global_ny = ctx->cfg.global_ny; h.i_offset = ctx->cfg.i_offset; h.j_offset = ctx->cfg.j_offset;
  h.periodic_x = ctx->cfg.periodic_x; h.periodic_y = ctx->cfg.periodic_y;
  h.kmax_ktype = ctx->kmax_ktype; h.any_nonflat = ctx->any_nonflat ? 1 : 0;
  h.cold = v.cold.n ? 1 : 0; h.n_opq = v.n_opq;
  h.n_particles = N;
  if (fwrite(&h, sizeof(h), 1, f) != 1 || fwrite(counts.data(), sizeof(int), nc, f) != nc) {
    fclose(f);
    return PMCGPU_ERR_ARG;
  }
  {  // particle records, packed in cell order like pmcgpu_get_particles
    const size_t n = (size_t)std::max<long long>(N, 1);
    std::vector<double> vol(n * v.S), lct(n), gct(n);
    std::vector<int> wg(n), wc(n), np(n), hy(n);
    std::vector<int64_t> id(n);
    CK(pmcgpu_get_particles(ctx, N, vol.data(), wg.data(), wc.data(), id.data(), np.data(), hy.data(), lct.data(), gct.data()));
    const size_t m = (size_t)N;
    bool ok = fwrite(vol.data(), sizeof(double), m * v.S, f) == m * v.S && fwrite(wg.data(), sizeof(int), m, f) == m &&
              fwrite(wc.data(), sizeof(int), m, f) == m && fwrite(id.data(), sizeof(int64_t), m, f) == m &&
              fwrite(np.data(), sizeof(int), m, f) == m && fwrite(hy.data(), sizeof(int), m, f) == m &&
              fwrite(lct.data(), sizeof(double), m, f) == m && fwrite(gct.data(), sizeof(double), m, f) == m;
    if (!ok) {
      fclose(f);
      return PMCGPU_ERR_ARG;
    }
    if (v.cold.n) {  // cold part of the records: [N] counts, [N][20] sources and times, [N][n_opq] opaque
      std::vector<int> ncmp(n), src(n * PMC_MAX_COMP);
      std::vector<double> tim(n * PMC_MAX_COMP), opq(std::max<size_t>(n * v.n_opq, 1));
      CK(pmcgpu_get_components(ctx, N, ncmp.data(), src.data(), tim.data(), opq.data()));
      ok = fwrite(ncmp.data(), sizeof(int), m, f) == m && fwrite(src.data(), sizeof(int), m * PMC_MAX_COMP, f) == m * PMC_MAX_COMP &&
           fwrite(tim.data(), sizeof(double), m * PMC_MAX_COMP, f) == m * PMC_MAX_COMP &&
           fwrite(opq.data(), sizeof(double), m * v.n_opq, f) == m * v.n_opq;
      if (!ok) {
        fclose(f);
        return PMCGPU_ERR_ARG;
      }
    }
  }
  const size_t ncg = nc * v.C * v.G;
  CK(put(ctx, f, v.wtype, ncg)); CK(put(ctx, f, v.wmag, ncg)); CK(put(ctx, f, v.wexp, ncg));
  CK(put(ctx, f, v.ideal_gc, (size_t)v.C * v.G));
  CK(put(ctx, f, v.temp, nc)); CK(put(ctx, f, v.pres, nc)); CK(put(ctx, f, v.rh, nc)); CK(put(ctx, f, v.rrho, nc));
  CK(put(ctx, f, v.cellvol, nc));
  CK(put(ctx, f, v.next_id, nc)); CK(put(ctx, f, v.status, nc));
  CK(put(ctx, f, v.grid_nb, nc)); CK(put(ctx, f, v.grid_min, nc)); CK(put(ctx, f, v.grid_max, nc));
  CK(put(ctx, f, v.bounds_valid, nc));
  CK(put(ctx, f, v.kmax, nc * v.nbcap * v.nbcap));
  fclose(f);
  return PMCGPU_OK;
}

extern "C" int pmcgpu_checkpoint_read(pmcgpu_ctx* ctx, const char* path) {
  if (!ctx || !path) return PMCGPU_ERR_ARG;
  if (cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;
  DevView& v = ctx->v;
  FILE* f = fopen(path, "rb");
  if (!f) {
    ctx->err = std::string("cannot open ") + path;
    return PMCGPU_ERR_ARG;
  }
  CkHeader h;
  if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, CK_MAGIC, 8) != 0) {
    fclose(f);
    ctx->err = "not a pmcgpu checkpoint";
    return PMCGPU_ERR_ARG;
  }
  if (h.n_cells != v.n_cells || h.cap > v.cap || h.S != v.S || h.G != v.G || h.C != v.C || h.nbcap != v.nbcap ||
      h.nx != ctx->cfg.nx || h.nz != ctx->cfg.nz || h.ny != ctx->cfg.ny || h.global_nx != ctx->cfg.global_nx ||
      h.global_ny != ctx->cfg.global_ny || h.i_offset != ctx->cfg.i_offset || h.j_offset != ctx->cfg.j_offset ||
      h.periodic_x != ctx->cfg.periodic_x || h.periodic_y != ctx->cfg.periodic_y) {
    fclose(f);
    ctx->err = "checkpoint was written by a context with another geometry";
    return PMCGPU_ERR_ARG;
  }
  const size_t nc = (size_t)v.n_cells;
  std::vector<int> counts(nc);
  if (fread(counts.data(), sizeof(int), nc, f) != nc) {
    fclose(f);
    return PMCGPU_ERR_ARG;
  }
  {
    const size_t m = (size_t)h.n_particles, n = std::max<size_t>(m, 1);
    std::vector<double> vol(n * v.S), lct(n), gct(n);
    std::vector<int> wg(n), wc(n), np(n), hy(n);
    std::vector<int64_t> id(n);
    bool ok = fread(vol.data(), sizeof(double), m * v.S, f) == m * v.S && fread(wg.data(), sizeof(int), m, f) == m &&
              fread(wc.data(), sizeof(int), m, f) == m && fread(id.data(), sizeof(int64_t), m, f) == m &&
              fread(np.data(), sizeof(int), m, f) == m && fread(hy.data(), sizeof(int), m, f) == m &&
              fread(lct.data(), sizeof(double), m, f) == m && fread(gct.data(), sizeof(double), m, f) == m;
    if (!ok) {
      fclose(f);
      return PMCGPU_ERR_ARG;
    }
    CK(pmcgpu_set_particles(ctx, counts.data(), vol.data(), wg.data(), wc.data(), id.data(), np.data(), hy.data(),
                            lct.data(), gct.data()));
    if (h.cold) {
      CK(pmcgpu_track_components(ctx, h.n_opq));  // no-op when already enabled with the same width
      std::vector<int> ncmp(n), src(n * PMC_MAX_COMP);
      std::vector<double> tim(n * PMC_MAX_COMP), opq(std::max<size_t>(n * h.n_opq, 1));
      ok = fread(ncmp.data(), sizeof(int), m, f) == m && fread(src.data(), sizeof(int), m * PMC_MAX_COMP, f) == m * PMC_MAX_COMP &&
           fread(tim.data(), sizeof(double), m * PMC_MAX_COMP, f) == m * PMC_MAX_COMP &&
           fread(opq.data(), sizeof(double), m * h.n_opq, f) == m * h.n_opq;
      if (!ok) {
        fclose(f);
        return PMCGPU_ERR_ARG;
      }
      CK(pmcgpu_set_components(ctx, (long long)m, ncmp.data(), src.data(), tim.data(), opq.data()));
    }
  }
  const size_t ncg = nc * v.C * v.G;
  CK(get(ctx, f, v.wtype, ncg)); CK(get(ctx, f, v.wmag, ncg)); CK(get(ctx, f, v.wexp, ncg));
  CK(get(ctx, f, v.ideal_gc, (size_t)v.C * v.G));
  CK(get(ctx, f, v.temp, nc)); CK(get(ctx, f, v.pres, nc)); CK(get(ctx, f, v.rh, nc)); CK(get(ctx, f, v.rrho, nc));
  CK(get(ctx, f, v.cellvol, nc));
  CK(get(ctx, f, v.next_id, nc)); CK(get(ctx, f, v.status, nc));
  CK(get(ctx, f, v.grid_nb, nc)); CK(get(ctx, f, v.grid_min, nc)); CK(get(ctx, f, v.grid_max, nc));
  CK(get(ctx, f, v.bounds_valid, nc));
  CK(get(ctx, f, v.kmax, nc * v.nbcap * v.nbcap));
  fclose(f);
  ctx->kmax_ktype = h.kmax_ktype;
  ctx->any_nonflat = h.any_nonflat != 0;
  ctx->tot_valid = false;   // the first particle messages after a restart use the full pitch
  ctx->w2p_ready = false;
  return PMCGPU_OK;
}
