// step.cu -- partmc_timestep (interface/wrf_pmc_driver.F90:46-290) for every
// grid cell of the patch: one asynchronous chain of launches on the context's
// stream, one status read at the end.
//
// Per cell the reference runs (wrf_pmc_driver.F90:169-254):
//   wrf_to_partmc (:120, new env_state, old one kept in old_env_states)
//   mc_coag(BROWN, env, chem_dt)            when mod(elapsed_time, chem_dt) == 0 and do_coagulation (:189-197)
//   scenario_update_aero_state              WRF branch: awa *= old_rrho / rrho (scenario.F90:331-334)
//   mosaic_timestep                         (:235; MOSAIC is a private submodule that is not part of the
//                                            reference tree -- its water-uptake slot is filled by
//                                            condense_particles(old_env, env, chem_dt), condense.F90:147,
//                                            as BASELINE.json's north star prescribes)
//   aero_state_rebalance (:248)
//   partmc_to_wrf (:1704-1719)              number concentration per class back to WRF
#include "pmcgpu_internal.cuh"

int pmc_sort_async(pmcgpu_ctx* ctx, int edge_recheck);
int pmc_mc_coag_async(pmcgpu_ctx* ctx, int ktype, double del_t, unsigned step);
int pmc_condense_async(pmcgpu_ctx* ctx, const double* dT0, const double* dp0, const double* dT1, const double* dp1,
                       double del_t, double reltol, unsigned step);
int pmc_rebalance_async(pmcgpu_ctx* ctx, int allow_doubling, int allow_halving, unsigned step);
int pmc_update_density_async(pmcgpu_ctx* ctx, const double* d_new_rrho);
int pmc_num_conc_async(pmcgpu_ctx* ctx, float* d_out);
int pmc_count_particles_async(pmcgpu_ctx* ctx);

// wrf_to_partmc (wrf_pmc_driver.F90:576-644), rel_hum (:649-666), get_grid_cell_volume (:2002-2020): one thread
// per cell; every intermediate has the type the Fortran gives it (see include/pmcgpu.h)
__global__ void k_wrf_to_partmc(int nx, int nz, int ny, const float* __restrict__ p, const float* __restrict__ pb,
                                const float* __restrict__ t_2, const float* __restrict__ qv,
                                const float* __restrict__ alt, const float* __restrict__ z_at_w,
                                const float* __restrict__ msftx, const float* __restrict__ msfty, float rdx, float rdy,
                                double* __restrict__ out) {
  const int n = nx * nz * ny;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  const int i = c % nx, k = (c / nx) % nz, j = c / (nx * nz);
  const float t0 = 300.f, p0 = 100000.f;
  const float r_d = 287.f, cp = 7.f * r_d / 2.f;
  const float rcp = r_d / cp;
  const double pressure = (double)__fadd_rn(p[c], pb[c]);
  const double potential_temp = (double)__fadd_rn(t_2[c], t0);
  const double temperature = potential_temp * pow(pressure / (double)p0, (double)rcp);
  const double q = (double)qv[c];
  const double rh = q / (3.8 * exp((double)17.27f * (temperature - 273.0) / (temperature - 36.0)) / ((double).01f * pressure));
  const double rel = fmax(0.1, fmin(rh, 0.95));
  const size_t w = (size_t)i + (size_t)nx * ((size_t)k + (size_t)(nz + 1) * j);
  const double dz = (double)__fsub_rn(z_at_w[w + nx], z_at_w[w]);
  const double dx = (1.0 / (double)rdx) / (double)(msftx ? msftx[i + nx * j] : 1.f);
  const double dy = (1.0 / (double)rdy) / (double)(msfty ? msfty[i + nx * j] : 1.f);
  out[c] = temperature;
  out[(size_t)n + c] = pressure;
  out[(size_t)2 * n + c] = rel;
  out[(size_t)3 * n + c] = (double)alt[c];
  out[(size_t)4 * n + c] = __dmul_rn(__dmul_rn(dx, dy), dz);
  out[(size_t)5 * n + c] = dz;
}

extern "C" int pmcgpu_wrf_to_partmc(pmcgpu_ctx* ctx, const pmcgpu_wrf_fields* f) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !f || !f->p || !f->pb || !f->t_2 || !f->qv || !f->alt || !f->z_at_w || !(f->rdx > 0.f) || !(f->rdy > 0.f))
    return PMCGPU_ERR_ARG;
  CUDA_OK(ctx, cudaSetDevice(ctx->cfg.device));
  const int nx = ctx->cfg.nx, nz = ctx->cfg.nz, ny = ctx->cfg.ny;
  const size_t n = (size_t)nx * nz * ny, nw = (size_t)nx * (nz + 1) * ny, n2 = (size_t)nx * ny;
  cudaStream_t st = ctx->stream;
  TmpDev<float> buf;  // the WRF fields, staged on the device when they live in host memory
  CUDA_OK(ctx, buf.alloc(5 * n + nw + 2 * n2));
  float* d = buf.p;
  const float* src[8] = {f->p, f->pb, f->t_2, f->qv, f->alt, f->z_at_w, f->msftx, f->msfty};
  const size_t len[8] = {n, n, n, n, n, nw, n2, n2};
  const float* dev[8];
  for (int q = 0; q < 8; q++) {
    dev[q] = nullptr;
    if (src[q]) {
      CUDA_OK(ctx, cudaMemcpyAsync(d, src[q], len[q] * sizeof(float), cudaMemcpyDefault, st));
      dev[q] = d;
    }
    d += len[q];
  }
  k_wrf_to_partmc<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(nx, nz, ny, dev[0], dev[1], dev[2], dev[3], dev[4], dev[5],
                                                               dev[6], dev[7], f->rdx, f->rdy, ctx->v.w2p);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaStreamSynchronize(st));  // the staging buffer is freed on return
  ctx->w2p_ready = true;
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_env_staging(pmcgpu_ctx* ctx, double* temp, double* pressure, double* rel_humid,
                                      double* inverse_density, double* cell_volume, double* height) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !ctx->w2p_ready) return PMCGPU_ERR_ARG;
  const size_t n = (size_t)ctx->v.n_cells;
  double* out[6] = {temp, pressure, rel_humid, inverse_density, cell_volume, height};
  for (int q = 0; q < 6; q++)
    if (out[q]) CUDA_OK(ctx, cudaMemcpyAsync(out[q], ctx->v.w2p + q * n, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

extern "C" int pmcgpu_partmc_timestep(pmcgpu_ctx* ctx, const pmcgpu_timestep_args* a0) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !a0) return PMCGPU_ERR_ARG;
  if (!(a0->chem_dt > 0.0)) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  // env_state fields staged by pmcgpu_wrf_to_partmc stand in for NULL pointers (once)
  pmcgpu_timestep_args a_ = *a0;
  if (ctx->w2p_ready && !a_.temp && !a_.pressure && !a_.rel_humid && !a_.inverse_density && !a_.cell_volume) {
    const size_t nc = (size_t)v.n_cells;
    a_.temp = v.w2p;
    a_.pressure = v.w2p + nc;
    a_.rel_humid = v.w2p + 2 * nc;
    a_.inverse_density = v.w2p + 3 * nc;
    a_.cell_volume = v.w2p + 4 * nc;
    ctx->w2p_ready = false;
  }
  const pmcgpu_timestep_args* a = &a_;
  const size_t n = (size_t)v.n_cells, nb = n * sizeof(double);
  cudaStream_t st = ctx->stream;
  // ---- wrf_to_partmc: keep the old (T, p) as env_state_initial, install the new env_state
  const bool chem = fmod(a->elapsed_time, a->chem_dt) == 0.0;  // wrf_pmc_driver.F90:189
  const bool cond = chem && a->do_condensation;
  if (cond) PMC_SPHERICAL_ONLY(ctx, "pmcgpu_partmc_timestep with do_condensation");
  if (cond) {
    CUDA_OK(ctx, cudaMemcpyAsync(v.cond_T0, v.temp, nb, cudaMemcpyDeviceToDevice, st));
    CUDA_OK(ctx, cudaMemcpyAsync(v.cond_p0, v.pres, nb, cudaMemcpyDeviceToDevice, st));
  }
  if (a->temp) CUDA_OK(ctx, cudaMemcpyAsync(v.temp, a->temp, nb, cudaMemcpyDefault, st));
  if (a->pressure) CUDA_OK(ctx, cudaMemcpyAsync(v.pres, a->pressure, nb, cudaMemcpyDefault, st));
  // condense_particles starts from env_state_initial%rel_humid (condense.F90:237), i.e. from old_env_states
  // (the copy made before wrf_to_partmc refreshes rel_humid, wrf_pmc_driver.F90:615-624), and overwrites the
  // final state's rel_humid (condense.F90:263): WRF's new value only takes effect in steps without condensation
  if (a->rel_humid && !cond) CUDA_OK(ctx, cudaMemcpyAsync(v.rh, a->rel_humid, nb, cudaMemcpyDefault, st));
  if (a->cell_volume) CUDA_OK(ctx, cudaMemcpyAsync(v.cellvol, a->cell_volume, nb, cudaMemcpyDefault, st));
  if (a->inverse_density)
    CUDA_OK(ctx, cudaMemcpyAsync(v.stage, a->inverse_density, nb, cudaMemcpyDefault, st));
  int rc;
  if ((rc = pmc_count_particles_async(ctx))) return rc;
  // ---- mc_coag
  if (chem && a->do_coagulation) {
    if ((rc = pmc_sort_async(ctx, 1))) return rc;
    if ((rc = pmc_mc_coag_async(ctx, a->coag_kernel_type, a->chem_dt, a->step))) return rc;
  }
  // ---- scenario_update_aero_state (WRF branch)
  if (a->inverse_density)
    if ((rc = pmc_update_density_async(ctx, v.stage))) return rc;
  // ---- water uptake
  if (cond) {
    if (v.i_water < 1) {
      ctx->err = "do_condensation needs a water species (aero_data%i_water)";
      return PMCGPU_ERR_ARG;
    }
    if ((rc = pmc_condense_async(ctx, v.cond_T0, v.cond_p0, v.temp, v.pres, a->chem_dt, 0.0, a->step))) return rc;
  }
  // ---- aero_state_rebalance
  if ((rc = pmc_rebalance_async(ctx, a->allow_doubling, a->allow_halving, a->step))) return rc;
  // ---- partmc_to_wrf
  if (a->num_conc_out) {
    if ((rc = pmc_num_conc_async(ctx, v.numconc))) return rc;
    CUDA_OK(ctx, cudaMemcpyAsync(a->num_conc_out, v.numconc, n * v.C * sizeof(float), cudaMemcpyDefault, st));
  }
  if (a->rel_humid_out) CUDA_OK(ctx, cudaMemcpyAsync(a->rel_humid_out, v.rh, nb, cudaMemcpyDefault, st));
  if (a->no_wait) return PMCGPU_OK;
  return pmc_first_status(ctx);
}

extern "C" int pmcgpu_check_status(pmcgpu_ctx* ctx) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  return pmc_first_status(ctx);
}
