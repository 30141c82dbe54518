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

extern "C" int pmcgpu_partmc_timestep(pmcgpu_ctx* ctx, const pmcgpu_timestep_args* a) {
  if (!ctx || !a) return PMCGPU_ERR_ARG;
  if (!(a->chem_dt > 0.0)) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  const size_t n = (size_t)v.n_cells, nb = n * sizeof(double);
  cudaStream_t st = ctx->stream;
  // ---- wrf_to_partmc: keep the old (T, p) as env_state_initial, install the new env_state
  const bool chem = fmod(a->elapsed_time, a->chem_dt) == 0.0;  // wrf_pmc_driver.F90:189
  const bool cond = chem && a->do_condensation;
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
  if (!ctx) return PMCGPU_ERR_ARG;
  return pmc_first_status(ctx);
}
