// pmc_oracle_ext_c.cpp -- C entry points of the oracle: condensation (and transport).
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
#include "pmc_oracle_domain.hpp"
#include <cstring>
#include <thread>

extern "C" {

// condense_rates (condense.F90:371-502): in[9] = T,Tdot,H,p,pdot,V_comp,D,D_dry,kappa; out[9] as the Fortran type
int pmco_condense_rates(int n, const double* in, double* out) {
  for (int i = 0; i < n; i++) {
    CondenseRatesInputs ci{in[9 * i + 0], in[9 * i + 1], in[9 * i + 2], in[9 * i + 3], in[9 * i + 4],
                           in[9 * i + 5], in[9 * i + 6], in[9 * i + 7], in[9 * i + 8]};
    CondenseRatesOutputs co{};
    condense_rates(ci, co);
    double* o = out + 9 * i;
    o[0] = co.Ddot; o[1] = co.Hdot_i; o[2] = co.Hdot_env; o[3] = co.dDdot_dD; o[4] = co.dDdot_dH;
    o[5] = co.dHdoti_dD; o[6] = co.dHdoti_dH; o[7] = co.dHdotenv_dD; o[8] = co.dHdotenv_dH;
  }
  return 0;
}
// wrf_to_partmc (interface/wrf_pmc_driver.F90:576-644), rel_hum (:649-666), get_grid_cell_volume (:2002-2020)
// for every cell of a patch; out[6][n] = temp, pressure, rel_humid, inverse_density, cell_volume, height.
// Every intermediate carries the type the Fortran gives it: WRF fields are `real`, `grid%p + grid%pb` and
// `grid%t_2 + t0` are single-precision sums assigned to real(dp) variables, t0 = 300., p0 = 100000.,
// rcp = r_d / cp = 287. / (7. * 287. / 2.) are single-precision parameters (module_model_constants.F:23-43)
// promoted to double where they meet a double; the literals 17.27 and .01 of rel_hum are default reals.
int pmco_wrf_to_partmc(int nx, int nz, int ny, const float* p, const float* pb, const float* t_2, const float* qv,
                       const float* alt, const float* z_at_w, const float* msftx, const float* msfty, float rdx,
                       float rdy, double* out) {
  const size_t n = (size_t)nx * nz * ny;
  const float t0 = 300.f, p0 = 100000.f, r_d = 287.f;
  const float cp = 7.f * r_d / 2.f;
  const float rcp = r_d / cp;
  for (int j = 0; j < ny; j++)
    for (int k = 0; k < nz; k++)
      for (int i = 0; i < nx; i++) {
        const size_t c = (size_t)i + (size_t)nx * ((size_t)k + (size_t)nz * j);
        volatile float psum = p[c] + pb[c];    // :617 single-precision sum
        volatile float thsum = t_2[c] + t0;    // :619
        const double pressure = (double)psum;
        const double potential_temp = (double)thsum;
        const double temperature = potential_temp * std::pow(pressure / (double)p0, (double)rcp);  // :620
        const double q = (double)qv[c];
        double rh = q / (3.8 * std::exp((double)17.27f * (temperature - 273.0) / (temperature - 36.0)) /
                         ((double).01f * pressure));                                                   // :662
        rh = std::max(0.1, std::min(rh, 0.95));                                                        // :664
        const size_t w = (size_t)i + (size_t)nx * ((size_t)k + (size_t)(nz + 1) * j);
        volatile float dzf = z_at_w[w + nx] - z_at_w[w];  // :627, :2015
        const double dz = (double)dzf;
        const double dx = (1.0 / (double)rdx) / (double)(msftx ? msftx[i + (size_t)nx * j] : 1.f);  // :2016
        const double dy = (1.0 / (double)rdy) / (double)(msfty ? msfty[i + (size_t)nx * j] : 1.f);  // :2017
        out[c] = temperature;
        out[n + c] = pressure;
        out[2 * n + c] = rh;
        out[3 * n + c] = (double)alt[c];  // :635
        out[4 * n + c] = dx * dy * dz;    // :2019
        out[5 * n + c] = dz;
      }
  return 0;
}
// time derivatives at fixed (D, H): out[3] = dDdot/dt, dHdot_i/dt, dHdot_env/dt
int pmco_condense_rates_dt(int n, const double* in, double* out) {
  for (int i = 0; i < n; i++) {
    CondenseRatesInputs ci{in[9 * i + 0], in[9 * i + 1], in[9 * i + 2], in[9 * i + 3], in[9 * i + 4],
                           in[9 * i + 5], in[9 * i + 6], in[9 * i + 7], in[9 * i + 8]};
    condense_rates_dt(ci, out[3 * i + 0], out[3 * i + 1], out[3 * i + 2]);
  }
  return 0;
}
// bare integrator on arrays; stats: n_steps, n_reject, n_rates
int pmco_condense_integrate(int n, double* D, double* H, const double* D_dry, const double* kappa,
                            const double* num_conc, double T0, double p0, double Tdot, double pdot, double del_t,
                            double reltol, double abstol_scale, double abstol_H, double h_init, int method, int* stats) {
  CondenseOpts opt;
  opt.method = method;
  opt.reltol = reltol;
  opt.abstol_scale = abstol_scale;
  opt.abstol_H = abstol_H;
  opt.h_init = h_init;
  CondenseStats st;
  const int rc = condense_integrate(n, D, H, D_dry, kappa, num_conc, T0, p0, Tdot, pdot, del_t, opt, st);
  if (stats) {
    stats[0] = st.n_steps;
    stats[1] = st.n_reject;
    stats[2] = st.n_rates;
  }
  return rc;
}
// condense_equilib_particles (condense.F90:744-770) with the cell's env_state
int pmco_condense_equilib(void* h, int cell, unsigned step) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  RunCtx rc{d->rng.get(), (uint32_t)d->global_cell[cell], step};
  condense_equilib_particles(d->env[cell], d->data, d->cells[cell], rc);
  ORACLE_CATCH(d)
}
// condense_particles (condense.F90:147-325): env_state_initial = the cell's env, final = (T1, p1);
// on return the cell's env is the final one with rel_humid from the integrator.
// stats_out: n_steps, n_reject, n_rates ; water_rel_error returned through *wre
int pmco_set_condense_method(void* h, int method) {
  ((OracleDomain*)h)->cond_method = method;
  return 0;
}
int pmco_condense_particles(void* h, int cell, double T1, double p1, double del_t, unsigned step, double reltol,
                            int method, int* stats_out, double* wre) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  RunCtx rc{d->rng.get(), (uint32_t)d->global_cell[cell], step};
  EnvState e0 = d->env[cell];
  EnvState e1 = e0;
  e1.temp = T1;
  e1.pressure = p1;
  CondenseOpts opt;
  if (reltol > 0) opt.reltol = reltol;
  if (d->cond_method > 0) opt.method = d->cond_method;
  if (method > 0) opt.method = method;
  if (d->cond_h.size() != d->cells.size()) d->cond_h.assign(d->cells.size(), 0.0);
  CondenseStats st;
  condense_particles(d->cells[cell], d->data, e0, e1, del_t, rc, opt, st);
  d->cond_h[cell] = st.h_last;
  d->env[cell] = e1;
  if (stats_out) {
    stats_out[0] = st.n_steps;
    stats_out[1] = st.n_reject;
    stats_out[2] = st.n_rates;
  }
  if (wre) *wre = st.water_rel_error;
  ORACLE_CATCH(d)
}
double pmco_sat_vapor_pressure(double T) { return sat_vapor_pressure(T); }

// partmc_timestep (interface/wrf_pmc_driver.F90:46-290) for the cells [cell_lo, cell_hi): per cell
// wrf_to_partmc (new T, p[, RH, rrho]; the previous env_state becomes old_env_states, :120), mc_coag when
// mod(elapsed_time, chem_dt) == 0 and do_coag (:189-197), scenario_update_aero_state's WRF branch
// (scenario.F90:331-334), condense_particles(old_env, env, chem_dt) in the slot of mosaic_timestep (:235),
// aero_state_rebalance (:248).  n_threads host threads work on contiguous cell ranges (cells are
// independent; one MPI rank per core in WRF).  The arrays are indexed by cell number (full length).
// tot[0..3] = samples, coagulation events, condensation steps, condensation rate evaluations.
int pmco_partmc_timestep(void* h, int cell_lo, int cell_hi, const double* T, const double* p, const double* RH,
                         const double* rrho, double elapsed_time, double chem_dt, int do_coag, int kernel_type,
                         int do_cond, int allow_doubling, int allow_halving, unsigned step, int n_threads,
                         long long* tot) {
  OracleDomain* d = (OracleDomain*)h;
  const int nc = cell_hi - cell_lo;
  if (n_threads < 1) n_threads = 1;
  if (d->cond_h.size() != d->cells.size()) d->cond_h.assign(d->cells.size(), 0.0);
  std::vector<long long> acc((size_t)n_threads * 4, 0);
  std::vector<int> err((size_t)n_threads, 0);
  const bool chem = std::fmod(elapsed_time, chem_dt) == 0.0;
  auto work = [&](int t) {
    const int lo = cell_lo + (int)((long long)nc * t / n_threads);
    const int hi = cell_lo + (int)((long long)nc * (t + 1) / n_threads);
    for (int c = lo; c < hi; c++) {
      try {
        RunCtx rc{d->rng.get(), (uint32_t)d->global_cell[c], step};
        const EnvState old_env = d->env[c];
        EnvState& env = d->env[c];
        if (T) env.temp = T[c];
        if (p) env.pressure = p[c];
        if (RH) env.rel_humid = RH[c];
        if (rrho) env.inverse_density = rrho[c];
        env.elapsed_time = elapsed_time;
        AeroState& st = d->cells[c];
        if (d->trans.open) {
          // open domain: wrf_to_partmc covers every cell, the aerosol step only the interior ones
          // (wrf_pmc_driver.F90:157-167).  The device sorts every cell before mc_coag (particle ORDER only), so the
          // boundary cells are sorted here too.
          const int gi = c % d->trans.gnx + 1, gj = c / (d->trans.gnx * d->trans.nz) + 1;
          if (d->trans.kind(gi, gj) != 0) {
            if (chem && do_coag) {
              st.valid_sort = false;
              aero_state_sort(st);
              aero_state_sort(st);
              aero_state_physical_sort(st);
            }
            continue;
          }
        }
        if (chem && do_coag) {
          CoagStats cs;
          // the device keeps no sorting between steps: full sort (valid_sort = .false.), the bin-edge
          // safety test a valid sorting gets (aero_sorted.F90:258-276), and the particle array in the
          // device's (class, bin, index) order -- exactly what pmco_sort(force, physical, edge_recheck) does
          st.valid_sort = false;
          aero_state_sort(st);
          aero_state_sort(st);
          aero_state_physical_sort(st);
          mc_coag(kernel_type, env, d->data, st, chem_dt, rc, cs);
          acc[4 * t + 0] += cs.tot_n_samp;
          acc[4 * t + 1] += cs.tot_n_coag;
        }
        if (rrho) {
          const double f = old_env.inverse_density * (1.0 / env.inverse_density);
          for (auto& w : st.awa.w) w.magnitude *= f;
        }
        if (chem && do_cond) {
          // env_state_initial is old_env_states, the full copy made before wrf_to_partmc refreshed rel_humid
          // (wrf_pmc_driver.F90:615-624): the integrator starts from the OLD relative humidity (condense.F90:237)
          EnvState e0 = old_env;
          EnvState e1 = env;
          CondenseOpts opt;
          if (d->cond_method > 0) opt.method = d->cond_method;
          CondenseStats cst;
          condense_particles(st, d->data, e0, e1, chem_dt, rc, opt, cst);
          d->cond_h[c] = cst.h_last;
          env.rel_humid = e1.rel_humid;
          acc[4 * t + 2] += cst.n_steps;
          acc[4 * t + 3] += (long long)cst.n_rates;
        }
        aero_state_rebalance(st, allow_doubling != 0, allow_halving != 0, rc);
      } catch (const PmcError& e) {
        err[t] = e.code ? e.code : -1;
      } catch (const std::exception&) {
        err[t] = -1;
      }
    }
  };
  if (n_threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  int e = 0;
  long long s[4] = {0, 0, 0, 0};
  for (int t = 0; t < n_threads; t++) {
    for (int k = 0; k < 4; k++) s[k] += acc[4 * t + k];
    if (err[t]) e = err[t];
  }
  if (tot)
    for (int k = 0; k < 4; k++) tot[k] = s[k];
  return e;
}

}  // extern "C"
