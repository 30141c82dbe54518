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
  if (method > 0) opt.method = method;
  if (d->cond_h.size() != d->cells.size()) d->cond_h.assign(d->cells.size(), 0.0);
  opt.h_init = d->cond_h[cell];
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

}  // extern "C"
