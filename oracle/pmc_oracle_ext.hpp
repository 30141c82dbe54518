// pmc_oracle_ext.hpp -- oracle part 2: condensation and transport declarations.
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
#pragma once
#include "pmc_oracle.hpp"

namespace pmco {

// ---------------------------------------------------------------- condense.F90
struct CondenseRatesInputs {  // condense.F90:62-72
  double T, Tdot, H, p, pdot, V_comp, D, D_dry, kappa;
};
struct CondenseRatesOutputs {  // condense.F90:74-84
  double Ddot, Hdot_i, Hdot_env, dDdot_dD, dDdot_dH, dHdoti_dD, dHdoti_dH, dHdotenv_dD, dHdotenv_dH;
};
void condense_rates(const CondenseRatesInputs& in, CondenseRatesOutputs& out);  // condense.F90:371-502
double sat_vapor_pressure(double temp);                                         // env_state.F90:163-172
void condense_equilib_particle(const EnvState& env, const AeroData& d, Particle& p);  // condense.F90:701-739

// Integrator options.  The reference hands the system to SUNDIALS CVODE (BDF,
// Newton + SPGMR, condense_solver.c:84-181), which is not in the reference tree
// and not installed; the tolerances are the reference's (condense.F90:239-255:
// reltol 1e-8, abstol_i = max(1e-30, 1e-8 (D_i - D_dry_i)), abstol_H = 1e-10,
// mxsteps 1e5).  The time stepper restated here is the one the device runs: a
// stiffly accurate, L-stable Rosenbrock method (RODAS3 / RODAS4) with embedded
// error estimate and the exact arrow-head Jacobian of
// condense_jac (:566-616), solved by the Schur complement on H exactly as the
// reference's preconditioner condense_jac_solve_f (:625-695) does.
struct CondenseOpts {
  double reltol = 1e-8;
  double abstol_scale = 1e-8;
  double abstol_H = 1e-10;
  int max_steps = 100000;
  double h_init = 0.0;  // 0: start with the whole interval
  int method = 4;       // 3: RODAS3, 4: RODAS4 (what the device runs)
};
struct CondenseStats {
  int n_steps = 0, n_reject = 0, n_rates = 0;
  double water_rel_error = 0.0;
  double h_last = 0.0;
};
// condense_particles (condense.F90:147-325).  env_final.rel_humid is overwritten.
void condense_particles(AeroState& st, const AeroData& d, const EnvState& env_initial, EnvState& env_final,
                        double del_t, const RunCtx& rc, const CondenseOpts& opt, CondenseStats& stats);
void condense_equilib_particles(const EnvState& env, const AeroData& d, AeroState& st, const RunCtx& rc);  // :744-770
// the bare integrator on arrays (used by condense_particles and by the unit tests)
int condense_integrate(int n, double* D, double* H, const double* D_dry, const double* kappa, const double* num_conc,
                       double T0, double p0, double Tdot, double pdot, double del_t, const CondenseOpts& opt,
                       CondenseStats& stats);

struct TransDomain {};  // filled in by the transport restatement

}  // namespace pmco
