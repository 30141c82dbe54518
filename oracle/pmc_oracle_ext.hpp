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
void condense_rates_dt(const CondenseRatesInputs& in, double& dDdot_dt, double& dHdoti_dt, double& dHdotenv_dt);
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
  int method = 4;       // 4: RODAS4 (what the device runs by default), 5: CVODE-style BDF, 3: RODAS3
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

// ---------------------------------------------------------------- transport (pmc_oracle_trans.cpp)
// Global doubly periodic domain of gnx x nz x gny cells, emulated px x py patch decomposition.
// Indices are 1-based like the Fortran; fields are WRF `real` (float), class index slowest.
struct TransDomain {
  int gnx = 0, nz = 0, gny = 0, px = 1, py = 1, n_class = 1;
  float dx = 100.f, dy = 100.f, dt = 1.f;
  int vertmix = 1;
  std::vector<float> u_flux, v_flux, w_flux, xkhh, alt, z, z_at_w, exch_h;
  // probabilities per (cell, class): adv -x +x -y +y -z +z; diff -x +x -y +y; vert[nz] (row of B* A)
  std::vector<double> adv, diff, vert;
  void setup(int gnx_, int nz_, int gny_, int px_, int py_, int n_class_);
  size_t cell(int i, int k, int j) const { return (size_t)(i - 1) + (size_t)gnx * ((size_t)(k - 1) + (size_t)nz * (j - 1)); }
  size_t cell_wrap(int i, int k, int j) const {
    const int iw = ((i - 1) % gnx + gnx) % gnx + 1, jw = ((j - 1) % gny + gny) % gny + 1;
    return cell(iw, k, jw);
  }
  size_t uidx(int i, int k, int j, int c) const {
    return (size_t)(i - 1) + (size_t)(gnx + 1) * ((size_t)(k - 1) + (size_t)nz * ((size_t)(j - 1) + (size_t)gny * (c - 1)));
  }
  size_t vidx(int i, int k, int j, int c) const {
    return (size_t)(i - 1) + (size_t)gnx * ((size_t)(k - 1) + (size_t)nz * ((size_t)(j - 1) + (size_t)(gny + 1) * (c - 1)));
  }
  size_t widx(int i, int k, int j, int c) const {
    return (size_t)(i - 1) + (size_t)gnx * ((size_t)(k - 1) + (size_t)(nz + 1) * ((size_t)(j - 1) + (size_t)gny * (c - 1)));
  }
  size_t widx1(int i, int k, int j) const { return (size_t)(i - 1) + (size_t)gnx * ((size_t)(k - 1) + (size_t)(nz + 1) * (j - 1)); }
};

}  // namespace pmco
