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

// ---------------------------------------------------------------- emission / inflow sampling (pmc_oracle_emit.cpp)
enum { MODE_LOG_NORMAL = 1, MODE_EXP = 2, MODE_MONO = 3, MODE_SAMPLED = 4 };  // aero_mode.F90:26-37
struct AeroMode {
  int type = MODE_LOG_NORMAL;
  int weight_class = 1;
  double char_radius = 1e-7, log10_std_dev_radius = 0.2, num_conc = 0.0;
  std::vector<double> vol_frac, vol_frac_std;
  // sampled mode (:70-74): bin-edge radii [n + 1] and number concentrations [n]; num_conc is a factor on the table
  std::vector<double> sample_radius, sample_num_conc;
};
double aero_mode_sample_radius(const AeroMode& m, const Weight& w, double prob_threshold, Stream& s);  // aero_mode.F90:549
void aero_mode_sample_vols(const AeroMode& m, double total_vol, std::vector<double>& vols, Stream& s);  // :633
int aero_state_add_aero_dist_sample(AeroState& st, const AeroData& d, const std::vector<AeroMode>& modes,
                                    double sample_prop, double characteristic_factor, double create_time,
                                    bool allow_doubling, bool allow_halving, const RunCtx& rc,
                                    uint32_t stream = 0u);                                              // aero_state.F90:758
// seasalt_emissions (interface/wrf_pmc_driver.F90:1767-1924) for one surface ocean cell
int seasalt_emissions(AeroState& st, const AeroData& d, double wind_speed, double height, double elapsed_time,
                      double delta_t, bool allow_doubling, bool allow_halving, int sea_salt_opt, int i_na, int i_cl,
                      const int* weight_class, const RunCtx& rc);
int seasalt_n_modes(int sea_salt_opt);
void seasalt_modes(int sea_salt_opt, int n_spec, int i_na, int i_cl, const int* weight_class,
                   std::vector<AeroMode>& modes);

// ---------------------------------------------------------------- transport (pmc_oracle_trans.cpp)
// Global doubly periodic domain of gnx x nz x gny cells, emulated px x py patch decomposition.
// Indices are 1-based like the Fortran; fields are WRF `real` (float), class index slowest.
struct TransDomain {
  int gnx = 0, nz = 0, gny = 0, px = 1, py = 1, n_class = 1;
  float dx = 100.f, dy = 100.f, dt = 1.f;
  int vertmix = 1;
  // open (non-periodic) domain: the outermost ring of cells holds the boundary condition
  // (wrf_pmc_driver.F90:157-167, wrf_pmc_trans.F90:116-136, wrf_pmc_trans_aero.F90:689-766, 811-955)
  bool open = false;
  std::vector<AeroMode> bnd_modes;     // background aerosol (scenario%aero_background at the current time)
  std::vector<double> bnd_num_conc;    // [cell][mode] number concentration of every mode in every boundary cell
  std::vector<int> bnd_outflow;        // [cell] 1: the wind blows out of the domain there (dup from the neighbour)
  std::vector<int> bnd_throw;          // [cell] 1: throw_away_previous (wind reversed / background changed)
  // 0 interior, 1 west, 2 east, 3 south, 4 north edge, 5 corner (1-based global indices)
  int kind(int i, int j) const {
    if (!open) return 0;
    const bool bi = (i == 1 || i == gnx), bj = (j == 1 || j == gny);
    if (bi && bj) return 5;
    if (i == 1) return 1;
    if (i == gnx) return 2;
    if (j == 1) return 3;
    if (j == gny) return 4;
    return 0;
  }
  std::vector<float> u_flux, v_flux, w_flux, xkhh, alt, z, z_at_w, exch_h;
  // probabilities per (cell, class): adv -x +x -y +y -z +z; diff -x +x -y +y; vert[nz] (row of B* A)
  std::vector<double> adv, diff, vert;
  void setup(int gnx_, int nz_, int gny_, int px_, int py_, int n_class_);
  size_t cell(int i, int k, int j) const { return (size_t)(i - 1) + (size_t)gnx * ((size_t)(k - 1) + (size_t)nz * (j - 1)); }
  size_t cell_wrap(int i, int k, int j) const {
    if (open) {  // fields beyond the domain: the host hands the nearest cell's value (zero-gradient halo)
      const int ic = i < 1 ? 1 : (i > gnx ? gnx : i), jc = j < 1 ? 1 : (j > gny ? gny : j);
      return cell(ic, k, jc);
    }
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
