// pmc_oracle_cond.cpp -- oracle restatement of water condensation
// (partmc/src/condense.F90).  TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
#include "pmc_oracle_ext.hpp"
#include <algorithm>
#include <cstdio>

namespace pmco {

double sat_vapor_pressure(double temp) {
  return Const::water_eq_vap_press * std::pow(10.0, 7.45 * (temp - Const::water_freeze_temp) / (temp - 38.0));
}

// condense.F90:371-502, statement for statement
void condense_rates(const CondenseRatesInputs& in, CondenseRatesOutputs& out) {
  const double rho_w = Const::water_density;
  const double M_w = Const::water_molec_weight;
  const double P_0 =
      Const::water_eq_vap_press * std::pow(10.0, 7.45 * (in.T - Const::water_freeze_temp) / (in.T - 38.0));
  const double dP0_dT_div_P0 =
      7.45 * std::log(10.0) * (Const::water_freeze_temp - 38.0) / ((in.T - 38.0) * (in.T - 38.0));
  const double rho_air = Const::air_molec_weight * in.p / (Const::univ_gas_const * in.T);
  const double k_a = 1e-3 * (4.39 + 0.071 * in.T);
  const double D_v = 0.211e-4 / (in.p / Const::air_std_press) * std::pow(in.T / 273.0, 1.94);
  const double U = Const::water_latent_heat * rho_w / (4.0 * in.T);
  const double V = 4.0 * M_w * P_0 / (rho_w * Const::univ_gas_const * in.T);
  const double W = Const::water_latent_heat * M_w / (Const::univ_gas_const * in.T);
  const double X = 4.0 * M_w * Const::water_surf_eng / (Const::univ_gas_const * in.T * rho_w);
  const double Y = 2.0 * k_a / (Const::accom_coeff * rho_air * Const::air_spec_heat) *
                   std::sqrt(2.0 * Const::pi * Const::air_molec_weight / (Const::univ_gas_const * in.T));
  const double Z =
      2.0 * D_v / Const::accom_coeff * std::sqrt(2.0 * Const::pi * M_w / (Const::univ_gas_const * in.T));

  out.Hdot_env = -dP0_dT_div_P0 * in.Tdot * in.H + in.H * in.pdot / in.p;
  out.dHdotenv_dD = 0.0;
  out.dHdotenv_dH = -dP0_dT_div_P0 * in.Tdot + in.pdot / in.p;

  if (in.D <= in.D_dry) {
    const double k_ap = k_a / (1.0 + Y / in.D_dry);
    const double D_vp = D_v / (1.0 + Z / in.D_dry);
    const double delta_star = U * V * D_vp * in.H / k_ap;
    out.Ddot = k_ap * delta_star / (U * in.D_dry);
    out.Hdot_i = -2.0 * Const::pi / (V * in.V_comp) * (in.D_dry * in.D_dry) * out.Ddot;
    const double dh_ddelta = k_ap;
    const double dh_dH = -U * V * D_vp;
    const double ddeltastar_dH = -dh_dH / dh_ddelta;
    out.dDdot_dD = 0.0;
    out.dDdot_dH = k_ap / (U * in.D_dry) * ddeltastar_dH;
    out.dHdoti_dD = -2.0 * Const::pi / (V * in.V_comp) * (in.D_dry * in.D_dry) * out.dDdot_dD;
    out.dHdoti_dH = -2.0 * Const::pi / (V * in.V_comp) * (in.D_dry * in.D_dry) * out.dDdot_dH;
    return;
  }
  const double D3 = in.D * in.D * in.D, Dd3 = in.D_dry * in.D_dry * in.D_dry;
  const double k_ap = k_a / (1.0 + Y / in.D);
  const double dkap_dD = k_a * Y / ((in.D + Y) * (in.D + Y));
  const double D_vp = D_v / (1.0 + Z / in.D);
  const double dDvp_dD = D_v * Z / ((in.D + Z) * (in.D + Z));
  const double a_w = (D3 - Dd3) / (D3 + (in.kappa - 1.0) * Dd3);
  const double den = D3 + (in.kappa - 1.0) * Dd3;
  const double daw_dD = 3.0 * (in.D * in.D) * in.kappa * Dd3 / (den * den);
  double delta_star = 0.0, h = 0.0, dh_ddelta = 1.0;
  for (int newton_step = 1; newton_step <= 5; newton_step++) {
    delta_star = delta_star - h / dh_ddelta;
    const double ex = std::exp(W * delta_star / (1.0 + delta_star) + (X / in.D) / (1.0 + delta_star));
    h = k_ap * delta_star - U * V * D_vp * (in.H - a_w / (1.0 + delta_star) * ex);
    dh_ddelta = k_ap - U * V * D_vp * a_w / ((1.0 + delta_star) * (1.0 + delta_star)) *
                           (1.0 - W / (1.0 + delta_star) + (X / in.D) / (1.0 + delta_star)) * ex;
  }
  out.Ddot = k_ap * delta_star / (U * in.D);
  out.Hdot_i = -2.0 * Const::pi / (V * in.V_comp) * (in.D * in.D) * out.Ddot;
  const double ex = std::exp((W * delta_star) / (1.0 + delta_star) + (X / in.D) / (1.0 + delta_star));
  const double dh_dD = dkap_dD * delta_star - U * V * dDvp_dD * in.H +
                       U * V * (a_w * dDvp_dD + D_vp * daw_dD - D_vp * a_w * (X / (in.D * in.D)) / (1.0 + delta_star)) *
                           (1.0 / (1.0 + delta_star)) * ex;
  const double dh_dH = -U * V * D_vp;
  const double ddeltastar_dD = -dh_dD / dh_ddelta;
  const double ddeltastar_dH = -dh_dH / dh_ddelta;
  out.dDdot_dD = dkap_dD * delta_star / (U * in.D) + k_ap * ddeltastar_dD / (U * in.D) -
                 k_ap * delta_star / (U * (in.D * in.D));
  out.dDdot_dH = k_ap / (U * in.D) * ddeltastar_dH;
  out.dHdoti_dD = -2.0 * Const::pi / (V * in.V_comp) * (2.0 * in.D * out.Ddot + (in.D * in.D) * out.dDdot_dD);
  out.dHdoti_dH = -2.0 * Const::pi / (V * in.V_comp) * (in.D * in.D) * out.dDdot_dH;
}

// Partial derivatives with respect to TIME of the outputs of condense_rates at fixed (D, H): T = T0 + t Tdot and
// p = p0 + t pdot enter through the cell-level terms k_a, D_v, U, V, W, X, Y, Z (condense.F90:391-416), V_comp
// through (T p0) / (T0 p).  delta* is defined implicitly by h(delta*; T, p) = 0 (:449-466), so
// d delta* / dt = -(dh/dt) / (dh/ddelta).  Not a reference routine: the Rosenbrock integrator needs df/dt of the
// non-autonomous system; the reference's CVODE does not.  Checked against a central difference of
// condense_rates in tests/test_oracle_condense.py.
void condense_rates_dt(const CondenseRatesInputs& in, double& dDdot_dt, double& dHdoti_dt, double& dHdotenv_dt) {
  const double rho_w = Const::water_density;
  const double M_w = Const::water_molec_weight;
  const double T = in.T, p = in.p;
  const double P_0 = Const::water_eq_vap_press * std::pow(10.0, 7.45 * (T - Const::water_freeze_temp) / (T - 38.0));
  const double dP0_dT_div_P0 = 7.45 * std::log(10.0) * (Const::water_freeze_temp - 38.0) / ((T - 38.0) * (T - 38.0));
  const double rho_air = Const::air_molec_weight * p / (Const::univ_gas_const * T);
  const double k_a = 1e-3 * (4.39 + 0.071 * T);
  const double D_v = 0.211e-4 / (p / Const::air_std_press) * std::pow(T / 273.0, 1.94);
  const double U = Const::water_latent_heat * rho_w / (4.0 * T);
  const double V = 4.0 * M_w * P_0 / (rho_w * Const::univ_gas_const * T);
  const double W = Const::water_latent_heat * M_w / (Const::univ_gas_const * T);
  const double X = 4.0 * M_w * Const::water_surf_eng / (Const::univ_gas_const * T * rho_w);
  const double Y = 2.0 * k_a / (Const::accom_coeff * rho_air * Const::air_spec_heat) *
                   std::sqrt(2.0 * Const::pi * Const::air_molec_weight / (Const::univ_gas_const * T));
  const double Z = 2.0 * D_v / Const::accom_coeff * std::sqrt(2.0 * Const::pi * M_w / (Const::univ_gas_const * T));
  // logarithmic time derivatives of the cell-level terms
  const double lT = in.Tdot / T, lp = in.pdot / p;
  const double ka_t = 1e-3 * 0.071 * in.Tdot;
  const double lDv = 1.94 * lT - lp;
  const double lU = -lT;
  const double lV = dP0_dT_div_P0 * in.Tdot - lT;
  const double W_t = -W * lT, X_t = -X * lT;
  const double Y_t = Y * (ka_t / k_a + 0.5 * lT - lp);
  const double Z_t = Z * (lDv - 0.5 * lT);
  const double lG = -(lV + lT - lp);  // G = -2 pi / (V V_comp), V_comp ~ T / p
  dHdotenv_dt = 2.0 * dP0_dT_div_P0 / (T - 38.0) * in.Tdot * in.Tdot * in.H - in.H * lp * lp;
  const double G = -2.0 * Const::pi / (V * in.V_comp);
  if (in.D <= in.D_dry) {
    const double D_vp = D_v / (1.0 + Z / in.D_dry);
    const double Ddot = V * D_vp * in.H / in.D_dry;
    dDdot_dt = Ddot * (lV + lDv - Z_t / (in.D_dry + Z));
    dHdoti_dt = G * (in.D_dry * in.D_dry) * (Ddot * lG + dDdot_dt);
    return;
  }
  CondenseRatesOutputs out{};
  condense_rates(in, out);
  const double D = in.D;
  const double D3 = D * D * D, Dd3 = in.D_dry * in.D_dry * in.D_dry;
  const double k_ap = k_a / (1.0 + Y / D);
  const double D_vp = D_v / (1.0 + Z / D);
  const double a_w = (D3 - Dd3) / (D3 + (in.kappa - 1.0) * Dd3);
  const double delta = out.Ddot * (U * D) / k_ap;
  const double q = 1.0 / (1.0 + delta);
  const double E = std::exp((W * delta + X / D) * q);
  const double UVD = U * V * D_vp;
  const double kap_t = k_ap * (ka_t / k_a - Y_t / (D + Y));
  const double UVD_t = UVD * (lU + lV + lDv - Z_t / (D + Z));
  const double E_t = E * q * (W_t * delta + X_t / D);
  const double h_t = kap_t * delta - UVD_t * (in.H - a_w * q * E) + UVD * a_w * q * E_t;
  const double dh_ddelta = k_ap - UVD * a_w * (q * q) * (1.0 - W * q + (X / D) * q) * E;
  const double delta_t = -h_t / dh_ddelta;
  dDdot_dt = out.Ddot * (kap_t / k_ap - lU) + k_ap * delta_t / (U * D);
  dHdoti_dt = G * (D * D) * (out.Ddot * lG + dDdot_dt);
}

// condense.F90:701-739
void condense_equilib_particle(const EnvState& env, const AeroData& d, Particle& p) {
  const double X = 4.0 * Const::water_molec_weight * Const::water_surf_eng /
                   (Const::univ_gas_const * env.temp * Const::water_density);
  const double kappa = particle_solute_kappa(p, d);
  const double D_dry = 2.0 * sphere_vol2rad(particle_solute_volume(p, d));
  double D = D_dry, g = 0.0, dg_dD = 1.0;
  const double Dd3 = D_dry * D_dry * D_dry;
  for (int newton_step = 1; newton_step <= 20; newton_step++) {
    D = D - g / dg_dD;
    const double D3 = D * D * D;
    const double a_w = (D3 - Dd3) / (D3 + (kappa - 1.0) * Dd3);
    const double den = D3 + (kappa - 1.0) * Dd3;
    const double daw_dD = 3.0 * (D * D) * kappa * Dd3 / (den * den);
    g = env.rel_humid - a_w * std::exp(X / D);
    dg_dD = -daw_dD * std::exp(X / D) + a_w * std::exp(X / D) * (X / (D * D));
  }
  p.vol[d.i_water - 1] = sphere_rad2vol(D / 2.0) - sphere_rad2vol(D_dry / 2.0);
}
void condense_equilib_particles(const EnvState& env, const AeroData& d, AeroState& st, const RunCtx& rc) {
  st.valid_sort = false;
  std::vector<double> rw((size_t)st.n_part());
  for (int i = 0; i < st.n_part(); i++) rw[i] = awa_single_num_conc(st.awa, st.apa[i]);
  for (int i = st.n_part(); i >= 1; i--) condense_equilib_particle(env, d, st.apa[i - 1]);
  aero_state_reweight(st, rw, rc, 1000u);
}

// ------------------------------------------------------------------ integrator
namespace {
struct CondSys {
  int n;
  const double *D_dry, *kappa, *num_conc;
  double T0, p0, Tdot, pdot;
  // f and the arrow-head Jacobian at (t, D, H): condense_vf_f / condense_jac (condense.F90:509-616)
  void eval(double t, const double* D, double H, double* fD, double& fH, double* a, double* b, double* c, double& dd,
            int& n_rates, bool jac, double* dfdt = nullptr, double* dfdtH = nullptr) const {
    CondenseRatesInputs in;
    CondenseRatesOutputs out{};
    in.T = T0 + t * Tdot;
    in.p = p0 + t * pdot;
    in.Tdot = Tdot;
    in.pdot = pdot;
    in.H = H;
    double Hdot = 0.0, dHdot_dH = 0.0, Hdot_t = 0.0, env_t = 0.0;
    for (int i = 0; i < n; i++) {
      in.D = D[i];
      in.D_dry = D_dry[i];
      in.V_comp = (in.T * p0) / (T0 * in.p) / num_conc[i];
      in.kappa = kappa[i];
      condense_rates(in, out);
      fD[i] = out.Ddot;
      Hdot += out.Hdot_i;
      if (dfdt) {  // analytic df/dt at fixed (D, H) for the Rosenbrock stages
        double hi_t;
        condense_rates_dt(in, dfdt[i], hi_t, env_t);
        Hdot_t += hi_t;
      }
      if (jac) {
        a[i] = out.dDdot_dD;
        b[i] = out.dDdot_dH;
        c[i] = out.dHdoti_dD + out.dHdotenv_dD;
        dHdot_dH += out.dHdoti_dH;
      }
    }
    if (n == 0) {  // Hdot_env does not depend on the particle
      in.D = 1e-7;
      in.D_dry = 1e-7;
      in.V_comp = 1.0;
      in.kappa = 0.0;
      condense_rates(in, out);
      if (dfdtH) {  // (not `dfdt`: the per-particle array of an empty cell is a null pointer)
        double d0, h0;
        condense_rates_dt(in, d0, h0, env_t);
      }
    }
    n_rates += n;
    fH = Hdot + out.Hdot_env;
    if (dfdtH) *dfdtH = Hdot_t + env_t;
    if (jac) dd = dHdot_dH + out.dHdotenv_dH;  // (an empty cell has no arrays but still has dH'/dH)
  }
};
}  // namespace

// Rosenbrock tableaux in the (A, C, M, E, alpha, gamma_i) form:
//   (1/(h gamma) - J) K_i = f(t + alpha_i h, y + sum_j A_ij K_j) + sum_j (C_ij / h) K_j + gamma_i h df/dt
//   y_new = y + sum_i M_i K_i,  err = sum_i E_i K_i
struct RosTableau {
  int S;
  double gamma;
  double A[6][6], C[6][6], M[6], E[6], alpha[6], gam[6];
  bool newf[6];
  double elo;  // order used by the step-size controller
};
static RosTableau tableau_rodas3() {
  // RODAS3 (Sandu, Verwer, Blom, Spee, Carmichael, Potra 1997): 4 stages, order 3(2),
  // stiffly accurate, L-stable.
  RosTableau t{};
  t.S = 4;
  t.gamma = 0.5;
  t.A[2][0] = 2.0;
  t.A[3][0] = 2.0; t.A[3][2] = 1.0;
  t.C[1][0] = 4.0;
  t.C[2][0] = 1.0; t.C[2][1] = -1.0;
  t.C[3][0] = 1.0; t.C[3][1] = -1.0; t.C[3][2] = -8.0 / 3.0;
  t.M[0] = 2.0; t.M[1] = 0.0; t.M[2] = 1.0; t.M[3] = 1.0;
  t.E[3] = 1.0;
  t.alpha[0] = 0.0; t.alpha[1] = 0.0; t.alpha[2] = 1.0; t.alpha[3] = 1.0;
  t.gam[0] = 0.5; t.gam[1] = 1.5; t.gam[2] = 0.0; t.gam[3] = 0.0;
  t.newf[0] = true; t.newf[1] = false; t.newf[2] = true; t.newf[3] = true;
  t.elo = 3.0;
  return t;
}
static RosTableau tableau_rodas4() {
  // RODAS4 (Hairer & Wanner, Solving ODEs II, sect. IV.7): 6 stages, order 4(3),
  // stiffly accurate, L-stable.
  RosTableau t{};
  t.S = 6;
  t.gamma = 0.25;
  t.alpha[0] = 0.0; t.alpha[1] = 0.386; t.alpha[2] = 0.210; t.alpha[3] = 0.630; t.alpha[4] = 1.0; t.alpha[5] = 1.0;
  t.gam[0] = 0.25; t.gam[1] = -0.1043; t.gam[2] = 0.1035; t.gam[3] = -0.3620000000000023e-01; t.gam[4] = 0.0; t.gam[5] = 0.0;
  t.A[1][0] = 0.1544000000000000e+01;
  t.A[2][0] = 0.9466785280815826; t.A[2][1] = 0.2557011698983284;
  t.A[3][0] = 0.3314825187068521e+01; t.A[3][1] = 0.2896124015972201e+01; t.A[3][2] = 0.9986419139977817;
  t.A[4][0] = 0.1221224509226641e+01; t.A[4][1] = 0.6019134481288629e+01; t.A[4][2] = 0.1253708332932087e+02;
  t.A[4][3] = -0.6878860361058950;
  t.A[5][0] = t.A[4][0]; t.A[5][1] = t.A[4][1]; t.A[5][2] = t.A[4][2]; t.A[5][3] = t.A[4][3]; t.A[5][4] = 1.0;
  t.C[1][0] = -0.5668800000000000e+01;
  t.C[2][0] = -0.2430093356833875e+01; t.C[2][1] = -0.2063599157091915;
  t.C[3][0] = -0.1073529058151375; t.C[3][1] = -0.9594562251023355e+01; t.C[3][2] = -0.2047028614809616e+02;
  t.C[4][0] = 0.7496443313967647e+01; t.C[4][1] = -0.1024680431464352e+02; t.C[4][2] = -0.3399990352819905e+02;
  t.C[4][3] = 0.1170890893206160e+02;
  t.C[5][0] = 0.8083246795921522e+01; t.C[5][1] = -0.7981132988064893e+01; t.C[5][2] = -0.3152159432874371e+02;
  t.C[5][3] = 0.1631930543123136e+02; t.C[5][4] = -0.6058818238834054e+01;
  t.M[0] = t.A[4][0]; t.M[1] = t.A[4][1]; t.M[2] = t.A[4][2]; t.M[3] = t.A[4][3]; t.M[4] = 1.0; t.M[5] = 1.0;
  t.E[5] = 1.0;
  for (int i = 0; i < 6; i++) t.newf[i] = true;
  t.elo = 4.0;
  return t;
}

// ------------------------------------------------------------------ BDF integrator (method 5)
// Restatement of the algorithm the reference hands the system to: SUNDIALS CVODE with CV_BDF
// (condense_solver.c:131-166).  SUNDIALS is not in the reference tree and its version is not pinned
// upstream (condense_solver.c:96-155 compiles against major versions < 6 and >= 6 alike), so what is
// restated is CVODE's published method (Cohen & Hindmarsh, "CVODE, a stiff/nonstiff ODE solver in C",
// Computers in Physics 10 (1996); Hindmarsh et al., ACM TOMS 31 (2005), sect. 2.1): variable-order
// (1..5), variable-step BDF in fixed-leading-coefficient Nordsieck form, with CVODE's local error test
// (weighted RMS norm, weights 1 / (reltol |y_i| + abstol_i)), its step-size / order selection
// heuristics and constants, and its initial step estimate.  The nonlinear system of every step is
// solved by Newton iteration with CVODE's convergence test; each linear system (I - gamma J) x = r
// is solved exactly by the Schur complement on H, which is what the reference's preconditioner
// condense_jac_solve_f (condense.F90:625-695) does inside SPGMR (with an exact preconditioner the
// Krylov iteration has nothing left to do).  J is evaluated at the predicted state of every step.
// The answer at t_final is interpolated from the Nordsieck array like CVode(..., CV_NORMAL).
namespace {
struct BVec {
  std::vector<double> d;
  double H = 0.0;
  explicit BVec(int n = 0) : d((size_t)n, 0.0) {}
};
struct Bdf {
  static constexpr int QMAX = 5;
  static constexpr double ADDON = 1e-6, BIAS1 = 6.0, BIAS2 = 6.0, BIAS3 = 10.0, ETAMX1 = 10000.0, ETAMX2 = 10.0,
                          ETAMX3 = 10.0, ETAMXF = 0.2, ETAMIN = 0.1, ETACF = 0.25, THRESH = 1.5, CRDOWN = 0.3,
                          RDIV = 2.0, CORTES = 0.1;
  static constexpr int SMALL_NST = 10, MXNCF = 10, MXNEF = 7, MXNEF1 = 3, SMALL_NEF = 2, LONG_WAIT = 10, MAXCOR = 3;

  const CondSys& sys;
  const int n;
  const double reltol, atolH;
  const std::vector<double>& atol;
  BVec zn[QMAX + 1], acor, ewt, y, f, tmp;
  std::vector<double> ja, jb, jc, inv;
  double jdd = 0.0, lhsH = 0.0;
  int q = 1, L = 2, qprime = 1, qwait = 2, nst = 0, n_rates = 0, n_fail = 0;
  double h = 0, hprime = 0, eta = 1, hscale = 0, tn = 0, etamax = ETAMX1, tau[QMAX + 2] = {0}, tq[6] = {0},
         l[QMAX + 1] = {0}, rl1 = 1, gamma = 0, saved_tq5 = 0, acnrm = 0, crate = 1;

  Bdf(const CondSys& s, double rt, double aH, const std::vector<double>& at)
      : sys(s), n(s.n), reltol(rt), atolH(aH), atol(at), acor(s.n), ewt(s.n), y(s.n), f(s.n), tmp(s.n),
        ja((size_t)s.n), jb((size_t)s.n), jc((size_t)s.n), inv((size_t)s.n) {
    for (auto& z : zn) z = BVec(s.n);
  }
  double wrms(const BVec& v) const {
    double s = 0.0;
    for (int i = 0; i < n; i++) s += (v.d[i] * ewt.d[i]) * (v.d[i] * ewt.d[i]);
    s += (v.H * ewt.H) * (v.H * ewt.H);
    return std::sqrt(s / (double)(n + 1));
  }
  void set_ewt() {  // cvEwtSetSV
    for (int i = 0; i < n; i++) ewt.d[i] = 1.0 / (reltol * std::fabs(zn[0].d[i]) + atol[i]);
    ewt.H = 1.0 / (reltol * std::fabs(zn[0].H) + atolH);
  }
  void rhs(double t, const BVec& yy, BVec& out, bool jac) {
    double dummy;
    sys.eval(t, yy.d.data(), yy.H, out.d.data(), out.H, jac ? ja.data() : nullptr, jac ? jb.data() : nullptr,
             jac ? jc.data() : nullptr, jac ? jdd : dummy, n_rates, jac);
  }
  // ---- cvHin: initial step from ||y''|| (cvode.c cvHin / cvUpperBoundH0 / cvYddNorm)
  double ydd_norm(double hg) {
    for (int i = 0; i < n; i++) y.d[i] = hg * zn[1].d[i] + zn[0].d[i];
    y.H = hg * zn[1].H + zn[0].H;
    rhs(tn + hg, y, tmp, false);
    for (int i = 0; i < n; i++) tmp.d[i] = (tmp.d[i] - zn[1].d[i]) / hg;
    tmp.H = (tmp.H - zn[1].H) / hg;
    return wrms(tmp);
  }
  double hin(double t0, double tout) {
    const double uround = 2.220446049250313e-16;
    const double tdist = std::fabs(tout - t0);
    const double hlb = 100.0 * uround * std::max(std::fabs(t0), std::fabs(tout));
    double hub_inv = 0.0;  // cvUpperBoundH0
    for (int i = 0; i < n; i++)
      hub_inv = std::max(hub_inv, std::fabs(zn[1].d[i]) / (0.1 * std::fabs(zn[0].d[i]) + 1.0 / ewt.d[i]));
    hub_inv = std::max(hub_inv, std::fabs(zn[1].H) / (0.1 * std::fabs(zn[0].H) + 1.0 / ewt.H));
    double hub = 0.1 * tdist;
    if (hub * hub_inv > 1.0) hub = 1.0 / hub_inv;
    double hg = std::sqrt(hlb * hub), hnew = hg;
    if (hub < hlb) return hg;
    for (int count = 1; count <= 4; count++) {
      const double yddnrm = ydd_norm(hg);
      hnew = (yddnrm * hub * hub > 2.0) ? std::sqrt(2.0 / yddnrm) : std::sqrt(hg * hub);
      if (count == 4) break;
      const double hrat = hnew / hg;
      if (hrat > 0.5 && hrat < 2.0) break;
      if (count > 1 && hrat > 2.0) {
        hnew = hg;
        break;
      }
      hg = hnew;
    }
    return std::min(hub, std::max(hlb, 0.5 * hnew));
  }
  // ---- Nordsieck array manipulation
  void rescale() {  // cvRescale
    double factor = eta;
    for (int j = 1; j <= q; j++) {
      for (int i = 0; i < n; i++) zn[j].d[i] *= factor;
      zn[j].H *= factor;
      factor *= eta;
    }
    h = hscale * eta;
    hscale = h;
  }
  void axpy(double a, const BVec& x, BVec& yv) {
    for (int i = 0; i < n; i++) yv.d[i] += a * x.d[i];
    yv.H += a * x.H;
  }
  void increase_order() {  // cvIncreaseBDF
    for (int i = 0; i <= QMAX; i++) l[i] = 0.0;
    l[2] = 1.0;
    double alpha1 = 1.0, prod = 1.0, xiold = 1.0, alpha0 = -1.0, hsum = hscale;
    if (q > 1) {
      for (int j = 1; j < q; j++) {
        hsum += tau[j + 1];
        const double xi = hsum / hscale;
        prod *= xi;
        alpha0 -= 1.0 / (j + 1);
        alpha1 += 1.0 / xi;
        for (int i = j + 2; i >= 2; i--) l[i] = l[i] * xiold + l[i - 1];
        xiold = xi;
      }
    }
    const double A1 = (-alpha0 - alpha1) / prod;
    for (int i = 0; i < n; i++) zn[L].d[i] = A1 * zn[QMAX].d[i];
    zn[L].H = A1 * zn[QMAX].H;
    for (int j = 2; j <= q; j++) axpy(l[j], zn[L], zn[j]);
  }
  void decrease_order() {  // cvDecreaseBDF
    for (int i = 0; i <= QMAX; i++) l[i] = 0.0;
    l[2] = 1.0;
    double hsum = 0.0;
    for (int j = 1; j <= q - 2; j++) {
      hsum += tau[j];
      const double xi = hsum / hscale;
      for (int i = j + 2; i >= 2; i--) l[i] = l[i] * xi + l[i - 1];
    }
    for (int j = 2; j < q; j++) axpy(-l[j], zn[q], zn[j]);
  }
  void adjust_params() {  // cvAdjustParams
    if (qprime != q) {
      if (qprime > q) increase_order(); else decrease_order();
      q = qprime;
      L = q + 1;
      qwait = L;
    }
    rescale();
  }
  void predict() {  // cvPredict
    tn += h;
    for (int k = 1; k <= q; k++)
      for (int j = q; j >= k; j--) axpy(1.0, zn[j], zn[j - 1]);
  }
  void restore(double saved_t) {  // cvRestore
    tn = saved_t;
    for (int k = 1; k <= q; k++)
      for (int j = q; j >= k; j--) axpy(-1.0, zn[j], zn[j - 1]);
  }
  void set_coeffs() {  // cvSetBDF + cvSetTqBDF
    l[0] = l[1] = 1.0;
    double xi_inv = 1.0, xistar_inv = 1.0;
    for (int i = 2; i <= q; i++) l[i] = 0.0;
    double alpha0 = -1.0, alpha0_hat = -1.0, hsum = h;
    if (q > 1) {
      for (int j = 2; j < q; j++) {
        hsum += tau[j - 1];
        xi_inv = h / hsum;
        alpha0 -= 1.0 / j;
        for (int i = j; i >= 1; i--) l[i] += l[i - 1] * xi_inv;
      }
      alpha0 -= 1.0 / q;
      xistar_inv = -l[1] - alpha0;
      hsum += tau[q - 1];
      xi_inv = h / hsum;
      alpha0_hat = -l[1] - xi_inv;
      for (int i = q; i >= 1; i--) l[i] += l[i - 1] * xistar_inv;
    }
    const double A1 = 1.0 - alpha0_hat + alpha0, A2 = 1.0 + q * A1;
    tq[2] = std::fabs(A1 / (alpha0 * A2));
    tq[5] = std::fabs(A2 * xistar_inv / (l[q] * xi_inv));
    if (qwait == 1) {
      if (q > 1) {
        const double C = xistar_inv / l[q], A3 = alpha0 + 1.0 / q, A4 = alpha0_hat + xi_inv;
        const double Cpinv = (1.0 - A4 + A3) / A3;
        tq[1] = std::fabs(C * Cpinv);
      } else {
        tq[1] = 1.0;
      }
      hsum += tau[q];
      xi_inv = h / hsum;
      const double A5 = alpha0 - 1.0 / (q + 1), A6 = alpha0_hat - xi_inv;
      const double Cppinv = (1.0 - A6 + A5) / A2;
      tq[3] = std::fabs(Cppinv / (xi_inv * (q + 2) * A5));
    }
    tq[4] = CORTES / tq[2];
    rl1 = 1.0 / l[1];
    gamma = h * rl1;
  }
  // ---- Newton iteration of one step (cvNls with the exact arrow-head solve)
  void solve(BVec& b) {  // (I - gamma J) x = b, x overwrites b (condense.F90:665-680)
    double rhs_n = b.H;
    for (int i = 0; i < n; i++) rhs_n += gamma * jc[i] * b.d[i] * inv[i];
    b.H = rhs_n / lhsH;
    for (int i = 0; i < n; i++) b.d[i] = (b.d[i] + gamma * jb[i] * b.H) * inv[i];
  }
  bool nls() {
    crate = 1.0;
    for (int i = 0; i < n; i++) { y.d[i] = zn[0].d[i]; acor.d[i] = 0.0; }
    y.H = zn[0].H;
    acor.H = 0.0;
    rhs(tn, y, f, true);
    lhsH = 1.0 - gamma * jdd;
    for (int i = 0; i < n; i++) {
      inv[i] = 1.0 / (1.0 - gamma * ja[i]);
      lhsH -= (gamma * jb[i]) * (gamma * jc[i]) * inv[i];
    }
    double delp = 0.0;
    for (int m = 0;;) {
      for (int i = 0; i < n; i++) tmp.d[i] = gamma * f.d[i] - (rl1 * zn[1].d[i] + acor.d[i]);
      tmp.H = gamma * f.H - (rl1 * zn[1].H + acor.H);
      solve(tmp);
      const double del = wrms(tmp);
      if (!std::isfinite(del)) return false;
      axpy(1.0, tmp, acor);
      for (int i = 0; i < n; i++) y.d[i] = zn[0].d[i] + acor.d[i];
      y.H = zn[0].H + acor.H;
      if (m > 0) crate = std::max(CRDOWN * crate, del / delp);
      const double dcon = del * std::min(1.0, crate) / tq[4];
      if (dcon <= 1.0) {
        acnrm = (m == 0) ? del : wrms(acor);
        return true;
      }
      m++;
      if (m == MAXCOR || (m >= 2 && del > RDIV * delp)) return false;
      delp = del;
      rhs(tn, y, f, false);
    }
  }
  void complete_step() {  // cvCompleteStep
    nst++;
    for (int i = q; i >= 2; i--) tau[i] = tau[i - 1];
    if (q == 1 && nst > 1) tau[2] = tau[1];
    tau[1] = h;
    for (int j = 0; j <= q; j++) axpy(l[j], acor, zn[j]);
    qwait--;
    if (qwait == 1 && q != QMAX) {
      zn[QMAX] = acor;
      saved_tq5 = tq[5];
    }
  }
  void set_eta() {  // cvSetEta (no hmax)
    if (eta < THRESH) {
      eta = 1.0;
      hprime = h;
    } else {
      eta = std::min(eta, etamax);
      hprime = h * eta;
    }
  }
  void prepare_next_step(double dsm) {  // cvPrepareNextStep, cvComputeEtaqm1/qp1, cvChooseEta
    if (etamax == 1.0) {
      qwait = std::max(qwait, 2);
      qprime = q;
      hprime = h;
      eta = 1.0;
      return;
    }
    const double etaq = 1.0 / (std::pow(BIAS2 * dsm, 1.0 / L) + ADDON);
    if (qwait != 0) {
      eta = etaq;
      qprime = q;
      set_eta();
      return;
    }
    qwait = 2;
    double etaqm1 = 0.0, etaqp1 = 0.0;
    if (q > 1) {
      const double ddn = wrms(zn[q]) * tq[1];
      etaqm1 = 1.0 / (std::pow(BIAS1 * ddn, 1.0 / q) + ADDON);
    }
    if (q != QMAX && saved_tq5 != 0.0) {
      const double cquot = (tq[5] / saved_tq5) * std::pow(h / tau[2], (double)L);
      for (int i = 0; i < n; i++) tmp.d[i] = acor.d[i] - cquot * zn[QMAX].d[i];
      tmp.H = acor.H - cquot * zn[QMAX].H;
      const double dup = wrms(tmp) * tq[3];
      etaqp1 = 1.0 / (std::pow(BIAS3 * dup, 1.0 / (L + 1)) + ADDON);
    }
    const double etam = std::max(etaqm1, std::max(etaq, etaqp1));
    if (etam < THRESH) {
      eta = 1.0;
      qprime = q;
    } else if (etam == etaq) {
      eta = etaq;
      qprime = q;
    } else if (etam == etaqm1) {
      eta = etaqm1;
      qprime = q - 1;
    } else {
      eta = etaqp1;
      qprime = q + 1;
      zn[QMAX] = acor;
    }
    set_eta();
  }
  // one internal step (cvStep); 0 = ok, 7 = repeated convergence / error test failures
  int step() {
    const double saved_t = tn;
    int ncf = 0, nef = 0;
    if (nst > 0 && hprime != h) adjust_params();
    double dsm = 0.0;
    for (;;) {
      predict();
      set_coeffs();
      if (!nls()) {  // cvHandleNFlag
        ncf++;
        n_fail++;
        etamax = 1.0;
        if (ncf == MXNCF || !(std::fabs(h) > 1e-300)) return 7;
        eta = ETACF;
        restore(saved_t);
        rescale();
        continue;
      }
      dsm = acnrm * tq[2];  // cvDoErrorTest
      if (dsm <= 1.0) break;
      nef++;
      n_fail++;
      restore(saved_t);
      if (nef == MXNEF) return 7;
      etamax = 1.0;
      if (nef <= MXNEF1) {
        eta = 1.0 / (std::pow(BIAS2 * dsm, 1.0 / L) + ADDON);
        eta = std::max(ETAMIN, eta);
        if (nef >= SMALL_NEF) eta = std::min(eta, ETAMXF);
        rescale();
        continue;
      }
      eta = ETAMIN;
      if (q > 1) {
        decrease_order();
        L = q;
        q--;
        qwait = L;
        rescale();
        continue;
      }
      h *= eta;
      hscale = h;
      qwait = LONG_WAIT;
      rhs(tn, zn[0], tmp, false);
      for (int i = 0; i < n; i++) zn[1].d[i] = h * tmp.d[i];
      zn[1].H = h * tmp.H;
    }
    complete_step();
    prepare_next_step(dsm);
    etamax = (nst <= SMALL_NST) ? ETAMX2 : ETAMX3;
    for (int i = 0; i < n; i++) acor.d[i] *= tq[2];
    acor.H *= tq[2];
    return 0;
  }
};
}  // namespace

static int condense_integrate_bdf(int n, double* D, double* Hio, const double* D_dry, const double* kappa,
                                  const double* num_conc, double T0, double p0, double Tdot, double pdot,
                                  double del_t, const CondenseOpts& opt, CondenseStats& stats) {
  CondSys sys{n, D_dry, kappa, num_conc, T0, p0, Tdot, pdot};
  std::vector<double> atol((size_t)n);
  for (int i = 0; i < n; i++) atol[i] = std::max(1e-30, opt.abstol_scale * (D[i] - D_dry[i]));
  Bdf s(sys, opt.reltol, opt.abstol_H, atol);
  for (int i = 0; i < n; i++) s.zn[0].d[i] = D[i];
  s.zn[0].H = *Hio;
  s.tn = 0.0;
  s.set_ewt();
  s.rhs(0.0, s.zn[0], s.zn[1], false);
  s.h = s.hin(0.0, del_t);
  s.hscale = s.h;
  s.hprime = s.h;
  for (int i = 0; i < n; i++) s.zn[1].d[i] *= s.h;
  s.zn[1].H *= s.h;
  stats.n_steps = stats.n_reject = stats.n_rates = 0;
  int rc = 0;
  for (;;) {  // CVode(..., CV_NORMAL)
    if (s.nst > 0) s.set_ewt();
    if (s.nst >= opt.max_steps) { rc = 6; break; }
    rc = s.step();
    if (rc) break;
    if (s.tn - del_t >= 0.0) break;
  }
  stats.n_steps = s.nst;
  stats.n_reject = s.n_fail;
  stats.n_rates = s.n_rates;
  stats.h_last = s.h;
  if (rc) return rc;
  // CVodeGetDky(t_final, 0): y = sum_j zn[j] s^j, s = (t_final - tn) / h
  const double sfac = (del_t - s.tn) / s.h;
  for (int i = 0; i < n; i++) {
    double v = s.zn[s.q].d[i];
    for (int j = s.q - 1; j >= 0; j--) v = v * sfac + s.zn[j].d[i];
    D[i] = v;
  }
  double v = s.zn[s.q].H;
  for (int j = s.q - 1; j >= 0; j--) v = v * sfac + s.zn[j].H;
  *Hio = v;
  return 0;
}

int condense_integrate(int n, double* D, double* Hio, const double* D_dry, const double* kappa,
                       const double* num_conc, double T0, double p0, double Tdot, double pdot, double del_t,
                       const CondenseOpts& opt, CondenseStats& stats) {
  if (opt.method == 5)
    return condense_integrate_bdf(n, D, Hio, D_dry, kappa, num_conc, T0, p0, Tdot, pdot, del_t, opt, stats);
  const RosTableau tb = (opt.method == 4) ? tableau_rodas4() : tableau_rodas3();
  const double FAC_MIN = 0.2, FAC_MAX = 6.0, FAC_SAFE = 0.95, FAC_REJ = 0.1;
  CondSys sys{n, D_dry, kappa, num_conc, T0, p0, Tdot, pdot};
  const int S = tb.S;
  std::vector<double> atol((size_t)n), f0((size_t)n), a((size_t)n), b((size_t)n), c((size_t)n),
      y((size_t)n), fs((size_t)n), dfdt((size_t)n), r((size_t)n);
  std::vector<std::vector<double>> K((size_t)S, std::vector<double>((size_t)n));
  double KH[6];
  for (int i = 0; i < n; i++) atol[i] = std::max(1e-30, opt.abstol_scale * (D[i] - D_dry[i]));
  double H = *Hio;
  double t = 0.0;
  stats.n_steps = stats.n_reject = stats.n_rates = 0;
  double h = (opt.h_init > 0.0) ? std::min(opt.h_init, del_t) : del_t;  // else: estimated before the first step
  double h_natural = h;  // last step size chosen by the controller (not by hitting the end of the interval)
  bool reject_last = false, reject_more = false;
  while (t < del_t) {
    if (stats.n_steps + stats.n_reject >= opt.max_steps) return 6;  // PMC_CONDENSE_SOLVER_SET_MAX_STEPS-like
    // f, the arrow-head Jacobian and the analytic df/dt (T and p are linear in t) at (t, y): ONE rate evaluation
    double f0H, dfdtH, dd, dummy;
    sys.eval(t, D, H, f0.data(), f0H, a.data(), b.data(), c.data(), dd, stats.n_rates, true, dfdt.data(), &dfdtH);
    if (stats.n_steps + stats.n_reject == 0 && !(opt.h_init > 0.0)) {
      // first step size, after cvHin (cvode.c): h = 1/2 sqrt(2 / ||y''||), bounded by hub, times 4 for a
      // 4th-order method; y'' = J f + df/dt from the Jacobian of this step (no differencing noise)
      double s1 = 0.0, hub_inv = 0.0, sq = 0.0;
      for (int i = 0; i < n; i++) {
        const double ew = 1.0 / (opt.reltol * std::fabs(D[i]) + atol[i]);
        s1 += c[i] * f0[i];
        hub_inv = std::max(hub_inv, std::fabs(f0[i]) / (0.1 * std::fabs(D[i]) + 1.0 / ew));
        const double w = (a[i] * f0[i] + b[i] * f0H + dfdt[i]) * ew;
        sq += w * w;
      }
      const double ewH = 1.0 / (opt.reltol * std::fabs(H) + opt.abstol_H);
      const double yddH = s1 + dd * f0H + dfdtH;
      hub_inv = std::max(hub_inv, std::fabs(f0H) / (0.1 * std::fabs(H) + 1.0 / ewH));
      const double yddnrm = std::sqrt((sq + (yddH * ewH) * (yddH * ewH)) / (double)(n + 1));
      double hub = 0.1 * del_t;
      if (hub * hub_inv > 1.0) hub = 1.0 / hub_inv;
      const double hnew = (yddnrm * hub * hub > 2.0) ? std::sqrt(2.0 / yddnrm) : hub;
      // rounded down to a power of two, which does not carry the rounding noise of the estimate
      h = 4.0 * 0.5 * hnew;
      {
        int ex;
        std::frexp(h, &ex);
        h = std::ldexp(0.5, ex);
      }
      h = std::min(del_t, h);
      h_natural = h;
    }
    bool truncated = false;
    if (t + h > del_t * (1.0 - 1e-14)) {
      h_natural = h;
      h = del_t - t;
      truncated = true;
    }
    for (;;) {
      const double eps = 1.0 / (tb.gamma * h);
      double sden = 0.0;
      for (int i = 0; i < n; i++) sden += c[i] * b[i] / (eps - a[i]);
      const double lhs_H = eps - dd - sden;
      double fsH = f0H;
      for (int i = 0; i < n; i++) fs[i] = f0[i];
      for (int st = 0; st < S; st++) {
        if (st > 0 && tb.newf[st]) {
          double Hs = H;
          for (int i = 0; i < n; i++) y[i] = D[i];
          for (int j = 0; j < st; j++) {
            if (tb.A[st][j] == 0.0) continue;
            for (int i = 0; i < n; i++) y[i] += tb.A[st][j] * K[j][i];
            Hs += tb.A[st][j] * KH[j];
          }
          sys.eval(t + tb.alpha[st] * h, y.data(), Hs, fs.data(), fsH, nullptr, nullptr, nullptr, dummy,
                   stats.n_rates, false);
        }
        double rH = fsH + h * tb.gam[st] * dfdtH;
        for (int i = 0; i < n; i++) r[i] = fs[i] + h * tb.gam[st] * dfdt[i];
        for (int j = 0; j < st; j++) {
          if (tb.C[st][j] == 0.0) continue;
          const double cj = tb.C[st][j] / h;
          for (int i = 0; i < n; i++) r[i] += cj * K[j][i];
          rH += cj * KH[j];
        }
        double ssum = 0.0;
        for (int i = 0; i < n; i++) ssum += c[i] * r[i] / (eps - a[i]);
        KH[st] = (rH + ssum) / lhs_H;
        for (int i = 0; i < n; i++) K[st][i] = (r[i] + b[i] * KH[st]) / (eps - a[i]);
      }
      // weighted RMS error norm with CVODE-style weights (condense.F90:239-255 tolerances)
      double esum = 0.0;
      for (int i = 0; i < n; i++) {
        double ynew = D[i], e = 0.0;
        for (int j = 0; j < S; j++) {
          ynew += tb.M[j] * K[j][i];
          e += tb.E[j] * K[j][i];
        }
        y[i] = ynew;
        const double w = opt.reltol * std::max(std::fabs(D[i]), std::fabs(ynew)) + atol[i];
        esum += (e / w) * (e / w);
      }
      double Hnew = H, eH = 0.0;
      for (int j = 0; j < S; j++) {
        Hnew += tb.M[j] * KH[j];
        eH += tb.E[j] * KH[j];
      }
      {
        const double w = opt.reltol * std::max(std::fabs(H), std::fabs(Hnew)) + opt.abstol_H;
        esum += (eH / w) * (eH / w);
      }
      double err = std::sqrt(esum / (double)(n + 1));
      if (!std::isfinite(err)) err = 1e10;
      err = std::max(err, 1e-10);
      if (getenv("PMCO_DEBUG")) fprintf(stderr, "cpu t %.17g h %.17g err %.17g\n", t, h, err);
      const double fac = std::min(FAC_MAX, std::max(FAC_MIN, FAC_SAFE / std::pow(err, 1.0 / tb.elo)));
      double hnew = h * fac;
      if (err <= 1.0) {
        for (int i = 0; i < n; i++) D[i] = y[i];
        H = Hnew;
        t += h;
        stats.n_steps++;
        if (reject_last) hnew = std::min(hnew, h);  // no growth right after a rejection
        reject_last = false;
        reject_more = false;
        h = truncated ? std::max(h_natural, hnew) : hnew;
        h_natural = h;
        stats.h_last = h_natural;
        break;
      }
      stats.n_reject++;
      if (stats.n_steps + stats.n_reject >= opt.max_steps) return 6;
      if (reject_more) hnew = h * FAC_REJ;
      reject_more = reject_last;
      reject_last = true;
      h = hnew;
      truncated = false;
      if (!(h > 1e-13 * del_t)) return 7;  // PMC_CONDENSE_SOLVER_FAIL-like
    }
  }
  *Hio = H;
  return 0;
}

void condense_particles(AeroState& st, const AeroData& d, const EnvState& env_initial, EnvState& env_final,
                        double del_t, const RunCtx& rc, const CondenseOpts& opt, CondenseStats& stats) {
  const int n = st.n_part();
  const int iw = d.i_water - 1;
  double water_vol_conc_initial = 0.0;
  for (int i = 0; i < n; i++) water_vol_conc_initial += st.apa[i].vol[iw] * awa_num_conc(st.awa, st.apa[i]);
  const double Tdot = (env_final.temp - env_initial.temp) / del_t;
  const double pdot = (env_final.pressure - env_initial.pressure) / del_t;
  std::vector<double> kappa((size_t)n), D_dry((size_t)n), nc((size_t)n), D((size_t)n);
  for (int i = n - 1; i >= 0; i--) {
    kappa[i] = particle_solute_kappa(st.apa[i], d);
    D_dry[i] = 2.0 * sphere_vol2rad(particle_solute_volume(st.apa[i], d));
    nc[i] = awa_num_conc(st.awa, st.apa[i]);
    D[i] = 2.0 * particle_radius(st.apa[i]);
  }
  double H = env_initial.rel_humid;
  const int rcode = condense_integrate(n, D.data(), &H, D_dry.data(), kappa.data(), nc.data(), env_initial.temp,
                                       env_initial.pressure, Tdot, pdot, del_t, opt, stats);
  if (rcode != 0) die_msg(rcode == 6 ? 592214019 : 592432510, "condense_solver failed");  // condense.F90:331-363
  env_final.rel_humid = H;
  double water_vol_conc_final = 0.0;
  std::vector<double> rw((size_t)n);
  for (int i = 0; i < n; i++) rw[i] = awa_single_num_conc(st.awa, st.apa[i]);
  for (int i = 0; i < n; i++) {
    const double num_conc = awa_num_conc(st.awa, st.apa[i]);
    st.apa[i].vol[iw] = sphere_rad2vol(D[i] / 2.0) - particle_solute_volume(st.apa[i], d);
    st.apa[i].vol[iw] = std::max(0.0, st.apa[i].vol[iw]);
    water_vol_conc_final += st.apa[i].vol[iw] * num_conc;
  }
  // sizes changed: the reference leaves valid_sort alone here (the lists are
  // stale until the next sort); the device always re-sorts, so mark it invalid
  st.valid_sort = false;
  aero_state_reweight(st, rw, rc, 0u);
  const double V_comp_ratio = env_final.temp * env_initial.pressure / (env_initial.temp * env_final.pressure);
  const double rho_w = d.density[iw];
  const double M_w = d.molec_weight[iw];
  const double vapor_i = M_w / (Const::univ_gas_const * env_initial.temp) * sat_vapor_pressure(env_initial.temp) *
                         env_initial.rel_humid / rho_w;
  const double vapor_f = M_w / (Const::univ_gas_const * env_final.temp) * sat_vapor_pressure(env_final.temp) *
                         env_final.rel_humid * V_comp_ratio / rho_w;
  stats.water_rel_error =
      ((vapor_f - vapor_i) + (water_vol_conc_final - water_vol_conc_initial)) / (vapor_f + water_vol_conc_final);
}

}  // namespace pmco
