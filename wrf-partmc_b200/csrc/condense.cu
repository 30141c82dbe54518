// condense.cu -- condense_particles (partmc/src/condense.F90:147-325) for every
// grid cell of the patch.
//
// The reference integrates, per cell, the state (D_1..D_n, H) with SUNDIALS
// CVODE (BDF + Newton + SPGMR, condense_solver.c:84-181) using the rate
// function condense_rates (:371-502) and, as preconditioner, the exact solve of
// the arrow-head system (I - gamma J) by the Schur complement on H
// (condense_jac_solve_f, :625-695).  SUNDIALS is not part of the reference tree;
// here every particle carries its own implicit stage equations and the cell
// couples them only through H:
//   * time stepper: 4-stage / 4th-order Rosenbrock method (Kaps-Rentrop, Shampine
//     parameters) with embedded 3rd-order error estimate, the reference's
//     tolerances (reltol 1e-8, abstol_i = max(1e-30, 1e-8 (D_i - D_dry_i)),
//     abstol_H = 1e-10) in a CVODE-style weighted RMS norm;
//   * Jacobian: the analytic partials of condense_rates, evaluated once per step;
//   * linear algebra: each stage solves (1/(gamma h) - J) x = r exactly --
//     per particle x_i = (r_i + b_i x_H) / (eps - a_i), with x_H from two
//     cell-wide sums -- so the only cross-particle work is a block reduction;
//   * T and p are linear in time over the step (condense_vf_f, :525-528); df/dt
//     comes from one extra rate evaluation.
// One CTA per cell, particle state in registers (PPT particles per thread),
// FP64-pipe bound: ~6 rate evaluations x (5 Newton steps with one exp each) per
// particle per internal step.  Fused multiply-add is allowed in this file.
#include "pmcgpu_internal.cuh"

#define COND_THREADS 256

struct CondEnv {  // cell-level terms of condense_rates that depend on (T, p) only
  double T, p, Tdot, pdot;
  double k_a, D_v, U, V, W, X, Y, Z, UV;
  double dP0_dT_div_P0;
  double vc;  // V_comp * num_conc = (T * p0) / (T0 * p)
  // scalars every thread would otherwise divide out again in every rate evaluation
  double g_coef;      // -2 pi / (V vc): Hdot_i = g_coef * num_conc * D^2 * Ddot
  double invU;        // 1 / U
  double pdot_div_p;  // pdot / p
  // time derivatives of the terms above at fixed (D, H) (T and p are linear in t): what the analytic df/dt of the
  // Rosenbrock stages needs (the oracle's condense_rates_dt states the same formulas)
  double lka;     // (dk_a/dt) / k_a
  double Y_t, Z_t, W_t, X_t;
  double luvd;    // d ln(U V D_v) / dt
  double lU;      // d ln U / dt
  double lG;      // d ln(-2 pi / (V vc)) / dt
  double envH_t;  // (d Hdot_env / dt) / H
};
__device__ __forceinline__ void cond_env_dt(CondEnv& e) {
  const double lT = e.Tdot / e.T, lp = e.pdot / e.p;
  const double lDv = 1.94 * lT - lp;
  const double lV = e.dP0_dT_div_P0 * e.Tdot - lT;
  e.lka = 1e-3 * 0.071 * e.Tdot / e.k_a;
  e.lU = -lT;
  e.W_t = -e.W * lT;
  e.X_t = -e.X * lT;
  e.Y_t = e.Y * (e.lka + 0.5 * lT - lp);
  e.Z_t = e.Z * (lDv - 0.5 * lT);
  e.luvd = e.lU + lV + lDv;
  e.lG = -(lV + lT - lp);
  e.envH_t = 2.0 * e.dP0_dT_div_P0 / (e.T - 38.0) * e.Tdot * e.Tdot - lp * lp;
}
__device__ __forceinline__ CondEnv cond_env(double T0, double p0, double Tdot, double pdot, double t) {
  CondEnv e;
  e.T = T0 + t * Tdot;
  e.p = p0 + t * pdot;
  e.Tdot = Tdot;
  e.pdot = pdot;
  const double rho_w = PmcConst::water_density, M_w = PmcConst::water_molec_weight;
  const double P_0 =
      PmcConst::water_eq_vap_press * pow(10.0, 7.45 * (e.T - PmcConst::water_freeze_temp) / (e.T - 38.0));
  e.dP0_dT_div_P0 = 7.45 * log(10.0) * (PmcConst::water_freeze_temp - 38.0) / ((e.T - 38.0) * (e.T - 38.0));
  const double rho_air = PmcConst::air_molec_weight * e.p / (PmcConst::univ_gas_const * e.T);
  e.k_a = 1e-3 * (4.39 + 0.071 * e.T);
  e.D_v = 0.211e-4 / (e.p / PmcConst::air_std_press) * pow(e.T / 273.0, 1.94);
  e.U = PmcConst::water_latent_heat * rho_w / (4.0 * e.T);
  e.V = 4.0 * M_w * P_0 / (rho_w * PmcConst::univ_gas_const * e.T);
  e.W = PmcConst::water_latent_heat * M_w / (PmcConst::univ_gas_const * e.T);
  e.X = 4.0 * M_w * PmcConst::water_surf_eng / (PmcConst::univ_gas_const * e.T * rho_w);
  e.Y = 2.0 * e.k_a / (PmcConst::accom_coeff * rho_air * PmcConst::air_spec_heat) *
        sqrt(2.0 * PmcConst::pi * PmcConst::air_molec_weight / (PmcConst::univ_gas_const * e.T));
  e.Z = 2.0 * e.D_v / PmcConst::accom_coeff * sqrt(2.0 * PmcConst::pi * M_w / (PmcConst::univ_gas_const * e.T));
  e.UV = e.U * e.V;
  e.vc = (e.T * p0) / (T0 * e.p);
  e.g_coef = -2.0 * PmcConst::pi / (e.V * e.vc);
  e.invU = 1.0 / e.U;
  e.pdot_div_p = pdot / e.p;
  cond_env_dt(e);
  return e;
}

// condense_rates (condense.F90:371-502) for one particle.  WITH_JAC also returns
// a = dDdot/dD, b = dDdot/dH, c = dHdot_i/dD, dH = dHdot_i/dH.
template <bool WITH_JAC>
__device__ __forceinline__ void cond_rates(const CondEnv& e, double H, double D, double D_dry, double kappa,
                                           double num_conc, double& Ddot, double& Hdot_i, double& a, double& b,
                                           double& c, double& dH) {
  const double hfac = -2.0 * PmcConst::pi / (e.V * (e.vc / num_conc));
  if (D <= D_dry) {
    const double k_ap = e.k_a / (1.0 + e.Y / D_dry);
    const double D_vp = e.D_v / (1.0 + e.Z / D_dry);
    const double delta_star = e.UV * D_vp * H / k_ap;
    Ddot = k_ap * delta_star / (e.U * D_dry);
    Hdot_i = hfac * (D_dry * D_dry) * Ddot;
    if (WITH_JAC) {
      const double ddeltastar_dH = e.UV * D_vp / k_ap;
      a = 0.0;
      b = k_ap / (e.U * D_dry) * ddeltastar_dH;
      c = 0.0;
      dH = hfac * (D_dry * D_dry) * b;
    }
    return;
  }
  const double D2 = D * D, D3 = D2 * D, Dd3 = D_dry * D_dry * D_dry;
  const double k_ap = e.k_a / (1.0 + e.Y / D);
  const double D_vp = e.D_v / (1.0 + e.Z / D);
  const double den = D3 + (kappa - 1.0) * Dd3;
  const double a_w = (D3 - Dd3) / den;
  const double XD = e.X / D;
  const double UVD = e.UV * D_vp;
  double delta_star = 0.0, h = 0.0, dh_ddelta = 1.0, ex = 0.0, q = 1.0;
#pragma unroll
  for (int newton_step = 1; newton_step <= 5; newton_step++) {
    delta_star = delta_star - h / dh_ddelta;
    q = 1.0 / (1.0 + delta_star);
    ex = exp((e.W * delta_star + XD) * q);
    h = k_ap * delta_star - UVD * (H - a_w * q * ex);
    dh_ddelta = k_ap - UVD * a_w * (q * q) * (1.0 - e.W * q + XD * q) * ex;
  }
  Ddot = k_ap * delta_star / (e.U * D);
  Hdot_i = hfac * D2 * Ddot;
  if (WITH_JAC) {
    const double dkap_dD = e.k_a * e.Y / ((D + e.Y) * (D + e.Y));
    const double dDvp_dD = e.D_v * e.Z / ((D + e.Z) * (D + e.Z));
    const double daw_dD = 3.0 * D2 * kappa * Dd3 / (den * den);
    const double dh_dD = dkap_dD * delta_star - e.UV * dDvp_dD * H +
                         e.UV * (a_w * dDvp_dD + D_vp * daw_dD - D_vp * a_w * (e.X / D2) * q) * q * ex;
    const double dh_dH = -UVD;
    const double ddeltastar_dD = -dh_dD / dh_ddelta;
    const double ddeltastar_dH = -dh_dH / dh_ddelta;
    a = dkap_dD * delta_star / (e.U * D) + k_ap * ddeltastar_dD / (e.U * D) - k_ap * delta_star / (e.U * D2);
    b = k_ap / (e.U * D) * ddeltastar_dH;
    c = hfac * (2.0 * D * Ddot + D2 * a);
    dH = hfac * D2 * b;
  }
}

struct CondParams {
  double del_t, reltol, abstol_scale, abstol_H;
  int max_steps;
  int debug_cell;  // -1: off; else the cell whose step sequence thread 0 prints (PMCGPU_DEBUG_CELL)
};

// RODAS4 (Hairer & Wanner, Solving ODEs II, sect. IV.7): 6 stages, order 4(3), stiffly
// accurate and L-stable, in the (A, C, M, E, alpha, gamma_i) form
//   (1/(h gamma) - J) K_i = f(t + alpha_i h, y + sum_j A_ij K_j) + sum_j (C_ij / h) K_j + gamma_i h df/dt
#define ROS_S 6
__constant__ double c_rosA[ROS_S][ROS_S] = {
    {0, 0, 0, 0, 0, 0},
    {0.1544000000000000e+01, 0, 0, 0, 0, 0},
    {0.9466785280815826, 0.2557011698983284, 0, 0, 0, 0},
    {0.3314825187068521e+01, 0.2896124015972201e+01, 0.9986419139977817, 0, 0, 0},
    {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950, 0, 0},
    {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950, 1.0, 0}};
__constant__ double c_rosC[ROS_S][ROS_S] = {
    {0, 0, 0, 0, 0, 0},
    {-0.5668800000000000e+01, 0, 0, 0, 0, 0},
    {-0.2430093356833875e+01, -0.2063599157091915, 0, 0, 0, 0},
    {-0.1073529058151375, -0.9594562251023355e+01, -0.2047028614809616e+02, 0, 0, 0},
    {0.7496443313967647e+01, -0.1024680431464352e+02, -0.3399990352819905e+02, 0.1170890893206160e+02, 0, 0},
    {0.8083246795921522e+01, -0.7981132988064893e+01, -0.3152159432874371e+02, 0.1631930543123136e+02,
     -0.6058818238834054e+01, 0}};
__constant__ double c_rosM[ROS_S] = {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02,
                                     -0.6878860361058950, 1.0, 1.0};
__constant__ double c_rosAlpha[ROS_S] = {0.0, 0.386, 0.210, 0.630, 1.0, 1.0};
__constant__ double c_rosGam[ROS_S] = {0.25, -0.1043, 0.1035, -0.3620000000000023e-01, 0.0, 0.0};
#define ROS_GAMMA 0.25
#define ROS_ELO 4.0

// ---- branch-free FP64 primitives for the rate evaluation.  The library versions of 1/x and exp
// carry special-case branches and immediate-operand constants that showed up as 16 % UMOV and 10 %
// BRA/BSSY/BSYNC of all executed instructions (profiles/r01_s9_condense_ncu.md); the arguments here
// are always finite, normal and of moderate size, so the fast paths alone are enough.
__constant__ double c_expc[14] = {1.0, 1.0, 1.0 / 2, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040,
                                  1.0 / 40320, 1.0 / 362880, 1.0 / 3628800, 1.0 / 39916800, 1.0 / 479001600,
                                  1.0 / 6227020800.0};
// 1/x for normal x: MUFU.RCP64H seed (about 20 bits), one cubic and one quadratic correction (< 1 ulp)
__device__ __forceinline__ double fast_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  e = fma(e, e, e);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}
// exp(x) for |x| < 700: x = n ln2 + r, |r| <= ln2 / 2, degree-13 Taylor polynomial (remainder 4e-18),
// 2^n by an add into the exponent field
__device__ __forceinline__ double fast_exp(double x) {
  const double SHIFT = 6755399441055744.0;  // 1.5 * 2^52
  const double kd = fma(x, 1.4426950408889634, SHIFT);
  const int n = __double2loint(kd);
  const double nf = kd - SHIFT;
  double r = fma(nf, -6.93147180369123816490e-01, x);
  r = fma(nf, -1.90821492927058770002e-10, r);
  double p = c_expc[13];
#pragma unroll
  for (int i = 12; i >= 0; i--) p = fma(p, r, c_expc[i]);
  return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}

// exp(x) with a 32-entry table of 2^(j/32) (staged in shared memory by the caller): x = (32 n + j) ln2/32 + r,
// |r| <= ln2/64, exp(r) - 1 by a degree-6 polynomial (remainder 3e-18): 11 FP64 instructions and a short
// dependency chain instead of 17 in a row
__constant__ double c_exp2tab[32] = {
    1.0, 1.0218971486541166, 1.0442737824274138, 1.0671404006768237,
    1.0905077326652577, 1.1143867425958924, 1.1387886347566916, 1.1637248587775775,
    1.189207115002721, 1.215247359980469, 1.241857812073484, 1.2690509571917332,
    1.2968395546510096, 1.3252366431597413, 1.3542555469368927, 1.383909881963832,
    1.4142135623730951, 1.4451808069770467, 1.4768261459394993, 1.5091644275934228,
    1.5422108254079407, 1.5759808451078865, 1.6104903319492543, 1.645755478153965,
    1.681792830507429, 1.718619298122478, 1.7562521603732995, 1.7947090750031072,
    1.8340080864093424, 1.8741676341103, 1.9152065613971474, 1.9571441241754002};
__device__ __forceinline__ double fast_exp_t(double x, const double* __restrict__ tab) {
  const double SHIFT = 6755399441055744.0;  // 1.5 * 2^52
  const double kd = fma(x, 46.16624130844683, SHIFT);
  const int ki = __double2loint(kd);
  const double nf = kd - SHIFT;
  double r = fma(nf, -0.02166084938653512, x);
  r = fma(nf, -5.9631716539705866e-12, r);
  const double t = tab[ki & 31];
  const double r2 = r * r;
  double p = fma(r, 1.0 / 720, 1.0 / 120);
  p = fma(p, r, 1.0 / 24);
  p = fma(p, r, 1.0 / 6);
  p = fma(p, r, 0.5);
  const double q = fma(r2, p, r);  // exp(r) - 1
  const double y = fma(t, q, t);
  return __hiloint2double(__double2hiint(y) + ((ki >> 5) << 20), __double2loint(y));
}

// deterministic block reduction of NV values (sum or max) with one barrier: the per-warp partials go to
// one of two alternating buffers, so the next reduction cannot overwrite values still being read
template <int NV, int NT, bool IS_MAX>
__device__ __forceinline__ void block_red(double (&x)[NV], double* s_red /*[2][4][NT/32]*/, int& par) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* buf = s_red + par * (4 * (NT / 32));
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double t = x[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double u = __shfl_down_sync(0xffffffffu, t, o);
      t = IS_MAX ? fmax(t, u) : t + u;
    }
    if (lane == 0) buf[k * (NT / 32) + warp] = t;
  }
  __syncthreads();
  // the NT/32 partials of every value are added pairwise (depth log2 instead of a serial chain of NT/32 adds
  // that every thread would wait for)
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double pw[NT / 32];
#pragma unroll
    for (int w = 0; w < NT / 32; w++) pw[w] = buf[k * (NT / 32) + w];
#pragma unroll
    for (int sd = 1; sd < NT / 32; sd *= 2)
#pragma unroll
      for (int i = 0; i + sd < NT / 32; i += 2 * sd) pw[i] = IS_MAX ? fmax(pw[i], pw[i + sd]) : pw[i] + pw[i + sd];
    x[k] = pw[0];
  }
  par ^= 1;
}

// cond_env with the branch-free primitives (same terms, condense.F90:391-416)
__device__ __forceinline__ CondEnv cond_env_fast(double T0, double p0, double Tdot, double pdot, double t) {
  CondEnv e;
  e.T = T0 + t * Tdot;
  e.p = p0 + t * pdot;
  e.Tdot = Tdot;
  e.pdot = pdot;
  const double R = PmcConst::univ_gas_const, rho_w = PmcConst::water_density, M_w = PmcConst::water_molec_weight;
  const double LN10 = 2.302585092994045684;
  const double invT = fast_rcp(e.T), invp = fast_rcp(e.p), invT38 = fast_rcp(e.T - 38.0);
  const double P_0 = PmcConst::water_eq_vap_press * fast_exp(LN10 * 7.45 * (e.T - PmcConst::water_freeze_temp) * invT38);
  e.dP0_dT_div_P0 = 7.45 * LN10 * (PmcConst::water_freeze_temp - 38.0) * (invT38 * invT38);
  const double inv_rho_air = R * e.T * invp / PmcConst::air_molec_weight;
  e.k_a = 1e-3 * (4.39 + 0.071 * e.T);
  e.D_v = 0.211e-4 * PmcConst::air_std_press * invp * fast_exp(1.94 * log(e.T / 273.0));
  e.U = PmcConst::water_latent_heat * rho_w * 0.25 * invT;
  e.V = 4.0 * M_w * P_0 * invT / (rho_w * R);
  e.W = PmcConst::water_latent_heat * M_w * invT / R;
  e.X = 4.0 * M_w * PmcConst::water_surf_eng * invT / (R * rho_w);
  const double rsT = sqrt(invT);
  e.Y = 2.0 * e.k_a * inv_rho_air / (PmcConst::accom_coeff * PmcConst::air_spec_heat) *
        (sqrt(2.0 * PmcConst::pi * PmcConst::air_molec_weight / R) * rsT);
  e.Z = 2.0 * e.D_v / PmcConst::accom_coeff * (sqrt(2.0 * PmcConst::pi * M_w / R) * rsT);
  e.UV = e.U * e.V;
  e.vc = (e.T * p0) * invp / T0;
  e.g_coef = -2.0 * PmcConst::pi / (e.V * e.vc);
  e.invU = 1.0 / e.U;
  e.pdot_div_p = pdot / e.p;
  cond_env_dt(e);
  return e;
}

// condense_rates (condense.F90:371-502) for PPT particles of one thread at once.  The PPT
// evaluations are independent dependency chains that the unrolled inner loops interleave, which is
// what keeps the FP64 pipe busy with few warps per SM.  Per particle one reciprocal yields
// 1/D, 1/(D+Y), 1/(D+Z), 1/den.
// The reference finds delta* by five fixed Newton steps from zero on
//   h(delta) = k_a' delta - U V D_v' (H - a_w exp((W delta + X/D)/(1+delta)) / (1+delta));
// its iterates reach rounding level after the third update (profiles/r01_newton_convergence.md).  Here
// the same root is found in the variable u = delta / (1 + delta), which removes the division from the
// iteration: (1 - u) h = k_a' u - (1 - u) (B - g), g = A (1 - u) exp(X/D + (W - X/D) u), A = U V D_v' a_w,
// B = U V D_v' H.  The iteration stops when every lane's update is below 1e-7 |u| (the next error is
// then below 1e-16 |u|), after at most five evaluations like the reference.
// Per-particle constants (D_dry, kappa, num_conc) are read from shared memory (stride COND_THREADS).
// WITH_JAC also returns a = dDdot/dD, b = dDdot/dH, c = dHdot_i/dD, dH = dHdot_i/dH.
// Newton iteration on u = delta* / (1 + delta*) (condense.F90:449-466 does five fixed updates from 0).  Newton's
// error recursion is e' = K e^2 with K = h'' / (2 h'); here h'' ~ 280 A and h' >= 17 A (W = L_v M_w / (R T) ~ 19
// enters as exp(W u)), so K < 10 in absolute terms (profiles/r01_newton_convergence.md shows the resulting
// relative constants 1e-3 .. 2e-2).  An update of size du therefore leaves an error below K du^2, and the loop
// stops when NEWTON_K du^2 <= NEWTON_TOL max(|u|, NEWTON_FLOOR) with a tenfold margin on K: the root is then
// exact to 1e-14 relative (rounding level; activation amplifies differences by 1e4 or more).  The floor is there because particles in equilibrium have delta* -> 0, where a
// purely relative test would never pass (their rounding noise is ~1e-16 absolute).
#define NEWTON_K 100.0
#define NEWTON_TOL 1e-14
#define NEWTON_FLOOR 1e-6
// MODE (where the iteration starts; sw = shared arrays u0, du/dD, du/dH, du/dt, uN of PPT * NT doubles each):
//   0  cold, from 0 (the BDF kernel)
//   1  f at (t, y) with the Jacobian: starts from uN (root at the end of the previous step); leaves u0 and
//      the sensitivities du/dD, du/dH of the root (from d delta* / dD, d delta* / dH, which the Jacobian needs anyway)
//      and du/dt from the analytic time derivative
//   3  stage values: starts from the first-order prediction u0 + du/dD (y - D) + du/dH (Hs - H) + du/dt dts
//      (the error of the start is second order in the stage increment, so one update usually suffices);
//      write_uN: the last stage (state at t + h) leaves its root in uN
//   5  plain warm start from / to sw[0] (8 particles per thread: no room for the sensitivities)
template <int PPT, bool WITH_JAC, int NT = COND_THREADS, int MODE = 0>
__device__ __forceinline__ void cond_rates_vec(const double* __restrict__ etab, const CondEnv& e, double H,
                                               const double (&D)[PPT], const double* __restrict__ sDdry, const double* __restrict__ sKap,
                                               const double* __restrict__ sNc, double (&Ddot)[PPT],
                                               double (&Hd)[PPT], double (&a)[PPT], double (&b)[PPT],
                                               double (&c)[PPT], double (&dH)[PPT], double* sw = nullptr,
                                               const double* D0 = nullptr, double H0 = 0.0, double dts = 0.0,
                                               bool write_uN = false, double* ft = nullptr, double* hdt = nullptr) {
  constexpr int AS = PPT * NT;
  const double g = e.g_coef;  // Hdot_i = g * num_conc * D^2 * Ddot
  const double invU = e.invU;
  double k_ap[PPT], UVD[PPT], a_w[PPT], XD[PPT], invD[PPT];
  double inv_dY[PPT], inv_dZ[PPT], inv_den[PPT];
  double u[PPT], A[PPT], kB[PPT], cW[PPT];
#pragma unroll
  for (int k = 0; k < PPT; k++) {
    const double d = D[k], dd = sDdry[k * NT], kp = sKap[k * NT];
    const double D3 = d * d * d, Dd3 = dd * dd * dd;
    const double dY = d + e.Y, dZ = d + e.Z, den = fma(kp - 1.0, Dd3, D3);
    const double pYZ = dY * dZ, pDn = d * den;
    const double r = fast_rcp(pYZ * pDn);
    invD[k] = r * pYZ * den;
    inv_dY[k] = r * dZ * pDn;
    inv_dZ[k] = r * dY * pDn;
    inv_den[k] = r * pYZ * d;
    k_ap[k] = e.k_a * d * inv_dY[k];
    UVD[k] = e.UV * (e.D_v * d * inv_dZ[k]);
    a_w[k] = (D3 - Dd3) * inv_den[k];
    XD[k] = e.X * invD[k];
    A[k] = UVD[k] * a_w[k];
    kB[k] = fma(UVD[k], H, k_ap[k]);
    cW[k] = e.W - XD[k];
    u[k] = 0.0;
    if (MODE == 1 || MODE == 5) {
      const double up = sw[(MODE == 1 ? 4 : 0) * AS + k * NT];
      if (fabs(up) < 0.5) u[k] = up;
    }
    if (MODE == 3) {
      double up = fma(sw[1 * AS + k * NT], d - D0[k], sw[0 * AS + k * NT]);
      up = fma(sw[2 * AS + k * NT], H - H0, up);
      up = fma(sw[3 * AS + k * NT], dts, up);
      if (fabs(up) < 0.5) u[k] = up;
    }
  }
  // The loop body is written operation by operation across the PPT particles of the thread (not particle
  // by particle): the FP64 pipe needs several independent instructions between two dependent ones, and the
  // instruction scheduler keeps this order, while it serialises PPT long per-particle chains.
#pragma unroll 1
  for (int newton_step = 1; newton_step <= 5; newton_step++) {
    bool small = true;
    double s_[PPT], ex[PPT], gg[PPT], dh[PPT], hh[PPT], rc[PPT];
    {
      const double SHIFT = 6755399441055744.0;  // fast_exp_t, see above
      double x[PPT], kd[PPT], nf[PPT], r[PPT], pl[PPT], tb[PPT];
      int ki[PPT];
#pragma unroll
      for (int k = 0; k < PPT; k++) x[k] = fma(cW[k], u[k], XD[k]);
#pragma unroll
      for (int k = 0; k < PPT; k++) kd[k] = fma(x[k], 46.16624130844683, SHIFT);
#pragma unroll
      for (int k = 0; k < PPT; k++) {
        ki[k] = __double2loint(kd[k]);
        nf[k] = kd[k] - SHIFT;
      }
#pragma unroll
      for (int k = 0; k < PPT; k++) r[k] = fma(nf[k], -0.02166084938653512, x[k]);
#pragma unroll
      for (int k = 0; k < PPT; k++) tb[k] = etab[ki[k] & 31];
#pragma unroll
      for (int k = 0; k < PPT; k++) r[k] = fma(nf[k], -5.9631716539705866e-12, r[k]);
#pragma unroll
      for (int k = 0; k < PPT; k++) pl[k] = fma(r[k], 1.0 / 720, 1.0 / 120);
#pragma unroll
      for (int k = 0; k < PPT; k++) s_[k] = 1.0 - u[k];
#pragma unroll
      for (int k = 0; k < PPT; k++) pl[k] = fma(pl[k], r[k], 1.0 / 24);
#pragma unroll
      for (int k = 0; k < PPT; k++) x[k] = r[k] * r[k];
#pragma unroll
      for (int k = 0; k < PPT; k++) pl[k] = fma(pl[k], r[k], 1.0 / 6);
#pragma unroll
      for (int k = 0; k < PPT; k++) gg[k] = A[k] * s_[k];
#pragma unroll
      for (int k = 0; k < PPT; k++) pl[k] = fma(pl[k], r[k], 0.5);
#pragma unroll
      for (int k = 0; k < PPT; k++) dh[k] = fma(s_[k], cW[k], -2.0);
#pragma unroll
      for (int k = 0; k < PPT; k++) pl[k] = fma(x[k], pl[k], r[k]);  // exp(r) - 1
#pragma unroll
      for (int k = 0; k < PPT; k++) pl[k] = fma(tb[k], pl[k], tb[k]);
#pragma unroll
      for (int k = 0; k < PPT; k++)
        ex[k] = __hiloint2double(__double2hiint(pl[k]) + ((ki[k] >> 5) << 20), __double2loint(pl[k]));
    }
#pragma unroll
    for (int k = 0; k < PPT; k++) gg[k] = gg[k] * ex[k];                    // A (1 - u) exp(..)
#pragma unroll
    for (int k = 0; k < PPT; k++) dh[k] = fma(gg[k], dh[k], kB[k]);
#pragma unroll
    for (int k = 0; k < PPT; k++) hh[k] = fma(UVD[k], H, -gg[k]);           // B - g
    // fast_rcp(dh), see above
#pragma unroll
    for (int k = 0; k < PPT; k++) asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(rc[k]) : "d"(dh[k]));
#pragma unroll
    for (int k = 0; k < PPT; k++) hh[k] = fma(k_ap[k], u[k], -s_[k] * hh[k]);  // (1 - u) h
    {
      double e1[PPT];
#pragma unroll
      for (int k = 0; k < PPT; k++) e1[k] = fma(-dh[k], rc[k], 1.0);
#pragma unroll
      for (int k = 0; k < PPT; k++) e1[k] = fma(e1[k], e1[k], e1[k]);
#pragma unroll
      for (int k = 0; k < PPT; k++) rc[k] = fma(rc[k], e1[k], rc[k]);
#pragma unroll
      for (int k = 0; k < PPT; k++) e1[k] = fma(-dh[k], rc[k], 1.0);
#pragma unroll
      for (int k = 0; k < PPT; k++) rc[k] = fma(rc[k], e1[k], rc[k]);
    }
#pragma unroll
    for (int k = 0; k < PPT; k++) {
      const double du = hh[k] * rc[k];
      // NEWTON_K du^2 <= NEWTON_TOL max(|u|, NEWTON_FLOOR), written as two comparisons (no FP64 max)
      const double du2 = du * du;
      small = small && ((du2 <= (NEWTON_TOL / NEWTON_K) * fabs(u[k])) || (du2 <= (NEWTON_TOL / NEWTON_K) * NEWTON_FLOOR));
      u[k] -= du;
    }
    if (__all_sync(0xffffffffu, small)) break;
  }
  if (MODE == 1 || MODE == 5) {
#pragma unroll
    for (int k = 0; k < PPT; k++) sw[k * NT] = u[k];
  }
  if (MODE == 3 && write_uN) {
#pragma unroll
    for (int k = 0; k < PPT; k++) sw[4 * AS + k * NT] = u[k];
  }
#pragma unroll
  for (int k = 0; k < PPT; k++) {
    const double d = D[k];
    const double hf = g * sNc[k * NT];
    const double kU = k_ap[k] * invD[k] * invU;  // k_ap / (U D)
    const double q = 1.0 - u[k];                 // 1 / (1 + delta*)
    const double delta = u[k] * fast_rcp(q);
    Ddot[k] = kU * delta;
    Hd[k] = hf * (d * d) * Ddot[k];
    if (WITH_JAC) {
      const double dd = sDdry[k * NT], kp = sKap[k * NT];
      const double Dd3 = dd * dd * dd;
      const double ex = fast_exp_t(fma(cW[k], u[k], XD[k]), etab);  // exp((W delta + X/D) / (1 + delta))
      const double dhd = k_ap[k] - A[k] * (q * q) * (1.0 - e.W * q + XD[k] * q) * ex;  // dh/ddelta (:462)
      const double D_vp = e.D_v * d * inv_dZ[k];
      const double dkap_dD = e.k_a * e.Y * (inv_dY[k] * inv_dY[k]);
      const double dDvp_dD = e.D_v * e.Z * (inv_dZ[k] * inv_dZ[k]);
      const double daw_dD = 3.0 * (d * d) * kp * Dd3 * (inv_den[k] * inv_den[k]);
      const double dh_dD = dkap_dD * delta - e.UV * dDvp_dD * H +
                           e.UV * (a_w[k] * dDvp_dD + D_vp * daw_dD - D_vp * a_w[k] * (XD[k] * invD[k]) * q) * q * ex;
      const double inv_dh = fast_rcp(dhd);
      const double ddeltastar_dD = -dh_dD * inv_dh;
      const double ddeltastar_dH = UVD[k] * inv_dh;
      // analytic df/dt at fixed (D, H): d delta*/dt = -(dh/dt) / (dh/ddelta), see cond_env_dt
      if (ft != nullptr) {  // (the BDF kernel has no use for df/dt)
        const double rk = e.lka - e.Y_t * inv_dY[k];     // d ln k_a' / dt
        const double ru = e.luvd - e.Z_t * inv_dZ[k];    // d ln(U V D_v') / dt
        const double qE = q * ex;
        const double h_t = k_ap[k] * rk * delta - UVD[k] * ru * (H - a_w[k] * qE) +
                           A[k] * qE * (q * fma(e.W_t, delta, e.X_t * invD[k]));
        const double delta_t = -h_t * inv_dh;
        const double Ddot_t = fma(Ddot[k], rk - e.lU, kU * delta_t);
        ft[k] = Ddot_t;
        hdt[k] = hf * (d * d) * fma(Ddot[k], e.lG, Ddot_t);
        if (MODE == 1) sw[3 * AS + k * NT] = (q * q) * delta_t;
      }
      if (MODE == 1) {  // du = q^2 d delta
        sw[1 * AS + k * NT] = (q * q) * ddeltastar_dD;
        sw[2 * AS + k * NT] = (q * q) * ddeltastar_dH;
      }
      a[k] = dkap_dD * delta * invD[k] * invU + kU * ddeltastar_dD - kU * delta * invD[k];
      b[k] = kU * ddeltastar_dH;
      c[k] = hf * (2.0 * d * Ddot[k] + (d * d) * a[k]);
      dH[k] = hf * (d * d) * b[k];
    }
  }
  // dry particles (D <= D_dry, condense.F90:418-436): rare; the whole block sits behind a warp-uniform branch so
  // that its divisions are not issued (predicated off) in every evaluation
  bool any_dry = false;
#pragma unroll
  for (int k = 0; k < PPT; k++) any_dry = any_dry || (D[k] <= sDdry[k * NT]);
  if (__any_sync(0xffffffffu, any_dry)) {
#pragma unroll
  for (int k = 0; k < PPT; k++) {
    const double dd = sDdry[k * NT];
    if (D[k] <= dd) {
      const double hf = g * sNc[k * NT];
      const double kd = e.k_a / (1.0 + e.Y / dd);
      const double Dvd = e.D_v / (1.0 + e.Z / dd);
      const double delta_star = e.UV * Dvd * H / kd;
      Ddot[k] = kd * delta_star / (e.U * dd);
      Hd[k] = hf * (dd * dd) * Ddot[k];
      if (WITH_JAC) {
        a[k] = 0.0;
        b[k] = kd / (e.U * dd) * (e.UV * Dvd / kd);
        c[k] = 0.0;
        dH[k] = hf * (dd * dd) * b[k];
        if (ft != nullptr) {
          const double Ddot_t = Ddot[k] * ((e.luvd - e.lU) - e.Z_t / (dd + e.Z));
          ft[k] = Ddot_t;
          hdt[k] = hf * (dd * dd) * fma(Ddot[k], e.lG, Ddot_t);
        }
      }
    }
  }
  }
}

// shared arrays per particle: 6 stage vectors, abstol, D_dry, kappa, num_conc, then the Newton start values
// (u0, du/dD, du/dH, du/dt, uN with the predictor; one array without)
#define COND_PRED(PPT) ((PPT) <= 4)
#define COND_NARR(PPT) (ROS_S + 4 + (COND_PRED(PPT) ? 5 : 1))
#define ROS_NENV 6             // (T, p) terms of one step: t, then t + alpha_i h of the stages i = 2..6
#define ROS_HIN_FACTOR 4.0     // first step = 4 x CVODE's first-order estimate (a 4th-order method), rounded down to 2^k
// ---- cells larger than one CTA holds (single-box runs, 1e4..1e5 particles): the same integrator with the
// particles of ONE cell spread over all CTAs of a co-resident (cooperatively launched) grid.  Every cell-wide
// sum becomes: block reduction -> one partial per CTA in global memory -> grid barrier -> every CTA adds the
// partials in the same fixed order (so all CTAs hold bit-identical sums and take the same branches).
#define COND_GRID_MAXB 1024
struct GridRed {
  double* buf;     // [2][4][COND_GRID_MAXB] partials, two alternating sets
  unsigned* bar;   // arrival counter (zeroed before the launch)
  unsigned gen;    // barriers passed so far
  int par;
};
__device__ __forceinline__ void grid_barrier(GridRed& g) {
  g.gen++;
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(g.bar, 1u);
    const unsigned target = g.gen * gridDim.x;
    unsigned seen;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(g.bar) : "memory");
    } while (seen < target);
  }
  __syncthreads();
}
template <int NV, int NT, bool IS_MAX, bool GRID>
__device__ __forceinline__ void cell_red(double (&x)[NV], double* s_red, int& par, GridRed& g, double* s_g /*[4]*/) {
  block_red<NV, NT, IS_MAX>(x, s_red, par);
  if (GRID) {
    double* gb = g.buf + (size_t)g.par * 4 * COND_GRID_MAXB;
    if (threadIdx.x == 0) {
#pragma unroll
      for (int k = 0; k < NV; k++) __stcg(&gb[k * COND_GRID_MAXB + blockIdx.x], x[k]);
    }
    grid_barrier(g);
    if (threadIdx.x < 32) {
#pragma unroll
      for (int k = 0; k < NV; k++) {
        double t = IS_MAX ? -1e300 : 0.0;
        for (int bq = threadIdx.x; bq < (int)gridDim.x; bq += 32) {
          const double u = __ldcg(&gb[k * COND_GRID_MAXB + bq]);
          t = IS_MAX ? fmax(t, u) : t + u;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const double u = __shfl_xor_sync(0xffffffffu, t, o);
          t = IS_MAX ? fmax(t, u) : t + u;
        }
        if (threadIdx.x == 0) s_g[k] = t;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NV; k++) x[k] = s_g[k];
    g.par ^= 1;
  }
}

// One stage's linear combination of the earlier stage vectors, stage index ST a compile-time constant:
//   C_FORM false: y += sum_j A[ST][j] K_j,              H part: Hs += sum_j A[ST][j] KH[j]
//   C_FORM true : f += sum_j (C[ST][j] * inv_h) K_j,    H part: rH += sum_j (C[ST][j] * inv_h) KH[j]
template <int ST, int PPT, int NT, bool C_FORM>
__device__ __forceinline__ void ros_combo(const double* __restrict__ s_K, int tid, const double (&KH)[ROS_S], double inv_h,
                                          double (&y)[PPT], double& yH) {
#pragma unroll
  for (int j = 0; j < ST; j++) {
    const double cj = C_FORM ? c_rosC[ST][j] * inv_h : c_rosA[ST][j];
#pragma unroll
    for (int k = 0; k < PPT; k++) y[k] += cj * s_K[(j * PPT + k) * NT + tid];
    yH += cj * KH[j];
  }
}

template <int PPT, int NT, bool GRID = false>
__global__ void __launch_bounds__(NT, (NT == 160 || NT == 192) ? 2 : 1) k_condense(DevView v, const double* T0a /* may alias v.temp */,
                                                           const double* p0a,
                                                           const double* T1 /* may alias v.temp */,
                                                           const double* p1, CondParams prm,
                                                           int* __restrict__ stats, double* __restrict__ wre_out,
                                                           int n_limit_lo, int n_limit_hi, GridRed gr) {
  // GRID: all CTAs walk the cells together and share each big cell; otherwise one CTA per cell
  const int poff = GRID ? blockIdx.x * PPT : 0;
  const bool writer = !GRID || blockIdx.x == 0;
  __shared__ double s_g[4];
  if (GRID) {  // nothing to do unless some cell is larger than the per-CTA kernels handle (uniform over the grid)
    int any = 0;
    for (int c = threadIdx.x; c < v.n_cells; c += NT) any |= (v.count[c] > n_limit_lo);
    if (!__syncthreads_or(any)) return;
  }
  for (int cell = GRID ? 0 : blockIdx.x; cell < (GRID ? v.n_cells : (int)blockIdx.x + 1); cell++) {
  const int n = v.count[cell];
  if (n <= n_limit_lo || n > n_limit_hi) continue;  // another instantiation owns this cell
  if (v.status[cell] != 0 || cell_is_boundary(v, cell)) continue;
  const int tid = threadIdx.x;
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  __shared__ double s_red[2 * 4 * (NT / 32)];
  __shared__ CondEnv s_env[ROS_NENV];
  extern __shared__ double s_K[];  // [COND_NARR(PPT)][PPT][NT]
  constexpr bool PRED = COND_PRED(PPT);
#define SK(st, k) s_K[((st) * PPT + (k)) * NT + tid]
#define S_ATOL(k) SK(ROS_S, k)
  const double* sDdry = &SK(ROS_S + 1, 0);
  const double* sKap = &SK(ROS_S + 2, 0);
  const double* sNc = &SK(ROS_S + 3, 0);
  double* sUw = &SK(ROS_S + 4, 0);
  int par = 0;
  __shared__ double s_etab[32];
  if (tid < 32) s_etab[tid] = c_exp2tab[tid];
  __syncthreads();
  // aero_state_reweight (aero_state.F90:597-653) after the solve is the identity for flat weights
  // (ratio 1, prob_round(1) = 1).  For size-dependent weights this kernel leaves each particle's
  // number concentration before the solve (aero_state_num_conc_for_reweight :564-587) in v.lnr and
  // its new radius in v.gct_alt; k_reweight (rebalance.cu) then duplicates / removes particles.
  bool nonflat = false;
  for (int q = 0; q < v.C * v.G; q++)
    if (v.wtype[(size_t)cell * v.C * v.G + q] != PMCGPU_WEIGHT_NONE) nonflat = true;
  const int iw = v.i_water - 1;
  const double T0 = T0a[cell], p0 = p0a[cell];  // env_state_initial
  const double del_t = prm.del_t;
  const double Tdot = (T1[cell] - T0) / del_t, pdot = (p1[cell] - p0) / del_t;
  const double H0 = v.rh[cell];
  double H = H0;

  // ---- set-up (condense.F90:196-237): kappa, D_dry, num_conc, D, abstol per particle.
  // Slots past the population hold a harmless wet dummy particle with number concentration 0: its Hdot_i and
  // every Jacobian entry that couples it to H vanish, so it enters no cell-wide sum and needs no masking there.
  double D[PPT];
  double wsum = 0.0;
  {
    // species rows: the loads of all PPT particles and of four species at a time are in flight together
    // (one serial chain of S dependent DRAM round trips per particle before); sums in species order as before
    double vt[PPT], vs[PPT], ks[PPT], wv[PPT];
#pragma unroll
    for (int k = 0; k < PPT; k++) { vt[k] = 0.0; vs[k] = 0.0; ks[k] = 0.0; wv[k] = 0.0; }
    const double M_w = v.molw[iw], rho_w = v.density[iw];
#pragma unroll 4
    for (int s = 0; s < v.S; s++) {
      // aero_particle_solute_kappa (aero_particle.F90:724-755)
      const double kp = (v.nions[s] > 0) ? M_w * v.density[s] / (v.molw[s] * rho_w) * (double)v.nions[s] : v.kappa[s];
      double x[PPT];
#pragma unroll
      for (int k = 0; k < PPT; k++) {
        const int p = tid + (poff + k) * NT;
        x[k] = (p < n) ? v.vol[(size_t)s * NS + base + p] : 0.0;
      }
#pragma unroll
      for (int k = 0; k < PPT; k++) {
        vt[k] += x[k];
        if (s != iw) {
          vs[k] += x[k];
          ks[k] += x[k] * kp;
        } else {
          wv[k] = x[k];
        }
      }
    }
#pragma unroll
    for (int k = 0; k < PPT; k++) {
      const int p = tid + (poff + k) * NT;
      D[k] = 2.0;
      S_ATOL(k) = 1.0;
      SK(ROS_S + 1, k) = 1.0;
      SK(ROS_S + 2, k) = 0.5;
      SK(ROS_S + 3, k) = 0.0;
      SK(ROS_S + 4, k) = 0.0;
      if (PRED) {
        SK(ROS_S + 5, k) = 0.0;
        SK(ROS_S + 6, k) = 0.0;
        SK(ROS_S + 7, k) = 0.0;
        SK(ROS_S + 8, k) = 0.0;
      }
      if (p < n) {
        const int gc = v.wgc[base + p];
        const double r = vol2rad(vt[k]);
        const double ddry = 2.0 * vol2rad(vs[k]);
        const double ncp = awa_nc_at_radius(v, cell, gc / v.G + 1, r);
        SK(ROS_S + 1, k) = ddry;
        SK(ROS_S + 2, k) = ks[k] / vs[k];
        SK(ROS_S + 3, k) = ncp;
        D[k] = 2.0 * r;
        S_ATOL(k) = fmax(1e-30, prm.abstol_scale * (D[k] - ddry));
        wsum += wv[k] * ncp;
        if (nonflat) {
          const size_t wi = (size_t)cell * v.C * v.G + gc;
          v.lnr[base + p] = weight_nc(v.wtype[wi], v.wmag[wi], v.wexp[wi], r);
        }
      }
    }
  }
  {
    double x[1] = {wsum};
    cell_red<1, NT, false, GRID>(x, s_red, par, gr, s_g);
    wsum = x[0];  // water_vol_conc_initial
  }

  const double FAC_MIN = 0.2, FAC_MAX = 6.0, FAC_SAFE = 0.95, FAC_REJ = 0.1;
  double u1[PPT], u2[PPT], u3[PPT], u4[PPT];  // unused outputs of the no-Jacobian evaluations
  const int warp = tid >> 5, lane = tid & 31;
  // the (T, p) terms of condense_rates at the ROS_NENV times of a step: warp w evaluates entry w
  auto make_envs = [&](double t, double h, int first, int last) {
    for (int w = first + warp; w <= last; w += NT / 32) {  // entry w by warp (w - first) mod (NT / 32)
      const double tt = (w == 0) ? t : t + c_rosAlpha[w] * h;
      const CondEnv e = cond_env_fast(T0, p0, Tdot, pdot, tt);
      if (lane == 0) s_env[w] = e;
    }
    __syncthreads();
  };
  auto env_H = [&](const CondEnv& e, double Hs) { return -e.dP0_dT_div_P0 * Tdot * Hs + Hs * e.pdot_div_p; };

  double h = del_t;  // replaced by the estimate below before the first step
  int n_evals = 0;
  double t = 0.0;
  double h_natural = h;
  int n_steps = 0, n_reject = 0, fail = 0;
  bool reject_last = false, reject_more = false;

  while (t < del_t && !fail) {
    if (n_steps + n_reject >= prm.max_steps) { fail = 6; break; }
    // After the first step h is known here, so the stage environments are built together with those of the
    // Jacobian evaluation (one entry per warp, one barrier) instead of in a second round.
    const bool first_step = (n_steps + n_reject == 0);
    bool truncated = false;
    if (!first_step && t + h > del_t * (1.0 - 1e-14)) { h_natural = h; h = del_t - t; truncated = true; }
    make_envs(t, h, 0, first_step ? 0 : ROS_NENV - 1);
    bool stage_envs_ready = !first_step;
    // f, the Jacobian and the analytic df/dt at (t, y): one rate evaluation
    double f0[PPT], dfdt[PPT], ja[PPT], jb[PPT], jc[PPT];
    double f0H, dfdtH, dd;
    {
      double hd[PPT], dh[PPT], hdt[PPT];
      cond_rates_vec<PPT, true, NT, PRED ? 1 : 5>(s_etab, s_env[0], H, D, sDdry, sKap, sNc, f0, hd, ja, jb, jc, dh, sUw,
                                                   nullptr, 0.0, 0.0, false, dfdt, hdt);
      n_evals += 1;
      double acc[3] = {0.0, 0.0, 0.0};
#pragma unroll
      for (int k = 0; k < PPT; k++) {
        const bool live = tid + (poff + k) * NT < n;
        if (!live) { f0[k] = 0.0; dfdt[k] = 0.0; ja[k] = 0.0; jb[k] = 0.0; jc[k] = 0.0; }
        acc[0] += hd[k];  // (exactly zero for the dummy slots: their number concentration is 0)
        acc[1] += dh[k];
        acc[2] += hdt[k];
      }
      cell_red<3, NT, false, GRID>(acc, s_red, par, gr, s_g);
      f0H = acc[0] + env_H(s_env[0], H);
      dd = acc[1] + (-s_env[0].dP0_dT_div_P0 * Tdot + s_env[0].pdot_div_p);
      dfdtH = acc[2] + s_env[0].envH_t * H;
    }
    if (first_step) {
      // first step size, after cvHin (cvode.c): h = 1/2 sqrt(2 / ||y''||), bounded by hub, times ROS_HIN_FACTOR
      // for a 4th-order method (small enough that the first step is never near rejection).  CVODE differences f along the solution; here y'' = J f + df/dt comes from the
      // Jacobian of this step, which carries no cancellation noise (a noisy first step would make two
      // implementations take different step sequences).
      double a1[2] = {0.0, 0.0};
      double ew[PPT];
#pragma unroll
      for (int k = 0; k < PPT; k++) {
        ew[k] = 1.0 / (prm.reltol * fabs(D[k]) + S_ATOL(k));
        if (tid + (poff + k) * NT < n) {
          a1[0] += jc[k] * f0[k];
          a1[1] = fmax(a1[1], fabs(f0[k]) / (0.1 * fabs(D[k]) + 1.0 / ew[k]));
        }
      }
      double a1s[1] = {a1[0]}, a1m[1] = {a1[1]};
      cell_red<1, NT, false, GRID>(a1s, s_red, par, gr, s_g);
      cell_red<1, NT, true, GRID>(a1m, s_red, par, gr, s_g);
      const double ewH = 1.0 / (prm.reltol * fabs(H) + prm.abstol_H);
      const double yddH = a1s[0] + dd * f0H + dfdtH;
      const double hub_inv = fmax(a1m[0], fabs(f0H) / (0.1 * fabs(H) + 1.0 / ewH));
      double a2[1] = {0.0};
#pragma unroll
      for (int k = 0; k < PPT; k++)
        if (tid + (poff + k) * NT < n) {
          const double w = (ja[k] * f0[k] + jb[k] * f0H + dfdt[k]) * ew[k];
          a2[0] += w * w;
        }
      cell_red<1, NT, false, GRID>(a2, s_red, par, gr, s_g);
      const double yddnrm = sqrt((a2[0] + (yddH * ewH) * (yddH * ewH)) / (double)(n + 1));
      double hub = 0.1 * del_t;
      if (hub * hub_inv > 1.0) hub = 1.0 / hub_inv;
      const double hnew = (yddnrm * hub * hub > 2.0) ? sqrt(2.0 / yddnrm) : hub;
      // rounded down to a power of two: the estimate inherits the cancellation noise of the rates of
      // particles in equilibrium (a few 1e-4 between two implementations); the power of two does not
      h = ROS_HIN_FACTOR * 0.5 * hnew;
      h = __longlong_as_double(__double_as_longlong(h) & 0x7FF0000000000000ll);
      h = fmin(del_t, h);
      h_natural = h;
      if (t + h > del_t * (1.0 - 1e-14)) { h_natural = h; h = del_t - t; truncated = true; }
    }
    for (;;) {
      if (!stage_envs_ready) make_envs(t, h, 1, ROS_NENV - 1);  // stage times of this attempt
      stage_envs_ready = false;
      const double eps = 1.0 / (ROS_GAMMA * h);
      const double inv_h = 1.0 / h;
      double inv[PPT];
      double KH[ROS_S];  // indexed by compile-time constants only (registers)
      // Stage 1 needs no rate evaluation (f at (t, y) is known), so its Schur sum travels in the same
      // reduction as the Schur complement of the attempt.
      double fs[PPT];
      {
        double r2[2] = {0.0, 0.0};
        const double hg = h * c_rosGam[0];
#pragma unroll
        for (int k = 0; k < PPT; k++) {
          inv[k] = fast_rcp(eps - ja[k]);
          fs[k] = f0[k] + hg * dfdt[k];
          r2[0] += jc[k] * jb[k] * inv[k];  // jc = 0 in the dummy slots
          r2[1] += jc[k] * fs[k] * inv[k];
        }
        cell_red<2, NT, false, GRID>(r2, s_red, par, gr, s_g);
        const double inv_lhs_H = 1.0 / (eps - dd - r2[0]);  // one division per attempt, not one per stage
        KH[0] = ((f0H + hg * dfdtH) + r2[1]) * inv_lhs_H;
#pragma unroll
        for (int k = 0; k < PPT; k++) SK(0, k) = (fs[k] + jb[k] * KH[0]) * inv[k];
#pragma unroll 1
        for (int st = 1; st < ROS_S; st++) {
          double hd[PPT], y[PPT];
          double acc[2] = {0.0, 0.0};
          double Hs = H;
#pragma unroll
          for (int k = 0; k < PPT; k++) y[k] = D[k];
          // y + sum_j A_ij K_j with the stage index as a compile-time constant: the K_j loads of a stage are
          // issued together and the coefficients are constant-bank operands
          switch (st) {
            case 1: ros_combo<1, PPT, NT, false>(s_K, tid, KH, 1.0, y, Hs); break;
            case 2: ros_combo<2, PPT, NT, false>(s_K, tid, KH, 1.0, y, Hs); break;
            case 3: ros_combo<3, PPT, NT, false>(s_K, tid, KH, 1.0, y, Hs); break;
            case 4: ros_combo<4, PPT, NT, false>(s_K, tid, KH, 1.0, y, Hs); break;
            default: ros_combo<5, PPT, NT, false>(s_K, tid, KH, 1.0, y, Hs); break;
          }
          cond_rates_vec<PPT, false, NT, PRED ? 3 : 5>(s_etab, s_env[st], Hs, y, sDdry, sKap, sNc, fs, hd, u1, u2, u3, u4, sUw,
                                                         D, H, c_rosAlpha[st] * h, st == ROS_S - 1);
          n_evals++;
          const double hgs = h * c_rosGam[st];
          double rH = 0.0;
#pragma unroll
          for (int k = 0; k < PPT; k++) fs[k] = fs[k] + hgs * dfdt[k];
          switch (st) {
            case 1: ros_combo<1, PPT, NT, true>(s_K, tid, KH, inv_h, fs, rH); break;
            case 2: ros_combo<2, PPT, NT, true>(s_K, tid, KH, inv_h, fs, rH); break;
            case 3: ros_combo<3, PPT, NT, true>(s_K, tid, KH, inv_h, fs, rH); break;
            case 4: ros_combo<4, PPT, NT, true>(s_K, tid, KH, inv_h, fs, rH); break;
            default: ros_combo<5, PPT, NT, true>(s_K, tid, KH, inv_h, fs, rH); break;
          }
#pragma unroll
          for (int k = 0; k < PPT; k++) {
            acc[0] += hd[k];                     // zero in the dummy slots (number concentration 0)
            acc[1] += jc[k] * fs[k] * inv[k];    // jc = 0 there
          }
          cell_red<2, NT, false, GRID>(acc, s_red, par, gr, s_g);
          const double fsH = acc[0] + env_H(s_env[st], Hs);
          const double kh = (((fsH + hgs * dfdtH) + rH) + acc[1]) * inv_lhs_H;
          switch (st) {
            case 1: KH[1] = kh; break;
            case 2: KH[2] = kh; break;
            case 3: KH[3] = kh; break;
            case 4: KH[4] = kh; break;
            default: KH[5] = kh; break;
          }
#pragma unroll
          for (int k = 0; k < PPT; k++) SK(st, k) = (fs[k] + jb[k] * kh) * inv[k];
        }
      }
      // new state and weighted RMS error norm (E = last stage for RODAS4)
      double esum[1] = {0.0};
      double Dn[PPT];
#pragma unroll
      for (int k = 0; k < PPT; k++) {
        const int p = tid + (poff + k) * NT;
        double yn = D[k];
#pragma unroll
        for (int j = 0; j < ROS_S; j++) yn += c_rosM[j] * SK(j, k);
        Dn[k] = yn;
        if (p < n) {
          const double er = SK(ROS_S - 1, k);
          const double ew = er * fast_rcp(fma(prm.reltol, fmax(fabs(D[k]), fabs(yn)), S_ATOL(k)));
          esum[0] = fma(ew, ew, esum[0]);
        }
      }
      cell_red<1, NT, false, GRID>(esum, s_red, par, gr, s_g);
      double Hn = H;
#pragma unroll
      for (int j = 0; j < ROS_S; j++) Hn += c_rosM[j] * KH[j];
      {
        const double er = KH[ROS_S - 1];
        const double w = prm.reltol * fmax(fabs(H), fabs(Hn)) + prm.abstol_H;
        esum[0] += (er / w) * (er / w);
      }
      double err = sqrt(esum[0] / (double)(n + 1));
      if (!isfinite(err)) err = 1e10;
      err = fmax(err, 1e-10);
      if (cell == prm.debug_cell && tid == 0 && writer) printf("dev t %.17g h %.17g err %.17g\n", t, h, err);
      const double fac = fmin(FAC_MAX, fmax(FAC_MIN, FAC_SAFE / sqrt(sqrt(err))));
      double hnew = h * fac;
      if (err <= 1.0) {
#pragma unroll
        for (int k = 0; k < PPT; k++)
          if (tid + (poff + k) * NT < n) D[k] = Dn[k];
        H = Hn;
        t += h;
        n_steps++;
        if (reject_last) hnew = fmin(hnew, h);
        reject_last = false;
        reject_more = false;
        h = truncated ? fmax(h_natural, hnew) : hnew;
        h_natural = h;
        break;
      }
      n_reject++;
      if (n_steps + n_reject >= prm.max_steps) { fail = 6; break; }
      if (reject_more) hnew = h * FAC_REJ;
      reject_more = reject_last;
      reject_last = true;
      h = hnew;
      truncated = false;
      if (!(h > 1e-13 * del_t)) { fail = 7; break; }
    }
  }
  // ---- write back (condense.F90:266-293): water volume, clamp at zero, water balance
  double wfin = 0.0;
#pragma unroll
  for (int k = 0; k < PPT; k++) {
    const int p = tid + (poff + k) * NT;
    if (p < n && !fail) {
      double vs = 0.0;  // aero_particle_solute_volume
      for (int s = 0; s < v.S; s++)
        if (s != iw) vs += v.vol[(size_t)s * NS + base + p];
      double wv = rad2vol(D[k] / 2.0) - vs;
      wv = fmax(0.0, wv);
      v.vol[(size_t)iw * NS + base + p] = wv;
      wfin += wv * sNc[k * NT];
      if (nonflat) v.gct_alt[base + p] = vol2rad(vs + wv);
    }
  }
  {
    double x[1] = {wfin};
    cell_red<1, NT, false, GRID>(x, s_red, par, gr, s_g);
    wfin = x[0];
  }
  if (tid == 0 && writer) {
    if (fail) {
      if (v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_CONDENSE;
    } else {
      const double Tf = T1[cell], pf = p1[cell];
      const double V_comp_ratio = Tf * p0 / (T0 * pf);
      const double rho_w = v.density[iw], M_w = v.molw[iw];
      const double P0i = PmcConst::water_eq_vap_press * pow(10.0, 7.45 * (T0 - PmcConst::water_freeze_temp) / (T0 - 38.0));
      const double P0f = PmcConst::water_eq_vap_press * pow(10.0, 7.45 * (Tf - PmcConst::water_freeze_temp) / (Tf - 38.0));
      const double vap_i = M_w / (PmcConst::univ_gas_const * T0) * P0i * H0 / rho_w;
      const double vap_f = M_w / (PmcConst::univ_gas_const * Tf) * P0f * H * V_comp_ratio / rho_w;
      wre_out[cell] = ((vap_f - vap_i) + (wfin - wsum)) / (vap_f + wfin);
      v.rh[cell] = H;
      v.temp[cell] = Tf;
      v.pres[cell] = pf;
    }
    stats[3 * cell + 0] = n_steps;
    stats[3 * cell + 1] = n_reject;
    stats[3 * cell + 2] = n_evals;  // rate evaluations per particle
    atomicAdd(&v.work[0], (unsigned long long)n_evals * (unsigned long long)n);
    atomicAdd(&v.work[1], (unsigned long long)n);
  }
    if (GRID) __syncthreads();
  }
#undef SK
#undef S_ATOL
}

// ===================================================================== BDF integrator
// condense_solver (condense_solver.c:84-181) hands the cell's system to SUNDIALS CVODE with CV_BDF.
// k_condense_bdf restates CVODE's published method (variable order 1..5, variable step, fixed-leading-
// coefficient Nordsieck form, CVODE's error test, step / order selection and initial step; DESIGN.md
// section 3 names the sources) with one CTA per cell:
//   * the Nordsieck vectors zn[0..5] of the particles live in shared memory (PPT particles per
//     thread), the H components and all step-control scalars are kept redundantly by every thread;
//   * every Newton iteration is one evaluation of condense_rates per particle plus the exact solve of
//     the arrow-head system (I - gamma J) x = r by the Schur complement on H (the reference's
//     preconditioner, condense.F90:625-695): two cell-wide sums; J is evaluated at the predicted state;
//   * norms are CVODE's weighted RMS norms with weights 1 / (reltol |y| + abstol): one cell-wide sum;
//   * the (T, p)-dependent terms of condense_rates are evaluated once per step (all Newton iterations
//     of a step share t_n).
#define BDF_QMAX 5
#define BDF_NARR 10  // shared arrays per particle: zn[0..5], abstol, D_dry, kappa, num_conc

// cvPredict (SIGN = +1) / cvRestore (SIGN = -1) for order Q: the Pascal triangle on registers
template <int Q, int SIGN>
__device__ __forceinline__ void bdf_pascal(double* __restrict__ zn0, int stride) {
  double z[Q + 1];
#pragma unroll
  for (int j = 0; j <= Q; j++) z[j] = zn0[j * stride];
#pragma unroll
  for (int kk = 1; kk <= Q; kk++)
#pragma unroll
    for (int j = Q; j >= kk; j--) z[j - 1] = (SIGN > 0) ? z[j - 1] + z[j] : z[j - 1] - z[j];
#pragma unroll
  for (int j = 0; j < Q; j++) zn0[j * stride] = z[j];
}
template <int SIGN>
__device__ __forceinline__ void bdf_pascal_q(int q, double* __restrict__ zn0, int stride) {
  switch (q) {
    case 1: bdf_pascal<1, SIGN>(zn0, stride); break;
    case 2: bdf_pascal<2, SIGN>(zn0, stride); break;
    case 3: bdf_pascal<3, SIGN>(zn0, stride); break;
    case 4: bdf_pascal<4, SIGN>(zn0, stride); break;
    default: bdf_pascal<5, SIGN>(zn0, stride); break;
  }
}
// x^(1/n) for the step-size ratios
__device__ __forceinline__ double bdf_root(double x, int n) { return fast_exp(log(fmax(x, 1e-300)) / (double)n); }

template <int PPT, int NT>
__global__ void __launch_bounds__(NT) k_condense_bdf(DevView v, const double* T0a /* may alias v.temp */,
                                                     const double* p0a, const double* T1 /* may alias v.temp */,
                                                     const double* p1, CondParams prm, int* __restrict__ stats,
                                                     double* __restrict__ wre_out, int n_limit_lo, int n_limit_hi) {
  const int cell = blockIdx.x;
  const int n = v.count[cell];
  if (n <= n_limit_lo || n > n_limit_hi) return;  // another instantiation owns this cell
  if (v.status[cell] != 0 || cell_is_boundary(v, cell)) return;
  const int tid = threadIdx.x;
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  __shared__ double s_red[2 * 4 * (NT / 32)];
  extern __shared__ double s_K[];  // [BDF_NARR][PPT][NT]
#define ZN(j, k) s_K[((j) * PPT + (k)) * NT + tid]
#define S_ATOL(k) ZN(6, k)
  const double* sDdry = &ZN(7, 0);
  const double* sKap = &ZN(8, 0);
  const double* sNc = &ZN(9, 0);
  int par = 0;
  __shared__ double s_etab[32];
  if (tid < 32) s_etab[tid] = c_exp2tab[tid];
  __syncthreads();
  // aero_state_reweight after the solve is the identity for flat weights (see k_condense)
  bool nonflat = false;
  for (int q = 0; q < v.C * v.G; q++)
    if (v.wtype[(size_t)cell * v.C * v.G + q] != PMCGPU_WEIGHT_NONE) nonflat = true;
  const int iw = v.i_water - 1;
  const double T0 = T0a[cell], p0 = p0a[cell];  // env_state_initial
  const double del_t = prm.del_t;
  const double Tdot = (T1[cell] - T0) / del_t, pdot = (p1[cell] - p0) / del_t;
  const double H0 = v.rh[cell];

  // ---- set-up (condense.F90:196-237): kappa, D_dry, num_conc, D, abstol per particle.
  // Slots past the population hold a harmless wet dummy particle that never enters a sum.
  double wsum = 0.0;
#pragma unroll
  for (int k = 0; k < PPT; k++) {
    const int p = tid + k * NT;
#pragma unroll
    for (int j = 1; j <= BDF_QMAX; j++) ZN(j, k) = 0.0;
    ZN(0, k) = 2.0;
    S_ATOL(k) = 1.0;
    ZN(7, k) = 1.0;
    ZN(8, k) = 0.5;
    ZN(9, k) = 1.0;
    if (p < n) {
      double vt = 0.0, vs = 0.0, ks = 0.0;  // aero_particle_solute_kappa (aero_particle.F90:724-755)
      const double M_w = v.molw[iw], rho_w = v.density[iw];
      for (int s = 0; s < v.S; s++) {
        const double x = v.vol[(size_t)s * NS + base + p];
        vt += x;
        if (s != iw) {
          vs += x;
          const double kp = (v.nions[s] > 0) ? M_w * v.density[s] / (v.molw[s] * rho_w) * (double)v.nions[s] : v.kappa[s];
          ks += x * kp;
        }
      }
      const int gc = v.wgc[base + p];
      const double r = vol2rad(vt);
      const double ddry = 2.0 * vol2rad(vs);
      const double ncp = awa_nc_at_radius(v, cell, gc / v.G + 1, r);
      ZN(7, k) = ddry;
      ZN(8, k) = ks / vs;
      ZN(9, k) = ncp;
      ZN(0, k) = 2.0 * r;
      S_ATOL(k) = fmax(1e-30, prm.abstol_scale * (2.0 * r - ddry));
      wsum += v.vol[(size_t)iw * NS + base + p] * ncp;
      if (nonflat) {
        const size_t wi = (size_t)cell * v.C * v.G + gc;
        v.lnr[base + p] = weight_nc(v.wtype[wi], v.wmag[wi], v.wexp[wi], r);
      }
    }
  }
  {
    double x[1] = {wsum};
    block_red<1, NT, false>(x, s_red, par);
    wsum = x[0];  // water_vol_conc_initial
  }

  // ---- integrator state (names follow cvode_impl.h)
  const double ADDON = 1e-6, BIAS1 = 6.0, BIAS2 = 6.0, BIAS3 = 10.0, ETAMX1 = 10000.0, ETAMX2 = 10.0, ETAMX3 = 10.0,
               ETAMXF = 0.2, ETAMIN = 0.1, ETACF = 0.25, THRESH = 1.5, CRDOWN = 0.3, RDIV = 2.0, CORTES = 0.1;
  const int SMALL_NST = 10, MXNCF = 10, MXNEF = 7, MXNEF1 = 3, SMALL_NEF = 2, LONG_WAIT = 10, MAXCOR = 3;
  double zH[BDF_QMAX + 1] = {H0, 0.0, 0.0, 0.0, 0.0, 0.0};
  double tau[BDF_QMAX + 2] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  double l[BDF_QMAX + 1] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  double tq[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  int q = 1, L = 2, qprime = 1, qwait = 2, nst = 0, n_fail = 0, n_evals = 0, fail = 0;
  double h = 0.0, hprime = 0.0, eta = 1.0, hscale = 0.0, tn = 0.0, etamax = ETAMX1, rl1 = 1.0, gamma = 0.0,
         saved_tq5 = 0.0, acnrm = 0.0, crate = 1.0;
  double ewt[PPT], ewtH = 0.0;
  double acor[PPT], acorH = 0.0;
  double u1[PPT], u2[PPT], u3[PPT], u4[PPT];  // unused outputs of the no-Jacobian evaluations

  auto set_ewt = [&]() {  // cvEwtSetSV
#pragma unroll
    for (int k = 0; k < PPT; k++) ewt[k] = fast_rcp(fma(prm.reltol, fabs(ZN(0, k)), S_ATOL(k)));
    ewtH = 1.0 / (prm.reltol * fabs(zH[0]) + prm.abstol_H);
  };
  auto wrms = [&](const double (&vD)[PPT], double vH) -> double {
    double acc[1] = {0.0};
#pragma unroll
    for (int k = 0; k < PPT; k++)
      if (tid + k * NT < n) acc[0] = fma(vD[k] * ewt[k], vD[k] * ewt[k], acc[0]);
    block_red<1, NT, false>(acc, s_red, par);
    return sqrt((acc[0] + (vH * ewtH) * (vH * ewtH)) / (double)(n + 1));
  };
  // condense_vf_f (condense.F90:509-557): f = (Ddot_i, Hdot)
  auto eval_f = [&](const CondEnv& e, double Hs, const double (&yD)[PPT], double (&fD)[PPT], double& fH) {
    double hd[PPT];
    cond_rates_vec<PPT, false, NT>(s_etab, e, Hs, yD, sDdry, sKap, sNc, fD, hd, u1, u2, u3, u4);
    double acc[1] = {0.0};
#pragma unroll
    for (int k = 0; k < PPT; k++) {
      if (tid + k * NT < n) acc[0] += hd[k]; else fD[k] = 0.0;
    }
    block_red<1, NT, false>(acc, s_red, par);
    fH = acc[0] + (-e.dP0_dT_div_P0 * Tdot * Hs + Hs * pdot / e.p);
    n_evals++;
  };
  // f and the arrow-head Jacobian (condense_jac, condense.F90:566-616), folded with gamma into what the
  // Schur-complement solve needs: inv = 1 / (1 - gamma a), P = gamma c inv, Q = gamma b inv,
  // lhsH = 1 - gamma dH'/dH - sum (gamma b)(gamma c) inv  (condense.F90:665-672)
  auto eval_fj = [&](const CondEnv& e, double Hs, const double (&yD)[PPT], double (&fD)[PPT], double& fH,
                     double (&inv)[PPT], double (&Pc)[PPT], double (&Qb)[PPT], double& lhsH) {
    double hd[PPT], dh[PPT];
    cond_rates_vec<PPT, true, NT>(s_etab, e, Hs, yD, sDdry, sKap, sNc, fD, hd, inv, Qb, Pc, dh);  // inv = a, Qb = b, Pc = c
    double acc[3] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int k = 0; k < PPT; k++) {
      const double iv = fast_rcp(fma(-gamma, inv[k], 1.0));
      const double gb = gamma * Qb[k], gc = gamma * Pc[k];
      if (tid + k * NT < n) {
        acc[0] += hd[k];
        acc[1] += dh[k];
        acc[2] += gb * gc * iv;
        inv[k] = iv;
        Pc[k] = gc * iv;
        Qb[k] = gb * iv;
      } else {
        fD[k] = 0.0; inv[k] = 1.0; Pc[k] = 0.0; Qb[k] = 0.0;
      }
    }
    block_red<3, NT, false>(acc, s_red, par);
    fH = acc[0] + (-e.dP0_dT_div_P0 * Tdot * Hs + Hs * pdot / e.p);
    const double jdd = acc[1] + (-e.dP0_dT_div_P0 * Tdot + pdot / e.p);
    lhsH = 1.0 - gamma * jdd - acc[2];
    n_evals++;
  };
  auto rescale = [&]() {  // cvRescale
    double factor = eta;
    for (int j = 1; j <= q; j++) {
#pragma unroll
      for (int k = 0; k < PPT; k++) ZN(j, k) *= factor;
      zH[j] *= factor;
      factor *= eta;
    }
    h = hscale * eta;
    hscale = h;
  };
  auto predict = [&]() {  // cvPredict
#pragma unroll
    for (int k = 0; k < PPT; k++) bdf_pascal_q<1>(q, &ZN(0, k), PPT * NT);
    bdf_pascal_q<1>(q, zH, 1);
  };
  auto restore = [&]() {  // cvRestore
#pragma unroll
    for (int k = 0; k < PPT; k++) bdf_pascal_q<-1>(q, &ZN(0, k), PPT * NT);
    bdf_pascal_q<-1>(q, zH, 1);
  };
  auto set_coeffs = [&]() {  // cvSetBDF + cvSetTqBDF
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
    tq[2] = fabs(A1 / (alpha0 * A2));
    tq[5] = fabs(A2 * xistar_inv / (l[q] * xi_inv));
    if (qwait == 1) {
      if (q > 1) {
        const double Cc = xistar_inv / l[q], A3 = alpha0 + 1.0 / q, A4 = alpha0_hat + xi_inv;
        const double Cpinv = (1.0 - A4 + A3) / A3;
        tq[1] = fabs(Cc * Cpinv);
      } else {
        tq[1] = 1.0;
      }
      hsum += tau[q];
      xi_inv = h / hsum;
      const double A5 = alpha0 - 1.0 / (q + 1), A6 = alpha0_hat - xi_inv;
      const double Cppinv = (1.0 - A6 + A5) / A2;
      tq[3] = fabs(Cppinv / (xi_inv * (q + 2) * A5));
    }
    tq[4] = CORTES / tq[2];
    rl1 = 1.0 / l[1];
    gamma = h * rl1;
  };
  auto increase_order = [&]() {  // cvIncreaseBDF: zn[L] = A1 * (saved acor in zn[QMAX])
    for (int i = 0; i <= BDF_QMAX; i++) l[i] = 0.0;
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
#pragma unroll
    for (int k = 0; k < PPT; k++) ZN(L, k) = A1 * ZN(BDF_QMAX, k);
    zH[L] = A1 * zH[BDF_QMAX];
    for (int j = 2; j <= q; j++) {
#pragma unroll
      for (int k = 0; k < PPT; k++) ZN(j, k) = fma(l[j], ZN(L, k), ZN(j, k));
      zH[j] = fma(l[j], zH[L], zH[j]);
    }
  };
  auto decrease_order = [&]() {  // cvDecreaseBDF
    for (int i = 0; i <= BDF_QMAX; i++) l[i] = 0.0;
    l[2] = 1.0;
    double hsum = 0.0;
    for (int j = 1; j <= q - 2; j++) {
      hsum += tau[j];
      const double xi = hsum / hscale;
      for (int i = j + 2; i >= 2; i--) l[i] = l[i] * xi + l[i - 1];
    }
    for (int j = 2; j < q; j++) {
#pragma unroll
      for (int k = 0; k < PPT; k++) ZN(j, k) = fma(-l[j], ZN(q, k), ZN(j, k));
      zH[j] = fma(-l[j], zH[q], zH[j]);
    }
  };

  // ---- first call of CVode: f(t0, y0), initial step (cvHin), zn[1] = h f
  set_ewt();
  {
    double y0[PPT], f0[PPT], f0H;
#pragma unroll
    for (int k = 0; k < PPT; k++) y0[k] = ZN(0, k);
    eval_f(cond_env_fast(T0, p0, Tdot, pdot, 0.0), zH[0], y0, f0, f0H);
#pragma unroll
    for (int k = 0; k < PPT; k++) ZN(1, k) = f0[k];
    zH[1] = f0H;
    // cvUpperBoundH0
    double hub_inv[1] = {fabs(f0H) / (0.1 * fabs(zH[0]) + 1.0 / ewtH)};
#pragma unroll
    for (int k = 0; k < PPT; k++)
      if (tid + k * NT < n) hub_inv[0] = fmax(hub_inv[0], fabs(f0[k]) / (0.1 * fabs(y0[k]) + 1.0 / ewt[k]));
    block_red<1, NT, true>(hub_inv, s_red, par);
    const double hlb = 100.0 * 2.220446049250313e-16 * del_t;
    double hub = 0.1 * del_t;
    if (hub * hub_inv[0] > 1.0) hub = 1.0 / hub_inv[0];
    double hg = sqrt(hlb * hub), hnew = hg;
    if (!(hub < hlb)) {
      for (int count = 1; count <= 4; count++) {
        // cvYddNorm: || (f(t0 + hg, y0 + hg f0) - f0) / hg ||
        double yy[PPT], ff[PPT], ffH;
#pragma unroll
        for (int k = 0; k < PPT; k++) yy[k] = fma(hg, f0[k], y0[k]);
        eval_f(cond_env_fast(T0, p0, Tdot, pdot, hg), fma(hg, f0H, zH[0]), yy, ff, ffH);
#pragma unroll
        for (int k = 0; k < PPT; k++) ff[k] = (ff[k] - f0[k]) / hg;
        const double yddnrm = wrms(ff, (ffH - f0H) / hg);
        hnew = (yddnrm * hub * hub > 2.0) ? sqrt(2.0 / yddnrm) : sqrt(hg * hub);
        if (count == 4) break;
        const double hrat = hnew / hg;
        if (hrat > 0.5 && hrat < 2.0) break;
        if (count > 1 && hrat > 2.0) {
          hnew = hg;
          break;
        }
        hg = hnew;
      }
      h = fmin(hub, fmax(hlb, 0.5 * hnew));
    } else {
      h = hg;
    }
    hscale = h;
    hprime = h;
#pragma unroll
    for (int k = 0; k < PPT; k++) ZN(1, k) = h * f0[k];
    zH[1] = h * f0H;
  }

  // ---- CVode(..., CV_NORMAL): internal steps until tn >= t_final
  for (;;) {
    if (nst > 0) set_ewt();
    if (nst >= prm.max_steps) { fail = 6; break; }
    // ------------------------------------------------------------ cvStep
    const double saved_t = tn;
    int ncf = 0, nef = 0;
    if (nst > 0 && hprime != h) {  // cvAdjustParams
      if (qprime != q) {
        if (qprime > q) increase_order(); else decrease_order();
        q = qprime;
        L = q + 1;
        qwait = L;
      }
      rescale();
    }
    double dsm = 0.0;
    for (;;) {
      tn += h;
      predict();
      set_coeffs();
      // ---- cvNls: Newton iteration with the exact arrow-head solve
      bool converged = false;
      const CondEnv env_n = cond_env_fast(T0, p0, Tdot, pdot, tn);
      {
        crate = 1.0;
        double y[PPT], f[PPT], inv[PPT], Pc[PPT], Qb[PPT], fH, lhsH;
#pragma unroll
        for (int k = 0; k < PPT; k++) { y[k] = ZN(0, k); acor[k] = 0.0; }
        double yH = zH[0];
        acorH = 0.0;
        eval_fj(env_n, yH, y, f, fH, inv, Pc, Qb, lhsH);
        double delp = 0.0;
        for (int m = 0;;) {
          double b[PPT];
          double acc[1] = {0.0};
#pragma unroll
          for (int k = 0; k < PPT; k++) {
            b[k] = fma(gamma, f[k], -fma(rl1, ZN(1, k), acor[k]));
            acc[0] = fma(Pc[k], b[k], acc[0]);  // P = 0 in the slots past the population
          }
          block_red<1, NT, false>(acc, s_red, par);
          const double bH = (fma(gamma, fH, -fma(rl1, zH[1], acorH)) + acc[0]) / lhsH;
#pragma unroll
          for (int k = 0; k < PPT; k++) b[k] = fma(Qb[k], bH, b[k] * inv[k]);
          const double del = wrms(b, bH);
          if (!isfinite(del)) break;
#pragma unroll
          for (int k = 0; k < PPT; k++) {
            acor[k] += b[k];
            y[k] = ZN(0, k) + acor[k];
          }
          acorH += bH;
          yH = zH[0] + acorH;
          if (m > 0) crate = fmax(CRDOWN * crate, del / delp);
          const double dcon = del * fmin(1.0, crate) / tq[4];
          if (dcon <= 1.0) {
            acnrm = (m == 0) ? del : wrms(acor, acorH);
            converged = true;
            break;
          }
          m++;
          if (m == MAXCOR || (m >= 2 && del > RDIV * delp)) break;
          delp = del;
          eval_f(env_n, yH, y, f, fH);
        }
      }
      if (!converged) {  // cvHandleNFlag
        ncf++;
        n_fail++;
        etamax = 1.0;
        if (ncf == MXNCF || !(fabs(h) > 1e-300)) { fail = 7; break; }
        eta = ETACF;
        tn = saved_t;
        restore();
        rescale();
        continue;
      }
      dsm = acnrm * tq[2];  // cvDoErrorTest
      if (dsm <= 1.0) break;
      nef++;
      n_fail++;
      tn = saved_t;
      restore();
      if (nef == MXNEF) { fail = 7; break; }
      etamax = 1.0;
      if (nef <= MXNEF1) {
        eta = 1.0 / (bdf_root(BIAS2 * dsm, L) + ADDON);
        eta = fmax(ETAMIN, eta);
        if (nef >= SMALL_NEF) eta = fmin(eta, ETAMXF);
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
      {
        double y[PPT], f[PPT], fH;
#pragma unroll
        for (int k = 0; k < PPT; k++) y[k] = ZN(0, k);
        eval_f(env_n, zH[0], y, f, fH);
#pragma unroll
        for (int k = 0; k < PPT; k++) ZN(1, k) = h * f[k];
        zH[1] = h * fH;
      }
    }
    if (fail) break;
    // ------------------------------------------------------------ cvCompleteStep
    nst++;
    for (int i = q; i >= 2; i--) tau[i] = tau[i - 1];
    if (q == 1 && nst > 1) tau[2] = tau[1];
    tau[1] = h;
    for (int j = 0; j <= q; j++) {
#pragma unroll
      for (int k = 0; k < PPT; k++) ZN(j, k) = fma(l[j], acor[k], ZN(j, k));
      zH[j] = fma(l[j], acorH, zH[j]);
    }
    qwait--;
    if (qwait == 1 && q != BDF_QMAX) {
#pragma unroll
      for (int k = 0; k < PPT; k++) ZN(BDF_QMAX, k) = acor[k];
      zH[BDF_QMAX] = acorH;
      saved_tq5 = tq[5];
    }
    // ------------------------------------------------------------ cvPrepareNextStep
    if (etamax == 1.0) {
      qwait = max(qwait, 2);
      qprime = q;
      hprime = h;
      eta = 1.0;
    } else {
      const double etaq = 1.0 / (bdf_root(BIAS2 * dsm, L) + ADDON);
      if (qwait != 0) {
        eta = etaq;
        qprime = q;
      } else {
        qwait = 2;
        double etaqm1 = 0.0, etaqp1 = 0.0;
        if (q > 1) {  // cvComputeEtaqm1
          double t[PPT];
#pragma unroll
          for (int k = 0; k < PPT; k++) t[k] = ZN(q, k);
          const double ddn = wrms(t, zH[q]) * tq[1];
          etaqm1 = 1.0 / (bdf_root(BIAS1 * ddn, q) + ADDON);
        }
        if (q != BDF_QMAX && saved_tq5 != 0.0) {  // cvComputeEtaqp1
          double hrL = 1.0;
          for (int i = 0; i < L; i++) hrL *= h / tau[2];
          const double cquot = (tq[5] / saved_tq5) * hrL;
          double t[PPT];
#pragma unroll
          for (int k = 0; k < PPT; k++) t[k] = fma(-cquot, ZN(BDF_QMAX, k), acor[k]);
          const double dup = wrms(t, fma(-cquot, zH[BDF_QMAX], acorH)) * tq[3];
          etaqp1 = 1.0 / (bdf_root(BIAS3 * dup, L + 1) + ADDON);
        }
        const double etam = fmax(etaqm1, fmax(etaq, etaqp1));  // cvChooseEta
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
#pragma unroll
          for (int k = 0; k < PPT; k++) ZN(BDF_QMAX, k) = acor[k];
          zH[BDF_QMAX] = acorH;
        }
      }
      if (eta < THRESH) {  // cvSetEta (no hmax)
        eta = 1.0;
        hprime = h;
      } else {
        eta = fmin(eta, etamax);
        hprime = h * eta;
      }
    }
    etamax = (nst <= SMALL_NST) ? ETAMX2 : ETAMX3;
    if (tn - del_t >= 0.0) break;
  }

  // ---- CVodeGetDky(t_final, 0) and write back (condense.F90:266-293): water volume, clamp at zero, balance
  double wfin = 0.0;
  const double sfac = (del_t - tn) / h;
  double Hfin = zH[q];
  for (int j = q - 1; j >= 0; j--) Hfin = fma(Hfin, sfac, zH[j]);
#pragma unroll
  for (int k = 0; k < PPT; k++) {
    const int p = tid + k * NT;
    if (p < n && !fail) {
      double Df = ZN(q, k);
      for (int j = q - 1; j >= 0; j--) Df = fma(Df, sfac, ZN(j, k));
      double vs = 0.0;  // aero_particle_solute_volume
      for (int s = 0; s < v.S; s++)
        if (s != iw) vs += v.vol[(size_t)s * NS + base + p];
      double wv = rad2vol(Df / 2.0) - vs;
      wv = fmax(0.0, wv);
      v.vol[(size_t)iw * NS + base + p] = wv;
      wfin += wv * sNc[k * NT];
      if (nonflat) v.gct_alt[base + p] = vol2rad(vs + wv);
    }
  }
  {
    double x[1] = {wfin};
    block_red<1, NT, false>(x, s_red, par);
    wfin = x[0];
  }
  if (tid == 0) {
    if (fail) {
      if (v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_CONDENSE;
    } else {
      const double Tf = T1[cell], pf = p1[cell];
      const double V_comp_ratio = Tf * p0 / (T0 * pf);
      const double rho_w = v.density[iw], M_w = v.molw[iw];
      const double P0i = PmcConst::water_eq_vap_press * pow(10.0, 7.45 * (T0 - PmcConst::water_freeze_temp) / (T0 - 38.0));
      const double P0f = PmcConst::water_eq_vap_press * pow(10.0, 7.45 * (Tf - PmcConst::water_freeze_temp) / (Tf - 38.0));
      const double vap_i = M_w / (PmcConst::univ_gas_const * T0) * P0i * H0 / rho_w;
      const double vap_f = M_w / (PmcConst::univ_gas_const * Tf) * P0f * Hfin * V_comp_ratio / rho_w;
      wre_out[cell] = ((vap_f - vap_i) + (wfin - wsum)) / (vap_f + wfin);
      v.rh[cell] = Hfin;
      v.temp[cell] = Tf;
      v.pres[cell] = pf;
    }
    stats[3 * cell + 0] = nst;
    stats[3 * cell + 1] = n_fail;
    stats[3 * cell + 2] = n_evals;  // rate evaluations per particle
    atomicAdd(&v.work[0], (unsigned long long)n_evals * (unsigned long long)n);
    atomicAdd(&v.work[1], (unsigned long long)n);
  }
#undef ZN
#undef S_ATOL
}


// device pointers T1/p1 (final env per cell)
static size_t bdf_smem(int ppt) { return (size_t)BDF_NARR * ppt * COND_THREADS * sizeof(double); }
static size_t ros_smem(int ppt, int nt) { return (size_t)COND_NARR(ppt) * ppt * nt * sizeof(double); }

int pmc_reweight_async(pmcgpu_ctx* ctx, unsigned step, unsigned pass);

int pmc_condense_async(pmcgpu_ctx* ctx, const double* dT0, const double* dp0, const double* dT1, const double* dp1,
                       double del_t, double reltol, unsigned step) {
  if (!ctx->cfg_cond) {
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<4, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(4, 128)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<4, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(4, 64)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<4, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(4, 256)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<4, 160>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(4, 160)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<4, 192>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(4, 192)));

    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<4, 288>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(4, 288)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<8, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(8, 256)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense_bdf<2, COND_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bdf_smem(2)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense_bdf<4, COND_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bdf_smem(4)));
    CUDA_OK(ctx, cudaFuncSetAttribute(k_condense_bdf<8, COND_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bdf_smem(8)));
    ctx->cfg_cond = true;
  }
  DevView& v = ctx->v;
  DevView* cw = &v;
  CUDA_OK(ctx, cudaMemsetAsync(v.cond_stats, 0, 3 * v.n_cells * sizeof(int), ctx->stream));
  CUDA_OK(ctx, cudaMemsetAsync(v.cond_wre, 0, v.n_cells * sizeof(double), ctx->stream));
  if (v.i_water < 1) {
    ctx->err = "condense_particles needs a water species (aero_data%i_water)";
    return PMCGPU_ERR_ARG;
  }
  CondParams prm;
  prm.del_t = del_t;
  prm.reltol = reltol > 0.0 ? reltol : 1e-8;
  prm.abstol_scale = 1e-8;
  prm.abstol_H = 1e-10;
  prm.max_steps = 100000;
  prm.debug_cell = getenv("PMCGPU_DEBUG_CELL") ? atoi(getenv("PMCGPU_DEBUG_CELL")) : -1;
  {
  ProfScope ps_(ctx, "k_condense");
  // the population size of a cell picks the instantiation: 1..256 -> PPT 1, ..512 -> 4 x 128 threads (BDF: 2),
  // ..1024 -> 4, ..2048 -> 8
  const int cap = v.cap;
  if (ctx->cond_method == 5) {
#define BDF_LAUNCH(PPT, LO, HI)                                                                                   \
  k_condense_bdf<PPT, COND_THREADS><<<v.n_cells, COND_THREADS, bdf_smem(PPT), ctx->stream>>>(                    \
      v, dT0, dp0, dT1, dp1, prm, cw->cond_stats, cw->cond_wre, LO, HI);                                          \
  LAUNCH_CHECK(ctx)
    BDF_LAUNCH(1, -1, 256);
    if (cap > 256) { BDF_LAUNCH(2, 256, 512); }
    if (cap > 512) { BDF_LAUNCH(4, 512, 1024); }
    if (cap > 1024) { BDF_LAUNCH(8, 1024, 2048); }
#undef BDF_LAUNCH
  } else {
    // the variants touch disjoint cells: each goes to its own side stream between a fork and a join on the context's
    // stream (serialised, every launch ends with SMs idling behind its last ~0.9 ms cells: +1.1 ms per step on a
    // 4000-cell patch with seven launches)
    static const bool side = getenv("PMCGPU_COND_SERIAL") == nullptr;  // measurements only
    if (side && !ctx->cond_fork) {
      CUDA_OK(ctx, cudaEventCreateWithFlags(&ctx->cond_fork, cudaEventDisableTiming));
      for (int q = 0; q < 8; q++) {
        CUDA_OK(ctx, cudaStreamCreateWithFlags(&ctx->cond_stream[q], cudaStreamNonBlocking));
        CUDA_OK(ctx, cudaEventCreateWithFlags(&ctx->cond_join[q], cudaEventDisableTiming));
      }
    }
    int n_side = 0;
    if (side) CUDA_OK(ctx, cudaEventRecord(ctx->cond_fork, ctx->stream));
#define ROS_LAUNCH(PPT, NT, LO, HI)                                                                               \
  {                                                                                                               \
    cudaStream_t st_ = ctx->stream;                                                                               \
    if (side) {                                                                                                   \
      st_ = ctx->cond_stream[n_side];                                                                             \
      CUDA_OK(ctx, cudaStreamWaitEvent(st_, ctx->cond_fork, 0));                                                  \
    }                                                                                                             \
    k_condense<PPT, NT><<<v.n_cells, NT, ros_smem(PPT, NT), st_>>>(v, dT0, dp0, dT1, dp1, prm, cw->cond_stats,   \
                                                                   cw->cond_wre, LO, HI, GridRed{});             \
    LAUNCH_CHECK(ctx);                                                                                            \
    if (side) {                                                                                                   \
      CUDA_OK(ctx, cudaEventRecord(ctx->cond_join[n_side], st_));                                                 \
      CUDA_OK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->cond_join[n_side], 0));                                  \
      n_side++;                                                                                                   \
    }                                                                                                             \
  }
    // 256 threads x 4 particles per thread measured faster than 512 x 2 at ~1000 particles per cell
    // (300 vs 318 ms per K4 step): instruction-level parallelism beats the extra warps here
    ROS_LAUNCH(4, 64, -1, 256);  // small cells: four CTAs of 64 threads per SM
    // 257..512 particles: 128 threads x 4 particles, so that two cells share an SM (244 registers x 128
    // threads each) instead of one CTA of 256 threads x 2 particles leaving half the register file idle
    if (cap > 256) { ROS_LAUNCH(4, 128, 256, 512); }
    // (round 2: 3 particles x 384 threads -- a third warp per scheduler at 168 registers -- measured 224.7 ms per
    // K4 step, 2 x 512 236.8 ms, against 206.0 ms for 4 x 256: profiles/r02_condense_notes.md)
    // 513..768 particles (a 500-particle population between rebalances, config K5): five / six warps x 4 particles with
    // TWO cells per SM (launch bounds: 204 / 170 registers) instead of one cell on a 256-thread CTA at twice the cost
    static const bool mid = getenv("PMCGPU_COND_NO_MID") == nullptr;  // measurements only
    if (cap > 512 && mid) { ROS_LAUNCH(4, 160, 512, 640); }
    if (cap > 640 && mid) { ROS_LAUNCH(4, 192, 640, 768); }
    if (cap > 512) { ROS_LAUNCH(4, 256, mid ? 768 : 512, 1024); }
    // cells a little over 1024 particles (a 1000-particle population drifts there between rebalances):
    // nine warps x 4 particles at 168 registers (three warps share a scheduler's 16384 registers; 422 bytes
    // of spills) instead of 8 particles per thread (no predictor, 255 registers, 838 bytes of spills)
    if (cap > 1024) { ROS_LAUNCH(4, 288, 1024, 1152); }
    if (cap > 1152) { ROS_LAUNCH(8, 256, 1152, 2048); }
#undef ROS_LAUNCH
  }
  if (cap > 2048) {
    // single-box sized cells: the RODAS4 integrator with one cell spread over a co-resident grid
    // (both methods: the BDF restatement has no grid-wide variant)
    int& n_sm = ctx->cond_n_sm;
    int& max_blk4 = ctx->cond_max_blk4;
    int& max_blk8 = ctx->cond_max_blk8;
    if (!n_sm) {
      int dev = 0;
      CUDA_OK(ctx, cudaGetDevice(&dev));
      CUDA_OK(ctx, cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
      CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<4, 256, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(4, 256)));
      CUDA_OK(ctx, cudaFuncSetAttribute(k_condense<8, 256, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ros_smem(8, 256)));
      CUDA_OK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&max_blk4, k_condense<4, 256, true>, 256, ros_smem(4, 256)));
      CUDA_OK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&max_blk8, k_condense<8, 256, true>, 256, ros_smem(8, 256)));
    }
    if (!ctx->cond_grid_buf) {
      CUDA_OK(ctx, cudaMalloc((void**)&ctx->cond_grid_buf, 2 * 4 * COND_GRID_MAXB * sizeof(double) + 256));
      ctx->allocs.push_back(ctx->cond_grid_buf);
    }
    GridRed gr;
    gr.buf = ctx->cond_grid_buf;
    gr.bar = (unsigned*)(ctx->cond_grid_buf + 2 * 4 * COND_GRID_MAXB);
    gr.gen = 0;
    gr.par = 0;
    CUDA_OK(ctx, cudaMemsetAsync(gr.bar, 0, sizeof(unsigned), ctx->stream));
    const int ppt = (cap <= n_sm * max_blk4 * 4 * 256) ? 4 : 8;
    int nblk = n_sm * (ppt == 4 ? max_blk4 : max_blk8);
    if (nblk > COND_GRID_MAXB) nblk = COND_GRID_MAXB;
    if (nblk < 1 || (long long)cap > (long long)nblk * ppt * 256) {
      ctx->err = "condense_particles: cell_capacity exceeds what one co-resident grid integrates";
      return PMCGPU_ERR_CAPACITY;
    }
    int lo = 2048, hi = cap;
    void* args[] = {&v, &dT0, &dp0, &dT1, &dp1, &prm, &cw->cond_stats, &cw->cond_wre, &lo, &hi, &gr};
    if (ppt == 4)
      CUDA_OK(ctx, cudaLaunchCooperativeKernel((void*)k_condense<4, 256, true>, dim3(nblk), dim3(256), args, ros_smem(4, 256), ctx->stream));
    else
      CUDA_OK(ctx, cudaLaunchCooperativeKernel((void*)k_condense<8, 256, true>, dim3(nblk), dim3(256), args, ros_smem(8, 256), ctx->stream));
    ctx->launches++;
  }
  }
  if (ctx->any_nonflat) {  // aero_state_reweight (condense.F90:297), pass 0
    int rc = pmc_reweight_async(ctx, step, 0u);
    if (rc) return rc;
  }
  return PMCGPU_OK;
}

extern "C" int pmcgpu_set_condense_method(pmcgpu_ctx* ctx, int method) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || (method != 4 && method != 5)) return PMCGPU_ERR_ARG;
  ctx->cond_method = method;
  return PMCGPU_OK;
}

extern "C" int pmcgpu_condense_particles(pmcgpu_ctx* ctx, const double* temp_final, const double* pressure_final,
                                         double del_t, double reltol, double* rel_humid_out, int* stats_out,
                                         double* water_rel_error_out) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_condense_particles");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !temp_final || !pressure_final || !(del_t > 0.0)) return PMCGPU_ERR_ARG;
  DevView& v = ctx->v;
  DevView* cw = &v;
  int rc;
  CUDA_OK(ctx, cudaMemcpyAsync(cw->cond_T1, temp_final, v.n_cells * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(ctx, cudaMemcpyAsync(cw->cond_p1, pressure_final, v.n_cells * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  rc = pmc_condense_async(ctx, v.temp, v.pres, cw->cond_T1, cw->cond_p1, del_t, reltol, ctx->rng_step);
  if (rc) return rc;
  if (rel_humid_out)
    CUDA_OK(ctx, cudaMemcpyAsync(rel_humid_out, v.rh, v.n_cells * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (stats_out)
    CUDA_OK(ctx, cudaMemcpyAsync(stats_out, cw->cond_stats, 3 * v.n_cells * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if (water_rel_error_out)
    CUDA_OK(ctx, cudaMemcpyAsync(water_rel_error_out, cw->cond_wre, v.n_cells * sizeof(double), cudaMemcpyDeviceToHost,
                                 ctx->stream));
  return pmc_first_status(ctx);
}



// ---- condense_equilib_particles (condense.F90:701-770), one thread per particle
__global__ void k_condense_equilib(DevView v) {
  const int cell = blockIdx.x;
  const int n = v.count[cell];
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  const int iw = v.i_water - 1;
  const double X = 4.0 * PmcConst::water_molec_weight * PmcConst::water_surf_eng /
                   (PmcConst::univ_gas_const * v.temp[cell] * PmcConst::water_density);
  const double RH = v.rh[cell];
  bool nonflat = false;
  for (int q = 0; q < v.C * v.G; q++)
    if (v.wtype[(size_t)cell * v.C * v.G + q] != PMCGPU_WEIGHT_NONE) nonflat = true;
  for (int p = threadIdx.x; p < n; p += blockDim.x) {
    double vs = 0.0, ks = 0.0;
    const double M_w = v.molw[iw], rho_w = v.density[iw];
    for (int s = 0; s < v.S; s++) {
      if (s == iw) continue;
      const double x = v.vol[(size_t)s * NS + base + p];
      vs += x;
      const double kp = (v.nions[s] > 0) ? M_w * v.density[s] / (v.molw[s] * rho_w) * (double)v.nions[s] : v.kappa[s];
      ks += x * kp;
    }
    const double kappa = ks / vs;
    const double D_dry = 2.0 * vol2rad(vs);
    const double Dd3 = D_dry * D_dry * D_dry;
    double D = D_dry, g = 0.0, dg_dD = 1.0;
    for (int newton_step = 1; newton_step <= 20; newton_step++) {
      D = D - g / dg_dD;
      const double D3 = D * D * D;
      const double den = D3 + (kappa - 1.0) * Dd3;
      const double a_w = (D3 - Dd3) / den;
      const double daw_dD = 3.0 * (D * D) * kappa * Dd3 / (den * den);
      const double ex = exp(X / D);
      g = RH - a_w * ex;
      dg_dD = -daw_dD * ex + a_w * ex * (X / (D * D));
    }
    if (nonflat) {
      // aero_state_num_conc_for_reweight before, new radius after (k_reweight reads both)
      double vt = 0.0;
      for (int s = 0; s < v.S; s++) vt += v.vol[(size_t)s * NS + base + p];
      const size_t wi = (size_t)cell * v.C * v.G + v.wgc[base + p];
      v.lnr[base + p] = weight_nc(v.wtype[wi], v.wmag[wi], v.wexp[wi], vol2rad(vt));
    }
    const double wv = rad2vol(D / 2.0) - rad2vol(D_dry / 2.0);
    v.vol[(size_t)iw * NS + base + p] = wv;
    if (nonflat) {
      double vt = 0.0;
      for (int s = 0; s < v.S; s++) vt += v.vol[(size_t)s * NS + base + p];
      v.gct_alt[base + p] = vol2rad(vt);
    }
  }
}
extern "C" int pmcgpu_condense_equilib_particles(pmcgpu_ctx* ctx) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_condense_equilib_particles");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  if (ctx->v.i_water < 1) return PMCGPU_ERR_ARG;
  k_condense_equilib<<<ctx->v.n_cells, 128, 0, ctx->stream>>>(ctx->v);
  LAUNCH_CHECK(ctx);
  if (ctx->any_nonflat) {  // aero_state_reweight (condense.F90:767), pass 1000
    int rc = pmc_reweight_async(ctx, ctx->rng_step, 1000u);
    if (rc) return rc;
  }
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

__global__ void k_eval_cond_rates(int n, const double* __restrict__ in, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* q = in + 9 * i;
  // T = T0 at t = 0; V_comp is passed directly: choose num_conc so that vc / num_conc = V_comp
  CondEnv e = cond_env(q[0], q[3], q[1], q[4], 0.0);
  const double num_conc = e.vc / q[5];
  double Ddot, Hd, a, b, c, dH;
  cond_rates<true>(e, q[2], q[6], q[7], q[8], num_conc, Ddot, Hd, a, b, c, dH);
  double* o = out + 6 * i;
  o[0] = Ddot; o[1] = Hd; o[2] = a; o[3] = b; o[4] = c; o[5] = dH;
}
extern "C" int pmcgpu_eval_condense_rates(pmcgpu_ctx* ctx, int n, const double* in, double* out) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n < 0 || !in || !out) return PMCGPU_ERR_ARG;
  double *di = nullptr, *dout = nullptr;
  CUDA_OK(ctx, cudaMalloc((void**)&di, (size_t)n * 9 * sizeof(double)));
  CUDA_OK(ctx, cudaMalloc((void**)&dout, (size_t)n * 6 * sizeof(double)));
  CUDA_OK(ctx, cudaMemcpyAsync(di, in, (size_t)n * 9 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  k_eval_cond_rates<<<(n + 127) / 128, 128, 0, ctx->stream>>>(n, di, dout);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaMemcpyAsync(out, dout, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(di);
  cudaFree(dout);
  return PMCGPU_OK;
}
