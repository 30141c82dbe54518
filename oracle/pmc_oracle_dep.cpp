// pmc_oracle_dep.cpp -- CPU restatement of the reference's dry deposition of particles
// (SURVEY.md section 8f, row N2).  TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
//
//   wrf_pmc_dep_dry_driver   interface/wrf_pmc_dep.F90:31-97
//   depvel_partmc            interface/wrf_pmc_dep.F90:188-213
//   dep_aero_dry_driver      interface/wrf_pmc_dep_aero.F90:31-150  (land-use tables of Zhang et al. 2001)
//   dry_dep_aero_state       :154-199, compute_dep_vel :204-249, compute_dep_prob :254-275,
//   aero_particles_remove_by_dep :280-296, calculate_vs :301-315, calculate_rs :320-394,
//   calculate_vd :399-412, slip_correction_factor :417-431, calculate_st :436-451,
//   compute_brownian_diff :456-470
//
// The reference mixes default-real literals into double expressions (1.257, .4, 1.1, 0.74, 0.001, 4.7,
// 1.0e-4, the data tables, gamma = 2.0/3.0 ...).  Those constants have single-precision VALUES in the
// compiled reference, so they are written here as (double)<float literal>: the result is what the
// Fortran computes, not what the formula on paper would give.
#include "pmc_oracle_domain.hpp"
#include <algorithm>
#include <cmath>

namespace pmco {

static const double AIR_DYN_VISC = 1.78e-5;  // constants.F90:65
static const double GRAV = 9.81;             // constants.F90:33

// wrf_pmc_dep_aero.F90:48-74: usgs land use -> zhang type
static const int ZHANG_FROM_USGS[25] = {15, 7, 7, 7, 7, 10, 6, 6, 6, 6, 4, 3, 2, 1, 5, 14, 11, 10, 8, 9, 10, 9, 8, 12, 8};
// :77-88 (data statements with default-real literals)
static const float ZH_ALPHA[15] = {1.0f, 0.6f, 1.1f, 0.8f, 0.8f, 1.2f, 1.2f, 50.0f, 50.0f, 1.3f, 2.0f, 50.0f, 100.0f, 100.0f, 1.5f};
// :90-106, A(iseason = 1, luc)
static const float ZH_A1[15] = {0.002f, 0.005f, 0.002f, 0.005f, 0.005f, 0.002f, 0.002f, 9.999f, 9.999f, 0.010f, 0.010f, 9.999f, 9.999f, 9.999f, 0.010f};

int dep_zhang_type(int ivgtyp) {  // :115-124 with modesurftype = 2
  int isurftype = 7;
  if (ivgtyp >= 1 && ivgtyp <= 25) isurftype = ZHANG_FROM_USGS[ivgtyp - 1];
  return isurftype;
}
double dep_alpha(int isurftype) { return (double)ZH_ALPHA[isurftype - 1]; }
double dep_A(int isurftype) { return (double)ZH_A1[isurftype - 1]; }

// wrf_pmc_dep.F90:188-213.  rmol is a WRF real.
double depvel_partmc(float rmol, double zr, double z0, double ustar) {
  const double vk = 0.40;
  if (std::fabs(rmol) < 1.0e-6f) rmol = 0.f;
  double polint;
  const double c074 = (double)0.74f, c0001 = (double)0.001f;
  if (rmol < 0.0f) {
    const double ta = std::pow(1.0 - 9.0 * zr * (double)rmol, 0.25) + c0001;
    const double to = std::pow(1.0 - 9.0 * z0 * (double)rmol, 0.25) + c0001;
    const double ar = ta * ta, ao = to * to;
    polint = c074 * (std::log((ar - 1.0) / (ar + 1.0)) - std::log((ao - 1.0) / (ao + 1.0)));
  } else if (rmol == 0.0f) {
    polint = c074 * std::log(zr / z0);
  } else {
    const float f = 4.7f * rmol;  // real * real
    polint = c074 * std::log(zr / z0) + (double)f * (zr - z0);
  }
  return polint / (vk * std::max(ustar, (double)1.0e-4f));
}

// :417-431
double dep_slip_correction_factor(double D_p) {
  const double lambda = .065e-6;
  return 1.0 + ((2.0 * lambda) / D_p) *
                   ((double)1.257f + (double).4f * std::exp(-((double)1.1f * D_p) / (2.0 * lambda)));
}
// :301-315
double dep_calculate_vs(double diameter, double density) {
  const double c_c = dep_slip_correction_factor(diameter);
  return ((diameter * diameter) * density * GRAV * c_c) / (18.0 * AIR_DYN_VISC);
}
// :456-470
static double compute_brownian_diff(double diameter, double temp) {
  const double C_c = dep_slip_correction_factor(diameter);
  return (Const::boltzmann * temp * C_c) / (3.0 * Const::pi * AIR_DYN_VISC * diameter);
}
// :436-451
static double calculate_st(double vs, double rrho, double ustar, double A) {
  if (A > 1.0) return (vs * (ustar * ustar)) / (AIR_DYN_VISC * rrho);
  return (vs * ustar) / (GRAV * A);
}
// :320-394 with updated_version = .true. (gamma of the caller is overwritten by 2.0/3.0 in single precision)
double dep_calculate_rs(double diameter, double vs, double temperature, double rrho, double ustar, double alpha,
                        double A) {
  const double diff = compute_brownian_diff(diameter, temperature);
  const double Sc = 1.45e-5 / diff;
  const double gamma = (double)(2.0f / 3.0f);
  const double C_b = (double).2f, nu = (double)0.8f, C_IM = (double)0.4f;
  const double St = calculate_st(vs, rrho, ustar, A);
  const double EB = C_b * std::pow(Sc, -gamma);
  const double q = St / (alpha + St);
  const double EIM = C_IM * (q * q);  // beta = 2.0d0
  const double EIN = .5 * std::pow(diameter / A, nu);
  const double R1 = std::exp(-std::sqrt(St));  // St**.5
  const double eps_0 = 3.0;
  return 1.0 / (eps_0 * ustar * (EB + EIM + EIN) * R1);
}
// :399-412
double dep_calculate_vd(double ra, double rs, double vs) {
  const double tmp = ra + rs + ra * rs * vs;
  return (1.0 / tmp) + vs;
}

// removal probability of one particle (compute_dep_vel + compute_dep_prob)
double dep_remove_prob(double diameter, double density, double temp, double rrho, double ustar, double aer_res,
                       int isurftype, double dt, double dz, double* vd_out) {
  const double vs = dep_calculate_vs(diameter, density);
  const double rs = dep_calculate_rs(diameter, vs, temp, rrho, ustar, dep_alpha(isurftype), dep_A(isurftype));
  const double vd = dep_calculate_vd(aer_res, rs, vs);
  if (vd_out) *vd_out = vd;
  return (dt / dz) * vd;
}

// dry_dep_aero_state (:154-199) for one surface cell
void dry_dep_aero_state(AeroState& st, const AeroData& d, const EnvState& env, double height, double aer_res,
                        double ustar, int isurftype, double dt, const RunCtx& rc) {
  const int n = st.n_part();
  std::vector<double> prob((size_t)n);
  for (int i = 0; i < n; i++) {
    const double diam = 2.0 * particle_radius(st.apa[i]);
    const double dens = particle_density(st.apa[i], d);
    prob[i] = dep_remove_prob(diam, dens, env.temp, env.inverse_density, ustar, aer_res, isurftype, dt, height, nullptr);
  }
  // aero_particles_remove_by_dep (:280-296): backwards, swap-with-last removal
  for (int i = n; i >= 1; i--) {
    const double u = rc.rng->uniform(rc.key(PH_DEP, 0u, (uint32_t)i), 0);
    if (u < prob[i - 1]) aero_state_remove_particle(st, i);
  }
}

}  // namespace pmco

extern "C" {

// wrf_pmc_dep_dry_driver for a patch of nx x nz x ny cells (WRF order: i fastest, then k, then j) whose
// local (1,1) is global (i_offset + 1, j_offset + 1).  2-D fields [nx*ny], i fastest.
int pmco_dep_dry_driver(void* h, int nx, int nz, int ny, int i_offset, int j_offset, int global_nx, int global_ny,
                        const float* rmol, const float* znt, const float* ust, const int* ivgtyp,
                        const double* height, double dt, unsigned step) {
  OracleDomain* D = (OracleDomain*)h;
  ORACLE_TRY
  const double zr = 2.0;
  for (int j = 1; j <= ny; j++)
    for (int i = 1; i <= nx; i++) {
      const int gi = i + i_offset, gj = j + j_offset;
      // dep_aero_dry_driver :113-114: the outermost ring of the global domain is skipped
      if (gi < 2 || gi > global_nx - 1 || gj < 2 || gj > global_ny - 1) continue;
      const size_t col = (size_t)(i - 1) + (size_t)nx * (j - 1);
      const size_t c = (size_t)(i - 1) + (size_t)nx * ((size_t)0 + (size_t)nz * (j - 1));
      const double ustar = (double)ust[col];
      const double aer_res = depvel_partmc(rmol[col], zr, (double)znt[col], ustar);
      RunCtx rc{D->rng.get(), (uint32_t)D->global_cell[c], step};
      D->cells[c].valid_sort = false;
      dry_dep_aero_state(D->cells[c], D->data, D->env[c], height[col], aer_res, ustar, dep_zhang_type(ivgtyp[col]), dt, rc);
    }
  ORACLE_CATCH(D)
}

// unit evaluation: in[n][10] = diameter, density, temp, rrho, ustar, rmol, znt, ivgtyp, dt, dz
//                  out[n][3] = aer_res, vd, remove_prob
int pmco_dep_eval(int n, const double* in, double* out) {
  for (int k = 0; k < n; k++) {
    const double* q = in + 10 * k;
    const double aer_res = depvel_partmc((float)q[5], 2.0, (double)(float)q[6], (double)(float)q[4]);
    double vd;
    const double p = dep_remove_prob(q[0], q[1], q[2], q[3], (double)(float)q[4], aer_res, dep_zhang_type((int)q[7]), q[8], q[9], &vd);
    out[3 * k + 0] = aer_res;
    out[3 * k + 1] = vd;
    out[3 * k + 2] = p;
  }
  return 0;
}
}
