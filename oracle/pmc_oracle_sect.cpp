// pmc_oracle_sect.cpp -- CPU restatement of the reference's SECTIONAL coagulation solver, the model the reference
// itself verifies mc_coag(Brownian) against (partmc/test/brownian: run_sect vs run_part at 0.2 / 0.3 / 0.4):
//   run_sect            partmc/src/run_sect.F90:56-198   (Bott's flux method on a logarithmic mass grid)
//   coad                :203-290
//   courant             :295-330
//   smooth_bin_kernel   :335-362
//   bin_kernel          partmc/src/coag_kernel.F90:216-246
//   num_conc_log_normal partmc/src/aero_mode.F90:131-160 (aero_binned_add_aero_dist, aero_binned.F90:227)
//   find_1d             partmc/src/util.F90:592-617
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).  One species, no emissions / dilution (test/brownian has none).
#include "pmc_oracle_domain.hpp"
#include <algorithm>
#include <cmath>

namespace pmco {

// util.F90:592-617
static int find_1d(int n, const double* x_vals, double x) {
  if (n == 0) return 0;
  int p = 1;
  while (x >= x_vals[p - 1]) {
    p = p + 1;
    if (p > n) break;
  }
  return p - 1;
}

// run_sect.F90:295-330 (1-based arrays kept as [i-1][j-1])
static void courant(int n_bin, double log_width, const std::vector<double>& e, std::vector<int>& ima,
                    std::vector<double>& c) {
  c.assign((size_t)n_bin * n_bin, 0.0);
  ima.assign((size_t)n_bin * n_bin, 0);
  for (int i = 1; i <= n_bin; i++)
    for (int j = i; j <= n_bin; j++) {
      const double x0 = e[i - 1] + e[j - 1];
      int k = find_1d(n_bin, e.data(), x0);
      if (k < n_bin) {
        k = k + 1;
        int kk;
        if (c[(size_t)(i - 1) * n_bin + (j - 1)] < 1.0 - 1e-08) {
          kk = k - 1;
          c[(size_t)(i - 1) * n_bin + (j - 1)] = std::log(x0 / e[k - 2]) / (3.0 * log_width);
        } else {
          c[(size_t)(i - 1) * n_bin + (j - 1)] = 0.0;
          kk = k;
        }
        ima[(size_t)(i - 1) * n_bin + (j - 1)] = std::min(n_bin - 1, kk);
      }
      c[(size_t)(j - 1) * n_bin + (i - 1)] = c[(size_t)(i - 1) * n_bin + (j - 1)];
      ima[(size_t)(j - 1) * n_bin + (i - 1)] = ima[(size_t)(i - 1) * n_bin + (j - 1)];
    }
}

// run_sect.F90:335-362
static void smooth_bin_kernel(int n_bin, const std::vector<double>& k, std::vector<double>& ks) {
  ks.assign((size_t)n_bin * n_bin, 0.0);
  auto K = [&](int i, int j) { return k[(size_t)(i - 1) * n_bin + (j - 1)]; };
  for (int i = 1; i <= n_bin; i++)
    for (int j = 1; j <= n_bin; j++) {
      const int jm = std::max(j - 1, 1), im = std::max(i - 1, 1);
      const int jp = std::min(j + 1, n_bin), ip = std::min(i + 1, n_bin);
      double v = 0.125 * (K(i, jm) + K(im, j) + K(ip, j) + K(i, jp)) + 0.5 * K(i, j);
      if (i == j) v = 0.5 * v;
      ks[(size_t)(i - 1) * n_bin + (j - 1)] = v;
    }
}

// run_sect.F90:203-290; g, ck, c, ima as in the Fortran (g is updated in place)
static void coad(int n_bin, std::vector<double>& g, const std::vector<double>& e, const std::vector<double>& ck,
                 const std::vector<double>& c, const std::vector<int>& ima) {
  const double gmin = 1e-60;
  int i0, i1;
  for (i0 = 1; i0 <= n_bin - 1; i0++)
    if (g[i0 - 1] > gmin) break;
  for (i1 = n_bin - 1; i1 >= 1; i1--)
    if (g[i1 - 1] > gmin) break;
  for (int i = i0; i <= i1; i++)
    for (int j = i; j <= i1; j++) {
      const int k = ima[(size_t)(i - 1) * n_bin + (j - 1)];  // k = 0: i + j goes nowhere
      const int kp = k + 1;
      double x0 = ck[(size_t)(i - 1) * n_bin + (j - 1)] * g[i - 1] * g[j - 1];
      x0 = std::min(x0, g[i - 1] * e[j - 1]);
      if (j != k) x0 = std::min(x0, g[j - 1] * e[i - 1]);
      const double gsi = x0 / e[j - 1];
      const double gsj = x0 / e[i - 1];
      const double gsk = gsi + gsj;
      g[i - 1] = g[i - 1] - gsi;
      g[j - 1] = g[j - 1] - gsj;
      if (k > 0) {
        const double gk = g[k - 1] + gsk;
        if (gk > gmin) {
          const double x1 = std::log(g[kp - 1] / gk + 1e-60);
          double flux = gsk / x1 * (std::exp(0.5 * x1) - std::exp(x1 * (0.5 - c[(size_t)(i - 1) * n_bin + (j - 1)])));
          flux = std::min(flux, gk);
          g[k - 1] = gk - flux;
          g[kp - 1] = g[kp - 1] + flux;
        }
      }
    }
}

// run_sect (run_sect.F90:56-198) for one log-normal mode of a one-species aerosol.  num_out / mass_out
// [n_out][n_bin]: number and mass concentration DENSITIES per ln(r) at t = 0, t_output, 2 t_output, ...
// (aero_binned%num_conc, vol_conc * density: what extract_sectional_aero_size prints).
int run_sect(int n_bin, double d_min, double d_max, double density, double total_num_conc, double geom_mean_diam,
             double log10_sigma, int kernel_type, double temp, double pressure, double t_max, double del_t,
             double t_output, const AeroData& data, double* num_out, double* mass_out, int n_out_cap) {
  BinGrid grid;
  bin_grid_make_log(grid, n_bin, d_min / 2.0, d_max / 2.0);
  const double log_width = std::log(grid.edges[1]) - std::log(grid.edges[0]);  // bin_grid%widths(1)
  std::vector<double> e((size_t)n_bin), vol_conc((size_t)n_bin), num_conc((size_t)n_bin);
  for (int i = 0; i < n_bin; i++) e[i] = sphere_rad2vol(grid.centers[i]) * density * 1e6;  // m^3 -> mg
  // aero_binned_add_aero_dist: num_conc_log_normal / vol_conc_log_normal (aero_mode.F90:131-188)
  const double rg = geom_mean_diam / 2.0;
  for (int k = 0; k < n_bin; k++) {
    num_conc[k] = total_num_conc / (std::sqrt(2.0 * Const::pi) * log10_sigma) *
                  std::exp(-std::pow(std::log10(grid.centers[k]) - std::log10(rg), 2.0) /
                           (2.0 * std::pow(log10_sigma, 2.0))) /
                  std::log(10.0);
    vol_conc[k] = num_conc[k] * sphere_rad2vol(grid.centers[k]);
  }
  std::vector<int> ima;
  std::vector<double> c, k_bin((size_t)n_bin * n_bin), ck;
  courant(n_bin, log_width, e, ima, c);
  // bin_kernel (coag_kernel.F90:216-246)
  EnvState env;
  env.temp = temp;
  env.pressure = pressure;
  for (int i = 0; i < n_bin; i++)
    for (int j = 0; j < n_bin; j++) {
      Particle p1, p2;
      p1.vol.assign(1, sphere_rad2vol(grid.centers[i]));
      p2.vol.assign(1, sphere_rad2vol(grid.centers[j]));
      k_bin[(size_t)i * n_bin + j] = kernel(kernel_type, p1, p2, data, env);
    }
  smooth_bin_kernel(n_bin, k_bin, ck);
  for (size_t q = 0; q < ck.size(); q++) ck[q] = ck[q] * 1e6;                 // m^3/s -> cm^3/s
  for (size_t q = 0; q < ck.size(); q++) ck[q] = ck[q] * del_t * log_width;   // x timestep x grid distance
  int n_out = 0;
  auto output = [&]() {
    if (n_out >= n_out_cap) return;
    for (int k = 0; k < n_bin; k++) {
      num_out[(size_t)n_out * n_bin + k] = num_conc[k];
      mass_out[(size_t)n_out * n_bin + k] = vol_conc[k] * density;
    }
    n_out++;
  };
  output();
  const int num_t = (int)std::lround(t_max / del_t);
  const int out_every = (int)std::lround(t_output / del_t);
  std::vector<double> g((size_t)n_bin);
  for (int i_time = 1; i_time <= num_t; i_time++) {
    for (int k = 0; k < n_bin; k++) g[k] = vol_conc[k] * density;
    coad(n_bin, g, e, ck, c, ima);
    for (int k = 0; k < n_bin; k++) {
      vol_conc[k] = g[k] / density;
      num_conc[k] = vol_conc[k] / sphere_rad2vol(grid.centers[k]);
    }
    if (out_every > 0 && i_time % out_every == 0) output();
  }
  return n_out;
}

}  // namespace pmco

extern "C" {
// returns the number of output rows written (t = 0, t_output, ...)
int pmco_run_sect(int n_bin, double d_min, double d_max, double density, double total_num_conc, double geom_mean_diam,
                  double log10_sigma, int kernel_type, double temp, double pressure, double t_max, double del_t,
                  double t_output, double* num_out, double* mass_out, int n_out_cap) {
  AeroData data;
  data.n_spec = 1;
  data.density.assign(1, density);
  data.kappa.assign(1, 0.0);
  data.molec_weight.assign(1, 18e-3);
  data.num_ions.assign(1, 0);
  data.i_water = 1;
  try {
    return run_sect(n_bin, d_min, d_max, density, total_num_conc, geom_mean_diam, log10_sigma, kernel_type, temp,
                    pressure, t_max, del_t, t_output, data, num_out, mass_out, n_out_cap);
  } catch (...) {
    return -1;
  }
}
}
