// pmc_oracle_emit.cpp -- CPU restatement of particle emission / inflow sampling:
//   aero_state_add_aero_dist_sample     partmc/src/aero_state.F90:758-831
//   aero_state_prepare_weight_for_add   partmc/src/aero_state.F90:703-757 (PMC_USE_WRF branch)
//   aero_mode_number                    partmc/src/aero_mode.F90:496-544
//   aero_mode_sample_radius             partmc/src/aero_mode.F90:549-628 (log-normal, exp, mono modes)
//   aero_mode_sample_vols               partmc/src/aero_mode.F90:633-668
//   rand_normal with prob_threshold     partmc/src/rand.F90:410-474 (in-tree Box-Muller branch)
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
//
// Draw addresses (PH_EMIT): a = (group - 1) n_mode + (mode - 1), b = 0: the Poisson count; b = i_samp >= 1:
// the sequential stream of new particle i_samp (radius attempts, then volume-fraction attempts).  `stream` is added
// to a: callers that sample in the same cell and step (scenario emissions, then sea salt) use different streams.
// Deviation that changes particle ORDER only: every aero_state_scale_weight inside starts from a fresh
// aero_state_sort (the device keeps no lists between phases).
#include "pmc_oracle_domain.hpp"
#include <algorithm>
#include <cmath>

namespace pmco {

// aero_mode_weighted_sampled_num_conc (aero_mode.F90:456-491)
static std::vector<double> weighted_sampled_num_conc(const AeroMode& m, const Weight& w) {
  std::vector<double> out(m.sample_num_conc.size());
  for (size_t i = 0; i < out.size(); i++) {
    if (w.type == WEIGHT_NONE) {
      out[i] = m.sample_num_conc[i];
    } else {
      const double x0 = std::log(m.sample_radius[i]), x1 = std::log(m.sample_radius[i + 1]);
      out[i] = m.sample_num_conc[i] / w.exponent * (std::exp(-w.exponent * x0) - std::exp(-w.exponent * x1)) / (x1 - x0);
    }
  }
  return out;
}
// aero_mode.F90:496-544
static double aero_mode_number(const AeroMode& m, const Weight& w) {
  if (m.type == MODE_SAMPLED) {  // :533-538
    double sum = 0.0;
    for (double x : weighted_sampled_num_conc(m, w)) sum = sum + x;
    return m.num_conc * (sum / w.magnitude);
  }
  if (m.type == MODE_LOG_NORMAL) {
    if (w.type == WEIGHT_NONE) return m.num_conc / w.magnitude;
    const double x_mean_prime =
        std::log10(m.char_radius) - w.exponent * (m.log10_std_dev_radius * m.log10_std_dev_radius) * std::log(10.0);
    const double l10 = std::log10(m.char_radius);
    return m.num_conc / w.magnitude *
           std::exp((x_mean_prime * x_mean_prime - l10 * l10) / (2.0 * (m.log10_std_dev_radius * m.log10_std_dev_radius)));
  } else if (m.type == MODE_EXP) {
    if (w.type != WEIGHT_NONE) die_msg(822252601, "cannot use exponential modes with weighting");
    return m.num_conc / w.magnitude;
  } else if (m.type == MODE_MONO) {
    return m.num_conc / weight_num_conc_at_radius(w, m.char_radius);
  }
  die_msg(901140225, "unknown aero_mode type");
  return 0.0;
}

// rand.F90:410-474 with prob_threshold: redraw while the sample is further out in the tails than the threshold
static double rand_normal_threshold(double mean, double stddev, double prob_threshold, Stream& s) {
  const double sqrt2 = (double)std::sqrt(2.0f);  // 2.0**.5 is a default-real expression
  for (;;) {
    const double u1 = s.next();
    const double u2 = s.next();
    const double r = std::sqrt(-2.0 * std::log(u1));
    const double theta = 2.0 * Const::pi * u2;
    const double z0 = r * std::cos(theta);
    if (1.0 - std::fabs(std::erf(z0 / sqrt2)) >= prob_threshold) return stddev * z0 + mean;
  }
}

// aero_mode.F90:549-628
double aero_mode_sample_radius(const AeroMode& m, const Weight& w, double prob_threshold, Stream& s) {
  if (m.type == MODE_SAMPLED) {  // :581-608
    const std::vector<double> pdf = weighted_sampled_num_conc(m, w);
    const int i_sample = sample_cts_pdf(pdf, s);
    const double r = s.next();
    if (w.type == WEIGHT_NONE || w.exponent == 0.0) {
      const double x0 = std::log(m.sample_radius[i_sample - 1]), x1 = std::log(m.sample_radius[i_sample]);
      return std::exp((1.0 - r) * x0 + r * x1);
    }
    const double inv_nc0 = 1.0 / weight_num_conc_at_radius(w, m.sample_radius[i_sample - 1]);
    const double inv_nc1 = 1.0 / weight_num_conc_at_radius(w, m.sample_radius[i_sample]);
    const double inv_nc = (1.0 - r) * inv_nc0 + r * inv_nc1;
    return std::exp(std::log((1.0 / inv_nc) / w.magnitude) / w.exponent);  // aero_weight_radius_at_num_conc
  }
  if (m.type == MODE_LOG_NORMAL) {
    double x_mean_prime = std::log10(m.char_radius);
    if (w.type != WEIGHT_NONE)
      x_mean_prime = std::log10(m.char_radius) -
                     w.exponent * (m.log10_std_dev_radius * m.log10_std_dev_radius) * std::log(10.0);
    return std::pow(10.0, rand_normal_threshold(x_mean_prime, m.log10_std_dev_radius, prob_threshold, s));
  } else if (m.type == MODE_EXP) {
    if (w.type != WEIGHT_NONE) die_msg(678481276, "cannot use exponential modes with weighting");
    return aero_vol2rad(-aero_rad2vol(m.char_radius) * std::log(s.next()));  // aero_mode.F90:611
  }
  return m.char_radius;
}

// aero_mode.F90:633-668: accept-reject on non-negative volumes after projecting out the mean direction
void aero_mode_sample_vols(const AeroMode& m, double total_vol, std::vector<double>& vols, Stream& s) {
  const size_t S = m.vol_frac.size();
  vols.assign(S, 0.0);
  for (int i_sample = 1; i_sample <= 1000000; i_sample++) {
    for (size_t k = 0; k < S; k++) vols[k] = rand_normal(m.vol_frac[k], m.vol_frac_std[k], s);
    double sum = 0.0;
    for (size_t k = 0; k < S; k++) sum += vols[k];
    const double offset = 1.0 - sum;
    double mn = 1e300;
    for (size_t k = 0; k < S; k++) {
      vols[k] = vols[k] + offset * m.vol_frac[k];
      mn = std::min(mn, vols[k]);
    }
    if (mn >= 0.0) break;
    if (i_sample == 1000000) die_msg(549015143, "Unable to sample non-negative volumes");
  }
  double sum = 0.0;
  for (size_t k = 0; k < S; k++) sum += vols[k];
  for (size_t k = 0; k < S; k++) vols[k] = vols[k] / sum * total_vol;  // :663
}

// aero_state.F90:703-757, PMC_USE_WRF branch (per-cell counts, no division by the process count)
static void aero_state_prepare_weight_for_add(AeroState& st, int g, int c, double n_add, bool allow_doubling,
                                              bool allow_halving, const RunCtx& rc, uint32_t salt) {
  st.valid_sort = false;  // see the header comment
  const double mean_n_part = (double)aero_state_total_particles(st, g, c);
  const double n_part_new = mean_n_part + n_add;
  if (n_part_new == 0.0) return;
  const double ideal = st.ideal(g, c);
  if ((n_part_new < ideal / 2.0) || (n_part_new > ideal * 2.0)) {
    const double weight_ratio = n_part_new / ideal;
    aero_state_scale_weight(st, g, c, weight_ratio, allow_doubling, allow_halving, rc, salt);
  }
}

// aero_state.F90:758-831
int aero_state_add_aero_dist_sample(AeroState& st, const AeroData& d, const std::vector<AeroMode>& modes,
                                    double sample_prop, double characteristic_factor, double create_time,
                                    bool allow_doubling, bool allow_halving, const RunCtx& rc, uint32_t stream) {
  const int n_group = st.awa.n_group;
  const int n_mode = (int)modes.size();
  int n_part_add = 0;
  std::vector<double> vols;
  for (int g = 1; g <= n_group; g++)
    for (int im = 1; im <= n_mode; im++) {
      const AeroMode& m = modes[im - 1];
      const int c = m.weight_class;
      const uint32_t a = stream + (uint32_t)((g - 1) * n_mode + (im - 1));
      double n_samp_avg = sample_prop * aero_mode_number(m, st.awa.at(g, c));
      aero_state_prepare_weight_for_add(st, g, c, n_samp_avg, allow_doubling, allow_halving, rc, 1u + a);
      if (n_samp_avg == 0.0) continue;
      n_samp_avg = sample_prop * aero_mode_number(m, st.awa.at(g, c));
      Stream s0(rc.rng, rc.key(PH_EMIT, a, 0u));
      const int n_samp = rand_poisson(n_samp_avg, s0);
      const double size_factor = std::min(1.0 / (characteristic_factor * n_samp_avg), 0.32);
      n_part_add += n_samp;
      for (int i_samp = 1; i_samp <= n_samp; i_samp++) {
        Stream s(rc.rng, rc.key(PH_EMIT, a, (uint32_t)i_samp));
        const double radius = aero_mode_sample_radius(m, st.awa.at(g, c), size_factor, s);
        const double total_vol = aero_rad2vol(radius);  // aero_state.F90:817
        aero_mode_sample_vols(m, total_vol, vols, s);
        Particle p;  // aero_particle_zero
        p.vol = vols;
        p.id = st.idgen.new_id();
        p.weight_group = g;
        p.weight_class = c;
        p.least_create_time = create_time;  // aero_particle_set_component
        p.greatest_create_time = create_time;
        p.n_primary_parts = 1;
        p.water_hyst_leg = 0;
        p.comp_src.assign(1, c);  // aero_particle_set_component (:823); source = weight class (one per source in WRF)
        p.comp_time.assign(1, create_time);
        aero_state_add_particle(st, p, true);
      }
    }
  return n_part_add;
}

// ---- sea salt (interface/wrf_pmc_driver.F90:1767-1924).  The Fortran declares its tables as real(kind=dp) arrays
// initialised with DEFAULT-REAL literals, so every tabulated number is the single-precision value widened to double;
// `wind_speed**3.41` has a single-precision exponent as well.  match_mam3_modes is a compile-time .true. (:1816), so
// the 25-bin sampled Gong spectrum (seasalt_emissions_1bin) is dead code in the reference and not restated.
static const int SEASALT_STREAM = 1024;  // keeps the draws apart from scenario emissions in the same cell and step
int seasalt_n_modes(int sea_salt_opt) { return sea_salt_opt == 1 ? 2 : (sea_salt_opt == 2 ? 5 : 0); }
void seasalt_modes(int sea_salt_opt, int n_spec, int i_na, int i_cl, const int* weight_class,
                   std::vector<AeroMode>& modes) {
  const double mw_na = 22.990, mw_cl = 35.450;
  const int n_modes = seasalt_n_modes(sea_salt_opt);
  if (n_modes == 0) die_msg(1767, "invalid sea salt option");
  modes.assign((size_t)n_modes, AeroMode());
  for (int m = 0; m < n_modes; m++) {
    AeroMode& md = modes[m];
    md.type = MODE_LOG_NORMAL;
    md.vol_frac.assign((size_t)n_spec, 0.0);
    md.vol_frac_std.assign((size_t)n_spec, 0.0);
    md.vol_frac[i_na - 1] = mw_na / (mw_na + mw_cl);
    md.vol_frac[i_cl - 1] = 1.0 - (mw_na / (mw_na + mw_cl));
    md.weight_class = weight_class[m];
    if (sea_salt_opt == 1) {  // Gong (2003) flux matched to the two MAM3 modes (:1826-1848)
      static const float flux[2] = {73.7095947f, 10.1706810f};
      static const float diam[2] = {2.03347408e-07f, 1.09840811e-06f};
      static const float sd[2] = {2.0f, 2.2f};
      md.char_radius = (double)diam[m] / 2;
      md.log10_std_dev_radius = std::log10((double)sd[m]);
      md.num_conc = (double)flux[m];
    } else {  // Ovadnevaite et al. (2014), five modes (:1880-1909); Re is fixed to 3.1e6 by the reference
      static const float sigma[5] = {1.37f, 1.5f, 1.42f, 1.53f, 1.85f};
      static const float CMD[5] = {.018f, .041f, .09f, .23f, .83f};
      static const float coef[5] = {104.5f, .0442f, 149.6f, 2.96f, .51f};
      static const double offset[5] = {1e5, 1e5, 1e5, 1e5, 2e5};
      static const float expon[5] = {.556f, 1.08f, .545f, .79f, .87f};
      md.char_radius = .5 * (double)CMD[m] * 1e-6;
      md.log10_std_dev_radius = std::log10((double)sigma[m]);
      const double Re = 3.1e6;
      md.num_conc = (double)coef[m] * std::pow(Re - offset[m], (double)expon[m]);
    }
  }
}
int seasalt_emissions(AeroState& st, const AeroData& d, double wind_speed, double height, double elapsed_time,
                      double delta_t, bool allow_doubling, bool allow_halving, int sea_salt_opt, int i_na, int i_cl,
                      const int* weight_class, const RunCtx& rc) {
  std::vector<AeroMode> modes;
  seasalt_modes(sea_salt_opt, d.n_spec, i_na, i_cl, weight_class, modes);
  const double emission_rate_scale = (sea_salt_opt == 1) ? std::pow(wind_speed, (double)3.41f) : 1.0;
  const double p = emission_rate_scale * delta_t / height;
  const double sample_timescale = 3600.0;
  const double effective = std::max(1.0, std::min(sample_timescale, elapsed_time));
  const double characteristic_factor = effective / delta_t;
  return aero_state_add_aero_dist_sample(st, d, modes, p, characteristic_factor, elapsed_time, allow_doubling,
                                         allow_halving, rc, SEASALT_STREAM);
}

}  // namespace pmco

extern "C" {
// modes: type[n_mode], weight_class[n_mode], char_radius, log10_std, vol_frac[n_mode][S], vol_frac_std (may be NULL),
// num_conc[n_mode] (for this cell)
int pmco_add_aero_dist_sample(void* h, int cell, int n_mode, const int* type, const int* weight_class,
                              const double* char_radius, const double* log10_std, const double* vol_frac,
                              const double* vol_frac_std, const double* num_conc, double sample_prop,
                              double characteristic_factor, double create_time, int allow_doubling,
                              int allow_halving, unsigned step, int* n_part_add, int n_sample_bins,
                              const double* sample_radius, const double* sample_num_conc) {
  OracleDomain* D = (OracleDomain*)h;
  ORACLE_TRY
  const int S = D->data.n_spec;
  std::vector<AeroMode> modes((size_t)n_mode);
  for (int m = 0; m < n_mode; m++) {
    modes[m].type = type[m];
    modes[m].weight_class = weight_class[m];
    modes[m].char_radius = char_radius[m];
    modes[m].log10_std_dev_radius = log10_std[m];
    modes[m].num_conc = num_conc[m];
    if (type[m] == MODE_SAMPLED) {
      modes[m].sample_radius.assign(sample_radius + (size_t)m * (n_sample_bins + 1),
                                    sample_radius + (size_t)(m + 1) * (n_sample_bins + 1));
      modes[m].sample_num_conc.assign(sample_num_conc + (size_t)m * n_sample_bins,
                                      sample_num_conc + (size_t)(m + 1) * n_sample_bins);
    }
    modes[m].vol_frac.assign(vol_frac + (size_t)m * S, vol_frac + (size_t)(m + 1) * S);
    if (vol_frac_std)
      modes[m].vol_frac_std.assign(vol_frac_std + (size_t)m * S, vol_frac_std + (size_t)(m + 1) * S);
    else
      modes[m].vol_frac_std.assign((size_t)S, 0.0);
  }
  RunCtx rc{D->rng.get(), (uint32_t)D->global_cell[cell], step};
  const int n = aero_state_add_aero_dist_sample(D->cells[cell], D->data, modes, sample_prop, characteristic_factor,
                                                create_time, allow_doubling != 0, allow_halving != 0, rc);
  if (n_part_add) *n_part_add = n;
  ORACLE_CATCH(D)
}

// partmc_timestep's sea salt block (wrf_pmc_driver.F90:212-223) for the cells of this oracle: lowest level (k == 1),
// xland > 1.5 (water); wind = (u10^2 + v10^2)**.5 in single precision.  u10, v10, xland [ny][nx] float,
// height [ny][nx] double; cell index = i + nx (k + nz j).  n_add [n_cells] (may be NULL).
int pmco_seasalt_emissions(void* h, int nx, int nz, int ny, const float* u10, const float* v10, const float* xland,
                           const double* height, int sea_salt_opt, int i_na, int i_cl, const int* weight_class,
                           double dt, double elapsed_time, int allow_doubling, int allow_halving, unsigned step,
                           int* n_add) {
  OracleDomain* D = (OracleDomain*)h;
  ORACLE_TRY
  if (n_add)
    for (int c = 0; c < nx * nz * ny; c++) n_add[c] = 0;
  for (int j = 0; j < ny; j++)
    for (int i = 0; i < nx; i++) {
      if (!(xland[j * nx + i] > 1.5f)) continue;
      const float u = u10[j * nx + i], v = v10[j * nx + i];
      const float w2 = u * u + v * v;
      const double wind_speed = (double)std::sqrt(w2);  // **.5 of a default real
      const int cell = i + nx * (0 + nz * j);
      RunCtx rc{D->rng.get(), (uint32_t)D->global_cell[cell], step};
      const int n = seasalt_emissions(D->cells[cell], D->data, wind_speed, height[j * nx + i], elapsed_time, dt,
                                      allow_doubling != 0, allow_halving != 0, sea_salt_opt, i_na, i_cl, weight_class, rc);
      if (n_add) n_add[cell] = n;
    }
  ORACLE_CATCH(D)
}
// the mode table itself (char_radius, log10 sigma, flux, vol_frac [n_modes][n_spec]) for known-answer tests
int pmco_seasalt_modes(int sea_salt_opt, int n_spec, int i_na, int i_cl, double* char_radius, double* log10_std,
                       double* flux, double* vol_frac) {
  std::vector<AeroMode> modes;
  std::vector<int> wc((size_t)5, 1);
  try {
    seasalt_modes(sea_salt_opt, n_spec, i_na, i_cl, wc.data(), modes);
  } catch (...) {
    return -1;
  }
  for (size_t m = 0; m < modes.size(); m++) {
    char_radius[m] = modes[m].char_radius;
    log10_std[m] = modes[m].log10_std_dev_radius;
    flux[m] = modes[m].num_conc;
    for (int k = 0; k < n_spec; k++) vol_frac[m * n_spec + k] = modes[m].vol_frac[k];
  }
  return (int)modes.size();
}
}
