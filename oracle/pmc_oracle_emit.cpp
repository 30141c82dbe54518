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
// the sequential stream of new particle i_samp (radius attempts, then volume-fraction attempts).
// Deviation that changes particle ORDER only: every aero_state_scale_weight inside starts from a fresh
// aero_state_sort (the device keeps no lists between phases).
#include "pmc_oracle_domain.hpp"
#include <algorithm>
#include <cmath>

namespace pmco {

// aero_mode.F90:496-544
static double aero_mode_number(const AeroMode& m, const Weight& w) {
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
  if (m.type == MODE_LOG_NORMAL) {
    double x_mean_prime = std::log10(m.char_radius);
    if (w.type != WEIGHT_NONE)
      x_mean_prime = std::log10(m.char_radius) -
                     w.exponent * (m.log10_std_dev_radius * m.log10_std_dev_radius) * std::log(10.0);
    return std::pow(10.0, rand_normal_threshold(x_mean_prime, m.log10_std_dev_radius, prob_threshold, s));
  } else if (m.type == MODE_EXP) {
    if (w.type != WEIGHT_NONE) die_msg(678481276, "cannot use exponential modes with weighting");
    return sphere_vol2rad(-sphere_rad2vol(m.char_radius) * std::log(s.next()));
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
                                    bool allow_doubling, bool allow_halving, const RunCtx& rc) {
  const int n_group = st.awa.n_group;
  const int n_mode = (int)modes.size();
  int n_part_add = 0;
  std::vector<double> vols;
  for (int g = 1; g <= n_group; g++)
    for (int im = 1; im <= n_mode; im++) {
      const AeroMode& m = modes[im - 1];
      const int c = m.weight_class;
      const uint32_t a = (uint32_t)((g - 1) * n_mode + (im - 1));
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
        const double total_vol = sphere_rad2vol(radius);
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
        aero_state_add_particle(st, p, true);
      }
    }
  return n_part_add;
}

}  // namespace pmco

extern "C" {
// modes: type[n_mode], weight_class[n_mode], char_radius, log10_std, vol_frac[n_mode][S], vol_frac_std (may be NULL),
// num_conc[n_mode] (for this cell)
int pmco_add_aero_dist_sample(void* h, int cell, int n_mode, const int* type, const int* weight_class,
                              const double* char_radius, const double* log10_std, const double* vol_frac,
                              const double* vol_frac_std, const double* num_conc, double sample_prop,
                              double characteristic_factor, double create_time, int allow_doubling,
                              int allow_halving, unsigned step, int* n_part_add) {
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
}
