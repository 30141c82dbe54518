// pmc_oracle.cpp -- CPU restatement of the PartMC per-cell aerosol step.
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp header).  Compile with
// -O2 -ffp-contract=off so that no multiply-add is fused, like the default
// gfortran x86-64 build of the reference.
#include "pmc_oracle.hpp"
#include <algorithm>
#include <cstdio>
#include <cstring>

namespace pmco {

void die_msg(int code, const std::string& msg) { throw PmcError(code, msg); }

// ======================================================================= RNG
static inline void philox_round(uint32_t c[4], const uint32_t k[2]) {
  const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
  const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
  const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
  const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
  const uint32_t n0 = hi1 ^ c[1] ^ k[0];
  const uint32_t n1 = lo1;
  const uint32_t n2 = hi0 ^ c[3] ^ k[1];
  const uint32_t n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c[4] = {ctr[0], ctr[1], ctr[2], ctr[3]};
  uint32_t k[2] = {key[0], key[1]};
  for (int r = 0; r < 10; r++) {
    philox_round(c, k);
    k[0] += 0x9E3779B9u;
    k[1] += 0xBB67AE85u;
  }
  out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
}
// Addressing: counter = (j>>1, b, a, cell), key = (seed, step*16 + phase);
// the two 53-bit uniforms of one block serve j even / odd.
double philox_uniform(uint32_t seed, const RngKey& k, uint32_t j) {
  uint32_t ctr[4] = {j >> 1, k.b, k.a, k.cell};
  uint32_t key[2] = {seed, k.step * 16u + k.phase};
  uint32_t o[4];
  philox4x32_10(ctr, key, o);
  const uint32_t hi = (j & 1u) ? o[2] : o[0];
  const uint32_t lo = (j & 1u) ? o[3] : o[1];
  const uint64_t m = ((uint64_t)(hi >> 5) << 26) | (uint64_t)(lo >> 6);
  return (double)m * (1.0 / 9007199254740992.0);
}

RngMT::RngMT(uint32_t s) {
  // gsl mt19937 seeding (== the 2002 reference init_genrand); seed 0 -> 4357
  if (s == 0) s = 4357;
  mt[0] = s;
  for (mti = 1; mti < 624; mti++) mt[mti] = 1812433253u * (mt[mti - 1] ^ (mt[mti - 1] >> 30)) + (uint32_t)mti;
}
uint32_t RngMT::next_u32() {
  if (mti >= 624) {
    int kk;
    for (kk = 0; kk < 624 - 397; kk++) {
      uint32_t y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
      mt[kk] = mt[kk + 397] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    for (; kk < 623; kk++) {
      uint32_t y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
      mt[kk] = mt[kk + (397 - 624)] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    uint32_t y = (mt[623] & 0x80000000u) | (mt[0] & 0x7fffffffu);
    mt[623] = mt[396] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    mti = 0;
  }
  uint32_t y = mt[mti++];
  y ^= (y >> 11);
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= (y >> 18);
  return y;
}

// rand.F90:172-205 (non-GSL branch): mod(int(u * n), n) + 1
int pmc_rand_int(double u, int n) {
  pmc_assert(669532625, n >= 1);
  int v = (int)(u * (double)n);
  return (v % n) + 1;
}
// rand.F90:214-226
int prob_round(double val, double u) {
  if (u < std::ceil(val) - val) return (int)std::floor(val);
  return (int)std::ceil(val);
}
// rand.F90:410-474, Box-Muller branch (z1 discarded)
double rand_normal(double mean, double stddev, Stream& s) {
  pmc_assert(898978929, stddev >= 0.0);
  const double u1 = s.next();
  const double u2 = s.next();
  const double r = std::sqrt(-2.0 * std::log(u1));
  const double theta = 2.0 * Const::pi * u2;
  const double z0 = r * std::cos(theta);
  return stddev * z0 + mean;
}
static inline int f_nint(double x) { return (int)std::llround(x); }  // Fortran nint: half away from zero
// rand.F90:252-304, non-GSL branch
int rand_poisson(double mean, Stream& s) {
  pmc_assert(368397056, mean >= 0.0);
  if (mean <= 10.0) {
    const double L = std::exp(-mean);
    int k = 0;
    double p = 1.0;
    do {
      k = k + 1;
      p = p * s.next();
    } while (!(p < L));
    return k - 1;
  }
  const int k = f_nint(rand_normal(mean, std::sqrt(mean), s));
  return std::max(k, 0);
}
// rand.F90:319-404, non-GSL branch
int rand_binomial(int n, double p, Stream& s) {
  pmc_assert(130699849, n >= 0);
  pmc_assert(754379796, p >= 0.0);
  pmc_assert(678506029, p <= 1.0);
  const double np = (double)n * p;
  const double nomp = (double)n * (1.0 - p);
  if (np >= 10.0 && nomp >= 10.0) {
    const int k = f_nint(rand_normal(np, std::sqrt(np * (1.0 - p)), s));
    return std::min(std::max(k, 0), n);
  } else if (p < 1e-15) {
    return 0;
  } else if (np < 1e-200) {
    return 0;
  } else if (nomp < 1e-200) {
    return n;
  }
  const double q = (p <= 0.5) ? p : 1.0 - p;
  int k = 0, sum = 0;
  for (;;) {
    const double G_real = std::log(s.next()) / std::log(1.0 - q);
    if (G_real > (double)(n - sum)) break;
    const int G = (int)std::ceil(G_real);
    sum = sum + G;
    if (sum > n) break;
    k = k + 1;
  }
  const int r = (p <= 0.5) ? k : n - k;
  pmc_assert(359087410, r <= n);
  return r;
}
// rand.F90:550-575
int sample_cts_pdf(const std::vector<double>& pdf, Stream& s) {
  double pdf_max = pdf[0], pdf_min = pdf[0];
  for (double v : pdf) { pdf_max = std::max(pdf_max, v); pdf_min = std::min(pdf_min, v); }
  if (pdf_min < 0.0) die_msg(121838078, "pdf contains negative values");
  if (pdf_max <= 0.0) die_msg(119208863, "pdf is not positive");
  for (;;) {
    const int k = pmc_rand_int(s.next(), (int)pdf.size());
    if (s.next() < pdf[k - 1] / pdf_max) return k;
  }
}
// test/rand/test_poisson_sample.F90: pmf = exp(-mean) * mean^k / k!
double poisson_pmf(double mean, int k) {
  return std::exp(-mean + k * std::log(mean) - std::lgamma(k + 1.0));
}
double binomial_pmf(int n, double p, int k) {
  return std::exp(std::lgamma(n + 1.0) - std::lgamma(k + 1.0) - std::lgamma(n - k + 1.0) + k * std::log(p) +
                  (n - k) * std::log(1.0 - p));
}

// ======================================================================= util
double sphere_vol2rad(double v) { return std::pow(3.0 * v / 4.0 / Const::pi, 1.0 / 3.0); }
double sphere_rad2vol(double r) { return 4.0 * Const::pi * (r * r * r) / 3.0; }

// ======================================================================= fractal.F90 (geometric radius only)
static Fractal g_fractal;
void set_fractal(double frac_dim, double prime_radius, double vol_fill_factor) {
  g_fractal.frac_dim = frac_dim;
  g_fractal.prime_radius = prime_radius;
  g_fractal.vol_fill_factor = vol_fill_factor;
}
const Fractal& fractal() { return g_fractal; }
bool fractal_is_spherical() { return g_fractal.frac_dim == 3.0 && g_fractal.vol_fill_factor == 1.0; }
double fractal_vol_to_num_of_monomers(double v) {
  const double x = sphere_vol2rad(v) / g_fractal.prime_radius;
  return x * x * x;  // **3
}
double aero_vol2rad(double v) {
  if (fractal_is_spherical()) return sphere_vol2rad(v);
  return g_fractal.prime_radius *
         std::pow(fractal_vol_to_num_of_monomers(v) * g_fractal.vol_fill_factor, 1.0 / g_fractal.frac_dim);
}
double aero_rad2vol(double r) {
  if (fractal_is_spherical()) return sphere_rad2vol(r);
  return sphere_rad2vol(g_fractal.prime_radius) * std::pow(r / g_fractal.prime_radius, g_fractal.frac_dim) /
         g_fractal.vol_fill_factor;
}
static void require_spherical(const char* what) {
  if (!fractal_is_spherical()) die_msg(0, std::string(what) + ": the mobility radius of fractal particles is not restated");
}

std::vector<double> linspace(double min_x, double max_x, int n) {
  std::vector<double> x((size_t)std::max(n, 0));
  pmc_assert(999299119, n >= 0);
  for (int i = 2; i <= n - 1; i++) {
    const double a = (double)(i - 1) / (double)(n - 1);
    x[i - 1] = (1.0 - a) * min_x + a * max_x;
  }
  if (n > 0) {
    x[0] = min_x;
    x[n - 1] = max_x;
  }
  return x;
}
std::vector<double> logspace(double min_x, double max_x, int n) {
  pmc_assert(804623592, n >= 0);
  if (n == 0) return {};
  pmc_assert(548290438, min_x > 0.0);
  pmc_assert(805259035, max_x > 0.0);
  std::vector<double> x = linspace(std::log(min_x), std::log(max_x), n);
  for (auto& v : x) v = std::exp(v);
  x[0] = min_x;
  x[n - 1] = max_x;
  return x;
}
int linspace_find(double min_x, double max_x, int n, double x) {
  if (x == max_x) return n - 1;
  const double f = std::floor((x - min_x) / (max_x - min_x) * (double)(n - 1));
  // guard the int conversion (Fortran floor() of a huge value is undefined)
  long long v = (f > 2.0e9) ? 2000000000LL : (f < -2.0e9 ? -2000000000LL : (long long)f);
  v = v + 1;
  v = std::min<long long>(v, n);
  v = std::max<long long>(v, 0);
  return (int)v;
}
int logspace_find(double min_x, double max_x, int n, double x) {
  return linspace_find(std::log(min_x), std::log(max_x), n, std::log(x));
}
double interp_linear_disc(double x_1, double x_n, int n, int i) {
  if (x_1 == x_n) return x_1;
  const double alpha = (double)(i - 1) / (double)(n - 1);
  return (1.0 - alpha) * x_1 + alpha * x_n;
}
double student_t_95_coeff(int n_sample) {
  const double limit = 1.959963984540054;
  static const double values[15] = {12.7062047364, 4.30265272991, 3.18244630528, 2.7764451052,  2.57058183661,
                                    2.44691184879, 2.36462425101, 2.30600413503, 2.26215716274, 2.22813885196,
                                    2.20098516008, 2.17881282966, 2.16036865646, 2.14478668792, 2.13144954556};
  const int n_dof = n_sample - 1;
  pmc_assert(359779741, n_dof >= 1);
  if (n_dof <= 15) return values[n_dof - 1];
  if (n_dof <= 20) return (2.11990529922 - limit) * 16.0 / (double)n_dof + limit;
  return (2.07961384473 - limit) * 21.0 / (double)n_dof + limit;
}

// ======================================================================= bin grid
void bin_grid_make_log(BinGrid& g, int n_bin, double min, double max) {
  pmc_assert(538534122, n_bin >= 0);
  g.edges.clear();
  g.centers.clear();
  if (n_bin > 0) {
    pmc_assert(966541762, min > 0.0);
    pmc_assert(711537859, min < max);
  }
  if (n_bin == 0) return;
  g.edges = logspace(min, max, n_bin + 1);
  const double c1 = std::exp(interp_linear_disc(std::log(min), std::log(max), 2 * n_bin + 1, 2));
  const double c2 = std::exp(interp_linear_disc(std::log(min), std::log(max), 2 * n_bin + 1, 2 * n_bin));
  g.centers = logspace(c1, c2, n_bin);
}
int bin_grid_find(const BinGrid& g, double val) {
  if (g.size() <= 0) return -1;
  return logspace_find(g.edges[0], g.edges[g.size()], g.size() + 1, val);
}

// ======================================================================= weights
double weight_num_conc_at_radius(const Weight& w, double radius) {
  if (w.type == WEIGHT_NONE) return w.magnitude;
  if (w.type == WEIGHT_POWER || w.type == WEIGHT_MFA) return w.magnitude * std::pow(radius, w.exponent);
  die_msg(700421478, "unknown aero_weight type");
  return 0.0;
}
double awa_num_conc_at_radius(const WeightArray& a, int i_class, double radius) {
  double s = 0.0;
  for (int g = 1; g <= a.n_group; g++) s += 1.0 / weight_num_conc_at_radius(a.at(g, i_class), radius);
  return 1.0 / s;
}
void awa_check_monotonicity(const WeightArray& a, bool& inc, bool& dec) {
  inc = true;
  dec = true;
  for (const Weight& w : a.w) {
    if (w.type == WEIGHT_POWER) {
      if (w.exponent < 0.0) inc = false;
      if (w.exponent > 0.0) dec = false;
    }
    if (w.type == WEIGHT_MFA) inc = false;
  }
}
void awa_minmax_num_conc(const WeightArray& a, int i_class, double r1, double r2, double& mn, double& mx) {
  bool inc, dec;
  awa_check_monotonicity(a, inc, dec);
  pmc_assert(857727714, inc || dec);
  const double n1 = awa_num_conc_at_radius(a, i_class, r1);
  const double n2 = awa_num_conc_at_radius(a, i_class, r2);
  mn = std::min(n1, n2);
  mx = std::max(n1, n2);
}
int awa_rand_group(const WeightArray& a, int i_class, double radius, Stream& s) {
  std::vector<double> comp_vols((size_t)a.n_group);
  for (int g = 1; g <= a.n_group; g++) comp_vols[g - 1] = 1.0 / weight_num_conc_at_radius(a.at(g, i_class), radius);
  return sample_cts_pdf(comp_vols, s);
}

// ======================================================================= particle
double particle_volume(const Particle& p) {
  double s = 0.0;
  for (double v : p.vol) s += v;
  return s;
}
double particle_radius(const Particle& p) { return aero_vol2rad(particle_volume(p)); }
double particle_mass(const Particle& p, const AeroData& d) {
  double s = 0.0;
  for (int i = 0; i < d.n_spec; i++) s += p.vol[i] * d.density[i];
  return s;
}
double particle_density(const Particle& p, const AeroData& d) { return particle_mass(p, d) / particle_volume(p); }
static double total_solute_quantity(const Particle& p, const AeroData& d, const double* q) {
  double total = 0.0;
  for (int i = 1; i <= d.n_spec; i++)
    if (i != d.i_water) total = total + p.vol[i - 1] * (q ? q[i - 1] : 1.0);
  return total;
}
double particle_solute_volume(const Particle& p, const AeroData& d) { return total_solute_quantity(p, d, nullptr); }
double particle_solute_kappa(const Particle& p, const AeroData& d) {
  std::vector<double> kappa((size_t)d.n_spec);
  const double M_w = d.i_water ? d.molec_weight[d.i_water - 1] : Const::water_molec_weight;
  const double rho_w = d.i_water ? d.density[d.i_water - 1] : Const::water_density;
  for (int i = 1; i <= d.n_spec; i++) {
    if (i == d.i_water) {
      kappa[i - 1] = 0.0;
    } else if (d.num_ions[i - 1] > 0) {
      pmc_assert(123681459, d.kappa[i - 1] == 0.0);
      const double M_a = d.molec_weight[i - 1];
      const double rho_a = d.density[i - 1];
      kappa[i - 1] = M_w * rho_a / (M_a * rho_w) * (double)d.num_ions[i - 1];
    } else {
      kappa[i - 1] = d.kappa[i - 1];
    }
  }
  return total_solute_quantity(p, d, kappa.data()) / total_solute_quantity(p, d, nullptr);
}
// rand.F90:515-542 (selection sampling; 1-based indices in ascending order)
std::vector<int> sample_without_replacement(int sample_size, int population_size, Stream& s) {
  std::vector<int> sample((size_t)std::max(sample_size, 0));
  int t = 0, m = 0;
  while (m < sample_size && t < 64) {  // (the guard only matters for sample_size > population_size: out of bounds upstream)
    const double r = s.next();
    if ((double)(population_size - t) * r >= (double)(sample_size - m)) {
      t = t + 1;
    } else {
      t = t + 1;
      m = m + 1;
      sample[m - 1] = t;
    }
  }
  sample.resize((size_t)m);
  return sample;
}
void particle_coagulate(const Particle& p1, const Particle& p2, Particle& pn, Rng* rng, RngKey ckey, uint32_t j0) {
  pmc_assert(203741686, p1.vol.size() == p2.vol.size());
  // components (:940-972): concatenate, or keep MAX_AERO_COMPONENT_SIZE split in proportion to n_primary_parts
  std::vector<int> csrc;
  std::vector<double> ctime;
  {
    const int n1 = (int)p1.comp_src.size(), n2 = (int)p2.comp_src.size();
    if (n1 + n2 > MAX_AERO_COMPONENT_SIZE && rng) {
      ckey.phase = PH_COMP;
      Stream s0(rng, ckey, j0);
      int n1_new = prob_round((double)p1.n_primary_parts / (double)(p1.n_primary_parts + p2.n_primary_parts) *
                                  (double)MAX_AERO_COMPONENT_SIZE, s0.next());
      n1_new = std::min(n1_new, n1);  // holds upstream whenever n_components = min(n_primary_parts, 20)
      int n2_new = std::min(MAX_AERO_COMPONENT_SIZE - n1_new, n2);
      Stream sa(rng, ckey, j0 + 1), sb(rng, ckey, j0 + 33);
      const std::vector<int> pick1 = sample_without_replacement(n1_new, n1, sa);
      const std::vector<int> pick2 = sample_without_replacement(n2_new, n2, sb);
      for (int i : pick1) { csrc.push_back(p1.comp_src[i - 1]); ctime.push_back(p1.comp_time[i - 1]); }
      for (int i : pick2) { csrc.push_back(p2.comp_src[i - 1]); ctime.push_back(p2.comp_time[i - 1]); }
    } else {
      csrc = p1.comp_src;
      ctime = p1.comp_time;
      csrc.insert(csrc.end(), p2.comp_src.begin(), p2.comp_src.end());
      ctime.insert(ctime.end(), p2.comp_time.begin(), p2.comp_time.end());
    }
  }
  const size_t n_opq = std::max(p1.opq.size(), p2.opq.size());
  std::vector<double> v(p1.vol.size());
  for (size_t i = 0; i < v.size(); i++) v[i] = p1.vol[i] + p2.vol[i];
  const int hyst = (p1.water_hyst_leg == 1 && p2.water_hyst_leg == 1) ? 1 : 0;
  const double gct = std::max(p1.greatest_create_time, p2.greatest_create_time);
  const int npp = p1.n_primary_parts + p2.n_primary_parts;
  pn.vol = v;
  pn.weight_group = 0;
  pn.weight_class = 0;
  pn.water_hyst_leg = hyst;
  pn.id = 0;
  pn.greatest_create_time = gct;
  // least_create_time is not assigned by aero_particle_coagulate; a fresh
  // aero_particle_t carries the default (we use min, documented in DESIGN.md)
  pn.least_create_time = std::min(p1.least_create_time, p2.least_create_time);
  pn.n_primary_parts = npp;
  pn.comp_src = csrc;
  pn.comp_time = ctime;
  pn.opq.assign(n_opq, 0.0);  // optics of a coagulated particle are zero until recomputed (:920-929)
}
double awa_num_conc(const WeightArray& a, const Particle& p) {
  return awa_num_conc_at_radius(a, p.weight_class, particle_radius(p));
}
double awa_single_num_conc(const WeightArray& a, const Particle& p) {
  return weight_num_conc_at_radius(a.at(p.weight_group, p.weight_class), particle_radius(p));
}

// ======================================================================= kernels
double kernel_brown_helper(double vol_i, double den_i, double vol_j, double den_j, double temp, double pressure) {
  const double rhoair = (pressure * Const::air_molec_weight) / (Const::univ_gas_const * temp);
  const double viscosd = 1.8325e-05 * (416.16 / (temp + 120.0)) * std::pow(temp / 296.16, 1.5);
  const double viscosk = viscosd / rhoair;
  const double gasspeed =
      std::sqrt((8.0 * Const::boltzmann * temp * Const::avagadro) / (Const::pi * Const::air_molec_weight));
  const double gasfreepath = 2.0 * viscosk / gasspeed;

  const double rad_i = sphere_vol2rad(vol_i);
  const double Rme_i = rad_i;  // spherical: fractal.F90:330-333
  double knud = gasfreepath / Rme_i;
  double cunning = 1.0 + knud * (1.249 + 0.42 * std::exp(-0.87 / knud));
  const double diffus_i = (Const::boltzmann * temp * cunning) / (6.0 * Const::pi * Rme_i * viscosd);
  const double speedsq_i = 8.0 * Const::boltzmann * temp / (Const::pi * den_i * vol_i);
  double freepath = 8.0 * diffus_i / (Const::pi * std::sqrt(speedsq_i));
  double t = 2.0 * Rme_i + freepath;
  double tmp1 = t * t * t;
  double tmp2 = std::pow(4.0 * Rme_i * Rme_i + freepath * freepath, 1.5);
  double dd = (tmp1 - tmp2) / (6.0 * Rme_i * freepath) - 2.0 * Rme_i;
  const double deltasq_i = dd * dd;

  const double rad_j = sphere_vol2rad(vol_j);
  const double Rme_j = rad_j;
  knud = gasfreepath / Rme_j;
  cunning = 1.0 + knud * (1.249 + 0.42 * std::exp(-0.87 / knud));
  const double diffus_j = (Const::boltzmann * temp * cunning) / (6.0 * Const::pi * Rme_j * viscosd);
  const double speedsq_j = 8.0 * Const::boltzmann * temp / (Const::pi * den_j * vol_j);
  freepath = 8.0 * diffus_j / (Const::pi * std::sqrt(speedsq_j));
  t = 2.0 * Rme_j + freepath;
  tmp1 = t * t * t;
  tmp2 = std::pow(4.0 * Rme_j * Rme_j + freepath * freepath, 1.5);
  dd = (tmp1 - tmp2) / (6.0 * Rme_j * freepath) - 2.0 * Rme_j;
  const double deltasq_j = dd * dd;

  const double rad_sum = rad_i + rad_j;
  const double diffus_sum = diffus_i + diffus_j;
  tmp1 = rad_sum / (rad_sum + std::sqrt(deltasq_i + deltasq_j));
  tmp2 = 4.0 * diffus_sum / (rad_sum * std::sqrt(speedsq_i + speedsq_j));
  return 4.0 * Const::pi * rad_sum * diffus_sum / (tmp1 + tmp2);
}
// ---- sedimentation kernel (coag_kernel_sedi.F90): terminal velocities of Beard (1976) as coded in fall_g (:112-173),
// Hall (1980) collision efficiencies with two-dimensional linear interpolation as coded in effic (:180-291)
#include "pmc_oracle_sedi_table.hpp"
// x**n for an integer n >= 0 the way gfortran forms it (libgcc __powidf2: square and multiply)
static double sedi_powi(double x, int m) {
  unsigned n = (unsigned)m;
  double y = (n & 1u) ? x : 1.0;
  while (n >>= 1) {
    x = x * x;
    if (n & 1u) y = y * x;
  }
  return y;
}
static double sedi_fall_g(double r) {  // r in m, result in m/s
  static const double b[7] = {-0.318657e1, 0.992696e0, -0.153193e-2, -0.987059e-3, -0.578878e-3, 0.855176e-4, -0.327815e-5};
  static const double c[6] = {-0.500015e1, 0.523778e1, -0.204914e1, 0.475294e0, -0.542819e-1, 0.238449e-2};
  const double eta = 1.818e-4, xlamb = 6.62e-6, rhow = 1.0, rhoa = 1.225e-3, grav = 980.665;
  const double cunh = 1.257 * xlamb, t0 = 273.15;
  const double sigma = 76.1 - 0.155 * (293.15 - t0);
  const double stok = 2.0 * grav * (rhow - rhoa) / (9.0 * eta);
  const double stb = 32.0 * rhoa * (rhow - rhoa) * grav / (3.0 * eta * eta);
  const double phy = sigma * sigma * sigma * rhoa * rhoa / (sedi_powi(eta, 4) * grav * (rhow - rhoa));
  const double py = std::pow(phy, 1.0 / 6.0);
  const double rr = r * 1e2;  // cm
  double w_inf = 0.0;
  if (rr <= 1e-3) {
    w_inf = stok * (rr * rr + cunh * rr);
  } else if (rr > 1e-3 && rr <= 5.35e-2) {
    const double x = std::log(stb * rr * rr * rr);
    double y = 0.0;
    for (int i = 1; i <= 7; i++) y = y + b[i - 1] * sedi_powi(x, i - 1);
    const double xrey = (1.0 + cunh / rr) * std::exp(y);
    w_inf = xrey * eta / (2.0 * rhoa * rr);
  } else if (rr > 5.35e-2) {
    double bond = grav * (rhow - rhoa) * (rr * rr) / sigma;
    if (rr > 0.35) bond = grav * (rhow - rhoa) * (0.35 * 0.35) / sigma;
    const double x = std::log(16.0 * bond * py / 3.0);
    double y = 0.0;
    for (int i = 1; i <= 6; i++) y = y + c[i - 1] * sedi_powi(x, i - 1);
    const double xrey = py * std::exp(y);
    w_inf = xrey * eta / (2.0 * rhoa * rr);
    if (rr > 0.35) w_inf = xrey * eta / (2.0 * rhoa * 0.35);
  }
  return w_inf / 100.0;
}
static double sedi_effic(double r1, double r2) {
  const double r_small = std::min(r1 * 1e6, r2 * 1e6), r_big = std::max(r1 * 1e6, r2 * 1e6);
  const double rq = r_small / r_big;
  int ir = 1, iq = 1;
  for (int k = 1; k <= 15; k++)
    if (r_big > sedi_r0[k - 1]) ir = k + 1;
  for (int kk = 1; kk <= 21; kk++)
    if (rq > sedi_rat[kk - 1]) iq = kk + 1;
  auto E = [&](int i, int j) { return sedi_ecoll[(j - 1) * 15 + (i - 1)]; };
  double ec;
  if (ir < 16) {
    if (ir >= 2) {
      const double p = (r_big - sedi_r0[ir - 2]) / (sedi_r0[ir - 1] - sedi_r0[ir - 2]);
      const double q = (rq - sedi_rat[iq - 2]) / (sedi_rat[iq - 1] - sedi_rat[iq - 2]);
      ec = (1.0 - p) * (1.0 - q) * E(ir - 1, iq - 1) + p * (1.0 - q) * E(ir, iq - 1) + q * (1.0 - p) * E(ir - 1, iq) +
           p * q * E(ir, iq);
    } else {
      const double q = (rq - sedi_rat[iq - 2]) / (sedi_rat[iq - 1] - sedi_rat[iq - 2]);
      ec = (1.0 - q) * E(1, iq - 1) + q * E(1, iq);
    }
  } else {
    const double q = (rq - sedi_rat[iq - 2]) / (sedi_rat[iq - 1] - sedi_rat[iq - 2]);
    const double ek = (1.0 - q) * E(15, iq - 1) + q * E(15, iq);
    ec = std::min(ek, 1.0);
  }
  if (ec < 1e-20) die_msg(99, "effic: collision efficiency vanishes");  // :289 pmc_stop(99)
  return ec;
}
// kernel_sedi_helper (coag_kernel_sedi.F90:79-103); spherical particles: the mobility radius is the radius
double kernel_sedi_helper(double v1, double v2) {
  const double r1 = sphere_vol2rad(v1), r2 = sphere_vol2rad(v2);
  const double winf1 = sedi_fall_g(r1), winf2 = sedi_fall_g(r2);
  const double ec = sedi_effic(r1, r2);
  return ec * Const::pi * ((r1 + r2) * (r1 + r2)) * std::fabs(winf1 - winf2);
}
// kernel_brown_free_helper (coag_kernel_brown_free.F90:118-147), Eq. 4 of Vemury and Pratsinis [1995]
double kernel_brown_free_helper(double v1, double d1, double v2, double d2, double temp) {
  const Fractal& f = fractal();
  const double rho = (v1 / (v1 + v2)) * d1 + (v2 / (v1 + v2)) * d2;
  const double N_i = fractal_vol_to_num_of_monomers(v1), N_j = fractal_vol_to_num_of_monomers(v2);
  const double N_i_inv_df = std::pow(N_i, 1.0 / f.frac_dim), N_j_inv_df = std::pow(N_j, 1.0 / f.frac_dim);
  const double t = N_i_inv_df + N_j_inv_df;
  return std::sqrt(6.0 * Const::boltzmann * temp * f.prime_radius / rho * (1.0 / N_i + 1.0 / N_j)) * (t * t);
}
// kernel_brown_cont_helper (coag_kernel_brown_cont.F90:117-144), Eq. 6 of Vemury and Pratsinis [1995]
double kernel_brown_cont_helper(double v1, double d1, double v2, double d2, double temp) {
  (void)d1;
  (void)d2;
  const Fractal& f = fractal();
  const double N_i = fractal_vol_to_num_of_monomers(v1), N_j = fractal_vol_to_num_of_monomers(v2);
  const double N_i_inv_df = std::pow(N_i, 1.0 / f.frac_dim), N_j_inv_df = std::pow(N_j, 1.0 / f.frac_dim);
  return 2.0 * Const::boltzmann * temp / 3.0 / Const::air_dyn_visc * (1.0 / N_i_inv_df + 1.0 / N_j_inv_df) *
         (N_i_inv_df + N_j_inv_df);
}
static const double beta_1 = 1000.0;  // coag_kernel_additive.F90:23
static const double beta_0 = 0.25 / (60.0 * 2e8);  // coag_kernel_constant.F90:22

// coag_kernel_brown.F90:58-101
static void kernel_brown_minmax(double v1, double v2, const AeroData& d, const EnvState& e, double& k_min,
                                double& k_max) {
  const int n_sample = 3;
  double d_min = d.density[0], d_max = d.density[0];
  for (double x : d.density) { d_min = std::min(d_min, x); d_max = std::max(d_max, x); }
  bool first = true;
  for (int i = 1; i <= n_sample; i++)
    for (int j = 1; j <= n_sample; j++) {
      const double d1 = interp_linear_disc(d_min, d_max, n_sample, i);
      const double d2 = interp_linear_disc(d_min, d_max, n_sample, j);
      const double k = kernel_brown_helper(v1, d1, v2, d2, e.temp, e.pressure);
      if (first) {
        first = false;
        k_min = k;
        k_max = k;
      } else {
        k_min = std::min(k_min, k);
        k_max = std::max(k_max, k);
      }
    }
}
// kernel_brown_free_minmax / kernel_brown_cont_minmax (coag_kernel_brown_free.F90:62-108, _cont.F90:62-107): 3 x 3
// densities in [min rho, max rho]
static void kernel_brown_fc_minmax(int type, double v1, double v2, const AeroData& d, const EnvState& e, double& k_min,
                                   double& k_max) {
  const int n_sample = 3;
  double d_min = d.density[0], d_max = d.density[0];
  for (double x : d.density) { d_min = std::min(d_min, x); d_max = std::max(d_max, x); }
  bool first = true;
  for (int i = 1; i <= n_sample; i++)
    for (int j = 1; j <= n_sample; j++) {
      const double d1 = interp_linear_disc(d_min, d_max, n_sample, i);
      const double d2 = interp_linear_disc(d_min, d_max, n_sample, j);
      const double k = type == KERNEL_BROWN_FREE ? kernel_brown_free_helper(v1, d1, v2, d2, e.temp)
                                                 : kernel_brown_cont_helper(v1, d1, v2, d2, e.temp);
      if (first) {
        first = false;
        k_min = k;
        k_max = k;
      } else {
        k_min = std::min(k_min, k);
        k_max = std::max(k_max, k);
      }
    }
}
void kernel_minmax(int type, double v1, double v2, const AeroData& d, const EnvState& e, double& kmin, double& kmax) {
  if (type == KERNEL_BROWN || type == KERNEL_SEDI) require_spherical("kernel_minmax (brown / sedi)");
  if (type == KERNEL_BROWN_FREE || type == KERNEL_BROWN_CONT) {
    kernel_brown_fc_minmax(type, v1, v2, d, e, kmin, kmax);
  } else if (type == KERNEL_ADDITIVE) {
    kmin = beta_1 * (v1 + v2);  // coag_kernel_additive.F90:52-70
    kmax = kmin;
  } else if (type == KERNEL_CONSTANT) {
    kmin = beta_0;
    kmax = beta_0;
  } else if (type == KERNEL_BROWN) {
    kernel_brown_minmax(v1, v2, d, e, kmin, kmax);
  } else if (type == KERNEL_ZERO) {
    kmin = 0.0;
    kmax = 0.0;
  } else if (type == KERNEL_SEDI) {
    kmin = kernel_sedi_helper(v1, v2);  // kernel_sedi_minmax (coag_kernel_sedi.F90:52-73)
    kmax = kmin;
  } else {
    die_msg(330498208, "Unknown kernel type");
  }
}
double kernel(int type, const Particle& p1, const Particle& p2, const AeroData& d, const EnvState& e) {
  if (type == KERNEL_ADDITIVE) return beta_1 * (particle_volume(p1) + particle_volume(p2));
  if (type == KERNEL_CONSTANT) return beta_0;
  if (type == KERNEL_ZERO) return 0.0;
  if (type == KERNEL_BROWN || type == KERNEL_SEDI) require_spherical("kernel (brown / sedi)");
  if (type == KERNEL_SEDI) return kernel_sedi_helper(particle_volume(p1), particle_volume(p2));
  if (type == KERNEL_BROWN || type == KERNEL_BROWN_FREE || type == KERNEL_BROWN_CONT) {
    const double v1 = particle_volume(p1), v2 = particle_volume(p2);
    const double d1 = particle_density(p1, d), d2 = particle_density(p2, d);
    if (type == KERNEL_BROWN_FREE) return kernel_brown_free_helper(v1, d1, v2, d2, e.temp);
    if (type == KERNEL_BROWN_CONT) return kernel_brown_cont_helper(v1, d1, v2, d2, e.temp);
    return kernel_brown_helper(v1, d1, v2, d2, e.temp, e.pressure);
  }
  die_msg(200724934, "Unknown kernel type");
  return 0.0;
}
// coag_kernel.F90:286-341
static void est_k_minmax_for_bin_unweighted(const BinGrid& g, int type, int b1, int b2, const AeroData& d,
                                            const EnvState& e, double& k_min, double& k_max) {
  const int n_sample = 3;
  const double over_scale = 2.0;
  const double v1_low = aero_rad2vol(g.edges[b1 - 1]);
  const double v1_high = aero_rad2vol(g.edges[b1]);
  const double v2_low = aero_rad2vol(g.edges[b2 - 1]);
  const double v2_high = aero_rad2vol(g.edges[b2]);
  for (int i = 1; i <= n_sample; i++)
    for (int j = 1; j <= n_sample; j++) {
      const double v1 = interp_linear_disc(v1_low, v1_high, n_sample, i);
      const double v2 = interp_linear_disc(v2_low, v2_high, n_sample, j);
      double nmin, nmax;
      kernel_minmax(type, v1, v2, d, e, nmin, nmax);
      if (i == 1 && j == 1) {
        k_min = nmin;
        k_max = nmax;
      } else {
        k_min = std::min(k_min, nmin);
        k_max = std::max(k_max, nmax);
      }
    }
  k_max = k_max * over_scale;
}
void est_k_minmax_binned_unweighted(const BinGrid& g, int type, const AeroData& d, const EnvState& e,
                                    std::vector<double>& kmin, std::vector<double>& kmax) {
  const int nb = g.size();
  kmin.assign((size_t)nb * nb, 0.0);
  kmax.assign((size_t)nb * nb, 0.0);
  for (int i = 1; i <= nb; i++)
    for (int j = 1; j <= nb; j++)
      est_k_minmax_for_bin_unweighted(g, type, i, j, d, e, kmin[(size_t)(j - 1) * nb + (i - 1)],
                                      kmax[(size_t)(j - 1) * nb + (i - 1)]);
}
double coag_num_conc_factor(const WeightArray& a, double i_r, double j_r, int ic, int jc, int ijc) {
  const double ij_r = aero_vol2rad(aero_rad2vol(i_r) + aero_rad2vol(j_r));
  const double i_nc = awa_num_conc_at_radius(a, ic, i_r);
  const double j_nc = awa_num_conc_at_radius(a, jc, j_r);
  const double ij_nc = awa_num_conc_at_radius(a, ijc, ij_r);
  const double nc_min = std::min(std::min(i_nc, j_nc), ij_nc);
  return i_nc * j_nc / nc_min;
}
int coag_dest_class(const WeightArray& a, const BinGrid& g, int ib, int jb, int ic, int jc) {
  (void)jb;
  const double i_r = g.centers[ib - 1];
  const double j_r = g.centers[ib - 1];  // sic: coag_kernel.F90:403 uses i_bin for j_r
  const double ij_r = aero_vol2rad(aero_rad2vol(i_r) + aero_rad2vol(j_r));
  const double ij_nc_i = awa_num_conc_at_radius(a, ic, ij_r);
  const double ij_nc_j = awa_num_conc_at_radius(a, jc, ij_r);
  return (ij_nc_i < ij_nc_j) ? ic : jc;
}
double max_coag_num_conc_factor(const WeightArray& a, const BinGrid& g, int ib, int jb, int ic, int jc, int ijc) {
  const int n_sample = 5;
  const double i_r_min = g.edges[ib - 1], i_r_max = g.edges[ib];
  const double j_r_min = g.edges[jb - 1], j_r_max = g.edges[jb];
  double f_max = 0.0;
  for (int is = 1; is <= n_sample; is++)
    for (int js = 1; js <= n_sample; js++) {
      const double i_r = interp_linear_disc(i_r_min, i_r_max, n_sample, is);
      const double j_r = interp_linear_disc(j_r_min, j_r_max, n_sample, js);
      f_max = std::max(f_max, coag_num_conc_factor(a, i_r, j_r, ic, jc, ijc));
    }
  return f_max;
}
double num_conc_weighted_kernel(int type, const Particle& p1, const Particle& p2, int ic, int jc, int ijc,
                                const AeroData& d, const WeightArray& a, const EnvState& e) {
  const double unweighted_k = kernel(type, p1, p2, d, e);
  const double i_r = particle_radius(p1);
  const double j_r = particle_radius(p2);
  return unweighted_k * coag_num_conc_factor(a, i_r, j_r, ic, jc, ijc);
}

// ======================================================================= varray / rmap2
void IntVarray::remove_entry(int index) {
  pmc_assert(541032660, index >= 1);
  pmc_assert(385739765, index <= n_entry());
  if (index < n_entry()) entry[index - 1] = entry[n_entry() - 1];
  entry.pop_back();
}
void Rmap2::set_ranges(int a, int b) {
  n1 = a;
  n2 = b;
  inverse.assign((size_t)a * b, IntVarray());
  forward1.entry.clear();
  forward2.entry.clear();
  index.entry.clear();
}
void Rmap2::zero() {
  forward1.entry.clear();
  forward2.entry.clear();
  index.entry.clear();
  for (auto& v : inverse) v.entry.clear();
}
void Rmap2::append(int i1, int i2) {
  pmc_assert(708651144, i1 >= 1);
  pmc_assert(779828769, i1 <= n1);
  pmc_assert(978259336, i2 >= 1);
  pmc_assert(238981205, i2 <= n2);
  forward1.append(i1);
  forward2.append(i2);
  inv(i1, i2).append(forward1.n_entry());
  index.append(inv(i1, i2).n_entry());
}
void Rmap2::change(int i_domain, int i1, int i2) {
  pmc_assert(191141591, i_domain >= 1);
  pmc_assert(240079303, i_domain <= forward1.n_entry());
  pmc_assert(671426897, i1 >= 1);
  pmc_assert(311976942, i1 <= n1);
  pmc_assert(383129645, i2 >= 1);
  pmc_assert(771283685, i2 <= n2);
  const int i1_old = forward1.at(i_domain);
  const int i2_old = forward2.at(i_domain);
  // sic (integer_rmap2.F90:160): range 2 is not tested, so a change of class
  // alone inside the same bin returns early.
  if ((i1_old == i1) && (i1_old == i1)) return;
  const int i_index_old = index.at(i_domain);
  inv(i1_old, i2_old).remove_entry(i_index_old);
  if (i_index_old <= inv(i1_old, i2_old).n_entry()) {
    const int i_domain_shifted = inv(i1_old, i2_old).at(i_index_old);
    index.at(i_domain_shifted) = i_index_old;
  }
  forward1.at(i_domain) = i1;
  forward2.at(i_domain) = i2;
  inv(i1, i2).append(i_domain);
  index.at(i_domain) = inv(i1, i2).n_entry();
}
void Rmap2::remove(int i_domain) {
  pmc_assert(242566612, i_domain >= 1);
  pmc_assert(110569289, i_domain <= forward1.n_entry());
  const int i1_old = forward1.at(i_domain);
  const int i2_old = forward2.at(i_domain);
  const int i_index_old = index.at(i_domain);
  const int i_domain_shifted = forward1.n_entry();
  if (i_domain_shifted != i_domain) {
    const int i1_fix = forward1.at(i_domain_shifted);
    const int i2_fix = forward2.at(i_domain_shifted);
    const int i_index_fix = index.at(i_domain_shifted);
    inv(i1_fix, i2_fix).at(i_index_fix) = i_domain;
  }
  forward1.remove_entry(i_domain);
  forward2.remove_entry(i_domain);
  index.remove_entry(i_domain);
  const int i_index_fix = inv(i1_old, i2_old).n_entry();
  const int i_domain_fix = inv(i1_old, i2_old).at(i_index_fix);
  inv(i1_old, i2_old).remove_entry(i_index_old);
  if (i_index_fix != i_index_old) index.at(i_domain_fix) = i_index_old;
}
void Rmap2::check(int n_domain) const {
  pmc_assert(721004561, forward1.n_entry() == n_domain);
  pmc_assert(721004562, forward2.n_entry() == n_domain);
  pmc_assert(721004563, index.n_entry() == n_domain);
  for (int i = 1; i <= n_domain; i++) {
    const int a = forward1.at(i), b = forward2.at(i), k = index.at(i);
    pmc_assert(721004564, a >= 1 && a <= n1 && b >= 1 && b <= n2);
    pmc_assert(721004565, k >= 1 && k <= inv(a, b).n_entry());
    pmc_assert(721004566, inv(a, b).at(k) == i);
  }
  int total = 0;
  for (const auto& v : inverse) total += v.n_entry();
  pmc_assert(721004567, total == n_domain);
}

// ======================================================================= aero_sorted
static const double AERO_SORTED_BINS_PER_DECADE = 10.0;
static const double AERO_SORTED_BIN_OVER_FACTOR = 10.0;
static const double AERO_SORTED_BIN_SAFETY_FACTOR = 3.0;

void aero_sorted_set_bin_grid(AeroSorted& s, const BinGrid& g, int n_group, int n_class) {
  const int n_bin = g.size();
  s.size_class.set_ranges(n_bin, n_class);
  s.group_class.set_ranges(n_group, n_class);
  s.coag_kernel_bounds_valid = false;
  s.coag_kernel_min.assign((size_t)n_bin * n_bin, 0.0);
  s.coag_kernel_max.assign((size_t)n_bin * n_bin, 0.0);
  s.bin_grid = g;
}
int aero_sorted_particle_in_bin(const AeroSorted& s, const Particle& p) {
  return bin_grid_find(s.bin_grid, particle_radius(p));
}
void aero_sorted_sort_particles(AeroSorted& s, const std::vector<Particle>& apa) {
  s.size_class.zero();
  s.group_class.zero();
  for (size_t i = 0; i < apa.size(); i++) {
    const int i_bin = aero_sorted_particle_in_bin(s, apa[i]);
    s.size_class.append(i_bin, apa[i].weight_class);
    s.group_class.append(apa[i].weight_group, apa[i].weight_class);
  }
}
bool aero_sorted_remake_if_needed(AeroSorted& s, std::vector<Particle>& apa, bool valid_sort, int n_group,
                                  int n_class) {
  bool need_new_bin_grid = false;
  double r_min = 0.0, r_max = 0.0;
  if (valid_sort) {
    int i_bin_min = 0, i_bin_max = 0;
    for (int i_bin = 1; i_bin <= s.bin_grid.size(); i_bin++) {
      int tot = 0;
      for (int c = 1; c <= s.n_class(); c++) tot += s.size_class.inv(i_bin, c).n_entry();
      if (tot > 0) {
        if (i_bin_min == 0) i_bin_min = i_bin;
        i_bin_max = i_bin;
      }
    }
    if (i_bin_min == 0) {
      pmc_assert(333430891, i_bin_max == 0);
      if (s.bin_grid.size() > 0) {
        r_min = s.bin_grid.edges[s.bin_grid.size()];
        r_max = s.bin_grid.edges[0];
      }
    } else {
      r_min = s.bin_grid.edges[i_bin_min - 1];
      r_max = s.bin_grid.edges[i_bin_max];
    }
  } else {
    for (size_t i = 0; i < apa.size(); i++) {
      const double r = particle_radius(apa[i]);
      if (i == 0) {
        r_min = r;
        r_max = r;
      } else {
        r_min = std::min(r_min, r);
        r_max = std::max(r_max, r);
      }
    }
  }
  if (r_max == 0.0) {
    if (valid_sort) return false;
    BinGrid empty;
    aero_sorted_set_bin_grid(s, empty, n_group, n_class);
    return true;
  }
  if (s.bin_grid.size() < 1) {
    need_new_bin_grid = true;
  } else {
    const double grid_r_min = s.bin_grid.edges[0];
    const double grid_r_max = s.bin_grid.edges[s.bin_grid.size()];
    if ((r_min / grid_r_min < AERO_SORTED_BIN_SAFETY_FACTOR) ||
        (grid_r_max / r_max < AERO_SORTED_BIN_SAFETY_FACTOR))
      need_new_bin_grid = true;
  }
  if (need_new_bin_grid) {
    const double grid_r_min = r_min / AERO_SORTED_BIN_OVER_FACTOR;
    const double grid_r_max = r_max * AERO_SORTED_BIN_OVER_FACTOR;
    const int n_bin = (int)std::ceil((std::log10(grid_r_max) - std::log10(grid_r_min)) * AERO_SORTED_BINS_PER_DECADE);
    BinGrid ng;
    bin_grid_make_log(ng, n_bin, grid_r_min, grid_r_max);
    aero_sorted_set_bin_grid(s, ng, n_group, n_class);
    aero_sorted_sort_particles(s, apa);
    return true;
  }
  if (!valid_sort) aero_sorted_sort_particles(s, apa);
  return false;
}
void aero_sorted_check(const AeroSorted& s, const std::vector<Particle>& apa, int n_group, int n_class) {
  const int n = (int)apa.size();
  pmc_assert(577701401, s.size_class.n2 == n_class && s.group_class.n1 == n_group);
  s.size_class.check(n);
  s.group_class.check(n);
  for (int i = 1; i <= n; i++) {
    pmc_assert(577701402, s.size_class.forward1.at(i) == aero_sorted_particle_in_bin(s, apa[i - 1]));
    pmc_assert(577701403, s.size_class.forward2.at(i) == apa[i - 1].weight_class);
    pmc_assert(577701404, s.group_class.forward1.at(i) == apa[i - 1].weight_group);
  }
}

// ======================================================================= aero_state
void aero_state_sort(AeroState& st) {
  aero_sorted_remake_if_needed(st.sorted, st.apa, st.valid_sort, st.awa.n_group, st.awa.n_class);
  st.valid_sort = true;
}
void aero_state_physical_sort(AeroState& st) {
  pmc_assert(910000001, st.valid_sort);
  const int nb = st.sorted.n_bin(), nc = st.sorted.n_class();
  std::vector<Particle> out;
  out.reserve(st.apa.size());
  for (int c = 1; c <= nc; c++)
    for (int b = 1; b <= nb; b++) {
      const IntVarray& l = st.sorted.size_class.inv(b, c);
      for (int k = 1; k <= l.n_entry(); k++) out.push_back(st.apa[l.at(k) - 1]);
    }
  pmc_assert(910000002, out.size() == st.apa.size());
  st.apa.swap(out);
  aero_sorted_sort_particles(st.sorted, st.apa);
}
int aero_state_total_particles(const AeroState& st, int g, int c) {
  if (st.valid_sort) return st.sorted.group_class.inv(g, c).n_entry();
  int n = 0;
  for (const Particle& p : st.apa)
    if (p.weight_group == g && p.weight_class == c) n++;
  return n;
}
// aero_sorted.F90:397-449 + aero_state.F90:365-380
void aero_state_add_particle(AeroState& st, const Particle& p, bool allow_resort) {
  if (!st.valid_sort) {
    st.apa.push_back(p);
    return;
  }
  AeroSorted& s = st.sorted;
  const int i_bin = aero_sorted_particle_in_bin(s, p);
  const int n_bin = s.bin_grid.size();
  pmc_assert(417177855, p.weight_group >= 1 && p.weight_group <= s.n_group());
  pmc_assert(233133947, p.weight_class >= 1 && p.weight_class <= s.n_class());
  st.apa.push_back(p);
  if (i_bin < 1 || i_bin > n_bin) {
    if (!allow_resort) die_msg(134572570, "particle outside of bin_grid: try reducing the timestep del_t");
    aero_sorted_remake_if_needed(s, st.apa, false, s.n_group(), s.n_class());
  } else {
    s.size_class.append(i_bin, p.weight_class);
    s.group_class.append(p.weight_group, p.weight_class);
  }
}
// aero_particle_array.F90:178-200 + aero_sorted.F90:455-470
void aero_state_remove_particle(AeroState& st, int i_part) {
  pmc_assert(992946227, i_part >= 1 && i_part <= st.n_part());
  if (i_part < st.n_part()) st.apa[i_part - 1] = st.apa.back();
  st.apa.pop_back();
  if (st.valid_sort) {
    st.sorted.size_class.remove(i_part);
    st.sorted.group_class.remove(i_part);
  }
}
void aero_state_dup_particle(AeroState& st, int i_part, double n_part_mean, bool random_weight_group, Stream& s) {
  const int n_copies = prob_round(n_part_mean, s.next());
  if (n_copies == 0) {
    aero_state_remove_particle(st, i_part);
  } else if (n_copies > 1) {
    for (int i_dup = 1; i_dup <= n_copies - 1; i_dup++) {
      Particle np = st.apa[i_part - 1];
      np.id = st.idgen.new_id();
      if (random_weight_group) {
        np.weight_group = awa_rand_group(st.awa, st.apa[i_part - 1].weight_class, particle_radius(st.apa[i_part - 1]), s);
      }
      aero_state_add_particle(st, np);
    }
  }
}
void aero_state_double(AeroState& st, int g, int c) {
  const int n = st.n_part();
  for (int i = 1; i <= n; i++) {
    if (st.apa[i - 1].weight_group == g && st.apa[i - 1].weight_class == c) {
      Particle np = st.apa[i - 1];
      np.id = st.idgen.new_id();
      aero_state_add_particle(st, np);
    }
  }
  st.valid_sort = false;
  st.awa.at(g, c).magnitude *= 0.5;
}
void aero_state_halve(AeroState& st, int g, int c, const RunCtx& rc, uint32_t pass) {
  const uint32_t gc = (uint32_t)((c - 1) * st.awa.n_group + (g - 1));
  for (int i = st.n_part(); i >= 1; i--) {
    if (st.apa[i - 1].weight_group == g && st.apa[i - 1].weight_class == c) {
      const double u = rc.rng->uniform(rc.key(PH_HALVE, pass * 4096u + gc, (uint32_t)i), 0);
      if (u < 0.5) aero_state_remove_particle(st, i);
    }
  }
  st.awa.at(g, c).magnitude *= 2.0;
}
void aero_state_rebalance(AeroState& st, bool allow_doubling, bool allow_halving, const RunCtx& rc) {
  const int ng = st.awa.n_group, nc = st.awa.n_class;
  if (allow_doubling) {
    for (int g = 1; g <= ng; g++)
      for (int c = 1; c <= nc; c++) {
        int n = aero_state_total_particles(st, g, c);
        while (((double)n < st.ideal(g, c) / 2.0) && (n > 0)) {
          aero_state_double(st, g, c);
          n = aero_state_total_particles(st, g, c);
        }
      }
  }
  if (allow_halving) {
    for (int g = 1; g <= ng; g++)
      for (int c = 1; c <= nc; c++) {
        uint32_t pass = 0;
        while ((double)aero_state_total_particles(st, g, c) > st.ideal(g, c) * 2.0) {
          aero_state_halve(st, g, c, rc, pass);
          pass++;
        }
      }
  }
}
void aero_state_scale_weight(AeroState& st, int g, int c, double weight_ratio, bool allow_doubling,
                             bool allow_halving, const RunCtx& rc, uint32_t salt) {
  aero_state_sort(st);
  const uint32_t gc = (uint32_t)((c - 1) * st.awa.n_group + (g - 1)) + 4096u * salt;
  const int n_part = st.sorted.group_class.inv(g, c).n_entry();
  if ((weight_ratio > 1.0) && (allow_halving || (n_part == 0))) {
    st.awa.at(g, c).magnitude *= weight_ratio;
    const int n_remove = prob_round((double)n_part * (1.0 - 1.0 / weight_ratio), rc.rng->uniform(rc.key(PH_SCALE, gc, 0), 0));
    for (int i_remove = 1; i_remove <= n_remove; i_remove++) {
      const int ne = st.sorted.group_class.inv(g, c).n_entry();
      const int i_entry = pmc_rand_int(rc.rng->uniform(rc.key(PH_SCALE, gc, (uint32_t)i_remove), 0), ne);
      const int i_part = st.sorted.group_class.inv(g, c).at(i_entry);
      aero_state_remove_particle(st, i_part);
    }
  } else if ((weight_ratio < 1.0) && (allow_doubling || (n_part == 0))) {
    st.awa.at(g, c).magnitude *= weight_ratio;
    for (int i_entry = n_part; i_entry >= 1; i_entry--) {
      const int i_part = st.sorted.group_class.inv(g, c).at(i_entry);
      Stream s(rc.rng, rc.key(PH_SCALE, gc, 0x40000000u + (uint32_t)i_entry));
      aero_state_dup_particle(st, i_part, 1.0 / weight_ratio, false, s);
    }
  }
}
void aero_state_reweight(AeroState& st, const std::vector<double>& reweight_num_conc, const RunCtx& rc,
                         uint32_t pass) {
  const int ng = st.awa.n_group, nc = st.awa.n_class;
  std::vector<double> n_part_old((size_t)ng * nc, 0.0), n_part_new((size_t)ng * nc, 0.0);
  const int n = st.n_part();
  for (int i = 1; i <= n; i++) {
    const double old_nc = reweight_num_conc[i - 1];
    const double new_nc = awa_single_num_conc(st.awa, st.apa[i - 1]);
    const double n_part_mean = old_nc / new_nc;
    const size_t k = (size_t)(st.apa[i - 1].weight_class - 1) * ng + (st.apa[i - 1].weight_group - 1);
    n_part_new[k] += n_part_mean;
    n_part_old[k] += 1.0;
  }
  for (int g = 1; g <= ng; g++)
    for (int c = 1; c <= nc; c++) {
      const size_t k = (size_t)(c - 1) * ng + (g - 1);
      if (n_part_old[k] == 0.0) continue;
      st.awa.at(g, c).magnitude *= n_part_new[k] / n_part_old[k];
    }
  for (int i = n; i >= 1; i--) {
    const double old_nc = reweight_num_conc[i - 1];
    const double new_nc = awa_single_num_conc(st.awa, st.apa[i - 1]);
    Stream s(rc.rng, rc.key(PH_REWEIGHT, pass, (uint32_t)i));
    aero_state_dup_particle(st, i, old_nc / new_nc, false, s);
  }
}

// ======================================================================= coagulation.F90
static const double COAG_ACCEL_N_EVENT = 1.0;
static const double COAG_ACCEL_MAX_CV = 0.1;
static const double MAX_ALLOWABLE_GROWTH_FACTOR = 1.5;

uint32_t coag_pair_id(int n_class, int c1, int c2, int b1, int b2) {
  return ((uint32_t)((c1 - 1) * n_class + (c2 - 1)) * 256u + (uint32_t)b1) * 256u + (uint32_t)b2;
}

// coagulation.F90:247-307
static void determine_target_and_source(const WeightArray& awa, const BinGrid& g, int b1, int b2, int c1, int c2,
                                        int cc, int& bt, int& bs, int& ct, int& cs, bool& ok) {
  bool inc, dec;
  awa_check_monotonicity(awa, inc, dec);
  ok = false;
  if (!dec) return;
  double nc1_min, nc1_max, nc2_min, nc2_max;
  awa_minmax_num_conc(awa, c1, g.edges[b1 - 1], g.edges[b1], nc1_min, nc1_max);
  awa_minmax_num_conc(awa, c2, g.edges[b2 - 1], g.edges[b2], nc2_min, nc2_max);
  if ((nc1_max < nc2_min) && (cc == c1)) {
    bt = b1; bs = b2; ct = c1; cs = c2;
    ok = true;
  }
  if ((nc2_max < nc1_min) && (cc == c2)) {
    bt = b2; bs = b1; ct = c2; cs = c1;
    ok = true;
  }
}

// coagulation.F90:810-922.  Draw positions inside the sample stream:
// 3,4,5 = remove_1, remove_2, create_new; 6.. = rand_group (sample_cts_pdf).
static void coagulate_weighting(const Particle& pt1, const Particle& pt2, Particle& ptc, int c1, int c2, int cc,
                                const WeightArray& awa, IdGen& idgen, Stream& s, bool& remove_1, bool& remove_2,
                                bool& create_new, bool& id_1_lost, bool& id_2_lost) {
  pmc_assert(371947172, pt1.id != pt2.id);
  const double r1 = particle_radius(pt1);
  const double r2 = particle_radius(pt2);
  const double rc = aero_vol2rad(aero_rad2vol(r1) + aero_rad2vol(r2));
  const double nc1 = awa_num_conc_at_radius(awa, c1, r1);
  const double nc2 = awa_num_conc_at_radius(awa, c2, r2);
  const double ncc = awa_num_conc_at_radius(awa, cc, rc);
  Stream sg(s.rng, s.key, 6);
  const int new_group = awa_rand_group(awa, cc, rc, sg);
  const double nc_min = std::min(std::min(nc1, nc2), ncc);
  const double prob_remove_1 = nc_min / nc1;
  const double prob_remove_2 = nc_min / nc2;
  const double prob_create_new = nc_min / ncc;
  remove_1 = (s.at(3) < prob_remove_1);
  remove_2 = (s.at(4) < prob_remove_2);
  create_new = (s.at(5) < prob_create_new);
  if (create_new) {
    id_1_lost = false;
    id_2_lost = false;
    if (remove_1 && remove_2) {
      if (particle_volume(pt1) > particle_volume(pt2))
        id_2_lost = true;
      else
        id_1_lost = true;
    }
  } else {
    id_1_lost = remove_1;
    id_2_lost = remove_2;
  }
  if (create_new) {
    particle_coagulate(pt1, pt2, ptc, s.rng, s.key, 0u);
    ptc.weight_group = new_group;
    ptc.weight_class = cc;
    if (remove_1 && !id_1_lost) {
      ptc.id = pt1.id;
    } else if (remove_2 && !id_2_lost) {
      ptc.id = pt2.id;
    } else {
      ptc.id = idgen.new_id();
    }
  }
}
// coagulation.F90:928-985
static void coagulate(AeroState& st, int p1, int p2, int c1, int c2, int cc, Stream& s) {
  Particle ptc;
  bool remove_1, remove_2, create_new, id_1_lost, id_2_lost;
  coagulate_weighting(st.apa[p1 - 1], st.apa[p2 - 1], ptc, c1, c2, cc, st.awa, st.idgen, s, remove_1, remove_2,
                      create_new, id_1_lost, id_2_lost);
  if (p2 > p1) {
    if (remove_2) aero_state_remove_particle(st, p2);
    if (remove_1) aero_state_remove_particle(st, p1);
  } else {
    if (remove_1) aero_state_remove_particle(st, p1);
    if (remove_2) aero_state_remove_particle(st, p2);
  }
  if (create_new) aero_state_add_particle(st, ptc, /*allow_resort=*/false);
}
// coagulation.F90:766-803 with draws 0 (i1) and 1 (i2) of the sample stream
static void find_rand_pair(const AeroSorted& so, int b1, int b2, int c1, int c2, Stream& s, int& i1, int& i2) {
  const int n1 = so.size_class.inv(b1, c1).n_entry();
  const int n2 = so.size_class.inv(b2, c2).n_entry();
  pmc_assert(619608562, n1 >= 1);
  i1 = pmc_rand_int(s.at(0), n1);
  if (b1 == b2 && c1 == c2) {
    pmc_assert(956184336, n2 >= 2);
    i2 = pmc_rand_int(s.at(1), n2 - 1);
    if (i2 == i1) i2 = n2;
  } else {
    pmc_assert(271635751, n2 >= 1);
    i2 = pmc_rand_int(s.at(1), n2);
  }
}
// coagulation.F90:712-759
static bool maybe_coag_pair(const EnvState& env, const AeroData& d, AeroState& st, int b1, int b2, int c1, int c2,
                            int cc, int type, double accept_factor, Stream& s) {
  int i1, i2;
  find_rand_pair(st.sorted, b1, b2, c1, c2, s, i1, i2);
  const int p1 = st.sorted.size_class.inv(b1, c1).at(i1);
  const int p2 = st.sorted.size_class.inv(b2, c2).at(i2);
  const double k = num_conc_weighted_kernel(type, st.apa[p1 - 1], st.apa[p2 - 1], c1, c2, cc, d, st.awa, env);
  const double p = k * accept_factor;
  if (s.at(2) < p) {
    coagulate(st, p1, p2, c1, c2, cc, s);
    return true;
  }
  return false;
}
// coagulation.F90:648-702
static void compute_n_samp(int ni, int nj, bool same_bin, double k_max, double del_t, Stream& s,
                           double& n_samp_mean, int& n_samp, double& accept_factor) {
  double n_possible;
  if (same_bin)
    n_possible = (double)ni * ((double)nj - 1.0) / 2.0;
  else
    n_possible = (double)ni * (double)nj;
  const double r_samp = k_max * del_t;
  n_samp_mean = r_samp * n_possible;
  n_samp = rand_poisson(n_samp_mean, s);
  accept_factor = 1.0 / k_max;
}
// coagulation.F90:587-643
static void per_set_coag(int type, double k_max, const EnvState& env, const AeroData& d, AeroState& st,
                         double del_t, const RunCtx& rc, CoagStats& out, int b1, int b2, int c1, int c2, int cc) {
  const uint32_t pid = coag_pair_id(st.awa.n_class, c1, c2, b1, b2);
  int n1 = st.sorted.size_class.inv(b1, c1).n_entry();
  int n2 = st.sorted.size_class.inv(b2, c2).n_entry();
  double n_samp_mean, accept_factor;
  int n_samp;
  Stream s0(rc.rng, rc.key(PH_COAG, pid, 0));
  compute_n_samp(n1, n2, (b1 == b2) && (c1 == c2), k_max, del_t, s0, n_samp_mean, n_samp, accept_factor);
  out.tot_n_samp += n_samp;
  for (int i_samp = 1; i_samp <= n_samp; i_samp++) {
    n1 = st.sorted.size_class.inv(b1, c1).n_entry();
    n2 = st.sorted.size_class.inv(b2, c2).n_entry();
    if (((n1 < 2) && (b1 == b2) && (c1 == c2)) || (n1 < 1) || (n2 < 1)) break;
    Stream s(rc.rng, rc.key(PH_COAG, pid, (uint32_t)i_samp));
    if (maybe_coag_pair(env, d, st, b1, b2, c1, c2, cc, type, accept_factor, s)) out.tot_n_coag += 1;
  }
}

// coagulation.F90:333-517
static void sample_source_particle(AeroState& st, const AeroData& d, const EnvState& env, int type, int bs, int cs,
                                   const Particle& coag_particle, double n_samp_mean, double accept_factor,
                                   Stream& s, Stream& ss, int& n_samp, int& n_coag, int& n_remove, Particle& source) {
  // Draw addresses: `s` is the sequential stream of the target (Poisson counts, binomial, prob_round,
  // rand_group, dup); the three draws of sample i_samp sit at positions 3 (i_samp - 1) + {0, 1, 2} of the
  // sample stream `ss`, so that a parallel implementation can evaluate candidate samples ahead of time.
  if (st.sorted.size_class.inv(bs, cs).n_entry() == 0) {
    n_samp = 0; n_remove = 0; n_coag = 0;
    return;
  }
  const double num_conc_target = awa_num_conc(st.awa, coag_particle);
  const int ct = coag_particle.weight_class;
  const double num_conc_source_min = awa_num_conc_at_radius(st.awa, cs, st.sorted.bin_grid.edges[bs - 1]);
  const double prob_remove_source_max = num_conc_target / num_conc_source_min;
  pmc_assert(653606684, prob_remove_source_max <= 1.0);
  const int n_samp_remove = rand_poisson(prob_remove_source_max * n_samp_mean, s);
  const int n_samp_extra = rand_poisson((1.0 - prob_remove_source_max) * n_samp_mean, s);
  const int n_samp_total = n_samp_remove + n_samp_extra;
  int n_avg = 0;
  n_samp = 0;
  n_remove = 0;
  double prob_coag_tot = 0.0;
  const int S = d.n_spec;
  std::vector<double> vol_sq((size_t)S, 0.0);
  source = Particle();
  source.vol.assign((size_t)S, 0.0);
  source.n_primary_parts = 0;
  for (int i_samp = 1; i_samp <= n_samp_total; i_samp++) {
    if (st.sorted.size_class.inv(bs, cs).n_entry() == 0) break;
    if ((n_samp > n_samp_remove) && (n_avg >= 2)) {
      double vol_cv_max = 0.0;
      bool any = false;
      for (int k = 0; k < S; k++) {
        const double vol_mean = source.vol[k] / (double)n_avg;
        if (vol_mean > 0.0) {
          // the variance is clamped at zero: rounding can make it slightly negative when every accepted
          // sample is the same particle, and the reference's sqrt would then hand a NaN to maxval
          const double cv = std::sqrt(std::max(0.0, vol_sq[k] / (double)n_avg - vol_mean * vol_mean)) / vol_mean;
          vol_cv_max = any ? std::max(vol_cv_max, cv) : cv;
          any = true;
        }
      }
      const double mean_95_conf_cv = vol_cv_max / std::sqrt((double)n_avg) * student_t_95_coeff(n_avg);
      if (mean_95_conf_cv < COAG_ACCEL_MAX_CV) break;
    }
    n_samp = n_samp + 1;
    ss.j = 3u * (uint32_t)(i_samp - 1);
    const int ne = st.sorted.size_class.inv(bs, cs).n_entry();
    const int i_unif_entry = pmc_rand_int(ss.next(), ne);
    const int i_part = st.sorted.size_class.inv(bs, cs).at(i_unif_entry);
    const double k = num_conc_weighted_kernel(type, st.apa[i_part - 1], coag_particle, cs, ct, ct, d, st.awa, env);
    const double prob_coag = k * accept_factor;
    prob_coag_tot = prob_coag_tot + prob_coag;
    if (ss.next() < prob_coag) {
      n_avg = n_avg + 1;
      Particle tmp;
      {
        RngKey ck = ss.key;
        particle_coagulate(source, st.apa[i_part - 1], tmp, ss.rng, ck, 128u * (uint32_t)i_samp);
      }
      source = tmp;
      for (int q = 0; q < S; q++) vol_sq[q] += st.apa[i_part - 1].vol[q] * st.apa[i_part - 1].vol[q];
      if (i_samp <= n_samp_remove) {
        const double num_conc_i = awa_num_conc(st.awa, st.apa[i_part - 1]);
        const double prob_remove_i = num_conc_target / num_conc_i;
        if (ss.next() < prob_remove_i / prob_remove_source_max) {
          n_remove = n_remove + 1;
          aero_state_remove_particle(st, i_part);
        }
      }
    }
  }
  if (n_avg == 0) {
    n_coag = 0;
    return;
  }
  const double prob_coag_mean = prob_coag_tot / (double)n_samp;
  n_coag = rand_binomial(n_samp_total, prob_coag_mean, s);
  const double f = (double)n_coag / (double)n_avg;
  for (int q = 0; q < S; q++) source.vol[q] = source.vol[q] * f;
  source.n_primary_parts = prob_round((double)source.n_primary_parts * ((double)n_coag / n_avg), s.next());
  // components of the averaged source (:472-508): down to min(n_primary_parts, 20) by sampling without replacement,
  // or up to it by whole copies plus a sample of the rest (draws: PH_COMP, key of the sample stream, j = 1 + t)
  {
    const int n_comp = (int)source.comp_src.size();
    const int n_comp_new = std::min(source.n_primary_parts, MAX_AERO_COMPONENT_SIZE);
    RngKey ck = ss.key;
    ck.phase = PH_COMP;
    Stream sc(ss.rng, ck, 1u);
    std::vector<int> nsrc;
    std::vector<double> ntime;
    if (n_comp_new < n_comp) {
      for (int i : sample_without_replacement(n_comp_new, n_comp, sc)) {
        nsrc.push_back(source.comp_src[i - 1]);
        ntime.push_back(source.comp_time[i - 1]);
      }
      source.comp_src = nsrc;
      source.comp_time = ntime;
    } else if (n_comp < MAX_AERO_COMPONENT_SIZE && n_comp > 0) {
      int i_comp = 1;
      while (i_comp <= n_comp_new) {
        if (i_comp + n_comp <= n_comp_new) {
          for (int i = 1; i <= n_comp; i++) {
            nsrc.push_back(source.comp_src[i - 1]);
            ntime.push_back(source.comp_time[i - 1]);
          }
          i_comp = i_comp + n_comp;
        } else {
          const int sample_size = n_comp_new - i_comp + 1;
          for (int i : sample_without_replacement(sample_size, n_comp, sc)) {
            nsrc.push_back(source.comp_src[i - 1]);
            ntime.push_back(source.comp_time[i - 1]);
          }
          i_comp = n_comp_new + 1;
        }
      }
      source.comp_src = nsrc;
      source.comp_time = ntime;
    }
  }
}
// coagulation.F90:522-582
static void coag_target_with_source(AeroState& st, int bt, int ct, int target_unif_entry, const Particle& source,
                                    int cc, Stream& s, const RngKey& ckey) {
  const int target_part = st.sorted.size_class.inv(bt, ct).at(target_unif_entry);
  Particle& tp = st.apa[target_part - 1];
  const int64_t target_id = tp.id;
  const double old_num_conc_target = awa_num_conc(st.awa, tp);
  Particle merged;
  particle_coagulate(tp, source, merged, s.rng, ckey, 64u);
  merged.id = target_id;
  const int new_group = awa_rand_group(st.awa, cc, particle_radius(merged), s);
  merged.weight_group = new_group;
  merged.weight_class = cc;
  st.apa[target_part - 1] = merged;
  const int new_bin = aero_sorted_particle_in_bin(st.sorted, st.apa[target_part - 1]);
  if ((new_bin < 1) || (new_bin > st.sorted.bin_grid.size()))
    die_msg(765620746, "particle outside of bin_grid: try reducing the timestep del_t");
  st.sorted.size_class.change(target_part, new_bin, cc);
  st.sorted.group_class.change(target_part, new_group, cc);
  const double new_num_conc_target = awa_num_conc(st.awa, st.apa[target_part - 1]);
  aero_state_dup_particle(st, target_part, old_num_conc_target / new_num_conc_target, true, s);
}
// coagulation.F90:150-241
static bool try_per_particle_coag(int type, double k_max, const EnvState& env, const AeroData& d, AeroState& st,
                                  double del_t, const RunCtx& rc, CoagStats& out, int b1, int b2, int c1, int c2,
                                  int cc) {
  int bt = 0, bs = 0, ct = 0, cs = 0;
  bool ok;
  determine_target_and_source(st.awa, st.sorted.bin_grid, b1, b2, c1, c2, cc, bt, bs, ct, cs, ok);
  if (!ok) return false;
  // compute_n_source, coagulation.F90:310-329
  const double n_source_per_target = k_max * del_t * (double)st.sorted.size_class.inv(bs, cs).n_entry();
  const double accept_factor = 1.0 / k_max;
  if (n_source_per_target < COAG_ACCEL_N_EVENT) return false;
  const double min_target_vol = aero_rad2vol(st.sorted.bin_grid.edges[bt - 1]);
  const double max_source_vol = aero_rad2vol(st.sorted.bin_grid.edges[bs]);
  const double max_new_target_vol = min_target_vol + max_source_vol * n_source_per_target;
  const double max_target_growth_factor = max_new_target_vol / min_target_vol;
  if (max_target_growth_factor > MAX_ALLOWABLE_GROWTH_FACTOR) return false;
  const uint32_t pid = coag_pair_id(st.awa.n_class, c1, c2, b1, b2);
  for (int e = st.sorted.size_class.inv(bt, ct).n_entry(); e >= 1; e--) {
    const int target_part = st.sorted.size_class.inv(bt, ct).at(e);
    const Particle target_particle = st.apa[target_part - 1];
    Stream s(rc.rng, rc.key(PH_COAG_ACCEL, pid, (uint32_t)e));
    Stream ss(rc.rng, rc.key(PH_COAG_ACCEL, pid | 0x80000000u, (uint32_t)e));
    int n_samp, n_coag, n_remove;
    Particle source;
    sample_source_particle(st, d, env, type, bs, cs, target_particle, n_source_per_target, accept_factor, s, ss, n_samp,
                           n_coag, n_remove, source);
    if (n_coag > 0) coag_target_with_source(st, bt, ct, e, source, cc, s, ss.key);
    out.tot_n_samp += n_samp;
    out.tot_n_coag += n_coag;
  }
  out.n_accel_pairs += 1;
  return true;
}
// coagulation.F90:101-145
static void mc_coag_for_bin(int type, const EnvState& env, const AeroData& d, AeroState& st, double del_t,
                            const RunCtx& rc, CoagStats& out, int b1, int b2, int c1, int c2) {
  const int nb = st.sorted.n_bin();
  const int cc = coag_dest_class(st.awa, st.sorted.bin_grid, b1, b2, c1, c2);
  const double f_max = max_coag_num_conc_factor(st.awa, st.sorted.bin_grid, b1, b2, c1, c2, cc);
  const double k_max = st.sorted.coag_kernel_max[(size_t)(b2 - 1) * nb + (b1 - 1)] * f_max;
  if (try_per_particle_coag(type, k_max, env, d, st, del_t, rc, out, b1, b2, c1, c2, cc)) return;
  per_set_coag(type, k_max, env, d, st, del_t, rc, out, b1, b2, c1, c2, cc);
}
void mc_coag(int type, const EnvState& env, const AeroData& d, AeroState& st, double del_t, const RunCtx& rc,
             CoagStats& out) {
  aero_state_sort(st);
  if (!st.sorted.coag_kernel_bounds_valid) {
    est_k_minmax_binned_unweighted(st.sorted.bin_grid, type, d, env, st.sorted.coag_kernel_min,
                                   st.sorted.coag_kernel_max);
    st.sorted.coag_kernel_bounds_valid = true;
  }
  out.tot_n_samp = 0;
  out.tot_n_coag = 0;
  for (int c1 = 1; c1 <= st.sorted.n_class(); c1++)
    for (int c2 = 1; c2 <= c1; c2++)
      for (int b1 = 1; b1 <= st.sorted.n_bin(); b1++) {
        if (st.sorted.size_class.inv(b1, c1).n_entry() == 0) continue;
        const int b2_start = (c1 == c2) ? b1 : 1;
        for (int b2 = b2_start; b2 <= st.sorted.n_bin(); b2++) {
          if (st.sorted.size_class.inv(b2, c2).n_entry() == 0) continue;
          mc_coag_for_bin(type, env, d, st, del_t, rc, out, b1, b2, c1, c2);
        }
      }
}

}  // namespace pmco
