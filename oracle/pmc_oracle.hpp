// pmc_oracle.hpp -- CPU restatement ("oracle") of the PartMC per-grid-cell
// aerosol step that WRF-PartMC runs once per cell per chemistry timestep.
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, bench.py's
// cpu_baseline / --impl reference leg and __graft_entry__.smoke() may link or
// call it, and only as the checker.  The product (wrf-partmc_b200/csrc) never
// includes this header and has no CPU fallback.
//
// Every function cites the reference file:line it follows (paths relative to
// the reference tree: partmc/src/... and interface/...).  The reference is
// Fortran; no Fortran compiler, GSL, SUNDIALS or netCDF exists in the build
// container, so the reference itself cannot be compiled (SURVEY
// F5).  Parity pins: Golovin closed form (coag_kernel_additive.F90:101-167),
// test/condense/ref_condense_0001_env.txt, test/rand pmf formulas, and the
// structural invariants of aero_sorted_check / normalize_probs.
//
// Index conventions: bins, classes, groups and particle numbers are 1-based in
// every signature here, exactly as in the Fortran, so that call sites can be
// compared line by line.
#pragma once
#include <cstdint>
#include <cmath>
#include <vector>
#include <string>
#include <stdexcept>

namespace pmco {

// constants.F90:30-65
struct Const {
  static constexpr double pi = 3.14159265358979323846;
  static constexpr double boltzmann = 1.3806505e-23;
  static constexpr double avagadro = 6.02214179e23;
  static constexpr double air_dyn_visc = 1.78e-5;  // constants.F90:65
  static constexpr double univ_gas_const = 8.314472;
  static constexpr double accom_coeff = 1.0;
  static constexpr double water_eq_vap_press = 611.0;
  static constexpr double water_freeze_temp = 273.15;
  static constexpr double water_surf_eng = 0.073;
  static constexpr double water_latent_heat = 2.272e6;
  static constexpr double air_spec_heat = 1005.0;
  static constexpr double water_molec_weight = 18e-3;
  static constexpr double water_density = 1e3;
  static constexpr double air_molec_weight = 2.89644e-2;
  static constexpr double air_std_press = 101325.0;
};

// util.F90:77-145: assert / die_msg carry an integer code.
struct PmcError : std::runtime_error {
  int code;
  PmcError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
inline void pmc_assert(int code, bool cond) {
  if (!cond) throw PmcError(code, "assertion failed: " + std::to_string(code));
}
void die_msg(int code, const std::string& msg);

// ---------------------------------------------------------------- RNG
// rand.F90 has ONE global stream (MT19937 via GSL in the WRF build,
// rand_gsl.c:46-58).  A counter-based device generator cannot reproduce that
// consumption order, so every draw in the oracle is *addressed*: a key names
// the logical stream (cell, step, phase, a, b) and j is the position inside
// it.  The MT back end ignores the address and hands out the next value of
// the single global stream in program order (== reference order); the Philox
// back end is a pure function of (seed, key, j), which is what the CUDA
// kernels evaluate.
enum Phase : uint32_t {
  PH_COAG = 1,        // a = bin-pair id, b = 0 (n_samp) or sample number d>=1
  PH_COAG_ACCEL = 2,  // a = bin-pair id, b = target entry
  PH_REWEIGHT = 3,    // a = pass, b = particle index
  PH_HALVE = 4,       // a = pass*4096 + (group,class), b = particle index
  PH_DOUBLE = 5,
  PH_SCALE = 6,       // a = (group,class), b = 0 (count) / removal number
  PH_TRANS = 7,       // a = class*64 + direction slot, b = 0 (count) / transfer
  PH_INIT = 8,
  PH_DEP = 9,         // a = 0, b = particle index (dry deposition)
  PH_EMIT = 10,       // a = (group, mode), b = 0 (count) / sample number (aero_state_add_aero_dist_sample)
  PH_COMP = 12,       // subsampling of component lists: (a, b) of the coagulation event (accelerated path: a | 2^31),
                      // j = j0 + {0 prob_round, 1 + t first sample_without_replacement, 33 + t second}; j0 = 0 event,
                      // accelerated path: 128 i_samp per accepted sample, 0 the resampling, 64 the target merge
  PH_BND = 11         // aero_state_resample of a boundary cell: a = (group, class): b = 0 count, b >= 1 removal;
                      // a = 1000 + (group, class) n_mode + mode: b = 0 count of the mode, b >= 1 new particle
};
struct RngKey {
  uint32_t cell, step, phase, a, b;
};
struct Rng {
  virtual ~Rng() {}
  virtual double uniform(const RngKey& k, uint32_t j) = 0;
};
// Philox4x32-10 (Salmon et al., SC'11) -- published algorithm, restated.
void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
double philox_uniform(uint32_t seed, const RngKey& k, uint32_t j);
struct RngPhilox : Rng {
  uint32_t seed;
  explicit RngPhilox(uint32_t s) : seed(s) {}
  double uniform(const RngKey& k, uint32_t j) override { return philox_uniform(seed, k, j); }
};
// MT19937 in the GSL convention (gsl_rng_uniform: x / 2^32), rand_gsl.c:77-90.
struct RngMT : Rng {
  uint32_t mt[624];
  int mti;
  explicit RngMT(uint32_t s);
  uint32_t next_u32();
  double uniform(const RngKey&, uint32_t) override { return next_u32() / 4294967296.0; }
};
// sequential cursor over one addressed stream
struct Stream {
  Rng* rng;
  RngKey key;
  uint32_t j;
  Stream(Rng* r, RngKey k, uint32_t j0 = 0) : rng(r), key(k), j(j0) {}
  double next() { return rng->uniform(key, j++); }
  double at(uint32_t jj) { return rng->uniform(key, jj); }
};
// rand.F90 samplers (the in-tree, non-GSL definitions)
int pmc_rand_int(double u, int n);                 // rand.F90:172-205
int prob_round(double val, double u);              // rand.F90:214-226
int rand_poisson(double mean, Stream& s);          // rand.F90:252-304
int rand_binomial(int n, double p, Stream& s);     // rand.F90:319-404
double rand_normal(double mean, double stddev, Stream& s);  // rand.F90:410-474
int sample_cts_pdf(const std::vector<double>& pdf, Stream& s);  // rand.F90:550-575
double poisson_pmf(double mean, int k);            // test/rand/test_poisson_sample.F90
double binomial_pmf(int n, double p, int k);       // test/rand/test_binomial_sample.F90

// ---------------------------------------------------------------- util.F90
double sphere_vol2rad(double v);                   // util.F90:240
double sphere_rad2vol(double r);                   // util.F90:265
// ---------------------------------------------------------------- fractal.F90
// aero_data%fractal (fractal.F90:55-66): ONE setting per run in the reference; here one per process (set_fractal),
// read by every radius <-> volume conversion of the sort / coagulation path (aero_data_vol2rad, aero_data_rad2vol,
// aero_data.F90:101-140).  Defaults = fractal_set_spherical (fractal.F90:73-82).  The mobility-radius conversions
// (fractal.F90:239-440) that the Brownian and sedimentation kernels and dry deposition apply to non-spherical
// particles are NOT restated: those functions die when the setting is not spherical.
struct Fractal {
  double frac_dim = 3.0, prime_radius = 1e-8, vol_fill_factor = 1.0;
};
void set_fractal(double frac_dim, double prime_radius, double vol_fill_factor);
const Fractal& fractal();
bool fractal_is_spherical();                        // fractal.F90:296-307
double fractal_vol_to_num_of_monomers(double v);    // fractal.F90:90-100
double aero_vol2rad(double v);                      // fractal_vol2rad, fractal.F90:106-121
double aero_rad2vol(double r);                      // fractal_rad2vol, fractal.F90:142-158
std::vector<double> linspace(double min_x, double max_x, int n);  // util.F90:460
std::vector<double> logspace(double min_x, double max_x, int n);  // util.F90:492
int linspace_find(double min_x, double max_x, int n, double x);   // util.F90:532
int logspace_find(double min_x, double max_x, int n, double x);   // util.F90:566
double interp_linear_disc(double x_1, double x_n, int n, int i);  // util.F90:663
double student_t_95_coeff(int n_sample);           // stats.F90:510-546

// ---------------------------------------------------------------- data types
struct AeroData {  // aero_data.F90:55-93 (fields the path reads)
  int n_spec = 0;
  std::vector<double> density, kappa, molec_weight;
  std::vector<int> num_ions;
  int i_water = 0;  // 1-based, 0 = none
};
struct EnvState {  // env_state.F90:29-76
  double temp = 288.0, pressure = 1e5, rel_humid = 0.5;
  double inverse_density = 1.0, cell_volume = 1.0, elapsed_time = 0.0;
};
enum { WEIGHT_NONE = 1, WEIGHT_POWER = 2, WEIGHT_MFA = 3 };  // aero_weight.F90:25-30
struct Weight {
  int type = WEIGHT_NONE;
  double magnitude = 1.0, exponent = 0.0;
};
struct WeightArray {  // aero_weight_array.F90:35-39, weight(n_group, n_class)
  int n_group = 1, n_class = 1;
  std::vector<Weight> w;
  void set_sizes(int g, int c) { n_group = g; n_class = c; w.assign((size_t)g * c, Weight()); }
  Weight& at(int g, int c) { return w[(size_t)(c - 1) * n_group + (g - 1)]; }
  const Weight& at(int g, int c) const { return w[(size_t)(c - 1) * n_group + (g - 1)]; }
};
struct Particle {  // aero_particle.F90:30-61 (hot fields; optics/components are cold)
  std::vector<double> vol;
  int weight_group = 0, weight_class = 0;
  int water_hyst_leg = 0;
  int64_t id = 0;
  double least_create_time = 0.0, greatest_create_time = 0.0;
  int n_primary_parts = 1;
  // cold part (aero_particle.F90:38-54): component(:) = {source_id, create_time} pairs, at most
  // MAX_AERO_COMPONENT_SIZE = 20 (:22), and the optical fields as an opaque vector
  std::vector<int> comp_src;
  std::vector<double> comp_time;
  std::vector<double> opq;
};
static const int MAX_AERO_COMPONENT_SIZE = 20;
struct BinGrid {  // bin_grid.F90 (log type only, as aero_sorted uses)
  std::vector<double> edges, centers;
  int size() const { return edges.empty() ? 0 : (int)edges.size() - 1; }
};
struct IntVarray {  // integer_varray.F90
  std::vector<int> entry;
  int n_entry() const { return (int)entry.size(); }
  void append(int v) { entry.push_back(v); }
  void remove_entry(int index);  // integer_varray.F90:151-169
  int& at(int i) { return entry[i - 1]; }
  int at(int i) const { return entry[i - 1]; }
};
struct Rmap2 {  // integer_rmap2.F90
  IntVarray forward1, forward2, index;
  int n1 = 0, n2 = 0;
  std::vector<IntVarray> inverse;
  IntVarray& inv(int i1, int i2) { return inverse[(size_t)(i2 - 1) * n1 + (i1 - 1)]; }
  const IntVarray& inv(int i1, int i2) const { return inverse[(size_t)(i2 - 1) * n1 + (i1 - 1)]; }
  void set_ranges(int a, int b);
  void zero();
  void append(int i1, int i2);            // :107
  void change(int i_domain, int i1, int i2);  // :135-185 (keeps the :160 quirk)
  void remove(int i_domain);              // :190-243
  void check(int n_domain) const;         // :248
};
struct AeroSorted {  // aero_sorted.F90:47-66
  BinGrid bin_grid;
  Rmap2 size_class, group_class;
  bool coag_kernel_bounds_valid = false;
  std::vector<double> coag_kernel_min, coag_kernel_max;  // (b1,b2) -> [(b2-1)*nb + b1-1]
  int n_bin() const { return size_class.n1; }
  int n_class() const { return size_class.n2; }
  int n_group() const { return group_class.n1; }
};
// aero_particle.F90:141-149: id = (next_id - 1) * size + rank + 1.  The device
// keeps one counter per grid cell (size = number of global cells, rank = global
// cell index) so that ids are unique without a cross-cell atomic.
struct IdGen {
  int64_t next_id = 1, size = 1, rank = 0;
  int64_t new_id() { int64_t v = (next_id - 1) * size + rank + 1; next_id++; return v; }
};
struct AeroState {  // aero_state.F90:69-86
  std::vector<Particle> apa;
  AeroSorted sorted;
  bool valid_sort = false;
  WeightArray awa;
  std::vector<double> n_part_ideal;  // (g,c) like awa
  IdGen idgen;
  int n_part() const { return (int)apa.size(); }
  double& ideal(int g, int c) { return n_part_ideal[(size_t)(c - 1) * awa.n_group + (g - 1)]; }
};
struct RunCtx {  // what every stochastic routine needs to address its draws
  Rng* rng;
  uint32_t cell, step;
  RngKey key(uint32_t phase, uint32_t a, uint32_t b) const { return RngKey{cell, step, phase, a, b}; }
};

// ---------------------------------------------------------------- bin grid / weights
void bin_grid_make_log(BinGrid& g, int n_bin, double min, double max);  // bin_grid.F90:83-134
int bin_grid_find(const BinGrid& g, double val);                        // bin_grid.F90:143-166
double weight_num_conc_at_radius(const Weight& w, double radius);       // aero_weight.F90:127-146
double awa_num_conc_at_radius(const WeightArray& a, int i_class, double radius);  // aero_weight_array.F90:232-251
void awa_check_monotonicity(const WeightArray& a, bool& inc, bool& dec);          // :304-330
void awa_minmax_num_conc(const WeightArray& a, int i_class, double r1, double r2, double& mn, double& mx);  // :338
int awa_rand_group(const WeightArray& a, int i_class, double radius, Stream& s);  // :374-392

// ---------------------------------------------------------------- particle
double particle_volume(const Particle& p);                       // aero_particle.F90:316
double particle_radius(const Particle& p);                       // :388-399 (aero_data_vol2rad)
double particle_mass(const Particle& p, const AeroData& d);      // :252
double particle_density(const Particle& p, const AeroData& d);   // :474
double particle_solute_volume(const Particle& p, const AeroData& d);  // :687
double particle_solute_kappa(const Particle& p, const AeroData& d);   // :724-755
void particle_coagulate(const Particle& p1, const Particle& p2, Particle& pn, Rng* rng = nullptr,
                        RngKey ckey = RngKey{0, 0, 0, 0, 0}, uint32_t j0 = 0);  // :900-980
std::vector<int> sample_without_replacement(int sample_size, int population_size, Stream& s);  // rand.F90:515-542
double awa_num_conc(const WeightArray& a, const Particle& p);    // aero_weight_array.F90:257
double awa_single_num_conc(const WeightArray& a, const Particle& p);  // :218-227

// ---------------------------------------------------------------- coagulation kernels
enum { KERNEL_SEDI = 1, KERNEL_ADDITIVE = 2, KERNEL_CONSTANT = 3, KERNEL_BROWN = 4, KERNEL_ZERO = 5,
       KERNEL_BROWN_FREE = 6, KERNEL_BROWN_CONT = 7 };  // coag_kernel.F90:29-45
double kernel_brown_free_helper(double v1, double d1, double v2, double d2, double temp);  // coag_kernel_brown_free.F90:118-147
double kernel_brown_cont_helper(double v1, double d1, double v2, double d2, double temp);  // coag_kernel_brown_cont.F90:117-144
double kernel_sedi_helper(double v1, double v2);  // coag_kernel_sedi.F90:79-103
double kernel_brown_helper(double vol_i, double den_i, double vol_j, double den_j, double temp,
                           double pressure);  // coag_kernel_brown.F90:111-197
void kernel_minmax(int type, double v1, double v2, const AeroData& d, const EnvState& e, double& kmin,
                   double& kmax);  // coag_kernel.F90:128-165
double kernel(int type, const Particle& p1, const Particle& p2, const AeroData& d,
              const EnvState& e);  // coag_kernel.F90:83-125
void est_k_minmax_binned_unweighted(const BinGrid& g, int type, const AeroData& d, const EnvState& e,
                                    std::vector<double>& kmin, std::vector<double>& kmax);  // :253-280
double coag_num_conc_factor(const WeightArray& a, double i_r, double j_r, int ic, int jc, int ijc);  // :346-377
int coag_dest_class(const WeightArray& a, const BinGrid& g, int ib, int jb, int ic, int jc);  // :382-416
double max_coag_num_conc_factor(const WeightArray& a, const BinGrid& g, int ib, int jb, int ic, int jc,
                                int ijc);  // :422-465
double num_conc_weighted_kernel(int type, const Particle& p1, const Particle& p2, int ic, int jc, int ijc,
                                const AeroData& d, const WeightArray& a, const EnvState& e);  // :175-209
// Golovin solution, coag_kernel_additive.F90:101-167 (+ bessi1 :172-381)
double bessi1(double x);
void soln_additive_exp(const BinGrid& g, double time, double num_conc, double radius_at_mean_vol,
                       std::vector<double>& num_conc_per_lnr, std::vector<double>& vol_conc_per_lnr);

// ---------------------------------------------------------------- aero_sorted / aero_state
void aero_sorted_set_bin_grid(AeroSorted& s, const BinGrid& g, int n_group, int n_class);  // aero_sorted.F90:122-146
int aero_sorted_particle_in_bin(const AeroSorted& s, const Particle& p);                   // :378
void aero_sorted_sort_particles(AeroSorted& s, const std::vector<Particle>& apa);          // :185-209
// :214-373.  Returns true when a new bin grid was made.
bool aero_sorted_remake_if_needed(AeroSorted& s, std::vector<Particle>& apa, bool valid_sort, int n_group,
                                  int n_class);
void aero_sorted_check(const AeroSorted& s, const std::vector<Particle>& apa, int n_group, int n_class);  // :499
void aero_state_sort(AeroState& st);                                         // aero_state.F90:3218
// Device layout helper: permute apa into (class, bin, original index) order so
// that every size_class list is a contiguous run.  apa order carries no meaning
// in the reference (aero_particle_array.F90:178-200 reorders it freely).
void aero_state_physical_sort(AeroState& st);
int aero_state_total_particles(const AeroState& st, int g, int c);           // :287
void aero_state_add_particle(AeroState& st, const Particle& p, bool allow_resort = true);  // :365
void aero_state_remove_particle(AeroState& st, int i_part);                  // :389-446
void aero_state_dup_particle(AeroState& st, int i_part, double n_part_mean, bool random_weight_group,
                             Stream& s);                                     // :488-539
void aero_state_double(AeroState& st, int g, int c);                         // :1695
void aero_state_halve(AeroState& st, int g, int c, const RunCtx& rc, uint32_t pass);  // :1726
void aero_state_rebalance(AeroState& st, bool allow_doubling, bool allow_halving, const RunCtx& rc);  // :1765-1825
// salt: distinguishes the draw addresses of several calls for the same (group, class) inside one step
// (emissions call it once per mode); 0 for a stand-alone call
void aero_state_scale_weight(AeroState& st, int g, int c, double weight_ratio, bool allow_doubling,
                             bool allow_halving, const RunCtx& rc, uint32_t salt = 0);          // :1832-1895
void aero_state_reweight(AeroState& st, const std::vector<double>& reweight_num_conc, const RunCtx& rc,
                         uint32_t pass);                                     // :597-653

// ---------------------------------------------------------------- coagulation.F90
struct CoagStats {
  int tot_n_samp = 0, tot_n_coag = 0;
  int n_accel_pairs = 0;  // bin pairs handled by try_per_particle_coag
};
void mc_coag(int kernel_type, const EnvState& env, const AeroData& d, AeroState& st, double del_t,
             const RunCtx& rc, CoagStats& out);  // coagulation.F90:43-96
uint32_t coag_pair_id(int n_class, int c1, int c2, int b1, int b2);

}  // namespace pmco
