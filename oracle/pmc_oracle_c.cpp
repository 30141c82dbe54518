// pmc_oracle_c.cpp -- C entry points of the CPU oracle, for ctypes.
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).  The shape of this API mirrors
// include/pmcgpu.h so that the parity tests drive both with the same arrays.
#include "pmc_oracle.hpp"
#include "pmc_oracle_ext.hpp"
#include "pmc_oracle_domain.hpp"
#include <cstring>
#include <memory>
#include <thread>

using namespace pmco;

extern "C" {

void* pmco_create(int n_cells, int n_spec, const double* density, const double* kappa, const double* molec_weight,
                  const int* num_ions, int i_water, int n_group, int n_class, double n_part_ideal, int rng_kind,
                  unsigned seed, long long n_cells_global, const long long* global_cell) {
  OracleDomain* d = new OracleDomain();
  d->data.n_spec = n_spec;
  d->data.density.assign(density, density + n_spec);
  d->data.kappa.assign(kappa, kappa + n_spec);
  d->data.molec_weight.assign(molec_weight, molec_weight + n_spec);
  d->data.num_ions.assign(num_ions, num_ions + n_spec);
  d->data.i_water = i_water;
  d->cells.resize((size_t)n_cells);
  d->env.resize((size_t)n_cells);
  d->global_cell.resize((size_t)n_cells);
  for (int c = 0; c < n_cells; c++) {
    AeroState& st = d->cells[c];
    st.awa.set_sizes(n_group, n_class);
    // aero_state_set_n_part_ideal, aero_state.F90:241-256
    st.n_part_ideal.assign((size_t)n_group * n_class, n_part_ideal / (double)(n_group * n_class));
    d->global_cell[c] = global_cell ? global_cell[c] : c;
    st.idgen.size = n_cells_global;
    st.idgen.rank = d->global_cell[c];
    st.idgen.next_id = 1;
  }
  if (rng_kind == 0)
    d->rng.reset(new RngPhilox(seed));
  else
    d->rng.reset(new RngMT(seed));
  return d;
}
void pmco_destroy(void* h) { delete (OracleDomain*)h; }
const char* pmco_last_error(void* h) { return ((OracleDomain*)h)->last_error.c_str(); }

int pmco_set_cell(void* h, int cell, int n, const double* vol, const int* wgroup, const int* wclass,
                  const long long* id, const int* n_primary, const int* hyst, const double* lct,
                  const double* gct) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  AeroState& st = d->cells[cell];
  const int S = d->data.n_spec;
  st.apa.assign((size_t)n, Particle());
  for (int i = 0; i < n; i++) {
    Particle& p = st.apa[i];
    p.vol.resize((size_t)S);
    for (int s = 0; s < S; s++) p.vol[s] = vol[(size_t)s * n + i];
    p.weight_group = wgroup ? wgroup[i] : 1;
    p.weight_class = wclass ? wclass[i] : 1;
    p.id = id ? id[i] : 0;
    p.n_primary_parts = n_primary ? n_primary[i] : 1;
    p.water_hyst_leg = hyst ? hyst[i] : 0;
    p.least_create_time = lct ? lct[i] : 0.0;
    p.greatest_create_time = gct ? gct[i] : 0.0;
    p.comp_src.assign(1, p.weight_class);  // default cold part, like pmcgpu_set_particles
    p.comp_time.assign(1, p.least_create_time);
  }
  st.valid_sort = false;
  ORACLE_CATCH(d)
}
// components / opaque optics of a cell's particles: n_comp [n], source / create_time [n][20], opaque [n][n_opq]
int pmco_set_components(void* h, int cell, const int* n_comp, const int* source, const double* create_time,
                        int n_opq, const double* opaque) {
  AeroState& st = ((OracleDomain*)h)->cells[cell];
  for (int i = 0; i < st.n_part(); i++) {
    Particle& p = st.apa[i];
    const int nc = std::max(0, std::min(n_comp[i], MAX_AERO_COMPONENT_SIZE));
    p.comp_src.assign(source + (size_t)i * MAX_AERO_COMPONENT_SIZE, source + (size_t)i * MAX_AERO_COMPONENT_SIZE + nc);
    p.comp_time.assign(create_time + (size_t)i * MAX_AERO_COMPONENT_SIZE,
                       create_time + (size_t)i * MAX_AERO_COMPONENT_SIZE + nc);
    if (opaque && n_opq > 0) p.opq.assign(opaque + (size_t)i * n_opq, opaque + (size_t)(i + 1) * n_opq);
  }
  return 0;
}
int pmco_get_components(void* h, int cell, int* n_comp, int* source, double* create_time, int n_opq, double* opaque) {
  const AeroState& st = ((OracleDomain*)h)->cells[cell];
  for (int i = 0; i < st.n_part(); i++) {
    const Particle& p = st.apa[i];
    const int nc = (int)p.comp_src.size();
    n_comp[i] = nc;
    for (int q = 0; q < MAX_AERO_COMPONENT_SIZE; q++) {
      source[(size_t)i * MAX_AERO_COMPONENT_SIZE + q] = q < nc ? p.comp_src[q] : 0;
      create_time[(size_t)i * MAX_AERO_COMPONENT_SIZE + q] = q < nc ? p.comp_time[q] : 0.0;
    }
    if (opaque)
      for (int q = 0; q < n_opq; q++) opaque[(size_t)i * n_opq + q] = q < (int)p.opq.size() ? p.opq[q] : 0.0;
  }
  return 0;
}
int pmco_set_next_id(void* h, int cell, long long next_id) {
  ((OracleDomain*)h)->cells[cell].idgen.next_id = next_id;
  return 0;
}
long long pmco_get_next_id(void* h, int cell) { return ((OracleDomain*)h)->cells[cell].idgen.next_id; }
int pmco_cell_count(void* h, int cell) { return ((OracleDomain*)h)->cells[cell].n_part(); }
int pmco_get_cell(void* h, int cell, double* vol, int* wgroup, int* wclass, long long* id, int* n_primary, int* hyst,
                  double* lct, double* gct) {
  OracleDomain* d = (OracleDomain*)h;
  const AeroState& st = d->cells[cell];
  const int S = d->data.n_spec, n = st.n_part();
  for (int i = 0; i < n; i++) {
    const Particle& p = st.apa[i];
    if (vol)
      for (int s = 0; s < S; s++) vol[(size_t)s * n + i] = p.vol[s];
    if (wgroup) wgroup[i] = p.weight_group;
    if (wclass) wclass[i] = p.weight_class;
    if (id) id[i] = p.id;
    if (n_primary) n_primary[i] = p.n_primary_parts;
    if (hyst) hyst[i] = p.water_hyst_leg;
    if (lct) lct[i] = p.least_create_time;
    if (gct) gct[i] = p.greatest_create_time;
  }
  return 0;
}
// weights in Fortran (group, class) order: index (c-1)*G + (g-1)
int pmco_set_weights(void* h, int cell, const int* type, const double* magnitude, const double* exponent) {
  AeroState& st = ((OracleDomain*)h)->cells[cell];
  for (size_t k = 0; k < st.awa.w.size(); k++) {
    st.awa.w[k].type = type[k];
    st.awa.w[k].magnitude = magnitude[k];
    st.awa.w[k].exponent = exponent[k];
  }
  return 0;
}
int pmco_get_weights(void* h, int cell, int* type, double* magnitude, double* exponent) {
  const AeroState& st = ((OracleDomain*)h)->cells[cell];
  for (size_t k = 0; k < st.awa.w.size(); k++) {
    if (type) type[k] = st.awa.w[k].type;
    if (magnitude) magnitude[k] = st.awa.w[k].magnitude;
    if (exponent) exponent[k] = st.awa.w[k].exponent;
  }
  return 0;
}
int pmco_set_env(void* h, int cell, double T, double p, double RH, double rrho, double cell_volume,
                 double elapsed_time) {
  EnvState& e = ((OracleDomain*)h)->env[cell];
  e.temp = T;
  e.pressure = p;
  e.rel_humid = RH;
  e.inverse_density = rrho;
  e.cell_volume = cell_volume;
  e.elapsed_time = elapsed_time;
  return 0;
}
double pmco_get_rel_humid(void* h, int cell) { return ((OracleDomain*)h)->env[cell].rel_humid; }

// aero_state_sort (aero_state.F90:3218).  force_resort=1 models an entry state
// with valid_sort = .false. (what every device step starts from); physical=1
// additionally permutes apa into the device's (class, bin, index) order.
// edge_recheck=1 repeats aero_state_sort on the now valid state, which is what the
// aero_state_sort call at the top of mc_coag (coagulation.F90:63) does after a
// preceding full sort: the safety-factor test on occupied-bin edges
// (aero_sorted.F90:258-276), possibly a new grid and a re-append in apa order.
int pmco_sort(void* h, int cell, int force_resort, int physical, int edge_recheck) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  AeroState& st = d->cells[cell];
  if (force_resort) st.valid_sort = false;
  aero_state_sort(st);
  if (edge_recheck) aero_state_sort(st);
  if (physical) aero_state_physical_sort(st);
  aero_sorted_check(st.sorted, st.apa, st.awa.n_group, st.awa.n_class);
  ORACLE_CATCH(d)
}
int pmco_get_grid(void* h, int cell, int* n_bin, double* grid_min, double* grid_max, double* edges, int edges_cap) {
  const AeroSorted& s = ((OracleDomain*)h)->cells[cell].sorted;
  const int nb = s.bin_grid.size();
  *n_bin = nb;
  if (nb > 0) {
    *grid_min = s.bin_grid.edges[0];
    *grid_max = s.bin_grid.edges[nb];
    if (edges)
      for (int i = 0; i <= nb && i < edges_cap; i++) edges[i] = s.bin_grid.edges[i];
  }
  return 0;
}
int pmco_get_bins(void* h, int cell, int* bin_of_particle) {
  const AeroState& st = ((OracleDomain*)h)->cells[cell];
  if (!st.valid_sort) return -2;
  for (int i = 1; i <= st.n_part(); i++) bin_of_particle[i - 1] = st.sorted.size_class.forward1.at(i);
  return 0;
}
// counts(b, c) -> out[(c-1)*nb + (b-1)]
int pmco_get_bin_counts(void* h, int cell, int* out) {
  const AeroState& st = ((OracleDomain*)h)->cells[cell];
  if (!st.valid_sort) return -2;
  const int nb = st.sorted.n_bin(), nc = st.sorted.n_class();
  for (int c = 1; c <= nc; c++)
    for (int b = 1; b <= nb; b++) out[(size_t)(c - 1) * nb + (b - 1)] = st.sorted.size_class.inv(b, c).n_entry();
  return 0;
}
// list (b,c) entries (1-based particle numbers) in list order
int pmco_get_bin_list(void* h, int cell, int b, int c, int* out) {
  const AeroState& st = ((OracleDomain*)h)->cells[cell];
  const IntVarray& l = st.sorted.size_class.inv(b, c);
  for (int k = 1; k <= l.n_entry(); k++) out[k - 1] = l.at(k);
  return l.n_entry();
}
// est_k_minmax_binned_unweighted on the cell's current grid; out[(b2-1)*nb + b1-1]
int pmco_kernel_bounds(void* h, int cell, int kernel_type, double* kmin, double* kmax) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  AeroState& st = d->cells[cell];
  std::vector<double> mn, mx;
  est_k_minmax_binned_unweighted(st.sorted.bin_grid, kernel_type, d->data, d->env[cell], mn, mx);
  st.sorted.coag_kernel_min = mn;
  st.sorted.coag_kernel_max = mx;
  st.sorted.coag_kernel_bounds_valid = true;
  if (kmin) std::memcpy(kmin, mn.data(), mn.size() * sizeof(double));
  if (kmax) std::memcpy(kmax, mx.data(), mx.size() * sizeof(double));
  ORACLE_CATCH(d)
}
double pmco_kernel_sedi(double v1, double v2) { return kernel_sedi_helper(v1, v2); }
// process-wide fractal setting (aero_data%fractal); spherical = (3, 1e-8, 1)
void pmco_set_fractal(double frac_dim, double prime_radius, double vol_fill_factor) {
  set_fractal(frac_dim, prime_radius, vol_fill_factor);
}
double pmco_aero_vol2rad(double v) { return aero_vol2rad(v); }
double pmco_aero_rad2vol(double r) { return aero_rad2vol(r); }
double pmco_kernel_brown_free(double v1, double d1, double v2, double d2, double T) {
  return kernel_brown_free_helper(v1, d1, v2, d2, T);
}
double pmco_kernel_brown_cont(double v1, double d1, double v2, double d2, double T) {
  return kernel_brown_cont_helper(v1, d1, v2, d2, T);
}
double pmco_kernel_brown(double v1, double d1, double v2, double d2, double T, double p) {
  return kernel_brown_helper(v1, d1, v2, d2, T, p);
}
int pmco_bin_grid_find(int n_bin, double gmin, double gmax, int n, const double* r, int* out) {
  BinGrid g;
  bin_grid_make_log(g, n_bin, gmin, gmax);
  for (int i = 0; i < n; i++) out[i] = bin_grid_find(g, r[i]);
  return 0;
}
int pmco_bin_grid_make(int n_bin, double gmin, double gmax, double* edges, double* centers) {
  BinGrid g;
  bin_grid_make_log(g, n_bin, gmin, gmax);
  for (int i = 0; i <= n_bin; i++) edges[i] = g.edges[i];
  for (int i = 0; i < n_bin; i++) centers[i] = g.centers[i];
  return 0;
}

// mc_coag (coagulation.F90:43) for one cell.  stats: [tot_n_samp, tot_n_coag, n_accel_pairs]
int pmco_mc_coag(void* h, int cell, int kernel_type, double del_t, unsigned step, int* stats) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  RunCtx rc{d->rng.get(), (uint32_t)d->global_cell[cell], step};
  CoagStats cs;
  mc_coag(kernel_type, d->env[cell], d->data, d->cells[cell], del_t, rc, cs);
  if (stats) {
    stats[0] = cs.tot_n_samp;
    stats[1] = cs.tot_n_coag;
    stats[2] = cs.n_accel_pairs;
  }
  ORACLE_CATCH(d)
}
// all cells, n_threads host threads over contiguous cell ranges (cells are
// independent, like one MPI rank per core in WRF).  Only valid with the Philox
// generator (addressed draws) when n_threads > 1.
int pmco_mc_coag_all(void* h, int kernel_type, double del_t, unsigned step, int n_threads, long long* tot) {
  OracleDomain* d = (OracleDomain*)h;
  const int nc = (int)d->cells.size();
  std::vector<long long> samp((size_t)n_threads, 0), coag((size_t)n_threads, 0);
  std::vector<int> err((size_t)n_threads, 0);
  auto work = [&](int t) {
    const int lo = (int)((long long)nc * t / n_threads), hi = (int)((long long)nc * (t + 1) / n_threads);
    for (int c = lo; c < hi; c++) {
      try {
        RunCtx rc{d->rng.get(), (uint32_t)d->global_cell[c], step};
        CoagStats cs;
        d->cells[c].valid_sort = false;
        mc_coag(kernel_type, d->env[c], d->data, d->cells[c], del_t, rc, cs);
        samp[t] += cs.tot_n_samp;
        coag[t] += cs.tot_n_coag;
      } catch (const PmcError& e) {
        err[t] = e.code ? e.code : -1;
      }
    }
  };
  if (n_threads <= 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  long long s = 0, c = 0;
  int e = 0;
  for (int t = 0; t < n_threads; t++) {
    s += samp[t];
    c += coag[t];
    if (err[t]) e = err[t];
  }
  if (tot) {
    tot[0] = s;
    tot[1] = c;
  }
  return e;
}

int pmco_rebalance(void* h, int cell, int allow_doubling, int allow_halving, unsigned step) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  RunCtx rc{d->rng.get(), (uint32_t)d->global_cell[cell], step};
  aero_state_rebalance(d->cells[cell], allow_doubling != 0, allow_halving != 0, rc);
  ORACLE_CATCH(d)
}
int pmco_scale_weight(void* h, int cell, int g, int c, double ratio, int allow_doubling, int allow_halving,
                      unsigned step) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  RunCtx rc{d->rng.get(), (uint32_t)d->global_cell[cell], step};
  aero_state_scale_weight(d->cells[cell], g, c, ratio, allow_doubling != 0, allow_halving != 0, rc);
  ORACLE_CATCH(d)
}

// ---- samplers and generators (KATs)
double pmco_philox_uniform(unsigned seed, unsigned cell, unsigned step, unsigned phase, unsigned a, unsigned b,
                           unsigned j) {
  RngKey k{cell, step, phase, a, b};
  return philox_uniform(seed, k, j);
}
void pmco_philox_raw(const unsigned* ctr, const unsigned* key, unsigned* out) { philox4x32_10(ctr, key, out); }
// sample_without_replacement (rand.F90:515-542) with the MT generator: out[sample_size] 1-based indices; returns the count
int pmco_sample_without_replacement(int sample_size, int population_size, unsigned seed, int* out) {
  RngMT mt(seed);
  Stream s(&mt, RngKey{0, 0, 0, 0, 0});
  const std::vector<int> v = sample_without_replacement(sample_size, population_size, s);
  for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
  return (int)v.size();
}
// the component rule of aero_particle_coagulate (aero_particle.F90:940-972) for two particles with n1 / n2 components
// (sources 1000 + i and 2000 + i) and n_primary_parts np1 / np2; out_src[<= 20]; returns the new component count
int pmco_coagulate_components(int n1, int np1, int n2, int np2, unsigned seed, int* out_src) {
  Particle p1, p2, pn;
  p1.vol.assign(1, 1.0);
  p2.vol.assign(1, 2.0);
  p1.n_primary_parts = np1;
  p2.n_primary_parts = np2;
  for (int i = 0; i < n1; i++) { p1.comp_src.push_back(1000 + i); p1.comp_time.push_back((double)i); }
  for (int i = 0; i < n2; i++) { p2.comp_src.push_back(2000 + i); p2.comp_time.push_back((double)i); }
  RngPhilox rng(seed);
  particle_coagulate(p1, p2, pn, &rng, RngKey{1, 2, PH_COAG, 3, 4}, 0u);
  for (size_t i = 0; i < pn.comp_src.size(); i++) out_src[i] = pn.comp_src[i];
  return (int)pn.comp_src.size();
}
unsigned pmco_mt_u32(unsigned seed, int skip) {
  RngMT mt(seed);
  for (int i = 0; i < skip; i++) mt.next_u32();
  return mt.next_u32();
}
// histogram of n samples of rand_poisson(mean) / rand_binomial(n,p) drawn from one MT stream
int pmco_sample_poisson(double mean, int n_samp, unsigned seed, int rng_kind, long long* hist, int hist_len) {
  std::unique_ptr<Rng> r;
  if (rng_kind == 0) r.reset(new RngPhilox(seed)); else r.reset(new RngMT(seed));
  for (int i = 0; i < n_samp; i++) {
    Stream s(r.get(), RngKey{0, 0, PH_INIT, (uint32_t)i, 0});
    const int k = rand_poisson(mean, s);
    if (k < hist_len) hist[k]++;
  }
  return 0;
}
int pmco_sample_binomial(int n, double p, int n_samp, unsigned seed, int rng_kind, long long* hist, int hist_len) {
  std::unique_ptr<Rng> r;
  if (rng_kind == 0) r.reset(new RngPhilox(seed)); else r.reset(new RngMT(seed));
  for (int i = 0; i < n_samp; i++) {
    Stream s(r.get(), RngKey{0, 0, PH_INIT, (uint32_t)i, 1});
    const int k = rand_binomial(n, p, s);
    if (k < hist_len) hist[k]++;
  }
  return 0;
}
double pmco_poisson_pmf(double mean, int k) { return poisson_pmf(mean, k); }
double pmco_binomial_pmf(int n, double p, int k) { return binomial_pmf(n, p, k); }

}  // extern "C"

// condensation + transport entry points live in pmc_oracle_ext_c.cpp
void* oracle_domain_data(void* h) { return &((OracleDomain*)h)->data; }
