// pmc_oracle_trans.cpp -- oracle part 3: stochastic particle transport between grid cells.
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
//
// Restates interface/wrf_pmc_trans.F90 (wrf_pmc_trans_driver :37-200: compute_advect_probs_wrf :205-286,
// compute_diffusion_probs :358-409, compute_vertical_probs :414-562, stable_timestep :626-652,
// normalize_probs :692-773) and interface/wrf_pmc_trans_aero.F90 (trans_aero :40-971:
// trans_aero_preweight :1103-1423, compute_fluxes :1428-1624, aero_state_multisample :2351-2482,
// aero_state_sample_particles_by_class :2181-2322, the ghost exchanges :150-273 / :1256-1362 / :381-473 /
// :543-675 and add + rebalance + sort :777-804) for the doubly periodic ("ideal") domain.
//
// The reference runs one MPI rank per 2-D patch and moves ghost-cell contents with MPI; the order in which
// a cell receives particles (and flux sums) therefore depends on the patch decomposition.  The oracle
// emulates a px x py decomposition of the global domain inside one process: every "rank" walks its own
// cells in the reference order (j, k, i) and the four ghost exchanges are applied in the reference order
// (y -1, y +1, x -1, x +1; get_array_index_for_send_recv_from_ghost :1826-1933), so that a device run on
// the same decomposition can be compared bit for bit.
#include "pmc_oracle_domain.hpp"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>

namespace pmco {

// direction slots of one cell: 0 -x, 1 +x, 2 -y, 3 +y, then 4 + (k-1) for the vertical level k (1-based)
static inline int dir_di(int d) { return d == 0 ? -1 : (d == 1 ? 1 : 0); }
static inline int dir_dj(int d) { return d == 2 ? -1 : (d == 3 ? 1 : 0); }

void TransDomain::setup(int gnx_, int nz_, int gny_, int px_, int py_, int n_class_) {
  gnx = gnx_; nz = nz_; gny = gny_; px = px_; py = py_; n_class = n_class_;
  const size_t nc = (size_t)gnx * nz * gny;
  adv.assign(nc * n_class * 6, 0.0);
  diff.assign(nc * n_class * 4, 0.0);
  vert.assign(nc * n_class * nz, 0.0);
}

// ---------------------------------------------------------------- probabilities
// compute_advect_probs_wrf (wrf_pmc_trans.F90:205-286).  adv slots: 0 -x, 1 +x, 2 -y, 3 +y, 4 -z, 5 +z
static void compute_advect_probs_wrf(OracleDomain& d, int i, int k, int j) {
  TransDomain& t = d.trans;
  const size_t g = t.cell(i, k, j);
  AeroState& st = d.cells[g];
  const double rho_mass = 1.0 / (double)t.alt[g];
  for (int c = 1; c <= t.n_class; c++) {
    double* a = &t.adv[(g * t.n_class + (c - 1)) * 6];
    for (int q = 0; q < 6; q++) a[q] = 0.0;
    const int n_parts = aero_state_total_particles(st, 1, c);
    const double mix_rat = (double)n_parts * st.awa.at(1, c).magnitude / rho_mass;
    const float uf1 = t.u_flux[t.uidx(i + 1, k, j, c)], uf0 = t.u_flux[t.uidx(i, k, j, c)];
    const float vf1 = t.v_flux[t.vidx(i, k, j + 1, c)], vf0 = t.v_flux[t.vidx(i, k, j, c)];
    const float wf1 = t.w_flux[t.widx(i, k + 1, j, c)], wf0 = t.w_flux[t.widx(i, k, j, c)];
    const double min_mix_rat = 0.0;
    if (uf1 > 0.0f && mix_rat > min_mix_rat) a[1] = (double)uf1 / mix_rat;
    if (uf0 < 0.0f && mix_rat > min_mix_rat) a[0] = -1.0 * (double)uf0 / mix_rat;
    if (vf1 > 0.0f && mix_rat > min_mix_rat) a[3] = (double)vf1 / mix_rat;
    if (vf0 < 0.0f && mix_rat > min_mix_rat) a[2] = -1.0 * (double)vf0 / mix_rat;
    if (wf1 > 0.0f && mix_rat > min_mix_rat) a[5] = std::min((double)wf1 / mix_rat, 0.95);
    if (wf0 < 0.0f && mix_rat > min_mix_rat) a[4] = std::min(-1.0 * (double)wf0 / mix_rat, 0.95);
    // sum(prob_advection(:,:,:,i_class)) in array element order (i fastest, then k, then j): -y -z -x +x +z +y
    const double prob_sum = ((((a[2] + a[4]) + a[0]) + a[1]) + a[5]) + a[3];
    if (prob_sum > 1.0)
      for (int q = 0; q < 6; q++) a[q] = a[q] / (prob_sum + 1.0e-8);
  }
}
// compute_diffusion_probs (:358-409).  diff slots: 0 -x, 1 +x, 2 -y, 3 +y; the same for every class
static void compute_diffusion_probs(OracleDomain& d, int i, int k, int j) {
  TransDomain& t = d.trans;
  const size_t g = t.cell(i, k, j);
  const double dx = (double)t.dx, dy = (double)t.dy, dt = (double)t.dt;
  const double rho_mass = 1.0 / (double)t.alt[g];
  double f[4];
  const int di[4] = {-1, 1, 0, 0}, dj[4] = {0, 0, -1, 1};
  for (int q = 0; q < 4; q++) {
    const size_t gn = t.cell_wrap(i + di[q], k, j + dj[q]);
    const float ksum = t.xkhh[g] + t.xkhh[gn];  // single precision sum, as in the Fortran
    const double diff_coef = std::max(0.5 * (double)ksum, 0.0);
    const double rho_edge = 0.5 * (rho_mass + (1.0 / (double)t.alt[gn]));
    const double dd = (q < 2) ? dx : dy;
    f[q] = (dt * diff_coef * rho_edge) / (dd * dd * rho_mass);
  }
  for (int c = 0; c < t.n_class; c++)
    for (int q = 0; q < 4; q++) t.diff[(g * t.n_class + c) * 4 + q] = f[q];
}
// stable_timestep (:626-652)
static void stable_timestep(const TransDomain& t, int i, int j, double& small_dt, int& num_steps) {
  const double dt = (double)t.dt;
  const double min_dz = (double)(t.z[t.cell(i, 2, j)] - t.z[t.cell(i, 1, j)]);
  float mx = t.exch_h[t.widx1(i, 1, j)];
  for (int k = 2; k <= t.nz + 1; k++) mx = std::max(mx, t.exch_h[t.widx1(i, k, j)]);
  const double max_k = std::max((double)mx, 1.0e-6);
  const double max_time = 0.1 * (min_dz * min_dz) / max_k;
  num_steps = (int)std::ceil(dt / max_time);
  small_dt = dt / (double)num_steps;
}
// compute_vertical_probs (:414-562) for the column (i, j); matrices are row major M[r*nz + c], 0-based
static void compute_vertical_probs(OracleDomain& d, int i, int j) {
  TransDomain& t = d.trans;
  const int nz = t.nz;
  double small_dt;
  int num_steps;
  stable_timestep(t, i, j, small_dt, num_steps);
  std::vector<double> z_mass(nz), rho_mass(nz), z_w(nz + 1), k_w(nz + 1, 0.0), rho_w(nz + 1, 0.0);
  for (int k = 1; k <= nz; k++) {
    z_mass[k - 1] = (double)t.z[t.cell(i, k, j)];
    rho_mass[k - 1] = 1.0 / (double)t.alt[t.cell(i, k, j)];
  }
  for (int k = 1; k <= nz + 1; k++) z_w[k - 1] = (double)t.z_at_w[t.widx1(i, k, j)];
  for (int k = 1; k <= nz - 1; k++) k_w[k - 1] = std::max(1.0e-6, (double)t.exch_h[t.widx1(i, k + 1, j)]);
  for (int k = 1; k <= nz - 1; k++) {
    const double tmp = z_mass[k] - z_mass[k - 1];
    rho_w[k - 1] = ((z_mass[k] - z_w[k]) / tmp) * rho_mass[k - 1] + ((z_w[k] - z_mass[k - 1]) / tmp) * rho_mass[k];
  }
  std::vector<double> B((size_t)nz * nz, 0.0), Bt, R((size_t)nz * nz), A((size_t)nz * nz);
  auto Bm = [&](int r, int c) -> double& { return B[(size_t)(r - 1) * nz + (c - 1)]; };
  for (int k = 2; k <= nz - 1; k++) {
    Bm(k, k - 1) = (small_dt / ((z_w[k] - z_w[k - 1]) * (z_mass[k - 1] - z_mass[k - 2]))) *
                   (rho_w[k - 2] / rho_mass[k - 1]) * k_w[k - 2];
    Bm(k, k + 1) = (small_dt / ((z_w[k] - z_w[k - 1]) * (z_mass[k] - z_mass[k - 1]))) *
                   (rho_w[k - 1] / rho_mass[k - 1]) * k_w[k - 1];
    Bm(k, k) = 1.0 - Bm(k, k - 1) - Bm(k, k + 1);
  }
  Bm(1, 2) = (small_dt / ((z_w[1] - z_w[0]) * (z_mass[1] - z_mass[0]))) * (rho_w[0] / rho_mass[0]) * k_w[0];
  Bm(1, 1) = 1.0 - Bm(1, 2);
  Bm(nz, nz - 1) = (small_dt / ((z_w[nz] - z_w[nz - 1]) * (z_mass[nz - 1] - z_mass[nz - 2]))) *
                   (rho_w[nz - 2] / rho_mass[nz - 1]) * k_w[nz - 2];
  Bm(nz, nz) = 1.0 - Bm(nz, nz - 1);
  // B* = B^num_steps with row renormalisation (:496-505); matmul summed in ascending inner index
  Bt = B;
  for (int s = 2; s <= num_steps; s++) {
    for (int r = 0; r < nz; r++)
      for (int c = 0; c < nz; c++) {
        double acc = 0.0;
        for (int q = 0; q < nz; q++) acc += Bt[(size_t)r * nz + q] * B[(size_t)q * nz + c];
        R[(size_t)r * nz + c] = acc;
      }
    Bt = R;
    for (int r = 0; r < nz; r++) {
      double rs = 0.0;
      for (int c = 0; c < nz; c++) rs += Bt[(size_t)r * nz + c];
      for (int c = 0; c < nz; c++) Bt[(size_t)r * nz + c] = Bt[(size_t)r * nz + c] / rs;
    }
  }
  for (int cl = 1; cl <= t.n_class; cl++) {
    std::fill(A.begin(), A.end(), 0.0);
    auto Am = [&](int r, int c) -> double& { return A[(size_t)(r - 1) * nz + (c - 1)]; };
    auto up = [&](int k) { return t.adv[(t.cell(i, k, j) * t.n_class + (cl - 1)) * 6 + 5]; };
    auto dn = [&](int k) { return t.adv[(t.cell(i, k, j) * t.n_class + (cl - 1)) * 6 + 4]; };
    for (int k = 2; k <= nz - 1; k++) {
      Am(k, k) = 1.0 - (up(k) + dn(k));
      Am(k, k - 1) = dn(k);
      Am(k, k + 1) = up(k);
    }
    Am(1, 1) = 1.0 - up(1);
    Am(1, 2) = up(1);
    Am(nz, nz) = 1.0 - dn(nz);
    Am(nz, nz - 1) = dn(nz);
    for (int r = 0; r < nz; r++)
      for (int c = 0; c < nz; c++) {
        double v;
        if (t.vertmix > 0) {
          double acc = 0.0;
          for (int q = 0; q < nz; q++) acc += Bt[(size_t)r * nz + q] * A[(size_t)q * nz + c];
          v = acc;
        } else {
          v = A[(size_t)r * nz + c];
        }
        t.vert[(t.cell(i, r + 1, j) * t.n_class + (cl - 1)) * nz + c] = v;
      }
  }
}
// normalize_probs (:692-773)
static void normalize_probs(OracleDomain& d, int i, int k, int j, int cl) {
  TransDomain& t = d.trans;
  const size_t g = t.cell(i, k, j);
  double* a = &t.adv[(g * t.n_class + (cl - 1)) * 6];
  double* f = &t.diff[(g * t.n_class + (cl - 1)) * 4];
  double* v = &t.vert[(g * t.n_class + (cl - 1)) * t.nz];
  auto valid_lo = [&]() {
    for (int q = 0; q < 6; q++) if (!(a[q] >= 0.0)) return false;
    for (int q = 0; q < 4; q++) if (!(f[q] >= 0.0)) return false;
    for (int q = 0; q < t.nz; q++) if (!(v[q] >= 0.0)) return false;
    return true;
  };
  auto total = [&]() {
    // sum(prob_advection(:,0,:)) and sum(prob_diffusion(:,:,:)) in array element order: -y, -x, +x, +y
    double s = ((a[2] + a[0]) + a[1]) + a[3];
    s = s + (((f[2] + f[0]) + f[1]) + f[3]);
    for (int kk = 1; kk <= t.nz; kk++)
      if (kk != k) s = s + v[kk - 1];
    return s;
  };
  pmc_assert(414897882, valid_lo());
  const double prob_sum = total();
  if (prob_sum >= 1.0) {
    for (int q = 0; q < t.nz; q++) v[q] = v[q] / prob_sum;
    for (int q = 0; q < 4; q++) f[q] = f[q] / prob_sum;
    for (int q = 0; q < 4; q++) a[q] = a[q] / prob_sum;  // prob_advection(:,0,:): the horizontal entries
  }
  pmc_assert(184874367, valid_lo());
  bool hi = true;
  for (int q = 0; q < 6; q++) hi = hi && (a[q] <= 1.0);
  for (int q = 0; q < 4; q++) hi = hi && (f[q] <= 1.0);
  for (int q = 0; q < t.nz; q++) hi = hi && (v[q] <= 1.0);
  pmc_assert(184874369, hi);
  pmc_assert(414897881, total() < 1.0 + 1.0e-8);
}

void trans_compute_probs(OracleDomain& d) {
  TransDomain& t = d.trans;
  for (int j = 1; j <= t.gny; j++)
    for (int k = 1; k <= t.nz; k++)
      for (int i = 1; i <= t.gnx; i++) {
        // aero_state_total_particles needs no sorting; keep valid_sort untouched
        compute_advect_probs_wrf(d, i, k, j);
        compute_diffusion_probs(d, i, k, j);
      }
  std::fill(t.vert.begin(), t.vert.end(), 0.0);  // prob_vert_diffusion = 0 (wrf_pmc_trans.F90:104)
  for (int j = 1; j <= t.gny; j++)
    for (int i = 1; i <= t.gnx; i++)
      if (t.kind(i, j) == 0) compute_vertical_probs(d, i, j);  // interior columns only in open domains (:116-136)
  for (int j = 1; j <= t.gny; j++)
    for (int k = 1; k <= t.nz; k++)
      for (int i = 1; i <= t.gnx; i++)
        for (int c = 1; c <= t.n_class; c++) normalize_probs(d, i, k, j, c);
}

// ---------------------------------------------------------------- trans_aero
namespace {
struct Patch {
  int is, ie, js, je;  // owned global index ranges (1-based, inclusive)
};
Patch patch_of(const TransDomain& t, int r) {
  const int rx = r % t.px, ry = r / t.px;
  Patch p;
  p.is = rx * (t.gnx / t.px) + 1;
  p.ie = (rx == t.px - 1) ? t.gnx : (rx + 1) * (t.gnx / t.px);
  p.js = ry * (t.gny / t.py) + 1;
  p.je = (ry == t.py - 1) ? t.gny : (ry + 1) * (t.gny / t.py);
  return p;
}
// move probability of one horizontal direction slot
inline double hprob(const TransDomain& t, size_t g, int cl, int dslot) {
  return t.adv[(g * t.n_class + (cl - 1)) * 6 + dslot] + t.diff[(g * t.n_class + (cl - 1)) * 4 + dslot];
}
}  // namespace

// trans_aero_preweight (wrf_pmc_trans_aero.F90:1103-1423) for the patch of one rank.  ghost[4] receive the
// flux sums of the ghost ring (y-, y+, x-, x+ ghost rows / columns), added into the neighbour afterwards.
struct PreweightWork {
  std::vector<double> num_conc_new;             // [cell global][class], only owned cells used
  std::vector<double> ghost;                    // [4][edge cell][nz][class]
};

// aero_state_sample_particles_by_class (wrf_pmc_trans_aero.F90:2181-2322): particles of class c of cell g move
// to (or are copied into) `out`, the delta state of cell gd, with probability adjusted_prob each; draw address
// a = (c - 1) 64 + s.  Returns the number of records added to `out`.
static long long sample_particles_by_class(OracleDomain& d, size_t g, size_t gd, int c, double adjusted_prob, int s,
                                           std::vector<Particle>* out, const RunCtx& rc) {
  AeroState& from = d.cells[g];
  const double source_vol = d.env[g].cell_volume, dest_vol = d.env[gd].cell_volume;
  pmc_assert(721006963, adjusted_prob >= 0.0 && adjusted_prob <= 1.0);
  const double factor = source_vol / dest_vol;
  const double mag_to = d.cells[gd].awa.at(1, c).magnitude * (1.0 / factor);
  if (!from.valid_sort) aero_state_sort(from);
  const double ratio = (mag_to / 1.0) / (from.awa.at(1, c).magnitude / 1.0);
  const int n_part_from = aero_state_total_particles(from, 1, c);
  const uint32_t a = (uint32_t)(c - 1) * 64u + (uint32_t)s;
  Stream s0(rc.rng, rc.key(PH_TRANS, a, 0));
  int n_transfer;
  if (ratio >= 1.0) n_transfer = rand_binomial(n_part_from, adjusted_prob, s0);
  else n_transfer = rand_poisson((double)n_part_from * (adjusted_prob / ratio), s0);
  long long moved = 0;
  int i_transfer = 0;
  while (i_transfer < n_transfer) {
    IntVarray& lst = from.sorted.group_class.inv(1, c);
    if (lst.n_entry() <= 0) break;
    i_transfer++;
    const RngKey key = rc.key(PH_TRANS, a, (uint32_t)i_transfer);
    const int i_entry = pmc_rand_int(rc.rng->uniform(key, 0), lst.n_entry());
    const int i_part = lst.at(i_entry);
    const Particle& pt = from.apa[i_part - 1];
    // flat weights: num_conc = magnitude (aero_weight_array_num_conc, aero_weight_array.F90:257)
    const double mix_rat_from = awa_num_conc(from.awa, pt);
    Weight wto = d.cells[gd].awa.at(pt.weight_group, c);
    wto.magnitude = wto.magnitude * (1.0 / factor);  // aero_weight_array_scale(awa, 1 / factor)
    const double mix_rat_to = weight_num_conc_at_radius(wto, particle_radius(pt));
    bool do_add, do_remove;
    if (mix_rat_to == mix_rat_from) {
      do_add = true; do_remove = true;
    } else if (mix_rat_to < mix_rat_from) {
      do_add = true;
      do_remove = rc.rng->uniform(key, 1) < mix_rat_to / mix_rat_from;
    } else {
      do_add = rc.rng->uniform(key, 1) < mix_rat_from / mix_rat_to;
      do_remove = true;
    }
    if (do_add) {
      out->push_back(pt);
      if (!do_remove) out->back().id = from.idgen.new_id();
      moved++;
    }
    if (do_remove) aero_state_remove_particle(from, i_part);
  }
  return moved;
}

// aero_state_dup_all_particles (partmc/src/aero_state.F90:3529-3551): the whole state of `from` (particles,
// weights, sorting, n_part_ideal) becomes the state of `to`, every particle with a new id; the id counter of
// `to` is its own (one counter per cell: DESIGN.md section 2).
static void aero_state_dup_all_particles(const AeroState& from, AeroState& to) {
  const IdGen keep = to.idgen;
  if (from.n_part() > 0) {
    to = from;
    to.idgen = keep;
    for (Particle& p : to.apa) p.id = to.idgen.new_id();
  } else {
    to.apa.clear();  // aero_state_zero
    to.valid_sort = false;
    to.n_part_ideal = from.n_part_ideal;
  }
}

// aero_state_resample (wrf_pmc_trans_aero.F90:976-1098): bring a boundary cell back to the background
// distribution.  density = 1 / inverse_density, sample_prop = 1, create_time = 0, allow_doubling = allow_halving
// = .true. at all call sites (:839-842).  Starts from a fresh sort (the device keeps no lists between phases).
static int aero_state_resample(AeroState& st, const AeroData& data, const std::vector<AeroMode>& modes, double density,
                               double create_time, bool throw_away_previous, const RunCtx& rc) {
  const double sample_timescale = 1.0, factor = 0.1;
  int n_part_add = 0;
  if (throw_away_previous) {
    st.apa.clear();  // aero_state_zero keeps awa and n_part_ideal
    st.valid_sort = false;
    return aero_state_add_aero_dist_sample(st, data, modes, density, sample_timescale, create_time, true, true, rc);
  }
  st.valid_sort = false;
  aero_state_sort(st);
  const int n_group = st.awa.n_group, n_class = st.awa.n_class, n_mode = (int)modes.size();
  std::vector<double> vols;
  for (int g = 1; g <= n_group; g++)
    for (int c = 1; c <= n_class; c++) {
      const uint32_t gc = (uint32_t)((c - 1) * n_group + (g - 1));
      const int n_part_before = aero_state_total_particles(st, g, c);
      double total_num_conc = 0.0;
      for (const AeroMode& m : modes)
        if (m.weight_class == c) total_num_conc = total_num_conc + m.num_conc * density;  // aero_mode_total_num_conc
      const double n_parts_new = total_num_conc / st.awa.at(g, c).magnitude;
      Stream s0(rc.rng, rc.key(PH_BND, gc, 0u));
      const int n_samp = rand_poisson(n_parts_new, s0);
      const int n_parts_to_add = n_samp - n_part_before;
      if (n_parts_to_add > 0) {
        for (int im = 1; im <= n_mode; im++) {
          const AeroMode& m = modes[im - 1];
          if (m.weight_class != c) continue;
          const uint32_t am = 1000u + gc * (uint32_t)n_mode + (uint32_t)(im - 1);
          const double sample_proportion = (density * m.num_conc) / total_num_conc;
          Stream sm(rc.rng, rc.key(PH_BND, am, 0u));
          const int n_samp_mode = rand_poisson(sample_proportion * (double)n_parts_to_add, sm);
          const double n_samp_avg = sample_proportion * (double)n_parts_to_add;
          const double num_conc_threshold = (density * m.num_conc / n_samp_avg) * factor;
          const double size_factor = std::min(1.0 / (sample_timescale * n_samp_avg), 0.32);
          for (int i_samp = 1; i_samp <= n_samp_mode; i_samp++) {
            Stream s(rc.rng, rc.key(PH_BND, am, (uint32_t)i_samp));
            const double radius = aero_mode_sample_radius(m, st.awa.at(g, c), size_factor, s);
            const double total_vol = aero_rad2vol(radius);
            aero_mode_sample_vols(m, total_vol, vols, s);
            Particle p;
            p.vol = vols;
            p.id = st.idgen.new_id();  // (an id is consumed whether or not the particle is kept, :1067)
            p.weight_group = g;
            p.weight_class = c;
            p.least_create_time = create_time;
            p.greatest_create_time = create_time;
            p.n_primary_parts = 1;
            p.water_hyst_leg = 0;
            p.comp_src.assign(1, c);
            p.comp_time.assign(1, create_time);
            if (awa_num_conc(st.awa, p) > num_conc_threshold) {
              aero_state_add_particle(st, p, true);
              n_part_add++;
            }
          }
        }
      } else {
        const int n_part = st.sorted.group_class.inv(g, c).n_entry();
        const int n_rm = std::min(-n_parts_to_add, n_part);
        for (int i_samp = 1; i_samp <= n_rm; i_samp++) {
          IntVarray& lst = st.sorted.group_class.inv(g, c);
          const int i_entry = pmc_rand_int(rc.rng->uniform(rc.key(PH_BND, gc, (uint32_t)i_samp), 0), lst.n_entry());
          aero_state_remove_particle(st, lst.at(i_entry));
        }
      }
    }
  return n_part_add;
}

void trans_aero_run(OracleDomain& d, unsigned step, bool allow_doubling, bool allow_halving, int n_threads,
                    long long* n_moved_out) {
  TransDomain& t = d.trans;
  const int nz = t.nz, C = t.n_class, gnx = t.gnx, gny = t.gny;
  const int n_rank = t.px * t.py;
  const size_t n_cells = (size_t)gnx * nz * gny;
  auto par_ranks = [&](auto fn) {
    if (n_threads <= 1 || n_rank == 1) {
      for (int r = 0; r < n_rank; r++) fn(r);
    } else {
      std::vector<std::thread> th;
      const int nt = std::min(n_threads, n_rank);
      for (int w = 0; w < nt; w++)
        th.emplace_back([&, w]() {
          for (int r = w; r < n_rank; r += nt) fn(r);
        });
      for (auto& x : th) x.join();
    }
  };
  std::vector<int> err((size_t)n_rank, 0);

  // ---- preweight: sort (from scratch: the device keeps no sorting between phases), old number
  // concentrations, fluxes (:1200-1245).  num_new holds owned cells; gh[r][side] the ghost ring sums.
  std::vector<double> num_old(n_cells * C, 0.0), num_new(n_cells * C, 0.0);
  par_ranks([&](int r) {
    const Patch p = patch_of(t, r);
    for (int j = p.js; j <= p.je; j++)
      for (int k = 1; k <= nz; k++)
        for (int i = p.is; i <= p.ie; i++) {
          AeroState& st = d.cells[t.cell(i, k, j)];
          st.valid_sort = false;
          aero_state_sort(st);
          for (int c = 1; c <= C; c++) {
            const int n = aero_state_total_particles(st, 1, c);
            const double v = (double)n * st.awa.at(1, c).magnitude;
            num_old[t.cell(i, k, j) * C + (c - 1)] = v;
            num_new[t.cell(i, k, j) * C + (c - 1)] = v;
          }
        }
  });
  // ghost sums: side 0 = row js-1 (y-), 1 = row je+1 (y+), 2 = column is-1 (x-), 3 = column ie+1 (x+)
  std::vector<std::vector<double>> gh((size_t)n_rank * 4);
  par_ranks([&](int r) {
    const Patch p = patch_of(t, r);
    const int lx = p.ie - p.is + 1, ly = p.je - p.js + 1;
    gh[(size_t)r * 4 + 0].assign((size_t)lx * nz * C, 0.0);
    gh[(size_t)r * 4 + 1].assign((size_t)lx * nz * C, 0.0);
    gh[(size_t)r * 4 + 2].assign((size_t)ly * nz * C, 0.0);
    gh[(size_t)r * 4 + 3].assign((size_t)ly * nz * C, 0.0);
    for (int j = p.js; j <= p.je; j++)
      for (int k = 1; k <= nz; k++)
        for (int i = p.is; i <= p.ie; i++)
          for (int c = 1; c <= C; c++) {
            // compute_fluxes (:1428-1624): ideal branch, or the "real case" branch of an open domain, where a
            // boundary-edge cell only feeds its interior neighbour and a corner cell feeds nobody (:1517-1572)
            const size_t g = t.cell(i, k, j);
            const double nco = num_old[g * C + (c - 1)];
            const double source_vol = d.env[g].cell_volume;
            double& self = num_new[g * C + (c - 1)];
            const int kd = t.kind(i, j);
            auto horizontal = [&](int di, int dj) {
              const double dest_vol = d.env[t.cell_wrap(i + di, k, j + dj)].cell_volume;
              const double prob = hprob(t, g, c, di < 0 ? 0 : (di > 0 ? 1 : (dj < 0 ? 2 : 3)));
              const double flux = prob * nco;
              double* dst;
              if (j + dj < p.js) dst = &gh[(size_t)r * 4 + 0][((size_t)(i - p.is) * nz + (k - 1)) * C + (c - 1)];
              else if (j + dj > p.je) dst = &gh[(size_t)r * 4 + 1][((size_t)(i - p.is) * nz + (k - 1)) * C + (c - 1)];
              else if (i + di < p.is) dst = &gh[(size_t)r * 4 + 2][((size_t)(j - p.js) * nz + (k - 1)) * C + (c - 1)];
              else if (i + di > p.ie) dst = &gh[(size_t)r * 4 + 3][((size_t)(j - p.js) * nz + (k - 1)) * C + (c - 1)];
              else dst = &num_new[t.cell(i + di, k, j + dj) * C + (c - 1)];
              *dst = *dst + flux * (dest_vol / source_vol);
              self = self - flux;
            };
            if (kd == 1) { horizontal(1, 0); continue; }
            if (kd == 2) { horizontal(-1, 0); continue; }
            if (kd == 3) { horizontal(0, 1); continue; }
            if (kd == 4) { horizontal(0, -1); continue; }
            if (kd == 5) continue;
            horizontal(0, -1);
            horizontal(0, 1);
            horizontal(-1, 0);
            horizontal(1, 0);
            for (int kk = 1; kk <= nz; kk++) {
              if (kk == k) continue;
              const double prob = t.vert[(g * C + (c - 1)) * nz + (kk - 1)];
              const double dest_vol = d.env[t.cell(i, kk, j)].cell_volume;
              const double flux = prob * nco;
              double& dst = num_new[t.cell(i, kk, j) * C + (c - 1)];
              dst = dst + flux * (dest_vol / source_vol);
              self = self - flux;
            }
          }
  });
  // ghost sums are added into the neighbour's edge cells in the order y-, y+, x-, x+ (:1256-1362)
  for (int side = 0; side < 4; side++)
    for (int r = 0; r < n_rank; r++) {
      const Patch p = patch_of(t, r);
      const int lx = p.ie - p.is + 1, ly = p.je - p.js + 1;
      const std::vector<double>& G = gh[(size_t)r * 4 + side];
      if (side < 2) {
        const int jd = (side == 0) ? p.js - 1 : p.je + 1;
        if (t.open && (jd < 1 || jd > gny)) continue;
        for (int k = 1; k <= nz; k++)
          for (int i = p.is; i <= p.ie; i++)
            for (int c = 0; c < C; c++) {
              double& dst = num_new[t.cell_wrap(i, k, jd) * C + c];
              dst = dst + G[((size_t)(i - p.is) * nz + (k - 1)) * C + c];
            }
        (void)lx;
      } else {
        const int id = (side == 2) ? p.is - 1 : p.ie + 1;
        if (t.open && (id < 1 || id > gnx)) continue;
        for (int j = p.js; j <= p.je; j++)
          for (int k = 1; k <= nz; k++)
            for (int c = 0; c < C; c++) {
              double& dst = num_new[t.cell_wrap(id, k, j) * C + c];
              dst = dst + G[((size_t)(j - p.js) * nz + (k - 1)) * C + c];
            }
        (void)ly;
      }
    }
  // adjust the weights where the population would more than double (:1373-1400)
  par_ranks([&](int r) {
    const Patch p = patch_of(t, r);
    try {
      for (int j = p.js; j <= p.je; j++)
        for (int k = 1; k <= nz; k++)
          for (int i = p.is; i <= p.ie; i++) {
            const size_t g = t.cell(i, k, j);
            AeroState& st = d.cells[g];
            RunCtx rc{d.rng.get(), (uint32_t)d.global_cell[g], step};
            for (int c = 1; c <= C; c++) {
              const double n_parts_new = num_new[g * C + (c - 1)] / st.awa.at(1, c).magnitude;
              if (n_parts_new > 2.0 * st.ideal(1, c)) {
                const double weight_ratio = n_parts_new / st.ideal(1, c);
                aero_state_scale_weight(st, 1, c, weight_ratio, true, true, rc);
              }
            }
          }
    } catch (const PmcError& e) {
      err[r] = e.code ? e.code : -1;
    }
  });
  for (int r = 0; r < n_rank; r++)
    if (err[r]) throw PmcError(err[r], "trans_aero_preweight failed");

  // ---- sampling (:519-529): delta states per destination, local part and ghost part kept apart
  std::vector<std::vector<Particle>> delta(n_cells);            // local arrivals (same rank)
  std::vector<std::vector<Particle>> halo(n_cells * 4);         // arrivals through ghost side s of the SENDER
  std::vector<long long> moved((size_t)n_rank, 0);
  par_ranks([&](int r) {
    const Patch p = patch_of(t, r);
    std::vector<double> probs((size_t)nz + 4);
    std::vector<int> slot((size_t)nz + 4), perm((size_t)nz + 4);
    try {
      for (int j = p.js; j <= p.je; j++)
        for (int k = 1; k <= nz; k++)
          for (int i = p.is; i <= p.ie; i++) {
            if (t.kind(i, j) != 0) continue;  // open domain: interior cells only (:497-506)
            // aero_state_multisample (:2351-2482)
            const size_t g = t.cell(i, k, j);
            AeroState& from = d.cells[g];
            RunCtx rc{d.rng.get(), (uint32_t)d.global_cell[g], step};
            for (int c = 1; c <= C; c++) {
              double remainder_prob = 1.0;
              int count = 0;
              for (int q = 0; q < 4; q++) {
                const double pr = hprob(t, g, c, q);
                if (pr > 0.0) { probs[count] = pr; slot[count] = q; count++; }
              }
              for (int kk = 1; kk <= nz; kk++) {
                if (kk == k) continue;
                const double pr = t.vert[(g * C + (c - 1)) * nz + (kk - 1)];
                if (pr > 0.0) { probs[count] = pr; slot[count] = 4 + (kk - 1); count++; }
              }
              // real_sort (util.F90:1612 -> sort.c:74, qsort): ascending; ties keep the collection order
              for (int q = 0; q < count; q++) perm[q] = q;
              std::stable_sort(perm.begin(), perm.begin() + count,
                               [&](int a, int b) { return probs[a] < probs[b]; });
              for (int s = 0; s < count; s++) {
                const double prob = probs[perm[s]];
                const double adjusted_prob = std::min(prob / remainder_prob, 1.0);
                if (!(remainder_prob >= 1.0e-8)) continue;
                const int q = slot[perm[s]];
                int di = 0, dj = 0, dk = k;
                if (q < 4) { di = dir_di(q); dj = dir_dj(q); } else { dk = q - 4 + 1; }
                const size_t gd = t.cell_wrap(i + di, dk, j + dj);
                std::vector<Particle>* out;
                if (j + dj < p.js) out = &halo[gd * 4 + 0];
                else if (j + dj > p.je) out = &halo[gd * 4 + 1];
                else if (i + di < p.is) out = &halo[gd * 4 + 2];
                else if (i + di > p.ie) out = &halo[gd * 4 + 3];
                else out = &delta[gd];
                moved[r] += sample_particles_by_class(d, g, gd, c, adjusted_prob, s, out, rc);
                remainder_prob = remainder_prob - prob;
              }
            }
          }
    } catch (const PmcError& e) {
      err[r] = e.code ? e.code : -1;
    }
  });
  for (int r = 0; r < n_rank; r++)
    if (err[r]) throw PmcError(err[r], "aero_state_multisample failed");
  // ---- ghost particle exchange (:543-675): y-, y+, x-, x+, appended behind the local arrivals
  for (int side = 0; side < 4; side++)
    for (size_t g = 0; g < n_cells; g++) {
      std::vector<Particle>& h = halo[g * 4 + side];
      delta[g].insert(delta[g].end(), h.begin(), h.end());
      h.clear();
    }
  // ---- open domain: inflow from the boundary cells into their interior neighbours (:689-766), appended behind
  // the ghost arrivals, in the order west, east, south, north
  if (t.open) {
    for (int r = 0; r < n_rank; r++) {
      const Patch p = patch_of(t, r);
      auto inflow = [&](int i, int j, int di, int dj, int dslot) {
        for (int k = 1; k <= nz; k++) {
          const size_t g = t.cell(i, k, j), gd = t.cell(i + di, k, j + dj);
          RunCtx rc{d.rng.get(), (uint32_t)d.global_cell[g], step};
          for (int c = 1; c <= C; c++) {
            const double prob = std::min(hprob(t, g, c, dslot), 1.0);
            moved[r] += sample_particles_by_class(d, g, gd, c, prob, 0, &delta[gd], rc);
          }
        }
      };
      if (p.is == 1)
        for (int j = std::max(p.js, 2); j <= std::min(p.je, gny - 1); j++) inflow(1, j, 1, 0, 1);
      if (p.ie == gnx)
        for (int j = std::max(p.js, 2); j <= std::min(p.je, gny - 1); j++) inflow(gnx, j, -1, 0, 0);
      if (p.js == 1)
        for (int i = std::max(p.is, 2); i <= std::min(p.ie, gnx - 1); i++) inflow(i, 1, 0, 1, 3);
      if (p.je == gny)
        for (int i = std::max(p.is, 2); i <= std::min(p.ie, gnx - 1); i++) inflow(i, gny, 0, -1, 2);
    }
  }
  // ---- add, rebalance, sort (:777-804): interior cells (all cells of a periodic domain)
  par_ranks([&](int r) {
    const Patch p = patch_of(t, r);
    try {
      for (int j = p.js; j <= p.je; j++)
        for (int k = 1; k <= nz; k++)
          for (int i = p.is; i <= p.ie; i++) {
            if (t.kind(i, j) != 0) continue;
            const size_t g = t.cell(i, k, j);
            AeroState& st = d.cells[g];
            RunCtx rc{d.rng.get(), (uint32_t)d.global_cell[g], step};
            // the device appends to an unsorted particle array and sorts afterwards
            st.valid_sort = false;
            for (const Particle& pt : delta[g]) aero_state_add_particle(st, pt);
            aero_state_rebalance(st, allow_doubling, allow_halving, rc);
            st.valid_sort = false;
            aero_state_sort(st);
          }
    } catch (const PmcError& e) {
      err[r] = e.code ? e.code : -1;
    }
  });
  for (int r = 0; r < n_rank; r++)
    if (err[r]) throw PmcError(err[r], "trans_aero add/rebalance/sort failed");
  // ---- open domain: refresh of the boundary cells (:811-955): a copy of the interior neighbour where the wind
  // blows out of the domain, a resample of the background distribution where it blows in
  if (t.open) {
    if (t.bnd_outflow.size() != n_cells || t.bnd_throw.size() != n_cells ||
        t.bnd_num_conc.size() != n_cells * t.bnd_modes.size())
      throw PmcError(-2, "open domain: boundary conditions not set (pmco_trans_set_boundary)");
    try {
      for (int j = 1; j <= gny; j++)
        for (int k = 1; k <= nz; k++)
          for (int i = 1; i <= gnx; i++) {
            const int kd = t.kind(i, j);
            if (kd < 1 || kd > 4) continue;
            const size_t g = t.cell(i, k, j);
            const size_t gn = t.cell(i + (kd == 1 ? 1 : (kd == 2 ? -1 : 0)), k, j + (kd == 3 ? 1 : (kd == 4 ? -1 : 0)));
            RunCtx rc{d.rng.get(), (uint32_t)d.global_cell[g], step};
            if (t.bnd_outflow[g]) {
              aero_state_dup_all_particles(d.cells[gn], d.cells[g]);
            } else {
              std::vector<AeroMode> modes = t.bnd_modes;
              for (size_t m = 0; m < modes.size(); m++) modes[m].num_conc = t.bnd_num_conc[g * modes.size() + m];
              aero_state_resample(d.cells[g], d.data, modes, 1.0 / d.env[g].inverse_density, 0.0, t.bnd_throw[g] != 0, rc);
            }
            // the device decides the bin grid of every refreshed boundary cell from its final population
            d.cells[g].valid_sort = false;
            aero_state_sort(d.cells[g]);
          }
    } catch (const PmcError& e) {
      throw PmcError(e.code ? e.code : -1, "boundary refresh failed");
    }
  }
  if (n_moved_out) {
    long long m = 0;
    for (int r = 0; r < n_rank; r++) m += moved[r];
    *n_moved_out = m;
  }
}

}  // namespace pmco

// ---------------------------------------------------------------- C entry points
extern "C" {

int pmco_trans_setup(void* h, int gnx, int nz, int gny, int px, int py, float dx, float dy, int vertmix) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  if ((size_t)gnx * nz * gny != d->cells.size()) throw PmcError(-2, "domain size does not match the cells");
  if (gnx % px != 0 || gny % py != 0) throw PmcError(-2, "patch grid must divide the domain");
  d->trans.setup(gnx, nz, gny, px, py, d->cells[0].awa.n_class);
  d->trans.dx = dx;
  d->trans.dy = dy;
  d->trans.vertmix = vertmix;
  ORACLE_CATCH(d)
}
// global fields in WRF memory order (i fastest, then k, then j, class slowest):
// u_flux [(gnx+1), nz, gny, C], v_flux [gnx, nz, (gny+1), C], w_flux [gnx, (nz+1), gny, C],
// xkhh / alt / z [gnx, nz, gny], z_at_w / exch_h [gnx, (nz+1), gny]
int pmco_trans_set_met(void* h, const float* u_flux, const float* v_flux, const float* w_flux, const float* xkhh,
                       const float* alt, const float* z, const float* z_at_w, const float* exch_h) {
  OracleDomain* d = (OracleDomain*)h;
  TransDomain& t = d->trans;
  const size_t C = t.n_class, nx = t.gnx, ny = t.gny, nz = t.nz;
  t.u_flux.assign(u_flux, u_flux + (nx + 1) * nz * ny * C);
  t.v_flux.assign(v_flux, v_flux + nx * nz * (ny + 1) * C);
  t.w_flux.assign(w_flux, w_flux + nx * (nz + 1) * ny * C);
  t.xkhh.assign(xkhh, xkhh + nx * nz * ny);
  t.alt.assign(alt, alt + nx * nz * ny);
  t.z.assign(z, z + nx * nz * ny);
  t.z_at_w.assign(z_at_w, z_at_w + nx * (nz + 1) * ny);
  t.exch_h.assign(exch_h, exch_h + nx * (nz + 1) * ny);
  return 0;
}
// open domain: switch + boundary conditions.  Mode tables like pmco_add_aero_dist_sample; num_conc [n_cells][n_mode],
// outflow / throw_away [n_cells] (read for the boundary-edge cells only), all in global cell order.
int pmco_trans_set_boundary(void* h, int open, int n_mode, const int* type, const int* weight_class,
                            const double* char_radius, const double* log10_std, const double* vol_frac,
                            const double* vol_frac_std, const double* num_conc, const int* outflow,
                            const int* throw_away) {
  OracleDomain* d = (OracleDomain*)h;
  TransDomain& t = d->trans;
  t.open = open != 0;
  if (!t.open) return 0;
  const size_t S = (size_t)d->data.n_spec, nc = d->cells.size();
  t.bnd_modes.assign((size_t)n_mode, AeroMode());
  for (int m = 0; m < n_mode; m++) {
    AeroMode& am = t.bnd_modes[m];
    am.type = type[m];
    am.weight_class = weight_class[m];
    am.char_radius = char_radius[m];
    am.log10_std_dev_radius = log10_std[m];
    am.vol_frac.assign(vol_frac + (size_t)m * S, vol_frac + (size_t)(m + 1) * S);
    if (vol_frac_std) am.vol_frac_std.assign(vol_frac_std + (size_t)m * S, vol_frac_std + (size_t)(m + 1) * S);
    else am.vol_frac_std.assign(S, 0.0);
  }
  t.bnd_num_conc.assign(num_conc, num_conc + nc * (size_t)n_mode);
  t.bnd_outflow.assign(outflow, outflow + nc);
  t.bnd_throw.assign(throw_away, throw_away + nc);
  return 0;
}
int pmco_trans_probs(void* h, float dt) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  d->trans.dt = dt;
  trans_compute_probs(*d);
  ORACLE_CATCH(d)
}
// adv[6] (-x +x -y +y -z +z), diff[4] (-x +x -y +y), vert[nz] of (cell, class); cell is the 0-based global index
int pmco_trans_get_probs(void* h, int cell, int cl, double* adv, double* diff, double* vert) {
  OracleDomain* d = (OracleDomain*)h;
  const TransDomain& t = d->trans;
  const size_t b = (size_t)cell * t.n_class + (cl - 1);
  for (int q = 0; q < 6; q++) adv[q] = t.adv[b * 6 + q];
  for (int q = 0; q < 4; q++) diff[q] = t.diff[b * 4 + q];
  for (int q = 0; q < t.nz; q++) vert[q] = t.vert[b * t.nz + q];
  return 0;
}
int pmco_trans_aero(void* h, unsigned step, int allow_doubling, int allow_halving, int n_threads,
                    long long* n_moved) {
  OracleDomain* d = (OracleDomain*)h;
  ORACLE_TRY
  trans_aero_run(*d, step, allow_doubling != 0, allow_halving != 0, n_threads, n_moved);
  ORACLE_CATCH(d)
}

// partmc_to_wrf (interface/wrf_pmc_driver.F90:1704-1719): out[cell * C + c]
int pmco_num_conc_by_class(void* h, float* out) {
  OracleDomain* d = (OracleDomain*)h;
  for (size_t g = 0; g < d->cells.size(); g++) {
    const AeroState& st = d->cells[g];
    const int C = st.awa.n_class;
    for (int c = 1; c <= C; c++) {
      int n = 0;
      for (const Particle& p : st.apa)
        if (p.weight_group == 1 && p.weight_class == c) n++;
      const double nc = (double)n * st.awa.at(1, c).magnitude;
      out[g * C + (c - 1)] = std::max((float)nc * (float)d->env[g].inverse_density, 1e-16f);
    }
  }
  return 0;
}

}  // extern "C"
