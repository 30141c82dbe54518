// coag.cu -- mc_coag (partmc/src/coagulation.F90:43-96) for every grid cell.
//
// One warp owns one cell and reproduces the reference's *sequential* semantics
// exactly (bin pairs visited in the order of coagulation.F90:74-94; every
// sample sees the population left by the previous accepted event), so the
// result is bit-identical to the CPU oracle given the same addressed draws:
//   * the 32 lanes scan 32 candidate (b1, b2) pairs at once -- destination
//     class (coag_kernel.F90:382), weight factor majorant (:422), k_max and
//     the Poisson sample count (coagulation.F90:648) -- and the warp then
//     handles the first pair that has samples; pairs behind an accepted event
//     are rescanned with the new counts;
//   * inside a pair the lanes evaluate 32 candidate samples speculatively
//     (find_rand_pair :766, num_conc_weighted_kernel, accept test :753); the
//     first accepted one is applied (coagulate :928, coagulate_weighting :810)
//     and evaluation resumes right behind it.
// Every draw is a pure function Philox(seed; cell, step, phase, pair, sample, j),
// which is what makes the speculation exact.
//
// The size_class multimap of integer_rmap2.F90 is kept per cell as index lists
// in a 3*cap arena (lists start as the contiguous runs the sort produced and
// are relocated with doubled capacity when they outgrow their run), with the
// same swap-with-last bookkeeping as integer_rmap2_remove (:190-243).
#define PMC_RNG_NOINLINE
#include "pmcgpu_internal.cuh"
#include "sedi.cuh"

// This file is compiled twice.  Plain (coag.o): spherical particles, the conversions below are the sphere
// formulas and nothing of the fractal treatment is in the code of the WRF path (k_mc_coag is bound by instruction
// fetch).  With PMC_COAG_FRACTAL (coag_fractal.cu includes this file): aero_data_vol2rad / aero_data_rad2vol
// (aero_data.F90:101-140) with the fractal parameters of the context, plus the brown_free / brown_cont kernels;
// pmc_mc_coag_async hands over to that build when pmcgpu_set_fractal made the particles non-spherical or one of
// the two kernels is asked for.  Each build lives in its own namespace (the kernel templates would collide).
#ifdef PMC_COAG_FRACTAL
#define PMC_COAG_NS coag_fractal
#define CX_VOL2RAD(x) fr_vol2rad(*cx.v, (x))
#define CX_RAD2VOL(x) fr_rad2vol(*cx.v, (x))
#else
#define PMC_COAG_NS coag_sphere
#define CX_VOL2RAD(x) vol2rad(x)
#define CX_RAD2VOL(x) rad2vol(x)
#endif
namespace PMC_COAG_NS {

#ifndef COAG_MIN_CTAS
#define COAG_MIN_CTAS 12  // 40 registers: measured 28.2 (5) / 27.4 (6) / 23.3 (8) / 22.1 (10) / 20.2 (12) / 20.1 (16) ms per K4 step
#endif
#define COAG_WARPS 4
static const double BETA_1 = 1000.0;
static const double BETA_0 = 0.25 / (60.0 * 2e8);
#define COAG_ACCEL_N_EVENT 1.0
#define MAX_ALLOWABLE_GROWTH_FACTOR 1.5
#define COAG_ACCEL_MAX_CV 0.1

struct CellCtx {
  const DevView* v;
  int cell, lane;
  size_t base;   // cell * cap
  size_t NS;     // n_cells * cap
  int nb;
  double gmin, gmax, lnmin, lnmax;
  int* L;        // cell arena
  int* lptr; int* llen; int* lcap;  // per (class, bin): key = (c-1)*nbcap + (b-1)
  int n_part;
  int arena_top;
  int err;
};

__device__ __forceinline__ int lkey(const CellCtx& cx, int c, int b) { return (c - 1) * cx.v->nbcap + (b - 1); }

// aero_particle_volume / aero_particle_mass (aero_particle.F90:316, :252): in-order sums
__device__ __forceinline__ void part_vol_mass(const CellCtx& cx, int p, double& vol, double& mass) {
  const DevView& v = *cx.v;
  double vt = 0.0, mt = 0.0;
#pragma unroll 5
  for (int s = 0; s < v.S; s++) {
    const double x = v.vol[(size_t)s * cx.NS + cx.base + p];
    vt += x;
    mt += x * v.density[s];
  }
  vol = vt;
  mass = mt;
}

// coag_kernel.F90:346-377
__device__ __noinline__ double coag_num_conc_factor(const CellCtx& cx, double i_r, double j_r, int ic, int jc,
                                                       int ijc) {
  const double ij_r = CX_VOL2RAD(CX_RAD2VOL(i_r) + CX_RAD2VOL(j_r));
  const double i_nc = awa_nc_at_radius(*cx.v, cx.cell, ic, i_r);
  const double j_nc = awa_nc_at_radius(*cx.v, cx.cell, jc, j_r);
  const double ij_nc = awa_nc_at_radius(*cx.v, cx.cell, ijc, ij_r);
  const double nc_min = fmin(fmin(i_nc, j_nc), ij_nc);
  return i_nc * j_nc / nc_min;
}

// the cold part of a record, out of line: the instances without component tracking only carry the call
__device__ __noinline__ void move_record_cold(const DevView& v, size_t NS, size_t dst, size_t src, int lane) {
  cold_copy_warp(v.cold, NS, dst, v.cold, NS, src, v.n_opq, lane);
}
// move the record of slot `src` into slot `dst` (aero_particle_array.F90:178-200), all lanes.  COLD is a template
// argument down to here so that the instances without component tracking are the code they were before it existed
// (mc_coag is bound by instruction fetch: 4 KB more code cost 1.7 ms per K4 step).
template <bool COLD>
__device__ __forceinline__ void move_record(CellCtx& cx, int dst, int src) {
  const DevView& v = *cx.v;
  for (int s = cx.lane; s < v.S; s += 32) v.vol[(size_t)s * cx.NS + cx.base + dst] = v.vol[(size_t)s * cx.NS + cx.base + src];
  if (cx.lane == 0) v.id[cx.base + dst] = v.id[cx.base + src];
  if (cx.lane == 1) v.wgc[cx.base + dst] = v.wgc[cx.base + src];
  if (cx.lane == 2) v.nprim[cx.base + dst] = v.nprim[cx.base + src];
  if (cx.lane == 3) v.hyst[cx.base + dst] = v.hyst[cx.base + src];
  if (cx.lane == 4) v.lct[cx.base + dst] = v.lct[cx.base + src];
  if (cx.lane == 5) v.gct[cx.base + dst] = v.gct[cx.base + src];
  if (COLD) move_record_cold(v, cx.NS, cx.base + dst, cx.base + src, cx.lane);
}

// ---- component lists (aero_particle.F90:53-54) held by the lanes of the warp: lane q has component q (q < n).
// Only the COLD instances of the kernel (pmcgpu_track_components) reach this code.
struct LaneComp {
  int n;
  int src;
  double time;
};
__device__ __forceinline__ LaneComp comp_load(const CellCtx& cx, int slot) {
  const DevView& v = *cx.v;
  LaneComp c;
  c.n = v.cold.n[cx.base + slot];
  c.src = (cx.lane < c.n) ? v.cold.src[(size_t)cx.lane * cx.NS + cx.base + slot] : 0;
  c.time = (cx.lane < c.n) ? v.cold.time[(size_t)cx.lane * cx.NS + cx.base + slot] : 0.0;
  return c;
}
// the list of a coagulated particle; its optics are zero until the host recomputes them (aero_particle.F90:920-929)
__device__ __forceinline__ void comp_store(const CellCtx& cx, int slot, const LaneComp& c) {
  const DevView& v = *cx.v;
  if (cx.lane < c.n) {
    v.cold.src[(size_t)cx.lane * cx.NS + cx.base + slot] = c.src;
    v.cold.time[(size_t)cx.lane * cx.NS + cx.base + slot] = c.time;
  }
  for (int q = cx.lane; q < v.n_opq; q += 32) v.cold.opq[(size_t)q * cx.NS + cx.base + slot] = 0.0;
  if (cx.lane == 0) v.cold.n[cx.base + slot] = c.n;
}
// sample_without_replacement (rand.F90:515-542): the 0-based population index of sample number `want` (0-based),
// -1 when want is outside the sample; draws key, j0 + t for the t-th member of the population.  Uniform control flow.
__device__ __noinline__ int swr_pick(int sample_size, int population_size, const RKey& key, unsigned j0, int want) {
  int t = 0, m = 0, res = -1;
  while (m < sample_size && t < 64) {
    const double r = rng_uniform(key, j0 + (unsigned)t);
    if ((double)(population_size - t) * r >= (double)(sample_size - m)) {
      t = t + 1;
    } else {
      if (m == want) res = t;
      t = t + 1;
      m = m + 1;
    }
  }
  return res;
}
// aero_particle_coagulate's rule (:955-972): concatenation, or PMC_MAX_COMP components split between the two
// particles in proportion to their n_primary_parts.  Draws: j0 prob_round, j0 + 1 + t / j0 + 33 + t the two samples.
__device__ __noinline__ LaneComp comp_merge(const LaneComp& A, int nprim_a, const LaneComp& B, int nprim_b,
                                            const RKey& key, unsigned j0, int lane) {
  LaneComp R;
  int pick_a, pick_b, n_a;
  if (A.n + B.n <= PMC_MAX_COMP) {
    n_a = A.n;
    R.n = A.n + B.n;
    pick_a = lane;
    pick_b = lane - A.n;
  } else {
    n_a = prob_round((double)nprim_a / (double)(nprim_a + nprim_b) * (double)PMC_MAX_COMP, rng_uniform(key, j0));
    n_a = min(n_a, A.n);  // holds upstream whenever n_components = min(n_primary_parts, 20)
    const int n_b = min(PMC_MAX_COMP - n_a, B.n);
    R.n = n_a + n_b;
    pick_a = swr_pick(n_a, A.n, key, j0 + 1u, lane);
    pick_b = swr_pick(n_b, B.n, key, j0 + 33u, lane - n_a);
  }
  const int sa = __shfl_sync(0xffffffffu, A.src, pick_a & 31), sb = __shfl_sync(0xffffffffu, B.src, pick_b & 31);
  const double ta = __shfl_sync(0xffffffffu, A.time, pick_a & 31), tb = __shfl_sync(0xffffffffu, B.time, pick_b & 31);
  R.src = (lane < n_a) ? sa : sb;
  R.time = (lane < n_a) ? ta : tb;
  return R;
}
// the averaged source of the accelerated path after its n_primary_parts were rescaled (coagulation.F90:472-508):
// down to min(n_primary_parts, 20) components by sampling, or up to it by whole copies plus a sample of the rest
__device__ __noinline__ LaneComp comp_resample(const LaneComp& S, int nprim_new, const RKey& key, int lane) {
  const int n = S.n, n_new = min(nprim_new, PMC_MAX_COMP);
  int pick = lane;
  LaneComp R = S;
  if (n_new < n) {
    pick = swr_pick(n_new, n, key, 1u, lane);
    R.n = n_new;
  } else if (n < PMC_MAX_COMP && n > 0) {
    int i_comp = 1, n_whole = 0;
    while (i_comp + n <= n_new) {
      n_whole++;
      i_comp += n;
    }
    const int left = (i_comp <= n_new) ? n_new - i_comp + 1 : 0;
    const int lp = swr_pick(left, n, key, 1u, lane - n_whole * n);
    pick = (lane < n_whole * n) ? lane % n : lp;
    R.n = n_whole * n + left;
  } else {
    return R;
  }
  R.src = __shfl_sync(0xffffffffu, S.src, pick & 31);
  R.time = __shfl_sync(0xffffffffu, S.time, pick & 31);
  return R;
}

// aero_sorted_remove_particle (aero_sorted.F90:455-470): apa swap-with-last +
// integer_rmap2_remove on size_class.  p is a 0-based slot.
template <bool COLD>
__device__ __noinline__ void cell_remove_particle(CellCtx& cx, int p) {
  const DevView& v = *cx.v;
  const int last = cx.n_part - 1;
  if (cx.lane == 0) {
    const int b_old = v.pbin[cx.base + p];
    const int c_old = v.wgc[cx.base + p] / v.G + 1;
    const int pos_old = v.fidx[cx.base + p];
    if (last != p) {
      const int b_fix = v.pbin[cx.base + last];
      const int c_fix = v.wgc[cx.base + last] / v.G + 1;
      const int pos_fix = v.fidx[cx.base + last];
      cx.L[cx.lptr[lkey(cx, c_fix, b_fix)] + pos_fix] = p;
      v.pbin[cx.base + p] = b_fix;
      v.fidx[cx.base + p] = pos_fix;
    }
    const int k_old = lkey(cx, c_old, b_old);
    const int lp = cx.lptr[k_old];
    const int pos_last = cx.llen[k_old] - 1;
    const int dom_fix = cx.L[lp + pos_last];
    if (pos_old < pos_last) cx.L[lp + pos_old] = cx.L[lp + pos_last];
    cx.llen[k_old] = pos_last;
    if (pos_last != pos_old) v.fidx[cx.base + dom_fix] = pos_old;
  }
  __syncwarp();
  if (last != p) move_record<COLD>(cx, p, last);
  cx.n_part = last;
  __syncwarp();
}

// integer_varray_append on list (c, b) with pow2-style growth inside the arena
__device__ __noinline__ void cell_list_append(CellCtx& cx, int c, int b, int slot) {
  const DevView& v = *cx.v;
  const int k = lkey(cx, c, b);
  int len = cx.llen[k], capk = cx.lcap[k], lp = cx.lptr[k];
  if (len == capk) {
    const int ncap = max(4, 2 * capk);
    if (cx.arena_top + ncap > 3 * v.cap) {
      cx.err = PMCGPU_ERR_CAPACITY;
      return;
    }
    const int np = cx.arena_top;
    for (int q = cx.lane; q < len; q += 32) cx.L[np + q] = cx.L[lp + q];
    __syncwarp();
    cx.arena_top += ncap;
    lp = np;
    if (cx.lane == 0) {
      cx.lptr[k] = np;
      cx.lcap[k] = ncap;
    }
  }
  if (cx.lane == 0) {
    cx.L[lp + len] = slot;
    v.fidx[cx.base + slot] = len;
    cx.llen[k] = len + 1;
  }
  __syncwarp();
}

// KT: the kernel type when it is known at compile time (the Brownian instance carries no code of the other
// kernels: mc_coag is instruction-fetch bound), 0 = decided by ktype at run time
#ifdef PMC_COAG_FRACTAL
#define UNWEIGHTED_KERNEL(ktype, be, v1, m1, v2, m2) unweighted_kernel_fr(*cx.v, cx.cell, ktype, be, v1, m1, v2, m2)
#else
#define UNWEIGHTED_KERNEL(ktype, be, v1, m1, v2, m2) unweighted_kernel<KT>(ktype, be, v1, m1, v2, m2)
#endif
template <int KT>
__device__ __forceinline__ double unweighted_kernel(int ktype, const BrownEnv& be, double v1, double m1, double v2,
                                                    double m2) {
  if (KT == PMCGPU_KERNEL_BROWN) return brown_pair(brown_part(be, v1, m1 / v1), brown_part(be, v2, m2 / v2));
  if (ktype == PMCGPU_KERNEL_BROWN) {
    return brown_pair(brown_part(be, v1, m1 / v1), brown_part(be, v2, m2 / v2));
  } else if (ktype == PMCGPU_KERNEL_ADDITIVE) {
    return BETA_1 * (v1 + v2);
  } else if (ktype == PMCGPU_KERNEL_CONSTANT) {
    return BETA_0;
  } else if (ktype == PMCGPU_KERNEL_SEDI) {
    return kernel_sedi(v1, v2);
  }
  return 0.0;
}
#ifdef PMC_COAG_FRACTAL
// kernel (coag_kernel.F90:99-124) with the two kernels of Vemury and Pratsinis [1995]; densities = mass / volume
__device__ __noinline__ double unweighted_kernel_fr(const DevView& v, int cell, int ktype, const BrownEnv& be, double v1,
                                                    double m1, double v2, double m2) {
  if (ktype == PMCGPU_KERNEL_BROWN_FREE) return kernel_brown_free(v, v1, m1 / v1, v2, m2 / v2, v.temp[cell]);
  if (ktype == PMCGPU_KERNEL_BROWN_CONT) return kernel_brown_cont(v, v1, v2, v.temp[cell]);
  return unweighted_kernel<0>(ktype, be, v1, m1, v2, m2);
}
#endif

// coagulate (coagulation.F90:928-985) + coagulate_weighting (:810-922) for the
// particles in slots p1, p2 (0-based); warp-cooperative, uniform arguments.
template <bool COLD>
__device__ __noinline__ void cell_coagulate(CellCtx& cx, int p1, int p2, int c1, int c2, int cc, const RKey& key) {
  const DevView& v = *cx.v;
  double v1, m1, v2, m2;
  part_vol_mass(cx, p1, v1, m1);
  part_vol_mass(cx, p2, v2, m2);
  const double r1 = CX_VOL2RAD(v1), r2 = CX_VOL2RAD(v2);
  const double rc = CX_VOL2RAD(CX_RAD2VOL(r1) + CX_RAD2VOL(r2));
  const double nc1 = awa_nc_at_radius(v, cx.cell, c1, r1);
  const double nc2 = awa_nc_at_radius(v, cx.cell, c2, r2);
  const double ncc = awa_nc_at_radius(v, cx.cell, cc, rc);
  // aero_weight_array_rand_group (aero_weight_array.F90:374) -> sample_cts_pdf (rand.F90:550), draws 6..
  int new_group = 1;
  if (v.G > 1) {
    const size_t wb = ((size_t)cx.cell * v.C + (cc - 1)) * v.G;
    double pdf_max = 0.0;
    for (int g = 0; g < v.G; g++) pdf_max = fmax(pdf_max, 1.0 / weight_nc(v.wtype[wb + g], v.wmag[wb + g], v.wexp[wb + g], rc));
    RStream sg{key, 6};
    for (;;) {
      const int k = pmc_rand_int(sg.next(), v.G);
      const double pk = 1.0 / weight_nc(v.wtype[wb + k - 1], v.wmag[wb + k - 1], v.wexp[wb + k - 1], rc);
      if (sg.next() < pk / pdf_max) {
        new_group = k;
        break;
      }
    }
  }
  const double nc_min = fmin(fmin(nc1, nc2), ncc);
  const bool remove_1 = rng_uniform(key, 3) < nc_min / nc1;
  const bool remove_2 = rng_uniform(key, 4) < nc_min / nc2;
  const bool create_new = rng_uniform(key, 5) < nc_min / ncc;
  bool id_1_lost, id_2_lost;
  if (create_new) {
    id_1_lost = false;
    id_2_lost = false;
    if (remove_1 && remove_2) {
      if (v1 > v2) id_2_lost = true; else id_1_lost = true;
    }
  } else {
    id_1_lost = remove_1;
    id_2_lost = remove_2;
  }
  // aero_particle_coagulate (aero_particle.F90:900-980) into registers
  double volc[(PMC_MAX_SPEC + 31) / 32];
  long long idc = 0;
  int hystc = 0, nprimc = 0;
  double lctc = 0.0, gctc = 0.0;
  LaneComp compc{0, 0, 0.0};
  if (COLD && create_new) {
    RKey ckey = key;
    ckey.phase = PH_COMP;
    compc = comp_merge(comp_load(cx, p1), v.nprim[cx.base + p1], comp_load(cx, p2), v.nprim[cx.base + p2], ckey, 0u,
                       cx.lane);
  }
  if (create_new) {
#pragma unroll
    for (int q = 0; q < (PMC_MAX_SPEC + 31) / 32; q++) {
      const int s = q * 32 + cx.lane;
      volc[q] = (s < v.S) ? v.vol[(size_t)s * cx.NS + cx.base + p1] + v.vol[(size_t)s * cx.NS + cx.base + p2] : 0.0;
    }
    const long long id1 = v.id[cx.base + p1], id2 = v.id[cx.base + p2];
    hystc = (v.hyst[cx.base + p1] == 1 && v.hyst[cx.base + p2] == 1) ? 1 : 0;
    nprimc = v.nprim[cx.base + p1] + v.nprim[cx.base + p2];
    gctc = fmax(v.gct[cx.base + p1], v.gct[cx.base + p2]);
    lctc = fmin(v.lct[cx.base + p1], v.lct[cx.base + p2]);
    if (remove_1 && !id_1_lost) {
      idc = id1;
    } else if (remove_2 && !id_2_lost) {
      idc = id2;
    } else {
      // aero_particle_new_id (aero_particle.F90:141-149) with one counter per cell
      const long long nx = v.next_id[cx.cell];
      idc = (nx - 1) * v.n_cells_global + v.gcell[cx.cell] + 1;
      __syncwarp();
      if (cx.lane == 0) v.next_id[cx.cell] = nx + 1;
    }
  }
  __syncwarp();
  // remove old particles (coagulation.F90:948-970)
  if (p2 > p1) {
    if (remove_2) cell_remove_particle<COLD>(cx, p2);
    if (remove_1) cell_remove_particle<COLD>(cx, p1);
  } else {
    if (remove_1) cell_remove_particle<COLD>(cx, p1);
    if (remove_2) cell_remove_particle<COLD>(cx, p2);
  }
  // add new particle, allow_resort = .false. (aero_sorted.F90:397-449)
  if (create_new) {
    double vt = 0.0;
    for (int s = 0; s < v.S; s++) vt += __shfl_sync(0xffffffffu, volc[s >> 5], s & 31);
    const int b_new = log_grid_find(cx.lnmin, cx.lnmax, cx.nb, log(CX_VOL2RAD(vt)));
    if (b_new < 1 || b_new > cx.nb) {
      cx.err = PMCGPU_ERR_BIN_RANGE;
      return;
    }
    if (cx.n_part >= v.cap) {
      cx.err = PMCGPU_ERR_CAPACITY;
      return;
    }
    const int slot = cx.n_part;
#pragma unroll
    for (int q = 0; q < (PMC_MAX_SPEC + 31) / 32; q++) {
      const int s = q * 32 + cx.lane;
      if (s < v.S) v.vol[(size_t)s * cx.NS + cx.base + slot] = volc[q];
    }
    if (cx.lane == 0) {
      v.id[cx.base + slot] = idc;
      v.wgc[cx.base + slot] = (cc - 1) * v.G + (new_group - 1);
      v.nprim[cx.base + slot] = nprimc;
      v.hyst[cx.base + slot] = hystc;
      v.lct[cx.base + slot] = lctc;
      v.gct[cx.base + slot] = gctc;
      v.pbin[cx.base + slot] = b_new;
    }
    if (COLD) comp_store(cx, slot, compc);
    cx.n_part = slot + 1;
    __syncwarp();
    cell_list_append(cx, cc, b_new, slot);
  }
}

// stats.F90:510-546
__device__ __forceinline__ double student_t_95_coeff(int n_sample) {
  const double limit = 1.959963984540054;
  const int n_dof = n_sample - 1;
  if (n_dof <= 15) {
    switch (n_dof) {
      case 1: return 12.7062047364;
      case 2: return 4.30265272991;
      case 3: return 3.18244630528;
      case 4: return 2.7764451052;
      case 5: return 2.57058183661;
      case 6: return 2.44691184879;
      case 7: return 2.36462425101;
      case 8: return 2.30600413503;
      case 9: return 2.26215716274;
      case 10: return 2.22813885196;
      case 11: return 2.20098516008;
      case 12: return 2.17881282966;
      case 13: return 2.16036865646;
      case 14: return 2.14478668792;
      default: return 2.13144954556;
    }
  }
  if (n_dof <= 20) return (2.11990529922 - limit) * 16.0 / (double)n_dof + limit;
  return (2.07961384473 - limit) * 21.0 / (double)n_dof + limit;
}

// determine_target_and_source (coagulation.F90:247-307) for bins (b1, b2) of classes (c1, c2); the
// monotone_decreasing test over the whole weight array is done by the caller.
__device__ __forceinline__ bool accel_target_source(const CellCtx& cx, bool pair_flat, const double* hm, int b1,
                                                    int b2, int c1, int c2, int cc, int& bt, int& bs, int& ct,
                                                    int& cs) {
  const DevView& v = *cx.v;
  double nc1_min, nc1_max, nc2_min, nc2_max;
  if (pair_flat) {
    nc1_min = nc1_max = hm[c1 - 1];
    nc2_min = nc2_max = hm[c2 - 1];
  } else {
    const double a1 = awa_nc_at_radius(v, cx.cell, c1, grid_edge(cx.gmin, cx.gmax, cx.nb, b1));
    const double a2 = awa_nc_at_radius(v, cx.cell, c1, grid_edge(cx.gmin, cx.gmax, cx.nb, b1 + 1));
    const double e1 = awa_nc_at_radius(v, cx.cell, c2, grid_edge(cx.gmin, cx.gmax, cx.nb, b2));
    const double e2 = awa_nc_at_radius(v, cx.cell, c2, grid_edge(cx.gmin, cx.gmax, cx.nb, b2 + 1));
    nc1_min = fmin(a1, a2); nc1_max = fmax(a1, a2);
    nc2_min = fmin(e1, e2); nc2_max = fmax(e1, e2);
  }
  bool ok = false;
  if ((nc1_max < nc2_min) && (cc == c1)) { bt = b1; bs = b2; ct = c1; cs = c2; ok = true; }
  if ((nc2_max < nc1_min) && (cc == c2)) { bt = b2; bs = b1; ct = c2; cs = c1; ok = true; }
  return ok;
}

// aero_weight_array_rand_group (aero_weight_array.F90:374) -> sample_cts_pdf (rand.F90:550) on the
// sequential stream s: two draws per rejection round, also when there is a single group
__device__ __forceinline__ int rand_group_seq(const CellCtx& cx, int c, double radius, RStream& s) {
  const DevView& v = *cx.v;
  if (v.G == 1) {
    s.j += 2;
    return 1;
  }
  const size_t wb = ((size_t)cx.cell * v.C + (c - 1)) * v.G;
  double pdf_max = 0.0;
  for (int g = 0; g < v.G; g++) pdf_max = fmax(pdf_max, 1.0 / weight_nc(v.wtype[wb + g], v.wmag[wb + g], v.wexp[wb + g], radius));
  for (;;) {
    const int k = pmc_rand_int(s.next(), v.G);
    const double pk = 1.0 / weight_nc(v.wtype[wb + k - 1], v.wmag[wb + k - 1], v.wexp[wb + k - 1], radius);
    if (s.next() < pk / pdf_max) return k;
  }
}

// integer_rmap2_change (integer_rmap2.F90:139-185) on size_class for slot p of class c: only the bin
// changes on the accelerated path (ct == cc); an unchanged bin returns early like the reference (:160)
__device__ __noinline__ void cell_change_bin(CellCtx& cx, int p, int c, int new_bin) {
  const DevView& v = *cx.v;
  const int b_old = v.pbin[cx.base + p];
  if (b_old == new_bin) return;
  __syncwarp();
  if (cx.lane == 0) {
    const int k_old = lkey(cx, c, b_old);
    const int lp = cx.lptr[k_old];
    const int pos_old = v.fidx[cx.base + p];
    const int pos_last = cx.llen[k_old] - 1;
    if (pos_old < pos_last) {
      const int moved = cx.L[lp + pos_last];
      cx.L[lp + pos_old] = moved;
      v.fidx[cx.base + moved] = pos_old;
    }
    cx.llen[k_old] = pos_last;
    v.pbin[cx.base + p] = new_bin;
  }
  __syncwarp();
  cell_list_append(cx, c, new_bin, p);
}

// aero_state_dup_particle (aero_state.F90:488-539) for the sorted state of mc_coag: slot p of class c
template <bool COLD>
__device__ __noinline__ void cell_dup_particle(CellCtx& cx, int p, int c, double n_part_mean, bool random_weight_group,
                                  RStream& s) {
  const DevView& v = *cx.v;
  const int n_copies = prob_round(n_part_mean, s.next());
  if (n_copies == 0) {
    cell_remove_particle<COLD>(cx, p);
    return;
  }
  for (int i_dup = 1; i_dup <= n_copies - 1 && !cx.err; i_dup++) {
    int new_group = 0;
    if (random_weight_group) {
      double vt, mt;
      part_vol_mass(cx, p, vt, mt);
      new_group = rand_group_seq(cx, c, CX_VOL2RAD(vt), s);
    }
    if (cx.n_part >= v.cap) {
      cx.err = PMCGPU_ERR_CAPACITY;
      return;
    }
    const int slot = cx.n_part;
    const int b = v.pbin[cx.base + p];
    __syncwarp();
    move_record<COLD>(cx, slot, p);
    __syncwarp();
    if (cx.lane == 0) {
      // aero_particle_new_id (aero_particle.F90:141-149) with one counter per cell
      const long long nx = v.next_id[cx.cell];
      v.id[cx.base + slot] = (nx - 1) * v.n_cells_global + v.gcell[cx.cell] + 1;
      v.next_id[cx.cell] = nx + 1;
      if (new_group) v.wgc[cx.base + slot] = (c - 1) * v.G + (new_group - 1);
      v.pbin[cx.base + slot] = b;
    }
    cx.n_part = slot + 1;
    __syncwarp();
    cell_list_append(cx, c, b, slot);
  }
}

// try_per_particle_coag (coagulation.F90:150-241) for a bin pair the scan found eligible: every target
// particle (backwards through its bin) coagulates with the average of a sample of source particles
// (sample_source_particle :333-517, coag_target_with_source :522-582).  Warp-cooperative: the 32 lanes
// evaluate the kernel of 32 candidate samples ahead (their draws are addressed by sample number), the
// warp then consumes them in order with the running sums held one species per lane; a removal of a
// source particle invalidates the candidates behind it, which are re-evaluated.
template <int KT, bool COLD>
__device__ __noinline__ void cell_accel_pair(CellCtx& cx, int ktype, const BrownEnv& be, double k_max, double del_t, unsigned step,
                                unsigned pid, int bt, int bs, int ct, int cs, int cc, int& tot_n_samp,
                                int& tot_n_coag) {
  const DevView& v = *cx.v;
  constexpr int NQ = (PMC_MAX_SPEC + 31) / 32;
  const int kt = lkey(cx, ct, bt), ks = lkey(cx, cs, bs);
  const double n_samp_mean = k_max * del_t * (double)cx.llen[ks];  // compute_n_source :310-329
  const double accept_factor = 1.0 / k_max;
  for (int e = cx.llen[kt]; e >= 1 && !cx.err; e--) {
    __syncwarp();
    const int target0 = cx.L[cx.lptr[kt] + e - 1];
    double v_t, m_t;
    part_vol_mass(cx, target0, v_t, m_t);
    const double r_t = CX_VOL2RAD(v_t);
    RStream s{RKey{v.seed, (unsigned)v.gcell[cx.cell], step, PH_COAG_ACCEL, pid, (unsigned)e}, 0u};
    const RKey skey{v.seed, (unsigned)v.gcell[cx.cell], step, PH_COAG_ACCEL, pid | 0x80000000u, (unsigned)e};
    int n_samp = 0, n_coag = 0, n_avg = 0;
    if (cx.llen[ks] == 0) continue;
    const double num_conc_target = awa_nc_at_radius(v, cx.cell, ct, r_t);
    const double num_conc_source_min = awa_nc_at_radius(v, cx.cell, cs, grid_edge(cx.gmin, cx.gmax, cx.nb, bs));
    const double prob_remove_source_max = num_conc_target / num_conc_source_min;
    if (!(prob_remove_source_max <= 1.0)) {
      cx.err = PMCGPU_ERR_ACCEL;  // assert 653606684
      return;
    }
    const int n_samp_remove = rand_poisson(prob_remove_source_max * n_samp_mean, s);
    const int n_samp_extra = rand_poisson((1.0 - prob_remove_source_max) * n_samp_mean, s);
    const int n_samp_total = n_samp_remove + n_samp_extra;
    double prob_coag_tot = 0.0;
    double src_vol[NQ], vol_sq[NQ];
#pragma unroll
    for (int q = 0; q < NQ; q++) { src_vol[q] = 0.0; vol_sq[q] = 0.0; }
    int src_nprim = 0;
    double src_lct = 0.0, src_gct = 0.0;  // aero_particle_zero
    LaneComp src_comp{0, 0, 0.0};
    RKey ckey = skey;
    ckey.phase = PH_COMP;
    int i_samp = 1;
    bool done = false;
    while (i_samp <= n_samp_total && !done) {
      __syncwarp();
      const int ne = cx.llen[ks];
      if (ne == 0) break;
      // ---- candidates i_samp + lane
      const int my_i = i_samp + cx.lane;
      int i_part = 0;
      double prob_coag = 0.0, u_acc = 1.0, u_rem = 1.0, nc_i = 1.0;
      if (my_i <= n_samp_total) {
        const unsigned j0 = 3u * (unsigned)(my_i - 1);
        i_part = cx.L[cx.lptr[ks] + pmc_rand_int(rng_uniform(skey, j0), ne) - 1];
        double v_i, m_i;
        part_vol_mass(cx, i_part, v_i, m_i);
        const double r_i = CX_VOL2RAD(v_i);
        const double uk = UNWEIGHTED_KERNEL(ktype, be, v_i, m_i, v_t, m_t);
        prob_coag = (uk * coag_num_conc_factor(cx, r_i, r_t, cs, ct, ct)) * accept_factor;
        u_acc = rng_uniform(skey, j0 + 1);
        u_rem = rng_uniform(skey, j0 + 2);
        nc_i = awa_nc_at_radius(v, cx.cell, cs, r_i);
      }
      // ---- in-order consumption
      int l = 0;
      for (; l < 32 && i_samp + l <= n_samp_total; l++) {
        if ((n_samp > n_samp_remove) && (n_avg >= 2)) {
          double cv_max = 0.0;
#pragma unroll
          for (int q = 0; q < NQ; q++) {
            const double vol_mean = src_vol[q] / (double)n_avg;
            if (vol_mean > 0.0)
              cv_max = fmax(cv_max, sqrt(fmax(0.0, vol_sq[q] / (double)n_avg - vol_mean * vol_mean)) / vol_mean);
          }
          for (int o = 16; o > 0; o >>= 1) cv_max = fmax(cv_max, __shfl_xor_sync(0xffffffffu, cv_max, o));
          const double mean_95_conf_cv = cv_max / sqrt((double)n_avg) * student_t_95_coeff(n_avg);
          if (mean_95_conf_cv < COAG_ACCEL_MAX_CV) {
            done = true;
            break;
          }
        }
        n_samp++;
        const double pc = __shfl_sync(0xffffffffu, prob_coag, l);
        prob_coag_tot = prob_coag_tot + pc;
        if (__shfl_sync(0xffffffffu, u_acc, l) < pc) {
          n_avg++;
          const int ip = __shfl_sync(0xffffffffu, i_part, l);
          if (COLD)
            src_comp = comp_merge(src_comp, src_nprim, comp_load(cx, ip), v.nprim[cx.base + ip], ckey,
                                  128u * (unsigned)(i_samp + l), cx.lane);
#pragma unroll
          for (int q = 0; q < NQ; q++) {
            const int sp = q * 32 + cx.lane;
            const double x = (sp < v.S) ? v.vol[(size_t)sp * cx.NS + cx.base + ip] : 0.0;
            src_vol[q] = src_vol[q] + x;
            vol_sq[q] = vol_sq[q] + x * x;
          }
          src_nprim += v.nprim[cx.base + ip];
          src_gct = fmax(src_gct, v.gct[cx.base + ip]);
          src_lct = fmin(src_lct, v.lct[cx.base + ip]);
          if (i_samp + l <= n_samp_remove) {
            const double prob_remove_i = num_conc_target / __shfl_sync(0xffffffffu, nc_i, l);
            if (__shfl_sync(0xffffffffu, u_rem, l) < prob_remove_i / prob_remove_source_max) {
              cell_remove_particle<COLD>(cx, ip);
              l++;
              break;  // the source bin changed: candidates behind this one are stale
            }
          }
        }
      }
      i_samp += l;
    }
    if (n_avg > 0) {
      const double prob_coag_mean = prob_coag_tot / (double)n_samp;  // note not n_avg
      if (!(prob_coag_mean <= 1.0)) {
        cx.err = PMCGPU_ERR_ACCEL;  // rand_binomial asserts p <= 1 (rand.F90:333)
        return;
      }
      n_coag = rand_binomial(n_samp_total, prob_coag_mean, s);
      const double f = (double)n_coag / (double)n_avg;
#pragma unroll
      for (int q = 0; q < NQ; q++) src_vol[q] = src_vol[q] * f;
      src_nprim = prob_round((double)src_nprim * ((double)n_coag / n_avg), s.next());
      if (COLD) src_comp = comp_resample(src_comp, src_nprim, ckey, cx.lane);
    }
    if (n_coag > 0) {
      // ---- coag_target_with_source (coagulation.F90:522-582)
      __syncwarp();
      const int tp = cx.L[cx.lptr[kt] + e - 1];  // re-read: removals may have moved the target
      double v_old, m_old;
      part_vol_mass(cx, tp, v_old, m_old);
      const double old_num_conc_target = awa_nc_at_radius(v, cx.cell, ct, CX_VOL2RAD(v_old));
      __syncwarp();
      if (COLD) {
        const LaneComp merged = comp_merge(comp_load(cx, tp), v.nprim[cx.base + tp], src_comp, src_nprim, ckey, 64u, cx.lane);
        __syncwarp();
        comp_store(cx, tp, merged);
      }
#pragma unroll
      for (int q = 0; q < NQ; q++) {
        const int sp = q * 32 + cx.lane;
        if (sp < v.S) v.vol[(size_t)sp * cx.NS + cx.base + tp] = v.vol[(size_t)sp * cx.NS + cx.base + tp] + src_vol[q];
      }
      if (cx.lane == 0) {
        v.hyst[cx.base + tp] = 0;  // the averaged source starts from aero_particle_zero: leg 0
        v.nprim[cx.base + tp] = v.nprim[cx.base + tp] + src_nprim;
        v.gct[cx.base + tp] = fmax(v.gct[cx.base + tp], src_gct);
        v.lct[cx.base + tp] = fmin(v.lct[cx.base + tp], src_lct);
      }
      __syncwarp();
      double v_new, m_new;
      part_vol_mass(cx, tp, v_new, m_new);
      const double r_new = CX_VOL2RAD(v_new);
      const int new_group = rand_group_seq(cx, cc, r_new, s);
      if (cx.lane == 0) v.wgc[cx.base + tp] = (cc - 1) * v.G + (new_group - 1);
      const int new_bin = log_grid_find(cx.lnmin, cx.lnmax, cx.nb, log(r_new));
      if (new_bin < 1 || new_bin > cx.nb) {
        cx.err = PMCGPU_ERR_BIN_RANGE;
        return;
      }
      cell_change_bin(cx, tp, cc, new_bin);
      const double new_num_conc_target = awa_nc_at_radius(v, cx.cell, cc, r_new);
      cell_dup_particle<COLD>(cx, tp, cc, old_num_conc_target / new_num_conc_target, true, s);
    }
    tot_n_samp += n_samp;
    tot_n_coag += n_coag;
  }
}

// the reference's bin-pair order (coagulation.F90:74-94) as a flat sequence: step k pairs forward from
// (c1, c2, b1, b2), skipping rows whose first bin is empty; false past the end
__device__ __forceinline__ bool pair_advance(const CellCtx& cx, int C, int nb, int& c1, int& c2, int& b1, int& b2,
                                             int k) {
  for (;;) {
    if (cx.llen[lkey(cx, c1, b1)] > 0) {
      const int rem = nb - b2 + 1;
      if (k < rem) {
        b2 += k;
        return true;
      }
      k -= rem;
    }
    b1++;
    if (b1 > nb) {
      b1 = 1;
      c2++;
      if (c2 > c1) {
        c1++;
        c2 = 1;
        if (c1 > C) return false;
      }
    }
    b2 = (c1 == c2) ? b1 : 1;
  }
}
// one pair forward, whatever the occupancy of the row (the pair behind the one just handled)
__device__ __forceinline__ bool pair_advance_exact(int C, int nb, int& c1, int& c2, int& b1, int& b2) {
  b2++;
  if (b2 <= nb) return true;
  b1++;
  if (b1 > nb) {
    b1 = 1;
    c2++;
    if (c2 > c1) {
      c1++;
      c2 = 1;
      if (c1 > C) return false;
    }
  }
  b2 = (c1 == c2) ? b1 : 1;
  return true;
}

template <int KT, bool COLD, int MIN_CTAS>
__global__ void __launch_bounds__(COAG_WARPS * 32, MIN_CTAS) k_mc_coag(const __grid_constant__ DevView v, int ktype, double del_t, unsigned step) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * COAG_WARPS + warp;
  if (cell >= v.n_cells) return;
  extern __shared__ double s_dyn[];
  double* hm = s_dyn + (size_t)warp * v.C;                                  // harmonic magnitude per class
  int* flat = (int*)(s_dyn + (size_t)COAG_WARPS * v.C) + (size_t)warp * v.C;  // class has only flat weights

  CellCtx cx;
  cx.v = &v;
  cx.cell = cell;
  cx.lane = lane;
  cx.base = (size_t)cell * v.cap;
  cx.NS = (size_t)v.n_cells * v.cap;
  cx.nb = v.grid_nb[cell];
  cx.gmin = v.grid_min[cell];
  cx.gmax = v.grid_max[cell];
  cx.lnmin = cx.nb > 0 ? log(cx.gmin) : 0.0;
  cx.lnmax = cx.nb > 0 ? log(cx.gmax) : 0.0;
  cx.L = v.L + (size_t)cell * 3 * v.cap;
  cx.lptr = v.lptr + (size_t)cell * v.C * v.nbcap;
  cx.llen = v.llen + (size_t)cell * v.C * v.nbcap;
  cx.lcap = v.lcap + (size_t)cell * v.C * v.nbcap;
  cx.n_part = v.count[cell];
  cx.arena_top = v.cap;
  cx.err = 0;
  int tot_n_samp = 0, tot_n_coag = 0;
  const int nb = cx.nb;
  if (v.status[cell] != 0 || nb < 1 || cx.n_part == 0 || cell_is_boundary(v, cell)) {
    if (lane == 0) {
      v.coag_samp[cell] = 0;
      v.coag_ncoag[cell] = 0;
    }
    return;
  }
  // ---- list set-up: exclusive scan of bin_count in (class, bin) order
  {
    const int* bc = v.bin_count + (size_t)cell * v.C * v.nbcap;
    int carry = 0;
    const int nk = v.C * v.nbcap;
    for (int k0 = 0; k0 < nk; k0 += 32) {
      const int k = k0 + lane;
      const int x = (k < nk) ? bc[k] : 0;
      int incl = x;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (k < nk) {
        cx.lptr[k] = carry + incl - x;
        cx.llen[k] = x;
        cx.lcap[k] = x;
      }
      carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    __syncwarp();
    for (int p = lane; p < cx.n_part; p += 32) {
      cx.L[p] = p;
      const int k = (v.wgc[cx.base + p] / v.G) * v.nbcap + (v.pbin[cx.base + p] - 1);
      v.fidx[cx.base + p] = p - cx.lptr[k];
    }
    // per-class weight summary; monotone_decreasing over the whole array (aero_weight_array.F90:304)
    for (int c = lane; c < v.C; c += 32) {
      flat[c] = class_is_flat(v, cell, c + 1) ? 1 : 0;
      hm[c] = awa_nc_at_radius(v, cell, c + 1, 1.0);
    }
    __syncwarp();
  }
  bool mono_dec = true, all_flat = true;
  for (int q = 0; q < v.C * v.G; q++) {
    const size_t wi = (size_t)cell * v.C * v.G + q;
    if (v.wtype[wi] == PMCGPU_WEIGHT_POWER && v.wexp[wi] > 0.0) mono_dec = false;
    if (v.wtype[wi] != PMCGPU_WEIGHT_NONE) all_flat = false;
  }
  const BrownEnv be = brown_env(v.temp[cell], v.pres[cell]);
  const double* ktab = v.kmax + (size_t)cell * v.nbcap * v.nbcap;

  // The bin pairs are visited in the reference's order (coagulation.F90:74-94: c1, c2 <= c1, b1, b2 from b1
  // resp. 1) through one cursor over the flattened sequence: the 32 lanes scan the next 32 pairs -- across
  // row and class boundaries, rows with an empty first bin skipped like the reference's `cycle` -- and the
  // warp handles the first pair that has samples; after an event the scan resumes right behind that pair.
  int cu_c1 = 1, cu_c2 = 1, cu_b1 = 1, cu_b2 = 1;
  bool more = true;
  while (more && !cx.err) {
    {
      {
        {
          // ---- lane-parallel scan of 32 candidate bin pairs
          int c1 = cu_c1, c2 = cu_c2, b1 = cu_b1, my_b2 = cu_b2;
          const bool valid = pair_advance(cx, v.C, nb, c1, c2, b1, my_b2, lane);
          more = __shfl_sync(0xffffffffu, (int)valid, 0) != 0;
          if (!more) break;
          int n_samp = 0, cc = c2, accel = 0;
          double k_max = 0.0;
          int n1 = 0, n2 = 0;
          if (valid) {
            n1 = cx.llen[lkey(cx, c1, b1)];
            n2 = cx.llen[lkey(cx, c2, my_b2)];
          }
          if (n1 > 0 && n2 > 0) {
            double f_max;
            const bool pair_flat = all_flat || (flat[c1 - 1] && flat[c2 - 1]);
            if (pair_flat) {
              // flat weights: every radius gives the same factor (coag_kernel.F90:382-465)
              cc = (hm[c1 - 1] < hm[c2 - 1]) ? c1 : c2;
              const double ncm = fmin(fmin(hm[c1 - 1], hm[c2 - 1]), hm[cc - 1]);
              f_max = fmax(0.0, hm[c1 - 1] * hm[c2 - 1] / ncm);
            } else {
              // coag_dest_class: sic, both radii from bin b1 (coag_kernel.F90:402-403)
              const double i_r = grid_center(cx.gmin, cx.gmax, nb, b1);
              const double ij_r = CX_VOL2RAD(CX_RAD2VOL(i_r) + CX_RAD2VOL(i_r));
              cc = (awa_nc_at_radius(v, cell, c1, ij_r) < awa_nc_at_radius(v, cell, c2, ij_r)) ? c1 : c2;
              const double i_lo = grid_edge(cx.gmin, cx.gmax, nb, b1), i_hi = grid_edge(cx.gmin, cx.gmax, nb, b1 + 1);
              const double j_lo = grid_edge(cx.gmin, cx.gmax, nb, my_b2), j_hi = grid_edge(cx.gmin, cx.gmax, nb, my_b2 + 1);
              f_max = 0.0;
              for (int is = 1; is <= 5; is++)
                for (int js = 1; js <= 5; js++) {
                  const double ir = interp_linear_disc(i_lo, i_hi, 5, is);
                  const double jr = interp_linear_disc(j_lo, j_hi, 5, js);
                  f_max = fmax(f_max, coag_num_conc_factor(cx, ir, jr, c1, c2, cc));
                }
            }
            k_max = ktab[(size_t)(my_b2 - 1) * nb + (b1 - 1)] * f_max;
            // determine_target_and_source + compute_n_source (coagulation.F90:247-329)
            if (mono_dec) {
              int bt = 0, bs = 0, ct = 0, cs = 0;
              if (accel_target_source(cx, pair_flat, hm, b1, my_b2, c1, c2, cc, bt, bs, ct, cs)) {
                const double n_source_per_target = k_max * del_t * (double)((bs == b1 && cs == c1) ? n1 : n2);
                if (!(n_source_per_target < COAG_ACCEL_N_EVENT)) {
                  const double min_target_vol = CX_RAD2VOL(grid_edge(cx.gmin, cx.gmax, nb, bt));
                  const double max_source_vol = CX_RAD2VOL(grid_edge(cx.gmin, cx.gmax, nb, bs + 1));
                  const double growth = (min_target_vol + max_source_vol * n_source_per_target) / min_target_vol;
                  if (!(growth > MAX_ALLOWABLE_GROWTH_FACTOR)) accel = 1;
                }
              }
            }
            if (!accel) {
              // compute_n_samp (coagulation.F90:648-685)
              const bool same = (b1 == my_b2) && (c1 == c2);
              const double n_possible = same ? (double)n1 * ((double)n2 - 1.0) / 2.0 : (double)n1 * (double)n2;
              const double n_samp_mean = (k_max * del_t) * n_possible;
              const unsigned pid = ((unsigned)((c1 - 1) * v.C + (c2 - 1)) * 256u + (unsigned)b1) * 256u + (unsigned)my_b2;
              RStream s0{RKey{v.seed, (unsigned)v.gcell[cell], step, PH_COAG, pid, 0u}, 0u};
              n_samp = rand_poisson(n_samp_mean, s0);
            }
          }
          unsigned has = __ballot_sync(0xffffffffu, (n_samp > 0) || accel);
          bool rescan = false;
          while (has && !rescan && !cx.err) {
            const int f = __ffs(has) - 1;
            has &= has - 1;
            const int pb2 = __shfl_sync(0xffffffffu, my_b2, f);
            const int p_c1 = __shfl_sync(0xffffffffu, c1, f), p_c2 = __shfl_sync(0xffffffffu, c2, f);
            const int p_b1 = __shfl_sync(0xffffffffu, b1, f);
            const int p_nsamp = __shfl_sync(0xffffffffu, n_samp, f);
            const int p_cc = __shfl_sync(0xffffffffu, cc, f);
            const double p_kmax = __shfl_sync(0xffffffffu, k_max, f);
            const int p_accel = __shfl_sync(0xffffffffu, accel, f);
            if (p_accel) {
              const unsigned apid = ((unsigned)((p_c1 - 1) * v.C + (p_c2 - 1)) * 256u + (unsigned)p_b1) * 256u + (unsigned)pb2;
              const bool pflat = all_flat || (flat[p_c1 - 1] && flat[p_c2 - 1]);
              int bt = 0, bs = 0, ct = 0, cs = 0;
              accel_target_source(cx, pflat, hm, p_b1, pb2, p_c1, p_c2, p_cc, bt, bs, ct, cs);
              cell_accel_pair<KT, COLD>(cx, ktype, be, p_kmax, del_t, step, apid, bt, bs, ct, cs, p_cc, tot_n_samp, tot_n_coag);
              rescan = true;  // counts changed: pairs behind this one must be rescanned
              cu_c1 = p_c1; cu_c2 = p_c2; cu_b1 = p_b1; cu_b2 = pb2;
              more = pair_advance_exact(v.C, nb, cu_c1, cu_c2, cu_b1, cu_b2);
              continue;
            }
            // ---- per_set_coag (coagulation.F90:587-643)
            tot_n_samp += p_nsamp;
            const bool same = (p_b1 == pb2) && (p_c1 == p_c2);
            const double accept_factor = 1.0 / p_kmax;
            const unsigned pid = ((unsigned)((p_c1 - 1) * v.C + (p_c2 - 1)) * 256u + (unsigned)p_b1) * 256u + (unsigned)pb2;
            const int k1 = lkey(cx, p_c1, p_b1), k2 = lkey(cx, p_c2, pb2);
            int d = 1;
            while (d <= p_nsamp && !cx.err) {
              const int m1 = cx.llen[k1], m2 = cx.llen[k2];
              if (((m1 < 2) && same) || (m1 < 1) || (m2 < 1)) break;
              const int my_d = d + lane;
              bool accept = false;
              int p1 = 0, p2 = 0;
              if (my_d <= p_nsamp) {
                const RKey key{v.seed, (unsigned)v.gcell[cell], step, PH_COAG, pid, (unsigned)my_d};
                // find_rand_pair (coagulation.F90:766-803): draws 0, 1
                const int i1 = pmc_rand_int(rng_uniform(key, 0), m1);
                int i2;
                if (same) {
                  i2 = pmc_rand_int(rng_uniform(key, 1), m2 - 1);
                  if (i2 == i1) i2 = m2;
                } else {
                  i2 = pmc_rand_int(rng_uniform(key, 1), m2);
                }
                p1 = cx.L[cx.lptr[k1] + i1 - 1];
                p2 = cx.L[cx.lptr[k2] + i2 - 1];
                // num_conc_weighted_kernel (coag_kernel.F90:175-209)
                double v1, ms1, v2, ms2;
                part_vol_mass(cx, p1, v1, ms1);
                part_vol_mass(cx, p2, v2, ms2);
                const double uk = UNWEIGHTED_KERNEL(ktype, be, v1, ms1, v2, ms2);
                const double k = uk * coag_num_conc_factor(cx, CX_VOL2RAD(v1), CX_VOL2RAD(v2), p_c1, p_c2, p_cc);
                accept = rng_uniform(key, 2) < k * accept_factor;  // maybe_coag_pair :753
              }
              const unsigned acc = __ballot_sync(0xffffffffu, accept);
              if (acc == 0) {
                d += 32;
                continue;
              }
              const int fa = __ffs(acc) - 1;
              const int e1 = __shfl_sync(0xffffffffu, p1, fa);
              const int e2 = __shfl_sync(0xffffffffu, p2, fa);
              const RKey ekey{v.seed, (unsigned)v.gcell[cell], step, PH_COAG, pid, (unsigned)(d + fa)};
              cell_coagulate<COLD>(cx, e1, e2, p_c1, p_c2, p_cc, ekey);
              tot_n_coag += 1;
              rescan = true;  // counts changed: pairs behind this one must be rescanned
              d = d + fa + 1;
            }
            if (rescan) {  // resume right behind the pair that changed the counts
              cu_c1 = p_c1; cu_c2 = p_c2; cu_b1 = p_b1; cu_b2 = pb2;
              more = pair_advance_exact(v.C, nb, cu_c1, cu_c2, cu_b1, cu_b2);
            }
          }
          if (!rescan) more = pair_advance(cx, v.C, nb, cu_c1, cu_c2, cu_b1, cu_b2, 32);
        }
      }
    }
  }
  __syncwarp();
  // publish the per-(class, bin) counts and the population size
  {
    int* bc = v.bin_count + (size_t)cell * v.C * v.nbcap;
    for (int k = lane; k < v.C * v.nbcap; k += 32) bc[k] = cx.llen[k];
  }
  if (lane == 0) {
    v.count[cell] = cx.n_part;
    v.coag_samp[cell] = tot_n_samp;
    v.coag_ncoag[cell] = tot_n_coag;
    if (cx.err && v.status[cell] == 0) v.status[cell] = cx.err;
  }
}

}  // namespace PMC_COAG_NS
using namespace PMC_COAG_NS;

int pmc_kernel_bounds_async(pmcgpu_ctx* ctx, int ktype);

#ifdef PMC_COAG_FRACTAL
// mc_coag for fractal particles and / or the brown_free and brown_cont kernels (called by pmc_mc_coag_async)
int pmc_mc_coag_fractal_async(pmcgpu_ctx* ctx, int ktype, double del_t, unsigned step) {
  DevView& v = ctx->v;
  int rc = pmc_kernel_bounds_async(ctx, ktype);  // rejects the kernels that need the mobility radius
  if (rc) return rc;
  const size_t smem = (size_t)COAG_WARPS * v.C * (sizeof(double) + sizeof(int));
  const int blocks = (v.n_cells + COAG_WARPS - 1) / COAG_WARPS;
  ProfScope ps_(ctx, "k_mc_coag");
  if (v.cold.n)
    k_mc_coag<0, true, COAG_MIN_CTAS><<<blocks, COAG_WARPS * 32, smem, ctx->stream>>>(v, ktype, del_t, step);
  else
    k_mc_coag<0, false, COAG_MIN_CTAS><<<blocks, COAG_WARPS * 32, smem, ctx->stream>>>(v, ktype, del_t, step);
  LAUNCH_CHECK(ctx);
  return PMCGPU_OK;
}
#else
int pmc_mc_coag_fractal_async(pmcgpu_ctx* ctx, int ktype, double del_t, unsigned step);
int pmc_mc_coag_async(pmcgpu_ctx* ctx, int ktype, double del_t, unsigned step) {
  DevView& v = ctx->v;
  if (v.fr_on || ktype == PMCGPU_KERNEL_BROWN_FREE || ktype == PMCGPU_KERNEL_BROWN_CONT)
    return pmc_mc_coag_fractal_async(ctx, ktype, del_t, step);
  // coagulation.F90:64-70: bounds are (re)computed only when the grid was remade
  int rc = pmc_kernel_bounds_async(ctx, ktype);
  if (rc) return rc;
  const size_t smem = (size_t)COAG_WARPS * v.C * (sizeof(double) + sizeof(int));
  const int blocks = (v.n_cells + COAG_WARPS - 1) / COAG_WARPS;
  {
    ProfScope ps_(ctx, "k_mc_coag");
    // Resident CTAs per SM (= registers per thread) of the Brownian instance: the variant with the MOST registers whose
    // single wave still holds every CTA (8 CTAs per SM at 64 registers, 12 at 40, 16 at 32); 16 when nothing fits in
    // one wave.  Measured ms per K4-type step, cells = 4000 / 6000 / 8000 / 16 000 / 32 000:
    //   8 CTAs: 2.94 / 4.45 / 5.34 / 10.18 / 19.7;  12 CTAs: 3.21 / 3.92 / 5.51 / 9.55 / 17.9;  16 CTAs: 3.41 / 4.17 /
    //   4.90 / 9.02 / 17.3  (fewer spills against more warps to hide the instruction fetch, and wave quantisation)
    const int n_sm = 148;
    if (v.cold.n)  // component lists are tracked (pmcgpu_track_components): the instance that merges them
      k_mc_coag<0, true, COAG_MIN_CTAS><<<blocks, COAG_WARPS * 32, smem, ctx->stream>>>(v, ktype, del_t, step);
    else if (ktype == PMCGPU_KERNEL_BROWN && blocks <= n_sm * 8)
      k_mc_coag<PMCGPU_KERNEL_BROWN, false, 8><<<blocks, COAG_WARPS * 32, smem, ctx->stream>>>(v, ktype, del_t, step);
    else if (ktype == PMCGPU_KERNEL_BROWN && blocks > n_sm * 12)
      k_mc_coag<PMCGPU_KERNEL_BROWN, false, 16><<<blocks, COAG_WARPS * 32, smem, ctx->stream>>>(v, ktype, del_t, step);
    else if (ktype == PMCGPU_KERNEL_BROWN)
      k_mc_coag<PMCGPU_KERNEL_BROWN, false, COAG_MIN_CTAS><<<blocks, COAG_WARPS * 32, smem, ctx->stream>>>(v, ktype, del_t, step);
    else
      k_mc_coag<0, false, COAG_MIN_CTAS><<<blocks, COAG_WARPS * 32, smem, ctx->stream>>>(v, ktype, del_t, step);
    LAUNCH_CHECK(ctx);
  }
  return PMCGPU_OK;
}

int pmc_sort_async(pmcgpu_ctx* ctx, int edge_recheck);

extern "C" int pmcgpu_mc_coag(pmcgpu_ctx* ctx, int coag_kernel_type, double del_t, unsigned int step,
                              int* tot_n_samp, int* tot_n_coag) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  // mc_coag starts with aero_state_sort (coagulation.F90:63).  The device state is
  // unsorted between steps, so this is a full sort (valid_sort = .false.) followed by
  // the bin-edge safety test the reference applies to an already sorted state.
  int rc = pmc_sort_async(ctx, 1);
  if (rc) return rc;
  rc = pmc_mc_coag_async(ctx, coag_kernel_type, del_t, step);
  if (rc) return rc;
  if (tot_n_samp)
    CUDA_OK(ctx, cudaMemcpyAsync(tot_n_samp, ctx->v.coag_samp, ctx->v.n_cells * sizeof(int), cudaMemcpyDeviceToHost,
                                 ctx->stream));
  if (tot_n_coag)
    CUDA_OK(ctx, cudaMemcpyAsync(tot_n_coag, ctx->v.coag_ncoag, ctx->v.n_cells * sizeof(int), cudaMemcpyDeviceToHost,
                                 ctx->stream));
  return pmc_first_status(ctx);
}
#endif  // PMC_COAG_FRACTAL
