// trans.cu -- wrf_pmc_trans_driver (interface/wrf_pmc_trans.F90:37-200) and trans_aero
// (interface/wrf_pmc_trans_aero.F90:40-971) for the doubly periodic patch of this rank.
//
// Reference flow per WRF step                                  here
//   compute_advect_probs_wrf / compute_diffusion_probs          k_tr_counts, k_tr_hprobs   (thread per cell, class)
//   compute_vertical_probs (B^N A per column) + normalize_probs k_tr_vprobs               (CTA per column)
//   C1 ghost cell_volume                                        edge pack / exchange / unpack
//   trans_aero_preweight: compute_fluxes scatter, C2, scale     k_tr_edge_flux + exchange, k_tr_preweight (gather
//                                                               in the scatter's summation order)
//   C3 ghost weights                                            edge pack / exchange / unpack
//   aero_state_multisample -> sample_particles_by_class         k_tr_multisample (warp per source cell: draws in
//                                                               parallel, swap-with-last replay on indices in
//                                                               shared memory, records do not move yet)
//   C4 ghost particles (MPI_Pack)                               k_tr_pack_* -> SoA slabs -> exchange (NCCL send/recv
//                                                               between ranks, device copy when the neighbour is
//                                                               this rank itself)
//   add + rebalance + sort                                      k_tr_stage: the sampled records of every cell to the
//                                                               alternate slab; k_tr_gather: survivors stay in
//                                                               place (holes filled from the tail), every
//                                                               destination cell appends its arrivals in the
//                                                               reference's arrival order (no atomics); then
//                                                               k_rebalance and the bin-grid decision of
//                                                               aero_state_sort
// Sides: 0 = y-, 1 = y+, 2 = x-, 3 = x+ (the order of get_array_index_for_send_recv_from_ghost, :1826-1933).
// Direction slots of a cell: 0 -x, 1 +x, 2 -y, 3 +y, 4 + k for the vertical level k (0-based).
#include "pmcgpu_internal.cuh"
#include <algorithm>
#include <cstdio>
#include <cstring>

#define TR_ND(nz) (4 + (nz))

// ------------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int tr_ext(const TransView& t, int i, int k, int j) {  // i in [-1, nx], j in [-1, ny]
  return (i + 1) + (t.nx + 2) * (k + t.nz * (j + 1));
}
__device__ __forceinline__ int tr_cell(const TransView& t, int i, int k, int j) { return i + t.nx * (k + t.nz * j); }
__device__ __forceinline__ double tr_hprob(const TransView& t, int cell, int c, int q) {
  return t.padv[((size_t)cell * t.C + c) * 6 + q] + t.pdiff[((size_t)cell * t.C + c) * 4 + q];
}

// open domain (compute_fluxes "real case", wrf_pmc_trans_aero.F90:1517-1623; boundary inflow :689-766): a west-edge
// cell only feeds its east neighbour (slot 1), east -> slot 0, south -> slot 3, north -> slot 2, a corner nobody
__device__ __forceinline__ int tr_kind(const DevView& v, int cell) { return v.kind ? (int)v.kind[cell] : 0; }
__device__ __forceinline__ bool tr_allowed(int kind, int slot) {
  return kind == 0 || (kind == 1 && slot == 1) || (kind == 2 && slot == 0) || (kind == 3 && slot == 3) ||
         (kind == 4 && slot == 2);
}

// ---- particles of (group 1, class c) per cell: aero_state_total_particles(aero_state, 1, i_class)
__global__ void k_tr_counts(DevView v, TransView t) {
  const int lane = threadIdx.x & 31;
  const int cell = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (cell >= v.n_cells) return;
  const size_t base = (size_t)cell * v.cap;
  const int n = v.count[cell];
  for (int c = 0; c < v.C; c++) {
    int m = 0;
    for (int p = lane; p < n; p += 32) m += (v.wgc[base + p] == c * v.G);
    for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
    if (lane == 0) t.ccount[(size_t)cell * v.C + c] = m;
  }
}

// ---- compute_advect_probs_wrf (wrf_pmc_trans.F90:205-286) + compute_diffusion_probs (:358-409)
__global__ void k_tr_hprobs(DevView v, TransView t) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= v.n_cells * v.C) return;
  const int cell = q / v.C, c = q % v.C;
  const int i = cell % t.nx, k = (cell / t.nx) % t.nz, j = cell / (t.nx * t.nz);
  const size_t ah = (size_t)(i + 1) + (size_t)(t.nx + 2) * (k + (size_t)t.nz * (j + 1));
  const double rho_mass = 1.0 / (double)t.alt[ah];
  const double mag = v.wmag[((size_t)cell * v.C + c) * v.G];
  const int n_parts = t.ccount[(size_t)cell * v.C + c];
  t.num_old[(size_t)cell * v.C + c] = (double)n_parts * mag;  // trans_aero_preweight :1213
  const double mix_rat = (double)n_parts * mag / rho_mass;
  const size_t usz = (size_t)(t.nx + 1) * t.nz * t.ny, vsz = (size_t)t.nx * t.nz * (t.ny + 1),
               wsz = (size_t)t.nx * (t.nz + 1) * t.ny;
  const float uf0 = t.u_flux[c * usz + (size_t)i + (size_t)(t.nx + 1) * (k + (size_t)t.nz * j)];
  const float uf1 = t.u_flux[c * usz + (size_t)(i + 1) + (size_t)(t.nx + 1) * (k + (size_t)t.nz * j)];
  const float vf0 = t.v_flux[c * vsz + (size_t)i + (size_t)t.nx * (k + (size_t)t.nz * j)];
  const float vf1 = t.v_flux[c * vsz + (size_t)i + (size_t)t.nx * (k + (size_t)t.nz * (j + 1))];
  const float wf0 = t.w_flux[c * wsz + (size_t)i + (size_t)t.nx * (k + (size_t)(t.nz + 1) * j)];
  const float wf1 = t.w_flux[c * wsz + (size_t)i + (size_t)t.nx * ((k + 1) + (size_t)(t.nz + 1) * j)];
  double a[6] = {0, 0, 0, 0, 0, 0};
  if (uf1 > 0.0f && mix_rat > 0.0) a[1] = (double)uf1 / mix_rat;
  if (uf0 < 0.0f && mix_rat > 0.0) a[0] = -1.0 * (double)uf0 / mix_rat;
  if (vf1 > 0.0f && mix_rat > 0.0) a[3] = (double)vf1 / mix_rat;
  if (vf0 < 0.0f && mix_rat > 0.0) a[2] = -1.0 * (double)vf0 / mix_rat;
  if (wf1 > 0.0f && mix_rat > 0.0) a[5] = fmin((double)wf1 / mix_rat, 0.95);
  if (wf0 < 0.0f && mix_rat > 0.0) a[4] = fmin(-1.0 * (double)wf0 / mix_rat, 0.95);
  const double prob_sum = ((((a[2] + a[4]) + a[0]) + a[1]) + a[5]) + a[3];  // array element order
  if (prob_sum > 1.0)
    for (int m = 0; m < 6; m++) a[m] = a[m] / (prob_sum + 1.0e-8);
  for (int m = 0; m < 6; m++) t.padv[((size_t)cell * v.C + c) * 6 + m] = a[m];
  // diffusion
  const double dx = (double)t.dx, dy = (double)t.dy, dt = (double)t.dt;
  const int di[4] = {-1, 1, 0, 0}, dj[4] = {0, 0, -1, 1};
  for (int m = 0; m < 4; m++) {
    const size_t an = (size_t)(i + di[m] + 1) + (size_t)(t.nx + 2) * (k + (size_t)t.nz * (j + dj[m] + 1));
    const float ksum = t.xkhh[ah] + t.xkhh[an];
    const double diff_coef = fmax(0.5 * (double)ksum, 0.0);
    const double rho_edge = 0.5 * (rho_mass + (1.0 / (double)t.alt[an]));
    const double dd = (m < 2) ? dx : dy;
    t.pdiff[((size_t)cell * v.C + c) * 4 + m] = (dt * diff_coef * rho_edge) / (dd * dd * rho_mass);
  }
}

// ---- compute_vertical_probs (:414-562) + stable_timestep (:626-652) + normalize_probs (:692-773), CTA per column
__global__ void __launch_bounds__(256) k_tr_vprobs(DevView v, TransView t) {
  const int col = blockIdx.x;
  const int i = col % t.nx, j = col / t.nx;
  const int nz = t.nz, tid = threadIdx.x;
  extern __shared__ double s_m[];
  double* B = s_m;              // [nz][nz]
  double* Bt = B + nz * nz;
  double* R = Bt + nz * nz;
  double* A = R + nz * nz;
  double* z_mass = A + nz * nz;  // nz
  double* rho_mass = z_mass + nz;
  double* z_w = rho_mass + nz;   // nz + 1
  double* k_w = z_w + nz + 1;
  double* rho_w = k_w + nz + 1;
  __shared__ double s_small_dt;
  __shared__ int s_num_steps;
  auto cidx = [&](int k) { return (size_t)i + (size_t)t.nx * (k + (size_t)nz * j); };
  auto widx = [&](int k) { return (size_t)i + (size_t)t.nx * (k + (size_t)(nz + 1) * j); };
  const bool bcol = tr_kind(v, tr_cell(t, i, 0, j)) != 0;
  for (int k = tid; k < nz; k += blockDim.x) {
    z_mass[k] = (double)t.zc[cidx(k)];
    rho_mass[k] = 1.0 / (double)t.alt[(size_t)(i + 1) + (size_t)(t.nx + 2) * (k + (size_t)nz * (j + 1))];
  }
  for (int k = tid; k <= nz; k += blockDim.x) {
    z_w[k] = (double)t.z_at_w[widx(k)];
    k_w[k] = 0.0;
    rho_w[k] = 0.0;
  }
  for (int q = tid; q < nz * nz; q += blockDim.x) B[q] = 0.0;
  if (tid == 0) {
    const double dt = (double)t.dt;
    const double min_dz = (double)(t.zc[cidx(1)] - t.zc[cidx(0)]);
    float mx = t.exch_h[widx(0)];
    for (int k = 1; k <= nz; k++) mx = fmaxf(mx, t.exch_h[widx(k)]);
    const double max_k = fmax((double)mx, 1.0e-6);
    const double max_time = 0.1 * (min_dz * min_dz) / max_k;
    const int ns = (int)ceil(dt / max_time);
    s_num_steps = ns;
    s_small_dt = dt / (double)ns;
  }
  __syncthreads();
  for (int k = tid; k < nz - 1; k += blockDim.x) k_w[k] = fmax(1.0e-6, (double)t.exch_h[widx(k + 1)]);
  for (int k = tid; k < nz - 1; k += blockDim.x) {
    const double tmp = z_mass[k + 1] - z_mass[k];
    rho_w[k] = ((z_mass[k + 1] - z_w[k + 1]) / tmp) * rho_mass[k] + ((z_w[k + 1] - z_mass[k]) / tmp) * rho_mass[k + 1];
  }
  __syncthreads();
  const double small_dt = s_small_dt;
  const int num_steps = s_num_steps;
  // rows r (0-based); Fortran row k = r + 1
  for (int r = tid; r < nz; r += blockDim.x) {
    if (r >= 1 && r <= nz - 2) {
      const double dn = (small_dt / ((z_w[r + 1] - z_w[r]) * (z_mass[r] - z_mass[r - 1]))) * (rho_w[r - 1] / rho_mass[r]) * k_w[r - 1];
      const double up = (small_dt / ((z_w[r + 1] - z_w[r]) * (z_mass[r + 1] - z_mass[r]))) * (rho_w[r] / rho_mass[r]) * k_w[r];
      B[r * nz + r - 1] = dn;
      B[r * nz + r + 1] = up;
      B[r * nz + r] = 1.0 - dn - up;
    } else if (r == 0) {
      const double up = (small_dt / ((z_w[1] - z_w[0]) * (z_mass[1] - z_mass[0]))) * (rho_w[0] / rho_mass[0]) * k_w[0];
      B[1] = up;
      B[0] = 1.0 - up;
    } else {
      const double dn = (small_dt / ((z_w[nz] - z_w[nz - 1]) * (z_mass[nz - 1] - z_mass[nz - 2]))) *
                        (rho_w[nz - 2] / rho_mass[nz - 1]) * k_w[nz - 2];
      B[(nz - 1) * nz + nz - 2] = dn;
      B[(nz - 1) * nz + nz - 1] = 1.0 - dn;
    }
  }
  __syncthreads();
  for (int q = tid; q < nz * nz; q += blockDim.x) Bt[q] = B[q];
  __syncthreads();
  for (int s = 2; s <= num_steps; s++) {
    for (int q = tid; q < nz * nz; q += blockDim.x) {
      const int r = q / nz, c = q % nz;
      double acc = 0.0;
      for (int m = 0; m < nz; m++) acc += Bt[r * nz + m] * B[m * nz + c];
      R[q] = acc;
    }
    __syncthreads();
    for (int r = tid; r < nz; r += blockDim.x) {
      double rs = 0.0;
      for (int c = 0; c < nz; c++) rs += R[r * nz + c];
      for (int c = 0; c < nz; c++) Bt[r * nz + c] = R[r * nz + c] / rs;
    }
    __syncthreads();
  }
  for (int cl = 0; cl < v.C; cl++) {
    for (int q = tid; q < nz * nz; q += blockDim.x) A[q] = 0.0;
    __syncthreads();
    for (int r = tid; r < nz; r += blockDim.x) {
      const size_t pb = ((size_t)tr_cell(t, i, r, j) * v.C + cl) * 6;
      const double up = t.padv[pb + 5], dn = t.padv[pb + 4];
      if (r >= 1 && r <= nz - 2) {
        A[r * nz + r] = 1.0 - (up + dn);
        A[r * nz + r - 1] = dn;
        A[r * nz + r + 1] = up;
      } else if (r == 0) {
        A[0] = 1.0 - up;
        A[1] = up;
      } else {
        A[(nz - 1) * nz + nz - 1] = 1.0 - dn;
        A[(nz - 1) * nz + nz - 2] = dn;
      }
    }
    __syncthreads();
    for (int q = tid; q < nz * nz; q += blockDim.x) {
      const int r = q / nz, c = q % nz;
      double val;
      if (t.vertmix > 0) {
        double acc = 0.0;
        for (int m = 0; m < nz; m++) acc += Bt[r * nz + m] * A[m * nz + c];
        val = acc;
      } else {
        val = A[q];
      }
      // prob_vert_diffusion stays 0 in the boundary columns of an open domain (wrf_pmc_trans.F90:104, 116-136)
      t.pvert[((size_t)tr_cell(t, i, r, j) * v.C + cl) * nz + c] = bcol ? 0.0 : val;
    }
    __syncthreads();
    // normalize_probs for (k = r, class cl)
    for (int r = tid; r < nz; r += blockDim.x) {
      const int cell = tr_cell(t, i, r, j);
      double* a = t.padv + ((size_t)cell * v.C + cl) * 6;
      double* f = t.pdiff + ((size_t)cell * v.C + cl) * 4;
      double* pv = t.pvert + ((size_t)cell * v.C + cl) * nz;
      bool ok = true;
      for (int m = 0; m < 6; m++) ok = ok && (a[m] >= 0.0);
      for (int m = 0; m < 4; m++) ok = ok && (f[m] >= 0.0);
      for (int m = 0; m < nz; m++) ok = ok && (pv[m] >= 0.0);
      double s = ((a[2] + a[0]) + a[1]) + a[3];
      s = s + (((f[2] + f[0]) + f[1]) + f[3]);
      for (int m = 0; m < nz; m++)
        if (m != r) s = s + pv[m];
      if (s >= 1.0) {
        for (int m = 0; m < nz; m++) pv[m] = pv[m] / s;
        for (int m = 0; m < 4; m++) f[m] = f[m] / s;
        for (int m = 0; m < 4; m++) a[m] = a[m] / s;
      }
      for (int m = 0; m < 6; m++) ok = ok && (a[m] >= 0.0) && (a[m] <= 1.0);
      for (int m = 0; m < 4; m++) ok = ok && (f[m] >= 0.0) && (f[m] <= 1.0);
      for (int m = 0; m < nz; m++) ok = ok && (pv[m] >= 0.0) && (pv[m] <= 1.0);
      double s2 = ((a[2] + a[0]) + a[1]) + a[3];
      s2 = s2 + (((f[2] + f[0]) + f[1]) + f[3]);
      for (int m = 0; m < nz; m++)
        if (m != r) s2 = s2 + pv[m];
      ok = ok && (s2 < 1.0 + 1.0e-8);
      if (!ok && v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_PROB;
    }
    __syncthreads();
  }
}

// ---- edge <-> ghost plumbing for per-cell fields with a ghost ring ("ext" arrays, ncomp doubles per cell)
// pack the edge row / column next to side s of an ext array
__global__ void k_tr_pack_edge(TransView t, const double* __restrict__ ext, int ncomp, int side, double* __restrict__ out) {
  const int L = (side < 2) ? t.nx : t.ny;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= L * t.nz * ncomp) return;
  const int m = q % ncomp, k = (q / ncomp) % t.nz, e = q / (ncomp * t.nz);
  int i, j;
  if (side == 0) { i = e; j = 0; } else if (side == 1) { i = e; j = t.ny - 1; }
  else if (side == 2) { i = 0; j = e; } else { i = t.nx - 1; j = e; }
  out[q] = ext[(size_t)tr_ext(t, i, k, j) * ncomp + m];
}
// store what arrived from the neighbour on side s into the ghost row / column of side s
__global__ void k_tr_unpack_ghost(TransView t, double* __restrict__ ext, int ncomp, int side, const double* __restrict__ in) {
  const int L = (side < 2) ? t.nx : t.ny;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= L * t.nz * ncomp) return;
  const int m = q % ncomp, k = (q / ncomp) % t.nz, e = q / (ncomp * t.nz);
  int i, j;
  if (side == 0) { i = e; j = -1; } else if (side == 1) { i = e; j = t.ny; }
  else if (side == 2) { i = -1; j = e; } else { i = t.nx; j = e; }
  ext[(size_t)tr_ext(t, i, k, j) * ncomp + m] = in[q];
}
__global__ void k_tr_fill_ext(DevView v, TransView t, int what) {  // 0: cell volume, 1: weight magnitudes
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= v.n_cells) return;
  const int i = cell % t.nx, k = (cell / t.nx) % t.nz, j = cell / (t.nx * t.nz);
  const int e = tr_ext(t, i, k, j);
  if (what == 0) {
    t.cellvol_ext[e] = v.cellvol[cell];
  } else {
    for (int c = 0; c < v.C; c++) t.wmag_ext[(size_t)e * v.C + c] = v.wmag[((size_t)cell * v.C + c) * v.G];
  }
}

// ---- compute_fluxes (:1428-1624) towards the ghost ring: the value the neighbour adds to its edge cell (C2)
__global__ void k_tr_edge_flux(DevView v, TransView t, int side, double* __restrict__ out) {
  const int L = (side < 2) ? t.nx : t.ny;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= L * t.nz * v.C) return;
  const int c = q % v.C, k = (q / v.C) % t.nz, e = q / (v.C * t.nz);
  int i, j, gi, gj, slot;
  if (side == 0) { i = e; j = 0; gi = i; gj = -1; slot = 2; }
  else if (side == 1) { i = e; j = t.ny - 1; gi = i; gj = t.ny; slot = 3; }
  else if (side == 2) { i = 0; j = e; gi = -1; gj = j; slot = 0; }
  else { i = t.nx - 1; j = e; gi = t.nx; gj = j; slot = 1; }
  const int cell = tr_cell(t, i, k, j);
  const double source_vol = t.cellvol_ext[tr_ext(t, i, k, j)], dest_vol = t.cellvol_ext[tr_ext(t, gi, k, gj)];
  const double flux = tr_hprob(t, cell, c, slot) * t.num_old[(size_t)cell * v.C + c];
  out[q] = tr_allowed(tr_kind(v, cell), slot) ? flux * (dest_vol / source_vol) : 0.0;
}

// ---- trans_aero_preweight (:1103-1423): expected number concentration of every (cell, class) gathered in
// the summation order of the reference's scatter loop, then the weight adjustment decision (:1373-1400)
__global__ void k_tr_preweight(DevView v, TransView t) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= v.n_cells * v.C) return;
  const int cell = q / v.C, c = q % v.C;
  const int nx = t.nx, nz = t.nz, ny = t.ny;
  const int i = cell % nx, k = (cell / nx) % nz, j = cell / (nx * nz);
  const double volD = t.cellvol_ext[tr_ext(t, i, k, j)];
  const double ncoD = t.num_old[q];
  double acc = ncoD;
  const int kindD = tr_kind(v, cell);
  auto from_h = [&](int si, int sj, int slot) {
    const int s = tr_cell(t, si, k, sj);
    if (!tr_allowed(tr_kind(v, s), slot)) return;
    const double flux = tr_hprob(t, s, c, slot) * t.num_old[(size_t)s * v.C + c];
    acc = acc + flux * (volD / t.cellvol_ext[tr_ext(t, si, k, sj)]);
  };
  auto from_v = [&](int sk) {
    const int s = tr_cell(t, i, sk, j);
    const double flux = t.pvert[((size_t)s * v.C + c) * nz + k] * t.num_old[(size_t)s * v.C + c];
    acc = acc + flux * (volD / t.cellvol_ext[tr_ext(t, i, sk, j)]);
  };
  if (j - 1 >= 0) from_h(i, j - 1, 3);
  if (kindD == 0)
    for (int sk = 0; sk < k; sk++) from_v(sk);
  if (i - 1 >= 0) from_h(i - 1, j, 1);
  // the cell's own losses, in the order of compute_fluxes: -y, +y, -x, +x, then the vertical levels (a boundary
  // cell of an open domain only loses towards its interior neighbour, a corner nothing)
  if (tr_allowed(kindD, 2)) acc = acc - tr_hprob(t, cell, c, 2) * ncoD;
  if (tr_allowed(kindD, 3)) acc = acc - tr_hprob(t, cell, c, 3) * ncoD;
  if (tr_allowed(kindD, 0)) acc = acc - tr_hprob(t, cell, c, 0) * ncoD;
  if (tr_allowed(kindD, 1)) acc = acc - tr_hprob(t, cell, c, 1) * ncoD;
  if (kindD == 0)
    for (int kk = 0; kk < nz; kk++)
      if (kk != k) acc = acc - t.pvert[((size_t)cell * v.C + c) * nz + kk] * ncoD;
  if (i + 1 <= nx - 1) from_h(i + 1, j, 0);
  if (kindD == 0)
    for (int sk = k + 1; sk < nz; sk++) from_v(sk);
  if (j + 1 <= ny - 1) from_h(i, j + 1, 2);
  // ghost sums of the neighbours, in the order the reference receives them: sender side y-, y+, x-, x+
  if (j == ny - 1) acc = acc + t.rbuf[1][((size_t)i * nz + k) * v.C + c];  // neighbour y+ sent its y- ghost row
  if (j == 0) acc = acc + t.rbuf[0][((size_t)i * nz + k) * v.C + c];       // neighbour y- sent its y+ ghost row
  if (i == nx - 1) acc = acc + t.rbuf[3][((size_t)j * nz + k) * v.C + c];
  if (i == 0) acc = acc + t.rbuf[2][((size_t)j * nz + k) * v.C + c];
  const size_t wi = ((size_t)cell * v.C + c) * v.G;
  const double mag = v.wmag[wi];
  const double n_parts_new = acc / mag;
  double ratio = 0.0, new_mag = mag;
  const double ideal = v.ideal_gc[c * v.G];  // group 1
  if (n_parts_new > 2.0 * ideal) {
    ratio = n_parts_new / ideal;  // weight_ratio; aero_state_scale_weight, ratio > 1 branch
    new_mag = mag * ratio;
    v.wmag[wi] = new_mag;
  }
  t.scale_ratio[q] = ratio;
  t.wmag_ext[(size_t)tr_ext(t, i, k, j) * v.C + c] = new_mag;
}

// ---- aero_state_multisample (:2351-2482) + aero_state_sample_particles_by_class (:2181-2322) and the
// removals of aero_state_scale_weight (aero_state.F90:1832-1895), one warp per source cell.
// Index state in shared memory (cap ints each): P2O current particle position -> original slot,
// LST class lists (positions; class c at lst_off[c]), FXC position -> index in its list | class << 24.
struct TrIdx {
  int* P2O; int* LST; int* FXC;
  int* len;      // [C] list lengths
  int* off;      // [C] list offsets
  int n_cur;
};
// aero_particle_array swap-with-last + integer_rmap2_remove (integer_rmap2.F90:190-243) on group_class
__device__ __forceinline__ void tr_remove(TrIdx& x, int p) {
  const int last = x.n_cur - 1;
  const int fo = x.FXC[p];
  const int c_old = fo >> 24, pos_old = fo & 0xffffff;
  if (last != p) {
    const int ff = x.FXC[last];
    x.LST[x.off[ff >> 24] + (ff & 0xffffff)] = p;
    x.FXC[p] = ff;
    x.P2O[p] = x.P2O[last];
  }
  const int lp = x.off[c_old];
  const int pos_last = x.len[c_old] - 1;
  const int dom_fix = x.LST[lp + pos_last];
  if (pos_old < pos_last) x.LST[lp + pos_old] = dom_fix;
  x.len[c_old] = pos_last;
  if (pos_last != pos_old) x.FXC[dom_fix] = (c_old << 24) | pos_old;
  x.n_cur = last;
}

#define TR_CHUNK 32
// The index state of a cell is sized by lcap (>= its population), not by the slab capacity: the launch for
// the cells with n_lo < n <= n_hi = lcap keeps several CTAs per SM, a second launch (lcap = capacity) takes the
// few larger cells.
__global__ void k_tr_multisample(DevView v, TransView t, unsigned step, int warps_per_cta, int lcap, int n_lo,
                                 int n_hi) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * warps_per_cta + warp;
  if (cell >= v.n_cells) return;
  {
    const int nn = v.count[cell];
    if (nn <= n_lo || nn > n_hi) return;
  }
  extern __shared__ int s_i[];
  const int nd = TR_ND(t.nz);
  const int per_warp = (3 * lcap + 2 * v.C + 3 * TR_CHUNK + 3 * nd + 1) & ~1;
  int* sm = s_i + (size_t)warp * per_warp;
  TrIdx x;
  x.P2O = sm; x.LST = sm + lcap; x.FXC = sm + 2 * lcap;
  x.len = sm + 3 * lcap; x.off = x.len + v.C;
  int* ch_entry = x.off + v.C;           // [TR_CHUNK] list entry (1-based) of every transfer of the chunk
  int* ch_flags = ch_entry + TR_CHUNK;   // bit 0 add, bit 1 remove
  int* sdir = ch_flags + TR_CHUNK;       // [nd] direction slot in sorted order
  // (sdir + nd .. sdir + 3 nd: scratch, part of per_warp)
  double* sprob = (double*)(s_i + (size_t)warps_per_cta * per_warp) + (size_t)warp * nd;  // [nd] sorted probabilities

  const size_t base = (size_t)cell * v.cap;
  const int n0 = v.count[cell];
  const int nx = t.nx, nz = t.nz;
  const int i = cell % nx, k = (cell / nx) % nz, j = cell / (nx * nz);
  int err = 0;
  int* odir = t.out_dir + (size_t)cell * v.C * nd * 2;
  for (int q = lane; q < v.C * nd * 2; q += 32) odir[q] = 0;
  if (v.status[cell] != 0) {  // a failed cell neither sends nor loses particles
    int* surv0 = v.L + (size_t)cell * 3 * v.cap;
    for (int p = lane; p < n0; p += 32) surv0[p] = p;
    if (lane == 0) {
      t.n_surv[cell] = n0;
      t.n_out[cell] = 0;
    }
    return;
  }
  // ---- class lists: positions of class c in ascending order (a fresh aero_state_sort, aero_sorted.F90:185-209)
  {
    int run = 0;
    for (int c = 0; c < v.C; c++) {
      if (lane == 0) x.off[c] = run;
      int m = 0;
      for (int p0 = 0; p0 < n0; p0 += 32) {
        const int p = p0 + lane;
        const bool is = (p < n0) && (v.wgc[base + p] == c * v.G);
        const unsigned b = __ballot_sync(0xffffffffu, is);
        if (is) {
          const int idx = m + __popc(b & ((1u << lane) - 1u));
          x.LST[run + idx] = p;
          x.FXC[p] = (c << 24) | idx;
        }
        m += __popc(b);
      }
      if (lane == 0) x.len[c] = m;
      run += m;
    }
    for (int p = lane; p < n0; p += 32) x.P2O[p] = p;
    if (run != n0) err = PMCGPU_ERR_ARG;  // particles outside group 1 are not transported (WRF has one group)
    x.n_cur = n0;
  }
  __syncwarp();
  const unsigned gcell = (unsigned)v.gcell[cell];
  int n_out = 0;
  int* out_slot = t.out_slot + base;
  long long* out_newid = t.out_newid + base;
  __syncwarp();
  const double source_vol = t.cellvol_ext[tr_ext(t, i, k, j)];

  // ---- aero_state_scale_weight removals decided by the preweighting (magnitude already scaled): all classes first,
  // like trans_aero_preweight (:1373-1400), which runs before any particle is sampled (the order of the
  // swap-with-last removals decides the particle order the sampling sees)
  for (int c = 0; c < v.C && !err; c++) {
    const double sr = t.scale_ratio[(size_t)cell * v.C + c];
    if (sr > 0.0) {
      const unsigned gc = (unsigned)(c * v.G);
      const int n_part = x.len[c];
      const int n_remove = prob_round((double)n_part * (1.0 - 1.0 / sr), rng_uniform(RKey{v.seed, gcell, step, PH_SCALE, gc, 0u}, 0));
      if (lane == 0) {
        for (int r = 1; r <= n_remove; r++) {
          const int ne = x.len[c];
          const int ie = pmc_rand_int(rng_uniform(RKey{v.seed, gcell, step, PH_SCALE, gc, (unsigned)r}, 0), ne);
          tr_remove(x, x.LST[x.off[c] + ie - 1]);
        }
      }
      x.n_cur = __shfl_sync(0xffffffffu, x.n_cur, 0);
      __syncwarp();
    }
  }
  for (int c = 0; c < v.C && !err; c++) {
    // ---- move probabilities of this class, ascending, ties in collection order (real_sort, util.F90:1612)
    int count = 0;
    const int kindS = tr_kind(v, cell);
    if (lane == 0 && kindS != 0) {
      // boundary cell of an open domain: no multinomial sampling; the edge cells feed their interior neighbour with
      // probability min(advection + diffusion, 1) (boundary inflow, wrf_pmc_trans_aero.F90:689-766), corners nothing
      if (kindS <= 4) {
        const int q = (kindS == 1) ? 1 : (kindS == 2 ? 0 : (kindS == 3 ? 3 : 2));
        const double pr = fmin(tr_hprob(t, cell, c, q), 1.0);
        if (pr > 0.0) { sprob[0] = pr; sdir[0] = q; count = 1; }
      }
    }
    if (lane == 0 && kindS == 0) {
      for (int q = 0; q < 4; q++) {
        const double pr = tr_hprob(t, cell, c, q);
        if (pr > 0.0) { sprob[count] = pr; sdir[count] = q; count++; }
      }
      for (int kk = 0; kk < nz; kk++) {
        if (kk == k) continue;
        const double pr = t.pvert[((size_t)cell * v.C + c) * nz + kk];
        if (pr > 0.0) { sprob[count] = pr; sdir[count] = 4 + kk; count++; }
      }
      for (int a = 1; a < count; a++) {  // stable insertion sort
        const double pa = sprob[a];
        const int da = sdir[a];
        int b = a - 1;
        while (b >= 0 && sprob[b] > pa) { sprob[b + 1] = sprob[b]; sdir[b + 1] = sdir[b]; b--; }
        sprob[b + 1] = pa;
        sdir[b + 1] = da;
      }
    }
    count = __shfl_sync(0xffffffffu, count, 0);
    __syncwarp();
    const double mag_src = v.wmag[((size_t)cell * v.C + c) * v.G];
    double remainder_prob = 1.0;
    for (int s = 0; s < count && !err; s++) {
      const double prob = sprob[s];
      const double adjusted_prob = fmin(prob / remainder_prob, 1.0);
      if (!(remainder_prob >= 1.0e-8)) continue;
      const int d = sdir[s];
      int di = 0, dj = 0, dk = k;
      if (d == 0) di = -1; else if (d == 1) di = 1; else if (d == 2) dj = -1; else if (d == 3) dj = 1; else dk = d - 4;
      const int ed = tr_ext(t, i + di, dk, j + dj);
      const double dest_vol = t.cellvol_ext[ed];
      const double factor = source_vol / dest_vol;
      const double mag_to = t.wmag_ext[(size_t)ed * v.C + c] * (1.0 / factor);
      const double ratio = mag_to / mag_src;
      const int n_part_from = x.len[c];
      const unsigned a = (unsigned)c * 64u + (unsigned)s;
      RStream s0{RKey{v.seed, gcell, step, PH_TRANS, a, 0u}, 0u};
      int n_transfer;
      if (ratio >= 1.0) n_transfer = rand_binomial(n_part_from, adjusted_prob, s0);
      else n_transfer = rand_poisson((double)n_part_from * (adjusted_prob / ratio), s0);
      // flat weights: every particle of the class sees the same pair (mix_rat_from, mix_rat_to)
      const int mode = (mag_to == mag_src) ? 0 : (mag_to < mag_src ? 1 : 2);
      const double thr = (mode == 1) ? mag_to / mag_src : (mode == 2 ? mag_src / mag_to : 1.0);
      const int off0 = n_out;
      bool stop = false;
      for (int t0 = 0; t0 < n_transfer && !stop && !err; t0 += TR_CHUNK) {
        const int tt = t0 + lane + 1;  // 1-based transfer number
        const bool active = tt <= n_transfer;
        bool do_add = false, do_remove = false;
        double u1 = 0.0;
        if (active) {
          const RKey key{v.seed, gcell, step, PH_TRANS, a, (unsigned)tt};
          u1 = rng_uniform(key, 0);
          if (mode == 0) { do_add = true; do_remove = true; }
          else if (mode == 1) { do_add = true; do_remove = rng_uniform(key, 1) < thr; }
          else { do_add = rng_uniform(key, 1) < thr; do_remove = true; }
        }
        const unsigned rem_mask = __ballot_sync(0xffffffffu, do_remove);
        const int n_before = x.len[c] - __popc(rem_mask & ((1u << lane) - 1u));
        // the reference leaves the loop when the class runs empty (:2259-2260)
        const unsigned dead = __ballot_sync(0xffffffffu, active && n_before <= 0);
        const int first_dead = dead ? (__ffs(dead) - 1) : 32;
        const bool live = active && lane < first_dead;
        ch_entry[lane] = live ? pmc_rand_int(u1, n_before) : 0;
        ch_flags[lane] = live ? ((do_add ? 1 : 0) | (do_remove ? 2 : 0)) : 0;
        __syncwarp();
        if (lane == 0) {
          for (int l = 0; l < TR_CHUNK; l++) {
            const int fl = ch_flags[l];
            if (fl == 0) break;
            const int pos = x.LST[x.off[c] + ch_entry[l] - 1];
            if (fl & 1) {
              if (n_out >= v.cap) { err = PMCGPU_ERR_CAPACITY; break; }
              out_slot[n_out] = x.P2O[pos];
              long long nid = 0;
              if (!(fl & 2)) {  // a copy stays behind: the transported one gets a new id (:2297-2301)
                const long long nxid = v.next_id[cell];
                nid = (nxid - 1) * v.n_cells_global + v.gcell[cell] + 1;
                v.next_id[cell] = nxid + 1;
              }
              out_newid[n_out] = nid;
              n_out++;
            }
            if (fl & 2) tr_remove(x, pos);
          }
        }
        n_out = __shfl_sync(0xffffffffu, n_out, 0);
        err = __shfl_sync(0xffffffffu, err, 0);
        x.n_cur = __shfl_sync(0xffffffffu, x.n_cur, 0);
        __syncwarp();
        if (dead) stop = true;
      }
      if (lane == 0) {
        odir[(c * nd + d) * 2 + 0] = off0;
        odir[(c * nd + d) * 2 + 1] = n_out - off0;
      }
      remainder_prob = remainder_prob - prob;
    }
  }
  __syncwarp();
  // survivors in their final order
  int* surv = v.L + (size_t)cell * 3 * v.cap;
  for (int p = lane; p < x.n_cur; p += 32) surv[p] = x.P2O[p];
  if (lane == 0) {
    t.n_surv[cell] = x.n_cur;
    t.n_out[cell] = n_out;
    if (err && v.status[cell] == 0) v.status[cell] = err;
  }
}

// ---- ghost particles (C4, :543-675): count, scan and copy the out lists that point across side s
__global__ void k_tr_pack_count(DevView v, TransView t, int side, int* __restrict__ cnt) {
  const int L = (side < 2) ? t.nx : t.ny;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= L * t.nz) return;
  // ghost cells of a side in the reference's pack order: j, k, i loops -> (k, i) for rows, (j, k) for columns
  int i, j, k, slot;
  if (side < 2) { k = q / t.nx; i = q % t.nx; j = (side == 0) ? 0 : t.ny - 1; slot = (side == 0) ? 2 : 3; }
  else { j = q / t.nz; k = q % t.nz; i = (side == 2) ? 0 : t.nx - 1; slot = (side == 2) ? 0 : 1; }
  const int cell = tr_cell(t, i, k, j);
  const int nd = TR_ND(t.nz);
  int m = 0;
  for (int c = 0; c < v.C; c++) m += t.out_dir[(((size_t)cell * v.C + c) * nd + slot) * 2 + 1];
  cnt[q] = m;
}
__global__ void k_tr_scan(int n, const int* __restrict__ cnt, int* __restrict__ off) {  // single CTA exclusive scan
  __shared__ int s_carry;
  __shared__ int s_w[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < n; b0 += blockDim.x) {
    const int q = b0 + tid;
    const int x = (q < n) ? cnt[q] : 0;
    int incl = x;
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = (lane < (int)(blockDim.x >> 5)) ? s_w[lane] : 0;
      int wi = w;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += y;
      }
      s_w[lane] = wi - w;
    }
    __syncthreads();
    const int carry = s_carry;
    if (q < n) off[q] = carry + s_w[warp] + incl - x;
    __syncthreads();
    if (tid == blockDim.x - 1) s_carry = carry + s_w[warp] + incl;
    __syncthreads();
  }
  if (tid == 0) off[n] = s_carry;
}
__global__ void k_tr_pack_copy(DevView v, TransView t, int side, const int* __restrict__ off, TrSlab slab) {
  const int q = blockIdx.x;
  int i, j, k, slot;
  if (side < 2) { k = q / t.nx; i = q % t.nx; j = (side == 0) ? 0 : t.ny - 1; slot = (side == 0) ? 2 : 3; }
  else { j = q / t.nz; k = q % t.nz; i = (side == 2) ? 0 : t.nx - 1; slot = (side == 2) ? 0 : 1; }
  const int cell = tr_cell(t, i, k, j);
  const int nd = TR_ND(t.nz);
  const size_t base = (size_t)cell * v.cap, NS = (size_t)v.n_cells * v.cap;
  int o = off[q];
  if (off[q + 1] > slab.cap) {  // beyond the pitch of this message: reported, the receiver skips the cell too
    if (threadIdx.x == 0 && v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_CAPACITY;
    return;
  }
  for (int c = 0; c < v.C; c++) {
    const int lo = t.out_dir[(((size_t)cell * v.C + c) * nd + slot) * 2 + 0];
    const int m = t.out_dir[(((size_t)cell * v.C + c) * nd + slot) * 2 + 1];
    for (int e = threadIdx.x; e < m; e += blockDim.x) {
      const size_t src = base + lo + e;  // staged out list (k_tr_stage): contiguous entries, new ids applied
      const size_t dst = (size_t)o + e;
      for (int s = 0; s < v.S; s++) slab.vol[(size_t)s * slab.cap + dst] = v.vol_alt[(size_t)s * NS + src];
      slab.id[dst] = v.id_alt[src];
      slab.wgc[dst] = v.wgc_alt[src];
      slab.nprim[dst] = v.nprim_alt[src];
      slab.hyst[dst] = v.hyst_alt[src];
      slab.lct[dst] = v.lct_alt[src];
      slab.gct[dst] = v.gct_alt[src];
      if (v.cold.n) cold_copy(slab.cold, (size_t)slab.cap, dst, v.cold_alt, NS, src, v.n_opq);
    }
    o += m;
  }
}

// ---- add the arrivals (:777-804 aero_state_add_particles): every destination cell pulls survivors and
// arrivals into the alternate slab.  Arrival order = the order in which the reference's loops append to
// delta_aero_states(dest): local sources in (j, k, i) order, then the four ghost receives.
// candidate segments of one destination cell: survivors, (nz + 3) source cells x C classes, four ghost sides
// (+ 4 C boundary-inflow candidates of an open domain: west, east, south, north boundary neighbour x class)
static inline __host__ __device__ int tr_n_cand(int nz, int C) { return 1 + (nz + 3) * C + 4 + 4 * C; }
// one particle's species rows from one slab to another (the slabs never overlap): loads in batches of eight
// ahead of the stores
__device__ __forceinline__ void tr_copy_rows(double* __restrict__ dst, size_t dstride, const double* __restrict__ src,
                                             size_t sstride, int S) {
  for (int s = 0; s < S; s += 8) {  // the last batch is predicated, not serialised
    double a[8];
#pragma unroll
    for (int u = 0; u < 8; u++) a[u] = (s + u < S) ? src[(size_t)(s + u) * sstride] : 0.0;
#pragma unroll
    for (int u = 0; u < 8; u++)
      if (s + u < S) dst[(size_t)(s + u) * dstride] = a[u];
  }
}
// ---- the out list of every cell (sampled particles of all classes and directions, in list order) copied to the
// ALTERNATE slab at [cell][entry], ids of copies applied: the destination cells read their arrivals from there, so the
// main slab of a cell is touched by its own CTA only and survivors need not move (k_tr_gather)
// Species rows: when a third or more of the cell's particles were sampled, the scattered 8-byte reads would pull more
// 32-byte sectors than the rows hold, so the CTA streams every row once through shared memory (coalesced) and picks
// the sampled slots from there; sparse out lists read their slots directly.  `dense_ok`: the launch carries
// cap doubles of dynamic shared memory.
#define TR_STAGE_THREADS 256
__global__ void __launch_bounds__(TR_STAGE_THREADS) k_tr_stage(DevView v, TransView t, int dense_ok) {
  const int cell = blockIdx.x;
  const int n_out = t.n_out[cell];
  if (n_out == 0) return;
  const int n = v.count[cell];  // still the population before transport (k_tr_gather sets the new one)
  const size_t base = (size_t)cell * v.cap, NS = (size_t)v.n_cells * v.cap;
  extern __shared__ double s_row[];  // [cap] when dense_ok
  const bool dense = dense_ok && 3 * n_out >= n;
  if (dense) {
    for (int sp = 0; sp < v.S; sp++) {
      for (int p = threadIdx.x; p < n; p += TR_STAGE_THREADS) s_row[p] = v.vol[(size_t)sp * NS + base + p];
      __syncthreads();
      for (int e = threadIdx.x; e < n_out; e += TR_STAGE_THREADS)
        v.vol_alt[(size_t)sp * NS + base + e] = s_row[t.out_slot[base + e]];
      __syncthreads();
    }
  }
  for (int e = threadIdx.x; e < n_out; e += TR_STAGE_THREADS) {
    const size_t src = base + t.out_slot[base + e], dst = base + e;
    const long long nid = t.out_newid[base + e];
    const long long r_id = nid ? nid : v.id[src];
    const int r_wgc = v.wgc[src], r_nprim = v.nprim[src], r_hyst = v.hyst[src];
    const double r_lct = v.lct[src], r_gct = v.gct[src];
    if (!dense) tr_copy_rows(v.vol_alt + dst, NS, v.vol + src, NS, v.S);
    v.id_alt[dst] = r_id;
    v.wgc_alt[dst] = r_wgc;
    v.nprim_alt[dst] = r_nprim;
    v.hyst_alt[dst] = r_hyst;
    v.lct_alt[dst] = r_lct;
    v.gct_alt[dst] = r_gct;
    if (v.cold.n) cold_copy(v.cold_alt, NS, dst, v.cold, NS, src, v.n_opq);
  }
}
struct TrSeg {
  int kind;       // 0 survivors, 1 local out list, 2 halo slab
  int src;        // source cell (kind 0, 1) or side (kind 2)
  int off;        // offset in the list / slab
  int n;
};
// The four receive slabs travel as ONE __grid_constant__ parameter: the side is chosen at run time, and a reference
// picked among four by-value parameters made the compiler copy all of them (384 B) to every thread's stack at kernel
// entry -- 1.6 GB of local-memory stores per launch on K4, more than the records the kernel moves.
struct TrSlab4 {
  TrSlab s[4];
};
__global__ void __launch_bounds__(128, 8) k_tr_gather(DevView v, TransView t, const __grid_constant__ TrSlab4 rr) {
  const int cell = blockIdx.x;
  const int nx = t.nx, nz = t.nz, ny = t.ny, nd = TR_ND(t.nz);
  const int i = cell % nx, k = (cell / nx) % nz, j = cell / (nx * nz);
  extern __shared__ TrSeg seg[];  // [tr_n_cand(nz, C)]
  __shared__ int s_nseg, s_total;
  if (threadIdx.x < 32) {
    // Candidate segments in arrival order: e = 0 the survivors, then (source cell, class) pairs of the local
    // sources, then the four ghost receives.  Lanes read the candidates of a round side by side; the
    // non-empty ones are appended in order (ballot compaction).
    const int lane = threadIdx.x;
    const int C = v.C;
    const int n_src = nz + 3;  // y-, below (k), x-, x+, above (nz-1-k), y+
    const int n_main = 1 + n_src * C + 4;
    const int n_cand = n_main + 4 * C;
    const int kindD = tr_kind(v, cell);
    int ns = 0, tot = 0, fail = 0;
    const int cap = v.cap;
    for (int e0 = 0; e0 < n_cand; e0 += 32) {
      const int e = e0 + lane;
      int kind = 0, src = 0, off = 0, cnt = 0;
      if (e == 0) {
        kind = 0; src = cell; off = 0; cnt = t.n_surv[cell];
      } else if (kindD != 0) {
        // boundary cell of an open domain: what moves into it leaves the domain (never added, :777-804)
      } else if (e < 1 + n_src * C) {
        const int u = (e - 1) / C, c = (e - 1) % C;
        int si = i, sk = k, sj = j, slot = 0;
        bool ok = true;
        if (u == 0) { sj = j - 1; slot = 3; ok = (j - 1 >= 0); }
        else if (u <= k) { sk = u - 1; slot = 4 + k; }
        else if (u == k + 1) { si = i - 1; slot = 1; ok = (i - 1 >= 0); }
        else if (u == k + 2) { si = i + 1; slot = 0; ok = (i + 1 <= nx - 1); }
        else if (u <= nz + 1) { sk = u - 2; slot = 4 + k; }   // u = k+3 .. nz+1 -> sk = k+1 .. nz-1
        else { sj = j + 1; slot = 2; ok = (j + 1 <= ny - 1); }
        if (ok) {
          const int sc = tr_cell(t, si, sk, sj);
          // a boundary cell is not part of the multinomial sampling: its inflow arrives behind the ghost sides
          if (tr_kind(v, sc) == 0) {
            const int* od = t.out_dir + (((size_t)sc * C + c) * nd + slot) * 2;
            kind = 1; src = sc; off = od[0]; cnt = od[1];
          }
        }
      } else if (e < n_main) {
        const int g = e - (1 + n_src * C);  // 0: side 1 (j == ny-1), 1: side 0 (j == 0), 2: side 3, 3: side 2
        const int side = (g == 0) ? 1 : (g == 1 ? 0 : (g == 2 ? 3 : 2));
        const bool ok = (g == 0) ? (j == ny - 1) : (g == 1 ? (j == 0) : (g == 2 ? (i == nx - 1) : (i == 0)));
        if (ok) {
          const int q = (side < 2) ? (k * nx + i) : (j * nz + k);
          kind = 2; src = side; off = t.roff[side][q]; cnt = t.roff[side][q + 1] - off;
          const int pitch = rr.s[side].cap;
          if (off + cnt > pitch) { cnt = 0; fail = 1; }  // the sender could not pack this cell (message pitch)
        }
      } else if (e < n_cand && v.kind) {
        // boundary inflow (:689-766), after the ghost receives: west, east, south, north boundary neighbour
        const int b = (e - n_main) / C, c = (e - n_main) % C;
        int si = i, sj = j, slot = 0, want = 0;
        if (b == 0) { si = i - 1; slot = 1; want = 1; }
        else if (b == 1) { si = i + 1; slot = 0; want = 2; }
        else if (b == 2) { sj = j - 1; slot = 3; want = 3; }
        else { sj = j + 1; slot = 2; want = 4; }
        if (si >= 0 && si <= nx - 1 && sj >= 0 && sj <= ny - 1) {
          const int sc = tr_cell(t, si, k, sj);
          if (tr_kind(v, sc) == want) {
            const int* od = t.out_dir + (((size_t)sc * C + c) * nd + slot) * 2;
            kind = 1; src = sc; off = od[0]; cnt = od[1];
          }
        }
      }
      // a cell that would outgrow its slab keeps what fits (survivors first, then arrivals in order) and is
      // reported: the device state stays consistent (count = records written)
      int incl = cnt;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      const int room = cap - tot - (incl - cnt);
      if (cnt > room) { cnt = room > 0 ? room : 0; fail = 1; }
      const bool have = cnt > 0;
      const unsigned m = __ballot_sync(0xffffffffu, have);
      const int pos = ns + __popc(m & ((1u << lane) - 1u));
      if (have) {
        seg[pos].kind = kind; seg[pos].src = src; seg[pos].off = off; seg[pos].n = cnt;
      }
      int csum = have ? cnt : 0;
      for (int o = 16; o > 0; o >>= 1) csum += __shfl_xor_sync(0xffffffffu, csum, o);
      tot += csum;
      ns += __popc(m);
    }
    fail = __any_sync(0xffffffffu, fail);
    if (lane == 0) {
      s_nseg = ns;
      s_total = tot;
      if (fail && v.status[cell] == 0) v.status[cell] = PMCGPU_ERR_CAPACITY;
    }
  }
  __syncthreads();
  const size_t NS = (size_t)v.n_cells * v.cap;
  const size_t dbase = (size_t)cell * v.cap;
  const int total = s_total, nseg = s_nseg;
  // ---- survivors stay where they are.  The swap-with-last removals only ever move records out of the tail
  // [n_surv, n_old) into holes below n_surv (a record at slot s < n_surv is never the last one while a removal is
  // still to come), so the holes are filled in place: reads from the tail, writes below it, no overlap.
  const bool have_surv = nseg > 0 && seg[0].kind == 0;  // always the first segment when there are survivors
  const int n_surv = have_surv ? seg[0].n : 0;
  {
    const int* __restrict__ surv = v.L + (size_t)cell * 3 * v.cap;
    for (int idx = threadIdx.x; idx < n_surv; idx += blockDim.x) {
      const int from = surv[idx];
      if (from == idx) continue;
      const size_t src = dbase + from, dst = dbase + idx;
      const long long r_id = v.id[src];
      const int r_wgc = v.wgc[src], r_nprim = v.nprim[src], r_hyst = v.hyst[src];
      const double r_lct = v.lct[src], r_gct = v.gct[src];
      tr_copy_rows(v.vol + dst, NS, v.vol + src, NS, v.S);
      v.id[dst] = r_id;
      v.wgc[dst] = r_wgc;
      v.nprim[dst] = r_nprim;
      v.hyst[dst] = r_hyst;
      v.lct[dst] = r_lct;
      v.gct[dst] = r_gct;
      if (v.cold.n) cold_copy(v.cold, NS, dst, v.cold, NS, src, v.n_opq);
    }
  }
  __syncthreads();  // the tail has been read: arrivals may overwrite it
  // ---- arrivals behind the survivors, from the staged out lists of the source cells (alternate slab, written by
  // k_tr_stage: this cell's own main-slab records are read by nobody else) and from the halo messages
  for (int idx = n_surv + threadIdx.x; idx < total; idx += blockDim.x) {
    int sgi = have_surv ? 1 : 0, rel = idx - n_surv;
    while (sgi < nseg - 1 && rel >= seg[sgi].n) { rel -= seg[sgi].n; sgi++; }
    const TrSeg sg = seg[sgi];
    const size_t dst = dbase + idx;
    long long r_id;
    int r_wgc, r_nprim, r_hyst;
    double r_lct, r_gct;
    if (sg.kind == 2) {
      const TrSlab& sl = rr.s[sg.src];
      const size_t src = (size_t)sg.off + rel;
      r_id = sl.id[src]; r_wgc = sl.wgc[src]; r_nprim = sl.nprim[src]; r_hyst = sl.hyst[src];
      r_lct = sl.lct[src]; r_gct = sl.gct[src];
      tr_copy_rows(v.vol + dst, NS, sl.vol + src, (size_t)sl.cap, v.S);
      if (v.cold.n) cold_copy(v.cold, NS, dst, sl.cold, (size_t)sl.cap, src, v.n_opq);
    } else {
      const size_t src = (size_t)sg.src * v.cap + sg.off + rel;  // entry of the source cell's staged out list
      r_id = v.id_alt[src]; r_wgc = v.wgc_alt[src]; r_nprim = v.nprim_alt[src]; r_hyst = v.hyst_alt[src];
      r_lct = v.lct_alt[src]; r_gct = v.gct_alt[src];
      tr_copy_rows(v.vol + dst, NS, v.vol_alt + src, NS, v.S);
      if (v.cold.n) cold_copy(v.cold, NS, dst, v.cold_alt, NS, src, v.n_opq);
    }
    v.id[dst] = r_id;
    v.wgc[dst] = r_wgc;
    v.nprim[dst] = r_nprim;
    v.hyst[dst] = r_hyst;
    v.lct[dst] = r_lct;
    v.gct[dst] = r_gct;
  }
  if (threadIdx.x == 0) {
    v.count[cell] = total;
    t.n_moved[cell] = total - t.n_surv[cell];
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
int pmc_sort_grid_only_async(pmcgpu_ctx* ctx, int interior_only);
int pmc_boundary_refresh_async(pmcgpu_ctx* ctx, unsigned step);
int pmc_rebalance_async(pmcgpu_ctx* ctx, int allow_doubling, int allow_halving, unsigned step);
void pmc_swap_slabs(pmcgpu_ctx* ctx);
int pmc_exchange(pmcgpu_ctx* ctx, void* const send[4], void* const recv[4], const size_t bytes[4]);

template <typename T>
static int tr_alloc(pmcgpu_ctx* ctx, T** p, size_t n) {
  void* q = nullptr;
  CUDA_OK(ctx, cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
  CUDA_OK(ctx, cudaMemsetAsync(q, 0, std::max<size_t>(n, 1) * sizeof(T), ctx->stream));
  ctx->allocs.push_back(q);
  *p = (T*)q;
  return 0;
}
// allocate the transport work space on first use
static int tr_setup(pmcgpu_ctx* ctx) {
  if (ctx->trans_ready) return PMCGPU_OK;
  DevView& v = ctx->v;
  TransView& t = ctx->t;
  memset(&t, 0, sizeof(t));
  t.nx = ctx->cfg.nx; t.nz = ctx->cfg.nz; t.ny = ctx->cfg.ny; t.C = v.C;
  if (v.G != 1) {
    ctx->err = "transport samples weight group 1 only (wrf_pmc_trans_aero.F90:2229); n_group must be 1";
    return PMCGPU_ERR_ARG;
  }
  if (t.nz < 2 || t.nz > 60 || v.C > 127) return PMCGPU_ERR_ARG;
  if ((size_t)tr_n_cand(t.nz, v.C) * sizeof(TrSeg) > 160 * 1024) {
    ctx->err = "transport: (nz + 3) * n_class + 5 arrival segments per cell exceed the shared memory of k_tr_gather";
    return PMCGPU_ERR_ARG;
  }
  const size_t nc = v.n_cells, C = v.C, nz = t.nz, nx = t.nx, ny = t.ny;
  const size_t n_ext = (nx + 2) * nz * (ny + 2);
  const int nd = TR_ND(t.nz);
  int rc;
#define TA(p, n) if ((rc = tr_alloc(ctx, &(p), (n)))) return rc;
  TA(t.u_flux, (nx + 1) * nz * ny * C) TA(t.v_flux, nx * nz * (ny + 1) * C) TA(t.w_flux, nx * (nz + 1) * ny * C)
  TA(t.xkhh, n_ext) TA(t.alt, n_ext) TA(t.zc, nc) TA(t.z_at_w, nx * (nz + 1) * ny) TA(t.exch_h, nx * (nz + 1) * ny)
  TA(t.ccount, nc * C) TA(t.num_old, nc * C) TA(t.scale_ratio, nc * C)
  TA(t.padv, nc * C * 6) TA(t.pdiff, nc * C * 4) TA(t.pvert, nc * C * nz)
  TA(t.cellvol_ext, n_ext) TA(t.wmag_ext, n_ext * C)
  TA(t.out_dir, nc * C * nd * 2) TA(t.n_surv, nc) TA(t.n_out, nc) TA(t.n_moved, nc)
  for (int s = 0; s < 4; s++) {
    const size_t L = (s < 2) ? nx : ny;
    TA(t.sbuf[s], L * nz * C) TA(t.rbuf[s], L * nz * C)
    TA(t.scnt[s], L * nz + 1)
    // a quarter of the slots of every edge cell may cross one side in one step (pitch: a multiple of 4 records)
    int slab_cap = (int)std::min<size_t>(((size_t)L * nz * (size_t)std::max(v.cap / 4, 64) + 3) & ~(size_t)3, (size_t)1 << 30);
    if (const char* e = getenv("PMCGPU_HALO_SLAB_CAP")) slab_cap = std::max(4, atoi(e) & ~3);  // tests: force an overflow
    ctx->slab_cap[s] = slab_cap;
    ctx->hdr_ints[s] = (int)((L * nz + 1 + 3) & ~(size_t)3);
    const size_t bytes = (size_t)ctx->hdr_ints[s] * sizeof(int) + (size_t)slab_cap * tr_rec_bytes(v.S, v.cold.n != nullptr, v.n_opq);
    TA(ctx->smsg[s], bytes) TA(ctx->rmsg[s], bytes)
    t.soff[s] = (int*)ctx->smsg[s];  // the offsets travel as the header of the particle message
    t.roff[s] = (int*)ctx->rmsg[s];
  }
#undef TA
  t.out_slot = v.fidx;                // free outside mc_coag
  t.out_newid = (long long*)v.lnr;    // free outside aero_state_sort
  CUDA_OK(ctx, cudaMallocHost((void**)&ctx->h_tot, 8 * sizeof(int)));
  CUDA_OK(ctx, cudaEventCreateWithFlags(&ctx->ev_tot, cudaEventDisableTiming));
  ctx->tot_valid = false;
  ctx->trans_ready = true;
  return PMCGPU_OK;
}

extern "C" int pmcgpu_wrf_pmc_trans_driver(pmcgpu_ctx* ctx, const pmcgpu_trans_args* a) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_wrf_pmc_trans_driver");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !a) return PMCGPU_ERR_ARG;
  int rc = tr_setup(ctx);
  if (rc) return rc;
  DevView& v = ctx->v;
  TransView& t = ctx->t;
  cudaStream_t st = ctx->stream;
  const size_t C = v.C, nz = t.nz, nx = t.nx, ny = t.ny, nc = v.n_cells;
  const size_t n_ext = (nx + 2) * nz * (ny + 2);
  t.dx = a->dx; t.dy = a->dy; t.dt = a->dt; t.vertmix = a->vertmix_onoff;
#define UP(dst, src, n) \
  if (src) CUDA_OK(ctx, cudaMemcpyAsync(dst, src, (size_t)(n) * sizeof(float), cudaMemcpyDefault, st));
  UP(t.u_flux, a->u_flux, (nx + 1) * nz * ny * C) UP(t.v_flux, a->v_flux, nx * nz * (ny + 1) * C)
  UP(t.w_flux, a->w_flux, nx * (nz + 1) * ny * C) UP(t.xkhh, a->xkhh, n_ext) UP(t.alt, a->alt, n_ext)
  UP(t.zc, a->z, nc) UP(t.z_at_w, a->z_at_w, nx * (nz + 1) * ny) UP(t.exch_h, a->exch_h, nx * (nz + 1) * ny)
#undef UP
  const int TB = 256;
  auto blocks = [&](size_t n) { return (unsigned)((n + TB - 1) / TB); };
  // ---- aero_state_sort of every cell at the top of trans_aero_preweight (:1210): only its bin-grid
  // decision has an effect that outlives the call
  if (v.kind && !ctx->bnd_buf) {
    ctx->err = "open domain: call pmcgpu_set_boundary before pmcgpu_wrf_pmc_trans_driver";
    return PMCGPU_ERR_ARG;
  }
  if ((rc = pmc_sort_grid_only_async(ctx, 0))) return rc;
  // ---- probabilities
  {
    ProfScope ps_(ctx, "k_tr_probs");
    k_tr_counts<<<(unsigned)((nc + 7) / 8), 256, 0, st>>>(v, t);
    LAUNCH_CHECK(ctx);
    k_tr_hprobs<<<blocks(nc * C), TB, 0, st>>>(v, t);
    LAUNCH_CHECK(ctx);
    const size_t smem = (4 * nz * nz + 5 * (nz + 1)) * sizeof(double);
    if (smem > 48 * 1024) CUDA_OK(ctx, cudaFuncSetAttribute(k_tr_vprobs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_tr_vprobs<<<(unsigned)(nx * ny), 256, smem, st>>>(v, t);
    LAUNCH_CHECK(ctx);
  }
  // ---- C1: ghost cell volumes
  auto to_ghost = [&](double* ext, int ncomp) -> int {
    void* sp[4]; void* rp[4]; size_t by[4];
    for (int s = 0; s < 4; s++) {
      const size_t L = (s < 2) ? nx : ny;
      k_tr_pack_edge<<<blocks(L * nz * ncomp), TB, 0, st>>>(t, ext, ncomp, s, t.sbuf[s]);
      LAUNCH_CHECK(ctx);
      sp[s] = t.sbuf[s]; rp[s] = t.rbuf[s]; by[s] = L * nz * ncomp * sizeof(double);
    }
    int e = pmc_exchange(ctx, sp, rp, by);
    if (e) return e;
    for (int s = 0; s < 4; s++) {
      const size_t L = (s < 2) ? nx : ny;
      k_tr_unpack_ghost<<<blocks(L * nz * ncomp), TB, 0, st>>>(t, ext, ncomp, s, t.rbuf[s]);
      LAUNCH_CHECK(ctx);
    }
    return PMCGPU_OK;
  };
  {
    ProfScope ps_(ctx, "k_tr_preweight");
    k_tr_fill_ext<<<blocks(nc), TB, 0, st>>>(v, t, 0);
    LAUNCH_CHECK(ctx);
    if ((rc = to_ghost(t.cellvol_ext, 1))) return rc;
    // ---- C2: flux sums towards the ghost ring, then the preweighting
    void* sp[4]; void* rp[4]; size_t by[4];
    for (int s = 0; s < 4; s++) {
      const size_t L = (s < 2) ? nx : ny;
      k_tr_edge_flux<<<blocks(L * nz * C), TB, 0, st>>>(v, t, s, t.sbuf[s]);
      LAUNCH_CHECK(ctx);
      sp[s] = t.sbuf[s]; rp[s] = t.rbuf[s]; by[s] = L * nz * C * sizeof(double);
    }
    if ((rc = pmc_exchange(ctx, sp, rp, by))) return rc;
    k_tr_preweight<<<blocks(nc * C), TB, 0, st>>>(v, t);
    LAUNCH_CHECK(ctx);
    // ---- C3: ghost weights
    if ((rc = to_ghost(t.wmag_ext, (int)C))) return rc;
  }
  // ---- sampling
  {
    ProfScope ps_(ctx, "k_tr_multisample");
    const int nd = TR_ND(t.nz);
    // small cells (at most half the slab capacity) first, then the rest
    const int lsmall = (v.cap > 512) ? ((v.cap / 2 + 1) & ~1) : v.cap;
    for (int pass = 0; pass < 2; pass++) {
      const int lcap = pass == 0 ? lsmall : v.cap;
      const int n_lo = pass == 0 ? -1 : lsmall, n_hi = lcap;
      if (pass == 1 && lsmall >= v.cap) break;
      const size_t per_warp_i = (3 * (size_t)lcap + 2 * C + 3 * TR_CHUNK + 3 * nd + 1) & ~(size_t)1;
      const size_t per_warp_b = per_warp_i * sizeof(int) + nd * sizeof(double);
      // CTAs of at most 56 KB so that four or more share an SM when the lists are short
      int wpc = (int)std::min<size_t>(4, (56 * 1024) / per_warp_b);
      if (wpc < 1) wpc = 1;
      if (per_warp_b > 200 * 1024) {
        ctx->err = "cell_capacity too large for the shared-memory index state of the transport sampler";
        return PMCGPU_ERR_ARG;
      }
      // the sorted probabilities (doubles) start behind the int regions of all warps
      const size_t smem_total = (size_t)wpc * per_warp_i * sizeof(int) + (size_t)wpc * nd * sizeof(double);
      if (smem_total > 48 * 1024 && smem_total > ctx->cfg_msamp_smem) {
        CUDA_OK(ctx, cudaFuncSetAttribute(k_tr_multisample, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_total));
        ctx->cfg_msamp_smem = smem_total;
      }
      k_tr_multisample<<<(unsigned)((nc + wpc - 1) / wpc), wpc * 32, smem_total, st>>>(v, t, a->step, wpc, lcap, n_lo,
                                                                                       n_hi);
      LAUNCH_CHECK(ctx);
    }
  }
  {
    // the sampled records of every cell, in list order, to the alternate slab
    ProfScope ps_(ctx, "k_tr_stage");
    const size_t row_bytes = (size_t)v.cap * sizeof(double);
    const int dense_ok = row_bytes <= 48 * 1024;
    k_tr_stage<<<(unsigned)nc, TR_STAGE_THREADS, dense_ok ? row_bytes : 0, st>>>(v, t, dense_ok);
    LAUNCH_CHECK(ctx);
  }
  // ---- C4: ghost particles: one message per neighbour (header with the per-cell offsets + SoA records), its
  // row pitch derived from the counts of the previous step, so the host never waits for this step's counts
  TrSlab rview[4];
  {
    ProfScope ps_(ctx, "k_tr_halo_particles");
    int pitch_s[4], pitch_r[4];
    if (ctx->tot_valid) CUDA_OK(ctx, cudaEventSynchronize(ctx->ev_tot));  // recorded a whole step ago
    for (int s = 0; s < 4; s++) {
      auto pitch_of = [&](int n_prev) {
        if (!ctx->tot_valid || n_prev < 0) return ctx->slab_cap[s];
        const long long want = ((long long)n_prev + n_prev / 2 + 2048 + 255) & ~255ll;
        return (int)std::min<long long>(want, ctx->slab_cap[s]);
      };
      pitch_s[s] = pitch_of(ctx->h_tot[s]);
      pitch_r[s] = pitch_of(ctx->h_tot[4 + s]);
    }
    void* sp[4]; void* rp[4]; size_t sby[4], rby[4];
    for (int s = 0; s < 4; s++) {
      const size_t L = (s < 2) ? nx : ny;
      k_tr_pack_count<<<blocks(L * nz), TB, 0, st>>>(v, t, s, t.scnt[s]);
      LAUNCH_CHECK(ctx);
      k_tr_scan<<<1, 1024, 0, st>>>((int)(L * nz), t.scnt[s], t.soff[s]);
      LAUNCH_CHECK(ctx);
      k_tr_pack_copy<<<(unsigned)(L * nz), 128, 0, st>>>(v, t, s, t.soff[s],
                                                         tr_slab_view(ctx->smsg[s], ctx->hdr_ints[s], pitch_s[s], v.S, v.cold.n != nullptr, v.n_opq));
      LAUNCH_CHECK(ctx);
      sp[s] = ctx->smsg[s]; rp[s] = ctx->rmsg[s];
      sby[s] = (size_t)ctx->hdr_ints[s] * sizeof(int) + (size_t)pitch_s[s] * tr_rec_bytes(v.S, v.cold.n != nullptr, v.n_opq);
      rby[s] = (size_t)ctx->hdr_ints[s] * sizeof(int) + (size_t)pitch_r[s] * tr_rec_bytes(v.S, v.cold.n != nullptr, v.n_opq);
      rview[s] = tr_slab_view(ctx->rmsg[s], ctx->hdr_ints[s], pitch_r[s], v.S, v.cold.n != nullptr, v.n_opq);
    }
    if ((rc = pmc_exchange_sized(ctx, sp, rp, sby, rby))) return rc;
    // the counts of this step size the messages of the next one
    for (int s = 0; s < 4; s++) {
      const size_t L = (s < 2) ? nx : ny;
      CUDA_OK(ctx, cudaMemcpyAsync(ctx->h_tot + s, t.soff[s] + L * nz, sizeof(int), cudaMemcpyDeviceToHost, st));
      CUDA_OK(ctx, cudaMemcpyAsync(ctx->h_tot + 4 + s, t.roff[s] + L * nz, sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    CUDA_OK(ctx, cudaEventRecord(ctx->ev_tot, st));
    ctx->tot_valid = true;
  }
  // ---- add + rebalance + sort
  {
    ProfScope ps_(ctx, "k_tr_gather");
    const size_t seg_bytes = (size_t)tr_n_cand(t.nz, v.C) * sizeof(TrSeg);
    if (seg_bytes > 40 * 1024 && !ctx->cfg_trans_smem) {
      CUDA_OK(ctx, cudaFuncSetAttribute(k_tr_gather, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)seg_bytes));
      ctx->cfg_trans_smem = true;
    }
    TrSlab4 rr;
    for (int q = 0; q < 4; q++) rr.s[q] = rview[q];
    k_tr_gather<<<(unsigned)nc, 128, (size_t)tr_n_cand(t.nz, v.C) * sizeof(TrSeg), st>>>(v, t, rr);
    LAUNCH_CHECK(ctx);
  }
  if ((rc = pmc_rebalance_async(ctx, a->allow_doubling, a->allow_halving, a->step))) return rc;
  if ((rc = pmc_sort_grid_only_async(ctx, 1))) return rc;
  // ---- open domain: refresh of the boundary cells (wrf_pmc_trans_aero.F90:811-955)
  if (v.kind && (rc = pmc_boundary_refresh_async(ctx, a->step))) return rc;
  if (a->no_wait) return PMCGPU_OK;
  return pmc_first_status(ctx);
}

extern "C" int pmcgpu_get_trans_probs(pmcgpu_ctx* ctx, double* adv, double* diff, double* vert) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !ctx->trans_ready) return PMCGPU_ERR_ARG;
  const size_t n = (size_t)ctx->v.n_cells * ctx->v.C;
  if (adv) CUDA_OK(ctx, cudaMemcpyAsync(adv, ctx->t.padv, n * 6 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (diff) CUDA_OK(ctx, cudaMemcpyAsync(diff, ctx->t.pdiff, n * 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (vert) CUDA_OK(ctx, cudaMemcpyAsync(vert, ctx->t.pvert, n * ctx->t.nz * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
extern "C" int pmcgpu_get_trans_moved(pmcgpu_ctx* ctx, int* n_moved) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !ctx->trans_ready || !n_moved) return PMCGPU_ERR_ARG;
  CUDA_OK(ctx, cudaMemcpyAsync(n_moved, ctx->t.n_moved, ctx->v.n_cells * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}
