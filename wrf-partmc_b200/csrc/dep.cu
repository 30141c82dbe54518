// dep.cu -- dry deposition of particles in the lowest model layer
// (interface/wrf_pmc_dep.F90:31-97 wrf_pmc_dep_dry_driver, :188-213 depvel_partmc;
//  interface/wrf_pmc_dep_aero.F90:31-150 dep_aero_dry_driver, :154-199 dry_dep_aero_state,
//  :204-275 deposition velocity and removal probability, :280-296 removal, :301-470 Zhang et al. (2001)
//  resistances).  SURVEY.md section 8f, row N2: the step right after transport in wrf_pmc_trans_driver.
//
// One warp per surface cell.  Every particle's removal probability (dt / dz) v_d is evaluated in parallel,
// its draw is addressed (PH_DEP, 0, particle index), the reference's backward swap-with-last loop is
// replayed on indices only and the surviving records then move in parallel (same scheme as
// aero_state_halve in rebalance.cu), so the result equals the sequential CPU loop particle for particle.
//
// The reference mixes default-real literals into double expressions; those constants carry their
// single-precision VALUES here ((double)1.257f ...), as in the compiled Fortran.
#include "pmcgpu_internal.cuh"

#define DEP_WARPS 4
#define DEP_AIR_DYN_VISC 1.78e-5  // constants.F90:65
#define DEP_GRAV 9.81             // constants.F90:33

__constant__ int c_zhang_from_usgs[25] = {15, 7, 7, 7, 7, 10, 6, 6, 6, 6, 4, 3, 2, 1, 5, 14, 11, 10, 8, 9, 10, 9, 8, 12, 8};
__constant__ float c_zh_alpha[15] = {1.0f, 0.6f, 1.1f, 0.8f, 0.8f, 1.2f, 1.2f, 50.0f, 50.0f, 1.3f, 2.0f, 50.0f, 100.0f, 100.0f, 1.5f};
__constant__ float c_zh_A1[15] = {0.002f, 0.005f, 0.002f, 0.005f, 0.005f, 0.002f, 0.002f, 9.999f, 9.999f, 0.010f, 0.010f, 9.999f, 9.999f, 9.999f, 0.010f};

// wrf_pmc_dep.F90:188-213
__device__ __forceinline__ double depvel_partmc(float rmol, double zr, double z0, double ustar) {
  const double vk = 0.40;
  if (fabsf(rmol) < 1.0e-6f) rmol = 0.f;
  const double c074 = (double)0.74f, c0001 = (double)0.001f;
  double polint;
  if (rmol < 0.0f) {
    const double ta = pow(1.0 - 9.0 * zr * (double)rmol, 0.25) + c0001;
    const double to = pow(1.0 - 9.0 * z0 * (double)rmol, 0.25) + c0001;
    const double ar = ta * ta, ao = to * to;
    polint = c074 * (log((ar - 1.0) / (ar + 1.0)) - log((ao - 1.0) / (ao + 1.0)));
  } else if (rmol == 0.0f) {
    polint = c074 * log(zr / z0);
  } else {
    const float f = __fmul_rn(4.7f, rmol);
    polint = c074 * log(zr / z0) + (double)f * (zr - z0);
  }
  return polint / (vk * fmax(ustar, (double)1.0e-4f));
}
// wrf_pmc_dep_aero.F90:417-431
__device__ __forceinline__ double dep_slip(double D_p) {
  const double lambda = .065e-6;
  return 1.0 + ((2.0 * lambda) / D_p) * ((double)1.257f + (double).4f * exp(-((double)1.1f * D_p) / (2.0 * lambda)));
}
// removal probability of one particle: calculate_vs :301, calculate_rs :320 (updated_version), calculate_vd :399,
// compute_dep_prob :254
__device__ __forceinline__ double dep_remove_prob(double diameter, double density, double temp, double rrho,
                                                  double ustar, double aer_res, int isurftype, double dt, double dz,
                                                  double* vd_out) {
  const double alpha = (double)c_zh_alpha[isurftype - 1], A = (double)c_zh_A1[isurftype - 1];
  const double c_c = dep_slip(diameter);
  const double vs = ((diameter * diameter) * density * DEP_GRAV * c_c) / (18.0 * DEP_AIR_DYN_VISC);
  const double diff = (PmcConst::boltzmann * temp * c_c) / (3.0 * PmcConst::pi * DEP_AIR_DYN_VISC * diameter);
  const double Sc = 1.45e-5 / diff;
  const double gamma = (double)(2.0f / 3.0f);
  const double C_b = (double).2f, nu = (double)0.8f, C_IM = (double)0.4f;
  const double St = (A > 1.0) ? (vs * (ustar * ustar)) / (DEP_AIR_DYN_VISC * rrho) : (vs * ustar) / (DEP_GRAV * A);
  const double EB = C_b * pow(Sc, -gamma);
  const double q = St / (alpha + St);
  const double EIM = C_IM * (q * q);
  const double EIN = .5 * pow(diameter / A, nu);
  const double R1 = exp(-sqrt(St));
  const double rs = 1.0 / (3.0 * ustar * (EB + EIM + EIN) * R1);
  const double tmp = aer_res + rs + aer_res * rs * vs;
  const double vd = (1.0 / tmp) + vs;
  if (vd_out) *vd_out = vd;
  return (dt / dz) * vd;
}
__device__ __forceinline__ int dep_zhang_type(int ivgtyp) {
  return (ivgtyp >= 1 && ivgtyp <= 25) ? c_zhang_from_usgs[ivgtyp - 1] : 7;
}

__global__ void __launch_bounds__(DEP_WARPS * 32) k_dep_dry(DevView v, int nx, int nz, int ny, int i_offset,
                                                             int j_offset, int global_nx, int global_ny,
                                                             const float* __restrict__ rmol, const float* __restrict__ znt,
                                                             const float* __restrict__ ust, const int* __restrict__ ivgtyp,
                                                             const double* __restrict__ height, double dt, unsigned step) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * DEP_WARPS + warp;
  if (col >= nx * ny) return;
  const int i = col % nx + 1, j = col / nx + 1;
  const int gi = i + i_offset, gj = j + j_offset;
  // dep_aero_dry_driver :113-114: the outermost ring of the global domain is skipped
  if (gi < 2 || gi > global_nx - 1 || gj < 2 || gj > global_ny - 1) return;
  const int cell = (i - 1) + nx * (0 + nz * (j - 1));  // k = 1
  if (v.status[cell] != 0) return;
  const size_t base = (size_t)cell * v.cap;
  const size_t NS = (size_t)v.n_cells * v.cap;
  const int n_part = v.count[cell];
  int* map = v.L + (size_t)cell * 3 * v.cap;  // scratch: current slot -> original slot
  const double ustar = (double)ust[col];
  const double aer_res = depvel_partmc(rmol[col], 2.0, (double)znt[col], ustar);
  const int isurf = dep_zhang_type(ivgtyp[col]);
  const double temp = v.temp[cell], rrho = v.rrho[cell], dz = height[col];
  // removal decisions, all draws in parallel; map[p] = -1 marks a removal
  for (int p = lane; p < n_part; p += 32) {
    double vt = 0.0, mt = 0.0;
    for (int s = 0; s < v.S; s++) {
      const double x = v.vol[(size_t)s * NS + base + p];
      vt += x;
      mt += x * v.density[s];
    }
    const double diam = 2.0 * vol2rad(vt);
    const double prob = dep_remove_prob(diam, mt / vt, temp, rrho, ustar, aer_res, isurf, dt, dz, nullptr);
    const RKey key{v.seed, (unsigned)v.gcell[cell], step, PH_DEP, 0u, (unsigned)(p + 1)};
    map[p] = (rng_uniform(key, 0) < prob) ? -1 : p;
  }
  __syncwarp();
  // replay the backward swap-with-last loop (:288-293) on indices only
  int n_cur = n_part;
  if (lane == 0) {
    for (int q = n_part - 1; q >= 0; q--) {
      if (map[q] < 0) {
        if (q != n_cur - 1) map[q] = map[n_cur - 1];
        n_cur--;
      }
    }
  }
  n_cur = __shfl_sync(0xffffffffu, n_cur, 0);
  __syncwarp();
  // sources are original slots >= n_cur, destinations < n_cur: disjoint
  for (int p = lane; p < n_cur; p += 32) {
    const int src = map[p];
    if (src != p) {
      for (int s = 0; s < v.S; s++) v.vol[(size_t)s * NS + base + p] = v.vol[(size_t)s * NS + base + src];
      v.id[base + p] = v.id[base + src];
      v.wgc[base + p] = v.wgc[base + src];
      v.nprim[base + p] = v.nprim[base + src];
      v.hyst[base + p] = v.hyst[base + src];
      v.lct[base + p] = v.lct[base + src];
      v.gct[base + p] = v.gct[base + src];
      if (v.cold.n) cold_copy(v.cold, NS, base + p, v.cold, NS, base + src, v.n_opq);
    }
  }
  if (lane == 0) v.count[cell] = n_cur;
}

extern "C" int pmcgpu_wrf_pmc_dep_dry_driver(pmcgpu_ctx* ctx, const pmcgpu_dep_args* a) {
  PMC_SPHERICAL_ONLY(ctx, "pmcgpu_wrf_pmc_dep_dry_driver");
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !a || !a->rmol || !a->znt || !a->ust || !a->ivgtyp || !a->height || !(a->dt > 0.0)) return PMCGPU_ERR_ARG;
  const pmcgpu_config& cfg = ctx->cfg;
  const size_t n2 = (size_t)cfg.nx * cfg.ny;
  if (!ctx->dep_buf) {
    CUDA_OK(ctx, cudaMalloc((void**)&ctx->dep_buf, n2 * (3 * sizeof(float) + sizeof(int) + sizeof(double))));
    ctx->allocs.push_back((void*)ctx->dep_buf);
  }
  double* d_height = (double*)ctx->dep_buf;
  float* d_rmol = (float*)(d_height + n2);
  float* d_znt = d_rmol + n2;
  float* d_ust = d_znt + n2;
  int* d_ivg = (int*)(d_ust + n2);
  cudaStream_t st = ctx->stream;
  CUDA_OK(ctx, cudaMemcpyAsync(d_height, a->height, n2 * sizeof(double), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_rmol, a->rmol, n2 * sizeof(float), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_znt, a->znt, n2 * sizeof(float), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_ust, a->ust, n2 * sizeof(float), cudaMemcpyDefault, st));
  CUDA_OK(ctx, cudaMemcpyAsync(d_ivg, a->ivgtyp, n2 * sizeof(int), cudaMemcpyDefault, st));
  {
    ProfScope ps_(ctx, "k_dep_dry");
    const int blocks = (int)((n2 + DEP_WARPS - 1) / DEP_WARPS);
    k_dep_dry<<<blocks, DEP_WARPS * 32, 0, st>>>(ctx->v, cfg.nx, cfg.nz, cfg.ny, cfg.i_offset, cfg.j_offset, cfg.global_nx,
                                                 cfg.global_ny, d_rmol, d_znt, d_ust, d_ivg, d_height, a->dt, a->step);
    LAUNCH_CHECK(ctx);
  }
  if (a->no_wait) return PMCGPU_OK;
  return pmc_first_status(ctx);
}

// unit evaluation for the parity test: in[n][10] = diameter, density, temp, rrho, ustar, rmol, znt, ivgtyp, dt, dz;
// out[n][3] = aer_res, vd, remove_prob
__global__ void k_dep_eval(int n, const double* __restrict__ in, double* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const double* q = in + 10 * k;
  const double ustar = (double)(float)q[4];
  const double aer_res = depvel_partmc((float)q[5], 2.0, (double)(float)q[6], ustar);
  double vd;
  const double p = dep_remove_prob(q[0], q[1], q[2], q[3], ustar, aer_res, dep_zhang_type((int)q[7]), q[8], q[9], &vd);
  out[3 * k + 0] = aer_res;
  out[3 * k + 1] = vd;
  out[3 * k + 2] = p;
}
extern "C" int pmcgpu_eval_dep(pmcgpu_ctx* ctx, int n, const double* in, double* out) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n < 0 || !in || !out) return PMCGPU_ERR_ARG;
  double *di = nullptr, *dout = nullptr;
  CUDA_OK(ctx, cudaMalloc((void**)&di, (size_t)n * 10 * sizeof(double)));
  CUDA_OK(ctx, cudaMalloc((void**)&dout, (size_t)n * 3 * sizeof(double)));
  CUDA_OK(ctx, cudaMemcpyAsync(di, in, (size_t)n * 10 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  k_dep_eval<<<(n + 127) / 128, 128, 0, ctx->stream>>>(n, di, dout);
  LAUNCH_CHECK(ctx);
  CUDA_OK(ctx, cudaMemcpyAsync(out, dout, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(di);
  cudaFree(dout);
  return PMCGPU_OK;
}
