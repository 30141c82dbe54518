// kmax.cu -- per-cell coagulation-kernel majorant tables.
//
// est_k_minmax_binned_unweighted (partmc/src/coag_kernel.F90:253-280) ->
// est_k_minmax_for_bin_unweighted (:286-341: 3x3 linear-in-volume samples of
// the two bins, over_scale = 2) -> kernel_brown_minmax
// (coag_kernel_brown.F90:58-101: 3x3 densities in [min rho, max rho]); the additive, constant, zero and
// sedimentation (coag_kernel_sedi.F90:52-73, sedi.cuh) kernels take the generic 3x3 volume samples.
// Like the reference, a cell's table is rebuilt only when its bin grid was
// remade (coag_kernel_bounds_valid, aero_sorted.F90:132, coagulation.F90:64).
//
// One CTA per stale cell.  The Brownian kernel separates into a
// single-particle part (radius, diffusivity, speed^2, delta^2 -- all the
// exp/pow/sqrt work) and a cheap pair combination, so the CTA first fills a
// shared table of nb x 3 x 3 single-particle terms and then takes the
// 81-sample maxima from shared memory: nb*9 heavy evaluations instead of
// nb^2*81.  FP64-pipe bound; k_min is never read on the path and is not kept.
#include "pmcgpu_internal.cuh"
#include "sedi.cuh"

#define KMAX_THREADS 256
static const double BETA_1 = 1000.0;                 // coag_kernel_additive.F90:23
static const double BETA_0 = 0.25 / (60.0 * 2e8);    // coag_kernel_constant.F90:22

__global__ void __launch_bounds__(KMAX_THREADS) k_kernel_bounds(DevView v, int ktype) {
  const int cell = blockIdx.x;
  const int nb = v.grid_nb[cell];
  if (nb < 1 || v.bounds_valid[cell]) return;
  extern __shared__ double s_tab[];  // BrownPart[nb][3][3] (4 doubles each) or vol samples [nb][3]
  const double gmin = v.grid_min[cell], gmax = v.grid_max[cell];
  const int nd = (v.dens_min == v.dens_max) ? 1 : 3;
  double* kout = v.kmax + (size_t)cell * v.nbcap * v.nbcap;
  if (ktype == PMCGPU_KERNEL_BROWN) {
    const BrownEnv e = brown_env(v.temp[cell], v.pres[cell]);
    for (int q = threadIdx.x; q < nb * 3 * nd; q += blockDim.x) {
      const int b = q / (3 * nd) + 1, i = (q / nd) % 3 + 1, j = q % nd + 1;
      const double v_low = rad2vol(grid_edge(gmin, gmax, nb, b));
      const double v_high = rad2vol(grid_edge(gmin, gmax, nb, b + 1));
      const double vol = interp_linear_disc(v_low, v_high, 3, i);
      const double den = interp_linear_disc(v.dens_min, v.dens_max, 3, j);
      const BrownPart bp = brown_part(e, vol, den);
      s_tab[4 * q + 0] = bp.rad;
      s_tab[4 * q + 1] = bp.diffus;
      s_tab[4 * q + 2] = bp.speedsq;
      s_tab[4 * q + 3] = bp.deltasq;
    }
    __syncthreads();
    const int ns = 3 * nd;
    for (int q = threadIdx.x; q < nb * nb; q += blockDim.x) {
      const int b1 = q % nb, b2 = q / nb;
      double kmax = 0.0;
      for (int s1 = 0; s1 < ns; s1++) {
        const double* p1 = s_tab + 4 * (b1 * ns + s1);
        const BrownPart a{p1[0], p1[1], p1[2], p1[3]};
        for (int s2 = 0; s2 < ns; s2++) {
          const double* p2 = s_tab + 4 * (b2 * ns + s2);
          const BrownPart b{p2[0], p2[1], p2[2], p2[3]};
          kmax = fmax(kmax, brown_pair(a, b));
        }
      }
      kout[(size_t)b2 * nb + b1] = kmax * 2.0;  // over_scale
    }
  } else {
    for (int q = threadIdx.x; q < nb * nb; q += blockDim.x) {
      const int b1 = q % nb + 1, b2 = q / nb + 1;
      double kmax = 0.0;
      if (ktype == PMCGPU_KERNEL_ADDITIVE || ktype == PMCGPU_KERNEL_SEDI || ktype == PMCGPU_KERNEL_BROWN_FREE ||
          ktype == PMCGPU_KERNEL_BROWN_CONT) {
        // aero_data_rad2vol of the bin edges (coag_kernel.F90:316-321): fractal particles after pmcgpu_set_fractal
        const double v1l = fr_rad2vol(v, grid_edge(gmin, gmax, nb, b1)), v1h = fr_rad2vol(v, grid_edge(gmin, gmax, nb, b1 + 1));
        const double v2l = fr_rad2vol(v, grid_edge(gmin, gmax, nb, b2)), v2h = fr_rad2vol(v, grid_edge(gmin, gmax, nb, b2 + 1));
        const double temp = v.temp[cell];
        for (int i = 1; i <= 3; i++)
          for (int j = 1; j <= 3; j++) {
            const double v1 = interp_linear_disc(v1l, v1h, 3, i);
            const double v2 = interp_linear_disc(v2l, v2h, 3, j);
            double k;
            if (ktype == PMCGPU_KERNEL_SEDI) {
              k = kernel_sedi(v1, v2);
            } else if (ktype == PMCGPU_KERNEL_ADDITIVE) {
              k = BETA_1 * (v1 + v2);
            } else if (ktype == PMCGPU_KERNEL_BROWN_CONT) {
              k = kernel_brown_cont(v, v1, v2, temp);  // independent of the densities (coag_kernel_brown_cont.F90:62-107)
            } else {
              // kernel_brown_free_minmax (coag_kernel_brown_free.F90:62-108): 3 x 3 densities in [min rho, max rho]
              k = 0.0;
              for (int a = 1; a <= nd; a++)
                for (int b = 1; b <= nd; b++)
                  k = fmax(k, kernel_brown_free(v, v1, interp_linear_disc(v.dens_min, v.dens_max, 3, a), v2,
                                                interp_linear_disc(v.dens_min, v.dens_max, 3, b), temp));
            }
            kmax = (i == 1 && j == 1) ? k : fmax(kmax, k);
          }
      } else if (ktype == PMCGPU_KERNEL_CONSTANT) {
        kmax = BETA_0;
      } else {
        kmax = 0.0;
      }
      kout[(size_t)(b2 - 1) * nb + (b1 - 1)] = kmax * 2.0;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) v.bounds_valid[cell] = 1;
}

int pmc_kernel_bounds_async(pmcgpu_ctx* ctx, int ktype) {
  DevView& v = ctx->v;
  if (ktype != PMCGPU_KERNEL_BROWN && ktype != PMCGPU_KERNEL_ADDITIVE && ktype != PMCGPU_KERNEL_CONSTANT &&
      ktype != PMCGPU_KERNEL_ZERO && ktype != PMCGPU_KERNEL_SEDI && ktype != PMCGPU_KERNEL_BROWN_FREE &&
      ktype != PMCGPU_KERNEL_BROWN_CONT) {
    ctx->err = "unknown kernel type";  // coag_kernel.F90:121 die_msg(200724934)
    return 200724934;
  }
  if (v.fr_on && (ktype == PMCGPU_KERNEL_BROWN || ktype == PMCGPU_KERNEL_SEDI)) {
    ctx->err = "the Brownian and sedimentation kernels take the mobility radius of fractal particles "
               "(fractal.F90:315-362), which is not built: use brown_free / brown_cont / additive / constant";
    return PMCGPU_ERR_ARG;
  }
  const size_t smem = (size_t)v.nbcap * 9 * 4 * sizeof(double);
  if (smem > 48 * 1024 && smem > ctx->cfg_kmax_smem) {
    CUDA_OK(ctx, cudaFuncSetAttribute(k_kernel_bounds, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctx->cfg_kmax_smem = smem;
  }
  // the per-cell tables belong to one kernel type: a call with another type rebuilds every cell's table
  if (ctx->kmax_ktype != ktype) {
    CUDA_OK(ctx, cudaMemsetAsync(v.bounds_valid, 0, v.n_cells * sizeof(int), ctx->stream));
    ctx->kmax_ktype = ktype;
  }
  {
    ProfScope ps_(ctx, "k_kernel_bounds");
    k_kernel_bounds<<<v.n_cells, KMAX_THREADS, smem, ctx->stream>>>(v, ktype);
    LAUNCH_CHECK(ctx);
  }
  return PMCGPU_OK;
}

extern "C" int pmcgpu_kernel_bounds(pmcgpu_ctx* ctx, int coag_kernel_type) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  int rc = pmc_kernel_bounds_async(ctx, coag_kernel_type);
  if (rc) return rc;
  CUDA_OK(ctx, cudaStreamSynchronize(ctx->stream));
  return PMCGPU_OK;
}

// aero_data%fractal (fractal.F90:55-66; spec file: do_fractal, frac_dim, prime_radius, vol_fill_factor).  The bin
// grid and the majorant tables are functions of the geometric radius, so every cell gets a new grid at its next sort.
extern "C" int pmcgpu_set_fractal(pmcgpu_ctx* ctx, double frac_dim, double prime_radius, double vol_fill_factor) {
  if (!ctx) return PMCGPU_ERR_ARG;
  if (cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;
  if (!(frac_dim >= 1.0 && frac_dim <= 3.0) || !(prime_radius > 0.0) || !(vol_fill_factor >= 1.0)) {
    ctx->err = "pmcgpu_set_fractal: 1 <= frac_dim <= 3, prime_radius > 0, vol_fill_factor >= 1";
    return PMCGPU_ERR_ARG;
  }
  DevView& v = ctx->v;
  v.fr_df = frac_dim;
  v.fr_r0 = prime_radius;
  v.fr_fill = vol_fill_factor;
  v.fr_on = !(frac_dim == 3.0 && vol_fill_factor == 1.0);  // fractal_is_spherical (fractal.F90:296-307)
  CUDA_OK(ctx, cudaMemsetAsync(v.grid_nb, 0, v.n_cells * sizeof(int), ctx->stream));
  CUDA_OK(ctx, cudaMemsetAsync(v.bounds_valid, 0, v.n_cells * sizeof(int), ctx->stream));
  return PMCGPU_OK;
}
