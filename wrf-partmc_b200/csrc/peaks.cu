// peaks.cu -- FP64 FMA throughput of the device, measured in place.  MEASURED_PEAKS.json holds the HBM
// copy bandwidth and the bf16 GEMM rate only; the condensation and coagulation kernels are bounded by the
// FP64 pipe, so bench.py takes their roofline denominator from this microbenchmark.
#include "pmcgpu_internal.cuh"

__global__ void __launch_bounds__(256) k_dfma_peak(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 123.456) out[0] = s;  // never true: keeps the chain alive
}

// tflops = 2 * FMAs / best-of-5 event time, FP64
extern "C" int pmcgpu_measure_fp64_peak(pmcgpu_ctx* ctx, double* tflops) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || !tflops) return PMCGPU_ERR_ARG;
  cudaDeviceProp pr;
  CUDA_OK(ctx, cudaGetDeviceProperties(&pr, ctx->cfg.device));
  const int blocks = pr.multiProcessorCount * 8, iters = 20000;
  cudaEvent_t a, b;
  CUDA_OK(ctx, cudaEventCreate(&a));
  CUDA_OK(ctx, cudaEventCreate(&b));
  double best = 0.0;
  for (int r = 0; r < 6; r++) {
    CUDA_OK(ctx, cudaEventRecord(a, ctx->stream));
    k_dfma_peak<<<blocks, 256, 0, ctx->stream>>>((double*)ctx->v.stage, iters, 0.999999, 1e-9);
    LAUNCH_CHECK(ctx);
    CUDA_OK(ctx, cudaEventRecord(b, ctx->stream));
    CUDA_OK(ctx, cudaEventSynchronize(b));
    float ms = 0.f;
    CUDA_OK(ctx, cudaEventElapsedTime(&ms, a, b));
    const double tf = 2.0 * 8.0 * (double)iters * 256.0 * (double)blocks / (ms * 1e-3) / 1e12;
    if (r > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  *tflops = best;
  return PMCGPU_OK;
}
