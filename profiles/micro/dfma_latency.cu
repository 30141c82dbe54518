// DFMA issue model on one SM: throughput of dependent chains as a function of warps per scheduler and independent
// chains per warp (ILP).  nvcc -O3 -gencode arch=compute_100a,code=sm_100a dfma_latency.cu -o dfma_latency
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void k(double* out, int iters, double a, double b) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) x[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int r = 0; r < 16; r++) {
#pragma unroll
      for (int i = 0; i < ILP; i++) x[i] = fma(x[i], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
void run(int warps_per_sm, double* out) {
  const int iters = 4096;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k<ILP><<<148, warps_per_sm * 32>>>(out, 16, 1.0000001, 1e-9);
  cudaEventRecord(e0);
  k<ILP><<<148, warps_per_sm * 32>>>(out, iters, 1.0000001, 1e-9);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double n_warp_instr_per_sched = (double)iters * 16 * ILP * warps_per_sm / 4.0;
  int clk;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double cycles = ms * 1e-3 * clk * 1e3;
  printf("warps/sched %d ILP %d : %.2f cycles per DFMA warp-instr per scheduler (chain latency ~ %.1f cycles), %.1f TFLOP/s\n",
         warps_per_sm / 4, ILP, cycles / n_warp_instr_per_sched, cycles / (iters * 16.0),
         2.0 * iters * 16.0 * ILP * warps_per_sm * 32 * 148 / (ms * 1e-3) / 1e12);
}
int main() {
  double* out;
  cudaMalloc(&out, 148 * 1024 * sizeof(double));
  for (int w : {4, 8, 16, 32}) {
    run<1>(w, out);
    run<2>(w, out);
    run<4>(w, out);
    run<8>(w, out);
  }
  return 0;
}
