// FP64 peak microbenchmarks (roofline denominators for the FP64 kernels: MEASURED_PEAKS.json only
// carries HBM and bf16).  Two kernels, both register resident:
//   * DFMA: 8 independent fused-multiply-add chains per thread (vector FP64 pipe),
//   * DMMA: mma.sync.aligned.m8n8k4 f64 with 4 independent accumulator tiles per warp (SASS DMMA.8x8x4,
//     the native sm_100a FP64 tensor op).
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/msckf_b200.h"

namespace {

__global__ void __launch_bounds__(256) k_peak_dfma(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 4
  for (int i = 0; i < iters; ++i) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;     // never true; keeps the chains alive
}

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256) k_peak_dmma(double* out, int iters, double a, double b) {
  double c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const double av = a + threadIdx.x * 1e-9, bv = b - threadIdx.x * 1e-9;
#pragma unroll 2
  for (int i = 0; i < iters; ++i) {
    dmma(c[0], c[1], av, bv);
    dmma(c[2], c[3], av, bv);
    dmma(c[4], c[5], av, bv);
    dmma(c[6], c[7], av, bv);
  }
  const double s = c[0] + c[1] + c[2] + c[3] + c[4] + c[5] + c[6] + c[7];
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace

extern "C" int msckf_fp64_peak(int32_t device, double* dfma_tflops, double* dmma_tflops) {
  if (!dfma_tflops || !dmma_tflops) return MSCKF_ERR_ARG;
  if (cudaSetDevice(device) != cudaSuccess) return MSCKF_ERR_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MSCKF_ERR_CUDA;
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
  double* out = nullptr;
  if (cudaMalloc((void**)&out, sizeof(double) * blocks * threads) != cudaSuccess) return MSCKF_ERR_CUDA;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best[2] = {0, 0};
  for (int which = 0; which < 2; ++which) {
    for (int rep = 0; rep < 5; ++rep) {          // first repetition is the warm-up
      cudaEventRecord(e0);
      if (which == 0) k_peak_dfma<<<blocks, threads>>>(out, iters, 0.999999, 1e-7);
      else k_peak_dmma<<<blocks, threads>>>(out, iters, 0.999999, 1e-7);
      cudaEventRecord(e1);
      if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return MSCKF_ERR_CUDA; }
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double flops = which == 0 ? 2.0 * 8 * iters * (double)blocks * threads
                                      : 512.0 * 4 * iters * (double)blocks * (threads / 32);
      const double tf = flops / (ms * 1e-3) / 1e12;
      if (rep > 0 && tf > best[which]) best[which] = tf;
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  *dfma_tflops = best[0];
  *dmma_tflops = best[1];
  return cudaGetLastError() == cudaSuccess ? MSCKF_OK : MSCKF_ERR_CUDA;
}
