// Where do the cycles of the QR panel chain go?  One warp factors the same 128 x 8 panel over and over (shared memory, as in
// k_qr_dmma); variants remove one ingredient at a time.  Build + run on the GPU box:
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I visual-inertial-odometry-msckf-stereo_b200/csrc \
//        tools/probe/panel_probe.cu -o /tmp/panel_probe && /tmp/panel_probe
#include <cstdio>
#include <cuda_runtime.h>
#include "qr_panel.cuh"

using namespace msckf;

// mode 0: full panel with T   1: no T (need_T = false)   2: plain (unfolded) reductions, with T
// mode 3: reductions replaced by a no-op (wrong numbers, same arithmetic otherwise), with T
// mode 4: rsqrt / reciprocal replaced by cheap arithmetic, with T
template <int MODE>
__device__ __forceinline__ void panel_variant(double* sR, double* sC, double* sTm, int ldr, int lane) {
  if (MODE == 0) qr_panel<4, true>(sR, ldr, sR, ldr, sC, ldr, sTm, 12, 8, true, lane);
  if (MODE == 1) qr_panel<4, true>(sR, ldr, sR, ldr, sC, ldr, sTm, 12, 8, false, lane);
  if (MODE == 2) qr_panel<4, false>(sR, ldr, sR, ldr, sC, ldr, sTm, 12, 8, true, lane);
  if (MODE >= 3) {
    // a copy of the reflector loop of qr_panel with one ingredient removed (T as in the original)
    constexpr int RPL = 4;
    double x[RPL][8];
    for (int t = 0; t < RPL; ++t)
      for (int c = 0; c < 8; ++c) x[t][c] = sC[(lane + 32 * t) * ldr + c];
    double tau[8];
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
      double s[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        double a = 0.0;
        if (c >= jj)
          for (int t = 0; t < RPL; ++t) a += x[t][jj] * x[t][c];
        s[c] = a;
      }
      double rrow[8];
#pragma unroll
      for (int c = jj; c < 8; ++c) rrow[c] = sR[jj * ldr + c];
      if (MODE != 3) {
        if (jj < 4) warp_allsum8(s, lane);
        else
          for (int o = 16; o > 0; o >>= 1)
            for (int c = jj; c < 8; ++c) s[c] += __shfl_xor_sync(WR_FULL, s[c], o);
      }
      const double alpha = rrow[jj];
      const double n2 = alpha * alpha + s[jj] + 1.0;
      double rn, sc;
      if (MODE == 4) { rn = n2 * 0.25 + 0.5; sc = (alpha + rn) * 0.125; }
      else { rn = rsqrt(n2); }
      const double nrm = n2 * rn;
      const double beta = (alpha >= 0) ? -nrm : nrm;
      if (MODE != 4) sc = 1.0 / (alpha - beta);
      const double tj = (beta - alpha) * ((alpha >= 0) ? -rn : rn);
      tau[jj] = tj;
      for (int t = 0; t < RPL; ++t) x[t][jj] *= sc;
#pragma unroll
      for (int c = jj + 1; c < 8; ++c) {
        const double w = tj * (rrow[c] + s[c] * sc);
        if (lane == 0) sR[jj * ldr + c] = rrow[c] - w;
        for (int t = 0; t < RPL; ++t) x[t][c] -= w * x[t][jj];
      }
      if (lane == 0) sR[jj * ldr + jj] = beta;
    }
    for (int t = 0; t < RPL; ++t)
      for (int c = 0; c < 8; ++c) sC[(lane + 32 * t) * ldr + c] = x[t][c];
    double gv[32];
    for (int i = 0; i < 32; ++i) gv[i] = 0.0;
#pragma unroll
    for (int j = 1; j < 8; ++j)
#pragma unroll
      for (int l = 0; l < j; ++l) {
        double a = 0.0;
        for (int t = 0; t < RPL; ++t) a += x[t][l] * x[t][j];
        gv[j * (j - 1) / 2 + l] = a;
      }
    if (MODE != 3) warp_allsum32(gv, lane);
    double trow[8];
    const int ri = lane & 7;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      double acc = 0.0;
#pragma unroll
      for (int l = 0; l < j; ++l) acc += trow[l] * gv[j * (j - 1) / 2 + l];
      trow[j] = (ri == j) ? tau[j] : ((ri < j) ? -tau[j] * acc : 0.0);
    }
    if (lane < 8)
      for (int j = 0; j < 8; ++j) sTm[ri * 12 + j] = trow[j];
  }
}

template <int MODE>
__global__ void probe(long long* out, int reps, int nwarps_active) {
  extern __shared__ double sm[];
  const int ldr = 52, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* sR = sm + warp * (8 * ldr + 128 * ldr + 96);
  double* sC = sR + 8 * ldr;
  double* sTm = sC + 128 * ldr;
  for (int e = lane; e < 128 * ldr; e += 32) sC[e] = 0.001 * ((e * 37 + warp) % 101) - 0.05;
  for (int e = lane; e < 8 * ldr; e += 32) sR[e] = (e % (ldr + 1) == 0) ? 2.0 : 0.01 * (e % 7);
  __syncwarp();
  long long t0 = clock64();
  if (warp < nwarps_active)
    for (int r = 0; r < reps; ++r) {
      panel_variant<MODE>(sR, sC, sTm, ldr, lane);
      __syncwarp();
      if (MODE != 1) {   // keep the data bounded: rescale the chunk
        for (int e = lane; e < 128 * 8; e += 32) { const int i = e / 8, c = e % 8; sC[i * ldr + c] = 0.001 * ((e * 37 + r) % 101) - 0.05; }
      }
      __syncwarp();
    }
  long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = (t1 - t0) / reps;
}

template <int MODE>
static void run(const char* name, int warps) {
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  const int per_warp = (8 * 52 + 128 * 52 + 96) * 8;
  cudaFuncSetAttribute(probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  probe<MODE><<<148, 32 * warps, per_warp * warps>>>(d, 200, warps);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
  printf("%-44s warps/SM %d: %lld cycles per panel  (%s)\n", name, warps, h[0], cudaGetErrorString(cudaGetLastError()));
  cudaFree(d);
}

int main() {
  for (int warps : {1, 3}) {
    run<0>("full panel, folded reductions, T", warps);
    run<1>("no T, no V write-back", warps);
    run<2>("full panel, plain butterflies, T", warps);
    run<3>("reductions removed (reflectors + T)", warps);
    run<4>("rsqrt / reciprocal removed", warps);
  }
  return 0;
}
