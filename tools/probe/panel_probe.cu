// Where do the cycles of the QR panel chain go?  One warp factors the same 128 x 8 panel over and over (shared memory, as in
// k_qr_dmma); variants remove one ingredient at a time.  Build + run on the GPU box:
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I visual-inertial-odometry-msckf-stereo_b200/csrc \
//        tools/probe/panel_probe.cu -o /tmp/panel_probe && /tmp/panel_probe
#include <cstdio>
#include <cmath>
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

// one panel with T on known data, checked on the host: (I - U T U^T)^T [R; C] must equal [R_new; 0] in the panel columns
__global__ void check_kernel(double* out) {
  extern __shared__ double sm[];
  const int ldr = 52, lane = threadIdx.x;
  double* sR = sm;
  double* sC = sR + 8 * ldr;
  double* sTm = sC + 128 * ldr;
  for (int e = lane; e < 128 * ldr; e += 32) sC[e] = 0.001 * ((e * 37) % 101) - 0.05;
  for (int e = lane; e < 8 * ldr; e += 32) { const int i = e / ldr, c = e % ldr; sR[e] = (c < i) ? 0.0 : ((c == i) ? 2.0 + i : 0.01 * ((e * 13) % 7)); }
  __syncwarp();
  for (int e = lane; e < 8 * 8; e += 32) out[e] = sR[(e / 8) * ldr + e % 8];
  for (int e = lane; e < 128 * 8; e += 32) out[64 + e] = sC[(e / 8) * ldr + e % 8];
  __syncwarp();
  qr_panel<4, true>(sR, ldr, sR, ldr, sC, ldr, sTm, 12, 8, true, lane);
  __syncwarp();
  double* o2 = out + 64 + 1024;
  for (int e = lane; e < 8 * 8; e += 32) o2[e] = sR[(e / 8) * ldr + e % 8];
  for (int e = lane; e < 128 * 8; e += 32) o2[64 + e] = sC[(e / 8) * ldr + e % 8];
  for (int e = lane; e < 64; e += 32) o2[64 + 1024 + e] = sTm[(e / 8) * 12 + e % 8];
}
static void check() {
  double* d;
  const int n = 2 * (64 + 1024) + 64;
  cudaMalloc(&d, n * sizeof(double));
  cudaFuncSetAttribute(check_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  check_kernel<<<1, 32, (8 * 52 + 128 * 52 + 96) * 8>>>(d);
  cudaDeviceSynchronize();
  static double h[2 * (64 + 1024) + 64];
  cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
  const double *R0 = h, *C0 = h + 64, *R1 = h + 1088, *V = h + 1088 + 64, *T = h + 1088 + 64 + 1024;
  // A = [R0; C0] (136 x 8), U = [I; V];  B = A - U T^T (U^T A)
  static double A[136][8], U[136][8], W[8][8], Y[8][8];
  for (int i = 0; i < 8; ++i) for (int c = 0; c < 8; ++c) { A[i][c] = R0[i * 8 + c]; U[i][c] = (i == c) ? 1.0 : 0.0; }
  for (int i = 0; i < 128; ++i) for (int c = 0; c < 8; ++c) { A[8 + i][c] = C0[i * 8 + c]; U[8 + i][c] = V[i * 8 + c]; }
  for (int a = 0; a < 8; ++a) for (int c = 0; c < 8; ++c) { double s = 0; for (int i = 0; i < 136; ++i) s += U[i][a] * A[i][c]; W[a][c] = s; }
  for (int a = 0; a < 8; ++a) for (int c = 0; c < 8; ++c) { double s = 0; for (int l = 0; l < 8; ++l) s += T[l * 8 + a] * W[l][c]; Y[a][c] = s; }
  double eR = 0, eZ = 0, tmax = 0;
  for (int i = 0; i < 136; ++i) for (int c = 0; c < 8; ++c) {
    double b = A[i][c];
    for (int a = 0; a < 8; ++a) b -= U[i][a] * Y[a][c];
    if (i < 8) { const double e = fabs(b - (c >= i ? R1[i * 8 + c] : 0.0)); if (e > eR) eR = e; }
    else if (fabs(b) > eZ) eZ = fabs(b);
  }
  for (int e = 0; e < 64; ++e) if (fabs(T[e]) > tmax) tmax = fabs(T[e]);
  printf("panel check: |Q^T A - R| top %.3e, below %.3e, max |T| %.3e  (%s)\n", eR, eZ, tmax, cudaGetErrorString(cudaGetLastError()));
  cudaFree(d);
}

int main() {
  check();
  for (int warps : {1, 3}) {
    run<0>("full panel, folded reductions, T", warps);
    run<1>("no T, no V write-back", warps);
    run<2>("full panel, plain butterflies, T", warps);
    run<3>("reductions removed (reflectors + T)", warps);
    run<4>("rsqrt / reciprocal removed", warps);
  }
  return 0;
}
