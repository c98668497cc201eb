// Batched warp all-reductions for the Householder panels of the QR kernels.
//
// A panel step needs the sums over the 32 lanes of several per-lane partial dot products at once (one reflector: the
// column's own norm and its dots with the columns right of it; the compact-WY factor: all 28 dots of V^T V).  A plain
// butterfly moves every value through all five stages: 5 n shuffles for n values.  Here the values are folded while
// they travel -- at every stage a lane keeps half of its values and hands the other half to its partner -- so that
// after the five stages each total sits in a known lane and one more shuffle per value broadcasts it:
// about n + n/2 + n/4 + ... + n = 3 n shuffles' worth of 64-bit traffic instead of 5 n, in six dependent steps.
// The additions pair the same lanes in the same order as the plain butterfly (xor 16, 8, 4, 2, 1), so the sums are
// bit-identical to it.
#pragma once
#include <cuda_runtime.h>

namespace msckf {

constexpr unsigned WR_FULL = 0xffffffffu;

// One folding stage: 2 H values in, H out.  Lanes whose `bit` is clear keep v[0..H), the others keep v[H..2H).
template <int H>
__device__ __forceinline__ void wr_fold(const double* v, double* o, int dist, bool up) {
#pragma unroll
  for (int k = 0; k < H; ++k) {
    const double snd = up ? v[k] : v[k + H];
    const double keep = up ? v[k + H] : v[k];
    o[k] = keep + __shfl_xor_sync(WR_FULL, snd, dist);
  }
}

// s[0..8) <- sum over the warp of every s[c], on every lane.
__device__ __forceinline__ void warp_allsum8(double* s, int lane) {
  double t[4], u[2], w;
  wr_fold<4>(s, t, 16, (lane & 16) != 0);
  wr_fold<2>(t, u, 8, (lane & 8) != 0);
  {
    const bool up = (lane & 4) != 0;
    const double snd = up ? u[0] : u[1], keep = up ? u[1] : u[0];
    w = keep + __shfl_xor_sync(WR_FULL, snd, 4);
  }
  w += __shfl_xor_sync(WR_FULL, w, 2);
  w += __shfl_xor_sync(WR_FULL, w, 1);
  // the total of value c now sits in lanes 4 c .. 4 c + 3
#pragma unroll
  for (int c = 0; c < 8; ++c) s[c] = __shfl_sync(WR_FULL, w, 4 * c);
}

// s[0..n) <- sum over the warp of every s[c], on every lane, for n <= 16 values (s has 16 slots; the unused ones are zero).
__device__ __forceinline__ void warp_allsum16(double* s, int n, int lane) {
  double a[8], b[4], c[2], w;
  wr_fold<8>(s, a, 16, (lane & 16) != 0);
  wr_fold<4>(a, b, 8, (lane & 8) != 0);
  wr_fold<2>(b, c, 4, (lane & 4) != 0);
  {
    const bool up = (lane & 2) != 0;
    const double snd = up ? c[0] : c[1], keep = up ? c[1] : c[0];
    w = keep + __shfl_xor_sync(WR_FULL, snd, 2);
  }
  w += __shfl_xor_sync(WR_FULL, w, 1);
  // the total of value i sits in lanes 2 i, 2 i + 1
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if (i < n) s[i] = __shfl_sync(WR_FULL, w, 2 * i);
}

// v[0..32) <- sum over the warp of every v[i], on every lane (callers pad unused slots with zeros; the folding makes
// 32 values cost what 8 cost with a plain butterfly).
__device__ __forceinline__ void warp_allsum32(double* v, int lane) {
  double a[16], b[8], c[4], d[2], w;
  wr_fold<16>(v, a, 16, (lane & 16) != 0);
  wr_fold<8>(a, b, 8, (lane & 8) != 0);
  wr_fold<4>(b, c, 4, (lane & 4) != 0);
  wr_fold<2>(c, d, 2, (lane & 2) != 0);
  {
    const bool up = (lane & 1) != 0;
    const double snd = up ? d[0] : d[1], keep = up ? d[1] : d[0];
    w = keep + __shfl_xor_sync(WR_FULL, snd, 1);
  }
  // value i = 16 b4 + 8 b3 + 4 b2 + 2 b1 + b0 ended in the lane with exactly those bits: lane i
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __shfl_sync(WR_FULL, w, i);
}

}  // namespace msckf
