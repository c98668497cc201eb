// Synthetic EuRoC-MH_01-shaped message generator on the device: Monte-Carlo noise realisations of
// one noise-free stream, written straight into the argument layout of msckf_feature_callback_dev.  Filter f of a batch
// uses the noise stream of seed (seed0 + f); every random number is a pure function of (seed, frame / sample index,
// element), so any subset of filters or frames can be regenerated anywhere (counter-based: Philox4x32-10).
//
// The noise model is the one of the host generator (synth.py): white gyro / accelerometer noise on top of the
// clean signal + constant biases; pixel noise sigma_z on every coordinate plus, for a fraction of the observations, a
// gross outlier of sigma_out on all four coordinates.  HBM bound: 32 B written per observation, 48 B per IMU sample.
#include <cstring>
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

struct Philox {
  uint32_t k0, k1;
  __device__ __forceinline__ Philox(unsigned long long seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  // 128 random bits for the 128-bit counter (c0..c3)
  __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ a; c1 = lo1; c2 = hi0 ^ c3 ^ b; c3 = lo0;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};
__device__ __forceinline__ double u01(uint32_t x) { return ((double)x + 0.5) * (1.0 / 4294967296.0); }   // (0, 1)
// two independent standard normals from two 32-bit words (Box-Muller)
__device__ __forceinline__ void normal2(uint32_t x, uint32_t y, double& n0, double& n1) {
  const double r = sqrt(-2.0 * log(u01(x)));
  double s, c;
  sincospi(2.0 * u01(y), &s, &c);
  n0 = r * c;
  n1 = r * s;
}

struct SynthArgs {
  int F, K, n;
  unsigned long long seed0;
  long long frame, sample0, cam_id;
  double gyro_bias[3], acc_bias[3], sigma_g, sigma_a, sigma_z, outlier_frac, outlier_sigma;
  const double* ts;          // [K]       device copies of the frame's clean structure
  const double* gyro_clean;  // [K][3]
  const double* acc_clean;   // [K][3]
  const long long* ids;      // [n]
  const double* uv_clean;    // [n][4]
  int* imu_off; double* ts_out; double* gyro; double* acc; long long* cam_ids; int* feat_off; long long* ids_out; double* uv;
};

// stream ids inside the counter keep the IMU and the pixel noise of a seed apart
constexpr uint32_t STREAM_IMU = 1u, STREAM_UV = 2u, STREAM_OUT = 3u;

__global__ void __launch_bounds__(256) k_synth_frame(SynthArgs a) {
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
  // offsets / ids shared by the filters of a lock-step batch
  for (long long e = tid; e <= a.F; e += stride) {
    a.imu_off[e] = (int)(e * a.K);
    a.feat_off[e] = (int)(e * a.n);
    if (e < a.F) a.cam_ids[e] = a.cam_id;
  }
  // IMU samples: one thread per (filter, sample)
  for (long long e = tid; e < (long long)a.F * a.K; e += stride) {
    const int f = (int)(e / a.K), k = (int)(e - (long long)f * a.K);
    const Philox rng(a.seed0 + (unsigned long long)f);
    const unsigned long long s = (unsigned long long)(a.sample0 + k);
    const uint4 r0 = rng((uint32_t)s, (uint32_t)(s >> 32), 0u, STREAM_IMU), r1 = rng((uint32_t)s, (uint32_t)(s >> 32), 1u, STREAM_IMU);
    double g0, g1, g2, g3, g4, g5;
    normal2(r0.x, r0.y, g0, g1);
    normal2(r0.z, r0.w, g2, g3);
    normal2(r1.x, r1.y, g4, g5);
    a.ts_out[e] = a.ts[k];
    a.gyro[3 * e + 0] = a.gyro_clean[3 * k + 0] + a.gyro_bias[0] + a.sigma_g * g0;
    a.gyro[3 * e + 1] = a.gyro_clean[3 * k + 1] + a.gyro_bias[1] + a.sigma_g * g1;
    a.gyro[3 * e + 2] = a.gyro_clean[3 * k + 2] + a.gyro_bias[2] + a.sigma_g * g2;
    a.acc[3 * e + 0] = a.acc_clean[3 * k + 0] + a.acc_bias[0] + a.sigma_a * g3;
    a.acc[3 * e + 1] = a.acc_clean[3 * k + 1] + a.acc_bias[1] + a.sigma_a * g4;
    a.acc[3 * e + 2] = a.acc_clean[3 * k + 2] + a.acc_bias[2] + a.sigma_a * g5;
  }
  // observations: one thread per (filter, feature); the noise of a landmark is keyed by its persistent id
  for (long long e = tid; e < (long long)a.F * a.n; e += stride) {
    const int f = (int)(e / a.n), i = (int)(e - (long long)f * a.n);
    const Philox rng(a.seed0 + (unsigned long long)f);
    const unsigned long long id = (unsigned long long)a.ids[i];
    const uint4 r0 = rng((uint32_t)id, (uint32_t)(id >> 32) ^ ((uint32_t)a.frame << 8), (uint32_t)(a.frame >> 24), STREAM_UV);
    double z0, z1, z2, z3;
    normal2(r0.x, r0.y, z0, z1);
    normal2(r0.z, r0.w, z2, z3);
    double4 v = *(const double4*)(a.uv_clean + 4 * (size_t)i);
    v.x += a.sigma_z * z0; v.y += a.sigma_z * z1; v.z += a.sigma_z * z2; v.w += a.sigma_z * z3;
    const uint4 r1 = rng((uint32_t)id, (uint32_t)(id >> 32) ^ ((uint32_t)a.frame << 8), (uint32_t)(a.frame >> 24), STREAM_OUT);
    if (u01(r1.x) < a.outlier_frac) {
      const uint4 r2 = rng((uint32_t)id, (uint32_t)(id >> 32) ^ ((uint32_t)a.frame << 8), (uint32_t)(a.frame >> 24) | 0x80000000u, STREAM_OUT);
      double o0, o1, o2, o3;
      normal2(r2.x, r2.y, o0, o1);
      normal2(r2.z, r2.w, o2, o3);
      v.x += a.outlier_sigma * o0; v.y += a.outlier_sigma * o1; v.z += a.outlier_sigma * o2; v.w += a.outlier_sigma * o3;
    }
    a.ids_out[e] = (long long)id;
    *(double4*)(a.uv + 4 * (size_t)e) = v;
  }
}

}  // namespace msckf

using namespace msckf;

extern "C" int msckf_synth_frame(int32_t device, int32_t n_filters, uint64_t seed0, int64_t frame, int32_t n_imu, int64_t sample0,
                                 const double* ts, const double* gyro_clean, const double* acc_clean, const double* gyro_bias,
                                 const double* acc_bias, double sigma_gyro, double sigma_acc, int32_t n_features, const int64_t* ids,
                                 const double* uv_clean, double sigma_z, double outlier_frac, double outlier_sigma, int64_t cam_id,
                                 int32_t* imu_off, double* ts_out, double* gyro, double* acc, int64_t* cam_ids, int32_t* feat_off,
                                 int64_t* ids_out, double* uv) {
  if (n_filters < 1 || n_imu < 0 || n_features < 0 || !imu_off || !cam_ids || !feat_off || !gyro_bias || !acc_bias) return MSCKF_ERR_ARG;
  if ((n_imu > 0 && !(ts && gyro_clean && acc_clean && ts_out && gyro && acc)) || (n_features > 0 && !(ids && uv_clean && ids_out && uv)))
    return MSCKF_ERR_ARG;
  if ((long long)n_filters * n_imu >= (1LL << 31) || (long long)n_filters * n_features >= (1LL << 31)) return MSCKF_ERR_ARG;
  if (cudaSetDevice(device) != cudaSuccess) return MSCKF_ERR_CUDA;
  // the frame's clean structure (a few KB) goes to the device in one staged copy
  const size_t nb_imu = (size_t)n_imu * 7 * sizeof(double), nb_feat = (size_t)n_features * (sizeof(int64_t) + 4 * sizeof(double));
  char* dbuf = nullptr;
  if (cudaMalloc((void**)&dbuf, nb_imu + nb_feat + 64) != cudaSuccess) return MSCKF_ERR_CUDA;
  std::vector<char> h(nb_imu + nb_feat + 64);
  size_t o = 0;
  auto put = [&](const void* src, size_t bytes) { const size_t at = o; if (bytes) memcpy(h.data() + at, src, bytes); o += bytes; return at; };
  const size_t o_uv = put(uv_clean, (size_t)n_features * 4 * sizeof(double));      // first: 32-byte aligned rows for the double4 loads
  const size_t o_ts = put(ts, (size_t)n_imu * sizeof(double)), o_g = put(gyro_clean, (size_t)n_imu * 3 * sizeof(double));
  const size_t o_a = put(acc_clean, (size_t)n_imu * 3 * sizeof(double)), o_id = put(ids, (size_t)n_features * sizeof(int64_t));
  int rc = MSCKF_OK;
  if (cudaMemcpy(dbuf, h.data(), o, cudaMemcpyHostToDevice) != cudaSuccess) rc = MSCKF_ERR_CUDA;
  if (!rc) {
    SynthArgs a;
    a.F = n_filters; a.K = n_imu; a.n = n_features; a.seed0 = seed0; a.frame = frame; a.sample0 = sample0; a.cam_id = cam_id;
    for (int i = 0; i < 3; ++i) { a.gyro_bias[i] = gyro_bias[i]; a.acc_bias[i] = acc_bias[i]; }
    a.sigma_g = sigma_gyro; a.sigma_a = sigma_acc; a.sigma_z = sigma_z; a.outlier_frac = outlier_frac; a.outlier_sigma = outlier_sigma;
    a.uv_clean = (const double*)(dbuf + o_uv); a.ts = (const double*)(dbuf + o_ts); a.gyro_clean = (const double*)(dbuf + o_g);
    a.acc_clean = (const double*)(dbuf + o_a); a.ids = (const long long*)(dbuf + o_id);
    a.imu_off = imu_off; a.ts_out = ts_out; a.gyro = gyro; a.acc = acc; a.cam_ids = (long long*)cam_ids; a.feat_off = feat_off;
    a.ids_out = (long long*)ids_out; a.uv = uv;
    const long long work = std::max<long long>((long long)n_filters * std::max(n_features, n_imu), n_filters + 1);
    const int grid = (int)std::min<long long>((work + 255) / 256, 148LL * 16);
    k_synth_frame<<<grid, 256>>>(a);
    if (cudaDeviceSynchronize() != cudaSuccess || cudaGetLastError() != cudaSuccess) rc = MSCKF_ERR_CUDA;
  }
  cudaFree(dbuf);
  return rc;
}
