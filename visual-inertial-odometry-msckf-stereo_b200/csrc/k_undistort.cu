// Point undistortion + normalisation of the stereo feature message: the step right in front of the filter
// (reference MSCKF/image.py:406-438 `publish`, :640-674 `undistort_points`, which call cv2.undistortPoints /
// cv2.fisheye.undistortPoints).  One thread per point, FP64, 16 B in / 16 B out per point and camera: HBM bound.
//   radtan       five fixed-point iterations x <- (x0 - delta(x, y)) / (1 + k1 r2 + k2 r2^2)   (OpenCV default criteria)
//   equidistant  Newton on theta (1 + k1 theta^2 + ... + k4 theta^8) = theta_d, at most 10 steps, |step| < 1e-8
// then the rectification rotation and the new camera matrix (RR = K_new R).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "../../include/msckf_b200.h"

namespace {

struct UndistortJob {
  double fx, fy, cx, cy;
  double k[4];
  double RR[9];
  int model;
};

__global__ void __launch_bounds__(256) k_undistort(UndistortJob job, int n, const double2* __restrict__ in, double* __restrict__ out,
                                                    int out_stride, int out_off) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double2 p = in[i];
    double x, y;
    bool ok = true;
    if (job.model == 1) {
      const double pwx = (p.x - job.cx) / job.fx, pwy = (p.y - job.cy) / job.fy;
      double theta_d = sqrt(pwx * pwx + pwy * pwy);
      theta_d = fmin(fmax(-M_PI / 2., theta_d), M_PI / 2.);
      bool converged = false;
      double theta = theta_d, scale = 0.0;
      if (fabs(theta_d) > 1e-8) {
        for (int j = 0; j < 10; ++j) {
          const double t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t6 * t2;
          const double k0 = job.k[0] * t2, k1 = job.k[1] * t4, k2 = job.k[2] * t6, k3 = job.k[3] * t8;
          const double fix = (theta * (1 + k0 + k1 + k2 + k3) - theta_d) / (1 + 3 * k0 + 5 * k1 + 7 * k2 + 9 * k3);
          theta = theta - fix;
          if (fabs(fix) < 1e-8) { converged = true; break; }
        }
        scale = tan(theta) / theta_d;
      } else {
        converged = true;
      }
      const bool flipped = (theta_d < 0 && theta > 0) || (theta_d > 0 && theta < 0);
      ok = converged && !flipped;
      x = pwx * scale;
      y = pwy * scale;
    } else {
      const double ifx = 1.0 / job.fx, ify = 1.0 / job.fy;
      const double x0 = (p.x - job.cx) * ifx, y0 = (p.y - job.cy) * ify;
      x = x0;
      y = y0;
      for (int j = 0; j < 5; ++j) {
        const double r2 = x * x + y * y;
        const double icdist = 1.0 / (1 + (job.k[1] * r2 + job.k[0]) * r2);
        if (icdist < 0) { x = x0; y = y0; break; }
        const double dx = 2 * job.k[2] * x * y + job.k[3] * (r2 + 2 * x * x);
        const double dy = job.k[2] * (r2 + 2 * y * y) + 2 * job.k[3] * x * y;
        x = (x0 - dx) * icdist;
        y = (y0 - dy) * icdist;
      }
    }
    double ox = -1000000.0, oy = -1000000.0;                 // OpenCV's marker for a fisheye point that did not converge
    if (ok) {
      const double xx = job.RR[0] * x + job.RR[1] * y + job.RR[2];
      const double yy = job.RR[3] * x + job.RR[4] * y + job.RR[5];
      const double ww = 1.0 / (job.RR[6] * x + job.RR[7] * y + job.RR[8]);
      ox = xx * ww;
      oy = yy * ww;
    }
    out[(size_t)i * out_stride + out_off] = ox;
    out[(size_t)i * out_stride + out_off + 1] = oy;
  }
}

UndistortJob make_job(const double* intr, int model, const double* dist, const double* R, const double* newK) {
  UndistortJob j;
  j.fx = intr[0]; j.fy = intr[1]; j.cx = intr[2]; j.cy = intr[3];
  for (int i = 0; i < 4; ++i) j.k[i] = dist[i];
  j.model = model == 1 ? 1 : 0;
  const double nk[4] = {newK ? newK[0] : 1.0, newK ? newK[1] : 1.0, newK ? newK[2] : 0.0, newK ? newK[3] : 0.0};
  const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  // reference quirk (image.py:668-670): the positional call of cv2.fisheye.undistortPoints hands the rectification
  // matrix to the output slot and K_new to R, so the equidistant model ignores the rectification
  const double* Rm = (R && j.model == 0) ? R : I;
  const double Kn[9] = {nk[0], 0, nk[2], 0, nk[1], nk[3], 0, 0, 1};
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) j.RR[3 * r + c] = Kn[3 * r] * Rm[c] + Kn[3 * r + 1] * Rm[3 + c] + Kn[3 * r + 2] * Rm[6 + c];
  return j;
}

int run(int device, int n, const double* in0, const UndistortJob& j0, const double* in1, const UndistortJob* j1, double* out) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return MSCKF_ERR_CUDA;
  if (cudaSetDevice(device) != cudaSuccess) return MSCKF_ERR_CUDA;
  if (n == 0) return MSCKF_OK;
  const int ncam = j1 ? 2 : 1;
  const size_t in_bytes = (size_t)n * 2 * sizeof(double), out_bytes = (size_t)n * 2 * ncam * sizeof(double);
  double *d_in = nullptr, *d_out = nullptr;
  int rc = MSCKF_ERR_CUDA;
  if (cudaMalloc((void**)&d_in, in_bytes * ncam) == cudaSuccess && cudaMalloc((void**)&d_out, out_bytes) == cudaSuccess &&
      cudaMemcpy(d_in, in0, in_bytes, cudaMemcpyHostToDevice) == cudaSuccess &&
      (!j1 || cudaMemcpy(d_in + (size_t)n * 2, in1, in_bytes, cudaMemcpyHostToDevice) == cudaSuccess)) {
    const int grid = (n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8;
    k_undistort<<<grid, 256>>>(j0, n, (const double2*)d_in, d_out, 2 * ncam, 0);
    if (j1) k_undistort<<<grid, 256>>>(*j1, n, (const double2*)(d_in + (size_t)n * 2), d_out, 4, 2);
    if (cudaGetLastError() == cudaSuccess && cudaMemcpy(out, d_out, out_bytes, cudaMemcpyDeviceToHost) == cudaSuccess) rc = MSCKF_OK;
  }
  cudaFree(d_in);
  cudaFree(d_out);
  return rc;
}

}  // namespace

extern "C" int msckf_undistort_points(int32_t device, int32_t n, const double* pts_in, const double* intrinsics, int32_t model,
                                      const double* dist, const double* rectification, const double* new_intrinsics, double* pts_out) {
  if (n < 0 || (n > 0 && (!pts_in || !pts_out)) || !intrinsics || !dist) return MSCKF_ERR_ARG;
  return run(device, n, pts_in, make_job(intrinsics, model, dist, rectification, new_intrinsics), nullptr, nullptr, pts_out);
}

extern "C" int msckf_undistort_stereo(int32_t device, int32_t n, const double* cam0_pts, const double* cam1_pts, const double* intr0,
                                      int32_t model0, const double* dist0, const double* intr1, int32_t model1, const double* dist1,
                                      double* uv_out) {
  if (n < 0 || (n > 0 && (!cam0_pts || !cam1_pts || !uv_out)) || !intr0 || !dist0 || !intr1 || !dist1) return MSCKF_ERR_ARG;
  const UndistortJob j1 = make_job(intr1, model1, dist1, nullptr, nullptr);
  return run(device, n, cam0_pts, make_job(intr0, model0, dist0, nullptr, nullptr), cam1_pts, &j1, uv_out);
}
