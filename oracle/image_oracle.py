"""CPU oracle (TEST INFRASTRUCTURE, not product code) for the step right in front of the MSCKF back end:
point undistortion + normalisation of the stereo feature message (reference MSCKF/image.py:406-438 `publish`,
:640-674 `undistort_points`).

The reference delegates the arithmetic to OpenCV (`cv2.undistortPoints`, `cv2.fisheye.undistortPoints`), a third
party dependency it does not pin (README.md:5-11 lists bare package names); the build container has opencv 4.13.0.
This file restates OpenCV's published algorithms:
  * radtan  (calib3d undistortPoints, default TermCriteria(COUNT, 5, 0.01)): five fixed-point iterations of
      x <- (x0 - deltaX(x, y)) * icdist(r2),  icdist = 1 / (1 + k1 r2 + k2 r2^2)   (4-coefficient model k1 k2 p1 p2)
  * equidistant (fisheye::undistortPoints, default TermCriteria(COUNT + EPS, 10, 1e-8)): Newton on
      theta (1 + k1 theta^2 + k2 theta^4 + k3 theta^6 + k4 theta^8) = theta_d,  scale = tan(theta) / theta_d
followed by the rectification rotation and the new camera matrix.  Pinned against the reference's own
`ImageProcessor.undistort_points` run here: tests/golden/make_golden_image.py -> tests/golden/undistort.npz."""
import numpy as np

RADTAN, EQUIDISTANT = 0, 1


def model_code(name):
    return EQUIDISTANT if name == 'equidistant' else RADTAN      # image.py:668-673: anything else is radtan


def undistort_points(pts_in, intrinsics, distortion_model, distortion_coeffs, rectification_matrix=np.identity(3),
                     new_intrinsics=np.array([1, 1, 0, 0])):
    """image.py:640-674.  pts_in [n][2] pixel coordinates -> [n][2]."""
    if len(pts_in) == 0:
        return []
    pts = np.asarray(pts_in, dtype=np.float64).reshape(-1, 2)
    fx, fy, cx, cy = [float(v) for v in intrinsics]
    k = [float(v) for v in distortion_coeffs]
    Kn = np.array([[new_intrinsics[0], 0.0, new_intrinsics[2]], [0.0, new_intrinsics[1], new_intrinsics[3]], [0.0, 0.0, 1.0]], dtype=np.float64)
    RR = Kn @ np.asarray(rectification_matrix, dtype=np.float64)
    if model_code(distortion_model) == EQUIDISTANT:
        # reference quirk (image.py:668-670): cv2.fisheye.undistortPoints takes (distorted, K, D, undistorted, R, P), so
        # the positional call hands the rectification matrix to the OUTPUT slot and K_new to R: rectification is ignored
        RR = Kn
    out = np.zeros_like(pts)
    for i, (u, v) in enumerate(pts):
        if model_code(distortion_model) == EQUIDISTANT:
            pwx, pwy = (u - cx) / fx, (v - cy) / fy
            theta_d = np.sqrt(pwx * pwx + pwy * pwy)
            theta_d = min(max(-np.pi / 2., theta_d), np.pi / 2.)
            converged, theta, scale = False, theta_d, 0.0
            if abs(theta_d) > 1e-8:
                for _ in range(10):
                    t2 = theta * theta; t4 = t2 * t2; t6 = t4 * t2; t8 = t6 * t2
                    k0, k1, k2, k3 = k[0] * t2, k[1] * t4, k[2] * t6, k[3] * t8
                    fix = (theta * (1 + k0 + k1 + k2 + k3) - theta_d) / (1 + 3 * k0 + 5 * k1 + 7 * k2 + 9 * k3)
                    theta = theta - fix
                    if abs(fix) < 1e-8:
                        converged = True
                        break
                scale = np.tan(theta) / theta_d
            else:
                converged = True
            flipped = (theta_d < 0 and theta > 0) or (theta_d > 0 and theta < 0)
            if converged and not flipped:
                x, y = pwx * scale, pwy * scale
            else:
                out[i] = (-1000000.0, -1000000.0)
                continue
        else:
            ifx, ify = 1.0 / fx, 1.0 / fy
            x0 = x = (u - cx) * ifx
            y0 = y = (v - cy) * ify
            for _ in range(5):
                r2 = x * x + y * y
                icdist = 1.0 / (1 + (k[1] * r2 + k[0]) * r2)
                if icdist < 0:
                    x, y = x0, y0
                    break
                dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x)
                dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y
                x = (x0 - dx) * icdist
                y = (y0 - dy) * icdist
        xx = RR[0, 0] * x + RR[0, 1] * y + RR[0, 2]
        yy = RR[1, 0] * x + RR[1, 1] * y + RR[1, 2]
        ww = 1.0 / (RR[2, 0] * x + RR[2, 1] * y + RR[2, 2])
        out[i] = (xx * ww, yy * ww)
    return out


def publish_measurements(cam0_points, cam1_points, config):
    """image.py:406-438: the [u0 v0 u1 v1] rows of the feature message from raw pixel tracks of both cameras."""
    a = undistort_points(cam0_points, config._vio_cam0_intrinsics__, config._vio_cam0_distortion_model__, config._vio_cam0_distortion_coeffs__)
    b = undistort_points(cam1_points, config._vio_cam1_intrinsics__, config._vio_cam1_distortion_model__, config._vio_cam1_distortion_coeffs__)
    if len(a) == 0:
        return np.zeros((0, 4))
    return np.hstack([a, b])
