"""The arithmetic of the reference's ``ImageProcessor`` that sits right in front of the filter, on the GPU:
``undistort_points`` (image.py:640-674) and the feature-message assembly of ``publish`` (image.py:406-438).
The cv2 KLT / FAST front end itself (tracking, detection, RANSAC) stays out of scope; with these two calls a
caller can hand raw pixel tracks to the back end.  Same argument meaning as the reference methods."""
import ctypes as C
from collections import namedtuple

import numpy as np

from . import capi

RADTAN, EQUIDISTANT = 0, 1


def _model(name):
    return EQUIDISTANT if name == 'equidistant' else RADTAN       # image.py:668-673: anything else is radtan


def undistort_points(pts_in, intrinsics, distortion_model, distortion_coeffs, rectification_matrix=np.identity(3),
                     new_intrinsics=np.array([1, 1, 0, 0]), device=0):
    """image.py:640-674.  Returns [] for an empty input, else an (n, 2) float64 array."""
    if len(pts_in) == 0:
        return []
    lib = capi.load_library()
    pts = np.ascontiguousarray(np.reshape(np.asarray(pts_in, dtype=np.float64), (-1, 2)))
    intr = np.ascontiguousarray(intrinsics, dtype=np.float64)
    dist = np.ascontiguousarray(distortion_coeffs, dtype=np.float64)
    R = np.ascontiguousarray(rectification_matrix, dtype=np.float64)
    nk = np.ascontiguousarray(new_intrinsics, dtype=np.float64)
    out = np.zeros_like(pts)
    rc = lib.msckf_undistort_points(device, len(pts), pts.ctypes.data, intr.ctypes.data, _model(distortion_model), dist.ctypes.data,
                                    R.ctypes.data, nk.ctypes.data, out.ctypes.data)
    if rc != 0:
        raise capi.MsckfCudaError(rc, 'msckf_undistort_points failed (no usable sm_100 device?)')
    return out


def undistort_stereo(cam0_points, cam1_points, config, device=0):
    """The [u0 v0 u1 v1] rows of publish() (image.py:419-433) for n stereo tracks: (n, 4) float64."""
    p0 = np.ascontiguousarray(np.reshape(np.asarray(cam0_points, dtype=np.float64), (-1, 2)))
    p1 = np.ascontiguousarray(np.reshape(np.asarray(cam1_points, dtype=np.float64), (-1, 2)))
    assert p0.shape == p1.shape
    out = np.zeros((len(p0), 4))
    if len(p0) == 0:
        return out
    lib = capi.load_library()
    a = [np.ascontiguousarray(x, dtype=np.float64) for x in (config._vio_cam0_intrinsics__, config._vio_cam0_distortion_coeffs__,
                                                             config._vio_cam1_intrinsics__, config._vio_cam1_distortion_coeffs__)]
    rc = lib.msckf_undistort_stereo(device, len(p0), p0.ctypes.data, p1.ctypes.data, a[0].ctypes.data, _model(config._vio_cam0_distortion_model__),
                                    a[1].ctypes.data, a[2].ctypes.data, _model(config._vio_cam1_distortion_model__), a[3].ctypes.data,
                                    out.ctypes.data)
    if rc != 0:
        raise capi.MsckfCudaError(rc, 'msckf_undistort_stereo failed (no usable sm_100 device?)')
    return out


FeatureMeasurement = namedtuple('FeatureMeasurement', ['id', 'u0', 'v0', 'u1', 'v1'])


def publish(timestamp, ids, cam0_points, cam1_points, config, device=0):
    """image.py:406-438: the feature message (timestamp, vio_features) MSCKF.feature_callback consumes."""
    uv = undistort_stereo(cam0_points, cam1_points, config, device)
    feats = [FeatureMeasurement(int(i), *row) for i, row in zip(ids, uv)]
    return namedtuple('vio_feature_msg__', ['timestamp', 'vio_features'])(timestamp, feats)
