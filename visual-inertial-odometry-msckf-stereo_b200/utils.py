"""Host-side geometry helpers with the reference's names (utils.py:4-152).

Only message conversion and ``publish`` use these (a handful of 3x3 products per frame on
values the GPU already produced); the filter arithmetic itself -- including the device
versions of every function here -- lives in ``csrc/msckf_dev.cuh``.
"""
import numpy as np


def skew(vec):
    """utils.py:4-12"""
    x, y, z = vec
    return np.array([[0., -z, y], [z, 0., -x], [-y, x, 0.]])


def to_rotation(q):
    """utils.py:14-27  JPL [x y z w] -> rotation matrix (world -> body)."""
    q = np.asarray(q, dtype=np.float64)
    q = q / np.linalg.norm(q)
    vec, w = q[:3], q[3]
    return (2 * w * w - 1) * np.identity(3) - 2 * w * skew(vec) + 2 * vec[:, None] * vec


def to_quaternion(R):
    """utils.py:29-53"""
    if R[2, 2] < 0:
        if R[0, 0] > R[1, 1]:
            t = 1 + R[0, 0] - R[1, 1] - R[2, 2]
            q = [t, R[0, 1] + R[1, 0], R[2, 0] + R[0, 2], R[1, 2] - R[2, 1]]
        else:
            t = 1 - R[0, 0] + R[1, 1] - R[2, 2]
            q = [R[0, 1] + R[1, 0], t, R[2, 1] + R[1, 2], R[2, 0] - R[0, 2]]
    else:
        if R[0, 0] < -R[1, 1]:
            t = 1 - R[0, 0] - R[1, 1] + R[2, 2]
            q = [R[0, 2] + R[2, 0], R[2, 1] + R[1, 2], t, R[0, 1] - R[1, 0]]
        else:
            t = 1 + R[0, 0] + R[1, 1] + R[2, 2]
            q = [R[1, 2] - R[2, 1], R[2, 0] - R[0, 2], R[0, 1] - R[1, 0], t]
    q = np.array(q, dtype=np.float64)
    return q / np.linalg.norm(q)


def quaternion_normalize(q):
    """utils.py:55-59"""
    return q / np.linalg.norm(q)


def quaternion_conjugate(q):
    """utils.py:61-65"""
    return np.array([-q[0], -q[1], -q[2], q[3]])


def from_two_vectors(v0, v1):
    """utils.py:104-128  used once, by initialize_gravity_and_bias (msckf.py:257)."""
    v0 = v0 / np.linalg.norm(v0)
    v1 = v1 / np.linalg.norm(v1)
    d = v0 @ v1
    if d < -0.999999:
        axis = np.cross([1, 0, 0], v0)
        if np.linalg.norm(axis) < 0.000001:
            axis = np.cross([0, 1, 0], v0)
        q = np.array([axis[0], axis[1], axis[2], 0.])
    elif d > 0.999999:
        q = np.array([0., 0., 0., 1.])
    else:
        s = np.sqrt((1 + d) * 2)
        axis = np.cross(v0, v1)
        q = np.array([axis[0] / s, axis[1] / s, axis[2] / s, 0.5 * s])
    q = q / np.linalg.norm(q)
    return quaternion_conjugate(q)


class Isometry3d(object):
    """utils.py:132-152  (attribute names as in the reference)."""

    def __init__(self, R, t):
        self._vio_R__ = R
        self._vio_t__ = t

    def matrix(self):
        m = np.identity(4)
        m[:3, :3] = self._vio_R__
        m[:3, 3] = self._vio_t__
        return m

    def inverse(self):
        return Isometry3d(self._vio_R__.T, -self._vio_R__.T @ self._vio_t__)

    def __mul__(self, T1):
        return Isometry3d(self._vio_R__ @ T1._vio_R__, self._vio_R__ @ T1._vio_t__ + self._vio_t__)
