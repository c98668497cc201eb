"""CPU oracle for the MSCKF back-end hot path  --  TEST INFRASTRUCTURE ONLY.

This module is a plain-numpy restatement of the reference's algorithm
(``/root/reference/MSCKF/{msckf,feature,jit_utils,utils}.py``).  It exists so the
CUDA path can be checked; it is never on the product path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import it.

Parity pinning: the reference ships no tests or golden vectors, so this
oracle is pinned against the reference *itself*, run unmodified in the build
container by ``tests/golden/make_golden.py``; the resulting vectors are committed
under ``tests/golden/`` and ``tests/test_oracle_golden.py`` replays them.

Every function cites the reference lines it follows.  Attribute names that form
the reference's public surface (``state_server._vio_state_cov__`` ...) are kept so
a single harness can read any implementation; everything else is written in this
repo's own style.  numba is not used: the reference's jitted helpers lower to the
same LAPACK/BLAS calls numpy makes.

Quirks of the reference that are reproduced on purpose:
  * RK4 re-uses a stale skew matrix                       (jit_utils.py:80-104)
  * the LM loop accepts at most one step                  (feature.py:224-276)
  * chi-square gate uses the 5 % lower tail               (msckf.py:123)
  * covariance update is sym((I-KH)P), not Joseph         (msckf.py:598-604)
  * the "null" velocity/position alias the live arrays, so in-place "+=" in
    measurement_update also moves them                    (msckf.py:366-368,580-582,595)
"""
from collections import namedtuple

import numpy as np
from scipy.stats import chi2

# --------------------------------------------------------------------------- messages
ImuMsg = namedtuple('imu_msg', ['vio_timestamp__', 'angular_velocity', 'linear_acceleration'])
FeatureMsg = namedtuple('vio_feature_msg__', ['timestamp', 'vio_features'])
VioResult = namedtuple('vio_result', ['timestamp', 'pose', 'velocity', 'cam0_pose'])


class FeatureMeasurement(object):
    """image.py:22-33 -- one stereo measurement in normalised coordinates."""
    __slots__ = ('id', 'u0', 'v0', 'u1', 'v1')

    def __init__(self, fid, u0, v0, u1, v1):
        self.id, self.u0, self.v0, self.u1, self.v1 = fid, u0, v0, u1, v1


# --------------------------------------------------------------------------- geometry (utils.py:4-152)
def skew(v):
    """utils.py:4-12"""
    return np.array([[0., -v[2], v[1]], [v[2], 0., -v[0]], [-v[1], v[0], 0.]])


def to_rotation(q):
    """utils.py:14-27  JPL quaternion [x y z w] -> R (world->body); normalises q first."""
    q = q / np.linalg.norm(q)
    v, w = q[:3], q[3]
    return (2 * w * w - 1) * np.identity(3) - 2 * w * skew(v) + 2 * v[:, None] * v


def to_quaternion(R):
    """utils.py:29-53  four-branch extraction, normalised by its norm."""
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
    q = np.array(q)
    return q / np.linalg.norm(q)


def quaternion_conjugate(q):
    """utils.py:61-65"""
    return np.array([-q[0], -q[1], -q[2], q[3]])


def quaternion_multiplication(q1, q2):
    """utils.py:67-82  q1 (x) q2, inputs and output normalised."""
    a = q1 / np.linalg.norm(q1)
    b = q2 / np.linalg.norm(q2)
    L = np.array([[a[3], a[2], -a[1], a[0]],
                  [-a[2], a[3], a[0], a[1]],
                  [a[1], -a[0], a[3], a[2]],
                  [-a[0], -a[1], -a[2], a[3]]])
    q = L @ b
    return q / np.linalg.norm(q)


def small_angle_quaternion(dtheta):
    """utils.py:85-101"""
    dq = dtheta / 2.
    s = dq @ dq
    if s <= 1:
        return np.array([dq[0], dq[1], dq[2], np.sqrt(1 - s)])
    q = np.array([dq[0], dq[1], dq[2], 1.])
    return q / np.sqrt(1 + s)


def from_two_vectors(v0, v1):
    """utils.py:104-128  rotation v0->v1, returned as the JPL conjugate."""
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
    """utils.py:132-152  rigid transform; attribute names are the reference's."""

    def __init__(self, R, t):
        self._vio_R__ = R
        self._vio_t__ = t

    # readable aliases used inside this repo
    @property
    def R(self):
        return self._vio_R__

    @property
    def t(self):
        return self._vio_t__

    def matrix(self):
        m = np.identity(4)
        m[:3, :3] = self._vio_R__
        m[:3, 3] = self._vio_t__
        return m

    def inverse(self):
        return Isometry3d(self._vio_R__.T, -self._vio_R__.T @ self._vio_t__)

    def __mul__(self, other):
        return Isometry3d(self._vio_R__ @ other._vio_R__,
                          self._vio_R__ @ other._vio_t__ + self._vio_t__)


# --------------------------------------------------------------------------- jit_utils.py restated
def process_model_matrices(gyro, R_w_i, acc, dt):
    """jit_utils.py:6-43  -> (G 21x12, Phi 21x21) ; Phi is the 3rd-order series of exp(F dt)."""
    F = np.zeros((21, 21))
    G = np.zeros((21, 12))
    F[0:3, 0:3] = -skew(gyro)
    F[0:3, 3:6] = -np.identity(3)
    F[6:9, 0:3] = -R_w_i.T @ skew(acc)
    F[6:9, 9:12] = -R_w_i.T
    F[12:15, 6:9] = np.identity(3)
    G[0:3, 0:3] = -np.identity(3)
    G[3:6, 3:6] = np.identity(3)
    G[6:9, 6:9] = -R_w_i.T
    G[9:12, 9:12] = np.identity(3)
    Fdt = F * dt
    Fdt2 = Fdt @ Fdt
    Fdt3 = Fdt2 @ Fdt
    Phi = np.identity(21) + Fdt + Fdt2 / 2. + Fdt3 / 6.
    return G, Phi


def predict_new_state(dt, gyro, acc, q, v, p, gravity):
    """jit_utils.py:46-128  closed-form dq + RK4 on v, p, *including the stale-skew quirk*:
    the skew matrix is built once from the normalised dq(dt) (lines 80-84) and re-used
    for R(dq(dt/2)) (line 94) and R(q) (line 104)."""
    gnorm = np.linalg.norm(gyro)
    Omega = np.zeros((4, 4))
    Omega[:3, :3] = -skew(gyro)
    Omega[:3, 3] = gyro
    Omega[3, :3] = -gyro
    if gnorm > 1e-5:
        dq_dt = (np.cos(gnorm * dt * 0.5) * np.identity(4) + np.sin(gnorm * dt * 0.5) / gnorm * Omega) @ q
        dq_dt2 = (np.cos(gnorm * dt * 0.25) * np.identity(4) + np.sin(gnorm * dt * 0.25) / gnorm * Omega) @ q
    else:
        dq_dt = np.cos(gnorm * dt * 0.5) * (np.identity(4) + Omega * dt * 0.5) @ q
        dq_dt2 = np.cos(gnorm * dt * 0.25) * (np.identity(4) + Omega * dt * 0.25) @ q

    def rot_stale(quat, stale_skew):
        vec, w = quat[:3], quat[3]
        return (2 * w * w - 1) * np.identity(3) - 2 * w * stale_skew + 2 * vec[:, None] * vec

    dq_dt = dq_dt / np.linalg.norm(dq_dt)
    stale = skew(dq_dt[:3])
    dR_dt_T = rot_stale(dq_dt, stale).T
    dq_dt2 = dq_dt2 / np.linalg.norm(dq_dt2)
    dR_dt2_T = rot_stale(dq_dt2, stale).T
    qn = q / np.linalg.norm(q)
    R = rot_stale(qn, stale)

    k1_p, k1_v = v, R.T @ acc + gravity
    k2_p = v + k1_v * dt / 2.
    k2_v = dR_dt2_T @ acc + gravity
    k3_p = v + k2_v * dt / 2
    k3_v = dR_dt2_T @ acc + gravity
    k4_p = v + k3_v * dt
    k4_v = dR_dt_T @ acc + gravity
    q_new = dq_dt / np.linalg.norm(dq_dt)
    v_new = v + (k1_v + 2 * k2_v + 2 * k3_v + k4_v) * dt / 6.
    p_new = p + (k1_p + 2 * k2_p + 2 * k3_p + k4_p) * dt / 6.
    return q_new, v_new, p_new


def propagate_imu_covariance(Qc, P, G, Phi, dt):
    """jit_utils.py:130-135  mutates P[:21,:21] in place and returns P."""
    Q = Phi @ G @ Qc @ G.T @ Phi.T * dt
    P[:21, :21] = Phi @ P[:21, :21] @ Phi.T + Q
    return P


def augment_covariance(R_i_c, R_w_i, t_c_i, P):
    """jit_utils.py:137-167  grows P by the 6 rows/cols of the new camera state."""
    n = P.shape[0]
    J = np.zeros((6, 21))
    J[0:3, 0:3] = R_i_c
    J[0:3, 15:18] = np.identity(3)
    J[3:6, 0:3] = skew(R_w_i.T @ t_c_i)
    J[3:6, 12:15] = np.identity(3)
    J[3:6, 18:21] = np.identity(3)
    Pn = np.zeros((n + 6, n + 6))
    Pn[:n, :n] = P
    Pn[n:, :n] = J @ Pn[:21, :n]
    Pn[:n, n:] = Pn[n:, :n].T
    Pn[n:, n:] = J @ Pn[:21, :21] @ J.T
    return Pn


# --------------------------------------------------------------------------- config (config.py:5-124, hot-path fields only)
class OptimizationConfigEuRoC(object):
    """config.py:5-15"""

    def __init__(self):
        self._vio_translation_threshold__ = -1.0
        self._vio_huber_epsilon__ = 0.01
        self._vio_estimation_precision__ = 5e-7
        self._vio_initial_damping__ = 1e-3
        self._vio_outer_loop_max_iteration__ = 5
        self._vio_inner_loop_max_iteration__ = 5


class ConfigEuRoC(object):
    """config.py:17-124 without the cv2-only front-end fields."""

    def __init__(self):
        self._vio_optimization_config__ = OptimizationConfigEuRoC()
        self._vio_gravity_acc__ = 9.81
        self._vio_gravity__ = np.array([0.0, 0.0, -self._vio_gravity_acc__])
        self._vio_frame_rate__ = 20
        self._vio_max_cam_state_size__ = 20
        self._vio_position_std_threshold__ = 8.0
        self._vio_rotation_threshold__ = 0.2618
        self._vio_translation_threshold__ = 0.4
        self._vio_tracking_rate_threshold__ = 0.5
        self._vio_gyro_noise__ = 0.005 ** 2
        self._vio_acc_noise__ = 0.05 ** 2
        self._vio_gyro_bias_noise__ = 0.001 ** 2
        self._vio_acc_bias_noise__ = 0.01 ** 2
        self._vio_observation_noise__ = 0.035 ** 2
        self._vio_velocity__ = np.zeros(3)
        self._vio_velocity_cov__ = 0.25
        self._vio_gyro_bias_cov__ = 0.01
        self._vio_acc_bias_cov__ = 0.01
        self._vio_extrinsic_rotation_cov__ = 3.0462e-4
        self._vio_extrinsic_translation_cov__ = 2.5e-5
        self._vio_T_imu_cam0__ = np.array([
            [0.014865542981794, 0.999557249008346, -0.025774436697440, 0.065222909535531],
            [-0.999880929698575, 0.014967213324719, 0.003756188357967, -0.020706385492719],
            [0.004140296794224, 0.025715529947966, 0.999660727177902, -0.008054602460030],
            [0, 0, 0, 1.000000000000000]])
        self._vio_T_cn_cnm1__ = np.array([
            [0.999997256477881, 0.002312067192424, 0.000376008102415, -0.110073808127187],
            [-0.002317135723281, 0.999898048506644, 0.014089835846648, 0.000399121547014],
            [-0.000343393120525, -0.014090668452714, 0.999900662637729, -0.000853702503357],
            [0, 0, 0, 1.000000000000000]])
        self._vio_T_imu_body__ = np.eye(4)


# --------------------------------------------------------------------------- state types (msckf.py:16-100)
class IMUState(object):
    """msckf.py:16-59.  next-id / gravity are per filter here (the reference keeps them as
    class globals; that is identical for filters run one after another)."""

    def __init__(self, new_id=None):
        self.id = new_id
        self.timestamp = None
        self.orientation = np.array([0., 0., 0., 1.])
        self._vio_position__ = np.zeros(3)
        self.velocity = np.zeros(3)
        self._vio_gyro_bias__ = np.zeros(3)
        self.acc_bias = np.zeros(3)
        self.orientation_null = np.array([0., 0., 0., 1.])
        self.position_null = np.zeros(3)
        self.velocity_null = np.zeros(3)
        self.R_imu_cam0 = np.identity(3)
        self.t_cam0_imu = np.zeros(3)


class CAMState(object):
    """msckf.py:62-86"""

    def __init__(self, new_id=None):
        self.id = new_id
        self.timestamp = None
        self.orientation = np.array([0., 0., 0., 1.])
        self._vio_position__ = np.zeros(3)
        self.orientation_null = np.array([0., 0., 0., 1.])
        self.position_null = np.zeros(3)


class StateServer(object):
    """msckf.py:89-100"""

    def __init__(self):
        self._vio_imu_state__ = IMUState()
        self._vio_cam_states__ = dict()
        self._vio_state_cov__ = np.zeros((21, 21))
        self._vio_continuous_noise_cov__ = np.zeros((12, 12))


# --------------------------------------------------------------------------- landmark (feature.py:7-295)
class Feature(object):
    def __init__(self, new_id=0, optimization_config=None, R_cam0_cam1=None, t_cam0_cam1=None):
        self.id = new_id
        self.observations = dict()
        self._vio_position__ = np.zeros(3)
        self.is_initialized = False
        self.optimization_config = optimization_config
        self.R_cam0_cam1 = R_cam0_cam1
        self.t_cam0_cam1 = t_cam0_cam1
        self.lm_solves = 0   # instrumentation only

    @staticmethod
    def _h(T, x):
        """feature.py:46-48 / 72-75"""
        return T._vio_R__ @ np.array([x[0], x[1], 1.0]) + x[2] * T._vio_t__

    def cost(self, T_c0_ci, x, z):
        """feature.py:33-55"""
        h = self._h(T_c0_ci, x)
        return ((h[:2] / h[2] - z) ** 2).sum()

    def jacobian(self, T_c0_ci, x, z):
        """feature.py:57-97 -> (J 2x3, r 2, huber weight)"""
        h1, h2, h3 = self._h(T_c0_ci, x)
        W = np.zeros((3, 3))
        W[:, :2] = T_c0_ci._vio_R__[:, :2]
        W[:, 2] = T_c0_ci._vio_t__
        J = np.zeros((2, 3))
        J[0] = W[0] / h3 - W[2] * h1 / (h3 * h3)
        J[1] = W[1] / h3 - W[2] * h2 / (h3 * h3)
        r = np.array([h1 / h3, h2 / h3]) - z
        e = np.linalg.norm(r)
        eps = self.optimization_config._vio_huber_epsilon__
        w = 1.0 if e <= eps else eps / (2 * e)
        return J, r, w

    def generate_initial_guess(self, T_c1_c2, z1, z2):
        """feature.py:99-122  two-view depth by least squares."""
        m = T_c1_c2._vio_R__ @ np.array([z1[0], z1[1], 1.0])
        a = m[:2] - z2 * m[2]
        b = z2 * T_c1_c2._vio_t__[2] - T_c1_c2._vio_t__[:2]
        depth = a @ b / (a @ a)
        return np.array([z1[0], z1[1], 1.0]) * depth

    def check_motion(self, cam_states):
        """feature.py:124-165"""
        thr = self.optimization_config._vio_translation_threshold__
        if thr < 0:
            return True
        ids = list(self.observations.keys())
        first, last = cam_states[ids[0]], cam_states[ids[-1]]
        R_first = to_rotation(first.orientation).T
        d = np.array([self.observations[ids[0]][0], self.observations[ids[0]][1], 1.0])
        d = R_first @ (d / np.linalg.norm(d))
        tr = last._vio_position__ - first._vio_position__
        ortho = tr - (tr @ d) * d
        return np.linalg.norm(ortho) > thr

    def initialize_position(self, cam_states):
        """feature.py:167-295  LM triangulation in inverse-depth coordinates of the first cam0."""
        T_c1_c0 = Isometry3d(self.R_cam0_cam1, self.t_cam0_cam1).inverse()
        poses, zs = [], []
        for cam_id, m in self.observations.items():
            cs = cam_states.get(cam_id)
            if cs is None:                                    # feature.py:189-192
                continue
            zs += [m[:2], m[2:]]
            c0 = Isometry3d(to_rotation(cs.orientation).T, cs._vio_position__)
            poses += [c0, c0 * T_c1_c0]
        T_c0_w = poses[0]
        poses = [T.inverse() * T_c0_w for T in poses]         # feature.py:209-213

        guess = self.generate_initial_guess(poses[-2], zs[0], zs[-2])
        x = np.array([guess[0], guess[1], 1.0]) / guess[2]

        cfg = self.optimization_config
        lam = cfg._vio_initial_damping__
        inner = outer = 0
        reduced = False                                       # never reset to False after a success
        delta_norm = float('inf')
        total = 0.0
        for T, z in zip(poses, zs):
            total += self.cost(T, x, z)

        while outer < cfg._vio_outer_loop_max_iteration__ and delta_norm > cfg._vio_estimation_precision__:
            A = np.zeros((3, 3))
            b = np.zeros(3)
            for T, z in zip(poses, zs):
                J, r, w = self.jacobian(T, x, z)
                if w == 1.0:
                    A += J.T @ J
                    b += J.T @ r
                else:
                    A += w * w * J.T @ J
                    b += w * w * J.T @ r
            while inner < cfg._vio_inner_loop_max_iteration__ and not reduced:
                delta = np.linalg.solve(A + lam * np.identity(3), b)
                self.lm_solves += 1
                x_new = x - delta
                delta_norm = np.linalg.norm(delta)
                new_cost = 0.0
                for T, z in zip(poses, zs):
                    new_cost += self.cost(T, x_new, z)
                if new_cost < total:
                    reduced, x, total = True, x_new, new_cost
                    lam = max(lam / 10., 1e-10)
                else:
                    reduced = False
                    lam = min(lam * 10., 1e12)
                inner += 1
            inner = 0
            outer += 1

        final = np.array([x[0], x[1], 1.0]) / x[2]
        valid = True
        for T in poses:
            if (T._vio_R__ @ final + T._vio_t__)[2] <= 0:
                valid = False
                break
        self._vio_position__ = T_c0_w._vio_R__ @ final + T_c0_w._vio_t__
        self.is_initialized = valid
        return valid


# --------------------------------------------------------------------------- the filter (msckf.py:104-908)
class MSCKF(object):
    ROW_CAP = 1500            # msckf.py:678 (literal in the reference)

    def __init__(self, config):
        self.config = config
        self.optimization_config = config._vio_optimization_config__
        self.imu_msg_buffer = []
        self.state_server = StateServer()
        self.map_server = dict()
        self.chi_squared_test_table = {i: chi2.ppf(0.05, i) for i in range(1, 100)}   # msckf.py:121-123
        self.state_server._vio_imu_state__.velocity = config._vio_velocity__
        self.reset_state_cov()
        Qc = np.identity(12)
        Qc[0:3, 0:3] *= config._vio_gyro_noise__
        Qc[3:6, 3:6] *= config._vio_gyro_bias_noise__
        Qc[6:9, 6:9] *= config._vio_acc_noise__
        Qc[9:, 9:] *= config._vio_acc_bias_noise__
        self.state_server._vio_continuous_noise_cov__ = Qc
        self.gravity = config._vio_gravity__                      # msckf.py:140 (class global there)
        T_cam0_imu = np.linalg.inv(config._vio_T_imu_cam0__)      # msckf.py:143-145
        self.state_server._vio_imu_state__.R_imu_cam0 = T_cam0_imu[:3, :3].T
        self.state_server._vio_imu_state__.t_cam0_imu = T_cam0_imu[:3, 3]
        self.R_cam0_cam1 = config._vio_T_cn_cnm1__[:3, :3]         # msckf.py:148-152
        self.t_cam0_cam1 = config._vio_T_cn_cnm1__[:3, 3]
        self.T_imu_body = Isometry3d(config._vio_T_imu_body__[:3, :3], config._vio_T_imu_body__[:3, 3])
        self.next_imu_id = 0                                       # msckf.py:18
        self.tracking_rate = None
        self.is_gravity_set = False
        self.is_first_img = True
        # instrumentation read by the parity harness; no effect on the algorithm
        self.log = None

    # ---- logging helper
    def _log(self, key, value):
        if self.log is not None:
            self.log.setdefault(key, []).append(value)

    # ---- msckf.py:166-178
    def imu_callback(self, imu_msg):
        self.imu_msg_buffer.append(imu_msg)
        if not self.is_gravity_set and len(self.imu_msg_buffer) >= 200:
            self.initialize_gravity_and_bias()
            self.is_gravity_set = True

    # ---- msckf.py:180-233
    def feature_callback(self, feature_msg):
        if not self.is_gravity_set:
            return None
        if self.is_first_img:
            self.is_first_img = False
            self.state_server._vio_imu_state__.timestamp = feature_msg.timestamp
        self.batch_imu_processing(feature_msg.timestamp)
        self.state_augmentation(feature_msg.timestamp)
        self.add_feature_observations(feature_msg)
        self.remove_lost_features()
        self.prune_cam_state_buffer()
        try:
            return self.publish(feature_msg.timestamp)
        finally:
            self.online_reset()

    # ---- msckf.py:235-258
    def initialize_gravity_and_bias(self):
        sum_w = np.zeros(3)
        sum_a = np.zeros(3)
        for m in self.imu_msg_buffer:
            sum_w += m.angular_velocity
            sum_a += m.linear_acceleration
        imu = self.state_server._vio_imu_state__
        imu._vio_gyro_bias__ = sum_w / len(self.imu_msg_buffer)
        g_imu = sum_a / len(self.imu_msg_buffer)
        self.gravity = np.array([0., 0., -np.linalg.norm(g_imu)])
        imu.orientation = from_two_vectors(-self.gravity, g_imu)

    # ---- msckf.py:262-287
    def batch_imu_processing(self, time_bound):
        imu = self.state_server._vio_imu_state__
        used = 0
        for m in self.imu_msg_buffer:
            t = m.vio_timestamp__
            if t < imu.timestamp:
                used += 1
                continue
            if t > time_bound:
                break
            self.process_model(t, m.angular_velocity, m.linear_acceleration)
            used += 1
            imu.timestamp = t
        imu.id = self.next_imu_id
        self.next_imu_id += 1
        self.imu_msg_buffer = self.imu_msg_buffer[used:]

    # ---- msckf.py:291-368
    def process_model(self, time, m_gyro, m_acc):
        ss = self.state_server
        imu = ss._vio_imu_state__
        dt = time - imu.timestamp
        gyro = m_gyro - imu._vio_gyro_bias__
        acc = m_acc - imu.acc_bias
        G, Phi = process_model_matrices(gyro, to_rotation(imu.orientation), acc, dt)
        imu.orientation, imu.velocity, imu._vio_position__ = predict_new_state(
            dt, gyro, acc, imu.orientation, imu.velocity, imu._vio_position__, self.gravity)

        R_kk_1 = to_rotation(imu.orientation_null)                         # msckf.py:329-344
        Phi[0:3, 0:3] = to_rotation(imu.orientation) @ R_kk_1.T
        u = R_kk_1 @ self.gravity
        s = u / (u @ u)
        A1 = Phi[6:9, 0:3]
        w1 = skew(imu.velocity_null - imu.velocity) @ self.gravity
        Phi[6:9, 0:3] = A1 - (A1 @ u - w1)[:, None] * s
        A2 = Phi[12:15, 0:3]
        w2 = skew(dt * imu.velocity_null + imu.position_null - imu._vio_position__) @ self.gravity
        Phi[12:15, 0:3] = A2 - (A2 @ u - w2)[:, None] * s

        ss._vio_state_cov__[:21, :21] = propagate_imu_covariance(
            ss._vio_continuous_noise_cov__, ss._vio_state_cov__, G, Phi, dt)[:21, :21]
        if len(ss._vio_cam_states__) > 0:
            ss._vio_state_cov__[:21, 21:] = Phi @ ss._vio_state_cov__[:21, 21:]
            ss._vio_state_cov__[21:, :21] = ss._vio_state_cov__[21:, :21] @ Phi.T
        ss._vio_state_cov__ = (ss._vio_state_cov__ + ss._vio_state_cov__.T) / 2.
        # msckf.py:366-368: plain assignment -> the null members ALIAS the live arrays
        imu.orientation_null = imu.orientation
        imu.position_null = imu._vio_position__
        imu.velocity_null = imu.velocity

    # ---- msckf.py:385-407
    def state_augmentation(self, time):
        ss = self.state_server
        imu = ss._vio_imu_state__
        R_i_c, t_c_i = imu.R_imu_cam0, imu.t_cam0_imu
        R_w_i = to_rotation(imu.orientation)
        cam = CAMState(imu.id)
        cam.timestamp = time
        cam.orientation = to_quaternion(R_i_c @ R_w_i)
        cam._vio_position__ = imu._vio_position__ + R_w_i.T @ t_c_i
        cam.orientation_null = cam.orientation          # aliases (msckf.py:399-400)
        cam.position_null = cam._vio_position__
        ss._vio_cam_states__[imu.id] = cam
        Pn = augment_covariance(R_i_c, R_w_i, t_c_i, ss._vio_state_cov__)
        ss._vio_state_cov__ = (Pn + Pn.T) / 2.

    # ---- msckf.py:409-427
    def add_feature_observations(self, feature_msg):
        sid = self.state_server._vio_imu_state__.id
        before = len(self.map_server)
        tracked = 0
        for f in feature_msg.vio_features:
            z = np.array([f.u0, f.v0, f.u1, f.v1])
            if f.id not in self.map_server:
                feat = Feature(f.id, self.optimization_config, self.R_cam0_cam1, self.t_cam0_cam1)
                feat.observations[sid] = z
                self.map_server[f.id] = feat
            else:
                self.map_server[f.id].observations[sid] = z
                tracked += 1
        self.tracking_rate = tracked / (before + 1e-5)

    # ---- msckf.py:429-498
    def measurement_jacobian(self, cam_state_id, feature_id):
        cam = self.state_server._vio_cam_states__[cam_state_id]
        feat = self.map_server[feature_id]
        R_w_c0 = to_rotation(cam.orientation)
        t_c0_w = cam._vio_position__
        R_w_c1 = self.R_cam0_cam1 @ R_w_c0
        t_c1_w = t_c0_w - R_w_c1.T @ self.t_cam0_cam1
        p_w = feat._vio_position__
        z = feat.observations[cam_state_id]
        p_c0 = R_w_c0 @ (p_w - t_c0_w)
        p_c1 = R_w_c1 @ (p_w - t_c1_w)

        dz_dpc0 = np.zeros((4, 3))
        dz_dpc0[0, 0] = 1 / p_c0[2]
        dz_dpc0[1, 1] = 1 / p_c0[2]
        dz_dpc0[0, 2] = -p_c0[0] / (p_c0[2] * p_c0[2])
        dz_dpc0[1, 2] = -p_c0[1] / (p_c0[2] * p_c0[2])
        dz_dpc1 = np.zeros((4, 3))
        dz_dpc1[2, 0] = 1 / p_c1[2]
        dz_dpc1[3, 1] = 1 / p_c1[2]
        dz_dpc1[2, 2] = -p_c1[0] / (p_c1[2] * p_c1[2])
        dz_dpc1[3, 2] = -p_c1[1] / (p_c1[2] * p_c1[2])
        dpc0_dxc = np.zeros((3, 6))
        dpc0_dxc[:, :3] = skew(p_c0)
        dpc0_dxc[:, 3:] = -R_w_c0
        dpc1_dxc = np.zeros((3, 6))
        dpc1_dxc[:, :3] = self.R_cam0_cam1 @ skew(p_c0)
        dpc1_dxc[:, 3:] = -R_w_c1
        H_x = dz_dpc0 @ dpc0_dxc + dz_dpc1 @ dpc1_dxc

        u = np.zeros(6)                                                    # msckf.py:483-490
        u[:3] = to_rotation(cam.orientation_null) @ self.gravity
        u[3:] = skew(p_w - cam.position_null) @ self.gravity
        H_x = H_x - (H_x @ u)[:, None] * u / (u @ u)
        H_f = -H_x[:4, 3:6]
        r = z - np.array([p_c0[0] / p_c0[2], p_c0[1] / p_c0[2], p_c1[0] / p_c1[2], p_c1[1] / p_c1[2]])
        return H_x, H_f, r

    # ---- msckf.py:500-541
    def feature_jacobian(self, feature_id, cam_state_ids):
        feat = self.map_server[feature_id]
        valid = [c for c in cam_state_ids if c in feat.observations]
        rows = 4 * len(valid)
        order = list(self.state_server._vio_cam_states__.keys())
        H_xj = np.zeros((rows, 21 + 6 * len(order)))
        H_fj = np.zeros((rows, 3))
        r_j = np.zeros(rows)
        for k, cam_id in enumerate(valid):
            H_xi, H_fi, r_i = self.measurement_jacobian(cam_id, feat.id)
            idx = order.index(cam_id)
            H_xj[4 * k:4 * k + 4, 21 + 6 * idx:27 + 6 * idx] = H_xi
            H_fj[4 * k:4 * k + 4, :3] = H_fi
            r_j[4 * k:4 * k + 4] = r_i
        U, _, _ = np.linalg.svd(H_fj)               # full U, LAPACK dgesdd (jit_utils.py:169-171)
        A = U[:, 3:]
        return A.T @ H_xj, A.T @ r_j

    # ---- msckf.py:543-604
    def measurement_update(self, H, r):
        if len(H) == 0 or len(r) == 0:
            return
        if H.shape[0] > H.shape[1]:
            Q, R = np.linalg.qr(H)
            H_thin, r_thin = R, Q.T @ r
        else:
            H_thin, r_thin = H, r
        ss = self.state_server
        P = ss._vio_state_cov__
        S = H_thin @ P @ H_thin.T + self.config._vio_observation_noise__ * np.identity(len(H_thin))
        K = np.linalg.solve(S, H_thin @ P).T
        dx = K @ r_thin
        self._log('delta_x', dx.copy())

        imu = ss._vio_imu_state__
        d = dx[:21]
        imu.orientation = quaternion_multiplication(small_angle_quaternion(d[0:3]), imu.orientation)
        imu._vio_gyro_bias__ += d[3:6]          # in place, as in the reference
        imu.velocity += d[6:9]
        imu.acc_bias += d[9:12]
        imu._vio_position__ += d[12:15]
        imu.R_imu_cam0 = to_rotation(small_angle_quaternion(d[15:18])) @ imu.R_imu_cam0
        imu.t_cam0_imu += d[18:21]
        for i, cam in enumerate(ss._vio_cam_states__.values()):
            dc = dx[21 + 6 * i:27 + 6 * i]
            cam.orientation = quaternion_multiplication(small_angle_quaternion(dc[:3]), cam.orientation)
            cam._vio_position__ += dc[3:]

        I_KH = np.identity(len(K)) - K @ H_thin
        Pn = I_KH @ ss._vio_state_cov__
        ss._vio_state_cov__ = (Pn + Pn.T) / 2.

    # ---- msckf.py:606-614
    def gating_test(self, H, r, dof):
        P1 = H @ self.state_server._vio_state_cov__ @ H.T
        P2 = self.config._vio_observation_noise__ * np.identity(len(H))
        gamma = r @ np.linalg.solve(P1 + P2, r)
        ok = bool(gamma < self.chi_squared_test_table[dof])
        self._log('gate', (gamma, self.chi_squared_test_table[dof], ok))
        return ok

    # ---- msckf.py:616-689
    def remove_lost_features(self):
        ss = self.state_server
        rows = 0
        invalid, processed = [], []
        for feat in self.map_server.values():
            if ss._vio_imu_state__.id in feat.observations:
                continue
            if len(feat.observations) < 3:
                invalid.append(feat.id)
                continue
            if not feat.is_initialized:
                if not feat.check_motion(ss._vio_cam_states__):
                    invalid.append(feat.id)
                    continue
                if feat.initialize_position(ss._vio_cam_states__) is False:
                    invalid.append(feat.id)
                    continue
            rows += 4 * len(feat.observations) - 3
            processed.append(feat.id)
        self._log('lost_invalid', list(invalid))
        self._log('lost_processed', list(processed))
        for fid in invalid:
            del self.map_server[fid]
        if len(processed) == 0:
            return
        H_x = np.zeros((rows, 21 + 6 * len(ss._vio_cam_states__)))
        r = np.zeros(rows)
        stack = 0
        for fid in processed:
            feat = self.map_server[fid]
            cam_ids = list(feat.observations.keys())
            H_xj, r_j = self.feature_jacobian(feat.id, cam_ids)
            if self.gating_test(H_xj, r_j, len(cam_ids) - 1):
                H_x[stack:stack + H_xj.shape[0], :H_xj.shape[1]] = H_xj
                r[stack:stack + len(r_j)] = r_j
                stack += H_xj.shape[0]
            if stack > self.ROW_CAP:
                break
        self.measurement_update(H_x[:stack], r[:stack])
        for fid in processed:
            del self.map_server[fid]

    # ---- msckf.py:691-727
    def find_redundant_cam_states(self):
        pairs = list(self.state_server._vio_cam_states__.items())
        key_idx = len(pairs) - 4
        idx = key_idx + 1
        first_idx = 0
        key_pos = pairs[key_idx][1]._vio_position__
        key_rot = to_rotation(pairs[key_idx][1].orientation)
        rm = []
        for _ in range(2):
            pos = pairs[idx][1]._vio_position__
            rot = to_rotation(pairs[idx][1].orientation)
            distance = np.linalg.norm(pos - key_pos)
            angle = 2 * np.arccos(to_quaternion(rot @ key_rot.T)[-1])
            if angle < 0.2618 and distance < 0.4 and self.tracking_rate > 0.5:
                rm.append(pairs[idx][0])
                idx += 1
            else:
                rm.append(pairs[first_idx][0])
                first_idx += 1
                idx += 1
        return sorted(rm)

    # ---- msckf.py:730-818
    def prune_cam_state_buffer(self):
        ss = self.state_server
        if len(ss._vio_cam_states__) < self.config._vio_max_cam_state_size__:
            return
        rm_ids = self.find_redundant_cam_states()
        self._log('rm_cam_ids', list(rm_ids))
        rows = 0
        for feat in self.map_server.values():
            involved = [c for c in rm_ids if c in feat.observations]
            if len(involved) == 0:
                continue
            if len(involved) == 1:
                del feat.observations[involved[0]]
                continue
            if not feat.is_initialized:
                if (not feat.check_motion(ss._vio_cam_states__)
                        or feat.initialize_position(ss._vio_cam_states__) is False):
                    for c in involved:
                        del feat.observations[c]
                    continue
            rows += 4 * len(involved) - 3
        H_x = np.zeros((rows, 21 + 6 * len(ss._vio_cam_states__)))
        r = np.zeros(rows)
        stack = 0
        used = []
        for feat in self.map_server.values():
            involved = [c for c in rm_ids if c in feat.observations]
            if len(involved) == 0:
                continue
            used.append(feat.id)
            H_xj, r_j = self.feature_jacobian(feat.id, involved)
            if self.gating_test(H_xj, r_j, len(involved)):
                H_x[stack:stack + H_xj.shape[0], :H_xj.shape[1]] = H_xj
                r[stack:stack + len(r_j)] = r_j
                stack += H_xj.shape[0]
            for c in involved:
                del feat.observations[c]
        self._log('prune_processed', used)
        self.measurement_update(H_x[:stack], r[:stack])
        for cam_id in rm_ids:
            idx = list(ss._vio_cam_states__.keys()).index(cam_id)
            a, b = 21 + 6 * idx, 27 + 6 * idx
            P = ss._vio_state_cov__.copy()
            if b < P.shape[0]:
                P[a:-6, :] = P[b:, :]
                P[:, a:-6] = P[:, b:]
            ss._vio_state_cov__ = P[:-6, :-6]
            del ss._vio_cam_states__[cam_id]

    # ---- msckf.py:820-830
    def reset_state_cov(self):
        c = self.config
        P = np.zeros((21, 21))
        P[3:6, 3:6] = c._vio_gyro_bias_cov__ * np.identity(3)
        P[6:9, 6:9] = c._vio_velocity_cov__ * np.identity(3)
        P[9:12, 9:12] = c._vio_acc_bias_cov__ * np.identity(3)
        P[15:18, 15:18] = c._vio_extrinsic_rotation_cov__ * np.identity(3)
        P[18:21, 18:21] = c._vio_extrinsic_translation_cov__ * np.identity(3)
        self.state_server._vio_state_cov__ = P

    # ---- msckf.py:832-857
    def reset(self):
        old = self.state_server._vio_imu_state__
        imu = IMUState()
        imu.id, imu.R_imu_cam0, imu.t_cam0_imu = old.id, old.R_imu_cam0, old.t_cam0_imu
        self.state_server._vio_imu_state__ = imu
        self.state_server._vio_cam_states__.clear()
        self.reset_state_cov()
        self.map_server.clear()
        self.imu_msg_buffer.clear()
        self.is_gravity_set = False
        self.is_first_img = True

    # ---- msckf.py:859-886
    def online_reset(self):
        thr = self.config._vio_position_std_threshold__
        if thr <= 0:
            return
        P = self.state_server._vio_state_cov__
        if max(np.sqrt(P[12, 12]), np.sqrt(P[13, 13]), np.sqrt(P[14, 14])) < thr:
            return
        self._log('online_reset', True)
        self.state_server._vio_cam_states__.clear()
        self.map_server.clear()
        self.reset_state_cov()

    # ---- msckf.py:888-908
    def publish(self, time):
        imu = self.state_server._vio_imu_state__
        T_i_w = Isometry3d(to_rotation(imu.orientation).T, imu._vio_position__)
        T_b_w = self.T_imu_body * T_i_w * self.T_imu_body.inverse()
        body_velocity = self.T_imu_body._vio_R__ @ imu.velocity
        R_w_c = imu.R_imu_cam0 @ T_i_w._vio_R__.T
        t_c_w = imu._vio_position__ + T_i_w._vio_R__ @ imu.t_cam0_imu
        return VioResult(time, T_b_w, body_velocity, Isometry3d(R_w_c.T, t_c_w))
