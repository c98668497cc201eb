"""Drop-in ``Feature`` (feature.py:7-295): the landmark container with its triangulation
methods routed to the CUDA kernels of ``libmsckf_b200.so``.

``check_motion``, ``generate_initial_guess`` and ``initialize_position`` keep the reference's
signatures; each packs the landmark's observations with the poses of the observing camera states
and makes one C-ABI call (``msckf_check_motion`` / ``msckf_initial_guess`` / ``msckf_triangulate``).
``triangulate_batch`` is the form the hardware wants: any number of landmarks in one launch, one
warp per landmark.  Inside ``MSCKF.feature_callback`` none of this is used -- the device-resident
map triangulates in place -- these entry points exist for callers that use ``Feature`` directly.
"""
import numpy as np

from . import capi
from .utils import to_quaternion

_shared_ctx = {}


def shared_context(optimization_config, R_cam0_cam1, t_cam0_cam1, device=0):
    """A small context (one filter, no map) used only for its stateless kernels."""
    key = (device, optimization_config._vio_translation_threshold__, optimization_config._vio_huber_epsilon__,
           optimization_config._vio_estimation_precision__, optimization_config._vio_initial_damping__,
           optimization_config._vio_outer_loop_max_iteration__, optimization_config._vio_inner_loop_max_iteration__,
           tuple(np.asarray(R_cam0_cam1).ravel()), tuple(np.asarray(t_cam0_cam1).ravel()))
    ctx = _shared_ctx.get(key)
    if ctx is None:
        from .config import ConfigEuRoC
        cfg = ConfigEuRoC()
        cfg._vio_optimization_config__ = optimization_config
        T = np.identity(4)
        T[:3, :3], T[:3, 3] = R_cam0_cam1, t_cam0_cam1
        cfg._vio_T_cn_cnm1__ = T
        cfg._vio_max_cam_state_size__ = 8
        ctx = capi.Context(capi.make_cconfig(cfg, n_filters=1, feature_capacity=8, device=device))
        _shared_ctx[key] = ctx
    return ctx


def pack_observations(features, cam_states):
    """CSR packing of landmarks for the stateless kernels: observations whose camera state is gone
    are skipped, as in feature.py:189-192."""
    off = [0]
    uv, q, p = [], [], []
    for ft in features:
        for cam_id, z in ft.observations.items():
            cs = cam_states.get(cam_id)
            if cs is None:
                continue
            uv.append(z)
            q.append(cs.orientation)
            p.append(cs._vio_position__)
        off.append(len(uv))
    n = len(uv)
    return (np.array(off, dtype=np.int32), np.array(uv, dtype=np.float64).reshape(n, 4),
            np.array(q, dtype=np.float64).reshape(n, 4), np.array(p, dtype=np.float64).reshape(n, 3))


class Feature(object):
    _vio_next_id__ = 0
    # cam0 -> cam1 extrinsics, set by MSCKF.__init__ in the reference (msckf.py:148-152)
    _vio_R_cam0_cam1__ = None
    _vio_t_cam0_cam1__ = None
    device = 0

    def __init__(self, new_id=0, optimization_config=None):
        self.id = new_id
        self.observations = dict()
        self._vio_position__ = np.zeros(3)
        self.is_initialized = False
        self.optimization_config = optimization_config

    def _ctx(self):
        return shared_context(self.optimization_config, Feature._vio_R_cam0_cam1__, Feature._vio_t_cam0_cam1__, Feature.device)

    def check_motion(self, cam_states):
        """feature.py:124-165"""
        off, uv, q, p = pack_observations([self], cam_states)
        return bool(self._ctx().check_motion(off, uv, q, p)[0])

    def generate_initial_guess(self, T_c1_c2, z1, z2):
        """feature.py:99-122: two-view depth of the landmark in the c1 frame.  T_c1_c2 maps c1
        coordinates to c2 coordinates; the kernel takes camera poses, so c1 is placed at the world
        origin and c2 at T_c1_c2^-1."""
        R, t = np.asarray(T_c1_c2._vio_R__, dtype=np.float64), np.asarray(T_c1_c2._vio_t__, dtype=np.float64)
        q = np.array([[0., 0., 0., 1.], to_quaternion(R)])
        p = np.array([np.zeros(3), -R.T @ t])
        uv = np.array([[z1[0], z1[1], 0., 0.], [z2[0], z2[1], 0., 0.]])
        return self._ctx().initial_guess([0, 2], uv, q, p)[0]

    def initialize_position(self, cam_states):
        """feature.py:167-295: writes ``_vio_position__`` (also when the result is invalid, as the
        reference does) and ``is_initialized``."""
        off, uv, q, p = pack_observations([self], cam_states)
        pos, valid, _ = self._ctx().triangulate(off, uv, q, p, check_motion=False)
        self._vio_position__ = pos[0]
        self.is_initialized = bool(valid[0])
        return self.is_initialized


def triangulate_batch(features, cam_states, ctx=None):
    """check_motion + initialize_position for many landmarks in one launch.  Returns the validity
    flags; positions / flags are written into the Feature objects like the per-object calls."""
    if not features:
        return np.zeros(0, dtype=bool)
    ctx = ctx or features[0]._ctx()
    off, uv, q, p = pack_observations(features, cam_states)
    pos, valid, _ = ctx.triangulate(off, uv, q, p)
    for k, ft in enumerate(features):
        ft._vio_position__ = pos[k]
        ft.is_initialized = bool(valid[k])
    return valid
