"""msckf_gating_test_noalloc (include/msckf_b200.h): the gating test with its scratch in a static device array -- no
allocation inside the call -- against the existing msckf_gating_test, through the C ABI.  Two calls in a row of different
shapes (the second one runs over the first one's left-overs in the shared array), then the size limit.

Collected last on purpose (file name)."""
import numpy as np
import pytest

from msckf_b200 import capi
from msckf_b200.config import ConfigEuRoC

pytestmark = pytest.mark.gpu

TOL = 1e-9     # the tolerance of the other GPU tests; the two entry points run the same kernel, so they agree to the bit


def make_context(n_cams):
    """One filter with a window of n_cams camera states and an SPD covariance with cross terms."""
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = 20
    ctx = capi.Context(capi.make_cconfig(cfg, n_filters=1, feature_capacity=64))
    rng = np.random.default_rng(7)
    cams = np.zeros((n_cams, capi.CAM_STRIDE))
    for c in range(n_cams):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        cams[c, 0:4] = q
        cams[c, 4:7] = rng.normal(size=3)
        cams[c, 7:11] = q
        cams[c, 11:14] = cams[c, 4:7]
    ctx.set_cam_states(0, np.arange(n_cams, dtype=np.int64), cams)
    n = 21 + 6 * n_cams
    A = rng.normal(size=(n, n))
    ctx.set_cov(0, 1e-3 * (A @ A.T / n + np.eye(n)))
    return ctx, n


def test_gating_test_noalloc_equals_gating_test():
    import torch
    ctx, n = make_context(6)
    try:
        rng = np.random.default_rng(11)
        dev = torch.device('cuda', 0)
        gam = torch.zeros(1, dtype=torch.float64, device=dev)
        pas = torch.full((1,), -7, dtype=torch.int32, device=dev)
        seen = []
        # two calls in a row: a larger shape first, then a smaller one on top of its left-overs; then the first shape again
        for m, dof, scale in ((17, 9, 0.02), (5, 5, 0.5), (17, 9, 0.02)):
            H = rng.normal(size=(m, n))
            r = scale * rng.normal(size=m)
            g_ref, p_ref = ctx.gating_test(0, H, r, dof)
            Hd = torch.from_numpy(H).to(dev)
            rd = torch.from_numpy(r).to(dev)
            torch.cuda.synchronize()
            ctx.gating_test_noalloc(0, Hd.data_ptr(), rd.data_ptr(), m, n, dof, gam.data_ptr(), pas.data_ptr())
            ctx.sync()
            g, p = float(gam.cpu()[0]), int(pas.cpu()[0])
            assert np.isfinite(g_ref) and g_ref > 0
            assert abs(g - g_ref) <= TOL * abs(g_ref), (m, g, g_ref)
            assert p == int(p_ref), (m, p, p_ref)
            # and against the definition (msckf.py:606-614): gamma = r^T (H P H^T + sigma^2 I)^-1 r
            P = ctx.get_cov(0)
            S = H @ P @ H.T + ctx.cfg.observation_noise * np.eye(m)
            g_np = float(r @ np.linalg.solve(S, r))
            assert abs(g - g_np) <= 1e-8 * abs(g_np), (g, g_np)
            seen.append(p)
        assert 0 in seen or 1 in seen
    finally:
        ctx.close()


def test_gating_test_noalloc_rejects_sizes_beyond_the_static_array():
    import torch
    ctx, n = make_context(6)
    try:
        dev = torch.device('cuda', 0)
        m = capi.GATE_NOALLOC_MAX_DOUBLES // n + 1           # m * n alone exceeds the array
        Hd = torch.zeros(8, dtype=torch.float64, device=dev)  # never read: the call must refuse before it launches
        gam = torch.zeros(1, dtype=torch.float64, device=dev)
        pas = torch.zeros(1, dtype=torch.int32, device=dev)
        with pytest.raises(capi.MsckfCudaError) as e:
            ctx.gating_test_noalloc(0, Hd.data_ptr(), Hd.data_ptr(), m, n, 5, gam.data_ptr(), pas.data_ptr())
        assert e.value.code == capi.ERR_CAPACITY
        assert 'static array' in str(e.value)
        # the largest shape that fits at this state size is accepted
        m_ok = 1
        while (m_ok + 1) * n + (m_ok + 1) ** 2 + (m_ok + 1) <= capi.GATE_NOALLOC_MAX_DOUBLES:
            m_ok += 1
        rng = np.random.default_rng(3)
        H = rng.normal(size=(m_ok, n))
        r = 0.05 * rng.normal(size=m_ok)
        Hd, rd = torch.from_numpy(H).to(dev), torch.from_numpy(r).to(dev)
        torch.cuda.synchronize()
        ctx.gating_test_noalloc(0, Hd.data_ptr(), rd.data_ptr(), m_ok, n, 50, gam.data_ptr(), pas.data_ptr())
        ctx.sync()
        g_ref, p_ref = ctx.gating_test(0, H, r, 50)
        assert abs(float(gam.cpu()[0]) - g_ref) <= TOL * abs(g_ref)
        assert int(pas.cpu()[0]) == int(p_ref)
    finally:
        ctx.close()
