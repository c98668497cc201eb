"""CPU-side checks: the C ABI library exports every symbol include/msckf_b200.h declares, the host partition logic,
and the world_size-2 gather over gloo."""
import ctypes
import os
import re
import socket
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def _declared_symbols():
    text = open(os.path.join(REPO, 'include', 'msckf_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(msckf_[a-z0-9_]+)\s*\(', text)))


def test_header_symbols_are_exported():
    import __graft_entry__ as g
    g.build()
    from msckf_b200 import capi
    lib = ctypes.CDLL(capi.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 40, names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    lib.msckf_version.restype = ctypes.c_char_p
    assert b'sm_100a' in lib.msckf_version()
    assert lib.msckf_kernel_count() > 10


def test_no_gpu_is_loud():
    """Without a CUDA device the product path fails loudly instead of falling back to the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from msckf_b200 import capi
    from msckf_b200.config import ConfigEuRoC
    with pytest.raises(capi.MsckfCudaError):
        capi.Context(capi.make_cconfig(ConfigEuRoC()))


def test_package_does_not_import_the_oracle():
    pkg = os.path.join(REPO, 'visual-inertial-odometry-msckf-stereo_b200')
    for fn in os.listdir(pkg):
        if fn.endswith('.py'):
            src = open(os.path.join(pkg, fn)).read()
            assert 'import oracle' not in src and 'from oracle' not in src, fn


def test_filter_partition():
    from msckf_b200.partition import filter_range, owner_of
    for total in (1, 7, 1024, 1000):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                lo, hi = filter_range(r, world, total)
                assert 0 <= lo <= hi <= total
                seen.extend(range(lo, hi))
                for k in range(lo, hi):
                    assert owner_of(k, world, total) == r
            assert seen == list(range(total))
    assert filter_range(3, 8, 1024) == (384, 512)
    with pytest.raises(ValueError):
        filter_range(2, 2, 10)


def _gloo_worker(rank, world, port, total, out_q):
    import torch
    import torch.distributed as dist
    from msckf_b200.partition import filter_range, gather_blocks, reduce_max, reduce_sum
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    lo, hi = filter_range(rank, world, total)
    # a fake trajectory block: [frames, filters, 7] would be gathered along filters -> keep filters leading
    local = torch.arange(lo, hi, dtype=torch.float64).reshape(-1, 1, 1) * torch.ones((1, 5, 7), dtype=torch.float64)
    full = gather_blocks(local, world, rank)
    tmax = reduce_max(1.0 + rank, world)
    tsum = reduce_sum(float(hi - lo), world)
    if rank == 0:
        out_q.put((full.numpy(), tmax, tsum))
    dist.barrier()
    dist.destroy_process_group()


def test_gather_world_size_2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    total, world = 11, 2
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    full, tmax, tsum = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert full.shape == (total, 5, 7)
    assert np.array_equal(full[:, 0, 0], np.arange(total, dtype=np.float64))
    assert tmax == 2.0 and tsum == float(total)


def test_context_group_splits_ragged_batches():
    """capi.ContextGroup (a batch spread over several contexts): ragged batch arguments are cut at the group boundaries with
    rebased offsets, per-filter calls reach the owning context with the local index, read-backs are concatenated in filter
    order.  Host logic only: the contexts are stand-ins that record what they receive."""
    from msckf_b200 import capi

    class Part(object):
        def __init__(self, F, tag):
            self.F, self.tag, self.calls = F, tag, []

        def feature_callback_async(self, *a):
            self.calls.append(('frame',) + tuple(np.array(x, copy=True) for x in a))

        def collect(self):
            return np.full((self.F, capi.IMU_STRIDE), float(self.tag)), np.zeros(self.F, dtype=bool)

        def get_cov(self, f, n=None):
            return (self.tag, f)

        def counters(self):
            return np.full((self.F, capi.N_COUNTERS), self.tag, dtype=np.int64)

        def kernel_launches(self):
            return 10 * (self.tag + 1)

        def close(self):
            pass

    g = object.__new__(capi.ContextGroup)
    g.parts = [Part(3, 0), Part(2, 1)]
    g.bounds = np.array([0, 3, 5], dtype=np.int64)
    g.F = 5
    for name in capi.ContextGroup.PER_FILTER:
        if hasattr(Part, name):
            setattr(g, name, g._route(name))
    assert g.ranges() == [(0, 3), (3, 5)]
    # ragged: filter k has k + 1 IMU samples and 2 k features
    imu_n, feat_n = [1, 2, 3, 4, 5], [0, 2, 4, 6, 8]
    imu_off = np.concatenate([[0], np.cumsum(imu_n)]).astype(np.int32)
    feat_off = np.concatenate([[0], np.cumsum(feat_n)]).astype(np.int32)
    ts = np.arange(imu_off[-1], dtype=np.float64)
    gyro = np.arange(3 * imu_off[-1], dtype=np.float64).reshape(-1, 3)
    acc = -gyro
    ids = np.arange(feat_off[-1], dtype=np.int64) + 100
    uv = np.arange(4 * feat_off[-1], dtype=np.float64).reshape(-1, 4)
    cam_ids = np.array([7, 7, 7, 8, 8], dtype=np.int64)
    g.feature_callback_async(imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv)
    a, b = g.parts[0].calls[0], g.parts[1].calls[0]
    assert a[1].tolist() == [0, 1, 3, 6] and b[1].tolist() == [0, 4, 9]                 # rebased IMU offsets
    assert np.array_equal(a[2], ts[:6]) and np.array_equal(b[2], ts[6:])
    assert np.array_equal(a[3], gyro[:6]) and np.array_equal(b[4], acc[6:])
    assert a[5].tolist() == [7, 7, 7] and b[5].tolist() == [8, 8]
    assert a[6].tolist() == [0, 0, 2, 6] and b[6].tolist() == [0, 6, 14]                # rebased feature offsets
    assert np.array_equal(a[7], ids[:6]) and np.array_equal(b[7], ids[6:])
    assert np.array_equal(a[8], uv[:6]) and np.array_equal(b[8], uv[6:])
    imu, rst = g.collect()
    assert imu.shape == (5, capi.IMU_STRIDE) and imu[:3].max() == 0.0 and imu[3:].min() == 1.0 and rst.shape == (5,)
    assert g.get_cov(0) == (0, 0) and g.get_cov(2) == (0, 2) and g.get_cov(3) == (1, 0) and g.get_cov(4) == (1, 1)
    with pytest.raises(capi.MsckfCudaError):
        g.get_cov(5)
    assert g.counters()[:, 0].tolist() == [0, 0, 0, 1, 1]
    assert g.kernel_launches() == 30


def test_device_frame_parts_addresses():
    """MSCKFBatch.device_frame_parts: the per-group addresses of a frame resident in (here: host) tensors are the slices of the
    ragged arrays at the group boundaries, with the offsets rebased per group."""
    import torch
    from msckf_b200.msckf import MSCKFBatch

    class Ctx(object):
        def ranges(self):
            return [(0, 2), (2, 5)]

    b = object.__new__(MSCKFBatch)
    b.groups, b.ctx, b.F = 2, Ctx(), 5
    keys = ['imu_off', 'ts', 'gyro', 'acc', 'cam_ids', 'feat_off', 'ids', 'uv']
    imu_off = torch.tensor([0, 2, 4, 7, 7, 9], dtype=torch.int32)
    feat_off = torch.tensor([0, 3, 3, 5, 9, 10], dtype=torch.int32)
    frame = {'imu_off': imu_off, 'feat_off': feat_off, 'ts': torch.arange(9, dtype=torch.float64),
             'gyro': torch.zeros((9, 3), dtype=torch.float64), 'acc': torch.ones((9, 3), dtype=torch.float64),
             'cam_ids': torch.arange(5, dtype=torch.int64), 'ids': torch.arange(10, dtype=torch.int64),
             'uv': torch.zeros((10, 4), dtype=torch.float64)}
    parts = b.device_frame_parts(frame, keys)
    assert len(parts) == 2
    g0, g1 = [dict(zip(keys, p)) for p in parts]
    assert g0['ts'] == frame['ts'].data_ptr() and g1['ts'] == frame['ts'].data_ptr() + 4 * 8
    assert g1['gyro'] == frame['gyro'].data_ptr() + 4 * 3 * 8 and g1['acc'] == frame['acc'].data_ptr() + 4 * 3 * 8
    assert g1['cam_ids'] == frame['cam_ids'].data_ptr() + 2 * 8
    assert g0['ids'] == frame['ids'].data_ptr() and g1['ids'] == frame['ids'].data_ptr() + 3 * 8
    assert g1['uv'] == frame['uv'].data_ptr() + 3 * 4 * 8
    (o0, f0), (o1, f1) = frame['_group_offsets']
    assert o0.tolist() == [0, 2, 4] and o1.tolist() == [0, 3, 3, 5]
    assert f0.tolist() == [0, 3, 3] and f1.tolist() == [0, 2, 6, 7]
    assert g0['imu_off'] == o0.data_ptr() and g1['feat_off'] == f1.data_ptr()
    # one group: the frame's own addresses
    b.groups = 1
    assert b.device_frame_parts(frame, keys) == [tuple(frame[k].data_ptr() for k in keys)]
