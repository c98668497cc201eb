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
