"""Scratch: device-resident frame loop timed three ways."""
import os, sys, time
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch
import bench
from msckf_b200 import capi, synth
from msckf_b200.config import ConfigEuRoC
from msckf_b200.msckf import MSCKFBatch
F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
NF = 160
cfg = ConfigEuRoC(); cfg._vio_max_cam_state_size__ = 20
st = synth.make_structure(NF, 300, structure_seed=0)
gyro, acc, uvs = bench.build_frames(st, np.random.default_rng(1000), F, NF)
b = MSCKFBatch(cfg, F, device=0, feature_capacity=1024)
b.initialize_gravity_and_bias(gyro[:, :200], acc[:, :200])
frames = [bench.frame_arrays(b, st, gyro, acc, uvs, f) for f in range(NF)]
dev = torch.device('cuda', 0)
dfr = [dict((k, torch.from_numpy(fr[k]).to(dev)) for k in bench.KEYS) for fr in frames]
torch.cuda.synchronize()
ctx = b.ctx
def run(f): ctx.feature_callback_dev(*[dfr[f][k].data_ptr() for k in bench.KEYS])
for f in range(40): run(f)
ctx.sync()
def timed(lo, hi, sync_each):
    t0 = time.perf_counter(); ctx.mark(0)
    for f in range(lo, hi):
        run(f)
        if sync_each: ctx.sync()
    ctx.mark(1); ms = ctx.mark_elapsed_ms(); ctx.sync()
    print('frames %d-%d sync_each=%s: dev %.3f ms/frame wall %.3f' % (lo, hi, sync_each, ms / (hi - lo), (time.perf_counter() - t0) * 1e3 / (hi - lo)))
def prof(lo, hi):
    ctx.set_kernel_timing(True)
    k0 = ctx.kernel_times()
    for f in range(lo, hi): run(f)
    ctx.sync()
    k1 = ctx.kernel_times()
    ctx.set_kernel_timing(False)
    tot = sum(k1[n][2] - k0.get(n, (0, 0, 0))[2] for n in k1)
    print('frames %d-%d kernel sum %.3f ms/frame' % (lo, hi, tot / (hi - lo)))
    for n in sorted(k1, key=lambda n: -(k1[n][2] - k0.get(n, (0, 0, 0))[2]))[:8]:
        print('   %-16s %.3f' % (n, (k1[n][2] - k0.get(n, (0, 0, 0))[2]) / (hi - lo)))
prof(40, 55)
prof(55, 70)
prof(70, 85)
timed(85, 115, False)
timed(115, 145, False)
