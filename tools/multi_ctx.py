"""Scratch: F filters split over G contexts (one CUDA stream each) driven round-robin from one host thread.
usage: python tools/multi_ctx.py F G [G ...]"""
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
GS = [int(x) for x in sys.argv[2:]] or [1, 2, 4, 8]
NF = 150
cfg = ConfigEuRoC(); cfg._vio_max_cam_state_size__ = 20
st = synth.make_structure(NF, 300, structure_seed=0)
gyro, acc, uvs = bench.build_frames(st, np.random.default_rng(1000), F, NF)
dev = torch.device('cuda', 0)
for G in GS:
    Fg = F // G
    batches, dfr = [], []
    for g in range(G):
        b = MSCKFBatch(cfg, Fg, device=0, feature_capacity=1024)
        sl = slice(g * Fg, (g + 1) * Fg)
        b.initialize_gravity_and_bias(gyro[sl, :200], acc[sl, :200])
        frames = [bench.frame_arrays(b, st, gyro[sl], acc[sl], [u[sl] for u in uvs], f) for f in range(NF)]
        dfr.append([dict((k, torch.from_numpy(fr[k]).to(dev)) for k in bench.KEYS) for fr in frames])
        batches.append(b)
    torch.cuda.synchronize()

    def run(f):
        for g in range(G):
            batches[g].ctx.feature_callback_dev(*[dfr[g][f][k].data_ptr() for k in bench.KEYS])

    def sync():
        for b in batches:
            b.ctx.sync()
    for f in range(90):
        run(f)
    sync()
    for lo, hi in ((90, 120), (120, 150)):
        t0 = time.perf_counter()
        for f in range(lo, hi):
            run(f)
        t_enq = time.perf_counter() - t0
        sync()
        dt = time.perf_counter() - t0
        print('F=%d G=%d frames %d-%d: %.3f ms/frame wall (enqueue %.3f ms/frame)' % (F, G, lo, hi, dt * 1e3 / (hi - lo), t_enq * 1e3 / (hi - lo)), flush=True)
    for b in batches:
        b.close()
    del batches, dfr
    torch.cuda.empty_cache()
