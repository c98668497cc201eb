"""Scratch: F filters split over G contexts (one CUDA stream each) driven round-robin from one host thread.
usage: python tools/multi_ctx.py F G [G ...]"""
import sys
import time

from _stream import make_batch, dev_ptrs
import torch

F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
GS = [int(x) for x in sys.argv[2:]] or [1, 2, 4]
NF = 150
st = None
for G in GS:
    Fg = F // G
    batches, dfr = [], []
    for g in range(G):
        b, st, frames = make_batch(Fg, NF, structure=st, seed0=g * Fg)
        batches.append(b)
        dfr.append(frames)

    def run(f):
        for g in range(G):
            batches[g].ctx.feature_callback_dev(*dev_ptrs(dfr[g][f]))

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
