"""Shared by the scratch tools: the headline workload's messages for one context, generated on the GPU (bench.DeviceStream)."""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch  # noqa: E402

import bench  # noqa: E402
from msckf_b200 import synth  # noqa: E402
from msckf_b200.config import ConfigEuRoC  # noqa: E402
from msckf_b200.msckf import MSCKFBatch  # noqa: E402

KEYS = bench.KEYS


def make_batch(F, n_frames, structure=None, seed0=0, max_cam=20, feats=300, device=0, groups=1):
    """(batch, structure, device frames): a context of F filters (noise seeds seed0 ..) with gravity initialised and the
    messages of n_frames frames resident in HBM."""
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    st = structure if structure is not None else synth.make_structure(n_frames, feats, structure_seed=0)
    batch = MSCKFBatch(cfg, F, device=device, feature_capacity=1024, groups=groups)
    ds = bench.DeviceStream(st, batch, F, seed0, device, torch)
    g0, a0 = ds.static_samples()
    batch.initialize_gravity_and_bias(g0, a0)
    frames = [ds.frame(f) for f in range(n_frames)]
    torch.cuda.synchronize()
    return batch, st, frames


def host_copy(frame):
    """numpy copies of a frame's arrays, in the order the C ABI takes them"""
    return [frame[k].cpu().numpy().reshape(-1, *frame[k].shape[2:]) if frame[k].dim() == 3 else frame[k].cpu().numpy() for k in KEYS]


def dev_ptrs(frame):
    return [frame[k].data_ptr() for k in KEYS]
