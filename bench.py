#!/usr/bin/env python
"""bench.py -- MSCKF back-end hot path on B200: feature-updates/s (BASELINE.json metric).

A "step" is one frame (feature_callback: propagate, augment, observe, remove_lost_features,
prune_cam_state_buffer, online_reset) over EVERY filter of the job.  Workload = BASELINE.json
configs[3]: independent filters (Monte-Carlo noise realisations of one synthetic
EuRoC-MH_01-shaped stream: 200 Hz IMU, 20 Hz stereo, ~300 features/frame,
max_cam_state_size = 20); filter k of the job uses noise seed k.  No collective on the data
path, one final gather of the trajectories.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

  value   feature-updates/s with every input of the timed frames already in HBM (device pointers
          through msckf_feature_callback_dev).  The timed region is R blocks of K frames (R chosen so
          that it lasts >= ~0.5 s), each block bracketed by barrier + synchronize and timed with CUDA
          events on the context's stream, max over ranks; value / ms_per_step are the MEDIAN block,
          min / max under "blocks".
  e2e     the same metric through the C ABI with HOST buffers (msckf_feature_callback_async /
          msckf_collect): H2D of every frame's message arrays from page-locked host memory, kernels,
          D2H of every frame's publish() snapshot; two frames in flight.  Same block scheme.
  scaling the headline keeps a fixed number of filters PER GPU (weak); at N > 1 a second leg runs
          BASELINE's 1024 filters IN TOTAL, block-partitioned over the ranks (partition.filter_range,
          filter k on rank floor(k N / 1024)) and is printed under "strong".
  parity_spot  after the device-timed frames, 8 of the filters (first, last, random) are downloaded
          and compared with the CPU oracle run on exactly the messages those filters received.
  roofline  dominant kernel family of a per-kernel CUDA-event pass over further frames of the same
          run: algorithmic FLOPs (device work counters, formulas in DESIGN.md) / its device time,
          against the FP64 peak measured on this GPU (msckf_fp64_peak; MEASURED_PEAKS.json has no FP64).
  c3, c5  BASELINE configs[2] (stress frame: 2000 lost landmarks, ~1.2e5 x 201 stacked Jacobian, one
          filter) and configs[4] (64 camera states, n = 405, 4096 features/frame, a batch of filters):
          ms per frame, per-family roofline fractions, the CPU oracle's time for the same frame.
  cpu_baseline  the CPU oracle (plain-numpy port of the reference, oracle/msckf_oracle.py) on one
          core of this box over a bounded sample of the same stream.
  latency single-filter (configs[1]: max_cam_state_size = 30) p50 / p99 of feature_callback, with the
          CPU oracle's latency on the same stream beside it.

All messages are generated on the GPU (msckf_synth_frame: counter-based noise on top of the shared
noise-free structure), so no per-filter data is produced by the host.

--impl reference: the reference algorithm's CPU port (the reference itself is Python and is not on
the GPU box) on all host cores, one filter per process, same stream shape, same metric.
"""
import argparse
import json
import os

os.environ.setdefault('OPENBLAS_NUM_THREADS', '1')   # before numpy loads: the CPU legs run one filter per core
os.environ.setdefault('OMP_NUM_THREADS', '1')
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

METRIC = 'msckf_feature_updates_per_s'
UNIT = 'feature-updates/s'
WINDOW_FILL = 80      # untimed frames that fill the 20-state window so the timed frames are steady state
PROFILE_FRAMES = 20   # frames of the per-kernel event pass
SIGMA_Z, OUTLIER_FRAC, OUTLIER_SIGMA = 0.001, 0.05, 0.05
TOTAL_FILTERS = 1024  # BASELINE configs[3]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=100)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--filters-per-gpu', type=int, default=TOTAL_FILTERS)
    ap.add_argument('--feats', type=int, default=300)
    ap.add_argument('--max-cam', type=int, default=20)
    ap.add_argument('--timed-seconds', type=float, default=0.5, help='lower bound of each timed region (sets the number of blocks)')
    ap.add_argument('--clock-warm-seconds', type=float, default=0.5, help='own frames run before anything is timed')
    ap.add_argument('--no-latency', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-parity-spot', action='store_true')
    ap.add_argument('--no-c3', action='store_true')
    ap.add_argument('--no-c5', action='store_true')
    ap.add_argument('--no-strong', action='store_true')
    ap.add_argument('--c5-filters', type=int, default=128)
    ap.add_argument('--groups', type=int, default=0, help='contexts (streams) a GPU\'s filters are spread over; 0 = the library default (2 from 512 filters on)')
    ap.add_argument('--only', default='', help='development: comma list of c3,c5 -- run just these legs and print them')
    return ap.parse_args()


def workload_config(args, n_gpus, filters_total=None):
    total = args.filters_per_gpu * n_gpus if filters_total is None else filters_total
    return {'workload': 'configs[3]: %d independent filters (Monte-Carlo noise seeds 0..%d of one synthetic EuRoC-MH_01-shaped '
                        'stream), %d features/frame, max_cam_state_size=%d, 10 IMU samples/frame; step = one frame over all filters'
                        % (total, total - 1, args.feats, args.max_cam),
            'filters_per_gpu': total // n_gpus, 'filters_total': total,
            'features_per_frame': args.feats, 'max_cam_state_size': args.max_cam, 'lead_in_frames': WINDOW_FILL,
            'l2': 'inputs of successive frames are distinct buffers (12.9 MB per frame, several GB in total: far beyond the 126 MB L2); the '
                  'filter state (P, map, Jacobian pool: > 1 GB) is resident and re-read every frame as in production',
            'partition': 'filters block-partitioned over GPUs, no data-path collective'}


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler(object):
    def __init__(self, device):
        self.device = device
        self.rows = []
        self.proc = None

    def start(self):
        q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.device), '--query-gpu=' + q, '--format=csv,noheader,nounits',
                                          '-lms', '100'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            p = [x.strip() for x in r.split(',')]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for nm, v in zip(names, p[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(nm)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': float(max(mx)) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


# ------------------------------------------------------------------------------------------ messages (generated on the GPU)
KEYS = ['imu_off', 'ts', 'gyro', 'acc', 'cam_ids', 'feat_off', 'ids', 'uv']


class DeviceStream(object):
    """Lock-step messages of F Monte-Carlo filters (noise seeds seed0 .. seed0+F-1) of the structure `st`, generated frame
    by frame on the device (msckf_synth_frame).  The host only walks the shared structure: which IMU samples belong to
    which frame (batch_imu_processing's selection, msckf.py:267-281) and the tracked ids."""

    def __init__(self, st, batch, F, seed0, device, torch):
        from msckf_b200 import synth
        from msckf_b200.config import ConfigEuRoC
        cfg = ConfigEuRoC()
        self.st, self.batch, self.F, self.seed0, self.device, self.torch = st, batch, F, seed0, device, torch
        self.dev = torch.device('cuda', device)
        self.sg = float(np.sqrt(cfg._vio_gyro_noise__ / synth.IMU_DT))
        self.sa = float(np.sqrt(cfg._vio_acc_noise__ / synth.IMU_DT))
        self.plan = {}

    def _alloc(self, K, n):
        t, F, dev = self.torch, self.F, self.dev
        return {'imu_off': t.empty(F + 1, dtype=t.int32, device=dev), 'ts': t.empty(F * K, dtype=t.float64, device=dev),
                'gyro': t.empty((F, K, 3), dtype=t.float64, device=dev), 'acc': t.empty((F, K, 3), dtype=t.float64, device=dev),
                'cam_ids': t.empty(F, dtype=t.int64, device=dev), 'feat_off': t.empty(F + 1, dtype=t.int32, device=dev),
                'ids': t.empty(F * n, dtype=t.int64, device=dev), 'uv': t.empty((F, n, 4), dtype=t.float64, device=dev)}

    def static_samples(self, count=200):
        """The first `count` IMU samples of every filter (the static lead-in of gravity / bias initialisation)."""
        from msckf_b200 import capi
        st = self.st
        d = self._alloc(count, 0)
        capi.synth_frame(self.device, self.F, self.seed0, -1, 0, st.imu_ts[:count], st.gyro_clean[:count], st.acc_clean[:count],
                         st.gyro_bias, st.acc_bias, self.sg, self.sa, np.zeros(0, dtype=np.int64), np.zeros((0, 4)), SIGMA_Z,
                         OUTLIER_FRAC, OUTLIER_SIGMA, -1, [d[k].data_ptr() for k in KEYS])
        return d['gyro'].cpu().numpy(), d['acc'].cpu().numpy()

    def frame(self, f):
        """Device tensors of frame f (frames must be requested in order: the IMU selection is stateful)."""
        from msckf_b200 import capi
        st, batch = self.st, self.batch
        b = int(np.searchsorted(st.imu_ts, st.frame_ts[f], side='right'))
        a = 0 if f == 0 else int(np.searchsorted(st.imu_ts, st.frame_ts[f - 1], side='right'))
        take = np.flatnonzero(batch.select_imu(float(st.frame_ts[f]), st.imu_ts[a:b])) + a
        cam_id = batch.next_imu_id
        batch.next_imu_id += 1
        ids = st.frame_ids[f].astype(np.int64)
        d = self._alloc(len(take), len(ids))
        sample0 = int(take[0]) if len(take) else a
        assert len(take) == 0 or np.array_equal(take, np.arange(sample0, sample0 + len(take)))
        capi.synth_frame(self.device, self.F, self.seed0, f, sample0, st.imu_ts[take], st.gyro_clean[take], st.acc_clean[take],
                         st.gyro_bias, st.acc_bias, self.sg, self.sa, ids, st.frame_uv[f], SIGMA_Z, OUTLIER_FRAC, OUTLIER_SIGMA,
                         cam_id, [d[k].data_ptr() for k in KEYS])
        self.plan[f] = (st.imu_ts[take].copy(), ids)
        return d


def to_pinned(d, torch):
    """Host copy of a frame's arrays in ONE page-locked slab (what a feeder thread would hand to the C ABI)."""
    from msckf_b200 import capi
    dt = {'imu_off': np.int32, 'feat_off': np.int32, 'cam_ids': np.int64, 'ids': np.int64}
    offs, total = {}, 0
    for k in KEYS:
        offs[k] = total
        total += (d[k].numel() * d[k].element_size() + 255) & ~255
    slab = capi.pinned_empty(max(total, 256), np.uint8)
    out = {}
    for k in KEYS:
        n = d[k].numel()
        h = slab[offs[k]:offs[k] + n * d[k].element_size()].view(dt.get(k, np.float64)).reshape(tuple(d[k].shape))
        if n:
            torch.from_numpy(h).copy_(d[k])
        out[k] = h
    out['_slab'] = slab
    return out


_STRUCTURES = {}


def get_structure(n_frames, feats):
    """The shared noise-free structure (seed 0); a longer one already built serves a shorter request (frames are generated
    sequentially, so the first n frames do not depend on how many follow)."""
    from msckf_b200 import synth
    for (nf, ft), st in _STRUCTURES.items():
        if ft == feats and nf >= n_frames:
            return st
    st = synth.make_structure(n_frames, feats, structure_seed=0)
    _STRUCTURES[(n_frames, feats)] = st
    return st


# ------------------------------------------------------------------------------------------ CPU arms
def oracle_filter_run(seed, n_lead, n_warm, n_timed, feats, max_cam, structure_frames):
    """One oracle filter over lead-in + warm-up + timed frames; returns (feature-updates, seconds, per-frame seconds)."""
    os.environ.setdefault('OPENBLAS_NUM_THREADS', '1')
    from oracle import msckf_oracle as orc
    from msckf_b200 import synth
    st = synth.make_structure(structure_frames, feats, structure_seed=0)
    stream = synth.realize(st, noise_seed=seed)
    cfg = orc.ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    filt = orc.MSCKF(cfg)
    msgs = stream.imu_msgs()
    k = 0
    updates, t_total, per_frame = 0, 0.0, []
    for f in range(n_lead + n_warm + n_timed):
        while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
            filt.imu_callback(msgs[k])
            k += 1
        msg = stream.feature_msg(f)
        filt.log = {}
        t0 = time.perf_counter()
        filt.feature_callback(msg)
        dt = time.perf_counter() - t0
        if f >= n_lead + n_warm:
            updates += len(filt.log.get('gate', []))
            t_total += dt
            per_frame.append(dt)
    return updates, t_total, per_frame


def _oracle_worker(a):
    return oracle_filter_run(*a)


def cpu_baseline_leg(args):
    """Bounded sample on one core: one oracle filter, WINDOW_FILL + 3 + 100 frames of the same stream shape."""
    n_timed = 100
    t0 = time.perf_counter()
    upd, sec, per = oracle_filter_run(0, WINDOW_FILL, 3, n_timed, args.feats, args.max_cam, WINDOW_FILL + 3 + n_timed)
    return {'value': upd / sec, 'unit': UNIT, 'cores': 1, 'kind': 'port',
            'sample': '1 filter x %d frames after %d lead-in frames of the same stream shape (numpy noise seed 0), oracle/msckf_oracle.py, '
                      'OPENBLAS_NUM_THREADS=1; %d feature-updates in %.1f s (whole leg %.1f s)' % (n_timed, WINDOW_FILL + 3, upd, sec,
                                                                                                 time.perf_counter() - t0),
            'p50_ms_per_frame': float(np.median(per) * 1e3), 'p99_ms_per_frame': float(np.percentile(per, 99) * 1e3),
            'filter_frames_per_s': n_timed / sec}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 64))
    n_timed = max(1, min(args.steps, 60))           # bounded sample: each step = one frame of every worker's filter
    n_warm = max(0, min(args.warmup, 5))
    total = WINDOW_FILL + n_warm + n_timed
    jobs = [(k, WINDOW_FILL, n_warm, n_timed, args.feats, args.max_cam, total) for k in range(workers)]
    t0 = time.perf_counter()
    with mp.get_context('fork').Pool(workers) as pool:
        res = pool.map(_oracle_worker, jobs)
    wall = time.perf_counter() - t0
    upd = sum(r[0] for r in res)
    tmax = max(r[1] for r in res)
    value = upd / tmax
    cfg = workload_config(args, args.gpus)
    cfg['workload'] = ('configs[3] stream shape on the host cores: %d oracle filters (noise seeds 0..%d), one per process -- a bounded sample '
                       'of the %d-filter job (throughput metric: filters are independent); ' % (workers, workers - 1, cfg['filters_total'])
                       + cfg['workload'].split('stream), ', 1)[1])
    cfg['filters_run'] = workers
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': n_timed,
            'warmup': n_warm, 'ms_per_step': tmax / n_timed * 1e3, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f64', 'data': 'synthetic', 'config': cfg,
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': workers, 'kind': 'port',
                             'sample': '%d oracle filters (one per process, host has %d cores), %d timed frames each after %d lead-in '
                                       'frames; whole run %.1f s' % (workers, cores, n_timed, WINDOW_FILL + n_warm, wall)},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'filter_frames_per_s': workers * n_timed / tmax, 'gpu_launches': 0}
    print(json.dumps(line))


def spot_oracle_worker(job):
    """The CPU oracle on exactly the messages one filter of the GPU run received; returns its final snapshot."""
    os.environ.setdefault('OPENBLAS_NUM_THREADS', '1')
    from collections import namedtuple
    from oracle import msckf_oracle as orc
    from msckf_b200 import synth
    max_cam, init_ts, init_g, init_a, frames = job
    cfg = orc.ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    filt = orc.MSCKF(cfg)
    for t, g, a in zip(init_ts, init_g, init_a):
        filt.imu_callback(synth.ImuMsg(float(t), g.copy(), a.copy()))
    PackedMsg = namedtuple('PackedMsg', ['timestamp', 'vio_features'])
    for (t_frame, ts, g, a, ids, uv) in frames:
        for k in range(len(ts)):
            filt.imu_callback(synth.ImuMsg(float(ts[k]), g[k].copy(), a[k].copy()))
        feats = [synth.FeatureMeasurement(int(ids[i]), float(uv[i, 0]), float(uv[i, 1]), float(uv[i, 2]), float(uv[i, 3]))
                 for i in range(len(ids))]
        filt.feature_callback(PackedMsg(float(t_frame), feats))
    ss = filt.state_server
    imu = ss._vio_imu_state__
    return {'imu': np.concatenate([imu.orientation, imu._vio_gyro_bias__, imu.velocity, imu.acc_bias, imu._vio_position__]),
            'P': np.array(ss._vio_state_cov__), 'cam_ids': np.array(list(ss._vio_cam_states__.keys()), dtype=np.int64),
            'map_ids': np.array(list(filt.map_server.keys()), dtype=np.int64)}


# ------------------------------------------------------------------------------------------ roofline accounting
FAMILIES = {
    'triangulate': ['k_triangulate'],
    'jacobian_nullspace_gate': ['k_jacobian_gate', 'k_jacobian_rows', 'k_jacobian_big'],
    'stacked_qr': ['k_qr', 'k_qr_wide', 'k_qr_tall', 'k_qr_d16', 'k_qr_d32', 'k_qr_d48', 'k_qr_d64', 'k_qr_d80'],
    'gain_update': ['k_upd_dmma16', 'k_upd_dmma32', 'k_upd_dmma48', 'k_upd_dmma64', 'k_upd_dmma80', 'k_upd_T', 'k_upd_S', 'k_upd_chol',
                    'k_upd_trsm', 'k_upd_inject', 'k_upd_downdate'],
}


def family_stats(kt0, kt1, dc, dfma, dmma):
    """Per-family device time (CUDA events around every launch) against the algorithmic work of the device counters."""
    from msckf_b200 import capi
    kt = dict((n, (kt1[n][0] - kt0[n][0], kt1[n][1] - kt0[n][1], kt1[n][2] - kt0[n][2])) for n in kt1)
    work = {
        'triangulate': (632.0 * dc[capi.CNT_TRI_OBS] + 150.0 * dc[capi.CNT_TRIANGULATIONS]
                        + (dc[capi.CNT_TRI_SOLVES] - dc[capi.CNT_TRIANGULATIONS]) * 40.0,
                        36.0 * dc[capi.CNT_TRI_OBS] + 25.0 * dc[capi.CNT_TRIANGULATIONS], None),
        'jacobian_nullspace_gate': (dc[capi.CNT_K2_FLOPS], dc[capi.CNT_K2_BYTES], None),
        'stacked_qr': (dc[capi.CNT_QR_FLOPS], dc[capi.CNT_QR_BYTES], None),
        'gain_update': (dc[capi.CNT_K4_FLOPS], dc[capi.CNT_K4_BYTES], dc[capi.CNT_K4_EXEC_FLOPS]),
    }
    total_ms = sum(v[2] for v in kt.values())
    fams = {}
    for name, names in FAMILIES.items():
        flops, bts, exec_flops = work[name]
        ms = sum(kt[n][2] for n in names)
        nl = sum(kt[n][1] for n in names)
        peak = dmma if name in ('gain_update', 'stacked_qr') else dfma
        tf = flops / (ms * 1e-3) / 1e12 if ms else 0.0
        fams[name] = {'ms': ms, 'launches': nl, 'share': ms / total_ms if total_ms else 0.0, 'tflops': tf, 'gflops': tf * 1e3,
                      'frac': tf / peak if peak else None, 'peak_tflops': peak, 'peak_kind': 'DMMA' if peak == dmma else 'DFMA',
                      'gbs': bts / (ms * 1e-3) / 1e9 if ms else 0.0,
                      'flops_per_launch': flops / nl if nl else 0.0, 'bytes_per_launch': bts / nl if nl else 0.0}
        if exec_flops is not None:
            te = exec_flops / (ms * 1e-3) / 1e12 if ms else 0.0
            fams[name]['executed_tflops'] = te
            fams[name]['executed_frac'] = te / peak if peak else None
    other = dict((n, {'ms': v[2], 'launches': v[1], 'share': v[2] / total_ms if total_ms else 0.0}) for n, v in kt.items() if v[1])
    return fams, other, total_ms


# ------------------------------------------------------------------------------------------ B200 arm: single-filter latency
def single_filter_latency(device, feats, max_cam=30, n_frames=150, with_cpu=True):
    """configs[1]: one filter, max_cam_state_size=30, per-frame wall latency of MSCKF.feature_callback
    (reference-format message in, vio_result out); the CPU oracle on the same stream beside it."""
    from msckf_b200 import synth
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKF
    stream = synth.make_stream(n_frames, feats, structure_seed=0, noise_seed=0)
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    filt = MSCKF(cfg, device=device, feature_capacity=2048)
    msgs = stream.imu_msgs()

    def run(flt):
        k = 0
        lat = []
        for f in range(n_frames):
            while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
                flt.imu_callback(msgs[k])
                k += 1
            msg = stream.feature_msg(f)
            t0 = time.perf_counter()
            flt.feature_callback(msg)
            lat.append(time.perf_counter() - t0)
        return np.array(lat[45:]) * 1e3

    steady = run(filt)
    upd = int(filt.ctx.counters()[0, 0])
    filt.close()
    out = {'workload': 'configs[1]: 1 filter, max_cam_state_size=%d, %d features/frame, %d frames (first 45 excluded)' % (max_cam, feats, n_frames),
           'p50_ms': float(np.median(steady)), 'p99_ms': float(np.percentile(steady, 99)), 'mean_ms': float(steady.mean()),
           'feature_updates': upd}
    if with_cpu:
        from oracle import msckf_oracle as orc
        ocfg = orc.ConfigEuRoC()
        ocfg._vio_max_cam_state_size__ = max_cam
        cpu = run(orc.MSCKF(ocfg))
        out['cpu_oracle'] = {'p50_ms': float(np.median(cpu)), 'p99_ms': float(np.percentile(cpu, 99)), 'mean_ms': float(cpu.mean()),
                             'cores': 1, 'kind': 'port', 'same_stream': True}
    return out


# ------------------------------------------------------------------------------------------ B200 arm: configs[2] stress frame
def c3_leg(device, with_cpu, peaks):
    """BASELINE configs[2]: one filter, 30 camera states, 2000 lost landmarks (M ~ U{3..29}), row cap lifted: one
    remove_lost_features call = batched triangulation + Jacobians + null-space projection + gate + compression of the
    ~1.2e5 x 201 stacked Jacobian + update.  The state is restored between repetitions (untimed)."""
    from msckf_b200 import capi, synth
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKF
    max_cam, n_lead, n_lm = 30, 29, 2000
    stream = synth.make_stream(n_lead + 1, 40, structure_seed=8, noise_seed=3)
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    cfg.row_cap = 0
    filt = MSCKF(cfg, device=device, feature_capacity=2304)
    ctx = filt.ctx
    msgs = stream.imu_msgs()
    k = 0
    for f in range(n_lead):
        while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
            filt.imu_callback(msgs[k])
            k += 1
        filt.feature_callback(stream.feature_msg(f))
    t = stream.frame_ts[n_lead]
    while k < len(msgs) and msgs[k].vio_timestamp__ <= t:
        filt.imu_callback(msgs[k])
        k += 1
    filt.batch_imu_processing(t)
    filt.state_augmentation(t)
    cam_ids, cams = ctx.get_cam_states(0)
    T01 = np.asarray(cfg._vio_T_cn_cnm1__, dtype=np.float64)
    ids, masks, obs = synth.stress_tracks(cams[:, 0:4], cams[:, 4:7], T01[:3, :3], T01[:3, 3], n_lm, np.random.default_rng(42),
                                          outlier_frac=0.005)
    imu0, P0 = ctx.get_imu_state(0), ctx.get_cov(0)

    def restore():
        ctx.set_imu_state(0, imu0)
        ctx.set_cam_states(0, cam_ids, cams)
        ctx.set_cov(0, P0)
        ctx.set_features(0, ids, masks, None, None, obs)

    times = []
    for rep in range(6):
        restore()
        ctx.sync()
        ctx.mark(0)
        ctx.remove_lost_features()
        ctx.mark(1)
        times.append(ctx.mark_elapsed_ms())
        ctx.sync()
    m, na, kr, path = [int(x) for x in ctx.update_shapes()[0]]
    log = ctx.stage_log(0, 0)
    gates = int(np.sum((log['status'] == capi.FEAT_GATE_PASS) | (log['status'] == capi.FEAT_GATE_FAIL)))
    # per-kernel pass
    restore()
    c0 = ctx.counters().sum(axis=0)
    ctx.set_kernel_timing(True)
    kt0 = ctx.kernel_times()
    ctx.remove_lost_features()
    ctx.sync()
    kt1 = ctx.kernel_times()
    ctx.set_kernel_timing(False)
    dc = (ctx.counters().sum(axis=0) - c0).astype(np.float64)
    fams, other, total_ms = family_stats(kt0, kt1, dc, *peaks)
    filt.close()
    steady = np.array(times[1:])
    out = {'workload': 'configs[2]: 1 filter, 30 camera states (n=201), %d lost landmarks with 3..29 stereo observations, row cap lifted; '
                       'step = one remove_lost_features call' % n_lm,
           'stacked_rows': m, 'active_cols': na, 'state_dim': 201, 'feature_updates': gates,
           'ms_per_frame': float(np.median(steady)), 'ms_min': float(steady.min()), 'ms_max': float(steady.max()), 'reps': len(steady),
           'feature_updates_per_s': gates / (float(np.median(steady)) * 1e-3),
           'kernel_ms_in_order': total_ms, 'families': fams,
           'kernels': dict((n, round(v['ms'], 4)) for n, v in sorted(other.items(), key=lambda x: -x[1]['ms'])[:12])}
    if with_cpu:
        from oracle import msckf_oracle as orc
        ocfg = orc.ConfigEuRoC()
        ocfg._vio_max_cam_state_size__ = max_cam
        of = orc.MSCKF(ocfg)
        of.ROW_CAP = 10 ** 9
        k = 0
        for f in range(n_lead):
            while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
                of.imu_callback(msgs[k])
                k += 1
            of.feature_callback(stream.feature_msg(f))
        while k < len(msgs) and msgs[k].vio_timestamp__ <= t:
            of.imu_callback(msgs[k])
            k += 1
        of.batch_imu_processing(t)
        of.state_augmentation(t)
        ocam = [int(c) for c in of.state_server._vio_cam_states__.keys()]
        of.map_server.clear()
        for j in range(n_lm):
            ft = orc.Feature(int(ids[j]), of.optimization_config, of.R_cam0_cam1, of.t_cam0_cam1)
            for c in range(30):
                if (int(masks[j]) >> c) & 1:
                    ft.observations[ocam[c]] = obs[j, c].copy()
            of.map_server[int(ids[j])] = ft
        t0 = time.perf_counter()
        of.remove_lost_features()
        sec = time.perf_counter() - t0
        out['cpu_oracle'] = {'ms_per_frame': sec * 1e3, 'cores': 1, 'kind': 'port', 'same_frame': True}
        out['speedup_vs_cpu_oracle'] = sec * 1e3 / out['ms_per_frame']
    return out


# ------------------------------------------------------------------------------------------ B200 arm: configs[4] large window
def c5_leg(device, F, with_cpu, peaks, torch):
    """BASELINE configs[4]: max_cam_state_size = 64 (n = 405), 4096 features per frame, a batch of F filters (noise
    seeds); fast platform (synth speed 10) so that every frame carries a capped 1500-row x 405 update on the full
    covariance -- the covariance-update-dominated FP64 tensor-pipe path."""
    from msckf_b200 import capi, synth
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKFBatch
    n_lead, n_warm, n_timed, n_prof = 64, 4, 20, 10
    n_frames = n_lead + n_warm + n_timed + n_prof
    target = np.array([300] * n_lead + [4096] * (n_frames - n_lead))
    st = synth.make_structure(n_frames, target, structure_seed=12, n_landmarks=150000, speed=10.0)
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = 64
    batch = MSCKFBatch(cfg, F, device=device, feature_capacity=8192)
    ctx = batch.ctx
    ds = DeviceStream(st, batch, F, 0, device, torch)
    g0, a0 = ds.static_samples()
    batch.initialize_gravity_and_bias(g0, a0)
    frames = [ds.frame(f) for f in range(n_frames)]

    def run(f):
        d = frames[f]
        ctx.feature_callback_dev(*[d[k].data_ptr() for k in KEYS])

    for f in range(n_lead + n_warm):
        run(f)
    ctx.sync()
    c0 = ctx.counters().sum(axis=0)
    ctx.mark(0)
    for f in range(n_lead + n_warm, n_lead + n_warm + n_timed):
        run(f)
    ctx.mark(1)
    ms = ctx.mark_elapsed_ms()
    ctx.sync()
    c1 = ctx.counters().sum(axis=0)
    ctx.set_kernel_timing(True)
    kt0 = ctx.kernel_times()
    for f in range(n_lead + n_warm + n_timed, n_frames):
        run(f)
    ctx.sync()
    kt1 = ctx.kernel_times()
    ctx.set_kernel_timing(False)
    dc = (ctx.counters().sum(axis=0) - c1).astype(np.float64)
    fams, other, total_ms = family_stats(kt0, kt1, dc, *peaks)
    batch.close()
    upd = float(c1[capi.CNT_FEATURE_UPDATES] - c0[capi.CNT_FEATURE_UPDATES])
    out = {'workload': 'configs[4]: %d filters, max_cam_state_size=64 (n=405), 4096 features/frame, fast platform (capped 1500-row update on '
                       'the full window every frame); step = one frame over all filters' % F,
           'filters': F, 'frames_timed': n_timed, 'ms_per_frame': ms / n_timed, 'ms_per_filter_frame': ms / n_timed / F,
           'feature_updates_per_s': upd / (ms * 1e-3), 'filter_frames_per_s': F * n_timed / (ms * 1e-3),
           'kernel_ms_in_order_per_frame': total_ms / n_prof, 'families': fams,
           'kernels': dict((n, round(v['ms'] / n_prof, 4)) for n, v in sorted(other.items(), key=lambda x: -x[1]['ms'])[:14]),
           }
    if with_cpu:
        from oracle import msckf_oracle as orc
        stream = synth.realize(st, noise_seed=0)
        ocfg = orc.ConfigEuRoC()
        ocfg._vio_max_cam_state_size__ = 64
        of = orc.MSCKF(ocfg)
        msgs = stream.imu_msgs()
        k = 0
        per = []
        n_cpu = n_lead + n_warm + 6
        for f in range(n_cpu):
            while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
                of.imu_callback(msgs[k])
                k += 1
            msg = stream.feature_msg(f)
            t0 = time.perf_counter()
            of.feature_callback(msg)
            per.append(time.perf_counter() - t0)
        cpu_ms = float(np.mean(per[n_lead + n_warm:]) * 1e3)
        out['cpu_oracle'] = {'ms_per_filter_frame': cpu_ms, 'cores': 1, 'kind': 'port', 'frames': 6,
                             'sample': '1 filter, frames %d..%d of the same structure (numpy noise seed 0)' % (n_lead + n_warm, n_cpu - 1)}
        out['speedup_vs_cpu_oracle_per_filter_frame'] = cpu_ms / out['ms_per_filter_frame']
    return out


# ------------------------------------------------------------------------------------------ B200 arm: configs[3]
def measure_batch(args, torch, dist, world, rank, local, F, seed0, K, W, R, with_profile, with_spot, label):
    """One context of F filters with noise seeds seed0 .. seed0+F-1: lead-in, device-timed blocks, e2e blocks and
    (optionally) the per-kernel pass and the parity spot check.  Collective calls inside are matched on every rank."""
    from msckf_b200 import capi, synth
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKFBatch
    dev = torch.device('cuda', local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def reduce(x, op):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=op)
        return float(t.item())

    rmax = lambda x: reduce(x, dist.ReduceOp.MAX if world > 1 else None)   # noqa: E731
    rsum = lambda x: reduce(x, dist.ReduceOp.SUM if world > 1 else None)   # noqa: E731

    W2 = 3
    est_ms = 1.3 * F / 1024.0 + 0.25                      # rough frame time, only used to size the clock warm-up
    n_clock = int(max(20, min(400, args.clock_warm_seconds * 1e3 / est_ms)))
    n_lead = WINDOW_FILL + n_clock
    n_prof = PROFILE_FRAMES if with_profile else 0
    n_frames = n_lead + W + K * R + W2 + K * R + n_prof
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = args.max_cam
    st = get_structure(n_frames, args.feats)
    batch = MSCKFBatch(cfg, F, device=local, feature_capacity=1024, groups=(args.groups or None))
    ctx = batch.ctx
    ds = DeviceStream(st, batch, F, seed0, local, torch)
    g0, a0 = ds.static_samples()
    batch.initialize_gravity_and_bias(g0, a0)

    first_dev = n_lead + W
    first_e2e = first_dev + K * R
    first_prof = first_e2e + W2 + K * R
    rng = np.random.default_rng(12345 + seed0)
    spot = sorted(set([0, F - 1] + [int(x) for x in rng.integers(0, F, 6)])) if with_spot else []
    spot_frames = dict((s, []) for s in spot)
    dev_frames, host_frames = {}, {}
    for f in range(n_frames):
        d = ds.frame(f)
        if first_e2e <= f < first_prof:
            host_frames[f] = to_pinned(d, torch)      # e2e frames live in page-locked HOST memory
        else:
            dev_frames[f] = (d, batch.device_frame_parts(d, KEYS))   # everything else stays in HBM
        if spot and f < first_e2e:
            ts, ids = ds.plan[f]
            gy, ac, uv = d['gyro'][spot].cpu().numpy(), d['acc'][spot].cpu().numpy(), d['uv'][spot].cpu().numpy()
            for j, s in enumerate(spot):
                spot_frames[s].append((float(st.frame_ts[f]), ts, gy[j], ac[j], ids, uv[j]))
    traj = torch.zeros((K * R, F, capi.IMU_STRIDE), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def run_dev(f, out=None):
        batch.feature_callback_dev(dev_frames[f][1], out)

    # ---- lead-in: fills the camera window, then keeps running the job's own frames until the clocks have settled
    sampler = ClockSampler(local)
    sampler.start()
    for f in range(n_lead + W):
        run_dev(f)
    ctx.sync()

    # ---- device-resident blocks
    c0 = ctx.counters().sum(axis=0)
    l0 = ctx.kernel_launches()
    blocks = []
    for b in range(R):
        cb0 = ctx.counters()[:, capi.CNT_FEATURE_UPDATES].sum()
        barrier()
        ctx.mark(0)
        for i in range(K):
            run_dev(first_dev + b * K + i, traj[b * K + i])
        ctx.mark(1)
        ms = ctx.mark_elapsed_ms()
        ctx.sync()
        torch.cuda.synchronize()
        barrier()
        upd = float(ctx.counters()[:, capi.CNT_FEATURE_UPDATES].sum() - cb0)
        blocks.append((rmax(ms), rsum(upd)))
    l1 = ctx.kernel_launches()
    c1 = ctx.counters().sum(axis=0)

    # ---- parity spot check on the state after the device-timed frames (outside every timed region)
    spot_state = {}
    for s in spot:
        cid, _ = ctx.get_cam_states(s)
        spot_state[s] = {'imu': ctx.get_imu_state(s), 'P': ctx.get_cov(s), 'cam_ids': cid,
                         'map_ids': ctx.get_features(s, with_obs=False)[0]}

    # ---- e2e blocks: host buffers through the C ABI, H2D + kernels + D2H of every frame, two frames in flight
    def e2e_run(lo, n):
        last = None
        resets = 0
        ctx.feature_callback_async(*[host_frames[lo][k] for k in KEYS])
        for i in range(1, n):
            ctx.feature_callback_async(*[host_frames[lo + i][k] for k in KEYS])
            last = ctx.collect()
            resets += int(last[1].sum())
        last = ctx.collect()
        resets += int(last[1].sum())
        return last, resets

    e2e_run(first_e2e, W2)
    e2e_blocks = []
    resets = 0
    for b in range(R):
        cb0 = ctx.counters()[:, capi.CNT_FEATURE_UPDATES].sum()
        barrier()
        t0 = time.perf_counter()
        (imu_last, _), rs = e2e_run(first_e2e + W2 + b * K, K)
        sec = time.perf_counter() - t0
        barrier()
        assert np.isfinite(imu_last).all()
        resets += rs
        upd = float(ctx.counters()[:, capi.CNT_FEATURE_UPDATES].sum() - cb0)
        e2e_blocks.append((rmax(sec), rsum(upd)))
    clocks = sampler.stop()
    fr = host_frames[first_e2e + W2]
    h2d = int(sum(fr[k].nbytes for k in KEYS))
    d2h = int(F * capi.IMU_STRIDE * 8 + F + 32)

    # ---- per-kernel event pass
    prof = None
    if with_profile:
        ctx.set_kernel_timing(True)
        kt0 = ctx.kernel_times()
        c4 = ctx.counters().sum(axis=0)
        for f in range(first_prof, n_frames):
            if batch.groups == 1:
                run_dev(f)
            else:                                     # per-kernel event times want the GPU to themselves: one context at a time
                for g, part in enumerate(ctx.parts):
                    part.feature_callback_dev(*dev_frames[f][1][g])
                    part.sync()
        ctx.sync()
        kt1 = ctx.kernel_times()
        ctx.set_kernel_timing(False)
        prof = (kt0, kt1, (ctx.counters().sum(axis=0) - c4).astype(np.float64))
    cnt_final = ctx.counters().sum(axis=0)

    # ---- final gather of the trajectories (the only collective of the job) + ATE against the known trajectory
    pos = traj[:, :, capi.IMU_P:capi.IMU_P + 3]
    truth = torch.from_numpy(st.truth_p[first_dev:first_dev + K * R]).to(dev)
    err2 = ((pos - truth[:, None, :]) ** 2).sum(dim=2)                       # [frames][F]
    assert bool(torch.isfinite(err2).all())
    if world > 1:
        gathered = [torch.zeros_like(err2) for _ in range(world)] if rank == 0 else None
        dist.gather(err2, gathered, dst=0)
        err2_all = torch.cat(gathered, dim=1) if rank == 0 else err2
    else:
        err2_all = err2
    ate = {'rmse_m': float(torch.sqrt(err2_all.mean()).item()), 'worst_filter_rmse_m': float(torch.sqrt(err2_all.mean(dim=0).max()).item()),
           'final_frame_mean_m': float(torch.sqrt(err2_all[-1]).mean().item()), 'filters': int(err2_all.shape[1]),
           'frames': int(err2_all.shape[0]), 'sim_seconds_at_end': float(st.frame_ts[first_dev + K * R - 1] - synth.T_START),
           'truth': 'synth.truth_position at the frame timestamps (world frame = IMU pose at the first image)'}
    batch.close()

    vals = np.array([u / (t * 1e-3) for t, u in blocks])
    tms = np.array([t for t, _ in blocks])
    evals = np.array([u / t for t, u in e2e_blocks])
    ets = np.array([t for t, _ in e2e_blocks])
    med = int(np.argsort(tms)[len(tms) // 2])
    emed = int(np.argsort(ets)[len(ets) // 2])
    Ftot = rsum(float(F))
    res = {'label': label, 'filters_local': F, 'filters_total': int(Ftot), 'seed0': seed0, 'contexts_per_gpu': batch.groups,
           'value': float(vals[med]), 'ms_per_step': float(tms[med] / K),
           'blocks': {'n': R, 'steps_per_block': K, 'timed_seconds': float(tms.sum() * 1e-3), 'ms_per_step_min': float(tms.min() / K),
                      'ms_per_step_max': float(tms.max() / K), 'value_min': float(vals.min()), 'value_max': float(vals.max()),
                      'value_mean': float(sum(u for _, u in blocks) / (tms.sum() * 1e-3))},
           'feature_updates_per_step': float(blocks[med][1] / K), 'filter_frames_per_s': float(Ftot * K / (tms[med] * 1e-3)),
           'e2e': {'value': float(evals[emed]), 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                   'ms_per_step': float(ets[emed] / K * 1e3), 'filter_frames_per_s': float(Ftot * K / ets[emed]),
                   'blocks': {'n': R, 'timed_seconds': float(ets.sum()), 'value_min': float(evals.min()), 'value_max': float(evals.max())}},
           'gpu_launches': int((l1 - l0) // R), 'gpu_launches_all_blocks': int(l1 - l0), 'clocks': clocks, 'ate': ate,
           'online_resets': resets, 'gate_not_spd': int(cnt_final[capi.CNT_GATE_NOT_SPD]),
           'lead_in_frames': n_lead, 'prof': prof, 'spot': (spot, spot_state, spot_frames, (st.imu_ts[:200], g0, a0))}
    return res


def parity_spot(args, res):
    """Oracle vs device on the spot filters (rank 0): worst relative error of the state vector and of P, id equality."""
    import multiprocessing as mp
    spot, state, frames, (its, g0, a0) = res['spot']
    if not spot:
        return None
    t0 = time.perf_counter()
    jobs = [(args.max_cam, its, g0[s], a0[s], frames[s]) for s in spot]
    with mp.get_context('spawn').Pool(min(len(jobs), max(1, (os.cpu_count() or 2) - 1))) as pool:
        out = pool.map(spot_oracle_worker, jobs)
    worst, ids_equal = 0.0, True
    from msckf_b200 import capi
    for s, o in zip(spot, out):
        b = state[s]['imu']
        dimu = np.concatenate([b[capi.IMU_Q:capi.IMU_Q + 4], b[capi.IMU_BG:capi.IMU_BG + 3], b[capi.IMU_V:capi.IMU_V + 3],
                               b[capi.IMU_BA:capi.IMU_BA + 3], b[capi.IMU_P:capi.IMU_P + 3]])
        ids_equal = ids_equal and np.array_equal(state[s]['cam_ids'], o['cam_ids']) and np.array_equal(state[s]['map_ids'], o['map_ids'])
        if state[s]['P'].shape == o['P'].shape:
            worst = max(worst, float(np.linalg.norm(dimu - o['imu']) / np.linalg.norm(o['imu'])),
                        float(np.linalg.norm(state[s]['P'] - o['P']) / np.linalg.norm(o['P'])))
        else:
            worst, ids_equal = float('inf'), False
    return {'filters': [int(s + res['seed0']) for s in spot], 'frames': len(frames[spot[0]]), 'worst_rel': worst, 'ids_equal': bool(ids_equal),
            'tolerance': 1e-9, 'ok': bool(ids_equal and worst < 1e-9), 'oracle_seconds': time.perf_counter() - t0,
            'yardstick': 'the oracle against itself with every measurement moved by one ulp: 4.7e-13 median, 2.9e-9 worst over 96 seeds x 100 frames '
                         '(tools/oracle_sensitivity.py)'}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from msckf_b200 import capi
    from msckf_b200.partition import filter_range

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device -- the B200 arm has no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    if args.only:
        peaks = capi.fp64_peak(local)
        out = {'peaks_tflops': {'dfma': peaks[0], 'dmma': peaks[1]}}
        if 'c3' in args.only:
            out['c3'] = c3_leg(local, False, peaks)
        if 'c5' in args.only:
            out['c5'] = c5_leg(local, args.c5_filters, False, peaks, torch)
        print(json.dumps(out))
        return
    K, W = args.steps, max(args.warmup, 3)
    F = args.filters_per_gpu
    est = 1.3 * F / 1024.0 + 0.1
    R = int(max(3, min(25, np.ceil(args.timed_seconds * 1e3 / (K * est)))))
    res = measure_batch(args, torch, dist, world, rank, local, F, rank * F, K, W, R, True, not args.no_parity_spot and rank == 0, 'weak')
    strong = None
    if world > 1 and not args.no_strong:
        lo, hi = filter_range(rank, world, TOTAL_FILTERS)
        est_s = 1.3 * (hi - lo) / 1024.0 + 0.25
        Rs = int(max(3, min(25, np.ceil(args.timed_seconds * 1e3 / (K * est_s)))))
        strong = measure_batch(args, torch, dist, world, rank, local, hi - lo, lo, K, W, Rs, False, False, 'strong')

    if rank == 0:
        dfma, dmma = capi.fp64_peak(local)
        kt0, kt1, dc = res.pop('prof')
        fams, other, total_ms = family_stats(kt0, kt1, dc, dfma, dmma)
        dom = max(fams, key=lambda k: fams[k]['ms'])
        dk = fams[dom]
        traffic = None
        for name in ('traffic_r2.json', 'traffic_r1.json'):
            tpath = os.path.join(REPO, 'profiles', name)
            if os.path.exists(tpath):   # dram bytes per launch of each family, from the committed ncu --set full capture of this workload
                traffic = json.load(open(tpath)).get('dram_bytes_per_launch', {}).get(dom)
                break
        roofline = {'kernel': dom, 'bound': 'tensor', 'achieved': dk['tflops'], 'peak': dk['peak_tflops'], 'unit': 'TFLOP/s',
                    'frac': dk['frac'], 'traffic': traffic,
                    'peak_source': 'FP64 peaks measured on this GPU by msckf_fp64_peak (MEASURED_PEAKS.json carries no FP64 figure): '
                                   'DFMA %.1f / DMMA %.1f TFLOP/s; QR and update families are rated against DMMA, the others against DFMA'
                                   % (dfma, dmma),
                    'avg_launch_ms': dk['ms'] / dk['launches'] if dk['launches'] else None,
                    'algorithmic_flops_per_launch': dk['flops_per_launch'], 'algorithmic_bytes_per_launch': dk['bytes_per_launch'],
                    'hbm_gbs_achieved': dk['gbs'], 'families': fams, 'all_kernels': other, 'profile_frames': PROFILE_FRAMES}
        spot = res.pop('spot')
        line = {'metric': METRIC, 'value': res['value'], 'unit': UNIT, 'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': res['ms_per_step'],
                'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic (generated on the GPU)',
                'config': dict(workload_config(args, world), contexts_per_gpu=res['contexts_per_gpu']), 'blocks': res['blocks'],
                'filter_frames_per_s': res['filter_frames_per_s'], 'feature_updates_per_step': res['feature_updates_per_step'],
                'e2e': res['e2e'], 'gpu_launches': res['gpu_launches'], 'clocks': res['clocks'], 'roofline': roofline, 'ate': res['ate'],
                'online_resets': res['online_resets'], 'gate_not_spd': res['gate_not_spd'],
                'clock_warmup': '%d frames of the job itself before anything is timed (no library GEMM spin)' % res['lead_in_frames']}
        if strong is not None:
            strong.pop('prof')
            strong.pop('spot')
            line['strong'] = {'filters_total': strong['filters_total'], 'filters_per_gpu': strong['filters_local'], 'value': strong['value'],
                              'ms_per_step': strong['ms_per_step'], 'e2e': strong['e2e'], 'blocks': strong['blocks'],
                              'filter_frames_per_s': strong['filter_frames_per_s'], 'ate': strong['ate'],
                              'contexts_per_gpu': strong['contexts_per_gpu'],
                              'config': workload_config(args, world, TOTAL_FILTERS)['workload']}
        elif world == 1 and F == TOTAL_FILTERS:
            line['strong'] = {'filters_total': TOTAL_FILTERS, 'filters_per_gpu': F, 'value': res['value'], 'ms_per_step': res['ms_per_step'],
                              'e2e': res['e2e'], 'note': 'at one GPU the strong and the weak configuration are the same run'}
        res['spot'] = spot
        if not args.no_parity_spot:
            line['parity_spot'] = parity_spot(args, res)
        if not args.no_latency:
            line['latency'] = single_filter_latency(local, args.feats, with_cpu=not args.no_cpu_baseline)
        if not args.no_cpu_baseline:
            line['cpu_baseline'] = cpu_baseline_leg(args)
        if world == 1 and not args.no_c3:
            line['c3'] = c3_leg(local, not args.no_cpu_baseline, (dfma, dmma))
        if world == 1 and not args.no_c5:
            line['c5'] = c5_leg(local, args.c5_filters, not args.no_cpu_baseline, (dfma, dmma), torch)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_b200(a)
