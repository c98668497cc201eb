#!/usr/bin/env python
"""bench.py -- MSCKF back-end hot path on B200: feature-updates/s (BASELINE.json metric).

A "step" is one frame (feature_callback: propagate, augment, observe, remove_lost_features,
prune_cam_state_buffer, online_reset) over EVERY filter of the job.  Workload = BASELINE.json
configs[3] shape: independent filters (Monte-Carlo noise realisations of one synthetic
EuRoC-MH_01-shaped stream: 200 Hz IMU, 20 Hz stereo, ~300 features/frame,
max_cam_state_size = 20), a fixed number of filters PER GPU (weak scaling), no collective on the
data path, one final gather of the trajectories.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

  value   feature-updates/s with every input of the timed frames already in HBM (device pointers
          through msckf_feature_callback_dev), K frames enqueued back to back, CUDA events on the
          context's stream, max over ranks.
  e2e     the same metric through the C ABI with HOST buffers (msckf_feature_callback_async / msckf_collect):
          H2D of every frame's message arrays from page-locked host memory, kernels, D2H of the publish()
          snapshot of every frame; two frames in flight.
  roofline  dominant kernel family of a per-kernel CUDA-event pass over further frames of the same
          run: algorithmic FLOPs (device work counters, formulas in DESIGN.md) / its device time,
          against the FP64 peak measured on this GPU (msckf_fp64_peak; MEASURED_PEAKS.json has no FP64).
  cpu_baseline  the CPU oracle (plain-numpy port of the reference, oracle/msckf_oracle.py) on one
          core of this box over a bounded sample of the same stream.
  latency single-filter (configs[1]: max_cam_state_size = 30) p50 / p99 of feature_callback.

--impl reference: the reference algorithm's CPU port (the reference itself is Python and is not on
the GPU box) on all host cores, one filter per process, same stream, same metric.
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
LEAD_IN = 80          # untimed frames that fill the 20-state window so the timed frames are steady state
PROFILE_FRAMES = 20   # frames of the per-kernel event pass


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=100)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--filters-per-gpu', type=int, default=1024)
    ap.add_argument('--feats', type=int, default=300)
    ap.add_argument('--max-cam', type=int, default=20)
    ap.add_argument('--no-latency', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    return ap.parse_args()


def workload_config(args, n_gpus):
    return {'workload': 'configs[3]: %d independent filters per GPU (Monte-Carlo noise seeds of one synthetic '
                        'EuRoC-MH_01-shaped stream), %d features/frame, max_cam_state_size=%d, 10 IMU samples/frame; '
                        'step = one frame over all filters' % (args.filters_per_gpu, args.feats, args.max_cam),
            'filters_per_gpu': args.filters_per_gpu, 'filters_total': args.filters_per_gpu * n_gpus,
            'features_per_frame': args.feats, 'max_cam_state_size': args.max_cam, 'lead_in_frames': LEAD_IN,
            'l2': 'working set per frame (P + map + Jacobians of all filters) exceeds nothing by design: state is '
                  'resident and re-read every frame as in production; inputs of successive frames are distinct buffers',
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


# ------------------------------------------------------------------------------------------ data
def build_frames(st, rng, F, n_frames, sigma_z=0.001, outlier_frac=0.05, outlier_sigma=0.05):
    """Noisy realisations for F filters sharing the structure `st` (ids, timestamps): per-frame host arrays."""
    from msckf_b200 import synth
    from msckf_b200.config import ConfigEuRoC
    cfg = ConfigEuRoC()
    sg = np.sqrt(cfg._vio_gyro_noise__ / synth.IMU_DT)
    sa = np.sqrt(cfg._vio_acc_noise__ / synth.IMU_DT)
    K = len(st.imu_ts)
    gyro = st.gyro_clean[None] + st.gyro_bias + sg * rng.standard_normal((F, K, 3))
    acc = st.acc_clean[None] + st.acc_bias + sa * rng.standard_normal((F, K, 3))
    uvs = []
    for f in range(n_frames):
        uv = st.frame_uv[f]
        n = sigma_z * rng.standard_normal((F,) + uv.shape)
        out = rng.random((F, len(uv))) < outlier_frac
        n[out] += outlier_sigma * rng.standard_normal((int(out.sum()), 4))
        uvs.append(uv[None] + n)
    return gyro, acc, uvs


def frame_arrays(batch, st, gyro, acc, uvs, f):
    """The C-ABI argument arrays of frame f (host, contiguous)."""
    from msckf_b200 import synth  # noqa: F401
    F = batch.F
    b = int(np.searchsorted(st.imu_ts, st.frame_ts[f], side='right'))
    a = 0 if f == 0 else int(np.searchsorted(st.imu_ts, st.frame_ts[f - 1], side='right'))
    take = batch.select_imu(float(st.frame_ts[f]), st.imu_ts[a:b])
    ts = st.imu_ts[a:b][take]
    K, n = len(ts), len(st.frame_ids[f])
    cam_id = batch.next_imu_id
    batch.next_imu_id += 1
    return {'imu_off': (np.arange(F + 1) * K).astype(np.int32), 'ts': np.ascontiguousarray(np.tile(ts, F)),
            'gyro': np.ascontiguousarray(gyro[:, a:b][:, take]), 'acc': np.ascontiguousarray(acc[:, a:b][:, take]),
            'cam_ids': np.full(F, cam_id, dtype=np.int64), 'feat_off': (np.arange(F + 1) * n).astype(np.int32),
            'ids': np.ascontiguousarray(np.tile(st.frame_ids[f].astype(np.int64), F)), 'uv': np.ascontiguousarray(uvs[f])}


KEYS = ['imu_off', 'ts', 'gyro', 'acc', 'cam_ids', 'feat_off', 'ids', 'uv']


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
    """Bounded sample on one core: one oracle filter, LEAD_IN + 3 + 100 frames of the same stream shape."""
    n_timed = 100
    t0 = time.perf_counter()
    upd, sec, per = oracle_filter_run(0, LEAD_IN, 3, n_timed, args.feats, args.max_cam, LEAD_IN + 3 + n_timed)
    return {'value': upd / sec, 'unit': UNIT, 'cores': 1, 'kind': 'port',
            'sample': '1 filter x %d frames after %d lead-in frames of the same stream (noise seed 0), oracle/msckf_oracle.py, '
                      'OPENBLAS_NUM_THREADS=1; %d feature-updates in %.1f s (whole leg %.1f s)' % (n_timed, LEAD_IN + 3, upd, sec,
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
    total = LEAD_IN + n_warm + n_timed
    jobs = [(k, LEAD_IN, n_warm, n_timed, args.feats, args.max_cam, total) for k in range(workers)]
    t0 = time.perf_counter()
    with mp.get_context('fork').Pool(workers) as pool:
        res = pool.map(_oracle_worker, jobs)
    wall = time.perf_counter() - t0
    upd = sum(r[0] for r in res)
    tmax = max(r[1] for r in res)
    value = upd / tmax
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': n_timed,
            'warmup': n_warm, 'ms_per_step': tmax / n_timed * 1e3, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f64', 'data': 'synthetic', 'config': workload_config(args, args.gpus),
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': workers, 'kind': 'port',
                             'sample': '%d oracle filters (one per process, host has %d cores), %d timed frames each after %d lead-in '
                                       'frames; whole run %.1f s' % (workers, cores, n_timed, LEAD_IN + n_warm, wall)},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'filter_frames_per_s': workers * n_timed / tmax, 'gpu_launches': 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------ B200 arm
def single_filter_latency(device, feats, max_cam=30, n_frames=150):
    """configs[1]: one filter, max_cam_state_size=30, per-frame wall latency of MSCKF.feature_callback
    (reference-format message in, vio_result out)."""
    from msckf_b200 import synth
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKF
    stream = synth.make_stream(n_frames, feats, structure_seed=0, noise_seed=0)
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    filt = MSCKF(cfg, device=device, feature_capacity=2048)
    msgs = stream.imu_msgs()
    k = 0
    lat = []
    for f in range(n_frames):
        while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
            filt.imu_callback(msgs[k])
            k += 1
        msg = stream.feature_msg(f)
        t0 = time.perf_counter()
        filt.feature_callback(msg)
        lat.append(time.perf_counter() - t0)
    upd = int(filt.ctx.counters()[0, 0])
    filt.close()
    steady = np.array(lat[45:]) * 1e3
    return {'workload': 'configs[1]: 1 filter, max_cam_state_size=%d, %d features/frame, %d frames (first 45 excluded)' % (max_cam, feats, n_frames),
            'p50_ms': float(np.median(steady)), 'p99_ms': float(np.percentile(steady, 99)), 'mean_ms': float(steady.mean()),
            'feature_updates': upd}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from msckf_b200 import capi, synth
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKFBatch

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device -- the B200 arm has no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def reduce_max(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    F, K, W = args.filters_per_gpu, args.steps, max(args.warmup, 3)
    W2 = 3
    n_frames = LEAD_IN + W + K + W2 + K + PROFILE_FRAMES
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = args.max_cam
    st = synth.make_structure(n_frames, args.feats, structure_seed=0)
    rng = np.random.default_rng(1000 + rank)
    gyro, acc, uvs = build_frames(st, rng, F, n_frames)
    batch = MSCKFBatch(cfg, F, device=local, feature_capacity=1024)
    ctx = batch.ctx
    batch.initialize_gravity_and_bias(gyro[:, :200], acc[:, :200])
    frames = [frame_arrays(batch, st, gyro, acc, uvs, f) for f in range(n_frames)]

    # ---- every input of the device-resident frames goes to HBM before anything is timed
    first_e2e = LEAD_IN + W + K
    first_prof = first_e2e + W2 + K
    dev_frames = {}
    for f in list(range(0, first_e2e)) + list(range(first_prof, n_frames)):
        dev_frames[f] = dict((k, torch.from_numpy(frames[f][k]).to(dev)) for k in KEYS)
    traj = torch.zeros((K, F, capi.IMU_STRIDE), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def run_dev(f, out_ptr=0):
        d = dev_frames[f]
        ctx.feature_callback_dev(*[d[k].data_ptr() for k in KEYS], imu_out=out_ptr)

    for f in range(LEAD_IN):                           # lead-in (fills the camera window), untimed
        run_dev(f)
    ctx.sync()
    # bring the SM clocks up from idle before anything is timed (the first ~0.3 s after idle run at reduced clocks)
    sampler = ClockSampler(local)      # samples every 100 ms from the clock warm-up to the end of the e2e leg: all of it is under load
    sampler.start()
    wa = torch.randn((4096, 4096), device=dev, dtype=torch.bfloat16)
    t_spin = time.perf_counter()
    while time.perf_counter() - t_spin < 1.0:
        for _ in range(20):
            wa @ wa
        torch.cuda.synchronize()
    for f in range(LEAD_IN, LEAD_IN + W):              # warm-up steps, untimed
        run_dev(f)
    ctx.sync()
    c0 = ctx.counters().sum(axis=0)
    l0 = ctx.kernel_launches()
    barrier()
    t0 = time.perf_counter()
    ctx.mark(0)
    for i in range(K):
        run_dev(LEAD_IN + W + i, traj[i].data_ptr())
    ctx.mark(1)
    dev_ms = ctx.mark_elapsed_ms()
    ctx.sync()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    c1 = ctx.counters().sum(axis=0)
    l1 = ctx.kernel_launches()
    upd_dev = float(c1[capi.CNT_FEATURE_UPDATES] - c0[capi.CNT_FEATURE_UPDATES])
    t_dev = reduce_max(dev_ms)
    total_upd = reduce_sum(upd_dev)
    value = total_upd / (t_dev * 1e-3)

    # ---- e2e: host buffers through the C ABI, sync + D2H of the result every frame
    # message buffers in page-locked host memory (msckf_host_alloc), as a feeder thread would fill them; every frame is
    # copied host->device inside the timed region and its publish() snapshot is read back (msckf_collect); up to two
    # frames are in flight, so the copy of frame k+1 overlaps the kernels of frame k
    for f in range(first_e2e, first_prof):
        frames[f] = dict((k, capi.pinned_copy(frames[f][k])) for k in KEYS)

    def e2e_run(lo, n):
        last = None
        ctx.feature_callback_async(*[frames[lo][k] for k in KEYS])
        for i in range(1, n):
            ctx.feature_callback_async(*[frames[lo + i][k] for k in KEYS])
            last = ctx.collect()
        last = ctx.collect()
        return last

    e2e_run(first_e2e, W2)
    c2 = ctx.counters().sum(axis=0)
    barrier()
    t0 = time.perf_counter()
    imu_last, _ = e2e_run(first_e2e + W2, K)
    e2e_s = time.perf_counter() - t0
    assert np.isfinite(imu_last).all()
    clocks = sampler.stop()
    barrier()
    c3 = ctx.counters().sum(axis=0)
    upd_e2e = reduce_sum(float(c3[capi.CNT_FEATURE_UPDATES] - c2[capi.CNT_FEATURE_UPDATES]))
    e2e_t = reduce_max(e2e_s)
    fr = frames[first_e2e + W2]
    h2d = int(sum(fr[k].nbytes for k in KEYS))
    d2h = int(F * capi.IMU_STRIDE * 8 + F + 4)

    # ---- per-kernel event pass (rank 0 reports it)
    ctx.set_kernel_timing(True)
    c4 = ctx.counters().sum(axis=0)
    for f in range(first_prof, n_frames):
        run_dev(f)
    ctx.sync()
    kt = ctx.kernel_times()
    ctx.set_kernel_timing(False)
    c5 = ctx.counters().sum(axis=0)
    dc = (c5 - c4).astype(np.float64)

    # ---- final gather of the trajectories (the only collective of the job)
    if world > 1:
        gathered = [torch.zeros_like(traj) for _ in range(world)] if rank == 0 else None
        dist.gather(traj, gathered, dst=0)
    torch.cuda.synchronize()
    pos_final = traj[-1, :, capi.IMU_P:capi.IMU_P + 3].cpu().numpy()
    assert np.isfinite(pos_final).all()

    if rank == 0:
        dfma, dmma = capi.fp64_peak(local)
        fam = {
            'triangulate': (['k_triangulate'], 632.0 * dc[capi.CNT_TRI_OBS] + 150.0 * dc[capi.CNT_TRIANGULATIONS]
                            + (dc[capi.CNT_TRI_SOLVES] - dc[capi.CNT_TRIANGULATIONS]) * 40.0, 36.0 * dc[capi.CNT_TRI_OBS] + 25.0 * dc[capi.CNT_TRIANGULATIONS]),
            'jacobian_nullspace_gate': (['k_jacobian_gate', 'k_jacobian_rows', 'k_jacobian_big'], dc[capi.CNT_K2_FLOPS], dc[capi.CNT_K2_BYTES]),
            'stacked_qr': (['k_qr', 'k_qr_tall', 'k_qr_d16', 'k_qr_d32', 'k_qr_d48', 'k_qr_d64', 'k_qr_d80'], dc[capi.CNT_QR_FLOPS], dc[capi.CNT_QR_BYTES]),
            'gain_update': (['k_upd_dmma16', 'k_upd_dmma32', 'k_upd_dmma48', 'k_upd_dmma64', 'k_upd_dmma80', 'k_upd_T', 'k_upd_S', 'k_upd_chol', 'k_upd_trsm', 'k_upd_inject', 'k_upd_downdate'], dc[capi.CNT_K4_FLOPS],
                            dc[capi.CNT_K4_BYTES]),
        }
        total_ms = sum(v[2] for v in kt.values())
        kernels = {}
        for name, (names, flops, bts) in fam.items():
            ms = sum(kt[n][2] for n in names)
            nl = sum(kt[n][1] for n in names)
            kernels[name] = {'ms': ms, 'launches': nl, 'share': ms / total_ms if total_ms else 0.0,
                             'gflops': flops / (ms * 1e-3) / 1e9 if ms else 0.0, 'gbs': bts / (ms * 1e-3) / 1e9 if ms else 0.0,
                             'flops_per_launch': flops / nl if nl else 0.0, 'bytes_per_launch': bts / nl if nl else 0.0}
        other = dict((n, {'ms': v[2], 'launches': v[1], 'share': v[2] / total_ms if total_ms else 0.0}) for n, v in kt.items() if v[1])
        dom = max(kernels, key=lambda k: kernels[k]['ms'])
        dk = kernels[dom]
        peak = dmma if dom == 'gain_update' else dfma
        traffic = None
        tpath = os.path.join(REPO, 'profiles', 'traffic_r1.json')
        if os.path.exists(tpath):       # dram bytes per launch of each family, from the committed ncu --set full capture of this workload
            traffic = json.load(open(tpath)).get('dram_bytes_per_launch', {}).get(dom)
        roofline = {'kernel': dom, 'bound': 'tensor', 'achieved': dk['gflops'] / 1e3, 'peak': peak, 'unit': 'TFLOP/s',
                    'frac': (dk['gflops'] / 1e3) / peak if peak else None, 'traffic': traffic,
                    'peak_source': 'FP64 %s peak measured on this GPU by msckf_fp64_peak (MEASURED_PEAKS.json carries no FP64 figure); '
                                   'DFMA %.1f / DMMA %.1f TFLOP/s' % ('DMMA' if dom == 'gain_update' else 'DFMA', dfma, dmma),
                    'avg_launch_ms': dk['ms'] / dk['launches'] if dk['launches'] else None,
                    'algorithmic_flops_per_launch': dk['flops_per_launch'], 'algorithmic_bytes_per_launch': dk['bytes_per_launch'],
                    'hbm_gbs_achieved': dk['gbs'], 'families': kernels, 'all_kernels': other,
                    'profile_frames': PROFILE_FRAMES}
        line = {'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': t_dev / K,
                'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
                'config': workload_config(args, world), 'wall_ms_per_step': wall_ms / K,
                'filter_frames_per_s': F * world * K / (t_dev * 1e-3), 'feature_updates_per_step': total_upd / K,
                'e2e': {'value': upd_e2e / e2e_t, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                        'ms_per_step': e2e_t / K * 1e3, 'filter_frames_per_s': F * world * K / e2e_t},
                'gpu_launches': int(l1 - l0), 'clocks': clocks, 'roofline': roofline}
        if not args.no_latency:
            line['latency'] = single_filter_latency(local, args.feats)
        if not args.no_cpu_baseline:
            line['cpu_baseline'] = cpu_baseline_leg(args)
        print(json.dumps(line))
    batch.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_b200(a)
