"""Scratch (CPU only): how far two runs of the CPU oracle drift apart when the feature measurements of one of them are
moved by one unit in the last place -- the sensitivity of the reference algorithm itself to rounding, as a yardstick for the
CUDA-vs-oracle differences reported by bench.py's parity_spot.
usage: python tools/oracle_sensitivity.py [n_seeds] [n_frames]"""
import multiprocessing as mp
import os
import sys

os.environ.setdefault('OPENBLAS_NUM_THREADS', '1')     # before numpy: one BLAS thread per worker process
import numpy as np  # noqa: E402

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def run(job):
    from oracle import msckf_oracle as orc
    import msckf_b200  # noqa: F401
    from msckf_b200 import synth
    seed, n_frames, ulps = job
    st = synth.make_structure(n_frames, 300, structure_seed=0)
    stream = synth.realize(st, noise_seed=seed)
    if ulps:
        stream.frame_uv = [np.nextafter(uv, np.inf) for uv in stream.frame_uv]
    cfg = orc.ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = 20
    f = orc.MSCKF(cfg)
    snaps = []

    def on_frame(fi, flt, res):
        if fi % 10 == 9:
            ss = flt.state_server
            imu = ss._vio_imu_state__
            snaps.append((np.concatenate([imu.orientation, imu._vio_gyro_bias__, imu.velocity, imu.acc_bias, imu._vio_position__]),
                          np.array(ss._vio_state_cov__), list(flt.map_server.keys())))
    synth.drive(f, stream, n_frames, on_frame=on_frame)
    return snaps


if __name__ == '__main__':
    n_seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    n_frames = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    jobs = [(s, n_frames, u) for s in range(n_seeds) for u in (0, 1)]
    with mp.get_context('spawn').Pool(max(1, (os.cpu_count() or 2) - 1)) as pool:
        out = pool.map(run, jobs)
    worst = []
    for s in range(n_seeds):
        a, b = out[2 * s], out[2 * s + 1]
        errs = []
        for (ia, Pa, ma), (ib, Pb, mb) in zip(a, b):
            if Pa.shape != Pb.shape or ma != mb:
                errs.append(float('inf'))
            else:
                errs.append(max(np.linalg.norm(ia - ib) / np.linalg.norm(ia), np.linalg.norm(Pa - Pb) / np.linalg.norm(Pa)))
        worst.append(max(errs))
        print('seed %2d  ' % s + ' '.join('%.0e' % e for e in errs), flush=True)
    w = np.array(worst)
    print('one ulp in the measurements -> relative difference of state / P after %d frames: median %.1e, max %.1e; %d of %d seeds above 1e-11, %d above 1e-10'
          % (n_frames, np.median(w), w.max(), int((w > 1e-11).sum()), n_seeds, int((w > 1e-10).sum())))
