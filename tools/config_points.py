"""Development aid: times the BASELINE.json configurations C1..C5 (or bounded samples of them) on one GPU through the
C ABI (NativeWaveNet) and prints one line per point.  Runs on the GPU box only.

  python tools/config_points.py [--quick]
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200')):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--quick', action='store_true')
    args = ap.parse_args()
    import torch
    import hparams as hp_mod
    import paper_hparams
    from wavenet_vocoder import weights as W
    from wavenet_vocoder._native import NativeWaveNet
    from wavenet_vocoder.models.wavenet import make_noise

    def run(name, hp, B, frames, engine='auto', step=False, note=''):
        w = W.init_weights(hp, seed=5339, kind='tf-init')
        net = NativeWaveNet(hp, w, engine=engine)
        T = frames * hp.hop_size
        rng = np.random.default_rng(0)
        c = rng.uniform(0, 1, (B, frames, hp.num_mels)).astype(np.float32)
        t0 = time.perf_counter()
        if step:
            x = rng.uniform(-1, 1, (B, T)).astype(np.float32)
            dx, dc = torch.from_numpy(x).cuda(), torch.from_numpy(c).cuda()
            raw = torch.empty((B, T, hp.out_channels), dtype=torch.float32, device='cuda')
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(2):   # the first call also allocates the scratch
                t0 = time.perf_counter()
                net.step_full(dx, dc, B, frames, raw, stream=torch.cuda.current_stream().cuda_stream)
                torch.cuda.synchronize()
            ok = bool(torch.isfinite(raw).all())
        else:
            nz = make_noise(hp, T, B, rng)
            t0 = time.perf_counter()
            a = net.synthesize_host(c, nz)
            ok = bool(np.isfinite(a).all())
        wall = time.perf_counter() - t0
        info = net.info()
        loop_ms = info['last_loop_ms']
        print('{:34s} engine={:8s} B={:3d} T={:7d}  loop {:10.1f} ms  {:8.2f} us/step  {:10.0f} samples/s  RTF {:7.2f}  wall {:7.2f} s  finite={} {}'.format(
            name, str(info['engine']), B, T, loop_ms, 1e3 * loop_ms / T, B * T / (loop_ms / 1e3),
            B * T / (loop_ms / 1e3) / hp.sample_rate, wall, ok, note), flush=True)
        net.close()

    d = hp_mod.default_22k()
    pp = paper_hparams.hparams.copy() if hasattr(paper_hparams.hparams, 'copy') else paper_hparams.hparams
    q = args.quick
    run('C1 default B=1 x 1 s', d, 1, 81)
    run('C1 default B=1 x 1 s', d, 1, 81, engine='generic')
    run('C3/8 default B=32 x 10 s', d, 32, 802 if not q else 80)
    run('C2 default B=64 x 5 s', d, 64, 401 if not q else 40)
    run('C3/2 default B=128 x 10 s', d, 128, 802 if not q else 80, note='(sample of the 10 s case)' if q else '')
    run('C5 default step() B=8 x 10 s', d, 8, 802 if not q else 80, step=True)
    run('C4 paper B=16, 1 s sample of 30 s', pp, 16, 81 if not q else 8, note='(bounded sample)')


if __name__ == '__main__':
    main()
