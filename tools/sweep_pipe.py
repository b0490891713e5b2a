"""Development aid: time the loop kernel for a list of (B, plan) points and optionally dump the per-CTA clock64
counters of the pipelined engine (WN_PIPE_STATS).  Runs on the GPU box only.

  python tools/sweep_pipe.py --frames 40 --points "64:P=4,NU=4 32:P=4,NU=4 64:P=4,NU=8,UT=2" [--stats]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200')):
    if p not in sys.path:
        sys.path.insert(0, p)


LAYER_STAMPS = ['D poll+store', 'sync1', 'prefetch issue', 'mac ph1', 'reduce16', 'gate', 'sync3', 'chain', 'st.async',
                'X first (resid/rs poll)', 'commit', 'sync4', 'X wait+combine+send', 'phase0']


def summarize_stats(path, clock_mhz=1965.0):
    rows = [l.split() for l in open(path) if l.strip() and not l.startswith('#')]
    a = np.array([[float(x) for x in r] for r in rows])
    out = []
    for th, name in ((0, 'thread 0 (dot16 + chain)'), (1, 'exchange thread')):
        lay = a[(a[:, 1] == 0) & (a[:, 2] == th)]
        if not len(lay):
            continue
        items = lay[:, 18].mean()
        per = lay[:, 3:17].mean(axis=0) / items / clock_mhz   # us per item
        out.append('layer CTAs, {}: total {:.2f} us/item'.format(name, per.sum()))
        out.append('    ' + ' | '.join('{} {:.2f}'.format(n, v) for n, v in zip(LAYER_STAMPS[:7], per[:7])))
        out.append('    ' + ' | '.join('{} {:.2f}'.format(n, v) for n, v in zip(LAYER_STAMPS[7:], per[7:])))
    head = a[a[:, 1] == 1]
    if len(head):
        items = head[:, 18].mean()
        per = head[:, 3:7].mean(axis=0) / items / clock_mhz
        out.append('head CTAs us/step: wait+relu {:.2f} conv1 {:.2f} conv2 {:.2f} sample {:.2f} | total {:.2f}'.format(*per, head[:, 17].mean() / items / clock_mhz))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--frames', type=int, default=40)
    ap.add_argument('--points', default='64:P=4,NU=4')
    ap.add_argument('--stats', action='store_true')
    ap.add_argument('--hparams', default='default_22k')
    ap.add_argument('--reps', type=int, default=2)
    ap.add_argument('--profile-range', action='store_true', help='cudaProfilerStart/Stop around the last repetition (ncu --replay-mode app-range)')
    args = ap.parse_args()
    import torch
    import hparams as hp_mod
    from wavenet_vocoder import weights as W
    from wavenet_vocoder._native import NativeWaveNet
    from wavenet_vocoder.models.wavenet import make_noise
    if args.hparams == 'paper':
        import paper_hparams
        hp = paper_hparams.hparams.copy()
    else:
        hp = hp_mod.default_22k()
    w = W.init_weights(hp, seed=5339, kind='tf-init')
    Tc = args.frames
    T = Tc * hp.hop_size
    for pt in args.points.split():
        b, _, plan = pt.partition(':')
        B = int(b)
        for k in list(os.environ):
            if k.startswith('WN_PIPE_'):
                del os.environ[k]
        for kv in filter(None, plan.split(',')):
            k, v = kv.split('=')
            os.environ['WN_PIPE_' + k] = v
        stats_path = os.path.join(ROOT, 'gpurun_out', 'pipe_stats_{}.txt'.format(pt.replace(':', '_').replace(',', '_').replace('=', '')))
        try:
            net = NativeWaveNet(hp, w, engine='pipeline')
        except Exception as e:
            print('{:40s} plan rejected: {}'.format(pt, e), flush=True)
            continue
        rng = np.random.default_rng(0)
        c = rng.uniform(0, 1, (B, Tc, hp.num_mels)).astype(np.float32)
        nz = make_noise(hp, T, B, rng)
        best = 1e30
        for r in range(args.reps):
            if args.stats and r == args.reps - 1:
                if os.path.exists(stats_path):
                    os.remove(stats_path)
                os.environ['WN_PIPE_STATS'] = stats_path
            if args.profile_range and r == args.reps - 1:
                torch.cuda.synchronize(); torch.cuda.profiler.start()
            a = net.synthesize_host(c, nz)
            if args.profile_range and r == args.reps - 1:
                torch.cuda.synchronize(); torch.cuda.profiler.stop()
            best = min(best, net.info()['last_loop_ms'])
        os.environ.pop('WN_PIPE_STATS', None)
        finite = bool(np.isfinite(a).all())   # ablated builds (tools/ablate.sh) may overflow
        us = 1e3 * best / T
        print('{:40s} loop {:9.2f} ms  {:7.2f} us/step  {:10.0f} samples/s  RTF {:6.1f}{}'.format(
            pt, best, us, B * T / (best / 1e3), B * T / (best / 1e3) / 22050, '' if finite else '  (non-finite output)'), flush=True)
        if args.stats and os.path.exists(stats_path):
            for line in summarize_stats(stats_path):
                print('    ' + line, flush=True)
        net.close()


if __name__ == '__main__':
    main()
