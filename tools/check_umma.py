"""Development aid: WaveNet.step on the tensor-core path (wn_step_umma.cu) against the fp32 FFMA GEMMs and timing at C5."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200'), os.path.join(ROOT, 'tests')):
    if p not in sys.path:
        sys.path.insert(0, p)
import common
from wavenet_vocoder.models import create_model


def run(hp, B, Tc, label, serial=True):
    w = common.weights_for(hp)
    T = Tc * hp.hop_size
    c = common.cond_for(hp, B, Tc)
    x = common.teacher_for(hp, B, T)
    model = create_model('WaveNet', hp)
    model.load_weights(w)
    ct = np.transpose(c, [0, 2, 1])
    os.environ.pop('WN_STEP_FFMA', None); os.environ.pop('WN_STEP_SERIAL', None)
    y = model.step(x[:, None, :], c=ct)
    os.environ['WN_STEP_FFMA'] = '1'
    y_f = model.step(x[:, None, :], c=ct)
    os.environ.pop('WN_STEP_FFMA')
    msg = '{}: B={} T={} finite={} umma vs ffma max rel {:.3e}'.format(label, B, T, bool(np.isfinite(y).all()), common.max_rel(y, y_f))
    if serial:
        os.environ['WN_STEP_SERIAL'] = '1'
        y_s = model.step(x[:, None, :], c=ct)
        os.environ.pop('WN_STEP_SERIAL')
        msg += '  umma vs serial {:.3e}  ffma vs serial {:.3e}'.format(common.max_rel(y, y_s), common.max_rel(y_f, y_s))
    print(msg, flush=True)
    return model


run(common.default_22k(), 3, 3, 'default_22k')
run(common.small(residual_channels=64, gate_channels=128, skip_out_channels=64), 2, 40, 'R64 G128 S64')
run(common.small(common.paper(), layers=6, stacks=3, residual_channels=256, gate_channels=512, skip_out_channels=256), 2, 20, 'paper block x 6 layers')
if '--time' in sys.argv:
    import torch
    from wavenet_vocoder import weights as W
    from wavenet_vocoder._native import NativeWaveNet
    hp = common.default_22k()
    B, frames = 8, 802
    T = frames * hp.hop_size
    net = NativeWaveNet(hp, W.init_weights(hp, seed=5339, kind='tf-init'))
    rng = np.random.default_rng(0)
    dc = torch.from_numpy(rng.uniform(0, 1, (B, frames, hp.num_mels)).astype(np.float32)).cuda()
    dx = torch.from_numpy(rng.uniform(-1, 1, (B, T)).astype(np.float32)).cuda()
    raws = {}
    for ffma in (1, 0, 0):
        os.environ['WN_STEP_FFMA'] = str(ffma)
        raw = torch.empty((B, T, hp.out_channels), dtype=torch.float32, device='cuda')
        for _ in range(2):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            net.step_full(dx, dc, B, frames, raw, stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize(); dt = time.perf_counter() - t0
        raws[ffma] = raw.cpu().numpy()
        print('C5 (8 x 10 s) {}: {:.1f} ms wall, loop_ms {:.1f}'.format('fp32 FFMA2' if ffma else 'tcgen05 bf16x3', dt * 1e3, net.info()['last_loop_ms']), flush=True)
    print('C5 umma vs ffma max rel {:.3e}'.format(common.max_rel(raws[0], raws[1])), flush=True)
