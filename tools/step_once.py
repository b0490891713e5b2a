"""Development aid: one WaveNet.step call (default_22k, B utterances x frames) through the C ABI - a target for ncu."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200')):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402
import hparams as hp_mod  # noqa: E402
from wavenet_vocoder import weights as W  # noqa: E402
from wavenet_vocoder._native import NativeWaveNet  # noqa: E402

B, frames = int(sys.argv[1]) if len(sys.argv) > 1 else 8, int(sys.argv[2]) if len(sys.argv) > 2 else 81
hp = hp_mod.default_22k()
net = NativeWaveNet(hp, W.init_weights(hp, seed=5339, kind='tf-init'))
T = frames * hp.hop_size
rng = np.random.default_rng(0)
dx = torch.from_numpy(rng.uniform(-1, 1, (B, T)).astype(np.float32)).cuda()
dc = torch.from_numpy(rng.uniform(0, 1, (B, frames, hp.num_mels)).astype(np.float32)).cuda()
raw = torch.empty((B, T, hp.out_channels), dtype=torch.float32, device='cuda')
for _ in range(2):
    net.step_full(dx, dc, B, frames, raw, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
print('step ms', net.info()['last_loop_ms'], 'TFLOP/s', B * T * 6095616 / (net.info()['last_loop_ms'] / 1e3) / 1e12)
net.close()
