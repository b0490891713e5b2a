"""Development aid: one short synthesis of the paper block on the generic engine - a target for ncu."""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200')):
    if p not in sys.path:
        sys.path.insert(0, p)
import paper_hparams
from wavenet_vocoder import weights as W
from wavenet_vocoder._native import NativeWaveNet
from wavenet_vocoder.models.wavenet import make_noise
hp = paper_hparams.hparams.copy()
B, frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2, int(sys.argv[2]) if len(sys.argv) > 2 else 1
T = frames * hp.hop_size
rng = np.random.default_rng(0)
net = NativeWaveNet(hp, W.init_weights(hp, seed=5339, kind='stress'), engine='generic')
c = rng.uniform(0, 1, (B, frames, hp.num_mels)).astype(np.float32)
net.synthesize_host(c, make_noise(hp, T, B, rng))
print('loop ms', net.info()['last_loop_ms'], 'us/step', 1e3 * net.info()['last_loop_ms'] / T)
net.close()
