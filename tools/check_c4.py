"""Development aid: the generic engine on the paper block (C4 sample) for several cluster sizes, with a bit check between them."""
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
w = W.init_weights(hp, seed=5339, kind='stress')
B, frames = 16, int(sys.argv[1]) if len(sys.argv) > 1 else 8
T = frames * hp.hop_size
rng = np.random.default_rng(0)
c = rng.uniform(0, 1, (B, frames, hp.num_mels)).astype(np.float32)
nz = make_noise(hp, T, B, rng)
ref = None
for C in (1, 2, 4, 8, 0):
    if C:
        os.environ['WN_GEN_CLUSTER'] = str(C)
    else:
        os.environ.pop('WN_GEN_CLUSTER', None)
    net = NativeWaveNet(hp, w, engine='generic')
    a = net.synthesize_host(c, nz)
    ms = net.info()['last_loop_ms']
    if ref is None:
        ref = a
    print('C4 sample B={} T={} cluster={}: loop {:.1f} ms = {:.1f} us/step, {:.0f} samples/s (RTF {:.2f}), identical to C=1: {}'.format(
        B, T, C or 'auto', ms, 1e3 * ms / T, B * T / ms * 1e3, B * T / ms * 1e3 / hp.sample_rate, np.array_equal(a, ref)), flush=True)
    net.close()
