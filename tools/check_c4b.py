"""Development aid: generic engine, paper block, step time against the batch and the cluster size."""
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
frames = 4
T = frames * hp.hop_size
for B, C in [(int(x.split(':')[0]), int(x.split(':')[1])) for x in sys.argv[1:]]:
    rng = np.random.default_rng(0)
    c = rng.uniform(0, 1, (B, frames, hp.num_mels)).astype(np.float32)
    nz = make_noise(hp, T, B, rng)
    os.environ['WN_GEN_CLUSTER'] = str(C)
    net = NativeWaveNet(hp, w, engine='generic')
    net.synthesize_host(c, nz)
    ms = net.info()['last_loop_ms']
    print('paper B={} cluster={}: {:.1f} us/step'.format(B, C, 1e3 * ms / T), flush=True)
    net.close()
