"""Development aid: the pipelined engine with groups of 8 utterances (WN_PIPE_NU=8) against the C oracle, bit for bit."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200')):
    if p not in sys.path:
        sys.path.insert(0, p)
os.environ['WN_PIPE_NU'] = sys.argv[1] if len(sys.argv) > 1 else '8'
import hparams as hp_mod
from wavenet_vocoder import weights as W
from wavenet_vocoder._native import NativeWaveNet
from oracle.c_oracle import COracle

hp = hp_mod.default_22k()
w = W.init_weights(hp, seed=5339, kind='stress')
for B, Tc in ((11, 2), (8, 1), (1, 1), (19, 1)):
    T = Tc * hp.hop_size
    rng = np.random.default_rng(B)
    c = rng.uniform(0, 1, (B, Tc, hp.num_mels)).astype(np.float32)
    noise = rng.standard_normal((T, B, 1)).astype(np.float32)
    net = NativeWaveNet(hp, w, engine='pipeline')
    info = net.info()
    a_ref, _, raw_ref, _ = COracle(hp, w, ksplit=info['ksplit']).synthesize(c, noise=noise)
    audio, raw = net.synthesize_host(c, noise, want_raw=True)
    print('NU={} B={} T={} group_size={} audio equal={} raw equal={}'.format(
        os.environ['WN_PIPE_NU'], B, T, info.get('pipeline_group_size'), np.array_equal(audio, a_ref), np.array_equal(raw, raw_ref)), flush=True)
    net.close()
