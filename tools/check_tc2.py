"""Development aid: the 256-channel tensor-core engine against the C oracle, teacher forced, per-utterance error.
  python tools/check_tc2.py B [Tc]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'tacotron-2_b200'), os.path.join(ROOT, 'tests')]
import numpy as np
import common
from oracle.c_oracle import COracle
from wavenet_vocoder._native import NativeWaveNet

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
Tc = int(sys.argv[2]) if len(sys.argv) > 2 else 1
hp = common.paper()
w = common.weights_for(hp)
net = NativeWaveNet(hp, w, engine='tc')
T = Tc * hp.hop_size
c, nz, teacher = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B), common.teacher_for(hp, B, T)
_, raw = net.synthesize_host(c, nz, teacher=teacher, want_raw=True)
_, _, raw_c, _ = COracle(hp, w).synthesize(c, noise=nz, teacher=teacher)
err = np.abs(raw - raw_c) / np.maximum(np.abs(raw_c), 1e-2)
print('info', net.info())
print('per-utterance max rel err:', ' '.join('{:.1e}'.format(e) for e in err.max(axis=(1, 2))))
bad = np.argwhere(err > 1e-3)
print('first bad (b, t, o):', bad[:5].tolist())
