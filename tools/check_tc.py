"""Development aid: the tensor-core loop engine against the fp32 pipelined engine (teacher forced and free running)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'tacotron-2_b200'), os.path.join(ROOT, 'tests')]
import numpy as np
import common
from wavenet_vocoder._native import NativeWaveNet

B = int(sys.argv[1]) if len(sys.argv) > 1 else 5
Tc = int(sys.argv[2]) if len(sys.argv) > 2 else 2
hp = common.default_22k()
w = common.weights_for(hp)
T = Tc * hp.hop_size
c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B)
teacher = common.teacher_for(hp, B, T)
ref = NativeWaveNet(hp, w, engine='pipeline')
tc = NativeWaveNet(hp, w, engine='tc')
t0 = time.time()
a_tc, r_tc = tc.synthesize_host(c, nz, teacher=teacher, want_raw=True)
print('tc teacher-forced call: %.2f s, loop %.2f ms = %.2f us/step, info %s' % (time.time() - t0, tc.info()['last_loop_ms'], tc.info()['last_loop_ms'] * 1e3 / T, tc.info()))
a_ref, r_ref = ref.synthesize_host(c, nz, teacher=teacher, want_raw=True)
print('teacher forced: raw max rel err %.3e (floor 1e-2), max abs %.3e; finite %s' % (common.max_rel(r_tc, r_ref), np.abs(r_tc - r_ref).max(), np.isfinite(r_tc).all()))
for b in range(min(B, 3)):
    print('  utt', b, 'raw[0..2]', r_tc[b, :3].ravel(), 'ref', r_ref[b, :3].ravel())
a_tc, r_tc = tc.synthesize_host(c, nz, want_raw=True)
a_ref, r_ref = ref.synthesize_host(c, nz, want_raw=True)
d = np.abs(a_tc - a_ref)
first = [int(np.argmax(d[b] > 1e-3)) if (d[b] > 1e-3).any() else T for b in range(B)]
print('free running: max |audio diff| %.3e; first sample with |diff| > 1e-3 per utterance: %s' % (d.max(), first))
print('loop ms', tc.info()['last_loop_ms'], 'us/step', tc.info()['last_loop_ms'] * 1e3 / T, ' ref (pipeline) us/step', ref.info()['last_loop_ms'] * 1e3 / T)
