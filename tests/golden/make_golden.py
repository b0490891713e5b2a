"""Generates tests/golden/*.npz - committed fixtures that pin the oracles and the CUDA path.

The reference itself cannot produce vectors here (TensorFlow 1.x is not installable in this image and the reference
ships no golden vectors), so parity with TF stays "unpinned"; what these fixtures pin is OUR
restatement: the op-for-op numpy oracle (oracle/wavenet_oracle.py, teacher-forced distribution parameters, the
upsampling network, WaveNet.step) and the fixed-order C oracle (free-running audio, bit exact on any IEEE-754 + fma
machine).  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200'), os.path.join(ROOT, 'tests')):
    if p not in sys.path:
        sys.path.insert(0, p)

import common  # noqa: E402
from oracle import wavenet_oracle as O  # noqa: E402
from oracle.c_oracle import COracle  # noqa: E402

CASES = {
    # name: (hparams builder, B, Tc)
    'small_gauss': (lambda: common.small(), 2, 3),
    'small_mol': (lambda: common.small_variants()['mol_2d_nolegacy'], 2, 3),
    'default_22k_prefix': (lambda: common.default_22k(), 1, 2),
}


def build(name):
    mk, B, Tc = CASES[name]
    hp = mk()
    w = common.weights_for(hp)
    T = Tc * hp.hop_size
    c, nz, teacher = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B), common.teacher_for(hp, B, T)
    orc = O.WaveNetOracle(hp, w)
    c_up = orc.upsample(np.transpose(c, [0, 2, 1]))                       # [B, cin, T]
    _, raw_tf = orc.incremental(c=c, noise=nz, test_inputs=teacher)      # [B, O, T] teacher forced
    step = orc.step(teacher[:, None, :], c=np.transpose(c, [0, 2, 1]))   # [B, O, T]
    out = dict(c=c, noise=nz, teacher=teacher, c_up=c_up.astype(np.float32), raw_teacher_forced=raw_tf.astype(np.float32),
               step=step.astype(np.float32))
    for ks in (1, 4):
        a, _, raw, _ = COracle(hp, w, ksplit=ks).synthesize(c, noise=nz)
        out['audio_free_ksplit{}'.format(ks)] = a
        out['raw_free_ksplit{}'.format(ks)] = raw
    return out


if __name__ == '__main__':
    for name in CASES:
        d = build(name)
        np.savez_compressed(os.path.join(HERE, name + '.npz'), **d)
        print(name, {k: v.shape for k, v in d.items()})
