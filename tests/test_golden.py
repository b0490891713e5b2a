"""Committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py).
CPU: both oracles reproduce the fixtures.  GPU: the sm_100a kernels, through the C ABI, reproduce them -
free-running audio bit for bit (fixed-order fp32), teacher-forced distribution parameters within
|a-b| <= 1e-3 * max(|b|, 1e-2) of the op-for-op numpy oracle (BASELINE.json's fp32 tolerance)."""
import os
import sys

import numpy as np
import pytest

import common
from oracle import wavenet_oracle as O
from oracle.c_oracle import COracle

sys.path.insert(0, common.GOLDEN_DIR)
import make_golden  # noqa: E402

CASES = list(make_golden.CASES)


def _load(name):
    d = np.load(os.path.join(common.GOLDEN_DIR, name + '.npz'))
    mk, B, Tc = make_golden.CASES[name]
    hp = mk()
    return hp, common.weights_for(hp), d


@pytest.mark.parametrize('name', CASES)
def test_c_oracle_reproduces_golden(name):
    hp, w, d = _load(name)
    for ks in (1, 4):
        a, _, raw, c_up = COracle(hp, w, ksplit=ks).synthesize(d['c'], noise=d['noise'])
        assert np.array_equal(a, d['audio_free_ksplit{}'.format(ks)])       # bit exact: fixed fp32 order
        assert np.array_equal(raw, d['raw_free_ksplit{}'.format(ks)])
    assert common.max_rel(np.transpose(c_up, [0, 2, 1]), d['c_up'], floor=1.0) <= 1e-5
    _, _, raw_t, _ = COracle(hp, w).synthesize(d['c'], noise=d['noise'], teacher=d['teacher'])
    assert np.all(common.rel_close(np.transpose(raw_t, [0, 2, 1]), d['raw_teacher_forced']))


@pytest.mark.parametrize('name', CASES[:2])
def test_numpy_oracle_reproduces_golden(name):
    hp, w, d = _load(name)
    orc = O.WaveNetOracle(hp, w)
    _, raw = orc.incremental(c=d['c'], noise=d['noise'], test_inputs=d['teacher'])
    assert np.all(common.rel_close(raw, d['raw_teacher_forced'], rtol=1e-5))     # BLAS order may differ between hosts
    assert np.allclose(orc.upsample(np.transpose(d['c'], [0, 2, 1])), d['c_up'], atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
@pytest.mark.parametrize('engine', ['generic', 'pipeline'])
def test_cuda_reproduces_golden(name, engine):
    from wavenet_vocoder._native import NativeWaveNet
    hp, w, d = _load(name)
    for ks in (1, 4):
        try:
            net = NativeWaveNet(hp, w, engine=engine, ksplit=ks)
        except ValueError:
            assert engine == 'pipeline' and ks == 1 and name.startswith('default')   # one CTA per layer cannot hold a default layer
            continue
        assert net.info()['ksplit'] == ks and net.info()['engine'] == engine
        a, raw = net.synthesize_host(d['c'], d['noise'], want_raw=True)
        assert np.array_equal(a, d['audio_free_ksplit{}'.format(ks)])
        assert np.array_equal(raw, d['raw_free_ksplit{}'.format(ks)])
        _, raw_t = net.synthesize_host(d['c'], d['noise'], teacher=d['teacher'], want_raw=True)
        assert np.all(common.rel_close(np.transpose(raw_t, [0, 2, 1]), d['raw_teacher_forced']))
        net.close()
