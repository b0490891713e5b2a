"""The C restatement (fixed accumulation order, include/wn_math.h) against the op-for-op numpy oracle."""
import numpy as np
import pytest

import common
from oracle import wavenet_oracle as O
from oracle.c_oracle import COracle, math_fn


def _ulp_err(y, ref64):
    sp = np.spacing(np.abs(ref64.astype(np.float32))).astype(np.float64)
    return float(np.max(np.abs(y.astype(np.float64) - ref64) / sp))


def test_wn_math_accuracy():
    x = np.linspace(-87, 88, 400001).astype(np.float32)
    assert _ulp_err(math_fn('exp', x), np.exp(x.astype(np.float64))) <= 1.5
    x = np.linspace(-12, 12, 400001).astype(np.float32)
    assert _ulp_err(math_fn('tanh', x), np.tanh(x.astype(np.float64))) <= 2.0
    x = np.linspace(-30, 30, 400001).astype(np.float32)
    assert _ulp_err(math_fn('sigmoid', x), 1 / (1 + np.exp(-x.astype(np.float64)))) <= 3.0
    x = np.exp(np.linspace(-12, 3, 400001)).astype(np.float32)
    assert _ulp_err(math_fn('log', x), np.log(x.astype(np.float64))) <= 1.5
    # edges
    assert math_fn('exp', np.array([-200., 100., 0.], np.float32)).tolist() == [0.0, np.inf, 1.0]
    assert math_fn('sigmoid', np.array([-200., 200.], np.float32)).tolist() == [0.0, 1.0]
    assert math_fn('tanh', np.array([-50., 50., 0.], np.float32)).tolist() == [-1.0, 1.0, 0.0]
    y = np.linspace(-1, 1, 2001).astype(np.float32)
    assert np.max(np.abs(math_fn('inv_mulaw', y) - O.inv_mulaw(y.astype(np.float64)))) < 1e-6


@pytest.mark.parametrize('name', list(common.small_variants()))
def test_c_oracle_matches_numpy_oracle(name):
    hp = common.small_variants()[name]
    w = common.weights_for(hp)
    orc, co = O.WaveNetOracle(hp, w), COracle(hp, w)
    B, Tc = 3, 3
    T = Tc * hp.hop_size
    c, nz, g = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B), common.speaker_ids(hp, B)
    teacher = common.teacher_for(hp, B, T)
    # upsampling net
    cup = co.upsample(c)
    assert common.max_rel(cup, np.transpose(orc.upsample(np.transpose(c, [0, 2, 1])), [0, 2, 1]), floor=1.0) <= 1e-5  # fp32 rounding of up to 80-term sums, two layers
    # teacher forced: distribution parameters, BASELINE.json criterion (1e-3 relative) with a wide margin
    _, raw_np = orc.incremental(c=c, g=g, noise=nz, test_inputs=teacher)
    raw_np = raw_np if hp.input_type == 'mulaw-quantize' else np.transpose(raw_np, [0, 2, 1])
    audio, samples, raw, _ = co.synthesize(c, g=g, noise=nz, teacher=teacher)
    assert common.max_rel(raw, raw_np) < 1e-4
    # free running over a short prefix: identical draws -> same audio up to rounding (categorical: same classes)
    wav_np, _, _ = orc.synthesize(c, g=g, noise=nz)
    audio, samples, raw, _ = co.synthesize(c, g=g, noise=nz)
    if hp.input_type == 'mulaw-quantize':
        assert np.mean(np.abs(audio - wav_np) < 1e-6) > 0.98
    else:
        assert np.max(np.abs(audio - wav_np)) < 1e-4
    # the reference-faithful shift-and-concat queue gives bit-identical results to the ring buffer
    a2, s2, r2, _ = co.synthesize(c, g=g, noise=nz, shiftqueue=True)
    assert np.array_equal(audio, a2) and np.array_equal(raw, r2)


def test_c_oracle_threads_do_not_change_results():
    hp = common.small()
    co = COracle(hp, common.weights_for(hp))
    B, Tc = 5, 2
    c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, Tc * hp.hop_size, B)
    a1 = co.synthesize(c, noise=nz, nthreads=1)[0]
    a4 = co.synthesize(c, noise=nz, nthreads=4)[0]
    assert np.array_equal(a1, a4)
