"""Pins the oracles to the reference's OWN Python.  tests/golden/ref_*.npz were produced by tests/golden/make_ref_golden.py,
which imports /root/reference/code (util.py, models/{wavenet,modules,gaussian,mixture}.py, datasets/audio.py, hparams.py,
paper_hparams.py) over the numpy `tf` stand-in tests/golden/tf_shim.py and runs create_model -> WaveNet.initialize
(synthesis branch) -> WaveNet.incremental (free running and teacher forced) and WaveNet.step on seeded small models.

Here oracle/wavenet_oracle.py (numpy restatement) and oracle/c (fixed-order C restatement, the bit-level checker of the CUDA
kernels) must reproduce those outputs: mu-law / quantisation / samplers' discrete choices exactly, float results to 1e-5
(both sides are float32 numpy; only the association of a few sums differs).  What this does NOT pin: the TF kernels behind
matmul / conv2d / conv2d_transpose / batch_to_space_nd / resize_images - the stand-in restates those from their published
semantics (header of tf_shim.py).
"""
import json
import os

import numpy as np
import pytest

import common
import hparams as hp_mod
import paper_hparams
from oracle import wavenet_oracle as O
from oracle.c_oracle import COracle

G = common.GOLDEN_DIR
VARIANTS = ['gauss_subpixel', 'mol_2d_nolegacy', 'mulaw_quantize', 'mulaw_1d', 'nearest_nobias_k2', 'subpixel_no_nninit_odd']
TOL = 1e-5


def _hp(name):
    v = common.small_variants()
    if name == 'mulaw_1d':   # the reference's synthesis branch rejects global conditioning (wavenet.py:430); run without it
        return v['mulaw_gin_1d'].override(gin_channels=-1, use_speaker_embedding=False, wavenet_dropout=0.0)
    return v[name].override(wavenet_dropout=0.0)


def _close(a, b, tol=TOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0))) <= tol


def test_util_helpers_match_reference_execution():
    d = np.load(os.path.join(G, 'ref_util.npz'))
    from wavenet_vocoder import util as U
    x, q = d['x'], d['q']
    for mod in (U, O):
        assert np.array_equal(np.asarray(mod.mulaw_quantize(x, 256)), d['mulaw_quantize'])       # integer classes: exact
        assert np.allclose(mod.mulaw(x, 256), d['mulaw'], rtol=0, atol=1e-7)
        assert np.allclose(mod.inv_mulaw(x, 256), d['inv_mulaw'], rtol=0, atol=1e-7)
        assert np.allclose(mod.inv_mulaw_quantize(q, 256), d['inv_mulaw_quantize'], rtol=0, atol=1e-7)
        assert int(mod.mulaw_quantize(0, 256)) == int(d['mulaw_quantize_0']) == 127
    # Synthesizer prologue: _interp (feeder.py:426-428)
    from wavenet_vocoder.synthesizer import _interp
    assert np.array_equal(_interp(d['interp_in'], (-4.0, 4.0)), d['interp_out'])
    # save_wavenet_wav (datasets/audio.py:17-20)
    assert np.array_equal(O.wav_to_int16(d['wav_in']), d['wav_int16'])
    from wavenet_vocoder.weights import receptive_field_size
    assert [receptive_field_size(20, 2, 3), receptive_field_size(24, 4, 3)] == list(d['receptive_field']) == [4093, 505]


def test_hparams_blocks_match_reference_modules():
    ref = json.load(open(os.path.join(G, 'ref_hparams.json')))
    mine = {'default': hp_mod.hparams, 'paper': paper_hparams.hparams}
    for tag in ('default', 'paper'):
        vals = mine[tag].values()
        for k, v in ref[tag].items():
            assert k in vals, (tag, k)
            if isinstance(v, float):
                assert abs(vals[k] - v) <= 1e-12 * max(1.0, abs(v)), (tag, k, vals[k], v)
            else:
                assert vals[k] == v, (tag, k, vals[k], v)
    # default_22k = the reference block with the two audio keys of SURVEY.md F5
    d22 = hp_mod.default_22k().values()
    for k, v in ref['default'].items():
        if k not in ('sample_rate', 'hop_size'):
            assert d22[k] == v or (isinstance(v, float) and abs(d22[k] - v) < 1e-12), k
    assert d22['sample_rate'] == 22050 and d22['hop_size'] == 275


def test_samplers_match_reference_execution():
    d = np.load(os.path.join(G, 'ref_samplers.npz'))
    hp = common.small().override(log_scale_min_gauss=-7.0, log_scale_min=-7.0)
    o = O.WaveNetOracle(hp, common.weights_for(hp))
    assert _close(o.sample_gaussian(d['y'][:, :, 0], d['eps']).reshape(-1), d['gauss'].reshape(-1), 1e-6)
    hpm = common.small_variants()['mol_2d_nolegacy'].override(log_scale_min=-7.0)
    om = O.WaveNetOracle(hpm, common.weights_for(hpm))
    u = np.concatenate([d['u1'][:, 0, :], d['u2']], axis=1)
    assert _close(om.sample_mol(d['ym'][:, :, 0], u).reshape(-1), d['mol'].reshape(-1), 1e-6)


@pytest.mark.parametrize('name', VARIANTS)
def test_numpy_oracle_matches_reference_execution(name):
    hp = _hp(name)
    w = common.weights_for(hp)
    d = np.load(os.path.join(G, 'ref_model_{}.npz'.format(name)))
    o = O.WaveNetOracle(hp, w)
    y, raw, c_up = o.synthesize(d['c'], noise=d['noise'])
    assert _close(raw, d['free_raw']) and _close(y, d['free_y_hat'])
    if hp.input_type == 'mulaw-quantize':   # same classes sampled, exactly
        assert np.array_equal(y, d['free_y_hat'])
    if 'c_up' in d.files:
        assert _close(c_up, d['c_up'])
    y_t, raw_t, _ = o.synthesize(d['c'], noise=d['noise'], test_inputs=d['teacher'])
    assert _close(raw_t, d['forced_raw']) and _close(y_t, d['forced_y_hat'])
    # WaveNet.step (wavenet.py:650-721)
    if hp.input_type != 'mulaw-quantize':
        x = d['teacher'][:, None, :].astype(np.float32)
    else:
        x = np.transpose(np.eye(hp.quantize_channels, dtype=np.float32)[d['teacher'].astype(np.int64)], [0, 2, 1])
    assert _close(o.step(x, np.transpose(d['c'], [0, 2, 1])), d['step_raw'])


@pytest.mark.parametrize('name', VARIANTS)
def test_c_oracle_matches_reference_execution(name):
    """The fixed-order C restatement (what the CUDA kernels are compared with bit for bit) against the reference-executed
    outputs: teacher forced within the BASELINE.json criterion, free-running audio to 1e-4 over the fixture's 45 samples."""
    hp = _hp(name)
    w = common.weights_for(hp)
    d = np.load(os.path.join(G, 'ref_model_{}.npz'.format(name)))
    co = COracle(hp, w)
    a_t, _, raw_t, _ = co.synthesize(d['c'], noise=d['noise'], teacher=d['teacher'])
    ref_raw = d['forced_raw'] if hp.input_type == 'mulaw-quantize' else np.transpose(d['forced_raw'], [0, 2, 1])
    assert np.all(common.rel_close(raw_t, ref_raw))
    a, _, raw, _ = co.synthesize(d['c'], noise=d['noise'])
    assert float(np.max(np.abs(a - d['free_y_hat']))) <= 1e-4
