"""hparams shim: parse semantics, presets, eager validation of the reference's asserts."""
import numpy as np
import pytest

import hparams as hp_mod
import paper_hparams
from wavenet_vocoder import weights as W


def test_parse_overrides_types():
    hp = hp_mod.hparams.copy()
    hp.parse('sample_rate=22050,hop_size=275,upsample_scales=[5, 5, 11],legacy=False,input_type=mulaw,NN_scaler=0.5')
    assert hp.sample_rate == 22050 and hp.hop_size == 275
    assert hp.upsample_scales == [5, 5, 11] and hp.legacy is False
    assert hp.input_type == 'mulaw' and hp.NN_scaler == 0.5
    with pytest.raises(ValueError):
        hp.parse('no_such_key=1')
    assert hp.parse('') is hp


def test_reference_default_is_inconsistent_and_caught_early():
    # code/hparams.py ships hop 200 with upsample_scales [11, 25]; the reference only fails at run time (wavenet.py:800)
    with pytest.raises(ValueError, match='upsample_scales'):
        hp_mod.validate_wavenet_hparams(hp_mod.hparams)
    hp_mod.validate_wavenet_hparams(hp_mod.default_22k())
    hp_mod.validate_wavenet_hparams(paper_hparams.hparams)


def test_known_answers_param_count_and_receptive_field():
    d, p = hp_mod.default_22k(), paper_hparams.hparams
    assert W.count_params(d) == 3063786
    assert W.count_params(p) == 13676896
    assert W.receptive_field_size(d.layers, d.stacks, d.kernel_size) == 4093
    assert W.receptive_field_size(p.layers, p.stacks, p.kernel_size) == 505
    assert W.layer_dilations(d) == [2 ** (i % 10) for i in range(20)]


def test_constructor_asserts():
    d = hp_mod.default_22k()
    with pytest.raises(RuntimeError, match='quantize_chennels'):
        hp_mod.validate_wavenet_hparams(d.override(input_type='mulaw-quantize', quantize_channels=256, out_channels=30))
    with pytest.raises(AssertionError):
        hp_mod.validate_wavenet_hparams(d.override(layers=21))
    with pytest.raises(AssertionError):
        hp_mod.validate_wavenet_hparams(d.override(num_mels=64))
    with pytest.raises(AssertionError):
        hp_mod.validate_wavenet_hparams(d.override(input_type='pcm'))
    with pytest.raises(AssertionError):
        hp_mod.validate_wavenet_hparams(d.override(out_channels=7))


def test_init_weights_deterministic_and_shaped():
    d = hp_mod.default_22k()
    a, b = W.init_weights(d, seed=1), W.init_weights(d, seed=1)
    assert list(a) == list(W.weight_specs(d))
    assert all(np.array_equal(a[k], b[k]) for k in a)
    assert W.check_weights(d, a)
    assert a['final_convolution_2/bias'][1] == -4.0
    t = W.init_weights(d, kind='tf-init')
    k = t['local_conditioning_upsampling_1/kernel']  # NN init: centre tap = NN_scaler**(1/2), everything else zero
    assert np.allclose(k[1, 1, 0, :], 0.3 ** 0.5) and np.count_nonzero(k) == k.shape[-1]
    bad = dict(a)
    bad.pop('input_convolution/bias')
    with pytest.raises(ValueError):
        W.check_weights(d, bad)
