"""Likelihood metrics of the reference (models/mixture.py:18-74, models/gaussian.py:5-37) restated on the host: closed
forms and hand-built parameters (SURVEY.md 8c check vii), plus the eval-mode loss through the CUDA step() path."""
import math

import numpy as np
import pytest

import common
from wavenet_vocoder.models.gaussian import gaussian_maximum_likelihood_estimation_loss as gauss_loss
from wavenet_vocoder.models.mixture import discretized_mix_logistic_loss as mol_loss, log_prob_from_logits, log_sum_exp


def test_log_helpers():
    x = np.array([[1000.0, 1000.0], [-1000.0, 0.0]])
    assert np.allclose(log_sum_exp(x), [1000.0 + math.log(2), 0.0])
    lp = log_prob_from_logits(np.array([[0.0, 0.0, 0.0, 0.0]]))
    assert np.allclose(lp, math.log(0.25))


def test_gaussian_density_and_bin_mass():
    B, T = 2, 7
    rng = np.random.default_rng(0)
    mean, ls = rng.uniform(-0.5, 0.5, (B, T)), rng.uniform(-3, -1, (B, T))
    y = (mean + rng.uniform(-2, 2, (B, T)) * np.exp(ls))[..., None]   # within two standard deviations
    y_hat = np.stack([mean, ls], axis=1)
    pdf = gauss_loss(y_hat, y, -16.0, 65536, use_cdf=False, reduce=False)[..., 0]
    want = 0.5 * math.log(2 * math.pi) + ls + 0.5 * (y[..., 0] - mean) ** 2 / np.exp(2 * ls)
    assert np.allclose(pdf, want)
    assert math.isclose(gauss_loss(y_hat, y, -16.0, 65536, use_cdf=False), want.sum())
    # the mass of a narrow bin is density * width
    cdf = gauss_loss(y_hat, y, -16.0, 65536, use_cdf=True, reduce=False)[..., 0]
    assert np.allclose(cdf, want - math.log(2.0 / 65535), atol=1e-4)
    # log_scale_min clamps
    clamped = gauss_loss(np.stack([mean, ls - 100], axis=1), y, -2.0, 65536, use_cdf=False, reduce=False)[..., 0]
    assert np.allclose(clamped, 0.5 * math.log(2 * math.pi) + np.maximum(ls - 100, -2.0) + 0.5 * (y[..., 0] - mean) ** 2 / np.exp(-4.0))
    with pytest.raises(ValueError):
        gauss_loss(np.zeros((1, 3, 4)), np.zeros((1, 4, 1)), -7.0, 256)


def test_mixture_single_component_is_a_discretized_logistic():
    # one dominant component: P(bin) = sigmoid((y - m + h) / s) - sigmoid((y - m - h) / s), h = 1 / (classes - 1)
    K, T = 3, 5
    y = np.linspace(-0.8, 0.8, T).reshape(1, T, 1)
    m, ls = 0.1, -2.0
    y_hat = np.zeros((1, 3 * K, T))
    y_hat[0, 0] = 50.0          # logits: component 0 wins
    y_hat[0, K:2 * K] = m
    y_hat[0, 2 * K:] = ls
    h, s = 1.0 / 255, math.exp(ls)
    sig = lambda v: 1.0 / (1.0 + np.exp(-v))
    want = -np.log(sig((y[0, :, 0] - m + h) / s) - sig((y[0, :, 0] - m - h) / s))
    got = mol_loss(y_hat, y, num_classes=256, log_scale_min=-7.0, reduce=False)[0, :, 0]
    assert np.allclose(got, want, atol=1e-9)
    assert math.isclose(mol_loss(y_hat, y, num_classes=256), want.sum(), rel_tol=1e-9)
    # edge bins: y = -1 takes the whole lower tail, y = +1 the whole upper tail
    ye = np.array([-1.0, 1.0]).reshape(1, 2, 1)
    got = mol_loss(y_hat[:, :, :2], ye, num_classes=256, reduce=False)[0, :, 0]
    assert np.allclose(got, [-np.log(sig((-1 - m + h) / s)), -np.log(1 - sig((1 - m - h) / s))], atol=1e-9)
    # equal mixture of identical components = the component
    y_hat2 = y_hat.copy(); y_hat2[0, :K] = 0.0
    assert np.allclose(mol_loss(y_hat2, y, reduce=False), mol_loss(y_hat, y, reduce=False), atol=1e-9)
    with pytest.raises(ValueError):
        mol_loss(np.zeros((1, 4, 4)), np.zeros((1, 4, 1)))


@pytest.mark.gpu
@pytest.mark.parametrize('name', ['gauss_subpixel', 'mol_2d_nolegacy', 'mulaw_quantize'])
def test_eval_loss_through_the_cuda_step_path(name):
    """WaveNet.eval_loss: parameters from step() on the GPU, likelihood on the host; against the numpy oracle's step()."""
    from oracle import wavenet_oracle as O
    from wavenet_vocoder.models import create_model
    hp = common.small_variants()[name]
    w = common.weights_for(hp)
    B, Tc = 2, 3
    T = Tc * hp.hop_size
    c = np.transpose(common.cond_for(hp, B, Tc), [0, 2, 1])
    y = common.teacher_for(hp, B, T)
    model = create_model('WaveNet', hp)
    model.load_weights(w)
    got = model.eval_loss(y, c=c, reduce=False)
    assert got.shape == (B, T - 1) and np.isfinite(got).all()
    x = y[:, None, :] if hp.input_type != 'mulaw-quantize' else np.eye(hp.quantize_channels, dtype=np.float32)[y.astype(np.int64)].transpose(0, 2, 1)
    y_hat = O.WaveNetOracle(hp, w).step(x, c=c)[:, :, :-1]
    tgt = y[:, 1:, None]
    if hp.input_type == 'mulaw-quantize':
        want = -np.take_along_axis(log_prob_from_logits(np.transpose(y_hat, [0, 2, 1])), tgt.astype(np.int64), axis=-1)[..., 0]
    elif hp.out_channels == 2:
        want = gauss_loss(y_hat, tgt, hp.log_scale_min_gauss, hp.quantize_channels, use_cdf=getattr(hp, 'cdf_loss', False), reduce=False)[..., 0]
    else:
        want = mol_loss(y_hat, tgt, num_classes=hp.quantize_channels, log_scale_min=hp.log_scale_min, reduce=False)[..., 0]
    assert np.allclose(got, want, rtol=2e-3, atol=2e-3)
    assert math.isclose(model.eval_loss(y, c=c), float(np.mean(got)), rel_tol=1e-9)
