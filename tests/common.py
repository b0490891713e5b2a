"""Shared builders for the tests: hparams variants, synthetic inputs, tolerance helper."""
import numpy as np

import hparams as hp_mod
import paper_hparams
from wavenet_vocoder import weights as W

GOLDEN_DIR = __import__('os').path.join(__import__('os').path.dirname(__import__('os').path.abspath(__file__)), 'golden')


def default_22k():
    return hp_mod.default_22k()


def paper():
    return paper_hparams.hparams.copy()


def small(base=None, **kw):
    """A few-layer model with the default block's structure (fast on the numpy oracle)."""
    base = base or default_22k()
    d = dict(layers=6, stacks=2, residual_channels=32, gate_channels=64, skip_out_channels=32,
             upsample_scales=[3, 5], hop_size=15)
    d.update(kw)
    return base.override(**d)


def small_variants():
    """name -> hparams covering every head / input type / upsampler / conditioning the reference accepts."""
    v = {}
    v['gauss_subpixel'] = small()
    v['mol_2d_nolegacy'] = small(paper(), layers=6, stacks=3)
    v['mulaw_quantize'] = small(input_type='mulaw-quantize', quantize_channels=256, out_channels=256,
                                upsample_type='Resize')
    v['mulaw_gin_1d'] = small(input_type='mulaw', quantize_channels=256, gin_channels=8, use_speaker_embedding=True,
                              n_speakers=5, upsample_type='1D', upsample_activation='LeakyRelu')
    v['nearest_nobias_k2'] = small(upsample_type='NearestNeighbor', use_bias=False, kernel_size=2,
                                   upsample_activation=None)
    v['subpixel_no_nninit_odd'] = small(NN_init=False, residual_channels=24, gate_channels=40, skip_out_channels=20,
                                        upsample_activation='LeakyRelu')
    return v


def noise_for(hp, T, B, seed=2024):
    rng = np.random.default_rng(seed)
    if hp.input_type != 'mulaw-quantize':
        if hp.out_channels == 2:
            return rng.standard_normal((T, B, 1)).astype(np.float32)
        K = hp.out_channels // 3
        return rng.uniform(1e-5, 1 - 1e-5, (T, B, K + 1)).astype(np.float32)
    return rng.uniform(0, 1, (T, B, 1)).astype(np.float32)


def mels_for(hp, B, Tc, ragged=False):
    """Un-normalised Tacotron-range mels, one per utterance (seed 1234+u)."""
    out = []
    for u in range(B):
        rng = np.random.default_rng(1234 + u)
        n = Tc if not ragged else int(rng.integers(max(1, Tc // 2), Tc + 1))
        out.append(rng.uniform(-4, 4, (n, hp.num_mels)).astype(np.float32))
    return out


def cond_for(hp, B, Tc):
    """Normalised [B, Tc, 80] conditioning (what Synthesizer feeds the model)."""
    mels = np.stack(mels_for(hp, B, Tc))
    return ((np.clip(mels, -4, 4) + 4) / 8).astype(np.float32)


def teacher_for(hp, B, T, seed=7):
    rng = np.random.default_rng(seed)
    if hp.input_type == 'mulaw-quantize':
        return rng.integers(0, hp.quantize_channels, (B, T)).astype(np.float32)
    x = rng.uniform(-1, 1, (B, T + 4))
    x = np.stack([x[:, i:i + T] for i in range(5)]).mean(0)  # 5-tap mean
    return x.astype(np.float32)


def speaker_ids(hp, B):
    if hp.gin_channels > 0:
        return (np.arange(B) % hp.n_speakers).astype(np.int32).reshape(B, 1)
    return None


def weights_for(hp, kind='stress', seed=5339):
    return W.init_weights(hp, seed=seed, kind=kind)


def rel_close(a, b, rtol=1e-3, floor=1e-2):
    """|a-b| <= rtol * max(|b|, floor): the 1e-3 relative criterion of BASELINE.json with a floor near zero."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b) <= rtol * np.maximum(np.abs(b), floor)


def max_rel(a, b, floor=1e-2):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))
