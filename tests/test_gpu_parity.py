"""Parity tests proper: the sm_100a kernels, called through the C ABI (ctypes -> libwavenet_b200.so), against the
oracles on the same seeded inputs.

Criteria (BASELINE.json / SURVEY.md 8d):
  * teacher-forced per-step distribution parameters vs the op-for-op numpy oracle: |a-b| <= 1e-3 * max(|b|, 1e-2) (fp32);
  * free-running audio vs the fixed-order C oracle under identical draws: sample-for-sample, BIT-EXACT, over the
    whole utterance (the fp32 kernels use the same fma order and the same include/wn_math.h);
  * upsampling network: bit-exact vs the C oracle, <= 1e-5 relative vs numpy.
"""
import numpy as np
import pytest

import common
from oracle import wavenet_oracle as O
from oracle.c_oracle import COracle

pytestmark = pytest.mark.gpu

ENGINES = ['generic', 'pipeline']


def _native(hp, w, engine='generic'):
    from wavenet_vocoder._native import NativeWaveNet
    return NativeWaveNet(hp, w, engine=engine)


def _upsample_gpu(net, c):
    import torch
    B, Tc, M = c.shape
    dc = torch.from_numpy(c).cuda()
    out = torch.empty((B, Tc * net.hop, M), dtype=torch.float32, device='cuda')
    net.upsample(dc, B, Tc, out, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize('utype', ['SubPixel', '2D', '1D', 'Resize', 'NearestNeighbor'])
@pytest.mark.parametrize('act', ['Relu', 'LeakyRelu', None])
def test_upsample_bit_exact_vs_c_oracle(utype, act):
    hp = common.small(upsample_type=utype, upsample_activation=act, upsample_scales=[3, 5], hop_size=15)
    w = common.weights_for(hp)
    c = common.cond_for(hp, 3, 7)
    net = _native(hp, w)
    got = _upsample_gpu(net, c)
    want = COracle(hp, w).upsample(c)
    assert got.shape == want.shape
    assert np.array_equal(got, want)
    ref = np.transpose(O.WaveNetOracle(hp, w).upsample(np.transpose(c, [0, 2, 1])), [0, 2, 1])
    assert common.max_rel(got, ref, floor=1.0) <= 1e-5


@pytest.mark.parametrize('cfg', ['default', 'paper'])
def test_upsample_shipped_configs(cfg):
    hp = common.default_22k() if cfg == 'default' else common.paper()
    w = common.weights_for(hp)
    c = common.cond_for(hp, 2, 9)
    got = _upsample_gpu(_native(hp, w), c)
    assert np.array_equal(got, COracle(hp, w).upsample(c))
    # NN-init known answer: NN_scaler * nearest-neighbour repeat (modules.py:642-654, :761-770)
    w2 = common.weights_for(hp, kind='tf-init')
    got2 = _upsample_gpu(_native(hp, w2), c)
    assert np.allclose(got2, hp.NN_scaler * np.repeat(c, hp.hop_size, axis=1), atol=1e-6)


def test_upsample_odd_band_count_uses_scalar_path():
    hp = common.small(num_mels=18, cin_channels=18)  # 18 % 4 != 0
    w = common.weights_for(hp)
    c = np.random.default_rng(0).uniform(0, 1, (2, 5, 18)).astype(np.float32)
    assert np.array_equal(_upsample_gpu(_native(hp, w), c), COracle(hp, w).upsample(c))


@pytest.mark.parametrize('name', list(common.small_variants()))
def test_generic_engine_small_variants(name):
    hp = common.small_variants()[name]
    w = common.weights_for(hp)
    net = _native(hp, w, 'generic')
    B, Tc = 3, 4
    T = Tc * hp.hop_size
    c, nz, g = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B), common.speaker_ids(hp, B)
    teacher = common.teacher_for(hp, B, T)
    orc, co = O.WaveNetOracle(hp, w), COracle(hp, w)
    # teacher forced vs numpy oracle: 1e-3 relative
    audio, raw = net.synthesize_host(c, nz, g_ids=g, teacher=teacher, want_raw=True)
    _, raw_np = orc.incremental(c=c, g=g, noise=nz, test_inputs=teacher)
    raw_np = raw_np if hp.input_type == 'mulaw-quantize' else np.transpose(raw_np, [0, 2, 1])
    assert np.all(common.rel_close(raw, raw_np)), common.max_rel(raw, raw_np)
    # ... and bit-exact vs the C oracle
    a_c, _, raw_c, _ = co.synthesize(c, g=g, noise=nz, teacher=teacher)
    assert np.array_equal(raw, raw_c) and np.array_equal(audio, a_c)
    # free running: bit-exact audio over the whole utterance
    audio, raw = net.synthesize_host(c, nz, g_ids=g, want_raw=True)
    a_c, _, raw_c, _ = co.synthesize(c, g=g, noise=nz)
    assert np.array_equal(audio, a_c) and np.array_equal(raw, raw_c)


def test_generic_engine_default_config_c1_prefix():
    """BASELINE config 1 geometry (default_22k, batch 1), first 10 frames: exact vs C oracle, 1e-3 vs numpy."""
    hp = common.default_22k()
    w = common.weights_for(hp)
    B, Tc = 1, 10
    T = Tc * hp.hop_size
    c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B)
    net = _native(hp, w, 'generic')
    audio, raw = net.synthesize_host(c, nz, want_raw=True)
    a_c, _, raw_c, _ = COracle(hp, w).synthesize(c, noise=nz)
    assert np.array_equal(audio, a_c) and np.array_equal(raw, raw_c)
    assert np.std(audio) > 1e-3 and np.mean(np.abs(audio) >= 1.0) < 0.01  # a live, unsaturated stream
    teacher = common.teacher_for(hp, B, 300)
    teacher = np.pad(teacher, [(0, 0), (0, T - 300)])
    _, raw = net.synthesize_host(c, nz, teacher=teacher, want_raw=True)
    # the numpy oracle gets 4 frames so that the first 2 frames see the same neighbours through the 3x3 upsampling convs
    _, raw_np = O.WaveNetOracle(hp, w).incremental(c=c[:, :4], noise=nz[:1100], test_inputs=teacher[:, :1100])
    assert np.all(common.rel_close(raw[:, :550], np.transpose(raw_np, [0, 2, 1])[:, :550]))


@pytest.mark.parametrize('path', ['gemm', 'serial-generic', 'serial-pipeline'])
def test_step_full_matches_oracle_step(path, monkeypatch):
    """WaveNet.step (wavenet.py:650-721): the time-parallel GEMM kernels and the loop engines in inputs mode."""
    hp = common.small()
    w = common.weights_for(hp)
    B, Tc = 2, 4
    T = Tc * hp.hop_size
    c = common.cond_for(hp, B, Tc)
    x = common.teacher_for(hp, B, T)
    from wavenet_vocoder.models import create_model
    if path != 'gemm':
        monkeypatch.setenv('WN_STEP_SERIAL', '1')
    model = create_model('WaveNet', hp, engine='auto' if path == 'gemm' else path.split('-')[1])
    model.load_weights(w)
    y = model.step(x[:, None, :], c=np.transpose(c, [0, 2, 1]))
    want = O.WaveNetOracle(hp, w).step(x[:, None, :], c=np.transpose(c, [0, 2, 1]))
    assert y.shape == (B, hp.out_channels, T)
    assert np.all(common.rel_close(y, want)), common.max_rel(y, want)


@pytest.mark.parametrize('name', ['default_22k', 'mol_2d_nolegacy'])
def test_step_gemm_equals_serial_step(name, monkeypatch):
    """The dense-GEMM step against the position-by-position engines (bit-identical to incremental synthesis) on shapes
    that do not fill the 128-position tiles; utterance starts exercise the causal zero padding of every dilation."""
    hp = common.default_22k() if name == 'default_22k' else common.small_variants()[name]
    w = common.weights_for(hp)
    B, Tc = 3, 3 if name == 'default_22k' else 9
    T = Tc * hp.hop_size
    c = common.cond_for(hp, B, Tc)
    x = common.teacher_for(hp, B, T)
    from wavenet_vocoder.models import create_model
    model = create_model('WaveNet', hp)
    model.load_weights(w)
    y = model.step(x[:, None, :], c=np.transpose(c, [0, 2, 1]))
    monkeypatch.setenv('WN_STEP_SERIAL', '1')
    y_serial = model.step(x[:, None, :], c=np.transpose(c, [0, 2, 1]))
    assert np.isfinite(y).all() and not np.array_equal(y, np.zeros_like(y))
    assert np.all(common.rel_close(y, y_serial)), common.max_rel(y, y_serial)


@pytest.mark.parametrize('name', ['gauss_subpixel', 'mulaw_gin_1d', 'subpixel_no_nninit_odd', 'mulaw_quantize'])
def test_generic_engine_cluster_sizes_bit_identical(name, monkeypatch):
    """The generic engine splits the rows of every inner product over a thread-block cluster of 1, 2, 4 or 8 CTAs per
    utterance (csrc/wn_loop_generic.cu); a row's arithmetic does not depend on who computes it, so every size returns
    the bits of the C oracle - also with channel counts that do not divide by the cluster size (24 / 40 / 20)."""
    hp = common.small_variants()[name]
    w = common.weights_for(hp)
    B, Tc = 3, 3
    T = Tc * hp.hop_size
    c, nz, g = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B), common.speaker_ids(hp, B)
    a_c, _, raw_c, _ = COracle(hp, w).synthesize(c, noise=nz, g=g)
    for C, minb in ((1, 1), (2, 1), (4, 2), (8, 1), (8, 2)):   # minb: the 128- / 64-register variant of the kernel
        monkeypatch.setenv('WN_GEN_CLUSTER', str(C))
        monkeypatch.setenv('WN_GEN_MINB', str(minb))
        net = _native(hp, w, 'generic')
        audio, raw = net.synthesize_host(c, nz, g_ids=g, want_raw=True)
        assert np.array_equal(audio, a_c) and np.array_equal(raw, raw_c), 'cluster of {}'.format(C)
        net.close()


_TENSOR_CASES = {
    'default_22k': (lambda: common.default_22k(), 3, 3),
    'r64_g128': (lambda: common.small(residual_channels=64, gate_channels=128, skip_out_channels=64), 2, 40),
    'paper_block': (lambda: common.small(common.paper(), layers=6, stacks=3, residual_channels=256, gate_channels=512,
                                         skip_out_channels=256), 2, 20),
}


@pytest.mark.parametrize('name', list(_TENSOR_CASES))
def test_step_tensor_core_path(name, monkeypatch):
    """WaveNet.step on the tensor cores (csrc/wn_step_umma.cu: tcgen05.mma, bf16 hi/lo split, three passes) against the fp32
    FFMA GEMMs (WN_STEP_FFMA=1) and the position-by-position engines: 1e-3 relative (BASELINE.json), observed ~2e-5
    (default block) / ~3e-4 (paper block, mixture head).  Column tiles of 256 and 128, K tails of 16, utterance starts inside
    a tile; the default block takes the fused layer kernel, the others the two-kernel form."""
    mk, B, Tc = _TENSOR_CASES[name]
    hp = mk()
    w = common.weights_for(hp)
    T = Tc * hp.hop_size
    c = np.transpose(common.cond_for(hp, B, Tc), [0, 2, 1])
    x = common.teacher_for(hp, B, T)[:, None, :]
    from wavenet_vocoder.models import create_model
    model = create_model('WaveNet', hp)
    model.load_weights(w)
    monkeypatch.delenv('WN_STEP_FFMA', raising=False)
    monkeypatch.delenv('WN_STEP_SERIAL', raising=False)
    y = model.step(x, c=c)
    monkeypatch.setenv('WN_STEP_FFMA', '1')
    y_f = model.step(x, c=c)
    monkeypatch.delenv('WN_STEP_FFMA')
    monkeypatch.setenv('WN_STEP_SERIAL', '1')
    y_s = model.step(x, c=c)
    monkeypatch.delenv('WN_STEP_SERIAL')
    monkeypatch.setenv('WN_STEP_UNFUSED', '1')   # two kernels per layer (z through HBM) instead of the fused layer kernel
    y_u = model.step(x, c=c)
    assert common.max_rel(y, y_u) < 1e-5, common.max_rel(y, y_u)
    assert np.isfinite(y).all()
    assert not np.array_equal(y, y_f), 'the tensor-core path did not run'
    assert np.all(common.rel_close(y, y_f)), common.max_rel(y, y_f)
    assert np.all(common.rel_close(y, y_s)), common.max_rel(y, y_s)
    assert common.max_rel(y_f, y_s) < 1e-4


def test_synthesizer_end_to_end_ragged(tmp_path):
    """Synthesizer.load / .synthesize (the drop-in boundary): ragged mels, padding, crop, wav files."""
    from wavenet_vocoder.synthesizer import Synthesizer
    hp = common.small()
    w = common.weights_for(hp)
    mels = common.mels_for(hp, 4, 6, ragged=True)
    s = Synthesizer().load(w, hp, engine='generic')
    wavs = s.synthesize(mels, None, list('abcd'), None, None, seed=11, return_wavs=True)
    c, lens = s.prepare_batch(mels)
    nz = __import__('wavenet_vocoder.models.wavenet', fromlist=['make_noise']).make_noise(
        hp, c.shape[1] * hp.hop_size, 4, np.random.default_rng(11))
    a_c, _, _, _ = COracle(hp, w).synthesize(c, noise=nz)
    for i in range(4):
        assert len(wavs[i]) == lens[i] and np.array_equal(wavs[i], a_c[i, :lens[i]])
    files = s.synthesize(mels, None, list('abcd'), str(tmp_path), None, seed=11)
    assert len(files) == 4 and all(__import__('os').path.getsize(f) > 44 for f in files)


def test_errors_surface_as_python_exceptions():
    hp = common.small()
    w = common.weights_for(hp)
    net = _native(hp, w)
    with pytest.raises(ValueError, match='local condition'):
        net.synthesize_host(np.zeros((2, 3, 7), np.float32), common.noise_for(hp, 45, 2))
    bad = dict(w)
    bad['input_convolution/kernel'] = np.zeros((1, 2, 32), np.float32)
    from wavenet_vocoder._native import NativeWaveNet
    with pytest.raises(ValueError, match='shape'):
        NativeWaveNet(hp, bad)
    info = net.info()
    assert info['receptive_field'] == 2 * 2 * (1 + 2 + 4) + 1 and info['n_params'] > 0 and info['sm_count'] >= 100
