"""The pipelined persistent engine (csrc/wn_loop_pipe.cu) through the C ABI: because it keeps the fixed fma order of
the C oracle, every comparison here is BIT-EXACT over the whole utterance, free running and teacher forced, for every
split of the network over CTAs (P), group size (NU) and utterances-per-thread (UT)."""
import os

import numpy as np
import pytest

import common
from oracle import wavenet_oracle as O
from oracle.c_oracle import COracle

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def _variants():
    v = common.small_variants()
    out = {k: v[k] for k in ('gauss_subpixel', 'mol_2d_nolegacy', 'subpixel_no_nninit_odd')}
    out['mulaw_scalar'] = common.small(input_type='mulaw', quantize_channels=256)
    return out


def _net(hp, w, **env):
    from wavenet_vocoder._native import NativeWaveNet
    old = {k: os.environ.get(k) for k in env}
    try:
        for k, val in env.items():
            os.environ[k] = str(val)
        net = NativeWaveNet(hp, w, engine='pipeline')
    finally:
        for k, val in old.items():
            if val is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = val
    assert net.info()['engine'] == 'pipeline'
    return net


def _check(net, hp, w, B, Tc):
    T = Tc * hp.hop_size
    c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B)
    teacher = common.teacher_for(hp, B, T)
    co = COracle(hp, w, ksplit=net.info()['ksplit'])   # the plan's CTAs per layer fix the summation order
    audio, raw = net.synthesize_host(c, nz, want_raw=True)
    a_c, _, raw_c, _ = co.synthesize(c, noise=nz)
    assert np.array_equal(raw[:, 0], raw_c[:, 0]), 'first step differs'
    bad = np.where(np.any(raw != raw_c, axis=(0, 2)))[0]
    assert bad.size == 0, 'free-running raw params differ first at step {}'.format(bad[:1])
    assert np.array_equal(audio, a_c)
    audio, raw = net.synthesize_host(c, nz, teacher=teacher, want_raw=True)
    a_c, _, raw_c, _ = co.synthesize(c, noise=nz, teacher=teacher)
    assert np.array_equal(raw, raw_c) and np.array_equal(audio, a_c)
    return raw


@pytest.mark.parametrize('name', list(_variants()))
@pytest.mark.parametrize('B', [1, 3, 9])
def test_pipeline_small_variants_bit_exact(name, B):
    hp = _variants()[name]
    w = common.weights_for(hp)
    net = _net(hp, w)
    raw = _check(net, hp, w, B, 4)
    if B == 3:  # and the 1e-3 relative criterion against the op-for-op numpy oracle
        T = 4 * hp.hop_size
        c, nz, teacher = common.cond_for(hp, B, 4), common.noise_for(hp, T, B), common.teacher_for(hp, B, T)
        _, raw_np = O.WaveNetOracle(hp, w).incremental(c=c, noise=nz, test_inputs=teacher)
        assert np.all(common.rel_close(raw, np.transpose(raw_np, [0, 2, 1])))


@pytest.mark.parametrize('env', [dict(WN_PIPE_P=1), dict(WN_PIPE_P=2), dict(WN_PIPE_P=3), 
                                 dict(WN_PIPE_P=7), dict(WN_PIPE_P=5, WN_PIPE_NU=4)])
def test_pipeline_plans_bit_exact(env):
    hp = common.small()
    w = common.weights_for(hp)
    _check(_net(hp, w, **env), hp, w, 7, 3)


def test_pipeline_default_config_bit_exact():
    """default_22k (L20 R128 G256 S128): 6 utterances x 4 frames, whole utterance bit-exact vs the C oracle."""
    hp = common.default_22k()
    w = common.weights_for(hp)
    net = _net(hp, w)
    info = net.info()
    assert info['pipeline_ctas'] == 20 * info['pipeline_ctas_per_layer'] and info['ksplit'] == info['pipeline_ctas_per_layer']
    _check(net, hp, w, 6, 4)


@pytest.mark.parametrize('P', [4, 5])
def test_pipeline_default_config_plans(P):
    hp = common.default_22k()
    w = common.weights_for(hp)
    net = _net(hp, w, WN_PIPE_P=P)
    assert net.info()['ksplit'] == P
    _check(net, hp, w, 5, 3)


def test_pipeline_groups_of_8_for_large_batches():
    """From WN_PIPE_NU8_MIN utterances (default 112) a launch groups the utterances by 8 instead of 4 (fewer hand-offs per
    step once the engine is throughput bound); the arithmetic per utterance is unchanged, so both stay bit-exact, on one
    handle, in either order."""
    hp = common.default_22k()
    w = common.weights_for(hp)
    net = _net(hp, w, WN_PIPE_NU8_MIN=9)
    _check(net, hp, w, 11, 1)
    assert net.info()['pipeline_group_size'] == 8
    _check(net, hp, w, 5, 2)
    assert net.info()['pipeline_group_size'] == 4
    _check(net, hp, w, 17, 1)
    assert net.info()['pipeline_group_size'] == 8
    forced = _net(hp, w, WN_PIPE_NU=8)
    _check(forced, hp, w, 3, 1)
    assert forced.info()['pipeline_group_size'] == 8


def test_generic_and_pipeline_agree_for_the_same_ksplit():
    """wn_config.ksplit pins the summation order: the two engines then return identical bits."""
    from wavenet_vocoder._native import NativeWaveNet
    hp = common.small()
    w = common.weights_for(hp)
    B, Tc = 5, 3
    T = Tc * hp.hop_size
    c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B)
    a_p, r_p = NativeWaveNet(hp, w, engine='pipeline', ksplit=3).synthesize_host(c, nz, want_raw=True)
    gen = NativeWaveNet(hp, w, engine='generic', ksplit=3)
    assert gen.info()['ksplit'] == 3
    a_g, r_g = gen.synthesize_host(c, nz, want_raw=True)
    assert np.array_equal(a_p, a_g) and np.array_equal(r_p, r_g)
    a_1 = NativeWaveNet(hp, w, engine='generic').synthesize_host(c, nz)
    assert np.allclose(a_1[:, :50], a_g[:, :50], atol=1e-4)   # a different order, the same function


def test_pipeline_repeated_calls_and_long_dilation_history():
    """Two calls on one handle (mailboxes are re-armed) and T > receptive field so every ring wraps (512*2+1 slots)."""
    hp = common.default_22k()
    w = common.weights_for(hp)
    net = _net(hp, w)
    B, Tc = 4, 17  # 4675 steps > 4093
    T = Tc * hp.hop_size
    c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B)
    a1 = net.synthesize_host(c, nz)
    a2 = net.synthesize_host(c, nz)
    a_c = COracle(hp, w, ksplit=net.info()['ksplit']).synthesize(c, noise=nz)[0]
    assert np.array_equal(a1, a_c) and np.array_equal(a2, a_c)


def test_auto_engine_picks_pipeline_and_falls_back():
    from wavenet_vocoder._native import NativeWaveNet
    hp = common.small()
    assert NativeWaveNet(hp, common.weights_for(hp), engine='auto').info()['engine'] == 'pipeline'
    hd = common.default_22k()   # the 128-channel block: the tensor-core loop engine
    assert NativeWaveNet(hd, common.weights_for(hd), engine='auto').info()['engine'] == 'tc'
    hq = common.small_variants()['mulaw_quantize']
    assert NativeWaveNet(hq, common.weights_for(hq), engine='auto').info()['engine'] == 'generic'
    hp_paper = common.paper()  # the 256-channel block: the four-CTAs-per-layer tensor-core engine; with global conditioning (which
    assert NativeWaveNet(hp_paper, common.weights_for(hp_paper), engine='auto').info()['engine'] == 'tc'   # no fast engine covers) the
    hg = hp_paper.override(gin_channels=16, n_speakers=4, use_speaker_embedding=True)   # fp32 weights do not fit shared memory either -> generic
    assert NativeWaveNet(hg, common.weights_for(hg), engine='auto').info()['engine'] == 'generic'
    with pytest.raises(ValueError, match='pipeline engine unavailable'):
        NativeWaveNet(hq, common.weights_for(hq), engine='pipeline')
