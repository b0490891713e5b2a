"""Self-consistency of the numpy oracle (the reference ships no tests or vectors for this path):
two independent code paths of the reference must agree, and each upsampler must match a brute-force loop."""
import numpy as np
import pytest

import common
from oracle import wavenet_oracle as O


@pytest.mark.parametrize('name', list(common.small_variants()))
def test_ring_buffer_equals_literal_shift_queue(name):
    hp = common.small_variants()[name]
    orc = O.WaveNetOracle(hp, common.weights_for(hp))
    B, Tc = 2, 2
    T = Tc * hp.hop_size
    c, nz, g = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B), common.speaker_ids(hp, B)
    o1, r1 = orc.incremental(c=c, g=g, noise=nz, literal_queue=True)
    o2, r2 = orc.incremental(c=c, g=g, noise=nz, literal_queue=False)
    assert np.array_equal(o1, o2) and np.array_equal(r1, r2)


@pytest.mark.parametrize('name', list(common.small_variants()))
def test_incremental_teacher_forced_equals_step(name):
    """incremental() fed x as teacher == step(x): raw[t] is computed from inputs [x0, teacher[0..t-1]] and c[t]."""
    hp = common.small_variants()[name]
    orc = O.WaveNetOracle(hp, common.weights_for(hp))
    B, Tc = 2, 3
    T = Tc * hp.hop_size
    c, nz, g = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B), common.speaker_ids(hp, B)
    teacher = common.teacher_for(hp, B, T)
    _, raw = orc.incremental(c=c, g=g, noise=nz, test_inputs=teacher)
    init = orc.initial_input(B)
    if orc.scalar_input:
        x = np.concatenate([init, teacher[:, :-1]], axis=1)[:, None, :]
        raw_btO = np.transpose(raw, [0, 2, 1])
    else:
        first = np.argmax(init, axis=1)[:, None]
        ids = np.concatenate([first, teacher[:, :-1].astype(np.int64)], axis=1)
        x = np.transpose(np.eye(hp.quantize_channels, dtype=np.float32)[ids], [0, 2, 1])
        raw_btO = raw
    y = orc.step(x, c=np.transpose(c, [0, 2, 1]), g=g)
    assert np.max(np.abs(np.transpose(y, [0, 2, 1]) - raw_btO)) <= 2e-5


def _brute_upsample(hp, w, c):
    """Nested loops straight from the formulas of SURVEY.md H4a-H4d. c [B,80,W] -> [B,80,W*prod]."""
    x = c.astype(np.float64)
    kh = hp.freq_axis_kernel_size
    pt = (kh - 1) // 2
    for li, s in enumerate(hp.upsample_scales):
        K = w['local_conditioning_upsampling_%d/kernel' % (li + 1)].astype(np.float64)
        b = w['local_conditioning_upsampling_%d/bias' % (li + 1)].astype(np.float64)
        if hp.upsample_type == 'SubPixel' and not hp.NN_init:
            K = np.tile(K[..., :1], [1, 1, 1, K.shape[-1]])
        B, H, Wd = x.shape
        y = np.zeros((B, H, Wd * s))
        for h in range(H):
            for u in range(Wd * s):
                wq, f = divmod(u, s)
                acc = 0.0
                if hp.upsample_type == 'SubPixel':
                    for i in range(kh):
                        for j in range(3):
                            hh, ww = h + i - pt, wq + j - 1
                            if 0 <= hh < H and 0 <= ww < Wd:
                                acc = acc + x[:, hh, ww] * K[i, j, 0, f]
                    acc = acc + b[f]
                elif hp.upsample_type == '2D':
                    for i in range(kh):
                        hh = h + pt - i
                        if 0 <= hh < H:
                            acc = acc + x[:, hh, wq] * K[i, f, 0, 0]
                    acc = acc + b[0]
                elif hp.upsample_type == '1D':
                    acc = x[:, :, wq] @ K[0, f, h, :] + b[h]
                else:  # Resize
                    pl = (s - 1) // 2
                    for i in range(kh):
                        for j in range(s):
                            hh, uu = h + i - pt, u + j - pl
                            if 0 <= hh < H and 0 <= uu < Wd * s:
                                acc = acc + x[:, hh, uu // s] * K[i, j, 0, 0]
                    acc = acc + b[0]
                y[:, h, u] = acc
        if hp.upsample_activation == 'Relu':
            y = np.maximum(y, 0)
        elif hp.upsample_activation == 'LeakyRelu':
            y = np.where(y > 0, y, hp.leaky_alpha * y)
        x = y
    return x


@pytest.mark.parametrize('utype', ['SubPixel', '2D', '1D', 'Resize'])
@pytest.mark.parametrize('act', ['Relu', 'LeakyRelu', None])
def test_upsamplers_match_brute_force(utype, act):
    hp = common.small(upsample_type=utype, upsample_activation=act, upsample_scales=[2, 3], hop_size=6)
    w = common.weights_for(hp)
    orc = O.WaveNetOracle(hp, w)
    c = np.transpose(common.cond_for(hp, 2, 4), [0, 2, 1])
    got = orc.upsample(c)
    want = _brute_upsample(hp, w, c)
    assert got.shape == (2, 80, 24)
    assert np.max(np.abs(got - want)) < 1e-5


@pytest.mark.parametrize('utype', ['SubPixel', '2D', '1D', 'Resize', 'NearestNeighbor'])
def test_nn_init_reduces_to_scaled_nearest_neighbour(utype):
    """With the reference's NN-init kernels every upsampler is NN_scaler * repeat(c, hop) (modules.py:642-770)."""
    scales = [3, 5]  # odd scales: the even-width Resize init is not an exact repeat in the reference either
    hp = common.small(upsample_type=utype, upsample_scales=scales, hop_size=15, NN_scaler=0.3)
    w = common.weights_for(hp, kind='tf-init')
    c = np.transpose(common.cond_for(hp, 1, 5), [0, 2, 1])
    got = O.WaveNetOracle(hp, w).upsample(c)
    scale = 1.0 if utype == 'NearestNeighbor' else 0.3
    assert np.allclose(got, scale * np.repeat(c, 15, axis=2), atol=1e-6)


def test_receptive_field_causality():
    """Output at step t does not change when inputs older than receptive_field-1 steps change (wavenet.py:54-71)."""
    hp = common.small(layers=4, stacks=2)  # dilations 1,2,1,2 -> receptive field 13
    orc = O.WaveNetOracle(hp, common.weights_for(hp))
    assert orc.receptive_field == 13
    B, Tc = 1, 2
    T = Tc * hp.hop_size
    c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B)
    t1 = common.teacher_for(hp, B, T)
    t2 = t1.copy()
    t2[:, 5] += 0.5  # teacher[5] is the input of step 6; it can reach outputs of steps 6 .. 6+12
    _, r1 = orc.incremental(c=c, noise=nz, test_inputs=t1)
    _, r2 = orc.incremental(c=c, noise=nz, test_inputs=t2)
    changed = np.where(np.any(r1 != r2, axis=(0, 1)))[0]
    assert changed.min() == 6 and changed.max() == 6 + 12


def test_mulaw_known_answers():
    assert O.mulaw_quantize(0) == 127  # wavenet.py:433-445 start frame
    assert O.mulaw(0.0) == 0.0
    x = np.linspace(-1, 1, 1001)
    assert np.allclose(O.inv_mulaw(O.mulaw(x)), x, atol=1e-12)
    q = O.mulaw_quantize(x)
    assert q.min() == 0 and q.max() == 255
    assert np.max(np.abs(O.inv_mulaw_quantize(q) - x)) < 0.045  # one truncated 8-bit mu-law step near +-1
    # mu is forced to 255 whatever is passed (util.py:48)
    assert O.mulaw(0.3, 65536) == O.mulaw(0.3, 2)


def test_samplers_closed_forms():
    hp = common.small()
    orc = O.WaveNetOracle(hp, common.weights_for(hp))
    out = np.array([[0.25, np.log(0.1)], [0.9, 5.0], [0.0, -100.0]], np.float32)
    got = orc.sample_gaussian(out, np.array([2.0, 1.0, 3.0], np.float32))
    assert np.allclose(got, [0.25 + 0.2, 1.0, 3e-7], atol=1e-6)  # clip at +1; log-scale floor ln(1e-7)
    hpm = common.small(common.paper(), layers=6, stacks=3)
    om = O.WaveNetOracle(hpm, common.weights_for(hpm))
    K = 10
    outm = np.zeros((1, 3 * K), np.float32)
    outm[0, :K] = -50
    outm[0, 3] = 50  # component 3 always wins
    outm[0, K + 3] = 0.1
    outm[0, 2 * K + 3] = np.log(0.05)
    u = np.full((1, K + 1), 0.5, np.float32)
    u[0, K] = 0.75
    assert np.allclose(om.sample_mol(outm, u), 0.1 + 0.05 * np.log(3.0), atol=1e-6)
    hq = common.small(input_type='mulaw-quantize', quantize_channels=256, out_channels=256)
    oq = O.WaveNetOracle(hq, common.weights_for(hq))
    logits = np.full((2, 256), -30.0, np.float32)
    logits[0, 17] = logits[1, 200] = 10.0
    assert list(oq.sample_categorical(logits, np.array([0.5, 0.01], np.float32))) == [17, 200]
    logits = np.zeros((1, 256), np.float32)
    assert oq.sample_categorical(logits, np.array([0.5], np.float32))[0] == 128
