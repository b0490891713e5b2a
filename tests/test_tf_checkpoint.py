"""TensorFlow checkpoint import without TensorFlow (wavenet_vocoder/tf_checkpoint.py): the table / proto
reader against bundles written to the same specification, and the mapping of the reference's variable names
(wavenet_vocoder/train.py:67-86: EMA shadow names) to the weight table."""
import os
import struct

import numpy as np
import pytest

import common
from wavenet_vocoder import tf_checkpoint as T
from wavenet_vocoder import weights as W


def _tf_names(hp, w):
    """Plausible reference names: model scope, layer scopes, Keras-style repeated layer names, EMA suffix; plus optimizer
    slots and the raw (non-averaged) variables that a real checkpoint also holds."""
    out = {}
    rng = np.random.default_rng(1)
    for name, a in w.items():
        parts = name.split('/')
        if parts[0].startswith('ResidualConv1DGLU_'):
            tf = 'WaveNet_model/inference/{0}/{1}_{0}/{2}'.format(parts[0], parts[1], parts[2])
        elif name == 'gc_embedding':
            tf = 'WaveNet_model/inference/gc_embedding'
        else:
            tf = 'WaveNet_model/inference/{0}/{0}_conv/{1}'.format(parts[0], parts[1])
        out[tf + '/ExponentialMovingAverage'] = a
        out[tf] = (a + rng.standard_normal(a.shape).astype(np.float32))          # raw weights: must lose against the EMA
        out[tf + '/Adam'] = np.zeros_like(a)
        out[tf + '/Adam_1'] = np.zeros_like(a)
    out['global_step'] = np.array(123456, np.int64)
    out['WaveNet_model/beta1_power'] = np.array(0.5, np.float32)
    return out


def test_varint_and_block_round_trip():
    for v in (0, 1, 127, 128, 300, 2 ** 31, 2 ** 40 + 5):
        assert T._varint(T._put_varint(v), 0) == (v, len(T._put_varint(v)))
    items = [(b'a/b/c', b'1'), (b'a/b/d', b'22'), (b'a/x', b''), (b'b', b'4444'), (b'b/long/key/name', b'5')]
    assert list(T._block_entries(T._build_block(items, restart_interval=2))) == items


@pytest.mark.parametrize('name', ['gauss_subpixel', 'mol_2d_nolegacy', 'mulaw_gin_1d'])
def test_checkpoint_round_trip_and_name_mapping(name, tmp_path):
    hp = common.small_variants()[name]
    w = common.weights_for(hp)
    prefix = str(tmp_path / 'wavenet_model.ckpt-1000')
    T.write_bundle(prefix, _tf_names(hp, w))
    header, entries = T.read_index(prefix + '.index')
    assert header['num_shards'] == 1 and 'global_step' in entries and entries['global_step']['shape'] == []
    raw = T.read_bundle(prefix)
    assert int(raw['global_step']) == 123456
    table = T.load_tf_checkpoint(prefix, hp)
    assert list(table) == list(W.weight_specs(hp))
    for n in table:
        assert table[n].dtype == np.float32 and np.array_equal(table[n], w[n]), n     # the EMA copy, bit for bit
    # the drop-in boundary resolves a checkpoint prefix
    from wavenet_vocoder.synthesizer import resolve_weights
    got = resolve_weights(prefix, hp)
    assert all(np.array_equal(got[n], w[n]) for n in w)


def test_checkpoint_errors(tmp_path):
    hp = common.small()
    w = common.weights_for(hp)
    prefix = str(tmp_path / 'm.ckpt')
    names = _tf_names(hp, w)
    missing = {k: v for k, v in names.items() if 'final_convolution_2' not in k}
    T.write_bundle(prefix, missing)
    with pytest.raises(RuntimeError, match='Failed to load checkpoint'):
        T.load_tf_checkpoint(prefix, hp)
    bad = dict(names)
    k = [k for k in bad if k.endswith('input_convolution/input_convolution_conv/kernel/ExponentialMovingAverage')][0]
    bad[k] = np.zeros((1, 2, 7), np.float32)
    T.write_bundle(prefix, bad)
    with pytest.raises(ValueError, match='shape'):
        T.load_tf_checkpoint(prefix, hp)
    with open(prefix + '.index', 'r+b') as f:
        f.seek(-8, os.SEEK_END)
        f.write(struct.pack('<Q', 1))
    with pytest.raises(ValueError, match='magic'):
        T.read_index(prefix + '.index')
    from wavenet_vocoder.synthesizer import resolve_weights
    with pytest.raises(RuntimeError, match='Failed to load checkpoint'):
        resolve_weights(str(tmp_path / 'nothing_here'), hp)


def test_logical_names():
    f = T.logical_name
    assert f('WaveNet_model/inference/ResidualConv1DGLU_13/residual_block_skip_conv_ResidualConv1DGLU_13/bias/ExponentialMovingAverage') == \
        'ResidualConv1DGLU_13/residual_block_skip_conv/bias'
    assert f('model/inference/final_convolution_2/kernel') == 'final_convolution_2/kernel'
    assert f('a/local_conditioning_upsampling_2/SubPixelConvolution/kernel/ExponentialMovingAverage') == 'local_conditioning_upsampling_2/kernel'
    assert f('global_step') is None and f('x/ResidualConv1DGLU_1/unknown/kernel') is None
