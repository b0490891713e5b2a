"""Host-side logic of the boundary: Synthesizer pre/post-processing (synthesizer.py:56-70, :100-110), wav writer,
the batching driver's file handling, CLI argument checks.  No GPU needed: the model object is stubbed."""
import os
import struct

import numpy as np
import pytest

import common
from datasets import audio
from oracle import wavenet_oracle as O
from wavenet_vocoder import synthesizer as S
from wavenet_vocoder import util


class _FakeModel:
    def __init__(self, hp):
        self.hp = hp
        self.calls = []

    def initialize(self, y, c, g, input_lengths, test_inputs=None, noise=None, want_raw=True):
        self.calls.append((c, g, noise))
        B, Tc, _ = c.shape
        T = Tc * self.hp.hop_size
        return np.tile(np.linspace(-0.5, 0.5, T, dtype=np.float32), [B, 1])


def _synth(hp):
    s = S.Synthesizer()
    s._hparams = hp
    s.synth_debug = False
    s.model = _FakeModel(hp)
    s._rng = np.random.default_rng(0)
    return s


def test_prepare_batch_matches_oracle_prologue():
    hp = common.small()
    mels = common.mels_for(hp, 4, 7, ragged=True)
    mels[0][0, 0] = 9.0  # out of range: must be clipped to max_abs_value
    s = _synth(hp)
    c, lens = s.prepare_batch(mels)
    c0, lens0 = O.prepare_local_conditions(mels, hp)
    assert np.array_equal(c, c0) and lens == lens0
    assert c.dtype == np.float32 and c.min() >= 0 and c.max() <= 1
    n = len(mels[1])
    assert np.all(c[1, n:, :] == 0.0) or n == c.shape[1]  # padding = range minimum -> 0 after rescale
    hp2 = hp.override(symmetric_mels=False, normalize_for_wavenet=False)
    c2, _ = _synth(hp2).prepare_batch(mels)
    assert c2.min() >= 0 and c2.max() <= 4


def test_synthesize_crops_and_writes_peak_normalised_wavs(tmp_path):
    hp = common.small()
    mels = common.mels_for(hp, 3, 6, ragged=True)
    s = _synth(hp)
    names = ['a', 'b', 'c']
    files = s.synthesize(mels, None, names, str(tmp_path), None, seed=3)
    assert [os.path.basename(f) for f in files] == ['wavenet-audio-%s.wav' % n for n in names]
    for f, m in zip(files, mels):
        raw = open(f, 'rb').read()
        assert raw[:4] == b'RIFF' and raw[8:16] == b'WAVEfmt '
        fmt = struct.unpack('<HHIIHH', raw[20:36])
        assert fmt == (1, 1, hp.sample_rate, hp.sample_rate * 2, 2, 16)
        data = np.frombuffer(raw[44:], dtype='<i2')
        assert len(data) == len(m) * hp.hop_size  # cropped to the true length (synthesizer.py:103)
        assert np.max(np.abs(data)) == 32767     # save_wavenet_wav peak normalisation (datasets/audio.py:19)
    noise = s.model.calls[0][2]
    assert noise.shape == (6 * hp.hop_size, 3, 1)
    wavs = s.synthesize(mels, None, names, None, None, seed=3, return_wavs=True)
    assert [len(w) for w in wavs] == [len(m) * hp.hop_size for m in mels]
    assert np.array_equal(s.model.calls[1][2], noise)  # same seed, same draws


def test_wav_int16_conversion_matches_oracle():
    rng = np.random.default_rng(0)
    x = rng.uniform(-0.3, 0.3, 1000).astype(np.float32)
    assert np.array_equal(audio.float_to_int16_peak_normalized(x), O.wav_to_int16(x))
    z = np.zeros(10, np.float32)
    assert np.array_equal(audio.float_to_int16_peak_normalized(z), np.zeros(10, np.int16))  # silence: peak floored at 0.01


def test_util_matches_oracle_mulaw():
    x = np.linspace(-1, 1, 513)
    assert np.array_equal(util.mulaw(x), O.mulaw(x))
    assert np.array_equal(util.mulaw_quantize(x), O.mulaw_quantize(x))
    assert np.array_equal(util.inv_mulaw_quantize(np.arange(256)), O.inv_mulaw_quantize(np.arange(256)))
    assert util.mulaw_quantize(0) == 127 and util.is_scalar_input('raw') and not util.is_scalar_input('mulaw-quantize')
    with pytest.raises(AssertionError):
        util.is_raw('pcm')


def test_resolve_weights(tmp_path):
    from wavenet_vocoder import weights as W
    hp = common.small()
    w = S.resolve_weights('random:7:stress', hp)
    assert np.array_equal(w['input_convolution/kernel'], W.init_weights(hp, seed=7, kind='stress')['input_convolution/kernel'])
    p = str(tmp_path / 'w.npz')
    W.save_weights(p, w)
    w2 = S.resolve_weights(p, hp)
    assert list(w2) == list(w) and all(np.array_equal(w[k], w2[k]) for k in w)
    with pytest.raises(RuntimeError, match='Failed to load checkpoint'):
        S.resolve_weights(str(tmp_path / 'model.ckpt-1000'), hp)


def test_cli_argument_checks():
    import synthesize
    with pytest.raises(ValueError, match='valid model'):
        synthesize.main(['--model', 'Foo'])
    with pytest.raises(ValueError, match='accepted modes'):
        synthesize.main(['--model', 'WaveNet', '--mode', 'x'])
    with pytest.raises(ValueError, match='GTA'):
        synthesize.main(['--model', 'WaveNet', '--GTA', 'maybe'])
    with pytest.raises(ValueError, match='world might end'):
        synthesize.main(['--model', 'Tacotron-2', '--mode', 'synthesis'])


def test_run_synthesis_batches_and_map_file(tmp_path, monkeypatch):
    """Driver: sorted *.npy listing, chunks of wavenet_synthesis_batch_size, wavs/map.txt rows mel|wav."""
    from wavenet_vocoder import synthesize as drv
    hp = common.small(wavenet_synthesis_batch_size=2)
    mel_dir = tmp_path / 'mels'
    mel_dir.mkdir()
    for i, m in enumerate(common.mels_for(hp, 5, 4, ragged=True)):
        np.save(str(mel_dir / ('mel-%d.npy' % i)), m)
    batches = []

    class FakeSynth(S.Synthesizer):
        def load(self, checkpoint_path, hparams, model_name='WaveNet', engine='auto'):
            self._hparams = hparams
            self.synth_debug = False
            self.model = _FakeModel(hparams)
            self._rng = np.random.default_rng(0)

        def synthesize(self, mels, spk, names, out_dir, log_dir, **kw):
            batches.append(list(names))
            return super().synthesize(mels, spk, names, out_dir, log_dir, **kw)

    monkeypatch.setattr(drv, 'Synthesizer', FakeSynth)

    class Args:
        model = 'WaveNet'
        mels_dir = str(mel_dir)
        speaker_id = None
        output_dir = 'out/'
    out = str(tmp_path / 'wavenet_out')
    drv.run_synthesis(Args, 'random:1', out, hp)
    assert [len(b) for b in batches] == [2, 2, 1]
    rows = open(os.path.join(out, 'wavs', 'map.txt')).read().strip().split('\n')
    assert len(rows) == 5 and all(r.split('|')[1].endswith('.wav') and os.path.exists(r.split('|')[1]) for r in rows)


def test_resolve_checkpoint_forms(tmp_path):
    """random / .npz / TF prefix / directory with a TF `checkpoint` state file / directory of weight tables
    (reference wavenet_vocoder/synthesize.py:69-78 asks tf.train.get_checkpoint_state)."""
    from wavenet_vocoder import synthesize as drv
    assert drv.resolve_checkpoint('random:3') == 'random:3'
    assert drv.resolve_checkpoint(None) is None
    npz = tmp_path / 'w.npz'
    npz.write_bytes(b'x')
    assert drv.resolve_checkpoint(str(npz)) == str(npz)
    d = tmp_path / 'logs'
    d.mkdir()
    assert drv.resolve_checkpoint(str(d)) is None
    (d / 'model.ckpt-500.index').write_bytes(b'')
    (d / 'model.ckpt-1000.index').write_bytes(b'')
    os.utime(str(d / 'model.ckpt-500.index'), (1, 1))
    assert drv.resolve_checkpoint(str(d / 'model.ckpt-500')) == str(d / 'model.ckpt-500')      # a prefix
    assert drv.resolve_checkpoint(str(d)) == str(d / 'model.ckpt-1000')                        # newest prefix of the directory
    (d / 'checkpoint').write_text('model_checkpoint_path: "model.ckpt-500"\nall_model_checkpoint_paths: "model.ckpt-500"\n')
    assert drv.resolve_checkpoint(str(d)) == str(d / 'model.ckpt-500')                         # the TF state file wins

    class Args:
        output_dir = 'o'
    with pytest.raises(RuntimeError, match='Failed to load checkpoint'):
        drv.wavenet_synthesize(Args, common.small(), str(tmp_path / 'missing'))


def test_tf_checkpoint_prefix_reaches_the_weight_table(tmp_path):
    """A TF V2 checkpoint written under EMA names, found through a `checkpoint` state file, loads into the weight table."""
    from wavenet_vocoder import synthesize as drv
    from wavenet_vocoder import tf_checkpoint as T
    hp = common.small()
    w = common.weights_for(hp)
    tensors = {}
    for name, arr in w.items():
        parts = name.split('/')
        if parts[0].startswith('ResidualConv1DGLU_'):
            role = parts[1].replace('residual_block_', '')
            tf_name = 'WaveNet_model/inference/{0}/residual_block_{1}_{0}/{2}'.format(parts[0], role, parts[2])
        else:
            tf_name = 'WaveNet_model/inference/' + name
        tensors[tf_name + '/ExponentialMovingAverage'] = arr
    prefix = str(tmp_path / 'wavenet_model.ckpt-7')
    T.write_bundle(prefix, tensors)
    (tmp_path / 'checkpoint').write_text('model_checkpoint_path: "wavenet_model.ckpt-7"\n')
    path = drv.resolve_checkpoint(str(tmp_path))
    assert path == prefix
    w2 = S.resolve_weights(path, hp)
    assert list(w2) == list(w) and all(np.array_equal(w[k], w2[k]) for k in w)


def test_weight_norm_fold_is_what_the_reference_computes_at_synthesis():
    """wavenet_weight_normalization=True (modules.py:44-181): g is re-assigned to ||v|| before every use, so the effective
    upsampling kernel is l2_normalize(v) * ||v|| (= v to a few ulp) and the incremental convolutions read v itself."""
    from wavenet_vocoder import weights as W
    hp = common.small(wavenet_weight_normalization=True)
    w = common.weights_for(hp)
    w_g = dict(w)
    w_g['ResidualConv1DGLU_0/residual_block_causal_conv/g'] = np.full(hp.gate_channels, 3.0, np.float32)   # a stale gain
    folded = W.fold_weight_norm(hp, w_g)
    assert not any(k.endswith('/g') for k in folded)
    W.check_weights(hp, folded)
    for k in w:
        if k.startswith('local_conditioning_upsampling_') and k.endswith('/kernel'):
            assert np.allclose(folded[k], w[k], rtol=3e-7, atol=0)
        else:
            assert np.array_equal(folded[k], w[k])
    assert W.fold_weight_norm(common.small(), w) is w
