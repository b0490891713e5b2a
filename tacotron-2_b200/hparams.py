"""WaveNet-vocoder hyper-parameters (host side of the B200 hot path).

Mirrors the configuration surface the reference exposes through one global
``tf.contrib.training.HParams`` object (reference ``code/hparams.py:12-402``),
restricted to the keys the WaveNet synthesis path reads (SURVEY.md section 8a,
row H0).  TensorFlow is not used: ``HParams`` below is a dependency-free shim
with the same ``.parse('k=v,...')`` / ``.values()`` / attribute-access contract
(``code/synthesize.py:15``, ``code/hparams.py:404-407``).

``hparams`` (module global, like the reference) holds the reference's default
WaveNet block *as shipped* (16 kHz / hop 200, which is inconsistent with
``upsample_scales=[11, 25]``).  ``default_22k()`` returns the
configuration BASELINE.json quotes ("hparams.py default" at 22.05 kHz, hop 275).
``validate_wavenet_hparams`` raises eagerly what the reference only trips over
as a late graph assert (``wavenet.py:800``).
"""
import ast
import copy

import numpy as np


class HParams(object):
	"""Minimal stand-in for tf.contrib.training.HParams (attribute bag + parse)."""

	def __init__(self, **kwargs):
		object.__setattr__(self, '_keys', [])
		for k, v in kwargs.items():
			self.add_hparam(k, v)

	def add_hparam(self, name, value):
		if name.startswith('_') or hasattr(HParams, name):
			raise ValueError('Invalid hyper-parameter name: {}'.format(name))
		if name not in self._keys:
			self._keys.append(name)
		object.__setattr__(self, name, value)

	def set_hparam(self, name, value):
		if name not in self._keys:
			raise KeyError('Unknown hyper-parameter: {}'.format(name))
		object.__setattr__(self, name, value)

	def __setattr__(self, name, value):
		if name in self._keys:
			object.__setattr__(self, name, value)
		else:
			self.add_hparam(name, value)

	def __contains__(self, name):
		return name in self._keys

	def values(self):
		return {k: getattr(self, k) for k in self._keys}

	def get(self, name, default=None):
		return getattr(self, name) if name in self._keys else default

	def copy(self):
		return HParams(**copy.deepcopy(self.values()))

	def override(self, **kwargs):
		"""Return a copy with the given keys replaced (unknown keys are added)."""
		out = self.copy()
		for k, v in kwargs.items():
			out.add_hparam(k, v)
		return out

	@staticmethod
	def _split_top_level(s):
		parts, depth, cur = [], 0, []
		for ch in s:
			if ch in '[(':
				depth += 1
			elif ch in '])':
				depth -= 1
			if ch == ',' and depth == 0:
				parts.append(''.join(cur))
				cur = []
			else:
				cur.append(ch)
		if cur:
			parts.append(''.join(cur))
		return [p for p in (x.strip() for x in parts) if p]

	@staticmethod
	def _coerce(text, like):
		text = text.strip()
		if isinstance(like, bool):
			low = text.lower()
			if low in ('true', '1'):
				return True
			if low in ('false', '0'):
				return False
			raise ValueError('Could not parse boolean: {}'.format(text))
		if isinstance(like, int):
			return int(text)
		if isinstance(like, float):
			return float(text)
		if isinstance(like, (list, tuple)):
			val = ast.literal_eval(text)
			if not isinstance(val, (list, tuple)):
				val = [val]
			return list(val)
		if isinstance(like, str):
			if len(text) >= 2 and text[0] == text[-1] and text[0] in '\'"':
				return text[1:-1]
			return text
		# None-valued default: best effort literal, else keep the string
		if text in ('None', 'none'):
			return None
		try:
			return ast.literal_eval(text)
		except (ValueError, SyntaxError):
			return text

	def parse(self, values):
		"""Override from a 'name=value,name2=[1,2]' string; returns self (like TF)."""
		if not values:
			return self
		for item in self._split_top_level(values):
			if '=' not in item:
				raise ValueError('Could not parse hparam override: {!r}'.format(item))
			name, text = item.split('=', 1)
			name = name.strip()
			if name not in self._keys:
				raise ValueError('Unknown hyper-parameter: {}'.format(name))
			self.set_hparam(name, self._coerce(text, getattr(self, name)))
		return self


def _wavenet_block_default():
	"""Key/value pairs of the reference default WaveNet block and the audio/mel keys the
	synthesis path reads (code/hparams.py:45-46,71,88-91,103-104,117-124,207-253,322-366,397-400)."""
	return dict(
		#audio / mel keys consumed by the WaveNet path
		num_mels=80,
		hop_size=200,
		win_size=800,
		sample_rate=16000,
		frame_shift_ms=None,
		symmetric_mels=True,
		max_abs_value=4.,
		normalize_for_wavenet=True,
		clip_for_wavenet=True,
		wavenet_pad_sides=1,
		preemphasize=True,
		preemphasis=0.97,
		#model type
		input_type='raw',
		quantize_channels=2 ** 16,
		use_bias=True,
		legacy=True,
		residual_legacy=True,
		log_scale_min=float(np.log(1e-14)),
		log_scale_min_gauss=float(np.log(1e-7)),
		cdf_loss=False,
		#model size
		out_channels=2,
		layers=20,
		stacks=2,
		residual_channels=128,
		gate_channels=256,
		skip_out_channels=128,
		kernel_size=3,
		#local conditioning / upsampling
		cin_channels=80,
		upsample_type='SubPixel',
		upsample_activation='Relu',
		upsample_scales=[11, 25],
		freq_axis_kernel_size=3,
		leaky_alpha=0.4,
		NN_init=True,
		NN_scaler=0.3,
		#global conditioning
		gin_channels=-1,
		use_speaker_embedding=False,
		n_speakers=5,
		#run-time knobs of the synthesis path
		wavenet_random_seed=5339,
		wavenet_data_random_state=1234,
		wavenet_num_gpus=1,
		split_on_cpu=True,
		wavenet_swap_with_cpu=False,
		wavenet_synthesis_batch_size=10 * 2,
		wavenet_ema_decay=0.9999,
		wavenet_weight_normalization=False,
		wavenet_init_scale=1.,
		wavenet_dropout=0.05,
		wavenet_natural_eval=False,
		train_with_GTA=True,
		wavenet_synth_debug=False,
		wavenet_debug_wavs=['training_data/audio/audio-LJ001-0008.npy'],
		wavenet_debug_mels=['training_data/mels/mel-LJ001-0008.npy'],
		#B200 build additions (not in the reference): see DESIGN.md
		wavenet_compute_dtype='fp32',	#'fp32' is the parity path
		wavenet_synthesis_seed=None,	#None = fresh entropy (the reference's synthesis draws are unseeded)
	)


#The reference's global object (code/hparams.py:12).
hparams = HParams(**_wavenet_block_default())


def default_22k():
	"""'hparams.py default' of BASELINE.json: the default WaveNet block at 22.05 kHz / hop 275
	(the in-line alternatives of code/hparams.py:88-90)."""
	return hparams.copy().override(sample_rate=22050, hop_size=275, win_size=1100)


def hparams_debug_string(hp=None):
	values = (hp or hparams).values()
	hp_lines = ['  %s: %s' % (name, values[name]) for name in sorted(values) if name != 'sentences']
	return 'Hyperparameters:\n' + '\n'.join(hp_lines)


def get_hop_size(hp):
	"""Same rule as code/datasets/audio.py:54-59."""
	hop_size = hp.hop_size
	if hop_size is None:
		assert hp.frame_shift_ms is not None
		hop_size = int(hp.frame_shift_ms / 1000 * hp.sample_rate)
	return hop_size


_INPUT_TYPES = ('raw', 'mulaw', 'mulaw-quantize')
_UPSAMPLE_TYPES = ('1D', '2D', 'Resize', 'SubPixel', 'NearestNeighbor')
_UPSAMPLE_ACTIVATIONS = ('Relu', 'LeakyRelu', None)


def validate_wavenet_hparams(hp):
	"""Eager versions of the reference's constructor/graph asserts:
	models/__init__.py:6-9, wavenet.py:93-97, wavenet.py:800 (upsampled length), gaussian.py:41,
	mixture.py:83.  Raises the same exception types the reference would."""
	if hp.input_type not in _INPUT_TYPES:
		raise AssertionError('input_type must be one of {}'.format(_INPUT_TYPES))
	if hp.input_type == 'mulaw-quantize' and hp.out_channels != hp.quantize_channels:
		raise RuntimeError(
			"out_channels must equal to quantize_chennels if input_type is 'mulaw-quantize'")
	if hp.cin_channels > 0 and hp.num_mels != hp.cin_channels:
		raise AssertionError('num_mels ({}) must equal cin_channels ({})'.format(hp.num_mels, hp.cin_channels))
	if hp.layers % hp.stacks != 0:
		raise AssertionError('layers ({}) must be a multiple of stacks ({})'.format(hp.layers, hp.stacks))
	if hp.gate_channels % 2 != 0:
		raise AssertionError('gate_channels must be even (split into tanh / sigmoid halves)')
	if hp.input_type != 'mulaw-quantize':
		if hp.out_channels != 2 and hp.out_channels % 3 != 0:
			raise AssertionError('scalar input needs out_channels == 2 (Gaussian) or a multiple of 3 (MoL)')
	if hp.cin_channels > 0:
		if hp.upsample_type not in _UPSAMPLE_TYPES:
			raise AssertionError('upsample_type must be one of {}'.format(_UPSAMPLE_TYPES))
		if hp.upsample_activation not in _UPSAMPLE_ACTIVATIONS:
			raise AssertionError('upsample_activation must be one of {}'.format(_UPSAMPLE_ACTIVATIONS))
		hop = get_hop_size(hp)
		if hp.upsample_type != 'NearestNeighbor' and int(np.prod(hp.upsample_scales)) != hop:
			raise ValueError(
				'prod(upsample_scales)={} must equal hop_size={} (the reference fails late on the '
				'upsampled-length assert, wavenet.py:800); e.g. pass '
				'--hparams "sample_rate=22050,hop_size=275"'.format(int(np.prod(hp.upsample_scales)), hop))
	if hp.gin_channels > 0 and hp.use_speaker_embedding and hp.n_speakers is None:
		raise AssertionError('n_speakers must be set when use_speaker_embedding is on')
	return hp
