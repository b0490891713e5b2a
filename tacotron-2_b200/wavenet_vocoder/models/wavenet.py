"""WaveNet vocoder model object - host-side mirror of the reference's ``WaveNet`` class
(``code/wavenet_vocoder/models/wavenet.py:86-923``) for the synthesis path only.

Same constructor checks, attribute names (``receptive_field``, ``tower_y_hat``, ``tower_y_hat_eval``,
``tower_synth_upsampled_local_features``) and method names (``initialize``, ``incremental``, ``step``) as the
reference, but eager: the reference builds a TF graph over placeholders and runs it later through a session, here
the calls take numpy arrays and run the sm_100a kernels of libwavenet_b200.so immediately.  Training members
(``add_loss``, ``add_optimizer``, EMA) are out of scope (disabled in this fork).

The random draws the reference takes inside its ``tf.while_loop`` (gaussian.py:49, mixture.py:91,104, wavenet.py:865)
are an explicit ``noise`` input here so the output is a pure function of (weights, c, noise, teacher).
"""
import numpy as np

from hparams import get_hop_size, validate_wavenet_hparams
from infolog import log
from wavenet_vocoder import util
from wavenet_vocoder import weights as wn_weights
from wavenet_vocoder._native import NativeWaveNet
from wavenet_vocoder.util import is_mulaw, is_mulaw_quantize, is_scalar_input

receptive_field_size = wn_weights.receptive_field_size


def make_noise(hp, T, B, rng):
	"""Draws for T steps of B utterances in the layout the kernels consume, [T, B, NZ]."""
	if is_scalar_input(hp.input_type):
		if hp.out_channels == 2:
			return rng.standard_normal((T, B, 1), dtype=np.float32)
		K = hp.out_channels // 3
		return rng.uniform(1e-5, 1. - 1e-5, (T, B, K + 1)).astype(np.float32)
	return rng.uniform(0., 1., (T, B, 1)).astype(np.float32)


class _LazyTower(object):
	"""A one-tower list whose element is produced on first access."""
	def __init__(self, make):
		self._make, self._val, self._done = make, None, False
	def __len__(self):
		return 1
	def __getitem__(self, i):
		if i not in (0, -1):
			raise IndexError(i)
		if not self._done:
			self._val, self._done = self._make(), True
		return self._val
	def __iter__(self):
		yield self[0]


class WaveNet(object):
	"""Tacotron-2 Wavenet Vocoder model (synthesis path)."""

	def __init__(self, hparams, init=False, engine='auto'):
		self._hparams = hparams
		validate_wavenet_hparams(hparams)	#wavenet.py:93-97 + models/__init__.py:6-9 + the late assert of :800, eagerly
		self.scalar_input = is_scalar_input(hparams.input_type)
		self.receptive_field = receptive_field_size(hparams.layers, hparams.stacks, hparams.kernel_size)
		self._engine = engine
		self._native = None
		self._rng = np.random.default_rng(hparams.get('wavenet_synthesis_seed'))
		self.tower_y_hat = []
		self.tower_y_hat_eval = []
		self.tower_synth_upsampled_local_features = []
		self.variables = None

	#Sanity check functions (wavenet.py:637-647)
	def has_speaker_embedding(self):
		return self._hparams.gin_channels > 0 and self._hparams.use_speaker_embedding

	def local_conditioning_enabled(self):
		return self._hparams.cin_channels > 0

	def global_conditioning_enabled(self):
		return self._hparams.gin_channels > 0

	# ---- weights ("restore") ------------------------------------------------------------------------
	def load_weights(self, weights):
		"""weights: dict name -> array in TF layouts (wavenet_vocoder/weights.py).  Uploads to the current CUDA device."""
		weights = wn_weights.fold_weight_norm(self._hparams, weights)	#modules.py:44-181 at synthesis time (row H4e)
		wn_weights.check_weights(self._hparams, weights)
		if self._native is not None:
			self._native.close()
		self._native = NativeWaveNet(self._hparams, weights, engine=self._engine)
		self.variables = list(weights.keys())
		info = self._native.info()
		log('Initialized Wavenet model. Dimensions (? = dynamic shape): ')
		log('  Receptive Field:           ({} samples / {:.1f} ms)'.format(self.receptive_field,
			self.receptive_field / self._hparams.sample_rate * 1000.))
		log('  WaveNet Parameters:        {:.3f} Million.'.format(info['n_params'] / 1000000))
		log('  Engine:                    {} ({} SMs)'.format(info['engine'], info['sm_count']))
		return self

	@property
	def native(self):
		if self._native is None:
			raise RuntimeError('WaveNet weights not loaded: call load_weights() (Synthesizer.load does it)')
		return self._native

	def set_seed(self, seed):
		self._rng = np.random.default_rng(seed)

	# ---- wavenet.py:218-473, synthesis branch only ---------------------------------------------------
	def initialize(self, y, c, g, input_lengths, x=None, synthesis_length=None, test_inputs=None, split_infos=None,
			noise=None, want_raw=True):
		"""Synthesis branch of the reference's initialize() (y is None, x is None): c [B, Tc, num_mels] normalised
		local conditions, g [B, 1] speaker ids or None, test_inputs [B, T, 1] teacher-forcing waveform or None.
		Fills tower_y_hat ([B, T] waveform), tower_y_hat_eval (raw per-step parameters) and
		tower_synth_upsampled_local_features like the reference (one tower: utterances shard across processes)."""
		hp = self._hparams
		if y is not None or x is not None:
			raise NotImplementedError('training / evaluation towers are out of scope (disabled in the reference fork)')
		if c is None:
			#wavenet.py:409-411: without local conditioning the length comes from synthesis_length
			if synthesis_length is None:
				raise AssertionError('synthesis_length is required when c is None (wavenet.py:410)')
			if hp.cin_channels > 0:
				raise ValueError('this model was built with local conditioning (cin_channels > 0): c is required')
			T = int(synthesis_length)
			if noise is not None:
				B = np.asarray(noise).shape[1]
			elif g is not None:
				B = np.asarray(g).shape[0]
			elif test_inputs is not None:
				B = np.asarray(test_inputs).shape[0]
			else:
				B = 1
		else:
			c = np.asarray(c, np.float32)
			if c.ndim != 3:
				raise ValueError('Expected 3 dimension shape [batch_size(1), time_length, {}] for local condition features '
					'but found {}'.format(hp.cin_channels, c.shape))	#wavenet.py:414-417
			B, Tc, _ = c.shape
			T = Tc * get_hop_size(hp)	#wavenet.py:419-423
		teacher = None
		if test_inputs is not None:
			teacher = np.asarray(test_inputs, np.float32).reshape(B, -1)
			if teacher.shape[1] < T:
				teacher = np.pad(teacher, [(0, 0), (0, T - teacher.shape[1])])
			teacher = teacher[:, :T]
		if noise is None:
			noise = make_noise(hp, T, B, self._rng)
		res = self.native.synthesize_host(c, noise, g_ids=g, teacher=teacher, want_raw=want_raw)
		audio, raw = res if want_raw else (res, None)
		self.tower_y_hat = [audio]
		if raw is not None:
			#scalar input: [B, O, T]; mulaw-quantize: [B, T, C]  (wavenet.py:904-908)
			self.tower_y_hat_eval = [raw if is_mulaw_quantize(hp.input_type) else np.transpose(raw, [0, 2, 1])]
		#wavenet.py:458-459: [B, cin, T], 80 x the audio and only ever plotted -> computed when somebody reads it
		self.tower_synth_upsampled_local_features = _LazyTower((lambda: self.upsample(c)) if c is not None else (lambda: None))
		return audio

	def upsample(self, c):
		"""The conditioning network alone (wavenet.py:781-803): c [B, Tc, 80] -> [B, 80, T] like
		upsampled_local_features."""
		import torch
		c = np.ascontiguousarray(c, np.float32)
		B, Tc, M = c.shape
		T = Tc * get_hop_size(self._hparams)
		dc = torch.from_numpy(c).cuda()
		out = torch.empty((B, T, M), dtype=torch.float32, device='cuda')
		self.native.upsample(dc, B, Tc, out, stream=torch.cuda.current_stream().cuda_stream)
		torch.cuda.synchronize()
		return out.permute(0, 2, 1).contiguous().cpu().numpy()

	def incremental(self, initial_input=None, c=None, g=None, time_length=100, test_inputs=None, softmax=False,
			quantize=True, log_scale_min=None, log_scale_min_gauss=None, noise=None):
		"""Reference signature (wavenet.py:724).  c is [B, cin, Tc] (channels first, as the reference passes it) or None with
		time_length steps (models built with cin_channels <= 0); returns the generated samples [B, C, T] (scalar input: C = 1,
		the value fed back, i.e. still mu-law companded for input_type 'mulaw'; mulaw-quantize: one-hot of the sampled class).
		The kernels implement what initialize() asks of this function (wavenet.py:447-448): start from the silence frame,
		softmax=False, quantize=True, scale floors from hparams.  Other argument values are rejected, not ignored."""
		hp = self._hparams
		if softmax and not self.scalar_input:
			raise NotImplementedError('softmax=True re-normalises the logits before tf.multinomial (wavenet.py:861-865); '
				'initialize() passes softmax=False and so does this build')
		if not quantize and not self.scalar_input:
			raise NotImplementedError('quantize=False (feeding back the un-sampled distribution) is not built')
		for name, val, ref in (('log_scale_min', log_scale_min, hp.log_scale_min),
				('log_scale_min_gauss', log_scale_min_gauss, hp.log_scale_min_gauss)):
			if val is not None and abs(float(val) - float(ref)) > 1e-6 * max(1.0, abs(float(ref))):
				raise ValueError('{}={} differs from hparams.{}={}: the scale floors are fixed when the model is created'.format(
					name, val, name, ref))
		if initial_input is not None:
			ii = np.asarray(initial_input, np.float32)
			if self.scalar_input:
				ok = np.all(ii == 0.0)	#raw 0.0 / mulaw(0.0) = 0.0 (wavenet.py:433-445)
			else:
				ok = np.all(np.argmax(ii.reshape(ii.shape[0], -1), axis=-1) == util.mulaw_quantize(0, hp.quantize_channels))
			if not ok:
				raise NotImplementedError('only the reference silence start frame is supported as initial_input (wavenet.py:433-445)')
		if c is None:
			audio = self.initialize(None, None, g, None, synthesis_length=time_length, test_inputs=test_inputs, noise=noise)
		else:
			c_btc = np.ascontiguousarray(np.transpose(np.asarray(c, np.float32), [0, 2, 1]))
			audio = self.initialize(None, c_btc, g, None, test_inputs=test_inputs, noise=noise)
		B = audio.shape[0]
		if is_mulaw_quantize(hp.input_type):
			#the kernel decoded class k to inv_mulaw_quantize(k): recover k by nearest table entry (re-quantising the decoded
			#value is NOT the identity: truncation moves 77 of the 256 classes by one)
			table = util.inv_mulaw_quantize(np.arange(hp.quantize_channels), hp.quantize_channels).astype(np.float64)
			pos = np.clip(np.searchsorted(table, audio.astype(np.float64)), 1, len(table) - 1)
			cls = np.where(np.abs(audio - table[pos - 1]) <= np.abs(audio - table[pos]), pos - 1, pos)
			return np.transpose(np.eye(hp.quantize_channels, dtype=np.float32)[cls], [0, 2, 1])
		if is_mulaw(hp.input_type):
			return util.mulaw(audio, hp.quantize_channels).astype(np.float32).reshape(B, 1, -1)
		return audio.reshape(B, 1, -1)

	def step(self, x, c=None, g=None, softmax=False):
		"""Teacher-forced full-sequence pass (wavenet.py:650-721), dropout off.  x [B, C_in, T] (scalar: C_in = 1;
		mulaw-quantize: one-hot), c [B, cin, Tc] -> [B, out_channels, T]."""
		import torch
		hp = self._hparams
		x = np.asarray(x, np.float32)
		B, Cin, T = x.shape
		xin = x[:, 0, :] if self.scalar_input else np.argmax(x, axis=1).astype(np.float32)
		if c is None:
			if hp.cin_channels > 0:
				raise ValueError('this model was built with local conditioning (cin_channels > 0): c is required')
			Tc, dc = T, None
		else:
			c_btc = np.ascontiguousarray(np.transpose(np.asarray(c, np.float32), [0, 2, 1]))
			Tc = c_btc.shape[1]
			if Tc * get_hop_size(hp) != T:
				raise ValueError('upsampled conditioning length {} != input length {} (wavenet.py:697)'.format(
					Tc * get_hop_size(hp), T))
			dc = torch.from_numpy(c_btc).cuda()
		dx = torch.from_numpy(np.ascontiguousarray(xin)).cuda()
		dg = None if g is None else torch.from_numpy(np.asarray(g, np.int32).reshape(B)).cuda()
		raw = torch.empty((B, T, hp.out_channels), dtype=torch.float32, device='cuda')
		self.native.step_full(dx, dc, B, Tc, raw, g_ids=dg, stream=torch.cuda.current_stream().cuda_stream)
		torch.cuda.synchronize()
		out = raw.permute(0, 2, 1).contiguous().cpu().numpy()
		if softmax:
			e = np.exp(out - out.max(axis=1, keepdims=True))
			out = e / e.sum(axis=1, keepdims=True)
		return out

	def eval_loss(self, y, c=None, g=None, reduce=True):
		"""Likelihood of a waveform under the model, the reference's eval-mode loss (wavenet.py:497-506 with
		modules.py:800-836): one teacher-forced pass over y [B, T] (in [-1, 1]; mulaw input types: already companded)
		on the GPU (step()), then the negative log-likelihood of y[t+1] under the parameters predicted at t, averaged
		over all predicted samples (reduce=True) or returned per sample [B, T-1].  Gaussian head:
		gaussian_maximum_likelihood_estimation_loss (CDF form, as hparams.cdf_loss), mixture head:
		discretized_mix_logistic_loss, categorical head: cross entropy."""
		from wavenet_vocoder.models.gaussian import gaussian_maximum_likelihood_estimation_loss
		from wavenet_vocoder.models.mixture import discretized_mix_logistic_loss, log_prob_from_logits
		hp = self._hparams
		y = np.asarray(y, np.float32)
		B, T = y.shape
		if self.scalar_input:
			x = y[:, None, :]
		else:
			x = np.eye(hp.quantize_channels, dtype=np.float32)[y.astype(np.int64)].transpose(0, 2, 1)
		y_hat = self.step(x, c=c, g=g)[:, :, :-1]          # parameters for samples 1 .. T-1
		tgt = y[:, 1:, None]
		if not self.scalar_input:
			logp = log_prob_from_logits(np.transpose(y_hat, [0, 2, 1]))
			nll = -np.take_along_axis(logp, tgt.astype(np.int64), axis=-1)
		elif hp.out_channels == 2:
			nll = gaussian_maximum_likelihood_estimation_loss(y_hat, tgt, hp.log_scale_min_gauss, hp.quantize_channels,
				use_cdf=getattr(hp, 'cdf_loss', False), reduce=False)
		else:
			nll = discretized_mix_logistic_loss(y_hat, tgt, num_classes=hp.quantize_channels,
				log_scale_min=hp.log_scale_min, reduce=False)
		return float(np.mean(nll)) if reduce else nll[..., 0]
