"""Generates tests/golden/ref_*.npz / ref_hparams.json by EXECUTING THE REFERENCE'S OWN PYTHON from /root/reference/code over
the numpy `tf` stand-in of tests/golden/tf_shim.py (TensorFlow cannot be installed here).  Run in the build container:

    python tests/golden/make_ref_golden.py

The fixtures travel to the GPU box; /root/reference does not.  tests/test_oracle_vs_reference_shim.py compares
oracle/wavenet_oracle.py (and the repo's host helpers) with them.  See tf_shim.py for what is executed from the reference
and what is restated there from TF op semantics.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = '/root/reference/code'
sys.path[:0] = [HERE, os.path.join(ROOT, 'tests'), ROOT, os.path.join(ROOT, 'tacotron-2_b200')]

import tf_shim	# noqa: E402

np.int = int	# the reference predates numpy 1.24 (util.py uses np.int)
_expand_dims = np.expand_dims


def _lenient_expand_dims(a, axis):
	"""numpy < 1.18 clipped an axis beyond ndim instead of raising; SubPixelConvolution._init_kernel (modules.py:652) relies on it
	(SURVEY.md appendix A).  Only the initialiser value is affected, which a restored checkpoint overrides anyway."""
	a = np.asarray(a)
	if isinstance(axis, int) and axis > a.ndim:
		axis = a.ndim
	return _expand_dims(a, axis)


np.expand_dims = _lenient_expand_dims
tf = tf_shim.install()

import common	# noqa: E402  (repo test helpers: seeded hparams variants, weights, inputs)


def _import_reference():
	"""The reference package shadows the repo's `wavenet_vocoder`, `datasets`, `hparams`: import it under its own sys.path
	and hand the modules back, then restore."""
	saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k.split('.')[0] in ('wavenet_vocoder', 'datasets', 'hparams', 'infolog', 'paper_hparams')}
	sys.path.insert(0, REF)
	try:
		from wavenet_vocoder import util as r_util
		from wavenet_vocoder.models import wavenet as r_wavenet, gaussian as r_gauss, mixture as r_mix, modules as r_modules
		from wavenet_vocoder.models import create_model as r_create
		from datasets import audio as r_audio
		import hparams as r_hparams
		import paper_hparams as r_paper
		mods = dict(util=r_util, wavenet=r_wavenet, gaussian=r_gauss, mixture=r_mix, modules=r_modules, create_model=r_create,
			audio=r_audio, hparams=r_hparams, paper=r_paper)
	finally:
		sys.path.remove(REF)
		for k in [k for k in list(sys.modules) if k.split('.')[0] in ('wavenet_vocoder', 'datasets', 'hparams', 'infolog', 'paper_hparams')]:
			sys.modules.pop(k)
		sys.modules.update(saved)
	return mods


R = _import_reference()


def util_fixture():
	x = np.linspace(-1, 1, 401).astype(np.float32)
	q = np.arange(256)
	out = dict(x=x, q=q,
		mulaw=R['util'].mulaw(x, 256), inv_mulaw=R['util'].inv_mulaw(x, 256),
		mulaw_quantize=R['util'].mulaw_quantize(x, 256), inv_mulaw_quantize=R['util'].inv_mulaw_quantize(q, 256),
		mulaw_quantize_0=np.asarray(R['util'].mulaw_quantize(0, 256)), mulaw_0=np.asarray(R['util'].mulaw(0.0, 256)))
	# feeder._interp (feeder.py:426-428) is one line; the module itself imports sklearn/keras feeders, so it is exec'd from source
	src = open(os.path.join(REF, 'wavenet_vocoder', 'feeder.py')).read()
	start = src.index('def _interp(')
	ns = {}
	exec(src[start:], ns)
	feats = np.random.default_rng(3).uniform(-4, 4, (7, 80)).astype(np.float32)
	out['interp_in'] = feats
	out['interp_out'] = ns['_interp'](feats, (-4.0, 4.0))
	out['interp_out_asym'] = ns['_interp'](np.abs(feats), (0.0, 4.0))
	# save_wavenet_wav (datasets/audio.py:17-20): peak normalisation to int16
	from scipy.io import wavfile
	wav = (np.random.default_rng(4).uniform(-0.3, 0.3, 2000)).astype(np.float32)
	path = '/tmp/_ref_wav.wav'
	R['audio'].save_wavenet_wav(wav.copy(), path, 22050, True, 0.97)
	out['wav_in'] = wav
	out['wav_int16'] = wavfile.read(path)[1]
	out['receptive_field'] = np.asarray([R['wavenet'].receptive_field_size(20, 2, 3), R['wavenet'].receptive_field_size(24, 4, 3)])
	np.savez_compressed(os.path.join(HERE, 'ref_util.npz'), **out)


WAVENET_KEYS = ['input_type', 'quantize_channels', 'log_scale_min', 'log_scale_min_gauss', 'out_channels', 'layers', 'stacks',
	'residual_channels', 'gate_channels', 'skip_out_channels', 'kernel_size', 'cin_channels', 'upsample_type', 'upsample_activation',
	'upsample_scales', 'freq_axis_kernel_size', 'leaky_alpha', 'NN_init', 'NN_scaler', 'gin_channels', 'use_speaker_embedding',
	'n_speakers', 'use_bias', 'legacy', 'residual_legacy', 'wavenet_weight_normalization', 'wavenet_init_scale', 'wavenet_dropout',
	'wavenet_synthesis_batch_size', 'wavenet_swap_with_cpu', 'wavenet_ema_decay', 'num_mels', 'max_abs_value', 'symmetric_mels',
	'clip_for_wavenet', 'normalize_for_wavenet', 'sample_rate', 'hop_size', 'wavenet_num_gpus']


def hparams_fixture():
	out = {}
	for tag, mod in (('default', R['hparams']), ('paper', R['paper'])):
		v = mod.hparams.values()
		out[tag] = {k: v[k] for k in WAVENET_KEYS if k in v}
	json.dump(out, open(os.path.join(HERE, 'ref_hparams.json'), 'w'), indent=1, sort_keys=True, default=float)


def sampler_fixture():
	rng = np.random.default_rng(21)
	B = 6
	y = rng.normal(0, 1, (B, 2, 1)).astype(np.float32)
	y[:, 1] = rng.uniform(-9, -1, (B, 1))
	eps = rng.standard_normal(B).astype(np.float32)
	tf_shim.feed_noise(normal=[eps])
	g = R['gaussian'].sample_from_gaussian(y, log_scale_min_gauss=np.float32(-7.0))
	K = 10
	ym = rng.normal(0, 1, (B, 3 * K, 1)).astype(np.float32)
	ym[:, 2 * K:] = rng.uniform(-9, -1, (B, K, 1))
	u1 = rng.uniform(1e-5, 1 - 1e-5, (B, 1, K)).astype(np.float32)
	u2 = rng.uniform(1e-5, 1 - 1e-5, (B, 1)).astype(np.float32)
	tf_shim.feed_noise(uniform_k=[u1], uniform_1=[u2])
	m = R['mixture'].sample_from_discretized_mix_logistic(ym, log_scale_min=np.float32(-7.0))
	np.savez_compressed(os.path.join(HERE, 'ref_samplers.npz'), y=y, eps=eps, gauss=g, ym=ym, u1=u1, u2=u2, mol=m)


def run_reference_model(hp, w, c, noise, teacher=None, g=None):
	"""Synthesizer.load + one session.run, eagerly: create_model -> initialize (synthesis branch) (synthesizer.py:27-30, :97)."""
	tf_shim.install_weights(w)
	T = noise.shape[0]
	if hp.input_type != 'mulaw-quantize':
		if hp.out_channels == 2:
			tf_shim.feed_noise(normal=[noise[t, :, 0] for t in range(T)])
		else:
			K = hp.out_channels // 3
			tf_shim.feed_noise(uniform_k=[noise[t, :, None, :K] for t in range(T)], uniform_1=[noise[t, :, K:K + 1] for t in range(T)])
	else:
		tf_shim.feed_noise(categorical=[noise[t, :, 0] for t in range(T)])
	model = R['create_model']('WaveNet', hp)
	test_inputs = None
	if teacher is not None:
		test_inputs = teacher[:, :, None].astype(np.float32)	#[B, T, 1] (synthesizer.py:23: targets placeholder)
	model.initialize(None, c, g, input_lengths=None, synthesis_length=None, test_inputs=test_inputs)
	out = dict(y_hat=np.asarray(model.tower_y_hat[0]), raw=np.asarray(model.tower_y_hat_eval[0]))
	if hp.cin_channels > 0:
		out['c_up'] = np.asarray(model.tower_synth_upsampled_local_features[0])
	return model, out


def model_fixtures():
	variants = common.small_variants()
	for name, hp in variants.items():
		hp = hp.override(wavenet_dropout=0.0)
		if hp.gin_channels > 0:
			# the reference's own synthesis branch rejects global conditioning: `assert tower_g[i].shape == (batch_size, 1)`
			# (wavenet.py:430) compares a TensorShape with a tuple holding a Tensor, which is False under TF-1 as under this
			# stand-in.  The variant is run without it; the global-conditioning arithmetic stays restated-only in the oracle.
			hp = hp.override(gin_channels=-1, use_speaker_embedding=False)
			name = name.replace('_gin', '')
		w = common.weights_for(hp)
		B, Tc = 2, 3
		T = Tc * hp.hop_size
		c = common.cond_for(hp, B, Tc)
		noise = common.noise_for(hp, T, B)
		g = common.speaker_ids(hp, B)
		teacher = common.teacher_for(hp, B, T)
		_, free = run_reference_model(hp, w, c, noise, None, g)
		model, forced = run_reference_model(hp, w, c, noise, teacher, g)
		out = dict(c=c, noise=noise, teacher=teacher, free_y_hat=free['y_hat'], free_raw=free['raw'], forced_y_hat=forced['y_hat'],
			forced_raw=forced['raw'])
		if 'c_up' in free:
			out['c_up'] = free['c_up']
		if g is not None:
			out['g'] = g
		# WaveNet.step (wavenet.py:650-721) on the same weights: x [B, C_in, T] channels first, c [B, cin, Tc]
		if hp.input_type != 'mulaw-quantize':
			x = teacher[:, None, :].astype(np.float32)
		else:
			x = np.transpose(np.eye(hp.quantize_channels, dtype=np.float32)[teacher.astype(np.int64)], [0, 2, 1])
		model.set_mode(True)	#step() is the training-graph path: channels-first convolutions (modules.py:256)
		tf_shim.install_weights(w)
		model2 = R['create_model']('WaveNet', hp)
		model2.set_mode(True)
		out['step_raw'] = np.asarray(model2.step(x, np.transpose(c, [0, 2, 1]), g, softmax=False))
		np.savez_compressed(os.path.join(HERE, 'ref_model_{}.npz'.format(name)), **out)
		print('ref_model_{}: y_hat {} raw {} step {}'.format(name, out['free_y_hat'].shape, out['free_raw'].shape, out['step_raw'].shape))


if __name__ == '__main__':
	util_fixture()
	hparams_fixture()
	sampler_fixture()
	model_fixtures()
	print('done')
