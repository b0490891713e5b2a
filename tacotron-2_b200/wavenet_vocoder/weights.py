"""Weight table of the WaveNet vocoder: logical names, TensorFlow layouts, deterministic random init, npz IO.

Layouts are the ones the reference's layers create:
``tf.layers.Conv1D`` kernel ``[k, C_in, C_out]`` + bias ``[C_out]`` (``modules.py:206-224``); upsamplers
``Conv2D`` kernel ``[kh, kw, 1, filters]`` (``modules.py:539-695``) / ``Conv2DTranspose`` kernel
``[kh, kw, out, in]`` (``modules.py:697-770``); embedding ``[n_speakers, gin]`` (``modules.py:10-21``).
Names follow the reference's scope strings (``wavenet.py:103-205``, ``modules.py:407-450``) without the
auto-generated TF prefixes, which cannot be reproduced without TensorFlow.

The same arrays feed the CUDA library (through ``wn_load_weights``) and the test oracle.
"""
from collections import OrderedDict

import numpy as np

from wavenet_vocoder.util import is_scalar_input


def layer_dilations(hp):
	"""dilation of residual layer l = 2**(l mod layers_per_stack)  (wavenet.py:118-125)."""
	per_stack = hp.layers // hp.stacks
	return [2 ** (l % per_stack) for l in range(hp.layers)]


def receptive_field_size(total_layers, num_cycles, kernel_size, dilation=lambda x: 2 ** x):
	"""(k-1)*sum(dilations)+1  (wavenet.py:54-71)."""
	assert total_layers % num_cycles == 0
	layers_per_cycle = total_layers // num_cycles
	dilations = [dilation(i % layers_per_cycle) for i in range(total_layers)]
	return (kernel_size - 1) * sum(dilations) + 1


def upsample_layer_specs(hp):
	"""[(kernel_shape, bias_shape, scale)] per learnable upsampling layer (wavenet.py:163-192)."""
	if hp.cin_channels <= 0 or hp.upsample_type == 'NearestNeighbor':
		return []
	kh = hp.freq_axis_kernel_size
	out = []
	for s in hp.upsample_scales:
		if hp.upsample_type == 'SubPixel':
			out.append(((kh, 3, 1, s), (s,), s))
		elif hp.upsample_type == 'Resize':
			out.append(((kh, s, 1, 1), (1,), s))
		elif hp.upsample_type == '2D':
			out.append(((kh, s, 1, 1), (1,), s))
		elif hp.upsample_type == '1D':
			out.append(((1, s, hp.cin_channels, hp.cin_channels), (hp.cin_channels,), s))
		else:
			raise AssertionError('unknown upsample_type {}'.format(hp.upsample_type))
	return out


def weight_specs(hp):
	"""OrderedDict name -> shape of every tensor the synthesis path reads."""
	R, G, S, O, k = hp.residual_channels, hp.gate_channels, hp.skip_out_channels, hp.out_channels, hp.kernel_size
	in_ch = 1 if is_scalar_input(hp.input_type) else hp.quantize_channels
	specs = OrderedDict()

	def conv(name, kw, cin, cout, bias=None):
		specs[name + '/kernel'] = (kw, cin, cout)
		if hp.use_bias if bias is None else bias:
			specs[name + '/bias'] = (cout,)

	# hparams.use_bias reaches the residual blocks only (wavenet.py:118-133); the input and the two final 1x1s are built
	# without the argument and always carry a bias (wavenet.py:103-115, :136-149; Conv1D1x1 default use_bias=True)
	conv('input_convolution', 1, in_ch, R, bias=True)
	for l in range(hp.layers):
		p = 'ResidualConv1DGLU_{}/'.format(l)
		conv(p + 'residual_block_causal_conv', k, R, G)
		if hp.cin_channels > 0:
			conv(p + 'residual_block_cin_conv', 1, hp.cin_channels, G)
		if hp.gin_channels > 0:
			conv(p + 'residual_block_gin_conv', 1, hp.gin_channels, G)
		conv(p + 'residual_block_out_conv', 1, G // 2, R)
		conv(p + 'residual_block_skip_conv', 1, G // 2, S)
	conv('final_convolution_1', 1, S, S, bias=True)
	conv('final_convolution_2', 1, S, O, bias=True)
	if hp.gin_channels > 0 and hp.use_speaker_embedding:
		specs['gc_embedding'] = (hp.n_speakers, hp.gin_channels)
	for i, (kshape, bshape, _) in enumerate(upsample_layer_specs(hp)):
		specs['local_conditioning_upsampling_{}/kernel'.format(i + 1)] = kshape
		specs['local_conditioning_upsampling_{}/bias'.format(i + 1)] = bshape
	return specs


def count_params(hp):
	return int(sum(int(np.prod(s)) for s in weight_specs(hp).values()))


def nn_init_upsample_kernel(hp, layer_index):
	"""The reference's 'nearest-neighbour' initial upsampling kernels
	(modules.py:642-654 SubPixel, :685-695 Resize, :724-733 1D, :761-770 2D)."""
	kshape, _, s = upsample_layer_specs(hp)[layer_index]
	n_layers = len(hp.upsample_scales)
	scale = hp.NN_scaler ** (1 / n_layers)
	kh, kw = kshape[0], kshape[1]
	overlap = kw // s
	if hp.upsample_type == '1D':
		eye = np.eye(hp.cin_channels, dtype=np.float32).reshape(1, 1, hp.cin_channels, hp.cin_channels)
		kern = np.tile(eye, [kh, kw, 1, 1])
		if kw % 2 == 0:
			kern = kern / max(float(overlap), 1.)
		return (kern * scale).astype(np.float32)
	base = np.zeros((kh, kw), dtype=np.float32)
	i = kh // 2
	if hp.upsample_type == '2D':
		cols = range(kw)
	else:
		cols = [kw // 2 - 1, kw // 2] if kw % 2 == 0 else [kw // 2]
	for j in cols:
		base[i, j] = 1. / max(overlap, 1.) if kw % 2 == 0 else 1.
	kern = np.tile(base[:, :, None, None], [1, 1, kshape[2], kshape[3]])
	return (kern * scale).astype(np.float32)


def _glorot_uniform(rng, shape):
	"""tf glorot_uniform for a conv kernel [.., C_in, C_out]: receptive field folds into both fans."""
	rf = int(np.prod(shape[:-2])) if len(shape) > 2 else 1
	fan_in, fan_out = rf * shape[-2], rf * shape[-1]
	limit = np.sqrt(6.0 / (fan_in + fan_out))
	return rng.uniform(-limit, limit, size=shape).astype(np.float32)


def init_weights(hp, seed=5339, kind='stress'):
	"""Deterministic random-init weights (there is no network access for trained checkpoints).

	kind='tf-init': what a freshly built reference graph would hold in distribution - Glorot-uniform kernels,
	zero biases, NN-init upsampling kernels when hp.NN_init, truncated-normal(0.1) embedding.
	kind='stress' (parity tests): same kernels, biases ~N(0, 0.05), upsampling kernels ~U(-0.5, 0.5) with biases
	~N(0, 0.05), and the head bias of every log-scale channel set to -4 so free-running samples are not
	pinned to the +-1 clip (a saturated stream would hide errors).  SURVEY.md section 8d.
	"""
	assert kind in ('tf-init', 'stress')
	rng = np.random.default_rng(seed)
	specs = weight_specs(hp)
	w = OrderedDict()
	for name, shape in specs.items():
		if name.startswith('local_conditioning_upsampling_'):
			idx = int(name.split('_')[-1].split('/')[0]) - 1
			if name.endswith('/kernel'):
				if kind == 'tf-init' and hp.NN_init:
					w[name] = nn_init_upsample_kernel(hp, idx)
					rng.uniform(size=shape)	#keep the stream position independent of NN_init
				elif kind == 'tf-init':
					w[name] = _glorot_uniform(rng, shape)
				else:
					w[name] = rng.uniform(-0.5, 0.5, size=shape).astype(np.float32)
			else:
				w[name] = (rng.normal(0., 0.05, size=shape).astype(np.float32) if kind == 'stress'
					else np.zeros(shape, np.float32))
		elif name == 'gc_embedding':
			w[name] = np.clip(rng.normal(0., 0.1, size=shape), -0.2, 0.2).astype(np.float32)
		elif name.endswith('/kernel'):
			w[name] = _glorot_uniform(rng, shape)
		else:
			w[name] = (rng.normal(0., 0.05, size=shape).astype(np.float32) if kind == 'stress'
				else np.zeros(shape, np.float32))
	if kind == 'stress' and is_scalar_input(hp.input_type):
		b = w['final_convolution_2/bias']
		if hp.out_channels == 2:
			b[1] = -4.
		else:
			K = hp.out_channels // 3
			b[2 * K:3 * K] = -4.
	return w


def fold_weight_norm(hp, weights):
	"""What `wavenet_weight_normalization=True` means at synthesis time in the reference (modules.py:44-181).

	WeightNorm keeps the layer's `kernel` variable as the direction v plus a gain g, but re-assigns g := ||v|| under a
	control dependency every time the effective kernel is computed (:162-164), so the effective kernel is
	l2_normalize(v) * ||v|| = v up to rounding, whatever g a checkpoint holds.  The incremental convolutions do not even do
	that: CausalConv1D reads `self.layer.kernel`, which on the wrapper is the raw v captured before normalisation
	(:151, :245).  Hence: residual / input / final convolutions use `kernel` as stored; the upsampling layers, which run
	through WeightNorm.call, use l2_normalize(v) * ||v|| in float32 (TF: x * rsqrt(max(sum x^2, 1e-12)); norm over all axes
	but the last); `g` tensors are ignored.  Returns a new table."""
	if not getattr(hp, 'wavenet_weight_normalization', False):
		return weights
	out = OrderedDict()
	for name, arr in weights.items():
		if name.endswith('/g'):
			continue
		a = np.asarray(arr, np.float32)
		if name.startswith('local_conditioning_upsampling_') and name.endswith('/kernel'):
			sq = np.sum(np.square(a, dtype=np.float32), axis=tuple(range(a.ndim - 1)), keepdims=True, dtype=np.float32)
			unit = a * (np.float32(1.0) / np.sqrt(np.maximum(sq, np.float32(1e-12)))).astype(np.float32)
			a = (unit * np.sqrt(sq).astype(np.float32)).astype(np.float32)
		out[name] = a
	return out


def check_weights(hp, weights):
	"""Raise ValueError unless `weights` holds exactly the tensors of weight_specs(hp) with those shapes."""
	specs = weight_specs(hp)
	missing = [n for n in specs if n not in weights]
	extra = [n for n in weights if n not in specs]
	if missing or extra:
		raise ValueError('weight table mismatch: missing {} unexpected {}'.format(missing[:4], extra[:4]))
	for n, shape in specs.items():
		if tuple(weights[n].shape) != tuple(shape):
			raise ValueError('weight {} has shape {} expected {}'.format(n, tuple(weights[n].shape), tuple(shape)))
	return True


def save_weights(path, weights):
	np.savez(path, **{k.replace('/', '.'): np.asarray(v, np.float32) for k, v in weights.items()})


def load_weights(path):
	with np.load(path) as z:
		return OrderedDict((k.replace('.', '/'), z[k].astype(np.float32)) for k in z.files)
