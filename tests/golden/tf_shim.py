"""TEST INFRASTRUCTURE - a numpy stand-in for the slice of the TensorFlow 1.x API that the reference's WaveNet synthesis
path touches, so that the reference's OWN Python (/root/reference/code/wavenet_vocoder/{util,models/gaussian,models/mixture,
models/modules,models/wavenet}.py, datasets/audio.py, hparams.py) can be imported and executed in this container, where
TensorFlow, keras, librosa and matplotlib do not exist.

What runs unmodified from the reference: every line of graph-building Python on the path (WaveNet.__init__, .initialize
(synthesis branch), .incremental incl. its while_loop body, .step, ResidualConv1DGLU.step, CausalConv1D.call (incremental
and full-sequence branches), the samplers, the mu-law helpers, the upsampler classes incl. their NN-init kernels and the
periodic shuffle built from transposes / batch_to_space_nd).
What is RESTATED here from the published TF op semantics (there is no TensorFlow to ask): the kernels behind the ops -
matmul, bias_add, Conv1D (VALID, dilated), Conv2D (SAME), Conv2DTranspose (SAME), batch_to_space_nd, resize_images
(nearest), one_hot, split, pad, ... - written as plain numpy, independently of oracle/.

Eager evaluation: tensors are numpy arrays; `tf.while_loop` is a Python loop; `TensorArray` a list; variables come from a
weight provider (install(weights)); random draws come from an explicit noise feed (feed_noise) in the order the reference
consumes them.  `tf.shape(x)` returns an object whose elements compare by identity with Python numbers, as a TF-1 Tensor
does, so that source lines like `if tf.shape(test_inputs)[1] == 1:` take the branch they take under TensorFlow.
"""
import contextlib
import sys
import types

import numpy as np

F32 = np.float32
_state = {'weights': None, 'noise': None, 'variables': [], 'used': set()}


def install_weights(weights):
	_state['weights'] = weights
	_state['variables'] = []
	_state['used'] = set()


def feed_noise(normal=None, uniform_k=None, uniform_1=None, categorical=None):
	"""Per-step draws, consumed in call order: Normal.sample() takes normal[t] ([B]); random_uniform of shape [B,1,K] takes
	uniform_k[t], of shape [B,1] uniform_1[t]; multinomial takes categorical[t] ([B] uniforms)."""
	_state['noise'] = {'normal': normal, 'uk': uniform_k, 'u1': uniform_1, 'cat': categorical, 'i': {'normal': 0, 'uk': 0, 'u1': 0, 'cat': 0}}


def _next(kind):
	nz = _state['noise']
	i = nz['i'][kind]
	nz['i'][kind] = i + 1
	return nz[kind][i]


class _Dim(int):
	"""An element of tf.shape(x): arithmetic like an int, == / != with a Python number by identity (TF-1 Tensor.__eq__)."""
	def __eq__(self, other):
		return self is other
	def __ne__(self, other):
		return self is not other
	__hash__ = int.__hash__


class _ShapeT(tuple):
	"""tf.shape(x): indexable; comparing it with a tuple is an identity comparison (modules.py:240 is therefore False)."""
	def __eq__(self, other):
		return self is other
	def __ne__(self, other):
		return self is not other
	__hash__ = tuple.__hash__
	def __getitem__(self, i):
		v = tuple.__getitem__(self, i)
		return _Dim(v) if isinstance(v, int) else v


class _Var(np.ndarray):
	def get_shape(self):
		return _TensorShape(self.shape)


class _TensorShape(object):
	def __init__(self, dims):
		self.dims = [None if d is None else int(d) for d in dims]
	def as_list(self):
		return list(self.dims)
	def __len__(self):
		return len(self.dims)


def _variable(name, shape):
	w = _state['weights']
	if name not in w:
		raise KeyError('tf_shim: no weight for variable {!r}'.format(name))
	a = np.asarray(w[name], dtype=F32)
	assert tuple(a.shape) == tuple(int(s) for s in shape), (name, a.shape, tuple(shape))
	v = a.view(_Var)
	_state['variables'].append(v)
	_state['used'].add(name)
	return v


def _logical(layer_name):
	"""reference layer name -> logical weight name of tacotron-2_b200/wavenet_vocoder/weights.py"""
	import re
	m = re.match(r'residual_block_(causal|cin|gin|out|skip)_conv_(ResidualConv1DGLU_\d+)$', layer_name)
	if m:
		return '{}/residual_block_{}_conv'.format(m.group(2), m.group(1))
	m = re.match(r'(SubPixelConvolution|ResizeConvolution|ConvTranspose1D|ConvTranspose2D)_layer_(\d+)$', layer_name)
	if m:
		return 'local_conditioning_upsampling_{}'.format(int(m.group(2)) + 1)
	return layer_name	#input_convolution, final_convolution_1/2


# ---------------------------------------------------------------------------------------------------------------------
# ops
# ---------------------------------------------------------------------------------------------------------------------
def _a(x):
	return np.asarray(x)


def shape(x):
	return _ShapeT(int(d) for d in np.shape(x))


def rank(x):
	return np.ndim(x)


def reshape(x, shp, name=None):
	return np.reshape(_a(x), [int(s) for s in shp])


def transpose(x, perm=None):
	return np.transpose(_a(x), perm)


def concat(values, axis):
	return np.concatenate([_a(v) for v in values], axis=axis)


def expand_dims(x, axis):
	if isinstance(axis, (list, tuple)):
		axis = axis[0]
	return np.expand_dims(_a(x), axis)


def squeeze(x, axis=None):
	return np.squeeze(_a(x), axis=None if axis is None else tuple(axis))


def split(x, num_or_size_splits, axis=0):
	return np.split(_a(x), num_or_size_splits, axis=axis)


def tile(x, multiples):
	return np.tile(_a(x), [int(m) for m in multiples])


def cast(x, dtype):
	return _a(x).astype(dtype)


def pad(x, paddings):
	return np.pad(_a(x), [(int(a), int(b)) for a, b in _a(paddings)], mode='constant')


def constant(v, dtype=None):
	return np.asarray(v, dtype=dtype)


def identity(x, name=None):
	return x


def zeros(shp, dtype=F32, name=None):
	return np.zeros([int(s) for s in shp], dtype=dtype)


def ones(shp, dtype=F32):
	return np.ones([int(s) for s in shp], dtype=dtype)


def one_hot(indices, depth, dtype=F32):
	idx = _a(indices).astype(np.int64)
	out = np.zeros(idx.shape + (int(depth),), dtype=dtype)
	np.put_along_axis(out, idx[..., None], 1, axis=-1)
	return out


def argmax(x, axis=None):
	return np.argmax(_a(x), axis=axis)	#first maximum, like TF


def maximum(a, b):
	return np.maximum(a, F32(b) if np.isscalar(b) else b)


def minimum(a, b):
	return np.minimum(a, F32(b) if np.isscalar(b) else b)


def exp(x):
	return np.exp(_a(x))


def log(x):
	return np.log(_a(x))


def less(a, b):
	return a < b


def mod(a, b):
	return a % b


def reduce_sum(x, axis=None, keepdims=False):
	return np.sum(_a(x), axis=axis, keepdims=keepdims)


def reduce_max(x, axis=None, keepdims=False):
	return np.max(_a(x), axis=axis, keepdims=keepdims)


def sign(x):
	return np.sign(x)


def matmul(a, b):
	return np.matmul(_a(a), _a(b))


def assert_equal(a, b, message=None):
	if int(a) != int(b):
		raise AssertionError(message or 'tf.assert_equal({}, {})'.format(a, b))
	return None


@contextlib.contextmanager
def _ctx(*a, **k):
	yield types.SimpleNamespace()


def Print(x, data):
	return x


def while_loop(cond, body, loop_vars, parallel_iterations=None, swap_memory=None):
	v = list(loop_vars)
	while bool(cond(*v)):
		v = list(body(*v))
	return v


class TensorArray(object):
	def __init__(self, dtype=None, size=0, dynamic_size=True):
		self.items = {}
	def write(self, i, v):
		self.items[int(i)] = np.array(v)
		return self
	def stack(self):
		return np.stack([self.items[i] for i in range(len(self.items))])


def random_uniform(shp, minval=0., maxval=1.):
	shp = tuple(int(s) for s in shp)
	u = _a(_next('uk' if len(shp) == 3 else 'u1')).astype(F32)
	return u.reshape(shp)


def multinomial(logits, num_samples):
	"""Inverse-CDF restatement with ONE fed uniform per row (TF's own Philox stream cannot be reproduced): smallest k with
	cumsum(exp(x - max))[k] > u * sum, sequential float32 sums - the definition SURVEY.md 8c-10 gives the oracle."""
	x = _a(logits).astype(F32)
	u = _a(_next('cat')).astype(F32)
	out = np.zeros((x.shape[0], 1), np.int64)
	for b in range(x.shape[0]):
		e = np.exp(x[b] - x[b].max()).astype(F32)
		tot = F32(0)
		for v in e:
			tot = F32(tot + v)
		thr, acc, k = F32(u[b] * tot), F32(0), len(e) - 1
		for i, v in enumerate(e):
			acc = F32(acc + v)
			if acc > thr:
				k = i
				break
		out[b, 0] = k
	return out


class _Normal(object):
	def __init__(self, loc, scale, allow_nan_stats=False):
		self.loc, self.scale = loc, scale
	def sample(self):
		eps = _a(_next('normal')).astype(F32).reshape(np.shape(self.loc))
		return (self.loc + self.scale * eps).astype(F32)


# ---------------------------------------------------------------------------------------------------------------------
# nn / image / misc namespaces
# ---------------------------------------------------------------------------------------------------------------------
def _bias_add(x, b, data_format=None):
	if data_format == 'NCHW':
		return (_a(x) + _a(b).reshape((1, -1) + (1,) * (np.ndim(x) - 2))).astype(F32)
	return (_a(x) + _a(b)).astype(F32)


def _softmax(x, axis=-1):
	e = np.exp(x - np.max(x, axis=axis, keepdims=True))
	return e / np.sum(e, axis=axis, keepdims=True)


def _resize_images(x, size, method=1):
	"""tf.image.resize_images, NEAREST_NEIGHBOR, align_corners=False: src = floor(dst * in / out).  NHWC."""
	assert method == 1
	x = _a(x)
	H, W = x.shape[1], x.shape[2]
	Ho, Wo = int(size[0]), int(size[1])
	hi = np.minimum(np.floor(np.arange(Ho) * (F32(H) / F32(Ho))).astype(np.int64), H - 1)
	wi = np.minimum(np.floor(np.arange(Wo) * (F32(W) / F32(Wo))).astype(np.int64), W - 1)
	return x[:, hi][:, :, wi]


def batch_to_space_nd(x, block_shape, crops):
	"""tf.batch_to_space_nd with one spatial dimension and zero crops: input [b*batch, s, ...] -> [batch, s*b, ...] with
	out[n, i*b + k, ...] = in[k*batch + n, i, ...] (the batch index is block-major)."""
	x = _a(x)
	b = int(block_shape[0])
	assert len(block_shape) == 1 and all(int(c) == 0 for c in np.ravel(crops))
	nb, s = x.shape[0], x.shape[1]
	batch = nb // b
	y = x.reshape((b, batch, s) + x.shape[2:])
	y = np.moveaxis(y, 0, 2)	#[batch, s, b, ...]
	return y.reshape((batch, s * b) + x.shape[2:])


def _dropout(x, rate=0.5, training=False):
	if training and rate != 0:
		raise RuntimeError('tf_shim: dropout with a non-zero rate is random; run step() with wavenet_dropout=0')
	return x


def _same_pads(k):
	"""TF SAME padding at stride 1: total k - 1, the smaller half in front."""
	return (k - 1) // 2, (k - 1) - (k - 1) // 2


class _Layer(object):
	"""tf.layers.Layer: variables are created by build() from the weight provider, keyed by the layer name."""
	def __init__(self, filters=None, kernel_size=None, strides=1, padding='valid', data_format='channels_last', dilation_rate=1,
			activation=None, use_bias=True, kernel_initializer=None, bias_initializer=None, name=None, trainable=True, **kw):
		tup = lambda v, n: tuple(v) if isinstance(v, (tuple, list)) else (v,) * n
		self.rank = self._rank
		self.filters, self.kernel_size = filters, tup(kernel_size, self._rank)
		self.strides, self.dilation_rate = tup(strides, self._rank), tup(dilation_rate, self._rank)
		self.padding, self.data_format, self.use_bias, self.name = padding, data_format, use_bias, name
		self.activation = activation
		self.built = False
		self.kernel = self.bias = None

	def _kernel_shape(self, cin):
		return self.kernel_size + (cin, self.filters)

	def build(self, input_shape):
		shp = input_shape.as_list() if hasattr(input_shape, 'as_list') else list(input_shape)
		cin = shp[1] if self.data_format == 'channels_first' else shp[-1]
		base = _logical(self.name)
		self.kernel = _variable(base + '/kernel', self._kernel_shape(int(cin)))
		self.bias = _variable(base + '/bias', (self._n_bias(),)) if self.use_bias else None
		self.built = True

	def _n_bias(self):
		return self.filters

	def __call__(self, inputs, *a, **k):
		if not self.built:
			self.build(np.shape(inputs))
		return self.call(inputs, *a, **k)


class Conv1D(_Layer):
	_rank = 1
	def call(self, inputs):
		"""VALID dilated convolution; NCW or NWC.  out[b, t, f] = sum_{k, c} x[b, t + k*d, c] K[k, c, f] (+ bias)."""
		x = _a(inputs)
		if self.data_format == 'channels_first':
			x = np.transpose(x, [0, 2, 1])
		kw, d = self.kernel_size[0], self.dilation_rate[0]
		T = x.shape[1] - (kw - 1) * d
		out = np.zeros((x.shape[0], T, self.filters), F32)
		for k in range(kw):
			out += np.matmul(x[:, k * d:k * d + T, :], _a(self.kernel)[k])
		if self.use_bias:
			out = out + _a(self.bias)
		out = out.astype(F32)
		return np.transpose(out, [0, 2, 1]) if self.data_format == 'channels_first' else out


class Conv2D(_Layer):
	_rank = 2
	def call(self, inputs):
		"""SAME, stride 1, NHWC.  out[b, h, w, f] = bias[f] + sum_{i, j, c} x[b, h + i - ph, w + j - pw, c] K[i, j, c, f]."""
		assert self.data_format == 'channels_last' and self.padding == 'same' and self.strides == (1, 1)
		x = _a(inputs)
		kh, kw = self.kernel_size
		(pt, pb), (pl, pr) = _same_pads(kh), _same_pads(kw)
		xp = np.pad(x, [(0, 0), (pt, pb), (pl, pr), (0, 0)], mode='constant')
		H, W = x.shape[1], x.shape[2]
		out = np.zeros(x.shape[:3] + (int(np.shape(self.kernel)[-1]),), F32)
		for i in range(kh):
			for j in range(kw):
				out += np.matmul(xp[:, i:i + H, j:j + W, :], _a(self.kernel)[i, j])
		if self.use_bias:
			out = out + _a(self.bias)
		return out.astype(F32)


class Conv2DTranspose(_Layer):
	_rank = 2
	def _kernel_shape(self, cin):
		return self.kernel_size + (self.filters, cin)	#[kh, kw, out, in]
	def call(self, inputs):
		"""SAME, NCHW, strides (1, s) with kernel width == s (every use in the reference).  The transposed convolution is the
		gradient of a forward SAME conv: out[b, co, y, x*s + j] = bias[co] + sum_{i, ci} in[b, ci, y + ph - i, x] K[i, j, co, ci],
		ph = (kh - 1) // 2 (height stride 1, SAME); width: stride s = kernel width, SAME output width W*s, no overlap, no crop."""
		assert self.data_format == 'channels_first' and self.padding == 'same'
		x = _a(inputs)
		kh, kw = self.kernel_size
		sh, sw = self.strides
		assert sh == 1 and sw == kw
		ph = (kh - 1) // 2
		B, Ci, H, W = x.shape
		K = _a(self.kernel)
		out = np.zeros((B, self.filters, H, W * sw), F32)
		for i in range(kh):
			for y in range(H):
				ys = y + ph - i
				if ys < 0 or ys >= H:
					continue
				#[B, Ci, W] x [kw, Co, Ci] -> [B, Co, W, kw]
				contrib = np.einsum('bcw,joc->bowj', x[:, :, ys, :], K[i])
				out[:, :, y, :] += contrib.reshape(B, self.filters, W * sw)
		if self.use_bias:
			out = out + _a(self.bias).reshape(1, -1, 1, 1)
		return out.astype(F32)


class _Wrapper(object):
	"""tf.keras.layers.Wrapper as the reference uses it: holds .layer, builds on first call, forwards to call()."""
	def __init__(self, layer, name=None, **kwargs):
		self.layer = layer
		self.name = name
		self.built = False
	def _track_checkpointable(self, *a, **k):
		pass
	def build(self, input_shape=None):
		self.built = True
	def __call__(self, inputs, *a, **k):
		if not self.built:
			self.build(np.shape(inputs))
		return self.call(inputs, *a, **k)


class HParams(object):
	"""tf.contrib.training.HParams: attribute bag with .values() and .parse('k=v,...')."""
	def __init__(self, **kw):
		self.__dict__['_v'] = dict(kw)
	def __getattr__(self, k):
		try:
			return self.__dict__['_v'][k]
		except KeyError:
			raise AttributeError(k)
	def __setattr__(self, k, v):
		self._v[k] = v
	def values(self):
		return dict(self._v)
	def parse(self, s):
		import ast
		for part in [p for p in s.split(',') if p.strip()]:
			k, v = part.split('=', 1)
			k = k.strip()
			old = self._v[k]
			try:
				val = ast.literal_eval(v.strip())
			except Exception:
				val = v.strip()
			self._v[k] = type(old)(val) if old is not None and not isinstance(old, (list, type(None))) and not isinstance(val, str) else val
		return self
	def set_hparam(self, k, v):
		self._v[k] = v


def install():
	"""Put the stand-in modules into sys.modules (tensorflow, keras.utils.np_utils, librosa, matplotlib, sklearn bits)."""
	ns = types.SimpleNamespace
	tf = types.ModuleType('tensorflow')
	for name, obj in list(globals().items()):
		if not name.startswith('_') and callable(obj) and name not in ('install', 'install_weights', 'feed_noise'):
			setattr(tf, name, obj)
	tf.float32, tf.int32, tf.int64 = np.float32, np.int32, np.int64
	tf.variable_scope = tf.control_dependencies = tf.device = tf.name_scope = _ctx
	tf.nn = ns(bias_add=_bias_add, tanh=lambda x: np.tanh(x).astype(F32), sigmoid=lambda x: (F32(1) / (F32(1) + np.exp(-_a(x)))).astype(F32),
		relu=lambda x, name=None: np.maximum(x, F32(0)), leaky_relu=lambda x, alpha=0.2, name=None: np.where(x > 0, x, F32(alpha) * x).astype(F32),
		softmax=_softmax, embedding_lookup=lambda table, ids: _a(table)[_a(ids)])
	tf.layers = ns(Conv1D=Conv1D, Conv2D=Conv2D, Conv2DTranspose=Conv2DTranspose, dropout=_dropout,
		InputSpec=lambda shape=None, **k: ns(shape=shape))
	tf.keras = ns(layers=ns(Wrapper=_Wrapper))
	tf.image = ns(resize_images=_resize_images)
	tf.TensorShape = _TensorShape
	tf.zeros_initializer = lambda *a, **k: None
	tf.constant_initializer = lambda value=None, dtype=None, **k: ns(value=value)
	tf.truncated_normal_initializer = lambda *a, **k: None
	tf.get_variable = lambda name, shape=None, dtype=None, initializer=None, **k: _variable(name, shape)
	tf.trainable_variables = lambda: list(_state['variables'])
	tf.train = ns(replica_device_setter=lambda **k: None, ExponentialMovingAverage=lambda decay=None: ns(decay=decay))
	tf.contrib = ns(distributions=ns(Normal=_Normal), training=ns(HParams=HParams))
	tf.logging = ns(set_verbosity=lambda *a: None, ERROR=0)
	tf.softmax = _softmax
	sys.modules['tensorflow'] = tf
	keras = types.ModuleType('keras')
	keras.utils = types.ModuleType('keras.utils')
	keras.utils.np_utils = types.ModuleType('keras.utils.np_utils')
	keras.utils.np_utils.to_categorical = lambda y, num_classes=None: np.eye(num_classes or (int(np.max(y)) + 1), dtype=F32)[np.asarray(y, np.int64)]
	sys.modules.update({'keras': keras, 'keras.utils': keras.utils, 'keras.utils.np_utils': keras.utils.np_utils})
	for name in ('librosa', 'librosa.filters', 'librosa.display', 'matplotlib', 'matplotlib.pyplot'):
		m = types.ModuleType(name)
		m.use = lambda *a, **k: None
		sys.modules[name] = m
	sys.modules['librosa'].filters = sys.modules['librosa.filters']
	sys.modules['librosa'].display = sys.modules['librosa.display']
	sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
	return tf
