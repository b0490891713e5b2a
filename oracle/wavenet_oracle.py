"""TEST INFRASTRUCTURE - CPU oracle for the WaveNet incremental synthesis hot path (numpy, float32).

This file restates, op for op, what the reference's TF-1.x graph computes on the path
``Synthesizer.synthesize -> WaveNet.initialize (synthesis branch) -> WaveNet.incremental / step``.
It is the checker the CUDA path is compared with.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the product never does.

PARITY STATUS: pinned to the reference's own Python, executed here; the TensorFlow kernels themselves stay unpinned.
The reference ships no golden vectors, unit tests or fixtures for this path and TensorFlow 1.x cannot be
installed here (F3).  What is pinned, and how:
(a) tests/golden/make_ref_golden.py imports /root/reference/code and EXECUTES the reference's functions over a numpy
    stand-in for the `tf` namespace (tests/golden/tf_shim.py): util.py mu-law helpers, feeder._interp,
    datasets/audio.save_wavenet_wav, receptive_field_size, hparams.py / paper_hparams.py (WaveNet keys),
    sample_from_gaussian, sample_from_discretized_mix_logistic, and - on six seeded small models covering every input
    type, head and upsampler - create_model -> WaveNet.initialize (synthesis branch) -> WaveNet.incremental (its
    while_loop body, CausalConv1D.call(incremental=True), ResidualConv1DGLU.step; free running and teacher forced) and
    WaveNet.step.  The outputs are committed as tests/golden/ref_*.npz / ref_hparams.json and
    tests/test_oracle_vs_reference_shim.py holds this oracle and oracle/c to them (1e-5; discrete results exactly).
(b) NOT pinned: the kernels behind the TF ops (matmul, Conv1D / Conv2D SAME / Conv2DTranspose, batch_to_space_nd,
    resize_images) - the stand-in restates those from their published semantics, independently of this file; and global
    conditioning, which the reference's own synthesis branch rejects (wavenet.py:430 compares a TensorShape with a tuple
    holding a Tensor) and is therefore restated only.
(c) self-consistency on top: incremental(teacher forced) == step() with the one-sample conditioning shift, ring buffer ==
    literal shift queue, every upsampler == brute-force loops, NN-init == scaled nearest-neighbour repeat.

Citations are file:line into /root/reference/code/wavenet_vocoder/.
"""
import numpy as np

F32 = np.float32
SQRT_HALF = F32(np.sqrt(0.5))	#np.sqrt(0.5) is float64 in the reference, cast to f32 by the TF multiply


def _f32(x):
	return np.asarray(x, dtype=F32)


# --------------------------------------------------------------------------------------------
# util.py:13-129  input types and mu-law
# --------------------------------------------------------------------------------------------
def is_mulaw_quantize(s):
	assert s in ('mulaw-quantize', 'mulaw', 'raw')
	return s == 'mulaw-quantize'


def is_mulaw(s):
	assert s in ('mulaw-quantize', 'mulaw', 'raw')
	return s == 'mulaw'


def is_scalar_input(s):
	return s in ('raw', 'mulaw')


def mulaw(x, mu=256):
	mu = 255	#util.py:48 overrides the argument
	return np.sign(x) * np.log1p(mu * np.abs(x)) / np.log1p(mu)


def inv_mulaw(y, mu=256):
	mu = 255	#util.py:67
	return np.sign(y) * (1.0 / mu) * ((1.0 + mu) ** np.abs(y) - 1.0)


def mulaw_quantize(x, mu=256):
	mu = 255	#util.py:99
	y = mulaw(x, mu)
	y = (y + 1) / 2 * mu
	return y.astype(np.int64) if isinstance(y, np.ndarray) else int(y)	#truncation (util.py:102, _asint)


def inv_mulaw_quantize(y, mu=256):
	mu = 255	#util.py:127
	y = 2 * _f32(y) / mu - 1	#_asfloat -> float32 (util.py:128, :159-163)
	return inv_mulaw(y, mu)


# --------------------------------------------------------------------------------------------
# synthesizer.py:56-70  host prologue;  :100-104 crop;  datasets/audio.py:17-20 wav conversion
# --------------------------------------------------------------------------------------------
def prepare_local_conditions(mels, hp):
	"""list of [Tc_i, num_mels] -> (c_batch [B, max Tc, num_mels] f32, audio_lengths)."""
	hop = hp.hop_size if hp.hop_size is not None else int(hp.frame_shift_ms / 1000 * hp.sample_rate)
	audio_lengths = [len(x) * hop for x in mels]
	maxlen = max(len(x) for x in mels)
	rng_ = (-hp.max_abs_value, hp.max_abs_value) if hp.symmetric_mels else (0, hp.max_abs_value)
	if hp.clip_for_wavenet:
		mels = [np.clip(x, rng_[0], rng_[1]) for x in mels]
	c_batch = np.stack([np.pad(x, [(0, maxlen - len(x)), (0, 0)], mode='constant', constant_values=rng_[0])
		for x in mels]).astype(F32)
	if hp.normalize_for_wavenet:
		c_batch = ((c_batch - rng_[0]) / (rng_[1] - rng_[0])).astype(F32)	#feeder.py:426-428
	return c_batch, audio_lengths


def wav_to_int16(wav):
	wav = _f32(wav).copy()
	wav *= 32767 / max(0.01, np.max(np.abs(wav)))
	return wav.astype(np.int16)


# --------------------------------------------------------------------------------------------
# TF op restatements used by the upsamplers (published TF semantics; see module docstring)
# --------------------------------------------------------------------------------------------
def _same_pad(k):
	total = k - 1
	return total // 2, total - total // 2


def conv2d_same_nhwc(x, kernel, bias):
	"""tf.layers.Conv2D(strides 1, padding 'same', channels_last).  x [B,H,W,Cin], kernel [kh,kw,Cin,F]."""
	kh, kw, cin, f = kernel.shape
	pt, pb = _same_pad(kh)
	pl, pr = _same_pad(kw)
	xp = np.pad(x, [(0, 0), (pt, pb), (pl, pr), (0, 0)])
	B, H, W, _ = x.shape
	y = np.zeros((B, H, W, f), F32)
	for i in range(kh):
		for j in range(kw):
			y += xp[:, i:i + H, j:j + W, :] @ kernel[i, j]
	return (y + bias).astype(F32)


def resize_nearest_width(x, s):
	"""tf.image.resize_images(method=NEAREST, align_corners=False) along W of [B,H,W,C] by integer factor s:
	in_x = min(floor(out_x * (in_w / float(out_w))), in_w - 1), evaluated in float32 like the TF kernel."""
	W = x.shape[2]
	out_w = W * s
	scale = F32(W) / F32(out_w)
	src = np.minimum(np.floor(np.arange(out_w, dtype=F32) * scale).astype(np.int64), W - 1)
	return x[:, :, src, :]


# --------------------------------------------------------------------------------------------
# modules.py:524-770  the five upsamplers.  All take/return [B, 80, W] (freq-major) float32.
# --------------------------------------------------------------------------------------------
def _activation(x, hp):
	if hp.upsample_activation == 'Relu':
		return np.maximum(x, F32(0))	#modules.py:23-30
	if hp.upsample_activation == 'LeakyRelu':
		return np.where(x > 0, x, F32(hp.leaky_alpha) * x).astype(F32)	#modules.py:33-41
	assert hp.upsample_activation is None
	return x


def effective_upsample_kernel(hp, kernel):
	"""SubPixelConvolution.build (modules.py:584-592): without NN_init every output filter is filter 0."""
	if hp.upsample_type == 'SubPixel' and not hp.NN_init:
		return np.tile(kernel[:, :, :, 0:1], [1, 1, 1, kernel.shape[3]])
	return kernel


def upsample_layer(x, kernel, bias, s, hp):
	"""One learnable upsampling layer, x [B,80,W] -> [B,80,W*s]."""
	t = hp.upsample_type
	if t == 'SubPixel':
		#modules.py:596-640: conv2d SAME on [B,H,W,1] with s filters, then periodic shuffle out[..., w*s+f] = y[..., w, f]
		y = conv2d_same_nhwc(x[:, :, :, None], effective_upsample_kernel(hp, kernel), bias)	#[B,H,W,s]
		B, H, W, _ = y.shape
		return y.reshape(B, H, W * s)
	if t == 'Resize':
		#modules.py:678-683: NN resize then conv2d SAME (kh, s), one filter
		r = resize_nearest_width(x[:, :, :, None], s)
		return conv2d_same_nhwc(r, kernel, bias)[:, :, :, 0]
	if t == '2D':
		#modules.py:736-759: conv2d_transpose, kernel [kh, s, 1, 1], strides (1, s), SAME, input [B,1,80,W]
		kh = kernel.shape[0]
		pt = (kh - 1) // 2
		B, H, W = x.shape
		xp = np.pad(x, [(0, 0), (kh, kh), (0, 0)])
		out = np.zeros((B, H, W, s), F32)
		for i in range(kh):
			#out[h, w*s+j] += x[h + pt - i, w] * K[i, j]
			rows = xp[:, kh + pt - i: kh + pt - i + H, :]
			out += rows[:, :, :, None] * kernel[i, :, 0, 0][None, None, None, :]
		return (out.reshape(B, H, W * s) + bias[0]).astype(F32)
	if t == '1D':
		#modules.py:697-722: conv2d_transpose, kernel [1, s, out=80, in=80], input [B,80,1,W]
		#out[b, co, w*s+j] = bias[co] + sum_ci x[b, ci, w] * K[0, j, co, ci]
		B, C, W = x.shape
		out = np.einsum('biw,joi->bowj', x, kernel[0]).astype(F32)
		return (out.reshape(B, kernel.shape[2], W * s) + bias[None, :, None]).astype(F32)
	raise AssertionError(t)


# --------------------------------------------------------------------------------------------
# The model
# --------------------------------------------------------------------------------------------
class WaveNetOracle(object):
	"""hp: any object with the WaveNet hparams as attributes; weights: dict name -> float32 array in TF layouts
	(names: tacotron-2_b200/wavenet_vocoder/weights.py)."""

	def __init__(self, hp, weights):
		self.hp = hp
		self.w = {k: _f32(v) for k, v in weights.items()}
		assert hp.layers % hp.stacks == 0	#wavenet.py:97
		per_stack = hp.layers // hp.stacks
		self.dilations = [2 ** (l % per_stack) for l in range(hp.layers)]	#wavenet.py:125
		self.scalar_input = is_scalar_input(hp.input_type)
		self.receptive_field = (hp.kernel_size - 1) * sum(self.dilations) + 1	#wavenet.py:54-71

	# -- small accessors ------------------------------------------------------------------
	def _lin(self, name):
		"""_get_linearized_weight (modules.py:238-251): [kw, in, filters] -> [kw*in, filters], row = tap*in + ci."""
		k = self.w[name + '/kernel']
		return k.reshape(-1, k.shape[-1])

	def _bias(self, name):
		return self.w.get(name + '/bias')

	def _dense(self, x, name):
		"""Conv1D1x1.incremental_step on [B, C] (modules.py:294-297): matmul + bias_add."""
		y = x @ self._lin(name)
		b = self._bias(name)
		return (y + b).astype(F32) if b is not None else y.astype(F32)

	# -- wavenet.py:781-803 (same code at :680-702) ---------------------------------------
	def upsample(self, c):
		"""c [B, 80, Tc] -> [B, 80, T]."""
		hp = self.hp
		c = _f32(c)
		if hp.upsample_type == 'NearestNeighbor':
			hop = hp.hop_size if hp.hop_size is not None else int(hp.frame_shift_ms / 1000 * hp.sample_rate)
			return resize_nearest_width(c[:, :, :, None], hop)[:, :, :, 0]	#modules.py:524-536, wavenet.py:165-167
		for i, s in enumerate(hp.upsample_scales):
			p = 'local_conditioning_upsampling_{}'.format(i + 1)
			c = upsample_layer(c, self.w[p + '/kernel'], self.w[p + '/bias'], s, hp)
			c = _activation(c, hp)	#wavenet.py:197-203
		return c

	def _global_embedding(self, g, B):
		"""wavenet.py:771-778: ids [B,1] -> table rows [B, gin] (constant over time)."""
		if g is None:
			return None
		g = np.asarray(g)
		if 'gc_embedding' in self.w:
			return self.w['gc_embedding'][g.reshape(B).astype(np.int64)]
		return _f32(g).reshape(B, -1)

	def initial_input(self, B):
		"""wavenet.py:433-445."""
		hp = self.hp
		if is_mulaw_quantize(hp.input_type):
			v = mulaw_quantize(0, hp.quantize_channels)
			assert 0 <= v < hp.quantize_channels
			x = np.zeros((B, hp.quantize_channels), F32)
			x[:, v] = 1
			return x
		v = mulaw(0.0, hp.quantize_channels) if is_mulaw(hp.input_type) else 0.0
		return np.full((B, 1), v, F32)

	# -- samplers -------------------------------------------------------------------------
	def sample_gaussian(self, out, eps):
		"""gaussian.py:39-52.  out [B,2]; eps [B] standard normal draws."""
		mean = out[:, 0]
		log_scale = np.maximum(out[:, 1], F32(self.hp.log_scale_min_gauss))
		scale = np.exp(log_scale)
		x = (mean + scale * _f32(eps)).astype(F32)
		return np.minimum(np.maximum(x, F32(-1.)), F32(1.))

	def sample_mol(self, out, u):
		"""mixture.py:76-107.  out [B,3K]; u [B,K+1] uniforms in [1e-5, 1-1e-5]: K for the mixture pick, 1 for the logistic."""
		K = out.shape[1] // 3
		logit_probs = out[:, :K]
		temp = logit_probs - np.log(-np.log(_f32(u[:, :K])))
		argmax = np.argmax(temp, axis=-1)	#lowest index on ties, like tf.argmax
		one_hot = np.eye(K, dtype=F32)[argmax]
		means = np.sum(out[:, K:2 * K] * one_hot, axis=-1)
		log_scales = np.maximum(np.sum(out[:, 2 * K:3 * K] * one_hot, axis=-1), F32(self.hp.log_scale_min))
		u2 = _f32(u[:, K])
		x = (means + np.exp(log_scales) * (np.log(u2) - np.log(F32(1.) - u2))).astype(F32)
		return np.minimum(np.maximum(x, F32(-1.)), F32(1.))

	def sample_categorical(self, logits, u):
		"""wavenet.py:860-869 draws with tf.multinomial, whose stream cannot be fed; the oracle *defines* the draw
		from one uniform u[b] by inverse CDF: e_k = exp(x_k - max), S = sequential sum,
		pick the smallest k with cumsum_k > u*S (else the last class)."""
		B, C = logits.shape
		idx = np.zeros(B, np.int64)
		for b in range(B):
			m = np.max(logits[b])
			e = np.exp(logits[b] - m).astype(F32)
			S = F32(0)
			for k in range(C):
				S = F32(S + e[k])
			thr = F32(_f32(u[b]) * S)
			acc = F32(0)
			pick = C - 1
			for k in range(C):
				acc = F32(acc + e[k])
				if acc > thr:
					pick = k
					break
			idx[b] = pick
		return idx

	# -- wavenet.py:724-911  the incremental loop ------------------------------------------
	def incremental(self, c=None, g=None, noise=None, test_inputs=None, time_length=None,
			softmax=False, literal_queue=False, initial_input=None, return_states=False):
		"""Fast-wavenet loop.

		c [B, Tc, 80] already normalised (what Synthesizer feeds), g ids [B,1] or None,
		noise [T, B, NZ] (NZ=1 Gaussian eps / categorical uniform; K+1 for MoL), test_inputs [B, T] scalar
		teacher (or int classes for mulaw-quantize).  Returns (outputs, raw):
		  outputs [B, C, T]  - sampled values ([B,1,T]) or one-hots ([B,256,T])   (wavenet.py:911)
		  raw     [B, O, T] (scalar input) or [B, T, C] (mulaw-quantize)          (tower_y_hat_eval, :904-908)
		literal_queue=True shifts/concats the whole queue every step exactly like modules.py:285-292; False uses the
		mathematically identical ring indexing (tested equal).
		"""
		hp = self.hp
		R, L, k = hp.residual_channels, hp.layers, hp.kernel_size
		if c is not None:
			c = _f32(c)
			B = c.shape[0]
			c_up = self.upsample(np.transpose(c, [0, 2, 1]))	#wavenet.py:427, 781-803
			T_c = c_up.shape[-1]
			if time_length is None:
				time_length = T_c
			c_btc = np.transpose(c_up, [0, 2, 1])	#self.c, wavenet.py:801
		else:
			assert time_length is not None
			B = len(test_inputs) if test_inputs is not None else (len(g) if g is not None else (np.shape(noise)[1] if noise is not None else 1))
			c_btc = None
		if test_inputs is not None:
			test_inputs = np.asarray(test_inputs)
			time_length = max(time_length, test_inputs.shape[1])	#wavenet.py:765-768
		if c_btc is not None:
			assert c_btc.shape[1] == time_length, 'upsampled length {} != time_length {} (wavenet.py:800)'.format(
				c_btc.shape[1], time_length)
		T = int(time_length)
		g_emb = self._global_embedding(g, B)

		current = self.initial_input(B) if initial_input is None else _f32(initial_input)
		qlen = [k + (k - 1) * (d - 1) for d in self.dilations]	#wavenet.py:815
		if literal_queue:
			queues = [np.zeros((B, qlen[l], R), F32) for l in range(L)]
		else:
			hist = [np.zeros((B, qlen[l], R), F32) for l in range(L)]	#ring: slot t % qlen
		O = hp.out_channels
		raw = np.zeros((T, B, O), F32)
		outs = np.zeros((T, B, current.shape[1]), F32)
		lin = [self._lin('ResidualConv1DGLU_{}/residual_block_causal_conv'.format(l)) for l in range(L)]

		for t in range(T):
			ct = c_btc[:, t, :] if c_btc is not None else None	#wavenet.py:823
			x = self._dense(current, 'input_convolution')	#wavenet.py:826
			skips = None
			for l in range(L):
				p = 'ResidualConv1DGLU_{}/'.format(l)
				d = self.dilations[l]
				residual = x
				#modules.py:273-297 (dropout is identity at inference, :484)
				if k > 1:
					if literal_queue:
						q = np.concatenate([queues[l][:, 1:, :], x[:, None, :]], axis=1)
						queues[l] = q
						taps = q[:, 0::d, :] if d > 1 else q
					else:
						h = hist[l]
						h[:, t % qlen[l], :] = x
						idx = [(t - (k - 1 - j) * d) for j in range(k)]
						taps = np.stack([h[:, i % qlen[l], :] if i >= 0 else np.zeros((B, R), F32) for i in idx], axis=1)
					inp = taps.reshape(B, -1)
				else:
					inp = x
				pre = inp @ lin[l]
				b = self._bias(p + 'residual_block_causal_conv')
				if b is not None:
					pre = pre + b
				half = pre.shape[1] // 2
				a, bg = pre[:, :half], pre[:, half:]	#modules.py:494
				if ct is not None:
					cc = self._dense(ct, p + 'residual_block_cin_conv')	#modules.py:497-501
					a, bg = a + cc[:, :half], bg + cc[:, half:]
				if g_emb is not None:
					gg = self._dense(g_emb, p + 'residual_block_gin_conv')	#modules.py:504-508
					a, bg = a + gg[:, :half], bg + gg[:, half:]
				z = (np.tanh(a) * (F32(1) / (F32(1) + np.exp(-bg)))).astype(F32)	#modules.py:510
				s = self._dense(z, p + 'residual_block_skip_conv')	#modules.py:512
				o = self._dense(z, p + 'residual_block_out_conv')	#modules.py:515
				x = ((o + residual) * SQRT_HALF).astype(F32) if hp.residual_legacy else (o + residual).astype(F32)	#:517-520
				if hp.legacy:	#wavenet.py:833-836
					skips = s if skips is None else ((skips + s) * SQRT_HALF).astype(F32)
				else:
					skips = s if skips is None else (skips + s).astype(F32)
			x = skips
			x = np.maximum(x, F32(0))	#wavenet.py:840-844, head
			x = self._dense(x, 'final_convolution_1')
			x = np.maximum(x, F32(0))
			x = self._dense(x, 'final_convolution_2')
			raw[t] = x	#wavenet.py:847

			nz = noise[t] if noise is not None else None
			if self.scalar_input:
				if O == 2:
					smp = self.sample_gaussian(x, nz[:, 0])	#wavenet.py:851-854
				else:
					smp = self.sample_mol(x, nz)	#wavenet.py:856-857
				nxt = smp[:, None]
				outs[t] = nxt
			else:
				probs = x
				if softmax:
					e = np.exp(x - np.max(x, axis=1, keepdims=True))
					probs = (e / np.sum(e, axis=1, keepdims=True)).astype(F32)
				idx = self.sample_categorical(probs, nz[:, 0])
				nxt = np.eye(hp.quantize_channels, dtype=F32)[idx]
				outs[t] = nxt
			if test_inputs is not None:	#wavenet.py:877-878
				if self.scalar_input:
					nxt = _f32(test_inputs[:, t])[:, None]
				else:
					nxt = np.eye(hp.quantize_channels, dtype=F32)[np.asarray(test_inputs[:, t]).astype(np.int64)]
			current = nxt

		outputs = np.transpose(outs, [1, 2, 0])	#wavenet.py:911
		if is_mulaw_quantize(hp.input_type):
			raw_out = np.transpose(raw, [1, 0, 2])	#wavenet.py:905-906
		else:
			raw_out = np.transpose(raw, [1, 2, 0])	#:907-908
		if return_states:
			return outputs, raw_out, c_btc
		return outputs, raw_out

	# -- wavenet.py:407-456  synthesis branch of initialize() -------------------------------
	def synthesize(self, c, g=None, noise=None, test_inputs=None, literal_queue=False):
		"""c [B,Tc,80] normalised -> (wav [B,T] f32, raw params, upsampled features [B,80,T])."""
		hp = self.hp
		outputs, raw, c_btc = self.incremental(c=c, g=g, noise=noise, test_inputs=test_inputs,
			softmax=False, literal_queue=literal_queue, return_states=True)
		B = outputs.shape[0]
		if is_mulaw_quantize(hp.input_type):
			y = np.argmax(outputs, axis=1).reshape(B, -1)
			y = inv_mulaw_quantize(y, hp.quantize_channels).astype(F32)
		elif is_mulaw(hp.input_type):
			y = inv_mulaw(outputs.reshape(B, -1), hp.quantize_channels).astype(F32)
		else:
			y = outputs.reshape(B, -1)
		return y, raw, (np.transpose(c_btc, [0, 2, 1]) if c_btc is not None else None)

	# -- wavenet.py:650-721  teacher-forced full-sequence pass -------------------------------
	def _conv_full(self, x, name, d):
		"""CausalConv1D.call normal branch (modules.py:305-325), channels_first.  x [B,Cin,T] -> [B,Cout,T]."""
		kern = self.w[name + '/kernel']	#[kw, Cin, Cout]
		kw = kern.shape[0]
		pad = (kw - 1) * d
		xp = np.pad(x, [(0, 0), (0, 0), (pad, 0)])
		T = x.shape[-1]
		y = np.zeros((x.shape[0], kern.shape[2], T), F32)
		for j in range(kw):
			#VALID dilated correlation: out[t] = sum_j xp[t + j*d] * K[j]
			y += np.einsum('bct,co->bot', xp[:, :, j * d: j * d + T], kern[j]).astype(F32)
		b = self._bias(name)
		return (y + b[None, :, None]).astype(F32) if b is not None else y

	def step(self, x, c=None, g=None, softmax=False):
		"""x [B, C_in, T] (scalar: C_in=1; mulaw-quantize: one-hot), c [B,80,Tc] -> [B, O, T].  Dropout is treated as
		rate 0 (the reference always applies it in step(), modules.py:484; logit equivalence needs it off)."""
		hp = self.hp
		x = _f32(x)
		B, _, T = x.shape
		g_emb = self._global_embedding(g, B)
		g_bct = None if g_emb is None else np.tile(g_emb[:, :, None], [1, 1, T])	#wavenet.py:14-51
		if c is not None:
			c = self.upsample(_f32(c))
			assert c.shape[-1] == T	#wavenet.py:697
		x = self._conv_full(x, 'input_convolution', 1)
		skips = None
		for l in range(hp.layers):
			p = 'ResidualConv1DGLU_{}/'.format(l)
			residual = x
			y = self._conv_full(x, p + 'residual_block_causal_conv', self.dilations[l])[:, :, :T]
			half = y.shape[1] // 2
			a, b = y[:, :half], y[:, half:]	#splitdim=1, modules.py:489-494
			if c is not None:
				cc = self._conv_full(c, p + 'residual_block_cin_conv', 1)
				a, b = a + cc[:, :half], b + cc[:, half:]
			if g_bct is not None:
				gg = self._conv_full(g_bct, p + 'residual_block_gin_conv', 1)
				a, b = a + gg[:, :half], b + gg[:, half:]
			z = (np.tanh(a) * (F32(1) / (F32(1) + np.exp(-b)))).astype(F32)
			s = self._conv_full(z, p + 'residual_block_skip_conv', 1)
			o = self._conv_full(z, p + 'residual_block_out_conv', 1)
			x = ((o + residual) * SQRT_HALF).astype(F32) if hp.residual_legacy else (o + residual).astype(F32)
			if skips is None:	#wavenet.py:707-715
				skips = s
			else:
				skips = skips + s
				if hp.legacy:
					skips = (skips * SQRT_HALF).astype(F32)
		x = np.maximum(skips, F32(0))
		x = self._conv_full(x, 'final_convolution_1', 1)
		x = np.maximum(x, F32(0))
		x = self._conv_full(x, 'final_convolution_2', 1)
		if softmax:
			e = np.exp(x - np.max(x, axis=1, keepdims=True))
			x = (e / np.sum(e, axis=1, keepdims=True)).astype(F32)
		return x
