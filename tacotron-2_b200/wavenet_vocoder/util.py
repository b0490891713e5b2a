"""input_type predicates and mu-law helpers (reference ``code/wavenet_vocoder/util.py:13-129``).

numpy only: in the reference these functions are numpy-or-TF polymorphic; the in-graph (TF) uses
of the synthesis path are done on the device by the head of the loop kernels (csrc/wn_loop_tc.cu, wn_loop_pipe.cu, wn_loop_generic.cu), the numpy
uses stay here.  ``mu`` is hard-wired to 255 whatever is passed, exactly like the reference
(:48, :67, :99, :127)."""
import numpy as np


def _assert_valid_input_type(s):
	assert s == 'mulaw-quantize' or s == 'mulaw' or s == 'raw'


def is_mulaw_quantize(s):
	_assert_valid_input_type(s)
	return s == 'mulaw-quantize'


def is_mulaw(s):
	_assert_valid_input_type(s)
	return s == 'mulaw'


def is_raw(s):
	_assert_valid_input_type(s)
	return s == 'raw'


def is_scalar_input(s):
	return is_raw(s) or is_mulaw(s)


def mulaw(x, mu=256):
	mu = 255
	return np.sign(x) * np.log1p(mu * np.abs(x)) / np.log1p(mu)


def inv_mulaw(y, mu=256):
	mu = 255
	return np.sign(y) * (1.0 / mu) * ((1.0 + mu) ** np.abs(y) - 1.0)


def mulaw_quantize(x, mu=256):
	mu = 255
	y = mulaw(x, mu)
	#[-1, 1] -> [0, mu], truncated
	y = (y + 1) / 2 * mu
	return y.astype(np.int64) if isinstance(y, np.ndarray) else int(y)


def inv_mulaw_quantize(y, mu=256):
	mu = 255
	y = 2 * np.asarray(y).astype(np.float32) / mu - 1
	return inv_mulaw(y, mu)


def sequence_mask(input_lengths, max_len=None, expand=True):
	input_lengths = np.asarray(input_lengths)
	if max_len is None:
		max_len = int(input_lengths.max())
	mask = (np.arange(max_len)[None, :] < input_lengths[:, None]).astype(np.float32)
	return mask[..., None] if expand else mask
