"""Discretized mixture-of-logistics likelihood (reference: wavenet_vocoder/models/mixture.py:5-74), host side, numpy.

The reference evaluates this loss in TensorFlow on the distribution parameters the network returns (`y_hat`, what
`WaveNet.step` / `tower_y_hat_eval` produce).  Here the parameters come from the CUDA path and the scalar metric is
computed on the host in float64 - it is an evaluation figure (a reference-defined quality number for
comparing weight / arithmetic formats), not part of the synthesis hot path.  Same argument names, shapes and return
conventions as the reference; sampling from the mixture lives in the CUDA loop (csrc/wn_loop_*.cu)."""
import numpy as np


def log_sum_exp(x):
	"""numerically stable log-sum-exp over the last axis (mixture.py:5-10)"""
	x = np.asarray(x, np.float64)
	m = np.max(x, axis=-1)
	return m + np.log(np.sum(np.exp(x - m[..., None]), axis=-1))


def log_prob_from_logits(x):
	"""numerically stable log-softmax over the last axis (mixture.py:12-16)"""
	x = np.asarray(x, np.float64)
	m = np.max(x, axis=-1, keepdims=True)
	return x - m - np.log(np.sum(np.exp(x - m), axis=-1, keepdims=True))


def _softplus(x):
	return np.logaddexp(0.0, x)


def _sigmoid(x):
	return np.exp(-_softplus(-x))


def discretized_mix_logistic_loss(y_hat, y, num_classes=256, log_scale_min=-7.0, reduce=True):
	"""y_hat [B, 3K, T] (logits | means | log-scales), y [B, T, 1] in [-1, 1] -> summed negative log-likelihood
	(reduce=True) or [B, T, 1] per-sample values (mixture.py:18-74)."""
	y_hat = np.asarray(y_hat, np.float64)
	y = np.asarray(y, np.float64)
	if y_hat.ndim != 3 or y_hat.shape[1] % 3 != 0:
		raise ValueError('y_hat must be [batch, 3 * mixtures, time], got {}'.format(y_hat.shape))
	nr_mix = y_hat.shape[1] // 3
	y_hat = np.transpose(y_hat, [0, 2, 1])
	logit_probs = y_hat[:, :, :nr_mix]
	means = y_hat[:, :, nr_mix:2 * nr_mix]
	log_scales = np.maximum(y_hat[:, :, 2 * nr_mix:3 * nr_mix], log_scale_min)
	y = y * np.ones((1, 1, nr_mix))
	centered_y = y - means
	inv_stdv = np.exp(-log_scales)
	plus_in = inv_stdv * (centered_y + 1. / (num_classes - 1))
	cdf_plus = _sigmoid(plus_in)
	min_in = inv_stdv * (centered_y - 1. / (num_classes - 1))
	cdf_min = _sigmoid(min_in)
	log_cdf_plus = plus_in - _softplus(plus_in)            # edge case y = -1 (class 0)
	log_one_minus_cdf_min = -_softplus(min_in)             # edge case y = +1 (last class)
	cdf_delta = cdf_plus - cdf_min
	mid_in = inv_stdv * centered_y
	log_pdf_mid = mid_in - log_scales - 2. * _softplus(mid_in)
	log_probs = np.where(y < -0.999, log_cdf_plus,
		np.where(y > 0.999, log_one_minus_cdf_min,
			np.where(cdf_delta > 1e-5, np.log(np.maximum(cdf_delta, 1e-12)), log_pdf_mid - np.log((num_classes - 1) / 2))))
	log_probs = log_probs + log_prob_from_logits(logit_probs)
	if reduce:
		return float(-np.sum(log_sum_exp(log_probs)))
	return -log_sum_exp(log_probs)[..., None]
