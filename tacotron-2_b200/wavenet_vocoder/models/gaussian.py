"""Single-Gaussian likelihood (reference: wavenet_vocoder/models/gaussian.py:5-37), host side, numpy float64.

`y_hat` are the distribution parameters the network returns (mean, log-scale); see mixture.py for why this metric is
evaluated on the host.  Sampling from the Gaussian lives in the CUDA loop (csrc/wn_loop_*.cu)."""
import math

import numpy as np


def _normal_cdf(x, loc, scale):
	z = (x - loc) / (scale * math.sqrt(2.0))
	return 0.5 * (1.0 + np.vectorize(math.erf, otypes=[np.float64])(z))


def gaussian_maximum_likelihood_estimation_loss(y_hat, y, log_scale_min_gauss, num_classes, use_cdf=True, reduce=True):
	"""y_hat [B, 2, T] (mean | log-scale), y [B, T, 1] -> summed negative log-likelihood (reduce=True) or [B, T, 1]
	(gaussian.py:5-37).  use_cdf: probability mass of the quantisation bin around y (the reference's "CDF trick"),
	else the log density."""
	y_hat = np.asarray(y_hat, np.float64)
	if y_hat.ndim != 3 or y_hat.shape[1] != 2:
		raise ValueError('y_hat must be [batch, 2, time], got {}'.format(y_hat.shape))
	y_hat = np.transpose(y_hat, [0, 2, 1])
	mean = y_hat[:, :, 0]
	log_scale = np.maximum(y_hat[:, :, 1], log_scale_min_gauss)
	y = np.asarray(y, np.float64)[..., 0]
	if use_cdf:
		scale = np.exp(log_scale)
		cdf_plus = _normal_cdf(y + 1. / (num_classes - 1), mean, scale)
		cdf_min = _normal_cdf(y - 1. / (num_classes - 1), mean, scale)
		log_prob = np.log(np.maximum(cdf_plus - cdf_min, 1e-12))
	else:
		log_prob = -0.5 * (np.log(2. * np.pi) + 2. * log_scale + np.square(y - mean) * np.exp(-2. * log_scale))
	if reduce:
		return float(-np.sum(log_prob))
	return -log_prob[..., None]
