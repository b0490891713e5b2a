"""The 'paper' WaveNet configuration (reference ``code/paper_hparams.py:175-231, 290-335``).

The reference never imports that file (every entry point does ``from hparams import hparams``;
its README says to use it by substitution), and it lacks ``split_on_cpu`` which
``wavenet.py:227`` reads.  Here it is an alternative ``hparams`` object that every entry point accepts:
24 layers / 4 stacks, R=256 G=512 S=256, 10-component mixture of logistics, ConvTranspose2D
upsampling [5, 5, 11], no legacy scaling, 22.05 kHz / hop 275.
"""
import numpy as np

from hparams import HParams, _wavenet_block_default


def _wavenet_block_paper():
	d = _wavenet_block_default()
	d.update(
		hop_size=275,
		win_size=1100,
		sample_rate=22050,
		legacy=False,
		residual_legacy=False,
		log_scale_min=float(np.log(1e-14)),
		log_scale_min_gauss=float(np.log(9.1188196 * 1e-4)),
		cdf_loss=True,
		out_channels=30,
		layers=24,
		stacks=4,
		residual_channels=256,
		gate_channels=512,
		skip_out_channels=256,
		kernel_size=3,
		cin_channels=80,
		upsample_type='2D',
		upsample_activation='Relu',
		upsample_scales=[5, 5, 11],
		freq_axis_kernel_size=3,
		leaky_alpha=0.4,
		NN_init=True,
		NN_scaler=0.1,
		gin_channels=-1,
		use_speaker_embedding=True,
		n_speakers=5,
	)
	return d


hparams = HParams(**_wavenet_block_paper())
