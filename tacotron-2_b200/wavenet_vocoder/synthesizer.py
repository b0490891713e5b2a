"""``Synthesizer`` - the drop-in boundary of the hot path (reference ``code/wavenet_vocoder/synthesizer.py``).

Same ``load(checkpoint_path, hparams, model_name)`` / ``synthesize(mel_spectrograms, speaker_ids, basenames,
out_dir, log_dir)`` contract: a list of un-normalised ``[frames, num_mels]`` mels in, peak-normalised int16 wav files
out.  Where the reference does one ``session.run`` on a TF graph (:97) this does one ``wn_synthesize_host`` call into
libwavenet_b200.so.  Differences, all stated in DESIGN.md: weights come from an ``.npz`` weight table or
``random:<seed>`` or a TensorFlow V2 checkpoint prefix (wavenet_vocoder/tf_checkpoint.py; see its STATUS note); PNG plots are skipped (no matplotlib here,
out of scope); an optional ``seed`` makes the draws reproducible (the reference's are unseeded).
"""
import os

import numpy as np

from datasets.audio import get_hop_size, save_wavenet_wav
from infolog import log
from wavenet_vocoder import weights as wn_weights
from wavenet_vocoder.models import create_model
from wavenet_vocoder.models.wavenet import make_noise


def _interp(feats, in_range):
	#rescales from [-max, max] (or [0, max]) to [0, 1]  (feeder.py:426-428)
	return (feats - in_range[0]) / (in_range[1] - in_range[0])


def _pad_inputs(x, maxlen, _pad=0):
	return np.pad(x, [(0, maxlen - len(x)), (0, 0)], mode='constant', constant_values=_pad)


def resolve_weights(checkpoint_path, hparams):
	"""'random:<seed>[:kind]' -> deterministic random init; '*.npz' -> weight table; anything else is a TF checkpoint,
	read without TensorFlow by wavenet_vocoder/tf_checkpoint.py."""
	if isinstance(checkpoint_path, dict):
		return checkpoint_path
	if checkpoint_path is None or str(checkpoint_path).startswith('random'):
		parts = str(checkpoint_path or 'random').split(':')
		seed = int(parts[1]) if len(parts) > 1 and parts[1] else hparams.wavenet_random_seed
		kind = parts[2] if len(parts) > 2 else 'tf-init'
		return wn_weights.init_weights(hparams, seed=seed, kind=kind)
	if str(checkpoint_path).endswith('.npz') and os.path.exists(checkpoint_path):
		return wn_weights.load_weights(checkpoint_path)
	if os.path.exists(str(checkpoint_path) + '.index'):      # a TensorFlow V2 checkpoint prefix (train.py:67-86)
		from wavenet_vocoder.tf_checkpoint import load_tf_checkpoint
		return load_tf_checkpoint(str(checkpoint_path), hparams)
	raise RuntimeError('Failed to load checkpoint at {} (expected a TensorFlow checkpoint prefix, an .npz weight table or '
		'random:<seed>)'.format(checkpoint_path))


class Synthesizer:
	"""Multi-GPU (reference: in-graph towers over a tf.split of the batch, wavenet.py:227-239, lists flattened on the host,
	synthesizer.py:100-101).  Utterances never interact, so they are sharded, with wavenet_vocoder.sharding:
	  * under torchrun (torch.distributed initialised, one rank per GPU): every rank calls synthesize() with the same batch,
	    synthesises its shard and rank 0 gathers the audio (the only collective of the path) and writes the files;
	  * in one process with hparams.wavenet_num_gpus > 1: one model per visible device, one host thread per device
	    (the library call releases the GIL), results merged in utterance order.
	The draws of an utterance depend on the seed and on its index in the batch only, never on the number of GPUs."""

	def load(self, checkpoint_path, hparams, model_name='WaveNet', engine='auto'):
		log('Constructing model: {}'.format(model_name))
		self._hparams = hparams
		self.synth_debug = hparams.wavenet_synth_debug
		weights = resolve_weights(checkpoint_path, hparams)
		log('Loading checkpoint: {}'.format(checkpoint_path))
		self.models = []
		n_towers = self._tower_count()
		if n_towers > 1:
			import torch
			keep = torch.cuda.current_device()
			for d in range(n_towers):
				torch.cuda.set_device(d)	#wn_create binds the handle to the current device
				self.models.append(create_model(model_name, hparams, engine=engine).load_weights(weights))
			torch.cuda.set_device(keep)
		else:
			self.models.append(create_model(model_name, hparams, engine=engine).load_weights(weights))
		self.model = self.models[0]
		self._rng = np.random.default_rng(hparams.get('wavenet_synthesis_seed'))
		return self

	def _tower_count(self):
		n = int(self._hparams.get('wavenet_num_gpus') or 1)
		if n <= 1 or self._dist()[1] > 1:
			return 1
		import torch
		return max(1, min(n, torch.cuda.device_count()))

	@staticmethod
	def _dist():
		"""(rank, world size) of the torch.distributed job this process belongs to, (0, 1) outside one."""
		try:
			import torch.distributed as dist
			if dist.is_available() and dist.is_initialized():
				return dist.get_rank(), dist.get_world_size()
		except ImportError:
			pass
		return 0, 1

	@property
	def is_writer(self):
		return self._dist()[0] == 0

	def prepare_batch(self, mel_spectrograms):
		"""synthesizer.py:56-70: true lengths, clip, pad with the range minimum, stack, rescale to [0, 1]."""
		hp = self._hparams
		audio_lengths = [len(x) * get_hop_size(hp) for x in mel_spectrograms]
		maxlen = max([len(x) for x in mel_spectrograms])
		T2_output_range = (-hp.max_abs_value, hp.max_abs_value) if hp.symmetric_mels else (0, hp.max_abs_value)
		if hp.clip_for_wavenet:
			mel_spectrograms = [np.clip(x, T2_output_range[0], T2_output_range[1]) for x in mel_spectrograms]
		c_batch = np.stack([_pad_inputs(x, maxlen, _pad=T2_output_range[0]) for x in mel_spectrograms]).astype(np.float32)
		if hp.normalize_for_wavenet:
			c_batch = _interp(c_batch, T2_output_range).astype(np.float32)
		return c_batch, audio_lengths

	def synthesize(self, mel_spectrograms, speaker_ids, basenames, out_dir, log_dir, seed=None, noise=None,
			return_wavs=False):
		hp = self._hparams
		local_cond, global_cond = self._check_conditions()
		if not local_cond:	#synthesizer.py:76-77: a fixed 100 samples per utterance, no mels
			B = len(mel_spectrograms)
			rng = self._rng if seed is None else np.random.default_rng(seed)
			nz = make_noise(hp, 100, B, rng) if noise is None else noise
			g = None if (speaker_ids is None or not global_cond) else np.asarray(speaker_ids, dtype=np.int32).reshape(B, 1)
			wavs = list(self.model.initialize(None, None, g, None, synthesis_length=100, noise=nz, want_raw=False))
			if return_wavs:
				return wavs
			return self._write(wavs, basenames, out_dir)

		if self.synth_debug:	#synthesizer.py:51-53
			assert len(hp.wavenet_debug_mels) == len(hp.wavenet_debug_wavs)
			mel_spectrograms = [np.load(mel_file) for mel_file in hp.wavenet_debug_mels]

		c_batch, audio_lengths = self.prepare_batch(mel_spectrograms)
		g = None if (speaker_ids is None or not global_cond) else np.asarray(speaker_ids, dtype=np.int32).reshape(len(c_batch), 1)

		teacher = None
		if self.synth_debug:	#synthesizer.py:83-94 teacher forcing with the target waveforms
			test_wavs = [np.load(debug_wav).reshape(-1, 1) for debug_wav in hp.wavenet_debug_wavs]
			max_test_len = max([len(x) for x in test_wavs])
			teacher = np.stack([_pad_inputs(x, max_test_len) for x in test_wavs]).astype(np.float32)

		B, Tc = c_batch.shape[0], c_batch.shape[1]
		T = Tc * get_hop_size(hp)
		if noise is None:
			rng = self._rng if seed is None else np.random.default_rng(seed)
			noise = make_noise(hp, T, B, rng)

		#Generate wavs and clip extra padding to select Real speech parts  (synthesizer.py:97-104)
		generated_wavs = self._run_sharded(c_batch, g, teacher, noise, audio_lengths)
		if generated_wavs is None:	#a non-writing rank of a torch.distributed job
			return []
		if return_wavs:
			return generated_wavs

		return self._write(generated_wavs, basenames, out_dir)

	def _write(self, generated_wavs, basenames, out_dir):
		hp = self._hparams
		audio_filenames = []
		for i, generated_wav in enumerate(generated_wavs):
			#Save wav to disk (plots of the reference, :115-125, are out of scope)
			audio_filename = os.path.join(out_dir, 'wavenet-audio-{}.wav'.format(basenames[i]))
			save_wavenet_wav(generated_wav, audio_filename, sr=hp.sample_rate, inv_preemphasize=hp.preemphasize, k=hp.preemphasis)
			audio_filenames.append(audio_filename)
		return audio_filenames

	def _run_sharded(self, c_batch, g, teacher, noise, audio_lengths):
		"""The session.run of the reference (:97) over however many GPUs take part; returns the cropped wavs in utterance
		order (None on the ranks of a distributed job that do not hold the result)."""
		from wavenet_vocoder import sharding
		rank, world = self._dist()
		models = getattr(self, 'models', None) or [self.model]
		towers = len(models)

		def run(model, idx):
			if not idx:
				return []
			sel = np.asarray(idx)
			frames = max(audio_lengths[i] for i in idx) // get_hop_size(self._hparams)   # a shard stops at its own longest utterance
			T = frames * get_hop_size(self._hparams)
			out = model.initialize(None, np.ascontiguousarray(c_batch[sel, :frames]), None if g is None else g[sel], None,
				test_inputs=None if teacher is None else teacher[sel], noise=np.ascontiguousarray(noise[:T, sel]), want_raw=False)
			return [out[k, :audio_lengths[i]] for k, i in enumerate(idx)]

		if world > 1:
			shards = sharding.shard_utterances(audio_lengths, world)
			local = run(self.model, shards[rank])
			return sharding.gather_wavs(local, shards[rank], len(audio_lengths))
		if towers > 1:
			import torch
			from concurrent.futures import ThreadPoolExecutor
			shards = sharding.shard_utterances(audio_lengths, towers)

			def tower(d):
				torch.cuda.set_device(d)
				return run(models[d], shards[d])
			with ThreadPoolExecutor(max_workers=towers) as pool:
				parts = list(pool.map(tower, range(towers)))
			wavs = [None] * len(audio_lengths)
			for d in range(towers):
				for i, w in zip(shards[d], parts[d]):
					wavs[i] = w
			return wavs
		return run(self.model, list(range(len(audio_lengths))))

	def _check_conditions(self):
		local_condition = self._hparams.cin_channels > 0
		global_condition = self._hparams.gin_channels > 0
		return local_condition, global_condition
