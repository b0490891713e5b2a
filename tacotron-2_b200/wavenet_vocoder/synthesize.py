"""Batching driver of WaveNet synthesis (reference ``code/wavenet_vocoder/synthesize.py:12-78``): lists ``*.npy`` mels
(or reads Tacotron's ``map.txt``), chunks them by ``wavenet_synthesis_batch_size`` and writes ``wavs/map.txt``."""
import os
import time

import numpy as np

from hparams import hparams_debug_string
from infolog import log
from wavenet_vocoder.synthesizer import Synthesizer


def run_synthesis(args, checkpoint_path, output_dir, hparams):
	log_dir = os.path.join(output_dir, 'plots')
	wav_dir = os.path.join(output_dir, 'wavs')

	log(hparams_debug_string(hparams))
	synth = Synthesizer()
	synth.load(checkpoint_path, hparams)

	if args.model == 'Tacotron-2':
		#mels evaluated by Tacotron: map.txt rows are text|mel_path|speaker_id
		metadata_filename = os.path.join(args.mels_dir, 'map.txt')
		with open(metadata_filename, encoding='utf-8') as f:
			metadata = np.array([line.strip().split('|') for line in f])
		speaker_ids = metadata[:, 2]
		mel_files = metadata[:, 1]
		texts = metadata[:, 0]
		speaker_ids = None if (speaker_ids == '<no_g>').all() else speaker_ids
	else:
		mel_files = sorted([os.path.join(args.mels_dir, f) for f in os.listdir(args.mels_dir) if f.split('.')[-1] == 'npy'])
		speaker_ids = None if args.speaker_id is None else args.speaker_id.replace(' ', '').split(',')
		if speaker_ids is not None:
			assert len(speaker_ids) == len(mel_files)
		texts = None

	log('Starting synthesis! (this will take a while..)')
	os.makedirs(log_dir, exist_ok=True)
	os.makedirs(wav_dir, exist_ok=True)

	bs = hparams.wavenet_synthesis_batch_size
	mel_files = [mel_files[i: i + bs] for i in range(0, len(mel_files), bs)]
	speaker_ids = None if speaker_ids is None else [speaker_ids[i: i + bs] for i in range(0, len(speaker_ids), bs)]
	texts = None if texts is None else [texts[i: i + bs] for i in range(0, len(texts), bs)]

	with open(os.path.join(wav_dir, 'map.txt'), 'w') as file:
		for i, mel_batch in enumerate(mel_files):
			mel_spectros = [np.load(mel_file) for mel_file in mel_batch]
			basenames = [os.path.basename(mel_file).replace('.npy', '') for mel_file in mel_batch]
			speaker_id_batch = None if speaker_ids is None else speaker_ids[i]
			start = time.time()
			audio_files = synth.synthesize(mel_spectros, speaker_id_batch, basenames, wav_dir, log_dir)
			n = sum(len(m) for m in mel_spectros) * hparams.hop_size
			log('Generation of {} audio frames took {:.3f} sec ({:.3f} frames/sec)'.format(n, time.time() - start, n / (time.time() - start)))
			speaker_logs = ['<no_g>'] * len(mel_batch) if speaker_id_batch is None else speaker_id_batch
			for j, mel_file in enumerate(mel_batch):
				if texts is None:
					file.write('{}|{}\n'.format(mel_file, audio_files[j], speaker_logs[j]))
				else:
					file.write('{}|{}|{}\n'.format(texts[i][j], mel_file, audio_files[j], speaker_logs[j]))

	log('synthesized audio waveforms at {}'.format(wav_dir))


def wavenet_synthesize(args, hparams, checkpoint):
	output_dir = 'wavenet_' + args.output_dir
	checkpoint_path = None
	if checkpoint is not None and str(checkpoint).startswith('random'):
		checkpoint_path = checkpoint
	elif checkpoint is not None and os.path.isfile(str(checkpoint)):
		checkpoint_path = checkpoint
	elif checkpoint is not None and os.path.isdir(str(checkpoint)):
		#newest weight table in the directory plays the role of tf.train.get_checkpoint_state (:72-76)
		cands = sorted([os.path.join(checkpoint, f) for f in os.listdir(checkpoint) if f.endswith('.npz')], key=os.path.getmtime)
		checkpoint_path = cands[-1] if cands else None
	if checkpoint_path is None:
		raise RuntimeError('Failed to load checkpoint at {}'.format(checkpoint))
	log('loaded model at {}'.format(checkpoint_path))
	run_synthesis(args, checkpoint_path, output_dir, hparams)
