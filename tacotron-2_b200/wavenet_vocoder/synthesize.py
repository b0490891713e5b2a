"""WaveNet synthesis driver: what `synthesize.py --model='WaveNet'` runs (reference code/wavenet_vocoder/synthesize.py:12-78).

Contract kept from the reference: mels are the `*.npy` files of --mels_dir (or the rows `text|mel_path|speaker_id` of
Tacotron's `map.txt` when --model='Tacotron-2', `<no_g>` meaning no speaker ids), they are synthesised in chunks of
`hparams.wavenet_synthesis_batch_size`, wavs land in `<output_dir>/wavs/` and `wavs/map.txt` lists `mel_path|wav_path`
(`text|mel_path|wav_path` for Tacotron-2 input).  The checkpoint argument accepts everything `resolve_checkpoint` names.
Under torchrun every rank runs this driver; Synthesizer.synthesize shards each chunk over the ranks and rank 0 writes.
"""
import os
import re
import time

import numpy as np

from hparams import get_hop_size, hparams_debug_string
from infolog import log
from wavenet_vocoder.synthesizer import Synthesizer


def resolve_checkpoint(checkpoint):
	"""The weights a run should load, from what the command line gave (reference :69-78 asks tf.train.get_checkpoint_state):
	  random[:seed[:kind]]          deterministic random initialisation
	  <file>.npz                    a weight table
	  <prefix> with <prefix>.index  a TensorFlow V2 checkpoint prefix (model.ckpt-1000)
	  <dir>                         its TF `checkpoint` state file (model_checkpoint_path: "..."), else its newest
	                                *.index prefix, else its newest *.npz"""
	if checkpoint is None:
		return None
	ck = str(checkpoint)
	if ck.startswith('random') or os.path.isfile(ck) or os.path.exists(ck + '.index'):
		return ck
	if os.path.isdir(ck):
		state = os.path.join(ck, 'checkpoint')
		if os.path.isfile(state):
			m = re.search(r'^model_checkpoint_path:\s*"(.*)"\s*$', open(state).read(), flags=re.M)
			if m:
				path = m.group(1) if os.path.isabs(m.group(1)) else os.path.join(ck, m.group(1))
				if os.path.exists(path + '.index') or os.path.isfile(path):
					return path
		newest = lambda suffix: sorted((os.path.join(ck, f) for f in os.listdir(ck) if f.endswith(suffix)), key=os.path.getmtime)
		if newest('.index'):
			return newest('.index')[-1][:-len('.index')]
		if newest('.npz'):
			return newest('.npz')[-1]
	return None


def _collect_inputs(args):
	"""(mel paths, speaker ids or None, texts or None)"""
	if args.model == 'Tacotron-2':
		with open(os.path.join(args.mels_dir, 'map.txt'), encoding='utf-8') as f:
			rows = [line.strip().split('|') for line in f if line.strip()]
		texts, mels, speakers = [r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows]
		return mels, (None if all(s == '<no_g>' for s in speakers) else speakers), texts
	mels = sorted(os.path.join(args.mels_dir, f) for f in os.listdir(args.mels_dir) if f.endswith('.npy'))
	speakers = None
	if args.speaker_id is not None:
		speakers = args.speaker_id.replace(' ', '').split(',')
		assert len(speakers) == len(mels), '--speaker_id needs one id per mel file'
	return mels, speakers, None


def run_synthesis(args, checkpoint_path, output_dir, hparams):
	wav_dir, log_dir = os.path.join(output_dir, 'wavs'), os.path.join(output_dir, 'plots')
	log(hparams_debug_string(hparams))
	synth = Synthesizer()
	synth.load(checkpoint_path, hparams)
	mels, speakers, texts = _collect_inputs(args)
	os.makedirs(wav_dir, exist_ok=True)
	os.makedirs(log_dir, exist_ok=True)
	log('Starting synthesis! (this will take a while..)')
	step, hop = hparams.wavenet_synthesis_batch_size, get_hop_size(hparams)
	rows = []
	for lo in range(0, len(mels), step):
		chunk = mels[lo:lo + step]
		spectros = [np.load(path) for path in chunk]
		names = [os.path.basename(path).replace('.npy', '') for path in chunk]
		t0 = time.time()
		wavs = synth.synthesize(spectros, None if speakers is None else speakers[lo:lo + step], names, wav_dir, log_dir)
		took, n = time.time() - t0, sum(len(m) for m in spectros) * hop
		log('Generation of {} audio frames took {:.3f} sec ({:.3f} frames/sec, {:.1f} x real time)'.format(
			n, took, n / took, n / took / hparams.sample_rate))
		for j, wav in enumerate(wavs):	#empty on the ranks that do not write
			rows.append('|'.join(([texts[lo + j]] if texts is not None else []) + [chunk[j], wav]))
	if synth.is_writer:
		with open(os.path.join(wav_dir, 'map.txt'), 'w') as f:
			f.write(''.join(r + '\n' for r in rows))
	log('synthesized audio waveforms at {}'.format(wav_dir))


def wavenet_synthesize(args, hparams, checkpoint):
	checkpoint_path = resolve_checkpoint(checkpoint)
	if checkpoint_path is None:
		raise RuntimeError('Failed to load checkpoint at {}'.format(checkpoint))
	log('loaded model at {}'.format(checkpoint_path))
	run_synthesis(args, checkpoint_path, 'wavenet_' + args.output_dir, hparams)
