"""Top-level synthesis CLI with the reference's flags (``code/synthesize.py:47-96``), WaveNet part only.

  python synthesize.py --model='WaveNet' --mels_dir=<dir of [frames, 80] .npy> --checkpoint=<weights.npz | random:5339> \
      --hparams='sample_rate=22050,hop_size=275'

``--model='Tacotron'`` (spectrogram prediction) is a different model and out of scope; ``--model='Tacotron-2'`` runs
only its second half: it reads the ``map.txt`` Tacotron left in ``--mels_dir`` (``tacotron/synthesize.py:53-83``).
"""
import argparse
import os

from hparams import hparams
from infolog import log
from wavenet_vocoder.synthesize import wavenet_synthesize


def prepare_run(args):
	modified_hp = hparams.parse(args.hparams)
	run_name = args.name or args.wavenet_name or args.model
	wave_checkpoint = os.path.join('logs-' + run_name, 'wave_' + args.checkpoint)
	if args.checkpoint.startswith('random') or os.path.exists(args.checkpoint):
		wave_checkpoint = args.checkpoint
	return wave_checkpoint, modified_hp


def main(argv=None):
	accepted_modes = ['eval', 'synthesis', 'live']
	parser = argparse.ArgumentParser()
	parser.add_argument('--checkpoint', default='pretrained/', help='Path to model checkpoint')
	parser.add_argument('--hparams', default='',
		help='Hyperparameter overrides as a comma-separated list of name=value pairs')
	parser.add_argument('--name', help='Name of logging directory if the two models were trained together.')
	parser.add_argument('--tacotron_name', help='Name of logging directory of Tacotron. If trained separately')
	parser.add_argument('--wavenet_name', help='Name of logging directory of WaveNet. If trained separately')
	parser.add_argument('--model', default='Tacotron-2')
	parser.add_argument('--input_dir', default='training_data/', help='folder to contain inputs sentences/targets')
	parser.add_argument('--mels_dir', default='tacotron_output/eval/', help='folder to contain mels to synthesize audio from using the Wavenet')
	parser.add_argument('--output_dir', default='output/', help='folder to contain synthesized mel spectrograms')
	parser.add_argument('--mode', default='eval', help='mode of run: can be one of {}'.format(accepted_modes))
	parser.add_argument('--GTA', default='True', help='Ground truth aligned synthesis, defaults to True, only considered in synthesis mode')
	parser.add_argument('--text_list', default='', help='Text file contains list of texts to be synthesized. Valid if mode=eval')
	parser.add_argument('--speaker_id', default=None, help='Defines the speakers ids to use when running standalone Wavenet on a folder of mels. this variable must be a comma-separated list of ids')
	args = parser.parse_args(argv)

	accepted_models = ['Tacotron', 'WaveNet', 'Tacotron-2']

	if args.model not in accepted_models:
		raise ValueError('please enter a valid model to synthesize with: {}'.format(accepted_models))

	if args.mode not in accepted_modes:
		raise ValueError('accepted modes are: {}, found {}'.format(accepted_modes, args.mode))

	if args.mode == 'live' and args.model == 'Wavenet':
		raise RuntimeError('Wavenet vocoder cannot be tested live due to its slow generation. Live only works with Tacotron!')

	if args.GTA not in ('True', 'False'):
		raise ValueError('GTA option must be either True or False')

	if args.model == 'Tacotron-2' and args.mode == 'synthesis':
		raise ValueError('I don\'t recommend running WaveNet on entire dataset.. The world might end before the synthesis :) (only eval allowed)')

	if args.model == 'Tacotron':
		raise NotImplementedError('Tacotron spectrogram prediction is a different model and out of scope of this build')

	wave_checkpoint, hp = prepare_run(args)
	wavenet_synthesize(args, hp, wave_checkpoint)
	if args.model == 'Tacotron-2':
		log('Tacotron-2 TTS synthesis complete!')


if __name__ == '__main__':
	main()
