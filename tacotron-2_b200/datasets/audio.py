"""Host-side audio trivia of the WaveNet synthesis path (reference ``code/datasets/audio.py``):
only ``get_hop_size`` (:54-59) and ``save_wavenet_wav`` (:17-20) are on the path.
The STFT / mel / Griffin-Lim part of that file is preprocessing and out of scope."""
import struct

import numpy as np


def get_hop_size(hparams):
	hop_size = hparams.hop_size
	if hop_size is None:
		assert hparams.frame_shift_ms is not None
		hop_size = int(hparams.frame_shift_ms / 1000 * hparams.sample_rate)
	return hop_size


def float_to_int16_peak_normalized(wav):
	"""The sample conversion of save_wavenet_wav: scale so the peak hits 32767 (peak floored at
	0.01 so silence is not blown up), then truncate toward zero to int16."""
	wav = np.asarray(wav, dtype=np.float32)
	peak = np.max(np.abs(wav)) if wav.size else 0.
	wav = wav * (32767 / max(0.01, peak))
	return wav.astype(np.int16)


def write_wav_int16(path, sr, samples):
	"""Minimal RIFF/WAVE PCM16 mono writer (what scipy.io.wavfile.write emits for an int16 vector)."""
	samples = np.ascontiguousarray(samples, dtype='<i2')
	data = samples.tobytes()
	header = b'RIFF' + struct.pack('<I', 36 + len(data)) + b'WAVE'
	header += b'fmt ' + struct.pack('<IHHIIHH', 16, 1, 1, int(sr), int(sr) * 2, 2, 16)
	header += b'data' + struct.pack('<I', len(data))
	with open(path, 'wb') as f:
		f.write(header)
		f.write(data)


def save_wavenet_wav(wav, path, sr, inv_preemphasize=None, k=None):
	"""Same signature as the reference; inverse pre-emphasis is disabled there too (commented out, :18)."""
	write_wav_int16(path, sr, float_to_int16_peak_normalized(wav))
