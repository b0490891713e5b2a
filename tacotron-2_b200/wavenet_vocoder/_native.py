"""ctypes binding of libwavenet_b200.so (include/wavenet_b200.h) - the only way into the CUDA kernels.

There is no CPU path: if the shared library is missing or no B200 is visible this module raises, loudly
(``NativeLibraryError``), instead of computing anything on the host.
"""
import ctypes as C
import os

import numpy as np

from wavenet_vocoder.util import is_scalar_input

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('WN_B200_LIB') or os.path.join(_HERE, 'libwavenet_b200.so')   # override: development builds

WN_ABI_VERSION = 2
MAX_UP = 8
INPUT_TYPES = {'raw': 0, 'mulaw': 1, 'mulaw-quantize': 2}
UP_TYPES = {'SubPixel': 0, '2D': 1, '1D': 2, 'Resize': 3, 'NearestNeighbor': 4}
ACTS = {None: 0, 'Relu': 1, 'LeakyRelu': 2}
ENGINES = {'auto': 0, 'generic': 1, 'pipeline': 2, 'tc': 3}
ENGINE_NAMES = {v: k for k, v in ENGINES.items()}

EXPORTS = ('wn_create', 'wn_load_weights', 'wn_finalize', 'wn_upsample', 'wn_synthesize', 'wn_synthesize_host',
	'wn_step_full', 'wn_query', 'wn_last_error', 'wn_destroy', 'wn_upsample_static')
STATIC_SCRATCH_FLOATS = 4 * 1024 * 1024   # WN_STATIC_SCRATCH_FLOATS of include/wavenet_b200.h


class NativeLibraryError(RuntimeError):
	pass


class WnConfig(C.Structure):
	_fields_ = [
		('abi_version', C.c_int32), ('input_type', C.c_int32), ('quantize_channels', C.c_int32),
		('out_channels', C.c_int32), ('layers', C.c_int32), ('stacks', C.c_int32),
		('residual_channels', C.c_int32), ('gate_channels', C.c_int32), ('skip_out_channels', C.c_int32),
		('kernel_size', C.c_int32), ('use_bias', C.c_int32), ('legacy', C.c_int32), ('residual_legacy', C.c_int32),
		('log_scale_min', C.c_float), ('log_scale_min_gauss', C.c_float),
		('cin_channels', C.c_int32), ('upsample_type', C.c_int32), ('upsample_activation', C.c_int32),
		('n_upsample_scales', C.c_int32), ('upsample_scales', C.c_int32 * MAX_UP),
		('freq_axis_kernel_size', C.c_int32), ('leaky_alpha', C.c_float), ('nn_init', C.c_int32),
		('hop_size', C.c_int32), ('gin_channels', C.c_int32), ('use_speaker_embedding', C.c_int32),
		('n_speakers', C.c_int32), ('engine', C.c_int32), ('ksplit', C.c_int32)]


class WnInfo(C.Structure):
	_fields_ = [
		('receptive_field', C.c_int32), ('n_params', C.c_int64), ('flops_per_sample', C.c_int64),
		('bytes_per_sample', C.c_int64), ('weight_bytes', C.c_int64), ('engine', C.c_int32), ('ksplit', C.c_int32),
		('pipeline_ctas', C.c_int32), ('pipeline_ctas_per_layer', C.c_int32), ('pipeline_group_size', C.c_int32),
		('sm_count', C.c_int32), ('kernel_launches', C.c_int64),
		('last_upsample_ms', C.c_float), ('last_loop_ms', C.c_float), ('aborted', C.c_int32)]


_lib = None


def load_library(path=None):
	"""dlopen the library and declare the prototypes.  Raises NativeLibraryError when it is missing."""
	global _lib
	if _lib is not None and path is None:
		return _lib
	path = path or LIB_PATH
	if not os.path.exists(path):
		raise NativeLibraryError(
			'{} not found: build it with `python tacotron-2_b200/csrc/build.py` (or __graft_entry__.build()). '
			'There is no CPU fallback for the WaveNet synthesis path.'.format(path))
	try:
		lib = C.CDLL(path)
	except OSError as e:
		raise NativeLibraryError('cannot load {}: {}'.format(path, e))
	vp, ip, fp = C.c_void_p, C.c_int, C.c_void_p
	lib.wn_create.restype = ip
	lib.wn_create.argtypes = [C.POINTER(WnConfig), C.POINTER(vp)]
	lib.wn_load_weights.restype = ip
	lib.wn_load_weights.argtypes = [vp, C.c_char_p, fp, C.POINTER(C.c_int64), ip]
	lib.wn_finalize.restype = ip
	lib.wn_finalize.argtypes = [vp]
	lib.wn_upsample.restype = ip
	lib.wn_upsample.argtypes = [vp, fp, ip, ip, fp, vp]
	lib.wn_upsample_static.restype = ip
	lib.wn_upsample_static.argtypes = [vp, fp, ip, ip, fp, vp]
	lib.wn_synthesize.restype = ip
	lib.wn_synthesize.argtypes = [vp, fp, fp, ip, ip, fp, fp, fp, fp, fp, vp]
	lib.wn_synthesize_host.restype = ip
	lib.wn_synthesize_host.argtypes = [vp, fp, fp, ip, ip, fp, fp, fp, fp, vp]
	lib.wn_step_full.restype = ip
	lib.wn_step_full.argtypes = [vp, fp, fp, fp, ip, ip, fp, vp]
	lib.wn_query.restype = ip
	lib.wn_query.argtypes = [vp, C.POINTER(WnInfo)]
	lib.wn_last_error.restype = C.c_char_p
	lib.wn_last_error.argtypes = [vp]
	lib.wn_destroy.restype = None
	lib.wn_destroy.argtypes = [vp]
	if path == LIB_PATH:
		_lib = lib
	return lib


def config_from_hparams(hp, engine='auto', ksplit=0):
	cfg = WnConfig()
	cfg.abi_version = WN_ABI_VERSION
	cfg.input_type = INPUT_TYPES[hp.input_type]
	cfg.quantize_channels = hp.quantize_channels
	cfg.out_channels = hp.out_channels
	cfg.layers, cfg.stacks = hp.layers, hp.stacks
	cfg.residual_channels, cfg.gate_channels = hp.residual_channels, hp.gate_channels
	cfg.skip_out_channels, cfg.kernel_size = hp.skip_out_channels, hp.kernel_size
	cfg.use_bias, cfg.legacy, cfg.residual_legacy = int(hp.use_bias), int(hp.legacy), int(hp.residual_legacy)
	cfg.log_scale_min, cfg.log_scale_min_gauss = hp.log_scale_min, hp.log_scale_min_gauss
	cfg.cin_channels = hp.cin_channels
	cfg.upsample_type = UP_TYPES[hp.upsample_type]
	cfg.upsample_activation = ACTS[hp.upsample_activation]
	scales = list(hp.upsample_scales)
	cfg.n_upsample_scales = len(scales)
	for i, s in enumerate(scales[:MAX_UP]):
		cfg.upsample_scales[i] = int(s)
	cfg.freq_axis_kernel_size = hp.freq_axis_kernel_size
	cfg.leaky_alpha = hp.leaky_alpha
	cfg.nn_init = int(hp.NN_init)
	hop = hp.hop_size if hp.hop_size is not None else int(hp.frame_shift_ms / 1000 * hp.sample_rate)
	cfg.hop_size = hop
	cfg.gin_channels = hp.gin_channels
	cfg.use_speaker_embedding = int(hp.use_speaker_embedding)
	cfg.n_speakers = hp.n_speakers if hp.n_speakers is not None else 0
	cfg.engine = ENGINES[engine]
	cfg.ksplit = int(ksplit)
	return cfg


def _ptr(x):
	"""device pointer of a torch tensor / int, host pointer of a numpy array, None -> NULL."""
	if x is None:
		return None
	if isinstance(x, int):
		return C.c_void_p(x)
	if isinstance(x, np.ndarray):
		return C.c_void_p(x.ctypes.data)
	return C.c_void_p(x.data_ptr())


class NativeWaveNet(object):
	"""One model on the current CUDA device (one wn_handle)."""

	def __init__(self, hp, weights, engine='auto', ksplit=0):
		self.lib = load_library()
		self.hp = hp
		self._h = C.c_void_p()
		cfg = config_from_hparams(hp, engine, ksplit)
		rc = self.lib.wn_create(C.byref(cfg), C.byref(self._h))
		if rc != 0:
			msg = self.lib.wn_last_error(None).decode()
			self._h = C.c_void_p()
			if rc == -1:
				raise ValueError(msg)
			raise NativeLibraryError('wn_create failed ({}): {}'.format(rc, msg))
		for name, arr in weights.items():
			a = np.ascontiguousarray(arr, dtype=np.float32)
			shape = (C.c_int64 * a.ndim)(*a.shape)
			self._check(self.lib.wn_load_weights(self._h, name.encode(), C.c_void_p(a.ctypes.data), shape, a.ndim))
		self._check(self.lib.wn_finalize(self._h))
		self.noise_channels = (hp.out_channels // 3 + 1) if (is_scalar_input(hp.input_type) and hp.out_channels != 2) else 1
		self.hop = cfg.hop_size if hp.cin_channels > 0 else 1

	def _check(self, rc):
		if rc != 0:
			msg = self.lib.wn_last_error(self._h).decode()
			if rc == -1:
				raise ValueError(msg)
			raise RuntimeError('libwavenet_b200 error {}: {}'.format(rc, msg))

	def close(self):
		if self._h:
			self.lib.wn_destroy(self._h)
			self._h = C.c_void_p()

	def __del__(self):
		try:
			self.close()
		except Exception:
			pass

	def info(self):
		inf = WnInfo()
		self._check(self.lib.wn_query(self._h, C.byref(inf)))
		d = {f[0]: getattr(inf, f[0]) for f in WnInfo._fields_}
		d['engine'] = ENGINE_NAMES.get(d['engine'], d['engine'])
		return d

	# -- device-pointer calls (torch CUDA tensors or raw ints); asynchronous on `stream` ----------------
	def upsample(self, c, B, Tc, c_up, stream=0):
		self._check(self.lib.wn_upsample(self._h, _ptr(c), B, Tc, _ptr(c_up), C.c_void_p(stream)))

	def upsample_static(self, c, B, Tc, c_up, stream=0):
		"""wn_upsample without allocation (intermediates in the library's static device array; see the header)."""
		self._check(self.lib.wn_upsample_static(self._h, _ptr(c), B, Tc, _ptr(c_up), C.c_void_p(stream)))

	def synthesize(self, c, B, Tc, noise, audio, g_ids=None, teacher=None, raw=None, c_up=None, stream=0):
		self._check(self.lib.wn_synthesize(self._h, _ptr(c), _ptr(g_ids), B, Tc, _ptr(noise), _ptr(teacher),
			_ptr(audio), _ptr(raw), _ptr(c_up), C.c_void_p(stream)))

	def step_full(self, x, c, B, Tc, raw, g_ids=None, stream=0):
		self._check(self.lib.wn_step_full(self._h, _ptr(x), _ptr(c), _ptr(g_ids), B, Tc, _ptr(raw), C.c_void_p(stream)))

	# -- host-buffer call (numpy in / numpy out, copies inside; synchronous) ---------------------------
	def synthesize_host(self, c, noise, g_ids=None, teacher=None, want_raw=False, stream=0, audio_out=None):
		"""c [B,Tc,cin] float32 (or None with local conditioning off), noise [T,B,NZ] -> audio [B,T] (, raw [B,T,O])."""
		hp = self.hp
		noise = np.ascontiguousarray(noise, np.float32)
		T, B, NZ = noise.shape
		if NZ != self.noise_channels:
			raise ValueError('noise must be [T, B, {}], got {}'.format(self.noise_channels, noise.shape))
		if c is not None:
			c = np.ascontiguousarray(c, np.float32)
			if c.ndim != 3 or c.shape[0] != B or c.shape[2] != hp.cin_channels:
				raise ValueError('Expected 3 dimension shape [batch_size({}), time_length, {}] for local condition '
					'features but found {}'.format(B, hp.cin_channels, c.shape))
			Tc = c.shape[1]
			if Tc * self.hop != T:
				raise ValueError('noise covers {} steps but the conditioning upsamples to {}'.format(T, Tc * self.hop))
		else:
			Tc = T
		g = None if g_ids is None else np.ascontiguousarray(np.asarray(g_ids).reshape(B), np.int32)
		if teacher is not None:
			teacher = np.ascontiguousarray(teacher, np.float32)
			assert teacher.shape == (B, T), teacher.shape
		audio = audio_out if audio_out is not None else np.empty((B, T), np.float32)
		raw = np.empty((B, T, hp.out_channels), np.float32) if want_raw else None
		self._check(self.lib.wn_synthesize_host(self._h, _ptr(c), _ptr(g), B, Tc, _ptr(noise), _ptr(teacher),
			_ptr(audio), _ptr(raw), C.c_void_p(stream)))
		return (audio, raw) if want_raw else audio
