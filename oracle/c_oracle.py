"""TEST INFRASTRUCTURE - ctypes front end of oracle/c/wn_oracle.c (the fixed-accumulation-order C restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os

import numpy as np

from . import build_oracle

MAX_LAYERS = 64
MAX_UP = 8
F32P = C.POINTER(C.c_float)

INPUT_TYPES = {'raw': 0, 'mulaw': 1, 'mulaw-quantize': 2}
UP_TYPES = {'SubPixel': 0, '2D': 1, '1D': 2, 'Resize': 3, 'NearestNeighbor': 4}
ACTS = {None: 0, 'Relu': 1, 'LeakyRelu': 2}


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        'layers', 'residual_channels', 'gate_channels', 'skip_channels', 'out_channels', 'kernel_size',
        'cin_channels', 'gin_channels', 'in_channels', 'input_type', 'legacy', 'residual_legacy', 'use_bias')] + [
        ('log_scale_min', C.c_float), ('log_scale_min_gauss', C.c_float),
        ('dilations', C.c_int32 * MAX_LAYERS),
        ('upsample_type', C.c_int32), ('upsample_activation', C.c_int32), ('n_upsample', C.c_int32),
        ('freq_axis_kernel_size', C.c_int32), ('hop_size', C.c_int32),
        ('upsample_scales', C.c_int32 * MAX_UP), ('leaky_alpha', C.c_float), ('ksplit', C.c_int32)]


class Weights(C.Structure):
    _fields_ = [('first_w', F32P), ('first_b', F32P)] + [
        (n, F32P * MAX_LAYERS) for n in ('conv_w', 'conv_b', 'cin_w', 'cin_b', 'gin_w', 'gin_b',
                                         'out_w', 'out_b', 'skip_w', 'skip_b')] + [
        ('head1_w', F32P), ('head1_b', F32P), ('head2_w', F32P), ('head2_b', F32P),
        ('up_w', F32P * MAX_UP), ('up_b', F32P * MAX_UP)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = build_oracle.build()
        _lib = C.CDLL(path)
        for name in ('wno_incremental', 'wno_incremental_shiftqueue'):
            fn = getattr(_lib, name)
            fn.restype = C.c_int
            fn.argtypes = [C.POINTER(Config), C.POINTER(Weights), F32P, F32P, F32P, F32P, C.c_int, C.c_int,
                           F32P, F32P, F32P, C.c_int]
        _lib.wno_upsample.restype = C.c_int
        _lib.wno_upsample.argtypes = [C.POINTER(Config), C.POINTER(Weights), F32P, C.c_int, C.c_int, F32P, C.c_int]
        _lib.wno_math.restype = None
        _lib.wno_math.argtypes = [C.c_int, F32P, F32P, C.c_int64]
    return _lib


def _ptr(a):
    return a.ctypes.data_as(F32P) if a is not None else None


def _hop(hp):
    return hp.hop_size if hp.hop_size is not None else int(hp.frame_shift_ms / 1000 * hp.sample_rate)


class COracle(object):
    """Same inputs as oracle.wavenet_oracle.WaveNetOracle (hp + weight dict in TF layouts)."""

    def __init__(self, hp, weights, ksplit=1):
        """ksplit: number of blocks of the out / skip 1x1 reductions - pass NativeWaveNet.info()['ksplit'] to mirror
        the CUDA plan bit for bit (the value only changes the fp32 summation order)."""
        self.hp = hp
        self._keep = []
        cfg = Config()
        cfg.layers = hp.layers
        cfg.residual_channels = hp.residual_channels
        cfg.gate_channels = hp.gate_channels
        cfg.skip_channels = hp.skip_out_channels
        cfg.out_channels = hp.out_channels
        cfg.kernel_size = hp.kernel_size
        cfg.cin_channels = hp.cin_channels
        cfg.gin_channels = hp.gin_channels
        cfg.in_channels = 1 if hp.input_type != 'mulaw-quantize' else hp.quantize_channels
        cfg.input_type = INPUT_TYPES[hp.input_type]
        cfg.legacy = int(hp.legacy)
        cfg.residual_legacy = int(hp.residual_legacy)
        cfg.use_bias = int(hp.use_bias)
        cfg.log_scale_min = hp.log_scale_min
        cfg.log_scale_min_gauss = hp.log_scale_min_gauss
        per_stack = hp.layers // hp.stacks
        for l in range(hp.layers):
            cfg.dilations[l] = 2 ** (l % per_stack)
        cfg.upsample_type = UP_TYPES[hp.upsample_type]
        cfg.upsample_activation = ACTS[hp.upsample_activation]
        learnable = hp.cin_channels > 0 and hp.upsample_type != 'NearestNeighbor'
        cfg.n_upsample = len(hp.upsample_scales) if learnable else 0
        cfg.freq_axis_kernel_size = hp.freq_axis_kernel_size
        cfg.hop_size = _hop(hp)
        for i, s in enumerate(hp.upsample_scales[:MAX_UP]):
            cfg.upsample_scales[i] = s
        cfg.leaky_alpha = hp.leaky_alpha
        cfg.ksplit = int(ksplit)
        self.cfg = cfg

        w = Weights()

        def arr(name):
            if name not in weights:
                return None
            a = np.ascontiguousarray(weights[name], dtype=np.float32)
            self._keep.append(a)
            return _ptr(a)

        w.first_w, w.first_b = arr('input_convolution/kernel'), arr('input_convolution/bias')
        for l in range(hp.layers):
            p = 'ResidualConv1DGLU_{}/residual_block_'.format(l)
            for field, nm in (('conv', 'causal_conv'), ('cin', 'cin_conv'), ('gin', 'gin_conv'),
                              ('out', 'out_conv'), ('skip', 'skip_conv')):
                getattr(w, field + '_w')[l] = arr(p + nm + '/kernel')
                getattr(w, field + '_b')[l] = arr(p + nm + '/bias')
        w.head1_w, w.head1_b = arr('final_convolution_1/kernel'), arr('final_convolution_1/bias')
        w.head2_w, w.head2_b = arr('final_convolution_2/kernel'), arr('final_convolution_2/bias')
        for i in range(cfg.n_upsample):
            kname = 'local_conditioning_upsampling_{}/kernel'.format(i + 1)
            kern = np.asarray(weights[kname], np.float32)
            if hp.upsample_type == 'SubPixel' and not hp.NN_init:  # modules.py:584-592
                kern = np.tile(kern[:, :, :, 0:1], [1, 1, 1, kern.shape[3]])
            kern = np.ascontiguousarray(kern)
            self._keep.append(kern)
            w.up_w[i] = _ptr(kern)
            w.up_b[i] = arr('local_conditioning_upsampling_{}/bias'.format(i + 1))
        self.w = w
        self.emb = np.ascontiguousarray(weights['gc_embedding'], np.float32) if 'gc_embedding' in weights else None

    @property
    def noise_channels(self):
        hp = self.hp
        if hp.input_type != 'mulaw-quantize' and hp.out_channels != 2:
            return hp.out_channels // 3 + 1
        return 1

    def upsample(self, c, nthreads=0):
        """c [B, Tc, cin] -> [B, T, cin] (time-major, unlike the numpy oracle's [B, cin, T])."""
        c = np.ascontiguousarray(c, np.float32)
        B, Tc, cin = c.shape
        out = np.empty((B, Tc * self.cfg.hop_size, cin), np.float32)
        rc = lib().wno_upsample(C.byref(self.cfg), C.byref(self.w), _ptr(c), B, Tc, _ptr(out), nthreads)
        if rc:
            raise RuntimeError('wno_upsample failed: {}'.format(rc))
        return out

    def incremental(self, c_up=None, g=None, noise=None, teacher=None, T=None, want_raw=True, nthreads=0,
                    shiftqueue=False):
        """c_up [B,T,cin] upsampled conditioning (or None), g ids [B] -> (audio [B,T], samples [B,T], raw [B,T,O])."""
        hp = self.hp
        if c_up is not None:
            c_up = np.ascontiguousarray(c_up, np.float32)
            B, T = c_up.shape[0], c_up.shape[1]
        else:
            B = teacher.shape[0] if teacher is not None else noise.shape[1]
        g_emb = None
        if g is not None and hp.gin_channels > 0:
            g_emb = np.ascontiguousarray(self.emb[np.asarray(g).reshape(B).astype(np.int64)], np.float32) \
                if self.emb is not None else np.ascontiguousarray(g, np.float32).reshape(B, -1)
        noise = np.ascontiguousarray(noise, np.float32) if noise is not None else None
        if noise is not None:
            assert noise.shape == (T, B, self.noise_channels), (noise.shape, (T, B, self.noise_channels))
        teacher = np.ascontiguousarray(teacher, np.float32) if teacher is not None else None
        audio = np.empty((B, T), np.float32)
        samples = np.empty((B, T), np.float32)
        raw = np.empty((B, T, hp.out_channels), np.float32) if want_raw else None
        fn = lib().wno_incremental_shiftqueue if shiftqueue else lib().wno_incremental
        rc = fn(C.byref(self.cfg), C.byref(self.w), _ptr(c_up), _ptr(g_emb), _ptr(noise), _ptr(teacher), B, T,
                _ptr(audio), _ptr(samples), _ptr(raw), nthreads)
        if rc:
            raise RuntimeError('wno_incremental failed: {}'.format(rc))
        return audio, samples, raw

    def synthesize(self, c, g=None, noise=None, teacher=None, nthreads=0, shiftqueue=False):
        """c [B,Tc,cin] normalised mels -> (audio, samples, raw, c_up)."""
        c_up = self.upsample(c, nthreads)
        audio, samples, raw = self.incremental(c_up, g, noise, teacher, nthreads=nthreads, shiftqueue=shiftqueue)
        return audio, samples, raw, c_up


def math_fn(fn, x):
    x = np.ascontiguousarray(x, np.float32)
    y = np.empty_like(x)
    lib().wno_math({'exp': 0, 'log': 1, 'tanh': 2, 'sigmoid': 3, 'inv_mulaw': 4}[fn], _ptr(x), _ptr(y), x.size)
    return y
