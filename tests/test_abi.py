"""The C-ABI library: loads, exports every symbol include/wavenet_b200.h declares, and refuses to run without a GPU.
No compute calls here (this file runs in the CPU container)."""
import ctypes
import os
import re

import pytest

import common
from wavenet_vocoder import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    text = open(os.path.join(ROOT, 'include', 'wavenet_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(wn_[a-z_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    lib = _native.load_library()
    names = _declared_functions()
    assert set(names) == set(_native.EXPORTS)
    for n in names:
        assert hasattr(lib, n), n


def test_config_struct_matches_header_field_order():
    text = open(os.path.join(ROOT, 'include', 'wavenet_b200.h')).read()
    body = re.search(r'typedef struct wn_config \{(.*?)\} wn_config;', text, flags=re.S).group(1)
    body = re.sub(r'/\*.*?\*/', '', body, flags=re.S)
    fields = []
    for decl in body.split(';'):
        decl = decl.strip()
        if not decl:
            continue
        names = decl.split(None, 1)[1]
        fields += [re.sub(r'\[.*\]', '', n).strip() for n in names.split(',')]
    assert fields == [f[0] for f in _native.WnConfig._fields_]
    assert ctypes.sizeof(_native.WnConfig) == 4 * (len(fields) - 1 + _native.MAX_UP)


def test_invalid_config_is_rejected_with_reference_message():
    lib = _native.load_library()
    hp = common.default_22k().override(hop_size=200)
    cfg = _native.config_from_hparams(hp)
    h = ctypes.c_void_p()
    rc = lib.wn_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == -1 and b'upsample_scales' in lib.wn_last_error(None)
    hp = common.default_22k().override(input_type='mulaw-quantize', quantize_channels=256, out_channels=30)
    cfg = _native.config_from_hparams(hp)
    rc = lib.wn_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == -1 and b'quantize_chennels' in lib.wn_last_error(None)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    hp = common.small()
    with pytest.raises(_native.NativeLibraryError, match='no CPU path'):
        _native.NativeWaveNet(hp, common.weights_for(hp))


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(_native.NativeLibraryError, match='no CPU fallback'):
        _native.load_library(str(tmp_path / 'libwavenet_b200.so'))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'tacotron-2_b200')
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(d, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), os.path.join(d, f)
                assert not re.search(r'#include\s+"[^"]*oracle', src), os.path.join(d, f)
                assert 'libwn_oracle' not in src, os.path.join(d, f)
