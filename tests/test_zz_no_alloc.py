"""wn_upsample_static (include/wavenet_b200.h): the upsampling network with its intermediates in the library's
file-scope device array, for callers that cannot allocate.  Called through the C ABI and compared with wn_upsample
(the existing entry point) bit for bit - the tolerance of tests/test_gpu_parity.py for the upsampling network - on two
calls in a row (the array is shared by all calls and keeps the previous call's contents).

Collected last on purpose: a failure here hides nothing after it.
"""
import ctypes as C

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


def _both(net, c):
    """(wn_upsample, wn_upsample_static) outputs for conditioning c [B, Tc, M], through ctypes."""
    import torch
    B, Tc, M = c.shape
    dc = torch.from_numpy(c).cuda()
    st = torch.cuda.current_stream().cuda_stream
    want = torch.zeros((B, Tc * net.hop, M), dtype=torch.float32, device='cuda')
    got = torch.full_like(want, float('nan'))
    rc = net.lib.wn_upsample(net._h, C.c_void_p(dc.data_ptr()), B, Tc, C.c_void_p(want.data_ptr()), C.c_void_p(st))
    assert rc == 0, net.lib.wn_last_error(net._h)
    rc = net.lib.wn_upsample_static(net._h, C.c_void_p(dc.data_ptr()), B, Tc, C.c_void_p(got.data_ptr()), C.c_void_p(st))
    assert rc == 0, net.lib.wn_last_error(net._h)
    torch.cuda.synchronize()
    return want, got


@pytest.mark.parametrize('cfg', ['default', 'paper', 'resize_small'])
def test_static_scratch_upsample_matches_wn_upsample_twice(cfg):
    import torch
    from wavenet_vocoder._native import NativeWaveNet
    hp = {'default': common.default_22k, 'paper': common.paper,
          'resize_small': lambda: common.small(upsample_type='Resize', upsample_activation='LeakyRelu')}[cfg]()
    net = NativeWaveNet(hp, common.weights_for(hp))
    rng = np.random.default_rng(11)
    # two calls in a row with different shapes and data: the second one finds the first one's values in the array
    for B, Tc in ((3, 9), (2, 5)):
        c = rng.uniform(0, 1, (B, Tc, hp.num_mels)).astype(np.float32)
        want, got = _both(net, c)
        assert bool(torch.isfinite(got).all())
        assert torch.equal(got, want)          # bit for bit
    net.close()


def test_static_scratch_rejects_oversize_and_reports_the_limit():
    import torch
    from wavenet_vocoder import _native
    hp = common.default_22k()
    net = _native.NativeWaveNet(hp, common.weights_for(hp))
    s0 = hp.upsample_scales[0]
    Tc = _native.STATIC_SCRATCH_FLOATS // (s0 * hp.num_mels) + 1      # one column too many at B = 1
    c = torch.zeros((1, Tc, hp.num_mels), device='cuda')
    out = torch.zeros((1, 8, hp.num_mels), device='cuda')            # never written: the call must fail before any launch
    rc = net.lib.wn_upsample_static(net._h, C.c_void_p(c.data_ptr()), 1, Tc, C.c_void_p(out.data_ptr()), None)
    assert rc == -1
    msg = net.lib.wn_last_error(net._h).decode()
    assert 'static scratch' in msg and str(_native.STATIC_SCRATCH_FLOATS) in msg
    # the largest size that fits still runs and still matches
    Tc_ok = _native.STATIC_SCRATCH_FLOATS // (s0 * hp.num_mels)
    cc = np.random.default_rng(3).uniform(0, 1, (1, Tc_ok, hp.num_mels)).astype(np.float32)
    want, got = _both(net, cc)
    assert torch.equal(got, want)
    net.close()
