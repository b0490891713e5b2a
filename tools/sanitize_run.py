"""Development aid: small runs of every hand-rolled inter-CTA protocol, for
  compute-sanitizer --tool racecheck|synccheck python tools/sanitize_run.py
(the tensor-core loop engine, the pipelined fp32 engine, the cluster generic engine, the fused tcgen05 step kernel)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'tacotron-2_b200'), os.path.join(ROOT, 'tests')]
import numpy as np
import torch
import common
from wavenet_vocoder._native import NativeWaveNet

which = sys.argv[1:] or ['tc', 'pipeline', 'generic', 'step']
hp = common.default_22k()
w = common.weights_for(hp)
for name in which:
    if name in ('tc', 'pipeline'):
        B, Tc = 5, 1
        net = NativeWaveNet(hp, w, engine=name)
        T = Tc * hp.hop_size
        c, nz = common.cond_for(hp, B, Tc), common.noise_for(hp, T, B)
        a = net.synthesize_host(c, nz)
        print(name, 'engine', net.info()['engine'], 'finite', bool(np.isfinite(a).all()), 'aborted', net.info()['aborted'], flush=True)
        net.close()
    elif name == 'generic':
        hs = common.small()
        ws = common.weights_for(hs)
        net = NativeWaveNet(hs, ws, engine='generic')
        T = 2 * hs.hop_size
        a = net.synthesize_host(common.cond_for(hs, 3, 2), common.noise_for(hs, T, 3))
        print('generic finite', bool(np.isfinite(a).all()), flush=True)
        net.close()
    elif name == 'step':
        net = NativeWaveNet(hp, w)
        B, Tc = 1, 2
        T = Tc * hp.hop_size
        dx = torch.from_numpy(common.teacher_for(hp, B, T)).cuda()
        dc = torch.from_numpy(common.cond_for(hp, B, Tc)).cuda()
        raw = torch.empty((B, T, hp.out_channels), dtype=torch.float32, device='cuda')
        net.step_full(dx, dc, B, Tc, raw, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        print('step finite', bool(torch.isfinite(raw).all()), flush=True)
        net.close()
