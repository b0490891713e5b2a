"""Development aid: BASELINE.json configs[3] at full length - paper_hparams.py, 16 utterances x 30 s (2406 frames -> 661 650
samples each), device-resident buffers, one pass through wn_synthesize; prints the loop time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'tacotron-2_b200'), os.path.join(ROOT, 'tests')]
import numpy as np
import torch
import common
from wavenet_vocoder._native import NativeWaveNet
from wavenet_vocoder.models.wavenet import make_noise

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
Tc = int(sys.argv[2]) if len(sys.argv) > 2 else 2406
hp = common.paper()
net = NativeWaveNet(hp, common.weights_for(hp, kind='tf-init'))
T = Tc * hp.hop_size
d_c = torch.from_numpy(common.cond_for(hp, B, Tc)).cuda()
d_noise = torch.from_numpy(make_noise(hp, T, B, np.random.default_rng(1))).cuda()
d_audio = torch.empty((B, T), dtype=torch.float32, device='cuda')
t0 = time.time()
net.synthesize(d_c, B, Tc, d_noise, d_audio, stream=torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
info = net.info()
print('paper_hparams B={} x {:.1f} s (T={}): engine {} wall {:.1f} s, loop {:.1f} ms = {:.1f} us/step, {:.0f} samples/s, RTF {:.2f}, finite {}, aborted {}'.format(
    B, T / hp.sample_rate, T, info['engine'], time.time() - t0, info['last_loop_ms'], 1e3 * info['last_loop_ms'] / T,
    B * T / (info['last_loop_ms'] / 1e3), B * T / (info['last_loop_ms'] / 1e3) / hp.sample_rate, bool(torch.isfinite(d_audio).all()), info['aborted']))
