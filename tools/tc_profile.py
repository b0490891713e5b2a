"""Development aid: one call of the tensor-core loop engine inside a cudaProfilerStart/Stop range, for
  ncu --replay-mode app-range --set full --clock-control none python tools/tc_profile.py 64 4
(the loop kernel is a cooperative launch of thread-block clusters, which ncu's kernel replay cannot launch; range replay runs it natively)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'tacotron-2_b200'), os.path.join(ROOT, 'tests')]
import numpy as np
import torch
import common
from wavenet_vocoder._native import NativeWaveNet

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
Tc = int(sys.argv[2]) if len(sys.argv) > 2 else 4
engine = sys.argv[3] if len(sys.argv) > 3 else 'tc'
hp = common.paper() if (len(sys.argv) > 4 and sys.argv[4] == 'paper') else common.default_22k()   # argv[4]: 'paper' = the 256-channel block
net = NativeWaveNet(hp, common.weights_for(hp, kind='tf-init'), engine=engine)
T = Tc * hp.hop_size
d_c = torch.from_numpy(common.cond_for(hp, B, Tc)).cuda()
d_noise = torch.from_numpy(common.noise_for(hp, T, B)).cuda()
d_audio = torch.empty((B, T), dtype=torch.float32, device='cuda')
sp = torch.cuda.current_stream().cuda_stream
net.synthesize(d_c, B, Tc, d_noise, d_audio, stream=sp)      # warm-up: scratch allocation
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
net.synthesize(d_c, B, Tc, d_noise, d_audio, stream=sp)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
info = net.info()
assert bool(torch.isfinite(d_audio).all()) and not info['aborted']
print('engine {} B {} T {} loop {:.3f} ms = {:.2f} us/step, replicas {} x {} utterances'.format(
    info['engine'], B, T, info['last_loop_ms'], 1e3 * info['last_loop_ms'] / T, info['pipeline_ctas_per_layer'], info['pipeline_group_size']))
