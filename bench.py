"""bench.py - audio samples/s and real-time factor of WaveNet incremental (fast-wavenet) synthesis on B200.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference ...                     (the CPU arm: the oracle port on the host cores)

A "step" is one pass of the hot path over one batch of synthetic utterances: upsampling network + the autoregressive
loop + decode (wn_synthesize), random-init weights (seed 5339, 'tf-init'), synthetic mels and noise.
Workload: BASELINE.json configs[1] - default_22k hparams, 64 utterances x 5 s (401 frames -> 110 275 samples each)
PER GPU (weak scaling: utterances are independent, every rank synthesises its own 64 and rank 0 gathers the audio).
`--workload c3` runs configs[2] instead: 256 utterances x 10 s sharded over the ranks (strong scaling).

One JSON line on stdout (rank 0).  `value` is device-resident (inputs in HBM before the timed region, CUDA events on
the launch stream, max over ranks); `e2e` goes through the host-buffer C-ABI call wn_synthesize_host (pinned host
buffers, H2D + D2H inside the timed region).  Secondary objects on the same line, each measured outside the headline's timed
region: `strong_c3` (configs[2], 256 x 10 s over the ranks, host buffers + gather, at every N), and at N = 1 `step_full`
(configs[4], WaveNet.step on 8 x 10 s) and `c4_paper` (configs[3], paper_hparams, 16 utterances, 2 s sample).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# stdout carries exactly ONE JSON line.  Libraries print there too (NCCL's version banner under NCCL_DEBUG=VERSION, for
# one), so file descriptor 1 is pointed at stderr for the whole run and the JSON line is written to the saved descriptor.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + '\n').encode())

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, 'tacotron-2_b200')):
    if p not in sys.path:
        sys.path.insert(0, p)

SAMPLE_RATE = 22050
METRIC = 'audio samples/sec (real-time factor = value / 22050), WaveNet incremental synthesis'
UNIT = 'samples/s'
WORKLOADS = {
    # name: (utterances, frames, weak scaling?)
    'c2': (64, 401, True),    # BASELINE.json configs[1]: batch 64 x 5 s on 1 B200 (per GPU when N > 1)
    'c3': (256, 802, False),  # configs[2]: 256 x 10 s sharded over the ranks
    'c1': (1, 81, True),      # configs[0]: batch 1 x 1 s
    'tiny': (8, 8, True),     # development only
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='c2', choices=sorted(WORKLOADS))
    ap.add_argument('--engine', default='auto', choices=['auto', 'generic', 'pipeline', 'tc'])
    ap.add_argument('--hparams', default='default_22k', choices=['default_22k', 'paper'])
    ap.add_argument('--no-step-full', action='store_true', help='skip the secondary WaveNet.step (configs[4]) figure')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-c4', action='store_true', help='skip the configs[3] figure (paper_hparams, 16 utterances, 2 s sample)')
    ap.add_argument('--no-strong-c3', action='store_true', help='skip the configs[2] strong-scaling pass (256 x 10 s over the ranks)')
    return ap.parse_args()


def get_hparams(name):
    import hparams as hp_mod
    if name == 'paper':
        import paper_hparams
        return paper_hparams.hparams.copy()
    return hp_mod.default_22k()


def synth_inputs(hp, utt_ids, Tc, with_noise=True):
    """Synthetic normalised mels [B, Tc, 80] (seed 1234+u per utterance) and Gaussian / uniform noise [T, B, NZ]."""
    from wavenet_vocoder.models.wavenet import make_noise
    B = len(utt_ids)
    c = np.empty((B, Tc, hp.num_mels), np.float32)
    for i, u in enumerate(utt_ids):
        mel = np.random.default_rng(1234 + u).uniform(-4, 4, (Tc, hp.num_mels)).astype(np.float32)
        c[i] = (np.clip(mel, -4, 4) + 4) / 8
    T = Tc * hp.hop_size
    noise = make_noise(hp, T, B, np.random.default_rng(2024 + utt_ids[0])) if with_noise else None
    return c, noise


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (C restatement, OpenMP over utterances) on the box's host cores
# ---------------------------------------------------------------------------------------------------------------
def cpu_sample(hp, weights, cores, frames=None):
    """One bounded sample of the workload: `cores` utterances x `frames` frames through upsample + loop.
    Returns (samples, seconds)."""
    from oracle.c_oracle import COracle
    frames = frames or 24   # 24 frames = 6 600 samples/utterance, ~5 s per core at ~1.3 k samples/s/core
    co = COracle(hp, weights)
    c, noise = synth_inputs(hp, list(range(cores)), frames)
    t0 = time.perf_counter()
    co.synthesize(c, noise=noise, nthreads=cores)
    dt = time.perf_counter() - t0
    return cores * frames * hp.hop_size, dt


def cpu_shiftqueue_sample(hp, weights, frames=2):
    """The reference-faithful CPU row of BASELINE.md: one utterance with the LITERAL convolution queues of modules.py:285-292
    (the whole [2d+1, R] queue of every layer is shifted and re-concatenated every sample), scalar C, one core."""
    from oracle.c_oracle import COracle
    co = COracle(hp, weights)
    c, noise = synth_inputs(hp, [0], frames)
    t0 = time.perf_counter()
    co.synthesize(c, noise=noise, nthreads=1, shiftqueue=True)
    return frames * hp.hop_size, time.perf_counter() - t0


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from wavenet_vocoder import weights as W
    hp = get_hparams(args.hparams)
    weights = W.init_weights(hp, seed=5339, kind='tf-init')
    cores = os.cpu_count() or 1
    B, Tc, weak = WORKLOADS[args.workload]
    frames = 24   # 6 600 samples per utterance, ~5 s per core
    for _ in range(args.warmup):
        cpu_sample(hp, weights, cores, frames=2)
    n_tot, t_tot = 0, 0.0
    for _ in range(args.steps):
        n, dt = cpu_sample(hp, weights, cores, frames=frames)
        n_tot += n
        t_tot += dt
    v = n_tot / t_tot
    sample = ('{} utterances x {} frames ({} samples) per step: a bounded sample of the workload (the CPU cost per sample does not '
              'depend on batch or length), C restatement of the reference math (oracle/c/wn_oracle.c, ring-buffer queues), OpenMP '
              'over utterances, upsample included').format(cores, frames, cores * frames * hp.hop_size)
    cfg = workload_config(args, hp, B, Tc, args.gpus)
    cfg['measured_sample'] = sample
    cfg['note'] = ('this arm times the bounded sample above on ONE host (rank 0) whatever --gpus is: the ratio against an N-GPU '
                   'line compares N GPUs with one host')
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': UNIT, 'rtf': v / SAMPLE_RATE, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t_tot / args.steps, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': cfg,
        'cpu_baseline': {'value': v, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample,
                         'note': 'the reference is TensorFlow-1.x Python and cannot run on this box; this is the CPU '
                                 'restatement of its math (oracle/c/wn_oracle.c), pinned to the reference\'s own Python by '
                                 'tests/test_oracle_vs_reference_shim.py'},
        'e2e': {'value': v, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    try:
        n, dt = cpu_shiftqueue_sample(hp, weights)
        line['cpu_baseline_shiftqueue'] = {'value': n / dt, 'unit': UNIT, 'cores': 1, 'kind': 'port',
                                           'sample': '1 utterance x 2 frames ({} samples), literal shift-and-concat queues (modules.py:285-292)'.format(n)}
    except Exception as e:
        line['cpu_baseline_shiftqueue'] = {'error': '{}: {}'.format(type(e).__name__, e)}
    emit(line)


def workload_config(args, hp, B, Tc, n_gpus):
    weak = WORKLOADS[args.workload][2]
    return {
        'workload': 'WaveNet incremental synthesis, {} hparams (L{} R{} G{} S{} out{} {}), {} utterances x {:.1f} s '
                    '({} frames -> {} samples){}'.format(
                        args.hparams, hp.layers, hp.residual_channels, hp.gate_channels, hp.skip_out_channels,
                        hp.out_channels, hp.upsample_type, B, Tc * hp.hop_size / hp.sample_rate, Tc, Tc * hp.hop_size,
                        ' per GPU' if weak else ' total, sharded over the ranks'),
        'baseline_config': {'c2': 'configs[1]', 'c3': 'configs[2]', 'c1': 'configs[0]'}.get(args.workload, 'dev'),
        'utterances_per_gpu': B if weak else B // n_gpus, 'frames': Tc, 'samples_per_utterance': Tc * hp.hop_size,
        'weights': 'random-init seed 5339 (tf-init)', 'parallelism': 'dp{} (utterances, no data-path collective; final gather)'.format(n_gpus),
        'l2': 'inputs+scratch larger than L2 (upsampled conditioning {:.2f} GB per pass) and an L2 flush between steps'.format(
            (B if weak else B // n_gpus) * Tc * hp.hop_size * hp.num_mels * 4 / 1e9),
    }


# ---------------------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ---------------------------------------------------------------------------------------------------------------
class ClockSampler(object):
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '200'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2])); power.append(float(r[3]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith('active'):
                        reasons.add(n)
            except Exception:
                pass
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(smax) if smax else None,
                'power_w_max': max(power) if power else None, 'samples': len(sm), 'reasons': sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
def step_full_point(torch, net, hp, peaks):
    """BASELINE.json configs[4]: WaveNet.step (teacher forced, full sequence) on 8 x 10 s - the dense-GEMM formulation on the
    tensor cores (csrc/wn_step_umma.cu: tcgen05.mma, every fp32 operand split into two bf16 numbers, three passes).  A
    secondary figure: outside the timed region of the headline metric, device-resident buffers, CUDA events."""
    B, frames = 8, 802
    T = frames * hp.hop_size
    rng = np.random.default_rng(0)
    dc = torch.from_numpy(rng.uniform(0, 1, (B, frames, hp.num_mels)).astype(np.float32)).cuda()
    dx = torch.from_numpy(rng.uniform(-1, 1, (B, T)).astype(np.float32)).cuda()
    raw = torch.empty((B, T, hp.out_channels), dtype=torch.float32, device='cuda')
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sp = torch.cuda.current_stream().cuda_stream
    best = None
    for i in range(4):   # the first call allocates the scratch
        ev0.record()
        net.step_full(dx, dc, B, frames, raw, stream=sp)
        ev1.record()
        torch.cuda.synchronize()
        if i:
            best = ev0.elapsed_time(ev1) if best is None else min(best, ev0.elapsed_time(ev1))
    assert bool(torch.isfinite(raw).all())
    flops = float(net.info()['flops_per_sample']) * B * T
    peak = peaks.get('bf16_tflops', 1590.0)
    tensor_tflops = 3 * flops / (best / 1e3) / 1e12
    return {'workload': 'configs[4]: WaveNet.step, default_22k, 8 utterances x 10 s (1 764 400 positions)', 'ms': best,
            'fp32_equivalent_tflops': flops / (best / 1e3) / 1e12, 'bf16_tensor_tflops': tensor_tflops,
            'roofline': {'bound': 'tensor', 'achieved': flops / (best / 1e3) / 1e12, 'peak': peak, 'unit': 'TFLOP/s',
                         'frac': flops / (best / 1e3) / 1e12 / peak, 'executed_bf16_tflops': tensor_tflops, 'executed_frac': tensor_tflops / peak,
                         'peak_source': 'measured burst (MEASURED_PEAKS.json)' if 'bf16_tflops' in peaks else 'fallback (B200_PROFILING.md)',
                         'note': 'three bf16 passes (hi*hi, hi*lo, lo*hi) per fp32 product; see DESIGN.md 4.5'}}


def c4_point(torch):
    """BASELINE.json configs[3]: paper_hparams (R 256, G 512, S 256, 24 layers, 10 logistics, ConvTranspose2D [5, 5, 11]), 16
    utterances, on a 2 s sample of its 30 s (the step time does not depend on the length; the full-length pass is
    profiles/r2_c4_full_run.txt).  Its own model object (AUTO: csrc/wn_loop_tc2.cu), device-resident buffers, one warm-up call,
    one timed call; loop time from the library's CUDA events.  A secondary figure outside the headline's timed region."""
    from wavenet_vocoder._native import NativeWaveNet
    from wavenet_vocoder import weights as W
    hp = get_hparams('paper')
    B, frames = 16, 161
    T = frames * hp.hop_size
    net = NativeWaveNet(hp, W.init_weights(hp, seed=5339, kind='tf-init'))
    c_np, noise_np = synth_inputs(hp, list(range(B)), frames)
    d_c, d_noise = torch.from_numpy(c_np).cuda(), torch.from_numpy(noise_np).cuda()
    d_audio = torch.empty((B, T), dtype=torch.float32, device='cuda')
    sp = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        net.synthesize(d_c, B, frames, d_noise, d_audio, stream=sp)
        torch.cuda.synchronize()
    info = net.info()
    assert bool(torch.isfinite(d_audio).all()) and not info['aborted']
    ms = info['last_loop_ms'] + info['last_upsample_ms']
    v = B * T / (ms / 1e3)
    out = {'workload': 'configs[3]: paper_hparams, 16 utterances, 2 s sample of 30 s ({} frames -> {} samples each)'.format(frames, T),
           'value': v, 'unit': UNIT, 'rtf': v / hp.sample_rate, 'ms': ms, 'us_per_step': 1e3 * info['last_loop_ms'] / T,
           'engine': info['engine'], 'ctas': info['pipeline_ctas'], 'utterances_per_group': info['pipeline_group_size'],
           'round1_us_per_step': 329.5}
    net.close()
    return out


def strong_c3_point(torch, dist, net, hp, world, rank):
    """BASELINE.json configs[2]: 256 utterances x 10 s (802 frames) sharded over the ranks with wavenet_vocoder.sharding (the
    product's own functions), one pass, end to end: host buffers through wn_synthesize_host on every rank, then the final
    gather of the audio onto rank 0.  Wall clock between barriers, max over ranks.  Strong scaling: total work is fixed."""
    from wavenet_vocoder import sharding
    n_utt, frames = WORKLOADS['c3'][0], WORKLOADS['c3'][1]
    T = frames * hp.hop_size
    mine = sharding.shard_utterances([T] * n_utt, world)[rank]
    c_np, noise_np = synth_inputs(hp, mine, frames)
    h_audio = np.empty((len(mine), T), np.float32)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    sync()
    t0 = time.perf_counter()
    net.synthesize_host(c_np, noise_np, audio_out=h_audio)
    info = net.info()
    if world > 1:
        out = sharding.gather_wavs(list(h_audio), mine, n_utt)
    else:
        out = list(h_audio)
    sync()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    dt = float(dt.item())
    if rank != 0:
        return None
    assert len(out) == n_utt and all(len(w) == T and np.isfinite(w).all() for w in out) and not info['aborted']
    v = n_utt * T / dt
    return {'workload': 'configs[2]: default_22k, 256 utterances x 10 s (802 frames -> 220 550 samples each) sharded over {} GPU(s) '
                        'with sharding.shard_utterances, final gather with sharding.gather_wavs included'.format(world),
            'scaling': 'strong', 'value': v, 'unit': UNIT, 'rtf': v / SAMPLE_RATE, 'rtf_target': 1000.0, 'seconds': dt, 'n_gpus': world,
            'utterances_per_gpu': len(mine), 'engine': info['engine'], 'loop_ms_rank0': info['last_loop_ms'],
            'us_per_step_rank0': 1e3 * info['last_loop_ms'] / T / max(1, -(-len(mine) // 336)),   # per launch: one launch takes 336 utterances
            'layer_chains': info['pipeline_ctas_per_layer'], 'utterances_per_group': info['pipeline_group_size'],
            'limiter': 'serial stage latency of the layer chain (us_per_step x 220 550 steps); no collective on the data path',
            'api': 'wn_synthesize_host per rank (host buffers) + one gather'}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from wavenet_vocoder import weights as W
    from wavenet_vocoder._native import NativeWaveNet

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    assert torch.cuda.is_available(), 'bench.py --impl ours needs a B200 (there is no CPU path)'
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    n_gpus = world

    hp = get_hparams(args.hparams)
    B_total, Tc, weak = WORKLOADS[args.workload]
    B = B_total if weak else B_total // world
    utt0 = rank * B
    T = Tc * hp.hop_size
    weights = W.init_weights(hp, seed=5339, kind='tf-init')
    net = NativeWaveNet(hp, weights, engine=args.engine)
    info0 = net.info()
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream

    # synthetic inputs, resident in HBM before the timed region
    c_np, noise_np = synth_inputs(hp, list(range(utt0, utt0 + B)), Tc)
    d_c = torch.from_numpy(c_np).cuda()
    d_noise = torch.from_numpy(noise_np).cuda()
    d_audio = torch.empty((B, T), dtype=torch.float32, device='cuda')
    gathered = [torch.empty_like(d_audio) for _ in range(world)] if (world > 1 and rank == 0) else None
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device='cuda')  # 256 MB > 126 MB L2

    def one_step():
        net.synthesize(d_c, B, Tc, d_noise, d_audio, stream=sp)
        if world > 1:   # the final gather of the path (synthesizer.py:100-101): audio of every rank onto rank 0
            dist.gather(d_audio, gathered, dst=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_step()
    barrier()
    assert bool(torch.isfinite(d_audio).all()), 'non-finite audio'

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = net.info()['kernel_launches']
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    loop_ms, up_ms = [], []
    barrier()
    for s, e in ev:
        flush.fill_(1.0)           # L2 flush, outside the event pair
        s.record(stream)
        one_step()
        e.record(stream)
        e.synchronize()
        inf = net.info()
        loop_ms.append(inf['last_loop_ms'])
        up_ms.append(inf['last_upsample_ms'])
    barrier()
    assert not net.info()['aborted'], 'the persistent loop kernel hit its watchdog (WN_ERR_WATCHDOG): the numbers would be meaningless'
    assert bool(torch.isfinite(d_audio).all()), 'non-finite audio after the timed steps'
    total_ms = sum(s.elapsed_time(e) for s, e in ev)
    launches = net.info()['kernel_launches'] - launches0   # this library's kernels only (the L2 flush fills are torch's)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    samples_per_step = B * T * world
    value = samples_per_step * args.steps / (total_ms / 1e3)

    # e2e: the host-buffer C-ABI call (pinned host memory; H2D of mels+noise and D2H of audio inside the timed region)
    e2e = None
    if not args.no_e2e:
        h_c = torch.from_numpy(c_np).pin_memory()
        h_noise = torch.from_numpy(noise_np).pin_memory()
        h_audio = torch.empty((B, T), dtype=torch.float32).pin_memory()
        net.synthesize_host(h_c.numpy(), h_noise.numpy(), audio_out=h_audio.numpy(), stream=sp)  # warm-up (staging buffers)
        n_e2e = max(1, min(args.steps, 3))
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            net.synthesize_host(h_c.numpy(), h_noise.numpy(), audio_out=h_audio.numpy(), stream=sp)
            if world > 1:   # the final gather belongs to the path: rank 0 ends up holding every rank's audio
                d_audio.copy_(h_audio, non_blocking=True)
                dist.gather(d_audio, gathered, dst=0)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {'value': samples_per_step * n_e2e / float(dt.item()), 'unit': UNIT,
               'h2d_bytes_per_step': int(c_np.nbytes + noise_np.nbytes), 'd2h_bytes_per_step': int(B * T * 4),
               'steps': n_e2e, 'api': 'wn_synthesize_host (libwavenet_b200.so C ABI), pinned host buffers'}
        assert bool(np.isfinite(h_audio.numpy()).all())

    strong = None
    if not args.no_strong_c3 and args.workload == 'c2' and args.hparams == 'default_22k':
        try:
            strong = strong_c3_point(torch, dist, net, hp, world, rank)
        except Exception as e:   # a secondary figure must never cost the headline line
            strong = {'error': '{}: {}'.format(type(e).__name__, e)}
    if rank == 0:
        info = net.info()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:
            pass
        hbm_peak = peaks.get('hbm_gbs', 6650.0)
        peak_src = 'measured (MEASURED_PEAKS.json)' if 'hbm_gbs' in peaks else 'fallback (B200_PROFILING.md)'
        lm = float(np.mean(loop_ms))
        alg_bytes = float(info['bytes_per_sample']) * B * T
        achieved = alg_bytes / (lm / 1e3) / 1e9
        # DRAM traffic per launch: bytes/sample of the committed `ncu --set full` capture (same kernel, same batch, shorter
        # utterances) scaled to this launch's sample count
        traffic, traffic_src = None, None
        try:
            prof = {}
            for ln in open(os.path.join(ROOT, 'profiles', 'r1_pipe_ncu_full_summary.csv')):
                f = ln.strip().split(',')
                if len(f) == 3 and f[0].startswith('dram__bytes'):
                    prof[f[0]] = float(f[2]) * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}[f[1]]
            per_sample = (prof['dram__bytes_read.sum'] + prof['dram__bytes_write.sum']) / (64 * 1100)
            traffic = per_sample * B * T
            traffic_src = 'profiles/r1_pipe_ncu_full_summary.csv: {:.0f} B/sample measured at B=64, T=1100, scaled to this launch'.format(per_sample)
        except Exception:
            pass
        fp32_peak = 148 * 128 * 2 * 1.965e9 / 1e12   # CUDA-core fp32 peak (FFMA2), nominal: MEASURED_PEAKS.json has no fp32 figure
        alg_tflops = info['flops_per_sample'] * B * T / (lm / 1e3) / 1e12
        common_fields = {
            'kernel': 'autoregressive loop kernel ({} engine)'.format(info['engine']),
            'kernel_ms': lm, 'upsample_ms': float(np.mean(up_ms)), 'us_per_step': 1e3 * lm / T,
            'us_per_step_budget': 11.6, 'algorithmic_flops_per_launch': float(info['flops_per_sample']) * B * T,
            'algorithmic_bytes_per_launch': alg_bytes, 'bytes_per_sample': int(info['bytes_per_sample']),
            'hbm_gbs_algorithmic': achieved, 'hbm_frac': achieved / hbm_peak, 'hbm_peak_gbs': hbm_peak, 'peak_source': peak_src,
            'upsample_gbs': B * T * hp.num_mels * 4 * (1 + 1.0 / hp.hop_size) / (float(np.mean(up_ms)) / 1e3) / 1e9,
            'note': 'the loop is a serial chain of layers + 1 stages per sample: us_per_step (budget 11.6 for the 1000x target) is '
                    'the figure of merit; bytes and flops are far from their roofs; see DESIGN.md 4.2 / 4.7',
        }
        if info['engine'] == 'tc':
            # tcgen05 engine: a batched step is tensor work (north_star); the fraction is on ALGORITHMIC flops against the
            # measured sustained bf16 GEMM rate (fp16 MMAs run at the same rate); the kernel executes 2-3 fp16 MMA passes per fp32 product
            t_peak = peaks.get('bf16_tflops_sustained', peaks.get('bf16_tflops', 1590.0))
            roofline = dict(common_fields, bound='tensor', achieved=alg_tflops, peak=t_peak, unit='TFLOP/s', frac=alg_tflops / t_peak,
                            traffic=None, traffic_source='not captured on the timed shape (see profiles/ for the range capture)',
                            executed_fp16_tflops=3 * alg_tflops, layer_chains=info['pipeline_ctas_per_layer'], utterances_per_group=info['pipeline_group_size'])
        else:
            roofline = dict(common_fields, bound='hbm', achieved=achieved, peak=hbm_peak, unit='GB/s', frac=achieved / hbm_peak,
                            traffic=traffic, traffic_source=None if traffic is None else 'EXTRAPOLATED, not measured on this launch: ' + traffic_src,
                            fp32_peak_tflops_nominal=fp32_peak, fp32_frac_nominal=alg_tflops / fp32_peak, fp32_tflops=alg_tflops)
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'rtf': value / SAMPLE_RATE, 'n_gpus': n_gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': total_ms / args.steps, 'higher_is_better': True,
            'scaling': 'weak' if weak else 'strong', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': workload_config(args, hp, B_total, Tc, n_gpus), 'engine': info['engine'],
            'clocks': clocks, 'gpu_launches': int(launches), 'roofline': roofline,
        }
        if e2e is not None:
            line['e2e'] = e2e
        if strong is not None:
            line['strong_c3'] = strong
        if world == 1 and not args.no_step_full and args.hparams == 'default_22k':
            try:   # a secondary figure must never cost the headline line
                line['step_full'] = step_full_point(torch, net, hp, peaks)
            except Exception as e:
                line['step_full'] = {'error': '{}: {}'.format(type(e).__name__, e)}
        if world == 1 and not args.no_c4 and args.workload == 'c2' and args.hparams == 'default_22k':
            try:
                line['c4_paper'] = c4_point(torch)
            except Exception as e:
                line['c4_paper'] = {'error': '{}: {}'.format(type(e).__name__, e)}
        if not args.no_cpu_baseline and world == 1:
            cores = os.cpu_count() or 1
            n, dt = cpu_sample(hp, weights, cores)
            line['cpu_baseline'] = {
                'value': n / dt, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                'sample': '{} utterances x 24 frames ({} samples), C oracle port (oracle/c/wn_oracle.c), OpenMP over '
                          'utterances, upsample included'.format(cores, n)}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
