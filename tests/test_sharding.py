"""Multi-GPU host logic on CPU: world_size-2 gloo run of the shard -> synthesise -> gather path (the per-rank
synthesis is replaced by a deterministic stand-in: there is no CPU implementation of the kernels)."""
import os
import socket

import numpy as np
import pytest

from wavenet_vocoder.sharding import shard_utterances


def test_shard_utterances_balanced_and_complete():
    lengths = [50, 10, 40, 30, 20, 60, 5]
    for world in (1, 2, 3, 4, 8):
        shards = shard_utterances(lengths, world)
        assert sorted(i for s in shards for i in s) == list(range(len(lengths)))
        sizes = [len(s) for s in shards]
        assert max(sizes) - min(sizes) <= 1
    s2 = shard_utterances(lengths, 2)
    loads = [sum(lengths[i] for i in s) for s in s2]
    assert abs(loads[0] - loads[1]) <= 25


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, lengths, q):
    import torch.distributed as dist
    from wavenet_vocoder.sharding import gather_wavs, shard_utterances as shard
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        mine = shard(lengths, world)[rank]
        wavs = [np.full(lengths[i], float(i), np.float32) + np.arange(lengths[i], dtype=np.float32) / 1024 for i in mine]
        out = gather_wavs(wavs, mine, len(lengths))
        if rank == 0:
            ok = all(len(out[i]) == lengths[i] and out[i][0] == float(i) and
                     np.array_equal(out[i], np.full(lengths[i], float(i), np.float32) + np.arange(lengths[i], dtype=np.float32) / 1024)
                     for i in range(len(lengths)))
            q.put(ok)
        else:
            assert out is None
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_gather_world_size_2_gloo():
    import torch.multiprocessing as mp
    lengths = [45, 30, 15, 45, 7]
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(100)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


# ---- the product path: Synthesizer.synthesize under a world_size-2 torch.distributed job (gloo, CPU) -------------------
class _StandInModel:
    """Deterministic stand-in for the GPU model: wav[b, t] = mean(c[b]) + noise[t, b, 0] / 16, so the result depends on the
    utterance's own conditioning and its own draws only."""
    def __init__(self, hp):
        self.hp = hp

    def initialize(self, y, c, g, input_lengths, test_inputs=None, noise=None, want_raw=True):
        B, Tc, _ = c.shape
        T = Tc * self.hp.hop_size
        return (c.mean(axis=(1, 2))[:, None] + noise[:T, :, 0].T / 16).astype(np.float32)


def _make_synth(hp):
    from wavenet_vocoder import synthesizer as S
    s = S.Synthesizer()
    s._hparams = hp
    s.synth_debug = False
    s.model = _StandInModel(hp)
    s.models = [s.model]
    s._rng = np.random.default_rng(0)
    return s


def _synth_worker(rank, world, port, out_dir, q):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path[:0] = [here]
    import torch.distributed as dist
    import common
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        hp = common.small()
        mels = common.mels_for(hp, 5, 6, ragged=True)
        s = _make_synth(hp)
        files = s.synthesize(mels, None, ['u%d' % i for i in range(5)], out_dir, out_dir, seed=11)
        assert s.is_writer == (rank == 0)
        q.put((rank, len(files)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_synthesizer_shards_and_gathers_world_size_2_gloo(tmp_path):
    """Same files, same samples as the single-process run: the draws of an utterance do not depend on the number of ranks."""
    import torch.multiprocessing as mp
    import common
    from scipy.io import wavfile
    hp = common.small()
    mels = common.mels_for(hp, 5, 6, ragged=True)
    single = tmp_path / 'single'
    multi = tmp_path / 'multi'
    single.mkdir()
    multi.mkdir()
    names = ['u%d' % i for i in range(5)]
    ref_files = _make_synth(hp).synthesize(mels, None, names, str(single), str(single), seed=11)
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_synth_worker, args=(r, 2, port, str(multi), q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(100)
        assert p.exitcode == 0
    counts = dict(q.get(timeout=5) for _ in range(2))
    assert counts == {0: 5, 1: 0}          # rank 0 writes, the other rank returns nothing
    for f in ref_files:
        a = wavfile.read(f)[1]
        b = wavfile.read(os.path.join(str(multi), os.path.basename(f)))[1]
        assert len(a) > 0 and np.array_equal(a, b)
