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
