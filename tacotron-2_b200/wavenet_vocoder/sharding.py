"""Utterance-level data parallelism of the synthesis path over the GPUs of one box.

The reference splits the batch with ``tf.split`` over in-graph towers (``wavenet.py:227-239``) and flattens the
per-tower lists on the host (``synthesizer.py:100-101``).  Here every GPU is its own process (torchrun, one rank per
GPU); utterances never interact, so there is no data-path collective: each rank synthesises its shard and rank 0
gathers the audio once at the end (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
import numpy as np


def shard_utterances(lengths, world_size):
	"""Longest-processing-time-first assignment of utterances to ranks (the step count of a rank is the longest
	utterance it holds, its work the sum).  Returns a list of index lists, one per rank; every rank gets
	ceil/floor(n / world_size) utterances so that batch shapes stay balanced."""
	n = len(lengths)
	order = sorted(range(n), key=lambda i: (-int(lengths[i]), i))
	cap = [n // world_size + (1 if r < n % world_size else 0) for r in range(world_size)]
	shards = [[] for _ in range(world_size)]
	load = [0] * world_size
	for i in order:
		cand = [r for r in range(world_size) if len(shards[r]) < cap[r]]
		r = min(cand, key=lambda r: (load[r], r))
		shards[r].append(i)
		load[r] += int(lengths[i])
	return [sorted(s) for s in shards]


def gather_wavs(local_wavs, local_ids, n_total, group=None, dst=0):
	"""Final gather of the path: every rank contributes its [wav, ...] (ragged float32 arrays) for utterance ids
	``local_ids``; rank ``dst`` returns the full list in utterance order, the others None."""
	import torch
	import torch.distributed as dist
	world = dist.get_world_size(group)
	rank = dist.get_rank(group)
	backend = dist.get_backend(group)
	dev = torch.device('cuda', torch.cuda.current_device()) if backend == 'nccl' else torch.device('cpu')
	# 1) lengths and ids (small), 2) one padded [n_local_max, T_max] float tensor per rank
	meta = torch.full((n_total, 2), -1, dtype=torch.int64, device=dev)
	for j, (i, w) in enumerate(zip(local_ids, local_wavs)):
		meta[j, 0] = i
		meta[j, 1] = len(w)
	metas = [torch.empty_like(meta) for _ in range(world)]
	dist.all_gather(metas, meta, group=group)
	all_meta = torch.stack(metas).cpu().numpy()
	n_max = int((all_meta[:, :, 0] >= 0).sum(axis=1).max())
	t_max = int(max(1, all_meta[:, :, 1].max()))
	buf = torch.zeros((max(1, n_max), t_max), dtype=torch.float32, device=dev)
	for j, w in enumerate(local_wavs):
		buf[j, :len(w)] = torch.as_tensor(np.asarray(w, np.float32), device=dev)
	bufs = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
	dist.gather(buf, bufs, dst=dst, group=group)
	if rank != dst:
		return None
	out = [None] * n_total
	for r in range(world):
		host = bufs[r].cpu().numpy()
		for j in range(n_total):
			i, ln = int(all_meta[r, j, 0]), int(all_meta[r, j, 1])
			if i >= 0:
				out[i] = host[j, :ln].copy()
	return out
