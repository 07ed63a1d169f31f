"""Multi-GPU decode: sequences are sharded, nothing else.

Every (sequence, head) output of paged attention depends only on that sequence's own KV pages
(attention_kernels.cu:111-115,207,257-258), so the path partitions by sequence: one process per
GPU, each rank owns its KV pools, block tables and the sequences assigned to it; physical block
ids are rank-local.  This mirrors the reference's only multi-GPU mechanism -- independent
replicas fed from one queue (prefill_only/executor/gpu_data_parallelism_executor.py:41-72).
There is no data-path collective; the only optional exchange is the final gather of the per-rank
outputs ``[B_rank, H, D]`` (a few MB) over NCCL (NVLink 5 / NVSwitch) -- or gloo in CPU tests.
"""
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


def partition_sequences(seq_lens: Sequence[int], world_size: int) -> List[List[int]]:
    """Assign sequence indices to ranks, balancing KV tokens (= bytes the rank must read per step).

    Longest-processing-time greedy: sequences in decreasing length go to the currently lightest
    rank; ties broken by rank id, indices kept sorted per rank, so every rank computes the same
    assignment from the same lengths without communicating."""
    assert world_size >= 1
    order = sorted(range(len(seq_lens)), key=lambda i: (-int(seq_lens[i]), i))
    load = [0] * world_size
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], len(shards[k]), k))
        shards[r].append(i)
        load[r] += int(seq_lens[i])
    return [sorted(s) for s in shards]


def gather_outputs(local_out: torch.Tensor, shards: List[List[int]], num_seqs: int,
                   group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """All-gather the per-rank outputs ``[len(shards[rank]), ...]`` into global sequence order.

    Ranks may hold different numbers of sequences: shards are padded to the largest one so a plain
    (NCCL / gloo) all_gather into one tensor suffices."""
    if not (dist.is_available() and dist.is_initialized()):
        assert len(shards) == 1
        return local_out
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    assert world == len(shards) and local_out.shape[0] == len(shards[rank])
    width = max(len(s) for s in shards)
    padded = local_out.new_zeros((width,) + tuple(local_out.shape[1:]))
    padded[:local_out.shape[0]] = local_out
    flat = local_out.new_empty((world * width,) + tuple(local_out.shape[1:]))
    dist.all_gather_into_tensor(flat, padded, group=group)
    gathered = flat.view((world, width) + tuple(local_out.shape[1:]))
    out = local_out.new_empty((num_seqs,) + tuple(local_out.shape[1:]))
    for r, idx in enumerate(shards):
        if idx:
            out[torch.as_tensor(idx, device=out.device)] = gathered[r, :len(idx)]
    return out
