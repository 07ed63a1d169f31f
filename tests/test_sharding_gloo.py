"""CPU, world_size 2 over gloo: the multi-GPU path's host logic -- every rank derives the same
sequence partition without communicating, runs its shard independently (the C oracle stands in for
the device kernel here; the path has no data-path collective), and the optional final gather
restores global order."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from light_vllm_b200.sharding import gather_outputs, partition_sequences


def test_partition_is_balanced_and_deterministic():
    lens = [4096, 100, 32768, 7, 2048, 2048, 9000, 1]
    shards = partition_sequences(lens, 3)
    assert sorted(i for s in shards for i in s) == list(range(len(lens)))
    loads = [sum(lens[i] for i in s) for s in shards]
    assert max(loads) == 32768  # the longest sequence alone bounds the step
    assert partition_sequences(lens, 3) == shards
    assert partition_sequences(lens, 1) == [list(range(len(lens)))]
    assert partition_sequences([], 2) == [[], []]
    eq = partition_sequences([2048] * 64, 8)
    assert all(len(s) == 8 for s in eq)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, tmpdir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    from kv_util import make_case
    from oracle import oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        lens = [40, 700, 3, 130, 513]
        full = make_case(len(lens), 4, 2, 64, 16, lens, torch.float32, seed=5)
        shards = partition_sequences(lens, world)
        mine = shards[rank]
        # this rank's shard: its own rows of the block table / seq_lens (pool shared read-only
        # here only because the test builds one global case; on GPUs each rank owns its pool)
        idx = torch.tensor(mine, dtype=torch.long)
        q = full["query"][idx].contiguous()
        out = torch.empty_like(q)
        oracle.paged_attention_v1(out, q, full["key_cache"], full["value_cache"], 2, full["scale"],
                                  full["block_tables"][idx].contiguous(),
                                  full["seq_lens"][idx].contiguous(), 16, max(lens), None, "auto",
                                  1.0, 1.0)
        merged = gather_outputs(out, shards, len(lens))
        want = torch.empty_like(full["query"])
        oracle.paged_attention_v1(want, full["query"], full["key_cache"], full["value_cache"], 2,
                                  full["scale"], full["block_tables"], full["seq_lens"], 16,
                                  max(lens), None, "auto", 1.0, 1.0)
        assert torch.equal(merged, want), f"rank {rank}: gathered output differs"
        # max-over-ranks timing reduction used by bench.py
        t = torch.tensor([float(rank + 1)])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == world
        open(os.path.join(tmpdir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


def test_two_rank_sequence_sharding(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
