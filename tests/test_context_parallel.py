"""Context-sharded decode (one sequence's pages on several GPUs).

CPU: the host policy (who owns which logical block, local lengths / slots) and, over gloo with two
ranks, that attention over the rank-local slices merged with the v2 formula
(attention_kernels.cu:565-669) equals attention over the whole context.
GPU: the product path (partition pass + NVLink exchange/merge kernel through the C ABI) against the
C oracle run on the unsharded case: virtual ranks on one device (one CUDA stream each), and real
one-process-per-GPU ranks over CUDA IPC when the box has two GPUs."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from kv_util import make_case, naive_attention, to_device
from light_vllm_b200.context_parallel import ContextSharding


# ---------------------------------------------------------------- host policy (CPU) ----
@pytest.mark.parametrize("world,chunk", [(1, 1), (2, 1), (3, 2), (4, 1), (8, 4)])
def test_policy_partitions_every_block_once(world, chunk):
    bs = 16
    sh = ContextSharding(world, bs, chunk)
    for seq_len in (0, 1, 15, 16, 17, 31 * bs, 31 * bs + 5, 1000, 4096 + 3):
        nb = (seq_len + bs - 1) // bs
        table = list(range(100, 100 + nb))
        parts = sh.split_block_table(table, seq_len)
        assert sorted(b for p in parts for b in p) == table
        for r, p in enumerate(parts):
            assert len(p) == sh.local_num_blocks(nb, r)
            assert p == sorted(p)  # local order preserves global order
            for lb, phys in enumerate(p):
                b = phys - 100
                assert sh.owner_of_block(b) == r and sh.local_block(b) == lb
        lens = [sh.local_len(seq_len, r) for r in range(world)]
        assert sum(lens) == seq_len
        for r in range(world):
            assert (lens[r] + bs - 1) // bs == len(parts[r])
        if seq_len:
            pos = seq_len - 1
            r = sh.owner_of_token(pos)
            # the newest token lands in the last slot of its owner's local slice
            assert sh.local_slot(pos, parts[r]) == parts[r][-1] * bs + pos % bs
            assert sh.local_block(pos // bs) == len(parts[r]) - 1


def test_exchange_bytes():
    from light_vllm_b200 import _lib
    f = _lib.lib.b200_ctx_exchange_bytes
    assert f(0, 8, 128) == -1 and f(17, 8, 128) == -1 and f(2, 0, 128) == -1
    n2 = f(2, 64, 128)
    assert n2 >= 2 * 2 * 64 * 132 * 4 and n2 % 16 == 0
    assert f(8, 64, 128) > n2


def _merge(triples):
    """v2 merge of [(m, l, o)] over ranks, fp64 (attention_kernels.cu:643-667 without the 1e-6)."""
    m = torch.stack([t[0] for t in triples])
    l = torch.stack([t[1] for t in triples])
    o = torch.stack([t[2] for t in triples])
    M = m.max(dim=0).values
    w = l * torch.exp(m - M)
    w = torch.where(l > 0, w, torch.zeros_like(w))
    return (o * w.unsqueeze(-1)).sum(0) / w.sum(0).clamp_min(1e-300).unsqueeze(-1)


def _local_triple(case, bt, lens):
    """(max logit, exp sum, normalised out) per (seq, head) over a rank-local slice, fp64."""
    q = case["query"].double()
    B, H, D = q.shape
    KH, BS = case["num_kv_heads"], case["block_size"]
    m = torch.full((B, H), float("-inf"), dtype=torch.float64)
    l = torch.zeros(B, H, dtype=torch.float64)
    o = torch.zeros(B, H, D, dtype=torch.float64)
    for b in range(B):
        n = int(lens[b])
        if n == 0:
            continue
        nblk = (n + BS - 1) // BS
        blks = bt[b, :nblk].long()
        k = case["key_cache"][blks].double().permute(1, 0, 3, 2, 4).reshape(KH, nblk * BS, D)[:, :n]
        v = case["value_cache"][blks].double().permute(1, 0, 3, 2).reshape(KH, nblk * BS, D)[:, :n]
        for h in range(H):
            kvh = h // (H // KH)
            logits = (k[kvh] @ q[b, h]) * case["scale"]
            m[b, h] = logits.max()
            p = torch.exp(logits - m[b, h])
            l[b, h] = p.sum()
            o[b, h] = (p @ v[kvh]) / l[b, h]
    return m, l, o


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, tmpdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        lens = [5, 16, 100, 700, 0]
        case = make_case(len(lens), 4, 2, 64, 16, lens, torch.float32, seed=11)
        sh = ContextSharding(world, 16, chunk_blocks=2)
        bt, ll, _ = sh.shard_batch(case["block_tables"], lens, rank)
        mine = torch.cat([t.reshape(len(lens), 4, -1) for t in _local_triple(case, bt, ll)], dim=-1)
        every = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)  # stands in for the NVLink peer stores of the device kernel
        got = _merge([(e[..., 0], e[..., 1], e[..., 2:]) for e in every])
        want = naive_attention(case)
        assert (got - want).abs().max().item() < 1e-12
        open(os.path.join(tmpdir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


def test_two_rank_context_sharding_gloo(tmp_path):
    mp.spawn(_gloo_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(2))


# ---------------------------------------------------------------- product path (GPU) ----
def _oracle_out(oracle, case, kv_cache_dtype="auto", k_scale=1.0, v_scale=1.0):
    want = torch.empty_like(case["query"])
    oracle.paged_attention_v1(want, case["query"], case["key_cache"], case["value_cache"],
                              case["num_kv_heads"], case["scale"], case["block_tables"],
                              case["seq_lens"], case["block_size"], case["max_seq_len"], None,
                              kv_cache_dtype, k_scale, v_scale)
    return want


def _run_virtual(case, world, chunk, calls=1, kv_cache_dtype="auto", k_scale=1.0, v_scale=1.0,
                 fused=False):
    from light_vllm_b200.context_parallel import ContextShardGroup
    dev = torch.device("cuda:0")
    d = to_device(case, dev)
    B, H, D = case["query"].shape
    sh = ContextSharding(world, case["block_size"], chunk)
    groups = ContextShardGroup.create_local(world, B * H, D, 0)
    streams = [torch.cuda.Stream(dev) for _ in range(world)]
    local = []
    for r in range(world):
        bt, ll, mx = sh.shard_batch(case["block_tables"], case["seq_lens"].tolist(), r)
        local.append((bt.to(dev), ll.to(dev), mx))
    outs = [torch.full_like(d["query"], float("nan")) for _ in range(world)]
    torch.cuda.synchronize()
    try:
        # Virtual ranks share ONE device: a fused publish+wait kernel of one rank spins until the
        # other rank's kernels have run, and those may be queued behind it (shared hardware queue)
        # or be denied the SMs it occupies (observed with a 148-CTA exchange grid).  So virtual
        # ranks use the split-phase form -- every publish is queued before any wait; the fused
        # kernel is exercised by the real one-process-per-GPU ranks (test_two_gpu_ipc_ranks).
        for _ in range(calls):
            for ph in ((3,) if fused else (1, 2)):
                for r in range(world):
                    with torch.cuda.stream(streams[r]):
                        bt, ll, mx = local[r]
                        groups[r].paged_attention(outs[r], d["query"], d["key_cache"],
                                                  d["value_cache"], case["num_kv_heads"],
                                                  case["scale"], bt, ll, case["block_size"], mx,
                                                  kv_cache_dtype, k_scale, v_scale, phases=ph)
        torch.cuda.synchronize()
        assert [g.status() for g in groups] == [0] * world
    finally:
        for g in groups:
            g.close()
    return [o.cpu() for o in outs]


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [(torch.float16, 2e-3), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("world,chunk", [(2, 1), (4, 2), (8, 1)])
def test_virtual_ranks_match_oracle(oracle, dtype, tol, world, chunk):
    lens = [1, 15, 33, 700, 2049, 0, 5000]
    case = make_case(len(lens), 32, 8, 128, 16, lens, dtype, seed=21, nan_padding=True)
    want = _oracle_out(oracle, case).float()
    outs = _run_virtual(case, world, chunk, calls=3)  # 3 calls: both data slots + epoch reuse
    for r, o in enumerate(outs):
        err = (o.float() - want).abs().max().item()
        assert err <= tol, f"rank {r}: max-abs {err}"
    for o in outs[1:]:  # every rank merges the same triples in the same order
        assert torch.equal(o, outs[0])


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(12))
def test_virtual_ranks_fuzz_against_fp64_truth(seed):
    """Random worlds / chunk sizes / head shapes / ragged lengths (many sequences own no block on
    some rank): every rank's merged output against the fp64 softmax over the whole context."""
    import random
    r = random.Random(9000 + seed)
    world = r.choice([2, 3, 4, 8])
    chunk = r.choice([1, 2, 4])
    D = r.choice([64, 80, 96, 112, 120, 128, 192, 256])
    BS = r.choice([8, 16, 16, 32])
    G = r.choice([1, 2, 4, 7, 8])
    KH = r.choice([1, 2, 4])
    dtype = r.choice([torch.float16, torch.bfloat16])
    B = r.choice([1, 2, 5, 9])
    longest = r.choice([60, 900, 4000, 9000])
    lens = [r.choice([0, 1, r.randint(1, 3 * BS), r.randint(1, longest)]) for _ in range(B)]
    lens[r.randrange(B)] = longest
    case = make_case(B, G * KH, KH, D, BS, lens, dtype, seed=seed, nan_padding=True)
    want = naive_attention(case)
    outs = _run_virtual(case, world, chunk, calls=2)
    live = torch.tensor(lens) > 0
    tol = 2e-3 if dtype == torch.float16 else 1e-2
    for rank, o in enumerate(outs):
        err = (o.double()[live] - want[live]).abs().max().item()
        assert err <= tol, (seed, rank, world, chunk, D, BS, G, KH, dtype, lens, err)
    for o in outs[1:]:
        assert torch.equal(o[live], outs[0][live])


@pytest.mark.gpu
def test_virtual_ranks_fp8_and_small_heads(oracle):
    lens = [40, 1300, 513]
    case = make_case(len(lens), 28, 4, 128, 16, lens, torch.bfloat16, seed=3)
    from light_vllm_b200 import _custom_ops as ops
    dev = "cuda:0"
    kc8 = torch.empty(case["key_cache"].shape, dtype=torch.uint8, device=dev)
    vc8 = torch.empty(case["value_cache"].shape, dtype=torch.uint8, device=dev)
    ops.convert_fp8(kc8, case["key_cache"].to(dev), 0.5, "fp8_e4m3")
    ops.convert_fp8(vc8, case["value_cache"].to(dev), 0.25, "fp8_e4m3")
    c8 = dict(case, key_cache=kc8.cpu(), value_cache=vc8.cpu())
    want = _oracle_out(oracle, c8, "fp8_e4m3", 0.5, 0.25).float()
    outs = _run_virtual(c8, 2, 1, kv_cache_dtype="fp8_e4m3", k_scale=0.5, v_scale=0.25)
    assert (outs[0].float() - want).abs().max().item() <= 1e-2
    c64 = make_case(2, 14, 2, 64, 16, [100, 3000], torch.float16, seed=4)
    want = _oracle_out(oracle, c64).float()
    outs = _run_virtual(c64, 4, 1)
    assert (outs[3].float() - want).abs().max().item() <= 2e-3


@pytest.mark.gpu
def test_missing_peer_times_out_instead_of_hanging():
    from light_vllm_b200.context_parallel import ContextShardGroup
    case = make_case(1, 8, 2, 128, 16, [64], torch.float16, seed=1)
    d = to_device(case, "cuda:0")
    sh = ContextSharding(2, 16, 1)
    groups = ContextShardGroup.create_local(2, 8, 128, 0)
    try:
        bt, ll, mx = sh.shard_batch(case["block_tables"], [64], 0)
        out = torch.zeros_like(d["query"])
        groups[0].set_timeout(50)  # per-group, lives in the exchange buffer (default 2 s)
        args = (out, d["query"], d["key_cache"], d["value_cache"], 2, case["scale"], bt.cuda(),
                ll.cuda(), 16, mx)
        groups[0].paged_attention(*args)  # rank 1 never calls
        torch.cuda.synchronize()
        assert groups[0].status() == 2  # 1 + the rank that never published
        # loud in the data: the stale peer slot is NOT merged, the rows come back NaN ...
        assert torch.isnan(out.float()).all()
        # ... and loud on the host: the next call (or check()) raises, naming the rank and the epoch
        with pytest.raises(RuntimeError, match="timed out waiting for rank 1 to publish epoch 1"):
            groups[0].paged_attention(*args)
        with pytest.raises(RuntimeError, match="timed out"):
            groups[0].check(wait=True)
        groups[0].clear_status()
        groups[0].check(wait=True)
        assert groups[0].status() == 0
    finally:
        for g in groups:
            g.close()


@pytest.mark.gpu
def test_exchange_trace_is_monotonic():
    """b200_ctx_exchange_trace: CTA 0's eight timestamps of the last exchange launch, in program order."""
    import ctypes
    from light_vllm_b200 import _lib
    from light_vllm_b200.context_parallel import ContextShardGroup
    case = make_case(2, 8, 2, 128, 16, [700, 90], torch.float16, seed=3)
    d = to_device(case, "cuda:0")
    sh = ContextSharding(1, 16, 1)
    grp = ContextShardGroup.create_local(1, 16, 128, 0)[0]
    try:
        bt, ll, mx = sh.shard_batch(case["block_tables"], [700, 90], 0)
        out = torch.zeros_like(d["query"])
        for _ in range(2):
            grp.paged_attention(out, d["query"], d["key_cache"], d["value_cache"], 2, case["scale"], bt.cuda(),
                                ll.cuda(), 16, mx)
        buf = (ctypes.c_uint64 * 32)()
        assert _lib.lib.b200_ctx_exchange_trace(grp._own, 1, buf, 0) == 0
        for call in range(2):
            ns = [buf[call * 16 + 2 * i] for i in range(8)]
            assert ns[0] > 0 and all(b >= a for a, b in zip(ns, ns[1:])), ns
            assert ns[7] - ns[0] < 1_000_000_000
        grp.check(wait=True)
    finally:
        grp.close()


def _ipc_worker(rank, world, port, tmpdir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    from light_vllm_b200.context_parallel import ContextShardGroup
    from oracle import oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        lens = [3000, 17, 9000, 0, 700]  # 160 (seq, head) pairs: more than the 148-CTA exchange grid
        case = make_case(len(lens), 32, 8, 128, 16, lens, torch.float16, seed=8)
        sh = ContextSharding(world, 16, 4)
        bt, ll, mx = sh.shard_batch(case["block_tables"], lens, rank)
        dev = torch.device("cuda", rank)
        d = to_device(case, dev)  # each rank reads only the pages its local table names
        grp = ContextShardGroup.create(len(lens) * 32, 128, rank)
        out = torch.empty_like(d["query"])
        for _ in range(3):
            grp.paged_attention(out, d["query"], d["key_cache"], d["value_cache"], 8, case["scale"],
                                bt.to(dev), ll.to(dev), 16, mx)
        torch.cuda.synchronize()
        assert grp.status() == 0
        # the same call replayed from a CUDA graph (device-managed epochs)
        dist.barrier()
        first = out.clone()
        out.fill_(float("nan"))
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(dev)
        btd, lld = bt.to(dev), ll.to(dev)
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                grp.paged_attention(out, d["query"], d["key_cache"], d["value_cache"], 8,
                                    case["scale"], btd, lld, 16, mx)
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        assert grp.status() == 0 and torch.equal(out, first)
        want = torch.empty_like(case["query"])
        oracle.paged_attention_v1(want, case["query"], case["key_cache"], case["value_cache"], 8,
                                  case["scale"], case["block_tables"], case["seq_lens"], 16,
                                  max(lens), None, "auto", 1.0, 1.0)
        err = (out.cpu().float() - want.float()).abs().max().item()
        assert err <= 2e-3, err
        dist.barrier()
        grp.close()
        open(os.path.join(tmpdir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_two_gpu_ipc_ranks(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run under gpurun --gpus 2)")
    mp.spawn(_ipc_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(2))
