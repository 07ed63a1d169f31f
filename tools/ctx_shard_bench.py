"""Strong-scaling microbench of context-sharded decode (light_vllm_b200/context_parallel.py).

A few very long sequences (default 4 x 128 Ki tokens, Llama-3-8B-shaped bf16) are decoded one token
per step.  Baseline: the whole context on ONE GPU (paged_attention_v2).  Sharded: the logical blocks
are dealt round-robin to N ranks, each rank streams 1/N of the pages and the ranks exchange one
(max, sum, out) triple per (seq, head) over NVLink inside the merge kernel.

    torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/ctx_shard_bench.py --steps 50
    python tools/ctx_shard_bench.py --local 2      # virtual ranks on one GPU (protocol check only)

Prints one JSON line (rank 0): per-step device time (CUDA events, max over ranks), aggregate KV
GB/s, speed-up over the single-GPU pass measured in the same process, max-abs difference between the
two outputs.  Every rank builds the same full pool from one seed (so rank 0 can run the unsharded
baseline) but reads only the pages its local block table names.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from light_vllm_b200 import _custom_ops as ops  # noqa: E402
from light_vllm_b200.context_parallel import ContextShardGroup, ContextSharding  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--ctx", type=int, default=131072)
    ap.add_argument("--layers", type=int, default=2, help="distinct pools cycled per step (> L2)")
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--chunk-blocks", type=int, default=4)
    ap.add_argument("--local", type=int, default=0, help="N virtual ranks on one GPU")
    ap.add_argument("--graph", action="store_true",
                    help="replay each step (one per layer pool) from a CUDA graph (one-process-per-GPU mode)")
    a = ap.parse_args()

    H, KH, D, BS = 32, 8, 128, 16
    dtype = torch.bfloat16
    distributed = "RANK" in os.environ and not a.local
    if distributed:
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        local_rank = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        rank, world, local_rank = 0, max(1, a.local), 0
        torch.cuda.set_device(0)
    dev = torch.device("cuda", local_rank)

    B, ctx = a.batch, a.ctx
    nblk = (ctx + BS - 1) // BS
    NB = B * nblk + 8
    g = torch.Generator(device=dev).manual_seed(1234)
    pools = []
    for _ in range(a.layers):
        kc = (torch.rand(NB, KH, D // 8, BS, 8, device=dev, generator=g, dtype=torch.float32) - 0.5).to(dtype)
        vc = (torch.rand(NB, KH, D, BS, device=dev, generator=g, dtype=torch.float32) - 0.5).to(dtype)
        pools.append((kc, vc))
    q = ((torch.rand(B, H, D, device=dev, generator=g) - 0.5) * 0.5).to(dtype)
    gc = torch.Generator().manual_seed(7)
    perm = torch.randperm(NB, generator=gc)[:B * nblk].to(torch.int32).view(B, nblk)
    lens = [ctx] * B
    scale = D ** -0.5
    kv_bytes = B * ctx * KH * D * 2 * 2  # K + V, bf16, per layer pool

    # ---- single-GPU baseline (whole context on this GPU) ----
    bt_full = perm.to(dev)
    sl_full = torch.tensor(lens, dtype=torch.int32, device=dev)
    P = (ctx + 511) // 512
    es = torch.empty(B, H, P, dtype=torch.float32, device=dev)
    ml = torch.empty_like(es)
    tmp = torch.empty(B, H, P, D, dtype=dtype, device=dev)
    out_ref = torch.empty_like(q)

    def base_step(i):
        kc, vc = pools[i % a.layers]
        ops.paged_attention_v2(out_ref, es, ml, tmp, q, kc, vc, KH, scale, bt_full, sl_full, BS, ctx,
                               None, "auto", 1.0, 1.0)

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize()
        if distributed:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        if distributed:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    base_ms = timed(base_step, a.steps, a.warmup)

    # ---- context-sharded ----
    sh = ContextSharding(world, BS, a.chunk_blocks)
    if distributed:
        groups = [ContextShardGroup.create(B * H, D, local_rank)]
        ranks = [rank]
    else:
        groups = ContextShardGroup.create_local(world, B * H, D, 0)
        ranks = list(range(world))
    streams = [torch.cuda.current_stream(dev)] if distributed else [torch.cuda.Stream(dev) for _ in ranks]
    local = []
    for r in ranks:
        bt, ll, mx = sh.shard_batch(perm, lens, r)
        local.append((bt.to(dev), ll.to(dev), mx))
    outs = [torch.empty_like(q) for _ in ranks]

    def shard_step(i):
        kc, vc = pools[i % a.layers]
        if distributed:
            bt, ll, mx = local[0]
            groups[0].paged_attention(outs[0], q, kc, vc, KH, scale, bt, ll, BS, mx)
            return
        cur = torch.cuda.current_stream(dev)
        for ph in (1, 2):  # virtual ranks: split-phase (see tests/test_context_parallel.py)
            for j, r in enumerate(ranks):
                streams[j].wait_stream(cur)
                with torch.cuda.stream(streams[j]):
                    bt, ll, mx = local[j]
                    groups[j].paged_attention(outs[j], q, kc, vc, KH, scale, bt, ll, BS, mx, phases=ph)
        for s in streams:
            cur.wait_stream(s)

    step_fn = shard_step
    if a.graph and distributed:
        # device-managed epochs make the sharded call graph-capturable: one graph per layer pool
        for i in range(a.layers):
            shard_step(i)
        torch.cuda.synchronize()
        dist.barrier()
        graphs = []
        side = torch.cuda.Stream(dev)
        for i in range(a.layers):
            g_ = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                with torch.cuda.graph(g_, stream=side):
                    shard_step(i)
            graphs.append(g_)
        torch.cuda.synchronize()
        dist.barrier()

        def step_fn(i):
            graphs[i % a.layers].replay()
    shard_ms = timed(step_fn, a.steps, a.warmup)
    status = [grp.status() for grp in groups]
    # same pool for the comparison
    base_step(0)
    step_fn(0)
    torch.cuda.synchronize()
    err = max((o.float() - out_ref.float()).abs().max().item() for o in outs)
    if distributed:
        t = torch.tensor([err], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        err = t.item()
    if rank == 0:
        print(json.dumps({
            "metric": "ctx_sharded_decode_kv_read_GBps", "unit": "GB/s", "n_gpus": world if distributed else 1,
            "virtual_ranks": 0 if distributed else world, "scaling": "strong",
            "cuda_graph": bool(a.graph and distributed),
            "value": round(kv_bytes / shard_ms / 1e6, 1), "ms_per_step": round(shard_ms, 4),
            "single_gpu_ms_per_step": round(base_ms, 4),
            "single_gpu_GBps": round(kv_bytes / base_ms / 1e6, 1),
            "speedup_vs_single_gpu": round(base_ms / shard_ms, 3),
            "max_abs_vs_single_gpu": err, "exchange_status": status, "steps": a.steps, "warmup": a.warmup,
            "dtype": "bf16",
            "config": {"workload": f"{B} x {ctx}-token sequences, H{H}/KH{KH}/D{D}/BS{BS}, context-sharded over "
                                   f"{world} ranks (chunk {a.chunk_blocks} pages)", "layer_pools": a.layers,
                       "kv_bytes_per_pool": kv_bytes}}), flush=True)
    if distributed:
        dist.barrier()
    for grp in groups:
        grp.close()
    if distributed:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
