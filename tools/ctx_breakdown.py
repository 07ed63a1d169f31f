"""Where the context-sharded decode step spends its time (run under torchrun, one rank per GPU).

Per rank, on ITS shard of the pages (all CUDA-graph replays of 8 calls, device-timed, max over ranks):
  T_v2      ops.paged_attention_v2 on the local shard only           (partition pass + reduce, no exchange)
  T_self    ContextShardGroup of world 1 on the local shard           (partition pass + fold/publish/merge to self)
  T_shard   the real group                                           (+ NVLink publish, flag wait, skew)
  T_full    ops.paged_attention_v2 on the whole context (one GPU)
T_self - T_v2 = what the exchange kernel costs beyond the v2 reduce; T_shard - T_self = NVLink + skew.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/ctx_breakdown.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from light_vllm_b200 import _custom_ops as ops  # noqa: E402
from light_vllm_b200.context_parallel import ContextShardGroup, ContextSharding  # noqa: E402


def trace(grp):
    """Per-phase microseconds of CTA 0 in the last exchange launch + the gap since the previous launch's end."""
    import ctypes

    from light_vllm_b200 import _lib
    buf = (ctypes.c_uint64 * 32)()
    assert _lib.lib.b200_ctx_exchange_trace(grp._own, grp.world, buf, grp.device) == 0
    calls = [[(buf[c * 16 + 2 * i], buf[c * 16 + 2 * i + 1]) for i in range(8)] for c in range(2)]
    last, prev = sorted(calls, key=lambda c: c[0][0], reverse=True)
    names = ["parked", "stats_loaded", "folded", "published", "peers_arrived", "merged", "tail"]
    row = {n: round((last[i + 1][1] - last[i][1]) / 1.965e3, 2) for i, n in enumerate(names)}
    row["kernel_us(globaltimer)"] = round((last[7][0] - last[0][0]) / 1e3, 2)
    row["entry_after_prev_end_us"] = round((last[0][0] - prev[7][0]) / 1e3, 2)
    row["depwait_end_after_prev_end_us"] = round((last[1][0] - prev[7][0]) / 1e3, 2)
    return row


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    H, KH, D, BS, layers = 32, 8, 128, 16, 2
    ctx = int(sys.argv[1]) if len(sys.argv) > 1 else 131072   # e.g. 32768 at N = 2: the per-rank work of N = 8
    dtype, scale = torch.bfloat16, D ** -0.5
    rows = {}
    for B in (4, 1):
        nblk = ctx // BS
        NB = B * nblk + 8
        g = torch.Generator(device=dev).manual_seed(1234)
        pools = []
        for _ in range(layers):
            kc = torch.empty(NB, KH, D // 8, BS, 8, device=dev, dtype=dtype).uniform_(-0.5, 0.5, generator=g)
            vc = torch.empty(NB, KH, D, BS, device=dev, dtype=dtype).uniform_(-0.5, 0.5, generator=g)
            pools.append((kc, vc))
        q = torch.empty(B, H, D, device=dev, dtype=dtype).uniform_(-0.25, 0.25, generator=g)
        perm = torch.randperm(NB, generator=torch.Generator().manual_seed(7))[:B * nblk].int().view(B, nblk)
        lens = [ctx] * B
        bt_full, sl_full = perm.to(dev), torch.tensor(lens, dtype=torch.int32, device=dev)
        P = ctx // 512
        es = torch.empty(B, H, P, dtype=torch.float32, device=dev)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, H, P, D, dtype=dtype, device=dev)
        out = torch.empty_like(q)
        sh = ContextSharding(world, BS, 4)
        bt, ll, mx = sh.shard_batch(perm, lens, rank)
        bt, ll = bt.to(dev), ll.to(dev)
        grp = ContextShardGroup.create(B * H, D, dev.index)
        solo = ContextShardGroup.create_local(1, B * H, D, dev.index)[0]

        def f_full(i):
            kc, vc = pools[i % layers]
            ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, KH, scale, bt_full, sl_full, BS, ctx, None,
                                   "auto", 1.0, 1.0)

        def f_v2(i):
            kc, vc = pools[i % layers]
            ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, KH, scale, bt, ll, BS, mx, None, "auto", 1.0, 1.0)

        def f_self(i):
            kc, vc = pools[i % layers]
            solo.paged_attention(out, q, kc, vc, KH, scale, bt, ll, BS, mx)

        def f_shard(i):
            kc, vc = pools[i % layers]
            grp.paged_attention(out, q, kc, vc, KH, scale, bt, ll, BS, mx)

        def timed_graph(fn, n=8, reps=40):
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                fn(0)
                side.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=side):
                    for i in range(n):
                        fn(i)
            torch.cuda.current_stream(dev).wait_stream(side)
            for _ in range(5):
                gr.replay()
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                gr.replay()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / reps / n * 1e3], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return round(t.item(), 2)

        rows[f"{B}x{ctx}"] = {"T_full_us": timed_graph(f_full), "T_v2_local_us": timed_graph(f_v2),
                              "T_self_us": timed_graph(f_self), "trace_self": trace(solo),
                              "T_shard_us": timed_graph(f_shard), "trace_shard": trace(grp)}
        grp.close()
        solo.close()
        del pools, es, ml, tmp
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps({"world": world, "rows": rows}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
