#!/usr/bin/env python
"""Tuning / comparison harness (not part of the product): times paged_attention for a config under
different launch plans (warps x ring stages x split [x q-heads per CTA]; split = v1 cluster size or v2
partitions per sequence) and, when oracle/_ref/libref_cuda.so
is present, the reference's own CUDA kernel recompiled for sm_100a on the same inputs.

    python tools/sweep_attn.py --config c2 [--plans "4,3,4;4,3,1;8,2,2"] [--ref] [--reps 30]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import CONFIGS, algorithmic_bytes_attention  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2")
    ap.add_argument("--plans", default="0,0,0")
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--version", type=int, default=0)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--ctx", type=int, default=0)
    ap.add_argument("--graph", type=int, default=0,
                    help="also time the calls issued N per CUDA graph (no host / launch-API time): graph_us")
    args = ap.parse_args()
    from light_vllm_b200 import _custom_ops as ops
    from light_vllm_b200 import _lib
    from light_vllm_b200.attention import PagedAttention
    c = dict(CONFIGS[args.config])
    if args.version:
        c["version"] = args.version
    if args.batch:
        c["B"] = args.batch
    if args.ctx:
        c["ctx"] = args.ctx
    dev = torch.device("cuda:0")
    dtype = getattr(torch, c["dtype"])
    kv = c["kv"]
    H, KH, D, BS, B, ctx, L = c["H"], c["KH"], c["D"], c["BS"], c["B"], c["ctx"], args.layers
    scale = D ** -0.5
    ks = vs = 1.0 if kv == "auto" else 0.05
    bps = (ctx + BS - 1) // BS
    NB = B * bps + 64
    torch.manual_seed(0)
    pools = []
    for _ in range(L):
        src = torch.empty((2, NB, BS * KH * D), dtype=dtype, device=dev).uniform_(-scale, scale)
        if kv != "auto":
            p = torch.empty(src.shape, dtype=torch.uint8, device=dev)
            ops.convert_fp8(p, src, ks, "fp8")
            src = p
        pools.append(src)
    split = [PagedAttention.split_kv_cache(p, KH, D) for p in pools]
    bt = torch.randperm(NB)[:B * bps].view(B, bps).int().to(dev)
    sl = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    q = torch.empty(B, H, D, dtype=dtype, device=dev).uniform_(-scale, scale)
    out = torch.empty_like(q)
    P = (ctx + 511) // 512
    scratch = (torch.empty(B, H, P, dtype=torch.float32, device=dev),
               torch.empty(B, H, P, dtype=torch.float32, device=dev),
               torch.empty(B, H, P, D, dtype=dtype, device=dev))

    def call(mod, l):
        kc, vc = split[l]
        if c["version"] == 1:
            mod.paged_attention_v1(out, q, kc, vc, KH, scale, bt, sl, BS, ctx, None, kv, ks, vs,
                                   0, 0, 0, 64, 0)
        else:
            mod.paged_attention_v2(out, *scratch, q, kc, vc, KH, scale, bt, sl, BS, ctx, None, kv,
                                   ks, vs, 0, 0, 0, 64, 0)

    def tuned_call(tuning):
        def f(_mod, l):
            kc, vc = split[l]
            _lib.paged_attention_tuned(c["version"], out, scratch if c["version"] == 2 else None, q,
                                       kc, vc, KH, scale, bt, sl, BS, ctx, kv, ks, vs, tuning)
        return f

    def time_it(mod, call=call):
        for l in range(L):
            call(mod, l)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.reps):
            call(mod, i % L)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.reps * 1e3

    def time_graph(mod, call, n):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            call(mod, 0)
            side.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=side):
                for i in range(n):
                    call(mod, i % L)
        torch.cuda.current_stream().wait_stream(side)
        reps = max(3, args.reps // n)
        for _ in range(2):
            gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps / n * 1e3

    ab = algorithmic_bytes_attention(c)
    for plan in args.plans.split(";"):
        tn = tuple(int(x) for x in plan.split(","))  # warps,stages,split[,head_rows]
        try:
            us = time_it(ops, tuned_call(tn))
            info = _lib.attention_plan(c["version"], B, H, KH, D, BS, ctx, dtype, kv, 0, tn)
            row = {"impl": "b200", "config": args.config, "v": c["version"], "B": B,
                   "ctx": ctx, "plan": plan, "us": round(us, 2),
                   "GBps": round(ab / us / 1e3, 1), "resolved": info}
            if args.graph:
                gus = time_graph(ops, tuned_call(tn), args.graph)
                row["graph_us"] = round(gus, 2)
                row["graph_GBps"] = round(ab / gus / 1e3, 1)
            print(json.dumps(row), flush=True)
        except RuntimeError as e:
            print(json.dumps({"impl": "b200", "plan": plan, "error": str(e)[:200]}), flush=True)
    if args.ref:
        from oracle import ref
        rC, _ = ref.ref_cuda()
        ref_out = torch.empty_like(q)
        us = time_it(rC)
        print(json.dumps({"impl": "reference-cuda-sm100a", "config": args.config, "v": c["version"],
                          "B": B, "ctx": ctx, "us": round(us, 2), "GBps": round(ab / us / 1e3, 1)}),
              flush=True)


if __name__ == "__main__":
    main()
