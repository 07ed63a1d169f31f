"""Microbench of prefix / chunked-prefill attention over the paged cache
(light_vllm_b200.attention.prefix_prefill.context_attention_fwd, csrc/prefix_prefill.cu).

Workload (defaults): Llama-3-8B-shaped bf16 (32 q / 8 kv heads, head 128, block 16), B sequences
with `ctx` cached tokens each and `qlen` new tokens.  Reports device time (CUDA events), attention
TFLOP/s = 4*D*H*sum(qlen*ctx + qlen*(qlen+1)/2) / time, the fraction of the measured dense bf16
peak (MEASURED_PEAKS.json), for the tile loop picked with --path, and the same call served by
torch SDPA on a gathered contiguous copy of the context (a library baseline on the same box; the
gather is excluded from its time).  Prints one JSON line.
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from light_vllm_b200 import _lib  # noqa: E402
from light_vllm_b200.attention.prefix_prefill import context_attention_fwd  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--qlen", type=int, default=512)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--kv-heads", type=int, default=8)
    ap.add_argument("--head-size", type=int, default=128)
    ap.add_argument("--block-size", type=int, default=16)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--dtype", default="bfloat16")
    ap.add_argument("--no-scratch", action="store_true", help="CTA-local staging of the new k/v")
    ap.add_argument("--path", default="auto", choices=["auto", "mma_sync", "tcgen05", "tcgen05_gather"],
                    help="tile loop: library choice, the mma.sync kernel, or the tcgen05 / TMEM kernel")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    dtype = getattr(torch, a.dtype)
    B, ctx, ql, H, KH, D, BS = a.batch, a.ctx, a.qlen, a.heads, a.kv_heads, a.head_size, a.block_size
    T = B * ql
    g = torch.Generator(device=dev).manual_seed(3)

    def rnd(*shape):
        return ((torch.rand(*shape, device=dev, generator=g) - 0.5) * 0.5).to(dtype)

    q, k, v = rnd(T, H, D), rnd(T, KH, D), rnd(T, KH, D)
    nblk = (ctx + ql + BS - 1) // BS
    NB = B * nblk + 4
    kc, vc = rnd(NB, KH, D // 8, BS, 8), rnd(NB, KH, D, BS)
    bt = torch.randperm(NB, generator=torch.Generator().manual_seed(1))[:B * nblk].to(torch.int32).view(B, nblk).to(dev)
    start = torch.arange(0, T + 1, ql, dtype=torch.int32, device=dev)
    seq = torch.full((B,), ctx + ql, dtype=torch.int32, device=dev)
    cl = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    out = torch.empty_like(q)

    def ours():
        context_attention_fwd(q, k, v, out, kc, vc, bt, start[:-1], seq, cl, ql, None, None,
                              use_scratch=not a.no_scratch, path=a.path)

    # library baseline: contiguous K/V per sequence (context gathered from the pages + new tokens)
    nctx = (ctx + BS - 1) // BS
    kctx = kc[bt[:, :nctx].long()].permute(0, 2, 1, 4, 3, 5).reshape(B, KH, nctx * BS, D)[:, :, :ctx]
    vctx = vc[bt[:, :nctx].long()].permute(0, 2, 1, 4, 3).reshape(B, KH, nctx * BS, D)[:, :, :ctx]
    kall = torch.cat([kctx, k.view(B, ql, KH, D).transpose(1, 2)], dim=2).contiguous()
    vall = torch.cat([vctx, v.view(B, ql, KH, D).transpose(1, 2)], dim=2).contiguous()
    qb = q.view(B, ql, H, D).transpose(1, 2).contiguous()
    mask = torch.ones(ql, ctx + ql, dtype=torch.bool, device=dev).tril(diagonal=ctx)

    def sdpa():
        return F.scaled_dot_product_attention(qb, kall, vall, attn_mask=mask, enable_gqa=True)

    def timed(fn):
        for _ in range(a.warmup):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / a.steps

    n0 = _lib.launch_count()
    ms = timed(ours)
    launches = _lib.launch_count() - n0
    ref = sdpa().transpose(1, 2).reshape(T, H, D)
    err = (out.float() - ref.float()).abs().max().item()
    ms_sdpa = timed(sdpa)
    flops = 4.0 * D * H * B * (ql * ctx + ql * (ql + 1) / 2)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("bf16_tflops", 2250.0))
    tf = flops / ms / 1e9
    kv_bytes = B * (ctx + ql) * KH * D * 2 * 2
    print(json.dumps({
        "metric": "prefix_prefill_attention_TFLOPs", "value": round(tf, 1), "unit": "TFLOP/s",
        "ms_per_call": round(ms, 4), "tokens_per_s": round(T / ms * 1e3, 1),
        "roofline": {"bound": "tensor", "achieved": round(tf, 1), "peak": peak, "unit": "TFLOP/s",
                     "frac": round(tf / peak, 4),
                     "peak_source": "MEASURED_PEAKS.json bf16_tflops" if peaks else "datasheet",
                     "path": a.path},
        "kv_bytes_algorithmic": kv_bytes, "kv_GBps_algorithmic": round(kv_bytes / ms / 1e6, 1),
        "sdpa_library_ms": round(ms_sdpa, 4), "speedup_vs_sdpa": round(ms_sdpa / ms, 3),
        "max_abs_vs_sdpa": err, "gpu_launches": int(launches), "steps": a.steps, "warmup": a.warmup,
        "dtype": a.dtype, "staged_new_kv": not a.no_scratch,
        "config": {"workload": f"{B} seqs x (ctx {ctx} cached + {ql} new tokens), H{H}/KH{KH}/D{D}/BS{BS}"}}),
        flush=True)


if __name__ == "__main__":
    main()
