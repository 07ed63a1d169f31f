#!/usr/bin/env python
"""BASELINE.json configs[0] at full size: Qwen2-0.5B-shaped random-init greedy decode, batch 8,
128-token synthetic prompts, 64 new tokens.

  CPU arm  : the reference's own CPU kernels (oracle/_ref/libref_cpu.so: csrc/cpu paged attention +
             reshape_and_cache), fp32, all host threads -- or the C oracle port if the library did
             not travel.  light-vllm has no CPU worker, so this is a harness around the ops
             (tests/qwen2_harness.py), not the reference engine.
  B200 arm : the same model through PagedB200AttentionImpl.forward in fp32 / fp16 / bf16.

Prints one JSON object: tokens/s of both arms, whether the fp32 B200 continuation equals the CPU
one token for token, the first divergence step of the free-running 16-bit runs, and the per-step
max-abs difference of the last layer's attention output under teacher forcing.

    python tools/c1_decode.py [--layers 24] [--batch 8] [--prompt 128] [--new 64]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from qwen2_harness import (B200Engine, CpuEngine, Qwen2Shape, graph_decode, greedy_decode,  # noqa: E402
                           init_weights, make_block_tables, make_prompts)


def first_divergence(a, b):
    neq = (a != b).any(dim=0).nonzero()
    return int(neq[0]) if neq.numel() else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=24)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--prompt", type=int, default=128)
    ap.add_argument("--new", type=int, default=64)
    ap.add_argument("--block-size", type=int, default=16)
    args = ap.parse_args()
    from oracle import oracle, ref
    shape = Qwen2Shape(layers=args.layers)
    BS = args.block_size
    w = init_weights(shape, 0)
    prompts = make_prompts(args.batch, args.prompt, shape.vocab, seed=1)
    bt, nb = make_block_tables(args.batch, args.prompt + args.new, BS, seed=2)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    mod, kind = (ref.ref_cpu(), "reference") if ref.have_ref_cpu() else (oracle, "port")

    cpu = greedy_decode(shape, w, CpuEngine(mod, shape, nb, BS, torch.float32), prompts, args.new,
                        BS, bt, record=True)
    n_tok = args.batch * args.new
    res = {"config": "Qwen2-0.5B-shaped random-init greedy decode, batch %d, %d-token synthetic "
                     "prompts, %d new tokens" % (args.batch, args.prompt, args.new),
           "layers": shape.layers, "block_size": BS,
           "cpu": {"kind": kind, "cores": cores, "dtype": "float32",
                   "prefill_s": round(cpu["prefill_s"], 4), "decode_s": round(cpu["decode_s"], 4),
                   "decode_tokens_per_s": round(n_tok / cpu["decode_s"], 1)}}
    if not torch.cuda.is_available():
        print(json.dumps(res))
        return
    from light_vllm_b200 import _lib
    for dtype in (torch.float32, torch.float16, torch.bfloat16):
        wd = {k: v.to("cuda", dtype) for k, v in w.items()}
        eng = B200Engine(shape, nb, BS, dtype, "cuda")
        greedy_decode(shape, wd, eng, prompts, 4, BS, bt, device="cuda")          # warm-up
        n0 = _lib.launch_count()
        free = greedy_decode(shape, wd, eng, prompts, args.new, BS, bt, device="cuda")
        launches = _lib.launch_count() - n0
        forced = greedy_decode(shape, wd, eng, prompts, args.new, BS, bt, device="cuda",
                               forced=cpu["tokens"], record=True)
        errs = [(a - b).abs().max().item() for a, b in zip(forced["attn"], cpu["attn"])]
        res[str(dtype).replace("torch.", "b200_")] = {
            "prefill_s": round(free["prefill_s"], 4), "decode_s": round(free["decode_s"], 4),
            "decode_tokens_per_s": round(n_tok / free["decode_s"], 1),
            "tokens_equal_cpu": bool(torch.equal(free["tokens"], cpu["tokens"])),
            "first_divergence_step": first_divergence(free["tokens"], cpu["tokens"]),
            "teacher_forced_token_agreement": round(
                (forced["tokens"] == cpu["tokens"]).float().mean().item(), 4),
            "attn_out_max_abs_vs_cpu_per_step": {"max": max(errs), "first": errs[0], "last": errs[-1]},
            "b200_kernel_launches": int(launches)}
        # the same decode with the whole step in ONE CUDA graph (fused rotary+cache write, paged
        # attention, advance_step on the GPU): no host work per token
        gr = graph_decode(shape, wd, prompts, args.new, BS, bt, nb, dtype)
        res[str(dtype).replace("torch.", "b200_") + "_cuda_graph"] = {
            "decode_s": round(gr["decode_s"], 4), "steps": gr["steps"],
            "decode_tokens_per_s": round(args.batch * gr["steps"] / gr["decode_s"], 1),
            "tokens_equal_eager": bool(torch.equal(gr["tokens"], free["tokens"])),
            "tokens_equal_cpu": bool(torch.equal(gr["tokens"], cpu["tokens"]))}
        del wd, eng
    print(json.dumps(res))


if __name__ == "__main__":
    main()
