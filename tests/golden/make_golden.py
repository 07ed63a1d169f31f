#!/usr/bin/env python
"""Generates tests/golden/*.pt from the REFERENCE's own compiled CPU kernels.

Run in the build container (needs oracle/_ref/libref_cpu.so, i.e. /root/reference present):
    make -C oracle ref_cpu && python tests/golden/make_golden.py
Each fixture stores the inputs and the outputs of torch.ops._ref_cpu.* (the unmodified
csrc/cpu/attention.cpp / cache.cpp) so that the C oracle -- and through it the CUDA
path -- stays pinned to the reference on machines where the reference cannot be built.
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))
from kv_util import make_case, v2_scratch  # noqa: E402
from oracle import ref  # noqa: E402


def main():
    rc = ref.ref_cpu()
    torch.set_num_threads(1)
    for name, dtype in (("f32", torch.float32), ("bf16", torch.bfloat16)):
        lens = [3, 17, 530]
        c = make_case(len(lens), 4, 2, 64, 16, lens, dtype, seed=42)
        args = (c["query"], c["key_cache"], c["value_cache"], 2, c["scale"], c["block_tables"],
                c["seq_lens"], 16, c["max_seq_len"], None, "auto", 1.0, 1.0, 0, 0, 0, 64, 0)
        out1 = torch.empty_like(c["query"])
        rc.paged_attention_v1(out1, *args)
        out2 = torch.empty_like(c["query"])
        es, ml, tmp = v2_scratch(c)
        rc.paged_attention_v2(out2, es, ml, tmp, *args)
        alibi = torch.tensor([0.5, 0.25, 0.125, 0.0625])
        out_alibi = torch.empty_like(c["query"])
        rc.paged_attention_v1(out_alibi, *args[:9], alibi, *args[10:])
        torch.save({"case": c, "out_v1": out1, "out_v2": out2, "alibi": alibi,
                    "out_v1_alibi": out_alibi}, os.path.join(HERE, f"attn_{name}.pt"))

        # reshape_and_cache + copy_blocks on a small pool
        g = torch.Generator().manual_seed(7)
        T, KH, D, BS, NB = 11, 2, 64, 16, 6
        x = 16 // torch.empty((), dtype=dtype).element_size()
        key = torch.randn(T, KH * D + 32, generator=g).to(dtype)[:, :KH * D].view(T, KH, D)
        value = torch.randn(T, KH * D + 32, generator=g).to(dtype)[:, :KH * D].view(T, KH, D)
        kc = torch.randn(NB, KH, D // x, BS, x, generator=g).to(dtype)
        vc = torch.randn(NB, KH, D, BS, generator=g).to(dtype)
        slots = torch.randperm(NB * BS, generator=g)[:T].to(torch.int64)
        kc_in, vc_in = kc.clone(), vc.clone()
        rc.reshape_and_cache(key, value, kc, vc, slots, "auto", 1.0, 1.0)
        kc2, vc2 = kc.clone(), vc.clone()
        mapping = torch.tensor([[0, 3], [1, 5], [0, 4]], dtype=torch.int64)
        rc.copy_blocks([kc2], [vc2], mapping)
        torch.save({"key": key.contiguous(), "value": value.contiguous(), "kc_in": kc_in,
                    "vc_in": vc_in, "slots": slots, "kc_out": kc, "vc_out": vc, "mapping": mapping,
                    "kc_copied": kc2, "vc_copied": vc2}, os.path.join(HERE, f"cache_{name}.pt"))
    print("wrote", sorted(f for f in os.listdir(HERE) if f.endswith(".pt")))


if __name__ == "__main__":
    main()
