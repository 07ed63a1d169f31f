#!/usr/bin/env python
"""Generates tests/golden/prefix.pt from the REFERENCE's own prefix-prefill kernel.

``context_attention_fwd`` (light_vllm/decoding/backends/attention/ops/prefix_prefill.py:674-811) is
a Triton kernel; it is run here unmodified on CPU tensors through Triton's interpreter
(``TRITON_INTERPRET=1``), so the committed vectors are outputs of the reference itself.  The file is
loaded on its own: its one package import, ``light_vllm.platforms.current_platform`` (used only for
``get_device_capability()`` to pick the tile size, :687-688), needs NVML and is satisfied by a stub
that reports (10, 0).

    python tests/golden/make_golden_prefix.py      # needs /root/reference; ~1 min
"""
import importlib.util
import os
import sys
import types

os.environ["TRITON_INTERPRET"] = "1"

import torch  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200_REFERENCE", "/root/reference")


def load_reference_prefix():
    pkg = types.ModuleType("light_vllm")
    pkg.__path__ = []
    plat = types.ModuleType("light_vllm.platforms")

    class _Platform:
        @staticmethod
        def get_device_capability(device_id: int = 0):
            return (10, 0)

    plat.current_platform = _Platform()
    sys.modules.update({"light_vllm": pkg, "light_vllm.platforms": plat})
    full = "light_vllm_ref_prefix_prefill"
    spec = importlib.util.spec_from_file_location(
        full, os.path.join(REF, "light_vllm", "decoding", "backends", "attention", "ops",
                           "prefix_prefill.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[full] = mod
    spec.loader.exec_module(mod)
    return mod


def make_prefix_case(ctx_lens, query_lens, H, KH, D, BS, dtype, seed, q_pad=0):
    """CPU tensors for one context_attention_fwd call: packed q/k/v of the new tokens, a paged cache
    holding ctx_len tokens per sequence (scattered blocks; the rest of the pool is random data that
    must not be read), block tables, start locations, lengths."""
    g = torch.Generator().manual_seed(seed)
    B = len(ctx_lens)
    T = sum(query_lens)
    s = D ** -0.5
    x = 16 // torch.empty((), dtype=dtype).element_size()

    def rnd(*shape):
        return ((torch.rand(*shape, generator=g) * 2 - 1) * s * 4).to(dtype)

    qfull = rnd(T, H * D + q_pad)
    q = qfull[:, :H * D].view(T, H, D)
    k, v = rnd(T, KH, D), rnd(T, KH, D)
    nblk = [(c + ql + BS - 1) // BS for c, ql in zip(ctx_lens, query_lens)]
    NB = sum(nblk) + 3
    kc, vc = rnd(NB, KH, D // x, BS, x), rnd(NB, KH, D, BS)
    perm = torch.randperm(NB, generator=g)
    bt = torch.zeros(B, max(1, max(nblk)), dtype=torch.int32)
    pos = 0
    for i, n in enumerate(nblk):
        bt[i, :n] = perm[pos:pos + n].to(torch.int32)
        pos += n
    start = torch.zeros(B + 1, dtype=torch.int32)
    start[1:] = torch.cumsum(torch.tensor(query_lens), 0)
    return dict(query=q, key=k, value=v, key_cache=kc, value_cache=vc, block_tables=bt,
                query_start_loc=start,
                seq_lens=torch.tensor([c + ql for c, ql in zip(ctx_lens, query_lens)], dtype=torch.int32),
                context_lens=torch.tensor(ctx_lens, dtype=torch.int32),
                max_query_len=max(query_lens), num_kv_heads=KH, block_size=BS)


CASES = [
    # (ctx_lens, query_lens, H, KH, D, BS, dtype, alibi, sliding_window)
    ([20, 0, 37], [5, 9, 1], 4, 2, 64, 16, torch.float32, False, 0),
    ([16, 3], [17, 33], 4, 1, 128, 16, torch.float16, False, 0),
    ([40, 7], [12, 20], 8, 2, 64, 8, torch.float16, False, 0),  # (the interpreter mishandles bf16)
    ([33, 0], [6, 25], 4, 2, 64, 16, torch.float32, True, 0),
    ([50, 5], [10, 30], 4, 2, 64, 16, torch.float32, False, 16),
    ([130], [140], 2, 2, 80, 32, torch.float16, False, 0),
]


def main():
    mod = load_reference_prefix()
    out = []
    for i, (ctx, ql, H, KH, D, BS, dtype, alibi, win) in enumerate(CASES):
        c = make_prefix_case(ctx, ql, H, KH, D, BS, dtype, seed=100 + i)
        slopes = None
        if alibi:
            slopes = torch.tensor([2.0 ** (-(j + 1) * 8.0 / H) for j in range(H)], dtype=torch.float32)
        o = torch.zeros_like(c["query"])
        mod.context_attention_fwd(c["query"], c["key"], c["value"], o, c["key_cache"], c["value_cache"],
                                  c["block_tables"], c["query_start_loc"][:-1], c["seq_lens"],
                                  c["context_lens"], c["max_query_len"], slopes, win if win else None)
        assert torch.isfinite(o.float()).all()
        out.append(dict(ctx_lens=ctx, query_lens=ql, H=H, KH=KH, D=D, BS=BS, dtype=dtype, seed=100 + i,
                        alibi_slopes=slopes, sliding_window=win, out=o.clone()))
        print(f"case {i}: ctx={ctx} q={ql} H{H}/KH{KH}/D{D}/BS{BS} {dtype} alibi={alibi} win={win} "
              f"|out|={o.float().abs().mean().item():.4f}", flush=True)
    torch.save(out, os.path.join(HERE, "prefix.pt"))
    print("wrote", os.path.join(HERE, "prefix.pt"))


if __name__ == "__main__":
    main()
