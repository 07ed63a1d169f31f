#!/usr/bin/env python
"""Generates tests/golden/rope.pt from the REFERENCE's own Python rotary embedding.

``RotaryEmbedding.forward_native`` (light_vllm/backends/rotary_embedding.py:116-163) is the
reference's PyTorch statement of what its CUDA kernel (csrc/pos_encoding_kernels.cu) computes,
including the cos/sin cache construction (:90-114).  The file is loaded on its own (importing the
``light_vllm`` package pulls in NVML / CUDA-only modules that cannot load in the CPU container):
its two imports, ``light_vllm.backends.custom_op`` (also loaded from the reference) and
``light_vllm.utils`` (only the ``is_*`` platform predicates are used), are satisfied by stubs.

    python tests/golden/make_golden_rope.py      # needs /root/reference
"""
import importlib.util
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200_REFERENCE", "/root/reference")


def load_reference_rope():
    pkg = types.ModuleType("light_vllm")
    pkg.__path__ = []
    backends = types.ModuleType("light_vllm.backends")
    backends.__path__ = []
    utils = types.ModuleType("light_vllm.utils")
    for name in ("is_tpu", "is_cpu", "is_hip", "is_xpu"):
        setattr(utils, name, lambda: False)
    sys.modules.update({"light_vllm": pkg, "light_vllm.backends": backends,
                        "light_vllm.utils": utils})
    mods = {}
    for name in ("custom_op", "rotary_embedding"):
        full = f"light_vllm.backends.{name}"
        spec = importlib.util.spec_from_file_location(
            full, os.path.join(REF, "light_vllm", "backends", f"{name}.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
        mods[name] = mod
    return mods["rotary_embedding"]


def main():
    rope = load_reference_rope()
    g = torch.Generator().manual_seed(7)
    T, H, KH, D, MAXPOS = 6, 4, 2, 64, 128
    cases = []
    for dtype in (torch.float16, torch.bfloat16, torch.float32):
        for is_neox in (True, False):
            for rot_dim in (64, 32):
                for batched in (False, True):
                    mod = rope.RotaryEmbedding(D, rot_dim, MAXPOS, 10000, is_neox, dtype)
                    q = torch.randn(T, H * D, generator=g).to(dtype)
                    k = torch.randn(T, KH * D, generator=g).to(dtype)
                    pos = torch.randint(0, MAXPOS // 2, (T,), generator=g)
                    off = torch.randint(0, MAXPOS // 2, (T,), generator=g) if batched else None
                    # forward_native assumes positions [batch, seq_len] (:143-146)
                    qo, ko = mod.forward_native(pos[None], q.clone()[None], k.clone()[None],
                                                off[None] if batched else None)
                    qo, ko = qo[0], ko[0]
                    cases.append(dict(dtype=dtype, is_neox=is_neox, rot_dim=rot_dim, head_size=D,
                                      positions=pos, offsets=off, query=q, key=k,
                                      cos_sin_cache=mod.cos_sin_cache.clone(),
                                      query_out=qo.contiguous(), key_out=ko.contiguous()))
    torch.save(cases, os.path.join(HERE, "rope.pt"))
    print(f"wrote {len(cases)} cases to tests/golden/rope.pt")


if __name__ == "__main__":
    main()
