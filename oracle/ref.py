"""TEST INFRASTRUCTURE ONLY -- loaders for the reference's OWN compiled kernels.

``oracle/Makefile`` compiles the reference sources where they lie (never copied) into
``oracle/_ref/``: ``libref_cpu.so`` (csrc/cpu/attention.cpp + cache.cpp -> torch.ops._ref_cpu) and
``libref_cuda.so`` (csrc/attention/attention_kernels.cu + csrc/cache_kernels.cu for sm_100a ->
torch.ops._ref_C / _ref_C_cache_ops).  On the GPU box /root/reference does not exist; only these
prebuilt files travel.
"""
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_CPU = os.path.join(_HERE, "_ref", "libref_cpu.so")
REF_CUDA = os.path.join(_HERE, "_ref", "libref_cuda.so")
_loaded = {}


def have_ref_cpu() -> bool:
    return os.path.exists(REF_CPU)


def have_ref_cuda() -> bool:
    return os.path.exists(REF_CUDA)


def ref_cpu():
    """torch.ops._ref_cpu: paged_attention_v1/v2, reshape_and_cache, copy_blocks (CPU tensors)."""
    if "cpu" not in _loaded:
        torch.ops.load_library(REF_CPU)
        _loaded["cpu"] = torch.ops._ref_cpu
    return _loaded["cpu"]


def ref_cuda():
    """(torch.ops._ref_C, torch.ops._ref_C_cache_ops): the reference CUDA kernels on sm_100a."""
    if "cuda" not in _loaded:
        torch.ops.load_library(REF_CUDA)
        _loaded["cuda"] = (torch.ops._ref_C, torch.ops._ref_C_cache_ops)
    return _loaded["cuda"]
