"""Drop-in boundary, checked with the reference's OWN Python: light_vllm/backends/_custom_ops.py is imported
unmodified (from /root/reference, CPU box only) with ``light_vllm._C`` served by this repository's
``light_vllm_b200/_C.so``; every hot-path wrapper (_custom_ops.py:72-130 paged attention, :134-178 rotary /
advance_step, :424-480 cache ops and device attributes) must resolve its ``torch.ops`` target and bind its
arguments to the registered schema.  The probe: call the wrapper with well-typed CPU tensors -- torch parses
the arguments against the schema first (a mismatch raises a schema error) and only then fails to find a CPU
kernel, which is the product's contract (CUDA dispatch key only, no fallback).

Runs in a subprocess: importing light_vllm must not leak into the other tests' module table.
"""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"

SCRIPT = textwrap.dedent(r'''
    import sys, types
    sys.path.insert(0, %(root)r)
    sys.path.insert(0, %(root)r + "/tests/golden")
    from make_golden_sched import _stub_pynvml
    _stub_pynvml()
    import torch
    from light_vllm_b200 import _lib            # loads libb200_paged.so + _C.so (registers torch.ops._C*)
    sys.modules["light_vllm._C"] = types.ModuleType("light_vllm._C")   # what `import light_vllm._C` finds
    sys.path.insert(0, %(ref)r)
    from light_vllm.backends import _custom_ops as ref_ops
    assert ref_ops.__file__.startswith(%(ref)r), ref_ops.__file__

    f16 = torch.float16
    B, H, KH, D, BS, NB, P = 2, 4, 2, 64, 16, 4, 1
    q = torch.zeros(B, H, D, dtype=f16); out = torch.zeros_like(q)
    kc = torch.zeros(NB, KH, D // 8, BS, 8, dtype=f16); vc = torch.zeros(NB, KH, D, BS, dtype=f16)
    bt = torch.zeros(B, 2, dtype=torch.int32); sl = torch.ones(B, dtype=torch.int32)
    k = torch.zeros(B, KH, D, dtype=f16); v = torch.zeros_like(k)
    slots = torch.zeros(B, dtype=torch.int64)
    es = torch.zeros(B, H, P); ml = torch.zeros(B, H, P); tmp = torch.zeros(B, H, P, D, dtype=f16)
    pos = torch.zeros(B, dtype=torch.int64); cs = torch.zeros(32, D, dtype=f16)
    kcf = torch.zeros(NB, BS, KH, D, dtype=f16)
    calls = {
        "paged_attention_v1": lambda: ref_ops.paged_attention_v1(out, q, kc, vc, KH, 0.1, bt, sl, BS, 16, None, "auto", 1.0, 1.0),
        "paged_attention_v2": lambda: ref_ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, KH, 0.1, bt, sl, BS, 16, None, "auto", 1.0, 1.0),
        "rotary_embedding": lambda: ref_ops.rotary_embedding(pos, q.view(B, -1), k.view(B, -1), D, cs, True),
        "batched_rotary_embedding": lambda: ref_ops.batched_rotary_embedding(pos, q.view(B, -1), k.view(B, -1), D, cs, True, D, pos),
        "advance_step": lambda: ref_ops.advance_step(B, B, BS, pos, pos.view(B, 1), pos, sl, slots, bt),
        "reshape_and_cache": lambda: ref_ops.reshape_and_cache(k, v, kc, vc, slots, "auto", 1.0, 1.0),
        "reshape_and_cache_flash": lambda: ref_ops.reshape_and_cache_flash(k, v, kcf, kcf.clone(), slots, "auto", 1.0, 1.0),
        "copy_blocks": lambda: ref_ops.copy_blocks([kc], [vc], torch.zeros(1, 2, dtype=torch.int64)),
        "swap_blocks": lambda: ref_ops.swap_blocks(kc, kc.clone(), torch.zeros(1, 2, dtype=torch.int64)),
        "convert_fp8": lambda: ref_ops.convert_fp8(torch.zeros(8, dtype=torch.uint8), torch.zeros(8, dtype=f16), 1.0, "fp8"),
    }
    ok = []
    for name, fn in calls.items():
        try:
            fn()
        except NotImplementedError as e:           # arguments bound; no CPU kernel by contract
            assert "CPU" in str(e) and name in str(e), (name, str(e)[:300])
            ok.append(name)
        except RuntimeError as e:                   # binding-level device check of the product
            # (advance_step validates its tensors itself: "tensor: name = ... is not as expected")
            assert "CUDA" in str(e) or "cuda" in str(e) or "tensor: name" in str(e), (name, str(e)[:300])
            ok.append(name)
        else:
            raise AssertionError(name + ": ran on CPU tensors -- a CPU fallback exists?")
    # schema-level mismatch must look different: proves the probe can fail
    try:
        torch.ops._C.paged_attention_v1(out, q, kc)
    except NotImplementedError:
        raise AssertionError("a truncated argument list reached dispatch")
    except (RuntimeError, TypeError):
        pass
    for name in ("get_device_attribute", "get_max_shared_memory_per_block_device_attribute"):
        assert callable(getattr(ref_ops, name)) and hasattr(torch.ops._C_cuda_utils, name)
    print("BOUND", ",".join(ok))
''')


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "light_vllm")), reason="needs /root/reference (CPU box)")
def test_reference_custom_ops_bind_to_this_library():
    r = subprocess.run([sys.executable, "-c", SCRIPT % {"root": ROOT, "ref": REF}], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("BOUND ")][-1]
    assert set(line.split()[1].split(",")) == {
        "paged_attention_v1", "paged_attention_v2", "rotary_embedding", "batched_rotary_embedding", "advance_step",
        "reshape_and_cache", "reshape_and_cache_flash", "copy_blocks", "swap_blocks", "convert_fp8"}


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "light_vllm")), reason="needs /root/reference (CPU box)")
def test_scheduler_stream_recorder_runs_against_the_reference():
    """The recorder still runs against the reference checkout and yields the same workload shape (steps,
    preemption counts) as the committed streams.  Block IDS are not compared: block_manager_v1 frees blocks
    in set-iteration order of objects, which differs between processes."""
    import json
    import tempfile
    script = textwrap.dedent(r'''
        import sys
        sys.path.insert(0, %(root)r + "/tests/golden")
        import make_golden_sched as m
        m.HERE = sys.argv[1]
        m.main()
    ''') % {"root": ROOT}
    with tempfile.TemporaryDirectory() as tmp:
        r = subprocess.run([sys.executable, "-c", script, tmp], capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stderr[-3000:]
        for name in ("recompute_v1", "swap_v1", "swap_v2"):
            a = json.load(open(os.path.join(tmp, f"sched_stream_{name}.json")))
            b = json.load(open(os.path.join(ROOT, "tests", "golden", f"sched_stream_{name}.json")))
            assert a["counts"]["steps"] == b["counts"]["steps"], name
            assert a["counts"]["cumulative_preemptions"] == b["counts"]["cumulative_preemptions"], name
            assert a["final"] == b["final"] and a["prompts"] == b["prompts"], name
