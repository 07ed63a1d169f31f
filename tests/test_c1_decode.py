"""BASELINE.json configs[0] (Qwen2-0.5B-shaped greedy decode over the paged cache) at unit-test
size: the harness against a paging-free dense recomputation (CPU), and the B200 plugin against
the reference's CPU kernels (GPU)."""
import pytest
import torch

from qwen2_harness import (B200Engine, CpuEngine, Qwen2Shape, dense_reference, greedy_decode,
                           init_weights, make_block_tables, make_prompts)

BS = 16


def _setup(batch=4, prompt_len=37, new=12):
    shape = Qwen2Shape.tiny()
    w = init_weights(shape, 0)
    prompts = make_prompts(batch, prompt_len, shape.vocab, seed=1)
    prompts[1] = prompts[1][:9]            # ragged prompts
    prompts[2] = prompts[2][:BS]           # exactly one block
    bt, nb = make_block_tables(batch, prompt_len + new, BS, seed=2)
    return shape, w, prompts, bt, nb, new


def _cpu_mods():
    from oracle import oracle, ref
    mods = [("oracle", oracle)]
    if ref.have_ref_cpu():
        mods.append(("reference_cpu", ref.ref_cpu()))
    return mods


def test_paged_cpu_engines_match_dense_recompute():
    shape, w, prompts, bt, nb, new = _setup()
    want = dense_reference(shape, w, prompts, new)
    for name, mod in _cpu_mods():
        got = greedy_decode(shape, w, CpuEngine(mod, shape, nb, BS, torch.float32), prompts, new,
                            BS, bt)["tokens"]
        assert torch.equal(got, want), name


@pytest.mark.gpu
def test_b200_fp32_matches_cpu_reference_tokens_and_attention():
    shape, w, prompts, bt, nb, new = _setup()
    name, mod = _cpu_mods()[-1]
    cpu = greedy_decode(shape, w, CpuEngine(mod, shape, nb, BS, torch.float32), prompts, new, BS,
                        bt, record=True)
    wd = {k: v.cuda() for k, v in w.items()}
    gpu = greedy_decode(shape, wd, B200Engine(shape, nb, BS, torch.float32, "cuda"), prompts, new,
                        BS, bt, device="cuda", record=True)
    assert torch.equal(gpu["tokens"], cpu["tokens"]), f"greedy tokens differ from {name}"
    for a, b in zip(gpu["attn"], cpu["attn"]):
        assert (a - b).abs().max().item() <= 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol,agree", [(torch.float16, 1e-2, 0.9), (torch.bfloat16, 5e-2, 0.75)])
def test_b200_16bit_teacher_forced(dtype, tol, agree):
    """16-bit weights/activations drift from the fp32 CPU run layer by layer, so the continuation
    is teacher-forced with the CPU tokens and compared per step."""
    shape, w, prompts, bt, nb, new = _setup()
    _, mod = _cpu_mods()[-1]
    cpu = greedy_decode(shape, w, CpuEngine(mod, shape, nb, BS, torch.float32), prompts, new, BS,
                        bt, record=True)
    wd = {k: v.to("cuda", dtype) for k, v in w.items()}
    gpu = greedy_decode(shape, wd, B200Engine(shape, nb, BS, dtype, "cuda"), prompts, new, BS, bt,
                        device="cuda", forced=cpu["tokens"], record=True)
    for a, b in zip(gpu["attn"], cpu["attn"]):
        assert (a - b).abs().max().item() <= tol
    assert (gpu["tokens"] == cpu["tokens"]).float().mean().item() >= agree


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_cuda_graph_decode_matches_eager(dtype):
    """Whole decode step in one CUDA graph (fused rotary+cache write, paged attention, on-GPU
    advance_step) == the eager plugin path, token for token."""
    from qwen2_harness import graph_decode
    shape, w, prompts, bt, nb, new = _setup()
    wd = {k: v.to("cuda", dtype) for k, v in w.items()}
    eager = greedy_decode(shape, wd, B200Engine(shape, nb, BS, dtype, "cuda"), prompts, new, BS, bt,
                          device="cuda")
    graph = graph_decode(shape, wd, prompts, new, BS, bt, nb, dtype)
    if dtype == torch.float32:
        assert torch.equal(graph["tokens"], eager["tokens"])
    else:  # fp16: rope runs in the op (3 roundings) vs torch (same), matmuls identical -> equal too,
        # but allow a late near-tie flip
        agree = (graph["tokens"] == eager["tokens"]).float().mean().item()
        assert agree >= 0.9, agree


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [(torch.float16, 2e-3), (torch.bfloat16, 1e-2)])
def test_identical_input_op_parity(oracle, dtype, tol):
    """Op-level parity INSIDE the c1 decode loop: at every layer of every decode step the oracle is
    fed the B200 arm's exact q and the exact pool the product just wrote, so the difference is the
    op's alone (BASELINE.json's 2e-3 / 1e-2 budget) -- unlike the teacher-forced model-level deltas
    of test_b200_16bit_teacher_forced, whose inputs drift layer by layer."""
    from light_vllm_b200.attention import PagedAttention
    shape, w, prompts, bt, nb, new = _setup(new=6)
    KH, D = shape.kv_heads, shape.head_dim

    class Probe(B200Engine):
        worst = 0.0
        calls = 0

        def decode(self, layer, q, k, v, slots, block_tables, seq_lens, max_len):
            o = super().decode(layer, q, k, v, slots, block_tables, seq_lens, max_len)
            kc, vc = PagedAttention.split_kv_cache(self.pools[layer], KH, D)
            qc = q.detach().cpu().contiguous()
            want = torch.empty_like(qc)
            oracle.paged_attention_v1(want, qc, kc.cpu().contiguous(), vc.cpu().contiguous(), KH,
                                      D ** -0.5, self._meta.block_tables.cpu().contiguous(),
                                      self._meta.seq_lens_tensor.cpu(), BS, max_len, None, "auto",
                                      1.0, 1.0)
            Probe.worst = max(Probe.worst, (o.float().cpu() - want.float()).abs().max().item())
            Probe.calls += 1
            return o

    wd = {k: v.to("cuda", dtype) for k, v in w.items()}
    greedy_decode(shape, wd, Probe(shape, nb, BS, dtype, "cuda"), prompts, new, BS, bt, device="cuda")
    assert Probe.calls == new * shape.layers
    assert Probe.worst <= tol, f"identical-input op parity {Probe.worst} > {tol}"
