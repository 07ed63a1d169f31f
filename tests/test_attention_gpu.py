"""GPU parity: paged_attention_v1/v2 (product, through torch.ops._C -> C ABI) vs the CPU oracle
(oracle/paged_oracle.c), the fp64 ground truth, and -- when oracle/_ref/libref_cuda.so travelled
with the snapshot -- the reference's own CUDA kernels recompiled for sm_100a.

Tolerances are BASELINE.json's: max-abs 2e-3 (fp16), 1e-2 (bf16) against the reference op.
"""
import pytest
import torch

from kv_util import make_case, naive_attention, to_device, v2_scratch

pytestmark = pytest.mark.gpu
# float32 (no tolerance in BASELINE.json): exact fp32 products, ex2.approx softmax -> 2e-5 is ample
TOL = {torch.float16: 2e-3, torch.bfloat16: 1e-2, torch.float32: 2e-5}
TAIL = (0, 0, 0, 64, 0)  # tp_rank, local_blocks, vert_stride, bs_block_size, head_sliding_step


def attn_args(c, kv="auto", ks=1.0, vs=1.0, alibi=None, tail=TAIL):
    return (c["query"], c["key_cache"], c["value_cache"], c["num_kv_heads"], c["scale"],
            c["block_tables"], c["seq_lens"], c["block_size"], c["max_seq_len"], alibi, kv, ks, vs,
            *tail)


def run_v1(mod, c, **kw):
    out = torch.full_like(c["query"].contiguous(), float("nan"))
    mod.paged_attention_v1(out, *attn_args(c, **kw))
    return out


def run_v2(mod, c, **kw):
    out = torch.full_like(c["query"].contiguous(), float("nan"))
    es, ml, tmp = v2_scratch(c, device=c["query"].device)
    mod.paged_attention_v2(out, es, ml, tmp, *attn_args(c, **kw))
    return out


def check(ops, oracle, c, dtype, version, **kw):
    g = to_device(c, "cuda")
    run_gpu = run_v1 if version == 1 else run_v2
    got = run_gpu(ops, g, **{k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()})
    torch.cuda.synchronize()
    want = (run_v1 if version == 1 else run_v2)(oracle, c, **kw)
    live = c["seq_lens"] > 0
    assert not torch.isnan(got.float().cpu()[live]).any(), "NaN in output"
    err = (got.float().cpu()[live] - want.float()[live]).abs().max().item() if live.any() else 0.0
    assert err <= TOL[dtype], f"v{version} {dtype}: max-abs {err} vs oracle"
    return got


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("version", [1, 2])
def test_mixed_lengths_gqa(ops, oracle, dtype, version):
    # config-2 shape (32 q / 8 kv heads, D 128, BS 16) at oracle-friendly batch; ragged lengths
    lens = [1, 2, 15, 16, 17, 100, 511, 512, 513, 1024, 1500, 2048]
    c = make_case(len(lens), 32, 8, 128, 16, lens, dtype, seed=3, nan_padding=True)
    got = check(ops, oracle, c, dtype, version)
    gt = naive_attention(c)
    assert (got.double().cpu() - gt).abs().max().item() <= TOL[dtype]


@pytest.mark.parametrize("heads", [(8, 8), (8, 1), (28, 4), (14, 2), (40, 2)])
@pytest.mark.parametrize("version", [1, 2])
def test_head_groupings(ops, oracle, heads, version):
    H, KH = heads  # MHA, MQA, GQA-7 (Qwen2-7B), GQA-7 (Qwen2-0.5B), GQA-20 (> 16 rows)
    lens = [33, 700, 129, 1, 600]
    c = make_case(len(lens), H, KH, 128 if H != 14 else 64, 16, lens, torch.float16, seed=H)
    check(ops, oracle, c, torch.float16, version)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_strided_query_and_alibi(ops, oracle, dtype):
    lens = [40, 333, 1000]
    c = make_case(len(lens), 16, 4, 128, 16, lens, dtype, seed=5, q_pad=256)
    alibi = torch.tensor([2.0 ** (-(i + 1) / 2) for i in range(16)], dtype=torch.float32)
    for version in (1, 2):
        check(ops, oracle, c, dtype, version, alibi=alibi)


def test_single_dominant_logit(ops, oracle):
    # adversarial softmax: one token's logit dwarfs the rest
    lens = [700, 64]
    c = make_case(len(lens), 8, 2, 128, 16, lens, torch.float16, seed=6)
    blk = int(c["block_tables"][0, 20])
    c["key_cache"][blk, :, :, 3, :] = 4.0
    c["query"][0] = 1.0
    for version in (1, 2):
        check(ops, oracle, c, torch.float16, version)


def test_empty_batch_and_zero_len(ops, oracle):
    c = make_case(3, 8, 2, 128, 16, [0, 5, 0], torch.float16, seed=7)
    check(ops, oracle, c, torch.float16, 1)
    check(ops, oracle, c, torch.float16, 2)
    g = to_device(make_case(1, 8, 2, 128, 16, [5], torch.float16), "cuda")
    empty = dict(g)
    empty["query"] = g["query"][:0]
    empty["block_tables"] = g["block_tables"][:0]
    empty["seq_lens"] = g["seq_lens"][:0]
    out = torch.empty_like(empty["query"])
    ops.paged_attention_v1(out, *attn_args(empty))  # must be a no-op, not an error


def test_vs_reference_cuda_kernels(ops):
    """Differential test against the reference's own CUDA kernels (authoritative for fp16)."""
    from oracle import ref
    if not ref.have_ref_cuda():
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    rC, _ = ref.ref_cuda()
    for dtype in (torch.float16, torch.bfloat16):
        lens = [1, 17, 512, 513, 2048, 1999, 640, 3000]
        g = to_device(make_case(len(lens), 32, 8, 128, 16, lens, dtype, seed=11, nan_padding=True), "cuda")
        for version in (1, 2):
            run = run_v1 if version == 1 else run_v2
            got, want = run(ops, g), run(rC, g)
            torch.cuda.synchronize()
            err = (got.float() - want.float()).abs().max().item()
            assert err <= TOL[dtype], f"v{version} {dtype}: max-abs {err} vs reference CUDA kernel"


def test_unsupported_raises(ops):
    g = to_device(make_case(1, 8, 2, 128, 16, [5], torch.float16), "cuda")
    out = torch.empty_like(g["query"])
    with pytest.raises(RuntimeError, match="Unsupported"):
        ops.paged_attention_v1(out, *attn_args(g, kv="int4"))
    bad = dict(g, block_size=24)
    with pytest.raises(RuntimeError, match="block size"):
        ops.paged_attention_v1(out, *attn_args(bad))


def make_fp8_case(oracle, lens, H, KH, D, BS, dtype, kv, k_scale, v_scale, seed):
    """Quantise a 16-bit case's caches with the oracle's (pinned, bit-exact) fp8 codec."""
    c = make_case(len(lens), H, KH, D, BS, lens, dtype, seed=seed, cache_dtype=dtype)
    x = 16
    kc16 = c["key_cache"]                       # [nb, KH, D/x16, BS, x16] with x16 = 8
    nb = kc16.shape[0]
    # re-tile K to the fp8 layout (x = 16 elements per 16-byte chunk)
    k_logical = kc16.permute(0, 1, 3, 2, 4).reshape(nb, KH, BS, D)          # [nb, KH, BS, D]
    k8_src = k_logical.view(nb, KH, BS, D // x, x).permute(0, 1, 3, 2, 4).contiguous()
    kc8 = torch.empty(k8_src.shape, dtype=torch.uint8)
    vc8 = torch.empty(c["value_cache"].shape, dtype=torch.uint8)
    oracle.convert_fp8(kc8, k8_src, k_scale, kv)
    oracle.convert_fp8(vc8, c["value_cache"].contiguous(), v_scale, kv)
    c["key_cache"], c["value_cache"] = kc8, vc8
    return c


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("kv", ["fp8", "fp8_e5m2"])
@pytest.mark.parametrize("version", [1, 2])
def test_fp8_kv_cache(ops, oracle, dtype, kv, version):
    lens = [1, 9, 16, 31, 500, 513, 1100]
    for (H, KH, ks, vs) in ((28, 4, 0.05, 0.02), (8, 8, 1.0, 1.0), (32, 2, 0.3, 2.0)):
        c = make_fp8_case(oracle, lens, H, KH, 128, 16, dtype, kv, ks, vs, seed=H)
        check(ops, oracle, c, dtype, version, kv=kv, ks=ks, vs=vs)


def test_fp8_vs_reference_cuda_kernels(ops, oracle):
    from oracle import ref
    if not ref.have_ref_cuda():
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    rC, _ = ref.ref_cuda()
    lens = [3, 700, 2048, 1023, 64]
    for dtype in (torch.float16, torch.bfloat16):
        for kv in ("fp8", "fp8_e5m2"):
            c = make_fp8_case(oracle, lens, 28, 4, 128, 16, dtype, kv, 0.05, 0.03, seed=21)
            g = to_device(c, "cuda")
            for version in (1, 2):
                run = run_v1 if version == 1 else run_v2
                got = run(ops, g, kv=kv, ks=0.05, vs=0.03)
                want = run(rC, g, kv=kv, ks=0.05, vs=0.03)
                torch.cuda.synchronize()
                err = (got.float() - want.float()).abs().max().item()
                assert err <= TOL[dtype], f"{kv} v{version} {dtype}: max-abs {err} vs reference CUDA"


HEADS = [64, 80, 96, 112, 120, 128, 192, 256]   # attention_kernels.cu:738-769
BLOCKS = [8, 16, 32]                             # attention_kernels.cu:792-806


@pytest.mark.parametrize("head_size", HEADS)
@pytest.mark.parametrize("block_size", BLOCKS)
def test_dispatch_surface_auto(ops, oracle, head_size, block_size):
    lens = [1, block_size, block_size + 1, 257, 600]
    for dtype, (H, KH) in ((torch.float16, (8, 2)), (torch.bfloat16, (12, 1))):
        c = make_case(len(lens), H, KH, head_size, block_size, lens, dtype, seed=head_size + block_size,
                      nan_padding=True)
        for version in (1, 2):
            check(ops, oracle, c, dtype, version)


@pytest.mark.parametrize("head_size", [64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("block_size", BLOCKS)
def test_dispatch_surface_fp8(ops, oracle, head_size, block_size):
    lens = [1, block_size + 3, 300, 520]
    for dtype, kv, (H, KH) in ((torch.float16, "fp8", (8, 2)), (torch.bfloat16, "fp8_e5m2", (16, 1))):
        c = make_fp8_case(oracle, lens, H, KH, head_size, block_size, dtype, kv, 0.1, 0.2,
                          seed=head_size * 3 + block_size)
        for version in (1, 2):
            check(ops, oracle, c, dtype, version, kv=kv, ks=0.1, vs=0.2)


@pytest.mark.parametrize("tail", [(0, 2, 4, 64, 1), (1, 1, 3, 32, -1), (0, 16, 8, 16, 0), (0, 0, 2, 64, 3)])
def test_blocksparse(ops, oracle, tail):
    # tail = (tp_rank, local_blocks, vert_stride, blocksparse_block_size, head_sliding_step)
    lens = [70, 640, 1500, 129]
    c = make_case(len(lens), 16, 4, 128, 16, lens, torch.float16, seed=31)
    for version in (1, 2):
        check(ops, oracle, c, torch.float16, version, tail=tail)
    from oracle import ref
    if ref.have_ref_cuda():
        rC, _ = ref.ref_cuda()
        g = to_device(c, "cuda")
        for version in (1, 2):
            run = run_v1 if version == 1 else run_v2
            got, want = run(ops, g, tail=tail), run(rC, g, tail=tail)
            torch.cuda.synchronize()
            assert (got.float() - want.float()).abs().max().item() <= TOL[torch.float16]


def test_dispatch_vs_reference_cuda(ops):
    """Every (head size, block size) against the reference CUDA kernels, fp16 'auto'."""
    from oracle import ref
    if not ref.have_ref_cuda():
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    rC, _ = ref.ref_cuda()
    for D in HEADS:
        for BS in BLOCKS:
            lens = [5, 1000, 513]
            g = to_device(make_case(len(lens), 8, 2, D, BS, lens, torch.float16, seed=D + BS), "cuda")
            for version in (1, 2):
                run = run_v1 if version == 1 else run_v2
                got, want = run(ops, g), run(rC, g)
                torch.cuda.synchronize()
                err = (got.float() - want.float()).abs().max().item()
                assert err <= 2e-3, f"D={D} BS={BS} v{version}: {err}"


# ---- float32 queries (attention_kernels.cu:738-769 with scalar_t = float; CUDA-core kernel) ----
@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("heads", [(32, 8), (8, 8), (28, 4), (6, 1)])
def test_float32_auto(ops, oracle, heads, version):
    H, KH = heads
    lens = [1, 15, 16, 17, 511, 512, 513, 1200]
    c = make_case(len(lens), H, KH, 128, 16, lens, torch.float32, seed=41 + H, nan_padding=True)
    got = check(ops, oracle, c, torch.float32, version)
    gt = naive_attention(c)
    assert (got.double().cpu() - gt).abs().max().item() <= TOL[torch.float32]


@pytest.mark.parametrize("head_size", HEADS)
@pytest.mark.parametrize("block_size", BLOCKS)
def test_float32_dispatch_surface(ops, oracle, head_size, block_size):
    lens = [1, block_size, block_size + 1, 257, 600]
    alibi = torch.tensor([2.0 ** (-(i + 1) / 2) for i in range(8)], dtype=torch.float32)
    c = make_case(len(lens), 8, 2, head_size, block_size, lens, torch.float32,
                  seed=head_size + block_size, nan_padding=True, q_pad=64)
    for version in (1, 2):
        check(ops, oracle, c, torch.float32, version)
        check(ops, oracle, c, torch.float32, version, alibi=alibi)
    if head_size % 16 == 0:
        c8 = make_fp8_case(oracle, lens, 8, 2, head_size, block_size, torch.float32, "fp8", 0.1, 0.2,
                           seed=head_size)
        for version in (1, 2):
            check(ops, oracle, c8, torch.float32, version, kv="fp8", ks=0.1, vs=0.2)


def test_float32_blocksparse_and_e5m2(ops, oracle):
    lens = [70, 640, 1500, 129]
    c = make_case(len(lens), 16, 4, 128, 16, lens, torch.float32, seed=43)
    for version in (1, 2):
        check(ops, oracle, c, torch.float32, version, tail=(0, 2, 4, 64, 1))
    c8 = make_fp8_case(oracle, lens, 16, 16, 64, 16, torch.float32, "fp8_e5m2", 0.5, 1.5, seed=44)
    for version in (1, 2):
        check(ops, oracle, c8, torch.float32, version, kv="fp8_e5m2", ks=0.5, vs=1.5)


def test_float32_vs_reference_cuda_kernels(ops, oracle):
    from oracle import ref
    if not ref.have_ref_cuda():
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    rC, _ = ref.ref_cuda()
    lens = [1, 17, 512, 513, 2048, 1999]
    g = to_device(make_case(len(lens), 32, 8, 128, 16, lens, torch.float32, seed=45), "cuda")
    g8 = to_device(make_fp8_case(oracle, lens, 28, 4, 128, 16, torch.float32, "fp8", 0.05, 0.03,
                                 seed=46), "cuda")
    for version in (1, 2):
        run = run_v1 if version == 1 else run_v2
        got, want = run(ops, g), run(rC, g)
        torch.cuda.synchronize()
        assert (got - want).abs().max().item() <= 2e-5
        # float queries over an fp8 cache: the reference's float2 dequantisation goes through a
        # half2 (quant_utils.cuh:419-427 -> :304-315 rounds value*scale to fp16), so ITS result
        # carries 2^-11 relative rounding of every K/V element; the product kernel (and the oracle,
        # which follows the scalar conversion :406-415) keep fp32.  Checked at 1e-4 instead.
        got, want = run(ops, g8, kv="fp8", ks=0.05, vs=0.03), run(rC, g8, kv="fp8", ks=0.05, vs=0.03)
        torch.cuda.synchronize()
        assert (got - want).abs().max().item() <= 1e-4


@pytest.mark.parametrize("seed", range(32))
def test_fuzz_against_fp64_truth(ops, seed):
    """Random decode batches against the fp64 softmax: every head size / block size, group sizes the
    row tile does not divide, batch sizes that flip the launch plan between clusters, splits and
    plain CTAs, ragged lengths from 1 token to tens of thousands, ALiBi, strided queries; v1 and v2."""
    import random
    r = random.Random(5000 + seed)
    D = r.choice(HEADS)
    BS = r.choice([8, 16, 16, 32])
    G = r.choice([1, 2, 3, 4, 5, 7, 8, 16])
    KH = r.choice([1, 2, 4])
    H = G * KH
    dtype = r.choice([torch.float16, torch.bfloat16])
    B = r.choice([1, 2, 3, 7, 19, 40])
    longest = r.choice([40, 600, 3000, 20000 if B <= 3 else 3000])
    lens = [r.choice([1, r.randint(1, 2 * BS), r.randint(1, longest)]) for _ in range(B)]
    lens[r.randrange(B)] = longest
    slopes = None
    if r.random() < 0.3:
        slopes = torch.tensor([2.0 ** (-(j + 1) * 8.0 / H) for j in range(H)], dtype=torch.float32)
    c = make_case(B, H, KH, D, BS, lens, dtype, seed=seed, nan_padding=True, q_pad=r.choice([0, 8]))
    want = naive_attention(c, alibi_slopes=slopes)
    g = to_device(c, "cuda")
    for run in (run_v1, run_v2):
        got = run(ops, g, alibi=slopes.cuda() if slopes is not None else None)
        torch.cuda.synchronize()
        err = (got.double().cpu() - want).abs().max().item()
        assert err <= TOL[dtype], (seed, run.__name__, D, BS, G, KH, dtype, B, longest,
                                   slopes is not None, err)


@pytest.mark.parametrize("seed", range(16))
def test_fuzz_fp8_against_oracle(ops, oracle, seed):
    """Random decode batches over fp8 (e4m3 / e5m2) caches with random scales, v1 and v2, against the
    oracle (whose fp8 codec is pinned bit-exact on the reference's tables)."""
    import random
    r = random.Random(7000 + seed)
    D = r.choice([64, 80, 96, 112, 128, 192, 256])
    BS = r.choice([8, 16, 16, 32])
    G = r.choice([1, 2, 3, 4, 7, 8, 16])
    KH = r.choice([1, 2, 4])
    dtype = r.choice([torch.float16, torch.bfloat16])
    kv = r.choice(["fp8", "fp8_e5m2"])
    B = r.choice([1, 2, 5, 19, 40])
    longest = r.choice([40, 600, 3000, 12000 if B <= 2 else 3000])
    lens = [r.choice([1, r.randint(1, 2 * BS), r.randint(1, longest)]) for _ in range(B)]
    lens[r.randrange(B)] = longest
    ks, vs = r.choice([1.0, 0.05, 0.3]), r.choice([1.0, 0.02, 2.0])
    c = make_fp8_case(oracle, lens, G * KH, KH, D, BS, dtype, kv, ks, vs, seed=seed)
    for version in (1, 2):
        check(ops, oracle, c, dtype, version, kv=kv, ks=ks, vs=vs)


@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("H,KH,D,BS", [(32, 8, 128, 16), (7, 1, 80, 8), (16, 1, 256, 32)])
def test_writes_only_its_output_rows(ops, version, H, KH, D, BS):
    """Guard rows around the output keep their sentinel and the inputs (query, caches, block tables,
    lengths) are bit-identical after the call; zero-length sequences' rows are not required to be
    written and everything else matches the fp64 truth."""
    lens = [1, 0, 3 * BS, 700, 2300]
    c = make_case(len(lens), H, KH, D, BS, lens, torch.float16, seed=77, nan_padding=True)
    g = to_device(c, "cuda")
    B = len(lens)
    big = torch.full((B + 4, H, D), 777.0, dtype=torch.float16, device="cuda")
    out = big[2:-2]
    before = {n: g[n].clone() for n in ("query", "key_cache", "value_cache", "block_tables", "seq_lens")}
    if version == 1:
        ops.paged_attention_v1(out, *attn_args(g))
    else:
        es, ml, tmp = v2_scratch(c, device="cuda")
        ops.paged_attention_v2(out, es, ml, tmp, *attn_args(g))
    torch.cuda.synchronize()
    assert (big[:2] == 777.0).all() and (big[-2:] == 777.0).all()
    for n, t in before.items():
        same = torch.equal(t, g[n]) if t.dtype in (torch.int32, torch.int64) else \
            torch.equal(t.view(torch.uint8), g[n].view(torch.uint8))
        assert same, n
    live = torch.tensor(lens) > 0
    want = naive_attention(c)
    assert (out.double().cpu()[live] - want[live]).abs().max().item() <= 2e-3


@pytest.mark.parametrize("version", [1, 2])
def test_graph_capture_behind_copy_and_cache_write(ops, version):
    """The attention kernels are launched as programmatic dependents (griddepcontrol.wait at entry) and
    reshape_and_cache releases its dependents at entry.  Captured in a CUDA graph behind (a) a host-to-device copy
    node and (b) the cache write of the newest token, the replay must equal the same sequence run eagerly -- with
    the query and the new token changed between capture and replay."""
    lens = [1, 100, 513, 1024, 2047]
    B, H, KH, D, BS = len(lens), 32, 8, 128, 16
    c = to_device(make_case(B, H, KH, D, BS, lens, torch.float16, seed=21), "cuda")
    g = torch.Generator().manual_seed(5)
    s = c["scale"]
    q_host = [((torch.rand(B, H, D, generator=g) * 2 - 1) * s).half().pin_memory() for _ in range(2)]
    kv_host = [((torch.rand(2, B, KH, D, generator=g) * 2 - 1) * s).half().pin_memory() for _ in range(2)]
    slots = torch.tensor([int(c["block_tables"][i, (n - 1) // BS]) * BS + (n - 1) % BS for i, n in enumerate(lens)],
                         dtype=torch.int64, device="cuda")
    q_dev = torch.empty(B, H, D, dtype=torch.float16, device="cuda")
    kv_dev = torch.empty(2, B, KH, D, dtype=torch.float16, device="cuda")
    out = torch.empty_like(q_dev)
    es, ml, tmp = v2_scratch(c, device="cuda")

    def step(qh, kvh):
        q_dev.copy_(qh, non_blocking=True)      # memcpy node in front of the first kernel
        kv_dev.copy_(kvh, non_blocking=True)
        ops.reshape_and_cache(kv_dev[0], kv_dev[1], c["key_cache"], c["value_cache"], slots, "auto", 1.0, 1.0)
        args = (q_dev, c["key_cache"], c["value_cache"], KH, s, c["block_tables"], c["seq_lens"], BS,
                c["max_seq_len"], None, "auto", 1.0, 1.0, *TAIL)
        if version == 1:
            ops.paged_attention_v1(out, *args)
        else:
            ops.paged_attention_v2(out, es, ml, tmp, *args)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step(q_host[0], kv_host[0])
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            step(q_host[0], kv_host[0])
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    # replay on other data (the graph reads the same pinned buffers: overwrite them in place)
    q_host[0].copy_(q_host[1])
    kv_host[0].copy_(kv_host[1])
    graph.replay()
    torch.cuda.synchronize()
    got = out.clone()
    step(q_host[1], kv_host[1])
    torch.cuda.synchronize()
    assert not torch.isnan(got.float()).any()
    assert torch.equal(got, out)
