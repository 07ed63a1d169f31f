"""GPU parity (bit-exact) for the cache ops: reshape_and_cache[_flash] incl. fused fp8
quantisation, copy_blocks, swap_blocks, convert_fp8 -- product vs CPU oracle, and vs the
reference CUDA kernels when oracle/_ref/libref_cuda.so is present.  Whole pools are compared
byte for byte, so untouched blocks are covered too."""
import pytest
import torch

from kv_util import elem_x

pytestmark = pytest.mark.gpu


def raw(t):
    return t.contiguous().view(torch.uint8).cpu()


def make_kv_inputs(T, KH, D, BS, NB, dtype, cache_dtype, seed=0, outliers=True, stride_pad=0):
    g = torch.Generator().manual_seed(seed)
    s = D ** -0.5
    width = KH * D
    # key/value are slices of a fused qkv projection in the engine: token stride != KH*D
    kfull = ((torch.rand(T, width + stride_pad, generator=g) * 2 - 1) * s)
    vfull = ((torch.rand(T, width + stride_pad, generator=g) * 2 - 1) * s)
    if outliers:  # exercise fp8 saturation and the subnormal range
        kfull[torch.rand(T, width + stride_pad, generator=g) < 0.01] *= 3000.0
        vfull[torch.rand(T, width + stride_pad, generator=g) < 0.01] *= 1e-4
    key = kfull.to(dtype)[:, :width].view(T, KH, D)
    value = vfull.to(dtype)[:, :width].view(T, KH, D)
    x = elem_x(cache_dtype)
    kc = torch.randint(0, 255, (NB, KH, D // x, BS, x), dtype=torch.uint8, generator=g) \
        if cache_dtype == torch.uint8 else torch.randn(NB, KH, D // x, BS, x, generator=g).to(cache_dtype)
    vc = torch.randint(0, 255, (NB, KH, D, BS), dtype=torch.uint8, generator=g) \
        if cache_dtype == torch.uint8 else torch.randn(NB, KH, D, BS, generator=g).to(cache_dtype)
    slots = torch.randperm(NB * BS, generator=g)[:T].to(torch.int64)
    slots[::7] = -1  # padding tokens must be skipped
    return key, value, kc, vc, slots


CASES = [
    (torch.float16, "auto", 1.0, 1.0), (torch.bfloat16, "auto", 1.0, 1.0),
    (torch.float32, "auto", 1.0, 1.0),
    (torch.float16, "fp8", 1.0, 1.0), (torch.bfloat16, "fp8", 0.05, 0.02),
    (torch.bfloat16, "fp8_e4m3", 2.5, 0.3), (torch.float16, "fp8_e5m2", 0.05, 1.0),
    (torch.float32, "fp8", 0.01, 0.01),
]


@pytest.mark.parametrize("dtype,kv,ks,vs", CASES)
@pytest.mark.parametrize("shape", [(37, 8, 128, 16), (5, 2, 64, 16), (64, 4, 128, 32), (9, 3, 80, 8)])
def test_reshape_and_cache(ops, oracle, dtype, kv, ks, vs, shape):
    T, KH, D, BS = shape
    cache_dtype = dtype if kv == "auto" else torch.uint8
    if D % elem_x(cache_dtype):
        pytest.skip("head_size not divisible by x")
    key, value, kc, vc, slots = make_kv_inputs(T, KH, D, BS, 11, dtype, cache_dtype, seed=T, stride_pad=64)
    kc_o, vc_o = kc.clone(), vc.clone()
    oracle.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, ks, vs)
    kc_g, vc_g = kc.cuda(), vc.cuda()
    kg = torch.empty(T, KH * D + 64, dtype=dtype, device="cuda")[:, :KH * D].view(T, KH, D)
    vg = torch.empty(T, KH * D + 64, dtype=dtype, device="cuda")[:, :KH * D].view(T, KH, D)
    kg.copy_(key)
    vg.copy_(value)
    ops.reshape_and_cache(kg, vg, kc_g, vc_g, slots.cuda(), kv, ks, vs)
    assert torch.equal(raw(kc_g), raw(kc_o)), "key cache bytes differ from oracle"
    assert torch.equal(raw(vc_g), raw(vc_o)), "value cache bytes differ from oracle"


@pytest.mark.parametrize("dtype,kv,ks,vs", CASES)
def test_reshape_and_cache_flash(ops, oracle, dtype, kv, ks, vs):
    T, KH, D, BS, NB = 41, 4, 128, 16, 9
    cache_dtype = dtype if kv == "auto" else torch.uint8
    key, value, _, _, slots = make_kv_inputs(T, KH, D, BS, NB, dtype, cache_dtype, seed=3, stride_pad=32)
    g = torch.Generator().manual_seed(1)
    # the engine's flash pool is kv_cache[0] / kv_cache[1] of a [2, NB, BS, KH, D] tensor
    pool = torch.randint(0, 255, (2, NB, BS, KH, D), dtype=torch.uint8, generator=g) \
        if cache_dtype == torch.uint8 else torch.randn(2, NB, BS, KH, D, generator=g).to(cache_dtype)
    pool_o = pool.clone()
    oracle.reshape_and_cache_flash(key, value, pool_o[0], pool_o[1], slots, kv, ks, vs)
    pool_g = pool.cuda()
    ops.reshape_and_cache_flash(key.cuda(), value.cuda(), pool_g[0], pool_g[1], slots.cuda(), kv, ks, vs)
    assert torch.equal(raw(pool_g), raw(pool_o))


@pytest.mark.parametrize("cache_dtype", [torch.float16, torch.bfloat16, torch.uint8, torch.float32])
@pytest.mark.parametrize("pinned_mapping", [False, True])
def test_copy_blocks(ops, oracle, cache_dtype, pinned_mapping):
    L, NB, BS, KH, D = 5, 23, 16, 4, 128
    g = torch.Generator().manual_seed(2)
    pools = [torch.randint(0, 255, (2, NB, BS * KH * D * torch.empty((), dtype=cache_dtype).element_size()),
                           dtype=torch.uint8, generator=g) for _ in range(L)]
    perm = torch.randperm(NB, generator=g)
    mapping = torch.stack([perm[:6], perm[6:12]], dim=1).to(torch.int64)
    mapping = torch.cat([mapping, torch.tensor([[int(perm[0]), int(perm[13])]])])  # one src, two dsts
    want = [p.clone() for p in pools]
    oracle.copy_blocks([p[0] for p in want], [p[1] for p in want], mapping)
    got = [p.cuda().view(cache_dtype) for p in pools]
    m = mapping.pin_memory() if pinned_mapping else mapping.cuda()
    ops.copy_blocks([p[0] for p in got], [p[1] for p in got], m)
    torch.cuda.synchronize()
    for a, b in zip(got, want):
        assert torch.equal(raw(a), b)
    ops.copy_blocks([p[0] for p in got], [p[1] for p in got], mapping[:0].cuda())  # empty mapping: no-op
    ops.copy_blocks([], [], mapping.cuda())  # no layers: no-op


@pytest.mark.parametrize("direction", ["d2h", "h2d", "d2d"])
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("pinned", [True, False])
def test_swap_blocks(ops, oracle, direction, mode, pinned):
    from light_vllm_b200 import _lib
    if mode == 2 and not pinned and direction != "d2d":
        pytest.skip("kernel mode needs pinned host memory")
    NB_S, NB_D, BB = 40, 31, 16 * 8 * 128 * 2
    g = torch.Generator().manual_seed(4)
    src = torch.randint(0, 255, (NB_S, BB), dtype=torch.uint8, generator=g)
    dst = torch.randint(0, 255, (NB_D, BB), dtype=torch.uint8, generator=g)
    sp, dp = torch.randperm(NB_S, generator=g)[:17], torch.randperm(NB_D, generator=g)[:17]
    sp[3:9] = torch.arange(10, 16)  # a run of consecutive (src+1, dst+1) pairs
    dp[3:9] = torch.arange(20, 26)
    keep = [i for i in range(17) if i in range(3, 9) or (int(sp[i]) not in range(10, 16) and int(dp[i]) not in range(20, 26))]
    mapping = torch.stack([sp[keep], dp[keep]], dim=1).to(torch.int64)
    want = dst.clone()
    oracle.swap_blocks(src, want, mapping)
    host = (lambda t: t.pin_memory()) if pinned else (lambda t: t.clone())
    s = src.cuda() if direction in ("d2h", "d2d") else host(src)
    d = dst.cuda() if direction in ("h2d", "d2d") else host(dst)
    if mode == 0:    # the op, as the engine calls it
        ops.swap_blocks(s.view(torch.float16), d.view(torch.float16), mapping)
    else:            # copy mechanism forced: a per-call argument of the C ABI (no global state)
        _lib.swap_blocks_mode(s.view(torch.float16), d.view(torch.float16), mapping, mode)
    torch.cuda.synchronize()
    assert torch.equal(d.cpu(), want)
    with pytest.raises(RuntimeError, match="block_mapping must be on CPU"):
        ops.swap_blocks(src.cuda(), dst.cuda(), mapping.cuda())


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("scale", [1.0, 0.05, 3.0])
def test_convert_fp8_roundtrip(ops, oracle, dtype, scale):
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(64, 1031, generator=g) * torch.logspace(-5, 3, 1031)).to(dtype)
    q_o = torch.empty(x.shape, dtype=torch.uint8)
    oracle.convert_fp8(q_o, x, scale, "fp8")
    q_g = torch.empty(x.shape, dtype=torch.uint8, device="cuda")
    ops.convert_fp8(q_g, x.cuda(), scale, "fp8")
    assert torch.equal(q_g.cpu(), q_o), "quantised bytes differ"
    allcodes = torch.arange(256, dtype=torch.uint8).repeat(5)
    allcodes = allcodes[(allcodes & 0x7f) != 0x7f]  # NaN payloads are not part of the contract
    d_o = torch.empty(allcodes.shape, dtype=dtype)
    oracle.convert_fp8(d_o, allcodes, scale, "fp8_e4m3")
    d_g = torch.empty(allcodes.shape, dtype=dtype, device="cuda")
    ops.convert_fp8(d_g, allcodes.cuda(), scale, "fp8_e4m3")
    assert torch.equal(raw(d_g), raw(d_o)), "dequantised bytes differ"
    with pytest.raises(RuntimeError, match="Unsupported data type"):
        ops.convert_fp8(q_g, x.cuda(), scale, "int8")


def test_vs_reference_cuda_cache_kernels(ops):
    """Byte-for-byte against the reference's own cache_kernels.cu compiled for sm_100a."""
    from oracle import ref
    if not ref.have_ref_cuda():
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    _, rcache = ref.ref_cuda()
    for dtype, kv, ks, vs in CASES:
        if kv == "fp8_e5m2" or dtype == torch.float32 and kv != "auto":
            pass
        cache_dtype = dtype if kv == "auto" else torch.uint8
        key, value, kc, vc, slots = make_kv_inputs(53, 8, 128, 16, 13, dtype, cache_dtype, seed=9)
        a = [t.cuda() for t in (key, value, kc, vc, slots)]
        b = [t.cuda() for t in (key, value, kc, vc, slots)]
        ops.reshape_and_cache(*a, kv, ks, vs)
        rcache.reshape_and_cache(*b, kv, ks, vs)
        assert torch.equal(raw(a[2]), raw(b[2])) and torch.equal(raw(a[3]), raw(b[3])), (dtype, kv)
    x = (torch.randn(32, 4096) * torch.logspace(-5, 3, 4096)).half().cuda()
    q1 = torch.empty(x.shape, dtype=torch.uint8, device="cuda")
    q2 = torch.empty_like(q1)
    ops.convert_fp8(q1, x, 0.07, "fp8")
    rcache.convert_fp8(q2, x, 0.07, "fp8")
    assert torch.equal(q1, q2)
    d1, d2 = torch.empty_like(x), torch.empty_like(x)
    ops.convert_fp8(d1, q1, 0.07, "fp8")
    rcache.convert_fp8(d2, q1, 0.07, "fp8")
    assert torch.equal(raw(d1), raw(d2))


@pytest.mark.parametrize("seed", range(20))
def test_reshape_and_cache_fuzz_bit_exact(ops, oracle, seed):
    """Random token counts (1 .. thousands), head counts / sizes, block sizes, row paddings, dtypes
    and fp8 scales: cache bytes identical to the oracle's, untouched slots included."""
    import random
    r = random.Random(3000 + seed)
    dtype, kv, ks, vs = r.choice(CASES)
    cache_dtype = dtype if kv == "auto" else torch.uint8
    x = elem_x(cache_dtype)
    D = r.choice([d for d in (64, 80, 96, 112, 120, 128, 192, 256) if d % x == 0])
    KH = r.choice([1, 2, 3, 8])
    BS = r.choice([8, 16, 32])
    T = r.choice([1, 2, r.randint(3, 100), r.randint(500, 3000)])
    NB = (T + BS - 1) // BS + r.randint(1, 5)
    pad = r.choice([0, 8, 64])
    key, value, kc, vc, slots = make_kv_inputs(T, KH, D, BS, NB, dtype, cache_dtype, seed=seed,
                                               stride_pad=pad)
    kc_o, vc_o = kc.clone(), vc.clone()
    oracle.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, ks, vs)
    kc_g, vc_g = kc.cuda(), vc.cuda()
    kg = torch.empty(T, KH * D + pad, dtype=dtype, device="cuda")[:, :KH * D].view(T, KH, D)
    vg = torch.empty(T, KH * D + pad, dtype=dtype, device="cuda")[:, :KH * D].view(T, KH, D)
    kg.copy_(key)
    vg.copy_(value)
    ops.reshape_and_cache(kg, vg, kc_g, vc_g, slots.cuda(), kv, ks, vs)
    assert torch.equal(raw(kc_g), raw(kc_o)), (seed, dtype, kv, D, KH, BS, T)
    assert torch.equal(raw(vc_g), raw(vc_o)), (seed, dtype, kv, D, KH, BS, T)
