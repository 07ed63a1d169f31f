"""CPU: pins the C oracle (oracle/paged_oracle.c) -- the checker every GPU parity test relies on.

1. against golden vectors produced by the reference's OWN CPU kernels (tests/golden/*.pt, made by
   tests/golden/make_golden.py from oracle/_ref/libref_cpu.so);
2. live against that library when it is present (build container);
3. the fp8 codecs exhaustively against torch's float8 dtypes (independent implementation), with
   the saturating behaviour of __NV_SATFINITE (quant_utils.cuh:458-489);
4. against an fp64 ground truth, incl. v1 == v2 and NaN-poisoned padding.
"""
import os

import pytest
import torch

from kv_util import make_case, naive_attention, v2_scratch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TAIL = (0, 0, 0, 64, 0)


def _args(c, alibi=None, kv="auto", ks=1.0, vs=1.0, tail=TAIL):
    return (c["query"], c["key_cache"], c["value_cache"], c["num_kv_heads"], c["scale"],
            c["block_tables"], c["seq_lens"], c["block_size"], c["max_seq_len"], alibi, kv, ks, vs,
            *tail)


def _v1(mod, c, **kw):
    out = torch.empty_like(c["query"])
    mod.paged_attention_v1(out, *_args(c, **kw))
    return out


def _v2(mod, c, **kw):
    out = torch.empty_like(c["query"])
    mod.paged_attention_v2(out, *v2_scratch(c), *_args(c, **kw))
    return out


@pytest.mark.parametrize("name,tol", [("f32", 2e-6), ("bf16", 4e-3)])
def test_attention_golden(oracle, name, tol):
    g = torch.load(os.path.join(GOLDEN, f"attn_{name}.pt"))
    c = g["case"]
    assert (_v1(oracle, c).float() - g["out_v1"].float()).abs().max() <= tol
    assert (_v2(oracle, c).float() - g["out_v2"].float()).abs().max() <= tol
    assert (_v1(oracle, c, alibi=g["alibi"]).float() - g["out_v1_alibi"].float()).abs().max() <= tol


@pytest.mark.parametrize("name", ["f32", "bf16"])
def test_cache_ops_golden(oracle, name):
    g = torch.load(os.path.join(GOLDEN, f"cache_{name}.pt"))
    kc, vc = g["kc_in"].clone(), g["vc_in"].clone()
    oracle.reshape_and_cache(g["key"], g["value"], kc, vc, g["slots"], "auto", 1.0, 1.0)
    assert torch.equal(kc.view(torch.uint8), g["kc_out"].view(torch.uint8))
    assert torch.equal(vc.view(torch.uint8), g["vc_out"].view(torch.uint8))
    oracle.copy_blocks([kc], [vc], g["mapping"])
    assert torch.equal(kc.view(torch.uint8), g["kc_copied"].view(torch.uint8))
    assert torch.equal(vc.view(torch.uint8), g["vc_copied"].view(torch.uint8))


def test_live_against_reference_cpu(oracle):
    from oracle import ref
    if not ref.have_ref_cpu():
        pytest.skip("oracle/_ref/libref_cpu.so not built (needs /root/reference)")
    try:
        rc = ref.ref_cpu()
    except OSError as e:  # e.g. a host without AVX512
        pytest.skip(f"cannot load libref_cpu.so: {e}")
    for dtype, tol in ((torch.float32, 2e-6), (torch.bfloat16, 4e-3)):
        for D in (64, 128, 256):
            lens = [1, 16, 31, 513, 700]
            c = make_case(len(lens), 8, 2, D, 16, lens, dtype, seed=D)
            assert (_v1(oracle, c).float() - _v1(rc, c).float()).abs().max() <= tol
            assert (_v2(oracle, c).float() - _v2(rc, c).float()).abs().max() <= tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-6), (torch.float16, 3e-4), (torch.bfloat16, 3e-3)])
def test_against_ground_truth(oracle, dtype, tol):
    lens = [1, 15, 16, 17, 511, 512, 513, 1030]
    c = make_case(len(lens), 8, 2, 64, 16, lens, dtype, seed=2, nan_padding=True, q_pad=32)
    gt = naive_attention(c)
    o1, o2 = _v1(oracle, c), _v2(oracle, c)
    assert (o1.double() - gt).abs().max() <= tol
    assert (o2.double() - gt).abs().max() <= tol
    alibi = torch.tensor([2.0 ** -(i + 1) for i in range(8)])
    gta = naive_attention(c, alibi_slopes=alibi)
    assert (_v1(oracle, c, alibi=alibi).double() - gta).abs().max() <= tol


@pytest.mark.parametrize("kv,tdt,mx", [("fp8_e4m3", torch.float8_e4m3fn, 448.0),
                                       ("fp8_e5m2", torch.float8_e5m2, 57344.0)])
def test_fp8_codec_against_torch(oracle, kv, tdt, mx):
    codes = torch.arange(256, dtype=torch.uint8)
    finite = ~codes.view(tdt).float().isnan() & ~codes.view(tdt).float().isinf()
    dec = torch.empty(256, dtype=torch.float32)
    oracle.convert_fp8(dec, codes, 1.0, kv)
    assert torch.equal(dec[finite], codes.view(tdt).float()[finite])
    g = torch.Generator().manual_seed(0)
    x = torch.randn(200000, generator=g) * torch.logspace(-6, 3.2, 200000)
    x[:4] = torch.tensor([0.0, -0.0, 1e9, -1e9])
    for scale in (1.0, 0.37):
        enc = torch.empty(x.shape, dtype=torch.uint8)
        oracle.convert_fp8(enc, x, scale, kv)
        want = (x / scale).clamp(-mx, mx).to(tdt).view(torch.uint8)   # saturate, then RN-even
        assert torch.equal(enc, want)
    # dequantise to 16-bit: half(fp8) * scale rounded once to T
    for dt in (torch.float16, torch.bfloat16):
        d = torch.empty(256, dtype=dt)
        oracle.convert_fp8(d, codes, 0.3, kv)
        want = (codes.view(tdt).float() * torch.tensor(0.3, dtype=torch.float32)).to(dt)
        assert torch.equal(d[finite].view(torch.int16), want[finite].view(torch.int16))


def test_fp8_attention_oracle_consistency(oracle):
    """fp8 attention in the oracle == 16-bit attention over the dequantised cache."""
    lens = [9, 300, 600]
    c = make_case(len(lens), 8, 2, 64, 16, lens, torch.float16, seed=4)
    kc8 = torch.empty(c["key_cache"].shape, dtype=torch.uint8)
    vc8 = torch.empty(c["value_cache"].shape, dtype=torch.uint8)
    # K layout for fp8 has x = 16: regroup [nb,kh,D/8,BS,8] -> [nb,kh,D/16,BS,16]
    nb, kh, _, bs, _ = c["key_cache"].shape
    klog = c["key_cache"].permute(0, 1, 3, 2, 4).reshape(nb, kh, bs, 64)
    k16 = klog.view(nb, kh, bs, 4, 16).permute(0, 1, 3, 2, 4).contiguous()
    oracle.convert_fp8(kc8, k16, 0.05, "fp8")
    oracle.convert_fp8(vc8, c["value_cache"].contiguous(), 0.02, "fp8")
    kd, vd = torch.empty_like(k16), torch.empty_like(c["value_cache"])
    oracle.convert_fp8(kd, kc8, 0.05, "fp8")
    oracle.convert_fp8(vd, vc8, 0.02, "fp8")
    c8 = dict(c, key_cache=kc8, value_cache=vc8)
    kd8 = kd.permute(0, 1, 3, 2, 4).reshape(nb, kh, bs, 64).view(nb, kh, bs, 8, 8).permute(0, 1, 3, 2, 4).contiguous()
    cd = dict(c, key_cache=kd8, value_cache=vd)
    a = _v1(oracle, c8, kv="fp8", ks=0.05, vs=0.02)
    b = _v1(oracle, cd)
    assert torch.equal(a, b)


def test_blocksparse_oracle_matches_masked_ground_truth(oracle):
    lens = [200, 70]
    c = make_case(len(lens), 4, 2, 64, 16, lens, torch.float32, seed=8)
    tail = (0, 2, 4, 32, 1)  # tp_rank, local_blocks, vert_stride, block, head_sliding_step
    got = _v1(oracle, c, tail=tail)
    # brute force with an explicit token mask (attention_kernels.cu:209-256)
    B, H, D = c["query"].shape
    out = torch.zeros(B, H, D, dtype=torch.float64)
    for b in range(B):
        n = int(c["seq_lens"][b])
        blks = c["block_tables"][b, :(n + 15) // 16].long()
        k = c["key_cache"][blks].permute(1, 0, 3, 2, 4).reshape(2, -1, D)[:, :n].double()
        v = c["value_cache"][blks].permute(1, 0, 3, 2).reshape(2, -1, D)[:, :n].double()
        for h in range(H):
            qb = (n - 1) // 32
            off = (0 * H + h) * 1 + 1
            kb = (torch.arange(n) // 16) * 16 // 32
            keep = ((kb + off) % 4 == 0) | (kb > qb - 2)
            logits = (k[h // 2] @ c["query"][b, h].double()) * c["scale"]
            logits[~keep] = float("-inf")
            out[b, h] = torch.softmax(logits, 0) @ v[h // 2]
    assert (got.double() - out).abs().max() <= 1e-5
