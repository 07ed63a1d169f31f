#!/usr/bin/env python
"""bench.py -- paged-attention decode hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c2|c3|c4] [--impl b200|reference]

A *step* is one pass of the hot path over one decode batch: for each of ``--layers`` KV pools
(the layers of a model; each pool is larger than L2, so KV reads always come from HBM):
``reshape_and_cache`` of the batch's newest token, then ``paged_attention`` over the whole context.
Default workload = BASELINE.json configs[1] (c2): fp16, 32 q / 8 kv heads, head 128, block 16,
batch 64, ctx 2048, paged_attention_v1.

Printed JSON (one line, rank 0):
  value      KV bytes read per second, whole job, inputs resident in HBM (device timed, max over ranks)
  e2e        same metric through the attention plugin with HOST (pinned) inputs: per step the
             H2D copy of q/k/v + metadata and the D2H read of the output are inside the timed region;
             the step is replayed as one CUDA graph (e2e.eager_* = the same step enqueued from Python)
  roofline   the attention kernel alone: algorithmic bytes per launch / CUDA-event launch time
  cpu_baseline  the reference's own CPU kernels (oracle/_ref/libref_cpu.so) or the C oracle port,
             timed on this box's host cores on a bounded sample
Multi-GPU: launched by torchrun, one process per GPU; sequences are sharded (each rank owns its
pools and block tables), no data-path collective; ``scaling`` = "weak".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: dtype, kv_cache_dtype, H, KH, D, BS, B, ctx, version, label
    "c2": dict(dtype="float16", kv="auto", H=32, KH=8, D=128, BS=16, B=64, ctx=2048, version=1,
               label="paged_attention_v1 microbench fp16: 32 q-heads/8 kv-heads, head_dim 128, "
                     "block_size 16, batch 64, ctx 2048"),
    "c3": dict(dtype="bfloat16", kv="auto", H=32, KH=8, D=128, BS=16, B=256, ctx=4096, version=2,
               label="Llama-3-8B-shaped decode bf16, batch 256, ctx 4096, paged_attention_v2"),
    "c4": dict(dtype="bfloat16", kv="fp8", H=28, KH=4, D=128, BS=16, B=512, ctx=8192, version=1,
               label="fp8_e4m3 KV cache Qwen2-7B-shaped decode, batch 512, ctx 8192"),
    # configs[4]: per GPU 16 sequences x 128 Ki tokens (512 MiB of KV per sequence per layer pool);
    # --ctx 32768|65536|131072 and --batch rescale it; sequences are sharded over the ranks.
    "c5": dict(dtype="bfloat16", kv="auto", H=32, KH=8, D=128, BS=16, B=16, ctx=131072, version=2,
               preempt=True,
               label="long-context decode bf16 (Llama-3-8B-shaped), 128 Ki-token sequences, "
                     "sequence-sharded KV pools, copy_blocks/swap_blocks under preemption"),
}


def kv_bytes_per_layer(c):
    es = 1 if c["kv"] != "auto" else 2
    return c["B"] * c["ctx"] * c["KH"] * c["D"] * 2 * es


def algorithmic_bytes_attention(c):
    """KV read + q + out + block-table + seq_lens bytes of one attention launch."""
    es = 1 if c["kv"] != "auto" else 2
    kv = c["B"] * c["ctx"] * c["KH"] * c["D"] * 2 * es
    qo = 2 * c["B"] * c["H"] * c["D"] * 2
    bt = 4 * c["B"] * ((c["ctx"] + c["BS"] - 1) // c["BS"])
    return kv + qo + bt + 4 * c["B"]


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def measured_peak_tflops():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["bf16_tflops"]), "MEASURED_PEAKS.json bf16_tflops (burst: kernel timed alone)"
    except Exception:
        return 1590.0, "fallback (B200_PROFILING.md)"


def prefix_prefill_record(dev, steps):
    """SURVEY 8f rank 4 on the driver-run line: context_attention_fwd (prefix / chunked prefill over the
    paged cache) at 16 sequences x (4096 cached + 512 new tokens), Llama-3-8B-shaped bf16 -- the tcgen05 /
    TMEM tile loop against the round-1 mma.sync tile loop of the same library, CUDA events, inputs
    (300 MB of K/V) larger than L2.  Parity: both tile loops against each other on the timed shape and
    against the CPU oracle on a small ragged case (the one place outside cpu_baseline where this file
    executes oracle/, as the checker)."""
    import torch
    from light_vllm_b200.attention.prefix_prefill import context_attention_fwd
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from make_golden_prefix import make_prefix_case
    from oracle import oracle

    B, ctx, ql, H, KH, D, BS = 16, 4096, 512, 32, 8, 128, 16
    dtype = torch.bfloat16
    g = torch.Generator(device=dev).manual_seed(3)

    def rnd(*shape):
        return ((torch.rand(*shape, device=dev, generator=g) - 0.5) * 0.5).to(dtype)

    T = B * ql
    q, k, v = rnd(T, H, D), rnd(T, KH, D), rnd(T, KH, D)
    nblk = (ctx + ql + BS - 1) // BS
    NB = B * nblk + 4
    kc, vc = rnd(NB, KH, D // 8, BS, 8), rnd(NB, KH, D, BS)
    bt = torch.randperm(NB, generator=torch.Generator().manual_seed(1))[:B * nblk].int().view(B, nblk).to(dev)
    start = torch.arange(0, T + 1, ql, dtype=torch.int32, device=dev)
    seq = torch.full((B,), ctx + ql, dtype=torch.int32, device=dev)
    cl = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    outs = {}
    ms = {}
    for path in ("tcgen05", "mma_sync"):
        out = torch.empty_like(q)

        def call():
            context_attention_fwd(q, k, v, out, kc, vc, bt, start[:-1], seq, cl, ql, None, None, path=path)

        for _ in range(3):
            call()
        torch.cuda.synchronize()
        n = max(5, steps if path == "tcgen05" else steps // 2)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            call()
        e1.record()
        torch.cuda.synchronize()
        ms[path] = e0.elapsed_time(e1) / n
        outs[path] = out
    cross = (outs["tcgen05"].float() - outs["mma_sync"].float()).abs().max().item()
    # oracle check on a small ragged case (seconds of scalar C)
    c = make_prefix_case([0, 77, 300], [130, 64, 257], 8, 2, D, BS, dtype, seed=5)
    d = {n: (t.to(dev) if isinstance(t, torch.Tensor) else t) for n, t in c.items()}
    got = torch.empty_like(d["query"])
    context_attention_fwd(d["query"], d["key"], d["value"], got, d["key_cache"], d["value_cache"],
                          d["block_tables"], d["query_start_loc"][:-1], d["seq_lens"], d["context_lens"],
                          c["max_query_len"], None, None, path="tcgen05")
    want = torch.zeros_like(c["query"])
    oracle.context_attention_fwd(c["query"], c["key"], c["value"], want, c["key_cache"], c["value_cache"],
                                 c["block_tables"], c["query_start_loc"][:-1], c["seq_lens"],
                                 c["context_lens"], c["max_query_len"], None, None)
    err = (got.float().cpu() - want.float()).abs().max().item()
    flops = 4.0 * D * H * B * (ql * ctx + ql * (ql + 1) / 2)
    peak, src = measured_peak_tflops()
    tf = flops / ms["tcgen05"] / 1e9
    if err > 1e-2 or cross > 1e-2:
        raise RuntimeError(f"prefix prefill parity failed: oracle {err}, tile loops {cross}")
    return {"workload": f"{B} seqs x ({ctx} cached + {ql} new tokens), bf16 H{H}/KH{KH}/D{D}/BS{BS}",
            "kernel": "context_attention_umma_kernel (tcgen05.mma, TMEM accumulators)",
            "ms_per_call": round(ms["tcgen05"], 4), "TFLOPs": round(tf, 1),
            "roofline": {"bound": "tensor", "achieved": round(tf, 1), "peak": peak, "unit": "TFLOP/s",
                         "frac": round(tf / peak, 4), "peak_source": src},
            "mma_sync_ms_per_call": round(ms["mma_sync"], 4),
            "speedup_vs_mma_sync": round(ms["mma_sync"] / ms["tcgen05"], 3),
            "parity": {"max_abs_vs_oracle_small_case": err, "max_abs_tcgen05_vs_mma_sync": cross, "tol": 1e-2}}


def other_config_record(name, dev, steps):
    """One more BASELINE.json config on the driver-run line: the attention launch alone at the config's
    full shape (one pool, larger than L2, cycled halves are not needed: a pool is >= 4 GiB), CUDA events."""
    import torch
    from light_vllm_b200 import _custom_ops as ops
    from light_vllm_b200 import _lib
    from light_vllm_b200.attention import PagedAttention, PagedB200AttentionBackend

    c = CONFIGS[name]
    dtype = getattr(torch, c["dtype"])
    kv = c["kv"]
    H, KH, D, BS, B, ctx = c["H"], c["KH"], c["D"], c["BS"], c["B"], c["ctx"]
    scale = D ** -0.5
    k_scale = v_scale = 1.0 if kv == "auto" else 0.05
    bps = (ctx + BS - 1) // BS
    NB = B * bps + 64
    shape = PagedB200AttentionBackend.get_kv_cache_shape(NB, BS, KH, D)
    if kv == "auto":
        pool = torch.empty(shape, dtype=dtype, device=dev).uniform_(-scale, scale)
    else:
        pool = torch.empty(shape, dtype=torch.uint8, device=dev)
        half = NB // 2  # convert in two halves: the 16-bit source of a whole pool would double the footprint
        for lo, hi in ((0, half), (half, NB)):
            src = torch.empty((2, hi - lo) + tuple(shape[2:]), dtype=dtype, device=dev).uniform_(-scale, scale)
            for t in range(2):
                ops.convert_fp8(pool[t, lo:hi], src[t], k_scale, "fp8")
            del src
    kc, vc = PagedAttention.split_kv_cache(pool, KH, D)
    bt = torch.randperm(NB - 64)[:B * bps].view(B, bps).int().to(dev)
    sl = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    q = torch.empty(B, H, D, dtype=dtype, device=dev).uniform_(-scale, scale)
    out = torch.empty_like(q)
    P = (ctx + 511) // 512
    scratch = (torch.empty(B, H, P, dtype=torch.float32, device=dev),
               torch.empty(B, H, P, dtype=torch.float32, device=dev),
               torch.empty(B, H, P, D, dtype=dtype, device=dev))

    def call():
        if c["version"] == 1:
            ops.paged_attention_v1(out, q, kc, vc, KH, scale, bt, sl, BS, ctx, None, kv, k_scale, v_scale,
                                   0, 0, 0, 64, 0)
        else:
            ops.paged_attention_v2(out, *scratch, q, kc, vc, KH, scale, bt, sl, BS, ctx, None, kv, k_scale,
                                   v_scale, 0, 0, 0, 64, 0)

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    n = max(5, steps // 2)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    if not torch.isfinite(out.float()).all():
        raise RuntimeError(f"{name}: non-finite attention output")
    peak, src = measured_peak_gbs()
    alg = algorithmic_bytes_attention(c)
    gbs = alg / ms / 1e6
    del pool, kc, vc, scratch
    torch.cuda.empty_cache()
    return {"workload": c["label"], "kv_cache_dtype": kv, "us_per_launch": round(ms * 1e3, 1),
            "algorithmic_bytes": alg, "decode_tokens_per_s_per_layer": round(B / ms * 1e3, 1),
            "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(gbs / peak, 4), "peak_source": src},
            "l2": f"one {alg >> 20} MiB pass per launch, far larger than L2",
            "plan": _lib.attention_plan(c["version"], B, H, KH, D, BS, ctx, dtype, kv, dev.index or 0)}


def ncu_traffic(config_name):
    """dram__bytes_read.sum + dram__bytes_write.sum of one attention launch, from the committed
    `ncu --set full` capture of this config (profiles/ncu_traffic.json); None if never captured."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return float(json.load(f)[config_name]["traffic"])
    except Exception:
        return None


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (NVML every ~0.5 ms in a thread -- the
    default timed region lasts only ~7 ms;
    falls back to `nvidia-smi -lms` when the NVML binding is unavailable)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []
        self.sm, self.mx, self.reasons, self._stop, self.thread = [], None, set(), False, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v for v in vis.split(",") if v.strip().isdigit()]
            phys = int(ids[self.index]) if self.index < len(ids) else self.index
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            bits = {"hw_slowdown": pynvml.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": pynvml.nvmlClocksThrottleReasonSwPowerCap}

            def poll():
                while not self._stop:
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        for n, b in bits.items():
                            if r & b:
                                self.reasons.add(n)
                    except Exception:
                        pass
                    time.sleep(0.0005)

            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.thread = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                 "100", "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.thread is not None:
            self._stop = True
            self.thread.join(timeout=1.0)
            sm = sorted(self.sm)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx,
                    "reasons": sorted(self.reasons), "samples": len(sm), "source": "nvml"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


# --------------------------------------------------------------------------------------------
# CPU side: the reference's own CPU kernels (oracle/_ref/libref_cpu.so) or the C oracle port.
# The only places bench.py executes anything under oracle/: cpu_baseline and --impl reference.
# --------------------------------------------------------------------------------------------
def cpu_step_fn(c, sample_B):
    """Returns (step_fn, kind, dtype_name, cores).  One step = reshape_and_cache + attention on a
    batch slice of ``sample_B`` sequences of the same ctx (the CPU path has no fp16/fp8: those
    configs run in bf16, same bytes per element as fp16; stated in `sample`)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from kv_util import make_case, v2_scratch
    from oracle import ref
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    use_ref = ref.have_ref_cpu()
    dtype = torch.bfloat16 if use_ref else getattr(torch, c["dtype"])
    H, KH, D, BS, ctx = c["H"], c["KH"], c["D"], c["BS"], c["ctx"]
    case = make_case(sample_B, H, KH, D, BS, [ctx] * sample_B, dtype, seed=0)
    g = torch.Generator().manual_seed(1)
    key = ((torch.rand(sample_B, KH, D, generator=g) * 2 - 1) * case["scale"]).to(dtype)
    value = ((torch.rand(sample_B, KH, D, generator=g) * 2 - 1) * case["scale"]).to(dtype)
    slots = torch.tensor([int(case["block_tables"][i, (ctx - 1) // BS]) * BS + (ctx - 1) % BS
                          for i in range(sample_B)], dtype=torch.int64)
    out = torch.empty_like(case["query"])
    if use_ref:
        mod, kind = ref.ref_cpu(), "reference"
    else:
        from oracle import oracle as mod
        kind = "port"
    attn_tail = (case["query"], case["key_cache"], case["value_cache"], KH, case["scale"],
                 case["block_tables"], case["seq_lens"], BS, ctx, None, "auto", 1.0, 1.0, 0, 0, 0,
                 64, 0)
    scratch = v2_scratch(case)

    def step():
        mod.reshape_and_cache(key, value, case["key_cache"], case["value_cache"], slots, "auto",
                              1.0, 1.0)
        if c["version"] == 1:
            mod.paged_attention_v1(out, *attn_tail)
        else:
            mod.paged_attention_v2(out, *scratch, *attn_tail)

    return step, kind, str(dtype).replace("torch.", ""), cores


def time_cpu(c, budget_s, sample_B=None, steps=None, warmup=1):
    # ~512 MiB of KV per step (well above the host's last-level cache), never more than the batch
    per_seq = c["ctx"] * c["KH"] * c["D"] * 2 * 2
    sample_B = sample_B or max(1, min(c["B"], (512 << 20) // per_seq))
    step, kind, dt, cores = cpu_step_fn(c, sample_B)
    for _ in range(warmup):
        step()
    times = []
    t_end = time.perf_counter() + budget_s
    while (steps is None and time.perf_counter() < t_end and len(times) < 50) or \
            (steps is not None and len(times) < steps):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
    es = 2
    kv_bytes = sample_B * c["ctx"] * c["KH"] * c["D"] * 2 * es
    best, mean = min(times), sum(times) / len(times)
    return dict(kind=kind, cores=cores, dtype=dt, sample_B=sample_B, best_s=best, mean_s=mean,
                n=len(times), kv_bytes=kv_bytes)


def run_reference_arm(args, c):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # >= 3 untimed steps: the first calls pay OpenMP thread start-up and first-touch page faults
    r = time_cpu(c, budget_s=1e9, steps=args.steps, warmup=max(args.warmup, 3))
    gbs = r["kv_bytes"] / r["mean_s"] / 1e9
    sample = (f"{r['sample_B']} of {c['B']} sequences x ctx {c['ctx']}, 1 layer pool, {r['dtype']} "
              f"(CPU path has no fp16/fp8), reshape_and_cache + paged_attention_v{c['version']} "
              f"per step, {r['n']} steps, all {r['cores']} host threads")
    line = {
        "impl": "reference", "metric": "paged_attn_kv_read_GBps", "value": round(gbs, 3),
        "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": round(r["mean_s"] * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": r["dtype"], "data": "synthetic",
        "config": {"workload": c["label"], "name": args.config},
        "decode_tokens_per_s_per_layer": round(r["sample_B"] / r["mean_s"], 1),
        "cpu_baseline": {"value": round(gbs, 3), "unit": "GB/s", "cores": r["cores"],
                         "kind": r["kind"], "sample": sample},
        "e2e": {"value": round(gbs, 3), "unit": "GB/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)



# --------------------------------------------------------------------------------------------
# Checker legs (untimed region; the only other places bench.py touches oracle/): the step's output
# against the CPU oracle on a slice of the batch, and the reference's own CUDA kernels
# (oracle/_ref/libref_cuda.so, recompiled for sm_100a) timed on the same pools -- SURVEY 0.3 / 8d's bar.
# --------------------------------------------------------------------------------------------
def oracle_slice_check(c, q, out, kc, vc, block_tables, seq_lens, scale, k_scale, v_scale):
    """max-abs of the product's `out` rows vs oracle/paged_oracle.c on up to 4 sequences of the
    batch (their pages gathered out of the pool and renumbered); asserts BASELINE.json's budget."""
    import torch
    from oracle import oracle
    B, BS, KH, ctx = c["B"], c["BS"], c["KH"], c["ctx"]
    per_seq = ctx * KH * c["D"] * 2 * kc.element_size()
    n = max(1, min(4, B, (1 << 30) // max(per_seq, 1)))
    rows = sorted({0, B // 3, (2 * B) // 3, B - 1})[:n]
    bps = (ctx + BS - 1) // BS
    blks = block_tables[rows].long().reshape(-1)
    case_kc, case_vc = kc[blks].cpu().contiguous(), vc[blks].cpu().contiguous()
    bt = torch.arange(len(rows) * bps, dtype=torch.int32).view(len(rows), bps)
    qs = q[rows].cpu().contiguous()
    want = torch.empty_like(qs)
    lens = seq_lens[rows].cpu()
    tail = (qs, case_kc, case_vc, KH, scale, bt, lens, BS, ctx, None, c["kv"], k_scale, v_scale)
    if c["version"] == 1:
        oracle.paged_attention_v1(want, *tail)
    else:
        H, D = qs.shape[1], qs.shape[2]
        P = (ctx + 511) // 512
        es = torch.empty(len(rows), H, P, dtype=torch.float32)
        oracle.paged_attention_v2(want, es, torch.empty_like(es),
                                  torch.empty(len(rows), H, P, D, dtype=qs.dtype), *tail)
    err = (out[rows].float().cpu() - want.float()).abs().max().item()
    tol = 2e-3 if c["dtype"] == "float16" else 1e-2
    if not err <= tol:
        raise RuntimeError(f"bench step output differs from the oracle: max-abs {err} > {tol}")
    return {"max_abs": err, "tol": tol, "sequences": rows, "against": "oracle/paged_oracle.c (CPU)"}


def reference_cuda_leg(c, timed, run_attention_with, L, reps, ms_attn):
    """Times torch.ops._ref_C.paged_attention_v1/v2 -- the reference's attention_kernels.cu built for
    sm_100a by oracle/Makefile -- with the product's own CUDA-event loop, pools and block tables."""
    from oracle import ref
    if not ref.have_ref_cuda():
        return {"unavailable": "oracle/_ref/libref_cuda.so not built"}
    rC, _ = ref.ref_cuda()
    for l in range(L):
        run_attention_with(rC, l)
    counter = [0]

    def one():
        run_attention_with(rC, counter[0] % L)
        counter[0] += 1

    ms = timed(one, reps) / reps
    return {"us_per_launch": round(ms * 1e3, 2),
            "GBps": round(algorithmic_bytes_attention(c) / (ms / 1e3) / 1e9, 1),
            "speedup": round(ms / ms_attn, 3),
            "kernel": f"reference paged_attention_v{c['version']} (csrc/attention/attention_kernels.cu "
                      f"recompiled for sm_100a), {reps} launches, same pools / block tables / events"}


def ctx_shard_record(dev, rank, world, dist, steps):
    """--gpus N >= 2, after the timed region: ONE sequence batch's context split over the N ranks
    (real one-process-per-GPU ranks over CUDA IPC, NVLink peer stores inside the merge kernel; no
    NCCL on the data path) vs the same decode step on one GPU (paged_attention_v2).  Strong scaling;
    device-timed, max over ranks.  4 x 128 Ki tokens and 1 x 128 Ki tokens, Llama-3-8B-shaped bf16."""
    import torch
    from light_vllm_b200 import _custom_ops as ops
    from light_vllm_b200.context_parallel import ContextShardGroup, ContextSharding
    H, KH, D, BS, ctx, layers = 32, 8, 128, 16, 131072, 2
    dtype = torch.bfloat16
    scale = D ** -0.5
    rec = {}
    for B in (4, 1):
        nblk = ctx // BS
        NB = B * nblk + 8
        g = torch.Generator(device=dev).manual_seed(1234)   # same pool on every rank
        pools = []
        for _ in range(layers):
            kc = torch.empty(NB, KH, D // 8, BS, 8, device=dev, dtype=dtype).uniform_(-0.5, 0.5, generator=g)
            vc = torch.empty(NB, KH, D, BS, device=dev, dtype=dtype).uniform_(-0.5, 0.5, generator=g)
            pools.append((kc, vc))
        q = torch.empty(B, H, D, device=dev, dtype=dtype).uniform_(-0.25, 0.25, generator=g)
        perm = torch.randperm(NB, generator=torch.Generator().manual_seed(7))[:B * nblk].int().view(B, nblk)
        lens = [ctx] * B
        bt_full = perm.to(dev)
        sl_full = torch.tensor(lens, dtype=torch.int32, device=dev)
        P = ctx // 512
        es = torch.empty(B, H, P, dtype=torch.float32, device=dev)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, H, P, D, dtype=dtype, device=dev)
        out_ref, out = torch.empty_like(q), torch.empty_like(q)

        def base_step(i):
            kc, vc = pools[i % layers]
            ops.paged_attention_v2(out_ref, es, ml, tmp, q, kc, vc, KH, scale, bt_full, sl_full, BS,
                                   ctx, None, "auto", 1.0, 1.0)

        sh = ContextSharding(world, BS, 4)
        grp = ContextShardGroup.create(B * H, D, dev.index)
        bt, ll, mx = sh.shard_batch(perm, lens, rank)
        bt, ll = bt.to(dev), ll.to(dev)

        def shard_step(i):
            kc, vc = pools[i % layers]
            grp.paged_attention(out, q, kc, vc, KH, scale, bt, ll, BS, mx)

        def timed(fn):
            for i in range(5):
                fn(i)
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()

        def as_graph(fn, n=8):
            """n consecutive calls (alternating pools, like n layers of one decode step) as ONE CUDA graph:
            how a serving loop issues decode attention, and the only way N one-per-GPU host threads do not
            add their launch jitter to every cross-rank wait."""
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                fn(0)
                side.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=side):
                    for i in range(n):
                        fn(i)
            torch.cuda.current_stream(dev).wait_stream(side)
            return gr, n

        base_ms = timed(base_step)
        shard_ms = timed(shard_step)
        g_base, n_g = as_graph(base_step)
        g_shard, _ = as_graph(shard_step)
        base_graph_ms = timed(lambda i: g_base.replay()) / n_g
        shard_graph_ms = timed(lambda i: g_shard.replay()) / n_g
        del g_base, g_shard
        base_step(0)
        shard_step(0)
        torch.cuda.synchronize()
        st = torch.tensor([float(grp.status()), (out.float() - out_ref.float()).abs().max().item()],
                          device=dev)
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
        kv_bytes = B * ctx * KH * D * 2 * 2
        rec[f"{B}x{ctx}"] = {
            "ms": round(shard_ms, 4), "single_gpu_ms": round(base_ms, 4),
            "speedup": round(base_ms / shard_ms, 3),
            "graph_ms": round(shard_graph_ms, 4), "single_gpu_graph_ms": round(base_graph_ms, 4),
            "graph_speedup": round(base_graph_ms / shard_graph_ms, 3), "max_abs": st[1].item(),
            "exchange_status": int(st[0].item()),
            "aggregate_kv_GBps": round(kv_bytes / shard_ms / 1e6, 1)}
        if not st[1].item() <= 1e-2 or st[0].item() != 0:
            raise RuntimeError(f"context-sharded decode disagrees with the single-GPU result: {rec}")
        grp.close()
        del pools, es, ml, tmp
        torch.cuda.empty_cache()
    rec["note"] = (f"one decode step, bf16 H{H}/KH{KH}/D{D}/BS{BS}, pages dealt round-robin (4-page "
                   f"chunks) to {world} one-process-per-GPU ranks; CUDA-IPC exchange buffers, peer "
                   "stores + epoch flags inside the merge kernel; strong scaling vs the whole context "
                   "on ONE GPU (paged_attention_v2); device-timed, max over ranks; max_abs vs that result; "
                   "graph_*: the same calls issued 8 per CUDA graph (both arms)")
    return rec

# --------------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------------
def run_b200_arm(args, c):
    import torch
    import torch.distributed as dist

    from light_vllm_b200 import _custom_ops as ops
    from light_vllm_b200 import _lib
    from light_vllm_b200.attention import (PagedAttention, PagedB200AttentionBackend,
                                           PagedB200AttentionMetadata)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback exists)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    dtype = getattr(torch, c["dtype"])
    kv = c["kv"]
    cache_dtype = dtype if kv == "auto" else torch.uint8
    H, KH, D, BS, B, ctx, L = c["H"], c["KH"], c["D"], c["BS"], c["B"], c["ctx"], args.layers
    scale = D ** -0.5
    k_scale = v_scale = 1.0 if kv == "auto" else 0.05
    blocks_per_seq = (ctx + BS - 1) // BS
    preempt = bool(c.get("preempt")) and not args.no_preempt
    # 64 spare blocks; plus one whole sequence's worth past them for the preemption exercise (CoW targets)
    n_live = B * blocks_per_seq + 64          # block ids the block tables draw from
    NB = n_live + (blocks_per_seq if preempt else 0)
    torch.manual_seed(1234 + rank)

    # ---- this rank's shard: its own pools, block tables and sequences (sequence sharding) ----
    shape = PagedB200AttentionBackend.get_kv_cache_shape(NB, BS, KH, D)
    pools = []
    for _ in range(L):
        if cache_dtype == torch.uint8:
            src = torch.empty(shape, dtype=dtype, device=dev).uniform_(-scale, scale)
            pool = torch.empty(shape, dtype=torch.uint8, device=dev)
            ops.convert_fp8(pool, src, k_scale, "fp8")
            del src
        else:
            pool = torch.empty(shape, dtype=dtype, device=dev).uniform_(-scale, scale)
        pools.append(pool)
    split = [PagedAttention.split_kv_cache(p, KH, D) for p in pools]
    # Host side of the e2e legs: ONE pinned staging buffer per step -- q/k/v, slot mapping, sequence lengths and
    # block tables are views into it, so a step's inputs cross PCIe as a single H2D copy.
    def _al(n):
        return (n + 255) // 256 * 256
    esz = torch.empty((), dtype=dtype).element_size()
    n_qkv, n_slots, n_lens, n_bt = B * (H + 2 * KH) * D * esz, B * 8, B * 4, B * blocks_per_seq * 4
    o_slots = _al(n_qkv)
    o_lens = o_slots + _al(n_slots)
    o_bt = o_lens + _al(n_lens)
    stage_h = torch.zeros(o_bt + _al(n_bt), dtype=torch.uint8).pin_memory()

    def stage_views(buf):
        return (buf[:n_qkv].view(dtype).view(B, (H + 2 * KH) * D), buf[o_slots:o_slots + n_slots].view(torch.int64),
                buf[o_lens:o_lens + n_lens].view(torch.int32),
                buf[o_bt:o_bt + n_bt].view(torch.int32).view(B, blocks_per_seq))

    qkv_h, slots_h, seq_lens_h, block_tables_h = stage_views(stage_h)
    block_tables_h.copy_(torch.randperm(n_live)[:B * blocks_per_seq].view(B, blocks_per_seq).int())
    seq_lens_h.fill_(ctx)
    slots_h.copy_(block_tables_h[:, (ctx - 1) // BS].long() * BS + (ctx - 1) % BS)
    qkv_h.uniform_(-scale, scale)
    out_h = torch.empty(B, H * D, dtype=dtype).pin_memory()

    block_tables, seq_lens, slots = block_tables_h.to(dev), seq_lens_h.to(dev), slots_h.to(dev)
    qkv = qkv_h.to(dev)
    q = qkv[:, :H * D].view(B, H, D)
    k = qkv[:, H * D:(H + KH) * D].view(B, KH, D)
    v = qkv[:, (H + KH) * D:].view(B, KH, D)
    out = torch.empty(B, H, D, dtype=dtype, device=dev)
    P = (ctx + 511) // 512
    scratch = (torch.empty(B, H, P, dtype=torch.float32, device=dev),
               torch.empty(B, H, P, dtype=torch.float32, device=dev),
               torch.empty(B, H, P, D, dtype=dtype, device=dev))

    def attention_with(mod, l):
        kc, vc = split[l]
        if c["version"] == 1:
            mod.paged_attention_v1(out, q, kc, vc, KH, scale, block_tables, seq_lens, BS, ctx, None,
                                   kv, k_scale, v_scale, 0, 0, 0, 64, 0)
        else:
            mod.paged_attention_v2(out, *scratch, q, kc, vc, KH, scale, block_tables, seq_lens, BS,
                                   ctx, None, kv, k_scale, v_scale, 0, 0, 0, 64, 0)

    def attention(l):
        attention_with(ops, l)

    def device_step():
        for l in range(L):
            kc, vc = split[l]
            ops.reshape_and_cache(k, v, kc, vc, slots, kv, k_scale, v_scale)
            attention(l)

    # ---- the same step through the plugin, from HOST buffers ----
    impl = PagedB200AttentionBackend.get_impl_cls()(H, D, scale, KH, None, None, kv)

    def e2e_forward(x):
        """The plugin call of one step on device views of the staged inputs: L x PagedB200AttentionImpl.forward."""
        xqkv, xslots, xlens, xbt = stage_views(x)
        meta = PagedB200AttentionMetadata(
            num_prefills=0, num_prefill_tokens=0, num_decode_tokens=B, slot_mapping=xslots,
            seq_lens=None, seq_lens_tensor=xlens, max_query_len=1, max_prefill_seq_len=0,
            max_decode_seq_len=ctx, query_start_loc=None, seq_start_loc=None,
            context_lens_tensor=None, block_tables=xbt, use_cuda_graph=False)
        xq, xk, xv = xqkv[:, :H * D], xqkv[:, H * D:(H + KH) * D], xqkv[:, (H + KH) * D:]
        o = None
        for l in range(L):
            o = impl.forward(xq, xk, xv, pools[l], meta, k_scale, v_scale)
        return o

    def e2e_body():
        out_h.copy_(e2e_forward(stage_h.to(dev, non_blocking=True)), non_blocking=True)

    def e2e_step():  # eager and serial: one H2D, L forwards, the D2H read, then the host waits for the result
        e2e_body()
        torch.cuda.current_stream().synchronize()

    h2d = stage_h.numel()  # the whole staging buffer (payload + < 1 KiB of alignment padding)
    d2h = out_h.numel() * out_h.element_size()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n, finish=None):
        """n calls of fn between a barrier+sync on both sides; device time by CUDA events on the
        launching (current) stream; max over ranks.  ``finish`` (pipelined legs) joins the side
        streams into the launching stream before the closing event."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        if finish is not None:
            finish()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    for _ in range(max(args.warmup, 3)):
        device_step()
    sampler = ClockSampler(local_rank)
    n0 = _lib.launch_count()
    sampler.start()
    ms_total = timed(device_step, args.steps)
    clocks = sampler.stop()
    launches = _lib.launch_count() - n0
    # the same device step as ONE CUDA graph (no launch gaps between the 2 L kernels): what the e2e legs replay
    gside = torch.cuda.Stream(device=dev)
    gside.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(gside):
        device_step()
    torch.cuda.current_stream().wait_stream(gside)
    torch.cuda.synchronize()
    dgraph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(dgraph):
        device_step()
    for _ in range(3):
        dgraph.replay()
    ms_total_graph = timed(dgraph.replay, args.steps)
    del dgraph

    # attention kernel alone, for the roofline (cycling pools: every launch reads cold KV)
    reps = max(20, args.steps)
    for l in range(L):
        attention(l)
    counter = [0]

    def attn_only():
        attention(counter[0] % L)
        counter[0] += 1

    ms_attn = timed(attn_only, reps) / reps

    # ---- checker legs (untimed): oracle on a slice of the step's output, reference CUDA kernels ----
    device_step()
    torch.cuda.synchronize()
    parity = None
    if rank == 0:
        kc_l, vc_l = split[L - 1]
        parity = oracle_slice_check(c, q, out, kc_l, vc_l, block_tables, seq_lens, scale, k_scale,
                                    v_scale)
    ref_cuda = None
    if not args.no_reference_cuda:
        ref_cuda = reference_cuda_leg(c, timed, attention_with, L, max(5, min(reps, 10)), ms_attn)
    for _ in range(max(args.warmup, 3)):
        e2e_step()
    ms_e2e_eager = timed(e2e_step, args.steps)
    out_eager = out_h.clone()

    # The same step (one pinned-host H2D copy, L plugin forwards, D2H read of the result) captured
    # once in a CUDA graph and replayed per step, the way a decode engine runs a fixed-shape decode
    # step: the host refreshes the pinned buffer, launches ONE graph and waits for the result.
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        e2e_body()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        e2e_body()

    def e2e_graph_step():
        graph.replay()
        torch.cuda.current_stream().synchronize()

    out_h.zero_()
    for _ in range(max(args.warmup, 3)):
        e2e_graph_step()
    if not torch.equal(out_h, out_eager):
        raise RuntimeError("graph-replayed e2e step disagrees with the eager e2e step")
    ms_e2e_serial = timed(e2e_graph_step, args.steps)

    # Two steps in flight, as the reference's executor keeps them (core/executor.py:62-185,
    # max_num_on_the_fly = 2): double-buffered staging / output buffers; the H2D copy of step t+1 and the
    # D2H read of step t-1 run on copy streams beside the kernels of step t; the host blocks on the
    # result of step t-1 after it has enqueued step t.  Every step still copies its inputs in and its
    # result out inside the timed region.
    main = torch.cuda.current_stream()
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    x_d = [torch.empty_like(stage_h, device=dev) for _ in range(2)]
    o_d = [torch.empty(B, H * D, dtype=dtype, device=dev) for _ in range(2)]
    o_h = [torch.empty(B, H * D, dtype=dtype).pin_memory() for _ in range(2)]
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_k = [torch.cuda.Event() for _ in range(2)]
    ev_out = [torch.cuda.Event() for _ in range(2)]
    for b in range(2):
        x_d[b].copy_(stage_h, non_blocking=True)
    torch.cuda.synchronize()
    graphs = []
    for b in range(2):
        side.wait_stream(main)
        with torch.cuda.stream(side):
            o_d[b].copy_(e2e_forward(x_d[b]))
        main.wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            o_d[b].copy_(e2e_forward(x_d[b]))
        graphs.append(g)
    state = {"t": 0, "primed": False}

    def stage_in(b):  # H2D of the next step's inputs, behind the kernels that last read this buffer
        s_in.wait_event(ev_k[b])
        with torch.cuda.stream(s_in):
            x_d[b].copy_(stage_h, non_blocking=True)
            ev_in[b].record(s_in)

    def make_pipelined(run_kernels):
        def step():
            t = state["t"]
            b = t & 1
            if not state["primed"]:
                ev_k[0].record(main)
                ev_k[1].record(main)
                stage_in(b)
                state["primed"] = True
            main.wait_event(ev_in[b])
            main.wait_event(ev_out[b])       # the D2H that last read o_d[b] is done
            run_kernels(b)
            ev_k[b].record(main)
            stage_in(b ^ 1)                  # prefetch step t+1's inputs
            s_out.wait_event(ev_k[b])
            with torch.cuda.stream(s_out):
                o_h[b].copy_(o_d[b], non_blocking=True)
                ev_out[b].record(s_out)
            if t > 0:
                ev_out[b ^ 1].synchronize()  # the host reads the result of step t-1
            state["t"] = t + 1
        return step

    def pipeline_finish():
        main.wait_stream(s_in)
        main.wait_stream(s_out)
        state["primed"] = False

    for b in range(2):
        ev_out[b].record(main)
    pipe_graph = make_pipelined(lambda b: graphs[b].replay())
    pipe_eager = make_pipelined(lambda b: o_d[b].copy_(e2e_forward(x_d[b])))
    for fn in (pipe_graph, pipe_eager):
        for _ in range(max(args.warmup, 3)):
            fn()
        pipeline_finish()
        torch.cuda.synchronize()
        if not (torch.equal(o_h[0], out_eager) and torch.equal(o_h[1], out_eager)):
            raise RuntimeError("pipelined e2e step disagrees with the eager e2e step")
    ms_e2e = timed(pipe_graph, args.steps, pipeline_finish)
    ms_e2e_eager_pipe = timed(pipe_eager, args.steps, pipeline_finish)
    del graphs

    pre = None
    if preempt:
        pre = preemption_exercise(args, c, dev, pools, block_tables_h, NB, device_step, timed)

    extra = {}
    if world == 1 and rank == 0 and not args.no_extra_records and args.config == "c2":
        del graph
        pools.clear()
        split.clear()
        torch.cuda.empty_cache()
        for name, fn in (("prefix_prefill", lambda: prefix_prefill_record(dev, args.steps)),
                         ("c3", lambda: other_config_record("c3", dev, args.steps)),
                         ("c4", lambda: other_config_record("c4", dev, args.steps)),
                         # c5's attention launch at full shape (16 x 128 Ki tokens, 16 GiB of K/V in one pool); its
                         # preemption leg (copy_blocks / swap_blocks) is `--config c5`
                         ("c5", lambda: other_config_record("c5", dev, max(6, args.steps // 4)))):
            try:
                extra[name] = fn()
            except Exception as e:  # reported beside the headline, never allowed to take it down
                extra[name] = {"error": repr(e)}
            torch.cuda.empty_cache()

    ctx_shard = None
    if world > 1 and not args.no_ctx_shard:
        del graph
        pools.clear()
        split.clear()
        torch.cuda.empty_cache()
        ctx_shard = ctx_shard_record(dev, rank, world, dist, max(10, args.steps))

    return dict(rank=rank, world=world, ms_total=ms_total, ms_e2e=ms_e2e, ms_e2e_eager=ms_e2e_eager,
                ms_e2e_serial=ms_e2e_serial, ms_e2e_eager_pipe=ms_e2e_eager_pipe, ms_total_graph=ms_total_graph,
                ms_attn=ms_attn, parity=parity, reference_cuda=ref_cuda, ctx_shard=ctx_shard, extra=extra,
                clocks=clocks, launches=launches, h2d=h2d, d2h=d2h, L=L, preemption=pre,
                plan=_lib.attention_plan(c["version"], B, H, KH, D, BS, ctx, dtype, kv, local_rank),
                dist=dist if world > 1 else None)


def preemption_exercise(args, c, dev, pools, block_tables_h, NB, device_step, timed):
    """configs[4]: what the scheduler's preemption does to one victim sequence.

    swap_blocks all blocks of sequence 0 out to a pinned host pool (K and V of every layer pool),
    scrub them on the GPU (the allocator would hand them to someone else), copy_blocks every
    block of sequence 1 to spare blocks (a fork's copy-on-write, at scale), swap the victim back
    in; the pools must be bit-exact afterwards.  Then the three ops are timed on their own, and
    the decode step is timed with one such cycle every ``--preempt-every`` steps on the same
    stream (the reference issues swap/copy before the model step, gpu_worker.py:206-221)."""
    import torch
    from light_vllm_b200.attention import PagedB200AttentionBackend as Backend
    BS, KH, D, L = c["BS"], c["KH"], c["D"], len(pools)
    bps = block_tables_h.shape[1]
    esz = pools[0].element_size()
    block_bytes = BS * KH * D * esz                       # one block of one layer's K (== V)
    cpu_pools = [torch.zeros((2, bps, BS * KH * D), dtype=pools[0].dtype).pin_memory()
                 for _ in range(L)]
    victim = block_tables_h[0].long()
    idx = torch.arange(bps, dtype=torch.int64)
    out_map = torch.stack([victim, idx], dim=1).contiguous()          # gpu block -> cpu block (CPU tensor)
    in_map = torch.stack([idx, victim], dim=1).contiguous()
    spare = torch.arange(NB - bps, NB, dtype=torch.int64)
    cow_map = torch.stack([block_tables_h[1].long(), spare], dim=1).contiguous().pin_memory()
    vdev = victim.to(dev)

    # ---- bit-exact round trip (untimed) ----
    before = [p[:, vdev].clone() for p in pools[:1]] + [None] * (L - 1)   # 1 GiB: layer 0 in full
    sums = [torch.sum(p[:, vdev].view(torch.int32), dtype=torch.int64).item() for p in pools]
    for l in range(L):
        Backend.swap_blocks(pools[l], cpu_pools[l], out_map)
    for p in pools:
        p[:, vdev] = 0
    Backend.copy_blocks(pools, cow_map)
    for l in range(L):
        Backend.swap_blocks(cpu_pools[l], pools[l], in_map)
    torch.cuda.synchronize()
    ok = torch.equal(pools[0][:, vdev], before[0])
    ok = ok and all(torch.sum(p[:, vdev].view(torch.int32), dtype=torch.int64).item() == s
                    for p, s in zip(pools, sums))
    src1 = block_tables_h[1].long().to(dev)
    ok = ok and all(torch.equal(p[:, src1], p[:, spare.to(dev)]) for p in pools)
    del before
    assert ok, "preemption round trip is not bit-exact"

    # ---- the three ops on their own ----
    def swap_out():
        for l in range(L):
            Backend.swap_blocks(pools[l], cpu_pools[l], out_map)

    def swap_in():
        for l in range(L):
            Backend.swap_blocks(cpu_pools[l], pools[l], in_map)

    def cow():
        Backend.copy_blocks(pools, cow_map)

    reps = 3
    for f in (swap_out, swap_in, cow):
        f()
    ms_out = timed(swap_out, reps) / reps
    ms_in = timed(swap_in, reps) / reps
    ms_cow = timed(cow, reps) / reps
    moved = L * 2 * bps * block_bytes                      # bytes one direction, K and V
    k = max(1, args.preempt_every)
    state = [0]

    def step_under_preemption():
        if state[0] % k == 0:
            swap_out()
            cow()
            swap_in()
        state[0] += 1
        device_step()

    step_under_preemption()
    state[0] = 0
    ms_mixed = timed(step_under_preemption, args.steps) / args.steps
    return dict(bit_exact=bool(ok), blocks=int(bps), layer_pools=L, bytes_per_direction=int(moved),
                swap_out_GBps=round(moved / ms_out / 1e6, 2), swap_in_GBps=round(moved / ms_in / 1e6, 2),
                swap_out_ms=round(ms_out, 3), swap_in_ms=round(ms_in, 3),
                copy_blocks_GBps=round(2 * moved / ms_cow / 1e6, 1), copy_blocks_ms=round(ms_cow, 3),
                copy_pairs=int(bps), preempt_every=k,
                ms_per_step_under_preemption=round(ms_mixed, 4),
                decode_tokens_per_s_per_layer_under_preemption=round(c["B"] * L / (ms_mixed / 1e3), 1),
                note="swap = whole victim sequence, all layer pools, K+V, pinned host pool; copy = "
                     "read+write bytes of a whole-sequence copy-on-write; 'under preemption' = one "
                     "swap-out + CoW + swap-in cycle every preempt_every decode steps, same stream")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--layers", type=int, default=4,
                    help="KV pools cycled per step (each > L2; a model's layers)")
    ap.add_argument("--cpu-budget", type=float, default=12.0,
                    help="seconds of CPU work for the cpu_baseline leg (0 disables)")
    ap.add_argument("--ctx", type=int, default=0, help="override the context length (c5: 32768..131072)")
    ap.add_argument("--batch", type=int, default=0, help="override sequences per GPU")
    ap.add_argument("--preempt-every", type=int, default=8,
                    help="c5: decode steps between preemption cycles in the mixed measurement")
    ap.add_argument("--no-preempt", action="store_true", help="c5: skip the preemption exercise")
    ap.add_argument("--no-reference-cuda", action="store_true",
                    help="skip the (untimed-region) leg that times the reference's CUDA kernels")
    ap.add_argument("--no-extra-records", action="store_true",
                    help="N = 1, c2: skip the prefix_prefill / c3 / c4 records printed beside the headline")
    ap.add_argument("--no-ctx-shard", action="store_true",
                    help="--gpus N >= 2: skip the context-sharded decode record")
    args = ap.parse_args()
    c = dict(CONFIGS[args.config])
    if args.ctx:
        c["ctx"] = args.ctx
    if args.batch:
        c["B"] = args.batch
    if args.ctx or args.batch:
        c["label"] += f" [ctx {c['ctx']}, batch {c['B']} per GPU]"
    if args.impl == "reference":
        run_reference_arm(args, c)
        return

    r = run_b200_arm(args, c)
    world = r["world"]
    cpu = None
    if r["rank"] == 0 and world == 1 and args.cpu_budget > 0:
        try:
            cpu = time_cpu(c, budget_s=args.cpu_budget, warmup=3)
        except Exception as e:  # the baseline is reported, never required for the GPU numbers
            cpu = {"error": repr(e)}
    if r["rank"] == 0:
        L = r["L"]
        kvb = kv_bytes_per_layer(c)
        step_s = r["ms_total"] / 1e3 / args.steps
        e2e_s = r["ms_e2e"] / 1e3 / args.steps
        value = world * L * kvb / step_s / 1e9
        e2e = world * L * kvb / e2e_s / 1e9
        peak, peak_src = measured_peak_gbs()
        achieved = algorithmic_bytes_attention(c) / (r["ms_attn"] / 1e3) / 1e9
        line = {
            "metric": "paged_attn_kv_read_GBps", "value": round(value, 1), "unit": "GB/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(step_s * 1e3, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16" if c["dtype"] == "float16" else "bf16",
            "data": "synthetic",
            "config": {"workload": c["label"], "name": args.config, "kv_cache_dtype": c["kv"],
                       "layer_pools_per_step": L, "kv_bytes_per_pool": kvb,
                       "l2": f"inputs larger than L2: {L} pools x {kvb >> 20} MiB cycled",
                       "sharding": "by sequence; each rank owns its pools and block tables",
                       "plan": r["plan"]},
            "frac_of_hbm_peak": round(value / world / peak, 4),
            "device_step_as_cuda_graph": {
                "ms_per_step": round(r["ms_total_graph"] / args.steps, 4),
                "value": round(world * L * kvb / (r["ms_total_graph"] / 1e3 / args.steps) / 1e9, 1),
                "note": "`value` enqueues the step's 2 L kernels from Python; replayed as one graph (as the e2e "
                        "legs do) the launch gaps disappear, which is why e2e can exceed `value`"},
            "decode_tokens_per_s_per_layer": round(world * c["B"] * L / step_s, 1),
            "e2e": {"value": round(e2e, 1), "unit": "GB/s", "h2d_bytes_per_step": r["h2d"],
                    "d2h_bytes_per_step": r["d2h"], "ms_per_step": round(e2e_s * 1e3, 4),
                    "api": "PagedB200AttentionImpl.forward per layer pool; per step ONE pinned-host staging "
                           "buffer (q/k/v + slot mapping + seq_lens + block tables) is copied in and the "
                           "output is read back; the forwards are one CUDA graph; two steps in flight like "
                           "the reference executor (double-buffered staging, H2D of step t+1 and D2H of "
                           "step t-1 on copy streams beside step t, host waits for step t-1's result)",
                    "serial_ms_per_step": round(r["ms_e2e_serial"] / args.steps, 4),
                    "serial_value": round(world * L * kvb / (r["ms_e2e_serial"] / 1e3 / args.steps) / 1e9, 1),
                    "serial_note": "same graph-replayed step, one at a time: H2D -> forwards -> D2H -> host sync",
                    "eager_ms_per_step": round(r["ms_e2e_eager_pipe"] / args.steps, 4),
                    "eager_value": round(world * L * kvb / (r["ms_e2e_eager_pipe"] / 1e3 / args.steps) / 1e9, 1),
                    "eager_note": "two steps in flight, every copy and launch enqueued from Python (no graph)",
                    "eager_serial_ms_per_step": round(r["ms_e2e_eager"] / args.steps, 4)},
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak,
                         "unit": "GB/s", "frac": round(achieved / peak, 4),
                         "traffic": None if (args.ctx or args.batch) else ncu_traffic(args.config),
                         "kernel": f"paged_attention_v{c['version']} (attention launch alone, "
                                   f"{algorithmic_bytes_attention(c)} algorithmic bytes)",
                         "us_per_launch": round(r["ms_attn"] * 1e3, 2), "peak_source": peak_src,
                         "frac_of_8TBps_datasheet": round(achieved / 8000.0, 4)},
            "clocks": r["clocks"], "gpu_launches": r["launches"],
        }
        if r.get("preemption") is not None:
            line["preemption"] = r["preemption"]
        if r.get("parity") is not None:
            line["parity"] = r["parity"]
        if r.get("reference_cuda") is not None:
            line["reference_cuda"] = r["reference_cuda"]
        if r.get("ctx_shard") is not None:
            line["ctx_shard"] = r["ctx_shard"]
        if r.get("extra"):
            if "prefix_prefill" in r["extra"]:
                line["prefix_prefill"] = r["extra"]["prefix_prefill"]
            line["other_configs"] = {k: v for k, v in r["extra"].items() if k != "prefix_prefill"}
        if cpu is not None and "error" not in cpu:
            # same statistic as the --impl reference arm: MEAN over the timed steps (>= 3 untimed)
            gbs = cpu["kv_bytes"] / cpu["mean_s"] / 1e9
            line["cpu_baseline"] = {
                "value": round(gbs, 3), "unit": "GB/s", "cores": cpu["cores"], "kind": cpu["kind"],
                "sample": f"{cpu['sample_B']} of {c['B']} sequences x ctx {c['ctx']}, 1 layer pool, "
                          f"{cpu['dtype']} (CPU path has no fp16/fp8), reshape_and_cache + "
                          f"paged_attention_v{c['version']}, mean of {cpu['n']} steps after 3 untimed "
                          f"({cpu['mean_s'] * 1e3:.1f} ms/step; best {cpu['best_s'] * 1e3:.1f} ms)"}
        elif cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if r["dist"] is not None:
        r["dist"].barrier()
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
