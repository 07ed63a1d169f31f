"""Engine integration: the reference's OWN decoding scheduler + block managers drive
PagedB200AttentionBackend + light_vllm_b200.cache_engine.CacheEngine through real preemption.

The streams under tests/golden/sched_stream_*.json were recorded on the CPU box from the unmodified
light_vllm/decoding/scheduler.py (:234-1075) and core/block_manager_v1/v2.py with
VLLM_TEST_ENABLE_ARTIFICIAL_PREEMPT=1 (scheduler.py:22-26) by tests/golden/make_golden_sched.py -- the
reference cannot be imported on the GPU box.  A stream is, per step: the sequences that run (prefill or
decode, block tables, lengths) and blocks_to_swap_in / blocks_to_swap_out / blocks_to_copy.  The replay
issues the swaps and copies in the order of Worker.non_blocking_h2d (decoding/worker/gpu_worker.py:206-221),
builds the step's metadata with the plugin's metadata builder, runs a Qwen2-shaped model greedily, and
must produce exactly the tokens of a run that was never preempted (recompute mode re-prefills prompt +
generated tokens; swap mode moves the blocks to the pinned CPU pool and back).

Also here: two decode steps in flight on two pooled non-default streams (core/executor.py:62-185).
"""
import json
import os
import types

import pytest
import torch
import torch.nn.functional as F

from qwen2_harness import Qwen2Shape, _layers_forward, init_weights, rmsnorm

HERE = os.path.dirname(os.path.abspath(__file__))
STREAMS = ["recompute_v1", "swap_v1", "swap_v2"]


def load(name):
    with open(os.path.join(HERE, "golden", f"sched_stream_{name}.json")) as f:
        return json.load(f)


# ------------------------------------------------------------------ CPU: the streams themselves ----
@pytest.mark.parametrize("name", STREAMS)
def test_stream_is_a_real_preemption_workload(name):
    """Sanity of the recorded data: every request finishes with max_tokens outputs, preemptions happened,
    swap streams move blocks both ways, block ids stay inside the pools, a decode step never shares a
    physical block between two sequences, swap-ins return exactly what was swapped out."""
    d = load(name)
    assert d["counts"]["cumulative_preemptions"] > 20
    if d["preemption_mode"] == "swap":
        assert d["counts"]["swapped_block_pairs"] > 200
    else:
        assert d["counts"]["recompute_prefills"] > 20 and d["counts"]["swapped_block_pairs"] == 0
    for rid, seqs in d["final"].items():
        for s in seqs.values():
            assert s["output_len"] == d["max_tokens"]
    on_cpu = {}
    for st in d["steps"]:
        for c, g in st["swap_in"]:
            assert 0 <= c < d["num_cpu_blocks"] and 0 <= g < d["num_gpu_blocks"]
            assert on_cpu.pop(c, None) is not None, "swap-in of a block that was never swapped out"
        for g, c in st["swap_out"]:
            assert 0 <= c < d["num_cpu_blocks"] and 0 <= g < d["num_gpu_blocks"]
            assert c not in on_cpu
            on_cpu[c] = g
        used = set()
        for grp in st["groups"]:
            for s in grp["seqs"]:
                n_needed = (s["len"] + d["block_size"] - 1) // d["block_size"]
                assert len(s["block_table"]) >= n_needed
                assert all(0 <= b < d["num_gpu_blocks"] for b in s["block_table"])
                assert not (used & set(s["block_table"]))
                used |= set(s["block_table"])
    assert not on_cpu


# ------------------------------------------------------------------ GPU: replay ----
class Replay:
    def __init__(self, doc, shape, weights, dtype, device="cuda"):
        from light_vllm_b200.attention import PagedB200AttentionBackend as Backend
        from light_vllm_b200.cache_engine import CacheEngine
        self.d, self.shape, self.w, self.dev = doc, shape, weights, device
        self.bs = doc["block_size"]
        self.cache = CacheEngine(shape.layers, shape.kv_heads, shape.head_dim, self.bs, doc["num_gpu_blocks"],
                                 doc["num_cpu_blocks"], dtype, "auto", device, Backend)
        self.impl = Backend.get_impl_cls()(shape.heads, shape.head_dim, shape.head_dim ** -0.5, shape.kv_heads,
                                           None, None, "auto")
        self.builder_cls = Backend.get_builder_cls()
        self.tokens = {}  # seq_id -> prompt + generated token ids
        self.prompt_len = {}

    def _input_builder(self, groups):
        return types.SimpleNamespace(
            sliding_window=None, block_size=self.bs,
            scheduler_config=types.SimpleNamespace(use_v2_block_manager=self.d["block_manager"] == "v2"),
            chunked_prefill_enabled=False, inter_data_list=groups)

    @torch.inference_mode()
    def step(self, st):
        t = lambda pairs: torch.tensor(pairs, dtype=torch.int64).view(-1, 2)  # noqa: E731
        # Worker.non_blocking_h2d order: swap in, swap out, copy; all on the current stream
        self.cache.apply(t(st["swap_in"]), t(st["swap_out"]), t(st["copy"]).to(self.dev))
        if not st["groups"]:
            return
        inter, toks, pos, seq_lens, query_lens, sample_rows, row = [], [], [], [], [], [], 0
        for g in st["groups"]:
            for s in g["seqs"]:
                sid = s["seq_id"]
                if sid not in self.tokens:
                    rid = int(g["request_id"][1:])
                    self.tokens[sid] = list(self.d["prompts"][rid])
                    self.prompt_len[sid] = len(self.tokens[sid])
                have = self.tokens[sid]
                assert len(have) == s["len"], (g["request_id"], len(have), s["len"])
                if g["is_prompt"]:
                    ctx, q = s["num_computed"], g["token_chunk_size"]
                    assert ctx == 0 and q == s["len"]  # the default policy never chunks
                else:
                    ctx, q = s["len"] - 1, 1
                ids = have[ctx:ctx + q]
                inter.append(types.SimpleNamespace(
                    is_prompt=g["is_prompt"], block_tables={sid: s["block_table"]}, seq_ids=[sid],
                    input_tokens=[ids], orig_seq_lens=[s["len"]], seq_lens=[ctx + q], query_lens=[q],
                    context_lens=[ctx], curr_sliding_window_blocks=[len(s["block_table"])],
                    prefix_cache_hit=False))
                toks += ids
                pos += list(range(ctx, ctx + q))
                seq_lens.append(ctx + q)
                query_lens.append(q)
                row += q
                sample_rows.append((sid, row - 1, g["do_sample"]))
        meta = self.builder_cls(self._input_builder(inter)).build(seq_lens, query_lens, -1, len(seq_lens))
        meta.to(self.dev, non_blocking=True)
        x = self.w["emb"][torch.tensor(toks, device=self.dev)]
        positions = torch.tensor(pos, dtype=torch.long, device=self.dev)

        def attn(l, q, k, v):
            T = q.shape[0]
            return self.impl.forward(q.reshape(T, -1), k.reshape(T, -1), v.reshape(T, -1), self.cache.gpu_cache[l],
                                     meta).view_as(q)

        x = _layers_forward(self.shape, self.w, None, x, positions, attn, None)
        rows = torch.tensor([r for _, r, _ in sample_rows], device=self.dev)
        nxt = F.linear(rmsnorm(x[rows], self.w["norm"], self.shape.eps), self.w["head"]).float().argmax(-1).tolist()
        for (sid, _, do_sample), tok in zip(sample_rows, nxt):
            if do_sample:
                self.tokens[sid].append(int(tok))

    def run(self):
        for st in self.d["steps"]:
            self.step(st)
        torch.cuda.synchronize()
        return {sid: t[self.prompt_len[sid]:] for sid, t in self.tokens.items()}


@torch.inference_mode()
def unpreempted_tokens(doc, shape, weights, dtype, device="cuda"):
    """The same requests, one at a time, never preempted: prefill + max_tokens decode steps through the
    same plugin over a private pool."""
    from light_vllm_b200.attention import PagedB200AttentionBackend as Backend
    bs = doc["block_size"]
    out = {}
    impl = Backend.get_impl_cls()(shape.heads, shape.head_dim, shape.head_dim ** -0.5, shape.kv_heads, None, None,
                                  "auto")
    builder_cls = Backend.get_builder_cls()
    for rid, prompt in enumerate(doc["prompts"]):
        total = len(prompt) + doc["max_tokens"]
        nblk = (total + bs - 1) // bs
        pools = [torch.zeros(Backend.get_kv_cache_shape(nblk, bs, shape.kv_heads, shape.head_dim), dtype=dtype,
                             device=device) for _ in range(shape.layers)]
        table = list(range(nblk))[::-1]
        toks = list(prompt)
        for step in range(doc["max_tokens"]):
            is_prompt = step == 0
            ctx = 0 if is_prompt else len(toks) - 1
            ids = toks[ctx:]
            g = types.SimpleNamespace(
                is_prompt=is_prompt, block_tables={0: table}, seq_ids=[0], input_tokens=[ids],
                orig_seq_lens=[len(toks)], seq_lens=[len(toks)], query_lens=[len(ids)], context_lens=[ctx],
                curr_sliding_window_blocks=[nblk], prefix_cache_hit=False)
            ib = types.SimpleNamespace(sliding_window=None, block_size=bs,
                                       scheduler_config=types.SimpleNamespace(use_v2_block_manager=False),
                                       chunked_prefill_enabled=False, inter_data_list=[g])
            meta = builder_cls(ib).build([len(toks)], [len(ids)], -1, 1)
            meta.to(device, non_blocking=True)
            x = weights["emb"][torch.tensor(ids, device=device)]
            positions = torch.arange(ctx, len(toks), device=device)

            def attn(l, q, k, v):
                T = q.shape[0]
                return impl.forward(q.reshape(T, -1), k.reshape(T, -1), v.reshape(T, -1), pools[l], meta).view_as(q)

            x = _layers_forward(shape, weights, None, x, positions, attn, None)
            logits = F.linear(rmsnorm(x[-1:], weights["norm"], shape.eps), weights["head"])
            toks.append(int(logits.float().argmax(-1)))
        out[rid] = toks[len(prompt):]
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("name", STREAMS)
def test_replay_matches_unpreempted_run(name):
    doc = load(name)
    shape = Qwen2Shape.tiny()
    w = {k: v.cuda() for k, v in init_weights(shape, 0).items()}
    want = unpreempted_tokens(doc, shape, w, torch.float32)
    r = Replay(doc, shape, w, torch.float32)
    got = r.run()
    assert len(got) == len(doc["prompts"])
    seq_to_req = {}
    for st in doc["steps"]:
        for g in st["groups"]:
            for s in g["seqs"]:
                seq_to_req[s["seq_id"]] = int(g["request_id"][1:])
    for sid, toks in got.items():
        assert len(toks) == doc["max_tokens"]
        assert toks == want[seq_to_req[sid]], (name, sid)


@pytest.mark.gpu
def test_two_decode_steps_in_flight_on_pooled_streams():
    """core/executor.py:62-185 keeps >= 2 steps in flight on non-default streams taken from a pool: two
    independent decode batches (own pools, own metadata, own outputs) enqueued back to back on two streams
    without a host synchronisation in between must give what each gives alone on the default stream."""
    from kv_util import make_case, to_device
    from light_vllm_b200 import _custom_ops as ops
    cases = [make_case(12, 32, 8, 128, 16, [900 + 37 * i for i in range(12)], torch.float16, seed=s) for s in (1, 2)]
    devs = [to_device(c, "cuda") for c in cases]

    def run(d, out):
        ops.paged_attention_v1(out, d["query"], d["key_cache"], d["value_cache"], 8, d["scale"], d["block_tables"],
                               d["seq_lens"], 16, d["max_seq_len"], None, "auto", 1.0, 1.0)

    alone = []
    for d in devs:
        o = torch.empty_like(d["query"])
        run(d, o)
        alone.append(o)
    torch.cuda.synchronize()
    pool = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [torch.zeros_like(d["query"]) for d in devs]
    for _ in range(8):  # several rounds: the same stream pool is reused step after step
        for s, d, o in zip(pool, devs, outs):
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                o.zero_()
                run(d, o)
    for s in pool:
        torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    for o, a in zip(outs, alone):
        assert torch.equal(o, a)
