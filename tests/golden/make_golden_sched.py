"""Records block-id streams of the reference's OWN decoding scheduler + block managers under artificial
preemption, for replay on the GPU box (where /root/reference does not exist).

Runs here (CPU box): imports light_vllm.decoding.scheduler (DecodingScheduler, scheduler.py:234-1075) and the
v1 / v2 block managers unmodified -- `pynvml` is stubbed because the package probes NVML at import -- sets
VLLM_TEST_ENABLE_ARTIFICIAL_PREEMPT (scheduler.py:22-26) and drives schedule() step by step the way the
engine does: scheduler output -> (dummy) sampler output -> the reference's SingleStepOutputProcessor
(fork / free / stop checks) -> free_finished_request.  Token VALUES do not influence scheduling (ignore_eos,
max_tokens), so the recorded stream -- per step: which sequences run as prefill / decode with which block
tables, and blocks_to_swap_in / blocks_to_swap_out / blocks_to_copy -- is valid for any model.

    python tests/golden/make_golden_sched.py        # writes tests/golden/sched_stream_*.json

tests/test_engine_replay.py replays the streams over PagedB200AttentionBackend + light_vllm_b200.cache_engine
and checks that greedy tokens equal the run without preemption.
"""
import json
import os
import random
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def _stub_pynvml():
    stub = types.ModuleType("pynvml")
    stub.__file__ = "pynvml.py"
    stub.NVMLError = type("NVMLError", (Exception,), {})
    stub.nvmlInit = lambda: None
    stub.nvmlShutdown = lambda: None
    stub.nvmlDeviceGetCount = lambda: 1
    stub.nvmlDeviceGetHandleByIndex = lambda i: i
    stub.nvmlDeviceGetCudaComputeCapability = lambda h: (10, 0)
    stub.nvmlDeviceGetName = lambda h: "NVIDIA B200"
    sys.modules["pynvml"] = stub


def import_reference():
    os.environ["VLLM_TEST_ENABLE_ARTIFICIAL_PREEMPT"] = "1"
    _stub_pynvml()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import light_vllm  # noqa: F401
    from light_vllm.decoding import scheduler as sched_mod
    return sched_mod


def record(name, *, block_manager, preemption_mode, n, prompts, max_tokens, block_size, num_gpu_blocks,
           num_cpu_blocks, max_num_seqs, seed):
    sched_mod = import_reference()
    from light_vllm.decoding.backends.sampling_params import SamplingParams
    from light_vllm.decoding.config import CacheConfig, SchedulerConfig
    from light_vllm.decoding.processor.utils.single_step import SingleStepOutputProcessor
    from light_vllm.decoding.processor.utils.stop_checker import StopChecker
    from light_vllm.decoding.schema.execute_io import CompletionSequenceGroupOutput, SequenceOutput
    from light_vllm.decoding.schema.sequence import Logprob, Sequence, SequenceGroup, SequenceStatus
    from light_vllm.utils import Counter

    assert sched_mod.ENABLE_ARTIFICIAL_PREEMPT
    random.seed(seed)  # the artificial preemption draws from `random` (scheduler.py:824-842)
    max_model_len = max(len(p) for p in prompts) + max_tokens + 8
    sc = SchedulerConfig(max_num_batched_tokens=4096, max_num_seqs=max_num_seqs, max_model_len=max_model_len,
                         use_v2_block_manager=(block_manager == "v2"), preemption_mode=preemption_mode,
                         scheduling="sync")
    cc = CacheConfig.__new__(CacheConfig)
    cc.block_size = block_size
    cc.num_gpu_blocks = num_gpu_blocks
    cc.num_cpu_blocks = num_cpu_blocks
    cc.sliding_window = None
    cc.enable_prefix_caching = False
    sched = sched_mod.DecodingScheduler(sc, cc, request_processor=None)
    # need_scheduling() (scheduler.py:843-855) answers False when every group is swapped out and nothing
    # waits -- the engine would stall there; the recorder keeps calling the unmodified schedule() instead.
    sched.need_scheduling = lambda: bool(sched.waiting or sched.running or sched.swapped)
    seq_counter = Counter()
    out_proc = SingleStepOutputProcessor(None, seq_counter, StopChecker(max_model_len, None), max_model_len)
    sp = SamplingParams(n=n, best_of=n, temperature=1.0 if n > 1 else 0.0, max_tokens=max_tokens, ignore_eos=True,
                        detokenize=False)

    class Inp:
        def __init__(self, ids):
            self.prompt = None
            self.prompt_token_ids = ids

    groups = {}
    for i, ids in enumerate(prompts):
        seq = Sequence(next(seq_counter), Inp(list(ids)), block_size, eos_token_id=None)
        g = SequenceGroup(request_id=f"r{i}", seqs=[seq], arrival_time=float(i), sampling_params=sp)
        groups[g.request_id] = g
        sched.waiting.append(g)
        if hasattr(sched, "requests"):
            sched.requests.add(g.request_id)

    steps = []
    n_swap = n_copy = n_recompute = 0
    for _ in range(20000):
        out = sched.schedule()
        if out is None or (not out.scheduled_seq_groups and not sched.waiting and not sched.running
                           and not sched.swapped):
            break
        rec = {"swap_in": [list(map(int, p)) for p in out.blocks_to_swap_in],
               "swap_out": [list(map(int, p)) for p in out.blocks_to_swap_out],
               "copy": [list(map(int, p)) for p in out.blocks_to_copy], "groups": []}
        n_swap += len(rec["swap_in"]) + len(rec["swap_out"])
        n_copy += len(rec["copy"])
        for ssg, meta in zip(out.scheduled_seq_groups, out.seq_group_metadata_list):
            g = {"request_id": meta.request_id, "is_prompt": bool(meta.is_prompt),
                 "token_chunk_size": int(ssg.token_chunk_size), "do_sample": bool(meta.do_sample), "seqs": []}
            for sid, data in meta.seq_data.items():
                g["seqs"].append({"seq_id": int(sid), "len": int(data.get_len()),
                                  "num_computed": int(data.get_num_computed_tokens()),
                                  "prompt_len": int(data.get_prompt_len()),
                                  "block_table": [int(b) for b in meta.block_tables[sid]]})
            if meta.is_prompt and g["seqs"][0]["len"] > g["seqs"][0]["prompt_len"]:
                n_recompute += 1
            rec["groups"].append(g)
        # ---- what the engine does with the sampler output (output_processor.py:67-110) ----
        finished = []
        for ssg, meta in zip(out.scheduled_seq_groups, out.seq_group_metadata_list):
            sg = ssg.seq_group
            sg.update_num_computed_tokens(ssg.token_chunk_size)
            if meta.do_sample:
                running = sg.get_seqs(status=SequenceStatus.RUNNING)
                samples = []
                if meta.is_prompt:
                    parent = running[0]
                    for _k in range(sp.n):
                        samples.append(SequenceOutput(parent.seq_id, 7, {7: Logprob(0.0)}))
                else:
                    for s in running:
                        samples.append(SequenceOutput(s.seq_id, 7, {7: Logprob(0.0)}))
                fork, free = out_proc.process_outputs(sg, [CompletionSequenceGroupOutput(samples, None)])
                rec.setdefault("forks", []).extend([[int(p.seq_id), int(c.seq_id)] for p, c in fork])
                for parent, child in fork:
                    sched.fork_seq(parent, child)
                for s in free:
                    sched.free_seq(s)
            if sg.is_finished():
                finished.append(sg)
        steps.append(rec)

        class _RO:
            def __init__(self, rid):
                self.request_id = rid

        sched.free_finished_request([_RO(ssg.seq_group.request_id) for ssg in out.scheduled_seq_groups])
    assert all(g.is_finished() for g in groups.values()), "scheduler did not drain"
    final = {rid: {str(s.seq_id): {"prompt_len": s.get_prompt_len(), "output_len": s.get_output_len()}
                   for s in g.get_seqs()} for rid, g in groups.items()}
    doc = {"name": name, "block_manager": block_manager, "preemption_mode": preemption_mode, "n": n,
           "prompts": [list(p) for p in prompts], "max_tokens": max_tokens, "block_size": block_size,
           "num_gpu_blocks": num_gpu_blocks, "num_cpu_blocks": num_cpu_blocks, "seed": seed,
           "counts": {"steps": len(steps), "swapped_block_pairs": n_swap, "copied_block_pairs": n_copy,
                      "recompute_prefills": n_recompute,
                      "cumulative_preemptions": int(sched.num_cumulative_preemption)},
           "final": final, "steps": steps,
           "source": "light_vllm/decoding/scheduler.py + core/block_manager_%s.py, unmodified, "
                     "VLLM_TEST_ENABLE_ARTIFICIAL_PREEMPT=1" % block_manager}
    path = os.path.join(HERE, f"sched_stream_{name}.json")
    with open(path, "w") as f:
        json.dump(doc, f, separators=(",", ":"))
    print(name, doc["counts"], os.path.getsize(path), "bytes")


def main():
    rng = random.Random(0)
    prompts = [[rng.randrange(3, 150) for _ in range(rng.choice([5, 17, 33, 64, 90, 128]))] for _ in range(8)]
    common = dict(prompts=prompts, max_tokens=48, block_size=16, num_gpu_blocks=96, num_cpu_blocks=96,
                  max_num_seqs=8)
    record("recompute_v1", block_manager="v1", preemption_mode="recompute", n=1, seed=11, **common)
    record("swap_v1", block_manager="v1", preemption_mode="swap", n=1, seed=12, **common)
    record("swap_v2", block_manager="v2", preemption_mode="swap", n=1, seed=13, **common)
    # n > 1 (fork -> copy-on-write -> blocks_to_copy) is not recordable: the reference's
    # SingleStepOutputProcessor only implements n == 1 (single_step.py:82-99), so its live paths never emit
    # blocks_to_copy without beam search / prefix caching.  copy_blocks is covered by tests/test_cache_ops_gpu.py.


if __name__ == "__main__":
    main()
