#!/bin/bash
# Round-1 evidence run (one B200): tests, bench lines for c2/c3/c4, sweeps (untruncated), then the
# ncu launch list and one full capture of the attention kernel -- each only after the same command
# exited 0 without ncu.
set -x
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest.log 2>&1; echo pytest rc=$?
python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python bench.py --config c3 --cpu-budget 8 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
python bench.py --config c4 --cpu-budget 8 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_c2.json 2> gpurun_out/bench_ref_c2.err
bash tools/sweep_main.sh > gpurun_out/sweep_final.jsonl 2>&1
python bench.py --steps 5 --warmup 3 --cpu-budget 0 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv \
    python bench.py --steps 5 --warmup 3 --cpu-budget 0 > gpurun_out/ncu_launch.log 2>&1
echo launches rc=$?
ncu --set full --clock-control none --import-source on -k regex:paged_attention_kernel -s 8 -c 2 \
    -o gpurun_out/prof_attn_c2_r1 python bench.py --steps 5 --warmup 3 --cpu-budget 0 > gpurun_out/ncu_full.log 2>&1
echo full rc=$?
