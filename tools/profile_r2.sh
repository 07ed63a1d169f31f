#!/bin/bash
# Round-2 evidence (one B200): launch list of the bench command and one full capture of the decode attention
# kernel at c2 -- each only after the same command exited 0 without ncu.  (The tcgen05 prefill kernel's capture
# is taken with tools/prefix_bench.py, see profiles/r2_prefix.md.)
CMD="python bench.py --steps 3 --warmup 3 --cpu-budget 0 --no-extra-records --no-reference-cuda"
$CMD > gpurun_out/plain_r2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2.csv $CMD > gpurun_out/ncu_launch_r2.log 2>&1
echo launches rc=$?
ncu --set full --clock-control none --import-source on -k regex:paged_attention_kernel -s 8 -c 1 -f \
    -o gpurun_out/prof_attn_c2_r2 $CMD > gpurun_out/ncu_full_r2.log 2>&1
echo full rc=$?
ncu -i gpurun_out/prof_attn_c2_r2.ncu-rep --page raw --csv > gpurun_out/prof_attn_c2_r2.raw.csv 2>/dev/null
