#!/bin/bash
# One `ncu --set full` capture of the attention kernel per config (after the same command exited 0
# without ncu); the raw page is exported to CSV on the box so only summaries travel back.
for c in c2 c3 c4; do
  python bench.py --config $c --steps 3 --warmup 3 --cpu-budget 0 > gpurun_out/plain_$c.log 2>&1 || { echo "plain $c failed"; continue; }
  ncu --set full --clock-control none --import-source on -k regex:paged_attention_kernel -s 8 -c 1 \
      -f -o gpurun_out/prof_attn_$c python bench.py --config $c --steps 3 --warmup 3 --cpu-budget 0 > gpurun_out/ncu_full_$c.log 2>&1
  echo "full $c rc=$?"
  ncu -i gpurun_out/prof_attn_$c.ncu-rep --page raw --csv > gpurun_out/prof_attn_$c.raw.csv 2>/dev/null
done
