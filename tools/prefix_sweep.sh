#!/bin/bash
# The four prefix-prefill shapes of profiles/r1_prefix.md, both tile loops.  Usage: tools/prefix_sweep.sh <out.jsonl>
out=${1:-gpurun_out/prefix_sweep.jsonl}
: > "$out"
for path in ${PATHS:-tcgen05 mma_sync}; do
  timeout 300 python tools/prefix_bench.py --path $path --batch 16 --ctx 4096 --qlen 512 >> "$out" 2>&1
  timeout 300 python tools/prefix_bench.py --path $path --batch 8 --ctx 0 --qlen 2048 >> "$out" 2>&1
  timeout 300 python tools/prefix_bench.py --path $path --batch 64 --ctx 2048 --qlen 64 >> "$out" 2>&1
  timeout 300 python tools/prefix_bench.py --path $path --batch 8 --ctx 16384 --qlen 256 >> "$out" 2>&1
done
cat "$out"
