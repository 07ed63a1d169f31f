#!/bin/bash
# Small-batch decode: eager vs graph-issued vs kernel-only (ncu gpu__time_duration) per shape.
# Usage: tools/small_probe.sh  (writes gpurun_out/small_probe.txt and gpurun_out/small_probe_ncu.csv)
out=gpurun_out/small_probe.txt
: > $out
for v in 1 2; do
  for shape in "1 2048" "1 8192" "1 32768" "2 8192" "2 32768" "4 8192" "8 4096" "16 2048" "16 4096" "32 2048"; do
    set -- $shape
    timeout 200 python tools/sweep_attn.py --config c2 --version $v --batch $1 --ctx $2 --plans "0,0,0" --reps 48 --graph 8 2>&1 | grep "^{" | python -c "
import json,sys
for l in sys.stdin:
    r=json.loads(l); g=r['resolved']; print(r['v'],r['B'],r['ctx'],'eager',r['us'],'graph',r['graph_us'],r['graph_GBps'],'grid',g['grid_x'],g['grid_y'],g['grid_z'],'c',g['cluster_size'],'w',g['num_warps'],'s',g['num_stages'],'smem',g['smem_bytes'],'occ',g['ctas_per_sm'])" >> $out
  done
done
cat $out
