#!/bin/bash
# Plan sweep (warps, stages, context split) for the small-batch shapes, graph-timed. Usage: tools/small_plan_sweep.sh
out=gpurun_out/small_plan_sweep.txt
: > $out
plans="0,0,0"
for w in 4 8; do for s in 2 3 4; do for c in 1 2 3 4 6 8; do plans="$plans;$w,$s,$c"; done; done; done
for v in 1 2; do
  for shape in "4 8192" "8 4096" "16 2048" "16 4096"; do
    set -- $shape
    timeout 280 python tools/sweep_attn.py --config c2 --version $v --batch $1 --ctx $2 --plans "$plans" --reps 32 --graph 8 2>&1 | grep "^{" | python -c "
import json,sys
rows=[json.loads(l) for l in sys.stdin]
rows=[r for r in rows if 'graph_us' in r]
base=[r for r in rows if r['plan']=='0,0,0']
rows.sort(key=lambda r:r['graph_us'])
for r in base+rows[:6]:
    g=r['resolved']; print(r['v'],r['B'],r['ctx'],r['plan'],'graph',r['graph_us'],r['graph_GBps'],'eager',r['us'],'grid',g['grid_x'],g['grid_y'],g['grid_z'],'c',g['cluster_size'],'w',g['num_warps'],'s',g['num_stages'],'occ',g['ctas_per_sm'])" >> $out
    echo >> $out
  done
done
cat $out
