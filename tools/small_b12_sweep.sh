#!/bin/bash
# B = 1, 2: is one 8-warp CTA per SM also better than the head-split 4-warp plan?  (plans: warps,stages,split,head_rows)
out=gpurun_out/small_b12_sweep.txt
: > $out
plans="0,0,0;8,2,8,4;8,3,8,4;8,2,8,2;8,3,8,2;8,2,4,1;8,2,8,1;8,2,4,2;4,2,8,2;4,2,8,4;4,3,8,1"
for v in 1 2; do
  for shape in "1 8192" "1 32768" "2 8192" "2 32768"; do
    set -- $shape
    p="$plans"; [ $v = 2 ] && p="0,0,0;8,2,16;8,3,16;8,2,8;8,3,8;8,2,32;4,2,16;4,2,32;4,3,16;8,2,64;4,2,64"
    timeout 280 python tools/sweep_attn.py --config c2 --version $v --batch $1 --ctx $2 --plans "$p" --reps 32 --graph 8 2>&1 | grep "^{" | python -c "
import json,sys
rows=[json.loads(l) for l in sys.stdin]
rows=[r for r in rows if 'graph_us' in r]
base=[r for r in rows if r['plan']=='0,0,0']
rows.sort(key=lambda r:r['graph_us'])
for r in base+rows[:5]:
    g=r['resolved']; print(r['v'],r['B'],r['ctx'],r['plan'],'graph',r['graph_us'],r['graph_GBps'],'eager',r['us'],'grid',g['grid_x'],g['grid_y'],g['grid_z'],'c',g['cluster_size'],'w',g['num_warps'],'s',g['num_stages'],'occ',g['ctas_per_sm'])" >> $out
    echo >> $out
  done
done
cat $out
