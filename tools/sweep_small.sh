#!/bin/bash
# Small-batch decode grid: library plan vs the same call with the GQA group kept in one CTA
# (head_rows = 4) vs the reference CUDA kernel, c2 geometry (fp16 H32/KH8/D128/BS16).  Usage: tools/sweep_small.sh <out.jsonl>
out=${1:-gpurun_out/sweep_small.jsonl}
: > "$out"
for v in 1 2; do
  for shape in "1 2048" "1 8192" "1 32768" "2 8192" "4 8192" "8 4096" "16 4096" "16 2048" "32 2048"; do
    set -- $shape
    timeout 300 python tools/sweep_attn.py --config c2 --version $v --batch $1 --ctx $2 --plans "0,0,0;0,0,0,4" --ref --reps 50 >> "$out" 2>&1
  done
done
python - "$out" <<'PY'
import json, sys
rows = [json.loads(l) for l in open(sys.argv[1]) if l.startswith("{")]
print("| v | B | ctx | plan | us | GB/s | frac of 6550 | grid |")
print("|---|---|---|---|---|---|---|---|")
for r in rows:
    if "error" in r:
        print("| error |", r); continue
    g = r.get("resolved", {})
    grid = f'{g.get("grid_x")}x{g.get("grid_y")}x{g.get("grid_z")} c{g.get("cluster_size")}' if g else "-"
    print(f'| {r["v"]} | {r["B"]} | {r["ctx"]} | {r.get("plan", r["impl"])} | {r["us"]} | {r["GBps"]} | {r["GBps"]/6550:.3f} | {grid} |')
PY
