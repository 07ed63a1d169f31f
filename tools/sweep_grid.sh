#!/bin/bash
# small-batch / long-context grid: which v1 cluster size (and v2) wins where
for B in 1 4 16 32; do for CTX in 2048 8192; do
  python tools/sweep_attn.py --config c2 --batch $B --ctx $CTX --layers 2 --plans "0,0,0;4,3,1;4,3,2;4,3,4;4,3,8"
  python tools/sweep_attn.py --config c2 --batch $B --ctx $CTX --layers 2 --version 2 --plans "0,0,0" --ref
done; done
python tools/sweep_attn.py --config c3 --ref --plans "0,0,0;8,3,0;4,2,0"
python tools/sweep_attn.py --config c3 --version 1 --ref --plans "0,0,0"
python tools/sweep_attn.py --config c4 --ref --plans "0,0,0;8,3,0;4,6,0;4,4,0"
python tools/sweep_attn.py --config c4 --version 2 --plans "0,0,0"
