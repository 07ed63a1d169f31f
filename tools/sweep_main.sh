#!/bin/bash
python tools/sweep_attn.py --config c2 --plans "0,0,0;4,3,1;4,2,1"
python tools/sweep_attn.py --config c3 --plans "0,0,0;4,3,0"
python tools/sweep_attn.py --config c3 --version 1 --plans "0,0,0;4,3,0"
python tools/sweep_attn.py --config c4 --plans "0,0,0"
python tools/sweep_attn.py --config c4 --version 2 --plans "0,0,0"
for B in 1 8 16 32; do python tools/sweep_attn.py --config c2 --batch $B --ctx 4096 --layers 2 --plans "0,0,0;4,3,0"; done
python tools/sweep_attn.py --config c2 --batch 8 --ctx 32768 --layers 2 --version 2 --plans "0,0,0;4,3,0" --ref
python tools/sweep_attn.py --config c2 --batch 2 --ctx 131072 --layers 2 --version 2 --plans "0,0,0;4,3,0" --ref
