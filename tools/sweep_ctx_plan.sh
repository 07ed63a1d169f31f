for shape in "4 16384" "4 32768" "1 16384" "1 65536" "4 131072"; do
  set -- $shape
  python tools/sweep_attn.py --config c3 --version 2 --batch $1 --ctx $2 --layers 2 --reps 40 --plans "0,0,0;0,0,7;0,0,10;0,0,13;0,0,14;0,0,16;0,0,21;0,0,27;0,0,28;0,0,32;0,0,41;0,0,55;0,0,64" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l)
        if 'error' in d: print(d); continue
        g=d['resolved']; print(d['B'], d['ctx'], d['plan'], d['us'], d['GBps'], f\"{g['grid_x']}x{g['grid_y']}x{g['grid_z']}\")
"
done
