for shape in "1 2048" "1 8192" "1 32768" "2 8192" "4 8192" "8 4096" "16 4096" "16 2048" "32 2048"; do
  set -- $shape
  timeout 300 python tools/sweep_attn.py --config c2 --version 2 --batch $1 --ctx $2 --plans "0,0,0" --reps 50 2>&1 | grep "^{" | python -c "
import json,sys
for l in sys.stdin:
    r=json.loads(l); print(r['v'],r['B'],r['ctx'],r['us'],r['GBps'])"
done
