import sys, json
src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin  # a file argument, else stdin
for l in src:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(d["roofline"]["path"], d["config"]["workload"][:48], "ms", d["ms_per_call"], "TF", d["value"], "frac", d["roofline"]["frac"], "sdpa_ms", d["sdpa_library_ms"], "x", d["speedup_vs_sdpa"], "err", d["max_abs_vs_sdpa"])
