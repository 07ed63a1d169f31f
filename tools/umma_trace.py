"""Debug: per-step clock64 timeline of one CTA of the tcgen05 prefill kernel.  Needs a trace build:
    make -C light_vllm_b200/csrc all EXTRA_NVFLAGS=-DB200_UMMA_TRACE   (touch prefix_prefill.cu first)
Prints, for steps 20..39, per M tile: softmax phases (wait S, tmem ld, max/mask, exp, st+fence) and when each
phase started relative to M tile 0's step start; and the MMA issuer's waits (P, V, K) and issue time."""
import ctypes
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "tools")
import torch  # noqa: E402
from light_vllm_b200 import _lib  # noqa: E402
import prefix_bench  # noqa: E402

sys.argv = ["prefix_bench.py", "--path", "tcgen05", "--steps", "2", "--warmup", "1"]
prefix_bench.main()
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (4 * 128 * 8))()
assert _lib.lib.b200_debug_umma_trace(buf) == 0
T = lambda r, s, i: buf[(r * 128 + s) * 8 + i]
print("step | tile: waitS ld max exp st | start offset of tile1 vs tile0 | period")
for s in range(20, 40):
    row = []
    for w in (0, 1):
        d = [T(w, s, i + 1) - T(w, s, i) for i in range(5)]
        row.append(" ".join(f"{x:5d}" for x in d))
    off = T(1, s, 0) - T(0, s, 0)
    exp_off = T(1, s, 3) - T(0, s, 3)
    per = T(0, s + 1, 0) - T(0, s, 0)
    print(f"{s:3d} | t0: {row[0]} | t1: {row[1]} | step-start off {off:6d} exp-start off {exp_off:6d} | period {per}")
print("MMA issuer: step | tile: waitP waitV waitK issue | P-ready -> issue done")
for s in range(20, 40):
    row = []
    for w in (0, 1):
        d = [T(2 + w, s, i + 1) - T(2 + w, s, i) for i in range(4)]
        lat = T(2 + w, s, 4) - T(w, s, 5)
        row.append(" ".join(f"{x:5d}" for x in d) + f" | after P {lat:5d}")
    print(f"{s:3d} | t0: {row[0]} | t1: {row[1]}")
