import sys, time, os
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
os.environ["SB200_TRACE_INSTANCE"] = "1"
import zkifgen, bench
import otti_b200 as sb
from otti_b200 import zkif
sb.lib().sb200_init(0)
g = zkifgen.OttiShaped(1 << 20, 1 << 20, 10, seed=11)
t = time.perf_counter()
r = zkif.R1cs.from_pointers(g.file_ptr(0), g.file_ptr(1), g.file_ptr(2))
print("parse %.3f s" % (time.perf_counter() - t))
for _ in range(2):
    t = time.perf_counter()
    inst = r.instance()
    print("instance %.3f s" % (time.perf_counter() - t))
    del inst
# kernel profile of the main workload pieces: spmvT3 before/after
n = 1 << 22
inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, 10, 0x0771, pinned=True)
gens = sb.NIZKGens.new(n, n, 10)
for _ in range(2):
    sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(b"nizk_example"), bytes(range(64)))
kp = bench.kernel_profile(sb, lambda: sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(b"nizk_example"), bytes(range(64))), 3)
for k, c, ms in kp[:12]:
    print("%-28s %5.1f launches %8.4f ms" % (k, c, ms))
