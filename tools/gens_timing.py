"""Developer aid: NIZKGens::new wall time and the following first/steady prove, by size and proofs_expected hint."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import otti_b200 as sb
assert sb.lib().sb200_init(0) == 0
for s in [int(x) for x in sys.argv[1:]] or [16, 20, 22]:
    n = 1 << s
    t0 = time.perf_counter()
    inst, v, i = sb.Instance.produce_synthetic_r1cs(n, n, 10, 0x0771)
    t_inst = time.perf_counter() - t0
    for hint in (1, 8, None):
        t0 = time.perf_counter()
        gens = sb.NIZKGens.new(n, n, 10, proofs_expected=hint)
        t1 = time.perf_counter()
        times = []
        for it in range(4):
            ta = time.perf_counter()
            p = sb.NIZK.prove(inst, v, i, gens, sb.Transcript.new(b"nizk_example"), bytes(range(64)))
            times.append(time.perf_counter() - ta)
        ta = time.perf_counter()
        p.verify(inst, i, sb.Transcript.new(b"nizk_example"), gens)
        tv = time.perf_counter() - ta
        print("2^%d hint=%s c=%d  instance %.3fs  gens %.4fs  first prove %.2f ms  best %.2f ms  verify %.2f ms" % (
            s, hint, gens.window_bits, t_inst, t1 - t0, times[0] * 1e3, min(times) * 1e3, tv * 1e3), flush=True)
        del gens
