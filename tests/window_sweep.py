"""Developer aid: prove time vs fixed-base window width."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import otti_b200 as sb
assert sb.lib().sb200_init(0) == 0
s = int(sys.argv[1])
n = 1 << s
inst, v, i = sb.Instance.produce_synthetic_r1cs(n, n, 10, 0x0771)
for c in [int(x) for x in sys.argv[2:]]:
    os.environ["SB200_MSM_WINDOW"] = str(c)
    t0 = time.time()
    gens = sb.NIZKGens.new(n, n, 10)
    t1 = time.time()
    best = None
    for it in range(4):
        p = sb.NIZK.prove(inst, v, i, gens, sb.Transcript.new(b"nizk_example"), bytes(range(64)))
        pr = sb.NIZK.last_profile()
        if best is None or pr["total"] < best["total"]:
            best = pr
    print("2^%d c=%2d gens %.2fs  total %.3f ms  polycommit %.3f  msm kernel %.3f  polyeval %.3f" % (
        s, c, t1 - t0, best["total"] * 1e3, best["polycommit"] * 1e3, best["msm_commit_kernel"] * 1e3, best["polyeval"] * 1e3), flush=True)
    del gens
