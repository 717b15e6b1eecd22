#!/usr/bin/env python
"""NIZK::prove timing at a few sizes with the per-kernel profile (developer aid): python tools/quick_prove_bench.py 16 22"""
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import otti_b200 as sb  # noqa: E402

assert sb.lib().sb200_init(0) == 0
for s in [int(a) for a in sys.argv[1:]] or [16, 22]:
    n = 1 << s
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, bench.NUM_INPUTS, bench.SEED, pinned=True)
    gens = sb.NIZKGens.new(n, n, bench.NUM_INPUTS)
    for _ in range(3):
        sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(bench.LABEL), bench.TAPE)
    inst.upload_witness(gens, vars_)
    ts, profs = [], []
    for _ in range(20 if s <= 18 else 8):
        t0 = time.perf_counter()
        sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(bench.LABEL), bench.TAPE, resident=True)
        ts.append(1e3 * (time.perf_counter() - t0))
        profs.append(sb.NIZK.last_profile())
    p = profs[ts.index(min(ts))]
    print("2^%d prove (resident, wall clock): median %.3f ms, min %.3f | polycommit %.3f sc1 %.3f sc2 %.3f polyeval %.3f host_fs %.3f" % (
        s, statistics.median(ts), min(ts), 1e3 * p["polycommit"], 1e3 * p["prove_sc_phase_one"], 1e3 * p["prove_sc_phase_two"],
        1e3 * p["polyeval"], 1e3 * p["host_fiat_shamir"]), flush=True)
    kp = bench.kernel_profile(sb, lambda: sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(bench.LABEL), bench.TAPE, resident=True), 3)
    print("   " + ", ".join("%s %.0fx%.1fus" % (k.replace("_kernel", ""), c, 1e3 * ms / max(c, 1)) for k, c, ms in kp[:9]), flush=True)
    del gens, inst
