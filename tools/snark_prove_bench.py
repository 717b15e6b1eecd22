#!/usr/bin/env python
"""SNARK::prove / verify timings (the snark_mode entries of bench.py) on their own: python tools/snark_prove_bench.py [log sizes]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import otti_b200 as sb  # noqa: E402

for s in [int(a) for a in sys.argv[1:]] or [16, 20]:
    n = 1 << s
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, bench.NUM_INPUTS, bench.SEED, pinned=True)
    gens = sb.SNARKGens.new(n, n, bench.NUM_INPUTS, n)
    out = {"log_size": s}
    out.update(bench.snark_prove_measurements(sb, inst, gens, n, s, vars_, inputs, oracle_check=s <= 20))
    print(json.dumps(out), flush=True)
    del gens, inst
