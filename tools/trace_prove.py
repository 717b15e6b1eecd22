#!/usr/bin/env python
"""One traced NIZK::prove per size (SB200_TRACE=1: host-side marks of prover.inc on stderr): python tools/trace_prove.py 22"""
import os
import sys

os.environ.setdefault("SB200_TRACE", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import otti_b200 as sb  # noqa: E402

assert sb.lib().sb200_init(0) == 0
for s in [int(a) for a in sys.argv[1:]] or [22]:
    n = 1 << s
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, bench.NUM_INPUTS, bench.SEED, pinned=True)
    gens = sb.NIZKGens.new(n, n, bench.NUM_INPUTS)
    inst.upload_witness(gens, vars_)
    for k in range(4):
        sys.stderr.write("==== 2^%d proof %d\n" % (s, k))
        sys.stderr.flush()
        sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(bench.LABEL), bench.TAPE, resident=True)
