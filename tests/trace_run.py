"""Developer aid: one traced NIZK::prove per size (SB200_TRACE=1 timeline on stderr)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import otti_b200 as sb
assert sb.lib().sb200_init(0) == 0
s = int(sys.argv[1]) if len(sys.argv) > 1 else 16
n = 1 << s
inst, v, i = sb.Instance.produce_synthetic_r1cs(n, n, 10, 0x0771)
gens = sb.NIZKGens.new(n, n, 10)
for it in range(3):
    if it == 2:
        os.environ["SB200_TRACE"] = os.environ.get("TRACE_LEVEL", "1")
    p = sb.NIZK.prove(inst, v, i, gens, sb.Transcript.new(b"nizk_example"), bytes(range(64)))
    print({k: round(x * 1e3, 3) for k, x in sb.NIZK.last_profile().items()}, file=sys.stderr)
