"""One rank of a sharded NIZK::prove (launched by tests/test_sharded.py and usable by hand):
  python tests/shard_worker.py <rank> <world> <shm name> <device> <case> <out file>
case = "syn:<log2 n>" (produce_synthetic_r1cs) or "rag:<nc>,<nv>,<ni>" (tests/common.ragged_r1cs).
Writes the proof bytes to <out file>.<rank>.  Several ranks may share one GPU (device index the same):
the kernels never wait on each other, only the hosts do."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def main():
    rank, world, name, device, case, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4]), sys.argv[5], sys.argv[6]
    import otti_b200 as sb
    from otti_b200 import comm as sbcomm
    from common import LABEL, TAPE_SEED, canon, ragged_r1cs
    assert sb.lib().sb200_init(device) == 0, sb.lib().sb200_last_error()
    kind, arg = case.split(":")
    if kind == "syn":
        n = 1 << int(arg)
        ni = min(10, n - 1)
        inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, ni, 0x0771, rank=rank, world=world)
        gens = sb.NIZKGens.new(n, n, ni)
    else:
        nc, nv, ni = (int(x) for x in arg.split(","))
        A, B, Cm, v, i = ragged_r1cs(nc, nv, ni)
        inst = sb.Instance.new(nc, nv, ni, A, B, Cm, rank=rank, world=world)
        vars_, inputs = sb.VarsAssignment.new(canon(v)), sb.InputsAssignment.new(canon(i))
        gens = sb.NIZKGens.new(nc, nv, ni)
    c = sbcomm.ShmComm(name, rank, world) if world > 1 else None
    tr = sb.Transcript.new(LABEL)
    proof = sb.NIZK.prove(inst, vars_, inputs, gens, tr, TAPE_SEED, comm=c)
    # the caller's transcript must end in the same state on every rank
    tail = tr.challenge_bytes(b"after", 16)
    with open("%s.%d" % (out, rank), "wb") as f:
        f.write(tail + proof.to_bytes())
    if c is not None:
        c.close()


if __name__ == "__main__":
    main()
