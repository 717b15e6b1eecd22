#!/usr/bin/env python
"""Timeline of ONE sharded proof on rank 0 and the last rank (developer aid):
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 tools/shard_trace.py 22"""
import os
import sys
import time
import uuid

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402
import otti_b200 as sb  # noqa: E402
from otti_b200 import comm as sbcomm  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
assert sb.lib().sb200_init(local) == 0
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
s = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n = 1 << s
name = ["/sb200-trace-" + uuid.uuid4().hex[:12]]
dist.broadcast_object_list(name, src=0)
comm = sbcomm.ShmComm(name[0], rank, world)
inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, bench.NUM_INPUTS, bench.SEED, pinned=True, rank=rank, world=world)
gens = sb.NIZKGens.new(n, n, bench.NUM_INPUTS)
for _ in range(3):
    sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(bench.LABEL), bench.TAPE, comm=comm)
inst.upload_witness(gens, vars_)
for it in range(4):
    dist.barrier()
    torch.cuda.synchronize()
    if it == 3 and rank in (0, world - 1):
        os.environ["SB200_TRACE"] = "1"
    t0 = time.perf_counter()
    sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(bench.LABEL), bench.TAPE, resident=True, comm=comm)
    dt = 1e3 * (time.perf_counter() - t0)
    p = sb.NIZK.last_profile()
    if rank in (0, world - 1):
        print("rank %d iter %d: %.3f ms  polycommit %.3f sc1 %.3f sc2 %.3f polyeval %.3f exchange %.3f host_fs %.3f" % (
            rank, it, dt, 1e3 * p["polycommit"], 1e3 * p["prove_sc_phase_one"], 1e3 * p["prove_sc_phase_two"], 1e3 * p["polyeval"],
            1e3 * p["exchange"], 1e3 * p["host_fiat_shamir"]), file=sys.stderr, flush=True)
comm.close()
dist.destroy_process_group()
