"""One proof sharded over several ranks.

CPU (`-m "not gpu"`): the exchange hooks -- the shared-memory transport of libspartan_b200 with 2 and 4
processes, and the torch.distributed transport with a world_size-2 gloo group -- through the library's
host-only self-test (all-gather + modular sums + transcript, as a sharded sum-check round does).
GPU (`-m gpu`): 2, 4 and 8 ranks (processes sharing the one GPU of the test box) prove the same R1CS;
every rank must return the oracle's proof bytes, and leave the caller's transcript in the same state."""
import os
import subprocess
import sys
import uuid

import pytest

import oracle_lib as O
from common import LABEL, TAPE_SEED, canon, ragged_r1cs

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _shm_rank(rank, world, name, q):
    sys.path.insert(0, ROOT)
    from otti_b200 import comm
    c = comm.ShmComm(name, rank, world)
    got, want = comm.selftest(c, 40)
    q.put((rank, got, want))
    c.close()


@pytest.mark.parametrize("world", [1, 2, 4])
def test_shm_exchange(world):
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    name = "/sb200-t-" + uuid.uuid4().hex[:12]
    q = ctx.Queue()
    ps = [ctx.Process(target=_shm_rank, args=(r, world, name, q)) for r in range(world)]
    [p.start() for p in ps]
    res = [q.get(timeout=120) for _ in range(world)]
    [p.join(60) for p in ps]
    assert sorted(r[0] for r in res) == list(range(world))
    for _, got, want in res:
        assert got == want


def _gloo_rank(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    from otti_b200 import comm
    dist.init_process_group("gloo", rank=rank, world_size=world)
    c = comm.DistComm()
    got, want = comm.selftest(c, 12)
    q.put((rank, got, want, c.calls))
    dist.destroy_process_group()


def test_torch_distributed_exchange_gloo_world2():
    import multiprocessing as mp
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_gloo_rank, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = [q.get(timeout=300) for _ in range(2)]
    [p.join(60) for p in ps]
    for _, got, want, calls in res:
        assert got == want and calls == 12


def _run_sharded(world, case, tmp_path, env=None):
    name = "/sb200-g-" + uuid.uuid4().hex[:12]
    out = str(tmp_path / "proof")
    e = dict(os.environ)
    e.update(env or {})
    ps = [subprocess.Popen([sys.executable, os.path.join(HERE, "shard_worker.py"), str(r), str(world), name, "0", case, out],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, env=e) for r in range(world)]
    logs = [p.communicate(timeout=600)[0].decode() for p in ps]
    assert all(p.returncode == 0 for p in ps), "\n".join(logs)
    return [open("%s.%d" % (out, r), "rb").read() for r in range(world)]


@pytest.mark.gpu
@pytest.mark.parametrize("world,s", [(2, 2), (2, 5), (2, 10), (4, 4), (4, 9), (8, 6), (8, 11), (4, 13)])
def test_sharded_prove_synthetic(world, s, tmp_path):
    n = 1 << s
    ni = min(10, n - 1)
    oi, ov, oin = O.Instance.synthetic(n, n, ni, 0x0771)
    og = O.Gens(n, n, ni)
    want = O.nizk_prove(oi, ov, oin, og, LABEL, TAPE_SEED)
    res = _run_sharded(world, "syn:%d" % s, tmp_path)
    single = _run_sharded(1, "syn:%d" % s, tmp_path)[0]
    for r in res:
        assert r[16:] == want
        assert r[:16] == single[:16]  # transcript state after the proof, as on one GPU


@pytest.mark.gpu
@pytest.mark.parametrize("world,dims", [(2, (37, 20, 5)), (4, (300, 100, 64)), (2, (64, 600, 2)), (8, (1000, 700, 17))])
def test_sharded_prove_ragged(world, dims, tmp_path):
    nc, nv, ni = dims
    A, B, Cm, v, i = ragged_r1cs(nc, nv, ni)
    oi = O.Instance.new(nc, nv, ni, A, B, Cm)
    og = O.Gens(nc, nv, ni)
    want = O.nizk_prove(oi, canon(v), canon(i), og, LABEL, TAPE_SEED)
    for r in _run_sharded(world, "rag:%d,%d,%d" % dims, tmp_path):
        assert r[16:] == want


@pytest.mark.gpu
def test_sharded_too_small_is_refused(tmp_path):
    name = "/sb200-g-" + uuid.uuid4().hex[:12]
    p = subprocess.run([sys.executable, os.path.join(HERE, "shard_worker.py"), "0", "8", name, "0", "syn:2", str(tmp_path / "x")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=300)
    assert p.returncode != 0 and b"too small" in p.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world,tail,s", [(1, 0, 9), (1, 3, 9), (1, 8, 9), (1, 8, 5), (2, 0, 9), (2, 8, 9), (4, 5, 12), (1, 6, 2)])
def test_host_tail_rounds_do_not_change_the_proof(world, tail, s, tmp_path):
    """The last SB200_HOST_TAIL_LOG rounds of each sum-check run on the host over the short tables (prover.inc);
    whatever the split between device and host rounds, the proof bytes are the oracle's."""
    n = 1 << s
    ni = min(10, n - 1)
    oi, ov, oin = O.Instance.synthetic(n, n, ni, 0x0771)
    og = O.Gens(n, n, ni)
    want = O.nizk_prove(oi, ov, oin, og, LABEL, TAPE_SEED)
    for r in _run_sharded(world, "syn:%d" % s, tmp_path, {"SB200_HOST_TAIL_LOG": str(tail)}):
        assert r[16:] == want


@pytest.mark.gpu
@pytest.mark.parametrize("world,case", [(2, "syn:16"), (4, "syn:16"), (4, "rag:48,40000,3"), (8, "rag:48,33000,3"), (2, "syn:18"), (4, "syn:18")])
def test_sharded_witness_exchange(world, case, tmp_path):
    """Witnesses of >= 1 MiB: each rank uploads only its block of commitment rows and reads the other blocks from the
    peers' HBM (CUDA IPC, prover.inc p2p_setup) -- also when the zero padding of a non-power-of-two witness leaves some
    ranks a short or empty block.  Same bytes as the oracle, with and without the exchange."""
    kind, arg = case.split(":")
    if kind == "syn":
        n = 1 << int(arg)
        oi, ov, oin = O.Instance.synthetic(n, n, 10, 0x0771)
        og = O.Gens(n, n, 10)
    else:
        nc, nv, ni = (int(x) for x in arg.split(","))
        A, B, Cm, v, i = ragged_r1cs(nc, nv, ni)
        oi, ov, oin = O.Instance.new(nc, nv, ni, A, B, Cm), canon(v), canon(i)
        og = O.Gens(nc, nv, ni)
    want = O.nizk_prove(oi, ov, oin, og, LABEL, TAPE_SEED, 8)
    for env in (None, {"SB200_P2P": "0"}):
        for r in _run_sharded(world, case, tmp_path, env):
            assert r[16:] == want


@pytest.mark.gpu
@pytest.mark.parametrize("world,s,tail", [(2, 13, None), (4, 14, "0"), (2, 14, "6"), (8, 16, None), (4, 13, "fallback")])
def test_sharded_factored_first_sumcheck(world, s, tail, tmp_path):
    """Every device round of phase one through the factored kernel (SB200_SC_LAT_PAIRS=0; kernels.cu sc_cubic_eq_kernel):
    each rank's suffix eq tables carry the factor of the index bits its rank fixes, the per-rank (q(0), leading coefficient)
    pairs are summed before the host rebuilds s(0), s(2), s(3).  Same bytes as the oracle."""
    n = 1 << s
    oi, ov, oin = O.Instance.synthetic(n, n, 10, 0x0771)
    want = O.nizk_prove(oi, ov, oin, O.Gens(n, n, 10), LABEL, TAPE_SEED, 8)
    env = {"SB200_SC_LAT_PAIRS": "0", "SB200_SC_HALF_ZERO": "2"}  # and phase two's half-empty z from the first size on
    if tail == "fallback":
        env["SB200_SC_HALF_ZERO_FALLBACK"] = "1"  # the path of the challenge r_0 = 1: ordinary bind over cleared upper halves
    elif tail is not None:
        env["SB200_HOST_TAIL_LOG"] = tail
    for r in _run_sharded(world, "syn:%d" % s, tmp_path, env):
        assert r[16:] == want
