"""Exchange hooks (`sb200_comm_vt`, include/spartan_b200.h) for one proof sharded over several GPUs.

One process per GPU.  The sharded prover needs an all-gather of a few dozen HOST bytes per
Fiat-Shamir round; two transports are provided:

  ShmComm   ranks on one box: POSIX shared memory inside libspartan_b200 (~1 us per exchange)
  DistComm  any torch.distributed process group (NCCL on GPUs, gloo on CPU): a ctypes callback
            around dist.all_gather_into_tensor

Both present the same C struct to `NIZK.prove(..., comm=...)`.
"""
import ctypes as C

from . import lib, _check

_ALLGATHER = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)


class CommVT(C.Structure):
    _fields_ = [("ctx", C.c_void_p), ("rank", C.c_int), ("world", C.c_int), ("allgather", _ALLGATHER)]


class ShmComm:
    """`sb200_comm_shm_open`: collective over the ranks of one box; `name` must be unique to the job."""

    def __init__(self, name, rank, world):
        self.vt = CommVT()
        self.rank, self.world = rank, world
        _check(lib().sb200_comm_shm_open(name.encode(), C.c_int(rank), C.c_int(world), C.byref(self.vt)))

    def close(self):
        if self.vt.ctx:
            lib().sb200_comm_shm_close(C.byref(self.vt))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class NcclNativeComm:
    """`sb200_comm_nccl_open`: ncclAllGather called from inside libspartan_b200 (no Python in the loop).  The unique id
    travels through torch.distributed once, at construction."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        ids = [None]
        if self.rank == 0:
            buf = (C.c_uint8 * 128)()
            _check(lib().sb200_comm_nccl_unique_id(buf))
            ids = [bytes(buf)]
        dist.broadcast_object_list(ids, src=0, group=group)
        self.vt = CommVT()
        idbuf = (C.c_uint8 * 128).from_buffer_copy(ids[0])
        _check(lib().sb200_comm_nccl_open(idbuf, C.c_int(self.rank), C.c_int(self.world), C.byref(self.vt)))

    def close(self):
        if self.vt.ctx:
            lib().sb200_comm_nccl_close(C.byref(self.vt))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DistComm:
    """All-gather through `torch.distributed` (group=None: the default group).

    With a NCCL group the bytes are staged through a device tensor; with gloo they stay on the host.
    """

    def __init__(self, group=None, device=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        backend = dist.get_backend(group)
        self.device = torch.device(device) if device is not None else (
            torch.device("cuda", torch.cuda.current_device()) if "nccl" in backend else torch.device("cpu"))
        self.calls = 0
        self._cb = _ALLGATHER(self._allgather)  # keep the thunk alive
        self.vt = CommVT(None, self.rank, self.world, self._cb)

    def _allgather(self, ctx, send, nbytes, recv):
        try:
            torch = self.torch
            src = torch.frombuffer((C.c_uint8 * nbytes).from_address(send), dtype=torch.uint8)
            dst = torch.frombuffer((C.c_uint8 * (nbytes * self.world)).from_address(recv), dtype=torch.uint8)
            if self.device.type == "cpu":
                self.dist.all_gather_into_tensor(dst, src, group=self.group)
            else:
                out = torch.empty(nbytes * self.world, dtype=torch.uint8, device=self.device)
                self.dist.all_gather_into_tensor(out, src.to(self.device), group=self.group)
                dst.copy_(out)  # synchronises
            self.calls += 1
            return 0
        except Exception as e:  # never let an exception cross the C boundary
            import sys
            print("DistComm.allgather failed: %r" % (e,), file=sys.stderr)
            return -1

    def close(self):
        pass


def selftest(comm, rounds=8):
    """Host-only check of a hook: returns (digest, expected digest)."""
    got, want = (C.c_uint8 * 32)(), (C.c_uint8 * 32)()
    _check(lib().sb200_comm_selftest(C.byref(comm.vt), C.c_int(rounds), got))
    _check(lib().sb200_comm_selftest_expected(C.c_int(comm.world), C.c_int(rounds), want))
    return bytes(got), bytes(want)
