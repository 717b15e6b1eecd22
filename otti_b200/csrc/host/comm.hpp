// In-node exchange for one proof sharded over several GPUs (one process per GPU, one box).
//
// What the sharded prover exchanges is tiny and lands on the HOST anyway (the Fiat-Shamir transcript
// needs it): 96 bytes per rank per sum-check round, one table entry per rank at the end of a
// sum-check, and each rank's block of compressed commitment rows.  The latency of that exchange is
// on the critical path of every round, so within a box it goes through a POSIX shared-memory
// segment that every rank maps: a rank publishes its contribution in its own slot and bumps a
// sequence number; every rank then reads all slots.  No system call, no copy engine, no stream
// operation per round (~1 us).  Across processes that do not share a box the caller supplies its own
// sb200_comm_vt (NCCL / torch.distributed, see otti_b200/comm.py).
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>

#include "../../../include/spartan_b200.h"
#include "pool.hpp"

namespace sbh {

struct ShmComm {
  static constexpr size_t kMaxBytes = 256 << 10;  // per contribution (the commitment rows of one rank: <= 4096*32/2)
  static constexpr int kMaxWorld = 8;
  struct alignas(64) Slot {
    std::atomic<uint64_t> seq;
    uint8_t pad[56];
    uint8_t data[2][kMaxBytes];  // double buffered by sequence parity: a rank can be at most one exchange ahead
  };
  struct Header {
    std::atomic<uint64_t> hello[kMaxWorld];  // rank r writes a fresh random value here ...
    std::atomic<uint64_t> echo[kMaxWorld];   // ... and rank 0 of THIS job answers with the same value
  };
  std::string name;
  int rank = 0, world = 1;
  uint8_t* base = nullptr;
  size_t bytes = 0;
  uint64_t seq = 0;
  Header* hdr() const { return reinterpret_cast<Header*>(base); }
  Slot* slot(int q) const { return reinterpret_cast<Slot*>(base + sizeof(Header)) + q; }

  static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
  static uint64_t fresh_random() {
    uint64_t v = 0;
    int fd = ::open("/dev/urandom", O_RDONLY);
    if (fd >= 0) { if (::read(fd, &v, 8) != 8) v = 0; ::close(fd); }
    if (!v) v = (uint64_t)(now() * 1e9) ^ ((uint64_t)getpid() << 32) ^ 0x9e3779b97f4a7c15ull;
    return v | 1;
  }
  void unmap() { if (base) munmap(base, bytes); base = nullptr; }

  // Collective: every rank calls with the same name.  A segment of that name left behind by a crashed job must not be
  // mistaken for this job's: rank 0 unlinks whatever exists and creates the object anew (O_EXCL); the other ranks map
  // what they find under the name, write a fresh random `hello` and wait for rank 0 to echo it -- a stale segment
  // never answers (nobody serves it), so after a short wait they drop it and look the name up again.
  int open(const char* nm, int rank_, int world_) {
    if (world_ < 1 || world_ > kMaxWorld || rank_ < 0 || rank_ >= world_) return -1;
    name = nm;
    rank = rank_;
    world = world_;
    bytes = sizeof(Header) + sizeof(Slot) * (size_t)world;
    const double t0 = now();
    if (rank == 0) {
      shm_unlink(nm);
      int fd = shm_open(nm, O_CREAT | O_EXCL | O_RDWR, 0600);
      if (fd < 0) return -2;
      if (ftruncate(fd, (off_t)bytes) != 0) { ::close(fd); shm_unlink(nm); return -3; }  // a fresh object reads as zeros
      void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
      ::close(fd);
      if (p == MAP_FAILED) { shm_unlink(nm); return -4; }
      base = (uint8_t*)p;
      int greeted = 1;
      bool seen[kMaxWorld] = {true};
      while (greeted < world) {
        for (int q = 1; q < world; q++) {
          uint64_t h = hdr()->hello[q].load(std::memory_order_acquire);
          if (h && !seen[q]) {
            hdr()->echo[q].store(h, std::memory_order_release);
            seen[q] = true;
            greeted++;
          }
        }
        SB_CPU_RELAX();
        if (now() - t0 > 120.0) { unmap(); shm_unlink(nm); return -5; }
      }
      return 0;
    }
    const uint64_t mine = fresh_random();
    for (;;) {
      if (now() - t0 > 120.0) return -5;
      int fd = shm_open(nm, O_RDWR, 0600);
      if (fd < 0) { usleep(500); continue; }  // rank 0 has not created it yet
      struct stat sb;
      if (fstat(fd, &sb) != 0 || (size_t)sb.st_size != bytes) { ::close(fd); usleep(500); continue; }  // not sized yet, or another job's
      void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
      ::close(fd);
      if (p == MAP_FAILED) return -4;
      base = (uint8_t*)p;
      hdr()->hello[rank].store(mine, std::memory_order_release);
      const double t1 = now();
      bool ok = false;
      while (now() - t1 < 0.25) {
        if (hdr()->echo[rank].load(std::memory_order_acquire) == mine) { ok = true; break; }
        SB_CPU_RELAX();
      }
      if (ok) return 0;
      unmap();  // stale segment (or rank 0 very slow: the next look-up finds the same, live, object again)
    }
  }
  void close() {
    if (!base) return;
    // the name can go as soon as everybody has mapped it (open() returned on this rank => all attached)
    if (rank == 0) shm_unlink(name.c_str());
    unmap();
  }
  int allgather(const void* send, size_t n, void* recv) {
    if (!base || n > kMaxBytes) return -1;
    const uint64_t s = ++seq;
    Slot* me = slot(rank);
    memcpy(me->data[s & 1], send, n);
    me->seq.store(s, std::memory_order_release);
    for (int q = 0; q < world; q++) {
      Slot* o = slot(q);
      uint64_t spins = 0;
      double t0 = 0;
      while (o->seq.load(std::memory_order_acquire) < s) {
        SB_CPU_RELAX();
        if ((++spins & 0xfffff) == 0) {
          double t = now();
          if (t0 == 0) t0 = t;
          if (t - t0 > 60.0) return -2;  // a peer died
        }
      }
      memcpy((uint8_t*)recv + (size_t)q * n, o->data[s & 1], n);
    }
    return 0;
  }
};

static int shm_allgather_thunk(void* ctx, const void* send, size_t n, void* recv) {
  return static_cast<ShmComm*>(ctx)->allgather(send, n, recv);
}

}  // namespace sbh
