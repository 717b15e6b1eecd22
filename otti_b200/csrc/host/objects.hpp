// Host-side objects behind the opaque C-ABI handles: the R1CS instance (device
// CSR + CSC of A, B, C), the NIZK generators (host fixed-base tables for the
// Fiat-Shamir tail + device window tables for the commitment kernel) and the
// per-instance device workspace.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../dev.hpp"
#include "hge.hpp"
#include "hmerlin.hpp"

namespace sbh {

// thread-local error text for sb200_last_error()
void set_error(const std::string& s);
const char* get_error();

#define SB_CUDA(expr)                                                                   \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      sbh::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));               \
      return SB200_ERR_CUDA;                                                            \
    }                                                                                   \
  } while (0)

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  cudaError_t alloc(size_t count) {
    release();
    n = count;
    if (!count) return cudaSuccess;
    return cudaMalloc((void**)&p, count * sizeof(T));
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
};

struct SparseDev {  // CSR (by row) or CSC (by column) of the three matrices
  DevBuf<uint32_t> ptr[3], idx[3];
  DevBuf<sb::fl_t> val[3];
  // long segments (dev.hpp LongSegs): lists built once by upload_sparse, scratch reused by every call on the instance
  DevBuf<uint32_t> lg_seg[3], lg_first[3], lg_lo[3], lg_hi[3], lg_big[3];
  DevBuf<sb::fl_t> lg_partial[3], lg_reduced[3];
  uint32_t nlong[3] = {0, 0, 0}, nchunks[3] = {0, 0, 0}, nbig[3] = {0, 0, 0};
  sb::Csr view() const {
    sb::Csr c;
    for (int m = 0; m < 3; m++) {
      c.ptr[m] = ptr[m].p; c.idx[m] = idx[m].p; c.val[m] = val[m].p;
      sb::LongSegs& l = c.lg[m];
      l.seg = lg_seg[m].p; l.first = lg_first[m].p; l.chunk_lo = lg_lo[m].p; l.chunk_hi = lg_hi[m].p; l.big = lg_big[m].p;
      l.partial = lg_partial[m].p; l.reduced = lg_reduced[m].p;
      l.nlong = nlong[m]; l.nchunks = nchunks[m]; l.nbig = nbig[m];
    }
    return c;
  }
};

struct Workspace {  // sized by (num_cons, num_vars); allocated on first prove
  bool ready = false;
  uint64_t resident_vars = ~0ull;  // witness length placed by sb200_witness_upload
  DevBuf<sb::fl_t> vars_canon, varsM, z, zloc, abc, T, A, B, C, eq_tmp, sc_partials, small, blinds, bound_tmp, LZ, bl_a, bl_b,
      coeff0, coeff1, srows;
  DevBuf<unsigned int> counter;
  DevBuf<uint32_t> bell_relay;  // 16 words: flag + challenge passed on by block 0 of a kernel waiting for a doorbell (dev.hpp Bell)
  DevBuf<uint32_t> flag;
  DevBuf<uint8_t> msm_scratch, pts;
  uint8_t* pinned = nullptr;  // host staging
  uint8_t* blind_stage = nullptr;  // pinned staging of the commitment blinds (canonical | Montgomery)
  size_t pinned_bytes = 0;
  uint8_t* mail = nullptr;      // mapped pinned mailbox the kernels write round results + flags into (prover.inc Mail)
  uint8_t* mail_dev = nullptr;  // the same memory as seen from the device
  uint32_t seq = 0;             // mailbox sequence number, never reused
  cudaStream_t stream = nullptr;
  int device = 0;  // the CUDA device of this workspace (helper threads must select it before launching)
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};  // [0,1] commit MSM kernel, [2,3] first cubic round
  // pipelined witness upload (prover.inc): pieces are copied on copy_stream and committed on msm_stream[0/1]
  static constexpr int kMaxPieces = 8;
  cudaStream_t copy_stream = nullptr, msm_stream[2] = {nullptr, nullptr};
  cudaEvent_t ev_piece[kMaxPieces] = {}, ev_fork = nullptr, ev_join[2] = {nullptr, nullptr};
  // sharded proofs: the peers' vars_canon buffers mapped through CUDA IPC (prover.inc p2p_setup)
  bool p2p_tried = false, p2p_ok = false;
  std::vector<sb::fl_t*> peer_vars;
  void close_peers() {
    for (sb::fl_t*& p : peer_vars) {
      if (p) cudaIpcCloseMemHandle(p);
      p = nullptr;
    }
  }
  ~Workspace() {
    close_peers();
    for (int i = 0; i < 4; i++) if (ev[i]) cudaEventDestroy(ev[i]);
    for (int i = 0; i < kMaxPieces; i++) if (ev_piece[i]) cudaEventDestroy(ev_piece[i]);
    if (ev_fork) cudaEventDestroy(ev_fork);
    for (int i = 0; i < 2; i++) {
      if (ev_join[i]) cudaEventDestroy(ev_join[i]);
      if (msm_stream[i]) cudaStreamDestroy(msm_stream[i]);
    }
    if (copy_stream) cudaStreamDestroy(copy_stream);
    if (pinned) cudaFreeHost(pinned);
    if (blind_stage) cudaFreeHost(blind_stage);
    if (mail) cudaFreeHost(mail);
    if (stream) cudaStreamDestroy(stream);
  }
};

struct Instance {
  uint64_t num_cons = 0, num_vars = 0, num_inputs = 0;  // padded, lib.rs:130-158
  uint64_t nnz[3] = {0, 0, 0};
  // One proof sharded over world = 2^gamma GPUs: this rank holds the CSR of the constraint
  // rows i = rank (mod world) (local row i >> gamma, global column indices) and the CSC of the z columns
  // j = rank (mod world) (local column j >> gamma, global row indices).  world == 1: the whole matrices.
  int rank = 0, world = 1;
  uint32_t gamma = 0;
  SparseDev csr, csc;
  // SNARK::encode needs the ops in insertion order: position in the (stable) CSR / CSC -> index of the entry in the
  // caller's list (world == 1 only).  pad_entries: lib.rs:191-197 appended zero-valued entries, which are not kept.
  DevBuf<uint32_t> op_csr[3], op_csc[3];
  bool pad_entries = false;
  mutable Workspace ws;
  // NIZK::verify / is_sat / prove take `&Instance` in the reference and may be called from several threads; here they
  // share the workspace, the stream and the mailbox of the handle, so every entry point that touches them locks this.
  mutable std::mutex mu;
};

struct Gens {
  uint64_t num_vars_padded = 0;
  uint32_t R = 0;  // gens_n.n = 2^(ell - ell/2)
  // host points: G[0..R-1], h, Q = gens_1.G[0]; sum-check generators G4[0..3], h4 (h3 == G4[3])
  std::vector<Ge> G;
  Ge h, Q, G4[4], h4;
  // fixed-base tables for the 7 generators of the Fiat-Shamir tail
  FixedBase fb_G4[4], fb_h4, fb_Q, fb_h;
  std::unique_ptr<uint8_t[]> wide_store[7];  // host copies of the device table slices behind the fb_* above (FixedBase::wide)
  sb::MsmTable table;  // device: bases [0,R) = G, R = h, R+1 = Q
  ~Gens() { sb::msm_table_free(table); }
};

}  // namespace sbh
