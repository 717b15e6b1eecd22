// The exchange hook of a sharded proof over NCCL, called natively (no Python in the loop): the per-round all-gather of a
// few dozen host bytes goes host -> device staging -> ncclAllGather over NVLink -> host.  NCCL is loaded with dlopen
// (the copy torch ships, or a system one), so libspartan_b200 has no link-time dependency on it.  This transport exists to be
// MEASURED against the shared-memory hook (host/comm.hpp): for 96-byte payloads a collective's launch and completion
// latency is the whole cost, and the sums are needed by the host transcript anyway.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdint>
#include <cstring>
#include <string>

namespace sbh {

struct NcclApi {
  typedef struct { char internal[128]; } UniqueId;
  typedef void* Comm;
  int (*GetUniqueId)(UniqueId*) = nullptr;
  int (*CommInitRank)(Comm*, int, UniqueId, int) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, Comm, cudaStream_t) = nullptr;
  int (*CommDestroy)(Comm) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  void* handle = nullptr;
  bool load(std::string& err) {
    if (handle) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (handle) break;
    }
    if (!handle) { err = "libnccl.so.2 not found (import torch first, or put NCCL on the loader path)"; return false; }
    GetUniqueId = (int (*)(UniqueId*))dlsym(handle, "ncclGetUniqueId");
    CommInitRank = (int (*)(Comm*, int, UniqueId, int))dlsym(handle, "ncclCommInitRank");
    AllGather = (int (*)(const void*, void*, size_t, int, Comm, cudaStream_t))dlsym(handle, "ncclAllGather");
    CommDestroy = (int (*)(Comm))dlsym(handle, "ncclCommDestroy");
    GetErrorString = (const char* (*)(int))dlsym(handle, "ncclGetErrorString");
    if (!GetUniqueId || !CommInitRank || !AllGather || !CommDestroy) { err = "libnccl lacks the expected symbols"; return false; }
    return true;
  }
  static NcclApi& get() {
    static NcclApi api;
    return api;
  }
};

struct NcclComm {
  static constexpr size_t kMaxBytes = 1 << 20;  // per rank and call
  NcclApi::Comm comm = nullptr;
  cudaStream_t stream = nullptr;
  uint8_t *d_send = nullptr, *d_recv = nullptr, *h_stage = nullptr;
  int rank = 0, world = 1;
  ~NcclComm() {
    if (comm) NcclApi::get().CommDestroy(comm);
    if (d_send) cudaFree(d_send);
    if (d_recv) cudaFree(d_recv);
    if (h_stage) cudaFreeHost(h_stage);
    if (stream) cudaStreamDestroy(stream);
  }
  // every rank contributes `bytes` bytes; recv holds world * bytes in rank order
  int allgather(const void* send, size_t bytes, void* recv) {
    if (bytes > kMaxBytes) return -2;
    memcpy(h_stage, send, bytes);
    if (cudaMemcpyAsync(d_send, h_stage, bytes, cudaMemcpyHostToDevice, stream) != cudaSuccess) return -3;
    if (NcclApi::get().AllGather(d_send, d_recv, bytes, /*ncclChar*/ 0, comm, stream) != 0) return -4;
    if (cudaMemcpyAsync(h_stage + kMaxBytes, d_recv, bytes * (size_t)world, cudaMemcpyDeviceToHost, stream) != cudaSuccess) return -5;
    if (cudaStreamSynchronize(stream) != cudaSuccess) return -6;
    memcpy(recv, h_stage + kMaxBytes, bytes * (size_t)world);
    return 0;
  }
};
static int nccl_allgather_thunk(void* ctx, const void* send, size_t bytes, void* recv) {
  return static_cast<NcclComm*>(ctx)->allgather(send, bytes, recv);
}

}  // namespace sbh
