// sb200_poly_bound_noalloc: DensePolynomial::bound (dense_mlpoly.rs:205-212) for callers that cannot allocate -- code under
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It is sb200_k_poly_bound's two-pass kernel
// (kernels.cu bound_partial_kernel / bound_sum_kernel) with the per-chunk partial sums in a file-scope device array of
// fixed size instead of a buffer allocated per call; inputs and output are the CALLER's device buffers.  The entry point
// calls no cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine, and does not synchronise.
//
// The scratch array is shared by every call of the process on a device and keeps the last call's contents.  Nothing in it
// has to start at zero: pass one writes every partial that pass two reads.  Two calls that overlap on the GPU (different
// streams, or one stream from two threads without ordering) race on it -- callers serialise.
//
// No counterpart in the reference.
#include <mutex>
#include <string>

#include "../../include/spartan_b200.h"
#include "../../include/spartan_b200_kernels.h"
#include "dev.hpp"

namespace sbh {
void set_error(const std::string& s);  // spartan_b200.cu (thread-local text behind sb200_last_error)
}

namespace sb {

// limit: at most kNoAllocMaxCols columns (n / Ls); the rows are cut into at most kNoAllocMaxChunks chunks
static constexpr uint32_t kNoAllocMaxCols = 4096;
static constexpr uint32_t kNoAllocMaxChunks = 64;
static constexpr size_t kNoAllocScratchBytes = (size_t)kNoAllocMaxChunks * kNoAllocMaxCols * sizeof(fl_t);  // 8 MiB per device
// In host code this name is a host-side stand-in, not a device address: the launches below receive the pointer that
// cudaGetSymbolAddress returned on the current device (noalloc_scratch()).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

__device__ __forceinline__ fl_t na_ld(const fl_t* p) {
  const uint4* q = reinterpret_cast<const uint4*>(p);
  const uint4 a = q[0], b = q[1];
  fl_t x;
  x.v[0] = a.x; x.v[1] = a.y; x.v[2] = a.z; x.v[3] = a.w; x.v[4] = b.x; x.v[5] = b.y; x.v[6] = b.z; x.v[7] = b.w;
  return x;
}
__device__ __forceinline__ void na_st(fl_t* p, const fl_t& x) {
  uint4* q = reinterpret_cast<uint4*>(p);
  q[0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]);
  q[1] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
}
// grid (ceil(R/128), chunks): scratch[jc*R + i] = sum_{j in chunk jc} Lv[j] * Z[j*R + i]
__global__ void na_bound_partial_kernel(const fl_t* __restrict__ Z, const fl_t* __restrict__ Lv, uint32_t Lsize, uint32_t Rsize,
                                        uint32_t jper, fl_t* __restrict__ scratch) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Rsize) return;
  const uint32_t j0 = blockIdx.y * jper, j1 = min(Lsize, j0 + jper);
  fl_t acc = fl_zero();
  for (uint32_t j = j0; j < j1; j++) acc = fl_add(acc, fl_mul(na_ld(Lv + j), na_ld(Z + (size_t)j * Rsize + i)));
  na_st(scratch + (size_t)blockIdx.y * Rsize + i, acc);
}
__global__ void na_bound_sum_kernel(const fl_t* __restrict__ scratch, uint32_t Rsize, uint32_t jc, fl_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Rsize) return;
  fl_t acc = fl_zero();
  for (uint32_t c = 0; c < jc; c++) acc = fl_add(acc, na_ld(scratch + (size_t)c * Rsize + i));
  na_st(out + i, acc);
}

// Device address of the scratch array on the CURRENT device: looked up once per device (the symbol has one instance per
// device that loads this library), with that device current.
static constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static fl_t* g_noalloc_ptr[kNoAllocMaxDevices];
static cudaError_t noalloc_scratch(fl_t** out, int* dev_out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_ptr[dev] = static_cast<fl_t*>(p);
  }
  *out = g_noalloc_ptr[dev];
  *dev_out = dev;
  return cudaSuccess;
}

}  // namespace sb

extern "C" int sb200_poly_bound_noalloc(const void* d_Z, size_t n, const void* d_L, size_t Ls, void* d_out, void* stream) {
  if (d_Z == nullptr || d_L == nullptr || d_out == nullptr || Ls == 0 || n == 0 || n % Ls != 0 || n >= ((size_t)1 << 32)) {
    sbh::set_error("sb200_poly_bound_noalloc: null pointer, or n is not a positive multiple of Ls below 2^32");
    return SB200_ERR_ARG;
  }
  const size_t Rs = n / Ls;
  if (Rs > sb::kNoAllocMaxCols) {
    sbh::set_error("sb200_poly_bound_noalloc: n / Ls = " + std::to_string(Rs) + " columns exceed the fixed scratch (at most " +
                   std::to_string(sb::kNoAllocMaxCols) + "); use sb200_k_poly_bound");
    return SB200_ERR_ARG;
  }
  if ((((uintptr_t)d_Z) | ((uintptr_t)d_L) | ((uintptr_t)d_out)) & 15) {
    sbh::set_error("sb200_poly_bound_noalloc: device buffers must be 16-byte aligned");
    return SB200_ERR_ARG;
  }
  sb::fl_t* scratch = nullptr;
  int dev = -1;
  cudaError_t e = sb::noalloc_scratch(&scratch, &dev);
  if (e != cudaSuccess) {
    sbh::set_error(std::string("sb200_poly_bound_noalloc: cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e));
    return e == cudaErrorNoDevice || e == cudaErrorInvalidDevice ? SB200_ERR_NO_DEVICE : SB200_ERR_CUDA;
  }
  // chunks of rows: enough blocks to fill the SMs, never more than the scratch holds
  const uint32_t L32 = (uint32_t)Ls, R32 = (uint32_t)Rs, xblocks = (R32 + 127) / 128;
  uint32_t jc = (8 * 148 + xblocks - 1) / xblocks;
  if (jc > sb::kNoAllocMaxChunks) jc = sb::kNoAllocMaxChunks;
  if (jc > L32) jc = L32;
  const uint32_t jper = (L32 + jc - 1) / jc;
  jc = (L32 + jper - 1) / jper;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  sb::na_bound_partial_kernel<<<dim3(xblocks, jc), 128, 0, s>>>(static_cast<const sb::fl_t*>(d_Z), static_cast<const sb::fl_t*>(d_L), L32, R32,
                                                               jper, scratch);
  e = cudaGetLastError();
  if (e == cudaSuccess) {
    sb::na_bound_sum_kernel<<<xblocks, 128, 0, s>>>(scratch, R32, jc, static_cast<sb::fl_t*>(d_out));
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    sbh::set_error(std::string("sb200_poly_bound_noalloc: launch: ") + cudaGetErrorString(e));
    return SB200_ERR_CUDA;
  }
  return SB200_OK;
}
