// Instance::new on the device (lib.rs:121-225 builds the three SparseMatPolynomials; this library keeps each matrix as a
// CSR by constraint row and a CSC by z column, both STABLE in the caller's entry order -- SNARK mode reads the position
// inside a segment as the op's rank, kernels.cu k_snark_dense).  The host used to counting-sort 40-byte entries six times
// (0.3 s for the 7.3 M non-zeros of an Otti-shaped 2^20 instance); here the entries are uploaded once per matrix, converted
// to Montgomery form by a kernel, and ordered by a stable radix sort of (key, entry index) pairs; the segment pointers are
// binary searches into the sorted keys.  The sort is cub::DeviceRadixSort (CUDA toolkit) -- ingestion plumbing, not a hot op.
#include <cub/device/device_radix_sort.cuh>

#include "dev.hpp"

namespace sb {

__global__ void sb_iota_kernel(uint32_t* out, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = i;
}
// canonical -> Montgomery with the index of the first scalar >= l (Scalar::from_bytes fails there, lib.rs:175-187)
__global__ void sb_to_mont_first_bad_kernel(const fl_t* __restrict__ in, fl_t* __restrict__ out, uint32_t n, unsigned int* first_bad) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint4* q = reinterpret_cast<const uint4*>(in + i);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    fl_t x;
    x.v[0] = a.x; x.v[1] = a.y; x.v[2] = a.z; x.v[3] = a.w; x.v[4] = b.x; x.v[5] = b.y; x.v[6] = b.z; x.v[7] = b.w;
    if (!fl_is_canonical(x)) atomicMin(first_bad, i);
    const fl_t m = fl_to_mont(x);
    uint4* o = reinterpret_cast<uint4*>(out + i);
    o[0] = make_uint4(m.v[0], m.v[1], m.v[2], m.v[3]);
    o[1] = make_uint4(m.v[4], m.v[5], m.v[6], m.v[7]);
  }
}
// ptr[s] = first position whose sorted key is >= s, s in [0, nseg]
__global__ void sb_segment_ptr_kernel(const uint32_t* __restrict__ keys, uint32_t kept, uint32_t nseg, uint32_t* __restrict__ ptr) {
  for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s <= nseg; s += gridDim.x * blockDim.x) {
    uint32_t lo = 0, hi = kept;
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (__ldg(keys + mid) < s) lo = mid + 1; else hi = mid;
    }
    ptr[s] = lo;
  }
}
__global__ void sb_gather_kernel(const uint32_t* __restrict__ perm, uint32_t kept, const uint32_t* __restrict__ other,
                                 const fl_t* __restrict__ vals, uint32_t* __restrict__ idx, fl_t* __restrict__ val, uint32_t* __restrict__ op) {
  for (uint32_t at = blockIdx.x * blockDim.x + threadIdx.x; at < kept; at += gridDim.x * blockDim.x) {
    const uint32_t j = __ldg(perm + at);
    idx[at] = __ldg(other + j);
    const uint4* q = reinterpret_cast<const uint4*>(vals + j);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    uint4* o = reinterpret_cast<uint4*>(val + at);
    o[0] = a;
    o[1] = b;
    if (op != nullptr) op[at] = j;
  }
}
static inline unsigned sbb_grid(size_t n) {
  size_t g = (n + 255) / 256;
  return (unsigned)(g < 1 ? 1 : (g > 148 * 8 ? 148 * 8 : g));
}

cudaError_t k_to_mont_first_bad(const fl_t* in, fl_t* out, uint32_t n, unsigned int* first_bad, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(first_bad, 0xff, sizeof(unsigned int), s);
  if (e != cudaSuccess || n == 0) return e;
  sb_to_mont_first_bad_kernel<<<sbb_grid(n), 256, 0, s>>>(in, out, n, first_bad);
  return cudaGetLastError();
}

size_t sparse_build_tmp_bytes(uint32_t nnz) {
  size_t cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (int)nnz, 0, 32, (cudaStream_t)0);
  return ((cub_bytes + 255) & ~(size_t)255) + 3 * (((size_t)nnz * 4 + 255) & ~(size_t)255) + 256;
}
// keys[i]: segment of entry i on this rank, or 0xFFFFFFFF if the entry belongs to another rank; kept = number of entries
// with a real key (they sort to the front).  key_bits: bits of the largest real key.
cudaError_t k_sparse_build(const uint32_t* keys, const uint32_t* other, const fl_t* vals_mont, uint32_t nnz, uint32_t kept, uint32_t nseg,
                           uint32_t key_bits, uint32_t* ptr, uint32_t* idx, fl_t* val, uint32_t* op, void* tmp, size_t tmp_bytes, cudaStream_t s) {
  if (nnz == 0 || kept == 0) {
    sb_segment_ptr_kernel<<<sbb_grid((size_t)nseg + 1), 256, 0, s>>>(keys, 0, nseg, ptr);
    return cudaGetLastError();
  }
  size_t cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (int)nnz, 0, 32, s);
  const size_t a = (cub_bytes + 255) & ~(size_t)255, w = ((size_t)nnz * 4 + 255) & ~(size_t)255;
  if (a + 3 * w > tmp_bytes) return cudaErrorInvalidValue;
  uint8_t* base = static_cast<uint8_t*>(tmp);
  uint32_t* iota = reinterpret_cast<uint32_t*>(base + a);
  uint32_t* keys_out = reinterpret_cast<uint32_t*>(base + a + w);
  uint32_t* perm = reinterpret_cast<uint32_t*>(base + a + 2 * w);
  sb_iota_kernel<<<sbb_grid(nnz), 256, 0, s>>>(iota, nnz);
  const int end_bit = kept == nnz ? (int)(key_bits < 1 ? 1 : key_bits) : 32;  // entries of other ranks carry 0xFFFFFFFF
  cudaError_t e = cub::DeviceRadixSort::SortPairs(base, cub_bytes, keys, keys_out, iota, perm, (int)nnz, 0, end_bit, s);
  if (e != cudaSuccess) return e;
  sb_segment_ptr_kernel<<<sbb_grid((size_t)nseg + 1), 256, 0, s>>>(keys_out, kept, nseg, ptr);
  sb_gather_kernel<<<sbb_grid(kept), 256, 0, s>>>(perm, kept, other, vals_mont, idx, val, op);
  return cudaGetLastError();
}

}  // namespace sb
