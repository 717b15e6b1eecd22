// Microbenchmark: latency of DEPENDENT field / point operations for a lone warp (the regime of msm_small_kernel's and
// msm_finish_kernel's reduction trees, where one warp per SM sub-partition works through a chain of additions).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o latency_bench latency_bench.cu && ./latency_bench
#include <cstdio>
#include <cuda_runtime.h>
#include "../../otti_b200/csrc/ge.cuh"
using namespace sb;

static __device__ __noinline__ fp_t fp_mul_call2(fp_t a, fp_t b) { return fp_mul(a, b); }

template <int MODE>
__global__ void chain(fp_t* io, int iters, long long* cycles) {
  fp_t x = io[threadIdx.x], y = io[32 + threadIdx.x];
  ge_t p;
  p.X = x; p.Y = y; p.Z = fp_one(); p.T = fp_mul(x, y);
  ge_t q = p;
  niels_t nq;
  nq.yplusx = fp_add(y, x); nq.yminusx = fp_sub(y, x); nq.xy2d = fp_mul(p.T, fp_2d());
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) {
    if (MODE == 0) x = fp_mul(x, y);                 // inlined dependent products
    if (MODE == 1) x = fp_mul_call2(x, y);           // the same through a call
    if (MODE == 2) x = fp_sqr(x);                    // dependent squarings (point compression)
    if (MODE == 3) p = ge_add(p, q);                 // full addition, one thread per addition
    if (MODE == 4) p = ge_madd(p, nq, false);        // mixed addition
    if (MODE == 5) { x = fp_add(x, y); y = fp_sub(y, x); }  // add / sub pair
    if (MODE == 6) {                                 // 32 shuffles of one word each (one ge_t through the warp)
#pragma unroll
      for (int k = 0; k < 8; k++) {
        p.X.v[k] = __shfl_xor_sync(0xffffffffu, p.X.v[k], 1); p.Y.v[k] = __shfl_xor_sync(0xffffffffu, p.Y.v[k], 1);
        p.Z.v[k] = __shfl_xor_sync(0xffffffffu, p.Z.v[k], 1); p.T.v[k] = __shfl_xor_sync(0xffffffffu, p.T.v[k], 1);
      }
    }
  }
  long long t1 = clock64();
  if (MODE >= 3) x = fp_add(fp_add(p.X, p.Y), fp_add(p.Z, p.T));
  io[threadIdx.x] = fp_add(x, y);
  if (threadIdx.x == 0) *cycles = t1 - t0;
}

int main() {
  fp_t* io;
  long long* cyc;
  cudaMalloc(&io, 64 * sizeof(fp_t));
  cudaMalloc(&cyc, 8);
  fp_t h[64];
  for (int i = 0; i < 64; i++) for (int k = 0; k < 8; k++) h[i].v[k] = 0x9e3779b9u * (i * 8 + k + 1);
  const char* names[] = {"fp_mul inlined", "fp_mul call", "fp_sqr", "ge_add (9M)", "ge_madd (7M)", "fp_add+fp_sub", "32 x SHFL"};
  const int iters = 2000;
#define RUN(M)                                                                         \
  {                                                                                    \
    cudaMemcpy(io, h, sizeof h, cudaMemcpyHostToDevice);                               \
    chain<M><<<1, 32>>>(io, iters, cyc);                                               \
    chain<M><<<1, 32>>>(io, iters, cyc);                                               \
    long long c = 0;                                                                   \
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);                                    \
    printf("%-16s %8.1f cycles per dependent step (one warp alone)\n", names[M], (double)c / iters); \
  }
  RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6)
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
