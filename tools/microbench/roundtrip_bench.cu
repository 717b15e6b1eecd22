// Microbenchmark: the floor of one host <-> device round trip of a Fiat-Shamir round -- cudaLaunchKernel of a one-block
// kernel that writes a flag into mapped pinned host memory, the host spinning on it -- with 0, 4 and 16 dependent F_l
// products in the kernel, and the same through a kernel launched ahead that waits for a host "doorbell".
#include <chrono>
#include <cstdio>
#include <cuda_runtime.h>
#include "../../otti_b200/csrc/fl.cuh"
using namespace sb;

__global__ void flag_kernel(int nmul, fl_t x, fl_t* sink, volatile uint32_t* host_flag, uint32_t seq) {
  fl_t a = x;
  for (int i = 0; i < nmul; i++) a = fl_mul(a, x);
  if (threadIdx.x == 0) {
    if (nmul) *sink = a;
    __threadfence_system();
    *host_flag = seq;
  }
}
__global__ void bell_kernel(int nmul, fl_t x, fl_t* sink, volatile uint32_t* bell, uint32_t want, volatile uint32_t* host_flag, uint32_t seq) {
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    while (*bell != want && clock64() - t0 < 2000000000LL) {}
  }
  __syncthreads();
  fl_t a = x;
  for (int i = 0; i < nmul; i++) a = fl_mul(a, x);
  if (threadIdx.x == 0) {
    if (nmul) *sink = a;
    __threadfence_system();
    *host_flag = seq;
  }
}
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
  uint32_t *flag_h, *flag_d, *bell_h, *bell_d;
  cudaHostAlloc((void**)&flag_h, 64, cudaHostAllocMapped);
  cudaHostAlloc((void**)&bell_h, 64, cudaHostAllocMapped);
  cudaHostGetDevicePointer((void**)&flag_d, flag_h, 0);
  cudaHostGetDevicePointer((void**)&bell_d, bell_h, 0);
  fl_t* sink;
  cudaMalloc(&sink, 32);
  cudaStream_t st;
  cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  fl_t x = fl_one();
  x.v[0] ^= 5;
  *flag_h = 0;
  *bell_h = 0;
  uint32_t seq = 0;
  const int iters = 2000;
  for (int nmul : {0, 4, 16}) {
    for (int threads : {32, 256}) {
      for (int w = 0; w < 100; w++) { flag_kernel<<<1, threads, 0, st>>>(nmul, x, sink, flag_d, ++seq); while (*(volatile uint32_t*)flag_h != seq) {} }
      double t0 = now();
      for (int i = 0; i < iters; i++) {
        flag_kernel<<<1, threads, 0, st>>>(nmul, x, sink, flag_d, ++seq);
        while (*(volatile uint32_t*)flag_h != seq) {}
      }
      printf("launch -> flag, %2d products, %3d threads: %.2f us per round trip\n", nmul, threads, (now() - t0) / iters * 1e6);
    }
  }
  // doorbell: the kernel of round i+1 is launched before round i's result is read; it starts when the host rings
  for (int nmul : {0, 4, 16}) {
    uint32_t bell = 0;
    bell_kernel<<<1, 256, 0, st>>>(nmul, x, sink, bell_d, bell + 1, flag_d, ++seq);
    double t0 = now();
    for (int i = 0; i < iters; i++) {
      const uint32_t this_seq = seq;
      bell_kernel<<<1, 256, 0, st>>>(nmul, x, sink, bell_d, bell + 2, flag_d, ++seq);  // next round, launched ahead
      *(volatile uint32_t*)bell_h = ++bell;                                             // ring: the waiting kernel runs
      while (*(volatile uint32_t*)flag_h != this_seq) {}
    }
    *(volatile uint32_t*)bell_h = ++bell;
    cudaStreamSynchronize(st);
    printf("doorbell -> flag, %2d products, 256 threads: %.2f us per round trip\n", nmul, (now() - t0) / iters * 1e6);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
