// Microbenchmark: cycles per Edwards mixed addition for one warp / several warps per SM sub-partition,
// with the sequential ge_madd and with the row-interleaved multi-multiplication variant.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../otti_b200/csrc/ge.cuh"
using namespace sb;

template <int N>
__device__ __forceinline__ void mp_mul8xN(uint32_t (&t)[N][16], const fp_t* a, const fp_t* b) {
  uint32_t E[N][16], O[N][16];
#pragma unroll
  for (int n = 0; n < N; n++) {
    mp_mul_row4(&E[n][0], a[n].v[0], a[n].v[2], a[n].v[4], a[n].v[6], b[n].v[0]);
    mp_mul_row4(&O[n][1], a[n].v[1], a[n].v[3], a[n].v[5], a[n].v[7], b[n].v[0]);
    E[n][8] = 0;
  }
#pragma unroll
  for (int i = 1; i < 8; i += 2) {
#pragma unroll
    for (int n = 0; n < N; n++) {
      E[n][i + 8] = 0;
      mp_mad_row4_nc(&E[n][i + 1], a[n].v[1], a[n].v[3], a[n].v[5], a[n].v[7], b[n].v[i]);
      mp_mad_row4_cf(&O[n][i], a[n].v[0], a[n].v[2], a[n].v[4], a[n].v[6], b[n].v[i]);
    }
    if (i + 1 < 8) {
#pragma unroll
      for (int n = 0; n < N; n++) {
        mp_mad_row4_cf(&E[n][i + 1], a[n].v[0], a[n].v[2], a[n].v[4], a[n].v[6], b[n].v[i + 1]);
        O[n][i + 9] = 0;
        mp_mad_row4_nc(&O[n][i + 2], a[n].v[1], a[n].v[3], a[n].v[5], a[n].v[7], b[n].v[i + 1]);
      }
    }
  }
#pragma unroll
  for (int n = 0; n < N; n++) {
    t[n][0] = E[n][0];
    uint32_t c;
    asm("add.cc.u32 %0,%9,%17;\n\taddc.cc.u32 %1,%10,%18;\n\taddc.cc.u32 %2,%11,%19;\n\taddc.cc.u32 %3,%12,%20;\n\t"
        "addc.cc.u32 %4,%13,%21;\n\taddc.cc.u32 %5,%14,%22;\n\taddc.cc.u32 %6,%15,%23;\n\taddc.cc.u32 %7,%16,%24;\n\taddc.u32 %8,0,0;"
        : "=r"(t[n][1]), "=r"(t[n][2]), "=r"(t[n][3]), "=r"(t[n][4]), "=r"(t[n][5]), "=r"(t[n][6]), "=r"(t[n][7]), "=r"(t[n][8]), "=r"(c)
        : "r"(E[n][1]), "r"(E[n][2]), "r"(E[n][3]), "r"(E[n][4]), "r"(E[n][5]), "r"(E[n][6]), "r"(E[n][7]), "r"(E[n][8]),
          "r"(O[n][1]), "r"(O[n][2]), "r"(O[n][3]), "r"(O[n][4]), "r"(O[n][5]), "r"(O[n][6]), "r"(O[n][7]), "r"(O[n][8]));
    asm("add.cc.u32 %7,%7,0xffffffff;\n\taddc.cc.u32 %0,%8,%15;\n\taddc.cc.u32 %1,%9,%16;\n\taddc.cc.u32 %2,%10,%17;\n\t"
        "addc.cc.u32 %3,%11,%18;\n\taddc.cc.u32 %4,%12,%19;\n\taddc.cc.u32 %5,%13,%20;\n\taddc.u32 %6,%14,%21;"
        : "=r"(t[n][9]), "=r"(t[n][10]), "=r"(t[n][11]), "=r"(t[n][12]), "=r"(t[n][13]), "=r"(t[n][14]), "=r"(t[n][15]), "+r"(c)
        : "r"(E[n][9]), "r"(E[n][10]), "r"(E[n][11]), "r"(E[n][12]), "r"(E[n][13]), "r"(E[n][14]), "r"(E[n][15]),
          "r"(O[n][9]), "r"(O[n][10]), "r"(O[n][11]), "r"(O[n][12]), "r"(O[n][13]), "r"(O[n][14]), "r"(O[n][15]));
  }
}
template <int N>
__device__ __forceinline__ void fp_mulN(fp_t* r, const fp_t* a, const fp_t* b) {
  uint32_t t[N][16];
  mp_mul8xN<N>(t, a, b);
#pragma unroll
  for (int n = 0; n < N; n++) r[n] = fp_reduce16_dev(t[n]);
}
__device__ __forceinline__ ge_t ge_madd_x(const ge_t& p, const niels_t& q) {
  fp_t a[4], b[4], r[4];
  a[0] = fp_sub(p.Y, p.X); b[0] = q.yminusx;
  a[1] = fp_add(p.Y, p.X); b[1] = q.yplusx;
  a[2] = p.T; b[2] = q.xy2d;
  fp_mulN<3>(r, a, b);
  fp_t D = fp_add(p.Z, p.Z);
  fp_t E = fp_sub(r[1], r[0]), H = fp_add(r[1], r[0]), F = fp_sub(D, r[2]), G = fp_add(D, r[2]);
  a[0] = E; b[0] = F; a[1] = G; b[1] = H; a[2] = F; b[2] = G; a[3] = E; b[3] = H;
  fp_mulN<4>(r, a, b);
  ge_t o;
  o.X = r[0]; o.Y = r[1]; o.Z = r[2]; o.T = r[3];
  return o;
}
template <int MODE>
__global__ void __launch_bounds__(128) bench(const niels_t* tbl, ge_t* out, long long* cyc, int iters) {
  ge_t acc = ge_identity();
  niels_t e = tbl[threadIdx.x & 7];
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) {
    if (MODE == 0) acc = ge_madd(acc, e, false);
    else if (MODE == 1) acc = ge_madd_x(acc, e);
    else { fp_t x = fp_mul(acc.X, e.yplusx); acc.X = x; }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  niels_t h[8];
  for (int i = 0; i < 8; i++) { h[i].yplusx = fp_set(3 + i, 5, 7, 9, 11, 13, 15, 1); h[i].yminusx = fp_set(2, 4 + i, 6, 8, 10, 12, 14, 3); h[i].xy2d = fp_set(1, 2, 3 + i, 4, 5, 6, 7, 5); }
  niels_t* d; ge_t* o; long long* c;
  cudaMalloc(&d, sizeof h); cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice);
  cudaMalloc(&o, sizeof(ge_t) * 148 * 16 * 128); cudaMalloc(&c, 8);
  const int iters = 2000;
  ge_t ref[2];
  for (int mode = 0; mode < 3; mode++)
    for (int blocks_per_sm = 0; blocks_per_sm <= 4; blocks_per_sm++) {
      int grid = blocks_per_sm == 0 ? 1 : 148 * blocks_per_sm, block = blocks_per_sm == 0 ? 32 : 128;
      cudaEvent_t ev0, ev1;
      cudaEventCreate(&ev0); cudaEventCreate(&ev1);
      for (int rep = 0; rep < 2; rep++) {
        if (rep == 1) cudaEventRecord(ev0);
        if (mode == 0) bench<0><<<grid, block>>>(d, o, c, iters);
        else if (mode == 1) bench<1><<<grid, block>>>(d, o, c, iters);
        else bench<2><<<grid, block>>>(d, o, c, iters);
        if (rep == 1) cudaEventRecord(ev1);
      }
      cudaError_t e = cudaDeviceSynchronize();
      float kms = 0; cudaEventElapsedTime(&kms, ev0, ev1);
      // whole-kernel figure: the clock64 figure below is block 0 alone, which overstates throughput when not every block is resident
      if (blocks_per_sm > 0) printf("  kernel %.3f ms -> %.0f ns per warp-op per SMSP\n", kms, kms * 1e6 / ((double)iters * blocks_per_sm));
      long long cyc; cudaMemcpy(&cyc, c, 8, cudaMemcpyDeviceToHost);
      if (blocks_per_sm == 0 && mode < 2) cudaMemcpy(&ref[mode], o, sizeof(ge_t), cudaMemcpyDeviceToHost);
      printf("mode %d (%s) warps/SMSP %s: %.1f cycles per op  [%s]\n", mode, mode == 0 ? "ge_madd" : mode == 1 ? "ge_madd interleaved" : "fp_mul chain",
             blocks_per_sm == 0 ? "1 warp alone" : (blocks_per_sm == 1 ? "1" : blocks_per_sm == 2 ? "2" : blocks_per_sm == 3 ? "3" : "4"), (double)cyc / iters, cudaGetErrorString(e));
    }
  printf("results equal: %d\n", memcmp(&ref[0], &ref[1], sizeof(ge_t)) == 0);
  { unsigned long long h = 1469598103934665603ull; const unsigned char* q = (const unsigned char*)&ref[0]; for (size_t i = 0; i < sizeof(ge_t); i++) h = (h ^ q[i]) * 1099511628211ull; printf("checksum %016llx\n", h); }
  return 0;
}
