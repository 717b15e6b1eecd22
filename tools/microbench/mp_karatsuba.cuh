// EXPERIMENT RECORD (not part of the product): one level of subtractive Karatsuba for the 8x8-limb product.
// 48 instead of 64 IMAD.WIDE.U32 per product, ~70 more additions.  Measured on B200 with fpmul_bench.cu
// (-DSB_FP_KARATSUBA): a mixed addition takes 1389 ns per warp per SM sub-partition against 1238 ns with the
// schoolbook product at 4 warps per sub-partition -- the adder pipe and the issue slots become the limit before
// the multiplier pipe is relieved.  Kept so the measurement can be repeated; fp.cuh does not use it.
#pragma once
#include "../../otti_b200/csrc/mp.cuh"
#if defined(__CUDA_ARCH__)
namespace sb {
// ---- one level of (subtractive) Karatsuba: 48 instead of 64 wide multiplications, ~70 more additions ---------
// For the point arithmetic, whose kernels are bound by the multiplier pipe while the adder pipe idles.
// acc[0..4) = (a0, a1) x b: two aligned products
__device__ __forceinline__ void mp_mul_row2(uint32_t* acc, uint32_t a0, uint32_t a1, uint32_t b) {
  asm("mul.lo.u32 %0,%4,%6;\n\tmul.hi.u32 %1,%4,%6;\n\tmul.lo.u32 %2,%5,%6;\n\tmul.hi.u32 %3,%5,%6;"
      : "=r"(acc[0]), "=r"(acc[1]), "=r"(acc[2]), "=r"(acc[3])
      : "r"(a0), "r"(a1), "r"(b));
}
__device__ __forceinline__ void mp_mad_row2_nc(uint32_t* acc, uint32_t a0, uint32_t a1, uint32_t b) {
  asm("mad.lo.cc.u32 %0,%4,%6,%0;\n\tmadc.hi.cc.u32 %1,%4,%6,%1;\n\tmadc.lo.cc.u32 %2,%5,%6,%2;\n\tmadc.hi.u32 %3,%5,%6,%3;"
      : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3])
      : "r"(a0), "r"(a1), "r"(b));
}
__device__ __forceinline__ void mp_mad_row2_cf(uint32_t* acc, uint32_t a0, uint32_t a1, uint32_t b) {
  asm("mad.lo.cc.u32 %0,%5,%7,%0;\n\tmadc.hi.cc.u32 %1,%5,%7,%1;\n\tmadc.lo.cc.u32 %2,%6,%7,%2;\n\tmadc.hi.cc.u32 %3,%6,%7,%3;\n\t"
      "addc.u32 %4,0,0;"
      : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "=r"(acc[4])
      : "r"(a0), "r"(a1), "r"(b));
}
// t[0..8) = a[0..4) * b[0..4), same even/odd accumulator scheme as mp_mul8
__device__ __forceinline__ void mp_mul4(uint32_t* t, const uint32_t* a, const uint32_t* b) {
  uint32_t E[8], O[8];  // O[k] is limb k (O[0] unused)
  mp_mul_row2(&E[0], a[0], a[2], b[0]);
  mp_mul_row2(&O[1], a[1], a[3], b[0]);
  E[4] = 0;
  E[5] = 0;
  mp_mad_row2_nc(&E[2], a[1], a[3], b[1]);
  mp_mad_row2_cf(&O[1], a[0], a[2], b[1]);  // O[5] = carry
  mp_mad_row2_cf(&E[2], a[0], a[2], b[2]);  // E[6] = carry
  O[6] = 0;
  mp_mad_row2_nc(&O[3], a[1], a[3], b[2]);
  E[7] = 0;
  mp_mad_row2_nc(&E[4], a[1], a[3], b[3]);
  mp_mad_row2_cf(&O[3], a[0], a[2], b[3]);  // O[7] = carry
  t[0] = E[0];
  asm("add.cc.u32 %0,%7,%14;\n\taddc.cc.u32 %1,%8,%15;\n\taddc.cc.u32 %2,%9,%16;\n\taddc.cc.u32 %3,%10,%17;\n\t"
      "addc.cc.u32 %4,%11,%18;\n\taddc.cc.u32 %5,%12,%19;\n\taddc.u32 %6,%13,%20;"
      : "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7])
      : "r"(E[1]), "r"(E[2]), "r"(E[3]), "r"(E[4]), "r"(E[5]), "r"(E[6]), "r"(E[7]),
        "r"(O[1]), "r"(O[2]), "r"(O[3]), "r"(O[4]), "r"(O[5]), "r"(O[6]), "r"(O[7]));
}
// d[0..4) = |x - y|, returns 0xffffffff if x < y else 0
__device__ __forceinline__ uint32_t mp_absdiff4(uint32_t* d, const uint32_t* x, const uint32_t* y) {
  uint32_t m, tmp;
  asm("sub.cc.u32 %0,%5,%9;\n\tsubc.cc.u32 %1,%6,%10;\n\tsubc.cc.u32 %2,%7,%11;\n\tsubc.cc.u32 %3,%8,%12;\n\tsubc.u32 %4,0,0;"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(m)
      : "r"(x[0]), "r"(x[1]), "r"(x[2]), "r"(x[3]), "r"(y[0]), "r"(y[1]), "r"(y[2]), "r"(y[3]));
  // two's complement negation when m is all ones: (d ^ m) + 1; the carry-in comes from m + m
  d[0] ^= m; d[1] ^= m; d[2] ^= m; d[3] ^= m;
  asm("add.cc.u32 %4,%5,%5;\n\taddc.cc.u32 %0,%0,0;\n\taddc.cc.u32 %1,%1,0;\n\taddc.cc.u32 %2,%2,0;\n\taddc.u32 %3,%3,0;"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]), "=r"(tmp)
      : "r"(m));
  return m;
}
// t[0..16) = a[0..8) * b[0..8):  a = a0 + a1 W, b = b0 + b1 W (W = 2^128),
//   a*b = a0 b0 + (a0 b0 + a1 b1 - (a0 - a1)(b0 - b1)) W + a1 b1 W^2
__device__ __forceinline__ void mp_mul8_kara(uint32_t* t, const uint32_t* a, const uint32_t* b) {
  uint32_t da[4], db[4], zm[8], s[9], tmp;
  const uint32_t ma = mp_absdiff4(da, a, a + 4), mb = mp_absdiff4(db, b, b + 4);
  mp_mul4(t, a, b);
  mp_mul4(t + 8, a + 4, b + 4);
  mp_mul4(zm, da, db);
  // s = z0 + z2 -+ zm (nine limbs; never negative): n = all ones when (a0 - a1)(b0 - b1) = +zm, which is subtracted
  const uint32_t n = ~(ma ^ mb);
  asm("add.cc.u32 %0,%9,%17;\n\taddc.cc.u32 %1,%10,%18;\n\taddc.cc.u32 %2,%11,%19;\n\taddc.cc.u32 %3,%12,%20;\n\t"
      "addc.cc.u32 %4,%13,%21;\n\taddc.cc.u32 %5,%14,%22;\n\taddc.cc.u32 %6,%15,%23;\n\taddc.cc.u32 %7,%16,%24;\n\taddc.u32 %8,0,0;"
      : "=r"(s[0]), "=r"(s[1]), "=r"(s[2]), "=r"(s[3]), "=r"(s[4]), "=r"(s[5]), "=r"(s[6]), "=r"(s[7]), "=r"(s[8])
      : "r"(t[0]), "r"(t[1]), "r"(t[2]), "r"(t[3]), "r"(t[4]), "r"(t[5]), "r"(t[6]), "r"(t[7]),
        "r"(t[8]), "r"(t[9]), "r"(t[10]), "r"(t[11]), "r"(t[12]), "r"(t[13]), "r"(t[14]), "r"(t[15]));
#pragma unroll
  for (int k = 0; k < 8; k++) zm[k] ^= n;
  asm("add.cc.u32 %9,%18,%18;\n\taddc.cc.u32 %0,%0,%10;\n\taddc.cc.u32 %1,%1,%11;\n\taddc.cc.u32 %2,%2,%12;\n\taddc.cc.u32 %3,%3,%13;\n\t"
      "addc.cc.u32 %4,%4,%14;\n\taddc.cc.u32 %5,%5,%15;\n\taddc.cc.u32 %6,%6,%16;\n\taddc.cc.u32 %7,%7,%17;\n\taddc.u32 %8,%8,%18;"
      : "+r"(s[0]), "+r"(s[1]), "+r"(s[2]), "+r"(s[3]), "+r"(s[4]), "+r"(s[5]), "+r"(s[6]), "+r"(s[7]), "+r"(s[8]), "=r"(tmp)
      : "r"(zm[0]), "r"(zm[1]), "r"(zm[2]), "r"(zm[3]), "r"(zm[4]), "r"(zm[5]), "r"(zm[6]), "r"(zm[7]), "r"(n));
  // t += s W
  asm("add.cc.u32 %0,%0,%12;\n\taddc.cc.u32 %1,%1,%13;\n\taddc.cc.u32 %2,%2,%14;\n\taddc.cc.u32 %3,%3,%15;\n\t"
      "addc.cc.u32 %4,%4,%16;\n\taddc.cc.u32 %5,%5,%17;\n\taddc.cc.u32 %6,%6,%18;\n\taddc.cc.u32 %7,%7,%19;\n\t"
      "addc.cc.u32 %8,%8,%20;\n\taddc.cc.u32 %9,%9,0;\n\taddc.cc.u32 %10,%10,0;\n\taddc.u32 %11,%11,0;"
      : "+r"(t[4]), "+r"(t[5]), "+r"(t[6]), "+r"(t[7]), "+r"(t[8]), "+r"(t[9]), "+r"(t[10]), "+r"(t[11]), "+r"(t[12]), "+r"(t[13]),
        "+r"(t[14]), "+r"(t[15])
      : "r"(s[0]), "r"(s[1]), "r"(s[2]), "r"(s[3]), "r"(s[4]), "r"(s[5]), "r"(s[6]), "r"(s[7]), "r"(s[8]));
}

}  // namespace sb
#endif
