// 256-bit multi-precision building blocks for sm_100a, written as PTX carry chains.
// ptxas fuses each `mad.lo.cc / madc.hi.cc` pair into one IMAD.WIDE.U32(.X) with a carry
// predicate (checked with cuobjdump), so a row of four 32x32->64 products that land on
// aligned limb pairs costs four instructions on the fma pipe plus one carry add.
//
// The 8x8-limb product keeps two accumulators: E takes the partial products whose limb
// position i+j is even, O the odd ones.  Inside one row the four products of a parity
// land on disjoint aligned pairs, so a row is ONE carry chain; across rows each
// accumulator is an ordinary integer that never exceeds a * (b mod 2^(32(i+1))), which is
// what proves the "no carry out of the top limb" cases below.
#pragma once
#include <stdint.h>

#if defined(__CUDA_ARCH__)
namespace sb {

// acc[0..8): four aligned products (a0,a1,a2,a3) x b
__device__ __forceinline__ void mp_mul_row4(uint32_t* acc, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
  asm("mul.lo.u32 %0,%8,%12;\n\tmul.hi.u32 %1,%8,%12;\n\tmul.lo.u32 %2,%9,%12;\n\tmul.hi.u32 %3,%9,%12;\n\t"
      "mul.lo.u32 %4,%10,%12;\n\tmul.hi.u32 %5,%10,%12;\n\tmul.lo.u32 %6,%11,%12;\n\tmul.hi.u32 %7,%11,%12;"
      : "=r"(acc[0]), "=r"(acc[1]), "=r"(acc[2]), "=r"(acc[3]), "=r"(acc[4]), "=r"(acc[5]), "=r"(acc[6]), "=r"(acc[7])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}
// acc[0..8) += (a0,a1,a2,a3) x b, the carry out of limb 7 is known to be zero
__device__ __forceinline__ void mp_mad_row4_nc(uint32_t* acc, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
  asm("mad.lo.cc.u32 %0,%8,%12,%0;\n\tmadc.hi.cc.u32 %1,%8,%12,%1;\n\tmadc.lo.cc.u32 %2,%9,%12,%2;\n\tmadc.hi.cc.u32 %3,%9,%12,%3;\n\t"
      "madc.lo.cc.u32 %4,%10,%12,%4;\n\tmadc.hi.cc.u32 %5,%10,%12,%5;\n\tmadc.lo.cc.u32 %6,%11,%12,%6;\n\tmadc.hi.u32 %7,%11,%12,%7;"
      : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]), "+r"(acc[7])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}
// acc[0..8) += (a0,a1,a2,a3) x b, acc[8] = carry out (acc[8] was untouched so far)
__device__ __forceinline__ void mp_mad_row4_cf(uint32_t* acc, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
  asm("mad.lo.cc.u32 %0,%9,%13,%0;\n\tmadc.hi.cc.u32 %1,%9,%13,%1;\n\tmadc.lo.cc.u32 %2,%10,%13,%2;\n\tmadc.hi.cc.u32 %3,%10,%13,%3;\n\t"
      "madc.lo.cc.u32 %4,%11,%13,%4;\n\tmadc.hi.cc.u32 %5,%11,%13,%5;\n\tmadc.lo.cc.u32 %6,%12,%13,%6;\n\tmadc.hi.cc.u32 %7,%12,%13,%7;\n\t"
      "addc.u32 %8,0,0;"
      : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]), "+r"(acc[7]), "=r"(acc[8])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}

// t[0..16) = a[0..8) * b[0..8)
__device__ __forceinline__ void mp_mul8(uint32_t* t, const uint32_t* a, const uint32_t* b) {
  uint32_t E[16], O[16];  // O[k] is limb k (O[0] unused)
  mp_mul_row4(&E[0], a[0], a[2], a[4], a[6], b[0]);
  mp_mul_row4(&O[1], a[1], a[3], a[5], a[7], b[0]);
  E[8] = 0;
#pragma unroll
  for (int i = 1; i < 8; i += 2) {
    // odd row i: even positions are a_odd x b[i] at limbs i+1.., odd positions a_even x b[i] at limbs i..
    E[i + 8] = 0;
    mp_mad_row4_nc(&E[i + 1], a[1], a[3], a[5], a[7], b[i]);
    mp_mad_row4_cf(&O[i], a[0], a[2], a[4], a[6], b[i]);
    if (i + 1 < 8) {
      // even row i+1: even positions a_even x b at limbs i+1.., odd positions a_odd x b at limbs i+2..
      mp_mad_row4_cf(&E[i + 1], a[0], a[2], a[4], a[6], b[i + 1]);
      O[i + 9] = 0;
      mp_mad_row4_nc(&O[i + 2], a[1], a[3], a[5], a[7], b[i + 1]);
    }
  }
  // t = E + (O << 32)
  t[0] = E[0];
  uint32_t c;
  asm("add.cc.u32 %0,%9,%17;\n\taddc.cc.u32 %1,%10,%18;\n\taddc.cc.u32 %2,%11,%19;\n\taddc.cc.u32 %3,%12,%20;\n\t"
      "addc.cc.u32 %4,%13,%21;\n\taddc.cc.u32 %5,%14,%22;\n\taddc.cc.u32 %6,%15,%23;\n\taddc.cc.u32 %7,%16,%24;\n\taddc.u32 %8,0,0;"
      : "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7]), "=r"(t[8]), "=r"(c)
      : "r"(E[1]), "r"(E[2]), "r"(E[3]), "r"(E[4]), "r"(E[5]), "r"(E[6]), "r"(E[7]), "r"(E[8]),
        "r"(O[1]), "r"(O[2]), "r"(O[3]), "r"(O[4]), "r"(O[5]), "r"(O[6]), "r"(O[7]), "r"(O[8]));
  asm("add.cc.u32 %7,%7,0xffffffff;\n\taddc.cc.u32 %0,%8,%15;\n\taddc.cc.u32 %1,%9,%16;\n\taddc.cc.u32 %2,%10,%17;\n\t"
      "addc.cc.u32 %3,%11,%18;\n\taddc.cc.u32 %4,%12,%19;\n\taddc.cc.u32 %5,%13,%20;\n\taddc.u32 %6,%14,%21;"
      : "=r"(t[9]), "=r"(t[10]), "=r"(t[11]), "=r"(t[12]), "=r"(t[13]), "=r"(t[14]), "=r"(t[15]), "+r"(c)
      : "r"(E[9]), "r"(E[10]), "r"(E[11]), "r"(E[12]), "r"(E[13]), "r"(E[14]), "r"(E[15]),
        "r"(O[9]), "r"(O[10]), "r"(O[11]), "r"(O[12]), "r"(O[13]), "r"(O[14]), "r"(O[15]));
}

}  // namespace sb
#endif
