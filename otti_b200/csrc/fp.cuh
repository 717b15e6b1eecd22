// F_p device arithmetic, p = 2^255 - 19, for Edwards25519 / ristretto255
// coordinates (the reference gets these from curve25519-dalek 3.2; call sites
// Spartan/src/group.rs:6-7,98-117, commitments.rs:79-84).
//
// Representation: 8 x u32 little-endian limbs holding ANY value in [0, 2^256)
// congruent to the element ("loose" form).  2^256 = 38 (mod p), so carries out
// of bit 256 fold back in as +38 and never need a compare.  fp_canon() gives
// the unique representative in [0, p) for compress / equality / sign.
// Plain C, __host__ __device__: the limb logic is unit-tested on the CPU.
#pragma once
#include "fl.cuh"
#include "mp.cuh"
#ifdef SB_FP_KARATSUBA
#include "../../tools/microbench/mp_karatsuba.cuh"
#endif

namespace sb {

struct alignas(16) fp_t {
  uint32_t v[8];
};

SB_HD fp_t fp_zero() {
  fp_t r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = 0;
  return r;
}
SB_HD fp_t fp_one() {
  fp_t r = fp_zero();
  r.v[0] = 1;
  return r;
}
SB_HD fp_t fp_set(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t a4, uint32_t a5, uint32_t a6, uint32_t a7) {
  fp_t r;
  r.v[0] = a0; r.v[1] = a1; r.v[2] = a2; r.v[3] = a3; r.v[4] = a4; r.v[5] = a5; r.v[6] = a6; r.v[7] = a7;
  return r;
}
// 2d mod p, d = -121665/121666
SB_HD fp_t fp_2d() { return fp_set(0x26b2f159u, 0xebd69b94u, 0x8283b156u, 0x00e0149au, 0xeef3d130u, 0x198e80f2u, 0x56dffce7u, 0x2406d9dcu); }
SB_HD fp_t fp_d() { return fp_set(0x135978a3u, 0x75eb4dcau, 0x4141d8abu, 0x00700a4du, 0x7779e898u, 0x8cc74079u, 0x2b6ffe73u, 0x52036ceeu); }

// r = x + 38*k (k in {0,1,2...small}), folding the (at most one) further carry
SB_HD fp_t fp_fold_small(const fp_t& x, uint32_t k38) {
  fp_t r;
  uint64_t c = k38;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += x.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  // a second wrap can only happen when the low limbs are tiny, so +38 cannot carry past limb 0's neighbours
  uint32_t k2 = (uint32_t)c * 38u;
  c = k2;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += r.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  return r;
}
SB_HD fp_t fp_add_portable(const fp_t& a, const fp_t& b) {
  fp_t t;
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)a.v[i] + b.v[i];
    t.v[i] = (uint32_t)c;
    c >>= 32;
  }
  return fp_fold_small(t, (uint32_t)c * 38u);
}
SB_HD fp_t fp_sub_portable(const fp_t& a, const fp_t& b) {
  fp_t d;
  uint64_t borrow = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t x = (uint64_t)a.v[i] - b.v[i] - borrow;
    d.v[i] = (uint32_t)x;
    borrow = x >> 63;
  }
  // a - b + 2^256 == a - b + 38 (mod p): subtract 38 when we wrapped
  uint64_t k = borrow * 38u;
  borrow = 0;
  fp_t r;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t x = (uint64_t)d.v[i] - (i == 0 ? k : 0) - borrow;
    r.v[i] = (uint32_t)x;
    borrow = x >> 63;
  }
  k = borrow * 38u;  // wrapped again (d was < 38): one more -38 cannot wrap a third time
  borrow = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t x = (uint64_t)r.v[i] - (i == 0 ? k : 0) - borrow;
    r.v[i] = (uint32_t)x;
    borrow = x >> 63;
  }
  return r;
}

// 16-limb product folded to 8 limbs: lo + 38*hi
SB_HD fp_t fp_reduce16(const uint32_t t[16]) {
  uint32_t r[9];
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)t[i + 8] * 38u + t[i];
    r[i] = (uint32_t)c;
    c >>= 32;
  }
  fp_t x;
#pragma unroll
  for (int i = 0; i < 8; i++) x.v[i] = r[i];
  return fp_fold_small(x, (uint32_t)c * 38u);  // c < 39 -> c*38 < 2^11
}
SB_HD fp_t fp_mul_portable(const fp_t& a, const fp_t& b) {
  uint32_t t[16];
#pragma unroll
  for (int i = 0; i < 16; i++) t[i] = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t c = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      c += (uint64_t)a.v[j] * b.v[i] + t[i + j];
      t[i + j] = (uint32_t)c;
      c >>= 32;
    }
    t[i + 8] = (uint32_t)c;
  }
  return fp_reduce16(t);
}
SB_HD fp_t fp_sqr_portable(const fp_t& a) { return fp_mul_portable(a, a); }
#if defined(__CUDA_ARCH__)
// ---- sm_100a versions: PTX carry chains (mp.cuh).  Same loose-form contract as the portable code.
__device__ __forceinline__ fp_t fp_add(const fp_t& a, const fp_t& b) {
  fp_t r;
  uint32_t c;
  asm("add.cc.u32 %0,%9,%17;\n\taddc.cc.u32 %1,%10,%18;\n\taddc.cc.u32 %2,%11,%19;\n\taddc.cc.u32 %3,%12,%20;\n\t"
      "addc.cc.u32 %4,%13,%21;\n\taddc.cc.u32 %5,%14,%22;\n\taddc.cc.u32 %6,%15,%23;\n\taddc.cc.u32 %7,%16,%24;\n\taddc.u32 %8,0,0;"
      : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7]), "=r"(c)
      : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
        "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
  // 2^256 = 38: fold the carry; a second wrap leaves a value < 38, so the last +38 touches limb 0 only
  asm("mad.lo.cc.u32 %0,%8,38,%0;\n\taddc.cc.u32 %1,%1,0;\n\taddc.cc.u32 %2,%2,0;\n\taddc.cc.u32 %3,%3,0;\n\t"
      "addc.cc.u32 %4,%4,0;\n\taddc.cc.u32 %5,%5,0;\n\taddc.cc.u32 %6,%6,0;\n\taddc.cc.u32 %7,%7,0;\n\taddc.u32 %8,0,0;\n\t"
      "mad.lo.u32 %0,%8,38,%0;"
      : "+r"(r.v[0]), "+r"(r.v[1]), "+r"(r.v[2]), "+r"(r.v[3]), "+r"(r.v[4]), "+r"(r.v[5]), "+r"(r.v[6]), "+r"(r.v[7]), "+r"(c));
  return r;
}
__device__ __forceinline__ fp_t fp_sub(const fp_t& a, const fp_t& b) {
  fp_t r;
  uint32_t bw;
  asm("sub.cc.u32 %0,%9,%17;\n\tsubc.cc.u32 %1,%10,%18;\n\tsubc.cc.u32 %2,%11,%19;\n\tsubc.cc.u32 %3,%12,%20;\n\t"
      "subc.cc.u32 %4,%13,%21;\n\tsubc.cc.u32 %5,%14,%22;\n\tsubc.cc.u32 %6,%15,%23;\n\tsubc.cc.u32 %7,%16,%24;\n\tsubc.u32 %8,0,0;"
      : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7]), "=r"(bw)
      : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
        "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
  // wrapped by 2^256 = 38 (mod p): take 38 off again; a second wrap leaves >= 2^256 - 38, so the last -38 touches limb 0 only
  uint32_t k = bw & 38u;
  asm("sub.cc.u32 %0,%0,%8;\n\tsubc.cc.u32 %1,%1,0;\n\tsubc.cc.u32 %2,%2,0;\n\tsubc.cc.u32 %3,%3,0;\n\t"
      "subc.cc.u32 %4,%4,0;\n\tsubc.cc.u32 %5,%5,0;\n\tsubc.cc.u32 %6,%6,0;\n\tsubc.cc.u32 %7,%7,0;\n\tsubc.u32 %8,0,0;"
      : "+r"(r.v[0]), "+r"(r.v[1]), "+r"(r.v[2]), "+r"(r.v[3]), "+r"(r.v[4]), "+r"(r.v[5]), "+r"(r.v[6]), "+r"(r.v[7]), "+r"(k));
  r.v[0] -= k & 38u;
  return r;
}
// t[0..16) -> lo + 38 * hi, folded into 8 limbs
__device__ __forceinline__ fp_t fp_reduce16_dev(const uint32_t* t) {
  uint32_t r0, r1, r2, r3, r4, r5, r6, r7, r8;
  asm("mad.lo.cc.u32 %0,%9,38,%13;\n\tmadc.hi.cc.u32 %1,%9,38,%14;\n\tmadc.lo.cc.u32 %2,%10,38,%15;\n\tmadc.hi.cc.u32 %3,%10,38,%16;\n\t"
      "madc.lo.cc.u32 %4,%11,38,%17;\n\tmadc.hi.cc.u32 %5,%11,38,%18;\n\tmadc.lo.cc.u32 %6,%12,38,%19;\n\tmadc.hi.cc.u32 %7,%12,38,%20;\n\t"
      "addc.u32 %8,0,0;"
      : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7), "=r"(r8)
      : "r"(t[8]), "r"(t[10]), "r"(t[12]), "r"(t[14]), "r"(t[0]), "r"(t[1]), "r"(t[2]), "r"(t[3]), "r"(t[4]), "r"(t[5]), "r"(t[6]), "r"(t[7]));
  asm("mad.lo.cc.u32 %0,%8,38,%0;\n\tmadc.hi.cc.u32 %1,%8,38,%1;\n\tmadc.lo.cc.u32 %2,%9,38,%2;\n\tmadc.hi.cc.u32 %3,%9,38,%3;\n\t"
      "madc.lo.cc.u32 %4,%10,38,%4;\n\tmadc.hi.cc.u32 %5,%10,38,%5;\n\tmadc.lo.cc.u32 %6,%11,38,%6;\n\tmadc.hi.u32 %7,%11,38,%7;"
      : "+r"(r1), "+r"(r2), "+r"(r3), "+r"(r4), "+r"(r5), "+r"(r6), "+r"(r7), "+r"(r8)
      : "r"(t[9]), "r"(t[11]), "r"(t[13]), "r"(t[15]));
  // r8 < 39 + 1: fold it (r8 * 38 < 2^11); a second wrap leaves a tiny value, so the last +38 touches limb 0 only
  asm("mad.lo.cc.u32 %0,%8,38,%0;\n\taddc.cc.u32 %1,%1,0;\n\taddc.cc.u32 %2,%2,0;\n\taddc.cc.u32 %3,%3,0;\n\t"
      "addc.cc.u32 %4,%4,0;\n\taddc.cc.u32 %5,%5,0;\n\taddc.cc.u32 %6,%6,0;\n\taddc.cc.u32 %7,%7,0;\n\taddc.u32 %8,0,0;\n\t"
      "mad.lo.u32 %0,%8,38,%0;"
      : "+r"(r0), "+r"(r1), "+r"(r2), "+r"(r3), "+r"(r4), "+r"(r5), "+r"(r6), "+r"(r7), "+r"(r8));
  fp_t x;
  x.v[0] = r0; x.v[1] = r1; x.v[2] = r2; x.v[3] = r3; x.v[4] = r4; x.v[5] = r5; x.v[6] = r6; x.v[7] = r7;
  return x;
}
__device__ __forceinline__ fp_t fp_mul(const fp_t& a, const fp_t& b) {
  uint32_t t[16];
#ifdef SB_FP_KARATSUBA  // tools/microbench only: measured slower, see tools/microbench/mp_karatsuba.cuh
  mp_mul8_kara(t, a.v, b.v);
#else
  mp_mul8(t, a.v, b.v);
#endif
  return fp_reduce16_dev(t);
}
__device__ __forceinline__ fp_t fp_sqr(const fp_t& a) { return fp_mul(a, a); }
#else
SB_HD fp_t fp_add(const fp_t& a, const fp_t& b) { return fp_add_portable(a, b); }
SB_HD fp_t fp_sub(const fp_t& a, const fp_t& b) { return fp_sub_portable(a, b); }
SB_HD fp_t fp_mul(const fp_t& a, const fp_t& b) { return fp_mul_portable(a, b); }
SB_HD fp_t fp_sqr(const fp_t& a) { return fp_mul_portable(a, a); }
#endif
SB_HD fp_t fp_neg(const fp_t& a) { return fp_sub(fp_zero(), a); }
SB_HD fp_t fp_mul_small(const fp_t& a, uint32_t k) {  // k < 2^26
  fp_t x;
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)a.v[i] * k;
    x.v[i] = (uint32_t)c;
    c >>= 32;
  }
  return fp_fold_small(x, (uint32_t)c * 38u);
}

// unique representative in [0, p)
SB_HD fp_t fp_canon(const fp_t& a) {
  // fold bit 255: x = (a mod 2^255) + 19*(a >> 255)   -> x < 2^255 + 19
  fp_t x = a;
  uint32_t top = x.v[7] >> 31;
  x.v[7] &= 0x7fffffffu;
  uint64_t c = top * 19u;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += x.v[i];
    x.v[i] = (uint32_t)c;
    c >>= 32;
  }
  // now x < 2^255 + 19 <= 2p + ...; subtract p while x >= p (at most twice)
#pragma unroll
  for (int rep = 0; rep < 2; rep++) {
    // x >= p  <=>  x + 19 >= 2^255
    uint64_t cc = 19;
    uint32_t y[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
      cc += x.v[i];
      y[i] = (uint32_t)cc;
      cc >>= 32;
    }
    uint32_t ge = (y[7] >> 31) | (uint32_t)cc;  // bit 255 (or beyond) set
    y[7] &= 0x7fffffffu;                        // x + 19 - 2^255 = x - p
    if (ge) {
#pragma unroll
      for (int i = 0; i < 8; i++) x.v[i] = y[i];
    }
  }
  return x;
}
SB_HD bool fp_is_zero(const fp_t& a) {
  fp_t c = fp_canon(a);
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) o |= c.v[i];
  return o == 0;
}
SB_HD bool fp_eq(const fp_t& a, const fp_t& b) { return fp_is_zero(fp_sub(a, b)); }
SB_HD bool fp_is_neg(const fp_t& a) { return fp_canon(a).v[0] & 1u; }
SB_HD fp_t fp_abs(const fp_t& a) { return fp_is_neg(a) ? fp_neg(a) : a; }

// Field-operation policies: Inl expands every multiplication in place (throughput kernels, the
// loop body is reused); Call goes through one out-of-line copy per kernel, which keeps the code of
// the latency-bound kernels (one pass over a long dependent chain) inside the instruction cache.
struct FpInl {
  static SB_HD fp_t mul(const fp_t& a, const fp_t& b) { return fp_mul(a, b); }
  static SB_HD fp_t sqr(const fp_t& a) { return fp_sqr(a); }
};
#if defined(__CUDACC__)
static __device__ __noinline__ fp_t fp_mul_call(fp_t a, fp_t b) { return fp_mul(a, b); }
struct FpCall {
  static __device__ __forceinline__ fp_t mul(const fp_t& a, const fp_t& b) { return fp_mul_call(a, b); }
  static __device__ __forceinline__ fp_t sqr(const fp_t& a) { return fp_mul_call(a, a); }
};
// products through the call, squarings inlined: the squarings of an inversion chain sit in a handful of loops (small
// code), and a lone warp pays 92 of 559 cycles per dependent product for the call (tools/microbench/latency_bench.cu)
struct FpSqrInl {
  static __device__ __forceinline__ fp_t mul(const fp_t& a, const fp_t& b) { return fp_mul_call(a, b); }
  static __device__ __forceinline__ fp_t sqr(const fp_t& a) { return fp_mul(a, a); }
};
#endif

template <class F>
SB_HD fp_t fp_pow2k_t(fp_t a, int k) {
#pragma unroll 1
  for (int i = 0; i < k; i++) a = F::sqr(a);
  return a;
}
// a^((p-5)/8) = a^(2^252 - 3)
template <class F>
SB_HD fp_t fp_pow_p58_t(const fp_t& z) {
  fp_t z2 = F::sqr(z), z9 = F::mul(fp_pow2k_t<F>(z2, 2), z), z11 = F::mul(z9, z2);
  fp_t z2_5_0 = F::mul(F::sqr(z11), z9);
  fp_t z2_10_0 = F::mul(fp_pow2k_t<F>(z2_5_0, 5), z2_5_0);
  fp_t z2_20_0 = F::mul(fp_pow2k_t<F>(z2_10_0, 10), z2_10_0);
  fp_t z2_40_0 = F::mul(fp_pow2k_t<F>(z2_20_0, 20), z2_20_0);
  fp_t z2_50_0 = F::mul(fp_pow2k_t<F>(z2_40_0, 10), z2_10_0);
  fp_t z2_100_0 = F::mul(fp_pow2k_t<F>(z2_50_0, 50), z2_50_0);
  fp_t z2_200_0 = F::mul(fp_pow2k_t<F>(z2_100_0, 100), z2_100_0);
  fp_t z2_250_0 = F::mul(fp_pow2k_t<F>(z2_200_0, 50), z2_50_0);
  return F::mul(fp_pow2k_t<F>(z2_250_0, 2), z);
}
// a^(p-2)
template <class F>
SB_HD fp_t fp_invert_t(const fp_t& z) {
  // a^(2^255-21) = (a^(2^252-3))^8 * a^3
  fp_t t = fp_pow2k_t<F>(fp_pow_p58_t<F>(z), 3);
  return F::mul(t, F::mul(F::sqr(z), z));
}
SB_HD fp_t fp_pow2k(fp_t a, int k) { return fp_pow2k_t<FpInl>(a, k); }
SB_HD fp_t fp_pow_p58(const fp_t& z) { return fp_pow_p58_t<FpInl>(z); }
SB_HD fp_t fp_invert(const fp_t& z) { return fp_invert_t<FpInl>(z); }

}  // namespace sb
