// F_l device arithmetic for sm_100a: l = 2^252 + 27742317777372353535851937790883648493,
// Montgomery form with R = 2^256 on 8 x u32 limbs (little-endian), always fully
// reduced (< l) so that the 32 bytes in HBM are exactly the reference's
// serialised `Scalar([u64;4])` (Spartan/src/scalar/ristretto255.rs:199-200).
// Same values as Scalar::{add,sub,neg,mul,montgomery_reduce}
// (ristretto255.rs:654-775); the word size differs (32-bit limbs feed the
// IMAD.WIDE.U32 pipe), the result bits do not.
//
// Everything here is __host__ __device__ plain C so the limb logic is unit
// tested on the CPU build box (tests/test_device_math_host.py) as well as on
// the GPU.
#pragma once
#include <stdint.h>
#include "mp.cuh"

#ifdef __CUDACC__
#define SB_HD __host__ __device__ __forceinline__
#else
#define SB_HD inline
#endif

namespace sb {

struct alignas(16) fl_t {
  uint32_t v[8];
};

#define SB_FL_L0 0x5cf5d3edu
#define SB_FL_L1 0x5812631au
#define SB_FL_L2 0xa2f79cd6u
#define SB_FL_L3 0x14def9deu
#define SB_FL_L7 0x10000000u
#define SB_FL_INV32 0x12547e1bu  // -l^{-1} mod 2^32 (low word of INV, ristretto255.rs:303)

SB_HD uint32_t fl_modl(int i) {
  return i == 0 ? SB_FL_L0 : i == 1 ? SB_FL_L1 : i == 2 ? SB_FL_L2 : i == 3 ? SB_FL_L3 : i == 7 ? SB_FL_L7 : 0u;
}
SB_HD fl_t fl_zero() {
  fl_t r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = 0;
  return r;
}
// R = 2^256 mod l (ristretto255.rs:306-311) = Scalar::one()
SB_HD fl_t fl_one() {
  fl_t r;
  r.v[0] = 0x8d98951du; r.v[1] = 0xd6ec3174u; r.v[2] = 0x737dcf70u; r.v[3] = 0xc6ef5bf4u;
  r.v[4] = 0xfffffffeu; r.v[5] = 0xffffffffu; r.v[6] = 0xffffffffu; r.v[7] = 0x0fffffffu;
  return r;
}
// R^2 mod l (ristretto255.rs:314-319)
SB_HD fl_t fl_r2() {
  fl_t r;
  r.v[0] = 0x449c0f01u; r.v[1] = 0xa40611e3u; r.v[2] = 0x68859347u; r.v[3] = 0xd00e1ba7u;
  r.v[4] = 0x17f5be65u; r.v[5] = 0xceec73d2u; r.v[6] = 0x7c309a3du; r.v[7] = 0x0399411bu;
  return r;
}
SB_HD bool fl_is_zero(const fl_t& a) {
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) o |= a.v[i];
  return o == 0;
}
SB_HD bool fl_eq(const fl_t& a, const fl_t& b) {
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) o |= a.v[i] ^ b.v[i];
  return o == 0;
}

// t - l if t >= l else t   (t < 2l < 2^254)
SB_HD fl_t fl_cond_sub_l(const fl_t& t) {
  fl_t d;
  uint64_t borrow = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t x = (uint64_t)t.v[i] - fl_modl(i) - borrow;
    d.v[i] = (uint32_t)x;
    borrow = (x >> 63);
  }
  fl_t r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = borrow ? t.v[i] : d.v[i];
  return r;
}
SB_HD fl_t fl_add_portable(const fl_t& a, const fl_t& b) {
  fl_t t;
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)a.v[i] + b.v[i];
    t.v[i] = (uint32_t)c;
    c >>= 32;
  }
  return fl_cond_sub_l(t);
}
SB_HD fl_t fl_sub_portable(const fl_t& a, const fl_t& b) {
  fl_t d;
  uint64_t borrow = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t x = (uint64_t)a.v[i] - b.v[i] - borrow;
    d.v[i] = (uint32_t)x;
    borrow = (x >> 63);
  }
  uint32_t mask = (uint32_t)0 - (uint32_t)borrow;
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)d.v[i] + (fl_modl(i) & mask);
    d.v[i] = (uint32_t)c;
    c >>= 32;
  }
  return d;
}

// Montgomery product a*b*2^-256 mod l, CIOS over 32-bit words.  The zero limbs
// of l (4..6) drop out of the reduction at compile time.
SB_HD fl_t fl_mul_portable(const fl_t& a, const fl_t& b) {
  uint32_t t[10];
#pragma unroll
  for (int i = 0; i < 10; i++) t[i] = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t c = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      c += (uint64_t)a.v[j] * b.v[i] + t[j];
      t[j] = (uint32_t)c;
      c >>= 32;
    }
    c += t[8];
    t[8] = (uint32_t)c;
    t[9] = (uint32_t)(c >> 32);
    uint32_t m = t[0] * SB_FL_INV32;
    c = (uint64_t)m * SB_FL_L0 + t[0];
    c >>= 32;
#pragma unroll
    for (int j = 1; j < 8; j++) {
      uint32_t lj = fl_modl(j);
      if (lj != 0) c += (uint64_t)m * lj;
      c += t[j];
      t[j - 1] = (uint32_t)c;
      c >>= 32;
    }
    c += t[8];
    t[7] = (uint32_t)c;
    t[8] = t[9] + (uint32_t)(c >> 32);
  }
  fl_t r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = t[i];
  return fl_cond_sub_l(r);
}
#if defined(__CUDA_ARCH__)
// ---- sm_100a versions: PTX carry chains (mp.cuh); bit-identical results to the portable code.
__device__ __forceinline__ fl_t fl_cond_sub_l_dev(const uint32_t* t) {
  uint32_t d0, d1, d2, d3, d4, d5, d6, d7, bw;
  asm("sub.cc.u32 %0,%9,%17;\n\tsubc.cc.u32 %1,%10,%18;\n\tsubc.cc.u32 %2,%11,%19;\n\tsubc.cc.u32 %3,%12,%20;\n\t"
      "subc.cc.u32 %4,%13,0;\n\tsubc.cc.u32 %5,%14,0;\n\tsubc.cc.u32 %6,%15,0;\n\tsubc.cc.u32 %7,%16,%21;\n\tsubc.u32 %8,0,0;"
      : "=r"(d0), "=r"(d1), "=r"(d2), "=r"(d3), "=r"(d4), "=r"(d5), "=r"(d6), "=r"(d7), "=r"(bw)
      : "r"(t[0]), "r"(t[1]), "r"(t[2]), "r"(t[3]), "r"(t[4]), "r"(t[5]), "r"(t[6]), "r"(t[7]),
        "r"(SB_FL_L0), "r"(SB_FL_L1), "r"(SB_FL_L2), "r"(SB_FL_L3), "r"(SB_FL_L7));
  fl_t r;
  r.v[0] = bw ? t[0] : d0; r.v[1] = bw ? t[1] : d1; r.v[2] = bw ? t[2] : d2; r.v[3] = bw ? t[3] : d3;
  r.v[4] = bw ? t[4] : d4; r.v[5] = bw ? t[5] : d5; r.v[6] = bw ? t[6] : d6; r.v[7] = bw ? t[7] : d7;
  return r;
}
__device__ __forceinline__ fl_t fl_add(const fl_t& a, const fl_t& b) {
  uint32_t t[8];
  asm("add.cc.u32 %0,%8,%16;\n\taddc.cc.u32 %1,%9,%17;\n\taddc.cc.u32 %2,%10,%18;\n\taddc.cc.u32 %3,%11,%19;\n\t"
      "addc.cc.u32 %4,%12,%20;\n\taddc.cc.u32 %5,%13,%21;\n\taddc.cc.u32 %6,%14,%22;\n\taddc.u32 %7,%15,%23;"
      : "=r"(t[0]), "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7])
      : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
        "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
  return fl_cond_sub_l_dev(t);
}
__device__ __forceinline__ fl_t fl_sub(const fl_t& a, const fl_t& b) {
  fl_t d;
  uint32_t bw;
  asm("sub.cc.u32 %0,%9,%17;\n\tsubc.cc.u32 %1,%10,%18;\n\tsubc.cc.u32 %2,%11,%19;\n\tsubc.cc.u32 %3,%12,%20;\n\t"
      "subc.cc.u32 %4,%13,%21;\n\tsubc.cc.u32 %5,%14,%22;\n\tsubc.cc.u32 %6,%15,%23;\n\tsubc.cc.u32 %7,%16,%24;\n\tsubc.u32 %8,0,0;"
      : "=r"(d.v[0]), "=r"(d.v[1]), "=r"(d.v[2]), "=r"(d.v[3]), "=r"(d.v[4]), "=r"(d.v[5]), "=r"(d.v[6]), "=r"(d.v[7]), "=r"(bw)
      : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
        "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
  asm("add.cc.u32 %0,%0,%8;\n\taddc.cc.u32 %1,%1,%9;\n\taddc.cc.u32 %2,%2,%10;\n\taddc.cc.u32 %3,%3,%11;\n\t"
      "addc.cc.u32 %4,%4,0;\n\taddc.cc.u32 %5,%5,0;\n\taddc.cc.u32 %6,%6,0;\n\taddc.u32 %7,%7,%12;"
      : "+r"(d.v[0]), "+r"(d.v[1]), "+r"(d.v[2]), "+r"(d.v[3]), "+r"(d.v[4]), "+r"(d.v[5]), "+r"(d.v[6]), "+r"(d.v[7])
      : "r"(bw & SB_FL_L0), "r"(bw & SB_FL_L1), "r"(bw & SB_FL_L2), "r"(bw & SB_FL_L3), "r"(bw & SB_FL_L7));
  return d;
}
// Montgomery product a*b*2^-256 mod l.  Full 16-limb product (mp_mul8), then word-by-word
// reduction t += m_i * l * 2^(32 i) with m_i = t_i * (-l^-1) mod 2^32, split by the shape
// of l = L0 + L1 2^32 + L2 2^64 + L3 2^96 + 2^252:
//   * the L0 product and the running carry stay in one 64-bit chain (it makes limb i zero);
//   * the L1, L3 products (limbs i+1.., i+3..) and the L2 product (limbs i+2..) are
//     accumulated lazily in two parity-split accumulators Ev / Od, whose integer values
//     never exceed (m_0..m_i) * 2^125 -- hence no carry out of limb i+4;
//   * the 2^252 term is a 4-bit shift of (m_7..m_0), added once at the end; only m_0's low
//     nibble reaches limb 7 and so feeds m_7.
__device__ __forceinline__ fl_t fl_mul(const fl_t& a, const fl_t& b) {
  uint32_t t[16];
  mp_mul8(t, a.v, b.v);
  // a*b < l^2 < 2^506, so this ORs in zero -- but ptxas cannot know, and therefore starts the reduction only after
  // the whole product.  Without it the scheduler interleaves the reduction's carry chains with the product rows,
  // runs out of carry predicates and spills them through P2R / ISETP: 327 instead of ~230 instructions per product.
  t[0] |= t[15] >> 26;
  uint32_t acc[2][12];  // acc[p][k]: limb k of the accumulator for positions of parity p
#pragma unroll
  for (int k = 0; k < 12; k++) acc[0][k] = acc[1][k] = 0;
  uint32_t m[9];
  m[8] = 0;
  uint64_t carry = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t cur = carry + t[i] + acc[0][i] + acc[1][i];
    if (i == 7) cur += (uint64_t)(m[0] << 28);
    m[i] = (uint32_t)cur * SB_FL_INV32;
    cur += (uint64_t)m[i] * SB_FL_L0;
    carry = cur >> 32;
    uint32_t* X = acc[(i + 1) & 1];
    uint32_t* Y = acc[i & 1];
    asm("mad.lo.cc.u32 %0,%4,%5,%0;\n\tmadc.hi.cc.u32 %1,%4,%5,%1;\n\tmadc.lo.cc.u32 %2,%4,%6,%2;\n\tmadc.hi.u32 %3,%4,%6,%3;"
        : "+r"(X[i + 1]), "+r"(X[i + 2]), "+r"(X[i + 3]), "+r"(X[i + 4])
        : "r"(m[i]), "r"(SB_FL_L1), "r"(SB_FL_L3));
    asm("mad.lo.cc.u32 %0,%3,%4,%0;\n\tmadc.hi.cc.u32 %1,%3,%4,%1;\n\taddc.u32 %2,%2,0;"
        : "+r"(Y[i + 2]), "+r"(Y[i + 3]), "+r"(Y[i + 4])
        : "r"(m[i]), "r"(SB_FL_L2));
  }
  // r = t_hi + Ev_hi + Od_hi + ((m_7..m_0) >> 4) + carry   (< 2l)
  uint32_t r[8];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    uint32_t msh = (m[k] >> 4) | (m[k + 1] << 28);
    carry += (uint64_t)t[8 + k] + msh;
    if (k < 4) carry += (uint64_t)acc[0][8 + k] + acc[1][8 + k];
    r[k] = (uint32_t)carry;
    carry >>= 32;
  }
  return fl_cond_sub_l_dev(r);
}
#else
SB_HD fl_t fl_add(const fl_t& a, const fl_t& b) { return fl_add_portable(a, b); }
SB_HD fl_t fl_sub(const fl_t& a, const fl_t& b) { return fl_sub_portable(a, b); }
SB_HD fl_t fl_mul(const fl_t& a, const fl_t& b) { return fl_mul_portable(a, b); }
#endif
SB_HD fl_t fl_neg(const fl_t& a) { return fl_sub(fl_zero(), a); }
SB_HD fl_t fl_dbl(const fl_t& a) { return fl_add(a, a); }
SB_HD fl_t fl_sqr(const fl_t& a) { return fl_mul(a, a); }
// multiplication policies, see FpInl / FpCall in fp.cuh
struct FlInl {
  static SB_HD fl_t mul(const fl_t& a, const fl_t& b) { return fl_mul(a, b); }
};
#if defined(__CUDACC__)
static __device__ __noinline__ fl_t fl_mul_call(fl_t a, fl_t b) { return fl_mul(a, b); }
struct FlCall {
  static __device__ __forceinline__ fl_t mul(const fl_t& a, const fl_t& b) { return fl_mul_call(a, b); }
};
#endif

// canonical integer (< l) -> Montgomery: Scalar::from_bytes multiplies by R2 (ristretto255.rs:417-419)
SB_HD fl_t fl_to_mont(const fl_t& canon) { return fl_mul(canon, fl_r2()); }
// Montgomery -> canonical integer: Scalar::to_bytes (ristretto255.rs:426-438)
SB_HD fl_t fl_from_mont(const fl_t& m) {
  fl_t one = fl_zero();
  one.v[0] = 1;
  return fl_mul(m, one);
}
// x < l ?  (Scalar::from_bytes canonicity check, ristretto255.rs:406-414)
SB_HD bool fl_is_canonical(const fl_t& x) {
  uint64_t borrow = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t d = (uint64_t)x.v[i] - fl_modl(i) - borrow;
    borrow = d >> 63;
  }
  return borrow != 0;
}

}  // namespace sb
