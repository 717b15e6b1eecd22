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
SB_HD fl_t fl_add(const fl_t& a, const fl_t& b) {
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
SB_HD fl_t fl_sub(const fl_t& a, const fl_t& b) {
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
SB_HD fl_t fl_neg(const fl_t& a) { return fl_sub(fl_zero(), a); }
SB_HD fl_t fl_dbl(const fl_t& a) { return fl_add(a, a); }

// Montgomery product a*b*2^-256 mod l, CIOS over 32-bit words.  The zero limbs
// of l (4..6) drop out of the reduction at compile time.
SB_HD fl_t fl_mul(const fl_t& a, const fl_t& b) {
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
SB_HD fl_t fl_sqr(const fl_t& a) { return fl_mul(a, a); }

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
