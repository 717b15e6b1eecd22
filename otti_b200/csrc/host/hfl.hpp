// Host-side F_l for the Fiat-Shamir tail of the prover (transcript scalars,
// UniPoly interpolation, sigma-protocol responses).  Same field and same
// Montgomery form (R = 2^256, fully reduced) as the device code in ../fl.cuh and
// as the reference's Scalar (Spartan/src/scalar/ristretto255.rs:200), here on
// 4 x u64 limbs with a fused (CIOS) multiply-reduce.
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

namespace sbh {

typedef unsigned __int128 u128;

struct Fl {
  uint64_t v[4];
  bool operator==(const Fl& o) const { return ((v[0] ^ o.v[0]) | (v[1] ^ o.v[1]) | (v[2] ^ o.v[2]) | (v[3] ^ o.v[3])) == 0; }
  bool is_zero() const { return (v[0] | v[1] | v[2] | v[3]) == 0; }
};

namespace flc {
static const uint64_t L[4] = {0x5812631a5cf5d3edULL, 0x14def9dea2f79cd6ULL, 0, 0x1000000000000000ULL};
static const uint64_t NINV = 0xd2b51da312547e1bULL;  // -l^-1 mod 2^64
static const Fl ONE = {{0xd6ec31748d98951dULL, 0xc6ef5bf4737dcf70ULL, 0xfffffffffffffffeULL, 0x0fffffffffffffffULL}};
static const Fl RR = {{0xa40611e3449c0f01ULL, 0xd00e1ba768859347ULL, 0xceec73d217f5be65ULL, 0x0399411b7c309a3dULL}};
static const Fl RRR = {{0x2a9e49687b83a2dbULL, 0x278324e6aef7f3ecULL, 0x8065dc6c04ec5b65ULL, 0x0e530b773599cec7ULL}};
}  // namespace flc

static inline Fl fl_zero() { return Fl{{0, 0, 0, 0}}; }
static inline Fl fl_one() { return flc::ONE; }

// x - l if x >= l
static inline Fl fl_reduce_once(const uint64_t x[4]) {
  uint64_t d[4];
  u128 b = 0;
  for (int i = 0; i < 4; i++) {
    u128 t = (u128)x[i] - flc::L[i] - (uint64_t)b;
    d[i] = (uint64_t)t;
    b = (t >> 64) & 1;
  }
  Fl r;
  for (int i = 0; i < 4; i++) r.v[i] = b ? x[i] : d[i];
  return r;
}
static inline Fl fl_add(const Fl& a, const Fl& b) {
  uint64_t s[4];
  u128 c = 0;
  for (int i = 0; i < 4; i++) {
    c += (u128)a.v[i] + b.v[i];
    s[i] = (uint64_t)c;
    c >>= 64;
  }
  return fl_reduce_once(s);
}
static inline Fl fl_sub(const Fl& a, const Fl& b) {
  uint64_t d[4];
  u128 bo = 0;
  for (int i = 0; i < 4; i++) {
    u128 t = (u128)a.v[i] - b.v[i] - (uint64_t)bo;
    d[i] = (uint64_t)t;
    bo = (t >> 64) & 1;
  }
  uint64_t mask = (uint64_t)0 - (uint64_t)bo;
  Fl r;
  u128 c = 0;
  for (int i = 0; i < 4; i++) {
    c += (u128)d[i] + (flc::L[i] & mask);
    r.v[i] = (uint64_t)c;
    c >>= 64;
  }
  return r;
}
static inline Fl fl_neg(const Fl& a) { return fl_sub(fl_zero(), a); }
static inline Fl fl_mul(const Fl& a, const Fl& b) {
  uint64_t t[6] = {0, 0, 0, 0, 0, 0};
  for (int i = 0; i < 4; i++) {
    u128 c = 0;
    for (int j = 0; j < 4; j++) {
      c += (u128)a.v[j] * b.v[i] + t[j];
      t[j] = (uint64_t)c;
      c >>= 64;
    }
    c += t[4];
    t[4] = (uint64_t)c;
    t[5] = (uint64_t)(c >> 64);
    uint64_t m = t[0] * flc::NINV;
    c = (u128)m * flc::L[0] + t[0];
    c >>= 64;
    c += (u128)m * flc::L[1] + t[1]; t[0] = (uint64_t)c; c >>= 64;
    c += t[2]; t[1] = (uint64_t)c; c >>= 64;  // L[2] == 0
    c += (u128)m * flc::L[3] + t[3]; t[2] = (uint64_t)c; c >>= 64;
    c += t[4]; t[3] = (uint64_t)c;
    t[4] = t[5] + (uint64_t)(c >> 64);
  }
  return fl_reduce_once(t);
}
static inline Fl fl_sqr(const Fl& a) { return fl_mul(a, a); }
static inline Fl fl_from_u64(uint64_t x) { return fl_mul(Fl{{x, 0, 0, 0}}, flc::RR); }
// canonical LE bytes -> Montgomery; false if >= l (Scalar::from_bytes, ristretto255.rs:398-422)
static inline bool fl_from_bytes(const uint8_t b[32], Fl& out) {
  Fl t;
  memcpy(t.v, b, 32);
  u128 bo = 0;
  for (int i = 0; i < 4; i++) {
    u128 d = (u128)t.v[i] - flc::L[i] - (uint64_t)bo;
    bo = (d >> 64) & 1;
  }
  out = fl_mul(t, flc::RR);
  return bo != 0;
}
static inline void fl_to_bytes(const Fl& a, uint8_t out[32]) {  // Scalar::to_bytes
  Fl t = fl_mul(a, Fl{{1, 0, 0, 0}});
  memcpy(out, t.v, 32);
}
static inline Fl fl_from_bytes_wide(const uint8_t b[64]) {  // Scalar::from_bytes_wide, ristretto255.rs:442-470
  Fl lo, hi;
  memcpy(lo.v, b, 32);
  memcpy(hi.v, b + 32, 32);
  return fl_add(fl_mul(lo, flc::RR), fl_mul(hi, flc::RRR));
}
// a^(l-2): plain square-and-multiply (same value as the reference's addition chain, ristretto255.rs:548-602)
static inline Fl fl_invert_fermat(const Fl& a) {
  static const uint64_t e[4] = {0x5812631a5cf5d3ebULL, 0x14def9dea2f79cd6ULL, 0, 0x1000000000000000ULL};
  Fl r = fl_one();
  for (int i = 252; i >= 0; i--) {
    r = fl_sqr(r);
    if ((e[i >> 6] >> (i & 63)) & 1) r = fl_mul(r, a);
  }
  return r;
}
// The same inverse by a binary extended Euclid on the raw limbs, halvings batched up to 62 bits at a time (x / 2^k mod l is
// (x + m l) >> k with m = x * (-l^-1) mod 2^k): about 190 subtract-and-shift steps instead of 320 Montgomery products
// (9.5 -> ~1.5 us).  Every bullet round inverts its challenge on the critical path (nizk/bullet.rs:100-101); the inputs are
// public transcript challenges, so the data-dependent running time gives nothing away.  Invariants: u = x1 * a, v = x2 * a
// (mod l) on the raw integers, v odd; the result a^-1 (raw: (a R)^-1) goes back to Montgomery form through R^3.
static inline Fl fl_invert(const Fl& a) {
  if (a.is_zero()) return fl_zero();  // as Fermat's: 0^(l-2) = 0
  uint64_t ua[4], va[4];
  Fl xa = Fl{{1, 0, 0, 0}}, xb = fl_zero();
  for (int i = 0; i < 4; i++) { ua[i] = a.v[i]; va[i] = flc::L[i]; }
  uint64_t *u = ua, *v = va;  // the roles swap by pointer
  Fl *x1 = &xa, *x2 = &xb;
  for (;;) {
    while ((u[0] & 1) == 0) {
      unsigned k = u[0] ? (unsigned)__builtin_ctzll(u[0]) : 62u;
      if (k > 62) k = 62;
      for (int i = 0; i < 3; i++) u[i] = (u[i] >> k) | (u[i + 1] << (64 - k));
      u[3] >>= k;
      const uint64_t m = (x1->v[0] * flc::NINV) & (((uint64_t)1 << k) - 1);  // x1 + m l = 0 (mod 2^k), and < l 2^k
      uint64_t t[5];
      u128 c = 0;
      for (int i = 0; i < 4; i++) {
        c += (u128)m * flc::L[i] + x1->v[i];
        t[i] = (uint64_t)c;
        c >>= 64;
      }
      t[4] = (uint64_t)c;
      for (int i = 0; i < 4; i++) x1->v[i] = (t[i] >> k) | (t[i + 1] << (64 - k));
    }
    if (u[0] == 1 && (u[1] | u[2] | u[3]) == 0) break;
    bool lt = false;  // u < v ?
    for (int i = 3; i >= 0; i--)
      if (u[i] != v[i]) { lt = u[i] < v[i]; break; }
    if (lt) {
      uint64_t* w = u; u = v; v = w;
      Fl* y = x1; x1 = x2; x2 = y;
    }
    u128 b = 0;
    for (int i = 0; i < 4; i++) {  // u -= v: both odd, u > v (they are coprime), so u becomes even and stays positive
      u128 d = (u128)u[i] - v[i] - (uint64_t)b;
      u[i] = (uint64_t)d;
      b = (d >> 64) & 1;
    }
    *x1 = fl_sub(*x1, *x2);
  }
  return fl_mul(*x1, flc::RRR);
}

typedef std::vector<Fl> FlVec;

}  // namespace sbh
