// ORACLE (test infrastructure only; never linked or included by otti_b200/).
// F_l, l = 2^252 + 27742317777372353535851937790883648493, in Montgomery form
// R = 2^256 on 4 x u64 limbs.  Restates the algorithms of
// /root/reference/Spartan/src/scalar/ristretto255.rs (line numbers cited per
// function).  Values are always fully reduced (< l) because the raw limbs are
// what the reference serialises (ristretto255.rs:199-200).
#pragma once
#include <cstdint>
#include <cstring>

namespace orc {

typedef unsigned __int128 u128;

// ristretto255.rs:20-38
static inline uint64_t adc(uint64_t a, uint64_t b, uint64_t& carry) {
  u128 r = (u128)a + b + carry;
  carry = (uint64_t)(r >> 64);
  return (uint64_t)r;
}
static inline uint64_t sbb(uint64_t a, uint64_t b, uint64_t& borrow) {
  u128 r = (u128)a - ((u128)b + (borrow >> 63));
  borrow = (uint64_t)(r >> 64);
  return (uint64_t)r;
}
static inline uint64_t mac(uint64_t a, uint64_t b, uint64_t c, uint64_t& carry) {
  u128 r = (u128)a + (u128)b * c + carry;
  carry = (uint64_t)(r >> 64);
  return (uint64_t)r;
}

struct Fl {
  uint64_t v[4];
  bool operator==(const Fl& o) const { return !memcmp(v, o.v, 32); }
  bool operator!=(const Fl& o) const { return !(*this == o); }
  bool is_zero() const { return (v[0] | v[1] | v[2] | v[3]) == 0; }
};

// ristretto255.rs:239-244, 303, 306-329
static const Fl FL_MODULUS = {{0x5812631a5cf5d3edULL, 0x14def9dea2f79cd6ULL, 0, 0x1000000000000000ULL}};
static const uint64_t FL_INV = 0xd2b51da312547e1bULL;
static const Fl FL_R = {{0xd6ec31748d98951dULL, 0xc6ef5bf4737dcf70ULL, 0xfffffffffffffffeULL, 0x0fffffffffffffffULL}};
static const Fl FL_R2 = {{0xa40611e3449c0f01ULL, 0xd00e1ba768859347ULL, 0xceec73d217f5be65ULL, 0x0399411b7c309a3dULL}};
static const Fl FL_R3 = {{0x2a9e49687b83a2dbULL, 0x278324e6aef7f3ecULL, 0x8065dc6c04ec5b65ULL, 0x0e530b773599cec7ULL}};

static inline Fl fl_zero() { return Fl{{0, 0, 0, 0}}; }
static inline Fl fl_one() { return FL_R; }

// Scalar::sub, ristretto255.rs:726-742
static inline Fl fl_sub(const Fl& a, const Fl& b) {
  uint64_t bo = 0;
  uint64_t d0 = sbb(a.v[0], b.v[0], bo), d1 = sbb(a.v[1], b.v[1], bo), d2 = sbb(a.v[2], b.v[2], bo),
           d3 = sbb(a.v[3], b.v[3], bo);
  uint64_t c = 0;
  Fl r;
  r.v[0] = adc(d0, FL_MODULUS.v[0] & bo, c);
  r.v[1] = adc(d1, FL_MODULUS.v[1] & bo, c);
  r.v[2] = adc(d2, FL_MODULUS.v[2] & bo, c);
  r.v[3] = adc(d3, FL_MODULUS.v[3] & bo, c);
  return r;
}
// Scalar::add, ristretto255.rs:745-755
static inline Fl fl_add(const Fl& a, const Fl& b) {
  uint64_t c = 0;
  Fl t;
  t.v[0] = adc(a.v[0], b.v[0], c);
  t.v[1] = adc(a.v[1], b.v[1], c);
  t.v[2] = adc(a.v[2], b.v[2], c);
  t.v[3] = adc(a.v[3], b.v[3], c);
  return fl_sub(t, FL_MODULUS);
}
// Scalar::neg, ristretto255.rs:758-775
static inline Fl fl_neg(const Fl& a) {
  uint64_t bo = 0;
  Fl r;
  r.v[0] = sbb(FL_MODULUS.v[0], a.v[0], bo);
  r.v[1] = sbb(FL_MODULUS.v[1], a.v[1], bo);
  r.v[2] = sbb(FL_MODULUS.v[2], a.v[2], bo);
  r.v[3] = sbb(FL_MODULUS.v[3], a.v[3], bo);
  uint64_t mask = (uint64_t)(a.is_zero() ? 1 : 0) - 1;
  for (int i = 0; i < 4; i++) r.v[i] &= mask;
  return r;
}
// Scalar::montgomery_reduce, ristretto255.rs:654-698 (HAC 14.32)
static inline Fl fl_montgomery_reduce(uint64_t r0, uint64_t r1, uint64_t r2, uint64_t r3, uint64_t r4, uint64_t r5,
                                      uint64_t r6, uint64_t r7) {
  const uint64_t* m = FL_MODULUS.v;
  uint64_t k, carry, carry2;
  k = r0 * FL_INV; carry = 0;
  (void)mac(r0, k, m[0], carry); r1 = mac(r1, k, m[1], carry); r2 = mac(r2, k, m[2], carry); r3 = mac(r3, k, m[3], carry);
  carry2 = 0; { uint64_t c = carry; r4 = adc(r4, 0, c); carry2 = c; }
  k = r1 * FL_INV; carry = 0;
  (void)mac(r1, k, m[0], carry); r2 = mac(r2, k, m[1], carry); r3 = mac(r3, k, m[2], carry); r4 = mac(r4, k, m[3], carry);
  { u128 t = (u128)r5 + carry2 + carry; r5 = (uint64_t)t; carry2 = (uint64_t)(t >> 64); }
  k = r2 * FL_INV; carry = 0;
  (void)mac(r2, k, m[0], carry); r3 = mac(r3, k, m[1], carry); r4 = mac(r4, k, m[2], carry); r5 = mac(r5, k, m[3], carry);
  { u128 t = (u128)r6 + carry2 + carry; r6 = (uint64_t)t; carry2 = (uint64_t)(t >> 64); }
  k = r3 * FL_INV; carry = 0;
  (void)mac(r3, k, m[0], carry); r4 = mac(r4, k, m[1], carry); r5 = mac(r5, k, m[2], carry); r6 = mac(r6, k, m[3], carry);
  { u128 t = (u128)r7 + carry2 + carry; r7 = (uint64_t)t; }
  return fl_sub(Fl{{r4, r5, r6, r7}}, FL_MODULUS);
}
// Scalar::mul, ristretto255.rs:702-723 (schoolbook then reduce)
static inline Fl fl_mul(const Fl& a, const Fl& b) {
  uint64_t r0, r1, r2, r3, r4, r5, r6, r7, c;
  c = 0; r0 = mac(0, a.v[0], b.v[0], c); r1 = mac(0, a.v[0], b.v[1], c); r2 = mac(0, a.v[0], b.v[2], c); r3 = mac(0, a.v[0], b.v[3], c); r4 = c;
  c = 0; r1 = mac(r1, a.v[1], b.v[0], c); r2 = mac(r2, a.v[1], b.v[1], c); r3 = mac(r3, a.v[1], b.v[2], c); r4 = mac(r4, a.v[1], b.v[3], c); r5 = c;
  c = 0; r2 = mac(r2, a.v[2], b.v[0], c); r3 = mac(r3, a.v[2], b.v[1], c); r4 = mac(r4, a.v[2], b.v[2], c); r5 = mac(r5, a.v[2], b.v[3], c); r6 = c;
  c = 0; r3 = mac(r3, a.v[3], b.v[0], c); r4 = mac(r4, a.v[3], b.v[1], c); r5 = mac(r5, a.v[3], b.v[2], c); r6 = mac(r6, a.v[3], b.v[3], c); r7 = c;
  return fl_montgomery_reduce(r0, r1, r2, r3, r4, r5, r6, r7);
}
// Scalar::square, ristretto255.rs:474-506 (value == mul(a,a))
static inline Fl fl_square(const Fl& a) { return fl_mul(a, a); }

// Scalar::from_bytes, ristretto255.rs:398-422; returns false if >= l
static inline bool fl_from_bytes(const uint8_t b[32], Fl& out) {
  Fl t;
  memcpy(t.v, b, 32);
  uint64_t bo = 0;
  (void)sbb(t.v[0], FL_MODULUS.v[0], bo); (void)sbb(t.v[1], FL_MODULUS.v[1], bo);
  (void)sbb(t.v[2], FL_MODULUS.v[2], bo); (void)sbb(t.v[3], FL_MODULUS.v[3], bo);
  bool ok = (bo & 1) != 0;
  out = fl_mul(t, FL_R2);
  return ok;
}
// Scalar::to_bytes, ristretto255.rs:426-438
static inline void fl_to_bytes(const Fl& a, uint8_t out[32]) {
  Fl t = fl_montgomery_reduce(a.v[0], a.v[1], a.v[2], a.v[3], 0, 0, 0, 0);
  memcpy(out, t.v, 32);
}
// Scalar::from_u512 / from_bytes_wide, ristretto255.rs:442-470
static inline Fl fl_from_u512(const uint64_t l[8]) {
  Fl d0{{l[0], l[1], l[2], l[3]}}, d1{{l[4], l[5], l[6], l[7]}};
  return fl_add(fl_mul(d0, FL_R2), fl_mul(d1, FL_R3));
}
static inline Fl fl_from_bytes_wide(const uint8_t b[64]) {
  uint64_t l[8];
  memcpy(l, b, 64);
  return fl_from_u512(l);
}
// From<u64>, ristretto255.rs:215-219
static inline Fl fl_from_u64(uint64_t x) { return fl_mul(Fl{{x, 0, 0, 0}}, FL_R2); }

// Scalar::invert, ristretto255.rs:548-602 (addition chain); zero -> zero (CtOption none)
static inline Fl fl_invert(const Fl& a) {
  Fl _1 = a, _10 = fl_square(_1), _100 = fl_square(_10), _11 = fl_mul(_10, _1), _101 = fl_mul(_10, _11),
     _111 = fl_mul(_10, _101), _1001 = fl_mul(_10, _111), _1011 = fl_mul(_10, _1001), _1111 = fl_mul(_100, _1011);
  Fl y = fl_mul(_1111, _1);
  auto sm = [&](int sq, const Fl& x) {
    for (int i = 0; i < sq; i++) y = fl_square(y);
    y = fl_mul(y, x);
  };
  sm(123 + 3, _101); sm(2 + 2, _11); sm(1 + 4, _1111); sm(1 + 4, _1111); sm(4, _1001); sm(2, _11); sm(1 + 4, _1111);
  sm(1 + 3, _101); sm(3 + 3, _101); sm(3, _111); sm(1 + 4, _1111); sm(2 + 3, _111); sm(2 + 2, _11); sm(1 + 4, _1011);
  sm(2 + 4, _1011); sm(6 + 4, _1001); sm(2 + 2, _11); sm(3 + 2, _11); sm(3 + 2, _11); sm(1 + 4, _1001); sm(1 + 3, _111);
  sm(2 + 4, _1111); sm(1 + 4, _1011); sm(3, _101); sm(2 + 4, _1111); sm(3, _101); sm(1 + 2, _11);
  return y;
}

// ScalarFromPrimitives for usize, scalar/mod.rs:10-15 ((0..n) sum of one)
static inline Fl fl_from_usize(size_t n) { return fl_from_u64((uint64_t)n); }

}  // namespace orc
