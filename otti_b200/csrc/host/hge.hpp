// Host-side F_p / Edwards25519 / ristretto255 for the prover's Fiat-Shamir
// tail: generator derivation (Spartan/src/commitments.rs:15-33), the tiny
// Pedersen commitments of every sum-check round and sigma protocol
// (sumcheck.rs:659-706, nizk/mod.rs:39-45,101-109,183-205,333-345) and their
// compression.  The reference gets this arithmetic from curve25519-dalek 3.2;
// this is a from-scratch 5 x 51-bit implementation with per-generator
// fixed-base window tables (the generators are public parameters, fixed for
// the life of an NIZKGens).
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

#include "hfl.hpp"

namespace sbh {

struct Fe {
  uint64_t v[5];
};
static const uint64_t MASK51 = (1ULL << 51) - 1;

static inline Fe fe_zero() { return Fe{{0, 0, 0, 0, 0}}; }
static inline Fe fe_one() { return Fe{{1, 0, 0, 0, 0}}; }
static inline Fe fe_weak(Fe r) {
  uint64_t c0 = r.v[0] >> 51, c1 = r.v[1] >> 51, c2 = r.v[2] >> 51, c3 = r.v[3] >> 51, c4 = r.v[4] >> 51;
  r.v[0] = (r.v[0] & MASK51) + c4 * 19;
  r.v[1] = (r.v[1] & MASK51) + c0;
  r.v[2] = (r.v[2] & MASK51) + c1;
  r.v[3] = (r.v[3] & MASK51) + c2;
  r.v[4] = (r.v[4] & MASK51) + c3;
  return r;
}
static inline Fe fe_add(const Fe& a, const Fe& b) {
  Fe r;
  for (int i = 0; i < 5; i++) r.v[i] = a.v[i] + b.v[i];
  return fe_weak(r);
}
static inline Fe fe_sub(const Fe& a, const Fe& b) {
  Fe r;  // + 16p keeps every limb positive
  r.v[0] = a.v[0] + 0x7ffffffffffed0ULL - b.v[0];
  r.v[1] = a.v[1] + 0x7ffffffffffff0ULL - b.v[1];
  r.v[2] = a.v[2] + 0x7ffffffffffff0ULL - b.v[2];
  r.v[3] = a.v[3] + 0x7ffffffffffff0ULL - b.v[3];
  r.v[4] = a.v[4] + 0x7ffffffffffff0ULL - b.v[4];
  return fe_weak(r);
}
static inline Fe fe_neg(const Fe& a) { return fe_sub(fe_zero(), a); }
static inline Fe fe_mul(const Fe& f, const Fe& g) {
  const uint64_t f0 = f.v[0], f1 = f.v[1], f2 = f.v[2], f3 = f.v[3], f4 = f.v[4];
  const uint64_t g0 = g.v[0], g1 = g.v[1], g2 = g.v[2], g3 = g.v[3], g4 = g.v[4];
  const uint64_t g1_19 = g1 * 19, g2_19 = g2 * 19, g3_19 = g3 * 19, g4_19 = g4 * 19;
  u128 h0 = (u128)f0 * g0 + (u128)f1 * g4_19 + (u128)f2 * g3_19 + (u128)f3 * g2_19 + (u128)f4 * g1_19;
  u128 h1 = (u128)f0 * g1 + (u128)f1 * g0 + (u128)f2 * g4_19 + (u128)f3 * g3_19 + (u128)f4 * g2_19;
  u128 h2 = (u128)f0 * g2 + (u128)f1 * g1 + (u128)f2 * g0 + (u128)f3 * g4_19 + (u128)f4 * g3_19;
  u128 h3 = (u128)f0 * g3 + (u128)f1 * g2 + (u128)f2 * g1 + (u128)f3 * g0 + (u128)f4 * g4_19;
  u128 h4 = (u128)f0 * g4 + (u128)f1 * g3 + (u128)f2 * g2 + (u128)f3 * g1 + (u128)f4 * g0;
  Fe r;
  h1 += (uint64_t)(h0 >> 51); r.v[0] = (uint64_t)h0 & MASK51;
  h2 += (uint64_t)(h1 >> 51); r.v[1] = (uint64_t)h1 & MASK51;
  h3 += (uint64_t)(h2 >> 51); r.v[2] = (uint64_t)h2 & MASK51;
  h4 += (uint64_t)(h3 >> 51); r.v[3] = (uint64_t)h3 & MASK51;
  uint64_t top = (uint64_t)(h4 >> 51);
  r.v[4] = (uint64_t)h4 & MASK51;
  r.v[0] += top * 19;
  r.v[1] += r.v[0] >> 51;
  r.v[0] &= MASK51;
  return r;
}
static inline Fe fe_sqr(const Fe& f) {
  const uint64_t f0 = f.v[0], f1 = f.v[1], f2 = f.v[2], f3 = f.v[3], f4 = f.v[4];
  const uint64_t f0_2 = f0 * 2, f1_2 = f1 * 2, f3_19 = f3 * 19, f4_19 = f4 * 19;
  u128 h0 = (u128)f0 * f0 + (u128)f1_2 * f4_19 + (u128)(f2 * 2) * f3_19;
  u128 h1 = (u128)f0_2 * f1 + (u128)(f2 * 2) * f4_19 + (u128)f3 * f3_19;
  u128 h2 = (u128)f0_2 * f2 + (u128)f1 * f1 + (u128)(f3 * 2) * f4_19;
  u128 h3 = (u128)f0_2 * f3 + (u128)f1_2 * f2 + (u128)f4 * f4_19;
  u128 h4 = (u128)f0_2 * f4 + (u128)f1_2 * f3 + (u128)f2 * f2;
  Fe r;
  h1 += (uint64_t)(h0 >> 51); r.v[0] = (uint64_t)h0 & MASK51;
  h2 += (uint64_t)(h1 >> 51); r.v[1] = (uint64_t)h1 & MASK51;
  h3 += (uint64_t)(h2 >> 51); r.v[2] = (uint64_t)h2 & MASK51;
  h4 += (uint64_t)(h3 >> 51); r.v[3] = (uint64_t)h3 & MASK51;
  uint64_t top = (uint64_t)(h4 >> 51);
  r.v[4] = (uint64_t)h4 & MASK51;
  r.v[0] += top * 19;
  r.v[1] += r.v[0] >> 51;
  r.v[0] &= MASK51;
  return r;
}
static inline Fe fe_sqn(Fe a, int n) {
  while (n-- > 0) a = fe_sqr(a);
  return a;
}
static inline void fe_tobytes(const Fe& a, uint8_t out[32]) {
  Fe t = fe_weak(fe_weak(a));
  uint64_t q = (t.v[0] + 19) >> 51;
  q = (t.v[1] + q) >> 51; q = (t.v[2] + q) >> 51; q = (t.v[3] + q) >> 51; q = (t.v[4] + q) >> 51;
  t.v[0] += 19 * q;
  t.v[1] += t.v[0] >> 51; t.v[0] &= MASK51;
  t.v[2] += t.v[1] >> 51; t.v[1] &= MASK51;
  t.v[3] += t.v[2] >> 51; t.v[2] &= MASK51;
  t.v[4] += t.v[3] >> 51; t.v[3] &= MASK51;
  t.v[4] &= MASK51;
  uint64_t w[4] = {t.v[0] | (t.v[1] << 51), (t.v[1] >> 13) | (t.v[2] << 38), (t.v[2] >> 26) | (t.v[3] << 25),
                   (t.v[3] >> 39) | (t.v[4] << 12)};
  memcpy(out, w, 32);
}
static inline Fe fe_frombytes(const uint8_t b[32]) {  // bit 255 ignored
  uint64_t w[4];
  memcpy(w, b, 32);
  Fe r;
  r.v[0] = w[0] & MASK51;
  r.v[1] = ((w[0] >> 51) | (w[1] << 13)) & MASK51;
  r.v[2] = ((w[1] >> 38) | (w[2] << 26)) & MASK51;
  r.v[3] = ((w[2] >> 25) | (w[3] << 39)) & MASK51;
  r.v[4] = (w[3] >> 12) & MASK51;
  return r;
}
static inline bool fe_equal(const Fe& a, const Fe& b) {
  uint8_t x[32], y[32];
  fe_tobytes(a, x);
  fe_tobytes(b, y);
  return memcmp(x, y, 32) == 0;
}
static inline bool fe_isneg(const Fe& a) {
  uint8_t x[32];
  fe_tobytes(a, x);
  return x[0] & 1;
}
static inline bool fe_iszero(const Fe& a) {
  uint8_t x[32];
  static const uint8_t z[32] = {0};
  fe_tobytes(a, x);
  return memcmp(x, z, 32) == 0;
}
static inline Fe fe_abs(const Fe& a) { return fe_isneg(a) ? fe_neg(a) : a; }
static inline Fe fe_pow22523(const Fe& z) {  // z^(2^252-3)
  Fe t0 = fe_sqr(z), t1 = fe_mul(z, fe_sqn(t0, 2));
  t0 = fe_mul(t0, t1);                 // z^11
  t0 = fe_mul(t1, fe_sqr(t0));         // z^31 = 2^5-1
  t1 = fe_mul(fe_sqn(t0, 5), t0);      // 2^10-1
  Fe t2 = fe_mul(fe_sqn(t1, 10), t1);  // 2^20-1
  t2 = fe_mul(fe_sqn(t2, 20), t2);     // 2^40-1
  t1 = fe_mul(fe_sqn(t2, 10), t1);     // 2^50-1
  t2 = fe_mul(fe_sqn(t1, 50), t1);     // 2^100-1
  t2 = fe_mul(fe_sqn(t2, 100), t2);    // 2^200-1
  t1 = fe_mul(fe_sqn(t2, 50), t1);     // 2^250-1
  return fe_mul(fe_sqn(t1, 2), z);     // 2^252-3
}
static inline Fe fe_invert(const Fe& z) { return fe_mul(fe_sqn(fe_pow22523(z), 3), fe_mul(fe_sqr(z), z)); }
static inline Fe fe_from_le_hex(const char* h) {
  uint8_t b[32];
  for (int i = 0; i < 32; i++) {
    auto nib = [](char c) { return c <= '9' ? c - '0' : c - 'a' + 10; };
    b[i] = (uint8_t)(nib(h[2 * i]) << 4 | nib(h[2 * i + 1]));
  }
  return fe_frombytes(b);
}

struct CurveConsts {
  Fe d, d2, sqrtm1, invsqrt_a_minus_d, one_minus_d_sq, d_minus_one_sq, sqrt_ad_minus_one, bx, by;
  CurveConsts() {
    d = fe_from_le_hex("a3785913ca4deb75abd841414d0a700098e879777940c78c73fe6f2bee6c0352");
    d2 = fe_add(d, d);
    sqrtm1 = fe_from_le_hex("b0a00e4a271beec478e42fad0618432fa7d7fb3d99004d2b0bdfc14f8024832b");
    invsqrt_a_minus_d = fe_from_le_hex("ea405d80aafdc899be72415a17162f9d40d801fe917bc216a2fcafcf05896c78");
    sqrt_ad_minus_one = fe_from_le_hex("1b2e7b49a0f6977ebd54781b0c8e9daffdd1f531c9fc3c0fac48832bbf316937");
    one_minus_d_sq = fe_sub(fe_one(), fe_sqr(d));
    d_minus_one_sq = fe_sqr(fe_sub(d, fe_one()));
    bx = fe_from_le_hex("1ad5258f602d56c9b2a7259560c72c695cdcd6fd31e2a4c0fe536ecdd3366921");
    by = fe_from_le_hex("5866666666666666666666666666666666666666666666666666666666666666");
  }
};
static inline const CurveConsts& CC() {
  static const CurveConsts c;
  return c;
}

// (was_square, abs(sqrt(u/v)))  -- RFC 9496 4.2
static inline bool fe_sqrt_ratio(const Fe& u, const Fe& v, Fe& out) {
  Fe v3 = fe_mul(fe_sqr(v), v), v7 = fe_mul(fe_sqr(v3), v);
  Fe r = fe_mul(fe_mul(u, v3), fe_pow22523(fe_mul(u, v7)));
  Fe chk = fe_mul(v, fe_sqr(r)), nu = fe_neg(u);
  bool ok = fe_equal(chk, u), flip = fe_equal(chk, nu), flip_i = fe_equal(chk, fe_mul(nu, CC().sqrtm1));
  if (flip || flip_i) r = fe_mul(r, CC().sqrtm1);
  out = fe_abs(r);
  return ok || flip;
}

struct Ge {  // extended
  Fe X, Y, Z, T;
};
struct GeNiels {  // affine (y+x, y-x, 2dxy)
  Fe ypx, ymx, t2d;
};
static inline Ge ge_identity() { return Ge{fe_zero(), fe_one(), fe_one(), fe_zero()}; }
static inline Ge ge_add(const Ge& p, const Ge& q) {
  Fe A = fe_mul(fe_sub(p.Y, p.X), fe_sub(q.Y, q.X)), B = fe_mul(fe_add(p.Y, p.X), fe_add(q.Y, q.X));
  Fe C = fe_mul(fe_mul(p.T, q.T), CC().d2), D = fe_mul(p.Z, q.Z);
  D = fe_add(D, D);
  Fe E = fe_sub(B, A), F = fe_sub(D, C), G = fe_add(D, C), H = fe_add(B, A);
  return Ge{fe_mul(E, F), fe_mul(G, H), fe_mul(F, G), fe_mul(E, H)};
}
static inline Ge ge_madd(const Ge& p, const GeNiels& q) {
  Fe A = fe_mul(fe_sub(p.Y, p.X), q.ymx), B = fe_mul(fe_add(p.Y, p.X), q.ypx), C = fe_mul(p.T, q.t2d);
  Fe D = fe_add(p.Z, p.Z);
  Fe E = fe_sub(B, A), F = fe_sub(D, C), G = fe_add(D, C), H = fe_add(B, A);
  return Ge{fe_mul(E, F), fe_mul(G, H), fe_mul(F, G), fe_mul(E, H)};
}
static inline Ge ge_double(const Ge& p) {
  Fe A = fe_sqr(p.X), B = fe_sqr(p.Y), C = fe_sqr(p.Z);
  C = fe_add(C, C);
  Fe D = fe_neg(A), E = fe_sub(fe_sub(fe_sqr(fe_add(p.X, p.Y)), A), B);
  Fe G = fe_add(D, B), F = fe_sub(G, C), H = fe_sub(D, B);
  return Ge{fe_mul(E, F), fe_mul(G, H), fe_mul(F, G), fe_mul(E, H)};
}
static inline Ge ge_neg(const Ge& p) { return Ge{fe_neg(p.X), p.Y, p.Z, fe_neg(p.T)}; }
static inline bool ge_equal(const Ge& a, const Ge& b) {
  return fe_equal(fe_mul(a.X, b.Y), fe_mul(a.Y, b.X)) || fe_equal(fe_mul(a.X, b.X), fe_mul(a.Y, b.Y));
}
static inline void ge_compress(const Ge& p, uint8_t out[32]) {
  const CurveConsts& k = CC();
  Fe u1 = fe_mul(fe_add(p.Z, p.Y), fe_sub(p.Z, p.Y)), u2 = fe_mul(p.X, p.Y);
  Fe inv;
  fe_sqrt_ratio(fe_one(), fe_mul(u1, fe_sqr(u2)), inv);
  Fe d1 = fe_mul(inv, u1), d2 = fe_mul(inv, u2), zinv = fe_mul(fe_mul(d1, d2), p.T);
  Fe x = p.X, y = p.Y, dn = d2;
  if (fe_isneg(fe_mul(p.T, zinv))) {
    x = fe_mul(p.Y, k.sqrtm1);
    y = fe_mul(p.X, k.sqrtm1);
    dn = fe_mul(d1, k.invsqrt_a_minus_d);
  }
  if (fe_isneg(fe_mul(x, zinv))) y = fe_neg(y);
  fe_tobytes(fe_abs(fe_mul(dn, fe_sub(p.Z, y))), out);
}
static inline bool ge_decompress(const uint8_t b[32], Ge& out) {
  Fe s = fe_frombytes(b);
  uint8_t chk[32];
  fe_tobytes(s, chk);
  if (memcmp(chk, b, 32) != 0 || (b[0] & 1)) return false;
  Fe ss = fe_sqr(s), u1 = fe_sub(fe_one(), ss), u2 = fe_add(fe_one(), ss), u2s = fe_sqr(u2);
  Fe v = fe_sub(fe_neg(fe_mul(CC().d, fe_sqr(u1))), u2s), inv;
  bool ok = fe_sqrt_ratio(fe_one(), fe_mul(v, u2s), inv);
  Fe dx = fe_mul(inv, u2), dy = fe_mul(fe_mul(inv, dx), v);
  Fe x = fe_abs(fe_mul(fe_add(s, s), dx)), y = fe_mul(u1, dy), t = fe_mul(x, y);
  if (!ok || fe_isneg(t) || fe_iszero(y)) return false;
  out = Ge{x, y, fe_one(), t};
  return true;
}
static inline Ge ge_elligator(const Fe& t) {
  const CurveConsts& k = CC();
  Fe r = fe_mul(k.sqrtm1, fe_sqr(t)), u = fe_mul(fe_add(r, fe_one()), k.one_minus_d_sq);
  Fe v = fe_mul(fe_sub(fe_neg(fe_one()), fe_mul(r, k.d)), fe_add(r, k.d)), s;
  bool sq = fe_sqrt_ratio(u, v, s);
  Fe sprime = fe_neg(fe_abs(fe_mul(s, t)));
  if (!sq) s = sprime;
  Fe c = sq ? fe_neg(fe_one()) : r;
  Fe N = fe_sub(fe_mul(fe_mul(c, fe_sub(r, fe_one())), k.d_minus_one_sq), v);
  Fe w0 = fe_mul(fe_add(s, s), v), w1 = fe_mul(N, k.sqrt_ad_minus_one), s2 = fe_sqr(s);
  Fe w2 = fe_sub(fe_one(), s2), w3 = fe_add(fe_one(), s2);
  return Ge{fe_mul(w0, w3), fe_mul(w2, w1), fe_mul(w1, w3), fe_mul(w0, w2)};
}
static inline Ge ge_from_uniform_bytes(const uint8_t b[64]) {
  return ge_add(ge_elligator(fe_frombytes(b)), ge_elligator(fe_frombytes(b + 32)));
}
// variable-base k*P, k canonical 32 bytes, 4-bit windows (verify side only)
static inline Ge ge_scalarmul(const uint8_t k[32], const Ge& p) {
  Ge tbl[16];
  tbl[0] = ge_identity();
  for (int i = 1; i < 16; i++) tbl[i] = ge_add(tbl[i - 1], p);
  Ge acc = ge_identity();
  for (int i = 63; i >= 0; i--) {
    for (int j = 0; j < 4; j++) acc = ge_double(acc);
    int d = (k[i >> 1] >> ((i & 1) * 4)) & 15;
    if (d) acc = ge_add(acc, tbl[d]);
  }
  return acc;
}
static inline Ge ge_scalarmul(const Fl& k, const Ge& p) {
  uint8_t kb[32];
  fl_to_bytes(k, kb);
  return ge_scalarmul(kb, p);
}

// a*P + b*Q with one shared doubling chain (Straus, 4-bit windows): the verifier's comm_target = w0*C1 + w1*C2
static inline Ge ge_double_scalarmul(const Fl& a, const Ge& p, const Fl& b, const Ge& q) {
  uint8_t ka[32], kb[32];
  fl_to_bytes(a, ka);
  fl_to_bytes(b, kb);
  Ge tp[16], tq[16];
  tp[0] = tq[0] = ge_identity();
  for (int i = 1; i < 16; i++) { tp[i] = ge_add(tp[i - 1], p); tq[i] = ge_add(tq[i - 1], q); }
  Ge acc = ge_identity();
  for (int i = 63; i >= 0; i--) {
    for (int j = 0; j < 4; j++) acc = ge_double(acc);
    const int da = (ka[i >> 1] >> ((i & 1) * 4)) & 15, db = (kb[i >> 1] >> ((i & 1) * 4)) & 15;
    if (da) acc = ge_add(acc, tp[da]);
    if (db) acc = ge_add(acc, tq[db]);
  }
  return acc;
}

// Batch-normalise extended points to affine Niels with one inversion.
static inline void ge_batch_to_niels(const Ge* pts, size_t n, GeNiels* out) {
  if (!n) return;
  std::vector<Fe> pre(n);
  Fe acc = fe_one();
  for (size_t i = 0; i < n; i++) {
    pre[i] = acc;
    acc = fe_mul(acc, pts[i].Z);
  }
  acc = fe_invert(acc);
  for (size_t i = n; i-- > 0;) {
    Fe zi = fe_mul(acc, pre[i]);
    acc = fe_mul(acc, pts[i].Z);
    Fe x = fe_mul(pts[i].X, zi), y = fe_mul(pts[i].Y, zi);
    out[i].ypx = fe_add(y, x);
    out[i].ymx = fe_sub(y, x);
    out[i].t2d = fe_mul(fe_mul(x, y), CC().d2);
  }
}

// Fixed-base table: 32 windows of 8 bits, entries d*256^w*P for d = 1..255.
// device fp_t (8 LE words, any representative < 2^256) -> host field element
static inline Fe fe_from_loose(const uint8_t* p) {
  Fe r = fe_frombytes(p);  // ignores bit 255
  if (p[31] & 0x80) r.v[0] += 19;  // 2^255 = 19 (mod p); limbs have 13 bits of headroom
  return r;
}

// k * P for one of the seven generators of the Fiat-Shamir tail.  Two tables:
//   narrow  32 windows x 255 multiples (unsigned 8-bit digits), built on the host, ~1 MB, cache resident;
//   wide    a host copy of this base's slice of the commitment kernel's window table (device format: ww windows x
//           2^(wc-1) entries of 96 bytes = (y+x, y-x, 2dxy) as 8 LE words each, signed radix-2^wc digits).  Half as
//           many additions per multiplication at wc = 16; the entries are prefetched before the additions start, so
//           the table need not fit any cache.
struct FixedBase {
  std::vector<GeNiels> tbl;  // [32][255]
  Ge point;
  const uint8_t* wide = nullptr;
  uint32_t wc = 0, ww = 0;
  void build(const Ge& p) {
    point = p;
    std::vector<Ge> ext(32 * 255);
    Ge base = p;
    for (int w = 0; w < 32; w++) {
      Ge cur = base;
      for (int d = 0; d < 255; d++) {
        ext[w * 255 + d] = cur;
        cur = ge_add(cur, base);
      }
      base = cur;  // 256 * base
    }
    tbl.resize(32 * 255);
    ge_batch_to_niels(ext.data(), ext.size(), tbl.data());
  }
  void set_wide(const uint8_t* entries, uint32_t c, uint32_t windows) {
    wide = entries;
    wc = c;
    ww = windows;
  }
  // acc += k * P, k canonical 32 bytes
  void mul_acc(Ge& acc, const uint8_t k[32]) const { mul_acc_range(acc, k, 0, 32); }
  // the part [lo, hi) / 32 of the windows only: k*P is a sum over windows, so one multiplication can be split over threads
  void mul_acc_range(Ge& acc, const uint8_t k[32], int lo, int hi) const {
    if (wide == nullptr) {
      for (int w = lo; w < hi; w++)
        if (k[w]) acc = ge_madd(acc, tbl[w * 255 + k[w] - 1]);
      return;
    }
    const uint32_t w0 = ww * (uint32_t)lo / 32, w1 = ww * (uint32_t)hi / 32;
    uint32_t kw[8];
    memcpy(kw, k, 32);
    // signed digits in (-2^(c-1), 2^(c-1)] with the running carry: the recoding of sb::msm_digit (ge.cuh)
    const uint32_t half = 1u << (wc - 1), mask = (1u << wc) - 1;
    const uint8_t* e[64];
    bool neg[64];
    uint32_t carry = 0;
    for (uint32_t w = 0; w < w1 && w < 64; w++) {
      const uint32_t bit = w * wc, limb = bit >> 5, sh = bit & 31;
      const uint64_t two = ((uint64_t)(limb < 7 ? kw[limb + 1] : 0u) << 32) | kw[limb];
      uint32_t d = ((uint32_t)(two >> sh) & mask) + carry;
      const bool n = d > half;
      carry = n ? 1u : 0u;
      if (n) d = (1u << wc) - d;
      e[w] = nullptr;
      if (w >= w0 && d != 0) {
        e[w] = wide + (((size_t)w << (wc - 1)) + (d - 1)) * 96;
        neg[w] = n;
        __builtin_prefetch(e[w]);
        __builtin_prefetch(e[w] + 64);
      }
    }
    for (uint32_t w = w0; w < w1 && w < 64; w++) {
      if (e[w] == nullptr) continue;
      GeNiels q{fe_from_loose(e[w]), fe_from_loose(e[w] + 32), fe_from_loose(e[w] + 64)};
      if (neg[w]) {  // -P = (y-x, y+x, -2dxy)
        std::swap(q.ypx, q.ymx);
        q.t2d = fe_neg(q.t2d);
      }
      acc = ge_madd(acc, q);
    }
  }
};

}  // namespace sbh
