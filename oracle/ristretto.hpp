// ORACLE (test infrastructure only; never linked or included by otti_b200/).
// F_p (p = 2^255-19) on 5 x 51-bit limbs, Edwards25519 extended coordinates and
// ristretto255 encode/decode/Elligator.  The reference takes these from the
// un-vendored crate curve25519-dalek 3.2.x (Spartan/Cargo.toml:14; call sites
// Spartan/src/group.rs:6-7,23-24,98-117, commitments.rs:25); this restates the
// published algorithms (RFC 9496 section 4).
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

#include "fl.hpp"

namespace orc {

struct Fp {
  uint64_t v[5];
};
static const uint64_t M51 = (1ULL << 51) - 1;

static inline Fp fp_from_u64(uint64_t x) { return Fp{{x & M51, x >> 51, 0, 0, 0}}; }
static inline Fp fp_zero() { return Fp{{0, 0, 0, 0, 0}}; }
static inline Fp fp_one() { return Fp{{1, 0, 0, 0, 0}}; }

static inline Fp fp_carry(const Fp& a) {  // weak reduction: limbs < 2^51 + small
  Fp r = a;
  uint64_t c;
  c = r.v[0] >> 51; r.v[0] &= M51; r.v[1] += c;
  c = r.v[1] >> 51; r.v[1] &= M51; r.v[2] += c;
  c = r.v[2] >> 51; r.v[2] &= M51; r.v[3] += c;
  c = r.v[3] >> 51; r.v[3] &= M51; r.v[4] += c;
  c = r.v[4] >> 51; r.v[4] &= M51; r.v[0] += c * 19;
  return r;
}
static inline Fp fp_add(const Fp& a, const Fp& b) {
  Fp r;
  for (int i = 0; i < 5; i++) r.v[i] = a.v[i] + b.v[i];
  return fp_carry(r);
}
static inline Fp fp_sub(const Fp& a, const Fp& b) {
  // a + 16p - b  (limbs of 16p: 2^55-304, 2^55-16, ...)
  Fp r;
  r.v[0] = a.v[0] + 36028797018963664ULL - b.v[0];
  for (int i = 1; i < 5; i++) r.v[i] = a.v[i] + 36028797018963952ULL - b.v[i];
  return fp_carry(r);
}
static inline Fp fp_neg(const Fp& a) { return fp_sub(fp_zero(), a); }
static inline Fp fp_mul(const Fp& a, const Fp& b) {
  const uint64_t *x = a.v, *y = b.v;
  uint64_t y1 = y[1] * 19, y2 = y[2] * 19, y3 = y[3] * 19, y4 = y[4] * 19;
  u128 c0 = (u128)x[0] * y[0] + (u128)x[4] * y1 + (u128)x[3] * y2 + (u128)x[2] * y3 + (u128)x[1] * y4;
  u128 c1 = (u128)x[1] * y[0] + (u128)x[0] * y[1] + (u128)x[4] * y2 + (u128)x[3] * y3 + (u128)x[2] * y4;
  u128 c2 = (u128)x[2] * y[0] + (u128)x[1] * y[1] + (u128)x[0] * y[2] + (u128)x[4] * y3 + (u128)x[3] * y4;
  u128 c3 = (u128)x[3] * y[0] + (u128)x[2] * y[1] + (u128)x[1] * y[2] + (u128)x[0] * y[3] + (u128)x[4] * y4;
  u128 c4 = (u128)x[4] * y[0] + (u128)x[3] * y[1] + (u128)x[2] * y[2] + (u128)x[1] * y[3] + (u128)x[0] * y[4];
  Fp r;
  c1 += (uint64_t)(c0 >> 51); r.v[0] = (uint64_t)c0 & M51;
  c2 += (uint64_t)(c1 >> 51); r.v[1] = (uint64_t)c1 & M51;
  c3 += (uint64_t)(c2 >> 51); r.v[2] = (uint64_t)c2 & M51;
  c4 += (uint64_t)(c3 >> 51); r.v[3] = (uint64_t)c3 & M51;
  uint64_t carry = (uint64_t)(c4 >> 51); r.v[4] = (uint64_t)c4 & M51;
  r.v[0] += carry * 19;
  r.v[1] += r.v[0] >> 51; r.v[0] &= M51;
  return r;
}
static inline Fp fp_sq(const Fp& a) { return fp_mul(a, a); }
static inline Fp fp_pow2k(Fp a, int k) {
  for (int i = 0; i < k; i++) a = fp_sq(a);
  return a;
}
// canonical 32-byte LE encoding
static inline void fp_to_bytes(const Fp& a, uint8_t out[32]) {
  Fp t = fp_carry(fp_carry(a));
  // compute t + 19 >> 255 to decide whether t >= p
  uint64_t q = (t.v[0] + 19) >> 51;
  q = (t.v[1] + q) >> 51; q = (t.v[2] + q) >> 51; q = (t.v[3] + q) >> 51; q = (t.v[4] + q) >> 51;
  t.v[0] += 19 * q;
  uint64_t c;
  c = t.v[0] >> 51; t.v[0] &= M51; t.v[1] += c;
  c = t.v[1] >> 51; t.v[1] &= M51; t.v[2] += c;
  c = t.v[2] >> 51; t.v[2] &= M51; t.v[3] += c;
  c = t.v[3] >> 51; t.v[3] &= M51; t.v[4] += c;
  t.v[4] &= M51;
  uint64_t w[4];
  w[0] = t.v[0] | (t.v[1] << 51);
  w[1] = (t.v[1] >> 13) | (t.v[2] << 38);
  w[2] = (t.v[2] >> 26) | (t.v[3] << 25);
  w[3] = (t.v[3] >> 39) | (t.v[4] << 12);
  memcpy(out, w, 32);
}
// loads the low 255 bits (top bit ignored), value may be >= p (non-canonical)
static inline Fp fp_from_bytes(const uint8_t b[32]) {
  uint64_t w[4];
  memcpy(w, b, 32);
  Fp r;
  r.v[0] = w[0] & M51;
  r.v[1] = ((w[0] >> 51) | (w[1] << 13)) & M51;
  r.v[2] = ((w[1] >> 38) | (w[2] << 26)) & M51;
  r.v[3] = ((w[2] >> 25) | (w[3] << 39)) & M51;
  r.v[4] = (w[3] >> 12) & M51;
  return r;
}
static inline bool fp_eq(const Fp& a, const Fp& b) {
  uint8_t x[32], y[32];
  fp_to_bytes(a, x);
  fp_to_bytes(b, y);
  return !memcmp(x, y, 32);
}
static inline bool fp_is_neg(const Fp& a) {
  uint8_t x[32];
  fp_to_bytes(a, x);
  return x[0] & 1;
}
static inline bool fp_is_zero(const Fp& a) { return fp_eq(a, fp_zero()); }
static inline Fp fp_abs(const Fp& a) { return fp_is_neg(a) ? fp_neg(a) : a; }

// a^(2^252-3) = a^((p-5)/8)
static inline Fp fp_pow_p58(const Fp& z) {
  Fp z2 = fp_sq(z), z9 = fp_mul(fp_pow2k(z2, 2), z), z11 = fp_mul(z9, z2);
  Fp z2_5_0 = fp_mul(fp_sq(z11), z9);
  Fp z2_10_0 = fp_mul(fp_pow2k(z2_5_0, 5), z2_5_0);
  Fp z2_20_0 = fp_mul(fp_pow2k(z2_10_0, 10), z2_10_0);
  Fp z2_40_0 = fp_mul(fp_pow2k(z2_20_0, 20), z2_20_0);
  Fp z2_50_0 = fp_mul(fp_pow2k(z2_40_0, 10), z2_10_0);
  Fp z2_100_0 = fp_mul(fp_pow2k(z2_50_0, 50), z2_50_0);
  Fp z2_200_0 = fp_mul(fp_pow2k(z2_100_0, 100), z2_100_0);
  Fp z2_250_0 = fp_mul(fp_pow2k(z2_200_0, 50), z2_50_0);
  return fp_mul(fp_pow2k(z2_250_0, 2), z);
}

static inline Fp fp_from_hex_le(const char* hex) {  // 64 hex chars, little-endian bytes
  uint8_t b[32];
  for (int i = 0; i < 32; i++) {
    auto h = [](char c) -> int { return c <= '9' ? c - '0' : c - 'a' + 10; };
    b[i] = (uint8_t)(h(hex[2 * i]) * 16 + h(hex[2 * i + 1]));
  }
  return fp_from_bytes(b);
}

// Curve constants (decimal values in SURVEY.md App. A.3/A.6), as LE byte hex.
struct Consts {
  Fp d, d2, sqrt_m1, invsqrt_a_minus_d, one_minus_d_sq, d_minus_one_sq, sqrt_ad_minus_one, bx, by;
  Consts() {
    d = fp_from_hex_le("a3785913ca4deb75abd841414d0a700098e879777940c78c73fe6f2bee6c0352");
    d2 = fp_add(d, d);
    sqrt_m1 = fp_from_hex_le("b0a00e4a271beec478e42fad0618432fa7d7fb3d99004d2b0bdfc14f8024832b");
    invsqrt_a_minus_d = fp_from_hex_le("ea405d80aafdc899be72415a17162f9d40d801fe917bc216a2fcafcf05896c78");
    one_minus_d_sq = fp_sub(fp_one(), fp_sq(d));
    Fp dm1 = fp_sub(d, fp_one());
    d_minus_one_sq = fp_sq(dm1);
    sqrt_ad_minus_one = fp_from_hex_le("1b2e7b49a0f6977ebd54781b0c8e9daffdd1f531c9fc3c0fac48832bbf316937");
    bx = fp_from_hex_le("1ad5258f602d56c9b2a7259560c72c695cdcd6fd31e2a4c0fe536ecdd3366921");
    by = fp_from_hex_le("5866666666666666666666666666666666666666666666666666666666666666");
  }
};
static inline const Consts& K() {
  static Consts k;
  return k;
}

// SQRT_RATIO_M1 (RFC 9496 4.2)
static inline bool sqrt_ratio_m1(const Fp& u, const Fp& v, Fp& out) {
  Fp v3 = fp_mul(fp_sq(v), v), v7 = fp_mul(fp_sq(v3), v);
  Fp r = fp_mul(fp_mul(u, v3), fp_pow_p58(fp_mul(u, v7)));
  Fp c = fp_mul(v, fp_sq(r));
  Fp nu = fp_neg(u);
  bool ok = fp_eq(c, u), f = fp_eq(c, nu), fi = fp_eq(c, fp_mul(nu, K().sqrt_m1));
  if (f || fi) r = fp_mul(r, K().sqrt_m1);
  out = fp_abs(r);
  return ok || f;
}

struct Pt {
  Fp X, Y, Z, T;
};
static inline Pt pt_identity() { return Pt{fp_zero(), fp_one(), fp_one(), fp_zero()}; }
static inline Pt pt_basepoint() { return Pt{K().bx, K().by, fp_one(), fp_mul(K().bx, K().by)}; }
static inline Pt pt_add(const Pt& p, const Pt& q) {
  Fp A = fp_mul(fp_sub(p.Y, p.X), fp_sub(q.Y, q.X));
  Fp B = fp_mul(fp_add(p.Y, p.X), fp_add(q.Y, q.X));
  Fp C = fp_mul(fp_mul(p.T, q.T), K().d2);
  Fp D = fp_mul(p.Z, q.Z);
  D = fp_add(D, D);
  Fp E = fp_sub(B, A), F = fp_sub(D, C), G = fp_add(D, C), H = fp_add(B, A);
  return Pt{fp_mul(E, F), fp_mul(G, H), fp_mul(F, G), fp_mul(E, H)};
}
static inline Pt pt_double(const Pt& p) {  // dbl-2008-hwcd, a = -1
  Fp A = fp_sq(p.X), B = fp_sq(p.Y), C = fp_sq(p.Z);
  C = fp_add(C, C);
  Fp D = fp_neg(A);
  Fp xy = fp_add(p.X, p.Y);
  Fp E = fp_sub(fp_sub(fp_sq(xy), A), B);
  Fp G = fp_add(D, B), F = fp_sub(G, C), H = fp_sub(D, B);
  return Pt{fp_mul(E, F), fp_mul(G, H), fp_mul(F, G), fp_mul(E, H)};
}
static inline Pt pt_neg(const Pt& p) { return Pt{fp_neg(p.X), p.Y, p.Z, fp_neg(p.T)}; }
static inline Pt pt_sub(const Pt& p, const Pt& q) { return pt_add(p, pt_neg(q)); }
static inline bool pt_eq(const Pt& a, const Pt& b) {  // ristretto equality
  return fp_eq(fp_mul(a.X, b.Y), fp_mul(a.Y, b.X)) || fp_eq(fp_mul(a.X, b.X), fp_mul(a.Y, b.Y));
}

// RistrettoPoint::compress
static inline void pt_compress(const Pt& p, uint8_t out[32]) {
  Fp u1 = fp_mul(fp_add(p.Z, p.Y), fp_sub(p.Z, p.Y));
  Fp u2 = fp_mul(p.X, p.Y);
  Fp inv;
  sqrt_ratio_m1(fp_one(), fp_mul(u1, fp_sq(u2)), inv);
  Fp d1 = fp_mul(inv, u1), d2 = fp_mul(inv, u2);
  Fp zinv = fp_mul(fp_mul(d1, d2), p.T);
  Fp x, y, dn;
  if (fp_is_neg(fp_mul(p.T, zinv))) {
    x = fp_mul(p.Y, K().sqrt_m1);
    y = fp_mul(p.X, K().sqrt_m1);
    dn = fp_mul(d1, K().invsqrt_a_minus_d);
  } else {
    x = p.X; y = p.Y; dn = d2;
  }
  if (fp_is_neg(fp_mul(x, zinv))) y = fp_neg(y);
  Fp s = fp_abs(fp_mul(dn, fp_sub(p.Z, y)));
  fp_to_bytes(s, out);
}
// CompressedRistretto::decompress; false on failure
static inline bool pt_decompress(const uint8_t b[32], Pt& out) {
  Fp s = fp_from_bytes(b);
  uint8_t chk[32];
  fp_to_bytes(s, chk);
  if (memcmp(chk, b, 32) != 0 || (b[0] & 1)) return false;  // non-canonical or negative
  Fp ss = fp_sq(s), u1 = fp_sub(fp_one(), ss), u2 = fp_add(fp_one(), ss), u2s = fp_sq(u2);
  Fp v = fp_sub(fp_neg(fp_mul(K().d, fp_sq(u1))), u2s);
  Fp inv;
  bool ok = sqrt_ratio_m1(fp_one(), fp_mul(v, u2s), inv);
  Fp dx = fp_mul(inv, u2), dy = fp_mul(fp_mul(inv, dx), v);
  Fp x = fp_abs(fp_mul(fp_add(s, s), dx));
  Fp y = fp_mul(u1, dy);
  Fp t = fp_mul(x, y);
  if (!ok || fp_is_neg(t) || fp_is_zero(y)) return false;
  out = Pt{x, y, fp_one(), t};
  return true;
}
// Elligator MAP
static inline Pt pt_elligator(const Fp& t) {
  const Consts& k = K();
  Fp r = fp_mul(k.sqrt_m1, fp_sq(t));
  Fp u = fp_mul(fp_add(r, fp_one()), k.one_minus_d_sq);
  Fp v = fp_mul(fp_sub(fp_neg(fp_one()), fp_mul(r, k.d)), fp_add(r, k.d));
  Fp s;
  bool sq = sqrt_ratio_m1(u, v, s);
  Fp sp = fp_neg(fp_abs(fp_mul(s, t)));
  if (!sq) s = sp;
  Fp c = sq ? fp_neg(fp_one()) : r;
  Fp N = fp_sub(fp_mul(fp_mul(c, fp_sub(r, fp_one())), k.d_minus_one_sq), v);
  Fp w0 = fp_mul(fp_add(s, s), v), w1 = fp_mul(N, k.sqrt_ad_minus_one);
  Fp ss = fp_sq(s), w2 = fp_sub(fp_one(), ss), w3 = fp_add(fp_one(), ss);
  return Pt{fp_mul(w0, w3), fp_mul(w2, w1), fp_mul(w1, w3), fp_mul(w0, w2)};
}
// RistrettoPoint::from_uniform_bytes
static inline Pt pt_from_uniform_bytes(const uint8_t b[64]) {
  return pt_add(pt_elligator(fp_from_bytes(b)), pt_elligator(fp_from_bytes(b + 32)));
}

// scalar (canonical 32 bytes) * point, simple 4-bit fixed window, vartime
static inline Pt pt_mul_bytes(const uint8_t k[32], const Pt& p) {
  Pt tbl[16];
  tbl[0] = pt_identity();
  for (int i = 1; i < 16; i++) tbl[i] = pt_add(tbl[i - 1], p);
  Pt acc = pt_identity();
  bool started = false;
  for (int i = 63; i >= 0; i--) {
    int d = (k[i / 2] >> ((i & 1) * 4)) & 15;
    if (started)
      for (int j = 0; j < 4; j++) acc = pt_double(acc);
    if (d) {
      acc = started ? pt_add(acc, tbl[d]) : tbl[d];
      started = true;
    }
  }
  return acc;
}
static inline Pt pt_mul(const Fl& k, const Pt& p) {
  uint8_t kb[32];
  fl_to_bytes(k, kb);
  return pt_mul_bytes(kb, p);
}

// GroupElement::vartime_multiscalar_mul (group.rs:98-117).  dalek picks Straus
// below 190 terms and Pippenger above; only the group element matters for
// parity, so this uses per-term windowed mul below 32 terms and a bucket
// (Pippenger) method above, window chosen by size like dalek's.
static inline Pt pt_msm(const Fl* scalars, const Pt* points, size_t n) {
  if (n == 0) return pt_identity();
  if (n < 32) {
    Pt acc = pt_identity();
    for (size_t i = 0; i < n; i++) acc = pt_add(acc, pt_mul(scalars[i], points[i]));
    return acc;
  }
  int w = n < 500 ? 6 : (n < 800 ? 7 : 8);
  std::vector<uint8_t> kb(32 * n);
  for (size_t i = 0; i < n; i++) fl_to_bytes(scalars[i], &kb[32 * i]);
  int nwin = (253 + w - 1) / w;
  size_t nb = ((size_t)1 << w) - 1;
  std::vector<Pt> buckets(nb);
  std::vector<uint8_t> used(nb);
  Pt total = pt_identity();
  for (int win = nwin - 1; win >= 0; win--) {
    for (int j = 0; j < w; j++) total = pt_double(total);
    std::fill(used.begin(), used.end(), 0);
    int bit0 = win * w;
    for (size_t i = 0; i < n; i++) {
      const uint8_t* k = &kb[32 * i];
      uint32_t d = 0;
      for (int b = 0; b < w; b++) {
        int bit = bit0 + b;
        if (bit < 256) d |= (uint32_t)((k[bit >> 3] >> (bit & 7)) & 1) << b;
      }
      if (!d) continue;
      if (used[d - 1]) buckets[d - 1] = pt_add(buckets[d - 1], points[i]);
      else { buckets[d - 1] = points[i]; used[d - 1] = 1; }
    }
    Pt run = pt_identity(), sum = pt_identity();
    for (size_t b = nb; b-- > 0;) {
      if (used[b]) run = pt_add(run, buckets[b]);
      sum = pt_add(sum, run);
    }
    total = pt_add(total, sum);
  }
  return total;
}

}  // namespace orc
