// Edwards25519 point arithmetic + ristretto255 compress for sm_100a.
// The reference reaches these through curve25519-dalek's RistrettoPoint
// (Spartan/src/group.rs:6-7): `+`, `compress()` and `vartime_multiscalar_mul`
// (group.rs:98-117).  Only group-element VALUES matter for parity (compress is
// canonical, RFC 9496 4.3.2), so the coordinate system and the MSM algorithm
// are free; here: extended coordinates (X:Y:Z:T), a = -1, and affine "Niels"
// table entries (y+x, y-x, 2dxy) for the fixed generator bases.
#pragma once
#include "fp.cuh"

namespace sb {

struct ge_t {  // extended
  fp_t X, Y, Z, T;
};
struct niels_t {  // affine precomputed: 96 bytes
  fp_t yplusx, yminusx, xy2d;
};

SB_HD ge_t ge_identity() {
  ge_t r;
  r.X = fp_zero(); r.Y = fp_one(); r.Z = fp_one(); r.T = fp_zero();
  return r;
}
SB_HD niels_t niels_identity() {
  niels_t n;
  n.yplusx = fp_one(); n.yminusx = fp_one(); n.xy2d = fp_zero();
  return n;
}
// full addition, 9M (add-2008-hwcd-3 with k = 2d)
template <class F>
SB_HD ge_t ge_add_t(const ge_t& p, const ge_t& q) {
  fp_t A = F::mul(fp_sub(p.Y, p.X), fp_sub(q.Y, q.X));
  fp_t B = F::mul(fp_add(p.Y, p.X), fp_add(q.Y, q.X));
  fp_t C = F::mul(F::mul(p.T, q.T), fp_2d());
  fp_t D = F::mul(p.Z, q.Z);
  D = fp_add(D, D);
  fp_t E = fp_sub(B, A), Fv = fp_sub(D, C), G = fp_add(D, C), H = fp_add(B, A);
  ge_t r;
  r.X = F::mul(E, Fv); r.Y = F::mul(G, H); r.Z = F::mul(Fv, G); r.T = F::mul(E, H);
  return r;
}
// mixed addition with an affine Niels entry, 7M; sign = true adds the negation
template <class F>
SB_HD ge_t ge_madd_t(const ge_t& p, const niels_t& q, bool neg) {
  fp_t ypx = fp_add(p.Y, p.X), ymx = fp_sub(p.Y, p.X);
  // -Q = (y-x, y+x, -2dxy): limb-wise selects (a select between the two structs would go through local memory)
  fp_t qa, qb;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    qa.v[i] = neg ? q.yplusx.v[i] : q.yminusx.v[i];
    qb.v[i] = neg ? q.yminusx.v[i] : q.yplusx.v[i];
  }
  fp_t A = F::mul(ymx, qa);
  fp_t B = F::mul(ypx, qb);
  fp_t C = F::mul(p.T, q.xy2d);
  fp_t D = fp_add(p.Z, p.Z);
  fp_t E = fp_sub(B, A), H = fp_add(B, A);
  const fp_t dmc = fp_sub(D, C), dpc = fp_add(D, C);
  fp_t Fv, G;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    Fv.v[i] = neg ? dpc.v[i] : dmc.v[i];
    G.v[i] = neg ? dmc.v[i] : dpc.v[i];
  }
  ge_t r;
  r.X = F::mul(E, Fv); r.Y = F::mul(G, H); r.Z = F::mul(Fv, G); r.T = F::mul(E, H);
  return r;
}
// doubling (dbl-2008-hwcd, a = -1), 4S + 4M
template <class F>
SB_HD ge_t ge_double_t(const ge_t& p) {
  fp_t A = F::sqr(p.X), B = F::sqr(p.Y), C = F::sqr(p.Z);
  C = fp_add(C, C);
  fp_t D = fp_neg(A);
  fp_t xy = fp_add(p.X, p.Y);
  fp_t E = fp_sub(fp_sub(F::sqr(xy), A), B);
  fp_t G = fp_add(D, B), Fv = fp_sub(G, C), H = fp_sub(D, B);
  ge_t r;
  r.X = F::mul(E, Fv); r.Y = F::mul(G, H); r.Z = F::mul(Fv, G); r.T = F::mul(E, H);
  return r;
}
SB_HD ge_t ge_add(const ge_t& p, const ge_t& q) { return ge_add_t<FpInl>(p, q); }
SB_HD ge_t ge_madd(const ge_t& p, const niels_t& q, bool neg) { return ge_madd_t<FpInl>(p, q, neg); }
SB_HD ge_t ge_double(const ge_t& p) { return ge_double_t<FpInl>(p); }
#if defined(__CUDACC__)
// one out-of-line copy per kernel (see FpCall): for the tree reductions and table builds
static __device__ __noinline__ ge_t ge_add_call(ge_t p, ge_t q) { return ge_add_t<FpInl>(p, q); }
static __device__ __noinline__ ge_t ge_madd_call(ge_t p, niels_t q, bool neg) { return ge_madd_t<FpInl>(p, q, neg); }
static __device__ __noinline__ ge_t ge_double_call(ge_t p) { return ge_double_t<FpInl>(p); }
#endif
// affine (x, y) -> Niels
SB_HD niels_t niels_from_affine(const fp_t& x, const fp_t& y) {
  niels_t n;
  n.yplusx = fp_add(y, x);
  n.yminusx = fp_sub(y, x);
  n.xy2d = fp_mul(fp_mul(x, y), fp_2d());
  return n;
}

SB_HD fp_t fp_sqrt_m1() { return fp_set(0x4a0ea0b0u, 0xc4ee1b27u, 0xad2fe478u, 0x2f431806u, 0x3dfbd7a7u, 0x2b4d0099u, 0x4fc1df0bu, 0x2b832480u); }
SB_HD fp_t fp_invsqrt_a_minus_d() { return fp_set(0x805d40eau, 0x99c8fdaau, 0x5a4172beu, 0x9d2f1617u, 0xfe01d840u, 0x16c27b91u, 0xcfaffca2u, 0x786c8905u); }

// SQRT_RATIO_M1(1, v): returns abs(r) with r^2 * v in {1, -1, -sqrt(-1)...}; only the
// "inverse square root" value is needed by compress.
template <class F>
SB_HD fp_t fp_invsqrt_t(const fp_t& v) {
  fp_t v3 = F::mul(F::sqr(v), v), v7 = F::mul(F::sqr(v3), v);
  fp_t r = F::mul(v3, fp_pow_p58_t<F>(v7));
  fp_t c = F::mul(v, F::sqr(r));
  fp_t one = fp_one(), m1 = fp_neg(one);
  bool f = fp_eq(c, m1), fi = fp_eq(c, fp_neg(fp_sqrt_m1()));
  if (f || fi) r = F::mul(r, fp_sqrt_m1());
  return fp_abs(r);
}
// RistrettoPoint::compress -> canonical 32 bytes (8 LE words)
template <class F>
SB_HD void ge_compress_t(const ge_t& p, uint32_t out[8]) {
  fp_t u1 = F::mul(fp_add(p.Z, p.Y), fp_sub(p.Z, p.Y));
  fp_t u2 = F::mul(p.X, p.Y);
  fp_t inv = fp_invsqrt_t<F>(F::mul(u1, F::sqr(u2)));
  fp_t d1 = F::mul(inv, u1), d2 = F::mul(inv, u2);
  fp_t zinv = F::mul(F::mul(d1, d2), p.T);
  fp_t x, y, dn;
  if (fp_is_neg(F::mul(p.T, zinv))) {
    x = F::mul(p.Y, fp_sqrt_m1());
    y = F::mul(p.X, fp_sqrt_m1());
    dn = F::mul(d1, fp_invsqrt_a_minus_d());
  } else {
    x = p.X; y = p.Y; dn = d2;
  }
  if (fp_is_neg(F::mul(x, zinv))) y = fp_neg(y);
  fp_t s = fp_canon(fp_abs(F::mul(dn, fp_sub(p.Z, y))));
#pragma unroll
  for (int i = 0; i < 8; i++) out[i] = s.v[i];
}
SB_HD fp_t fp_invsqrt(const fp_t& v) { return fp_invsqrt_t<FpInl>(v); }
SB_HD void ge_compress(const ge_t& p, uint32_t out[8]) { ge_compress_t<FpInl>(p, out); }

// Signed radix-2^c recoding used by the fixed-base MSM: digit w of the canonical scalar at kp
// (8 LE words), in (-2^(c-1), 2^(c-1)], with the running carry.  floor(252/c)+1 windows cover
// every scalar < l without a final carry (l = 2^252 + small).
SB_HD int32_t msm_digit(const uint32_t* kp, uint32_t w, uint32_t c, uint32_t& carry) {
  const uint32_t half = 1u << (c - 1), mask = (1u << c) - 1;
  uint32_t bit = w * c, limb = bit >> 5, sh = bit & 31;
  uint32_t lo = kp[limb], hi = limb < 7 ? kp[limb + 1] : 0u;
  uint32_t raw = (uint32_t)((((uint64_t)hi << 32) | lo) >> sh) & mask;
  uint32_t d = raw + carry;
  bool neg = d > half;
  carry = neg ? 1u : 0u;
  return neg ? -(int32_t)((1u << c) - d) : (int32_t)d;
}
SB_HD uint32_t msm_windows(uint32_t c) { return 252u / c + 1u; }

}  // namespace sb
