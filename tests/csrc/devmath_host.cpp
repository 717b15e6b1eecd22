// Host build of the __host__ __device__ limb arithmetic in otti_b200/csrc/{fl,fp,ge}.cuh so the
// exact code the sm_100a kernels run can be checked against the oracle on a box without a GPU.
// Test shim only; not part of the product library.
#include <cstring>
#include "../../otti_b200/csrc/ge.cuh"
using namespace sb;
static fl_t ldl(const uint8_t* p) { fl_t x; memcpy(x.v, p, 32); return x; }
static fp_t ldp(const uint8_t* p) { fp_t x; memcpy(x.v, p, 32); return x; }
extern "C" {
void dm_fl_binop(int op, const uint8_t* a, const uint8_t* b, uint8_t* out) {
  fl_t r = op == 0 ? fl_add(ldl(a), ldl(b)) : op == 1 ? fl_sub(ldl(a), ldl(b)) : fl_mul(ldl(a), ldl(b));
  memcpy(out, r.v, 32);
}
void dm_fl_to_mont(const uint8_t* a, uint8_t* out) { fl_t r = fl_to_mont(ldl(a)); memcpy(out, r.v, 32); }
void dm_fl_from_mont(const uint8_t* a, uint8_t* out) { fl_t r = fl_from_mont(ldl(a)); memcpy(out, r.v, 32); }
int dm_fl_is_canonical(const uint8_t* a) { return fl_is_canonical(ldl(a)); }
void dm_fl_neg(const uint8_t* a, uint8_t* out) { fl_t r = fl_neg(ldl(a)); memcpy(out, r.v, 32); }
// F_p on loose 256-bit inputs; outputs canonical
void dm_fp_binop(int op, const uint8_t* a, const uint8_t* b, uint8_t* out) {
  fp_t r = op == 0 ? fp_add(ldp(a), ldp(b)) : op == 1 ? fp_sub(ldp(a), ldp(b)) : fp_mul(ldp(a), ldp(b));
  r = fp_canon(r);
  memcpy(out, r.v, 32);
}
void dm_fp_invert(const uint8_t* a, uint8_t* out) { fp_t r = fp_canon(fp_invert(ldp(a))); memcpy(out, r.v, 32); }
void dm_fp_canon(const uint8_t* a, uint8_t* out) { fp_t r = fp_canon(ldp(a)); memcpy(out, r.v, 32); }
// k * P through the MSM digit recoding with window c, table entries made on the fly exactly as
// msm_table_kernel makes them (multiples of 2^(c*w) P, normalised to affine Niels); compress.
void dm_fixed_base_mul(const uint8_t* xy64, const uint8_t* k32, uint32_t c, uint8_t* out32) {
  ge_t P; P.X = ldp(xy64); P.Y = ldp(xy64 + 32); P.Z = fp_one(); P.T = fp_mul(P.X, P.Y);
  uint32_t kw[8]; memcpy(kw, k32, 32);
  ge_t acc = ge_identity();
  uint32_t carry = 0, W = msm_windows(c);
  for (uint32_t w = 0; w < W; w++) {
    int32_t d = msm_digit(kw, w, c, carry);
    if (d != 0) {
      uint32_t mag = d < 0 ? -d : d;
      ge_t M = ge_identity();
      for (int bit = 16; bit >= 0; bit--) { M = ge_double(M); if ((mag >> bit) & 1u) M = ge_add(M, P); }
      fp_t zi = fp_invert(M.Z);
      niels_t e = niels_from_affine(fp_mul(M.X, zi), fp_mul(M.Y, zi));
      acc = ge_madd(acc, e, d < 0);
    }
    for (uint32_t i = 0; i < c; i++) P = ge_double(P);
  }
  uint32_t o[8]; ge_compress(acc, o); memcpy(out32, o, 32);
}
int dm_final_carry(const uint8_t* k32, uint32_t c) {
  uint32_t kw[8]; memcpy(kw, k32, 32);
  uint32_t carry = 0, W = msm_windows(c);
  for (uint32_t w = 0; w < W; w++) msm_digit(kw, w, c, carry);
  return (int)carry;
}
// (P + Q) compressed, P and Q affine
void dm_add_compress(const uint8_t* p64, const uint8_t* q64, uint8_t* out32) {
  ge_t P; P.X = ldp(p64); P.Y = ldp(p64 + 32); P.Z = fp_one(); P.T = fp_mul(P.X, P.Y);
  ge_t Q; Q.X = ldp(q64); Q.Y = ldp(q64 + 32); Q.Z = fp_one(); Q.T = fp_mul(Q.X, Q.Y);
  uint32_t o[8]; ge_compress(ge_add(P, Q), o); memcpy(out32, o, 32);
}
}
