// Keccak-f[1600] with AVX-512 (compiled with -mavx512f -mavx512vl, called only after a CPUID check: hmerlin.hpp).
// Every Fiat-Shamir challenge and every RandomTape draw is one or two of these permutations, strictly one after the
// other -- a proof sharded over 8 GPUs is bounded by the 2^(s/2) blind draws of its commitment (host/tape.hpp) -- so the
// single-state latency is what matters, not multi-buffer throughput.
// Layout: plane y = (A[0][y] .. A[4][y]) in the low five qwords of one register.  theta and rho work on planes; the planes
// are then transposed to columns Q_x = (A[x][0] .. A[x][4]), which turns pi into pure blends (plane y' of B takes lane x'
// from column (x' + 3 y') mod 5, at the same lane); chi is three-operand logic on rotated copies of a plane.
#include <immintrin.h>
#include <stdint.h>

namespace sbh {

void keccakf_avx512(uint64_t s[25]) {
  static const uint64_t RC[24] = {0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
                                  0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
                                  0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
                                  0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
                                  0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
                                  0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
  const __mmask8 m5 = 0x1f;
  __m512i P0 = _mm512_maskz_loadu_epi64(m5, s), P1 = _mm512_maskz_loadu_epi64(m5, s + 5), P2 = _mm512_maskz_loadu_epi64(m5, s + 10),
          P3 = _mm512_maskz_loadu_epi64(m5, s + 15), P4 = _mm512_maskz_loadu_epi64(m5, s + 20);
  const __m512i idx_m1 = _mm512_setr_epi64(4, 0, 1, 2, 3, 5, 6, 7), idx_p1 = _mm512_setr_epi64(1, 2, 3, 4, 0, 5, 6, 7),
                idx_p2 = _mm512_setr_epi64(2, 3, 4, 0, 1, 5, 6, 7);
  // rho offsets r[x][y], one vector per plane y
  const __m512i R0 = _mm512_setr_epi64(0, 1, 62, 28, 27, 0, 0, 0), R1 = _mm512_setr_epi64(36, 44, 6, 55, 20, 0, 0, 0),
                R2 = _mm512_setr_epi64(3, 10, 43, 25, 39, 0, 0, 0), R3 = _mm512_setr_epi64(41, 45, 15, 21, 8, 0, 0, 0),
                R4 = _mm512_setr_epi64(18, 2, 61, 56, 14, 0, 0, 0);
  const __m512i ia = _mm512_setr_epi64(0, 1, 8, 9, 0, 0, 0, 0), ib = _mm512_setr_epi64(2, 3, 10, 11, 0, 0, 0, 0),
                ic = _mm512_setr_epi64(4, 5, 12, 13, 0, 0, 0, 0);
  const __m512i k0 = _mm512_set1_epi64(0), k1 = _mm512_set1_epi64(1), k2 = _mm512_set1_epi64(2), k3 = _mm512_set1_epi64(3),
                k4 = _mm512_set1_epi64(4);
#define SB_BL5(a, b, c, d, e) \
  _mm512_mask_blend_epi64(0x10, _mm512_mask_blend_epi64(0x08, _mm512_mask_blend_epi64(0x04, _mm512_mask_blend_epi64(0x02, a, b), c), d), e)
#define SB_CHI(B) _mm512_ternarylogic_epi64(B, _mm512_permutexvar_epi64(idx_p1, B), _mm512_permutexvar_epi64(idx_p2, B), 0xD2)
  for (int round = 0; round < 24; round++) {
    // theta: C = column parities, D[x] = C[x-1] ^ rol(C[x+1], 1)
    const __m512i C = _mm512_ternarylogic_epi64(_mm512_ternarylogic_epi64(P0, P1, P2, 0x96), P3, P4, 0x96);
    const __m512i D = _mm512_xor_si512(_mm512_permutexvar_epi64(idx_m1, C), _mm512_rol_epi64(_mm512_permutexvar_epi64(idx_p1, C), 1));
    // rho
    P0 = _mm512_rolv_epi64(_mm512_xor_si512(P0, D), R0);
    P1 = _mm512_rolv_epi64(_mm512_xor_si512(P1, D), R1);
    P2 = _mm512_rolv_epi64(_mm512_xor_si512(P2, D), R2);
    P3 = _mm512_rolv_epi64(_mm512_xor_si512(P3, D), R3);
    P4 = _mm512_rolv_epi64(_mm512_xor_si512(P4, D), R4);
    // planes -> columns: Q_x[y] = P_y[x]
    const __m512i t0 = _mm512_unpacklo_epi64(P0, P1), t1 = _mm512_unpackhi_epi64(P0, P1);
    const __m512i t2 = _mm512_unpacklo_epi64(P2, P3), t3 = _mm512_unpackhi_epi64(P2, P3);
    __m512i Q0 = _mm512_permutex2var_epi64(t0, ia, t2), Q1 = _mm512_permutex2var_epi64(t1, ia, t3),
            Q2 = _mm512_permutex2var_epi64(t0, ib, t2), Q3 = _mm512_permutex2var_epi64(t1, ib, t3),
            Q4 = _mm512_permutex2var_epi64(t0, ic, t2);
    Q0 = _mm512_mask_permutexvar_epi64(Q0, 0x10, k0, P4);
    Q1 = _mm512_mask_permutexvar_epi64(Q1, 0x10, k1, P4);
    Q2 = _mm512_mask_permutexvar_epi64(Q2, 0x10, k2, P4);
    Q3 = _mm512_mask_permutexvar_epi64(Q3, 0x10, k3, P4);
    Q4 = _mm512_mask_permutexvar_epi64(Q4, 0x10, k4, P4);
    // pi: B[y][2x+3y] = A[x][y], i.e. plane y' of B takes lane x' from column (x' + 3 y') mod 5
    const __m512i B0 = SB_BL5(Q0, Q1, Q2, Q3, Q4), B1 = SB_BL5(Q3, Q4, Q0, Q1, Q2), B2 = SB_BL5(Q1, Q2, Q3, Q4, Q0),
                  B3 = SB_BL5(Q4, Q0, Q1, Q2, Q3), B4 = SB_BL5(Q2, Q3, Q4, Q0, Q1);
    // chi, iota
    P0 = SB_CHI(B0); P1 = SB_CHI(B1); P2 = SB_CHI(B2); P3 = SB_CHI(B3); P4 = SB_CHI(B4);
    P0 = _mm512_xor_si512(P0, _mm512_maskz_set1_epi64(0x01, (long long)RC[round]));
  }
#undef SB_BL5
#undef SB_CHI
  _mm512_mask_storeu_epi64(s, m5, P0);
  _mm512_mask_storeu_epi64(s + 5, m5, P1);
  _mm512_mask_storeu_epi64(s + 10, m5, P2);
  _mm512_mask_storeu_epi64(s + 15, m5, P3);
  _mm512_mask_storeu_epi64(s + 20, m5, P4);
}

}  // namespace sbh
