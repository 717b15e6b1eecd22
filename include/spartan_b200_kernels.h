/*
 * spartan_b200_kernels.h -- per-kernel entry points of libspartan_b200.so.
 *
 * The reference exposes its hot loops only as Rust-internal functions; these
 * hooks run ONE sm_100a kernel (or one fused step) each on host buffers so the
 * parity tests can compare every kernel with the oracle restatement of the
 * loop it replaces.  All buffers are HOST pointers; scalars are 32-byte
 * Montgomery limbs (the reference's in-memory `Scalar`, ristretto255.rs:200)
 * unless the name says "canon".  Return codes as in spartan_b200.h.
 */
#ifndef SPARTAN_B200_KERNELS_H
#define SPARTAN_B200_KERNELS_H

#include "spartan_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Scalar::invert of one canonical scalar on the HOST (no device needed): the inversion the prover's Fiat-Shamir tail uses
 * (fermat = 0: extended Euclid) or a^(l-2) as the reference computes it (fermat != 0); 0 -> 0.    ristretto255.rs:548-602 */
int sb200_k_host_invert(const uint8_t* in_canon, uint8_t* out_canon, int fermat);
/* Scalar add / sub / mul (op 0/1/2), elementwise over n                 ristretto255.rs:702-755 */
int sb200_k_fl_binop(int op, const uint8_t* a, const uint8_t* b, uint8_t* out, size_t n);
/* Coordinate field F_p, p = 2^255-19 (curve25519-dalek FieldElement behind group.rs:6-7): out = canonical
 * bytes of a op b over n elements given as ANY 256-bit little-endian integers (op 0 add, 1 sub, 2 mul,
 * 3 a^2, 4 a^-1) */
int sb200_k_fp_binop(int op, const uint8_t* a, const uint8_t* b, uint8_t* out, size_t n);
/* Scalar::from_bytes over n canonical values; *any_invalid = 1 if one is >= l   ristretto255.rs:398-422 */
int sb200_k_to_mont(const uint8_t* canon, uint8_t* out, size_t n, int* any_invalid);
/* Scalar::to_bytes over n values                                           ristretto255.rs:426-438 */
int sb200_k_from_mont(const uint8_t* in, uint8_t* canon_out, size_t n);
/* EqPolynomial::new(r).evals(), out holds 2^ell scalars                    dense_mlpoly.rs:67-83 */
int sb200_k_eq_evals(const uint8_t* r, uint32_t ell, uint8_t* out);
/* R1CSInstance::multiply_vec: z has 2*num_vars scalars, outs num_cons each r1csinstance.rs:248-262 */
int sb200_k_multiply_vec(const sb200_instance* inst, const uint8_t* z, uint8_t* Az, uint8_t* Bz, uint8_t* Cz);
/* compute_eval_table_sparse x3 combined with r_A,r_B,r_C (rabc = 3 scalars); out 2*num_vars
 *                                                    r1csinstance.rs:264-278, r1csproof.rs:268-279 */
int sb200_k_eval_table_abc(const sb200_instance* inst, const uint8_t* rx_evals, const uint8_t* rabc, uint8_t* out);
/* One fused sum-check round over ntab (4: cubic with additive term, 2: quadratic) tables of `len`
 * scalars each: if r != NULL the tables are first bound at r (bound_poly_var_top) to len/2, then
 * the round evaluations e0,e2,e3 of the resulting tables are written to evals3 (e3 = 0 for ntab 2).
 * Tables are updated in place.                          sumcheck.rs:453-484,618-674, dense_mlpoly.rs:214-221 */
int sb200_k_sumcheck_round(uint8_t* const* tables, int ntab, size_t len, const uint8_t* r, uint8_t* evals3);
/* DensePolynomial::bound: Z has n scalars viewed as Ls x (n/Ls); out n/Ls    dense_mlpoly.rs:205-212 */
int sb200_k_poly_bound(const uint8_t* Z, size_t n, const uint8_t* Lv, size_t Ls, uint8_t* out);
/* The same DensePolynomial::bound for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * UNLIKE the hooks of this header, d_Z (n scalars), d_L (Ls scalars) and d_out (n/Ls scalars) are DEVICE pointers of the
 * caller, 16-byte aligned, Montgomery limbs as above; the two kernels are enqueued on `stream` (a cudaStream_t, NULL = the
 * default stream) and the call returns without synchronising.  It calls no cudaMalloc / cudaFree or other allocating
 * routine: the per-chunk partial sums live in a fixed 8 MiB array in the library's device image (held by every GPU that
 * loads the library until the process exits), whose address is looked up once per device on the first call there.
 * Limit: n / Ls <= 4096 columns (any Ls; n < 2^32) -- larger shapes get SB200_ERR_ARG and a message naming the limit.
 * Every call on a device shares that array: calls that overlap on the GPU race and must be serialised by the caller.
 * Nothing in it needs clearing between calls.  Output bytes equal sb200_k_poly_bound's. */
int sb200_poly_bound_noalloc(const void* d_Z, size_t n, const void* d_L, size_t Ls, void* d_out, void* stream);
/* DotProductProofLog::compute_dotproduct, n <= 65536                        nizk/mod.rs:430-433 */
int sb200_k_dot(const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out);
/* DensePolynomial::commit_inner: L rows of n/L canonical scalars, blinds canonical; out L x 32
 * compressed points                                 dense_mlpoly.rs:163-176, commitments.rs:79-84 */
int sb200_k_commit_rows(const sb200_gens* gens, const uint8_t* Z_canon, size_t n, const uint8_t* blinds_canon, size_t L,
                        uint8_t* out32);
/* Generators as derived by MultiCommitGens::new (compressed).  which: 0 gens_n.G[i], 1 gens_n.h,
 * 2 gens_1.G[0], 3 gens_3.G[i], 4 gens_3.h, 5 gens_4.G[i], 6 gens_4.h         commitments.rs:15-58 */
int sb200_k_gens_point(const sb200_gens* gens, int which, size_t i, uint8_t out32[32]);
size_t sb200_k_gens_n(const sb200_gens* gens);
/* window width (bits) of the device fixed-base tables built for these generators */
uint32_t sb200_k_gens_window(const sb200_gens* gens);
/* Host Pedersen commitment used in the Fiat-Shamir tail: x*G1[0] + blind*h (canonical inputs) commitments.rs:65-70 */
int sb200_k_commit_scalar(const sb200_gens* gens, const uint8_t x_canon[32], const uint8_t blind_canon[32], uint8_t out32[32]);
/* merlin: Transcript::new(label); append_message(mlabel,msg); challenge_bytes(clabel,out)   transcript.rs:13-37 */
int sb200_k_merlin(const uint8_t* label, size_t ll, const char* mlabel, const uint8_t* msg, size_t ml, const char* clabel,
                   uint8_t* out, size_t n);
/* MultiSparseMatPolynomialAsDense: comb_ops (16 N scalars) and comb_mem (2 num_mem_cells), canonical 32-byte values
 * sparse_mlpoly.rs:371-422; dims_out = {N, num_mem_cells}; ops_out / mem_out may be NULL */
int sb200_k_snark_dense(const sb200_instance* inst, uint64_t dims_out[2], uint8_t* ops_out, uint8_t* mem_out);
/* kernels launched since the last call to sb200_nizk_prove began (for bench.py's gpu_launches) */
uint64_t sb200_k_launch_count(void);
/* Per-kernel device time: while enabled, every kernel launch of the library is bracketed by CUDA events on its stream
 * (a few microseconds per launch, so never inside a timed region).  sb200_k_kprof_report writes "kernel,launches,total_ms"
 * lines, longest first, for everything launched since the last report, and clears the records. */
int sb200_k_kprof_enable(int on);
int sb200_k_kprof_report(char* out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
