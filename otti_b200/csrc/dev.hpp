// Host-callable launchers of the sm_100a kernels (kernels.cu).  Pointers are
// device pointers unless the name says host.  Everything runs on the stream
// passed in; nothing here synchronises unless stated.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "fl.cuh"

namespace sb {

struct niels_t;

// ---- fixed-base tables (one per NIZKGens) -------------------------------------------------
struct MsmTable {
  niels_t* tbl = nullptr;  // [nbases][windows][1 << (c-1)] affine Niels entries
  uint32_t nbases = 0;     // R + 2 : G[0..R-1], h, Q (= gens_1.G[0])
  uint32_t c = 0;          // window bits
  uint32_t windows = 0;    // floor(252/c) + 1
  size_t entries_per_base() const { return (size_t)windows << (c - 1); }
};
// affine_xy: host array, nbases x 64 bytes (x, y canonical LE).  Allocates and fills t.tbl.
cudaError_t msm_table_build(MsmTable& t, const uint8_t* affine_xy_host, uint32_t nbases, uint32_t c, cudaStream_t s);
void msm_table_free(MsmTable& t);

// rows x n fixed-base MSM over bases [0, n) plus optional per-row scalars for base n_h (h) and
// n_q (Q).  scalars: canonical integers < l, row-major with row stride `stride` (in scalars).
// out32: rows x 32-byte compressed ristretto points.  `scratch` must hold rows*splits*128 bytes.
struct MsmPlan {
  uint32_t rows, n, splits, block, tpt;  // grid (splits, rows) x block threads, tpt terms per thread
};
MsmPlan msm_plan(uint32_t rows, uint32_t n);
cudaError_t msm_fixed(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t rows, uint32_t n,
                      const fl_t* extra_h, const fl_t* extra_q, uint32_t h_index, uint32_t q_index, void* scratch,
                      uint8_t* out32, cudaStream_t s, cudaEvent_t* ev = nullptr);  // ev[0..1] bracket the main kernel
// Two-step form: row sums without any blind term, then (partials + extra_h[row]*base[h_index]) -> compressed.
cudaError_t msm_fixed_rows(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t rows, uint32_t n, void* scratch,
                           cudaStream_t s, cudaEvent_t* ev = nullptr);
// rows [row0, row0 + nrows) of a total_rows-row job (same plan and scratch layout as msm_fixed_rows over all rows)
cudaError_t msm_fixed_rows_part(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t total_rows, uint32_t row0,
                                uint32_t nrows, uint32_t n, void* scratch, cudaStream_t s);
cudaError_t msm_fixed_finish(const MsmTable& t, uint32_t rows, uint32_t n, const fl_t* extra_h, uint32_t h_index,
                             const void* scratch, uint8_t* out32, cudaStream_t s);
size_t msm_scratch_bytes(uint32_t rows, uint32_t n);
// Latency-oriented variant for a few rows (bullet reduction, one-row commitments): every (term, window)
// pair is a unit of work, the sum is left UNCOMPRESSED in out_ext[rows] (extended coordinates X,Y,Z,T as
// 4 x 8 LE words each, any representative < 2^256).  counters: rows + 1 zeroed words, self-resetting.
// If host_flag != NULL, out_ext may be mapped host memory and *host_flag = seq is written after all rows.
cudaError_t msm_small(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t rows, uint32_t n, void* scratch,
                      unsigned int* counters, void* out_ext, uint32_t* host_flag, uint32_t seq, cudaStream_t s);
size_t msm_small_scratch_bytes(uint32_t rows, uint32_t n, uint32_t windows);
// first element of up to four device vectors -> host mailbox (mapped memory), then *host_flag = seq
cudaError_t k_mail(const fl_t* p0, const fl_t* p1, const fl_t* p2, const fl_t* p3, int n, fl_t* host, uint32_t* host_flag,
                   uint32_t seq, cudaStream_t s);
// first `count` entries of ntab <= 4 tables -> host mailbox (table k at host + k*count), then *host_flag = seq
cudaError_t k_mail_tables(const fl_t* p0, const fl_t* p1, const fl_t* p2, const fl_t* p3, int ntab, uint32_t count, fl_t* host,
                          uint32_t* host_flag, uint32_t seq, cudaStream_t s);

// ---- F_l vector kernels -------------------------------------------------------------------
// canonical -> Montgomery, flag[0] |= 1 if any input >= l            (lib.rs:65-89)
cudaError_t k_to_mont(const fl_t* in, fl_t* out, size_t n, uint32_t* flag, cudaStream_t s, fl_t* out2 = nullptr);  // out2: optional second copy
cudaError_t k_from_mont(const fl_t* in, fl_t* out, size_t n, cudaStream_t s);
// out[i] = a[i] op b[i]  (op 0 add, 1 sub, 2 mul) -- parity-test hook for the limb arithmetic
cudaError_t k_fl_binop(int op, const fl_t* a, const fl_t* b, fl_t* out, size_t n, cudaStream_t s);
// F_p (2^255-19) parity-test hook: out[i] = canon(a[i] op b[i]), op 0 add, 1 sub, 2 mul, 3 sqr(a), 4 invert(a)
struct fp_t;
cudaError_t k_fp_binop(int op, const fp_t* a, const fp_t* b, fp_t* out, size_t n, cudaStream_t s);
// flag[0] |= 1 unless A[i]*B[i] == C[i] for all i                     (r1csinstance.rs:238-245)
cudaError_t k_check_sat(const fl_t* A, const fl_t* B, const fl_t* C, size_t n, uint32_t* flag, cudaStream_t s);
// eq(r) table of length 2^ell                                         (dense_mlpoly.rs:67-83)
// r: device array of ell Montgomery scalars; tmp: >= 2^ceil(ell/2)+2^floor(ell/2) scalars
cudaError_t k_eq_table(const fl_t* r, uint32_t ell, fl_t* out, fl_t* tmp, cudaStream_t s);
// the same table with every entry multiplied by `init` (Montgomery)
cudaError_t k_eq_table_scaled(const fl_t* r, uint32_t ell, fl_t init, fl_t* out, fl_t* tmp, cudaStream_t s);
// out[i] = in[i*step + first], i < n
cudaError_t k_gather_stride(const fl_t* in, size_t n, size_t step, size_t first, fl_t* out, cudaStream_t s);

// Segments (rows of a CSR, columns of a CSC) longer than kLongSeg entries are not walked by one thread: they are cut
// into chunks of at most kSegChunk entries, a warp sums each chunk (sp_long_kernel), segments of more than
// kSumInline chunks are then reduced by one warp each (sp_big_kernel), and the thread that owns the segment in the
// main kernel picks the sum up.  CirC output has such segments: the constant-1 column holds a quarter of all
// entries, decomposition rows have 32-64 terms.  F_l addition is exact, so the split cannot change a bit.
static constexpr uint32_t kLongSeg = 16, kSegChunk = 512, kSumInline = 8;
struct LongSegs {
  const uint32_t* seg = nullptr;       // [nlong] ids of the long segments, ascending
  const uint32_t* first = nullptr;     // [nlong + 1] first chunk of each
  const uint32_t* chunk_lo = nullptr;  // [nchunks] entry range of a chunk
  const uint32_t* chunk_hi = nullptr;
  const uint32_t* big = nullptr;       // [nbig] indices into seg[] of the segments with more than kSumInline chunks
  fl_t* partial = nullptr;             // [nchunks] scratch: chunk sums
  fl_t* reduced = nullptr;             // [nlong] scratch: sums of the big segments
  uint32_t nlong = 0, nchunks = 0, nbig = 0;
};
struct Csr {  // three matrices share the layout
  const uint32_t* ptr[3];
  const uint32_t* idx[3];
  const fl_t* val[3];
  LongSegs lg[3];
};
// SNARK::encode: the dense representation of the three matrices (sparse_mlpoly.rs:371-422) as canonical integers:
// comb_ops (16 N scalars) and comb_mem (2 cells).  op_csr / op_csc: position -> op index (stable CSR / CSC).
cudaError_t k_snark_dense(const Csr& csr, const uint32_t* const op_csr[3], uint32_t nrows, const Csr& csc,
                          const uint32_t* const op_csc[3], uint32_t ncols, const uint32_t nnz[3], uint32_t N, uint32_t cells,
                          fl_t* comb_ops, fl_t* comb_mem, cudaStream_t s);
// Derefs.comb (8 N canonical scalars) from the address blocks of comb_ops and the two eq tables (Montgomery)
cudaError_t k_snark_deref(const fl_t* comb_ops, const fl_t* mem_rx, const fl_t* mem_ry, uint32_t N, fl_t* out, cudaStream_t s);
// Hash layer of one side of PolyEvalNetwork (sparse_mlpoly.rs:531-606) into layer 0 of eight product-tree buffers:
// trees[0] = init, [1..3] = read, [4..6] = write, [7] = audit.  mem Montgomery; addr / ts / audit_ts canonical integers.
cudaError_t k_snark_hash_layer(const fl_t* mem, const fl_t* addr, const fl_t* ts, const fl_t* audit_ts, uint32_t N, uint32_t cells,
                               fl_t r_hash, fl_t r_hash_sqr, fl_t r_ms, fl_t* const trees[8], cudaStream_t s);
// ProductCircuit::new + evaluate (product_tree.rs:16-64) of up to 16 circuits: bufs[i] holds 2 lens[i] scalars, layer k
// (lens[i] >> k entries, Montgomery) at offset 2 len - (2 len >> k); products[i] canonical.  lens are powers of two >= 2.
cudaError_t k_prod_trees(fl_t* const bufs[], const uint32_t lens[], int ntrees, fl_t* products, cudaStream_t s);
// y[m][row] = sum val * z[col]                                        (sparse_mlpoly.rs:452-466)
cudaError_t k_spmv3(const Csr& csr, uint32_t num_rows, const fl_t* z, fl_t* yA, fl_t* yB, fl_t* yC, cudaStream_t s);
// out[col] = rA*sum_A e[row]*val + rB*... + rC*...                    (sparse_mlpoly.rs:468-483, r1csproof.rs:268-279)
cudaError_t k_spmvT3_combine(const Csr& csc, uint32_t num_cols, const fl_t* e, fl_t rA, fl_t rB, fl_t rC, fl_t* out,
                             cudaStream_t s);

// Sum-check workspace for the cross-block reduction
struct ScWork {
  fl_t* partials;      // [max_blocks][3]
  unsigned int* counter;
  fl_t* result;        // device [3]
  uint32_t max_blocks;
  // optional mailbox in mapped pinned host memory: the last block also stores the three sums there,
  // fences (system scope) and then writes `seq` to host_flag, so the host can spin on the flag
  // instead of paying a copy + stream synchronise per Fiat-Shamir round
  fl_t* host_result = nullptr;
  uint32_t* host_flag = nullptr;
  uint32_t seq = 0;
};
// One fused round of the cubic ZK sum-check (sumcheck.rs:618-674):
//  if do_bind: tables (length n_old) are bound at r in place -> length n_old/2   (dense_mlpoly.rs:214-221)
//  then e0,e2,e3 of the (new) tables are reduced into w.result.
cudaError_t k_sc_cubic(fl_t* T, fl_t* A, fl_t* B, fl_t* C, size_t n_old, bool do_bind, fl_t r, const ScWork& w,
                       cudaStream_t s);
// Same for the quadratic sum-check (sumcheck.rs:453-484): result = e0, e2, (0)
// phase one with eq(tau) factored out (kernels.cu sc_cubic_eq_kernel): pyramids of suffix eq tables, one factored round
bool sc_round_is_long(size_t pairs, const ScWork& w);
cudaError_t k_eq_pyramid(const fl_t* tau, uint32_t m, uint32_t lo_bits, fl_t init, fl_t* lo, fl_t* hi, cudaStream_t s);
cudaError_t k_sc_cubic_eq(const fl_t* lo, const fl_t* hi, uint32_t lo_bits, fl_t* A, fl_t* B, fl_t* C, size_t n_old, bool do_bind,
                          fl_t r, const ScWork& w, cudaStream_t s);
// phase two on a z whose upper half is zero past its k0 leading entries (kernels.cu sc_quad_half0_kernel)
cudaError_t k_sc_quad_half0(const fl_t* Z, const fl_t* Q, size_t h, uint32_t k0, const ScWork& w, cudaStream_t s);
cudaError_t k_sc_half_fix(fl_t* Z, fl_t* Q, size_t h, uint32_t k0, fl_t t, cudaStream_t s);
cudaError_t k_sc_quad(fl_t* Z, fl_t* Q, size_t n_old, bool do_bind, fl_t r, const ScWork& w, cudaStream_t s);

// LZ[i] = sum_j Lv[j] * Z[j*R + i]                                    (dense_mlpoly.rs:205-212)
cudaError_t k_bound_L(const fl_t* Z, const fl_t* Lv, uint32_t Lsize, uint32_t Rsize, fl_t* out, fl_t* tmp,
                      cudaStream_t s);
// the same for a polynomial stored as canonical integers (comb_ops, comb_mem, Derefs.comb): Montgomery Lv times canonical Z
// sums to the canonical LZ (out_canon: the MSM's input form); out_mont receives its Montgomery form
cudaError_t k_bound_L_canon(const fl_t* Z_canon, const fl_t* Lv, uint32_t Lsize, uint32_t Rsize, fl_t* out_canon, fl_t* out_mont,
                            fl_t* tmp, cudaStream_t s);
size_t bound_L_tmp_scalars(uint32_t Lsize, uint32_t Rsize);
// out[0] = <a, b>   (n <= 2^16, one block)                            (nizk/mod.rs:305-308)
cudaError_t k_dot_small(const fl_t* a, const fl_t* b, uint32_t n, fl_t* out, cudaStream_t s);

// w.result[m] = <e, y_m> for the three vectors (verifier's multi_evaluate)   (sparse_mlpoly.rs:412-450)
cudaError_t k_dot3(const fl_t* e, const fl_t* yA, const fl_t* yB, const fl_t* yC, size_t n, const ScWork& w, cudaStream_t s);

// Bullet reduction scalar side (nizk/bullet.rs:71-118), see kernels.cu
// (a_in, b_in) are read only; with do_fold the folded vectors go to (a_out, b_out) -- a different pair of buffers
// pos_h / pos_q: where a row keeps the scalars of H (blind) and Q (inner product) -- the positions of those two bases in
// the window table; 0 = the table of a NIZKGens ([R] = H, [R+1] = Q)
cudaError_t k_bullet_step(const fl_t* a_in, const fl_t* b_in, fl_t* a_out, fl_t* b_out, const fl_t* coeff_in, fl_t* coeff_out,
                          uint32_t n_cur, uint32_t Rsize, bool do_fold, fl_t u, fl_t u_inv, fl_t blind_L, fl_t blind_R, fl_t* sL,
                          fl_t* sR, cudaStream_t s, uint32_t pos_h = 0, uint32_t pos_q = 0);
// s[g] = from_mont(d * coeff[g]) for g < R                            (nizk/mod.rs:483-490)
cudaError_t k_scale_from_mont(const fl_t* coeff, fl_t d, uint32_t n, fl_t* out, cudaStream_t s);

// ---- SNARK::prove: the batched sum-checks of ProductCircuitEvalProofBatched (product_tree.rs:257-382) -----------------
// Instance k of a layer: tables A_k, B_k (left / right half of a product-circuit layer, or left / right of a dot-product
// circuit) and C_k = the instance's own weight table, or NULL for the eq table all product circuits share (Cin; with a
// bind its bound form is written to Cout, a different buffer, by the units of instance `cwriter`).
static constexpr int kPcMax = 18;
struct PcBatch {
  fl_t* A[kPcMax];
  fl_t* B[kPcMax];
  fl_t* C[kPcMax];
  fl_t coeff[kPcMax];  // rand_coeffs_next_layer, Montgomery
  const fl_t* Cin;
  fl_t* Cout;
  uint32_t ninst, cwriter;
};
// One round of SumcheckInstanceProof::prove_cubic_batched (sumcheck.rs:252-422) on tables of length n_old: if do_bind
// every table is first bound at r (-> n_old / 2); then the round's combined evaluations at 0, 2, 3 go to w.result and
// the mailbox.  Lengths are powers of two, length after the bind >= 2.
// Doorbell: a round kernel can be launched AHEAD of its challenge (while the host still works on the previous round) and
// then waits, on the device, until the host has stored r into *r and `want` into *flag (both in mapped pinned host
// memory).  That takes the launch latency out of the round trip: ~9.5 -> ~4.5 us of fixed cost per Fiat-Shamir round
// (tools/microbench/roundtrip_bench.cu).  flag == NULL: r comes as a launch argument as before.  A kernel whose bell is
// never rung gives up after ~2 s without touching the tables or the mailbox.
struct Bell {
  const uint32_t* flag = nullptr;  // mapped host memory: the doorbell
  const fl_t* r = nullptr;         // mapped host memory: the challenge
  uint32_t* relay = nullptr;       // device memory, 16 words: only block 0 polls the host (a grid of blocks polling over
                                   // PCIe starves the very writes it waits for); it passes flag and r on through here
  uint32_t want = 0;
};
cudaError_t k_pc_round(const PcBatch& b, size_t n_old, bool do_bind, fl_t r, const ScWork& w, cudaStream_t s,
                       const Bell* bell = nullptr);
// host_out[3k + t] = final value of table t (A, B, own C / zero) of instance k, after a last bind at r if do_bind
cudaError_t k_pc_final(const PcBatch& b, bool do_bind, fl_t r, fl_t* host_out, uint32_t* host_flag, uint32_t seq, cudaStream_t s);
// out[j] = sum_i X_j[i] * Y_j[i] (* W_j[i] if W_j != NULL), j < nvec; tmp: multi_dot_tmp_scalars(nvec).
// final_to_mont: Y canonical integers -> the sum is canonical, converted to Montgomery at the end.
static constexpr int kDotMax = 24;
struct DotBatch {
  const fl_t* X[kDotMax];
  const fl_t* Y[kDotMax];
  const fl_t* W[kDotMax];
  uint32_t nvec;
};
size_t multi_dot_tmp_scalars(uint32_t nvec);
cudaError_t k_multi_dot(const DotBatch& b, size_t n, bool final_to_mont, fl_t* tmp, fl_t* out, cudaStream_t s);

// ---- Instance::new on the device (sparse_build.cu) ------------------------------------------------------------------
// canonical -> Montgomery; *first_bad = index of the first scalar >= l, 0xFFFFFFFF if none
cudaError_t k_to_mont_first_bad(const fl_t* in, fl_t* out, uint32_t n, unsigned int* first_bad, cudaStream_t s);
// Stable ordering of one matrix by `keys` (row for the CSR, column for the CSC; 0xFFFFFFFF = entry of another rank; `kept`
// real keys, all below 2^key_bits): ptr[nseg + 1], and per position the other coordinate, the Montgomery value and (op !=
// NULL) the index of the entry in the caller's list.  tmp: sparse_build_tmp_bytes(nnz).
size_t sparse_build_tmp_bytes(uint32_t nnz);
cudaError_t k_sparse_build(const uint32_t* keys, const uint32_t* other, const fl_t* vals_mont, uint32_t nnz, uint32_t kept, uint32_t nseg,
                           uint32_t key_bits, uint32_t* ptr, uint32_t* idx, fl_t* val, uint32_t* op, void* tmp, size_t tmp_bytes,
                           cudaStream_t s);

// per-kernel device time of every launch while enabled (CUDA events around each launch); report = CSV lines
// "kernel,launches,total_ms" sorted by time, and clears the records
void kprof_enable(bool on);
std::string kprof_report();
// counter of kernel launches issued through this file (bench.py gpu_launches)
uint64_t launch_count();
void launch_count_reset();

}  // namespace sb
