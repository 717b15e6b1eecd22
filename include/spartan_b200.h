/*
 * spartan_b200.h -- C ABI of libspartan_b200.so, the B200-native backend for the
 * Spartan NIZK prove step that Otti's run.py drives
 * (run.py:58,100 -> spartan-zkinterface/src/main.rs:57-100 -> libspartan).
 *
 * Each entry point replaces one item of libspartan's public Rust API
 * (/root/reference/Spartan/src/lib.rs); the citation after each declaration is
 * the reference interface it stands in for.  INTEGRATION.md shows the Rust shim
 * that binds these symbols and keeps the Instance / Assignment / NIZKGens / NIZK
 * types unchanged for callers.
 *
 * Conventions
 *  - plain pointers and sizes only; no CUDA or torch types cross the boundary;
 *  - scalars cross as 32-byte canonical little-endian integers < l
 *    (Scalar::from_bytes, scalar/ristretto255.rs:398-422); a value >= l yields
 *    SB200_ERR_INVALID_SCALAR exactly where the reference returns
 *    R1CSError::InvalidScalar;
 *  - proofs cross as the bincode 1.3 serialisation of the reference's `NIZK`
 *    struct (lib.rs:481-485), byte-identical to the reference for the same
 *    instance, assignment, generators, transcript label and tape seed;
 *  - return value 0 = success, negative = error (codes below); no panic, abort
 *    or C++ exception crosses the boundary.  sb200_last_error() gives the text
 *    for the calling thread;
 *  - there is NO CPU fallback: every compute entry point fails with
 *    SB200_ERR_NO_DEVICE when no sm_100 GPU is usable;
 *  - handles may be shared between threads, as the reference's `&Instance` / `&NIZKGens` are: every call that uses an
 *    instance's device workspace (prove, verify, is_sat, witness upload, encode) holds that instance's mutex, and the two
 *    provers also serialise on one process-wide mutex (one process drives one GPU, and its host worker pool has a single
 *    posting thread).  Concurrent calls are therefore safe and run one after the other; freeing a handle while a call
 *    on it is in flight is not.
 */
#ifndef SPARTAN_B200_H
#define SPARTAN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sb200_instance sb200_instance; /* libspartan::Instance      (lib.rs:115-118) */
typedef struct sb200_gens sb200_gens;         /* libspartan::NIZKGens      (lib.rs:459-462) */

enum {
  SB200_OK = 0,
  /* R1CSError, errors.rs:19-32 */
  SB200_ERR_NON_POW2_CONS = -1,
  SB200_ERR_NON_POW2_VARS = -2,
  SB200_ERR_INVALID_NUMBER_OF_INPUTS = -3,
  SB200_ERR_INVALID_NUMBER_OF_VARS = -4,
  SB200_ERR_INVALID_SCALAR = -5,
  SB200_ERR_INVALID_INDEX = -6,
  /* ProofVerifyError, errors.rs:4-10 (and the reference's verify-side panics) */
  SB200_ERR_PROOF_VERIFY = -20,
  SB200_ERR_PROOF_MALFORMED = -21,
  /* conditions on which the reference panics (assert!) */
  SB200_ERR_ASSERT = -30,
  /* backend */
  SB200_ERR_NO_DEVICE = -100,
  SB200_ERR_CUDA = -101,
  SB200_ERR_ARG = -102,
  SB200_ERR_COMM = -103,
  SB200_ERR_ZKIF = -104 /* the reference unwrap()s / panics on such input */
};

/*
 * Caller-owned Fiat-Shamir transcript (merlin::Transcript keeps its state
 * private, so the Rust side passes two thunks over its own object; hashing
 * stays on the host as north_star requires).  NULL selects the library's
 * built-in Merlin, seeded with Transcript::new(label).
 *   append_message  == merlin::Transcript::append_message(label, msg)
 *   challenge_bytes == merlin::Transcript::challenge_bytes(label, out)
 */
typedef struct sb200_transcript_vt {
  void* ctx;
  void (*append_message)(void* ctx, const uint8_t* label, size_t label_len, const uint8_t* msg, size_t msg_len);
  void (*challenge_bytes)(void* ctx, const uint8_t* label, size_t label_len, uint8_t* out, size_t out_len);
} sb200_transcript_vt;

/* merlin::Transcript for callers without their own (C, C++, Python hosts).
 *   sb200_transcript_new             == Transcript::new(label)                  (spzk main.rs:78)
 *   sb200_transcript_append_message  == Transcript::append_message
 *   sb200_transcript_challenge_bytes == Transcript::challenge_bytes
 *   sb200_transcript_vtable fills a vtable bound to the object, to pass to prove/verify:
 *   the object's state advances exactly as the reference mutates `&mut Transcript`. */
typedef struct sb200_transcript sb200_transcript;
int sb200_transcript_new(const uint8_t* label, size_t label_len, sb200_transcript** out);
void sb200_transcript_append_message(sb200_transcript* t, const uint8_t* label, size_t label_len, const uint8_t* msg, size_t msg_len);
void sb200_transcript_challenge_bytes(sb200_transcript* t, const uint8_t* label, size_t label_len, uint8_t* out, size_t out_len);
void sb200_transcript_vtable(sb200_transcript* t, sb200_transcript_vt* out);
void sb200_transcript_free(sb200_transcript* t);

/*
 * Multi-GPU exchange hook (one process per GPU).  The prover shards the
 * hypercube / commitment rows across `world` ranks and needs only an
 * all-gather of a few dozen bytes per Fiat-Shamir round; the host side
 * supplies it (torch.distributed NCCL/gloo from Python, NCCL from Rust).
 * NULL = single GPU.
 *   allgather: every rank contributes `bytes` bytes from HOST memory `send`;
 *   `recv` (HOST, world*bytes) receives the contributions in rank order.
 */
typedef struct sb200_comm_vt {
  void* ctx;
  int rank;
  int world;
  int (*allgather)(void* ctx, const void* send, size_t bytes, void* recv);
} sb200_comm_vt;

/* Exchange over a POSIX shared-memory segment for ranks on ONE box (the 8 x B200 node of north_star): fills
 * *out with a hook whose allgather costs ~1 us.  Collective: every rank calls it with the same name, which should
 * be unique to the job.  Rank 0 removes whatever a crashed job left under that name and creates the segment anew; the
 * other ranks only accept a segment in which rank 0 of THIS call answers them.  Hosts spanning boxes supply their own
 * hook (NCCL / torch.distributed: otti_b200/comm.py). */
int sb200_comm_shm_open(const char* name, int rank, int world, sb200_comm_vt* out);
void sb200_comm_shm_close(sb200_comm_vt* comm);
/* The same hook over NCCL, called natively: the bytes are staged through device memory and gathered with ncclAllGather
 * over NVLink.  libnccl.so.2 is dlopen'ed (no link-time dependency).  Rank 0 makes the id, the host program hands it to
 * the other ranks (128 bytes), every rank opens.  Provided so that the shared-memory hook can be measured against it. */
int sb200_comm_nccl_unique_id(uint8_t id128[128]);
int sb200_comm_nccl_open(const uint8_t id128[128], int rank, int world, sb200_comm_vt* out);
void sb200_comm_nccl_close(sb200_comm_vt* comm);
/* Host-only self-test of a hook (no GPU): `rounds` all-gathers of 3 scalars per rank, summed mod l as a sharded
 * sum-check round does; the digest must equal sb200_comm_selftest_expected(world, rounds) on every rank. */
int sb200_comm_selftest(const sb200_comm_vt* comm, int rounds, uint8_t digest32_out[32]);
int sb200_comm_selftest_expected(int world, int rounds, uint8_t digest32_out[32]);

/* Selects the CUDA device for the calling process (one process per GPU).
 * Fails with SB200_ERR_NO_DEVICE if it is not compute capability 10.x.
 * With LOCAL_WORLD_SIZE > 1 (one rank per GPU under torchrun) the calling thread -- and the host worker threads the
 * library creates from it later -- are bound to a slice of the cores next to the GPU (its PCI local_cpulist, split
 * among the ranks that share it); SB200_BIND=0 turns that off, SB200_BIND=1 forces it for a single process.
 * sb200_bound_cpus() = how many CPUs the last sb200_init bound the caller to (0 = affinity left alone). */
int sb200_init(int device);
int sb200_bound_cpus(void);

/* Assignment::new(&[[u8;32]]) -> Result<Assignment, R1CSError>            lib.rs:65-89
 * SB200_ERR_INVALID_SCALAR if any of the n values is not a canonical scalar. */
int sb200_assignment_check(const uint8_t* vals32, uint64_t n);

/* Instance::new(num_cons, num_vars, num_inputs, &A, &B, &C)              lib.rs:121-225
 * rows/cols/vals32: three COO matrices; vals32[m] holds nnz[m] x 32 bytes. */
int sb200_instance_new(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, const uint64_t* const rows[3],
                       const uint64_t* const cols[3], const uint8_t* const vals32[3], const uint64_t nnz[3],
                       sb200_instance** out);

/* The same constructor for ONE PROOF SHARDED over world = 2, 4 or 8 GPUs (one process per GPU; every rank passes
 * the same matrices).  This rank keeps the constraint rows i = rank (mod world) as CSR and the z columns
 * j = rank (mod world) as CSC.  sb200_nizk_prove on such an instance needs a sb200_comm_vt
 * with the same rank/world; all ranks call it with identical arguments and obtain identical proofs.
 * SB200_ERR_ARG if the instance is too small to give every rank two rows, two columns and one commitment row. */
int sb200_instance_new_sharded(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, const uint64_t* const rows[3],
                               const uint64_t* const cols[3], const uint8_t* const vals32[3], const uint64_t nnz[3],
                               int rank, int world, sb200_instance** out);
int sb200_instance_synthetic_sharded(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t seed, int rank,
                                     int world, sb200_instance** out, uint8_t* vars32_out, uint8_t* inputs32_out);

/* Instance::produce_synthetic_r1cs(num_cons, num_vars, num_inputs)       lib.rs:260-271
 * OsRng is replaced by SplitMix64(seed) so runs are reproducible.  vars32_out
 * (num_vars x 32 B) and inputs32_out (num_inputs x 32 B) receive the
 * satisfying assignment. */
int sb200_instance_synthetic(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t seed,
                             sb200_instance** out, uint8_t* vars32_out, uint8_t* inputs32_out);

/* Instance::is_sat(&vars, &inputs) -> Result<bool, R1CSError>            lib.rs:228-257 */
int sb200_instance_is_sat(const sb200_instance* inst, const uint8_t* vars32, uint64_t n_vars, const uint8_t* inputs32,
                          uint64_t n_inputs, int* sat_out);

/* padded sizes: out[0]=num_cons, out[1]=num_vars, out[2]=num_inputs (r1csinstance.rs:124-134) */
int sb200_instance_dims(const sb200_instance* inst, uint64_t out[3]);
void sb200_instance_free(sb200_instance* inst);

/*
 * zkinterface ingestion: R1csReader::new + R1cs::from + R1cs::instance / vars_assignment / inputs_assignment
 * (spartan-zkinterface/src/lib.rs:75-260, main.rs:36-60).  Host only; parses the three flatbuffers spzk is given
 * (<circuit.zkif> <inputs.zkif> <witness.zkif>, main.rs:18) into COO triples and canonical assignment bytes.
 * Variable ids map as in R1cs::translate (lib.rs:96-107): witness position, else input position + num_vars + 1,
 * else the constant column num_vars; the first occurrence of a duplicated id wins.
 * SB200_ERR_ZKIF where the reference unwrap()s or panics (wrong message type, missing field, no constraints,
 * value wider than 32 bytes).
 *   sb200_r1cs_dims    out = num_cons, num_vars, num_inputs, nnz(A), nnz(B), nnz(C), non_zero_entries (lib.rs:250)
 *   sb200_r1cs_matrix  m = 0,1,2 -> A,B,C as (rows, cols, 32-byte LE values), owned by the object
 *   sb200_r1cs_instance == R1cs::instance (Instance::new on the parsed matrices; rank/world as in *_new_sharded) */
typedef struct sb200_r1cs sb200_r1cs;
int sb200_zkif_parse(const uint8_t* circuit, size_t circuit_len, const uint8_t* header, size_t header_len, const uint8_t* witness,
                     size_t witness_len, sb200_r1cs** out);
int sb200_r1cs_dims(const sb200_r1cs* r, uint64_t out[7]);
int sb200_r1cs_matrix(const sb200_r1cs* r, int m, const uint64_t** rows, const uint64_t** cols, const uint8_t** vals32);
const uint8_t* sb200_r1cs_vars(const sb200_r1cs* r);
const uint8_t* sb200_r1cs_inputs(const sb200_r1cs* r);
int sb200_r1cs_instance(const sb200_r1cs* r, int rank, int world, sb200_instance** out);
void sb200_r1cs_free(sb200_r1cs* r);

/*
 * rust-circ adapter: r1cs_to_spartan + lc_to_v + int_to_scalar        rust-circ/src/target/r1cs/spartan.rs:31-186
 * A CirC `R1cs` (signals by name, linear combinations over arbitrary-precision integers) -> the same sb200_r1cs
 * object the zkinterface reader returns, so the SGD pipeline (rust-circ/examples/circ.rs:246-273) can hand its circuit
 * over in memory.  Host only.
 *   signals      r1cs.values in the caller's iteration order (the reference iterates a HashMap, i.e. in no defined
 *                order): CirC id, name (names[name_off[s] .. name_off[s+1])) and value.  A name containing
 *                "main_f0_lex0_w" is a witness, else one containing "main_f0_lex0_" an input, else a witness
 *                (spartan.rs:52-75); witnesses / inputs are numbered in the order they are met.
 *   integers     little-endian MAGNITUDE bytes of any length: integer k of a list is bytes[off[k] .. off[k+1]).
 *                int_to_scalar reads rug's limbs, which carry no sign (spartan.rs:141-160): value = |i| mod l.
 *   n_public     r1cs.public_idxs.len(), which the reference asserts to be 0 (spartan.rs:42)
 *   next_idx     r1cs.next_idx; the reference asserts wit.len() + inp.len() == next_idx (spartan.rs:80)
 *   constraints  for m = 0,1,2 (A,B,C): monomials of constraint i are [lc_off[m][i], lc_off[m][i+1]) of
 *                mono_ids[m] / coeff_*[m] (caller's order); const_*[m] holds lc.constant of every constraint,
 *                added on the constant column num_vars when non-zero (spartan.rs:177-184).
 * Column mapping (spartan.rs:96-99,168-175): witness position, input position + num_vars + 1.
 * SB200_ERR_ASSERT where the reference asserts, unwrap()s or panics (public inputs present, next_idx mismatch,
 * monomial id >= num_vars "variable number off", unknown id).
 * Continue with sb200_r1cs_instance / sb200_r1cs_vars / sb200_r1cs_inputs; the reference then asserts
 * Instance::is_sat (spartan.rs:135-136), which is sb200_instance_is_sat.
 */
int sb200_circ_r1cs(uint64_t n_signals, const uint64_t* signal_ids, const uint8_t* names, const uint64_t* name_off,
                    const uint8_t* value_bytes, const uint64_t* value_off, uint64_t n_public, uint64_t next_idx, uint64_t n_cons,
                    const uint64_t* const lc_off[3], const uint64_t* const mono_ids[3], const uint8_t* const coeff_bytes[3],
                    const uint64_t* const coeff_off[3], const uint8_t* const const_bytes[3], const uint64_t* const const_off[3],
                    sb200_r1cs** out);
/* int_to_scalar (spartan.rs:141-160) on one integer: |i| mod l as 32 canonical little-endian bytes */
int sb200_circ_int_to_scalar(const uint8_t* magnitude_le, size_t len, uint8_t out32[32]);

/* NIZKGens::new(num_cons, num_vars, num_inputs)                          lib.rs:466-477
 * label is b"gens_r1cs_sat" in the reference (lib.rs:475); NULL selects it.
 * Besides deriving the generators this builds the fixed-base window tables in
 * HBM that the commitment kernel reads (one-off, outside the timed prove, as
 * the reference builds its gens outside the bench loop, benches/nizk.rs:27-39). */
int sb200_gens_new(const uint8_t* label, size_t label_len, uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs,
                   sb200_gens** out);
/* The same with a hint: how many proofs the caller expects to make with these generators.  The window width of the
 * tables is then chosen to minimise table build + proofs_expected x commitment instead of filling the HBM: a one-shot
 * caller like spzk, which times NIZKGens::new together with the prove (spartan-zkinterface/src/main.rs:71-85), passes 1
 * and gets generators in milliseconds.  0 = many (what sb200_gens_new does, unless SB200_PROOFS_EXPECTED is set).
 * The generators -- and so every proof byte -- do not depend on the hint. */
int sb200_gens_new_for(const uint8_t* label, size_t label_len, uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs,
                       uint64_t proofs_expected, sb200_gens** out);
void sb200_gens_free(sb200_gens* gens);

/* NIZK::prove(&inst, vars, &inputs, &gens, &mut transcript) -> NIZK      lib.rs:493-535
 *  transcript      caller transcript, or NULL to use Transcript::new(label)
 *  label           transcript label when transcript == NULL (b"nizk_example" in spzk, main.rs:78)
 *  tape_seed64     64 uniform bytes -> Scalar::from_bytes_wide, standing in for the
 *                  Scalar::random(OsRng) that seeds RandomTape (random.rs:11-19);
 *                  NULL draws them from the OS like the reference.
 *  comm            NULL, or the exchange hook when the proof is sharded over several GPUs
 *  vars32          host pointer, or NULL to use the witness placed by sb200_witness_upload
 *  proof_out/len   bincode(NIZK); release with sb200_free
 * SB200_ERR_ASSERT where the reference asserts (inputs.len() >= vars.len(), r1csproof.rs:153). */
int sb200_nizk_prove(const sb200_instance* inst, const uint8_t* vars32, uint64_t n_vars, const uint8_t* inputs32,
                     uint64_t n_inputs, const sb200_gens* gens, const sb200_transcript_vt* transcript,
                     const uint8_t* label, size_t label_len, const uint8_t* tape_seed64, const sb200_comm_vt* comm,
                     uint8_t** proof_out, size_t* proof_len);

/* NIZK::verify(&self, &inst, &inputs, &mut transcript, &gens)            lib.rs:538-574
 * 0 = accepted; SB200_ERR_PROOF_VERIFY where the reference returns Err or panics;
 * SB200_ERR_PROOF_MALFORMED if the bytes do not parse as bincode(NIZK). */
int sb200_nizk_verify(const sb200_instance* inst, const uint8_t* inputs32, uint64_t n_inputs, const sb200_gens* gens,
                      const sb200_transcript_vt* transcript, const uint8_t* label, size_t label_len,
                      const uint8_t* proof, size_t proof_len);

/* Timer scopes of the last prove on this thread, seconds               timer.rs, r1csproof.rs:155-307
 * out[0..9] = polycommit, prove_sc_phase_one, prove_sc_phase_two, polyeval, total,
 *             witness h2d (inside total), critical-path host Fiat-Shamir time (inside total),
 *             kernel launches, host->device bytes, device->host bytes,
 *             device time of the commitment MSM kernel, device time of the first cubic sum-check kernel
 *             (both from CUDA events on the instance's stream),
 *             [12] seconds inside the exchange hook, [13] exchange calls, [14] world, [15] reserved */
int sb200_last_prove_profile(double out[16]);

/* ---- SNARK mode: SNARKGens, SNARK::encode, SNARK::prove, SNARK::verify.
 * SNARKGens::new(num_cons, num_vars, num_inputs, num_nz_entries)          lib.rs:279-303
 *   only the part SNARK::encode reads: gens_r1cs_eval = R1CSCommitmentGens::new(b"gens_r1cs_eval", ...)
 *   (r1csinstance.rs:29-43, sparse_mlpoly.rs:301-327) and its window tables in HBM.  Arguments as in the reference
 *   (unpadded; num_nz_entries = the maximum number of non-zero entries of A, B, C).
 *   SB200_ERR_ASSERT where the reference asserts (num_inputs < num_vars, r1csinstance.rs:36). */
typedef struct sb200_snark_gens sb200_snark_gens; /* libspartan::SNARKGens (lib.rs:273-277) */
int sb200_snark_gens_new(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t num_nz_entries,
                         sb200_snark_gens** out);
/* The same with the hint of sb200_gens_new_for (how many proofs the caller expects to make: window widths that
 * minimise table build + proofs x commitments; 0 = many).  The generators do not depend on the hint. */
int sb200_snark_gens_new_for(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t num_nz_entries,
                             uint64_t proofs_expected, sb200_snark_gens** out);
void sb200_snark_gens_free(sb200_snark_gens* gens);
/* SNARK::encode(&inst, &gens) -> (ComputationCommitment, ComputationDecommitment)    lib.rs:321-333
 *   -> R1CSInstance::commit r1csinstance.rs:285-297 -> SparseMatPolynomial::multi_commit sparse_mlpoly.rs:485-506.
 * Public and deterministic (no transcript, no randomness).  comm_out (release with sb200_free): the fields of
 * R1CSCommitment (r1csinstance.rs:46-51; not Serialize in the reference) as three little-endian u64 -- num_cons,
 * num_vars, num_inputs, padded -- followed by the bincode 1.3 serialisation of SparseMatPolyCommitment
 * { batch_size, num_ops, num_mem_cells, comm_comb_ops, comm_comb_mem } (sparse_mlpoly.rs:329-337).
 * The decommitment (the dense representation) is rebuilt from the instance when needed and is not returned.
 * SB200_ERR_ASSERT where the reference asserts (generators made for other sizes, commitments.rs:81);
 * SB200_ERR_ARG for a sharded instance or one with fewer than two constraints. */
int sb200_snark_encode(const sb200_instance* inst, const sb200_snark_gens* gens, uint8_t** comm_out, size_t* comm_len);
/* First step of SNARK::prove's R1CSEvalProof (SparseMatPolyEvalProof::prove, sparse_mlpoly.rs:1485-1503, the
 * "commit_nondet_witness" timer): the eq tables of the equalized point (rx, ry) (:1453-1469), what every op of the
 * dense representation reads from them (Derefs, :47-62, 268-284, 508-514) and its unblinded commitment
 * (Derefs::commit, :64-67).  Deterministic given the point.  rx32 / ry32: canonical scalars.
 * points_out (release with sb200_free): the L compressed points of DerefsCommitment.comm_ops_val, in order;
 * comb_out: NULL, or a host buffer of next_power_of_two(6 N) x 32 bytes that receives Derefs.comb (canonical),
 * N = next_power_of_two(num_nz_entries).  sb200_last_encode_profile then reports out[0] = eq tables + dense
 * representation + gather, out[1] = the commitment.
 * SB200_ERR_ASSERT where the reference panics (generators made for other sizes; a point shorter than the memory). */
int sb200_snark_commit_derefs(const sb200_instance* inst, const sb200_snark_gens* gens, const uint8_t* rx32, uint64_t n_rx,
                              const uint8_t* ry32, uint64_t n_ry, uint8_t** points_out, size_t* points_len, uint8_t* comb_out);
/* Second step of SparseMatPolyEvalProof::prove (sparse_mlpoly.rs:1505-1519, the "build_layered_network" timer):
 * PolyEvalNetwork::new (:652-668) -- the hash layers (build_hash_layer, :531-606) and the product circuits
 * (product_tree.rs:16-64) of the row and the column side, built in HBM -- reduced to the sixteen values
 * ProductCircuit::evaluate returns (the claims ProductLayerProof::prove starts from): out16 = row side init, read[0..3),
 * write[0..3), audit, then the column side, canonical scalars.  r_hash32 / r_multiset32: the two "challenge_r_hash"
 * challenges (:1507).  sb200_last_encode_profile then reports out[0] = eq tables + dense representation, out[1] = hash
 * layers, out[2] = product circuits.  SB200_ERR_ASSERT where the reference panics (point / memory size mismatch). */
int sb200_snark_network_products(const sb200_instance* inst, const uint8_t* rx32, uint64_t n_rx, const uint8_t* ry32, uint64_t n_ry,
                                 const uint8_t r_hash32[32], const uint8_t r_multiset32[32], uint8_t out16[512]);
/* SNARK::prove(&inst, &comm, &decomm, vars, &inputs, &gens, &mut transcript) -> SNARK      lib.rs:335-403
 *   = R1CSProof::prove over gens_r1cs_sat (r1csproof.rs:141-345), inst.evaluate(rx, ry) (lib.rs:374-385), and
 *   R1CSEvalProof::prove (r1csinstance.rs:310-329 -> SparseMatPolyEvalProof::prove, sparse_mlpoly.rs:1471-1538:
 *   Derefs commitment, hash layers + product circuits, ProductLayerProof (the batched product-circuit sum-checks of
 *   product_tree.rs:257-382), HashLayerProof with its three PolyEvalProofs).
 * The commitment does not enter the prover's transcript and the decommitment (the dense representation of the matrices) is
 * rebuilt on the device from `inst`, so neither is an argument.  transcript / label / tape_seed64 / vars32 as for
 * sb200_nizk_prove (b"snark_example" is the label of the reference's example).  proof_out: bincode(SNARK), release with
 * sb200_free.  SB200_ERR_ASSERT where the reference asserts or panics (generators made for other sizes, an unsatisfied
 * memory check); SB200_ERR_ARG for a sharded instance or one with fewer than two constraints. */
int sb200_snark_prove(const sb200_instance* inst, const sb200_snark_gens* gens, const uint8_t* vars32, uint64_t n_vars,
                      const uint8_t* inputs32, uint64_t n_inputs, const sb200_transcript_vt* transcript, const uint8_t* label,
                      size_t label_len, const uint8_t* tape_seed64, uint8_t** proof_out, size_t* proof_len);
/* SNARK::verify(&self, &comm, &inputs, &mut transcript, &gens)                               lib.rs:405-457
 *   comm: the bytes sb200_snark_encode returned (R1CSCommitment: sizes + SparseMatPolyCommitment).  Needs no instance:
 *   R1CSProof::verify against the claimed evaluations, then R1CSEvalProof::verify (sparse_mlpoly.rs:1540-1582).  The
 *   size-R generator MSMs of the four bullet checks run on the device, everything else on the host.
 * 0 = accepted; SB200_ERR_PROOF_VERIFY where the reference returns Err or panics; SB200_ERR_PROOF_MALFORMED if the bytes
 * do not parse as bincode(SNARK) / as a commitment. */
int sb200_snark_verify(const uint8_t* comm, size_t comm_len, const sb200_snark_gens* gens, const uint8_t* inputs32, uint64_t n_inputs,
                       const sb200_transcript_vt* transcript, const uint8_t* label, size_t label_len, const uint8_t* proof,
                       size_t proof_len);
/* Timer scopes of the last sb200_snark_prove, seconds (the reference's profiler lines, README.md:305-345):
 * out[0] R1CSProof::prove, [1] eval_sparse_polys, [2] commit_nondet_witness, [3] build_layered_network,
 * [4] product layer proofs, [5] hash layer proof ([4] + [5] = evalproof_layered_network), [6] whole call, [7] kernel launches */
int sb200_last_snark_profile(double out[8]);
/* milliseconds of the last sb200_snark_encode: out[0] dense representation, out[1] commitment of comb_ops,
 * out[2] commitment of comb_mem (CUDA events), out[3] the whole call on the host clock */
int sb200_last_encode_profile(double out[4]);

/* Measurement helpers (no reference counterpart).
 * sb200_witness_upload places vars in HBM so that a following sb200_nizk_prove(vars32 = NULL, same
 * n_vars) starts with its input resident; sb200_instance_stream returns the cudaStream_t (as an
 * integer) the instance's kernels are launched on, for CUDA-event timing on that stream. */
int sb200_witness_upload(const sb200_instance* inst, const sb200_gens* gens, const uint8_t* vars32, uint64_t n_vars);
uint64_t sb200_instance_stream(const sb200_instance* inst);

void sb200_free(void* p);
const char* sb200_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* SPARTAN_B200_H */
