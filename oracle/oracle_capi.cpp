// ORACLE (test infrastructure only).  C entry points over the CPU restatement
// so that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs can drive it through ctypes.  Nothing under
// otti_b200/ links, includes or loads this.
#include <algorithm>
#include <chrono>
#include <cstdlib>

#include "snark.hpp"

using namespace orc;

static Fl ld(const uint8_t* p) { Fl x; memcpy(x.v, p, 32); return x; }
static void st(uint8_t* p, const Fl& x) { memcpy(p, x.v, 32); }
static FlVec ldv(const uint8_t* p, size_t n) { FlVec v(n); if (n) memcpy(v.data(), p, 32 * n); return v; }

struct OrcInst { Instance inst; };
struct OrcGens { NIZKGens g; OrcGens(size_t a, size_t b, size_t c, const std::string& l) : g(a, b, c, l) {} };

extern "C" {

// ---- F_l primitives on raw Montgomery limbs (scalar/ristretto255.rs)
int orc_fl_from_bytes(const uint8_t in[32], uint8_t out[32]) { Fl x; bool ok = fl_from_bytes(in, x); st(out, x); return ok; }
void orc_fl_to_bytes(const uint8_t in[32], uint8_t out[32]) { fl_to_bytes(ld(in), out); }
void orc_fl_from_bytes_wide(const uint8_t in[64], uint8_t out[32]) { st(out, fl_from_bytes_wide(in)); }
void orc_fl_binop(int op, const uint8_t* a, const uint8_t* b, uint8_t* out) {
  Fl x = ld(a), y = ld(b);
  st(out, op == 0 ? fl_add(x, y) : op == 1 ? fl_sub(x, y) : fl_mul(x, y));
}
void orc_fl_unop(int op, const uint8_t* a, uint8_t* out) {
  Fl x = ld(a);
  st(out, op == 0 ? fl_neg(x) : op == 1 ? fl_square(x) : fl_invert(x));
}
// batch canonical bytes -> Montgomery limbs; returns index of first invalid or -1
long orc_fl_batch_from_bytes(const uint8_t* in, size_t n, uint8_t* out) {
  long bad = -1;
  for (size_t i = 0; i < n; i++) { Fl x; if (!fl_from_bytes(in + 32 * i, x) && bad < 0) bad = (long)i; st(out + 32 * i, x); }
  return bad;
}
void orc_fl_batch_to_bytes(const uint8_t* in, size_t n, uint8_t* out) { for (size_t i = 0; i < n; i++) fl_to_bytes(ld(in + 32 * i), out + 32 * i); }

// ---- group
void orc_basepoint(uint8_t out[32]) { pt_compress(pt_basepoint(), out); }
int orc_pt_mul(const uint8_t pc[32], const uint8_t k_canon[32], uint8_t out[32]) {
  Pt p; if (!pt_decompress(pc, p)) return 0;
  pt_compress(pt_mul_bytes(k_canon, p), out); return 1;
}
int orc_pt_add(const uint8_t a[32], const uint8_t b[32], uint8_t out[32]) {
  Pt p, q; if (!pt_decompress(a, p) || !pt_decompress(b, q)) return 0;
  pt_compress(pt_add(p, q), out); return 1;
}
// affine (x, y) canonical LE of a compressed point (the ristretto representative decompress() yields)
int orc_pt_affine(const uint8_t a[32], uint8_t out[64]) {
  Pt p; if (!pt_decompress(a, p)) return 0;
  fp_to_bytes(p.X, out); fp_to_bytes(p.Y, out + 32); return 1;
}
int orc_pt_decompress_ok(const uint8_t a[32]) { Pt p; return pt_decompress(a, p); }
void orc_from_uniform_bytes(const uint8_t in[64], uint8_t out[32]) { pt_compress(pt_from_uniform_bytes(in), out); }
// sum_i k_i * P_i (k Montgomery limbs, P compressed)
int orc_pt_msm(const uint8_t* k, const uint8_t* pc, size_t n, uint8_t out[32]) {
  std::vector<Pt> P(n); FlVec s = ldv(k, n);
  for (size_t i = 0; i < n; i++) if (!pt_decompress(pc + 32 * i, P[i])) return 0;
  pt_compress(pt_msm(s.data(), P.data(), n), out); return 1;
}

// ---- transcript
void orc_merlin_test(const uint8_t* label, size_t ll, const uint8_t* mlabel, const uint8_t* msg, size_t ml,
                     const uint8_t* clabel, uint8_t* out, size_t n) {
  Transcript t(std::string((const char*)label, ll));
  t.append_message((const char*)mlabel, msg, ml);
  t.challenge_bytes((const char*)clabel, out, n);
}
void orc_shake256(const uint8_t* in, size_t n, uint8_t* out, size_t m) { Shake256 s; s.absorb(in, n); s.squeeze(out, m); }

// ---- generators
void* orc_gens_new(size_t num_cons, size_t num_vars, size_t num_inputs, const uint8_t* label, size_t ll) {
  return new OrcGens(num_cons, num_vars, num_inputs, std::string((const char*)label, ll));
}
void orc_gens_free(void* g) { delete (OrcGens*)g; }
size_t orc_gens_n(void* g) { return ((OrcGens*)g)->g.gens_n.n; }
// which: 0 gens_n.G[i], 1 gens_n.h, 2 gens_1.G[0], 3 gens_3.G[i], 4 gens_3.h, 5 gens_4.G[i], 6 gens_4.h
void orc_gens_point(void* gp, int which, size_t i, uint8_t out[32]) {
  NIZKGens& g = ((OrcGens*)gp)->g;
  const Pt* p = which == 0 ? &g.gens_n.G[i] : which == 1 ? &g.gens_n.h : which == 2 ? &g.gens_1.G[0]
              : which == 3 ? &g.gens_3.G[i] : which == 4 ? &g.gens_3.h : which == 5 ? &g.gens_4.G[i] : &g.gens_4.h;
  pt_compress(*p, out);
}

// ---- hot loops (arrays of Montgomery limbs)
void orc_eq_evals(const uint8_t* r, size_t ell, uint8_t* out) {
  FlVec e = eq_evals(ldv(r, ell));
  memcpy(out, e.data(), 32 * e.size());
}
void orc_bind_top(uint8_t* Z, size_t len, const uint8_t* r) {
  FlVec v = ldv(Z, len); bound_poly_var_top(v, ld(r)); memcpy(Z, v.data(), 32 * v.size());
}
void orc_sumcheck_round(const uint8_t* const* tables, int ntab, size_t len, uint8_t* out3) {
  std::vector<FlVec> t(ntab); std::vector<FlVec*> tp;
  for (int i = 0; i < ntab; i++) { t[i] = ldv(tables[i], len); tp.push_back(&t[i]); }
  Fl e0, e2, e3; sumcheck_round_evals(tp, ntab == 4 ? 3 : 2, e0, e2, e3);
  st(out3, e0); st(out3 + 32, e2); st(out3 + 64, e3);
}
void orc_dot(const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) {
  FlVec x = ldv(a, n), y = ldv(b, n); st(out, dotproduct(x.data(), y.data(), n));
}
void orc_poly_bound(const uint8_t* Z, size_t n, const uint8_t* Lv, size_t Ls, uint8_t* out) {
  FlVec z = ldv(Z, n), l = ldv(Lv, Ls); FlVec o = poly_bound(z, l, n / Ls); memcpy(out, o.data(), 32 * o.size());
}
void orc_unipoly_from_evals(const uint8_t* evals, size_t n, uint8_t* out) {
  FlVec c = unipoly_from_evals(ldv(evals, n)); memcpy(out, c.data(), 32 * c.size());
}
void orc_unipoly_eval(const uint8_t* coeffs, size_t n, const uint8_t* r, uint8_t* out) { st(out, unipoly_eval(ldv(coeffs, n), ld(r))); }
void orc_commit_rows(void* gp, const uint8_t* Z, size_t n, const uint8_t* blinds, size_t L, int threads, uint8_t* out) {
  std::vector<CPt> C = poly_commit_rows(ldv(Z, n), ldv(blinds, L), ((OrcGens*)gp)->g.gens_n, threads);
  for (size_t i = 0; i < L; i++) memcpy(out + 32 * i, C[i].b, 32);
}

// ---- instance
void* orc_instance_new(size_t num_cons, size_t num_vars, size_t num_inputs, const uint64_t* const rows[3],
                       const uint64_t* const cols[3], const uint8_t* const vals[3], const uint64_t nnz[3], int* err) {
  OrcInst* h = new OrcInst();
  int e = Instance::make(num_cons, num_vars, num_inputs, rows, cols, vals, nnz, h->inst);
  if (err) *err = e;
  if (e) { delete h; return nullptr; }
  return h;
}
// produce_synthetic_r1cs; vars_out/inputs_out: canonical 32-byte values
void* orc_synthetic_new(size_t num_cons, size_t num_vars, size_t num_inputs, uint64_t seed, uint8_t* vars_out,
                        uint8_t* inputs_out) {
  OrcInst* h = new OrcInst();
  FlVec v, in;
  synthetic_r1cs(num_cons, num_vars, num_inputs, seed, h->inst, v, in);
  for (size_t i = 0; i < v.size(); i++) fl_to_bytes(v[i], vars_out + 32 * i);
  for (size_t i = 0; i < in.size(); i++) fl_to_bytes(in[i], inputs_out + 32 * i);
  return h;
}
void orc_instance_free(void* h) { delete (OrcInst*)h; }
void orc_instance_dims(void* h, uint64_t out[6]) {
  Instance& I = ((OrcInst*)h)->inst;
  out[0] = I.num_cons; out[1] = I.num_vars; out[2] = I.num_inputs; out[3] = I.M[0].size(); out[4] = I.M[1].size(); out[5] = I.M[2].size();
}
// internal (padded-column) COO of matrix m; vals canonical bytes
void orc_instance_coo(void* h, int m, uint64_t* rows, uint64_t* cols, uint8_t* vals) {
  Instance& I = ((OrcInst*)h)->inst;
  for (size_t k = 0; k < I.M[m].size(); k++) { rows[k] = I.M[m][k].row; cols[k] = I.M[m][k].col; fl_to_bytes(I.M[m][k].val, vals + 32 * k); }
}
static bool load_assign(const uint8_t* b, size_t n, FlVec& out) {
  out.resize(n);
  for (size_t i = 0; i < n; i++) if (!fl_from_bytes(b + 32 * i, out[i])) return false;
  return true;
}
int orc_instance_is_sat(void* h, const uint8_t* vars, size_t nv, const uint8_t* inputs, size_t ni) {
  FlVec v, in;
  if (!load_assign(vars, nv, v) || !load_assign(inputs, ni, in)) return InvalidScalar;
  return ((OrcInst*)h)->inst.is_sat(v, in);
}
// z (len 2*num_vars, Montgomery) -> Az|Bz|Cz (3*num_cons)
void orc_multiply_vec(void* h, const uint8_t* z, uint8_t* out) {
  Instance& I = ((OrcInst*)h)->inst;
  FlVec zz = ldv(z, 2 * I.num_vars);
  for (int m = 0; m < 3; m++) { FlVec r = I.multiply_vec(m, zz); memcpy(out + 32 * I.num_cons * m, r.data(), 32 * I.num_cons); }
}
// r_A*A^T e + r_B*B^T e + r_C*C^T e  (r1csproof.rs:268-279); out len 2*num_vars
void orc_eval_table_abc(void* h, const uint8_t* rx_evals, const uint8_t* rabc, uint8_t* out) {
  Instance& I = ((OrcInst*)h)->inst;
  FlVec e = ldv(rx_evals, I.num_cons);
  size_t nc = 2 * I.num_vars;
  FlVec A = I.eval_table(0, e, nc), B = I.eval_table(1, e, nc), C = I.eval_table(2, e, nc);
  Fl rA = ld(rabc), rB = ld(rabc + 32), rC = ld(rabc + 64);
  for (size_t i = 0; i < nc; i++) st(out + 32 * i, fl_add(fl_add(fl_mul(rA, A[i]), fl_mul(rB, B[i])), fl_mul(rC, C[i])));
}

// ---- NIZK
// tape_seed64: 64 uniform bytes -> Scalar::from_bytes_wide (stands in for Scalar::random(OsRng), random.rs:13-15)
// times5 (optional): polycommit, phase1, phase2, polyeval, total seconds.  Returns malloc'd bincode of NIZK.
int orc_nizk_prove(void* h, const uint8_t* vars, size_t nv, const uint8_t* inputs, size_t ni, void* gp,
                   const uint8_t* label, size_t ll, const uint8_t tape_seed64[64], int threads, uint8_t** out,
                   size_t* out_len, double* times5) {
  FlVec v, in;
  if (!load_assign(vars, nv, v) || !load_assign(inputs, ni, in)) return InvalidScalar;
  try {
    Transcript tr(std::string((const char*)label, ll));
    PhaseTimes pt;
    NIZK nz = nizk_prove(((OrcInst*)h)->inst, v, in, ((OrcGens*)gp)->g, tr, fl_from_bytes_wide(tape_seed64), threads, &pt);
    Writer w; w.nizk(nz);
    *out = (uint8_t*)malloc(w.b.size());
    memcpy(*out, w.b.data(), w.b.size());
    *out_len = w.b.size();
    if (times5) { times5[0] = pt.polycommit; times5[1] = pt.phase1; times5[2] = pt.phase2; times5[3] = pt.polyeval; times5[4] = pt.total; }
    return 0;
  } catch (const std::exception&) { return -1; }
}
// 0 = accepted, -1 = rejected, -2 = malformed bytes
int orc_nizk_verify(void* h, const uint8_t* inputs, size_t ni, void* gp, const uint8_t* label, size_t ll,
                    const uint8_t* proof, size_t len) {
  FlVec in;
  if (!load_assign(inputs, ni, in)) return InvalidScalar;
  Reader r(proof, len);
  NIZK nz = r.nizk();
  if (!r.ok) return -2;
  Transcript tr(std::string((const char*)label, ll));
  return nizk_verify(nz, ((OrcInst*)h)->inst, in, ((OrcGens*)gp)->g, tr) ? 0 : -1;
}
// SNARK::encode (lib.rs:321-333).  num_nz_entries is SNARKGens::new's fourth argument (lib.rs:283).  Returns malloc'd
// bytes: num_cons, num_vars, num_inputs, then bincode of SparseMatPolyCommitment; -1 where the reference would panic.
// dense_ops / dense_mem (optional): canonical bytes of comb_ops (16 N scalars) / comb_mem (2 num_mem_cells).
int orc_snark_encode(void* h, size_t num_nz_entries, int threads, uint8_t** out, size_t* out_len, uint8_t* dense_ops,
                     uint8_t* dense_mem) {
  try {
    const Instance& inst = ((OrcInst*)h)->inst;
    R1CSCommitmentGens gens(inst.num_cons, inst.num_vars, inst.num_inputs, num_nz_entries);
    SparseAsDense d;
    R1CSCommitment c = snark_encode(inst, gens, threads, &d);
    Writer w; w.r1cs_commitment(c);
    *out = (uint8_t*)malloc(w.b.size());
    memcpy(*out, w.b.data(), w.b.size());
    *out_len = w.b.size();
    if (dense_ops) for (size_t i = 0; i < d.comb_ops.size(); i++) fl_to_bytes(d.comb_ops[i], dense_ops + 32 * i);
    if (dense_mem) for (size_t i = 0; i < d.comb_mem.size(); i++) fl_to_bytes(d.comb_mem[i], dense_mem + 32 * i);
    return 0;
  } catch (const std::exception&) { return -1; }
}
// the dense representation alone (no commitments): sizes out[0] = N, out[1] = num_mem_cells
int orc_snark_dense(void* h, uint64_t out2[2], uint8_t* dense_ops, uint8_t* dense_mem) {
  try {
    SparseAsDense d = sparse_to_dense(((OrcInst*)h)->inst);
    out2[0] = d.N; out2[1] = d.num_mem_cells;
    if (dense_ops) for (size_t i = 0; i < d.comb_ops.size(); i++) fl_to_bytes(d.comb_ops[i], dense_ops + 32 * i);
    if (dense_mem) for (size_t i = 0; i < d.comb_mem.size(); i++) fl_to_bytes(d.comb_mem[i], dense_mem + 32 * i);
    return 0;
  } catch (const std::exception&) { return -1; }
}
// commit_nondet_witness step of SparseMatPolyEvalProof::prove: rx / ry canonical scalars; out32: the L compressed
// points of DerefsCommitment.comm_ops_val (caller sizes it with cap_points; returns L, or -1 on a reference panic);
// comb_out (optional): canonical bytes of Derefs.comb (8 N scalars)
long orc_snark_commit_derefs(void* h, size_t num_nz_entries, const uint8_t* rx, size_t nrx, const uint8_t* ry, size_t nry,
                             int threads, uint8_t* out32, size_t cap_points, uint8_t* comb_out) {
  try {
    const Instance& inst = ((OrcInst*)h)->inst;
    FlVec vx, vy;
    if (!load_assign(rx, nrx, vx) || !load_assign(ry, nry, vy)) return InvalidScalar;
    R1CSCommitmentGens gens(inst.num_cons, inst.num_vars, inst.num_inputs, num_nz_entries);
    FlVec comb;
    std::vector<CPt> C = snark_commit_derefs(inst, gens, vx, vy, threads, &comb);
    if (C.size() > cap_points) return -1;
    for (size_t i = 0; i < C.size(); i++) memcpy(out32 + 32 * i, C[i].b, 32);
    if (comb_out) for (size_t i = 0; i < comb.size(); i++) fl_to_bytes(comb[i], comb_out + 32 * i);
    return (long)C.size();
  } catch (const std::exception&) { return -1; }
}
// build_layered_network of SparseMatPolyEvalProof::prove reduced to the 16 product-circuit evaluations
// (row side: init, read x3, write x3, audit; then the column side).  All scalars canonical.  -1 on a reference panic.
int orc_snark_network_products(void* h, const uint8_t* rx, size_t nrx, const uint8_t* ry, size_t nry, const uint8_t r_hash[32],
                               const uint8_t r_multiset[32], uint8_t out16[512]) {
  try {
    FlVec vx, vy, rr;
    uint8_t two[64];
    memcpy(two, r_hash, 32);
    memcpy(two + 32, r_multiset, 32);
    if (!load_assign(rx, nrx, vx) || !load_assign(ry, nry, vy) || !load_assign(two, 2, rr)) return InvalidScalar;
    Fl out[16];
    snark_network_products(((OrcInst*)h)->inst, vx, vy, rr[0], rr[1], out);
    for (int i = 0; i < 16; i++) fl_to_bytes(out[i], out16 + 32 * i);
    return 0;
  } catch (const std::exception&) { return -1; }
}

// SNARK::prove (lib.rs:335-403): the instance handle doubles as the decommitment (R1CSDecommitment is the dense form of the same
// matrices, rebuilt here).  Returns malloc'd bincode of SNARK.  times3 (optional): sat proof, eval proof, total seconds.
int orc_snark_prove(void* h, size_t num_nz_entries, const uint8_t* vars, size_t nv, const uint8_t* inputs, size_t ni, const uint8_t* label,
                    size_t ll, const uint8_t tape_seed64[64], int threads, uint8_t** out, size_t* out_len, double* times3) {
  FlVec v, in;
  if (!load_assign(vars, nv, v) || !load_assign(inputs, ni, in)) return InvalidScalar;
  try {
    const Instance& inst = ((OrcInst*)h)->inst;
    SNARKGens gens(inst.num_cons, inst.num_vars, inst.num_inputs, num_nz_entries);
    Transcript tr(std::string((const char*)label, ll));
    auto t0 = std::chrono::steady_clock::now();
    SNARK s = snark_prove(inst, v, in, gens, tr, fl_from_bytes_wide(tape_seed64), threads, times3);
    SnarkWriter w; w.snark(s);
    *out = (uint8_t*)malloc(w.b.size());
    memcpy(*out, w.b.data(), w.b.size());
    *out_len = w.b.size();
    if (times3) times3[2] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return 0;
  } catch (const std::exception& e) { if (getenv("ORC_VERBOSE")) fprintf(stderr, "orc_snark_prove: %s\n", e.what()); return -1; }
}
// SNARK::verify (lib.rs:405-457) against the commitment bytes orc_snark_encode returned.  0 = accepted, -1 = rejected, -2 = malformed
int orc_snark_verify(const uint8_t* comm, size_t comm_len, size_t num_nz_entries, const uint8_t* inputs, size_t ni, const uint8_t* label,
                     size_t ll, const uint8_t* proof, size_t len) {
  FlVec in;
  if (!load_assign(inputs, ni, in)) return InvalidScalar;
  try {
    Reader rc(comm, comm_len);
    R1CSCommitment c = rc.r1cs_commitment();
    if (!rc.ok) return -2;
    SnarkReader r(proof, len);
    SNARK s = r.snark();
    if (!r.ok) return -2;
    SNARKGens gens(c.num_cons, c.num_vars, c.num_inputs, num_nz_entries);
    Transcript tr(std::string((const char*)label, ll));
    return snark_verify(s, c, in, gens, tr) ? 0 : -1;
  } catch (const std::exception&) { return -1; }
}
void orc_free(void* p) { free(p); }

}  // extern "C"
