// ORACLE (test infrastructure only; never linked or included by otti_b200/).
// CPU restatement of the Spartan NIZK prove/verify path that Otti's run.py
// drives (run.py:58,100 -> spartan-zkinterface/src/main.rs:74-100 ->
// Spartan/src/lib.rs:493-574).  Every function cites the reference file:line
// (relative to /root/reference/Spartan/src) whose algorithm it follows.
//
// PARITY UNPINNED against a cargo build of the reference: no Rust toolchain in
// this image.  Pinned instead by the reference's F_l known answers
// (scalar/ristretto255.rs:784-1214), UniPoly / MLE KATs (unipoly.rs:123-182,
// dense_mlpoly.rs:430-449), merlin + RFC 9496 vectors, the README proof
// length (README.md:376) and the independent big-int restatement in
// oracle/pyspec/ via tests/golden/.
#pragma once
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "transcript.hpp"

namespace orc {

typedef std::vector<Fl> FlVec;
struct CPt {  // CompressedGroup
  uint8_t b[32];
  bool operator==(const CPt& o) const { return !memcmp(b, o.b, 32); }
};
static inline CPt compress(const Pt& p) {
  CPt c;
  pt_compress(p, c.b);
  return c;
}
static inline Pt unpack(const CPt& c) {  // group.rs:14-21; throws like .unwrap()
  Pt p;
  if (!pt_decompress(c.b, p)) throw std::runtime_error("DecompressionError");
  return p;
}
static inline size_t ilog2(size_t n) {
  size_t l = 0;
  while ((size_t(1) << (l + 1)) <= n) l++;
  return l;
}

// ---------------------------------------------------------------- generators
// commitments.rs:7-58
struct MultiCommitGens {
  size_t n;
  std::vector<Pt> G;
  Pt h;
  static MultiCommitGens make(size_t n, const std::string& label) {  // commitments.rs:15-33
    Shake256 sh;
    sh.absorb((const uint8_t*)label.data(), label.size());
    uint8_t base[32];
    pt_compress(pt_basepoint(), base);  // GROUP_BASEPOINT_COMPRESSED, group.rs:23-24
    sh.absorb(base, 32);
    std::vector<Pt> g(n + 1);
    for (size_t i = 0; i < n + 1; i++) {
      uint8_t ub[64];
      sh.squeeze(ub, 64);
      g[i] = pt_from_uniform_bytes(ub);
    }
    MultiCommitGens r;
    r.n = n;
    r.h = g[n];
    g.resize(n);
    r.G = g;
    return r;
  }
  void split_at(size_t mid, MultiCommitGens& a, MultiCommitGens& b) const {  // commitments.rs:43-58
    a.n = mid; a.G.assign(G.begin(), G.begin() + mid); a.h = h;
    b.n = n - mid; b.G.assign(G.begin() + mid, G.end()); b.h = h;
  }
};
// commitments.rs:65-84
static inline Pt commit_scalar(const Fl& x, const Fl& blind, const MultiCommitGens& g) {
  Fl s[2] = {x, blind};
  Pt p[2] = {g.G[0], g.h};
  return pt_msm(s, p, 2);
}
static inline Pt commit_vec(const Fl* x, size_t n, const Fl& blind, const MultiCommitGens& g) {
  return pt_add(pt_msm(x, g.G.data(), n), pt_mul(blind, g.h));
}

// lib.rs:459-477, r1csproof.rs:38-70, dense_mlpoly.rs:24-35, nizk/mod.rs:402-414
struct NIZKGens {
  MultiCommitGens gens_n, gens_1, gens_3, gens_4;
  size_t num_vars_padded;
  NIZKGens(size_t num_cons, size_t num_vars, size_t num_inputs, const std::string& label = "gens_r1cs_sat") {
    (void)num_cons;
    size_t nv = std::max(num_vars, num_inputs + 1);
    size_t p2 = 1;
    while (p2 < nv) p2 <<= 1;
    num_vars_padded = p2;
    size_t ell = ilog2(p2), right = ell - ell / 2, n = size_t(1) << right;
    MultiCommitGens::make(n + 1, label).split_at(n, gens_n, gens_1);
    gens_3 = MultiCommitGens::make(3, label);
    gens_4 = MultiCommitGens::make(4, label);
  }
};

// --------------------------------------------------------------- polynomials
// EqPolynomial::evals, dense_mlpoly.rs:67-83
static inline FlVec eq_evals(const FlVec& r) {
  size_t ell = r.size();
  FlVec ev(size_t(1) << ell, fl_one());
  size_t size = 1;
  for (size_t j = 0; j < ell; j++) {
    size *= 2;
    for (size_t i = size; i-- > 0;) {
      if (!(i & 1)) continue;
      Fl s = ev[i / 2];
      ev[i] = fl_mul(s, r[j]);
      ev[i - 1] = fl_sub(s, ev[i]);
    }
  }
  return ev;
}
// DensePolynomial::bound_poly_var_top, dense_mlpoly.rs:214-221
static inline void bound_poly_var_top(FlVec& Z, const Fl& r) {
  size_t n = Z.size() / 2;
  for (size_t i = 0; i < n; i++) Z[i] = fl_add(Z[i], fl_mul(r, fl_sub(Z[i + n], Z[i])));
  Z.resize(n);
}
static inline Fl dotproduct(const Fl* a, const Fl* b, size_t n) {  // nizk/mod.rs:305-308
  Fl s = fl_zero();
  for (size_t i = 0; i < n; i++) s = fl_add(s, fl_mul(a[i], b[i]));
  return s;
}
// UniPoly::from_evals, unipoly.rs:23-54
static inline FlVec unipoly_from_evals(const FlVec& e) {
  Fl two_inv = fl_invert(fl_from_usize(2));
  if (e.size() == 3) {
    Fl c = e[0];
    Fl a = fl_mul(two_inv, fl_add(fl_sub(fl_sub(e[2], e[1]), e[1]), c));
    Fl b = fl_sub(fl_sub(e[1], c), a);
    return {c, b, a};
  }
  Fl six_inv = fl_invert(fl_from_usize(6));
  Fl d = e[0];
  Fl t = fl_sub(e[3], e[2]); t = fl_sub(t, e[2]); t = fl_sub(t, e[2]);
  t = fl_add(t, e[1]); t = fl_add(t, e[1]); t = fl_add(t, e[1]); t = fl_sub(t, e[0]);
  Fl a = fl_mul(six_inv, t);
  Fl u = fl_add(e[0], e[0]);
  for (int i = 0; i < 5; i++) u = fl_sub(u, e[1]);
  for (int i = 0; i < 4; i++) u = fl_add(u, e[2]);
  u = fl_sub(u, e[3]);
  Fl b = fl_mul(two_inv, u);
  Fl c = fl_sub(fl_sub(fl_sub(e[1], d), a), b);
  return {d, c, b, a};
}
// UniPoly::evaluate, unipoly.rs:72-80
static inline Fl unipoly_eval(const FlVec& co, const Fl& r) {
  Fl ev = co[0], power = r;
  for (size_t i = 1; i < co.size(); i++) {
    ev = fl_add(ev, fl_mul(power, co[i]));
    power = fl_mul(power, r);
  }
  return ev;
}

// ------------------------------------------------------------------ instance
struct Entry {  // SparseMatEntry, sparse_mlpoly.rs:19-24
  size_t row, col;
  Fl val;
};
enum R1CSError { OK = 0, InvalidScalar = -5, InvalidIndex = -6, InvalidNumberOfInputs = -3, InvalidNumberOfVars = -4 };

struct Instance {
  size_t num_cons, num_vars, num_inputs;  // padded (lib.rs:130-158)
  std::vector<Entry> M[3];

  // Instance::new, lib.rs:121-225.  rows/cols/vals: COO triples, vals canonical 32 B.
  static int make(size_t num_cons, size_t num_vars, size_t num_inputs, const uint64_t* const rows[3],
                  const uint64_t* const cols[3], const uint8_t* const vals[3], const uint64_t nnz[3], Instance& out) {
    size_t nvp = std::max(num_vars, num_inputs + 1), p2 = 1;
    while (p2 < nvp) p2 <<= 1;
    nvp = p2;
    size_t ncp = num_cons;
    if (ncp == 0 || ncp == 1) ncp = 2;
    size_t npt = 1;
    while (npt < num_cons) npt <<= 1;  // usize::next_power_of_two (0 -> 1)
    if (npt != num_cons) ncp = npt;
    out.num_cons = ncp; out.num_vars = nvp; out.num_inputs = num_inputs;
    for (int m = 0; m < 3; m++) {
      out.M[m].clear();
      for (uint64_t k = 0; k < nnz[m]; k++) {
        size_t row = rows[m][k], col = cols[m][k];
        if (row >= num_cons) return InvalidIndex;
        if (col >= num_vars + 1 + num_inputs) return InvalidIndex;
        Fl v;
        if (!fl_from_bytes(vals[m] + 32 * k, v)) return InvalidScalar;
        out.M[m].push_back(Entry{row, col >= num_vars ? col + nvp - num_vars : col, v});
      }
      if (num_cons == 0 || num_cons == 1)
        for (size_t i = nnz[m]; i < ncp; i++) out.M[m].push_back(Entry{i, num_vars, fl_zero()});
    }
    return OK;
  }
  // SparseMatPolynomial::multiply_vec, sparse_mlpoly.rs:452-466
  FlVec multiply_vec(int m, const FlVec& z) const {
    FlVec out(num_cons, fl_zero());
    for (const Entry& e : M[m]) out[e.row] = fl_add(out[e.row], fl_mul(e.val, z[e.col]));
    return out;
  }
  // SparseMatPolynomial::compute_eval_table_sparse, sparse_mlpoly.rs:468-483
  FlVec eval_table(int m, const FlVec& rx, size_t num_cols) const {
    FlVec out(num_cols, fl_zero());
    for (const Entry& e : M[m]) out[e.col] = fl_add(out[e.col], fl_mul(rx[e.row], e.val));
    return out;
  }
  // R1CSInstance::evaluate -> multi_evaluate, r1csinstance.rs:280-283, sparse_mlpoly.rs:425-450
  void evaluate(const FlVec& rx, const FlVec& ry, Fl out[3]) const {
    FlVec ex = eq_evals(rx), ey = eq_evals(ry);
    for (int m = 0; m < 3; m++) {
      Fl s = fl_zero();
      for (const Entry& e : M[m]) s = fl_add(s, fl_mul(fl_mul(ex[e.row], ey[e.col]), e.val));
      out[m] = s;
    }
  }
  // Instance::is_sat lib.rs:228-257 + R1CSInstance::is_sat r1csinstance.rs:216-246
  int is_sat(const FlVec& vars, const FlVec& inputs) const {
    if (vars.size() > num_vars) return InvalidNumberOfInputs;
    if (inputs.size() != num_inputs) return InvalidNumberOfInputs;
    FlVec z = vars;
    z.resize(num_vars, fl_zero());
    z.push_back(fl_one());
    z.insert(z.end(), inputs.begin(), inputs.end());
    z.resize(2 * num_vars, fl_zero());
    FlVec Az = multiply_vec(0, z), Bz = multiply_vec(1, z), Cz = multiply_vec(2, z);
    for (size_t i = 0; i < num_cons; i++)
      if (fl_mul(Az[i], Bz[i]) != Cz[i]) return 0;
    return 1;
  }
};

static inline uint64_t splitmix64(uint64_t& st) {
  st += 0x9E3779B97F4A7C15ULL;
  uint64_t z = st;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
// R1CSInstance::produce_synthetic_r1cs, r1csinstance.rs:136-214; OsRng replaced by
// SplitMix64(seed), Scalar::random = from_u512(8 x next_u64) (ristretto255.rs:381-387).
// The per-row inversion is done with one batch inversion (value-identical).
static inline void synthetic_r1cs(size_t num_cons, size_t num_vars, size_t num_inputs, uint64_t seed, Instance& inst,
                                  FlVec& vars, FlVec& inputs) {
  size_t size_z = num_vars + num_inputs + 1;
  FlVec Z(size_z);
  uint64_t st = seed;
  for (size_t i = 0; i < size_z; i++) {
    uint64_t l[8];
    for (int k = 0; k < 8; k++) l[k] = splitmix64(st);
    Z[i] = fl_from_u512(l);
  }
  Z[num_vars] = fl_one();
  // batch-invert Z (zeros skipped)
  FlVec Zinv(size_z), pre(size_z);
  Fl acc = fl_one();
  for (size_t i = 0; i < size_z; i++) {
    pre[i] = acc;
    if (!Z[i].is_zero()) acc = fl_mul(acc, Z[i]);
  }
  acc = fl_invert(acc);
  for (size_t i = size_z; i-- > 0;) {
    if (Z[i].is_zero()) { Zinv[i] = fl_zero(); continue; }
    Zinv[i] = fl_mul(acc, pre[i]);
    acc = fl_mul(acc, Z[i]);
  }
  inst.num_cons = num_cons; inst.num_vars = num_vars; inst.num_inputs = num_inputs;
  for (int m = 0; m < 3; m++) { inst.M[m].clear(); inst.M[m].reserve(num_cons); }
  for (size_t i = 0; i < num_cons; i++) {
    size_t ai = i % size_z, bi = (i + 2) % size_z, ci = (i + 3) % size_z;
    inst.M[0].push_back(Entry{i, ai, fl_one()});
    inst.M[1].push_back(Entry{i, bi, fl_one()});
    Fl ab = fl_mul(Z[ai], Z[bi]);
    if (Z[ci].is_zero()) inst.M[2].push_back(Entry{i, num_vars, ab});
    else inst.M[2].push_back(Entry{i, ci, fl_mul(ab, Zinv[ci])});
  }
  vars.assign(Z.begin(), Z.begin() + num_vars);
  inputs.assign(Z.begin() + num_vars + 1, Z.end());
}

// ----------------------------------------------------------- proof structures
struct KnowledgeProof { CPt alpha; Fl z1, z2; };                          // nizk/mod.rs:14-19
struct EqualityProof { CPt alpha; Fl z; };                                // nizk/mod.rs:76-80
struct ProductProof { CPt alpha, beta, delta; Fl z[5]; };                 // nizk/mod.rs:145-151
struct DotProductProof { CPt delta, beta; FlVec z; Fl z_delta, z_beta; }; // nizk/mod.rs:291-298
struct BulletReductionProof { std::vector<CPt> L_vec, R_vec; };           // nizk/bullet.rs:14-18
struct DotProductProofLog { BulletReductionProof bullet; CPt delta, beta; Fl z1, z2; };  // nizk/mod.rs:416-423
struct ZKSumcheckProof { std::vector<CPt> comm_polys, comm_evals; std::vector<DotProductProof> proofs; };  // sumcheck.rs:64-69
struct R1CSProof {  // r1csproof.rs:20-36
  std::vector<CPt> comm_vars;
  ZKSumcheckProof sc_proof_phase1;
  CPt claims_phase2[4];
  KnowledgeProof pok_Cz;
  ProductProof proof_prod;
  EqualityProof proof_eq_sc_phase1;
  ZKSumcheckProof sc_proof_phase2;
  CPt comm_vars_at_ry;
  DotProductProofLog proof_eval_vars_at_ry;
  EqualityProof proof_eq_sc_phase2;
};
struct NIZK {  // lib.rs:481-485
  R1CSProof r1cs_sat_proof;
  FlVec rx, ry;
};

// ------------------------------------------------------------ sigma protocols
// KnowledgeProof::prove, nizk/mod.rs:26-51
static inline KnowledgeProof knowledge_prove(const MultiCommitGens& g, Transcript& tr, RandomTape& tape, const Fl& x,
                                             const Fl& r, CPt& C) {
  tr.append_protocol_name("knowledge proof");
  Fl t1 = tape.random_scalar("t1"), t2 = tape.random_scalar("t2");
  C = compress(commit_scalar(x, r, g));
  tr.append_point("C", C.b);
  KnowledgeProof p;
  p.alpha = compress(commit_scalar(t1, t2, g));
  tr.append_point("alpha", p.alpha.b);
  Fl c = tr.challenge_scalar("c");
  p.z1 = fl_add(fl_mul(x, c), t1);
  p.z2 = fl_add(fl_mul(r, c), t2);
  return p;
}
// KnowledgeProof::verify, nizk/mod.rs:53-73
static inline bool knowledge_verify(const KnowledgeProof& p, const MultiCommitGens& g, Transcript& tr, const CPt& C) {
  tr.append_protocol_name("knowledge proof");
  tr.append_point("C", C.b);
  tr.append_point("alpha", p.alpha.b);
  Fl c = tr.challenge_scalar("c");
  CPt lhs = compress(commit_scalar(p.z1, p.z2, g));
  CPt rhs = compress(pt_add(pt_mul(c, unpack(C)), unpack(p.alpha)));
  return lhs == rhs;
}
// EqualityProof::prove, nizk/mod.rs:87-115
static inline EqualityProof equality_prove(const MultiCommitGens& g, Transcript& tr, RandomTape& tape, const Fl& v1,
                                           const Fl& s1, const Fl& v2, const Fl& s2) {
  tr.append_protocol_name("equality proof");
  Fl r = tape.random_scalar("r");
  CPt C1 = compress(commit_scalar(v1, s1, g));
  tr.append_point("C1", C1.b);
  CPt C2 = compress(commit_scalar(v2, s2, g));
  tr.append_point("C2", C2.b);
  EqualityProof p;
  p.alpha = compress(pt_mul(r, g.h));
  tr.append_point("alpha", p.alpha.b);
  Fl c = tr.challenge_scalar("c");
  p.z = fl_add(fl_mul(c, fl_sub(s1, s2)), r);
  return p;
}
// EqualityProof::verify, nizk/mod.rs:117-142
static inline bool equality_verify(const EqualityProof& p, const MultiCommitGens& g, Transcript& tr, const CPt& C1,
                                   const CPt& C2) {
  tr.append_protocol_name("equality proof");
  tr.append_point("C1", C1.b);
  tr.append_point("C2", C2.b);
  tr.append_point("alpha", p.alpha.b);
  Fl c = tr.challenge_scalar("c");
  Pt C = pt_sub(unpack(C1), unpack(C2));
  CPt rhs = compress(pt_add(pt_mul(c, C), unpack(p.alpha)));
  CPt lhs = compress(pt_mul(p.z, g.h));
  return lhs == rhs;
}
// ProductProof::prove, nizk/mod.rs:158-228
static inline ProductProof product_prove(const MultiCommitGens& g, Transcript& tr, RandomTape& tape, const Fl& x,
                                         const Fl& rX, const Fl& y, const Fl& rY, const Fl& z, const Fl& rZ, CPt& X,
                                         CPt& Y, CPt& Z) {
  tr.append_protocol_name("product proof");
  Fl b1 = tape.random_scalar("b1"), b2 = tape.random_scalar("b2"), b3 = tape.random_scalar("b3"),
     b4 = tape.random_scalar("b4"), b5 = tape.random_scalar("b5");
  X = compress(commit_scalar(x, rX, g)); tr.append_point("X", X.b);
  Y = compress(commit_scalar(y, rY, g)); tr.append_point("Y", Y.b);
  Z = compress(commit_scalar(z, rZ, g)); tr.append_point("Z", Z.b);
  ProductProof p;
  p.alpha = compress(commit_scalar(b1, b2, g)); tr.append_point("alpha", p.alpha.b);
  p.beta = compress(commit_scalar(b3, b4, g)); tr.append_point("beta", p.beta.b);
  MultiCommitGens gX;
  gX.n = 1; gX.G = {unpack(X)}; gX.h = g.h;
  p.delta = compress(commit_scalar(b3, b5, gX)); tr.append_point("delta", p.delta.b);
  Fl c = tr.challenge_scalar("c");
  p.z[0] = fl_add(b1, fl_mul(c, x));
  p.z[1] = fl_add(b2, fl_mul(c, rX));
  p.z[2] = fl_add(b3, fl_mul(c, y));
  p.z[3] = fl_add(b4, fl_mul(c, rY));
  p.z[4] = fl_add(b5, fl_mul(c, fl_sub(rZ, fl_mul(rX, y))));
  return p;
}
// ProductProof::verify, nizk/mod.rs:230-288
static inline bool product_verify(const ProductProof& p, const MultiCommitGens& g, Transcript& tr, const CPt& X,
                                  const CPt& Y, const CPt& Z) {
  tr.append_protocol_name("product proof");
  tr.append_point("X", X.b); tr.append_point("Y", Y.b); tr.append_point("Z", Z.b);
  tr.append_point("alpha", p.alpha.b); tr.append_point("beta", p.beta.b); tr.append_point("delta", p.delta.b);
  Fl c = tr.challenge_scalar("c");
  auto chk = [&](const CPt& P, const CPt& Xc, const MultiCommitGens& gg, const Fl& za, const Fl& zb) {
    CPt lhs = compress(pt_add(unpack(P), pt_mul(c, unpack(Xc))));
    return lhs == compress(commit_scalar(za, zb, gg));
  };
  MultiCommitGens gX;
  gX.n = 1; gX.G = {unpack(X)}; gX.h = g.h;
  return chk(p.alpha, X, g, p.z[0], p.z[1]) && chk(p.beta, Y, g, p.z[2], p.z[3]) && chk(p.delta, Z, gX, p.z[2], p.z[4]);
}
// DotProductProof::prove, nizk/mod.rs:310-367
static inline DotProductProof dotproduct_prove(const MultiCommitGens& g1, const MultiCommitGens& gn, Transcript& tr,
                                               RandomTape& tape, const FlVec& x, const Fl& blind_x, const FlVec& a,
                                               const Fl& y, const Fl& blind_y) {
  tr.append_protocol_name("dot product proof");
  size_t n = x.size();
  FlVec d = tape.random_vector("d_vec", n);
  Fl r_delta = tape.random_scalar("r_delta"), r_beta = tape.random_scalar("r_beta");
  CPt Cx = compress(commit_vec(x.data(), n, blind_x, gn)); tr.append_point("Cx", Cx.b);
  CPt Cy = compress(commit_scalar(y, blind_y, g1)); tr.append_point("Cy", Cy.b);
  DotProductProof p;
  p.delta = compress(commit_vec(d.data(), n, r_delta, gn)); tr.append_point("delta", p.delta.b);
  Fl dp = dotproduct(a.data(), d.data(), n);
  p.beta = compress(commit_scalar(dp, r_beta, g1)); tr.append_point("beta", p.beta.b);
  Fl c = tr.challenge_scalar("c");
  p.z.resize(n);
  for (size_t i = 0; i < n; i++) p.z[i] = fl_add(fl_mul(c, x[i]), d[i]);
  p.z_delta = fl_add(fl_mul(c, blind_x), r_delta);
  p.z_beta = fl_add(fl_mul(c, blind_y), r_beta);
  return p;
}
// DotProductProof::verify, nizk/mod.rs:369-400
static inline bool dotproduct_verify(const DotProductProof& p, const MultiCommitGens& g1, const MultiCommitGens& gn,
                                     Transcript& tr, const FlVec& a, const CPt& Cx, const CPt& Cy) {
  if (gn.n != a.size() || p.z.size() != a.size()) return false;
  tr.append_protocol_name("dot product proof");
  tr.append_point("Cx", Cx.b); tr.append_point("Cy", Cy.b);
  tr.append_point("delta", p.delta.b); tr.append_point("beta", p.beta.b);
  Fl c = tr.challenge_scalar("c");
  bool ok = pt_eq(pt_add(pt_mul(c, unpack(Cx)), unpack(p.delta)), commit_vec(p.z.data(), p.z.size(), p.z_delta, gn));
  Fl dz = dotproduct(p.z.data(), a.data(), a.size());
  ok &= pt_eq(pt_add(pt_mul(c, unpack(Cy)), unpack(p.beta)), commit_scalar(dz, p.z_beta, g1));
  return ok;
}
// BulletReductionProof::prove, nizk/bullet.rs:31-131
static inline BulletReductionProof bullet_prove(Transcript& tr, const Pt& Q, const std::vector<Pt>& G_vec, const Pt& H,
                                                const FlVec& a_vec, const FlVec& b_vec, const Fl& blind,
                                                const std::vector<std::pair<Fl, Fl>>& blinds_vec, Fl& a_out, Fl& b_out,
                                                Pt& G_out, Fl& blind_out) {
  std::vector<Pt> G = G_vec;
  FlVec a = a_vec, b = b_vec;
  size_t n = G.size(), round = 0;
  BulletReductionProof pf;
  Fl blind_fin = blind;
  while (n != 1) {
    n /= 2;
    Fl cL = dotproduct(a.data(), b.data() + n, n), cR = dotproduct(a.data() + n, b.data(), n);
    const Fl& blind_L = blinds_vec[round].first;
    const Fl& blind_R = blinds_vec[round].second;
    round++;
    FlVec sL(a.begin(), a.begin() + n), sR(a.begin() + n, a.begin() + 2 * n);
    sL.push_back(cL); sL.push_back(blind_L);
    sR.push_back(cR); sR.push_back(blind_R);
    std::vector<Pt> pL(G.begin() + n, G.begin() + 2 * n), pR(G.begin(), G.begin() + n);
    pL.push_back(Q); pL.push_back(H);
    pR.push_back(Q); pR.push_back(H);
    Pt L = pt_msm(sL.data(), pL.data(), n + 2), R = pt_msm(sR.data(), pR.data(), n + 2);
    CPt Lc = compress(L), Rc = compress(R);
    tr.append_point("L", Lc.b);
    tr.append_point("R", Rc.b);
    Fl u = tr.challenge_scalar("u"), u_inv = fl_invert(u);
    for (size_t i = 0; i < n; i++) {
      a[i] = fl_add(fl_mul(a[i], u), fl_mul(u_inv, a[n + i]));
      b[i] = fl_add(fl_mul(b[i], u_inv), fl_mul(u, b[n + i]));
      G[i] = pt_add(pt_mul(u_inv, G[i]), pt_mul(u, G[n + i]));
    }
    blind_fin = fl_add(fl_add(blind_fin, fl_mul(fl_mul(blind_L, u), u)), fl_mul(fl_mul(blind_R, u_inv), u_inv));
    pf.L_vec.push_back(Lc);
    pf.R_vec.push_back(Rc);
    a.resize(n); b.resize(n); G.resize(n);
  }
  a_out = a[0]; b_out = b[0]; G_out = G[0]; blind_out = blind_fin;
  return pf;
}
// BulletReductionProof::verify (+verification_scalars), nizk/bullet.rs:136-224
static inline bool bullet_verify(const BulletReductionProof& pf, size_t n, const FlVec& a, Transcript& tr,
                                 const Pt& Gamma, const std::vector<Pt>& G, Pt& G_hat, Pt& Gamma_hat, Fl& a_hat) {
  size_t lg_n = pf.L_vec.size();
  if (lg_n >= 32 || pf.R_vec.size() != lg_n) return false;
  if (n != (size_t(1) << lg_n)) return false;
  FlVec ch(lg_n), ch_inv(lg_n);
  for (size_t i = 0; i < lg_n; i++) {
    tr.append_point("L", pf.L_vec[i].b);
    tr.append_point("R", pf.R_vec[i].b);
    ch[i] = tr.challenge_scalar("u");
  }
  Fl allinv = fl_one();
  for (size_t i = 0; i < lg_n; i++) {
    ch_inv[i] = fl_invert(ch[i]);  // Scalar::batch_invert: same values
    allinv = fl_mul(allinv, ch_inv[i]);
  }
  FlVec u_sq(lg_n), u_inv_sq(lg_n);
  for (size_t i = 0; i < lg_n; i++) { u_sq[i] = fl_square(ch[i]); u_inv_sq[i] = fl_square(ch_inv[i]); }
  FlVec s(n);
  s[0] = allinv;
  for (size_t i = 1; i < n; i++) {
    size_t lg_i = ilog2(i), k = size_t(1) << lg_i;
    s[i] = fl_mul(s[i - k], u_sq[(lg_n - 1) - lg_i]);
  }
  std::vector<Pt> pts;
  FlVec sc;
  for (size_t i = 0; i < lg_n; i++) { Pt p; if (!pt_decompress(pf.L_vec[i].b, p)) return false; pts.push_back(p); sc.push_back(u_sq[i]); }
  for (size_t i = 0; i < lg_n; i++) { Pt p; if (!pt_decompress(pf.R_vec[i].b, p)) return false; pts.push_back(p); sc.push_back(u_inv_sq[i]); }
  pts.push_back(Gamma); sc.push_back(fl_one());
  G_hat = pt_msm(s.data(), G.data(), n);
  a_hat = dotproduct(a.data(), s.data(), n);
  Gamma_hat = pt_msm(sc.data(), pts.data(), pts.size());
  return true;
}
// DotProductProofLog::prove, nizk/mod.rs:435-512
static inline DotProductProofLog dotproductlog_prove(const MultiCommitGens& gens_n, const MultiCommitGens& gens_1,
                                                     Transcript& tr, RandomTape& tape, const FlVec& x, const Fl& blind_x,
                                                     const FlVec& a, const Fl& y, const Fl& blind_y, CPt& Cx, CPt& Cy) {
  tr.append_protocol_name("dot product proof (log)");
  size_t n = x.size();
  Fl d = tape.random_scalar("d"), r_delta = tape.random_scalar("r_delta");
  Fl r_beta = tape.random_scalar("r_delta");  // sic, nizk/mod.rs:454
  FlVec v1 = tape.random_vector("blinds_vec_1", 2 * ilog2(n)), v2 = tape.random_vector("blinds_vec_2", 2 * ilog2(n));
  std::vector<std::pair<Fl, Fl>> blinds_vec;
  for (size_t i = 0; i < v1.size(); i++) blinds_vec.push_back({v1[i], v2[i]});
  Cx = compress(commit_vec(x.data(), n, blind_x, gens_n)); tr.append_point("Cx", Cx.b);
  Cy = compress(commit_scalar(y, blind_y, gens_1)); tr.append_point("Cy", Cy.b);
  Fl blind_Gamma = fl_add(blind_x, blind_y);
  DotProductProofLog p;
  Fl x_hat, a_hat, rhat;
  Pt g_hat;
  p.bullet = bullet_prove(tr, gens_1.G[0], gens_n.G, gens_n.h, x, a, blind_Gamma, blinds_vec, x_hat, a_hat, g_hat, rhat);
  Fl y_hat = fl_mul(x_hat, a_hat);
  MultiCommitGens gh;
  gh.n = 1; gh.G = {g_hat}; gh.h = gens_1.h;
  p.delta = compress(commit_scalar(d, r_delta, gh)); tr.append_point("delta", p.delta.b);
  p.beta = compress(commit_scalar(d, r_beta, gens_1)); tr.append_point("beta", p.beta.b);
  Fl c = tr.challenge_scalar("c");
  p.z1 = fl_add(d, fl_mul(c, y_hat));
  p.z2 = fl_add(fl_mul(a_hat, fl_add(fl_mul(c, rhat), r_beta)), r_delta);
  return p;
}
// DotProductProofLog::verify, nizk/mod.rs:514-558
static inline bool dotproductlog_verify(const DotProductProofLog& p, size_t n, const MultiCommitGens& gens_n,
                                        const MultiCommitGens& gens_1, Transcript& tr, const FlVec& a, const CPt& Cx,
                                        const CPt& Cy) {
  tr.append_protocol_name("dot product proof (log)");
  tr.append_point("Cx", Cx.b);
  tr.append_point("Cy", Cy.b);
  Pt Gamma = pt_add(unpack(Cx), unpack(Cy));
  Pt g_hat, Gamma_hat;
  Fl a_hat;
  if (!bullet_verify(p.bullet, n, a, tr, Gamma, gens_n.G, g_hat, Gamma_hat, a_hat)) return false;
  tr.append_point("delta", p.delta.b);
  tr.append_point("beta", p.beta.b);
  Fl c = tr.challenge_scalar("c");
  CPt lhs = compress(pt_add(pt_mul(a_hat, pt_add(pt_mul(c, Gamma_hat), unpack(p.beta))), unpack(p.delta)));
  CPt rhs = compress(pt_add(pt_mul(p.z1, pt_add(g_hat, pt_mul(a_hat, gens_1.G[0]))), pt_mul(p.z2, gens_1.h)));
  return lhs == rhs;
}

// ------------------------------------------------------------------ sum-check
// One round's evaluation points.  degree 3: sumcheck.rs:622-650 with
// comb = A*(B*C-D) (r1csproof.rs:84-88); degree 2: sumcheck.rs:454-470 with comb = A*B.
static inline void sumcheck_round_evals(const std::vector<FlVec*>& T, int degree, Fl& e0, Fl& e2, Fl& e3) {
  size_t len = T[0]->size() / 2;
  e0 = e2 = e3 = fl_zero();
  if (degree == 3) {
    const FlVec &A = *T[0], &B = *T[1], &C = *T[2], &D = *T[3];
    for (size_t i = 0; i < len; i++) {
      e0 = fl_add(e0, fl_mul(A[i], fl_sub(fl_mul(B[i], C[i]), D[i])));
      Fl a2 = fl_sub(fl_add(A[len + i], A[len + i]), A[i]), b2 = fl_sub(fl_add(B[len + i], B[len + i]), B[i]),
         c2 = fl_sub(fl_add(C[len + i], C[len + i]), C[i]), d2 = fl_sub(fl_add(D[len + i], D[len + i]), D[i]);
      e2 = fl_add(e2, fl_mul(a2, fl_sub(fl_mul(b2, c2), d2)));
      Fl a3 = fl_sub(fl_add(a2, A[len + i]), A[i]), b3 = fl_sub(fl_add(b2, B[len + i]), B[i]),
         c3 = fl_sub(fl_add(c2, C[len + i]), C[i]), d3 = fl_sub(fl_add(d2, D[len + i]), D[i]);
      e3 = fl_add(e3, fl_mul(a3, fl_sub(fl_mul(b3, c3), d3)));
    }
  } else {
    const FlVec &A = *T[0], &B = *T[1];
    for (size_t i = 0; i < len; i++) {
      e0 = fl_add(e0, fl_mul(A[i], B[i]));
      Fl a2 = fl_sub(fl_add(A[len + i], A[len + i]), A[i]), b2 = fl_sub(fl_add(B[len + i], B[len + i]), B[i]);
      e2 = fl_add(e2, fl_mul(a2, b2));
    }
  }
}
// ZKSumcheckInstanceProof::prove_quad / prove_cubic_with_additive_term, sumcheck.rs:426-774
static inline ZKSumcheckProof zk_sumcheck_prove(const Fl& claim, const Fl& blind_claim, size_t num_rounds,
                                                std::vector<FlVec*> T, int degree, const MultiCommitGens& g1,
                                                const MultiCommitGens& gn, Transcript& tr, RandomTape& tape, FlVec& r,
                                                FlVec& final_claims, Fl& final_blind) {
  FlVec blinds_poly = tape.random_vector("blinds_poly", num_rounds);
  FlVec blinds_evals = tape.random_vector("blinds_evals", num_rounds);
  Fl claim_per_round = claim;
  CPt comm_claim_per_round = compress(commit_scalar(claim, blind_claim, g1));
  ZKSumcheckProof pf;
  r.clear();
  for (size_t j = 0; j < num_rounds; j++) {
    Fl e0, e2, e3;
    sumcheck_round_evals(T, degree, e0, e2, e3);
    FlVec evals = {e0, fl_sub(claim_per_round, e0), e2};
    if (degree == 3) evals.push_back(e3);
    FlVec poly = unipoly_from_evals(evals);
    CPt comm_poly = compress(commit_vec(poly.data(), poly.size(), blinds_poly[j], gn));
    tr.append_point("comm_poly", comm_poly.b);
    pf.comm_polys.push_back(comm_poly);
    Fl r_j = tr.challenge_scalar("challenge_nextround");
    for (FlVec* t : T) bound_poly_var_top(*t, r_j);
    Fl eval = unipoly_eval(poly, r_j);
    CPt comm_eval = compress(commit_scalar(eval, blinds_evals[j], g1));
    tr.append_point("comm_claim_per_round", comm_claim_per_round.b);
    tr.append_point("comm_eval", comm_eval.b);
    FlVec w = tr.challenge_vector("combine_two_claims_to_one", 2);
    Fl target = fl_add(fl_mul(w[0], claim_per_round), fl_mul(w[1], eval));
    Pt cts[2] = {unpack(comm_claim_per_round), unpack(comm_eval)};
    CPt comm_target = compress(pt_msm(w.data(), cts, 2));
    const Fl& blind_sc = j == 0 ? blind_claim : blinds_evals[j - 1];
    Fl blind = fl_add(fl_mul(w[0], blind_sc), fl_mul(w[1], blinds_evals[j]));
    if (!(compress(commit_scalar(target, blind, g1)) == comm_target)) throw std::runtime_error("sumcheck target assert");
    size_t nc = poly.size();
    FlVec a(nc);
    Fl pw = fl_one();
    for (size_t k = 0; k < nc; k++) {
      Fl a_sc = k == 0 ? fl_add(fl_one(), fl_one()) : fl_one();
      a[k] = fl_add(fl_mul(w[0], a_sc), fl_mul(w[1], pw));
      pw = fl_mul(pw, r_j);
    }
    pf.proofs.push_back(dotproduct_prove(g1, gn, tr, tape, poly, blinds_poly[j], a, target, blind));
    claim_per_round = eval;
    comm_claim_per_round = comm_eval;
    r.push_back(r_j);
    pf.comm_evals.push_back(comm_eval);
  }
  final_claims.clear();
  for (FlVec* t : T) final_claims.push_back((*t)[0]);
  final_blind = blinds_evals[num_rounds - 1];
  return pf;
}
// ZKSumcheckInstanceProof::verify, sumcheck.rs:84-177
static inline bool zk_sumcheck_verify(const ZKSumcheckProof& pf, const CPt& comm_claim, size_t num_rounds, size_t degree,
                                      const MultiCommitGens& g1, const MultiCommitGens& gn, Transcript& tr,
                                      CPt& comm_final, FlVec& r) {
  if (gn.n != degree + 1) return false;
  if (pf.comm_polys.size() != num_rounds || pf.comm_evals.size() != num_rounds || pf.proofs.size() != num_rounds)
    return false;
  r.clear();
  for (size_t i = 0; i < num_rounds; i++) {
    tr.append_point("comm_poly", pf.comm_polys[i].b);
    Fl r_i = tr.challenge_scalar("challenge_nextround");
    const CPt& ccpr = i == 0 ? comm_claim : pf.comm_evals[i - 1];
    tr.append_point("comm_claim_per_round", ccpr.b);
    tr.append_point("comm_eval", pf.comm_evals[i].b);
    FlVec w = tr.challenge_vector("combine_two_claims_to_one", 2);
    Pt cts[2] = {unpack(ccpr), unpack(pf.comm_evals[i])};
    CPt comm_target = compress(pt_msm(w.data(), cts, 2));
    FlVec a(degree + 1);
    Fl pw = fl_one();
    for (size_t k = 0; k <= degree; k++) {
      Fl a_sc = k == 0 ? fl_add(fl_one(), fl_one()) : fl_one();
      a[k] = fl_add(fl_mul(w[0], a_sc), fl_mul(w[1], pw));
      pw = fl_mul(pw, r_i);
    }
    if (!dotproduct_verify(pf.proofs[i], g1, gn, tr, a, pf.comm_polys[i], comm_target)) return false;
    r.push_back(r_i);
  }
  comm_final = pf.comm_evals[num_rounds - 1];
  return true;
}

// ------------------------------------------------------------------- commit
// DensePolynomial::commit_inner, dense_mlpoly.rs:147-176 (threads > 1 mirrors the
// `multicore` feature: parallel over rows only)
static inline std::vector<CPt> poly_commit_rows(const FlVec& Z, const FlVec& blinds, const MultiCommitGens& gn,
                                                int threads) {
  size_t L = blinds.size(), R = Z.size() / L;
  std::vector<CPt> C(L);
  auto work = [&](size_t lo, size_t hi) {
    for (size_t i = lo; i < hi; i++) C[i] = compress(commit_vec(&Z[R * i], R, blinds[i], gn));
  };
  if (threads <= 1 || L < 2) work(0, L);
  else {
    std::vector<std::thread> th;
    size_t nt = std::min<size_t>(threads, L);
    for (size_t t = 0; t < nt; t++) th.emplace_back(work, L * t / nt, L * (t + 1) / nt);
    for (auto& t : th) t.join();
  }
  return C;
}
// DensePolynomial::bound, dense_mlpoly.rs:205-212
static inline FlVec poly_bound(const FlVec& Z, const FlVec& Lv, size_t R) {
  FlVec out(R, fl_zero());
  for (size_t j = 0; j < Lv.size(); j++)
    for (size_t i = 0; i < R; i++) out[i] = fl_add(out[i], fl_mul(Lv[j], Z[j * R + i]));
  return out;
}

struct PhaseTimes { double polycommit, phase1, phase2, polyeval, total; };

// R1CSProof::prove r1csproof.rs:141-345 (shared by NIZK::prove and SNARK::prove; `vars` already padded).  Returns the proof, rx, ry.
static inline R1CSProof r1cs_prove(const Instance& inst, const FlVec& vars, const FlVec& inputs, const NIZKGens& gens, Transcript& tr,
                                   RandomTape& tape, int threads, FlVec& rx, FlVec& ry, PhaseTimes* times = nullptr) {
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  double t0 = now();
  tr.append_protocol_name("R1CS proof");
  if (!(inputs.size() < vars.size())) throw std::runtime_error("assert input.len() < vars.len()");  // r1csproof.rs:153
  R1CSProof pf;
  size_t nv = vars.size(), ell = ilog2(nv), lnv = ell / 2, Ls = size_t(1) << lnv, Rs = size_t(1) << (ell - lnv);
  // polycommit, dense_mlpoly.rs:178-203
  FlVec blinds_vars = tape.random_vector("poly_blinds", Ls);
  pf.comm_vars = poly_commit_rows(vars, blinds_vars, gens.gens_n, threads);
  tr.append_message("poly_commitment", "poly_commitment_begin");  // dense_mlpoly.rs:289-297
  for (const CPt& c : pf.comm_vars) tr.append_point("poly_commitment_share", c.b);
  tr.append_message("poly_commitment", "poly_commitment_end");
  double t1 = now();
  // phase 1, r1csproof.rs:169-205
  FlVec z = vars;
  z.push_back(fl_one());
  z.insert(z.end(), inputs.begin(), inputs.end());
  z.resize(2 * nv, fl_zero());
  size_t nrx = ilog2(inst.num_cons), nry = ilog2(z.size());
  FlVec tau = tr.challenge_vector("challenge_tau", nrx);
  FlVec poly_tau = eq_evals(tau);
  FlVec Az = inst.multiply_vec(0, z), Bz = inst.multiply_vec(1, z), Cz = inst.multiply_vec(2, z);
  FlVec claims1;
  Fl blind_post1;
  pf.sc_proof_phase1 = zk_sumcheck_prove(fl_zero(), fl_zero(), nrx, {&poly_tau, &Az, &Bz, &Cz}, 3, gens.gens_1,
                                         gens.gens_4, tr, tape, rx, claims1, blind_post1);
  double t2 = now();
  Fl tau_claim = claims1[0], Az_claim = claims1[1], Bz_claim = claims1[2], Cz_claim = claims1[3];
  Fl Az_blind = tape.random_scalar("Az_blind"), Bz_blind = tape.random_scalar("Bz_blind"),
     Cz_blind = tape.random_scalar("Cz_blind"), prod_blind = tape.random_scalar("prod_Az_Bz_blind");
  CPt comm_Cz, comm_Az, comm_Bz, comm_prod;
  pf.pok_Cz = knowledge_prove(gens.gens_1, tr, tape, Cz_claim, Cz_blind, comm_Cz);
  Fl prod = fl_mul(Az_claim, Bz_claim);
  pf.proof_prod = product_prove(gens.gens_1, tr, tape, Az_claim, Az_blind, Bz_claim, Bz_blind, prod, prod_blind, comm_Az,
                                comm_Bz, comm_prod);
  tr.append_point("comm_Az_claim", comm_Az.b);
  tr.append_point("comm_Bz_claim", comm_Bz.b);
  tr.append_point("comm_Cz_claim", comm_Cz.b);
  tr.append_point("comm_prod_Az_Bz_claims", comm_prod.b);
  pf.claims_phase2[0] = comm_Az; pf.claims_phase2[1] = comm_Bz; pf.claims_phase2[2] = comm_Cz; pf.claims_phase2[3] = comm_prod;
  Fl blind_exp1 = fl_mul(tau_claim, fl_sub(prod_blind, Cz_blind));
  Fl claim_post1 = fl_mul(fl_sub(fl_mul(Az_claim, Bz_claim), Cz_claim), tau_claim);
  pf.proof_eq_sc_phase1 = equality_prove(gens.gens_1, tr, tape, claim_post1, blind_exp1, claim_post1, blind_post1);
  // phase 2, r1csproof.rs:260-292
  Fl r_A = tr.challenge_scalar("challenege_Az"), r_B = tr.challenge_scalar("challenege_Bz"),
     r_C = tr.challenge_scalar("challenege_Cz");
  Fl claim2 = fl_add(fl_add(fl_mul(r_A, Az_claim), fl_mul(r_B, Bz_claim)), fl_mul(r_C, Cz_claim));
  Fl blind2 = fl_add(fl_add(fl_mul(r_A, Az_blind), fl_mul(r_B, Bz_blind)), fl_mul(r_C, Cz_blind));
  FlVec evals_rx = eq_evals(rx);
  FlVec eA = inst.eval_table(0, evals_rx, z.size()), eB = inst.eval_table(1, evals_rx, z.size()),
        eC = inst.eval_table(2, evals_rx, z.size());
  FlVec evals_ABC(z.size());
  for (size_t i = 0; i < z.size(); i++)
    evals_ABC[i] = fl_add(fl_add(fl_mul(r_A, eA[i]), fl_mul(r_B, eB[i])), fl_mul(r_C, eC[i]));
  FlVec zz = z, claims2;
  Fl blind_post2;
  pf.sc_proof_phase2 = zk_sumcheck_prove(claim2, blind2, nry, {&zz, &evals_ABC}, 2, gens.gens_1, gens.gens_3, tr, tape,
                                         ry, claims2, blind_post2);
  double t3 = now();
  // polyeval, r1csproof.rs:294-307, dense_mlpoly.rs:233-239,309-362
  FlVec ry1(ry.begin() + 1, ry.end());
  FlVec chis = eq_evals(ry1);
  Fl eval_vars = dotproduct(vars.data(), chis.data(), nv);
  Fl blind_eval = tape.random_scalar("blind_eval");
  tr.append_protocol_name("polynomial evaluation proof");
  FlVec Lv = eq_evals(FlVec(ry1.begin(), ry1.begin() + lnv)), Rv = eq_evals(FlVec(ry1.begin() + lnv, ry1.end()));
  FlVec LZ = poly_bound(vars, Lv, Rs);
  Fl LZ_blind = dotproduct(blinds_vars.data(), Lv.data(), Ls);
  CPt C_LR;
  pf.proof_eval_vars_at_ry = dotproductlog_prove(gens.gens_n, gens.gens_1, tr, tape, LZ, LZ_blind, Rv, eval_vars,
                                                 blind_eval, C_LR, pf.comm_vars_at_ry);
  double t4 = now();
  Fl blind_eval_Z = fl_mul(fl_sub(fl_one(), ry[0]), blind_eval);
  Fl blind_exp2 = fl_mul(claims2[1], blind_eval_Z);
  Fl claim_post2 = fl_mul(claims2[0], claims2[1]);
  pf.proof_eq_sc_phase2 = equality_prove(gens.gens_1, tr, tape, claim_post2, blind_exp2, claim_post2, blind_post2);
  if (times) { times->polycommit = t1 - t0; times->phase1 = t2 - t1; times->phase2 = t3 - t2; times->polyeval = t4 - t3; times->total = now() - t0; }
  return pf;
}
// NIZK::prove lib.rs:493-535
static inline NIZK nizk_prove(const Instance& inst, const FlVec& vars_in, const FlVec& inputs, const NIZKGens& gens,
                              Transcript& tr, const Fl& tape_seed, int threads = 1, PhaseTimes* times = nullptr) {
  RandomTape tape("proof", tape_seed);
  tr.append_protocol_name("Spartan NIZK proof");
  FlVec vars = vars_in;
  if (inst.num_vars > vars.size()) vars.resize(inst.num_vars, fl_zero());  // Assignment::pad, lib.rs:92-105
  NIZK out;
  out.r1cs_sat_proof = r1cs_prove(inst, vars, inputs, gens, tr, tape, threads, out.rx, out.ry, times);
  return out;
}

// SparsePolynomial::evaluate, sparse_mlpoly.rs:1605-1628
static inline Fl sparse_poly_evaluate(size_t num_vars, const std::vector<std::pair<size_t, Fl>>& Z, const FlVec& r) {
  Fl tot = fl_zero();
  for (auto& e : Z) {
    Fl chi = fl_one();
    for (size_t j = 0; j < num_vars; j++) {
      bool bit = (e.first >> (num_vars - j - 1)) & 1;
      chi = fl_mul(chi, bit ? r[j] : fl_sub(fl_one(), r[j]));
    }
    tot = fl_add(tot, fl_mul(chi, e.second));
  }
  return tot;
}

// NIZK::verify lib.rs:538-574 + R1CSProof::verify r1csproof.rs:347-498.
// The reference's asserts/unwraps (panics) become `false` here.
// R1CSProof::verify r1csproof.rs:347-498: `ev` = the claimed (A, B, C)(rx, ry); returns the rx, ry of the two sum-checks
static inline bool r1cs_verify(const R1CSProof& pf, size_t num_vars, size_t num_cons, const FlVec& inputs, const Fl ev[3],
                               const NIZKGens& gens, Transcript& tr, FlVec& rx, FlVec& ry) {
  struct { size_t num_cons, num_vars; } inst{num_cons, num_vars};
  tr.append_protocol_name("R1CS proof");
  size_t n = inst.num_vars;
  tr.append_message("poly_commitment", "poly_commitment_begin");
  for (const CPt& c : pf.comm_vars) tr.append_point("poly_commitment_share", c.b);
  tr.append_message("poly_commitment", "poly_commitment_end");
  size_t nrx = ilog2(inst.num_cons), nry = ilog2(2 * n);
  FlVec tau = tr.challenge_vector("challenge_tau", nrx);
  CPt claim1 = compress(commit_scalar(fl_zero(), fl_zero(), gens.gens_1));
  CPt comm_post1, comm_post2;
  if (!zk_sumcheck_verify(pf.sc_proof_phase1, claim1, nrx, 3, gens.gens_1, gens.gens_4, tr, comm_post1, rx)) return false;
  const CPt &comm_Az = pf.claims_phase2[0], &comm_Bz = pf.claims_phase2[1], &comm_Cz = pf.claims_phase2[2],
            &comm_prod = pf.claims_phase2[3];
  if (!knowledge_verify(pf.pok_Cz, gens.gens_1, tr, comm_Cz)) return false;
  if (!product_verify(pf.proof_prod, gens.gens_1, tr, comm_Az, comm_Bz, comm_prod)) return false;
  tr.append_point("comm_Az_claim", comm_Az.b);
  tr.append_point("comm_Bz_claim", comm_Bz.b);
  tr.append_point("comm_Cz_claim", comm_Cz.b);
  tr.append_point("comm_prod_Az_Bz_claims", comm_prod.b);
  Fl taus = fl_one();
  for (size_t i = 0; i < rx.size(); i++)
    taus = fl_mul(taus, fl_add(fl_mul(rx[i], tau[i]), fl_mul(fl_sub(fl_one(), rx[i]), fl_sub(fl_one(), tau[i]))));
  CPt exp1 = compress(pt_mul(taus, pt_sub(unpack(comm_prod), unpack(comm_Cz))));
  if (!equality_verify(pf.proof_eq_sc_phase1, gens.gens_1, tr, exp1, comm_post1)) return false;
  Fl r_A = tr.challenge_scalar("challenege_Az"), r_B = tr.challenge_scalar("challenege_Bz"),
     r_C = tr.challenge_scalar("challenege_Cz");
  Fl rs[3] = {r_A, r_B, r_C};
  Pt cs[3] = {unpack(comm_Az), unpack(comm_Bz), unpack(comm_Cz)};
  CPt comm_claim2 = compress(pt_msm(rs, cs, 3));
  if (!zk_sumcheck_verify(pf.sc_proof_phase2, comm_claim2, nry, 2, gens.gens_1, gens.gens_3, tr, comm_post2, ry)) return false;
  // PolyEvalProof::verify, dense_mlpoly.rs:364-386
  tr.append_protocol_name("polynomial evaluation proof");
  FlVec ry1(ry.begin() + 1, ry.end());
  size_t lnv = ry1.size() / 2;
  FlVec Lv = eq_evals(FlVec(ry1.begin(), ry1.begin() + lnv)), Rv = eq_evals(FlVec(ry1.begin() + lnv, ry1.end()));
  if (pf.comm_vars.size() != Lv.size()) return false;
  std::vector<Pt> Cd;
  for (const CPt& c : pf.comm_vars) Cd.push_back(unpack(c));
  CPt C_LZ = compress(pt_msm(Lv.data(), Cd.data(), Lv.size()));
  if (gens.gens_n.n != Rv.size()) return false;
  if (!dotproductlog_verify(pf.proof_eval_vars_at_ry, Rv.size(), gens.gens_n, gens.gens_1, tr, Rv, C_LZ, pf.comm_vars_at_ry))
    return false;
  std::vector<std::pair<size_t, Fl>> entries;
  entries.push_back({0, fl_one()});
  for (size_t i = 0; i < inputs.size(); i++) entries.push_back({i + 1, inputs[i]});
  Fl poly_input_eval = sparse_poly_evaluate(ilog2(n), entries, ry1);
  Fl ws[2] = {fl_sub(fl_one(), ry[0]), ry[0]};
  Pt ps[2] = {unpack(pf.comm_vars_at_ry), commit_scalar(poly_input_eval, fl_zero(), gens.gens_1)};
  Pt comm_eval_Z = pt_msm(ws, ps, 2);
  Fl k = fl_add(fl_add(fl_mul(r_A, ev[0]), fl_mul(r_B, ev[1])), fl_mul(r_C, ev[2]));
  CPt exp2 = compress(pt_mul(k, comm_eval_Z));
  return equality_verify(pf.proof_eq_sc_phase2, gens.gens_1, tr, exp2, comm_post2);
}
static inline bool nizk_verify_inner(const NIZK& nz, const Instance& inst, const FlVec& inputs, const NIZKGens& gens,
                                     Transcript& tr) {  // lib.rs:538-574
  tr.append_protocol_name("Spartan NIZK proof");
  if (nz.rx.size() != ilog2(inst.num_cons) || nz.ry.size() != ilog2(2 * inst.num_vars)) return false;
  Fl ev[3];
  inst.evaluate(nz.rx, nz.ry, ev);
  if (inputs.size() != inst.num_inputs) return false;
  FlVec rx, ry;
  if (!r1cs_verify(nz.r1cs_sat_proof, inst.num_vars, inst.num_cons, inputs, ev, gens, tr, rx, ry)) return false;
  return rx == nz.rx && ry == nz.ry;
}
static inline bool nizk_verify(const NIZK& nz, const Instance& inst, const FlVec& inputs, const NIZKGens& gens,
                               Transcript& tr) {
  try {
    return nizk_verify_inner(nz, inst, inputs, gens, tr);
  } catch (const std::exception&) {
    return false;
  }
}

// ------------------------------------------------ SNARK::encode (computation commitment)
// The public, deterministic half of SNARK mode: lib.rs:321-333 -> R1CSInstance::commit
// r1csinstance.rs:285-297 -> SparseMatPolynomial::multi_commit sparse_mlpoly.rs:485-506.  No randomness, no transcript.
struct SparseMatPolyCommitment {  // sparse_mlpoly.rs:330-337
  size_t batch_size, num_ops, num_mem_cells;
  std::vector<CPt> comm_comb_ops, comm_comb_mem;
};
struct R1CSCommitment {  // r1csinstance.rs:46-51
  size_t num_cons, num_vars, num_inputs;
  SparseMatPolyCommitment comm;
};
// the dense representation the commitment is taken over (MultiSparseMatPolynomialAsDense, sparse_mlpoly.rs:286-293)
struct SparseAsDense {
  size_t N, num_mem_cells;
  FlVec comb_ops, comb_mem;
};
static inline size_t next_pow2(size_t n) {  // usize::next_power_of_two (0 -> 1)
  size_t p = 1;
  while (p < n) p <<= 1;
  return p;
}
// DensePolynomial::from_usize uses Scalar::from(u64) = [val,0,0,0] * R2 (scalar/ristretto255.rs:213-217): fl_from_u64, fl.hpp
// AddrTimestamps::new, sparse_mlpoly.rs:233-266: one audit counter per memory cell, carried across the three matrices
static inline void addr_timestamps(size_t num_cells, size_t num_ops, const std::vector<size_t> ops_addr[3],
                                   std::vector<size_t> read_ts[3], std::vector<size_t>& audit_ts) {
  audit_ts.assign(num_cells, 0);
  for (int m = 0; m < 3; m++) {
    if (ops_addr[m].size() != num_ops) throw std::runtime_error("assert_eq!(item.len(), num_ops)");
    read_ts[m].assign(num_ops, 0);
    for (size_t i = 0; i < num_ops; i++) {
      size_t addr = ops_addr[m][i];
      if (addr >= num_cells) throw std::runtime_error("assert!(addr < num_cells)");
      size_t r_ts = audit_ts[addr];
      read_ts[m][i] = r_ts;
      audit_ts[addr] = r_ts + 1;
    }
  }
}
// SparseMatPolynomial::multi_sparse_to_dense_rep, sparse_mlpoly.rs:371-422 (+ sparse_to_dense_vecs :353-369,
// DensePolynomial::merge dense_mlpoly.rs:256-269, from_usize :271-277)
static inline SparseAsDense sparse_to_dense(const Instance& inst) {
  SparseAsDense d;
  size_t N = 0;
  for (int m = 0; m < 3; m++) N = std::max(N, next_pow2(inst.M[m].size()));
  N = next_pow2(N);
  d.N = N;
  size_t nvx = ilog2(inst.num_cons), nvy = ilog2(2 * inst.num_vars);  // r1csinstance.rs:100-101
  d.num_mem_cells = size_t(1) << std::max(nvx, nvy);
  std::vector<size_t> rows[3], cols[3], rts[3], cts[3], raudit, caudit;
  FlVec val[3];
  for (int m = 0; m < 3; m++) {
    rows[m].assign(N, 0); cols[m].assign(N, 0); val[m].assign(N, fl_zero());
    for (size_t i = 0; i < inst.M[m].size(); i++) {
      rows[m][i] = inst.M[m][i].row; cols[m][i] = inst.M[m][i].col; val[m][i] = inst.M[m][i].val;
    }
  }
  addr_timestamps(d.num_mem_cells, N, rows, rts, raudit);
  addr_timestamps(d.num_mem_cells, N, cols, cts, caudit);
  // comb_ops = row.ops_addr ++ row.read_ts ++ col.ops_addr ++ col.read_ts ++ val, zero-padded to a power of two
  auto put = [&](const std::vector<size_t>& v) { for (size_t x : v) d.comb_ops.push_back(fl_from_u64(x)); };
  for (int m = 0; m < 3; m++) put(rows[m]);
  for (int m = 0; m < 3; m++) put(rts[m]);
  for (int m = 0; m < 3; m++) put(cols[m]);
  for (int m = 0; m < 3; m++) put(cts[m]);
  for (int m = 0; m < 3; m++) d.comb_ops.insert(d.comb_ops.end(), val[m].begin(), val[m].end());
  d.comb_ops.resize(next_pow2(d.comb_ops.size()), fl_zero());
  // comb_mem = row.audit_ts extended by col.audit_ts (DensePolynomial::extend, dense_mlpoly.rs:245-254)
  for (size_t x : raudit) d.comb_mem.push_back(fl_from_u64(x));
  for (size_t x : caudit) d.comb_mem.push_back(fl_from_u64(x));
  return d;
}
// SparseMatPolyCommitmentGens::new, sparse_mlpoly.rs:301-327: the three generator sets read the same SHAKE stream
// (same label), so each is a prefix of the longest; only gens_ops and gens_mem are used by encode.
struct R1CSCommitmentGens {  // r1csinstance.rs:25-44
  size_t num_vars_ops, num_vars_mem, num_vars_derefs;
  MultiCommitGens ops_n, mem_n, derefs_n;
  R1CSCommitmentGens(size_t num_cons, size_t num_vars, size_t num_inputs, size_t num_nz_entries,
                     const std::string& label = "gens_r1cs_eval") {
    if (!(num_inputs < num_vars)) throw std::runtime_error("assert!(num_inputs < num_vars)");
    size_t nvx = ilog2(num_cons), nvy = ilog2(2 * num_vars), batch = 3;
    num_vars_ops = ilog2(next_pow2(num_nz_entries)) + ilog2(next_pow2(batch * 5));
    num_vars_mem = std::max(nvx, nvy) + 1;
    num_vars_derefs = ilog2(next_pow2(num_nz_entries)) + ilog2(next_pow2(batch * 2));
    auto poly_gens = [&](size_t nv) {  // PolyCommitmentGens::new -> DotProductProofGens::new, dense_mlpoly.rs:28-35
      size_t n = size_t(1) << (nv - nv / 2);
      MultiCommitGens a, b;
      MultiCommitGens::make(n + 1, label).split_at(n, a, b);
      return a;
    };
    ops_n = poly_gens(num_vars_ops);
    mem_n = poly_gens(num_vars_mem);
    derefs_n = poly_gens(num_vars_derefs);
  }
};
// DensePolynomial::commit(gens, None), dense_mlpoly.rs:178-203: zero blinds; [Scalar]::commit asserts gens_n.n == R
static inline std::vector<CPt> poly_commit_unblinded(const FlVec& Z, const MultiCommitGens& gn, int threads) {
  size_t ell = ilog2(Z.size());
  if ((size_t(1) << ell) != Z.size()) throw std::runtime_error("assert_eq!(n, ell.pow2())");
  size_t L = size_t(1) << (ell / 2), R = Z.size() / L;
  if (gn.n != R) throw std::runtime_error("assert_eq!(gens_n.n, self.len())");
  return poly_commit_rows(Z, FlVec(L, fl_zero()), gn, threads);
}
// SNARKGens::new pads num_vars (lib.rs:283-291); R1CSInstance::commit, r1csinstance.rs:285-297
static inline R1CSCommitment snark_encode(const Instance& inst, const R1CSCommitmentGens& gens, int threads,
                                          SparseAsDense* dense_out = nullptr) {
  SparseAsDense d = sparse_to_dense(inst);
  R1CSCommitment c;
  c.num_cons = inst.num_cons; c.num_vars = inst.num_vars; c.num_inputs = inst.num_inputs;
  c.comm.batch_size = 3;
  c.comm.num_mem_cells = d.num_mem_cells;
  c.comm.num_ops = d.N;
  c.comm.comm_comb_ops = poly_commit_unblinded(d.comb_ops, gens.ops_n, threads);
  c.comm.comm_comb_mem = poly_commit_unblinded(d.comb_mem, gens.mem_n, threads);
  if (dense_out) *dense_out = std::move(d);
  return c;
}

// First step of SparseMatPolyEvalProof::prove (sparse_mlpoly.rs:1485-1503, timer "commit_nondet_witness"): the eq
// tables of the equalized point, the values every op reads from them (AddrTimestamps::deref :268-284,
// MultiSparseMatPolynomialAsDense::deref :508-514), merged (Derefs::new :47-62) and committed without blinds
// (Derefs::commit :64-67).  Deterministic given (rx, ry).
static inline std::vector<CPt> snark_commit_derefs(const Instance& inst, const R1CSCommitmentGens& gens, const FlVec& rx,
                                                   const FlVec& ry, int threads, FlVec* comb_out = nullptr) {
  FlVec rxe = rx, rye = ry;  // equalize, :1453-1469: the shorter point is extended with leading zeros
  if (rxe.size() < rye.size()) rxe.insert(rxe.begin(), rye.size() - rxe.size(), fl_zero());
  if (rye.size() < rxe.size()) rye.insert(rye.begin(), rxe.size() - rye.size(), fl_zero());
  FlVec mem_rx = eq_evals(rxe), mem_ry = eq_evals(rye);
  size_t N = 0;
  for (int m = 0; m < 3; m++) N = std::max(N, next_pow2(inst.M[m].size()));
  N = next_pow2(N);
  FlVec comb;
  for (int side = 0; side < 2; side++)
    for (int m = 0; m < 3; m++)
      for (size_t i = 0; i < N; i++) {
        size_t addr = i < inst.M[m].size() ? (side == 0 ? inst.M[m][i].row : inst.M[m][i].col) : 0;
        const FlVec& mem = side == 0 ? mem_rx : mem_ry;
        if (addr >= mem.size()) throw std::runtime_error("index out of bounds: mem_val[a]");
        comb.push_back(mem[addr]);
      }
  comb.resize(next_pow2(comb.size()), fl_zero());
  std::vector<CPt> C = poly_commit_unblinded(comb, gens.derefs_n, threads);
  if (comb_out) *comb_out = std::move(comb);
  return C;
}

// PolyEvalNetwork::new (sparse_mlpoly.rs:652-668, timer "build_layered_network"): Layers::new for the row and the
// column side -- build_hash_layer (:531-606) and one ProductCircuit (product_tree.rs:16-64) per hashed polynomial --
// reduced to what the proof then claims: ProductCircuit::evaluate of init, read[0..3), write[0..3), audit.
static inline Fl product_circuit_evaluate(const FlVec& poly) {  // ProductCircuit::new + evaluate, product_tree.rs:35-64
  if (poly.size() < 2) throw std::runtime_error("ProductCircuit::new of a polynomial with fewer than two entries");
  size_t num_layers = ilog2(poly.size());
  FlVec left(poly.begin(), poly.begin() + poly.size() / 2), right(poly.begin() + poly.size() / 2, poly.end());
  for (size_t i = 0; i + 1 < num_layers; i++) {  // compute_layer, :17-33
    size_t len = left.size() + right.size();
    FlVec ol, orr;
    for (size_t k = 0; k < len / 4; k++) ol.push_back(fl_mul(left[k], right[k]));
    for (size_t k = len / 4; k < len / 2; k++) orr.push_back(fl_mul(left[k], right[k]));
    left.swap(ol);
    right.swap(orr);
  }
  return fl_mul(left[0], right[0]);
}
// out[8] = init, read[0], read[1], read[2], write[0], write[1], write[2], audit
static inline void layers_products(const FlVec& eval_table, const std::vector<size_t> addrs[3], const FlVec derefs[3],
                                   const std::vector<size_t> read_ts[3], const std::vector<size_t>& audit_ts, const Fl& r_hash,
                                   const Fl& r_multiset_check, Fl out[8]) {
  const Fl r_hash_sqr = fl_mul(r_hash, r_hash);
  auto hash = [&](const Fl& addr, const Fl& val, const Fl& ts) {  // ts * r_hash_sqr + val * r_hash + addr, :548-550
    return fl_add(fl_add(fl_mul(ts, r_hash_sqr), fl_mul(val, r_hash)), addr);
  };
  size_t cells = eval_table.size();
  FlVec init(cells), audit(cells);
  for (size_t i = 0; i < cells; i++) {
    init[i] = fl_sub(hash(fl_from_u64(i), eval_table[i], fl_zero()), r_multiset_check);
    audit[i] = fl_sub(hash(fl_from_u64(i), eval_table[i], fl_from_u64(audit_ts[i])), r_multiset_check);
  }
  out[0] = product_circuit_evaluate(init);
  out[7] = product_circuit_evaluate(audit);
  for (int m = 0; m < 3; m++) {
    size_t n = addrs[m].size();
    FlVec rd(n), wr(n);
    for (size_t i = 0; i < n; i++) {
      Fl a = fl_from_u64(addrs[m][i]), ts = fl_from_u64(read_ts[m][i]);
      rd[i] = fl_sub(hash(a, derefs[m][i], ts), r_multiset_check);
      wr[i] = fl_sub(hash(a, derefs[m][i], fl_add(ts, fl_one())), r_multiset_check);  // write_ts = read_ts + 1, :592-597
    }
    out[1 + m] = product_circuit_evaluate(rd);
    out[4 + m] = product_circuit_evaluate(wr);
  }
}
// out[16]: the eight products of the row side, then of the column side
static inline void snark_network_products(const Instance& inst, const FlVec& rx, const FlVec& ry, const Fl& r_hash,
                                          const Fl& r_multiset_check, Fl out[16]) {
  FlVec rxe = rx, rye = ry;
  if (rxe.size() < rye.size()) rxe.insert(rxe.begin(), rye.size() - rxe.size(), fl_zero());
  if (rye.size() < rxe.size()) rye.insert(rye.begin(), rxe.size() - rye.size(), fl_zero());
  FlVec mem_rx = eq_evals(rxe), mem_ry = eq_evals(rye);
  size_t N = 0;
  for (int m = 0; m < 3; m++) N = std::max(N, next_pow2(inst.M[m].size()));
  N = next_pow2(N);
  size_t cells = size_t(1) << std::max(ilog2(inst.num_cons), ilog2(2 * inst.num_vars));
  if (mem_rx.size() != cells) throw std::runtime_error("assert_eq!(eval_table.len(), audit_ts.len())");
  for (int side = 0; side < 2; side++) {
    const FlVec& mem = side == 0 ? mem_rx : mem_ry;
    std::vector<size_t> addrs[3], rts[3], audit;
    FlVec derefs[3];
    for (int m = 0; m < 3; m++) {
      addrs[m].assign(N, 0);
      for (size_t i = 0; i < inst.M[m].size(); i++) addrs[m][i] = side == 0 ? inst.M[m][i].row : inst.M[m][i].col;
    }
    addr_timestamps(cells, N, addrs, rts, audit);
    for (int m = 0; m < 3; m++) {
      derefs[m].resize(N);
      for (size_t i = 0; i < N; i++) derefs[m][i] = mem[addrs[m][i]];
    }
    layers_products(mem, addrs, derefs, rts, audit, r_hash, r_multiset_check, out + 8 * side);
  }
}

// ------------------------------------------------ bincode 1.3
struct Writer {
  std::vector<uint8_t> b;
  void u64(uint64_t x) { for (int i = 0; i < 8; i++) b.push_back((uint8_t)(x >> (8 * i))); }
  void pt(const CPt& c) { b.insert(b.end(), c.b, c.b + 32); }
  void fl(const Fl& x) { const uint8_t* p = (const uint8_t*)x.v; b.insert(b.end(), p, p + 32); }  // raw Montgomery limbs (F6)
  void pts(const std::vector<CPt>& v) { u64(v.size()); for (auto& c : v) pt(c); }
  void fls(const FlVec& v) { u64(v.size()); for (auto& x : v) fl(x); }
  void sc(const ZKSumcheckProof& s) {
    pts(s.comm_polys); pts(s.comm_evals); u64(s.proofs.size());
    for (auto& p : s.proofs) { pt(p.delta); pt(p.beta); fls(p.z); fl(p.z_delta); fl(p.z_beta); }
  }
  void eq(const EqualityProof& e) { pt(e.alpha); fl(e.z); }
  void r1cs(const R1CSProof& p) {
    pts(p.comm_vars); sc(p.sc_proof_phase1);
    for (int i = 0; i < 4; i++) pt(p.claims_phase2[i]);
    pt(p.pok_Cz.alpha); fl(p.pok_Cz.z1); fl(p.pok_Cz.z2);
    pt(p.proof_prod.alpha); pt(p.proof_prod.beta); pt(p.proof_prod.delta);
    for (int i = 0; i < 5; i++) fl(p.proof_prod.z[i]);
    eq(p.proof_eq_sc_phase1); sc(p.sc_proof_phase2); pt(p.comm_vars_at_ry);
    const DotProductProofLog& d = p.proof_eval_vars_at_ry;
    pts(d.bullet.L_vec); pts(d.bullet.R_vec); pt(d.delta); pt(d.beta); fl(d.z1); fl(d.z2);
    eq(p.proof_eq_sc_phase2);
  }
  void nizk(const NIZK& n) { r1cs(n.r1cs_sat_proof); fls(n.rx); fls(n.ry); }
  // R1CSCommitment's own fields (it is not Serialize in the reference), then bincode of the Serialize
  // SparseMatPolyCommitment in declaration order (sparse_mlpoly.rs:329-337)
  void r1cs_commitment(const R1CSCommitment& c) {
    u64(c.num_cons); u64(c.num_vars); u64(c.num_inputs);
    u64(c.comm.batch_size); u64(c.comm.num_ops); u64(c.comm.num_mem_cells);
    pts(c.comm.comm_comb_ops); pts(c.comm.comm_comb_mem);
  }
};
struct Reader {
  const uint8_t* p;
  size_t n, pos;
  bool ok;
  Reader(const uint8_t* d, size_t len) : p(d), n(len), pos(0), ok(true) {}
  bool need(size_t k) { if (pos + k > n) { ok = false; return false; } return true; }
  uint64_t u64() { if (!need(8)) return 0; uint64_t x; memcpy(&x, p + pos, 8); pos += 8; return x; }
  CPt pt() { CPt c; memset(c.b, 0, 32); if (need(32)) { memcpy(c.b, p + pos, 32); pos += 32; } return c; }
  Fl fl() {
    Fl x = fl_zero();
    if (need(32)) {
      memcpy(x.v, p + pos, 32); pos += 32;
      // serde gives no range check; treat limbs >= l as malformed
      uint64_t bo = 0;
      (void)sbb(x.v[0], FL_MODULUS.v[0], bo); (void)sbb(x.v[1], FL_MODULUS.v[1], bo);
      (void)sbb(x.v[2], FL_MODULUS.v[2], bo); (void)sbb(x.v[3], FL_MODULUS.v[3], bo);
      if (!(bo & 1)) ok = false;
    }
    return x;
  }
  std::vector<CPt> pts() { uint64_t k = u64(); std::vector<CPt> v; if (k > n) { ok = false; return v; } for (uint64_t i = 0; i < k && ok; i++) v.push_back(pt()); return v; }
  FlVec fls() { uint64_t k = u64(); FlVec v; if (k > n) { ok = false; return v; } for (uint64_t i = 0; i < k && ok; i++) v.push_back(fl()); return v; }
  ZKSumcheckProof sc() {
    ZKSumcheckProof s;
    s.comm_polys = pts(); s.comm_evals = pts();
    uint64_t k = u64();
    if (k > n) { ok = false; return s; }
    for (uint64_t i = 0; i < k && ok; i++) { DotProductProof d; d.delta = pt(); d.beta = pt(); d.z = fls(); d.z_delta = fl(); d.z_beta = fl(); s.proofs.push_back(d); }
    return s;
  }
  EqualityProof eq() { EqualityProof e; e.alpha = pt(); e.z = fl(); return e; }
  R1CSProof r1cs() {
    R1CSProof q;
    q.comm_vars = pts(); q.sc_proof_phase1 = sc();
    for (int i = 0; i < 4; i++) q.claims_phase2[i] = pt();
    q.pok_Cz.alpha = pt(); q.pok_Cz.z1 = fl(); q.pok_Cz.z2 = fl();
    q.proof_prod.alpha = pt(); q.proof_prod.beta = pt(); q.proof_prod.delta = pt();
    for (int i = 0; i < 5; i++) q.proof_prod.z[i] = fl();
    q.proof_eq_sc_phase1 = eq(); q.sc_proof_phase2 = sc(); q.comm_vars_at_ry = pt();
    DotProductProofLog& d = q.proof_eval_vars_at_ry;
    d.bullet.L_vec = pts(); d.bullet.R_vec = pts(); d.delta = pt(); d.beta = pt(); d.z1 = fl(); d.z2 = fl();
    q.proof_eq_sc_phase2 = eq();
    return q;
  }
  NIZK nizk() {
    NIZK nz;
    nz.r1cs_sat_proof = r1cs();
    nz.rx = fls(); nz.ry = fls();
    if (pos != n) ok = false;
    return nz;
  }
  R1CSCommitment r1cs_commitment() {
    R1CSCommitment c;
    c.num_cons = u64(); c.num_vars = u64(); c.num_inputs = u64();
    c.comm.batch_size = u64(); c.comm.num_ops = u64(); c.comm.num_mem_cells = u64();
    c.comm.comm_comb_ops = pts(); c.comm.comm_comb_mem = pts();
    if (pos != n) ok = false;
    return c;
  }
};

}  // namespace orc
