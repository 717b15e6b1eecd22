// ORACLE (test infrastructure, parity unpinned against a cargo build like the rest): SNARK::prove / SNARK::verify,
// the second half of Spartan's SNARK mode -- the proof that the committed R1CS matrices evaluate to the claimed
// (A(rx,ry), B(rx,ry), C(rx,ry)).  CPU restatement, statement by statement, of
//   SNARK::prove / verify                       Spartan/src/lib.rs:335-457
//   R1CSEvalProof                               r1csinstance.rs:300-358
//   SparseMatPolyEvalProof::prove / verify      sparse_mlpoly.rs:1441-1582
//   PolyEvalNetworkProof                        sparse_mlpoly.rs:1332-1438
//   ProductLayerProof::prove / verify           sparse_mlpoly.rs:1040-1330
//   HashLayerProof::prove / verify              sparse_mlpoly.rs:671-1038
//   DerefsEvalProof                             sparse_mlpoly.rs:69-215
//   ProductCircuit, DotProductCircuit,
//   ProductCircuitEvalProofBatched              product_tree.rs:10-115, 177-486
//   SumcheckInstanceProof::verify,
//   prove_cubic_batched                         sumcheck.rs:17-61, 252-422
//   UniPoly::compress / decompress / append     unipoly.rs:82-120
//   PolyEvalProof::prove / verify / verify_plain dense_mlpoly.rs:304-401
//   IdentityPolynomial, bound_poly_var_bot      dense_mlpoly.rs:99-115, 223-230
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may use this file.
#pragma once
#include "spartan.hpp"

namespace orc {

// ------------------------------------------------ proof objects (serde field order = declaration order in the reference)
struct SumcheckProof { std::vector<FlVec> compressed_polys; };  // CompressedUniPoly.coeffs_except_linear_term, unipoly.rs:15-19
struct LayerProofBatched {                                      // product_tree.rs:147-152
  SumcheckProof proof;
  FlVec claims_prod_left, claims_prod_right;
};
struct ProductCircuitEvalProofBatched {  // product_tree.rs:177-181
  std::vector<LayerProofBatched> proof;
  FlVec claims_dotp[3];
};
struct PolyEvalProof { DotProductProofLog proof; };  // dense_mlpoly.rs:299-302
struct HashLayerProof {                              // sparse_mlpoly.rs:671-680
  FlVec eval_row_addr, eval_row_read_ts; Fl eval_row_audit_ts;
  FlVec eval_col_addr, eval_col_read_ts; Fl eval_col_audit_ts;
  FlVec eval_val;
  FlVec eval_derefs_row, eval_derefs_col;
  PolyEvalProof proof_ops, proof_mem, proof_derefs;
};
struct ProductLayerProof {  // sparse_mlpoly.rs:1040-1047
  Fl row_init; FlVec row_read, row_write; Fl row_audit;
  Fl col_init; FlVec col_read, col_write; Fl col_audit;
  FlVec eval_dotp_left, eval_dotp_right;
  ProductCircuitEvalProofBatched proof_mem, proof_ops;
};
struct SparseMatPolyEvalProof {  // sparse_mlpoly.rs:1441-1445 (+ PolyEvalNetworkProof :1332-1336)
  std::vector<CPt> comm_derefs;
  ProductLayerProof prod;
  HashLayerProof hash;
};
struct SNARK {  // lib.rs:307-312
  R1CSProof r1cs_sat_proof;
  Fl inst_evals[3];
  SparseMatPolyEvalProof r1cs_eval_proof;
};

// gens_ops / gens_mem / gens_derefs of SparseMatPolyCommitmentGens with their gens_1 halves (dense_mlpoly.rs:28-35,
// nizk/mod.rs:409-414): the R1CSCommitmentGens of spartan.hpp keeps only the gens_n halves, which is all encode needs
struct PolyGens { MultiCommitGens gens_n, gens_1; };
static inline PolyGens poly_gens(size_t num_vars, const std::string& label) {
  size_t n = size_t(1) << (num_vars - num_vars / 2);
  PolyGens g;
  MultiCommitGens::make(n + 1, label).split_at(n, g.gens_n, g.gens_1);
  return g;
}
struct SNARKGens {  // lib.rs:273-303
  NIZKGens sat;     // R1CSGens::new(b"gens_r1cs_sat", num_cons, num_vars_padded) has the same content
  R1CSCommitmentGens eval;
  PolyGens ops, mem, derefs;
  SNARKGens(size_t num_cons, size_t num_vars, size_t num_inputs, size_t num_nz_entries)
      : sat(num_cons, num_vars, num_inputs), eval(num_cons, pad_vars(num_vars, num_inputs), num_inputs, num_nz_entries) {
    ops = poly_gens(eval.num_vars_ops, "gens_r1cs_eval");
    mem = poly_gens(eval.num_vars_mem, "gens_r1cs_eval");
    derefs = poly_gens(eval.num_vars_derefs, "gens_r1cs_eval");
  }
  static size_t pad_vars(size_t num_vars, size_t num_inputs) { return next_pow2(std::max(num_vars, num_inputs + 1)); }
};

// ------------------------------------------------ small pieces
static inline void append_scalars(Transcript& tr, const char* label, const FlVec& v) {  // impl AppendToTranscript for Vec<Scalar>
  tr.append_message(label, "begin_append_vector");
  for (const Fl& x : v) tr.append_scalar(label, x);
  tr.append_message(label, "end_append_vector");
}
static inline void append_poly_commitment(Transcript& tr, const char* label, const std::vector<CPt>& C) {  // dense_mlpoly.rs:289-297
  tr.append_message(label, "poly_commitment_begin");
  for (const CPt& c : C) tr.append_point("poly_commitment_share", c.b);
  tr.append_message(label, "poly_commitment_end");
}
static inline void bound_poly_var_bot(FlVec& Z, const Fl& r) {  // dense_mlpoly.rs:223-230
  size_t n = Z.size() / 2;
  for (size_t i = 0; i < n; i++) Z[i] = fl_add(Z[2 * i], fl_mul(r, fl_sub(Z[2 * i + 1], Z[2 * i])));
  Z.resize(n);
}
static inline Fl poly_evaluate(const FlVec& Z, const FlVec& r) {  // DensePolynomial::evaluate, dense_mlpoly.rs:233-239
  FlVec chis = eq_evals(r);
  if (chis.size() != Z.size()) throw std::runtime_error("assert_eq!(chis.len(), self.Z.len())");
  return dotproduct(Z.data(), chis.data(), Z.size());
}
static inline FlVec compress_unipoly(const FlVec& coeffs) {  // unipoly.rs:82-88
  FlVec c{coeffs[0]};
  c.insert(c.end(), coeffs.begin() + 2, coeffs.end());
  return c;
}
static inline FlVec decompress_unipoly(const FlVec& c, const Fl& hint) {  // unipoly.rs:98-110
  Fl lin = fl_sub(fl_sub(hint, c[0]), c[0]);
  for (size_t i = 1; i < c.size(); i++) lin = fl_sub(lin, c[i]);
  FlVec coeffs{c[0], lin};
  coeffs.insert(coeffs.end(), c.begin() + 1, c.end());
  return coeffs;
}
static inline void append_unipoly(Transcript& tr, const char* label, const FlVec& coeffs) {  // unipoly.rs:112-120
  tr.append_message(label, "UniPoly_begin");
  for (const Fl& x : coeffs) tr.append_scalar("coeff", x);
  tr.append_message(label, "UniPoly_end");
}
// n-to-1 reduction of claimed evaluations (sparse_mlpoly.rs:99-114 and :768-785): returns the joint claim, r_joint = challenges ++ r
static inline Fl combine_claims(Transcript& tr, const char* challenge_label, FlVec evals, const FlVec& r, FlVec& r_joint) {
  FlVec ch = tr.challenge_vector(challenge_label, ilog2(evals.size()));
  for (size_t i = ch.size(); i-- > 0;) bound_poly_var_bot(evals, ch[i]);
  if (evals.size() != 1) throw std::runtime_error("assert_eq!(poly_evals.len(), 1)");
  r_joint = ch;
  r_joint.insert(r_joint.end(), r.begin(), r.end());
  return evals[0];
}

// ------------------------------------------------ PolyEvalProof, dense_mlpoly.rs:304-401 (blinds None, blind_Zr None)
static inline PolyEvalProof polyeval_prove_plain(const FlVec& poly, const FlVec& r, const Fl& Zr, const PolyGens& g, Transcript& tr,
                                                 RandomTape& tape) {
  tr.append_protocol_name("polynomial evaluation proof");
  if ((size_t(1) << r.size()) != poly.size()) throw std::runtime_error("assert_eq!(poly.get_num_vars(), r.len())");
  size_t lnv = r.size() / 2, Rs = size_t(1) << (r.size() - lnv);
  FlVec Lv = eq_evals(FlVec(r.begin(), r.begin() + lnv)), Rv = eq_evals(FlVec(r.begin() + lnv, r.end()));
  FlVec LZ = poly_bound(poly, Lv, Rs);
  PolyEvalProof p;
  CPt Cx, Cy;
  p.proof = dotproductlog_prove(g.gens_n, g.gens_1, tr, tape, LZ, fl_zero(), Rv, Zr, fl_zero(), Cx, Cy);
  return p;
}
static inline bool polyeval_verify_plain(const PolyEvalProof& p, const PolyGens& g, Transcript& tr, const FlVec& r, const Fl& Zr,
                                         const std::vector<CPt>& comm) {
  CPt C_Zr = compress(commit_scalar(Zr, fl_zero(), g.gens_1));
  tr.append_protocol_name("polynomial evaluation proof");
  size_t lnv = r.size() / 2;
  FlVec Lv = eq_evals(FlVec(r.begin(), r.begin() + lnv)), Rv = eq_evals(FlVec(r.begin() + lnv, r.end()));
  if (comm.size() != Lv.size()) return false;
  std::vector<Pt> Cd;
  for (const CPt& c : comm) Cd.push_back(unpack(c));
  CPt C_LZ = compress(pt_msm(Lv.data(), Cd.data(), Lv.size()));
  if (g.gens_n.n != Rv.size()) return false;
  return dotproductlog_verify(p.proof, Rv.size(), g.gens_n, g.gens_1, tr, Rv, C_LZ, C_Zr);
}

// ------------------------------------------------ product circuits, product_tree.rs:10-115
struct ProductCircuit {
  std::vector<FlVec> left, right;
  explicit ProductCircuit(const FlVec& poly) {
    if (poly.size() < 2) throw std::runtime_error("ProductCircuit::new of a polynomial with fewer than two entries");
    size_t num_layers = ilog2(poly.size());
    left.push_back(FlVec(poly.begin(), poly.begin() + poly.size() / 2));
    right.push_back(FlVec(poly.begin() + poly.size() / 2, poly.end()));
    for (size_t i = 0; i + 1 < num_layers; i++) {  // compute_layer, :17-33
      const FlVec &l = left[i], &r = right[i];
      size_t len = l.size() + r.size();
      FlVec ol, orr;
      for (size_t k = 0; k < len / 4; k++) ol.push_back(fl_mul(l[k], r[k]));
      for (size_t k = len / 4; k < len / 2; k++) orr.push_back(fl_mul(l[k], r[k]));
      left.push_back(ol);
      right.push_back(orr);
    }
  }
  Fl evaluate() const { return fl_mul(left.back()[0], right.back()[0]); }
};
struct DotProductCircuit {  // product_tree.rs:66-115
  FlVec left, right, weight;
  Fl evaluate() const {
    Fl s = fl_zero();
    for (size_t i = 0; i < left.size(); i++) s = fl_add(s, fl_mul(fl_mul(left[i], right[i]), weight[i]));
    return s;
  }
};

// SumcheckInstanceProof::prove_cubic_batched, sumcheck.rs:252-422, comb_func = A * B * C (product_tree.rs:276-279)
static inline SumcheckProof prove_cubic_batched(const Fl& claim, size_t num_rounds, std::vector<FlVec*>& A_par, std::vector<FlVec*>& B_par,
                                                FlVec& C_par, std::vector<FlVec*>& A_seq, std::vector<FlVec*>& B_seq,
                                                std::vector<FlVec*>& C_seq, const FlVec& coeffs, Transcript& tr, FlVec& r,
                                                FlVec claims_prod[2], FlVec claims_dotp[3]) {
  Fl e = claim;
  SumcheckProof pf;
  r.clear();
  auto evals3 = [](const FlVec& A, const FlVec& B, const FlVec& C, Fl out[3]) {
    out[0] = out[1] = out[2] = fl_zero();
    size_t len = A.size() / 2;
    for (size_t i = 0; i < len; i++) {
      out[0] = fl_add(out[0], fl_mul(fl_mul(A[i], B[i]), C[i]));
      Fl a = fl_sub(fl_add(A[len + i], A[len + i]), A[i]), b = fl_sub(fl_add(B[len + i], B[len + i]), B[i]),
         c = fl_sub(fl_add(C[len + i], C[len + i]), C[i]);
      out[1] = fl_add(out[1], fl_mul(fl_mul(a, b), c));
      a = fl_sub(fl_add(a, A[len + i]), A[i]); b = fl_sub(fl_add(b, B[len + i]), B[i]); c = fl_sub(fl_add(c, C[len + i]), C[i]);
      out[2] = fl_add(out[2], fl_mul(fl_mul(a, b), c));
    }
  };
  for (size_t j = 0; j < num_rounds; j++) {
    Fl c0 = fl_zero(), c2 = fl_zero(), c3 = fl_zero();
    size_t k = 0;
    for (size_t i = 0; i < A_par.size(); i++, k++) {
      Fl ev[3];
      evals3(*A_par[i], *B_par[i], C_par, ev);
      c0 = fl_add(c0, fl_mul(ev[0], coeffs[k])); c2 = fl_add(c2, fl_mul(ev[1], coeffs[k])); c3 = fl_add(c3, fl_mul(ev[2], coeffs[k]));
    }
    for (size_t i = 0; i < A_seq.size(); i++, k++) {
      Fl ev[3];
      evals3(*A_seq[i], *B_seq[i], *C_seq[i], ev);
      c0 = fl_add(c0, fl_mul(ev[0], coeffs[k])); c2 = fl_add(c2, fl_mul(ev[1], coeffs[k])); c3 = fl_add(c3, fl_mul(ev[2], coeffs[k]));
    }
    FlVec poly = unipoly_from_evals({c0, fl_sub(e, c0), c2, c3});
    append_unipoly(tr, "poly", poly);
    Fl r_j = tr.challenge_scalar("challenge_nextround");
    r.push_back(r_j);
    for (size_t i = 0; i < A_par.size(); i++) { bound_poly_var_top(*A_par[i], r_j); bound_poly_var_top(*B_par[i], r_j); }
    bound_poly_var_top(C_par, r_j);
    for (size_t i = 0; i < A_seq.size(); i++) {
      bound_poly_var_top(*A_seq[i], r_j); bound_poly_var_top(*B_seq[i], r_j); bound_poly_var_top(*C_seq[i], r_j);
    }
    e = unipoly_eval(poly, r_j);
    pf.compressed_polys.push_back(compress_unipoly(poly));
  }
  claims_prod[0].clear(); claims_prod[1].clear();
  for (size_t i = 0; i < A_par.size(); i++) { claims_prod[0].push_back((*A_par[i])[0]); claims_prod[1].push_back((*B_par[i])[0]); }
  for (int t = 0; t < 3; t++) claims_dotp[t].clear();
  for (size_t i = 0; i < A_seq.size(); i++) {
    claims_dotp[0].push_back((*A_seq[i])[0]); claims_dotp[1].push_back((*B_seq[i])[0]); claims_dotp[2].push_back((*C_seq[i])[0]);
  }
  return pf;
}
// SumcheckInstanceProof::verify, sumcheck.rs:22-61
static inline bool sumcheck_verify(const SumcheckProof& pf, const Fl& claim, size_t num_rounds, size_t degree, Transcript& tr, Fl& e_out,
                                   FlVec& r) {
  Fl e = claim;
  r.clear();
  if (pf.compressed_polys.size() != num_rounds) return false;
  for (const FlVec& cp : pf.compressed_polys) {
    if (cp.size() != degree) return false;  // poly.degree() == degree_bound
    FlVec poly = decompress_unipoly(cp, e);
    Fl at1 = fl_zero();
    for (const Fl& x : poly) at1 = fl_add(at1, x);
    if (fl_add(poly[0], at1) != e) return false;
    append_unipoly(tr, "poly", poly);
    Fl r_i = tr.challenge_scalar("challenge_nextround");
    r.push_back(r_i);
    e = unipoly_eval(poly, r_i);
  }
  e_out = e;
  return true;
}

// ProductCircuitEvalProofBatched::prove, product_tree.rs:257-382
static inline ProductCircuitEvalProofBatched product_batched_prove(std::vector<ProductCircuit*>& prods, std::vector<DotProductCircuit*>& dotps,
                                                                   Transcript& tr, FlVec& rand) {
  ProductCircuitEvalProofBatched out;
  size_t num_layers = prods[0]->left.size();
  FlVec claims_to_verify;
  for (ProductCircuit* c : prods) claims_to_verify.push_back(c->evaluate());
  rand.clear();
  for (size_t layer_id = num_layers; layer_id-- > 0;) {
    size_t len = prods[0]->left[layer_id].size() + prods[0]->right[layer_id].size();
    FlVec C_par = eq_evals(rand);
    if (C_par.size() != len / 2) throw std::runtime_error("assert_eq!(poly_C_par.len(), len / 2)");
    size_t num_rounds = ilog2(C_par.size());
    std::vector<FlVec*> A_par, B_par, A_seq, B_seq, C_seq;
    for (ProductCircuit* c : prods) { A_par.push_back(&c->left[layer_id]); B_par.push_back(&c->right[layer_id]); }
    if (layer_id == 0 && !dotps.empty()) {
      for (DotProductCircuit* d : dotps) {
        claims_to_verify.push_back(d->evaluate());
        if (d->left.size() != len / 2) throw std::runtime_error("assert_eq!(len / 2, item.left.len())");
      }
      for (DotProductCircuit* d : dotps) { A_seq.push_back(&d->left); B_seq.push_back(&d->right); C_seq.push_back(&d->weight); }
    }
    FlVec coeff_vec = tr.challenge_vector("rand_coeffs_next_layer", claims_to_verify.size());
    Fl claim = fl_zero();
    for (size_t i = 0; i < claims_to_verify.size(); i++) claim = fl_add(claim, fl_mul(claims_to_verify[i], coeff_vec[i]));
    LayerProofBatched lp;
    FlVec rand_prod, claims_prod[2], claims_dotp[3];
    lp.proof = prove_cubic_batched(claim, num_rounds, A_par, B_par, C_par, A_seq, B_seq, C_seq, coeff_vec, tr, rand_prod, claims_prod,
                                   claims_dotp);
    for (size_t i = 0; i < prods.size(); i++) {
      tr.append_scalar("claim_prod_left", claims_prod[0][i]);
      tr.append_scalar("claim_prod_right", claims_prod[1][i]);
    }
    if (layer_id == 0 && !dotps.empty()) {
      for (size_t i = 0; i < dotps.size(); i++) {
        tr.append_scalar("claim_dotp_left", claims_dotp[0][i]);
        tr.append_scalar("claim_dotp_right", claims_dotp[1][i]);
        tr.append_scalar("claim_dotp_weight", claims_dotp[2][i]);
      }
      for (int t = 0; t < 3; t++) out.claims_dotp[t] = claims_dotp[t];
    }
    Fl r_layer = tr.challenge_scalar("challenge_r_layer");
    claims_to_verify.clear();
    for (size_t i = 0; i < prods.size(); i++)
      claims_to_verify.push_back(fl_add(claims_prod[0][i], fl_mul(r_layer, fl_sub(claims_prod[1][i], claims_prod[0][i]))));
    FlVec ext{r_layer};
    ext.insert(ext.end(), rand_prod.begin(), rand_prod.end());
    rand = ext;
    lp.claims_prod_left = claims_prod[0];
    lp.claims_prod_right = claims_prod[1];
    out.proof.push_back(lp);
  }
  return out;
}
// ProductCircuitEvalProofBatched::verify, product_tree.rs:384-485
static inline bool product_batched_verify(const ProductCircuitEvalProofBatched& pf, const FlVec& claims_prod_vec, const FlVec& claims_dotp_vec,
                                          size_t len, Transcript& tr, FlVec& claims_out, FlVec& claims_dotp_out, FlVec& rand) {
  size_t num_layers = ilog2(len);
  rand.clear();
  if (pf.proof.size() != num_layers) return false;
  FlVec claims_to_verify = claims_prod_vec;
  claims_dotp_out.clear();
  for (size_t i = 0; i < num_layers; i++) {
    const size_t num_rounds = i;
    if (i == num_layers - 1) claims_to_verify.insert(claims_to_verify.end(), claims_dotp_vec.begin(), claims_dotp_vec.end());
    FlVec coeff_vec = tr.challenge_vector("rand_coeffs_next_layer", claims_to_verify.size());
    Fl claim = fl_zero();
    for (size_t k = 0; k < claims_to_verify.size(); k++) claim = fl_add(claim, fl_mul(claims_to_verify[k], coeff_vec[k]));
    Fl claim_last;
    FlVec rand_prod;
    if (!sumcheck_verify(pf.proof[i].proof, claim, num_rounds, 3, tr, claim_last, rand_prod)) return false;
    const FlVec &cl = pf.proof[i].claims_prod_left, &cr = pf.proof[i].claims_prod_right;
    if (cl.size() != claims_prod_vec.size() || cr.size() != claims_prod_vec.size()) return false;
    for (size_t k = 0; k < claims_prod_vec.size(); k++) {
      tr.append_scalar("claim_prod_left", cl[k]);
      tr.append_scalar("claim_prod_right", cr[k]);
    }
    if (rand.size() != rand_prod.size()) return false;
    Fl eq = fl_one();
    for (size_t k = 0; k < rand.size(); k++)
      eq = fl_mul(eq, fl_add(fl_mul(rand[k], rand_prod[k]), fl_mul(fl_sub(fl_one(), rand[k]), fl_sub(fl_one(), rand_prod[k]))));
    Fl claim_expected = fl_zero();
    for (size_t k = 0; k < claims_prod_vec.size(); k++)
      claim_expected = fl_add(claim_expected, fl_mul(coeff_vec[k], fl_mul(fl_mul(cl[k], cr[k]), eq)));
    if (i == num_layers - 1) {
      const size_t np = claims_prod_vec.size();
      if (pf.claims_dotp[1].size() != pf.claims_dotp[0].size() || pf.claims_dotp[2].size() != pf.claims_dotp[0].size()) return false;
      if (np + pf.claims_dotp[0].size() > coeff_vec.size()) return false;
      for (size_t k = 0; k < pf.claims_dotp[0].size(); k++) {
        tr.append_scalar("claim_dotp_left", pf.claims_dotp[0][k]);
        tr.append_scalar("claim_dotp_right", pf.claims_dotp[1][k]);
        tr.append_scalar("claim_dotp_weight", pf.claims_dotp[2][k]);
        claim_expected = fl_add(claim_expected,
                                fl_mul(fl_mul(fl_mul(coeff_vec[k + np], pf.claims_dotp[0][k]), pf.claims_dotp[1][k]), pf.claims_dotp[2][k]));
      }
    }
    if (claim_expected != claim_last) return false;
    Fl r_layer = tr.challenge_scalar("challenge_r_layer");
    claims_to_verify.clear();
    for (size_t k = 0; k < cl.size(); k++) claims_to_verify.push_back(fl_add(cl[k], fl_mul(r_layer, fl_sub(cr[k], cl[k]))));
    if (i == num_layers - 1) {
      if (2 * (claims_dotp_vec.size() / 2) > pf.claims_dotp[0].size()) return false;
      for (size_t k = 0; k < claims_dotp_vec.size() / 2; k++)
        for (int t = 0; t < 3; t++) {
          const FlVec& c = pf.claims_dotp[t];
          claims_dotp_out.push_back(fl_add(c[2 * k], fl_mul(r_layer, fl_sub(c[2 * k + 1], c[2 * k]))));
        }
    }
    FlVec ext{r_layer};
    ext.insert(ext.end(), rand_prod.begin(), rand_prod.end());
    rand = ext;
  }
  claims_out = claims_to_verify;
  return true;
}

// ------------------------------------------------ the dense representation with everything the eval proof touches
struct DenseFull {  // MultiSparseMatPolynomialAsDense, sparse_mlpoly.rs:286-293
  size_t N, cells;
  std::vector<size_t> row_addr[3], col_addr[3], row_ts[3], col_ts[3], row_audit, col_audit;
  FlVec val[3], comb_ops, comb_mem;
};
static inline DenseFull dense_full(const Instance& inst) {
  DenseFull d;
  size_t N = 0;
  for (int m = 0; m < 3; m++) N = std::max(N, next_pow2(inst.M[m].size()));
  d.N = N = next_pow2(N);
  d.cells = size_t(1) << std::max(ilog2(inst.num_cons), ilog2(2 * inst.num_vars));
  for (int m = 0; m < 3; m++) {
    d.row_addr[m].assign(N, 0); d.col_addr[m].assign(N, 0); d.val[m].assign(N, fl_zero());
    for (size_t i = 0; i < inst.M[m].size(); i++) {
      d.row_addr[m][i] = inst.M[m][i].row; d.col_addr[m][i] = inst.M[m][i].col; d.val[m][i] = inst.M[m][i].val;
    }
  }
  addr_timestamps(d.cells, N, d.row_addr, d.row_ts, d.row_audit);
  addr_timestamps(d.cells, N, d.col_addr, d.col_ts, d.col_audit);
  SparseAsDense s = sparse_to_dense(inst);
  d.comb_ops = s.comb_ops;
  d.comb_mem = s.comb_mem;
  return d;
}
static inline FlVec from_usizes(const std::vector<size_t>& v) {
  FlVec o;
  for (size_t x : v) o.push_back(fl_from_u64(x));
  return o;
}
// Layers::build_hash_layer + Layers::new, sparse_mlpoly.rs:531-649: circuits init, read[3], write[3], audit
struct Layers { std::vector<ProductCircuit> c; };
static inline Layers layers_new(const FlVec& eval_table, const std::vector<size_t> addrs[3], const FlVec derefs[3],
                                const std::vector<size_t> read_ts[3], const std::vector<size_t>& audit_ts, const Fl& r_hash, const Fl& r_ms) {
  const Fl r_hash_sqr = fl_mul(r_hash, r_hash);
  auto hash = [&](const Fl& addr, const Fl& val, const Fl& ts) { return fl_add(fl_add(fl_mul(ts, r_hash_sqr), fl_mul(val, r_hash)), addr); };
  size_t cells = eval_table.size();
  FlVec init(cells), audit(cells);
  for (size_t i = 0; i < cells; i++) {
    init[i] = fl_sub(hash(fl_from_u64(i), eval_table[i], fl_zero()), r_ms);
    audit[i] = fl_sub(hash(fl_from_u64(i), eval_table[i], fl_from_u64(audit_ts[i])), r_ms);
  }
  Layers L;
  L.c.emplace_back(init);
  std::vector<FlVec> rd(3), wr(3);
  for (int m = 0; m < 3; m++) {
    size_t n = addrs[m].size();
    rd[m].resize(n); wr[m].resize(n);
    for (size_t i = 0; i < n; i++) {
      Fl a = fl_from_u64(addrs[m][i]), ts = fl_from_u64(read_ts[m][i]);
      rd[m][i] = fl_sub(hash(a, derefs[m][i], ts), r_ms);
      wr[m][i] = fl_sub(hash(a, derefs[m][i], fl_add(ts, fl_one())), r_ms);
    }
  }
  for (int m = 0; m < 3; m++) L.c.emplace_back(rd[m]);
  for (int m = 0; m < 3; m++) L.c.emplace_back(wr[m]);
  L.c.emplace_back(audit);
  return L;
}

// ------------------------------------------------ SparseMatPolyEvalProof::prove, sparse_mlpoly.rs:1471-1538
static inline void equalize(const FlVec& rx, const FlVec& ry, FlVec& rxe, FlVec& rye) {  // :1453-1469
  rxe = rx; rye = ry;
  if (rxe.size() < rye.size()) rxe.insert(rxe.begin(), rye.size() - rxe.size(), fl_zero());
  if (rye.size() < rxe.size()) rye.insert(rye.begin(), rxe.size() - rye.size(), fl_zero());
}
static inline SparseMatPolyEvalProof sparse_eval_prove(const Instance& inst, const FlVec& rx, const FlVec& ry, const Fl evals[3],
                                                       const SNARKGens& gens, Transcript& tr, RandomTape& tape, int threads) {
  SparseMatPolyEvalProof out;
  tr.append_protocol_name("Sparse polynomial evaluation proof");
  DenseFull d = dense_full(inst);
  const size_t N = d.N;
  FlVec rxe, rye;
  equalize(rx, ry, rxe, rye);
  FlVec mem_rx = eq_evals(rxe), mem_ry = eq_evals(rye);
  FlVec row_ops_val[3], col_ops_val[3];  // Derefs, :268-284, 508-514
  for (int m = 0; m < 3; m++) {
    row_ops_val[m].resize(N); col_ops_val[m].resize(N);
    for (size_t i = 0; i < N; i++) {
      if (d.row_addr[m][i] >= mem_rx.size() || d.col_addr[m][i] >= mem_ry.size()) throw std::runtime_error("index out of bounds: mem_val[a]");
      row_ops_val[m][i] = mem_rx[d.row_addr[m][i]];
      col_ops_val[m][i] = mem_ry[d.col_addr[m][i]];
    }
  }
  FlVec derefs_comb;  // Derefs::new -> DensePolynomial::merge(row ++ col), :47-62
  for (int m = 0; m < 3; m++) derefs_comb.insert(derefs_comb.end(), row_ops_val[m].begin(), row_ops_val[m].end());
  for (int m = 0; m < 3; m++) derefs_comb.insert(derefs_comb.end(), col_ops_val[m].begin(), col_ops_val[m].end());
  derefs_comb.resize(next_pow2(derefs_comb.size()), fl_zero());
  out.comm_derefs = poly_commit_unblinded(derefs_comb, gens.derefs.gens_n, threads);  // Derefs::commit, :64-67
  tr.append_message("derefs_commitment", "begin_derefs_commitment");                  // :217-223
  append_poly_commitment(tr, "comm_poly_row_col_ops_val", out.comm_derefs);
  tr.append_message("derefs_commitment", "end_derefs_commitment");
  FlVec r_mem_check = tr.challenge_vector("challenge_r_hash", 2);
  if (mem_rx.size() != d.cells) throw std::runtime_error("assert_eq!(eval_table.len(), audit_ts.len())");
  Layers row = layers_new(mem_rx, d.row_addr, row_ops_val, d.row_ts, d.row_audit, r_mem_check[0], r_mem_check[1]);
  Layers col = layers_new(mem_ry, d.col_addr, col_ops_val, d.col_ts, d.col_audit, r_mem_check[0], r_mem_check[1]);

  tr.append_protocol_name("Sparse polynomial evaluation proof");  // PolyEvalNetworkProof::protocol_name, :1339-1341
  // ---- ProductLayerProof::prove, :1052-1230
  ProductLayerProof& pl = out.prod;
  tr.append_protocol_name("Sparse polynomial product layer proof");
  auto side_claims = [&](Layers& L, Fl& init, FlVec& read, FlVec& write, Fl& audit, const char* names[4]) {
    init = L.c[0].evaluate();
    audit = L.c[7].evaluate();
    read.clear(); write.clear();
    for (int m = 0; m < 3; m++) { read.push_back(L.c[1 + m].evaluate()); write.push_back(L.c[4 + m].evaluate()); }
    Fl ws = fl_one(), rs = fl_one();
    for (int m = 0; m < 3; m++) { ws = fl_mul(ws, write[m]); rs = fl_mul(rs, read[m]); }
    if (fl_mul(init, ws) != fl_mul(rs, audit)) throw std::runtime_error("assert_eq!(eval_init * ws, rs * eval_audit)");
    tr.append_scalar(names[0], init);
    append_scalars(tr, names[1], read);
    append_scalars(tr, names[2], write);
    tr.append_scalar(names[3], audit);
  };
  const char* rn[4] = {"claim_row_eval_init", "claim_row_eval_read", "claim_row_eval_write", "claim_row_eval_audit"};
  const char* cn[4] = {"claim_col_eval_init", "claim_col_eval_read", "claim_col_eval_write", "claim_col_eval_audit"};
  side_claims(row, pl.row_init, pl.row_read, pl.row_write, pl.row_audit, rn);
  side_claims(col, pl.col_init, pl.col_read, pl.col_write, pl.col_audit, cn);
  std::vector<DotProductCircuit> dl(3), dr(3);
  for (int m = 0; m < 3; m++) {  // DotProductCircuit::new(left = row_ops_val, right = col_ops_val, weight = val) and split, :1112-1135
    size_t h = N / 2;
    dl[m].left.assign(row_ops_val[m].begin(), row_ops_val[m].begin() + h); dr[m].left.assign(row_ops_val[m].begin() + h, row_ops_val[m].end());
    dl[m].right.assign(col_ops_val[m].begin(), col_ops_val[m].begin() + h); dr[m].right.assign(col_ops_val[m].begin() + h, col_ops_val[m].end());
    dl[m].weight.assign(d.val[m].begin(), d.val[m].begin() + h); dr[m].weight.assign(d.val[m].begin() + h, d.val[m].end());
    Fl el = dl[m].evaluate(), er = dr[m].evaluate();
    tr.append_scalar("claim_eval_dotp_left", el);
    tr.append_scalar("claim_eval_dotp_right", er);
    if (fl_add(el, er) != evals[m]) throw std::runtime_error("assert_eq!(eval_dotp_left + eval_dotp_right, eval[i])");
    pl.eval_dotp_left.push_back(el);
    pl.eval_dotp_right.push_back(er);
  }
  FlVec rand_ops, rand_mem;
  {
    std::vector<ProductCircuit*> pc;
    for (int k = 1; k <= 6; k++) pc.push_back(&row.c[k]);  // row read A,B,C, row write A,B,C
    for (int k = 1; k <= 6; k++) pc.push_back(&col.c[k]);
    std::vector<DotProductCircuit*> dc;
    for (int m = 0; m < 3; m++) { dc.push_back(&dl[m]); dc.push_back(&dr[m]); }
    pl.proof_ops = product_batched_prove(pc, dc, tr, rand_ops);
  }
  {
    std::vector<ProductCircuit*> pc{&row.c[0], &row.c[7], &col.c[0], &col.c[7]};
    std::vector<DotProductCircuit*> dc;
    pl.proof_mem = product_batched_prove(pc, dc, tr, rand_mem);
  }
  // ---- HashLayerProof::prove, :723-837
  HashLayerProof& hl = out.hash;
  tr.append_protocol_name("Sparse polynomial hash layer proof");
  for (int m = 0; m < 3; m++) hl.eval_derefs_row.push_back(poly_evaluate(row_ops_val[m], rand_ops));
  for (int m = 0; m < 3; m++) hl.eval_derefs_col.push_back(poly_evaluate(col_ops_val[m], rand_ops));
  {  // DerefsEvalProof::prove, :129-152, prove_single :81-127
    tr.append_protocol_name("Derefs evaluation proof");
    FlVec ev = hl.eval_derefs_row;
    ev.insert(ev.end(), hl.eval_derefs_col.begin(), hl.eval_derefs_col.end());
    ev.resize(next_pow2(ev.size()), fl_zero());
    append_scalars(tr, "evals_ops_val", ev);
    FlVec r_joint;
    Fl joint = combine_claims(tr, "challenge_combine_n_to_one", ev, rand_ops, r_joint);
    tr.append_scalar("joint_claim_eval", joint);
    hl.proof_derefs = polyeval_prove_plain(derefs_comb, r_joint, joint, gens.derefs, tr, tape);
  }
  for (int m = 0; m < 3; m++) hl.eval_row_addr.push_back(poly_evaluate(from_usizes(d.row_addr[m]), rand_ops));
  for (int m = 0; m < 3; m++) hl.eval_row_read_ts.push_back(poly_evaluate(from_usizes(d.row_ts[m]), rand_ops));
  hl.eval_row_audit_ts = poly_evaluate(from_usizes(d.row_audit), rand_mem);
  for (int m = 0; m < 3; m++) hl.eval_col_addr.push_back(poly_evaluate(from_usizes(d.col_addr[m]), rand_ops));
  for (int m = 0; m < 3; m++) hl.eval_col_read_ts.push_back(poly_evaluate(from_usizes(d.col_ts[m]), rand_ops));
  hl.eval_col_audit_ts = poly_evaluate(from_usizes(d.col_audit), rand_mem);
  for (int m = 0; m < 3; m++) hl.eval_val.push_back(poly_evaluate(d.val[m], rand_ops));
  {
    FlVec ev = hl.eval_row_addr;
    ev.insert(ev.end(), hl.eval_row_read_ts.begin(), hl.eval_row_read_ts.end());
    ev.insert(ev.end(), hl.eval_col_addr.begin(), hl.eval_col_addr.end());
    ev.insert(ev.end(), hl.eval_col_read_ts.begin(), hl.eval_col_read_ts.end());
    ev.insert(ev.end(), hl.eval_val.begin(), hl.eval_val.end());
    ev.resize(next_pow2(ev.size()), fl_zero());
    append_scalars(tr, "claim_evals_ops", ev);
    FlVec r_joint;
    Fl joint = combine_claims(tr, "challenge_combine_n_to_one", ev, rand_ops, r_joint);
    tr.append_scalar("joint_claim_eval_ops", joint);
    hl.proof_ops = polyeval_prove_plain(d.comb_ops, r_joint, joint, gens.ops, tr, tape);
  }
  {
    FlVec ev{hl.eval_row_audit_ts, hl.eval_col_audit_ts};
    append_scalars(tr, "claim_evals_mem", ev);
    FlVec r_joint;
    Fl joint = combine_claims(tr, "challenge_combine_two_to_one", ev, rand_mem, r_joint);
    tr.append_scalar("joint_claim_eval_mem", joint);
    hl.proof_mem = polyeval_prove_plain(d.comb_mem, r_joint, joint, gens.mem, tr, tape);
  }
  return out;
}

// HashLayerProof::verify_helper, sparse_mlpoly.rs:839-894
static inline bool hash_verify_helper(const FlVec& rand_mem, const Fl& claim_init, const FlVec& claim_read, const FlVec& claim_write,
                                      const Fl& claim_audit, const FlVec& eval_ops_val, const FlVec& eval_ops_addr, const FlVec& eval_read_ts,
                                      const Fl& eval_audit_ts, const FlVec& r, const Fl& r_hash, const Fl& r_ms) {
  const Fl r_hash_sqr = fl_mul(r_hash, r_hash);
  auto hash = [&](const Fl& addr, const Fl& val, const Fl& ts) { return fl_add(fl_add(fl_mul(ts, r_hash_sqr), fl_mul(val, r_hash)), addr); };
  Fl eval_init_addr = fl_zero();  // IdentityPolynomial::evaluate, dense_mlpoly.rs:108-114
  const size_t len = rand_mem.size();
  for (size_t i = 0; i < len; i++) eval_init_addr = fl_add(eval_init_addr, fl_mul(fl_from_u64(uint64_t(1) << (len - i - 1)), rand_mem[i]));
  if (r.size() != rand_mem.size()) return false;
  Fl eval_init_val = fl_one();  // EqPolynomial::evaluate, dense_mlpoly.rs:59-65
  for (size_t i = 0; i < r.size(); i++)
    eval_init_val = fl_mul(eval_init_val, fl_add(fl_mul(r[i], rand_mem[i]), fl_mul(fl_sub(fl_one(), r[i]), fl_sub(fl_one(), rand_mem[i]))));
  if (fl_sub(hash(eval_init_addr, eval_init_val, fl_zero()), r_ms) != claim_init) return false;
  if (claim_read.size() != eval_ops_addr.size() || claim_write.size() != eval_ops_addr.size() || eval_ops_val.size() != eval_ops_addr.size() ||
      eval_read_ts.size() != eval_ops_addr.size())
    return false;
  for (size_t i = 0; i < eval_ops_addr.size(); i++) {
    if (fl_sub(hash(eval_ops_addr[i], eval_ops_val[i], eval_read_ts[i]), r_ms) != claim_read[i]) return false;
    if (fl_sub(hash(eval_ops_addr[i], eval_ops_val[i], fl_add(eval_read_ts[i], fl_one())), r_ms) != claim_write[i]) return false;
  }
  return fl_sub(hash(eval_init_addr, eval_init_val, eval_audit_ts), r_ms) == claim_audit;
}

// SparseMatPolyEvalProof::verify, sparse_mlpoly.rs:1540-1582 (+ PolyEvalNetworkProof::verify :1375-1438,
// ProductLayerProof::verify :1232-1330, HashLayerProof::verify :896-1037).  asserts / unwraps become `false`.
static inline bool sparse_eval_verify(const SparseMatPolyEvalProof& pf, const SparseMatPolyCommitment& comm, const FlVec& rx, const FlVec& ry,
                                      const Fl evals[3], const SNARKGens& gens, Transcript& tr) {
  tr.append_protocol_name("Sparse polynomial evaluation proof");
  FlVec rxe, rye;
  equalize(rx, ry, rxe, rye);
  const size_t nz = comm.num_ops, num_mem_cells = comm.num_mem_cells;
  if ((size_t(1) << rxe.size()) != num_mem_cells) return false;
  tr.append_message("derefs_commitment", "begin_derefs_commitment");
  append_poly_commitment(tr, "comm_poly_row_col_ops_val", pf.comm_derefs);
  tr.append_message("derefs_commitment", "end_derefs_commitment");
  FlVec r_mem_check = tr.challenge_vector("challenge_r_hash", 2);
  tr.append_protocol_name("Sparse polynomial evaluation proof");
  const size_t num_instances = 3, num_ops = next_pow2(nz), num_cells = size_t(1) << rxe.size();
  // ---- ProductLayerProof::verify
  const ProductLayerProof& pl = pf.prod;
  tr.append_protocol_name("Sparse polynomial product layer proof");
  auto side = [&](const Fl& init, const FlVec& read, const FlVec& write, const Fl& audit, const char* names[4]) {
    if (read.size() != num_instances || write.size() != num_instances) return false;
    Fl ws = fl_one(), rs = fl_one();
    for (size_t m = 0; m < num_instances; m++) { ws = fl_mul(ws, write[m]); rs = fl_mul(rs, read[m]); }
    if (fl_mul(init, ws) != fl_mul(rs, audit)) return false;
    tr.append_scalar(names[0], init);
    append_scalars(tr, names[1], read);
    append_scalars(tr, names[2], write);
    tr.append_scalar(names[3], audit);
    return true;
  };
  const char* rn[4] = {"claim_row_eval_init", "claim_row_eval_read", "claim_row_eval_write", "claim_row_eval_audit"};
  const char* cn[4] = {"claim_col_eval_init", "claim_col_eval_read", "claim_col_eval_write", "claim_col_eval_audit"};
  if (!side(pl.row_init, pl.row_read, pl.row_write, pl.row_audit, rn)) return false;
  if (!side(pl.col_init, pl.col_read, pl.col_write, pl.col_audit, cn)) return false;
  if (pl.eval_dotp_left.size() != num_instances || pl.eval_dotp_right.size() != num_instances) return false;
  FlVec claims_dotp_circuit;
  for (size_t m = 0; m < num_instances; m++) {
    if (fl_add(pl.eval_dotp_left[m], pl.eval_dotp_right[m]) != evals[m]) return false;
    tr.append_scalar("claim_eval_dotp_left", pl.eval_dotp_left[m]);
    tr.append_scalar("claim_eval_dotp_right", pl.eval_dotp_right[m]);
    claims_dotp_circuit.push_back(pl.eval_dotp_left[m]);
    claims_dotp_circuit.push_back(pl.eval_dotp_right[m]);
  }
  FlVec claims_prod_circuit = pl.row_read;
  claims_prod_circuit.insert(claims_prod_circuit.end(), pl.row_write.begin(), pl.row_write.end());
  claims_prod_circuit.insert(claims_prod_circuit.end(), pl.col_read.begin(), pl.col_read.end());
  claims_prod_circuit.insert(claims_prod_circuit.end(), pl.col_write.begin(), pl.col_write.end());
  FlVec claims_ops, claims_dotp, rand_ops, claims_mem, claims_mem_dotp, rand_mem;
  if (!product_batched_verify(pl.proof_ops, claims_prod_circuit, claims_dotp_circuit, num_ops, tr, claims_ops, claims_dotp, rand_ops)) return false;
  if (!product_batched_verify(pl.proof_mem, {pl.row_init, pl.row_audit, pl.col_init, pl.col_audit}, {}, num_cells, tr, claims_mem, claims_mem_dotp,
                              rand_mem))
    return false;
  if (claims_mem.size() != 4 || claims_ops.size() != 4 * num_instances || claims_dotp.size() != 3 * num_instances) return false;
  FlVec row_read(claims_ops.begin(), claims_ops.begin() + 3), row_write(claims_ops.begin() + 3, claims_ops.begin() + 6),
      col_read(claims_ops.begin() + 6, claims_ops.begin() + 9), col_write(claims_ops.begin() + 9, claims_ops.end());
  // ---- HashLayerProof::verify
  const HashLayerProof& hl = pf.hash;
  tr.append_protocol_name("Sparse polynomial hash layer proof");
  if (hl.eval_derefs_row.size() != hl.eval_derefs_col.size()) return false;
  {  // DerefsEvalProof::verify, :195-215
    tr.append_protocol_name("Derefs evaluation proof");
    FlVec ev = hl.eval_derefs_row;
    ev.insert(ev.end(), hl.eval_derefs_col.begin(), hl.eval_derefs_col.end());
    ev.resize(next_pow2(ev.size()), fl_zero());
    append_scalars(tr, "evals_ops_val", ev);
    FlVec r_joint;
    Fl joint = combine_claims(tr, "challenge_combine_n_to_one", ev, rand_ops, r_joint);
    tr.append_scalar("joint_claim_eval", joint);
    if (!polyeval_verify_plain(hl.proof_derefs, gens.derefs, tr, r_joint, joint, pf.comm_derefs)) return false;
  }
  if (claims_dotp.size() != 3 * hl.eval_derefs_row.size() || hl.eval_val.size() != hl.eval_derefs_row.size()) return false;
  for (size_t i = 0; i < claims_dotp.size() / 3; i++) {
    if (claims_dotp[3 * i] != hl.eval_derefs_row[i] || claims_dotp[3 * i + 1] != hl.eval_derefs_col[i] || claims_dotp[3 * i + 2] != hl.eval_val[i])
      return false;
  }
  {
    FlVec ev = hl.eval_row_addr;
    ev.insert(ev.end(), hl.eval_row_read_ts.begin(), hl.eval_row_read_ts.end());
    ev.insert(ev.end(), hl.eval_col_addr.begin(), hl.eval_col_addr.end());
    ev.insert(ev.end(), hl.eval_col_read_ts.begin(), hl.eval_col_read_ts.end());
    ev.insert(ev.end(), hl.eval_val.begin(), hl.eval_val.end());
    ev.resize(next_pow2(ev.size()), fl_zero());
    append_scalars(tr, "claim_evals_ops", ev);
    FlVec r_joint;
    Fl joint = combine_claims(tr, "challenge_combine_n_to_one", ev, rand_ops, r_joint);
    tr.append_scalar("joint_claim_eval_ops", joint);
    if (!polyeval_verify_plain(hl.proof_ops, gens.ops, tr, r_joint, joint, comm.comm_comb_ops)) return false;
  }
  {
    FlVec ev{hl.eval_row_audit_ts, hl.eval_col_audit_ts};
    append_scalars(tr, "claim_evals_mem", ev);
    FlVec r_joint;
    Fl joint = combine_claims(tr, "challenge_combine_two_to_one", ev, rand_mem, r_joint);
    tr.append_scalar("joint_claim_eval_mem", joint);
    if (!polyeval_verify_plain(hl.proof_mem, gens.mem, tr, r_joint, joint, comm.comm_comb_mem)) return false;
  }
  if (!hash_verify_helper(rand_mem, claims_mem[0], row_read, row_write, claims_mem[1], hl.eval_derefs_row, hl.eval_row_addr, hl.eval_row_read_ts,
                          hl.eval_row_audit_ts, rxe, r_mem_check[0], r_mem_check[1]))
    return false;
  return hash_verify_helper(rand_mem, claims_mem[2], col_read, col_write, claims_mem[3], hl.eval_derefs_col, hl.eval_col_addr, hl.eval_col_read_ts,
                            hl.eval_col_audit_ts, rye, r_mem_check[0], r_mem_check[1]);
}

// ------------------------------------------------ SNARK::prove / verify, lib.rs:335-457
static inline SNARK snark_prove(const Instance& inst, const FlVec& vars_in, const FlVec& inputs, const SNARKGens& gens, Transcript& tr,
                                const Fl& tape_seed, int threads = 1, double* times2 = nullptr) {
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  double t0 = now();
  RandomTape tape("proof", tape_seed);
  tr.append_protocol_name("Spartan SNARK proof");
  FlVec vars = vars_in;
  if (inst.num_vars > vars.size()) vars.resize(inst.num_vars, fl_zero());
  SNARK out;
  FlVec rx, ry;
  out.r1cs_sat_proof = r1cs_prove(inst, vars, inputs, gens.sat, tr, tape, threads, rx, ry);
  double t1 = now();
  inst.evaluate(rx, ry, out.inst_evals);
  tr.append_scalar("Ar_claim", out.inst_evals[0]);
  tr.append_scalar("Br_claim", out.inst_evals[1]);
  tr.append_scalar("Cr_claim", out.inst_evals[2]);
  out.r1cs_eval_proof = sparse_eval_prove(inst, rx, ry, out.inst_evals, gens, tr, tape, threads);
  if (times2) { times2[0] = t1 - t0; times2[1] = now() - t1; }
  return out;
}
static inline bool snark_verify(const SNARK& pf, const R1CSCommitment& comm, const FlVec& inputs, const SNARKGens& gens, Transcript& tr) {
  try {
    tr.append_protocol_name("Spartan SNARK proof");
    if (inputs.size() != comm.num_inputs) return false;
    FlVec rx, ry;
    if (!r1cs_verify(pf.r1cs_sat_proof, comm.num_vars, comm.num_cons, inputs, pf.inst_evals, gens.sat, tr, rx, ry)) return false;
    tr.append_scalar("Ar_claim", pf.inst_evals[0]);
    tr.append_scalar("Br_claim", pf.inst_evals[1]);
    tr.append_scalar("Cr_claim", pf.inst_evals[2]);
    return sparse_eval_verify(pf.r1cs_eval_proof, comm.comm, rx, ry, pf.inst_evals, gens, tr);
  } catch (const std::exception&) {
    return false;
  }
}

// ------------------------------------------------ bincode of SNARK (field order of the serde derives)
struct SnarkWriter : Writer {
  void sumcheck_plain(const SumcheckProof& s) { u64(s.compressed_polys.size()); for (const FlVec& c : s.compressed_polys) fls(c); }
  void pcepb(const ProductCircuitEvalProofBatched& p) {
    u64(p.proof.size());
    for (const LayerProofBatched& l : p.proof) { sumcheck_plain(l.proof); fls(l.claims_prod_left); fls(l.claims_prod_right); }
    for (int t = 0; t < 3; t++) fls(p.claims_dotp[t]);
  }
  void dlog(const DotProductProofLog& d) { pts(d.bullet.L_vec); pts(d.bullet.R_vec); pt(d.delta); pt(d.beta); fl(d.z1); fl(d.z2); }
  void snark(const SNARK& s) {
    r1cs(s.r1cs_sat_proof);
    for (int i = 0; i < 3; i++) fl(s.inst_evals[i]);
    const SparseMatPolyEvalProof& e = s.r1cs_eval_proof;
    pts(e.comm_derefs);
    const ProductLayerProof& p = e.prod;
    fl(p.row_init); fls(p.row_read); fls(p.row_write); fl(p.row_audit);
    fl(p.col_init); fls(p.col_read); fls(p.col_write); fl(p.col_audit);
    fls(p.eval_dotp_left); fls(p.eval_dotp_right);
    pcepb(p.proof_mem); pcepb(p.proof_ops);
    const HashLayerProof& h = e.hash;
    fls(h.eval_row_addr); fls(h.eval_row_read_ts); fl(h.eval_row_audit_ts);
    fls(h.eval_col_addr); fls(h.eval_col_read_ts); fl(h.eval_col_audit_ts);
    fls(h.eval_val);
    fls(h.eval_derefs_row); fls(h.eval_derefs_col);
    dlog(h.proof_ops.proof); dlog(h.proof_mem.proof); dlog(h.proof_derefs.proof);
  }
};
struct SnarkReader : Reader {
  SnarkReader(const uint8_t* d, size_t len) : Reader(d, len) {}
  SumcheckProof sumcheck_plain() {
    SumcheckProof s;
    uint64_t k = u64();
    if (k > n) { ok = false; return s; }
    for (uint64_t i = 0; i < k && ok; i++) s.compressed_polys.push_back(fls());
    return s;
  }
  ProductCircuitEvalProofBatched pcepb() {
    ProductCircuitEvalProofBatched p;
    uint64_t k = u64();
    if (k > n) { ok = false; return p; }
    for (uint64_t i = 0; i < k && ok; i++) {
      LayerProofBatched l;
      l.proof = sumcheck_plain(); l.claims_prod_left = fls(); l.claims_prod_right = fls();
      p.proof.push_back(l);
    }
    for (int t = 0; t < 3; t++) p.claims_dotp[t] = fls();
    return p;
  }
  DotProductProofLog dlog() {
    DotProductProofLog d;
    d.bullet.L_vec = pts(); d.bullet.R_vec = pts(); d.delta = pt(); d.beta = pt(); d.z1 = fl(); d.z2 = fl();
    return d;
  }
  SNARK snark() {
    SNARK s;
    s.r1cs_sat_proof = r1cs();
    for (int i = 0; i < 3; i++) s.inst_evals[i] = fl();
    SparseMatPolyEvalProof& e = s.r1cs_eval_proof;
    e.comm_derefs = pts();
    ProductLayerProof& p = e.prod;
    p.row_init = fl(); p.row_read = fls(); p.row_write = fls(); p.row_audit = fl();
    p.col_init = fl(); p.col_read = fls(); p.col_write = fls(); p.col_audit = fl();
    p.eval_dotp_left = fls(); p.eval_dotp_right = fls();
    p.proof_mem = pcepb(); p.proof_ops = pcepb();
    HashLayerProof& h = e.hash;
    h.eval_row_addr = fls(); h.eval_row_read_ts = fls(); h.eval_row_audit_ts = fl();
    h.eval_col_addr = fls(); h.eval_col_read_ts = fls(); h.eval_col_audit_ts = fl();
    h.eval_val = fls();
    h.eval_derefs_row = fls(); h.eval_derefs_col = fls();
    h.proof_ops.proof = dlog(); h.proof_mem.proof = dlog(); h.proof_derefs.proof = dlog();
    if (pos != n) ok = false;
    return s;
  }
};

}  // namespace orc
