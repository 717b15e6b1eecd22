// Proof objects and their bincode 1.3 wire form.  Field order and framing
// follow the reference's serde derives:
//   NIZK            lib.rs:481-485        R1CSProof        r1csproof.rs:20-36
//   ZKSumcheck      sumcheck.rs:64-69     DotProductProof  nizk/mod.rs:291-298
//   KnowledgeProof  nizk/mod.rs:14-19     EqualityProof    nizk/mod.rs:76-80
//   ProductProof    nizk/mod.rs:145-151   DotProductProofLog nizk/mod.rs:416-423
//   BulletReduction nizk/bullet.rs:14-18  PolyCommitment   dense_mlpoly.rs:41-44
// Scalars are written as their raw Montgomery limbs (scalar/ristretto255.rs:199-200).
#pragma once
#include <array>
#include <cstdint>
#include <vector>

#include "hfl.hpp"

namespace sbh {

typedef std::array<uint8_t, 32> Cpt;  // CompressedRistretto

struct DotProd {
  Cpt delta, beta;
  FlVec z;
  Fl z_delta, z_beta;
};
struct ZkSumcheck {
  std::vector<Cpt> comm_polys, comm_evals;
  std::vector<DotProd> proofs;
};
struct Knowledge {
  Cpt alpha;
  Fl z1, z2;
};
struct Equality {
  Cpt alpha;
  Fl z;
};
struct Product {
  Cpt alpha, beta, delta;
  Fl z[5];
};
struct DotProdLog {
  std::vector<Cpt> L_vec, R_vec;
  Cpt delta, beta;
  Fl z1, z2;
};
struct NizkProof {
  std::vector<Cpt> comm_vars;
  ZkSumcheck sc1;
  Cpt claims_phase2[4];
  Knowledge pok_Cz;
  Product prod;
  Equality eq1;
  ZkSumcheck sc2;
  Cpt comm_vars_at_ry;
  DotProdLog eval_proof;
  Equality eq2;
  FlVec rx, ry;
};

// ---- SNARK (lib.rs:307-312): R1CSProof + the three claimed evaluations + R1CSEvalProof
struct PlainSumcheck { std::vector<FlVec> compressed_polys; };  // SumcheckInstanceProof of CompressedUniPoly, sumcheck.rs:17-20, unipoly.rs:15-19
struct LayerProofBatched {                                      // product_tree.rs:147-152
  PlainSumcheck proof;
  FlVec claims_prod_left, claims_prod_right;
};
struct ProductBatched {  // ProductCircuitEvalProofBatched, product_tree.rs:177-181
  std::vector<LayerProofBatched> proof;
  FlVec claims_dotp[3];
};
struct HashLayer {  // HashLayerProof, sparse_mlpoly.rs:671-680
  FlVec eval_row_addr, eval_row_read_ts;
  Fl eval_row_audit_ts;
  FlVec eval_col_addr, eval_col_read_ts;
  Fl eval_col_audit_ts;
  FlVec eval_val, eval_derefs_row, eval_derefs_col;
  DotProdLog proof_ops, proof_mem, proof_derefs;
};
struct ProductLayer {  // ProductLayerProof, sparse_mlpoly.rs:1040-1047
  Fl row_init;
  FlVec row_read, row_write;
  Fl row_audit;
  Fl col_init;
  FlVec col_read, col_write;
  Fl col_audit;
  FlVec eval_dotp_left, eval_dotp_right;
  ProductBatched proof_mem, proof_ops;
};
struct SnarkProof {
  NizkProof sat;  // rx, ry are not part of the wire form here
  Fl inst_evals[3];
  std::vector<Cpt> comm_derefs;  // DerefsCommitment, sparse_mlpoly.rs:42-44
  ProductLayer prod;
  HashLayer hash;
};

class BincodeOut {
  std::vector<uint8_t> b_;
  void u64(uint64_t x) {
    for (int i = 0; i < 8; i++) b_.push_back((uint8_t)(x >> (8 * i)));
  }
  void pt(const Cpt& c) { b_.insert(b_.end(), c.begin(), c.end()); }
  void sc(const Fl& x) {
    const uint8_t* p = (const uint8_t*)x.v;
    b_.insert(b_.end(), p, p + 32);
  }
  void pts(const std::vector<Cpt>& v) {
    u64(v.size());
    for (const Cpt& c : v) pt(c);
  }
  void scs(const FlVec& v) {
    u64(v.size());
    for (const Fl& x : v) sc(x);
  }
  void sumcheck(const ZkSumcheck& s) {
    pts(s.comm_polys);
    pts(s.comm_evals);
    u64(s.proofs.size());
    for (const DotProd& d : s.proofs) {
      pt(d.delta); pt(d.beta); scs(d.z); sc(d.z_delta); sc(d.z_beta);
    }
  }

  void dlog(const DotProdLog& d) {
    pts(d.L_vec); pts(d.R_vec);
    pt(d.delta); pt(d.beta); sc(d.z1); sc(d.z2);
  }
  void batched(const ProductBatched& p) {
    u64(p.proof.size());
    for (const LayerProofBatched& l : p.proof) {
      u64(l.proof.compressed_polys.size());
      for (const FlVec& c : l.proof.compressed_polys) scs(c);
      scs(l.claims_prod_left);
      scs(l.claims_prod_right);
    }
    for (int t = 0; t < 3; t++) scs(p.claims_dotp[t]);
  }

 public:
  // bincode(SNARK): r1cs_sat_proof, inst_evals, r1cs_eval_proof (field order of the serde derives)
  std::vector<uint8_t>& write(const SnarkProof& s) {
    r1cs(s.sat);
    for (int i = 0; i < 3; i++) sc(s.inst_evals[i]);
    pts(s.comm_derefs);
    const ProductLayer& p = s.prod;
    sc(p.row_init); scs(p.row_read); scs(p.row_write); sc(p.row_audit);
    sc(p.col_init); scs(p.col_read); scs(p.col_write); sc(p.col_audit);
    scs(p.eval_dotp_left); scs(p.eval_dotp_right);
    batched(p.proof_mem); batched(p.proof_ops);
    const HashLayer& h = s.hash;
    scs(h.eval_row_addr); scs(h.eval_row_read_ts); sc(h.eval_row_audit_ts);
    scs(h.eval_col_addr); scs(h.eval_col_read_ts); sc(h.eval_col_audit_ts);
    scs(h.eval_val);
    scs(h.eval_derefs_row); scs(h.eval_derefs_col);
    dlog(h.proof_ops); dlog(h.proof_mem); dlog(h.proof_derefs);
    return b_;
  }
  std::vector<uint8_t>& write(const NizkProof& p) {
    r1cs(p);
    scs(p.rx);
    scs(p.ry);
    return b_;
  }

 private:
  void r1cs(const NizkProof& p) {
    pts(p.comm_vars);
    sumcheck(p.sc1);
    for (int i = 0; i < 4; i++) pt(p.claims_phase2[i]);
    pt(p.pok_Cz.alpha); sc(p.pok_Cz.z1); sc(p.pok_Cz.z2);
    pt(p.prod.alpha); pt(p.prod.beta); pt(p.prod.delta);
    for (int i = 0; i < 5; i++) sc(p.prod.z[i]);
    pt(p.eq1.alpha); sc(p.eq1.z);
    sumcheck(p.sc2);
    pt(p.comm_vars_at_ry);
    pts(p.eval_proof.L_vec); pts(p.eval_proof.R_vec);
    pt(p.eval_proof.delta); pt(p.eval_proof.beta); sc(p.eval_proof.z1); sc(p.eval_proof.z2);
    pt(p.eq2.alpha); sc(p.eq2.z);
  }
};

}  // namespace sbh
