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

 public:
  std::vector<uint8_t>& write(const NizkProof& p) {
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
    scs(p.rx);
    scs(p.ry);
    return b_;
  }
};

}  // namespace sbh
