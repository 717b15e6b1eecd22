// RandomTape (Spartan/src/random.rs:11-27) drawn AHEAD of the protocol.
//
// The tape is a private merlin transcript seeded once; every draw is
// `challenge_scalar(label)`, and both the labels and their order are fixed by the instance
// dimensions alone -- no draw depends on protocol data.  So the whole
// sequence can be produced by one background thread while the GPU commits to the witness,
// and every commitment whose scalars are tape-only can be computed before its round.
// The values are exactly what the reference's in-line draws return.
#pragma once
#include <atomic>
#include <vector>

#include "hmerlin.hpp"
#include "pool.hpp"

namespace sbh {

struct TapeLayout {
  size_t Ls, s_x, s_y, lgR;
  // offsets into the draw sequence
  size_t poly_blinds;                  // Ls                         dense_mlpoly.rs:194
  size_t sc1_blinds_poly, sc1_blinds_evals;  // s_x each             sumcheck.rs:603-606
  size_t sc1_rounds;                   // s_x x (d_vec[4], r_delta, r_beta)   nizk/mod.rs:329-331
  size_t claim_blinds;                 // Az, Bz, Cz, prod           r1csproof.rs:209-214
  size_t pok;                          // t1, t2                     nizk/mod.rs:36-37
  size_t prod;                         // b1..b5                     nizk/mod.rs:177-181
  size_t eq1_r;                        // r                          nizk/mod.rs:99
  size_t sc2_blinds_poly, sc2_blinds_evals;  // s_y each             sumcheck.rs:441-444
  size_t sc2_rounds;                   // s_y x (d_vec[3], r_delta, r_beta)
  size_t blind_eval;                   //                            r1csproof.rs:296
  size_t dlog;                         // d, r_delta, r_beta (label "r_delta", sic)  nizk/mod.rs:452-454
  size_t bv1, bv2;                     // 2 lgR each                 nizk/mod.rs:456-457
  size_t eq2_r;
  // SNARK::prove continues on the same tape: one DotProductProofLog::prove draw sequence (d, r_delta, r_beta, 2 lg n
  // blinds_vec_1, 2 lg n blinds_vec_2) per polynomial evaluation proof of the R1CSEvalProof, in proving order
  // (derefs, ops, mem: sparse_mlpoly.rs:723-837); extra_lg[k] = lg n of proof k, extra_off[k] its first draw
  std::vector<size_t> extra_lg, extra_off;
  size_t total;
  TapeLayout(size_t Ls_, size_t s_x_, size_t s_y_, size_t lgR_, const std::vector<size_t>& extra = {})
      : Ls(Ls_), s_x(s_x_), s_y(s_y_), lgR(lgR_), extra_lg(extra) {
    size_t o = 0;
    poly_blinds = o; o += Ls;
    sc1_blinds_poly = o; o += s_x;
    sc1_blinds_evals = o; o += s_x;
    sc1_rounds = o; o += 6 * s_x;
    claim_blinds = o; o += 4;
    pok = o; o += 2;
    prod = o; o += 5;
    eq1_r = o; o += 1;
    sc2_blinds_poly = o; o += s_y;
    sc2_blinds_evals = o; o += s_y;
    sc2_rounds = o; o += 5 * s_y;
    blind_eval = o; o += 1;
    dlog = o; o += 3;
    bv1 = o; o += 2 * lgR;
    bv2 = o; o += 2 * lgR;
    eq2_r = o; o += 1;
    for (size_t lg : extra_lg) { extra_off.push_back(o); o += 3 + 4 * lg; }
    total = o;
  }
};

class TapeAhead {
  MerlinTranscript t_;
  std::vector<Fl> vals_;
  std::atomic<size_t> ready_{0};

  void draw(const char* label, size_t n, size_t& k) {
    for (size_t i = 0; i < n; i++) {
      vals_[k] = t_.challenge_scalar(label);
      ready_.store(++k, std::memory_order_release);
    }
  }

 public:
  const TapeLayout lay;
  TapeAhead(const Fl& seed, const TapeLayout& l) : t_("proof", 5), vals_(l.total), lay(l) {
    t_.append_scalar("init_randomness", seed);  // random.rs:14-15
  }
  // The whole draw sequence, in the reference's order, in two parts: the commitment blinds, which the caller of
  // NIZK::prove is about to wait for and may as well draw itself (a sharded proof on a box with four cores per rank
  // saw 0.8 - 1.6 ms for them on a worker competing with the spinning caller, against 0.73 ms of sequential hashing),
  // and everything after them, which only the tape-only commitments of the worker pool read.
  void run() {
    run_blinds();
    run_rest();
  }
  void run_blinds() {
    size_t k = 0;
    draw("poly_blinds", lay.Ls, k);
  }
  void run_rest() {
    size_t k = lay.Ls;
    draw("blinds_poly", lay.s_x, k);
    draw("blinds_evals", lay.s_x, k);
    for (size_t j = 0; j < lay.s_x; j++) { draw("d_vec", 4, k); draw("r_delta", 1, k); draw("r_beta", 1, k); }
    draw("Az_blind", 1, k); draw("Bz_blind", 1, k); draw("Cz_blind", 1, k); draw("prod_Az_Bz_blind", 1, k);
    draw("t1", 1, k); draw("t2", 1, k);
    draw("b1", 1, k); draw("b2", 1, k); draw("b3", 1, k); draw("b4", 1, k); draw("b5", 1, k);
    draw("r", 1, k);
    draw("blinds_poly", lay.s_y, k);
    draw("blinds_evals", lay.s_y, k);
    for (size_t j = 0; j < lay.s_y; j++) { draw("d_vec", 3, k); draw("r_delta", 1, k); draw("r_beta", 1, k); }
    draw("blind_eval", 1, k);
    draw("d", 1, k); draw("r_delta", 1, k); draw("r_delta", 1, k);  // sic: r_beta is drawn under the label "r_delta"
    draw("blinds_vec_1", 2 * lay.lgR, k);
    draw("blinds_vec_2", 2 * lay.lgR, k);
    draw("r", 1, k);
    for (size_t lg : lay.extra_lg) {
      draw("d", 1, k); draw("r_delta", 1, k); draw("r_delta", 1, k);
      draw("blinds_vec_1", 2 * lg, k);
      draw("blinds_vec_2", 2 * lg, k);
    }
  }
  // value #i of the sequence; waits until the drawing thread has produced it
  const Fl& at(size_t i) const {
    while (ready_.load(std::memory_order_acquire) <= i) SB_CPU_RELAX();
    return vals_[i];
  }
  void wait_until(size_t n) const {
    while (ready_.load(std::memory_order_acquire) < n) SB_CPU_RELAX();
  }
};

}  // namespace sbh
