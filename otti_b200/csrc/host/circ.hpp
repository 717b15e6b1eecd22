// rust-circ adapter: a CirC R1cs (linear combinations over arbitrary-precision integers, signals by name) -> the
// COO triples and assignments libspartan's Instance::new / Assignment::new take.  Replaces `r1cs_to_spartan`,
// `lc_to_v` and `int_to_scalar` of rust-circ/src/target/r1cs/spartan.rs:31-186 in front of the prove path, so that
// the SGD pipeline can hand its circuit over in memory instead of through zkinterface files.
//
// The reference iterates two Rust HashMaps (`r1cs.values`, `lc.monomials`), whose order is unspecified; here the
// caller's order is THE order (signals and monomials are taken as listed).  Everything else follows the reference
// statement by statement, including what it asserts / unwraps (reported as an error string instead of a panic).
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "hfl.hpp"
#include "zkif.hpp"

namespace sbh {

// int_to_scalar, spartan.rs:141-160: Horner over the 64-bit limbs of |i|, most significant first, with
// m = 2^63 * 2.  rug's `as_ref()` exposes the limbs of the magnitude, so the sign never enters.
// `mag` = little-endian magnitude bytes, any length.
static inline Fl circ_int_to_scalar(const uint8_t* mag, size_t len) {
  const Fl m = fl_mul(fl_from_u64(1ull << 63), fl_from_u64(2));
  Fl acc = fl_zero();
  const size_t limbs = (len + 7) / 8;
  for (size_t k = limbs; k-- > 0;) {
    uint64_t d = 0;
    for (size_t b = 0; b < 8 && 8 * k + b < len; b++) d |= (uint64_t)mag[8 * k + b] << (8 * b);
    acc = fl_add(fl_mul(acc, m), fl_from_u64(d));
  }
  return acc;
}

struct CircBigInts {  // integer k = bytes[off[k], off[k+1])
  const uint8_t* bytes;
  const uint64_t* off;
  Fl at(uint64_t k) const { return circ_int_to_scalar(bytes + off[k], (size_t)(off[k + 1] - off[k])); }
  bool is_zero(uint64_t k) const {
    for (uint64_t i = off[k]; i < off[k + 1]; i++)
      if (bytes[i]) return false;
    return true;
  }
};

struct CircLcs {  // the A, B or C side of every constraint
  const uint64_t* lc_off;    // n_cons + 1: monomials of constraint i are [lc_off[i], lc_off[i+1])
  const uint64_t* mono_ids;  // CirC variable id of each monomial
  CircBigInts coeff;         // one per monomial
  CircBigInts constant;      // one per constraint (lc.constant)
};

static inline bool name_contains(const uint8_t* s, size_t n, const char* pat) {
  const size_t m = strlen(pat);
  if (m > n) return false;
  for (size_t i = 0; i + m <= n; i++)
    if (memcmp(s + i, pat, m) == 0) return true;
  return false;
}

// r1cs_to_spartan, spartan.rs:31-139, into the same container the zkinterface reader fills
static bool circ_r1cs_to_spartan(uint64_t n_signals, const uint64_t* signal_ids, const uint8_t* names, const uint64_t* name_off,
                                 const CircBigInts& values, uint64_t n_public, uint64_t next_idx, uint64_t n_cons,
                                 const CircLcs lcs[3], ZkR1cs& out) {
  if (n_public != 0) { out.error = "assert r1cs.public_idxs.len() == 0"; return false; }  // spartan.rs:42
  std::unordered_map<uint64_t, uint64_t> trans, itrans;
  trans.reserve(n_signals * 2);
  uint64_t nwit = 0, ninp = 0;
  uint8_t b[32];
  for (uint64_t s = 0; s < n_signals; s++) {  // spartan.rs:47-75
    const uint8_t* nm = names + name_off[s];
    const size_t nl = (size_t)(name_off[s + 1] - name_off[s]);
    fl_to_bytes(values.at(s), b);
    const bool input = !name_contains(nm, nl, "main_f0_lex0_w") && name_contains(nm, nl, "main_f0_lex0_");
    if (input) {
      itrans[signal_ids[s]] = ninp++;  // HashMap::insert: a repeated id keeps the last position
      out.inputs.insert(out.inputs.end(), b, b + 32);
    } else {
      trans[signal_ids[s]] = nwit++;
      out.vars.insert(out.vars.end(), b, b + 32);
    }
  }
  if (nwit + ninp != next_idx) { out.error = "assert_eq!(wit.len() + inp.len(), r1cs.next_idx)"; return false; }  // spartan.rs:80
  const uint64_t const_id = nwit;
  for (const auto& kv : itrans) trans[kv.first] = kv.second + const_id + 1;  // spartan.rs:96-99
  for (uint64_t i = 0; i < n_cons; i++) {  // spartan.rs:110-130, lc_to_v :164-186
    for (int m = 0; m < 3; m++) {
      const CircLcs& L = lcs[m];
      for (uint64_t t = L.lc_off[i]; t < L.lc_off[i + 1]; t++) {
        const uint64_t k = L.mono_ids[t];
        if (k >= const_id) { out.error = "Error: variable number off"; return false; }  // spartan.rs:168
        auto it = trans.find(k);
        if (it == trans.end()) { out.error = "trans.get(k).unwrap() on an unknown variable id"; return false; }
        fl_to_bytes(L.coeff.at(t), b);
        out.rows[m].push_back(i);
        out.cols[m].push_back(it->second);
        out.vals[m].insert(out.vals[m].end(), b, b + 32);
      }
      if (!L.constant.is_zero(i)) {  // spartan.rs:177-184
        fl_to_bytes(L.constant.at(i), b);
        out.rows[m].push_back(i);
        out.cols[m].push_back(const_id);
        out.vals[m].insert(out.vals[m].end(), b, b + 32);
      }
    }
  }
  out.num_cons = n_cons;
  out.num_vars = nwit;
  out.num_inputs = ninp;
  uint64_t nz = 0;
  for (int m = 0; m < 3; m++) nz = std::max<uint64_t>(nz, out.rows[m].size());
  out.non_zero_entries = nz;
  memset(out.field_max, 0, 32);
  return true;
}

}  // namespace sbh
