// ORACLE (test infrastructure only; never linked or included by otti_b200/).
// Keccak-f[1600], SHAKE256, STROBE-128 and the Merlin transcript.  The
// reference takes these from un-vendored crates merlin 3.0.0 / sha3 0.8.2 /
// keccak 0.1.0 (spartan-zkinterface/Cargo.lock); call sites
// Spartan/src/transcript.rs:13-37, random.rs:11-27, commitments.rs:15-33.
// Restated from the published algorithms.
#pragma once
#include <cassert>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "ristretto.hpp"

namespace orc {

static inline uint64_t rol64(uint64_t x, int n) { return n ? (x << n) | (x >> (64 - n)) : x; }

static inline void keccak_f1600(uint64_t A[25]) {
  static const uint64_t RC[24] = {
      0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808AULL, 0x8000000080008000ULL,
      0x000000000000808BULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
      0x000000000000008AULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000AULL,
      0x000000008000808BULL, 0x800000000000008BULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
      0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800AULL, 0x800000008000000AULL,
      0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
  // ROT[x][y]
  static const int ROT[5][5] = {{0, 36, 3, 41, 18}, {1, 44, 10, 45, 2}, {62, 6, 43, 15, 61}, {28, 55, 25, 21, 56}, {27, 20, 39, 8, 14}};
  for (int rnd = 0; rnd < 24; rnd++) {
    uint64_t C[5], D[5], B[25];
    for (int x = 0; x < 5; x++) C[x] = A[x] ^ A[x + 5] ^ A[x + 10] ^ A[x + 15] ^ A[x + 20];
    for (int x = 0; x < 5; x++) D[x] = C[(x + 4) % 5] ^ rol64(C[(x + 1) % 5], 1);
    for (int i = 0; i < 25; i++) A[i] ^= D[i % 5];
    for (int x = 0; x < 5; x++)
      for (int y = 0; y < 5; y++) B[y + 5 * ((2 * x + 3 * y) % 5)] = rol64(A[x + 5 * y], ROT[x][y]);
    for (int x = 0; x < 5; x++)
      for (int y = 0; y < 5; y++) A[x + 5 * y] = B[x + 5 * y] ^ (~B[(x + 1) % 5 + 5 * y] & B[(x + 2) % 5 + 5 * y]);
    A[0] ^= RC[rnd];
  }
}

// SHAKE256 XOF (rate 136)
struct Shake256 {
  uint64_t st[25];
  size_t pos;
  bool squeezing;
  Shake256() : pos(0), squeezing(false) { memset(st, 0, sizeof st); }
  void absorb(const uint8_t* d, size_t n) {
    uint8_t* s = (uint8_t*)st;
    for (size_t i = 0; i < n; i++) {
      s[pos++] ^= d[i];
      if (pos == 136) { keccak_f1600(st); pos = 0; }
    }
  }
  void squeeze(uint8_t* out, size_t n) {
    uint8_t* s = (uint8_t*)st;
    if (!squeezing) {
      s[pos] ^= 0x1f;
      s[135] ^= 0x80;
      keccak_f1600(st);
      pos = 0;
      squeezing = true;
    }
    for (size_t i = 0; i < n; i++) {
      if (pos == 136) { keccak_f1600(st); pos = 0; }
      out[i] = s[pos++];
    }
  }
};

struct Strobe128 {
  static const int R = 166;
  enum { FLAG_I = 1, FLAG_A = 2, FLAG_C = 4, FLAG_T = 8, FLAG_M = 16, FLAG_K = 32 };
  uint64_t st64[25];
  uint8_t pos, pos_begin, cur_flags;
  uint8_t* st() { return (uint8_t*)st64; }
  explicit Strobe128(const char* label) {
    memset(st64, 0, sizeof st64);
    const uint8_t init[6] = {1, R + 2, 1, 0, 1, 96};
    memcpy(st(), init, 6);
    memcpy(st() + 6, "STROBEv1.0.2", 12);
    keccak_f1600(st64);
    pos = pos_begin = cur_flags = 0;
    meta_ad((const uint8_t*)label, strlen(label), false);
  }
  void run_f() {
    st()[pos] ^= pos_begin;
    st()[pos + 1] ^= 0x04;
    st()[R + 1] ^= 0x80;
    keccak_f1600(st64);
    pos = 0;
    pos_begin = 0;
  }
  void absorb(const uint8_t* d, size_t n) {
    for (size_t i = 0; i < n; i++) {
      st()[pos++] ^= d[i];
      if (pos == R) run_f();
    }
  }
  void squeeze(uint8_t* d, size_t n) {
    for (size_t i = 0; i < n; i++) {
      d[i] = st()[pos];
      st()[pos++] = 0;
      if (pos == R) run_f();
    }
  }
  void begin_op(uint8_t flags, bool more) {
    if (more) { assert(cur_flags == flags); return; }
    assert(!(flags & FLAG_T));
    uint8_t old_begin = pos_begin;
    pos_begin = pos + 1;
    cur_flags = flags;
    uint8_t hdr[2] = {old_begin, flags};
    absorb(hdr, 2);
    if ((flags & (FLAG_C | FLAG_K)) && pos != 0) run_f();
  }
  void meta_ad(const uint8_t* d, size_t n, bool more) { begin_op(FLAG_M | FLAG_A, more); absorb(d, n); }
  void ad(const uint8_t* d, size_t n, bool more) { begin_op(FLAG_A, more); absorb(d, n); }
  void prf(uint8_t* d, size_t n, bool more) { begin_op(FLAG_I | FLAG_A | FLAG_C, more); squeeze(d, n); }
};

// merlin::Transcript + ProofTranscript (transcript.rs:13-37)
struct Transcript {
  Strobe128 s;
  explicit Transcript(const std::string& label) : s("Merlin v1.0") {
    append_message("dom-sep", (const uint8_t*)label.data(), label.size());
  }
  void append_message(const char* label, const uint8_t* m, size_t n) {
    uint32_t len = (uint32_t)n;
    s.meta_ad((const uint8_t*)label, strlen(label), false);
    s.meta_ad((const uint8_t*)&len, 4, true);
    s.ad(m, n, false);
  }
  void append_message(const char* label, const char* m) { append_message(label, (const uint8_t*)m, strlen(m)); }
  void challenge_bytes(const char* label, uint8_t* out, size_t n) {
    uint32_t len = (uint32_t)n;
    s.meta_ad((const uint8_t*)label, strlen(label), false);
    s.meta_ad((const uint8_t*)&len, 4, true);
    s.prf(out, n, false);
  }
  void append_protocol_name(const char* name) { append_message("protocol-name", name); }
  void append_scalar(const char* label, const Fl& x) {
    uint8_t b[32];
    fl_to_bytes(x, b);
    append_message(label, b, 32);
  }
  void append_point(const char* label, const uint8_t c[32]) { append_message(label, c, 32); }
  Fl challenge_scalar(const char* label) {
    uint8_t b[64];
    challenge_bytes(label, b, 64);
    return fl_from_bytes_wide(b);
  }
  std::vector<Fl> challenge_vector(const char* label, size_t n) {
    std::vector<Fl> v(n);
    for (size_t i = 0; i < n; i++) v[i] = challenge_scalar(label);
    return v;
  }
};

// random.rs:11-27 with the OsRng-drawn seed scalar injected
struct RandomTape {
  Transcript t;
  RandomTape(const char* name, const Fl& seed) : t(name) { t.append_scalar("init_randomness", seed); }
  Fl random_scalar(const char* l) { return t.challenge_scalar(l); }
  std::vector<Fl> random_vector(const char* l, size_t n) { return t.challenge_vector(l, n); }
};

}  // namespace orc
