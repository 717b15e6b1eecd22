// Host-side Fiat-Shamir: Keccak-f[1600], SHAKE256, STROBE-128 and the Merlin
// transcript, framed exactly as Spartan's ProofTranscript / AppendToTranscript
// (Spartan/src/transcript.rs:13-63).  north_star keeps transcript hashing on
// the host; across the C ABI a caller may instead pass a vtable over its own
// merlin::Transcript (include/spartan_b200.h, sb200_transcript_vt).
#pragma once
#include <cstdint>
#include <cstring>
#include <string>

#include "hfl.hpp"

namespace sbh {

static inline uint64_t rotl64(uint64_t x, unsigned n) { return (x << n) | (x >> (64 - n)); }

static inline void keccakf(uint64_t s[25]) {
  static const uint64_t RC[24] = {0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
                                  0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
                                  0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
                                  0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
                                  0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
                                  0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
  static const unsigned ROTC[24] = {1, 3, 6, 10, 15, 21, 28, 36, 45, 55, 2, 14, 27, 41, 56, 8, 25, 43, 62, 18, 39, 61, 20, 44};
  static const unsigned PILN[24] = {10, 7, 11, 17, 18, 3, 5, 16, 8, 21, 24, 4, 15, 23, 19, 13, 12, 2, 20, 14, 22, 9, 6, 1};
  for (int round = 0; round < 24; round++) {
    uint64_t bc[5];
    for (int i = 0; i < 5; i++) bc[i] = s[i] ^ s[i + 5] ^ s[i + 10] ^ s[i + 15] ^ s[i + 20];
    for (int i = 0; i < 5; i++) {
      uint64_t t = bc[(i + 4) % 5] ^ rotl64(bc[(i + 1) % 5], 1);
      for (int j = 0; j < 25; j += 5) s[j + i] ^= t;
    }
    uint64_t t = s[1];
    for (int i = 0; i < 24; i++) {
      unsigned j = PILN[i];
      uint64_t b = s[j];
      s[j] = rotl64(t, ROTC[i]);
      t = b;
    }
    for (int j = 0; j < 25; j += 5) {
      for (int i = 0; i < 5; i++) bc[i] = s[j + i];
      for (int i = 0; i < 5; i++) s[j + i] ^= (~bc[(i + 1) % 5]) & bc[(i + 2) % 5];
    }
    s[0] ^= RC[round];
  }
}

class Shake256Xof {
  uint64_t s_[25];
  size_t pos_ = 0;
  bool out_ = false;

 public:
  Shake256Xof() { memset(s_, 0, sizeof s_); }
  void update(const void* data, size_t n) {
    const uint8_t* d = (const uint8_t*)data;
    uint8_t* b = (uint8_t*)s_;
    while (n--) {
      b[pos_++] ^= *d++;
      if (pos_ == 136) { keccakf(s_); pos_ = 0; }
    }
  }
  void read(uint8_t* out, size_t n) {
    uint8_t* b = (uint8_t*)s_;
    if (!out_) {
      b[pos_] ^= 0x1f;
      b[135] ^= 0x80;
      keccakf(s_);
      pos_ = 0;
      out_ = true;
    }
    while (n--) {
      if (pos_ == 136) { keccakf(s_); pos_ = 0; }
      *out++ = b[pos_++];
    }
  }
};

// STROBE-128/1600 restricted to the three operations merlin uses.
class Strobe {
  static const unsigned RATE = 166;
  uint64_t s_[25];
  uint8_t pos_ = 0, begin_ = 0, flags_ = 0;
  uint8_t* b() { return (uint8_t*)s_; }
  void permute() {
    b()[pos_] ^= begin_;
    b()[pos_ + 1] ^= 0x04;
    b()[RATE + 1] ^= 0x80;
    keccakf(s_);
    pos_ = 0;
    begin_ = 0;
  }
  void absorb(const uint8_t* d, size_t n) {
    while (n--) {
      b()[pos_++] ^= *d++;
      if (pos_ == RATE) permute();
    }
  }
  void begin(uint8_t flags, bool more) {
    if (more) return;
    uint8_t hdr[2] = {begin_, flags};
    begin_ = pos_ + 1;
    flags_ = flags;
    absorb(hdr, 2);
    if ((flags & (4 | 32)) && pos_ != 0) permute();
  }

 public:
  explicit Strobe(const char* proto) {
    memset(s_, 0, sizeof s_);
    const uint8_t hdr[6] = {1, RATE + 2, 1, 0, 1, 96};
    memcpy(b(), hdr, 6);
    memcpy(b() + 6, "STROBEv1.0.2", 12);
    keccakf(s_);
    meta_ad(proto, strlen(proto), false);
  }
  void meta_ad(const void* d, size_t n, bool more) { begin(16 | 2, more); absorb((const uint8_t*)d, n); }
  void ad(const void* d, size_t n, bool more) { begin(2, more); absorb((const uint8_t*)d, n); }
  void prf(uint8_t* out, size_t n) {
    begin(1 | 2 | 4, false);
    while (n--) {
      *out++ = b()[pos_];
      b()[pos_++] = 0;
      if (pos_ == RATE) permute();
    }
  }
};

// Abstract transcript so the C ABI can route to a caller-owned merlin::Transcript.
struct TranscriptIface {
  virtual ~TranscriptIface() {}
  virtual void append_message(const char* label, const uint8_t* m, size_t n) = 0;
  virtual void challenge_bytes(const char* label, uint8_t* out, size_t n) = 0;
  // ProofTranscript, transcript.rs:13-37
  void append_protocol_name(const char* name) { append_message("protocol-name", (const uint8_t*)name, strlen(name)); }
  void append_str(const char* label, const char* m) { append_message(label, (const uint8_t*)m, strlen(m)); }
  void append_scalar(const char* label, const Fl& x) {
    uint8_t buf[32];
    fl_to_bytes(x, buf);
    append_message(label, buf, 32);
  }
  void append_point(const char* label, const uint8_t c[32]) { append_message(label, c, 32); }
  Fl challenge_scalar(const char* label) {
    uint8_t buf[64];
    challenge_bytes(label, buf, 64);
    return fl_from_bytes_wide(buf);
  }
  FlVec challenge_vector(const char* label, size_t n) {
    FlVec v(n);
    for (size_t i = 0; i < n; i++) v[i] = challenge_scalar(label);
    return v;
  }
};

class MerlinTranscript : public TranscriptIface {
  Strobe st_;

 public:
  MerlinTranscript(const void* label, size_t n) : st_("Merlin v1.0") { append_message_n("dom-sep", 7, (const uint8_t*)label, n); }
  void append_message_n(const void* label, size_t ll, const uint8_t* m, size_t n) {
    uint32_t len = (uint32_t)n;
    st_.meta_ad(label, ll, false);
    st_.meta_ad(&len, 4, true);
    st_.ad(m, n, false);
  }
  void challenge_bytes_n(const void* label, size_t ll, uint8_t* out, size_t n) {
    uint32_t len = (uint32_t)n;
    st_.meta_ad(label, ll, false);
    st_.meta_ad(&len, 4, true);
    st_.prf(out, n);
  }
  void append_message(const char* label, const uint8_t* m, size_t n) override { append_message_n(label, strlen(label), m, n); }
  void challenge_bytes(const char* label, uint8_t* out, size_t n) override { challenge_bytes_n(label, strlen(label), out, n); }
};

}  // namespace sbh
