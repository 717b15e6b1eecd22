// Otti-shaped R1CS generator + zkinterface writer: INPUT infrastructure for tests/ and bench.py (not product, not oracle).
//
// Real Otti circuits (CirC output for the LP / SDP / SGD checkers) cannot be produced in this image (no GHC / cargo /
// z3), so the sizes BASELINE.json names for configs[2..3] are covered by shape-substituted instances (SURVEY.md
// section 8(d) item ii): what CirC emits is dominated by bit decompositions of fixed-point values
// (compiler/src/IR/SMT/ToPf.hs:558-560,644-646) and by products of affine forms with constants, so
//   * 80 % of the witness are bits, the rest 32-bit words assembled from them; inputs are 32-bit words;
//   * rows (period 20): 6 booleanity rows  b * (b - 1) = 0;  1 decomposition row  (sum 2^k b_k) * 1 = w  (34 entries);
//     11 affine products (x + c1) * (y + c2) = t + c3;  2 short random combinations (1..8 terms per side);
//   * the constant-1 column therefore holds ~30 % of all non-zero entries -- ONE very heavy column, the case a
//     column-per-thread sparse kernel cannot survive ("What's weak" 5).
// Every value is a small signed integer (|v| < 2^110), encoded mod l as v or l - |v|, so no field arithmetic is needed
// here and the generator shares no code with either the product or the oracle.
//
// The same constraints are returned as COO triples (what Instance::new takes) AND as the three zkinterface flatbuffers
// spzk reads (schema: spartan-zkinterface/schema/zkinterface.fbs; forward layout: a table precedes what it points to).
//
//   g++ -O2 -shared -fPIC -o tools/_build/libzkifgen.so tools/zkifgen.cpp
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

typedef __int128 i128;
typedef unsigned __int128 u128;

const uint64_t ELL[4] = {0x5812631a5cf5d3edull, 0x14def9dea2f79cd6ull, 0, 0x1000000000000000ull};

void enc(i128 v, uint8_t out[32]) {  // v mod l, canonical little-endian
  uint64_t w[4] = {0, 0, 0, 0};
  if (v >= 0) {
    w[0] = (uint64_t)v;
    w[1] = (uint64_t)((u128)v >> 64);
  } else {
    u128 m = (u128)(-v);
    uint64_t lo = (uint64_t)m, hi = (uint64_t)(m >> 64);
    u128 d = (u128)ELL[0] - lo;
    w[0] = (uint64_t)d;
    uint64_t borrow = (ELL[0] < lo) ? 1 : 0;
    u128 d1 = (u128)ELL[1] - hi - borrow;
    w[1] = (uint64_t)d1;
    borrow = ((u128)ELL[1] < (u128)hi + borrow) ? 1 : 0;
    w[2] = ELL[2] - borrow;  // 0 - borrow
    uint64_t b2 = borrow;    // ELL[2] == 0: borrows again exactly when borrow == 1
    w[3] = ELL[3] - b2;
  }
  memcpy(out, w, 32);
}

struct Rng {  // SplitMix64
  uint64_t s;
  uint64_t next() {
    uint64_t z = (s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
  }
  uint64_t below(uint64_t n) { return next() % n; }
};

struct Term {
  uint64_t col;
  i128 coef;
};

struct Buf {  // forward-layout flatbuffer writer
  std::vector<uint8_t> b;
  void align(size_t n) { while (b.size() % n) b.push_back(0); }
  void put32(size_t at, uint32_t v) { memcpy(&b[at], &v, 4); }
  // table with `nf` fields, all uoffsets (kind 0), or u8 (1) / u64 (2); returns table position; field positions in at[]
  size_t table(int nf, const int* kind, size_t* at) {
    size_t off = 4, offs[4];
    for (int i = 0; i < nf; i++) {
      size_t size = kind[i] == 1 ? 1 : kind[i] == 2 ? 8 : 4;
      off = (off + size - 1) / size * size;
      offs[i] = off;
      off += size;
    }
    const size_t tsize = off, vtsize = 4 + 2 * (size_t)nf;
    align(8);
    size_t pad = (8 - (b.size() + vtsize) % 8) % 8;
    b.insert(b.end(), pad, 0);
    const size_t vpos = b.size();
    uint16_t h[2] = {(uint16_t)vtsize, (uint16_t)tsize};
    b.insert(b.end(), (uint8_t*)h, (uint8_t*)h + 4);
    for (int i = 0; i < nf; i++) {
      uint16_t o = (uint16_t)offs[i];
      b.insert(b.end(), (uint8_t*)&o, (uint8_t*)&o + 2);
    }
    const size_t tpos = b.size();
    b.insert(b.end(), tsize, 0);
    int32_t so = (int32_t)(tpos - vpos);
    memcpy(&b[tpos], &so, 4);
    for (int i = 0; i < nf; i++) at[i] = tpos + offs[i];
    return tpos;
  }
  void link(size_t at, size_t child) { put32(at, (uint32_t)(child - at)); }
  size_t vec_begin(size_t elem, uint32_t count) {
    align(4);
    if (elem == 8) while ((b.size() + 4) % 8) b.push_back(0);
    size_t pos = b.size();
    b.insert(b.end(), (uint8_t*)&count, (uint8_t*)&count + 4);
    return pos;
  }
};

// Variables { variable_ids: [u64], values: [u8] } with 32-byte values
size_t write_variables(Buf& B, const std::vector<uint64_t>& ids, const std::vector<uint8_t>& vals) {
  const int kinds[2] = {0, 0};
  size_t at[2];
  size_t t = B.table(2, kinds, at);
  size_t v0 = B.vec_begin(8, (uint32_t)ids.size());
  B.b.insert(B.b.end(), (const uint8_t*)ids.data(), (const uint8_t*)ids.data() + 8 * ids.size());
  B.link(at[0], v0);
  size_t v1 = B.vec_begin(1, (uint32_t)vals.size());
  B.b.insert(B.b.end(), vals.begin(), vals.end());
  B.link(at[1], v1);
  return t;
}

struct Gen {
  uint64_t nc, nv, ni;
  std::vector<std::vector<Term>> lc[3];
  std::vector<int64_t> z;  // vars || 1 || inputs, all < 2^32
  std::vector<uint8_t> vars, inputs, file[3];
  uint64_t const_entries = 0;

  uint64_t wit_id(uint64_t i) const { return 1000 + 3 * i; }
  uint64_t in_id(uint64_t i) const { return 1 + i; }
  uint64_t var_id(uint64_t col) const { return col < nv ? wit_id(col) : col == nv ? 0 : in_id(col - nv - 1); }

  i128 eval(const std::vector<Term>& l) const {
    i128 s = 0;
    for (const Term& t : l) s += t.coef * (i128)z[t.col];
    return s;
  }

  void generate(uint64_t seed) {
    Rng rng{seed * 0x2545f4914f6cdd1dull + 99};
    const uint64_t nb = nv * 4 / 5 ? nv * 4 / 5 : 1, nw = nv - nb, ONE = nv, ncols = nv + 1 + ni;
    z.assign(ncols, 0);
    for (uint64_t i = 0; i < nb; i++) z[i] = (int64_t)(rng.next() & 1);
    for (uint64_t j = 0; j < nw; j++) {
      int64_t w = 0;
      for (int k = 0; k < 32; k++) w |= z[(j * 32 + k) % nb] << k;
      z[nb + j] = w;
    }
    z[ONE] = 1;
    for (uint64_t i = 0; i < ni; i++) z[nv + 1 + i] = (int64_t)(rng.next() & 0xffffffffu);
    for (int m = 0; m < 3; m++) lc[m].resize(nc);
    auto target = [&](uint64_t r, i128 p) {  // C = t + c3 with c3 = p - z[t]
      uint64_t t = nw ? nb + rng.below(nw) : rng.below(nb);
      lc[2][r] = {{t, 1}, {ONE, p - (i128)z[t]}};
    };
    for (uint64_t r = 0; r < nc; r++) {
      const uint64_t k = r % 20;
      if (k < 6 || (k == 6 && nw == 0)) {
        uint64_t b = (r * 7 + k) % nb;
        lc[0][r] = {{b, 1}};
        lc[1][r] = {{b, 1}, {ONE, -1}};
      } else if (k == 6) {
        uint64_t j = (r / 20) % nw;
        for (int q = 0; q < 32; q++) lc[0][r].push_back({(j * 32 + q) % nb, (i128)1 << q});
        lc[1][r] = {{ONE, 1}};
        lc[2][r] = {{nb + j, 1}};
      } else if (k < 18) {
        uint64_t x = rng.below(ncols), y = rng.below(ncols);
        i128 c1 = (i128)(rng.next() & 0xffffffffu) - 0x7fffffff, c2 = (i128)(rng.next() & 0xffff) + 1;
        lc[0][r] = {{x, 1}, {ONE, c1}};
        lc[1][r] = {{y, 1}, {ONE, c2}};
        target(r, eval(lc[0][r]) * eval(lc[1][r]));
      } else {
        for (int side = 0; side < 2; side++) {
          int n = 1 + (int)rng.below(8);
          for (int q = 0; q < n; q++) {
            uint64_t sel = rng.below(4);
            i128 coef = sel == 0 ? 1 : sel == 1 ? -1 : sel == 2 ? 3 : (i128)1 << rng.below(16);
            lc[side][r].push_back({rng.below(ncols), coef});
          }
        }
        target(r, eval(lc[0][r]) * eval(lc[1][r]));
      }
    }
    for (int m = 0; m < 3; m++)
      for (auto& l : lc[m])
        for (auto& t : l) const_entries += t.col == ONE;
    vars.resize(32 * nv);
    for (uint64_t i = 0; i < nv; i++) enc(z[i], &vars[32 * i]);
    inputs.resize(32 * ni);
    for (uint64_t i = 0; i < ni; i++) enc(z[nv + 1 + i], &inputs[32 * i]);
  }

  void root(Buf& B, uint8_t msg_type, size_t& msg_at) {
    B.b.assign(8, 0);
    memcpy(&B.b[4], "zkif", 4);
    const int kinds[2] = {1, 0};
    size_t at[2];
    size_t t = B.table(2, kinds, at);
    B.b[at[0]] = msg_type;
    B.put32(0, (uint32_t)t);
    msg_at = at[1];
  }

  void write_files() {
    {  // ConstraintSystem (message type 2): constraints: [BilinearConstraint { a, b, c: Variables }]
      Buf B;
      size_t msg_at;
      root(B, 2, msg_at);
      const int k1[1] = {0};
      size_t at1[1];
      size_t cs = B.table(1, k1, at1);
      B.link(msg_at, cs);
      size_t v = B.vec_begin(4, (uint32_t)nc);
      B.b.insert(B.b.end(), 4 * nc, 0);
      B.link(at1[0], v);
      std::vector<uint64_t> ids;
      std::vector<uint8_t> vals;
      for (uint64_t r = 0; r < nc; r++) {
        const int k3[3] = {0, 0, 0};
        size_t at3[3];
        size_t t = B.table(3, k3, at3);
        B.link(v + 4 + 4 * r, t);
        for (int m = 0; m < 3; m++) {
          const auto& l = lc[m][r];
          ids.resize(l.size());
          vals.resize(32 * l.size());
          for (size_t q = 0; q < l.size(); q++) {
            ids[q] = var_id(l[q].col);
            enc(l[q].coef, &vals[32 * q]);
          }
          B.link(at3[m], write_variables(B, ids, vals));
        }
      }
      file[0].swap(B.b);
    }
    {  // CircuitHeader (1): instance_variables, free_variable_id, field_maximum
      Buf B;
      size_t msg_at;
      root(B, 1, msg_at);
      const int k3[3] = {0, 2, 0};
      size_t at3[3];
      size_t h = B.table(3, k3, at3);
      B.link(msg_at, h);
      std::vector<uint64_t> ids(ni);
      for (uint64_t i = 0; i < ni; i++) ids[i] = in_id(i);
      B.link(at3[0], write_variables(B, ids, inputs));
      uint64_t free_id = 0;
      memcpy(&B.b[at3[1]], &free_id, 8);
      size_t fm = B.vec_begin(1, 32);
      B.b.insert(B.b.end(), (const uint8_t*)ELL, (const uint8_t*)ELL + 32);
      B.link(at3[2], fm);
      file[1].swap(B.b);
    }
    {  // Witness (3): assigned_variables
      Buf B;
      size_t msg_at;
      root(B, 3, msg_at);
      const int k1[1] = {0};
      size_t at1[1];
      size_t w = B.table(1, k1, at1);
      B.link(msg_at, w);
      std::vector<uint64_t> ids(nv);
      for (uint64_t i = 0; i < nv; i++) ids[i] = wit_id(i);
      B.link(at1[0], write_variables(B, ids, vars));
      file[2].swap(B.b);
    }
  }
};

}  // namespace

extern "C" {
void* zg_new(uint64_t nc, uint64_t nv, uint64_t ni, uint64_t seed, int want_files) {
  if (nc == 0 || nv < 5 || ni >= 999) return nullptr;
  Gen* g = new Gen();
  g->nc = nc; g->nv = nv; g->ni = ni;
  g->generate(seed);
  if (want_files) g->write_files();
  return g;
}
// out = num_cons, num_vars, num_inputs, nnz(A), nnz(B), nnz(C), entries in the constant column
void zg_dims(void* h, uint64_t out[7]) {
  Gen* g = (Gen*)h;
  out[0] = g->nc; out[1] = g->nv; out[2] = g->ni;
  for (int m = 0; m < 3; m++) {
    uint64_t n = 0;
    for (auto& l : g->lc[m]) n += l.size();
    out[3 + m] = n;
  }
  out[6] = g->const_entries;
}
// COO of matrix m in the order R1cs::instance pushes entries (constraint by constraint)
void zg_coo(void* h, int m, uint64_t* rows, uint64_t* cols, uint8_t* vals32) {
  Gen* g = (Gen*)h;
  size_t k = 0;
  for (uint64_t r = 0; r < g->nc; r++)
    for (const Term& t : g->lc[m][r]) {
      rows[k] = r;
      cols[k] = t.col;
      enc(t.coef, vals32 + 32 * k);
      k++;
    }
}
const uint8_t* zg_vars(void* h) { return ((Gen*)h)->vars.data(); }
const uint8_t* zg_inputs(void* h) { return ((Gen*)h)->inputs.data(); }
// which = 0 circuit, 1 header (inputs), 2 witness -- the argument order of `spzk prove --nizk` (main.rs:18)
size_t zg_file(void* h, int which, const uint8_t** data) {
  Gen* g = (Gen*)h;
  *data = g->file[which].data();
  return g->file[which].size();
}
void zg_free(void* h) { delete (Gen*)h; }
}
