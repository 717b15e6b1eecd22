// zkinterface ingestion: three flatbuffers (circuit header, constraint system, witness) -> the COO triples and
// assignments libspartan's Instance::new / Assignment::new take.  Replaces spartan-zkinterface's R1csReader / R1cs
// (spartan-zkinterface/src/lib.rs:75-260, schema/zkinterface.fbs) for the prove path.
//
// No flatc here, so the reader walks the wire format directly (root offset -> table -> vtable -> fields), with every
// offset bounds-checked: the files are untrusted input and nothing may crash or read out of bounds.  Where the
// reference `unwrap()`s or panics (missing field, no constraints, value wider than 32 bytes) parse() fails with a
// message instead.  Variable-id translation uses a direct table or hash maps (first occurrence wins, as `position` does in
// lib.rs:96-107) instead of the reference's linear scan per matrix entry, which is O(nnz * n); the constraints are walked
// by several threads (they are independent tables), the output keeps file order.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace sbh {

class FbView {  // bounds-checked little-endian reads over one buffer
 public:
  const uint8_t* p;
  size_t n;
  bool ok = true;
  FbView(const uint8_t* p_, size_t n_) : p(p_), n(n_) {}
  bool in(size_t off, size_t len) const { return off <= n && len <= n - off; }
  uint16_t u16(size_t off) { if (!in(off, 2)) { ok = false; return 0; } uint16_t v; memcpy(&v, p + off, 2); return v; }
  uint32_t u32(size_t off) { if (!in(off, 4)) { ok = false; return 0; } uint32_t v; memcpy(&v, p + off, 4); return v; }
  int32_t i32(size_t off) { return (int32_t)u32(off); }
  uint64_t u64(size_t off) { if (!in(off, 8)) { ok = false; return 0; } uint64_t v; memcpy(&v, p + off, 8); return v; }
  // position of field `id` inside the table at `t`, or 0 when the field is absent
  size_t field(size_t t, unsigned id) {
    int64_t vt = (int64_t)t - (int64_t)i32(t);
    if (!ok || vt < 0 || !in((size_t)vt, 4)) { ok = false; return 0; }
    uint16_t vsize = u16((size_t)vt);
    size_t slot = 4 + 2 * (size_t)id;
    if (slot + 2 > vsize) return 0;
    uint16_t off = u16((size_t)vt + slot);
    if (!ok || off == 0) return 0;
    if (!in(t + off, 1)) { ok = false; return 0; }
    return t + off;
  }
  // follow the uoffset stored at `at` (tables, vectors, strings)
  size_t indirect(size_t at) {
    uint32_t o = u32(at);
    if (!ok || !in(at + o, 4)) { ok = false; return 0; }
    return at + o;
  }
  // vector at `v`: element count and start of the elements
  bool vec(size_t v, size_t elem, size_t& count, size_t& data) {
    count = u32(v);
    data = v + 4;
    if (!ok || count > (n - data) / (elem ? elem : 1) || !in(data, count * elem)) { ok = false; return false; }
    return true;
  }
};

// vector storage that resize() leaves uninitialised: the reader's threads fill every element (350 MB for a 2^20 circuit,
// whose zero-fill by one thread would cost as much as the parse itself)
template <typename T>
struct NoInitAlloc : std::allocator<T> {
  template <typename U> struct rebind { typedef NoInitAlloc<U> other; };
  NoInitAlloc() = default;
  template <typename U> NoInitAlloc(const NoInitAlloc<U>&) {}
  template <typename U> void construct(U* p) { ::new ((void*)p) U; }
  template <typename U, typename... A> void construct(U* p, A&&... a) { ::new ((void*)p) U(std::forward<A>(a)...); }
};

struct ZkVar {
  uint64_t id;
  uint8_t value[32];
};

struct ZkR1cs {
  // COO triples in the reference's order (constraint by constraint, a then b then c entries in file order)
  std::vector<uint64_t, NoInitAlloc<uint64_t>> rows[3], cols[3];
  std::vector<uint8_t, NoInitAlloc<uint8_t>> vals[3];
  std::vector<uint8_t> vars, inputs;  // n x 32 canonical bytes
  uint64_t num_cons = 0, num_vars = 0, num_inputs = 0, non_zero_entries = 0;
  uint8_t field_max[32];
  std::string error;

  static const unsigned MSG_HEADER = 1, MSG_CS = 2, MSG_WITNESS = 3;  // union Message, zkinterface.fbs

  // root_type Root { message: Message }: returns the table of the expected union member
  static size_t root_message(FbView& b, unsigned want, std::string& err, const char* what) {
    if (b.n < 8) { err = std::string(what) + ": buffer too short"; return 0; }
    size_t root = b.indirect(0);
    size_t ftype = b.ok ? b.field(root, 0) : 0, fmsg = b.ok ? b.field(root, 1) : 0;
    if (!b.ok || !ftype || !fmsg || b.p[ftype] != want) {
      err = std::string("Input file is not a flatbuffer ") + what;  // lib.rs:166-185
      return 0;
    }
    size_t t = b.indirect(fmsg);
    if (!b.ok) { err = std::string(what) + ": malformed"; return 0; }
    return t;
  }
  // get_variables, lib.rs:194-219: ids and values are required; each value is values.len()/ids.len() bytes, zero-extended
  static bool variables(FbView& b, size_t field_pos, std::vector<ZkVar>& out, std::string& err) {
    out.clear();
    if (!field_pos) { err = "missing Variables table"; return false; }
    size_t t = b.indirect(field_pos);
    size_t fid = b.ok ? b.field(t, 0) : 0, fval = b.ok ? b.field(t, 1) : 0;
    if (!b.ok || !fid || !fval) { err = "Variables without variable_ids / values"; return false; }
    size_t nid, idp, nval, valp;
    if (!b.vec(b.indirect(fid), 8, nid, idp) || !b.vec(b.indirect(fval), 1, nval, valp)) { err = "malformed Variables vectors"; return false; }
    if (nid == 0) return true;
    size_t ba = nval / nid;
    if (ba > 32) { err = "variable value wider than 32 bytes"; return false; }
    out.resize(nid);
    for (size_t i = 0; i < nid; i++) {
      out[i].id = b.u64(idp + 8 * i);
      memset(out[i].value, 0, 32);
      memcpy(out[i].value, b.p + valp + i * ba, ba);
    }
    return b.ok;
  }

  bool parse(const uint8_t* cs, size_t cs_len, const uint8_t* hd, size_t hd_len, const uint8_t* wt, size_t wt_len, unsigned threads = 1) {
    FbView bh(hd, hd_len), bc(cs, cs_len), bw(wt, wt_len);
    size_t th = root_message(bh, MSG_HEADER, error, "Circuit Header");
    if (!th) return false;
    size_t tc = root_message(bc, MSG_CS, error, "Constraint System");
    if (!tc) return false;
    size_t tw = root_message(bw, MSG_WITNESS, error, "Witness");
    if (!tw) return false;
    std::vector<ZkVar> in, wit;
    if (!variables(bh, bh.field(th, 0), in, error)) return false;   // header.instance_variables
    {
      size_t ff = bh.field(th, 2), cnt, dp;                          // header.field_maximum, lib.rs:222-223
      if (!ff || !bh.vec(bh.indirect(ff), 1, cnt, dp) || cnt != 32) { error = "field_maximum must be 32 bytes"; return false; }
      memcpy(field_max, bh.p + dp, 32);
    }
    if (!variables(bw, bw.field(tw, 0), wit, error)) return false;  // witness.assigned_variables
    num_inputs = in.size();
    num_vars = wit.size();
    // translate, lib.rs:96-107: witness position, else inputs position + num_vars + 1, else the constant column num_vars.
    // CirC numbers its variables densely, so a direct table serves unless the ids are sparse (then hash maps);
    // either way the first occurrence of a duplicated id wins, as `position` does.
    static const uint64_t kNone = ~0ull;
    uint64_t max_id = 0;
    for (const ZkVar& v : wit) max_id = std::max(max_id, v.id);
    for (const ZkVar& v : in) max_id = std::max(max_id, v.id);
    const bool dense = max_id < 4 * (uint64_t)(wit.size() + in.size()) + 1024;
    std::vector<uint64_t> table;
    std::unordered_map<uint64_t, uint64_t> wpos, ipos;
    if (dense) {
      table.assign(max_id + 1, kNone);
      for (size_t i = 0; i < wit.size(); i++) if (table[wit[i].id] == kNone) table[wit[i].id] = i;
      for (size_t i = 0; i < in.size(); i++) if (table[in[i].id] == kNone) table[in[i].id] = i + num_vars + 1;
    } else {
      wpos.reserve(wit.size() * 2);
      for (size_t i = 0; i < wit.size(); i++) wpos.emplace(wit[i].id, i);  // emplace keeps the first occurrence
      for (size_t i = 0; i < in.size(); i++) ipos.emplace(in[i].id, i);
    }
    auto translate = [&](uint64_t id) -> uint64_t {
      if (dense) return id <= max_id && table[id] != kNone ? table[id] : num_vars;
      auto a = wpos.find(id);
      if (a != wpos.end()) return a->second;
      auto c = ipos.find(id);
      if (c != ipos.end()) return c->second + num_vars + 1;
      return num_vars;
    };
    size_t fc = bc.field(tc, 0), ncons = 0, cp = 0;
    if (!fc || !bc.vec(bc.indirect(fc), 4, ncons, cp)) { error = "ConstraintSystem without constraints"; return false; }
    if (ncons == 0) { error = "No constraints given!"; return false; }  // lib.rs:235-237
    num_cons = ncons;
    // Constraints are independent tables: slices of them are walked by `threads` threads, first to count the entries of
    // each matrix (so that every slice knows where its output starts), then to fill.  Output order = file order.
    const unsigned T = ncons < 4096 || threads < 2 ? 1 : threads;
    std::vector<uint64_t> cnt(3 * (size_t)T, 0), nzs(3 * (size_t)T, 0);
    std::vector<std::string> errs(T);
    auto run = [&](auto&& f) {
      std::vector<std::thread> th;
      for (unsigned t = 1; t < T; t++) th.emplace_back([&f, t] { f(t); });
      f(0);
      for (auto& x : th) x.join();
    };
    run([&](unsigned t) {  // pass 1: number of entries per matrix in this slice (the length of each variable_ids vector)
      FbView b(cs, cs_len);
      for (size_t i = ncons * t / T; i < ncons * (t + 1) / T; i++) {
        size_t tb = b.indirect(cp + 4 * i);
        if (!b.ok) { errs[t] = "malformed constraint"; return; }
        for (unsigned m = 0; m < 3; m++) {
          size_t fp = b.field(tb, m);
          if (!fp) { errs[t] = "missing Variables table"; return; }
          size_t vt = b.indirect(fp);
          size_t fid = b.ok ? b.field(vt, 0) : 0, nid = 0, idp = 0;
          if (!b.ok || !fid || !b.vec(b.indirect(fid), 8, nid, idp)) { errs[t] = "Variables without variable_ids / values"; return; }
          cnt[3 * (size_t)t + m] += nid;
        }
      }
    });
    for (unsigned t = 0; t < T; t++) if (!errs[t].empty()) { error = errs[t]; return false; }
    std::vector<uint64_t> start(3 * (size_t)T, 0);
    for (unsigned m = 0; m < 3; m++) {
      uint64_t o = 0;
      for (unsigned t = 0; t < T; t++) { start[3 * (size_t)t + m] = o; o += cnt[3 * (size_t)t + m]; }
      rows[m].resize(o);
      cols[m].resize(o);
      vals[m].resize(32 * o);
    }
    run([&](unsigned t) {  // pass 2: fill
      FbView b(cs, cs_len);
      std::vector<ZkVar> lc;
      uint64_t at[3] = {start[3 * (size_t)t], start[3 * (size_t)t + 1], start[3 * (size_t)t + 2]};
      for (size_t i = ncons * t / T; i < ncons * (t + 1) / T; i++) {
        size_t tb = b.indirect(cp + 4 * i);
        for (unsigned m = 0; m < 3; m++) {
          if (!variables(b, b.field(tb, m), lc, errs[t])) return;
          if (at[m] + lc.size() > start[3 * (size_t)t + m] + cnt[3 * (size_t)t + m]) { errs[t] = "malformed constraint"; return; }
          for (const ZkVar& v : lc) {
            rows[m][at[m]] = i;
            cols[m][at[m]] = translate(v.id);
            memcpy(&vals[m][32 * at[m]], v.value, 32);
            bool nonzero = false;
            for (int k = 0; k < 32; k++) nonzero |= v.value[k] != 0;
            nzs[3 * (size_t)t + m] += nonzero;
            at[m]++;
          }
        }
      }
      if (!b.ok) errs[t] = "malformed flatbuffer";
    });
    for (unsigned t = 0; t < T; t++) if (!errs[t].empty()) { error = errs[t]; return false; }
    uint64_t nz[3] = {0, 0, 0};
    for (unsigned t = 0; t < T; t++) for (unsigned m = 0; m < 3; m++) nz[m] += nzs[3 * (size_t)t + m];
    non_zero_entries = std::max(nz[0], std::max(nz[1], nz[2]));  // lib.rs:250
    vars.resize(32 * wit.size());
    for (size_t i = 0; i < wit.size(); i++) memcpy(&vars[32 * i], wit[i].value, 32);
    inputs.resize(32 * in.size());
    for (size_t i = 0; i < in.size(); i++) memcpy(&inputs[32 * i], in[i].value, 32);
    if (!bh.ok || !bc.ok || !bw.ok) { error = "malformed flatbuffer"; return false; }
    return true;
  }
};

}  // namespace sbh
