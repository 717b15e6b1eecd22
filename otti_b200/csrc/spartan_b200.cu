// libspartan_b200: C-ABI entry points (include/spartan_b200.h) and the host
// driver of the B200 Spartan NIZK prover.  The host keeps the Fiat-Shamir
// transcript, the RandomTape and the constant-size sigma protocols; every loop
// over the witness, the constraint hypercube or the generators runs in the
// sm_100a kernels of kernels.cu.  Protocol order follows
// Spartan/src/lib.rs:493-535 and r1csproof.rs:141-345 statement by statement
// (transcript framing SURVEY.md App. A.1, tape draw order App. A.2).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>

#include "../../include/spartan_b200.h"
#include "../../include/spartan_b200_kernels.h"
#include "host/objects.hpp"
#include "host/proof.hpp"

namespace sbh {

static thread_local std::string g_err;
void set_error(const std::string& s) { g_err = s; }
const char* get_error() { return g_err.c_str(); }
static thread_local double g_profile[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
static thread_local uint64_t g_h2d_bytes = 0, g_d2h_bytes = 0;

static bool g_device_ok = false;
static int ensure_device() {
  if (g_device_ok) return SB200_OK;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    set_error("no CUDA device: libspartan_b200 has no CPU fallback");
    return SB200_ERR_NO_DEVICE;
  }
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, dev) != cudaSuccess || p.major != 10) {
    set_error("device is not compute capability 10.x (sm_100a kernels only)");
    return SB200_ERR_NO_DEVICE;
  }
  g_device_ok = true;
  return SB200_OK;
}

static inline size_t ilog2(size_t n) {
  size_t l = 0;
  while (((size_t)1 << (l + 1)) <= n) l++;
  return l;
}
static inline double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
static inline sb::fl_t to_dev(const Fl& x) {
  sb::fl_t r;
  memcpy(r.v, x.v, 32);
  return r;
}
static inline Fl from_dev_bytes(const uint8_t* p) {
  Fl r;
  memcpy(r.v, p, 32);
  return r;
}
static inline Cpt cpt_from(const uint8_t* p) {
  Cpt c;
  memcpy(c.data(), p, 32);
  return c;
}

// ------------------------------------------------------------------------------ instance
struct CooEntry {
  uint32_t row, col;
  Fl val;  // Montgomery
};

static int upload_sparse(const std::vector<CooEntry> M[3], uint64_t nrows, uint64_t ncols, Instance& inst) {
  for (int orient = 0; orient < 2; orient++) {
    SparseDev& sd = orient == 0 ? inst.csr : inst.csc;
    uint64_t nseg = orient == 0 ? nrows : ncols;
    for (int m = 0; m < 3; m++) {
      size_t nnz = M[m].size();
      std::vector<uint32_t> ptr(nseg + 1, 0), idx(nnz);
      std::vector<Fl> val(nnz);
      for (const CooEntry& e : M[m]) ptr[(orient == 0 ? e.row : e.col) + 1]++;
      for (uint64_t i = 0; i < nseg; i++) ptr[i + 1] += ptr[i];
      std::vector<uint32_t> fill(ptr.begin(), ptr.end() - 1);
      for (const CooEntry& e : M[m]) {
        uint32_t k = fill[orient == 0 ? e.row : e.col]++;
        idx[k] = orient == 0 ? e.col : e.row;
        val[k] = e.val;
      }
      SB_CUDA(sd.ptr[m].alloc(nseg + 1));
      SB_CUDA(sd.idx[m].alloc(nnz ? nnz : 1));
      SB_CUDA(sd.val[m].alloc(nnz ? nnz : 1));
      SB_CUDA(cudaMemcpy(sd.ptr[m].p, ptr.data(), (nseg + 1) * 4, cudaMemcpyHostToDevice));
      if (nnz) {
        SB_CUDA(cudaMemcpy(sd.idx[m].p, idx.data(), nnz * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(sd.val[m].p, val.data(), nnz * 32, cudaMemcpyHostToDevice));
      }
    }
  }
  return SB200_OK;
}

// padded sizes, lib.rs:130-158
static void padded_dims(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t& ncp, uint64_t& nvp) {
  nvp = std::max(num_vars, num_inputs + 1);
  uint64_t p2 = 1;
  while (p2 < nvp) p2 <<= 1;
  nvp = p2;
  ncp = num_cons;
  if (ncp == 0 || ncp == 1) ncp = 2;
  uint64_t npt = 1;
  while (npt < num_cons) npt <<= 1;  // usize::next_power_of_two; 0 -> 1
  if (npt != num_cons) ncp = npt;
}

}  // namespace sbh

using namespace sbh;

// =============================================================================== C ABI
extern "C" {

const char* sb200_last_error(void) { return get_error(); }
void sb200_free(void* p) { free(p); }

int sb200_init(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    set_error("no CUDA device: libspartan_b200 has no CPU fallback");
    return SB200_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= n) {
    set_error("device index out of range");
    return SB200_ERR_ARG;
  }
  SB_CUDA(cudaSetDevice(device));
  g_device_ok = false;
  return ensure_device();
}

int sb200_instance_new(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, const uint64_t* const rows[3],
                       const uint64_t* const cols[3], const uint8_t* const vals32[3], const uint64_t nnz[3],
                       sb200_instance** out) {
  if (!out) return SB200_ERR_ARG;
  *out = nullptr;
  uint64_t ncp, nvp;
  padded_dims(num_cons, num_vars, num_inputs, ncp, nvp);
  if (ncp >= (1ull << 31) || nvp >= (1ull << 30)) {
    set_error("instance too large for 32-bit indices");
    return SB200_ERR_ARG;
  }
  std::vector<CooEntry> M[3];
  for (int m = 0; m < 3; m++) {
    if (nnz[m] >= (1ull << 32)) return SB200_ERR_ARG;
    M[m].reserve(nnz[m]);
    for (uint64_t k = 0; k < nnz[m]; k++) {
      uint64_t row = rows[m][k], col = cols[m][k];
      if (row >= num_cons) { set_error("InvalidIndex: row"); return SB200_ERR_INVALID_INDEX; }                      // lib.rs:166-168
      if (col >= num_vars + 1 + num_inputs) { set_error("InvalidIndex: col"); return SB200_ERR_INVALID_INDEX; }     // lib.rs:171-173
      Fl v;
      if (!fl_from_bytes(vals32[m] + 32 * k, v)) { set_error("InvalidScalar"); return SB200_ERR_INVALID_SCALAR; }   // lib.rs:175-187
      if (col >= num_vars) col += nvp - num_vars;  // 1 and the inputs live in the second half of z, lib.rs:180-181
      M[m].push_back(CooEntry{(uint32_t)row, (uint32_t)col, v});
    }
    // lib.rs:191-197 appends zero-valued entries when num_cons is 0 or 1; they contribute nothing to any product.
  }
  int rc = ensure_device();
  if (rc) return rc;
  Instance* inst = new Instance();
  inst->num_cons = ncp; inst->num_vars = nvp; inst->num_inputs = num_inputs;
  for (int m = 0; m < 3; m++) inst->nnz[m] = M[m].size();
  rc = upload_sparse(M, ncp, 2 * nvp, *inst);
  if (rc) { delete inst; return rc; }
  *out = reinterpret_cast<sb200_instance*>(inst);
  return SB200_OK;
}

static inline uint64_t splitmix64(uint64_t& s) {
  s += 0x9E3779B97F4A7C15ULL;
  uint64_t z = s;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

// R1CSInstance::produce_synthetic_r1cs, r1csinstance.rs:136-214
int sb200_instance_synthetic(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t seed,
                             sb200_instance** out, uint8_t* vars32_out, uint8_t* inputs32_out) {
  if (!out) return SB200_ERR_ARG;
  *out = nullptr;
  auto pow2 = [](uint64_t x) { return x && !(x & (x - 1)); };
  if (!pow2(num_cons)) { set_error("NonPowerOfTwoCons"); return SB200_ERR_NON_POW2_CONS; }   // :146
  if (!pow2(num_vars)) { set_error("NonPowerOfTwoVars"); return SB200_ERR_NON_POW2_VARS; }   // :147
  if (!(num_inputs < num_vars)) { set_error("assert num_inputs < num_vars"); return SB200_ERR_ASSERT; }  // :150
  if (num_cons >= (1ull << 31) || num_vars >= (1ull << 30)) return SB200_ERR_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  const uint64_t size_z = num_vars + num_inputs + 1;
  FlVec Z(size_z);
  uint64_t st = seed;
  for (uint64_t i = 0; i < size_z; i++) {  // Scalar::random = from_u512(8 x next_u64), ristretto255.rs:381-387
    uint8_t wide[64];
    for (int k = 0; k < 8; k++) {
      uint64_t w = splitmix64(st);
      memcpy(wide + 8 * k, &w, 8);
    }
    Z[i] = fl_from_bytes_wide(wide);
  }
  Z[num_vars] = fl_one();
  // Z^-1 for the C coefficients with one shared inversion
  FlVec pre(size_z), Zinv(size_z);
  Fl acc = fl_one();
  for (uint64_t i = 0; i < size_z; i++) {
    pre[i] = acc;
    if (!Z[i].is_zero()) acc = fl_mul(acc, Z[i]);
  }
  acc = fl_invert(acc);
  for (uint64_t i = size_z; i-- > 0;) {
    if (Z[i].is_zero()) { Zinv[i] = fl_zero(); continue; }
    Zinv[i] = fl_mul(acc, pre[i]);
    acc = fl_mul(acc, Z[i]);
  }
  std::vector<CooEntry> M[3];
  for (int m = 0; m < 3; m++) M[m].reserve(num_cons);
  for (uint64_t i = 0; i < num_cons; i++) {
    uint64_t ai = i % size_z, bi = (i + 2) % size_z, ci = (i + 3) % size_z;
    M[0].push_back(CooEntry{(uint32_t)i, (uint32_t)ai, fl_one()});
    M[1].push_back(CooEntry{(uint32_t)i, (uint32_t)bi, fl_one()});
    Fl ab = fl_mul(Z[ai], Z[bi]);
    if (Z[ci].is_zero()) M[2].push_back(CooEntry{(uint32_t)i, (uint32_t)num_vars, ab});
    else M[2].push_back(CooEntry{(uint32_t)i, (uint32_t)ci, fl_mul(ab, Zinv[ci])});
  }
  Instance* inst = new Instance();
  inst->num_cons = num_cons; inst->num_vars = num_vars; inst->num_inputs = num_inputs;
  for (int m = 0; m < 3; m++) inst->nnz[m] = M[m].size();
  rc = upload_sparse(M, num_cons, 2 * num_vars, *inst);
  if (rc) { delete inst; return rc; }
  for (uint64_t i = 0; i < num_vars; i++) fl_to_bytes(Z[i], vars32_out + 32 * i);
  for (uint64_t i = 0; i < num_inputs; i++) fl_to_bytes(Z[num_vars + 1 + i], inputs32_out + 32 * i);
  *out = reinterpret_cast<sb200_instance*>(inst);
  return SB200_OK;
}

int sb200_assignment_check(const uint8_t* vals32, uint64_t n) {
  for (uint64_t i = 0; i < n; i++) {
    Fl v;
    memcpy(v.v, vals32 + 32 * i, 32);
    u128 bo = 0;
    for (int k = 0; k < 4; k++) {
      u128 d = (u128)v.v[k] - flc::L[k] - (uint64_t)bo;
      bo = (d >> 64) & 1;
    }
    if (!bo) { set_error("InvalidScalar"); return SB200_ERR_INVALID_SCALAR; }
  }
  return SB200_OK;
}

struct sb200_transcript { MerlinTranscript t; sb200_transcript(const void* l, size_t n) : t(l, n) {} };
int sb200_transcript_new(const uint8_t* label, size_t label_len, sb200_transcript** out) {
  if (!out) return SB200_ERR_ARG;
  *out = new sb200_transcript(label, label_len);
  return SB200_OK;
}
void sb200_transcript_append_message(sb200_transcript* t, const uint8_t* label, size_t ll, const uint8_t* msg, size_t ml) {
  t->t.append_message_n(label, ll, msg, ml);
}
void sb200_transcript_challenge_bytes(sb200_transcript* t, const uint8_t* label, size_t ll, uint8_t* out, size_t n) {
  t->t.challenge_bytes_n(label, ll, out, n);
}
static void vt_append(void* ctx, const uint8_t* label, size_t ll, const uint8_t* msg, size_t ml) {
  ((sb200_transcript*)ctx)->t.append_message_n(label, ll, msg, ml);
}
static void vt_challenge(void* ctx, const uint8_t* label, size_t ll, uint8_t* out, size_t n) {
  ((sb200_transcript*)ctx)->t.challenge_bytes_n(label, ll, out, n);
}
void sb200_transcript_vtable(sb200_transcript* t, sb200_transcript_vt* out) {
  out->ctx = t;
  out->append_message = vt_append;
  out->challenge_bytes = vt_challenge;
}
void sb200_transcript_free(sb200_transcript* t) { delete t; }

// Instance::is_sat, lib.rs:228-257 -> R1CSInstance::is_sat, r1csinstance.rs:216-246 (on the device)
int sb200_instance_is_sat(const sb200_instance* h, const uint8_t* vars32, uint64_t n_vars, const uint8_t* inputs32,
                          uint64_t n_inputs, int* sat_out) {
  if (!h || !sat_out) return SB200_ERR_ARG;
  const Instance& inst = *reinterpret_cast<const Instance*>(h);
  if (n_vars > inst.num_vars) { set_error("InvalidNumberOfInputs"); return SB200_ERR_INVALID_NUMBER_OF_INPUTS; }      // lib.rs:233-235
  if (n_inputs != inst.num_inputs) { set_error("InvalidNumberOfInputs"); return SB200_ERR_INVALID_NUMBER_OF_INPUTS; }  // lib.rs:237-239
  int rc = ensure_device();
  if (rc) return rc;
  const size_t Nv = inst.num_vars, Nc = inst.num_cons;
  DevBuf<sb::fl_t> zc, z, y;
  DevBuf<uint32_t> flag;
  SB_CUDA(zc.alloc(2 * Nv)); SB_CUDA(z.alloc(2 * Nv)); SB_CUDA(y.alloc(3 * Nc)); SB_CUDA(flag.alloc(2));
  SB_CUDA(cudaMemset(zc.p, 0, 64 * Nv));
  SB_CUDA(cudaMemset(flag.p, 0, 8));
  if (n_vars) SB_CUDA(cudaMemcpy(zc.p, vars32, 32 * n_vars, cudaMemcpyHostToDevice));
  uint8_t one[32] = {1};
  SB_CUDA(cudaMemcpy(zc.p + Nv, one, 32, cudaMemcpyHostToDevice));
  if (n_inputs) SB_CUDA(cudaMemcpy(zc.p + Nv + 1, inputs32, 32 * n_inputs, cudaMemcpyHostToDevice));
  SB_CUDA(sb::k_to_mont(zc.p, z.p, 2 * Nv, flag.p, 0));
  SB_CUDA(sb::k_spmv3(inst.csr.view(), (uint32_t)Nc, z.p, y.p, y.p + Nc, y.p + 2 * Nc, 0));
  SB_CUDA(sb::k_check_sat(y.p, y.p + Nc, y.p + 2 * Nc, Nc, flag.p + 1, 0));
  uint32_t f[2];
  SB_CUDA(cudaMemcpy(f, flag.p, 8, cudaMemcpyDeviceToHost));
  if (f[0]) { set_error("InvalidScalar"); return SB200_ERR_INVALID_SCALAR; }
  *sat_out = f[1] ? 0 : 1;
  return SB200_OK;
}

int sb200_instance_dims(const sb200_instance* h, uint64_t out[3]) {
  if (!h) return SB200_ERR_ARG;
  const Instance* inst = reinterpret_cast<const Instance*>(h);
  out[0] = inst->num_cons; out[1] = inst->num_vars; out[2] = inst->num_inputs;
  return SB200_OK;
}
void sb200_instance_free(sb200_instance* h) { delete reinterpret_cast<Instance*>(h); }

// ------------------------------------------------------------------------------ generators
int sb200_gens_new(const uint8_t* label, size_t label_len, uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs,
                   sb200_gens** out) {
  (void)num_cons;
  if (!out) return SB200_ERR_ARG;
  *out = nullptr;
  int rc = ensure_device();
  if (rc) return rc;
  static const char kDefault[] = "gens_r1cs_sat";  // lib.rs:475
  if (!label) { label = (const uint8_t*)kDefault; label_len = sizeof(kDefault) - 1; }
  uint64_t nvp = std::max(num_vars, num_inputs + 1), p2 = 1;  // lib.rs:467-473
  while (p2 < nvp) p2 <<= 1;
  nvp = p2;
  size_t ell = ilog2(nvp), right = ell - ell / 2;  // dense_mlpoly.rs:30-32, 85-87
  uint32_t R = 1u << right;
  Gens* g = new Gens();
  g->num_vars_padded = nvp;
  g->R = R;
  // MultiCommitGens::new(R + 1, label): R + 2 points from SHAKE256(label || basepoint), commitments.rs:15-33;
  // split_at(R) -> gens_n = (G[0..R), h), gens_1 = (G[R], h)            nizk/mod.rs:409-414
  // gens_3 / gens_4 re-read the same stream: (G[0..3), G[3]) and (G[0..4), G[4])  r1csproof.rs:45-56
  static const uint8_t kBase[32] = {0xe2, 0xf2, 0xae, 0x0a, 0x6a, 0xbc, 0x4e, 0x71, 0xa8, 0x84, 0xa9, 0x61, 0xc5, 0x00, 0x51, 0x5f,
                                    0x58, 0xe3, 0x0b, 0x6a, 0xa5, 0x82, 0xdd, 0x8d, 0xb6, 0xa6, 0x59, 0x45, 0xe0, 0x8d, 0x2d, 0x76};
  size_t npts = std::max<size_t>((size_t)R + 2, 5);
  std::vector<Ge> pts(npts);
  {
    Shake256Xof xof;
    xof.update(label, label_len);
    xof.update(kBase, 32);  // GROUP_BASEPOINT_COMPRESSED, group.rs:23-24
    for (size_t i = 0; i < npts; i++) {
      uint8_t ub[64];
      xof.read(ub, 64);
      pts[i] = ge_from_uniform_bytes(ub);
    }
  }
  g->G.assign(pts.begin(), pts.begin() + R);
  g->Q = pts[R];
  g->h = pts[R + 1];
  for (int i = 0; i < 4; i++) g->G4[i] = pts[i];
  g->h4 = pts[4];
  for (int i = 0; i < 4; i++) g->fb_G4[i].build(g->G4[i]);
  g->fb_h4.build(g->h4);
  g->fb_Q.build(g->Q);
  g->fb_h.build(g->h);
  // device tables over [G..., h, Q]: affine coordinates with one shared inversion
  std::vector<Ge> bases(g->G);
  bases.push_back(g->h);
  bases.push_back(g->Q);
  std::vector<uint8_t> xy(bases.size() * 64);
  {
    std::vector<Fe> pre(bases.size());
    Fe acc = fe_one();
    for (size_t i = 0; i < bases.size(); i++) { pre[i] = acc; acc = fe_mul(acc, bases[i].Z); }
    acc = fe_invert(acc);
    for (size_t i = bases.size(); i-- > 0;) {
      Fe zi = fe_mul(acc, pre[i]);
      acc = fe_mul(acc, bases[i].Z);
      fe_tobytes(fe_mul(bases[i].X, zi), &xy[64 * i]);
      fe_tobytes(fe_mul(bases[i].Y, zi), &xy[64 * i + 32]);
    }
  }
  uint32_t c = 8;
  if (const char* env = getenv("SB200_MSM_WINDOW")) {
    int v = atoi(env);
    if (v >= 2 && v <= 16) c = (uint32_t)v;
  }
  cudaError_t ce = sb::msm_table_build(g->table, xy.data(), (uint32_t)bases.size(), c, 0);
  if (ce != cudaSuccess) {
    set_error(std::string("msm_table_build: ") + cudaGetErrorString(ce));
    delete g;
    return SB200_ERR_CUDA;
  }
  *out = reinterpret_cast<sb200_gens*>(g);
  return SB200_OK;
}
void sb200_gens_free(sb200_gens* h) { delete reinterpret_cast<Gens*>(h); }

}  // extern "C"

// =============================================================================== prover
namespace sbh {

struct PinnedLayout {
  static constexpr size_t kResult = 0;      // 256 B of kernel results
  static constexpr size_t kUpload = 1024;   // 4 KB of small uploads
  static constexpr size_t kPoints = 8192;   // compressed points (L x 32)
};

static int workspace_init(const Instance& inst, uint32_t R) {
  Workspace& w = inst.ws;
  if (w.ready) return SB200_OK;
  const size_t Nc = inst.num_cons, Nv = inst.num_vars;
  const size_t ell = ilog2(Nv), Ls = (size_t)1 << (ell / 2);
  SB_CUDA(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
  for (int i = 0; i < 4; i++) SB_CUDA(cudaEventCreate(&w.ev[i]));
  SB_CUDA(w.vars_canon.alloc(Nv));
  SB_CUDA(w.varsM.alloc(Nv));
  SB_CUDA(w.z.alloc(2 * Nv));
  SB_CUDA(w.abc.alloc(2 * Nv));
  SB_CUDA(w.T.alloc(Nc));
  SB_CUDA(w.A.alloc(Nc));
  SB_CUDA(w.B.alloc(Nc));
  SB_CUDA(w.C.alloc(Nc));
  size_t big = std::max(Nc, 2 * Nv);
  SB_CUDA(w.eq_tmp.alloc(2 * ((size_t)1 << ((ilog2(big) + 1) / 2)) + 256));
  const uint32_t max_blocks = 148 * 4;
  SB_CUDA(w.sc_partials.alloc((size_t)max_blocks * 3));
  SB_CUDA(w.counter.alloc(1));
  SB_CUDA(cudaMemset(w.counter.p, 0, sizeof(unsigned int)));
  SB_CUDA(w.flag.alloc(1));
  SB_CUDA(w.small.alloc(256));
  SB_CUDA(w.blinds.alloc(2 * Ls));  // canonical | Montgomery
  SB_CUDA(w.bound_tmp.alloc(sb::bound_L_tmp_scalars((uint32_t)Ls, R) + 2 * (size_t)R + 2 * Ls));
  SB_CUDA(w.LZ.alloc(R));
  SB_CUDA(w.bl_a.alloc(R));
  SB_CUDA(w.bl_b.alloc(R));
  SB_CUDA(w.coeff0.alloc(R));
  SB_CUDA(w.coeff1.alloc(R));
  SB_CUDA(w.srows.alloc(2 * ((size_t)R + 2)));
  size_t scratch = std::max(sb::msm_scratch_bytes((uint32_t)Ls, R), sb::msm_scratch_bytes(2, R + 2));
  SB_CUDA(w.msm_scratch.alloc(scratch));
  SB_CUDA(w.pts.alloc(std::max<size_t>(Ls, 2) * 32));
  w.pinned_bytes = PinnedLayout::kPoints + std::max<size_t>(Ls, 2) * 32;
  SB_CUDA(cudaMallocHost((void**)&w.pinned, w.pinned_bytes));
  w.ready = true;
  return SB200_OK;
}

// Pedersen commitments of the Fiat-Shamir tail, on the host with fixed-base tables.
struct HostCommit {
  const Gens& g;
  // impl Commitments for Scalar: x*G1[0] + blind*h            commitments.rs:65-70
  Ge scalar_pt(const Fl& x, const Fl& blind) const {
    uint8_t k[32];
    Ge acc = ge_identity();
    fl_to_bytes(x, k); g.fb_Q.mul_acc(acc, k);
    fl_to_bytes(blind, k); g.fb_h.mul_acc(acc, k);
    return acc;
  }
  Cpt scalar(const Fl& x, const Fl& blind) const {
    Cpt c;
    ge_compress(scalar_pt(x, blind), c.data());
    return c;
  }
  // impl Commitments for [Scalar] over gens_3 (n = 3) or gens_4 (n = 4)   commitments.rs:79-84
  Cpt poly(const FlVec& v, const Fl& blind) const {
    uint8_t k[32];
    Ge acc = ge_identity();
    for (size_t i = 0; i < v.size(); i++) { fl_to_bytes(v[i], k); g.fb_G4[i].mul_acc(acc, k); }
    fl_to_bytes(blind, k);
    if (v.size() == 3) g.fb_G4[3].mul_acc(acc, k);  // gens_3.h is the 4th point of the stream
    else g.fb_h4.mul_acc(acc, k);
    Cpt c;
    ge_compress(acc, c.data());
    return c;
  }
  Cpt h_times(const Fl& r) const {
    uint8_t k[32];
    Ge acc = ge_identity();
    fl_to_bytes(r, k); g.fb_h.mul_acc(acc, k);
    Cpt c;
    ge_compress(acc, c.data());
    return c;
  }
};

struct Tape {  // RandomTape, random.rs:11-27
  MerlinTranscript t;
  explicit Tape(const Fl& seed) : t("proof", 5) { t.append_scalar("init_randomness", seed); }
  Fl scalar(const char* l) { return t.challenge_scalar(l); }
  FlVec vector(const char* l, size_t n) { return t.challenge_vector(l, n); }
};

// UniPoly::from_evals, unipoly.rs:23-54
static FlVec unipoly_from_evals(const FlVec& e, const Fl& two_inv, const Fl& six_inv) {
  if (e.size() == 3) {
    Fl c = e[0];
    Fl a = fl_mul(two_inv, fl_add(fl_sub(fl_sub(e[2], e[1]), e[1]), c));
    Fl b = fl_sub(fl_sub(e[1], c), a);
    return {c, b, a};
  }
  Fl d = e[0];
  Fl e1x3 = fl_add(fl_add(e[1], e[1]), e[1]), e2x3 = fl_add(fl_add(e[2], e[2]), e[2]);
  Fl a = fl_mul(six_inv, fl_sub(fl_add(fl_sub(e[3], e2x3), e1x3), e[0]));
  Fl e1x5 = fl_add(fl_add(e1x3, e[1]), e[1]), e2x4 = fl_add(e2x3, e[2]);
  Fl b = fl_mul(two_inv, fl_sub(fl_add(fl_sub(fl_add(e[0], e[0]), e1x5), e2x4), e[3]));
  Fl c = fl_sub(fl_sub(fl_sub(e[1], d), a), b);
  return {d, c, b, a};
}
static Fl unipoly_eval(const FlVec& co, const Fl& r) {  // unipoly.rs:72-80
  Fl ev = co[0], pw = r;
  for (size_t i = 1; i < co.size(); i++) {
    ev = fl_add(ev, fl_mul(pw, co[i]));
    pw = fl_mul(pw, r);
  }
  return ev;
}

// DotProductProof::prove, nizk/mod.rs:310-367 (n = 3 or 4; Cx is the caller's comm_poly: same value)
static DotProd dotproduct_prove(const HostCommit& hc, TranscriptIface& tr, Tape& tape, const FlVec& x, const Fl& blind_x,
                                const FlVec& a, const Fl& y, const Fl& blind_y, const Cpt& Cx) {
  tr.append_protocol_name("dot product proof");
  size_t n = x.size();
  FlVec d = tape.vector("d_vec", n);
  Fl r_delta = tape.scalar("r_delta"), r_beta = tape.scalar("r_beta");
  tr.append_point("Cx", Cx.data());
  Cpt Cy = hc.scalar(y, blind_y);
  tr.append_point("Cy", Cy.data());
  DotProd p;
  p.delta = hc.poly(d, r_delta);
  tr.append_point("delta", p.delta.data());
  Fl dp = fl_zero();
  for (size_t i = 0; i < n; i++) dp = fl_add(dp, fl_mul(a[i], d[i]));
  p.beta = hc.scalar(dp, r_beta);
  tr.append_point("beta", p.beta.data());
  Fl c = tr.challenge_scalar("c");
  p.z.resize(n);
  for (size_t i = 0; i < n; i++) p.z[i] = fl_add(fl_mul(c, x[i]), d[i]);
  p.z_delta = fl_add(fl_mul(c, blind_x), r_delta);
  p.z_beta = fl_add(fl_mul(c, blind_y), r_beta);
  return p;
}

struct ProverCtx {
  const Instance& inst;
  const Gens& gens;
  Workspace& w;
  TranscriptIface& tr;
  Tape& tape;
  HostCommit hc;
  Fl two_inv, six_inv;
  double host_fs = 0;  // seconds spent in host Fiat-Shamir work while the GPU is idle
  sb::ScWork scw;
  ProverCtx(const Instance& i, const Gens& g, TranscriptIface& t, Tape& tp)
      : inst(i), gens(g), w(i.ws), tr(t), tape(tp), hc{g} {
    two_inv = fl_invert(fl_from_u64(2));
    six_inv = fl_invert(fl_from_u64(6));
    scw.partials = w.sc_partials.p;
    scw.counter = w.counter.p;
    scw.result = w.small.p;  // small[0..3)
    scw.max_blocks = 148 * 4;
  }
  int fetch(const void* dev, size_t bytes, size_t pinned_off = PinnedLayout::kResult) {
    g_d2h_bytes += bytes;
    SB_CUDA(cudaMemcpyAsync(w.pinned + pinned_off, dev, bytes, cudaMemcpyDeviceToHost, w.stream));
    SB_CUDA(cudaStreamSynchronize(w.stream));
    return SB200_OK;
  }
  int upload_scalars(const FlVec& v, sb::fl_t* dev) {
    if (v.empty()) return SB200_OK;
    g_h2d_bytes += v.size() * 32;
    memcpy(w.pinned + PinnedLayout::kUpload, v.data(), v.size() * 32);
    SB_CUDA(cudaMemcpyAsync(dev, w.pinned + PinnedLayout::kUpload, v.size() * 32, cudaMemcpyHostToDevice, w.stream));
    SB_CUDA(cudaStreamSynchronize(w.stream));
    return SB200_OK;
  }
};

// ZKSumcheckInstanceProof::prove_cubic_with_additive_term / prove_quad, sumcheck.rs:426-774.
// `launch(do_bind, r)` runs the fused bind+evaluate kernel; the round evaluations come back
// through ctx.scw.result.  Host work that does not feed r_j runs after the next launch.
template <typename Launch>
static int zk_sumcheck(ProverCtx& ctx, int degree, size_t num_rounds, const Fl& claim, const Fl& blind_claim, Launch launch,
                       ZkSumcheck& pf, FlVec& r, Fl& final_blind) {
  FlVec blinds_poly = ctx.tape.vector("blinds_poly", num_rounds);
  FlVec blinds_evals = ctx.tape.vector("blinds_evals", num_rounds);
  Fl claim_per_round = claim;
  Cpt comm_claim_per_round = ctx.hc.scalar(claim, blind_claim);
  r.clear();
  SB_CUDA(launch(false, fl_zero()));
  for (size_t j = 0; j < num_rounds; j++) {
    int rc = ctx.fetch(ctx.scw.result, 96);
    if (rc) return rc;
    double t0 = now_s();
    Fl e0 = from_dev_bytes(ctx.w.pinned), e2 = from_dev_bytes(ctx.w.pinned + 32), e3 = from_dev_bytes(ctx.w.pinned + 64);
    FlVec evals = {e0, fl_sub(claim_per_round, e0), e2};
    if (degree == 3) evals.push_back(e3);
    FlVec poly = unipoly_from_evals(evals, ctx.two_inv, ctx.six_inv);
    Cpt comm_poly = ctx.hc.poly(poly, blinds_poly[j]);
    ctx.tr.append_point("comm_poly", comm_poly.data());
    pf.comm_polys.push_back(comm_poly);
    Fl r_j = ctx.tr.challenge_scalar("challenge_nextround");
    ctx.host_fs += now_s() - t0;
    SB_CUDA(launch(true, r_j));  // bind at r_j, evaluate the next round (overlaps the host work below)
    Fl eval = unipoly_eval(poly, r_j);
    Cpt comm_eval = ctx.hc.scalar(eval, blinds_evals[j]);
    ctx.tr.append_point("comm_claim_per_round", comm_claim_per_round.data());
    ctx.tr.append_point("comm_eval", comm_eval.data());
    FlVec wv = ctx.tr.challenge_vector("combine_two_claims_to_one", 2);
    Fl target = fl_add(fl_mul(wv[0], claim_per_round), fl_mul(wv[1], eval));
    const Fl& blind_sc = j == 0 ? blind_claim : blinds_evals[j - 1];
    Fl blind = fl_add(fl_mul(wv[0], blind_sc), fl_mul(wv[1], blinds_evals[j]));
    // (the reference also recomputes comm_target from the two decompressed commitments and asserts
    //  it equals Com(target; blind), sumcheck.rs:700-720: an identity, nothing enters the proof)
    size_t nc = poly.size();
    FlVec a(nc);
    Fl pw = fl_one(), two = fl_add(fl_one(), fl_one());
    for (size_t k = 0; k < nc; k++) {
      a[k] = fl_add(fl_mul(wv[0], k == 0 ? two : fl_one()), fl_mul(wv[1], pw));
      pw = fl_mul(pw, r_j);
    }
    pf.proofs.push_back(dotproduct_prove(ctx.hc, ctx.tr, ctx.tape, poly, blinds_poly[j], a, target, blind, comm_poly));
    claim_per_round = eval;
    comm_claim_per_round = comm_eval;
    r.push_back(r_j);
    pf.comm_evals.push_back(comm_eval);
  }
  final_blind = blinds_evals[num_rounds - 1];
  return SB200_OK;
}

static int prove_impl(const Instance& inst, const uint8_t* vars32, uint64_t n_vars, const uint8_t* inputs32,
                      uint64_t n_inputs, const Gens& gens, TranscriptIface& tr, const Fl& tape_seed, NizkProof& pf) {
  const size_t Nc = inst.num_cons, Nv = inst.num_vars;
  // conditions on which the reference panics
  if (n_vars > Nv) { set_error("vars longer than the instance's num_vars"); return SB200_ERR_ASSERT; }
  if (n_inputs != inst.num_inputs) { set_error("assert z.len() == num_cols (wrong number of inputs)"); return SB200_ERR_ASSERT; }
  if (!(n_inputs < Nv)) { set_error("assert input.len() < vars.len()"); return SB200_ERR_ASSERT; }  // r1csproof.rs:153
  if (gens.num_vars_padded != Nv) { set_error("gens built for a different num_vars"); return SB200_ERR_ASSERT; }
  const size_t s_x = ilog2(Nc), s_y = ilog2(2 * Nv), ell = ilog2(Nv), lnv = ell / 2;
  const size_t Ls = (size_t)1 << lnv, Rs = (size_t)1 << (ell - lnv);
  if (s_x == 0) { set_error("zero sum-check rounds (num_cons padded to 1): the reference indexes blinds_evals[-1]"); return SB200_ERR_ASSERT; }
  int rc = workspace_init(inst, (uint32_t)Rs);
  if (rc) return rc;
  Workspace& w = inst.ws;
  cudaStream_t st = w.stream;
  sb::launch_count_reset();
  g_h2d_bytes = g_d2h_bytes = 0;
  double t_begin = now_s(), t_h2d = 0;

  Tape tape(tape_seed);
  ProverCtx ctx(inst, gens, tr, tape);
  tr.append_protocol_name("Spartan NIZK proof");  // lib.rs:504
  tr.append_protocol_name("R1CS proof");          // r1csproof.rs:150

  // ---- witness to HBM: canonical copy feeds the MSM digits, Montgomery copy everything else
  {
    double t0 = now_s();
    if (vars32) {
      g_h2d_bytes += n_vars * 32;
      SB_CUDA(cudaMemcpyAsync(w.vars_canon.p, vars32, n_vars * 32, cudaMemcpyHostToDevice, st));
    } else if (w.resident_vars != n_vars) {
      set_error("vars32 == NULL but no witness of that length is resident (sb200_witness_upload)");
      return SB200_ERR_ARG;
    }
    if (Nv > n_vars) SB_CUDA(cudaMemsetAsync(w.vars_canon.p + n_vars, 0, (Nv - n_vars) * 32, st));  // Assignment::pad, lib.rs:92-105
    SB_CUDA(cudaMemsetAsync(w.flag.p, 0, 4, st));
    SB_CUDA(sb::k_to_mont(w.vars_canon.p, w.varsM.p, Nv, w.flag.p, st));
    // z = vars || 1 || inputs || 0...                         r1csproof.rs:172-180
    SB_CUDA(cudaMemcpyAsync(w.z.p, w.varsM.p, Nv * 32, cudaMemcpyDeviceToDevice, st));
    SB_CUDA(cudaMemsetAsync(w.z.p + Nv, 0, Nv * 32, st));
    FlVec tail(1 + n_inputs);
    tail[0] = fl_one();
    for (uint64_t i = 0; i < n_inputs; i++)
      if (!fl_from_bytes(inputs32 + 32 * i, tail[1 + i])) { set_error("InvalidScalar in inputs"); return SB200_ERR_INVALID_SCALAR; }
    for (size_t off = 0; off < tail.size(); off += 64) {
      FlVec part(tail.begin() + off, tail.begin() + std::min(tail.size(), off + 64));
      rc = ctx.upload_scalars(part, w.z.p + Nv + off);
      if (rc) return rc;
    }
    rc = ctx.fetch(w.flag.p, 4);
    if (rc) return rc;
    uint32_t bad;
    memcpy(&bad, w.pinned, 4);
    if (bad) { set_error("InvalidScalar in vars"); return SB200_ERR_INVALID_SCALAR; }  // Assignment::new, lib.rs:65-89
    t_h2d = now_s() - t0;
  }

  // ---- polycommit                                           r1csproof.rs:155-167, dense_mlpoly.rs:178-203
  double t_commit0 = now_s();
  FlVec blinds_vars = tape.vector("poly_blinds", Ls);
  {
    std::vector<uint8_t> bb(64 * Ls);
    for (size_t i = 0; i < Ls; i++) {
      fl_to_bytes(blinds_vars[i], &bb[32 * i]);           // canonical, for the MSM
      memcpy(&bb[32 * (Ls + i)], blinds_vars[i].v, 32);   // Montgomery, for LZ_blind
    }
    g_h2d_bytes += bb.size();
    SB_CUDA(cudaMemcpyAsync(w.blinds.p, bb.data(), bb.size(), cudaMemcpyHostToDevice, st));
    SB_CUDA(cudaStreamSynchronize(st));
  }
  SB_CUDA(sb::msm_fixed(gens.table, w.vars_canon.p, Rs, (uint32_t)Ls, (uint32_t)Rs, w.blinds.p, nullptr, (uint32_t)Rs, 0,
                        w.msm_scratch.p, w.pts.p, st, w.ev));
  rc = ctx.fetch(w.pts.p, Ls * 32, PinnedLayout::kPoints);
  if (rc) return rc;
  pf.comm_vars.resize(Ls);
  for (size_t i = 0; i < Ls; i++) pf.comm_vars[i] = cpt_from(w.pinned + PinnedLayout::kPoints + 32 * i);
  tr.append_str("poly_commitment", "poly_commitment_begin");  // dense_mlpoly.rs:289-297
  for (size_t i = 0; i < Ls; i++) tr.append_point("poly_commitment_share", pf.comm_vars[i].data());
  tr.append_str("poly_commitment", "poly_commitment_end");
  double t_commit1 = now_s();

  // ---- phase 1                                              r1csproof.rs:169-205
  FlVec tau = tr.challenge_vector("challenge_tau", s_x);
  rc = ctx.upload_scalars(tau, w.small.p + 8);
  if (rc) return rc;
  SB_CUDA(sb::k_eq_table(w.small.p + 8, (uint32_t)s_x, w.T.p, w.eq_tmp.p, st));
  SB_CUDA(sb::k_spmv3(inst.csr.view(), (uint32_t)Nc, w.z.p, w.A.p, w.B.p, w.C.p, st));
  FlVec rx;
  Fl blind_post1;
  {
    size_t len = Nc;
    auto launch = [&](bool bind, const Fl& r) {
      if (!bind) cudaEventRecord(w.ev[2], st);
      cudaError_t e = sb::k_sc_cubic(w.T.p, w.A.p, w.B.p, w.C.p, len, bind, to_dev(r), ctx.scw, st);
      if (!bind) cudaEventRecord(w.ev[3], st);
      if (bind) len /= 2;
      return e;
    };
    rc = zk_sumcheck(ctx, 3, s_x, fl_zero(), fl_zero(), launch, pf.sc1, rx, blind_post1);
    if (rc) return rc;
  }
  // final claims poly_tau[0], poly_Az[0], poly_Bz[0], poly_Cz[0]
  g_d2h_bytes += 128 + 64 + 64;  // final claims of both phases, a[0]/b[0]
  SB_CUDA(cudaMemcpyAsync(w.pinned + 0, w.T.p, 32, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaMemcpyAsync(w.pinned + 32, w.A.p, 32, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaMemcpyAsync(w.pinned + 64, w.B.p, 32, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaMemcpyAsync(w.pinned + 96, w.C.p, 32, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaStreamSynchronize(st));
  Fl tau_claim = from_dev_bytes(w.pinned), Az_claim = from_dev_bytes(w.pinned + 32), Bz_claim = from_dev_bytes(w.pinned + 64),
     Cz_claim = from_dev_bytes(w.pinned + 96);
  double t_phase1 = now_s();

  double tfs0 = now_s();
  Fl Az_blind = tape.scalar("Az_blind"), Bz_blind = tape.scalar("Bz_blind"), Cz_blind = tape.scalar("Cz_blind"),
     prod_blind = tape.scalar("prod_Az_Bz_blind");  // r1csproof.rs:209-214
  Cpt comm_Cz, comm_Az, comm_Bz, comm_prod;
  {  // KnowledgeProof::prove, nizk/mod.rs:26-51
    tr.append_protocol_name("knowledge proof");
    Fl t1 = tape.scalar("t1"), t2 = tape.scalar("t2");
    comm_Cz = ctx.hc.scalar(Cz_claim, Cz_blind);
    tr.append_point("C", comm_Cz.data());
    pf.pok_Cz.alpha = ctx.hc.scalar(t1, t2);
    tr.append_point("alpha", pf.pok_Cz.alpha.data());
    Fl c = tr.challenge_scalar("c");
    pf.pok_Cz.z1 = fl_add(fl_mul(Cz_claim, c), t1);
    pf.pok_Cz.z2 = fl_add(fl_mul(Cz_blind, c), t2);
  }
  {  // ProductProof::prove, nizk/mod.rs:158-228
    const Fl &x = Az_claim, &rX = Az_blind, &y = Bz_claim, &rY = Bz_blind, &rZ = prod_blind;
    Fl z = fl_mul(Az_claim, Bz_claim);
    tr.append_protocol_name("product proof");
    Fl b1 = tape.scalar("b1"), b2 = tape.scalar("b2"), b3 = tape.scalar("b3"), b4 = tape.scalar("b4"), b5 = tape.scalar("b5");
    comm_Az = ctx.hc.scalar(x, rX); tr.append_point("X", comm_Az.data());
    comm_Bz = ctx.hc.scalar(y, rY); tr.append_point("Y", comm_Bz.data());
    comm_prod = ctx.hc.scalar(z, rZ); tr.append_point("Z", comm_prod.data());
    pf.prod.alpha = ctx.hc.scalar(b1, b2); tr.append_point("alpha", pf.prod.alpha.data());
    pf.prod.beta = ctx.hc.scalar(b3, b4); tr.append_point("beta", pf.prod.beta.data());
    // delta = b3*X + b5*h with X = x*G + rX*h  ==  (b3*x)*G + (b3*rX + b5)*h      nizk/mod.rs:198-206
    pf.prod.delta = ctx.hc.scalar(fl_mul(b3, x), fl_add(fl_mul(b3, rX), b5));
    tr.append_point("delta", pf.prod.delta.data());
    Fl c = tr.challenge_scalar("c");
    pf.prod.z[0] = fl_add(b1, fl_mul(c, x));
    pf.prod.z[1] = fl_add(b2, fl_mul(c, rX));
    pf.prod.z[2] = fl_add(b3, fl_mul(c, y));
    pf.prod.z[3] = fl_add(b4, fl_mul(c, rY));
    pf.prod.z[4] = fl_add(b5, fl_mul(c, fl_sub(rZ, fl_mul(rX, y))));
  }
  tr.append_point("comm_Az_claim", comm_Az.data());  // r1csproof.rs:241-244
  tr.append_point("comm_Bz_claim", comm_Bz.data());
  tr.append_point("comm_Cz_claim", comm_Cz.data());
  tr.append_point("comm_prod_Az_Bz_claims", comm_prod.data());
  pf.claims_phase2[0] = comm_Az; pf.claims_phase2[1] = comm_Bz; pf.claims_phase2[2] = comm_Cz; pf.claims_phase2[3] = comm_prod;
  auto equality_prove = [&](const Fl& v1, const Fl& s1, const Fl& v2, const Fl& s2, Equality& out) {  // nizk/mod.rs:87-115
    tr.append_protocol_name("equality proof");
    Fl r = tape.scalar("r");
    Cpt C1 = ctx.hc.scalar(v1, s1); tr.append_point("C1", C1.data());
    Cpt C2 = ctx.hc.scalar(v2, s2); tr.append_point("C2", C2.data());
    out.alpha = ctx.hc.h_times(r); tr.append_point("alpha", out.alpha.data());
    Fl c = tr.challenge_scalar("c");
    out.z = fl_add(fl_mul(c, fl_sub(s1, s2)), r);
  };
  {
    Fl blind_exp1 = fl_mul(tau_claim, fl_sub(prod_blind, Cz_blind));                          // r1csproof.rs:247-249
    Fl claim_post1 = fl_mul(fl_sub(fl_mul(Az_claim, Bz_claim), Cz_claim), tau_claim);
    equality_prove(claim_post1, blind_exp1, claim_post1, blind_post1, pf.eq1);
  }

  // ---- phase 2                                              r1csproof.rs:260-292
  Fl r_A = tr.challenge_scalar("challenege_Az"), r_B = tr.challenge_scalar("challenege_Bz"), r_C = tr.challenge_scalar("challenege_Cz");
  Fl claim2 = fl_add(fl_add(fl_mul(r_A, Az_claim), fl_mul(r_B, Bz_claim)), fl_mul(r_C, Cz_claim));
  Fl blind2 = fl_add(fl_add(fl_mul(r_A, Az_blind), fl_mul(r_B, Bz_blind)), fl_mul(r_C, Cz_blind));
  ctx.host_fs += now_s() - tfs0;
  rc = ctx.upload_scalars(rx, w.small.p + 8);
  if (rc) return rc;
  SB_CUDA(sb::k_eq_table(w.small.p + 8, (uint32_t)s_x, w.T.p, w.eq_tmp.p, st));  // evals_rx reuses the tau buffer
  SB_CUDA(sb::k_spmvT3_combine(inst.csc.view(), (uint32_t)(2 * Nv), w.T.p, to_dev(r_A), to_dev(r_B), to_dev(r_C), w.abc.p, st));
  FlVec ry;
  Fl blind_post2;
  {
    size_t len = 2 * Nv;
    auto launch = [&](bool bind, const Fl& r) {
      cudaError_t e = sb::k_sc_quad(w.z.p, w.abc.p, len, bind, to_dev(r), ctx.scw, st);
      if (bind) len /= 2;
      return e;
    };
    rc = zk_sumcheck(ctx, 2, s_y, claim2, blind2, launch, pf.sc2, ry, blind_post2);
    if (rc) return rc;
  }
  SB_CUDA(cudaMemcpyAsync(w.pinned + 0, w.z.p, 32, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaMemcpyAsync(w.pinned + 32, w.abc.p, 32, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaStreamSynchronize(st));
  Fl claims2_0 = from_dev_bytes(w.pinned), claims2_1 = from_dev_bytes(w.pinned + 32);
  double t_phase2 = now_s();

  // ---- polyeval                                             r1csproof.rs:294-307, dense_mlpoly.rs:233-239,309-362
  FlVec ry1(ry.begin() + 1, ry.end());
  sb::fl_t* d_ry = w.small.p + 8;
  sb::fl_t* d_Lv = w.bound_tmp.p;            // Ls
  sb::fl_t* d_Rv = w.bound_tmp.p + Ls;       // Rs (kept intact; bl_b is the working copy)
  sb::fl_t* d_btmp = w.bound_tmp.p + Ls + Rs;
  rc = ctx.upload_scalars(ry1, d_ry);
  if (rc) return rc;
  SB_CUDA(sb::k_eq_table(d_ry, (uint32_t)lnv, d_Lv, w.eq_tmp.p, st));                 // L = eq(r[..l/2]), dense_mlpoly.rs:89-97
  SB_CUDA(sb::k_eq_table(d_ry + lnv, (uint32_t)(ell - lnv), d_Rv, w.eq_tmp.p, st));
  SB_CUDA(sb::k_bound_L(w.varsM.p, d_Lv, (uint32_t)Ls, (uint32_t)Rs, w.LZ.p, d_btmp, st));  // LZ = L * Z
  SB_CUDA(sb::k_dot_small(w.LZ.p, d_Rv, (uint32_t)Rs, w.small.p + 4, st));                 // Z(r) = <LZ, R>
  SB_CUDA(sb::k_dot_small(w.blinds.p + Ls, d_Lv, (uint32_t)Ls, w.small.p + 5, st));        // LZ_blind, dense_mlpoly.rs:347
  rc = ctx.fetch(w.small.p + 4, 64);
  if (rc) return rc;
  Fl eval_vars = from_dev_bytes(w.pinned), LZ_blind = from_dev_bytes(w.pinned + 32);
  Fl blind_eval = tape.scalar("blind_eval");
  tr.append_protocol_name("polynomial evaluation proof");  // dense_mlpoly.rs:319
  {  // DotProductProofLog::prove, nizk/mod.rs:435-512
    DotProdLog& dl = pf.eval_proof;
    tr.append_protocol_name("dot product proof (log)");
    const size_t n = Rs, lg = ilog2(n);
    Fl d = tape.scalar("d"), r_delta = tape.scalar("r_delta"), r_beta = tape.scalar("r_delta");  // sic, nizk/mod.rs:454
    FlVec bv1 = tape.vector("blinds_vec_1", 2 * lg), bv2 = tape.vector("blinds_vec_2", 2 * lg);
    // Cx = Com(LZ; LZ_blind) over gens_n: one fixed-base MSM row
    SB_CUDA(sb::k_from_mont(w.LZ.p, w.srows.p, n, st));
    FlVec ex(1);
    { uint8_t b[32]; fl_to_bytes(LZ_blind, b); memcpy(ex[0].v, b, 32); }
    rc = ctx.upload_scalars(ex, w.small.p + 6);
    if (rc) return rc;
    SB_CUDA(sb::msm_fixed(gens.table, w.srows.p, n, 1, (uint32_t)n, w.small.p + 6, nullptr, (uint32_t)Rs, 0, w.msm_scratch.p, w.pts.p, st));
    rc = ctx.fetch(w.pts.p, 32);
    if (rc) return rc;
    Cpt Cx = cpt_from(w.pinned);
    tr.append_point("Cx", Cx.data());
    pf.comm_vars_at_ry = ctx.hc.scalar(eval_vars, blind_eval);  // Cy
    tr.append_point("Cy", pf.comm_vars_at_ry.data());
    Fl blind_fin = fl_add(LZ_blind, blind_eval);  // blind_Gamma
    // BulletReductionProof::prove, nizk/bullet.rs:31-131.  The folded generators are never built: every L, R
    // and the final G_hat are fixed-base MSMs over the original generators with folded scalar coefficients.
    SB_CUDA(cudaMemcpyAsync(w.bl_a.p, w.LZ.p, n * 32, cudaMemcpyDeviceToDevice, st));
    SB_CUDA(cudaMemcpyAsync(w.bl_b.p, d_Rv, n * 32, cudaMemcpyDeviceToDevice, st));
    FlVec one1(1, fl_one());
    rc = ctx.upload_scalars(one1, w.coeff0.p);
    if (rc) return rc;
    sb::fl_t* sL = w.srows.p;
    sb::fl_t* sR = w.srows.p + (Rs + 2);
    sb::fl_t *cin = w.coeff0.p, *cout = w.coeff1.p;
    size_t cur = n;
    if (lg >= 1)
      SB_CUDA(sb::k_bullet_step(w.bl_a.p, w.bl_b.p, cin, cin, (uint32_t)cur, (uint32_t)Rs, false, to_dev(fl_zero()), to_dev(fl_zero()),
                                to_dev(bv1[0]), to_dev(bv2[0]), sL, sR, st));
    for (size_t round = 0; round < lg; round++) {
      SB_CUDA(sb::msm_fixed(gens.table, w.srows.p, Rs + 2, 2, (uint32_t)(Rs + 2), nullptr, nullptr, 0, 0, w.msm_scratch.p, w.pts.p, st));
      rc = ctx.fetch(w.pts.p, 64);
      if (rc) return rc;
      Cpt Lc = cpt_from(w.pinned), Rc = cpt_from(w.pinned + 32);
      tr.append_point("L", Lc.data());
      tr.append_point("R", Rc.data());
      Fl u = tr.challenge_scalar("u"), u_inv = fl_invert(u);
      const Fl& bL = bv1[round];
      const Fl& bR = bv2[round];
      blind_fin = fl_add(fl_add(blind_fin, fl_mul(fl_mul(bL, u), u)), fl_mul(fl_mul(bR, u_inv), u_inv));
      Fl nbL = round + 1 < lg ? bv1[round + 1] : fl_zero(), nbR = round + 1 < lg ? bv2[round + 1] : fl_zero();
      SB_CUDA(sb::k_bullet_step(w.bl_a.p, w.bl_b.p, cin, cout, (uint32_t)cur, (uint32_t)Rs, true, to_dev(u), to_dev(u_inv), to_dev(nbL),
                                to_dev(nbR), sL, sR, st));
      std::swap(cin, cout);
      cur /= 2;
      dl.L_vec.push_back(Lc);
      dl.R_vec.push_back(Rc);
    }
    // x_hat = a[0], a_hat = b[0]; g_hat = sum_g coeff[g]*G[g]
    SB_CUDA(cudaMemcpyAsync(w.pinned + 0, w.bl_a.p, 32, cudaMemcpyDeviceToHost, st));
    SB_CUDA(cudaMemcpyAsync(w.pinned + 32, w.bl_b.p, 32, cudaMemcpyDeviceToHost, st));
    // delta = d*g_hat + r_delta*h                             nizk/mod.rs:483-490
    SB_CUDA(sb::k_scale_from_mont(cin, to_dev(d), (uint32_t)Rs, w.srows.p, st));
    SB_CUDA(cudaStreamSynchronize(st));
    Fl x_hat = from_dev_bytes(w.pinned), a_hat = from_dev_bytes(w.pinned + 32);
    { uint8_t b[32]; fl_to_bytes(r_delta, b); memcpy(ex[0].v, b, 32); }
    rc = ctx.upload_scalars(ex, w.small.p + 6);
    if (rc) return rc;
    SB_CUDA(sb::msm_fixed(gens.table, w.srows.p, Rs, 1, (uint32_t)Rs, w.small.p + 6, nullptr, (uint32_t)Rs, 0, w.msm_scratch.p, w.pts.p, st));
    rc = ctx.fetch(w.pts.p, 32);
    if (rc) return rc;
    dl.delta = cpt_from(w.pinned);
    tr.append_point("delta", dl.delta.data());
    dl.beta = ctx.hc.scalar(d, r_beta);
    tr.append_point("beta", dl.beta.data());
    Fl c = tr.challenge_scalar("c");
    Fl y_hat = fl_mul(x_hat, a_hat);
    dl.z1 = fl_add(d, fl_mul(c, y_hat));
    dl.z2 = fl_add(fl_mul(a_hat, fl_add(fl_mul(c, blind_fin), r_beta)), r_delta);
  }
  double t_polyeval = now_s();
  {  // r1csproof.rs:309-321
    Fl blind_eval_Z = fl_mul(fl_sub(fl_one(), ry[0]), blind_eval);
    Fl blind_exp2 = fl_mul(claims2_1, blind_eval_Z);
    Fl claim_post2 = fl_mul(claims2_0, claims2_1);
    equality_prove(claim_post2, blind_exp2, claim_post2, blind_post2, pf.eq2);
  }
  pf.rx = rx;
  pf.ry = ry;
  double t_end = now_s();
  g_profile[0] = t_commit1 - t_commit0;
  g_profile[1] = t_phase1 - t_commit1;
  g_profile[2] = t_phase2 - t_phase1;
  g_profile[3] = t_polyeval - t_phase2;
  g_profile[4] = t_end - t_begin;
  g_profile[5] = t_h2d;
  g_profile[6] = ctx.host_fs;
  g_profile[7] = (double)sb::launch_count();
  {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, w.ev[0], w.ev[1]) == cudaSuccess) g_profile[10] = ms * 1e-3;
    if (cudaEventElapsedTime(&ms, w.ev[2], w.ev[3]) == cudaSuccess) g_profile[11] = ms * 1e-3;
  }
  g_profile[8] = (double)g_h2d_bytes;
  g_profile[9] = (double)g_d2h_bytes;
  return SB200_OK;
}

// caller-owned transcript through the vtable
class VtTranscript : public TranscriptIface {
  const sb200_transcript_vt* vt_;

 public:
  explicit VtTranscript(const sb200_transcript_vt* vt) : vt_(vt) {}
  void append_message(const char* label, const uint8_t* m, size_t n) override {
    vt_->append_message(vt_->ctx, (const uint8_t*)label, strlen(label), m, n);
  }
  void challenge_bytes(const char* label, uint8_t* out, size_t n) override {
    vt_->challenge_bytes(vt_->ctx, (const uint8_t*)label, strlen(label), out, n);
  }
};

}  // namespace sbh

extern "C" {

int sb200_nizk_prove(const sb200_instance* hi, const uint8_t* vars32, uint64_t n_vars, const uint8_t* inputs32,
                     uint64_t n_inputs, const sb200_gens* hg, const sb200_transcript_vt* transcript, const uint8_t* label,
                     size_t label_len, const uint8_t* tape_seed64, const sb200_comm_vt* comm, uint8_t** proof_out,
                     size_t* proof_len) {
  if (!hi || !hg || !proof_out || !proof_len || (!inputs32 && n_inputs)) return SB200_ERR_ARG;
  if (comm && comm->world > 1) { set_error("sharded prove: use sb200_nizk_prove with world == 1 in this build"); return SB200_ERR_ARG; }
  int rc = ensure_device();
  if (rc) return rc;
  const Instance& inst = *reinterpret_cast<const Instance*>(hi);
  const Gens& gens = *reinterpret_cast<const Gens*>(hg);
  uint8_t seed[64];
  if (tape_seed64) memcpy(seed, tape_seed64, 64);
  else {  // Scalar::random(&mut OsRng), random.rs:13-15
    std::random_device rd;
    for (int i = 0; i < 64; i += 4) { uint32_t v = rd(); memcpy(seed + i, &v, 4); }
  }
  Fl seed_scalar = fl_from_bytes_wide(seed);
  NizkProof pf;
  try {
    if (transcript) {
      VtTranscript tr(transcript);
      rc = prove_impl(inst, vars32, n_vars, inputs32, n_inputs, gens, tr, seed_scalar, pf);
    } else {
      MerlinTranscript tr(label, label_len);
      rc = prove_impl(inst, vars32, n_vars, inputs32, n_inputs, gens, tr, seed_scalar, pf);
    }
  } catch (const std::exception& e) {
    set_error(std::string("exception: ") + e.what());
    return SB200_ERR_ASSERT;
  }
  if (rc) return rc;
  BincodeOut w;
  std::vector<uint8_t>& b = w.write(pf);
  *proof_out = (uint8_t*)malloc(b.size());
  memcpy(*proof_out, b.data(), b.size());
  *proof_len = b.size();
  return SB200_OK;
}

// Places the witness in HBM ahead of time; a following sb200_nizk_prove with vars32 == NULL and the same
// n_vars uses it (the "inputs already resident" measurement of bench.py).
int sb200_witness_upload(const sb200_instance* hi, const sb200_gens* hg, const uint8_t* vars32, uint64_t n_vars) {
  if (!hi || !hg || (!vars32 && n_vars)) return SB200_ERR_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  const Instance& inst = *reinterpret_cast<const Instance*>(hi);
  if (n_vars > inst.num_vars) return SB200_ERR_ARG;
  rc = workspace_init(inst, reinterpret_cast<const Gens*>(hg)->R);
  if (rc) return rc;
  Workspace& w = inst.ws;
  SB_CUDA(cudaMemcpyAsync(w.vars_canon.p, vars32, n_vars * 32, cudaMemcpyHostToDevice, w.stream));
  SB_CUDA(cudaStreamSynchronize(w.stream));
  w.resident_vars = n_vars;
  return SB200_OK;
}
// cudaStream_t the instance's kernels run on (for CUDA-event timing by the caller); 0 before the first prove
uint64_t sb200_instance_stream(const sb200_instance* hi) {
  return hi ? (uint64_t)(uintptr_t) reinterpret_cast<const Instance*>(hi)->ws.stream : 0;
}

int sb200_last_prove_profile(double out[12]) {
  for (int i = 0; i < 12; i++) out[i] = g_profile[i];
  return SB200_OK;
}

}  // extern "C"

#include "verify.inc"
#include "kernel_hooks.inc"
