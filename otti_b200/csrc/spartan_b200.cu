// libspartan_b200: C-ABI entry points (include/spartan_b200.h) and the host
// driver of the B200 Spartan NIZK prover.  The host keeps the Fiat-Shamir
// transcript, the RandomTape and the constant-size sigma protocols; every loop
// over the witness, the constraint hypercube or the generators runs in the
// sm_100a kernels of kernels.cu.  Protocol order follows
// Spartan/src/lib.rs:493-535 and r1csproof.rs:141-345 statement by statement
// (transcript framing SURVEY.md App. A.1, tape draw order App. A.2).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <random>
#include <thread>

#include "../../include/spartan_b200.h"
#include "../../include/spartan_b200_kernels.h"
#include "host/objects.hpp"
#include "host/proof.hpp"
#include "host/affinity.hpp"

namespace sbh {

static thread_local std::string g_err;
void set_error(const std::string& s) { g_err = s; }
const char* get_error() { return g_err.c_str(); }
static thread_local double g_profile[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
static std::atomic<uint64_t> g_h2d_bytes{0}, g_d2h_bytes{0};  // one proof at a time per process (prove_mu); counted from two threads

static bool g_device_ok = false;
static std::mutex g_prove_mu;  // NIZK::prove and SNARK::prove: one proof at a time per process (the worker pool has a single posting thread)
static thread_local int g_window_override = 0;  // set around sb200_gens_new_for by sb200_snark_gens_new_for (snark.inc)
static int g_bound_cpus = 0;
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
// vector storage that resize() leaves uninitialised (every element is written by the conversion threads)
template <typename T>
struct RawAlloc : std::allocator<T> {
  template <typename U> struct rebind { typedef RawAlloc<U> other; };
  RawAlloc() = default;
  template <typename U> RawAlloc(const RawAlloc<U>&) {}
  template <typename U> void construct(U* p) { ::new ((void*)p) U; }
  template <typename U, typename... A> void construct(U* p, A&&... a) { ::new ((void*)p) U(std::forward<A>(a)...); }
};

// threads for the host-side loops of Instance::new and the zkinterface reader (independent of the prover's worker pool)
static unsigned ingest_threads(size_t items) {
  unsigned t = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
  if (g_bound_cpus > 0) t = std::min<unsigned>(t, (unsigned)g_bound_cpus);
  if (items < (size_t)1 << 15) t = 1;
  return t;
}
template <typename F>
static void run_threads(unsigned n, F&& f) {  // f(t) for t < n, f(0) on the caller
  std::vector<std::thread> th;
  for (unsigned t = 1; t < n; t++) th.emplace_back([&f, t] { f(t); });
  f(0);
  for (auto& x : th) x.join();
}

// One matrix in the caller's entry order: 32-bit global coordinates (the columns already moved to the padded z layout) and
// the values, either the caller's canonical 32-byte scalars or Montgomery limbs (the synthetic generator).
typedef std::vector<uint32_t, RawAlloc<uint32_t>> U32Vec;
static_assert(sizeof(Fl) == 32 && sizeof(sb::fl_t) == 32, "scalars travel to the device as 8 x u32");
struct CooHost {
  U32Vec rows, cols;
  const uint8_t* canon = nullptr;
  const Fl* mont = nullptr;
  size_t nnz = 0;
};
// long segments of one sorted (orientation, matrix) -> chunk lists (dev.hpp LongSegs), from the segment pointers
static int upload_long_segs(const std::vector<uint32_t>& ptr, uint64_t nseg, SparseDev& sd, int m) {
  std::vector<uint32_t> lseg, lfirst, lo, hi, big;
  for (uint64_t sgm = 0; sgm < nseg; sgm++) {
    const uint32_t a = ptr[sgm], b = ptr[sgm + 1];
    if (b - a <= sb::kLongSeg) continue;
    lseg.push_back((uint32_t)sgm);
    lfirst.push_back((uint32_t)lo.size());
    for (uint32_t q = a; q < b; q += sb::kSegChunk) { lo.push_back(q); hi.push_back(std::min(b, q + sb::kSegChunk)); }
    if (lo.size() - lfirst.back() > sb::kSumInline) big.push_back((uint32_t)lseg.size() - 1);
  }
  lfirst.push_back((uint32_t)lo.size());
  sd.nlong[m] = (uint32_t)lseg.size(); sd.nchunks[m] = (uint32_t)lo.size(); sd.nbig[m] = (uint32_t)big.size();
  auto put = [](DevBuf<uint32_t>& d, const std::vector<uint32_t>& v) -> cudaError_t {
    cudaError_t e = d.alloc(v.size() ? v.size() : 1);
    if (e != cudaSuccess || v.empty()) return e;
    return cudaMemcpy(d.p, v.data(), v.size() * 4, cudaMemcpyHostToDevice);
  };
  SB_CUDA(put(sd.lg_seg[m], lseg)); SB_CUDA(put(sd.lg_first[m], lfirst)); SB_CUDA(put(sd.lg_lo[m], lo));
  SB_CUDA(put(sd.lg_hi[m], hi)); SB_CUDA(put(sd.lg_big[m], big));
  SB_CUDA(sd.lg_partial[m].alloc(lo.size() ? lo.size() : 1));
  SB_CUDA(sd.lg_reduced[m].alloc(lseg.size() ? lseg.size() : 1));
  return SB200_OK;
}
// The device form of the three matrices: CSR by constraint row and CSC by z column, both stable in the caller's entry
// order, built by sparse_build.cu from one upload of each matrix.  bad_scalar[m] receives the index of the first scalar
// >= l of matrix m (0xFFFFFFFF if none); the build stops at the first matrix that has one.
static int upload_sparse(const CooHost M[3], uint64_t nrows, uint64_t ncols, Instance& inst, uint32_t bad_scalar[3]) {
  const uint32_t g = (uint32_t)inst.world, k = (uint32_t)inst.rank, sh = inst.gamma;
  for (int m = 0; m < 3; m++) bad_scalar[m] = 0xFFFFFFFFu;
  cudaStream_t s = 0;
  for (int m = 0; m < 3; m++) {
    const size_t nnz = M[m].nnz;
    const uint32_t n32 = (uint32_t)nnz;
    DevBuf<uint32_t> d_rows, d_cols, d_key, d_bad;
    DevBuf<sb::fl_t> d_in, d_mont;
    DevBuf<uint8_t> d_tmp;
    const size_t tmp_bytes = sb::sparse_build_tmp_bytes(n32);
    SB_CUDA(d_rows.alloc(nnz ? nnz : 1)); SB_CUDA(d_cols.alloc(nnz ? nnz : 1)); SB_CUDA(d_bad.alloc(1));
    SB_CUDA(d_mont.alloc(nnz ? nnz : 1)); SB_CUDA(d_tmp.alloc(tmp_bytes));
    if (g > 1) SB_CUDA(d_key.alloc(nnz ? nnz : 1));
    if (nnz) {
      SB_CUDA(cudaMemcpyAsync(d_rows.p, M[m].rows.data(), nnz * 4, cudaMemcpyHostToDevice, s));
      SB_CUDA(cudaMemcpyAsync(d_cols.p, M[m].cols.data(), nnz * 4, cudaMemcpyHostToDevice, s));
    }
    if (M[m].canon != nullptr) {
      SB_CUDA(d_in.alloc(nnz ? nnz : 1));
      if (nnz) SB_CUDA(cudaMemcpyAsync(d_in.p, M[m].canon, nnz * 32, cudaMemcpyHostToDevice, s));
      SB_CUDA(sb::k_to_mont_first_bad(d_in.p, d_mont.p, n32, d_bad.p, s));
      SB_CUDA(cudaMemcpyAsync(&bad_scalar[m], d_bad.p, 4, cudaMemcpyDeviceToHost, s));
    } else if (nnz) {
      SB_CUDA(cudaMemcpyAsync(d_mont.p, M[m].mont, nnz * 32, cudaMemcpyHostToDevice, s));
    }
    for (int orient = 0; orient < 2; orient++) {
      SparseDev& sd = orient == 0 ? inst.csr : inst.csc;
      const uint64_t nseg = (orient == 0 ? nrows : ncols) >> sh;
      const U32Vec& coord = orient == 0 ? M[m].rows : M[m].cols;
      const uint32_t* d_coord = orient == 0 ? d_rows.p : d_cols.p;
      const uint32_t* d_other = orient == 0 ? d_cols.p : d_rows.p;
      size_t kept = nnz;
      const uint32_t* d_keys = d_coord;
      U32Vec keys;
      if (g > 1) {  // this rank keeps the segments congruent to its rank; the other entries sort behind them
        keys.resize(nnz);
        const unsigned T = ingest_threads(nnz);
        std::vector<size_t> cnt(T, 0);
        run_threads(T, [&](unsigned t) {
          size_t c = 0;
          for (size_t i = nnz * t / T, e = nnz * (t + 1) / T; i < e; i++) {
            const uint32_t key = coord[i];
            const bool mine = (key & (g - 1)) == k;
            keys[i] = mine ? key >> sh : 0xFFFFFFFFu;
            c += mine;
          }
          cnt[t] = c;
        });
        kept = 0;
        for (size_t c : cnt) kept += c;
        if (nnz) SB_CUDA(cudaMemcpyAsync(d_key.p, keys.data(), nnz * 4, cudaMemcpyHostToDevice, s));
        d_keys = d_key.p;
      }
      uint32_t key_bits = 1;
      while (key_bits < 32 && ((uint64_t)1 << key_bits) < nseg) key_bits++;
      uint32_t* d_op = nullptr;
      if (g == 1) {  // SNARK mode reads the op each position came from
        DevBuf<uint32_t>& od = orient == 0 ? inst.op_csr[m] : inst.op_csc[m];
        SB_CUDA(od.alloc(kept ? kept : 1));
        d_op = od.p;
      }
      SB_CUDA(sd.ptr[m].alloc(nseg + 1));
      SB_CUDA(sd.idx[m].alloc(kept ? kept : 1));
      SB_CUDA(sd.val[m].alloc(kept ? kept : 1));
      SB_CUDA(sb::k_sparse_build(d_keys, d_other, d_mont.p, n32, (uint32_t)kept, (uint32_t)nseg, key_bits, sd.ptr[m].p, sd.idx[m].p,
                                 sd.val[m].p, d_op, d_tmp.p, tmp_bytes, s));
      std::vector<uint32_t> ptr(nseg + 1);
      SB_CUDA(cudaMemcpyAsync(ptr.data(), sd.ptr[m].p, (nseg + 1) * 4, cudaMemcpyDeviceToHost, s));
      SB_CUDA(cudaStreamSynchronize(s));  // also keeps `keys` alive until its copy is done
      if (bad_scalar[m] != 0xFFFFFFFFu) return SB200_OK;  // the caller reports InvalidScalar
      int rc = upload_long_segs(ptr, nseg, sd, m);
      if (rc) return rc;
    }
  }
  return SB200_OK;
}
// world must be 1, 2, 4 or 8 and leave every rank at least two constraint rows, two z columns and one commitment row
static int shard_check(int rank, int world, uint64_t ncp, uint64_t nvp, uint32_t& gamma) {
  gamma = 0;
  if (world < 1 || (world & (world - 1)) || world > 8 || rank < 0 || rank >= world) {
    set_error("sharding: world must be 1, 2, 4 or 8 and 0 <= rank < world");
    return SB200_ERR_ARG;
  }
  while ((1 << gamma) < world) gamma++;
  uint64_t ell = 0;
  while (((uint64_t)1 << (ell + 1)) <= nvp) ell++;
  if (world > 1 && (ncp < 2ull * world || 2 * nvp < 2ull * world || ((uint64_t)1 << (ell / 2)) < (uint64_t)world)) {
    set_error("sharding: instance too small for this many GPUs");
    return SB200_ERR_ARG;
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
  g_bound_cpus = sbh::bind_near_device(device, n);  // one process per GPU: host threads next to the GPU (host/affinity.hpp)
  return ensure_device();
}
// CPUs the calling thread was bound to by the last sb200_init (0 = affinity left alone)
int sb200_bound_cpus(void) { return g_bound_cpus; }

int sb200_instance_new_sharded(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, const uint64_t* const rows[3],
                               const uint64_t* const cols[3], const uint8_t* const vals32[3], const uint64_t nnz[3],
                               int rank, int world, sb200_instance** out) {
  if (!out) return SB200_ERR_ARG;
  *out = nullptr;
  uint64_t ncp, nvp;
  padded_dims(num_cons, num_vars, num_inputs, ncp, nvp);
  if (ncp >= (1ull << 31) || nvp >= (1ull << 30)) {
    set_error("instance too large for 32-bit indices");
    return SB200_ERR_ARG;
  }
  const bool itrace = getenv("SB200_TRACE_INSTANCE") != nullptr;  // stage times of this call on stderr
  double it0 = now_s();
  auto imark = [&](const char* what) {
    if (!itrace) return;
    double t = now_s();
    fprintf(stderr, "[sb200-instance] %8.3f ms  %s\n", (t - it0) * 1e3, what);
    it0 = t;
  };
  CooHost M[3];
  uint64_t bad_at[3] = {~0ull, ~0ull, ~0ull};
  int bad_code[3] = {0, 0, 0};
  bool any_bad = false;
  for (int m = 0; m < 3; m++) {
    if (nnz[m] >= (1ull << 31)) { set_error("more than 2^31 - 1 non-zero entries in one matrix"); return SB200_ERR_ARG; }  // the device sort counts in int
    M[m].nnz = nnz[m];
    M[m].canon = vals32[m];
    M[m].rows.resize(nnz[m]);
    M[m].cols.resize(nnz[m]);
    // the indices are checked and narrowed in slices on the host; the scalars are checked by the kernel that converts them
    const unsigned T = ingest_threads(nnz[m]);
    std::vector<uint64_t> t_bad(T, ~0ull);
    std::vector<int> t_code(T, 0);
    run_threads(T, [&](unsigned t) {
      const uint64_t k0 = nnz[m] * t / T, k1 = nnz[m] * (t + 1) / T;
      for (uint64_t k = k0; k < k1; k++) {
        uint64_t row = rows[m][k], col = cols[m][k];
        int code = 0;
        if (row >= num_cons) code = 1;                                    // lib.rs:166-168
        else if (col >= num_vars + 1 + num_inputs) code = 2;              // lib.rs:171-173
        if (code) { t_bad[t] = k; t_code[t] = code; return; }
        if (col >= num_vars) col += nvp - num_vars;  // 1 and the inputs live in the second half of z, lib.rs:180-181
        M[m].rows[k] = (uint32_t)row;
        M[m].cols[k] = (uint32_t)col;
      }
    });
    for (unsigned t = 0; t < T; t++) {
      if (t_bad[t] == ~0ull) continue;
      bad_at[m] = t_bad[t]; bad_code[m] = t_code[t]; any_bad = true;
      break;
    }
    // lib.rs:191-197 appends zero-valued entries when num_cons is 0 or 1; they contribute nothing to any product
    // (SNARK::encode would see them as ops: Instance::pad_entries makes sb200_snark_encode refuse such an instance).
  }
  if (any_bad) {
    // the error reported is the one the reference meets first: matrices in order, entries in order, and for one entry the
    // row, the column, then the scalar (lib.rs:163-187).  Only this path reads the scalars on the host.
    for (int m = 0; m < 3; m++) {
      const uint64_t end = bad_at[m] == ~0ull ? nnz[m] : bad_at[m];
      for (uint64_t k = 0; k < end; k++) {
        Fl v;
        if (!fl_from_bytes(vals32[m] + 32 * k, v)) { set_error("InvalidScalar"); return SB200_ERR_INVALID_SCALAR; }
      }
      if (bad_at[m] == ~0ull) continue;
      set_error(bad_code[m] == 1 ? "InvalidIndex: row" : "InvalidIndex: col");
      return SB200_ERR_INVALID_INDEX;
    }
  }
  imark("indices checked and narrowed");
  uint32_t gamma = 0;
  int rc = shard_check(rank, world, ncp, nvp, gamma);
  if (rc) return rc;
  rc = ensure_device();
  if (rc) return rc;
  Instance* inst = new Instance();
  inst->num_cons = ncp; inst->num_vars = nvp; inst->num_inputs = num_inputs;
  inst->rank = rank; inst->world = world; inst->gamma = gamma;
  for (int m = 0; m < 3; m++) inst->nnz[m] = M[m].nnz;
  inst->pad_entries = num_cons == 0 || num_cons == 1;
  uint32_t bad_scalar[3];
  rc = upload_sparse(M, ncp, 2 * nvp, *inst, bad_scalar);
  if (rc) { delete inst; return rc; }
  for (int m = 0; m < 3; m++)
    if (bad_scalar[m] != 0xFFFFFFFFu) {  // lib.rs:175-187
      delete inst;
      set_error("InvalidScalar");
      return SB200_ERR_INVALID_SCALAR;
    }
  imark("scalars converted, CSR / CSC ordered on the device");
  *out = reinterpret_cast<sb200_instance*>(inst);
  return SB200_OK;
}
int sb200_instance_new(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, const uint64_t* const rows[3],
                       const uint64_t* const cols[3], const uint8_t* const vals32[3], const uint64_t nnz[3],
                       sb200_instance** out) {
  return sb200_instance_new_sharded(num_cons, num_vars, num_inputs, rows, cols, vals32, nnz, 0, 1, out);
}

static inline uint64_t splitmix64(uint64_t& s) {
  s += 0x9E3779B97F4A7C15ULL;
  uint64_t z = s;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

// R1CSInstance::produce_synthetic_r1cs, r1csinstance.rs:136-214
int sb200_instance_synthetic_sharded(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t seed, int rank,
                                     int world, sb200_instance** out, uint8_t* vars32_out, uint8_t* inputs32_out) {
  if (!out) return SB200_ERR_ARG;
  *out = nullptr;
  auto pow2 = [](uint64_t x) { return x && !(x & (x - 1)); };
  if (!pow2(num_cons)) { set_error("NonPowerOfTwoCons"); return SB200_ERR_NON_POW2_CONS; }   // :146
  if (!pow2(num_vars)) { set_error("NonPowerOfTwoVars"); return SB200_ERR_NON_POW2_VARS; }   // :147
  if (!(num_inputs < num_vars)) { set_error("assert num_inputs < num_vars"); return SB200_ERR_ASSERT; }  // :150
  if (num_cons >= (1ull << 31) || num_vars >= (1ull << 30)) return SB200_ERR_ARG;
  uint32_t gamma = 0;
  int rc = shard_check(rank, world, num_cons, num_vars, gamma);
  if (rc) return rc;
  rc = ensure_device();
  if (rc) return rc;
  const uint64_t size_z = num_vars + num_inputs + 1;
  // Every loop below runs in slices on the ingestion threads: SplitMix64's state after k draws is seed + k * increment, so a
  // slice starts its stream where the sequential generator would be; the inverses come from one inversion per slice.
  FlVec Z(size_z), Zinv(size_z);
  const unsigned T = ingest_threads(size_z);
  run_threads(T, [&](unsigned t) {
    const uint64_t i0 = size_z * t / T, i1 = size_z * (t + 1) / T;
    uint64_t st = seed + 8 * i0 * 0x9E3779B97F4A7C15ULL;
    for (uint64_t i = i0; i < i1; i++) {  // Scalar::random = from_u512(8 x next_u64), ristretto255.rs:381-387
      uint8_t wide[64];
      for (int k = 0; k < 8; k++) {
        uint64_t w = splitmix64(st);
        memcpy(wide + 8 * k, &w, 8);
      }
      Z[i] = fl_from_bytes_wide(wide);
    }
    if (num_vars >= i0 && num_vars < i1) Z[num_vars] = fl_one();
    // Z^-1 for the C coefficients with one shared inversion per slice
    FlVec pre(i1 - i0);
    Fl acc = fl_one();
    for (uint64_t i = i0; i < i1; i++) {
      pre[i - i0] = acc;
      if (!Z[i].is_zero()) acc = fl_mul(acc, Z[i]);
    }
    acc = fl_invert(acc);
    for (uint64_t i = i1; i-- > i0;) {
      if (Z[i].is_zero()) { Zinv[i] = fl_zero(); continue; }
      Zinv[i] = fl_mul(acc, pre[i - i0]);
      acc = fl_mul(acc, Z[i]);
    }
  });
  CooHost M[3];
  FlVec V[3];
  for (int m = 0; m < 3; m++) {
    M[m].nnz = num_cons;
    M[m].rows.resize(num_cons); M[m].cols.resize(num_cons); V[m].resize(num_cons);
  }
  const unsigned Tc = ingest_threads(num_cons);
  run_threads(Tc, [&](unsigned t) {
    for (uint64_t i = num_cons * t / Tc, e = num_cons * (t + 1) / Tc; i < e; i++) {
      uint64_t ai = i % size_z, bi = (i + 2) % size_z, ci = (i + 3) % size_z;
      for (int m = 0; m < 3; m++) M[m].rows[i] = (uint32_t)i;
      M[0].cols[i] = (uint32_t)ai; V[0][i] = fl_one();
      M[1].cols[i] = (uint32_t)bi; V[1][i] = fl_one();
      Fl ab = fl_mul(Z[ai], Z[bi]);
      if (Z[ci].is_zero()) { M[2].cols[i] = (uint32_t)num_vars; V[2][i] = ab; }
      else { M[2].cols[i] = (uint32_t)ci; V[2][i] = fl_mul(ab, Zinv[ci]); }
    }
  });
  for (int m = 0; m < 3; m++) M[m].mont = V[m].data();
  Instance* inst = new Instance();
  inst->num_cons = num_cons; inst->num_vars = num_vars; inst->num_inputs = num_inputs;
  inst->rank = rank; inst->world = world; inst->gamma = gamma;
  for (int m = 0; m < 3; m++) inst->nnz[m] = M[m].nnz;
  uint32_t bad_scalar[3];
  rc = upload_sparse(M, num_cons, 2 * num_vars, *inst, bad_scalar);
  if (rc) { delete inst; return rc; }
  run_threads(T, [&](unsigned t) {
    for (uint64_t i = num_vars * t / T, e = num_vars * (t + 1) / T; i < e; i++) fl_to_bytes(Z[i], vars32_out + 32 * i);
  });
  for (uint64_t i = 0; i < num_inputs; i++) fl_to_bytes(Z[num_vars + 1 + i], inputs32_out + 32 * i);
  *out = reinterpret_cast<sb200_instance*>(inst);
  return SB200_OK;
}
int sb200_instance_synthetic(uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs, uint64_t seed,
                             sb200_instance** out, uint8_t* vars32_out, uint8_t* inputs32_out) {
  return sb200_instance_synthetic_sharded(num_cons, num_vars, num_inputs, seed, 0, 1, out, vars32_out, inputs32_out);
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
  std::lock_guard<std::mutex> inst_lock(inst.mu);  // the handle's workspace, stream and scratch serve one call at a time
  if (n_vars > inst.num_vars) { set_error("InvalidNumberOfInputs"); return SB200_ERR_INVALID_NUMBER_OF_INPUTS; }      // lib.rs:233-235
  if (n_inputs != inst.num_inputs) { set_error("InvalidNumberOfInputs"); return SB200_ERR_INVALID_NUMBER_OF_INPUTS; }  // lib.rs:237-239
  if (inst.world > 1) { set_error("is_sat needs an unsharded instance"); return SB200_ERR_ARG; }
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
  uint64_t expected = 0;  // 0 = many proofs: the widest tables that fit
  if (const char* env = getenv("SB200_PROOFS_EXPECTED")) expected = strtoull(env, nullptr, 10);
  return sb200_gens_new_for(label, label_len, num_cons, num_vars, num_inputs, expected, out);
}
int sb200_gens_new_for(const uint8_t* label, size_t label_len, uint64_t num_cons, uint64_t num_vars, uint64_t num_inputs,
                       uint64_t proofs_expected, sb200_gens** out) {
  (void)num_cons;
  if (!out) return SB200_ERR_ARG;
  *out = nullptr;
  int rc = ensure_device();
  if (rc) return rc;
  const bool gtrace = getenv("SB200_TRACE_GENS") != nullptr;  // stage times of this call on stderr
  double gt0 = now_s();
  auto gmark = [&](const char* what) {
    if (!gtrace) return;
    double t = now_s();
    fprintf(stderr, "[sb200-gens] %8.3f ms  %s\n", (t - gt0) * 1e3, what);
    gt0 = t;
  };
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
    std::vector<uint8_t> stream(64 * npts);
    xof.read(stream.data(), stream.size());
    // from_uniform_bytes is two Elligator maps (~15 us a point): spread over the host cores
    unsigned nthr = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
    if (g_bound_cpus > 0) nthr = std::min<unsigned>(nthr, (unsigned)g_bound_cpus);
    if (npts < 64) nthr = 1;
    auto work = [&](unsigned t) { for (size_t i = t; i < npts; i += nthr) pts[i] = ge_from_uniform_bytes(&stream[64 * i]); };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nthr; t++) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
  }
  gmark("generator points (SHAKE256 + from_uniform_bytes)");
  g->G.assign(pts.begin(), pts.begin() + R);
  g->Q = pts[R];
  g->h = pts[R + 1];
  for (int i = 0; i < 4; i++) g->G4[i] = pts[i];
  g->h4 = pts[4];
  for (int i = 0; i < 4; i++) g->fb_G4[i].point = g->G4[i];
  g->fb_h4.point = g->h4;
  g->fb_Q.point = g->Q;
  g->fb_h.point = g->h;
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
  // Window width of the device tables: every extra bit removes madds from the commitment kernel
  // (floor(252/c)+1 per term) and doubles the table.  Largest c whose table fits SB200_TABLE_MB
  // (default: 112 GiB of the 180 GB HBM, less if the device has not got 24 GiB more than that free for the
  // instance and the workspace: c = 16/16/16/15 at 2^16/2^20/2^22/2^24); SB200_MSM_WINDOW forces a value.
  uint32_t c = 8;
  {
    size_t budget = (size_t)114688 << 20;
    {
      size_t free_b = 0, total_b = 0;
      const size_t reserve = (size_t)24 << 30;
      if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess)
        budget = std::min(budget, free_b > reserve + ((size_t)1 << 30) ? free_b - reserve : (size_t)1 << 30);
    }
    if (const char* env = getenv("SB200_TABLE_MB")) {
      long v = atol(env);
      if (v > 0) budget = (size_t)v << 20;
    }
    // proofs_expected > 0: the width that minimises  table build + proofs_expected x commitment,  in mixed additions:
    // an entry of the table costs ~3 (a full addition + its share of the chunk's inversion and normalisation), a
    // term of the witness one per window.  A one-shot caller (spzk: gens + one proof per process, main.rs:71-85) gets
    // a table that is built in milliseconds instead of the widest one that fits the HBM.
    double best = 0;
    for (uint32_t cc = 8; cc <= 16; cc++) {
      const size_t W = 252 / cc + 1, entries = bases.size() * W * ((size_t)1 << (cc - 1));
      if (entries * 96 > budget && cc > 8) continue;
      if (proofs_expected == 0) { c = cc; continue; }
      const double terms = (double)nvp + (double)((size_t)1 << (ell / 2));
      const double cost = 3.0 * (double)entries + (double)proofs_expected * terms * (double)W;
      if (best == 0 || cost < best) { best = cost; c = cc; }
    }
  }
  if (g_window_override >= 2 && g_window_override <= 16) c = (uint32_t)g_window_override;  // SNARKGens: widths chosen jointly
  if (const char* env = getenv("SB200_MSM_WINDOW")) {
    int v = atoi(env);
    if (v >= 2 && v <= 16) c = (uint32_t)v;
  }
  gmark("affine bases + window choice");
  cudaError_t ce = sb::msm_table_build(g->table, xy.data(), (uint32_t)bases.size(), c, 0);
  gmark("device window table");
  if (ce != cudaSuccess) {
    set_error(std::string("msm_table_build: ") + cudaGetErrorString(ce));
    delete g;
    return SB200_ERR_CUDA;
  }
  // Host copies of the table slices of the seven generators of the Fiat-Shamir tail (FixedBase::wide, hge.hpp): the
  // stream points 0..4 (gens_3 / gens_4), Q = point R and h = point R+1, wherever they are among the device bases
  // [G_0..G_{R-1}, h, Q].  SB200_HOST_WIDE=0 keeps the narrow host tables.
  {
    const char* hw = getenv("SB200_HOST_WIDE");
    const bool want_wide = !(hw && atoi(hw) == 0);
    FixedBase* fbs[7] = {&g->fb_G4[0], &g->fb_G4[1], &g->fb_G4[2], &g->fb_G4[3], &g->fb_h4, &g->fb_Q, &g->fb_h};
    const size_t point_of[7] = {0, 1, 2, 3, 4, (size_t)R, (size_t)R + 1};
    if (want_wide && c < 13) {
      // narrow device windows (one-shot generators): the seven tail generators get a table of their own, 10- or 13-bit
      // windows (9 / 55 MB), built by the same kernels and copied to the host as a whole
      const uint32_t ch = proofs_expected > 0 && proofs_expected <= 4 ? 10 : 13;
      std::vector<uint8_t> xy7(7 * 64);
      for (int k = 0; k < 7; k++) {
        const size_t pi = point_of[k];
        const size_t bi = pi < R ? pi : (pi == R ? (size_t)R + 1 : (size_t)R);  // position among [G..., h, Q]
        if (pi <= (size_t)R + 1) memcpy(&xy7[64 * k], &xy[64 * bi], 64);
        else {  // R < 5: a stream point beyond the bases (gens_3 / gens_4 of a tiny instance)
          Fe zi = fe_invert(pts[pi].Z);
          fe_tobytes(fe_mul(pts[pi].X, zi), &xy7[64 * k]);
          fe_tobytes(fe_mul(pts[pi].Y, zi), &xy7[64 * k + 32]);
        }
      }
      sb::MsmTable t7;
      cudaError_t e7 = sb::msm_table_build(t7, xy7.data(), 7, ch, 0);
      const size_t bytes = t7.entries_per_base() * (size_t)96;
      if (e7 == cudaSuccess) {
        g->wide_store[0].reset(new uint8_t[7 * bytes]);  // uninitialised: a vector would zero-fill what the copy overwrites
        e7 = cudaMemcpy(g->wide_store[0].get(), t7.tbl, 7 * bytes, cudaMemcpyDeviceToHost);
      }
      sb::msm_table_free(t7);
      if (e7 != cudaSuccess) {
        set_error(std::string("host tables of the tail generators: ") + cudaGetErrorString(e7));
        delete g;
        return SB200_ERR_CUDA;
      }
      for (int k = 0; k < 7; k++) fbs[k]->set_wide(g->wide_store[0].get() + (size_t)k * bytes, ch, 252u / ch + 1u);
    } else if (want_wide) {
      const size_t bytes = g->table.entries_per_base() * (size_t)96;  // sizeof(niels_t)
      SB_CUDA(cudaDeviceSynchronize());
      for (int k = 0; k < 7; k++) {
        const size_t pi = point_of[k];
        long di = pi < R ? (long)pi : (pi == R ? (long)R + 1 : (pi == (size_t)R + 1 ? (long)R : -1));
        if (di < 0) continue;
        g->wide_store[k].reset(new uint8_t[bytes]);
        cudaError_t e2 = cudaMemcpy(g->wide_store[k].get(), reinterpret_cast<const uint8_t*>(g->table.tbl) + (size_t)di * bytes, bytes,
                                    cudaMemcpyDeviceToHost);
        if (e2 != cudaSuccess) {
          set_error(std::string("copying a table slice to the host: ") + cudaGetErrorString(e2));
          delete g;
          return SB200_ERR_CUDA;
        }
        fbs[k]->set_wide(g->wide_store[k].get(), g->table.c, g->table.windows);
      }
    }
    for (int k = 0; k < 7; k++)
      if (fbs[k]->wide == nullptr) fbs[k]->build(fbs[k]->point);  // cache-resident 8-bit tables (SB200_HOST_WIDE=0, or a missing slice)
  }
  gmark("host tables of the tail generators");
  *out = reinterpret_cast<sb200_gens*>(g);
  return SB200_OK;
}
void sb200_gens_free(sb200_gens* h) { delete reinterpret_cast<Gens*>(h); }

}  // extern "C"

// =============================================================================== prover
#include "host/comm.hpp"
#include "host/comm_nccl.hpp"
#include "host/zkif.hpp"
#include "host/circ.hpp"
#include "host/pool.hpp"
#include "host/tape.hpp"
#include "prover.inc"

namespace sbh {
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


// synchronous device I/O helpers of the verifier
struct DevIo {
  const Instance& inst;
  Workspace& w;
  sb::ScWork scw;
  explicit DevIo(const Instance& i) : inst(i), w(i.ws) {
    scw.partials = w.sc_partials.p;
    scw.counter = w.counter.p;
    scw.result = w.small.p;
    scw.max_blocks = 148 * 4;
  }
  int fetch(const void* dev, size_t bytes, size_t pinned_off = 0) {
    g_d2h_bytes += bytes;
    SB_CUDA(cudaMemcpyAsync(w.pinned + pinned_off, dev, bytes, cudaMemcpyDeviceToHost, w.stream));
    SB_CUDA(cudaStreamSynchronize(w.stream));
    return SB200_OK;
  }
  int upload_scalars(const FlVec& v, sb::fl_t* dev) {
    if (v.empty()) return SB200_OK;
    if (v.size() * 32 > 4096) { set_error("upload_scalars: too many scalars"); return SB200_ERR_ARG; }
    g_h2d_bytes += v.size() * 32;
    memcpy(w.pinned + 1024, v.data(), v.size() * 32);
    SB_CUDA(cudaMemcpyAsync(dev, w.pinned + 1024, v.size() * 32, cudaMemcpyHostToDevice, w.stream));
    SB_CUDA(cudaStreamSynchronize(w.stream));
    return SB200_OK;
  }
};

}  // namespace sbh

extern "C" {

int sb200_nizk_prove(const sb200_instance* hi, const uint8_t* vars32, uint64_t n_vars, const uint8_t* inputs32,
                     uint64_t n_inputs, const sb200_gens* hg, const sb200_transcript_vt* transcript, const uint8_t* label,
                     size_t label_len, const uint8_t* tape_seed64, const sb200_comm_vt* comm, uint8_t** proof_out,
                     size_t* proof_len) {
  if (!hi || !hg || !proof_out || !proof_len || (!inputs32 && n_inputs)) return SB200_ERR_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  const Instance& inst = *reinterpret_cast<const Instance*>(hi);
  std::lock_guard<std::mutex> inst_lock(inst.mu);  // the handle's workspace, stream and scratch serve one call at a time
  const Gens& gens = *reinterpret_cast<const Gens*>(hg);
  uint8_t seed[64];
  if (tape_seed64) memcpy(seed, tape_seed64, 64);
  else {  // Scalar::random(&mut OsRng), random.rs:13-15
    std::random_device rd;
    for (int i = 0; i < 64; i += 4) { uint32_t v = rd(); memcpy(seed + i, &v, 4); }
  }
  Fl seed_scalar = fl_from_bytes_wide(seed);
  NizkProof pf;
  std::lock_guard<std::mutex> prove_lock(g_prove_mu);  // one proof at a time per process: the host worker pool has a single posting thread
  try {
    if (transcript) {
      VtTranscript tr(transcript);
      rc = prove_impl(inst, vars32, n_vars, inputs32, n_inputs, gens, tr, seed_scalar, comm, pf);
    } else {
      MerlinTranscript tr(label, label_len);
      rc = prove_impl(inst, vars32, n_vars, inputs32, n_inputs, gens, tr, seed_scalar, comm, pf);
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
  std::lock_guard<std::mutex> inst_lock(inst.mu);  // the handle's workspace, stream and scratch serve one call at a time
  if (n_vars > inst.num_vars) return SB200_ERR_ARG;
  rc = workspace_init(inst, *reinterpret_cast<const Gens*>(hg));
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

// ---- zkinterface ingestion (spartan-zkinterface/src/lib.rs:75-260); host only ---------------
struct sb200_r1cs { ZkR1cs r; };
int sb200_zkif_parse(const uint8_t* circuit, size_t circuit_len, const uint8_t* header, size_t header_len, const uint8_t* witness,
                     size_t witness_len, sb200_r1cs** out) {
  if (!circuit || !header || !witness || !out) return SB200_ERR_ARG;
  *out = nullptr;
  sb200_r1cs* r = new sb200_r1cs();
  bool ok = false;
  try {
    ok = r->r.parse(circuit, circuit_len, header, header_len, witness, witness_len, ingest_threads(circuit_len / 64));
  } catch (const std::exception& e) {
    r->r.error = std::string("exception: ") + e.what();
  }
  if (!ok) {
    set_error("zkif: " + r->r.error);
    delete r;
    return SB200_ERR_ZKIF;
  }
  *out = r;
  return SB200_OK;
}
int sb200_r1cs_dims(const sb200_r1cs* r, uint64_t out[7]) {
  if (!r || !out) return SB200_ERR_ARG;
  out[0] = r->r.num_cons; out[1] = r->r.num_vars; out[2] = r->r.num_inputs;
  for (int m = 0; m < 3; m++) out[3 + m] = r->r.rows[m].size();
  out[6] = r->r.non_zero_entries;
  return SB200_OK;
}
int sb200_r1cs_matrix(const sb200_r1cs* r, int m, const uint64_t** rows, const uint64_t** cols, const uint8_t** vals32) {
  if (!r || m < 0 || m > 2 || !rows || !cols || !vals32) return SB200_ERR_ARG;
  *rows = r->r.rows[m].data(); *cols = r->r.cols[m].data(); *vals32 = r->r.vals[m].data();
  return SB200_OK;
}
const uint8_t* sb200_r1cs_vars(const sb200_r1cs* r) { return r ? r->r.vars.data() : nullptr; }
const uint8_t* sb200_r1cs_inputs(const sb200_r1cs* r) { return r ? r->r.inputs.data() : nullptr; }
int sb200_r1cs_instance(const sb200_r1cs* r, int rank, int world, sb200_instance** out) {
  if (!r || !out) return SB200_ERR_ARG;
  const uint64_t* rows[3]; const uint64_t* cols[3]; const uint8_t* vals[3]; uint64_t nnz[3];
  for (int m = 0; m < 3; m++) { rows[m] = r->r.rows[m].data(); cols[m] = r->r.cols[m].data(); vals[m] = r->r.vals[m].data(); nnz[m] = r->r.rows[m].size(); }
  return sb200_instance_new_sharded(r->r.num_cons, r->r.num_vars, r->r.num_inputs, rows, cols, vals, nnz, rank, world, out);
}
void sb200_r1cs_free(sb200_r1cs* r) { delete r; }

// ---- rust-circ adapter (rust-circ/src/target/r1cs/spartan.rs:31-186); host only -------------
int sb200_circ_r1cs(uint64_t n_signals, const uint64_t* signal_ids, const uint8_t* names, const uint64_t* name_off,
                    const uint8_t* value_bytes, const uint64_t* value_off, uint64_t n_public, uint64_t next_idx, uint64_t n_cons,
                    const uint64_t* const lc_off[3], const uint64_t* const mono_ids[3], const uint8_t* const coeff_bytes[3],
                    const uint64_t* const coeff_off[3], const uint8_t* const const_bytes[3], const uint64_t* const const_off[3],
                    sb200_r1cs** out) {
  if (!out || !name_off || !value_off || !lc_off || !mono_ids || !coeff_off || !const_off) return SB200_ERR_ARG;
  if (n_signals && (!signal_ids || !names || !value_bytes)) return SB200_ERR_ARG;
  *out = nullptr;
  CircLcs lcs[3];
  for (int m = 0; m < 3; m++) {
    if (!lc_off[m] || !coeff_off[m] || !const_off[m]) return SB200_ERR_ARG;
    lcs[m] = CircLcs{lc_off[m], mono_ids[m], CircBigInts{coeff_bytes ? coeff_bytes[m] : nullptr, coeff_off[m]},
                     CircBigInts{const_bytes ? const_bytes[m] : nullptr, const_off[m]}};
  }
  sb200_r1cs* r = new sb200_r1cs();
  bool ok = false;
  try {
    ok = circ_r1cs_to_spartan(n_signals, signal_ids, names, name_off, CircBigInts{value_bytes, value_off}, n_public, next_idx, n_cons,
                              lcs, r->r);
  } catch (const std::exception& e) {
    r->r.error = std::string("exception: ") + e.what();
  }
  if (!ok) {
    set_error("circ: " + r->r.error);
    delete r;
    return SB200_ERR_ASSERT;
  }
  *out = r;
  return SB200_OK;
}
int sb200_circ_int_to_scalar(const uint8_t* magnitude_le, size_t len, uint8_t out32[32]) {
  if ((!magnitude_le && len) || !out32) return SB200_ERR_ARG;
  fl_to_bytes(circ_int_to_scalar(magnitude_le, len), out32);
  return SB200_OK;
}

// ---- exchange for sharded proofs ------------------------------------------------------------
int sb200_comm_shm_open(const char* name, int rank, int world, sb200_comm_vt* out) {
  if (!name || !out) return SB200_ERR_ARG;
  ShmComm* c = new ShmComm();
  int rc = c->open(name, rank, world);
  if (rc) {
    set_error("sb200_comm_shm_open failed (" + std::to_string(rc) + ")");
    delete c;
    return SB200_ERR_COMM;
  }
  out->ctx = c;
  out->rank = rank;
  out->world = world;
  out->allgather = shm_allgather_thunk;
  return SB200_OK;
}
// ---- native NCCL exchange (host/comm_nccl.hpp)
int sb200_comm_nccl_unique_id(uint8_t id128[128]) {
  std::string err;
  if (!id128) return SB200_ERR_ARG;
  if (!NcclApi::get().load(err)) { set_error(err); return SB200_ERR_COMM; }
  NcclApi::UniqueId id;
  if (NcclApi::get().GetUniqueId(&id) != 0) { set_error("ncclGetUniqueId failed"); return SB200_ERR_COMM; }
  memcpy(id128, id.internal, 128);
  return SB200_OK;
}
int sb200_comm_nccl_open(const uint8_t id128[128], int rank, int world, sb200_comm_vt* out) {
  if (!id128 || !out || world < 1 || rank < 0 || rank >= world) return SB200_ERR_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  std::string err;
  if (!NcclApi::get().load(err)) { set_error(err); return SB200_ERR_COMM; }
  NcclComm* c = new NcclComm();
  c->rank = rank;
  c->world = world;
  NcclApi::UniqueId id;
  memcpy(id.internal, id128, 128);
  cudaError_t ce = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (ce == cudaSuccess) ce = cudaMalloc((void**)&c->d_send, NcclComm::kMaxBytes);
  if (ce == cudaSuccess) ce = cudaMalloc((void**)&c->d_recv, NcclComm::kMaxBytes * (size_t)world);
  if (ce == cudaSuccess) ce = cudaMallocHost((void**)&c->h_stage, NcclComm::kMaxBytes * ((size_t)world + 1));
  if (ce != cudaSuccess) { set_error(std::string("nccl exchange buffers: ") + cudaGetErrorString(ce)); delete c; return SB200_ERR_CUDA; }
  int nrc = NcclApi::get().CommInitRank(&c->comm, world, id, rank);
  if (nrc != 0) {
    set_error(std::string("ncclCommInitRank: ") + (NcclApi::get().GetErrorString ? NcclApi::get().GetErrorString(nrc) : "error"));
    delete c;
    return SB200_ERR_COMM;
  }
  out->ctx = c;
  out->rank = rank;
  out->world = world;
  out->allgather = nccl_allgather_thunk;
  return SB200_OK;
}
void sb200_comm_nccl_close(sb200_comm_vt* comm) {
  if (!comm || !comm->ctx) return;
  delete static_cast<NcclComm*>(comm->ctx);
  comm->ctx = nullptr;
}
void sb200_comm_shm_close(sb200_comm_vt* comm) {
  if (!comm || !comm->ctx) return;
  ShmComm* c = static_cast<ShmComm*>(comm->ctx);
  c->close();
  delete c;
  comm->ctx = nullptr;
}
// Host-only check of an exchange hook (no GPU work): `rounds` all-gathers of three scalars per rank, summed mod l the
// way a sharded sum-check round does and absorbed into a transcript; digest32_out must be the same on every rank
// and equal to sb200_comm_selftest_expected(world, rounds).
static Fl selftest_scalar(int rank, int round, int k) {
  uint8_t wide[64];
  for (int i = 0; i < 64; i++) wide[i] = (uint8_t)(i * 131 + rank * 17 + round * 29 + k * 43 + 1);
  return fl_from_bytes_wide(wide);
}
int sb200_comm_selftest(const sb200_comm_vt* comm, int rounds, uint8_t digest32_out[32]) {
  if (!comm || !comm->allgather || !digest32_out || comm->world < 1) return SB200_ERR_ARG;
  MerlinTranscript tr("comm selftest", 13);
  std::vector<uint8_t> recv((size_t)comm->world * 96);
  for (int j = 0; j < rounds; j++) {
    Fl mine[3];
    for (int k = 0; k < 3; k++) mine[k] = selftest_scalar(comm->rank, j, k);
    if (comm->allgather(comm->ctx, mine, 96, recv.data()) != 0) { set_error("allgather failed"); return SB200_ERR_COMM; }
    for (int k = 0; k < 3; k++) {
      Fl sum = fl_zero();
      for (int q = 0; q < comm->world; q++) sum = fl_add(sum, from_dev_bytes(&recv[96 * (size_t)q + 32 * k]));
      tr.append_scalar("sum", sum);
    }
  }
  tr.challenge_bytes("digest", digest32_out, 32);
  return SB200_OK;
}
int sb200_comm_selftest_expected(int world, int rounds, uint8_t digest32_out[32]) {
  if (world < 1 || !digest32_out) return SB200_ERR_ARG;
  MerlinTranscript tr("comm selftest", 13);
  for (int j = 0; j < rounds; j++)
    for (int k = 0; k < 3; k++) {
      Fl sum = fl_zero();
      for (int q = 0; q < world; q++) sum = fl_add(sum, selftest_scalar(q, j, k));
      tr.append_scalar("sum", sum);
    }
  tr.challenge_bytes("digest", digest32_out, 32);
  return SB200_OK;
}

int sb200_last_prove_profile(double out[16]) {
  for (int i = 0; i < 16; i++) out[i] = g_profile[i];
  return SB200_OK;
}

}  // extern "C"

#include "verify.inc"
#include "kernel_hooks.inc"
#include "snark.inc"
#include "snark_prove.inc"
