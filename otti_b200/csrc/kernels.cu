// sm_100a kernels for the Spartan NIZK prover hot loops.
// All arithmetic is exact 253/255-bit modular integer work on the IMAD pipe -- no
// tensor cores by design.  Kernel <-> reference map:
//   k_to_mont          Assignment::new / Scalar::from_bytes        lib.rs:65-89
//   k_eq_table         EqPolynomial::evals                         dense_mlpoly.rs:67-83
//   k_spmv3            R1CSInstance::multiply_vec                  r1csinstance.rs:248-262, sparse_mlpoly.rs:452-466
//   k_sc_cubic         prove_cubic_with_additive_term round + 4x bound_poly_var_top
//                                                                  sumcheck.rs:618-674, dense_mlpoly.rs:214-221
//   k_spmvT3_combine   compute_eval_table_sparse x3 + r_A/r_B/r_C  sparse_mlpoly.rs:468-483, r1csproof.rs:268-279
//   k_sc_quad          prove_quad round + 2x bound_poly_var_top    sumcheck.rs:453-484
//   k_bound_L          DensePolynomial::bound                      dense_mlpoly.rs:205-212
//   k_dot_small        DotProductProofLog::compute_dotproduct      nizk/mod.rs:430-433
//   k_bullet_step      BulletReductionProof::prove scalar side     nizk/bullet.rs:71-118
//   msm_fixed          [Scalar]::commit -> vartime_multiscalar_mul commitments.rs:79-84, dense_mlpoly.rs:163-176
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "dev.hpp"
#include "ge.cuh"

namespace sb {

static std::atomic<uint64_t> g_launches{0};
uint64_t launch_count() { return g_launches.load(); }
void launch_count_reset() { g_launches.store(0); }
#define SB_LAUNCHED() (g_launches.fetch_add(1), cudaGetLastError())

// Per-kernel device time (bench.py: which kernel dominates a step, and for how long): when switched on, every launch
// of this file is bracketed by two CUDA events on its stream.  Off during every timed region -- the brackets cost a few
// microseconds per launch; the totals come from a separate profiling pass over the same workload.
struct KProfRec { const char* name; cudaEvent_t a, b; };
static std::atomic<int> g_kprof_on{0};
static std::mutex g_kprof_mu;
static std::vector<KProfRec> g_kprof_recs;
static std::vector<cudaEvent_t> g_kprof_free;
struct KProf {
  cudaStream_t s_;
  cudaEvent_t a_ = nullptr, b_ = nullptr;
  const char* name_;
  static cudaEvent_t get() {
    std::lock_guard<std::mutex> lk(g_kprof_mu);
    if (!g_kprof_free.empty()) { cudaEvent_t e = g_kprof_free.back(); g_kprof_free.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
  }
  KProf(const char* name, cudaStream_t s) : s_(s), name_(name) {
    if (!g_kprof_on.load(std::memory_order_relaxed)) return;
    a_ = get(); b_ = get();
    cudaEventRecord(a_, s_);
  }
  ~KProf() {
    if (!a_) return;
    cudaEventRecord(b_, s_);
    std::lock_guard<std::mutex> lk(g_kprof_mu);
    g_kprof_recs.push_back(KProfRec{name_, a_, b_});
  }
};
void kprof_enable(bool on) { g_kprof_on.store(on ? 1 : 0); }
// "name,launches,total_ms\n" per kernel, sorted by total time; clears the records
std::string kprof_report() {
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lk(g_kprof_mu);
  std::vector<std::pair<std::string, std::pair<uint64_t, double>>> tot;
  for (const KProfRec& r : g_kprof_recs) {
    float ms = 0;
    cudaEventElapsedTime(&ms, r.a, r.b);
    size_t i = 0;
    for (; i < tot.size(); i++) if (tot[i].first == r.name) break;
    if (i == tot.size()) tot.push_back({r.name, {0, 0.0}});
    tot[i].second.first++;
    tot[i].second.second += ms;
    g_kprof_free.push_back(r.a);
    g_kprof_free.push_back(r.b);
  }
  g_kprof_recs.clear();
  std::sort(tot.begin(), tot.end(), [](const auto& x, const auto& y) { return x.second.second > y.second.second; });
  std::string out;
  char line[256];
  for (auto& t : tot) {
    snprintf(line, sizeof line, "%s,%llu,%.6f\n", t.first.c_str(), (unsigned long long)t.second.first, t.second.second);
    out += line;
  }
  return out;
}

static constexpr int kSMs = 148;  // B200

// ------------------------------------------------------------------ 128-bit memory helpers
__device__ __forceinline__ fl_t ld_fl(const fl_t* p) {
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 a = q[0], b = q[1];
  fl_t r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ fl_t ldg_fl(const fl_t* p) {  // read-only path
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 a = __ldg(q), b = __ldg(q + 1);
  fl_t r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ fl_t ldcg_fl(const fl_t* p) {  // L2 only (cross-block visibility)
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 a = __ldcg(q), b = __ldcg(q + 1);
  fl_t r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ void st_fl(fl_t* p, const fl_t& x) {
  uint4* q = reinterpret_cast<uint4*>(p);
  q[0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]);
  q[1] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
}
__device__ __forceinline__ fp_t ldg_fp(const fp_t* p) {
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 a = __ldg(q), b = __ldg(q + 1);
  fp_t r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ void st_fp(fp_t* p, const fp_t& x) {
  uint4* q = reinterpret_cast<uint4*>(p);
  q[0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]);
  q[1] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
}

// ------------------------------------------------------------------ reductions over F_l
__device__ __forceinline__ fl_t warp_sum_fl(fl_t x) {
#pragma unroll 1
  for (int o = 16; o > 0; o >>= 1) {
    fl_t y;
#pragma unroll
    for (int i = 0; i < 8; i++) y.v[i] = __shfl_down_sync(0xffffffffu, x.v[i], o);
    x = fl_add(x, y);
  }
  return x;
}
// Sum NV values per thread over the block; valid in thread 0.  smem: fl_t[NV * 32].
template <int NV>
__device__ __forceinline__ void block_sum_fl(fl_t (&x)[NV], fl_t* smem) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; k++) x[k] = warp_sum_fl(x[k]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) smem[wid * NV + k] = x[k];
  }
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) {
      fl_t v = lane < nw ? smem[lane * NV + k] : fl_zero();
      x[k] = warp_sum_fl(v);
    }
  }
}
// Grid-wide sum by "last block done": every block deposits NV partials, the last one to
// arrive adds them up (F_l addition is exact, so arrival order cannot change a bit).
struct Mailbox {  // see ScWork
  fl_t* host_result;
  uint32_t* host_flag;
  uint32_t seq;
};
template <int NV>
__device__ __forceinline__ void grid_sum_fl(fl_t (&x)[NV], fl_t* smem, fl_t* partials, unsigned int* counter, fl_t* result,
                                            Mailbox mb = Mailbox{nullptr, nullptr, 0}) {
  __shared__ bool s_last;
  block_sum_fl<NV>(x, smem);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) st_fl(&partials[(size_t)blockIdx.x * NV + k], x[k]);
    __threadfence();
    unsigned int t = atomicAdd(counter, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    fl_t acc[NV];
#pragma unroll
    for (int k = 0; k < NV; k++) acc[k] = fl_zero();
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
#pragma unroll
      for (int k = 0; k < NV; k++) acc[k] = fl_add(acc[k], ldcg_fl(&partials[(size_t)b * NV + k]));
    }
    block_sum_fl<NV>(acc, smem);
    if (threadIdx.x == 0) {
#pragma unroll
      for (int k = 0; k < NV; k++) st_fl(&result[k], acc[k]);
      *counter = 0;
      if (mb.host_flag != nullptr) {
#pragma unroll
        for (int k = 0; k < NV; k++) st_fl(&mb.host_result[k], acc[k]);
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(mb.host_flag) = mb.seq;
      }
    }
  }
}

// ------------------------------------------------------------------ elementwise
// out2 (optional): a second copy of the result -- the witness in Montgomery form and the lower half of z are the same values
__global__ void to_mont_kernel(const fl_t* __restrict__ in, fl_t* __restrict__ out, fl_t* __restrict__ out2, size_t n, uint32_t* flag) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  bool bad = false;
  for (; i < n; i += stride) {
    fl_t x = ldg_fl(in + i);
    bad |= !fl_is_canonical(x);
    const fl_t m = fl_to_mont(x);
    st_fl(out + i, m);
    if (out2 != nullptr) st_fl(out2 + i, m);
  }
  if (bad) atomicOr(flag, 1u);
}
__global__ void from_mont_kernel(const fl_t* __restrict__ in, fl_t* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) st_fl(out + i, fl_from_mont(ldg_fl(in + i)));
}
__global__ void fl_binop_kernel(int op, const fl_t* a, const fl_t* b, fl_t* out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    fl_t x = ld_fl(a + i), y = ld_fl(b + i);
    st_fl(out + i, op == 0 ? fl_add(x, y) : op == 1 ? fl_sub(x, y) : fl_mul(x, y));
  }
}
static inline unsigned grid_for(size_t n, int block, int per_sm = 8) {
  size_t g = (n + block - 1) / block;
  size_t cap = (size_t)kSMs * per_sm;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}
__global__ void fp_binop_kernel(int op, const fp_t* a, const fp_t* b, fp_t* out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    fp_t x = ldg_fp(a + i), y = ldg_fp(b + i);
    fp_t r = op == 0 ? fp_add(x, y) : op == 1 ? fp_sub(x, y) : op == 2 ? fp_mul(x, y) : op == 3 ? fp_sqr(x) : fp_invert(x);
    st_fp(out + i, fp_canon(r));
  }
}
cudaError_t k_fp_binop(int op, const fp_t* a, const fp_t* b, fp_t* out, size_t n, cudaStream_t s) {
  if (!n) return cudaSuccess;
  { KProf kp_("fp_binop_kernel", s); fp_binop_kernel<<<grid_for(n, 128), 128, 0, s>>>(op, a, b, out, n); }
  return SB_LAUNCHED();
}
cudaError_t k_to_mont(const fl_t* in, fl_t* out, size_t n, uint32_t* flag, cudaStream_t s, fl_t* out2) {
  if (!n) return cudaSuccess;
  { KProf kp_("to_mont_kernel", s); to_mont_kernel<<<grid_for(n, 256), 256, 0, s>>>(in, out, out2, n, flag); }
  return SB_LAUNCHED();
}
cudaError_t k_from_mont(const fl_t* in, fl_t* out, size_t n, cudaStream_t s) {
  if (!n) return cudaSuccess;
  { KProf kp_("from_mont_kernel", s); from_mont_kernel<<<grid_for(n, 256), 256, 0, s>>>(in, out, n); }
  return SB_LAUNCHED();
}
cudaError_t k_fl_binop(int op, const fl_t* a, const fl_t* b, fl_t* out, size_t n, cudaStream_t s) {
  if (!n) return cudaSuccess;
  { KProf kp_("fl_binop_kernel", s); fl_binop_kernel<<<grid_for(n, 256), 256, 0, s>>>(op, a, b, out, n); }
  return SB_LAUNCHED();
}

// flag |= 1 if Az[i]*Bz[i] != Cz[i] for some i        (r1csinstance.rs:238-245)
__global__ void check_sat_kernel(const fl_t* __restrict__ A, const fl_t* __restrict__ B, const fl_t* __restrict__ C, size_t n,
                                 uint32_t* flag) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  bool bad = false;
  for (; i < n; i += stride) bad |= !fl_eq(fl_mul(ld_fl(A + i), ld_fl(B + i)), ld_fl(C + i));
  if (bad) atomicOr(flag, 1u);
}
cudaError_t k_check_sat(const fl_t* A, const fl_t* B, const fl_t* C, size_t n, uint32_t* flag, cudaStream_t s) {
  if (!n) return cudaSuccess;
  { KProf kp_("check_sat_kernel", s); check_sat_kernel<<<grid_for(n, 256), 256, 0, s>>>(A, B, C, n, flag); }
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ eq table
// E[i] = prod_j (bit_{ell-1-j}(i) ? r[j] : 1 - r[j])  -- thread per entry, used for the two
// half-size factor tables (<= 2^12 entries each).
// `init` (Montgomery) multiplies every entry: 1 for a plain table; a sharded prover passes the factor
// contributed by the index bits its rank fixes.
__global__ void eq_small_kernel(const fl_t* __restrict__ r, uint32_t ell, fl_t init, fl_t* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (1u << ell)) return;
  fl_t acc = init, one = fl_one();
  for (uint32_t j = 0; j < ell; j++) {
    fl_t rj = ldg_fl(r + j);
    bool bit = (i >> (ell - 1 - j)) & 1u;
    acc = fl_mul(acc, bit ? rj : fl_sub(one, rj));
  }
  st_fl(out + i, acc);
}
// out[(ih << b) | il] = hi[ih] * lo[il]: one multiplication and one 32-byte store per entry.
__global__ void eq_outer_kernel(const fl_t* __restrict__ hi, const fl_t* __restrict__ lo, uint32_t b, size_t n,
                                fl_t* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  const size_t mask = ((size_t)1 << b) - 1;
  for (; i < n; i += stride) st_fl(out + i, fl_mul(ldg_fl(hi + (i >> b)), ldg_fl(lo + (i & mask))));
}
cudaError_t k_eq_table_scaled(const fl_t* r, uint32_t ell, fl_t init, fl_t* out, fl_t* tmp, cudaStream_t s) {
  if (ell <= 8) {
    { KProf kp_("eq_small_kernel", s); eq_small_kernel<<<((1u << ell) + 127) / 128, 128, 0, s>>>(r, ell, init, out); }
    return SB_LAUNCHED();
  }
  uint32_t a = ell / 2, b = ell - a;  // high a variables (r[0..a)), low b variables
  fl_t* hi = tmp;
  fl_t* lo = tmp + ((size_t)1 << a);
  { KProf kp_("eq_small_kernel", s); eq_small_kernel<<<((1u << a) + 127) / 128, 128, 0, s>>>(r, a, init, hi); }
  cudaError_t e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  { KProf kp_("eq_small_kernel", s); eq_small_kernel<<<((1u << b) + 127) / 128, 128, 0, s>>>(r + a, b, fl_one(), lo); }
  e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  size_t n = (size_t)1 << ell;
  { KProf kp_("eq_outer_kernel", s); eq_outer_kernel<<<grid_for(n, 256), 256, 0, s>>>(hi, lo, b, n, out); }
  return SB_LAUNCHED();
}
cudaError_t k_eq_table(const fl_t* r, uint32_t ell, fl_t* out, fl_t* tmp, cudaStream_t s) {
  return k_eq_table_scaled(r, ell, fl_one(), out, tmp, s);
}
// out[i] = in[i * step + first]: a rank's cyclic slice of a full-length table
__global__ void gather_stride_kernel(const fl_t* __restrict__ in, size_t n, size_t step, size_t first, fl_t* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) st_fl(out + i, ldg_fl(in + i * step + first));
}
cudaError_t k_gather_stride(const fl_t* in, size_t n, size_t step, size_t first, fl_t* out, cudaStream_t s) {
  if (!n) return cudaSuccess;
  { KProf kp_("gather_stride_kernel", s); gather_stride_kernel<<<grid_for(n, 256), 256, 0, s>>>(in, n, step, first, out); }
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ sparse mat-vec
// R1CS matrices out of a circuit compiler are mostly +1 / -1 (copy constraints, booleanity, bit decompositions):
// those entries cost an addition instead of a Montgomery product (136 IMAD.WIDE).
__device__ __forceinline__ fl_t fl_minus_one() {  // l - R
  fl_t r;
  r.v[0] = 0xcf5d3ed0u; r.v[1] = 0x812631a5u; r.v[2] = 0x2f79cd65u; r.v[3] = 0x4def9deau;
  r.v[4] = 0x00000001u; r.v[5] = 0x00000000u; r.v[6] = 0x00000000u; r.v[7] = 0x00000000u;
  return r;
}
__device__ __forceinline__ fl_t coef_times(const fl_t& val, const fl_t& x) {
  if (fl_eq(val, fl_one())) return x;
  if (fl_eq(val, fl_minus_one())) return fl_sub(fl_zero(), x);
  return fl_mul_call(val, x);
}
// sum over the entries [k0, k1) of one segment, by the thread that owns the segment: short segments are walked, long
// ones were summed by sp_long_kernel / sp_big_kernel (dev.hpp LongSegs)
__device__ __forceinline__ fl_t seg_sum(const uint32_t* __restrict__ idx, const fl_t* __restrict__ val, const LongSegs& lg,
                                        const fl_t* __restrict__ x, uint32_t seg, uint32_t k0, uint32_t k1) {
  fl_t acc = fl_zero();
  if (k1 - k0 <= kLongSeg) {
    for (uint32_t k = k0; k < k1; k++) acc = fl_add(acc, coef_times(ldg_fl(val + k), ldg_fl(x + __ldg(idx + k))));
    return acc;
  }
  uint32_t lo = 0, hi = lg.nlong;  // position of seg in the ascending list of long segments
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (__ldg(lg.seg + mid) <= seg) lo = mid; else hi = mid;
  }
  const uint32_t c0 = __ldg(lg.first + lo), c1 = __ldg(lg.first + lo + 1);
  if (c1 - c0 > kSumInline) return ldcg_fl(lg.reduced + lo);
  for (uint32_t c = c0; c < c1; c++) acc = fl_add(acc, ldcg_fl(lg.partial + c));
  return acc;
}
// one warp per chunk of a long segment (grid.y = matrix)
__global__ void __launch_bounds__(256) sp_long_kernel(Csr mat, const fl_t* __restrict__ x) {
  const int m = blockIdx.y;
  const LongSegs& lg = mat.lg[m];
  const uint32_t lane = threadIdx.x & 31;
  for (uint32_t c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < lg.nchunks; c += gridDim.x * (blockDim.x >> 5)) {
    const uint32_t k0 = __ldg(lg.chunk_lo + c), k1 = __ldg(lg.chunk_hi + c);
    const uint32_t* __restrict__ idx = mat.idx[m];
    const fl_t* __restrict__ val = mat.val[m];
    fl_t acc = fl_zero();
    for (uint32_t k = k0 + lane; k < k1; k += 32) acc = fl_add(acc, coef_times(ldg_fl(val + k), ldg_fl(x + __ldg(idx + k))));
    acc = warp_sum_fl(acc);
    if (lane == 0) st_fl(lg.partial + c, acc);
  }
}
// one warp per segment of more than kSumInline chunks: reduced[j] = sum of its chunk sums
__global__ void __launch_bounds__(256) sp_big_kernel(Csr mat) {
  const int m = blockIdx.y;
  const LongSegs& lg = mat.lg[m];
  const uint32_t lane = threadIdx.x & 31;
  for (uint32_t b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); b < lg.nbig; b += gridDim.x * (blockDim.x >> 5)) {
    const uint32_t j = __ldg(lg.big + b), c0 = __ldg(lg.first + j), c1 = __ldg(lg.first + j + 1);
    fl_t acc = fl_zero();
    for (uint32_t c = c0 + lane; c < c1; c += 32) acc = fl_add(acc, ldcg_fl(lg.partial + c));
    acc = warp_sum_fl(acc);
    if (lane == 0) st_fl(lg.reduced + j, acc);
  }
}
static cudaError_t sp_long_pass(const Csr& mat, const fl_t* x, cudaStream_t s) {
  uint32_t maxc = 0, maxb = 0;
  for (int m = 0; m < 3; m++) { maxc = std::max(maxc, mat.lg[m].nchunks); maxb = std::max(maxb, mat.lg[m].nbig); }
  if (!maxc) return cudaSuccess;
  dim3 grid(grid_for(maxc, 8, 16), 3);  // 8 chunks per block
  { KProf kp_("sp_long_kernel", s); sp_long_kernel<<<grid, 256, 0, s>>>(mat, x); }
  cudaError_t e = SB_LAUNCHED();
  if (e != cudaSuccess || !maxb) return e;
  dim3 gridb(grid_for(maxb, 8, 16), 3);
  { KProf kp_("sp_big_kernel", s); sp_big_kernel<<<gridb, 256, 0, s>>>(mat); }
  return SB_LAUNCHED();
}
__global__ void spmv3_kernel(Csr csr, uint32_t num_rows, const fl_t* __restrict__ z, fl_t* __restrict__ yA,
                             fl_t* __restrict__ yB, fl_t* __restrict__ yC) {
  const int m = blockIdx.y;
  const uint32_t* __restrict__ ptr = csr.ptr[m];
  fl_t* y = m == 0 ? yA : m == 1 ? yB : yC;
  for (uint32_t row = blockIdx.x * blockDim.x + threadIdx.x; row < num_rows; row += gridDim.x * blockDim.x) {
    const uint32_t k0 = __ldg(ptr + row), k1 = __ldg(ptr + row + 1);
    st_fl(y + row, seg_sum(csr.idx[m], csr.val[m], csr.lg[m], z, row, k0, k1));
  }
}
cudaError_t k_spmv3(const Csr& csr, uint32_t num_rows, const fl_t* z, fl_t* yA, fl_t* yB, fl_t* yC, cudaStream_t s) {
  cudaError_t e = sp_long_pass(csr, z, s);
  if (e != cudaSuccess) return e;
  dim3 grid(grid_for(num_rows, 128), 3);
  { KProf kp_("spmv3_kernel", s); spmv3_kernel<<<grid, 128, 0, s>>>(csr, num_rows, z, yA, yB, yC); }
  return SB_LAUNCHED();
}
__global__ void spmvT3_kernel(Csr csc, uint32_t num_cols, const fl_t* __restrict__ e, fl_t rA, fl_t rB, fl_t rC,
                              fl_t* __restrict__ out) {
  for (uint32_t col = blockIdx.x * blockDim.x + threadIdx.x; col < num_cols; col += gridDim.x * blockDim.x) {
    // the six segment bounds are requested together; most columns of a padded z are empty in all three matrices
    uint32_t k0[3], k1[3];
#pragma unroll
    for (int m = 0; m < 3; m++) { k0[m] = __ldg(csc.ptr[m] + col); k1[m] = __ldg(csc.ptr[m] + col + 1); }
    // A column's three segments are independent chains of dependent loads (index -> value and gathered e -> product).
    // Walked one after the other they leave one request in flight per thread; here the first entry of every short
    // segment is fetched for all three matrices at once (the synthetic shape and most circuit columns have one or two
    // entries), and only what is left goes through the sequential walk.
    uint32_t i0[3];
    fl_t v0[3], e0[3];
    bool first[3];
#pragma unroll
    for (int m = 0; m < 3; m++) {
      first[m] = k1[m] != k0[m] && k1[m] - k0[m] <= kLongSeg;
      if (first[m]) { i0[m] = __ldg(csc.idx[m] + k0[m]); v0[m] = ldg_fl(csc.val[m] + k0[m]); }
    }
#pragma unroll
    for (int m = 0; m < 3; m++)
      if (first[m]) e0[m] = ldg_fl(e + i0[m]);
    fl_t tot = fl_zero();
#pragma unroll
    for (int m = 0; m < 3; m++) {
      if (k0[m] == k1[m]) continue;
      fl_t acc;
      if (first[m]) {
        acc = coef_times(v0[m], e0[m]);
        if (k1[m] - k0[m] > 1) acc = fl_add(acc, seg_sum(csc.idx[m], csc.val[m], csc.lg[m], e, col, k0[m] + 1, k1[m]));
      } else {
        acc = seg_sum(csc.idx[m], csc.val[m], csc.lg[m], e, col, k0[m], k1[m]);
      }
      tot = fl_add(tot, fl_mul_call(m == 0 ? rA : m == 1 ? rB : rC, acc));
    }
    st_fl(out + col, tot);
  }
}
cudaError_t k_spmvT3_combine(const Csr& csc, uint32_t num_cols, const fl_t* e, fl_t rA, fl_t rB, fl_t rC, fl_t* out,
                             cudaStream_t s) {
  cudaError_t err = sp_long_pass(csc, e, s);
  if (err != cudaSuccess) return err;
  { KProf kp_("spmvT3_kernel", s); spmvT3_kernel<<<grid_for(num_cols, 128), 128, 0, s>>>(csc, num_cols, e, rA, rB, rC, out); }
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ sum-check rounds
template <class M>
__device__ __forceinline__ fl_t bind1(const fl_t& lo, const fl_t& hi, const fl_t& r) {
  return fl_add(lo, M::mul(r, fl_sub(hi, lo)));  // Z[i] + r*(Z[i+n] - Z[i]), dense_mlpoly.rs:217
}
// M = FlInl for the long rounds (the loop body is reused), FlCall for the short ones, where a round is
// one pass over ~14 multiplications per thread and the kernel must stay inside the instruction cache.
template <bool BIND, class M>
__global__ void __launch_bounds__(256, 2) sc_cubic_kernel(fl_t* __restrict__ T, fl_t* __restrict__ A, fl_t* __restrict__ B,
                                                       fl_t* __restrict__ C, size_t newlen, fl_t r, fl_t* partials,
                                                       unsigned int* counter, fl_t* result, Mailbox mb) {
  __shared__ fl_t smem[3 * 32];
  const size_t h = newlen >> 1;
  fl_t e[3];
  e[0] = e[1] = e[2] = fl_zero();
  fl_t* tabs[4] = {T, A, B, C};
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < h; i += (size_t)gridDim.x * blockDim.x) {
    fl_t lo[4], hi[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      fl_t* X = tabs[k];
      if (BIND) {
        lo[k] = bind1<M>(ld_fl(X + i), ld_fl(X + i + newlen), r);
        hi[k] = bind1<M>(ld_fl(X + i + h), ld_fl(X + i + h + newlen), r);
        st_fl(X + i, lo[k]);
        st_fl(X + i + h, hi[k]);
      } else {
        lo[k] = ld_fl(X + i);
        hi[k] = ld_fl(X + i + h);
      }
    }
    // comb(T,A,B,C) = T*(A*B - C) at the points 0, 2, 3 (sumcheck.rs:622-650, r1csproof.rs:84-88)
    e[0] = fl_add(e[0], M::mul(lo[0], fl_sub(M::mul(lo[1], lo[2]), lo[3])));
    fl_t d[4], p[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      d[k] = fl_sub(hi[k], lo[k]);
      p[k] = fl_add(hi[k], d[k]);  // 2*hi - lo
    }
    e[1] = fl_add(e[1], M::mul(p[0], fl_sub(M::mul(p[1], p[2]), p[3])));
#pragma unroll
    for (int k = 0; k < 4; k++) p[k] = fl_add(p[k], d[k]);  // 3*hi - 2*lo
    e[2] = fl_add(e[2], M::mul(p[0], fl_sub(M::mul(p[1], p[2]), p[3])));
  }
  if (BIND && newlen == 1 && blockIdx.x == 0 && threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 4; k++) st_fl(tabs[k], bind1<M>(ld_fl(tabs[k]), ld_fl(tabs[k] + 1), r));
  }
  grid_sum_fl<3>(e, smem, partials, counter, result, mb);
}
template <bool BIND, class M>
__global__ void __launch_bounds__(256, 2) sc_quad_kernel(fl_t* __restrict__ Z, fl_t* __restrict__ Q, size_t newlen, fl_t r,
                                                      fl_t* partials, unsigned int* counter, fl_t* result, Mailbox mb) {
  __shared__ fl_t smem[3 * 32];
  const size_t h = newlen >> 1;
  fl_t e[3];
  e[0] = e[1] = e[2] = fl_zero();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < h; i += (size_t)gridDim.x * blockDim.x) {
    fl_t zl, zh, ql, qh;
    if (BIND) {
      zl = bind1<M>(ld_fl(Z + i), ld_fl(Z + i + newlen), r);
      zh = bind1<M>(ld_fl(Z + i + h), ld_fl(Z + i + h + newlen), r);
      ql = bind1<M>(ld_fl(Q + i), ld_fl(Q + i + newlen), r);
      qh = bind1<M>(ld_fl(Q + i + h), ld_fl(Q + i + h + newlen), r);
      st_fl(Z + i, zl); st_fl(Z + i + h, zh); st_fl(Q + i, ql); st_fl(Q + i + h, qh);
    } else {
      zl = ld_fl(Z + i); zh = ld_fl(Z + i + h); ql = ld_fl(Q + i); qh = ld_fl(Q + i + h);
    }
    e[0] = fl_add(e[0], M::mul(zl, ql));  // sumcheck.rs:459-467
    fl_t z2 = fl_sub(fl_add(zh, zh), zl), q2 = fl_sub(fl_add(qh, qh), ql);
    e[1] = fl_add(e[1], M::mul(z2, q2));
  }
  if (BIND && newlen == 1 && blockIdx.x == 0 && threadIdx.x == 0) {
    st_fl(Z, bind1<M>(ld_fl(Z), ld_fl(Z + 1), r));
    st_fl(Q, bind1<M>(ld_fl(Q), ld_fl(Q + 1), r));
  }
  grid_sum_fl<3>(e, smem, partials, counter, result, mb);
}
// ---- phase two, first round on a z whose upper half is almost empty ---------------------------------------------------
// z = (w, 1, io, 0, ..., 0) (r1csproof.rs:150-160): of the upper half only the first k0 = 1 + |io| entries can be non-zero,
// and the same holds for the combined table (rA A + rB B + rC C)^T eq(rx), whose columns past the inputs have no entries.
// For a pair i >= k0 the round polynomial's value at 2 is (2*0 - z_i)(2*0 - q_i) = z_i q_i, its value at 0: one product and
// two reads instead of two and four.  Binding r_0 then gives z'_i = (1 - r_0) z_i for i >= k0: the prover keeps the factor
// outside the tables (prover.inc HalfZero) instead of running the bind, after sc_half_fix_kernel has put the k0 leading
// entries on the same scale.
template <class M>
__global__ void __launch_bounds__(256, 2) sc_quad_half0_kernel(const fl_t* __restrict__ Z, const fl_t* __restrict__ Q, size_t h, uint32_t k0,
                                                            fl_t* partials, unsigned int* counter, fl_t* result, Mailbox mb) {
  __shared__ fl_t smem[3 * 32];
  fl_t e[3];
  e[0] = e[1] = e[2] = fl_zero();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < h; i += (size_t)gridDim.x * blockDim.x) {
    const fl_t zl = ld_fl(Z + i), ql = ld_fl(Q + i);
    const fl_t p = M::mul(zl, ql);  // sumcheck.rs:459-467
    e[0] = fl_add(e[0], p);
    if (i < k0) {
      const fl_t zh = ld_fl(Z + i + h), qh = ld_fl(Q + i + h);
      e[1] = fl_add(e[1], M::mul(fl_sub(fl_add(zh, zh), zl), fl_sub(fl_add(qh, qh), ql)));
    } else {
      e[1] = fl_add(e[1], p);
    }
  }
  grid_sum_fl<3>(e, smem, partials, counter, result, mb);
}
// X[i] += t * X[i + h] for i < k0, both tables: with t = r/(1-r) the lower half times (1-r) is the table bound at r
__global__ void sc_half_fix_kernel(fl_t* __restrict__ Z, fl_t* __restrict__ Q, size_t h, uint32_t k0, fl_t t) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * k0; i += gridDim.x * blockDim.x) {
    fl_t* X = (i & 1) ? Q : Z;
    const uint32_t k = i >> 1;
    st_fl(X + k, fl_add(ld_fl(X + k), fl_mul_call(t, ld_fl(X + k + h))));
  }
}
// ---- phase one with eq(tau) factored out of the round polynomial ------------------------------------------------------
// Round j of prove_cubic_with_additive_term (sumcheck.rs:618-674) sums T*(A*B - C) over the table halves with
// T = eq(tau) bound at r_0..r_{j-1}.  T is a product: T[b*h + x] = S_j * eq(tau_j, b) * E_j[x], S_j = prod_k eq(tau_k, r_k)
// and E_j = eq(tau_{j+1}.., x), so the round polynomial is s_j(X) = S_j * eq(tau_j, X) * q_j(X) with the QUADRATIC
//     q_j(X) = sum_x E_j[x] * (A(X,x) * B(X,x) - C(X,x)).
// The kernel returns q_j(0) and q_j's leading coefficient sum_x E_j[x] * dA * dB; the host gets q_j(1) from the running
// claim and evaluates s_j at 0, 2, 3 (prover.inc EqFactor) -- the same field elements the reference computes, with 4
// instead of 6 products per pair, and T is neither read nor bound (8 -> 6 bind products, 24h -> 18h scalars moved).
// E_j[x] = HI[x >> lo_bits] * LO[x & mask]: a thread's pairs all share x & mask (lo_bits <= 8, 256 threads per block), so
// it multiplies by HI inside the loop and by LO once at the end.  The tables of every round come from one small launch:
// eq_pyramid_kernel writes, for t = 0..bits, eq over the LAST t variables of its range at offset 2^t - 1.
__global__ void eq_pyramid_kernel(const fl_t* __restrict__ tau, uint32_t m, uint32_t lo_bits, uint32_t hi_bits, fl_t init,
                                  fl_t* __restrict__ lo, fl_t* __restrict__ hi) {
  const uint32_t nlo = (2u << lo_bits) - 1, nhi = (2u << hi_bits) - 1;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nlo + nhi) return;
  const bool is_hi = p >= nlo;
  if (is_hi) p -= nlo;
  const uint32_t t = 31 - __clz(p + 1), x = p + 1 - (1u << t);
  const uint32_t end = is_hi ? m - lo_bits : m;  // the table covers tau[end - t .. end)
  fl_t acc = is_hi ? fl_one() : init;
  const fl_t one = fl_one();
  for (uint32_t k = 0; k < t; k++) {
    const fl_t rk = ldg_fl(tau + end - t + k);
    const bool bit = (x >> (t - 1 - k)) & 1u;
    acc = fl_mul(acc, bit ? rk : fl_sub(one, rk));
  }
  st_fl((is_hi ? hi : lo) + p, acc);
}
template <bool BIND, class M>
__global__ void __launch_bounds__(256, 2) sc_cubic_eq_kernel(const fl_t* __restrict__ lo_tab, const fl_t* __restrict__ hi_tab,
                                                          uint32_t lo_bits, fl_t* __restrict__ A, fl_t* __restrict__ B,
                                                          fl_t* __restrict__ C, size_t newlen, fl_t r, fl_t* partials,
                                                          unsigned int* counter, fl_t* result, Mailbox mb) {
  __shared__ fl_t smem[3 * 32];
  const size_t h = newlen >> 1;
  fl_t e[3];
  e[0] = e[1] = e[2] = fl_zero();
  fl_t* tabs[3] = {A, B, C};
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < h; i += (size_t)gridDim.x * blockDim.x) {
    fl_t lo[3], hi[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      fl_t* X = tabs[k];
      if (BIND) {
        lo[k] = bind1<M>(ld_fl(X + i), ld_fl(X + i + newlen), r);
        hi[k] = bind1<M>(ld_fl(X + i + h), ld_fl(X + i + h + newlen), r);
        st_fl(X + i, lo[k]);
        st_fl(X + i + h, hi[k]);
      } else {
        lo[k] = ld_fl(X + i);
        hi[k] = ld_fl(X + i + h);
      }
    }
    const fl_t hy = ldg_fl(hi_tab + (i >> lo_bits));
    e[0] = fl_add(e[0], M::mul(hy, fl_sub(M::mul(lo[0], lo[1]), lo[2])));
    e[1] = fl_add(e[1], M::mul(hy, M::mul(fl_sub(hi[0], lo[0]), fl_sub(hi[1], lo[1]))));
  }
  const fl_t lx = ldg_fl(lo_tab + (threadIdx.x & ((1u << lo_bits) - 1)));
  e[0] = M::mul(e[0], lx);
  e[1] = M::mul(e[1], lx);
  grid_sum_fl<3>(e, smem, partials, counter, result, mb);
}
// ---- latency variants for the short rounds --------------------------------------------------
// A short round is one pass of ~14 dependent field multiplications per thread in the kernels above: ~25 us however
// little data there is.  Here FOUR threads share a pair: each binds one table (cubic) or one table half (quadratic),
// the bound values meet in shared memory, and three (two) of the four threads evaluate one point each -- the
// dependent chain shrinks to 3-4 multiplications.  Lane l works on value index l & 3, so sums over equal indices
// are shuffles by 16, 8, 4.
__device__ __forceinline__ fl_t shfl_down_fl(const fl_t& x, int o) {
  fl_t y;
#pragma unroll
  for (int i = 0; i < 8; i++) y.v[i] = __shfl_down_sync(0xffffffffu, x.v[i], o);
  return y;
}
// e: this thread's term of value index (threadIdx.x & 3); sums over the grid end up in result[0..3) (index 3 unused)
__device__ __forceinline__ void quad_grid_sum(fl_t e, fl_t (*sred)[4], fl_t* partials, unsigned int* counter, fl_t* result, Mailbox mb) {
  __shared__ bool s_last_q;
  const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll 1
  for (int o = 16; o >= 4; o >>= 1) e = fl_add(e, shfl_down_fl(e, o));
  if (lane < 4) sred[wid][lane] = e;
  __syncthreads();
  if (threadIdx.x < 4) {
    fl_t acc = sred[0][threadIdx.x];
    for (uint32_t w = 1; w < nw; w++) acc = fl_add(acc, sred[w][threadIdx.x]);
    st_fl(&partials[(size_t)blockIdx.x * 4 + threadIdx.x], acc);
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(counter, 1u);
    s_last_q = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last_q) return;
  __threadfence();
  fl_t acc = fl_zero();
  for (uint32_t b = threadIdx.x >> 2; b < gridDim.x; b += blockDim.x >> 2) acc = fl_add(acc, ldcg_fl(&partials[(size_t)b * 4 + (threadIdx.x & 3)]));
#pragma unroll 1
  for (int o = 16; o >= 4; o >>= 1) acc = fl_add(acc, shfl_down_fl(acc, o));
  __syncthreads();
  if (lane < 4) sred[wid][lane] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    if (threadIdx.x < 3) {
      fl_t tot = sred[0][threadIdx.x];
      for (uint32_t w = 1; w < nw; w++) tot = fl_add(tot, sred[w][threadIdx.x]);
      st_fl(&result[threadIdx.x], tot);
      if (mb.host_flag != nullptr) {
        st_fl(&mb.host_result[threadIdx.x], tot);
        __threadfence_system();
      }
    }
    __syncwarp();
    if (threadIdx.x == 0) {
      *counter = 0;
      if (mb.host_flag != nullptr) {
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(mb.host_flag) = mb.seq;
      }
    }
  }
}
template <bool BIND>
__global__ void __launch_bounds__(256) sc_cubic_lat_kernel(fl_t* __restrict__ T, fl_t* __restrict__ A, fl_t* __restrict__ B,
                                                           fl_t* __restrict__ C, size_t newlen, fl_t r, fl_t* partials,
                                                           unsigned int* counter, fl_t* result, Mailbox mb) {
  __shared__ fl_t sv[64][3][4];  // [pair of the block][point 0,2,3][table]
  __shared__ fl_t sred[8][4];
  const size_t h = newlen >> 1;
  const uint32_t k = threadIdx.x & 3, pl = threadIdx.x >> 2;
  fl_t* X = k == 0 ? T : k == 1 ? A : k == 2 ? B : C;
  if (BIND && newlen == 1 && blockIdx.x == 0 && threadIdx.x < 4) st_fl(X, bind1<FlInl>(ld_fl(X), ld_fl(X + 1), r));
  fl_t e = fl_zero();
  // blockDim / 4 pairs per block and pass; the pass count is the same for every thread of a block (barriers inside)
  const size_t upb = blockDim.x >> 2;
  for (size_t base = (size_t)blockIdx.x * upb; base < h; base += (size_t)gridDim.x * upb) {
    const size_t pair = base + pl;
    fl_t lo = fl_zero(), hi = fl_zero();
    if (pair < h) {
      if (BIND) {
        fl_t a0 = ld_fl(X + pair), a1 = ld_fl(X + pair + newlen), b0 = ld_fl(X + pair + h), b1 = ld_fl(X + pair + h + newlen);
        lo = bind1<FlInl>(a0, a1, r);
        hi = bind1<FlInl>(b0, b1, r);
        st_fl(X + pair, lo);
        st_fl(X + pair + h, hi);
      } else {
        lo = ld_fl(X + pair);
        hi = ld_fl(X + pair + h);
      }
    }
    const fl_t d = fl_sub(hi, lo), p2 = fl_add(hi, d), p3 = fl_add(p2, d);  // values at 0, 2, 3 of this table's line
    if (base != (size_t)blockIdx.x * upb) __syncthreads();  // the previous pass has read sv
    sv[pl][0][k] = lo;
    sv[pl][1][k] = p2;
    sv[pl][2][k] = p3;
    __syncthreads();
    if (k < 3 && pair < h)  // comb(T,A,B,C) = T*(A*B - C) at point k (sumcheck.rs:622-650, r1csproof.rs:84-88)
      e = fl_add(e, fl_mul(sv[pl][k][0], fl_sub(fl_mul(sv[pl][k][1], sv[pl][k][2]), sv[pl][k][3])));
  }
  quad_grid_sum(e, sred, partials, counter, result, mb);
}
template <bool BIND>
__global__ void __launch_bounds__(256) sc_quad_lat_kernel(fl_t* __restrict__ Z, fl_t* __restrict__ Q, size_t newlen, fl_t r,
                                                          fl_t* partials, unsigned int* counter, fl_t* result, Mailbox mb) {
  __shared__ fl_t sv[64][4];  // [pair of the block][zl, ql, zh, qh]
  __shared__ fl_t sred[8][4];
  const size_t h = newlen >> 1;
  const uint32_t k = threadIdx.x & 3, pl = threadIdx.x >> 2;
  fl_t* X = (k & 1) ? Q : Z;
  if (BIND && newlen == 1 && blockIdx.x == 0 && threadIdx.x < 2) st_fl(X, bind1<FlInl>(ld_fl(X), ld_fl(X + 1), r));
  fl_t e = fl_zero();
  const size_t upb = blockDim.x >> 2;
  for (size_t base = (size_t)blockIdx.x * upb; base < h; base += (size_t)gridDim.x * upb) {
    const size_t pair = base + pl;
    const size_t at = pair + ((k & 2) ? h : 0);
    fl_t v = fl_zero();
    if (pair < h) {
      if (BIND) {
        v = bind1<FlInl>(ld_fl(X + at), ld_fl(X + at + newlen), r);
        st_fl(X + at, v);
      } else {
        v = ld_fl(X + at);
      }
    }
    if (base != (size_t)blockIdx.x * upb) __syncthreads();  // the previous pass has read sv
    sv[pl][k] = v;
    __syncthreads();
    if (pair < h) {  // sumcheck.rs:459-467
      if (k == 0) e = fl_add(e, fl_mul(sv[pl][0], sv[pl][1]));
      else if (k == 1) e = fl_add(e, fl_mul(fl_sub(fl_add(sv[pl][2], sv[pl][2]), sv[pl][0]), fl_sub(fl_add(sv[pl][3], sv[pl][3]), sv[pl][1])));
    }
  }
  quad_grid_sum(e, sred, partials, counter, result, mb);
}
static constexpr size_t kLatPairs = 8192;  // rounds with at most this many pairs run the latency variants
static size_t lat_pairs_setting() {  // SB200_SC_LAT_PAIRS: experiment knob (the four-threads-per-pair kernels loop over 64-pair passes)
  static const size_t v = [] {
    const char* e = getenv("SB200_SC_LAT_PAIRS");
    return e ? (size_t)strtoull(e, nullptr, 10) : kLatPairs;
  }();
  return v;
}
// Threads per block of the four-lanes-per-pair kernels.  128 = one warp per SM sub-partition: a dependent F_l product costs
// a warp about half as much when it does not share the sub-partition's integer pipe (tools/microbench/roundtrip_bench.cu).
// SB200_LAT_THREADS overrides (256 = the shape before).
static unsigned lat_threads() {
  static const unsigned v = [] {
    const char* e = getenv("SB200_LAT_THREADS");
    int t = e ? atoi(e) : 128;
    return (unsigned)(t == 64 || t == 128 || t == 256 ? t : 128);
  }();
  return v;
}
static inline unsigned lat_grid(size_t pairs, const ScWork& w) {
  const size_t upb = lat_threads() / 4;
  size_t g = (pairs + upb - 1) / upb, cap = ((size_t)w.max_blocks * 3) / 4;  // the partials buffer holds max_blocks * 3 scalars, four per block here
  if (const char* e = getenv("SB200_SC_LAT_BLOCKS")) cap = std::min<size_t>(cap, (size_t)atoi(e));
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}
static constexpr unsigned kCompactBlocks = 148;  // rounds with at most one block per SM run the compact kernels
static unsigned compact_blocks() {  // SB200_SC_COMPACT_BLOCKS: experiment knob (a large value = the compact kernels everywhere)
  static const unsigned v = [] {
    const char* e = getenv("SB200_SC_COMPACT_BLOCKS");
    return e ? (unsigned)strtoul(e, nullptr, 10) : kCompactBlocks;
  }();
  return v;
}
static inline unsigned sc_grid(size_t pairs, const ScWork& w) {
  size_t g = (pairs + 255) / 256;
  size_t cap = (size_t)kSMs * 4;
  if (cap > w.max_blocks) cap = w.max_blocks;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}
cudaError_t k_sc_cubic(fl_t* T, fl_t* A, fl_t* B, fl_t* C, size_t n_old, bool do_bind, fl_t r, const ScWork& w,
                       cudaStream_t s) {
  size_t newlen = do_bind ? n_old / 2 : n_old;
  unsigned g = sc_grid(newlen / 2, w);
  Mailbox mb{w.host_result, w.host_flag, w.seq};
  if (newlen / 2 <= lat_pairs_setting() && w.max_blocks >= kLatPairs / 64) {
    unsigned gl = lat_grid(newlen / 2, w);
    if (do_bind) { KProf kp_("sc_cubic_lat_kernel", s); sc_cubic_lat_kernel<true><<<gl, lat_threads(), 0, s>>>(T, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
    else { KProf kp_("sc_cubic_lat_kernel", s); sc_cubic_lat_kernel<false><<<gl, lat_threads(), 0, s>>>(T, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
    return SB_LAUNCHED();
  }
  const bool compact = g <= compact_blocks();
  if (do_bind && compact) { KProf kp_("sc_cubic_kernel", s); sc_cubic_kernel<true, FlCall><<<g, 256, 0, s>>>(T, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  else if (do_bind) { KProf kp_("sc_cubic_kernel", s); sc_cubic_kernel<true, FlInl><<<g, 256, 0, s>>>(T, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  else if (compact) { KProf kp_("sc_cubic_kernel", s); sc_cubic_kernel<false, FlCall><<<g, 256, 0, s>>>(T, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  else { KProf kp_("sc_cubic_kernel", s); sc_cubic_kernel<false, FlInl><<<g, 256, 0, s>>>(T, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  return SB_LAUNCHED();
}
// round 0 of phase two over tables of 2h entries whose upper halves are zero from index k0 on (sc_quad_half0_kernel)
cudaError_t k_sc_quad_half0(const fl_t* Z, const fl_t* Q, size_t h, uint32_t k0, const ScWork& w, cudaStream_t s) {
  unsigned g = sc_grid(h, w);
  Mailbox mb{w.host_result, w.host_flag, w.seq};
  if (g <= compact_blocks()) { KProf kp_("sc_quad_half0_kernel", s); sc_quad_half0_kernel<FlCall><<<g, 256, 0, s>>>(Z, Q, h, k0, w.partials, w.counter, w.result, mb); }
  else { KProf kp_("sc_quad_half0_kernel", s); sc_quad_half0_kernel<FlInl><<<g, 256, 0, s>>>(Z, Q, h, k0, w.partials, w.counter, w.result, mb); }
  return SB_LAUNCHED();
}
cudaError_t k_sc_half_fix(fl_t* Z, fl_t* Q, size_t h, uint32_t k0, fl_t t, cudaStream_t s) {
  if (!k0) return cudaSuccess;
  { KProf kp_("sc_half_fix_kernel", s); sc_half_fix_kernel<<<(2 * k0 + 127) / 128, 128, 0, s>>>(Z, Q, h, k0, t); }
  return SB_LAUNCHED();
}
// rounds with this many pairs run sc_cubic_kernel / sc_quad_kernel (the others the four-lane latency variants)
bool sc_round_is_long(size_t pairs, const ScWork& w) { return !(pairs <= lat_pairs_setting() && w.max_blocks >= kLatPairs / 64); }
cudaError_t k_eq_pyramid(const fl_t* tau, uint32_t m, uint32_t lo_bits, fl_t init, fl_t* lo, fl_t* hi, cudaStream_t s) {
  const uint32_t hi_bits = m - 1 - lo_bits, n = (2u << lo_bits) + (2u << hi_bits) - 2;
  { KProf kp_("eq_pyramid_kernel", s); eq_pyramid_kernel<<<(n + 127) / 128, 128, 0, s>>>(tau, m, lo_bits, hi_bits, init, lo, hi); }
  return SB_LAUNCHED();
}
// One factored round: the tables hold n_old entries (m_now = log2 of the length after this launch's bind); lo / hi are the
// pyramids of k_eq_pyramid built with the same lo_bits over the m variables of the unbound tables.
cudaError_t k_sc_cubic_eq(const fl_t* lo, const fl_t* hi, uint32_t lo_bits, fl_t* A, fl_t* B, fl_t* C, size_t n_old, bool do_bind,
                          fl_t r, const ScWork& w, cudaStream_t s) {
  const size_t newlen = do_bind ? n_old / 2 : n_old, h = newlen / 2;
  uint32_t nv = 0;  // variables of E: h = 2^nv
  while (((size_t)1 << nv) < h) nv++;
  const uint32_t lb = nv < lo_bits ? nv : lo_bits, u = nv - lb;
  const fl_t* lo_tab = lo + (((size_t)1 << lb) - 1);
  const fl_t* hi_tab = hi + (((size_t)1 << u) - 1);
  unsigned g = sc_grid(h, w);
  Mailbox mb{w.host_result, w.host_flag, w.seq};
  const bool compact = g <= compact_blocks();
  if (do_bind && compact) { KProf kp_("sc_cubic_eq_kernel", s); sc_cubic_eq_kernel<true, FlCall><<<g, 256, 0, s>>>(lo_tab, hi_tab, lb, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  else if (do_bind) { KProf kp_("sc_cubic_eq_kernel", s); sc_cubic_eq_kernel<true, FlInl><<<g, 256, 0, s>>>(lo_tab, hi_tab, lb, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  else if (compact) { KProf kp_("sc_cubic_eq_kernel", s); sc_cubic_eq_kernel<false, FlCall><<<g, 256, 0, s>>>(lo_tab, hi_tab, lb, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  else { KProf kp_("sc_cubic_eq_kernel", s); sc_cubic_eq_kernel<false, FlInl><<<g, 256, 0, s>>>(lo_tab, hi_tab, lb, A, B, C, newlen, r, w.partials, w.counter, w.result, mb); }
  return SB_LAUNCHED();
}
cudaError_t k_sc_quad(fl_t* Z, fl_t* Q, size_t n_old, bool do_bind, fl_t r, const ScWork& w, cudaStream_t s) {
  size_t newlen = do_bind ? n_old / 2 : n_old;
  unsigned g = sc_grid(newlen / 2, w);
  Mailbox mb{w.host_result, w.host_flag, w.seq};
  if (newlen / 2 <= lat_pairs_setting() && w.max_blocks >= kLatPairs / 64) {
    unsigned gl = lat_grid(newlen / 2, w);
    if (do_bind) { KProf kp_("sc_quad_lat_kernel", s); sc_quad_lat_kernel<true><<<gl, lat_threads(), 0, s>>>(Z, Q, newlen, r, w.partials, w.counter, w.result, mb); }
    else { KProf kp_("sc_quad_lat_kernel", s); sc_quad_lat_kernel<false><<<gl, lat_threads(), 0, s>>>(Z, Q, newlen, r, w.partials, w.counter, w.result, mb); }
    return SB_LAUNCHED();
  }
  const bool compact = g <= compact_blocks();
  if (do_bind && compact) { KProf kp_("sc_quad_kernel", s); sc_quad_kernel<true, FlCall><<<g, 256, 0, s>>>(Z, Q, newlen, r, w.partials, w.counter, w.result, mb); }
  else if (do_bind) { KProf kp_("sc_quad_kernel", s); sc_quad_kernel<true, FlInl><<<g, 256, 0, s>>>(Z, Q, newlen, r, w.partials, w.counter, w.result, mb); }
  else if (compact) { KProf kp_("sc_quad_kernel", s); sc_quad_kernel<false, FlCall><<<g, 256, 0, s>>>(Z, Q, newlen, r, w.partials, w.counter, w.result, mb); }
  else { KProf kp_("sc_quad_kernel", s); sc_quad_kernel<false, FlInl><<<g, 256, 0, s>>>(Z, Q, newlen, r, w.partials, w.counter, w.result, mb); }
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ bound(L), dot
// grid (ceil(R/128), JC): tmp[jc*R + i] = sum_{j in chunk jc} Lv[j] * Z[j*R + i]
__global__ void bound_partial_kernel(const fl_t* __restrict__ Z, const fl_t* __restrict__ Lv, uint32_t Lsize,
                                     uint32_t Rsize, uint32_t jper, fl_t* __restrict__ tmp) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Rsize) return;
  uint32_t j0 = blockIdx.y * jper, j1 = min(Lsize, j0 + jper);
  fl_t acc = fl_zero();
  for (uint32_t j = j0; j < j1; j++) acc = fl_add(acc, fl_mul(ldg_fl(Lv + j), ldg_fl(Z + (size_t)j * Rsize + i)));
  st_fl(tmp + (size_t)blockIdx.y * Rsize + i, acc);
}
__global__ void bound_sum_kernel(const fl_t* __restrict__ tmp, uint32_t Rsize, uint32_t jc, fl_t* __restrict__ out,
                                 fl_t* __restrict__ out_mont = nullptr) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Rsize) return;
  // four independent partial sums: the loads of a step do not wait for the previous step's addition
  fl_t acc[4];
#pragma unroll
  for (int k = 0; k < 4; k++) acc[k] = fl_zero();
  uint32_t c = 0;
  for (; c + 4 <= jc; c += 4) {
    fl_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = ldg_fl(tmp + (size_t)(c + k) * Rsize + i);
#pragma unroll
    for (int k = 0; k < 4; k++) acc[k] = fl_add(acc[k], v[k]);
  }
  for (; c < jc; c++) acc[0] = fl_add(acc[0], ldg_fl(tmp + (size_t)c * Rsize + i));
  const fl_t tot = fl_add(fl_add(acc[0], acc[1]), fl_add(acc[2], acc[3]));
  st_fl(out + i, tot);
  if (out_mont != nullptr) st_fl(out_mont + i, fl_to_mont(tot));
}
// chunks of rows: enough blocks for ~8 per SM, but few enough (<= 64) that the second pass stays short
static inline uint32_t bound_chunks(uint32_t Lsize, uint32_t Rsize) {
  uint32_t xblocks = (Rsize + 127) / 128;
  uint32_t want = (uint32_t)(8 * kSMs + xblocks - 1) / xblocks;
  if (want > 64) want = 64;
  if (want < 1) want = 1;
  if (want > Lsize) want = Lsize;
  return want;
}
size_t bound_L_tmp_scalars(uint32_t Lsize, uint32_t Rsize) { return (size_t)bound_chunks(Lsize, Rsize) * Rsize; }
cudaError_t k_bound_L(const fl_t* Z, const fl_t* Lv, uint32_t Lsize, uint32_t Rsize, fl_t* out, fl_t* tmp, cudaStream_t s) {
  uint32_t jc = bound_chunks(Lsize, Rsize), jper = (Lsize + jc - 1) / jc;
  jc = (Lsize + jper - 1) / jper;
  dim3 grid((Rsize + 127) / 128, jc);
  { KProf kp_("bound_partial_kernel", s); bound_partial_kernel<<<grid, 128, 0, s>>>(Z, Lv, Lsize, Rsize, jper, tmp); }
  cudaError_t e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  { KProf kp_("bound_sum_kernel", s); bound_sum_kernel<<<(Rsize + 127) / 128, 128, 0, s>>>(tmp, Rsize, jc, out); }
  return SB_LAUNCHED();
}
cudaError_t k_bound_L_canon(const fl_t* Z_canon, const fl_t* Lv, uint32_t Lsize, uint32_t Rsize, fl_t* out_canon, fl_t* out_mont,
                            fl_t* tmp, cudaStream_t s) {
  uint32_t jc = bound_chunks(Lsize, Rsize), jper = (Lsize + jc - 1) / jc;
  jc = (Lsize + jper - 1) / jper;
  dim3 grid((Rsize + 127) / 128, jc);
  { KProf kp_("bound_partial_kernel", s); bound_partial_kernel<<<grid, 128, 0, s>>>(Z_canon, Lv, Lsize, Rsize, jper, tmp); }
  cudaError_t e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  { KProf kp_("bound_sum_kernel", s); bound_sum_kernel<<<(Rsize + 127) / 128, 128, 0, s>>>(tmp, Rsize, jc, out_canon, out_mont); }
  return SB_LAUNCHED();
}
__global__ void dot_small_kernel(const fl_t* __restrict__ a, const fl_t* __restrict__ b, uint32_t n, fl_t* out) {
  __shared__ fl_t smem[32];
  fl_t acc[1];
  acc[0] = fl_zero();
  for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) acc[0] = fl_add(acc[0], fl_mul(ld_fl(a + i), ld_fl(b + i)));
  block_sum_fl<1>(acc, smem);
  if (threadIdx.x == 0) st_fl(out, acc[0]);
}
cudaError_t k_dot_small(const fl_t* a, const fl_t* b, uint32_t n, fl_t* out, cudaStream_t s) {
  { KProf kp_("dot_small_kernel", s); dot_small_kernel<<<1, 1024, 0, s>>>(a, b, n, out); }
  return SB_LAUNCHED();
}

// result[m] = <e, y_m>, m = 0..2, over n entries (multi-block)        (sparse_mlpoly.rs:412-423 evaluate_with_tables)
__global__ void __launch_bounds__(256) dot3_kernel(const fl_t* __restrict__ e, const fl_t* __restrict__ yA, const fl_t* __restrict__ yB,
                                                   const fl_t* __restrict__ yC, size_t n, fl_t* partials, unsigned int* counter,
                                                   fl_t* result) {
  __shared__ fl_t smem[3 * 32];
  fl_t acc[3];
  acc[0] = acc[1] = acc[2] = fl_zero();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    fl_t x = ldg_fl(e + i);
    acc[0] = fl_add(acc[0], fl_mul(x, ldg_fl(yA + i)));
    acc[1] = fl_add(acc[1], fl_mul(x, ldg_fl(yB + i)));
    acc[2] = fl_add(acc[2], fl_mul(x, ldg_fl(yC + i)));
  }
  grid_sum_fl<3>(acc, smem, partials, counter, result);
}
cudaError_t k_dot3(const fl_t* e, const fl_t* yA, const fl_t* yB, const fl_t* yC, size_t n, const ScWork& w, cudaStream_t s) {
  { KProf kp_("dot3_kernel", s); dot3_kernel<<<sc_grid(n, w), 256, 0, s>>>(e, yA, yB, yC, n, w.partials, w.counter, w.result); }
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ bullet reduction, scalar side
// One block.  State: a, b (length n_cur, Montgomery), coeff (length R/n_cur) with
//   G_k[i] = sum_t coeff[t] * G[t*n_cur + i]           (the folded generators, never materialised)
// Step:  (1) if do_fold: a_L*u + u^-1*a_R, b_L*u^-1 + u*b_R, coeff' = coeff (x) (u^-1, u)   bullet.rs:103-108
//        (2) if n > 1: c_L = <a_L,b_R>, c_R = <a_R,b_L> and the two MSM scalar rows over the ORIGINAL
//            bases:  L = sum a_L[i]*G_R[i] + c_L*Q + blind_L*H,  R = sum a_R[i]*G_L[i] + c_R*Q + blind_R*H
//            (bullet.rs:79-96).  Rows have stride R+2: [0,R) bases G, [R] = H, [R+1] = Q; canonical form.
__device__ __forceinline__ fl_t fl_from_mont_c(const fl_t& m) {
  fl_t one = fl_zero();
  one.v[0] = 1;
  return fl_mul_call(m, one);
}
// Grid: ceil(R / 512) "row" blocks + one "fold" block, so that a round costs five dependent multiplications instead
// of one SM working through every loop.  A row block derives what it needs from the UNFOLDED vectors (the fold of one
// entry of a and of one coefficient are three multiplications), the fold block writes the folded a, b, coeff for the
// next round into the other buffer of a ping-pong pair -- the row blocks never read what it writes -- and adds the
// two inner products and the blind rows.
static constexpr int kBulletThreads = 512;
__global__ void __launch_bounds__(kBulletThreads) bullet_step_kernel(const fl_t* __restrict__ a_in, const fl_t* __restrict__ b_in,
                                                                     fl_t* __restrict__ a_out, fl_t* __restrict__ b_out,
                                                                     const fl_t* __restrict__ coeff_in, fl_t* __restrict__ coeff_out,
                                                                     uint32_t n_cur, uint32_t Rsize, bool do_fold, fl_t u, fl_t u_inv,
                                                                     fl_t blind_L, fl_t blind_R, fl_t* __restrict__ sL, fl_t* __restrict__ sR,
                                                                     uint32_t pos_h, uint32_t pos_q) {
  __shared__ fl_t smem[2 * 32];
  const uint32_t n = do_fold ? n_cur / 2 : n_cur, half = n / 2;
  if (blockIdx.x + 1 < gridDim.x) {  // row block: the two MSM scalar rows over the ORIGINAL bases
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (n <= 1 || g >= Rsize) return;
    const uint32_t t = g / n, pos = g % n, j = pos >= half ? pos - half : half + pos;
    fl_t co, aj;
    if (do_fold) {
      co = FlCall::mul(ld_fl(coeff_in + (t >> 1)), (t & 1) ? u : u_inv);
      aj = fl_add(FlCall::mul(ld_fl(a_in + j), u), FlCall::mul(u_inv, ld_fl(a_in + n + j)));
    } else {
      co = ld_fl(coeff_in + t);
      aj = ld_fl(a_in + j);
    }
    const fl_t sc = fl_from_mont_c(FlCall::mul(aj, co));
    st_fl(sL + g, pos >= half ? sc : fl_zero());
    st_fl(sR + g, pos >= half ? fl_zero() : sc);
    return;
  }
  // fold block
  const fl_t *a = a_in, *b = b_in;
  if (do_fold) {
    const uint32_t nco_in = Rsize / n_cur;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
      fl_t al = ld_fl(a_in + i), ar = ld_fl(a_in + n + i), bl = ld_fl(b_in + i), br = ld_fl(b_in + n + i);
      st_fl(a_out + i, fl_add(FlCall::mul(al, u), FlCall::mul(u_inv, ar)));
      st_fl(b_out + i, fl_add(FlCall::mul(bl, u_inv), FlCall::mul(u, br)));
    }
    for (uint32_t t = threadIdx.x; t < nco_in; t += blockDim.x) {
      fl_t c = ld_fl(coeff_in + t);
      st_fl(coeff_out + 2 * t, FlCall::mul(c, u_inv));
      st_fl(coeff_out + 2 * t + 1, FlCall::mul(c, u));
    }
    __syncthreads();
    a = a_out;
    b = b_out;
  }
  if (n <= 1) return;
  fl_t c[2];
  c[0] = c[1] = fl_zero();
  for (uint32_t i = threadIdx.x; i < half; i += blockDim.x) {
    fl_t al = ld_fl(a + i), ar = ld_fl(a + half + i), bl = ld_fl(b + i), br = ld_fl(b + half + i);
    c[0] = fl_add(c[0], FlCall::mul(al, br));
    c[1] = fl_add(c[1], FlCall::mul(ar, bl));
  }
  block_sum_fl<2>(c, smem);
  if (threadIdx.x < 4) {  // rows have stride R + 2: [R] = H (blind), [R + 1] = Q (inner product); canonical form
    const fl_t v = threadIdx.x == 0 ? blind_L : threadIdx.x == 1 ? blind_R : fl_zero();
    if (threadIdx.x < 2) st_fl((threadIdx.x == 0 ? sL : sR) + pos_h, fl_from_mont_c(v));
    if (threadIdx.x == 0) {
      st_fl(sL + pos_q, fl_from_mont_c(c[0]));
      st_fl(sR + pos_q, fl_from_mont_c(c[1]));
    }
  }
}
cudaError_t k_bullet_step(const fl_t* a_in, const fl_t* b_in, fl_t* a_out, fl_t* b_out, const fl_t* coeff_in, fl_t* coeff_out,
                          uint32_t n_cur, uint32_t Rsize, bool do_fold, fl_t u, fl_t u_inv, fl_t blind_L, fl_t blind_R, fl_t* sL,
                          fl_t* sR, cudaStream_t s, uint32_t pos_h, uint32_t pos_q) {
  const unsigned row_blocks = (Rsize + kBulletThreads - 1) / kBulletThreads;
  if (pos_h == 0 && pos_q == 0) { pos_h = Rsize; pos_q = Rsize + 1; }
  { KProf kp_("bullet_step_kernel", s); bullet_step_kernel<<<row_blocks + 1, kBulletThreads, 0, s>>>(a_in, b_in, a_out, b_out, coeff_in, coeff_out, n_cur, Rsize, do_fold, u, u_inv,
                                                              blind_L, blind_R, sL, sR, pos_h, pos_q); }
  return SB_LAUNCHED();
}
__global__ void scale_from_mont_kernel(const fl_t* __restrict__ coeff, fl_t d, uint32_t n, fl_t* __restrict__ out) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    st_fl(out + i, fl_from_mont(fl_mul(ld_fl(coeff + i), d)));
}
cudaError_t k_scale_from_mont(const fl_t* coeff, fl_t d, uint32_t n, fl_t* out, cudaStream_t s) {
  { KProf kp_("scale_from_mont_kernel", s); scale_from_mont_kernel<<<grid_for(n, 128), 128, 0, s>>>(coeff, d, n, out); }
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ SNARK::encode: dense representation
// SparseMatPolynomial::multi_sparse_to_dense_rep (sparse_mlpoly.rs:371-422) without the sequential timestamp loop of
// AddrTimestamps::new (:233-266).  That loop walks the three matrices in order and, per memory cell, counts the
// accesses so far: the read timestamp of op i of matrix m at address a is (accesses to a by matrices before m) +
// (rank of i among the ops of m at a, in op order), the audit timestamp the total.  The CSR / CSC of the instance are
// stable by construction (spartan_b200.cu upload_sparse), so the rank is the position inside the segment, and the op a
// position came from is kept beside it (Instance::op_csr / op_csc).  Ops past the last entry of a matrix read cell 0.
// All outputs are canonical integers (the MSM's input form), not Montgomery.
struct DenseRepArgs {
  const uint32_t* ptr[3];  // segment starts (rows or columns), nseg + 1 entries
  const uint32_t* idx[3];  // unused here (the other coordinate), kept for symmetry with Csr
  const uint32_t* op[3];   // position -> op index within the matrix
  const fl_t* val[3];      // Montgomery values by position (NULL: do not write the val block)
  uint32_t nnz[3];
  uint32_t nseg, N;
};
__device__ __forceinline__ fl_t fl_small(uint64_t x) {
  fl_t r = fl_zero();
  r.v[0] = (uint32_t)x;
  r.v[1] = (uint32_t)(x >> 32);
  return r;
}
__device__ __forceinline__ uint32_t seg_count(const DenseRepArgs& a, int m, uint32_t addr) {
  uint32_t c = addr < a.nseg ? __ldg(a.ptr[m] + addr + 1) - __ldg(a.ptr[m] + addr) : 0u;
  return addr == 0 ? c + (a.N - a.nnz[m]) : c;
}
// audit_ts[addr] for addr < cells                                            sparse_mlpoly.rs:249-252, 262
__global__ void snark_audit_kernel(DenseRepArgs a, uint32_t cells, fl_t* __restrict__ audit) {
  for (uint32_t addr = blockIdx.x * blockDim.x + threadIdx.x; addr < cells; addr += gridDim.x * blockDim.x)
    st_fl(audit + addr, fl_small((uint64_t)seg_count(a, 0, addr) + seg_count(a, 1, addr) + seg_count(a, 2, addr)));
}
// grid.y = matrix.  addr_out / ts_out: 3 blocks of N scalars each (ops_addr and read_ts of this orientation);
// val_out: 3 blocks of N (written when a.val[m] != NULL).
__global__ void snark_ops_kernel(DenseRepArgs a, fl_t* __restrict__ addr_out, fl_t* __restrict__ ts_out, fl_t* __restrict__ val_out) {
  const int m = blockIdx.y;
  const uint32_t nnz = a.nnz[m], N = a.N;
  const uint32_t* __restrict__ ptr = a.ptr[m];
  uint32_t before0 = 0;  // accesses to cell 0 by the matrices before m
  for (int k = 0; k < m; k++) before0 += seg_count(a, k, 0);
  const uint32_t own0 = a.nseg ? __ldg(ptr + 1) - __ldg(ptr) : 0u;
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < N; p += gridDim.x * blockDim.x) {
    uint32_t i, addr, ts;
    if (p < nnz) {
      uint32_t lo = 0, hi = a.nseg;  // the segment holding position p: largest s with ptr[s] <= p
      while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(ptr + mid) <= p) lo = mid; else hi = mid;
      }
      addr = lo;
      i = __ldg(a.op[m] + p);
      ts = p - __ldg(ptr + addr);
      for (int k = 0; k < m; k++) ts += seg_count(a, k, addr);
      if (a.val[m] != nullptr) st_fl(val_out + (size_t)m * N + i, fl_from_mont(ldg_fl(a.val[m] + p)));
    } else {  // padding op: reads cell 0 after every real op of this matrix        sparse_mlpoly.rs:353-369
      i = p;
      addr = 0;
      ts = before0 + own0 + (p - nnz);
      if (a.val[m] != nullptr) st_fl(val_out + (size_t)m * N + i, fl_zero());
    }
    st_fl(addr_out + (size_t)m * N + i, fl_small(addr));
    st_fl(ts_out + (size_t)m * N + i, fl_small(ts));
  }
}
__global__ void fl_fill_zero_kernel(fl_t* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) st_fl(out + i, fl_zero());
}
cudaError_t k_snark_dense(const Csr& csr, const uint32_t* const op_csr[3], uint32_t nrows, const Csr& csc,
                          const uint32_t* const op_csc[3], uint32_t ncols, const uint32_t nnz[3], uint32_t N, uint32_t cells,
                          fl_t* comb_ops, fl_t* comb_mem, cudaStream_t s) {
  DenseRepArgs r, c;
  for (int m = 0; m < 3; m++) {
    r.ptr[m] = csr.ptr[m]; r.idx[m] = csr.idx[m]; r.op[m] = op_csr[m]; r.val[m] = csr.val[m]; r.nnz[m] = nnz[m];
    c.ptr[m] = csc.ptr[m]; c.idx[m] = csc.idx[m]; c.op[m] = op_csc[m]; c.val[m] = nullptr; c.nnz[m] = nnz[m];
  }
  r.nseg = nrows; c.nseg = ncols;
  r.N = c.N = N;
  // comb_ops = row addr | row read_ts | col addr | col read_ts | val | zero padding   (5 x 3 blocks of N, then N zeros)
  dim3 grid(grid_for(N, 256), 3);
  { KProf kp_("snark_ops_kernel", s); snark_ops_kernel<<<grid, 256, 0, s>>>(r, comb_ops, comb_ops + (size_t)3 * N, comb_ops + (size_t)12 * N); }
  cudaError_t e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  { KProf kp_("snark_ops_kernel", s); snark_ops_kernel<<<grid, 256, 0, s>>>(c, comb_ops + (size_t)6 * N, comb_ops + (size_t)9 * N, nullptr); }
  e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  { KProf kp_("fl_fill_zero_kernel", s); fl_fill_zero_kernel<<<grid_for(N, 256), 256, 0, s>>>(comb_ops + (size_t)15 * N, N); }
  e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  { KProf kp_("snark_audit_kernel", s); snark_audit_kernel<<<grid_for(cells, 256), 256, 0, s>>>(r, cells, comb_mem); }
  e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  { KProf kp_("snark_audit_kernel", s); snark_audit_kernel<<<grid_for(cells, 256), 256, 0, s>>>(c, cells, comb_mem + cells); }
  return SB_LAUNCHED();
}

// Derefs.comb (sparse_mlpoly.rs:47-62, 268-284, 508-514): what every op reads from the eq tables of the evaluation point,
// row side then column side, three matrices of N ops each, then 2 N zeros.  The addresses are the low words of the
// address blocks of comb_ops; mem_rx / mem_ry are Montgomery, the output canonical (the MSM's input form).
__global__ void snark_deref_kernel(const fl_t* __restrict__ comb_ops, const fl_t* __restrict__ mem_rx, const fl_t* __restrict__ mem_ry,
                                   uint32_t N, fl_t* __restrict__ out) {
  const size_t total = (size_t)8 * N;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    const uint32_t blk = (uint32_t)(t / N);
    fl_t v = fl_zero();
    if (blk < 6) {
      const size_t i = t - (size_t)blk * N;
      const fl_t* addr = comb_ops + (size_t)(blk < 3 ? blk : 3 + blk) * N + i;  // row addr blocks 0..2, col addr blocks 6..8
      const uint32_t a = __ldg(&addr->v[0]);
      v = fl_from_mont(ldg_fl((blk < 3 ? mem_rx : mem_ry) + a));
    }
    st_fl(out + t, v);
  }
}
cudaError_t k_snark_deref(const fl_t* comb_ops, const fl_t* mem_rx, const fl_t* mem_ry, uint32_t N, fl_t* out, cudaStream_t s) {
  { KProf kp_("snark_deref_kernel", s); snark_deref_kernel<<<grid_for((size_t)8 * N, 256), 256, 0, s>>>(comb_ops, mem_rx, mem_ry, N, out); }
  return SB_LAUNCHED();
}

// PolyEvalNetwork::new (sparse_mlpoly.rs:652-668): the hash layer (build_hash_layer, :531-606) and the product circuits
// (product_tree.rs:16-64) of one side.  hash(addr, val, ts) = ts * r_hash^2 + val * r_hash + addr - r_multiset_check.
// A product circuit's layer k+1 is the element-wise product of the two halves of layer k (left_vec / right_vec ARE those
// halves), so a tree is one buffer of 2 len scalars: layer k (len >> k entries) at offset 2 len - (2 len >> k).
// All values Montgomery; addresses and timestamps arrive as the canonical integers of comb_ops / comb_mem.
__global__ void snark_hash_mem_kernel(const fl_t* __restrict__ mem, const fl_t* __restrict__ audit_ts, uint32_t cells, fl_t r_hash,
                                      fl_t r_hash_sqr, fl_t r_ms, fl_t* __restrict__ init_out, fl_t* __restrict__ audit_out) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cells; i += gridDim.x * blockDim.x) {
    const fl_t a = fl_to_mont(fl_small(i));
    const fl_t base = fl_sub(fl_add(fl_mul(ldg_fl(mem + i), r_hash), a), r_ms);  // ts = 0 at init time
    st_fl(init_out + i, base);
    st_fl(audit_out + i, fl_add(fl_mul(fl_to_mont(ldg_fl(audit_ts + i)), r_hash_sqr), base));
  }
}
// grid.y = matrix; addr / ts: 3 blocks of N canonical integers; read_out / write_out: the 3 tree buffers' layer 0
struct HashOpsOut {
  fl_t* read[3];
  fl_t* write[3];
};
__global__ void snark_hash_ops_kernel(const fl_t* __restrict__ mem, const fl_t* __restrict__ addr, const fl_t* __restrict__ ts,
                                      uint32_t N, fl_t r_hash, fl_t r_hash_sqr, fl_t r_ms, HashOpsOut out) {
  const int m = blockIdx.y;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const fl_t a = ldg_fl(addr + (size_t)m * N + i);
    const fl_t v = ldg_fl(mem + a.v[0]);  // the value the op reads (Derefs)
    const fl_t base = fl_sub(fl_add(fl_mul(v, r_hash), fl_to_mont(a)), r_ms);
    const fl_t rd = fl_add(fl_mul(fl_to_mont(ldg_fl(ts + (size_t)m * N + i)), r_hash_sqr), base);
    st_fl(out.read[m] + i, rd);
    st_fl(out.write[m] + i, fl_add(rd, r_hash_sqr));  // write_ts = read_ts + 1
  }
}
static constexpr int kMaxTrees = 16;
struct ProdTrees {
  fl_t* buf[kMaxTrees];
  uint32_t len[kMaxTrees];
};
__device__ __forceinline__ size_t tree_off(uint32_t len, uint32_t k) { return 2 * (size_t)len - ((2 * (size_t)len) >> k); }
// layer k -> k + 1 of every tree that still has more than `tail` entries in layer k (grid.y = tree)
__global__ void prod_layer_kernel(ProdTrees t, uint32_t k, uint32_t tail) {
  const uint32_t len = t.len[blockIdx.y], cur = len >> k;
  if (cur <= tail) return;
  const uint32_t h = cur >> 1;
  const fl_t* __restrict__ in = t.buf[blockIdx.y] + tree_off(len, k);
  fl_t* __restrict__ out = t.buf[blockIdx.y] + tree_off(len, k + 1);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < h; i += gridDim.x * blockDim.x)
    st_fl(out + i, fl_mul(ld_fl(in + i), ld_fl(in + h + i)));
}
// the remaining layers (at most `tail` = 2 blockDim.x entries) of one tree per block; products[tree] canonical
__global__ void __launch_bounds__(512) prod_tail_kernel(ProdTrees t, uint32_t tail, fl_t* __restrict__ products) {
  const uint32_t len = t.len[blockIdx.x];
  uint32_t k = 0;
  while ((len >> k) > tail) k++;
  fl_t* buf = t.buf[blockIdx.x];
  for (uint32_t cur = len >> k; cur > 1; cur >>= 1, k++) {
    const uint32_t h = cur >> 1;
    const fl_t* in = buf + tree_off(len, k);
    fl_t* out = buf + tree_off(len, k + 1);
    for (uint32_t i = threadIdx.x; i < h; i += blockDim.x) st_fl(out + i, fl_mul(ld_fl(in + i), ld_fl(in + h + i)));
    __syncthreads();
  }
  if (threadIdx.x == 0) st_fl(products + blockIdx.x, fl_from_mont(ld_fl(buf + tree_off(len, k))));
}
// one side: trees[0] = init, [1..3] = read, [4..6] = write, [7] = audit (buffers of 2 len scalars, layer 0 written here)
cudaError_t k_snark_hash_layer(const fl_t* mem, const fl_t* addr, const fl_t* ts, const fl_t* audit_ts, uint32_t N, uint32_t cells,
                               fl_t r_hash, fl_t r_hash_sqr, fl_t r_ms, fl_t* const trees[8], cudaStream_t s) {
  { KProf kp_("snark_hash_mem_kernel", s); snark_hash_mem_kernel<<<grid_for(cells, 256), 256, 0, s>>>(mem, audit_ts, cells, r_hash, r_hash_sqr, r_ms, trees[0], trees[7]); }
  cudaError_t e = SB_LAUNCHED();
  if (e != cudaSuccess) return e;
  HashOpsOut o;
  for (int m = 0; m < 3; m++) { o.read[m] = trees[1 + m]; o.write[m] = trees[4 + m]; }
  dim3 grid(grid_for(N, 256), 3);
  { KProf kp_("snark_hash_ops_kernel", s); snark_hash_ops_kernel<<<grid, 256, 0, s>>>(mem, addr, ts, N, r_hash, r_hash_sqr, r_ms, o); }
  return SB_LAUNCHED();
}
// ntrees <= 16 product circuits at once; products[ntrees] canonical
cudaError_t k_prod_trees(fl_t* const bufs[], const uint32_t lens[], int ntrees, fl_t* products, cudaStream_t s) {
  if (ntrees < 1 || ntrees > kMaxTrees) return cudaErrorInvalidValue;
  ProdTrees t;
  uint32_t maxlen = 0;
  for (int i = 0; i < kMaxTrees; i++) {
    t.buf[i] = i < ntrees ? bufs[i] : nullptr;
    t.len[i] = i < ntrees ? lens[i] : 0;
    maxlen = t.len[i] > maxlen ? t.len[i] : maxlen;
  }
  const uint32_t tail = 1024;
  for (uint32_t k = 0; (maxlen >> k) > tail; k++) {
    dim3 grid(grid_for((maxlen >> k) / 2, 256, 4), ntrees);
    { KProf kp_("prod_layer_kernel", s); prod_layer_kernel<<<grid, 256, 0, s>>>(t, k, tail); }
    cudaError_t e = SB_LAUNCHED();
    if (e != cudaSuccess) return e;
  }
  { KProf kp_("prod_tail_kernel", s); prod_tail_kernel<<<ntrees, 512, 0, s>>>(t, tail, products); }
  return SB_LAUNCHED();
}

#include "snark_kernels.inc"

// ------------------------------------------------------------------ fixed-base MSM
__device__ __forceinline__ niels_t ldg_niels(const niels_t* p) {
  niels_t n;
  n.yplusx = ldg_fp(&p->yplusx);
  n.yminusx = ldg_fp(&p->yminusx);
  n.xy2d = ldg_fp(&p->xy2d);
  return n;
}
// digit `w` of the signed radix-2^c recoding of the scalar at kp (carry recomputed from the lower windows)
__device__ __forceinline__ int32_t msm_digit_at(const uint32_t* kp, uint32_t w, uint32_t c) {
  uint32_t carry = 0;
  int32_t d = 0;
  for (uint32_t i = 0; i <= w; i++) d = msm_digit(kp, i, c, carry);
  return d;
}
__device__ __forceinline__ ge_t shfl_down_ge(const ge_t& x, int o) {
  ge_t y;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    y.X.v[i] = __shfl_down_sync(0xffffffffu, x.X.v[i], o);
    y.Y.v[i] = __shfl_down_sync(0xffffffffu, x.Y.v[i], o);
    y.Z.v[i] = __shfl_down_sync(0xffffffffu, x.Z.v[i], o);
    y.T.v[i] = __shfl_down_sync(0xffffffffu, x.T.v[i], o);
  }
  return y;
}
// ---- point sums with four lanes per addition
// A tree level of a point sum leaves most lanes idle, and a lone warp pays the full latency of the nine dependent
// field multiplications of an addition.  Here the four lanes of a quad (lanes 4g..4g+3) hold the same P and Q and
// share one addition: each lane computes one of the four products of either half of add-2008-hwcd-3, the halves meet
// through shuffles -- three multiplications deep instead of nine.  Every lane of the warp must call.
__device__ __forceinline__ fp_t shfl_fp(const fp_t& x, int src) {
  fp_t y;
#pragma unroll
  for (int i = 0; i < 8; i++) y.v[i] = __shfl_sync(0xffffffffu, x.v[i], src);
  return y;
}
__device__ __forceinline__ ge_t shfl_ge(const ge_t& x, int src) {
  ge_t y;
  y.X = shfl_fp(x.X, src); y.Y = shfl_fp(x.Y, src); y.Z = shfl_fp(x.Z, src); y.T = shfl_fp(x.T, src);
  return y;
}
__device__ __forceinline__ fp_t fp_sel(bool c, const fp_t& a, const fp_t& b) {
  fp_t r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = c ? a.v[i] : b.v[i];
  return r;
}
static __device__ __noinline__ ge_t ge_add_quad(ge_t p, ge_t q) {
  const int lane = threadIdx.x & 31, role = lane & 3, base = lane & ~3;
  // first half: role 0: A = (Y1-X1)(Y2-X2), 1: B = (Y1+X1)(Y2+X2), 2: D = 2 Z1 Z2, 3: C = 2d T1 T2
  fp_t a = role == 0 ? fp_sub(p.Y, p.X) : role == 1 ? fp_add(p.Y, p.X) : role == 2 ? p.Z : p.T;
  fp_t b = role == 0 ? fp_sub(q.Y, q.X) : role == 1 ? fp_add(q.Y, q.X) : role == 2 ? q.Z : q.T;
  // products inlined: through a call a dependent product costs a lone warp 559 cycles instead of 467 (tools/microbench/latency_bench.cu)
  fp_t m = fp_mul(a, b);
  m = fp_mul(m, fp_sel(role == 3, fp_2d(), fp_one()));
  m = fp_sel(role == 2, fp_add(m, m), m);
  const fp_t A = shfl_fp(m, base), B = shfl_fp(m, base + 1), D = shfl_fp(m, base + 2), C = shfl_fp(m, base + 3);
  const fp_t E = fp_sub(B, A), Fv = fp_sub(D, C), G = fp_add(D, C), H = fp_add(B, A);
  // second half: role 0: X = E F, 1: Y = G H, 2: Z = F G, 3: T = E H
  a = fp_sel((role & 1) == 0, fp_sel(role == 0, E, Fv), fp_sel(role == 1, G, E));
  b = fp_sel((role & 1) == 0, fp_sel(role == 0, Fv, G), H);
  m = fp_mul(a, b);
  ge_t r;
  r.X = shfl_fp(m, base); r.Y = shfl_fp(m, base + 1); r.Z = shfl_fp(m, base + 2); r.T = shfl_fp(m, base + 3);
  return r;
}
// Sum of the values held by lanes [0, count) of a warp (the other lanes must hold the identity); valid in lane 0
// (in every lane of quad 0).  Every lane of the warp must call, with the same count.
__device__ __forceinline__ ge_t warp_sum_ge_n(const ge_t& x, uint32_t count) {
  const int g = (threadIdx.x & 31) >> 2;
  ge_t y;
  if (count > 16) {  // 32 -> 16 in two passes of eight additions, then each quad adds its own two sums
    ge_t r0 = ge_add_quad(shfl_ge(x, g), shfl_ge(x, g + 16));
    y = ge_add_quad(shfl_ge(x, g + 8), shfl_ge(x, g + 24));
    y = ge_add_quad(r0, y);
  } else if (count > 8) {
    y = ge_add_quad(shfl_ge(x, g), shfl_ge(x, g + 8));
  } else {
    y = shfl_ge(x, g);
  }
  // quads [0, min(count, 8)) hold one value each; the levels that could only add identities are skipped
  const uint32_t q = count < 8 ? count : 8;
#pragma unroll 1
  for (int o = 16; o >= 4; o >>= 1)
    if ((uint32_t)(o >> 2) < q) y = ge_add_quad(y, shfl_down_ge(y, o));
  return y;
}
// Commitment rows: block (split, row) sums terms [split*B*tpt, (split+1)*B*tpt) of its row.  A thread
// walks the windows of one term at a time, fetching the table entry of window w+1 while it adds the
// entry of window w; partial sums are combined by warp shuffles, then across the (<= 8) warps.  The
// blind*h term of a row is spread over the lanes of warp 0 of split 0, one window per lane.
__global__ void __launch_bounds__(256, 2) msm_fixed_kernel(const niels_t* __restrict__ tbl, uint32_t c, uint32_t windows,
                                                           const fl_t* __restrict__ scalars, size_t stride, uint32_t n,
                                                           const fl_t* __restrict__ extra_h, uint32_t h_index, uint32_t tpt,
                                                           ge_t* partial) {
  __shared__ ge_t sm[8];
  const uint32_t row = blockIdx.y, split = blockIdx.x, nsplit = gridDim.x;
  const size_t epb = (size_t)windows << (c - 1);
  const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  ge_t acc = ge_identity();
  const uint32_t j0 = split * blockDim.x * tpt;
  for (uint32_t k = 0; k < tpt; k++) {
    const uint32_t j = j0 + k * blockDim.x + threadIdx.x;
    if (j >= n) break;
    uint32_t kw[8];
    {
      const uint4* q = reinterpret_cast<const uint4*>(scalars + (size_t)row * stride + j);
      uint4 a = __ldg(q), b = __ldg(q + 1);
      kw[0] = a.x; kw[1] = a.y; kw[2] = a.z; kw[3] = a.w; kw[4] = b.x; kw[5] = b.y; kw[6] = b.z; kw[7] = b.w;
    }
    if ((kw[0] | kw[1] | kw[2] | kw[3] | kw[4] | kw[5] | kw[6] | kw[7]) == 0) continue;
    const niels_t* tb = tbl + (size_t)j * epb;
    // The entry of window w+1 is requested while window w is added.  Nothing may touch the loaded words before
    // their window comes up (a sign fix-up at load time made every request wait for its own data: 14 % of the
    // kernel's stall samples in profiles/r01_ncu_full_2p22_v5), so the sign of the digit travels to ge_madd.
    // Windows above the scalar's top bit (plus one for the recoding carry) hold zero digits: a short scalar -- a bit, a
    // 32-bit word, an address or a timestamp of the SNARK-mode polynomials -- stops after its own windows.
    uint32_t bitlen = 1;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (kw[i] != 0) bitlen = 32u * i + 32u - (uint32_t)__clz(kw[i]);
    uint32_t wn = (bitlen - 1) / c + 2;
    wn = wn < windows ? wn : windows;
    uint32_t carry = 0;
    int32_t d_next = msm_digit(kw, 0, c, carry);
    niels_t e_next;
    if (d_next != 0) e_next = ldg_niels(tb + ((uint32_t)(d_next < 0 ? -d_next : d_next) - 1));
    for (uint32_t w = 0; w < wn; w++) {
      const int32_t d = d_next;
      const niels_t e = e_next;
      if (w + 1 < wn) {
        d_next = msm_digit(kw, w + 1, c, carry);
        if (d_next != 0) e_next = ldg_niels(tb + ((size_t)(w + 1) << (c - 1)) + ((uint32_t)(d_next < 0 ? -d_next : d_next) - 1));
      }
      if (d != 0) acc = ge_madd(acc, e, d < 0);
    }
  }
  if (extra_h != nullptr && split == 0 && wid == 0 && lane < windows) {
    uint32_t kw[8];
    const uint4* q = reinterpret_cast<const uint4*>(extra_h + row);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    kw[0] = a.x; kw[1] = a.y; kw[2] = a.z; kw[3] = a.w; kw[4] = b.x; kw[5] = b.y; kw[6] = b.z; kw[7] = b.w;
    for (uint32_t w = lane; w < windows; w += 32) {
      int32_t d = msm_digit_at(kw, w, c);
      if (d != 0)
        acc = ge_madd(acc, ldg_niels(tbl + (size_t)h_index * epb + ((size_t)w << (c - 1)) + ((uint32_t)(d < 0 ? -d : d) - 1)), d < 0);
    }
  }
  acc = warp_sum_ge_n(acc, 32);
  if (nw > 1) {
    if (lane == 0) sm[wid] = acc;
    __syncthreads();
    if (wid != 0) return;
    acc = lane < nw ? sm[lane] : ge_identity();
    acc = warp_sum_ge_n(acc, nw);
  }
  if (threadIdx.x == 0) partial[(size_t)row * nsplit + split] = acc;
}
// Sums the per-split partials of a row, optionally adds the row's blind*h term (one table look-up per window --
// kept out of the row kernel so that the commitment can start before the RandomTape has produced the blinds),
// and compresses.  One warp per row: the windows of blind*h and the partials are spread over the lanes and summed
// by the quad tree, so the chain in front of the compression is ~25 field multiplications instead of 7 per window.
// The compression is a chain of ~280 dependent F_p products that costs a warp the same with one active lane as with 32,
// so the W row sums of a block meet in shared memory and ONE warp compresses them, a row per lane: with 2048 rows the
// chip runs 128 such chains side by side instead of 2048 (3.5 per SM sub-partition, each slowing the others down).
// W grows with the row count so that the blocks stay at about one per SM (finish_warps()).
template <int W>
__global__ void __launch_bounds__(32 * W) msm_finish_kernel(const ge_t* __restrict__ partial, uint32_t rows, uint32_t nsplit,
                                                            uint8_t* __restrict__ out32, const niels_t* __restrict__ tbl, uint32_t c,
                                                            uint32_t windows, const fl_t* __restrict__ extra_h, uint32_t h_index) {
  __shared__ ge_t srow[W];
  const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5, row = blockIdx.x * W + wid;
  if (row < rows) {  // whole warps
    const uint32_t nh = extra_h != nullptr ? windows : 0, items = nh + nsplit;
    ge_t acc = ge_identity();
    if (lane < nh) {
      const size_t epb = (size_t)windows << (c - 1);
      uint32_t kw[8];
      const uint4* q = reinterpret_cast<const uint4*>(extra_h + row);
      uint4 a = __ldg(q), b = __ldg(q + 1);
      kw[0] = a.x; kw[1] = a.y; kw[2] = a.z; kw[3] = a.w; kw[4] = b.x; kw[5] = b.y; kw[6] = b.z; kw[7] = b.w;
      for (uint32_t w = lane; w < nh; w += 32) {
        int32_t d = msm_digit_at(kw, w, c);
        if (d != 0) {
          bool neg = d < 0;
          niels_t e = ldg_niels(tbl + (size_t)h_index * epb + ((size_t)w << (c - 1)) + ((uint32_t)(neg ? -d : d) - 1));
          acc = ge_madd_call(acc, e, neg);
        }
      }
    }
    // item i (windows first, then the partials) belongs to lane i mod 32
    for (uint32_t i = nh + ((lane + 32 - (nh & 31)) & 31); i < items; i += 32) acc = ge_add_call(acc, partial[(size_t)row * nsplit + (i - nh)]);
    acc = warp_sum_ge_n(acc, items < 32 ? items : 32);
    if (lane == 0) srow[wid] = acc;
  }
  __syncthreads();
  // co-resident blocks (b and b + 148 under round-robin placement) pick different warps, i.e. different sub-partitions
  const uint32_t cw = (blockIdx.x / kSMs) % W;
  if (wid != cw || lane >= W) return;
  const uint32_t crow = blockIdx.x * W + lane;
  if (crow >= rows) return;
  uint32_t w[8];
  ge_compress_t<FpSqrInl>(srow[lane], w);
  uint4* o = reinterpret_cast<uint4*>(out32 + (size_t)crow * 32);
  o[0] = make_uint4(w[0], w[1], w[2], w[3]);
  o[1] = make_uint4(w[4], w[5], w[6], w[7]);
}
// warps (= rows) per block of msm_finish_kernel: 4 up to 592 rows, then 8, then 16.  SB200_FINISH_WARPS overrides.
static int finish_warps(uint32_t rows) {
  static const int forced = [] { const char* e = getenv("SB200_FINISH_WARPS"); int v = e ? atoi(e) : 0; return v == 4 || v == 8 || v == 16 ? v : 0; }();
  if (forced) return forced;
  return rows <= 4u * kSMs ? 4 : rows <= 8u * kSMs ? 8 : 16;
}
static cudaError_t launch_msm_finish(const ge_t* partial, uint32_t rows, uint32_t nsplit, uint8_t* out32, const niels_t* tbl, uint32_t c,
                                     uint32_t windows, const fl_t* extra_h, uint32_t h_index, cudaStream_t s) {
  const int W = finish_warps(rows);
  KProf kp_("msm_finish_kernel", s);
  if (W == 4) msm_finish_kernel<4><<<(rows + 3) / 4, 128, 0, s>>>(partial, rows, nsplit, out32, tbl, c, windows, extra_h, h_index);
  else if (W == 8) msm_finish_kernel<8><<<(rows + 7) / 8, 256, 0, s>>>(partial, rows, nsplit, out32, tbl, c, windows, extra_h, h_index);
  else msm_finish_kernel<16><<<(rows + 15) / 16, 512, 0, s>>>(partial, rows, nsplit, out32, tbl, c, windows, extra_h, h_index);
  return cudaSuccess;
}
MsmPlan msm_plan(uint32_t rows, uint32_t n) {
  MsmPlan p;
  p.rows = rows;
  p.n = n;
  uint32_t block = 32;
  while (block < 256 && block < n) block <<= 1;
  p.block = block;
  // terms per thread: enough blocks for ~6 waves of 2 x 256 threads per SM, at most 8 terms per thread
  size_t total = (size_t)rows * n, per_wave = (size_t)kSMs * 512 * 6;
  uint32_t tpt = 1;
  while (tpt < 8 && (size_t)tpt * 2 * per_wave <= total) tpt <<= 1;
  p.tpt = tpt;
  p.splits = (n + block * tpt - 1) / (block * tpt);
  if (p.splits < 1) p.splits = 1;
  return p;
}
size_t msm_scratch_bytes(uint32_t rows, uint32_t n) {
  MsmPlan p = msm_plan(rows, n);
  return (size_t)rows * p.splits * sizeof(ge_t);
}
cudaError_t msm_fixed(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t rows, uint32_t n, const fl_t* extra_h,
                      const fl_t* extra_q, uint32_t h_index, uint32_t q_index, void* scratch, uint8_t* out32,
                      cudaStream_t s, cudaEvent_t* ev) {
  (void)extra_q;
  (void)q_index;
  if (!rows) return cudaSuccess;
  MsmPlan p = msm_plan(rows, n);
  dim3 grid(p.splits, rows);
  if (ev) cudaEventRecord(ev[0], s);
  { KProf kp_("msm_fixed_kernel", s); msm_fixed_kernel<<<grid, p.block, 0, s>>>(t.tbl, t.c, t.windows, scalars, stride, n, extra_h, h_index, p.tpt,
                                             reinterpret_cast<ge_t*>(scratch)); }
  cudaError_t e = SB_LAUNCHED();
  if (ev) cudaEventRecord(ev[1], s);
  if (e != cudaSuccess) return e;
  launch_msm_finish(reinterpret_cast<const ge_t*>(scratch), rows, p.splits, out32, t.tbl, t.c, t.windows, nullptr, 0, s);
  return SB_LAUNCHED();
}
// The same in two halves: the row sums first (no blinds needed), then blind*h + compression.
cudaError_t msm_fixed_rows(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t rows, uint32_t n, void* scratch,
                           cudaStream_t s, cudaEvent_t* ev) {
  if (!rows) return cudaSuccess;
  MsmPlan p = msm_plan(rows, n);
  dim3 grid(p.splits, rows);
  if (ev) cudaEventRecord(ev[0], s);
  { KProf kp_("msm_fixed_kernel", s); msm_fixed_kernel<<<grid, p.block, 0, s>>>(t.tbl, t.c, t.windows, scalars, stride, n, nullptr, 0, p.tpt,
                                             reinterpret_cast<ge_t*>(scratch)); }
  cudaError_t e = SB_LAUNCHED();
  if (ev) cudaEventRecord(ev[1], s);
  return e;
}
// Rows [row0, row0 + nrows) of a `total_rows`-row job, with the launch plan and the scratch layout of the whole job:
// the witness can then be committed in pieces, each launched when its part of the host->device copy has landed.
cudaError_t msm_fixed_rows_part(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t total_rows, uint32_t row0,
                                uint32_t nrows, uint32_t n, void* scratch, cudaStream_t s) {
  if (!nrows) return cudaSuccess;
  MsmPlan p = msm_plan(total_rows, n);
  dim3 grid(p.splits, nrows);
  { KProf kp_("msm_fixed_kernel", s); msm_fixed_kernel<<<grid, p.block, 0, s>>>(t.tbl, t.c, t.windows, scalars + (size_t)row0 * stride, stride, n, nullptr, 0, p.tpt,
                                             reinterpret_cast<ge_t*>(scratch) + (size_t)row0 * p.splits); }
  return SB_LAUNCHED();
}
cudaError_t msm_fixed_finish(const MsmTable& t, uint32_t rows, uint32_t n, const fl_t* extra_h, uint32_t h_index, const void* scratch,
                             uint8_t* out32, cudaStream_t s) {
  if (!rows) return cudaSuccess;
  MsmPlan p = msm_plan(rows, n);
  launch_msm_finish(reinterpret_cast<const ge_t*>(scratch), rows, p.splits, out32, t.tbl, t.c, t.windows, extra_h, h_index, s);
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ small fixed-base MSM (latency path)
// The bullet reduction and the one-row commitments of the evaluation proof are a handful of MSMs
// of R+2 terms on the critical path of a Fiat-Shamir round: what matters is latency, not
// throughput.  Here every (term, window) pair is its own unit of work -- a thread does one or two
// table look-ups -- and the points are summed by a shuffle tree, a shared-memory tree and a
// last-block-done tree.  The result is left in EXTENDED coordinates (optionally in mapped host
// memory, followed by a flag): one point compression is ~280 dependent field multiplications,
// which a host core does in ~4 us and a lone GPU thread in ~40 us.
// valid in thread 0; smem: ge_t[blockDim.x / 32].  The second level only runs the shuffle steps that can
// still carry data (log2 of the number of warps): every skipped step is a full point addition of latency.
__device__ __forceinline__ ge_t block_sum_ge(ge_t x, ge_t* smem) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  x = warp_sum_ge_n(x, 32);
  __syncthreads();
  if (lane == 0) smem[wid] = x;
  __syncthreads();
  if (wid == 0) {
    x = lane < nw ? smem[lane] : ge_identity();
    x = warp_sum_ge_n(x, (uint32_t)nw);
  }
  return x;
}
// sum of `count` <= blockDim.x values, one per thread (identity elsewhere); valid in thread 0
__device__ __forceinline__ ge_t block_sum_ge_n(ge_t x, uint32_t count, ge_t* smem) {
  if (count <= 32) {  // all of them sit in warp 0
    if (threadIdx.x < 32) x = warp_sum_ge_n(x, count);
    return x;
  }
  return block_sum_ge(x, smem);
}
__global__ void __launch_bounds__(256) msm_small_kernel(const niels_t* __restrict__ tbl, uint32_t c, uint32_t windows,
                                                        const fl_t* __restrict__ scalars, size_t stride, uint32_t n,
                                                        ge_t* partial, unsigned int* counters, ge_t* out,
                                                        uint32_t* host_flag, uint32_t seq) {
  __shared__ ge_t smem[8];
  __shared__ bool s_last;
  const uint32_t row = blockIdx.y, rows = gridDim.y;
  const size_t epb = (size_t)windows << (c - 1);
  const uint32_t pairs = n * windows;
  ge_t acc = ge_identity();
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < pairs; p += gridDim.x * blockDim.x) {
    uint32_t j = p / windows, w = p - j * windows;
    uint32_t k[8];
    const uint4* q = reinterpret_cast<const uint4*>(scalars + (size_t)row * stride + j);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    k[0] = a.x; k[1] = a.y; k[2] = a.z; k[3] = a.w; k[4] = b.x; k[5] = b.y; k[6] = b.z; k[7] = b.w;
    if ((k[0] | k[1] | k[2] | k[3] | k[4] | k[5] | k[6] | k[7]) == 0) continue;
    int32_t d = msm_digit_at(k, w, c);
    if (d != 0) {
      bool neg = d < 0;
      uint32_t mag = (uint32_t)(neg ? -d : d);
      niels_t e = ldg_niels(tbl + (size_t)j * epb + ((size_t)w << (c - 1)) + (mag - 1));
      acc = ge_madd_call(acc, e, neg);
    }
  }
  acc = block_sum_ge(acc, smem);
  if (threadIdx.x == 0) {
    partial[(size_t)row * gridDim.x + blockIdx.x] = acc;
    __threadfence();
    unsigned int t = atomicAdd(&counters[row], 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  acc = ge_identity();
  for (uint32_t b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
    const uint4* src = reinterpret_cast<const uint4*>(&partial[(size_t)row * gridDim.x + b]);
    ge_t tmp;
    uint32_t* tw = reinterpret_cast<uint32_t*>(&tmp);
#pragma unroll
    for (int i = 0; i < 8; i++) {
      uint4 v = __ldcg(src + i);
      tw[4 * i] = v.x; tw[4 * i + 1] = v.y; tw[4 * i + 2] = v.z; tw[4 * i + 3] = v.w;
    }
    acc = b < blockDim.x && b == threadIdx.x ? tmp : ge_add_call(acc, tmp);
  }
  acc = block_sum_ge_n(acc, gridDim.x, smem);
  if (threadIdx.x == 0) {
    out[row] = acc;
    counters[row] = 0;
    if (host_flag != nullptr) {
      __threadfence_system();
      unsigned int t = atomicAdd(&counters[rows], 1u);  // rows finished
      if (t == rows - 1) {
        counters[rows] = 0;
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(host_flag) = seq;
      }
    }
  }
}
// Threads per block of the small MSM.  One warp per SM sub-partition: a dependent F_p product costs a warp 0.23 us when
// it has the sub-partition's integer pipe to itself and 0.50 us when two warps share it (tools/microbench/roundtrip_bench.cu:
// 16 dependent products in a 32-thread block against a 256-thread block), and this kernel is a tree of dependent
// additions.  SB200_MSM_SMALL_THREADS overrides (256 = the shape before).
static uint32_t msm_small_threads() {
  static const uint32_t v = [] {
    const char* e = getenv("SB200_MSM_SMALL_THREADS");
    int t = e ? atoi(e) : 128;
    return (uint32_t)(t == 32 || t == 64 || t == 128 || t == 256 ? t : 128);
  }();
  return v;
}
uint32_t msm_small_blocks(uint32_t rows, uint32_t n, uint32_t windows) {
  uint32_t pairs = n * windows;
  // SB200_MSM_SMALL_UNITS (experiment knob): (term, window) units per thread, default 2; SB200_MSM_SMALL_CAP: blocks per SM, default 2
  static const uint32_t upt = [] { const char* e = getenv("SB200_MSM_SMALL_UNITS"); int v = e ? atoi(e) : 2; return (uint32_t)(v >= 1 && v <= 16 ? v : 2); }();
  static const uint32_t bps = [] { const char* e = getenv("SB200_MSM_SMALL_CAP"); int v = e ? atoi(e) : 2; return (uint32_t)(v >= 1 && v <= 4 ? v : 2); }();
  const uint32_t per_block = upt * msm_small_threads();
  uint32_t want = (pairs + per_block - 1) / per_block;  // ~2 look-ups per thread
  uint32_t cap = (uint32_t)(kSMs * bps) / (rows ? rows : 1);
  if (cap < 1) cap = 1;
  if (want < 1) want = 1;
  return want < cap ? want : cap;
}
size_t msm_small_scratch_bytes(uint32_t rows, uint32_t n, uint32_t windows) {
  return (size_t)rows * msm_small_blocks(rows, n, windows) * sizeof(ge_t);
}
cudaError_t msm_small(const MsmTable& t, const fl_t* scalars, size_t stride, uint32_t rows, uint32_t n, void* scratch,
                      unsigned int* counters, void* out_ext, uint32_t* host_flag, uint32_t seq, cudaStream_t s) {
  if (!rows) return cudaSuccess;
  dim3 grid(msm_small_blocks(rows, n, t.windows), rows);
  { KProf kp_("msm_small_kernel", s); msm_small_kernel<<<grid, msm_small_threads(), 0, s>>>(t.tbl, t.c, t.windows, scalars, stride, n, reinterpret_cast<ge_t*>(scratch), counters,
                                        reinterpret_cast<ge_t*>(out_ext), host_flag, seq); }
  return SB_LAUNCHED();
}
// up to four scalars (the first element of each source) to a host mailbox, then the flag
__global__ void mail_kernel(const fl_t* p0, const fl_t* p1, const fl_t* p2, const fl_t* p3, int n, fl_t* host, uint32_t* host_flag,
                            uint32_t seq) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const fl_t* src[4] = {p0, p1, p2, p3};
    for (int i = 0; i < n; i++) st_fl(host + i, ld_fl(src[i]));
    __threadfence_system();
    *reinterpret_cast<volatile uint32_t*>(host_flag) = seq;
  }
}
cudaError_t k_mail(const fl_t* p0, const fl_t* p1, const fl_t* p2, const fl_t* p3, int n, fl_t* host, uint32_t* host_flag,
                   uint32_t seq, cudaStream_t s) {
  { KProf kp_("mail_kernel", s); mail_kernel<<<1, 32, 0, s>>>(p0, p1, p2, p3, n, host, host_flag, seq); }
  return SB_LAUNCHED();
}
// the first `count` entries of up to four tables to a host mailbox (table k at host + k*count), then the flag
__global__ void mail_tables_kernel(const fl_t* p0, const fl_t* p1, const fl_t* p2, const fl_t* p3, int ntab, uint32_t count,
                                   fl_t* host, uint32_t* host_flag, uint32_t seq) {
  const fl_t* src[4] = {p0, p1, p2, p3};
  for (uint32_t i = threadIdx.x; i < (uint32_t)ntab * count; i += blockDim.x) st_fl(host + i, ld_fl(src[i / count] + (i % count)));
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    *reinterpret_cast<volatile uint32_t*>(host_flag) = seq;
  }
}
cudaError_t k_mail_tables(const fl_t* p0, const fl_t* p1, const fl_t* p2, const fl_t* p3, int ntab, uint32_t count, fl_t* host,
                          uint32_t* host_flag, uint32_t seq, cudaStream_t s) {
  { KProf kp_("mail_tables_kernel", s); mail_tables_kernel<<<1, 256, 0, s>>>(p0, p1, p2, p3, ntab, count, host, host_flag, seq); }
  return SB_LAUNCHED();
}

// ------------------------------------------------------------------ table build
// tbl[base][w][m-1] = m * 2^(c*w) * G_base as affine Niels, in two kernels:
//   msm_winbase_kernel  thread per base: B_w = 2^(c*w) * G for every window w (252 doublings once per base, instead of
//                       c*w doublings in every thread of the second kernel -- that was 40 % of the build);
//   msm_table_kernel    thread per (base, window, chunk of kTblChunk multiples): m0 * B_w by double-and-add from the
//                       top set bit of m0, then the run of multiples, one shared inversion per chunk.
static constexpr int kTblChunk = 64;
__global__ void __launch_bounds__(64) msm_winbase_kernel(const fp_t* __restrict__ affine_xy, uint32_t nbases, uint32_t c,
                                                         uint32_t windows, ge_t* __restrict__ winbase) {
  const uint32_t base = blockIdx.x * blockDim.x + threadIdx.x;
  if (base >= nbases) return;
  ge_t P;
  P.X = ldg_fp(affine_xy + 2 * (size_t)base);
  P.Y = ldg_fp(affine_xy + 2 * (size_t)base + 1);
  P.Z = fp_one();
  P.T = fp_mul_call(P.X, P.Y);
  for (uint32_t w = 0; w < windows; w++) {
    winbase[(size_t)base * windows + w] = P;
    if (w + 1 < windows)
      for (uint32_t k = 0; k < c; k++) P = ge_double_call(P);
  }
}
__global__ void __launch_bounds__(64) msm_table_kernel(const ge_t* __restrict__ winbase, uint32_t nbases, uint32_t c,
                                                       uint32_t windows, niels_t* __restrict__ tbl) {
  const uint32_t epw = 1u << (c - 1);
  const uint32_t ch = epw < (uint32_t)kTblChunk ? epw : (uint32_t)kTblChunk;
  const uint32_t chunks = epw / ch;
  size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)nbases * windows * chunks;
  if (tid >= total) return;
  uint32_t chunk = (uint32_t)(tid % chunks);
  uint32_t w = (uint32_t)((tid / chunks) % windows);
  uint32_t base = (uint32_t)(tid / ((size_t)chunks * windows));
  const ge_t P = winbase[(size_t)base * windows + w];
  // M = (chunk*ch + 1) * P
  uint32_t m0 = chunk * ch + 1;
  ge_t M = P;
  for (int bit = 30 - __clz(m0); bit >= 0; bit--) {
    M = ge_double_call(M);
    if ((m0 >> bit) & 1u) M = ge_add_call(M, P);
  }
  niels_t* slot = tbl + ((size_t)base * windows + w) * epw + (size_t)chunk * ch;
  fp_t pre[kTblChunk];
  fp_t acc = fp_one();
  for (uint32_t k = 0; k < ch; k++) {
    pre[k] = acc;
    acc = fp_mul_call(acc, M.Z);
    st_fp(&slot[k].yplusx, M.X);
    st_fp(&slot[k].yminusx, M.Y);
    st_fp(&slot[k].xy2d, M.Z);
    if (k + 1 < ch) M = ge_add_call(M, P);
  }
  fp_t inv = fp_invert_t<FpCall>(acc);
  for (int k = (int)ch - 1; k >= 0; k--) {
    fp_t X = slot[k].yplusx, Y = slot[k].yminusx, Z = slot[k].xy2d;
    fp_t zi = fp_mul_call(inv, pre[k]);
    inv = fp_mul_call(inv, Z);
    niels_t e = niels_from_affine(fp_mul_call(X, zi), fp_mul_call(Y, zi));
    st_fp(&slot[k].yplusx, e.yplusx);
    st_fp(&slot[k].yminusx, e.yminusx);
    st_fp(&slot[k].xy2d, e.xy2d);
  }
}
cudaError_t msm_table_build(MsmTable& t, const uint8_t* affine_xy_host, uint32_t nbases, uint32_t c, cudaStream_t s) {
  t.nbases = nbases;
  t.c = c;
  t.windows = msm_windows(c);
  size_t bytes = (size_t)nbases * t.entries_per_base() * sizeof(niels_t);
  cudaError_t e = cudaMalloc(&t.tbl, bytes);
  if (e != cudaSuccess) return e;
  fp_t* d_xy = nullptr;
  ge_t* d_wb = nullptr;
  e = cudaMalloc(&d_xy, (size_t)nbases * 64);
  if (e != cudaSuccess) return e;
  e = cudaMalloc(&d_wb, (size_t)nbases * t.windows * sizeof(ge_t));
  if (e != cudaSuccess) { cudaFree(d_xy); return e; }
  e = cudaMemcpyAsync(d_xy, affine_xy_host, (size_t)nbases * 64, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) {
    { KProf kp_("msm_winbase_kernel", s); msm_winbase_kernel<<<(nbases + 63) / 64, 64, 0, s>>>(d_xy, nbases, c, t.windows, d_wb); }
    e = SB_LAUNCHED();
  }
  if (e == cudaSuccess) {
    uint32_t epw = 1u << (c - 1), ch = epw < (uint32_t)kTblChunk ? epw : (uint32_t)kTblChunk;
    size_t total = (size_t)nbases * t.windows * (epw / ch);
    { KProf kp_("msm_table_kernel", s); msm_table_kernel<<<(unsigned)((total + 63) / 64), 64, 0, s>>>(d_wb, nbases, c, t.windows, t.tbl); }
    e = SB_LAUNCHED();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  cudaFree(d_xy);
  cudaFree(d_wb);
  return e;
}
void msm_table_free(MsmTable& t) {
  if (t.tbl) cudaFree(t.tbl);
  t.tbl = nullptr;
}

}  // namespace sb
