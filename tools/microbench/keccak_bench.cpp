// Host microbenchmark: Keccak-f[1600] (portable and AVX-512) and one RandomTape draw (merlin challenge_scalar), the unit a
// sharded proof's commitment waits for 2^(s/2) times.  g++ -O3 -march=x86-64-v3 -o keccak_bench keccak_bench.cpp ../../otti_b200/csrc/host/keccak_avx512.cpp -mavx512f -mavx512vl (second file only)
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include "../../otti_b200/csrc/host/hfl.hpp"
#include "../../otti_b200/csrc/host/hmerlin.hpp"
using namespace sbh;
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
  uint64_t a[25], b[25];
  for (int i = 0; i < 25; i++) a[i] = b[i] = 0x9E3779B97F4A7C15ULL * (i + 1);
  const int N = 200000;
  double t0 = now();
  for (int i = 0; i < N; i++) keccakf_scalar(a);
  double t1 = now();
  bool wide = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512vl");
  if (wide) for (int i = 0; i < N; i++) keccakf_avx512(b);
  double t2 = now();
  printf("keccak-f portable %.3f us, avx512 %.3f us (%s), equal %d\n", (t1 - t0) / N * 1e6, (t2 - t1) / N * 1e6, wide ? "run" : "not supported",
         (int)(!wide || memcmp(a, b, 200) == 0));
  MerlinTranscript t("proof", 5);
  t.append_scalar("init_randomness", fl_one());
  Fl acc = fl_zero();
  t0 = now();
  for (int i = 0; i < 4096; i++) acc = fl_add(acc, t.challenge_scalar("poly_blinds"));
  t1 = now();
  printf("RandomTape draw (SB200_KECCAK=%s): %.3f us each, 2048 draws %.3f ms  [%lx]\n", getenv("SB200_KECCAK") ? getenv("SB200_KECCAK") : "",
         (t1 - t0) / 4096 * 1e6, (t1 - t0) / 2 * 1e3, (unsigned long)acc.v[0]);
  return 0;
}
