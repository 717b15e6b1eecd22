// One process per GPU on a multi-socket host: keep this process's host threads (the caller, the Fiat-Shamir worker
// pool it creates later, the RandomTape drawer) on the cores next to its GPU, and give the processes that share a
// socket disjoint slices of it.  The per-round host work is a few microseconds between mailbox polls; a thread that
// migrates to the other socket pays for every poll of the pinned mailbox and every cache line of the fixed-base tables.
// Applied by sb200_init when LOCAL_WORLD_SIZE > 1 (torchrun / one rank per GPU); SB200_BIND=0 turns it off,
// SB200_BIND=1 forces it for a single process.  Linux only; any failure leaves the affinity untouched.
#pragma once
#include <cuda_runtime.h>
#include <sched.h>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

namespace sbh {

static inline std::string pci_local_cpulist(int device) {
  char bus[32] = {0};
  if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) return "";
  for (char* p = bus; *p; p++) *p = (char)tolower(*p);
  std::string path = std::string("/sys/bus/pci/devices/") + bus + "/local_cpulist";
  FILE* f = fopen(path.c_str(), "r");
  if (!f) return "";
  char buf[4096] = {0};
  size_t n = fread(buf, 1, sizeof buf - 1, f);
  fclose(f);
  std::string s(buf, n);
  while (!s.empty() && isspace((unsigned char)s.back())) s.pop_back();
  return s;
}
static inline std::vector<int> parse_cpulist(const std::string& s) {  // "0-31,64-95"
  std::vector<int> out;
  size_t i = 0;
  while (i < s.size()) {
    char* end = nullptr;
    long a = strtol(s.c_str() + i, &end, 10);
    if (end == s.c_str() + i) break;
    long b = a;
    i = (size_t)(end - s.c_str());
    if (i < s.size() && s[i] == '-') {
      b = strtol(s.c_str() + i + 1, &end, 10);
      i = (size_t)(end - s.c_str());
    }
    for (long c = a; c <= b && c - a < 4096; c++) out.push_back((int)c);
    if (i < s.size() && s[i] == ',') i++;
  }
  return out;
}
// Returns the number of CPUs the calling thread was bound to (0 = left alone).
static inline int bind_near_device(int device, int ndevices) {
  const char* mode = getenv("SB200_BIND");
  if (mode && atoi(mode) == 0) return 0;
  const char* lw = getenv("LOCAL_WORLD_SIZE");
  const int ranks = lw ? atoi(lw) : 1;
  if (ranks <= 1 && !(mode && atoi(mode) == 1)) return 0;
  const std::string mine = pci_local_cpulist(device);
  std::vector<int> cpus = parse_cpulist(mine);
  if (cpus.empty()) return 0;
  // the devices that share this cpulist split it evenly, in device order (one rank per device)
  int group = 0, index = 0;
  for (int d = 0; d < ndevices && d < ranks; d++) {
    if (pci_local_cpulist(d) == mine) {
      if (d < device) index++;
      group++;
    }
  }
  if (group < 1) group = 1;
  // only CPUs this process may already use
  cpu_set_t cur;
  CPU_ZERO(&cur);
  if (sched_getaffinity(0, sizeof cur, &cur) != 0) return 0;
  std::vector<int> usable;
  for (int c : cpus)
    if (c < CPU_SETSIZE && CPU_ISSET(c, &cur)) usable.push_back(c);
  if (usable.size() < (size_t)group * 2) return 0;
  // "0-31,64-95" lists cores and then their hyper-thread siblings: deal the list out round-robin by position inside each
  // half so that a slice holds whole cores where the two halves are siblings, and contiguous cores otherwise
  const size_t per = usable.size() / (size_t)group;
  cpu_set_t set;
  CPU_ZERO(&set);
  int count = 0;
  const size_t half = usable.size() / 2, per_half = half / (size_t)group;
  if (per_half >= 1 && mine.find(',') != std::string::npos) {
    for (size_t k = 0; k < per_half; k++) {
      CPU_SET(usable[(size_t)index * per_half + k], &set);
      CPU_SET(usable[half + (size_t)index * per_half + k], &set);
      count += 2;
    }
  } else {
    for (size_t k = 0; k < per; k++) { CPU_SET(usable[(size_t)index * per + k], &set); count++; }
  }
  if (count < 2 || sched_setaffinity(0, sizeof set, &set) != 0) return 0;
  return count;
}

}  // namespace sbh
