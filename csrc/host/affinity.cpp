// CPU affinity and physical-core discovery for the host runtime (reference: C/native-dnn/src/main/c/affinity.c —
// sched_setaffinity/getaffinity over JNI — and native_utils.c's physical core count, used by Engine to size and pin its
// thread pools; S/utils/DistriParameterSynchronizer.scala:128-144 pins the communication pools to the last cores).
// Here the pinned threads are the data-loader / record-IO workers that feed one GPU per process.
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include <sched.h>
#include <unistd.h>

#include <cstdio>
#include <set>
#include <utility>

#include "bigdl_b200_host.h"

extern "C" {

// Pin the calling thread to `n` logical CPUs. Returns 0 on success, -1 on failure (errno set by the kernel).
int bdlh_set_affinity(const int* cpus, int n) {
  cpu_set_t set;
  CPU_ZERO(&set);
  for (int i = 0; i < n; ++i)
    if (cpus[i] >= 0 && cpus[i] < CPU_SETSIZE) CPU_SET(cpus[i], &set);
  return sched_setaffinity(0, sizeof(set), &set);
}

// Writes up to `cap` logical CPU ids the calling thread may run on; returns how many there are.
int bdlh_get_affinity(int* cpus, int cap) {
  cpu_set_t set;
  CPU_ZERO(&set);
  if (sched_getaffinity(0, sizeof(set), &set) != 0) return -1;
  int n = 0;
  for (int c = 0; c < CPU_SETSIZE; ++c)
    if (CPU_ISSET(c, &set)) {
      if (n < cap) cpus[n] = c;
      ++n;
    }
  return n;
}

int bdlh_logical_cores(void) {
  const long n = sysconf(_SC_NPROCESSORS_ONLN);
  return n > 0 ? static_cast<int>(n) : 1;
}

// Distinct (package, core) pairs from sysfs; falls back to the logical count when the topology files are missing
// (containers often hide them).
int bdlh_physical_cores(void) {
  std::set<std::pair<int, int>> cores;
  const int logical = bdlh_logical_cores();
  for (int c = 0; c < logical; ++c) {
    int ids[2] = {-1, -1};
    const char* names[2] = {"physical_package_id", "core_id"};
    for (int k = 0; k < 2; ++k) {
      char path[128];
      snprintf(path, sizeof(path), "/sys/devices/system/cpu/cpu%d/topology/%s", c, names[k]);
      FILE* f = fopen(path, "r");
      if (!f) return logical;
      const int got = fscanf(f, "%d", &ids[k]);
      fclose(f);
      if (got != 1) return logical;
    }
    cores.insert({ids[0], ids[1]});
  }
  return cores.empty() ? logical : static_cast<int>(cores.size());
}

}  // extern "C"
