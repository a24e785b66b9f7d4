// chq_numa_bind_to_device: puts the calling thread — its CPU affinity and the memory policy of everything it allocates
// from now on, pinned staging buffers included — on the NUMA node the GPU's PCIe root hangs off.  A broker that runs
// one routing thread (or process) per GPU calls it once before chq_table_create: with eight GPUs on a two-socket
// box, result buffers that all live on socket 0 make half of the device-to-host copies cross the socket link.
// No counterpart in the reference (it has no device code).  Plain Linux: sysfs + the set_mempolicy system call
// (no libnuma in this image).
#include <cuda_runtime.h>
#include <sched.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/chanamq_gpu.h"

namespace {

bool read_line(const std::string& path, std::string& out) {
  FILE* f = fopen(path.c_str(), "r");
  if (!f) return false;
  char buf[4096];
  const bool ok = fgets(buf, sizeof(buf), f) != nullptr;
  fclose(f);
  if (ok) {
    out = buf;
    while (!out.empty() && isspace((unsigned char)out.back())) out.pop_back();
  }
  return ok;
}

}  // namespace

extern "C" int32_t chq_numa_bind_to_device(int32_t device, int32_t* node_out, int32_t* n_cpus_out) {
  if (node_out) *node_out = -1;
  if (n_cpus_out) *n_cpus_out = 0;
  char bus[64] = {};
  if (cudaDeviceGetPCIBusId(bus, sizeof(bus), device) != cudaSuccess) { cudaGetLastError(); return CHQ_E_CUDA; }
  for (char* c = bus; *c; ++c) *c = (char)tolower((unsigned char)*c);
  std::string line;
  if (!read_line(std::string("/sys/bus/pci/devices/") + bus + "/numa_node", line)) return CHQ_OK;   // nothing known: leave as is
  const int node = atoi(line.c_str());
  if (node < 0 || node >= 1024) return CHQ_OK;   // single-node machine (or a VM that hides the topology)
  // the node's CPUs, intersected with what this thread may use (a container's cpuset)
  cpu_set_t allowed, want;
  CPU_ZERO(&want);
  if (sched_getaffinity(0, sizeof(allowed), &allowed) != 0) return CHQ_OK;
  if (read_line("/sys/devices/system/node/node" + std::to_string(node) + "/cpulist", line)) {
    const char* p = line.c_str();
    while (*p) {   // "0-31,64-95"
      char* end = nullptr;
      long a = strtol(p, &end, 10), b = a;
      if (end == p) break;
      p = end;
      if (*p == '-') { b = strtol(p + 1, &end, 10); p = end; }
      for (long c = a; c <= b && c < CPU_SETSIZE; ++c)
        if (CPU_ISSET((int)c, &allowed)) CPU_SET((int)c, &want);
      if (*p == ',') ++p;
    }
  }
  const int n_cpus = CPU_COUNT(&want);
  if (n_cpus > 0) sched_setaffinity(0, sizeof(want), &want);
  // MPOL_PREFERRED (1): allocate on that node while it has memory, fall back instead of failing
  unsigned long mask[16] = {};
  mask[node / (8 * sizeof(unsigned long))] |= 1ul << (node % (8 * sizeof(unsigned long)));
  syscall(SYS_set_mempolicy, 1, mask, (unsigned long)(sizeof(mask) * 8));
  if (node_out) *node_out = node;
  if (n_cpus_out) *n_cpus_out = n_cpus;
  return CHQ_OK;
}
