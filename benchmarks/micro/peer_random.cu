// Random 8-byte reads of a table in ANOTHER GPU's memory (a cudaIpc mapping, as the symmetric arena
// of this framework uses) against the same reads of a local table, as a function of the table
// size: what one probe of a remote open-addressing table costs, and whether that depends on how
// much memory the probes spread over.  Also the same number of keys resolved by 32-byte reads.
// Needs 2 GPUs.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/micro/peer_random benchmarks/micro/peer_random.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <unistd.h>
#include <sys/wait.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}
// n threads, each `chain` dependent reads at random slots of table[cap] (cap a power of two)
__global__ void k_probe(const unsigned long long* table, uint64_t cap, int n, int chain, uint64_t seed, unsigned long long* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t h = mix64(seed + i);
  unsigned long long acc = 0;
  for (int c = 0; c < chain; ++c) {
    const unsigned long long v = *reinterpret_cast<const volatile unsigned long long*>(table + (h & (cap - 1)));
    acc += v;
    h = mix64(h + v + 1);
  }
  if (acc == 0x1234567) out[0] = acc;
}
__global__ void k_probe32(const ulonglong4* table, uint64_t cap4, int n, int chain, uint64_t seed, unsigned long long* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t h = mix64(seed + i);
  unsigned long long acc = 0;
  for (int c = 0; c < chain; ++c) {
    const ulonglong2* q = reinterpret_cast<const ulonglong2*>(table + (h & (cap4 - 1)));
    const ulonglong2 v = __ldcv(q), w = __ldcv(q + 1);  // one 32-byte sector
    acc += v.x + v.y + w.x + w.y;
    h = mix64(h + acc + 1);
  }
  if (acc == 0x1234567) out[0] = acc;
}

// the probing loop of the framework's key tables (csrc/cuda/sgd.cu k_kv_lookup_insert): read the home
// slot, claim it with a system-scope CAS when empty, walk on otherwise
__global__ void k_lookup_insert(unsigned long long* table, uint64_t cap, int n, uint64_t seed, long long* slot_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long key = mix64(seed + i) >> 2;  // never the empty marker
  uint64_t h = mix64(key) & (cap - 1);
  long long found = -1;
  for (uint64_t probe = 0; probe < cap; ++probe) {
    unsigned long long k = *reinterpret_cast<volatile unsigned long long*>(table + h);
    if (k == ~0ull) k = atomicCAS_system(table + h, ~0ull, key);
    if (k == ~0ull || k == key) { found = (long long)h; break; }
    h = (h + 1) & (cap - 1);
  }
  slot_out[i] = found;
}

static float time_probe(const void* table, uint64_t bytes, int n, int chain, bool wide, unsigned long long* out) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  float best = 1e30f;
  for (int rep = 0; rep < 6; ++rep) {
    CK(cudaEventRecord(a));
    if (wide) k_probe32<<<(n + 255) / 256, 256>>>((const ulonglong4*)table, bytes / 32, n, chain, 1000 + rep, out);
    else k_probe<<<(n + 255) / 256, 256>>>((const unsigned long long*)table, bytes / 8, n, chain, 1000 + rep, out);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    if (rep > 0 && ms < best) best = ms;
  }
  return best * 1e3f;  // microseconds
}

int main() {
  const uint64_t kMax = 8ull << 30;
  int to_parent[2], to_child[2];
  if (pipe(to_parent) || pipe(to_child)) return 1;
  // fork BEFORE the first CUDA call: a child of a process that has initialised CUDA cannot use it
  pid_t pid = fork();
  if (pid == 0) {  // child: owns the table on GPU 1
    close(to_parent[0]); close(to_child[1]);
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) { fprintf(stderr, "needs 2 GPUs\n"); return 1; }
    CK(cudaSetDevice(1));
    // PEER_ODD=1: an allocation that is not a multiple of the 2 MB page (as 2 GB + 4 KB would be)
    const uint64_t extra = getenv("PEER_ODD") ? 4096 : 0;
    void* p; CK(cudaMalloc(&p, kMax + extra)); CK(cudaMemset(p, 0, kMax)); CK(cudaDeviceSynchronize());
    cudaIpcMemHandle_t h; CK(cudaIpcGetMemHandle(&h, p));
    if (write(to_parent[1], &h, sizeof h) != sizeof h) return 1;
    char c; if (read(to_child[0], &c, 1) != 1) return 1;
    cudaFree(p);
    return 0;
  }
  close(to_parent[1]); close(to_child[0]);  // a child that died gives EOF here, not a hang
  cudaIpcMemHandle_t h;
  if (read(to_parent[0], &h, sizeof h) != sizeof h) { fprintf(stderr, "the owner process failed\n"); return 1; }
  CK(cudaSetDevice(0));
  void* peer; CK(cudaIpcOpenMemHandle(&peer, h, cudaIpcMemLazyEnablePeerAccess));
  void* local; CK(cudaMalloc(&local, kMax)); CK(cudaMemset(local, 0, kMax));
  unsigned long long* out; CK(cudaMalloc(&out, 8));
  printf("%-6s %-9s %8s %6s %5s %10s %12s\n", "where", "table", "threads", "chain", "bytes", "us", "ns/round");
  for (int where = 0; where < 2; ++where)
    for (uint64_t bytes : {64ull << 20, 512ull << 20, 2ull << 30, 8ull << 30})
      for (int n : {25000, 200000})
        for (int chain : {1, 8})
          for (int wide = 0; wide < 2; ++wide) {
            const float us = time_probe(where ? peer : local, bytes, n, chain, wide, out);
            printf("%-6s %6llu MB %8d %6d %5d %10.1f %12.1f\n", where ? "peer" : "local", (unsigned long long)(bytes >> 20),
                   n, chain, wide ? 32 : 8, us, us * 1e3 / chain);
          }
  // the table code: every repetition printed (nothing discarded), new keys, then the same keys again
  long long* slots; CK(cudaMalloc(&slots, 200000 * sizeof(long long)));
  printf("\nlookup/insert of 200000 keys, table of 2048 MB (us per call, every call shown)\n");
  for (int where = 0; where < 2; ++where) {
    unsigned long long* t = (unsigned long long*)(where ? peer : local);
    const uint64_t cap = (2ull << 30) / 8;
    CK(cudaMemset(local, 0xff, 2ull << 30));
    if (where) CK(cudaMemcpy(peer, local, 2ull << 30, cudaMemcpyDefault));  // an empty table on the peer
    CK(cudaDeviceSynchronize());
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int round = 0; round < 3; ++round)        // round r: keys of seed r, first new, then present (twice)
      for (int again = 0; again < 3; ++again) {
        CK(cudaEventRecord(a));
        k_lookup_insert<<<(200000 + 255) / 256, 256>>>(t, cap, 200000, 777 + round * 1000003ull, slots);
        CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b));
        printf("%-6s round %d %-8s %10.1f\n", where ? "peer" : "local", round, again ? "present" : "new", ms * 1e3f);
      }
  }
  // does an idle gap before the call matter (clocks, link power state)?  present keys of round 0
  printf("\npeer lookup of 200000 present keys after an idle gap (us per call)\n");
  {
    unsigned long long* t = (unsigned long long*)peer;
    const uint64_t cap = (2ull << 30) / 8;
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int gap_ms : {0, 1, 10, 100, 1000, 0, 0})
      for (int rep = 0; rep < 2; ++rep) {
        CK(cudaDeviceSynchronize());
        usleep(gap_ms * 1000);
        CK(cudaEventRecord(a));
        k_lookup_insert<<<(200000 + 255) / 256, 256>>>(t, cap, 200000, 777, slots);
        CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b));
        printf("gap %5d ms  %10.1f\n", gap_ms, ms * 1e3f);
      }
  }
  CK(cudaIpcCloseMemHandle(peer));
  char c = 1; if (write(to_child[1], &c, 1) != 1) return 1;
  waitpid(pid, nullptr, 0);
  return 0;
}
