// Micro-benchmark (tuning aid, not product): dependent-load latency of 32-byte random slots, the floor
// of one hop of the walk kernel.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/chase.cu -o /tmp/chase
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
struct __align__(32) Slot { unsigned next, a, b, c, d, e, f, g; };
template <int MODE>
__global__ void chase(const Slot* t, const unsigned* start, int hops, unsigned* out, int active) {
  int lane = threadIdx.x;
  if (lane >= active) return;
  unsigned i = start[blockIdx.x * 32 + lane];
  unsigned acc = 0;
  for (int h = 0; h < hops; ++h) {
    unsigned n, x;
    if (MODE == 0) { int4 v = __ldg(reinterpret_cast<const int4*>(t + i)); n = v.x; x = v.y; }
    else if (MODE == 1) {
      unsigned r[8];
      asm volatile("ld.global.nc.L2::evict_last.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(t + i));
      n = r[0]; x = r[7];
    } else { int4 v = __ldcg(reinterpret_cast<const int4*>(t + i)); n = v.x; x = v.y; }
    acc += x;
    i = n;
  }
  out[blockIdx.x * 32 + lane] = acc + i;
}
int main() {
  const size_t N = 1u << 22;   // 128 MB table
  std::vector<Slot> h(N);
  std::vector<unsigned> perm(N);
  for (size_t i = 0; i < N; ++i) perm[i] = i;
  srand(1);
  for (size_t i = N - 1; i > 0; --i) { size_t j = ((size_t)rand() * 32768 + rand()) % (i + 1); std::swap(perm[i], perm[j]); }
  for (size_t i = 0; i < N; ++i) h[perm[i]].next = perm[(i + 1) % N];
  // small hot ring of 64K slots (2 MB) for L2/L1-resident chase
  Slot* d; cudaMalloc(&d, N * sizeof(Slot)); cudaMemcpy(d, h.data(), N * sizeof(Slot), cudaMemcpyHostToDevice);
  const size_t M = 1u << 16;
  std::vector<Slot> hs(N);
  for (size_t i = 0; i < M; ++i) hs[perm[i]].next = perm[(i + 1) % M];
  Slot* ds; cudaMalloc(&ds, N * sizeof(Slot)); cudaMemcpy(ds, hs.data(), N * sizeof(Slot), cudaMemcpyHostToDevice);
  unsigned* st; cudaMalloc(&st, 4096 * 32 * 4);
  std::vector<unsigned> s(4096 * 32);
  for (auto& v : s) v = perm[((size_t)rand() * 32768 + rand()) % M];
  cudaMemcpy(st, s.data(), s.size() * 4, cudaMemcpyHostToDevice);
  unsigned* out; cudaMalloc(&out, 4096 * 32 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int hops = 2000;
  auto run = [&](const char* name, void (*k)(const Slot*, const unsigned*, int, unsigned*, int), const Slot* tab, int blocks, int active) {
    k<<<blocks, 32>>>(tab, st, hops, out, active); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<<<blocks, 32>>>(tab, st, hops, out, active); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s blocks %5d lanes %2d : %7.1f ns/hop\n", name, blocks, active, ms * 1e6 / hops);
  };
  run("ldg128 L2-hot(2MB ring)", chase<0>, ds, 1, 1);
  run("ldg128 L2-hot(2MB ring)", chase<0>, ds, 1, 32);
  run("ldg256.evict_last hot", chase<1>, ds, 1, 1);
  run("ldg256.evict_last hot", chase<1>, ds, 1, 32);
  run("ldcg128 hot", chase<2>, ds, 1, 32);
  run("ldg256 hot 2048 warps", chase<1>, ds, 2048, 32);
  run("ldg128 hot 2048 warps", chase<0>, ds, 2048, 32);
  run("ldg128 cold(128MB)", chase<0>, d, 1, 32);
  run("ldg128 cold 2048 warps", chase<0>, d, 2048, 32);
  return 0;
}
