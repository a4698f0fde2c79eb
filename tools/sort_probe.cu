// Times rm_sort_pairs (radixmesh_b200/csrc/rm_sort.cuh) alone for a range of sizes, with the item count read on
// the device as in a match batch, and checks the result against std::stable_sort.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Iradixmesh_b200/csrc -o build/sort_probe tools/sort_probe.cu
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <vector>

#include "rm_sort.cuh"

unsigned long long g_rm_launches = 0;

__global__ void set_n(int* p, int n) { *p = n; }

int main(int argc, char** argv) {
  const int n_max = argc > 1 ? atoi(argv[1]) : 65536;
  const int bits = argc > 2 ? atoi(argv[2]) : 30;
  u64 *k0, *k1;
  unsigned *v0, *v1, *hist;
  int* n_dev;
  cudaMalloc(&k0, 8ull * n_max);
  cudaMalloc(&k1, 8ull * n_max);
  cudaMalloc(&v0, 4ull * n_max);
  cudaMalloc(&v1, 4ull * n_max);
  size_t words = rm_sort_scratch_words(n_max);
  cudaMalloc(&hist, 4 * words);
  cudaMemset(hist, 0, 4 * words);
  cudaMalloc(&n_dev, 4);
  std::vector<u64> hk(n_max);
  std::vector<unsigned> hv(n_max);
  srand(1);
  for (int i = 0; i < n_max; ++i) {
    hk[i] = (((u64)rand() << 16) ^ (u64)rand()) & ((1ull << bits) - 1);
    if (i % 3 == 0 && i) hk[i] = hk[i - 1];   // duplicates: stability matters
    hv[i] = i;
  }
  cudaStream_t s;
  cudaStreamCreate(&s);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int sizes[] = {0, 1, 100, 500, 1000, 2000, 4096, 4097, 8000, 16384, 40000, 65536, 200000, 600000, 1000000};
  for (int n : sizes) {
    if (n > n_max) break;
    const int reps = 50;
    float tot = 0;
    int w = 0;
    for (int r = 0; r < reps + 3; ++r) {
      cudaMemcpyAsync(k0, hk.data(), 8ull * n_max, cudaMemcpyHostToDevice, s);
      cudaMemcpyAsync(v0, hv.data(), 4ull * n_max, cudaMemcpyHostToDevice, s);
      set_n<<<1, 1, 0, s>>>(n_dev, n);
      cudaEventRecord(e0, s);
      w = rm_sort_pairs(k0, v0, k1, v1, n_max, bits, hist, s, n_dev);
      cudaEventRecord(e1, s);
      cudaStreamSynchronize(s);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (r >= 3) tot += ms;
    }
    std::vector<u64> rk(n);
    std::vector<unsigned> rv(n);
    cudaMemcpy(rk.data(), w ? k1 : k0, 8ull * n, cudaMemcpyDeviceToHost);
    cudaMemcpy(rv.data(), w ? v1 : v0, 4ull * n, cudaMemcpyDeviceToHost);
    std::vector<unsigned> idx(n);
    std::iota(idx.begin(), idx.end(), 0u);
    std::stable_sort(idx.begin(), idx.end(), [&](unsigned a, unsigned b) { return hk[a] < hk[b]; });
    int bad = 0;
    for (int i = 0; i < n; ++i) bad += (rv[i] != idx[i]) || (rk[i] != hk[idx[i]]);
    cudaError_t err = cudaGetLastError();
    printf("n %8d of %8d  bits %d  %8.2f us  %s %s\n", n, n_max, bits, 1e3 * tot / reps, bad ? "MISMATCH" : "ok",
           err ? cudaGetErrorString(err) : "");
  }
  return 0;
}
