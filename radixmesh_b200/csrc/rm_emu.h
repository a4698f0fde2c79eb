// Host emulation of the small CUDA subset the tree kernels use (TEST HARNESS ONLY).
//
// Built only by tests/emu/build_emu.py into tests/emu/_rm_emu.so, with -DRM_EMU, so that the
// sequential-semantics logic of the kernels (walks, splits, in-order insert apply, lock and
// evict bookkeeping) can be checked against the oracle in the GPU-less build container.
// Every "group" of lanes is emulated with ONE lane (RM_G == 1) and blocks/threads run one
// after another, so nothing here exercises real parallelism.  The product library is never
// built with RM_EMU and never loads this; bench.py and the package cannot reach it.
#pragma once
#ifndef RM_EMU
#error "rm_emu.h is only for the host emulation build"
#endif
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <stdio.h>
using std::max;
using std::min;

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static   /* threads run one after another: a static buffer indexed by threadIdx works */
#define __align__(n) __attribute__((aligned(n)))

struct rm_emu_dim3 { unsigned x, y, z; };
extern thread_local rm_emu_dim3 threadIdx, blockIdx, blockDim, gridDim;

typedef int cudaError_t;
typedef void* cudaStream_t;
#define cudaSuccess 0
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
static inline cudaError_t cudaMalloc(void** p, size_t n) { *p = malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaFree(void* p) { free(p); return 0; }
static inline cudaError_t cudaMallocHost(void** p, size_t n) { *p = malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaFreeHost(void* p) { free(p); return 0; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t) { memset(d, v, n); return 0; }
static inline cudaError_t cudaMemset(void* d, int v, size_t n) { memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaSetDevice(int) { return 0; }
static inline cudaError_t cudaGetDevice(int* d) { *d = 0; return 0; }
static inline cudaError_t cudaGetLastError() { return 0; }
static inline const char* cudaGetErrorString(cudaError_t) { return "emu"; }
template <class T> static inline cudaError_t cudaGetSymbolAddress(void** p, T& sym) { *p = (void*)&sym; return 0; }

template <class T> static inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
template <class T> static inline T atomicSub(T* p, T v) { T o = *p; *p = o - v; return o; }
template <class T> static inline T atomicMax(T* p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <class T> static inline T atomicMin(T* p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <class T> static inline T atomicExch(T* p, T v) { T o = *p; *p = v; return o; }
template <class T> static inline T atomicCAS(T* p, T c, T v) { T o = *p; if (o == c) *p = v; return o; }
template <class T> static inline T atomicOr(T* p, T v) { T o = *p; *p = o | v; return o; }

static inline unsigned __ballot_sync(unsigned, int pred) { return pred ? 1u : 0u; }
template <class T> static inline T __shfl_sync(unsigned, T v, int, int = 32) { return v; }
template <class T> static inline T __shfl_down_sync(unsigned, T v, int, int = 32) { return v; }
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int, int = 32) { return v; }
static inline int __ffs(unsigned v) { return v ? __builtin_ffs((int)v) : 0; }
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline void __syncwarp(unsigned = 0xffffffffu) {}
static inline void __threadfence() {}
static inline void __syncthreads() {}
template <class T> static inline T __ldg(const T* p) { return *p; }
template <class T> static inline T __ldcg(const T* p) { return *p; }
template <class T> static inline void __stcg(T* p, T v) { *p = v; }
struct int4 { int x, y, z, w; };
struct int2 { int x, y; };
static inline int2 make_int2(int x, int y) { return int2{x, y}; }
struct uint2 { unsigned x, y; };
struct uint4 { unsigned x, y, z, w; };

#define RM_LAUNCH_PDL RM_LAUNCH
static inline void rm_grid_dependency_wait() {}
static inline void rm_grid_dependency_trigger() {}

// Sequential launcher: one emulated thread per GROUP (RM_G == 1), block by block.
#define RM_LAUNCH(kernel, grid, block, stream, ...)                                  \
  do {                                                                               \
    gridDim = rm_emu_dim3{(unsigned)(grid), 1, 1};                                   \
    blockDim = rm_emu_dim3{(unsigned)(block), 1, 1};                                 \
    for (unsigned _b = 0; _b < (unsigned)(grid); ++_b)                               \
      for (unsigned _t = 0; _t < (unsigned)(block); ++_t) {                          \
        blockIdx = rm_emu_dim3{_b, 0, 0};                                            \
        threadIdx = rm_emu_dim3{_t, 0, 0};                                           \
        kernel(__VA_ARGS__);                                                         \
      }                                                                              \
  } while (0)
