// rm_match_batch_static: the batched prefix match with its scratch in a fixed, file-scope device
// array instead of memory the index allocated.
//
// Who needs it.  An index is normally created once, with scratch sized for rm_config.max_batch, and
// every later call reuses that scratch.  Some hosts of this library cannot size that up front and
// cannot call cudaMalloc afterwards either: a serving engine hands almost all of HBM to its KV pool
// at start-up, so a later device allocation can simply fail; and a scheduler or an oplog
// receive thread on the request path cannot afford cudaMalloc's implicit device-wide
// synchronisation.  Such a caller creates its index with a
// minimal max_batch and sends its match batches (up to RM_STATIC_MAX_BATCH queries) through this
// entry point: no cudaMalloc, cudaFree or other allocating call is made, here or below.
//
// The price.  The array is part of the library image: every GPU that loads the library holds
// RM_STATIC_SCRATCH_BYTES of it until the process exits, used or not.  All calls, on all indexes
// and streams of a device, share it, so two calls that overlap on the GPU race; issue them on one
// stream or order them yourself.  It keeps the previous call's bytes and the entry point clears nothing:
// each kernel writes every word it later reads (K1 fills the request arrays and chain links of every
// request it counts; the split stage reads only those, and its rare sorted path writes its lists before
// reading them), so nothing in the array has to start at zero.
//
// First call on a device: looks up the array's device address (cudaGetSymbolAddress) and caches it
// per device.  In host code the array's name is only a handle for that lookup, never an address a
// kernel may be given.  Make this first call before any stream capture begins.
//
// Like every entry point, calls on one index must be serialised by the caller: for the duration of the
// call the index's scratch view points into the static array.
#include <mutex>

#include "rm_internal.h"
#include "rm_sort.cuh"

#define RM_STATIC_MAX_BATCH 4096
#define RM_STATIC_SCRATCH_BYTES (256 * 1024)
#define RM_STATIC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_rm_static_scratch[RM_STATIC_SCRATCH_BYTES];

static std::mutex g_static_mu;
static unsigned char* g_static_ptr[RM_STATIC_MAX_DEVICES];

// Device address of the scratch array on `device` (which must be current).
static int rm_static_base(int device, unsigned char** out) {
  if (device < 0 || device >= RM_STATIC_MAX_DEVICES) {
    rm_set_error("rm_match_batch_static: device ordinal %d out of range", device);
    return RM_E_INVALID;
  }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_rm_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      rm_set_error("rm_match_batch_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return RM_E_CUDA;
    }
    g_static_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[device];
  return RM_OK;
}

template <class T>
static T* rm_carve(unsigned char*& cur, size_t count) {
  T* p = reinterpret_cast<T*>(cur);
  size_t bytes = (sizeof(T) * count + 255) & ~(size_t)255;
  cur += bytes;
  return p;
}

extern "C" int rm_match_batch_static(rm_index* idx, const int32_t* tokens, const int64_t* offsets, int32_t B,
                                     int32_t* out_len, int32_t* out_node, void* stream) {
  if (!idx || B < 0) {
    rm_set_error("rm_match_batch_static: bad arguments");
    return RM_E_INVALID;
  }
  if (B > RM_STATIC_MAX_BATCH) {
    rm_set_error("rm_match_batch_static: batch of %d queries exceeds the fixed scratch (at most %d per call; "
                 "split the batch or use rm_match_batch)", B, RM_STATIC_MAX_BATCH);
    return RM_E_TOO_LARGE;
  }
  if (idx->cfg.mode == RM_MODE_ROUTER) {
    rm_set_error("rm_match_batch_static: a ROUTER index answers through rm_route_batch");
    return RM_E_MODE;
  }
  if (B == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  unsigned char* base = nullptr;
  int rc = rm_static_base(idx->cfg.device, &base);
  if (rc) return rc;
  // lay the match pipeline's scratch out inside the static array
  unsigned char* cur = base;
  RmScratch view = idx->sc;
  view.key[0] = rm_carve<u64>(cur, RM_STATIC_MAX_BATCH);
  view.key[1] = rm_carve<u64>(cur, RM_STATIC_MAX_BATCH);
  view.val[0] = rm_carve<unsigned>(cur, RM_STATIC_MAX_BATCH);
  view.val[1] = rm_carve<unsigned>(cur, RM_STATIC_MAX_BATCH);
  view.sufmax = rm_carve<u64>(cur, RM_STATIC_MAX_BATCH);
  view.krank = rm_carve<int>(cur, RM_STATIC_MAX_BATCH);
  view.uniq_pos = rm_carve<int>(cur, RM_STATIC_MAX_BATCH);
  view.uniq_min = rm_carve<unsigned>(cur, RM_STATIC_MAX_BATCH);
  view.link = rm_carve<int>(cur, RM_STATIC_MAX_BATCH);
  if ((size_t)(cur - base) > (size_t)RM_STATIC_SCRATCH_BYTES) {
    rm_set_error("rm_match_batch_static: internal layout exceeds the static array");
    return RM_E_INVALID;
  }
  // run the unchanged pipeline on that view, then hand the index its own scratch back
  const RmScratch saved = idx->sc;
  idx->sc = view;
  rc = rm_launch_match(idx, tokens, (const i64*)offsets, B, out_len, out_node, (cudaStream_t)stream);
  idx->sc = saved;
  return rc;
}
