// K8 router pick + consistent-hash fallback (sm_100a).
//
// Replaces RadixMesh.match_prefix in ROUTER mode (radix_mesh.py:219-238), CacheAwareRouter.
// cache_aware_route (router/cache_aware_router.py:23-39) and ConsistentHash.get_node/_hash
// (:97-118).  The walk is K1 (rm_match.cu) with the O(1)-per-hop rank state; this file turns the
// (prefill, decode) ranks into address indices and, for the -1 cases, hashes the key the way the
// reference does: MD5 over Python's str(list) rendering "[t0, t1, ...]", first four digest bytes
// big-endian, bisect-right on the precomputed ring of 3 virtual nodes per address.
#include "rm_internal.h"

#ifdef RM_EMU
#define RM_RT_TPB 64
#else
#define RM_RT_TPB 128
#endif

// match kernel launcher for ROUTER indexes (rm_match.cu)
int rm_launch_router_match(rm_index* idx, const int* tokens, const i64* offsets, int B, int* out_len, int* out_node,
                           int* out_prefill, int* out_decode, unsigned char* out_err, cudaStream_t s);

struct RmMd5 {
  unsigned a, b, c, d;
  unsigned w[16];   // current 64-byte block as little-endian words
  unsigned fill;    // bytes in the block
  u64 total;        // bytes hashed so far
};

__device__ __forceinline__ unsigned rm_rotl(unsigned x, int s) { return (x << s) | (x >> (32 - s)); }

__device__ void rm_md5_block(RmMd5& m) {
  // RFC 1321 round constants and shifts
  const unsigned K[64] = {
      0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
      0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
      0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
      0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
      0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
      0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
      0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
      0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
  const int S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9,  14, 20, 5, 9,
                     14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                     4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
  unsigned a = m.a, b = m.b, c = m.c, d = m.d;
#pragma unroll
  for (int i = 0; i < 64; ++i) {
    unsigned f;
    int g;
    if (i < 16) {
      f = (b & c) | (~b & d);
      g = i;
    } else if (i < 32) {
      f = (d & b) | (~d & c);
      g = (5 * i + 1) & 15;
    } else if (i < 48) {
      f = b ^ c ^ d;
      g = (3 * i + 5) & 15;
    } else {
      f = c ^ (b | ~d);
      g = (7 * i) & 15;
    }
    unsigned tmp = d;
    d = c;
    c = b;
    b = b + rm_rotl(a + f + K[i] + m.w[g], S[i]);
    a = tmp;
  }
  m.a += a;
  m.b += b;
  m.c += c;
  m.d += d;
}

__device__ __forceinline__ void rm_md5_init(RmMd5& m) {
  m.a = 0x67452301u;
  m.b = 0xefcdab89u;
  m.c = 0x98badcfeu;
  m.d = 0x10325476u;
  m.fill = 0;
  m.total = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) m.w[i] = 0;
}

__device__ __forceinline__ void rm_md5_put(RmMd5& m, unsigned char ch) {
  const unsigned wi = m.fill >> 2, sh = (m.fill & 3) * 8;
  // w[] is indexed with a runtime value only here; keep it simple (local memory is fine for a
  // fallback path that runs once per missed request)
  m.w[wi] |= (unsigned)ch << sh;
  ++m.fill;
  ++m.total;
  if (m.fill == 64) {
    rm_md5_block(m);
    m.fill = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) m.w[i] = 0;
  }
}

__device__ unsigned rm_md5_finish_u32be(RmMd5& m) {
  const u64 bits = m.total * 8ull;
  rm_md5_put(m, 0x80);
  while (m.fill != 56) rm_md5_put(m, 0);
  m.w[14] = (unsigned)(bits & 0xffffffffu);
  m.w[15] = (unsigned)(bits >> 32);
  rm_md5_block(m);
  // digest bytes 0..3 are the little-endian bytes of `a`; the reference reads them big-endian
  const unsigned a = m.a;
  return ((a & 0xffu) << 24) | ((a & 0xff00u) << 8) | ((a >> 8) & 0xff00u) | (a >> 24);
}

// str(int) for a Python int, appended to the hash
__device__ __forceinline__ void rm_md5_put_int(RmMd5& m, int v) {
  char buf[12];
  int n = 0;
  long long x = v;
  if (x < 0) {
    rm_md5_put(m, '-');
    x = -x;
  }
  do {
    buf[n++] = (char)('0' + (int)(x % 10));
    x /= 10;
  } while (x);
  while (n) rm_md5_put(m, (unsigned char)buf[--n]);
}

// ConsistentHash._hash(key) for key = list of ints (cache_aware_router.py:110-118)
__device__ unsigned rm_hash_token_list(const int* key, int n) {
  RmMd5 m;
  rm_md5_init(m);
  rm_md5_put(m, '[');
  for (int i = 0; i < n; ++i) {
    if (i) {
      rm_md5_put(m, ',');
      rm_md5_put(m, ' ');
    }
    rm_md5_put_int(m, key[i]);
  }
  rm_md5_put(m, ']');
  return rm_md5_finish_u32be(m);
}

// bisect.bisect(ring, h) % len(ring) (cache_aware_router.py:106-108)
__device__ __forceinline__ int rm_ring_owner(const unsigned* ring, const int* owner, int n, unsigned h) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (h < ring[mid]) hi = mid;
    else lo = mid + 1;
  }
  return owner[lo % n];
}

__global__ void __launch_bounds__(RM_RT_TPB)
rm_route_finish_kernel(RmTree t, const int* __restrict__ qtok, const i64* __restrict__ qoff, int B, int warm,
                       int* __restrict__ prefill, int* __restrict__ decode, unsigned char* __restrict__ err,
                       int* __restrict__ paddr, int* __restrict__ daddr, const unsigned* ring_p, const int* ring_p_owner,
                       int ring_p_n, const unsigned* ring_d, const int* ring_d_owner, int ring_d_n) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= B) return;
  int pf = -1, dc = -1;
  unsigned char e = 0;
  if (warm) {
    pf = prefill[q];
    dc = decode[q];
    e = err[q];
  } else {
    prefill[q] = -1;
    decode[q] = -1;
  }
  int pa = -1, da = -1;
  if (e == 0) {
    const bool need_hash = pf == -1 || dc == -1;
    unsigned h = 0;
    if (need_hash) h = rm_hash_token_list(qtok + qoff[q], (int)(qoff[q + 1] - qoff[q]));
    if (pf != -1) {
      // prefill_cache_nodes[rank]: a decode rank here is the reference's IndexError (radix_mesh.py:451)
      if (pf >= t.n_prefill) e = 2;
      else pa = pf;
    } else if (ring_p_n > 0) {
      pa = rm_ring_owner(ring_p, ring_p_owner, ring_p_n, h);
    }
    if (e == 0) {
      if (dc != -1) {
        // decode_cache_nodes[local_node_rank(rank)] (radix_mesh.py:453-457, cache_config.py:30-35)
        da = dc - t.n_prefill;
        if (da < 0 || da >= t.n_decode) e = 2;
      } else if (ring_d_n > 0) {
        da = rm_ring_owner(ring_d, ring_d_owner, ring_d_n, h);
      }
    }
  }
  paddr[q] = pa;
  daddr[q] = da;
  err[q] = e;
}

int rm_launch_route(rm_index* idx, const int* tokens, const i64* offsets, int B, int warm, int* out_prefill,
                    int* out_decode, int* out_paddr, int* out_daddr, unsigned char* out_err, cudaStream_t s) {
  if (B <= 0) return RM_OK;
  if (warm) {
    int rc = rm_launch_router_match(idx, tokens, offsets, B, idx->sc.h_out_a, idx->sc.h_out_b, out_prefill, out_decode,
                                    out_err, s);
    if (rc) return rc;
  } else {
    RM_CUDA(cudaMemsetAsync(out_err, 0, (size_t)B, s));
  }
  RM_LAUNCH(rm_route_finish_kernel, (B + RM_RT_TPB - 1) / RM_RT_TPB, RM_RT_TPB, s, idx->t, tokens, offsets, B, warm,
            out_prefill, out_decode, out_err, out_paddr, out_daddr, idx->ring_p, idx->ring_p_owner, idx->ring_p_n,
            idx->ring_d, idx->ring_d_owner, idx->ring_d_n);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

extern "C" int rm_route_batch(rm_index* idx, const int32_t* tokens, const int64_t* offsets, int32_t B, int32_t warm,
                              int32_t* out_prefill, int32_t* out_decode, int32_t* out_paddr, int32_t* out_daddr,
                              uint8_t* out_err, void* stream) {
  if (!idx || B < 0) return RM_E_INVALID;
  if (B > idx->cfg.max_batch) {
    rm_set_error("rm_route_batch: batch of %d exceeds rm_config.max_batch = %d", B, idx->cfg.max_batch);
    return RM_E_TOO_LARGE;
  }
  if (idx->cfg.mode != RM_MODE_ROUTER) {
    rm_set_error("rm_route_batch: index is not in ROUTER mode");
    return RM_E_MODE;
  }
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_route(idx, tokens, (const i64*)offsets, B, warm, out_prefill, out_decode, out_paddr, out_daddr,
                         out_err, (cudaStream_t)stream);
}

// K9 pick: after the all-gather of per-GPU match depths (GPU g's tree holds the keys rank g inserted),
// choose the prefill / decode rank the way the reference's router tree would:
// deepest owner among prefill ranks; a decode rank only if it is strictly deeper; lowest rank on ties.
__global__ void __launch_bounds__(RM_RT_TPB)
rm_pick_kernel(const int* __restrict__ depths, int n_ranks, int B, int n_prefill, int n_decode, int* __restrict__ prefill,
               int* __restrict__ decode, unsigned char* __restrict__ err) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= B) return;
  int bp = 0, bd = 0, pr = -1, dr = -1;
  for (int g = 0; g < n_ranks; ++g) {
    const int d = depths[(size_t)g * B + q];
    if (g < n_prefill) {
      if (d > bp) {
        bp = d;
        pr = g;
      }
    } else if (g < n_prefill + n_decode) {
      if (d > bd) {
        bd = d;
        dr = g;
      }
    }
  }
  prefill[q] = bp > 0 ? pr : -1;
  decode[q] = (bd > 0 && bd > bp) ? dr : -1;
  err[q] = 0;
}

extern "C" int rm_pick_batch(rm_index* idx, const int32_t* depths, int32_t n_ranks, const int32_t* tokens,
                             const int64_t* offsets, int32_t B, int32_t* out_prefill, int32_t* out_decode,
                             int32_t* out_paddr, int32_t* out_daddr, uint8_t* out_err, void* stream) {
  if (!idx || B < 0 || n_ranks < 1) return RM_E_INVALID;
  if (B == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RM_LAUNCH(rm_pick_kernel, (B + RM_RT_TPB - 1) / RM_RT_TPB, RM_RT_TPB, s, depths, n_ranks, B, idx->cfg.n_prefill,
            idx->cfg.n_decode, out_prefill, out_decode, out_err);
  RM_LAUNCH(rm_route_finish_kernel, (B + RM_RT_TPB - 1) / RM_RT_TPB, RM_RT_TPB, s, idx->t, tokens, (const i64*)offsets, B, 1,
            out_prefill, out_decode, out_err, out_paddr, out_daddr, idx->ring_p, idx->ring_p_owner, idx->ring_p_n,
            idx->ring_d, idx->ring_d_owner, idx->ring_d_n);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

static int rm_set_ring(unsigned** dpts, int** down, int* dn, const uint32_t* pts_h, const int32_t* own_h, int n) {
  if (*dpts) cudaFree(*dpts);
  if (*down) cudaFree(*down);
  *dpts = nullptr;
  *down = nullptr;
  *dn = 0;
  if (n <= 0) return RM_OK;
  RM_CUDA(cudaMalloc(reinterpret_cast<void**>(dpts), sizeof(unsigned) * n));
  RM_CUDA(cudaMalloc(reinterpret_cast<void**>(down), sizeof(int) * n));
  RM_CUDA(cudaMemcpy(*dpts, pts_h, sizeof(unsigned) * n, cudaMemcpyHostToDevice));
  RM_CUDA(cudaMemcpy(*down, own_h, sizeof(int) * n, cudaMemcpyHostToDevice));
  *dn = n;
  return RM_OK;
}

extern "C" int rm_route_set_rings(rm_index* idx, const uint32_t* p_points_h, const int32_t* p_owner_h, int32_t n_p,
                                  const uint32_t* d_points_h, const int32_t* d_owner_h, int32_t n_d) {
  if (!idx) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  int rc = rm_set_ring(&idx->ring_p, &idx->ring_p_owner, &idx->ring_p_n, p_points_h, p_owner_h, n_p);
  if (rc) return rc;
  return rm_set_ring(&idx->ring_d, &idx->ring_d_owner, &idx->ring_d_n, d_points_h, d_owner_h, n_d);
}
