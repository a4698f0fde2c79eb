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

// MD5 (RFC 1321) streamed per thread.  The message words of the current 64-byte block live in shared
// memory, one column per thread (wsm[word][thread]: conflict free), because they are written at a
// run-time index while the text is rendered and read at compile-time indices by the unrolled rounds; the
// first version kept them in a per-thread array (local memory, one read-modify-write per BYTE) and took
// 4.7 ms for 16K 2K-token requests.  Bytes are gathered four at a time in a 64-bit shift register.
struct RmMd5 {
  unsigned a, b, c, d;
};

__device__ __forceinline__ unsigned rm_rotl(unsigned x, int s) { return (x << s) | (x >> (32 - s)); }

// message word + round constant.  An MD5 step is a chain through the word computed by the step before; written
// as one sum the compiler emits LOP3 -> IADD3(f, a, x) -> IADD(k) -> LEA.HI(rotate, + b), four dependent
// instructions.  x + k does not depend on the chain: kept apart (the asm hides it from reassociation) the chain
// is LOP3 -> IADD3(f, a, x + k) -> LEA.HI.
__device__ __forceinline__ unsigned rm_md5_xk(unsigned x, unsigned k) {
#ifdef RM_EMU
  return x + k;
#else
  unsigned r;
  asm("add.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(k));
  return r;
#endif
}

#define RM_MD5_STEP(f, a, b, c, d, x, k, s) \
  a = b + rm_rotl(a + (f) + rm_md5_xk(x, k), s)

template <class W>
__device__ __forceinline__ void rm_md5_block(RmMd5& m, const W& w) {
  unsigned a = m.a, b = m.b, c = m.c, d = m.d;
#define F1(x, y, z) ((x & y) | (~x & z))
#define F2(x, y, z) ((z & x) | (~z & y))
#define F3(x, y, z) (x ^ y ^ z)
#define F4(x, y, z) (y ^ (x | ~z))
  RM_MD5_STEP(F1(b, c, d), a, b, c, d, w(0), 0xd76aa478u, 7);
  RM_MD5_STEP(F1(a, b, c), d, a, b, c, w(1), 0xe8c7b756u, 12);
  RM_MD5_STEP(F1(d, a, b), c, d, a, b, w(2), 0x242070dbu, 17);
  RM_MD5_STEP(F1(c, d, a), b, c, d, a, w(3), 0xc1bdceeeu, 22);
  RM_MD5_STEP(F1(b, c, d), a, b, c, d, w(4), 0xf57c0fafu, 7);
  RM_MD5_STEP(F1(a, b, c), d, a, b, c, w(5), 0x4787c62au, 12);
  RM_MD5_STEP(F1(d, a, b), c, d, a, b, w(6), 0xa8304613u, 17);
  RM_MD5_STEP(F1(c, d, a), b, c, d, a, w(7), 0xfd469501u, 22);
  RM_MD5_STEP(F1(b, c, d), a, b, c, d, w(8), 0x698098d8u, 7);
  RM_MD5_STEP(F1(a, b, c), d, a, b, c, w(9), 0x8b44f7afu, 12);
  RM_MD5_STEP(F1(d, a, b), c, d, a, b, w(10), 0xffff5bb1u, 17);
  RM_MD5_STEP(F1(c, d, a), b, c, d, a, w(11), 0x895cd7beu, 22);
  RM_MD5_STEP(F1(b, c, d), a, b, c, d, w(12), 0x6b901122u, 7);
  RM_MD5_STEP(F1(a, b, c), d, a, b, c, w(13), 0xfd987193u, 12);
  RM_MD5_STEP(F1(d, a, b), c, d, a, b, w(14), 0xa679438eu, 17);
  RM_MD5_STEP(F1(c, d, a), b, c, d, a, w(15), 0x49b40821u, 22);
  RM_MD5_STEP(F2(b, c, d), a, b, c, d, w(1), 0xf61e2562u, 5);
  RM_MD5_STEP(F2(a, b, c), d, a, b, c, w(6), 0xc040b340u, 9);
  RM_MD5_STEP(F2(d, a, b), c, d, a, b, w(11), 0x265e5a51u, 14);
  RM_MD5_STEP(F2(c, d, a), b, c, d, a, w(0), 0xe9b6c7aau, 20);
  RM_MD5_STEP(F2(b, c, d), a, b, c, d, w(5), 0xd62f105du, 5);
  RM_MD5_STEP(F2(a, b, c), d, a, b, c, w(10), 0x02441453u, 9);
  RM_MD5_STEP(F2(d, a, b), c, d, a, b, w(15), 0xd8a1e681u, 14);
  RM_MD5_STEP(F2(c, d, a), b, c, d, a, w(4), 0xe7d3fbc8u, 20);
  RM_MD5_STEP(F2(b, c, d), a, b, c, d, w(9), 0x21e1cde6u, 5);
  RM_MD5_STEP(F2(a, b, c), d, a, b, c, w(14), 0xc33707d6u, 9);
  RM_MD5_STEP(F2(d, a, b), c, d, a, b, w(3), 0xf4d50d87u, 14);
  RM_MD5_STEP(F2(c, d, a), b, c, d, a, w(8), 0x455a14edu, 20);
  RM_MD5_STEP(F2(b, c, d), a, b, c, d, w(13), 0xa9e3e905u, 5);
  RM_MD5_STEP(F2(a, b, c), d, a, b, c, w(2), 0xfcefa3f8u, 9);
  RM_MD5_STEP(F2(d, a, b), c, d, a, b, w(7), 0x676f02d9u, 14);
  RM_MD5_STEP(F2(c, d, a), b, c, d, a, w(12), 0x8d2a4c8au, 20);
  RM_MD5_STEP(F3(b, c, d), a, b, c, d, w(5), 0xfffa3942u, 4);
  RM_MD5_STEP(F3(a, b, c), d, a, b, c, w(8), 0x8771f681u, 11);
  RM_MD5_STEP(F3(d, a, b), c, d, a, b, w(11), 0x6d9d6122u, 16);
  RM_MD5_STEP(F3(c, d, a), b, c, d, a, w(14), 0xfde5380cu, 23);
  RM_MD5_STEP(F3(b, c, d), a, b, c, d, w(1), 0xa4beea44u, 4);
  RM_MD5_STEP(F3(a, b, c), d, a, b, c, w(4), 0x4bdecfa9u, 11);
  RM_MD5_STEP(F3(d, a, b), c, d, a, b, w(7), 0xf6bb4b60u, 16);
  RM_MD5_STEP(F3(c, d, a), b, c, d, a, w(10), 0xbebfbc70u, 23);
  RM_MD5_STEP(F3(b, c, d), a, b, c, d, w(13), 0x289b7ec6u, 4);
  RM_MD5_STEP(F3(a, b, c), d, a, b, c, w(0), 0xeaa127fau, 11);
  RM_MD5_STEP(F3(d, a, b), c, d, a, b, w(3), 0xd4ef3085u, 16);
  RM_MD5_STEP(F3(c, d, a), b, c, d, a, w(6), 0x04881d05u, 23);
  RM_MD5_STEP(F3(b, c, d), a, b, c, d, w(9), 0xd9d4d039u, 4);
  RM_MD5_STEP(F3(a, b, c), d, a, b, c, w(12), 0xe6db99e5u, 11);
  RM_MD5_STEP(F3(d, a, b), c, d, a, b, w(15), 0x1fa27cf8u, 16);
  RM_MD5_STEP(F3(c, d, a), b, c, d, a, w(2), 0xc4ac5665u, 23);
  RM_MD5_STEP(F4(b, c, d), a, b, c, d, w(0), 0xf4292244u, 6);
  RM_MD5_STEP(F4(a, b, c), d, a, b, c, w(7), 0x432aff97u, 10);
  RM_MD5_STEP(F4(d, a, b), c, d, a, b, w(14), 0xab9423a7u, 15);
  RM_MD5_STEP(F4(c, d, a), b, c, d, a, w(5), 0xfc93a039u, 21);
  RM_MD5_STEP(F4(b, c, d), a, b, c, d, w(12), 0x655b59c3u, 6);
  RM_MD5_STEP(F4(a, b, c), d, a, b, c, w(3), 0x8f0ccc92u, 10);
  RM_MD5_STEP(F4(d, a, b), c, d, a, b, w(10), 0xffeff47du, 15);
  RM_MD5_STEP(F4(c, d, a), b, c, d, a, w(1), 0x85845dd1u, 21);
  RM_MD5_STEP(F4(b, c, d), a, b, c, d, w(8), 0x6fa87e4fu, 6);
  RM_MD5_STEP(F4(a, b, c), d, a, b, c, w(15), 0xfe2ce6e0u, 10);
  RM_MD5_STEP(F4(d, a, b), c, d, a, b, w(6), 0xa3014314u, 15);
  RM_MD5_STEP(F4(c, d, a), b, c, d, a, w(13), 0x4e0811a1u, 21);
  RM_MD5_STEP(F4(b, c, d), a, b, c, d, w(4), 0xf7537e82u, 6);
  RM_MD5_STEP(F4(a, b, c), d, a, b, c, w(11), 0xbd3af235u, 10);
  RM_MD5_STEP(F4(d, a, b), c, d, a, b, w(2), 0x2ad7d2bbu, 15);
  RM_MD5_STEP(F4(c, d, a), b, c, d, a, w(9), 0xeb86d391u, 21);
#undef F1
#undef F2
#undef F3
#undef F4
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
}

// word column of the calling thread in the block's shared message buffer
struct RmMd5Words {
  unsigned* col;     // &wsm[0][threadIdx.x]
  int stride;        // threads per block
  __device__ __forceinline__ unsigned operator()(int i) const { return col[i * stride]; }
  __device__ __forceinline__ void set(int i, unsigned v) const { col[i * stride] = v; }
};

// The common case, 0 <= v < 1000000 (the token ids of every vocabulary in use: 32 K to 262 K entries), without a
// loop: the top two digits and the lower four as two-digit numbers (n / 10 == n * 103 >> 10 for n < 100; the lower
// two pairs share one register: neither the products nor the differences carry from one half into the other),
// leading zeros shifted out.  Returns the length; text in s, first character lowest (at most 8 bytes with the
// separator).
__device__ __forceinline__ int rm_render_small(unsigned v, int lead, u64& s) {
  const unsigned top = v / 10000u;                          // 0..99
  const unsigned r = v - top * 10000u;
  const unsigned h = r / 100u;
  const unsigned p = h | ((r - h * 100u) << 16);
  const unsigned t = ((p * 103u) >> 10) & 0x000f000fu;      // tens of both halves
  const unsigned o = p - t * 10u;                           // ones of both halves
  const unsigned w = (t | (o << 8)) + 0x30303030u;          // d3 d2 d1 d0
  const unsigned tt = (top * 103u) >> 10;
  const unsigned w2 = (tt | ((top - tt * 10u) << 8)) + 0x3030u;   // d5 d4
  u64 t6 = (u64)w2 | ((u64)w << 16);
  const int nd = 1 + (v >= 10u) + (v >= 100u) + (v >= 1000u) + (v >= 10000u) + (v >= 100000u);
  t6 >>= 8 * (6 - nd);
  if (lead == 1) {
    s = (t6 << 16) | ((u64)' ' << 8) | (u64)',';
    return nd + 2;
  }
  if (lead == 2) {
    s = (t6 << 8) | (u64)'[';
    return nd + 1;
  }
  s = t6;
  return nd;
}

// Text of one list item, first character in the lowest byte of (lo, hi): the decimal digits of v ('-' in front
// of a negative one), preceded by ", " (lead 1) or "[" (lead 2) or nothing (lead 0).  At most 13 bytes.
__device__ __forceinline__ int rm_render_token(int v, int lead, u64& tlo, u64& thi) {
  if ((unsigned)v < 1000000u) {
    thi = 0;
    return rm_render_small((unsigned)v, lead, tlo);
  }
  // decimal digits, least significant first; shifting what is there up by a byte puts the most
  // significant digit in the lowest byte, i.e. first in the text
  int len = 0;
  tlo = 0;
  thi = 0;
  unsigned ux = v < 0 ? (unsigned)(-(long long)v) : (unsigned)v;
  do {
    thi = (thi << 8) | (tlo >> 56);
    tlo = (tlo << 8) | (u64)('0' + (ux % 10u));
    ux /= 10u;
    ++len;
  } while (ux);
  if (v < 0) {
    thi = (thi << 8) | (tlo >> 56);
    tlo = (tlo << 8) | (u64)'-';
    ++len;
  }
  if (lead == 1) {
    thi = (thi << 16) | (tlo >> 48);
    tlo = (tlo << 16) | ((u64)' ' << 8) | (u64)',';
    len += 2;
  } else if (lead == 2) {
    thi = (thi << 8) | (tlo >> 56);
    tlo = (tlo << 8) | (u64)'[';
    ++len;
  }
  return len;
}

// ConsistentHash._hash(key) for key = list of ints (cache_aware_router.py:110-118): MD5 of Python's
// str(list) rendering "[t0, t1, ...]", first four digest bytes big-endian.
//
// Shape of the loop (it matters: the 32 requests of a warp render numbers of different widths, and a
// version that compressed "whenever this lane's block filled up" ran the 64-step compression once per
// LANE instead of once per warp, 4.2 ms for 16K 2K-token requests).  Each lane owns a byte queue of up
// to 16 pending bytes (mlo/mhi).  The inner loop only renders tokens and moves whole words from the
// queue into the lane's block; it ends when the block is full.  The compression sits after it, at a
// point where the lanes of a warp have reconverged.
__device__ unsigned rm_hash_token_list(const int* __restrict__ key, int n, const RmMd5Words& w) {
  RmMd5 m;
  rm_md5_init(m);
  u64 mlo = (u64)'[', mhi = 0;   // pending bytes, little-endian (first byte lowest)
  int pend = 1;                  // number of pending bytes
  u64 total = 1;
  int i = 0;
  bool closed = false;
  int widx = 0;
  while (true) {
    // One body for every lane and every iteration: (1) if fewer than 4 bytes are pending and input is
    // left, render one item (a token with its separator, or the closing bracket); (2) move as many whole
    // words as the block still takes.  No lane-dependent branch apart from the digit count.
    while (widx < 16) {
      if (pend < 4) {
        u64 tlo = 0, thi = 0;
        int len = 0;
        if (i < n) {
          len = rm_render_token(key[i], i ? 1 : 0, tlo, thi);
          ++i;
        } else if (!closed) {
          tlo = (u64)']';
          len = 1;
          closed = true;
        } else {
          break;                 // input exhausted, fewer than 4 bytes pending
        }
        const int sh = 8 * pend;
        mhi = (thi << sh) | (sh ? (tlo >> (64 - sh)) : 0ull);
        mlo |= tlo << sh;
        pend += len;
        total += (u64)len;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (pend >= 4 && widx < 16) {
          w.set(widx++, (unsigned)mlo);
          mlo = (mlo >> 32) | (mhi << 32);
          mhi >>= 32;
          pend -= 4;
        }
      }
    }
    if (widx < 16) break;
    rm_md5_block(m, w);
    widx = 0;
  }
  // padding: 0x80, zeros to 56 mod 64, the bit length (RFC 1321 3.1-3.2)
  const u64 bits = total * 8ull;
  mlo |= (u64)0x80 << (8 * pend);
  w.set(widx++, (unsigned)mlo);          // pend < 4: the 0x80 completes at most one word
  if (widx > 14) {
    while (widx < 16) w.set(widx++, 0u);
    rm_md5_block(m, w);
    widx = 0;
  }
  while (widx < 14) w.set(widx++, 0u);
  w.set(14, (unsigned)(bits & 0xffffffffu));
  w.set(15, (unsigned)(bits >> 32));
  rm_md5_block(m, w);
  // digest bytes 0..3 are the little-endian bytes of `a`; the reference reads them big-endian
  const unsigned a = m.a;
  return ((a & 0xffu) << 24) | ((a & 0xff00u) << 8) | ((a >> 8) & 0xff00u) | (a >> 24);
}

// ---- two-step form of the same hash, for long requests --------------------------------------------------
// One thread rendering AND compressing a 2K-token request is ~250 dependent instructions per token at an
// occupancy of a few warps per SM: 1.4 ms for a batch, however small the batch.  The text of a request is
// independent per token, so a warp renders it in parallel (item lengths -> warp prefix sum -> whole words) into
// a scratch area, padded per RFC 1321; the compression, the only sequential part, then reads 64-byte blocks.
// Place of request q's message: 64 * ((qoff[q] - qoff[0] + 3) / 4 + 2 q) bytes, which leaves room for 13 bytes
// per token + brackets + padding without a prefix sum over the requests.  A request that does not fit the
// scratch keeps nblk = 0 and is hashed by the streamed form above.
#define RM_RENDER_TPB 256
#define RM_RENDER_WPB (RM_RENDER_TPB / 32)
#define RM_RENDER_BUF (128 * 8 + 16)

__device__ __forceinline__ u64 rm_msg_off(i64 rel_tok, int q) { return 64ull * ((u64)((rel_tok + 3) >> 2) + 2ull * (u64)q); }

__device__ __forceinline__ void rm_put_bytes(unsigned char* dst, u64 lo, u64 hi, int len) {
#pragma unroll
  for (int b = 0; b < 8; ++b)
    if (b < len) dst[b] = (unsigned char)(lo >> (8 * b));
  for (int b = 8; b < len; ++b) dst[b] = (unsigned char)(hi >> (8 * (b - 8)));
}

// closing bracket, 0x80, zeros to 56 mod 64, bit length; returns the number of 64-byte blocks
__device__ __forceinline__ int rm_msg_finish(unsigned char* dst, u64 total) {
  dst[total++] = (unsigned char)']';
  const u64 bits = total * 8ull;
  dst[total++] = 0x80;
  while ((total & 63ull) != 56ull) dst[total++] = 0;
  for (int b = 0; b < 8; ++b) dst[total++] = (unsigned char)((bits >> (8 * b)) & 0xffu);
  return (int)(total >> 6);
}

// ---- four items per lane (rm_route_render_kernel's usual chunk) ---------------------------------------------
// A lane's text starts sh = (its first byte's place) & 3 bytes into a word; those sh bytes are the end of the text
// of the lane before it.  rm_quad_rest: the last `rest` (0..3) bytes of a lane's text, which are the tail of its
// last item because an item with its separator is at least 3 bytes long.
__device__ __forceinline__ unsigned rm_quad_rest(u64 s3, int l3, unsigned rest) {
  return (unsigned)(s3 >> (8 * ((unsigned)l3 - rest)));
}

// The first 32 bytes of cin (low sh bytes) ++ s0 ++ s1 ++ s2 ++ s3 in q0..q3, first byte lowest.  The text is at most
// 3 + 4 * 8 = 35 bytes; what lies beyond the last whole 32-bit word is never stored by this lane (rm_quad_rest hands
// it to the next), so bytes 32..34 need no home.  An item lands in the 64-bit word its first byte falls in and spills
// into the next (the split shift is written so that a byte offset of 0 shifts by 56 + 8, never by 64).
__device__ __forceinline__ void rm_quad_pack(unsigned cin, unsigned sh, u64 s0, int l0, u64 s1, int l1, u64 s2, int l2, u64 s3,
                                             int l3, u64& q0, u64& q1, u64& q2, u64& q3) {
  (void)l3;
  cin &= (1u << (8 * sh)) - 1u;
  const unsigned r0 = sh, r1 = r0 + (unsigned)l0, r2 = r1 + (unsigned)l1, r3 = r2 + (unsigned)l2;
  q0 = (u64)cin | (s0 << (8 * r0));
  q1 = (s0 >> 8) >> (56 - 8 * r0);
  q2 = 0;
  q3 = 0;
  {
    const unsigned rr = 8 * (r1 & 7u);
    const u64 lo = s1 << rr, hi = (s1 >> 8) >> (56 - rr);
    if (r1 < 8) { q0 |= lo; q1 |= hi; } else { q1 |= lo; q2 |= hi; }
  }
  {
    const unsigned rr = 8 * (r2 & 7u), ix = r2 >> 3;
    const u64 lo = s2 << rr, hi = (s2 >> 8) >> (56 - rr);
    if (ix == 0) { q0 |= lo; q1 |= hi; } else if (ix == 1) { q1 |= lo; q2 |= hi; } else { q2 |= lo; q3 |= hi; }
  }
  {
    const unsigned rr = 8 * (r3 & 7u), ix = r3 >> 3;
    const u64 lo = s3 << rr, hi = (s3 >> 8) >> (56 - rr);
    if (ix == 0) { q0 |= lo; q1 |= hi; } else if (ix == 1) { q1 |= lo; q2 |= hi; }
    else if (ix == 2) { q2 |= lo; q3 |= hi; } else { q3 |= lo; }
  }
}

// A WARP per request (a block per request with four block barriers per 128-token chunk was 1.4 ms for 64K
// 2K-token requests, six times its instruction count: the barriers, not the issue slots, set the pace).
__global__ void __launch_bounds__(RM_RENDER_TPB)
rm_route_render_kernel(const int* __restrict__ qtok, const i64* __restrict__ qoff, int B, int warm,
                       const int* __restrict__ prefill, const int* __restrict__ decode, const unsigned char* __restrict__ err,
                       unsigned char* __restrict__ msg, u64 msg_bytes, int* __restrict__ nblk) {
#ifdef RM_EMU
  const int lane = 0, q = blockIdx.x;           // one thread per request
#else
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int q = blockIdx.x * RM_RENDER_WPB + wid;
#endif
  if (q >= B) return;
  const bool need = warm ? ((prefill[q] == -1 || decode[q] == -1) && err[q] == 0) : true;
  const i64 o0 = qoff[q];
  const int n = (int)(qoff[q + 1] - o0);
  const u64 base = rm_msg_off(o0 - qoff[0], q);
  const bool fits = msg != nullptr && rm_msg_off(qoff[q + 1] - qoff[0], q + 1) <= msg_bytes;
  if (!need || !fits) {
    if (lane == 0) nblk[q] = 0;
    return;
  }
  const int* __restrict__ key = qtok + o0;
  unsigned char* dst = msg + base;
#ifdef RM_EMU
  if (lane == 0) {
    // the device's chunking, lanes visited in order: full chunks of 128 small ids through rm_quad_pack (whole words
    // out, the rest handed to the next lane), anything else 32 items at a time byte by byte
    u64 total = 0;
    if (n == 0) dst[total++] = (unsigned char)'[';
    int c0 = 0;
    while (c0 < n) {
      bool quad = c0 + 128 <= n;
      for (int i = c0; quad && i < c0 + 128; ++i) quad = (unsigned)key[i] < 1000000u;
      if (quad) {
        unsigned cout = 0;
        const unsigned left = (unsigned)(total & 3ull);
        for (unsigned b2 = 0; b2 < left; ++b2) cout |= (unsigned)dst[total - left + b2] << (8 * b2);
        for (int l = 0; l < 32; ++l) {
          u64 s0, s1, s2, s3, qq[4];
          const int* kp = key + c0 + 4 * l;
          const int l0 = rm_render_small((unsigned)kp[0], (c0 | l) ? 1 : 2, s0);
          const int l1 = rm_render_small((unsigned)kp[1], 1, s1);
          const int l2 = rm_render_small((unsigned)kp[2], 1, s2);
          const int l3 = rm_render_small((unsigned)kp[3], 1, s3);
          const unsigned tot = (unsigned)(l0 + l1 + l2 + l3);
          const unsigned sh = (unsigned)(total & 3ull);
          const unsigned cin = cout;
          cout = rm_quad_rest(s3, l3, (unsigned)((total + tot) & 3ull));
          rm_quad_pack(cin, sh, s0, l0, s1, l1, s2, l2, s3, l3, qq[0], qq[1], qq[2], qq[3]);
          const unsigned nw = (sh + tot) >> 2;
          unsigned char* o = dst + (total - sh);
          for (unsigned w = 0; w < nw; ++w)
            for (int b2 = 0; b2 < 4; ++b2) o[4 * w + b2] = (unsigned char)(qq[w >> 1] >> (32 * (w & 1) + 8 * b2));
          total += tot;
        }
        const unsigned rest = (unsigned)(total & 3ull);
        for (unsigned b2 = 0; b2 < rest; ++b2) dst[total - rest + b2] = (unsigned char)(cout >> (8 * b2));
        c0 += 128;
        continue;
      }
      for (int i = c0; i < n && i < c0 + 32; ++i) {
        u64 lo, hi;
        const int len = rm_render_token(key[i], i ? 1 : 2, lo, hi);
        rm_put_bytes(dst + total, lo, hi, len);
        total += (u64)len;
      }
      c0 += 32;
    }
    nblk[q] = rm_msg_finish(dst, total);
  }
#else
  // Text leaves the warp's shared buffer as whole 32-bit words (byte stores straight to global memory: 250 M
  // partial-sector writes per batch, 0.37 ms); the 0-3 bytes that do not fill a word wait at the front of the
  // buffer for the next chunk.
  //
  // Full chunks of 128 small tokens, four per lane (the usual case): a lane strings its four items together in
  // registers behind the 0-3 bytes that the lane before it could not fit into a whole word (those are the tail
  // of that lane's last item, at least 3 bytes long: one shuffle), stores its whole words, and hands its own
  // remainder on.  No byte stores, one prefix sum and one flush per 128 tokens: ~55 instructions per token
  // against ~240 for the 32-token form below, which remains for the last chunk and for ids outside [0, 1000000).
  __shared__ __align__(16) unsigned char bufs[RM_RENDER_WPB][RM_RENDER_BUF];
  unsigned char* buf = bufs[wid];
  unsigned* bw = reinterpret_cast<unsigned*>(buf);
  u64 flushed = 0;       // bytes already in global memory (multiple of 4)
  unsigned left = 0;     // bytes waiting in buf[0..left)
  if (n == 0) {
    if (lane == 0) buf[0] = (unsigned char)'[';
    left = 1;
  }
  int c0 = 0;
  while (c0 < n) {
    bool quad = false;
    int v0 = 0, v1 = 0, v2 = 0, v3 = 0;
    if (c0 + 128 <= n) {
      const int* kp = key + c0 + 4 * lane;
      v0 = __ldg(kp);
      v1 = __ldg(kp + 1);
      v2 = __ldg(kp + 2);
      v3 = __ldg(kp + 3);
      quad = __all_sync(0xffffffffu, ((unsigned)v0 < 1000000u) & ((unsigned)v1 < 1000000u) & ((unsigned)v2 < 1000000u) &
                                         ((unsigned)v3 < 1000000u));
    }
    if (quad) {
      u64 s0, s1, s2, s3;
      const int l0 = rm_render_small((unsigned)v0, (c0 | lane) ? 1 : 2, s0);
      const int l1 = rm_render_small((unsigned)v1, 1, s1);
      const int l2 = rm_render_small((unsigned)v2, 1, s2);
      const int l3 = rm_render_small((unsigned)v3, 1, s3);
      const unsigned tot = (unsigned)(l0 + l1 + l2 + l3);
      unsigned inc = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += x;
      }
      const unsigned chunk_total = __shfl_sync(0xffffffffu, inc, 31);
      const unsigned p0 = left + inc - tot;                  // this lane's first byte in the buffer
      const unsigned sh = p0 & 3u;                           // bytes of the word it starts in that come from before
      const unsigned sh_out = (p0 + tot) & 3u;               // bytes this lane leaves to the next
      const unsigned cout = rm_quad_rest(s3, l3, sh_out);
      unsigned cin = __shfl_up_sync(0xffffffffu, cout, 1);
      if (lane == 0) cin = bw[0];                            // carried over from the chunk before
      u64 q0, q1, q2, q3;
      rm_quad_pack(cin, sh, s0, l0, s1, l1, s2, l2, s3, l3, q0, q1, q2, q3);
      const unsigned nw = (sh + tot) >> 2;                   // whole words: 2..8
      unsigned* o = bw + (p0 >> 2);                          // word 0 is lane 0's alone: read above, written here
      o[0] = (unsigned)q0;
      o[1] = (unsigned)(q0 >> 32);
      if (nw > 2) o[2] = (unsigned)q1;
      if (nw > 3) o[3] = (unsigned)(q1 >> 32);
      if (nw > 4) o[4] = (unsigned)q2;
      if (nw > 5) o[5] = (unsigned)(q2 >> 32);
      if (nw > 6) o[6] = (unsigned)q3;
      if (nw > 7) o[7] = (unsigned)(q3 >> 32);
      __syncwarp();
      const unsigned avail = left + chunk_total;
      const unsigned words = avail >> 2;
      unsigned* __restrict__ gw = reinterpret_cast<unsigned*>(dst + flushed);
      for (unsigned w = lane; w < words; w += 32) gw[w] = bw[w];
      __syncwarp();
      if (lane == 31) bw[0] = cout;                          // avail & 3 bytes, the low ones
      flushed += 4ull * words;
      left = avail & 3u;
      __syncwarp();
      c0 += 128;
      continue;
    }
    const int i = c0 + lane;
    u64 lo = 0, hi = 0;
    int len = 0;
    if (i < n) len = rm_render_token(__ldg(key + i), i ? 1 : 2, lo, hi);
    unsigned inc = (unsigned)len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    const unsigned chunk_total = __shfl_sync(0xffffffffu, inc, 31);
    rm_put_bytes(buf + left + inc - (unsigned)len, lo, hi, len);
    __syncwarp();
    const unsigned avail = left + chunk_total;
    const unsigned words = avail >> 2;
    unsigned* __restrict__ gw = reinterpret_cast<unsigned*>(dst + flushed);
    for (unsigned w = lane; w < words; w += 32) gw[w] = bw[w];
    const unsigned rest = avail & 3u;
    unsigned char carry = 0;
    if ((unsigned)lane < rest) carry = buf[4 * words + lane];
    __syncwarp();
    if ((unsigned)lane < rest) buf[lane] = carry;
    flushed += 4ull * words;
    left = rest;
    __syncwarp();
    c0 += 32;
  }
  // closing bracket, 0x80, zeros to 56 mod 64, bit length (RFC 1321): at most 3 + 73 bytes, built in the buffer
  const u64 text = flushed + left + 1;                           // with the closing bracket
  const u64 padded = (text + 1 + 8 + 63) & ~63ull;
  const unsigned tail = (unsigned)(padded - flushed);            // bytes of the buffer to flush, multiple of 4
  const u64 bits = text * 8ull;
  for (unsigned j = left + lane; j < tail; j += 32) {
    unsigned char v = 0;
    if (j == left) v = (unsigned char)']';
    else if (j == left + 1) v = 0x80;
    else if (j >= tail - 8) v = (unsigned char)(bits >> (8 * (j - (tail - 8))));
    buf[j] = v;
  }
  __syncwarp();
  unsigned* __restrict__ gw = reinterpret_cast<unsigned*>(dst + flushed);
  for (unsigned w = lane; w < (tail >> 2); w += 32) gw[w] = bw[w];
  if (lane == 0) nblk[q] = (int)(padded >> 6);
#endif
}

struct RmMd5Regs {
  unsigned m[16];
  __device__ __forceinline__ unsigned operator()(int i) const { return m[i]; }
};

// MD5 of a rendered, padded message of nb 64-byte blocks (64-byte aligned).  One thread per request and 50-250
// blocks per request: a chain of ~3 dependent instructions per step with one or two warps per scheduler, so a
// block that waits for its 64 bytes waits in full.  Loads written as "next block into registers while this one
// is compressed" were scheduled two thirds of the way through the compression (the words of a block stay live
// until its last round), ~350 cycles ahead of their use against ~800 of DRAM latency.  The blocks therefore
// come through a ring of RM_MD5_STAGES asynchronous 64-byte copies per thread in shared memory
// (ring[stage][quarter][thread], 16 bytes each: conflict free), RM_MD5_STAGES - 1 compressions ahead.
#define RM_MD5_STAGES 4

__device__ __forceinline__ unsigned rm_digest_word(const RmMd5& m) {
  const unsigned a = m.a;
  return ((a & 0xffu) << 24) | ((a & 0xff00u) << 8) | ((a >> 8) & 0xff00u) | (a >> 24);
}

#ifdef RM_EMU
__device__ unsigned rm_hash_rendered(const unsigned char* __restrict__ msg, int nb, uint4* ring) {
  (void)ring;
  RmMd5 m;
  rm_md5_init(m);
  const unsigned* p = reinterpret_cast<const unsigned*>(msg);
  for (int b = 0; b < nb; ++b) {
    RmMd5Regs w;
    for (int i = 0; i < 16; ++i) w.m[i] = p[16 * b + i];
    rm_md5_block(m, w);
  }
  return rm_digest_word(m);
}
#else
__device__ __forceinline__ void rm_md5_fetch(uint4* ring, int stage, const unsigned char* src) {
  const unsigned dst = (unsigned)__cvta_generic_to_shared(ring + (size_t)stage * 4 * RM_RT_TPB);
#pragma unroll
  for (int j = 0; j < 4; ++j)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (unsigned)(j * RM_RT_TPB * 16)), "l"(src + 16 * j)
                 : "memory");
}

// ring: &ring[0][0][threadIdx.x]
__device__ unsigned rm_hash_rendered(const unsigned char* __restrict__ msg, int nb, uint4* ring) {
  RmMd5 m;
  rm_md5_init(m);
#pragma unroll
  for (int s = 0; s < RM_MD5_STAGES - 1; ++s) {
    if (s < nb) rm_md5_fetch(ring, s, msg + 64 * s);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  int stage = 0, ahead = RM_MD5_STAGES - 1;
  for (int b = 0; b < nb; ++b) {
    if (b + RM_MD5_STAGES - 1 < nb) rm_md5_fetch(ring, ahead, msg + 64 * (size_t)(b + RM_MD5_STAGES - 1));
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group %0;" ::"n"(RM_MD5_STAGES - 1) : "memory");
    const uint4* r = ring + (size_t)stage * 4 * RM_RT_TPB;
    const uint4 x0 = r[0], x1 = r[RM_RT_TPB], x2 = r[2 * RM_RT_TPB], x3 = r[3 * RM_RT_TPB];
    RmMd5Regs w;
    w.m[0] = x0.x; w.m[1] = x0.y; w.m[2] = x0.z; w.m[3] = x0.w;
    w.m[4] = x1.x; w.m[5] = x1.y; w.m[6] = x1.z; w.m[7] = x1.w;
    w.m[8] = x2.x; w.m[9] = x2.y; w.m[10] = x2.z; w.m[11] = x2.w;
    w.m[12] = x3.x; w.m[13] = x3.y; w.m[14] = x3.z; w.m[15] = x3.w;
    rm_md5_block(m, w);
    stage = stage + 1 == RM_MD5_STAGES ? 0 : stage + 1;
    ahead = ahead + 1 == RM_MD5_STAGES ? 0 : ahead + 1;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  return rm_digest_word(m);
}
#endif

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
                       int ring_p_n, const unsigned* ring_d, const int* ring_d_owner, int ring_d_n,
                       const unsigned char* __restrict__ msg, const int* __restrict__ nblk) {
  __shared__ unsigned wsm[16 * RM_RT_TPB];
#ifdef RM_EMU
  uint4* ring = nullptr;
#else
  __shared__ uint4 ring_sm[RM_MD5_STAGES * 4 * RM_RT_TPB];
  uint4* ring = ring_sm + threadIdx.x;
#endif
  const RmMd5Words words = {wsm + threadIdx.x, RM_RT_TPB};
#ifdef RM_EMU
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
#else
  // The compression runs as long as the longest request of the warp, and requests that picked both ranks in the
  // walk need none: with lengths of 1K-4K tokens and a third of the lanes idle a warp did 2.4x the work of its
  // lanes, and at 64K requests the kernel is bound by the integer pipe.  The block's requests are dealt to its
  // threads in order of message length (rank by counting in shared memory), so the lanes of a warp finish together
  // and the warps left with nothing to hash leave at once: 406 -> 333 us at 64K requests.
  __shared__ int skey[RM_RT_TPB];
  __shared__ int sperm[RM_RT_TPB];
  int q;
  {
    const int q0 = blockIdx.x * blockDim.x + threadIdx.x;
    int k = -1;
    if (q0 < B) {
      const bool need = warm ? ((prefill[q0] == -1 || decode[q0] == -1) && err[q0] == 0) : true;
      const int nb = nblk ? nblk[q0] : 0;
      k = !need ? 0 : nb > 0 ? nb : (int)((qoff[q0 + 1] - qoff[q0]) >> 3) + 1;     // streamed form: ~8 tokens a block
    }
    skey[threadIdx.x] = k;
    __syncthreads();
    int rank = 0;
    for (int j = 0; j < RM_RT_TPB; ++j) {
      const int kj = skey[j];
      rank += (kj > k) | ((kj == k) & (j < (int)threadIdx.x));
    }
    sperm[rank] = q0;
    __syncthreads();
    q = sperm[threadIdx.x];        // (giving the long quarter to a different warp in every block: 333 -> 380 us)
  }
#endif
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
    if (need_hash) {
      const int nb = nblk ? nblk[q] : 0;
      if (nb > 0) h = rm_hash_rendered(msg + rm_msg_off(qoff[q] - qoff[0], q), nb, ring);     // rendered by rm_route_render_kernel
      else h = rm_hash_token_list(qtok + qoff[q], (int)(qoff[q + 1] - qoff[q]), words);
    }
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

// renders the requests that will need the hash (all of them when warm == 0) into the route scratch, if the
// index has one (rm_route_set_rings allocates it) and the batch is within max_batch
static bool rm_launch_render(rm_index* idx, const int* tokens, const i64* offsets, int B, int warm, const int* prefill,
                             const int* decode, const unsigned char* err, cudaStream_t s) {
  if (!idx->route_nblk || !idx->route_msg || B > idx->cfg.max_batch) return false;
#ifdef RM_EMU
  const int tpb = 1, grid = B;
#else
  const int tpb = RM_RENDER_TPB, grid = (B + RM_RENDER_WPB - 1) / RM_RENDER_WPB;
#endif
  RM_LAUNCH(rm_route_render_kernel, grid, tpb, s, tokens, offsets, B, warm, prefill, decode, err, idx->route_msg,
            (u64)idx->route_msg_bytes, idx->route_nblk);
  return true;
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
  // (Rendering every request on a second stream beside the walk, for small batches: no gain at 2K-16K requests, the
  // two kernels take turns on the SMs rather than share them; 8K requests 0.375 ms either way.)
  const bool rendered = rm_launch_render(idx, tokens, offsets, B, warm, out_prefill, out_decode, out_err, s);
  RM_LAUNCH(rm_route_finish_kernel, (B + RM_RT_TPB - 1) / RM_RT_TPB, RM_RT_TPB, s, idx->t, tokens, offsets, B, warm,
            out_prefill, out_decode, out_err, out_paddr, out_daddr, idx->ring_p, idx->ring_p_owner, idx->ring_p_n,
            idx->ring_d, idx->ring_d_owner, idx->ring_d_n, rendered ? idx->route_msg : nullptr,
            rendered ? idx->route_nblk : nullptr);
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
  const bool rendered = rm_launch_render(idx, tokens, (const i64*)offsets, B, 1, out_prefill, out_decode, out_err, s);
  RM_LAUNCH(rm_route_finish_kernel, (B + RM_RT_TPB - 1) / RM_RT_TPB, RM_RT_TPB, s, idx->t, tokens, (const i64*)offsets, B, 1,
            out_prefill, out_decode, out_err, out_paddr, out_daddr, idx->ring_p, idx->ring_p_owner, idx->ring_p_n,
            idx->ring_d, idx->ring_d_owner, idx->ring_d_n, rendered ? idx->route_msg : nullptr,
            rendered ? idx->route_nblk : nullptr);
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
  rc = rm_set_ring(&idx->ring_d, &idx->ring_d_owner, &idx->ring_d_n, d_points_h, d_owner_h, n_d);
  if (rc) return rc;
  // scratch of the rendered hash messages: 16 bytes per token + 128 per request of the largest batch
  // (rm_config.max_batch_tokens; without it max_batch * max_key_len capped at RM_ROUTE_SCRATCH_MB, default 1024);
  // requests that do not fit are hashed by the streamed form
  if (!idx->route_nblk) {
    size_t cap_mb = 1024;
    if (const char* e = getenv("RM_ROUTE_SCRATCH_MB")) cap_mb = (size_t)atoll(e);
    size_t want = (size_t)idx->cfg.max_batch * ((size_t)idx->cfg.max_key_len * 16 + 128) + 64;
    if (idx->cfg.max_batch_tokens > 0)      // the caller stated its largest batch: room for all of it, no cap
      want = (size_t)idx->cfg.max_batch_tokens * 16 + (size_t)idx->cfg.max_batch * 128 + 256;
    else if (want > (cap_mb << 20)) want = cap_mb << 20;
    RM_CUDA(cudaMalloc(reinterpret_cast<void**>(&idx->route_nblk), sizeof(int) * (size_t)idx->cfg.max_batch));
    idx->route_msg = nullptr;
    idx->route_msg_bytes = 0;
    if (want >= 256) {
      RM_CUDA(cudaMalloc(reinterpret_cast<void**>(&idx->route_msg), want));
      idx->route_msg_bytes = want;
    }
  }
  return RM_OK;
}
