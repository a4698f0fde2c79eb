// Paged keys on the device: RadixCache(page_size > 1) (radix_cache.py:23-32, 109-114, 148-150).
//
// With page_size P the reference keys children by their first PAGE (tuple(key[:P])), compares keys a page at a time
// (_key_match_paged) and therefore splits only at page multiples.  A paged radix tree over tokens is exactly a radix
// tree over page SYMBOLS, so the index (walk, split, insert, evict kernels) runs unchanged once pages are interned:
//   rm_pages_intern     tokens -> symbols, one thread per page: 64-bit hash of the page, open-addressed table, every hit
//                       verified against the stored page (an unseen page gets the next symbol when create != 0; in
//                       lookup mode it and everything after it in its key become the "no such page" symbol, which no
//                       edge holds, so the walk stops there exactly like the reference's dict miss);
//   rm_pages_rows_*     the value of a symbol position is the id of a ROW of P KV indices in a device row store
//                       (put: rows for the pages of an insert; free: the rows of evicted runs and of pages an insert
//                       found already present; get: expand row ids to the KV indices match_prefix / evict hand back).
//   rm_pages_retain     drops the pages no edge of the tree holds any more (evicted since they were interned): their
//                       symbols are handed out again and the table is rebuilt from the live ones.  The reference's
//                       dict entries die with their nodes; without this the symbol space only grows.
// Keys are taken in whole pages; a short last page arrives filled up with a reserved token (include/radixmesh_b200.h).
#include "rm_internal.h"

#define RM_PG_TPB 128
#define RM_PAGE_NONE (-2)          // symbol of a page the table does not hold (lookup mode)

struct rm_pages {
  int device;
  int P;
  i64 max_pages, max_rows;
  unsigned mask;                   // table capacity - 1
  u64* tag;                        // 0 = empty
  int* sym;                        // -1 until the creator has published it
  int* store;                      // page tokens by symbol: store[sym * P + j]
  i64* vrows;                      // value rows: vrows[row * P + j]
  i64* rows_free;                  // ring of free row ids
  u64* ctr;                        // [0] symbols handed out, [1] rows popped, [2] rows pushed, [3] sticky error,
                                   // [4] free symbols taken, [5] free symbols there are (rm_pages_retain)
  int* sym_free;                   // symbols rm_pages_retain found unused, handed out again before fresh ones
  unsigned char* live;             // rm_pages_retain: symbol is held by an edge of the tree
};

__device__ __forceinline__ u64 rm_page_hash(const int* __restrict__ p, int P) {
  u64 h = 0x9E3779B97F4A7C15ull ^ (u64)P;
  for (int j = 0; j < P; ++j) {
    h ^= (u64)(unsigned)p[j] + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
    h *= 0xFF51AFD7ED558CCDull;
    h ^= h >> 33;
  }
  return h | 1ull;                 // never the empty tag
}

__device__ __forceinline__ bool rm_page_same(const int* __restrict__ a, const int* __restrict__ b, int P) {
  for (int j = 0; j < P; ++j)
    if (a[j] != b[j]) return false;
  return true;
}

// pass A: one thread per page.  pend[i] = slot to look at again in pass B (a page another thread of THIS launch is
// creating: its symbol is not published yet), else -1.
__global__ void rm_pages_claim_kernel(rm_pages pg, const int* __restrict__ tokens, const i64* __restrict__ offsets, int B,
                                      int create, const i64* __restrict__ sym_offsets, int* __restrict__ out_syms,
                                      int* __restrict__ pend, int total_pages_bound) {
  const i64 gi = (i64)blockIdx.x * blockDim.x + threadIdx.x;     // global page index in the sym_offsets layout
  if (gi >= total_pages_bound) return;
  // key q with sym_offsets[q] <= gi < sym_offsets[q + 1]: binary search
  int lo = 0, hi = B;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (sym_offsets[mid] <= gi) lo = mid;
    else hi = mid;
  }
  const int q = lo;
  if (gi >= sym_offsets[q + 1]) return;                          // (padding between keys, if the caller left any)
  const int k = (int)(gi - sym_offsets[q]);
  const int* __restrict__ page = tokens + offsets[q] + (i64)k * pg.P;
  const u64 tg = rm_page_hash(page, pg.P);
  unsigned i = (unsigned)(tg >> 20) & pg.mask;
  int res = RM_PAGE_NONE, pd = -1;
  for (unsigned steps = 0; steps <= pg.mask; ++steps) {
    u64 cur;
    if (create) cur = atomicCAS(&pg.tag[i], 0ull, tg);
    else cur = *reinterpret_cast<volatile u64*>(&pg.tag[i]);
    if (cur == 0ull) {
      if (create) {                                              // the slot is mine: next symbol, page stored, then published
        // a symbol rm_pages_retain freed, if one is left, else the next fresh one ([4] may run past [5]: both are
        // set again by the next retain)
        u64 s;
        const u64 fr = pg.ctr[5] ? atomicAdd(&pg.ctr[4], 1ull) : ~0ull;
        if (fr < pg.ctr[5]) s = (u64)pg.sym_free[fr];
        else s = atomicAdd(&pg.ctr[0], 1ull);
        if ((i64)s >= pg.max_pages) {
          atomicMax(&pg.ctr[3], (u64)RM_E_NOSPACE);
          break;
        }
        for (int j = 0; j < pg.P; ++j) pg.store[(i64)s * pg.P + j] = page[j];
        __threadfence();
        *reinterpret_cast<volatile int*>(&pg.sym[i]) = (int)s;
        res = (int)s;
      }
      break;                                                     // lookup: no such page
    }
    if (cur == tg) {
      const int s = *reinterpret_cast<volatile int*>(&pg.sym[i]);
      if (s < 0) {                                               // being created by another thread of this launch
        pd = (int)i;
        break;
      }
      if (rm_page_same(pg.store + (i64)s * pg.P, page, pg.P)) {
        res = s;
        break;
      }
    }
    i = (i + 1) & pg.mask;                                       // another page with this slot (or, 2^-64, this hash)
  }
  out_syms[gi] = res;
  pend[gi] = pd;
}

// pass B: pages that met a same-launch creator read its symbol now and verify the content.  Two DIFFERENT new pages of
// one batch with the same 64-bit hash cannot both be placed by this scheme: that raises the sticky error instead of
// returning a wrong symbol.
__global__ void rm_pages_resolve_kernel(rm_pages pg, const int* __restrict__ tokens, const i64* __restrict__ offsets, int B,
                                        const i64* __restrict__ sym_offsets, int* __restrict__ out_syms,
                                        const int* __restrict__ pend, int total_pages_bound) {
  const i64 gi = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= total_pages_bound) return;
  const int pd = pend[gi];
  if (pd < 0) return;
  int lo = 0, hi = B;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (sym_offsets[mid] <= gi) lo = mid;
    else hi = mid;
  }
  const int* __restrict__ page = tokens + offsets[lo] + (gi - sym_offsets[lo]) * pg.P;
  const int s = pg.sym[pd];
  if (s >= 0 && rm_page_same(pg.store + (i64)s * pg.P, page, pg.P)) out_syms[gi] = s;
  else atomicMax(&pg.ctr[3], (u64)RM_E_INVALID);
}

// lookup mode: everything after the first unknown page of a key is unknown too (the walk must not resume beyond it)
__global__ void rm_pages_truncate_kernel(int B, const i64* __restrict__ sym_offsets, int* __restrict__ out_syms,
                                         int* __restrict__ out_npages) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= B) return;
  const i64 a = sym_offsets[q], b = sym_offsets[q + 1];
  i64 k = a;
  while (k < b && out_syms[k] >= 0) ++k;
  if (out_npages) out_npages[q] = (int)(k - a);
  for (i64 x = k; x < b; ++x) out_syms[x] = RM_PAGE_NONE;
}

__global__ void rm_pages_reset_kernel(rm_pages pg) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 x = (i64)blockIdx.x * blockDim.x + threadIdx.x; x < pg.max_rows; x += stride) pg.rows_free[x] = x;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    pg.ctr[0] = 0;
    pg.ctr[1] = 0;
    pg.ctr[2] = (u64)pg.max_rows;   // ring: rows [popped, pushed) are free
    pg.ctr[3] = 0;
    pg.ctr[4] = 0;
    pg.ctr[5] = 0;
  }
}

__global__ void rm_pages_rows_put_kernel(rm_pages pg, const i64* __restrict__ values, i64 n, const u64* __restrict__ base_p,
                                         i64* __restrict__ out_rows) {
  const i64 x = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n * pg.P) return;
  const i64 r = x / pg.P;
  const i64 base = (i64)*base_p;
  const i64 row = pg.rows_free[(base + r) % pg.max_rows];
  pg.vrows[row * pg.P + x % pg.P] = values[x];
  if (x % pg.P == 0) out_rows[r] = row;
}

__global__ void rm_pages_rows_pop_kernel(rm_pages pg, i64 n, u64* __restrict__ out_base) {
  if (blockIdx.x || threadIdx.x) return;
  const u64 pos = pg.ctr[1];
  if (pos + (u64)n > pg.ctr[2]) {
    atomicMax(&pg.ctr[3], (u64)RM_E_NOSPACE);
    *out_base = pos;              // (the rows written below are then live ones: the error is sticky and loud)
    return;
  }
  pg.ctr[1] = pos + (u64)n;
  *out_base = pos;
}

__global__ void rm_pages_rows_free_kernel(rm_pages pg, const i64* __restrict__ rows, i64 n) {
  const i64 x = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n) return;
  const u64 pos = atomicAdd(&pg.ctr[2], 1ull);
  pg.rows_free[pos % (u64)pg.max_rows] = rows[x];
}

__global__ void rm_pages_rows_get_kernel(rm_pages pg, const i64* __restrict__ rows, i64 n, i64* __restrict__ out) {
  const i64 x = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n * pg.P) return;
  const i64 row = rows[x / pg.P];
  out[x] = (row >= 0 && row < pg.max_rows) ? pg.vrows[row * pg.P + x % pg.P] : -1;
}

__global__ void rm_pages_tokens_kernel(rm_pages pg, const int* __restrict__ syms, i64 n, int* __restrict__ out) {
  const i64 x = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n * pg.P) return;
  const int s = syms[x / pg.P];
  out[x] = (s >= 0 && (i64)s < pg.max_pages) ? pg.store[(i64)s * pg.P + x % pg.P] : -1;
}

static int rm_pg_grid(i64 n) { return (int)((n + RM_PG_TPB - 1) / RM_PG_TPB); }

// ---- rm_pages_retain -----------------------------------------------------------------------------------------------
// live[s] = 1 for every symbol an owned arena position of the index holds (evicted runs have their owner words cleared)
__global__ void rm_pages_mark_kernel(rm_pages pg, RmTree t) {
  const u64 top = t.ctr->arena_top;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < (i64)top; i += stride) {
    const int2 r = t.ts[i];
    if (r.y >= 0 && r.x >= 0 && (i64)r.x < pg.max_pages) pg.live[r.x] = 1;
  }
}

// every symbol handed out so far: a live one goes back into the (cleared) table, a dead one onto the free list
__global__ void rm_pages_sweep_kernel(rm_pages pg) {
  const i64 n = (i64)pg.ctr[0] < pg.max_pages ? (i64)pg.ctr[0] : pg.max_pages;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 s = (i64)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += stride) {
    if (!pg.live[s]) {
      const u64 at = atomicAdd(&pg.ctr[5], 1ull);
      pg.sym_free[at] = (int)s;
      continue;
    }
    const u64 tg = rm_page_hash(pg.store + s * pg.P, pg.P);
    unsigned i = (unsigned)(tg >> 20) & pg.mask;
    for (unsigned steps = 0; steps <= pg.mask; ++steps) {
      if (atomicCAS(&pg.tag[i], 0ull, tg) == 0ull) {
        pg.sym[i] = (int)s;
        break;
      }
      i = (i + 1) & pg.mask;
    }
  }
}

extern "C" {

int rm_pages_retain(rm_pages* p, rm_index* idx, void* stream) {
  if (!p || !idx) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  RM_CUDA(cudaMemsetAsync(p->live, 0, (size_t)p->max_pages, s));
  RM_LAUNCH(rm_pages_mark_kernel, 148 * 8, RM_PG_TPB, s, *p, idx->t);
  RM_CUDA(cudaMemsetAsync(p->tag, 0, sizeof(u64) * ((size_t)p->mask + 1), s));
  RM_CUDA(cudaMemsetAsync(p->sym, 0xFF, sizeof(int) * ((size_t)p->mask + 1), s));
  RM_CUDA(cudaMemsetAsync(p->ctr + 4, 0, 2 * sizeof(u64), s));
  RM_LAUNCH(rm_pages_sweep_kernel, 148 * 8, RM_PG_TPB, s, *p);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_pages_create(int32_t device, int32_t page_size, int64_t max_pages, int64_t max_rows, rm_pages** out) {
  if (!out || page_size < 2 || page_size > 4096 || max_pages < 1 || max_pages > (1ll << 30) || max_rows < 1) {
    rm_set_error("rm_pages_create: page_size in [2, 4096], max_pages in [1, 2^30], max_rows >= 1");
    return RM_E_INVALID;
  }
  RM_CUDA(cudaSetDevice(device));
  rm_pages* p = new rm_pages();
  memset(p, 0, sizeof(*p));
  p->device = device;
  p->P = page_size;
  p->max_pages = max_pages;
  p->max_rows = max_rows;
  size_t cap = 16;
  while (cap < 2 * (size_t)max_pages) cap <<= 1;
  p->mask = (unsigned)(cap - 1);
  cudaError_t e = cudaSuccess;
#define A(ptr, T, n) \
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&(ptr)), sizeof(T) * (size_t)(n))
  A(p->tag, u64, cap);
  A(p->sym, int, cap);
  A(p->store, int, (size_t)max_pages * page_size);
  A(p->vrows, i64, (size_t)max_rows * page_size);
  A(p->rows_free, i64, max_rows);
  A(p->ctr, u64, 8);
  A(p->sym_free, int, max_pages);
  A(p->live, unsigned char, max_pages);
#undef A
  if (e != cudaSuccess) {
    rm_set_error("rm_pages_create: device allocation failed: %s", cudaGetErrorString(e));
    rm_pages_destroy(p);
    return RM_E_NOSPACE;
  }
  int rc = rm_pages_reset(p, nullptr);
  if (rc == RM_OK && cudaStreamSynchronize(0) != cudaSuccess) rc = RM_E_CUDA;
  if (rc != RM_OK) {
    rm_pages_destroy(p);
    return rc;
  }
  *out = p;
  return RM_OK;
}

int rm_pages_destroy(rm_pages* p) {
  if (!p) return RM_OK;
  cudaSetDevice(p->device);
  void* ptrs[] = {p->tag, p->sym, p->store, p->vrows, p->rows_free, p->ctr, p->sym_free, p->live};
  for (void* q : ptrs)
    if (q) cudaFree(q);
  delete p;
  return RM_OK;
}

int rm_pages_reset(rm_pages* p, void* stream) {
  if (!p) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  RM_CUDA(cudaMemsetAsync(p->tag, 0, sizeof(u64) * ((size_t)p->mask + 1), s));
  RM_CUDA(cudaMemsetAsync(p->sym, 0xFF, sizeof(int) * ((size_t)p->mask + 1), s));
  int grid = rm_pg_grid(p->max_rows);
  if (grid > 1184) grid = 1184;
  RM_LAUNCH(rm_pages_reset_kernel, grid, RM_PG_TPB, s, *p);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_pages_intern(rm_pages* p, const int32_t* tokens, const int64_t* offsets, int32_t B, int32_t create,
                    const int64_t* sym_offsets, int64_t total_pages, int32_t* out_syms, int32_t* out_npages, int32_t* scratch,
                    void* stream) {
  if (!p || B < 0 || total_pages < 0 || (B > 0 && (!tokens || !offsets || !sym_offsets || !out_syms || !scratch))) return RM_E_INVALID;
  if (B == 0 || total_pages == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = rm_pg_grid(total_pages);
  RM_LAUNCH(rm_pages_claim_kernel, grid, RM_PG_TPB, s, *p, tokens, (const i64*)offsets, B, create ? 1 : 0, (const i64*)sym_offsets,
            out_syms, scratch, (int)total_pages);
  if (create)
    RM_LAUNCH(rm_pages_resolve_kernel, grid, RM_PG_TPB, s, *p, tokens, (const i64*)offsets, B, (const i64*)sym_offsets, out_syms,
              scratch, (int)total_pages);
  RM_LAUNCH(rm_pages_truncate_kernel, rm_pg_grid(B), RM_PG_TPB, s, B, (const i64*)sym_offsets, out_syms, out_npages);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_pages_rows_put(rm_pages* p, const int64_t* values, int64_t n_rows, int64_t* out_rows, void* stream) {
  if (!p || n_rows < 0 || (n_rows > 0 && (!values || !out_rows))) return RM_E_INVALID;
  if (n_rows == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  RM_LAUNCH(rm_pages_rows_pop_kernel, 1, 1, s, *p, (i64)n_rows, p->ctr + 4);      // ring position of the rows, kept on the device
  RM_LAUNCH(rm_pages_rows_put_kernel, rm_pg_grid(n_rows * p->P), RM_PG_TPB, s, *p, (const i64*)values, (i64)n_rows, p->ctr + 4,
            (i64*)out_rows);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_pages_rows_free(rm_pages* p, const int64_t* rows, int64_t n, void* stream) {
  if (!p || n < 0 || (n > 0 && !rows)) return RM_E_INVALID;
  if (n == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(p->device));
  RM_LAUNCH(rm_pages_rows_free_kernel, rm_pg_grid(n), RM_PG_TPB, (cudaStream_t)stream, *p, (const i64*)rows, (i64)n);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_pages_rows_get(rm_pages* p, const int64_t* rows, int64_t n, int64_t* out_values, void* stream) {
  if (!p || n < 0 || (n > 0 && (!rows || !out_values))) return RM_E_INVALID;
  if (n == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(p->device));
  RM_LAUNCH(rm_pages_rows_get_kernel, rm_pg_grid(n * p->P), RM_PG_TPB, (cudaStream_t)stream, *p, (const i64*)rows, (i64)n,
            (i64*)out_values);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_pages_tokens(rm_pages* p, const int32_t* syms, int64_t n, int32_t* out_tokens, void* stream) {
  if (!p || n < 0 || (n > 0 && (!syms || !out_tokens))) return RM_E_INVALID;
  if (n == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(p->device));
  RM_LAUNCH(rm_pages_tokens_kernel, rm_pg_grid(n * p->P), RM_PG_TPB, (cudaStream_t)stream, *p, syms, (i64)n, out_tokens);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_pages_status(rm_pages* p, int64_t* out4_h, void* stream) {
  if (!p || !out4_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  u64 c[6];
  RM_CUDA(cudaMemcpyAsync(c, p->ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  const u64 taken = c[4] < c[5] ? c[4] : c[5];
  out4_h[0] = (int64_t)(c[0] - (c[5] - taken));   // symbols in use: handed out so far minus what the free list still holds
  out4_h[1] = (int64_t)(c[2] - c[1]);        // free rows
  out4_h[2] = (int64_t)c[3];                 // sticky device error
  out4_h[3] = p->P;
  return RM_OK;
}

}  // extern "C"
