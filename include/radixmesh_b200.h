/*
 * radixmesh_b200 — C ABI of the B200-native radix-tree KV-cache index.
 *
 * The reference (zzt93/RadixMesh) has no FFI: its boundary for this path is the Python
 * class surface RadixCache / RadixMesh / CacheAwareRouter.  Each entry point below names
 * the reference method(s) it replaces (paths relative to /root/reference/python/src).
 * INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - every function returns an rm_status (0 = RM_OK); nothing throws across the ABI;
 *   - pointers are DEVICE pointers unless the parameter name ends in _h (HOST pointer);
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  All
 *     calls on one index must be issued on ONE stream: stream order = operation order =
 *     last-access stamp order.  Calls enqueue work and return without synchronising unless
 *     the comment says "synchronises";
 *   - the caller owns every input / output buffer; the index owns the node pool, the edge
 *     arenas, the child table and its scratch;
 *   - node handles are int32 node ids; id 0 is the root.  A node's identity for parity is
 *     its full key from the root, never its id;
 *   - one host thread at a time per index (the Python shim holds a mutex per index).
 */
#ifndef RADIXMESH_B200_H
#define RADIXMESH_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rm_index rm_index;
typedef struct rm_node_info_t rm_node_info_t;   /* defined with rm_node_info below */

typedef enum {
  RM_OK = 0,
  RM_E_INVALID = 1,     /* bad argument */
  RM_E_CUDA = 2,        /* a CUDA runtime call failed; rm_last_error() has the text */
  RM_E_NOSPACE = 3,     /* node pool / edge arena / scratch exhausted (the reference cannot produce this) */
  RM_E_TOO_LARGE = 4,   /* batch larger than rm_config.max_batch */
  RM_E_MODE = 5,        /* operation not allowed in this mode (e.g. insert on a ROUTER index,
                           radix/radix_mesh.py:194) */
  RM_E_TABLE = 6        /* a child-table probe visited every slot: no empty slot is left (tombstones of evicted
                           nodes took them).  rm_evict rebuilds the table long before that; rm_compact does too.
                           Raised on the device (sticky, rm_sizes.device_error) instead of spinning. */
} rm_status;

/* Tree semantics selector.  BASE = RadixCache (radix/sglang/srt/mem_cache/radix_cache.py:87);
 * the other three are RadixMesh's modes (radix/core_enum.py:4-7, radix/radix_mesh.py:76-84). */
typedef enum { RM_MODE_BASE = 0, RM_MODE_PREFILL = 1, RM_MODE_DECODE = 2, RM_MODE_ROUTER = 3 } rm_mode;

typedef struct {
  int32_t struct_size;   /* sizeof(rm_config), for forward compatibility */
  int32_t device;        /* CUDA device ordinal that hosts this mesh node's tree */
  int32_t mode;          /* rm_mode */
  int32_t self_rank;     /* global node rank of this replica (config/cache_config.py:50-75) */
  int32_t n_prefill;     /* len(prefill_cache_nodes): ranks [0, P) are prefill */
  int32_t n_decode;      /* len(decode_cache_nodes):  ranks [P, P+D) are decode */
  int32_t max_batch;     /* largest B / N any *_batch call will pass (scratch is sized from it) */
  int32_t max_dups;      /* capacity of the duplicate-node table (RadixMesh.dup_nodes) */
  int32_t max_key_len;   /* longest key (tokens) any insert / query will pass; bounds edge lengths */
  int32_t max_batch_tokens; /* token budget of one *_batch call (sum of key lengths), 0 = unknown.  When given, the
                               device staging of the *_h entry points and the consistent-hash message scratch are
                               sized from it once, at rm_create / rm_route_set_rings; with 0 the staging grows on
                               demand (a cudaFree + cudaMalloc on the calling path the first time a larger batch
                               arrives) and the hash scratch is capped at 1 GiB (RM_ROUTE_SCRATCH_MB) */
  int64_t max_nodes;     /* node pool capacity, including the root */
  int64_t max_tokens;    /* edge arena capacity in tokens (keys i32 + values i64 per token) */
} rm_config;

/* Per-index counters, filled by rm_sizes (synchronises). */
typedef struct {
  int64_t evictable;     /* RadixCache.evictable_size()  radix_cache.py:232-233 */
  int64_t protected_;    /* RadixCache.protected_size()  radix_cache.py:235-237 */
  int64_t n_nodes;       /* live nodes, root excluded */
  int64_t arena_used;    /* tokens handed out of the edge arena (monotonic until rm_compact) */
  int64_t n_dups;        /* live entries in the duplicate-node table */
  int64_t op_seq;        /* operations applied so far (the logical clock) */
  int32_t device_error;  /* sticky rm_status raised inside a kernel (RM_E_NOSPACE, RM_E_TABLE) */
  int32_t node_hwm;      /* high-water node id + 1 */
  int32_t table_slots;   /* child-table capacity (a power of two >= 2 * max_nodes) */
  int32_t table_tombs;   /* of those, tombstones left by evicted nodes (rm_evict rebuilds the table at a quarter) */
} rm_sizes_t;

const char* rm_last_error(void);
const char* rm_version(void);

/* RadixCache.__init__ / RadixMesh.__init__ tree part (radix_cache.py:89-115, radix_mesh.py:74-89). */
int rm_create(const rm_config* cfg, rm_index** out);
int rm_destroy(rm_index* idx);
/* RadixCache.reset / RadixMesh.reset (radix_cache.py:118-125, radix_mesh.py:240-245). */
int rm_reset(rm_index* idx, void* stream);

/*
 * Batched longest-prefix match: RadixCache.match_prefix + _match_prefix_helper + _split_node
 * (radix_cache.py:127-162, 252-294, 14-20) and RadixMesh.match_prefix/_match_prefix_helper
 * (radix_mesh.py:203-218, 247-271).  Equivalent to calling match_prefix on queries 0..B-1 in
 * order.  Query q is tokens[offsets[q] .. offsets[q+1]).  BASE/PREFILL split a partially
 * matched edge and return the new upper node; DECODE/ROUTER never split and return the
 * partially matched child.  Zero-length queries return (0, root) (the mesh shim raises
 * IndexError before calling, radix_mesh.py:253).
 *   out_len[q]   matched prefix length (= len(device_indices))
 *   out_node[q]  last_device_node
 */
int rm_match_batch(rm_index* idx, const int32_t* tokens, const int64_t* offsets, int32_t B,
                   int32_t* out_len, int32_t* out_node, void* stream);
/* Same, HOST buffers: H2D of tokens/offsets, match, D2H of both outputs; synchronises.  Pinned
 * host memory makes the copies asynchronous with respect to each other. */
int rm_match_batch_h(rm_index* idx, const int32_t* tokens_h, const int64_t* offsets_h, int32_t B,
                     int32_t* out_len_h, int32_t* out_node_h, void* stream);

/*
 * RadixCache.match_prefix / RadixMesh.match_prefix (P/D) for ONE key in HOST memory, end to end with one synchronisation:
 * H2D of the key, walk (+ split), gather of the values along the matched path (radix_cache.py:154-157), the record of
 * last_device_node (rm_node_info), D2H.  out_values_h must hold n entries (the first *out_len_h are written);
 * out_info_h may be NULL.  The single-key surface of the drop-in classes goes through this call: ~10 separate copies and
 * launches otherwise.
 */
int rm_match_prefix_h(rm_index* idx, const int32_t* key_h, int32_t n, int32_t* out_len_h, int32_t* out_node_h,
                      int64_t* out_values_h, rm_node_info_t* out_info_h, void* stream);

/*
 * rm_match_batch for callers that must not allocate device memory.  No counterpart in the reference
 * (a Python dict tree has no device scratch).  Same arguments and results as rm_match_batch, but the
 * scratch of the match pipeline (split-request arrays, chain links, the lists of its sorted path) lives in a fixed
 * 256 KiB array inside the library image instead of memory sized by rm_config.max_batch, so an index
 * created with max_batch = 1 can still serve batches here.  For engines that gave all of HBM to their
 * KV pool (a later cudaMalloc can fail) and for request-path threads that cannot take cudaMalloc's
 * device-wide synchronisation.
 *   - at most 4096 queries per call, else RM_E_TOO_LARGE;
 *   - makes no cudaMalloc / cudaFree / allocating call;
 *   - the first call on each device looks the array's address up (cudaGetSymbolAddress): make it
 *     before any stream capture starts;
 *   - every GPU that loads the library holds the 256 KiB until exit; all calls on a device share
 *     it, so calls that overlap on the GPU race: keep them on one stream or order them.
 */
int rm_match_batch_static(rm_index* idx, const int32_t* tokens, const int64_t* offsets, int32_t B,
                          int32_t* out_len, int32_t* out_node, void* stream);

/*
 * torch.cat of the values along the matched path (radix_cache.py:154-157, radix_mesh.py:209-213):
 * for query q writes out_len[q] int64 KV indices to out_values[out_offsets[q] ...].
 * `nodes`/`lens` are rm_match_batch's outputs (call before any later mutation of the tree).
 */
int rm_gather_values(rm_index* idx, const int32_t* nodes, const int32_t* lens, const int64_t* out_offsets,
                     int32_t B, int64_t* out_values, void* stream);

/*
 * Ordered batch of inserts: RadixCache.insert/_insert_helper (radix_cache.py:164-170, 296-327);
 * in the mesh modes RadixMesh._insert_helper + NodeRankConflictResolver.keep + _swap_node
 * (radix_mesh.py:273-323, 466-495, policy/conflict_resolve.py:1-6), i.e. what
 * oplog_received does with N INSERT oplogs in arrival order (radix_mesh.py:411-416).
 * Record r: key = tokens[offsets[r] .. offsets[r+1]), values[same range] (ignored for ROUTER),
 * owner rank = ranks[r] (ignored for BASE; NULL = self_rank for every record).
 * Equivalent to applying records 0..N-1 in order.  out_prefix_len[r] = the reference's return
 * value (matched prefix length at the time record r is applied).
 */
int rm_insert_batch(rm_index* idx, const int32_t* tokens, const int64_t* values, const int64_t* offsets,
                    const int32_t* ranks, int32_t N, int32_t* out_prefix_len, void* stream);
int rm_insert_batch_h(rm_index* idx, const int32_t* tokens_h, const int64_t* values_h, const int64_t* offsets_h,
                      const int32_t* ranks_h, int32_t N, int32_t* out_prefix_len_h, void* stream);

/*
 * inc_lock_ref / dec_lock_ref (radix_cache.py:204-230; mesh override radix_mesh.py:168-179) for
 * B nodes, all with the same sign (dir = +1 or -1), equivalent to the calls in order.
 * out_delta[i] = the reference's return value for call i.
 * gens (may be NULL): the generation each handle was taken at (rm_dump's gen array at that time).
 * A tree id whose object was parked by a locked _swap_node (radix_mesh.py:478-495) after the handle
 * was taken resolves to the parked node, as the reference's TreeNode object would.
 */
int rm_lock_batch(rm_index* idx, const int32_t* nodes, const int32_t* gens, int32_t B, int32_t dir,
                  int64_t* out_delta, void* stream);

/*
 * RadixCache.evict(num_tokens) (radix_cache.py:179-202 with _collect_leaves :366-377 and
 * _delete_leaf :347-352).  Evicted nodes are reported in the reference's eviction order:
 * out_nodes_h[k] = node id, out_offs_h[k] / out_lens_h[k] = its value run in the value arena
 * (rm_read_values reads it) so the caller can call token_to_kv_pool_allocator.free exactly as
 * radix_cache.py:195 does.  *out_n_h = number evicted (<= cap).  Synchronises ONCE in the usual call: the host asks
 * the device nothing before the work (the node high-water mark is bounded from its last reading, the candidate count
 * stays on the device); when a bound turns out too small the pass evicts nothing and the call repeats it with exact
 * sizes.  Also rebuilds the child table when a quarter of its slots are tombstones.
 */
int rm_evict(rm_index* idx, int64_t num_tokens, int32_t cap, int32_t* out_nodes_h, int64_t* out_offs_h,
             int32_t* out_lens_h, int32_t* out_n_h, void* stream);

/*
 * Maintenance (no counterpart in the reference, whose Python objects are garbage collected): slide all
 * live key/value runs to the front of fresh edge arenas and rebuild the child table, reclaiming the
 * space of evicted nodes and dropping tombstones.  Node ids and everything observable stay the same.
 * Allocates the new arenas, frees the old ones, synchronises.
 */
int rm_compact(rm_index* idx, void* stream);

/* RadixMesh._gc candidate scan (radix_mesh.py:151-156): entries of dup_nodes whose parked node is
 * not None and has lock_ref == 0.  out_h rows are (tree node id, rank, parked node id).  Synchronises. */
int rm_gc_candidates(rm_index* idx, int32_t cap, int32_t* out_h, int32_t* out_n_h, void* stream);
/* All dup_nodes entries (tree node id, rank, parked node id or -1, parked lock_ref or 0).  Synchronises. */
int rm_dup_entries(rm_index* idx, int32_t cap, int32_t* out_h, int32_t* out_n_h, void* stream);
/*
 * Duplicate-node GC between replicas: RadixMesh._gc_handle (radix_mesh.py:362-389).  A key crosses replicas as
 * (full key tokens, rank) = ImmutableNodeKey (cache_oplog.py:25-40); every call here takes HOST pointers, locates
 * the keys read-only and synchronises (control plane: a handful of keys every gc_interval_second).
 *   rm_gc_keys_h   full keys of tree nodes (the node ids rm_gc_candidates returns): row i of out_tokens_h
 *                  (stride tokens per row) and out_lens_h[i]; a key longer than stride only reports its length.
 *   rm_gc_agree_h  out_agree_h[i] = 1 iff the key is in dup_nodes here and its entry is None or has
 *                  lock_ref == 0 (:386-387), else 0 (also when the key's node does not exist here).
 *   rm_gc_exec_h   GC_EXEC: dup_nodes.pop(key, None) (:363-366).  With free_parked != 0 (the replica that asked,
 *                  :371-377) a parked node held by the entry is released and its arena run returned in
 *                  out_off_h[i] / out_len_h[i] (else -1 / 0): read the KV indices with rm_read_arena and hand them
 *                  to token_to_kv_pool_allocator.free, as :376-377 does.
 * The all-reduce that replaces the GC_QUERY ring pass on one box is host code (radixmesh_b200/mesh_gc.py); the
 * reference's own ring pass cannot complete because CacheOplog.to_dict drops gc_query / gc_exec.
 */
int rm_gc_keys_h(rm_index* idx, const int32_t* nodes_h, int32_t n, int32_t stride, int32_t* out_tokens_h,
                 int32_t* out_lens_h, void* stream);
int rm_gc_agree_h(rm_index* idx, const int32_t* tokens_h, const int64_t* offsets_h, const int32_t* ranks_h, int32_t n,
                  uint8_t* out_agree_h, void* stream);
int rm_gc_exec_h(rm_index* idx, const int32_t* tokens_h, const int64_t* offsets_h, const int32_t* ranks_h, int32_t n,
                 int32_t free_parked, int64_t* out_off_h, int32_t* out_len_h, void* stream);

/*
 * Router pick: RadixMesh.match_prefix in ROUTER mode (radix_mesh.py:219-238) fused with
 * CacheAwareRouter.cache_aware_route (router/cache_aware_router.py:23-39) and the ConsistentHash
 * fallback (:42-118; ring built once by rm_route_set_rings).
 *   out_prefill[q] / out_decode[q]  ranks as RouterMatchResult holds them (-1 = none)
 *   out_paddr[q] / out_daddr[q]     index into prefill_cache_nodes / decode_cache_nodes after the
 *                                   fallback (the strings stay on the host)
 *   out_err[q]   0 ok, 1 = the reference raises AssertionError (radix_mesh.py:233-237),
 *                2 = the reference raises IndexError (radix_mesh.py:451)
 * warm = 0 reproduces the warm-up gate (cache_aware_router.py:24-25): no match, hash only.
 */
int rm_route_batch(rm_index* idx, const int32_t* tokens, const int64_t* offsets, int32_t B, int32_t warm,
                   int32_t* out_prefill, int32_t* out_decode, int32_t* out_paddr, int32_t* out_daddr,
                   uint8_t* out_err, void* stream);
/* Hash rings for the fallback: sorted u32 points and the address index that owns each point
 * (ConsistentHash.add_node, cache_aware_router.py:58-73).  HOST pointers; copied. */
int rm_route_set_rings(rm_index* idx, const uint32_t* p_points_h, const int32_t* p_owner_h, int32_t n_p,
                       const uint32_t* d_points_h, const int32_t* d_owner_h, int32_t n_d);

/*
 * Mesh routing across the GPUs of one box (north_star item 4; no single counterpart in the reference,
 * whose router reads owner ranks from its own replicated tree, radix_mesh.py:219-238).  GPU g hosts the
 * tree of mesh node g holding the keys rank g inserted; after an all-gather of the per-GPU matched
 * lengths, depths[g*B + q], this picks prefill = lowest prefill rank with the deepest match (> 0),
 * decode = lowest decode rank with the deepest match if strictly deeper than the prefill match, then
 * applies the same address lookup / consistent-hash fallback as rm_route_batch.  `idx` supplies
 * n_prefill / n_decode and the rings.
 */
int rm_pick_batch(rm_index* idx, const int32_t* depths, int32_t n_ranks, const int32_t* tokens,
                  const int64_t* offsets, int32_t B, int32_t* out_prefill, int32_t* out_decode,
                  int32_t* out_paddr, int32_t* out_daddr, uint8_t* out_err, void* stream);

/* evictable_size / protected_size / node count (radix_cache.py:232-237).  Synchronises. */
int rm_sizes(rm_index* idx, rm_sizes_t* out_h, void* stream);

/*
 * Parity / TreeNode-proxy support.  rm_dump copies the per-node arrays of ids [0, n) to HOST
 * (any pointer may be NULL): parent (-1 = free id; the root also has -1), edge offset, edge length,
 * lock_ref, last-access stamp (op sequence number), owner rank, full-key length, child count,
 * generation (< 0 = parked by a locked swap), parked-chain link.
 * *out_n_h = high-water node id + 1.  Synchronises.
 */
int rm_dump(rm_index* idx, int32_t cap, int32_t* parent_h, int64_t* edge_off_h, int32_t* edge_len_h,
            int32_t* lock_ref_h, uint64_t* stamp_h, int32_t* rank_h, int32_t* depth_h, int32_t* nchild_h,
            int32_t* gen_h, int32_t* parked_h, int32_t* out_n_h, void* stream);
/*
 * TreeNode attribute reads for the handles a caller holds (radix_cache.py:35-64; test/test_util.py:10-13 reads
 * .id .key .value .lock_ref .parent .children): rows for n node ids given in a HOST array, independent of the tree size.
 * parent = -1 for the root and for a free id, -2 for an id outside the pool.  Synchronises.
 */
struct rm_node_info_t {
  int32_t parent, edge_len, lock_ref, rank, depth_end, nchild, gen, parked;
  int64_t edge_off;      /* start of the node's key / value run in the arenas (rm_read_arena) */
  uint64_t stamp;        /* raw last-access stamp of the node itself (its effective stamp is the subtree maximum) */
};
int rm_node_info(rm_index* idx, const int32_t* ids_h, int32_t n, rm_node_info_t* out_h, void* stream);
/* TreeNode.children of one node (radix_cache.py:39): child id, first token of its edge (the dict key), generation and
 * dict-order key (ascending = the reference's dict insertion order) per child, to HOST arrays of `cap` entries;
 * *out_n_h = number of children.  One pass over the node pool on the device, a few bytes back.  Synchronises. */
int rm_node_children(rm_index* idx, int32_t node, int32_t cap, int32_t* ids_h, int32_t* first_tokens_h, int32_t* gens_h,
                     uint64_t* order_h, int32_t* out_n_h, void* stream);
/* RadixCache.total_size (radix_cache.py:239-241, 354-364): sum of len(value) over the tree, reduced on the device.
 * Synchronises. */
int rm_total_size(rm_index* idx, int64_t* out_h, void* stream);
/* Dict-order keys of node ids [0, cap) to HOST (with rm_dump: all_values_flatten in the reference's pre-order,
 * children in dict insertion order, radix_cache.py:239-248).  Synchronises. */
int rm_dump_order(rm_index* idx, int32_t cap, uint64_t* order_h, void* stream);

/* Read `n` tokens / values of the edge arenas starting at token offset `off` into HOST buffers. */
int rm_read_arena(rm_index* idx, int64_t off, int64_t n, int32_t* tokens_h, int64_t* values_h, void* stream);
/* The VALUE runs [offs_h[i], offs_h[i] + lens_h[i]) of the edge arena, i < n (n <= rm_config.max_nodes), one after
 * another into out_values_h (sum of lens_h entries): what RadixCache.evict hands to token_to_kv_pool_allocator.free
 * node by node (radix_cache.py:195), for all evicted nodes of an rm_evict call in one gather and one copy instead of a
 * synchronising read per node.  Synchronises. */
int rm_read_runs(rm_index* idx, const int64_t* offs_h, const int32_t* lens_h, int32_t n, int64_t* out_values_h, void* stream);

/* Parity aid: the per-position owner map of the edge arena (node id, bit 30 = contiguous child of the node
 * owning the previous position, -1 = free) for `n` positions from `off`, to HOST.  Synchronises. */
int rm_read_segments(rm_index* idx, int64_t off, int64_t n, int32_t* seg_h, void* stream);

/*
 * Paged keys: RadixCache(page_size > 1) (radix_cache.py:23-32 _key_match_paged, :109-114 child key = first page,
 * :148-150 page-aligned match).  A paged radix tree over tokens is a radix tree over page symbols, so an rm_index in
 * BASE mode runs unchanged on symbols; an rm_pages object beside it turns tokens into symbols on the device and keeps
 * the KV indices of a page as one row (the index's value of a symbol position is the row id).  All pointers are
 * DEVICE pointers; keys are taken in whole pages (len / page_size).  A key with a short last page (the reference stores
 * it in the leaf, radix_cache.py:296-326) is passed with that page filled up by a token no key contains: it then is a
 * page of its own, equal only to the same short page, and the caller remembers how many tokens its value row really
 * holds (radixmesh_b200/radix_cache.py does exactly this).
 *   rm_pages_intern    key q = tokens[offsets[q] .. offsets[q+1]); its symbols go to out_syms[sym_offsets[q] ..
 *                      sym_offsets[q+1]) (sym_offsets[q+1] - sym_offsets[q] = whole pages of key q; total_pages =
 *                      sym_offsets[B]).  create != 0 (insert): unseen pages get new symbols.  create == 0 (match): the
 *                      first unseen page of a key and everything after it become a symbol no edge holds, so the walk
 *                      ends there as the reference's dict lookup would; out_npages[q] (may be NULL) = pages resolved.
 *                      scratch: total_pages int32.  Every table hit is verified against the stored page tokens.
 *   rm_pages_rows_put  n_rows rows of page_size KV indices each (values, row-major) -> out_rows[n_rows] row ids: the
 *                      `values` array of rm_insert_batch on the symbols.
 *   rm_pages_rows_free hand rows back: those of evicted runs, and those of the pages an insert found already present
 *                      (the first out_prefix_len[r] rows of record r).
 *   rm_pages_rows_get  row ids (rm_gather_values / rm_read_arena output) -> KV indices, page_size per row.
 *   rm_pages_tokens    symbols -> page tokens (TreeNode.key).
 *   rm_pages_retain    drop the pages that no edge of `idx` (the BASE index running on this object's symbols) holds
 *                      any more - evicted since they were interned - so that their symbols are handed out again; the
 *                      table is rebuilt from the live pages.  Call it between operations (every symbol interned with
 *                      create != 0 must have reached the tree through rm_insert_batch, or it is dropped).  The
 *                      reference's dict keys die with their nodes (radix_cache.py:347-352); no other counterpart.
 *   rm_pages_status    HOST out: symbols in use, free rows, sticky device error (RM_E_NOSPACE: table or rows exhausted;
 *                      RM_E_INVALID: two different new pages of one batch with the same 64-bit hash), page_size.
 *                      Synchronises.
 */
typedef struct rm_pages rm_pages;
int rm_pages_create(int32_t device, int32_t page_size, int64_t max_pages, int64_t max_rows, rm_pages** out);
int rm_pages_destroy(rm_pages* p);
int rm_pages_reset(rm_pages* p, void* stream);
int rm_pages_intern(rm_pages* p, const int32_t* tokens, const int64_t* offsets, int32_t B, int32_t create,
                    const int64_t* sym_offsets, int64_t total_pages, int32_t* out_syms, int32_t* out_npages, int32_t* scratch,
                    void* stream);
int rm_pages_rows_put(rm_pages* p, const int64_t* values, int64_t n_rows, int64_t* out_rows, void* stream);
int rm_pages_rows_free(rm_pages* p, const int64_t* rows, int64_t n, void* stream);
int rm_pages_rows_get(rm_pages* p, const int64_t* rows, int64_t n, int64_t* out_values, void* stream);
int rm_pages_tokens(rm_pages* p, const int32_t* syms, int64_t n, int32_t* out_tokens, void* stream);
int rm_pages_retain(rm_pages* p, rm_index* idx, void* stream);
int rm_pages_status(rm_pages* p, int64_t* out4_h, void* stream);

/*
 * Measurement aids (no counterpart in the reference, which has no instrumentation).
 * rm_launch_count: kernels this library has launched in this process.
 * rm_profile(idx, 1): record a CUDA-event pair around the walk kernel (K1) of every following
 * rm_match_batch; rm_profile_read returns the summed device time of those kernels and their count,
 * and clears the record (at most 128 batches are kept).  Synchronises.
 * rm_profile(idx, 2) + rm_profile_stages (call it before rm_profile_read, which clears): the same sums for the
 * stages of a splitting match batch (one more event per batch): stage_ms_h[0] walk (K1), [1] split stage (K3);
 * [2] and [3] are reserved and read 0.
 */
unsigned long long rm_launch_count(void);
int rm_profile(rm_index* idx, int enable);
int rm_profile_read(rm_index* idx, double* walk_ms_h, int* n_h);
int rm_profile_stages(rm_index* idx, double* stage_ms_h, int* n_h);

#ifdef __cplusplus
}
#endif
#endif /* RADIXMESH_B200_H */
