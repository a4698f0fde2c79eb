// C ABI (include/radixmesh_b200.h): index lifetime, argument checks, host-buffer variants.
#include <stdarg.h>
#include <string.h>

#include "rm_internal.h"
#include "rm_sort.cuh"

#ifdef RM_EMU
thread_local rm_emu_dim3 threadIdx, blockIdx, blockDim, gridDim;
#endif

#ifndef RM_EMU
unsigned long long g_rm_launches = 0;
#endif
static thread_local char g_err[512] = "";

void rm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int rm_bits(u64 v) {   // bits needed to hold values 0..v
  int b = 1;
  while (b < 63 && (v >> b) != 0) ++b;
  return b;
}

template <class T>
static cudaError_t rm_dalloc(T** p, size_t n) {
  return cudaMalloc(reinterpret_cast<void**>(p), sizeof(T) * (n ? n : 1));
}

#ifndef RM_EMU
static __global__ void rm_policy_kernel(u64* out) { *out = rm_policy_evict_first(); }
#endif

extern "C" {

const char* rm_last_error(void) { return g_err; }
const char* rm_version(void) {
#ifdef RM_EMU
  return "radixmesh_b200 0.1 (host emulation, tests only)";
#else
  return "radixmesh_b200 0.1 (sm_100a)";
#endif
}

unsigned long long rm_launch_count(void) {
#ifdef RM_EMU
  return 0;
#else
  return g_rm_launches;
#endif
}

int rm_profile(rm_index* idx, int enable) {
  if (!idx) return RM_E_INVALID;
#ifndef RM_EMU
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  if (enable && !idx->prof_ev[0]) {
    for (int i = 0; i < RM_PROF_EVENTS; ++i) RM_CUDA(cudaEventCreate(&idx->prof_ev[i]));
  }
  idx->prof_on = enable < 0 ? 0 : enable > 2 ? 2 : enable;   // 1: walk kernel only, 2: walk and split stage
  idx->prof_n = 0;
#endif
  return RM_OK;
}

// per batch: [0] before K1, [1] after K1, [2] after the split kernel
static int rm_profile_sum(rm_index* idx, double* stage_ms, int* n_h) {
  for (int k = 0; k < 4; ++k) stage_ms[k] = 0.0;
  *n_h = 0;
#ifndef RM_EMU
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  const int cap = RM_PROF_EVENTS / RM_PROF_STRIDE;
  const int n = idx->prof_n < cap ? idx->prof_n : cap;
  const int last = idx->prof_on > 1 ? RM_PROF_STRIDE - 1 : 1;
  for (int i = 0; i < n; ++i) {
    cudaEvent_t* e = idx->prof_ev + RM_PROF_STRIDE * i;
    RM_CUDA(cudaEventSynchronize(e[last]));
    for (int k = 0; k < last; ++k) {
      float ms = 0.f;
      RM_CUDA(cudaEventElapsedTime(&ms, e[k], e[k + 1]));
      stage_ms[k] += ms;
    }
  }
  *n_h = n;
#endif
  return RM_OK;
}

int rm_profile_stages(rm_index* idx, double* stage_ms_h, int* n_h) {
  if (!idx || !stage_ms_h || !n_h) return RM_E_INVALID;
  return rm_profile_sum(idx, stage_ms_h, n_h);
}

int rm_profile_read(rm_index* idx, double* walk_ms_h, int* n_h) {
  if (!idx || !walk_ms_h || !n_h) return RM_E_INVALID;
  double st[4];
  int rc = rm_profile_sum(idx, st, n_h);
  *walk_ms_h = st[0];
#ifndef RM_EMU
  idx->prof_n = 0;
#endif
  return rc;
}

int rm_destroy(rm_index* idx) {
  if (!idx) return RM_OK;
  cudaSetDevice(idx->cfg.device);
  RmTree& t = idx->t;
  RmScratch& sc = idx->sc;
  void* ptrs[] = {t.table, t.parent, t.edge_off, t.edge_len, t.lock_ref, t.stamp, t.rank, t.depth_end, t.nchild,
                  t.gen, t.parked, t.born, t.ord, t.ts, t.val, t.req_cnt, t.free_ids, t.ctr, t.dup_key, t.dup_parked, sc.key[0], sc.key[1],
                  sc.val[0], sc.val[1], sc.sufmax, sc.krank, sc.uniq_pos, sc.uniq_min, sc.link, t.req_head, sc.hist, sc.ev_key[0], sc.ev_key[1],
                  sc.ev_val[0], sc.ev_val[1], sc.ev_hist, sc.ev_lens, sc.ev_eff, sc.ev_rem, sc.ev_selw, sc.ev_selc, sc.ins_mark, sc.lock_first, sc.lock_cnt, sc.lock_flag,
                  sc.dup_out, sc.dup_out_n, sc.h_tokens, sc.h_values, sc.h_offsets, sc.h_ranks, sc.h_out_a, sc.h_out_b, sc.info_ids, sc.info_rows,
                  idx->ring_p, idx->ring_p_owner, idx->ring_d, idx->ring_d_owner, idx->route_msg, idx->route_nblk};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  if (idx->mail) cudaFreeHost(idx->mail);
  delete idx;
  return RM_OK;
}

int rm_create(const rm_config* cfg, rm_index** out) {
  if (!cfg || !out || cfg->struct_size != (int32_t)sizeof(rm_config)) {
    rm_set_error("rm_create: bad config (struct_size %d, expected %d)", cfg ? cfg->struct_size : -1, (int)sizeof(rm_config));
    return RM_E_INVALID;
  }
  if (cfg->max_nodes < 2 || cfg->max_nodes > (1ll << 29) || cfg->max_tokens < 1 || cfg->max_batch < 1 ||
      cfg->max_key_len < 1 || cfg->max_key_len > RM_MAX_EDGE_LEN || cfg->max_batch_tokens < 0 || cfg->n_prefill < 0 || cfg->n_decode < 0 ||
      cfg->n_prefill + cfg->n_decode > 250 || cfg->self_rank > 253 || cfg->mode < RM_MODE_BASE || cfg->mode > RM_MODE_ROUTER) {
    rm_set_error("rm_create: max_nodes/max_tokens/max_batch/max_key_len/mode out of range");
    return RM_E_INVALID;
  }
  RM_CUDA(cudaSetDevice(cfg->device));
  rm_index* idx = new rm_index();
  memset(idx, 0, sizeof(*idx));
  idx->cfg = *cfg;
  RmTree& t = idx->t;
  RmScratch& sc = idx->sc;
  const size_t N = (size_t)cfg->max_nodes;
  size_t cap = 1;
  while (cap < 2 * N) cap <<= 1;   // load factor <= 0.5: one slot per non-root node
  size_t dcap = 16;
  while (dcap < 2 * (size_t)(cfg->max_dups > 0 ? cfg->max_dups : 1)) dcap <<= 1;
  t.table_mask = (unsigned)(cap - 1);
  {
    int lg = 0;
    while ((1ull << lg) < cap) ++lg;
    t.table_shift = (unsigned)(32 - lg);
  }
  t.dup_mask = (unsigned)(dcap - 1);
  t.max_nodes = (int)N;
  t.max_tokens = cfg->max_tokens;
  t.max_key_len = cfg->max_key_len;
  t.mode = cfg->mode;
  t.n_prefill = cfg->n_prefill;
  t.n_decode = cfg->n_decode;
  t.self_rank = cfg->self_rank;
  idx->node_bits = rm_bits((u64)N);
  idx->len_bits = rm_bits((u64)cfg->max_key_len);
  const size_t B = (size_t)cfg->max_batch;
  sc.max_batch = cfg->max_batch;
  sc.hist_words = rm_sort_scratch_words((int)B);
  const bool with_values = cfg->mode != RM_MODE_ROUTER;
  cudaError_t e = cudaSuccess;
#define A(call) \
  if (e == cudaSuccess) e = (call)
  A(rm_dalloc(&t.table, cap));
  A(rm_dalloc(&t.parent, N));
  A(rm_dalloc(&t.edge_off, N));
  A(rm_dalloc(&t.edge_len, N));
  A(rm_dalloc(&t.lock_ref, N));
  A(rm_dalloc(&t.stamp, N));
  A(rm_dalloc(&t.rank, N));
  A(rm_dalloc(&t.depth_end, N));
  A(rm_dalloc(&t.nchild, N));
  A(rm_dalloc(&t.gen, N));
  A(rm_dalloc(&t.parked, N));
  A(rm_dalloc(&t.born, N));
  A(rm_dalloc(&t.ord, N));
  A(rm_dalloc(&t.ts, (size_t)cfg->max_tokens + RM_ARENA_PAD));
  A(rm_dalloc(&t.val, with_values ? (size_t)cfg->max_tokens + RM_ARENA_PAD : 1));
  A(rm_dalloc(&t.free_ids, N));
  A(rm_dalloc(&t.ctr, 1));
  A(rm_dalloc(&t.dup_key, dcap));
  A(rm_dalloc(&t.dup_parked, dcap));
  A(rm_dalloc(&sc.key[0], B));
  A(rm_dalloc(&sc.key[1], B));
  A(rm_dalloc(&sc.val[0], B));
  A(rm_dalloc(&sc.val[1], B));
  A(rm_dalloc(&sc.sufmax, B));
  A(rm_dalloc(&sc.krank, B));
  A(rm_dalloc(&sc.uniq_pos, B));
  A(rm_dalloc(&sc.uniq_min, B));
  A(rm_dalloc(&sc.link, B));
  A(rm_dalloc(&t.req_head, N));
  A(rm_dalloc(&t.req_cnt, N));
  A(rm_dalloc(&sc.hist, sc.hist_words));
  A(cudaMemset(sc.hist, 0, sizeof(unsigned) * RM_SORT_TABLE_OFF));      // grid-barrier words of the one-launch sort
  A(rm_dalloc(&sc.ev_key[0], N));
  A(rm_dalloc(&sc.ev_key[1], N));
  A(rm_dalloc(&sc.ev_val[0], N));
  A(rm_dalloc(&sc.ev_val[1], N));
  A(rm_dalloc(&sc.ev_hist, rm_sort_scratch_words((int)N)));
  A(cudaMemset(sc.ev_hist, 0, sizeof(unsigned) * RM_SORT_TABLE_OFF));
  A(rm_dalloc(&sc.ev_lens, N));
  A(rm_dalloc(&sc.ev_eff, N));
  A(rm_dalloc(&sc.ev_rem, N));
  A(rm_dalloc(&sc.ev_selw, 4096 + 8));
  A(rm_dalloc(&sc.ev_selc, 4096));
  A(rm_dalloc(&sc.ins_mark, N));
  A(rm_dalloc(&sc.lock_first, N));
  A(rm_dalloc(&sc.lock_cnt, N));
  A(rm_dalloc(&sc.lock_flag, 1));
  A(rm_dalloc(&sc.dup_out, 4 * dcap));
  A(rm_dalloc(&sc.dup_out_n, 1));
  A(rm_dalloc(&sc.h_offsets, B + 1));
  A(rm_dalloc(&sc.h_ranks, B));
  A(rm_dalloc(&sc.h_out_a, B));
  A(rm_dalloc(&sc.h_out_b, B));
  A(rm_dalloc(&sc.info_ids, 256));
  A(rm_dalloc(&sc.info_rows, 256));
  if (cfg->max_batch_tokens > 0) {          // staging of the *_h entry points sized once, here (else it grows on demand)
    A(rm_dalloc(&sc.h_tokens, (size_t)cfg->max_batch_tokens));
    if (with_values) A(rm_dalloc(&sc.h_values, (size_t)cfg->max_batch_tokens));
    sc.h_tokens_cap = cfg->max_batch_tokens;
  }
#undef A
  if (e == cudaSuccess) e = cudaMallocHost(&idx->mail, RM_MAIL_BYTES);
  if (e != cudaSuccess) {
    rm_set_error("rm_create: device allocation failed: %s", cudaGetErrorString(e));
    rm_destroy(idx);
    return RM_E_NOSPACE;
  }
#ifndef RM_EMU
  {
    // the counters block is free until the reset below fills it: borrow its first word
    rm_policy_kernel<<<1, 1>>>(reinterpret_cast<u64*>(t.ctr));
    if (cudaMemcpy(&t.qpol, t.ctr, sizeof(u64), cudaMemcpyDeviceToHost) != cudaSuccess) {
      rm_set_error("rm_create: cache policy set-up failed: %s", cudaGetErrorString(cudaGetLastError()));
      rm_destroy(idx);
      return RM_E_CUDA;
    }
  }
#endif
  int rc = rm_launch_reset(idx, 0);
  if (rc == RM_OK && cudaStreamSynchronize(0) != cudaSuccess) rc = RM_E_CUDA;
  if (rc != RM_OK) {
    rm_destroy(idx);
    return rc;
  }
  *out = idx;
  return RM_OK;
}

int rm_reset(rm_index* idx, void* stream) {
  if (!idx) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_reset(idx, (cudaStream_t)stream);
}

static int rm_check_batch(rm_index* idx, int n, const char* who) {
  if (!idx || n < 0) {
    rm_set_error("%s: bad arguments", who);
    return RM_E_INVALID;
  }
  if (n > idx->cfg.max_batch) {
    rm_set_error("%s: batch of %d exceeds rm_config.max_batch = %d", who, n, idx->cfg.max_batch);
    return RM_E_TOO_LARGE;
  }
  return RM_OK;
}

int rm_match_batch(rm_index* idx, const int32_t* tokens, const int64_t* offsets, int32_t B, int32_t* out_len,
                   int32_t* out_node, void* stream) {
  int rc = rm_check_batch(idx, B, "rm_match_batch");
  if (rc) return rc;
  if (idx->cfg.mode == RM_MODE_ROUTER) {
    rm_set_error("rm_match_batch: a ROUTER index answers through rm_route_batch");
    return RM_E_MODE;
  }
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_match(idx, tokens, (const i64*)offsets, B, out_len, out_node, (cudaStream_t)stream);
}

// grow the device staging buffers of the *_h entry points
static int rm_stage_reserve(rm_index* idx, i64 n_tokens, bool values) {
  RmScratch& sc = idx->sc;
  if (n_tokens > sc.h_tokens_cap) {
    if (sc.h_tokens) cudaFree(sc.h_tokens);
    if (sc.h_values) cudaFree(sc.h_values);
    sc.h_tokens = nullptr;
    sc.h_values = nullptr;
    i64 cap = n_tokens + n_tokens / 4 + 1024;
    if (rm_dalloc(&sc.h_tokens, (size_t)cap) != cudaSuccess) {
      rm_set_error("staging allocation of %lld tokens failed", (long long)cap);
      return RM_E_NOSPACE;
    }
    sc.h_tokens_cap = cap;
  }
  if (values && !sc.h_values) {
    if (rm_dalloc(&sc.h_values, (size_t)sc.h_tokens_cap) != cudaSuccess) {
      rm_set_error("staging allocation failed");
      return RM_E_NOSPACE;
    }
  }
  return RM_OK;
}

int rm_match_batch_h(rm_index* idx, const int32_t* tokens_h, const int64_t* offsets_h, int32_t B, int32_t* out_len_h,
                     int32_t* out_node_h, void* stream) {
  int rc = rm_check_batch(idx, B, "rm_match_batch_h");
  if (rc) return rc;
  if (B == 0) return RM_OK;
  if (idx->cfg.mode == RM_MODE_ROUTER) return RM_E_MODE;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RmScratch& sc = idx->sc;
  const i64 nt = offsets_h[B] - offsets_h[0];
  if (offsets_h[0] != 0) {
    rm_set_error("rm_match_batch_h: offsets must start at 0");
    return RM_E_INVALID;
  }
  rc = rm_stage_reserve(idx, nt, false);
  if (rc) return rc;
  RM_CUDA(cudaMemcpyAsync(sc.h_offsets, offsets_h, sizeof(i64) * ((size_t)B + 1), cudaMemcpyHostToDevice, s));
  RM_CUDA(cudaMemcpyAsync(sc.h_tokens, tokens_h, sizeof(int) * (size_t)nt, cudaMemcpyHostToDevice, s));
  rc = rm_launch_match(idx, sc.h_tokens, sc.h_offsets, B, sc.h_out_a, sc.h_out_b, s);
  if (rc) return rc;
  RM_CUDA(cudaMemcpyAsync(out_len_h, sc.h_out_a, sizeof(int) * (size_t)B, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaMemcpyAsync(out_node_h, sc.h_out_b, sizeof(int) * (size_t)B, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

int rm_match_prefix_h(rm_index* idx, const int32_t* key_h, int32_t n, int32_t* out_len_h, int32_t* out_node_h,
                      int64_t* out_values_h, rm_node_info_t* out_info_h, void* stream) {
  if (!idx || n < 0 || !out_len_h || !out_node_h || (n > 0 && (!key_h || !out_values_h))) return RM_E_INVALID;
  if (idx->cfg.mode == RM_MODE_ROUTER) return RM_E_MODE;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RmScratch& sc = idx->sc;
  int rc = rm_stage_reserve(idx, n > 0 ? n : 1, true);
  if (rc) return rc;
  // Through the pinned mailbox when the key fits (8192 tokens): every copy is then asynchronous and the call waits
  // once.  (Straight from / into the caller's pageable memory each of the six copies blocks the host.)
  const bool mailed = n <= RM_EVICT_FIRST;
  char* mail = static_cast<char*>(idx->mail);
  i64* m_offs = reinterpret_cast<i64*>(mail);                               // [0, 16)
  int* m_out = reinterpret_cast<int*>(mail + 16);                           // length, node
  rm_node_info_t* m_info = reinterpret_cast<rm_node_info_t*>(mail + 64);
  int* m_key = reinterpret_cast<int*>(mail + 4096);                         // 4 n bytes
  i64* m_val = reinterpret_cast<i64*>(mail + 4096 + 4 * RM_EVICT_FIRST);    // 8 n bytes
  static_assert(4096 + 12 * RM_EVICT_FIRST <= RM_MAIL_BYTES && 64 + sizeof(rm_node_info_t) <= 4096, "mailbox layout");
  const i64 offs[2] = {0, (i64)n};
  if (mailed) {
    m_offs[0] = 0;
    m_offs[1] = (i64)n;
    if (n > 0) memcpy(m_key, key_h, sizeof(int) * (size_t)n);
  }
  RM_CUDA(cudaMemcpyAsync(sc.h_offsets, mailed ? (const void*)m_offs : (const void*)offs, sizeof(offs), cudaMemcpyHostToDevice, s));
  if (n > 0) RM_CUDA(cudaMemcpyAsync(sc.h_tokens, mailed ? (const void*)m_key : (const void*)key_h, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, s));
  rc = rm_launch_match(idx, sc.h_tokens, sc.h_offsets, 1, sc.h_out_a, sc.h_out_b, s);
  if (rc) return rc;
  // values of the matched path into the staging array (written from position h_offsets[0] = 0), node record of the last node
  rc = rm_launch_gather(idx, sc.h_out_b, sc.h_out_a, sc.h_offsets, 1, sc.h_values, s);
  if (rc) return rc;
  rc = rm_launch_node_info_dev(idx, sc.h_out_b, 1, s);
  if (rc) return rc;
  RM_CUDA(cudaMemcpyAsync(mailed ? (void*)m_out : (void*)out_len_h, sc.h_out_a, sizeof(int), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaMemcpyAsync(mailed ? (void*)(m_out + 1) : (void*)out_node_h, sc.h_out_b, sizeof(int), cudaMemcpyDeviceToHost, s));
  if (n > 0) RM_CUDA(cudaMemcpyAsync(mailed ? (void*)m_val : (void*)out_values_h, sc.h_values, sizeof(i64) * (size_t)n, cudaMemcpyDeviceToHost, s));
  if (out_info_h) RM_CUDA(cudaMemcpyAsync(mailed ? m_info : out_info_h, sc.info_rows, sizeof(rm_node_info_t), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  if (mailed) {
    *out_len_h = m_out[0];
    *out_node_h = m_out[1];
    if (m_out[0] > 0) memcpy(out_values_h, m_val, sizeof(i64) * (size_t)m_out[0]);
    if (out_info_h) *out_info_h = *m_info;
  }
  return RM_OK;
}

int rm_gather_values(rm_index* idx, const int32_t* nodes, const int32_t* lens, const int64_t* out_offsets, int32_t B,
                     int64_t* out_values, void* stream) {
  int rc = rm_check_batch(idx, B, "rm_gather_values");
  if (rc) return rc;
  if (idx->cfg.mode == RM_MODE_ROUTER) return RM_E_MODE;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_gather(idx, nodes, lens, (const i64*)out_offsets, B, (i64*)out_values, (cudaStream_t)stream);
}

int rm_insert_batch(rm_index* idx, const int32_t* tokens, const int64_t* values, const int64_t* offsets,
                    const int32_t* ranks, int32_t N, int32_t* out_prefix_len, void* stream) {
  int rc = rm_check_batch(idx, N, "rm_insert_batch");
  if (rc) return rc;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_insert(idx, tokens, (const i64*)values, (const i64*)offsets, ranks, N, out_prefix_len,
                          (cudaStream_t)stream);
}

int rm_insert_batch_h(rm_index* idx, const int32_t* tokens_h, const int64_t* values_h, const int64_t* offsets_h,
                      const int32_t* ranks_h, int32_t N, int32_t* out_prefix_len_h, void* stream) {
  int rc = rm_check_batch(idx, N, "rm_insert_batch_h");
  if (rc) return rc;
  if (N == 0) return RM_OK;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RmScratch& sc = idx->sc;
  if (offsets_h[0] != 0) {
    rm_set_error("rm_insert_batch_h: offsets must start at 0");
    return RM_E_INVALID;
  }
  const i64 nt = offsets_h[N];
  const bool with_values = idx->cfg.mode != RM_MODE_ROUTER && values_h != nullptr;
  rc = rm_stage_reserve(idx, nt, with_values);
  if (rc) return rc;
  RM_CUDA(cudaMemcpyAsync(sc.h_offsets, offsets_h, sizeof(i64) * ((size_t)N + 1), cudaMemcpyHostToDevice, s));
  RM_CUDA(cudaMemcpyAsync(sc.h_tokens, tokens_h, sizeof(int) * (size_t)nt, cudaMemcpyHostToDevice, s));
  if (with_values) RM_CUDA(cudaMemcpyAsync(sc.h_values, values_h, sizeof(i64) * (size_t)nt, cudaMemcpyHostToDevice, s));
  if (ranks_h) RM_CUDA(cudaMemcpyAsync(sc.h_ranks, ranks_h, sizeof(int) * (size_t)N, cudaMemcpyHostToDevice, s));
  idx->hint_tokens = nt;
  rc = rm_launch_insert(idx, sc.h_tokens, sc.h_values, sc.h_offsets, ranks_h ? sc.h_ranks : nullptr, N, sc.h_out_a, s);
  if (rc) return rc;
  if (out_prefix_len_h)
    RM_CUDA(cudaMemcpyAsync(out_prefix_len_h, sc.h_out_a, sizeof(int) * (size_t)N, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

int rm_lock_batch(rm_index* idx, const int32_t* nodes, const int32_t* gens, int32_t B, int32_t dir, int64_t* out_delta,
                  void* stream) {
  int rc = rm_check_batch(idx, B, "rm_lock_batch");
  if (rc) return rc;
  if (dir != 1 && dir != -1) {
    rm_set_error("rm_lock_batch: dir must be +1 or -1");
    return RM_E_INVALID;
  }
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_lock_g(idx, nodes, gens, B, dir, (i64*)out_delta, (cudaStream_t)stream);
}

int rm_evict(rm_index* idx, int64_t num_tokens, int32_t cap, int32_t* out_nodes_h, int64_t* out_offs_h,
             int32_t* out_lens_h, int32_t* out_n_h, void* stream) {
  if (!idx || !out_n_h) return RM_E_INVALID;
  if (idx->cfg.mode != RM_MODE_BASE) {
    *out_n_h = 0;   // RadixMesh.evict is a no-op (radix_mesh.py:349-351)
    return RM_OK;
  }
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_evict(idx, num_tokens, cap, out_nodes_h, (i64*)out_offs_h, out_lens_h, out_n_h, (cudaStream_t)stream);
}

int rm_gc_candidates(rm_index* idx, int32_t cap, int32_t* out_h, int32_t* out_n_h, void* stream) {
  if (!idx || !out_n_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_dups(idx, 1, cap, out_h, out_n_h, (cudaStream_t)stream);
}

int rm_dup_entries(rm_index* idx, int32_t cap, int32_t* out_h, int32_t* out_n_h, void* stream) {
  if (!idx || !out_n_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_dups(idx, 0, cap, out_h, out_n_h, (cudaStream_t)stream);
}

int rm_gc_keys_h(rm_index* idx, const int32_t* nodes_h, int32_t n, int32_t stride, int32_t* out_tokens_h,
                 int32_t* out_lens_h, void* stream) {
  if (!idx || n < 0 || stride <= 0 || (n > 0 && (!nodes_h || !out_tokens_h || !out_lens_h))) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_gc_keys(idx, nodes_h, n, stride, out_tokens_h, out_lens_h, (cudaStream_t)stream);
}

int rm_gc_agree_h(rm_index* idx, const int32_t* tokens_h, const int64_t* offsets_h, const int32_t* ranks_h, int32_t n,
                  uint8_t* out_agree_h, void* stream) {
  if (!idx || n < 0 || (n > 0 && (!tokens_h || !offsets_h || !ranks_h || !out_agree_h))) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_gc_agree(idx, tokens_h, (const i64*)offsets_h, ranks_h, n, out_agree_h, (cudaStream_t)stream);
}

int rm_gc_exec_h(rm_index* idx, const int32_t* tokens_h, const int64_t* offsets_h, const int32_t* ranks_h, int32_t n,
                 int32_t free_parked, int64_t* out_off_h, int32_t* out_len_h, void* stream) {
  if (!idx || n < 0 || (n > 0 && (!tokens_h || !offsets_h || !ranks_h || !out_off_h || !out_len_h))) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_gc_exec(idx, tokens_h, (const i64*)offsets_h, ranks_h, n, free_parked, (i64*)out_off_h, out_len_h,
                           (cudaStream_t)stream);
}

int rm_sizes(rm_index* idx, rm_sizes_t* out_h, void* stream) {
  if (!idx || !out_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RmCounters c;
  RM_CUDA(cudaMemcpyAsync(&c, idx->t.ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  out_h->evictable = c.evictable;
  out_h->protected_ = c.protected_;
  out_h->n_nodes = c.n_nodes;
  out_h->arena_used = (int64_t)c.arena_top;
  out_h->n_dups = c.n_dups;
  out_h->op_seq = (int64_t)idx->op_seq;
  out_h->device_error = c.err;
  out_h->node_hwm = c.node_hwm;
  out_h->table_slots = (int32_t)(idx->t.table_mask + 1);
  out_h->table_tombs = c.n_tomb;
  if (getenv("RM_DEBUG_COUNTERS")) fprintf(stderr, "[rm] n_req %d on_shared_edges %d groups %d max_group %d  insert records ordered %llu resumed %llu\n", c.n_req, c.n_big, c.dbg_groups, c.dbg_max_group, c.ins_ordered, c.ins_resumed);
  return RM_OK;
}

int rm_dump(rm_index* idx, int32_t cap, int32_t* parent_h, int64_t* edge_off_h, int32_t* edge_len_h, int32_t* lock_ref_h,
            uint64_t* stamp_h, int32_t* rank_h, int32_t* depth_h, int32_t* nchild_h, int32_t* gen_h, int32_t* parked_h,
            int32_t* out_n_h, void* stream) {
  if (!idx || !out_n_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RmCounters c;
  RM_CUDA(cudaMemcpyAsync(&c, idx->t.ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  int n = c.node_hwm;
  *out_n_h = n;
  if (n > cap) n = cap;
  RmTree& t = idx->t;
#define D(dst, src, T) \
  if (dst) RM_CUDA(cudaMemcpyAsync(dst, src, sizeof(T) * (size_t)n, cudaMemcpyDeviceToHost, s))
  D(parent_h, t.parent, int);
  D(edge_off_h, t.edge_off, i64);
  D(edge_len_h, t.edge_len, int);
  D(lock_ref_h, t.lock_ref, int);
  D(stamp_h, t.stamp, u64);
  D(rank_h, t.rank, int);
  D(depth_h, t.depth_end, int);
  D(nchild_h, t.nchild, int);
  D(gen_h, t.gen, int);
  D(parked_h, t.parked, int);
#undef D
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

// one component (0 = token, 1 = owner word) of `n` arena records from `off`, to HOST.  Tokens and owner words are
// interleaved on the device; the records are copied as they are and taken apart here.
static int rm_read_arena_component(rm_index* idx, i64 off, i64 n, int which, int* out_h, cudaStream_t s) {
  const i64 chunk = 1 << 20;
  int2* tmp = (int2*)malloc(sizeof(int2) * (size_t)(n < chunk ? (n > 0 ? n : 1) : chunk));
  if (!tmp) return RM_E_NOSPACE;
  for (i64 a = 0; a < n; a += chunk) {
    const i64 m = n - a < chunk ? n - a : chunk;
    cudaError_t e = cudaMemcpyAsync(tmp, idx->t.ts + off + a, sizeof(int2) * (size_t)m, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
      free(tmp);
      rm_set_error("arena read failed: %s", cudaGetErrorString(e));
      return RM_E_CUDA;
    }
    for (i64 k = 0; k < m; ++k) out_h[a + k] = which ? tmp[k].y : tmp[k].x;
  }
  free(tmp);
  return RM_OK;
}

int rm_read_segments(rm_index* idx, int64_t off, int64_t n, int32_t* seg_h, void* stream) {
  if (!idx || off < 0 || n < 0 || off + n > idx->t.max_tokens || !seg_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_read_arena_component(idx, off, n, 1, seg_h, (cudaStream_t)stream);
}

int rm_node_info(rm_index* idx, const int32_t* ids_h, int32_t n, rm_node_info_t* out_h, void* stream) {
  if (!idx || n < 0 || (n > 0 && (!ids_h || !out_h))) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_node_info(idx, ids_h, n, out_h, (cudaStream_t)stream);
}

int rm_node_children(rm_index* idx, int32_t node, int32_t cap, int32_t* ids_h, int32_t* first_tokens_h, int32_t* gens_h,
                     uint64_t* order_h, int32_t* out_n_h, void* stream) {
  if (!idx || !out_n_h || cap < 0 || (cap > 0 && (!ids_h || !first_tokens_h || !gens_h || !order_h))) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_node_children(idx, node, cap, ids_h, first_tokens_h, gens_h, (u64*)order_h, out_n_h, (cudaStream_t)stream);
}

int rm_total_size(rm_index* idx, int64_t* out_h, void* stream) {
  if (!idx || !out_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  return rm_launch_total_size(idx, (i64*)out_h, (cudaStream_t)stream);
}

int rm_dump_order(rm_index* idx, int32_t cap, uint64_t* order_h, void* stream) {
  if (!idx || cap < 0 || !order_h) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  const int n = cap < idx->t.max_nodes ? cap : idx->t.max_nodes;
  RM_CUDA(cudaMemcpyAsync(order_h, idx->t.ord, sizeof(u64) * (size_t)n, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

int rm_read_runs(rm_index* idx, const int64_t* offs_h, const int32_t* lens_h, int32_t n, int64_t* out_values_h, void* stream) {
  if (!idx || n < 0 || (n > 0 && (!offs_h || !lens_h || !out_values_h))) return RM_E_INVALID;
  if (idx->cfg.mode == RM_MODE_ROUTER) return RM_E_MODE;
  if (n == 0) return RM_OK;
  if (n > idx->t.max_nodes) return RM_E_TOO_LARGE;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RmScratch& sc = idx->sc;
  // where each run starts in the packed output: prefix sums on the host (n is a few thousand)
  i64* at_h = (i64*)malloc(sizeof(i64) * ((size_t)n + 1));
  if (!at_h) return RM_E_NOSPACE;
  i64 total = 0;
  for (int i = 0; i < n; ++i) {
    if (offs_h[i] < 0 || lens_h[i] < 0 || offs_h[i] + lens_h[i] > idx->t.max_tokens) {
      free(at_h);
      return RM_E_INVALID;
    }
    at_h[i] = total;
    total += lens_h[i];
  }
  at_h[n] = total;
  int rc = rm_stage_reserve(idx, total > 0 ? total : 1, true);
  if (rc) {
    free(at_h);
    return rc;
  }
  // the eviction scratch is free between calls: run starts, lengths, output positions
  i64* d_off = reinterpret_cast<i64*>(sc.ev_key[0]);
  i64* d_at = reinterpret_cast<i64*>(sc.ev_key[1]);
  int* d_len = sc.ev_rem;
  cudaError_t e = cudaMemcpyAsync(d_off, offs_h, sizeof(i64) * (size_t)n, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_len, lens_h, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_at, at_h, sizeof(i64) * (size_t)n, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);      // at_h is pageable: the copies are done before it is freed
  free(at_h);
  if (e != cudaSuccess) {
    rm_set_error("rm_read_runs: %s", cudaGetErrorString(e));
    return RM_E_CUDA;
  }
  if (total > 0) {
    rc = rm_launch_copy_runs(idx, d_off, d_len, d_at, n, sc.h_values, s);
    if (rc) return rc;
    RM_CUDA(cudaMemcpyAsync(out_values_h, sc.h_values, sizeof(i64) * (size_t)total, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaStreamSynchronize(s));
  }
  return RM_OK;
}

int rm_read_arena(rm_index* idx, int64_t off, int64_t n, int32_t* tokens_h, int64_t* values_h, void* stream) {
  if (!idx || off < 0 || n < 0 || off + n > idx->t.max_tokens) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  if (tokens_h) {
    int rc = rm_read_arena_component(idx, off, n, 0, tokens_h, s);
    if (rc) return rc;
  }
  if (values_h && idx->cfg.mode != RM_MODE_ROUTER)
    RM_CUDA(cudaMemcpyAsync(values_h, idx->t.val + off, sizeof(i64) * (size_t)n, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

}  // extern "C"
