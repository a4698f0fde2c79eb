// Host-side state of one index and the launchers each .cu file exports to rm_api.cu.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include "rm_common.cuh"

struct RmScratch {
  int max_batch;
  u64* key[2];        // sort ping-pong: split requests of a match batch / group keys of an insert batch
  unsigned* val[2];
  u64* sufmax;        // per sorted split request: max op seq of requests at or beyond it on the same edge
  int* krank;         // per sorted split request: rank of its split point among the edge's distinct points
  int* uniq_pos;      // per edge group: sorted positions of the distinct split points
  unsigned* uniq_min; // per distinct split point: lowest query index that asked for it
  int* link;          // per query: predecessor in its node's request chain (RmTree.req_link points here)
  unsigned* hist;     // radix-sort histogram table
  size_t hist_words;
  // eviction (sized by max_nodes)
  u64* ev_key[2];
  unsigned* ev_val[2];
  unsigned* ev_hist;
  int* ev_lens;       // value-run lengths of the evicted nodes, in eviction order
  u64* ev_eff;        // effective (subtree-max) stamps
  int* ev_rem;        // children still to report, per node
  u64* ev_selw;       // radix select of the eviction frontier: token weight per digit (4096) + its state behind them
  unsigned* ev_selc;  // ... and key count per digit
  int* ins_mark;      // per node, zero between insert batches: root child whose first-token group stays on the ordered path
  // lock batches (sized by max_nodes)
  int* lock_first;
  int* lock_cnt;
  int* lock_flag;
  // duplicate-node table scan output: rows of (tree node, rank, parked node, parked lock_ref)
  int* dup_out;
  int* dup_out_n;
  // staging for the *_h entry points
  int* h_tokens;      // device copies of host inputs
  i64* h_values;
  i64* h_offsets;
  int* h_ranks;
  int* h_out_a;
  int* h_out_b;
  i64 h_tokens_cap;
  // rm_node_info staging (fixed size)
  int* info_ids;
  rm_node_info_t* info_rows;
};

#define RM_EVICT_FIRST 8192        // evicted nodes whose (id, offset, length) travel with the count
#define RM_MAIL_BYTES (4096 + RM_EVICT_FIRST * 16)

struct rm_index {
  rm_config cfg;
  RmTree t;           // device pointers, passed by value to kernels
  RmScratch sc;
  u64 op_seq;         // logical clock: operations applied so far
  u64 ins_records;    // insert records applied since the last reset (a cold index applies its first batches in chunks)
  int hwm_exact;      // node high-water mark the host last read from the device ...
  u64 hwm_seq;        // ... and the logical clock at that moment: an operation creates at most three nodes, which bounds
                      // the mark without asking the device (rm_evict sizes its grids from the bound)
  void* mail;         // pinned host memory, RM_MAIL_BYTES: small results come back through it in one asynchronous
                      // copy + one synchronisation (a copy into pageable memory blocks the host once per call)
  i64 hint_tokens;    // tokens of the batch being inserted when the caller knows them (the *_h entry points), else 0
  int node_bits;      // bits of a node id
  int len_bits;       // bits of an edge length / split point
  unsigned* ring_p;   // consistent-hash rings (device)
  int* ring_p_owner;
  int ring_p_n;
  unsigned* ring_d;
  int* ring_d_owner;
  int ring_d_n;
  unsigned char* route_msg;   // rendered str(list) messages of a route / pick batch (rm_route.cu); NULL until rings are set
  size_t route_msg_bytes;
  int* route_nblk;            // 64-byte blocks per rendered request, 0 = hash it with the streamed form
  // optional CUDA events around the stages of each match batch (rm_profile)
#define RM_PROF_EVENTS 384
#define RM_PROF_STRIDE 3
#ifndef RM_EMU
  cudaEvent_t prof_ev[RM_PROF_EVENTS];
  int prof_on;
  int prof_n;
#endif
};

void rm_set_error(const char* fmt, ...);
#define RM_CUDA(call)                                                                         \
  do {                                                                                        \
    cudaError_t _e = (call);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      rm_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return RM_E_CUDA;                                                                       \
    }                                                                                         \
  } while (0)

// rm_match.cu
int rm_launch_match(rm_index* idx, const int* tokens, const i64* offsets, int B, int* out_len, int* out_node,
                    cudaStream_t s);
int rm_launch_gather(rm_index* idx, const int* nodes, const int* lens, const i64* out_offsets, int B, i64* out_values,
                     cudaStream_t s);
int rm_launch_route(rm_index* idx, const int* tokens, const i64* offsets, int B, int warm, int* out_prefill,
                    int* out_decode, int* out_paddr, int* out_daddr, unsigned char* out_err, cudaStream_t s);
// rm_insert.cu
int rm_launch_insert(rm_index* idx, const int* tokens, const i64* values, const i64* offsets, const int* ranks, int N,
                     int* out_prefix_len, cudaStream_t s);
// rm_maint.cu
int rm_launch_reset(rm_index* idx, cudaStream_t s);
int rm_launch_lock(rm_index* idx, const int* nodes, int B, int dir, i64* out_delta, cudaStream_t s);
int rm_launch_lock_g(rm_index* idx, const int* nodes, const int* gens, int B, int dir, i64* out_delta, cudaStream_t s);
int rm_launch_copy_runs(rm_index* idx, const i64* d_off, const int* d_len, const i64* d_at, int n, i64* d_out, cudaStream_t s);
int rm_launch_evict(rm_index* idx, i64 num_tokens, int cap, int* out_nodes_h, i64* out_offs_h, int* out_lens_h,
                    int* out_n_h, cudaStream_t s);
int rm_launch_node_info(rm_index* idx, const int* ids_h, int n, rm_node_info_t* out_h, cudaStream_t s);
int rm_launch_node_children(rm_index* idx, int node, int cap, int* ids_h, int* toks_h, int* gens_h, u64* ords_h, int* out_n_h,
                            cudaStream_t s);
int rm_launch_total_size(rm_index* idx, i64* out_h, cudaStream_t s);
int rm_launch_node_info_dev(rm_index* idx, const int* ids_d, int n, cudaStream_t s);   // rows into sc.info_rows (n <= 256)
int rm_table_rebuild(rm_index* idx, int n_nodes_hwm, cudaStream_t s);
int rm_launch_dups(rm_index* idx, int only_candidates, int cap, int* out_h, int* out_n_h, cudaStream_t s);
int rm_launch_gc_agree(rm_index* idx, const int* tokens_h, const i64* offsets_h, const int* ranks_h, int n,
                       unsigned char* out_agree_h, cudaStream_t s);
int rm_launch_gc_exec(rm_index* idx, const int* tokens_h, const i64* offsets_h, const int* ranks_h, int n, int free_parked,
                      i64* out_off_h, int* out_len_h, cudaStream_t s);
int rm_launch_gc_keys(rm_index* idx, const int* nodes_h, int n, int stride, int* out_tokens_h, int* out_lens_h,
                      cudaStream_t s);
