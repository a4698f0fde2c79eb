"""BASELINE configs[2]: oplog-apply + LRU eviction on a tree kept at 90 % pool occupancy under streaming
inserts.  Secondary measurement (bench.py stays the headline contract); prints one JSON line.

    python tools/bench_c3.py [--nodes 1000000] [--batches 40] [--cpu-inserts 4096]

The tree is grown to `--nodes` nodes; the pool capacity C (tokens) is then set so that the tree is at 90 %
of C; from there every 4096-record insert batch is followed by evict(batch tokens) while occupancy is
above 0.9*C, and by rm_compact when the bump arena runs low.  Reports inserts/s, evicted nodes/s and the
share of time in each call, with the oracle (pure-Python port of the reference) on a sample beside it.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from radixmesh_b200 import capi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=1_000_000)
    ap.add_argument("--batches", type=int, default=40)
    ap.add_argument("--cpu-inserts", type=int, default=4096)
    args = ap.parse_args()
    wl = dict(name="C3", n_docs=32768, L=512, tree_nodes=args.nodes, B=65536, seed=1)
    L, NB = wl["L"], bench.INSERT_BATCH
    dev = torch.device("cuda")
    docs = bench.gen_docs(wl)
    arena = int(args.nodes * L * 0.6) + (1 << 24)
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, max_nodes=int(args.nodes * 1.5) + (1 << 16), max_tokens=arena,
                          max_batch=NB, max_key_len=1024)
    offs = torch.arange(0, (NB + 1) * L, L, dtype=torch.int64, device=dev)
    out = torch.empty(NB, dtype=torch.int32, device=dev)
    stream = bench.insert_stream(wl, docs)
    vc = 0

    def insert(batch):
        nonlocal vc
        tok = torch.from_numpy(batch.reshape(-1)).to(dev)
        val = torch.arange(vc, vc + NB * L, dtype=torch.int64, device=dev)
        vc += NB * L
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ix.insert(tok, val, offs, None, NB, out)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    t_build = 0.0
    n_build = 0
    while ix.sizes().n_nodes < args.nodes:
        t_build += insert(next(stream))
        n_build += NB
    s0 = ix.sizes()
    capacity = int((s0.evictable + s0.protected_) / 0.9)
    t_ins = t_ev = t_cp = 0.0
    n_ev = n_cp = n_ev_calls = 0
    for _ in range(args.batches):
        if ix.sizes().arena_used + NB * L > arena:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ix.compact()
            torch.cuda.synchronize()
            t_cp += time.perf_counter() - t0
            n_cp += 1
        t_ins += insert(next(stream))
        while True:
            s = ix.sizes()
            if s.evictable + s.protected_ <= 0.9 * capacity:
                break
            t0 = time.perf_counter()
            nodes, _, _ = ix.evict(NB * L // 4)
            t_ev += time.perf_counter() - t0
            n_ev += len(nodes)
            n_ev_calls += 1
            if len(nodes) == 0:
                break
    ix.check_device_error()
    s1 = ix.sizes()
    # oracle on a sample: same stream from an empty tree (a 1M-node Python tree takes minutes to build)
    from oracle.radix_port import PortRadixCache
    t = PortRadixCache(1)
    rows = next(bench.insert_stream(wl, docs))[:args.cpu_inserts]
    lists = [r.tolist() for r in rows]
    vals = np.arange(L, dtype=np.int64)
    t0 = time.perf_counter()
    for k in lists:
        t.insert(k, vals)
    cpu_ins = len(lists) / (time.perf_counter() - t0)
    t0 = time.perf_counter()
    before = len(t.freed)
    t.evict(len(lists) * L // 8)
    cpu_ev = (len(t.freed) - before) / (time.perf_counter() - t0)
    total = t_ins + t_ev + t_cp
    print(json.dumps({
        "workload": f"C3 streaming insert + LRU evict at 90% occupancy, {args.nodes}-node tree, {L}-token keys, {NB}-record batches",
        "build": {"inserts": n_build, "inserts_per_s": n_build / t_build},
        "steady": {"batches": args.batches, "inserts_per_s": args.batches * NB / t_ins,
                   "inserts_per_s_incl_evict_compact": args.batches * NB / total,
                   "evicted_nodes": n_ev, "evict_calls": n_ev_calls, "evicted_nodes_per_s": n_ev / t_ev if t_ev else None,
                   "ms_per_evict_call": 1e3 * t_ev / max(n_ev_calls, 1), "compactions": n_cp,
                   "ms_per_compaction": 1e3 * t_cp / max(n_cp, 1),
                   "time_share": {"insert": t_ins / total, "evict": t_ev / total, "compact": t_cp / total}},
        "tree": {"nodes_start": int(s0.n_nodes), "nodes_end": int(s1.n_nodes), "capacity_tokens": capacity,
                 "occupancy_end": (s1.evictable + s1.protected_) / capacity},
        "cpu_port_1core": {"inserts_per_s": cpu_ins, "evicted_nodes_per_s": cpu_ev,
                           "sample": f"{len(lists)} inserts into an empty tree, then evict 1/8 of the tokens"}}))


if __name__ == "__main__":
    main()
