"""rm_evict alone: the C3 loop of bench.py (4,096-record insert batches, evict what is over 90 % occupancy) on a
tree of --nodes nodes, eviction calls timed on the host clock.  For profiling the eviction kernels (ncu launch list)
without the rest of the bench.  Prints one JSON line.

    python tools/bench_evict.py [--nodes 1000000] [--batches 12]
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from radixmesh_b200 import capi  # noqa: E402


def run(nodes, batches, warmup=2):
    wl = dict(bench.WORKLOADS["target"], tree_nodes=nodes)
    L, NB = wl["L"], bench.INSERT_BATCH
    docs = bench.gen_docs(wl)
    dev = torch.device("cuda")
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, device=0, max_nodes=int(nodes * 1.3) + (1 << 16),
                          max_tokens=int(nodes * L * 0.6) + (1 << 24), max_batch=NB, max_key_len=1024)
    offs = torch.arange(0, (NB + 1) * L, L, dtype=torch.int64, device=dev)
    out = torch.empty(NB, dtype=torch.int32, device=dev)
    vc = 0
    stream = bench.insert_stream(wl, docs)
    for batch in stream:
        ix.insert(torch.from_numpy(batch.reshape(-1)).to(dev), torch.arange(vc, vc + NB * L, dtype=torch.int64, device=dev), offs, None, NB, out)
        vc += NB * L
        if ix.sizes().n_nodes >= nodes:
            break
    s0 = ix.sizes()
    capacity = int((s0.evictable + s0.protected_) / 0.9)
    calls, secs, n_nodes, asked = 0, 0.0, 0, 0
    for b in range(batches + warmup):
        batch = next(stream)
        ix.insert(torch.from_numpy(batch.reshape(-1)).to(dev), torch.arange(vc, vc + NB * L, dtype=torch.int64, device=dev), offs, None, NB, out)
        vc += NB * L
        while True:
            s = ix.sizes()
            over = s.evictable + s.protected_ - int(0.9 * capacity)
            if over <= 0:
                break
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            nodes_e, _, lens = ix.evict(over)
            dt = time.perf_counter() - t0
            if b >= warmup:
                calls += 1
                secs += dt
                n_nodes += len(nodes_e)
                asked += over
            if len(nodes_e) == 0:
                break
    ix.check_device_error()
    s1 = ix.sizes()
    ix.close()
    return {"tree_nodes": int(s0.n_nodes), "node_hwm": int(s1.node_hwm), "evict_calls": calls, "ms_per_evict_call": secs / max(calls, 1) * 1e3,
            "nodes_per_call": n_nodes / max(calls, 1), "tokens_asked_per_call": asked / max(calls, 1)}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=1_000_000)
    ap.add_argument("--batches", type=int, default=12)
    a = ap.parse_args()
    print(json.dumps(run(a.nodes, a.batches)))
