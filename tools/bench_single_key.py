"""Single-key latency of the drop-in RadixCache at tree sizes where round 1 copied the whole node pool per call.

    python tools/bench_single_key.py [--nodes 100000 1000000]

Builds the tree through `RadixCache.insert_batch` (bench.py's target-shape generator, 512-token keys), then times
`match_prefix(key)` - walk, split, value gather, `MatchResult.last_device_node` - and reads of `.lock_ref` / `.parent` on
the returned handle, one key at a time, next to the oracle port on a bounded replica (first 8192 inserts of the same
stream).  What it shows: the call costs a few launches + a few dozen bytes of D2H whatever the tree size
(rm_node_info), i.e. it is launch-latency bound, not O(tree).  Prints one JSON line."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from radixmesh_b200 import RadixCache  # noqa: E402


def run(nodes, n_keys=300):
    wl = dict(bench.WORKLOADS["target"], tree_nodes=nodes)
    L, NB = wl["L"], bench.INSERT_BATCH
    docs = bench.gen_docs(wl)
    dev = torch.device("cuda")
    t = RadixCache(None, None, 1, max_nodes=int(nodes * 1.3) + (1 << 16), max_tokens=int(nodes * L * 0.6) + (1 << 24),
                   max_batch=NB, max_key_len=1024)
    offs = torch.arange(0, (NB + 1) * L, L, dtype=torch.int64, device=dev)
    vc = 0
    for batch in bench.insert_stream(wl, docs):
        t.insert_batch(torch.from_numpy(batch.reshape(-1)).to(dev), torch.arange(vc, vc + NB * L, dtype=torch.int64, device=dev), offs)
        vc += NB * L
        if t._ix.sizes().n_nodes >= nodes:
            break
    keys = bench.gen_batch(bench.query_seed(wl, 0, 5, 0), n_keys, L, docs)
    lists = [k.tolist() for k in keys]
    for k in lists[:20]:
        t.match_prefix(k)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in lists[20:]:
        m = t.match_prefix(k)
    t_match = (time.perf_counter() - t0) / (n_keys - 20)
    t0 = time.perf_counter()
    for k in lists[20:]:
        m = t.match_prefix(k)
        _ = m.last_device_node.lock_ref
        _ = m.last_device_node.parent
    t_handle = (time.perf_counter() - t0) / (n_keys - 20)
    return {"tree_nodes": int(t._ix.sizes().n_nodes), "match_prefix_us": 1e6 * t_match, "match_prefix_plus_handle_reads_us": 1e6 * t_handle}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, nargs="+", default=[100_000, 1_000_000])
    a = ap.parse_args()
    out = {"what": "RadixCache.match_prefix(key), one 512-token key per call, through the drop-in class", "device": [run(n) for n in a.nodes]}
    from oracle.radix_port import PortRadixCache
    wl = bench.WORKLOADS["target"]
    docs = bench.gen_docs(wl)
    p = PortRadixCache(1)
    vals = np.arange(wl["L"], dtype=np.int64)
    for b, batch in enumerate(bench.insert_stream(wl, docs)):
        if b >= 2:
            break
        for r in batch:
            p.insert(r.tolist(), vals)
    keys = [k.tolist() for k in bench.gen_batch(bench.query_seed(wl, 0, 5, 0), 300, wl["L"], docs)]
    t0 = time.perf_counter()
    for k in keys:
        p.match_prefix(k)
    out["cpu_port"] = {"match_prefix_us": 1e6 * (time.perf_counter() - t0) / len(keys), "replica": "first 8192 inserts of the same stream"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
