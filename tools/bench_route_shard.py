"""Replicated mesh route of ONE rank's shard at several shard sizes (the per-GPU latency floor of rm_route_batch).

    python tools/bench_route_shard.py [--sizes 8192,65536] [--steps 10]

Same C4 workload as bench.py's mesh_route block (tools/bench_blocks.py): an 8-rank router index, 2K-token requests.
A shard of 64K requests is the N=1 case of the block, 8K the N=8 case.  Prints one JSON line; under ncu the launch list
splits a shard into walk / render / MD5 (profiles/README.md).
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench_blocks as bb  # noqa: E402
from radixmesh_b200 import capi  # noqa: E402
from radixmesh_b200.mesh_router import ReplicatedMeshRouter  # noqa: E402
from radixmesh_b200.router import ConsistentHash  # noqa: E402


def run(sizes, steps):
    dev = torch.device("cuda")
    world, P, D = 8, 4, 4
    wk = bb.c4_workload(world)
    pa = [f"p{i}:1" for i in range(P)]
    da = [f"d{i}:1" for i in range(D)]
    rings = (*ConsistentHash(pa).device_arrays(pa), *ConsistentHash(da).device_arrays(da))
    bmax = max(sizes)
    rp = capi.DeviceIndex(mode=capi.MODE_ROUTER, self_rank=P + D, max_dups=1 << 20, device=0, n_prefill=P, n_decode=D,
                          max_nodes=1 << 19, max_tokens=1 << 25, max_batch=bmax, max_key_len=wk["Lmax"],
                          max_batch_tokens=int(bmax * 2600))
    rp.set_rings(*rings)
    bb.c4_apply_inserts(rp, wk, bb.c4_inserts(wk), with_values=False)
    rp.compact()
    pool_d = torch.from_numpy(wk["pool"]).to(dev)
    r = ReplicatedMeshRouter(rp, 0, 1)
    rows = []
    for B in sizes:
        bb.C4["B"] = B
        tok, offs, _, _ = bb.c4_batch(torch, wk, pool_d, 0, dev)
        for _ in range(3):
            out = r.route_shard(tok, offs, B)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = r.route_shard(tok, offs, B)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        rows.append({"requests": B, "ms": ms, "requests_per_s": B / (ms * 1e-3), "tokens": int(tok.numel()),
                     "hashed_fraction": float(((out[0] == -1) | (out[1] == -1)).float().mean())})
    rp.close()
    return {"workload": "C4 replicated route, one rank's shard", "shards": rows}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="8192,65536")
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    print(json.dumps(run([int(x) for x in a.sizes.split(",")], a.steps)))
