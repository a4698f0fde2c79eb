"""Experiment (not part of the product): walk-kernel time on a fresh batch vs the same batch again
(tree and queries then L2 resident, no new splits), for several node-pool sizes."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from radixmesh_b200 import capi

wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "c2"]
B, L = wl["B"], wl["L"]
docs = bench.gen_docs(wl)
dev = torch.device("cuda")
for max_nodes in [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "600000,4000000").split(",")]:
    ix = capi.DeviceIndex(mode=0, max_nodes=max_nodes, max_tokens=int(wl["tree_nodes"] * L * 0.9) + (1 << 22), max_batch=B, max_key_len=1024)
    offs_ins = torch.arange(0, (bench.INSERT_BATCH + 1) * L, L, dtype=torch.int64, device=dev)
    out_ins = torch.empty(bench.INSERT_BATCH, dtype=torch.int32, device=dev)
    vc = 0
    for batch in bench.insert_stream(wl, docs):
        tok = torch.from_numpy(batch.reshape(-1)).to(dev)
        val = torch.arange(vc, vc + bench.INSERT_BATCH * L, dtype=torch.int64, device=dev)
        ix.insert(tok, val, offs_ins, None, bench.INSERT_BATCH, out_ins)
        vc += bench.INSERT_BATCH * L
        if ix.sizes().n_nodes >= wl["tree_nodes"]:
            break
    offs_q = torch.arange(0, (B + 1) * L, L, dtype=torch.int64, device=dev)
    ol = torch.empty(B, dtype=torch.int32, device=dev); on = torch.empty(B, dtype=torch.int32, device=dev)
    batches = [torch.from_numpy(bench.gen_batch(bench.query_seed(wl, 0, 0, s), B, L, docs).reshape(-1)).to(dev) for s in range(6)]
    ix.profile(True)
    res = []
    for s in range(6):
        ix.match(batches[s], offs_q, B, ol, on); torch.cuda.synchronize(); a = ix.profile_read()[0]
        ix.match(batches[s], offs_q, B, ol, on); torch.cuda.synchronize(); b = ix.profile_read()[0]
        ix.match(batches[s], offs_q, B, ol, on); torch.cuda.synchronize(); c = ix.profile_read()[0]
        res.append((round(a, 4), round(b, 4), round(c, 4)))
    print("max_nodes", max_nodes, "nodes", ix.sizes().n_nodes, "walk ms (fresh, again, again):", res)
    ix.check_device_error()
    ix.close()
