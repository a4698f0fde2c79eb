"""Experiment: walk-kernel time vs batch size (same tree, prefixes of the same query batch, tree L2-warm)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from radixmesh_b200 import capi
wl = bench.WORKLOADS["c2"]; B, L = wl["B"], wl["L"]
docs = bench.gen_docs(wl); dev = torch.device("cuda")
ix = capi.DeviceIndex(mode=0, max_nodes=2_000_000, max_tokens=int(wl["tree_nodes"] * L * 0.9) + (1 << 22), max_batch=B, max_key_len=1024)
offs_ins = torch.arange(0, (bench.INSERT_BATCH + 1) * L, L, dtype=torch.int64, device=dev)
out_ins = torch.empty(bench.INSERT_BATCH, dtype=torch.int32, device=dev); vc = 0
for batch in bench.insert_stream(wl, docs):
    tok = torch.from_numpy(batch.reshape(-1)).to(dev)
    val = torch.arange(vc, vc + bench.INSERT_BATCH * L, dtype=torch.int64, device=dev)
    ix.insert(tok, val, offs_ins, None, bench.INSERT_BATCH, out_ins); vc += bench.INSERT_BATCH * L
    if ix.sizes().n_nodes >= wl["tree_nodes"]: break
ol = torch.empty(B, dtype=torch.int32, device=dev); on = torch.empty(B, dtype=torch.int32, device=dev)
qs = [torch.from_numpy(bench.gen_batch(bench.query_seed(wl, 0, 0, s), B, L, docs).reshape(-1)).to(dev) for s in range(4)]
offs_q = torch.arange(0, (B + 1) * L, L, dtype=torch.int64, device=dev)
for s in range(4):   # warm: all splits of these batches done
    ix.match(qs[s], offs_q, B, ol, on)
torch.cuda.synchronize(); ix.profile(True)
for b in ([int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else (65536, 32768, 16384, 8192, 4096, 1024, 256, 32)):
    ts = []
    for s in range(4):
        ix.match(qs[s], offs_q[:b + 1], b, ol, on); torch.cuda.synchronize(); ts.append(ix.profile_read()[0])
    print("B", b, "walk ms", [round(t, 4) for t in ts], "max len", int(ol[:b].max()))
