"""2..8-GPU check + timing of the mesh routers over NCCL (run under torchrun; one process per GPU).

GPU g hosts mesh node g's tree with the keys rank g inserted (first half of the ranks are prefill
nodes, second half decode nodes).  Requests = 2K-token system prompts + turns (BASELINE configs[3]
shape, reduced counts).  Prints one JSON line from rank 0."""
import json, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radixmesh_b200 import capi
from radixmesh_b200.mesh_router import ShardedMeshRouter
from radixmesh_b200.router import ConsistentHash

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
P = max(1, world // 2); D = world - P
rng = np.random.default_rng(2)
n_prompts, Lp, B = 64, 2048, 16384
prompts = rng.integers(0, 32000, (n_prompts, Lp), dtype=np.int32)
ix = capi.DeviceIndex(mode=capi.MODE_DECODE, device=local, self_rank=rank, n_prefill=P, n_decode=D, max_nodes=1 << 20,
                      max_tokens=1 << 26, max_batch=B, max_key_len=4096)
pa = [f"p{i}:1" for i in range(P)]; da = [f"d{i}:1" for i in range(max(D, 1))]
ix.set_rings(*ConsistentHash(pa).device_arrays(pa), *ConsistentHash(da).device_arrays(da))
# every rank inserts the prompts it owns (+ one turn), ownership spread over ranks
mine = [p for p in range(n_prompts) if p % world == rank]
L = Lp + 256
keys = np.concatenate([np.concatenate([prompts[p], rng.integers(0, 32000, 256, dtype=np.int32)]) for p in mine]).astype(np.int32)
offs = np.arange(0, (len(mine) + 1) * L, L, dtype=np.int64)
ix.insert_host(keys, np.arange(len(keys), dtype=np.int64), offs, np.full(len(mine), rank, np.int32))
# request batch (valid on rank 0, broadcast to the rest)
w = 1.0 / np.arange(1, n_prompts + 1) ** 1.1; w /= w.sum()
qd = np.random.default_rng(7).choice(n_prompts, size=B, p=w)
q = np.concatenate([prompts[qd], np.random.default_rng(8).integers(0, 32000, (B, 128), dtype=np.int32)], axis=1)
Lq = q.shape[1]
tok = torch.from_numpy(q.reshape(-1).copy()).to(dev) if rank == 0 else torch.zeros(B * Lq, dtype=torch.int32, device=dev)
qo = torch.arange(0, (B + 1) * Lq, Lq, dtype=torch.int64, device=dev)
def timed(router, K=10):
    for _ in range(3):
        out = router.route(tok, qo, B)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        out = router.route(tok, qo, B)
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return out, ms.item() / K

out, ms_batch = timed(ShardedMeshRouter(ix, rank, world))
out_full, ms_full = timed(ShardedMeshRouter(ix, rank, world, shard_pick=False))
pf = out[0].cpu().numpy(); dc = out[1].cpu().numpy()
same = all(bool((a.cpu() == b.cpu()).all()) for a, b in zip(out, out_full))
exp_owner = qd % world
ok_p = bool(np.all(np.where(exp_owner < P, pf == exp_owner, pf == -1)))
ok_d = bool(np.all(np.where(exp_owner >= P, dc == exp_owner, dc == -1)))
# phase breakdown on rank 0 (one extra batch, events between the phases; the sharded-pick pipeline)
import numpy as _np
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
be = ix.be
Bs = (B + world - 1) // world
bounds = [min(j * Bs, B) for j in range(world + 1)]
lo, n = bounds[rank], bounds[rank + 1] - bounds[rank]
ev[0].record()
dist.broadcast(tok, src=0); dist.broadcast(qo, src=0)
ev[1].record()
lens = be.empty(B, _np.int32); nodes = be.empty(B, _np.int32)
ix.match(tok, qo, B, lens, nodes)
ev[2].record()
depths = be.empty(world * n, _np.int32)
dist.all_to_all_single(depths, lens, [n] * world, [bounds[j + 1] - bounds[j] for j in range(world)])
ev[3].record()
loc = be.empty(4 * Bs, _np.int32); e8 = be.empty(Bs, _np.uint8)
ix.pick(depths, world, tok, qo[lo:], n, loc[0:Bs], loc[Bs:2 * Bs], loc[2 * Bs:3 * Bs], loc[3 * Bs:4 * Bs], e8)
ev[4].record()
g = be.empty(world * 4 * Bs, _np.int32); ge = be.empty(world * Bs, _np.uint8)
dist.all_gather_into_tensor(g, loc); dist.all_gather_into_tensor(ge, e8)
ev[5].record(); torch.cuda.synchronize()
phases = {nm: round(ev[i].elapsed_time(ev[i + 1]), 3) for i, nm in enumerate(["broadcast_ms", "walk_ms", "all_to_all_ms", "pick_md5_slice_ms", "allgather_results_ms"])}
if rank == 0:
    print(json.dumps({"phases": phases, "mesh_route": "sharded trees + all-to-all of matched lengths + sharded pick + all-gather of results",
                      "n_gpus": world, "requests_per_s": B / (ms_batch * 1e-3), "tree_lookups_per_s": world * B / (ms_batch * 1e-3),
                      "ms_per_batch": ms_batch, "ms_per_batch_unsharded_pick": ms_full, "same_result_as_unsharded_pick": same,
                      "batch": B, "request_len": Lq, "owner_pick_correct": ok_p and ok_d}))
dist.destroy_process_group()
