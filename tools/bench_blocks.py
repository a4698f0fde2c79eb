"""Secondary blocks of bench.py's JSON line (the headline stays C2): the north_star target shape, C3 (streaming insert +
LRU evict at 90 % occupancy), the C4 mesh-route path under --gpus N, and a pinned host-to-device copy probe.

Every block carries its own roofline figures (SURVEY.md 8d byte formulas over CUDA-event kernel time), its own bounded
CPU-port sample and its own parity check.  The oracle (oracle/) is used here only as the checker and as the timed
CPU sample, never on the measured path.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _events(torch, n):
    return [torch.cuda.Event(enable_timing=True) for _ in range(n)]


def node_hops(ix, nodes):
    """How many nodes lie on the root path of each node id in `nodes` (= the hops a walk that ends there makes), by
    climbing the device tree's parent array.  Used for the 8d byte model where the Python port cannot hold the tree."""
    parent = ix.dump()["parent"].astype(np.int64)
    cur = np.asarray(nodes, dtype=np.int64).copy()
    hops = np.zeros(len(cur), dtype=np.int64)
    while True:
        live = cur > 0
        if not live.any():
            return hops
        hops += live
        cur = np.where(live, parent[np.maximum(cur, 0)], 0)


def build_tree(bench, wl, docs, ix, dev, torch, keep_rows):
    L, nb = wl["L"], bench.INSERT_BATCH
    offs = torch.arange(0, (nb + 1) * L, L, dtype=torch.int64, device=dev)
    out = torch.empty(nb, dtype=torch.int32, device=dev)
    rows, vc, n_ins = [], 0, 0
    stream = bench.insert_stream(wl, docs)
    for batch in stream:
        tok = torch.from_numpy(batch.reshape(-1)).to(dev)
        val = torch.arange(vc, vc + nb * L, dtype=torch.int64, device=dev)
        ix.insert(tok, val, offs, None, nb, out)
        if keep_rows:
            rows.append(batch)
        vc += nb * L
        n_ins += nb
        if ix.sizes().n_nodes >= wl["tree_nodes"]:
            break
    ix.check_device_error()
    return (np.concatenate(rows) if keep_rows else None), n_ins, vc, stream


def port_sample_tree(bench, wl, docs, n_batches):
    """A bounded replica for the CPU port: the first n_batches insert batches of the same stream."""
    from oracle.radix_port import PortRadixCache
    t = PortRadixCache(1)
    L = wl["L"]
    vals = np.arange(L, dtype=np.int64)
    n = 0
    t0 = time.perf_counter()
    for b, batch in enumerate(bench.insert_stream(wl, docs)):
        if b >= n_batches:
            break
        for row in batch:
            t.insert(row.tolist(), vals)
            n += 1
    return t, n, time.perf_counter() - t0


def target_block(bench, torch, capi, peak, steps=6, warmup=3, cpu_queries=2048, cpu_batches=4, check_queries=8192):
    """north_star target shape: 1 M-node tree, 65,536 x 512-token queries (the >= 100 M lookups/s at >= 50 % roofline
    bar is quoted on it).  Returns (block dict, live index, state for the C3 block that continues on the same tree)."""
    from oracle.lcp_oracle import longest_prefix_lengths
    wl = bench.WORKLOADS["target"]
    L, B = wl["L"], wl["B"]
    dev = torch.device("cuda", torch.cuda.current_device())
    docs = bench.gen_docs(wl)
    c3_batches = 12
    max_nodes = int(wl["tree_nodes"] * 1.3) + (steps + warmup + 2) * (B // 2) + c3_batches * 2 * bench.INSERT_BATCH + (1 << 16)
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, device=dev.index, max_nodes=max_nodes, max_batch=B, max_key_len=1024,
                          max_tokens=int(wl["tree_nodes"] * L * 0.6) + c3_batches * bench.INSERT_BATCH * L + (1 << 24))
    t0 = time.perf_counter()
    ins, n_ins, vc, stream = build_tree(bench, wl, docs, ix, dev, torch, keep_rows=True)
    ix.compact()                       # heavy paths contiguous (index maintenance, outside the timed region)
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    nodes_start = int(ix.sizes().n_nodes)
    offs_q = torch.arange(0, (B + 1) * L, L, dtype=torch.int64, device=dev)
    host_q = [bench.gen_batch(bench.query_seed(wl, 0, 3, s), B, L, docs) for s in range(steps + warmup)]
    dev_q = [torch.from_numpy(q.reshape(-1)).to(dev) for q in host_q]
    out_len = torch.empty((steps + warmup, B), dtype=torch.int32, device=dev)
    out_node = torch.empty((steps + warmup, B), dtype=torch.int32, device=dev)
    for s in range(warmup):
        ix.match(dev_q[s], offs_q, B, out_len[s], out_node[s])
    torch.cuda.synchronize()
    ix.profile(2)
    e0, e1 = _events(torch, 2)
    e0.record()
    for s in range(warmup, warmup + steps):
        ix.match(dev_q[s], offs_q, B, out_len[s], out_node[s])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    stage_ms, nprof = ix.profile_stages()
    ix.profile_read()
    ix.profile(False)
    ix.check_device_error()
    walk_ms, split_ms = stage_ms[0] / max(nprof, 1), stage_ms[1] / max(nprof, 1)
    # parity: every checked query's matched length = max LCP with the inserted keys (numpy oracle, full tree)
    s_chk = warmup
    got = out_len[s_chk].cpu().numpy()
    exp = longest_prefix_lengths(ins, host_q[s_chk][:check_queries])
    mism = int((got[:check_queries] != exp).sum())
    # 8d byte model: m from the oracle, hops from the device tree (the Python port cannot hold 1 M nodes)
    hops = node_hops(ix, out_node[s_chk].cpu().numpy())
    m = got.astype(np.int64)
    bpl = float((8 * np.minimum(L, m + 1) + 40 * hops + 8).mean())
    achieved = bpl * B / (walk_ms * 1e-3) / 1e9
    # CPU port on a bounded replica (first cpu_batches insert batches), queries of the same stream
    t, n_t, _ = port_sample_tree(bench, wl, docs, cpu_batches)
    lists = [r.tolist() for r in host_q[s_chk][:cpu_queries]]
    t1 = time.perf_counter()
    for k in lists:
        t.match_prefix(k)
    cpu_v = len(lists) / (time.perf_counter() - t1)
    block = {
        "workload": wl["name"], "value": B / (ms * 1e-3), "unit": "lookups/s", "ms_per_step": ms, "steps": steps, "warmup": warmup,
        "tree_nodes_start": nodes_start, "tree_nodes_end": int(ix.sizes().n_nodes), "inserts_to_build": n_ins,
        "build_s": round(build_s, 2), "mean_matched_len": float(m.mean()), "matched_tokens_per_s": B / (ms * 1e-3) * float(m.mean()),
        "layout": "rm_compact after the build (heavy paths contiguous)",
        "stages_ms_per_step": {"walk": walk_ms, "split_stage": split_ms},
        "parity_check": {"queries": int(check_queries), "mismatches": mism,
                         "oracle": "max LCP with all inserted keys (oracle/lcp_oracle.py, numpy)"},
        "roofline": {"bound": "hbm", "kernel": "rm_match_rung_kernel (K1 walk)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "bytes_per_lookup": bpl, "mean_hops": float(hops.mean()), "kernel_ms": walk_ms,
                     "hops_source": "device tree (parity-checked shape; the Python port cannot hold 1 M nodes)",
                     "step_frac": bpl * B / (ms * 1e-3) / 1e9 / peak},
        "cpu_baseline": {"value": cpu_v, "unit": "lookups/s", "cores": 1, "kind": "port",
                         "sample": f"{len(lists)} queries of the first timed batch on a replica holding the first {n_t} inserts"},
    }
    state = {"wl": wl, "docs": docs, "stream": stream, "vc": vc, "port": t, "port_inserts": n_t, "port_batches": cpu_batches}
    return block, ix, state


def c3_block(bench, torch, capi, peak, ix, state, mean_hops_per_token, batches=8, warmup=2, cpu_inserts=1024):
    """BASELINE configs[2]: streaming 4,096-record insert batches + LRU eviction at 90 % pool occupancy, on the
    1 M-node tree the target block leaves behind (capacity C is set so that the tree is at 90 % of C;
    after every batch, while evictable + protected > 0.9 C, evict what is over)."""
    wl, docs, stream = state["wl"], state["docs"], state["stream"]
    L, NB = wl["L"], bench.INSERT_BATCH
    dev = torch.device("cuda", torch.cuda.current_device())
    offs = torch.arange(0, (NB + 1) * L, L, dtype=torch.int64, device=dev)
    outs = torch.empty((batches + warmup, NB), dtype=torch.int32, device=dev)
    s0 = ix.sizes()
    capacity = int((s0.evictable + s0.protected_) / 0.9)
    vc = state["vc"]
    ins_ms, ev_s, ev_calls, ev_nodes, ev_tokens, ev_asked, ev_bad = [], 0.0, 0, 0, 0, 0, 0
    ev_scan_bytes = 0
    for b in range(batches + warmup):
        batch = next(stream)
        tok = torch.from_numpy(batch.reshape(-1)).to(dev)
        val = torch.arange(vc, vc + NB * L, dtype=torch.int64, device=dev)
        vc += NB * L
        e0, e1 = _events(torch, 2)
        torch.cuda.synchronize()
        e0.record()
        ix.insert(tok, val, offs, None, NB, outs[b])
        e1.record()
        torch.cuda.synchronize()
        if b >= warmup:
            ins_ms.append(e0.elapsed_time(e1))
        while True:
            s = ix.sizes()
            over = s.evictable + s.protected_ - int(0.9 * capacity)
            if over <= 0:
                break
            t0 = time.perf_counter()
            nodes, _, lens = ix.evict(over)
            dt = time.perf_counter() - t0
            if b >= warmup:
                ev_s += dt
                ev_calls += 1
                ev_nodes += len(nodes)
                ev_tokens += int(lens.sum())
                ev_asked += over
                ev_scan_bytes += 16 * int(s.node_hwm) + 2 * 5 * 12 * len(nodes)
            # the evicted set is the shortest LRU prefix that frees `over` tokens (radix_cache.py:183-202) ...
            tot = int(lens.sum())
            if len(nodes) == 0 or tot < over or tot - int(lens[-1]) >= over:
                ev_bad += 1
            s2 = ix.sizes()
            if s.evictable - s2.evictable != tot or s.n_nodes - s2.n_nodes != len(nodes):      # ... and the books agree
                ev_bad += 1
            if len(nodes) == 0:
                break
    ix.check_device_error()
    p = outs[warmup:].to(torch.int64)
    p_mean = float(p.to(torch.float64).mean().item())
    n_mean = L - p_mean
    h_mean = mean_hops_per_token * p_mean + 1.0
    bytes_per_insert = 12.0 * n_mean + 8.0 * p_mean + 40.0 * h_mean
    ms_ins = float(np.mean(ins_ms))
    ins_gbs = bytes_per_insert * NB / (ms_ins * 1e-3) / 1e9
    total_s = ms_ins * 1e-3 * batches + ev_s
    # CPU port: the same stream's next records into the bounded replica, then one eviction of what they added
    import itertools
    t = state["port"]
    rows = next(itertools.islice(bench.insert_stream(wl, docs), state["port_batches"], None))[:cpu_inserts]   # the batch after the replica's
    lists = [r.tolist() for r in rows]
    vals = np.arange(L, dtype=np.int64)
    ev0 = t.evictable_size()
    t0 = time.perf_counter()
    for k in lists:
        t.insert(k, vals)
    cpu_ins = len(lists) / (time.perf_counter() - t0)
    added = t.evictable_size() - ev0
    t0 = time.perf_counter()
    before = len(t.freed)
    t.evict(max(added, 1))
    cpu_ev_s = time.perf_counter() - t0
    cpu_ev_nodes = len(t.freed) - before
    s1 = ix.sizes()
    # the first batch into an EMPTY index (applied in chunks of 256 records: csrc/rm_insert.cu)
    cold = capi.DeviceIndex(mode=capi.MODE_BASE, device=dev.index, max_nodes=1 << 16, max_tokens=NB * L + (1 << 16), max_batch=NB, max_key_len=1024)
    first = next(bench.insert_stream(wl, docs))
    tok0 = torch.from_numpy(first.reshape(-1)).to(dev)
    val0 = torch.arange(NB * L, dtype=torch.int64, device=dev)
    e0, e1 = _events(torch, 2)
    torch.cuda.synchronize()
    e0.record()
    cold.insert(tok0, val0, offs, None, NB, outs[0])
    e1.record()
    torch.cuda.synchronize()
    cold_ms = e0.elapsed_time(e1)
    cold.check_device_error()
    cold.close()
    return {
        "cold_first_batch_ms": cold_ms,
        "workload": f"C3 streaming insert + LRU evict at 90% occupancy: 1M-node tree, {NB}-record batches of {L}-token keys",
        "value": NB * batches / total_s, "unit": "inserts/s (insert + evict time)", "batches": batches, "warmup": warmup,
        "inserts_per_s_insert_only": NB / (ms_ins * 1e-3), "ms_per_insert_batch": ms_ins,
        "evict_calls": ev_calls, "ms_per_evict_call": 1e3 * ev_s / max(ev_calls, 1), "evicted_nodes": ev_nodes,
        "evicted_nodes_per_s": ev_nodes / ev_s if ev_s else None, "evicted_tokens": ev_tokens, "tokens_asked": ev_asked,
        "capacity_tokens": capacity, "occupancy_end": (s1.evictable + s1.protected_) / capacity,
        "tree_nodes_end": int(s1.n_nodes), "mean_matched_prefix": p_mean,
        "parity_check": {"evict_calls_checked": ev_calls, "violations": ev_bad,
                         "property": "each call frees the shortest eviction-order prefix reaching the requested tokens; "
                                     "evictable and node count move by exactly what was handed back (evicted-set equality "
                                     "with the port: tests/cases.py run_stream_evict)"},
        "roofline": {"bound": "hbm", "kernel": "insert pipeline (pre-pass walk + group sort + rm_insert_kernel)",
                     "achieved": ins_gbs, "peak": peak, "unit": "GB/s", "frac": ins_gbs / peak,
                     "bytes_per_insert": bytes_per_insert,
                     "formula": "12*n + 8*p + 40*h (SURVEY 8d); h estimated from the match block's hops per matched token",
                     "evict": {"achieved": (ev_scan_bytes / ev_s / 1e9) if ev_s else None, "unit": "GB/s",
                               "frac": (ev_scan_bytes / ev_s / 1e9 / peak) if ev_s else None,
                               "formula": "16 B per node slot scanned + 2*passes*12 B per sorted candidate, over host wall "
                                          "time of rm_evict (it synchronises)"}},
        "cpu_baseline": {"value": cpu_ins, "unit": "inserts/s", "cores": 1, "kind": "port",
                         "evict_ms": 1e3 * cpu_ev_s, "evicted_nodes": cpu_ev_nodes,
                         "sample": f"{len(lists)} inserts into a replica holding the first {state['port_inserts']} inserts, then one evict "
                                   "of the tokens they added"},
    }


def h2d_probe(torch, world, mb=256, reps=6):
    """Aggregate pinned host-to-device copy rate with every rank copying at once (the ceiling of the e2e figure)."""
    import torch.distributed as dist
    dev = torch.device("cuda", torch.cuda.current_device())
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = _events(torch, 2)
    e0.record()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return world * reps * n / (float(t.item()) * 1e-3) / 1e9


# ---------------------------------------------------------------------------------------------- C4: mesh routing
C4 = dict(seed=2, n_prompts=64, prompt_len=2048, n_conv=16384, max_turns=8, turn_lo=64, turn_hi=256, B=65536)


def c4_workload(world):
    """BASELINE configs[3]: 64 distinct 2K-token system prompts (Zipf), conversations of 1-8 turns of
    64-256 tokens.  Conversation c has reached turn t_c: the prefixes through its earlier turns were inserted, each by
    the rank that served that turn (ownership spread over the ranks), and its next request is the prefix through turn t_c."""
    w = C4
    rng = np.random.default_rng(w["seed"])
    Lp, Lmax = w["prompt_len"], w["prompt_len"] + w["max_turns"] * w["turn_hi"]
    prompts = rng.integers(0, 32000, (w["n_prompts"], Lp), dtype=np.int32)
    zw = 1.0 / np.arange(1, w["n_prompts"] + 1) ** 1.1
    zw /= zw.sum()
    pid = rng.choice(w["n_prompts"], size=w["n_conv"], p=zw)
    T = rng.integers(1, w["max_turns"] + 1, size=w["n_conv"])
    tl = rng.integers(w["turn_lo"], w["turn_hi"] + 1, size=(w["n_conv"], w["max_turns"]))
    ends = Lp + np.cumsum(tl, axis=1)                                   # ends[c, j] = length of the prefix through turn j
    pool = rng.integers(0, 32000, (w["n_conv"], Lmax), dtype=np.int32)
    pool[:, :Lp] = prompts[pid]
    # KV indices: one per token of the workload, the same whoever inserts it (a prompt's are shared by its conversations)
    vals = np.arange(w["n_conv"], dtype=np.int64)[:, None] * Lmax + np.arange(Lmax, dtype=np.int64)[None, :]
    vals[:, :Lp] = (w["n_conv"] + pid.astype(np.int64))[:, None] * Lmax + np.arange(Lp, dtype=np.int64)[None, :]
    t_c = (rng.random(w["n_conv"]) * T).astype(np.int64)                # turn the conversation is at (0 = first turn)
    owner = rng.integers(0, world, size=(w["n_conv"], w["max_turns"]))  # rank that served (inserted) turn j
    req_len = ends[np.arange(w["n_conv"]), t_c]
    return dict(pool=pool, vals=vals, pid=pid, ends=ends, t_c=t_c, owner=owner, req_len=req_len, Lmax=Lmax)


def c4_inserts(wk, only_rank=None, convs=None):
    """(conversation, turn, owner) of every inserted prefix, turn-major (the order the mesh applies them in)."""
    out = []
    n_conv = wk["pool"].shape[0]
    cs = range(n_conv) if convs is None else convs
    for j in range(C4["max_turns"]):
        for c in cs:
            if j < wk["t_c"][c] and (only_rank is None or wk["owner"][c, j] == only_rank):
                out.append((c, j, int(wk["owner"][c, j])))
    return out


def c4_apply_inserts(ix, wk, recs, with_values, chunk=1024):
    pool, ends = wk["pool"], wk["ends"]
    for a in range(0, len(recs), chunk):
        part = recs[a:a + chunk]
        lens = np.asarray([ends[c, j] for c, j, _ in part], dtype=np.int64)
        offs = np.zeros(len(part) + 1, dtype=np.int64)
        np.cumsum(lens, out=offs[1:])
        tok = np.empty(int(offs[-1]), dtype=np.int32)
        vals = np.empty(int(offs[-1]), dtype=np.int64) if with_values else None
        for i, (c, j, _) in enumerate(part):
            tok[offs[i]:offs[i + 1]] = pool[c, :lens[i]]
            if with_values:
                vals[offs[i]:offs[i + 1]] = wk["vals"][c, :lens[i]]
        ranks = np.asarray([r for _, _, r in part], dtype=np.int32)
        ix.insert_host(tok, vals, offs, ranks)
    ix.check_device_error()


def c4_batch(torch, wk, pool_d, step, dev):
    """One 64K-request batch on the device: (tokens, offsets, conversation ids)."""
    B = C4["B"]
    g = np.random.default_rng(C4["seed"] * 7919 + 100 + step)
    conv = g.integers(0, wk["pool"].shape[0], size=B)
    lens = wk["req_len"][conv]
    offs = np.zeros(B + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    conv_d = torch.from_numpy(conv).to(dev)
    lens_d = torch.from_numpy(lens).to(dev)
    offs_d = torch.from_numpy(offs).to(dev)
    total = int(offs[-1])
    row = torch.repeat_interleave(conv_d * wk["Lmax"] - offs_d[:-1], lens_d)     # pool index = row base + (flat pos - query start)
    row += torch.arange(total, dtype=torch.int64, device=dev)
    tok = pool_d.view(-1)[row].contiguous()
    del row
    return tok, offs_d, conv, offs


def mesh_route_block(torch, capi, rank, world, steps=4, warmup=2, check_requests=384):
    """north_star (4) on the C4 shape, both arrangements of radixmesh_b200/mesh_router.py:
    sharded    GPU g hosts the tree of the keys rank g inserted; the request batch is broadcast (pipelined in chunks),
               every GPU walks its own tree, matched lengths cross by all-to-all, picks by slice, results all-gathered;
    replicated every GPU holds the ROUTER-mode tree and routes its 1/N of the batch; one all-gather of the results.
    Requests/s, phase times of the sharded form, and a check of a sample against the reference router's rule
    (oracle PortRadixMesh in ROUTER mode on the closure of a few prompts)."""
    import torch.distributed as dist
    from radixmesh_b200.mesh_router import ReplicatedMeshRouter, ShardedMeshRouter
    from radixmesh_b200.router import ConsistentHash
    dev = torch.device("cuda", torch.cuda.current_device())
    P = max(1, world // 2)
    D = world - P
    wk = c4_workload(world)
    B = C4["B"]
    pa = [f"p{i}:1" for i in range(P)]
    da = [f"d{i}:1" for i in range(max(D, 1))]
    rings = (*ConsistentHash(pa).device_arrays(pa), *ConsistentHash(da).device_arrays(da))
    kw = dict(device=dev.index, n_prefill=P, n_decode=D, max_nodes=1 << 19, max_tokens=1 << 25, max_batch=B, max_key_len=wk["Lmax"],
              max_batch_tokens=int(B * float(np.mean(wk["req_len"])) * 1.08))
    t0 = time.perf_counter()
    sh = capi.DeviceIndex(mode=capi.MODE_DECODE, self_rank=rank, **kw)
    sh.set_rings(*rings)
    c4_apply_inserts(sh, wk, c4_inserts(wk, only_rank=rank), with_values=True)
    rp = capi.DeviceIndex(mode=capi.MODE_ROUTER, self_rank=P + D, max_dups=1 << 20, **kw)   # ownership conflicts on shared prompts are booked as duplicates
    rp.set_rings(*rings)
    all_recs = c4_inserts(wk)
    c4_apply_inserts(rp, wk, all_recs, with_values=False)
    sh.compact()
    rp.compact()
    pool_d = torch.from_numpy(wk["pool"]).to(dev)
    batches = [c4_batch(torch, wk, pool_d, s, dev) for s in range(steps + warmup)]
    del pool_d
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    grp_b = dist.new_group(list(range(world))) if world > 1 else None
    sharded = ShardedMeshRouter(sh, rank, world, bcast_group=grp_b)
    replicated = ReplicatedMeshRouter(rp, rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_ms(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- sharded form: only rank 0 holds the batch before the call
    work = [(b[0].clone() if rank == 0 else torch.zeros_like(b[0]), b[1].clone() if rank == 0 else torch.zeros_like(b[1])) for b in batches]
    def time_sharded(fn):
        first = None
        for s in range(warmup):
            fn(work[s][0], work[s][1], B)
        barrier()
        e0, e1 = _events(torch, 2)
        e0.record()
        for s in range(warmup, warmup + steps):
            out = fn(work[s][0], work[s][1], B)
            if s == warmup:
                first = out
        e1.record()
        barrier()
        return max_ms(e0.elapsed_time(e1)) / steps, first

    ms_sh, res_first_sh = time_sharded(sharded.route_pipelined)
    if world > 1:                     # the same batches again (every rank holds them now; the broadcast still runs)
        ms_sh_plain, _ = time_sharded(sharded.route)
    else:
        ms_sh_plain = ms_sh
    # phases of the unpipelined sequence, one extra batch (rank 0's view)
    ph = {}
    if world > 1:
        tok, offs = work[0][0].clone() if rank == 0 else torch.zeros_like(work[0][0]), work[0][1]
        ev = _events(torch, 3)
        barrier()
        ev[0].record()
        dist.broadcast(tok, src=0)
        ev[1].record()
        lens = torch.empty(B, dtype=torch.int32, device=dev)
        nodes = torch.empty(B, dtype=torch.int32, device=dev)
        sh.match(tok, offs, B, lens, nodes)
        ev[2].record()
        torch.cuda.synchronize()
        bytes_b = tok.numel() * 4
        ph = {"broadcast_ms": ev[0].elapsed_time(ev[1]), "walk_ms": ev[1].elapsed_time(ev[2]),
              "broadcast_GBps": bytes_b / (ev[0].elapsed_time(ev[1]) * 1e-3) / 1e9, "request_bytes_per_batch": bytes_b}
        del tok
    del work
    # ---- replicated form: rank r routes requests [r*B/W, (r+1)*B/W)
    Bl = B // world
    shards = []
    for tok, offs_d, _, offs_h in batches:
        a, b = rank * Bl, (rank + 1) * Bl
        shards.append((tok[int(offs_h[a]):int(offs_h[b])], (offs_d[a:b + 1] - int(offs_h[a])).contiguous()))
    for s in range(warmup):
        replicated.route_shard(shards[s][0], shards[s][1], Bl)
    barrier()
    e0, e1 = _events(torch, 2)
    e0.record()
    for s in range(warmup, warmup + steps):
        out = replicated.route_shard(shards[s][0], shards[s][1], Bl)
        if s == warmup:
            res_first_rp = out
    e1.record()
    barrier()
    ms_rp = max_ms(e0.elapsed_time(e1)) / steps
    sh.check_device_error()
    rp.check_device_error()
    if rank != 0:
        return None
    # ---- checks on the first timed batch
    _, _, conv, offs_h = batches[warmup]
    n_rp = world * Bl
    r_pf, r_dc, r_pa, r_da, r_err = [x.cpu().numpy()[:n_rp] for x in res_first_rp]
    s_pf, s_dc, s_pa, s_da, s_err = [x.cpu().numpy()[:n_rp] for x in res_first_sh]
    ok_rows = r_err == 0
    agree = float(((s_pf == r_pf) & (s_dc == r_dc) & (s_pa == r_pa) & (s_da == r_da))[ok_rows].mean()) if ok_rows.any() else None
    # oracle: ROUTER-mode port fed every insert under the rarest prompts (their closure), the requests of batch 0 under them
    from oracle.radix_port import MODE_ROUTER, PortRadixMesh, RankValue, RouterPickError
    counts = np.bincount(wk["pid"], minlength=C4["n_prompts"])
    rare = set(np.argsort(counts)[:6].tolist())
    convs = [c for c in range(len(wk["pid"])) if wk["pid"][c] in rare]
    port = PortRadixMesh(MODE_ROUTER, P + D, P, D)
    for c, j, r in c4_inserts(wk, convs=convs):
        port.apply_insert(wk["pool"][c, :wk["ends"][c, j]].tolist(), RankValue(r), r)
    cset = set(convs)
    idxs = [i for i in range(n_rp) if int(conv[i]) in cset][:check_requests]
    mism = 0
    for i in idxs:
        key = wk["pool"][conv[i], :wk["req_len"][conv[i]]].tolist()
        try:
            pf, dc = port.match_prefix(key)
            exp = (pf, dc, 0)
        except RouterPickError:
            exp = (None, None, 1)
        got = (int(r_pf[i]), int(r_dc[i]), int(r_err[i]))
        if exp[2] == 1:
            mism += got[2] != 1
        else:
            mism += (got[0], got[1]) != (exp[0], exp[1]) or got[2] == 1
    mean_len = float(np.mean(wk["req_len"][conv]))
    return {
        "workload": f"C4 mesh route: {C4['n_prompts']} x {C4['prompt_len']}-token prompts (Zipf), {C4['n_conv']} conversations of 1-8 turns of "
                    f"{C4['turn_lo']}-{C4['turn_hi']} tokens, {B}-request batches (mean {mean_len:.0f} tokens), {P} prefill + {D} decode ranks",
        "n_gpus": world, "steps": steps, "warmup": warmup, "setup_s": round(setup_s, 1),
        "replicated": {"requests_per_s": n_rp / (ms_rp * 1e-3), "ms_per_batch": ms_rp,
                       "what": "ROUTER tree on every GPU, batch sharded, one all-gather of the results (throughput form)"},
        "sharded": {"requests_per_s": B / (min(ms_sh, ms_sh_plain) * 1e-3), "tree_lookups_per_s": world * B / (min(ms_sh, ms_sh_plain) * 1e-3),
                    "ms_per_batch": min(ms_sh, ms_sh_plain), "ms_per_batch_one_broadcast": ms_sh_plain, "ms_per_batch_chunked_broadcast": ms_sh,
                    "what": "tree of rank g's own keys on GPU g (memory-saving form: 1/N of the index per GPU): NCCL broadcast of the "
                            "request tokens, non-splitting walk on every GPU, all-to-all of the matched lengths, pick by 1/N slice, "
                            "one gather; timed with one broadcast per batch and with the broadcast cut into 8 chunks on a "
                            "communicator of their own, each chunk walked as it lands (the better of the two is the headline)",
                    "phases_unpipelined": ph,
                    "cap": "the broadcast of the request tokens (4 B per token to every GPU over NVLink)" if world > 1 else "none (one GPU)"},
        "parity_check": {"requests": len(idxs), "mismatches": int(mism),
                         "oracle": "PortRadixMesh (ROUTER mode) fed every insert under the 6 rarest prompts; replicated form",
                         "sharded_equals_replicated_frac": agree, "error_rows": int((~ok_rows).sum())},
    }


def oplog_exchange_block(torch, capi, rank, world, records=1024, L=512, rounds=4, warmup=1):
    """every rank stages `records` local inserts per round (Zipf-shared prefixes, half of them carrying
    rank-specific values: conflicts for the lower-rank-wins rule), all ranks all-gather the packed records over NCCL
    and apply rank 0's, rank 1's, ... in that order.  Replicas are compared by a digest of their canonical trees."""
    import hashlib
    import torch.distributed as dist
    from radixmesh_b200.mesh_oplog import NvlinkOplogExchange
    from tests.parity import device_digest
    dev = torch.device("cuda", torch.cuda.current_device())
    P = max(1, world // 2)
    D = world - P
    docs = np.random.default_rng(3).integers(0, 32000, (256, L), dtype=np.int32)
    mode = capi.MODE_PREFILL if rank < P else capi.MODE_DECODE
    tot = world * records * (rounds + warmup)
    ix = capi.DeviceIndex(mode=mode, device=dev.index, self_rank=rank, n_prefill=P, n_decode=max(D, 1), max_nodes=4 * tot + (1 << 16),
                          max_tokens=tot * L + (1 << 22), max_batch=4096, max_key_len=max(L, 64), max_dups=2 * tot + 1024)
    ex = NvlinkOplogExchange([ix], rank, world)
    rng = np.random.default_rng(100 + rank)
    w = 1.0 / np.arange(1, 257) ** 1.1
    w /= w.sum()
    times = []
    for rd in range(rounds + warmup):
        d = rng.choice(256, size=records, p=w)
        cut = rng.integers(L // 8, L + 1, size=records)
        keys = rng.integers(0, 32000, (records, L), dtype=np.int32)
        mask = np.arange(L)[None, :] < cut[:, None]
        keys[mask] = docs[d][mask]
        base = (rd * world + rank) * records * L
        for i in range(records):
            v = np.arange(base + i * L, base + (i + 1) * L, dtype=np.int64) if i % 2 else np.arange(L, dtype=np.int64) + int(d[i]) * L
            ex.stage(keys[i], v)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        ex.exchange_and_apply()
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if rd >= warmup:
            times.append(float(dt.item()))
    ix.check_device_error()
    dd = device_digest(ix)
    rk = dd["dump"]["rank"]
    h = hashlib.blake2b(digest_size=16)
    d = dd["dump"]
    for dg, ln, lk, r in sorted((dd["by_id"][x], int(d["edge_len"][x]), int(d["lock_ref"][x]), int(rk[x])) for x in dd["by_id"] if x):
        h.update(dg)
        h.update(np.asarray([ln, lk, r], dtype=np.int64).tobytes())
    s = ix.sizes()
    sig = torch.tensor(list(h.digest()) + [int(s.n_nodes) & 0xFF, int(s.n_dups) & 0xFF], dtype=torch.uint8, device=dev)
    allsig = [torch.zeros_like(sig) for _ in range(world)]
    if world > 1:
        dist.all_gather(allsig, sig)
    else:
        allsig = [sig]
    same = all(bool((a == allsig[0]).all()) for a in allsig)
    ix.close()
    if rank != 0:
        return None
    per = sum(times) / len(times)
    return {"what": "all-gather of packed INSERT records over NCCL + one ordered apply on every replica (mesh_oplog.py)",
            "n_gpus": world, "records_per_round_per_rank": records, "tokens_per_record": L, "rounds": rounds,
            "ms_per_round": 1e3 * per, "records_per_s_applied_per_replica": world * records / per,
            "bytes_gathered_per_round_per_gpu": world * records * L * 12,
            "replicas_identical": same, "nodes": int(s.n_nodes), "dups": int(s.n_dups),
            "check": "digest over every replica's canonical tree (keys, values, boundaries, owner ranks, locks) + node and duplicate counts"}
