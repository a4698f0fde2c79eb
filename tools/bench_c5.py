"""BASELINE configs[4] at full tree size: 10 M-node tree, 4K-token agentic queries, one B200 (opt-in: minutes of set-up).

    python tools/bench_c5.py [--nodes 10000000] [--steps 6] [--warmup 3]

Generator (stated, because SURVEY.md 8d's fills every insert to L = 4096 tokens - 29 KB of arenas per insert, 144 GB for
the 5 M inserts a 10 M-node tree takes): an agent trajectory keeps extending a long shared context by short steps.
  docs      65,536 trajectories x 4096 tokens over token ids [0, 24000)
  insert    docs[d][:cut] ++ a fresh tail of 64-256 tokens over [24000, 32000);  d ~ Zipf(1.1), cut ~ U{512..4096}
  query     docs[d][:cut] ++ tokens >= 32000 up to 4096 tokens: same Zipf, same cut law
Tails and trajectories use disjoint token ranges and query padding is outside the vocabulary, so the matched length of
every query is known in closed form: min(cut, longest cut inserted for d).  That is the parity check, at full size, on
every query of every timed batch (a 10 M-node Python tree is ~10 GB and an hour to build).  Everything is generated on
the device (80 GB of insert keys would not fit a host generator in minutes).  Prints one JSON line.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from radixmesh_b200 import capi  # noqa: E402

N_DOCS, L, V_DOC, V_TAIL, NB = 65536, 4096, 24000, 32000, 4096


def zipf_sampler(n, s, dev):
    w = 1.0 / torch.arange(1, n + 1, dtype=torch.float64, device=dev) ** s
    return torch.cumsum(w / w.sum(), 0)


def ragged(lens, dev):
    offs = torch.zeros(lens.numel() + 1, dtype=torch.int64, device=dev)
    torch.cumsum(lens, 0, out=offs[1:])
    total = int(offs[-1].item())
    row = torch.repeat_interleave(torch.arange(lens.numel(), device=dev), lens)
    pos = torch.arange(total, device=dev) - offs[:-1][row]
    return offs, row, pos, total


def run(nodes=10_000_000, steps=6, warmup=3, B=16384, peak=None, seed=3):
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    docs = torch.randint(0, V_DOC, (N_DOCS, L), dtype=torch.int32, device=dev, generator=g)
    cdf = zipf_sampler(N_DOCS, 1.1, dev)
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, device=dev.index, max_nodes=int(nodes * 1.15) + (steps + warmup + 1) * B + (1 << 16),
                          max_tokens=int(nodes * 152) + (1 << 24), max_batch=B, max_key_len=L + 256)   # measured: ~142 arena tokens per node
    maxcut = torch.zeros(N_DOCS, dtype=torch.int64, device=dev)
    out = torch.empty(NB, dtype=torch.int32, device=dev)
    t0 = time.perf_counter()
    n_ins, vc, ins_s = 0, 0, 0.0
    while True:
        d = torch.searchsorted(cdf, torch.rand(NB, dtype=torch.float64, device=dev, generator=g)).clamp_(max=N_DOCS - 1)
        cut = torch.randint(512, L + 1, (NB,), device=dev, generator=g)
        tail = torch.randint(64, 257, (NB,), device=dev, generator=g)
        offs, row, pos, total = ragged(cut + tail, dev)
        tok = torch.randint(V_DOC, V_TAIL, (total,), dtype=torch.int32, device=dev, generator=g)
        in_doc = pos < cut[row]
        tok[in_doc] = docs.view(-1)[(d[row] * L + pos)[in_doc]]
        val = torch.arange(vc, vc + total, dtype=torch.int64, device=dev)
        vc += total
        del row, pos, in_doc
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        ix.insert(tok, val, offs, None, NB, out)
        torch.cuda.synchronize()
        ins_s += time.perf_counter() - t1
        maxcut.scatter_reduce_(0, d, cut, reduce="amax")
        n_ins += NB
        if n_ins % (16 * NB) == 0 or n_ins == NB:
            sz = ix.sizes()
            if sz.device_error:
                raise RuntimeError(f"device error {sz.device_error} after {n_ins} inserts: nodes {sz.n_nodes} (hwm {sz.node_hwm}, pool "
                                   f"{ix.cfg.max_nodes}), arena {sz.arena_used} of {ix.cfg.max_tokens}")
            if sz.n_nodes >= nodes:
                break
    ix.compact()
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    s0 = ix.sizes()
    # ---- query batches
    qb = []
    for s in range(steps + warmup):
        d = torch.searchsorted(cdf, torch.rand(B, dtype=torch.float64, device=dev, generator=g)).clamp_(max=N_DOCS - 1)
        cut = torch.randint(512, L + 1, (B,), device=dev, generator=g)
        q = torch.full((B, L), 40000, dtype=torch.int32, device=dev)
        cols = torch.arange(L, device=dev)[None, :]
        m = cols < cut[:, None]
        q[m] = docs[d][m]
        qb.append((q.view(-1), d, cut))
    offs_q = torch.arange(0, (B + 1) * L, L, dtype=torch.int64, device=dev)
    out_len = torch.empty((steps + warmup, B), dtype=torch.int32, device=dev)
    out_node = torch.empty((steps + warmup, B), dtype=torch.int32, device=dev)
    for s in range(warmup):
        ix.match(qb[s][0], offs_q, B, out_len[s], out_node[s])
    torch.cuda.synchronize()
    ix.profile(2)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(warmup, warmup + steps):
        ix.match(qb[s][0], offs_q, B, out_len[s], out_node[s])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    stage_ms, nprof = ix.profile_stages()
    ix.profile_read()
    ix.profile(False)
    ix.check_device_error()
    walk_ms, split_ms = stage_ms[0] / max(nprof, 1), stage_ms[1] / max(nprof, 1)
    # ---- parity: closed-form matched length on every query of every batch (timed and warm-up)
    checked = mism = 0
    for s in range(steps + warmup):
        _, d, cut = qb[s]
        mc = maxcut[d]
        known = mc > 0
        exp = torch.minimum(cut, mc)
        checked += int(known.sum().item())
        mism += int(((out_len[s].to(torch.int64) != exp) & known).sum().item())
    # ---- 8d byte model: m in closed form, hops by climbing the device tree from each last node
    dmp = ix.dump()
    parent = dmp["parent"].astype(np.int64)
    cur = out_node[warmup].cpu().numpy().astype(np.int64)
    hops = np.zeros(B, dtype=np.int64)
    while True:
        live = cur > 0
        if not live.any():
            break
        hops += live
        cur = np.where(live, parent[np.maximum(cur, 0)], 0)
    m = out_len[warmup].cpu().numpy().astype(np.int64)
    bpl = float((8 * np.minimum(L, m + 1) + 40 * hops + 8).mean())
    s1 = ix.sizes()
    block = {
        "workload": f"C5: {int(s0.n_nodes)}-node tree, {B} queries x {L} tokens per batch, agentic generator (tools/bench_c5.py)",
        "value": B / (ms * 1e-3), "unit": "lookups/s", "ms_per_step": ms, "steps": steps, "warmup": warmup,
        "tree_nodes_start": int(s0.n_nodes), "tree_nodes_end": int(s1.n_nodes), "arena_tokens": int(s0.arena_used),
        "inserts_to_build": n_ins, "build_s": round(build_s, 1), "inserts_per_s_build": n_ins / ins_s,
        "mean_matched_len": float(m.mean()), "matched_tokens_per_s": B / (ms * 1e-3) * float(m.mean()),
        "layout": "rm_compact after the build (heavy paths contiguous)",
        "stages_ms_per_step": {"walk": walk_ms, "split_stage": split_ms},
        "parity_check": {"queries": checked, "mismatches": mism,
                         "oracle": "closed form min(cut, longest inserted cut of the trajectory): disjoint token ranges make it exact"},
    }
    if peak:
        ach = bpl * B / (walk_ms * 1e-3) / 1e9
        block["roofline"] = {"bound": "hbm", "kernel": "rm_match_rung_kernel (K1 walk)", "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "bytes_per_lookup": bpl, "mean_hops": float(hops.mean()), "kernel_ms": walk_ms,
                             "hops_source": "device tree"}
    ix.close()
    return block


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = float(json.load(f)["hbm_gbs"])
    except Exception:
        pk = 6650.0
    print(json.dumps(run(a.nodes, a.steps, a.warmup, peak=pk)))
