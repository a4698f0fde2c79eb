"""Headline benchmark: batched prefix-match lookups/s on the device-resident radix index.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--with-c5] [--no-blocks]

One "step" = one pass of the hot path (rm_match_batch: walk kernel + split stage, two launches)
over one batch of synthetic Zipf shared-prefix queries (SURVEY.md 8d generator).  Every step
gets a batch the tree has not seen before, so the splitting work the reference does on first sight
of a prefix (radix_cache.py:262-266) is inside the timed region.

  value      lookups/s, whole job, query batches already resident in HBM
  e2e        the same metric through rm_match_batch_h with pinned HOST buffers (H2D of the tokens
             and D2H of both result arrays inside the timed region)
  roofline   walk kernel (K1): algorithmic bytes per launch / its mean CUDA-event duration
  cpu_baseline  the oracle port (the reference's algorithm and data structures, pure Python like
             the reference) timed on the host cores on a bounded sample of the same workload

N > 1 (torchrun): every rank hosts a replica of the tree and its own shard of query batches; no
data-path collective (weak scaling); time = max over ranks.

The same line carries secondary blocks (tools/bench_blocks.py), each with its own parity check - the line is not
printed if one fails: `parity_check` (first timed batch against the CPU port), `target` (1 M-node tree, 64 K x 512
tokens), `c3` (streaming insert + LRU evict at 90 % occupancy), `mesh_route` (the C4 routing path on N GPUs, replicated
and sharded forms), `oplog_exchange` (packed INSERT records all-gathered and applied on every replica), and with
--with-c5 `c5` (10 M-node tree, 4,096-token queries).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1]
    "c2": dict(name="C2 single-tree batched match_prefix: 100K-node tree, 64K queries x 256 tokens, Zipf s=1.1 prefix sharing",
               n_docs=4096, L=256, tree_nodes=100_000, B=65536, seed=0),
    # north_star target shape
    "target": dict(name="north_star target: 1M-node tree, 64K queries x 512 tokens, Zipf s=1.1 prefix sharing",
                   n_docs=32768, L=512, tree_nodes=1_000_000, B=65536, seed=1),
    # BASELINE configs[4] at reduced size: 4K-token queries.  The full shape (10 M-node trees) does not fit this
    # generator: every insert appends ~1,800 fresh tokens (29 KB of arenas), i.e. 144 GB for 5 M inserts.
    "c5r": dict(name="C5 reduced: 0.5M-node tree, 16K queries x 4096 tokens, Zipf s=1.1 prefix sharing",
                n_docs=8192, L=4096, tree_nodes=500_000, B=16384, seed=3),
}
INSERT_BATCH = 4096


def gen_batch(seed, n, L, docs, s=1.1, vocab=32000):
    """sequence = docs[d][:cut] ++ fresh random tail; d ~ Zipf(s); cut ~ U{L/8..L}."""
    rng = np.random.default_rng(seed)
    n_docs = docs.shape[0]
    w = 1.0 / np.arange(1, n_docs + 1) ** s
    w /= w.sum()
    d = rng.choice(n_docs, size=n, p=w)
    cut = rng.integers(L // 8, L + 1, size=n)
    out = rng.integers(0, vocab, (n, L), dtype=np.int32)
    mask = np.arange(L)[None, :] < cut[:, None]
    out[mask] = docs[d][mask]
    return out


def gen_docs(wl):
    return np.random.default_rng(wl["seed"] ^ 0x5EED).integers(0, 32000, (wl["n_docs"], wl["L"]), dtype=np.int32)


def insert_stream(wl, docs):
    """Deterministic stream of insert batches (same for the CUDA path and the CPU baseline)."""
    i = 0
    while True:
        yield gen_batch(wl["seed"] * 1000 + 17 + i, INSERT_BATCH, wl["L"], docs)
        i += 1


def query_seed(wl, rank, kind, step):
    return wl["seed"] * 100000 + 50000 + rank * 10007 + kind * 1009 + step


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False
        self.proc = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc is not None:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 3 + k and r[3 + k].lower() == "active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def build_oracle_tree(wl, docs, n_nodes):
    """The oracle fed the same insert stream until it holds >= n_nodes nodes."""
    from oracle.radix_port import PortRadixCache
    t = PortRadixCache(1)
    L = wl["L"]
    count = 0
    vc = 0
    for batch in insert_stream(wl, docs):
        for row in batch:
            t.insert(row.tolist(), np.arange(vc, vc + L, dtype=np.int64))
            vc += L
        count += len(batch)
        n = sum(1 for _ in t.iter_nodes())
        if n >= n_nodes:
            return t, count


def cpu_sample(wl, docs, n_queries, rank=0, step=0):
    """(lookups/s of the oracle on one core, mean algorithmic bytes per lookup, mean hops, mean match, matched lengths)
    on the first n_queries queries of device batch `step`."""
    t, _ = build_oracle_tree(wl, docs, wl["tree_nodes"])
    qs = gen_batch(query_seed(wl, rank, 0, step), wl["B"], wl["L"], docs)[:n_queries]
    lists = [r.tolist() for r in qs]
    L = wl["L"]
    bytes_tot = hops_tot = m_tot = 0
    t0 = time.perf_counter()
    res = []
    for k in lists:
        h0 = t.hops
        m = len(t.match_prefix(k).indices)
        res.append((m, t.hops - h0))
    dt = time.perf_counter() - t0
    for m, h in res:
        bytes_tot += 8 * min(L, m + 1) + 40 * h + 8
        hops_tot += h
        m_tot += m
    n = len(lists)
    return n / dt, bytes_tot / n, hops_tot / n, m_tot / n, np.asarray([m for m, _ in res], dtype=np.int32)


def _ref_worker(args):
    wl, n_queries, shard, steps, warmup = args
    docs = gen_docs(wl)
    t, _ = build_oracle_tree(wl, docs, wl["tree_nodes"])
    times = []
    for s in range(steps + warmup):
        qs = gen_batch(query_seed(wl, shard, 2, s), n_queries, wl["L"], docs)
        lists = [r.tolist() for r in qs]
        t0 = time.perf_counter()
        for k in lists:
            t.match_prefix(k)
        times.append(time.perf_counter() - t0)
    return times[warmup:]


def run_reference(args, wl):
    """Reference arm: the oracle port (the reference is pure Python; the port keeps its algorithm and
    data structures) on all host cores, one tree replica per process, query stream sharded."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = max(1, (os.cpu_count() or 1))
    procs = min(cores, 64)
    per_step = 512            # queries per process per step: a bounded sample of the 64K-query batch
    steps, warmup = args.steps, args.warmup
    ctx = mp.get_context("fork")
    with ctx.Pool(procs) as pool:
        res = pool.map(_ref_worker, [(wl, per_step, i, steps, warmup) for i in range(procs)])
    step_times = [max(r[s] for r in res) for s in range(steps)]
    total = sum(step_times)
    value = procs * per_step * steps / total
    line = {"metric": "prefix-match lookups/sec", "value": value, "unit": "lookups/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * total / steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic", "impl": "reference",
            "config": {"workload": wl["name"], "tree_nodes": wl["tree_nodes"], "query_len": wl["L"]},
            "cpu_baseline": {"value": value, "unit": "lookups/s", "cores": procs, "kind": "port",
                             "sample": f"{procs} processes x {per_step} queries per step of the same Zipf query stream, "
                                       f"each on its own replica of the {wl['tree_nodes']}-node tree"},
            "e2e": {"value": value, "unit": "lookups/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=4096, help="queries timed on the CPU oracle (rank 0, N=1)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--stages", action="store_true", help="also time the split stage (one more CUDA event per step)")
    ap.add_argument("--no-compact", action="store_true", help="skip rm_compact after the build (A/B of the arena layout)")
    ap.add_argument("--no-blocks", action="store_true", help="headline only: no target / c3 / mesh_route blocks")
    ap.add_argument("--with-c5", action="store_true", help="also run the 10 M-node / 4K-token block (tools/bench_c5.py; ~1 min more)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.workload != "c2":
        args.no_cpu = True     # the port cannot hold these trees (minutes to hours): the default run's `target` block and
                               # tools/bench_c5.py carry the parity checks and CPU samples for these shapes
    if args.impl == "reference":
        run_reference(args, wl)
        return

    import torch
    import torch.distributed as dist
    from radixmesh_b200 import capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    K, W, B, L = args.steps, max(args.warmup, 0), wl["B"], wl["L"]
    assert W >= 3 or os.environ.get("RM_BENCH_ALLOW_SHORT_WARMUP"), "timing rules: at least 3 warm-up steps"

    n_batches = 2 * (K + W)
    max_nodes = int(wl["tree_nodes"] * 1.3) + n_batches * (B // 2) + (1 << 16)   # ~15 % of first-sight queries add a node
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, device=local, max_nodes=max_nodes,
                          max_tokens=int(wl["tree_nodes"] * L * 0.9) + (1 << 22), max_batch=B, max_key_len=max(L, 1024),
                          max_batch_tokens=B * L)      # staging of the host-buffer entry point sized once, at rm_create
    docs = gen_docs(wl)

    # ---- build the tree on the device (untimed) -----------------------------------------------
    offs_ins = torch.arange(0, (INSERT_BATCH + 1) * L, L, dtype=torch.int64, device=dev)
    out_ins = torch.empty(INSERT_BATCH, dtype=torch.int32, device=dev)
    vc = 0
    n_ins = 0
    t_build0 = time.perf_counter()
    for batch in insert_stream(wl, docs):
        tok = torch.from_numpy(batch.reshape(-1)).to(dev)
        val = torch.arange(vc, vc + INSERT_BATCH * L, dtype=torch.int64, device=dev)
        ix.insert(tok, val, offs_ins, None, INSERT_BATCH, out_ins)
        vc += INSERT_BATCH * L
        n_ins += INSERT_BATCH
        if ix.sizes().n_nodes >= wl["tree_nodes"]:
            break
    if not args.no_compact:
        ix.compact()       # index maintenance, untimed: reclaims nothing here, lays heavy paths out contiguously (csrc/rm_maint.cu)
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t_build0
    ix.check_device_error()
    nodes_start = int(ix.sizes().n_nodes)

    # ---- query batches: device-resident for `value`, pinned host for `e2e` ---------------------
    offs_q = torch.arange(0, (B + 1) * L, L, dtype=torch.int64, device=dev)
    dev_batches = [torch.from_numpy(gen_batch(query_seed(wl, rank, 0, s), B, L, docs).reshape(-1)).to(dev)
                   for s in range(K + W)]
    host_batches = [torch.from_numpy(gen_batch(query_seed(wl, rank, 1, s), B, L, docs).reshape(-1)).pin_memory()
                    for s in range(K + W)]
    offs_h = torch.arange(0, (B + 1) * L, L, dtype=torch.int64).pin_memory()
    out_lens = torch.empty((K, B), dtype=torch.int32, device=dev)   # every timed step keeps its result (parity_check reads step 0)
    out_len = out_lens[K - 1]
    out_node = torch.empty(B, dtype=torch.int32, device=dev)
    out_len_h = torch.empty(B, dtype=torch.int32).pin_memory()
    out_node_h = torch.empty(B, dtype=torch.int32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing ---------------------------------------------------------------
    for s in range(W):
        ix.match(dev_batches[s], offs_q, B, out_len, out_node)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    ix.profile(2 if args.stages else 1)
    launches0 = ix.lib.rm_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    matched_tokens = torch.zeros((), dtype=torch.int64, device=dev)
    barrier()
    ev0.record()
    for s in range(W, W + K):
        ix.match(dev_batches[s], offs_q, B, out_lens[s - W], out_node)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = int(ix.lib.rm_launch_count() - launches0)
    stage_ms, _ = ix.profile_stages() if args.stages else (None, 0)
    walk_ms, walk_n = ix.profile_read()
    ix.profile(False)
    clocks = sampler.finish()
    ix.check_device_error()

    # ---- end to end through the host-buffer C ABI ----------------------------------------------
    for s in range(W):
        ix.match_host(host_batches[s].numpy(), offs_h.numpy(), out_len_h.numpy(), out_node_h.numpy())
    barrier()
    t0 = time.perf_counter()
    for s in range(W, W + K):
        ix.match_host(host_batches[s].numpy(), offs_h.numpy(), out_len_h.numpy(), out_node_h.numpy())
    barrier()
    e2e_s = time.perf_counter() - t0
    ix.check_device_error()

    t_ms = torch.tensor([ms, e2e_s * 1e3, walk_ms / max(walk_n, 1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms, e2e_ms, walk_ms_mean = [float(x) for x in t_ms.cpu()]

    # ---- secondary blocks (tools/bench_blocks.py): every rank takes part in the collective ones
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_blocks as blocks
    h2d_gbs = blocks.h2d_probe(torch, world)
    mean_match = float(out_len.to(torch.float64).mean().item())
    nodes_end = int(ix.sizes().n_nodes)
    first_lens = out_lens[0, :args.cpu_sample].cpu().numpy()
    ix.close()
    del dev_batches, host_batches
    torch.cuda.empty_cache()
    mesh = oplog = None
    if not args.no_blocks and args.workload == "c2":
        mesh = blocks.mesh_route_block(torch, capi, rank, world)
        oplog = blocks.oplog_exchange_block(torch, capi, rank, world)

    if rank == 0:
        value = world * B * K / (ms * 1e-3)
        e2e = world * B * K / (e2e_ms * 1e-3)
        peak, peak_src = measured_peak()
        line = {"metric": "prefix-match lookups/sec", "value": value, "unit": "lookups/s", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                "config": {"workload": wl["name"], "tree_nodes_start": nodes_start, "tree_nodes_end": nodes_end,
                           "layout": "bump-allocated arenas as built" if args.no_compact else "rm_compact after the build (heavy paths contiguous)",
                           "inserts_to_build": n_ins, "build_s": round(build_s, 2), "batch": B, "query_len": L,
                           "mode": "RadixCache (splitting match)", "walk": os.environ.get("RM_WALK", "run"), "lanes_per_query": int(os.environ.get("RM_RUN_G", "32")),
                           "l2": f"every step reads a fresh {B * L * 4 >> 20} MiB query batch ({(K + W) * B * L * 4 >> 20} MiB "
                                 "of distinct batches > 126 MB L2); the tree stays warm as in steady-state serving",
                           "parallelism": f"replica x{world}, query batches sharded, no collective"},
                "matched_tokens_per_s": value * mean_match, "mean_matched_len": mean_match,
                "e2e": {"value": e2e, "unit": "lookups/s", "h2d_bytes_per_step": B * L * 4 + (B + 1) * 8,
                        "d2h_bytes_per_step": B * 8,
                        "h2d_GBps": world * (B * L * 4 + (B + 1) * 8) / (e2e_ms * 1e-3 / K) / 1e9,
                        "h2d_ceiling_GBps": h2d_gbs, "frac_of_h2d_ceiling": world * (B * L * 4 + (B + 1) * 8) / (e2e_ms * 1e-3 / K) / 1e9 / h2d_gbs,
                        "h2d_ceiling_how": f"pinned 256 MiB cudaMemcpyAsync host-to-device on all {world} GPU(s) at once, aggregate"},
                "gpu_launches": launches, "clocks": clocks}
        if stage_ms:
            line["stages_ms_per_step"] = dict(zip(("walk", "split_stage"), (x / max(walk_n, 1) for x in stage_ms[:2])))
        cpu = None
        if world == 1 and not args.no_cpu:
            v, bpl, hops, mm, cpu_lens = cpu_sample(wl, docs, args.cpu_sample, rank, W)
            got = first_lens
            mism = int((got != cpu_lens).sum())
            line["parity_check"] = {"queries": int(len(cpu_lens)), "mismatches": mism,
                                    "what": "matched lengths of the first timed batch against the CPU port on its replica of the tree"}
            if mism:
                sys.stderr.write(json.dumps({"error": "parity_check failed: no value reported", "parity_check": line["parity_check"]}) + "\n")
                raise SystemExit(3)
            cpu = {"value": v, "unit": "lookups/s", "cores": 1, "kind": "port",
                   "sample": f"first {args.cpu_sample} queries of the first timed batch on the oracle's replica of the tree "
                             f"(pure-Python port of the reference's RadixCache, like the reference itself)"}
            line["cpu_baseline"] = cpu
            achieved = bpl * B / (walk_ms_mean * 1e-3) / 1e9
            traffic = None
            traffic_src = None
            try:
                with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                    tj = json.load(f).get(args.workload, {})
                    traffic = tj.get("walk_kernel_dram_bytes_per_launch")
                    traffic_src = "profiles/traffic.json <- " + str(tj.get("source"))
            except Exception:
                pass
            line["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                                "traffic": traffic, "traffic_source": traffic_src, "kernel": "rm_match_rung_kernel (K1 walk)", "peak_source": peak_src,
                                "bytes_per_lookup": bpl, "mean_hops": hops, "kernel_ms": walk_ms_mean,
                                "kernel_share_of_step": walk_ms_mean / (ms / K),
                                "lookups_per_s_kernel_only": B / (walk_ms_mean * 1e-3),
                                "dram_achieved": (traffic / (walk_ms_mean * 1e-3) / 1e9) if traffic else None,
                                "dram_frac": (traffic / (walk_ms_mean * 1e-3) / 1e9 / peak) if traffic else None,
                                "note": "achieved = SURVEY 8d algorithmic bytes (8 B per compared position + 40 B of node "
                                        "metadata per hop + 8 B result) / kernel time; the walk follows contiguous arena runs "
                                        "and reads a 4-byte owner word per position instead of node records, and the tree is "
                                        "largely L2 resident, so DRAM traffic (traffic, dram_achieved, dram_frac; ncu) is well "
                                        "below the algorithmic figure and frac can pass 1"}
        if mesh is not None:
            line["mesh_route"] = mesh
            if mesh["parity_check"]["mismatches"]:
                sys.stderr.write(json.dumps({"error": "mesh_route parity_check failed", "parity_check": mesh["parity_check"]}) + "\n")
                raise SystemExit(3)
        if oplog is not None:
            line["oplog_exchange"] = oplog
        if world == 1 and not args.no_blocks and not args.no_cpu and args.workload == "c2":
            tb, tix, tstate = blocks.target_block(sys.modules[__name__], torch, capi, peak)
            line["target"] = tb
            if tb["parity_check"]["mismatches"]:
                sys.stderr.write(json.dumps({"error": "target parity_check failed", "parity_check": tb["parity_check"]}) + "\n")
                raise SystemExit(3)
            hpt = tb["roofline"]["mean_hops"] / max(tb["mean_matched_len"], 1.0)
            line["c3"] = blocks.c3_block(sys.modules[__name__], torch, capi, peak, tix, tstate, hpt)
            tix.close()
        if world == 1 and args.with_c5:
            import bench_c5
            line["c5"] = bench_c5.run(peak=peak)
            if line["c5"]["parity_check"]["mismatches"]:
                sys.stderr.write(json.dumps({"error": "c5 parity_check failed", "parity_check": line["c5"]["parity_check"]}) + "\n")
                raise SystemExit(3)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
