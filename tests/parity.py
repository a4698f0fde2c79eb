"""Shared parity helpers: canonical views of a device index and of the oracle, workloads."""
import numpy as np


def device_canonical(ix, with_stamps=True, mesh=False):
    """Canonical dump of a DeviceIndex (same row shape as oracle canonical())."""
    d = ix.dump()
    s = ix.sizes()
    n = d["n"]
    used = int(s.arena_used)
    tok, val = ix.read_arena(0, used) if used else (np.zeros(0, np.int32), np.zeros(0, np.int64))
    full = {0: ()}
    alive = [x for x in range(1, n) if d["parent"][x] >= 0 and d["gen"][x] >= 0]

    def fk(x):
        if x in full:
            return full[x]
        chain = []
        y = x
        while y not in full:
            chain.append(y)
            y = int(d["parent"][y])
        base = full[y]
        for z in reversed(chain):
            o, l = int(d["edge_off"][z]), int(d["edge_len"][z])
            base = base + tuple(int(t) for t in tok[o:o + l])
            full[z] = base
        return full[x]

    check_owner_map(ix, d, used)
    # lazy stamps: the stamp the reference holds for a node is the maximum raw stamp in its subtree
    eff = {x: int(d["stamp"][x]) for x in alive}
    for x in sorted(alive, key=lambda z: -int(d["depth"][z])):
        p = int(d["parent"][x])
        if p > 0:
            eff[p] = max(eff[p], eff[x])
    rows = []
    for x in alive:
        o, l = int(d["edge_off"][x]), int(d["edge_len"][x])
        v = tuple(int(t) for t in val[o:o + l]) if val is not None else ()
        row = (fk(x), l, v, int(d["rank"][x]), int(d["lock_ref"][x]))
        if not mesh:
            row = row + (eff[x] >> 1,)   # 2*op (+1): compare the operation only
        rows.append(row)
    rows.sort()
    out = {"nodes": rows, "evictable": int(s.evictable), "protected": int(s.protected_)}
    out["_ids"] = {fk(x): x for x in alive}
    out["_dump"] = d
    return out


CONT = 0x40000000


def check_owner_map(ix, d, used):
    """Invariants of the per-position owner map (csrc/rm_common.cuh RM_SEG_CONT): every live run (tree nodes and
    parked duplicates) is owned by its node, the contiguity flag sits only on the first position of a node
    whose parent's run ends right there, and everything else is free."""
    if not used:
        return
    seg = ix.read_segments(0, used)
    n = d["n"]
    covered = np.zeros(used, dtype=bool)
    for x in range(1, n):
        live = (d["parent"][x] >= 0 or d["gen"][x] < 0) and d["edge_len"][x] > 0
        if not live:
            continue
        o, l = int(d["edge_off"][x]), int(d["edge_len"][x])
        run = seg[o:o + l]
        assert np.all((run & 0x3FFFFFFF) == x) and np.all(run >= 0), f"owner map of node {x}: {run.tolist()}"
        assert not np.any(run[1:] & CONT), f"contiguity flag inside node {x}"
        if run[0] & CONT:
            p = int(d["parent"][x])
            assert d["gen"][x] >= 0 and p > 0 and int(d["edge_off"][p]) + int(d["edge_len"][p]) == o, \
                f"node {x} is flagged contiguous but its parent {p} does not end at {o}"
        covered[o:o + l] = True
    assert np.all(seg[~covered] == -1), "free arena positions must be unowned"


def rank_stamps(rows):
    """Replace the stamp column (last) by its dense rank: only the order matters (eviction)."""
    if rows and rows[0][-1] >= (1 << 20):          # oracle stamps: (op << 20) + k
        rows = [r[:-1] + (r[-1] >> 20,) for r in rows]
    stamps = sorted({r[-1] for r in rows})
    pos = {s: i for i, s in enumerate(stamps)}
    return [r[:-1] + (pos[r[-1]],) for r in rows]


def assert_same_tree(dev, orc, stamps=True):
    a, b = dev["nodes"], orc["nodes"]
    if stamps:
        a, b = rank_stamps(a), rank_stamps(b)
    else:
        a, b = [r[:5] for r in a], [r[:5] for r in b]
    assert len(a) == len(b), f"node count {len(a)} != {len(b)}"
    for x, y in zip(a, b):
        assert x == y, f"node differs:\n dev {x}\n orc {y}"
    assert dev["evictable"] == orc["evictable"], (dev["evictable"], orc["evictable"])
    assert dev["protected"] == orc["protected"], (dev["protected"], orc["protected"])


def pack(keys, dtype=np.int32):
    """list of lists -> (flat, offsets[int64])"""
    offs = np.zeros(len(keys) + 1, dtype=np.int64)
    for i, k in enumerate(keys):
        offs[i + 1] = offs[i] + len(k)
    flat = np.zeros(int(offs[-1]), dtype=dtype)
    for i, k in enumerate(keys):
        flat[offs[i]:offs[i + 1]] = k
    return flat, offs


def zipf_prefix(seed, n_docs, n_seq, L, s=1.1, vocab=32000, docs=None):
    """SURVEY.md 8d generator: sequences = docs[d][:cut] ++ fresh random tail, d ~ Zipf(s)."""
    rng = np.random.default_rng(seed)
    if docs is None:
        docs = np.random.default_rng(seed ^ 0x5EED).integers(0, vocab, (n_docs, L), dtype=np.int32)
    w = 1.0 / np.arange(1, n_docs + 1) ** s
    w /= w.sum()
    d = rng.choice(n_docs, size=n_seq, p=w)
    cut = rng.integers(L // 8, L + 1, size=n_seq)
    out = rng.integers(0, vocab, (n_seq, L), dtype=np.int32)
    cols = np.arange(L)[None, :]
    mask = cols < cut[:, None]
    out[mask] = docs[d][mask]
    return out, docs


# ---- large trees: digests instead of full-key tuples ---------------------------------------------------------------
# A node's digest chains its parent's digest with its own key and value bytes, so two trees have the same multiset
# of (digest, edge length, lock_ref, stamp order) rows iff they have the same nodes with the same boundaries, keys,
# values, locks and eviction order.  Costs O(tokens) instead of O(nodes * depth).
def _dg(parent_digest: bytes, tok: np.ndarray, val) -> bytes:
    import hashlib
    h = hashlib.blake2b(parent_digest, digest_size=16)
    h.update(np.ascontiguousarray(tok, dtype=np.int32).tobytes())
    if val is not None:
        h.update(np.ascontiguousarray(val, dtype=np.int64).tobytes())
    return h.digest()


def device_digest(ix):
    """{"rows": sorted (digest, len, lock_ref, effective stamp >> 1), "by_id": {node id: digest}, sizes} of a DeviceIndex."""
    d = ix.dump()
    s = ix.sizes()
    n = d["n"]
    used = int(s.arena_used)
    tok, val = ix.read_arena(0, used) if used else (np.zeros(0, np.int32), np.zeros(0, np.int64))
    parent, off, ln, depth = d["parent"][:n], d["edge_off"][:n], d["edge_len"][:n], d["depth"][:n]
    alive = np.nonzero((parent >= 0) & (d["gen"][:n] >= 0))[0]
    alive = alive[alive != 0]
    order = alive[np.argsort(depth[alive], kind="stable")]        # parents before children
    dig = {0: b""}
    for x in order.tolist():
        o, l = int(off[x]), int(ln[x])
        dig[x] = _dg(dig[int(parent[x])], tok[o:o + l], val[o:o + l] if val is not None else None)
    eff = d["stamp"][:n].astype(np.uint64).copy()
    for x in order[::-1].tolist():                                 # children before parents
        p = int(parent[x])
        if p > 0 and eff[x] > eff[p]:
            eff[p] = eff[x]
    rows = sorted((dig[x], int(ln[x]), int(d["lock_ref"][x]), int(eff[x]) >> 1) for x in order.tolist())
    return {"rows": rows, "by_id": dig, "evictable": int(s.evictable), "protected": int(s.protected_),
            "val": val, "dump": d}


def oracle_digest(orc):
    """The same rows for a PortRadixCache, plus {Node: digest}."""
    from oracle.radix_port import SUB_BITS
    dig = {orc.root: b""}
    rows = []
    stack = [orc.root]
    while stack:
        nd = stack.pop()
        for c in nd.children.values():
            dig[c] = _dg(dig[nd], np.asarray(c.key, dtype=np.int32), np.asarray(c.value, dtype=np.int64))
            rows.append((dig[c], len(c.key), c.lock_ref, c.stamp >> SUB_BITS))
            stack.append(c)
    rows.sort()
    return {"rows": rows, "by_node": dig, "evictable": orc.evictable, "protected": orc.protected}


def assert_same_digest(dev, orc):
    a, b = dev["rows"], orc["rows"]
    assert len(a) == len(b), f"node count {len(a)} != {len(b)}"
    # stamps: only the order matters; nodes compare in (digest) order, stamps by dense rank
    ra = {s: i for i, s in enumerate(sorted({r[3] for r in a}))}
    rb = {s: i for i, s in enumerate(sorted({r[3] for r in b}))}
    for x, y in zip(a, b):
        assert x[:3] == y[:3], f"node differs: dev {x[0].hex()} len {x[1]} lock {x[2]} / orc {y[0].hex()} len {y[1]} lock {y[2]}"
        assert ra[x[3]] == rb[y[3]], f"stamp order differs at node {x[0].hex()}: {ra[x[3]]} vs {rb[y[3]]}"
    assert dev["evictable"] == orc["evictable"] and dev["protected"] == orc["protected"]
