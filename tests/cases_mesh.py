"""Mesh scenarios shared by the host-emulation tests (CPU) and the GPU parity tests: a 3P+2D+1R mesh of
device indexes next to the oracle's PortMesh, fed the same INSERT stream in ring order."""
import random

import numpy as np

from oracle.radix_port import MODE_DECODE, MODE_PREFILL, MODE_ROUTER, PortConsistentHash, PortMesh, RouterPickError
from tests.parity import device_canonical, pack

P, D, R = 3, 2, 1


def mesh_canonical(ix):
    c = device_canonical(ix, mesh=True)
    inv = {v: k for k, v in c["_ids"].items()}
    inv[0] = ()
    dup = {}
    for node, rank, parked, lk in ix.dup_entries().tolist():
        dup[(inv[node], rank)] = None if parked < 0 else lk
    c["dup"] = dup
    c["_inv"] = inv
    return c


def rings(addrs):
    h = PortConsistentHash(addrs)
    pts = np.asarray(h.ring, dtype=np.uint32)
    own = np.asarray([addrs.index(h.owner[p]) for p in h.ring], dtype=np.int32)
    return pts, own


def run_mesh_sequence(make_index, seed, n_steps=25, vocab=3, maxlen=7, max_recs=5, max_batch=64, docs=0):
    """docs > 0: keys are prefixes of `docs` fixed random strings of maxlen tokens (plus a short tail now and then), so
    that large batches pile onto a few long shared paths instead of spreading over a bushy tree."""
    rnd = random.Random(seed)
    port = PortMesh(P, D, R)
    ixs = []
    big = max_batch > 64
    doc_keys = [[rnd.randrange(vocab) for _ in range(maxlen)] for _ in range(docs)]

    def new_key():
        if not docs:
            return [rnd.randrange(vocab) for _ in range(rnd.randrange(1, maxlen))]
        d = rnd.choice(doc_keys)
        k = d[:rnd.randrange(1, maxlen + 1)]
        if rnd.random() < 0.3:
            k = k + [rnd.randrange(vocab, vocab + 2) for _ in range(rnd.randrange(1, 4))]
        return k

    for r in range(P + D + R):
        mode = MODE_PREFILL if r < P else MODE_DECODE if r < P + D else MODE_ROUTER
        ixs.append(make_index(mode=mode, self_rank=r, n_prefill=P, n_decode=D, max_nodes=1 << 16 if big else 4096,
                              max_tokens=1 << 20 if big else 1 << 15, max_batch=max_batch, max_key_len=64,
                              max_dups=1 << 15 if big else 512))
    pp, po = rings(port.p_addrs)
    dp, do = rings(port.d_addrs)
    ixs[5].set_rings(pp, po, dp, do)
    held = []   # (rank, device node id, gen, oracle node)
    for step in range(n_steps):
        kind = rnd.choice(["ins", "ins", "ins", "match", "route", "lock", "unlock"])
        if kind == "ins":
            recs = []
            for _ in range(rnd.randrange(1, max_recs + 1)):
                rank = rnd.randrange(P + D)
                key = new_key()
                variant = rnd.randrange(2)
                val = [100 * (rank + 1) * variant + 10 * i + t for i, t in enumerate(key)]
                recs.append((rank, key, val))
            exp = [port.insert(rank, key, np.asarray(val, dtype=np.int64)) for rank, key, val in recs]
            tk, offs = pack([k for _, k, _ in recs])
            vv, _ = pack([v for _, _, v in recs], np.int64)
            rk = np.asarray([r for r, _, _ in recs], dtype=np.int32)
            # every replica applies the whole batch in arrival order (oplog_received -> _insert)
            for r, ix in enumerate(ixs):
                got = ix.insert_host(tk, vv, offs, rk)
                ix.check_device_error()
                # replicas hold the same logical tree, so every one reports what the origin's insert returned
                assert got.tolist() == exp, f"insert return values on rank {r}"
        elif kind == "match":
            rank = rnd.randrange(P + D)
            keys = [[rnd.randrange(vocab) for _ in range(rnd.randrange(1, maxlen))] for _ in range(rnd.randrange(1, 5))]
            ix = ixs[rank]
            tk, offs = pack(keys)
            ol = np.zeros(len(keys), np.int32)
            on = np.zeros(len(keys), np.int32)
            ix.match_host(tk, offs, ol, on)
            exp = [port.match(rank, k) for k in keys]
            assert ol.tolist() == [len(m.indices) for m in exp]
            oo = np.zeros(len(keys) + 1, np.int64)
            oo[1:] = np.cumsum(ol)
            d_out = ix.be.empty(int(oo[-1]) + 1, np.int64)
            ix.gather(ix.be.to_device(on), ix.be.to_device(ol), ix.be.to_device(oo), len(keys), d_out)
            outv = ix.be.to_host(d_out)
            c = mesh_canonical(ix)
            gens = c["_dump"]["gen"]
            for i, m in enumerate(exp):
                assert outv[oo[i]:oo[i + 1]].tolist() == m.indices.tolist()
                assert c["_inv"][int(on[i])] == tuple(port.nodes[rank].full_key(m.last_node)), f"last node, query {i}"
                held.append((rank, int(on[i]), int(gens[int(on[i])]), m.last_node))
        elif kind == "route":
            keys = [[rnd.randrange(vocab) for _ in range(rnd.randrange(1, maxlen))] for _ in range(rnd.randrange(1, 5))]
            warm = rnd.random() < 0.85
            ix = ixs[5]
            tk, offs = pack(keys)
            n = len(keys)
            be = ix.be
            outs = [be.empty(n, np.int32) for _ in range(4)]
            err = be.empty(n, np.uint8)
            ix.route(be.to_device(tk), be.to_device(offs), n, warm, outs[0], outs[1], outs[2], outs[3], err)
            pf, dc, pa, da = [be.to_host(o).tolist() for o in outs]
            er = be.to_host(err).tolist()
            for i, k in enumerate(keys):
                if warm:
                    try:
                        exp_ranks = port.match(5, k)
                    except RouterPickError:
                        assert er[i] == 1, f"query {i}: the oracle asserts"
                        continue
                    assert (pf[i], dc[i]) == tuple(exp_ranks)
                else:
                    assert (pf[i], dc[i]) == (-1, -1)
                try:
                    ea = port.route(k, warm=warm) if not warm else None
                    if warm:
                        # port.route would run the match a second time (one more op on the oracle clock only)
                        from oracle.radix_port import port_route  # noqa: F401
                        p_addr = port.p_addrs[exp_ranks[0]] if exp_ranks[0] != -1 else PortConsistentHash(port.p_addrs).get_node(k)
                        d_addr = (port.d_addrs[port.nodes[5].local_rank(exp_ranks[1])] if exp_ranks[1] != -1
                                  else PortConsistentHash(port.d_addrs).get_node(k))
                        ea = (p_addr, d_addr)
                except IndexError:
                    assert er[i] == 2, f"query {i}: the oracle raises IndexError"
                    continue
                assert er[i] == 0
                assert (port.p_addrs[pa[i]], port.d_addrs[da[i]]) == tuple(ea)
        elif kind in ("lock", "unlock") and held:
            rank, node, gen, onode = held[rnd.randrange(len(held))]
            ix = ixs[rank]
            d_out = ix.be.empty(1, np.int64)
            sign = +1 if kind == "lock" else -1
            ix.lock(ix.be.to_device(np.asarray([node], np.int32)), ix.be.to_device(np.asarray([gen], np.int32)), 1, sign, d_out)
            exp = port.nodes[rank].inc_lock_ref(onode) if sign > 0 else port.nodes[rank].dec_lock_ref(onode)
            assert int(ix.be.to_host(d_out)[0]) == exp
        if rnd.random() < 0.15:
            # maintenance between operations: nothing observable may change (parked nodes and the duplicate table keep
            # pointing at the same values)
            for ix in ixs:
                ix.compact()
        for r, ix in enumerate(ixs):
            ix.check_device_error()
            dc_ = mesh_canonical(ix)
            oc = port.canonical(r)
            assert dc_["nodes"] == oc["nodes"], f"rank {r} tree differs at step {step} ({kind})\n dev {dc_['nodes']}\n orc {oc['nodes']}"
            assert dc_["evictable"] == oc["evictable"] and dc_["protected"] == oc["protected"], f"rank {r} sizes"
            assert dc_["dup"] == oc["dup"], f"rank {r} dup table differs at step {step} ({kind})\n dev {dc_['dup']}\n orc {oc['dup']}"
    for r in range(P + D):
        cand = sorted((mesh_canonical(ixs[r])["_inv"][n], rk) for n, rk, _, _ in ixs[r].dup_entries(only_candidates=True).tolist())
        assert cand == sorted(port.nodes[r].gc_candidates())
    stats = {"dups": sum(len(port.nodes[r].dup_nodes) for r in range(P + D)),
             "parked": sum(1 for r in range(P + D) for n in port.nodes[r].dup_nodes.values() if n is not None)}
    for ix in ixs:
        ix.close()
    return stats


def md5_awkward_keys():
    """Negative and 10-digit ints, lengths around the 64-byte block and padding edges, the empty list, and enough
    2K-token requests to overflow a 1 MB message scratch (the rest then takes the streamed form)."""
    import random
    rnd = random.Random(5)
    pool = [0, 7, 10, 99, 100, 31999, 32000, 123456, 2147483647, -1, -2147483648, 999999, 1000000]
    lens = list(range(0, 40)) + [55, 56, 57, 63, 64, 65, 119, 120, 121, 127, 128, 129, 500] + [2176] * 40
    keys = [[rnd.choice(pool) for _ in range(n)] for n in lens]
    # runs of ordinary token ids (< 1000000), which the render kernel takes 128 at a time, four per lane: every digit
    # count, the shortest and the longest text a lane can hold, lengths around the 128-token chunk, and runs broken
    # by an id that sends one chunk through the general path
    small = [0, 5, 9, 10, 42, 99, 100, 999, 1000, 9999, 10000, 31999, 99999]
    for n in (128, 129, 131, 159, 160, 255, 256, 257, 300, 511, 640, 1000, 2176, 4000):
        keys.append([rnd.choice(small) for _ in range(n)])
    for v in (0, 7, 10, 99, 100, 1234, 99999):
        keys.append([v] * 384)
    keys.append([rnd.randrange(0, 100000) for _ in range(3000)])
    broken = [rnd.randrange(0, 32000) for _ in range(700)]
    broken[10], broken[200], broken[699] = -5, 123456, 2147483647
    keys.append(broken)
    keys.append([1000000] + [rnd.randrange(0, 100000) for _ in range(511)])
    # six-digit ids (vocabularies of 128 K - 262 K entries) stay on the four-per-lane path; seven digits leave it
    keys.append([rnd.randrange(100000, 1000000) for _ in range(640)])
    keys.append([999999] * 256)
    keys.append([rnd.choice([0, 9, 99999, 100000, 151643, 262143, 999999]) for _ in range(1500)])
    keys.append([rnd.randrange(0, 300000) for _ in range(255)] + [1000000] + [rnd.randrange(0, 300000) for _ in range(400)])
    return keys
