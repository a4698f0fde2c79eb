"""rm_match_batch_static: the prefix match for callers that must not allocate device memory.

Engines that pre-commit HBM to their KV pool and request-path threads that cannot absorb cudaMalloc's
device-wide synchronisation send their match batches through this entry point; its scratch is a fixed
array inside the library.  The test builds the same tree in two indexes and in the CPU oracle, sends
the same two query batches through the regular entry point on one index and through the static-scratch
entry point on the other (whose own scratch is far too small for these batches), and requires
identical results: matched lengths bit-exact against each other and against the oracle, the same last
nodes (compared by full key, since node ids are handed out in parallel) and the same final tree as the
oracle's.  Collected last on purpose.
"""
import numpy as np
import pytest

from tests.parity import assert_same_tree, device_canonical, zipf_prefix

pytestmark = pytest.mark.gpu


def _build(ix, ins, L):
    import torch
    dev = torch.device("cuda")
    step = 128
    for b0 in range(0, len(ins), step):
        chunk = ins[b0:b0 + step]
        n = len(chunk)
        tok = torch.from_numpy(chunk.reshape(-1).copy()).to(dev)
        val = torch.arange(b0 * L, (b0 + n) * L, dtype=torch.int64, device=dev)
        offs = torch.arange(0, (n + 1) * L, L, dtype=torch.int64, device=dev)
        out = torch.empty(n, dtype=torch.int32, device=dev)
        ix.insert(tok, val, offs, None, n, out)
    ix.check_device_error()


def test_static_scratch_match_equals_regular_match_twice():
    import torch
    from radixmesh_b200 import capi
    L, n_ins, n_q = 64, 1024, 3000
    ins, docs = zipf_prefix(21, 48, n_ins, L)
    regular = capi.DeviceIndex(mode=capi.MODE_BASE, max_nodes=1 << 16, max_tokens=1 << 20, max_batch=4096, max_key_len=256)
    frugal = capi.DeviceIndex(mode=capi.MODE_BASE, max_nodes=1 << 16, max_tokens=1 << 20, max_batch=128, max_key_len=256)
    _build(regular, ins, L)
    _build(frugal, ins, L)
    from oracle.radix_port import PortRadixCache
    orc = PortRadixCache(1)
    for i, row in enumerate(ins):
        orc.insert(row.tolist(), np.arange(i * L, (i + 1) * L, dtype=np.int64))
    dev = torch.device("cuda")
    offs = torch.arange(0, (n_q + 1) * L, L, dtype=torch.int64, device=dev)
    for call in range(2):   # two calls in a row: the second reuses whatever the first left in the array
        qs, _ = zipf_prefix(30 + call, 48, n_q, L, docs=docs)
        tok = torch.from_numpy(qs.reshape(-1).copy()).to(dev)
        la, na = (torch.empty(n_q, dtype=torch.int32, device=dev) for _ in range(2))
        lb, nb = (torch.empty(n_q, dtype=torch.int32, device=dev) for _ in range(2))
        regular.match(tok, offs, n_q, la, na)
        frugal.match_static(tok, offs, n_q, lb, nb)
        torch.cuda.synchronize()
        assert torch.equal(la, lb), f"matched lengths differ on call {call}"
        exp = [orc.match_prefix(r.tolist()) for r in qs]
        assert lb.cpu().tolist() == [len(m.indices) for m in exp], f"static entry point differs from the oracle on call {call}"
        ca, cb = device_canonical(regular), device_canonical(frugal)
        ka = {v: k for k, v in ca["_ids"].items()}
        kb = {v: k for k, v in cb["_ids"].items()}
        ka[0] = kb[0] = ()
        assert [ka[x] for x in na.cpu().tolist()] == [kb[x] for x in nb.cpu().tolist()], f"last nodes differ on call {call}"
        assert_same_tree(cb, {"nodes": ca["nodes"], "evictable": ca["evictable"], "protected": ca["protected"]})
        assert [kb[x] for x in nb.cpu().tolist()] == [tuple(orc.full_key(m.last_node)) for m in exp], f"last nodes differ from the oracle on call {call}"
        assert_same_tree(cb, orc.canonical())
    regular.check_device_error()
    frugal.check_device_error()
    regular.close()
    frugal.close()


def test_static_scratch_rejects_oversized_batch():
    import torch
    from radixmesh_b200 import capi
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, max_nodes=1 << 10, max_tokens=1 << 12, max_batch=16, max_key_len=64)
    dev = torch.device("cuda")
    n = 4097
    tok = torch.zeros(n, dtype=torch.int32, device=dev)
    offs = torch.arange(0, n + 1, dtype=torch.int64, device=dev)
    out = torch.empty(n, dtype=torch.int32, device=dev)
    with pytest.raises(capi.RadixMeshError) as e:
        ix.match_static(tok, offs, n, out, out)
    assert e.value.status == capi.RM_E_TOO_LARGE and "4096" in str(e.value)
    ix.close()
