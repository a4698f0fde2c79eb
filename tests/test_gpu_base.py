"""GPU parity tests proper: the CUDA library through its C ABI against the CPU oracle (RadixCache semantics)."""
import numpy as np
import pytest

from tests import cases
from tests.parity import assert_same_tree, device_canonical, pack, zipf_prefix

pytestmark = pytest.mark.gpu


def make_index(**kw):
    from radixmesh_b200 import capi
    return capi.DeviceIndex(**kw)


@pytest.mark.parametrize("seed", range(12))
def test_gpu_random_sequences(seed):
    cases.run_sequence(make_index, seed)


@pytest.mark.parametrize("seed", range(3))
def test_gpu_longer_keys(seed):
    cases.run_sequence(make_index, 1000 + seed, n_steps=25, vocab=4, maxlen=40, batch=12)


@pytest.mark.parametrize("seed", range(12))
def test_gpu_batched_match_splits_then_evict_order(seed):
    cases.run_split_evict(make_index, seed)


@pytest.mark.parametrize("seed", range(3))
def test_gpu_evict_second_pass_with_exact_sizes(seed, monkeypatch):
    """rm_evict sizes its sort from a bound on the candidate count and keeps the count on the device; with more
    candidates than the bound it must evict nothing and run again with exact sizes (forced: a bound of 2)."""
    monkeypatch.setenv("RM_EVICT_M_OPT", "2")
    cases.run_split_evict(make_index, 40 + seed)
    cases.run_stream_evict(make_index, 40 + seed)


WALKS = [("run", 4), ("run", 8), ("run", 16), ("run", 32), ("table", 2), ("table", 8), ("table", 32)]


@pytest.mark.parametrize("walk,group", WALKS, ids=[f"{w}{g}" for w, g in WALKS])
def test_gpu_zipf_match_all_walk_kernels(walk, group):
    """Zipf shared-prefix workload at reduced size, device-pointer entry points, every form of the
    walk kernel (run-following with 4..32 lanes per query - 32 is the product path - and hop-by-hop through the
    child table with 2..32 lanes per query): lengths, last nodes, values and final tree."""
    import os
    import subprocess
    import sys
    # the kernel choice is read once per process: run the check in a child with the environment set
    code = ("import sys; sys.path.insert(0, '.'); from tests.test_gpu_base import zipf_check; "
            "zipf_check(seed=5, n_docs=64, L=96, n_ins=600, n_q=3000)")
    env = dict(os.environ, RM_WALK=walk, RM_RUN_G=str(group), RM_MATCH_G=str(group))
    subprocess.run([sys.executable, "-c", code], check=True, env=env,
                   cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@pytest.mark.parametrize("env", [{"RM_RUN_DEPTH": "2"}, {"RM_RUN_DEPTH": "4"}, {"RM_SPLIT_ROLES": "0"}, {"RM_SPLIT_THREAD_MAX": "1"},
                                 {"RM_SPLIT_SMALL_MAX": "1"}, {"RM_SPLIT_SMALL_MAX": "3"}],
                         ids=lambda e: "-".join(f"{k}={v}" for k, v in e.items()))
def test_gpu_zipf_match_kernel_options(env):
    """The same check with the product walk's options (2 or 4 steps in flight) and the split stage's: singles and
    groups in the same blocks, handfuls of requests ordered by the warp instead of one thread, and the in-register
    paths cut down so that edges with more than 1 or 3 requests go through the sorted path (run by the last block of
    the same launch)."""
    import os
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, '.'); from tests.test_gpu_base import zipf_check; "
            "zipf_check(seed=7, n_docs=64, L=96, n_ins=600, n_q=3000)")
    subprocess.run([sys.executable, "-c", code], check=True, env=dict(os.environ, **env),
                   cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@pytest.mark.parametrize("small_max", [None, "1", "5"])
def test_gpu_split_stage_paths(small_max):
    """Split/evict scenarios (tens of requests per edge) with every request-count class of the split stage in play."""
    import os
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, '.'); from tests import cases; from tests.test_gpu_base import make_index; "
            "[cases.run_split_evict(make_index, 100 + s) for s in range(6)]")
    env = dict(os.environ) if small_max is None else dict(os.environ, RM_SPLIT_SMALL_MAX=small_max)
    subprocess.run([sys.executable, "-c", code], check=True, env=env,
                   cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def test_gpu_many_requests_on_one_edge():
    """A batch that meets a long unsplit edge for the first time: thousands of requests on ONE edge (and a second
    edge with a few hundred), far beyond what a warp orders in registers - the sorted path of the split stage, more
    than one tile of its single-block sort."""
    import torch
    from oracle.radix_port import PortRadixCache
    from tests.parity import assert_same_digest, device_digest, oracle_digest
    rng = np.random.default_rng(5)
    L = 600
    docs = rng.integers(0, 32000, (3, L), dtype=np.int32)
    ix = make_index(mode=0, max_nodes=1 << 14, max_tokens=1 << 16, max_batch=8192, max_key_len=1024)
    orc = PortRadixCache(1)
    vals = np.arange(3 * L, dtype=np.int64)
    got = ix.insert_host(docs.reshape(-1).copy(), vals, np.arange(0, 4 * L, L, dtype=np.int64))
    assert got.tolist() == [orc.insert(d.tolist(), vals[i * L:(i + 1) * L]) for i, d in enumerate(docs)]
    for rnd in range(2):
        qs = []
        for _ in range(6000):
            d = 0 if rng.random() < 0.9 else (1 if rng.random() < 0.9 else 2)
            cut = int(rng.integers(1, L + 1))
            qs.append(docs[d, :cut].tolist() + [40000])
        tk, offs = pack(qs)
        ol = np.zeros(len(qs), np.int32)
        on = np.zeros(len(qs), np.int32)
        ix.match_host(tk, offs, ol, on)
        exp = [orc.match_prefix(k) for k in qs]
        assert ol.tolist() == [len(m.indices) for m in exp]
        dd, od = device_digest(ix), oracle_digest(orc)
        assert_same_digest(dd, od)
        assert [dd["by_id"][x] for x in on.tolist()] == [od["by_node"][m.last_node] for m in exp]
    ix.check_device_error()
    _, eo, el = ix.evict(1 << 30)
    orc.evict(1 << 30)
    assert el.tolist() == [len(r) for r in orc.freed]
    ix.close()


def test_gpu_zipf_match_without_dependent_launch():
    """The same check with programmatic dependent launch turned off (RM_PDL=0): the split-stage kernels then start
    only after their predecessor has finished, as plain stream order gives."""
    import os
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, '.'); from tests.test_gpu_base import zipf_check; "
            "zipf_check(seed=6, n_docs=64, L=96, n_ins=600, n_q=3000)")
    subprocess.run([sys.executable, "-c", code], check=True, env=dict(os.environ, RM_PDL="0"),
                   cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def zipf_check(seed, n_docs, L, n_ins, n_q, batch=4096):
    import torch
    from oracle.radix_port import PortRadixCache
    ix = make_index(mode=0, max_nodes=1 << 21, max_tokens=1 << 26, max_batch=1 << 16, max_key_len=max(L, 1024))
    orc = PortRadixCache(1)
    ins, docs = zipf_prefix(seed, n_docs, n_ins, L)
    qs, _ = zipf_prefix(seed + 1, n_docs, n_q, L, docs=docs)
    dev = torch.device("cuda")
    vc = 0
    for b0 in range(0, n_ins, batch):
        chunk = ins[b0:b0 + batch]
        n = len(chunk)
        tok = torch.from_numpy(chunk.reshape(-1)).to(dev)
        val = torch.arange(vc, vc + n * L, dtype=torch.int64, device=dev)
        offs = torch.arange(0, (n + 1) * L, L, dtype=torch.int64, device=dev)
        out = torch.empty(n, dtype=torch.int32, device=dev)
        ix.insert(tok, val, offs, None, n, out)
        exp = [orc.insert(row.tolist(), np.arange(vc + i * L, vc + (i + 1) * L, dtype=np.int64)) for i, row in enumerate(chunk)]
        assert out.cpu().tolist() == exp
        vc += n * L
    ix.check_device_error()
    assert_same_tree(device_canonical(ix), orc.canonical())
    for b0 in range(0, n_q, batch):
        chunk = qs[b0:b0 + batch]
        n = len(chunk)
        tok = torch.from_numpy(chunk.reshape(-1)).to(dev)
        offs = torch.arange(0, (n + 1) * L, L, dtype=torch.int64, device=dev)
        ol = torch.empty(n, dtype=torch.int32, device=dev)
        on = torch.empty(n, dtype=torch.int32, device=dev)
        ix.match(tok, offs, n, ol, on)
        exp = [orc.match_prefix(row.tolist()) for row in chunk]
        assert ol.cpu().tolist() == [len(m.indices) for m in exp]
        oo = torch.zeros(n + 1, dtype=torch.int64, device=dev)
        oo[1:] = torch.cumsum(ol.to(torch.int64), 0)
        vals = torch.empty(int(oo[-1].item()) + 1, dtype=torch.int64, device=dev)
        ix.gather(on, ol, oo, n, vals)
        hv, ho = vals.cpu().numpy(), oo.cpu().numpy()
        d = device_canonical(ix)
        inv = {v: k for k, v in d["_ids"].items()}
        inv[0] = ()
        onh = on.cpu().tolist()
        for i, m in enumerate(exp):
            assert hv[ho[i]:ho[i + 1]].tolist() == m.indices.tolist()
            assert inv[onh[i]] == tuple(orc.full_key(m.last_node))
    ix.check_device_error()
    assert_same_tree(device_canonical(ix), orc.canonical())
    ix.close()


def test_gpu_zipf_c2_reduced():
    """configs[1] shape (Zipf, 4096 docs, L=256) at a size the oracle finishes in seconds."""
    zipf_check(seed=0, n_docs=4096, L=256, n_ins=6000, n_q=8192)


@pytest.mark.parametrize("seed", range(3))
def test_gpu_streaming_insert_evict_compact(seed):
    cases.run_stream_evict(make_index, seed, n_batches=16, batch=96, L=32, n_docs=24, capacity=6000)


@pytest.mark.parametrize("seed", range(2))
def test_gpu_cold_batches_are_applied_in_chunks(seed):
    cases.run_cold_batch(make_index, seed, n=2500, L=96, n_docs=12)


@pytest.mark.parametrize("seed", range(2))
def test_gpu_tombstone_pressure_rebuilds_the_table(seed):
    cases.run_tombstone_pressure(make_index, seed)


def test_gpu_radix_sort_all_forms():
    """tools/sort_probe.cu (built by __graft_entry__.build()): the device radix sort against std::stable_sort for
    item counts from 0 to 1 M with the count read on the device - one block alone, the one-launch grid form, and
    the three-kernels-per-pass form (bound > 148 tiles)."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "build", "sort_probe")
    if not os.path.exists(exe):
        pytest.skip("build/sort_probe not built (run __graft_entry__.build())")
    for args in (["65536", "30"], ["1100000", "40"]):
        out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr
        lines = [ln for ln in out.stdout.splitlines() if ln.startswith("n ")]
        assert len(lines) >= 10 and all(ln.rstrip().endswith("ok") for ln in lines), out.stdout
