"""The numpy matched-length oracle (oracle/lcp_oracle.py) against the pure-Python restatement of RadixCache."""
import numpy as np
import pytest

from oracle.lcp_oracle import longest_prefix_lengths
from oracle.radix_port import PortRadixCache
from tests.parity import zipf_prefix


def port_lengths(ins, qs):
    t = PortRadixCache(1)
    for r in ins:
        t.insert(r.tolist(), np.zeros(len(r), dtype=np.int64))
    return np.asarray([len(t.match_prefix(r.tolist()).indices) for r in qs], dtype=np.int32)


@pytest.mark.parametrize("seed", range(4))
def test_zipf_workload(seed):
    ins, docs = zipf_prefix(100 + seed, 24, 600, 48)
    qs, _ = zipf_prefix(200 + seed, 24, 800, 48, docs=docs)
    assert longest_prefix_lengths(ins, qs).tolist() == port_lengths(ins, qs).tolist()


@pytest.mark.parametrize("vocab", [2, 3, 40000])
def test_small_vocabulary_and_zero_tokens(vocab):
    # zeros everywhere (trailing NUL bytes in the sort keys), duplicates, chance collisions between documents
    rng = np.random.default_rng(vocab)
    ins = rng.integers(0, vocab, (500, 12), dtype=np.int32)
    ins[::7] = 0
    qs = rng.integers(0, vocab, (700, 12), dtype=np.int32)
    qs[::5, 6:] = 0
    assert longest_prefix_lengths(ins, qs).tolist() == port_lengths(ins, qs).tolist()


def test_empty_sides():
    z = np.zeros((0, 8), dtype=np.int32)
    q = np.ones((3, 8), dtype=np.int32)
    assert longest_prefix_lengths(z, q).tolist() == [0, 0, 0]
    assert longest_prefix_lengths(q, z).tolist() == []
