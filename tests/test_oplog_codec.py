"""Oplog ingest: wire formats and the staging buffer, on CPU.

* JSON compatibility with the reference's JsonSerializer: golden messages recorded from the reference
  (tools/make_golden.py -> tests/golden/oplog_wire.json), byte for byte on the way out, field for field on the way in.
* Packed format: round trip, header patching, length checks.
* IngestBuffer / RadixMesh.oplog_received_wire: a mesh node fed packed, JSON and object messages in one stream ends
  with the same tree as the oracle's mesh node fed the same records (host emulation build of the kernels).
"""
import random

import numpy as np
import pytest

from radixmesh_b200.host import CacheOplog, CacheOplogType
from radixmesh_b200.oplog_codec import (HEADER_BYTES, IngestBuffer, JsonSerializer, PackedSerializer, deserialize_any,
                                        is_packed, with_header)
from tests.golden_util import load


def test_json_wire_matches_reference_golden():
    ser = JsonSerializer()
    for case in load("oplog_wire"):
        f = case["fields"]
        op = ser.deserialize(case["wire"].encode("utf-8"))
        assert (int(op.oplog_type), op.node_rank, op.local_logic_id, op.key, op.value, op.ttl) == \
               (f["oplog_type"], f["node_rank"], f["local_logic_id"], f["key"], f["value"], f["ttl"])
        # and the way out: the reference's bytes exactly (json.dumps of to_dict, serializer.py:24-27)
        assert ser.serialize(op) == case["wire"].encode("utf-8")
        # sized-buffer form of the interface (the communicator hands a larger bytearray and a size)
        buf = bytearray(case["wire"].encode("utf-8")) + b"garbage"
        assert ser.deserialize(buf, len(buf) - 7).key == f["key"]


def test_packed_round_trip_of_the_golden_messages():
    ser = PackedSerializer()
    for case in load("oplog_wire"):
        f = case["fields"]
        op = CacheOplog(oplog_type=CacheOplogType(f["oplog_type"]), node_rank=f["node_rank"], local_logic_id=f["local_logic_id"],
                        key=f["key"], value=f["value"], ttl=f["ttl"])
        wire = ser.serialize(op)
        assert is_packed(wire) and len(wire) == HEADER_BYTES + 4 * len(f["key"]) + 8 * len(f["value"])
        back = deserialize_any(wire)
        assert back == op
        fwd = deserialize_any(with_header(wire, 9, 1234))
        assert (fwd.ttl, fwd.local_logic_id, fwd.key, fwd.value) == (9, 1234, op.key, op.value)
    with pytest.raises(ValueError):
        ser.deserialize(wire[:-3])
    with pytest.raises(ValueError):
        ser.deserialize(b"XXXX" + wire[4:])


def test_ingest_buffer_gate_and_layout():
    ser = PackedSerializer()
    ib = IngestBuffer(with_values=True, tokens=4, records=1)      # tiny: growth is exercised
    msgs = []
    rnd = random.Random(3)
    for i in range(50):
        n = rnd.randrange(0, 40)
        msgs.append(CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=rnd.randrange(0, 4), local_logic_id=i,
                               key=[rnd.randrange(100) for _ in range(n)], value=[rnd.randrange(1 << 40) for _ in range(n)],
                               ttl=rnd.randrange(1, 4)))
    kept = []
    for m in msgs:
        what, op, ttl = ib.feed_wire(ser.serialize(m), None, self_rank=2)
        if m.node_rank == 2 or m.ttl - 1 <= 0:            # radix_mesh.py:408-409
            assert what == "drop"
        else:
            assert what == "insert" and ttl == m.ttl - 1
            kept.append(m)
    what, op, _ = ib.feed_wire(ser.serialize(CacheOplog(oplog_type=CacheOplogType.TICK, node_rank=1, ttl=3)), None, 2)
    assert what == "other" and int(op.oplog_type) == CacheOplogType.TICK and op.ttl == 3
    extra = CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=1, local_logic_id=77, key=[5, -6, 7], value=[8, 9, -10], ttl=3)
    what, op, ttl = ib.feed_wire(JsonSerializer().serialize(extra), None, 2)       # the reference's JSON: fast path
    assert what == "insert" and ttl == 2
    kept.append(extra)
    empty = CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=0, local_logic_id=78, key=[], value=[], ttl=3)
    assert ib.feed_wire(JsonSerializer().serialize(empty), None, 2)[0] == "insert"
    kept.append(empty)
    odd = b'{"node_rank": 1, "oplog_type": 1, "ttl": 3, "key": [1], "value": [2]}'   # other key order: json.loads
    what, op, _ = ib.feed_wire(odd, None, 2)
    assert what == "other" and op.key == [1] and op.ttl == 3
    fwd = deserialize_any(with_header(JsonSerializer().serialize(extra), 6, 99))
    assert (fwd.ttl, fwd.local_logic_id, fwd.key, fwd.value, fwd.node_rank) == (6, 99, extra.key, extra.value, 1)
    tok, val, offs, ranks = ib.take()
    assert len(ib) == 0 and len(ranks) == len(kept)
    assert ranks.tolist() == [m.node_rank for m in kept]
    for i, m in enumerate(kept):
        assert tok[offs[i]:offs[i + 1]].tolist() == m.key and val[offs[i]:offs[i + 1]].tolist() == m.value
    with pytest.raises(ValueError):
        ib.add_arrays(np.arange(3, dtype=np.int32), np.arange(2, dtype=np.int64), 0)


def _decode_node(kw):
    from radixmesh_b200 import RadixMesh, ServerArgs
    from tests import dropin_cases as dc
    args = ServerArgs(prefill_cache_nodes=dc.P_ADDRS, decode_cache_nodes=dc.D_ADDRS, router_cache_nodes=dc.R_ADDRS)
    args.decode_node_rank = 3
    return RadixMesh(args, communicator=None, max_nodes=1 << 12, max_tokens=1 << 15, max_batch=16, max_key_len=64,
                     max_dups=512, **kw)


@pytest.mark.parametrize("seed", range(6))
def test_mesh_node_fed_from_the_wire_matches_object_path_and_oracle(seed):
    from tests.emu.emu_backend import NumpyBackend, emu_lib
    run_wire_case(seed, lambda: dict(_backend=NumpyBackend(), _library=emu_lib()))


def run_wire_case(seed, make_kw):
    """A decode node receives the INSERT stream of its peers as packed bytes, JSON bytes and objects, interleaved;
    its tree must equal that of a node fed the same records as objects (the path the reference goldens pin) and
    its matches those of the oracle's node.  (Host emulation here; tests/test_gpu_mesh.py runs it on CUDA indexes.)"""
    from oracle.radix_port import MODE_DECODE, PortRadixMesh
    from tests import dropin_cases as dc
    rnd = random.Random(seed)
    wire_node = _decode_node(make_kw())
    obj_node = _decode_node(make_kw())
    port = PortRadixMesh(MODE_DECODE, 3, 3, 2)
    packed, js = PackedSerializer(), JsonSerializer()
    for step in range(60):                                  # > max_batch records between flushes: chunked apply
        rank = rnd.choice([0, 1, 2, 4, 3])                  # rank 3 = this node: dropped (radix_mesh.py:408)
        key = [rnd.randrange(3) for _ in range(rnd.randrange(1, 7))]
        variant = rnd.randrange(2)
        val = [100 * (rank + 1) * variant + 10 * i + t for i, t in enumerate(key)]
        ttl = rnd.choice([4, 4, 4, 1])                      # ttl 1 expires on arrival (radix_mesh.py:409)
        op = CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=rank, local_logic_id=step, key=key, value=val, ttl=ttl)
        how = rnd.randrange(3)
        if how == 0:
            wire_node.oplog_received_wire(packed.serialize(op))
        elif how == 1:
            w = js.serialize(op)
            wire_node.oplog_received_wire(bytearray(w) + b"xx", len(w))
        else:
            wire_node.oplog_received(CacheOplog(**op.to_dict()))
        obj_node.oplog_received(CacheOplog(**op.to_dict()))
        if rank != 3 and ttl - 1 > 0:
            port.apply_insert(list(key), np.asarray(val, dtype=np.int64), rank)
        if rnd.random() < 0.15:
            q = [rnd.randrange(3) for _ in range(rnd.randrange(1, 7))]
            got = wire_node.match_prefix(q)
            assert got.device_indices.tolist() == obj_node.match_prefix(q).device_indices.tolist()
            assert got.device_indices.tolist() == port.match_prefix(list(q)).indices.tolist()
    assert dc.canon(wire_node) == dc.canon(obj_node)
