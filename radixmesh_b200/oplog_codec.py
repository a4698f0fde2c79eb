"""Oplog ingest: wire messages decoded straight into the staging arrays of one ordered `rm_insert_batch`.

SURVEY.md 8f-1, the step immediately upstream of K4.  The reference decodes every inbound message with
`JsonSerializer.deserialize` = UTF-8 decode + `CacheOplog.model_validate_json` (pydantic) into Python lists
(`communication/serializer.py:20-35`, `radix/cache_oplog.py:48-66`) and hands the object to
`RadixMesh.oplog_received` (`radix/radix_mesh.py:391-423`), one tree insert per message.  Here:

* `PackedSerializer` - a binary wire format for the same fields (the "pb serializer" of the reference's roadmap,
  README.md:55): 28-byte header + int32 key + int64 value.  Same `serialize` / `deserialize(buf, size)` interface
  as the reference's `Serializer` (`serializer.py:9-15`), so a `Communicator` takes it unchanged.
* `JsonSerializer` - the reference's own wire format, byte for byte on the way out, parsed with `json.loads` on the
  way in (no pydantic), for meshes whose peers still speak JSON.
* `IngestBuffer` - growable flat arrays (tokens i32, values i64, offsets i64, ranks i32).  `feed_wire` applies the
  gate of `oplog_received` (ttl decrement, own-rank and expired drop, `radix_mesh.py:392-409`) on the header alone
  and copies an INSERT's payload with two `np.frombuffer` slices; nothing is turned into Python ints.  `take()`
  hands the arrays to `DeviceIndex.insert_host` in arrival order.

Host code only; the device side is `rm_insert_batch_h` (include/radixmesh_b200.h).
"""
from __future__ import annotations

import json
import re
import struct
import warnings
from typing import Optional, Tuple

import numpy as np

from .host import CacheOplog, CacheOplogType

MAGIC = b"RMO1"
# magic, oplog_type u8, pad, node_rank i16, local_logic_id i64, ttl i32, n_key u32, n_value u32
_HDR = struct.Struct("<4sBxhqiII")
HEADER_BYTES = _HDR.size
_TTL_OFF = 16           # byte offset of ttl inside the header
_TTL = struct.Struct("<i")
_LID_OFF = 8            # byte offset of local_logic_id
_LID = struct.Struct("<q")


class PackedSerializer:
    """Binary wire format; interface of the reference's Serializer (communication/serializer.py:9-15)."""

    def serialize(self, value) -> bytes:
        key = np.asarray(value.key, dtype=np.int32)
        val = np.asarray(value.value.tolist() if hasattr(value.value, "tolist") else value.value, dtype=np.int64)
        return _HDR.pack(MAGIC, int(value.oplog_type), int(value.node_rank), int(value.local_logic_id), int(value.ttl),
                         key.size, val.size) + key.tobytes() + val.tobytes()

    def deserialize(self, buf, size: Optional[int] = None) -> CacheOplog:
        mv = memoryview(buf)[:size] if size is not None else memoryview(buf)
        if len(mv) < HEADER_BYTES or bytes(mv[:4]) != MAGIC:
            raise ValueError("not a packed oplog message")
        magic, typ, rank, lid, ttl, nk, nv = _HDR.unpack_from(mv, 0)
        if len(mv) != HEADER_BYTES + 4 * nk + 8 * nv:
            raise ValueError(f"packed oplog: {len(mv)} bytes, header announces {HEADER_BYTES + 4 * nk + 8 * nv}")
        key = np.frombuffer(mv, dtype=np.int32, count=nk, offset=HEADER_BYTES).tolist()
        val = np.frombuffer(mv, dtype=np.int64, count=nv, offset=HEADER_BYTES + 4 * nk).tolist()
        return CacheOplog(oplog_type=CacheOplogType(typ), node_rank=rank, local_logic_id=lid, key=key, value=val, ttl=ttl)


class JsonSerializer:
    """The reference's wire format (communication/serializer.py:20-35): json.dumps(oplog.to_dict()) in UTF-8."""

    def serialize(self, value) -> bytes:
        return json.dumps(value.to_dict()).encode("utf-8")

    def deserialize(self, buf, size: Optional[int] = None) -> CacheOplog:
        raw = bytes(memoryview(buf)[:size] if size is not None else buf)
        d = json.loads(raw.decode("utf-8"))
        # model_validate_json: oplog_type, node_rank and ttl are required (cache_oplog.py:48-56)
        return CacheOplog(oplog_type=CacheOplogType(d["oplog_type"]), node_rank=int(d["node_rank"]),
                          local_logic_id=int(d.get("local_logic_id", 0)), key=list(d.get("key", [])),
                          value=list(d.get("value", [])), ttl=int(d["ttl"]))


# The reference's JSON message is json.dumps of a dict with a fixed key order (cache_oplog.py:58-66), so the two
# integer lists can be cut out of the text and parsed by numpy's C reader instead of json.loads + np.asarray
# (6.4 K -> ~40 K messages/s at 512 tokens).  Anything that does not match goes through json.loads.
_JSON_INSERT = re.compile(rb'^\{"oplog_type": 1, "node_rank": (-?\d+), "local_logic_id": (-?\d+), "key": \[([0-9, -]*)\], '
                          rb'"value": \[([0-9, -]*)\], "ttl": (-?\d+)\}$')


def _int_list(text: bytes, dtype) -> np.ndarray:
    if not text:
        return np.empty(0, dtype=dtype)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", DeprecationWarning)
        return np.fromstring(text.decode("ascii"), dtype=np.int64, sep=",").astype(dtype, copy=False)


def is_packed(buf) -> bool:
    return bytes(memoryview(buf)[:4]) == MAGIC


def deserialize_any(buf, size: Optional[int] = None) -> CacheOplog:
    return (PackedSerializer() if is_packed(buf) else JsonSerializer()).deserialize(buf, size)


def with_header(buf, ttl: int, local_logic_id: Optional[int] = None) -> bytes:
    """The same INSERT message with another ttl (and logic id): re-forwarding (radix_mesh.py:411-416, 325-347)
    without decoding the payload.  Packed messages are patched in place, JSON ones re-assembled around their lists."""
    if not is_packed(buf):
        raw = bytes(buf)
        m = _JSON_INSERT.match(raw)
        if m is None:
            op = JsonSerializer().deserialize(raw)
            op.ttl = int(ttl)
            if local_logic_id is not None:
                op.local_logic_id = int(local_logic_id)
            return JsonSerializer().serialize(op)
        lid = int(m.group(2)) if local_logic_id is None else int(local_logic_id)
        return (b'{"oplog_type": 1, "node_rank": %d, "local_logic_id": %d, "key": [%s], "value": [%s], "ttl": %d}'
                % (int(m.group(1)), lid, m.group(3), m.group(4), int(ttl)))
    out = bytearray(buf)
    _TTL.pack_into(out, _TTL_OFF, int(ttl))
    if local_logic_id is not None:
        _LID.pack_into(out, _LID_OFF, int(local_logic_id))
    return bytes(out)


class IngestBuffer:
    """INSERT records in arrival order, as the flat arrays rm_insert_batch_h takes."""

    def __init__(self, with_values: bool = True, tokens: int = 1 << 16, records: int = 1 << 10, alloc=None):
        """alloc(n, dtype) -> 1-D array for the token and value arrays (default numpy.empty); a caller that copies them
        to a GPU passes an allocator of pinned memory (mesh_oplog.py) so that the copy is asynchronous."""
        self.with_values = with_values
        self._alloc = alloc if alloc is not None else (lambda n, dt: np.empty(n, dtype=dt))
        self._tok = self._alloc(tokens, np.int32)
        self._val = self._alloc(tokens if with_values else 1, np.int64)
        self._offs = np.zeros(records + 1, dtype=np.int64)
        self._rank = np.empty(records, dtype=np.int32)
        self.n = 0          # records
        self.ntok = 0       # tokens

    def __len__(self) -> int:
        return self.n

    def _room(self, nk: int) -> None:
        if self.ntok + nk > self._tok.size:
            size = max(2 * self._tok.size, self.ntok + nk)
            tok = self._alloc(size, np.int32)
            tok[:self.ntok] = self._tok[:self.ntok]
            self._tok = tok
            if self.with_values:
                val = self._alloc(size, np.int64)
                val[:self.ntok] = self._val[:self.ntok]
                self._val = val
        if self.n + 1 > self._rank.size:
            cap = 2 * self._rank.size
            self._rank = np.resize(self._rank, cap)
            self._offs = np.resize(self._offs, cap + 1)

    def add_arrays(self, key: np.ndarray, value: Optional[np.ndarray], rank: int) -> None:
        nk = int(key.size)
        if self.with_values and (value is None or int(value.size) != nk):
            raise ValueError(f"insert record: {nk} key tokens but {0 if value is None else int(value.size)} values")
        self._room(nk)
        a = self.ntok
        self._tok[a:a + nk] = key
        if self.with_values:
            self._val[a:a + nk] = value
        self._rank[self.n] = rank
        self.n += 1
        self.ntok = a + nk
        self._offs[self.n] = self.ntok

    def add(self, key, value, rank: int) -> None:
        """Object path (a decoded CacheOplog, or a local insert): lists / tensors / arrays."""
        k = np.asarray(key, dtype=np.int32)
        v = None
        if self.with_values:
            v = value.cpu().numpy() if hasattr(value, "cpu") else np.asarray(value)
            v = v.astype(np.int64, copy=False)
        self.add_arrays(k, v, rank)

    def feed_wire(self, buf, size: Optional[int], self_rank: int) -> Tuple[str, Optional[CacheOplog], int]:
        """One inbound message.  Returns (what, oplog, ttl):
        ("insert", None, ttl)  an INSERT (packed, or the reference's JSON), staged; ttl is the decremented ttl (the
                               caller re-forwards)
        ("drop", None, ttl)    an INSERT from this node or expired (radix_mesh.py:408-409)
        ("other", oplog, ttl)  anything else, decoded, ttl NOT yet decremented: goes through oplog_received
        """
        mv = memoryview(buf)[:size] if size is not None else memoryview(buf)
        if bytes(mv[:4]) != MAGIC:
            m = _JSON_INSERT.match(mv)
            if m is None:
                return "other", JsonSerializer().deserialize(mv), 0
            rank, ttl = int(m.group(1)), int(m.group(5)) - 1
            if rank == self_rank or ttl <= 0:
                return "drop", None, ttl
            self.add_arrays(_int_list(m.group(3), np.int32), _int_list(m.group(4), np.int64) if self.with_values else None, rank)
            return "insert", None, ttl
        if len(mv) < HEADER_BYTES:
            raise ValueError("packed oplog: shorter than its header")
        magic, typ, rank, lid, ttl, nk, nv = _HDR.unpack_from(mv, 0)
        if typ != CacheOplogType.INSERT:
            return "other", PackedSerializer().deserialize(mv), 0
        if len(mv) != HEADER_BYTES + 4 * nk + 8 * nv:
            raise ValueError(f"packed oplog: {len(mv)} bytes, header announces {HEADER_BYTES + 4 * nk + 8 * nv}")
        ttl -= 1
        if rank == self_rank or ttl <= 0:
            return "drop", None, ttl
        key = np.frombuffer(mv, dtype=np.int32, count=nk, offset=HEADER_BYTES)
        val = np.frombuffer(mv, dtype=np.int64, count=nv, offset=HEADER_BYTES + 4 * nk) if self.with_values else None
        self.add_arrays(key, val, rank)
        return "insert", None, ttl

    def take(self):
        """(tokens, values or None, offsets, ranks) of everything staged, and an empty buffer again."""
        n, nt = self.n, self.ntok
        # the caller keeps the filled arrays (views of them) and the buffer continues on fresh ones of the same size:
        # no copy of the staged tokens and values (6 MB for a thousand 512-token records)
        out = (self._tok[:nt], self._val[:nt] if self.with_values else None, self._offs[:n + 1], self._rank[:n])
        self._tok = self._alloc(self._tok.size, np.int32)
        self._val = self._alloc(self._val.size, np.int64)
        self._offs = np.zeros_like(self._offs)
        self._rank = np.empty_like(self._rank)
        self.n = 0
        self.ntok = 0
        return out
