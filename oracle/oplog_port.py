"""TEST / MEASUREMENT INFRASTRUCTURE ONLY - never imported by radixmesh_b200/.

CPU restatement of the reference's inbound oplog decode, for timing beside the packed ingest path
(tools/bench_ingest.py) on machines where /root/reference does not exist:

  JsonSerializer.deserialize  (communication/serializer.py:29-35): buf[:size].decode('utf-8') then
  CacheOplog.model_validate_json(data)  (radix/cache_oplog.py:48-56: a pydantic BaseModel with
  oplog_type: IntEnum, node_rank: int, local_logic_id: int = 0, key: List = [], value: List = [], ttl: int,
  gc_query: List[GCQuery] = [], gc_exec: List[ImmutableNodeKey] = []).

Parity pinned: tests/test_oracle_vs_reference.py::test_oplog_wire_matches_reference compares this model, the
reference's and radixmesh_b200.oplog_codec.JsonSerializer on the same bytes; tests/golden/oplog_wire.json holds
messages recorded from the reference.
"""
import enum
from typing import List, Tuple

from pydantic import BaseModel


class PortOplogType(enum.IntEnum):        # cache_oplog.py:13-22
    INSERT = 1
    DELETE = 2
    RESET = 3
    GC_QUERY = 4
    GC_EXEC = 5
    TICK = 10


class PortNodeKey(BaseModel):             # cache_oplog.py:25-28
    key: Tuple[int, ...]
    node_rank: int
    key_hash: int


class PortGCQuery(BaseModel):             # cache_oplog.py:43-45
    agree: int = 1
    key: PortNodeKey


class PortCacheOplog(BaseModel):          # cache_oplog.py:48-56
    oplog_type: PortOplogType
    node_rank: int
    local_logic_id: int = 0
    key: List = []
    value: List = []
    ttl: int
    gc_query: List[PortGCQuery] = []
    gc_exec: List[PortNodeKey] = []


def port_deserialize(buf, size: int) -> PortCacheOplog:
    """serializer.py:29-35"""
    data = bytes(buf[:size]).decode("utf-8")
    return PortCacheOplog.model_validate_json(data)
