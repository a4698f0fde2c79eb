"""Oplog ingest throughput, host side only: inbound INSERT messages -> staged flat arrays.

  reference   the reference's decode, restated (oracle/oplog_port.py): UTF-8 decode + pydantic model_validate_json
              per message, then what RadixMesh.oplog_received does with the object (ttl gate, list(key), value list)
  json        the product's JSON path for unmodified peers: the two integer lists cut out of the text and parsed by
              numpy's C reader (IngestBuffer.feed_wire)
  packed      the product's binary path: header gate + two np.frombuffer copies (IngestBuffer.feed_wire)

Each arm ends with the same arrays (checked).  With --device the packed arm also applies the batch on cuda:0 through
rm_insert_batch_h, to show decode and apply side by side.  Prints one JSON line.

    python tools/bench_ingest.py [--messages 4096] [--tokens 512] [--device]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--messages", type=int, default=4096)
    ap.add_argument("--tokens", type=int, default=512)
    ap.add_argument("--device", action="store_true")
    ap.add_argument("--repeat", type=int, default=3)
    args = ap.parse_args()
    from oracle.oplog_port import port_deserialize
    from radixmesh_b200.host import CacheOplog, CacheOplogType
    from radixmesh_b200.oplog_codec import IngestBuffer, JsonSerializer, PackedSerializer

    rng = np.random.default_rng(0)
    n, L = args.messages, args.tokens
    docs = rng.integers(0, 32000, (64, L), dtype=np.int32)
    ops = []
    for i in range(n):
        cut = int(rng.integers(L // 8, L + 1))
        key = np.concatenate([docs[int(rng.integers(0, 64))][:cut], rng.integers(0, 32000, L - cut, dtype=np.int32)])
        ops.append(CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=int(rng.integers(0, 4)), local_logic_id=i,
                              key=key.tolist(), value=list(range(i * L, (i + 1) * L)), ttl=4))
    js, pk = JsonSerializer(), PackedSerializer()
    wire_json = [js.serialize(o) for o in ops]
    wire_packed = [pk.serialize(o) for o in ops]
    SELF = 7

    def arm_reference():
        ib = IngestBuffer()
        for w in wire_json:
            o = port_deserialize(bytearray(w), len(w))
            o.ttl -= 1
            if o.node_rank == SELF or o.ttl <= 0:
                continue
            ib.add(list(o.key), o.value, o.node_rank)
        return ib.take()

    def arm_json():
        ib = IngestBuffer()
        for w in wire_json:
            ib.feed_wire(w, None, SELF)
        return ib.take()

    def arm_packed():
        ib = IngestBuffer()
        for w in wire_packed:
            ib.feed_wire(w, None, SELF)
        return ib.take()

    res, out = {}, {}
    for name, fn in (("reference", arm_reference), ("json", arm_json), ("packed", arm_packed)):
        best = 1e9
        for _ in range(args.repeat):
            t0 = time.perf_counter()
            out[name] = fn()
            best = min(best, time.perf_counter() - t0)
        res[name] = {"messages_per_s": n / best, "ms_total": 1e3 * best}
    for name in ("json", "packed"):
        for a, b in zip(out["reference"], out[name]):
            assert np.array_equal(a, b), name
    line = {"workload": f"{n} INSERT messages x {L} tokens (i32 key + i64 value), one core",
            "wire_bytes_per_message": {"json": len(wire_json[0]), "packed": len(wire_packed[0])}, "decode": res}
    if args.device:
        import torch
        from radixmesh_b200 import capi
        assert torch.cuda.is_available()
        ix = capi.DeviceIndex(mode=capi.MODE_DECODE, self_rank=SELF, n_prefill=4, n_decode=4, max_nodes=1 << 18,
                              max_tokens=n * L + (1 << 20), max_batch=n, max_key_len=max(L, 64), max_dups=1 << 16)
        tok, val, offs, ranks = out["packed"]
        best = 1e9
        for _ in range(args.repeat):
            ix.reset()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ix.insert_host(tok, val, offs, ranks)
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        ix.check_device_error()
        line["apply_on_device"] = {"records_per_s": n / best, "ms_total": 1e3 * best,
                                   "note": "rm_insert_batch_h: pageable host arrays in, conflict rule + duplicate table on the device"}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
