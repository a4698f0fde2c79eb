"""ctypes binding of include/radixmesh_b200.h plus a thin array-level wrapper (`DeviceIndex`).

The shared library is the product: if it is missing or was not built for this GPU the import of
the package fails loudly — there is no CPU fallback anywhere in the package.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "_lib", "libradixmesh_b200.so")

RM_OK, RM_E_INVALID, RM_E_CUDA, RM_E_NOSPACE, RM_E_TOO_LARGE, RM_E_MODE, RM_E_TABLE = range(7)
MODE_BASE, MODE_PREFILL, MODE_DECODE, MODE_ROUTER = range(4)


class RmConfig(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("device", C.c_int32), ("mode", C.c_int32), ("self_rank", C.c_int32),
                ("n_prefill", C.c_int32), ("n_decode", C.c_int32), ("max_batch", C.c_int32), ("max_dups", C.c_int32),
                ("max_key_len", C.c_int32), ("max_batch_tokens", C.c_int32), ("max_nodes", C.c_int64),
                ("max_tokens", C.c_int64)]


class RmSizes(C.Structure):
    _fields_ = [("evictable", C.c_int64), ("protected_", C.c_int64), ("n_nodes", C.c_int64),
                ("arena_used", C.c_int64), ("n_dups", C.c_int64), ("op_seq", C.c_int64),
                ("device_error", C.c_int32), ("node_hwm", C.c_int32), ("table_slots", C.c_int32),
                ("table_tombs", C.c_int32)]


class RmNodeInfo(C.Structure):
    _fields_ = [("parent", C.c_int32), ("edge_len", C.c_int32), ("lock_ref", C.c_int32), ("rank", C.c_int32),
                ("depth_end", C.c_int32), ("nchild", C.c_int32), ("gen", C.c_int32), ("parked", C.c_int32),
                ("edge_off", C.c_int64), ("stamp", C.c_uint64)]


class RadixMeshError(RuntimeError):
    def __init__(self, status: int, text: str):
        super().__init__(f"radixmesh_b200 status {status}: {text}")
        self.status = status


_P = C.c_void_p
_PROTOS = {
    "rm_last_error": (C.c_char_p, []),
    "rm_version": (C.c_char_p, []),
    "rm_create": (C.c_int, [C.POINTER(RmConfig), C.POINTER(_P)]),
    "rm_destroy": (C.c_int, [_P]),
    "rm_reset": (C.c_int, [_P, _P]),
    "rm_match_batch": (C.c_int, [_P, _P, _P, C.c_int32, _P, _P, _P]),
    "rm_match_batch_h": (C.c_int, [_P, _P, _P, C.c_int32, _P, _P, _P]),
    "rm_match_batch_static": (C.c_int, [_P, _P, _P, C.c_int32, _P, _P, _P]),
    "rm_match_prefix_h": (C.c_int, [_P, _P, C.c_int32, _P, _P, _P, _P, _P]),
    "rm_gather_values": (C.c_int, [_P, _P, _P, _P, C.c_int32, _P, _P]),
    "rm_insert_batch": (C.c_int, [_P, _P, _P, _P, _P, C.c_int32, _P, _P]),
    "rm_insert_batch_h": (C.c_int, [_P, _P, _P, _P, _P, C.c_int32, _P, _P]),
    "rm_lock_batch": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, _P, _P]),
    "rm_evict": (C.c_int, [_P, C.c_int64, C.c_int32, _P, _P, _P, _P, _P]),
    "rm_compact": (C.c_int, [_P, _P]),
    "rm_gc_candidates": (C.c_int, [_P, C.c_int32, _P, _P, _P]),
    "rm_dup_entries": (C.c_int, [_P, C.c_int32, _P, _P, _P]),
    "rm_route_batch": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P]),
    "rm_pick_batch": (C.c_int, [_P, _P, C.c_int32, _P, _P, C.c_int32, _P, _P, _P, _P, _P, _P]),
    "rm_route_set_rings": (C.c_int, [_P, _P, _P, C.c_int32, _P, _P, C.c_int32]),
    "rm_sizes": (C.c_int, [_P, C.POINTER(RmSizes), _P]),
    "rm_dump": (C.c_int, [_P, C.c_int32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "rm_node_info": (C.c_int, [_P, _P, C.c_int32, _P, _P]),
    "rm_node_children": (C.c_int, [_P, C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P]),
    "rm_total_size": (C.c_int, [_P, _P, _P]),
    "rm_dump_order": (C.c_int, [_P, C.c_int32, _P, _P]),
    "rm_pages_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.POINTER(_P)]),
    "rm_pages_destroy": (C.c_int, [_P]),
    "rm_pages_reset": (C.c_int, [_P, _P]),
    "rm_pages_intern": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, _P, C.c_int64, _P, _P, _P, _P]),
    "rm_pages_rows_put": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "rm_pages_rows_free": (C.c_int, [_P, _P, C.c_int64, _P]),
    "rm_pages_rows_get": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "rm_pages_tokens": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "rm_read_runs": (C.c_int, [_P, _P, _P, C.c_int32, _P, _P]),
    "rm_pages_retain": (C.c_int, [_P, _P, _P]),
    "rm_pages_status": (C.c_int, [_P, _P, _P]),
    "rm_read_arena": (C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, _P]),
    "rm_read_segments": (C.c_int, [_P, C.c_int64, C.c_int64, _P, _P]),
    "rm_launch_count": (C.c_uint64, []),
    "rm_gc_keys_h": (C.c_int, [_P, _P, C.c_int32, C.c_int32, _P, _P, _P]),
    "rm_gc_agree_h": (C.c_int, [_P, _P, _P, _P, C.c_int32, _P, _P]),
    "rm_gc_exec_h": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P]),
    "rm_profile": (C.c_int, [_P, C.c_int]),
    "rm_profile_read": (C.c_int, [_P, C.POINTER(C.c_double), C.POINTER(C.c_int)]),
    "rm_profile_stages": (C.c_int, [_P, C.POINTER(C.c_double), C.POINTER(C.c_int)]),
}
EXPORTS = tuple(_PROTOS)


def load_library(path: Optional[str] = None) -> C.CDLL:
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise ImportError(
            f"radixmesh_b200: {path} is missing. Build it with `python radixmesh_b200/build.py` "
            "(nvcc, sm_100a); this package has no CPU fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)   # AttributeError here = the library does not export the ABI
        fn.restype = res
        fn.argtypes = args
    return lib


_lib_lock = threading.Lock()
_lib: Optional[C.CDLL] = None


def lib() -> C.CDLL:
    global _lib
    with _lib_lock:
        if _lib is None:
            _lib = load_library()
        return _lib


def ptr(x) -> Optional[int]:
    """Raw address of a torch tensor / numpy array (None passes NULL)."""
    if x is None:
        return None
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if isinstance(x, int):
        return x
    raise TypeError(type(x))


class TorchCuda:
    """Device-memory plumbing through torch (allocation, streams, copies)."""

    name = "cuda"

    def __init__(self, device: int = 0):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("radixmesh_b200 needs a CUDA device (sm_100a); none is visible")
        self.torch = torch
        self.device = torch.device("cuda", device)
        self.ordinal = device

    def empty(self, n, dtype):
        return self.torch.empty(int(n), dtype=getattr(self.torch, np.dtype(dtype).name), device=self.device)

    def to_device(self, a: np.ndarray):
        return self.torch.from_numpy(np.ascontiguousarray(a)).to(self.device)

    def to_host(self, t) -> np.ndarray:
        return t.cpu().numpy()

    def stream(self) -> int:
        return self.torch.cuda.current_stream(self.device).cuda_stream


class DeviceIndex:
    """One rm_index plus array-level helpers.  Arrays are whatever the backend hands out (torch CUDA
    tensors in the product; tests/emu uses numpy with the host emulation library)."""

    def __init__(self, *, mode=MODE_BASE, max_nodes=1 << 16, max_tokens=1 << 20, max_batch=1 << 12,
                 max_key_len=1 << 13, max_dups=1 << 10, self_rank=0, n_prefill=0, n_decode=0, device=0,
                 max_batch_tokens=0, backend=None, library=None):
        self.be = backend if backend is not None else TorchCuda(device)
        self.lib = library if library is not None else lib()
        self.cfg = RmConfig(C.sizeof(RmConfig), device, mode, self_rank, n_prefill, n_decode, max_batch, max_dups,
                            max_key_len, int(max_batch_tokens), max_nodes, max_tokens)
        self.mode = mode
        self._h = _P()
        self._mutex = threading.RLock()   # one host thread at a time per index (SURVEY 8b threading)
        self._check(self.lib.rm_create(C.byref(self.cfg), C.byref(self._h)))

    # -- plumbing ----------------------------------------------------------
    def _check(self, st: int) -> None:
        if st != RM_OK:
            raise RadixMeshError(st, (self.lib.rm_last_error() or b"").decode())

    def close(self) -> None:
        if self._h:
            self.lib.rm_destroy(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _s(self):
        return self.be.stream()

    # -- operations on device arrays ---------------------------------------
    def reset(self) -> None:
        with self._mutex:
            self._check(self.lib.rm_reset(self._h, self._s()))

    def match(self, tokens, offsets, B, out_len, out_node) -> None:
        with self._mutex:
            self._check(self.lib.rm_match_batch(self._h, ptr(tokens), ptr(offsets), B, ptr(out_len), ptr(out_node), self._s()))

    def match_static(self, tokens, offsets, B, out_len, out_node) -> None:
        """rm_match_batch_static: same as match(), scratch in the library's fixed device array (<= 4096 queries)."""
        with self._mutex:
            self._check(self.lib.rm_match_batch_static(self._h, ptr(tokens), ptr(offsets), B, ptr(out_len), ptr(out_node), self._s()))

    def match_host(self, tokens_h: np.ndarray, offsets_h: np.ndarray, out_len_h: np.ndarray, out_node_h: np.ndarray) -> None:
        B = len(offsets_h) - 1
        with self._mutex:
            self._check(self.lib.rm_match_batch_h(self._h, ptr(tokens_h), ptr(offsets_h), B, ptr(out_len_h), ptr(out_node_h), self._s()))

    def match_prefix_host(self, key_h: np.ndarray):
        """rm_match_prefix_h: one key end to end -> (matched length, last node id, values[:length], RmNodeInfo of the node)."""
        n = len(key_h)
        ln, nd = C.c_int32(0), C.c_int32(0)
        vals = np.empty(max(n, 1), dtype=np.int64)
        info = RmNodeInfo()
        with self._mutex:
            self._check(self.lib.rm_match_prefix_h(self._h, ptr(key_h), n, C.addressof(ln), C.addressof(nd), ptr(vals),
                                                   C.addressof(info), self._s()))
        return ln.value, nd.value, vals[:ln.value], info

    def gather(self, nodes, lens, out_offsets, B, out_values) -> None:
        with self._mutex:
            self._check(self.lib.rm_gather_values(self._h, ptr(nodes), ptr(lens), ptr(out_offsets), B, ptr(out_values), self._s()))

    def insert(self, tokens, values, offsets, ranks, N, out_prefix) -> None:
        with self._mutex:
            self._check(self.lib.rm_insert_batch(self._h, ptr(tokens), ptr(values), ptr(offsets), ptr(ranks), N, ptr(out_prefix), self._s()))

    def insert_host(self, tokens_h, values_h, offsets_h, ranks_h=None) -> np.ndarray:
        N = len(offsets_h) - 1
        out = np.zeros(max(N, 1), dtype=np.int32)
        with self._mutex:
            self._check(self.lib.rm_insert_batch_h(self._h, ptr(tokens_h), ptr(values_h), ptr(offsets_h), ptr(ranks_h), N, ptr(out), self._s()))
        return out[:N]

    def lock(self, nodes, gens, B, direction, out_delta) -> None:
        with self._mutex:
            self._check(self.lib.rm_lock_batch(self._h, ptr(nodes), ptr(gens), B, direction, ptr(out_delta), self._s()))

    def evict(self, num_tokens: int, cap: Optional[int] = None):
        cap = int(cap or self.cfg.max_nodes)
        n = C.c_int32(0)
        with self._mutex:
            # result buffers are kept (zeroing three max_nodes-sized arrays per call cost more than the eviction)
            bufs = getattr(self, "_evict_bufs", None)
            if bufs is None or len(bufs[0]) < cap:
                bufs = (np.empty(cap, dtype=np.int32), np.empty(cap, dtype=np.int64), np.empty(cap, dtype=np.int32))
                self._evict_bufs = bufs
            nodes, offs, lens = bufs
            self._check(self.lib.rm_evict(self._h, int(num_tokens), cap, ptr(nodes), ptr(offs), ptr(lens), C.addressof(n), self._s()))
            k = n.value
            return nodes[:k].copy(), offs[:k].copy(), lens[:k].copy()

    def compact(self) -> None:
        with self._mutex:
            self._check(self.lib.rm_compact(self._h, self._s()))

    def dup_entries(self, only_candidates: bool = False) -> np.ndarray:
        cap = int(self.cfg.max_dups) * 2 + 16
        out = np.zeros((cap, 4), dtype=np.int32)
        n = C.c_int32(0)
        fn = self.lib.rm_gc_candidates if only_candidates else self.lib.rm_dup_entries
        with self._mutex:
            self._check(fn(self._h, cap, ptr(out), C.addressof(n), self._s()))
        return out[:n.value]

    # ---- duplicate-node GC between replicas (host arrays; radix_mesh.py:362-389)
    def gc_keys(self, nodes: np.ndarray):
        """Full keys (token lists) of tree nodes, e.g. column 0 of dup_entries(only_candidates=True)."""
        nodes = np.ascontiguousarray(nodes, dtype=np.int32)
        n = len(nodes)
        if n == 0:
            return []
        stride = int(self.cfg.max_key_len)
        tok = np.zeros((n, stride), dtype=np.int32)
        lens = np.zeros(n, dtype=np.int32)
        with self._mutex:
            self._check(self.lib.rm_gc_keys_h(self._h, ptr(nodes), n, stride, ptr(tok), ptr(lens), self._s()))
        return [tok[i, :lens[i]].tolist() for i in range(n)]

    def _gc_pack(self, keys, ranks):
        offs = np.zeros(len(keys) + 1, dtype=np.int64)
        np.cumsum([len(k) for k in keys], out=offs[1:])
        tok = np.asarray([t for k in keys for t in k], dtype=np.int32) if offs[-1] else np.zeros(1, np.int32)
        return tok, offs, np.ascontiguousarray(ranks, dtype=np.int32)

    def gc_agree(self, keys, ranks) -> np.ndarray:
        n = len(keys)
        out = np.zeros(max(n, 1), dtype=np.uint8)
        if n:
            tok, offs, rk = self._gc_pack(keys, ranks)
            with self._mutex:
                self._check(self.lib.rm_gc_agree_h(self._h, ptr(tok), ptr(offs), ptr(rk), n, ptr(out), self._s()))
        return out[:n].astype(bool)

    def gc_exec(self, keys, ranks, free_parked: bool):
        """Pops the keys; returns (arena offsets, lengths) of the parked runs released (-1 / 0 where none)."""
        n = len(keys)
        off = np.full(max(n, 1), -1, dtype=np.int64)
        ln = np.zeros(max(n, 1), dtype=np.int32)
        if n:
            tok, offs, rk = self._gc_pack(keys, ranks)
            with self._mutex:
                self._check(self.lib.rm_gc_exec_h(self._h, ptr(tok), ptr(offs), ptr(rk), n, 1 if free_parked else 0,
                                                  ptr(off), ptr(ln), self._s()))
        return off[:n], ln[:n]

    def route(self, tokens, offsets, B, warm, out_prefill, out_decode, out_paddr, out_daddr, out_err) -> None:
        with self._mutex:
            self._check(self.lib.rm_route_batch(self._h, ptr(tokens), ptr(offsets), B, 1 if warm else 0, ptr(out_prefill),
                                                ptr(out_decode), ptr(out_paddr), ptr(out_daddr), ptr(out_err), self._s()))

    def pick(self, depths, n_ranks, tokens, offsets, B, out_prefill, out_decode, out_paddr, out_daddr, out_err) -> None:
        with self._mutex:
            self._check(self.lib.rm_pick_batch(self._h, ptr(depths), n_ranks, ptr(tokens), ptr(offsets), B, ptr(out_prefill),
                                               ptr(out_decode), ptr(out_paddr), ptr(out_daddr), ptr(out_err), self._s()))

    def set_rings(self, p_points: np.ndarray, p_owner: np.ndarray, d_points: np.ndarray, d_owner: np.ndarray) -> None:
        p_points = np.ascontiguousarray(p_points, dtype=np.uint32)
        d_points = np.ascontiguousarray(d_points, dtype=np.uint32)
        p_owner = np.ascontiguousarray(p_owner, dtype=np.int32)
        d_owner = np.ascontiguousarray(d_owner, dtype=np.int32)
        with self._mutex:
            self._check(self.lib.rm_route_set_rings(self._h, ptr(p_points), ptr(p_owner), len(p_points),
                                                    ptr(d_points), ptr(d_owner), len(d_points)))

    def sizes(self) -> RmSizes:
        s = RmSizes()
        with self._mutex:
            self._check(self.lib.rm_sizes(self._h, C.byref(s), self._s()))
        return s

    def check_device_error(self) -> None:
        s = self.sizes()
        if s.device_error:
            raise RadixMeshError(s.device_error, "raised inside a kernel (node pool / arena / dup table exhausted "
                                                 "or key longer than max_key_len)")

    def dump(self) -> dict:
        """Host copies of the per-node arrays for ids [0, high-water)."""
        n = self.sizes().node_hwm
        a = {"parent": np.zeros(n, np.int32), "edge_off": np.zeros(n, np.int64), "edge_len": np.zeros(n, np.int32),
             "lock_ref": np.zeros(n, np.int32), "stamp": np.zeros(n, np.uint64), "rank": np.zeros(n, np.int32),
             "depth": np.zeros(n, np.int32), "nchild": np.zeros(n, np.int32), "gen": np.zeros(n, np.int32),
             "parked": np.zeros(n, np.int32)}
        out_n = C.c_int32(0)
        with self._mutex:
            self._check(self.lib.rm_dump(self._h, n, ptr(a["parent"]), ptr(a["edge_off"]), ptr(a["edge_len"]),
                                         ptr(a["lock_ref"]), ptr(a["stamp"]), ptr(a["rank"]), ptr(a["depth"]),
                                         ptr(a["nchild"]), ptr(a["gen"]), ptr(a["parked"]), C.addressof(out_n), self._s()))
        a["n"] = n
        return a

    def node_info(self, ids):
        """rm_node_info: one RmNodeInfo per node id (a few dozen bytes each, whatever the size of the tree)."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n = len(ids)
        out = (RmNodeInfo * max(n, 1))()
        if n:
            with self._mutex:
                self._check(self.lib.rm_node_info(self._h, ptr(ids), n, C.addressof(out), self._s()))
        return [out[i] for i in range(n)]

    def node_children(self, node: int, cap: int = 1 << 16):
        """rm_node_children: [(child id, first token, generation)] in the reference's dict order."""
        while True:
            ids, tok, gen = (np.zeros(cap, np.int32) for _ in range(3))
            od = np.zeros(cap, np.uint64)
            n = C.c_int32(0)
            with self._mutex:
                self._check(self.lib.rm_node_children(self._h, int(node), cap, ptr(ids), ptr(tok), ptr(gen), ptr(od),
                                                      C.addressof(n), self._s()))
            if n.value <= cap:
                break
            cap = n.value
        k = n.value
        order = np.lexsort((ids[:k], od[:k]))
        return [(int(ids[i]), int(tok[i]), int(gen[i])) for i in order]

    def total_size(self) -> int:
        v = C.c_int64(0)
        with self._mutex:
            self._check(self.lib.rm_total_size(self._h, C.addressof(v), self._s()))
        return int(v.value)

    def dump_order(self, n: int) -> np.ndarray:
        od = np.zeros(max(n, 1), np.uint64)
        with self._mutex:
            self._check(self.lib.rm_dump_order(self._h, int(n), ptr(od), self._s()))
        return od[:n]

    def read_segments(self, off: int, n: int) -> np.ndarray:
        seg = np.zeros(n, np.int32)
        with self._mutex:
            self._check(self.lib.rm_read_segments(self._h, int(off), int(n), ptr(seg), self._s()))
        return seg

    def profile(self, enable) -> None:
        """False/0 off, True/1 time the walk kernel of each match batch, 2 time all four stages."""
        self._check(self.lib.rm_profile(self._h, int(enable)))

    def profile_stages(self):
        """(ms summed over the recorded batches for [walk, single splits, sort, shared splits], batches)."""
        ms, n = (C.c_double * 4)(), C.c_int(0)
        self._check(self.lib.rm_profile_stages(self._h, ms, C.byref(n)))
        return list(ms), n.value

    def profile_read(self):
        ms, n = C.c_double(0.0), C.c_int(0)
        self._check(self.lib.rm_profile_read(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def read_runs(self, offs: np.ndarray, lens: np.ndarray) -> np.ndarray:
        """Values of the arena runs (offs[i], lens[i]), concatenated: one gather, one copy (rm_read_runs)."""
        offs = np.ascontiguousarray(offs, dtype=np.int64)
        lens = np.ascontiguousarray(lens, dtype=np.int32)
        out = np.empty(int(lens.sum()), dtype=np.int64)
        if len(lens):
            with self._mutex:
                self._check(self.lib.rm_read_runs(self._h, ptr(offs), ptr(lens), len(lens), ptr(out), self._s()))
        return out

    def read_arena(self, off: int, n: int, values: bool = True):
        tok = np.zeros(n, np.int32)
        val = np.zeros(n, np.int64) if values and self.mode != MODE_ROUTER else None
        with self._mutex:
            self._check(self.lib.rm_read_arena(self._h, int(off), int(n), ptr(tok), ptr(val), self._s()))
        return tok, val


class DevicePages:
    """rm_pages: page interning table + value rows for RadixCache(page_size > 1), beside a BASE-mode DeviceIndex that
    runs on page symbols.  Arrays are the backend's (torch CUDA tensors in the product)."""

    def __init__(self, page_size: int, max_pages: int, max_rows: int, *, device=0, backend=None, library=None):
        self.be = backend if backend is not None else TorchCuda(device)
        self.lib = library if library is not None else lib()
        self.P = int(page_size)
        self._h = _P()
        self._mutex = threading.RLock()
        self._check(self.lib.rm_pages_create(int(device), self.P, int(max_pages), int(max_rows), C.byref(self._h)))

    def _check(self, st: int) -> None:
        if st != RM_OK:
            raise RadixMeshError(st, (self.lib.rm_last_error() or b"").decode())

    def _s(self):
        return self.be.stream()

    def close(self) -> None:
        if self._h:
            self.lib.rm_pages_destroy(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self) -> None:
        with self._mutex:
            self._check(self.lib.rm_pages_reset(self._h, self._s()))

    def intern(self, tokens, offsets, B, create, sym_offsets, total_pages, out_syms, out_npages, scratch) -> None:
        with self._mutex:
            self._check(self.lib.rm_pages_intern(self._h, ptr(tokens), ptr(offsets), B, 1 if create else 0, ptr(sym_offsets),
                                                 int(total_pages), ptr(out_syms), ptr(out_npages), ptr(scratch), self._s()))

    def rows_put(self, values, n_rows, out_rows) -> None:
        with self._mutex:
            self._check(self.lib.rm_pages_rows_put(self._h, ptr(values), int(n_rows), ptr(out_rows), self._s()))

    def rows_free(self, rows, n) -> None:
        with self._mutex:
            self._check(self.lib.rm_pages_rows_free(self._h, ptr(rows), int(n), self._s()))

    def rows_get(self, rows, n, out_values) -> None:
        with self._mutex:
            self._check(self.lib.rm_pages_rows_get(self._h, ptr(rows), int(n), ptr(out_values), self._s()))

    def tokens(self, syms, n, out_tokens) -> None:
        with self._mutex:
            self._check(self.lib.rm_pages_tokens(self._h, ptr(syms), int(n), ptr(out_tokens), self._s()))

    def retain(self, index: "DeviceIndex") -> None:
        """Drop the pages no edge of `index` holds any more; their symbols are reused (rm_pages_retain)."""
        with self._mutex:
            self._check(self.lib.rm_pages_retain(self._h, index._h, self._s()))

    def status(self):
        """(symbols in use, free rows, sticky device error, page_size); raises on a device error."""
        out = (C.c_int64 * 4)()
        with self._mutex:
            self._check(self.lib.rm_pages_status(self._h, out, self._s()))
        if out[2]:
            raise RadixMeshError(int(out[2]), "raised inside a page kernel (page table / value rows exhausted, or two new "
                                              "pages of one batch with the same 64-bit hash)")
        return int(out[0]), int(out[1]), int(out[2]), int(out[3])
