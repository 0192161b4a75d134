"""ctypes binding of include/dissem.h.  No fallback: a missing libdissem.so raises at load time."""
from __future__ import annotations

import ctypes
import os
from typing import Iterable, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CHUNK_BYTES = 262144
FILL_ZERO, FILL_PATTERN, FILL_POISON = 0, 1, 2

_LIB: Optional[ctypes.CDLL] = None


class DissemError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"dissem error {code}: {msg}")
        self.code = code


class Xfer(ctypes.Structure):
    """dissem_xfer: one layerMsg (tempLayerInfo, distributor/transport.go:47-54)."""
    _fields_ = [("src_node", ctypes.c_int32), ("dst_node", ctypes.c_int32), ("layer", ctypes.c_int32),
                ("flags", ctypes.c_int32), ("off", ctypes.c_int64), ("size", ctypes.c_int64),
                ("total", ctypes.c_int64)]

    def __repr__(self):
        return f"Xfer({self.src_node}->{self.dst_node} L{self.layer} [{self.off},+{self.size})/{self.total})"


def lib_path() -> str:
    return os.path.join(_HERE, "libdissem.so")


def load_library() -> ctypes.CDLL:
    """Load libdissem.so (built in-tree by `make -C distributed-llm-dissemination_b200` or
    __graft_entry__.build()).  Raises if it is missing — there is no CPU implementation to fall back to."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise ImportError(f"{path} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          f"(the engine is CUDA-only; there is no CPU fallback)")
    L = ctypes.CDLL(path)
    c = ctypes
    P = c.c_void_p
    i64 = c.c_int64
    sig = {
        "dissem_version": (c.c_int, []),
        "dissem_last_error": (c.c_char_p, []),
        "dissem_device_count": (c.c_int, []),
        "dissem_init": (c.c_int, [c.c_int, c.POINTER(c.c_int), c.POINTER(P)]),
        "dissem_close": (c.c_int, [P]),
        "dissem_n_nodes": (c.c_int, [P]),
        "dissem_layer_alloc": (c.c_int, [P, c.c_int, c.c_int, i64]),
        "dissem_layer_free": (c.c_int, [P, c.c_int, c.c_int]),
        "dissem_layer_info": (c.c_int, [P, c.c_int, c.c_int, c.POINTER(i64), c.POINTER(P)]),
        "dissem_layer_fill": (c.c_int, [P, c.c_int, c.c_int, c.c_int, c.c_uint64]),
        "dissem_layer_reset": (c.c_int, [P, c.c_int, c.c_int]),
        "dissem_layer_write": (c.c_int, [P, c.c_int, c.c_int, P, i64, i64]),
        "dissem_layer_read": (c.c_int, [P, c.c_int, c.c_int, P, i64, i64]),
        "dissem_host_alloc": (c.c_int, [c.POINTER(P), i64]),
        "dissem_host_free": (c.c_int, [P]),
        "dissem_layer_reserve": (c.c_int, [P, c.c_int, i64, c.c_int]),
        "dissem_layer_attach_host": (c.c_int, [P, c.c_int, c.c_int, P, i64]),
        "dissem_layer_attach_file": (c.c_int, [P, c.c_int, c.c_int, c.c_char_p, i64, i64]),
        "dissem_layer_detach_file": (c.c_int, [P, c.c_int, c.c_int]),
        "dissem_layer_file_info": (c.c_int, [P, c.c_int, c.c_int, c.POINTER(i64)]),
        "dissem_layer_export": (c.c_int, [P, c.c_int, c.c_int, P]),
        "dissem_layer_import": (c.c_int, [P, c.c_int, c.c_int, i64, P]),
        "dissem_layer_export_fd": (c.c_int, [P, c.c_int, c.c_int, c.POINTER(c.c_int)]),
        "dissem_layer_import_fd": (c.c_int, [P, c.c_int, c.c_int, i64, c.c_int]),
        "dissem_mcast_create": (c.c_int, [P, c.c_int, c.c_int, c.POINTER(c.c_int), c.c_int, c.POINTER(c.c_int)]),
        "dissem_mcast_join": (c.c_int, [P, c.c_int, c.c_int, c.POINTER(c.c_int), c.c_int, c.c_int, c.c_int]),
        "dissem_mcast_bind": (c.c_int, [P, c.c_int, c.c_int, c.POINTER(c.c_int), c.c_int]),
        "dissem_send": (c.c_int, [P, c.c_int, c.c_int, c.c_int, i64, i64, i64, c.POINTER(i64)]),
        "dissem_send_file": (c.c_int, [P, c.c_int, c.c_int, c.c_int, i64, i64, i64, c.POINTER(i64)]),
        "dissem_schedule_submit": (c.c_int, [P, c.POINTER(Xfer), c.c_int, c.POINTER(i64)]),
        "dissem_schedule_compile": (c.c_int, [P, c.POINTER(Xfer), c.c_int, c.POINTER(i64)]),
        "dissem_schedule_launch": (c.c_int, [P, i64, c.POINTER(i64)]),
        "dissem_schedule_free": (c.c_int, [P, i64]),
        "dissem_schedule_info": (c.c_int, [P, i64, c.POINTER(i64), c.POINTER(c.c_int), c.POINTER(i64)]),
        "dissem_plan_directions": (c.c_int, [c.POINTER(Xfer), c.c_int, c.POINTER(c.c_int), c.c_int]),
        "dissem_plan_align": (c.c_int, [c.POINTER(Xfer), c.c_int, c.POINTER(c.c_int)]),
        "dissem_plan_localize": (c.c_int, [c.POINTER(Xfer), c.c_int, c.POINTER(c.c_int), c.c_int, c.POINTER(c.c_int), c.POINTER(c.c_int),
                                           c.c_int, c.POINTER(c.c_int)]),
        "dissem_plan_relay": (c.c_int, [c.POINTER(Xfer), c.c_int, c.c_int, c.POINTER(c.POINTER(Xfer)), c.POINTER(c.c_int)]),
        "dissem_plan_optimize": (c.c_int, [c.POINTER(Xfer), c.c_int, c.POINTER(c.c_int), c.c_int, c.POINTER(c.POINTER(Xfer)),
                                           c.POINTER(c.c_int), c.POINTER(c.c_int)]),
        "dissem_node_device": (c.c_int, [P, c.c_int, c.POINTER(c.c_int)]),
        "dissem_wait": (c.c_int, [P, i64, c.POINTER(c.c_double)]),
        "dissem_wait_detail": (c.c_int, [P, i64, c.POINTER(c.c_double), c.POINTER(c.c_double)]),
        "dissem_poll": (c.c_int, [P, i64, c.POINTER(c.c_int)]),
        "dissem_ticket_release": (c.c_int, [P, i64]),
        "dissem_layers_checksums": (c.c_int, [P, c.POINTER(c.c_int), c.POINTER(c.c_int), c.c_int, P, i64]),
        "dissem_layer_checksums": (c.c_int, [P, c.c_int, c.c_int, P, i64]),
        "dissem_layer_verify": (c.c_int, [P, c.c_int, c.c_int, P, i64]),
        "dissem_layer_received": (c.c_int, [P, c.c_int, c.c_int, c.POINTER(i64)]),
        "dissem_checksum_noalloc": (c.c_int, [c.c_int, P, i64, P, i64, P]),
        "dissem_set_option": (c.c_int, [P, c.c_char_p, i64]),
        "dissem_get_option": (c.c_int, [P, c.c_char_p, c.POINTER(i64)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _LIB = L
    return L


def lib() -> ctypes.CDLL:
    return load_library()


def _check(rc: int):
    if rc != 0:
        raise DissemError(rc, (lib().dissem_last_error() or b"").decode())


def make_xfers(xfers: Iterable) -> "ctypes.Array[Xfer]":
    """Accepts Xfer structs or any objects/tuples with (src, dst, layer, off, size, total)."""
    xs = list(xfers)
    arr = (Xfer * max(len(xs), 1))()
    for i, x in enumerate(xs):
        if isinstance(x, Xfer):
            arr[i] = x
        elif isinstance(x, (tuple, list)):
            s, d, l, off, size, total = x
            arr[i] = Xfer(s, d, l, 0, off, size, total)
        else:
            arr[i] = Xfer(x.src, x.dst, x.layer, getattr(x, "flags", 0), x.off, x.size, x.total)
    return arr


XFER_PULL = 1
XFER_FILE = 2


def plan_directions(xfers: List["Xfer"], gpu_of_node: Sequence[int]) -> List["Xfer"]:
    """dissem_plan_directions: set DISSEM_XFER_PULL on the transfers that are better pulled by their destination
    GPU (reverse NVLink direction idle), clear it on the others.  Returns new Xfer structs."""
    arr = make_xfers(xfers)
    g = (ctypes.c_int * len(gpu_of_node))(*gpu_of_node)
    _check(lib().dissem_plan_directions(arr, len(xfers), g, len(gpu_of_node)))
    return [Xfer(a.src_node, a.dst_node, a.layer, a.flags, a.off, a.size, a.total) for a in arr[:len(xfers)]]


def plan_align(xfers: List["Xfer"]) -> List["Xfer"]:
    """dissem_plan_align: fragment cuts moved down to checksum-chunk boundaries (empty fragments dropped)."""
    arr = make_xfers(xfers)
    m = ctypes.c_int()
    _check(lib().dissem_plan_align(arr, len(xfers), ctypes.byref(m)))
    return [Xfer(a.src_node, a.dst_node, a.layer, a.flags, a.off, a.size, a.total) for a in arr[:m.value]]


def plan_localize(xfers: List["Xfer"], gpu_of_node: Sequence[int], holders: Sequence) -> List["Xfer"]:
    """dissem_plan_localize: transfers that cross GPUs although a node on the destination's GPU holds the layer are
    re-sourced from that node.  holders: (node, layer) pairs of HBM-resident whole layers."""
    arr = make_xfers(xfers)
    holders = list(holders)
    g = (ctypes.c_int * len(gpu_of_node))(*gpu_of_node)
    hn = (ctypes.c_int * max(len(holders), 1))(*[h[0] for h in holders])
    hl = (ctypes.c_int * max(len(holders), 1))(*[h[1] for h in holders])
    m = ctypes.c_int()
    _check(lib().dissem_plan_localize(arr, len(xfers), g, len(gpu_of_node), hn, hl, len(holders), ctypes.byref(m)))
    return [Xfer(a.src_node, a.dst_node, a.layer, a.flags, a.off, a.size, a.total) for a in arr[:m.value]]


def xfer_round(x: "Xfer") -> int:
    return (x.flags >> 8) & 0xFFFF


def plan_relay(xfers: List["Xfer"], pieces: int = 16) -> List["Xfer"]:
    """dissem_plan_relay: multi-destination transfers become piece-pipelined chains over rounds."""
    arr = make_xfers(xfers)
    out = ctypes.POINTER(Xfer)()
    n_out = ctypes.c_int()
    L = lib()
    _check(L.dissem_plan_relay(arr, len(xfers), pieces, ctypes.byref(out), ctypes.byref(n_out)))
    res = [Xfer(out[i].src_node, out[i].dst_node, out[i].layer, out[i].flags, out[i].off, out[i].size, out[i].total)
           for i in range(n_out.value)]
    L.dissem_free.restype = None
    L.dissem_free.argtypes = [ctypes.c_void_p]
    L.dissem_free(ctypes.cast(out, ctypes.c_void_p))
    return res


def plan_optimize(xfers: List["Xfer"], gpu_of_node: Sequence[int]):
    """dissem_plan_optimize: relay chains when they unload the busiest port, then push/pull per transfer.
    Returns (transfers, relay_pieces) with relay_pieces == 0 when the plan was left as unicast."""
    arr = make_xfers(xfers)
    g = (ctypes.c_int * len(gpu_of_node))(*gpu_of_node)
    out = ctypes.POINTER(Xfer)()
    n_out, pieces = ctypes.c_int(), ctypes.c_int()
    L = lib()
    _check(L.dissem_plan_optimize(arr, len(xfers), g, len(gpu_of_node), ctypes.byref(out), ctypes.byref(n_out), ctypes.byref(pieces)))
    res = [Xfer(out[i].src_node, out[i].dst_node, out[i].layer, out[i].flags, out[i].off, out[i].size, out[i].total)
           for i in range(n_out.value)]
    L.dissem_free.restype = None
    L.dissem_free.argtypes = [ctypes.c_void_p]
    L.dissem_free(ctypes.cast(out, ctypes.c_void_p))
    return res, pieces.value


class HostBuffer:
    """Pinned, GPU-mapped host memory from dissem_host_alloc, viewable as a numpy array."""

    def __init__(self, size: int):
        self.size = size
        self._p = ctypes.c_void_p()
        _check(lib().dissem_host_alloc(ctypes.byref(self._p), size))
        self.ptr = self._p.value
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * max(size, 1)).from_address(self.ptr))[:size]

    def free(self):
        if self._p:
            self.array = None
            _check(lib().dissem_host_free(self._p))
            self._p = ctypes.c_void_p()


NOALLOC_MAX_CHUNKS = 8192


def checksum_noalloc(device: int, device_ptr: int, size: int, stream: Optional[int] = None) -> np.ndarray:
    """dissem_checksum_noalloc: per-chunk (A, B) of a device buffer the caller owns, computed without any
    device allocation (for hosts whose HBM is fully committed to their own pool, or that cannot tolerate
    cudaMalloc's stalls).  One call at a time per GPU; at most 2 GiB."""
    n = (size + CHUNK_BYTES - 1) // CHUNK_BYTES
    out = np.zeros((n, 2), dtype=np.uint64)
    _check(lib().dissem_checksum_noalloc(device, device_ptr, size, out.ctypes.data, n, stream))
    return out


class Engine:
    """One dissem_ctx.  `device_of_node[i]` is the CUDA ordinal of node i, or -1 for a node that lives in
    another process."""

    def __init__(self, device_of_node: Sequence[int]):
        L = lib()
        n = len(device_of_node)
        arr = (ctypes.c_int * n)(*device_of_node)
        self._ctx = ctypes.c_void_p()
        _check(L.dissem_init(n, arr, ctypes.byref(self._ctx)))
        self.n_nodes = n
        self.device_of_node = list(device_of_node)

    def close(self):
        if self._ctx:
            _check(lib().dissem_close(self._ctx))
            self._ctx = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- store ----
    def alloc(self, node: int, layer: int, size: int):
        _check(lib().dissem_layer_alloc(self._ctx, node, layer, size))

    def reserve(self, node: int, size: int, count: int = 1):
        """Set aside destination buffers so that later alloc() calls of this size skip cudaMalloc."""
        _check(lib().dissem_layer_reserve(self._ctx, node, size, count))

    def free(self, node: int, layer: int):
        _check(lib().dissem_layer_free(self._ctx, node, layer))

    def has(self, node: int, layer: int) -> bool:
        return lib().dissem_layer_info(self._ctx, node, layer, None, None) == 0

    def info(self, node: int, layer: int):
        size = ctypes.c_int64()
        ptr = ctypes.c_void_p()
        _check(lib().dissem_layer_info(self._ctx, node, layer, ctypes.byref(size), ctypes.byref(ptr)))
        return size.value, ptr.value

    def fill(self, node: int, layer: int, kind: int, seed: int = 0):
        _check(lib().dissem_layer_fill(self._ctx, node, layer, kind, seed & ((1 << 64) - 1)))

    def reset(self, node: int, layer: int):
        _check(lib().dissem_layer_reset(self._ctx, node, layer))

    def write(self, node: int, layer: int, data, off: int = 0, ptr: Optional[int] = None, size: Optional[int] = None):
        if ptr is None:
            data = np.ascontiguousarray(data, dtype=np.uint8)
            ptr, size = data.ctypes.data, data.size
        _check(lib().dissem_layer_write(self._ctx, node, layer, ptr, off, size))

    def read(self, node: int, layer: int, off: int = 0, size: Optional[int] = None) -> np.ndarray:
        if size is None:
            size = self.info(node, layer)[0] - off
        out = np.empty(size, dtype=np.uint8)
        _check(lib().dissem_layer_read(self._ctx, node, layer, out.ctypes.data, off, size))
        return out

    def attach_host(self, node: int, layer: int, ptr: int, size: int):
        """Register pinned, GPU-mapped host memory (dissem_host_alloc) as the node's copy of the layer (source only)."""
        _check(lib().dissem_layer_attach_host(self._ctx, node, layer, ptr, size))

    def attach_file(self, node: int, layer: int, path: str, size: int, file_off: int = 0):
        """The node holds the layer on disk (DiskLayer): `size` bytes of `path` from `file_off`."""
        _check(lib().dissem_layer_attach_file(self._ctx, node, layer, path.encode(), file_off, size))

    def detach_file(self, node: int, layer: int):
        _check(lib().dissem_layer_detach_file(self._ctx, node, layer))

    def file_size(self, node: int, layer: int) -> int:
        n = ctypes.c_int64()
        _check(lib().dissem_layer_file_info(self._ctx, node, layer, ctypes.byref(n)))
        return n.value

    def export(self, node: int, layer: int) -> bytes:
        buf = (ctypes.c_uint8 * 64)()
        _check(lib().dissem_layer_export(self._ctx, node, layer, buf))
        return bytes(buf)

    def import_(self, node: int, layer: int, size: int, handle: bytes):
        buf = (ctypes.c_uint8 * 64).from_buffer_copy(handle)
        _check(lib().dissem_layer_import(self._ctx, node, layer, size, buf))

    # ---- shareable VMM layers and cross-process multicast (file descriptors travel over a unix socket) ----
    def export_fd(self, node: int, layer: int) -> int:
        fd = ctypes.c_int(-1)
        _check(lib().dissem_layer_export_fd(self._ctx, node, layer, ctypes.byref(fd)))
        return fd.value

    def import_fd(self, node: int, layer: int, size: int, fd: int):
        _check(lib().dissem_layer_import_fd(self._ctx, node, layer, size, fd))

    def mcast_create(self, layer: int, src: int, dsts: Sequence[int]) -> int:
        arr = (ctypes.c_int * len(dsts))(*dsts)
        fd = ctypes.c_int(-1)
        _check(lib().dissem_mcast_create(self._ctx, layer, src, arr, len(dsts), ctypes.byref(fd)))
        return fd.value

    def mcast_join(self, layer: int, src: int, dsts: Sequence[int], my_node: int, fd: int):
        arr = (ctypes.c_int * len(dsts))(*dsts)
        _check(lib().dissem_mcast_join(self._ctx, layer, src, arr, len(dsts), my_node, fd))

    def mcast_bind(self, layer: int, src: int, dsts: Sequence[int]):
        arr = (ctypes.c_int * len(dsts))(*dsts)
        _check(lib().dissem_mcast_bind(self._ctx, layer, src, arr, len(dsts)))

    # ---- transfers ----
    def send(self, src: int, dst: int, layer: int, off: int, size: int, total: int) -> int:
        t = ctypes.c_int64()
        _check(lib().dissem_send(self._ctx, src, dst, layer, off, size, total, ctypes.byref(t)))
        return t.value

    def send_file(self, src: int, dst: int, layer: int, off: int, size: int, total: int) -> int:
        """dissem_send from the source node's disk copy; blocks while the file is read."""
        t = ctypes.c_int64()
        _check(lib().dissem_send_file(self._ctx, src, dst, layer, off, size, total, ctypes.byref(t)))
        return t.value

    def submit(self, xfers) -> int:
        arr = xfers if isinstance(xfers, ctypes.Array) else make_xfers(xfers)
        n = len(xfers) if not isinstance(xfers, ctypes.Array) else len(arr)
        t = ctypes.c_int64()
        _check(lib().dissem_schedule_submit(self._ctx, arr, n, ctypes.byref(t)))
        return t.value

    def compile(self, xfers) -> int:
        xs = list(xfers)
        arr = make_xfers(xs)
        s = ctypes.c_int64()
        _check(lib().dissem_schedule_compile(self._ctx, arr, len(xs), ctypes.byref(s)))
        return s.value

    def launch(self, sched: int) -> int:
        t = ctypes.c_int64()
        _check(lib().dissem_schedule_launch(self._ctx, sched, ctypes.byref(t)))
        return t.value

    def sched_free(self, sched: int):
        _check(lib().dissem_schedule_free(self._ctx, sched))

    def sched_info(self, sched: int):
        units, launches, nbytes = ctypes.c_int64(), ctypes.c_int(), ctypes.c_int64()
        _check(lib().dissem_schedule_info(self._ctx, sched, ctypes.byref(units), ctypes.byref(launches),
                                          ctypes.byref(nbytes)))
        return units.value, launches.value, nbytes.value

    def wait(self, ticket: int) -> float:
        ms = ctypes.c_double()
        _check(lib().dissem_wait(self._ctx, ticket, ctypes.byref(ms)))
        return ms.value

    def wait_detail(self, ticket: int):
        """-> (batch ms, copy-kernel ms), both device time from CUDA events on the engine's stream."""
        ms, cms = ctypes.c_double(), ctypes.c_double()
        _check(lib().dissem_wait_detail(self._ctx, ticket, ctypes.byref(ms), ctypes.byref(cms)))
        return ms.value, cms.value

    def poll(self, ticket: int) -> bool:
        d = ctypes.c_int()
        _check(lib().dissem_poll(self._ctx, ticket, ctypes.byref(d)))
        return bool(d.value)

    # ---- verification ----
    def _n_chunks(self, node, layer):
        return (self.info(node, layer)[0] + CHUNK_BYTES - 1) // CHUNK_BYTES

    def checksums(self, node: int, layer: int) -> np.ndarray:
        n = self._n_chunks(node, layer)
        out = np.zeros((n, 2), dtype=np.uint64)
        _check(lib().dissem_layer_checksums(self._ctx, node, layer, out.ctypes.data, n))
        return out

    def checksums_many(self, pairs: Sequence) -> List[np.ndarray]:
        """dissem_layers_checksums: the fused tables of several local (node, layer) pairs with one read-back per GPU."""
        pairs = list(pairs)
        counts = [self._n_chunks(n, l) for (n, l) in pairs]
        out = np.zeros((sum(counts), 2), dtype=np.uint64)
        nodes = (ctypes.c_int * max(len(pairs), 1))(*[n for (n, _l) in pairs])
        layers = (ctypes.c_int * max(len(pairs), 1))(*[l for (_n, l) in pairs])
        _check(lib().dissem_layers_checksums(self._ctx, nodes, layers, len(pairs), out.ctypes.data, sum(counts)))
        res, pos = [], 0
        for c in counts:
            res.append(out[pos:pos + c])
            pos += c
        return res

    def release(self, ticket: int):
        _check(lib().dissem_ticket_release(self._ctx, ticket))

    def verify(self, node: int, layer: int) -> np.ndarray:
        n = self._n_chunks(node, layer)
        out = np.zeros((n, 2), dtype=np.uint64)
        _check(lib().dissem_layer_verify(self._ctx, node, layer, out.ctypes.data, n))
        return out

    def received(self, node: int, layer: int) -> int:
        v = ctypes.c_int64()
        _check(lib().dissem_layer_received(self._ctx, node, layer, ctypes.byref(v)))
        return v.value

    def set_option(self, key: str, value: int):
        _check(lib().dissem_set_option(self._ctx, key.encode(), value))

    def get_option(self, key: str) -> int:
        v = ctypes.c_int64()
        _check(lib().dissem_get_option(self._ctx, key.encode(), ctypes.byref(v)))
        return v.value
