"""ctypes shim over oracle/liboracle.so (CPU ORACLE — test infrastructure, not product code;
see the header of oracle.c).  Built by `make -C oracle` (also done by __graft_entry__.build())."""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class OracleXfer(ctypes.Structure):
    _fields_ = [("src", ctypes.c_int32), ("dst", ctypes.c_int32), ("layer", ctypes.c_int32),
                ("_pad", ctypes.c_int32), ("off", ctypes.c_int64), ("size", ctypes.c_int64),
                ("total", ctypes.c_int64)]


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            subprocess.check_call(["make", "-C", _HERE, "-s"])
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_fill.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_int64, ctypes.c_uint64, ctypes.c_void_p]
        _LIB.oracle_checksum.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]
        _LIB.oracle_fill_range.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_int64, ctypes.c_int64, ctypes.c_uint64,
                                           ctypes.c_void_p]
        _LIB.oracle_fill_checksum.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_int64, ctypes.c_uint64, ctypes.c_int64,
                                              ctypes.c_void_p]
        _LIB.oracle_tcp_run.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(OracleXfer),
                                        ctypes.c_int, ctypes.POINTER(ctypes.c_void_p), ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)]
        _LIB.oracle_tcp_run_ex.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(OracleXfer),
                                           ctypes.c_int, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_char_p),
                                           ctypes.POINTER(ctypes.c_int64), ctypes.c_int,
                                           ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)]
    return _LIB


def fill_bytes(kind: int, layer_id: int, size: int, seed: int = 0) -> np.ndarray:
    out = np.empty(max(size, 0), dtype=np.uint8)
    rc = lib().oracle_fill(kind, layer_id, size, seed & ((1 << 64) - 1), out.ctypes.data)
    if rc != 0:
        raise ValueError("oracle_fill failed")
    return out


def checksum(data: np.ndarray, chunk_bytes: int = 256 * 1024) -> np.ndarray:
    data = np.ascontiguousarray(data, dtype=np.uint8)
    n_chunks = (data.size + chunk_bytes - 1) // chunk_bytes
    out = np.zeros((n_chunks, 2), dtype=np.uint64)
    rc = lib().oracle_checksum(data.ctypes.data, data.size, chunk_bytes, out.ctypes.data)
    if rc != 0:
        raise ValueError("oracle_checksum failed")
    return out


def fill_range(kind: int, layer_id: int, off: int, length: int, seed: int = 0) -> np.ndarray:
    """fill_bytes(kind, layer_id, off + length, seed)[off:] for off % 8 == 0, without building the bytes before `off`."""
    out = np.empty(max(length, 0), dtype=np.uint8)
    rc = lib().oracle_fill_range(kind, layer_id, off, max(length, 0), seed & ((1 << 64) - 1), out.ctypes.data)
    if rc != 0:
        raise ValueError("oracle_fill_range failed (offset must be a multiple of 8)")
    return out


def fill_checksum(kind: int, layer_id: int, size: int, seed: int = 0, chunk_bytes: int = 256 * 1024) -> np.ndarray:
    """checksum(fill_bytes(kind, layer_id, size, seed)) without building the layer in memory (one chunk at a time)."""
    n_chunks = (max(size, 0) + chunk_bytes - 1) // chunk_bytes
    out = np.zeros((n_chunks, 2), dtype=np.uint64)
    rc = lib().oracle_fill_checksum(kind, layer_id, max(size, 0), seed & ((1 << 64) - 1), chunk_bytes, out.ctypes.data)
    if rc != 0:
        raise ValueError("oracle_fill_checksum failed")
    return out


def fill_checksum_many(kind: int, layers: Sequence[int], size: int, seed: int = 0, threads: int = 0) -> dict:
    """{layer: fill_checksum(...)} computed on `threads` host threads (ctypes releases the GIL inside the C call)."""
    from concurrent.futures import ThreadPoolExecutor
    layers = list(layers)
    if not layers:
        return {}
    threads = threads or min(len(layers), max(1, (os.cpu_count() or 2) // 2))
    with ThreadPoolExecutor(max_workers=threads) as ex:
        tabs = list(ex.map(lambda l: fill_checksum(kind, l, size, seed), layers))
    return dict(zip(layers, tabs))


def tcp_run(n_nodes: int, leader: int, xfers: Sequence, src_layers: Sequence[np.ndarray],
            base_port: int = 18080, verify: bool = True, src_paths: Optional[Sequence[Optional[str]]] = None,
            rates: Optional[Sequence[int]] = None):
    """Run the transfer list once over loopback TCP.  `src_layers[i]` is the WHOLE source layer of
    transfer i (uint8 array).  `src_paths[i]` (optional) names a file that holds that layer instead (DiskLayer: the
    sender copies the section out of the file; `src_layers[i]` is then only what verification compares with, and may
    be None).  `rates[i]` (optional) is the transfer's LimitRate in bytes/s (<= 0: no limiter).
    Returns (seconds, bytes_moved)."""
    n = len(xfers)
    arr = (OracleXfer * max(n, 1))()
    ptrs = (ctypes.c_void_p * max(n, 1))()
    paths = (ctypes.c_char_p * max(n, 1))()
    lims = (ctypes.c_int64 * max(n, 1))()
    for i, x in enumerate(xfers):
        arr[i] = OracleXfer(x.src, x.dst, x.layer, 0, x.off, x.size, x.total)
        ptrs[i] = src_layers[i].ctypes.data if src_layers[i] is not None else None
        paths[i] = src_paths[i].encode() if src_paths and src_paths[i] else None
        lims[i] = int(rates[i]) if rates else 0
        if ptrs[i] is None and not paths[i]:
            raise ValueError(f"transfer {i} has neither source bytes nor a source file")
    sec = ctypes.c_double(0.0)
    moved = ctypes.c_int64(0)
    # a port of the block may be taken on a shared box: try a few blocks before giving up
    rc = -1
    for attempt in range(6):
        port = base_port + attempt * 997
        if port + n_nodes >= 65000:
            port = 20000 + (port % 20000)
        rc = lib().oracle_tcp_run_ex(n_nodes, leader, port, arr, n, ptrs, paths, lims, int(verify),
                                     ctypes.byref(sec), ctypes.byref(moved))
        if rc == 0 or rc == -2:
            break
    if rc != 0:
        raise RuntimeError(f"oracle_tcp_run failed rc={rc}" + (" (delivered bytes differ from the source)" if rc == -2 else ""))
    return sec.value, moved.value
