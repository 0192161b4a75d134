"""ctypes binding of include/dissem_host.h: the C++ mirror of the reference's plan generators and role state
machines.  Plans are pure host logic (no GPU); run_distribution needs a GPU."""
from __future__ import annotations

import ctypes
import json
from typing import List, Optional, Sequence

from ._capi import DissemError, Xfer, lib

POLICY_RANDOM, POLICY_BALANCED = 0, 1
PLAN_ZERO_RATE_UNLIMITED, PLAN_SIZES_FROM_ANNOUNCES, PLAN_MULTI_DEST = 1, 2, 4


class HostError(DissemError):
    pass


def _bind():
    L = lib()
    if not getattr(L, "_host_bound", False):
        c = ctypes
        L.dissem_plan_json.restype = c.c_int
        L.dissem_plan_json.argtypes = [c.c_char_p, c.c_int, c.c_uint64, c.c_int, c.c_int64, c.c_int,
                                       c.POINTER(c.c_void_p)]
        L.dissem_run_json.restype = c.c_int
        L.dissem_run_json.argtypes = [c.c_char_p, c.c_int, c.POINTER(c.c_int), c.c_char_p, c.POINTER(c.c_void_p)]
        L.dissem_msg_roundtrip_json.restype = c.c_int
        L.dissem_msg_roundtrip_json.argtypes = [c.c_char_p, c.POINTER(c.c_void_p), c.POINTER(c.c_void_p)]
        L.dissem_transport_script_json.restype = c.c_int
        L.dissem_transport_script_json.argtypes = [c.c_char_p, c.POINTER(c.c_void_p)]
        L.dissem_pace_schedule.restype = c.c_int
        L.dissem_pace_schedule.argtypes = [c.c_int64, c.c_int64, c.c_int64, c.POINTER(c.c_int64), c.POINTER(c.c_int64),
                                           c.POINTER(c.c_int64), c.c_int, c.POINTER(c.c_int)]
        L.dissem_free.restype = None
        L.dissem_free.argtypes = [c.c_void_p]
        L._host_bound = True
    return L


def _take(L, p: ctypes.c_void_p) -> str:
    s = ctypes.string_at(p.value).decode()
    L.dissem_free(p)
    return s


def plan(config_json: str, mode: int, seed: int = 0, policy: int = POLICY_RANDOM, time_scale: int = 1,
         flags: int = 0) -> dict:
    """The plan the leader of `mode` issues (cmd/main.go:153-165 -> node.go:326-352 / :554-626 / :1200-1288)."""
    L = _bind()
    out = ctypes.c_void_p()
    rc = L.dissem_plan_json(config_json.encode(), mode, seed & ((1 << 64) - 1), policy, time_scale, flags,
                            ctypes.byref(out))
    if rc != 0:
        raise HostError(rc, (L.dissem_last_error() or b"").decode())
    return json.loads(_take(L, out))


def parse_plan(p: dict) -> List[Xfer]:
    return [Xfer(s, d, l, 0, off, size, total) for (s, d, l, off, size, total, _rate) in p["xfers"]]


def run_distribution(config_json: str, mode: int, node_device: Optional[Sequence[int]] = None, **opts) -> dict:
    """Leader + receivers as threads over the device transport, one process (needs a GPU)."""
    L = _bind()
    out = ctypes.c_void_p()
    arr = None
    if node_device is not None:
        arr = (ctypes.c_int * len(node_device))(*node_device)
    rc = L.dissem_run_json(config_json.encode(), mode, arr, json.dumps(opts).encode(), ctypes.byref(out))
    if rc != 0:
        raise HostError(rc, (L.dissem_last_error() or b"").decode())
    return json.loads(_take(L, out))


def transport_script(peers: Sequence[int], ops: Sequence[dict]) -> dict:
    """Drive the device transport's control plane (Send / Broadcast / Close / RegisterPipe) and return what every peer's
    Deliver() yields, in order.  Needs no GPU."""
    L = _bind()
    out = ctypes.c_void_p()
    rc = L.dissem_transport_script_json(json.dumps({"peers": list(peers), "ops": list(ops)}).encode(), ctypes.byref(out))
    if rc != 0:
        raise HostError(rc, (L.dissem_last_error() or b"").decode())
    return json.loads(_take(L, out))


def pace_schedule(layer_off: int, size: int, rate: int):
    """[(piece_off, piece_size, release_ns)]: when a send limited to `rate` B/s may launch each piece (pure host logic)."""
    L = _bind()
    n = ctypes.c_int()
    rc = L.dissem_pace_schedule(layer_off, size, rate, None, None, None, 0, ctypes.byref(n))
    if rc != 0:
        raise HostError(rc, (L.dissem_last_error() or b"").decode())
    a, b, t = ((ctypes.c_int64 * max(n.value, 1))() for _ in range(3))
    rc = L.dissem_pace_schedule(layer_off, size, rate, a, b, t, n.value, ctypes.byref(n))
    if rc != 0:
        raise HostError(rc, (L.dissem_last_error() or b"").decode())
    return [(a[i], b[i], t[i]) for i in range(n.value)]


def msg_roundtrip(**fields):
    """Encode one control-plane message to the reference's wire envelope and decode it back."""
    L = _bind()
    wire, dec = ctypes.c_void_p(), ctypes.c_void_p()
    rc = L.dissem_msg_roundtrip_json(json.dumps(fields).encode(), ctypes.byref(wire), ctypes.byref(dec))
    if rc != 0:
        raise HostError(rc, (L.dissem_last_error() or b"").decode())
    return _take(L, wire), json.loads(_take(L, dec))
