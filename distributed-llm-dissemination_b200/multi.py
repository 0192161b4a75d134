"""One process per GPU: which rank owns which reference node, which transfers a rank launches, and which
destination buffers it must map from its peers.  Pure host logic plus a torch.distributed all-gather of 64-byte
handles (plumbing); the bytes themselves only ever move through the push kernels.

Sharding follows SURVEY §8(e): the units are independent (dest, layer) transfers, each executed by the rank that
owns the SOURCE node (push model); there is no data-path collective."""
from __future__ import annotations

import os

from typing import Callable, Dict, Iterable, List, Sequence, Tuple


def owner_round_robin(world: int) -> Callable[[int], int]:
    """node i -> rank i % world (at world == number of nodes every node is its own GPU, as in the north star)."""
    return lambda node: node % world


def device_of_node(n_nodes: int, owner: Callable[[int], int], rank: int, local_device: int) -> List[int]:
    """Argument for dissem_init: the local CUDA ordinal for nodes this rank owns, -1 for the others."""
    return [local_device if owner(n) == rank else -1 for n in range(n_nodes)]


def _sides(x, pull):
    """(executing node, other node) of a transfer: the source pushes unless the transfer is flagged PULL
    (DISSEM_XFER_PULL, set by dissem_plan_directions) or `pull` forces it."""
    flags = getattr(x, "flags", 0)
    is_pull = (bool(pull) or bool(flags & 1)) and not (flags & 2)        # a disk source (DISSEM_XFER_FILE) always pushes
    return (x.dst_node, x.src_node) if is_pull else (x.src_node, x.dst_node)


def my_transfers(xfers: Sequence, owner: Callable[[int], int], rank: int, pull: bool = False) -> List:
    """Transfers this rank launches: those whose executing side it owns."""
    return [x for x in xfers if owner(_sides(x, pull)[0]) == rank]


def exports_needed(xfers: Sequence, owner: Callable[[int], int], rank: int, pull: bool = False) -> List[Tuple[int, int, int]]:
    """(node, layer, total) of buffers this rank owns that ANOTHER rank's kernel touches: destination buffers it
    is pushed into, source layers that are pulled from it."""
    out = set()
    for x in xfers:
        ex, other = _sides(x, pull)
        if owner(other) == rank and owner(ex) != rank:
            out.add((other, x.layer, x.total))
    return sorted(out)


def imports_needed(xfers: Sequence, owner: Callable[[int], int], rank: int, pull: bool = False) -> List[Tuple[int, int, int]]:
    """(node, layer, total) of remote buffers this rank's kernels touch."""
    out = set()
    for x in xfers:
        ex, other = _sides(x, pull)
        if owner(ex) == rank and owner(other) != rank:
            out.add((other, x.layer, x.total))
    return sorted(out)


def exchange_handles(dist, world: int, rank: int, mine: Dict[Tuple[int, int], bytes]) -> Dict[Tuple[int, int], bytes]:
    """All-gather {(node, layer): 64-byte handle}.  In the reference's terms the handle rides in the announceMsg
    (message.go:31-41) that every node sends before distribution starts."""
    everyone = [None] * world
    dist.all_gather_object(everyone, {f"{n}:{l}": h for (n, l), h in mine.items()})
    merged: Dict[Tuple[int, int], bytes] = {}
    for r, items in enumerate(everyone):
        for key, h in items.items():
            n, l = key.split(":")
            merged[(int(n), int(l))] = h
    return merged


def link_bytes(xfers: Iterable, owner: Callable[[int], int], world: int):
    """Per-rank egress / ingress bytes that cross GPUs, and bytes that stay local: the inputs of the NVLink roofline
    T* = max over GPUs of max(egress, ingress) / link bandwidth."""
    egress, ingress, local = [0] * world, [0] * world, [0] * world
    for x in xfers:
        s, d = owner(x.src_node), owner(x.dst_node)
        if s == d:
            local[s] += x.size
        else:
            egress[s] += x.size
            ingress[d] += x.size
    return egress, ingress, local


def split_rounds(xfers: Sequence) -> List[List]:
    """Transfers grouped by their round (DISSEM_XFER_ROUND bits), ascending."""
    by: Dict[int, List] = {}
    for x in xfers:
        by.setdefault((getattr(x, "flags", 0) >> 8) & 0xFFFF, []).append(x)
    return [by[r] for r in sorted(by)]


class RoundedSchedule:
    """A plan with rounds (e.g. from dissem_plan_relay) in the one-process-per-GPU model.  Inside one process the engine
    orders rounds on the device; ACROSS processes nothing does, so each round is compiled as its own schedule and the
    ranks meet at `barrier()` (a torch.distributed barrier + device synchronise) between rounds."""

    def __init__(self, eng, xfers: Sequence):
        self.eng = eng
        self.scheds = [eng.compile(r) for r in split_rounds(xfers)]

    def run(self, barrier: Callable[[], None]) -> float:
        """Launch every round in turn; returns this rank's summed device milliseconds."""
        ms = 0.0
        for s in self.scheds:
            ms += self.eng.wait(self.eng.launch(s))
            barrier()
        return ms

    def free(self):
        for s in self.scheds:
            self.eng.sched_free(s)
        self.scheds = []


XFER_WAIT_SRC = 4


def flag_relay_ok(xfers: Sequence) -> bool:
    """Can the rounds of this plan order themselves on the device across processes?  Yes when every transfer that reads
    a copy delivered in an EARLIER round of the plan carries DISSEM_XFER_WAIT_SRC (what dissem_plan_relay emits for
    whole-layer chains)."""
    first_arrival: Dict[Tuple[int, int], int] = {}
    for x in xfers:
        r = (getattr(x, "flags", 0) >> 8) & 0xFFFF
        key = (x.dst_node, x.layer)
        first_arrival[key] = min(first_arrival.get(key, r), r)
    for x in xfers:
        r = (getattr(x, "flags", 0) >> 8) & 0xFFFF
        key = (x.src_node, x.layer)
        if key in first_arrival and first_arrival[key] < r and not (x.flags & XFER_WAIT_SRC):
            return False
        if key in first_arrival and first_arrival[key] >= r and r > 0 and key != (x.dst_node, x.layer):
            return False                    # reads a copy that arrives in the same or a later round: not a relay chain
    return True


class FlagRelaySchedule:
    """A plan with rounds in the one-process-per-GPU model WITHOUT a host barrier between rounds: every rank compiles all
    of its rounds into one schedule and launches them back to back; a hop that forwards a piece waits on the device for
    the source copy's received counter (DISSEM_XFER_WAIT_SRC).  The counters must start at zero: run() resets this rank's
    destination copies and meets the other ranks ONCE, before the dissemination starts."""

    def __init__(self, eng, xfers: Sequence, owner: Callable[[int], int], rank: int):
        self.eng = eng
        eng.set_option("wait_src", 1)
        self.sched = eng.compile(xfers)
        self.my_dsts = sorted({(x.dst_node, x.layer) for x in xfers if owner(x.dst_node) == rank})
        self.scheds = [self.sched]

    def run(self, barrier: Callable[[], None]):
        for (n, l) in self.my_dsts:
            self.eng.reset(n, l)
        barrier()
        return self.eng.wait_detail(self.eng.launch(self.sched))

    def free(self):
        self.eng.sched_free(self.sched)
        self.scheds = []


class FdExchange:
    """Passes file descriptors (shareable VMM / multicast handles) between the ranks of one box over unix sockets
    (SCM_RIGHTS).  Plumbing only.  Every descriptor travels with a text key; recv(key) blocks until it has arrived."""

    def __init__(self, rank: int, tag: str):
        import os
        import socket
        import threading
        self._socket, self._os = socket, os
        self.rank, self.tag = rank, tag
        self.path = self._path(rank)
        try:
            os.unlink(self.path)
        except FileNotFoundError:
            pass
        self._srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self._srv.bind(self.path)
        self._srv.listen(256)
        self._got: Dict[str, int] = {}
        self._cv = threading.Condition()
        self._stop = False
        self._thr = threading.Thread(target=self._serve, daemon=True)
        self._thr.start()

    def _path(self, rank: int) -> str:
        return f"/tmp/dissem-{self.tag}-{rank}.sock"

    def _serve(self):
        while not self._stop:
            try:
                conn, _ = self._srv.accept()
            except OSError:
                return
            with conn:
                msg, fds, _flags, _addr = self._socket.recv_fds(conn, 4096, 8)
                if fds:
                    with self._cv:
                        self._got[msg.decode()] = fds[0]
                        self._cv.notify_all()

    def send(self, to_rank: int, key: str, fd: int, keep: bool = False):
        import time
        s = self._socket.socket(self._socket.AF_UNIX, self._socket.SOCK_STREAM)
        for _ in range(200):
            try:
                s.connect(self._path(to_rank))
                break
            except (FileNotFoundError, ConnectionRefusedError):
                time.sleep(0.05)
        else:
            raise RuntimeError(f"rank {to_rank} is not listening for descriptors")
        with s:
            self._socket.send_fds(s, [key.encode()], [fd])
        if not keep:
            self._os.close(fd)

    def recv(self, key: str, timeout: float = 60.0) -> int:
        with self._cv:
            if not self._cv.wait_for(lambda: key in self._got, timeout=timeout):
                raise TimeoutError(f"descriptor {key!r} never arrived at rank {self.rank}")
            return self._got.pop(key)

    def close(self):
        self._stop = True
        try:
            self._srv.close()
            self._os.unlink(self.path)
        except OSError:
            pass


# ----------------------------------------------------------------------------------------------------------------
# Placements and store set-up shared by bench.py and tools/ (one process per GPU)
# ----------------------------------------------------------------------------------------------------------------
def fanout_plan(xfer_cls, n_nodes: int, n_layers: int, size: int, one_seed: bool):
    """BASELINE config 4: layer l -> {l, l+1, (l+2 if l%3>=1), (l+3 if l%3==2)} mod N, seeded on node
    (l+4) mod N — or every layer on node 0 (`one_seed`).  Returns ({layer: seed node}, [transfers])."""
    holders, xfers = {}, []
    for l in range(n_layers):
        seed = 0 if one_seed else (l + 4) % n_nodes
        holders[l] = seed
        dsts = [l % n_nodes, (l + 1) % n_nodes]
        if l % 3 >= 1:
            dsts.append((l + 2) % n_nodes)
        if l % 3 == 2:
            dsts.append((l + 3) % n_nodes)
        for dst in sorted(set(dsts)):
            if dst != seed:
                xfers.append(xfer_cls(seed, dst, l, 0, 0, size, size))
    return holders, xfers


def remote_touches(xfers: Sequence, owner: Callable[[int], int], pull: bool = False) -> List[Tuple[int, int, int, int]]:
    """(node, layer, total, executing rank) for every buffer that a kernel of ANOTHER rank than its owner touches."""
    out = set()
    for x in xfers:
        ex, other = _sides(x, pull)
        if owner(ex) != owner(other):
            out.add((other, x.layer, x.total, owner(ex)))
    return sorted(out)


class SharedStore:
    """Makes the buffers of this rank's nodes reachable from the ranks whose kernels touch them, and maps theirs here.
    IPC handles (cudaMalloc'ed layers, all-gathered) by default; with `fx` (an FdExchange) the layers are shareable VMM
    allocations and travel as file descriptors — the form NVLS multicast needs.  Every call is collective."""

    def __init__(self, eng, dist, world: int, rank: int, owner: Callable[[int], int], fx=None):
        self.eng, self.dist, self.world, self.rank, self.owner, self.fx = eng, dist, world, rank, owner, fx
        self.imported = set()
        self.sent = set()

    def share(self, xfers: Sequence, pull: bool = False):
        if self.dist is None:
            return
        touches = remote_touches(xfers, self.owner, pull)
        if self.fx is None:
            mine = {}
            for (n, l, _t, _r) in touches:
                if self.owner(n) == self.rank and (n, l) not in mine:
                    mine[(n, l)] = self.eng.export(n, l)
            handles = exchange_handles(self.dist, self.world, self.rank, mine)
            for (n, l, total, r) in touches:
                if r == self.rank and (n, l) not in self.imported:
                    self.eng.import_(n, l, total, handles[(n, l)])
                    self.imported.add((n, l))
        else:
            for (n, l, _t, r) in touches:
                if self.owner(n) == self.rank and (n, l, r) not in self.sent:
                    self.fx.send(r, f"buf:{n}:{l}:{r}", self.eng.export_fd(n, l))
                    self.sent.add((n, l, r))
            for (n, l, total, r) in touches:
                if r == self.rank and (n, l) not in self.imported:
                    fd = self.fx.recv(f"buf:{n}:{l}:{r}")
                    try:
                        self.eng.import_fd(n, l, total, fd)
                    finally:
                        os.close(fd)                 # the mapping keeps the allocation alive; the descriptor is done
                    self.imported.add((n, l))
        self.dist.barrier()


def mcast_groups(xfers: Sequence, owner: Callable[[int], int]) -> List[Tuple[int, int, List[int]]]:
    """(layer, source node, [destination nodes]) for the transfers of the same whole bytes from one source to two or more
    nodes, each on a GPU of its own — the groups the schedule compiler sends as one NVLS multicast stream."""
    by: Dict[Tuple[int, int, int, int], List[int]] = {}
    for x in xfers:
        by.setdefault((x.layer, x.src_node, x.off, x.size), []).append(x.dst_node)
    out = []
    for (layer, src, _off, _size), dsts in sorted(by.items()):
        ranks = [owner(src)] + [owner(d) for d in dsts]
        if len(dsts) >= 2 and len(set(ranks)) == len(ranks):
            out.append((layer, src, sorted(dsts)))
    return out


def setup_mcast(eng, fx, dist, groups: Sequence[Tuple[int, int, List[int]]], owner: Callable[[int], int], rank: int):
    """Create (source's rank), join (destinations' ranks) and bind (everybody) one multicast object per group.  The
    source's process must already have imported the destinations' copies (SharedStore with `fx`).  Collective: after
    each phase the ranks agree on whether it worked, so a box without NVLS support makes every rank return
    (False, reason) instead of leaving some of them waiting for descriptors that never come."""
    import os

    import torch

    def agreed(ok: bool) -> bool:
        f = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(f, op=dist.ReduceOp.MIN)
        return bool(f.item())

    why = ""
    ok = True
    try:
        for (layer, src, dsts) in groups:
            if owner(src) == rank:
                fd = eng.mcast_create(layer, src, dsts)
                for dn in dsts:
                    fx.send(owner(dn), f"mc:{layer}:{src}:{dn}", fd, keep=True)
                os.close(fd)
    except Exception as e:                                  # noqa: BLE001
        ok, why = False, f"create: {e}"
    if not agreed(ok):
        return False, why or "multicast object creation failed on another rank"
    try:
        for (layer, src, dsts) in groups:
            for dn in dsts:
                if owner(dn) == rank:
                    fd = fx.recv(f"mc:{layer}:{src}:{dn}")
                    try:
                        eng.mcast_join(layer, src, dsts, dn, fd)
                    finally:
                        os.close(fd)
    except Exception as e:                                  # noqa: BLE001
        ok, why = False, f"join: {e}"
    if not agreed(ok):
        return False, why or "joining the multicast object failed on another rank"
    try:
        for (layer, src, dsts) in groups:
            if owner(src) == rank or any(owner(dn) == rank for dn in dsts):
                eng.mcast_bind(layer, src, dsts)
    except Exception as e:                                  # noqa: BLE001
        ok, why = False, f"bind: {e}"
    if not agreed(ok):
        return False, why or "binding to the multicast object failed on another rank"
    return True, ""
