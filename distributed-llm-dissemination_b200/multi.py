"""One process per GPU: which rank owns which reference node, which transfers a rank launches, and which
destination buffers it must map from its peers.  Pure host logic plus a torch.distributed all-gather of 64-byte
handles (plumbing); the bytes themselves only ever move through the push kernels.

Sharding follows SURVEY §8(e): the units are independent (dest, layer) transfers, each executed by the rank that
owns the SOURCE node (push model); there is no data-path collective."""
from __future__ import annotations

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
