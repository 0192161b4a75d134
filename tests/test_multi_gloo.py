"""The N>1 host logic on CPU: two gloo ranks shard a plan, exchange (fake) buffer handles and agree on who pushes
what.  No GPU, no compute entry point."""
import importlib
import os
import socket

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import dissem_b200 as d
from oracle import dissem_oracle as o

multi = importlib.import_module("distributed-llm-dissemination_b200.multi")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, text, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        xfers = d.parse_plan(d.plan(text, 1, seed=5))
        owner = multi.owner_round_robin(world)
        mine = multi.my_transfers(xfers, owner, rank)
        exp = multi.exports_needed(xfers, owner, rank)
        imp = multi.imports_needed(xfers, owner, rank)
        handles = {(n, l): bytes([rank]) * 32 + n.to_bytes(16, "little") + l.to_bytes(16, "little") for (n, l, _t) in exp}
        merged = multi.exchange_handles(dist, world, rank, handles)
        ok = all((n, l) in merged and merged[(n, l)][0] == owner(n) and
                 int.from_bytes(merged[(n, l)][32:48], "little") == n for (n, l, _t) in imp)
        dev = multi.device_of_node(8, owner, rank, 0)
        q.put((rank, [(x.src_node, x.dst_node, x.layer) for x in mine], len(exp), len(imp), ok, dev))
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_the_plan_and_exchange_handles():
    world = 2
    text = o.config_stage_ring(8, 32, 1 << 20)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, text, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = d.parse_plan(d.plan(text, 1, seed=5))
    union = sorted(t for (_r, mine, *_rest) in got for t in mine)
    assert union == sorted((x.src_node, x.dst_node, x.layer) for x in full)          # every transfer launched exactly once
    assert all(ok for (_r, _m, _e, _i, ok, _d) in got)
    assert sum(e for (_r, _m, e, _i, _ok, _d) in got) == sum(i for (_r, _m, _e, i, _ok, _d) in got)
    assert got[0][5] == [0, -1, 0, -1, 0, -1, 0, -1] and got[1][5] == [-1, 0, -1, 0, -1, 0, -1, 0]


def test_link_bytes_roofline_inputs():
    text = o.config_stage_ring(8, 32, o.S2)
    xf = d.parse_plan(d.plan(text, 1, policy=1))                                     # balanced owners
    eg, ing, loc = multi.link_bytes(xf, multi.owner_round_robin(8), 8)
    assert sum(eg) == sum(ing) == 28 * o.S2 and sum(loc) == 0
    assert max(ing) == 4 * o.S2                                                      # each receiver takes its stage
    eg1, ing1, loc1 = multi.link_bytes(xf, multi.owner_round_robin(1), 1)
    assert eg1 == [0] and loc1 == [28 * o.S2]


def test_pull_flags_swap_exports_and_imports():
    text = o.config_stage_ring(8, 32, o.S2)
    xf = d.plan_directions(d.parse_plan(d.plan(text, 1, seed=0x5EED)), list(range(8)))
    owner = multi.owner_round_robin(8)
    pulled = [x for x in xf if x.flags & d.XFER_PULL]
    assert pulled
    for rank in range(8):
        mine = multi.my_transfers(xf, owner, rank)
        for x in mine:
            assert owner(x.dst_node if x.flags & 1 else x.src_node) == rank
        for (n, l, _t) in multi.imports_needed(xf, owner, rank):
            assert owner(n) != rank
        for (n, l, _t) in multi.exports_needed(xf, owner, rank):
            assert owner(n) == rank
    # every import is somebody's export
    imp = {t for r in range(8) for t in multi.imports_needed(xf, owner, r)}
    exp = {t for r in range(8) for t in multi.exports_needed(xf, owner, r)}
    assert imp == exp
    # the leader exports the SOURCE layers that are pulled from it, and imports nothing for them
    assert {(n, l) for (n, l, _t) in multi.exports_needed(xf, owner, 0)} == {(0, x.layer) for x in pulled}
    assert sum(len(multi.my_transfers(xf, owner, r)) for r in range(8)) == len(xf)


def _fd_sender(tag, path):
    fx = multi.FdExchange(1, tag)
    fd = os.open(path, os.O_RDONLY)
    fx.send(0, "buf:1:7", fd)
    fx.close()


def test_fd_exchange_passes_descriptors_between_processes(tmp_path):
    """The plumbing that carries shareable VMM / multicast handles between ranks (SCM_RIGHTS on unix sockets)."""
    p = tmp_path / "payload"
    p.write_bytes(b"layer bytes travel as descriptors")
    tag = f"test{os.getpid()}"
    fx = multi.FdExchange(0, tag)
    try:
        ctx = mp.get_context("spawn")
        child = ctx.Process(target=_fd_sender, args=(tag, str(p)))
        child.start()
        fd = fx.recv("buf:1:7", timeout=60)
        child.join(timeout=60)
        assert child.exitcode == 0
        assert os.read(fd, 100) == b"layer bytes travel as descriptors"
        os.close(fd)
        with pytest.raises(TimeoutError):
            fx.recv("never", timeout=0.2)
    finally:
        fx.close()


def test_flag_relay_ok_recognises_device_ordered_relay_plans():
    """multi.flag_relay_ok: dissem_plan_relay's whole-layer chains carry DISSEM_XFER_WAIT_SRC on every hop that reads a
    copy delivered in an earlier round, so their rounds can order themselves on the device across processes; a plan
    that forwards without the flag (or forwards a fragment chain, which gets none) needs the barrier per round."""
    size = 8 * d.CHUNK_BYTES
    base = [d.Xfer(0, dst, 5, 0, 0, size, size) for dst in (1, 2, 3)] + [d.Xfer(3, 1, 6, 0, 0, size, size)]
    relay = d.plan_relay(base, 4)
    hops = [x for x in relay if x.layer == 5 and x.src_node != 0]
    assert hops and all(x.flags & multi.XFER_WAIT_SRC for x in hops)
    assert not any(x.flags & multi.XFER_WAIT_SRC for x in relay if x.src_node == 0 or x.layer == 6)
    assert multi.flag_relay_ok(relay)
    assert multi.flag_relay_ok(d.plan_directions(relay, [0, 1, 2, 3]))          # direction choice keeps the flag
    assert all((a.flags & 4) == (b.flags & 4) for a, b in zip(relay, d.plan_directions(relay, [0, 1, 2, 3])))
    stripped = [d.Xfer(x.src_node, x.dst_node, x.layer, x.flags & ~multi.XFER_WAIT_SRC, x.off, x.size, x.total) for x in relay]
    assert not multi.flag_relay_ok(stripped)
    frag = d.plan_relay([d.Xfer(0, dst, 5, 0, d.CHUNK_BYTES, size - d.CHUNK_BYTES, size) for dst in (1, 2)], 2)
    assert not any(x.flags & multi.XFER_WAIT_SRC for x in frag) and not multi.flag_relay_ok(frag)
    assert multi.flag_relay_ok(base)                                            # no rounds at all: nothing to order


def test_store_sharing_and_multicast_group_helpers():
    """multi.remote_touches / mcast_groups / fanout_plan: which buffers another rank's kernel touches (push: the destination's,
    pull: the source's), which transfers form NVLS groups (same bytes, one source, >= 2 destinations on GPUs of their own),
    and the BASELINE config-4 placement (239 deliveries on 8 nodes, seed never in its own destination set)."""
    owner = lambda n: n
    xs = [d.Xfer(0, 1, 5, 0, 0, 100, 100), d.Xfer(0, 2, 5, d.XFER_PULL, 0, 100, 100), d.Xfer(3, 3, 6, 0, 0, 10, 10)]
    assert multi.remote_touches(xs, owner) == [(0, 5, 100, 2), (1, 5, 100, 0)]
    assert multi.remote_touches(xs, lambda n: 0) == []                       # one GPU: nothing crosses
    assert multi.mcast_groups(xs, owner) == [(5, 0, [1, 2])]
    assert multi.mcast_groups(xs, lambda n: n % 2) == []                     # source and a destination share a GPU
    holders, plan = multi.fanout_plan(d.Xfer, 8, 80, 1000, False)
    assert len(plan) == 239 and all(holders[x.layer] == x.src_node != x.dst_node for x in plan)
    assert sorted({len([x for x in plan if x.layer == l]) for l in range(80)}) == [2, 3, 4]
    seed_holders, seed_plan = multi.fanout_plan(d.Xfer, 8, 80, 1000, True)
    assert set(seed_holders.values()) == {0} and len(seed_plan) == 209
    groups = multi.mcast_groups(plan, owner)
    assert len(groups) == 80 and all(len(g[2]) >= 2 for g in groups)
