// Scheduling decisions on a plan, pure host logic behind the C ABI (include/dissem.h): which side executes each transfer
// (push or pull), relay chains over rounds, and the packaged choice between them.  None has a counterpart in the reference,
// whose TCP transport has one direction and no notion of a shared fabric; the port rates the rules encode are measured ones
// (profiles/r01_summary.md).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <tuple>
#include <vector>

#include "../../include/dissem.h"
#include "../csrc/dissem_engine.h"

using dissem::set_error;

extern "C" {

int dissem_plan_directions(dissem_xfer *xfers, int n, const int *gpu_of_node, int n_nodes) {
    if (n < 0 || (n > 0 && !xfers) || !gpu_of_node || n_nodes <= 0) return set_error(DISSEM_EINVAL, "dissem_plan_directions: bad arguments");
    std::map<int, int64_t> egress, ingress;
    for (int i = 0; i < n; ++i) {
        const dissem_xfer &x = xfers[i];
        if (x.src_node < 0 || x.src_node >= n_nodes || x.dst_node < 0 || x.dst_node >= n_nodes)
            return set_error(DISSEM_ENOENT, "transfer %d names a node outside [0,%d)", i, n_nodes);
        const int gs = gpu_of_node[x.src_node], gd = gpu_of_node[x.dst_node];
        if (gs == gd) continue;
        egress[gs] += x.size;
        ingress[gd] += x.size;
    }
    for (int i = 0; i < n; ++i) {
        dissem_xfer &x = xfers[i];
        const int gs = gpu_of_node[x.src_node], gd = gpu_of_node[x.dst_node];
        x.flags &= ~DISSEM_XFER_PULL;
        if (gs == gd || (x.flags & DISSEM_XFER_FILE)) continue;      // a disk copy is streamed by its own node's GPU
        const int64_t forward = std::max(egress[gs], ingress[gd]);
        const int64_t reverse = std::max(ingress[gs], egress[gd]);
        if (2 * reverse < forward) x.flags |= DISSEM_XFER_PULL;
    }
    return DISSEM_OK;
}

int dissem_plan_align(dissem_xfer *xfers, int n, int *n_out) {
    if (n < 0 || (n > 0 && !xfers) || !n_out) return set_error(DISSEM_EINVAL, "dissem_plan_align: bad arguments");
    // fragments of one (destination, layer, round), ascending by offset
    std::map<std::tuple<int, int, int>, std::vector<int>> groups;
    for (int i = 0; i < n; ++i)
        if (xfers[i].size > 0)
            groups[std::make_tuple(xfers[i].dst_node, xfers[i].layer, (xfers[i].flags >> DISSEM_XFER_ROUND_SHIFT) & 0xFFFF)].push_back(i);
    for (auto &kv : groups) {
        std::vector<int> &g = kv.second;
        if (g.size() < 2) continue;
        std::sort(g.begin(), g.end(), [&](int a, int b) { return xfers[a].off < xfers[b].off; });
        for (size_t k = 0; k + 1 < g.size(); ++k) {
            dissem_xfer &a = xfers[g[k]], &b = xfers[g[k + 1]];
            if (a.off + a.size != b.off || a.total != b.total) continue;          // not a cut of one contiguous range: leave it
            const int64_t cut = b.off / DISSEM_CHUNK_BYTES * DISSEM_CHUNK_BYTES;  // the boundary moves DOWN to a chunk start
            if (cut == b.off) continue;
            const int64_t new_cut = std::max(cut, a.off);                         // a fragment shorter than the move vanishes
            b.size += b.off - new_cut;
            b.off = new_cut;
            a.size = new_cut - a.off;
        }
    }
    int m = 0;
    for (int i = 0; i < n; ++i)
        if (xfers[i].size > 0 || xfers[i].total == 0) xfers[m++] = xfers[i];
    *n_out = m;
    return DISSEM_OK;
}

int dissem_plan_localize(dissem_xfer *xfers, int n, const int *gpu_of_node, int n_nodes, const int *holder_nodes,
                         const int *holder_layers, int n_holders, int *n_out) {
    if (n < 0 || (n > 0 && !xfers) || !gpu_of_node || n_nodes <= 0 || n_holders < 0 || (n_holders > 0 && (!holder_nodes || !holder_layers)) || !n_out)
        return set_error(DISSEM_EINVAL, "dissem_plan_localize: bad arguments");
    std::map<std::pair<int, int>, int> local_holder;          // (gpu, layer) -> lowest node id on that GPU holding the layer
    for (int i = 0; i < n_holders; ++i) {
        if (holder_nodes[i] < 0 || holder_nodes[i] >= n_nodes) return set_error(DISSEM_ENOENT, "holder %d names a node outside [0,%d)", i, n_nodes);
        auto key = std::make_pair(gpu_of_node[holder_nodes[i]], holder_layers[i]);
        auto it = local_holder.find(key);
        if (it == local_holder.end() || holder_nodes[i] < it->second) local_holder[key] = holder_nodes[i];
    }
    for (int i = 0; i < n; ++i) {
        dissem_xfer &x = xfers[i];
        if (x.src_node < 0 || x.src_node >= n_nodes || x.dst_node < 0 || x.dst_node >= n_nodes)
            return set_error(DISSEM_ENOENT, "transfer %d names a node outside [0,%d)", i, n_nodes);
        if ((x.flags & DISSEM_XFER_FILE) || gpu_of_node[x.src_node] == gpu_of_node[x.dst_node]) continue;
        auto it = local_holder.find(std::make_pair(gpu_of_node[x.dst_node], x.layer));
        if (it == local_holder.end() || it->second == x.dst_node) continue;
        x.src_node = it->second;                               // same bytes, read from the copy that shares the GPU
        x.flags &= ~DISSEM_XFER_PULL;
    }
    // fragments that now have the same source and touch each other become one transfer
    std::vector<int> order((size_t)n);
    for (int i = 0; i < n; ++i) order[(size_t)i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
        const dissem_xfer &x = xfers[a], &y = xfers[b];
        return std::make_tuple(x.dst_node, x.layer, x.flags, x.src_node, x.off) < std::make_tuple(y.dst_node, y.layer, y.flags, y.src_node, y.off);
    });
    std::vector<dissem_xfer> merged;
    for (int idx : order) {
        const dissem_xfer &x = xfers[idx];
        if (!merged.empty()) {
            dissem_xfer &m = merged.back();
            if (m.dst_node == x.dst_node && m.layer == x.layer && m.flags == x.flags && m.src_node == x.src_node && m.total == x.total &&
                m.off + m.size == x.off && x.size > 0 && m.size > 0) {
                m.size += x.size;
                continue;
            }
        }
        merged.push_back(x);
    }
    for (size_t i = 0; i < merged.size(); ++i) xfers[i] = merged[i];
    *n_out = (int)merged.size();
    return DISSEM_OK;
}

int dissem_plan_relay(const dissem_xfer *in, int n, int pieces, dissem_xfer **out, int *n_out) {
    if (n < 0 || (n > 0 && !in) || !out || !n_out || pieces < 1 || pieces > 4096)
        return set_error(DISSEM_EINVAL, "dissem_plan_relay: bad arguments");
    std::vector<dissem_xfer> res;
    std::vector<char> used((size_t)std::max(n, 1), 0);
    auto emit = [&](int src, int dst, const dissem_xfer &x, int hop) {
        // hops >= 1 read a copy that is still arriving: DISSEM_XFER_WAIT_SRC lets the executing kernel order itself behind the
        // upstream hop on the device (needed across processes; harmless where round events already order the hops)
        // (the counter threshold is off + size, which presumes a copy filled from offset 0: whole-layer chains only)
        const int wait = (hop > 0 && x.off == 0) ? DISSEM_XFER_WAIT_SRC : 0;
        // pieces are cut at multiples of the checksum chunk in LAYER space, so every piece owns whole chunks
        int64_t piece = (x.size + pieces - 1) / pieces;
        piece = (piece + DISSEM_CHUNK_BYTES - 1) / DISSEM_CHUNK_BYTES * DISSEM_CHUNK_BYTES;
        if (piece <= 0) piece = DISSEM_CHUNK_BYTES;
        int64_t pos = x.off;
        const int64_t end = x.off + x.size;
        for (int p = 0; pos < end; ++p) {
            int64_t stop = (pos / piece + 1) * piece;            // next piece boundary in layer space
            if (stop > end) stop = end;
            dissem_xfer y = x;
            y.src_node = src;
            y.dst_node = dst;
            y.off = pos;
            y.size = stop - pos;
            y.flags = (x.flags & ~(0xFFFF << DISSEM_XFER_ROUND_SHIFT) & ~DISSEM_XFER_PULL & ~DISSEM_XFER_WAIT_SRC) | DISSEM_XFER_ROUND(p + hop) | wait;
            res.push_back(y);
            pos = stop;
        }
    };
    for (int i = 0; i < n; ++i) {
        if (used[i]) continue;
        const dissem_xfer &x = in[i];
        used[i] = 1;
        if (x.size <= 0 || (x.flags & DISSEM_XFER_FILE)) { res.push_back(x); continue; }    // disk sources are not relayed
        std::vector<int> dsts{x.dst_node};
        for (int j = i + 1; j < n; ++j) {
            const dissem_xfer &y = in[j];
            if (!used[j] && !(y.flags & DISSEM_XFER_FILE) && y.src_node == x.src_node && y.layer == x.layer && y.off == x.off && y.size == x.size &&
                y.total == x.total && y.dst_node != x.dst_node) {
                dsts.push_back(y.dst_node);
                used[j] = 1;
            }
        }
        std::sort(dsts.begin(), dsts.end());
        // rotate so that the first hop (which forwards the most) is not always the same node
        std::rotate(dsts.begin(), dsts.begin() + (size_t)((unsigned)x.layer % dsts.size()), dsts.end());
        int prev = x.src_node;
        for (size_t h = 0; h < dsts.size(); ++h) {
            emit(prev, dsts[h], x, (int)h);                      // hop h forwards in round p+h what arrived in round p+h-1
            prev = dsts[h];
        }
    }
    dissem_xfer *buf = (dissem_xfer *)malloc(std::max<size_t>(res.size(), 1) * sizeof(dissem_xfer));
    if (!buf) return set_error(DISSEM_ENOMEM, "out of memory");
    memcpy(buf, res.data(), res.size() * sizeof(dissem_xfer));
    *out = buf;
    *n_out = (int)res.size();
    return DISSEM_OK;
}

int dissem_plan_optimize(const dissem_xfer *in, int n, const int *gpu_of_node, int n_nodes, dissem_xfer **out, int *n_out,
                         int *relay_pieces_out) {
    if (n < 0 || (n > 0 && !in) || !gpu_of_node || n_nodes <= 0 || !out || !n_out)
        return set_error(DISSEM_EINVAL, "dissem_plan_optimize: bad arguments");
    for (int i = 0; i < n; ++i)
        if (in[i].src_node < 0 || in[i].src_node >= n_nodes || in[i].dst_node < 0 || in[i].dst_node >= n_nodes)
            return set_error(DISSEM_ENOENT, "transfer %d names a node outside [0,%d)", i, n_nodes);
    // Disk sources (DISSEM_XFER_FILE): a layer that several nodes need is read from its file ONCE, into the first of
    // them; the others take it from that node's HBM, one round later (NVLink is ~16x the disk path).  Those forwards
    // are ordinary fan-out transfers, so the relay decision below applies to them as well.
    std::vector<dissem_xfer> staged(in, in + n);
    bool any_file = false;
    {
        std::vector<char> seen((size_t)std::max(n, 1), 0);
        for (int i = 0; i < n; ++i) {
            if (seen[i] || !(staged[i].flags & DISSEM_XFER_FILE)) continue;
            any_file = true;
            const dissem_xfer first = staged[i];
            if (first.size <= 0 || first.dst_node == first.src_node) continue;      // a node loading its own copy feeds nobody else
            for (int j = i + 1; j < n; ++j) {
                dissem_xfer &y = staged[j];
                if (!seen[j] && (y.flags & DISSEM_XFER_FILE) && y.src_node == first.src_node && y.layer == first.layer &&
                    y.off == first.off && y.size == first.size && y.total == first.total && y.dst_node != first.dst_node &&
                    y.dst_node != first.src_node) {
                    seen[j] = 1;
                    y.src_node = first.dst_node;
                    y.flags &= ~DISSEM_XFER_FILE;
                }
            }
        }
    }
    {
        // fragment cuts (a flow plan) move to checksum-chunk boundaries: no table entry is then shared by two senders
        int m = n;
        int rc = dissem_plan_align(staged.data(), n, &m);
        if (rc) return rc;
        staged.resize((size_t)m);
        n = m;
    }
    in = staged.data();
    // busiest-port bytes of a transfer list (transfers inside one GPU do not use a port)
    auto busiest = [&](const dissem_xfer *xs, int cnt) {
        std::map<int, int64_t> eg, ing;
        for (int i = 0; i < cnt; ++i) {
            const int gs = gpu_of_node[xs[i].src_node], gd = gpu_of_node[xs[i].dst_node];
            if (gs == gd) continue;
            eg[gs] += xs[i].size;
            ing[gd] += xs[i].size;
        }
        int64_t m = 0;
        for (auto &kv : eg) m = std::max(m, kv.second);
        for (auto &kv : ing) m = std::max(m, kv.second);
        return m;
    };
    // are there transfers of the same bytes from one source to several GPUs at all?
    bool fanout = false;
    int64_t biggest = 0;
    for (int i = 0; i < n && !fanout; ++i)
        for (int j = i + 1; j < n; ++j)
            if (in[j].src_node == in[i].src_node && in[j].layer == in[i].layer && in[j].off == in[i].off &&
                in[j].size == in[i].size && gpu_of_node[in[j].dst_node] != gpu_of_node[in[i].dst_node]) { fanout = true; break; }
    for (int i = 0; i < n; ++i) biggest = std::max(biggest, in[i].size);
    dissem_xfer *chosen = nullptr;
    int n_chosen = 0, pieces_used = 0;
    if (fanout && biggest >= (8ll << 20)) {
        // relay chains: the source sends once, destinations forward; costs (pieces + hops - 1) / pieces in pipeline fill.
        // Pieces of ~64 MiB keep a round long enough to hide its launch and barrier costs (profiles/r01_summary.md).
        int pieces = (int)std::min<int64_t>(32, std::max<int64_t>(4, biggest / (64ll << 20)));
        dissem_xfer *relay = nullptr;
        int n_relay = 0;
        int rc = dissem_plan_relay(in, n, pieces, &relay, &n_relay);
        if (rc) return rc;
        int max_round = 0;
        for (int i = 0; i < n_relay; ++i) max_round = std::max(max_round, (relay[i].flags >> DISSEM_XFER_ROUND_SHIFT) & 0xFFFF);
        const double t_uni = (double)busiest(in, n);
        const double t_relay = (double)busiest(relay, n_relay) * (double)(max_round + 1) / (double)pieces;
        if (t_relay < 0.9 * t_uni) {
            chosen = relay;
            n_chosen = n_relay;
            pieces_used = pieces;
        } else {
            free(relay);
        }
    }
    if (!chosen) {
        chosen = (dissem_xfer *)malloc(std::max<size_t>((size_t)n, 1) * sizeof(dissem_xfer));
        if (!chosen) return set_error(DISSEM_ENOMEM, "out of memory");
        memcpy(chosen, in, (size_t)n * sizeof(dissem_xfer));
        n_chosen = n;
    }
    int rc = dissem_plan_directions(chosen, n_chosen, gpu_of_node, n_nodes);
    if (rc) { free(chosen); return rc; }
    if (any_file)                 // everything that is not read from disk runs after the disk transfers (round 0)
        for (int i = 0; i < n_chosen; ++i)
            if (!(chosen[i].flags & DISSEM_XFER_FILE)) {
                const int r = (chosen[i].flags >> DISSEM_XFER_ROUND_SHIFT) & 0xFFFF;
                chosen[i].flags = (chosen[i].flags & ~(0xFFFF << DISSEM_XFER_ROUND_SHIFT)) | DISSEM_XFER_ROUND(r + 1);
            }
    *out = chosen;
    *n_out = n_chosen;
    if (relay_pieces_out) *relay_pieces_out = pieces_used;
    return DISSEM_OK;
}

}  // extern "C"
