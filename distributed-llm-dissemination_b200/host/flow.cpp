#include "flow.h"

#include <algorithm>
#include <limits>
#include <set>
#include <stdexcept>

namespace dissem {
namespace host {

static constexpr int64_t kInf = std::numeric_limits<int64_t>::max();   // math.MaxInt64, flow.go:252

std::vector<std::pair<LayerID, NodeID>> FlowGraph::layer_vertices(LayerID l) const {
    std::vector<std::pair<LayerID, NodeID>> v;
    for (auto it = layer_idx_.lower_bound({l, 0}); it != layer_idx_.end() && it->first.first == l; ++it) v.push_back(it->first);
    return v;
}

FlowGraph::FlowGraph(const Status &status, const std::map<LayerID, int64_t> &sizes,
                     const std::map<NodeID, int64_t> &network_bw, const Assignment &assignment, FlowOptions opt)
    : status_(status), sizes_(sizes), bw_(network_bw), assignment_(assignment), opt_(opt) {
    if (opt_.TimeScale <= 0) throw std::invalid_argument("TimeScale must be positive");
    std::set<std::pair<LayerID, NodeID>> need;
    for (auto &a : assignment_)
        for (auto &l : a.second) need.insert({l.first, opt_.PerDestLayers ? a.first : kAnyDest});
    needed_.assign(need.begin(), need.end());
    // vertex numbering = addIndex order of flow.go:79-122
    int n = 0;
    src_ = n++;
    for (auto &s : status_) sender_idx_[s.first] = n++;
    for (auto &s : status_) {
        std::set<int> kinds;
        for (auto &l : s.second) kinds.insert((int)l.second.Source);
        for (int k : kinds) client_idx_[{s.first, k}] = n++;
    }
    for (auto &v : needed_) layer_idx_[v] = n++;
    for (auto &a : assignment_) receiver_idx_[a.first] = n++;
    sink_ = n++;
    n_ = n;
    cap_.assign((size_t)n_ * n_, 0);
    // static topology -> neighbour lists (both directions, for the residual graph)
    std::vector<std::set<int>> nb(n_);
    auto link = [&](int a, int b) { nb[a].insert(b); nb[b].insert(a); };
    for (auto &s : status_) {
        link(src_, sender_idx_[s.first]);
        for (auto &l : s.second) {
            auto vs = layer_vertices(l.first);
            if (vs.empty()) continue;
            int c = client_idx_[{s.first, (int)l.second.Source}];
            link(sender_idx_[s.first], c);
            for (auto &v : vs) link(c, layer_idx_[v]);
        }
    }
    for (auto &a : assignment_) {
        int r = receiver_idx_[a.first];
        for (auto &l : a.second) link(layer_idx_[{l.first, opt_.PerDestLayers ? a.first : kAnyDest}], r);
        link(r, sink_);
    }
    nbr_.resize(n_);
    for (int v = 0; v < n_; ++v) nbr_[v].assign(nb[v].begin(), nb[v].end());
}

int64_t FlowGraph::Cap(int64_t rate, int64_t t) const {
    // TimeScale == 1 is `rate * time` as written at flow.go:236, :250, :267 (where Go would wrap on overflow)
    __int128 c = (__int128)rate * (__int128)t / (__int128)opt_.TimeScale;
    if (c > (__int128)kInf) throw std::overflow_error("flow capacity exceeds int64");
    return (int64_t)c;
}

void FlowGraph::BuildEdgeCapacity(int64_t t) {
    std::fill(cap_.begin(), cap_.end(), 0);
    auto C = [&](int a, int b) -> int64_t & { return cap_[(size_t)a * n_ + b]; };
    for (auto &s : status_) {
        auto bw = bw_.find(s.first);
        C(src_, sender_idx_[s.first]) = Cap(bw == bw_.end() ? 0 : bw->second, t);
        for (auto &l : s.second) {      // ascending layer id: "last write wins" made deterministic
            auto vs = layer_vertices(l.first);
            if (vs.empty()) continue;
            int c = client_idx_[{s.first, (int)l.second.Source}];
            C(sender_idx_[s.first], c) = (l.second.LimitRate == 0 && opt_.ZeroRateUnlimited) ? kInf : Cap(l.second.LimitRate, t);
            for (auto &v : vs) C(c, layer_idx_[v]) = kInf;
        }
    }
    for (auto &a : assignment_) {
        int r = receiver_idx_[a.first];
        for (auto &l : a.second) {
            auto sz = sizes_.find(l.first);
            C(layer_idx_[{l.first, opt_.PerDestLayers ? a.first : kAnyDest}], r) = sz == sizes_.end() ? 0 : sz->second;
        }
        auto bw = bw_.find(a.first);
        C(r, sink_) = Cap(bw == bw_.end() ? 0 : bw->second, t);
    }
}

bool FlowGraph::Bfs(int src, int dst, std::vector<int> *parent) {
    std::vector<char> seen(n_, 0);
    parent->assign(n_, 0);
    std::vector<int> queue;
    queue.reserve(n_);
    queue.push_back(src);
    seen[src] = 1;
    for (size_t head = 0; head < queue.size(); ++head) {
        int u = queue[head];
        for (int v : nbr_[u]) {
            if (seen[v] || cap_[(size_t)u * n_ + v] <= 0) continue;
            queue.push_back(v);
            seen[v] = 1;
            (*parent)[v] = u;
            if (v == dst) return true;
        }
    }
    return false;
}

int64_t FlowGraph::UpdateMaxFlow(int64_t t) {
    BuildEdgeCapacity(t);
    int64_t total = 0;
    std::vector<int> parent;
    while (Bfs(src_, sink_, &parent)) {
        int64_t push = kInf;
        for (int v = sink_; v != src_; v = parent[v]) push = std::min(push, cap_[(size_t)parent[v] * n_ + v]);
        total += push;
        for (int v = sink_; v != src_; v = parent[v]) {
            cap_[(size_t)parent[v] * n_ + v] -= push;
            cap_[(size_t)v * n_ + parent[v]] += push;
        }
    }
    return total;
}

int64_t FlowGraph::GetJobAssignment(std::vector<FlowJobInfo> *jobs) {
    int64_t required = 0;
    for (auto &v : needed_) {                         // each distinct layer once, flow.go:150-152 (per pair when extended)
        auto sz = sizes_.find(v.first);
        required += sz == sizes_.end() ? 0 : sz->second;
    }
    int64_t upper = 1;
    for (;;) {                                        // flow.go:155-167
        if (UpdateMaxFlow(upper) >= required) break;
        if (upper > kInf / 2) throw std::runtime_error("tUpper not found: no finite time satisfies the assignment");
        upper *= 2;
    }
    int64_t lo = 1, hi = upper, best = upper;
    while (lo <= hi) {                                // flow.go:176-187
        int64_t mid = lo + (hi - lo) / 2;
        if (UpdateMaxFlow(mid) < required) lo = mid + 1;
        else { best = std::min(best, mid); hi = mid - 1; }
    }
    UpdateMaxFlow(best);                              // flow.go:191
    jobs->clear();
    std::map<std::pair<LayerID, NodeID>, int64_t> offset;
    for (auto &s : status_)                           // Go iterates this map in random order (:197); ascending here
        for (auto &l : s.second)
            for (auto &v : layer_vertices(l.first)) {
                int c = client_idx_[{s.first, (int)l.second.Source}];
                int64_t flow = cap_[(size_t)layer_idx_[v] * n_ + c];   // reverse residual edge layer -> source, :205
                if (flow > 0) {
                    FlowJobInfo j;
                    j.SenderID = s.first;
                    j.Layer = l.first;
                    j.DataSize = flow;
                    j.Offset = offset[v];
                    j.Dest = v.second;
                    offset[v] += flow;
                    jobs->push_back(j);
                }
            }
    return best;
}

}  // namespace host
}  // namespace dissem
