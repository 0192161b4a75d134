#include "plan.h"

#include <algorithm>
#include <cstdint>

namespace dissem {
namespace host {

uint64_t Mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

Status StatusFromConfig(const Config &cfg) {
    Status st;
    for (auto &n : cfg.Nodes) {
        LayerIDs ids;
        for (auto &kv : CreateLayers(n)) {
            LayerMeta m;
            m.Location = kv.second.Meta.Location;       // Announce sends Location and LimitRate only,
            m.LimitRate = kv.second.Meta.LimitRate;     // node.go:1397-1401; SourceType stays 0
            m.DataSize = kv.second.DataSize;
            ids[kv.first] = m;
        }
        st[n.ID] = std::move(ids);
    }
    return st;
}

Plan PlanMode0(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a) {
    Plan p;
    for (auto &na : a)
        for (auto &la : na.second) {
            auto sn = status.find(na.first);
            if (sn != status.end()) {
                auto sl = sn->second.find(la.first);
                if (sl != sn->second.end() && sl->second.Location == InmemLayer) continue;     // node.go:335
            }
            auto ll = leader_layers.find(la.first);
            if (ll == leader_layers.end())      // node.go:339-342 only warns; the send then never completes
                throw PlanError("mode 0: the leader holds no layer " + std::to_string(la.first));
            PlannedXfer x;
            x.Src = leader;
            x.Dest = na.first;
            x.Layer = la.first;
            x.Offset = 0;
            x.DataSize = x.TotalSize = ll->second.DataSize;
            p.Xfers.push_back(x);
        }
    return p;
}

Plan PlanMode1(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a,
               const PlanOptions &opt) {
    (void)leader;
    (void)leader_layers;
    std::map<LayerID, std::vector<NodeID>> owners;       // layerOwners, node.go:559-568 (lazily keyed, :829-834)
    std::map<LayerID, int64_t> size_at;                  // the owner's own DataSize travels with its layerMsg (:1478)
    for (auto &sn : status)
        for (auto &sl : sn.second) {
            owners[sl.first].push_back(sn.first);
            if (sl.second.DataSize >= 0 && !size_at.count(sl.first)) size_at[sl.first] = sl.second.DataSize;
        }
    std::map<NodeID, int64_t> sent;
    Plan p;
    for (auto &na : a)
        for (auto &la : na.second) {
            auto ow = owners.find(la.first);
            if (ow == owners.end() || ow->second.empty())
                throw PlanError("mode 1: nobody holds layer " + std::to_string(la.first));
            const std::vector<NodeID> &own = ow->second;
            NodeID owner;
            if (std::find(own.begin(), own.end(), na.first) != own.end()) {
                // node.go:578-580 skips a destination that announced the layer, wherever it keeps it; a copy on DISK then
                // never reaches memory and assignmentSatisfied (node.go:440) never holds: the reference waits forever.
                // Fixed here: such a node is told to send the layer to itself, which loads the file into its HBM.
                if (status.at(na.first).at(la.first).Location == InmemLayer) continue;
                owner = na.first;
            } else if (opt.Policy == OwnerRandom) {
                uint64_t h = Mix64(opt.Seed + (uint64_t)(na.first + 1) * 0x9E3779B97F4A7C15ull +
                                   (uint64_t)(la.first + 1) * 0xD1B54A32D192ED03ull);
                owner = own[h % own.size()];
            } else {
                owner = own[0];
                for (NodeID o : own)
                    if (sent[o] < sent[owner]) owner = o;
            }
            PlannedXfer x;
            x.Src = owner;
            x.Dest = na.first;
            x.Layer = la.first;
            x.Offset = 0;
            // the owner slices its OWN copy: NewLayerMsg(..., ls, ls.DataSize), node.go:1478
            const LayerMeta &om = status.at(owner).at(la.first);
            auto ll = leader_layers.find(la.first);
            x.DataSize = x.TotalSize = om.DataSize >= 0 ? om.DataSize
                                        : (size_at.count(la.first) ? size_at[la.first]
                                           : (ll != leader_layers.end() ? ll->second.DataSize : 0));
            sent[owner] += x.DataSize;
            p.Xfers.push_back(x);
        }
    return p;
}

Plan PlanMode3(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a,
               const std::map<NodeID, int64_t> &network_bw, const PlanOptions &opt) {
    (void)leader;
    Plan p;
    std::map<LayerID, int64_t> sizes;
    for (auto &kv : leader_layers) sizes[kv.first] = kv.second.DataSize;                     // flow.go:151
    if (opt.SizesFromAnnounces)
        for (auto &sn : status)
            for (auto &sl : sn.second)
                if (!sizes.count(sl.first) && sl.second.DataSize >= 0) sizes[sl.first] = sl.second.DataSize;
    std::map<LayerID, NodeID> layer_dest;                 // LayerDests: ONE destination per layer, node.go:1087-1095
    for (auto &na : a)
        for (auto &la : na.second) layer_dest.emplace(la.first, na.first);
    Assignment modified;
    for (auto &na : a)                                    // node.go:1206-1218
        for (auto &la : na.second) {
            auto sn = status.find(na.first);
            if (sn != status.end() && sn->second.count(la.first)) {
                FlowJobInfo j;
                j.SenderID = na.first;
                j.Layer = la.first;
                j.DataSize = sizes.count(la.first) ? sizes[la.first] : 0;
                j.Offset = 0;
                p.SelfJobs.push_back(j);
            } else {
                if (!sizes.count(la.first))
                    throw PlanError("mode 3: the leader does not know the size of layer " + std::to_string(la.first));
                modified[na.first][la.first] = la.second;
            }
        }
    if (modified.empty()) return p;                       // node.go:1221-1224
    std::vector<FlowJobInfo> jobs;
    try {
        FlowGraph g(status, sizes, network_bw, modified, opt.Flow);
        p.MinTime = g.GetJobAssignment(&jobs);
    } catch (const std::exception &e) {
        throw PlanError(std::string("mode 3: ") + e.what());
    }
    for (auto &j : jobs) {                                // dispatchJob, node.go:1264-1288
        PlannedXfer x;
        x.Src = j.SenderID;
        x.Dest = j.Dest != kAnyDest ? j.Dest : layer_dest[j.Layer];
        x.Layer = j.Layer;
        x.Offset = j.Offset;
        x.DataSize = j.DataSize;
        x.TotalSize = sizes[j.Layer];
        // node.go:1281: rate = dataSize / minTime, bytes per second.  With a finer clock (TimeScale ticks per second)
        // minTime is in ticks, so the same quotient is scaled back to bytes per second.
        x.Rate = p.MinTime > 0 ? (int64_t)std::min<__int128>((__int128)j.DataSize * opt.Flow.TimeScale / p.MinTime, INT64_MAX) : 0;
        p.Xfers.push_back(x);
    }
    return p;
}

Plan PlanMode2Initial(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a) {
    (void)leader;
    std::map<LayerID, std::vector<NodeID>> owners;                    // node.go:826-838
    for (auto &sn : status)
        for (auto &sl : sn.second) owners[sl.first].push_back(sn.first);
    std::vector<LayerID> sorted;                                      // node.go:841-851
    for (auto &kv : owners) sorted.push_back(kv.first);
    std::stable_sort(sorted.begin(), sorted.end(), [&](LayerID x, LayerID y) {
        if (owners[x].size() != owners[y].size()) return owners[x].size() < owners[y].size();
        return x < y;
    });
    std::map<LayerID, std::vector<NodeID>> jobs;                      // node.go:854-867: (layer -> dests still lacking it in memory)
    for (auto &na : a)
        for (auto &la : na.second) {
            auto sn = status.find(na.first);
            bool has = false;
            if (sn != status.end()) {
                auto sl = sn->second.find(la.first);
                has = sl != sn->second.end() && sl->second.Location == InmemLayer;
            }
            if (!has) jobs[la.first].push_back(na.first);
        }
    std::map<NodeID, unsigned> load;                                  // node.go:870-874
    for (auto &sn : status) load[sn.first] = 0;
    Plan p;
    for (LayerID layer : sorted) {
        auto jl = jobs.find(layer);
        if (jl == jobs.end()) continue;
        for (NodeID dest : jl->second) {                              // Go: map order; ascending here
            NodeID best = 0;                                          // getMinLoadedSender, node.go:948-978
            int64_t best_rate = 0;
            unsigned min_count = ~0u;
            for (auto &lc : load) {
                auto sn = status.find(lc.first);
                auto sl = sn->second.find(layer);
                if (sl == sn->second.end()) continue;
                int64_t rate = sl->second.LimitRate == 0 ? INT64_MAX : sl->second.LimitRate;
                if (rate > best_rate || (rate == best_rate && (lc.second < min_count || (lc.second == min_count && lc.first < best)))) {
                    best = lc.first;
                    best_rate = rate;
                    min_count = lc.second;
                }
            }
            load[best]++;
            PlannedXfer x;
            x.Src = best;
            x.Dest = dest;
            x.Layer = layer;
            const LayerMeta &om = status.at(best).at(layer);
            auto ll = leader_layers.find(layer);
            x.DataSize = x.TotalSize = om.DataSize >= 0 ? om.DataSize : (ll != leader_layers.end() ? ll->second.DataSize : 0);
            p.Xfers.push_back(x);
        }
    }
    for (auto &jl : jobs)
        if (!owners.count(jl.first)) throw PlanError("mode 2: nobody holds layer " + std::to_string(jl.first));
    return p;
}

Plan MakePlan(int mode, NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a,
              const std::map<NodeID, int64_t> &network_bw, const PlanOptions &opt) {
    switch (mode) {                                       // cmd/main.go:153-165
        case 0: return PlanMode0(leader, leader_layers, status, a);
        case 1: return PlanMode1(leader, leader_layers, status, a, opt);
        case 3: return PlanMode3(leader, leader_layers, status, a, network_bw, opt);
        case 2: return PlanMode2Initial(leader, leader_layers, status, a);
        default: throw PlanError("unknown mode");         // cmd/main.go:163-164
    }
}

}  // namespace host
}  // namespace dissem
