#include "config.h"

#include <algorithm>
#include <cstdlib>

namespace dissem {
namespace host {

const NodeConf *Config::Leader() const {
    for (auto &n : Nodes)
        if (n.IsLeader) return &n;
    return nullptr;
}

const NodeConf *Config::Node(NodeID id) const {
    for (auto &n : Nodes)
        if (n.ID == id) return &n;
    return nullptr;
}

std::map<NodeID, int64_t> Config::NodeNetworkBW() const {
    std::map<NodeID, int64_t> m;
    for (auto &n : Nodes) m[n.ID] = n.NetworkBW;
    return m;
}

static uint32_t key_u32(const std::string &k) {
    char *end = nullptr;
    unsigned long v = strtoul(k.c_str(), &end, 10);
    if (k.empty() || *end != 0) throw JsonError("map key '" + k + "' is not an unsigned integer");
    return (uint32_t)v;
}

Config ParseConfig(const std::string &text) {
    Json root = Json::parse(text);
    if (!root.is_object()) throw JsonError("config: top level is not an object");
    Config cfg;
    if (const Json *ls = root.get("LayerSize")) cfg.LayerSize = ls->as_int();
    if (const Json *nodes = root.get("Nodes")) {
        for (auto &nd : nodes->arr) {
            if (!nd.is_object()) throw JsonError("config: Nodes entry is not an object");
            NodeConf nc;
            if (const Json *v = nd.get("ID")) nc.ID = (NodeID)v->as_int();
            if (const Json *v = nd.get("Addr")) nc.Addr = v->s;
            if (const Json *v = nd.get("NetworkBW")) nc.NetworkBW = v->as_int();
            if (const Json *v = nd.get("IsLeader")) nc.IsLeader = v->kind == Json::Bool ? v->b : v->as_int() != 0;
            if (const Json *v = nd.get("Sources"))
                for (auto &kv : v->obj) nc.Sources[(SourceType)key_u32(kv.first)] = kv.second.as_int();
            const Json *init = nd.get("InitialLayers");
            if (init && init->is_object() && !init->obj.empty()) {
                // flat legacy form: {"<layer>": {}} or {"<layer>": {"LayerSize": n}}
                bool flat = true;
                for (auto &kv : init->obj) {
                    const Json &v = kv.second;
                    if (!v.is_object()) { flat = false; break; }
                    if (v.obj.empty()) continue;
                    const Json *sz = v.get("LayerSize");
                    if (!(sz && sz->is_number())) { flat = false; break; }
                }
                // sort keys numerically so that "later wins" is deterministic (Go iterates maps randomly)
                std::vector<std::pair<uint32_t, const Json *>> outer;
                for (auto &kv : init->obj) outer.emplace_back(key_u32(kv.first), &kv.second);
                std::sort(outer.begin(), outer.end(), [](auto &a, auto &b) { return a.first < b.first; });
                if (flat) {
                    cfg.FlatSchema = true;
                    for (auto &o : outer) {
                        const Json *sz = o.second->get("LayerSize");
                        InitialLayerConf ic;
                        ic.LayerSize = std::max<int64_t>(sz ? sz->as_int() : cfg.LayerSize, 0);
                        ic.Source = SourceClient;
                        nc.InitialLayers[o.first] = ic;
                    }
                } else {
                    for (auto &o : outer) {
                        if (!o.second->is_object()) throw JsonError("config: InitialLayers[source] is not an object");
                        std::vector<std::pair<uint32_t, const Json *>> inner;
                        for (auto &kv : o.second->obj) inner.emplace_back(key_u32(kv.first), &kv.second);
                        std::sort(inner.begin(), inner.end(), [](auto &a, auto &b) { return a.first < b.first; });
                        for (auto &l : inner) {
                            const Json *sz = l.second->get("LayerSize");
                            InitialLayerConf ic;
                            ic.LayerSize = std::max<int64_t>(sz ? sz->as_int() : 0, 0);     // cmd/config.go:99-102
                            ic.Source = (SourceType)o.first;
                            nc.InitialLayers[l.first] = ic;
                        }
                    }
                }
            }
            cfg.Nodes.push_back(std::move(nc));
        }
    }
    if (const Json *a = root.get("Assignment")) {
        for (auto &kv : a->obj) {
            LayerIDs ids;
            for (auto &lv : kv.second.obj) ids[key_u32(lv.first)] = LayerMeta{};
            cfg.Assign[key_u32(kv.first)] = std::move(ids);
        }
    }
    return cfg;
}

std::string DiskLayerPath(const std::string &storage_path, NodeID node, LayerID layer) {
    std::string dir = storage_path;
    if (!dir.empty() && dir.back() != '/') dir += '/';
    return dir + "layers/" + std::to_string(node) + "/" + std::to_string(layer) + ".layer";
}

LayersSrc CreateLayers(const NodeConf &nc, bool save_disk, const std::string &storage_path) {
    LayersSrc layers;
    for (auto &kv : nc.InitialLayers) {
        LayerSrc ls;
        ls.DataSize = kv.second.LayerSize < 0 ? 0 : kv.second.LayerSize;   // cmd/config.go:99-102
        ls.Offset = 0;
        if (save_disk) {                                      // CreateDiskLayer, cmd/config.go:133-157
            ls.InDevice = false;
            ls.Fp = DiskLayerPath(storage_path, nc.ID, kv.first);
            ls.Meta.Location = DiskLayer;
        } else {                                              // CreateInmemLayer, cmd/config.go:159-171
            ls.InDevice = true;
            ls.Meta.Location = InmemLayer;
        }
        auto it = nc.Sources.find(kv.second.Source);
        ls.Meta.LimitRate = it == nc.Sources.end() ? 0 : it->second;   // cmd/config.go:111
        ls.Meta.DataSize = ls.DataSize;
        layers[kv.first] = ls;
    }
    return layers;
}

}  // namespace host
}  // namespace dissem
