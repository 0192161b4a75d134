// C ABI of the host mirror (include/dissem_host.h).
#include <errno.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>

#include "../../include/dissem_host.h"
#include "../csrc/dissem_engine.h"
#include "config.h"
#include "message.h"
#include "node.h"
#include "pace.h"
#include "plan.h"
#include "transport.h"

using namespace dissem;
using namespace dissem::host;

static char *dup_cstr(const std::string &s) {
    char *p = (char *)malloc(s.size() + 1);
    if (p) memcpy(p, s.c_str(), s.size() + 1);
    return p;
}

// mkdir -p for the directory part of `path` (os.MkdirAll, cmd/config.go:136)
static int make_parent_dirs(const std::string &path) {
    for (size_t i = 1; i < path.size(); ++i)
        if (path[i] == '/') {
            const std::string dir = path.substr(0, i);
            if (mkdir(dir.c_str(), 0755) != 0 && errno != EEXIST) return set_error(DISSEM_EIO, "mkdir %s: %s", dir.c_str(), strerror(errno));
        }
    return 0;
}

// CreateDiskLayer (cmd/config.go:133-157): the file is written only when it does not exist yet.  The reference writes
// zeros; here the contents follow the run's fill kind, produced by the fill kernel in a scratch HBM buffer of `node`
// (this library computes nothing on the CPU) and read back.
static int create_disk_layer(dissem_ctx *ctx, int node, int layer, int64_t size, int fill, uint64_t seed, const std::string &path) {
    struct stat st;
    if (stat(path.c_str(), &st) == 0) return 0;
    int rc = make_parent_dirs(path);
    if (rc) return rc;
    rc = dissem_layer_alloc(ctx, node, layer, size);
    if (rc) return rc;
    void *host = nullptr;
    rc = dissem_layer_fill(ctx, node, layer, fill, seed);
    if (!rc && size > 0) rc = dissem_host_alloc(&host, size);
    if (!rc && size > 0) rc = dissem_layer_read(ctx, node, layer, host, 0, size);
    if (!rc) {
        int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0644);
        if (fd < 0) rc = set_error(DISSEM_EIO, "create %s: %s", path.c_str(), strerror(errno));
        for (int64_t done = 0; !rc && done < size;) {
            ssize_t k = write(fd, (const char *)host + done, (size_t)(size - done));
            if (k < 0 && errno == EINTR) continue;
            if (k <= 0) rc = set_error(DISSEM_EIO, "write %s: %s", path.c_str(), strerror(errno));
            else done += k;
        }
        if (fd >= 0) close(fd);
    }
    if (host) dissem_host_free(host);
    int rc2 = dissem_layer_free(ctx, node, layer);
    return rc ? rc : rc2;
}

static PlanOptions make_opts(uint64_t seed, int policy, int64_t time_scale, int flags) {
    PlanOptions o;
    o.Seed = seed;
    o.Policy = policy == DISSEM_POLICY_BALANCED ? OwnerBalanced : OwnerRandom;
    o.Flow.TimeScale = time_scale > 0 ? time_scale : 1;
    o.Flow.ZeroRateUnlimited = (flags & DISSEM_PLAN_ZERO_RATE_UNLIMITED) != 0;
    o.SizesFromAnnounces = (flags & DISSEM_PLAN_SIZES_FROM_ANNOUNCES) != 0;
    o.Flow.PerDestLayers = (flags & DISSEM_PLAN_MULTI_DEST) != 0;
    return o;
}

static Json config_json(const Config &cfg) {
    Json nodes = Json::array();
    for (auto &n : cfg.Nodes) {
        Json layers = Json::object();
        for (auto &kv : n.InitialLayers) layers.set(std::to_string(kv.first), Json::of(kv.second.LayerSize));
        Json nd = Json::object();
        nd.set("id", Json::of((int64_t)n.ID));
        nd.set("leader", Json::boolean(n.IsLeader));
        nd.set("bw", Json::of(n.NetworkBW));
        nd.set("layers", std::move(layers));
        nodes.push(std::move(nd));
    }
    return nodes;
}

extern "C" {

void dissem_free(void *p) { free(p); }

int dissem_plan_json(const char *config_json_text, int mode, uint64_t seed, int policy, int64_t time_scale, int flags,
                     char **json_out) {
    if (!config_json_text || !json_out) return set_error(DISSEM_EINVAL, "dissem_plan_json: bad arguments");
    Config cfg;
    try {
        cfg = ParseConfig(config_json_text);
    } catch (const std::exception &e) {
        return set_error(DISSEM_EPARSE, "config: %s", e.what());
    }
    const NodeConf *leader = cfg.Leader();
    if (!leader) return set_error(DISSEM_EPLAN, "no leader found");         // cmd/config.go:71
    Plan plan;
    try {
        plan = MakePlan(mode, leader->ID, CreateLayers(*leader), StatusFromConfig(cfg), cfg.Assign, cfg.NodeNetworkBW(),
                        make_opts(seed, policy, time_scale, flags));
    } catch (const std::exception &e) {
        return set_error(DISSEM_EPLAN, "%s", e.what());
    }
    Json out = Json::object();
    out.set("mode", Json::of((int64_t)mode));
    out.set("leader", Json::of((int64_t)leader->ID));
    out.set("t", Json::of(plan.MinTime));
    out.set("n_nodes", Json::of((int64_t)cfg.Nodes.size()));
    Json xs = Json::array();
    for (auto &x : plan.Xfers) {
        Json r = Json::array();
        r.push(Json::of((int64_t)x.Src)).push(Json::of((int64_t)x.Dest)).push(Json::of((int64_t)x.Layer));
        r.push(Json::of(x.Offset)).push(Json::of(x.DataSize)).push(Json::of(x.TotalSize)).push(Json::of(x.Rate));
        xs.push(std::move(r));
    }
    out.set("xfers", std::move(xs));
    Json sj = Json::array();
    for (auto &j : plan.SelfJobs) {
        Json r = Json::array();
        r.push(Json::of((int64_t)j.SenderID)).push(Json::of((int64_t)j.Layer)).push(Json::of(j.DataSize));
        sj.push(std::move(r));
    }
    out.set("self_jobs", std::move(sj));
    out.set("nodes", config_json(cfg));
    Json as = Json::object();
    for (auto &na : cfg.Assign) {
        Json ls = Json::array();
        for (auto &la : na.second) ls.push(Json::of((int64_t)la.first));
        as.set(std::to_string(na.first), std::move(ls));
    }
    out.set("assignment", std::move(as));
    out.set("flat_schema", Json::boolean(cfg.FlatSchema));
    *json_out = dup_cstr(out.dump());
    return *json_out ? DISSEM_OK : set_error(DISSEM_ENOMEM, "out of memory");
}

// One message of any kind from its JSON description {"kind":..., fields...}; throws on an unknown kind.
static MessagePtr msg_from_json(const Json &k) {
    const Json *kind = k.get("kind");
    if (!kind) throw std::runtime_error("missing kind");
    auto geti = [&](const char *key) { const Json *v = k.get(key); return v ? v->as_int() : 0; };
    if (kind->s == "announce") {
        auto a = std::make_shared<AnnounceMsg>();
        a->SrcID = (NodeID)geti("SrcID");
        if (const Json *ids = k.get("LayerIDs"))
            for (auto &kv : ids->obj) {
                LayerMeta meta;
                if (const Json *v = kv.second.get("Location")) meta.Location = (LayerLocation)v->as_int();
                if (const Json *v = kv.second.get("LimitRate")) meta.LimitRate = v->as_int();
                if (const Json *v = kv.second.get("SourceType")) meta.Source = (SourceType)v->as_int();
                if (const Json *v = kv.second.get("DataSize")) meta.DataSize = v->as_int();
                a->Layers[(LayerID)strtoul(kv.first.c_str(), nullptr, 10)] = meta;
            }
        return a;
    }
    if (kind->s == "ack") {
        auto a = std::make_shared<AckMsg>();
        a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID"); a->location = (LayerLocation)geti("location");
        return a;
    }
    if (kind->s == "retransmit") {
        auto a = std::make_shared<RetransmitMsg>();
        a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID"); a->DestID = (NodeID)geti("DestID");
        return a;
    }
    if (kind->s == "flowRetransmit") {
        auto a = std::make_shared<FlowRetransmitMsg>();
        a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID"); a->DestID = (NodeID)geti("DestID");
        a->DataSize = geti("DataSize"); a->Offset = geti("Offset"); a->Rate = geti("Rate");
        return a;
    }
    if (kind->s == "startup") {
        auto a = std::make_shared<StartupMsg>();
        a->SrcID = (NodeID)geti("SrcID");
        return a;
    }
    if (kind->s == "simple") {
        auto a = std::make_shared<SimpleMsg>();
        if (const Json *v = k.get("SrcAddr")) a->SrcAddr = v->s;
        if (const Json *v = k.get("PayloadStr")) a->PayloadStr = v->s;
        return a;
    }
    if (kind->s == "layerHeader") {
        auto a = std::make_shared<LayerMsg>();
        a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID");
        a->Src_.InDevice = true; a->Src_.DataSize = geti("LayerSize"); a->Src_.Offset = geti("Offert");
        a->TotalSize = geti("TotalSize");
        return a;
    }
    throw std::runtime_error("unknown kind '" + kind->s + "'");
}

// What a receiver observes of a delivered message
static Json msg_to_json(const Message &m) {
    Json d = Json::object();
    d.set("type", Json::of((int64_t)m.Type()));
    d.set("src", Json::of(m.Src()));
    d.set("payload", m.Payload());
    if (auto *ack = dynamic_cast<const AckMsg *>(&m)) d.set("location_after_decode", Json::of((int64_t)ack->location));
    return d;
}

int dissem_msg_roundtrip_json(const char *kind_json, char **wire_out, char **decoded_out) {
    if (!kind_json || !wire_out || !decoded_out) return set_error(DISSEM_EINVAL, "dissem_msg_roundtrip_json: bad arguments");
    try {
        MessagePtr m = msg_from_json(Json::parse(kind_json));
        std::string wire = EncodeMsg(*m);
        MessagePtr back = DecodeMsg(wire);
        *wire_out = dup_cstr(wire);
        *decoded_out = dup_cstr(msg_to_json(*back).dump());
        return DISSEM_OK;
    } catch (const std::exception &e) {
        return set_error(DISSEM_EPARSE, "%s", e.what());
    }
}

int dissem_transport_script_json(const char *script_json, char **json_out) {
    if (!script_json || !json_out) return set_error(DISSEM_EINVAL, "dissem_transport_script_json: bad arguments");
    try {
        Json script = Json::parse(script_json);
        const Json *peers = script.get("peers");
        const Json *ops = script.get("ops");
        if (!peers || !ops) return set_error(DISSEM_EINVAL, "script needs \"peers\" and \"ops\"");
        DeviceFabric fabric(nullptr);                   // control plane only: no device context behind it
        std::map<NodeID, std::unique_ptr<DeviceTransport>> ts;
        for (auto &p : peers->arr) ts[(NodeID)p.as_int()].reset(new DeviceTransport(&fabric, (NodeID)p.as_int()));
        Json results = Json::array();
        for (auto &op : ops->arr) {
            const Json *name = op.get("op");
            auto geti = [&](const char *key) { const Json *v = op.get(key); return v ? v->as_int() : -1; };
            int rc = DISSEM_EINVAL;
            if (name && name->s == "send") {
                auto it = ts.find((NodeID)geti("from"));
                if (it != ts.end() && op.get("msg")) rc = it->second->Send((NodeID)geti("to"), msg_from_json(*op.get("msg")));
            } else if (name && name->s == "broadcast") {
                auto it = ts.find((NodeID)geti("from"));
                if (it != ts.end() && op.get("msg")) rc = it->second->Broadcast(msg_from_json(*op.get("msg")));
            } else if (name && name->s == "close") {
                auto it = ts.find((NodeID)geti("peer"));
                if (it != ts.end()) rc = it->second->Close();
            } else if (name && name->s == "pipe") {
                auto it = ts.find((NodeID)geti("from"));
                if (it != ts.end()) rc = it->second->RegisterPipe((LayerID)geti("layer"), (NodeID)geti("to"));
            }
            results.push(Json::of((int64_t)rc));
        }
        // Deliver(): what each peer's consumer would read, in order (channels are closed first so that Pop ends)
        Json delivered = Json::object();
        Json addrs = Json::object();
        for (auto &kv : ts) {
            addrs.set(std::to_string(kv.first), Json::of(kv.second->GetAddress()));
            kv.second->Close();
            Json list = Json::array();
            MessagePtr m;
            while (kv.second->Deliver().Pop(&m)) list.push(msg_to_json(*m));
            delivered.set(std::to_string(kv.first), std::move(list));
        }
        Json out = Json::object();
        out.set("results", std::move(results));
        out.set("delivered", std::move(delivered));
        out.set("addresses", std::move(addrs));
        *json_out = dup_cstr(out.dump());
        return *json_out ? DISSEM_OK : set_error(DISSEM_ENOMEM, "out of memory");
    } catch (const std::exception &e) {
        return set_error(DISSEM_EPARSE, "%s", e.what());
    }
}

int dissem_run_json(const char *config_json_text, int mode, const int *node_device, const char *opts_json, char **json_out) {
    if (!config_json_text || !json_out) return set_error(DISSEM_EINVAL, "dissem_run_json: bad arguments");
    Config cfg;
    Json opts = Json::object();
    try {
        cfg = ParseConfig(config_json_text);
        if (opts_json && *opts_json) opts = Json::parse(opts_json);
    } catch (const std::exception &e) {
        return set_error(DISSEM_EPARSE, "%s", e.what());
    }
    auto opt_i = [&](const char *k, int64_t dflt) { const Json *v = opts.get(k); return v ? v->as_int() : dflt; };
    const NodeConf *leader_conf = cfg.Leader();
    if (!leader_conf) return set_error(DISSEM_EPLAN, "no leader found");
    if (mode < 0 || mode > 3) return set_error(DISSEM_EPLAN, "unknown mode %d", mode);            // cmd/main.go:163-164
    const uint64_t seed = (uint64_t)opt_i("seed", 0);
    const int fill = (int)opt_i("fill", DISSEM_FILL_PATTERN);
    const bool batched = opt_i("batched", 0) != 0;
    const bool verify = opt_i("verify", 1) != 0;
    const int timeout_ms = (int)opt_i("timeout_ms", 60000);
    PlanOptions popt = make_opts(seed, (int)opt_i("policy", 0), opt_i("time_scale", 1), (int)opt_i("flags", 0));

    // node ids of the config -> dense engine node indices
    NodeID max_id = 0;
    for (auto &n : cfg.Nodes) max_id = std::max(max_id, n.ID);
    for (auto &na : cfg.Assign) max_id = std::max(max_id, na.first);
    if (max_id > 4096) return set_error(DISSEM_EINVAL, "node ids must be small integers");
    std::vector<int> dev(max_id + 1, 0);
    for (size_t i = 0; i < cfg.Nodes.size(); ++i) dev[cfg.Nodes[i].ID] = node_device ? node_device[i] : 0;
    dissem_ctx *ctx = nullptr;
    int rc = dissem_init((int)dev.size(), dev.data(), &ctx);
    if (rc) return rc;
    struct CtxGuard { dissem_ctx *c; ~CtxGuard() { dissem_close(c); } } guard{ctx};

    // CreateLayers for every node, outside the clock as in the reference (cmd/main.go:101 precedes :173).
    // "storage" (the -s flag): the initial layers are files under <storage>/layers/<node>/ instead of HBM buffers.
    const Json *storage_opt = opts.get("storage");
    const bool save_disk = storage_opt && storage_opt->kind == Json::String;
    const std::string storage = save_disk ? storage_opt->s : "";
    for (auto &n : cfg.Nodes)
        for (auto &kv : n.InitialLayers) {
            const int64_t size = kv.second.LayerSize < 0 ? 0 : kv.second.LayerSize;
            if (save_disk) {
                const std::string path = DiskLayerPath(storage, n.ID, kv.first);
                rc = create_disk_layer(ctx, (int)n.ID, (int)kv.first, size, fill, seed, path);
                if (rc) return rc;
                rc = dissem_layer_attach_file(ctx, (int)n.ID, (int)kv.first, path.c_str(), 0, size);
                if (rc) return rc;
                continue;
            }
            rc = dissem_layer_alloc(ctx, (int)n.ID, (int)kv.first, size);
            if (rc) return rc;
            rc = dissem_layer_fill(ctx, (int)n.ID, (int)kv.first, fill, seed);
            if (rc) return rc;
        }

    // Receive buffers for what the Assignment will deliver are set aside before the clock starts ("reserve":0 turns
    // this off and leaves the allocation inside the send, where the reference has it: transport.go:201).
    if (opt_i("reserve", 1) != 0) {
        std::map<LayerID, int64_t> size_of;
        for (auto &kv : leader_conf->InitialLayers) size_of[kv.first] = kv.second.LayerSize;
        for (auto &n : cfg.Nodes)
            for (auto &kv : n.InitialLayers) size_of.emplace(kv.first, kv.second.LayerSize);
        for (auto &na : cfg.Assign) {
            const NodeConf *nc = cfg.Node(na.first);
            for (auto &la : na.second) {
                const bool in_memory = !save_disk && nc && nc->InitialLayers.count(la.first);
                auto so = size_of.find(la.first);
                if (in_memory || so == size_of.end()) continue;
                rc = dissem_layer_reserve(ctx, (int)na.first, so->second, 1);
                if (rc) return rc;
            }
        }
    }

    DeviceFabric fabric(ctx);
    fabric.pace = opt_i("pace", 0) != 0;
    std::vector<std::unique_ptr<DeviceTransport>> transports;
    std::vector<std::unique_ptr<N>> nodes;
    std::unique_ptr<LeaderNode> leader;
    std::vector<std::unique_ptr<ReceiverNode>> receivers;
    for (auto &n : cfg.Nodes) {
        transports.emplace_back(new DeviceTransport(&fabric, n.ID));
        nodes.emplace_back(new N(n.ID, leader_conf->ID, transports.back().get()));
        LayersSrc layers = CreateLayers(n, save_disk, storage);
        if (n.ID == leader_conf->ID) {
            switch (mode) {                                   // RunLeader, cmd/main.go:153-165
                case 0: leader.reset(new LeaderNode(nodes.back().get(), layers, cfg.Assign)); break;
                case 1: leader.reset(new RetransmitLeaderNode(nodes.back().get(), layers, cfg.Assign, popt)); break;
                case 2: leader.reset(new PullRetransmitLeaderNode(nodes.back().get(), layers, cfg.Assign)); break;
                default: leader.reset(new FlowRetransmitLeaderNode(nodes.back().get(), layers, cfg.Assign, cfg.NodeNetworkBW(), popt)); break;
            }
            leader->SetBatched(batched);
        } else {
            switch (mode) {                                   // RunReceiver, cmd/main.go:187-197
                case 0: receivers.emplace_back(new ReceiverNode(nodes.back().get(), layers)); break;
                case 1:
                case 2: receivers.emplace_back(new RetransmitReceiverNode(nodes.back().get(), layers)); break;
                default: receivers.emplace_back(new FlowRetransmitReceiverNode(nodes.back().get(), layers)); break;
            }
        }
    }
    auto stop_all = [&] {
        for (auto &r : receivers) r->Stop();
        if (leader) leader->Stop();
    };
    for (auto &r : receivers) {                               // executeReceiver, cmd/main.go:204-215
        rc = r->Announce();
        if (rc) { stop_all(); return rc; }
    }
    std::string fail;
    if (cfg.Nodes.size() == 1 || cfg.Assign.empty()) fail = "nothing to distribute";
    Assignment got;
    double ttd_ms = 0.0;
    if (fail.empty()) {                                       // executeLeader, cmd/main.go:173-181
        if (!leader->StartDistribution().Wait(timeout_ms)) fail = "timeout waiting for announcements from receivers";
        else if (!leader->Ready().Wait(timeout_ms, &got)) fail = "timeout waiting for Ready()";
        else ttd_ms = std::chrono::duration<double, std::milli>(leader->Ready().when() - leader->StartDistribution().when()).count();
    }
    if (fail.empty())
        for (auto &r : receivers)
            if (!r->Ready().Wait(timeout_ms)) { fail = "a receiver never saw the startup message"; break; }
    std::string role_err = leader ? leader->LastError() : "";
    for (auto &r : receivers)
        if (role_err.empty()) role_err = r->LastError();
    Plan plan = leader ? leader->LastPlan() : Plan{};
    stop_all();
    if (!role_err.empty()) return set_error(DISSEM_EPLAN, "%s", role_err.c_str());
    if (!fail.empty()) return set_error(DISSEM_EPLAN, "%s", fail.c_str());

    // delivery check, all on the GPU: fused table at the destination == re-read of the destination == re-read of a source copy
    bool verified = true;
    int64_t checked = 0;
    if (verify) {
        for (auto &na : cfg.Assign)
            for (auto &la : na.second) {
                int64_t size = 0;
                rc = dissem_layer_info(ctx, (int)na.first, (int)la.first, &size, nullptr);
                if (rc) return set_error(DISSEM_EPLAN, "node %u ended without layer %u", na.first, la.first);
                int64_t nc = (size + DISSEM_CHUNK_BYTES - 1) / DISSEM_CHUNK_BYTES;
                std::vector<uint64_t> at_dst(2 * nc), src_ref(2 * nc), fused(2 * nc);
                rc = dissem_layer_verify(ctx, (int)na.first, (int)la.first, at_dst.data(), nc);
                if (rc) return rc;
                int src_node = -1;
                bool received_here = false;
                for (auto &x : plan.Xfers)
                    if (x.Dest == na.first && x.Layer == la.first) { received_here = true; src_node = (int)x.Src; }
                if (!received_here) continue;             // it held the layer from the start
                if (dissem_layer_info(ctx, src_node, (int)la.first, nullptr, nullptr) == DISSEM_ENOENT &&
                    dissem_layer_file_info(ctx, src_node, (int)la.first, nullptr) == DISSEM_OK) {
                    // the source holds the layer on disk only: load it into the source's HBM (outside the clock) to re-read it
                    dissem_ticket tk = 0;
                    rc = dissem_layer_alloc(ctx, src_node, (int)la.first, size);
                    if (!rc) rc = dissem_send_file(ctx, src_node, src_node, (int)la.first, 0, size, size, &tk);
                    if (!rc) rc = dissem_wait(ctx, tk, nullptr);
                    if (rc) return rc;
                }
                rc = dissem_layer_verify(ctx, src_node, (int)la.first, src_ref.data(), nc);
                if (rc) return rc;
                rc = dissem_layer_checksums(ctx, (int)na.first, (int)la.first, fused.data(), nc);
                if (rc) return rc;
                int64_t rcv = 0;
                rc = dissem_layer_received(ctx, (int)na.first, (int)la.first, &rcv);
                if (rc) return rc;
                if (at_dst != src_ref || fused != src_ref || rcv != size) verified = false;
                ++checked;
            }
    }
    // "dump_tables": the raw evidence of every assigned (node, layer) — fused checksum table, received counter, first and
    // last 4 KiB — so that a caller can hold the run against its own oracle instead of trusting `verified` above, which
    // only compares values this library computed with each other.
    Json tables = Json::array();
    if (opt_i("dump_tables", 0) != 0) {
        auto hex = [](const uint8_t *p, size_t n) {
            static const char *digits = "0123456789abcdef";
            std::string h(2 * n, '0');
            for (size_t i = 0; i < n; ++i) { h[2 * i] = digits[p[i] >> 4]; h[2 * i + 1] = digits[p[i] & 15]; }
            return h;
        };
        for (auto &na : cfg.Assign)
            for (auto &la : na.second) {
                int64_t size = 0;
                rc = dissem_layer_info(ctx, (int)na.first, (int)la.first, &size, nullptr);
                if (rc) return set_error(DISSEM_EPLAN, "node %u ended without layer %u", na.first, la.first);
                const int64_t nc = (size + DISSEM_CHUNK_BYTES - 1) / DISSEM_CHUNK_BYTES;
                std::vector<uint64_t> fused((size_t)std::max<int64_t>(2 * nc, 1));
                rc = dissem_layer_checksums(ctx, (int)na.first, (int)la.first, fused.data(), nc);
                if (rc) return rc;
                int64_t rcv = 0;
                rc = dissem_layer_received(ctx, (int)na.first, (int)la.first, &rcv);
                if (rc) return rc;
                const int64_t edge = std::min<int64_t>(size, 4096);
                std::vector<uint8_t> head((size_t)std::max<int64_t>(edge, 1)), tail((size_t)std::max<int64_t>(edge, 1));
                rc = dissem_layer_read(ctx, (int)na.first, (int)la.first, head.data(), 0, edge);
                if (!rc) rc = dissem_layer_read(ctx, (int)na.first, (int)la.first, tail.data(), size - edge, edge);
                if (rc) return rc;
                Json t = Json::object();
                t.set("node", Json::of((int64_t)na.first));
                t.set("layer", Json::of((int64_t)la.first));
                t.set("size", Json::of(size));
                t.set("received", Json::of(rcv));
                t.set("fused", Json::of(hex(reinterpret_cast<const uint8_t *>(fused.data()), (size_t)nc * 16)));   // little-endian (A, B) pairs
                t.set("head", Json::of(hex(head.data(), (size_t)edge)));
                t.set("tail", Json::of(hex(tail.data(), (size_t)edge)));
                tables.push(std::move(t));
            }
    }
    Json out = Json::object();
    out.set("mode", Json::of((int64_t)mode));
    out.set("batched", Json::boolean(batched));
    Json ttd; ttd.kind = Json::Double; ttd.d = ttd_ms;
    out.set("time_to_deliver_ms", ttd);
    Json dms; dms.kind = Json::Double; dms.d = fabric.device_ms;
    out.set("device_ms", dms);
    out.set("bytes", Json::of(fabric.layer_bytes));
    out.set("n_xfers", Json::of((int64_t)plan.Xfers.size()));
    out.set("layer_msgs", Json::of(fabric.layer_msgs));
    out.set("control_msgs", Json::of(fabric.control_msgs));
    out.set("launches", Json::of(fabric.launches));
    out.set("t", Json::of(plan.MinTime));
    out.set("ready_equals_assignment", Json::boolean(got.size() == cfg.Assign.size()));
    out.set("verified", Json::boolean(verified));
    out.set("layers_checked", Json::of(checked));
    Json xs = Json::array();
    for (auto &x : plan.Xfers) {
        Json r = Json::array();
        r.push(Json::of((int64_t)x.Src)).push(Json::of((int64_t)x.Dest)).push(Json::of((int64_t)x.Layer));
        r.push(Json::of(x.Offset)).push(Json::of(x.DataSize)).push(Json::of(x.TotalSize));
        xs.push(std::move(r));
    }
    out.set("xfers", std::move(xs));
    if (opt_i("dump_tables", 0) != 0) out.set("tables", std::move(tables));
    *json_out = dup_cstr(out.dump());
    return *json_out ? DISSEM_OK : set_error(DISSEM_ENOMEM, "out of memory");
}

int dissem_pace_schedule(int64_t layer_off, int64_t size, int64_t rate_bytes_per_s, int64_t *piece_off, int64_t *piece_size,
                         int64_t *release_ns, int cap, int *n_out) {
    if (layer_off < 0 || size < 0 || cap < 0 || !n_out) return set_error(DISSEM_EINVAL, "dissem_pace_schedule: bad arguments");
    const std::vector<PacePiece> pieces = PaceSchedule(layer_off, size, rate_bytes_per_s);
    *n_out = (int)pieces.size();
    for (int i = 0; i < cap && i < (int)pieces.size(); ++i) {
        if (piece_off) piece_off[i] = pieces[i].off;
        if (piece_size) piece_size[i] = pieces[i].size;
        if (release_ns) release_ns[i] = pieces[i].release_ns;
    }
    return DISSEM_OK;
}

}  // extern "C"
