// C ABI of the host mirror (include/dissem_host.h).
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "../../include/dissem_host.h"
#include "../csrc/dissem_engine.h"
#include "config.h"
#include "message.h"
#include "node.h"
#include "plan.h"
#include "transport.h"

using namespace dissem;
using namespace dissem::host;

static char *dup_cstr(const std::string &s) {
    char *p = (char *)malloc(s.size() + 1);
    if (p) memcpy(p, s.c_str(), s.size() + 1);
    return p;
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

int dissem_msg_roundtrip_json(const char *kind_json, char **wire_out, char **decoded_out) {
    if (!kind_json || !wire_out || !decoded_out) return set_error(DISSEM_EINVAL, "dissem_msg_roundtrip_json: bad arguments");
    try {
        Json k = Json::parse(kind_json);
        const Json *kind = k.get("kind");
        if (!kind) return set_error(DISSEM_EINVAL, "missing kind");
        auto geti = [&](const char *key) { const Json *v = k.get(key); return v ? v->as_int() : 0; };
        MessagePtr m;
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
            m = a;
        } else if (kind->s == "ack") {
            auto a = std::make_shared<AckMsg>();
            a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID"); a->location = (LayerLocation)geti("location");
            m = a;
        } else if (kind->s == "retransmit") {
            auto a = std::make_shared<RetransmitMsg>();
            a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID"); a->DestID = (NodeID)geti("DestID");
            m = a;
        } else if (kind->s == "flowRetransmit") {
            auto a = std::make_shared<FlowRetransmitMsg>();
            a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID"); a->DestID = (NodeID)geti("DestID");
            a->DataSize = geti("DataSize"); a->Offset = geti("Offset"); a->Rate = geti("Rate");
            m = a;
        } else if (kind->s == "startup") {
            auto a = std::make_shared<StartupMsg>();
            a->SrcID = (NodeID)geti("SrcID");
            m = a;
        } else if (kind->s == "simple") {
            auto a = std::make_shared<SimpleMsg>();
            if (const Json *v = k.get("SrcAddr")) a->SrcAddr = v->s;
            if (const Json *v = k.get("PayloadStr")) a->PayloadStr = v->s;
            m = a;
        } else if (kind->s == "layerHeader") {
            auto a = std::make_shared<LayerMsg>();
            a->SrcID = (NodeID)geti("SrcID"); a->Layer = (LayerID)geti("LayerID");
            a->Src_.InDevice = true; a->Src_.DataSize = geti("LayerSize"); a->Src_.Offset = geti("Offert");
            a->TotalSize = geti("TotalSize");
            m = a;
        } else return set_error(DISSEM_EINVAL, "unknown kind '%s'", kind->s.c_str());
        std::string wire = EncodeMsg(*m);
        MessagePtr back = DecodeMsg(wire);
        Json d = Json::object();
        d.set("type", Json::of((int64_t)back->Type()));
        d.set("src", Json::of(back->Src()));
        d.set("payload", back->Payload());
        if (auto *ack = dynamic_cast<AckMsg *>(back.get())) d.set("location_after_decode", Json::of((int64_t)ack->location));
        *wire_out = dup_cstr(wire);
        *decoded_out = dup_cstr(d.dump());
        return DISSEM_OK;
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

    // CreateLayers for every node, outside the clock as in the reference (cmd/main.go:101 precedes :173)
    for (auto &n : cfg.Nodes)
        for (auto &kv : n.InitialLayers) {
            rc = dissem_layer_alloc(ctx, (int)n.ID, (int)kv.first, kv.second.LayerSize);
            if (rc) return rc;
            rc = dissem_layer_fill(ctx, (int)n.ID, (int)kv.first, fill, seed);
            if (rc) return rc;
        }

    DeviceFabric fabric(ctx);
    std::vector<std::unique_ptr<DeviceTransport>> transports;
    std::vector<std::unique_ptr<N>> nodes;
    std::unique_ptr<LeaderNode> leader;
    std::vector<std::unique_ptr<ReceiverNode>> receivers;
    for (auto &n : cfg.Nodes) {
        transports.emplace_back(new DeviceTransport(&fabric, n.ID));
        nodes.emplace_back(new N(n.ID, leader_conf->ID, transports.back().get()));
        LayersSrc layers = CreateLayers(n);
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
    *json_out = dup_cstr(out.dump());
    return *json_out ? DISSEM_OK : set_error(DISSEM_ENOMEM, "out of memory");
}

}  // extern "C"
