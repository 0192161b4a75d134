#include "message.h"

#include <cstdlib>

namespace dissem {
namespace host {

static std::string layer_list(const LayerIDs &l) {
    std::string s = "[";
    bool first = true;
    for (auto &kv : l) {
        if (!first) s += " ";
        s += std::to_string(kv.first);
        first = false;
    }
    return s + "]";
}

std::string AnnounceMsg::String() const { return std::to_string(SrcID) + ": " + layer_list(Layers); }
Json AnnounceMsg::Payload() const {
    Json ids = Json::object();
    for (auto &kv : Layers) {
        Json m = Json::object();
        m.set("Location", Json::of((int64_t)kv.second.Location));
        m.set("LimitRate", Json::of(kv.second.LimitRate));
        m.set("SourceType", Json::of((int64_t)kv.second.Source));
        if (kv.second.DataSize >= 0) m.set("DataSize", Json::of(kv.second.DataSize));   // extension, see types.h
        ids.set(std::to_string(kv.first), std::move(m));
    }
    Json p = Json::object();
    p.set("SrcID", Json::of((int64_t)SrcID));
    p.set("LayerIDs", std::move(ids));
    return p;
}

std::string AckMsg::String() const { return "ack " + std::to_string(SrcID) + ": layer " + std::to_string(Layer); }
Json AckMsg::Payload() const {
    Json p = Json::object();
    p.set("SrcID", Json::of((int64_t)SrcID));
    p.set("LayerID", Json::of((int64_t)Layer));
    return p;
}

std::string RetransmitMsg::String() const {
    return "retransmit layer " + std::to_string(Layer) + " to " + std::to_string(DestID);
}
Json RetransmitMsg::Payload() const {
    Json p = Json::object();
    p.set("SrcID", Json::of((int64_t)SrcID));
    p.set("LayerID", Json::of((int64_t)Layer));
    p.set("DestID", Json::of((int64_t)DestID));
    return p;
}

std::string FlowRetransmitMsg::String() const {
    return "flow retransmit layer " + std::to_string(Layer) + " to " + std::to_string(DestID) + " [" +
           std::to_string(Offset) + ",+" + std::to_string(DataSize) + ")";
}
Json FlowRetransmitMsg::Payload() const {
    Json p = Json::object();
    p.set("SrcID", Json::of((int64_t)SrcID));
    p.set("LayerID", Json::of((int64_t)Layer));
    p.set("DestID", Json::of((int64_t)DestID));
    p.set("DataSize", Json::of(DataSize));
    p.set("Offset", Json::of(Offset));
    p.set("Rate", Json::of(Rate));
    return p;
}

std::string LayerMsg::String() const {
    return "layer " + std::to_string(Layer) + " from " + std::to_string(SrcID) + " (" + std::to_string(Src_.DataSize) +
           "/" + std::to_string(TotalSize) + " B)";
}
Json LayerMsg::Payload() const {
    Json p = Json::object();
    p.set("SrcID", Json::of((int64_t)SrcID));
    p.set("LayerID", Json::of((int64_t)Layer));
    p.set("LayerSize", Json::of(Src_.DataSize));
    p.set("TotalSize", Json::of(TotalSize));
    p.set("Offert", Json::of(Src_.Offset));       // sic: the reference's field name, transport.go:52
    return p;
}

std::string StartupMsg::String() const { return "startup from " + std::to_string(SrcID); }
Json StartupMsg::Payload() const {
    Json p = Json::object();
    p.set("SrcID", Json::of((int64_t)SrcID));
    return p;
}

Json SimpleMsg::Payload() const {
    Json p = Json::object();
    p.set("SrcAddr", Json::of(SrcAddr));
    p.set("PayloadStr", Json::of(PayloadStr));
    return p;
}

std::string EncodeMsg(const Message &m) {
    Json env = Json::object();
    env.set("type", Json::of((int64_t)m.Type()));
    env.set("src", Json::of(m.Src()));
    env.set("payload", m.Payload());
    return env.dump();
}

static int64_t geti(const Json &p, const char *k) {
    const Json *v = p.get(k);
    return v ? v->as_int() : 0;
}

MessagePtr DecodeMsg(const std::string &wire, size_t &pos) {
    Json env = Json::parse(wire, pos);
    const Json *type = env.get("type"), *payload = env.get("payload");
    if (!type || !payload || !payload->is_object()) throw JsonError("envelope lacks type/payload");
    const Json &p = *payload;
    switch ((MsgType)type->as_int()) {
        case MsgTypeAnnounce: {
            auto m = std::make_shared<AnnounceMsg>();
            m->SrcID = (NodeID)geti(p, "SrcID");
            if (const Json *ids = p.get("LayerIDs"))
                for (auto &kv : ids->obj) {
                    LayerMeta meta;
                    meta.Location = (LayerLocation)geti(kv.second, "Location");
                    meta.LimitRate = geti(kv.second, "LimitRate");
                    meta.Source = (SourceType)geti(kv.second, "SourceType");
                    const Json *ds = kv.second.get("DataSize");
                    meta.DataSize = ds ? ds->as_int() : -1;
                    m->Layers[(LayerID)strtoul(kv.first.c_str(), nullptr, 10)] = meta;
                }
            return m;
        }
        case MsgTypeAck: {
            auto m = std::make_shared<AckMsg>();
            m->SrcID = (NodeID)geti(p, "SrcID");
            m->Layer = (LayerID)geti(p, "LayerID");
            m->location = InmemLayer;                      // absent on the wire -> zero value
            return m;
        }
        case MsgTypeRetransmit: {
            auto m = std::make_shared<RetransmitMsg>();
            m->SrcID = (NodeID)geti(p, "SrcID");
            m->Layer = (LayerID)geti(p, "LayerID");
            m->DestID = (NodeID)geti(p, "DestID");
            return m;
        }
        case MsgTypeFlowRetransmit: {
            auto m = std::make_shared<FlowRetransmitMsg>();
            m->SrcID = (NodeID)geti(p, "SrcID");
            m->Layer = (LayerID)geti(p, "LayerID");
            m->DestID = (NodeID)geti(p, "DestID");
            m->DataSize = geti(p, "DataSize");
            m->Offset = geti(p, "Offset");
            m->Rate = geti(p, "Rate");
            return m;
        }
        case MsgTypeLayer: {                               // header only; Offset is dropped at the receiver
            auto m = std::make_shared<LayerMsg>();         // in the reference (transport.go:221) — kept here
            m->SrcID = (NodeID)geti(p, "SrcID");
            m->Layer = (LayerID)geti(p, "LayerID");
            m->Src_.InDevice = true;
            m->Src_.DataSize = geti(p, "LayerSize");
            m->Src_.Offset = geti(p, "Offert");
            m->Src_.Meta.Location = InmemLayer;
            m->TotalSize = geti(p, "TotalSize");
            return m;
        }
        case MsgTypeStartup: {
            auto m = std::make_shared<StartupMsg>();
            m->SrcID = (NodeID)geti(p, "SrcID");
            return m;
        }
        case MsgTypeSimple: {
            auto m = std::make_shared<SimpleMsg>();
            if (const Json *v = p.get("SrcAddr")) m->SrcAddr = v->s;
            if (const Json *v = p.get("PayloadStr")) m->PayloadStr = v->s;
            return m;
        }
        default: throw JsonError("unknown message type " + std::to_string(type->as_int()));
    }
}

MessagePtr DecodeMsg(const std::string &wire) {
    size_t pos = 0;
    return DecodeMsg(wire, pos);
}

}  // namespace host
}  // namespace dissem
