// Control-plane vocabulary: the eight message kinds of distributor/message.go and the JSON envelope they
// travel in (message.go:273-301; wire format in SURVEY Appendix A).  layerMsg is never encoded whole: only
// its header `tempLayerInfo` is (transport.go:47-54, :308-331); the bytes move on the device.
#pragma once
#include <memory>
#include <string>

#include "json.h"
#include "types.h"

namespace dissem {
namespace host {

enum MsgType : uint8_t {   // message.go:18-28
    MsgTypeAnnounce = 0, MsgTypeAck, MsgTypeLayer, MsgTypeRetransmit, MsgTypeFlowRetransmit,
    MsgTypeClientReq, MsgTypeStartup, MsgTypeSimple, MsgTypeTransport
};

struct Message {           // message.go:8-13
    virtual ~Message() = default;
    virtual std::string Src() const = 0;
    virtual MsgType Type() const = 0;
    virtual std::string String() const = 0;
    virtual Json Payload() const = 0;    // the exported fields, under their Go names
};
using MessagePtr = std::shared_ptr<Message>;

struct AnnounceMsg : Message {           // message.go:31-60
    NodeID SrcID = 0;
    LayerIDs Layers;
    std::string Src() const override { return std::to_string(SrcID); }
    MsgType Type() const override { return MsgTypeAnnounce; }
    std::string String() const override;
    Json Payload() const override;
};
struct AckMsg : Message {                // message.go:62-92; `location` is unexported: not serialised
    NodeID SrcID = 0;
    LayerID Layer = 0;
    LayerLocation location = InmemLayer;
    std::string Src() const override { return std::to_string(SrcID); }
    MsgType Type() const override { return MsgTypeAck; }
    std::string String() const override;
    Json Payload() const override;
};
struct RetransmitMsg : Message {         // message.go:94-119
    NodeID SrcID = 0;
    LayerID Layer = 0;
    NodeID DestID = 0;
    std::string Src() const override { return std::to_string(SrcID); }
    MsgType Type() const override { return MsgTypeRetransmit; }
    std::string String() const override;
    Json Payload() const override;
};
struct FlowRetransmitMsg : Message {     // message.go:121-152
    NodeID SrcID = 0;
    LayerID Layer = 0;
    NodeID DestID = 0;
    int64_t DataSize = 0, Offset = 0, Rate = 0;
    std::string Src() const override { return std::to_string(SrcID); }
    MsgType Type() const override { return MsgTypeFlowRetransmit; }
    std::string String() const override;
    Json Payload() const override;
};
struct LayerMsg : Message {              // message.go:154-191
    NodeID SrcID = 0;
    LayerID Layer = 0;
    LayerSrc Src_;                       // DataSize = bytes in this message, Offset = where they start
    int64_t TotalSize = 0;
    std::string Src() const override { return std::to_string(SrcID); }
    MsgType Type() const override { return MsgTypeLayer; }
    std::string String() const override;
    Json Payload() const override;       // tempLayerInfo: SrcID, LayerID, LayerSize, TotalSize, Offert
};
struct StartupMsg : Message {            // message.go:217-242
    NodeID SrcID = 0;
    std::string Src() const override { return std::to_string(SrcID); }
    MsgType Type() const override { return MsgTypeStartup; }
    std::string String() const override;
    Json Payload() const override;
};
struct SimpleMsg : Message {             // message.go:244-271 (tests only)
    std::string SrcAddr, PayloadStr;
    std::string Src() const override { return SrcAddr; }
    MsgType Type() const override { return MsgTypeSimple; }
    std::string String() const override { return PayloadStr; }
    Json Payload() const override;
};

// {"type":T,"src":"<id>","payload":{...}}   (TransportMsg, message.go:273-277)
std::string EncodeMsg(const Message &m);
// decodeMsg, message.go:280-301.  `pos` advances past one envelope, so concatenated envelopes can be read in turn.
MessagePtr DecodeMsg(const std::string &wire, size_t &pos);
MessagePtr DecodeMsg(const std::string &wire);

}  // namespace host
}  // namespace dissem
