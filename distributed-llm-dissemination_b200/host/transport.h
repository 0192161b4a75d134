// The drop-in boundary: distributor.Transport (distributor/transport.go:18-25) and its device
// implementation.  Control messages travel through in-process queues exactly like the reference's
// InmemoryTransport (transport.go:540-572) — but through the JSON envelope, so what a receiver sees is what
// TCP would have delivered; layerMsg bodies move on the GPU through the C ABI of include/dissem.h.
#pragma once
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/dissem.h"
#include "message.h"
#include "types.h"

namespace dissem {
namespace host {

// A Go channel of Messages: Deliver() returns one (transport.go:22, :467-469).
class MsgChan {
public:
    void Push(MessagePtr m);
    bool Pop(MessagePtr *out);   // blocks; false once closed and drained
    void Close();

private:
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<MessagePtr> q_;
    bool closed_ = false;
};

class Transport {                // transport.go:18-25
public:
    virtual ~Transport() = default;
    virtual int Send(NodeID dest, MessagePtr message) = 0;            // 0 or a DISSEM_E* code
    virtual int RegisterPipe(LayerID layer, NodeID dest) = 0;
    virtual int Broadcast(MessagePtr message) = 0;
    virtual MsgChan &Deliver() = 0;
    virtual std::string GetAddress() const = 0;
    virtual int Close() = 0;
};

class DeviceTransport;

// The process-wide registry the transports share (the role of InmemoryTransport's global `registry`,
// transport.go:507-511) plus the device context that holds every node's layer store.
class DeviceFabric {
public:
    explicit DeviceFabric(dissem_ctx *ctx) : ctx_(ctx) {}
    dissem_ctx *ctx() const { return ctx_; }
    void Register(NodeID id, DeviceTransport *t);
    void Unregister(NodeID id);
    DeviceTransport *Lookup(NodeID id);
    std::vector<NodeID> Ids();
    // LimitRate pacing (pace.h): off by default — NVLink is not the network the reference throttles for; switched on
    // to reproduce the reference's timing behaviour (`"pace":1` / --pace).
    bool pace = false;
    // counters for the run report
    std::mutex stat_mu;
    int64_t layer_msgs = 0, layer_bytes = 0, control_msgs = 0, launches = 0;
    double device_ms = 0.0;

private:
    dissem_ctx *ctx_;
    std::mutex mu_;
    std::map<NodeID, DeviceTransport *> peers_;
};

class DeviceTransport : public Transport {
public:
    DeviceTransport(DeviceFabric *fabric, NodeID me);
    ~DeviceTransport() override;
    // layerMsg: allocate the destination buffer if absent (make(LayerData, LayerSize), transport.go:201), push the
    // bytes with one kernel launch, wait, then deliver the header to the destination's channel.
    // anything else: envelope encode -> decode -> destination's channel.
    int Send(NodeID dest, MessagePtr message) override;
    int RegisterPipe(LayerID, NodeID) override;      // client tee-forwarding is out of scope: DISSEM_ENOTSUP
    int Broadcast(MessagePtr message) override;      // transport.go:290-306
    MsgChan &Deliver() override { return inbox_; }
    std::string GetAddress() const override { return "gpu:" + std::to_string(me_); }
    int Close() override;
    // Extension for plans compiled into one device schedule: sends every layerMsg of `xfers` whose source is any
    // local node with ONE dissem_schedule_submit, then delivers the headers.  device_ms_out may be NULL.
    int SendBatch(const std::vector<PlannedXfer> &xfers, double *device_ms_out);

private:
    int EnsureDestLayer(NodeID dest, LayerID layer, int64_t total);
    DeviceFabric *fabric_;
    NodeID me_;
    MsgChan inbox_;
    bool closed_ = false;
};

}  // namespace host
}  // namespace dissem
