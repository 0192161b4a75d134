#include "transport.h"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <thread>

#include "../csrc/dissem_engine.h"
#include "pace.h"

namespace dissem {
namespace host {

void MsgChan::Push(MessagePtr m) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (closed_) return;
        q_.push_back(std::move(m));
    }
    cv_.notify_one();
}

bool MsgChan::Pop(MessagePtr *out) {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return closed_ || !q_.empty(); });
    if (q_.empty()) return false;
    *out = std::move(q_.front());
    q_.pop_front();
    return true;
}

void MsgChan::Close() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        closed_ = true;
    }
    cv_.notify_all();
}

void DeviceFabric::Register(NodeID id, DeviceTransport *t) {
    std::lock_guard<std::mutex> lk(mu_);
    peers_[id] = t;
}
void DeviceFabric::Unregister(NodeID id) {
    std::lock_guard<std::mutex> lk(mu_);
    peers_.erase(id);
}
DeviceTransport *DeviceFabric::Lookup(NodeID id) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = peers_.find(id);
    return it == peers_.end() ? nullptr : it->second;
}
std::vector<NodeID> DeviceFabric::Ids() {
    std::lock_guard<std::mutex> lk(mu_);
    std::vector<NodeID> v;
    for (auto &kv : peers_) v.push_back(kv.first);
    return v;
}

DeviceTransport::DeviceTransport(DeviceFabric *fabric, NodeID me) : fabric_(fabric), me_(me) {
    fabric_->Register(me_, this);
}
DeviceTransport::~DeviceTransport() { Close(); }

int DeviceTransport::Close() {
    if (!closed_) {
        closed_ = true;
        fabric_->Unregister(me_);
        inbox_.Close();
    }
    return 0;
}

int DeviceTransport::RegisterPipe(LayerID, NodeID) {
    return set_error(DISSEM_ENOTSUP, "RegisterPipe: external-client tee forwarding is not part of the device transport");
}

int DeviceTransport::Broadcast(MessagePtr message) {
    for (NodeID id : fabric_->Ids()) Send(id, message);      // errors are logged and dropped, transport.go:298-303
    return 0;
}

int DeviceTransport::EnsureDestLayer(NodeID dest, LayerID layer, int64_t total) {
    int rc = dissem_layer_info(fabric_->ctx(), (int)dest, (int)layer, nullptr, nullptr);
    if (rc == DISSEM_ENOENT) {
        rc = dissem_layer_alloc(fabric_->ctx(), (int)dest, (int)layer, total);
        if (rc == DISSEM_EEXIST) rc = 0;          // another sender of the same layer got there first
    }
    return rc;
}

int DeviceTransport::Send(NodeID dest, MessagePtr message) {
    DeviceTransport *peer = fabric_->Lookup(dest);
    if (!peer) return set_error(DISSEM_ENOENT, "addr of %u does not exist", dest);     // transport.go:263-265
    if (auto *lm = dynamic_cast<LayerMsg *>(message.get())) {
        const bool in_device = lm->Src_.InDevice && lm->Src_.Meta.Location == InmemLayer;      // transport.go:333
        const bool on_disk = lm->Src_.Meta.Location == DiskLayer;                              // transport.go:351
        if (!in_device && !on_disk) return set_error(DISSEM_EINVAL, "unknown error sending layer %u", lm->Layer);   // transport.go:368-370
        if (on_disk && lm->Src_.Fp.empty()) return set_error(DISSEM_EINVAL, "no data source specified");         // transport.go:352-354
        int rc = EnsureDestLayer(dest, lm->Layer, lm->TotalSize);
        if (rc) return rc;
        dissem_ticket ticket = 0;
        if (on_disk) {
            // the section reader over the layer file (transport.go:356-364): the engine streams the file through pinned
            // staging buffers into the destination.  The file was attached when the node's layers were created.
            rc = dissem_send_file(fabric_->ctx(), (int)lm->SrcID, (int)dest, (int)lm->Layer, lm->Src_.Offset, lm->Src_.DataSize,
                                  lm->TotalSize, &ticket);
        } else if (fabric_->pace && lm->Src_.Meta.LimitRate > 0) {
            // writeWithLimit (transport.go:407-424): pieces are launched no earlier than the reference's token bucket
            // would have let their last byte go; the file branch above is not limited in the reference either (:364)
            const auto t0 = std::chrono::steady_clock::now();
            std::vector<dissem_ticket> in_flight;
            for (const PacePiece &pc : PaceSchedule(lm->Src_.Offset, lm->Src_.DataSize, lm->Src_.Meta.LimitRate)) {
                std::this_thread::sleep_until(t0 + std::chrono::nanoseconds(pc.release_ns));
                dissem_ticket tk = 0;
                rc = dissem_send(fabric_->ctx(), (int)lm->SrcID, (int)dest, (int)lm->Layer, pc.off, pc.size, lm->TotalSize, &tk);
                if (rc) break;
                in_flight.push_back(tk);
            }
            double paced_ms = 0.0;
            for (dissem_ticket tk : in_flight) {                 // every ticket is waited for, also after an error
                double one = 0.0;
                int rcw = dissem_wait(fabric_->ctx(), tk, &one);
                if (!rc) rc = rcw;
                paced_ms += one;
            }
            if (rc) return rc;
            {
                std::lock_guard<std::mutex> lk(fabric_->stat_mu);
                fabric_->device_ms += paced_ms;
                fabric_->launches += 2 * (int64_t)in_flight.size();
            }
        } else {
            rc = dissem_send(fabric_->ctx(), (int)lm->SrcID, (int)dest, (int)lm->Layer, lm->Src_.Offset, lm->Src_.DataSize,
                             lm->TotalSize, &ticket);
        }
        if (rc) return rc;
        double ms = 0.0;
        if (ticket) rc = dissem_wait(fabric_->ctx(), ticket, &ms);          // Send returns when the bytes are written
        if (rc) return rc;
        {
            std::lock_guard<std::mutex> lk(fabric_->stat_mu);
            fabric_->layer_msgs++;
            fabric_->layer_bytes += lm->Src_.DataSize;
            fabric_->device_ms += ms;
            fabric_->launches += (ticket && lm->Src_.DataSize > 0) ? 2 : 0;
        }
        // what handleIncomingMsg hands to the node (transport.go:221-222): header fields + an in-memory LayerSrc
        auto delivered = std::dynamic_pointer_cast<LayerMsg>(DecodeMsg(EncodeMsg(*lm)));
        peer->inbox_.Push(delivered);
        return 0;
    }
    {
        std::lock_guard<std::mutex> lk(fabric_->stat_mu);
        fabric_->control_msgs++;
    }
    peer->inbox_.Push(DecodeMsg(EncodeMsg(*message)));
    return 0;
}

int DeviceTransport::SendBatch(const std::vector<PlannedXfer> &xfers, double *device_ms_out) {
    std::vector<dissem_xfer> xs;
    // which sources are disk copies is settled before any destination buffer is created (a node may load its own file)
    std::vector<int> from_file;
    for (auto &x : xfers)
        from_file.push_back(dissem_layer_info(fabric_->ctx(), (int)x.Src, (int)x.Layer, nullptr, nullptr) == DISSEM_ENOENT &&
                            dissem_layer_file_info(fabric_->ctx(), (int)x.Src, (int)x.Layer, nullptr) == DISSEM_OK);
    // who holds which of the plan's layers in HBM right now (before destination buffers exist): lets the plan take a layer
    // from a node that shares the destination's GPU instead of across NVLink (dissem_plan_localize)
    const int n_nodes_all = dissem_n_nodes(fabric_->ctx());
    std::vector<int> holder_nodes, holder_layers;
    {
        std::vector<int> layers_seen;
        for (auto &x : xfers)
            if (std::find(layers_seen.begin(), layers_seen.end(), (int)x.Layer) == layers_seen.end()) layers_seen.push_back((int)x.Layer);
        for (int l : layers_seen)
            for (int nd = 0; nd < n_nodes_all; ++nd) {
                int64_t sz = 0;
                if (dissem_layer_info(fabric_->ctx(), nd, l, &sz, nullptr) == DISSEM_OK) {
                    holder_nodes.push_back(nd);
                    holder_layers.push_back(l);
                }
            }
    }
    for (auto &x : xfers) {
        int rc = EnsureDestLayer(x.Dest, x.Layer, x.TotalSize);
        if (rc) return rc;
        dissem_xfer d;
        d.src_node = (int32_t)x.Src;
        d.dst_node = (int32_t)x.Dest;
        d.layer = (int32_t)x.Layer;
        // a source that holds the layer on disk only (DiskLayer) is read from its file
        d.flags = from_file[xs.size()] ? DISSEM_XFER_FILE : 0;
        d.off = x.Offset;
        d.size = x.DataSize;
        d.total = x.TotalSize;
        xs.push_back(d);
    }
    // let the engine choose relay chains and push/pull directions for this placement of nodes on GPUs
    const int n_nodes = dissem_n_nodes(fabric_->ctx());
    std::vector<int> gpu_of_node((size_t)std::max(n_nodes, 0), 0);
    for (int i = 0; i < n_nodes; ++i) {
        int rcd = dissem_node_device(fabric_->ctx(), i, &gpu_of_node[i]);
        if (rcd) return rcd;
        if (gpu_of_node[i] < 0) gpu_of_node[i] = 1000 + i;            // a node of another process: its own GPU
    }
    dissem_xfer *opt = nullptr;
    int n_opt = 0;
    int rc = 0;
    dissem_ticket ticket = 0;
    // relay chains, push/pull directions, and for disk sources "read the file once, forward from HBM"
    {
        int m = (int)xs.size();
        rc = dissem_plan_localize(xs.data(), m, gpu_of_node.data(), n_nodes, holder_nodes.data(), holder_layers.data(),
                                  (int)holder_nodes.size(), &m);
        if (rc) return rc;
        xs.resize((size_t)m);
    }
    rc = dissem_plan_optimize(xs.data(), (int)xs.size(), gpu_of_node.data(), n_nodes, &opt, &n_opt, nullptr);
    if (rc) return rc;
    rc = dissem_schedule_submit(fabric_->ctx(), opt, n_opt, &ticket);
    free(opt);
    if (rc) return rc;
    double ms = 0.0;
    rc = dissem_wait(fabric_->ctx(), ticket, &ms);
    if (rc) return rc;
    if (device_ms_out) *device_ms_out = ms;
    {
        std::lock_guard<std::mutex> lk(fabric_->stat_mu);
        fabric_->device_ms += ms;
        fabric_->launches += 2;
    }
    for (auto &x : xfers) {
        DeviceTransport *peer = fabric_->Lookup(x.Dest);
        if (!peer) return set_error(DISSEM_ENOENT, "addr of %u does not exist", x.Dest);
        LayerMsg lm;
        lm.SrcID = x.Src;
        lm.Layer = x.Layer;
        lm.Src_.InDevice = true;
        lm.Src_.DataSize = x.DataSize;
        lm.Src_.Offset = x.Offset;
        lm.TotalSize = x.TotalSize;
        {
            std::lock_guard<std::mutex> lk(fabric_->stat_mu);
            fabric_->layer_msgs++;
            fabric_->layer_bytes += x.DataSize;
        }
        peer->inbox_.Push(DecodeMsg(EncodeMsg(lm)));
    }
    return 0;
}

}  // namespace host
}  // namespace dissem
