// Role state machines of the reference (distributor/node.go), kept thin: announce-gather barrier, plan,
// issue transfers / retransmit orders, store + ack, completion, startup.  One dispatcher thread per node
// consumes Transport.Deliver() (node.go:271-286, :1328-1342); sends that block run on worker threads where
// the reference spawns goroutines.
#pragma once
#include <atomic>
#include <chrono>
#include <functional>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "plan.h"
#include "transport.h"
#include "types.h"

namespace dissem {
namespace host {

// node / N, node.go:16-126
class N {
public:
    N(NodeID my_id, NodeID leader_id, Transport *t);         // NewNode, node.go:47-63
    NodeID GetMyID() const { return my_id_; }
    NodeID getLeader();
    bool getNextHop(NodeID goal, NodeID *next);
    void addNode(NodeID goal) { addRoutingTable(goal, goal, 1); }   // node.go:93-96
    void addRoutingTable(NodeID goal, NodeID next_hop, unsigned remaining_hops);
    bool updateLeader(NodeID leader);
    Transport *GetTransport() const { return t_; }

private:
    struct RoutingInfo { NodeID nextHop; unsigned remainingHops; };
    NodeID my_id_, leader_id_;
    Transport *t_;
    std::map<NodeID, RoutingInfo> routing_;
    std::mutex mu_;
};

// A Go `chan T` that is written once and read once (startDistributionChan / readyChan, node.go:233-236).
template <class T>
class OneShot {
public:
    void Fire(const T &v) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            if (fired_) return;
            fired_ = true;
            value_ = v;
            when_ = std::chrono::steady_clock::now();
        }
        cv_.notify_all();
    }
    bool Wait(int timeout_ms, T *out = nullptr) {
        std::unique_lock<std::mutex> lk(mu_);
        if (!cv_.wait_for(lk, std::chrono::milliseconds(timeout_ms), [&] { return fired_; })) return false;
        if (out) *out = value_;
        return true;
    }
    bool fired() { std::lock_guard<std::mutex> lk(mu_); return fired_; }
    std::chrono::steady_clock::time_point when() { std::lock_guard<std::mutex> lk(mu_); return when_; }

private:
    std::mutex mu_;
    std::condition_variable cv_;
    bool fired_ = false;
    T value_{};
    std::chrono::steady_clock::time_point when_;
};

class RoleBase {
public:
    RoleBase(N *node, LayersSrc layers);
    virtual ~RoleBase();
    void Stop();                      // closes the transport, joins the dispatcher and the workers
    LayersSrc Layers();
    std::string LastError();

protected:
    void Start();                     // spawn the dispatcher (the constructors' handleIncomingMsg())
    virtual void Dispatch(const MessagePtr &m) = 0;
    void Go(std::function<void()> fn);   // `go fn()`
    void Fail(const std::string &what);
    void StoreLayer(const LayerMsg &m, bool *complete);   // handleLayerMsg bodies, node.go:376-407, :1354-1384, :1520-1567
    int SendLayerTo(NodeID dest, LayerID layer, const LayerSrc &ls, int64_t total);

    N *node_;
    LayersSrc layers_;
    std::mutex mu_;
    std::string err_;

private:
    std::thread dispatcher_;
    std::mutex workers_mu_;
    std::vector<std::thread> workers_;
    bool stopped_ = false;
};

// Leader interface, node.go:214-226
class Leader {
public:
    virtual ~Leader() = default;
    virtual OneShot<Assignment> &StartDistribution() = 0;
    virtual OneShot<Assignment> &Ready() = 0;
};

// mode 0: LeaderNode, node.go:228-469
class LeaderNode : public RoleBase, public Leader {
public:
    LeaderNode(N *node, LayersSrc layers, Assignment assignment, bool autostart = true);   // NewLeaderNode, node.go:262
    OneShot<Assignment> &StartDistribution() override { return start_; }
    OneShot<Assignment> &Ready() override { return ready_; }
    void SetBatched(bool b) { batched_ = b; }   // hand the whole plan to the device as one schedule
    Plan LastPlan();

protected:
    void Dispatch(const MessagePtr &m) override;
    void handleAnnounceMsg(const AnnounceMsg &m);          // node.go:295-324
    virtual Plan makePlan();                               // mode 0: node.go:326-352
    virtual void sendLayers();                             // issues the plan
    virtual void handleAckMsg(const AckMsg &m);            // node.go:410-432
    void handleLayerMsg(const LayerMsg &m);                // node.go:376-407
    bool assignmentSatisfied();                            // node.go:435-446
    void sendStartup();                                    // node.go:456-469
    void issue(const Plan &p);

    Assignment assignment_;
    Status status_;
    OneShot<Assignment> start_, ready_;
    bool batched_ = false;
    bool started_ = false;
    Plan plan_;
};

// mode 1: RetransmitLeaderNode, node.go:471-626
class RetransmitLeaderNode : public LeaderNode {
public:
    RetransmitLeaderNode(N *node, LayersSrc layers, Assignment assignment, PlanOptions opt, bool autostart = true);

protected:
    Plan makePlan() override;
    void sendLayers() override;       // leader-owned layers are sent by the leader, others by retransmitMsg (:611-626)
    PlanOptions opt_;
};

// mode 2: PullRetransmitLeaderNode, node.go:628-1073 — pull-based work conservation: every sender has at most one
// job in flight; when the destination acks it, the sender is handed its next own job (rarest layer first) or steals a
// pending job of the sender expected to finish last.
class PullRetransmitLeaderNode : public RetransmitLeaderNode {
public:
    PullRetransmitLeaderNode(N *node, LayersSrc layers, Assignment assignment);   // NewPullRetransmitLeaderNode, node.go:674

protected:
    void sendLayers() override;                            // node.go:810-901
    void handleAckMsg(const AckMsg &m) override;           // node.go:741-807
    int assignNewJob(NodeID node);                         // node.go:906-945

private:
    enum JobStatus { Pending = 0, SendingRetransmit = 1 };                        // node.go:631-638
    struct JobInfo {                                       // node.go:640-645
        NodeID sender = 0;
        JobStatus status = Pending;
        std::chrono::steady_clock::time_point retransmitTime;
    };
    struct Perf { int64_t aveNs = 0; unsigned completed = 0; };                   // nodePerformance, node.go:657-660
    bool getRarestOwnJob(NodeID node, LayerID *layer, NodeID *dest);              // node.go:980-1010
    bool getRarestStealableJob(NodeID node, LayerID *layer, NodeID *dest, NodeID *from);   // node.go:1012-1073
    NodeID getMinLoadedSender(LayerID layer);              // node.go:948-978
    int sendRetransmit(LayerID layer, NodeID owner, NodeID dest);                 // node.go:611-626
    std::map<LayerID, std::map<NodeID, JobInfo>> jobs_;    // jobsInfoMap
    std::map<NodeID, unsigned> load_;                      // senderLoadCounter
    std::map<NodeID, Perf> perf_;
    std::map<LayerID, NodeIDs> owners_;                    // layerOwners
    std::map<NodeID, bool> completion_;                    // nodeCompletionStatus
};

// mode 3: FlowRetransmitLeaderNode, node.go:1075-1288
class FlowRetransmitLeaderNode : public RetransmitLeaderNode {
public:
    FlowRetransmitLeaderNode(N *node, LayersSrc layers, Assignment assignment, std::map<NodeID, int64_t> network_bw,
                             PlanOptions opt);

protected:
    void Dispatch(const MessagePtr &m) override;
    Plan makePlan() override;
    void sendLayers() override;       // flowRetransmitMsg per fragment to its sender (:1237-1288)
    std::map<NodeID, int64_t> bw_;
};

// Receiver interface, node.go:1291-1297
class Receiver {
public:
    virtual ~Receiver() = default;
    virtual int Announce() = 0;
    virtual OneShot<int> &Ready() = 0;
};

// mode 0: ReceiverNode, node.go:1299-1418
class ReceiverNode : public RoleBase, public Receiver {
public:
    ReceiverNode(N *node, LayersSrc layers, bool autostart = true);     // NewReceiverNode, node.go:1319
    int Announce() override;                                             // node.go:1392-1415
    OneShot<int> &Ready() override { return ready_; }

protected:
    void Dispatch(const MessagePtr &m) override;
    void handleLayerMsg(const LayerMsg &m);                              // node.go:1354-1384
    OneShot<int> ready_;
};

// modes 1/2: RetransmitReceiverNode, node.go:1420-1484
class RetransmitReceiverNode : public ReceiverNode {
public:
    RetransmitReceiverNode(N *node, LayersSrc layers, bool autostart = true);

protected:
    void Dispatch(const MessagePtr &m) override;
    void handleRetransmitMsg(const RetransmitMsg &m);                    // node.go:1462-1484
};

// mode 3: FlowRetransmitReceiverNode, node.go:1486-1643
class FlowRetransmitReceiverNode : public RetransmitReceiverNode {
public:
    FlowRetransmitReceiverNode(N *node, LayersSrc layers);

protected:
    void Dispatch(const MessagePtr &m) override;
};

// handleFlowRetransmit, node.go:1592-1643 (shared by leader and receivers)
int HandleFlowRetransmit(N *node, const LayersSrc &layers, const FlowRetransmitMsg &m);

}  // namespace host
}  // namespace dissem
