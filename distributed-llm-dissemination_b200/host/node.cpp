#include "node.h"

#include <algorithm>
#include <cstdint>
#include <functional>

#include "../csrc/dissem_engine.h"

namespace dissem {
namespace host {

// ---------------------------------------------------------------------------------------------- N
N::N(NodeID my_id, NodeID leader_id, Transport *t) : my_id_(my_id), leader_id_(leader_id), t_(t) {
    if (my_id != leader_id) addNode(leader_id);          // node.go:58-60
}
NodeID N::getLeader() {
    std::lock_guard<std::mutex> lk(mu_);
    return leader_id_;
}
bool N::getNextHop(NodeID goal, NodeID *next) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = routing_.find(goal);
    if (it == routing_.end()) return false;              // "routing entry ... does not exist", node.go:85-87
    *next = it->second.nextHop;
    return true;
}
void N::addRoutingTable(NodeID goal, NodeID next_hop, unsigned remaining_hops) {
    std::lock_guard<std::mutex> lk(mu_);
    routing_[goal] = RoutingInfo{next_hop, remaining_hops};
}
bool N::updateLeader(NodeID leader) {
    std::lock_guard<std::mutex> lk(mu_);
    if (!routing_.count(leader)) return false;           // node.go:115-118
    leader_id_ = leader;
    return true;
}

// --------------------------------------------------------------------------------------- RoleBase
RoleBase::RoleBase(N *node, LayersSrc layers) : node_(node), layers_(std::move(layers)) {}
RoleBase::~RoleBase() { Stop(); }

void RoleBase::Start() {
    dispatcher_ = std::thread([this] {
        MessagePtr m;
        while (node_->GetTransport()->Deliver().Pop(&m)) Dispatch(m);
    });
}

void RoleBase::Stop() {
    {
        std::lock_guard<std::mutex> lk(workers_mu_);
        if (stopped_) return;
        stopped_ = true;
    }
    node_->GetTransport()->Close();
    if (dispatcher_.joinable()) dispatcher_.join();
    std::vector<std::thread> ws;
    {
        std::lock_guard<std::mutex> lk(workers_mu_);
        ws.swap(workers_);
    }
    for (auto &w : ws)
        if (w.joinable()) w.join();
}

void RoleBase::Go(std::function<void()> fn) {
    std::lock_guard<std::mutex> lk(workers_mu_);
    if (stopped_) return;
    workers_.emplace_back(std::move(fn));
}

void RoleBase::Fail(const std::string &what) {
    std::lock_guard<std::mutex> lk(mu_);
    if (err_.empty()) err_ = what;
}
std::string RoleBase::LastError() {
    std::lock_guard<std::mutex> lk(mu_);
    return err_;
}
LayersSrc RoleBase::Layers() {
    std::lock_guard<std::mutex> lk(mu_);
    return layers_;
}

void RoleBase::StoreLayer(const LayerMsg &m, bool *complete) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = layers_.find(m.Layer);
    if (it == layers_.end()) {
        LayerSrc ls;
        ls.InDevice = true;
        ls.DataSize = 0;
        ls.Meta.Location = InmemLayer;
        it = layers_.emplace(m.Layer, ls).first;
    }
    LayerSrc &ls = it->second;
    if (ls.Meta.Location == DiskLayer) {       // the delivery replaces a disk copy with the in-memory one (node.go:1372)
        ls.DataSize = 0;
        ls.Fp.clear();
    }
    if (ls.DataSize < m.TotalSize) ls.DataSize += m.Src_.DataSize;       // node.go:1540-1543
    ls.InDevice = true;
    ls.Meta.Location = InmemLayer;
    ls.Meta.DataSize = m.TotalSize;
    *complete = ls.DataSize == m.TotalSize;                              // node.go:1555
}

int RoleBase::SendLayerTo(NodeID dest, LayerID layer, const LayerSrc &ls, int64_t total) {
    auto lm = std::make_shared<LayerMsg>();                              // NewLayerMsg, message.go:165
    lm->SrcID = node_->GetMyID();
    lm->Layer = layer;
    lm->Src_ = ls;
    lm->TotalSize = total;
    int rc = node_->GetTransport()->Send(dest, lm);
    if (rc) Fail("couldn't send layer " + std::to_string(layer) + " to " + std::to_string(dest) + ": " + last_error());
    return rc;
}

// ------------------------------------------------------------------------------------- LeaderNode
LeaderNode::LeaderNode(N *node, LayersSrc layers, Assignment assignment, bool autostart)
    : RoleBase(node, std::move(layers)), assignment_(std::move(assignment)) {
    LayerIDs mine;                                       // newLeaderNodeBase, node.go:251-257
    for (auto &kv : layers_) {
        LayerMeta m = kv.second.Meta;
        m.DataSize = kv.second.DataSize;
        mine[kv.first] = m;
    }
    status_[node_->GetMyID()] = std::move(mine);
    if (autostart) Start();
}

Plan LeaderNode::LastPlan() {
    std::lock_guard<std::mutex> lk(mu_);
    return plan_;
}

void LeaderNode::Dispatch(const MessagePtr &m) {
    if (auto *a = dynamic_cast<AnnounceMsg *>(m.get())) handleAnnounceMsg(*a);
    else if (auto *k = dynamic_cast<AckMsg *>(m.get())) handleAckMsg(*k);
    else if (auto *l = dynamic_cast<LayerMsg *>(m.get())) handleLayerMsg(*l);
}

void LeaderNode::handleAnnounceMsg(const AnnounceMsg &m) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (!status_.count(m.SrcID)) {
            status_[m.SrcID] = m.Layers;
            node_->addNode(m.SrcID);
        }
        for (auto &na : assignment_)                     // every node named by the assignment has announced?
            if (!status_.count(na.first)) return;
        if (started_) return;                            // single-shot (the reference can start twice, node.go:309-323)
        started_ = true;
    }
    start_.Fire(assignment_);
    Go([this] { sendLayers(); });
}

Plan LeaderNode::makePlan() {
    std::lock_guard<std::mutex> lk(mu_);
    return PlanMode0(node_->GetMyID(), layers_, status_, assignment_);
}

void LeaderNode::issue(const Plan &p) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        plan_ = p;
    }
    bool any = false;
    for (auto &x : p.Xfers) any = any || x.DataSize >= 0;
    if (!any) {
        // Nothing to send.  The reference reaches Ready only from handleAckMsg (node.go:421-428) and would
        // wait forever; an assignment that is already satisfied is reported ready at once.
        if (assignmentSatisfied()) {
            sendStartup();
            ready_.Fire(assignment_);
        }
    }
}

void LeaderNode::sendLayers() {
    Plan p;
    try {
        p = makePlan();
    } catch (const std::exception &e) {
        Fail(e.what());
        return;
    }
    issue(p);
    if (batched_) {
        auto *dt = dynamic_cast<DeviceTransport *>(node_->GetTransport());
        if (!dt) { Fail("batched plans need a DeviceTransport"); return; }
        if (!p.Xfers.empty() && dt->SendBatch(p.Xfers, nullptr)) Fail(std::string("batched send failed: ") + last_error());
        return;
    }
    for (auto &x : p.Xfers) {                            // go func(){ leader.sendLayer(...) }(), node.go:343-349
        LayerSrc ls;
        {
            std::lock_guard<std::mutex> lk(mu_);
            ls = layers_[x.Layer];
        }
        Go([this, x, ls] { SendLayerTo(x.Dest, x.Layer, ls, ls.DataSize); });
    }
}

void LeaderNode::handleLayerMsg(const LayerMsg &m) {
    bool complete = false;
    StoreLayer(m, &complete);
    if (!complete) return;
    auto ack = std::make_shared<AckMsg>();               // NewAckMsg, message.go:68
    ack->SrcID = node_->GetMyID();
    ack->Layer = m.Layer;
    if (node_->GetTransport()->Send(node_->getLeader(), ack)) Fail("failed to send ackMsg");
}

bool LeaderNode::assignmentSatisfied() {
    std::lock_guard<std::mutex> lk(mu_);
    for (auto &na : assignment_)
        for (auto &la : na.second) {
            auto sn = status_.find(na.first);
            if (sn == status_.end()) return false;
            auto sl = sn->second.find(la.first);
            if (sl == sn->second.end() || sl->second.Location != InmemLayer) return false;
        }
    return true;
}

void LeaderNode::handleAckMsg(const AckMsg &m) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        LayerMeta meta;
        meta.Location = m.location;
        status_[m.SrcID][m.Layer] = meta;
    }
    if (assignmentSatisfied() && !ready_.fired()) {
        sendStartup();                                   // "timer stop: startup", node.go:424
        ready_.Fire(assignment_);
    }
}

void LeaderNode::sendStartup() {
    std::vector<NodeID> ids;
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto &kv : status_) ids.push_back(kv.first);   // includes the leader itself, node.go:461-467
    }
    for (NodeID id : ids) {
        auto s = std::make_shared<StartupMsg>();
        s->SrcID = node_->GetMyID();
        node_->GetTransport()->Send(id, s);
    }
}

// --------------------------------------------------------------------------- RetransmitLeaderNode
RetransmitLeaderNode::RetransmitLeaderNode(N *node, LayersSrc layers, Assignment assignment, PlanOptions opt, bool autostart)
    : LeaderNode(node, std::move(layers), std::move(assignment), false), opt_(opt) {
    if (autostart) Start();
}

Plan RetransmitLeaderNode::makePlan() {
    std::lock_guard<std::mutex> lk(mu_);
    return PlanMode1(node_->GetMyID(), layers_, status_, assignment_, opt_);
}

void RetransmitLeaderNode::sendLayers() {
    Plan p;
    try {
        p = makePlan();
    } catch (const std::exception &e) {
        Fail(e.what());
        return;
    }
    issue(p);
    if (batched_) {
        auto *dt = dynamic_cast<DeviceTransport *>(node_->GetTransport());
        if (!dt) { Fail("batched plans need a DeviceTransport"); return; }
        if (!p.Xfers.empty() && dt->SendBatch(p.Xfers, nullptr)) Fail(std::string("batched send failed: ") + last_error());
        return;
    }
    for (auto &x : p.Xfers) {                            // sendRetransmit, node.go:611-626
        if (x.Src == node_->GetMyID()) {
            LayerSrc ls;
            {
                std::lock_guard<std::mutex> lk(mu_);
                ls = layers_[x.Layer];
            }
            SendLayerTo(x.Dest, x.Layer, ls, ls.DataSize);       // synchronous inside the loop, :614-621
        } else {
            auto r = std::make_shared<RetransmitMsg>();
            r->SrcID = node_->GetMyID();
            r->Layer = x.Layer;
            r->DestID = x.Dest;
            if (node_->GetTransport()->Send(x.Src, r)) Fail("couldn't send retransmit of " + std::to_string(x.Layer));
        }
    }
}

// ----------------------------------------------------------------------- PullRetransmitLeaderNode
PullRetransmitLeaderNode::PullRetransmitLeaderNode(N *node, LayersSrc layers, Assignment assignment)
    : RetransmitLeaderNode(node, std::move(layers), std::move(assignment), PlanOptions{}, false) {
    Start();
}

NodeID PullRetransmitLeaderNode::getMinLoadedSender(LayerID layer) {        // caller holds mu_
    NodeID best = 0;
    int64_t best_rate = 0;
    unsigned min_count = ~0u;
    for (auto &lc : load_) {
        auto sn = status_.find(lc.first);
        if (sn == status_.end()) continue;
        auto sl = sn->second.find(layer);
        if (sl == sn->second.end()) continue;
        const int64_t rate = sl->second.LimitRate == 0 ? INT64_MAX : sl->second.LimitRate;   // no rate limit, :962-966
        if (rate > best_rate || (rate == best_rate && (lc.second < min_count || (lc.second == min_count && lc.first < best)))) {
            best = lc.first;
            best_rate = rate;
            min_count = lc.second;
        }
    }
    return best;
}

void PullRetransmitLeaderNode::sendLayers() {
    std::vector<NodeID> dests;
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto &sn : status_)
            for (auto &sl : sn.second) owners_[sl.first].insert(sn.first);           // node.go:826-838
        std::vector<LayerID> sorted;
        for (auto &kv : owners_) sorted.push_back(kv.first);
        std::stable_sort(sorted.begin(), sorted.end(), [&](LayerID a, LayerID b) {   // rarest first, ties by id, :845-851
            if (owners_[a].size() != owners_[b].size()) return owners_[a].size() < owners_[b].size();
            return a < b;
        });
        for (auto &na : assignment_)                                                 // node.go:854-867
            for (auto &la : na.second) {
                auto sn = status_.find(na.first);
                bool has = false;
                if (sn != status_.end()) {
                    auto sl = sn->second.find(la.first);
                    has = sl != sn->second.end() && sl->second.Location == InmemLayer;
                }
                if (!has) jobs_[la.first][na.first] = JobInfo{};
            }
        for (auto &sn : status_) load_.emplace(sn.first, 0u);                        // node.go:870-874
        for (LayerID layer : sorted) {                                               // node.go:877-886
            auto jl = jobs_.find(layer);
            if (jl == jobs_.end()) continue;
            for (auto &dj : jl->second) {
                NodeID sender = getMinLoadedSender(layer);
                dj.second.sender = sender;
                dj.second.status = Pending;
                load_[sender]++;
            }
        }
        for (auto &jl : jobs_)
            if (!owners_.count(jl.first) && err_.empty()) err_ = "mode 2: nobody holds layer " + std::to_string(jl.first);
        for (auto &na : assignment_) dests.push_back(na.first);                      // ascending, node.go:891-896
    }
    {
        Plan empty;
        issue(empty);                        // an already satisfied assignment is ready at once (see LeaderNode::issue)
    }
    for (NodeID node : dests)
        Go([this, node] {
            if (assignNewJob(node)) Fail(std::string("failed to assign a new job: ") + last_error());
        });
}

bool PullRetransmitLeaderNode::getRarestOwnJob(NodeID node, LayerID *layer, NodeID *dest) {   // caller holds mu_
    size_t min_owners = ~(size_t)0;
    bool ok = false;
    auto sn = status_.find(node);
    if (sn == status_.end()) return false;
    for (auto &sl : sn->second) {
        auto jl = jobs_.find(sl.first);
        if (jl == jobs_.end() || jl->second.empty()) continue;
        for (auto &dj : jl->second) {
            if (dj.second.sender != node || dj.second.status != Pending) continue;
            const size_t cnt = owners_[sl.first].size();
            if (cnt < min_owners || (cnt == min_owners && sl.first < *layer)) {
                min_owners = cnt;
                *layer = sl.first;
                *dest = dj.first;
                ok = true;
            }
        }
    }
    return ok;
}

bool PullRetransmitLeaderNode::getRarestStealableJob(NodeID node, LayerID *layer, NodeID *dest, NodeID *from) {   // caller holds mu_
    bool have = false;
    size_t best_owners = 0;
    int64_t best_ttf = 0;
    auto sn = status_.find(node);
    if (sn == status_.end()) return false;
    for (auto &sl : sn->second) {
        auto jl = jobs_.find(sl.first);
        if (jl == jobs_.end()) continue;
        const size_t owner_count = owners_[sl.first].size();
        for (auto &dj : jl->second) {
            const NodeID sender = dj.second.sender;
            const int64_t sender_rate = status_[sender][sl.first].LimitRate;
            const int64_t node_rate = sl.second.LimitRate;
            if (sender == node || dj.second.status != Pending || load_[sender] == 0 ||
                (node_rate != 0 && node_rate < sender_rate))                         // node.go:1038-1044
                continue;
            int64_t ttf;
            auto pf = perf_.find(sender);
            if (pf == perf_.end()) ttf = INT64_MAX;      // the sender is still stuck at its first job, :1048-1051
            else ttf = pf->second.aveNs * (int64_t)load_[sender];
            if (!have || owner_count < best_owners || (owner_count == best_owners && ttf > best_ttf)) {
                have = true;
                best_owners = owner_count;
                best_ttf = ttf;
                *layer = sl.first;
                *dest = dj.first;
                *from = sender;
            }
        }
    }
    return have;
}

int PullRetransmitLeaderNode::sendRetransmit(LayerID layer, NodeID owner, NodeID dest) {
    {
        std::lock_guard<std::mutex> lk(mu_);             // what was actually sent, for the run report
        PlannedXfer x;
        x.Src = owner;
        x.Dest = dest;
        x.Layer = layer;
        const LayerMeta &om = status_[owner][layer];
        x.DataSize = x.TotalSize = om.DataSize >= 0 ? om.DataSize : layers_[layer].DataSize;
        plan_.Xfers.push_back(x);
    }
    if (owner == node_->GetMyID()) {                     // node.go:614-621
        LayerSrc ls;
        {
            std::lock_guard<std::mutex> lk(mu_);
            ls = layers_[layer];
        }
        return SendLayerTo(dest, layer, ls, ls.DataSize);
    }
    auto r = std::make_shared<RetransmitMsg>();
    r->SrcID = node_->GetMyID();
    r->Layer = layer;
    r->DestID = dest;
    return node_->GetTransport()->Send(owner, r);
}

int PullRetransmitLeaderNode::assignNewJob(NodeID node) {
    LayerID layer = 0;
    NodeID dest = 0, from = 0;
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (getRarestOwnJob(node, &layer, &dest)) {                                  // node.go:909-919
            JobInfo &j = jobs_[layer][dest];
            j.status = SendingRetransmit;
            j.retransmitTime = std::chrono::steady_clock::now();
            load_[node]--;
        } else if (getRarestStealableJob(node, &layer, &dest, &from)) {              // node.go:922-941
            load_[from]--;
            JobInfo &j = jobs_[layer][dest];
            j.sender = node;
            j.status = SendingRetransmit;
            j.retransmitTime = std::chrono::steady_clock::now();
        } else {
            return 0;                                    // "there is no job left to assign"
        }
    }
    return sendRetransmit(layer, node, dest);
}

void PullRetransmitLeaderNode::handleAckMsg(const AckMsg &m) {
    bool fire = false;
    {
        std::lock_guard<std::mutex> lk(mu_);
        LayerMeta meta;
        meta.Location = m.location;
        status_[m.SrcID][m.Layer] = meta;
    }
    if (!completion_[m.SrcID] && assignmentSatisfied()) {                            // node.go:752-759
        completion_[m.SrcID] = true;
        fire = true;
    }
    if (fire && !ready_.fired()) {
        sendStartup();
        ready_.Fire(assignment_);
    }
    NodeID sender;
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto jl = jobs_.find(m.Layer);
        if (jl == jobs_.end()) return;
        auto dj = jl->second.find(m.SrcID);
        if (dj == jl->second.end()) return;              // not one of our jobs (node.go:765-769)
        sender = dj->second.sender;
        const int64_t took = std::chrono::duration_cast<std::chrono::nanoseconds>(
                                 std::chrono::steady_clock::now() - dj->second.retransmitTime).count();
        Perf &pf = perf_[sender];                        // the reference's running "average", node.go:788-792, as written
        pf.aveNs = (took + pf.aveNs) / (int64_t)(pf.completed + 1);
        pf.completed++;
        jl->second.erase(dj);                            // node.go:797
    }
    // the sender is idle again: hand it its next job.  (assignNewJob may block in a layer send when the sender is the
    // leader itself, so it runs off the dispatcher as the reference's `go handleAckMsg` does.)
    Go([this, sender] {
        if (assignNewJob(sender)) Fail(std::string("failed to assign a new job after an ack: ") + last_error());
    });
}

// ----------------------------------------------------------------------- FlowRetransmitLeaderNode
FlowRetransmitLeaderNode::FlowRetransmitLeaderNode(N *node, LayersSrc layers, Assignment assignment,
                                                   std::map<NodeID, int64_t> network_bw, PlanOptions opt)
    : RetransmitLeaderNode(node, std::move(layers), std::move(assignment), opt, false), bw_(std::move(network_bw)) {
    Start();
}

void FlowRetransmitLeaderNode::Dispatch(const MessagePtr &m) {
    if (auto *f = dynamic_cast<FlowRetransmitMsg *>(m.get())) {
        FlowRetransmitMsg copy = *f;
        LayersSrc ls = Layers();
        Go([this, copy, ls] {
            if (HandleFlowRetransmit(node_, ls, copy)) Fail(std::string("flow retransmit failed: ") + last_error());
        });
        return;
    }
    LeaderNode::Dispatch(m);        // announce, ack — and layerMsg, which the reference's mode-3 leader drops (:1126-1127)
}

Plan FlowRetransmitLeaderNode::makePlan() {
    std::lock_guard<std::mutex> lk(mu_);
    return PlanMode3(node_->GetMyID(), layers_, status_, assignment_, bw_, opt_);
}

void FlowRetransmitLeaderNode::sendLayers() {
    Plan p;
    try {
        p = makePlan();                                  // assignJobs, node.go:1200-1234
    } catch (const std::exception &e) {
        Fail(e.what());
        return;
    }
    issue(p);
    if (batched_) {
        auto *dt = dynamic_cast<DeviceTransport *>(node_->GetTransport());
        if (!dt) { Fail("batched plans need a DeviceTransport"); return; }
        if (!p.Xfers.empty() && dt->SendBatch(p.Xfers, nullptr)) Fail(std::string("batched send failed: ") + last_error());
        return;
    }
    for (auto &j : p.SelfJobs) {                         // node.go:1239-1248
        auto f = std::make_shared<FlowRetransmitMsg>();
        f->SrcID = node_->GetMyID();
        f->Layer = j.Layer;
        f->DestID = j.SenderID;
        f->DataSize = j.DataSize;
        f->Offset = j.Offset;
        {
            std::lock_guard<std::mutex> lk(mu_);
            f->Rate = status_[j.SenderID][j.Layer].LimitRate;
        }
        node_->GetTransport()->Send(j.SenderID, f);
    }
    for (auto &x : p.Xfers) {                            // dispatchJob, node.go:1264-1288
        auto f = std::make_shared<FlowRetransmitMsg>();
        f->SrcID = node_->GetMyID();
        f->Layer = x.Layer;
        f->DestID = x.Dest;
        f->DataSize = x.DataSize;
        f->Offset = x.Offset;
        f->Rate = x.Rate;
        if (node_->GetTransport()->Send(x.Src, f)) Fail("couldn't dispatch job for layer " + std::to_string(x.Layer));
    }
}

// ----------------------------------------------------------------------------------- ReceiverNode
ReceiverNode::ReceiverNode(N *node, LayersSrc layers, bool autostart) : RoleBase(node, std::move(layers)) {
    if (autostart) Start();
}

void ReceiverNode::Dispatch(const MessagePtr &m) {
    if (auto *l = dynamic_cast<LayerMsg *>(m.get())) handleLayerMsg(*l);
    else if (dynamic_cast<StartupMsg *>(m.get())) ready_.Fire(1);        // handleStartupMsg, node.go:1387-1389
}

void ReceiverNode::handleLayerMsg(const LayerMsg &m) {
    bool complete = false;
    StoreLayer(m, &complete);
    if (!complete) return;
    auto ack = std::make_shared<AckMsg>();
    ack->SrcID = node_->GetMyID();
    ack->Layer = m.Layer;
    if (node_->GetTransport()->Send(node_->getLeader(), ack)) Fail("failed to send ackMsg");   // node.go:1379-1383
}

int ReceiverNode::Announce() {
    auto a = std::make_shared<AnnounceMsg>();
    a->SrcID = node_->GetMyID();
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto &kv : layers_) {
            LayerMeta m;                                  // Location and LimitRate only, node.go:1397-1401
            m.Location = kv.second.Meta.Location;
            m.LimitRate = kv.second.Meta.LimitRate;
            m.DataSize = kv.second.DataSize;              // extension (types.h)
            a->Layers[kv.first] = m;
        }
    }
    NodeID next;
    if (!node_->getNextHop(node_->getLeader(), &next)) return set_error(DISSEM_ENOENT, "leader not found in routing table");
    return node_->GetTransport()->Send(next, a);
}

// ------------------------------------------------------------------------- RetransmitReceiverNode
RetransmitReceiverNode::RetransmitReceiverNode(N *node, LayersSrc layers, bool autostart)
    : ReceiverNode(node, std::move(layers), false) {
    if (autostart) Start();
}

void RetransmitReceiverNode::Dispatch(const MessagePtr &m) {
    if (auto *r = dynamic_cast<RetransmitMsg *>(m.get())) {
        RetransmitMsg copy = *r;
        Go([this, copy] { handleRetransmitMsg(copy); });
        return;
    }
    ReceiverNode::Dispatch(m);
}

void RetransmitReceiverNode::handleRetransmitMsg(const RetransmitMsg &m) {
    LayerSrc ls;
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = layers_.find(m.Layer);
        if (it == layers_.end()) { err_ = err_.empty() ? "asked to retransmit a layer this node does not hold" : err_; return; }
        ls = it->second;
    }
    node_->addNode(m.DestID);                             // node.go:1467-1468
    SendLayerTo(m.DestID, m.Layer, ls, ls.DataSize);      // node.go:1478-1479
}

// --------------------------------------------------------------------- FlowRetransmitReceiverNode
FlowRetransmitReceiverNode::FlowRetransmitReceiverNode(N *node, LayersSrc layers)
    : RetransmitReceiverNode(node, std::move(layers), false) {
    Start();
}

void FlowRetransmitReceiverNode::Dispatch(const MessagePtr &m) {
    if (auto *f = dynamic_cast<FlowRetransmitMsg *>(m.get())) {
        FlowRetransmitMsg copy = *f;
        LayersSrc ls = Layers();
        Go([this, copy, ls] {
            if (HandleFlowRetransmit(node_, ls, copy)) Fail(std::string("flow retransmit failed: ") + last_error());
        });
        return;
    }
    RetransmitReceiverNode::Dispatch(m);
}

int HandleFlowRetransmit(N *node, const LayersSrc &layers, const FlowRetransmitMsg &m) {
    auto it = layers.find(m.Layer);
    if (it == layers.end()) return set_error(DISSEM_ENOENT, "node %u holds no layer %u", node->GetMyID(), m.Layer);
    LayerSrc partial = it->second;                        // node.go:1599-1606
    node->addNode(m.DestID);
    const int64_t total = it->second.DataSize;
    partial.DataSize = m.DataSize;
    partial.Offset = m.Offset;
    partial.Meta.LimitRate = m.Rate;
    auto lm = std::make_shared<LayerMsg>();
    lm->SrcID = node->GetMyID();
    lm->Layer = m.Layer;
    lm->Src_ = partial;
    lm->TotalSize = total;
    if (m.DestID == node->GetMyID() && partial.Meta.Location != DiskLayer) {
        // self job: the destination already holds the layer; the reference ships it to itself over TCP
        // (node.go:1239-1248); here only the header is delivered so that the ack still flows.  A layer held on DISK
        // does take the transport: the send to itself is what loads the file into the node's HBM.
        node->GetTransport()->Deliver().Push(DecodeMsg(EncodeMsg(*lm)));
        return 0;
    }
    return node->GetTransport()->Send(m.DestID, lm);
}

}  // namespace host
}  // namespace dissem
