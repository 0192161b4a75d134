// The three plan generators on the hot path, as pure functions of (status, layers, assignment):
//   mode 0  LeaderNode.sendLayers                         distributor/node.go:326-352
//   mode 1  RetransmitLeaderNode.sendLayers/sendRetransmit distributor/node.go:554-626
//   mode 3  FlowRetransmitLeaderNode.assignJobs/sendLayers/dispatchJob  distributor/node.go:1200-1288
// The role state machines (node.h) call these when every receiver has announced; dissem_plan_json exposes
// them for drivers that submit the whole plan as one device schedule.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "config.h"
#include "flow.h"
#include "types.h"

namespace dissem {
namespace host {

struct PlanError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

enum OwnerPolicy { OwnerRandom = 0, OwnerBalanced = 1 };

struct PlanOptions {
    uint64_t Seed = 0;                 // mode 1: the reference's owner pick is Go map order (node.go:582-587)
    OwnerPolicy Policy = OwnerRandom;  // OwnerBalanced: least bytes sent so far, ties -> lowest id (ours)
    FlowOptions Flow;                  // mode 3
    bool SizesFromAnnounces = false;   // mode 3: take sizes of layers the leader lacks from the announces
};

struct Plan {
    std::vector<PlannedXfer> Xfers;
    std::vector<FlowJobInfo> SelfJobs; // mode 3: (dest, layer) the destination already holds (node.go:1208-1209)
    int64_t MinTime = 0;               // mode 3: t
};

// status: node -> layers it announced (the leader's own entry included, node.go:252-257);
// leader_layers: the leader's LayersSrc.
Plan PlanMode0(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a);
Plan PlanMode1(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a,
               const PlanOptions &opt);
Plan PlanMode3(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a,
               const std::map<NodeID, int64_t> &network_bw, const PlanOptions &opt);
// mode 2, the deterministic part: PullRetransmitLeaderNode.sendLayers' initial job table (node.go:810-886) — layers
// taken rarest first (fewest owners, ties by id), each (layer, dest) job given to getMinLoadedSender (node.go:948-978:
// highest LimitRate, 0 = unlimited; ties by fewest jobs so far, then lowest id).  Which jobs are later STOLEN depends on
// measured completion times (node.go:980-1073) and is decided at run time by the role (node.cpp).
Plan PlanMode2Initial(NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a);
Plan MakePlan(int mode, NodeID leader, const LayersSrc &leader_layers, const Status &status, const Assignment &a,
              const std::map<NodeID, int64_t> &network_bw, const PlanOptions &opt);

// The status map the leader ends up with once every node of the config has announced.
Status StatusFromConfig(const Config &cfg);
uint64_t Mix64(uint64_t z);

}  // namespace host
}  // namespace dissem
