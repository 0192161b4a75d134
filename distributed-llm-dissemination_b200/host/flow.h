// Mode-3 planner: the minimum integer time t for which a max-flow through
//   source -> senders[NetworkBW*t] -> per-sender sources[LimitRate*t] -> layers[inf] -> receivers[layer size]
//          -> sink[NetworkBW*t]
// carries every assigned layer, and the per-(sender, layer) fragments read off that flow.
// Restates distributor/flow.go:55-353 (graph :55-144, capacities :221-270, search :146-219, Edmonds-Karp
// :283-353).  The augmenting-path ORDER is part of the result (it decides the fragment sizes), so vertices are
// numbered exactly as flow.go numbers them and the BFS visits neighbours in ascending vertex index.
#pragma once
#include <cstdint>
#include <map>
#include <vector>

#include "types.h"

namespace dissem {
namespace host {

constexpr NodeID kAnyDest = 0xFFFFFFFFu;

struct FlowJobInfo {          // flowJobInfo, flow.go:30-35
    NodeID SenderID = 0;
    LayerID Layer = 0;
    int64_t DataSize = 0;
    int64_t Offset = 0;
    NodeID Dest = kAnyDest;   // set only by the per-destination extension (FlowOptions::PerDestLayers)
};

struct FlowOptions {
    int64_t TimeScale = 1;            // planner ticks per second; 1 = the reference's integer seconds
    bool ZeroRateUnlimited = false;   // treat LimitRate == 0 as "no cap" (the reference would never terminate)
    // Extension (the reference accepts ONE destination per layer, node.go:1078, and counts each distinct layer once in
    // the required flow, flow.go:150-152): one layer vertex per (layer, destination) pair, demand = sum over pairs,
    // fragments and offsets per pair.  With one destination per layer the graph, and hence the plan, is unchanged.
    bool PerDestLayers = false;
};

class FlowGraph {
public:
    // status: what every node announced (layer -> meta); sizes: layer -> DataSize as the leader knows it;
    // assignment: only the (dest, layer) pairs that need network transfer (modifiedAssignment, node.go:1204).
    FlowGraph(const Status &status, const std::map<LayerID, int64_t> &sizes,
              const std::map<NodeID, int64_t> &network_bw, const Assignment &assignment, FlowOptions opt);

    int64_t UpdateMaxFlow(int64_t t);                              // flow.go:319-353
    // Returns the minimal t and fills `jobs` (ascending sender, then layer; offsets cumulative per layer).
    int64_t GetJobAssignment(std::vector<FlowJobInfo> *jobs);      // flow.go:146-219
    int NumVertex() const { return n_; }

private:
    int64_t Cap(int64_t rate, int64_t t) const;
    void BuildEdgeCapacity(int64_t t);                             // flow.go:221-270
    bool Bfs(int src, int dst, std::vector<int> *parent);         // flow.go:283-315

    Status status_;
    std::map<LayerID, int64_t> sizes_;
    std::map<NodeID, int64_t> bw_;
    Assignment assignment_;
    FlowOptions opt_;
    std::vector<std::pair<LayerID, NodeID>> needed_;     // layer vertices: (layer, kAnyDest) or (layer, dest)
    int n_ = 0, src_ = 0, sink_ = 0;
    std::map<NodeID, int> sender_idx_, receiver_idx_;
    std::map<std::pair<NodeID, int>, int> client_idx_;
    std::map<std::pair<LayerID, NodeID>, int> layer_idx_;
    std::vector<std::pair<LayerID, NodeID>> layer_vertices(LayerID l) const;
    std::vector<int64_t> cap_;                     // residual capacities, n x n
    std::vector<std::vector<int>> nbr_;            // ascending neighbour lists (either direction)
};

}  // namespace host
}  // namespace dissem
