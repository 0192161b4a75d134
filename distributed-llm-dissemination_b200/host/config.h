// Config JSON of the reference (cmd/config.go:14-45), both the nested schema the code parses and the flat
// legacy schema its README documents (readme.md:17-64).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "json.h"
#include "types.h"

namespace dissem {
namespace host {

struct InitialLayerConf {     // cmd/config.go:34-36 (+ the source it was listed under)
    int64_t LayerSize = 0;
    SourceType Source = SourceClient;
};

struct NodeConf {             // cmd/config.go:21-28
    NodeID ID = 0;
    std::string Addr;
    int64_t NetworkBW = 0;
    bool IsLeader = false;
    std::map<SourceType, int64_t> Sources;
    std::map<LayerID, InitialLayerConf> InitialLayers;   // flattened over sources; later source wins
};

struct Config {               // cmd/config.go:14-19
    std::vector<NodeConf> Nodes;
    Assignment Assign;
    int64_t LayerSize = 0;
    bool FlatSchema = false;

    const NodeConf *Leader() const;                  // GetLeaderConf, cmd/config.go:64-72
    const NodeConf *Node(NodeID id) const;           // GetNodeConf, cmd/config.go:74-82
    std::map<NodeID, int64_t> NodeNetworkBW() const; // cmd/main.go:136-139
};

// ReadJson (cmd/config.go:48-62) — but malformed JSON is an error here, not silently ignored (:59).
Config ParseConfig(const std::string &text);

// CreateLayers (cmd/config.go:94-117): the LayersSrc a node starts with.  All layers are in device memory
// (no -s path on the GPU box); Meta.LimitRate = Sources[sourceType] (:111).
LayersSrc CreateLayers(const NodeConf &nc);

}  // namespace host
}  // namespace dissem
