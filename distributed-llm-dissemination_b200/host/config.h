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

// CreateLayers (cmd/config.go:94-117): the LayersSrc a node starts with; Meta.LimitRate = Sources[sourceType] (:111).
// save_disk (the `-s` flag, :105-106): the layers are files `<storage_path>/layers/<node id>/<layer id>.layer`
// (CreateDiskLayer, :133-157) with Location = DiskLayer; otherwise they are in device memory (CreateInmemLayer).
// Only the descriptions are made here; the caller creates the files / HBM buffers.
LayersSrc CreateLayers(const NodeConf &nc, bool save_disk = false, const std::string &storage_path = "");
// The path CreateDiskLayer gives a layer file (cmd/config.go:135-139).
std::string DiskLayerPath(const std::string &storage_path, NodeID node, LayerID layer);

}  // namespace host
}  // namespace dissem
