// The data model of the reference's `distributor` package (distributor/node.go:128-211) with the layer
// store moved to HBM: LayerSrc carries a device handle (node, layer) instead of `InmemData *[]byte`.
#pragma once
#include <cstdint>
#include <map>
#include <set>
#include <string>
#include <vector>

namespace dissem {
namespace host {

using NodeID = uint32_t;    // node.go:128
using LayerID = uint32_t;   // node.go:129

enum LayerLocation : uint8_t { InmemLayer = 0, DiskLayer = 1, ClientLayer = 2 };   // node.go:187-191
enum SourceType : uint8_t { SourceClient = 0, SourceDisk = 1, SourceMem = 2 };     // node.go:196-200

struct LayerMeta {            // node.go:134-138
    LayerLocation Location = InmemLayer;
    int64_t LimitRate = 0;
    SourceType Source = SourceClient;
    // Extension: an announce also carries the layer size, so that a leader that does not hold the layer can
    // still plan its transfer (the reference reads sizes from the leader's own copy only, flow.go:151).
    int64_t DataSize = -1;
};

using LayerIDs = std::map<LayerID, LayerMeta>;          // node.go:141
using Assignment = std::map<NodeID, LayerIDs>;          // node.go:177
using Status = std::map<NodeID, LayerIDs>;              // node.go:179
using NodeIDs = std::set<NodeID>;                       // node.go:132

// LayerSrc, node.go:200-211.  `InmemData` is an HBM buffer owned by the engine's store and addressed by
// (node, layer); the fields the reference uses for slicing a fragment are kept.
struct LayerSrc {
    bool InDevice = false;   // stands for InmemData != nil
    std::string Fp;          // DiskLayer: the layer file, attached to the engine with dissem_layer_attach_file
    int64_t DataSize = 0;
    int64_t Offset = 0;
    LayerMeta Meta;
};
using LayersSrc = std::map<LayerID, LayerSrc>;          // node.go:172

// One planned layerMsg: bytes [Offset, Offset+DataSize) of a layer from Src to Dest.
struct PlannedXfer {
    NodeID Src = 0, Dest = 0;
    LayerID Layer = 0;
    int64_t Offset = 0, DataSize = 0, TotalSize = 0;
    int64_t Rate = 0;        // mode 3: dataSize / minTime (node.go:1281); 0 elsewhere
};

}  // namespace host
}  // namespace dissem
