// device_transport.go — a distributor.Transport whose layer branch is the B200 engine (libdissem.so).
//
// STATUS: shipped as source, NOT compiled or run: there is no Go toolchain in the image this repository is built and
// tested in.  The same call sequence is exercised through the C ABI from C++ (host/transport.cpp) and Python (tests/).
// Drop this file into the reference's `distributor` package (it uses the package's unexported message types) and build
// with
//   CGO_CFLAGS="-I$REPO/include" CGO_LDFLAGS="-L$REPO/distributed-llm-dissemination_b200 -ldissem" go build ./cmd
// then construct NewDeviceTransport where cmd/main.go:125 constructs NewTcpTransport.
package distributor

/*
#cgo LDFLAGS: -ldissem
#include <stdlib.h>
#include "dissem.h"
*/
import "C"

import (
	"encoding/json"
	"fmt"
	"sync"
	"unsafe"
)

// DeviceLayer marks a LayerSrc whose bytes live in the engine's HBM store under (node, layer).
const DeviceLayer LayerLocation = 3

// deviceFabric is shared by the transports of one process (the role of InmemoryTransport's registry,
// transport.go:507-531) and owns the dissem context.
type deviceFabric struct {
	ctx   *C.dissem_ctx
	mu    sync.RWMutex
	peers map[NodeID]*DeviceTransport
	// The C ABI names nodes by a dense index 0..n-1; the reference's NodeIDs are arbitrary uints (config "Id").
	index    map[NodeID]C.int
	uploaded map[[2]uint]bool // (node, layer) whose host InmemData already sits in HBM
	refs     int              // transports still open; the last Close releases the context
}

// NewDeviceFabric: ids[i] is the NodeID of the config's i-th node, deviceOfNode[i] the CUDA ordinal it lives on.
func NewDeviceFabric(ids []NodeID, deviceOfNode []int) (*deviceFabric, error) {
	if len(ids) == 0 || len(ids) != len(deviceOfNode) {
		return nil, fmt.Errorf("NewDeviceFabric: need one device per node and at least one node (got %d ids, %d devices)", len(ids), len(deviceOfNode))
	}
	devs := make([]C.int, len(deviceOfNode))
	index := make(map[NodeID]C.int, len(ids))
	for i, d := range deviceOfNode {
		devs[i] = C.int(d)
		index[ids[i]] = C.int(i)
	}
	var ctx *C.dissem_ctx
	if rc := C.dissem_init(C.int(len(devs)), &devs[0], &ctx); rc != 0 {
		return nil, fmt.Errorf("dissem_init: %s", C.GoString(C.dissem_last_error()))
	}
	return &deviceFabric{ctx: ctx, peers: make(map[NodeID]*DeviceTransport), index: index, uploaded: make(map[[2]uint]bool)}, nil
}

func (f *deviceFabric) idx(id NodeID) (C.int, error) {
	i, ok := f.index[id]
	if !ok {
		return 0, fmt.Errorf("node %d is not part of this fabric", id)
	}
	return i, nil
}

// ensureSource: the reference keeps an in-memory layer as a host slice (LayerSrc.InmemData, cmd/config.go:159-171).  The
// engine sends from HBM, so the first Send of a layer uploads the sender's slice once (dissem_layer_alloc +
// dissem_layer_write).  A deployment that never wants layers in host memory changes CreateInmemLayer instead to call
// dissem_layer_alloc / dissem_layer_fill and to leave InmemData empty; then there is nothing to upload here.
func (f *deviceFabric) ensureSource(node C.int, id NodeID, layer LayerID, src LayerSrc) error {
	if src.InmemData == nil || len(*src.InmemData) == 0 {
		return nil
	}
	key := [2]uint{uint(id), uint(layer)}
	f.mu.Lock()
	defer f.mu.Unlock()
	if f.uploaded[key] {
		return nil
	}
	if C.dissem_layer_info(f.ctx, node, C.int(layer), nil, nil) != 0 {
		if rc := C.dissem_layer_alloc(f.ctx, node, C.int(layer), C.int64_t(len(*src.InmemData))); rc != 0 {
			return derr("dissem_layer_alloc(source)")
		}
	}
	if rc := C.dissem_layer_write(f.ctx, node, C.int(layer), unsafe.Pointer(&(*src.InmemData)[0]), 0, C.int64_t(len(*src.InmemData))); rc != 0 {
		return derr("dissem_layer_write")
	}
	f.uploaded[key] = true
	return nil
}

type DeviceTransport struct {
	fabric          *deviceFabric
	me              NodeID
	incomingMsgChan chan Message
}

func NewDeviceTransport(f *deviceFabric, me NodeID, bufSize uint) *DeviceTransport {
	t := &DeviceTransport{fabric: f, me: me, incomingMsgChan: make(chan Message, bufSize)}
	f.mu.Lock()
	f.peers[me] = t
	f.refs++
	f.mu.Unlock()
	return t
}

func derr(what string) error { return fmt.Errorf("%s: %s", what, C.GoString(C.dissem_last_error())) }

// Send: the layer branch is the device path; every other message goes through the JSON envelope exactly as
// sendTransportMsg/decodeMsg do (transport.go:375-405, message.go:280-301), then into the peer's channel.
func (t *DeviceTransport) Send(destID NodeID, message Message) error {
	t.fabric.mu.RLock()
	peer, ok := t.fabric.peers[destID]
	t.fabric.mu.RUnlock()
	if !ok {
		return fmt.Errorf("addr of %d does not exist", destID)
	}
	if lm, ok := message.(*layerMsg); ok {
		ctx := t.fabric.ctx
		src, err := t.fabric.idx(lm.SrcID)
		if err != nil {
			return err
		}
		dst, err := t.fabric.idx(destID)
		if err != nil {
			return err
		}
		if lm.LayerSrc.Meta.Location != DiskLayer {
			if err := t.fabric.ensureSource(src, lm.SrcID, lm.LayerID, lm.LayerSrc); err != nil {
				return err
			}
		}
		// receiver side of transport.go:201: make sure the destination buffer exists
		if C.dissem_layer_info(ctx, dst, C.int(lm.LayerID), nil, nil) != 0 {
			if rc := C.dissem_layer_alloc(ctx, dst, C.int(lm.LayerID), C.int64_t(lm.TotalSize)); rc != 0 && rc != C.DISSEM_EEXIST {
				return derr("dissem_layer_alloc")
			}
		}
		var ticket C.dissem_ticket
		if lm.LayerSrc.Meta.Location == DiskLayer {
			// transport.go:351-367: the layer is a file; it was registered with dissem_layer_attach_file(node, layer,
			// LayerSrc.Fp, 0, DataSize) when CreateDiskLayer made it, and is streamed through pinned staging buffers
			if rc := C.dissem_send_file(ctx, src, dst, C.int(lm.LayerID),
				C.int64_t(lm.LayerSrc.Offset), C.int64_t(lm.LayerSrc.DataSize), C.int64_t(lm.TotalSize), &ticket); rc != 0 {
				return derr("dissem_send_file")
			}
		} else if rc := C.dissem_send(ctx, src, dst, C.int(lm.LayerID),
			C.int64_t(lm.LayerSrc.Offset), C.int64_t(lm.LayerSrc.DataSize), C.int64_t(lm.TotalSize), &ticket); rc != 0 {
			return derr("dissem_send")
		}
		if rc := C.dissem_wait(ctx, ticket, nil); rc != 0 { // Send returns once the bytes are written
			return derr("dissem_wait")
		}
		// what handleIncomingMsg hands to the node (transport.go:221-222); InmemData must be non-nil
		// (dereferenced at node.go:385, :1363): a zero-length slice stands for the device buffer.
		handle := LayerData{}
		peer.incomingMsgChan <- &layerMsg{lm.SrcID, lm.LayerID,
			LayerSrc{&handle, "", lm.LayerSrc.DataSize, lm.LayerSrc.Offset, LayerMeta{Location: InmemLayer}}, lm.TotalSize}
		return nil
	}
	// control plane: marshal -> unmarshal, so unexported fields are lost as over TCP
	raw, err := encodeEnvelope(message) // json.Marshal(TransportMsg{Type, Src, Payload}), transport.go:382-396
	if err != nil {
		return err
	}
	msg, err := decodeEnvelope(raw) // decodeMsg, message.go:280-301
	if err != nil {
		return err
	}
	peer.incomingMsgChan <- msg
	return nil
}

func (t *DeviceTransport) RegisterPipe(LayerID, NodeID) error { return fmt.Errorf("pipes are not supported") }
func (t *DeviceTransport) Broadcast(m Message) error {
	t.fabric.mu.RLock() // ids are copied first, as transport.go:290-297 does: Send takes the lock itself
	ids := make([]NodeID, 0, len(t.fabric.peers))
	for id := range t.fabric.peers {
		ids = append(ids, id)
	}
	t.fabric.mu.RUnlock()
	for _, id := range ids {
		_ = t.Send(id, m) // errors are logged and dropped, transport.go:299-303
	}
	return nil
}
func (t *DeviceTransport) Deliver() <-chan Message { return t.incomingMsgChan }
func (t *DeviceTransport) GetAddress() string      { return fmt.Sprintf("gpu:%d", t.me) }

// Close (transport.go:474-491): the transport leaves the fabric; the last one out releases the engine's context and with
// it every HBM buffer.
func (t *DeviceTransport) Close() error {
	f := t.fabric
	f.mu.Lock()
	defer f.mu.Unlock()
	if _, ok := f.peers[t.me]; !ok {
		return nil
	}
	delete(f.peers, t.me)
	close(t.incomingMsgChan)
	f.refs--
	if f.refs == 0 && f.ctx != nil {
		rc := C.dissem_close(f.ctx)
		f.ctx = nil
		if rc != 0 {
			return derr("dissem_close")
		}
	}
	return nil
}

// encodeEnvelope / decodeEnvelope: the control plane keeps the reference's wire format, so that what a node receives is
// what TCP would have delivered (unexported fields such as ackMsg.location are lost, message.go:62-66).
func encodeEnvelope(message Message) ([]byte, error) {
	payload, err := json.Marshal(message) // transport.go:382
	if err != nil {
		return nil, err
	}
	return json.Marshal(TransportMsg{Type: message.Type(), Src: message.Src(), Payload: payload}) // transport.go:387-393
}

func decodeEnvelope(raw []byte) (Message, error) {
	var m TransportMsg
	if err := json.Unmarshal(raw, &m); err != nil {
		return nil, err
	}
	return decodeMsg(m) // message.go:280-301
}
