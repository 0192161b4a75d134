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
)

// DeviceLayer marks a LayerSrc whose bytes live in the engine's HBM store under (node, layer).
const DeviceLayer LayerLocation = 3

// deviceFabric is shared by the transports of one process (the role of InmemoryTransport's registry,
// transport.go:507-531) and owns the dissem context.
type deviceFabric struct {
	ctx   *C.dissem_ctx
	mu    sync.RWMutex
	peers map[NodeID]*DeviceTransport
}

func NewDeviceFabric(deviceOfNode []int) (*deviceFabric, error) {
	devs := make([]C.int, len(deviceOfNode))
	for i, d := range deviceOfNode {
		devs[i] = C.int(d)
	}
	var ctx *C.dissem_ctx
	if rc := C.dissem_init(C.int(len(devs)), &devs[0], &ctx); rc != 0 {
		return nil, fmt.Errorf("dissem_init: %s", C.GoString(C.dissem_last_error()))
	}
	return &deviceFabric{ctx: ctx, peers: make(map[NodeID]*DeviceTransport)}, nil
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
		// receiver side of transport.go:201: make sure the destination buffer exists
		if C.dissem_layer_info(ctx, C.int(destID), C.int(lm.LayerID), nil, nil) != 0 {
			if rc := C.dissem_layer_alloc(ctx, C.int(destID), C.int(lm.LayerID), C.int64_t(lm.TotalSize)); rc != 0 && rc != C.DISSEM_EEXIST {
				return derr("dissem_layer_alloc")
			}
		}
		var ticket C.dissem_ticket
		if lm.LayerSrc.Meta.Location == DiskLayer {
			// transport.go:351-367: the layer is a file; it was registered with dissem_layer_attach_file(node, layer,
			// LayerSrc.Fp, 0, DataSize) when CreateDiskLayer made it, and is streamed through pinned staging buffers
			if rc := C.dissem_send_file(ctx, C.int(lm.SrcID), C.int(destID), C.int(lm.LayerID),
				C.int64_t(lm.LayerSrc.Offset), C.int64_t(lm.LayerSrc.DataSize), C.int64_t(lm.TotalSize), &ticket); rc != 0 {
				return derr("dissem_send_file")
			}
		} else if rc := C.dissem_send(ctx, C.int(lm.SrcID), C.int(destID), C.int(lm.LayerID),
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
func (t *DeviceTransport) Close() error            { return nil } // the fabric owns the context: C.dissem_close(f.ctx)

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
