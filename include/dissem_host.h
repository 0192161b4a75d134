/*
 * dissem_host.h — C ABI of the host-side mirror of the reference's `distributor` package, exported by the
 * same libdissem.so.  These entry points are pure host logic (no GPU needed) unless stated otherwise.
 *
 * Where Go exists the reference's own node.go / flow.go keep producing the plan and only Transport is
 * swapped; this header is for hosts without Go (this image): the C++ restatement of the
 * plan generators and role state machines, driven from C, C++ or ctypes.
 */
#ifndef DISSEM_HOST_H
#define DISSEM_HOST_H

#include <stdint.h>

#include "dissem.h"

#ifdef __cplusplus
extern "C" {
#endif

#define DISSEM_POLICY_RANDOM 0   /* mode 1: seeded stand-in for Go's map-order owner pick (node.go:582-587) */
#define DISSEM_POLICY_BALANCED 1 /* mode 1: least-loaded owner (ours) */

#define DISSEM_PLAN_ZERO_RATE_UNLIMITED 1 /* mode 3: LimitRate == 0 means "no cap" */
#define DISSEM_PLAN_SIZES_FROM_ANNOUNCES 2 /* mode 3: sizes of layers the leader lacks come from announces */
#define DISSEM_PLAN_MULTI_DEST 4 /* mode 3: one demand per (layer, destination) pair instead of the reference's one
                                    destination per layer (node.go:1078); unchanged plan when every layer has one */

/* Parse the reference's config JSON (cmd/config.go:14-45; nested or README-flat schema) and produce the plan
 * the leader of `mode` (cmd/main.go:153-165: 0 LeaderNode, 1 RetransmitLeaderNode, 3 FlowRetransmitLeaderNode)
 * issues once every node has announced.  *json_out receives a malloc'ed NUL-terminated JSON document:
 *   {"mode":M,"leader":L,"t":T,"n_nodes":N,
 *    "xfers":[[src,dst,layer,off,size,total,rate],...],
 *    "self_jobs":[[node,layer,size],...],
 *    "nodes":[{"id":I,"leader":bool,"bw":B,"layers":{"<layer>":size,...}},...],
 *    "assignment":{"<node>":[layer,...],...}}
 * time_scale: planner ticks per second for mode 3 (1 = the reference's integer seconds).
 * Free with dissem_free.  Errors: DISSEM_EPARSE, DISSEM_EPLAN, DISSEM_EINVAL. */
int dissem_plan_json(const char *config_json, int mode, uint64_t seed, int policy, int64_t time_scale,
                     int flags, char **json_out);
void dissem_free(void *p);

/* Run one whole distribution in this process, every node of the config being a (slice of a) local GPU:
 * leader and receivers as threads speaking the reference's control-plane messages (message.go) over an
 * in-process Transport whose layer branch is the device engine.  Needs a GPU.
 * node_device[i]: CUDA ordinal for the config's i-th node (NULL = all on device 0).
 * opts_json (may be NULL): {"seed":S,"policy":P,"time_scale":T,"flags":F,"fill":K,"batched":bool,"verify":bool,
 *                           "storage":"<path>","pace":bool,"reserve":bool,"dump_tables":bool}
 *   batched=false: one Transport.Send per layerMsg, as the reference issues them;
 *   batched=true : the leader hands the whole plan to dissem_schedule_submit (one launch per source GPU);
 *   storage      : the `-s` flag (cmd/main.go:17, :93-101): every node's initial layers are files
 *                  <path>/layers/<node id>/<layer id>.layer (CreateDiskLayer, cmd/config.go:133-157; written when missing,
 *                  with the run's fill) attached with dissem_layer_attach_file and sent as DiskLayer sources;
 *   reserve      : (default true) receive buffers for the Assignment are set aside with dissem_layer_reserve before the
 *                  clock starts; false leaves their allocation inside each send, where the reference has it;
 *   pace         : unbatched layer sends honour LayerSrc.Meta.LimitRate as writeWithLimit does (dissem_pace_schedule);
 *   dump_tables  : the report gains "tables": one entry per assigned (node, layer) with its size, received counter, fused
 *                  checksum table and first / last 4 KiB as hex strings — the evidence a caller needs to check the run
 *                  against an oracle of its own ("verified" only compares values this library computed with each other).
 * *json_out: {"time_to_deliver_ms":..,"device_ms":..,"bytes":..,"n_xfers":..,"acks":..,"verified":bool,...} */
int dissem_run_json(const char *config_json, int mode, const int *node_device, const char *opts_json,
                    char **json_out);

/* Wire format of the control plane (message.go:273-301): encode one message kind to the
 * envelope {"type":T,"src":"<id>","payload":{...}} and decode it back; used by the parity tests.
 * kind_json: {"kind":"announce|ack|retransmit|flowRetransmit|startup|simple|layerHeader", fields...}. */
int dissem_msg_roundtrip_json(const char *kind_json, char **wire_out, char **decoded_out);

/* The Transport interface (transport.go:18-25) of the device transport, driven from a script so that the reference's
 * transport tests (transport_test.go:18-186: send one, send three in order, broadcast) can be replayed without Go.
 * Control messages need no GPU.  script_json: {"peers":[1,2], "ops":[{"op":"send","from":1,"to":2,"msg":{kind_json}},
 * {"op":"broadcast","from":1,"msg":{..}}, {"op":"close","peer":2}, {"op":"pipe","from":1,"layer":3,"to":2}]}.
 * *json_out: {"results":[return code per op], "delivered":{"<peer>":[messages in Deliver() order]}, "addresses":{..}}. */
int dissem_transport_script_json(const char *script_json, char **json_out);

/* LimitRate pacing: TcpTransport.writeWithLimit (transport.go:407-424) sends a fragment in sub-writes of at most
 * 256 KiB, each after rate.Limiter.WaitN (golang.org/x/time/rate v0.14.0, go.mod:8; burst = 256 KiB, bucket full at
 * the start).  This returns the device-side schedule for the `size` bytes at `layer_off` of a layer: pieces that end on
 * 256 KiB boundaries of the layer, each with the earliest time (ns after the send began) at which the reference's
 * limiter would have released the sub-write holding the piece's last byte; the last piece is released when the
 * reference's last sub-write is.  rate <= 0 means unlimited (one piece at 0; the reference's limiter would never
 * finish).  Fills at most `cap` entries of each non-NULL array; *n_out = the number of pieces.  Pure host arithmetic. */
int dissem_pace_schedule(int64_t layer_off, int64_t size, int64_t rate_bytes_per_s, int64_t *piece_off,
                         int64_t *piece_size, int64_t *release_ns, int cap, int *n_out);

#ifdef __cplusplus
}
#endif
#endif /* DISSEM_HOST_H */
