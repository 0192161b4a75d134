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
 * opts_json (may be NULL): {"seed":S,"policy":P,"time_scale":T,"flags":F,"fill":K,"batched":bool,"verify":bool}
 *   batched=false: one Transport.Send per layerMsg, as the reference issues them;
 *   batched=true : the leader hands the whole plan to dissem_schedule_submit (one launch per source GPU).
 * *json_out: {"time_to_deliver_ms":..,"device_ms":..,"bytes":..,"n_xfers":..,"acks":..,"verified":bool,...} */
int dissem_run_json(const char *config_json, int mode, const int *node_device, const char *opts_json,
                    char **json_out);

/* Wire format of the control plane (message.go:273-301): encode one message kind to the
 * envelope {"type":T,"src":"<id>","payload":{...}} and decode it back; used by the parity tests.
 * kind_json: {"kind":"announce|ack|retransmit|flowRetransmit|startup|simple|layerHeader", fields...}. */
int dissem_msg_roundtrip_json(const char *kind_json, char **wire_out, char **decoded_out);

#ifdef __cplusplus
}
#endif
#endif /* DISSEM_HOST_H */
