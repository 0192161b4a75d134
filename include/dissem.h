/*
 * dissem.h — C ABI of libdissem.so, the B200-native layer-dissemination engine.
 *
 * This is the drop-in boundary for ONE path of ynishimi/distributed-llm-dissemination: moving layer
 * bytes from the nodes that hold them to the nodes the Assignment names.  Every "node" of the reference
 * is one GPU (or a slice of one GPU) and its layer store (`LayersSrc`, distributor/node.go:172) is a
 * table of HBM buffers.  The entry points are what a cgo `DeviceTransport` implementing
 * `distributor.Transport` (distributor/transport.go:18-25) binds — see INTEGRATION.md for the stub.
 *
 * Conventions: plain C, no CUDA or torch types.  Every function returns 0 on success or a negative
 * DISSEM_E* code; dissem_last_error() returns a thread-local description of the last failure.  All
 * functions are thread-safe (Transport.Send is called from many goroutines, node.go:343, :598).  Nothing
 * here falls back to the CPU: without a usable CUDA device the compute entry points fail with
 * DISSEM_ENODEV.
 *
 * Citations are paths under the reference tree (/root/reference).
 */
#ifndef DISSEM_H
#define DISSEM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DISSEM_VERSION 1

/* Checksum chunk: the reference's write quantum (BucketSize, distributor/transport.go:408). */
#define DISSEM_CHUNK_BYTES 262144

/* dissem_layer_fill kinds.  ZERO is what the reference's dummy layers contain (cmd/config.go:141, :161). */
#define DISSEM_FILL_ZERO 0
#define DISSEM_FILL_PATTERN 1 /* u64 word i of layer L = mix64(seed + (L+1)*0x9E3779B97F4A7C15 + i) */
#define DISSEM_FILL_POISON 2  /* 0xA5 */

#define DISSEM_OK 0
#define DISSEM_EINVAL -1   /* bad argument */
#define DISSEM_ENODEV -2   /* no usable CUDA device / driver */
#define DISSEM_ECUDA -3    /* a CUDA call failed (message in dissem_last_error) */
#define DISSEM_ENOENT -4   /* unknown node / layer / ticket / schedule */
#define DISSEM_EEXIST -5   /* layer already present */
#define DISSEM_ENOMEM -6
#define DISSEM_ERANGE -7   /* offset/size outside the layer; a size beyond a fixed scratch limit (no-allocation entry point) */
#define DISSEM_ENOTSUP -8  /* capability missing on this box (e.g. NVLS multicast) */
#define DISSEM_EPARSE -9   /* malformed config JSON */
#define DISSEM_EPLAN -10   /* the mode cannot serve this placement (e.g. mode 0, leader lacks a layer) */
#define DISSEM_EIO -11     /* a layer file could not be opened or read (disk tier) */

typedef struct dissem_ctx dissem_ctx;
typedef int64_t dissem_ticket;   /* one submitted batch of transfers */
typedef int64_t dissem_sched;    /* one compiled schedule */

/* One layerMsg on the wire (tempLayerInfo, distributor/transport.go:47-54): bytes [off, off+size) of
 * `layer` go from src_node's copy to the same offsets of dst_node's copy; total = full layer size
 * (LayerSize = size, TotalSize = total, Offert = off).  Modes 0/1 send whole layers (off = 0,
 * size = total; node.go:363, :1478); mode 3 sends fragments (node.go:1604-1606). */
/* dissem_xfer.flags: who executes the transfer.  Default (0): the GPU holding the source pushes.  PULL: the
 * destination's GPU loads from the source instead (its process needs the source layer imported). */
#define DISSEM_XFER_PULL 1
/* FILE: the source is src_node's DISK copy of the layer (dissem_layer_attach_file; LayerSrc with Location = DiskLayer,
 * transport.go:351-367), not its HBM copy.  Such transfers are streamed by dissem_schedule_submit / dissem_send_file. */
#define DISSEM_XFER_FILE 2
/* WAIT_SRC: the source copy of this transfer is itself a destination of an earlier round of the same dissemination (a
 * relay hop).  The kernel that executes the transfer waits, unit by unit, until the source copy's received-bytes
 * counter has reached off + size before it reads — the counter its upstream sender releases after every unit — so
 * the hops of a relay chain order themselves ON THE DEVICE even when they run in different processes, with no host
 * barrier between rounds.  Preconditions (the caller's): the source copy is filled in ascending offset order by ONE
 * upstream sender whose rounds run one after another (what dissem_plan_relay emits), and its counter was zeroed
 * (dissem_layer_reset, then a barrier between the processes) before the dissemination started.  The wait is bounded
 * (~4 s); "wait_timeouts" (dissem_get_option) counts units that gave up.  dissem_plan_relay sets the flag on hops >= 1. */
#define DISSEM_XFER_WAIT_SRC 4
/* Bits 8..23 of flags: the transfer's ROUND.  Within one dissem_schedule_* call the transfers of round r start
 * only after every transfer of round r-1 has completed on every GPU of this process (device-side event waits, no
 * host round trip).  A relay plan (dissem_plan_relay) forwards in round r the piece that arrived in round r-1.
 * Across processes rounds are not synchronised by the engine: either the transfers carry DISSEM_XFER_WAIT_SRC (device-side
 * waits on the received counters; submit all rounds in one call), or submit one round per call and separate them yourself. */
#define DISSEM_XFER_ROUND_SHIFT 8
#define DISSEM_XFER_ROUND(r) (((r) & 0xFFFF) << DISSEM_XFER_ROUND_SHIFT)

typedef struct dissem_xfer {
    int32_t src_node;
    int32_t dst_node;
    int32_t layer;
    int32_t flags;   /* DISSEM_XFER_* */
    int64_t off;
    int64_t size;
    int64_t total;
} dissem_xfer;

/* ---- library ------------------------------------------------------------------------------------- */
int dissem_version(void);
const char *dissem_last_error(void);
/* Number of CUDA devices visible to this process (0 when there is none; never an error). */
int dissem_device_count(void);

/* ---- context: stands for NewTcpTransport + the AddrRegistry (transport.go:56-95) ------------------ */
/* device_of_node[i] >= 0: node i lives in THIS process on that CUDA device (several nodes may share a
 * device: the 1-GPU configuration maps every node to device 0).  device_of_node[i] == -1: node i lives in
 * another process (one process per GPU); its layers become reachable through dissem_layer_import.
 * Peer access is enabled between all local devices. */
int dissem_init(int n_nodes, const int *device_of_node, dissem_ctx **out);
int dissem_close(dissem_ctx *ctx); /* Transport.Close, transport.go:474-491 */
int dissem_n_nodes(dissem_ctx *ctx);

/* ---- layer store: LayersSrc / LayerSrc (node.go:172, :200-211), CreateInmemLayer (config.go:159-171) - */
/* Allocates an HBM buffer for `layer` on a local node.  The buffer carries a trailer with the layer's
 * per-chunk checksum table and its received-bytes counter (both zeroed). */
int dissem_layer_alloc(dissem_ctx *ctx, int node, int layer, int64_t size);
int dissem_layer_free(dissem_ctx *ctx, int node, int layer);
/* Set aside `count` buffers for layers of `size` bytes on `node`'s GPU.  A later dissem_layer_alloc of that size on that
 * GPU (any of its nodes) takes one of them instead of calling cudaMalloc, which is most of what a small Transport.Send
 * costs (~0.3 ms against ~0.04 ms for moving 1 MiB; the reference pays make([]byte, LayerSize) inside its clock too,
 * transport.go:201).  A leader that knows the Assignment reserves before it starts the clock.  Unused buffers are
 * released by dissem_close.  Not used for VMM-backed layers ("alloc_vmm"). */
int dissem_layer_reserve(dissem_ctx *ctx, int node, int64_t size, int count);
/* size_out / device_ptr_out may be NULL.  DISSEM_ENOENT if the node does not hold the layer. */
int dissem_layer_info(dissem_ctx *ctx, int node, int layer, int64_t *size_out, void **device_ptr_out);
/* Synthesise contents on the GPU (DISSEM_FILL_*); asynchronous on the node's stream. */
int dissem_layer_fill(dissem_ctx *ctx, int node, int layer, int kind, uint64_t seed);
/* Zero the checksum table and the received counter (a destination does this before a new run). */
int dissem_layer_reset(dissem_ctx *ctx, int node, int layer);
/* Host <-> HBM.  `host` may be any host memory; pinned memory from dissem_host_alloc makes the copy
 * asynchronous and full-speed.  write is ordered before every later transfer that reads or writes the layer, whichever
 * GPU executes it (the same holds for alloc, fill and reset: each buffer carries an event the executing stream waits
 * for).  read waits for the transfers submitted so far that write THIS layer (not for the rest of the device), without
 * holding the context's lock: other threads keep sending meanwhile.  Do not free a layer that another thread reads. */
int dissem_layer_write(dissem_ctx *ctx, int node, int layer, const void *host, int64_t off, int64_t size);
int dissem_layer_read(dissem_ctx *ctx, int node, int layer, void *host, int64_t off, int64_t size);
int dissem_host_alloc(void **out, int64_t size); /* pinned, GPU-mapped host memory for layer bytes */
int dissem_host_free(void *p);
/* The layer stays where the reference keeps it: in host memory (LayerSrc.InmemData, node.go:200-211; the host tier of
 * SURVEY §8(f).4).  Registers `size` bytes of pinned, GPU-mapped host memory (from dissem_host_alloc; 16-byte aligned;
 * owned by the caller, who keeps it alive) as `node`'s copy of `layer` WITHOUT copying it: transfers that name it as
 * their source are read by the copy kernel straight over PCIe, so upload and dissemination are one pass.  Such a
 * copy can only be a source: it has no checksum table or received counter and cannot be exported. */
int dissem_layer_attach_host(dissem_ctx *ctx, int node, int layer, void *pinned, int64_t size);
/* The disk tier (DiskLayer / `-s`, cmd/config.go:133-157; sent with a section reader over the file,
 * transport.go:351-367).  Records that `node` holds `layer` as `size` bytes of the file `path` starting
 * at `file_off`.  The file is opened read-only and stays open until dissem_layer_detach_file / dissem_close; nothing is
 * read here.  This is independent of the node's HBM copy (dissem_layer_alloc), which may or may not exist: a transfer
 * names the disk copy with DISSEM_XFER_FILE, exactly as a layerMsg carries its LayerSrc.Meta.Location.  A transfer from
 * the disk copy is staged through a ring of pinned host buffers (4 x 16 MiB per GPU, allocated on first use): the
 * calling thread reads the next piece of the file while the copy kernel of the node's GPU moves the previous pieces
 * over PCIe into the destination and folds them into its checksum table.  The node may be its own destination
 * (loading its disk copy into HBM). */
int dissem_layer_attach_file(dissem_ctx *ctx, int node, int layer, const char *path, int64_t file_off, int64_t size);
int dissem_layer_detach_file(dissem_ctx *ctx, int node, int layer);
/* size_out: the attached size; DISSEM_ENOENT when the node has no disk copy of the layer. */
int dissem_layer_file_info(dissem_ctx *ctx, int node, int layer, int64_t *size_out);

/* ---- one process per GPU: make a local layer reachable from peer processes -------------------------- */
/* handle_out: 64 opaque bytes to ship over the control plane (e.g. inside the announceMsg, message.go:31-41). */
int dissem_layer_export(dissem_ctx *ctx, int node, int layer, uint8_t handle_out[64]);
/* Map a remote node's layer (allocated by its owner process) so that local nodes can push into it. */
int dissem_layer_import(dissem_ctx *ctx, int node, int layer, int64_t size, const uint8_t handle[64]);

/* The same for layers allocated with the VMM API ("alloc_vmm" = 2: shareable), which NVLS multicast needs: the handle
 * is a POSIX file descriptor that the host side passes to the peer process (SCM_RIGHTS on a unix socket) and closes. */
int dissem_layer_export_fd(dissem_ctx *ctx, int node, int layer, int *fd_out);
int dissem_layer_import_fd(dissem_ctx *ctx, int node, int layer, int64_t size, int fd);
/* NVLS multicast with one process per GPU, for the transfers of `layer` from src_node to dst_nodes[0..n_dst):
 *   1. the source's process:      dissem_mcast_create -> fd, sent to every destination's process;
 *   2. each destination process:  dissem_mcast_join(fd) with my_node = its member node;
 *   3. after ALL members have joined (a barrier of the host side), every member process: dissem_mcast_bind.
 * A later dissem_schedule_* call in the source's process that carries those transfers then uses the object
 * (subject to the "mcast" option).  Every member's copy, the source's included, must be a shareable VMM layer, and
 * the source's process must have imported the destinations' copies (dissem_layer_import_fd). */
int dissem_mcast_create(dissem_ctx *ctx, int layer, int src_node, const int *dst_nodes, int n_dst, int *fd_out);
int dissem_mcast_join(dissem_ctx *ctx, int layer, int src_node, const int *dst_nodes, int n_dst, int my_node, int fd);
int dissem_mcast_bind(dissem_ctx *ctx, int layer, int src_node, const int *dst_nodes, int n_dst);

/* ---- the hot path: TcpTransport.Send layer branch + sendLayerMsg + writeWithLimit on the sender
 *      (transport.go:258-275, :308-373, :407-424) and handleIncomingMsg layer branch on the receiver
 *      (transport.go:127-224), as one push kernel per source GPU with the checksum fused in ----------- */
/* Asynchronous.  The destination layer must exist (alloc or import).  Only transfers whose src_node is
 * local to this process are executed; others are skipped (their owner process runs them). */
int dissem_send(dissem_ctx *ctx, int src_node, int dst_node, int layer, int64_t off, int64_t size,
                int64_t total, dissem_ticket *out);
/* The same from src_node's disk copy (DISSEM_XFER_FILE).  Returns once the last piece of the file has been read and
 * its copy enqueued (sendLayerMsg blocks in io.Copy likewise, transport.go:364); the ticket completes when the bytes
 * have landed.  Transfers from disk cannot be part of a compiled schedule (dissem_schedule_compile: DISSEM_ENOTSUP)
 * or of a round other than 0; in dissem_schedule_submit they are streamed first, one after another, and the rest of
 * the plan is launched behind them. */
int dissem_send_file(dissem_ctx *ctx, int src_node, int dst_node, int layer, int64_t off, int64_t size,
                     int64_t total, dissem_ticket *out);
/* A whole plan at once (what LeaderNode.sendLayers / RetransmitLeaderNode.sendLayers / dispatchJob issue
 * one goroutine at a time, node.go:326-352, :554-626, :1237-1288): compiled into one launch per source
 * GPU.  dissem_schedule_submit = compile + launch + free. */
int dissem_schedule_submit(dissem_ctx *ctx, const dissem_xfer *xfers, int n, dissem_ticket *out);
int dissem_schedule_compile(dissem_ctx *ctx, const dissem_xfer *xfers, int n, dissem_sched *out);
/* A compiled schedule may be launched again and again, also with earlier launches still in flight: the checksum entries
 * it fills are stored or re-zeroed by the launch itself (partially covered chunks shared by several fragments of the
 * schedule get one owner; where those fragments run on different GPUs of this process, the launch first waits for the
 * other GPUs' previous launch).  Two limits: a chunk the schedule covers only in part keeps accumulating (it also takes
 * bytes from other submissions), and with ONE PROCESS PER GPU a chunk shared by fragments of different processes cannot
 * be re-zeroed by anybody — run dissem_plan_align on the plan first (dissem_plan_optimize does), or dissem_layer_reset the
 * destinations between launches.  The received counters are monotonic by design. */
int dissem_schedule_launch(dissem_ctx *ctx, dissem_sched sched, dissem_ticket *out);
int dissem_schedule_free(dissem_ctx *ctx, dissem_sched sched);
/* units_out: work units of the compiled schedule; launches_out: kernel launches one dissem_schedule_launch
 * issues; bytes_out: bytes the local sources deliver. */
int dissem_schedule_info(dissem_ctx *ctx, dissem_sched sched, int64_t *units_out, int *launches_out,
                         int64_t *bytes_out);
/* Per-link transfer direction for a plan (no counterpart in the reference, which has one TCP direction).
 * Measured on B200/NVLink 5 (profiles/): a PUSH stream carries ~710 GB/s whatever the reverse direction does;
 * a PULL stream carries ~785 GB/s while the reverse direction of the same ports is idle but only ~640-675 GB/s
 * when it is busy (read requests share it).  For each transfer this sets or clears DISSEM_XFER_PULL in place:
 * pull when the data load on the reverse ports, max(ingress of the source GPU, egress of the destination GPU),
 * is below half of the forward load max(egress of source, ingress of destination); transfers inside one GPU
 * are left as pushes.  gpu_of_node[i]: any id of the GPU that holds node i (e.g. the rank).  Pure host logic;
 * every process of a job computes the same answer from the same plan. */
int dissem_plan_directions(dissem_xfer *xfers, int n, const int *gpu_of_node, int n_nodes);

/* Fragment cuts on checksum-chunk boundaries (no counterpart in the reference, whose fragments are never checked).
 * flow.go cuts a layer wherever the max-flow solution falls (flow.go:193-212), so two senders usually share one
 * 256 KiB checksum chunk at every cut; their partial sums then meet in one table entry, which costs a zeroing owner,
 * cross-GPU ordering on every relaunch, and cannot be re-zeroed at all when the two senders are different processes.
 * Where consecutive fragments of one (destination, layer) are a cut of one contiguous range, this moves each internal
 * boundary DOWN to a multiple of DISSEM_CHUNK_BYTES (the later fragment's sender, which holds the whole layer as every
 * LayerSrc does, sends up to 256 KiB more; a fragment that shrinks to nothing is dropped).  In place; *n_out = the
 * number of transfers left.  Coverage and destinations are unchanged, so the delivered store is identical.
 * dissem_plan_optimize applies it first.  Pure host logic. */
int dissem_plan_align(dissem_xfer *xfers, int n, int *n_out);

/* Several nodes on one GPU (the 1-, 2- and 4-GPU runs of an 8-node config).  The reference's planners know nodes, not
 * GPUs: they may fetch a layer over NVLink although another node on the destination's own GPU holds it.  Given who holds
 * which layer in HBM (holder_nodes[i] holds holder_layers[i], whole), every transfer that crosses GPUs while a holder
 * shares the destination's GPU is re-sourced from that holder (an HBM-local copy, ~8x the link rate, and no port load);
 * touching fragments that end up with the same source are merged.  In place; *n_out = transfers left (the order
 * changes: sorted by destination, layer, source, offset).  The delivered store is unchanged.  Pure host logic. */
int dissem_plan_localize(dissem_xfer *xfers, int n, const int *gpu_of_node, int n_nodes, const int *holder_nodes,
                         const int *holder_layers, int n_holders, int *n_out);

/* Relay / pipelined forwarding (generalises the reference's tee-pipe, transport.go:145-196, and mode 1's peer
 * retransmission, node.go:1462-1484, to piece granularity).  Rewrites a plan: transfers of the same bytes from one
 * source to k destinations become a chain  source -> d0 -> d1 -> ...  cut into `pieces` pieces (boundaries at multiples
 * of DISSEM_CHUNK_BYTES); hop h sends piece p in round p+h, i.e. a GPU forwards a piece one round after it arrived,
 * so the source sends every byte once and all links work at the same time.  Every other transfer is cut into the
 * same pieces over rounds 0..pieces-1.  PULL flags are cleared (run dissem_plan_directions on the result).
 * *out is malloc'ed: release it with dissem_free (include/dissem_host.h).  Pure host logic. */
int dissem_plan_relay(const dissem_xfer *in, int n, int pieces, dissem_xfer **out, int *n_out);

/* The scheduling decisions above in one call, for a leader that just wants the fastest execution of its plan:
 * relay chains when the plan has fan-out and the busiest NVLink port gets at least 10 % less to carry that way
 * (pieces of about 64 MiB; the port load is scaled by the pipeline-fill factor (rounds / pieces)), then
 * dissem_plan_directions on whatever was chosen.  *relay_pieces_out (may be NULL): pieces used, 0 = no relay.
 * Disk sources (DISSEM_XFER_FILE): a layer that several nodes need keeps ONE transfer from the file (to the first of them,
 * round 0); the other destinations are fed from that node's HBM, and every transfer that is not read from disk is moved
 * one round later, so that a caller who submits one round per call (several processes) keeps the order too.
 * *out is malloc'ed (dissem_free).  Pure host logic. */
int dissem_plan_optimize(const dissem_xfer *in, int n, const int *gpu_of_node, int n_nodes, dissem_xfer **out, int *n_out,
                         int *relay_pieces_out);
/* CUDA ordinal of the device a local node lives on; -1 for a node of another process. */
int dissem_node_device(dissem_ctx *ctx, int node, int *device_out);

/* Block until the batch is complete (all bytes visible at their destinations).  ms_out (may be NULL):
 * device time of the batch, max over the source GPUs involved.  The ticket is consumed. */
int dissem_wait(dissem_ctx *ctx, dissem_ticket ticket, double *ms_out);
/* As dissem_wait; copy_ms_out (may be NULL) is the device time of the copy+checksum kernel alone (the batch
 * also runs a tiny kernel that zeroes the checksum entries it is about to fill). */
int dissem_wait_detail(dissem_ctx *ctx, dissem_ticket ticket, double *ms_out, double *copy_ms_out);
/* Non-blocking.  Every ticket must eventually be handed to dissem_wait (which consumes it) or, once dissem_poll has
 * reported it done, to dissem_ticket_release; otherwise its events stay allocated until dissem_close. */
int dissem_poll(dissem_ctx *ctx, dissem_ticket ticket, int *done_out);
int dissem_ticket_release(dissem_ctx *ctx, dissem_ticket ticket);   /* DISSEM_EINVAL while the batch is in flight */

/* ---- delivery verification (no counterpart in the reference, which never checks content) ----------- */
/* Per-chunk checksums accumulated by the copy pass at the DESTINATION copy of the layer:
 * out[2c] = A_c, out[2c+1] = B_c, n_chunks = ceil(size / DISSEM_CHUNK_BYTES).  Definition in DESIGN.md §4. */
int dissem_layer_checksums(dissem_ctx *ctx, int node, int layer, uint64_t *out, int64_t n_chunks);
/* The same for n local (node, layer) pairs at once: the tables back to back in `out`, in the order given, total_chunks =
 * the sum of their chunk counts.  One gather kernel and one device-to-host copy per GPU, instead of a synchronising copy
 * per layer (what a leader reads after a batch: one table per pair of its Assignment). */
int dissem_layers_checksums(dissem_ctx *ctx, const int *nodes, const int *layers, int n, uint64_t *out,
                            int64_t total_chunks);
/* Independent read-only pass over the layer as it now sits in HBM; same format.  Its scratch comes from an arena set
 * up by dissem_init (no allocation per call); layers beyond 128 Ki chunks (32 GiB) fail with DISSEM_ERANGE. */
int dissem_layer_verify(dissem_ctx *ctx, int node, int layer, uint64_t *out, int64_t n_chunks);
/* Bytes received so far for the layer (the counter FlowRetransmitReceiverNode keeps in
 * layerSrc.DataSize, node.go:1542; the ack condition is received == total, node.go:1555).  Monotonic: it is zeroed by
 * dissem_layer_alloc / dissem_layer_reset only, so launching a schedule twice counts twice.  An NVLS multicast stream
 * also adds to the SOURCE's own counter and checksum table (the source's copy is a member of the multicast object). */
int dissem_layer_received(dissem_ctx *ctx, int node, int layer, int64_t *bytes_out);

/* ---- for callers that cannot allocate device memory (no counterpart in the reference) ----------------- */
/* Same per-chunk checksum pass as dissem_layer_verify, over any 16-byte-aligned device buffer the caller
 * already owns, WITHOUT allocating or freeing anything: no cudaMalloc, no cudaFree, no pooled allocator.
 * It exists for hosts where an allocation at this point is not an option — a serving process that has given
 * all remaining HBM to its own pool (cudaMalloc would fail or fragment it), or a latency-critical caller for
 * whom cudaMalloc/cudaFree's driver locks and device-wide wait would stall every other stream.
 * The kernel's scratch (work-unit counters, one descriptor, the table it accumulates into) lives in a fixed
 * array inside the library image, about 129 KiB of HBM on each GPU that runs this library, held until the
 * process exits.  Limits and rules:
 *   - at most DISSEM_NOALLOC_MAX_CHUNKS chunks per call (2 GiB); larger sizes fail with DISSEM_ERANGE;
 *   - all calls on a GPU share the one array; each call clears the part it uses first, so results do not depend
 *     on earlier calls, but two calls whose GPU work may overlap (other streams or threads) corrupt each other:
 *     run them one after another;
 *   - calls on one GPU are serialised on the host (a per-GPU lock held until the stream has finished); calls on
 *     different GPUs run side by side, each GPU has its own copy of the array;
 *   - the first call on each GPU resolves the array's device address; make that first call at start-up, before
 *     any stream capture begins on that GPU.
 * device: CUDA ordinal that owns `data`.  data: device pointer, size bytes.  out: HOST array of 2*n_chunks
 * u64 (A_c, B_c), n_chunks = ceil(size / DISSEM_CHUNK_BYTES).  stream: a cudaStream_t, or NULL for the default
 * stream; the call returns after the stream has finished this work, with `out` filled. */
#define DISSEM_NOALLOC_MAX_CHUNKS 8192
int dissem_checksum_noalloc(int device, const void *data, int64_t size, uint64_t *out, int64_t n_chunks,
                            void *stream);

/* ---- tuning / introspection -------------------------------------------------------------------------- */
/* key: "ctas_per_sm" (0 = occupancy-derived), "variant" (0 = register path, 1 = TMA bulk-staged path),
 * "reg_cfg" (register-path shape, threads x 16-byte loads in flight: 0 256x4, 1 512x4 (default), 2 256x8,
 * 3 512x8, 4 1024x4, 5 1024x2, 6 512x2; 7/8/9 = 512x2, 512x4, 256x4 with 32-byte loads and stores),
 * "tma_cfg" (TMA ring shape, stage bytes x stages: 0 16Kx6 (default), 1 32Kx3, 2 8Kx12, 3 16Kx4, 4 8Kx6, 5 4Kx12;
 *  6 / 7 = 16Kx6 with 8 / 14 checksum warps, 8 = 8Kx6 with 14: probe shapes for multicast units),
 * "alloc_vmm" (1 = dissem_layer_alloc uses the VMM API so that layers can be bound to NVLS multicast objects;
 *  2 = the same, shareable with other processes through dissem_layer_export_fd),
 * "mcast" (transfers of the same bytes from one source to several GPUs can go as ONE NVLS multicast store stream
 *  when every buffer involved is VMM-backed and local.  1 (default) = do so when a cost model built on the measured
 *  port rates says the plan gets shorter (i.e. it is bound by a source's egress); 3 = always; 0 = never (unicast
 *  fan-out); 2 = always, and single-destination transfers to another GPU too (a testing aid)),
 * "wait_src" (0 (default) = DISSEM_XFER_WAIT_SRC flags are ignored: inside one process round events order relay hops, and
 *  a relaunch without dissem_layer_reset would find the counters already past their thresholds; 1 = honour them: the
 *  mode for one process per GPU, see DISSEM_XFER_WAIT_SRC.  A schedule compiled with flagged hops while the option is on
 *  is ordered by those waits ALONE, also between the GPUs of one process: no cross-GPU event separates its rounds, so
 *  the preconditions of DISSEM_XFER_WAIT_SRC are the caller's there too), read-only "wait_timeouts",
 * "mcast_variant" (kernel for schedule parts that carry multicast units: 0 (default) = register path whatever "variant"
 *  says, 1 = follow "variant"; transfers whose offset or size is not a multiple of 16 bytes are never multicast),
 * read-only: "mcast_supported" (do all local GPUs support multicast objects), "mcast_groups" (multicast
 *  descriptors compiled so far),
 * "unit_bytes" (0 = auto; else power of two in [4096, 262144]),
 * "file_threads" (1..64, default 8: threads that read one staged piece of a layer file in parallel, disk tier),
 * "pull" (0 = follow each transfer's DISSEM_XFER_PULL flag (default); 1 = pull every transfer: a process then
 *  executes the transfers whose DESTINATION node is local and needs the remote SOURCE layers imported;
 *  2 = push every transfer whatever its flag). */
int dissem_set_option(dissem_ctx *ctx, const char *key, int64_t value);
int dissem_get_option(dissem_ctx *ctx, const char *key, int64_t *value_out);

#ifdef __cplusplus
}
#endif
#endif /* DISSEM_H */
