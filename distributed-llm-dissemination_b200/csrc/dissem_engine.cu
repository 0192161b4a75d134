// Host engine behind the C ABI of include/dissem.h: the HBM layer store, the schedule compiler and the
// launch/wait machinery.  Reference citations are in the header; this file only implements them.
#include <cuda_runtime.h>
#include <errno.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/dissem.h"
#include "dissem_engine.h"
#include "dissem_kernels.cuh"

namespace dissem {

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;

int set_error(int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
const char *last_error() { return g_err.c_str(); }

#define CUDA_TRY(expr)                                                                                  \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess)                                                                          \
            return set_error(_e == cudaErrorNoDevice || _e == cudaErrorInsufficientDriver ? DISSEM_ENODEV \
                                                                                          : DISSEM_ECUDA, \
                             "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// ------------------------------------------------------------------------------------------------
// layer buffers
// ------------------------------------------------------------------------------------------------
// [ layer bytes | pad to 4 KiB | received counter (256 B slot) | checksum table: 16 B per chunk ]
static inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

int64_t LayerBuf::n_chunks() const { return (size + kChunkBytes - 1) / kChunkBytes; }
int64_t LayerBuf::trailer_off() const { return align_up(std::max<int64_t>(size, 1), 4096); }
// rounded to 2 MiB so that every layer owns its pages (an IPC handle exports whole pages)
int64_t LayerBuf::alloc_bytes() const { return align_up(trailer_off() + 256 + align_up(std::max<int64_t>(n_chunks(), 1) * 16, 256), 2 << 20); }
unsigned long long *LayerBuf::received() const { return reinterpret_cast<unsigned long long *>(base + trailer_off()); }
unsigned long long *LayerBuf::csum() const { return reinterpret_cast<unsigned long long *>(base + trailer_off() + 256); }
int64_t LayerBuf::trailer_bytes() const { return alloc_bytes() - trailer_off(); }

// Disk tier: one staging ring per GPU.  A slot holds one piece of a layer file between pread() and the copy kernel
// that moves it on; its event tells when that kernel has finished with the slot (and with the slot's descriptor).
constexpr int kStageSlots = 4;
constexpr int64_t kStageBytes = 16ll << 20;
struct StageRing {
    uint8_t *slot[kStageSlots] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t done[kStageSlots] = {nullptr, nullptr, nullptr, nullptr};
    bool in_flight[kStageSlots] = {false, false, false, false};
    XferDesc *d_descs = nullptr;      // one descriptor per slot
    int next = 0;
};

// A node's disk copy of a layer (LayerSrc.Fp with Location = DiskLayer, cmd/config.go:147-156)
struct FileSrc {
    int fd = -1;
    int64_t file_off = 0, size = 0;
    std::string path;
};

struct DeviceState {
    int dev = -1;
    int n_sm = 0;
    cudaStream_t stream = nullptr;
    SchedCounters *counters = nullptr;   // kernel scratch: ticket + drained-CTA count
    // Descriptor tables of small schedules (one Transport.Send = one descriptor) come from this pool instead of a
    // cudaMalloc/cudaFree pair per message (cudaFree waits for the whole device).
    XferDesc *desc_pool = nullptr;
    std::vector<int> free_slots;
    StageRing ring;
    // Recorded after every copy kernel this engine enqueues on `stream`.  A launch whose zeroing phase may wipe a
    // checksum entry that ANOTHER GPU's previous launch is still adding into waits for it first (launch_sched).
    cudaEvent_t last_end = nullptr;
    // Read-back side (verify pass, checksum tables, layer reads): its own stream, so that a reader never queues behind
    // transfers other threads enqueue meanwhile, and a scratch arena (device + pinned host) allocated at dissem_init, so
    // that no call on this path allocates or frees device memory.  arena_mu serialises the arena's users per GPU.
    cudaStream_t rd_stream = nullptr;
    uint8_t *arena_dev = nullptr, *arena_host = nullptr;
    size_t arena_bytes = 0;
    std::shared_ptr<std::mutex> arena_mu = std::make_shared<std::mutex>();
    std::shared_ptr<std::mutex> ring_mu = std::make_shared<std::mutex>();   // one file transfer streams through a ring at a time
};
constexpr size_t kArenaBytes = 2u << 20;      // verify scratch: descriptor + table of 128 Ki chunks (a 32 GiB layer)
constexpr int kPoolSlots = 256;
constexpr int kSlotDescs = 16;

struct SchedPart {
    int dev_slot = -1;
    int pool_slot = -1;               // >= 0: d_descs points into the device's descriptor pool
    int round = 0;
    bool host_src = false;            // reads pinned host memory: runs the register kernel
    bool needs_zero = true;           // false: every descriptor stores whole-chunk sums and owns no partial chunk -> no zeroing kernel
    bool has_mc = false;              // carries NVLS multicast units: runs the register kernel (multimem.st straight from the
                                      // loaded registers: 570 GB/s against 533 GB/s through the shared-memory ring, any ring
                                      // shape: profiles/r02_n2_mcast_probe.jsonl)
    XferDesc *d_descs = nullptr;
    int n_descs = 0;
    long long total_units = 0;
    int64_t bytes = 0;
    // (node, layer) of local buffers this part reads or writes that live on ANOTHER GPU: their `ready` events are
    // waited for before the part's kernels start
    std::vector<std::pair<int, int>> touch;
    // (node, layer) of every local destination buffer this part writes: remembered in the buffer (LayerBuf::writers) so
    // that a later read of it waits for exactly the streams that may still be writing
    std::vector<std::pair<int, int>> writes;
};

struct Sched {
    std::vector<SchedPart> parts;
    int unit_shift = kChunkShift;
    int64_t units = 0, bytes = 0;
    // some checksum entry that this schedule zeroes (an owned, partially covered chunk) is added into by kernels of
    // several GPUs: a relaunch must not zero it while another GPU's previous launch is still running
    bool cross_dev_shared = false;
    // "wait_src" is on and the schedule carries relay hops that wait for their source on the device: rounds are then NOT
    // separated by cross-GPU events (each GPU's stream runs its rounds in order; a hop waits only for its own upstream piece)
    bool flag_ordered = false;
};

struct TicketPart {
    int dev_slot;
    cudaEvent_t start, mid, end;      // start | zero_csum | mid | copy_csum | end
};
struct Ticket {
    std::vector<TicketPart> parts;
};

}  // namespace dissem

using namespace dissem;

struct dissem_ctx {
    std::mutex mu;
    int n_nodes = 0;
    std::vector<int> node_dev_slot;                    // node -> index into devs, or -1 (remote)
    std::vector<std::map<int, LayerBuf>> layers;       // node -> layer -> buffer
    std::vector<std::map<int, FileSrc>> files;         // node -> layer -> disk copy
    std::vector<DeviceState> devs;
    std::map<int64_t, Sched> scheds;
    std::map<int64_t, Ticket> tickets;
    int64_t next_id = 1;
    // NVLS multicast objects, keyed by (layer, team of nodes: source first, then destinations ascending)
    std::map<std::pair<int, std::vector<int>>, McObject> mcasts;
    int64_t mcast_groups_compiled = 0;
    std::string mcast_note;
    std::vector<cudaEvent_t> event_pool[64];   // per device ordinal: events are reused across launches
    // dissem_layer_reserve: destination buffers allocated ahead of time, keyed by (device ordinal, allocation bytes)
    std::map<std::pair<int, int64_t>, std::vector<uint8_t *>> reserved;
    // options
    int64_t opt_ctas_per_sm = 0, opt_variant = 1, opt_unit_bytes = 0, opt_reg_cfg = 1, opt_pull = 0, opt_tma_cfg = 0, opt_alloc_vmm = 0, opt_mcast = 1, opt_file_threads = 8, opt_mcast_variant = 0, opt_wait_src = 0;
};

namespace dissem {

static int check_node(dissem_ctx *ctx, int node) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    if (node < 0 || node >= ctx->n_nodes) return set_error(DISSEM_ENOENT, "node %d out of range [0,%d)", node, ctx->n_nodes);
    return 0;
}

static int find_layer(dissem_ctx *ctx, int node, int layer, LayerBuf **out) {
    int rc = check_node(ctx, node);
    if (rc) return rc;
    auto it = ctx->layers[node].find(layer);
    if (it == ctx->layers[node].end()) return set_error(DISSEM_ENOENT, "node %d holds no layer %d", node, layer);
    *out = &it->second;
    return 0;
}

static int local_dev(dissem_ctx *ctx, int node, DeviceState **out) {
    int rc = check_node(ctx, node);
    if (rc) return rc;
    int slot = ctx->node_dev_slot[node];
    if (slot < 0) return set_error(DISSEM_EINVAL, "node %d is not local to this process", node);
    *out = &ctx->devs[slot];
    return 0;
}

// Record `lb.ready` on the owner's stream (ds is current).  Called after the owner enqueued something that writes the
// buffer; kernels of other GPUs' streams wait for the event (launch_sched), the owner's own stream is ordered anyway.
static cudaError_t mark_ready(LayerBuf &lb, DeviceState &ds) {
    if (!lb.ready) {
        cudaEvent_t ev = nullptr;
        cudaError_t e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
        lb.ready = ev;
    }
    return cudaEventRecord(static_cast<cudaEvent_t>(lb.ready), ds.stream);
}

static void drop_ready(LayerBuf &lb) {
    if (lb.ready) cudaEventDestroy(static_cast<cudaEvent_t>(lb.ready));
    lb.ready = nullptr;
}

static void free_sched(dissem_ctx *ctx, Sched &s) {
    for (auto &p : s.parts) {
        if (p.pool_slot >= 0) ctx->devs[p.dev_slot].free_slots.push_back(p.pool_slot);
        else if (p.d_descs) {
            cudaSetDevice(ctx->devs[p.dev_slot].dev);
            cudaFree(p.d_descs);
        }
    }
    s.parts.clear();
}

static int pick_unit_shift(dissem_ctx *ctx, const std::vector<std::vector<dissem_xfer>> &by_dev) {
    if (ctx->opt_unit_bytes > 0) {
        int s = 12;
        while ((1ll << s) < ctx->opt_unit_bytes) ++s;
        return std::min(std::max(s, 12), kChunkShift);
    }
    // auto: 256 KiB units unless that leaves the persistent grid starved (small layers: the README case)
    int shift = kChunkShift;
    for (; shift > 14; --shift) {
        bool starved = false;
        for (size_t slot = 0; slot < by_dev.size(); ++slot) {
            if (by_dev[slot].empty()) continue;
            long long units = 0;
            for (auto &x : by_dev[slot]) units += units_of(x.off, x.size, shift);
            if (units < 4ll * ctx->devs[slot].n_sm) starved = true;
        }
        if (!starved) break;
    }
    return shift;
}

// NVLS: one multicast object per (layer, source, destination set), over the layer buffers of the team's GPUs.
// The source GPU must be a member to address the object, so its own copy of the layer is bound as well: the
// multicast store rewrites the source's bytes with themselves (and bumps its own trailer, which nobody reads).
// Returns NULL when the group cannot be multicast (capability or layout); the caller then sends unicast copies.
static McObject *get_or_make_mcast(dissem_ctx *ctx, int layer, int src_node, const std::vector<int> &dst_nodes) {
    std::vector<int> team{src_node};
    std::vector<int> sorted_dsts(dst_nodes);
    std::sort(sorted_dsts.begin(), sorted_dsts.end());
    team.insert(team.end(), sorted_dsts.begin(), sorted_dsts.end());
    auto key = std::make_pair(layer, team);
    auto it = ctx->mcasts.find(key);
    if (it != ctx->mcasts.end()) return it->second.valid() ? &it->second : nullptr;
    McObject mc;                                   // cached even when invalid: do not retry a failing set-up per compile
    std::vector<int> devices;
    std::vector<const VmmBlock *> blocks;
    bool ok = true;
    for (int node : team) {
        int slot = ctx->node_dev_slot[node];
        auto li = ctx->layers[node].find(layer);
        if (slot < 0 || li == ctx->layers[node].end() || !li->second.vmm) { ok = false; break; }
        int dev = ctx->devs[slot].dev;
        if (std::find(devices.begin(), devices.end(), dev) != devices.end()) { ok = false; break; }   // one block per GPU
        devices.push_back(dev);
        blocks.push_back(&li->second.blk);
    }
    std::string why;
    if (ok && !mcast_supported(devices, &why)) ok = false;
    if (ok) {
        why = mcast_create(devices, blocks, &mc);
        if (!why.empty()) ok = false;
    }
    if (!ok) {
        ctx->mcast_note = why.empty() ? "group layout cannot be multicast (non-VMM layer, remote node or shared GPU)" : why;
        mc = McObject();
    }
    auto ins = ctx->mcasts.emplace(key, mc);
    return ins.first->second.valid() ? &ins.first->second : nullptr;
}

// port rates behind the multicast cost model, GB/s (tools/mcast_probe.py on 2 and 4 B200s, profiles/r02_n*_mcast_probe.jsonl)
constexpr double kUnicastPushGBps = 713.0;
constexpr double kMulticastGBps = 560.0;

struct CompileItem {
    dissem_xfer x;
    McObject *mc = nullptr;      // set: one multicast stream standing for n_dsts transfers of the same bytes
    int n_dsts = 1;
    std::vector<int> mc_dsts;    // the destination nodes of a multicast stream
};

static int compile_sched(dissem_ctx *ctx, const dissem_xfer *xfers, int n, Sched *out) {
    if (n < 0 || (n > 0 && !xfers)) return set_error(DISSEM_EINVAL, "bad transfer list");
    for (int i = 0; i < n; ++i) {
        const dissem_xfer &x = xfers[i];
        int rc = check_node(ctx, x.src_node);
        if (rc) return rc;
        rc = check_node(ctx, x.dst_node);
        if (rc) return rc;
        if (x.size < 0 || x.off < 0 || x.off + x.size > x.total)
            return set_error(DISSEM_ERANGE, "transfer %d: [%lld,+%lld) outside layer of %lld bytes", i,
                             (long long)x.off, (long long)x.size, (long long)x.total);
    }
    // Rounds (DISSEM_XFER_ROUND): transfers of round r start after every transfer of round r-1 has completed on every
    // GPU of this process; a relay plan forwards in round r what arrived in round r-1.
    auto round_of = [](const dissem_xfer &x) { return (int)((x.flags >> DISSEM_XFER_ROUND_SHIFT) & 0xFFFF); };
    for (int i = 0; i < n; ++i)
        if (xfers[i].flags & DISSEM_XFER_FILE)
            return set_error(DISSEM_ENOTSUP, "layer %d from node %d's disk copy: a transfer from disk is streamed while its file is read and "
                             "cannot be compiled; use dissem_schedule_submit or dissem_send_file", xfers[i].layer, xfers[i].src_node);
    std::vector<int> rounds;
    for (int i = 0; i < n; ++i) rounds.push_back(round_of(xfers[i]));
    std::sort(rounds.begin(), rounds.end());
    rounds.erase(std::unique(rounds.begin(), rounds.end()), rounds.end());
    std::vector<std::vector<std::vector<CompileItem>>> assigned;          // [round index][device slot]
    std::vector<char> used((size_t)std::max(n, 1), 0);
    auto is_pull = [&](const dissem_xfer &x) { return ctx->opt_pull == 1 || (ctx->opt_pull == 0 && (x.flags & DISSEM_XFER_PULL)); };
    for (int cur_round : rounds) {
    std::vector<std::vector<CompileItem>> by_dev(ctx->devs.size());
    // Pass 1 (K2): the same bytes of one local source for several GPUs -> one NVLS multicast store stream.  This takes
    // precedence over a transfer's PULL flag: one pushed stream that lands k times beats k pulled streams.
    if (ctx->opt_mcast && ctx->opt_pull != 1) {
        struct Group { std::vector<int> members; };
        std::vector<Group> groups;
        std::vector<char> grouped((size_t)std::max(n, 1), 0);
        for (int i = 0; i < n; ++i) {
            const dissem_xfer &x = xfers[i];
            const int sslot = ctx->node_dev_slot[x.src_node];
            if (grouped[i] || round_of(x) != cur_round || x.size == 0 || sslot < 0 || x.off % 16 || x.size % 16) continue;
            Group g;
            g.members.push_back(i);
            for (int j = i + 1; j < n; ++j) {
                const dissem_xfer &y = xfers[j];
                if (!grouped[j] && round_of(y) == cur_round && y.src_node == x.src_node && y.layer == x.layer && y.off == x.off &&
                    y.size == x.size && y.total == x.total && y.dst_node != x.dst_node)
                    g.members.push_back(j);
            }
            const bool forced_single = ctx->opt_mcast == 2 && ctx->node_dev_slot[x.dst_node] != sslot;   // another GPU, possibly another process
            if (g.members.size() < 2 && !forced_single) continue;
            for (int m : g.members) grouped[m] = 1;
            groups.push_back(std::move(g));
        }
        // Cost model (opt_mcast == 1).  Measured on B200 / NVLink 5 with tools/mcast_probe.py (profiles/r02_n2_mcast_probe.jsonl,
        // r02_n4_mcast_probe.jsonl): a unicast push stream moves ~713 GB/s out of a port; an NVLS multicast stream leaves
        // its source at ~560 GB/s whatever the kernel shape and the number of destinations (1, 2 or 3), and each member
        // receives at that same rate.  Multicast therefore pays only when the plan is bound by a source's egress:
        // compare max-port time with every group sent as unicast copies against every group sent once.
        bool use_mc = !groups.empty();
        if (ctx->opt_mcast == 1 && use_mc) {
            auto gpu_of = [&](int node) { int sl = ctx->node_dev_slot[node]; return sl >= 0 ? ctx->devs[sl].dev : 1000 + node; };
            std::map<int, double> eg_uni, eg_mc, ing;
            for (int i = 0; i < n; ++i) {
                const dissem_xfer &x = xfers[i];
                const int gs = gpu_of(x.src_node), gd = gpu_of(x.dst_node);
                if (gs == gd || round_of(x) != cur_round) continue;
                eg_uni[gs] += (double)x.size;
                ing[gd] += (double)x.size;
                if (!grouped[i]) eg_mc[gs] += (double)x.size;
            }
            for (auto &g : groups) eg_mc[gpu_of(xfers[g.members[0]].src_node)] += (double)xfers[g.members[0]].size;
            auto mx = [](const std::map<int, double> &m) { double v = 0; for (auto &kv : m) v = std::max(v, kv.second); return v; };
            const double t_uni = std::max(mx(eg_uni), mx(ing)) / kUnicastPushGBps;
            const double t_mc = std::max(mx(eg_mc), mx(ing)) / kMulticastGBps;
            use_mc = t_mc < t_uni;
            if (!use_mc) ctx->mcast_note = "cost model: the plan is not egress-bound, unicast copies are faster than multicast";
        }
        if (use_mc)
            for (auto &g : groups) {
                const dissem_xfer &x = xfers[g.members[0]];
                const int sslot = ctx->node_dev_slot[x.src_node];
                LayerBuf *src = nullptr;
                if (find_layer(ctx, x.src_node, x.layer, &src) || src->size != x.total) continue;   // pass 2 reports it
                std::vector<int> dsts;
                bool ok = true;
                for (int m : g.members) {
                    LayerBuf *dl = nullptr;
                    if (find_layer(ctx, xfers[m].dst_node, x.layer, &dl) || dl->size != x.total) ok = false;
                    dsts.push_back(xfers[m].dst_node);
                }
                McObject *mc = ok ? get_or_make_mcast(ctx, x.layer, x.src_node, dsts) : nullptr;
                if (!mc) continue;
                CompileItem it;
                it.x = x;
                it.mc = mc;
                it.n_dsts = (int)g.members.size();
                it.mc_dsts = dsts;
                by_dev[sslot].push_back(it);
                for (int m : g.members) used[m] = 1;
                ctx->mcast_groups_compiled++;
            }
    }
    // Pass 2: everything else, executed by the source's GPU (push) or the destination's (pull)
    for (int i = 0; i < n; ++i) {
        if (used[i] || round_of(xfers[i]) != cur_round) continue;
        const dissem_xfer &x = xfers[i];
        int slot = ctx->node_dev_slot[is_pull(x) ? x.dst_node : x.src_node];
        if (slot < 0) continue;          // the process that owns the executing side runs it
        if (x.size == 0) continue;       // an empty layerMsg moves nothing (LayerSize 0)
        CompileItem it;
        it.x = x;
        by_dev[slot].push_back(it);
    }
    assigned.push_back(std::move(by_dev));
    }   // per round
    out->unit_shift = kChunkShift;
    for (auto &by_dev : assigned) {
        std::vector<std::vector<dissem_xfer>> plain(by_dev.size());
        for (size_t s = 0; s < by_dev.size(); ++s)
            for (auto &it : by_dev[s]) plain[s].push_back(it.x);
        out->unit_shift = std::min(out->unit_shift, pick_unit_shift(ctx, plain));
    }
    std::vector<std::vector<XferDesc>> part_descs;                        // parallel to out->parts
    for (size_t ri = 0; ri < assigned.size(); ++ri) {
    auto &by_dev = assigned[ri];
    for (size_t slot = 0; slot < by_dev.size(); ++slot) {
        if (by_dev[slot].empty()) continue;
        std::vector<XferDesc> descs;
        long long units = 0;
        int64_t bytes = 0;
        bool any_host_src = false, any_mc = false;
        std::vector<std::pair<int, int>> touch, writes;
        auto written = [&](int node, int layer) {
            if (ctx->node_dev_slot[node] < 0) return;                       // an imported buffer: its own process reads it
            auto key = std::make_pair(node, layer);
            if (std::find(writes.begin(), writes.end(), key) == writes.end()) writes.push_back(key);
        };
        auto touches = [&](int node, int layer, const LayerBuf *lb) {
            // a local buffer of another GPU: its owner's stream may still be zeroing / filling / uploading it
            if (lb->imported || lb->host || ctx->node_dev_slot[node] < 0 || ctx->node_dev_slot[node] == (int)slot) return;
            auto key = std::make_pair(node, layer);
            if (std::find(touch.begin(), touch.end(), key) == touch.end()) touch.push_back(key);
        };
        for (auto &it : by_dev[slot]) {
            const dissem_xfer &x = it.x;
            LayerBuf *src = nullptr, *dst = nullptr;
            int rc = find_layer(ctx, x.src_node, x.layer, &src);
            if (rc) return rc;
            touches(x.src_node, x.layer, src);
            XferDesc d;
            memset(&d, 0, sizeof d);
            d.src = src->base;
            d.off = x.off;
            d.size = x.size;
            d.total = x.total;
            d.first_unit = units;
            if ((x.flags & DISSEM_XFER_WAIT_SRC) && !src->host && ctx->opt_wait_src) {
                d.wait_ptr = src->received();
                d.wait_value = (unsigned long long)(x.off + x.size);
                out->flag_ordered = true;
            }
            if (it.mc) {
                uint8_t *mbase = reinterpret_cast<uint8_t *>(it.mc->va);
                d.dst = mbase;
                d.csum = reinterpret_cast<unsigned long long *>(mbase + src->trailer_off() + 256);
                d.received = reinterpret_cast<unsigned long long *>(mbase + src->trailer_off());
                d.flags = kFlagMulticast;
                any_mc = true;
                for (int dn : it.mc_dsts) {
                    LayerBuf *ml = nullptr;
                    if (find_layer(ctx, dn, x.layer, &ml) == 0) touches(dn, x.layer, ml);
                    written(dn, x.layer);
                }
                written(x.src_node, x.layer);          // the source's own copy is a member of the object
            } else {
                rc = find_layer(ctx, x.dst_node, x.layer, &dst);
                if (rc) return rc;
                if (src->size != x.total || dst->size != x.total)
                    return set_error(DISSEM_ERANGE, "layer %d: total %lld but src holds %lld and dst %lld bytes", x.layer,
                                     (long long)x.total, (long long)src->size, (long long)dst->size);
                if (src->base == dst->base) return set_error(DISSEM_EINVAL, "layer %d: source and destination are the same buffer", x.layer);
                if (dst->host) return set_error(DISSEM_ENOTSUP, "layer %d: node %d holds it in attached host memory, which can only be a source", x.layer, x.dst_node);
                d.dst = dst->base;
                d.csum = dst->csum();
                d.received = dst->received();
                d.flags = 0;
                touches(x.dst_node, x.layer, dst);
                written(x.dst_node, x.layer);
            }
            if (out->unit_shift == kChunkShift) d.flags |= kFlagStoreSums;       // units are whole chunks: sums stored, not added
            units += units_of(x.off, x.size, out->unit_shift);
            bytes += x.size * (int64_t)it.n_dsts;
            descs.push_back(d);
            if (src->host) any_host_src = true;
        }
        // stripe order (dissem_kernels.cuh): sort by unit count, then store each level's starting ticket
        std::stable_sort(descs.begin(), descs.end(), [&](const XferDesc &a, const XferDesc &b) {
            return units_of(a.off, a.size, out->unit_shift) < units_of(b.off, b.size, out->unit_shift);
        });
        {
            const long long m = (long long)descs.size();
            long long sum_before = 0, prev = 0;          // sum of c_i for i < j-1 ... maintained incrementally
            for (long long j = 0; j < m; ++j) {
                // L_j = sum_{i<j} min(c_i, c_{j-1}) + (m-j)*c_{j-1}  ==  sum_{i<j} c_i' where all c_i <= c_{j-1}
                descs[j].first_unit = sum_before + (m - j) * prev;
                descs[j].prev_units = (int)prev;
                const long long c = units_of(descs[j].off, descs[j].size, out->unit_shift);
                sum_before += c;
                prev = c;
            }
        }
        SchedPart part;
        part.dev_slot = (int)slot;
        part.round = rounds[ri];
        part.host_src = any_host_src;
        part.has_mc = any_mc;
        part.n_descs = (int)descs.size();
        part.total_units = units;
        part.bytes = bytes;
        part.touch = std::move(touch);
        part.writes = std::move(writes);
        out->parts.push_back(part);
        part_descs.push_back(std::move(descs));
        out->units += units;
        out->bytes += bytes;
    }
    }   // per round

    // Partially covered checksum chunks (fragment edges) may be shared by several descriptors of this schedule.  When the
    // schedule's fragments together cover such a chunk completely, every contribution to its entry comes from this
    // schedule, and the descriptor in the earliest round with the lowest offset owns the zeroing of the entry
    // (kFlagZeroHead / kFlagZeroTail), which makes relaunching idempotent.  A chunk this schedule covers only in part
    // also receives bytes from other submissions (one Transport.Send per fragment): it is left alone, its entry was
    // zeroed when the destination buffer was allocated or reset.
    {
        struct Owner { int round; long long off; size_t part, desc; int flag; long long covered, length; int first_slot; bool multi_slot; };
        std::map<std::pair<const void *, long long>, Owner> owners;
        for (size_t pi = 0; pi < part_descs.size(); ++pi)
            for (size_t di = 0; di < part_descs[pi].size(); ++di) {
                const XferDesc &d = part_descs[pi][di];
                if (!d.csum) continue;
                const long long end = d.off + d.size;
                const long long ch = d.off >> kChunkShift, ct = (end - 1) >> kChunkShift;
                auto chunk_end = [&](long long c) { return std::min<long long>((c + 1) << kChunkShift, d.total); };
                auto consider = [&](long long c, int flag) {
                    const long long c0 = c << kChunkShift, c1 = chunk_end(c);
                    const long long cov = std::min(end, c1) - std::max<long long>(d.off, c0);
                    auto key = std::make_pair((const void *)d.csum, c);
                    auto it = owners.find(key);
                    if (it == owners.end()) {
                        owners[key] = Owner{out->parts[pi].round, d.off, pi, di, flag, cov, c1 - c0, out->parts[pi].dev_slot, false};
                    } else {
                        it->second.covered += cov;
                        if (out->parts[pi].dev_slot != it->second.first_slot) it->second.multi_slot = true;
                        if (std::make_pair(out->parts[pi].round, (long long)d.off) < std::make_pair(it->second.round, it->second.off)) {
                            it->second.round = out->parts[pi].round;
                            it->second.off = d.off;
                            it->second.part = pi;
                            it->second.desc = di;
                            it->second.flag = flag;
                        }
                    }
                };
                if ((d.off & (kChunkBytes - 1)) != 0 || end < chunk_end(ch)) consider(ch, kFlagZeroHead);
                if (ct != ch && end < chunk_end(ct)) consider(ct, kFlagZeroTail);
            }
        for (auto &kv : owners)
            if (kv.second.covered == kv.second.length) {
                part_descs[kv.second.part][kv.second.desc].flags |= kv.second.flag;
                if (kv.second.multi_slot) out->cross_dev_shared = true;
            }
    }
    for (size_t pi = 0; pi < out->parts.size(); ++pi) {
        bool zero = false;
        for (auto &d : part_descs[pi])
            if (d.csum && (!(d.flags & kFlagStoreSums) || (d.flags & (kFlagZeroHead | kFlagZeroTail)))) zero = true;
        out->parts[pi].needs_zero = zero;
    }
    for (size_t pi = 0; pi < out->parts.size(); ++pi) {
        SchedPart &part = out->parts[pi];
        DeviceState &pds = ctx->devs[part.dev_slot];
        CUDA_TRY(cudaSetDevice(pds.dev));
        if ((int)part_descs[pi].size() <= kSlotDescs && !pds.free_slots.empty()) {
            // a slot is reused only by work enqueued later on the same stream, so the previous user has finished with it
            part.pool_slot = pds.free_slots.back();
            pds.free_slots.pop_back();
            part.d_descs = pds.desc_pool + (size_t)part.pool_slot * kSlotDescs;
        } else {
            CUDA_TRY(cudaMalloc(&part.d_descs, part_descs[pi].size() * sizeof(XferDesc)));
        }
        // stream-ordered upload from pageable memory: the copy is staged before the call returns
        CUDA_TRY(cudaMemcpyAsync(part.d_descs, part_descs[pi].data(), part_descs[pi].size() * sizeof(XferDesc),
                                 cudaMemcpyHostToDevice, ctx->devs[part.dev_slot].stream));
    }
    return 0;
}

static cudaError_t take_event(dissem_ctx *ctx, int dev, cudaEvent_t *out) {
    auto &pool = ctx->event_pool[dev & 63];
    if (!pool.empty()) {
        *out = pool.back();
        pool.pop_back();
        return cudaSuccess;
    }
    return cudaEventCreate(out);
}

static int launch_sched(dissem_ctx *ctx, const Sched &s, Ticket *t) {
    // Parts are ordered by round.  Per round, on the device (cudaStreamWaitEvent, no host round trip):
    //   phase A  every GPU waits for the previous round's end events, then zeroes the checksum entries it owns;
    //   phase B  every GPU waits for the other GPUs' phase-A events of this round, then runs its copy kernel.
    std::vector<cudaEvent_t> prev_round;
    size_t i = 0;
    while (i < s.parts.size()) {
        size_t j = i;
        while (j < s.parts.size() && s.parts[j].round == s.parts[i].round) ++j;
        const size_t first_tp = t->parts.size();
        for (size_t k = i; k < j; ++k) {                                  // phase A
            const SchedPart &p = s.parts[k];
            DeviceState &ds = ctx->devs[p.dev_slot];
            CUDA_TRY(cudaSetDevice(ds.dev));
            if (!s.flag_ordered)
                for (cudaEvent_t ev : prev_round) CUDA_TRY(cudaStreamWaitEvent(ds.stream, ev, 0));
            // buffers of other GPUs this part touches: wait until their owners' streams have finished writing them
            // (trailer zeroing at alloc / reset, fill, upload) -- streams of different GPUs are not ordered otherwise
            for (auto &nl : p.touch) {
                auto li = ctx->layers[nl.first].find(nl.second);
                if (li != ctx->layers[nl.first].end() && li->second.ready)
                    CUDA_TRY(cudaStreamWaitEvent(ds.stream, static_cast<cudaEvent_t>(li->second.ready), 0));
            }
            // a relaunch that re-zeroes a checksum entry shared with kernels of other GPUs must not overtake those GPUs'
            // previous launch (whose partial sums would land after the zeroing)
            if (s.cross_dev_shared && i == 0)
                for (auto &other : ctx->devs)
                    if (other.dev != ds.dev && other.last_end) CUDA_TRY(cudaStreamWaitEvent(ds.stream, other.last_end, 0));
            TicketPart tp;
            tp.dev_slot = p.dev_slot;
            tp.start = tp.mid = tp.end = nullptr;
            t->parts.push_back(tp);
            TicketPart &ntp = t->parts.back();
            CUDA_TRY(take_event(ctx, ds.dev, &ntp.start));
            CUDA_TRY(take_event(ctx, ds.dev, &ntp.mid));
            CUDA_TRY(take_event(ctx, ds.dev, &ntp.end));
            tp = ntp;
            CUDA_TRY(cudaEventRecord(tp.start, ds.stream));
            if (p.needs_zero) CUDA_TRY(launch_zero_csum(p.d_descs, p.n_descs, ds.stream));
            CUDA_TRY(cudaEventRecord(tp.mid, ds.stream));
        }
        std::vector<cudaEvent_t> this_round;
        for (size_t k = i; k < j; ++k) {                                  // phase B
            const SchedPart &p = s.parts[k];
            DeviceState &ds = ctx->devs[p.dev_slot];
            TicketPart &tp = t->parts[first_tp + (k - i)];
            CUDA_TRY(cudaSetDevice(ds.dev));
            for (size_t m = i; m < j; ++m)      // (a part that zeroes nothing has nothing the others must wait for)
                if (m != k && s.parts[m].needs_zero) CUDA_TRY(cudaStreamWaitEvent(ds.stream, t->parts[first_tp + (m - i)].mid, 0));
            for (auto &nl : p.writes) {
                auto li = ctx->layers[nl.first].find(nl.second);
                if (li != ctx->layers[nl.first].end()) li->second.writers |= 1ull << (p.dev_slot & 63);
            }
            const bool reg_path = p.host_src || (p.has_mc && ctx->opt_mcast_variant == 0);
            LaunchCfg cfg{ds.n_sm, (int)ctx->opt_ctas_per_sm, reg_path ? 0 : (int)ctx->opt_variant, (int)ctx->opt_reg_cfg, (int)ctx->opt_tma_cfg};
            CUDA_TRY(launch_copy_csum(p.d_descs, p.n_descs, p.total_units, s.unit_shift, ds.counters, cfg, ds.stream, nullptr));
            CUDA_TRY(cudaEventRecord(tp.end, ds.stream));
            if (ds.last_end) CUDA_TRY(cudaEventRecord(ds.last_end, ds.stream));
            this_round.push_back(tp.end);
        }
        prev_round.swap(this_round);
        i = j;
    }
    return 0;
}

// caller holds ctx->mu
static void drop_ticket(dissem_ctx *ctx, Ticket &t) {
    for (auto &p : t.parts) {
        auto &pool = ctx->event_pool[ctx->devs[p.dev_slot].dev & 63];
        for (cudaEvent_t ev : {p.start, p.mid, p.end})
            if (ev) pool.push_back(ev);
    }
    t.parts.clear();
}

// ------------------------------------------------------------------------------------------------
// disk tier: file -> pinned staging ring -> copy kernel (PCIe in, HBM or NVLink out) with the checksum fused
// ------------------------------------------------------------------------------------------------
static int ensure_ring(DeviceState &ds) {
    StageRing &r = ds.ring;
    if (r.d_descs) return 0;
    CUDA_TRY(cudaSetDevice(ds.dev));
    for (int i = 0; i < kStageSlots; ++i) {
        // portable + mapped: the kernel of any local GPU dereferences the host pointer itself (unified addressing)
        if (!r.slot[i]) CUDA_TRY(cudaHostAlloc((void **)&r.slot[i], (size_t)kStageBytes + 256, cudaHostAllocPortable | cudaHostAllocMapped));
        if (!r.done[i]) CUDA_TRY(cudaEventCreateWithFlags(&r.done[i], cudaEventDisableTiming));
    }
    CUDA_TRY(cudaMalloc(&r.d_descs, kStageSlots * sizeof(XferDesc)));
    return 0;
}

static int pread_range(const FileSrc &f, uint8_t *dst, int64_t len, int64_t layer_off, std::string *err) {
    int64_t got = 0;
    char buf[512];
    while (got < len) {
        ssize_t k = pread(f.fd, dst + got, (size_t)(len - got), (off_t)(f.file_off + layer_off + got));
        if (k < 0 && errno == EINTR) continue;
        if (k < 0) {
            snprintf(buf, sizeof buf, "reading %s at %lld: %s", f.path.c_str(), (long long)(f.file_off + layer_off + got), strerror(errno));
            *err = buf;
            return DISSEM_EIO;
        }
        if (k == 0) {
            snprintf(buf, sizeof buf, "%s ends at %lld, before the %lld bytes of the layer", f.path.c_str(),
                     (long long)(f.file_off + layer_off + got), (long long)f.size);
            *err = buf;
            return DISSEM_EIO;
        }
        got += k;
    }
    return 0;
}

// One piece of the file into its staging slot.  A single pread() stream copies out of the page cache at a few GB/s,
// far below what PCIe then carries; the piece is therefore read by `threads` threads, each its own 4 KiB-aligned share.
static int pread_full(const FileSrc &f, uint8_t *dst, int64_t len, int64_t layer_off, int threads) {
    constexpr int64_t kMinShare = 1 << 20;
    int n = (int)std::min<int64_t>(std::max(threads, 1), len / kMinShare);
    std::string err;
    if (n <= 1) {
        int rc = pread_range(f, dst, len, layer_off, &err);
        return rc ? set_error(rc, "%s", err.c_str()) : 0;
    }
    const int64_t share = align_up((len + n - 1) / n, 4096);
    std::vector<std::thread> pool;
    std::vector<int> rcs((size_t)n, 0);
    std::vector<std::string> errs((size_t)n);
    for (int i = 0; i < n; ++i) {
        const int64_t a = std::min(len, i * share), b = std::min(len, (i + 1) * share);
        pool.emplace_back([&, i, a, b] { rcs[i] = b > a ? pread_range(f, dst + a, b - a, layer_off + a, &errs[i]) : 0; });
    }
    for (auto &t : pool) t.join();
    for (int i = 0; i < n; ++i)
        if (rcs[i]) return set_error(rcs[i], "%s", errs[i].c_str());
    return 0;
}

// What stream_from_file needs from the store, copied out under ctx->mu so that the file can be read without it.
struct FileJob {
    dissem_xfer x;
    FileSrc file;
    int dev_slot = -1;
    uint8_t *dst = nullptr;
    unsigned long long *csum = nullptr, *received = nullptr;
    TicketPart tp{-1, nullptr, nullptr, nullptr};
    cudaEvent_t dst_ready = nullptr;  // set when the destination buffer lives on another GPU than the one that streams the file
    bool owns_fd = false;             // file.fd is a dup() made for this transfer; dissem_schedule_submit closes it
};

// caller holds ctx->mu
static int prepare_file_job(dissem_ctx *ctx, const dissem_xfer &x, FileJob *job) {
    int rc = check_node(ctx, x.src_node);
    if (rc) return rc;
    if ((x.flags >> DISSEM_XFER_ROUND_SHIFT) & 0xFFFF)
        return set_error(DISSEM_ENOTSUP, "layer %d from node %d's disk copy: transfers from disk belong to round 0", x.layer, x.src_node);
    auto fit = ctx->files[x.src_node].find(x.layer);
    if (fit == ctx->files[x.src_node].end()) return set_error(DISSEM_ENOENT, "node %d holds no disk copy of layer %d", x.src_node, x.layer);
    job->x = x;
    job->file = fit->second;
    // the transfer reads through its own descriptor: a concurrent dissem_layer_detach_file cannot pull the file away
    job->file.fd = dup(fit->second.fd);
    if (job->file.fd < 0) return set_error(DISSEM_EIO, "dup(%s): %s", fit->second.path.c_str(), strerror(errno));
    job->owns_fd = true;
    job->dev_slot = ctx->node_dev_slot[x.src_node];              // attach_file accepts local nodes only
    LayerBuf *dst = nullptr;
    rc = find_layer(ctx, x.dst_node, x.layer, &dst);
    if (rc) return rc;
    if (job->file.size != x.total || dst->size != x.total)
        return set_error(DISSEM_ERANGE, "layer %d: total %lld but the file holds %lld and dst %lld bytes", x.layer, (long long)x.total,
                         (long long)job->file.size, (long long)dst->size);
    if (x.off < 0 || x.size < 0 || x.off + x.size > x.total)
        return set_error(DISSEM_ERANGE, "layer %d: fragment [%lld,+%lld) outside %lld bytes", x.layer, (long long)x.off, (long long)x.size, (long long)x.total);
    if (dst->host) return set_error(DISSEM_ENOTSUP, "layer %d: node %d holds it in attached host memory, which can only be a source", x.layer, x.dst_node);
    job->dst = dst->base;
    job->csum = dst->csum();
    job->received = dst->received();
    if (!dst->imported && ctx->node_dev_slot[x.dst_node] != job->dev_slot) job->dst_ready = static_cast<cudaEvent_t>(dst->ready);
    dst->writers |= 1ull << (job->dev_slot & 63);
    DeviceState &ds = ctx->devs[job->dev_slot];
    CUDA_TRY(cudaSetDevice(ds.dev));
    job->tp.dev_slot = job->dev_slot;
    job->tp.start = job->tp.mid = job->tp.end = nullptr;
    CUDA_TRY(take_event(ctx, ds.dev, &job->tp.start));
    CUDA_TRY(take_event(ctx, ds.dev, &job->tp.mid));
    CUDA_TRY(take_event(ctx, ds.dev, &job->tp.end));
    return 0;
}

// Runs WITHOUT ctx->mu (other threads keep sending); the ring's own mutex serialises file transfers per GPU.
// Pieces are cut at multiples of kStageBytes in LAYER space, so a piece covers whole checksum chunks except at the
// fragment's own two ends; those entries are accumulated into, as for any single-fragment submission.
static int stream_from_file(dissem_ctx *ctx, FileJob &job, const LaunchCfg &cfg_in, int64_t opt_unit_bytes, int file_threads) {
    DeviceState &ds = ctx->devs[job.dev_slot];
    std::lock_guard<std::mutex> ring_lock(*ds.ring_mu);
    int rc = ensure_ring(ds);
    if (rc) return rc;
    StageRing &r = ds.ring;
    CUDA_TRY(cudaSetDevice(ds.dev));
    if (job.dst_ready) CUDA_TRY(cudaStreamWaitEvent(ds.stream, job.dst_ready, 0));   // the owner may still be zeroing its trailer
    CUDA_TRY(cudaEventRecord(job.tp.start, ds.stream));
    CUDA_TRY(cudaEventRecord(job.tp.mid, ds.stream));
    const dissem_xfer &x = job.x;
    LaunchCfg cfg = cfg_in;
    cfg.variant = 0;                                   // the source is host memory: register kernel (as for attached host layers)
    const int64_t end = x.off + x.size;
    for (int64_t pos = x.off; pos < end;) {
        const int64_t piece_end = std::min(end, (pos / kStageBytes + 1) * kStageBytes);
        const int64_t len = piece_end - pos;
        const int s = r.next;
        r.next = (r.next + 1) % kStageSlots;
        if (r.in_flight[s]) {
            CUDA_TRY(cudaEventSynchronize(r.done[s]));
            r.in_flight[s] = false;
        }
        uint8_t *stage = r.slot[s] + (pos & 255);      // keep the piece's alignment (the kernel vectorises on layer offsets)
        rc = pread_full(job.file, stage, len, pos, file_threads);
        if (rc) return rc;
        XferDesc d;
        memset(&d, 0, sizeof d);
        d.src = reinterpret_cast<const uint8_t *>(reinterpret_cast<uintptr_t>(stage) - (uintptr_t)pos);   // the kernel reads src + layer offset
        d.dst = job.dst;
        d.csum = job.csum;
        d.received = job.received;
        d.off = pos;
        d.size = len;
        d.total = x.total;
        // units: 256 KiB unless that starves the grid (one piece is at most 64 of them)
        int shift = kChunkShift;
        if (opt_unit_bytes > 0) {
            shift = 12;
            while ((1ll << shift) < opt_unit_bytes) ++shift;
            shift = std::min(shift, (int)kChunkShift);
        } else {
            while (shift > 14 && units_of(pos, len, shift) < 4ll * ds.n_sm) --shift;
        }
        const long long units = units_of(pos, len, shift);
        CUDA_TRY(cudaMemcpyAsync(r.d_descs + s, &d, sizeof d, cudaMemcpyHostToDevice, ds.stream));   // pageable: staged before return
        CUDA_TRY(launch_zero_csum(r.d_descs + s, 1, ds.stream));
        CUDA_TRY(launch_copy_csum(r.d_descs + s, 1, units, shift, ds.counters, cfg, ds.stream, nullptr));
        CUDA_TRY(cudaEventRecord(r.done[s], ds.stream));
        r.in_flight[s] = true;
        pos = piece_end;
    }
    CUDA_TRY(cudaEventRecord(job.tp.end, ds.stream));
    if (ds.last_end) CUDA_TRY(cudaEventRecord(ds.last_end, ds.stream));
    return 0;
}

}  // namespace dissem

// ====================================================================================================
// C ABI
// ====================================================================================================
extern "C" {

int dissem_version(void) { return DISSEM_VERSION; }
const char *dissem_last_error(void) { return dissem::last_error(); }

int dissem_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int dissem_init(int n_nodes, const int *device_of_node, dissem_ctx **out) {
    if (n_nodes <= 0 || !device_of_node || !out) return set_error(DISSEM_EINVAL, "dissem_init: bad arguments");
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        return set_error(DISSEM_ENODEV, "no usable CUDA device (%s); this library has no CPU path",
                         e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    std::unique_ptr<dissem_ctx> ctx(new dissem_ctx);
    ctx->n_nodes = n_nodes;
    ctx->node_dev_slot.assign(n_nodes, -1);
    ctx->layers.resize(n_nodes);
    ctx->files.resize(n_nodes);
    for (int i = 0; i < n_nodes; ++i) {
        int dev = device_of_node[i];
        if (dev < 0) continue;
        if (dev >= n_dev) return set_error(DISSEM_EINVAL, "node %d mapped to device %d but only %d visible", i, dev, n_dev);
        int slot = -1;
        for (size_t s = 0; s < ctx->devs.size(); ++s)
            if (ctx->devs[s].dev == dev) slot = (int)s;
        if (slot < 0) {
            DeviceState ds;
            ds.dev = dev;
            CUDA_TRY(cudaSetDevice(dev));
            CUDA_TRY(cudaDeviceGetAttribute(&ds.n_sm, cudaDevAttrMultiProcessorCount, dev));
            CUDA_TRY(cudaStreamCreateWithFlags(&ds.stream, cudaStreamNonBlocking));
            CUDA_TRY(cudaMalloc(&ds.counters, sizeof(SchedCounters)));
            CUDA_TRY(cudaMemset(ds.counters, 0, sizeof(SchedCounters)));
            CUDA_TRY(cudaMalloc(&ds.desc_pool, (size_t)kPoolSlots * kSlotDescs * sizeof(XferDesc)));
            for (int sl = kPoolSlots - 1; sl >= 0; --sl) ds.free_slots.push_back(sl);
            CUDA_TRY(cudaEventCreateWithFlags(&ds.last_end, cudaEventDisableTiming));
            CUDA_TRY(cudaStreamCreateWithFlags(&ds.rd_stream, cudaStreamNonBlocking));
            CUDA_TRY(cudaMalloc(&ds.arena_dev, kArenaBytes));
            CUDA_TRY(cudaHostAlloc((void **)&ds.arena_host, kArenaBytes, cudaHostAllocPortable));
            ds.arena_bytes = kArenaBytes;
            CUDA_TRY(warm_kernels());
            ctx->devs.push_back(ds);
            slot = (int)ctx->devs.size() - 1;
        }
        ctx->node_dev_slot[i] = slot;
    }
    // peer access between all local devices (single-process multi-GPU)
    for (auto &a : ctx->devs)
        for (auto &b : ctx->devs) {
            if (a.dev == b.dev) continue;
            int can = 0;
            CUDA_TRY(cudaDeviceCanAccessPeer(&can, a.dev, b.dev));
            if (!can) return set_error(DISSEM_ENOTSUP, "device %d cannot access device %d", a.dev, b.dev);
            CUDA_TRY(cudaSetDevice(a.dev));
            cudaError_t pe = cudaDeviceEnablePeerAccess(b.dev, 0);
            if (pe == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (pe != cudaSuccess) return set_error(DISSEM_ECUDA, "cudaDeviceEnablePeerAccess(%d->%d): %s", a.dev, b.dev, cudaGetErrorString(pe));
        }
    *out = ctx.release();
    return DISSEM_OK;
}

int dissem_close(dissem_ctx *ctx) {
    if (!ctx) return DISSEM_OK;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        for (auto &ds : ctx->devs) {
            cudaSetDevice(ds.dev);
            cudaStreamSynchronize(ds.stream);
        }
        for (auto &kv : ctx->tickets) drop_ticket(ctx, kv.second);
        for (auto &pool : ctx->event_pool) {
            for (cudaEvent_t ev : pool) cudaEventDestroy(ev);
            pool.clear();
        }
        for (auto &kv : ctx->scheds) free_sched(ctx, kv.second);
        for (auto &kv : ctx->mcasts) mcast_destroy(&kv.second);
        ctx->mcasts.clear();
        for (int n = 0; n < ctx->n_nodes; ++n)
            for (auto &kv : ctx->layers[n]) {
                LayerBuf &lb = kv.second;
                drop_ready(lb);
                if (lb.host) continue;
                if (lb.vmm) vmm_free(&lb.blk);
                else if (lb.imported) cudaIpcCloseMemHandle(lb.base);
                else {
                    cudaSetDevice(lb.device);
                    cudaFree(lb.base);
                }
            }
        for (auto &per_node : ctx->files)
            for (auto &kv : per_node)
                if (kv.second.fd >= 0) close(kv.second.fd);
        for (auto &kv : ctx->reserved) {
            cudaSetDevice(kv.first.first);
            for (uint8_t *p : kv.second) cudaFree(p);
        }
        for (auto &ds : ctx->devs) {
            cudaSetDevice(ds.dev);
            for (int i = 0; i < kStageSlots; ++i) {
                if (ds.ring.slot[i]) cudaFreeHost(ds.ring.slot[i]);
                if (ds.ring.done[i]) cudaEventDestroy(ds.ring.done[i]);
            }
            if (ds.ring.d_descs) cudaFree(ds.ring.d_descs);
            cudaFree(ds.counters);
            cudaFree(ds.desc_pool);
            if (ds.last_end) cudaEventDestroy(ds.last_end);
            if (ds.arena_dev) cudaFree(ds.arena_dev);
            if (ds.arena_host) cudaFreeHost(ds.arena_host);
            if (ds.rd_stream) cudaStreamDestroy(ds.rd_stream);
            cudaStreamDestroy(ds.stream);
        }
    }
    delete ctx;
    return DISSEM_OK;
}

int dissem_n_nodes(dissem_ctx *ctx) { return ctx ? ctx->n_nodes : set_error(DISSEM_EINVAL, "null context"); }

int dissem_layer_alloc(dissem_ctx *ctx, int node, int layer, int64_t size) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    DeviceState *ds = nullptr;
    int rc = local_dev(ctx, node, &ds);
    if (rc) return rc;
    if (size < 0) size = 0;                                   // cmd/config.go:100-102
    if (ctx->layers[node].count(layer)) return set_error(DISSEM_EEXIST, "node %d already holds layer %d", node, layer);
    LayerBuf lb;
    lb.size = size;
    lb.device = ds->dev;
    CUDA_TRY(cudaSetDevice(ds->dev));
    if (ctx->opt_alloc_vmm) {
        // VMM-backed: can be bound to an NVLS multicast object; mapped for every local GPU (peer stores)
        std::vector<int> access;
        for (auto &d : ctx->devs) access.push_back(d.dev);
        std::string err = vmm_alloc(ds->dev, access, (size_t)lb.alloc_bytes(), &lb.blk, ctx->opt_alloc_vmm == 2);
        if (!err.empty()) return set_error(DISSEM_ECUDA, "VMM allocation of layer %d on node %d: %s", layer, node, err.c_str());
        lb.vmm = true;
        lb.base = reinterpret_cast<uint8_t *>(lb.blk.va);
    } else {
        auto rit = ctx->reserved.find(std::make_pair(ds->dev, lb.alloc_bytes()));
        if (rit != ctx->reserved.end() && !rit->second.empty()) {
            lb.base = rit->second.back();                 // set aside by dissem_layer_reserve: no cudaMalloc on this path
            rit->second.pop_back();
        } else {
            void *p = nullptr;
            cudaError_t e = cudaMalloc(&p, (size_t)lb.alloc_bytes());
            if (e == cudaErrorMemoryAllocation) {
                cudaGetLastError();
                return set_error(DISSEM_ENOMEM, "out of HBM allocating %lld bytes for layer %d on node %d", (long long)lb.alloc_bytes(), layer, node);
            }
            CUDA_TRY(e);
            lb.base = static_cast<uint8_t *>(p);
        }
    }
    CUDA_TRY(cudaMemsetAsync(lb.base + lb.trailer_off(), 0, (size_t)lb.trailer_bytes(), ds->stream));
    CUDA_TRY(mark_ready(lb, *ds));
    ctx->layers[node][layer] = lb;
    return DISSEM_OK;
}

int dissem_layer_reserve(dissem_ctx *ctx, int node, int64_t size, int count) {
    if (!ctx || count < 0) return set_error(DISSEM_EINVAL, "dissem_layer_reserve: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    DeviceState *ds = nullptr;
    int rc = local_dev(ctx, node, &ds);
    if (rc) return rc;
    LayerBuf shape;
    shape.size = size < 0 ? 0 : size;
    CUDA_TRY(cudaSetDevice(ds->dev));
    auto &bucket = ctx->reserved[std::make_pair(ds->dev, shape.alloc_bytes())];
    for (int i = 0; i < count; ++i) {
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, (size_t)shape.alloc_bytes());
        if (e == cudaErrorMemoryAllocation) {
            cudaGetLastError();
            return set_error(DISSEM_ENOMEM, "out of HBM reserving %lld bytes on node %d's GPU", (long long)shape.alloc_bytes(), node);
        }
        CUDA_TRY(e);
        bucket.push_back(static_cast<uint8_t *>(p));
    }
    return DISSEM_OK;
}

int dissem_layer_free(dissem_ctx *ctx, int node, int layer) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *lb = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    for (auto &ds : ctx->devs) {          // nothing in flight may still touch it
        CUDA_TRY(cudaSetDevice(ds.dev));
        CUDA_TRY(cudaStreamSynchronize(ds.stream));
    }
    // multicast objects that bind this buffer go first
    for (auto it = ctx->mcasts.begin(); it != ctx->mcasts.end();) {
        const std::vector<int> &team = it->first.second;
        if (it->first.first == layer && std::find(team.begin(), team.end(), node) != team.end()) {
            mcast_destroy(&it->second);
            it = ctx->mcasts.erase(it);
        } else ++it;
    }
    drop_ready(*lb);
    if (lb->host) { /* caller-owned memory: nothing to release */ }
    else if (lb->vmm) vmm_free(&lb->blk);                       // owned or imported through a file descriptor
    else if (lb->imported) CUDA_TRY(cudaIpcCloseMemHandle(lb->base));
    else {
        CUDA_TRY(cudaSetDevice(lb->device));
        CUDA_TRY(cudaFree(lb->base));
    }
    ctx->layers[node].erase(layer);
    return DISSEM_OK;
}

int dissem_layer_info(dissem_ctx *ctx, int node, int layer, int64_t *size_out, void **device_ptr_out) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *lb = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    if (size_out) *size_out = lb->size;
    if (device_ptr_out) *device_ptr_out = lb->base;
    return DISSEM_OK;
}

int dissem_layer_fill(dissem_ctx *ctx, int node, int layer, int kind, uint64_t seed) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *lb = nullptr;
    DeviceState *ds = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    rc = local_dev(ctx, node, &ds);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(ds->dev));
    if (lb->size == 0) return DISSEM_OK;
    switch (kind) {
        case DISSEM_FILL_ZERO: CUDA_TRY(cudaMemsetAsync(lb->base, 0, (size_t)lb->size, ds->stream)); break;
        case DISSEM_FILL_POISON: CUDA_TRY(cudaMemsetAsync(lb->base, 0xA5, (size_t)lb->size, ds->stream)); break;
        case DISSEM_FILL_PATTERN:
            CUDA_TRY(launch_fill_pattern(lb->base, lb->size, (unsigned long long)(unsigned int)layer, seed, ds->n_sm, ds->stream));
            break;
        default: return set_error(DISSEM_EINVAL, "unknown fill kind %d", kind);
    }
    if (!lb->host) CUDA_TRY(mark_ready(*lb, *ds));
    return DISSEM_OK;
}

int dissem_layer_reset(dissem_ctx *ctx, int node, int layer) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *lb = nullptr;
    DeviceState *ds = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    rc = local_dev(ctx, node, &ds);
    if (rc) return rc;
    if (lb->host) return DISSEM_OK;
    CUDA_TRY(cudaSetDevice(ds->dev));
    CUDA_TRY(cudaMemsetAsync(lb->base + lb->trailer_off(), 0, (size_t)lb->trailer_bytes(), ds->stream));
    CUDA_TRY(mark_ready(*lb, *ds));
    return DISSEM_OK;
}

int dissem_layer_write(dissem_ctx *ctx, int node, int layer, const void *host, int64_t off, int64_t size) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *lb = nullptr;
    DeviceState *ds = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    rc = local_dev(ctx, node, &ds);
    if (rc) return rc;
    if (off < 0 || size < 0 || off + size > lb->size) return set_error(DISSEM_ERANGE, "write [%lld,+%lld) outside layer of %lld bytes", (long long)off, (long long)size, (long long)lb->size);
    if (size == 0) return DISSEM_OK;
    if (!host) return set_error(DISSEM_EINVAL, "null host pointer");
    CUDA_TRY(cudaSetDevice(ds->dev));
    CUDA_TRY(cudaMemcpyAsync(lb->base + off, host, (size_t)size, cudaMemcpyHostToDevice, ds->stream));
    if (!lb->host) CUDA_TRY(mark_ready(*lb, *ds));
    return DISSEM_OK;
}

// ---- read-back side ---------------------------------------------------------------------------------------------
// Reads of a layer (bytes, checksum table, received counter, verify pass) follow one discipline: look the buffer up under
// ctx->mu and copy what is needed out of the store; RELEASE ctx->mu; wait for the streams that may still be writing the
// buffer (its owner's and those recorded in LayerBuf::writers); then copy on the owner's read stream.  Other threads
// keep calling dissem_send meanwhile (the reference issues one goroutine per Send, node.go:343, :598).  The caller must
// not free a layer another thread is reading.
struct LayerView {
    uint8_t *base = nullptr;
    int64_t size = 0, n_chunks = 0, trailer_off = 0;
    bool host = false;
    int dev_slot = -1;
    std::vector<int> wait_slots;       // device slots whose transfer stream may still be writing the buffer
};

// caller holds ctx->mu
static int view_layer(dissem_ctx *ctx, int node, int layer, LayerView *v) {
    LayerBuf *lb = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    DeviceState *ds = nullptr;
    rc = local_dev(ctx, node, &ds);
    if (rc) return rc;
    v->base = lb->base;
    v->size = lb->size;
    v->n_chunks = lb->n_chunks();
    v->trailer_off = lb->trailer_off();
    v->host = lb->host;
    v->dev_slot = ctx->node_dev_slot[node];
    v->wait_slots.push_back(v->dev_slot);
    for (int sl = 0; sl < (int)ctx->devs.size() && sl < 64; ++sl)
        if (sl != v->dev_slot && ((lb->writers >> sl) & 1ull)) v->wait_slots.push_back(sl);
    return 0;
}

// no lock held (ctx->devs is fixed after dissem_init)
static int wait_writers(dissem_ctx *ctx, const LayerView &v) {
    for (int sl : v.wait_slots) {
        CUDA_TRY(cudaSetDevice(ctx->devs[sl].dev));
        CUDA_TRY(cudaStreamSynchronize(ctx->devs[sl].stream));
    }
    return 0;
}

int dissem_layer_read(dissem_ctx *ctx, int node, int layer, void *host, int64_t off, int64_t size) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    LayerView v;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        int rc = view_layer(ctx, node, layer, &v);
        if (rc) return rc;
    }
    if (off < 0 || size < 0 || off + size > v.size) return set_error(DISSEM_ERANGE, "read [%lld,+%lld) outside layer of %lld bytes", (long long)off, (long long)size, (long long)v.size);
    if (size == 0) return DISSEM_OK;
    if (!host) return set_error(DISSEM_EINVAL, "null host pointer");
    int rc = wait_writers(ctx, v);
    if (rc) return rc;
    DeviceState &ds = ctx->devs[v.dev_slot];
    CUDA_TRY(cudaSetDevice(ds.dev));
    CUDA_TRY(cudaMemcpyAsync(host, v.base + off, (size_t)size, cudaMemcpyDeviceToHost, ds.rd_stream));
    CUDA_TRY(cudaStreamSynchronize(ds.rd_stream));
    return DISSEM_OK;
}

int dissem_host_alloc(void **out, int64_t size) {
    if (!out || size < 0) return set_error(DISSEM_EINVAL, "dissem_host_alloc: bad arguments");
    CUDA_TRY(cudaHostAlloc(out, (size_t)std::max<int64_t>(size, 1), cudaHostAllocPortable | cudaHostAllocMapped));
    return DISSEM_OK;
}
int dissem_host_free(void *p) {
    if (!p) return DISSEM_OK;
    CUDA_TRY(cudaFreeHost(p));
    return DISSEM_OK;
}

int dissem_layer_attach_host(dissem_ctx *ctx, int node, int layer, void *pinned, int64_t size) {
    if (!ctx || (!pinned && size > 0) || size < 0) return set_error(DISSEM_EINVAL, "dissem_layer_attach_host: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    DeviceState *ds = nullptr;
    int rc = local_dev(ctx, node, &ds);
    if (rc) return rc;
    if (ctx->layers[node].count(layer)) return set_error(DISSEM_EEXIST, "node %d already holds layer %d", node, layer);
    if (reinterpret_cast<uintptr_t>(pinned) % 16) return set_error(DISSEM_EINVAL, "host layers must be 16-byte aligned");
    CUDA_TRY(cudaSetDevice(ds->dev));
    void *dptr = nullptr;
    cudaError_t e = cudaHostGetDevicePointer(&dptr, pinned, 0);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return set_error(DISSEM_EINVAL, "the buffer is not pinned, GPU-mapped host memory (use dissem_host_alloc): %s", cudaGetErrorString(e));
    }
    LayerBuf lb;
    lb.size = size;
    lb.base = static_cast<uint8_t *>(dptr);
    lb.device = ds->dev;
    lb.host = true;
    ctx->layers[node][layer] = lb;
    return DISSEM_OK;
}

int dissem_layer_attach_file(dissem_ctx *ctx, int node, int layer, const char *path, int64_t file_off, int64_t size) {
    if (!ctx || !path || file_off < 0 || size < 0) return set_error(DISSEM_EINVAL, "dissem_layer_attach_file: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    DeviceState *ds = nullptr;
    int rc = local_dev(ctx, node, &ds);               // the node's GPU runs the copy kernel that reads the staged pieces
    if (rc) return rc;
    if (ctx->files[node].count(layer)) return set_error(DISSEM_EEXIST, "node %d already has a disk copy of layer %d attached", node, layer);
    FileSrc f;
    f.fd = open(path, O_RDONLY | O_CLOEXEC);
    if (f.fd < 0) return set_error(DISSEM_EIO, "open %s: %s", path, strerror(errno));
    struct stat st;
    memset(&st, 0, sizeof st);
    if (fstat(f.fd, &st) != 0 || (int64_t)st.st_size < file_off + size) {
        const int64_t have = (int64_t)st.st_size;
        close(f.fd);
        return set_error(DISSEM_EIO, "%s holds %lld bytes, fewer than offset %lld + layer size %lld", path, (long long)have,
                         (long long)file_off, (long long)size);
    }
    f.file_off = file_off;
    f.size = size;
    f.path = path;
    ctx->files[node][layer] = f;
    return DISSEM_OK;
}

int dissem_layer_detach_file(dissem_ctx *ctx, int node, int layer) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    // a transfer that is being streamed holds its ring's mutex, not ctx->mu: wait for those first
    std::vector<std::shared_ptr<std::mutex>> rings;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        for (auto &ds : ctx->devs) rings.push_back(ds.ring_mu);
    }
    for (auto &m : rings) std::lock_guard<std::mutex> wait_for_stream(*m);
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = check_node(ctx, node);
    if (rc) return rc;
    auto it = ctx->files[node].find(layer);
    if (it == ctx->files[node].end()) return set_error(DISSEM_ENOENT, "node %d holds no disk copy of layer %d", node, layer);
    close(it->second.fd);
    ctx->files[node].erase(it);
    return DISSEM_OK;
}

int dissem_layer_file_info(dissem_ctx *ctx, int node, int layer, int64_t *size_out) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = check_node(ctx, node);
    if (rc) return rc;
    auto it = ctx->files[node].find(layer);
    if (it == ctx->files[node].end()) return set_error(DISSEM_ENOENT, "node %d holds no disk copy of layer %d", node, layer);
    if (size_out) *size_out = it->second.size;
    return DISSEM_OK;
}

int dissem_layer_export(dissem_ctx *ctx, int node, int layer, uint8_t handle_out[64]) {
    if (!ctx || !handle_out) return set_error(DISSEM_EINVAL, "dissem_layer_export: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *lb = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    if (lb->imported) return set_error(DISSEM_EINVAL, "layer %d of node %d is itself imported", layer, node);
    if (lb->host) return set_error(DISSEM_ENOTSUP, "layer %d of node %d is attached host memory; peers read pinned host memory of their own process only", layer, node);
    if (lb->vmm) return set_error(DISSEM_ENOTSUP, "layer %d of node %d is VMM-backed (alloc_vmm): export it with dissem_layer_export_fd", layer, node);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaSetDevice(lb->device));
    // the peer process cannot wait for this process's stream: whatever the owner enqueued for the buffer (trailer
    // zeroing, fill, upload) has completed before the handle leaves
    if (ctx->node_dev_slot[node] >= 0) CUDA_TRY(cudaStreamSynchronize(ctx->devs[ctx->node_dev_slot[node]].stream));
    CUDA_TRY(cudaIpcGetMemHandle(&h, lb->base));
    memcpy(handle_out, &h, 64);
    return DISSEM_OK;
}

int dissem_layer_import(dissem_ctx *ctx, int node, int layer, int64_t size, const uint8_t handle[64]) {
    if (!ctx || !handle) return set_error(DISSEM_EINVAL, "dissem_layer_import: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = check_node(ctx, node);
    if (rc) return rc;
    if (ctx->node_dev_slot[node] >= 0) return set_error(DISSEM_EINVAL, "node %d is local; import is for remote nodes", node);
    if (ctx->layers[node].count(layer)) return set_error(DISSEM_EEXIST, "layer %d of node %d already imported", layer, node);
    if (ctx->devs.empty()) return set_error(DISSEM_EINVAL, "no local device to map the layer on");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    // Map through the first local device; with peer access enabled by the open, every kernel of this
    // process that runs there can store into it.
    CUDA_TRY(cudaSetDevice(ctx->devs[0].dev));
    void *p = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    LayerBuf lb;
    lb.size = size < 0 ? 0 : size;
    lb.base = static_cast<uint8_t *>(p);
    lb.imported = true;
    lb.device = -1;
    ctx->layers[node][layer] = lb;
    return DISSEM_OK;
}

int dissem_layer_export_fd(dissem_ctx *ctx, int node, int layer, int *fd_out) {
    if (!ctx || !fd_out) return set_error(DISSEM_EINVAL, "dissem_layer_export_fd: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *lb = nullptr;
    int rc = find_layer(ctx, node, layer, &lb);
    if (rc) return rc;
    if (!lb->vmm || lb->imported) return set_error(DISSEM_EINVAL, "layer %d of node %d is not a locally owned shareable VMM layer (set alloc_vmm = 2 before allocating)", layer, node);
    if (ctx->node_dev_slot[node] >= 0) {             // as in dissem_layer_export: nothing of the owner's is still in flight
        CUDA_TRY(cudaSetDevice(ctx->devs[ctx->node_dev_slot[node]].dev));
        CUDA_TRY(cudaStreamSynchronize(ctx->devs[ctx->node_dev_slot[node]].stream));
    }
    std::string err = vmm_export_fd(lb->blk, fd_out);
    if (!err.empty()) return set_error(DISSEM_ECUDA, "%s", err.c_str());
    return DISSEM_OK;
}

int dissem_layer_import_fd(dissem_ctx *ctx, int node, int layer, int64_t size, int fd) {
    if (!ctx || fd < 0) return set_error(DISSEM_EINVAL, "dissem_layer_import_fd: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = check_node(ctx, node);
    if (rc) return rc;
    if (ctx->node_dev_slot[node] >= 0) return set_error(DISSEM_EINVAL, "node %d is local; import is for remote nodes", node);
    if (ctx->layers[node].count(layer)) return set_error(DISSEM_EEXIST, "layer %d of node %d already imported", layer, node);
    std::vector<int> access;
    for (auto &d : ctx->devs) access.push_back(d.dev);
    LayerBuf lb;
    lb.size = size < 0 ? 0 : size;
    std::string err = vmm_import_fd(fd, -1, access, (size_t)lb.alloc_bytes(), &lb.blk);
    if (!err.empty()) return set_error(DISSEM_ECUDA, "%s", err.c_str());
    lb.base = reinterpret_cast<uint8_t *>(lb.blk.va);
    lb.imported = true;
    lb.vmm = true;
    lb.device = -1;
    ctx->layers[node][layer] = lb;
    return DISSEM_OK;
}

static std::pair<int, std::vector<int>> mcast_key(int layer, int src_node, const int *dst_nodes, int n_dst) {
    std::vector<int> team{src_node};
    std::vector<int> d(dst_nodes, dst_nodes + n_dst);
    std::sort(d.begin(), d.end());
    team.insert(team.end(), d.begin(), d.end());
    return std::make_pair(layer, team);
}

int dissem_mcast_create(dissem_ctx *ctx, int layer, int src_node, const int *dst_nodes, int n_dst, int *fd_out) {
    if (!ctx || !dst_nodes || n_dst < 1 || !fd_out) return set_error(DISSEM_EINVAL, "dissem_mcast_create: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *src = nullptr;
    DeviceState *ds = nullptr;
    int rc = find_layer(ctx, src_node, layer, &src);
    if (rc) return rc;
    rc = local_dev(ctx, src_node, &ds);
    if (rc) return rc;
    if (!src->vmm || src->imported) return set_error(DISSEM_EINVAL, "the source copy of layer %d is not a shareable VMM layer of this process", layer);
    auto key = mcast_key(layer, src_node, dst_nodes, n_dst);
    if (ctx->mcasts.count(key)) return set_error(DISSEM_EEXIST, "multicast object for layer %d already exists", layer);
    McObject mc;
    std::string err = mcast_create_shared(n_dst + 1, src->blk.bytes, ds->dev, &mc, fd_out);
    if (!err.empty()) return set_error(DISSEM_ENOTSUP, "%s", err.c_str());
    ctx->mcasts.emplace(key, mc);
    return DISSEM_OK;
}

int dissem_mcast_join(dissem_ctx *ctx, int layer, int src_node, const int *dst_nodes, int n_dst, int my_node, int fd) {
    if (!ctx || !dst_nodes || n_dst < 1 || fd < 0) return set_error(DISSEM_EINVAL, "dissem_mcast_join: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    LayerBuf *mine = nullptr;
    DeviceState *ds = nullptr;
    int rc = find_layer(ctx, my_node, layer, &mine);
    if (rc) return rc;
    rc = local_dev(ctx, my_node, &ds);
    if (rc) return rc;
    if (!mine->vmm || mine->imported) return set_error(DISSEM_EINVAL, "node %d's copy of layer %d is not a shareable VMM layer of this process", my_node, layer);
    auto key = mcast_key(layer, src_node, dst_nodes, n_dst);
    if (ctx->mcasts.count(key)) return set_error(DISSEM_EEXIST, "multicast object for layer %d already joined", layer);
    McObject mc;
    std::string err = mcast_join(fd, mine->blk.bytes, ds->dev, &mc);
    if (!err.empty()) return set_error(DISSEM_ENOTSUP, "%s", err.c_str());
    ctx->mcasts.emplace(key, mc);
    return DISSEM_OK;
}

int dissem_mcast_bind(dissem_ctx *ctx, int layer, int src_node, const int *dst_nodes, int n_dst) {
    if (!ctx || !dst_nodes || n_dst < 1) return set_error(DISSEM_EINVAL, "dissem_mcast_bind: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto key = mcast_key(layer, src_node, dst_nodes, n_dst);
    auto it = ctx->mcasts.find(key);
    if (it == ctx->mcasts.end() || !it->second.handle) return set_error(DISSEM_ENOENT, "no multicast object for layer %d: create or join it first", layer);
    for (int node : key.second) {
        if (ctx->node_dev_slot[node] < 0) continue;                   // another process binds that member
        LayerBuf *lb = nullptr;
        int rc = find_layer(ctx, node, layer, &lb);
        if (rc) return rc;
        if (!lb->vmm || lb->imported) return set_error(DISSEM_EINVAL, "node %d's copy of layer %d cannot be bound", node, layer);
        std::string err = mcast_bind_block(&it->second, lb->blk);
        if (!err.empty()) return set_error(DISSEM_ENOTSUP, "%s", err.c_str());
    }
    if (ctx->node_dev_slot[src_node] >= 0 && !it->second.valid()) {
        std::vector<int> access{ctx->devs[ctx->node_dev_slot[src_node]].dev};
        std::string err = mcast_map(&it->second, access);
        if (!err.empty()) return set_error(DISSEM_ENOTSUP, "%s", err.c_str());
    }
    return DISSEM_OK;
}

int dissem_schedule_compile(dissem_ctx *ctx, const dissem_xfer *xfers, int n, dissem_sched *out) {
    if (!ctx || !out) return set_error(DISSEM_EINVAL, "dissem_schedule_compile: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Sched s;
    int rc = compile_sched(ctx, xfers, n, &s);
    if (rc) {
        free_sched(ctx, s);
        return rc;
    }
    int64_t id = ctx->next_id++;
    ctx->scheds[id] = std::move(s);
    *out = id;
    return DISSEM_OK;
}

int dissem_schedule_launch(dissem_ctx *ctx, dissem_sched sched, dissem_ticket *out) {
    if (!ctx || !out) return set_error(DISSEM_EINVAL, "dissem_schedule_launch: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->scheds.find(sched);
    if (it == ctx->scheds.end()) return set_error(DISSEM_ENOENT, "unknown schedule %lld", (long long)sched);
    Ticket t;
    int rc = launch_sched(ctx, it->second, &t);
    if (rc) {
        drop_ticket(ctx, t);
        return rc;
    }
    int64_t id = ctx->next_id++;
    ctx->tickets[id] = std::move(t);
    *out = id;
    return DISSEM_OK;
}

int dissem_schedule_free(dissem_ctx *ctx, dissem_sched sched) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->scheds.find(sched);
    if (it == ctx->scheds.end()) return set_error(DISSEM_ENOENT, "unknown schedule %lld", (long long)sched);
    for (auto &p : it->second.parts) {       // the kernels read the descriptor table
        CUDA_TRY(cudaSetDevice(ctx->devs[p.dev_slot].dev));
        CUDA_TRY(cudaStreamSynchronize(ctx->devs[p.dev_slot].stream));
    }
    free_sched(ctx, it->second);
    ctx->scheds.erase(it);
    return DISSEM_OK;
}

int dissem_schedule_info(dissem_ctx *ctx, dissem_sched sched, int64_t *units_out, int *launches_out, int64_t *bytes_out) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->scheds.find(sched);
    if (it == ctx->scheds.end()) return set_error(DISSEM_ENOENT, "unknown schedule %lld", (long long)sched);
    if (units_out) *units_out = it->second.units;
    if (launches_out) {                                                     // copy_csum (+ zero_csum where a part needs it) per executing GPU and round
        int n_launch = 0;
        for (auto &p : it->second.parts) n_launch += p.needs_zero ? 2 : 1;
        *launches_out = n_launch;
    }
    if (bytes_out) *bytes_out = it->second.bytes;
    return DISSEM_OK;
}

static int submit_compiled(dissem_ctx *ctx, const dissem_xfer *xfers, int n, dissem_ticket *out) {
    dissem_sched s = 0;
    int rc = dissem_schedule_compile(ctx, xfers, n, &s);
    if (rc) return rc;
    rc = dissem_schedule_launch(ctx, s, out);
    // the descriptor table must outlive the kernels: free it when the ticket completes
    if (rc == DISSEM_OK) {
        std::lock_guard<std::mutex> lk(ctx->mu);
        // keep the schedule id inside the ticket by parking it under the negative ticket key
        ctx->scheds[-(*out)] = std::move(ctx->scheds[s]);
        ctx->scheds.erase(s);
    } else {
        dissem_schedule_free(ctx, s);
    }
    return rc;
}

int dissem_schedule_submit(dissem_ctx *ctx, const dissem_xfer *xfers, int n, dissem_ticket *out) {
    if (!ctx || !out || n < 0 || (n > 0 && !xfers)) return set_error(DISSEM_EINVAL, "dissem_schedule_submit: bad arguments");
    bool any_file = false;
    for (int i = 0; i < n; ++i) any_file = any_file || (xfers[i].flags & DISSEM_XFER_FILE);
    if (!any_file) return submit_compiled(ctx, xfers, n, out);

    // Disk tier: the transfers from disk copies are streamed first (this call blocks while their files are read),
    // the rest of the plan is launched behind them, and one ticket covers both.
    std::vector<dissem_xfer> rest;
    std::vector<FileJob> jobs;
    struct CloseFds {
        std::vector<FileJob> &js;
        ~CloseFds() {
            for (auto &j : js)
                if (j.owns_fd) close(j.file.fd);
        }
    } close_fds{jobs};
    LaunchCfg cfg{};
    int64_t unit_bytes = 0;
    int file_threads = 1;
    auto give_back = [&](std::vector<FileJob> &js) {                  // caller holds ctx->mu
        Ticket t;
        for (auto &j : js)
            if (j.tp.dev_slot >= 0) t.parts.push_back(j.tp);
        drop_ticket(ctx, t);
    };
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        for (int i = 0; i < n; ++i) {
            if (!(xfers[i].flags & DISSEM_XFER_FILE)) { rest.push_back(xfers[i]); continue; }
            int rc = check_node(ctx, xfers[i].src_node);
            if (rc) { give_back(jobs); return rc; }
            if (ctx->node_dev_slot[xfers[i].src_node] < 0) continue;          // the process that holds the file streams it
            FileJob job;
            rc = prepare_file_job(ctx, xfers[i], &job);
            if (rc) {
                jobs.push_back(job);
                give_back(jobs);
                return rc;
            }
            jobs.push_back(job);
        }
        cfg = LaunchCfg{0, (int)ctx->opt_ctas_per_sm, 0, (int)ctx->opt_reg_cfg, (int)ctx->opt_tma_cfg};
        unit_bytes = ctx->opt_unit_bytes;
        file_threads = (int)ctx->opt_file_threads;
    }
    Ticket ft;
    int rc = DISSEM_OK;
    // Files held by nodes of different GPUs go up different PCIe links: one streaming thread per GPU, each working
    // through its own jobs in order; the reader threads are shared out so that the host cores are not oversubscribed.
    std::vector<int> slots;
    for (auto &job : jobs)
        if (std::find(slots.begin(), slots.end(), job.dev_slot) == slots.end()) slots.push_back(job.dev_slot);
    auto run_slot = [&](int slot, int readers, int *rc_out, std::string *err_out) {
        for (auto &job : jobs) {
            if (job.dev_slot != slot) continue;
            LaunchCfg c = cfg;
            c.n_sm = ctx->devs[slot].n_sm;
            const int r = stream_from_file(ctx, job, c, unit_bytes, readers);
            if (r) {                                   // the error text is thread-local: carry it to the caller
                *rc_out = r;
                *err_out = last_error();
                return;
            }
        }
    };
    if (slots.size() <= 1) {
        std::string err;
        if (!slots.empty()) run_slot(slots[0], file_threads, &rc, &err);
    } else {
        const int cores = (int)std::max(1u, std::thread::hardware_concurrency());
        const int readers = std::max(1, std::min(file_threads, cores / (int)slots.size()));
        std::vector<int> rcs(slots.size(), 0);
        std::vector<std::string> errs(slots.size());
        std::vector<std::thread> streamers;
        for (size_t i = 0; i < slots.size(); ++i) streamers.emplace_back(run_slot, slots[i], readers, &rcs[i], &errs[i]);
        for (auto &t : streamers) t.join();
        for (size_t i = 0; i < slots.size(); ++i)
            if (rcs[i] && !rc) rc = set_error(rcs[i], "%s", errs[i].c_str());
    }
    if (rc == DISSEM_OK && !rest.empty()) {
        // "behind them" on every GPU, not only on the streams that carried the files: a transfer of the rest may forward
        // what a disk transfer has just loaded into another node's HBM (dissem_plan_optimize plans exactly that)
        for (auto &ds : ctx->devs) {
            cudaSetDevice(ds.dev);
            for (auto &job : jobs)
                if (rc == DISSEM_OK && cudaStreamWaitEvent(ds.stream, job.tp.end, 0) != cudaSuccess)
                    rc = set_error(DISSEM_ECUDA, "cudaStreamWaitEvent: %s", cudaGetErrorString(cudaGetLastError()));
        }
        if (rc == DISSEM_OK) rc = submit_compiled(ctx, rest.data(), (int)rest.size(), out);
    }
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (rc) {
        // what was enqueued drains on its own; the events go back once the streams are idle
        for (auto &ds : ctx->devs) {
            cudaSetDevice(ds.dev);
            cudaStreamSynchronize(ds.stream);
        }
        give_back(jobs);
        return rc;
    }
    for (auto &job : jobs) ft.parts.push_back(job.tp);
    if (rest.empty()) {
        const int64_t id = ctx->next_id++;
        ctx->tickets[id] = std::move(ft);
        *out = id;
    } else {
        Ticket &t = ctx->tickets[*out];
        t.parts.insert(t.parts.begin(), ft.parts.begin(), ft.parts.end());
    }
    return DISSEM_OK;
}

int dissem_send(dissem_ctx *ctx, int src_node, int dst_node, int layer, int64_t off, int64_t size, int64_t total,
                dissem_ticket *out) {
    dissem_xfer x;
    memset(&x, 0, sizeof x);
    x.src_node = src_node;
    x.dst_node = dst_node;
    x.layer = layer;
    x.off = off;
    x.size = size;
    x.total = total;
    return dissem_schedule_submit(ctx, &x, 1, out);
}

int dissem_send_file(dissem_ctx *ctx, int src_node, int dst_node, int layer, int64_t off, int64_t size, int64_t total,
                     dissem_ticket *out) {
    dissem_xfer x;
    memset(&x, 0, sizeof x);
    x.src_node = src_node;
    x.dst_node = dst_node;
    x.layer = layer;
    x.flags = DISSEM_XFER_FILE;
    x.off = off;
    x.size = size;
    x.total = total;
    return dissem_schedule_submit(ctx, &x, 1, out);
}

int dissem_node_device(dissem_ctx *ctx, int node, int *device_out) {
    if (!ctx || !device_out) return set_error(DISSEM_EINVAL, "dissem_node_device: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = check_node(ctx, node);
    if (rc) return rc;
    const int slot = ctx->node_dev_slot[node];
    *device_out = slot < 0 ? -1 : ctx->devs[slot].dev;
    return DISSEM_OK;
}

int dissem_wait(dissem_ctx *ctx, dissem_ticket ticket, double *ms_out) { return dissem_wait_detail(ctx, ticket, ms_out, nullptr); }

int dissem_wait_detail(dissem_ctx *ctx, dissem_ticket ticket, double *ms_out, double *copy_ms_out) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    Ticket t;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        auto it = ctx->tickets.find(ticket);
        if (it == ctx->tickets.end()) return set_error(DISSEM_ENOENT, "unknown ticket %lld", (long long)ticket);
        t = std::move(it->second);
        ctx->tickets.erase(it);
    }
    // wait outside the lock: other goroutines keep calling Send meanwhile (node.go:343)
    double ms = 0.0, copy_ms = 0.0;
    int rc = DISSEM_OK;
    std::map<int, std::pair<cudaEvent_t, cudaEvent_t>> span;      // device slot -> (first start, last end)
    std::map<int, double> copy_sum;
    for (auto &p : t.parts) {
        cudaError_t e = cudaEventSynchronize(p.end);
        if (e != cudaSuccess) {
            rc = set_error(DISSEM_ECUDA, "transfer failed: %s", cudaGetErrorString(e));
            break;
        }
        auto it = span.find(p.dev_slot);
        if (it == span.end()) span[p.dev_slot] = std::make_pair(p.start, p.end);
        else it->second.second = p.end;
        float f = 0.f;
        if (cudaEventElapsedTime(&f, p.mid, p.end) == cudaSuccess) copy_sum[p.dev_slot] += (double)f;
    }
    if (rc == DISSEM_OK)
        for (auto &kv : span) {
            float f = 0.f;
            cudaError_t e = cudaEventElapsedTime(&f, kv.second.first, kv.second.second);
            if (e != cudaSuccess) {
                rc = set_error(DISSEM_ECUDA, "cudaEventElapsedTime: %s", cudaGetErrorString(e));
                break;
            }
            ms = std::max(ms, (double)f);
            copy_ms = std::max(copy_ms, copy_sum[kv.first]);
        }
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        drop_ticket(ctx, t);
        auto sit = ctx->scheds.find(-ticket);       // schedule parked by dissem_schedule_submit
        if (sit != ctx->scheds.end()) {
            free_sched(ctx, sit->second);
            ctx->scheds.erase(sit);
        }
    }
    if (rc == DISSEM_OK && ms_out) *ms_out = ms;
    if (rc == DISSEM_OK && copy_ms_out) *copy_ms_out = copy_ms;
    return rc;
}

int dissem_poll(dissem_ctx *ctx, dissem_ticket ticket, int *done_out) {
    if (!ctx || !done_out) return set_error(DISSEM_EINVAL, "dissem_poll: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->tickets.find(ticket);
    if (it == ctx->tickets.end()) return set_error(DISSEM_ENOENT, "unknown ticket %lld", (long long)ticket);
    int done = 1;
    for (auto &p : it->second.parts) {
        cudaError_t e = cudaEventQuery(p.end);
        if (e == cudaErrorNotReady) { done = 0; cudaGetLastError(); }
        else if (e != cudaSuccess) return set_error(DISSEM_ECUDA, "transfer failed: %s", cudaGetErrorString(e));
    }
    *done_out = done;
    if (done) {
        // the kernels are done with the descriptor table: give the parked schedule (dissem_schedule_submit / dissem_send)
        // back now, so that a caller who only ever polls does not exhaust the descriptor pool.  The ticket itself (its
        // events, its timings) stays until dissem_wait or dissem_ticket_release.
        auto sit = ctx->scheds.find(-ticket);
        if (sit != ctx->scheds.end()) {
            free_sched(ctx, sit->second);
            ctx->scheds.erase(sit);
        }
    }
    return DISSEM_OK;
}

int dissem_ticket_release(dissem_ctx *ctx, dissem_ticket ticket) {
    if (!ctx) return set_error(DISSEM_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->tickets.find(ticket);
    if (it == ctx->tickets.end()) return set_error(DISSEM_ENOENT, "unknown ticket %lld", (long long)ticket);
    for (auto &p : it->second.parts) {
        cudaError_t e = cudaEventQuery(p.end);
        if (e == cudaErrorNotReady) {
            cudaGetLastError();
            return set_error(DISSEM_EINVAL, "ticket %lld is still in flight: poll until done, or wait", (long long)ticket);
        }
    }
    drop_ticket(ctx, it->second);
    ctx->tickets.erase(it);
    auto sit = ctx->scheds.find(-ticket);
    if (sit != ctx->scheds.end()) {
        free_sched(ctx, sit->second);
        ctx->scheds.erase(sit);
    }
    return DISSEM_OK;
}

static int read_trailer(dissem_ctx *ctx, int node, int layer, int64_t byte_off, void *out, int64_t bytes, int64_t expect_chunks) {
    LayerView v;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        int rc = view_layer(ctx, node, layer, &v);
        if (rc) return rc;
    }
    if (v.host) return set_error(DISSEM_ENOTSUP, "an attached host layer has no checksum table or received counter (it is a source)");
    if (expect_chunks >= 0 && expect_chunks != v.n_chunks)
        return set_error(DISSEM_EINVAL, "layer %d has %lld chunks, caller passed %lld", layer, (long long)v.n_chunks, (long long)expect_chunks);
    if (bytes == 0) return DISSEM_OK;
    int rc = wait_writers(ctx, v);
    if (rc) return rc;
    DeviceState &ds = ctx->devs[v.dev_slot];
    CUDA_TRY(cudaSetDevice(ds.dev));
    CUDA_TRY(cudaMemcpyAsync(out, v.base + v.trailer_off + byte_off, (size_t)bytes, cudaMemcpyDeviceToHost, ds.rd_stream));
    CUDA_TRY(cudaStreamSynchronize(ds.rd_stream));
    return DISSEM_OK;
}

int dissem_layer_checksums(dissem_ctx *ctx, int node, int layer, uint64_t *out, int64_t n_chunks) {
    if (!ctx || (!out && n_chunks > 0) || n_chunks < 0) return set_error(DISSEM_EINVAL, "dissem_layer_checksums: bad arguments");
    return read_trailer(ctx, node, layer, 256, out, n_chunks * 16, n_chunks);
}

int dissem_layer_received(dissem_ctx *ctx, int node, int layer, int64_t *bytes_out) {
    if (!ctx || !bytes_out) return set_error(DISSEM_EINVAL, "dissem_layer_received: bad arguments");
    unsigned long long v = 0;
    int rc = read_trailer(ctx, node, layer, 0, &v, 8, -1);
    if (rc) return rc;
    *bytes_out = (int64_t)v;
    return DISSEM_OK;
}

// The fused tables of several local layers with one device-side gather and ONE device-to-host copy per GPU (what a
// leader reads after a batch: one table per (node, layer) of its Assignment).  out receives the tables back to back in
// the order given; total_chunks = sum of the layers' chunk counts.
int dissem_layers_checksums(dissem_ctx *ctx, const int *nodes, const int *layers, int n, uint64_t *out, int64_t total_chunks) {
    if (!ctx || n < 0 || (n > 0 && (!nodes || !layers)) || (total_chunks > 0 && !out) || total_chunks < 0)
        return set_error(DISSEM_EINVAL, "dissem_layers_checksums: bad arguments");
    std::vector<LayerView> views((size_t)n);
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        int64_t sum = 0;
        for (int i = 0; i < n; ++i) {
            int rc = view_layer(ctx, nodes[i], layers[i], &views[i]);
            if (rc) return rc;
            if (views[i].host) return set_error(DISSEM_ENOTSUP, "layer %d of node %d is attached host memory: no checksum table", layers[i], nodes[i]);
            sum += views[i].n_chunks;
        }
        if (sum != total_chunks) return set_error(DISSEM_EINVAL, "the %d layers have %lld chunks in all, caller passed %lld", n, (long long)sum, (long long)total_chunks);
    }
    std::vector<char> waited(ctx->devs.size(), 0);
    for (auto &v : views)
        for (int sl : v.wait_slots)
            if (!waited[sl]) {
                waited[sl] = 1;
                CUDA_TRY(cudaSetDevice(ctx->devs[sl].dev));
                CUDA_TRY(cudaStreamSynchronize(ctx->devs[sl].stream));
            }
    // per owning GPU: gather list -> arena, gather kernel -> arena, one copy -> pinned arena -> out
    for (size_t sl = 0; sl < ctx->devs.size(); ++sl) {
        DeviceState &ds = ctx->devs[sl];
        std::vector<GatherItem> items;
        std::vector<std::pair<int64_t, int64_t>> place;          // (offset in out, bytes) per item
        int64_t out_off = 0, arena_off = 0;
        for (int i = 0; i < n; ++i) {
            const int64_t bytes = views[i].n_chunks * 16;
            if (views[i].dev_slot == (int)sl && bytes > 0) {
                GatherItem g;
                g.src = reinterpret_cast<const unsigned long long *>(views[i].base + views[i].trailer_off + 256);
                g.dst_word = arena_off / 8;
                g.words = bytes / 8;
                items.push_back(g);
                place.emplace_back(out_off, bytes);
                arena_off += bytes;
            }
            out_off += bytes;
        }
        if (items.empty()) continue;
        std::lock_guard<std::mutex> arena_lock(*ds.arena_mu);
        CUDA_TRY(cudaSetDevice(ds.dev));
        const size_t list_bytes = align_up((int64_t)(items.size() * sizeof(GatherItem)), 256);
        if (list_bytes + (size_t)arena_off > ds.arena_bytes) {
            // beyond the arena (more than ~2 MiB of tables on one GPU): fall back to one copy per layer, still lock-free
            for (size_t k = 0; k < items.size(); ++k)
                CUDA_TRY(cudaMemcpyAsync(reinterpret_cast<uint8_t *>(out) + place[k].first, items[k].src, (size_t)place[k].second,
                                         cudaMemcpyDeviceToHost, ds.rd_stream));
            CUDA_TRY(cudaStreamSynchronize(ds.rd_stream));
            continue;
        }
        memcpy(ds.arena_host, items.data(), items.size() * sizeof(GatherItem));
        CUDA_TRY(cudaMemcpyAsync(ds.arena_dev, ds.arena_host, items.size() * sizeof(GatherItem), cudaMemcpyHostToDevice, ds.rd_stream));
        CUDA_TRY(launch_gather_tables(reinterpret_cast<const GatherItem *>(ds.arena_dev), (int)items.size(),
                                      reinterpret_cast<unsigned long long *>(ds.arena_dev + list_bytes), ds.n_sm, ds.rd_stream));
        CUDA_TRY(cudaMemcpyAsync(ds.arena_host + list_bytes, ds.arena_dev + list_bytes, (size_t)arena_off, cudaMemcpyDeviceToHost, ds.rd_stream));
        CUDA_TRY(cudaStreamSynchronize(ds.rd_stream));
        int64_t pos = 0;
        for (size_t k = 0; k < items.size(); ++k) {
            memcpy(reinterpret_cast<uint8_t *>(out) + place[k].first, ds.arena_host + list_bytes + pos, (size_t)place[k].second);
            pos += place[k].second;
        }
    }
    return DISSEM_OK;
}

int dissem_layer_verify(dissem_ctx *ctx, int node, int layer, uint64_t *out, int64_t n_chunks) {
    if (!ctx || (!out && n_chunks > 0)) return set_error(DISSEM_EINVAL, "dissem_layer_verify: bad arguments");
    LayerView v;
    LaunchCfg cfg{};
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        int rc = view_layer(ctx, node, layer, &v);
        if (rc) return rc;
        cfg = LaunchCfg{ctx->devs[v.dev_slot].n_sm, (int)ctx->opt_ctas_per_sm, v.host ? 0 : (int)ctx->opt_variant, (int)ctx->opt_reg_cfg, (int)ctx->opt_tma_cfg};
    }
    if (n_chunks != v.n_chunks) return set_error(DISSEM_EINVAL, "layer %d has %lld chunks, caller passed %lld", layer, (long long)v.n_chunks, (long long)n_chunks);
    if (n_chunks == 0) return DISSEM_OK;
    int rc = wait_writers(ctx, v);
    if (rc) return rc;
    DeviceState &ds = ctx->devs[v.dev_slot];
    // scratch (counters, one descriptor, a fresh table) comes from the GPU's arena: nothing is allocated or freed here
    const size_t table_bytes = (size_t)n_chunks * 16;
    if (512 + table_bytes > ds.arena_bytes)
        return set_error(DISSEM_ERANGE, "layer %d has %lld chunks; the verify arena holds %lld", layer, (long long)n_chunks,
                         (long long)((ds.arena_bytes - 512) / 16));
    std::lock_guard<std::mutex> arena_lock(*ds.arena_mu);
    CUDA_TRY(cudaSetDevice(ds.dev));
    uint8_t *scratch = ds.arena_dev;
    XferDesc d;
    memset(&d, 0, sizeof d);
    d.src = v.base;
    d.dst = nullptr;
    d.csum = reinterpret_cast<unsigned long long *>(scratch + 512);
    d.received = nullptr;
    d.off = 0;
    d.size = v.size;
    d.total = v.size;
    d.first_unit = 0;
    // its own counters (first 16 bytes of the arena) and the read stream: the pass neither waits for nor delays transfers
    memset(ds.arena_host, 0, 512);
    memcpy(ds.arena_host + 256, &d, sizeof d);
    cudaError_t e = cudaMemsetAsync(scratch + 512, 0, table_bytes, ds.rd_stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(scratch, ds.arena_host, 512, cudaMemcpyHostToDevice, ds.rd_stream);
    if (e == cudaSuccess)
        e = launch_copy_csum(reinterpret_cast<XferDesc *>(scratch + 256), 1, units_of(0, v.size, kChunkShift), kChunkShift,
                             reinterpret_cast<SchedCounters *>(scratch), cfg, ds.rd_stream, nullptr);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ds.arena_host + 512, scratch + 512, table_bytes, cudaMemcpyDeviceToHost, ds.rd_stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ds.rd_stream);
    if (e != cudaSuccess) return set_error(DISSEM_ECUDA, "verify pass failed: %s", cudaGetErrorString(e));
    memcpy(out, ds.arena_host + 512, table_bytes);
    return DISSEM_OK;
}

int dissem_set_option(dissem_ctx *ctx, const char *key, int64_t value) {
    if (!ctx || !key) return set_error(DISSEM_EINVAL, "dissem_set_option: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    std::string k(key);
    if (k == "ctas_per_sm") { if (value < 0 || value > 32) return set_error(DISSEM_EINVAL, "ctas_per_sm out of range"); ctx->opt_ctas_per_sm = value; }
    else if (k == "variant") { if (value < 0 || value > 1) return set_error(DISSEM_EINVAL, "variant must be 0 or 1"); ctx->opt_variant = value; }
    else if (k == "tma_cfg") { if (value < 0 || value > 8) return set_error(DISSEM_EINVAL, "tma_cfg must be 0..8"); ctx->opt_tma_cfg = value; }
    else if (k == "alloc_vmm") { if (value < 0 || value > 2) return set_error(DISSEM_EINVAL, "alloc_vmm must be 0, 1 or 2"); ctx->opt_alloc_vmm = value; }
    else if (k == "mcast") { if (value < 0 || value > 3) return set_error(DISSEM_EINVAL, "mcast must be 0..3"); ctx->opt_mcast = value; }
    else if (k == "wait_src") { if (value < 0 || value > 1) return set_error(DISSEM_EINVAL, "wait_src must be 0 or 1"); ctx->opt_wait_src = value; }
    else if (k == "mcast_variant") { if (value < 0 || value > 1) return set_error(DISSEM_EINVAL, "mcast_variant must be 0 or 1"); ctx->opt_mcast_variant = value; }
    else if (k == "pull") { if (value < 0 || value > 2) return set_error(DISSEM_EINVAL, "pull must be 0, 1 or 2"); ctx->opt_pull = value; }
    else if (k == "file_threads") { if (value < 1 || value > 64) return set_error(DISSEM_EINVAL, "file_threads must be 1..64"); ctx->opt_file_threads = value; }
    else if (k == "reg_cfg") { if (value < 0 || value > 9) return set_error(DISSEM_EINVAL, "reg_cfg must be 0..9"); ctx->opt_reg_cfg = value; }
    else if (k == "unit_bytes") {
        if (value != 0 && (value < 4096 || value > DISSEM_CHUNK_BYTES || (value & (value - 1)))) return set_error(DISSEM_EINVAL, "unit_bytes must be 0 or a power of two in [4096, 262144]");
        ctx->opt_unit_bytes = value;
    } else return set_error(DISSEM_EINVAL, "unknown option '%s'", key);
    return DISSEM_OK;
}

int dissem_get_option(dissem_ctx *ctx, const char *key, int64_t *value_out) {
    if (!ctx || !key || !value_out) return set_error(DISSEM_EINVAL, "dissem_get_option: bad arguments");
    std::lock_guard<std::mutex> lk(ctx->mu);
    std::string k(key);
    if (k == "ctas_per_sm") *value_out = ctx->opt_ctas_per_sm;
    else if (k == "variant") *value_out = ctx->opt_variant;
    else if (k == "reg_cfg") *value_out = ctx->opt_reg_cfg;
    else if (k == "pull") *value_out = ctx->opt_pull;
    else if (k == "tma_cfg") *value_out = ctx->opt_tma_cfg;
    else if (k == "alloc_vmm") *value_out = ctx->opt_alloc_vmm;
    else if (k == "mcast") *value_out = ctx->opt_mcast;
    else if (k == "mcast_variant") *value_out = ctx->opt_mcast_variant;
    else if (k == "wait_src") *value_out = ctx->opt_wait_src;
    else if (k == "wait_timeouts") {
        // units that gave up waiting for a relay source, over all local GPUs (waits for their streams)
        int64_t total = 0;
        for (auto &ds : ctx->devs) {
            SchedCounters c;
            CUDA_TRY(cudaSetDevice(ds.dev));
            CUDA_TRY(cudaStreamSynchronize(ds.stream));
            CUDA_TRY(cudaMemcpy(&c, ds.counters, sizeof c, cudaMemcpyDeviceToHost));
            total += c.wait_timeouts;
        }
        *value_out = total;
    }
    else if (k == "mcast_groups") *value_out = ctx->mcast_groups_compiled;     // read-only: multicast descriptors compiled so far
    else if (k == "mcast_supported") {
        std::vector<int> devs;
        for (auto &d : ctx->devs) devs.push_back(d.dev);
        std::string why;
        *value_out = (devs.size() >= 2 && mcast_supported(devs, &why)) ? 1 : 0;
        if (!*value_out) ctx->mcast_note = why;
    }
    else if (k == "unit_bytes") *value_out = ctx->opt_unit_bytes;
    else if (k == "file_threads") *value_out = ctx->opt_file_threads;
    else return set_error(DISSEM_EINVAL, "unknown option '%s'", key);
    return DISSEM_OK;
}

}  // extern "C"
