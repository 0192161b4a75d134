// Internal interface between the host engine (dissem_engine.cu) and the sm_100a kernels
// (dissem_kernels.cu).  Not part of the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dissem {

constexpr int kChunkShift = 18;                      // DISSEM_CHUNK_BYTES = 256 KiB
constexpr long long kChunkBytes = 1ll << kChunkShift;

enum : int {
    kFlagMulticast = 1,   // dst / csum / received are NVLS multicast addresses (multimem.* only)
    kFlagZeroHead = 2,    // this descriptor owns the zeroing of the partially covered chunk its fragment starts in
    kFlagZeroTail = 4,    // ... and of the partially covered chunk it ends in
    // Work units are whole checksum chunks (unit size == chunk size): a unit that covers its chunk completely is the ONLY
    // contribution to that table entry, so its sums are stored, not added, and the entry needs no zeroing beforehand
    // (set by the schedule compiler; partially covered chunks at a fragment's ends still accumulate with atomics).
    kFlagStoreSums = 8,
};

// One transfer as the kernels see it: bytes [off, off+size) of a layer, from the source copy to the
// destination copy (distributor/transport.go:342: data[Offset : Offset+DataSize]).
// Work units of a schedule are U-aligned slices of LAYER space (U = 1 << unit_shift <= 256 KiB), so a
// unit never straddles a checksum chunk.
// Ticket order is STRIPED over the table: round r hands out slice r of every descriptor that still has one, so at
// any moment a source GPU is spreading its CTAs over all of its destinations (no destination is incast while another
// idles).  The host sorts the table by unit count ascending; descriptor j then stores first_unit = L_j, the number
// of tickets issued before round c_{j-1}, and prev_units = c_{j-1}:  L_j = sum_{i<j-1} c_i + (m-j+1)*c_{j-1}.
struct __align__(16) XferDesc {
    const uint8_t *src;              // source layer base (local HBM)
    uint8_t *dst;                    // destination layer base: local, peer-mapped or multicast; NULL = checksum only
    unsigned long long *csum;        // destination checksum table (2 x u64 per chunk); NULL = none
    unsigned long long *received;    // destination received-bytes counter; NULL = none
    long long off;                   // fragment start within the layer
    long long size;                  // fragment bytes (> 0)
    long long total;                 // full layer size
    long long first_unit;            // stripe level start: tickets issued before round prev_units (see locate_unit)
    int flags;
    int prev_units;                  // unit count of the previous (shorter) descriptor in the table, 0 for the first
    int pad[2];
    // Relay across processes (DISSEM_XFER_WAIT_SRC): the source copy is itself being received during this dissemination.
    // Before a unit of this transfer is read, the kernel waits until *wait_ptr (the SOURCE copy's received-bytes
    // counter, which its upstream sender releases after every unit) has reached wait_value = off + size.  NULL = no wait.
    const unsigned long long *wait_ptr;
    unsigned long long wait_value;
};
static_assert(sizeof(XferDesc) == 96, "XferDesc layout");

// Scratch the copy kernels need beyond their inputs and outputs: the work-unit ticket and the count of
// CTAs that have drained (the last one re-arms both for the next launch).
struct SchedCounters {
    unsigned long long ticket;
    unsigned int done_ctas;
    unsigned int wait_timeouts;      // units that gave up waiting for their source (never re-armed: read by the host)
};

struct LaunchCfg {
    int n_sm;
    int ctas_per_sm;   // 0 = use the occupancy calculator
    int variant;       // 0 register path, 1 TMA bulk-staged path
    int reg_cfg;       // register-path shape (threads x loads in flight): 0 256x4, 1 512x4 (default), 2 256x8, 3 512x8, 4 1024x4, 5 1024x2, 6 512x2
    int tma_cfg = 0;   // TMA ring shape (stage bytes x stages, store lag): 0 16Kx6/3 (default), 1 32Kx3/1, 2 8Kx12/6, 3 16Kx4/2, 4 8Kx6/3, 5 4Kx12/6
};

// Zero the checksum entries a launch is about to accumulate into (so that launching the same schedule again does
// not add onto the previous run): every chunk a descriptor covers completely, plus the partially covered chunk at its
// head / tail when the schedule compiler made this descriptor the chunk's owner (kFlagZeroHead / kFlagZeroTail: of the
// fragments that share a chunk, the one in the earliest round with the lowest offset).  The engine runs all zeroing
// of a round before any copy of that round (device-side event waits), so owners never wipe a peer's contribution.
cudaError_t launch_zero_csum(const XferDesc *descs, int n_descs, cudaStream_t stream);

// The hot kernel: copy + fused per-chunk checksum (+ completion counter) over a descriptor table.
cudaError_t launch_copy_csum(const XferDesc *descs, int n_descs, long long total_units, int unit_shift,
                             SchedCounters *counters, const LaunchCfg &cfg, cudaStream_t stream,
                             int *grid_out);

// Gather of several checksum tables into one contiguous device buffer (dissem_layers_checksums: one device-to-host copy
// per GPU instead of one per layer).  dst_word / words are in u64 units.
struct GatherItem {
    const unsigned long long *src;
    long long dst_word;
    long long words;
    long long pad;
};
cudaError_t launch_gather_tables(const GatherItem *items, int n_items, unsigned long long *dst, int n_sm, cudaStream_t stream);

// Force the module with these kernels to load on the current device (CUDA loads lazily: without this the first
// transfer of a run would pay the load inside its timed region).
cudaError_t warm_kernels();

// Synthetic contents (DISSEM_FILL_PATTERN).
cudaError_t launch_fill_pattern(uint8_t *dst, long long size, unsigned long long layer_id,
                                unsigned long long seed, int n_sm, cudaStream_t stream);

#ifdef __CUDACC__
// Wait (one thread) until the source copy has received the bytes this transfer is about to forward.  The sender releases
// the counter at system scope after its data (fence + atomic), so an acquire load here orders the reads that follow;
// the proxy fence extends that to bulk (TMA) loads.  The wait is bounded (~4 s of SM clocks): a sender that never comes
// must not hang the GPU; the unit then proceeds, the launch is marked in SchedCounters::wait_timeouts and the delivered
// bytes will fail verification.
__device__ __forceinline__ void wait_for_source(const unsigned long long *p, unsigned long long want, SchedCounters *counters) {
    const long long t0 = clock64();
    for (;;) {
        unsigned long long got;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(got) : "l"(p) : "memory");
        if (got >= want) break;
        if (clock64() - t0 > 8000000000ll) {
            atomicAdd(&counters->wait_timeouts, 1u);
            break;
        }
        __nanosleep(256);
    }
    asm volatile("fence.proxy.async;" ::: "memory");
}

// ticket -> (descriptor, slice index within it).  Largest j with L_j <= t; within level j the (m - j) longest
// descriptors take turns.
__device__ __forceinline__ int locate_unit(const XferDesc *__restrict__ descs, int n, long long t, long long *slice_idx) {
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(&descs[mid].first_unit) <= t) lo = mid; else hi = mid - 1;
    }
    const long long q = t - __ldg(&descs[lo].first_unit);
    const int active = n - lo;
    *slice_idx = (long long)__ldg(&descs[lo].prev_units) + q / active;
    return lo + (int)(q % active);
}
#endif

// Number of units a fragment occupies for a given unit size.
static inline long long units_of(long long off, long long size, int unit_shift) {
    if (size <= 0) return 0;
    return ((off + size - 1) >> unit_shift) - (off >> unit_shift) + 1;
}

}  // namespace dissem
