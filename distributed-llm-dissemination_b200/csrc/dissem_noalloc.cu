// dissem_checksum_noalloc: the per-chunk checksum pass for callers that must not allocate device memory.
//
// Who needs this.  dissem_layer_verify() borrows a small piece of device memory for each call (a
// descriptor, the work-unit counters and a fresh checksum table) and gives it back afterwards.  Some
// hosts cannot afford that:
//   * a serving process that has already handed every free byte of HBM to its own pool (weights + KV
//     cache): a cudaMalloc there fails, or succeeds only by fragmenting what the pool thinks it owns;
//   * a caller on a latency-critical path: cudaMalloc/cudaFree take driver-wide locks and cudaFree
//     waits for the device, so one verification would stall every other stream of the process;
//   * a caller that runs under a memory-accounting wrapper that forbids allocation after start-up.
// For them the same kernel runs here with its scratch in a fixed array that is part of the library
// image: nothing is allocated, nothing is freed.
//
// The price, which the caller must know:
//   * the array costs kNoallocScratchBytes (about 129 KiB) of HBM on every GPU the process touches with
//     this library, from the moment the library's code is loaded on that GPU until the process exits;
//   * a layer may have at most DISSEM_NOALLOC_MAX_CHUNKS chunks (2 GiB); larger sizes are refused;
//   * there is ONE array per GPU and every call uses it.  It still holds whatever the previous call left,
//     so each call first clears the part it is going to accumulate into.  Two calls whose GPU work can
//     overlap (different streams, or different host threads) would trample each other's scratch: issue
//     them one after the other.
//   * the first call on each GPU looks up the array's device address.  Make that first call during
//     start-up, before any stream capture is begun on that GPU.
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>

#include "../../include/dissem.h"
#include "dissem_engine.h"
#include "dissem_kernels.cuh"

namespace {

constexpr long long kMaxChunks = DISSEM_NOALLOC_MAX_CHUNKS;
constexpr size_t kCountersOff = 0;       // SchedCounters, must be zero when the kernel starts
constexpr size_t kDescOff = 256;         // one XferDesc
constexpr size_t kTableOff = 512;        // checksum table, 16 bytes per chunk, accumulated into -> must start at zero
constexpr size_t kNoallocScratchBytes = kTableOff + (size_t)kMaxChunks * 16;

// The scratch itself.  In host code this NAME is only a handle for the runtime, not a device address:
// the address the kernel needs comes from cudaGetSymbolAddress (see device_scratch below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocScratchBytes];

constexpr int kMaxDevices = 64;
struct PerDevice {
    std::mutex mu;                      // one call at a time per GPU (each GPU has its own copy of the array)
    bool ready = false;
    unsigned char *scratch = nullptr;   // device address of g_noalloc_scratch on this GPU
    int n_sm = 0;
};
PerDevice g_dev[kMaxDevices];

// Once per GPU, with that GPU current and its PerDevice::mu held: resolve the array's device address and read the SM count.
int device_scratch(int device, PerDevice **out) {
    PerDevice &pd = g_dev[device];
    if (!pd.ready) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return dissem::set_error(DISSEM_ECUDA, "cudaGetSymbolAddress(no-allocation scratch) on device %d: %s", device,
                                     cudaGetErrorString(e));
        int n_sm = 0;
        e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || n_sm <= 0)
            return dissem::set_error(DISSEM_ECUDA, "cannot read the SM count of device %d: %s", device, cudaGetErrorString(e));
        pd.scratch = static_cast<unsigned char *>(ptr);
        pd.n_sm = n_sm;
        pd.ready = true;
    }
    *out = &pd;
    return DISSEM_OK;
}

}  // namespace

// Device code may name the array directly; this (never launched) kernel does, which also keeps the array in the
// device image whatever the compiler thinks of host-only references.
__global__ void dissem_noalloc_scratch_address(unsigned long long *where) {
    if (where) *where = reinterpret_cast<unsigned long long>(&g_noalloc_scratch[0]);
}

extern "C" int dissem_checksum_noalloc(int device, const void *data, int64_t size, uint64_t *out, int64_t n_chunks,
                                       void *stream) {
    using namespace dissem;
    if (size < 0 || n_chunks < 0 || (size > 0 && !data) || (n_chunks > 0 && !out))
        return set_error(DISSEM_EINVAL, "dissem_checksum_noalloc: bad arguments");
    const int64_t want_chunks = (size + kChunkBytes - 1) / kChunkBytes;
    if (n_chunks != want_chunks)
        return set_error(DISSEM_EINVAL, "%lld bytes are %lld chunks of %lld bytes, caller passed %lld", (long long)size,
                         (long long)want_chunks, (long long)kChunkBytes, (long long)n_chunks);
    if (want_chunks > kMaxChunks)
        return set_error(DISSEM_ERANGE,
                         "dissem_checksum_noalloc: %lld bytes need %lld chunk slots but the built-in scratch holds %lld "
                         "(at most %lld bytes per call); use dissem_layer_verify for larger layers",
                         (long long)size, (long long)want_chunks, kMaxChunks, kMaxChunks * kChunkBytes);
    if (reinterpret_cast<uintptr_t>(data) % 16 != 0)
        return set_error(DISSEM_EINVAL, "dissem_checksum_noalloc: data must be 16-byte aligned");
    if (size == 0) return DISSEM_OK;

    if (device < 0 || device >= kMaxDevices) return set_error(DISSEM_EINVAL, "device %d out of range", device);
    // Host-side serialisation per GPU: calls on different GPUs use different arrays and run side by side; calls on one
    // GPU queue here.  GPU-side overlap through other streams of the caller's remains the caller's to avoid.
    std::lock_guard<std::mutex> lk(g_dev[device].mu);
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return set_error(DISSEM_ENODEV, "no usable CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    }
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return set_error(DISSEM_ECUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    int rc = DISSEM_OK;
    PerDevice *pd = nullptr;
    rc = device_scratch(device, &pd);
    if (rc == DISSEM_OK) {
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        unsigned char *scratch = pd->scratch;
        XferDesc d;
        memset(&d, 0, sizeof d);
        d.src = static_cast<const uint8_t *>(data);
        d.dst = nullptr;                                  // checksum only
        d.csum = reinterpret_cast<unsigned long long *>(scratch + kTableOff);
        d.received = nullptr;
        d.off = 0;
        d.size = size;
        d.total = size;
        d.first_unit = 0;
        const size_t table_bytes = (size_t)n_chunks * 16;
        // the array still holds the previous call's counters, descriptor and sums: clear what this call uses
        e = cudaMemsetAsync(scratch, 0, kTableOff + table_bytes, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(scratch + kDescOff, &d, sizeof d, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) {
            LaunchCfg cfg{pd->n_sm, 0, 0, 1};
            e = launch_copy_csum(reinterpret_cast<const XferDesc *>(scratch + kDescOff), 1,
                                 units_of(0, size, kChunkShift), kChunkShift,
                                 reinterpret_cast<SchedCounters *>(scratch + kCountersOff), cfg, s, nullptr);
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(out, scratch + kTableOff, table_bytes, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) rc = set_error(DISSEM_ECUDA, "dissem_checksum_noalloc: %s", cudaGetErrorString(e));
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    return rc;
}
