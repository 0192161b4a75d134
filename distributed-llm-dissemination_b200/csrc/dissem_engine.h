// Internal host-side types shared by the engine and the host mirror.  Not part of the C ABI.
#pragma once
#include <stdint.h>

#include "dissem_vmm.h"

namespace dissem {

int set_error(int code, const char *fmt, ...) __attribute__((format(printf, 2, 3)));
const char *last_error();

// One layer of one node's store (LayerSrc, distributor/node.go:200-211) as an HBM buffer:
// [ layer bytes | pad to 4 KiB | received counter (256 B slot) | checksum table: 16 B per chunk ]
struct LayerBuf {
    int64_t size = 0;
    uint8_t *base = nullptr;
    int device = -1;        // CUDA ordinal of the owning device; -1 for an imported (remote) layer
    bool imported = false;
    bool host = false;      // a caller-owned pinned host buffer attached as this node's copy (source only, no trailer)
    bool vmm = false;       // allocated through the VMM API (can be bound to an NVLS multicast object)
    VmmBlock blk;
    // cudaEvent_t (kept opaque here): recorded on the OWNER's stream after everything the owner enqueued that writes
    // this buffer (trailer zeroing at alloc / reset, fill, host write).  A kernel of ANOTHER GPU's stream that reads or
    // writes the buffer waits for it first; streams of different GPUs are not ordered otherwise.  NULL for imported
    // and attached-host layers (their owner synchronises before handing out the handle / the memory).
    void *ready = nullptr;
    // bit s set: a kernel on local device slot s has been launched that writes this buffer (push from that GPU, pull or
    // disk load by the owner).  A host read of the buffer waits for those streams (and the owner's), not for every GPU.
    unsigned long long writers = 0;

    int64_t n_chunks() const;
    int64_t trailer_off() const;
    int64_t trailer_bytes() const;
    int64_t alloc_bytes() const;
    unsigned long long *received() const;
    unsigned long long *csum() const;
};

}  // namespace dissem
