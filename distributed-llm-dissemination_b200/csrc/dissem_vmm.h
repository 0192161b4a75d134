// Virtual-memory-management allocations and NVLS multicast objects (CUDA driver API, resolved at run time through
// cudaGetDriverEntryPoint so that libdissem.so still loads on a machine without libcuda).  Not part of the C ABI.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <string>
#include <vector>

namespace dissem {

// One physical allocation mapped at one virtual address, readable/writable from a set of local devices.
struct VmmBlock {
    unsigned long long handle = 0;   // CUmemGenericAllocationHandle
    unsigned long long va = 0;       // CUdeviceptr
    size_t bytes = 0;                // rounded to the allocation granularity
    int device = -1;
    bool valid() const { return va != 0; }
};

// A multicast object over the same-sized blocks of several devices: a store through `va` lands in every block.
struct McObject {
    unsigned long long handle = 0;
    unsigned long long va = 0;
    size_t bytes = 0;
    std::vector<int> devices;
    bool valid() const { return va != 0; }
};

// Empty string on success, else what failed.
// shareable: the allocation can be exported as a POSIX file descriptor (and bound to a shareable multicast object)
std::string vmm_alloc(int device, const std::vector<int> &access_devices, size_t bytes, VmmBlock *out, bool shareable = false);
// A file descriptor for the block, to be passed to another process (SCM_RIGHTS); the caller closes its copy.
std::string vmm_export_fd(const VmmBlock &b, int *fd_out);
// Map a block exported by another process; `bytes` as the exporter allocated it (rounded the same way).
std::string vmm_import_fd(int fd, int owner_device, const std::vector<int> &access_devices, size_t bytes, VmmBlock *out);
void vmm_free(VmmBlock *b);

// Does every device support multicast objects (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED)?
bool mcast_supported(const std::vector<int> &devices, std::string *why);
// blocks[i] lives on devices[i]; all have the same size.  The object is mapped for access from every device listed.
std::string mcast_create(const std::vector<int> &devices, const std::vector<const VmmBlock *> &blocks, McObject *out);
void mcast_destroy(McObject *m);

// One process per GPU.  The source's process creates the object for `n_devices` members, adds its own device and exports a
// file descriptor; every other member's process joins (imports the fd, adds its device).  After ALL members joined, each
// process binds its own block; the source's process also maps the object.  `m->devices` lists only the LOCAL devices.
std::string mcast_create_shared(int n_devices, size_t bytes, int own_device, McObject *m, int *fd_out);
std::string mcast_join(int fd, size_t bytes, int own_device, McObject *m);
std::string mcast_bind_block(McObject *m, const VmmBlock &b);
std::string mcast_map(McObject *m, const std::vector<int> &access_devices);

}  // namespace dissem
