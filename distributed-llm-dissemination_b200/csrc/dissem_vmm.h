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
std::string vmm_alloc(int device, const std::vector<int> &access_devices, size_t bytes, VmmBlock *out);
void vmm_free(VmmBlock *b);

// Does every device support multicast objects (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED)?
bool mcast_supported(const std::vector<int> &devices, std::string *why);
// blocks[i] lives on devices[i]; all have the same size.  The object is mapped for access from every device listed.
std::string mcast_create(const std::vector<int> &devices, const std::vector<const VmmBlock *> &blocks, McObject *out);
void mcast_destroy(McObject *m);

}  // namespace dissem
