#include "dissem_vmm.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>

namespace dissem {

namespace {

struct DriverApi {
    bool ok = false;
    std::string err;
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType) = nullptr;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags) = nullptr;
    CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
};

template <class F>
bool resolve(const char *name, F *fn, std::string *err) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
        cudaGetLastError();
        *err = std::string("driver entry point ") + name + " unavailable";
        return false;
    }
    *fn = reinterpret_cast<F>(p);
    return true;
}

const DriverApi &driver() {
    static DriverApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        std::string &e = api.err;
        api.ok = resolve("cuMemGetAllocationGranularity", &api.MemGetAllocationGranularity, &e) &&
                 resolve("cuMemCreate", &api.MemCreate, &e) && resolve("cuMemRelease", &api.MemRelease, &e) &&
                 resolve("cuMemAddressReserve", &api.MemAddressReserve, &e) &&
                 resolve("cuMemAddressFree", &api.MemAddressFree, &e) && resolve("cuMemMap", &api.MemMap, &e) &&
                 resolve("cuMemUnmap", &api.MemUnmap, &e) && resolve("cuMemSetAccess", &api.MemSetAccess, &e) &&
                 resolve("cuMemExportToShareableHandle", &api.MemExportToShareableHandle, &e) &&
                 resolve("cuMemImportFromShareableHandle", &api.MemImportFromShareableHandle, &e) &&
                 resolve("cuMulticastCreate", &api.MulticastCreate, &e) &&
                 resolve("cuMulticastAddDevice", &api.MulticastAddDevice, &e) &&
                 resolve("cuMulticastBindMem", &api.MulticastBindMem, &e) &&
                 resolve("cuMulticastUnbind", &api.MulticastUnbind, &e) &&
                 resolve("cuMulticastGetGranularity", &api.MulticastGetGranularity, &e) &&
                 resolve("cuDeviceGet", &api.DeviceGet, &e) &&
                 resolve("cuDeviceGetAttribute", &api.DeviceGetAttribute, &e) &&
                 resolve("cuGetErrorString", &api.GetErrorString, &e);
    });
    return api;
}

std::string cu_err(const char *what, CUresult r) {
    const char *s = nullptr;
    if (driver().GetErrorString) driver().GetErrorString(r, &s);
    return std::string(what) + ": " + (s ? s : "unknown driver error") + " (" + std::to_string((int)r) + ")";
}

CUmemAllocationProp alloc_prop(int device, bool shareable = false) {
    CUmemAllocationProp p;
    memset(&p, 0, sizeof p);
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = device;
    p.requestedHandleTypes = shareable ? CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR : CU_MEM_HANDLE_TYPE_NONE;
    return p;
}

size_t round_to_granularity(int device, bool shareable, size_t bytes, size_t *gran_out) {
    CUmemAllocationProp prop = alloc_prop(device, shareable);
    size_t gran = 0;
    if (driver().MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) gran = 0;
    if (gran < (2u << 20)) gran = 2u << 20;
    if (gran_out) *gran_out = gran;
    return (bytes + gran - 1) / gran * gran;
}

std::string set_access(CUdeviceptr va, size_t bytes, const std::vector<int> &devices) {
    std::vector<CUmemAccessDesc> acc(devices.size());
    for (size_t i = 0; i < devices.size(); ++i) {
        memset(&acc[i], 0, sizeof acc[i]);
        acc[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc[i].location.id = devices[i];
        acc[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    }
    CUresult r = driver().MemSetAccess(va, bytes, acc.data(), acc.size());
    return r == CUDA_SUCCESS ? std::string() : cu_err("cuMemSetAccess", r);
}

}  // namespace

std::string vmm_alloc(int device, const std::vector<int> &access_devices, size_t bytes, VmmBlock *out, bool shareable) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    cudaSetDevice(device);
    cudaFree(0);                                   // make sure the primary context exists
    CUmemAllocationProp prop = alloc_prop(device, shareable);
    size_t gran = 0;
    const size_t size = round_to_granularity(device, shareable, bytes, &gran);
    CUmemGenericAllocationHandle h = 0;
    CUresult r = d.MemCreate(&h, size, &prop, 0);
    if (r != CUDA_SUCCESS) return cu_err("cuMemCreate", r);
    CUdeviceptr va = 0;
    r = d.MemAddressReserve(&va, size, gran, 0, 0);
    if (r != CUDA_SUCCESS) { d.MemRelease(h); return cu_err("cuMemAddressReserve", r); }
    r = d.MemMap(va, size, 0, h, 0);
    if (r != CUDA_SUCCESS) { d.MemAddressFree(va, size); d.MemRelease(h); return cu_err("cuMemMap", r); }
    std::string e = set_access(va, size, access_devices);
    if (!e.empty()) { d.MemUnmap(va, size); d.MemAddressFree(va, size); d.MemRelease(h); return e; }
    out->handle = h;
    out->va = va;
    out->bytes = size;
    out->device = device;
    return std::string();
}

std::string vmm_export_fd(const VmmBlock &b, int *fd_out) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    int fd = -1;
    CUresult r = d.MemExportToShareableHandle(&fd, b.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) return cu_err("cuMemExportToShareableHandle", r);
    *fd_out = fd;
    return std::string();
}

std::string vmm_import_fd(int fd, int owner_device, const std::vector<int> &access_devices, size_t bytes, VmmBlock *out) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    if (access_devices.empty()) return "no local device to map the block on";
    cudaSetDevice(access_devices[0]);
    cudaFree(0);
    size_t gran = 0;
    const size_t size = round_to_granularity(access_devices[0], true, bytes, &gran);
    CUmemGenericAllocationHandle h = 0;
    CUresult r = d.MemImportFromShareableHandle(&h, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) return cu_err("cuMemImportFromShareableHandle", r);
    CUdeviceptr va = 0;
    r = d.MemAddressReserve(&va, size, gran, 0, 0);
    if (r != CUDA_SUCCESS) { d.MemRelease(h); return cu_err("cuMemAddressReserve", r); }
    r = d.MemMap(va, size, 0, h, 0);
    if (r != CUDA_SUCCESS) { d.MemAddressFree(va, size); d.MemRelease(h); return cu_err("cuMemMap(imported)", r); }
    std::string e = set_access(va, size, access_devices);
    if (!e.empty()) { d.MemUnmap(va, size); d.MemAddressFree(va, size); d.MemRelease(h); return e; }
    out->handle = h;
    out->va = va;
    out->bytes = size;
    out->device = owner_device;
    return std::string();
}

void vmm_free(VmmBlock *b) {
    const DriverApi &d = driver();
    if (!d.ok || !b->valid()) return;
    d.MemUnmap(b->va, b->bytes);
    d.MemAddressFree(b->va, b->bytes);
    d.MemRelease(b->handle);
    *b = VmmBlock();
}

bool mcast_supported(const std::vector<int> &devices, std::string *why) {
    const DriverApi &d = driver();
    if (!d.ok) { if (why) *why = d.err; return false; }
    for (int dev : devices) {
        CUdevice cd;
        int v = 0;
        if (d.DeviceGet(&cd, dev) != CUDA_SUCCESS ||
            d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS || !v) {
            if (why) *why = "device " + std::to_string(dev) + " does not support multicast objects";
            return false;
        }
    }
    return true;
}

std::string mcast_create(const std::vector<int> &devices, const std::vector<const VmmBlock *> &blocks, McObject *out) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    if (devices.size() < 2 || devices.size() != blocks.size()) return "multicast needs two or more devices, one block each";
    const size_t bytes = blocks[0]->bytes;
    for (auto *b : blocks)
        if (!b->valid() || b->bytes != bytes) return "multicast blocks must be VMM allocations of one size";
    CUmulticastObjectProp prop;
    memset(&prop, 0, sizeof prop);
    prop.numDevices = (unsigned)devices.size();
    prop.size = bytes;
    prop.handleTypes = CU_MEM_HANDLE_TYPE_NONE;
    size_t gran = 0, gran_min = 0;
    CUresult r = d.MulticastGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
    if (r == CUDA_SUCCESS) r = d.MulticastGetGranularity(&gran_min, &prop, CU_MULTICAST_GRANULARITY_MINIMUM);
    if (r != CUDA_SUCCESS) return cu_err("cuMulticastGetGranularity", r);
    if (gran_min && bytes % gran_min) return "block size " + std::to_string(bytes) + " is not a multiple of the multicast granularity " + std::to_string(gran_min);
    CUmemGenericAllocationHandle mc = 0;
    r = d.MulticastCreate(&mc, &prop);
    if (r != CUDA_SUCCESS) return cu_err("cuMulticastCreate", r);
    for (int dev : devices) {
        CUdevice cd;
        r = d.DeviceGet(&cd, dev);
        if (r == CUDA_SUCCESS) r = d.MulticastAddDevice(mc, cd);
        if (r != CUDA_SUCCESS) { d.MemRelease(mc); return cu_err("cuMulticastAddDevice", r); }
    }
    size_t bound = 0;
    for (; bound < blocks.size(); ++bound) {
        r = d.MulticastBindMem(mc, 0, blocks[bound]->handle, 0, bytes, 0);
        if (r != CUDA_SUCCESS) break;
    }
    std::string err;
    CUdeviceptr va = 0;
    if (r != CUDA_SUCCESS) err = cu_err("cuMulticastBindMem", r);
    if (err.empty()) {
        r = d.MemAddressReserve(&va, bytes, gran ? gran : (2u << 20), 0, 0);
        if (r != CUDA_SUCCESS) err = cu_err("cuMemAddressReserve(multicast)", r);
    }
    if (err.empty()) {
        r = d.MemMap(va, bytes, 0, mc, 0);
        if (r != CUDA_SUCCESS) { err = cu_err("cuMemMap(multicast)", r); d.MemAddressFree(va, bytes); va = 0; }
    }
    if (err.empty()) {
        err = set_access(va, bytes, devices);
        if (!err.empty()) { d.MemUnmap(va, bytes); d.MemAddressFree(va, bytes); va = 0; }
    }
    if (!err.empty()) {
        for (size_t i = 0; i < bound; ++i) {
            CUdevice cd;
            if (d.DeviceGet(&cd, devices[i]) == CUDA_SUCCESS) d.MulticastUnbind(mc, cd, 0, bytes);
        }
        d.MemRelease(mc);
        return err;
    }
    out->handle = mc;
    out->va = va;
    out->bytes = bytes;
    out->devices = devices;
    return std::string();
}

std::string mcast_create_shared(int n_devices, size_t bytes, int own_device, McObject *m, int *fd_out) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    cudaSetDevice(own_device);
    cudaFree(0);
    CUmulticastObjectProp prop;
    memset(&prop, 0, sizeof prop);
    prop.numDevices = (unsigned)n_devices;
    prop.size = bytes;
    prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t gran_min = 0;
    CUresult r = d.MulticastGetGranularity(&gran_min, &prop, CU_MULTICAST_GRANULARITY_MINIMUM);
    if (r != CUDA_SUCCESS) return cu_err("cuMulticastGetGranularity", r);
    if (gran_min && bytes % gran_min) return "block size is not a multiple of the multicast granularity";
    CUmemGenericAllocationHandle mc = 0;
    r = d.MulticastCreate(&mc, &prop);
    if (r != CUDA_SUCCESS) return cu_err("cuMulticastCreate(shareable)", r);
    CUdevice cd;
    r = d.DeviceGet(&cd, own_device);
    if (r == CUDA_SUCCESS) r = d.MulticastAddDevice(mc, cd);
    if (r != CUDA_SUCCESS) { d.MemRelease(mc); return cu_err("cuMulticastAddDevice", r); }
    int fd = -1;
    r = d.MemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) { d.MemRelease(mc); return cu_err("cuMemExportToShareableHandle(multicast)", r); }
    m->handle = mc;
    m->va = 0;
    m->bytes = bytes;
    m->devices.assign(1, own_device);
    *fd_out = fd;
    return std::string();
}

std::string mcast_join(int fd, size_t bytes, int own_device, McObject *m) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    cudaSetDevice(own_device);
    cudaFree(0);
    CUmemGenericAllocationHandle mc = 0;
    CUresult r = d.MemImportFromShareableHandle(&mc, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) return cu_err("cuMemImportFromShareableHandle(multicast)", r);
    CUdevice cd;
    r = d.DeviceGet(&cd, own_device);
    if (r == CUDA_SUCCESS) r = d.MulticastAddDevice(mc, cd);
    if (r != CUDA_SUCCESS) { d.MemRelease(mc); return cu_err("cuMulticastAddDevice(joined)", r); }
    m->handle = mc;
    m->va = 0;
    m->bytes = bytes;
    m->devices.assign(1, own_device);
    return std::string();
}

std::string mcast_bind_block(McObject *m, const VmmBlock &b) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    if (b.bytes != m->bytes) return "multicast block size mismatch";
    CUresult r = d.MulticastBindMem(m->handle, 0, b.handle, 0, m->bytes, 0);
    return r == CUDA_SUCCESS ? std::string() : cu_err("cuMulticastBindMem", r);
}

std::string mcast_map(McObject *m, const std::vector<int> &access_devices) {
    const DriverApi &d = driver();
    if (!d.ok) return d.err;
    CUdeviceptr va = 0;
    CUresult r = d.MemAddressReserve(&va, m->bytes, 2u << 20, 0, 0);
    if (r != CUDA_SUCCESS) return cu_err("cuMemAddressReserve(multicast)", r);
    r = d.MemMap(va, m->bytes, 0, m->handle, 0);
    if (r != CUDA_SUCCESS) { d.MemAddressFree(va, m->bytes); return cu_err("cuMemMap(multicast)", r); }
    std::string e = set_access(va, m->bytes, access_devices);
    if (!e.empty()) { d.MemUnmap(va, m->bytes); d.MemAddressFree(va, m->bytes); return e; }
    m->va = va;
    return std::string();
}

void mcast_destroy(McObject *m) {
    const DriverApi &d = driver();
    if (!d.ok || !m->handle) return;
    if (m->va) {
        d.MemUnmap(m->va, m->bytes);
        d.MemAddressFree(m->va, m->bytes);
    }
    for (int dev : m->devices) {
        CUdevice cd;
        if (d.DeviceGet(&cd, dev) == CUDA_SUCCESS) d.MulticastUnbind(m->handle, cd, 0, m->bytes);
    }
    d.MemRelease(m->handle);
    *m = McObject();
}

}  // namespace dissem
