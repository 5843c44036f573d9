// NVLS multicast objects for the single-process DataParallel broadcast (north-star path 2: "broadcast_coalesced of
// parameters and buffers to replicas, done as a multicast store").
//
// One multicast object spans `ndev` GPUs of this process.  Every GPU contributes a physical allocation that is bound to the
// object and also mapped at an ordinary (unicast) address, so replicas read their copy with plain loads; the object itself is
// mapped at one multicast address, and a `multimem.st` to that address is replicated by the NVSwitch into every bound
// allocation.  The broadcast kernel (collectives.cu, `mc_dst` argument of dmp_broadcast) then writes the coalesced buffer
// ONCE instead of once per replica.  Everything goes through the driver API (resolved at run time with
// cudaGetDriverEntryPoint, so the library does not link libcuda).  If the platform has no multicast support the create call
// fails cleanly and callers keep using peer stores.
//
// No counterpart in the reference: its DataParallel path is torch._C._broadcast_coalesced (torch/nn/parallel/comm.py:50-67).
#include "../common.cuh"
#include <cuda.h>
#include <vector>

namespace {

template <typename F>
bool load_drv(const char* name, F* out) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || p == nullptr) return false;
    *out = (F)p;
    return true;
}

struct Drv {
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
    CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
    bool ok = false;
    Drv() {
        ok = load_drv("cuMulticastCreate", &MulticastCreate) && load_drv("cuMulticastAddDevice", &MulticastAddDevice) &&
             load_drv("cuMulticastBindMem", &MulticastBindMem) && load_drv("cuMulticastGetGranularity", &MulticastGetGranularity) &&
             load_drv("cuMulticastUnbind", &MulticastUnbind) && load_drv("cuMemCreate", &MemCreate) &&
             load_drv("cuMemRelease", &MemRelease) && load_drv("cuMemAddressReserve", &MemAddressReserve) &&
             load_drv("cuMemAddressFree", &MemAddressFree) && load_drv("cuMemMap", &MemMap) && load_drv("cuMemUnmap", &MemUnmap) &&
             load_drv("cuMemSetAccess", &MemSetAccess) && load_drv("cuDeviceGet", &DeviceGet) &&
             load_drv("cuDeviceGetAttribute", &DeviceGetAttribute) &&
             load_drv("cuMemExportToShareableHandle", &MemExportToShareableHandle) &&
             load_drv("cuMemImportFromShareableHandle", &MemImportFromShareableHandle);
    }
};
Drv& drv() {
    static Drv d;
    return d;
}

struct McObject {
    int ndev = 0;
    size_t size = 0;
    std::vector<int> devs;
    std::vector<CUdevice> cudevs;
    CUmemGenericAllocationHandle mc = 0;
    std::vector<CUmemGenericAllocationHandle> mem;
    std::vector<CUdeviceptr> uc;  // unicast address of each device's allocation
    CUdeviceptr mc_va = 0;
    bool bound = false;
};

void destroy(McObject* o) {
    Drv& d = drv();
    if (o->mc_va) {
        d.MemUnmap(o->mc_va, o->size);
        d.MemAddressFree(o->mc_va, o->size);
    }
    for (size_t i = 0; i < o->uc.size(); ++i)
        if (o->uc[i]) {
            d.MemUnmap(o->uc[i], o->size);
            d.MemAddressFree(o->uc[i], o->size);
        }
    if (o->bound)
        for (size_t i = 0; i < o->cudevs.size(); ++i) d.MulticastUnbind(o->mc, o->cudevs[i], 0, o->size);
    for (auto h : o->mem)
        if (h) d.MemRelease(h);
    if (o->mc) d.MemRelease(o->mc);
    delete o;
}

#define MC_DRV(call, what)                                                                       \
    do {                                                                                         \
        CUresult _r = (call);                                                                    \
        if (_r != CUDA_SUCCESS) {                                                                \
            dmp_set_error("multicast: %s failed with CUresult %d", what, (int)_r);               \
            destroy(o);                                                                          \
            cudaSetDevice(prev);                                                                 \
            return -(int)_r - 1000;                                                              \
        }                                                                                        \
    } while (0)

}  // namespace

// 1 if every listed device reports CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, else 0.
DMP_API int dmp_mc_supported(const int* devs, int ndev) {
    Drv& d = drv();
    if (!d.ok) return 0;
    for (int i = 0; i < ndev; ++i) {
        CUdevice cd;
        int v = 0;
        cudaSetDevice(devs[i]);
        cudaFree(0);  // make sure the primary context exists
        if (d.DeviceGet(&cd, devs[i]) != CUDA_SUCCESS) return 0;
        if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS || !v) return 0;
    }
    return 1;
}

// Creates a multicast object of at least `bytes` over the devices, with one bound allocation per device.
// out_handle: opaque object; out_unicast[ndev]: per-device addresses of the bound memory; out_mc: the multicast address;
// out_size: the (granularity-rounded) size.  The caller's current device is restored.
DMP_API int dmp_mc_create(const int* devs, int ndev, long long bytes, void** out_handle, void** out_unicast, void** out_mc,
                          long long* out_size) {
    Drv& d = drv();
    DMP_REQUIRE(d.ok, "multicast: driver entry points unavailable");
    DMP_REQUIRE(ndev >= 2 && bytes > 0, "multicast: need at least two devices and a positive size");
    int prev = 0;
    cudaGetDevice(&prev);
    McObject* o = new McObject();
    o->ndev = ndev;
    o->devs.assign(devs, devs + ndev);
    o->cudevs.resize(ndev);
    o->mem.assign(ndev, 0);
    o->uc.assign(ndev, 0);
    for (int i = 0; i < ndev; ++i) {
        cudaSetDevice(devs[i]);
        cudaFree(0);
        MC_DRV(d.DeviceGet(&o->cudevs[i], devs[i]), "cuDeviceGet");
    }
    CUmulticastObjectProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.numDevices = (unsigned)ndev;
    prop.size = (size_t)bytes;
    prop.handleTypes = 0;
    size_t gran = 0;
    MC_DRV(d.MulticastGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity");
    o->size = ((size_t)bytes + gran - 1) / gran * gran;
    prop.size = o->size;
    MC_DRV(d.MulticastCreate(&o->mc, &prop), "cuMulticastCreate");
    for (int i = 0; i < ndev; ++i) MC_DRV(d.MulticastAddDevice(o->mc, o->cudevs[i]), "cuMulticastAddDevice");
    std::vector<CUmemAccessDesc> access(ndev);
    for (int i = 0; i < ndev; ++i) {
        access[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        access[i].location.id = devs[i];
        access[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    }
    for (int i = 0; i < ndev; ++i) {
        cudaSetDevice(devs[i]);
        CUmemAllocationProp ap;
        memset(&ap, 0, sizeof(ap));
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = devs[i];
        MC_DRV(d.MemCreate(&o->mem[i], o->size, &ap, 0), "cuMemCreate");
        MC_DRV(d.MulticastBindMem(o->mc, 0, o->mem[i], 0, o->size, 0), "cuMulticastBindMem");
        MC_DRV(d.MemAddressReserve(&o->uc[i], o->size, gran, 0, 0), "cuMemAddressReserve");
        MC_DRV(d.MemMap(o->uc[i], o->size, 0, o->mem[i], 0), "cuMemMap");
        MC_DRV(d.MemSetAccess(o->uc[i], o->size, access.data(), (size_t)ndev), "cuMemSetAccess");
    }
    o->bound = true;
    MC_DRV(d.MemAddressReserve(&o->mc_va, o->size, gran, 0, 0), "cuMemAddressReserve (multicast)");
    MC_DRV(d.MemMap(o->mc_va, o->size, 0, o->mc, 0), "cuMemMap (multicast)");
    MC_DRV(d.MemSetAccess(o->mc_va, o->size, access.data(), (size_t)ndev), "cuMemSetAccess (multicast)");
    for (int i = 0; i < ndev; ++i) {
        out_unicast[i] = (void*)o->uc[i];
        cudaSetDevice(devs[i]);
        cudaMemset((void*)o->uc[i], 0, o->size);
        cudaDeviceSynchronize();
    }
    *out_mc = (void*)o->mc_va;
    *out_size = (long long)o->size;
    *out_handle = o;
    cudaSetDevice(prev);
    return 0;
}

DMP_API int dmp_mc_destroy(void* handle) {
    if (handle) destroy((McObject*)handle);
    return 0;
}

// =====================================================================================================================
// Cross-process multicast (one process per GPU: DistributedDataParallel).  Rank 0 creates the multicast object with a POSIX
// file-descriptor handle type and exports it; the descriptor travels to the other ranks over a Unix socket (SCM_RIGHTS, done
// in Python: parallel/symm.py) and is imported there.  Every rank then adds its device, and -- after a barrier, because all
// devices must have been added before the first bind -- binds a local physical allocation, maps it at a unicast address and
// maps the object at the multicast address.  multimem.ld_reduce on that address returns the switch-side sum of all ranks'
// copies, multimem.st replicates a value into all of them: the NVLS all-reduce kernel in collectives.cu uses exactly that.
// No counterpart in the reference (it has no collective of its own; SURVEY.md section 5.8 names this mechanism).
// =====================================================================================================================
namespace {

struct XpObject {
    int dev = 0;
    CUdevice cudev = 0;
    size_t size = 0, gran = 0;
    CUmemGenericAllocationHandle mc = 0, mem = 0;
    CUdeviceptr uc = 0, mc_va = 0;
    bool bound = false;
};

void xp_destroy(XpObject* o) {
    Drv& d = drv();
    if (o->mc_va) {
        d.MemUnmap(o->mc_va, o->size);
        d.MemAddressFree(o->mc_va, o->size);
    }
    if (o->uc) {
        d.MemUnmap(o->uc, o->size);
        d.MemAddressFree(o->uc, o->size);
    }
    if (o->bound) d.MulticastUnbind(o->mc, o->cudev, 0, o->size);
    if (o->mem) d.MemRelease(o->mem);
    if (o->mc) d.MemRelease(o->mc);
    delete o;
}

#define XP_DRV(call, what)                                                                        \
    do {                                                                                          \
        CUresult _r = (call);                                                                     \
        if (_r != CUDA_SUCCESS) {                                                                 \
            dmp_set_error("multicast (cross-process): %s failed with CUresult %d", what, (int)_r); \
            return -(int)_r - 1000;                                                               \
        }                                                                                         \
    } while (0)

int xp_prop(int ndev_total, long long bytes, CUmulticastObjectProp* prop, size_t* gran) {
    Drv& d = drv();
    memset(prop, 0, sizeof(*prop));
    prop->numDevices = (unsigned)ndev_total;
    prop->size = (size_t)bytes;
    prop->handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    XP_DRV(d.MulticastGetGranularity(gran, prop, CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity");
    prop->size = ((size_t)bytes + *gran - 1) / *gran * *gran;
    return 0;
}

int xp_new(int dev, XpObject** out) {
    Drv& d = drv();
    DMP_REQUIRE(d.ok, "multicast: driver entry points unavailable");
    DMP_CHECK_CUDA(cudaSetDevice(dev));
    cudaFree(0);
    XpObject* o = new XpObject();
    o->dev = dev;
    if (d.DeviceGet(&o->cudev, dev) != CUDA_SUCCESS) {
        delete o;
        DMP_REQUIRE(false, "multicast: cuDeviceGet(%d) failed", dev);
    }
    *out = o;
    return 0;
}

}  // namespace

// Rank 0: creates the exportable multicast object for `ndev_total` devices; *out_fd is the descriptor to hand to the peers.
DMP_API int dmp_mc_xp_create(int dev, int ndev_total, long long bytes, void** out_handle, int* out_fd, long long* out_size) {
    DMP_REQUIRE(ndev_total >= 2 && bytes > 0 && out_handle && out_fd && out_size, "dmp_mc_xp_create: bad arguments");
    XpObject* o = nullptr;
    if (int rc = xp_new(dev, &o)) return rc;
    Drv& d = drv();
    CUmulticastObjectProp prop;
    if (int rc = xp_prop(ndev_total, bytes, &prop, &o->gran)) { xp_destroy(o); return rc; }
    o->size = prop.size;
    CUresult r = d.MulticastCreate(&o->mc, &prop);
    if (r != CUDA_SUCCESS) { xp_destroy(o); dmp_set_error("multicast: cuMulticastCreate failed with CUresult %d", (int)r); return -(int)r - 1000; }
    int fd = -1;
    r = d.MemExportToShareableHandle(&fd, o->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) { xp_destroy(o); dmp_set_error("multicast: cuMemExportToShareableHandle failed with CUresult %d", (int)r); return -(int)r - 1000; }
    *out_fd = fd;
    *out_size = (long long)o->size;
    *out_handle = o;
    return 0;
}

// Other ranks: import the object from the descriptor received from rank 0 (same ndev_total / bytes as the creator).
DMP_API int dmp_mc_xp_import(int dev, int fd, int ndev_total, long long bytes, void** out_handle, long long* out_size) {
    DMP_REQUIRE(fd >= 0 && ndev_total >= 2 && bytes > 0 && out_handle && out_size, "dmp_mc_xp_import: bad arguments");
    XpObject* o = nullptr;
    if (int rc = xp_new(dev, &o)) return rc;
    Drv& d = drv();
    CUmulticastObjectProp prop;
    if (int rc = xp_prop(ndev_total, bytes, &prop, &o->gran)) { xp_destroy(o); return rc; }
    o->size = prop.size;
    CUresult r = d.MemImportFromShareableHandle(&o->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) { xp_destroy(o); dmp_set_error("multicast: cuMemImportFromShareableHandle failed with CUresult %d", (int)r); return -(int)r - 1000; }
    *out_size = (long long)o->size;
    *out_handle = o;
    return 0;
}

// Every rank, once it holds the object.  ALL ranks must have returned from this before any rank calls dmp_mc_xp_bind_map.
DMP_API int dmp_mc_xp_add_device(void* handle) {
    XpObject* o = (XpObject*)handle;
    DMP_REQUIRE(o != nullptr, "dmp_mc_xp_add_device: null handle");
    DMP_CHECK_CUDA(cudaSetDevice(o->dev));
    XP_DRV(drv().MulticastAddDevice(o->mc, o->cudev), "cuMulticastAddDevice");
    return 0;
}

// Every rank: bind a zeroed local allocation and map both views.  *out_unicast is this rank's copy, *out_mc the multicast address.
DMP_API int dmp_mc_xp_bind_map(void* handle, void** out_unicast, void** out_mc) {
    XpObject* o = (XpObject*)handle;
    DMP_REQUIRE(o && out_unicast && out_mc, "dmp_mc_xp_bind_map: bad arguments");
    Drv& d = drv();
    DMP_CHECK_CUDA(cudaSetDevice(o->dev));
    CUmemAllocationProp ap;
    memset(&ap, 0, sizeof(ap));
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = o->dev;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    XP_DRV(d.MemCreate(&o->mem, o->size, &ap, 0), "cuMemCreate");
    XP_DRV(d.MulticastBindMem(o->mc, 0, o->mem, 0, o->size, 0), "cuMulticastBindMem");
    o->bound = true;
    CUmemAccessDesc access;
    access.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    access.location.id = o->dev;
    access.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    XP_DRV(d.MemAddressReserve(&o->uc, o->size, o->gran, 0, 0), "cuMemAddressReserve");
    XP_DRV(d.MemMap(o->uc, o->size, 0, o->mem, 0), "cuMemMap");
    XP_DRV(d.MemSetAccess(o->uc, o->size, &access, 1), "cuMemSetAccess");
    XP_DRV(d.MemAddressReserve(&o->mc_va, o->size, o->gran, 0, 0), "cuMemAddressReserve (multicast)");
    XP_DRV(d.MemMap(o->mc_va, o->size, 0, o->mc, 0), "cuMemMap (multicast)");
    XP_DRV(d.MemSetAccess(o->mc_va, o->size, &access, 1), "cuMemSetAccess (multicast)");
    DMP_CHECK_CUDA(cudaMemset((void*)o->uc, 0, o->size));
    DMP_CHECK_CUDA(cudaDeviceSynchronize());
    *out_unicast = (void*)o->uc;
    *out_mc = (void*)o->mc_va;
    return 0;
}

DMP_API int dmp_mc_xp_destroy(void* handle) {
    if (handle) xp_destroy((XpObject*)handle);
    return 0;
}
