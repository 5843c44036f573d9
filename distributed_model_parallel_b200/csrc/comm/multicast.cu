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
    bool ok = false;
    Drv() {
        ok = load_drv("cuMulticastCreate", &MulticastCreate) && load_drv("cuMulticastAddDevice", &MulticastAddDevice) &&
             load_drv("cuMulticastBindMem", &MulticastBindMem) && load_drv("cuMulticastGetGranularity", &MulticastGetGranularity) &&
             load_drv("cuMulticastUnbind", &MulticastUnbind) && load_drv("cuMemCreate", &MemCreate) &&
             load_drv("cuMemRelease", &MemRelease) && load_drv("cuMemAddressReserve", &MemAddressReserve) &&
             load_drv("cuMemAddressFree", &MemAddressFree) && load_drv("cuMemMap", &MemMap) && load_drv("cuMemUnmap", &MemUnmap) &&
             load_drv("cuMemSetAccess", &MemSetAccess) && load_drv("cuDeviceGet", &DeviceGet) &&
             load_drv("cuDeviceGetAttribute", &DeviceGetAttribute);
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
