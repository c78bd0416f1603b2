// rcull_mgpu.cpp -- single-process multi-GPU bootstrap ABOVE the per-rank C ABI (include/rcull.h): one context and
// one range-shard merge per GPU, the windows of all ranks mapped into each other, and -- when the driver offers it -- an
// NVLS multicast object bound over them, without any Python, torch or NCCL in the set-up.  The reference's caller is
// one process that steps one bevy schedule per frame (src/main.rs:907-910); this is what lets such a host stand up N
// ranks.  Multi-PROCESS hosts exchange rcull_merge_ipc_handle blobs over their own transport instead.
//
// Windows are cuMemCreate allocations mapped with cuMemMap and opened to every device with cuMemSetAccess (peer
// stores over NVLink); the multicast object (cuMulticastCreate / cuMulticastBindMem) is mapped once more as the address a
// multimem.st goes to.  The driver entry points are resolved at run time (cudaGetDriverEntryPoint), like
// cuTensorMapEncodeTiled in rcull.cu: librcull does not link libcuda.  Any failure on the multicast path falls back to
// plain peer pointers (cudaMalloc + cudaDeviceEnablePeerAccess) and says so in rcull_mgpu_transport().
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "rcull.h"

namespace {

struct Driver {
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
    bool ok = false;
};

template <typename F>
bool resolve(const char* name, F* out) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) return false;
    *out = reinterpret_cast<F>(fn);
    return true;
}

const Driver& driver() {
    static Driver d = [] {
        Driver x;
        x.ok = resolve("cuMemGetAllocationGranularity", &x.MemGetAllocationGranularity) && resolve("cuMemCreate", &x.MemCreate) &&
               resolve("cuMemRelease", &x.MemRelease) && resolve("cuMemAddressReserve", &x.MemAddressReserve) &&
               resolve("cuMemAddressFree", &x.MemAddressFree) && resolve("cuMemMap", &x.MemMap) && resolve("cuMemUnmap", &x.MemUnmap) &&
               resolve("cuMemSetAccess", &x.MemSetAccess) && resolve("cuMulticastCreate", &x.MulticastCreate) &&
               resolve("cuMulticastAddDevice", &x.MulticastAddDevice) && resolve("cuMulticastBindMem", &x.MulticastBindMem) &&
               resolve("cuMulticastGetGranularity", &x.MulticastGetGranularity) && resolve("cuDeviceGetAttribute", &x.DeviceGetAttribute);
        return x;
    }();
    return d;
}

thread_local char g_mgpu_err[512] = "";

}  // namespace

struct rcull_mgpu {
    uint32_t n = 0;
    std::vector<int> devices;
    std::vector<rcull_ctx*> ctx;
    std::vector<rcull_merge*> merge;
    std::vector<void*> stream;
    // virtual-memory windows (multicast path); empty on the peer-pointer path, where the merges own their windows
    size_t vm_size = 0;
    std::vector<CUmemGenericAllocationHandle> mem;
    std::vector<CUdeviceptr> va;
    CUmemGenericAllocationHandle mc = 0;
    CUdeviceptr mc_va = 0;
    bool mc_mapped = false;
    std::string transport;
};

namespace {

int mfail(const char* what) {
    snprintf(g_mgpu_err, sizeof g_mgpu_err, "%s: %s", what, rcull_last_error());
    return RCULL_ERR_CUDA;
}

void vm_teardown(rcull_mgpu* g) {
    const Driver& d = driver();
    if (!d.ok) return;
    if (g->mc_mapped) {
        d.MemUnmap(g->mc_va, g->vm_size);
        g->mc_mapped = false;
    }
    if (g->mc_va) d.MemAddressFree(g->mc_va, g->vm_size);
    g->mc_va = 0;
    for (size_t r = 0; r < g->va.size(); ++r)
        if (g->va[r]) {
            d.MemUnmap(g->va[r], g->vm_size);
            d.MemAddressFree(g->va[r], g->vm_size);
        }
    g->va.clear();
    for (CUmemGenericAllocationHandle h : g->mem)
        if (h) d.MemRelease(h);
    g->mem.clear();
    if (g->mc) d.MemRelease(g->mc);
    g->mc = 0;
}

// Windows as VMM allocations + one multicast object over them.  false: nothing usable was set up (and nothing is left behind).
bool vm_windows(rcull_mgpu* g, uint64_t window_bytes, std::string* why) {
    const Driver& d = driver();
    if (!d.ok) return *why = "driver entry points for virtual memory / multicast not available", false;
    for (int dev : g->devices) {
        int sup = 0;
        if (d.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS || !sup)
            return *why = "device " + std::to_string(dev) + " does not support multicast objects", false;
    }
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof mp);
    mp.numDevices = g->n;
    mp.handleTypes = 0;
    mp.size = window_bytes;
    size_t gran = 0;
    if (d.MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0)
        return *why = "cuMulticastGetGranularity failed", false;
    for (int dev : g->devices) {
        CUmemAllocationProp ap;
        memset(&ap, 0, sizeof ap);
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = dev;
        size_t g2 = 0;
        if (d.MemGetAllocationGranularity(&g2, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && g2 > gran) gran = g2;
    }
    g->vm_size = (window_bytes + gran - 1) / gran * gran;
    mp.size = g->vm_size;
    if (d.MulticastCreate(&g->mc, &mp) != CUDA_SUCCESS) return g->mc = 0, *why = "cuMulticastCreate failed", false;
    for (int dev : g->devices)
        if (d.MulticastAddDevice(g->mc, dev) != CUDA_SUCCESS) return vm_teardown(g), *why = "cuMulticastAddDevice failed", false;
    std::vector<CUmemAccessDesc> access(g->n);
    for (uint32_t r = 0; r < g->n; ++r) {
        access[r].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        access[r].location.id = g->devices[r];
        access[r].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    }
    g->mem.assign(g->n, 0);
    g->va.assign(g->n, 0);
    for (uint32_t r = 0; r < g->n; ++r) {
        cudaSetDevice(g->devices[r]);
        CUmemAllocationProp ap;
        memset(&ap, 0, sizeof ap);
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = g->devices[r];
        if (d.MemCreate(&g->mem[r], g->vm_size, &ap, 0) != CUDA_SUCCESS) return g->mem[r] = 0, vm_teardown(g), *why = "cuMemCreate failed", false;
        if (d.MemAddressReserve(&g->va[r], g->vm_size, gran, 0, 0) != CUDA_SUCCESS) return g->va[r] = 0, vm_teardown(g), *why = "cuMemAddressReserve failed", false;
        if (d.MemMap(g->va[r], g->vm_size, 0, g->mem[r], 0) != CUDA_SUCCESS) {
            d.MemAddressFree(g->va[r], g->vm_size);
            g->va[r] = 0;
            return vm_teardown(g), *why = "cuMemMap failed", false;
        }
        if (d.MemSetAccess(g->va[r], g->vm_size, access.data(), access.size()) != CUDA_SUCCESS) return vm_teardown(g), *why = "cuMemSetAccess failed", false;
        if (d.MulticastBindMem(g->mc, 0, g->mem[r], 0, g->vm_size, 0) != CUDA_SUCCESS) return vm_teardown(g), *why = "cuMulticastBindMem failed", false;
    }
    if (d.MemAddressReserve(&g->mc_va, g->vm_size, gran, 0, 0) != CUDA_SUCCESS) return g->mc_va = 0, vm_teardown(g), *why = "cuMemAddressReserve (multicast) failed", false;
    if (d.MemMap(g->mc_va, g->vm_size, 0, g->mc, 0) != CUDA_SUCCESS) return vm_teardown(g), *why = "cuMemMap (multicast) failed", false;
    g->mc_mapped = true;
    if (d.MemSetAccess(g->mc_va, g->vm_size, access.data(), access.size()) != CUDA_SUCCESS) return vm_teardown(g), *why = "cuMemSetAccess (multicast) failed", false;
    for (uint32_t r = 0; r < g->n; ++r) {  // fresh physical memory: the mailboxes must start from zero
        cudaSetDevice(g->devices[r]);
        if (cudaMemset(reinterpret_cast<void*>(g->va[r]), 0, g->vm_size) != cudaSuccess) return vm_teardown(g), *why = "clearing a window failed", false;
    }
    for (int dev : g->devices) {
        cudaSetDevice(dev);
        cudaDeviceSynchronize();
    }
    return true;
}

}  // namespace

extern "C" {

const char* rcull_mgpu_last_error(void) { return g_mgpu_err[0] ? g_mgpu_err : rcull_last_error(); }

void rcull_mgpu_destroy(rcull_mgpu* g) {
    if (!g) return;
    for (uint32_t r = 0; r < g->merge.size(); ++r)
        if (g->merge[r]) rcull_merge_destroy(g->merge[r]);
    vm_teardown(g);
    for (rcull_ctx* c : g->ctx)
        if (c) rcull_destroy(c);
    delete g;
}

int rcull_mgpu_create(const rcull_config* cfg, uint32_t n_gpus, const int32_t* devices, uint64_t capacity, uint64_t capacity_per_rank,
                      uint32_t use_multicast, rcull_mgpu** out) {
    g_mgpu_err[0] = 0;
    if (!cfg || !out || n_gpus < 1 || n_gpus > RCULL_MAX_RANKS) {
        snprintf(g_mgpu_err, sizeof g_mgpu_err, "rcull_mgpu_create: bad argument (1..%d GPUs)", RCULL_MAX_RANKS);
        return RCULL_ERR_INVALID;
    }
    *out = nullptr;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < (int)n_gpus) {
        snprintf(g_mgpu_err, sizeof g_mgpu_err, "rcull_mgpu_create: %u GPUs asked, %d visible (librcull has no CPU path)", n_gpus, n_dev);
        return RCULL_ERR_NO_DEVICE;
    }
    rcull_mgpu* g = new rcull_mgpu();
    g->n = n_gpus;
    for (uint32_t r = 0; r < n_gpus; ++r) g->devices.push_back(devices ? devices[r] : (int)r);
    g->ctx.assign(n_gpus, nullptr);
    g->merge.assign(n_gpus, nullptr);
    g->stream.assign(n_gpus, nullptr);
    for (uint32_t r = 0; r < n_gpus; ++r) {
        rcull_config c = *cfg;
        c.device = g->devices[r];
        if (rcull_create(&c, &g->ctx[r]) != RCULL_OK || rcull_context_stream(g->ctx[r], &g->stream[r]) != RCULL_OK) {
            int rc = mfail("rcull_mgpu_create: rcull_create");
            rcull_mgpu_destroy(g);
            return rc;
        }
    }
    // every device stores into every other device's window: peer access both ways
    for (uint32_t a = 0; a < n_gpus; ++a)
        for (uint32_t b = 0; b < n_gpus; ++b) {
            if (a == b) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, g->devices[a], g->devices[b]);
            if (!can) {
                snprintf(g_mgpu_err, sizeof g_mgpu_err, "rcull_mgpu_create: device %d cannot access device %d", g->devices[a], g->devices[b]);
                rcull_mgpu_destroy(g);
                return RCULL_ERR_CUDA;
            }
            cudaSetDevice(g->devices[a]);
            cudaError_t e = cudaDeviceEnablePeerAccess(g->devices[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                snprintf(g_mgpu_err, sizeof g_mgpu_err, "rcull_mgpu_create: cudaDeviceEnablePeerAccess(%d -> %d): %s", g->devices[a], g->devices[b], cudaGetErrorString(e));
                rcull_mgpu_destroy(g);
                return RCULL_ERR_CUDA;
            }
            cudaGetLastError();
        }
    uint64_t window_bytes = 0;
    if (rcull_merge_window_bytes(g->ctx[0], n_gpus, capacity_per_rank ? capacity_per_rank : capacity, &window_bytes) != RCULL_OK) {
        int rc = mfail("rcull_mgpu_create: rcull_merge_window_bytes");
        rcull_mgpu_destroy(g);
        return rc;
    }
    std::string why;
    const bool vm = use_multicast && n_gpus > 1 && vm_windows(g, window_bytes, &why);
    g->transport = vm ? "virtual-memory windows, NVLS multicast stores (cuMulticastCreate)"
                      : std::string("peer stores (cudaDeviceEnablePeerAccess)") + (use_multicast && n_gpus > 1 ? "; multicast unavailable: " + why : "");
    for (uint32_t r = 0; r < n_gpus; ++r) {
        int rc = vm ? rcull_merge_create_on(g->ctx[r], n_gpus, r, capacity, capacity_per_rank, reinterpret_cast<void*>(g->va[r]), &g->merge[r])
                    : rcull_merge_create(g->ctx[r], n_gpus, r, capacity, capacity_per_rank, &g->merge[r]);
        if (rc != RCULL_OK) {
            rc = mfail("rcull_mgpu_create: rcull_merge_create");
            rcull_mgpu_destroy(g);
            return rc;
        }
    }
    for (uint32_t r = 0; r < n_gpus; ++r) {
        for (uint32_t p = 0; p < n_gpus; ++p) {
            void* w = nullptr;
            if (rcull_merge_window(g->merge[p], &w) != RCULL_OK || rcull_merge_set_peer_ptr(g->merge[r], p, w) != RCULL_OK) {
                int rc = mfail("rcull_mgpu_create: naming the peers' windows");
                rcull_mgpu_destroy(g);
                return rc;
            }
        }
        if (vm) rcull_merge_set_multicast_ptr(g->merge[r], reinterpret_cast<void*>(g->mc_va));
    }
    *out = g;
    return RCULL_OK;
}

int rcull_mgpu_rank(rcull_mgpu* g, uint32_t rank, rcull_ctx** ctx, rcull_merge** merge, void** stream) {
    if (!g || rank >= g->n) {
        snprintf(g_mgpu_err, sizeof g_mgpu_err, "rcull_mgpu_rank: bad argument");
        return RCULL_ERR_INVALID;
    }
    if (ctx) *ctx = g->ctx[rank];
    if (merge) *merge = g->merge[rank];
    if (stream) *stream = g->stream[rank];
    return RCULL_OK;
}

const char* rcull_mgpu_transport(rcull_mgpu* g) { return g ? g->transport.c_str() : ""; }

int rcull_mgpu_frame(rcull_mgpu* g, const uint64_t* n, const uint32_t* draw_index_base, const rcull_view* views, uint32_t n_views,
                     const float cam_pos[3]) {
    g_mgpu_err[0] = 0;
    if (!g || !n || !draw_index_base) {
        snprintf(g_mgpu_err, sizeof g_mgpu_err, "rcull_mgpu_frame: null argument");
        return RCULL_ERR_INVALID;
    }
    // every rank's frame is enqueued before any rank's expand: the expand of a rank waits (on the device) for the
    // records of all ranks, and a host thread that issued them in rank order would otherwise deadlock itself
    for (uint32_t r = 0; r < g->n; ++r)
        if (rcull_frame(g->ctx[r], n[r], views, n_views, cam_pos, draw_index_base[r], 0, g->stream[r]) != RCULL_OK) return mfail("rcull_mgpu_frame: rcull_frame");
    for (uint32_t r = 0; r < g->n; ++r)
        if (rcull_merge_expand(g->merge[r], g->stream[r]) != RCULL_OK) return mfail("rcull_mgpu_frame: rcull_merge_expand");
    return RCULL_OK;
}

int rcull_mgpu_synchronize(rcull_mgpu* g) {
    if (!g) return RCULL_ERR_INVALID;
    for (uint32_t r = 0; r < g->n; ++r) {
        cudaSetDevice(g->devices[r]);
        cudaError_t e = cudaStreamSynchronize(static_cast<cudaStream_t>(g->stream[r]));
        if (e != cudaSuccess) {
            snprintf(g_mgpu_err, sizeof g_mgpu_err, "rcull_mgpu_synchronize: rank %u: %s", r, cudaGetErrorString(e));
            return RCULL_ERR_CUDA;
        }
    }
    return RCULL_OK;
}

}  // extern "C"
