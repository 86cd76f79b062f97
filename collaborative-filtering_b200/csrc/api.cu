// api.cu — library-level entry points: ABI version, thread-local error string,
// device probe.  See include/b200als.h.
#include <cuda.h>
#include <stdarg.h>

#include <map>
#include <mutex>
#include <string>

#include "common.cuh"

namespace b200als {

static thread_local char g_err[512] = "";
thread_local int g_launch_count = 0;

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

}  // namespace b200als

extern "C" {

int b200als_abi_version(void) { return B200ALS_ABI_VERSION; }

const char* b200als_last_error(void) { return b200als::g_err; }

int b200als_last_launch_count(void) { return b200als::g_launch_count; }

int b200als_device_ok(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        b200als::set_error("no CUDA device is current");
        return 0;
    }
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) {
        cudaGetLastError();
        b200als::set_error("cudaGetDeviceProperties failed");
        return 0;
    }
    if (p.major != 10) {
        b200als::set_error("device %s is sm_%d%d; libb200als is built for sm_100a only", p.name, p.major,
                           p.minor);
        return 0;
    }
    return 1;
}

int b200als_peer_enable(int32_t peer_device) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        b200als::set_error("no CUDA device is current");
        return B200ALS_E_CUDA;
    }
    if (peer_device == dev) return B200ALS_OK;
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, dev, peer_device) != cudaSuccess || !can) {
        cudaGetLastError();
        b200als::set_error("device %d cannot access device %d's memory", dev, peer_device);
        return B200ALS_E_UNSUPPORTED;
    }
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return B200ALS_OK;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        b200als::set_error("cudaDeviceEnablePeerAccess(%d) failed: %s", peer_device, cudaGetErrorString(e));
        return B200ALS_E_CUDA;
    }
    return B200ALS_OK;
}

// ---- replicas of other ranks (fused exchange): CUDA IPC, imported with the CALLER's device current ----------------
// The importing rank's kernels store into the peer's memory, so the mapping must live in the importing device's
// context: cudaIpcOpenMemHandle with that device current and cudaIpcMemLazyEnablePeerAccess (what NCCL's P2P
// transport does).  A mapping opened under the exporter's device context — what torch.multiprocessing's tensor
// rebuild does — is not reachable from a kernel of another device even with peer access enabled (measured: illegal
// address at N = 2).  A handle names a whole cudaMalloc block and may be opened once per context: opens are cached and
// reference counted per (device, handle).
namespace b200als {
struct IpcEntry { void* base; int refs; };
static std::mutex g_ipc_mu;
static std::map<std::string, IpcEntry> g_ipc_open;
static std::string ipc_key(int dev, const void* handle) {
    std::string k(reinterpret_cast<const char*>(handle), sizeof(cudaIpcMemHandle_t));
    k.push_back((char)dev);
    return k;
}
}  // namespace b200als

int b200als_ipc_export(const void* dev_ptr, void* handle64_host, int64_t* offset_host) {
    using namespace b200als;
    if (!dev_ptr || !handle64_host || !offset_host) {
        set_error("b200als_ipc_export: null argument");
        return B200ALS_E_INVALID;
    }
    // the driver entry point comes through the runtime (no link-time dependency on libcuda: the library must load on a
    // machine without a driver, where the CPU test-suite checks its exported symbols)
    typedef CUresult (*get_range_fn)(CUdeviceptr*, size_t*, CUdeviceptr);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t ge = cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qres);
    if (ge != cudaSuccess || !fn || qres != cudaDriverEntryPointSuccess) {
        cudaGetLastError();
        set_error("b200als_ipc_export: cuMemGetAddressRange is not available from this driver");
        return B200ALS_E_UNSUPPORTED;
    }
    CUdeviceptr base = 0;
    size_t size = 0;
    CUresult r = reinterpret_cast<get_range_fn>(fn)(&base, &size, (CUdeviceptr)(uintptr_t)dev_ptr);
    if (r != CUDA_SUCCESS) {
        set_error("b200als_ipc_export: cuMemGetAddressRange failed (%d)", (int)r);
        return B200ALS_E_CUDA;
    }
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, reinterpret_cast<void*>((uintptr_t)base));
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("b200als_ipc_export: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
        return B200ALS_E_CUDA;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    memcpy(handle64_host, &h, 64);
    *offset_host = (int64_t)((uintptr_t)dev_ptr - (uintptr_t)base);
    return B200ALS_OK;
}

int b200als_ipc_import(const void* handle64_host, int64_t offset, void** ptr_out_host) {
    using namespace b200als;
    if (!handle64_host || !ptr_out_host || offset < 0) {
        set_error("b200als_ipc_import: bad argument");
        return B200ALS_E_INVALID;
    }
    int dev = 0;
    ALS_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(g_ipc_mu);
    const std::string key = ipc_key(dev, handle64_host);
    auto it = g_ipc_open.find(key);
    if (it == g_ipc_open.end()) {
        cudaIpcMemHandle_t h;
        memcpy(&h, handle64_host, 64);
        void* base = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess || !base) {
            cudaGetLastError();
            set_error("b200als_ipc_import: cudaIpcOpenMemHandle failed on device %d: %s", dev, cudaGetErrorString(e));
            return B200ALS_E_CUDA;
        }
        it = g_ipc_open.emplace(key, IpcEntry{base, 0}).first;
    }
    it->second.refs++;
    *ptr_out_host = static_cast<char*>(it->second.base) + offset;
    return B200ALS_OK;
}

int b200als_ipc_release(const void* handle64_host) {
    using namespace b200als;
    if (!handle64_host) return B200ALS_OK;
    int dev = 0;
    ALS_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(g_ipc_mu);
    auto it = g_ipc_open.find(ipc_key(dev, handle64_host));
    if (it == g_ipc_open.end()) return B200ALS_OK;
    if (--it->second.refs <= 0) {
        cudaError_t e = cudaIpcCloseMemHandle(it->second.base);
        g_ipc_open.erase(it);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("b200als_ipc_release: cudaIpcCloseMemHandle failed: %s", cudaGetErrorString(e));
            return B200ALS_E_CUDA;
        }
    }
    return B200ALS_OK;
}

}  // extern "C"
