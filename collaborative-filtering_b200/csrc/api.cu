// api.cu — library-level entry points: ABI version, thread-local error string,
// device probe.  See include/b200als.h.
#include <stdarg.h>

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

}  // extern "C"
