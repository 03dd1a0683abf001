// Error reporting + device queries for libvlb_b200. No exceptions cross the C ABI: every entry point returns
// 0 on success and a non-zero code otherwise; the message is fetched with vlb_last_error_string().
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

namespace vlb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

int num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

}  // namespace vlb

extern "C" const char* vlb_last_error_string() { return vlb::g_err; }

extern "C" int vlb_abi_version() { return 1; }

// Hard requirement of the library: an sm_100 device. There is no fallback path.
extern "C" int vlb_check_device() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { vlb::set_error("cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
    int major = 0, minor = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
    if (major != 10) {
        vlb::set_error("libvlb_b200 needs an sm_100a device (B200); found sm_%d%d", major, minor);
        return (int)cudaErrorNoKernelImageForDevice;
    }
    return 0;
}
