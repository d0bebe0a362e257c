// api.cu -- library-wide state of libmsr.so: last-error string, device selection, launch counter.
#include <stdarg.h>

#include "common.cuh"

namespace msr {

static thread_local char g_err[1024] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int use_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        (void)cudaGetLastError();
        set_error("no CUDA device is usable (%s); libmsr has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return MSR_ERR_CUDA;
    }
    if (device < 0 || device >= n) {
        set_error("device %d out of range [0,%d)", device, n);
        return MSR_ERR_INVALID;
    }
    static thread_local int checked = -1;
    if (checked != device) {
        int major = 0;
        MSR_CUDA_CHECK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
        if (major != 10) {
            set_error("device %d has compute capability %d.x; libmsr is built for sm_100a (B200) only", device, major);
            return MSR_ERR_CUDA;
        }
        checked = device;
    }
    MSR_CUDA_CHECK(cudaSetDevice(device));
    return MSR_OK;
}

}  // namespace msr

extern "C" int msr_version(void) { return 100; }
extern "C" const char* msr_last_error(void) { return msr::g_err; }
extern "C" uint64_t msr_kernel_launches(void) { return msr::g_launches.load(); }
