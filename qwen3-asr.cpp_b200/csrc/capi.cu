// C-ABI plumbing shared by every translation unit of libq3a: per-thread error message, device
// selection (sm_100 only - there is no other code path), version.
#include "common.cuh"

#include <mutex>

namespace q3a {

static thread_local std::string t_error;
thread_local int g_launch_counter = 0;

void set_error(const std::string& msg) { t_error = msg; }
const char* last_error() { return t_error.c_str(); }

int use_device(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        set_error(std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                  " (libq3a has no CPU fallback)");
        cudaGetLastError();
        return Q3A_ERR_CUDA;
    }
    if (device < 0 || device >= count) { set_error("device index out of range"); return Q3A_ERR_ARG; }
    int major = 0;
    Q3A_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10) {
        set_error("device " + std::to_string(device) + " is not sm_100 (compute capability major " +
                  std::to_string(major) + "); libq3a is built for sm_100a only");
        return Q3A_ERR_CUDA;
    }
    Q3A_CUDA(cudaSetDevice(device));
    return Q3A_OK;
}

}  // namespace q3a

extern "C" const char* q3a_last_error(void) { return q3a::last_error(); }
extern "C" int q3a_version(void) { return 100; }
extern "C" int q3a_device_count(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int d = 0; d < count; ++d) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) ++ok;
    }
    return ok;
}
