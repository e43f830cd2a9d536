// C-ABI plumbing shared by every translation unit of libq3a: per-thread error message, device
// selection (sm_100 only - there is no other code path), version.
#include "common.cuh"

#include <cstdlib>
#include <map>
#include <mutex>
#include <vector>

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

int kernel_smem_optin(const void* kernel, int bytes) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, int> done;      // (kernel, device) -> bytes opted in
    int dev = 0;
    Q3A_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    auto it = done.find({kernel, dev});
    if (it != done.end() && it->second >= bytes) return Q3A_OK;
    Q3A_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    done[{kernel, dev}] = bytes;
    return Q3A_OK;
}

int sm_count_cached() {
    static std::mutex mu;
    static std::map<int, int> cache;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 148; }
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(dev);
    if (it != cache.end()) return it->second;
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 148; }
    cache[dev] = n;
    return n;
}

// ---- Q3A_GUARD: guard bands around every DevBuf allocation (see common.cuh) ------------------------------
static std::mutex g_guard_mu;
static std::map<void*, size_t> g_guards;            // cudaMalloc base -> user bytes between the bands

bool guard_enabled() {
    static const bool on = [] { const char* e = getenv("Q3A_GUARD"); return e && *e && *e != '0'; }();
    return on;
}
void guard_register(void* base, size_t user_bytes) {
    std::lock_guard<std::mutex> lk(g_guard_mu);
    g_guards[base] = user_bytes;
}
void guard_forget(void* base) {
    std::lock_guard<std::mutex> lk(g_guard_mu);
    g_guards.erase(base);
}

}  // namespace q3a

// Every live guarded allocation of the process: both bands copied back and compared with the fill pattern.
extern "C" int q3a_debug_check_guards(int* n_bands, int* n_damaged) {
    using namespace q3a;
    if (!n_bands || !n_damaged) { set_error("debug_check_guards: null argument"); return Q3A_ERR_ARG; }
    *n_bands = *n_damaged = 0;
    if (!guard_enabled()) { set_error("debug_check_guards: Q3A_GUARD was not set when the library allocated"); return Q3A_ERR_STATE; }
    Q3A_CUDA(cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lk(g_guard_mu);
    std::vector<unsigned char> host(kGuardBytes);
    std::string where;
    for (const auto& kv : g_guards) {
        for (int side = 0; side < 2; ++side) {
            const char* band = static_cast<const char*>(kv.first) + (side ? kGuardBytes + kv.second : 0);
            Q3A_CUDA(cudaMemcpy(host.data(), band, kGuardBytes, cudaMemcpyDefault));
            ++*n_bands;
            size_t bad = 0, first = 0;
            for (size_t i = 0; i < kGuardBytes; ++i)
                if (host[i] != 0xA5) { if (!bad) first = i; ++bad; }
            if (bad) {
                ++*n_damaged;
                where += (side ? " after" : " before") + std::string(" a ") + std::to_string(kv.second) + "-byte buffer (" +
                         std::to_string(bad) + " bytes, first at band offset " + std::to_string(first) + ");";
            }
        }
    }
    if (*n_damaged) set_error("guard bands damaged:" + where);
    return Q3A_OK;
}

// Detector self-test: one byte written just past the end of the first live guarded buffer.
extern "C" int q3a_debug_damage_guard(void) {
    using namespace q3a;
    std::lock_guard<std::mutex> lk(g_guard_mu);
    if (g_guards.empty()) { set_error("debug_damage_guard: no guarded buffer is live"); return Q3A_ERR_STATE; }
    const auto& kv = *g_guards.begin();
    Q3A_CUDA(cudaMemset(static_cast<char*>(kv.first) + kGuardBytes + kv.second, 0, 1));
    return Q3A_OK;
}

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
