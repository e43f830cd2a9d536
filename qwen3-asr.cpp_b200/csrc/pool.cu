// In-process multi-GPU executor behind the C ABI: one replica (frontend + encoder) and one host worker thread per device.
//
// The reference encodes one utterance at a time on one device (qwen3_asr.cpp:81-149, main.cpp:294-504 is a single process);
// a corpus therefore shards by utterance with no cross-shard reduction at all -- SURVEY 8(e).  q3a_pool_encode_pcm takes a
// whole corpus of host clips, assigns them to the replicas with the same longest-processing-time plan q3a_plan_shards
// exposes, lets every replica encode its shard through q3a_encode_pcm_batch (its own streams: upload | compute | download)
// and returns the features in utterance order, because every utterance writes into the caller's own out[i].  There is no
// collective and no device-to-device traffic; the kernels are batch-invariant, so an utterance's features are bit-identical
// whichever replica, shard or batch it lands in (tests/test_pool_gpu.py).
//
// Host placement: every worker thread binds itself to the CPUs the kernel reports as local to its GPU
// (/sys/bus/pci/devices/<bus id>/local_cpulist) before it touches pinned memory, and q3a_host_alloc hands out pinned,
// portable staging memory, so that a plain C++ caller needs neither NVML nor CUDA to place buffers and threads.
#include "encoder.h"
#include "mel.h"

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <mutex>
#include <sched.h>
#include <string>
#include <thread>
#include <vector>

using namespace q3a;

namespace {

// CPUs local to the GPU, from sysfs (no NVML dependency).  Empty when unknown.
std::vector<int> gpu_local_cpus(int device) {
    std::vector<int> cpus;
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof(bus), device) != cudaSuccess) { cudaGetLastError(); return cpus; }
    std::string id(bus);
    std::transform(id.begin(), id.end(), id.begin(), [](unsigned char c) { return (char)std::tolower(c); });
    std::ifstream f("/sys/bus/pci/devices/" + id + "/local_cpulist");
    std::string list;
    if (!f || !std::getline(f, list)) return cpus;
    size_t pos = 0;
    while (pos < list.size()) {                      // "0-31,64-95"
        size_t end = list.find(',', pos);
        if (end == std::string::npos) end = list.size();
        const std::string tok = list.substr(pos, end - pos);
        int a = 0, b = 0;
        if (sscanf(tok.c_str(), "%d-%d", &a, &b) == 2) { for (int c = a; c <= b; ++c) cpus.push_back(c); }
        else if (sscanf(tok.c_str(), "%d", &a) == 1) cpus.push_back(a);
        pos = end + 1;
    }
    return cpus;
}

// restrict the calling thread to the GPU's CPUs (intersected with what it may already run on); returns CPUs bound or 0
int bind_thread_to_gpu(int device) {
    const std::vector<int> local = gpu_local_cpus(device);
    if (local.empty()) return 0;
    cpu_set_t allowed, want;
    CPU_ZERO(&allowed); CPU_ZERO(&want);
    if (sched_getaffinity(0, sizeof(allowed), &allowed) != 0) return 0;
    int n = 0;
    for (int c : local) if (c >= 0 && c < CPU_SETSIZE && CPU_ISSET(c, &allowed)) { CPU_SET(c, &want); ++n; }
    if (n == 0 || n == CPU_COUNT(&allowed)) return 0;          // nothing to gain, or the cpuset excludes the GPU's cores
    return sched_setaffinity(0, sizeof(want), &want) == 0 ? n : 0;
}

}  // namespace

struct q3a_pool {
    struct Replica {
        int device = 0;
        q3a_frontend* fe = nullptr;
        q3a_encoder* enc = nullptr;
        std::thread th;
        // one job slot per replica, handed over under `mu`
        std::mutex mu;
        std::condition_variable cv;
        bool has_job = false, quit = false, done = false;
        std::vector<int> idx;                 // utterance indices of this replica's shard, ascending
        int rc = Q3A_OK;
        std::string err;
        float ms = 0.0f;
        int bound_cpus = 0;
    };
    std::vector<Replica*> reps;
    // the call being served (valid while a job is outstanding)
    const float* const* pcm = nullptr;
    const int* n_samples = nullptr;
    float* const* out = nullptr;
    int* n_tokens = nullptr;
    int mode = 0;
    std::mutex call_mu;                       // one q3a_pool_encode_pcm at a time per pool
};

static void pool_worker(q3a_pool* p, q3a_pool::Replica* r) {
    r->bound_cpus = bind_thread_to_gpu(r->device);
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(r->mu);
            r->cv.wait(lk, [&] { return r->has_job || r->quit; });
            if (r->quit) return;
        }
        const auto t0 = std::chrono::steady_clock::now();
        int rc = Q3A_OK;
        const int n = (int)r->idx.size();
        if (n > 0) {
            std::vector<const float*> in(n);
            std::vector<int> ns(n), nt(n);
            std::vector<float*> outs(n);
            for (int i = 0; i < n; ++i) { in[i] = p->pcm[r->idx[i]]; ns[i] = p->n_samples[r->idx[i]]; outs[i] = p->out[r->idx[i]]; }
            rc = q3a_encode_pcm_batch(r->fe, r->enc, in.data(), ns.data(), n, p->mode, outs.data(), nt.data());
            if (rc == Q3A_OK && p->n_tokens)
                for (int i = 0; i < n; ++i) p->n_tokens[r->idx[i]] = nt[i];
        }
        r->ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        {
            std::lock_guard<std::mutex> lk(r->mu);
            r->rc = rc;
            r->err = rc == Q3A_OK ? std::string() : std::string(q3a_last_error());   // the message is per thread: carry it over
            r->has_job = false; r->done = true;
        }
        r->cv.notify_all();
    }
}

extern "C" void q3a_pool_free(q3a_pool* p) {
    if (!p) return;
    for (auto* r : p->reps) {
        if (r->th.joinable()) {
            { std::lock_guard<std::mutex> lk(r->mu); r->quit = true; }
            r->cv.notify_all();
            r->th.join();
        }
        if (r->enc) q3a_encoder_free(r->enc);
        if (r->fe) q3a_frontend_free(r->fe);
        delete r;
    }
    delete p;
}

extern "C" int q3a_pool_create(const char* gguf_path, const int* devices, int n_devices, int operand_dtype, q3a_pool** out) {
    if (!out) { set_error("pool_create: out is null"); return Q3A_ERR_ARG; }
    *out = nullptr;
    if (!gguf_path || !devices || n_devices <= 0 || n_devices > 64) { set_error("pool_create: bad argument"); return Q3A_ERR_ARG; }
    q3a_pool* p = new q3a_pool();
    for (int i = 0; i < n_devices; ++i) {
        auto* r = new q3a_pool::Replica();
        r->device = devices[i];
        p->reps.push_back(r);
        int rc = q3a_frontend_create(r->device, nullptr, 0, 0, &r->fe);
        if (rc == Q3A_OK) rc = q3a_encoder_load(gguf_path, r->device, operand_dtype, &r->enc);
        if (rc != Q3A_OK) { const std::string msg = q3a_last_error(); q3a_pool_free(p); set_error(msg); return rc; }
    }
    for (auto* r : p->reps) r->th = std::thread(pool_worker, p, r);
    *out = p;
    return Q3A_OK;
}

extern "C" int q3a_pool_n_devices(const q3a_pool* p) { return p ? (int)p->reps.size() : 0; }

extern "C" int q3a_pool_encode_pcm(q3a_pool* p, const float* const* pcm, const int* n_samples, int n_clips, int mode,
                                   float* const* out, int* n_tokens, int* shard_of, float* shard_ms) {
    if (!p) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("encode: unknown mode"); return Q3A_ERR_ARG; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm || !n_samples || !out))) { set_error("pool_encode_pcm: null argument"); return Q3A_ERR_ARG; }
    for (int u = 0; u < n_clips; ++u)
        if (n_samples[u] < 0 || (n_samples[u] > 0 && !pcm[u]) || (n_samples[u] >= 160 && !out[u])) {
            set_error("pool_encode_pcm: bad clip " + std::to_string(u)); return Q3A_ERR_ARG;
        }
    std::lock_guard<std::mutex> call(p->call_mu);
    const int n_rep = (int)p->reps.size();
    std::vector<int> nf(n_clips), plan(n_clips, 0);
    for (int i = 0; i < n_clips; ++i) nf[i] = n_samples[i] / 160;
    Q3A_TRY(q3a_plan_shards(nf.data(), n_clips, mode, n_rep, plan.data()));
    if (shard_of) std::memcpy(shard_of, plan.data(), sizeof(int) * (size_t)n_clips);
    p->pcm = pcm; p->n_samples = n_samples; p->out = out; p->n_tokens = n_tokens; p->mode = mode;
    for (int s = 0; s < n_rep; ++s) {
        auto* r = p->reps[s];
        std::lock_guard<std::mutex> lk(r->mu);
        r->idx.clear();
        for (int i = 0; i < n_clips; ++i) if (plan[i] == s) r->idx.push_back(i);
        r->done = false; r->has_job = true;
    }
    for (auto* r : p->reps) r->cv.notify_all();
    int rc = Q3A_OK;
    std::string err;
    for (int s = 0; s < n_rep; ++s) {
        auto* r = p->reps[s];
        std::unique_lock<std::mutex> lk(r->mu);
        r->cv.wait(lk, [&] { return r->done; });
        if (r->rc != Q3A_OK && rc == Q3A_OK) { rc = r->rc; err = "replica " + std::to_string(s) + " (device " + std::to_string(r->device) + "): " + r->err; }
        if (shard_ms) shard_ms[s] = r->ms;
    }
    if (rc != Q3A_OK) set_error(err);
    return rc;
}

// Pinned, portable host memory for the host-pointer entry points (PCM in, features out): copies from it run at full
// PCIe rate and asynchronously.  Call it from the thread that will fill the buffer, after binding that thread
// (q3a_bind_thread_to_device) when the box has more than one NUMA node.
extern "C" int q3a_host_alloc(size_t bytes, void** out) {
    if (!out) { set_error("host_alloc: out is null"); return Q3A_ERR_ARG; }
    *out = nullptr;
    if (bytes == 0) return Q3A_OK;
    Q3A_CUDA(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
    return Q3A_OK;
}
extern "C" void q3a_host_free(void* p) { if (p) cudaFreeHost(p); }
// Restricts the calling thread to the CPUs local to `device`; returns the number of CPUs it is now bound to, 0 when
// nothing was changed (single NUMA node, unknown topology, or a cpuset that excludes those CPUs), < 0 never.
extern "C" int q3a_bind_thread_to_device(int device) { return bind_thread_to_gpu(device); }
