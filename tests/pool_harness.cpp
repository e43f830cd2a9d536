// A single-process C++ caller in the style of the reference's main.cpp:294-504 that uses every visible GPU through the
// pool executor, and the asynchronous submit / wait form on one encoder.  Plain C++17 against include/q3a.h:
//     g++ -std=c++17 -Iinclude tests/pool_harness.cpp -Lqwen3-asr.cpp_b200 -lq3a
// usage: pool_harness model.gguf n_replicas out.f32     (clips are synthesised here; features of all clips, in order)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "q3a.h"

static std::vector<float> make_clip(int seed, int n) {
    std::vector<float> v((size_t)n);
    unsigned s = 12345u + 977u * (unsigned)seed;
    for (int i = 0; i < n; ++i) {
        s = s * 1664525u + 1013904223u;
        const float noise = ((s >> 9) & 0xFFFF) / 65536.0f - 0.5f;
        v[i] = 0.3f * std::sin(0.002f * (float)(seed + 3) * (float)i) + 0.05f * noise;
    }
    return v;
}

int main(int argc, char** argv) {
    if (argc < 4) { std::fprintf(stderr, "usage: %s model.gguf n_replicas out.f32\n", argv[0]); return 1; }
    const int n_rep = std::atoi(argv[2]);
    const int n_dev = q3a_device_count();
    if (n_dev < 1) { std::fprintf(stderr, "no CUDA device (libq3a has no CPU fallback)\n"); return 2; }
    std::vector<int> devices(n_rep);
    for (int i = 0; i < n_rep; ++i) devices[i] = i % n_dev;
    q3a_pool* pool = nullptr;
    if (q3a_pool_create(argv[1], devices.data(), n_rep, Q3A_OP_F16, &pool) != Q3A_OK) { std::fprintf(stderr, "%s\n", q3a_last_error()); return 3; }

    const int lens[] = {16000 * 3 + 77, 0, 16000 * 7, 1600, 16000 * 2 + 5, 16000 * 11, 159, 16000 * 5 + 321};
    const int n = (int)(sizeof(lens) / sizeof(lens[0]));
    std::vector<std::vector<float>> clips;
    std::vector<const float*> pcm(n);
    std::vector<int> ns(n), ntok(n), shard(n);
    std::vector<float*> out(n);
    size_t total = 0;
    for (int i = 0; i < n; ++i) { clips.push_back(make_clip(i, lens[i])); ns[i] = lens[i]; total += (size_t)q3a_encoder_n_tokens(lens[i] / 160, Q3A_MODE_ASR); }
    // out_dim is not known to the pool's caller without an encoder handle: the harness is run on towers that project to 64
    const int width = 64;
    float* block = nullptr;
    if (q3a_host_alloc(sizeof(float) * total * width + 16, reinterpret_cast<void**>(&block)) != Q3A_OK) { std::fprintf(stderr, "%s\n", q3a_last_error()); return 3; }
    size_t row = 0;
    for (int i = 0; i < n; ++i) { pcm[i] = clips[i].data(); out[i] = block + row * width; row += (size_t)q3a_encoder_n_tokens(lens[i] / 160, Q3A_MODE_ASR); }
    std::vector<float> ms(n_rep);
    if (q3a_pool_encode_pcm(pool, pcm.data(), ns.data(), n, Q3A_MODE_ASR, out.data(), ntok.data(), shard.data(), ms.data()) != Q3A_OK) {
        std::fprintf(stderr, "%s\n", q3a_last_error()); return 4;
    }
    for (int i = 0; i < n; ++i) std::printf("utt %d: %d tokens on replica %d\n", i, ntok[i], shard[i]);

    // the same corpus through one encoder with two asynchronous calls in flight: must give the same bits
    q3a_frontend* fe = nullptr; q3a_encoder* enc = nullptr;
    if (q3a_frontend_create(0, nullptr, 0, 0, &fe) != Q3A_OK || q3a_encoder_load(argv[1], 0, Q3A_OP_F16, &enc) != Q3A_OK) { std::fprintf(stderr, "%s\n", q3a_last_error()); return 5; }
    std::vector<float> again(total * width, -1.0f);
    std::vector<float*> out2(n);
    row = 0;
    for (int i = 0; i < n; ++i) { out2[i] = again.data() + row * width; row += (size_t)ntok[i]; }
    const int half = n / 2;
    int64_t t0 = -1, t1 = -1;
    std::vector<int> nt2(n);
    if (q3a_encode_pcm_submit(fe, enc, pcm.data(), ns.data(), half, Q3A_MODE_ASR, out2.data(), nt2.data(), &t0) != Q3A_OK ||
        q3a_encode_pcm_submit(fe, enc, pcm.data() + half, ns.data() + half, n - half, Q3A_MODE_ASR, out2.data() + half, nt2.data() + half, &t1) != Q3A_OK ||
        q3a_encode_wait(enc, t0) != Q3A_OK || q3a_encode_wait(enc, t1) != Q3A_OK) { std::fprintf(stderr, "%s\n", q3a_last_error()); return 6; }
    if (std::memcmp(again.data(), block, sizeof(float) * total * width) != 0) { std::fprintf(stderr, "pool and asynchronous results differ\n"); return 7; }
    std::printf("pool == async: %zu rows identical\n", total);

    FILE* f = std::fopen(argv[3], "wb");
    if (!f || std::fwrite(block, sizeof(float), total * width, f) != total * width) return 1;
    std::fclose(f);
    q3a_encoder_free(enc); q3a_frontend_free(fe); q3a_host_free(block); q3a_pool_free(pool);
    return 0;
}
