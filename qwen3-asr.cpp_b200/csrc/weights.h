// Host-side weight preparation shared by the encoder (encoder.cu) and the aligner's decoder (decoder.cu): rounding to the
// 16-bit operand type, the LayerNorm / RMSNorm fold, and the one-arena upload plan.
#pragma once
#include "common.cuh"
#include "gguf_reader.h"

#include <cstring>
#include <string>
#include <vector>

namespace q3a {

// ---- 16-bit conversion on the host ---------------------------------------------------------------
inline uint16_t to_op(float v, int fmt) {
    if (fmt == OP_F16) { __half h = __float2half_rn(v); uint16_t u; std::memcpy(&u, &h, 2); return u; }
    __nv_bfloat16 b = __float2bfloat16_rn(v);
    uint16_t u; std::memcpy(&u, &b, 2);
    return u;
}

inline float op_to_float(uint16_t u, int fmt) {
    if (fmt == OP_F16) { __half h; std::memcpy(&h, &u, 2); return __half2float(h); }
    __nv_bfloat16 b; std::memcpy(&b, &u, 2);
    return __bfloat162float(b);
}

// LayerNorm fold at load time (epilogues.cuh): w [N][K] f32 is replaced by gamma o w; returns bias' = bias + beta . w
// (from the unfolded weights, double accumulation) and the column sums of the folded weights AFTER rounding to the
// operand type -- the epilogue subtracts mean * colsum from what the tensor cores computed with exactly those values.
inline void fold_layernorm(std::vector<float>& w, int N, int K, const std::vector<float>& gamma, const std::vector<float>& beta,
                           const std::vector<float>& bias, int fmt, std::vector<float>& bias_out, std::vector<float>& colsum) {
    bias_out.assign(N, 0.0f); colsum.assign(N, 0.0f);
    for (int n = 0; n < N; ++n) {
        double bw = 0.0, sw = 0.0;
        float* row = w.data() + (size_t)n * K;
        for (int k = 0; k < K; ++k) {
            bw += (double)beta[k] * row[k];
            row[k] *= gamma[k];
            sw += (double)op_to_float(to_op(row[k], fmt), fmt);
        }
        bias_out[n] = (float)((double)bias[n] + bw);
        colsum[n] = (float)sw;
    }
}

// ---- weight arena ----------------------------------------------------------------------------------
struct ArenaPlan {
    struct Item { const void** slot; std::vector<uint8_t> bytes; };
    std::vector<Item> items;
    void add_f32(const float** slot, const std::vector<float>& v) {
        Item it; it.slot = reinterpret_cast<const void**>(slot);
        it.bytes.resize(v.size() * 4);
        std::memcpy(it.bytes.data(), v.data(), it.bytes.size());
        items.push_back(std::move(it));
    }
    void add_raw(const void** slot, const std::vector<uint8_t>& v) {
        Item it; it.slot = slot; it.bytes = v;
        items.push_back(std::move(it));
    }
    void add_bytes(const void** slot, const void* src, size_t n) {
        Item it; it.slot = slot;
        it.bytes.assign(static_cast<const uint8_t*>(src), static_cast<const uint8_t*>(src) + n);
        items.push_back(std::move(it));
    }
    void add_op(const void** slot, const std::vector<float>& v, int fmt) {
        Item it; it.slot = slot;
        it.bytes.resize(v.size() * 2);
        uint16_t* d = reinterpret_cast<uint16_t*>(it.bytes.data());
        for (size_t i = 0; i < v.size(); ++i) d[i] = to_op(v[i], fmt);
        items.push_back(std::move(it));
    }
};

inline bool fetch(const GgufFile& f, const std::string& name, std::vector<float>& out, std::string& err,
                  int64_t expect = -1) {
    const GgufTensor* t = f.find(name);
    if (!t) { err = "missing tensor " + name; return false; }
    if (!f.read_f32(*t, out, err)) return false;
    if (expect >= 0 && (int64_t)out.size() != expect) {
        err = "tensor " + name + " has " + std::to_string(out.size()) + " elements, expected " + std::to_string(expect);
        return false;
    }
    return true;
}


// copies every planned item into one device arena (256-byte aligned) and patches the slots
inline int upload_arena(ArenaPlan& plan, DevBuf& arena, size_t* used) {
    size_t total = 0;
    for (auto& it : plan.items) total += (it.bytes.size() + 255) / 256 * 256;
    Q3A_TRY(arena.reserve(total));
    size_t off = 0;
    std::vector<uint8_t> host(total, 0);
    for (auto& it : plan.items) {
        std::memcpy(host.data() + off, it.bytes.data(), it.bytes.size());
        *it.slot = arena.as<uint8_t>() + off;
        off += (it.bytes.size() + 255) / 256 * 256;
    }
    Q3A_CUDA(cudaMemcpy(arena.p, host.data(), total, cudaMemcpyHostToDevice));
    if (used) *used = total;
    return Q3A_OK;
}

}  // namespace q3a
