// Minimal GGUF (v2/v3) reader: metadata KV + tensor directory + mmapped data, enough for the
// audio.encoder.* tensors the reference loads (gguf_loader.cpp:130-249, forced_aligner.cpp:204-266).
// Replaces the gguf_* / ggml-context plumbing of gguf_loader.cpp:17-116,255-322 with a plain parser:
// tensors are dequantised to f32 on the host and uploaded once.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace q3a {

enum GgmlType : uint32_t { GGML_F32 = 0, GGML_F16 = 1, GGML_Q8_0 = 8, GGML_BF16 = 30 };

struct GgufTensor {
    std::string name;
    std::vector<int64_t> ne;     // ggml order: ne[0] is the contiguous dimension
    uint32_t type = 0;
    uint64_t offset = 0;         // from the start of the data section
    int64_t n_elements() const { int64_t n = 1; for (auto d : ne) n *= d; return n; }
};

class GgufFile {
public:
    ~GgufFile();
    // returns false and fills err on failure
    bool open(const std::string& path, std::string& err);
    const GgufTensor* find(const std::string& name) const;
    bool has_u32(const std::string& key) const { return u32_.count(key) != 0; }
    uint32_t get_u32(const std::string& key, uint32_t dflt) const;
    float get_f32(const std::string& key, float dflt) const;
    // dequantise a tensor to f32 (F32 / F16 / BF16 / Q8_0); false if the type is unsupported or the data is truncated
    bool read_f32(const GgufTensor& t, std::vector<float>& out, std::string& err) const;
    // Q8_0 tensor as two planes: q[n_elements] int8 and d[n_elements / 32] f16 bit patterns (ggml block order kept)
    bool read_q8_planes(const GgufTensor& t, std::vector<int8_t>& q, std::vector<uint16_t>& d, std::string& err) const;
    const std::vector<GgufTensor>& tensors() const { return tensors_; }

private:
    void* map_ = nullptr;
    size_t map_size_ = 0;
    size_t data_start_ = 0;
    std::vector<GgufTensor> tensors_;
    std::map<std::string, size_t> index_;
    std::map<std::string, uint32_t> u32_;
    std::map<std::string, float> f32_;
};

}  // namespace q3a
