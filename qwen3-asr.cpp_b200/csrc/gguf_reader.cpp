#include "gguf_reader.h"

#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cuda_fp16.h>

namespace q3a {

namespace {

struct Cursor {
    const uint8_t* p;
    size_t n, pos = 0;
    bool ok = true;
    template <class T> T get() {
        T v{};
        if (pos + sizeof(T) > n) { ok = false; return v; }
        std::memcpy(&v, p + pos, sizeof(T));
        pos += sizeof(T);
        return v;
    }
    std::string str() {
        uint64_t len = get<uint64_t>();
        if (!ok || pos + len > n) { ok = false; return {}; }
        std::string s(reinterpret_cast<const char*>(p + pos), len);
        pos += len;
        return s;
    }
    void skip(size_t k) { if (pos + k > n) ok = false; else pos += k; }
};

// GGUF metadata value types
enum : uint32_t { T_U8, T_I8, T_U16, T_I16, T_U32, T_I32, T_F32, T_BOOL, T_STR, T_ARR, T_U64, T_I64, T_F64 };

size_t scalar_size(uint32_t t) {
    switch (t) {
        case T_U8: case T_I8: case T_BOOL: return 1;
        case T_U16: case T_I16: return 2;
        case T_U32: case T_I32: case T_F32: return 4;
        case T_U64: case T_I64: case T_F64: return 8;
        default: return 0;
    }
}

float half_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000) << 16;
    uint32_t exp = (h >> 10) & 0x1F, man = h & 0x3FF, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {
            int e = -1;
            do { man <<= 1; ++e; } while (!(man & 0x400));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3FF) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7F800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

}  // namespace

GgufFile::~GgufFile() {
    if (map_) munmap(map_, map_size_);
}

bool GgufFile::open(const std::string& path, std::string& err) {
    int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) { err = "cannot open " + path; return false; }
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size < 24) { ::close(fd); err = "cannot stat / file too small: " + path; return false; }
    map_size_ = (size_t)st.st_size;
    map_ = mmap(nullptr, map_size_, PROT_READ, MAP_PRIVATE, fd, 0);
    ::close(fd);
    if (map_ == MAP_FAILED) { map_ = nullptr; err = "mmap failed: " + path; return false; }

    Cursor c{static_cast<const uint8_t*>(map_), map_size_};
    if (std::memcmp(c.p, "GGUF", 4) != 0) { err = "not a GGUF file (bad magic): " + path; return false; }
    c.pos = 4;
    const uint32_t version = c.get<uint32_t>();
    if (version < 2 || version > 3) { err = "unsupported GGUF version " + std::to_string(version); return false; }
    const uint64_t n_tensors = c.get<uint64_t>();
    const uint64_t n_kv = c.get<uint64_t>();
    uint32_t alignment = 32;
    for (uint64_t i = 0; i < n_kv && c.ok; ++i) {
        std::string key = c.str();
        uint32_t type = c.get<uint32_t>();
        if (type == T_STR) { c.str(); continue; }
        if (type == T_ARR) {
            uint32_t et = c.get<uint32_t>();
            uint64_t cnt = c.get<uint64_t>();
            if (et == T_STR) { for (uint64_t k = 0; k < cnt && c.ok; ++k) c.str(); }
            else { size_t s = scalar_size(et); if (!s) { c.ok = false; break; } c.skip(s * cnt); }
            continue;
        }
        const size_t s = scalar_size(type);
        if (!s) { c.ok = false; break; }
        if (type == T_U32 || type == T_I32) u32_[key] = c.get<uint32_t>();
        else if (type == T_F32) f32_[key] = c.get<float>();
        else if (type == T_U64 || type == T_I64) u32_[key] = (uint32_t)c.get<uint64_t>();
        else if (type == T_U16 || type == T_I16) u32_[key] = c.get<uint16_t>();
        else if (type == T_U8 || type == T_I8 || type == T_BOOL) u32_[key] = c.get<uint8_t>();
        else c.skip(s);
        if (key == "general.alignment" && u32_.count(key)) alignment = u32_[key];
    }
    if (!c.ok) { err = "truncated or malformed GGUF metadata: " + path; return false; }
    tensors_.resize(n_tensors);
    for (uint64_t i = 0; i < n_tensors && c.ok; ++i) {
        GgufTensor& t = tensors_[i];
        t.name = c.str();
        const uint32_t nd = c.get<uint32_t>();
        if (nd > 8) { c.ok = false; break; }
        t.ne.resize(nd);
        for (uint32_t d = 0; d < nd; ++d) t.ne[d] = (int64_t)c.get<uint64_t>();
        t.type = c.get<uint32_t>();
        t.offset = c.get<uint64_t>();
        index_[t.name] = i;
    }
    if (!c.ok) { err = "truncated or malformed GGUF tensor directory: " + path; return false; }
    if (alignment == 0) alignment = 32;
    data_start_ = (c.pos + alignment - 1) / alignment * alignment;
    return true;
}

const GgufTensor* GgufFile::find(const std::string& name) const {
    auto it = index_.find(name);
    return it == index_.end() ? nullptr : &tensors_[it->second];
}

uint32_t GgufFile::get_u32(const std::string& key, uint32_t dflt) const {
    auto it = u32_.find(key);
    return it == u32_.end() ? dflt : it->second;
}

float GgufFile::get_f32(const std::string& key, float dflt) const {
    auto it = f32_.find(key);
    return it == f32_.end() ? dflt : it->second;
}

bool GgufFile::read_f32(const GgufTensor& t, std::vector<float>& out, std::string& err) const {
    const int64_t n = t.n_elements();
    size_t bytes = 0;
    switch (t.type) {
        case GGML_F32: bytes = (size_t)n * 4; break;
        case GGML_F16: case GGML_BF16: bytes = (size_t)n * 2; break;
        case GGML_Q8_0:
            if (t.ne.empty() || t.ne[0] % 32) { err = "Q8_0 tensor with a row length not a multiple of 32: " + t.name; return false; }
            bytes = (size_t)(n / 32) * 34;
            break;
        default: err = "unsupported tensor type " + std::to_string(t.type) + " for " + t.name; return false;
    }
    if (data_start_ + t.offset + bytes > map_size_) { err = "tensor data out of file bounds: " + t.name; return false; }
    const uint8_t* src = static_cast<const uint8_t*>(map_) + data_start_ + t.offset;
    out.resize((size_t)n);
    if (t.type == GGML_F32) std::memcpy(out.data(), src, bytes);
    else if (t.type == GGML_F16) {
        const uint16_t* h = reinterpret_cast<const uint16_t*>(src);
        for (int64_t i = 0; i < n; ++i) out[i] = half_to_float(h[i]);
    } else if (t.type == GGML_BF16) {
        const uint16_t* h = reinterpret_cast<const uint16_t*>(src);
        for (int64_t i = 0; i < n; ++i) { uint32_t b = (uint32_t)h[i] << 16; std::memcpy(&out[i], &b, 4); }
    } else {   // Q8_0: { f16 d; int8 q[32]; } per 32 values, value = d * q
        for (int64_t b = 0; b < n / 32; ++b) {
            const uint8_t* blk = src + b * 34;
            uint16_t dh;
            std::memcpy(&dh, blk, 2);
            const float d = half_to_float(dh);
            const int8_t* q = reinterpret_cast<const int8_t*>(blk + 2);
            for (int i = 0; i < 32; ++i) out[b * 32 + i] = d * (float)q[i];
        }
    }
    return true;
}

bool GgufFile::read_q8_planes(const GgufTensor& t, std::vector<int8_t>& q, std::vector<uint16_t>& d, std::string& err) const {
    if (t.type != GGML_Q8_0) { err = "tensor " + t.name + " is not Q8_0"; return false; }
    const int64_t n = t.n_elements();
    if (t.ne.empty() || t.ne[0] % 32) { err = "Q8_0 tensor with a row length not a multiple of 32: " + t.name; return false; }
    const size_t bytes = (size_t)(n / 32) * 34;
    if (data_start_ + t.offset + bytes > map_size_) { err = "tensor data out of file bounds: " + t.name; return false; }
    const uint8_t* src = static_cast<const uint8_t*>(map_) + data_start_ + t.offset;
    q.resize((size_t)n); d.resize((size_t)(n / 32));
    for (int64_t b = 0; b < n / 32; ++b) {                 // { f16 d; int8 q[32]; }
        std::memcpy(&d[b], src + b * 34, 2);
        std::memcpy(&q[b * 32], src + b * 34 + 2, 32);
    }
    return true;
}

}  // namespace q3a
