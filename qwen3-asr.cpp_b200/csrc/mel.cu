// Log-mel frontend for sm_100a: reflect-pad, periodic Hann window, 400-point real FFT in registers and
// shared memory, power spectrum, sparse mel filterbank, log10, and (second pass) the utterance-wide
// clamp(max-8) and (x+4)/4 scale.  Replaces log_mel_spectrogram (reference mel_spectrogram.cpp:484-628,
// whose non-Apple path is a double-precision O(N^2) DFT, :569-600) and generate_mel_filters (:361-415).
//
// Kernel 1 (mel_frames_kernel): one CTA = one tile of 32 consecutive frames of one clip, 288 threads.
//   load     5360 padded samples -> smem (float4, coalesced; reflect/zero rule of :499-515 at the edges)
//   stage 1  thread = (frame mod 11, n1): window and W400 twiddles of its n1 live in registers for its
//            three frames; 16 windowed samples -> rdft16 -> twiddle -> smem z[frame][k2][n1]
//   stage 2  warp = k2, lane = frame: 25 complex -> dft25 -> |X|^2 -> smem power[bin][frame]
//            (frame fastest, so stage 3 reads four frames of one bin with one LDS.128)
//   stage 3  thread = (mel row, 4 frames): compressed filterbank row . power -> log10(max(.,1e-10))
//            -> global [mel][frame] as float4, and the running per-clip maximum over frames < n_len
//            (:604-613) via one atomicMax per CTA.
// Kernel 2 (mel_finalize_kernel): x = (max(x, clipmax-8)+4)/4 in place (:615-625), float4.
// Algorithmic HBM bytes per clip: 4*n_samples (PCM read once) + 4*n_mel*n_len (mel written once).
#include "mel.h"
#include "mel_fft.cuh"
#include "tc.cuh"

#include <cmath>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace q3a {

constexpr int kNFft = 400;
constexpr int kHop = 160;
constexpr int kPad = kNFft / 2;
constexpr int kBins = 201;
constexpr int kFT = 32;                                  // frames per CTA
constexpr int kMelThreads = 288;                         // 9 warps: 288 = 9 k2 * 32 frames
constexpr int kTileSamples = (kFT - 1) * kHop + kNFft;   // 5360
constexpr int kUnionFloats = kFT * kBins;                // 6432 >= kTileSamples
static_assert(kUnionFloats >= kTileSamples, "power buffer aliases the sample buffer");
constexpr int kS1Frames = 11;                            // frames stage 1 covers per sweep (11 * 25 = 275 threads)
constexpr int kMaxTabWeights = 512, kMaxTabRows = 128;   // filterbank tables staged in smem up to this size

struct MelSmem {
    float u[2][kUnionFloats];         // per buffer: samples (load, stage 1) then power[bin][frame] (stage 2, 3);
                                      // the other buffer receives the next tile's samples meanwhile
    float2 z[kFT][9][25];             // stage-1 output
    float w[kMaxTabWeights];          // compressed filterbank rows
    int bin0[kMaxTabRows], ptr[kMaxTabRows + 1];
    float red[16];
    unsigned long long bar[2];        // mbarriers of the two sample buffers
};

__device__ __forceinline__ unsigned ordered_from_float(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float float_from_ordered(unsigned u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

// sample at padded position p of a clip (mel_spectrogram.cpp:493-515)
__device__ __forceinline__ float padded_sample(const float* __restrict__ pcm, int n, int p) {
    int src;
    if (p < kPad) { src = kPad - p; if (src >= n) return 0.0f; }
    else if (p < kPad + n) src = p - kPad;
    else { src = n - 2 - (p - kPad - n); if (src < 0) return 0.0f; }
    return __ldg(pcm + src);
}

struct MelTile {            // where tile `tile` lives (warp-uniform arithmetic, every thread computes it)
    int clip, f0, n_valid, n, n_len;
    long long pcm_off, mel_off;
    bool interior;          // its 5360 samples are one contiguous, 16-byte aligned run of the clip
};

__device__ __forceinline__ MelTile mel_tile_of(const MelClipDesc* __restrict__ clips, int n_clips, int tile) {
    int lo = 0, hi = n_clips - 1;                        // tile0 is ascending; n_clips is small
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(&clips[mid].tile0) <= tile) lo = mid; else hi = mid - 1;
    }
    const MelClipDesc cd = clips[lo];
    MelTile t;
    t.clip = lo; t.f0 = (tile - cd.tile0) * kFT; t.n = cd.n_samples; t.n_len = cd.n_len;
    t.n_valid = min(kFT, cd.n_len - t.f0);
    t.pcm_off = cd.pcm_off; t.mel_off = cd.mel_off;
    const int p0 = t.f0 * kHop;
    t.interior = p0 >= kPad && p0 + kTileSamples <= kPad + t.n;
    return t;
}

// Persistent: a CTA walks tiles blockIdx.x, +gridDim.x, ...  While it computes tile i out of one sample buffer, a
// bulk async copy (cp.async.bulk + mbarrier) already fills the other buffer with tile i+1, so the DRAM latency of
// the PCM read is off the critical path; clip-edge tiles (reflect / zero padding) are loaded by the threads.
__global__ void __launch_bounds__(kMelThreads, 2)
mel_frames_kernel(const float* __restrict__ pcm_base, const MelClipDesc* __restrict__ clips, int n_clips, int n_tiles,
                  const float* __restrict__ g_hann, const float2* __restrict__ g_tw,
                  const int* __restrict__ row_bin0, const int* __restrict__ row_ptr,
                  const float* __restrict__ weights, int n_mel, float* __restrict__ mel_base,
                  unsigned* __restrict__ clip_max) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    MelSmem& s = *reinterpret_cast<MelSmem*>(smem_raw);
    const int tid = threadIdx.x;
    pdl_trigger();

    // ---- once per CTA: barriers, filterbank tables, this thread's window and twiddles ------------------
    const uint32_t bar0 = tc::smem_u32(&s.bar[0]), bar1 = tc::smem_u32(&s.bar[1]);
    if (tid == 0) { tc::mbar_init(bar0, 1); tc::mbar_init(bar1, 1); tc::fence_barrier_init(); }
    // filterbank tables: smem when they fit (the standard 128 x 201 bank has 396 non-zeros), else global
    const bool tab_smem = n_mel <= kMaxTabRows && __ldg(row_ptr + n_mel) <= kMaxTabWeights;
    if (tab_smem) {
        const int nw = __ldg(row_ptr + n_mel);
        for (int i = tid; i < nw; i += kMelThreads) s.w[i] = __ldg(weights + i);
        for (int i = tid; i < n_mel; i += kMelThreads) s.bin0[i] = __ldg(row_bin0 + i);
        for (int i = tid; i <= n_mel; i += kMelThreads) s.ptr[i] = __ldg(row_ptr + i);
    }
    const float* wtab = tab_smem ? s.w : weights;
    const int* b0tab = tab_smem ? s.bin0 : row_bin0;
    const int* ptab = tab_smem ? s.ptr : row_ptr;
    const bool s1 = tid < kS1Frames * 25;
    const int fi = tid / 25, n1 = tid - fi * 25;
    float hw[16];
    float2 tw[9];
    if (s1) {
#pragma unroll
        for (int n2 = 0; n2 < 16; ++n2) hw[n2] = __ldg(g_hann + n1 + 25 * n2);
#pragma unroll
        for (int k2 = 1; k2 < 9; ++k2) tw[k2] = __ldg(g_tw + n1 * 9 + k2);
    }
    pdl_wait();                                           // PCM / descriptors may come from the previous kernel
    __syncthreads();

    auto prefetch = [&](const MelTile& t, int buf) {     // tid 0 only; t.interior
        const uint32_t bar = buf ? bar1 : bar0;
        tc::fence_proxy_async();                          // the buffer was last touched through the generic proxy
        tc::mbar_arrive_expect_tx(bar, kTileSamples * 4);
        const float* src = pcm_base + t.pcm_off + (t.f0 * kHop - kPad);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(tc::smem_u32(s.u[buf])), "l"(src), "r"(kTileSamples * 4), "r"(bar) : "memory");
    };

    int tile = blockIdx.x;
    if (tile >= n_tiles) return;
    MelTile t = mel_tile_of(clips, n_clips, tile);
    if (tid == 0 && t.interior) prefetch(t, 0);
    uint32_t ph0 = 0, ph1 = 0;

    for (int it = 0; tile < n_tiles; ++it) {
        const int cur = it & 1;
        float* u = s.u[cur];
        const int f0 = t.f0, n_valid = t.n_valid;
        const int next_tile = tile + gridDim.x;
        MelTile tn = t;
        if (next_tile < n_tiles) {
            tn = mel_tile_of(clips, n_clips, next_tile);
            if (tid == 0 && tn.interior) prefetch(tn, cur ^ 1);
        }
        // ---- this tile's samples ---------------------------------------------------------------------
        if (t.interior) {
            if (cur) { tc::mbar_wait(bar1, ph1); ph1 ^= 1; } else { tc::mbar_wait(bar0, ph0); ph0 ^= 1; }
        } else {
            const float* __restrict__ pcm = pcm_base + t.pcm_off;
            const int p0 = f0 * kHop;
            const int need = (n_valid - 1) * kHop + kNFft;        // samples the valid frames touch
            for (int i = tid; i < kTileSamples; i += kMelThreads)
                u[i] = (i < need) ? padded_sample(pcm, t.n, p0 + i) : 0.0f;
            __syncthreads();
        }

        // ---- stage 1: 25 real 16-point DFTs per frame + twiddle -----------------------------------
        if (s1) {
#pragma unroll 1
            for (int f = fi; f < n_valid; f += kS1Frames) {
                float x[16];
                const float* xs = u + f * kHop + n1;
#pragma unroll
                for (int n2 = 0; n2 < 16; ++n2) x[n2] = xs[25 * n2] * hw[n2];
                fft::cpx y[9];
                fft::rdft16(x, y);
                s.z[f][0][n1] = make_float2(y[0].re, y[0].im);      // W400^0 = 1
#pragma unroll
                for (int k2 = 1; k2 < 9; ++k2) {
                    fft::cpx zz = fft::cmul(y[k2], fft::mk(tw[k2].x, tw[k2].y));
                    s.z[f][k2][n1] = make_float2(zz.re, zz.im);
                }
            }
        }
        __syncthreads();

        // ---- stage 2: 9 complex 25-point DFTs per frame -> power spectrum, stored [bin][frame] -------
        {
            const int f = tid & 31, k2 = tid >> 5;            // warp = k2, lane = frame: both smem sides conflict-free
            if (f < n_valid) {
                fft::cpx a[25];
#pragma unroll
                for (int i = 0; i < 25; ++i) { float2 v = s.z[f][k2][i]; a[i] = fft::mk(v.x, v.y); }
                fft::dft25(a);
                // bin k = k2 + 16 k1 is direct for k1 <= 12 (k <= 200) and the mirror 400-k otherwise.  Mirrors go
                // first: for k2 = 0 and 8 they land on bins this same thread also computes directly, and the
                // direct value must win (program order within the thread); for other k2 they are unique.
                float* pw = u + f;
#pragma unroll
                for (int k1 = 13; k1 < 25; ++k1)
                    pw[(400 - 16 * k1 - k2) * kFT] = a[k1].re * a[k1].re + a[k1].im * a[k1].im;
#pragma unroll
                for (int k1 = 0; k1 < 13; ++k1)
                    pw[(16 * k1 + k2) * kFT] = a[k1].re * a[k1].re + a[k1].im * a[k1].im;
            }
        }
        __syncthreads();

        // ---- stage 3: filterbank, log10, store, running max ----------------------------------------
        float vmax = -INFINITY;
        {
            const int q = tid & 7;                            // frames 4q .. 4q+3 of the tile
            float* __restrict__ out = mel_base + t.mel_off;
            const bool vec_ok = ((t.n_len & 3) | (int)(reinterpret_cast<uintptr_t>(out) & 15)) == 0;   // rows stay 16-byte aligned
            const float4* pw4 = reinterpret_cast<const float4*>(u) + q;
            if (4 * q < n_valid) {
                for (int m = tid >> 3; m < n_mel; m += kMelThreads / 8) {
                    const int b0 = b0tab[m], w0 = ptab[m], w1 = ptab[m + 1];
                    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                    const float4* pp = pw4 + b0 * (kFT / 4);
#pragma unroll 4
                    for (int j = w0; j < w1; ++j) {
                        const float w = wtab[j];
                        const float4 p = pp[(j - w0) * (kFT / 4)];
                        acc.x = fmaf(w, p.x, acc.x); acc.y = fmaf(w, p.y, acc.y);
                        acc.z = fmaf(w, p.z, acc.z); acc.w = fmaf(w, p.w, acc.w);
                    }
                    const float kLog10Of2 = 0.30102999566398120f;
                    float4 v;
                    v.x = __log2f(fmaxf(acc.x, 1e-10f)) * kLog10Of2; v.y = __log2f(fmaxf(acc.y, 1e-10f)) * kLog10Of2;
                    v.z = __log2f(fmaxf(acc.z, 1e-10f)) * kLog10Of2; v.w = __log2f(fmaxf(acc.w, 1e-10f)) * kLog10Of2;
                    float* dst = out + (long long)m * t.n_len + f0 + 4 * q;
                    const int left = n_valid - 4 * q;         // >= 1
                    if (left >= 4 && vec_ok) {
                        *reinterpret_cast<float4*>(dst) = v;
                        vmax = fmaxf(vmax, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
                    } else {
                        dst[0] = v.x; vmax = fmaxf(vmax, v.x);
                        if (left > 1) { dst[1] = v.y; vmax = fmaxf(vmax, v.y); }
                        if (left > 2) { dst[2] = v.z; vmax = fmaxf(vmax, v.z); }
                        if (left > 3) { dst[3] = v.w; vmax = fmaxf(vmax, v.w); }
                    }
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
        if ((tid & 31) == 0) s.red[tid >> 5] = vmax;
        __syncthreads();                                  // also: every thread is done with u (power) and z
        if (tid == 0) {
            float v = s.red[0];
            for (int w = 1; w < kMelThreads / 32; ++w) v = fmaxf(v, s.red[w]);
            if (v > -INFINITY) atomicMax(clip_max + t.clip, ordered_from_float(v));
        }
        tile = next_tile;
        t = tn;
    }
}

__global__ void mel_max_decode_kernel(const unsigned* __restrict__ enc, float* __restrict__ out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = float_from_ordered(enc[i]);
}

// x = (max(x, max-8) + 4) / 4, in place.  One grid-stride pass over each clip's block.
__global__ void __launch_bounds__(256)
mel_finalize_kernel(float* __restrict__ mel_base, const MelClipDesc* __restrict__ clips,
                    const unsigned* __restrict__ clip_max, int n_mel, int blocks_per_clip) {
    const int c = blockIdx.x / blocks_per_clip, b = blockIdx.x - c * blocks_per_clip;
    const MelClipDesc cd = clips[c];
    const long long total = (long long)n_mel * cd.n_len;
    if (total == 0) return;
    const float thr = float_from_ordered(clip_max[c]) - 8.0f;
    float* __restrict__ p = mel_base + cd.mel_off;
    const long long stride = (long long)blocks_per_clip * blockDim.x;
    if ((cd.mel_off & 3) == 0) {
        float4* p4 = reinterpret_cast<float4*>(p);
        const long long n4 = total >> 2;
        for (long long i = (long long)b * blockDim.x + threadIdx.x; i < n4; i += stride) {
            float4 v = p4[i];
            v.x = (fmaxf(v.x, thr) + 4.0f) * 0.25f; v.y = (fmaxf(v.y, thr) + 4.0f) * 0.25f;
            v.z = (fmaxf(v.z, thr) + 4.0f) * 0.25f; v.w = (fmaxf(v.w, thr) + 4.0f) * 0.25f;
            p4[i] = v;
        }
        for (long long i = (n4 << 2) + (long long)b * blockDim.x + threadIdx.x; i < total; i += stride)
            p[i] = (fmaxf(p[i], thr) + 4.0f) * 0.25f;
    } else {
        for (long long i = (long long)b * blockDim.x + threadIdx.x; i < total; i += stride)
            p[i] = (fmaxf(p[i], thr) + 4.0f) * 0.25f;
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------

// Triangular HTK-scale, Slaney-normalised filterbank in single precision, the way the reference's
// generator evaluates it (mel_spectrogram.cpp:353-415): bin position (n_fft+1)*hz/sr, both triangle
// sides inclusive, rows narrower than one bin come out all-zero.
static void make_filterbank(float* out, int n_mels, int n_fft, int sample_rate) {
    const int n_bins = 1 + n_fft / 2;
    auto to_mel = [](float hz) { return 2595.0f * log10f(1.0f + hz / 700.0f); };
    auto to_hz = [](float mel) { return 700.0f * (powf(10.0f, mel / 2595.0f) - 1.0f); };
    const float lo = to_mel(0.0f), hi = to_mel(sample_rate / 2.0f);
    std::vector<float> hz(n_mels + 2), pos(n_mels + 2);
    for (int i = 0; i < n_mels + 2; ++i) {
        float mel = lo + (hi - lo) * i / (n_mels + 1);
        hz[i] = to_hz(mel);
        pos[i] = (n_fft + 1) * hz[i] / sample_rate;
    }
    for (int m = 0; m < n_mels; ++m) {
        const float l = pos[m], c = pos[m + 1], r = pos[m + 2];
        const float scale = 2.0f / (hz[m + 2] - hz[m]);
        float* row = out + (size_t)m * n_bins;
        for (int k = 0; k < n_bins; ++k) {
            float w = 0.0f;
            if (k >= l && k <= c) w = (k - l) / (c - l);
            else if (k >= c && k <= r) w = (r - k) / (r - c);
            row[k] = w * scale;
        }
    }
}

int mel_launch(q3a_frontend* fe, const float* pcm_dev, const int64_t* pcm_off, const int* n_samples, int n_clips,
               float* mel_dev, const int64_t* mel_off, float* max_dev, int finalize, cudaStream_t stream) {
    if (n_clips == 0) return Q3A_OK;
    // descriptors + max slots: [MelClipDesc x n][unsigned x n]
    const size_t desc_bytes = sizeof(MelClipDesc) * n_clips;
    void* hraw = nullptr;
    Q3A_TRY(fe->ring.acquire(desc_bytes, &hraw));
    MelClipDesc* h = reinterpret_cast<MelClipDesc*>(hraw);
    int tiles = 0;
    int max_len = 0;
    for (int b = 0; b < n_clips; ++b) {
        if (n_samples[b] < 0 || (pcm_off[b] & 3)) { set_error("mel: negative n_samples or pcm offset not a multiple of 4"); return Q3A_ERR_ARG; }
        h[b].pcm_off = pcm_off[b];
        h[b].mel_off = mel_off[b];
        h[b].n_samples = n_samples[b];
        h[b].n_len = n_samples[b] / kHop;
        h[b].tile0 = tiles;
        h[b].pad = 0;
        tiles += ceil_div(h[b].n_len, kFT);
        if (h[b].n_len > max_len) max_len = h[b].n_len;
    }
    Q3A_TRY(fe->desc.reserve(desc_bytes));
    Q3A_TRY(fe->maxbuf.reserve(sizeof(unsigned) * n_clips));
    Q3A_CUDA(cudaMemcpyAsync(fe->desc.p, h, desc_bytes, cudaMemcpyHostToDevice, stream));
    Q3A_TRY(fe->ring.mark(stream));
    Q3A_CUDA(cudaMemsetAsync(fe->maxbuf.p, 0, sizeof(unsigned) * n_clips, stream));
    const MelClipDesc* d_desc = fe->desc.as<MelClipDesc>();
    unsigned* d_max = fe->maxbuf.as<unsigned>();
    if (tiles > 0) {
        const int grid = tiles < 2 * fe->n_sms ? tiles : 2 * fe->n_sms;     // persistent: two CTAs per SM
        Q3A_CUDA(launch_pdl(mel_frames_kernel, dim3(grid), dim3(kMelThreads), sizeof(MelSmem), stream,
                            pcm_dev, d_desc, n_clips, tiles, (const float*)fe->d_hann, (const float2*)fe->d_twiddle,
                            (const int*)fe->d_row_bin0, (const int*)fe->d_row_ptr, (const float*)fe->d_weights, fe->n_mel,
                            mel_dev, d_max));
        ++g_launch_counter;
        Q3A_CUDA(cudaGetLastError());
    }
    if (max_dev) {
        mel_max_decode_kernel<<<ceil_div(n_clips, 128), 128, 0, stream>>>(d_max, max_dev, n_clips);
        ++g_launch_counter;
        Q3A_CUDA(cudaGetLastError());
    }
    if (finalize && tiles > 0) {
        // enough CTAs per clip to cover the longest clip with ~8 float4 per thread, grid a multiple of clips
        int bpc = ceil_div(ceil_div(fe->n_mel * max_len, 4), 256 * 8);
        if (bpc < 1) bpc = 1;
        mel_finalize_kernel<<<n_clips * bpc, 256, 0, stream>>>(mel_dev, d_desc, d_max, fe->n_mel, bpc);
        ++g_launch_counter;
        Q3A_CUDA(cudaGetLastError());
    }
    return Q3A_OK;
}

}  // namespace q3a

using namespace q3a;

extern "C" int q3a_generate_mel_filters(float* out, int n_mels, int n_fft, int sample_rate) {
    if (!out || n_mels <= 0 || n_fft <= 0 || sample_rate <= 0) { set_error("generate_mel_filters: bad argument"); return Q3A_ERR_ARG; }
    make_filterbank(out, n_mels, n_fft, sample_rate);
    return Q3A_OK;
}

extern "C" int q3a_mel_n_len(int n_samples) { return n_samples < 0 ? 0 : n_samples / kHop; }

extern "C" void q3a_frontend_free(q3a_frontend* fe) {
    if (!fe) return;
    cudaSetDevice(fe->device);
    cudaFree(fe->d_hann); cudaFree(fe->d_twiddle); cudaFree(fe->d_row_bin0); cudaFree(fe->d_row_ptr);
    cudaFree(fe->d_weights);
    fe->pcm.release(); fe->mel.release(); fe->desc.release(); fe->maxbuf.release();
    fe->ring.release();
    if (fe->stream) cudaStreamDestroy(fe->stream);
    delete fe;
}

extern "C" int q3a_frontend_create(int device, const float* filters, int n_mel, int n_bins, q3a_frontend** out) {
    if (!out) { set_error("frontend_create: out is null"); return Q3A_ERR_ARG; }
    *out = nullptr;
    if (filters == nullptr) { n_mel = 128; n_bins = kBins; }
    if (n_mel <= 0 || n_mel > 4096 || n_bins != kBins) {
        set_error("frontend_create: filters must be [n_mel][201] (n_fft = 400)");
        return Q3A_ERR_ARG;
    }
    Q3A_TRY(use_device(device));
    std::vector<float> bank((size_t)n_mel * n_bins);
    if (filters) std::memcpy(bank.data(), filters, bank.size() * sizeof(float));
    else make_filterbank(bank.data(), n_mel, kNFft, 16000);

    // compress each row to its non-zero span
    std::vector<int> bin0(n_mel), ptr(n_mel + 1);
    std::vector<float> w;
    for (int m = 0; m < n_mel; ++m) {
        const float* row = bank.data() + (size_t)m * n_bins;
        int a = 0, b = n_bins;
        while (a < n_bins && row[a] == 0.0f) ++a;
        while (b > a && row[b - 1] == 0.0f) --b;
        bin0[m] = (a < n_bins) ? a : 0;
        ptr[m] = (int)w.size();
        for (int k = a; k < b; ++k) w.push_back(row[k]);
    }
    ptr[n_mel] = (int)w.size();
    if (w.empty()) w.push_back(0.0f);

    std::vector<float> hann(kNFft);
    for (int i = 0; i < kNFft; ++i) hann[i] = (float)(0.5 * (1.0 - std::cos(2.0 * M_PI * i / kNFft)));
    std::vector<float2> tw(25 * 9);
    for (int n1 = 0; n1 < 25; ++n1)
        for (int k2 = 0; k2 < 9; ++k2) {
            double ang = -2.0 * M_PI * (double)(n1 * k2) / 400.0;
            tw[n1 * 9 + k2] = make_float2((float)std::cos(ang), (float)std::sin(ang));
        }

    q3a_frontend* fe = new q3a_frontend();
    fe->device = device; fe->n_mel = n_mel; fe->n_bins = n_bins;
    if (cudaDeviceGetAttribute(&fe->n_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || fe->n_sms <= 0) fe->n_sms = 148;
    auto fail = [&](int rc) { q3a_frontend_free(fe); return rc; };
#define FE_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { set_error(std::string(#x) + ": " + cudaGetErrorString(e_)); return fail(Q3A_ERR_CUDA); } } while (0)
    FE_CUDA(cudaMalloc(&fe->d_hann, sizeof(float) * kNFft));
    FE_CUDA(cudaMalloc(&fe->d_twiddle, sizeof(float2) * tw.size()));
    FE_CUDA(cudaMalloc(&fe->d_row_bin0, sizeof(int) * n_mel));
    FE_CUDA(cudaMalloc(&fe->d_row_ptr, sizeof(int) * (n_mel + 1)));
    FE_CUDA(cudaMalloc(&fe->d_weights, sizeof(float) * w.size()));
    FE_CUDA(cudaMemcpy(fe->d_hann, hann.data(), sizeof(float) * kNFft, cudaMemcpyHostToDevice));
    FE_CUDA(cudaMemcpy(fe->d_twiddle, tw.data(), sizeof(float2) * tw.size(), cudaMemcpyHostToDevice));
    FE_CUDA(cudaMemcpy(fe->d_row_bin0, bin0.data(), sizeof(int) * n_mel, cudaMemcpyHostToDevice));
    FE_CUDA(cudaMemcpy(fe->d_row_ptr, ptr.data(), sizeof(int) * (n_mel + 1), cudaMemcpyHostToDevice));
    FE_CUDA(cudaMemcpy(fe->d_weights, w.data(), sizeof(float) * w.size(), cudaMemcpyHostToDevice));
    FE_CUDA(cudaStreamCreateWithFlags(&fe->stream, cudaStreamNonBlocking));
    FE_CUDA(cudaFuncSetAttribute(mel_frames_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MelSmem)));
#undef FE_CUDA
    *out = fe;
    return Q3A_OK;
}

extern "C" int q3a_mel_batch_device(q3a_frontend* fe, const float* pcm_dev, const int64_t* pcm_off,
                                    const int* n_samples, int n_clips, float* mel_dev, const int64_t* mel_off,
                                    float* max_dev, int finalize, void* stream) {
    if (!fe) { set_error("mel_batch_device: frontend is null"); return Q3A_ERR_STATE; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm_dev || !pcm_off || !n_samples || !mel_dev || !mel_off))) {
        set_error("mel_batch_device: null argument"); return Q3A_ERR_ARG;
    }
    Q3A_TRY(use_device(fe->device));
    g_launch_counter = 0;
    return mel_launch(fe, pcm_dev, pcm_off, n_samples, n_clips, mel_dev, mel_off, max_dev, finalize,
                      reinterpret_cast<cudaStream_t>(stream));
}

// Packing used by every host-pointer entry point: clip b starts at a multiple of 64 samples.
int q3a_pack_offsets(const int* n, int count, int unit, int64_t align, std::vector<int64_t>& off) {
    off.resize(count + 1);
    int64_t cur = 0;
    for (int b = 0; b < count; ++b) { off[b] = cur; cur += round_up64((int64_t)n[b] * unit, align); }
    off[count] = cur;
    return Q3A_OK;
}

extern "C" int q3a_mel_batch(q3a_frontend* fe, const float* const* pcm, const int* n_samples, int n_clips,
                             float* const* out, int* n_len) {
    if (!fe) { set_error("mel_batch: frontend is null"); return Q3A_ERR_STATE; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm || !n_samples || !out))) { set_error("mel_batch: null argument"); return Q3A_ERR_ARG; }
    for (int b = 0; b < n_clips; ++b)
        if (n_samples[b] < 0 || (n_samples[b] > 0 && !pcm[b]) || !out[b]) { set_error("mel_batch: bad clip " + std::to_string(b)); return Q3A_ERR_ARG; }
    Q3A_TRY(use_device(fe->device));
    g_launch_counter = 0;
    std::vector<int64_t> poff, moff;
    std::vector<int> mel_elems(n_clips);
    for (int b = 0; b < n_clips; ++b) mel_elems[b] = fe->n_mel * (n_samples[b] / kHop);
    q3a_pack_offsets(n_samples, n_clips, 1, 64, poff);
    q3a_pack_offsets(mel_elems.data(), n_clips, 1, 64, moff);
    Q3A_TRY(fe->pcm.reserve(sizeof(float) * (size_t)(poff[n_clips] + 64)));
    Q3A_TRY(fe->mel.reserve(sizeof(float) * (size_t)(moff[n_clips] + 64)));
    cudaStream_t st = fe->stream;
    for (int b = 0; b < n_clips; ++b)
        if (n_samples[b] > 0)
            Q3A_CUDA(cudaMemcpyAsync(fe->pcm.as<float>() + poff[b], pcm[b], sizeof(float) * n_samples[b],
                                     cudaMemcpyHostToDevice, st));
    Q3A_TRY(mel_launch(fe, fe->pcm.as<float>(), poff.data(), n_samples, n_clips, fe->mel.as<float>(), moff.data(),
                       nullptr, 1, st));
    for (int b = 0; b < n_clips; ++b) {
        if (n_len) n_len[b] = n_samples[b] / kHop;
        if (mel_elems[b] > 0)
            Q3A_CUDA(cudaMemcpyAsync(out[b], fe->mel.as<float>() + moff[b], sizeof(float) * mel_elems[b],
                                     cudaMemcpyDeviceToHost, st));
    }
    Q3A_CUDA(cudaStreamSynchronize(st));
    return Q3A_OK;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of q3a_mel_batch_device (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).
//
// mel_launch keeps two pieces of scratch in grow-only device buffers: the clip descriptor table and the per-clip maximum
// slots that mel_frames_kernel's atomicMax needs zeroed.  Here both live in one file-scope __device__ array of a fixed
// size, so the call allocates nothing -- no cudaMalloc / cudaFree, no pinned ring, no host heap either: the clip table
// travels as a kernel argument (3 KB) to a set-up kernel that writes the descriptors and clears the maxima in place.
// Every GPU that loads the library holds the array (4.5 KB) until the process exits, every call on a device shares it and
// it keeps the last call's contents, so two calls that overlap on one GPU (different streams, or two host threads)
// race; calls in one stream are ordered and safe.
// ---------------------------------------------------------------------------------------------
namespace q3a {

constexpr int kNoAllocMaxClips = 128;                    // the stated limit of the entry point
struct MelNoAllocScratch {
    MelClipDesc desc[kNoAllocMaxClips];                  // 128 x 32 B
    unsigned clip_max[kNoAllocMaxClips];                 // 128 x 4 B, cleared by every call
};
static __device__ __align__(256) unsigned char g_mel_noalloc_scratch[sizeof(MelNoAllocScratch)];

struct MelNoAllocClips {                                 // by-value kernel argument: 128 x 24 B = 3 KB
    long long pcm_off[kNoAllocMaxClips];
    long long mel_off[kNoAllocMaxClips];
    int n_samples[kNoAllocMaxClips];
    int tile0[kNoAllocMaxClips];
};

// device code names the array directly; only the host needs cudaGetSymbolAddress
__global__ void mel_noalloc_setup_kernel(const __grid_constant__ MelNoAllocClips c, int n_clips) {
    MelNoAllocScratch& sc = *reinterpret_cast<MelNoAllocScratch*>(g_mel_noalloc_scratch);
    const int i = threadIdx.x;
    if (i >= kNoAllocMaxClips) return;
    if (i < n_clips) {
        MelClipDesc d;
        d.pcm_off = c.pcm_off[i]; d.mel_off = c.mel_off[i];
        d.n_samples = c.n_samples[i]; d.n_len = c.n_samples[i] / kHop;
        d.tile0 = c.tile0[i]; d.pad = 0;
        sc.desc[i] = d;
    }
    sc.clip_max[i] = 0u;                                 // what must start at zero: the atomicMax slots (all of them)
}

// The array's name in host code is a host-side stand-in, not a device address: resolve it once per device, with that
// device current, and hand the kernels the resolved pointer.
static int mel_noalloc_scratch(int device, MelNoAllocScratch** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static MelNoAllocScratch* ptr[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) { set_error("mel_batch_device_noalloc: device index out of range"); return Q3A_ERR_ARG; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        Q3A_CUDA(cudaGetSymbolAddress(&p, g_mel_noalloc_scratch));    // the caller made `device` current
        if (!p) { set_error("mel_batch_device_noalloc: cudaGetSymbolAddress returned null"); return Q3A_ERR_CUDA; }
        ptr[device] = static_cast<MelNoAllocScratch*>(p);
    }
    *out = ptr[device];
    return Q3A_OK;
}

}  // namespace q3a

extern "C" int q3a_mel_noalloc_max_clips(void) { return kNoAllocMaxClips; }

extern "C" int q3a_mel_batch_device_noalloc(q3a_frontend* fe, const float* pcm_dev, const int64_t* pcm_off,
                                            const int* n_samples, int n_clips, float* mel_dev, const int64_t* mel_off,
                                            float* max_dev, int finalize, void* stream) {
    if (!fe) { set_error("mel_batch_device_noalloc: frontend is null"); return Q3A_ERR_STATE; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm_dev || !pcm_off || !n_samples || !mel_dev || !mel_off))) {
        set_error("mel_batch_device_noalloc: null argument"); return Q3A_ERR_ARG;
    }
    if (n_clips > kNoAllocMaxClips) {
        set_error("mel_batch_device_noalloc: n_clips = " + std::to_string(n_clips) + " exceeds the static scratch's maximum of " +
                  std::to_string(kNoAllocMaxClips) + " clips per call (use q3a_mel_batch_device, or split the batch)");
        return Q3A_ERR_ARG;
    }
    if (n_clips == 0) return Q3A_OK;
    Q3A_TRY(use_device(fe->device));
    g_launch_counter = 0;
    MelNoAllocClips c;                                   // on the stack: no host allocation either
    int tiles = 0, max_len = 0;
    for (int b = 0; b < n_clips; ++b) {
        if (n_samples[b] < 0 || (pcm_off[b] & 3)) { set_error("mel: negative n_samples or pcm offset not a multiple of 4"); return Q3A_ERR_ARG; }
        const int n_len = n_samples[b] / kHop;
        c.pcm_off[b] = pcm_off[b]; c.mel_off[b] = mel_off[b]; c.n_samples[b] = n_samples[b]; c.tile0[b] = tiles;
        tiles += ceil_div(n_len, kFT);
        if (n_len > max_len) max_len = n_len;
    }
    for (int b = n_clips; b < kNoAllocMaxClips; ++b) { c.pcm_off[b] = 0; c.mel_off[b] = 0; c.n_samples[b] = 0; c.tile0[b] = tiles; }
    MelNoAllocScratch* sc = nullptr;
    Q3A_TRY(mel_noalloc_scratch(fe->device, &sc));
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    mel_noalloc_setup_kernel<<<1, kNoAllocMaxClips, 0, st>>>(c, n_clips);
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    if (tiles > 0) {
        const int grid = tiles < 2 * fe->n_sms ? tiles : 2 * fe->n_sms;
        Q3A_CUDA(launch_pdl(mel_frames_kernel, dim3(grid), dim3(kMelThreads), sizeof(MelSmem), st,
                            pcm_dev, (const MelClipDesc*)sc->desc, n_clips, tiles, (const float*)fe->d_hann,
                            (const float2*)fe->d_twiddle, (const int*)fe->d_row_bin0, (const int*)fe->d_row_ptr,
                            (const float*)fe->d_weights, fe->n_mel, mel_dev, (unsigned*)sc->clip_max));
        ++g_launch_counter;
        Q3A_CUDA(cudaGetLastError());
    }
    if (max_dev) {
        mel_max_decode_kernel<<<ceil_div(n_clips, 128), 128, 0, st>>>(sc->clip_max, max_dev, n_clips);
        ++g_launch_counter;
        Q3A_CUDA(cudaGetLastError());
    }
    if (finalize && tiles > 0) {
        int bpc = ceil_div(ceil_div(fe->n_mel * max_len, 4), 256 * 8);
        if (bpc < 1) bpc = 1;
        mel_finalize_kernel<<<n_clips * bpc, 256, 0, st>>>(mel_dev, sc->desc, sc->clip_max, fe->n_mel, bpc);
        ++g_launch_counter;
        Q3A_CUDA(cudaGetLastError());
    }
    return Q3A_OK;
}
