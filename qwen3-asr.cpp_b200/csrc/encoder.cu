// Audio-encoder forward for sm_100a: weight loading from GGUF and the launch sequence that replaces
// AudioEncoder::encode / encode_no_chunk / encode_conv_only (reference audio_encoder.cpp:312-913) and
// ForcedAligner::encode_audio (forced_aligner.cpp:591-924) for a packed batch of utterances.
//
//   stem       conv1 (CUDA cores) -> planes ; conv2, conv3 implicit GEMM (tcgen05) ; conv_out GEMM + PE
//   18/24 x    LN -> fused QKV GEMM(+bias) -> varlen attention -> out GEMM(+bias,+residual)
//              LN -> up GEMM(+bias, GELU) -> down GEMM(+bias,+residual)
//   post       LN -> proj1 GEMM(+bias, GELU) -> proj2 GEMM(+bias) -> [T][out_dim] f32
// The residual stream, LayerNorm statistics and softmax are fp32; tensor-core operands are 16-bit.
#include "encoder.h"
#include "gemm.h"
#include "gguf_reader.h"
#include "mel.h"
#include "weights.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <numeric>

namespace q3a {

constexpr int kChunkFrames = 100;        // n_window * 2 (audio_encoder.cpp:327-328)
constexpr int kWindowTokens = 104;       // 13 * (800 / 100) (forced_aligner.cpp:737-741)
constexpr int kMaxChunksPerPass = 4096;  // bounds the stem workspace of the host-pointer entry points (~17 GB)

int encoder_n_tokens(int n_frames, int mode) {
    if (n_frames <= 0) return 0;
    if (mode == Q3A_MODE_NO_CHUNK || mode == Q3A_MODE_CONV_ONLY) return stem_out_len(n_frames);
    int total = (n_frames / kChunkFrames) * stem_out_len(kChunkFrames);
    if (n_frames % kChunkFrames) total += stem_out_len(n_frames % kChunkFrames);
    return total;
}

static void sinusoidal_table(std::vector<float>& pe, int n_ctx, int d) {
    // same float arithmetic as compute_sinusoidal_pe (audio_encoder.cpp:12-22)
    const int half = d / 2;
    pe.assign((size_t)n_ctx * d, 0.0f);
    for (int i = 0; i < half; ++i) {
        const float div_term = expf(-logf(10000.0f) * i / (half - 1));
        for (int pos = 0; pos < n_ctx; ++pos) {
            const float angle = pos * div_term;
            pe[(size_t)pos * d + i] = sinf(angle);
            pe[(size_t)pos * d + half + i] = cosf(angle);
        }
    }
}

static int ensure_pe(q3a_encoder* e, int rows, cudaStream_t st) {
    if (rows <= e->pe_rows) return Q3A_OK;
    rows = std::max(rows, 16);
    std::vector<float> pe;
    sinusoidal_table(pe, rows, e->hp.d_model);
    Q3A_CUDA(cudaStreamSynchronize(st));          // nothing in flight may still read the old table
    Q3A_TRY(e->pe.reserve(pe.size() * 4));
    Q3A_CUDA(cudaMemcpy(e->pe.p, pe.data(), pe.size() * 4, cudaMemcpyHostToDevice));
    e->pe_rows = rows;
    return Q3A_OK;
}

// ---- profiling spans ---------------------------------------------------------------------------------
enum SpanKind { SP_STEM = 0, SP_XFMR = 1, SP_GEMM = 2, SP_ATTN = 3, SP_OTHER = 4, SP_MEL = 5 };

struct Prof {
    q3a_encoder* e; cudaStream_t st;
    int begin(int kind) {
        if (!e->profiling) return -1;
        while ((int)e->ev_pool.size() < e->ev_used + 2) {
            cudaEvent_t ev; if (cudaEventCreate(&ev) != cudaSuccess) return -1;
            e->ev_pool.push_back(ev);
        }
        const int i = e->ev_used; e->ev_used += 2;
        cudaEventRecord(e->ev_pool[i], st);
        e->spans.push_back({kind, i, i + 1});
        return (int)e->spans.size() - 1;
    }
    void end(int span) { if (span >= 0) cudaEventRecord(e->ev_pool[e->spans[span].e1], st); }
};

static void prof_reset(q3a_encoder* e) {
    e->spans.clear(); e->ev_used = 0;
    e->tm = q3a_timings{};
}

static int prof_collect(q3a_encoder* e, cudaStream_t st) {
    e->tm.total_launches = g_launch_counter;
    if (!e->profiling) return Q3A_OK;
    Q3A_CUDA(cudaStreamSynchronize(st));
    for (const auto& s : e->spans) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, e->ev_pool[s.e0], e->ev_pool[s.e1]) != cudaSuccess) { cudaGetLastError(); continue; }
        switch (s.kind) {
            case SP_STEM: e->tm.conv_chunk_ms += ms; break;
            case SP_XFMR: e->tm.transformer_ms += ms; break;
            case SP_GEMM: e->tm.gemm_ms += ms; break;
            case SP_ATTN: e->tm.attention_ms += ms; break;
            case SP_MEL: e->tm.mel_ms += ms; break;
            default: e->tm.other_ms += ms; break;
        }
    }
    return Q3A_OK;
}

// Q3A_STEM_GROUP=<n>: run conv1 -> conv2 in groups of n chunks over a two-slot ring of plane rows (experiment; 0 = off)
static int stem_group_chunks() {
    const char* v = getenv("Q3A_STEM_GROUP");
    const int g = v ? atoi(v) : 0;
    return g > 0 ? g : 0;
}

// ---- the forward pass --------------------------------------------------------------------------------
// Optional partition of one forward pass, used by the host-pointer entry points to overlap PCIe copies with
// kernels.  Inputs: the stem runs part by part (utterances [in_split[i], in_split[i+1])), calling before_stem
// first -- the hook makes the part's input resident (waits for its upload, runs its mel kernels).  Outputs: the
// last `tail_layers` layers and the projection run part by part over [out_split[i], out_split[i+1]), calling
// after_out when a part's rows are final, so its download overlaps the next part's tail.  Utterances are
// independent end to end and the kernels are batch-invariant, so the partition never changes a result.
struct PartHooks {
    std::vector<int> in_split, out_split;
    std::function<int(int part, int u0, int u1)> before_stem;
    std::function<int(int part, int u0, int u1, const int64_t* tok_off)> after_out;
    int tail_layers = 3;
};

template <class OpT>
static int encode_core(q3a_encoder* e, const float* mel_dev, const int64_t* mel_off, const float* max_dev,
                       const int* n_frames, int n_utts, int mode, float* out_dev, int64_t* tok_off_out,
                       cudaStream_t st, const PartHooks* hk) {
    const q3a_hparams& hp = e->hp;
    const int d = hp.d_model, C = hp.conv_channels, H = hp.n_attention_heads, ffn = hp.ffn_dim;
    const int fmt = e->fmt;
    const bool chunked = (mode == Q3A_MODE_ASR || mode == Q3A_MODE_ALIGNER);
    const bool pad_tail = (mode == Q3A_MODE_ALIGNER);
    Prof prof{e, st};

    // ---- host planning: chunks, token rows, attention segments ------------------------------------
    std::vector<StemChunk> chunks;
    std::vector<int> t2_valid, n_tok, tok_base, pos, cu;
    std::vector<int64_t> tok_off(n_utts + 1, 0);
    std::vector<int> utt_chunk0(n_utts + 1, 0), utt_seg0(n_utts + 1, 0);
    int T = 0, max_len = 0, max_pos = 0;
    cu.push_back(0);
    for (int u = 0; u < n_utts; ++u) {
        tok_off[u] = T;
        utt_chunk0[u] = (int)chunks.size();
        utt_seg0[u] = (int)cu.size() - 1;
        const int nf = n_frames[u];
        if (nf < 0) { set_error("encode: negative n_frames"); return Q3A_ERR_ARG; }
        const int step = chunked ? kChunkFrames : std::max(nf, 1);
        const int t_utt0 = T;
        for (int s = 0; s < nf; s += step) {
            const int len = std::min(step, nf - s);
            const int full = pad_tail ? kChunkFrames : len;       // frames the convolutions see
            const int l1 = (full - 1) / 2 + 1, l2 = (l1 - 1) / 2 + 1;
            StemChunk ck;
            ck.mel_off = mel_off[u] + s; ck.row_stride = nf; ck.len = len; ck.t1_valid = l1; ck.utt = u;
            chunks.push_back(ck);
            t2_valid.push_back(l2);
            const int nt = stem_out_len(len);
            n_tok.push_back(nt);
            tok_base.push_back(T);
            for (int t = 0; t < nt; ++t) pos.push_back(t);
            max_pos = std::max(max_pos, nt);
            T += nt;
            max_len = std::max(max_len, full);
        }
        if (mode == Q3A_MODE_ALIGNER) {
            for (int s = t_utt0; s < T; s += kWindowTokens) cu.push_back(std::min(s + kWindowTokens, T));
        } else if (T > t_utt0) {
            cu.push_back(T);
        }
    }
    tok_off[n_utts] = T;
    utt_chunk0[n_utts] = (int)chunks.size();
    utt_seg0[n_utts] = (int)cu.size() - 1;
    if (tok_off_out) std::memcpy(tok_off_out, tok_off.data(), sizeof(int64_t) * (n_utts + 1));
    const int n_chunks = (int)chunks.size();
    // partition (one part each way without hooks); a part that is empty still gets its hook calls
    std::vector<int> in_split = hk && hk->in_split.size() >= 2 ? hk->in_split : std::vector<int>{0, n_utts};
    std::vector<int> out_split = hk && hk->out_split.size() >= 2 ? hk->out_split : std::vector<int>{0, n_utts};
    if (T == 0) {
        if (hk && hk->before_stem)
            for (size_t i = 0; i + 1 < in_split.size(); ++i) Q3A_TRY(hk->before_stem((int)i, in_split[i], in_split[i + 1]));
        if (hk && hk->after_out)
            for (size_t i = 0; i + 1 < out_split.size(); ++i) Q3A_TRY(hk->after_out((int)i, out_split[i], out_split[i + 1], tok_off.data()));
        return Q3A_OK;
    }
    const int n_seg = (int)cu.size() - 1;
    std::vector<AttnTile> tiles;
    std::vector<int> seg_tile0(n_seg + 1, 0);             // first attention tile of every segment
    const int attn_v = attention_version();
    if (mode != Q3A_MODE_CONV_ONLY) {
        if (attn_v == 1) attention_tiles(cu.data(), n_seg, tiles); else attention2_items(cu.data(), n_seg, tiles);
        for (int sgi = 0; sgi < n_seg; ++sgi) {
            const int len = cu[sgi + 1] - cu[sgi];
            seg_tile0[sgi + 1] = seg_tile0[sgi] + (attn_v == 1 ? attention_tile_count(len) : attention2_item_count(len));
        }
    }

    const int maxL1 = (max_len - 1) / 2 + 1;
    const int P2 = 1 + (maxL1 + 1) / 2;           // rows per chunk of conv2's input planes / output row space
    const int P3 = 1 + P2 / 2;                    // same for conv3
    const int F2 = hp.n_mel_bins / 4, F3 = hp.n_mel_bins / 8;   // mel half-widths: 32, 16

    // ---- tables: one pinned block -> one device block -------------------------------------------------
    auto al = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t o_chunks = 0;
    const size_t o_t2 = al(o_chunks + sizeof(StemChunk) * n_chunks);
    const size_t o_ntok = al(o_t2 + 4 * (size_t)n_chunks);
    const size_t o_base = al(o_ntok + 4 * (size_t)n_chunks);
    const size_t o_pos = al(o_base + 4 * (size_t)n_chunks);
    const size_t o_tiles = al(o_pos + 4 * (size_t)T);
    const size_t tab_bytes = al(o_tiles + sizeof(AttnTile) * tiles.size());
    void* hraw = nullptr;
    Q3A_TRY(e->ring.acquire(tab_bytes, &hraw));
    uint8_t* hb = static_cast<uint8_t*>(hraw);
    std::memcpy(hb + o_chunks, chunks.data(), sizeof(StemChunk) * n_chunks);
    std::memcpy(hb + o_t2, t2_valid.data(), 4 * (size_t)n_chunks);
    std::memcpy(hb + o_ntok, n_tok.data(), 4 * (size_t)n_chunks);
    std::memcpy(hb + o_base, tok_base.data(), 4 * (size_t)n_chunks);
    std::memcpy(hb + o_pos, pos.data(), 4 * (size_t)T);
    if (!tiles.empty()) std::memcpy(hb + o_tiles, tiles.data(), sizeof(AttnTile) * tiles.size());
    Q3A_TRY(e->tables.reserve(tab_bytes));
    Q3A_CUDA(cudaMemcpyAsync(e->tables.p, hb, tab_bytes, cudaMemcpyHostToDevice, st));
    Q3A_TRY(e->ring.mark(st));
    const uint8_t* db = e->tables.as<uint8_t>();
    const StemChunk* d_chunks = reinterpret_cast<const StemChunk*>(db + o_chunks);
    const int* d_t2 = reinterpret_cast<const int*>(db + o_t2);
    const int* d_ntok = reinterpret_cast<const int*>(db + o_ntok);
    const int* d_base = reinterpret_cast<const int*>(db + o_base);
    const int* d_pos = reinterpret_cast<const int*>(db + o_pos);
    const AttnTile* d_tiles = reinterpret_cast<const AttnTile*>(db + o_tiles);

    // ---- workspace ----------------------------------------------------------------------------------------
    if (mode != Q3A_MODE_CONV_ONLY) Q3A_TRY(ensure_pe(e, max_pos, st));
    const int r1 = n_chunks * P2, r2 = n_chunks * P3;
    // Every chunk of a pass owns P2 plane rows, P2 taken from the longest one.  The chunked modes bound that at 100 frames;
    // the debug modes (one chunk = one utterance) do not, so a ragged batch there can ask for n_utts * max_len rows.
    if ((double)r1 * F2 * C * 2.0 * 4.0 > 48e9) {
        set_error("encode: the stem workspace of this batch would exceed 48 GB (no-chunk / conv-only modes size every utterance's "
                  "planes by the longest one): split the batch by length");
        return Q3A_ERR_ARG;
    }
    if (r1 > e->rcap1 || P2 != e->sigP2) {
        const int cap = std::max(r1 + r1 / 4, e->rcap1);
        Q3A_CUDA(cudaStreamSynchronize(st));
        Q3A_TRY(e->planes1.reserve((size_t)4 * cap * F2 * C * 2));
        Q3A_CUDA(cudaMemsetAsync(e->planes1.p, 0, (size_t)4 * cap * F2 * C * 2, st));
        e->rcap1 = cap; e->sigP2 = P2;
    }
    // planes2's pad rows depend on P2 as well: conv2 writes ot <= P2-2, so for an even P2 the last odd-time row of every
    // chunk (read by conv3 as zero padding) is never written, while an odd P2 with the same P3 fills it with data
    if (r2 > e->rcap2 || P3 != e->sigP3 || P2 != e->sigP2_planes2) {
        const int cap = std::max(r2 + r2 / 4, e->rcap2);
        Q3A_CUDA(cudaStreamSynchronize(st));
        Q3A_TRY(e->planes2.reserve((size_t)4 * cap * F3 * C * 2));
        Q3A_CUDA(cudaMemsetAsync(e->planes2.p, 0, (size_t)4 * cap * F3 * C * 2, st));
        e->rcap2 = cap; e->sigP3 = P3; e->sigP2_planes2 = P2;
    }
    const int feat_dim = F3 * C;
    Q3A_TRY(e->feat.reserve((size_t)T * feat_dim * 2));
    if (mode != Q3A_MODE_CONV_ONLY) {
        Q3A_TRY(e->x.reserve((size_t)T * d * 4));
        Q3A_TRY(e->h.reserve((size_t)T * d * 2));
        if (e->ln_fold) Q3A_TRY(e->stats.reserve((size_t)T * ln_partials(d) * sizeof(float2)));
        Q3A_TRY(e->qkv.reserve((size_t)T * 3 * d * 2));
        Q3A_TRY(e->att.reserve((size_t)T * d * 2));
        Q3A_TRY(e->up.reserve((size_t)T * ffn * 2));
    }
    OpT* planes1 = e->planes1.as<OpT>();
    OpT* planes2 = e->planes2.as<OpT>();
    const long long ps1 = (long long)e->rcap1 * F2 * C, ps2 = (long long)e->rcap2 * F3 * C;

    // ---- stem, part by part ---------------------------------------------------------------------------------
    // LayerNorm fold (epilogues.cuh): `out16` / `stats` are the 16-bit copy and the statistics partials of the residual
    // stream for the rows of this launch (producers write them, LNFOLD consumers read `stats` and `colsum`)
    const bool fold = e->ln_fold;
    const bool fold1 = e->fold_mask & 1, fold2 = e->fold_mask & 2, foldp = e->fold_mask & 4;
    const int n_part = fold ? ln_partials(d) : 0;
    auto linear = [&](EpiKind kind, const void* A, int K, const q3a_encoder::Wt& W, int N, void* out, const float* bias,
                      const float* res, const float* pe_tab, const int* pos_tab, int rows, void* out16 = nullptr,
                      float2* stats = nullptr, const float* colsum = nullptr) -> int {
        if (rows <= 0) return Q3A_OK;
        LinearArgs a;
        a.a = A; a.M = rows; a.K = K; a.lda = K; a.w = W.p; a.w_d = W.d; a.N = N; a.ldw = K; a.out = out; a.ldc = N;
        a.bias = bias; a.res = res; a.pe = pe_tab; a.pos = pos_tab;
        a.out16 = out16; a.stats = stats; a.n_part = n_part; a.colsum = colsum;
        a.ln_inv_d = 1.0f / (float)d; a.ln_eps = hp.layer_norm_eps;
        const int sp = prof.begin(SP_GEMM);
        const int rc = gemm_linear(kind, fmt, a, st);
        prof.end(sp);
        e->tm.gemm_flops += 2.0 * (double)rows * N * K; ++e->tm.gemm_launches;
        return rc;
    };
    auto rows16 = [](void* p, size_t cols, int t0) { return (void*)(static_cast<uint16_t*>(p) + (size_t)t0 * cols); };
    float* x = e->x.as<float>();
    float* stem_out = mode == Q3A_MODE_CONV_ONLY ? out_dev : x;
    const int sp_stem = prof.begin(SP_STEM);
    for (size_t part = 0; part + 1 < in_split.size(); ++part) {
        const int ua = in_split[part], ub = in_split[part + 1];
        if (hk && hk->before_stem) Q3A_TRY(hk->before_stem((int)part, ua, ub));
        const int c0 = utt_chunk0[ua], nc = utt_chunk0[ub] - c0;
        const int t0 = (int)tok_off[ua], tn = (int)(tok_off[ub] - tok_off[ua]);
        if (nc == 0) continue;
        const int row2 = c0 * P3;                           // first planes2 row of the part
        // conv1 -> conv2.  Default: the whole part at once (planes1 rows = chunk * P2).  Experiment of VERDICT r1 item 4
        // (Q3A_STEM_GROUP=<chunks>): groups of that many chunks over a two-slot ring of plane rows, so that the planes
        // conv2 reads were written a moment ago and the ring (2 * G * 3.07 MB at C = 480) can stay dirty in the 126 MB
        // L2 instead of travelling to HBM and back.  Results are identical either way (same kernels, same rows).
        const int G = stem_group_chunks() > 0 ? stem_group_chunks() : nc;
        int gi = 0;
        for (int cg = c0; cg < c0 + nc; cg += G, ++gi) {
            const int ncg = std::min(G, c0 + nc - cg);
            const int row1 = stem_group_chunks() > 0 ? (gi & 1) * G * P2 : cg * P2;
            {
                const int sp = prof.begin(SP_OTHER);
                if (e->conv1_wimg)
                    Q3A_TRY(conv1_tc_launch<OpT>(mel_dev, d_chunks + cg, ncg, P2, max_dev, e->conv1_wimg, e->conv1_b, C, planes1,
                                                 e->rcap1, row1, st));
                else
                    Q3A_TRY(conv1_launch<OpT>(mel_dev, d_chunks + cg, ncg, P2, max_dev, e->conv1_w, e->conv1_b, C,
                                              planes1 + (size_t)row1 * F2 * C, ps1, st));
                prof.end(sp);
            }
            {
                ConvArgs a;
                a.planes = planes1 + (size_t)row1 * F2 * C; a.R_cap = e->rcap1; a.R_map = e->rcap1 - row1; a.R_used = ncg * P2;
                a.Fh = F2; a.C = C;
                a.w = e->conv2_w; a.N = C; a.bias = e->conv2_b; a.P = P2;
                a.out_planes = planes2 + (size_t)cg * P3 * F3 * C; a.P_next = P3; a.out_plane_stride = ps2; a.ot_valid = d_t2 + cg;
                const int sp = prof.begin(SP_GEMM);
                Q3A_TRY(gemm_conv(false, fmt, a, st));
                prof.end(sp);
                double m = 0; for (int c = cg; c < cg + ncg; ++c) m += (double)t2_valid[c] * F2;
                e->tm.gemm_flops += 2.0 * m * C * 9.0 * C; ++e->tm.gemm_launches;
            }
        }
        {
            ConvArgs a;
            a.planes = planes2 + (size_t)row2 * F3 * C; a.R_cap = e->rcap2; a.R_map = e->rcap2 - row2; a.R_used = nc * P3;
            a.Fh = F3; a.C = C;
            a.w = e->conv3_w; a.N = C; a.bias = e->conv3_b; a.P = P3;
            a.feat = e->feat.p; a.n_tok = d_ntok + c0; a.tok_base = d_base + c0;
            const int sp = prof.begin(SP_GEMM);
            Q3A_TRY(gemm_conv(true, fmt, a, st));
            prof.end(sp);
            e->tm.gemm_flops += 2.0 * (double)tn * F3 * C * 9.0 * C; ++e->tm.gemm_launches;
        }
        if (mode == Q3A_MODE_CONV_ONLY)
            Q3A_TRY(linear(EPI_POSEMB_F32, rows16(e->feat.p, feat_dim, t0), feat_dim, q3a_encoder::Wt{e->conv_out_w, nullptr}, d, stem_out + (size_t)t0 * d,
                           nullptr, nullptr, nullptr, nullptr, tn));
        else if (fold1)
            Q3A_TRY(linear(EPI_POSEMB_F32_DUAL, rows16(e->feat.p, feat_dim, t0), feat_dim, q3a_encoder::Wt{e->conv_out_w, nullptr}, d, stem_out + (size_t)t0 * d,
                           nullptr, nullptr, e->pe.as<float>(), d_pos + t0, tn, rows16(e->h.p, d, t0),
                           e->stats.as<float2>() + (size_t)t0 * n_part));
        else
            Q3A_TRY(linear(EPI_POSEMB_F32, rows16(e->feat.p, feat_dim, t0), feat_dim, q3a_encoder::Wt{e->conv_out_w, nullptr}, d, stem_out + (size_t)t0 * d,
                           nullptr, nullptr, e->pe.as<float>(), d_pos + t0, tn));
    }
    prof.end(sp_stem);
    if (mode == Q3A_MODE_CONV_ONLY) {
        if (hk && hk->after_out)
            for (size_t i = 0; i + 1 < out_split.size(); ++i) Q3A_TRY(hk->after_out((int)i, out_split[i], out_split[i + 1], tok_off.data()));
        return Q3A_OK;
    }

    // ---- transformer layers: all but the tail over the whole pass, the tail + projection part by part ----------
    const int sp_x = prof.begin(SP_XFMR);
    OpT* h = e->h.as<OpT>();
    const double attn_fl_per_tile_unit = 4.0 * 64 * H;
    auto run_layers = [&](int l0, int l1, int ua, int ub, bool post) -> int {
        const int t0 = (int)tok_off[ua], tn = (int)(tok_off[ub] - tok_off[ua]);
        if (tn <= 0) return Q3A_OK;
        const int sg0 = utt_seg0[ua], sg1 = utt_seg0[ub];
        const int tile0 = seg_tile0[sg0], n_tiles = seg_tile0[sg1] - tile0;
        double attn_fl = 0.0;
        int kt_mask = 0;                                  // key-block sizes among this launch's segments (v1 kernel)
        for (int sgi = sg0; sgi < sg1; ++sgi) {
            const double len = cu[sgi + 1] - cu[sgi];
            attn_fl += attn_fl_per_tile_unit * len * len;
            if (len > 0) kt_mask |= attention_key_block((int)len) == 128 ? 2 : 1;
        }
        float* xr = x + (size_t)t0 * d;
        void* hr = rows16(h, d, t0);
        void* attr = rows16(e->att.p, d, t0);
        void* upr = rows16(e->up.p, ffn, t0);
        auto ln = [&](const float* g, const float* b) -> int {
            const int sp = prof.begin(SP_OTHER);
            const int rc = layernorm_launch<OpT>(xr, g, b, hp.layer_norm_eps, tn, d, static_cast<OpT*>(hr), st);
            prof.end(sp);
            return rc;
        };
        float2* str = fold ? e->stats.as<float2>() + (size_t)t0 * n_part : nullptr;
        auto attention = [&]() -> int {
            const int sp = prof.begin(SP_ATTN);
            if (attn_v == 1) Q3A_TRY(attention_launch(e->qkv.p, e->att.p, T, H, d_tiles + tile0, n_tiles, fmt, kt_mask, st));
            else Q3A_TRY(attention2_launch(e->qkv.p, e->att.p, T, H, d_tiles + tile0, n_tiles, fmt, st));
            prof.end(sp);
            e->tm.attention_flops += attn_fl; ++e->tm.attention_launches;
            return Q3A_OK;
        };
        void* qkvr = rows16(e->qkv.p, 3 * (size_t)d, t0);
        for (int li = l0; li < l1; ++li) {
            const auto& L = e->layers[li];
            // Folded sites take the residual stream's 16-bit copy (h) and its row statistics from the GEMM that produced it;
            // the others run the LayerNorm kernel into h.  A producer emits copy + statistics when the next consumer folds.
            const bool last = li + 1 == (int)e->layers.size();
            const bool next1 = last ? foldp : fold1;               // what consumes the stream after this layer
            if (fold1) {
                Q3A_TRY(linear(EPI_LNFOLD_OP, hr, d, L.w_qkv, 3 * d, qkvr, L.b_qkv, nullptr, nullptr, nullptr, tn, nullptr, str, L.s_qkv));
            } else {
                Q3A_TRY(ln(L.ln1_g, L.ln1_b));
                Q3A_TRY(linear(EPI_BIAS_OP, hr, d, L.w_qkv, 3 * d, qkvr, L.b_qkv, nullptr, nullptr, nullptr, tn));
            }
            Q3A_TRY(attention());
            if (fold2) {
                Q3A_TRY(linear(EPI_BIAS_RES_F32_DUAL, attr, d, L.w_o, d, xr, L.b_o, xr, nullptr, nullptr, tn, hr, str));
                Q3A_TRY(linear(EPI_LNFOLD_GELU_OP, hr, d, L.w_up, ffn, upr, L.b_up, nullptr, nullptr, nullptr, tn, nullptr, str, L.s_up));
            } else {
                Q3A_TRY(linear(EPI_BIAS_RES_F32, attr, d, L.w_o, d, xr, L.b_o, xr, nullptr, nullptr, tn));
                Q3A_TRY(ln(L.ln2_g, L.ln2_b));
                Q3A_TRY(linear(EPI_BIAS_GELU_OP, hr, d, L.w_up, ffn, upr, L.b_up, nullptr, nullptr, nullptr, tn));
            }
            if (next1) Q3A_TRY(linear(EPI_BIAS_RES_F32_DUAL, upr, ffn, L.w_down, d, xr, L.b_down, xr, nullptr, nullptr, tn, hr, str));
            else Q3A_TRY(linear(EPI_BIAS_RES_F32, upr, ffn, L.w_down, d, xr, L.b_down, xr, nullptr, nullptr, tn));
        }
        if (post) {
            if (foldp) {
                Q3A_TRY(linear(EPI_LNFOLD_GELU_OP, hr, d, e->w_p1, d, attr, e->b_p1, nullptr, nullptr, nullptr, tn, nullptr, str, e->s_p1));
            } else {
                Q3A_TRY(ln(e->lnp_g, e->lnp_b));
                Q3A_TRY(linear(EPI_BIAS_GELU_OP, hr, d, e->w_p1, d, attr, e->b_p1, nullptr, nullptr, nullptr, tn));
            }
            Q3A_TRY(linear(EPI_BIAS_F32, attr, d, e->w_p2, hp.out_dim, out_dev + (size_t)t0 * hp.out_dim, e->b_p2, nullptr, nullptr,
                           nullptr, tn));
        }
        return Q3A_OK;
    };
    const int n_layers = (int)e->layers.size();
    const int n_out = (int)out_split.size() - 1;
    const int tail = n_out > 1 ? std::min(std::max(hk->tail_layers, 0), n_layers) : 0;
    Q3A_TRY(run_layers(0, n_layers - tail, 0, n_utts, n_out <= 1));
    if (n_out > 1)
        for (int i = 0; i < n_out; ++i) {
            Q3A_TRY(run_layers(n_layers - tail, n_layers, out_split[i], out_split[i + 1], true));
            if (hk->after_out) Q3A_TRY(hk->after_out(i, out_split[i], out_split[i + 1], tok_off.data()));
        }
    else if (hk && hk->after_out)
        Q3A_TRY(hk->after_out(0, 0, n_utts, tok_off.data()));
    prof.end(sp_x);
    return Q3A_OK;
}

static int encode_dispatch(q3a_encoder* e, const float* mel_dev, const int64_t* mel_off, const float* max_dev,
                           const int* n_frames, int n_utts, int mode, float* out_dev, int64_t* tok_off,
                           cudaStream_t st, const PartHooks* hk = nullptr) {
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("encode: unknown mode"); return Q3A_ERR_ARG; }
    if (e->fmt == OP_F16)
        return encode_core<__half>(e, mel_dev, mel_off, max_dev, n_frames, n_utts, mode, out_dev, tok_off, st, hk);
    return encode_core<__nv_bfloat16>(e, mel_dev, mel_off, max_dev, n_frames, n_utts, mode, out_dev, tok_off, st, hk);
}

// ---- loading ------------------------------------------------------------------------------------------------
static int load_weights(q3a_encoder* e, const GgufFile& f) {
    std::string err;
    q3a_hparams& hp = e->hp;
    const std::string P = "audio.encoder.";
    // hyper-parameters: the keys the converter writes (convert_hf_to_gguf.py:441-447), else tensor shapes,
    // else the reference defaults (gguf_loader.h:15-25)
    const GgufTensor* t_co = f.find(P + "conv_out.weight");
    const GgufTensor* t_c1 = f.find(P + "conv1.weight");
    const GgufTensor* t_up = f.find(P + "blk.0.ffn_up.weight");
    const GgufTensor* t_p2 = f.find(P + "proj2.weight");
    if (!t_co || !t_c1 || !t_up || !t_p2 || t_co->ne.size() != 2 || t_c1->ne.size() != 4 || t_up->ne.size() != 2 ||
        t_p2->ne.size() != 2) {
        set_error("GGUF file has no audio.encoder.* tensors of the expected rank");
        return Q3A_ERR_FORMAT;
    }
    int n_layers = 0;
    while (f.find(P + "blk." + std::to_string(n_layers) + ".attn_q.weight")) ++n_layers;
    const std::string A = "qwen3-asr.audio.";
    hp.d_model = (int)f.get_u32(A + "encoder.embedding_length", (uint32_t)t_co->ne[1]);
    hp.n_encoder_layers = (int)f.get_u32(A + "encoder.layer_count", (uint32_t)n_layers);
    hp.ffn_dim = (int)f.get_u32(A + "encoder.feed_forward_length", (uint32_t)t_up->ne[1]);
    hp.n_attention_heads = (int)f.get_u32(A + "encoder.attention.head_count", (uint32_t)(hp.d_model / 64));
    hp.n_mel_bins = (int)f.get_u32(A + "num_mel_bins", 128);
    hp.conv_channels = (int)f.get_u32(A + "conv_channels", (uint32_t)t_c1->ne[3]);
    hp.conv_out_dim = hp.d_model;
    // the reference's loader reads these two under its own key names with the same defaults (gguf_loader.cpp:76-77);
    // the converter writes neither.  The window length must be the 800 the attention windows are built for.
    hp.n_window_infer = (int)f.get_u32("audio.n_window_infer", 800);
    hp.layer_norm_eps = f.get_f32("audio.layer_norm_eps", 1e-5f);
    if (hp.n_window_infer != 800) { set_error("only n_window_infer = 800 (104-token attention windows) is supported"); return Q3A_ERR_FORMAT; }
    hp.out_dim = (int)t_p2->ne[1];
    hp.weight_ftype = (int)t_up->type;
    // Q8_0 files keep their Linear weights quantised on the device unless Q3A_Q8_DEQUANT_AT_LOAD is set; gamma cannot be
    // folded into 32-element blocks that share one scale, so that path runs the separate LayerNorm kernel
    e->q8_fused = hp.weight_ftype == (int)GGML_Q8_0 && feature_enabled("Q3A_Q8_FUSED", "Q3A_Q8_DEQUANT_AT_LOAD", kDefaultQ8Fused);
    if (e->q8_fused) e->ln_fold = false;
    // Q3A_LN_FOLD_MASK=<1..7> folds only some of the three LayerNorm sites (experiments; the others keep the kernel)
    // Default: attn_norm -> QKV and ln_post -> proj1 folded, ffn_norm as a kernel.  The FFN-up GEMM is paced by its
    // epilogue (bias + GELU on 3584 columns); the fold's two extra FMAs and second column vector per element cost it 32 us
    // per layer (139 -> 171 us, ncu r2b) against the 24 us the LayerNorm kernel takes, while QKV absorbs them for free.
    e->fold_mask = e->ln_fold ? 5 : 0;
    if (e->ln_fold) if (const char* m = getenv("Q3A_LN_FOLD_MASK")) { e->fold_mask = atoi(m) & 7; e->ln_fold = e->fold_mask != 0; }
    const int d = hp.d_model, C = hp.conv_channels, ffn = hp.ffn_dim, F3 = hp.n_mel_bins / 8;
    if (hp.n_mel_bins != 128) { set_error("only 128 mel bins are supported"); return Q3A_ERR_FORMAT; }
    if (hp.n_attention_heads * 64 != d) { set_error("head_dim must be 64 (d_model = 64 * heads)"); return Q3A_ERR_FORMAT; }
    if (d % 8 || ffn % 8 || C % 8 || hp.out_dim % 8 || d > 1024 || n_layers < hp.n_encoder_layers || d < 4) {
        set_error("unsupported encoder shape (need multiples of 8, d_model <= 1024, all layers present)");
        return Q3A_ERR_FORMAT;
    }
    if (t_co->ne[0] != (int64_t)C * F3) { set_error("conv_out.weight does not match conv_channels * 16"); return Q3A_ERR_FORMAT; }

    ArenaPlan plan;
    std::vector<float> v, v2, v3, packed;
#define FETCH(name, vec, n) do { if (!fetch(f, P + (name), vec, err, (n))) { set_error(err); return Q3A_ERR_FORMAT; } } while (0)
    // conv1: [C][1][3 mel][3 time] -> f32 [C][9] unchanged (index df*3 + dt)
    FETCH("conv1.weight", v, (int64_t)C * 9);  plan.add_f32(&e->conv1_w, v);
    if (conv1_tc_supported(C)) {
        std::vector<uint8_t> img;
        conv1_tc_weight_image(v.data(), C, img);
        plan.add_raw(&e->conv1_wimg, img);
    }
    FETCH("conv1.bias", v, C);                 plan.add_f32(&e->conv1_b, v);
    // conv2/3: [OC][IC][3 mel (df)][3 time (dt)] -> [OC][tap = dt*3 + df][cpad] (gemm_tc.cuh LoaderConv)
    const int cpad = (C + 63) / 64 * 64;
    for (int ci = 2; ci <= 3; ++ci) {
        FETCH("conv" + std::to_string(ci) + ".weight", v, (int64_t)C * C * 9);
        packed.assign((size_t)C * 9 * cpad, 0.0f);
        for (int oc = 0; oc < C; ++oc)
            for (int ic = 0; ic < C; ++ic)
                for (int df = 0; df < 3; ++df)
                    for (int dt = 0; dt < 3; ++dt)
                        packed[((size_t)oc * 9 + dt * 3 + df) * cpad + ic] = v[(((size_t)oc * C + ic) * 3 + df) * 3 + dt];
        plan.add_op(ci == 2 ? &e->conv2_w : &e->conv3_w, packed, e->fmt);
        FETCH("conv" + std::to_string(ci) + ".bias", v, C);
        plan.add_f32(ci == 2 ? &e->conv2_b : &e->conv3_b, v);
    }
    // conv_out: [d][c*16 + f] (audio_encoder.cpp:139-142) -> [d][f*C + c] (the order conv3's epilogue writes)
    FETCH("conv_out.weight", v, (int64_t)d * C * F3);
    packed.assign(v.size(), 0.0f);
    for (int n = 0; n < d; ++n)
        for (int c = 0; c < C; ++c)
            for (int ff = 0; ff < F3; ++ff) packed[(size_t)n * C * F3 + ff * C + c] = v[(size_t)n * C * F3 + c * F3 + ff];
    plan.add_op(&e->conv_out_w, packed, e->fmt);

    // Linear weights.  16-bit path: dequantised to f32 here, rounded once to the operand type (optionally with the
    // LayerNorm folded in).  Fused Q8_0 path (Q8_0 file, K % 64 == 0): the ggml blocks go to the device as they are,
    // de-interleaved into an int8 plane and an f16 scale plane; the GEMM's dequant warps expand them (gemm_tc.cuh).
    auto q8_ok = [&](const std::string& name) {
        const GgufTensor* t = f.find(P + name);
        return e->q8_fused && t && t->type == GGML_Q8_0 && t->ne.size() == 2 && t->ne[0] % 64 == 0;
    };
    auto add_linear = [&](q3a_encoder::Wt* slot, const std::vector<std::string>& names, int64_t N_each, int64_t K,
                          const std::vector<float>* gamma, const std::vector<float>* beta, std::vector<float>* bias,
                          const float** colsum_slot) -> bool {
        bool all_q8 = true;
        for (const auto& nm : names) all_q8 = all_q8 && q8_ok(nm);
        if (all_q8) {
            std::vector<int8_t> q, qa; std::vector<uint16_t> dd, da;
            for (const auto& nm : names) {
                const GgufTensor* t = f.find(P + nm);
                if (t->n_elements() != N_each * K) { err = "tensor " + nm + " has an unexpected size"; return false; }
                if (!f.read_q8_planes(*t, q, dd, err)) return false;
                qa.insert(qa.end(), q.begin(), q.end()); da.insert(da.end(), dd.begin(), dd.end());
            }
            plan.add_bytes(&slot->p, qa.data(), qa.size());
            plan.add_bytes(&slot->d, da.data(), da.size() * 2);
            return true;
        }
        std::vector<float> w, one;
        for (const auto& nm : names) {
            if (!fetch(f, P + nm, one, err, N_each * K)) return false;
            w.insert(w.end(), one.begin(), one.end());
        }
        if (gamma) {
            std::vector<float> bias2, cs;
            fold_layernorm(w, (int)(N_each * (int64_t)names.size()), (int)K, *gamma, *beta, *bias, e->fmt, bias2, cs);
            *bias = bias2;
            plan.add_f32(colsum_slot, cs);
        }
        plan.add_op(&slot->p, w, e->fmt);
        return true;
    };
#define LINEAR(...) do { if (!add_linear(__VA_ARGS__)) { set_error(err); return Q3A_ERR_FORMAT; } } while (0)
    const bool fold1 = e->fold_mask & 1, fold2 = e->fold_mask & 2, foldp = e->fold_mask & 4;
    e->layers.assign(hp.n_encoder_layers, q3a_encoder::Layer{});
    for (int l = 0; l < hp.n_encoder_layers; ++l) {
        auto& L = e->layers[l];
        const std::string B = "blk." + std::to_string(l) + ".";
        std::vector<float> g1, b1, g2, b2, bias;
        FETCH(B + "attn_norm.weight", g1, d); plan.add_f32(&L.ln1_g, g1);
        FETCH(B + "attn_norm.bias", b1, d);   plan.add_f32(&L.ln1_b, b1);
        FETCH(B + "attn_q.bias", v, d); FETCH(B + "attn_k.bias", v2, d); FETCH(B + "attn_v.bias", v3, d);
        bias = v; bias.insert(bias.end(), v2.begin(), v2.end()); bias.insert(bias.end(), v3.begin(), v3.end());
        LINEAR(&L.w_qkv, {B + "attn_q.weight", B + "attn_k.weight", B + "attn_v.weight"}, d, d, fold1 ? &g1 : nullptr, &b1, &bias, &L.s_qkv);
        plan.add_f32(&L.b_qkv, bias);
        LINEAR(&L.w_o, {B + "attn_out.weight"}, d, d, nullptr, nullptr, nullptr, nullptr);
        FETCH(B + "attn_out.bias", v, d);                plan.add_f32(&L.b_o, v);
        FETCH(B + "ffn_norm.weight", g2, d); plan.add_f32(&L.ln2_g, g2);
        FETCH(B + "ffn_norm.bias", b2, d);   plan.add_f32(&L.ln2_b, b2);
        FETCH(B + "ffn_up.bias", bias, ffn);
        LINEAR(&L.w_up, {B + "ffn_up.weight"}, ffn, d, fold2 ? &g2 : nullptr, &b2, &bias, &L.s_up);
        plan.add_f32(&L.b_up, bias);
        LINEAR(&L.w_down, {B + "ffn_down.weight"}, d, ffn, nullptr, nullptr, nullptr, nullptr);
        FETCH(B + "ffn_down.bias", v, d);                  plan.add_f32(&L.b_down, v);
    }
    {
        std::vector<float> gp, bp, bias;
        FETCH("ln_post.weight", gp, d); plan.add_f32(&e->lnp_g, gp);
        FETCH("ln_post.bias", bp, d);   plan.add_f32(&e->lnp_b, bp);
        FETCH("proj1.bias", bias, d);
        LINEAR(&e->w_p1, {"proj1.weight"}, d, d, foldp ? &gp : nullptr, &bp, &bias, &e->s_p1);
        plan.add_f32(&e->b_p1, bias);
    }
    LINEAR(&e->w_p2, {"proj2.weight"}, hp.out_dim, d, nullptr, nullptr, nullptr, nullptr);
    FETCH("proj2.bias", v, hp.out_dim);                 plan.add_f32(&e->b_p2, v);
#undef LINEAR
#undef FETCH

    size_t total = 0;
    for (auto& it : plan.items) total += (it.bytes.size() + 255) / 256 * 256;
    Q3A_TRY(e->arena.reserve(total));
    size_t off = 0;
    std::vector<uint8_t> host(total, 0);
    for (auto& it : plan.items) {
        std::memcpy(host.data() + off, it.bytes.data(), it.bytes.size());
        *it.slot = e->arena.as<uint8_t>() + off;
        off += (it.bytes.size() + 255) / 256 * 256;
    }
    Q3A_CUDA(cudaMemcpy(e->arena.p, host.data(), total, cudaMemcpyHostToDevice));
    e->arena_used = total;
    return Q3A_OK;
}

}  // namespace q3a

using namespace q3a;

// q3a_pack_offsets is defined in mel.cu
int q3a_pack_offsets(const int* n, int count, int unit, int64_t align, std::vector<int64_t>& off);

extern "C" void q3a_encoder_free(q3a_encoder* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    e->arena.release(); e->pe.release();
    e->planes1.release(); e->planes2.release(); e->feat.release(); e->x.release(); e->h.release(); e->qkv.release();
    e->att.release(); e->up.release(); e->stats.release(); e->tables.release(); e->stage_in.release(); e->stage_out.release();
    e->stage_in2.release(); e->stage_out2.release(); e->stage_pcm.release();
    e->stage_mel.release(); e->stage_max.release();
    e->ring.release();
    for (auto ev : e->ev_pool) cudaEventDestroy(ev);
    for (int i = 0; i < 2; ++i) {
        if (e->ev_in[i]) cudaEventDestroy(e->ev_in[i]);
        if (e->ev_comp[i]) cudaEventDestroy(e->ev_comp[i]);
        if (e->ev_out[i]) cudaEventDestroy(e->ev_out[i]);
        for (auto ev : e->ev_part_in[i]) if (ev) cudaEventDestroy(ev);
        for (auto ev : e->ev_part_out[i]) if (ev) cudaEventDestroy(ev);
    }
    for (auto ev : e->ev_done) if (ev) cudaEventDestroy(ev);
    if (e->copy_in) cudaStreamDestroy(e->copy_in);
    if (e->copy_out) cudaStreamDestroy(e->copy_out);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

extern "C" int q3a_encoder_load(const char* gguf_path, int device, int operand_dtype, q3a_encoder** out) {
    if (!out) { set_error("encoder_load: out is null"); return Q3A_ERR_ARG; }
    *out = nullptr;
    if (!gguf_path) { set_error("encoder_load: path is null"); return Q3A_ERR_ARG; }
    if (operand_dtype != Q3A_OP_BF16 && operand_dtype != Q3A_OP_F16) { set_error("encoder_load: unknown operand dtype"); return Q3A_ERR_ARG; }
    GgufFile f;
    std::string err;
    if (!f.open(gguf_path, err)) {
        set_error(err);
        return (err.rfind("cannot open", 0) == 0 || err.rfind("cannot stat", 0) == 0 || err.rfind("mmap", 0) == 0) ? Q3A_ERR_IO : Q3A_ERR_FORMAT;
    }
    Q3A_TRY(use_device(device));
    q3a_encoder* e = new q3a_encoder();
    e->device = device; e->fmt = operand_dtype;
    e->ln_fold = feature_enabled("Q3A_LN_FOLD", "Q3A_NO_LN_FOLD", kDefaultLnFold);   // else separate layernorm_kernel launches
    int rc = load_weights(e, f);
    if (rc == Q3A_OK) {
        bool ok = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamCreateWithFlags(&e->copy_in, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamCreateWithFlags(&e->copy_out, cudaStreamNonBlocking) == cudaSuccess;
        for (int i = 0; i < 2 && ok; ++i)
            ok = cudaEventCreateWithFlags(&e->ev_in[i], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&e->ev_comp[i], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&e->ev_out[i], cudaEventDisableTiming) == cudaSuccess;
        for (int i = 0; i < 2 && ok; ++i) {
            for (auto& ev : e->ev_part_in[i]) ok = ok && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess;
            for (auto& ev : e->ev_part_out[i]) ok = ok && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess;
        }
        for (auto& ev : e->ev_done) ok = ok && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess;
        if (!ok) { set_error("cudaStreamCreate / cudaEventCreate failed"); rc = Q3A_ERR_CUDA; }
    }
    if (rc != Q3A_OK) { q3a_encoder_free(e); return rc; }
    *out = e;
    return Q3A_OK;
}

extern "C" int q3a_encoder_get_hparams(const q3a_encoder* e, q3a_hparams* out) {
    if (!e) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (!out) { set_error("get_hparams: out is null"); return Q3A_ERR_ARG; }
    *out = e->hp;
    return Q3A_OK;
}

extern "C" int q3a_encoder_n_tokens(int n_frames, int mode) { return encoder_n_tokens(n_frames, mode); }

extern "C" int q3a_encoder_set_profiling(q3a_encoder* e, int on) {
    if (!e) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    e->profiling = on != 0;
    return Q3A_OK;
}

extern "C" int q3a_encoder_get_timings(const q3a_encoder* e, q3a_timings* out) {
    if (!e) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (!out) { set_error("get_timings: out is null"); return Q3A_ERR_ARG; }
    *out = e->tm;
    return Q3A_OK;
}

extern "C" int64_t q3a_encoder_device_bytes(const q3a_encoder* e) {
    if (!e) return 0;
    return (int64_t)(e->arena.cap + e->pe.cap + e->planes1.cap + e->planes2.cap + e->feat.cap + e->x.cap + e->h.cap +
                     e->qkv.cap + e->att.cap + e->up.cap + e->stats.cap + e->tables.cap + e->stage_in.cap + e->stage_out.cap + e->stage_in2.cap + e->stage_out2.cap + e->stage_pcm.cap +
                     e->stage_mel.cap + e->stage_max.cap);
}

extern "C" int q3a_encoder_encode_device(q3a_encoder* e, const float* mel_dev, const int64_t* mel_off,
                                         const float* max_dev, const int* n_frames, int n_utts, int mode,
                                         float* out_dev, int64_t* tok_off, void* stream) {
    if (!e) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (n_utts < 0 || (n_utts > 0 && (!mel_dev || !mel_off || !n_frames || !out_dev))) { set_error("encode_device: null argument"); return Q3A_ERR_ARG; }
    if ((reinterpret_cast<uintptr_t>(mel_dev) & 15u) || (reinterpret_cast<uintptr_t>(out_dev) & 15u)) {
        set_error("encode_device: mel_dev and out_dev must be 16-byte aligned"); return Q3A_ERR_ARG;
    }
    Q3A_TRY(use_device(e->device));
    g_launch_counter = 0;
    prof_reset(e);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    Q3A_TRY(encode_dispatch(e, mel_dev, mel_off, max_dev, n_frames, n_utts, mode, out_dev, tok_off, st));
    return prof_collect(e, st);
}

// Splits [0, n) into passes.  A pass bounds the stem workspace (kMaxChunksPerPass); consecutive passes of the
// host-pointer entry points overlap as upload | compute | download.  Inside a pass the copies overlap the
// kernels at a finer grain (PartHooks), which measured better than cutting a batch that fits into two passes
// (each pass pays its own ramp-up and tail: 64 x 30 s ran 24.8 ms as 2 passes, 24.0 ms as one pass with parts).
// Q3A_PIPELINE_MIN_CHUNKS=<n> forces a split into two passes once a call has 2n chunks (experiments only).
// Results do not depend on the split (the kernels are batch-invariant).
static int pipeline_min_chunks() {
    static const int v = [] {
        const char* e = getenv("Q3A_PIPELINE_MIN_CHUNKS");
        const int x = e ? atoi(e) : 0;
        return x > 0 ? x : (1 << 30);
    }();
    return v;
}

static int chunks_of(int n_frames, int mode) {
    const int c = (n_frames + kChunkFrames - 1) / kChunkFrames;
    return (mode == Q3A_MODE_ASR || mode == Q3A_MODE_ALIGNER) ? c : c + 1;
}

static void plan_passes(const int* n_frames, int n, int mode, std::vector<int>& bounds) {
    long long total = 0;
    for (int i = 0; i < n; ++i) total += chunks_of(n_frames[i], mode);
    long long n_pass = (total + kMaxChunksPerPass - 1) / kMaxChunksPerPass;
    if (n_pass < 2 && total >= 2LL * pipeline_min_chunks()) n_pass = 2;
    if (n_pass < 1) n_pass = 1;
    const long long target = (total + n_pass - 1) / n_pass;
    bounds.assign(1, 0);
    long long chunks = 0;
    for (int i = 0; i < n; ++i) {
        const int c = chunks_of(n_frames[i], mode);
        if (chunks > 0 && (chunks + c > kMaxChunksPerPass || chunks >= target)) { bounds.push_back(i); chunks = 0; }
        chunks += c;
    }
    bounds.push_back(n);
}

// Cuts [0, n) into at most `parts` runs whose cumulative weight follows `frac` (ascending, last = 1).
static std::vector<int> weighted_split(const std::vector<long long>& weight, int parts, const double* frac) {
    const int n = (int)weight.size();
    long long total = 0;
    for (long long w : weight) total += w;
    std::vector<int> sp{0};
    long long acc = 0;
    for (int i = 0; i < n; ++i) {
        acc += weight[i];
        const int k = (int)sp.size();                       // next boundary to place
        if (k < parts && i + 1 < n && (double)acc >= frac[k - 1] * (double)total) sp.push_back(i + 1);
    }
    sp.push_back(n);
    return sp;
}
// uploads: a small first part (only its copy is exposed), then equal parts; downloads: halves
static const double kInFrac[4] = {0.10, 0.40, 0.70, 1.0};
static const double kOne[1] = {1.0};
// Download split and the number of trailing layers run per download part.  Tuning switches (not API):
// Q3A_OUT_FRAC="0.4,0.75,1" (ascending cumulative fractions, at most kOutParts), Q3A_TAIL_LAYERS=n.
struct OutSplit { double frac[q3a_encoder::kOutParts]; int parts; int tail_layers; };
static OutSplit out_split_config() {
    OutSplit c{{0.5, 1.0, 1.0, 1.0}, 2, 3};
    if (const char* f = getenv("Q3A_OUT_FRAC")) {
        int n = 0;
        double prev = 0.0;
        for (const char* p = f; *p && n < q3a_encoder::kOutParts;) {
            char* end = nullptr;
            const double v = strtod(p, &end);
            if (end == p) break;
            if (v > prev && v <= 1.0) { c.frac[n++] = v; prev = v; }
            p = *end == ',' ? end + 1 : end;
            if (*end != ',') break;
        }
        if (n > 0) { c.frac[n - 1] = 1.0; c.parts = n; }
    }
    if (const char* t = getenv("Q3A_TAIL_LAYERS")) c.tail_layers = std::max(0, atoi(t));
    return c;
}

// The pipeline itself.  `in[i]` has `in_elems[i]` floats on the host; `compute(u0, nu, in_dev, in_off, out_dev,
// hooks)` enqueues the pass on e->stream.  Inside a pass the upload is cut into kInParts and the download into
// kOutParts (PartHooks): the stem of part i runs while part i+1 is still on the wire, and the rows of the first
// output part travel back while the second runs its last layers.
constexpr long long kSplitMinBytes = 16ll << 20;

// Host-side plan of one host-pointer call: passes, staging offsets, upload / download parts of every pass.
struct HostPlan {
    std::vector<int> bounds;                       // pass p = utterances [bounds[p], bounds[p+1])
    std::vector<std::vector<int64_t>> ioff;        // per pass: element offset of every utterance in the staging buffer
    std::vector<std::vector<int>> in_split, out_split;   // per pass: part boundaries (pass-relative utterance indices)
    std::vector<long long> tokens;                 // per pass: output rows
};

static void plan_host_call(const int* in_elems, int elem_bytes, int64_t align_elems, const int* n_frames, int n, int mode,
                           int width, bool split_ok, HostPlan& plan) {
    plan_passes(n_frames, n, mode, plan.bounds);
    const int n_pass = (int)plan.bounds.size() - 1;
    plan.ioff.assign(n_pass, {}); plan.in_split.assign(n_pass, {}); plan.out_split.assign(n_pass, {});
    plan.tokens.assign(n_pass, 0);
    for (int p = 0; p < n_pass; ++p) {
        const int u0 = plan.bounds[p], nu = plan.bounds[p + 1] - u0;
        q3a_pack_offsets(in_elems + u0, nu, 1, align_elems, plan.ioff[p]);
        std::vector<long long> w_in(nu), w_out(nu);
        long long T = 0;
        for (int i = 0; i < nu; ++i) {
            w_in[i] = in_elems[u0 + i];
            w_out[i] = encoder_n_tokens(n_frames[u0 + i], mode);
            T += w_out[i];
        }
        plan.tokens[p] = T;
        const long long in_bytes = (long long)elem_bytes * plan.ioff[p][nu], out_bytes = 4ll * T * width;
        plan.in_split[p] = split_ok && in_bytes >= kSplitMinBytes ? weighted_split(w_in, q3a_encoder::kInParts, kInFrac)
                                                                  : weighted_split(w_in, 1, kOne);
        const OutSplit oc = out_split_config();
        plan.out_split[p] = split_ok && out_bytes >= kSplitMinBytes && mode != Q3A_MODE_CONV_ONLY
                                ? weighted_split(w_out, oc.parts, oc.frac) : weighted_split(w_out, 1, kOne);
    }
}

// `wait` = false: everything is enqueued and the call returns (q3a_encode_pcm_submit); the staging buffers and their events
// are picked by a pass counter that runs across calls, so the upload of the next call overlaps the tail of this one.
template <class Compute>
static int run_host_passes(q3a_encoder* e, const void* const* in, const int* in_elems, int elem_bytes, int64_t align_elems,
                           const int* n_frames, int n, int mode, int width, float* const* out, int* n_tokens, Compute compute,
                           bool wait = true) {
    static const bool trace = getenv("Q3A_TRACE") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    auto stamp = [&](const char* what) {
        if (trace) fprintf(stderr, "[q3a trace] %-28s %8.3f ms\n", what,
                           std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count());
    };
    HostPlan plan;
    // Intra-call parts overlap a call's own copies with its own kernels: what a synchronous caller needs.  With calls in
    // flight back to back (submit / wait) the next call's upload and the previous call's download already run under this
    // call's kernels, and the parts would only cut the launches into smaller, less efficient ones; Q3A_ASYNC_PARTS=1 keeps them.
    const bool parts = getenv("Q3A_NO_PART_OVERLAP") == nullptr && (wait || getenv("Q3A_ASYNC_PARTS") != nullptr);
    plan_host_call(in_elems, elem_bytes, align_elems, n_frames, n, mode, width, parts, plan);
    const std::vector<int>& bounds = plan.bounds;
    const std::vector<std::vector<int64_t>>& ioff = plan.ioff;
    const std::vector<std::vector<int>>&in_split = plan.in_split, &out_split = plan.out_split;
    const int n_pass = (int)bounds.size() - 1;
    const long long g0 = e->pass_seq;                      // global index of this call's first pass
    DevBuf* sin[2] = {&e->stage_in, &e->stage_in2};
    DevBuf* sout[2] = {&e->stage_out, &e->stage_out2};
    size_t need_in[2] = {0, 0}, need_out[2] = {0, 0};
    for (int p = 0; p < n_pass; ++p) {
        const int nu = bounds[p + 1] - bounds[p];
        need_in[(g0 + p) & 1] = std::max(need_in[(g0 + p) & 1], (size_t)elem_bytes * (size_t)(ioff[p][nu] + align_elems));
        need_out[(g0 + p) & 1] = std::max(need_out[(g0 + p) & 1], sizeof(float) * ((size_t)plan.tokens[p] * width + 64));
    }
    for (int b = 0; b < 2; ++b) {
        if (need_in[b]) Q3A_TRY(sin[b]->reserve(need_in[b]));
        if (need_out[b]) Q3A_TRY(sout[b]->reserve(need_out[b]));
    }
    auto drain = [&](int rc) {
        cudaStreamSynchronize(e->copy_in); cudaStreamSynchronize(e->stream); cudaStreamSynchronize(e->copy_out);
        return rc;
    };
#define PIPE_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { set_error(std::string(#x) + ": " + cudaGetErrorString(e_)); return drain(Q3A_ERR_CUDA); } } while (0)
#define HOOK_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { set_error(std::string(#x) + ": " + cudaGetErrorString(e_)); return (int)Q3A_ERR_CUDA; } } while (0)
    auto upload = [&](int p) -> int {
        const int b = (int)((g0 + p) & 1), u0 = bounds[p];
        if (g0 + p >= 2) PIPE_CUDA(cudaStreamWaitEvent(e->copy_in, e->ev_comp[b], 0));   // pass g-2 (maybe of an earlier call) has consumed the buffer
        const std::vector<int>& sp = in_split[p];
        for (size_t part = 0; part + 1 < sp.size(); ++part) {
            // one copy per run of clips that are contiguous on both sides (a caller that keeps its clips in one pinned
            // block gets one transfer per part instead of one per clip)
            for (int i = sp[part]; i < sp[part + 1];) {
                if (in_elems[u0 + i] <= 0) { ++i; continue; }
                const uint8_t* src = static_cast<const uint8_t*>(in[u0 + i]);
                const size_t dst0 = (size_t)elem_bytes * ioff[p][i];
                size_t bytes = (size_t)elem_bytes * in_elems[u0 + i];
                int j = i + 1;
                while (j < sp[part + 1] && in_elems[u0 + j] > 0 && static_cast<const uint8_t*>(in[u0 + j]) == src + bytes &&
                       (size_t)elem_bytes * ioff[p][j] == dst0 + bytes) {
                    bytes += (size_t)elem_bytes * in_elems[u0 + j];
                    ++j;
                }
                PIPE_CUDA(cudaMemcpyAsync(sin[b]->as<uint8_t>() + dst0, src, bytes, cudaMemcpyHostToDevice, e->copy_in));
                i = j;
            }
            PIPE_CUDA(cudaEventRecord(e->ev_part_in[b][part], e->copy_in));
        }
        return Q3A_OK;
    };
    stamp("planned");
    if (n_pass > 0) { const int rc = upload(0); if (rc != Q3A_OK) return rc; }
    stamp("upload 0 enqueued");
    for (int p = 0; p < n_pass; ++p) {
        const int b = (int)((g0 + p) & 1), u0 = bounds[p], nu = bounds[p + 1] - u0;
        if (p + 1 < n_pass) { const int rc = upload(p + 1); if (rc != Q3A_OK) return rc; }
        if (g0 + p >= 2) PIPE_CUDA(cudaStreamWaitEvent(e->stream, e->ev_out[b], 0));  // pass g-2 has been downloaded
        PartHooks hk;
        hk.in_split = in_split[p];
        hk.out_split = out_split[p];
        hk.tail_layers = out_split_config().tail_layers;
        hk.before_stem = [&, b](int part, int, int) -> int {
            if (part == 0) stamp("first before_stem");
            HOOK_CUDA(cudaStreamWaitEvent(e->stream, e->ev_part_in[b][part], 0));
            return Q3A_OK;
        };
        hk.after_out = [&, b, u0](int part, int ua, int ub, const int64_t* toff) -> int {
            HOOK_CUDA(cudaEventRecord(e->ev_part_out[b][part], e->stream));
            HOOK_CUDA(cudaStreamWaitEvent(e->copy_out, e->ev_part_out[b][part], 0));
            for (int i = ua; i < ub; ++i)
                if (n_tokens) n_tokens[u0 + i] = (int)(toff[i + 1] - toff[i]);
            for (int i = ua; i < ub;) {                       // rows are consecutive on the device: merge contiguous host targets
                const int64_t rows = toff[i + 1] - toff[i];
                if (rows <= 0) { ++i; continue; }
                float* dst = out[u0 + i];
                int64_t total = rows;
                int j = i + 1;
                while (j < ub && (toff[j + 1] - toff[j] <= 0 || out[u0 + j] == dst + total * width)) { total += toff[j + 1] - toff[j]; ++j; }
                HOOK_CUDA(cudaMemcpyAsync(dst, sout[b]->as<float>() + toff[i] * width, sizeof(float) * total * width,
                                          cudaMemcpyDeviceToHost, e->copy_out));
                i = j;
            }
            return Q3A_OK;
        };
        const int rc = compute(u0, nu, sin[b]->p, ioff[p].data(), sout[b]->as<float>(), &hk);
        if (rc != Q3A_OK) return drain(rc);
        stamp("compute enqueued");
        PIPE_CUDA(cudaEventRecord(e->ev_comp[b], e->stream));
        PIPE_CUDA(cudaEventRecord(e->ev_out[b], e->copy_out));
    }
    e->pass_seq += n_pass;
    // completion of this call = its last download (the download stream waits for the compute that feeds it); an empty call
    // completes behind whatever was submitted before it
    PIPE_CUDA(cudaStreamWaitEvent(e->copy_out, n_pass > 0 ? e->ev_comp[(g0 + n_pass - 1) & 1] : e->ev_comp[0], 0));
    PIPE_CUDA(cudaEventRecord(e->ev_done[e->ticket_seq % q3a_encoder::kTickets], e->copy_out));
    ++e->ticket_seq;
    if (!wait) { stamp("submitted"); return Q3A_OK; }
    if (trace) { cudaStreamSynchronize(e->copy_in); stamp("uploads done"); cudaStreamSynchronize(e->stream); stamp("compute done"); }
    PIPE_CUDA(cudaStreamSynchronize(e->copy_out));
    PIPE_CUDA(cudaStreamSynchronize(e->stream));
    stamp("all done");
#undef PIPE_CUDA
#undef HOOK_CUDA
    return Q3A_OK;
}

extern "C" int q3a_encoder_encode_batch(q3a_encoder* e, const float* const* mel, int n_mel, const int* n_frames,
                                        int n_utts, int mode, float* const* out, int* n_tokens) {
    if (!e) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (n_mel != e->hp.n_mel_bins) {
        set_error("Mel bins mismatch: expected " + std::to_string(e->hp.n_mel_bins) + ", got " + std::to_string(n_mel));
        return Q3A_ERR_ARG;
    }
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("encode: unknown mode"); return Q3A_ERR_ARG; }
    if (n_utts < 0 || (n_utts > 0 && (!mel || !n_frames || !out))) { set_error("encode_batch: null argument"); return Q3A_ERR_ARG; }
    for (int u = 0; u < n_utts; ++u)
        if (n_frames[u] < 0 || (n_frames[u] > 0 && (!mel[u] || !out[u]))) { set_error("encode_batch: bad utterance " + std::to_string(u)); return Q3A_ERR_ARG; }
    Q3A_TRY(use_device(e->device));
    g_launch_counter = 0;
    prof_reset(e);
    const int width = mode == Q3A_MODE_CONV_ONLY ? e->hp.d_model : e->hp.out_dim;
    std::vector<int> elems(n_utts);
    for (int i = 0; i < n_utts; ++i) elems[i] = n_mel * n_frames[i];
    Q3A_TRY(run_host_passes(e, reinterpret_cast<const void* const*>(mel), elems.data(), 4, 64, n_frames, n_utts, mode, width, out, n_tokens,
                            [&](int u0, int nu, const void* in_dev, const int64_t* in_off, float* out_dev, const PartHooks* hk) {
                                return encode_dispatch(e, static_cast<const float*>(in_dev), in_off, nullptr, n_frames + u0, nu, mode, out_dev, nullptr, e->stream, hk);
                            }));
    return prof_collect(e, e->stream);
}

// mel (raw log10 + per-clip max) and encoder (clamp fused into conv1's loads) for clips on the device
static int pcm_core(q3a_frontend* fe, q3a_encoder* e, const float* pcm_dev, const int64_t* pcm_off, const int* n_samples,
                    int n_clips, int mode, float* out_dev, int64_t* tok_off, cudaStream_t st, const PartHooks* hk = nullptr) {
    if (fe->n_mel != e->hp.n_mel_bins) {
        set_error("Mel bins mismatch: expected " + std::to_string(e->hp.n_mel_bins) + ", got " + std::to_string(fe->n_mel));
        return Q3A_ERR_ARG;
    }
    if (fe->device != e->device) { set_error("frontend and encoder live on different devices"); return Q3A_ERR_ARG; }
    std::vector<int> elems(n_clips), nf(n_clips);
    std::vector<int64_t> moff;
    for (int i = 0; i < n_clips; ++i) {
        if (n_samples[i] < 0) { set_error("negative n_samples"); return Q3A_ERR_ARG; }
        nf[i] = n_samples[i] / 160; elems[i] = fe->n_mel * nf[i];
    }
    q3a_pack_offsets(elems.data(), n_clips, 1, 64, moff);
    Q3A_TRY(e->stage_mel.reserve(sizeof(float) * (size_t)(moff[n_clips] + 64)));
    Q3A_TRY(e->stage_max.reserve(sizeof(float) * (size_t)(n_clips + 1)));
    Prof prof{e, st};
    // the mel kernels of a part run right before its stem (after the part's upload has landed)
    PartHooks mine;
    if (hk) { mine.in_split = hk->in_split; mine.out_split = hk->out_split; mine.after_out = hk->after_out; mine.tail_layers = hk->tail_layers; }
    mine.before_stem = [&](int part, int ua, int ub) -> int {
        if (hk && hk->before_stem) Q3A_TRY(hk->before_stem(part, ua, ub));
        if (ub <= ua) return Q3A_OK;
        const int sp = prof.begin(SP_MEL);
        Q3A_TRY(mel_launch(fe, pcm_dev, pcm_off + ua, n_samples + ua, ub - ua, e->stage_mel.as<float>(), moff.data() + ua,
                           e->stage_max.as<float>() + ua, 0, st));
        prof.end(sp);
        return Q3A_OK;
    };
    return encode_dispatch(e, e->stage_mel.as<float>(), moff.data(), e->stage_max.as<float>(), nf.data(), n_clips, mode,
                           out_dev, tok_off, st, &mine);
}

extern "C" int q3a_encode_pcm_device(q3a_frontend* fe, q3a_encoder* e, const float* pcm_dev, const int64_t* pcm_off,
                                     const int* n_samples, int n_clips, int mode, float* out_dev, int64_t* tok_off,
                                     void* stream) {
    if (!e || !fe) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("encode: unknown mode"); return Q3A_ERR_ARG; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm_dev || !pcm_off || !n_samples || !out_dev))) { set_error("encode_pcm_device: null argument"); return Q3A_ERR_ARG; }
    if ((reinterpret_cast<uintptr_t>(pcm_dev) & 15u) || (reinterpret_cast<uintptr_t>(out_dev) & 15u)) {
        set_error("encode_pcm_device: pcm_dev and out_dev must be 16-byte aligned (bulk copies / TMA)"); return Q3A_ERR_ARG;
    }
    for (int i = 0; i < n_clips; ++i)
        if (pcm_off[i] < 0 || (pcm_off[i] & 3)) { set_error("encode_pcm_device: pcm_off entries must be non-negative multiples of 4"); return Q3A_ERR_ARG; }
    Q3A_TRY(use_device(e->device));
    g_launch_counter = 0;
    prof_reset(e);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    Q3A_TRY(pcm_core(fe, e, pcm_dev, pcm_off, n_samples, n_clips, mode, out_dev, tok_off, st));
    return prof_collect(e, st);
}

extern "C" int q3a_encode_pcm_batch(q3a_frontend* fe, q3a_encoder* e, const float* const* pcm, const int* n_samples,
                                    int n_clips, int mode, float* const* out, int* n_tokens) {
    if (!e || !fe) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("encode: unknown mode"); return Q3A_ERR_ARG; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm || !n_samples || !out))) { set_error("encode_pcm_batch: null argument"); return Q3A_ERR_ARG; }
    for (int u = 0; u < n_clips; ++u)
        if (n_samples[u] < 0 || (n_samples[u] > 0 && !pcm[u]) || (n_samples[u] >= 160 && !out[u])) {
            set_error("encode_pcm_batch: bad clip " + std::to_string(u)); return Q3A_ERR_ARG;
        }
    Q3A_TRY(use_device(e->device));
    g_launch_counter = 0;
    prof_reset(e);
    const int width = mode == Q3A_MODE_CONV_ONLY ? e->hp.d_model : e->hp.out_dim;
    std::vector<int> nf(n_clips);
    for (int i = 0; i < n_clips; ++i) nf[i] = n_samples[i] / 160;
    Q3A_TRY(run_host_passes(e, reinterpret_cast<const void* const*>(pcm), n_samples, 4, 64, nf.data(), n_clips, mode, width, out, n_tokens,
                            [&](int u0, int nu, const void* in_dev, const int64_t* in_off, float* out_dev, const PartHooks* hk) {
                                return pcm_core(fe, e, static_cast<const float*>(in_dev), in_off, n_samples + u0, nu, mode, out_dev, nullptr, e->stream, hk);
                            }));
    return prof_collect(e, e->stream);
}

// Asynchronous twin of q3a_encode_pcm_batch: enqueues the whole call (uploads, kernels, downloads) on the encoder's three
// streams and returns a ticket; q3a_encode_wait blocks until that call -- and every earlier one -- has delivered its
// features.  Two calls can be in flight (the staging buffers are double-buffered across calls): while call k runs its last
// layers and downloads, call k+1 is already uploading and running its stem behind it on the compute stream.  pcm[i] and
// out[i] must stay valid and untouched until the wait returns; n_tokens is filled before submit returns.
extern "C" int q3a_encode_pcm_submit(q3a_frontend* fe, q3a_encoder* e, const float* const* pcm, const int* n_samples,
                                     int n_clips, int mode, float* const* out, int* n_tokens, int64_t* ticket) {
    if (!e || !fe) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (!ticket) { set_error("encode_pcm_submit: ticket is null"); return Q3A_ERR_ARG; }
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("encode: unknown mode"); return Q3A_ERR_ARG; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm || !n_samples || !out))) { set_error("encode_pcm_submit: null argument"); return Q3A_ERR_ARG; }
    for (int u = 0; u < n_clips; ++u)
        if (n_samples[u] < 0 || (n_samples[u] > 0 && !pcm[u]) || (n_samples[u] >= 160 && !out[u])) {
            set_error("encode_pcm_submit: bad clip " + std::to_string(u)); return Q3A_ERR_ARG;
        }
    Q3A_TRY(use_device(e->device));
    if (e->profiling) { set_error("encode_pcm_submit: per-stage profiling needs the synchronous call"); return Q3A_ERR_STATE; }
    g_launch_counter = 0;
    prof_reset(e);
    const int width = mode == Q3A_MODE_CONV_ONLY ? e->hp.d_model : e->hp.out_dim;
    std::vector<int> nf(n_clips);
    for (int i = 0; i < n_clips; ++i) nf[i] = n_samples[i] / 160;
    // the mel staging of pcm_core is single-buffered, but it is only touched by kernels on the compute stream, which
    // orders consecutive calls
    Q3A_TRY(run_host_passes(e, reinterpret_cast<const void* const*>(pcm), n_samples, 4, 64, nf.data(), n_clips, mode, width, out, n_tokens,
                            [&](int u0, int nu, const void* in_dev, const int64_t* in_off, float* out_dev, const PartHooks* hk) {
                                return pcm_core(fe, e, static_cast<const float*>(in_dev), in_off, n_samples + u0, nu, mode, out_dev, nullptr, e->stream, hk);
                            }, false));
    e->tm.total_launches = g_launch_counter;
    *ticket = e->ticket_seq - 1;
    return Q3A_OK;
}

extern "C" int q3a_encode_wait(q3a_encoder* e, int64_t ticket) {
    if (!e) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (ticket < 0 || ticket >= e->ticket_seq) { set_error("encode_wait: unknown ticket"); return Q3A_ERR_ARG; }
    Q3A_TRY(use_device(e->device));
    if (e->ticket_seq - ticket > q3a_encoder::kTickets) return Q3A_OK;     // so old that its event slot was reused: long done
    Q3A_CUDA(cudaEventSynchronize(e->ev_done[ticket % q3a_encoder::kTickets]));
    return Q3A_OK;
}

// int16 ingest: load_wav's conversion (mel_spectrogram.cpp:195-210) on the device, so the upload carries 2 bytes per
// sample instead of 4.  One launch per upload part: frame f of the packed mono buffer = average of the n_channels
// interleaved samples (float sum / n_channels, the reference's operation order) / 32768.
__global__ void __launch_bounds__(256)
pcm16_to_f32_kernel(const int16_t* __restrict__ in, float* __restrict__ out, long long n_frames, int n_channels) {
    pdl_trigger();
    pdl_wait();
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x; f < n_frames; f += stride) {
        float v;
        if (n_channels == 1) {
            v = (float)in[f] / 32768.0f;
        } else {
            float sum = 0.0f;
            for (int c = 0; c < n_channels; ++c) sum += (float)in[f * n_channels + c];
            v = (sum / (float)n_channels) / 32768.0f;
        }
        out[f] = v;
    }
}

extern "C" int q3a_encode_pcm16_batch(q3a_frontend* fe, q3a_encoder* e, const int16_t* const* pcm, const int* n_samples,
                                      int n_channels, int n_clips, int mode, float* const* out, int* n_tokens) {
    if (!e || !fe) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("encode: unknown mode"); return Q3A_ERR_ARG; }
    if (n_channels < 1 || n_channels > 8) { set_error("encode_pcm16_batch: n_channels must be 1..8"); return Q3A_ERR_ARG; }
    if (n_clips < 0 || (n_clips > 0 && (!pcm || !n_samples || !out))) { set_error("encode_pcm16_batch: null argument"); return Q3A_ERR_ARG; }
    for (int u = 0; u < n_clips; ++u)
        if (n_samples[u] < 0 || (n_samples[u] > 0 && !pcm[u]) || (n_samples[u] >= 160 && !out[u])) {
            set_error("encode_pcm16_batch: bad clip " + std::to_string(u)); return Q3A_ERR_ARG;
        }
    Q3A_TRY(use_device(e->device));
    g_launch_counter = 0;
    prof_reset(e);
    const int width = mode == Q3A_MODE_CONV_ONLY ? e->hp.d_model : e->hp.out_dim;
    std::vector<int> nf(n_clips), elems(n_clips);
    for (int i = 0; i < n_clips; ++i) { nf[i] = n_samples[i] / 160; elems[i] = n_samples[i] * n_channels; }
    // clips are packed at multiples of 64 frames, i.e. 64 * n_channels int16 elements: element offset = n_channels * frame offset
    Q3A_TRY(run_host_passes(e, reinterpret_cast<const void* const*>(pcm), elems.data(), 2, 64ll * n_channels, nf.data(), n_clips, mode,
                            width, out, n_tokens,
                            [&](int u0, int nu, const void* in_dev, const int64_t* in_off, float* out_dev, const PartHooks* hk) -> int {
                                std::vector<int64_t> foff(nu + 1);
                                for (int i = 0; i <= nu; ++i) foff[i] = in_off[i] / n_channels;
                                Q3A_TRY(e->stage_pcm.reserve(sizeof(float) * (size_t)(foff[nu] + 64)));
                                float* mono = e->stage_pcm.as<float>();
                                PartHooks mine = *hk;
                                mine.before_stem = [&](int part, int ua, int ub) -> int {
                                    if (hk->before_stem) Q3A_TRY(hk->before_stem(part, ua, ub));
                                    const long long f0 = foff[ua], n = foff[ub] - foff[ua];
                                    if (n <= 0) return Q3A_OK;
                                    const int blocks = (int)std::min<long long>((n + 256 * 8 - 1) / (256 * 8), 148 * 8);
                                    Q3A_CUDA(launch_pdl(pcm16_to_f32_kernel, dim3(blocks), dim3(256), 0, e->stream,
                                                        static_cast<const int16_t*>(in_dev) + f0 * n_channels, mono + f0, n, n_channels));
                                    ++g_launch_counter;
                                    return Q3A_OK;
                                };
                                return pcm_core(fe, e, mono, foff.data(), n_samples + u0, nu, mode, out_dev, nullptr, e->stream, &mine);
                            }));
    return prof_collect(e, e->stream);
}

// Test / inspection hook for the planner above: which pass, upload part and download part every utterance of a
// q3a_encode_pcm_batch call lands in.  Host only.
extern "C" int q3a_plan_host_call(const int* n_samples, int n_clips, int mode, int out_width, int* pass_of, int* in_part_of,
                                  int* out_part_of) {
    if (n_clips < 0 || (n_clips > 0 && (!n_samples || !pass_of || !in_part_of || !out_part_of)) || out_width <= 0) {
        set_error("plan_host_call: bad argument"); return Q3A_ERR_ARG;
    }
    if (mode < 0 || mode > Q3A_MODE_CONV_ONLY) { set_error("plan_host_call: unknown mode"); return Q3A_ERR_ARG; }
    std::vector<int> nf(n_clips);
    for (int i = 0; i < n_clips; ++i) {
        if (n_samples[i] < 0) { set_error("plan_host_call: negative n_samples"); return Q3A_ERR_ARG; }
        nf[i] = n_samples[i] / 160;
    }
    HostPlan plan;
    plan_host_call(n_samples, 4, 64, nf.data(), n_clips, mode, out_width, true, plan);
    for (size_t p = 0; p + 1 < plan.bounds.size(); ++p) {
        const int u0 = plan.bounds[p];
        for (int i = u0; i < plan.bounds[p + 1]; ++i) pass_of[i] = (int)p;
        for (size_t k = 0; k + 1 < plan.in_split[p].size(); ++k)
            for (int i = plan.in_split[p][k]; i < plan.in_split[p][k + 1]; ++i) in_part_of[u0 + i] = (int)k;
        for (size_t k = 0; k + 1 < plan.out_split[p].size(); ++k)
            for (int i = plan.out_split[p][k]; i < plan.out_split[p][k + 1]; ++i) out_part_of[u0 + i] = (int)k;
    }
    return Q3A_OK;
}

// Longest-processing-time-first assignment.  Cost model: per token ~0.71 GFLOP of dense work plus the
// attention term 4*64*H*L*len^2 (ASR constants); only the ratio between the two terms matters.
extern "C" int q3a_plan_shards(const int* n_frames, int n_utts, int mode, int n_shards, int* shard_of) {
    if (n_utts < 0 || n_shards <= 0 || (n_utts > 0 && (!n_frames || !shard_of))) { set_error("plan_shards: bad argument"); return Q3A_ERR_ARG; }
    std::vector<double> cost(n_utts);
    for (int i = 0; i < n_utts; ++i) {
        const double T = encoder_n_tokens(n_frames[i], mode);
        double seg2 = 0.0;
        if (mode == Q3A_MODE_ALIGNER) {
            const int full = (int)T / kWindowTokens, rem = (int)T % kWindowTokens;
            seg2 = (double)full * kWindowTokens * kWindowTokens + (double)rem * rem;
        } else seg2 = T * T;
        cost[i] = 0.71e9 * T + 4.0 * 64 * 14 * 18 * seg2;
    }
    std::vector<int> order(n_utts);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
    std::vector<double> load(n_shards, 0.0);
    for (int i : order) {
        int best = 0;
        for (int s = 1; s < n_shards; ++s) if (load[s] < load[best]) best = s;
        shard_of[i] = best;
        load[best] += cost[i];
    }
    return Q3A_OK;
}
