// Varlen tcgen05 attention (attention.cu): host interface.
#pragma once
#include "common.cuh"

namespace q3a {

struct AttnTile {
    int q0;          // first token row of this query tile (v1: 128 rows) / work item (v2: up to 256 rows)
    int seg_start;   // tokens [seg_start, seg_end) attend to each other
    int seg_end;
    int pad;         // v2: query rows of the item (1..256); v1: keys per block for this segment (64 or 128)
};

// keys per block the v1 kernel uses for a segment of this length (a function of the segment alone: batch-invariant)
int attention_key_block(int seg_len);
// one tile per 128 queries of every segment; cu_seqlens has n_seg+1 ascending entries.  Returns the block sizes present:
// bit 0 = some segment uses 64-key blocks, bit 1 = some uses 128 (attention_launch's kt_mask)
int attention_tiles(const int* cu_seqlens, int n_seg, std::vector<AttnTile>& tiles);
inline int attention_tile_count(int seg_len) { return (seg_len + 127) / 128; }   // tiles attention_tiles() emits per segment
// 4 * head_dim * H * sum(len^2): QK^T and PV, 2 flops per MAC, no padding counted
double attention_flops(const int* cu_seqlens, int n_seg, int n_heads);
// qkv: [T][3*H*64] 16-bit (q | k | v column blocks), out: [T][H*64] 16-bit
int attention_launch(const void* qkv, void* out, int T, int n_heads, const AttnTile* tiles_dev, int n_tiles, int fmt,
                     int kt_mask, cudaStream_t stream);

// ---- second-generation kernel (attention2.cu): persistent CTAs, two query tiles per CTA -------------------------------
// work items: ceil(len/256) per segment, equal row counts rounded up to whole warps
void attention2_items(const int* cu_seqlens, int n_seg, std::vector<AttnTile>& items);
int attention2_item_count(int seg_len);
int attention2_launch(const void* qkv, void* out, int T, int n_heads, const AttnTile* items_dev, int n_items, int fmt,
                      cudaStream_t stream);
// ---- the aligner decoder's attention (attention_dec.cu): causal, grouped-query, head_dim 128 --------------------------
// qkv: [n][(H + 2 Hkv) * 128] 16-bit (q | k | v column blocks); out: [n][H * 128] 16-bit
int attention_dec_launch(const void* qkv, void* out, int n_tokens, int n_heads, int n_kv_heads, int fmt, cudaStream_t stream);
// which kernel the encoder uses: 1 (kDefaultAttnV2 in common.cuh) unless Q3A_ATTN_V2 is set in the environment (A/B runs)
int attention_version();

}  // namespace q3a
