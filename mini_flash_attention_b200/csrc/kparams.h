// kparams.h -- kernel-side parameter blocks (internal; the public POD is include/mfa_b200.h).
#pragma once
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

namespace mfa {

constexpr int kDecodeTileRows = 64;   // KV rows per pipeline stage == the reference's kBlockN (traits.h, flash.cu:89)
constexpr int kDecodeThreads = 160;   // 4 consumer warps + 1 TMA producer warp
constexpr int kDecodeMaxGroup = 16;   // q heads per CTA (one m16 MMA tile)

struct DecodeKParams {
    const void* q;             // (B, 1, Hq, D)
    void* o;                   // final output (out dtype), or fp32 when partial_out
    float* lse;                // optional (B, Hq)
    float* o_part;             // (splits, B, Hq, D) fp32 when splits > 1
    float* lse_part;           // (splits, B, Hq)
    int64_t q_batch_stride, q_head_stride;
    int64_t o_batch_stride, o_head_stride;
    const int32_t* seqlens_k;  // may be null -> seqlen_k
    const int32_t* block_table;  // may be null -> contiguous cache, block index == batch index
    int32_t block_table_batch_stride;
    int32_t max_blocks_per_seq;
    int32_t page;              // rows per page (contiguous cache: capacity rows)
    int32_t box_rows;          // rows per TMA box (divides 64 and the page size when paged)
    int32_t heads, kv_heads, group, batch, head_dim;
    int32_t seqlen_k;          // capacity
    int32_t num_splits;
    int32_t head_tiles;        // ceil(group / 16)
    int32_t partial_out;
    int32_t window_left;       // >= 0: sliding window, the query (position len - 1) sees keys >= len - 1 - window_left; < 0: all
    float scale, scale_log2;
    // e4m3 KV cache: 1-byte elements, dequantised value = byte value * scale[kv_head * kv_scale_stride]
    int32_t kv_fp8;
    int32_t kv_scale_stride;   // 1 = one scale per KV head, 0 = one scale for the tensor
    const float* k_scale;      // device; null => 1
    const float* v_scale;
};

struct CombineKParams {
    const float* o_part;       // (splits, rows, D) or packed (splits, rows, D+1) when packed != 0
    const float* lse_part;     // (splits, rows) (unused when packed)
    void* out;                 // (rows, D) out dtype, row stride out_row_stride
    float* lse_out;            // optional (rows)
    int64_t out_row_stride;    // elements; rows are (b*H + h) => caller guarantees uniform stride
    int64_t out_batch_stride, out_head_stride;  // used when heads > 0 (row = b*heads + h)
    int32_t heads;
    int32_t num_splits, rows, head_dim, packed;
    int32_t out_f32;           // write fp32 instead of the 16-bit dtype (partials of a KV-sharded decode)
};

constexpr int kXchgMaxWorld = 8;   // ranks of one NVSwitch domain

// exchange + merge kernel of the KV-sharded decode (xchg_sm100.cu)
struct XchgKParams {
    const float* o_part;       // this rank's split partials (num_splits, rows, D) fp32, normalised
    const float* lse_part;     // (num_splits, rows)
    float* peer_data[kXchgMaxWorld];      // every rank's gather buffer [2 parities][world][max_rows][D + 4] (own included)
    uint32_t* peer_flags[kXchgMaxWorld];  // every rank's flags [world][max_rows], epoch valued
    uint32_t* ctrs;            // local {status (!= 0 after a wait timed out), epoch counter, warps done, pad}
    void* out;                 // (batch, heads, D) out dtype through out_batch_stride / out_head_stride
    float* lse_out;            // optional (rows)
    int64_t out_batch_stride, out_head_stride;
    long long timeout_cycles;
    int32_t num_splits, world, rank, rows, max_rows, head_dim, heads;
};

constexpr int kPrefillBlockM = 128;   // query rows per Q tile (one TMEM lane per row)
constexpr int kPrefillBlockN = 128;   // KV rows per tile
constexpr int kPrefillThreads = 384;  // 2 softmax warpgroups + 1 warpgroup {TMA warp, MMA warp, 2 spare}

struct PrefillKParams {
    const void* q;                 // (Q goes through its tensor map; kept for diagnostics)
    int64_t q_batch_stride, q_head_stride, q_row_stride;
    void* o;
    int64_t o_batch_stride, o_head_stride, o_row_stride;
    const int32_t* cu_seqlens_q;   // null => fixed length
    const int32_t* cu_seqlens_k;
    const int32_t* block_table;    // null => not paged
    int32_t block_table_batch_stride;
    int32_t max_blocks_per_seq;
    int32_t page;                  // rows per page when paged
    int32_t kv_box_rows;           // rows per K/V TMA box (128, or gcd(page,128) when paged)
    int32_t heads, kv_heads, group, batch, head_dim;
    int32_t seqlen_q, seqlen_k;    // fixed lengths or max lengths
    int32_t is_causal;
    int32_t causal_br;             // causal mask aligned to the bottom-right corner: col <= row + (len_k - len_q)
    // Band mask (sliding window): key col visible iff row + shift - win_left <= col <= row + shift +
    // win_right, shift = causal_br ? len_k - len_q : 0; a bound < 0 is absent.  is_causal is folded into win_right = 0
    // by the host, so the kernels only look at has_hi / has_lo.
    int32_t has_hi, win_right;
    int32_t has_lo, win_left;
    int32_t num_m_blocks;          // ceil(seqlen_q / 256)
    float scale, scale_log2;
    float* lse;                    // optional natural-log LSE: (B, H, Sq) fixed / (H, total_q) packed; -inf for empty rows
    int64_t lse_batch_stride, lse_head_stride;   // elements; the row index is q_off + row (q_off = 0 when fixed)
    int32_t snake;                 // packed batches run persistent with odd rounds of the item walk reversed (prefill_sm100.cu)
    int32_t compact;               // ragged causal packed batch: the item list is the device-built table of the items that exist
    long long* trace;              // MFA_TRACE builds only: clock64() stamps of CTA (0,0,0), [role][iter][8]
};

}  // namespace mfa
