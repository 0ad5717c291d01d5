// decode_sm100.cu -- q_len == 1 split-KV attention ("flash decoding") for sm_100a, plus the LSE combine.
//
// Replaces /root/reference/csrc/mfa/decode.cuh:523-662 (split-KV kernel) and :665-755 (combine).
// Not a port: the reference runs one CTA per *query* head with single-buffered cp.async tiles and
// CUDA-core dot products with a 5-step shuffle reduction per key row; here
//   * one CTA owns (batch, KV head, split) and serves the whole GQA group, so K/V are read from HBM
//     exactly once per KV head (the reference re-reads them `group` times, decode.cuh:48);
//   * a dedicated producer warp streams 64-row K and V tiles with TMA (cp.async.bulk.tensor, 128B
//     swizzle, per-page boxes gathered through block_table) into a multi-stage mbarrier ring, so
//     ~100 KB per CTA are in flight independent of the consumers' registers;
//   * four consumer warps each take 16 rows of a tile: S = Q K^T and O += P V run on the tensor cores
//     (mma.sync m16n8k16, the q heads of the group padded to M = 16), operands fetched with
//     ldmatrix from the swizzled tiles (conflict-free), online softmax in the exp2 domain in fp32;
//   * the four per-warp partial (max, sum, O) states are merged through shared memory, and the CTA
//     writes either the final output (1 split) or a normalised fp32 partial + LSE for the combine.
// The split partition is the reference's (decode.cuh:21-30): blocks of 64 rows,
// blocks_per_split = ceil(ceil(len/64)/num_splits); empty splits write LSE = -inf.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <type_traits>

#include "kparams.h"
#include "merge.cuh"
#include "ptx.cuh"

namespace mfa {

namespace {

template <typename T>
struct Pack2;
template <>
struct Pack2<__half> {
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    static __device__ __forceinline__ void store(__half* p, float v) { *p = __float2half_rn(v); }
};
template <>
struct Pack2<__nv_bfloat16> {
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

template <typename T>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma_16816<__half>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma_16816<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                                          uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}

// two e4m3 bytes (low 16 bits of x) -> packed f16x2 (exact: e4m3 is a subset of fp16)
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint32_t x) {
    uint32_t r;
    asm("{\n\t.reg .b16 t;\n\tcvt.u16.u32 t, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, t;\n\t}" : "=r"(r) : "r"(x));
    return r;
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}

// Byte offset of 16-byte chunk `chunk` (0 .. DPAD/8) of row `row` inside a [DPAD/64][64 rows][128 B] tile
// written by TMA with CU_TENSOR_MAP_SWIZZLE_128B (tile base 1024-byte aligned).
template <int ROWS>
__device__ __forceinline__ uint32_t swz_off(int row, int chunk) {
    return static_cast<uint32_t>((chunk >> 3) * (ROWS * 128) + row * 128 + (((chunk & 7) ^ (row & 7)) << 4));
}

#ifndef MFA_DECODE_STAGES
// Ring depth at head_dim 128/256.  Measured, not derived (scripts/decode_stages.py, B=24 sweep): 2 stages (64 KB per
// CTA) beat 3 stages by 1-3 % at every KV length (7045 -> 7262 GB/s at KV=8192, 4819 -> 4896 at KV=512).  Residency
// stays at TWO CTAs per SM either way -- the kernel uses 162 registers at head_dim 128 -- and forcing three with
// __launch_bounds__(160, 3) (128 registers, -DMFA_DECODE_MINBLOCKS=3) is 13 % slower (6313 GB/s at KV=8192).
#define MFA_DECODE_STAGES 2
#endif
template <int DPAD, bool KV8 = false>
struct DecodeCfg {
    static constexpr int kTile = kDecodeTileRows;
    static constexpr int kEB = KV8 ? 1 : 2;                      // bytes per cache element (e4m3 cache: 1)
    static constexpr int kSub = DPAD * kEB / 128;                // 128-byte wide sub-tiles per row
    static constexpr int kBoxElems = 128 / kEB;                  // elements per 128-byte box row
    static constexpr int kTileBytes = kTile * DPAD * kEB;        // one of K or V
    static constexpr int kStageBytes = 2 * kTileBytes;           // K then V
    static constexpr int kStages = (DPAD * kEB == 128) ? 2 * MFA_DECODE_STAGES : MFA_DECODE_STAGES;   // >= 32 KB per stage pair
    static constexpr int kRingBytes = kStages * kStageBytes;
    // cross-warp merge staging.  It has its own shared memory (not the ring's): the CTA is persistent, and while the
    // consumers merge item n the producer warp is already filling the ring with the first tiles of item n + 1
    static constexpr int kMergeBytes = 4 * 16 * DPAD * 4 + 512;
    static constexpr int kBarOffset = kRingBytes + kMergeBytes;
    static constexpr int kSmemBytes = kBarOffset + 1024 /*align slack*/ + 256 /*barriers*/;
};

// (tuning aid: -DMFA_DECODE_MINBLOCKS=3 adds a min-blocks launch bound at head_dim <= 128.  Even a bound of 1 changes
//  ptxas' allocation -- 180 instead of 162 registers -- and costs 28 % at KV=8192 (436 vs 340 us, profiles r01h vs r01g),
//  so the default build states NO min-blocks bound at all.)
#ifdef MFA_DECODE_MINBLOCKS
#define MFA_DECODE_BOUNDS(DP) __launch_bounds__(kDecodeThreads, ((DP) <= 128) ? MFA_DECODE_MINBLOCKS : 1)
#else
#define MFA_DECODE_BOUNDS(DP) __launch_bounds__(kDecodeThreads)
#endif
// KV8: the cache holds e4m3 bytes -- half the HBM bytes per key.  The tiles are converted to fp16 in
// registers (cvt.rn.f16x2.e4m3x2, exact) right before the MMAs, which then always run in fp16 (q is converted once per
// CTA; a bf16 q beyond the fp16 range saturates); the per-KV-head dequantisation scales are folded into the softmax
// scale (K) and into the final normalisation (V).  The MMA k / n indices are PERMUTED relative to head_dim so that one
// ldmatrix of the 1-byte tile feeds whole fragments: a thread's four bytes of a K row are d = 16c + 4tq + {0,1,2,3}
// (k-step c), and the transposed V load yields the kv-row pairs of an even and an odd head_dim column.
template <typename T, int DPAD, bool KV8>
__global__ void MFA_DECODE_BOUNDS(DPAD) decode_splitkv_kernel(const __grid_constant__ CUtensorMap tmK,
                                                                        const __grid_constant__ CUtensorMap tmV,
                                                                        const __grid_constant__ DecodeKParams p) {
    using Cfg = DecodeCfg<DPAD, KV8>;
    using TM = typename std::conditional<KV8, __half, T>::type;   // MMA operand type
    constexpr int kTile = Cfg::kTile;
    constexpr int kStages = Cfg::kStages;
    constexpr int KS = DPAD / 16;   // k-steps of QK^T
    constexpr int NT = DPAD / 8;    // n-tiles of PV
    constexpr int kSub = Cfg::kSub;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kBarOffset);
    uint64_t* empty_bar = full_bar + kStages;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // PERSISTENT: the grid is (resident CTAs per SM) x (SMs), or one CTA per work item when there are fewer items.  A work
    // item is (split, KV head x head tile, batch) -- what used to be one CTA -- and every CTA walks the items blockIdx.x,
    // blockIdx.x + gridDim.x, ... with both of its roles in the same order.  The barriers and the ring live across
    // items (phases come from a running tile counter), so the producer streams the next item's first tiles while the
    // consumers are still merging the current one: the ~3 us ramp of a fresh CTA (barrier init, tensor-map fetch, first
    // HBM round trip) is paid once per CTA instead of once per item -- a third of the kernel at KV = 512.
    struct Item {
        int split, kh, htile, b, len, first_key, t0, ntiles;
    };
    const int ny = p.kv_heads * p.head_tiles;
    const int total_items = p.num_splits * ny * p.batch;
    auto get_item = [&](int item) {
        Item w;
        w.split = item % p.num_splits;
        const int rest = item / p.num_splits;
        const int y = rest % ny;
        w.b = rest / ny;
        w.kh = y / p.head_tiles;
        w.htile = y % p.head_tiles;
        w.len = p.seqlens_k ? min(p.seqlens_k[w.b], p.seqlen_k) : p.seqlen_k;
        const int n_blocks = (w.len + kTile - 1) / kTile;
        // sliding window: the single query sits at position len - 1 and sees keys [first_key, len);
        // tiles in front of the window are not loaded and the splits partition only the tiles of the window
        w.first_key = (p.window_left >= 0) ? max(w.len - 1 - p.window_left, 0) : 0;
        const int tile_lo = w.first_key / kTile;
        const int per_split = (n_blocks - tile_lo + p.num_splits - 1) / p.num_splits;   // decode.cuh:27
        w.t0 = tile_lo + w.split * per_split;
        w.ntiles = max(min(n_blocks, w.t0 + per_split) - w.t0, 0);
        return w;
    };

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 4);
        }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == 4) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            prefetch_tensormap(&tmK);
            prefetch_tensormap(&tmV);
            const int chunk_bytes = p.box_rows * 128;   // one box = box_rows x 128 bytes
            int it = 0;   // ring position, runs across items
            for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
                const Item w = get_item(item);
                const int32_t* bt = p.block_table ? p.block_table + (int64_t)w.b * p.block_table_batch_stride : nullptr;
                for (int t = 0; t < w.ntiles; ++t, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    const int row0 = (w.t0 + t) * kTile;
                    const int rows_valid = min(w.len - row0, kTile);
                    const int nchunks = (rows_valid + p.box_rows - 1) / p.box_rows;
                    mbar_arrive_expect_tx(&full_bar[s], static_cast<uint32_t>(nchunks * chunk_bytes * kSub * 2));
                    uint8_t* kdst = smem + s * Cfg::kStageBytes;
                    uint8_t* vdst = kdst + Cfg::kTileBytes;
                    for (int c = 0; c < nchunks; ++c) {
                        const int row = row0 + c * p.box_rows;
                        int blk, in_page;
                        if (bt) {
                            const int pg = row / p.page;
                            in_page = row - pg * p.page;
                            blk = bt[pg];
                        } else {
                            in_page = row;
                            blk = w.b;
                        }
#pragma unroll
                        for (int hf = 0; hf < kSub; ++hf) {
                            tma_load_4d(kdst + hf * (kTile * 128) + c * chunk_bytes, &tmK, &full_bar[s], hf * Cfg::kBoxElems,
                                        in_page, w.kh, blk);
                            tma_load_4d(vdst + hf * (kTile * 128) + c * chunk_bytes, &tmV, &full_bar[s], hf * Cfg::kBoxElems,
                                        in_page, w.kh, blk);
                        }
                    }
                }
            }
        }
        return;
    }

    // ---------------------------------------------------------------- consumers (warps 0..3)
    const int g0 = lane >> 2;        // head row of c0,c1 / a0,a2
    const int g1 = g0 + 8;           // head row of c2,c3 / a1,a3
    const int tq = lane & 3;
    const int wrow = warp * 16;   // this warp's rows inside each tile
    float* stage_o = reinterpret_cast<float*>(smem + Cfg::kRingBytes);                    // [4 warps][16 heads][DPAD]
    float* stage_m = reinterpret_cast<float*>(smem + Cfg::kRingBytes + 4 * 16 * DPAD * 4);  // [4][16]
    float* stage_l = stage_m + 64;
    int it = 0;   // ring position, runs across items
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
    const Item w = get_item(item);
    const int split = w.split, kh = w.kh, b = w.b, len = w.len, first_key = w.first_key, t0 = w.t0, ntiles = w.ntiles;
    const int gvalid = min(p.group - w.htile * kDecodeMaxGroup, kDecodeMaxGroup);   // q heads served by this item
    const int head0 = kh * p.group + w.htile * kDecodeMaxGroup;
    const T* qbase = static_cast<const T*>(p.q) + (int64_t)b * p.q_batch_stride;

    uint32_t qa[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
        const T* q0 = qbase + (int64_t)(head0 + g0) * p.q_head_stride;
        const T* q1 = qbase + (int64_t)(head0 + g1) * p.q_head_stride;
        // 16-bit cache: MMA k index = head_dim index.  e4m3 cache: k-step ks covers d = 16 ks .. 16 ks + 15 with the
        // thread's k pairs (2tq, 2tq+1) / (2tq+8, 2tq+9) mapped to d = 16 ks + 4tq + {0,1} / {2,3} (see the K fragments)
        const int c = KV8 ? (ks * 16 + tq * 4) : (ks * 16 + tq * 2);
        const int c_hi = KV8 ? (c + 2) : (c + 8);
        qa[ks][0] = (g0 < gvalid && c < p.head_dim) ? *reinterpret_cast<const uint32_t*>(q0 + c) : 0u;
        qa[ks][1] = (g1 < gvalid && c < p.head_dim) ? *reinterpret_cast<const uint32_t*>(q1 + c) : 0u;
        qa[ks][2] = (g0 < gvalid && c_hi < p.head_dim) ? *reinterpret_cast<const uint32_t*>(q0 + c_hi) : 0u;
        qa[ks][3] = (g1 < gvalid && c_hi < p.head_dim) ? *reinterpret_cast<const uint32_t*>(q1 + c_hi) : 0u;
        if constexpr (KV8 && !std::is_same<T, __half>::value) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {   // bf16 pair -> fp16 pair (the MMAs of the e4m3 path run in fp16)
                const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&qa[ks][e]);
                const __half2 h = __floats2half2_rn(__low2float(b), __high2float(b));
                qa[ks][e] = *reinterpret_cast<const uint32_t*>(&h);
            }
        }
    }

    float o[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    // e4m3 cache: S = (q . k_bytes) * k_scale * softmax_scale -- fold the per-KV-head K scale into the exp2 scale
    const float k_descale = (KV8 && p.k_scale) ? p.k_scale[kh * p.kv_scale_stride] : 1.f;
    const float v_descale = (KV8 && p.v_scale) ? p.v_scale[kh * p.kv_scale_stride] : 1.f;
    const float c2 = p.scale_log2 * k_descale;

    for (int t = 0; t < ntiles; ++t, ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(&full_bar[s], ph);
        const uint32_t kaddr = smem_u32(smem + s * Cfg::kStageBytes);
        const uint32_t vaddr = kaddr + Cfg::kTileBytes;
        const int row0 = (t0 + t) * kTile;
        const int rows_valid = len - row0;   // >= 1

        if (rows_valid < wrow + 16) {
            // tail tile: rows >= len hold stale cache bytes (possibly NaN); zero this warp's V rows so that
            // P(=0) * V stays 0 (the reference zero-fills them with cp.async zfill, decode.cuh:243-249).
            uint8_t* vt = smem + s * Cfg::kStageBytes + Cfg::kTileBytes;
            for (int r = max(rows_valid - wrow, 0); r < 16; ++r) {
                const int row = wrow + r;
#pragma unroll
                for (int hf = 0; hf < kSub; ++hf)
                    *reinterpret_cast<uint32_t*>(vt + hf * (kTile * 128) + row * 128 + lane * 4) = 0u;
            }
            fence_proxy_async_smem();   // generic writes ordered before the next TMA refill of this stage
            __syncwarp();
        }

        // ---- S = Q K^T for this warp's 16 KV rows
        float sc[2][4];
#pragma unroll
        for (int i = 0; i < 2; ++i) sc[i][0] = sc[i][1] = sc[i][2] = sc[i][3] = 0.f;
        {
            const int mi = lane >> 3, r = lane & 7;
            const int row = wrow + (mi >> 1) * 8 + r;
            if constexpr (KV8) {
                // one ldmatrix.x4 = 16 kv rows x 32 bytes = two k-steps of both n-tiles; a thread's register of matrix
                // (n-tile, chunk) holds the row's bytes d = 16 chunk + 4tq + {0,1,2,3}: low pair -> b0, high pair -> b1
#pragma unroll
                for (int k2 = 0; k2 < KS / 2; ++k2) {
                    uint32_t kb[4];
                    ldsm_x4(kb, kaddr + swz_off<kTile>(row, 2 * k2 + (mi & 1)));
                    mma_16816<TM>(sc[0], qa[2 * k2], e4m3x2_to_f16x2(kb[0]), e4m3x2_to_f16x2(kb[0] >> 16));
                    mma_16816<TM>(sc[0], qa[2 * k2 + 1], e4m3x2_to_f16x2(kb[1]), e4m3x2_to_f16x2(kb[1] >> 16));
                    mma_16816<TM>(sc[1], qa[2 * k2], e4m3x2_to_f16x2(kb[2]), e4m3x2_to_f16x2(kb[2] >> 16));
                    mma_16816<TM>(sc[1], qa[2 * k2 + 1], e4m3x2_to_f16x2(kb[3]), e4m3x2_to_f16x2(kb[3] >> 16));
                }
            } else {
#pragma unroll
                for (int ks = 0; ks < KS; ++ks) {
                    uint32_t kb[4];
                    ldsm_x4(kb, kaddr + swz_off<kTile>(row, 2 * ks + (mi & 1)));
                    mma_16816<TM>(sc[0], qa[ks], kb[0], kb[1]);
                    mma_16816<TM>(sc[1], qa[ks], kb[2], kb[3]);
                }
            }
        }
        // ---- tail mask (decode.cuh:327-336)
        if (rows_valid < wrow + 16) {
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const int kv = wrow + nt * 8 + tq * 2;
                if (kv >= rows_valid) sc[nt][0] = sc[nt][2] = -INFINITY;
                if (kv + 1 >= rows_valid) sc[nt][1] = sc[nt][3] = -INFINITY;
            }
        }
        if (row0 < first_key) {   // head of the sliding window: keys in front of it
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const int kv = row0 + wrow + nt * 8 + tq * 2;
                if (kv < first_key) sc[nt][0] = sc[nt][2] = -INFINITY;
                if (kv + 1 < first_key) sc[nt][1] = sc[nt][3] = -INFINITY;
            }
        }
        // ---- online softmax (exp2 domain), two head rows per thread
        float tm0 = fmaxf(fmaxf(sc[0][0], sc[0][1]), fmaxf(sc[1][0], sc[1][1]));
        float tm1 = fmaxf(fmaxf(sc[0][2], sc[0][3]), fmaxf(sc[1][2], sc[1][3]));
        tm0 = fmaxf(tm0, __shfl_xor_sync(0xffffffffu, tm0, 1));
        tm0 = fmaxf(tm0, __shfl_xor_sync(0xffffffffu, tm0, 2));
        tm1 = fmaxf(tm1, __shfl_xor_sync(0xffffffffu, tm1, 1));
        tm1 = fmaxf(tm1, __shfl_xor_sync(0xffffffffu, tm1, 2));
        const float mn0 = fmaxf(m0, tm0), mn1 = fmaxf(m1, tm1);
        const float r0 = (mn0 == -INFINITY) ? 1.f : exp2f((m0 - mn0) * c2);
        const float r1 = (mn1 == -INFINITY) ? 1.f : exp2f((m1 - mn1) * c2);
        const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0 * c2;
        const float ms1 = (mn1 == -INFINITY) ? 0.f : mn1 * c2;
        m0 = mn0;
        m1 = mn1;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            sc[nt][0] = exp2f(sc[nt][0] * c2 - ms0);
            sc[nt][1] = exp2f(sc[nt][1] * c2 - ms0);
            sc[nt][2] = exp2f(sc[nt][2] * c2 - ms1);
            sc[nt][3] = exp2f(sc[nt][3] * c2 - ms1);
            ps0 += sc[nt][0] + sc[nt][1];
            ps1 += sc[nt][2] + sc[nt][3];
        }
        l0 = l0 * r0 + ps0;
        l1 = l1 * r1 + ps1;
        if (__any_sync(0xffffffffu, (r0 != 1.f) || (r1 != 1.f))) {
#pragma unroll
            for (int i = 0; i < NT; ++i) {
                o[i][0] *= r0;
                o[i][1] *= r0;
                o[i][2] *= r1;
                o[i][3] *= r1;
            }
        }
        // ---- O += P V
        uint32_t pa[4];
        pa[0] = Pack2<TM>::pack(sc[0][0], sc[0][1]);
        pa[1] = Pack2<TM>::pack(sc[0][2], sc[0][3]);
        pa[2] = Pack2<TM>::pack(sc[1][0], sc[1][1]);
        pa[3] = Pack2<TM>::pack(sc[1][2], sc[1][3]);
        {
            const int mi = lane >> 3, r = lane & 7;
            const int row = wrow + (mi & 1) * 8 + r;
            if constexpr (KV8) {
                // transposed load of the byte tile in 16-bit units: a thread's register of matrix (kv half, chunk) holds
                // {V[2tq][d], V[2tq][d+1], V[2tq+1][d], V[2tq+1][d+1]}, d = 16 chunk + 2 (lane / 4): bytes (0,2) are the
                // kv-row pair of the EVEN column, bytes (1,3) of the ODD one.  N-tile 4 n4 + 2 cc + e (chunk 2 n4 + cc,
                // parity e) therefore has MMA column n <-> d = 16 chunk + 2 n + e; the epilogue undoes the permutation.
#pragma unroll
                for (int n4 = 0; n4 < NT / 4; ++n4) {
                    uint32_t vb[4];
                    ldsm_x4_trans(vb, vaddr + swz_off<kTile>(row, 2 * n4 + (mi >> 1)));
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc) {
                        const uint32_t lo = vb[2 * cc], hi = vb[2 * cc + 1];   // kv rows 0-7 / 8-15 of the warp's 16
                        mma_16816<TM>(o[4 * n4 + 2 * cc], pa, e4m3x2_to_f16x2(prmt(lo, 0u, 0x4420u)), e4m3x2_to_f16x2(prmt(hi, 0u, 0x4420u)));
                        mma_16816<TM>(o[4 * n4 + 2 * cc + 1], pa, e4m3x2_to_f16x2(prmt(lo, 0u, 0x4431u)), e4m3x2_to_f16x2(prmt(hi, 0u, 0x4431u)));
                    }
                }
            } else {
#pragma unroll
                for (int n2 = 0; n2 < NT / 2; ++n2) {
                    uint32_t vb[4];
                    ldsm_x4_trans(vb, vaddr + swz_off<kTile>(row, 2 * n2 + (mi >> 1)));
                    mma_16816<TM>(o[2 * n2], pa, vb[0], vb[1]);
                    mma_16816<TM>(o[2 * n2 + 1], pa, vb[2], vb[3]);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    // ---------------------------------------------------------------- cross-warp merge
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

    named_bar_sync(1, 128);   // every consumer has finished reading the staging area of the previous item
    if (tq == 0) {
        stage_m[warp * 16 + g0] = m0;
        stage_m[warp * 16 + g1] = m1;
        stage_l[warp * 16 + g0] = l0;
        stage_l[warp * 16 + g1] = l1;
    }
    if constexpr (KV8) {
        // undo the n permutation of the e4m3 P.V: n-tiles (2c, 2c+1) hold the even / odd columns of d = 16c .. 16c+15
#pragma unroll
        for (int c = 0; c < NT / 2; ++c) {
            const int col = c * 16 + tq * 4;
            float* w0 = stage_o + (warp * 16 + g0) * DPAD + col;
            float* w1 = stage_o + (warp * 16 + g1) * DPAD + col;
            *reinterpret_cast<float4*>(w0) = make_float4(o[2 * c][0], o[2 * c + 1][0], o[2 * c][1], o[2 * c + 1][1]);
            *reinterpret_cast<float4*>(w1) = make_float4(o[2 * c][2], o[2 * c + 1][2], o[2 * c][3], o[2 * c + 1][3]);
        }
    } else {
#pragma unroll
        for (int i = 0; i < NT; ++i) {
            const int col = i * 8 + tq * 2;
            float* w0 = stage_o + (warp * 16 + g0) * DPAD + col;
            float* w1 = stage_o + (warp * 16 + g1) * DPAD + col;
            *reinterpret_cast<float2*>(w0) = make_float2(o[i][0], o[i][1]);
            *reinterpret_cast<float2*>(w1) = make_float2(o[i][2], o[i][3]);
        }
    }
    named_bar_sync(1, 128);

    const bool write_partial = (p.num_splits > 1);
    for (int idx = threadIdx.x; idx < gvalid * (DPAD / 4); idx += 128) {
        const int hr = idx / (DPAD / 4);
        const int col = (idx % (DPAD / 4)) * 4;
        float mw[4], mx = -INFINITY;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            mw[w] = stage_m[w * 16 + hr];
            mx = fmaxf(mx, mw[w]);
        }
        float L = 0.f, wgt[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            wgt[w] = (mw[w] == -INFINITY) ? 0.f : exp2f((mw[w] - mx) * c2);
            L += stage_l[w * 16 + hr] * wgt[w];
        }
        const float inv = (L > 0.f) ? v_descale / L : 0.f;   // (e4m3 cache: the V scale rides on the normalisation)
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const float4 v = *reinterpret_cast<const float4*>(stage_o + (w * 16 + hr) * DPAD + col);
            acc.x += v.x * wgt[w];
            acc.y += v.y * wgt[w];
            acc.z += v.z * wgt[w];
            acc.w += v.w * wgt[w];
        }
        acc.x *= inv;
        acc.y *= inv;
        acc.z *= inv;
        acc.w *= inv;
        const int head = head0 + hr;
        const float lse = (L > 0.f) ? mx * (p.scale * k_descale) + __logf(L) : -INFINITY;   // decode.cuh:649
        if (col < p.head_dim) {
            if (write_partial) {
                float* dst = p.o_part + (((int64_t)split * p.batch + b) * p.heads + head) * p.head_dim + col;
                *reinterpret_cast<float4*>(dst) = acc;
            } else if (p.partial_out) {
                float* dst = static_cast<float*>(p.o) + ((int64_t)b * p.heads + head) * p.head_dim + col;
                *reinterpret_cast<float4*>(dst) = acc;
            } else {
                T* dst = static_cast<T*>(p.o) + (int64_t)b * p.o_batch_stride + (int64_t)head * p.o_head_stride + col;
                uint2 pk;
                pk.x = Pack2<T>::pack(acc.x, acc.y);
                pk.y = Pack2<T>::pack(acc.z, acc.w);
                *reinterpret_cast<uint2*>(dst) = pk;
            }
        }
        if (col == 0) {
            if (write_partial) {
                p.lse_part[((int64_t)split * p.batch + b) * p.heads + head] = lse;
            } else if (p.lse) {
                p.lse[(int64_t)b * p.heads + head] = lse;
            }
        }
    }
    }   // items
}

// LSE merge of `num_splits` normalised partials (decode.cuh:665-755 restated with a max-subtracted
// log-sum-exp and no shared memory proportional to the split count).
// One CTA of 8 warps per output row (b, h): the splits are dealt over the warps (merge.cuh), warp 0 stores.
template <typename T, bool kPacked>
__global__ void __launch_bounds__(kMergeThreads) combine_kernel(const __grid_constant__ CombineKParams p) {
    __shared__ __align__(16) float red[kMergeWarps * 256];
    const int row = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int D = p.head_dim;
    float4 acc[2];
    float glse;
    block_merge_splits<kPacked>(p.o_part, p.lse_part, p.num_splits, p.rows, row, D, acc, glse, red);
    if (threadIdx.x >= 32) return;
    const int64_t out_off = (p.heads > 0) ? (int64_t)(row / p.heads) * p.out_batch_stride +
                                                (int64_t)(row % p.heads) * p.out_head_stride
                                          : (int64_t)row * p.out_row_stride;
    T* out = static_cast<T*>(p.out) + out_off;
    float* out32 = static_cast<float*>(p.out) + out_off;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int col = lane * 4 + 128 * i;
        if (col < D) {
            if (p.out_f32) {
                *reinterpret_cast<float4*>(out32 + col) = acc[i];
            } else {
                uint2 pk;
                pk.x = Pack2<T>::pack(acc[i].x, acc[i].y);
                pk.y = Pack2<T>::pack(acc[i].z, acc[i].w);
                *reinterpret_cast<uint2*>(out + col) = pk;
            }
        }
    }
    if (lane == 0 && p.lse_out) p.lse_out[row] = glse;
}

}  // namespace

// ---------------------------------------------------------------------------------- launchers
template <typename T, int DPAD, bool KV8 = false>
static cudaError_t launch_decode_t(const CUtensorMap& tmK, const CUtensorMap& tmV, const DecodeKParams& p,
                                   cudaStream_t stream) {
    using Cfg = DecodeCfg<DPAD, KV8>;
    auto kern = decode_splitkv_kernel<T, DPAD, KV8>;
    static bool attr_done[64] = {};   // per (T, DPAD) instantiation and device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    static int slots[64] = {};        // resident CTAs per SM x SMs, per (T, DPAD, KV8) instantiation and device
    if (dev < 64 && slots[dev] == 0) {
        int per_sm = 0, sms = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kDecodeThreads, Cfg::kSmemBytes) != cudaSuccess || per_sm < 1) per_sm = 1;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) sms = 148;
        slots[dev] = per_sm * sms;
    }
    const long long items = (long long)p.num_splits * p.kv_heads * p.head_tiles * p.batch;
    if (items <= 0 || items > 0x7fffffffLL) return cudaErrorInvalidValue;
    static const bool non_persistent = [] { const char* e = getenv("MFA_DECODE_NONPERSISTENT"); return e && e[0] == '1'; }();
    const long long cap = (dev < 64 && slots[dev] > 0 && !non_persistent) ? slots[dev] : items;
    dim3 grid((unsigned)(items < cap ? items : cap), 1, 1);
    kern<<<grid, kDecodeThreads, Cfg::kSmemBytes, stream>>>(tmK, tmV, p);
    return cudaGetLastError();
}

cudaError_t launch_decode(const CUtensorMap& tmK, const CUtensorMap& tmV, const DecodeKParams& p, bool is_bf16,
                          int dpad, cudaStream_t stream) {
#define MFA_DECODE_CASE(DP)                                                                  \
    case DP:                                                                                 \
        return is_bf16 ? launch_decode_t<__nv_bfloat16, DP>(tmK, tmV, p, stream)             \
                       : launch_decode_t<__half, DP>(tmK, tmV, p, stream);
    if (p.kv_fp8) {   // e4m3 cache: head_dim 128 / 256 (whole 128-byte rows)
        if (dpad == 128)
            return is_bf16 ? launch_decode_t<__nv_bfloat16, 128, true>(tmK, tmV, p, stream)
                           : launch_decode_t<__half, 128, true>(tmK, tmV, p, stream);
        if (dpad == 256)
            return is_bf16 ? launch_decode_t<__nv_bfloat16, 256, true>(tmK, tmV, p, stream)
                           : launch_decode_t<__half, 256, true>(tmK, tmV, p, stream);
        return cudaErrorInvalidValue;
    }
    switch (dpad) {
        MFA_DECODE_CASE(64)
        MFA_DECODE_CASE(128)
        MFA_DECODE_CASE(256)
        default:
            return cudaErrorInvalidValue;
    }
#undef MFA_DECODE_CASE
}

cudaError_t launch_combine(const CombineKParams& p, bool is_bf16, cudaStream_t stream) {
    dim3 grid(p.rows);
    if (p.packed) {
        if (is_bf16)
            combine_kernel<__nv_bfloat16, true><<<grid, kMergeThreads, 0, stream>>>(p);
        else
            combine_kernel<__half, true><<<grid, kMergeThreads, 0, stream>>>(p);
    } else {
        if (is_bf16)
            combine_kernel<__nv_bfloat16, false><<<grid, kMergeThreads, 0, stream>>>(p);
        else
            combine_kernel<__half, false><<<grid, kMergeThreads, 0, stream>>>(p);
    }
    return cudaGetLastError();
}

}  // namespace mfa
