// prefill_sm100.cu -- FlashAttention forward (prefill) for sm_100a: TMA + mbarrier ring + tcgen05.mma with
// accumulators in tensor memory + exp2 online softmax with one thread per query row.
//
// Replaces /root/reference/csrc/mfa/prefill.cuh:711-803 (kernel) and flash.cu:11-34 (launch).  It is not
// a port of that sm_80 mma.sync/cp.async kernel (64x64 tiles, 4 warps, single-buffered K/V); the
// structure here is
//
//   work item = (256 query rows = two 128-row Q tiles, one head, one sequence); persistent CTAs (one per SM, 384
//   threads) walk the items with TMEM, barriers and the K/V ring kept alive across them
//   warps 0-3   softmax warpgroup of Q tile 0   (thread t <-> query row t <-> TMEM lane t)
//   warps 4-7   softmax warpgroup of Q tile 1
//   warp  8     TMA producer: Q tiles once, then K_0,V_0,K_1,V_1,... through a kStages-deep smem ring
//   warp  9     MMA issuer (one thread): S_i = Q_i K_j^T (SS) and O_i += P_i V_j (TS, P read from TMEM)
//   warp 10     TMEM allocator / deallocator
//
//   TMEM (512 columns x 128 lanes x 32 bit):  S_0 [0,128)  S_1 [128,256)  O_0 [256,256+D)  O_1 [256+D, 256+2D)
//   P_i (fp16/bf16, 64 columns) aliases the first half of S_i: a softmax thread has its whole S row in
//   registers before it overwrites it, and tcgen05.mma instructions execute in issue order, so
//   S_i(j+1) is only written after PV_i(j) has consumed P_i(j).
//
// The two Q tiles ping-pong on the tensor pipe: while warpgroup 0 runs the softmax of tile j, the MMA
// warp computes PV_1(j-1) and S_1(j) and vice versa.  O stays in TMEM for the whole KV loop; the
// rescale O *= 2^(m_old-m_new) is done lazily by the owning softmax thread only when the running max
// grew by more than 2^8 (the final normalisation uses the same stale max for the row sum, so the
// result is unchanged).
//
// Semantics kept from the reference: scale = 1/sqrt(D) folded into exp2 (api.cpp:99-100); causal mask
// top-left aligned, key col masked iff col > row (prefill.cuh:416) with tile skipping (:749-752);
// key-length mask col >= len (:416); rows with an empty sum produce 0 (:604-605); GQA
// kv_head = head / group (:49); varlen through cu_seqlens (:19-33); paged K/V through block_table --
// gathered per page box, not per 64-row tile (fixes quirk Q3 of SURVEY.md).
#include <cstdlib>
#include <type_traits>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kparams.h"
#include "ptx.cuh"

// Register split over the CTA's 384 threads (setmaxnreg): 256 softmax threads and 128 data-movement / MMA / allocator
// threads.  setmaxnreg only moves registers INSIDE the pool the CTA was launched with -- 168 registers x 384 threads =
// 64512, not the SM's 65536 -- so 256 * SOFTMAX + 128 * HELPER must be <= 64512 (216 / 72 is exact; a split that needs
// more blocks in setmaxnreg.inc for ever).
#ifndef MFA_REG_SOFTMAX
#define MFA_REG_SOFTMAX 216
#endif
#ifndef MFA_REG_HELPER
#define MFA_REG_HELPER 72
#endif


#ifdef MFA_TRACE
#define MFA_TR(role, it, ev)                                                                              \
    do {                                                                                                  \
        if (tr_on && tr_base + (it) < 64) p.trace[((role) * 64 + tr_base + (it)) * 8 + (ev)] = clock64();  \
    } while (0)
#define MFA_TRI(role, ev)                                                                                  \
    do {                                                                                                  \
        if (tr_on && tr_item < 64) p.trace[((role) * 64 + tr_item) * 8 + (ev)] = clock64();                \
    } while (0)
#else
#define MFA_TR(role, it, ev) do { } while (0)
#define MFA_TRI(role, ev) do { } while (0)
#endif

namespace mfa {

namespace {

template <typename T>
struct Elem;
template <>
struct Elem<__half> {
    static constexpr uint32_t kFmt = 0;
    static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
        __half2 h = __floats2half2_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};
template <>
struct Elem<__nv_bfloat16> {
    static constexpr uint32_t kFmt = 1;
    static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
        __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};

template <int DPAD>
struct PrefillCfg {
    static constexpr int kBM = kPrefillBlockM;
    static constexpr int kBN = kPrefillBlockN;
    static constexpr int kNQ = (DPAD <= 128) ? 2 : 1;            // Q tiles per CTA (TMEM budget: 128 S + DPAD O columns per tile of 512)
    static constexpr int kSoftmaxWarps = kNQ * 4;                // one thread per query row
    static constexpr int kThreads = (kSoftmaxWarps + 4) * 32;    // + one warpgroup {TMA warp, MMA warp, TMEM warp, spare}
    static constexpr int kSub = DPAD / 64;                       // 64-column (128-byte) sub-tiles per row
    static constexpr int kSubBytes = 128 * 128;                  // 128 rows x 128 B
    static constexpr int kQTileBytes = kSub * kSubBytes;
    static constexpr int kKVTileBytes = kSub * kSubBytes;
#ifndef MFA_STAGES128
#define MFA_STAGES128 4
#endif
    static constexpr int kStages = (DPAD == 64) ? 8 : (DPAD == 128 ? MFA_STAGES128 : (DPAD == 192 ? 3 : 2));
    static constexpr int kQBytes = kNQ * kQTileBytes;
    static constexpr int kRingBytes = kStages * kKVTileBytes;
    static constexpr int kBarOffset = kQBytes + kRingBytes;
    static constexpr int kOStageWarpBytes = 32 * 128;            // epilogue: 32 rows x 64 columns of 16-bit output per softmax warp
    // item table of a ragged packed batch (p.compact): int len_q[kSchedMaxBatch], int cum[kSchedMaxMBlocks + 1]
    static constexpr int kSchedMaxBatch = 64, kSchedMaxMBlocks = 255;
    static constexpr int kSchedOffset = kBarOffset + 512 + kSoftmaxWarps * kOStageWarpBytes;
    static constexpr int kSchedBytes = (kSchedMaxBatch + kSchedMaxMBlocks + 1) * 4;
    static constexpr int kSmemBytes = kSchedOffset + kSchedBytes + 1024;   // barriers, output staging, item table, alignment slack
    static constexpr uint32_t kTmemCols = 512;
    static constexpr uint32_t kTmemS = 0;                        // S_i at kTmemS + i*128
    static constexpr uint32_t kTmemO = 256;                      // O_i at kTmemO + i*DPAD
    static constexpr float kRescaleThreshold = 8.0f;             // log2 domain
};

// smem descriptors -------------------------------------------------------------------------------
// K-major operand (Q or K tile): [sub][128 rows][64 elems], SWIZZLE_128B.  One UMMA K-step = 16 elems = 32 B.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_addr, int kstep) {
    const uint32_t addr = tile_addr + (kstep >> 2) * (128 * 128) + (kstep & 3) * 32;
    return umma_smem_desc(addr, /*LBO*/ 16, /*SBO: 8 rows x 128 B*/ 1024, /*SW128*/ 2);
}
// MN-major operand (V tile as B of P.V): [sub(d)][128 kv rows][64 elems(d)], SWIZZLE_128B.
// One UMMA K-step = 16 kv rows = 2048 B; LBO = distance between 64-wide d sub-tiles, SBO = 8 rows.
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_addr, int kstep) {
    const uint32_t addr = tile_addr + kstep * (16 * 128);
    return umma_smem_desc(addr, /*LBO*/ 128 * 128, /*SBO*/ 1024, /*SW128*/ 2);
}

// Opaque copy of a loop-invariant value.  The MMA warp lives on 72-80 registers; without this the compiler hoists every
// `base + constant` TMEM / shared-memory address of the unrolled MMA batches out of the KV loop, spills them to local
// memory and reloads them (LDL, ~40 cycles of exposed latency) in front of every batch of tcgen05.mma -- seen in the SASS
// of the round-1 kernel.  Re-deriving them from one laundered base costs one uniform add each.
__device__ __forceinline__ uint32_t opaque(uint32_t x) {
    asm volatile("" : "+r"(x));
    return x;
}

// One work item = (m-block of kNQ*128 query rows, head, sequence).
struct PrefillItem {
    int h, b, kh, q_off, k_off, len_q, len_k, m0;
    // band mask: key col visible iff row + shift_lo <= col <= row + shift_hi (each bound only when p.has_lo / p.has_hi);
    // causal is shift_hi = 0 (top-left like the reference) or len_k - len_q (bottom-right)
    int shift_hi, shift_lo;
    int lo_i[2], n_i[2];   // Q tile i works on KV tiles lo_i[i] <= j < n_i[i] (lo_i == n_i == 0: none)
    int j_lo, n_max;       // the item's KV tile range: the union over its Q tiles
    bool valid;
    __device__ __forceinline__ bool active(int i, int j) const { return j >= lo_i[i] && j < n_i[i]; }
    __device__ __forceinline__ int count(int i) const { return n_i[i] - lo_i[i]; }
};

// `sched` (p.compact only): shared-memory table {len_q[64], cum[num_m_blocks + 1]} built at kernel start, where cum[i] is
// the number of (m-block, sequence) pairs that exist among the TOP i m-blocks.  `item` is then a position in the list of
// the items that exist, ordered by descending m-block (= descending cost under the causal mask), then sequence, then head.
template <int kNQ, bool kCompact>
__device__ __forceinline__ PrefillItem prefill_item(const PrefillKParams& p, int item, const int* sched) {
    PrefillItem w;
    int mraw, rest, mblk;
    if constexpr (kCompact) {
        const int* slen = sched;
        const int* cum = sched + 64;
        const int M = p.num_m_blocks;
        const int pair = item / p.heads;            // which (m-block, sequence) pair
        const int h = item - pair * p.heads;
        int lo = 0, hi = M;                          // largest i with cum[i] <= pair
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (cum[mid] <= pair) lo = mid; else hi = mid;
        }
        mblk = M - 1 - lo;
        int r = pair - cum[lo];                      // the r-th sequence that has this m-block
        const int need = mblk * (kNQ * kPrefillBlockM);
        int b = 0;
        for (; b < p.batch - 1; ++b) {
            if (slen[b] > need) {
                if (r == 0) break;
                --r;
            }
        }
        rest = b * p.heads + h;
    } else if (p.cu_seqlens_q) {
        // Packed (ragged) batches launch one CTA per item and the hardware hands them out in blockIdx order: the m-block
        // index is the SLOWEST digit and runs downwards, so the items that exist only in the long sequences -- under the
        // causal mask also the heaviest ones -- start first and the short tail items fill the gaps at the end
        // (longest-processing-time order without knowing the lengths on the host).
        const int per_m = p.heads * p.batch;
        mraw = item / per_m;
        rest = item - mraw * per_m;
        mblk = p.num_m_blocks - 1 - mraw;
    } else {
        mraw = item % p.num_m_blocks;
        rest = item / p.num_m_blocks;
        mblk = (p.has_hi && !p.has_lo) ? (p.num_m_blocks - 1 - mraw) : mraw;   // causal: heaviest (largest m) blocks first
    }
    w.h = rest % p.heads;
    w.b = rest / p.heads;
    w.kh = w.h / p.group;
    w.q_off = 0;
    w.k_off = 0;
    w.len_q = p.seqlen_q;
    w.len_k = p.seqlen_k;
    if (p.cu_seqlens_q) {
        w.q_off = p.cu_seqlens_q[w.b];
        w.len_q = p.cu_seqlens_q[w.b + 1] - w.q_off;
        w.k_off = p.cu_seqlens_k[w.b];
        w.len_k = p.cu_seqlens_k[w.b + 1] - w.k_off;
    }
    const int shift = p.causal_br ? (w.len_k - w.len_q) : 0;
    w.shift_hi = shift + p.win_right;
    w.shift_lo = shift - p.win_left;
    w.m0 = mblk * (kNQ * kPrefillBlockM);
    w.valid = w.m0 < w.len_q;   // prefill.cuh:724-726: tiles past the sequence's own length do nothing
    const int n_tiles_k = (w.len_k + kPrefillBlockN - 1) / kPrefillBlockN;
    w.n_i[0] = w.n_i[1] = 0;
    w.lo_i[0] = w.lo_i[1] = 0;
    w.j_lo = 0x7fffffff;
    w.n_max = 0;
#pragma unroll
    for (int i = 0; i < kNQ; ++i) {
        const int r0 = w.m0 + i * kPrefillBlockM;
        if (r0 < w.len_q) {
            // upper bound: the tile's last row r0+127 sees cols <= r0+127+shift_hi (none at all when that is negative);
            // lower bound: its first row r0 sees cols >= r0+shift_lo
            const int hi = p.has_hi ? min(n_tiles_k, max(0, (r0 + kPrefillBlockM + w.shift_hi + kPrefillBlockN - 1) / kPrefillBlockN)) : n_tiles_k;
            const int lo = p.has_lo ? max(0, r0 + w.shift_lo) / kPrefillBlockN : 0;
            if (lo < hi) {
                w.lo_i[i] = lo;
                w.n_i[i] = hi;
                w.j_lo = min(w.j_lo, lo);
                w.n_max = max(w.n_max, hi);
            }
        }
    }
    if (w.n_max == 0) w.j_lo = 0;
    return w;
}

// PERSISTENT: the grid is one CTA per SM (or one per work item when there are fewer); every CTA walks the items
// blockIdx.x, blockIdx.x + gridDim.x, ... with all of its roles in the same order.  TMEM, the barriers and the K/V ring
// live across items, so the producer prefetches the next item's Q and first K/V tiles while the current item's last
// tiles are still in the softmax, and the epilogue of item n overlaps the first S MMAs of item n+1.  (Measured before:
// ~7 k cycles from CTA start to the first S and ~3 k of epilogue per ~115 k-cycle item at S=4096, 22 % at S=1024.)
// All mbarrier parities therefore come from RUNNING counters: items done (q_full / q_empty) and tiles done per Q tile.
// kCompact: a separate instantiation for ragged causal packed batches (p.compact) -- the item decode of that form costs
// the MMA-issuing warp registers and instructions, and the fixed-length kernel is sensitive to both (measured: with the
// branch at run time the headline lost 3 %, causal 7 %).
// kMode 0: fixed-length batches and packed batches with one CTA per item -- the plain strided item walk, nothing else;
// kMode 1: near-uniform packed batches, persistent, snake-order walk; kMode 2 (kCompact): ragged causal packed batches.
// Compile-time, not p.snake / p.compact at run time: the strided walk of mode 0 is the code the headline was tuned on.
template <typename T, int DPAD, int kMode>
__global__ void __launch_bounds__(PrefillCfg<DPAD>::kThreads, 1)
    prefill_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ PrefillKParams p) {
    using Cfg = PrefillCfg<DPAD>;
    constexpr bool kCompact = (kMode == 2);
    constexpr int kNQ = Cfg::kNQ;
    constexpr int kStages = Cfg::kStages;
    constexpr int kBM = Cfg::kBM, kBN = Cfg::kBN;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_smem = smem;
    uint8_t* ring = smem + Cfg::kQBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kBarOffset);
    uint64_t* q_full = bars;                 // [2]  Q tile i of the current item landed (TMA bytes)
    uint64_t* s_full = bars + 2;             // [2]
    uint64_t* p_lo = bars + 4;               // [2]  first half of P_i (KV rows 0-63) written
    uint64_t* o_full = bars + 6;             // [2]
    uint64_t* tail_bar = bars + 8;           // [1]
    uint64_t* p_hi = bars + 9;               // [2]  second half of P_i
    uint64_t* q_empty = bars + 11;           // [2]  every S MMA of the item that reads Q tile i has completed
    uint64_t* kv_full = bars + 13;           // [kStages]
    uint64_t* kv_empty = kv_full + kStages;  // [kStages]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(kv_empty + kStages);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int* sched = reinterpret_cast<const int*>(smem + Cfg::kSchedOffset);
    if (kCompact && warp == 1) {
        // Ragged packed batch: build the table of the items that exist (see prefill_item).  One warp, before the set-up
        // barrier below; every role then derives the same item list from it.
        int* slen = reinterpret_cast<int*>(smem + Cfg::kSchedOffset);
        int* cum = slen + 64;
        for (int b = lane; b < 64; b += 32) slen[b] = (b < p.batch) ? (p.cu_seqlens_q[b + 1] - p.cu_seqlens_q[b]) : 0;
        __syncwarp();
        const int M = p.num_m_blocks;
        for (int i = lane; i < M; i += 32) {   // i-th m-block from the top
            const int need = (M - 1 - i) * (kNQ * kPrefillBlockM);
            int c = 0;
            for (int b = 0; b < p.batch; ++b) c += (slen[b] > need) ? 1 : 0;
            cum[i + 1] = c;
        }
        __syncwarp();
        if (lane == 0) {
            int a = 0;
            cum[0] = 0;
            for (int i = 0; i < M; ++i) {
                a += cum[i + 1];
                cum[i + 1] = a;
            }
        }
    }
    constexpr int kWarpTma = Cfg::kSoftmaxWarps, kWarpMma = kWarpTma + 1, kWarpTmem = kWarpTma + 2;
#ifdef MFA_TRACE
    const bool tr_cta = p.trace && blockIdx.x == 0 && lane == 0 && (warp == kWarpMma || warp == kWarpTma || (warp < kWarpTma && (warp & 3) == 0));
    bool tr_on = false;
    int tr_base = 0;   // running KV-step index of this role over the CTA's items (the trace covers the first 64 steps)
    int tr_item = 0;   // items done by this role
#endif

    // ---- one-time setup
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&q_full[i], 1);
            mbar_init(&s_full[i], 1);
            mbar_init(&p_lo[i], 128);
            mbar_init(&p_hi[i], 128);
            mbar_init(&o_full[i], 1);
            mbar_init(&q_empty[i], 1);
        }
        mbar_init(tail_bar, 1);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        fence_barrier_init();
    }
    if (warp == kWarpTmem) {
        tmem_alloc(tmem_slot, Cfg::kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // The CTA allocates ALL 512 TMEM columns (one CTA per SM), so the allocation can only start at column 0, lane 0:
    // every TMEM address below is a compile-time constant (no register, no address arithmetic in the 72-register MMA
    // warp).  Checked, not assumed.
    if (*tmem_slot != 0u) __trap();
    // (after the set-up barrier: a ragged packed batch counts its items in the table warp 1 just built)
    const int total_items = kCompact ? sched[64 + p.num_m_blocks] * p.heads : p.num_m_blocks * p.heads * p.batch;
    // Item walk of a persistent CTA: round k hands out items k * grid ... k * grid + grid - 1.  Packed (ragged) batches are
    // listed by descending cost, so odd rounds run backwards (p.snake): the CTA that drew the heaviest item of one round gets
    // the lightest of the next, which bounds the imbalance by the cost drop inside ONE round instead of the whole list.
    const int n_rounds = (total_items + (int)gridDim.x - 1) / (int)gridDim.x;
    auto item_at = [&](int k) -> int {   // total_items = this CTA has no item in round k
        if (k >= n_rounds) return total_items;
        const int pos = (kMode != 0 && (k & 1)) ? (int)gridDim.x - 1 - (int)blockIdx.x : (int)blockIdx.x;
        const int it = k * (int)gridDim.x + pos;
        return it < total_items ? it : total_items;
    };
    // (only the last round can leave a CTA without an item, so a walk may stop at the first miss)
    auto next_item = [&](int k_next, int item) -> int {
        if constexpr (kMode == 0) return item + (int)gridDim.x;
        else return item_at(k_next);
    };
    constexpr uint32_t tmem_base = 0u;

    // Register re-balancing: the launch gives every thread 168 registers (65536 / 384 threads); the
    // data-movement warpgroup hands most of its share to the two softmax warpgroups (whole S row in registers).
    // (Each setmaxnreg sits at the top of its own role branch so that ptxas sees the budget per region.)
    if (warp >= kWarpTma) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(MFA_REG_HELPER));
    if (warp == kWarpTma) {
        // ======================================================================= TMA producer
        // Warp-uniform control flow with one ELECTED lane issuing (see the MMA warp): under a divergent
        // `if (lane == 0)` ptxas cannot keep the TMA operands in uniform registers.
        if (elect_one()) {
            prefetch_tensormap(&tmQ);
            prefetch_tensormap(&tmK);
            prefetch_tensormap(&tmV);
        }
        uint32_t tail_phase = 0;
        int it = 0;                  // K/V ring position, runs across items
        uint32_t q_ld[2] = {0, 0};   // loads of Q tile i so far (parity base of q_full / q_empty)
        for (int k = 0, item = blockIdx.x; item < total_items; ++k, item = next_item(k, item)) {
            const PrefillItem w = prefill_item<kNQ, kCompact>(p, item, sched);
            if (!w.valid) continue;
            const int q_row_coord = p.cu_seqlens_q ? (w.q_off + w.m0) : w.m0;
            const int q_b_coord = p.cu_seqlens_q ? 0 : w.b;
#ifdef MFA_TRACE
            tr_on = tr_cta;
#endif
            MFA_TRI(3, 4);   // producer reaches the item
            // Q tiles of this item -- only the tiles that have KV work (n_i > 0).  A tile without work is never read by
            // an MMA, so loading it would leave a bulk copy in flight that nobody awaits (it could land after the CTA
            // exited, or be re-armed before it landed).  q_full / q_empty phases therefore count LOADS of tile i
            // (q_ld[i]), not items: every load is awaited by the MMA warp before it commits q_empty, and the producer
            // observes that commit before the next load, so neither barrier ever runs more than one phase ahead.
            // (Tried: Q tile 0, the item's first K tile, then Q tile 1 -- tile 0's buffer is free half a period before tile
            // 1's -- with the MMA warp awaiting each Q tile in front of its own first S: S=256 +3 %, causal -3 %; and the
            // per-tile wait bookkeeping alone cost the MMA warp 8 % on the short causal shapes.  Removed.)
#pragma unroll
            for (int i = 0; i < kNQ; ++i) {
                if (w.n_i[i] == 0) continue;
                if (q_ld[i] > 0) mbar_wait(&q_empty[i], (q_ld[i] - 1) & 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&q_full[i], Cfg::kQTileBytes);
#pragma unroll
                    for (int sb = 0; sb < Cfg::kSub; ++sb)
                        tma_load_4d(q_smem + i * Cfg::kQTileBytes + sb * Cfg::kSubBytes, &tmQ, &q_full[i], sb * 64,
                                    q_row_coord + i * kBM, w.h, q_b_coord);
                }
                __syncwarp();
                ++q_ld[i];
            }
            MFA_TRI(3, 5);   // Q loads issued
            const int32_t* bt = p.block_table ? p.block_table + (int64_t)w.b * p.block_table_batch_stride : nullptr;
            const bool packed_k = (p.cu_seqlens_k != nullptr) && (bt == nullptr);
            // rows >= len_k of a V tile must be zero (P is 0 there, but 0 * NaN = NaN).  For fixed-length
            // batches TMA's out-of-bounds fill does it; for packed / paged K/V the producer warp zeroes them.
            const bool need_tail_fix = (bt != nullptr) || packed_k;
            for (int j = w.j_lo; j < w.n_max; ++j) {
                const int row0 = j * kBN;
                const int rows_valid = min(w.len_k - row0, kBN);
#pragma unroll 1
                for (int kv = 0; kv < 2; ++kv, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1;
                    const CUtensorMap* tm = kv ? &tmV : &tmK;
                    uint8_t* dst = ring + s * Cfg::kKVTileBytes;
                    const bool fix = kv && need_tail_fix && rows_valid < kBN;
                    uint64_t* bar = fix ? tail_bar : &kv_full[s];
                    mbar_wait(&kv_empty[s], ph ^ 1);
                    if (elect_one()) {
                        if (bt) {
                            const int nchunks = (rows_valid + p.kv_box_rows - 1) / p.kv_box_rows;
                            const int chunk_bytes = p.kv_box_rows * 128;
                            mbar_arrive_expect_tx(bar, (uint32_t)(nchunks * chunk_bytes * Cfg::kSub));
                            for (int c = 0; c < nchunks; ++c) {
                                const int row = row0 + c * p.kv_box_rows;
                                const int pg = row / p.page;
                                const int in_page = row - pg * p.page;
                                const int blk = bt[pg];
#pragma unroll
                                for (int sb = 0; sb < Cfg::kSub; ++sb)
                                    tma_load_4d(dst + sb * Cfg::kSubBytes + c * chunk_bytes, tm, bar, sb * 64, in_page,
                                                w.kh, blk);
                            }
                        } else {
                            mbar_arrive_expect_tx(bar, Cfg::kKVTileBytes);
#pragma unroll
                            for (int sb = 0; sb < Cfg::kSub; ++sb)
                                tma_load_4d(dst + sb * Cfg::kSubBytes, tm, bar, sb * 64, (packed_k ? w.k_off : 0) + row0,
                                            w.kh, packed_k ? 0 : w.b);
                        }
                    }
                    __syncwarp();
                    if (j == w.j_lo && kv == 0) MFA_TRI(3, 6);   // first K load of the item issued
                    if (fix) {
                        // wait for the bytes, zero rows [rows_valid, 128) of every sub-tile, publish to the MMA proxy
                        mbar_wait(tail_bar, tail_phase);
                        tail_phase ^= 1;
                        for (int r = rows_valid; r < kBN; ++r) {
#pragma unroll
                            for (int sb = 0; sb < Cfg::kSub; ++sb)
                                *reinterpret_cast<uint32_t*>(dst + sb * Cfg::kSubBytes + r * 128 + lane * 4) = 0u;
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&kv_full[s]);
                    }
                }
            }
            MFA_TRI(3, 7);   // last load of the item issued
#ifdef MFA_TRACE
            ++tr_item;
#endif
        }
    } else if (warp == kWarpMma) {
        // ======================================================================= MMA issuer
        // The whole warp runs the (warp-uniform) control flow and the barrier waits; one elected lane issues the
        // tcgen05 instructions.  Written this way -- not `if (lane == 0)` -- so that ptxas keeps descriptors and
        // addresses in uniform registers instead of wrapping every UTCHMMA in an ELECT/R2UR waterfall loop
        // (measured: ~110 cycles per issued MMA with the loop, which made the issue thread the bottleneck).
        {
            constexpr uint32_t idescS = umma_idesc_f16(Elem<T>::kFmt, kBM, kBN, 0, 0);
            constexpr uint32_t idescO = umma_idesc_f16(Elem<T>::kFmt, kBM, DPAD, 0, 1);
            const uint32_t q_addr = smem_u32(q_smem);
            const uint32_t ring_addr = smem_u32(ring);

            auto issue_S = [&](int i, int stage) {
                if (elect_one()) {
                    const uint32_t qa = opaque(q_addr) + i * Cfg::kQTileBytes;
                    const uint32_t ka = opaque(ring_addr) + stage * Cfg::kKVTileBytes;
                    constexpr uint32_t tb = tmem_base;
#pragma unroll
                    for (int ks = 0; ks < DPAD / 16; ++ks)
                        umma_ss(tb + Cfg::kTmemS + i * 128, desc_kmajor(qa, ks), desc_kmajor(ka, ks), idescS,
                                ks > 0 ? 1u : 0u);
                    umma_commit(&s_full[i]);
                }
                __syncwarp();
            };
            auto issue_PV = [&](int i, int stage, bool accumulate, int ks0, int ks1, bool commit_o) {
                if (elect_one()) {
                    const uint32_t va = opaque(ring_addr) + stage * Cfg::kKVTileBytes;
                    constexpr uint32_t tb = tmem_base;
#pragma unroll
                    for (int ks = ks0; ks < ks1; ++ks)
                        umma_ts(tb + Cfg::kTmemO + i * DPAD, tb + Cfg::kTmemS + i * 128 + ks * 8,
                                desc_mnmajor(va, ks), idescO, (accumulate || ks > 0) ? 1u : 0u);
                    if (commit_o) umma_commit(&o_full[i]);
                }
                __syncwarp();
            };
            auto commit = [&](uint64_t* bar) {
                if (elect_one()) umma_commit(bar);
                __syncwarp();
            };
            uint32_t q_ld[2] = {0, 0};    // loads of Q tile i so far (the producer loads a tile only when it has KV work)
            // The S MMA of a Q tile's LAST KV tile is the last reader of the tile's shared-memory buffer: commit q_empty
            // right behind it -- one KV step before the item ends -- so that the producer fetches the next item's Q under
            // this item's last P.V instead of after it (the Q reload used to be a ~2000-cycle bubble at every item
            // boundary: 2 % of an item at S = 4096, 20 % at S = 512).
            auto release_q = [&](const PrefillItem& wx, int i, int jn) {
                if (jn == wx.n_i[i] - 1) {
                    commit(&q_empty[i]);
                    ++q_ld[i];
                }
            };

            // MEASURED LIMITS OF THIS WARP (scripts/umma_rate.py and register-held clock() stamps):
            //  * tcgen05.mma issue BLOCKS once ~4 SS / ~2 TS instructions are queued, so this thread is never more
            //    than ~270 / ~150 cycles ahead of the tensor pipe, and whatever it executes between two batches of
            //    MMAs beyond that slack drains the pipe;
            //  * an mbarrier try_wait costs ~90 cycles even when the phase completed long ago; with the fence, the
            //    elect and the descriptor moves each of the 7 waits of a KV step shows up as a 150-250 cycle gap
            //    (register stamps: MMA issue 2470 cycles + gaps ~1000 = the 3470-cycle KV step).
            // Tried and measured NOT to help (each within +-5 % of this form): one issuing warp per Q tile with and
            // without an issue token, P.V + S as one 16-MMA batch (one wait per tile), early non-blocking probes
            // (test_wait: ptxas consumes the predicate at once, the latency stays exposed), splitting the softmax of
            // a tile by key columns over both warpgroups.
            int it = 0;                   // K/V ring position, runs across items
            uint32_t t_cnt[2] = {0, 0};   // KV tiles done so far per Q tile (parity base of s_full / p_* / o_full)
            // The descriptor of the NEXT item (a handful of integer divisions, ~1000 cycles of dependent latency for a lone
            // warp) is computed while this warp waits for the last P of the current one, not between the items.
            PrefillItem w_next = prefill_item<kNQ, kCompact>(p, blockIdx.x, sched);
            for (int k = 0, item = blockIdx.x; item < total_items; ++k, item = next_item(k, item)) {
                const PrefillItem w = w_next;
                const int nxt = next_item(k + 1, item);
                if (!w.valid) {
                    if (nxt < total_items) w_next = prefill_item<kNQ, kCompact>(p, nxt, sched);
                    continue;
                }
#ifdef MFA_TRACE
                tr_on = tr_cta;
#endif
                const int n_max = w.n_max;
#pragma unroll
                for (int i = 0; i < kNQ; ++i)
                    if (w.n_i[i] > 0) mbar_wait(&q_full[i], q_ld[i] & 1);
                if (n_max > 0) {
                    const int sK = it % kStages;
                    mbar_wait(&kv_full[sK], (it / kStages) & 1);
                    ++it;
                    tc_fence_after();
#pragma unroll
                    for (int i = 0; i < kNQ; ++i)
                        if (w.active(i, w.j_lo)) {
                            if (i == 0) MFA_TR(2, 0, 7);   // Q tiles and K(first) of the item landed
                            issue_S(i, sK);
                            release_q(w, i, w.j_lo);
                        }
                    commit(&kv_empty[sK]);
                }
                for (int j = w.j_lo; j < n_max; ++j) {
                    const int sV = it % kStages;
                    mbar_wait(&kv_full[sV], (it / kStages) & 1);
                    ++it;
                    const bool has_next = (j + 1 < n_max);
                    int sK = 0;
                    if (has_next) {
                        sK = it % kStages;
                        mbar_wait(&kv_full[sK], (it / kStages) & 1);
                        ++it;
                    }
                    tc_fence_after();
                    MFA_TR(2, j - w.j_lo, 0);
#ifndef MFA_NO_WNEXT
                    if (!has_next && nxt < total_items) w_next = prefill_item<kNQ, kCompact>(p, nxt, sched);
#endif
#pragma unroll
                    for (int i = 0; i < kNQ; ++i) {
                        if (w.active(i, j)) {
                            const uint32_t par = (t_cnt[i] + (uint32_t)(j - w.lo_i[i])) & 1u;
                            // (the first P.V of an item overwrites O_i: safe, because every softmax thread finished the
                            // previous item's epilogue -- its TMEM reads of O_i included -- before it released this P)
                            mbar_wait(&p_lo[i], par);
                            tc_fence_after();
                            MFA_TR(2, j - w.j_lo, 1 + 3 * i);
                            issue_PV(i, sV, j > w.lo_i[i], 0, kBN / 32, false);
                            mbar_wait(&p_hi[i], par);
                            tc_fence_after();
                            MFA_TR(2, j - w.j_lo, 2 + 3 * i);
                            issue_PV(i, sV, j > w.lo_i[i], kBN / 32, kBN / 16, true);
                        }
                        if (w.active(i, j + 1)) {
                            issue_S(i, sK);
                            release_q(w, i, j + 1);
                        }
                        MFA_TR(2, j - w.j_lo, 3 + 3 * i);
                    }
                    commit(&kv_empty[sV]);
                    if (has_next) commit(&kv_empty[sK]);
                }
#ifndef MFA_NO_WNEXT
                if (n_max <= w.j_lo && nxt < total_items) w_next = prefill_item<kNQ, kCompact>(p, nxt, sched);   // (item without KV work)
#else
                if (nxt < total_items) w_next = prefill_item<kNQ, kCompact>(p, nxt, sched);
#endif
#pragma unroll
                for (int i = 0; i < kNQ; ++i) t_cnt[i] += (uint32_t)w.count(i);
#ifdef MFA_TRACE
                tr_base += n_max - w.j_lo;
#endif
            }
        }
    }
    } else {
      asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MFA_REG_SOFTMAX));
      if (warp < Cfg::kSoftmaxWarps) {
        // ======================================================================= softmax warpgroups
        // warp -> (Q tile i, TMEM lane quarter): a warp may only touch TMEM lanes 32 * (warp % 4) ... + 31
        const int i = warp / 4;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
        const uint32_t tS = tmem_base + lane_sel + Cfg::kTmemS + i * 128;
        const uint32_t tO = tmem_base + lane_sel + Cfg::kTmemO + i * DPAD;
        const float c2 = p.scale_log2;
        uint32_t t_cnt = 0;   // KV tiles this warpgroup has processed so far (parity base)
        PrefillItem w_next = prefill_item<kNQ, kCompact>(p, blockIdx.x, sched);   // (computed under the epilogue's wait for the last P.V, see the MMA warp)
        for (int k = 0, item = blockIdx.x; item < total_items; ++k, item = next_item(k, item)) {
            const PrefillItem w = w_next;
            const int nxt = next_item(k + 1, item);
            if (!w.valid) {
                if (nxt < total_items) w_next = prefill_item<kNQ, kCompact>(p, nxt, sched);
                continue;
            }
#ifdef MFA_TRACE
            tr_on = tr_cta;
#endif
            const int m0 = w.m0, len_k = w.len_k, len_q = w.len_q;
            const int qrow = m0 + i * kBM + row_in_tile;            // row inside the sequence
            const int j_first = w.lo_i[i];
            const int n_mine = w.count(i);
            float m_run = -INFINITY, l_run = 0.f;

            for (int j = j_first; j < w.n_i[i]; ++j) {
                const uint32_t par = (t_cnt + (uint32_t)(j - j_first)) & 1u;
                MFA_TR(i, j - j_first, 0);
                mbar_wait(&s_full[i], par);
                tc_fence_after();
                MFA_TR(i, j - j_first, 1);
                const int col0 = j * kBN;
                const bool need_mask = (col0 + kBN > len_k) || (p.has_hi && (col0 + kBN - 1 > m0 + i * kBM + w.shift_hi));
                const bool need_mask_lo = p.has_lo && col0 < m0 + i * kBM + kBM - 1 + w.shift_lo;   // the tile's last row has the largest bound
                auto mask64 = [&](uint32_t (&a)[64], int cbase) {   // a = columns col0 + cbase .. + 63 of the row
                    if (need_mask) {
                        // 32-column chunks that every row of the warp sees whole skip the per-element compare + select (the
                        // masked tile's max phase was 1550 cycles against 400 for an unmasked one: 256 such instructions).
                        // A warp holds 32 consecutive rows and the bound grows with the row, so below the diagonal -- and in
                        // front of the end of K in a ragged last tile -- most chunks qualify.  (Filling whole-warp-masked
                        // chunks with -inf by a third branch was tried: ptxas then spills the S row, 1.5 KB.)
                        const int vis = (p.has_hi ? min(len_k, qrow + 1 + w.shift_hi) : len_k) - col0 - cbase;   // visible columns of a[]
                        auto chunk = [&](auto kc) {
                            constexpr int k = decltype(kc)::value;
                            if (__all_sync(0xffffffffu, vis >= 32 * k + 32)) return;   // visible to every row of the warp
#pragma unroll
                            for (int c = 0; c < 32; ++c)
                                if (32 * k + c >= vis) a[32 * k + c] = 0xff800000u;   // -inf
                        };
                        chunk(std::integral_constant<int, 0>{});
                        chunk(std::integral_constant<int, 1>{});
                    }
                    if (need_mask_lo) {
                        const int lim_lo = qrow + w.shift_lo;   // cols < lim_lo are masked
#pragma unroll
                        for (int c = 0; c < 64; ++c)
                            if (col0 + cbase + c < lim_lo) a[c] = 0xff800000u;
                    }
                };
                // P = exp2(S*c2 - ms), packed FFMA2 for the scale, MUFU.EX2, packed add for the row sum, pack to 16 bit.
                // (Every exponential on the MUFU: moving a share of them to the FMA pipe as a degree-3 polynomial, the
                // FlashAttention-4 trick, was measured twice -- round 1 and again in round 2 with 1/8 ... 1/2 of the pairs --
                // and never won: this warp is alone on its sub-partition for the whole phase and issues ~1 instruction per
                // 2.3 cycles, so the ~10 extra FMA-pipe instructions per pair cost what the saved MUFU slots return.)
                const uint64_t c2_2 = f2_pack(c2, c2);
                auto exp64 = [&](const uint32_t (&a)[64], float ms_, uint32_t (&pk)[32], uint64_t (&acc)[2]) {
                    const uint64_t nms_2 = f2_pack(-ms_, -ms_);
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const uint64_t x2 = f2_fma(f2_pack(__uint_as_float(a[2 * c]), __uint_as_float(a[2 * c + 1])), c2_2, nms_2);
                        float x0, x1;
                        f2_unpack(x2, x0, x1);
                        const float p0 = ex2_fast(x0), p1 = ex2_fast(x1);
                        acc[c & 1] = f2_add(acc[c & 1], f2_pack(p0, p1));
                        pk[c] = Elem<T>::pack(p0, p1);
                    }
                };
                // max of 64 columns: 4 independent 3-input max chains (the compiler fuses 2-input chains into FMNMX3 anyway;
                // most of this phase is the TMEM read of the row block, 64 B/clk, not the ALU work)
                auto max64 = [&](const uint32_t (&a)[64]) -> float {
                    float mx[4];
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) mx[q4] = fmaxf(__uint_as_float(a[q4 * 16]), __uint_as_float(a[q4 * 16 + 1]));
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
#pragma unroll
                        for (int c = 2; c < 16; c += 2)
                            mx[q4] = fmax3(mx[q4], __uint_as_float(a[q4 * 16 + c]), __uint_as_float(a[q4 * 16 + c + 1]));
                    }
                    return fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
                };
                // lazy rescale of this thread's O row by `factor`: O_i holds sum_{t<j} P_t V_t; wait for PV(j-1) first.
                // Must happen before any part of P(j) is released.
                auto rescale_o = [&](float factor) {
                    mbar_wait(&o_full[i], par ^ 1u);
                    tc_fence_after();
                    // (16 columns at a time: this rare path runs with the whole S row live in registers)
#pragma unroll 1
                    for (int c = 0; c < DPAD; c += 16) {
                        uint32_t o[16];
                        tmem_ld16(tO + c, o);
                        tmem_wait_ld();
#pragma unroll
                        for (int e = 0; e < 16; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * factor);
                        tmem_st16(tO + c, o);
                    }
                };
                uint64_t sum2[2] = {0ull, 0ull};
                uint32_t su_lo[64], su_hi[64];
                tmem_ld32(tS, *reinterpret_cast<uint32_t(*)[32]>(&su_lo[0]));
                tmem_ld32(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&su_lo[32]));
                const bool unmasked = !(need_mask || need_mask_lo);   // warp-uniform; all but the diagonal / last tiles
                float tmax;
                if (unmasked) {
                    // the second half of the row streams out of TMEM (64 B/clk per sub-partition) under the max of the first:
                    // one basic block, no mask code in between (measured: the max phase 472 -> 295 cycles, +2.4 %)
                    tmem_wait_ld();
                    tmem_ld32(tS + 64, *reinterpret_cast<uint32_t(*)[32]>(&su_hi[0]));
                    tmem_ld32(tS + 96, *reinterpret_cast<uint32_t(*)[32]>(&su_hi[32]));
                    MFA_TR(i, j - j_first, 2);
                    const float tmax_lo = max64(su_lo);
                    tmem_wait_ld();
                    tmax = fmaxf(tmax_lo, max64(su_hi));
                } else {
                    tmem_ld32(tS + 64, *reinterpret_cast<uint32_t(*)[32]>(&su_hi[0]));
                    tmem_ld32(tS + 96, *reinterpret_cast<uint32_t(*)[32]>(&su_hi[32]));
                    tmem_wait_ld();
                    MFA_TR(i, j - j_first, 2);
                    mask64(su_lo, 0);
                    mask64(su_hi, 64);
                    tmax = fmaxf(max64(su_lo), max64(su_hi));
                }
                MFA_TR(i, j - j_first, 6);
                const float m_new = fmaxf(m_run, tmax);
                // lazy rescale: keep the stale max unless it grew by more than 2^8
                const bool grow = (m_new - m_run) * c2 > Cfg::kRescaleThreshold;   // also true when m_run == -inf
                const float m_use = grow ? m_new : m_run;
                const float factor = grow ? ((m_run == -INFINITY) ? 0.f : exp2f((m_run - m_new) * c2)) : 1.f;
                const float ms = (m_use == -INFINITY) ? 0.f : m_use * c2;
                m_run = m_use;
                const bool any_grow = __any_sync(0xffffffffu, grow);
                MFA_TR(i, j - j_first, 7);
                if (j > j_first && any_grow) rescale_o(factor);
                MFA_TR(i, j - j_first, 3);
                // P is released to the MMA warp in two halves so that P.V k-steps 0-3 run while the second half is still
                // being exponentiated
                {
                    uint32_t pk0[32];
                    exp64(su_lo, ms, pk0, sum2);
                    tmem_st32(tS, pk0);
                    tmem_wait_st();
                    tc_fence_before();
                    mbar_arrive(&p_lo[i]);
                    MFA_TR(i, j - j_first, 4);
                }
                {
                    uint32_t pk1[32];
                    exp64(su_hi, ms, pk1, sum2);
                    tmem_st32(tS + 32, pk1);
                    tmem_wait_st();
                    tc_fence_before();
                    mbar_arrive(&p_hi[i]);
                    MFA_TR(i, j - j_first, 5);
                }
                float sa, sb, sc, sd;
                f2_unpack(sum2[0], sa, sb);
                f2_unpack(sum2[1], sc, sd);
                l_run = l_run * factor + ((sa + sb) + (sc + sd));
            }

            // ---- epilogue: O / l -> out dtype -> global.  tcgen05.ld is warp-collective (.sync.aligned), so every
            // lane takes part in the TMEM loads; only the global stores are guarded by row < len_q.
            {
                const bool row_ok = qrow < len_q;
                T* orow = static_cast<T*>(p.o) + (int64_t)w.h * p.o_head_stride +
                          (p.cu_seqlens_q ? (int64_t)(w.q_off + qrow) * p.o_row_stride
                                          : (int64_t)w.b * p.o_batch_stride + (int64_t)qrow * p.o_row_stride);
                MFA_TRI(3 + i, 0);
#ifndef MFA_NO_WNEXT
                if (nxt < total_items) w_next = prefill_item<kNQ, kCompact>(p, nxt, sched);
#endif
                if (n_mine > 0) {
                    mbar_wait(&o_full[i], (t_cnt + (uint32_t)n_mine - 1u) & 1u);
                    tc_fence_after();
                }
                MFA_TRI(3 + i, 1);
                const float l_row = l_run;
                const float inv = (l_row == 0.f || l_row != l_row) ? 1.f : 1.f / l_row;   // prefill.cuh:604-605
                if (p.lse != nullptr && row_ok)   // natural-log LSE of the scaled scores (the reference drops it, api.cpp:436)
                    p.lse[(int64_t)w.b * p.lse_batch_stride + (int64_t)w.h * p.lse_head_stride + w.q_off + qrow] =
                        (l_row == 0.f || l_row != l_row) ? -INFINITY : (m_run * c2 + log2f(l_row)) * 0.6931471805599453f;
#ifdef MFA_EPI_DIRECT
                constexpr bool kEpiDirect = true;
#else
                constexpr bool kEpiDirect = false;
#endif
                // (Tried and not kept: direct stores for a Q tile that ends before its item does -- causal: tile 0 has one KV tile
                // less than tile 1 --, and the direct form chosen per launch for short causal problems: the compile-time
                // direct build, -DMFA_EPI_DIRECT, is 1.5 % (S=1024) to 5 % (S=512) faster on those and 10-25 % slower on the
                // short non-causal ones; with both forms in one kernel the gain shrinks to 1 %.)
                if (kEpiDirect) {
#pragma unroll
                    for (int c = 0; c < DPAD / 32; ++c) {
                        uint32_t o[32];
                        if (n_mine > 0) {   // warp-uniform
                            tmem_ld32(tO + c * 32, o);
                            tmem_wait_ld();
                        } else {
#pragma unroll
                            for (int e = 0; e < 32; ++e) o[e] = 0u;
                        }
                        if (row_ok) {
#pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                if (c * 32 + v * 8 < p.head_dim) {
                                    uint4 x;
                                    x.x = Elem<T>::pack(__uint_as_float(o[v * 8 + 0]) * inv, __uint_as_float(o[v * 8 + 1]) * inv);
                                    x.y = Elem<T>::pack(__uint_as_float(o[v * 8 + 2]) * inv, __uint_as_float(o[v * 8 + 3]) * inv);
                                    x.z = Elem<T>::pack(__uint_as_float(o[v * 8 + 4]) * inv, __uint_as_float(o[v * 8 + 5]) * inv);
                                    x.w = Elem<T>::pack(__uint_as_float(o[v * 8 + 6]) * inv, __uint_as_float(o[v * 8 + 7]) * inv);
                                    *reinterpret_cast<uint4*>(orow + c * 32 + v * 8) = x;
                                }
                            }
                        }
                    }
                } else {
                    // A thread holds a whole output row, but 32 lanes storing 16 bytes each to 32 different rows is the worst
                    // pattern for the memory pipe.  Each warp transposes its 32 rows through a private 4 KB staging area, 64
                    // columns (128 bytes per row) at a time, and stores whole 128-byte lines: 8 lanes per row, 4 rows per
                    // instruction.  16-byte pieces are XOR-swizzled by the row so that both phases are bank-conflict free; the
                    // area is private to the warp, so __syncwarp is all the synchronisation there is.
                    const uint32_t stage = smem_u32(smem + Cfg::kBarOffset + 512 + warp * Cfg::kOStageWarpBytes);
                    const int row0_w = m0 + i * kBM + (warp & 3) * 32;   // sequence row of lane 0
                    T* obase = static_cast<T*>(p.o) + (int64_t)w.h * p.o_head_stride +
                               (p.cu_seqlens_q ? (int64_t)(w.q_off + row0_w) * p.o_row_stride
                                               : (int64_t)w.b * p.o_batch_stride + (int64_t)row0_w * p.o_row_stride);
                    // The O row is read from TMEM in groups of up to 128 columns with ONE wait per group (the S registers are
                    // free here): four dependent load -> wait round trips per row were ~600 cycles of every item boundary.
                    constexpr int kGroup = (DPAD % 128 == 0) ? 128 : 64;
#pragma unroll
                    for (int g0 = 0; g0 < DPAD; g0 += kGroup) {
                        uint32_t o[kGroup];
                        if (n_mine > 0) {   // warp-uniform
#pragma unroll
                            for (int c = 0; c < kGroup / 32; ++c)
                                tmem_ld32(tO + g0 + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&o[c * 32]));
                            tmem_wait_ld();
                        } else {
#pragma unroll
                            for (int e = 0; e < kGroup; ++e) o[e] = 0u;
                        }
#pragma unroll
                        for (int r = 0; r < kGroup / 64; ++r) {
#pragma unroll
                            for (int v = 0; v < 8; ++v) {
                                const uint32_t* ov = &o[r * 64 + v * 8];
                                uint4 x;
                                x.x = Elem<T>::pack(__uint_as_float(ov[0]) * inv, __uint_as_float(ov[1]) * inv);
                                x.y = Elem<T>::pack(__uint_as_float(ov[2]) * inv, __uint_as_float(ov[3]) * inv);
                                x.z = Elem<T>::pack(__uint_as_float(ov[4]) * inv, __uint_as_float(ov[5]) * inv);
                                x.w = Elem<T>::pack(__uint_as_float(ov[6]) * inv, __uint_as_float(ov[7]) * inv);
                                sts128(stage + lane * 128 + ((v ^ (lane & 7)) * 16), x);
                            }
                            __syncwarp();
#pragma unroll
                            for (int pass = 0; pass < 8; ++pass) {
                                const int row = pass * 4 + (lane >> 3);
                                const int piece = lane & 7;
                                const uint4 x = lds128(stage + row * 128 + ((piece ^ (row & 7)) * 16));
                                const int col = g0 + r * 64 + piece * 8;
                                if (row0_w + row < len_q && col < p.head_dim)
                                    *reinterpret_cast<uint4*>(obase + (int64_t)row * p.o_row_stride + col) = x;
                            }
                            __syncwarp();
                        }
                    }
                }
            }
            MFA_TRI(3 + i, 2);
#ifdef MFA_NO_WNEXT
            if (nxt < total_items) w_next = prefill_item<kNQ, kCompact>(p, nxt, sched);
#endif
            t_cnt += (uint32_t)n_mine;
#ifdef MFA_TRACE
            tr_base += n_mine;
            ++tr_item;
#endif
        }
      }
    }

    // ---- teardown
    tc_fence_before();
    __syncthreads();
    if (warp == kWarpTmem) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::kTmemCols);
    }
}

}  // namespace

template <typename T, int DPAD, int kMode>
static cudaError_t launch_prefill_t(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                                    const PrefillKParams& p, cudaStream_t stream) {
    using Cfg = PrefillCfg<DPAD>;
    // exp2 placement: every exponential on the MUFU.  (Round 1 carried a variant that put a quarter of them on the FMA
    // pipe as a degree-3 polynomial; it lost 2-4 % in every back-to-back measurement -- the board sits at its power cap
    // and the extra FMA work costs more clock than the relieved MUFU returns -- and was removed.  The packed
    // ex2.approx.f16x2 / bf16x2 forms are no way out either: ptxas splits them into two MUFU.EX2 per pair.)
    auto kern = prefill_fwd_kernel<T, DPAD, kMode>;
    static bool attr_done[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    // Fixed-length batches run PERSISTENT: one CTA per SM walks the work items (m-block fastest, so that the CTAs
    // running at the same time share K/V in L2).  With the causal mask an item's cost grows with its m-block, and a
    // CTA strides through the items by gridDim.x: the grid size is made coprime to the number of m-blocks so that every
    // CTA cycles through all of them (148 CTAs over 16 m-blocks would see only 4 each: measured 14 % slower).
    // Packed (varlen) batches have ragged items; they launch one CTA per item and leave the balancing to the hardware
    // CTA scheduler.  MFA_PREFILL_NONPERSISTENT=1 forces that form everywhere (tuning aid).
    static int sm_count[64] = {};
    if (dev < 64 && sm_count[dev] == 0) cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    static const bool non_persistent = [] { const char* e = getenv("MFA_PREFILL_NONPERSISTENT"); return e && e[0] == '1'; }();
    const long long items = (long long)p.num_m_blocks * p.heads * p.batch;
    if (items <= 0 || items > 0x7fffffffLL) return cudaErrorInvalidValue;
    const int sms = (dev < 64 && sm_count[dev] > 0) ? sm_count[dev] : 148;
    long long ctas = items;
    if (!non_persistent && (p.cu_seqlens_q == nullptr || p.snake) && items > sms) {
        ctas = sms;
        if (p.has_hi) {
            auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
            while (ctas > 1 && gcd(ctas, p.num_m_blocks) != 1) --ctas;
        }
    }
    dim3 grid((unsigned)ctas, 1, 1);
    kern<<<grid, Cfg::kThreads, Cfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
    return cudaGetLastError();
}

int prefill_rows_per_cta(int dpad) { return (dpad <= 128 ? 2 : 1) * kPrefillBlockM; }

cudaError_t launch_prefill(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                           const PrefillKParams& p, bool is_bf16, int dpad, cudaStream_t stream) {
#define MFA_PREFILL_CASE(DP)                                                                     \
    case DP:                                                                                     \
        if (p.compact)                                                                           \
            return is_bf16 ? launch_prefill_t<__nv_bfloat16, DP, 2>(tmQ, tmK, tmV, p, stream)        \
                           : launch_prefill_t<__half, DP, 2>(tmQ, tmK, tmV, p, stream);              \
        if (p.snake)                                                                             \
            return is_bf16 ? launch_prefill_t<__nv_bfloat16, DP, 1>(tmQ, tmK, tmV, p, stream)        \
                           : launch_prefill_t<__half, DP, 1>(tmQ, tmK, tmV, p, stream);              \
        return is_bf16 ? launch_prefill_t<__nv_bfloat16, DP, 0>(tmQ, tmK, tmV, p, stream)        \
                       : launch_prefill_t<__half, DP, 0>(tmQ, tmK, tmV, p, stream);
    switch (dpad) {
        MFA_PREFILL_CASE(64)
        MFA_PREFILL_CASE(128)
        MFA_PREFILL_CASE(192)   // head_dim 136 ... 192: one Q tile per CTA like 256, but 3 instead of 4 sub-tiles of every operand
        MFA_PREFILL_CASE(256)
        default:
            return cudaErrorInvalidValue;
    }
#undef MFA_PREFILL_CASE
}

}  // namespace mfa
