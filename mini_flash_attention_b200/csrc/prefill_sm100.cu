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
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kparams.h"
#include "ptx.cuh"

// Register split of the 64 K registers of an SM over the CTA's 384 threads (setmaxnreg): 256 softmax threads and
// 128 data-movement / MMA / allocator threads; 256 * SOFTMAX + 128 * HELPER <= 65536.
#ifndef MFA_REG_SOFTMAX
#define MFA_REG_SOFTMAX 216
#endif
#ifndef MFA_REG_HELPER
#define MFA_REG_HELPER 72
#endif
// P hand-off from the softmax threads to the MMA warp.  MFA_PSPLIT 1: in two halves (p_lo after KV rows 0-63, p_hi after
// 64-127) so that the first P.V k-steps run while the second half is exponentiated; 0: one hand-off per tile (one wait
// and one tcgen05.wait::st / fence / arrive per thread less).  MFA_PARRIVE_WARP 1: one arrival per WARP (lane 0 after
// __syncwarp; barrier count 4) instead of one per thread (count 128).
#ifndef MFA_PSPLIT
#define MFA_PSPLIT 1
#endif
// MFA_EARLY_S 1 (flat stream): a Q tile that finishes an item one KV step before the other (the causal mask: tile 0 has
// one KV tile less than tile 1) gets the NEXT item's first S in that step already, so its softmax warpgroup works
// through the other tile's lone last step instead of idling.
#ifndef MFA_EARLY_S
#define MFA_EARLY_S 1
#endif
#ifndef MFA_PARRIVE_WARP
#define MFA_PARRIVE_WARP 0
#endif
#ifndef MFA_POLY_MASK
#define MFA_POLY_MASK 0x8u   // pair 3 of every 4 -> a quarter of the exponentials off the MUFU (measured best)
#endif

#ifdef MFA_TRACE
#define MFA_TR(role, it, ev)                                                                              \
    do {                                                                                                  \
        if (tr_on && (it) < 64) p.trace[((role) * 64 + (it)) * 8 + (ev)] = clock64();                      \
    } while (0)
#else
#define MFA_TR(role, it, ev) do { } while (0)
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
    static constexpr int kNQ = (DPAD <= 128) ? 2 : 1;            // Q tiles per CTA (TMEM budget)
    static constexpr int kSub = DPAD / 64;                       // 64-column (128-byte) sub-tiles per row
    static constexpr int kSubBytes = 128 * 128;                  // 128 rows x 128 B
    static constexpr int kQTileBytes = kSub * kSubBytes;
    static constexpr int kKVTileBytes = kSub * kSubBytes;
#ifndef MFA_STAGES128
#define MFA_STAGES128 4
#endif
    static constexpr int kStages = (DPAD == 64) ? 8 : (DPAD == 128 ? MFA_STAGES128 : 2);
    // Flat K/V stream (kStages >= 4): the ring is used as kStages/2 DOUBLE SLOTS [V_j | K_{j+1}] with one full and one
    // empty barrier each -- the two tiles one KV step of the MMA warp needs arrive under one barrier, and the stream
    // runs across work items ([V_last(item a) | K_first(item a+1)]).  head_dim 256 has room for 2 slots only and keeps
    // the per-tile ring.
#ifdef MFA_NO_FLAT
    static constexpr bool kFlat = false;   // tuning aid: the per-tile ring of round 1 at every head_dim
#else
    static constexpr bool kFlat = kStages >= 4;
#endif
    static constexpr int kDS = kStages / 2;
    static constexpr int kQBytes = kNQ * kQTileBytes;
    static constexpr int kRingBytes = kStages * kKVTileBytes;
    static constexpr int kBarOffset = kQBytes + kRingBytes;
    static constexpr int kSmemBytes = kBarOffset + 512 + 1024;   // barriers + alignment slack
    static constexpr uint32_t kTmemCols = 512;
    static constexpr uint32_t kTmemS = 0;                        // S_i at kTmemS + i*128
    static constexpr uint32_t kTmemO = 256;                      // O_i at kTmemO + i*DPAD
    static constexpr float kRescaleThreshold = 8.0f;             // log2 domain
    static constexpr uint32_t kPolyMask = MFA_POLY_MASK;         // column pairs (c & 3) whose exp2 runs on the FMA pipe
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

// 2^x for two packed floats, x <= ~8, on the FMA pipe: clamp, round-to-nearest split x = n + f (|f| <= 0.5) with
// the 1.5*2^23 magic constant, 2^f by a degree-3 polynomial with p(0) == 1 exactly (max rel. error 1.0e-4, below
// the half-ulp of the fp16/bf16 P it is rounded to), exponent added with one integer shift-add.
// x <= -127 gives exactly +0 (masked columns hold -inf).
__device__ __forceinline__ uint64_t exp2_poly2(uint64_t x2) {
    float x0, x1;
    f2_unpack(x2, x0, x1);
    x0 = fmaxf(x0, -127.f);
    x1 = fmaxf(x1, -127.f);
    const uint64_t xc = f2_pack(x0, x1);
    const uint64_t magic = f2_pack(12582912.f, 12582912.f);
    const uint64_t nmagic = f2_pack(-12582912.f, -12582912.f);
    const uint64_t t = f2_add(xc, magic);
    const uint64_t n = f2_add(t, nmagic);
    const uint64_t f = f2_fma(n, f2_pack(-1.f, -1.f), xc);
    uint64_t q = f2_fma(f, f2_pack(0.05500871315598488f, 0.05500871315598488f), f2_pack(0.24221068620681763f, 0.24221068620681763f));
    q = f2_fma(q, f, f2_pack(0.6932829022407532f, 0.6932829022407532f));
    q = f2_fma(q, f, f2_pack(1.f, 1.f));
    float t0, t1, q0, q1;
    f2_unpack(t, t0, t1);
    f2_unpack(q, q0, q1);
    const float r0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
    const float r1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
    return f2_pack(r0, r1);
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

template <int kNQ>
__device__ __forceinline__ PrefillItem prefill_item(const PrefillKParams& p, int item) {
    PrefillItem w;
    const int mraw = item % p.num_m_blocks;
    const int rest = item / p.num_m_blocks;
    const int mblk = (p.has_hi && !p.has_lo) ? (p.num_m_blocks - 1 - mraw) : mraw;   // causal: heaviest (largest m) blocks first
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
template <typename T, int DPAD, bool kPoly>
__global__ void __launch_bounds__(kPrefillThreads, 1)
    prefill_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ PrefillKParams p) {
    using Cfg = PrefillCfg<DPAD>;
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
    const int total_items = p.num_m_blocks * p.heads * p.batch;
#ifdef MFA_TRACE
    const bool tr_cta = p.trace && blockIdx.x == 0 && lane == 0 && (warp == 9 || (warp & 3) == 0);
    bool tr_on = false;
#endif

    // ---- one-time setup
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&q_full[i], 1);
            mbar_init(&s_full[i], 1);
            mbar_init(&p_lo[i], MFA_PARRIVE_WARP ? 4 : 128);
            mbar_init(&p_hi[i], MFA_PARRIVE_WARP ? 4 : 128);
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
    if (warp == 10) {
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
    constexpr uint32_t tmem_base = 0u;

    // Register re-balancing: the launch gives every thread 168 registers (65536 / 384 threads); the
    // data-movement warpgroup hands most of its share to the two softmax warpgroups (whole S row in registers).
    // (Each setmaxnreg sits at the top of its own role branch so that ptxas sees the budget per region.)
    if (warp >= 8) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(MFA_REG_HELPER));
    if (warp == 8) {
        // ======================================================================= TMA producer
        // Warp-uniform control flow with one ELECTED lane issuing (see the MMA warp): under a divergent
        // `if (lane == 0)` ptxas cannot keep the TMA operands in uniform registers.
        if (elect_one()) {
            prefetch_tensormap(&tmQ);
            prefetch_tensormap(&tmK);
            prefetch_tensormap(&tmV);
        }
        if constexpr (Cfg::kFlat) {
            uint32_t tail_phase = 0;
            uint32_t ds = 0;             // double-slot counter, runs across items
            bool open = false;           // double slot `ds` holds a V tile (half A) and waits for the K tile that closes it
            uint32_t q_ld[2] = {0, 0};   // loads of Q tile i so far (parity base of q_full / q_empty)
            for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
                const PrefillItem w = prefill_item<kNQ>(p, item);
                if (!w.valid || w.n_max == 0) continue;   // nothing to load: the softmax warps write the zero rows alone
                const int32_t* bt = p.block_table ? p.block_table + (int64_t)w.b * p.block_table_batch_stride : nullptr;
                const bool packed_k = (p.cu_seqlens_k != nullptr) && (bt == nullptr);
                // rows >= len_k of a V tile must be zero (P is 0 there, but 0 * NaN = NaN).  For fixed-length
                // batches TMA's out-of-bounds fill does it; for packed / paged K/V the producer warp zeroes them.
                const bool need_tail_fix = (bt != nullptr) || packed_k;
                // one K or V tile of KV tile j into `dst`; its bytes are counted on `bar` (with the phase's one arrival
                // when `arrive`, as a bare expect_tx otherwise)
                auto load_tile = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* bar, bool arrive, int j) {
                    const int row0 = j * kBN;
                    if (elect_one()) {
                        if (bt) {
                            const int rows_valid = min(w.len_k - row0, kBN);
                            const int nchunks = (rows_valid + p.kv_box_rows - 1) / p.kv_box_rows;
                            const int chunk_bytes = p.kv_box_rows * 128;
                            const uint32_t bytes = (uint32_t)(nchunks * chunk_bytes * Cfg::kSub);
                            if (arrive) mbar_arrive_expect_tx(bar, bytes); else mbar_expect_tx(bar, bytes);
                            for (int c = 0; c < nchunks; ++c) {
                                const int row = row0 + c * p.kv_box_rows;
                                const int pg = row / p.page;
                                const int in_page = row - pg * p.page;
                                const int blk = bt[pg];
#pragma unroll
                                for (int sb = 0; sb < Cfg::kSub; ++sb)
                                    tma_load_4d(dst + sb * Cfg::kSubBytes + c * chunk_bytes, tm, bar, sb * 64, in_page, w.kh, blk);
                            }
                        } else {
                            if (arrive) mbar_arrive_expect_tx(bar, Cfg::kKVTileBytes); else mbar_expect_tx(bar, Cfg::kKVTileBytes);
#pragma unroll
                            for (int sb = 0; sb < Cfg::kSub; ++sb)
                                tma_load_4d(dst + sb * Cfg::kSubBytes, tm, bar, sb * 64, (packed_k ? w.k_off : 0) + row0, w.kh,
                                            packed_k ? 0 : w.b);
                        }
                    }
                    __syncwarp();
                };
                const int q_row_coord = p.cu_seqlens_q ? (w.q_off + w.m0) : w.m0;
                const int q_b_coord = p.cu_seqlens_q ? 0 : w.b;
                for (int j = w.j_lo; j < w.n_max; ++j) {
                    // ---- K_j: half B of the open double slot (closing it), or of a fresh one at the very start
                    {
                        const uint32_t st = ds % Cfg::kDS;
                        if (!open) mbar_wait(&kv_empty[st], ((ds / Cfg::kDS) & 1) ^ 1);
                        load_tile(&tmK, ring + (2 * st + 1) * Cfg::kKVTileBytes, &kv_full[st], true, j);
                        ++ds;
                        open = false;
                    }
                    if (j == w.j_lo) {
                        // Q tiles of this item, AFTER its first K tile (that one closes the previous item's last double
                        // slot, which must not wait for the Q buffers).  Only tiles with KV work are loaded: a tile
                        // without work is never read by an MMA, and a bulk copy nobody awaits could land after the CTA
                        // exited or be re-armed before it landed.  q_full / q_empty phases count LOADS of tile i; the
                        // MMA warp commits q_empty[i] right after the tile's last S MMA, about one KV step before the
                        // item ends, so this wait is normally over when the producer gets here.
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
                    }
                    // ---- V_j: half A of the next double slot, left open
                    {
                        const uint32_t st = ds % Cfg::kDS;
                        mbar_wait(&kv_empty[st], ((ds / Cfg::kDS) & 1) ^ 1);
                        uint8_t* dst = ring + (2 * st) * Cfg::kKVTileBytes;
                        const int rows_valid = min(w.len_k - j * kBN, kBN);
                        const bool fix = need_tail_fix && rows_valid < kBN;
                        load_tile(&tmV, dst, fix ? tail_bar : &kv_full[st], fix, j);
                        if (fix) {
                            // wait for the bytes, zero rows [rows_valid, 128) of every sub-tile, publish to the MMA proxy
                            // (the arrival that closes this double slot -- the next K tile or the final arrive -- follows)
                            mbar_wait(tail_bar, tail_phase);
                            tail_phase ^= 1;
                            for (int r = rows_valid; r < kBN; ++r) {
#pragma unroll
                                for (int sb = 0; sb < Cfg::kSub; ++sb)
                                    *reinterpret_cast<uint32_t*>(dst + sb * Cfg::kSubBytes + r * 128 + lane * 4) = 0u;
                            }
                            fence_proxy_async_smem();
                            __syncwarp();
                        }
                        open = true;
                    }
                }
            }
            if (open) {   // the CTA's last V tile: nothing follows, close its double slot
                if (elect_one()) mbar_arrive(&kv_full[ds % Cfg::kDS]);
                __syncwarp();
            }
        } else {
            uint32_t tail_phase = 0;
            int it = 0;                  // K/V ring position, runs across items
            uint32_t q_ld[2] = {0, 0};   // loads of Q tile i so far (parity base of q_full / q_empty)
            for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
                const PrefillItem w = prefill_item<kNQ>(p, item);
                if (!w.valid) continue;
                const int q_row_coord = p.cu_seqlens_q ? (w.q_off + w.m0) : w.m0;
                const int q_b_coord = p.cu_seqlens_q ? 0 : w.b;
                // Q tiles of this item -- only the tiles that have KV work (n_i > 0).  A tile without work is never read by
                // an MMA, so loading it would leave a bulk copy in flight that nobody awaits (it could land after the CTA
                // exited, or be re-armed before it landed).  q_full / q_empty phases therefore count LOADS of tile i
                // (q_ld[i]), not items: every load is awaited by the MMA warp before it commits q_empty, and the producer
                // observes that commit before the next load, so neither barrier ever runs more than one phase ahead.
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
            }
        }
    } else if (warp == 9) {
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
            auto issue_PV_at = [&](int i, uint32_t va, bool accumulate, int ks0, int ks1, bool commit_o) {
                if (elect_one()) {
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
            if constexpr (Cfg::kFlat) {
                // FLAT STREAM.  The CTA's KV steps form one sequence across its work items.  Step (item, j) waits for
                // ONE double slot [V_j | K_next] -- K_next is K_{j+1} of the item, or the first K tile of the next work
                // item when j is the item's last -- issues P.V(j) and S(next) for both Q tiles and releases the slot with
                // ONE commit: 5 barrier waits per step instead of 7, and the first S MMAs of item n+1 are in the queue
                // before the last P.V of item n has finished (no drain at the item boundary; with the causal mask the
                // lone last step of Q tile 1 overlaps the next item's first step of Q tile 0).
                uint32_t ds = 0;              // double-slot counter, runs across items
                uint32_t q_ld[2] = {0, 0};    // loads of Q tile i so far (the producer loads a tile only when it has KV work)
                uint32_t t_cnt[2] = {0, 0};   // KV tiles done so far per Q tile (parity base of s_full / p_* / o_full)
                // S_i(jn) of item wx from the K tile at `ka`: the tile's first S waits for its Q load, its last S
                // releases the Q buffer (so the next item's Q is loaded about one KV step before it is needed)
                auto s_step = [&](const PrefillItem& wx, int i, int jn, uint32_t ka) {
                    if (wx.active(i, jn)) {
                        if (jn == wx.lo_i[i]) {
                            mbar_wait(&q_full[i], q_ld[i] & 1);
                            tc_fence_after();
                        }
                        if (elect_one()) {
                            const uint32_t qa = opaque(q_addr) + i * Cfg::kQTileBytes;
                            constexpr uint32_t tb = tmem_base;
#pragma unroll
                            for (int ks = 0; ks < DPAD / 16; ++ks)
                                umma_ss(tb + Cfg::kTmemS + i * 128, desc_kmajor(qa, ks), desc_kmajor(ka, ks), idescS,
                                        ks > 0 ? 1u : 0u);
                            umma_commit(&s_full[i]);
                            if (jn == wx.n_i[i] - 1) umma_commit(&q_empty[i]);
                        }
                        __syncwarp();
                        if (jn == wx.n_i[i] - 1) ++q_ld[i];
                    }
                };
                auto next_work = [&](int item, PrefillItem& wx) -> int {
                    for (; item < total_items; item += gridDim.x) {
                        wx = prefill_item<kNQ>(p, item);
                        if (wx.valid && wx.n_max > 0) return item;
                    }
                    return -1;
                };
                PrefillItem w;
                int item = next_work(blockIdx.x, w);
                if (item >= 0) {
                    // the CTA's first double slot holds only K_first
                    mbar_wait(&kv_full[0], 0);
                    tc_fence_after();
#pragma unroll
                    for (int i = 0; i < kNQ; ++i) s_step(w, i, w.j_lo, ring_addr + Cfg::kKVTileBytes);
                    commit(&kv_empty[0]);
                    ds = 1;
                    mbar_wait(&kv_full[1 % Cfg::kDS], (1 / Cfg::kDS) & 1);   // the double slot of the first KV step
                    tc_fence_after();
                }
                while (item >= 0) {
#ifdef MFA_TRACE
                    tr_on = tr_cta && item == (int)blockIdx.x;
#endif
                    PrefillItem wn;
                    int nitem = -1;
                    bool looked = false;           // wn / nitem are valid
                    bool early[2] = {false, false};   // the next item's first S of tile i has been issued already
                    for (int j = w.j_lo; j < w.n_max; ++j) {
                        const bool last = (j + 1 == w.n_max);
                        const bool penult = MFA_EARLY_S && (j + 2 == w.n_max);
                        if ((last || penult) && !looked) {
                            nitem = next_work(item + (int)gridDim.x, wn);
                            looked = true;
                        }
                        // (this step's double slot was waited for in the MIDDLE of the previous step, right after Q tile 0's
                        // batch of S MMAs, when ~4 SS instructions are still queued: at the loop boundary the tensor pipe
                        // has only ~270 cycles of queued work, and a barrier wait there -- ~90 cycles even when the phase
                        // completed long ago -- plus the commit and the first P wait left it idle for ~300 cycles per step)
                        const uint32_t st = ds % Cfg::kDS;
                        ++ds;
                        const bool more = !last || nitem >= 0;   // another KV step follows (this item's or the next item's)
                        MFA_TR(2, j, 0);
                        const uint32_t va = opaque(ring_addr) + (2 * st) * Cfg::kKVTileBytes;
                        const uint32_t ka = va + Cfg::kKVTileBytes;
#pragma unroll
                        for (int i = 0; i < kNQ; ++i) {
                            if (w.active(i, j)) {
                                const uint32_t par = (t_cnt[i] + (uint32_t)(j - w.lo_i[i])) & 1u;
                                // (the first P.V of an item overwrites O_i: safe, because every softmax thread finished
                                // the previous item's epilogue -- its TMEM reads of O_i included -- before it released this P)
#if MFA_PSPLIT
                                mbar_wait(&p_lo[i], par);
                                tc_fence_after();
                                MFA_TR(2, j, 1 + 3 * i);
                                issue_PV_at(i, va, j > w.lo_i[i], 0, kBN / 32, false);
                                mbar_wait(&p_hi[i], par);
                                tc_fence_after();
                                MFA_TR(2, j, 2 + 3 * i);
                                issue_PV_at(i, va, j > w.lo_i[i], kBN / 32, kBN / 16, true);
#else
                                mbar_wait(&p_hi[i], par);
                                tc_fence_after();
                                MFA_TR(2, j, 2 + 3 * i);
                                issue_PV_at(i, va, j > w.lo_i[i], 0, kBN / 16, true);
#endif
                            }
                            if (!last) s_step(w, i, j + 1, ka);
                            if (i == 0 && more) {
                                mbar_wait(&kv_full[ds % Cfg::kDS], (ds / Cfg::kDS) & 1);
                                tc_fence_after();
                            }
                            MFA_TR(2, j, 3 + 3 * i);
                        }
                        // the item's last step: both P.V first, then the next item's first S (its Q tiles may still be
                        // in flight; the wait must not hold back this item's P.V)
                        if (last && nitem >= 0) {
#pragma unroll
                            for (int i = 0; i < kNQ; ++i)
                                if (!early[i]) s_step(wn, i, wn.j_lo, ka);
                        }
                        // the step before: a tile that is already done with this item (no S(j+1), e.g. Q tile 0 under the
                        // causal mask) takes the next item's first S now.  Its K tile is in the NEXT step's double slot,
                        // which was waited for above; that slot's release (the next step's commit) covers this MMA too.
                        if (penult && nitem >= 0) {
                            const uint32_t ka_next = opaque(ring_addr) + (2 * (ds % Cfg::kDS) + 1) * Cfg::kKVTileBytes;
#pragma unroll
                            for (int i = 0; i < kNQ; ++i) {
                                if (j + 1 >= w.n_i[i] && wn.active(i, wn.j_lo)) {
                                    s_step(wn, i, wn.j_lo, ka_next);
                                    early[i] = true;
                                }
                            }
                        }
                        commit(&kv_empty[st]);
                    }
#pragma unroll
                    for (int i = 0; i < kNQ; ++i) t_cnt[i] += (uint32_t)w.count(i);
                    w = wn;
                    item = nitem;
                }
            } else {
                int it = 0;                   // K/V ring position, runs across items
                uint32_t q_ld[2] = {0, 0};    // loads of Q tile i so far (the producer loads a tile only when it has KV work)
                uint32_t t_cnt[2] = {0, 0};   // KV tiles done so far per Q tile (parity base of s_full / p_* / o_full)
                for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
                    const PrefillItem w = prefill_item<kNQ>(p, item);
                    if (!w.valid) continue;
    #ifdef MFA_TRACE
                    tr_on = tr_cta && item == (int)blockIdx.x;
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
                            if (w.active(i, w.j_lo)) issue_S(i, sK);
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
                        MFA_TR(2, j, 0);
    #pragma unroll
                        for (int i = 0; i < kNQ; ++i) {
                            if (w.active(i, j)) {
                                const uint32_t par = (t_cnt[i] + (uint32_t)(j - w.lo_i[i])) & 1u;
                                // (the first P.V of an item overwrites O_i: safe, because every softmax thread finished the
                                // previous item's epilogue -- its TMEM reads of O_i included -- before it released this P)
#if MFA_PSPLIT
                                mbar_wait(&p_lo[i], par);
                                tc_fence_after();
                                MFA_TR(2, j, 1 + 3 * i);
                                issue_PV(i, sV, j > w.lo_i[i], 0, kBN / 32, false);
                                mbar_wait(&p_hi[i], par);
                                tc_fence_after();
                                MFA_TR(2, j, 2 + 3 * i);
                                issue_PV(i, sV, j > w.lo_i[i], kBN / 32, kBN / 16, true);
#else
                                mbar_wait(&p_hi[i], par);
                                tc_fence_after();
                                MFA_TR(2, j, 2 + 3 * i);
                                issue_PV(i, sV, j > w.lo_i[i], 0, kBN / 16, true);
#endif
                            }
                            if (w.active(i, j + 1)) issue_S(i, sK);
                            MFA_TR(2, j, 3 + 3 * i);
                        }
                        commit(&kv_empty[sV]);
                        if (has_next) commit(&kv_empty[sK]);
                    }
                    // every MMA that reads this item's Q tiles has been issued: release them to the producer
    #pragma unroll
                    for (int i = 0; i < kNQ; ++i) {
                        if (w.n_i[i] > 0) {
                            commit(&q_empty[i]);
                            ++q_ld[i];
                        }
                        t_cnt[i] += (uint32_t)w.count(i);
                    }
                }
            }
        }
    }
    } else {
      asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MFA_REG_SOFTMAX));
      if (warp < 4 * kNQ) {
        // ======================================================================= softmax warpgroups
        const int i = warp >> 2;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
        const uint32_t tS = tmem_base + lane_sel + Cfg::kTmemS + i * 128;
        const uint32_t tO = tmem_base + lane_sel + Cfg::kTmemO + i * DPAD;
        const float c2 = p.scale_log2;
        uint32_t t_cnt = 0;   // KV tiles this warpgroup has processed so far (parity base)
        for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
            const PrefillItem w = prefill_item<kNQ>(p, item);
            if (!w.valid) continue;
#ifdef MFA_TRACE
            tr_on = tr_cta && item == (int)blockIdx.x;
#endif
            const int m0 = w.m0, len_k = w.len_k, len_q = w.len_q;
            const int qrow = m0 + i * kBM + row_in_tile;            // row inside the sequence
            const int j_first = w.lo_i[i];
            const int n_mine = w.count(i);
            float m_run = -INFINITY, l_run = 0.f;

            for (int j = j_first; j < w.n_i[i]; ++j) {
                const uint32_t par = (t_cnt + (uint32_t)(j - j_first)) & 1u;
                MFA_TR(i, j, 0);
                mbar_wait(&s_full[i], par);
                tc_fence_after();
                MFA_TR(i, j, 1);
                uint32_t su[128];   // this thread's S row (fp32 bit patterns)
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&su[c * 32]));
                tmem_wait_ld();
                MFA_TR(i, j, 2);

                const int col0 = j * kBN;
                const bool need_mask = (col0 + kBN > len_k) || (p.has_hi && (col0 + kBN - 1 > m0 + i * kBM + w.shift_hi));
                if (need_mask) {
                    const int lim = p.has_hi ? min(len_k, qrow + 1 + w.shift_hi) : len_k;   // cols >= lim are masked
#pragma unroll
                    for (int c = 0; c < 128; ++c)
                        if (col0 + c >= lim) su[c] = 0xff800000u;   // -inf
                }
                if (p.has_lo && col0 < m0 + i * kBM + kBM - 1 + w.shift_lo) {   // the tile's last row has the largest bound
                    const int lim_lo = qrow + w.shift_lo;   // cols < lim_lo are masked
#pragma unroll
                    for (int c = 0; c < 128; ++c)
                        if (col0 + c < lim_lo) su[c] = 0xff800000u;
                }
                // row max: 8 independent 3-input max chains, then a tree
                float mx[8];
#pragma unroll
                for (int a = 0; a < 8; ++a) mx[a] = fmaxf(__uint_as_float(su[a * 16]), __uint_as_float(su[a * 16 + 1]));
#pragma unroll
                for (int a = 0; a < 8; ++a) {
#pragma unroll
                    for (int c = 2; c < 16; c += 2)
                        mx[a] = fmax3(mx[a], __uint_as_float(su[a * 16 + c]), __uint_as_float(su[a * 16 + c + 1]));
                }
                const float tmax = fmaxf(fmax3(mx[0], mx[1], mx[2]), fmax3(fmax3(mx[3], mx[4], mx[5]), mx[6], mx[7]));
                const float m_new = fmaxf(m_run, tmax);
                // lazy rescale: keep the stale max unless it grew by more than 2^8
                const bool grow = (m_new - m_run) * c2 > Cfg::kRescaleThreshold;   // also true when m_run == -inf
                const float m_use = grow ? m_new : m_run;
                const float factor = grow ? ((m_run == -INFINITY) ? 0.f : exp2f((m_run - m_new) * c2)) : 1.f;
                const float ms = (m_use == -INFINITY) ? 0.f : m_use * c2;
                m_run = m_use;

                if (j > j_first && __any_sync(0xffffffffu, grow)) {
                    // O_i currently holds sum_{t<j} P_t V_t: wait for PV(j-1), then rescale our row in place.
                    // Must happen before any part of P(j) is released to the MMA warp.
                    mbar_wait(&o_full[i], par ^ 1u);
                    tc_fence_after();
#pragma unroll
                    for (int c = 0; c < DPAD / 32; ++c) {
                        uint32_t o[32];
                        tmem_ld32(tO + c * 32, o);
                        tmem_wait_ld();
#pragma unroll
                        for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * factor);
                        tmem_st32(tO + c * 32, o);
                    }
                }

                // P = exp2(S*c2 - ms): packed FFMA2 for the scale; per pair of columns either two MUFU.EX2 or the
                // degree-3 Cody-Waite polynomial on the FMA pipe (pairs with (c & 3) in kPolyMask), because at 128x128
                // tiles the 16 MUFU lanes/clk cost as many cycles as the two MMAs.  Released to the MMA warp in two
                // halves so that P.V k-steps 0-3 run while the second half is still being exponentiated.
                MFA_TR(i, j, 3);
                const uint64_t c2_2 = f2_pack(c2, c2);
                const uint64_t nms_2 = f2_pack(-ms, -ms);
                uint64_t sum2[2] = {0ull, 0ull};
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    uint32_t pk[32];
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const int e = hf * 64 + 2 * c;
                        uint64_t x2 = f2_fma(f2_pack(__uint_as_float(su[e]), __uint_as_float(su[e + 1])), c2_2, nms_2);
                        uint64_t p2;
                        if (kPoly && ((Cfg::kPolyMask >> (c & 3)) & 1)) {
                            p2 = exp2_poly2(x2);
                        } else {
                            float x0, x1;
                            f2_unpack(x2, x0, x1);
                            p2 = f2_pack(ex2_fast(x0), ex2_fast(x1));
                        }
                        sum2[c & 1] = f2_add(sum2[c & 1], p2);
                        float p0, p1;
                        f2_unpack(p2, p0, p1);
                        pk[c] = Elem<T>::pack(p0, p1);
                    }
                    tmem_st32(tS + hf * 32, pk);
                    if (MFA_PSPLIT || hf == 1) {
                        tmem_wait_st();
                        tc_fence_before();
#if MFA_PARRIVE_WARP
                        __syncwarp();
                        if (lane == 0) mbar_arrive(hf == 0 ? &p_lo[i] : &p_hi[i]);
#else
                        mbar_arrive(hf == 0 ? &p_lo[i] : &p_hi[i]);
#endif
                    }
                    MFA_TR(i, j, 4 + hf);
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
                if (n_mine > 0) {
                    mbar_wait(&o_full[i], (t_cnt + (uint32_t)n_mine - 1u) & 1u);
                    tc_fence_after();
                }
                const float inv = (l_run == 0.f || l_run != l_run) ? 1.f : 1.f / l_run;   // prefill.cuh:604-605
                if (p.lse != nullptr && row_ok)   // natural-log LSE of the scaled scores (the reference drops it, api.cpp:436)
                    p.lse[(int64_t)w.b * p.lse_batch_stride + (int64_t)w.h * p.lse_head_stride + w.q_off + qrow] =
                        (l_run == 0.f || l_run != l_run) ? -INFINITY : (m_run * c2 + log2f(l_run)) * 0.6931471805599453f;
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
            }
            t_cnt += (uint32_t)n_mine;
        }
      }
    }

    // ---- teardown
    tc_fence_before();
    __syncthreads();
    if (warp == 10) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::kTmemCols);
    }
}

}  // namespace

template <typename T, int DPAD>
static cudaError_t launch_prefill_t(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                                    const PrefillKParams& p, cudaStream_t stream) {
    using Cfg = PrefillCfg<DPAD>;
    // exp2 placement.  Default: every exponential on the MUFU.  The kernel can put a quarter of them on the FMA pipe
    // (degree-3 polynomial, kPoly = true), which is 2 % faster for an ISOLATED launch on a cold, full-clock GPU (S=1024:
    // 913 vs 894 TFLOP/s with a sync between launches) -- but any sustained stream of launches sits at the board power
    // cap (sw_power_cap, 1.6-1.7 GHz), where the extra FMA work costs more clock than the relieved MUFU gives back:
    // back-to-back loops of >= 200 ms measure MUFU-only +1.5-2 % non-causal and +4-6 % causal at S=1024 and S=2048,
    // +1.7 % at the headline S=4096 (scripts/prefill_poly_ab.py, bench.py A/B on the same box).  Steady state is what a
    // serving loop sees, so MUFU-only is the default; MFA_PREFILL_POLY=1 selects the polynomial form.
    static const bool poly = [] { const char* e = getenv("MFA_PREFILL_POLY"); return e && e[0] == '1'; }();
    auto kern = poly ? prefill_fwd_kernel<T, DPAD, true> : prefill_fwd_kernel<T, DPAD, false>;
    static bool attr_done[64][2] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr_done[dev][poly]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[dev][poly] = true;
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
    if (!non_persistent && p.cu_seqlens_q == nullptr && items > sms) {
        ctas = sms;
        if (p.has_hi) {
            auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
            while (ctas > 1 && gcd(ctas, p.num_m_blocks) != 1) --ctas;
        }
    }
    dim3 grid((unsigned)ctas, 1, 1);
    kern<<<grid, kPrefillThreads, Cfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
    return cudaGetLastError();
}

int prefill_rows_per_cta(int dpad) { return (dpad <= 128 ? 2 : 1) * kPrefillBlockM; }

cudaError_t launch_prefill(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                           const PrefillKParams& p, bool is_bf16, int dpad, cudaStream_t stream) {
#define MFA_PREFILL_CASE(DP)                                                                     \
    case DP:                                                                                     \
        return is_bf16 ? launch_prefill_t<__nv_bfloat16, DP>(tmQ, tmK, tmV, p, stream)           \
                       : launch_prefill_t<__half, DP>(tmQ, tmK, tmV, p, stream);
    switch (dpad) {
        MFA_PREFILL_CASE(64)
        MFA_PREFILL_CASE(128)
        MFA_PREFILL_CASE(256)
        default:
            return cudaErrorInvalidValue;
    }
#undef MFA_PREFILL_CASE
}

}  // namespace mfa
