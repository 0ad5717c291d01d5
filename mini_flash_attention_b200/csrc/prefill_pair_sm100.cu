// prefill_pair_sm100.cu -- FlashAttention forward (prefill), head_dim <= 128, as a CTA-PAIR kernel for sm_100a.
//
// Replaces /root/reference/csrc/mfa/prefill.cuh:711-803 (kernel) and flash.cu:11-34 (launch) for head_dim in
// (64, 128].  Design (measured reasons in DESIGN.md 4.1):
//
//   cluster = 2 CTAs = 256 query rows of one (sequence, head); each CTA owns ONE 128-row Q tile.
//   Every tcgen05.mma is cta_group::2 (M = 256 over the pair): CTA r supplies its 128 rows of the A operand from
//   ITS tensor memory and half of the B operand from ITS shared memory, so each CTA streams only half of every
//   K and V tile (K rows [64r, 64r+64) of the tile, V head-dim columns [64r, 64r+64)).
//
//   Both GEMMs read A from TMEM (TS form): S = Q K^T with Q staged once into TMEM as packed 16-bit pairs, and
//   O += P V with P written over the S columns it came from.  (Measured: an SS-form M=128 N=128 K=16 MMA takes
//   ~102 cycles because of the shared-memory A read, the TS form 72; scripts/umma_rate.py.)
//
//   TMEM per CTA (512 columns): S_0 [0,128)  S_1 [128,256)  O [256,384)  Q [384,448)  P [448,512).  S is double-
//   buffered and P has its own columns, so S(j+2) = Q K_{j+2}^T is issued as soon as the softmax threads have READ
//   S(j); the S of a warpgroup's next tile is always computed in the shadow of its current softmax.
//
//   warps 0-3  softmax warpgroup A: KV tiles 0, 2, 4, ...   thread t <-> query row t <-> TMEM lane t, whole S row
//   warps 4-7  softmax warpgroup B: KV tiles 1, 3, 5, ...   same rows.  The groups run half a period apart (one in
//              its exp2 phase while the other loads / reduces); the running row maximum is handed from tile j to
//              tile j+1 through shared memory (bar.arrive / bar.sync, 64-thread named barrier per warp pair); each
//              group keeps its own partial row sum, merged in the epilogue.
//   warp  8    TMA producer (both CTAs; completion bytes are counted on the LEADER CTA's barriers)
//   warp  9    MMA issuer (leader CTA only, one elected lane)
//   warp 10    TMEM allocator
//
// Semantics kept from the reference: scale = 1/sqrt(D) folded into exp2 (api.cpp:99-100); causal mask top-left
// aligned, key col masked iff col > row (prefill.cuh:416) with tile skipping (:749-752); key-length mask
// col >= len (:416); rows with an empty sum produce 0 (:604-605); GQA kv_head = head / group (:49); varlen through
// cu_seqlens (:19-33); paged K/V through block_table, gathered per page box (fixes quirk Q3 of SURVEY.md).
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kparams.h"
#include "ptx.cuh"

#ifndef MFA_PAIR_POLY_MASK
#define MFA_PAIR_POLY_MASK 0x8u   // column pairs (c & 3) whose exp2 runs on the FMA pipe instead of the MUFU
#endif
#ifndef MFA_PAIR_STAGES
#define MFA_PAIR_STAGES 8         // 16 KB half-tiles in the K/V ring
#endif

#ifdef MFA_TRACE
#define MFA_TR(role, it, ev)                                                                              \
    do {                                                                                                  \
        if (tr_on && (it) < 64) p.trace[((role) * 64 + (it)) * 8 + (ev)] = clock64();                      \
    } while (0)
#define MFA_CT(ev, val)                                                                                   \
    do {                                                                                                  \
        if (ct_on) p.trace[4 * 64 * 8 + blockIdx.x * 8 + (ev)] = (long long)(val);                          \
    } while (0)
#else
#define MFA_TR(role, it, ev) do { } while (0)
#define MFA_CT(ev, val) do { } while (0)
#endif

namespace mfa {

namespace {

template <typename T>
struct PElem;
template <>
struct PElem<__half> {
    static constexpr uint32_t kFmt = 0;
    static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
        __half2 h = __floats2half2_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};
template <>
struct PElem<__nv_bfloat16> {
    static constexpr uint32_t kFmt = 1;
    static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
        __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};

struct PairCfg {
    static constexpr int kBM = 128;                       // query rows per CTA
    static constexpr int kBN = 128;                       // KV rows per tile
    static constexpr int kStages = MFA_PAIR_STAGES;
    static constexpr int kHalfBytes = 16384;              // one K or V half-tile
    static constexpr int kQBytes = 2 * 128 * 128;         // Q tile: [2 sub(d/64)][128 rows][64 elems], SWIZZLE_128B
    static constexpr int kRingBytes = kStages * kHalfBytes;
    static constexpr int kXchgOffset = kQBytes + kRingBytes;        // float xchg[2][2][128] (max) + float lsum[2][128]
    static constexpr int kXchgBytes = (2 * 2 * 128 + 2 * 128) * 4;
    static constexpr int kBarOffset = kXchgOffset + kXchgBytes;
    static constexpr int kSmemBytes = kBarOffset + 512 + 1024;
    static constexpr uint32_t kTmemCols = 512;
    static constexpr uint32_t kTmemS = 0;                 // S_b at b * 128
    static constexpr uint32_t kTmemO = 256;
    static constexpr uint32_t kTmemP = 384;               // P_b (16-bit pairs, 64 columns) at kTmemP + b * 64
    static constexpr float kRescaleThreshold = 8.0f;      // log2 domain
    static constexpr uint32_t kPolyMask = MFA_PAIR_POLY_MASK;
};

// Q tile as the K-major A operand of S = Q K^T: [sub(d/64)][128 rows][64 elems], SWIZZLE_128B; K-step = 32 B.
__device__ __forceinline__ uint64_t pair_desc_q(uint32_t tile_addr, int kstep) {
    const uint32_t addr = tile_addr + (kstep >> 2) * (128 * 128) + (kstep & 3) * 32;
    return umma_smem_desc(addr, /*LBO*/ 16, /*SBO: 8 rows x 128 B*/ 1024, /*SW128*/ 2);
}
// K half-tile as the K-major B operand of S = Q K^T: [sub(d/64)][64 kv rows][64 elems], SWIZZLE_128B; K-step = 32 B.
__device__ __forceinline__ uint64_t pair_desc_k(uint32_t tile_addr, int kstep) {
    const uint32_t addr = tile_addr + (kstep >> 2) * (64 * 128) + (kstep & 3) * 32;
    return umma_smem_desc(addr, /*LBO*/ 16, /*SBO: 8 rows x 128 B*/ 1024, /*SW128*/ 2);
}
// V half-tile as the MN-major B operand of O += P V: [128 kv rows][64 elems (d)], SWIZZLE_128B; K-step = 16 rows.
__device__ __forceinline__ uint64_t pair_desc_v(uint32_t tile_addr, int kstep) {
    const uint32_t addr = tile_addr + kstep * (16 * 128);
    return umma_smem_desc(addr, /*LBO (single 64-wide atom along N: unused)*/ 128 * 128, /*SBO*/ 1024, /*SW128*/ 2);
}

// 2^x for two packed floats on the FMA pipe (see prefill_sm100.cu): Cody-Waite split with the 1.5*2^23 magic
// constant, degree-3 polynomial with p(0) == 1 (max rel. error 1.0e-4 < half an ulp of the fp16/bf16 P).
__device__ __forceinline__ uint64_t pair_exp2_poly2(uint64_t x2) {
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

// Hand-off of the running row maximum between the two warpgroups (tile j -> tile j+1): the publisher stores its value to
// shared memory and does bar.arrive, the consumer does bar.sync on the same 64-thread named barrier and reads it.
// The two threads of a query row live in warps w and w + 4; ids 1-4 carry even tiles, 5-8 odd tiles.
__device__ __forceinline__ void pair_bar_arrive(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void pair_bar_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void pair_bar_sync_all_softmax() { asm volatile("bar.sync 9, 256;" ::: "memory"); }

template <typename T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPrefillThreads, 1)
    prefill_pair_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const __grid_constant__ PrefillKParams p) {
    using Cfg = PairCfg;
    constexpr int kStages = Cfg::kStages;
    constexpr int kBM = Cfg::kBM, kBN = Cfg::kBN;

    extern __shared__ uint8_t smem_raw[];
    // the dynamic smem base has the same offset in both CTAs of the pair, so every object below does too
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_smem = smem;
    uint8_t* ring = smem + Cfg::kQBytes;
    float* xm = reinterpret_cast<float*>(smem + Cfg::kXchgOffset);     // [2 (tile parity)][128] published running max
    float* lsum = xm + 2 * 128;                                        // [2 (wg)][128] partial row sums (epilogue)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kBarOffset);
    uint64_t* q_full = bars;                 // [1]  leader: its own arrive.expect_tx, Q tiles of BOTH CTAs landed
    uint64_t* s_full = bars + 1;             // [2]  per CTA: S(j) complete in buffer j&1 (multicast commit)
    uint64_t* s_free = bars + 3;             // [2]  leader: 8 warp arrivals, S(j) has been read into registers
    uint64_t* p_lo = bars + 5;               // [2]  leader: 8 warp arrivals, P(j) columns of KV rows 0-63 written (buffer j&1)
    uint64_t* p_hi = bars + 7;               // [2]  leader: KV rows 64-127
    uint64_t* o_full = bars + 9;             // [2]  per CTA: P V(j) complete, barrier j&1 (multicast commit)
    uint64_t* done_bar = bars + 11;          // [1]  per CTA: the last P V of this work item is complete
    uint64_t* tail_bar = bars + 12;          // [1]  per CTA: local TMA completion for V tiles that need zeroed tail rows
    uint64_t* vfix_bar = bars + 13;          // [1]  leader: 2 producer arrivals, the (last) V tile with zeroed tail rows is ready
    uint64_t* kv_full = bars + 14;           // [kStages] leader: its own arrive.expect_tx with the bytes of BOTH halves
    uint64_t* kv_empty = kv_full + kStages;  // [kStages] per CTA: stage consumed (multicast commit)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(kv_empty + kStages);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
#ifdef MFA_TRACE
    const bool tr_on = p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (warp == 9 || warp == 8 || (warp & 3) == 0);
    // CTA lifetime stamps of the CTAs (x < 64, 0, gridDim.z / 2): steady state, not the first wave
    const bool ct_on = p.trace && blockIdx.y == 0 && blockIdx.z == gridDim.z / 2 && blockIdx.x < 64 && threadIdx.x == 0;
    MFA_CT(0, clock64());
    MFA_CT(6, globaltimer_ns());
#endif

    // ---- work item.  Causal: heaviest (largest m) pairs first.
    const int npairs = p.num_m_blocks;
    const int pair = p.is_causal ? (npairs - 1 - (int)(blockIdx.x >> 1)) : (int)(blockIdx.x >> 1);
    const int h = blockIdx.y;
    const int b = blockIdx.z;
    const int kh = h / p.group;
    int q_off = 0, k_off = 0, len_q = p.seqlen_q, len_k = p.seqlen_k;
    if (p.cu_seqlens_q) {
        q_off = p.cu_seqlens_q[b];
        len_q = p.cu_seqlens_q[b + 1] - q_off;
        k_off = p.cu_seqlens_k[b];
        len_k = p.cu_seqlens_k[b + 1] - k_off;
    }
    const int m0_pair = pair * (2 * kBM);
    if (m0_pair >= len_q) return;   // the whole cluster exits (prefill.cuh:724-726)
    const int m0 = m0_pair + rank * kBM;

    // number of KV tiles the PAIR walks (both CTAs take part in every cta_group::2 MMA)
    int n_tiles = (len_k + kBN - 1) / kBN;
    if (p.is_causal) n_tiles = min(n_tiles, (min(len_q, m0_pair + 2 * kBM) - 1) / kBN + 1);

    // ---- one-time setup
    if (threadIdx.x == 0) {
        mbar_init(q_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&s_free[i], 8);
            mbar_init(&o_full[i], 1);
            mbar_init(&p_lo[i], 8);
            mbar_init(&p_hi[i], 8);
        }
        mbar_init(done_bar, 1);
        mbar_init(tail_bar, 1);
        mbar_init(vfix_bar, 2);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        fence_barrier_init();
    }
    if (warp == 10) {
        tmem_alloc_pair(tmem_slot, Cfg::kTmemCols);
        tmem_relinquish_pair();
    }
    tc_fence_before();
    cluster_sync();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    MFA_CT(1, clock64());

    // Register re-balancing (as in prefill_sm100.cu): the launch gives every thread 168 registers; the data-movement
    // warpgroup hands most of its share to the softmax warpgroups, which hold a whole 128-column S row per thread.
    if (warp >= 8) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    if (warp == 8) {
        // ======================================================================= TMA producer (both CTAs)
        // Warp-uniform control flow with one ELECTED lane issuing: under a divergent `if (lane == 0)` ptxas cannot keep
        // the TMA operands in uniform registers and every cp.async.bulk.tensor cost ~250 cycles to issue (measured: one
        // half-tile load per ~700 cycles, the MMA warp starved for K/V).
        if (elect_one()) {
            prefetch_tensormap(&tmQ);
            prefetch_tensormap(&tmK);
            prefetch_tensormap(&tmV);
            // this CTA's Q tile (rows past the tensor are zero-filled; rows past len_q of a packed batch belong to the
            // next sequence: finite, row-local, never stored)
            if (rank == 0) mbar_arrive_expect_tx(q_full, 2 * Cfg::kQBytes);
            const int q_row = p.cu_seqlens_q ? (q_off + m0) : m0;
            tma_load_4d_pair(q_smem, &tmQ, q_full, 0, q_row, h, p.cu_seqlens_q ? 0 : b);
            tma_load_4d_pair(q_smem + 16384, &tmQ, q_full, 64, q_row, h, p.cu_seqlens_q ? 0 : b);
        }
        const int32_t* bt = p.block_table ? p.block_table + (int64_t)b * p.block_table_batch_stride : nullptr;
        const bool packed_k = (p.cu_seqlens_k != nullptr) && (bt == nullptr);
        // rows >= len_k of a V tile must be zero (P is 0 there, but 0 * NaN = NaN).  For fixed-length batches TMA's
        // out-of-bounds fill does it; for packed / paged K/V the producer warp zeroes them.
        const bool need_tail_fix = (bt != nullptr) || packed_k;
        const int box = p.kv_box_rows;            // 64, or gcd(page, 64) when paged
        const int row_base = packed_k ? k_off : 0;
        const int b_coord = packed_k ? 0 : b;
        uint32_t tail_phase = 0;
        int it = 0;
        // consumption order of the MMA warp: K_0, K_1, then per step j = 0..n-1: K_{j+2}, V_{j-1}; finally V_{n-1}
        for (int step = 0; step < 2 * n_tiles + 4; ++step) {
            bool load_v;
            int j;
            if (step < 2) { load_v = false; j = step; }
            else if ((step & 1) == 0) { load_v = false; j = (step >> 1) + 1; }   // step 2,4,6.. -> K_2, K_3, ...
            else { load_v = true; j = (step >> 1) - 2; }                         // step 3,5,7.. -> V_-1 (none), V_0, V_1, ...
            if (j < 0 || j >= n_tiles) continue;
            const int s = it % kStages;
            const uint32_t ph = (it / kStages) & 1;
            ++it;
            uint8_t* dst = ring + s * Cfg::kHalfBytes;
            const int tile_row0 = j * kBN;
            MFA_TR(3, it - 1, 0);
            mbar_wait(&kv_empty[s], ph ^ 1);   // all lanes (warp-uniform), see the note on elect_one below
            __syncwarp();
            MFA_TR(3, it - 1, 1);
#ifdef MFA_TRACE
            if (tr_on && it - 1 < 64) p.trace[((3) * 64 + (it - 1)) * 8 + 2] = (load_v ? 1000 : 0) + j;
#endif
            // Barrier protocol of a stage: the LEADER alone arrives, posting the byte count of BOTH halves; the TMA
            // loads of both CTAs complete on the leader's barrier.
            if (!load_v) {
                // ---- K half: rows [tile_row0 + 64 rank, +64), all d
                const int r0 = tile_row0 + rank * 64;
                if (elect_one()) {
                    if (bt) {
                        auto nchunks = [&](int rk) {
                            const int valid = min(64, len_k - (tile_row0 + rk * 64));
                            return valid > 0 ? (valid + box - 1) / box : 0;
                        };
                        const int nch = nchunks(rank);
                        if (rank == 0) mbar_arrive_expect_tx(&kv_full[s], (uint32_t)((nchunks(0) + nchunks(1)) * box * 128 * 2));
                        for (int c = 0; c < nch; ++c) {
                            const int row = r0 + c * box;
                            const int pg = row / p.page;
                            const int blk = bt[pg];
                            tma_load_4d_pair(dst + c * box * 128, &tmK, &kv_full[s], 0, row - pg * p.page, kh, blk);
                            tma_load_4d_pair(dst + 8192 + c * box * 128, &tmK, &kv_full[s], 64, row - pg * p.page, kh, blk);
                        }
                    } else {
                        if (rank == 0) mbar_arrive_expect_tx(&kv_full[s], 2 * Cfg::kHalfBytes);
                        tma_load_4d_pair(dst, &tmK, &kv_full[s], 0, row_base + r0, kh, b_coord);
                        tma_load_4d_pair(dst + 8192, &tmK, &kv_full[s], 64, row_base + r0, kh, b_coord);
                    }
                }
            } else {
                // ---- V half: rows [tile_row0, +128), d columns [64 rank, +64)
                const int rows_valid = min(len_k - tile_row0, kBN);
                const bool fix = need_tail_fix && rows_valid < kBN;
                if (elect_one()) {
                    const int nch = bt ? (rows_valid + box - 1) / box : kBN / box;
                    const uint32_t bytes = (uint32_t)(nch * box * 128);
                    if (fix) mbar_arrive_expect_tx(tail_bar, bytes);
                    else if (rank == 0) mbar_arrive_expect_tx(&kv_full[s], 2 * bytes);
                    for (int c = 0; c < nch; ++c) {
                        const int row = tile_row0 + c * box;
                        int c1 = row_base + row, c3 = b_coord;
                        if (bt) {
                            const int pg = row / p.page;
                            c1 = row - pg * p.page;
                            c3 = bt[pg];
                        }
                        if (fix) tma_load_4d(dst + c * box * 128, &tmV, tail_bar, rank * 64, c1, kh, c3);
                        else tma_load_4d_pair(dst + c * box * 128, &tmV, &kv_full[s], rank * 64, c1, kh, c3);
                    }
                }
                if (fix) {
                    mbar_wait(tail_bar, tail_phase);
                    tail_phase ^= 1;
                    for (int r = rows_valid; r < kBN; ++r) *reinterpret_cast<uint32_t*>(dst + r * 128 + lane * 4) = 0u;
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_leader_release(vfix_bar);
                }
            }
        }
        // drain: the last multicast "stage consumed" arrivals must have landed in THIS CTA before it may exit
        if (lane == 0) {
            for (int i = max(0, it - kStages); i < it; ++i) mbar_wait(&kv_empty[i % kStages], (uint32_t)(i / kStages) & 1u);
        }
        __syncwarp();
    } else if (warp == 9) {
        // ======================================================================= MMA issuer (leader CTA only)
        // The whole warp runs the warp-uniform control flow; one elected lane issues (keeps descriptors in uniform
        // registers -- a divergent `if (lane == 0)` costs ~110 cycles per issued MMA).
        //
        // Issue order:  S(0) S(1) | S(2) | S(3) PV(0) | S(4) PV(1) | ... | PV(n-1).   S(j+2) needs only that the softmax
        // threads have READ S(j) (s_free), because P has its own TMEM columns; so the S of the next tile of each
        // warpgroup is always computed in the shadow of its current softmax.
        if (rank == 0 && n_tiles > 0) {
            constexpr uint32_t idescS = umma_idesc_f16(PElem<T>::kFmt, 2 * kBM, kBN, 0, 0);
            constexpr uint32_t idescO = umma_idesc_f16(PElem<T>::kFmt, 2 * kBM, 128, 0, 1);
            const uint32_t ring_addr = smem_u32(ring);
            const uint32_t q_addr = smem_u32(q_smem);
            const uint32_t tO = tmem_base + Cfg::kTmemO;

            auto issue_S = [&](int buf, int stage) {
                if (elect_one()) {
                    const uint32_t ka = ring_addr + stage * Cfg::kHalfBytes;
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks)
                        umma_ss_pair(tmem_base + Cfg::kTmemS + buf * 128, pair_desc_q(q_addr, ks), pair_desc_k(ka, ks), idescS, ks > 0 ? 1u : 0u);
                    umma_commit_pair(&s_full[buf]);
                    umma_commit_pair(&kv_empty[stage]);
                }
                __syncwarp();
            };
            auto issue_PV = [&](int j, int stage, int half, bool last) {
                if (elect_one()) {
                    const uint32_t va = ring_addr + stage * Cfg::kHalfBytes;
                    const uint32_t tP = tmem_base + Cfg::kTmemP + (j & 1) * 64;
#pragma unroll
                    for (int k4 = 0; k4 < 4; ++k4)
                        umma_ts_pair(tO, tP + (half * 4 + k4) * 8, pair_desc_v(va, half * 4 + k4), idescO, (j > 0 || half || k4 > 0) ? 1u : 0u);
                    if (half) {
                        umma_commit_pair(&o_full[j & 1]);
                        umma_commit_pair(&kv_empty[stage]);
                        if (last) umma_commit_pair(done_bar);
                    }
                }
                __syncwarp();
            };

            // the last V tile of a packed / paged sequence has its rows >= len_k zeroed by the producers (see there)
            const bool v_fix_last = (p.cu_seqlens_k != nullptr) && (len_k % kBN) != 0 && n_tiles == (len_k + kBN - 1) / kBN;
            mbar_wait(q_full, 0);
            tc_fence_after();
            int it = 0;
            for (int i = 0; i < 2 && i < n_tiles; ++i) {
                const int sK = it % kStages;
                mbar_wait(&kv_full[sK], (it / kStages) & 1);
                ++it;
                tc_fence_after();
                issue_S(i, sK);
            }
            auto do_PV = [&](int j) {
                const int sV = it % kStages;
                if (v_fix_last && j == n_tiles - 1) mbar_wait(vfix_bar, 0);   // zero-tailed V tile: signalled by both producers
                else mbar_wait(&kv_full[sV], (it / kStages) & 1);
                ++it;
                MFA_TR(2, j, 5);
                mbar_wait(&p_lo[j & 1], (uint32_t)(j >> 1) & 1u);
                tc_fence_after();
                MFA_TR(2, j, 1);
                issue_PV(j, sV, 0, false);
                mbar_wait(&p_hi[j & 1], (uint32_t)(j >> 1) & 1u);
                tc_fence_after();
                MFA_TR(2, j, 2);
                issue_PV(j, sV, 1, j == n_tiles - 1);
                MFA_TR(2, j, 3);
            };
            for (int j = 0; j < n_tiles; ++j) {
                if (j + 2 < n_tiles) {
                    const int sK = it % kStages;
                    mbar_wait(&kv_full[sK], (it / kStages) & 1);
                    ++it;
                    mbar_wait(&s_free[j & 1], (uint32_t)(j >> 1) & 1u);   // S(j) is in registers: its buffer may be overwritten
                    tc_fence_after();
                    MFA_TR(2, j, 0);
                    issue_S(j & 1, sK);
                    MFA_TR(2, j, 4);
                }
                if (j >= 1) do_PV(j - 1);
            }
            do_PV(n_tiles - 1);
        }
    }
    } else {
      asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
      {
        // ======================================================================= softmax warpgroups
        // Warpgroup g (warps 4g .. 4g+3) takes the KV tiles j = g, g+2, g+4, ...; thread t of either group owns query
        // row t (TMEM lane t) and a whole 128-column S row.  The two groups run half a period apart, so that one is in
        // its exp2 phase while the other loads / reduces; the only coupling is the running row maximum, handed from
        // tile j to tile j+1 through shared memory (bar.arrive / bar.sync on a 64-thread named barrier per warp pair).
        const int wg = warp >> 2;
        const int wq = warp & 3;
        const int row_in_tile = wq * 32 + lane;
        const int qrow = m0 + row_in_tile;                      // row inside the sequence
        const bool row_ok = qrow < len_q;
        const uint32_t lane_sel = (uint32_t)(wq * 32) << 16;
        const uint32_t tO = tmem_base + lane_sel + Cfg::kTmemO;
        const uint32_t tP = tmem_base + lane_sel + Cfg::kTmemP + wg * 64;   // P(j) lives in buffer j&1 == wg
        const float c2 = p.scale_log2;

        // l_part: this thread's share of the row sum, expressed relative to the reference maximum m_ref
        float m_ref = -INFINITY, l_part = 0.f;
        for (int j = wg; j < n_tiles; j += 2) {
            const int buf = j & 1;
            const uint32_t tS = tmem_base + lane_sel + Cfg::kTmemS + buf * 128;
            MFA_TR(wg, j, 0);
            mbar_wait(&s_full[buf], (uint32_t)(j >> 1) & 1u);
            tc_fence_after();
            MFA_TR(wg, j, 1);
            if (j == 0) MFA_CT(2, clock64());
#ifdef MFA_DBG_NOSOFTMAX
            // perf experiment only (wrong results): no softmax work -> pure TMA + MMA pipeline rate
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&s_free[buf]);
            if (j > 1) mbar_wait(&o_full[j & 1], (uint32_t)((j - 2) >> 1) & 1u);
            if (lane == 0) { mbar_arrive_leader(&p_lo[buf]); mbar_arrive_leader(&p_hi[buf]); }
            if (j > 0) mbar_wait(&o_full[(j - 1) & 1], (uint32_t)((j - 1) >> 1) & 1u);
            continue;
#endif
            uint32_t su[128];   // this thread's S row (fp32 bit patterns)
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&su[c * 32]));
            tmem_wait_ld();
            // S(j) is in registers: let the MMA warp overwrite this buffer with S(j+2)
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&s_free[buf]);
            MFA_TR(wg, j, 2);

            const int col0 = j * kBN;
            const bool need_mask = (col0 + kBN > len_k) || (p.is_causal && (col0 + kBN - 1 > m0));
            if (need_mask) {
                const int lim = p.is_causal ? min(len_k, qrow + 1) : len_k;   // cols >= lim are masked
#pragma unroll
                for (int c = 0; c < 128; ++c)
                    if (col0 + c >= lim) su[c] = 0xff800000u;   // -inf
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

            // running maximum after tile j-1, from the other warpgroup
            float m_prev = -INFINITY;
            if (j > 0) {
                pair_bar_sync(1 + ((j - 1) & 1) * 4 + wq);
                m_prev = xm[((j - 1) & 1) * 128 + row_in_tile];
            }
            const float m_new = fmaxf(m_prev, tmax);
            // lazy rescale: keep the stale max unless it grew by more than 2^8
            const bool grow = (m_new - m_prev) * c2 > Cfg::kRescaleThreshold;   // also true when m_prev == -inf < m_new
            const float m_use = grow ? m_new : m_prev;
            xm[buf * 128 + row_in_tile] = m_use;
            pair_bar_arrive(1 + buf * 4 + wq);
            const float ms = (m_use == -INFINITY) ? 0.f : m_use * c2;
            // bring my partial row sum to the new reference
            if (m_ref != m_use) {
                l_part = (m_ref == -INFINITY) ? 0.f : l_part * exp2f((m_ref - m_use) * c2);
                m_ref = m_use;
            }

            // o_full[1-wg] counts the OTHER group's P V(j-1), (j-3), ...: this thread observes each of its phases exactly
            // once per tile -- here when a rescale needs it early, else after P(j) has been released (parity waits
            // cannot skip phases).
            bool seen_prev = false;
            if (j > 0 && __any_sync(0xffffffffu, grow)) {
                seen_prev = true;
                // O holds sum_{t<j} P_t V_t relative to m_prev: wait for P V(j-1), then rescale my row in place.
                // Must happen before any part of P(j) is released to the MMA warp.
                const float factor = grow ? ((m_prev == -INFINITY) ? 0.f : exp2f((m_prev - m_new) * c2)) : 1.f;
                mbar_wait(&o_full[(j - 1) & 1], (uint32_t)((j - 1) >> 1) & 1u);
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t o[32];
                    tmem_ld32(tO + c * 32, o);
                    tmem_wait_ld();
#pragma unroll
                    for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * factor);
                    tmem_st32(tO + c * 32, o);
                }
                tmem_wait_st();
            }
            MFA_TR(wg, j, 3);

            // P = exp2(S*c2 - ms): packed FFMA2 for the scale; per pair of columns either two MUFU.EX2 or the degree-3
            // polynomial on the FMA pipe (pairs with (c & 3) in kPolyMask).  Released to the MMA warp in two halves so
            // that P.V k-steps 0-3 run while the second half is still being exponentiated.
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
                    if ((Cfg::kPolyMask >> (c & 3)) & 1) {
                        p2 = pair_exp2_poly2(x2);
                    } else {
                        float x0, x1;
                        f2_unpack(x2, x0, x1);
                        p2 = f2_pack(ex2_fast(x0), ex2_fast(x1));
                    }
                    sum2[c & 1] = f2_add(sum2[c & 1], p2);
                    float p0, p1;
                    f2_unpack(p2, p0, p1);
                    pk[c] = PElem<T>::pack(p0, p1);
                }
                // P buffer wg is re-used every second tile: P V(j-2) must have consumed P(j-2) (o_full[wg], whose every
                // phase this thread observes here; it completed about a full tile time ago)
                if (hf == 0 && j > 1) mbar_wait(&o_full[j & 1], (uint32_t)((j - 2) >> 1) & 1u);
                tmem_st32(tP + hf * 32, pk);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(hf == 0 ? &p_lo[buf] : &p_hi[buf]);
                MFA_TR(wg, j, 4 + hf);
            }
            float sa, sb, sc, sd;
            f2_unpack(sum2[0], sa, sb);
            f2_unpack(sum2[1], sc, sd);
            l_part += (sa + sb) + (sc + sd);
            if (j > 0 && !seen_prev) mbar_wait(&o_full[(j - 1) & 1], (uint32_t)((j - 1) >> 1) & 1u);
        }

        MFA_CT(3, clock64());
        // ---- epilogue: both partial sums at the final reference maximum, O / l -> out dtype -> global.
        // tcgen05.ld is warp-collective, so every lane takes part in the TMEM loads; only the global stores are guarded.
        {
            if (n_tiles > 0) {
                const int last = n_tiles - 1;
                if (wg != (last & 1)) {   // the other warpgroup ran the last tile: adopt its final maximum
                    pair_bar_sync(1 + (last & 1) * 4 + wq);
                    const float m_fin = xm[(last & 1) * 128 + row_in_tile];
                    if (m_ref != m_fin) l_part = (m_ref == -INFINITY) ? 0.f : l_part * exp2f((m_ref - m_fin) * c2);
                    m_ref = m_fin;
                }
            }
            lsum[wg * 128 + row_in_tile] = l_part;
            pair_bar_sync_all_softmax();
            const float l_tot = l_part + lsum[(wg ^ 1) * 128 + row_in_tile];
            T* orow = static_cast<T*>(p.o) + (int64_t)h * p.o_head_stride + wg * 64 +
                      (p.cu_seqlens_q ? (int64_t)(q_off + qrow) * p.o_row_stride
                                      : (int64_t)b * p.o_batch_stride + (int64_t)qrow * p.o_row_stride);
            if (n_tiles > 0) {
                mbar_wait(done_bar, 0);
                tc_fence_after();
            }
            const float inv = (l_tot == 0.f || l_tot != l_tot) ? 1.f : 1.f / l_tot;   // prefill.cuh:604-605
            if (p.lse != nullptr && row_ok && wg == 0)
                p.lse[(int64_t)b * p.lse_batch_stride + (int64_t)h * p.lse_head_stride + q_off + qrow] =
                    (l_tot == 0.f || l_tot != l_tot) ? -INFINITY : (m_ref * c2 + log2f(l_tot)) * 0.6931471805599453f;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t o[32];
                if (n_tiles > 0) {   // uniform
                    tmem_ld32(tO + wg * 64 + c * 32, o);
                    tmem_wait_ld();
                } else {
#pragma unroll
                    for (int e = 0; e < 32; ++e) o[e] = 0u;
                }
                if (row_ok) {
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        if (wg * 64 + c * 32 + v * 8 < p.head_dim) {
                            uint4 w;
                            w.x = PElem<T>::pack(__uint_as_float(o[v * 8 + 0]) * inv, __uint_as_float(o[v * 8 + 1]) * inv);
                            w.y = PElem<T>::pack(__uint_as_float(o[v * 8 + 2]) * inv, __uint_as_float(o[v * 8 + 3]) * inv);
                            w.z = PElem<T>::pack(__uint_as_float(o[v * 8 + 4]) * inv, __uint_as_float(o[v * 8 + 5]) * inv);
                            w.w = PElem<T>::pack(__uint_as_float(o[v * 8 + 6]) * inv, __uint_as_float(o[v * 8 + 7]) * inv);
                            *reinterpret_cast<uint4*>(orow + c * 32 + v * 8) = w;
                        }
                    }
                }
            }
        }
      }
    }

    if (warp < 8) MFA_CT(4, clock64());
    // ---- teardown: nobody may leave (or free TMEM) while the peer can still signal its barriers / read its smem
    tc_fence_before();
    cluster_sync();
    if (warp == 10) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
    }
    MFA_CT(5, clock64());
    MFA_CT(7, globaltimer_ns());
}

}  // namespace

int prefill_pair_smem_bytes() { return PairCfg::kSmemBytes; }

cudaError_t launch_prefill_pair(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                                const PrefillKParams& p, bool is_bf16, cudaStream_t stream) {
    static bool attr_done[64][2] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const void* kern = is_bf16 ? (const void*)prefill_pair_kernel<__nv_bfloat16> : (const void*)prefill_pair_kernel<__half>;
    if (dev < 64 && !attr_done[dev][is_bf16 ? 1 : 0]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PairCfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done[dev][is_bf16 ? 1 : 0] = true;
    }
    dim3 grid(2 * p.num_m_blocks, p.heads, p.batch);   // static __cluster_dims__(2,1,1): pairs along x
    if (is_bf16) prefill_pair_kernel<__nv_bfloat16><<<grid, kPrefillThreads, PairCfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
    else prefill_pair_kernel<__half><<<grid, kPrefillThreads, PairCfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
    return cudaGetLastError();
}

}  // namespace mfa
