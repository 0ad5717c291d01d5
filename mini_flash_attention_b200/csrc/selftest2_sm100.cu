// selftest2_sm100.cu -- CTA-pair (cluster of 2, tcgen05 cta_group::2) self tests and MMA-rate probes.
// They pin down the conventions the paired prefill kernel relies on:
//   mode 0:  D[256 x 128] = A[256 x 128] * B[128 x 128]^T   SS; CTA r holds A rows [128r,128r+128) and the B (N) rows
//            [64r, 64r+64); every CTA's TMEM receives its own 128 rows x all 128 columns.          (the Q K^T shape)
//   mode 1:  D[256 x 128] = P[256 x 128] * V[128 x 128]     TS; P rows of CTA r in its TMEM, V MN-major with the N
//            (head-dim) columns [64r, 64r+64) in CTA r.                                            (the P V shape)
//   mode 2/3: single-CTA (cta_group::1) SS / TS with the same per-CTA tile, for rate comparison only.
// `reps` > 1 repeats the 8 K-steps (accumulating) and reports clock64() cycles from first issue to completion.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "../../include/mfa_b200.h"
#include "ptx.cuh"
#include "tmap.h"

namespace mfa {
namespace {

struct SelfTest2Params {
    const uint16_t* p;   // mode 1/3: P row-major [256][128] 16-bit
    float* d;            // output row-major [256][128]
    long long* cycles;   // optional: [0] = cycles for reps*8 MMAs to complete, [1] = cycles until all were issued
    int mode, is_bf16, reps;
};

template <bool kPair>
__global__ void __launch_bounds__(128) umma_selftest2_kernel(const __grid_constant__ CUtensorMap tmA,
                                                             const __grid_constant__ CUtensorMap tmB,
                                                             const __grid_constant__ SelfTest2Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_smem = smem;                 // A: 2 sub-tiles [128 rows][64] = 32 KB
    uint8_t* b_smem = smem + 2 * 16384;     // B: up to 32 KB
    uint64_t* bar_tma = reinterpret_cast<uint64_t*>(smem + 4 * 16384);
    uint64_t* bar_mma = bar_tma + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_tma + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = kPair ? cluster_ctarank() : 0u;
    const bool ts = (p.mode & 1) != 0 || p.mode >= 4;   // A from TMEM
    const bool bk = p.mode >= 4;                        // modes 4/5: B is K-major (the S = Q K^T shape with Q in TMEM)

    if (threadIdx.x == 0) {
        mbar_init(bar_tma, kPair ? 2 : 1);
        mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        if (kPair) {
            tmem_alloc_pair(tmem_slot, 512);
            tmem_relinquish_pair();
        } else {
            tmem_alloc(tmem_slot, 512);
            tmem_relinquish();
        }
    }
    tc_fence_before();
    if (kPair) cluster_sync(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t lane_sel = (uint32_t)(warp * 32) << 16;
    const uint32_t fmt = p.is_bf16 ? 1u : 0u;
    const int nb = kPair ? 64 : 128;        // B rows (SS) held by this CTA

    if (ts) {
        // stage this CTA's P rows into TMEM columns [256, 320)
        const int row = (int)rank * 128 + warp * 32 + lane;
        const uint32_t* prow = reinterpret_cast<const uint32_t*>(p.p + (size_t)row * 128);
        uint32_t r[32];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
#pragma unroll
            for (int e = 0; e < 32; ++e) r[e] = prow[c * 32 + e];
            tmem_st32(tmem + lane_sel + 256 + c * 32, r);
        }
        tmem_wait_st();
        tc_fence_before();
        if (kPair) cluster_sync(); else __syncthreads();
        tc_fence_after();
    }

    if (threadIdx.x == 0) {
        // ---- loads: every CTA loads its own operand parts; completion is counted on the leader's barrier
        if (bk) {
            const uint32_t bytes = 2 * nb * 128;
            if (kPair) mbar_arrive_expect_tx_leader(bar_tma, bytes); else mbar_arrive_expect_tx(bar_tma, bytes);
            for (int s = 0; s < 2; ++s) {
                if (kPair) tma_load_4d_pair(b_smem + s * nb * 128, &tmB, bar_tma, s * 64, (int)rank * 64, 0, 0);
                else tma_load_4d(b_smem + s * nb * 128, &tmB, bar_tma, s * 64, 0, 0, 0);
            }
        } else if (!ts) {
            const uint32_t bytes = 2 * 16384 + 2 * nb * 128;
            if (kPair) mbar_arrive_expect_tx_leader(bar_tma, bytes); else mbar_arrive_expect_tx(bar_tma, bytes);
            for (int s = 0; s < 2; ++s) {
                if (kPair) {
                    tma_load_4d_pair(a_smem + s * 16384, &tmA, bar_tma, s * 64, (int)rank * 128, 0, 0);
                    tma_load_4d_pair(b_smem + s * nb * 128, &tmB, bar_tma, s * 64, (int)rank * 64, 0, 0);
                } else {
                    tma_load_4d(a_smem + s * 16384, &tmA, bar_tma, s * 64, 0, 0, 0);
                    tma_load_4d(b_smem + s * nb * 128, &tmB, bar_tma, s * 64, 0, 0, 0);
                }
            }
        } else {
            // V: [128 kv rows][64 d] sub-tiles; pair: the sub-tile of d columns [64 rank, +64); single: both
            const int nsub = kPair ? 1 : 2;
            const uint32_t bytes = nsub * 16384;
            if (kPair) mbar_arrive_expect_tx_leader(bar_tma, bytes); else mbar_arrive_expect_tx(bar_tma, bytes);
            for (int s = 0; s < nsub; ++s) {
                if (kPair) tma_load_4d_pair(b_smem, &tmB, bar_tma, (int)rank * 64, 0, 0, 0);
                else tma_load_4d(b_smem + s * 16384, &tmB, bar_tma, s * 64, 0, 0, 0);
            }
        }
    }
    // ---- MMA issue: the whole warp 0 of the leader runs the (warp-uniform) loop, one elected lane issues, so that
    // ptxas keeps the descriptors in uniform registers (a divergent `if (threadIdx.x == 0)` costs ~110 cycles/MMA)
    if (warp == 0 && rank == 0) {
        mbar_wait(bar_tma, 0);
        tc_fence_after();
        const uint32_t idesc = umma_idesc_f16(fmt, kPair ? 256 : 128, 128, 0, (ts && !bk) ? 1 : 0);
        const uint32_t a_addr = smem_u32(a_smem), b_addr = smem_u32(b_smem);
        const long long t0 = clock64();
        for (int rep = 0; rep < p.reps; ++rep) {
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    const uint32_t acc = (rep > 0 || ks > 0) ? 1u : 0u;
                    if (!ts) {
                        const uint64_t ad = umma_smem_desc(a_addr + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024, 2);
                        const uint64_t bd = umma_smem_desc(b_addr + (ks >> 2) * nb * 128 + (ks & 3) * 32, 16, 1024, 2);
                        if (kPair) umma_ss_pair(tmem, ad, bd, idesc, acc); else umma_ss(tmem, ad, bd, idesc, acc);
                    } else if (bk) {
                        const uint64_t bd = umma_smem_desc(b_addr + (ks >> 2) * nb * 128 + (ks & 3) * 32, 16, 1024, 2);
                        if (kPair) umma_ts_pair(tmem, tmem + 256 + ks * 8, bd, idesc, acc);
                        else umma_ts(tmem, tmem + 256 + ks * 8, bd, idesc, acc);
                    } else {
                        const uint64_t bd = umma_smem_desc(b_addr + ks * 2048, 16384, 1024, 2);
                        if (kPair) umma_ts_pair(tmem, tmem + 256 + ks * 8, bd, idesc, acc);
                        else umma_ts(tmem, tmem + 256 + ks * 8, bd, idesc, acc);
                    }
                }
            }
            __syncwarp();
        }
        const long long t_issued = clock64();   // all MMAs accepted by the tensor pipe's queue
        if (elect_one()) {
            if (kPair) umma_commit_pair(bar_mma); else umma_commit(bar_mma);
        }
        __syncwarp();
        mbar_wait(bar_mma, 0);
        const long long t1 = clock64();
        if (p.cycles && lane == 0) {
            p.cycles[0] = t1 - t0;
            p.cycles[1] = t_issued - t0;   // issue-side time: shows how far the issuing thread can run ahead
        }
    }
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    {
        const int row = (int)rank * 128 + warp * 32 + lane;
        for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            tmem_ld32(tmem + lane_sel + c * 32, r);
            tmem_wait_ld();
            for (int e = 0; e < 32; ++e) p.d[(size_t)row * 128 + c * 32 + e] = __uint_as_float(r[e]);
        }
    }
    tc_fence_before();
    if (kPair) cluster_sync(); else __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        if (kPair) tmem_dealloc_pair(tmem, 512); else tmem_dealloc(tmem, 512);
    }
}

}  // namespace
}  // namespace mfa

extern "C" int mfa_selftest_umma2(int mode, const void* a, const void* b, float* d, int is_bf16, int reps,
                                  long long* cycles, void* cuda_stream) {
    using namespace mfa;
    if (mode < 0 || mode > 5 || !a || !b || !d || reps < 1) return MFA_ERR_INVALID_ARG;
    const bool pair = mode < 2 || mode == 4;
    const bool bk = mode >= 4;
    const bool ts = (mode & 1) != 0 || bk;
    CUtensorMap tmA, tmB;
    const uint64_t dimsA[4] = {128, 256, 1, 1};
    const uint64_t dimsB[4] = {128, 128, 1, 1};
    const int64_t st[4] = {1, 128, 0, 0};
    const uint32_t boxA[4] = {64, 128, 1, 1};
    const uint32_t boxB[4] = {64, ((ts && !bk) || !pair) ? 128u : 64u, 1, 1};
    if (!ts && encode_tmap_4d(&tmA, a, is_bf16, dimsA, st, boxA)) return MFA_ERR_UNSUPPORTED;
    if (encode_tmap_4d(&tmB, b, is_bf16, dimsB, st, boxB)) return MFA_ERR_UNSUPPORTED;
    if (ts) tmA = tmB;
    SelfTest2Params p{};
    p.p = static_cast<const uint16_t*>(a);
    p.d = d;
    p.cycles = cycles;
    p.mode = mode;
    p.is_bf16 = is_bf16;
    p.reps = reps;
    const int smem = 4 * 16384 + 64 + 1024;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    if (pair) {
        auto kern = umma_selftest2_kernel<true>;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return MFA_ERR_CUDA;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(2, 1, 1);
        cfg.blockDim = dim3(128, 1, 1);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p) != cudaSuccess) return MFA_ERR_CUDA;
    } else {
        auto kern = umma_selftest2_kernel<false>;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return MFA_ERR_CUDA;
        kern<<<1, 128, smem, stream>>>(tmA, tmB, p);
    }
    return cudaGetLastError() == cudaSuccess ? MFA_OK : MFA_ERR_CUDA;
}
