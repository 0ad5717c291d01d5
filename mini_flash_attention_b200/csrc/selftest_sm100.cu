// selftest_sm100.cu -- single-tile tcgen05 self tests.  They isolate the UMMA descriptor / TMEM layout
// conventions the prefill kernel relies on (K-major SW128 operands from TMA, MN-major SW128 B operand,
// A operand read from TMEM as packed 16-bit pairs) from its pipeline logic, so that a wrong descriptor bit
// shows up as a failing 128x128 GEMM instead of a wrong attention output.
//   mode 0:  D[128 x 128] = A[128 x K] * B[128 x K]^T          (SS, both K-major; the Q K^T shape)
//   mode 1:  D[128 x N]   = P[128 x 128] * V[128 x N]          (TS, P in TMEM, V MN-major; the P V shape)
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "../../include/mfa_b200.h"
#include "ptx.cuh"
#include "tmap.h"

namespace mfa {
namespace {

struct SelfTestParams {
    const uint16_t* p;   // mode 1: P row-major [128][128] 16-bit
    float* d;            // output row-major [128][ncols]
    int mode, kdim, ncols, is_bf16;
    uint32_t lbo, sbo;   // mode 1 overrides (0 = defaults)
};

__global__ void __launch_bounds__(128) umma_selftest_kernel(const __grid_constant__ CUtensorMap tmA,
                                                            const __grid_constant__ CUtensorMap tmB,
                                                            const __grid_constant__ SelfTestParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_smem = smem;                 // up to 2 sub-tiles of 16 KB
    uint8_t* b_smem = smem + 4 * 16384;     // up to 4 sub-tiles of 16 KB
    uint64_t* bar_tma = reinterpret_cast<uint64_t*>(smem + 8 * 16384);
    uint64_t* bar_mma = bar_tma + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_tma + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(bar_tma, 1);
        mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t lane_sel = (uint32_t)(warp * 32) << 16;
    const uint32_t fmt = p.is_bf16 ? 1u : 0u;

    if (p.mode == 0) {
        const int sub = p.kdim / 64;
        if (threadIdx.x == 0) {
            mbar_arrive_expect_tx(bar_tma, (uint32_t)(2 * sub * 16384));
            for (int s = 0; s < sub; ++s) {
                tma_load_4d(a_smem + s * 16384, &tmA, bar_tma, s * 64, 0, 0, 0);
                tma_load_4d(b_smem + s * 16384, &tmB, bar_tma, s * 64, 0, 0, 0);
            }
            mbar_wait(bar_tma, 0);
            tc_fence_after();
            const uint32_t idesc = umma_idesc_f16(fmt, 128, 128, 0, 0);
            for (int ks = 0; ks < p.kdim / 16; ++ks) {
                const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
                umma_ss(tmem, umma_smem_desc(smem_u32(a_smem) + off, 16, 1024, 2),
                        umma_smem_desc(smem_u32(b_smem) + off, 16, 1024, 2), idesc, ks > 0);
            }
            umma_commit(bar_mma);
        }
    } else {
        // stage P into TMEM columns [256, 320) as packed 16-bit pairs: lane = row, column c = elements (2c, 2c+1)
        {
            const int row = warp * 32 + lane;
            const uint32_t* prow = reinterpret_cast<const uint32_t*>(p.p + (size_t)row * 128);
            uint32_t r[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
#pragma unroll
                for (int e = 0; e < 32; ++e) r[e] = prow[c * 32 + e];
                tmem_st32(tmem + lane_sel + 256 + c * 32, r);
            }
            tmem_wait_st();
        }
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        const int sub = p.ncols / 64;
        if (threadIdx.x == 0) {
            mbar_arrive_expect_tx(bar_tma, (uint32_t)(sub * 16384));
            for (int s = 0; s < sub; ++s) tma_load_4d(b_smem + s * 16384, &tmB, bar_tma, s * 64, 0, 0, 0);
            mbar_wait(bar_tma, 0);
            tc_fence_after();
            const uint32_t idesc = umma_idesc_f16(fmt, 128, (uint32_t)p.ncols, 0, 1);
            const uint32_t lbo = p.lbo ? p.lbo : 16384u;
            const uint32_t sbo = p.sbo ? p.sbo : 1024u;
            for (int ks = 0; ks < 8; ++ks)
                umma_ts(tmem, tmem + 256 + ks * 8, umma_smem_desc(smem_u32(b_smem) + ks * 2048, lbo, sbo, 2), idesc,
                        ks > 0);
            umma_commit(bar_mma);
        }
    }
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    {
        const int row = warp * 32 + lane;
        for (int c = 0; c < p.ncols / 32; ++c) {
            uint32_t r[32];
            tmem_ld32(tmem + lane_sel + c * 32, r);
            tmem_wait_ld();
            for (int e = 0; e < 32; ++e) p.d[(size_t)row * p.ncols + c * 32 + e] = __uint_as_float(r[e]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem, 512);
    }
}

}  // namespace
}  // namespace mfa

extern "C" int mfa_selftest_umma(int mode, const void* a, const void* b, float* d, int kdim, int ncols, int is_bf16,
                                 uint32_t lbo, uint32_t sbo, void* cuda_stream) {
    using namespace mfa;
    if (mode < 0 || mode > 1 || !b || !d) return MFA_ERR_INVALID_ARG;
    if (mode == 0 && (kdim != 64 && kdim != 128)) return MFA_ERR_INVALID_ARG;
    if (mode == 0) ncols = 128;
    if (mode == 1 && (ncols != 64 && ncols != 128 && ncols != 256)) return MFA_ERR_INVALID_ARG;
    CUtensorMap tmA, tmB;
    const uint32_t box[4] = {64, 128, 1, 1};
    if (mode == 0) {
        const uint64_t dims[4] = {(uint64_t)kdim, 128, 1, 1};
        const int64_t st[4] = {1, kdim, 0, 0};
        if (encode_tmap_4d(&tmA, a, is_bf16, dims, st, box)) return MFA_ERR_UNSUPPORTED;
        if (encode_tmap_4d(&tmB, b, is_bf16, dims, st, box)) return MFA_ERR_UNSUPPORTED;
    } else {
        const uint64_t dims[4] = {(uint64_t)ncols, 128, 1, 1};
        const int64_t st[4] = {1, ncols, 0, 0};
        if (encode_tmap_4d(&tmB, b, is_bf16, dims, st, box)) return MFA_ERR_UNSUPPORTED;
        tmA = tmB;
    }
    SelfTestParams p{};
    p.p = static_cast<const uint16_t*>(a);
    p.d = d;
    p.mode = mode;
    p.kdim = kdim;
    p.ncols = ncols;
    p.is_bf16 = is_bf16;
    p.lbo = lbo;
    p.sbo = sbo;
    const int smem = 8 * 16384 + 64 + 1024;
    static bool done = false;
    if (!done) {
        if (cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
            return MFA_ERR_CUDA;
        done = true;
    }
    umma_selftest_kernel<<<1, 128, smem, static_cast<cudaStream_t>(cuda_stream)>>>(tmA, tmB, p);
    return cudaGetLastError() == cudaSuccess ? MFA_OK : MFA_ERR_CUDA;
}
