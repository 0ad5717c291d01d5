// xchg_sm100.cu -- the exchange step of the KV-sharded (multi-GPU) decode, fused into the LSE-merge kernel.
//
// a long KV axis is cut over R GPUs; every rank computes a normalised partial (O_r, LSE_r) over its rows and
// the final result is  LSE = logsumexp_r LSE_r,  O = sum_r exp(LSE_r - LSE) O_r  (the arithmetic of the reference's
// combine kernel, /root/reference/csrc/mfa/decode.cuh:665-755, with a max-subtracted log-sum-exp).  The first form
// of this path packs the partial, calls ONE NCCL all-gather and then the combine kernel; that costs ~0.25 ms of fixed
// latency per step (profiles/multigpu_r01b.jsonl).  Here ONE kernel per rank does all of it over NVLink peer memory:
//
//   1. merge this rank's split-KV partials (S, rows, D) -> one normalised row + LSE        (registers)
//   2. PUSH the row into slot [rank] of every peer's gather buffer (plain 16-byte stores through the peer mapping),
//      fence.sys, then st.release.sys of the step's epoch into the peer's flag [rank][row]
//   3. spin (ld.acquire.sys, bounded by a clock64() timeout) until the R flags of this row carry the epoch
//   4. merge the R rows that the peers pushed into the LOCAL buffer, cast, store
//
// One warp owns one (batch, head) row end to end, flags are per (source rank, row), so there is no dependency between
// warps or CTAs of a grid and no co-residency requirement; a rank never blocks another rank's push.  The gather buffer
// is double-buffered by epoch parity: rank A can start step n+2 only after every peer pushed step n+1, i.e. after each
// peer's step-n kernel finished (stream order), so the parity-n rows are no longer being read when A overwrites them.
// Flags carry the monotonically increasing epoch and are never reset.  The epoch itself lives in DEVICE memory (ctrs[1]):
// every warp reads it on entry, and the last warp of the grid to finish bumps it for the next launch, so the launch
// parameters do not change from step to step and the whole step can be captured in a CUDA graph and replayed.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

#include "kparams.h"
#include "merge.cuh"

namespace mfa {
namespace {

template <typename T>
struct XPack;
template <>
struct XPack<__half> {
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};
template <>
struct XPack<__nv_bfloat16> {
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// polling load: relaxed at system scope (never served from a stale L1 line); the acquire is ONE fence behind the loop
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One CTA of 8 warps per row: the local splits are dealt over the warps (merge.cuh); warp 0 then owns the row -- push,
// flag, wait, cross-rank merge, store.  Lanes own columns lane*4 + 128*i (head_dim <= 256 => at most 2 float4 per lane).
template <typename T>
__global__ void __launch_bounds__(kMergeThreads) xchg_merge_kernel(const __grid_constant__ XchgKParams p) {
    __shared__ __align__(16) float red[kMergeWarps * 256];
    const int row = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int D = p.head_dim;
    const int DS = D + 4;   // row stride of the gather buffer: [O_0 .. O_{D-1}, LSE, pad x3] keeps rows 16-byte aligned
    // this step's epoch: launches completed so far + 1 (ctrs = {status, epoch counter, rows done})
    const uint32_t epoch = *reinterpret_cast<volatile const uint32_t*>(p.ctrs + 1) + 1u;

    // ---- 1. local merge over this rank's splits
    float4 mine[2];
    float my_lse;
    block_merge_splits<false>(p.o_part, p.lse_part, p.num_splits, p.rows, row, D, mine, my_lse, red);
    if (threadIdx.x >= 32) return;

    // ---- 2. push to every rank's gather buffer (own included), then publish the epoch
    const int parity = (int)(epoch & 1u);
    const int64_t slot_off = (((int64_t)parity * p.world + p.rank) * p.max_rows + row) * DS;
    for (int r = 0; r < p.world; ++r) {
        float* dst = p.peer_data[r] + slot_off;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int col = lane * 4 + 128 * i;
            if (col < D) *reinterpret_cast<float4*>(dst + col) = mine[i];
        }
        if (lane == 0) *reinterpret_cast<float4*>(dst + D) = make_float4(my_lse, 0.f, 0.f, 0.f);
    }
    // one system-scope fence per lane behind its own stores, the warp barrier, then the flags: the release store of lane r
    // is ordered behind every lane's data (fence + barrier are cumulative).  (Round 1 fenced three times per row; a
    // system fence waits for the NVLink write acknowledgements, and ncu showed this kernel at 25 us against 9 us for the
    // same merge without an exchange.)
    __threadfence_system();
    __syncwarp();
    if (lane < p.world) st_release_sys(p.peer_flags[lane] + (int64_t)p.rank * p.max_rows + row, epoch);

    // ---- 3. wait for the R pushes of this row into the local buffer.  The first polls are back to back (a peer that
    // runs in step arrives within a few microseconds: an NVLink round trip); only a late peer is waited for with
    // __nanosleep between polls.  A peer that never arrives within the timeout must not turn into silently wrong
    // attention: the row is written as NaN (output and LSE) and the sticky status word names the missing rank.
    bool timed_out = false;
    if (lane < p.world) {
        const uint32_t* f = p.peer_flags[p.rank] + (int64_t)lane * p.max_rows + row;
        const long long t0 = clock64();
        int polls = 0;
        while ((int32_t)(ld_relaxed_sys(f) - epoch) < 0) {
            if (++polls > 128) {
                if (clock64() - t0 > p.timeout_cycles) {
                    atomicExch(p.ctrs, 1u + (uint32_t)lane);   // reported (and cleared) through mfa_xchg_status / _clear_status
                    timed_out = true;
                    break;
                }
                __nanosleep(32);
            }
        }
    }
    asm volatile("fence.acq_rel.sys;" ::: "memory");   // acquire: the flags were seen, now the rows may be read
    timed_out = __any_sync(0xffffffffu, timed_out);   // (also the warp barrier behind the acquiring lanes)

    // ---- 4. merge the R partial rows (L2 loads: the lines were written by peers, never trust L1)
    const float* base = p.peer_data[p.rank] + ((int64_t)parity * p.world * p.max_rows + row) * DS;
    const int64_t rstride = (int64_t)p.max_rows * DS;
    const float l_r = (lane < p.world) ? __ldcg(base + lane * rstride + D) : -INFINITY;   // lane r: rank r's LSE
    float gmx = l_r;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) gmx = fmaxf(gmx, __shfl_xor_sync(0xffffffffu, gmx, o));
    float gtot = (l_r == -INFINITY) ? 0.f : __expf(l_r - gmx);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) gtot += __shfl_xor_sync(0xffffffffu, gtot, o);
    const float glse = (gmx == -INFINITY) ? -INFINITY : gmx + __logf(gtot);
    const float w_r = (l_r == -INFINITY) ? 0.f : __expf(l_r - glse);
    float4 acc[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
#pragma unroll 4
    for (int r = 0; r < p.world; ++r) {
        const float w = __shfl_sync(0xffffffffu, w_r, r);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int col = lane * 4 + 128 * i;
            if (col < D) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(base + r * rstride + col));
                if (w != 0.f) {
                    acc[i].x += v.x * w;
                    acc[i].y += v.y * w;
                    acc[i].z += v.z * w;
                    acc[i].w += v.w * w;
                }
            }
        }
    }
    if (timed_out) {   // visible in the data, not only in the status word
        const float qnan = __int_as_float(0x7fc00000);
        acc[0] = acc[1] = make_float4(qnan, qnan, qnan, qnan);
    }
    T* out = static_cast<T*>(p.out) + (int64_t)(row / p.heads) * p.out_batch_stride + (int64_t)(row % p.heads) * p.out_head_stride;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int col = lane * 4 + 128 * i;
        if (col < D) {
            uint2 pk;
            pk.x = XPack<T>::pack(acc[i].x, acc[i].y);
            pk.y = XPack<T>::pack(acc[i].z, acc[i].w);
            *reinterpret_cast<uint2*>(out + col) = pk;
        }
    }
    if (lane == 0 && p.lse_out) p.lse_out[row] = timed_out ? __int_as_float(0x7fc00000) : glse;

    // ---- 5. the last row of the grid closes the step: every CTA has read the epoch before it got here
    __syncwarp();
    if (lane == 0) {
        __threadfence();
        if (atomicAdd(p.ctrs + 2, 1u) == (uint32_t)p.rows - 1u) {
            p.ctrs[2] = 0u;
            __threadfence();
            atomicAdd(p.ctrs + 1, 1u);
        }
    }
}

}  // namespace

cudaError_t launch_xchg_merge(const XchgKParams& p, bool is_bf16, cudaStream_t stream) {
    // The kernel spins until the peers arrive, so it must never stand in the way of the kernels it is waiting for.  An SM
    // re-partitions L1 / shared memory only when idle: a grid of these CTAs (8 KB static smem, small carve-out) on every
    // SM would block a split-KV kernel (100+ KB dynamic smem) of ANOTHER rank that shares the GPU (ranks as streams of
    // one process, tests/test_kv_exchange_gpu.py: seen as a timeout with 8 ranks x 32 rows).  Asking for the maximum
    // shared-memory carve-out keeps both kernels in the same SM configuration.
    static bool attr_done[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr_done[dev]) {
        cudaFuncSetAttribute(xchg_merge_kernel<__nv_bfloat16>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(xchg_merge_kernel<__half>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr_done[dev] = true;
    }
    dim3 grid(p.rows);
    if (is_bf16)
        xchg_merge_kernel<__nv_bfloat16><<<grid, kMergeThreads, 0, stream>>>(p);
    else
        xchg_merge_kernel<__half><<<grid, kMergeThreads, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mfa
