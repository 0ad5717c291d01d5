// merge.cuh -- block-cooperative LSE merge of S normalised partial rows (the arithmetic of the reference's combine
// kernel, /root/reference/csrc/mfa/decode.cuh:665-755, with a max-subtracted log-sum-exp).
//
// One CTA of kMergeThreads (8 warps) per output row.  The merge of a long-context decode is pure latency: 37 splits x
// 32 rows is a few hundred KB, and the first form (one warp per row, split after split) spent 25 us on it -- more than
// the 17 us split-KV kernel in front of it (scripts/decode_fixed_cost.py).  Here the splits of a row are dealt over the
// 8 warps and every load of a warp is independent of the others (weights come from lanes by shuffle, addresses do not
// depend on data), so a row costs about three memory round trips: the LSEs, the rows, the shared-memory reduction.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace mfa {

constexpr int kMergeThreads = 256;
constexpr int kMergeWarps = kMergeThreads / 32;

// Result in WARP 0: acc[i] = merged columns lane*4 + 128*i .. +3 (i < 2, head_dim <= 256), glse in every thread.
// smem: kMergeWarps * 256 floats.  All threads of the CTA must call it.
template <bool kPacked>
__device__ __forceinline__ void block_merge_splits(const float* __restrict__ o_part, const float* __restrict__ lse_part,
                                                   int S, int rows, int row, int D, float4 (&acc)[2], float& glse,
                                                   float* smem) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pstride = kPacked ? (D + 1) : D;
    auto lse_of = [&](int s) -> float {
        return kPacked ? o_part[((int64_t)s * rows + row) * pstride + D] : lse_part[(int64_t)s * rows + row];
    };
    // every warp computes the row's global LSE itself (no block barrier for it)
    float mx = -INFINITY;
    for (int s = lane; s < S; s += 32) mx = fmaxf(mx, lse_of(s));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float tot = 0.f;
    for (int s = lane; s < S; s += 32) {
        const float l = lse_of(s);
        tot += (l == -INFINITY) ? 0.f : __expf(l - mx);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    glse = (mx == -INFINITY) ? -INFINITY : mx + __logf(tot);

    float4 part[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
    for (int base = 0; base < S; base += 128) {
        float wk[4];   // lane holds the weights of splits base + 32 k + lane
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int s = base + 32 * k + lane;
            const float l = (s < S) ? lse_of(s) : -INFINITY;
            wk[k] = (l == -INFINITY) ? 0.f : __expf(l - glse);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (base + 32 * k >= S) break;   // warp-uniform
#pragma unroll
            for (int jj = 0; jj < 32 / kMergeWarps; ++jj) {
                const int j = warp + kMergeWarps * jj;   // this warp's splits of the chunk
                const int s = base + 32 * k + j;
                const bool ok = s < S;
                const float w = ok ? __shfl_sync(0xffffffffu, wk[k], j) : 0.f;
                const float* src = o_part + ((int64_t)(ok ? s : 0) * rows + row) * pstride;   // always a valid address
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int col = lane * 4 + 128 * i;
                    if (col < D) {
                        float4 v;
                        if (kPacked) {   // rows of D+1 floats are not 16-byte aligned
                            v = make_float4(src[col], src[col + 1], src[col + 2], src[col + 3]);
                        } else {
                            v = *reinterpret_cast<const float4*>(src + col);
                        }
                        if (w != 0.f) {
                            part[i].x += v.x * w;
                            part[i].y += v.y * w;
                            part[i].z += v.z * w;
                            part[i].w += v.w * w;
                        }
                    }
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) *reinterpret_cast<float4*>(smem + warp * 256 + lane * 4 + 128 * i) = part[i];
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int w = 0; w < kMergeWarps; ++w) {
                const float4 v = *reinterpret_cast<const float4*>(smem + w * 256 + lane * 4 + 128 * i);
                a.x += v.x;
                a.y += v.y;
                a.z += v.z;
                a.w += v.w;
            }
            acc[i] = a;
        }
    }
}

}  // namespace mfa
