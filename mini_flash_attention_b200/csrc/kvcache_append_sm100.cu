// kvcache_append_sm100.cu -- in-place append of new K/V rows to a contiguous or paged KV cache, the step in front of a
// decode call.  The reference documents the semantics (interface.py:110-111, after flash-attn's
// flash_attn_with_kvcache(k=, v=)) but has no implementation; this is new code.  HBM-bound byte movement: one 16-byte
// vector per thread, the same thread moves the K and the V chunk; block (0,0) also publishes the new lengths.
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/mfa_b200.h"

namespace mfa {

namespace {

__global__ void __launch_bounds__(256) kvcache_append_kernel(const __grid_constant__ mfa_append_params_t p) {
    const int chunks = p.head_dim / 8;                       // 16-byte chunks per (row, head)
    const long long per_seq = (long long)p.s_new * p.kv_heads * chunks;
    const long long total = per_seq * p.batch;
    const int capacity = p.block_table ? p.max_blocks_per_seq * p.page_block_size : p.seqlen_k;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(idx / per_seq);
        long long r = idx - (long long)b * per_seq;
        const int t = (int)(r / ((long long)p.kv_heads * chunks));
        r -= (long long)t * p.kv_heads * chunks;
        const int h = (int)(r / chunks);
        const int c = (int)(r - (long long)h * chunks);
        const int pos = p.seqlens_k[b] + t;                  // logical row in the cache
        if (pos < 0 || pos >= capacity) continue;
        long long kb, vb;
        int row;
        if (p.block_table) {
            const int blk = p.block_table[(long long)b * p.block_table_batch_stride + pos / p.page_block_size];
            // a stale or negative page id must not become an out-of-bounds STORE (the decode's reads are bounds-checked by TMA)
            if (blk < 0 || (p.num_blocks > 0 && blk >= p.num_blocks)) continue;
            row = pos % p.page_block_size;
            kb = (long long)blk * p.k_batch_stride;
            vb = (long long)blk * p.v_batch_stride;
        } else {
            row = pos;
            kb = (long long)b * p.k_batch_stride;
            vb = (long long)b * p.v_batch_stride;
        }
        const uint16_t* ks = static_cast<const uint16_t*>(p.k_new) + (long long)b * p.knew_batch_stride +
                             (long long)t * p.knew_row_stride + (long long)h * p.knew_head_stride + c * 8;
        const uint16_t* vs = static_cast<const uint16_t*>(p.v_new) + (long long)b * p.vnew_batch_stride +
                             (long long)t * p.vnew_row_stride + (long long)h * p.vnew_head_stride + c * 8;
        uint16_t* kd = static_cast<uint16_t*>(p.k_cache) + kb + (long long)row * p.k_row_stride + (long long)h * p.k_head_stride + c * 8;
        uint16_t* vd = static_cast<uint16_t*>(p.v_cache) + vb + (long long)row * p.v_row_stride + (long long)h * p.v_head_stride + c * 8;
        const uint4 kx = *reinterpret_cast<const uint4*>(ks);
        const uint4 vx = *reinterpret_cast<const uint4*>(vs);
        *reinterpret_cast<uint4*>(kd) = kx;
        *reinterpret_cast<uint4*>(vd) = vx;
    }
    if (p.seqlens_out != nullptr && blockIdx.x == 0) {
        for (int b = threadIdx.x; b < p.batch; b += blockDim.x)
            p.seqlens_out[b] = max(0, min(p.seqlens_k[b] + p.s_new, capacity));
    }
}

}  // namespace

cudaError_t launch_kvcache_append(const mfa_append_params_t& p, int sm_count, cudaStream_t stream) {
    const long long total = (long long)p.batch * p.s_new * p.kv_heads * (p.head_dim / 8);
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)sm_count * 8;           // grid-stride above 8 resident blocks per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    kvcache_append_kernel<<<(unsigned)blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mfa
