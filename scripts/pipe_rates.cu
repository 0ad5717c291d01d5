// pipe_rates.cu -- which execution pipe do the softmax's instructions use on sm_100, and at what rate?
//
// The prefill trace (scripts/trace_prefill.py) shows a softmax warp spending ~543 cycles on 68 FMNMX3 and ~768 cycles on
// 64 MUFU.EX2 + 32 F2FP per half tile: 8 cycles per warp instruction in both cases, as if 3-input max and the
// fp32 -> bf16x2 pack shared the quarter-rate pipe of the MUFU.  This micro-benchmark measures it: one warp per SM
// sub-partition (4 warps of one CTA) issues independent chains of ONE instruction kind (or a mix of two) and reports
// cycles per warp instruction.  If a mix of A and B costs cost(A) + cost(B) they share a pipe; if it costs max() they
// do not.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/pipe_rates scripts/pipe_rates.cu && scripts/pipe_rates
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 512

enum Kind { MUFU = 0, F2FP_BF16, F2FP_F16, FMNMX3, FMNMX2, PRMT_, IADD_, LOP3_, FFMA2_, FADD2_, FMUL_, MIX_MUFU_F2FP, MIX_MUFU_FMNMX3,
            MIX_MUFU_FMNMX2, MIX_MUFU_PRMT, MIX_MUFU_IADD, MIX_F2FP_FMNMX3, NKINDS };
static const char* kNames[NKINDS] = {"MUFU.EX2", "F2FP.BF16 pack", "F2FP.F16 pack", "FMNMX3 (3-input max)", "FMNMX (2-input max)", "PRMT",
                                     "IADD3", "LOP3", "FFMA2", "FADD2", "FMUL", "mix MUFU + F2FP.BF16", "mix MUFU + FMNMX3",
                                     "mix MUFU + FMNMX", "mix MUFU + PRMT", "mix MUFU + IADD3", "mix F2FP.BF16 + FMNMX3"};

template <int K>
__device__ __forceinline__ void op(float& a, float& b, uint32_t& u, uint32_t& w) {
    if (K == MUFU) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a));
    if (K == F2FP_BF16) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(u) : "f"(a), "f"(b));
    if (K == F2FP_F16) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u) : "f"(a), "f"(b));
    if (K == FMNMX3) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a) : "f"(b), "f"(__uint_as_float(w)));
    if (K == FMNMX2) asm volatile("max.f32 %0, %0, %1;" : "+f"(a) : "f"(__uint_as_float(w)));
    if (K == PRMT_) asm volatile("prmt.b32 %0, %0, %1, 0x7632;" : "+r"(u) : "r"(w));
    if (K == IADD_) asm volatile("add.u32 %0, %0, %1;" : "+r"(u) : "r"(w));
    if (K == LOP3_) asm volatile("and.b32 %0, %0, %1;" : "+r"(u) : "r"(w));
    if (K == FMUL_) asm volatile("mul.f32 %0, %0, %1;" : "+f"(a) : "f"(b));
}

template <int K>
__global__ void bench(long long* out, float seed) {
    float a[CHAINS], b[CHAINS];
    uint32_t u[CHAINS], w[CHAINS];
    uint64_t p[CHAINS], q2;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
        a[c] = seed * (c + 1) * 1e-3f;
        b[c] = seed * (c + 2) * 1e-3f;
        u[c] = threadIdx.x + c;
        w[c] = threadIdx.x * 3 + c;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(p[c]) : "f"(a[c]), "f"(b[c]));
    }
    asm volatile("mov.b64 %0, {%1, %2};" : "=l"(q2) : "f"(seed), "f"(seed));
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (K < FFMA2_ || K == FMUL_) op<K>(a[c], b[c], u[c], w[c]);
            if (K == FFMA2_) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(p[c]) : "l"(q2));
            if (K == FADD2_) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[c]) : "l"(q2));
            if (K == MIX_MUFU_F2FP) { op<MUFU>(a[c], b[c], u[c], w[c]); op<F2FP_BF16>(b[c], b[c], u[c], w[c]); }
            if (K == MIX_MUFU_FMNMX3) { op<MUFU>(a[c], b[c], u[c], w[c]); op<FMNMX3>(b[c], b[c], u[c], w[c]); }
            if (K == MIX_MUFU_FMNMX2) { op<MUFU>(a[c], b[c], u[c], w[c]); op<FMNMX2>(b[c], b[c], u[c], w[c]); }
            if (K == MIX_MUFU_PRMT) { op<MUFU>(a[c], b[c], u[c], w[c]); op<PRMT_>(b[c], b[c], u[c], w[c]); }
            if (K == MIX_MUFU_IADD) { op<MUFU>(a[c], b[c], u[c], w[c]); op<IADD_>(b[c], b[c], u[c], w[c]); }
            if (K == MIX_F2FP_FMNMX3) { op<F2FP_BF16>(a[c], b[c], u[c], w[c]); op<FMNMX3>(b[c], b[c], u[c], w[c]); }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
    uint32_t x = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
        s += a[c] + b[c];
        x ^= u[c] ^ w[c] ^ (uint32_t)p[c];
    }
    if (s == 12345.678f && x == 42u) out[1] = 1;   // keep the chains alive
    if (threadIdx.x == 0) out[0] = t1 - t0;
}

template <int K>
static void run(long long* d_out, int warps) {
    bench<K><<<1, 32 * warps>>>(d_out, 1.0f);
    cudaDeviceSynchronize();
    bench<K><<<1, 32 * warps>>>(d_out, 1.0f);
    cudaDeviceSynchronize();
    long long cyc = 0;
    cudaMemcpy(&cyc, d_out, sizeof(cyc), cudaMemcpyDeviceToHost);
    const int per_iter = (K >= MIX_MUFU_F2FP) ? 2 * CHAINS : CHAINS;
    printf("%-26s warps/SMSP=%d  cycles per warp-instruction: %.2f%s\n", kNames[K], warps / 4, (double)cyc / ((double)ITERS * per_iter) * 1.0,
           (K >= MIX_MUFU_F2FP) ? "  (per instruction of the pair; x2 = per pair)" : "");
}

template <int K>
static void run_all(long long* d) {
    run<K>(d, 4);
    run<K>(d, 8);
    if constexpr (K + 1 < NKINDS) run_all<K + 1>(d);
}

int main() {
    long long* d = nullptr;
    cudaMalloc(&d, 64);
    cudaMemset(d, 0, 64);
    run_all<0>(d);
    printf("(one CTA; 4 warps = one warp per SM sub-partition, 8 warps = two per sub-partition; 8 independent chains per thread)\n");
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
