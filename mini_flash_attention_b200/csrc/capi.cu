// capi.cu -- the extern "C" boundary declared in include/mfa_b200.h.
//
// Host-side equivalent of the reference's L1/L2 layers: ForwardParams interpretation
// (/root/reference/csrc/mfa/api.cpp:30-101), dispatch on dtype x head_dim x {fixed | varlen | paged}
// (/root/reference/csrc/mfa/flash.cu:74-93, static_switch.h:3-52), the split-KV heuristic and workspace
// layout (api.cpp:269-340), and launch.  Every failure is reported through a return code +
// mfa_last_error(); nothing here aborts the process and nothing falls back to a CPU path.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>

#include "../../include/mfa_b200.h"
#include "kparams.h"
#include "tmap.h"

namespace mfa {
cudaError_t launch_decode(const CUtensorMap& tmK, const CUtensorMap& tmV, const DecodeKParams& p, bool is_bf16,
                          int dpad, cudaStream_t stream);
cudaError_t launch_combine(const CombineKParams& p, bool is_bf16, cudaStream_t stream);
cudaError_t launch_prefill(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                           const PrefillKParams& p, bool is_bf16, int dpad, cudaStream_t stream);
int prefill_rows_per_cta(int dpad);
cudaError_t launch_kvcache_append(const mfa_append_params_t& p, int sm_count, cudaStream_t stream);
cudaError_t launch_xchg_merge(const XchgKParams& p, bool is_bf16, cudaStream_t stream);
}  // namespace mfa

// Exchange buffer of the KV-sharded decode (include/mfa_b200.h; kernel in xchg_sm100.cu).  One cudaMalloc per rank:
//   floats [2 parities][world][max_rows][head_dim + 4]  |  uint32 flags [world][max_rows]  |  uint32 {status, epoch, done, pad}
struct mfa_xchg {
    int world, rank, max_rows, head_dim, device;
    size_t data_bytes, flag_bytes, bytes;
    void* local;
    void* peer[MFA_XCHG_MAX_WORLD];
    bool ipc_opened[MFA_XCHG_MAX_WORLD];
    bool connected;
    long long timeout_cycles;
};

namespace {

thread_local char g_err[512] = "";
long long* g_trace = nullptr;   // tuning aid, see mfa_debug_set_trace
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char* what) {
    return fail(MFA_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

int gcd_int(int a, int b) {
    while (b) {
        int t = a % b;
        a = b;
        b = t;
    }
    return a;
}

int sm_count() {
    static int cached[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

int check_device() {
    static int ok[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev >= 0 && dev < 64 && ok[dev]) return MFA_OK;
    int major = 0;
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
    if (major != 10) return fail(MFA_ERR_NO_DEVICE, "mfa_b200 kernels are sm_100a only; device %d is sm_%d", dev, major * 10);
    if (dev >= 0 && dev < 64) ok[dev] = 1;
    return MFA_OK;
}

int dpad_of(int d) { return d <= 64 ? 64 : (d <= 128 ? 128 : 256); }
int prefill_dpad_of(int d) {   // (the decode kernel has no 192 form)
    static const bool no192 = [] { const char* e = getenv("MFA_PREFILL_NO_D192"); return e && e[0] == '1'; }();   // A/B: pad 136..192 to 256 as before
    return d <= 64 ? 64 : (d <= 128 ? 128 : ((d <= 192 && !no192) ? 192 : 256));
}

// checks shared by every entry point (api.cpp:126-153)
int check_common(const mfa_params_t* p) {
    if (!p) return fail(MFA_ERR_INVALID_ARG, "params is NULL");
    if (!p->q_ptr || !p->k_ptr || !p->v_ptr || !p->o_ptr) return fail(MFA_ERR_INVALID_ARG, "q/k/v/o pointers must be non-NULL");
    if (p->head_dim <= 0 || p->head_dim > 256)
        return fail(MFA_ERR_INVALID_ARG, "head dimension must be less than or equal to 256 (got %d)", p->head_dim);
    if (p->head_dim % 8) return fail(MFA_ERR_UNSUPPORTED, "head dimension must be a multiple of 8 (got %d)", p->head_dim);
    if (p->heads <= 0 || p->kv_heads <= 0 || p->heads % p->kv_heads)
        return fail(MFA_ERR_INVALID_ARG, "number of key/value heads must be divisible by number of query heads (%d %% %d)",
                    p->heads, p->kv_heads);
    if (p->batch <= 0) return fail(MFA_ERR_INVALID_ARG, "batch must be positive");
    return MFA_OK;
}

bool mult8(int64_t x) { return (x % 8) == 0; }

}  // namespace

extern "C" {

int mfa_abi_version(void) { return MFA_B200_ABI_VERSION; }
void mfa_debug_set_trace(void* device_buffer) { g_trace = static_cast<long long*>(device_buffer); }
const char* mfa_last_error(void) { return g_err; }
uint64_t mfa_launch_count(void) { return g_launches.load(); }

int mfa_device_supported(void) {
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
    return major == 10;
}

// ------------------------------------------------------------------------------------------ prefill
int mfa_prefill_fwd(const mfa_params_t* p, void* cuda_stream) {
    int rc = check_common(p);
    if (rc) return rc;
    if ((rc = check_device())) return rc;
    if (p->kv_fp8) return fail(MFA_ERR_UNSUPPORTED, "an e4m3 K/V is supported by the decode entry points only");
    const bool varlen = p->cu_seqlens_q != nullptr;
    const bool paged = p->block_table != nullptr;
    if (varlen && !p->cu_seqlens_k) return fail(MFA_ERR_INVALID_ARG, "cu_seqlens_k must be given with cu_seqlens_q");
    if (paged && !varlen) return fail(MFA_ERR_INVALID_ARG, "block_table requires the varlen entry (cu_seqlens)");
    if (p->seqlen_q <= 0) return fail(MFA_ERR_INVALID_ARG, "seqlen_q must be positive");
    if (!mult8(p->q_row_stride) || !mult8(p->q_head_stride) || !mult8(p->k_row_stride) || !mult8(p->k_head_stride) ||
        !mult8(p->v_row_stride) || !mult8(p->v_head_stride) || !mult8(p->o_row_stride) || !mult8(p->o_head_stride) ||
        (!varlen && (!mult8(p->q_batch_stride) || !mult8(p->k_batch_stride) || !mult8(p->v_batch_stride) ||
                     !mult8(p->o_batch_stride))))
        return fail(MFA_ERR_UNSUPPORTED, "row/head/batch strides must be multiples of 8 elements (16 bytes)");
    if (reinterpret_cast<uintptr_t>(p->o_ptr) & 15) return fail(MFA_ERR_UNSUPPORTED, "output must be 16-byte aligned");
    if (reinterpret_cast<uintptr_t>(p->q_ptr) & 15) return fail(MFA_ERR_UNSUPPORTED, "q must be 16-byte aligned");

    const int D = p->head_dim;
    const int dpad = prefill_dpad_of(D);
    const bool bf16 = p->is_bf16 != 0;

    const bool window = p->local_window && (p->window_size_left >= 0 || p->window_size_right >= 0);
    const int kv_tile_box = mfa::kPrefillBlockN;

    CUtensorMap tmQ, tmK, tmV;
    const char* err;
    {
        const uint64_t rows = varlen ? (uint64_t)p->total_q : (uint64_t)p->seqlen_q;
        if (rows == 0) return fail(MFA_ERR_INVALID_ARG, "q has no rows");
        const uint64_t dims[4] = {(uint64_t)D, rows, (uint64_t)p->heads, varlen ? 1ull : (uint64_t)p->batch};
        const int64_t strides[4] = {1, p->q_row_stride, p->q_head_stride, varlen ? 0 : p->q_batch_stride};
        const uint32_t box[4] = {64, (uint32_t)mfa::kPrefillBlockM, 1, 1};
        if ((err = mfa::encode_tmap_4d(&tmQ, p->q_ptr, bf16, dims, strides, box))) return fail(MFA_ERR_UNSUPPORTED, "q: %s", err);
    }
    int kv_box_rows = kv_tile_box;
    if (paged) {
        if (p->page_block_size <= 0 || p->num_blocks <= 0 || p->max_blocks_per_seq <= 0)
            return fail(MFA_ERR_INVALID_ARG, "paged K/V needs page_block_size, num_blocks and max_blocks_per_seq");
        kv_box_rows = gcd_int(p->page_block_size, kv_tile_box);
        if (kv_box_rows < 8)
            return fail(MFA_ERR_UNSUPPORTED, "page_block_size must be a multiple of 8 (got %d)", p->page_block_size);
        if (!mult8(p->k_cache_block_stride) || !mult8(p->v_cache_block_stride))
            return fail(MFA_ERR_UNSUPPORTED, "cache block strides must be multiples of 8 elements");
        const uint64_t dims[4] = {(uint64_t)D, (uint64_t)p->page_block_size, (uint64_t)p->kv_heads, (uint64_t)p->num_blocks};
        const int64_t ks[4] = {1, p->k_row_stride, p->k_head_stride, p->k_cache_block_stride};
        const int64_t vs[4] = {1, p->v_row_stride, p->v_head_stride, p->v_cache_block_stride};
        const uint32_t box[4] = {64, (uint32_t)kv_box_rows, 1, 1};
        if ((err = mfa::encode_tmap_4d(&tmK, p->k_ptr, bf16, dims, ks, box))) return fail(MFA_ERR_UNSUPPORTED, "k: %s", err);
        if ((err = mfa::encode_tmap_4d(&tmV, p->v_ptr, bf16, dims, vs, box))) return fail(MFA_ERR_UNSUPPORTED, "v: %s", err);
    } else {
        const uint64_t rows = varlen ? (uint64_t)p->total_k : (uint64_t)p->seqlen_k;
        if (rows == 0) return fail(MFA_ERR_INVALID_ARG, "k has no rows");
        const uint64_t dims[4] = {(uint64_t)D, rows, (uint64_t)p->kv_heads, varlen ? 1ull : (uint64_t)p->batch};
        const int64_t ks[4] = {1, p->k_row_stride, p->k_head_stride, varlen ? 0 : p->k_batch_stride};
        const int64_t vs[4] = {1, p->v_row_stride, p->v_head_stride, varlen ? 0 : p->v_batch_stride};
        const uint32_t box[4] = {64, (uint32_t)kv_tile_box, 1, 1};
        if ((err = mfa::encode_tmap_4d(&tmK, p->k_ptr, bf16, dims, ks, box))) return fail(MFA_ERR_UNSUPPORTED, "k: %s", err);
        if ((err = mfa::encode_tmap_4d(&tmV, p->v_ptr, bf16, dims, vs, box))) return fail(MFA_ERR_UNSUPPORTED, "v: %s", err);
    }

    mfa::PrefillKParams kp{};
    kp.q = p->q_ptr;
    kp.q_batch_stride = p->q_batch_stride;
    kp.q_head_stride = p->q_head_stride;
    kp.q_row_stride = p->q_row_stride;
    kp.o = p->o_ptr;
    kp.o_batch_stride = p->o_batch_stride;
    kp.o_head_stride = p->o_head_stride;
    kp.o_row_stride = p->o_row_stride;
    kp.cu_seqlens_q = p->cu_seqlens_q;
    kp.cu_seqlens_k = p->cu_seqlens_k;
    kp.block_table = p->block_table;
    kp.block_table_batch_stride = p->block_table_batch_stride;
    kp.max_blocks_per_seq = p->max_blocks_per_seq;
    kp.page = p->page_block_size;
    kp.kv_box_rows = kv_box_rows;
    kp.heads = p->heads;
    kp.kv_heads = p->kv_heads;
    kp.group = p->heads / p->kv_heads;
    kp.batch = p->batch;
    kp.head_dim = D;
    kp.seqlen_q = p->seqlen_q;
    kp.seqlen_k = p->seqlen_k;
    kp.is_causal = p->is_causal ? 1 : 0;
    kp.causal_br = ((p->is_causal || window) && p->causal_bottom_right) ? 1 : 0;
    // band mask: causal is the upper bound 0; window_size_left / right are honoured only with local_window (the
    // reference accepts and ignores them, flash.h:33-34 / quirk Q10, and so do the plain entry points)
    kp.has_hi = (p->is_causal || (window && p->window_size_right >= 0)) ? 1 : 0;
    kp.win_right = p->is_causal ? 0 : (kp.has_hi ? p->window_size_right : 0);
    kp.has_lo = (window && p->window_size_left >= 0) ? 1 : 0;
    kp.win_left = kp.has_lo ? p->window_size_left : 0;
    const int rows_per_cta = mfa::prefill_rows_per_cta(dpad);
    kp.num_m_blocks = (p->seqlen_q + rows_per_cta - 1) / rows_per_cta;
    kp.scale = p->softmax_scale > 0.f ? p->softmax_scale : 1.0f / std::sqrt((float)D);   // api.cpp:99
    kp.scale_log2 = kp.scale * 1.4426950408889634f;                                      // api.cpp:100
    kp.lse = p->softmax_lse_ptr;   // (B, H, Sq) fixed / (H, total_q) packed, contiguous
    kp.lse_batch_stride = varlen ? 0 : (int64_t)p->heads * p->seqlen_q;
    kp.lse_head_stride = varlen ? (int64_t)p->total_q : (int64_t)p->seqlen_q;
    // Packed (varlen) batches whose sequences are (nearly) all of the maximum length run persistent like a fixed-length
    // batch, with odd rounds of the item walk reversed (snake order); ragged ones keep one CTA per item, because a static
    // walk cannot know which (m-block, sequence) items exist and the hardware scheduler balances them better
    // (profiles/varlen_persistent_ab_r02e.txt: 6 x 4096 1058 -> 1117 TFLOP/s, 48 x 512 406 -> 488; [128 ... 4096] 823 -> 744 if
    // it were persistent).  The fill ratio total_q / (batch x max_seqlen_q) is all the host knows without a device sync.
    // MFA_PREFILL_VARLEN_NONPERSISTENT=1: never persistent (A/B).
    static const bool varlen_np = [] { const char* e = getenv("MFA_PREFILL_VARLEN_NONPERSISTENT"); return e && e[0] == '1'; }();
    kp.snake = (varlen && !varlen_np && (long long)p->total_q * 10 >= 9ll * p->batch * p->seqlen_q) ? 1 : 0;
    // Ragged CAUSAL packed batches of up to 64 sequences: every CTA builds, from cu_seqlens, the table of the (m-block,
    // sequence) items that exist, ordered by descending m-block (= cost under the causal mask), and the persistent CTAs walk
    // THAT list in snake order (prefill_sm100.cu: p.compact) -- a static walk that does know which items exist.
    // MFA_PREFILL_NO_COMPACT=1: off (A/B).
    static const bool no_compact = [] { const char* e = getenv("MFA_PREFILL_NO_COMPACT"); return e && e[0] == '1'; }();
    // (a near-uniform batch is faster on the plain persistent walk above: 6 x 4096 1142 vs 1082 TFLOP/s)
    kp.compact = (varlen && !varlen_np && !no_compact && !kp.snake && kp.has_hi && !kp.has_lo && p->batch <= 64 && kp.num_m_blocks <= 255) ? 1 : 0;
    if (kp.compact) kp.snake = 1;
    kp.trace = g_trace;

    cudaError_t e = mfa::launch_prefill(tmQ, tmK, tmV, kp, bf16, dpad, static_cast<cudaStream_t>(cuda_stream));
    if (e != cudaSuccess) return cuda_fail(e, "prefill launch");
    g_launches.fetch_add(1);
    return MFA_OK;
}

// ------------------------------------------------------------------------------------------ decode
#ifndef MFA_DECODE_CTAS_PER_SM
// CTAs per SM the split heuristic plans for: two are resident (162 registers at head_dim 128, decode_sm100.cu).
// Planning for three was measured and is worse (scripts/decode_stages.py, B=24 GQA 32/8: KV=512 18.9 us vs 12.3 us,
// KV=8192 138 vs 127 us; B=1 long-context step at 16K rows 20.6 vs 16.5 us).
#define MFA_DECODE_CTAS_PER_SM 2
#endif
static int decode_splits_for(const mfa_params_t* p) {
    const int block_n = mfa::kDecodeTileRows;   // api.cpp:314-315
    const int num_n_blocks = std::max(1, (p->seqlen_k + block_n - 1) / block_n);
    int splits = p->num_splits;
    if (splits < 1) {
        // Heuristic in the place of api.cpp:269-302, fitted to this kernel (scripts/decode_splits_gqa.py, L2-cold sweep,
        // profiles/decode_splits_r02b.txt): one CTA serves a whole GQA group and two CTAs per SM are resident.  Splitting
        // pays only while ALL items still run as ONE wave (every extra item of a persistent CTA repeats the ~3 us ramp,
        // and the merge costs a few us), and only with >= 8 tiles (512 rows) per split:
        //   splits = min(slots / items_unsplit, tiles / 8).
        // B=24 GQA 32/8 (192 items) and B=24 MHA (576) stay unsplit at every KV length (KV=2048: 38 us unsplit vs 43-47
        // split); B=8 GQA KV=4096 takes 4 splits (29.9 us vs 52.5 unsplit and 38.6 with the 9 splits the wave-efficiency
        // rule of round 1 chose); a B=1 long-context step takes 37.
        const int group = p->heads / p->kv_heads;
        const int head_tiles = (group + mfa::kDecodeMaxGroup - 1) / mfa::kDecodeMaxGroup;
        const int base = std::max(1, p->batch * p->kv_heads * head_tiles);
        const int slots = MFA_DECODE_CTAS_PER_SM * sm_count();
        splits = std::max(1, std::min({slots / base, num_n_blocks / 8, 128}));
        // A partly filled single wave of LONG items (B=24 GQA at KV=8192: 192 items on 296 slots) gains ~3 % from being
        // cut into two full waves (3 splits of 43 tiles: 125.9 us vs 129.9); at 4096 rows it already loses (71.3 vs 68.5).
        if (splits == 1 && 2 * base > slots && base < slots) {
            const int two_waves = (2 * slots) / base;
            if (two_waves > 1 && num_n_blocks / two_waves >= 40) splits = two_waves;
        }
    }
    return std::max(1, std::min(splits, num_n_blocks));   // api.cpp:325-327
}

int mfa_decode_num_splits(const mfa_params_t* p) {
    if (check_common(p)) return -1;
    return decode_splits_for(p);
}

size_t mfa_decode_workspace_bytes(const mfa_params_t* p, int num_splits) {
    if (!p || num_splits <= 1) return 0;
    const size_t rows = (size_t)p->batch * (size_t)p->heads;
    return (size_t)num_splits * rows * ((size_t)p->head_dim + 1) * sizeof(float);   // api.cpp:333-334
}

static int decode_impl(const mfa_params_t* p, void* workspace, size_t workspace_bytes, mfa_xchg_t* x, void* cuda_stream);

int mfa_decode_fwd(const mfa_params_t* p, void* workspace, size_t workspace_bytes, void* cuda_stream) {
    return decode_impl(p, workspace, workspace_bytes, nullptr, cuda_stream);
}

size_t mfa_decode_kv_sharded_workspace_bytes(const mfa_params_t* p, int num_splits) {
    if (!p) return 0;
    const size_t rows = (size_t)p->batch * (size_t)p->heads;
    return (size_t)std::max(1, num_splits) * rows * ((size_t)p->head_dim + 1) * sizeof(float);
}

int mfa_decode_fwd_kv_sharded(const mfa_params_t* p, void* workspace, size_t workspace_bytes, mfa_xchg_t* x,
                              void* cuda_stream) {
    if (!x) return fail(MFA_ERR_INVALID_ARG, "null exchange");
    return decode_impl(p, workspace, workspace_bytes, x, cuda_stream);
}

static int decode_impl(const mfa_params_t* p, void* workspace, size_t workspace_bytes, mfa_xchg_t* x, void* cuda_stream) {
    int rc = check_common(p);
    if (rc) return rc;
    if ((rc = check_device())) return rc;
    if (p->seqlen_q != 1) return fail(MFA_ERR_INVALID_ARG, "flash decoding expects seqlen_q == 1, got %d", p->seqlen_q);
    if (p->seqlen_k <= 0) return fail(MFA_ERR_INVALID_ARG, "KV cache has no rows");
    const bool paged = p->block_table != nullptr;
    const bool fp8 = p->kv_fp8 != 0;
    const int64_t smul = fp8 ? 16 : 8;   // elements per 16 bytes
    auto mult8 = [smul](int64_t x) { return (x % smul) == 0; };   // (shadows the 16-bit helper: every cache stride check below)
    if (fp8 && (p->head_dim % 128))
        return fail(MFA_ERR_UNSUPPORTED, "an e4m3 KV cache needs head_dim to be a multiple of 128 (got %d)", p->head_dim);
    if (fp8 && p->kv_scale_stride != 0 && p->kv_scale_stride != 1)
        return fail(MFA_ERR_INVALID_ARG, "kv_scale_stride must be 0 (per tensor) or 1 (per KV head)");
    if (!mult8(p->k_row_stride) || !mult8(p->k_head_stride) || !mult8(p->v_row_stride) || !mult8(p->v_head_stride))
        return fail(MFA_ERR_UNSUPPORTED, "cache row/head strides must be multiples of 16 bytes");
    if ((p->q_head_stride % 2) || (p->q_batch_stride % 2) || (reinterpret_cast<uintptr_t>(p->q_ptr) & 3))
        return fail(MFA_ERR_UNSUPPORTED, "q strides must be even and q 4-byte aligned");
    if (p->partial_out) {
        if (!p->softmax_lse_ptr) return fail(MFA_ERR_INVALID_ARG, "partial_out requires softmax_lse_ptr");
        if (reinterpret_cast<uintptr_t>(p->o_ptr) & 15) return fail(MFA_ERR_UNSUPPORTED, "partial output must be 16-byte aligned");
    } else if ((p->o_head_stride % 4) || (p->o_batch_stride % 4) || (reinterpret_cast<uintptr_t>(p->o_ptr) & 7)) {
        return fail(MFA_ERR_UNSUPPORTED, "output strides must be multiples of 4 elements and 8-byte aligned");
    }

    const int D = p->head_dim;
    const int dpad = dpad_of(D);
    const bool bf16 = p->is_bf16 != 0;
    const int splits = decode_splits_for(p);
    if (x) {
        if (!x->connected) return fail(MFA_ERR_INVALID_ARG, "exchange is not connected (mfa_xchg_connect)");
        if (p->partial_out) return fail(MFA_ERR_INVALID_ARG, "partial_out and a fused exchange exclude each other");
        if (p->head_dim != x->head_dim || (int64_t)p->batch * p->heads > x->max_rows)
            return fail(MFA_ERR_INVALID_ARG, "exchange was created for head_dim %d, max_rows %d", x->head_dim, x->max_rows);
        int dev = -1;
        cudaGetDevice(&dev);
        if (dev != x->device) return fail(MFA_ERR_INVALID_ARG, "exchange lives on device %d, current device is %d", x->device, dev);
    }
    const size_t need = x ? mfa_decode_kv_sharded_workspace_bytes(p, splits) : mfa_decode_workspace_bytes(p, splits);
    if (need > 0 && (!workspace || workspace_bytes < need))
        return fail(MFA_ERR_INVALID_ARG, "decode workspace too small: need %zu bytes for %d splits, got %zu", need, splits,
                    workspace_bytes);
    if (need > 0 && (reinterpret_cast<uintptr_t>(workspace) & 15)) return fail(MFA_ERR_UNSUPPORTED, "workspace must be 16-byte aligned");

    CUtensorMap tmK, tmV;
    const char* err;
    int box_rows = mfa::kDecodeTileRows;
    int page = p->seqlen_k;
    if (paged) {
        if (p->page_block_size <= 0 || p->num_blocks <= 0 || p->max_blocks_per_seq <= 0)
            return fail(MFA_ERR_INVALID_ARG, "paged cache needs page_block_size, num_blocks and max_blocks_per_seq");
        page = p->page_block_size;
        box_rows = gcd_int(page, mfa::kDecodeTileRows);
        if (box_rows < 8) return fail(MFA_ERR_UNSUPPORTED, "page_block_size must be a multiple of 8 (got %d)", page);
        if (!mult8(p->k_cache_block_stride) || !mult8(p->v_cache_block_stride))
            return fail(MFA_ERR_UNSUPPORTED, "cache block strides must be multiples of 8 elements");
        const uint64_t dims[4] = {(uint64_t)D, (uint64_t)page, (uint64_t)p->kv_heads, (uint64_t)p->num_blocks};
        const int64_t ks[4] = {1, p->k_row_stride, p->k_head_stride, p->k_cache_block_stride};
        const int64_t vs[4] = {1, p->v_row_stride, p->v_head_stride, p->v_cache_block_stride};
        const uint32_t box[4] = {fp8 ? 128u : 64u, (uint32_t)box_rows, 1, 1};   // 128 bytes per box row either way
        if (fp8) {
            if ((err = mfa::encode_tmap_4d_u8(&tmK, p->k_ptr, dims, ks, box))) return fail(MFA_ERR_UNSUPPORTED, "k_cache: %s", err);
            if ((err = mfa::encode_tmap_4d_u8(&tmV, p->v_ptr, dims, vs, box))) return fail(MFA_ERR_UNSUPPORTED, "v_cache: %s", err);
        } else {
            if ((err = mfa::encode_tmap_4d(&tmK, p->k_ptr, bf16, dims, ks, box))) return fail(MFA_ERR_UNSUPPORTED, "k_cache: %s", err);
            if ((err = mfa::encode_tmap_4d(&tmV, p->v_ptr, bf16, dims, vs, box))) return fail(MFA_ERR_UNSUPPORTED, "v_cache: %s", err);
        }
    } else {
        if (!mult8(p->k_batch_stride) || !mult8(p->v_batch_stride))
            return fail(MFA_ERR_UNSUPPORTED, "cache batch strides must be multiples of 8 elements");
        const uint64_t dims[4] = {(uint64_t)D, (uint64_t)p->seqlen_k, (uint64_t)p->kv_heads, (uint64_t)p->batch};
        const int64_t ks[4] = {1, p->k_row_stride, p->k_head_stride, p->k_batch_stride};
        const int64_t vs[4] = {1, p->v_row_stride, p->v_head_stride, p->v_batch_stride};
        const uint32_t box[4] = {fp8 ? 128u : 64u, (uint32_t)box_rows, 1, 1};   // 128 bytes per box row either way
        if (fp8) {
            if ((err = mfa::encode_tmap_4d_u8(&tmK, p->k_ptr, dims, ks, box))) return fail(MFA_ERR_UNSUPPORTED, "k_cache: %s", err);
            if ((err = mfa::encode_tmap_4d_u8(&tmV, p->v_ptr, dims, vs, box))) return fail(MFA_ERR_UNSUPPORTED, "v_cache: %s", err);
        } else {
            if ((err = mfa::encode_tmap_4d(&tmK, p->k_ptr, bf16, dims, ks, box))) return fail(MFA_ERR_UNSUPPORTED, "k_cache: %s", err);
            if ((err = mfa::encode_tmap_4d(&tmV, p->v_ptr, bf16, dims, vs, box))) return fail(MFA_ERR_UNSUPPORTED, "v_cache: %s", err);
        }
    }

    const size_t rows = (size_t)p->batch * (size_t)p->heads;
    mfa::DecodeKParams kp{};
    kp.q = p->q_ptr;
    kp.o = p->o_ptr;
    kp.lse = p->softmax_lse_ptr;
    kp.o_part = splits > 1 ? static_cast<float*>(workspace) : nullptr;
    kp.lse_part = splits > 1 ? static_cast<float*>(workspace) + (size_t)splits * rows * D : nullptr;
    if (x && splits == 1) {   // one split: the kernel writes its normalised fp32 partial straight into the workspace
        kp.o = workspace;
        kp.lse = static_cast<float*>(workspace) + rows * D;
    }
    kp.q_batch_stride = p->q_batch_stride;
    kp.q_head_stride = p->q_head_stride;
    kp.o_batch_stride = p->o_batch_stride;
    kp.o_head_stride = p->o_head_stride;
    kp.seqlens_k = p->seqlens_k;
    kp.block_table = p->block_table;
    kp.block_table_batch_stride = p->block_table_batch_stride;
    kp.max_blocks_per_seq = p->max_blocks_per_seq;
    kp.page = page;
    kp.box_rows = box_rows;
    kp.heads = p->heads;
    kp.kv_heads = p->kv_heads;
    kp.group = p->heads / p->kv_heads;
    kp.batch = p->batch;
    kp.head_dim = D;
    kp.seqlen_k = p->seqlen_k;
    kp.num_splits = splits;
    kp.head_tiles = (kp.group + mfa::kDecodeMaxGroup - 1) / mfa::kDecodeMaxGroup;
    kp.partial_out = (p->partial_out || x) ? 1 : 0;
    kp.window_left = (p->local_window && p->window_size_left >= 0) ? p->window_size_left : -1;
    kp.scale = p->softmax_scale > 0.f ? p->softmax_scale : 1.0f / std::sqrt((float)D);
    kp.scale_log2 = kp.scale * 1.4426950408889634f;
    kp.kv_fp8 = fp8 ? 1 : 0;
    kp.kv_scale_stride = p->kv_scale_stride;
    kp.k_scale = fp8 ? p->k_scale_ptr : nullptr;
    kp.v_scale = fp8 ? p->v_scale_ptr : nullptr;

    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    cudaError_t e = mfa::launch_decode(tmK, tmV, kp, bf16, dpad, stream);
    if (e != cudaSuccess) return cuda_fail(e, "decode launch");
    g_launches.fetch_add(1);
    if (x) {
        // local split merge + push over NVLink + wait + cross-rank merge: one kernel (xchg_sm100.cu)
        mfa::XchgKParams xp{};
        xp.o_part = static_cast<const float*>(workspace);
        xp.lse_part = static_cast<const float*>(workspace) + (size_t)splits * rows * D;
        for (int r = 0; r < x->world; ++r) {
            xp.peer_data[r] = static_cast<float*>(x->peer[r]);
            xp.peer_flags[r] = reinterpret_cast<uint32_t*>(static_cast<char*>(x->peer[r]) + x->data_bytes);
        }
        xp.ctrs = reinterpret_cast<uint32_t*>(static_cast<char*>(x->local) + x->data_bytes + x->flag_bytes);
        xp.out = p->o_ptr;
        xp.lse_out = p->softmax_lse_ptr;
        xp.out_batch_stride = p->o_batch_stride;
        xp.out_head_stride = p->o_head_stride;
        xp.timeout_cycles = x->timeout_cycles;
        xp.num_splits = splits;
        xp.world = x->world;
        xp.rank = x->rank;
        xp.rows = (int)rows;
        xp.max_rows = x->max_rows;
        xp.head_dim = D;
        xp.heads = p->heads;
        e = mfa::launch_xchg_merge(xp, bf16, stream);
        if (e != cudaSuccess) return cuda_fail(e, "exchange-merge launch");
        g_launches.fetch_add(1);
        return MFA_OK;
    }
    if (splits > 1) {
        mfa::CombineKParams cp{};
        cp.out_f32 = p->partial_out ? 1 : 0;   // KV-sharded decode: the merged result stays a normalised fp32 partial
        cp.o_part = kp.o_part;
        cp.lse_part = kp.lse_part;
        cp.out = p->o_ptr;
        cp.lse_out = p->softmax_lse_ptr;
        cp.out_batch_stride = p->partial_out ? (int64_t)p->heads * D : p->o_batch_stride;
        cp.out_head_stride = p->partial_out ? (int64_t)D : p->o_head_stride;
        cp.heads = p->heads;
        cp.num_splits = splits;
        cp.rows = (int)rows;
        cp.head_dim = D;
        cp.packed = 0;
        e = mfa::launch_combine(cp, bf16, stream);
        if (e != cudaSuccess) return cuda_fail(e, "combine launch");
        g_launches.fetch_add(1);
    }
    return MFA_OK;
}

// ------------------------------------------------------------------------------------------ decode without allocation
// The split-KV scratch (fp32 partials + LSEs) of mfa_decode_fwd_static lives in the library image instead of a caller
// buffer: MFA_DECODE_STATIC_WS_BYTES of device memory on every GPU that loads the library, held until exit.  The name
// below is a device symbol; in HOST code it is only a stand-in, so the launch path gets the array's device address from
// cudaGetSymbolAddress -- once per device, with that device current -- and passes that pointer on.
static __device__ __align__(256) unsigned char g_decode_static_ws[MFA_DECODE_STATIC_WS_BYTES];

static int decode_static_ws_ptr(void** out) {
    static std::atomic<void*> ptr_of[64];   // per device; zero-initialised
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(MFA_ERR_UNSUPPORTED, "device ordinal %d beyond the 64 this entry point tracks", dev);
    void* ptr = ptr_of[dev].load(std::memory_order_acquire);
    if (!ptr) {
        e = cudaGetSymbolAddress(&ptr, g_decode_static_ws);   // the address on the CURRENT device
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(decode scratch)");
        if (!ptr) return fail(MFA_ERR_CUDA, "cudaGetSymbolAddress returned a null decode scratch address");
        ptr_of[dev].store(ptr, std::memory_order_release);
    }
    *out = ptr;
    return MFA_OK;
}

size_t mfa_decode_static_workspace_limit(void) { return (size_t)MFA_DECODE_STATIC_WS_BYTES; }

int mfa_decode_fwd_static(const mfa_params_t* p, void* cuda_stream) {
    int rc = check_common(p);
    if (rc) return rc;
    const int splits = decode_splits_for(p);
    const size_t need = mfa_decode_workspace_bytes(p, splits);
    if (need > (size_t)MFA_DECODE_STATIC_WS_BYTES)
        return fail(MFA_ERR_UNSUPPORTED,
                    "mfa_decode_fwd_static: %d splits x %d rows x head_dim %d need %zu bytes of scratch, the built-in array holds %zu; "
                    "lower num_splits or call mfa_decode_fwd with a caller workspace",
                    splits, p->batch * p->heads, p->head_dim, need, (size_t)MFA_DECODE_STATIC_WS_BYTES);
    void* ws = nullptr;
    if (need > 0 && (rc = decode_static_ws_ptr(&ws))) return rc;
    // The partials are fully written by the split-KV kernel before the combine reads them: nothing has to start at zero.
    return decode_impl(p, ws, need ? (size_t)MFA_DECODE_STATIC_WS_BYTES : 0, nullptr, cuda_stream);
}

// ------------------------------------------------------------------------------------------ KV-shard exchange
int mfa_xchg_create(int world, int rank, int max_rows, int head_dim, mfa_xchg_t** out) {
    if (!out) return fail(MFA_ERR_INVALID_ARG, "null out");
    if (world < 1 || world > MFA_XCHG_MAX_WORLD || rank < 0 || rank >= world)
        return fail(MFA_ERR_INVALID_ARG, "world must be 1..%d and 0 <= rank < world (got %d, %d)", MFA_XCHG_MAX_WORLD, world, rank);
    if (max_rows <= 0 || head_dim <= 0 || head_dim > 256 || head_dim % 8)
        return fail(MFA_ERR_INVALID_ARG, "max_rows must be positive and head_dim a multiple of 8, <= 256");
    int rc = check_device();
    if (rc) return rc;
    mfa_xchg_t* x = new mfa_xchg_t();
    x->world = world;
    x->rank = rank;
    x->max_rows = max_rows;
    x->head_dim = head_dim;
    cudaGetDevice(&x->device);
    x->data_bytes = (size_t)2 * world * max_rows * (head_dim + 4) * sizeof(float);
    x->flag_bytes = (size_t)world * max_rows * sizeof(uint32_t);
    x->bytes = x->data_bytes + x->flag_bytes + 16;
    for (int r = 0; r < MFA_XCHG_MAX_WORLD; ++r) {
        x->peer[r] = nullptr;
        x->ipc_opened[r] = false;
    }
    x->connected = false;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, x->device);
    // ~2 s of SM clocks by default (MFA_XCHG_TIMEOUT_MS overrides: a first call that loads modules on a peer, or a
    // peer held in a debugger, may need more); on expiry the row is written as NaN and the status word is set
    long long timeout_ms = 2000;
    if (const char* e = getenv("MFA_XCHG_TIMEOUT_MS")) {
        const long long v = atoll(e);
        if (v > 0) timeout_ms = v;
    }
    x->timeout_cycles = (long long)(khz > 0 ? khz : 2000000) * timeout_ms;
    cudaError_t e = cudaMalloc(&x->local, x->bytes);
    if (e != cudaSuccess) {
        delete x;
        return cuda_fail(e, "cudaMalloc(exchange buffer)");
    }
    e = cudaMemset(x->local, 0, x->bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        cudaFree(x->local);
        delete x;
        return cuda_fail(e, "cudaMemset(exchange buffer)");
    }
    x->peer[rank] = x->local;
    if (world == 1) x->connected = true;
    *out = x;
    return MFA_OK;
}

int mfa_xchg_export(mfa_xchg_t* x, void* handle_out) {
    if (!x || !handle_out) return fail(MFA_ERR_INVALID_ARG, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == MFA_XCHG_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, x->local);
    if (e != cudaSuccess) return cuda_fail(e, "cudaIpcGetMemHandle");
    memcpy(handle_out, &h, sizeof(h));
    return MFA_OK;
}

int mfa_xchg_connect(mfa_xchg_t* x, const void* handles) {
    if (!x || !handles) return fail(MFA_ERR_INVALID_ARG, "null argument");
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank || x->peer[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char*>(handles) + (size_t)r * MFA_XCHG_HANDLE_BYTES, sizeof(h));
        void* ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return cuda_fail(e, "cudaIpcOpenMemHandle (peer access between the GPUs is required)");
        x->peer[r] = ptr;
        x->ipc_opened[r] = true;
    }
    x->connected = true;
    return MFA_OK;
}

int mfa_xchg_connect_ptrs(mfa_xchg_t* x, void* const* peer_ptrs) {
    if (!x || !peer_ptrs) return fail(MFA_ERR_INVALID_ARG, "null argument");
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank) continue;
        if (!peer_ptrs[r]) return fail(MFA_ERR_INVALID_ARG, "null peer pointer for rank %d", r);
        x->peer[r] = peer_ptrs[r];
    }
    x->connected = true;
    return MFA_OK;
}

void* mfa_xchg_local_ptr(mfa_xchg_t* x) { return x ? x->local : nullptr; }

int mfa_xchg_status(mfa_xchg_t* x) {
    if (!x) return -1;
    uint32_t st = 0;
    cudaError_t e = cudaMemcpy(&st, static_cast<char*>(x->local) + x->data_bytes + x->flag_bytes, sizeof(st), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
        cuda_fail(e, "exchange status read");
        return -1;
    }
    return (int)st;
}

int mfa_xchg_clear_status(mfa_xchg_t* x) {
    const int st = mfa_xchg_status(x);
    if (st > 0) {
        const uint32_t zero = 0;
        cudaError_t e = cudaMemcpy(static_cast<char*>(x->local) + x->data_bytes + x->flag_bytes, &zero, sizeof(zero), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            cuda_fail(e, "exchange status clear");
            return -1;
        }
    }
    return st;
}

int mfa_xchg_destroy(mfa_xchg_t* x) {
    if (!x) return MFA_OK;
    for (int r = 0; r < x->world; ++r)
        if (x->ipc_opened[r]) cudaIpcCloseMemHandle(x->peer[r]);
    cudaFree(x->local);
    delete x;
    return MFA_OK;
}

// ------------------------------------------------------------------------------------------ KV-cache append
int mfa_kvcache_append(const mfa_append_params_t* p, void* cuda_stream) {
    if (!p) return fail(MFA_ERR_INVALID_ARG, "null params");
    if (!p->k_new || !p->v_new || !p->k_cache || !p->v_cache) return fail(MFA_ERR_INVALID_ARG, "null tensor pointer");
    if (!p->seqlens_k) return fail(MFA_ERR_INVALID_ARG, "kv-cache append needs seqlens_k (where to write)");
    if (p->batch < 0 || p->s_new < 0 || p->kv_heads <= 0) return fail(MFA_ERR_INVALID_ARG, "bad batch / s_new / kv_heads");
    if (p->head_dim <= 0 || p->head_dim > 256 || p->head_dim % 8) return fail(MFA_ERR_UNSUPPORTED, "head_dim must be a multiple of 8, <= 256");
    if (p->block_table && (p->page_block_size <= 0 || p->max_blocks_per_seq <= 0))
        return fail(MFA_ERR_INVALID_ARG, "paged append needs page_block_size and max_blocks_per_seq");
    auto mult8 = [](int64_t v) { return (v % 8) == 0; };
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if (!al16(p->k_new) || !al16(p->v_new) || !al16(p->k_cache) || !al16(p->v_cache) || !mult8(p->knew_batch_stride) ||
        !mult8(p->knew_row_stride) || !mult8(p->knew_head_stride) || !mult8(p->vnew_batch_stride) ||
        !mult8(p->vnew_row_stride) || !mult8(p->vnew_head_stride) || !mult8(p->k_batch_stride) ||
        !mult8(p->k_row_stride) || !mult8(p->k_head_stride) || !mult8(p->v_batch_stride) || !mult8(p->v_row_stride) ||
        !mult8(p->v_head_stride))
        return fail(MFA_ERR_INVALID_ARG, "append: pointers must be 16-byte aligned and strides multiples of 8 elements");
    int rc = check_device();
    if (rc) return rc;
    if (p->batch == 0 || p->s_new == 0) return MFA_OK;
    cudaError_t e = mfa::launch_kvcache_append(*p, sm_count(), static_cast<cudaStream_t>(cuda_stream));
    if (e != cudaSuccess) return cuda_fail(e, "kvcache append launch");
    g_launches.fetch_add(1);
    return MFA_OK;
}

int mfa_combine_partials(const float* packed, int num_partials, int rows, int head_dim, void* out,
                         int64_t out_row_stride, int is_bf16, float* lse_out, void* cuda_stream) {
    if (!packed || !out || num_partials <= 0 || rows <= 0 || head_dim <= 0 || head_dim % 4)
        return fail(MFA_ERR_INVALID_ARG, "combine: bad arguments");
    if ((out_row_stride % 4) || (reinterpret_cast<uintptr_t>(out) & 7))
        return fail(MFA_ERR_UNSUPPORTED, "combine: output rows must be 8-byte aligned");
    int rc = check_device();
    if (rc) return rc;
    mfa::CombineKParams cp{};
    cp.o_part = packed;
    cp.lse_part = nullptr;
    cp.out = out;
    cp.lse_out = lse_out;
    cp.out_row_stride = out_row_stride;
    cp.heads = 0;
    cp.num_splits = num_partials;
    cp.rows = rows;
    cp.head_dim = head_dim;
    cp.packed = 1;
    cudaError_t e = mfa::launch_combine(cp, is_bf16 != 0, static_cast<cudaStream_t>(cuda_stream));
    if (e != cudaSuccess) return cuda_fail(e, "combine launch");
    g_launches.fetch_add(1);
    return MFA_OK;
}

}  // extern "C"
