// torch_shim.cpp -- compiled host boundary: the three extension functions of the reference's pybind module
// `mini_flash_attention._C` (/root/reference/csrc/api.cpp:4-9, typed in _C.pyi:4-41), re-implemented as a THIN torch
// extension over the C-ABI of libmfa_b200.so (include/mfa_b200.h).  It replaces /root/reference/csrc/api.cpp +
// csrc/mfa/api.cpp: validation (api.cpp:125-162, 205-261, 357-403), output / workspace allocation through torch's
// caching allocator (:170, 229, 405, 333-334), current stream and device guard (:123, 182).  No arithmetic here.
// `_C.py` holds the same logic in Python (used when this module has not been built); both call the same C-ABI.
#include <torch/extension.h>

#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include <limits>
#include <optional>
#include <tuple>

#include "../../include/mfa_b200.h"

namespace {

void check_rc(int rc, const char* what) {
    TORCH_CHECK(rc == MFA_OK, what, " failed (", rc, "): ", mfa_last_error());
}

void check_qkv(const at::Tensor& q, const at::Tensor& k, const at::Tensor& v) {
    // api.cpp:125-137
    TORCH_CHECK(q.scalar_type() == at::kHalf || q.scalar_type() == at::kBFloat16,
                "FlashAttention only support fp16 and bf16 data type");
    TORCH_CHECK(k.scalar_type() == q.scalar_type(), "query and key must have the same dtype");
    TORCH_CHECK(v.scalar_type() == q.scalar_type(), "query and value must have the same dtype");
    TORCH_CHECK(q.is_cuda(), "q must be on CUDA");
    TORCH_CHECK(k.is_cuda(), "k must be on CUDA");
    TORCH_CHECK(v.is_cuda(), "v must be on CUDA");
    TORCH_CHECK(q.dim() >= 1 && q.stride(-1) == 1, "Input tensor must have contiguous last dimension");
    TORCH_CHECK(k.dim() >= 1 && k.stride(-1) == 1, "Input tensor must have contiguous last dimension");
    TORCH_CHECK(v.dim() >= 1 && v.stride(-1) == 1, "Input tensor must have contiguous last dimension");
}

// TMA needs a 16-byte aligned base and 16-byte multiples for every non-innermost stride
bool tma_ok(const at::Tensor& t) {
    if (reinterpret_cast<uintptr_t>(t.data_ptr()) % 16) return false;
    // (a broadcast view -- expand(): stride 0 on an extent > 1 -- cannot be expressed by a tensor map either)
    const int64_t per16 = 16 / (int64_t)t.element_size();   // elements per 16 bytes (8, or 16 for an e4m3 cache)
    for (int64_t d = 0; d + 1 < t.dim(); ++d)
        if (t.stride(d) % per16 || (t.stride(d) <= 0 && t.size(d) > 1)) return false;
    return t.dim() == 0 || t.stride(t.dim() - 1) == 1;
}
at::Tensor tma_ready(const at::Tensor& t) { return tma_ok(t) ? t : t.contiguous(); }

void check_head_dim(int64_t head_dim, int64_t heads, int64_t kv_heads) {
    TORCH_CHECK(head_dim <= 256, "head dimension must be less than or equal to 256");   // api.cpp:152
    TORCH_CHECK(kv_heads > 0 && heads % kv_heads == 0,
                "number of key/value heads must be divisible by number of query heads");   // api.cpp:153
    // the reference rounds head_dim up to a multiple of 32 and reads the neighbouring head
    TORCH_CHECK(head_dim % 8 == 0, "head dimension must be a multiple of 8");
}

void* stream_of(const at::Tensor& t) { return at::cuda::getCurrentCUDAStream(t.get_device()).stream(); }

#define SHAPE_IS(t, ...) ((t).sizes() == at::IntArrayRef({__VA_ARGS__}))

}  // namespace

// q (B,Sq,H,D); k, v (B,Sk,Hkv,D); returns (B,Sq,H,D) and, when asked, the natural-log LSE (B,H,Sq) fp32 that the
// reference computes and drops (api.cpp:436-437).  api.cpp:113-186
static std::tuple<at::Tensor, at::Tensor> forward_impl(const at::Tensor& q, const at::Tensor& k, const at::Tensor& v,
                                                       const std::optional<at::Tensor>& out_, bool is_causal,
                                                       int64_t window_left, int64_t window_right, bool want_lse,
                                                       bool bottom_right = false, bool local_window = false) {
    check_qkv(q, k, v);
    TORCH_CHECK(q.dim() == 4 && k.dim() == 4 && v.dim() == 4, "q, k, v must be 4-D (batch, seqlen, heads, head_dim)");
    const int64_t batch = q.size(0), seqlen_q = q.size(1), heads = q.size(2), head_dim = q.size(3);
    const int64_t seqlen_k = k.size(1), kv_heads = k.size(2);
    TORCH_CHECK(k.size(0) == batch, "batch size of q and k must be the same");
    TORCH_CHECK(v.size(0) == batch, "batch size of q and v must be the same");
    TORCH_CHECK(v.size(1) == seqlen_k, "sequence length of k must be the same as v");
    check_head_dim(head_dim, heads, kv_heads);
    TORCH_CHECK(SHAPE_IS(k, batch, seqlen_k, kv_heads, head_dim), "k must have shape (batch, seqlen_k, kv_num_heads, head_dim)");
    TORCH_CHECK(SHAPE_IS(v, batch, seqlen_k, kv_heads, head_dim), "v must have shape (batch, seqlen_k, kv_num_heads, head_dim)");
    c10::cuda::CUDAGuard guard(q.device());   // api.cpp:123
    at::Tensor out;
    if (out_.has_value()) {
        out = *out_;
        TORCH_CHECK(out.stride(-1) == 1, "Output tensor must have contiguous last dimension");
        TORCH_CHECK(SHAPE_IS(out, batch, seqlen_q, heads, head_dim), "out must have shape (batch, seqlen_q, num_heads, head_dim)");
        TORCH_CHECK(out.scalar_type() == q.scalar_type() && out.is_cuda(), "out must match q's dtype and device");
    } else {
        out = at::empty_like(q);
    }
    at::Tensor lse;
    if (want_lse) lse = at::empty({batch, heads, seqlen_q}, q.options().dtype(at::kFloat));
    if (batch == 0 || seqlen_q == 0 || heads == 0) return {out, lse};
    if (seqlen_k == 0) {
        if (want_lse) lse.fill_(-std::numeric_limits<float>::infinity());
        return {out.zero_(), lse};
    }
    const at::Tensor qc = tma_ready(q), kc = tma_ready(k), vc = tma_ready(v);
    at::Tensor oc = tma_ok(out) ? out : at::empty(out.sizes(), out.options());

    mfa_params_t P{};
    P.q_ptr = qc.data_ptr(); P.k_ptr = kc.data_ptr(); P.v_ptr = vc.data_ptr(); P.o_ptr = oc.data_ptr();
    P.q_batch_stride = qc.stride(0); P.q_row_stride = qc.stride(1); P.q_head_stride = qc.stride(2);
    P.k_batch_stride = kc.stride(0); P.k_row_stride = kc.stride(1); P.k_head_stride = kc.stride(2);
    P.v_batch_stride = vc.stride(0); P.v_row_stride = vc.stride(1); P.v_head_stride = vc.stride(2);
    P.o_batch_stride = oc.stride(0); P.o_row_stride = oc.stride(1); P.o_head_stride = oc.stride(2);
    P.is_causal = is_causal ? 1 : 0;   // api.cpp:157, :88
    P.causal_bottom_right = bottom_right ? 1 : 0;   // new; 0 = the reference's top-left mask
    P.local_window = local_window ? 1 : 0;          // new; 0 = window sizes ignored like the reference
    P.window_size_left = (int32_t)window_left; P.window_size_right = (int32_t)window_right;
    P.heads = (int32_t)heads; P.kv_heads = (int32_t)kv_heads; P.batch = (int32_t)batch;
    P.seqlen_q = (int32_t)seqlen_q; P.seqlen_k = (int32_t)seqlen_k; P.head_dim = (int32_t)head_dim;
    P.is_bf16 = q.scalar_type() == at::kBFloat16;
    if (want_lse) P.softmax_lse_ptr = lse.data_ptr<float>();
    check_rc(mfa_prefill_fwd(&P, stream_of(q)), "mfa_prefill_fwd");
    if (!oc.is_same(out)) out.copy_(oc);
    return {out, lse};
}

at::Tensor mini_flash_attention_forward(const at::Tensor& q, const at::Tensor& k, const at::Tensor& v,
                                        std::optional<at::Tensor> out_, bool is_causal, int64_t window_left,
                                        int64_t window_right) {
    return std::get<0>(forward_impl(q, k, v, out_, is_causal, window_left, window_right, false));
}
std::tuple<at::Tensor, at::Tensor> mini_flash_attention_forward_lse(const at::Tensor& q, const at::Tensor& k,
                                                                    const at::Tensor& v, std::optional<at::Tensor> out_,
                                                                    bool is_causal, int64_t window_left,
                                                                    int64_t window_right) {
    return forward_impl(q, k, v, out_, is_causal, window_left, window_right, true);
}
// superset entry point: optional LSE, mask aligned bottom-right, window_left / window_right HONOURED (sliding window)
std::tuple<at::Tensor, std::optional<at::Tensor>> mini_flash_attention_forward_ext(
    const at::Tensor& q, const at::Tensor& k, const at::Tensor& v, std::optional<at::Tensor> out_, bool is_causal,
    int64_t window_left, int64_t window_right, bool return_lse, bool bottom_right) {
    auto r = forward_impl(q, k, v, out_, is_causal, window_left, window_right, return_lse, bottom_right, true);
    return {std::get<0>(r), return_lse ? std::optional<at::Tensor>(std::get<1>(r)) : std::nullopt};
}

// Packed variable-length prefill, optionally with paged K/V.  api.cpp:189-267.
// q (total_q,H,D); k, v (total_k,Hkv,D) or, with block_table (batch,max_blocks) int32, (num_blocks,page,Hkv,D).
// With want_lse also the LSE, (H, total_q) fp32 (flash-attn's packed layout).
static std::tuple<at::Tensor, at::Tensor> varlen_impl(const at::Tensor& q, const at::Tensor& k, const at::Tensor& v,
                                                      const at::Tensor& cu_seqlens_q, const at::Tensor& cu_seqlens_k,
                                                      int64_t max_seqlen_q, int64_t max_seqlen_k, bool causal,
                                                      int64_t window_left, int64_t window_right,
                                                      const std::optional<at::Tensor>& block_table, bool want_lse,
                                                      bool bottom_right = false, bool local_window = false) {
    check_qkv(q, k, v);
    TORCH_CHECK(cu_seqlens_q.is_cuda(), "cu_seqlens_q must be on CUDA");
    TORCH_CHECK(cu_seqlens_k.is_cuda(), "cu_seqlens_k must be on CUDA");
    TORCH_CHECK(cu_seqlens_q.scalar_type() == at::kInt && cu_seqlens_k.scalar_type() == at::kInt, "cu_seqlens must be int32");
    TORCH_CHECK(q.dim() == 3, "q must have shape (total_q, num_heads, head_dim)");
    const int64_t total_q = q.size(0), heads = q.size(1), head_dim = q.size(2);
    const int64_t batch = cu_seqlens_q.numel() - 1;
    TORCH_CHECK(batch >= 0 && cu_seqlens_k.numel() == batch + 1, "cu_seqlens_q and cu_seqlens_k must have batch+1 entries");
    TORCH_CHECK(k.dim() >= 2, "k must have at least 2 dimensions");
    const int64_t kv_heads = k.size(-2);
    check_head_dim(head_dim, heads, kv_heads);
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor out = at::empty_like(q);
    at::Tensor lse;
    if (want_lse) lse = at::empty({heads, total_q}, q.options().dtype(at::kFloat));
    if (batch == 0 || total_q == 0 || heads == 0) return {out, lse};
    const at::Tensor cu_q = cu_seqlens_q.contiguous(), cu_k = cu_seqlens_k.contiguous();

    mfa_params_t P{};
    at::Tensor kc, vc, bt;
    if (block_table.has_value()) {
        TORCH_CHECK(k.dim() == 4 && v.dim() == 4, "paged k/v must have shape (num_blocks, page_block_size, kv_heads, head_dim)");
        const int64_t num_blocks = k.size(0), page = k.size(1);
        TORCH_CHECK(SHAPE_IS(k, num_blocks, page, kv_heads, head_dim), "k must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)");
        TORCH_CHECK(SHAPE_IS(v, num_blocks, page, kv_heads, head_dim), "v must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)");
        TORCH_CHECK(block_table->is_cuda(), "block_table must be on CUDA");
        TORCH_CHECK(block_table->scalar_type() == at::kInt, "block_table must be int32");
        TORCH_CHECK(block_table->dim() == 2 && block_table->size(0) == batch, "block_table must have the same batch size as q");
        bt = block_table->stride(1) == 1 ? *block_table : block_table->contiguous();
        kc = tma_ready(k); vc = tma_ready(v);
        P.block_table = bt.data_ptr<int32_t>();
        P.block_table_batch_stride = (int32_t)bt.stride(0);
        P.max_blocks_per_seq = (int32_t)bt.size(1);
        P.page_block_size = (int32_t)page; P.num_blocks = (int32_t)num_blocks;
        P.k_cache_block_stride = kc.stride(0); P.v_cache_block_stride = vc.stride(0);
        P.k_row_stride = kc.stride(1); P.k_head_stride = kc.stride(2);
        P.v_row_stride = vc.stride(1); P.v_head_stride = vc.stride(2);
    } else {
        TORCH_CHECK(k.dim() == 3 && v.dim() == 3, "k, v must have shape (total_k, kv_num_heads, head_dim)");
        const int64_t total_k = k.size(0);
        TORCH_CHECK(SHAPE_IS(k, total_k, kv_heads, head_dim), "k must have shape (total_k, kv_num_heads, head_dim)");
        TORCH_CHECK(SHAPE_IS(v, total_k, kv_heads, head_dim), "v must have shape (total_k, kv_num_heads, head_dim)");
        if (total_k == 0) {
            if (want_lse) lse.fill_(-std::numeric_limits<float>::infinity());
            return {out.zero_(), lse};
        }
        kc = tma_ready(k); vc = tma_ready(v);
        P.total_k = (int32_t)total_k;
        P.k_row_stride = kc.stride(0); P.k_head_stride = kc.stride(1);
        P.v_row_stride = vc.stride(0); P.v_head_stride = vc.stride(1);
    }
    const at::Tensor qc = tma_ready(q);
    P.q_ptr = qc.data_ptr(); P.k_ptr = kc.data_ptr(); P.v_ptr = vc.data_ptr(); P.o_ptr = out.data_ptr();
    P.q_row_stride = qc.stride(0); P.q_head_stride = qc.stride(1);
    P.o_row_stride = out.stride(0); P.o_head_stride = out.stride(1);
    P.is_causal = causal ? 1 : 0;   // api.cpp:231
    P.causal_bottom_right = bottom_right ? 1 : 0;
    P.local_window = local_window ? 1 : 0;
    P.window_size_left = (int32_t)window_left; P.window_size_right = (int32_t)window_right;
    P.heads = (int32_t)heads; P.kv_heads = (int32_t)kv_heads; P.batch = (int32_t)batch;
    P.seqlen_q = (int32_t)max_seqlen_q; P.seqlen_k = (int32_t)max_seqlen_k; P.head_dim = (int32_t)head_dim;
    P.is_bf16 = q.scalar_type() == at::kBFloat16;
    P.cu_seqlens_q = cu_q.data_ptr<int32_t>(); P.cu_seqlens_k = cu_k.data_ptr<int32_t>();
    P.total_q = (int32_t)total_q;
    if (want_lse) {
        // rows of q past cu_seqlens_q[batch] belong to no sequence and are not written by the kernel
        lse.fill_(-std::numeric_limits<float>::infinity());
        P.softmax_lse_ptr = lse.data_ptr<float>();
    }
    check_rc(mfa_prefill_fwd(&P, stream_of(q)), "mfa_prefill_fwd");
    return {out, lse};
}

at::Tensor mini_flash_attention_varlen_forward(const at::Tensor& q, const at::Tensor& k, const at::Tensor& v,
                                               const at::Tensor& cu_seqlens_q, const at::Tensor& cu_seqlens_k,
                                               int64_t max_seqlen_q, int64_t max_seqlen_k, bool causal,
                                               int64_t window_left, int64_t window_right,
                                               std::optional<at::Tensor> block_table) {
    return std::get<0>(varlen_impl(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal, window_left,
                                   window_right, block_table, false));
}
std::tuple<at::Tensor, at::Tensor> mini_flash_attention_varlen_forward_lse(
    const at::Tensor& q, const at::Tensor& k, const at::Tensor& v, const at::Tensor& cu_seqlens_q,
    const at::Tensor& cu_seqlens_k, int64_t max_seqlen_q, int64_t max_seqlen_k, bool causal, int64_t window_left,
    int64_t window_right, std::optional<at::Tensor> block_table) {
    return varlen_impl(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal, window_left, window_right,
                       block_table, true);
}
std::tuple<at::Tensor, std::optional<at::Tensor>> mini_flash_attention_varlen_forward_ext(
    const at::Tensor& q, const at::Tensor& k, const at::Tensor& v, const at::Tensor& cu_seqlens_q,
    const at::Tensor& cu_seqlens_k, int64_t max_seqlen_q, int64_t max_seqlen_k, bool causal, int64_t window_left,
    int64_t window_right, std::optional<at::Tensor> block_table, bool return_lse, bool bottom_right) {
    auto r = varlen_impl(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal, window_left,
                         window_right, block_table, return_lse, bottom_right, true);
    return {std::get<0>(r), return_lse ? std::optional<at::Tensor>(std::get<1>(r)) : std::nullopt};
}

namespace {

// Validation + parameter block shared by the decode entry points.  api.cpp:343-433.  The tensors the parameter block
// points into are returned through `keep`.
struct DecodeCall {
    mfa_params_t P{};
    at::Tensor q, k, v, bt, sl;
};

void decode_params(DecodeCall& c, const at::Tensor& q, const at::Tensor& k_cache, const at::Tensor& v_cache,
                   const std::optional<at::Tensor>& cache_seqlens, const std::optional<at::Tensor>& block_table,
                   int64_t num_splits, bool fp8 = false) {
    if (fp8) {
        // e4m3 cache (the reference gates the dtype at api.cpp:126-129): q keeps its 16-bit type
        TORCH_CHECK(q.scalar_type() == at::kHalf || q.scalar_type() == at::kBFloat16,
                    "FlashAttention only support fp16 and bf16 data type");
        TORCH_CHECK(k_cache.scalar_type() == at::kFloat8_e4m3fn && v_cache.scalar_type() == at::kFloat8_e4m3fn,
                    "k_cache and v_cache must both be torch.float8_e4m3fn");
        TORCH_CHECK(q.is_cuda() && k_cache.is_cuda() && v_cache.is_cuda(), "q, k_cache and v_cache must be on CUDA");
        TORCH_CHECK(q.dim() >= 1 && q.stride(-1) == 1 && k_cache.dim() >= 1 && k_cache.stride(-1) == 1 &&
                        v_cache.dim() >= 1 && v_cache.stride(-1) == 1,
                    "Input tensor must have contiguous last dimension");
        TORCH_CHECK(q.size(-1) % 128 == 0, "an e4m3 KV cache needs head_dim to be a multiple of 128");
    } else {
        check_qkv(q, k_cache, v_cache);
    }
    TORCH_CHECK(q.dim() == 4, "q must have shape (batch, 1, num_heads, head_dim)");
    const int64_t batch = q.size(0), seqlen_q = q.size(1), heads = q.size(2), head_dim = q.size(3);
    TORCH_CHECK(seqlen_q == 1, "flash decoding expects seqlen_q == 1, got ", seqlen_q);   // api.cpp:377
    TORCH_CHECK(k_cache.dim() == 4 && v_cache.dim() == 4, "k_cache / v_cache must be 4-D");
    const int64_t kv_heads = k_cache.size(-2);
    check_head_dim(head_dim, heads, kv_heads);
    mfa_params_t& P = c.P;
    c.k = tma_ready(k_cache);
    c.v = tma_ready(v_cache);
    int64_t seqlen_k;
    if (block_table.has_value()) {
        TORCH_CHECK(block_table->is_cuda(), "block_table must be on CUDA");
        TORCH_CHECK(block_table->scalar_type() == at::kInt, "block_table must be int32");
        TORCH_CHECK(block_table->dim() == 2 && block_table->size(0) == batch, "block_table must have the same batch size as q");
        const int64_t num_blocks = k_cache.size(0), page = k_cache.size(1);
        TORCH_CHECK(SHAPE_IS(k_cache, num_blocks, page, kv_heads, head_dim), "k_cache must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)");
        TORCH_CHECK(SHAPE_IS(v_cache, num_blocks, page, kv_heads, head_dim), "v_cache must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)");
        c.bt = block_table->stride(1) == 1 ? *block_table : block_table->contiguous();
        seqlen_k = c.bt.size(1) * page;   // api.cpp:392-394
        P.block_table = c.bt.data_ptr<int32_t>();
        P.block_table_batch_stride = (int32_t)c.bt.stride(0);
        P.max_blocks_per_seq = (int32_t)c.bt.size(1);
        P.page_block_size = (int32_t)page; P.num_blocks = (int32_t)num_blocks;
        P.k_cache_block_stride = c.k.stride(0); P.v_cache_block_stride = c.v.stride(0);
    } else {
        seqlen_k = k_cache.size(1);
        TORCH_CHECK(SHAPE_IS(k_cache, batch, seqlen_k, kv_heads, head_dim), "k_cache must have shape (batch, seqlen_k, kv_num_heads, head_dim)");
        TORCH_CHECK(SHAPE_IS(v_cache, batch, seqlen_k, kv_heads, head_dim), "v_cache must have shape (batch, seqlen_k, kv_num_heads, head_dim)");
        P.k_batch_stride = c.k.stride(0); P.v_batch_stride = c.v.stride(0);
    }
    P.k_row_stride = c.k.stride(1); P.k_head_stride = c.k.stride(2);
    P.v_row_stride = c.v.stride(1); P.v_head_stride = c.v.stride(2);
    // cache_seqlens: tensor (reference) or -- superset, SURVEY quirk Q5 -- None (whole cache); an int is broadcast
    // by the Python wrapper
    if (cache_seqlens.has_value()) {
        TORCH_CHECK(cache_seqlens->is_cuda(), "seqlens_k must be on CUDA");
        TORCH_CHECK(cache_seqlens->scalar_type() == at::kInt, "seqlens_k must be int32");   // api.cpp:430
        TORCH_CHECK(cache_seqlens->numel() == batch, "seqlens_k must have the same number of elements as batch size");
        c.sl = cache_seqlens->contiguous();
        P.seqlens_k = c.sl.data_ptr<int32_t>();
    }
    const bool q_ok = reinterpret_cast<uintptr_t>(q.data_ptr()) % 4 == 0 && q.stride(0) % 2 == 0 && q.stride(2) % 2 == 0;
    c.q = q_ok ? q : q.contiguous();
    P.q_ptr = c.q.data_ptr(); P.k_ptr = c.k.data_ptr(); P.v_ptr = c.v.data_ptr();
    P.q_batch_stride = c.q.stride(0); P.q_row_stride = c.q.stride(1); P.q_head_stride = c.q.stride(2);
    P.heads = (int32_t)heads; P.kv_heads = (int32_t)kv_heads; P.batch = (int32_t)batch;
    P.seqlen_q = 1; P.seqlen_k = (int32_t)seqlen_k; P.head_dim = (int32_t)head_dim;
    P.is_bf16 = q.scalar_type() == at::kBFloat16;
    P.num_splits = (int32_t)num_splits;
}

// split count, workspace from the caching allocator (api.cpp:333-334), launch
void run_decode(mfa_params_t& P, const at::Tensor& q) {
    const int splits = mfa_decode_num_splits(&P);
    TORCH_CHECK(splits >= 1, "mfa_decode_num_splits failed: ", mfa_last_error());
    P.num_splits = splits;
    const size_t ws_bytes = mfa_decode_workspace_bytes(&P, splits);
    at::Tensor ws;
    if (ws_bytes) ws = at::empty({(int64_t)(ws_bytes / 4)}, q.options().dtype(at::kFloat));
    check_rc(mfa_decode_fwd(&P, ws_bytes ? ws.data_ptr() : nullptr, ws_bytes, stream_of(q)), "mfa_decode_fwd");
}

}  // namespace

// q_len == 1 decode over a contiguous or paged KV cache.  api.cpp:343-446.  `causal` is accepted and ignored exactly
// like the reference (api.cpp:349, quirk Q10).
static std::tuple<at::Tensor, at::Tensor> kvcache_impl(const at::Tensor& q, const at::Tensor& k_cache,
                                                       const at::Tensor& v_cache,
                                                       const std::optional<at::Tensor>& cache_seqlens,
                                                       const std::optional<at::Tensor>& block_table, int64_t num_splits,
                                                       bool want_lse,
                                                       const std::optional<at::Tensor>& k_new = std::nullopt,
                                                       const std::optional<at::Tensor>& v_new = std::nullopt,
                                                       int64_t window_left = -1) {
    DecodeCall c;
    decode_params(c, q, k_cache, v_cache, cache_seqlens, block_table, num_splits);
    if (window_left >= 0) {   // sliding window: the last window_left + 1 keys
        c.P.local_window = 1;
        c.P.window_size_left = (int32_t)window_left;
        c.P.window_size_right = -1;
    }
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor seqlens_new;
    if (k_new.has_value() || v_new.has_value()) {
        // KV-cache append: write the new rows at cache_seqlens, attend to cache_seqlens + s_new
        TORCH_CHECK(k_new.has_value() && v_new.has_value(), "k and v must be given together");
        TORCH_CHECK(cache_seqlens.has_value(), "appending k/v needs cache_seqlens (where to write)");
        TORCH_CHECK(c.k.is_same(k_cache) && c.v.is_same(v_cache), "appending k/v needs a cache that can be written in place (16-byte aligned, strides multiples of 8)");
        const at::Tensor& kn = *k_new;
        const at::Tensor& vn = *v_new;
        TORCH_CHECK(kn.is_cuda() && vn.is_cuda(), "k and v must be on CUDA");
        TORCH_CHECK(kn.scalar_type() == q.scalar_type() && vn.scalar_type() == q.scalar_type(), "k and v must have q's dtype");
        TORCH_CHECK(kn.dim() == 4 && kn.size(0) == q.size(0) && kn.size(2) == c.P.kv_heads && kn.size(3) == c.P.head_dim,
                    "k must have shape (batch, seqlen_new, kv_num_heads, head_dim)");
        TORCH_CHECK(vn.sizes() == kn.sizes(), "v must have the same shape as k");
        const at::Tensor knc = tma_ready(kn.stride(-1) == 1 ? kn : kn.contiguous());
        const at::Tensor vnc = tma_ready(vn.stride(-1) == 1 ? vn : vn.contiguous());
        seqlens_new = at::empty({q.size(0)}, q.options().dtype(at::kInt));
        mfa_append_params_t A{};
        A.k_new = knc.data_ptr(); A.v_new = vnc.data_ptr();
        A.knew_batch_stride = knc.stride(0); A.knew_row_stride = knc.stride(1); A.knew_head_stride = knc.stride(2);
        A.vnew_batch_stride = vnc.stride(0); A.vnew_row_stride = vnc.stride(1); A.vnew_head_stride = vnc.stride(2);
        A.k_cache = c.k.data_ptr(); A.v_cache = c.v.data_ptr();
        A.k_batch_stride = c.k.stride(0); A.k_row_stride = c.k.stride(1); A.k_head_stride = c.k.stride(2);
        A.v_batch_stride = c.v.stride(0); A.v_row_stride = c.v.stride(1); A.v_head_stride = c.v.stride(2);
        A.seqlens_k = c.P.seqlens_k; A.seqlens_out = seqlens_new.data_ptr<int32_t>();
        A.block_table = c.P.block_table; A.block_table_batch_stride = c.P.block_table_batch_stride;
        A.max_blocks_per_seq = c.P.max_blocks_per_seq; A.page_block_size = c.P.page_block_size;
        A.seqlen_k = c.P.seqlen_k;
        A.batch = c.P.batch; A.s_new = (int32_t)kn.size(1); A.kv_heads = c.P.kv_heads; A.head_dim = c.P.head_dim;
        A.is_bf16 = c.P.is_bf16;
        A.num_blocks = c.P.num_blocks;
        check_rc(mfa_kvcache_append(&A, stream_of(q)), "mfa_kvcache_append");
        c.P.seqlens_k = seqlens_new.data_ptr<int32_t>();
    }
    at::Tensor out = at::empty_like(q);   // api.cpp:405
    at::Tensor lse;   // (B, H, 1) fp32, the tensor the reference allocates and drops (api.cpp:436-437)
    if (want_lse) lse = at::empty({q.size(0), q.size(2), 1}, q.options().dtype(at::kFloat));
    if (q.numel() == 0) return {out, lse};
    if (c.P.seqlen_k == 0) {
        if (want_lse) lse.fill_(-std::numeric_limits<float>::infinity());
        return {out.zero_(), lse};
    }
    const bool o_ok = reinterpret_cast<uintptr_t>(out.data_ptr()) % 8 == 0 && out.stride(0) % 4 == 0 && out.stride(2) % 4 == 0;
    at::Tensor oc = o_ok ? out : at::empty(out.sizes(), out.options());
    c.P.o_ptr = oc.data_ptr();
    c.P.o_batch_stride = oc.stride(0); c.P.o_row_stride = oc.stride(1); c.P.o_head_stride = oc.stride(2);
    if (want_lse) c.P.softmax_lse_ptr = lse.data_ptr<float>();
    run_decode(c.P, q);
    if (!oc.is_same(out)) out.copy_(oc);
    return {out, lse};
}

at::Tensor mini_flash_attention_with_kvcache(const at::Tensor& q, const at::Tensor& k_cache, const at::Tensor& v_cache,
                                             std::optional<at::Tensor> cache_seqlens,
                                             std::optional<at::Tensor> block_table, bool /*causal*/, int64_t num_splits) {
    return std::get<0>(kvcache_impl(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, false));
}
std::tuple<at::Tensor, at::Tensor> mini_flash_attention_with_kvcache_lse(
    const at::Tensor& q, const at::Tensor& k_cache, const at::Tensor& v_cache, std::optional<at::Tensor> cache_seqlens,
    std::optional<at::Tensor> block_table, bool /*causal*/, int64_t num_splits) {
    return kvcache_impl(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, true);
}
// decode over an e4m3 KV cache; k_scale / v_scale: optional (kv_heads,) fp32 device tensors (the Python side expands
// floats and per-tensor scales)
std::tuple<at::Tensor, std::optional<at::Tensor>> mini_flash_attention_with_kvcache_fp8(
    const at::Tensor& q, const at::Tensor& k_cache, const at::Tensor& v_cache, std::optional<at::Tensor> cache_seqlens,
    std::optional<at::Tensor> block_table, int64_t num_splits, bool return_lse, int64_t window_left,
    std::optional<at::Tensor> k_scale, std::optional<at::Tensor> v_scale) {
    DecodeCall c;
    decode_params(c, q, k_cache, v_cache, cache_seqlens, block_table, num_splits, true);
    at::Tensor ks, vs;
    c.P.kv_fp8 = 1;
    c.P.kv_scale_stride = 1;
    if (k_scale.has_value()) {
        TORCH_CHECK(k_scale->is_cuda() && k_scale->scalar_type() == at::kFloat && k_scale->numel() == c.P.kv_heads, "k_scale must be a (kv_num_heads,) fp32 CUDA tensor");
        ks = k_scale->contiguous();
        c.P.k_scale_ptr = ks.data_ptr<float>();
    }
    if (v_scale.has_value()) {
        TORCH_CHECK(v_scale->is_cuda() && v_scale->scalar_type() == at::kFloat && v_scale->numel() == c.P.kv_heads, "v_scale must be a (kv_num_heads,) fp32 CUDA tensor");
        vs = v_scale->contiguous();
        c.P.v_scale_ptr = vs.data_ptr<float>();
    }
    if (window_left >= 0) {
        c.P.local_window = 1;
        c.P.window_size_left = (int32_t)window_left;
        c.P.window_size_right = -1;
    }
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor out = at::empty_like(q);
    at::Tensor lse;
    if (return_lse) lse = at::empty({q.size(0), q.size(2), 1}, q.options().dtype(at::kFloat));
    auto ret = [&]() -> std::tuple<at::Tensor, std::optional<at::Tensor>> {
        return {out, return_lse ? std::optional<at::Tensor>(lse) : std::nullopt};
    };
    if (q.numel() == 0) return ret();
    if (c.P.seqlen_k == 0) {
        if (return_lse) lse.fill_(-std::numeric_limits<float>::infinity());
        out.zero_();
        return ret();
    }
    const bool o_ok = reinterpret_cast<uintptr_t>(out.data_ptr()) % 8 == 0 && out.stride(0) % 4 == 0 && out.stride(2) % 4 == 0;
    at::Tensor oc = o_ok ? out : at::empty(out.sizes(), out.options());
    c.P.o_ptr = oc.data_ptr();
    c.P.o_batch_stride = oc.stride(0); c.P.o_row_stride = oc.stride(1); c.P.o_head_stride = oc.stride(2);
    if (return_lse) c.P.softmax_lse_ptr = lse.data_ptr<float>();
    run_decode(c.P, q);
    if (!oc.is_same(out)) out.copy_(oc);
    return ret();
}

std::tuple<at::Tensor, std::optional<at::Tensor>> mini_flash_attention_with_kvcache_ext(
    const at::Tensor& q, const at::Tensor& k_cache, const at::Tensor& v_cache, std::optional<at::Tensor> cache_seqlens,
    std::optional<at::Tensor> block_table, bool /*causal*/, int64_t num_splits, bool return_lse,
    std::optional<at::Tensor> k_new, std::optional<at::Tensor> v_new, int64_t window_left) {
    auto r = kvcache_impl(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, return_lse, k_new, v_new, window_left);
    return {std::get<0>(r), return_lse ? std::optional<at::Tensor>(std::get<1>(r)) : std::nullopt};
}

// KV-shard building block of the multi-GPU decode: normalised fp32 partial O (B,H,D) and LSE (B,H) over this rank's
// part of the KV axis.  New -- the arithmetic is the reference's split-KV partial (decode.cuh:645-661).
std::tuple<at::Tensor, at::Tensor> decode_partial(const at::Tensor& q, const at::Tensor& k_cache,
                                                  const at::Tensor& v_cache, std::optional<at::Tensor> cache_seqlens,
                                                  std::optional<at::Tensor> block_table, int64_t num_splits) {
    DecodeCall c;
    decode_params(c, q, k_cache, v_cache, cache_seqlens, block_table, num_splits);
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor o = at::empty({q.size(0), q.size(2), q.size(3)}, q.options().dtype(at::kFloat));
    at::Tensor lse = at::empty({q.size(0), q.size(2)}, q.options().dtype(at::kFloat));
    c.P.o_ptr = o.data_ptr();
    c.P.softmax_lse_ptr = lse.data_ptr<float>();
    c.P.partial_out = 1;
    run_decode(c.P, q);
    return {o, lse};
}

// KV-sharded decode with the exchange fused into the merge kernel (include/mfa_b200.h mfa_decode_fwd_kv_sharded,
// csrc/xchg_sm100.cu): this rank's rows of the KV axis in, the full result (B,1,H,D) out on every rank.  `xchg` is a
// connected mfa_xchg_t* passed as an integer (parallel.KVShardExchange.handle).
at::Tensor decode_kv_sharded_fused(const at::Tensor& q, const at::Tensor& k_cache, const at::Tensor& v_cache,
                                   std::optional<at::Tensor> cache_seqlens, std::optional<at::Tensor> block_table,
                                   int64_t num_splits, int64_t xchg) {
    DecodeCall c;
    decode_params(c, q, k_cache, v_cache, cache_seqlens, block_table, num_splits);
    TORCH_CHECK(c.P.seqlen_k > 0 && q.numel() > 0, "KV-sharded decode needs a non-empty problem");
    TORCH_CHECK(xchg != 0, "null exchange handle");
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor out = at::empty(q.sizes(), q.options());
    c.P.o_ptr = out.data_ptr();
    c.P.o_batch_stride = out.stride(0); c.P.o_row_stride = out.stride(1); c.P.o_head_stride = out.stride(2);
    const int splits = mfa_decode_num_splits(&c.P);
    TORCH_CHECK(splits >= 1, "mfa_decode_num_splits failed: ", mfa_last_error());
    c.P.num_splits = splits;
    const size_t ws_bytes = mfa_decode_kv_sharded_workspace_bytes(&c.P, splits);
    at::Tensor ws = at::empty({(int64_t)(ws_bytes / 4)}, q.options().dtype(at::kFloat));
    check_rc(mfa_decode_fwd_kv_sharded(&c.P, ws.data_ptr(), ws_bytes, reinterpret_cast<mfa_xchg_t*>(xchg), stream_of(q)),
             "mfa_decode_fwd_kv_sharded");
    return out;
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.doc() = "mini_flash_attention_b200 compiled host shim over libmfa_b200.so";
    namespace py = pybind11;
    m.def("mini_flash_attention_forward", &mini_flash_attention_forward, py::arg("q"), py::arg("k"), py::arg("v"),
          py::arg("out") = py::none(), py::arg("is_causal") = false, py::arg("window_size_left") = -1,
          py::arg("window_size_right") = -1);
    m.def("mini_flash_attention_varlen_forward", &mini_flash_attention_varlen_forward, py::arg("q"), py::arg("k"),
          py::arg("v"), py::arg("cu_seqlens_q"), py::arg("cu_seqlens_k"), py::arg("max_seqlen_q"),
          py::arg("max_seqlen_k"), py::arg("causal"), py::arg("window_size_left"), py::arg("window_size_right"),
          py::arg("block_table") = py::none());
    m.def("mini_flash_attention_with_kvcache", &mini_flash_attention_with_kvcache, py::arg("q"), py::arg("k_cache"),
          py::arg("v_cache"), py::arg("cache_seqlens"), py::arg("block_table"), py::arg("causal"),
          py::arg("num_splits") = 0);
    m.def("mini_flash_attention_forward_lse", &mini_flash_attention_forward_lse, py::arg("q"), py::arg("k"), py::arg("v"),
          py::arg("out") = py::none(), py::arg("is_causal") = false, py::arg("window_size_left") = -1,
          py::arg("window_size_right") = -1);
    m.def("mini_flash_attention_varlen_forward_lse", &mini_flash_attention_varlen_forward_lse, py::arg("q"),
          py::arg("k"), py::arg("v"), py::arg("cu_seqlens_q"), py::arg("cu_seqlens_k"), py::arg("max_seqlen_q"),
          py::arg("max_seqlen_k"), py::arg("causal"), py::arg("window_size_left"), py::arg("window_size_right"),
          py::arg("block_table") = py::none());
    m.def("mini_flash_attention_with_kvcache_lse", &mini_flash_attention_with_kvcache_lse, py::arg("q"),
          py::arg("k_cache"), py::arg("v_cache"), py::arg("cache_seqlens"), py::arg("block_table"), py::arg("causal"),
          py::arg("num_splits") = 0);
    m.def("mini_flash_attention_forward_ext", &mini_flash_attention_forward_ext, py::arg("q"), py::arg("k"), py::arg("v"),
          py::arg("out"), py::arg("is_causal"), py::arg("window_size_left"), py::arg("window_size_right"),
          py::arg("return_lse"), py::arg("bottom_right"));
    m.def("mini_flash_attention_varlen_forward_ext", &mini_flash_attention_varlen_forward_ext, py::arg("q"),
          py::arg("k"), py::arg("v"), py::arg("cu_seqlens_q"), py::arg("cu_seqlens_k"), py::arg("max_seqlen_q"),
          py::arg("max_seqlen_k"), py::arg("causal"), py::arg("window_size_left"), py::arg("window_size_right"),
          py::arg("block_table"), py::arg("return_lse"), py::arg("bottom_right"));
    m.def("mini_flash_attention_with_kvcache_ext", &mini_flash_attention_with_kvcache_ext, py::arg("q"),
          py::arg("k_cache"), py::arg("v_cache"), py::arg("cache_seqlens"), py::arg("block_table"), py::arg("causal"),
          py::arg("num_splits"), py::arg("return_lse"), py::arg("k_new"), py::arg("v_new"), py::arg("window_left") = -1);
    m.def("mini_flash_attention_with_kvcache_fp8", &mini_flash_attention_with_kvcache_fp8, py::arg("q"), py::arg("k_cache"),
          py::arg("v_cache"), py::arg("cache_seqlens"), py::arg("block_table"), py::arg("num_splits"), py::arg("return_lse"),
          py::arg("window_left"), py::arg("k_scale"), py::arg("v_scale"));
    m.def("decode_kv_sharded_fused", &decode_kv_sharded_fused, py::arg("q"), py::arg("k_cache"), py::arg("v_cache"),
          py::arg("cache_seqlens"), py::arg("block_table"), py::arg("num_splits"), py::arg("xchg"));
    m.def("decode_partial", &decode_partial, py::arg("q"), py::arg("k_cache"), py::arg("v_cache"),
          py::arg("cache_seqlens"), py::arg("block_table"), py::arg("num_splits") = 0);
}
