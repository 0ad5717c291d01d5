/*
 * mfa_b200.h -- C-ABI of the B200-native (sm_100a) attention-forward library.
 *
 * This is the drop-in boundary for the reference's extension surface
 *   /root/reference/csrc/api.cpp:4-9            (pybind: the three _C functions)
 *   /root/reference/csrc/mfa/api.h:8-40         (their C++ prototypes)
 *   /root/reference/csrc/mfa/flash.h:8-77       (ForwardParams + the two launchers)
 * restated with plain pointers and sizes: no torch types, no C++ types.  The caller owns all
 * memory (inputs, output, workspaces); every call is asynchronous on the given CUDA stream and
 * returns an error code instead of aborting the process (the reference exit(1)s on CUDA errors,
 * csrc/mfa/api.cpp:16-24 -- deliberately not replicated).
 *
 * All strides are in ELEMENTS (fp16/bf16), exactly like ForwardParams (flash.h:17-31).
 * The last (head_dim) dimension must be contiguous (api.cpp:135-137).
 */
#ifndef MFA_B200_H_
#define MFA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MFA_B200_ABI_VERSION 4

/* error codes (0 == success).  mfa_last_error() returns the message of the last failure on the
 * calling thread. */
enum {
    MFA_OK = 0,
    MFA_ERR_INVALID_ARG = 1,   /* shape / dtype / alignment contract violated (TORCH_CHECK analogue) */
    MFA_ERR_UNSUPPORTED = 2,   /* e.g. head_dim not a multiple of 8 or > 256, page size not usable */
    MFA_ERR_CUDA = 3,          /* a CUDA runtime / driver call failed (message has the CUDA string) */
    MFA_ERR_NO_DEVICE = 4      /* not an sm_100 device */
};

/*
 * POD mirror of mfa::ForwardParams (/root/reference/csrc/mfa/flash.h:8-73).
 * Field meaning is identical unless noted.  Zero-initialise, then fill.
 */
typedef struct mfa_params {
    /* device pointers (fp16 or bf16, see is_bf16) -- flash.h:11-14 */
    const void* q_ptr;
    const void* k_ptr;
    const void* v_ptr;
    void*       o_ptr;

    /* element strides -- flash.h:16-30.  batch strides are ignored when cu_seqlens_q != NULL
     * (packed varlen) and, for K/V, when block_table != NULL (paged). */
    int64_t q_batch_stride, q_head_stride, q_row_stride;
    int64_t k_batch_stride, k_head_stride, k_row_stride;
    int64_t v_batch_stride, v_head_stride, v_row_stride;
    int64_t o_batch_stride, o_head_stride, o_row_stride;

    int32_t is_causal;          /* top-left aligned: key col masked iff col > row (prefill.cuh:416) */
    int32_t window_size_left;   /* accepted and ignored, like the reference (flash.h:33-34), unless local_window */
    int32_t window_size_right;

    int32_t heads;              /* query heads */
    int32_t kv_heads;           /* heads % kv_heads == 0; kv_head = head / (heads/kv_heads) (prefill.cuh:49) */
    int32_t batch;
    int32_t seqlen_q;           /* fixed-length: Sq ; varlen: max_seqlen_q ; decode: 1 */
    int32_t seqlen_k;           /* fixed-length: Sk ; varlen: max_seqlen_k ; decode: cache capacity in rows */
    int32_t head_dim;           /* <= 256, multiple of 8 */
    int32_t is_bf16;            /* 0 = fp16, 1 = bf16 */
    float   softmax_scale;      /* 0 => 1/sqrt(head_dim) (the reference's only choice, api.cpp:99) */

    /* packed variable-length batches -- flash.h:49-54 (int32, batch+1 entries, device memory) */
    const int32_t* cu_seqlens_q;
    const int32_t* cu_seqlens_k;
    int32_t total_q;            /* rows in the packed q / o tensors (varlen only) */
    int32_t total_k;            /* rows in the packed k / v tensors (varlen, non-paged only) */

    /* paged KV -- flash.h:57-61.  k/v are (num_blocks, page_block_size, kv_heads, head_dim) */
    const int32_t* block_table;          /* (batch, max_blocks_per_seq) int32, device */
    int32_t block_table_batch_stride;    /* elements */
    int32_t max_blocks_per_seq;
    int32_t page_block_size;
    int32_t num_blocks;
    int64_t k_cache_block_stride;        /* elements */
    int64_t v_cache_block_stride;

    /* decode -- flash.h:64-72 */
    const int32_t* seqlens_k;   /* (batch) int32 device; NULL => every row attends to seqlen_k keys */
    int32_t num_splits;         /* <= 0: heuristic ; clamped to the number of 64-row KV blocks (api.cpp:320-327) */
    float*  softmax_lse_ptr;    /* optional fp32 natural-log LSE output, may be NULL (-inf for rows that see no key):
                                 * decode (batch, heads); prefill fixed (batch, heads, seqlen_q), packed (heads, total_q) */

    /* multi-GPU KV-sharded decode: when partial_out != 0 the decode writes a NORMALISED fp32
     * partial O (batch, heads, head_dim) to o_ptr instead of fp16/bf16, and softmax_lse_ptr
     * (required) receives the partial LSE (-inf for empty shards).  New; no reference analogue. */
    int32_t partial_out;
    /* Prefill only, with is_causal: align the causal mask to the BOTTOM-RIGHT corner -- key col visible iff
     * col <= row + (len_k - len_q), per sequence (flash-attn >= 2.1 convention; what chunked prefill against a longer
     * cache needs).  0 = the reference's top-left mask (prefill.cuh:416).  New. */
    int32_t causal_bottom_right;
    /* HONOUR window_size_left / window_size_right as a sliding-window (band) mask (decode: the query sits at position
     * len - 1 and sees the last window_size_left + 1 keys; window_size_right is irrelevant).  Prefill: -- key col visible iff
     * row + shift - left <= col <= row + shift + right, a bound < 0 absent, is_causal forcing right = 0; shift = 0
     * (top-left, the reference's alignment) or len_k - len_q per sequence when causal_bottom_right is set (flash-attn's
     * window_size convention).  KV tiles outside the band are skipped.  0 = the reference's behaviour: the two window
     * fields are ignored (flash.h:33-34, README.md:153 "not implemented").  New. */
    int32_t local_window;
    /* e4m3 KV cache, DECODE only (the reference gates the dtype to fp16/bf16, api.cpp:126-129): half the
     * HBM bytes per key.  kv_fp8 != 0: k_ptr / v_ptr point at 1-byte e4m3 elements (the K/V strides stay in ELEMENTS, i.e.
     * bytes, and must be multiples of 16); the dequantised element is byte_value * scale with
     * scale = k_scale_ptr[kv_head * kv_scale_stride] (NULL => 1).  q / out keep the 16-bit type of is_bf16; the tiles are
     * converted to fp16 in registers (exact) and the MMAs run in fp16 (a bf16 q beyond the fp16 range saturates).
     * head_dim must be a multiple of 128.  mfa_prefill_fwd and mfa_kvcache_append reject kv_fp8.  New in ABI v4. */
    int32_t kv_fp8;
    int32_t kv_scale_stride;             /* 1: one scale per KV head; 0: one scale for the whole tensor */
    const float* k_scale_ptr;            /* device, fp32 */
    const float* v_scale_ptr;
} mfa_params_t;

/* --- introspection ------------------------------------------------------------------------ */
int         mfa_abi_version(void);
const char* mfa_last_error(void);
/* 1 if the current device is compute capability 10.x, else 0 */
int         mfa_device_supported(void);

/* --- prefill: replaces mfa::run_flash_attention_forward (flash.h:76, flash.cu:74-82) ---------
 * Fixed-length (cu_seqlens_q == NULL), packed varlen (cu_seqlens_* != NULL) and varlen with paged
 * K/V (block_table != NULL) in one entry point, like the reference. */
int mfa_prefill_fwd(const mfa_params_t* p, void* cuda_stream);

/* --- decode (q_len == 1): replaces mfa::run_flash_attention_with_kv_cache (flash.h:77,
 *     flash.cu:85-93) plus forward_params_set_split_kv (api.cpp:305-340). ----------------------
 * mfa_decode_num_splits: the split count the library will use for this problem
 *   (requested <= 0 => heuristic; result clamped to ceil(seqlen_k / 64) like api.cpp:325-327).
 * mfa_decode_workspace_bytes: bytes of scratch needed for that split count (0 when 1 split).
 * mfa_decode_fwd: launches split-KV kernel (+ LSE combine when splits > 1) on the stream.
 *   `workspace` must hold mfa_decode_workspace_bytes(...) bytes; it needs no initialisation. */
int    mfa_decode_num_splits(const mfa_params_t* p);
size_t mfa_decode_workspace_bytes(const mfa_params_t* p, int num_splits);
int    mfa_decode_fwd(const mfa_params_t* p, void* workspace, size_t workspace_bytes, void* cuda_stream);

/* --- decode without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 *     callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 *     it).  Same arguments, result and errors as mfa_decode_fwd, but the split-KV scratch is a file-scope
 *     `static __device__` array inside the library: MFA_DECODE_STATIC_WS_BYTES (8 MiB) of device memory on every GPU that
 *     loads the library, held until the process exits.  The call performs no cudaMalloc / cudaFree and calls no allocating
 *     library routine; inputs and outputs are still the caller's.  A problem whose scratch
 *     (mfa_decode_workspace_bytes for the split count in use) exceeds the array is REJECTED with MFA_ERR_UNSUPPORTED and a
 *     message naming both sizes -- pass a smaller p->num_splits, or use mfa_decode_fwd.  One split needs no scratch.
 *     All calls on a device share the one array and it keeps the previous call's contents (nothing in it has to start at
 *     zero: every partial is written before it is read).  Calls that OVERLAP on the GPU -- two streams, or two host
 *     threads on one device -- race on it; calls on one stream are ordered and safe. */
#define MFA_DECODE_STATIC_WS_BYTES (8u << 20)
size_t mfa_decode_static_workspace_limit(void);
int    mfa_decode_fwd_static(const mfa_params_t* p, void* cuda_stream);

/* --- KV-cache append (the reference's docstring mentions it, interface.py:110-111, but does not
 *     implement it): write s_new new K/V rows per sequence into the cache IN PLACE at logical positions
 *     seqlens_k[b] .. seqlens_k[b] + s_new - 1 (paged through block_table, or contiguous), and -- when
 *     seqlens_out != NULL -- store seqlens_k[b] + s_new there, which is the length the following decode call attends
 *     to (cache_seqlens itself is not modified, like flash-attn's flash_attn_with_kvcache(k=, v=)).  Rows that would
 *     fall beyond the cache capacity are not written and seqlens_out is clamped to the capacity. */
typedef struct mfa_append_params {
    const void* k_new;          /* (batch, s_new, kv_heads, head_dim) fp16/bf16 */
    const void* v_new;
    int64_t knew_batch_stride, knew_row_stride, knew_head_stride;   /* elements; last dim contiguous */
    int64_t vnew_batch_stride, vnew_row_stride, vnew_head_stride;
    void* k_cache;              /* (batch, seqlen_k, kv_heads, head_dim) or paged (num_blocks, page, kv_heads, head_dim) */
    void* v_cache;
    int64_t k_batch_stride, k_row_stride, k_head_stride;   /* batch stride = block stride when paged */
    int64_t v_batch_stride, v_row_stride, v_head_stride;
    const int32_t* seqlens_k;   /* (batch) int32 device, required */
    int32_t* seqlens_out;       /* (batch) int32 device, optional */
    const int32_t* block_table; /* NULL => contiguous cache */
    int32_t block_table_batch_stride;
    int32_t max_blocks_per_seq;
    int32_t page_block_size;
    int32_t seqlen_k;           /* capacity in rows of a contiguous cache (paged: max_blocks_per_seq * page) */
    int32_t batch, s_new, kv_heads, head_dim;   /* head_dim multiple of 8; all 16-byte aligned rows */
    int32_t is_bf16;            /* element size only (2 bytes either way) */
    int32_t num_blocks;         /* paged: pages in the cache tensors; a block_table entry outside [0, num_blocks) is never
                                 * written through (the row is skipped like a row beyond the capacity).  0 = unchecked. */
} mfa_append_params_t;
int mfa_kvcache_append(const mfa_append_params_t* p, void* cuda_stream);

/* --- LSE merge of R normalised partial results (the consumer of the multi-GPU all-gather and
 *     the same arithmetic as the reference's combine kernel, decode.cuh:665-755, but with a
 *     max-subtracted log-sum-exp).
 *   packed: (R, rows, head_dim + 1) fp32, per row [O_0 .. O_{D-1}, LSE]   (rows = batch*heads)
 *   out:    (rows, head_dim) fp16/bf16 with row stride out_row_stride (elements)
 *   lse_out: optional (rows) fp32 */
int mfa_combine_partials(const float* packed, int num_partials, int rows, int head_dim,
                         void* out, int64_t out_row_stride, int is_bf16, float* lse_out,
                         void* cuda_stream);

/* --- KV-sharded decode with the exchange fused into the merge kernel over NVLink peer memory (no reference
 *     analogue -- the reference is single-GPU).  Every rank owns a slice of the KV axis of every sequence; q is
 *     replicated.  mfa_decode_fwd_kv_sharded launches the split-KV kernel over the local rows and then ONE kernel that
 *     merges the local splits, pushes the normalised row + LSE into every peer's gather buffer (P2P stores + release
 *     flag), waits for the peers' rows and writes the final output (p->o_ptr, out dtype; optional p->softmax_lse_ptr):
 *     no NCCL call, no host synchronisation.  All ranks must call it the same number of times, in the same order.
 *     The step counter (epoch) lives in device memory, so the launches are identical from step to step and the call
 *     can be captured in a CUDA graph and replayed.
 *   mfa_xchg_create:  allocates this rank's gather buffer (cudaMalloc on the current device), max_rows >= batch*heads
 *   mfa_xchg_export:  writes the MFA_XCHG_HANDLE_BYTES-byte IPC handle of that buffer (exchange it out of band, e.g.
 *                     with torch.distributed.all_gather_object)
 *   mfa_xchg_connect: handles = world x MFA_XCHG_HANDLE_BYTES bytes in rank order (cudaIpcOpenMemHandle per peer)
 *   mfa_xchg_connect_ptrs: the same with raw device pointers (ranks that live in ONE process; tests)
 *   mfa_xchg_local_ptr: the local buffer's device pointer (for mfa_xchg_connect_ptrs)
 *   mfa_xchg_status:  0, or 1 + the rank whose push did not arrive within the timeout (synchronises the device).
 *                     Sticky.  A row whose peers timed out is written as NaN (output and LSE), never as a partial sum.
 *                     The timeout is ~2 s (MFA_XCHG_TIMEOUT_MS in the environment of mfa_xchg_create overrides it).
 *   mfa_xchg_clear_status: returns the status like mfa_xchg_status and resets it to 0
 *   mfa_decode_kv_sharded_workspace_bytes: scratch for num_splits (>= 1) splits: splits * batch*heads * (D+1) floats */
#define MFA_XCHG_HANDLE_BYTES 64
#define MFA_XCHG_MAX_WORLD 8
typedef struct mfa_xchg mfa_xchg_t;
int    mfa_xchg_create(int world, int rank, int max_rows, int head_dim, mfa_xchg_t** out);
int    mfa_xchg_export(mfa_xchg_t* x, void* handle_out);
int    mfa_xchg_connect(mfa_xchg_t* x, const void* handles);
int    mfa_xchg_connect_ptrs(mfa_xchg_t* x, void* const* peer_ptrs);
void*  mfa_xchg_local_ptr(mfa_xchg_t* x);
int    mfa_xchg_status(mfa_xchg_t* x);
int    mfa_xchg_clear_status(mfa_xchg_t* x);
int    mfa_xchg_destroy(mfa_xchg_t* x);
size_t mfa_decode_kv_sharded_workspace_bytes(const mfa_params_t* p, int num_splits);
int    mfa_decode_fwd_kv_sharded(const mfa_params_t* p, void* workspace, size_t workspace_bytes, mfa_xchg_t* x,
                                 void* cuda_stream);

/* tuning aid: libraries built with -DMFA_TRACE write clock64() stamps of prefill CTA (0,0,0) into this device
 * buffer ([3 roles][64 iterations][8 events] int64); a no-op in normal builds.  NULL disables. */
void mfa_debug_set_trace(void* device_buffer);
/* number of kernel launches issued by this library since load (bench.py reports it) */
uint64_t mfa_launch_count(void);

/* --- single-tile tcgen05 self test (tests/ only; isolates UMMA descriptor / TMEM layout conventions)
 *   mode 0: d[128 x 128] = a[128 x kdim] * b[128 x kdim]^T     (kdim 64 or 128; the Q K^T shape)
 *   mode 1: d[128 x ncols] = a[128 x 128] * b[128 x ncols]     (a staged in TMEM; the P V shape)
 *   a, b: 16-bit row-major device buffers; d: fp32 row-major; lbo/sbo: 0 = library defaults. */
int mfa_selftest_umma(int mode, const void* a, const void* b, float* d, int kdim, int ncols, int is_bf16,
                      uint32_t lbo, uint32_t sbo, void* cuda_stream);

/* CTA-pair (cluster of 2, tcgen05 cta_group::2) self test / MMA-rate probe (csrc/selftest2_sm100.cu):
 *   mode 0: d[256 x 128] = a[256 x 128] * b[128 x 128]^T  (pair, SS)   mode 1: d = a[256 x 128] * b[128 x 128] (pair, TS)
 *   mode 2 / 3: the same per-CTA tile with a single CTA (rows 0-127 only), for rate comparison.
 *   reps >= 1 repeats the 8 K-steps accumulating; cycles (optional, device int64[2]) receives clock64() ticks: [0] until completion, [1] until all MMAs were issued. */
int mfa_selftest_umma2(int mode, const void* a, const void* b, float* d, int is_bf16, int reps, long long* cycles,
                       void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* MFA_B200_H_ */
