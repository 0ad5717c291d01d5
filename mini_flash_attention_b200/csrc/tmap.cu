#include "tmap.h"

#include <cstdio>
#include <mutex>

namespace mfa {

namespace {
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
        if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}
}  // namespace

static const char* encode_impl(CUtensorMap* out, const void* base, CUtensorMapDataType dtype, int eb, const uint64_t dims[4],
                               const int64_t strides_elems[4], const uint32_t box[4], CUtensorMapSwizzle swizzle) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return "cuTensorMapEncodeTiled entry point not available";
    if (reinterpret_cast<uintptr_t>(base) & 15) return "tensor base address must be 16-byte aligned for TMA";
    cuuint64_t gdim[4];
    cuuint64_t gstride[3];
    cuuint32_t gbox[4];
    cuuint32_t estr[4] = {1, 1, 1, 1};
    for (int i = 0; i < 4; ++i) {
        gdim[i] = dims[i] ? dims[i] : 1;
        gbox[i] = box[i];
    }
    for (int i = 1; i < 4; ++i) {
        int64_t s = strides_elems[i];
        if (dims[i] <= 1) {
            // size-1 dimension: its stride is never applied, any legal value works; use the extent below it
            s = (i == 1) ? static_cast<int64_t>(gdim[0]) : static_cast<int64_t>(gstride[i - 2] / eb * gdim[i - 1]);
            const int64_t q = 16 / eb;   // elements per 16 bytes
            if (s % q) s = (s + q - 1) / q * q;
        } else if (s <= 0) {
            // a broadcast (expand()ed, stride 0) or reversed dimension of extent > 1: a tensor map cannot express it,
            // and inventing a stride would read the wrong rows -- the caller must materialise the tensor
            return "zero or negative stride on a dimension of extent > 1 (broadcast view): make the tensor contiguous";
        }
        if ((s * eb) % 16) return "tensor strides must be multiples of 16 bytes for TMA";
        gstride[i - 1] = static_cast<cuuint64_t>(s) * eb;
    }
    CUresult r = fn(out, dtype, 4,
                    const_cast<void*>(base), gdim, gstride, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        static thread_local char msg[256];
        snprintf(msg, sizeof(msg),
                 "cuTensorMapEncodeTiled failed (%d): dims=(%llu,%llu,%llu,%llu) strides_B=(%llu,%llu,%llu) "
                 "box=(%u,%u,%u,%u)",
                 (int)r, (unsigned long long)gdim[0], (unsigned long long)gdim[1], (unsigned long long)gdim[2],
                 (unsigned long long)gdim[3], (unsigned long long)gstride[0], (unsigned long long)gstride[1],
                 (unsigned long long)gstride[2], gbox[0], gbox[1], gbox[2], gbox[3]);
        return msg;
    }
    return nullptr;
}

const char* encode_tmap_4d(CUtensorMap* out, const void* base, bool is_bf16, const uint64_t dims[4],
                           const int64_t strides_elems[4], const uint32_t box[4], CUtensorMapSwizzle swizzle) {
    return encode_impl(out, base, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, dims,
                       strides_elems, box, swizzle);
}

const char* encode_tmap_4d_u8(CUtensorMap* out, const void* base, const uint64_t dims[4], const int64_t strides_elems[4],
                              const uint32_t box[4], CUtensorMapSwizzle swizzle) {
    return encode_impl(out, base, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, dims, strides_elems, box, swizzle);
}

}  // namespace mfa
