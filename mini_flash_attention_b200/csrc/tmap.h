// tmap.h -- host-side TMA tensor-map construction (cuTensorMapEncodeTiled fetched at run time through
// cudaGetDriverEntryPoint, so the library never links libcuda directly).
#pragma once
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

namespace mfa {

// Encode a 4-D tiled tensor map over a 16-bit tensor.
//   dims[0] is the innermost (contiguous) dimension; strides_elems[i] is the element stride of dims[i]
//   (strides_elems[0] must be 1).  box[i] is the tile extent.  swizzle: CU_TENSOR_MAP_SWIZZLE_128B etc.
// Returns nullptr on success, else a static error string.
// the same for a tensor of 1-byte elements (e4m3 KV cache): strides in elements = bytes, multiples of 16
const char* encode_tmap_4d_u8(CUtensorMap* out, const void* base, const uint64_t dims[4], const int64_t strides_elems[4],
                              const uint32_t box[4], CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B);
const char* encode_tmap_4d(CUtensorMap* out, const void* base, bool is_bf16, const uint64_t dims[4],
                           const int64_t strides_elems[4], const uint32_t box[4],
                           CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B);

}  // namespace mfa
