// Host-side TMA descriptor construction.  cuTensorMapEncodeTiled is resolved through the runtime
// (cudaGetDriverEntryPoint) so the library links against cudart only.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace llmfp8 {

// 2D row-major tensor: dim0 (inner, contiguous) x dim1 (outer) with a row pitch in bytes; box = box0 x box1 elements.
int make_tensor_map_2d(CUtensorMap* map, const void* base, CUtensorMapDataType dtype, int elem_bytes, int64_t dim0, int64_t dim1,
                       int64_t pitch_bytes, int box0, int box1, CUtensorMapSwizzle swizzle);

}  // namespace llmfp8
