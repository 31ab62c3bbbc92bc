// Host-side CUtensorMap construction without linking libcuda: the encode entry point is
// resolved through the runtime (cudaGetDriverEntryPoint), so the library also loads on a
// GPU-less build box.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

namespace dmpi {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    if (getenv("DMPI_TRACE")) fprintf(stderr, "[dmpi] resolving cuTensorMapEncodeTiled\n");
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
      fprintf(stderr, "[dmpi] cuTensorMapEncodeTiled entry point unavailable (%s)\n", cudaGetErrorString(e));
      return nullptr;
    }
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 tensor of rank `rank` (2..5); dims[0] is the contiguous dimension.  strides_elems[i] is the
// stride of dims[i+1] in elements.  box[] in elements.  128-byte swizzle, zero OOB fill.
inline int encode_bf16_map(CUtensorMap* out, const void* base, int rank, const int64_t* dims,
                           const int64_t* strides_elems, const int* box) {
  EncodeTiledFn fn = get_encode_tiled();
  if (!fn) return -1;
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = static_cast<cuuint64_t>(dims[i]);
    bx[i] = static_cast<cuuint32_t>(box[i]);
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = static_cast<cuuint64_t>(strides_elems[i]) * 2ull;
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(rank), const_cast<void*>(base),
                  gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr,
            "[dmpi] cuTensorMapEncodeTiled failed (%d): base=%p rank=%d dims=[%lld,%lld,%lld,%lld] "
            "strides=[%lld,%lld,%lld] box=[%d,%d,%d,%d]\n",
            static_cast<int>(r), base, rank, (long long)dims[0], (long long)(rank > 1 ? dims[1] : 0),
            (long long)(rank > 2 ? dims[2] : 0), (long long)(rank > 3 ? dims[3] : 0),
            (long long)(rank > 1 ? strides_elems[0] : 0), (long long)(rank > 2 ? strides_elems[1] : 0),
            (long long)(rank > 3 ? strides_elems[2] : 0), box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0,
            rank > 3 ? box[3] : 0);
    return -2;
  }
  return 0;
}

}  // namespace dmpi
