// Host-side tensor-map construction shared by the tcgen05 kernels' launchers (cuTensorMapEncodeTiled through the runtime's
// driver entry point: the library links no libcuda).
#pragma once
#include <mutex>

#include "common.cuh"

// ------------------------------------------------------------------------------------------------ tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// bf16 tensor map; dims/box innermost first; strides in ELEMENTS for dims 1..rank-1
static int make_tmap_bf16(CUtensorMap* tm, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_elems,
                          const uint32_t* box, int ck) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    ddimb_set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    return DDIMB_ERR_DRIVER;
  }
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) gs[i - 1] = strides_elems[i - 1] * 2;
  }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) {
    ddimb_set_error("tensor map base %p not 16-byte aligned", ptr);
    return DDIMB_ERR_ARG;
  }
  for (int i = 0; i < rank - 1; ++i)
    if (gs[i] % 16 != 0) {
      ddimb_set_error("tensor map stride %d = %llu bytes is not a multiple of 16", i, (unsigned long long)gs[i]);
      return DDIMB_ERR_ARG;
    }
  CUtensorMapSwizzle sw = (ck == 64) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    ddimb_set_error("cuTensorMapEncodeTiled failed: CUresult %d (rank %d dims %llu,%llu,%llu box %u,%u,%u)", (int)r, rank,
                    (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                    (unsigned long long)(rank > 2 ? dims[2] : 0), box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0);
    return DDIMB_ERR_TMAP;
  }
  return DDIMB_OK;
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

static int next_pow2_cols(int c) {
  int v = 32;
  while (v < c) v <<= 1;
  return v;
}

