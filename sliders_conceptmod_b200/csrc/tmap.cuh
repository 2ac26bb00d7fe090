// Host-side TMA descriptor construction.  cuTensorMapEncodeTiled is resolved through
// cudaGetDriverEntryPoint so the library has no link-time dependency on libcuda.so
// (it must load and export its symbols on a machine without a driver).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>
#include <string.h>
#include <unordered_map>
#include "common.cuh"

namespace slider {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// rank-`rank` tiled map; dims/box innermost first; strides in BYTES for dims 1..rank-1.
// elem_strides (optional, per dimension): the map visits every elem_strides[i]-th element along dimension i; a box of
// N * elem_strides[i] then lands N elements (stride-2 convolutions).
inline int make_tmap(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box, uint32_t inner_bytes,
                     const uint32_t* elem_strides = nullptr) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return SLIDER_ERR_CUDA;
  }
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t bdim[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = elem_strides ? elem_strides[i] : 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUtensorMapSwizzle sw = inner_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                          : inner_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : inner_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                              : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUtensorMapDataType dt = dtype == SLIDER_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                           : dtype == SLIDER_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bdim, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu,%llu,%llu,%llu box %u,%u,%u,%u)",
              (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0],
              rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return SLIDER_ERR_CUDA;
  }
  return SLIDER_OK;
}

// ---- descriptor cache ----------------------------------------------------------------------------------------
// A descriptor is a pure function of (pointer, dtype, shape, strides, box, swizzle), so it is encoded once per
// distinct key and per host thread; callers that repeat the same launches (the eager trainers: ~3 700 GEMM launches per
// step, 5-7 descriptors each) pay a hash lookup instead of cuTensorMapEncodeTiled.
struct TmapKey {
  const void* base;
  uint64_t dims[4], str[3];
  uint32_t box[4], est[4];
  uint32_t inner;
  int32_t dtype, rank;
  bool operator==(const TmapKey& o) const { return memcmp(this, &o, sizeof(TmapKey)) == 0; }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(TmapKey) / 8; ++i) { h ^= w[i]; h *= 1099511628211ull; }
    return (size_t)h;
  }
};
static_assert(sizeof(TmapKey) % 8 == 0, "TmapKey is hashed as 64-bit words");

inline std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash>& tmap_cache() {
  static thread_local std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  return cache;
}
// called at the START of an entry point (never between two lookups of one launch: the pointers handed out by
// cached_tmap stay valid until the next trim)
inline void tmap_cache_trim() {
  auto& c = tmap_cache();
  if (c.size() > 16384) c.clear();
}
// NULL on failure (the error text is set)
inline const CUtensorMap* cached_tmap(int dtype, int rank, const void* base, const uint64_t* dims,
                                      const uint64_t* strides_bytes, const uint32_t* box, uint32_t inner_bytes,
                                      const uint32_t* elem_strides = nullptr) {
  TmapKey k;
  memset(&k, 0, sizeof(k));
  k.base = base; k.inner = inner_bytes; k.dtype = dtype; k.rank = rank;
  for (int i = 0; i < rank; ++i) {
    k.dims[i] = dims[i]; k.box[i] = box[i]; k.est[i] = elem_strides ? elem_strides[i] : 1u;
    if (i > 0) k.str[i - 1] = strides_bytes[i - 1];
  }
  auto& c = tmap_cache();
  auto it = c.find(k);
  if (it != c.end()) return &it->second;
  CUtensorMap m;
  if (make_tmap(&m, dtype, rank, base, dims, strides_bytes, box, inner_bytes, elem_strides)) return nullptr;
  return &c.emplace(k, m).first->second;
}

}  // namespace slider
