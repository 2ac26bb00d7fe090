// Shared host/device helpers for libslider_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <atomic>

#include "../../include/slider_b200.h"

namespace slider {

// ---- error plumbing ---------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;

#define SLIDER_CHECK_ARG(cond, ...)                  \
  do {                                               \
    if (!(cond)) {                                   \
      ::slider::set_error(__VA_ARGS__);              \
      return SLIDER_ERR_BAD_ARG;                     \
    }                                                \
  } while (0)

#define SLIDER_CUDA(call)                                                     \
  do {                                                                        \
    cudaError_t e__ = (call);                                                 \
    if (e__ != cudaSuccess) {                                                 \
      ::slider::set_error("%s failed: %s (%s:%d)", #call,                     \
                          cudaGetErrorString(e__), __FILE__, __LINE__);       \
      return SLIDER_ERR_CUDA;                                                 \
    }                                                                         \
  } while (0)

inline int post_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("launch of %s failed: %s", what, cudaGetErrorString(e));
    return SLIDER_ERR_CUDA;
  }
  return SLIDER_OK;
}

int sm_count();

// ---- programmatic dependent launch (PDL) -----------------------------------------------------
// Every forward-path kernel starts with pdl_trigger() + pdl_wait(): the next kernel in the stream may be made
// resident while this one still runs (its launch latency and prologue overlap our tail), and nothing of ours
// touches global memory before the previous kernel has completed and flushed.  Both are no-ops for a launch
// without the attribute, which is set unless SLIDER_PDL=0 is in the environment (A/B: profiles/r01_ab_pdl_v2.txt).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---- element access ---------------------------------------------------------
template <typename T> struct Elem;
template <> struct Elem<float> {
  static constexpr int kVec = 4;  // elements per 16-byte access
  __device__ static inline void load(const float* p, float (&v)[4]) {
    float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ static inline void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
  __device__ static inline float to_f(float x) { return x; }
  __device__ static inline float from_f(float x) { return x; }
  // 16-bit pair helpers: unused for fp32 (present so that dtype-generic code compiles)
  __device__ static inline uint32_t pack2(float a, float) { return __float_as_uint(a); }
  __device__ static inline void unpack2(uint32_t w, float& a, float& b) { a = __uint_as_float(w); b = 0.f; }
  __device__ static inline float round(float x) { return x; }
};
template <> struct Elem<__nv_bfloat16> {
  static constexpr int kVec = 8;
  __device__ static inline void load(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static inline void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  __device__ static inline float to_f(__nv_bfloat16 x) { return __bfloat162float(x); }
  __device__ static inline __nv_bfloat16 from_f(float x) { return __float2bfloat16_rn(x); }
  __device__ static inline uint32_t pack2(float a, float b) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
  __device__ static inline void unpack2(uint32_t w, float& a, float& b) {
    a = __uint_as_float(w << 16); b = __uint_as_float(w & 0xffff0000u);
  }
  __device__ static inline float round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
};
template <> struct Elem<__half> {
  static constexpr int kVec = 8;
  __device__ static inline uint32_t pack2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
  __device__ static inline void unpack2(uint32_t w, float& a, float& b) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
    a = f.x; b = f.y;
  }
  __device__ static inline float round(float x) { return __half2float(__float2half_rn(x)); }
  __device__ static inline void load(const __half* p, float (&v)[8]) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) unpack2(w[i], v[2 * i], v[2 * i + 1]);
  }
  __device__ static inline void store(__half* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = pack2(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  __device__ static inline float to_f(__half x) { return __half2float(x); }
  __device__ static inline __half from_f(float x) { return __float2half_rn(x); }
};

// erf for bf16 outputs: Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, three orders below a bf16 ulp) on the fast
// exp / reciprocal units; the exact erff costs ~40 instructions per element and made the kernel ALU-bound
// (0.52 of HBM at 16384 x 5120).  fp32 tensors keep erff.
// The reciprocal and the exponential are the single-instruction approximations (MUFU.RCP / MUFU.EX2, 1-2 ulp, flush
// to zero): `__frcp_rn` and `__expf` each compile to a range test with a branch to a slow path around the MUFU, which
// fences the scheduler per element — the fused GEMM + GEGLU epilogue ran 40 evaluations per thread back to back with
// no overlap between them, ~10 000 cycles per 128 x 160 tile against ~1 650 cycles of MMAs (r02 ncu: 19 % tensor pipe).
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float erf_bf16_grade(float x) {
  const float ax = fabsf(x);
  const float t = rcp_approx(fmaf(0.3275911f, ax, 1.f));      // argument >= 1: no range fix-up needed
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float r = fmaf(-p * t, ex2_approx(ax * ax * -1.4426950408889634f), 1.f);      // underflow flushes to 0: r = 1
  return copysignf(r, x);
}

// gelu for 16-bit outputs without the erf detour: with x = |g| / sqrt(2) and A&S 7.1.26 erf(x) = 1 - poly(t) exp(-x^2),
//   gelu(g) = 0.5 g (1 + sign(g) erf(x)) = relu(g) - |g| * (0.5 poly(t)) * exp(-g^2 / 2),   t = 1 / (1 + 0.3275911 x)
// 12 FP32 instructions + 2 MUFU per value (the erf form took 19 + 2), and no cancellation for negative g.
__device__ __forceinline__ float gelu_bf16_grade(float g) {
  const float ag = fabsf(g);
  const float t = rcp_approx(fmaf(ag, 0.3275911f * 0.70710678118654752f, 1.f));
  const float e = ex2_approx(g * g * (-0.5f * 1.4426950408889634f));
  float p = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);
  p = fmaf(p, t, 0.5f * 1.421413741f);
  p = fmaf(p, t, 0.5f * -0.284496736f);
  p = fmaf(p, t, 0.5f * 0.254829592f);
  return fmaf(-ag, p * t * e, fmaxf(g, 0.f));
}

__device__ inline float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace slider
