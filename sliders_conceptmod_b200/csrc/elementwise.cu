// HBM-bound kernels of the slider path: guidance loss (+gradient), CFG + scheduler
// step, flat AdamW, gradient-norm clip coefficient.  One pass over the data each,
// 16-byte vector accesses, grid sized in multiples of the SM count.
#include <mutex>
#include "common.cuh"

namespace slider {

constexpr int kThreads = 256;
constexpr int kMaxBlocks = 148 * 8;  // partial-sum slots in the reduction workspace
constexpr size_t kReduceWsBytes = (kMaxBlocks + 4) * sizeof(float);

static inline int grid_for(int64_t n_vec, int per_thread) {
  int64_t need = (n_vec + (int64_t)kThreads * per_thread - 1) / ((int64_t)kThreads * per_thread);
  int64_t cap = (int64_t)sm_count() * 8;
  if (cap > kMaxBlocks) cap = kMaxBlocks;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

// Block-level sum -> ws[blockIdx]; last block to finish adds the partials in index
// order (deterministic), applies `finish`, and re-zeroes the ticket.
template <typename Finish>
__device__ inline void block_reduce_finish(float local, float* ws, float* out, Finish finish) {
  __shared__ float s_warp[kThreads / 32];
  __shared__ bool s_last;
  local = warp_sum(local);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) s_warp[warp] = local;
  __syncthreads();
  unsigned* ticket = reinterpret_cast<unsigned*>(ws + kMaxBlocks);
  if (threadIdx.x == 0) {
    float b = 0.f;
#pragma unroll
    for (int i = 0; i < kThreads / 32; ++i) b += s_warp[i];
    ws[blockIdx.x] = b;
    __threadfence();
    unsigned t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && warp == 0) {
    __threadfence();
    float acc = 0.f;
    for (int i = lane; i < (int)gridDim.x; i += 32) acc += __ldcg(ws + i);
    acc = warp_sum(acc);
    if (lane == 0) {
      out[0] = finish(acc);
      *ticket = 0u;
    }
  }
}

// ---- guidance loss -----------------------------------------------------------
template <typename T, bool kGuided, bool kGrad>
__global__ void __launch_bounds__(kThreads)
guidance_loss_kernel(float* __restrict__ loss_out, T* __restrict__ grad, const T* __restrict__ t,
                     const T* __restrict__ p, const T* __restrict__ n, const T* __restrict__ q,
                     int64_t n_elem, float sg, float* __restrict__ ws) {
  constexpr int V = Elem<T>::kVec;
  const int64_t n_vec = n_elem / V;
  const float gscale = 2.0f / (float)n_elem;
  float acc = 0.f;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_vec; i += stride) {
    float vt[V], vn[V], vp[V], vq[V], vg[V];
    Elem<T>::load(t + i * V, vt);
    Elem<T>::load(n + i * V, vn);
    if (kGuided) {
      Elem<T>::load(p + i * V, vp);
      Elem<T>::load(q + i * V, vq);
    }
#pragma unroll
    for (int k = 0; k < V; ++k) {
      float tgt = kGuided ? fmaf(sg, vp[k] - vq[k], vn[k]) : vn[k];
      float d = vt[k] - tgt;
      acc = fmaf(d, d, acc);
      vg[k] = d * gscale;
    }
    if (kGrad) Elem<T>::store(grad + i * V, vg);
  }
  // scalar tail (n_elem not a multiple of the vector width)
  if (blockIdx.x == 0) {
    for (int64_t i = n_vec * V + threadIdx.x; i < n_elem; i += kThreads) {
      float tgt = Elem<T>::to_f(n[i]);
      if (kGuided) tgt = fmaf(sg, Elem<T>::to_f(p[i]) - Elem<T>::to_f(q[i]), tgt);
      float d = Elem<T>::to_f(t[i]) - tgt;
      acc = fmaf(d, d, acc);
      if (kGrad) grad[i] = Elem<T>::from_f(d * gscale);
    }
  }
  const float inv_n = 1.0f / (float)n_elem;
  block_reduce_finish(acc, ws, loss_out, [inv_n](float s) { return s * inv_n; });
}

template <typename T>
static int launch_loss(float* loss_out, void* grad, const void* t, const void* p, const void* n,
                       const void* q, int64_t n_elem, float sg, float* ws, cudaStream_t st) {
  const int grid = grid_for(n_elem / Elem<T>::kVec, 2);
  const bool guided = p != nullptr, g = grad != nullptr;
#define LAUNCH(G, D)                                                                         \
  guidance_loss_kernel<T, G, D><<<grid, kThreads, 0, st>>>(loss_out, (T*)grad, (const T*)t, \
                                                           (const T*)p, (const T*)n,        \
                                                           (const T*)q, n_elem, sg, ws)
  if (guided && g) LAUNCH(true, true);
  else if (guided) LAUNCH(true, false);
  else if (g) LAUNCH(false, true);
  else LAUNCH(false, false);
#undef LAUNCH
  return post_launch("guidance_loss_kernel");
}

// ---- CFG + scheduler step ------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads)
cfg_step_kernel(T* __restrict__ out, const T* __restrict__ x, const T* __restrict__ eu,
                const T* __restrict__ ec, const T* __restrict__ nz, int64_t n_elem, float g,
                float cx, float ce, float cn, const float* __restrict__ dev_coef) {
  pdl_trigger();
  pdl_wait();
  constexpr int V = Elem<T>::kVec;
  if (dev_coef != nullptr) {
    g = __ldg(dev_coef); cx = __ldg(dev_coef + 1); ce = __ldg(dev_coef + 2); cn = __ldg(dev_coef + 3);
  }
  const bool has_x = x != nullptr, has_e = eu != nullptr, has_n = nz != nullptr;
  if (!has_x) { cx = 0.f; ce = 1.f; }
  const int64_t n_vec = n_elem / V;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_vec; i += stride) {
    float vx[V], vu[V], vc[V], vn[V], vo[V];
    if (has_x) Elem<T>::load(x + i * V, vx);
    if (has_e) { Elem<T>::load(eu + i * V, vu); Elem<T>::load(ec + i * V, vc); }
    if (has_n) Elem<T>::load(nz + i * V, vn);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      float o = has_x ? cx * vx[k] : 0.f;
      if (has_e) o = fmaf(ce, fmaf(g, vc[k] - vu[k], vu[k]), o);
      if (has_n) o = fmaf(cn, vn[k], o);
      vo[k] = o;
    }
    Elem<T>::store(out + i * V, vo);
  }
  if (blockIdx.x == 0) {
    for (int64_t i = n_vec * V + threadIdx.x; i < n_elem; i += kThreads) {
      float o = has_x ? cx * Elem<T>::to_f(x[i]) : 0.f;
      if (has_e) {
        float u = Elem<T>::to_f(eu[i]), c = Elem<T>::to_f(ec[i]);
        o = fmaf(ce, fmaf(g, c - u, u), o);
      }
      if (has_n) o = fmaf(cn, Elem<T>::to_f(nz[i]), o);
      out[i] = Elem<T>::from_f(o);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
scale_by_scalar_kernel(T* __restrict__ x, const float* __restrict__ sc, int64_t n_elem) {
  const float s = __ldg(sc);
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_elem;
       i += (int64_t)gridDim.x * kThreads)
    x[i] = Elem<T>::from_f(Elem<T>::to_f(x[i]) * s);
}

// ---- flat AdamW ----------------------------------------------------------------
__device__ __forceinline__ float adamw_one(float& pi, float g, float& mi, float& vi, float b1, float b2, float eps,
                                           float decay, float step, float rsqrt_bc2) {
  mi = fmaf(b1, mi, (1.f - b1) * g);
  vi = fmaf(b2, vi, (1.f - b2) * g * g);
  pi *= decay;
  pi -= step * (mi / (sqrtf(vi) * rsqrt_bc2 + eps));
  return pi;
}

// 4 elements per thread: 16-byte accesses to grad / m / v (fp32) and 8- or 16-byte accesses to p.
template <typename T>
__global__ void __launch_bounds__(kThreads)
adamw_flat_kernel(T* __restrict__ p, const float* __restrict__ grad, float* __restrict__ m,
                  float* __restrict__ v, int64_t n_elem, float lr, float b1, float b2, float eps,
                  float wd, float bc1, float rsqrt_bc2, const float* __restrict__ gs_dev) {
  const float gs = gs_dev ? __ldg(gs_dev) : 1.f;
  const float step = lr / bc1, decay = 1.f - lr * wd;
  const int64_t n4 = n_elem / 4;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kThreads) {
    const float4 g4 = __ldcs(reinterpret_cast<const float4*>(grad) + i);
    float4 m4 = reinterpret_cast<float4*>(m)[i], v4 = reinterpret_cast<float4*>(v)[i];
    float pv[4];
    if (sizeof(T) == 2) {
      const uint2 t = reinterpret_cast<const uint2*>(p)[i];
      pv[0] = __uint_as_float(t.x << 16); pv[1] = __uint_as_float(t.x & 0xffff0000u);
      pv[2] = __uint_as_float(t.y << 16); pv[3] = __uint_as_float(t.y & 0xffff0000u);
    } else {
      const float4 t = reinterpret_cast<const float4*>(p)[i];
      pv[0] = t.x; pv[1] = t.y; pv[2] = t.z; pv[3] = t.w;
    }
    adamw_one(pv[0], g4.x * gs, m4.x, v4.x, b1, b2, eps, decay, step, rsqrt_bc2);
    adamw_one(pv[1], g4.y * gs, m4.y, v4.y, b1, b2, eps, decay, step, rsqrt_bc2);
    adamw_one(pv[2], g4.z * gs, m4.z, v4.z, b1, b2, eps, decay, step, rsqrt_bc2);
    adamw_one(pv[3], g4.w * gs, m4.w, v4.w, b1, b2, eps, decay, step, rsqrt_bc2);
    reinterpret_cast<float4*>(m)[i] = m4;
    reinterpret_cast<float4*>(v)[i] = v4;
    if (sizeof(T) == 2) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pv[0], pv[1]), hi = __floats2bfloat162_rn(pv[2], pv[3]);
      reinterpret_cast<uint2*>(p)[i] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
    } else {
      reinterpret_cast<float4*>(p)[i] = make_float4(pv[0], pv[1], pv[2], pv[3]);
    }
  }
  if (blockIdx.x == 0) {
    for (int64_t i = n4 * 4 + threadIdx.x; i < n_elem; i += kThreads) {
      float pi = Elem<T>::to_f(p[i]), mi = m[i], vi = v[i];
      adamw_one(pi, grad[i] * gs, mi, vi, b1, b2, eps, decay, step, rsqrt_bc2);
      m[i] = mi; v[i] = vi;
      p[i] = Elem<T>::from_f(pi);
    }
  }
}

__global__ void __launch_bounds__(kThreads)
clip_coef_kernel(float* __restrict__ out, const float* __restrict__ grad, int64_t n_elem,
                 float max_norm, float pre, float* __restrict__ ws) {
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n_elem;
       i += (int64_t)gridDim.x * kThreads) {
    float g = grad[i];
    acc = fmaf(g, g, acc);
  }
  block_reduce_finish(acc, ws, out, [max_norm, pre](float s) {
    float norm = sqrtf(s) * pre;
    float c = max_norm / (norm + 1e-6f);
    return (c < 1.f ? c : 1.f) * pre;
  });
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- allocation-free guidance loss: scratch in a file-scope device array --------------------------
// Reduction scratch of slider_guidance_loss_noalloc: kMaxBlocks partial-sum slots + the ticket word, held by every
// GPU that loads the library until exit (4752 bytes).  In host code `g_noalloc_scratch` is only a stand-in for the
// symbol: its device address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) float g_noalloc_scratch[kMaxBlocks + 4];
constexpr int kNoAllocMaxDevices = 64;

static int noalloc_scratch_ptr(float** out) {
  static std::mutex mu;
  static float* per_device[kNoAllocMaxDevices] = {nullptr};
  int dev = -1;
  SLIDER_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) {
    set_error("guidance_loss_noalloc: device ordinal %d outside 0..%d", dev, kNoAllocMaxDevices - 1);
    return SLIDER_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (per_device[dev] == nullptr) {
    float* ptr = nullptr;
    SLIDER_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    if (ptr == nullptr) {
      set_error("guidance_loss_noalloc: cudaGetSymbolAddress returned a null device address");
      return SLIDER_ERR_CUDA;
    }
    per_device[dev] = ptr;
  }
  *out = per_device[dev];
  return SLIDER_OK;
}

}  // namespace slider

using namespace slider;

extern "C" {

size_t slider_loss_workspace_bytes(void) { return kReduceWsBytes; }

int slider_guidance_loss(float* loss_out, void* grad, const void* t, const void* p, const void* n,
                         const void* q, int64_t n_elem, float g, int sign, int dtype, void* ws,
                         size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(loss_out && t && n && ws, "guidance_loss: null pointer");
  SLIDER_CHECK_ARG((p == nullptr) == (q == nullptr), "guidance_loss: p and q must both be set or both NULL");
  SLIDER_CHECK_ARG(n_elem > 0, "guidance_loss: n_elem must be positive (got %lld)", (long long)n_elem);
  SLIDER_CHECK_ARG(sign == 1 || sign == -1, "guidance_loss: sign must be +1 or -1");
  SLIDER_CHECK_ARG(aligned16(t) && aligned16(n) && aligned16(p) && aligned16(q) && aligned16(grad),
                   "guidance_loss: tensors must be 16-byte aligned");
  if (ws_bytes < kReduceWsBytes) {
    set_error("guidance_loss: workspace %zu < %zu bytes", ws_bytes, kReduceWsBytes);
    return SLIDER_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const float sg = (float)sign * g;
  if (dtype == SLIDER_F32) return launch_loss<float>(loss_out, grad, t, p, n, q, n_elem, sg, (float*)ws, st);
  if (dtype == SLIDER_BF16)
    return launch_loss<__nv_bfloat16>(loss_out, grad, t, p, n, q, n_elem, sg, (float*)ws, st);
  set_error("guidance_loss: unsupported dtype %d", dtype);
  return SLIDER_ERR_UNSUPPORTED;
}

int64_t slider_guidance_loss_noalloc_max_elems(void) { return (int64_t)SLIDER_NOALLOC_MAX_ELEMS; }

int slider_guidance_loss_noalloc(float* loss_out, void* grad, const void* t, const void* p, const void* n,
                                 const void* q, int64_t n_elem, float g, int sign, int dtype, void* stream) {
  SLIDER_CHECK_ARG(loss_out && t && n, "guidance_loss_noalloc: null pointer");
  SLIDER_CHECK_ARG((p == nullptr) == (q == nullptr), "guidance_loss_noalloc: p and q must both be set or both NULL");
  SLIDER_CHECK_ARG(n_elem > 0, "guidance_loss_noalloc: n_elem must be positive (got %lld)", (long long)n_elem);
  if (n_elem > (int64_t)SLIDER_NOALLOC_MAX_ELEMS) {
    set_error("guidance_loss_noalloc: n_elem %lld exceeds the built-in scratch's limit of %lld elements; use "
              "slider_guidance_loss with a caller workspace", (long long)n_elem, (long long)SLIDER_NOALLOC_MAX_ELEMS);
    return SLIDER_ERR_UNSUPPORTED;
  }
  SLIDER_CHECK_ARG(sign == 1 || sign == -1, "guidance_loss_noalloc: sign must be +1 or -1");
  SLIDER_CHECK_ARG(aligned16(t) && aligned16(n) && aligned16(p) && aligned16(q) && aligned16(grad),
                   "guidance_loss_noalloc: tensors must be 16-byte aligned");
  if (dtype != SLIDER_F32 && dtype != SLIDER_BF16) {
    set_error("guidance_loss_noalloc: unsupported dtype %d", dtype);
    return SLIDER_ERR_UNSUPPORTED;
  }
  float* ws = nullptr;
  int rc = noalloc_scratch_ptr(&ws);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's contents: the ticket word must start at zero (the partial slots are
  // overwritten before they are read).  A stream-ordered 16-byte memset; nothing is allocated.
  SLIDER_CUDA(cudaMemsetAsync(ws + kMaxBlocks, 0, 4 * sizeof(float), st));
  const float sg = (float)sign * g;
  if (dtype == SLIDER_F32) return launch_loss<float>(loss_out, grad, t, p, n, q, n_elem, sg, ws, st);
  return launch_loss<__nv_bfloat16>(loss_out, grad, t, p, n, q, n_elem, sg, ws, st);
}

int slider_cfg_step(void* out, const void* x, const void* eps_u, const void* eps_c, const void* noise,
                    int64_t n_elem, float g, float c_x, float c_eps, float c_noise,
                    const float* dev_coef, int dtype, void* stream) {
  SLIDER_CHECK_ARG(out, "cfg_step: out is NULL");
  SLIDER_CHECK_ARG((eps_u == nullptr) == (eps_c == nullptr), "cfg_step: eps_u and eps_c must both be set or both NULL");
  SLIDER_CHECK_ARG(x || eps_u, "cfg_step: need x or eps");
  SLIDER_CHECK_ARG(n_elem > 0, "cfg_step: n_elem must be positive");
  SLIDER_CHECK_ARG(aligned16(out) && aligned16(x) && aligned16(eps_u) && aligned16(eps_c) && aligned16(noise),
                   "cfg_step: tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLIDER_F32) {
    const int grid = grid_for(n_elem / 4, 2);
    launch_pdl(cfg_step_kernel<float>, dim3(grid), dim3(kThreads), 0, st, (float*)out, (const float*)x,
               (const float*)eps_u, (const float*)eps_c, (const float*)noise, n_elem, g, c_x, c_eps, c_noise, dev_coef);
  } else if (dtype == SLIDER_BF16) {
    typedef __nv_bfloat16 T;
    const int grid = grid_for(n_elem / 8, 2);
    launch_pdl(cfg_step_kernel<T>, dim3(grid), dim3(kThreads), 0, st, (T*)out, (const T*)x, (const T*)eps_u,
               (const T*)eps_c, (const T*)noise, n_elem, g, c_x, c_eps, c_noise, dev_coef);
  } else if (dtype == SLIDER_F16) {
    typedef __half T;
    const int grid = grid_for(n_elem / 8, 2);
    launch_pdl(cfg_step_kernel<T>, dim3(grid), dim3(kThreads), 0, st, (T*)out, (const T*)x, (const T*)eps_u,
               (const T*)eps_c, (const T*)noise, n_elem, g, c_x, c_eps, c_noise, dev_coef);
  } else {
    set_error("cfg_step: unsupported dtype %d", dtype);
    return SLIDER_ERR_UNSUPPORTED;
  }
  return post_launch("cfg_step_kernel");
}

int slider_scale_by_device_scalar(void* x, const float* dev_scalar, int64_t n_elem, int dtype, void* stream) {
  SLIDER_CHECK_ARG(x && dev_scalar && n_elem > 0, "scale_by_device_scalar: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for(n_elem, 4);
  if (dtype == SLIDER_F32) scale_by_scalar_kernel<float><<<grid, kThreads, 0, st>>>((float*)x, dev_scalar, n_elem);
  else if (dtype == SLIDER_BF16)
    scale_by_scalar_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>((__nv_bfloat16*)x, dev_scalar, n_elem);
  else { set_error("scale_by_device_scalar: unsupported dtype %d", dtype); return SLIDER_ERR_UNSUPPORTED; }
  return post_launch("scale_by_scalar_kernel");
}

int slider_adamw_flat(void* p, const float* grad, float* m, float* v, int64_t n_elem, float lr, float beta1,
                      float beta2, float eps, float weight_decay, int step, const float* grad_scale_dev,
                      int dtype, void* stream) {
  SLIDER_CHECK_ARG(p && grad && m && v && n_elem > 0 && step >= 1, "adamw_flat: bad argument");
  SLIDER_CHECK_ARG(aligned16(p) && aligned16(grad) && aligned16(m) && aligned16(v), "adamw_flat: buffers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float rsqrt_bc2 = 1.f / sqrtf(1.f - powf(beta2, (float)step));
  const int grid = grid_for(n_elem / 4 + 1, 2);
  if (dtype == SLIDER_F32)
    adamw_flat_kernel<float><<<grid, kThreads, 0, st>>>((float*)p, grad, m, v, n_elem, lr, beta1, beta2, eps,
                                                        weight_decay, bc1, rsqrt_bc2, grad_scale_dev);
  else if (dtype == SLIDER_BF16)
    adamw_flat_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>((__nv_bfloat16*)p, grad, m, v, n_elem, lr, beta1,
                                                                beta2, eps, weight_decay, bc1, rsqrt_bc2,
                                                                grad_scale_dev);
  else { set_error("adamw_flat: unsupported dtype %d", dtype); return SLIDER_ERR_UNSUPPORTED; }
  return post_launch("adamw_flat_kernel");
}

int slider_clip_coef(float* out, const float* grad, int64_t n_elem, float max_norm, float pre_scale, void* ws,
                     size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(out && grad && ws && n_elem > 0, "clip_coef: bad argument");
  if (ws_bytes < kReduceWsBytes) { set_error("clip_coef: workspace too small"); return SLIDER_ERR_WORKSPACE; }
  const int grid = grid_for(n_elem, 4);
  clip_coef_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(out, grad, n_elem, max_norm, pre_scale, (float*)ws);
  return post_launch("clip_coef_kernel");
}

}  // extern "C"
