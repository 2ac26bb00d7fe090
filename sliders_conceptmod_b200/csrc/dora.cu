// DoRA (reference conceptmod/textsliders/dora.py:122-163) on top of the fused base + low-rank kernel.
//
// The reference forms W' = (W + B.A) * c with ONE factor per INPUT channel, c_i = m_i / ||(W + B.A)[:, i]|| (norm over
// outputs and taps, detached; m = dora_scale), and returns org(x) + s * ops(x, W' - W): a second dense GEMM per
// wrapped module.  Because c scales input channels,
//     y = x . (W * v)^T + bias + (x . (A * s*c)^T) . B^T,        v = 1 - s + s*c,
// which is ONE launch of lora_gemm_kernel on the column-scaled operands W_eff = W * v and A_s = A * (s c).
//   dora_colnorm_kernel   per-CTA partial sums of ||(W + B.A)[:, i]||^2 over an output-channel slice (one pass over W)
//   dora_apply_kernel     c, v, 1/n from the partials (summed in slice order), W_eff and A_s            (second pass)
//   dora_finish_kernel    backward tail: dx = v * (gy.W) + s*c * ((gy.B).A) and dm_i = s/n_i * sum_p x[p,i] * u[p,i],
//                         u = gy.W + (gy.B).A being the SAME input-gradient GEMM the LoRA path runs (s = 1); the
//                         reference's dense weight gradient (out x in per layer) is never formed.
// All three are bandwidth-class kernels (weights once / activations once); reductions are deterministic.
#include "common.cuh"

namespace slider {

constexpr int DORA_VL = 32;     // channel-vector lanes per block
constexpr int DORA_RL = 8;      // row / output-channel lanes per block

template <typename T>
__global__ void __launch_bounds__(DORA_VL * DORA_RL)
dora_colnorm_kernel(float* __restrict__ partial, const T* __restrict__ W, const T* __restrict__ A,
                    const T* __restrict__ Bup, int cout, int taps, int cin, int rank, int o_per_cta) {
  constexpr int V = Elem<T>::kVec;
  __shared__ float red[DORA_RL][DORA_VL * V + 1];
  const int vl = threadIdx.x % DORA_VL, rl = threadIdx.x / DORA_VL;
  const int c0 = (blockIdx.x * DORA_VL + vl) * V;
  const int o0 = blockIdx.y * o_per_cta, o1 = min(cout, o0 + o_per_cta);
  float acc[V];
#pragma unroll
  for (int k = 0; k < V; ++k) acc[k] = 0.f;
  if (c0 < cin) {
    for (int t = 0; t < taps; ++t) {
      for (int o = o0 + rl; o < o1; o += DORA_RL) {
        float w[V];
        Elem<T>::load(W + ((size_t)o * taps + t) * cin + c0, w);
        for (int j = 0; j < rank; ++j) {
          const float b = Elem<T>::to_f(Bup[(size_t)o * rank + j]);
          float a[V];
          Elem<T>::load(A + ((size_t)j * taps + t) * cin + c0, a);
#pragma unroll
          for (int k = 0; k < V; ++k) w[k] = fmaf(b, a[k], w[k]);
        }
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = fmaf(w[k], w[k], acc[k]);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < V; ++k) red[rl][vl * V + k] = acc[k];
  __syncthreads();
  for (int i = threadIdx.x; i < DORA_VL * V; i += blockDim.x) {
    const int c = blockIdx.x * DORA_VL * V + i;
    if (c < cin) {
      float s = 0.f;
#pragma unroll
      for (int r = 0; r < DORA_RL; ++r) s += red[r][i];
      partial[(size_t)blockIdx.y * cin + c] = s;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(DORA_VL * DORA_RL)
dora_apply_kernel(T* __restrict__ W_eff, T* __restrict__ A_s, float* __restrict__ c_out, float* __restrict__ v_out,
                  float* __restrict__ ninv_out, const float* __restrict__ partial, int nslices, const T* __restrict__ W,
                  const T* __restrict__ A, const T* __restrict__ m, float s, int cout, int taps, int cin, int rank,
                  int o_per_cta) {
  constexpr int V = Elem<T>::kVec;
  const int vl = threadIdx.x % DORA_VL, rl = threadIdx.x / DORA_VL;
  const int c0 = (blockIdx.x * DORA_VL + vl) * V;
  if (c0 >= cin) return;
  float cc[V], vv[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    float n2 = 0.f;
    for (int p = 0; p < nslices; ++p) n2 += partial[(size_t)p * cin + c0 + k];     // slice order: deterministic
    const float ninv = rsqrtf(n2);
    cc[k] = Elem<T>::to_f(m[c0 + k]) * ninv;
    vv[k] = 1.f - s + s * cc[k];
    if (blockIdx.y == 0 && rl == 0) { c_out[c0 + k] = cc[k]; v_out[c0 + k] = vv[k]; ninv_out[c0 + k] = ninv; }
  }
  const int o0 = blockIdx.y * o_per_cta, o1 = min(cout, o0 + o_per_cta);
  for (int o = o0 + rl; o < o1; o += DORA_RL)
    for (int t = 0; t < taps; ++t) {
      float w[V];
      const size_t off = ((size_t)o * taps + t) * cin + c0;
      Elem<T>::load(W + off, w);
#pragma unroll
      for (int k = 0; k < V; ++k) w[k] *= vv[k];
      Elem<T>::store(W_eff + off, w);
    }
  if (blockIdx.y == 0)
    for (int jt = rl; jt < rank * taps; jt += DORA_RL) {
      float a[V];
      const size_t off = (size_t)jt * cin + c0;
      Elem<T>::load(A + off, a);
#pragma unroll
      for (int k = 0; k < V; ++k) a[k] *= s * cc[k];
      Elem<T>::store(A_s + off, a);
    }
}

// dx[p,i] = v_i * T1 + s c_i * T2,   u = T1 + T2;   T1 given (3x3: a second input-gradient GEMM without the low-rank
// term), or T2 = sum_j g[p,j] A[j,i] formed here (1x1: g = gy.B comes out of the fused backward launch).
// dm_i = s/n_i * sum_p x[p,i] * u[p,i]: per-CTA column sums -> workspace -> the last CTA of a column block adds them in
// row-chunk order.
template <typename T>
__global__ void __launch_bounds__(DORA_VL * DORA_RL)
dora_finish_kernel(T* __restrict__ dx, float* __restrict__ dm, const T* __restrict__ u, const T* __restrict__ t1,
                   const T* __restrict__ x, const float* __restrict__ g, const T* __restrict__ A,
                   const float* __restrict__ c, const float* __restrict__ v, const float* __restrict__ ninv, float s,
                   long long M, int cin, int rank, int rows_per_cta, int accumulate, float* __restrict__ partial,
                   unsigned* __restrict__ tickets) {
  constexpr int V = Elem<T>::kVec;
  __shared__ float red[DORA_RL][DORA_VL * V + 1];
  __shared__ int s_last;
  const int vl = threadIdx.x % DORA_VL, rl = threadIdx.x / DORA_VL;
  const int c0 = (blockIdx.x * DORA_VL + vl) * V;
  const long long p0 = (long long)blockIdx.y * rows_per_cta, p1 = min(M, p0 + rows_per_cta);
  float acc[V];
#pragma unroll
  for (int k = 0; k < V; ++k) acc[k] = 0.f;
  if (c0 < cin) {
    float cv[V], vv[V];
#pragma unroll
    for (int k = 0; k < V; ++k) { cv[k] = s * c[c0 + k]; vv[k] = v[c0 + k]; }
    for (long long p = p0 + rl; p < p1; p += DORA_RL) {
      const size_t off = (size_t)p * cin + c0;
      float uu[V], a1[V], a2[V];
      Elem<T>::load(u + off, uu);
      if (t1 != nullptr) {
        Elem<T>::load(t1 + off, a1);
#pragma unroll
        for (int k = 0; k < V; ++k) a2[k] = uu[k] - a1[k];
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) a2[k] = 0.f;
        for (int j = 0; j < rank; ++j) {
          const float gj = g[(size_t)p * rank + j];
          float a[V];
          Elem<T>::load(A + (size_t)j * cin + c0, a);
#pragma unroll
          for (int k = 0; k < V; ++k) a2[k] = fmaf(gj, a[k], a2[k]);
        }
#pragma unroll
        for (int k = 0; k < V; ++k) a1[k] = uu[k] - a2[k];
      }
      if (dx != nullptr) {
        float o[V];
#pragma unroll
        for (int k = 0; k < V; ++k) o[k] = vv[k] * a1[k] + cv[k] * a2[k];
        Elem<T>::store(dx + off, o);
      }
      if (dm != nullptr) {
        float xx[V];
        Elem<T>::load(x + off, xx);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = fmaf(xx[k], uu[k], acc[k]);
      }
    }
  }
  if (dm == nullptr) return;
#pragma unroll
  for (int k = 0; k < V; ++k) red[rl][vl * V + k] = acc[k];
  __syncthreads();
  const int nchunks = (int)gridDim.y;
  for (int i = threadIdx.x; i < DORA_VL * V; i += blockDim.x) {
    const int cc = blockIdx.x * DORA_VL * V + i;
    if (cc < cin) {
      float sum = 0.f;
#pragma unroll
      for (int r = 0; r < DORA_RL; ++r) sum += red[r][i];
      __stcg(partial + (size_t)blockIdx.y * cin + cc, sum);
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(tickets + blockIdx.x, 1u) == (unsigned)(nchunks - 1)) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int i = threadIdx.x; i < DORA_VL * V; i += blockDim.x) {
    const int cc = blockIdx.x * DORA_VL * V + i;
    if (cc < cin) {
      float sum = 0.f;
      for (int ch0 = 0; ch0 < nchunks; ch0 += 8) {       // chunk order, eight loads in flight
        float t[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) t[u] = ch0 + u < nchunks ? __ldcg(partial + (size_t)(ch0 + u) * cin + cc) : 0.f;
#pragma unroll
        for (int u = 0; u < 8; ++u) sum += t[u];
      }
      dm[cc] = (accumulate ? dm[cc] : 0.f) + s * ninv[cc] * sum;
    }
  }
  if (threadIdx.x == 0) tickets[blockIdx.x] = 0u;
}

constexpr size_t DORA_TICKET_BYTES = 4096;
constexpr size_t DORA_PARTIAL_BYTES = (size_t)2 << 20;

template <typename T>
static int dora_prepare_impl(const slider_layer_geom* g, T* W_eff, T* A_s, float* c, float* v, float* ninv, const T* W,
                             const T* A, const T* B, const T* m, float s, void* ws, size_t ws_bytes, cudaStream_t st) {
  constexpr int V = 16 / (int)sizeof(T);
  const int taps = g->kh * g->kw;
  const int nvec = (g->cin + V - 1) / V;
  const int gx = (nvec + DORA_VL - 1) / DORA_VL;
  int slices = (2 * sm_count() + gx - 1) / gx;
  const int max_slices = (int)(DORA_PARTIAL_BYTES / ((size_t)g->cin * sizeof(float)));
  if (slices > max_slices) slices = max_slices;
  if (slices > (g->cout + DORA_RL - 1) / DORA_RL) slices = (g->cout + DORA_RL - 1) / DORA_RL;
  if (slices < 1) slices = 1;
  const int o_per = (g->cout + slices - 1) / slices;
  slices = (g->cout + o_per - 1) / o_per;
  if (ws == nullptr || ws_bytes < DORA_TICKET_BYTES + DORA_PARTIAL_BYTES) {
    set_error("dora_prepare: workspace %zu < %zu bytes", ws_bytes, DORA_TICKET_BYTES + DORA_PARTIAL_BYTES);
    return SLIDER_ERR_WORKSPACE;
  }
  float* partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + DORA_TICKET_BYTES);
  dim3 grid(gx, slices);
  dora_colnorm_kernel<T><<<grid, DORA_VL * DORA_RL, 0, st>>>(partial, W, A, B, g->cout, taps, g->cin, g->rank, o_per);
  int rc = post_launch("dora_colnorm_kernel");
  if (rc) return rc;
  dora_apply_kernel<T><<<grid, DORA_VL * DORA_RL, 0, st>>>(W_eff, A_s, c, v, ninv, partial, slices, W, A, m, s, g->cout,
                                                           taps, g->cin, g->rank, o_per);
  return post_launch("dora_apply_kernel");
}

template <typename T>
static int dora_finish_impl(const slider_layer_geom* g, T* dx, float* dm, const T* u, const T* t1, const T* x,
                            const float* gb, const T* A, const float* c, const float* v, const float* ninv, float s,
                            int accumulate, void* ws, size_t ws_bytes, cudaStream_t st) {
  constexpr int V = 16 / (int)sizeof(T);
  const long long M = (long long)g->batch * g->height * g->width;
  const int nvec = (g->cin + V - 1) / V;
  const int gx = (nvec + DORA_VL - 1) / DORA_VL;
  if (ws == nullptr || ws_bytes < DORA_TICKET_BYTES + DORA_PARTIAL_BYTES || gx > (int)(DORA_TICKET_BYTES / 4)) {
    set_error("dora_finish: workspace %zu < %zu bytes", ws_bytes, DORA_TICKET_BYTES + DORA_PARTIAL_BYTES);
    return SLIDER_ERR_WORKSPACE;
  }
  long long chunks = (2 * sm_count() + gx - 1) / gx;
  const long long max_chunks = (long long)(DORA_PARTIAL_BYTES / ((size_t)g->cin * sizeof(float)));
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks > (M + DORA_RL - 1) / DORA_RL) chunks = (M + DORA_RL - 1) / DORA_RL;
  if (chunks < 1) chunks = 1;
  const int rows_per = (int)((M + chunks - 1) / chunks);
  chunks = (M + rows_per - 1) / rows_per;
  dim3 grid(gx, (unsigned)chunks);
  dora_finish_kernel<T><<<grid, DORA_VL * DORA_RL, 0, st>>>(
      dx, dm, u, t1, x, gb, A, c, v, ninv, s, M, g->cin, g->rank, rows_per, accumulate,
      reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + DORA_TICKET_BYTES), reinterpret_cast<unsigned*>(ws));
  return post_launch("dora_finish_kernel");
}

}  // namespace slider

using namespace slider;

extern "C" {

int slider_dora_prepare(const slider_layer_geom* g, void* W_eff, void* A_s, float* c, float* v, float* ninv,
                        const void* W, const void* A, const void* B, const void* dora_scale, float s, void* ws,
                        size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(g && W_eff && A_s && c && v && ninv && W && A && B && dora_scale, "dora_prepare: null pointer");
  SLIDER_CHECK_ARG(g->dtype == SLIDER_F32 || g->dtype == SLIDER_BF16, "dora_prepare: unsupported dtype %d", g->dtype);
  SLIDER_CHECK_ARG(g->kh == g->kw && (g->kh == 1 || g->kh == 3) && g->cin > 0 && g->cout > 0 && g->cin % 8 == 0,
                   "dora_prepare: bad geometry");
  SLIDER_CHECK_ARG(g->rank >= 1 && g->rank <= 16, "dora_prepare: rank %d not in 1..16", g->rank);
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    return dora_prepare_impl<__nv_bfloat16>(g, (__nv_bfloat16*)W_eff, (__nv_bfloat16*)A_s, c, v, ninv,
                                            (const __nv_bfloat16*)W, (const __nv_bfloat16*)A, (const __nv_bfloat16*)B,
                                            (const __nv_bfloat16*)dora_scale, s, ws, ws_bytes, st);
  return dora_prepare_impl<float>(g, (float*)W_eff, (float*)A_s, c, v, ninv, (const float*)W, (const float*)A,
                                  (const float*)B, (const float*)dora_scale, s, ws, ws_bytes, st);
}

int slider_dora_finish(const slider_layer_geom* g, void* dx, float* dm, const void* u, const void* t1, const void* x,
                       const float* g_buf, const void* A, const float* c, const float* v, const float* ninv, float s,
                       int accumulate, void* ws, size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(g && u && c && v && ninv && (dx || dm), "dora_finish: null pointer");
  SLIDER_CHECK_ARG(g->dtype == SLIDER_F32 || g->dtype == SLIDER_BF16, "dora_finish: unsupported dtype %d", g->dtype);
  SLIDER_CHECK_ARG(t1 != nullptr || (g_buf != nullptr && A != nullptr && g->kh == 1),
                   "dora_finish: needs T1, or (g, A) of a 1x1 layer");
  SLIDER_CHECK_ARG(dm == nullptr || x != nullptr, "dora_finish: dm needs x");
  SLIDER_CHECK_ARG(g->cin > 0 && g->cin % 8 == 0, "dora_finish: channels must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    return dora_finish_impl<__nv_bfloat16>(g, (__nv_bfloat16*)dx, dm, (const __nv_bfloat16*)u, (const __nv_bfloat16*)t1,
                                           (const __nv_bfloat16*)x, g_buf, (const __nv_bfloat16*)A, c, v, ninv, s,
                                           accumulate, ws, ws_bytes, st);
  return dora_finish_impl<float>(g, (float*)dx, dm, (const float*)u, (const float*)t1, (const float*)x, g_buf,
                                 (const float*)A, c, v, ninv, s, accumulate, ws, ws_bytes, st);
}

}  // extern "C"
