// Row-wise elementwise / normalisation kernels of the UNet's transformer blocks: LayerNorm ahead of
// the LoRA-wrapped attn1 projections and the GEGLU gate of the feed-forward.  HBM-bound, 16-byte accesses.
#include "common.cuh"

namespace slider {

template <typename T> struct RowVec {
  static constexpr int V = Elem<T>::kVec;
};

// ---- LayerNorm over the last dimension: one warp per row, the row in registers ---------------------------------------
// Rows of up to 32 * MAXV 16-byte vectors (every row of the UNets of this path): gamma / beta are frozen parameters and
// are fetched BEFORE the programmatic-dependent-launch wait, x is read once, and the two reductions and the affine
// pass work on registers — one exposed global-load latency instead of three.  Same summation order as the streaming
// form below, so the results are bit-identical.
template <typename T> struct VecIO;
template <> struct VecIO<__nv_bfloat16> {
  __device__ static inline void unpack(const uint4& t, float (&v)[8]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { v[2 * i] = __uint_as_float(w[i] << 16); v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
  }
  __device__ static inline uint4 pack(const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <> struct VecIO<__half> {
  __device__ static inline void unpack(const uint4& t, float (&v)[8]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) Elem<__half>::unpack2(w[i], v[2 * i], v[2 * i + 1]);
  }
  __device__ static inline uint4 pack(const float (&v)[8]) {
    return make_uint4(Elem<__half>::pack2(v[0], v[1]), Elem<__half>::pack2(v[2], v[3]), Elem<__half>::pack2(v[4], v[5]),
                      Elem<__half>::pack2(v[6], v[7]));
  }
};
template <> struct VecIO<float> {
  __device__ static inline void unpack(const uint4& t, float (&v)[4]) {
    v[0] = __uint_as_float(t.x); v[1] = __uint_as_float(t.y); v[2] = __uint_as_float(t.z); v[3] = __uint_as_float(t.w);
  }
  __device__ static inline uint4 pack(const float (&v)[4]) {
    return make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3]));
  }
};

template <typename T, int MAXV>
__global__ void __launch_bounds__(256)
layernorm_kernel(T* __restrict__ y, const T* __restrict__ x, const T* __restrict__ gamma,
                 const T* __restrict__ beta, int64_t rows, int C, float eps, float* __restrict__ stats) {
  pdl_trigger();
  constexpr int V = Elem<T>::kVec;
  const int lane = threadIdx.x & 31;
  const int nv = C / V;
  uint4 gp[MAXV], bp[MAXV], xp[MAXV];
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int v = lane + 32 * j;
    gp[j] = make_uint4(0, 0, 0, 0); bp[j] = make_uint4(0, 0, 0, 0);
    if (v < nv) {
      if (gamma) gp[j] = reinterpret_cast<const uint4*>(gamma)[v];
      if (beta) bp[j] = reinterpret_cast<const uint4*>(beta)[v];
    }
  }
  pdl_wait();
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * C);
  uint4* yr = reinterpret_cast<uint4*>(y + row * C);
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int v = lane + 32 * j;
    xp[j] = make_uint4(0, 0, 0, 0);
    if (v < nv) xp[j] = xr[v];
  }
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    if (lane + 32 * j < nv) {
      float xv[V];
      VecIO<T>::unpack(xp[j], xv);
#pragma unroll
      for (int k = 0; k < V; ++k) sum += xv[k];
    }
  }
  const float mean = warp_sum(sum) / (float)C;
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    if (lane + 32 * j < nv) {
      float xv[V];
      VecIO<T>::unpack(xp[j], xv);
#pragma unroll
      for (int k = 0; k < V; ++k) { const float d = xv[k] - mean; sq = fmaf(d, d, sq); }
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / (float)C + eps);
  if (stats != nullptr && lane == 0) { stats[2 * row] = mean; stats[2 * row + 1] = rstd; }
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int v = lane + 32 * j;
    if (v < nv) {
      float xv[V], gv[V], bv[V], o[V];
      VecIO<T>::unpack(xp[j], xv);
      VecIO<T>::unpack(gp[j], gv);
      VecIO<T>::unpack(bp[j], bv);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        float u = (xv[k] - mean) * rstd;
        if (gamma) u *= gv[k];
        if (beta) u += bv[k];
        o[k] = u;
      }
      yr[v] = VecIO<T>::pack(o);
    }
  }
}

// ---- LayerNorm over the last dimension: one warp per row, rows of any length (three passes, the later ones hit L1) ----
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_stream_kernel(T* __restrict__ y, const T* __restrict__ x, const T* __restrict__ gamma,
                 const T* __restrict__ beta, int64_t rows, int C, float eps, float* __restrict__ stats) {
  pdl_trigger();
  pdl_wait();
  constexpr int V = Elem<T>::kVec;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const T* xr = x + row * C;
  T* yr = y + row * C;
  const int nv = C / V;
  float sum = 0.f;
  for (int v = lane; v < nv; v += 32) {
    float xv[V];
    Elem<T>::load(xr + v * V, xv);
#pragma unroll
    for (int k = 0; k < V; ++k) sum += xv[k];
  }
  const float mean = warp_sum(sum) / (float)C;
  float sq = 0.f;
  for (int v = lane; v < nv; v += 32) {     // second pass hits L1: centred sum of squares (no cancellation)
    float xv[V];
    Elem<T>::load(xr + v * V, xv);
#pragma unroll
    for (int k = 0; k < V; ++k) { const float d = xv[k] - mean; sq = fmaf(d, d, sq); }
  }
  const float rstd = rsqrtf(warp_sum(sq) / (float)C + eps);
  if (stats != nullptr && lane == 0) { stats[2 * row] = mean; stats[2 * row + 1] = rstd; }
  for (int v = lane; v < nv; v += 32) {
    float xv[V], gv[V], bv[V], o[V];
    Elem<T>::load(xr + v * V, xv);
    if (gamma) Elem<T>::load(gamma + v * V, gv);
    if (beta) Elem<T>::load(beta + v * V, bv);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      float u = (xv[k] - mean) * rstd;
      if (gamma) u *= gv[k];
      if (beta) u += bv[k];
      o[k] = u;
    }
    Elem<T>::store(yr + v * V, o);
  }
}

// ---- GEGLU: y[r, c] = x[r, c] * gelu(x[r, D + c]) (erf GELU) -------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
geglu_kernel(T* __restrict__ y, const T* __restrict__ x, int64_t rows, int D) {
  pdl_trigger();
  pdl_wait();
  constexpr int V = Elem<T>::kVec;
  const int nv = D / V;
  // blockIdx.x walks the row's vectors, blockIdx.y strides over rows: no 64-bit division per element
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  for (int64_t r = blockIdx.y; r < rows; r += gridDim.y) {
    float a[V], g[V], o[V];
    Elem<T>::load(x + r * 2 * D + v * V, a);
    Elem<T>::load(x + r * 2 * D + D + v * V, g);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      o[k] = a[k] * (sizeof(T) == 2 ? gelu_bf16_grade(g[k]) : 0.5f * g[k] * (1.f + erff(g[k] * 0.70710678118654752f)));
    }
    Elem<T>::store(y + r * D + v * V, o);
  }
}


// ---- LayerNorm backward for FROZEN gamma / beta (unet.requires_grad_(False), train_lora.py:69): dx only ----------------
//   ghat = gy * gamma,  xhat = (x - mean) * rstd,  dx = rstd * (ghat - mean(ghat) - xhat * mean(ghat * xhat))
// One warp per row, two passes over the row (the second one hits L1); mean / rstd come from the training forward.
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(T* __restrict__ dx, const T* __restrict__ gy, const T* __restrict__ x, const T* __restrict__ gamma,
                     const float* __restrict__ stats, int64_t rows, int C) {
  constexpr int V = Elem<T>::kVec;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const T* xr = x + row * C;
  const T* gr = gy + row * C;
  T* dr = dx + row * C;
  const int nv = C / V;
  const float mean = stats[2 * row], rstd = stats[2 * row + 1];
  float s1 = 0.f, s2 = 0.f;
  for (int v = lane; v < nv; v += 32) {
    float xv[V], gv[V], wv[V];
    Elem<T>::load(xr + v * V, xv);
    Elem<T>::load(gr + v * V, gv);
    if (gamma) Elem<T>::load(gamma + v * V, wv);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const float gh = gamma ? gv[k] * wv[k] : gv[k];
      s1 += gh;
      s2 = fmaf(gh, (xv[k] - mean) * rstd, s2);
    }
  }
  s1 = warp_sum(s1) / (float)C;
  s2 = warp_sum(s2) / (float)C;
  for (int v = lane; v < nv; v += 32) {
    float xv[V], gv[V], wv[V], o[V];
    Elem<T>::load(xr + v * V, xv);
    Elem<T>::load(gr + v * V, gv);
    if (gamma) Elem<T>::load(gamma + v * V, wv);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const float gh = gamma ? gv[k] * wv[k] : gv[k];
      o[k] = rstd * (gh - s1 - (xv[k] - mean) * rstd * s2);
    }
    Elem<T>::store(dr + v * V, o);
  }
}

// ---- GEGLU backward: y = a * gelu(g)  ->  da = gy * gelu(g),  dg = gy * a * gelu'(g),
//      gelu'(g) = 0.5 (1 + erf(g / sqrt 2)) + g * exp(-g^2 / 2) / sqrt(2 pi)
template <typename T>
__global__ void __launch_bounds__(256)
geglu_bwd_kernel(T* __restrict__ dx, const T* __restrict__ gy, const T* __restrict__ x, int64_t rows, int D) {
  constexpr int V = Elem<T>::kVec;
  const int nv = D / V;
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  for (int64_t r = blockIdx.y; r < rows; r += gridDim.y) {
    float a[V], g[V], u[V], da[V], dg[V];
    Elem<T>::load(x + r * 2 * D + v * V, a);
    Elem<T>::load(x + r * 2 * D + D + v * V, g);
    Elem<T>::load(gy + r * D + v * V, u);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const float e = sizeof(T) == 2 ? erf_bf16_grade(g[k] * 0.70710678118654752f) : erff(g[k] * 0.70710678118654752f);
      const float cdf = 0.5f * (1.f + e);
      const float pdf = 0.3989422804014327f * __expf(-0.5f * g[k] * g[k]);
      da[k] = u[k] * g[k] * cdf;
      dg[k] = u[k] * a[k] * fmaf(g[k], pdf, cdf);
    }
    Elem<T>::store(dx + r * 2 * D + v * V, da);
    Elem<T>::store(dx + r * 2 * D + D + v * V, dg);
  }
}


// ---- nearest-neighbour 2x upsample on NHWC (Upsample2D before its 3x3 convolution) -----------------------------------
// One thread per 16-byte INPUT vector: read once, stored to the 2 x 2 output pixels (a warp covers 512 contiguous bytes
// of each).  Pure copy (bit-exact); writes 4x what it reads.
__global__ void __launch_bounds__(256)
upsample2x_kernel(uint4* __restrict__ y, const uint4* __restrict__ x, int64_t rows_in /* B*H */, int W, int nv) {
  pdl_trigger();
  pdl_wait();
  // blockIdx.x walks the vectors of one input row (W * nv), blockIdx.y strides over the B*H input rows
  const int wv = W * nv;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= wv) return;
  const int w = i / nv, c = i - w * nv;                          // 32-bit, once per thread
  const int64_t row_out = (int64_t)2 * wv;                       // vectors per output row
  for (int64_t bh = blockIdx.y; bh < rows_in; bh += gridDim.y) {
    const uint4 v = x[bh * wv + i];
    uint4* o = y + (2 * bh) * row_out + (int64_t)(2 * w) * nv + c;
    o[0] = v; o[nv] = v; o[row_out] = v; o[row_out + nv] = v;
  }
}


// ---- channel concatenation of two NHWC tensors (the skip connections of the up path: cat([h, skip], dim=1)) ----------
// One thread per 16-byte OUTPUT vector; a pixel's row is the row of a followed by the row of b.  Pure copy.
__global__ void __launch_bounds__(256)
concat2_kernel(uint4* __restrict__ y, const uint4* __restrict__ a, const uint4* __restrict__ b, int64_t pixels, int nva,
               int nvb) {
  pdl_trigger();
  pdl_wait();
  // blockIdx.x walks a pixel's output vectors, blockIdx.y strides over pixels: no 64-bit division per element (the
  // first version's `i / nvo` made this copy 5x slower than the bandwidth allows), four pixels per thread in flight
  const int nvo = nva + nvb;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nvo) return;
  const bool from_a = c < nva;
  const uint4* src = from_a ? a + c : b + (c - nva);
  const int sstride = from_a ? nva : nvb;
  int64_t pix = blockIdx.y;
  for (; pix + 3 * (int64_t)gridDim.y < pixels; pix += 4 * (int64_t)gridDim.y) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = src[(pix + u * (int64_t)gridDim.y) * sstride];
#pragma unroll
    for (int u = 0; u < 4; ++u) y[(pix + u * (int64_t)gridDim.y) * nvo + c] = v[u];
  }
  for (; pix < pixels; pix += gridDim.y) y[pix * nvo + c] = src[pix * sstride];
}

}  // namespace slider

using namespace slider;

extern "C" {

int slider_layernorm(void* y, const void* x, const void* gamma, const void* beta, int64_t rows, int channels,
                     float eps, int dtype, void* stream) {
  return slider_layernorm_train(y, nullptr, x, gamma, beta, rows, channels, eps, dtype, stream);
}

int slider_layernorm_train(void* y, float* stats, const void* x, const void* gamma, const void* beta, int64_t rows,
                           int channels, float eps, int dtype, void* stream) {
  SLIDER_CHECK_ARG(y && x && rows > 0 && channels > 0, "layernorm: bad argument");
  SLIDER_CHECK_ARG(channels % 8 == 0, "layernorm: channels must be a multiple of 8 (got %d)", channels);
  SLIDER_CHECK_ARG((((uintptr_t)y | (uintptr_t)x | (uintptr_t)gamma | (uintptr_t)beta) & 15) == 0,
                   "layernorm: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)((rows + 7) / 8);
  if (dtype != SLIDER_BF16 && dtype != SLIDER_F32 && dtype != SLIDER_F16) { set_error("layernorm: unsupported dtype %d", dtype); return SLIDER_ERR_UNSUPPORTED; }
  SLIDER_CHECK_ARG(dtype != SLIDER_F16 || stats == nullptr, "layernorm_train: fp16 storage is forward / inference only");
  const int per_lane = (channels / (dtype == SLIDER_F32 ? 4 : 8) + 31) / 32;      // 16-byte vectors per lane
#define SLIDER_LN_LAUNCH(KERNEL, TYPE)                                                                              \
  launch_pdl(KERNEL, dim3(grid), dim3(256), 0, st, (TYPE*)y, (const TYPE*)x, (const TYPE*)gamma, (const TYPE*)beta, \
             rows, channels, eps, stats)
  if (dtype == SLIDER_BF16) {
    if (per_lane <= 2) SLIDER_LN_LAUNCH((layernorm_kernel<__nv_bfloat16, 2>), __nv_bfloat16);
    else if (per_lane <= 3) SLIDER_LN_LAUNCH((layernorm_kernel<__nv_bfloat16, 3>), __nv_bfloat16);
    else if (per_lane <= 5) SLIDER_LN_LAUNCH((layernorm_kernel<__nv_bfloat16, 5>), __nv_bfloat16);
    else SLIDER_LN_LAUNCH(layernorm_stream_kernel<__nv_bfloat16>, __nv_bfloat16);
  } else if (dtype == SLIDER_F16) {
    if (per_lane <= 2) SLIDER_LN_LAUNCH((layernorm_kernel<__half, 2>), __half);
    else if (per_lane <= 3) SLIDER_LN_LAUNCH((layernorm_kernel<__half, 3>), __half);
    else if (per_lane <= 5) SLIDER_LN_LAUNCH((layernorm_kernel<__half, 5>), __half);
    else SLIDER_LN_LAUNCH(layernorm_stream_kernel<__half>, __half);
  } else {
    if (per_lane <= 3) SLIDER_LN_LAUNCH((layernorm_kernel<float, 3>), float);
    else if (per_lane <= 5) SLIDER_LN_LAUNCH((layernorm_kernel<float, 5>), float);
    else SLIDER_LN_LAUNCH(layernorm_stream_kernel<float>, float);
  }
#undef SLIDER_LN_LAUNCH
  return post_launch("layernorm_kernel");
}

int slider_geglu(void* y, const void* x, int64_t rows, int d, int dtype, void* stream) {
  SLIDER_CHECK_ARG(y && x && rows > 0 && d > 0 && d % 8 == 0, "geglu: bad argument");
  SLIDER_CHECK_ARG((((uintptr_t)y | (uintptr_t)x) & 15) == 0, "geglu: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int V = dtype == SLIDER_F32 ? 4 : 8;
  const int nv = d / V;
  const int threads = nv >= 256 ? 256 : ((nv + 31) / 32) * 32;
  const unsigned gx = (unsigned)((nv + threads - 1) / threads);
  int64_t gy = ((int64_t)sm_count() * 16 + gx - 1) / gx;       // ~16 CTAs per SM in flight, rows strided over grid.y
  if (gy > rows) gy = rows;
  if (gy > 65535) gy = 65535;
  if (dtype == SLIDER_BF16)
    launch_pdl(geglu_kernel<__nv_bfloat16>, dim3(gx, (unsigned)gy), dim3(threads), 0, st, (__nv_bfloat16*)y,
               (const __nv_bfloat16*)x, rows, d);
  else if (dtype == SLIDER_F16)
    launch_pdl(geglu_kernel<__half>, dim3(gx, (unsigned)gy), dim3(threads), 0, st, (__half*)y, (const __half*)x, rows, d);
  else if (dtype == SLIDER_F32)
    launch_pdl(geglu_kernel<float>, dim3(gx, (unsigned)gy), dim3(threads), 0, st, (float*)y, (const float*)x, rows, d);
  else { set_error("geglu: unsupported dtype %d", dtype); return SLIDER_ERR_UNSUPPORTED; }
  return post_launch("geglu_kernel");
}

int slider_layernorm_backward(void* dx, const void* gy, const void* x, const void* gamma, const float* stats,
                              int64_t rows, int channels, int dtype, void* stream) {
  SLIDER_CHECK_ARG(dx && gy && x && stats && rows > 0 && channels > 0 && channels % 8 == 0, "layernorm_backward: bad argument");
  SLIDER_CHECK_ARG((((uintptr_t)dx | (uintptr_t)gy | (uintptr_t)x | (uintptr_t)gamma) & 15) == 0,
                   "layernorm_backward: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)((rows + 7) / 8);
  if (dtype == SLIDER_BF16)
    layernorm_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)dx, (const __nv_bfloat16*)gy,
                                                             (const __nv_bfloat16*)x, (const __nv_bfloat16*)gamma, stats, rows, channels);
  else if (dtype == SLIDER_F32)
    layernorm_bwd_kernel<float><<<grid, 256, 0, st>>>((float*)dx, (const float*)gy, (const float*)x, (const float*)gamma,
                                                     stats, rows, channels);
  else { set_error("layernorm_backward: unsupported dtype %d", dtype); return SLIDER_ERR_UNSUPPORTED; }
  return post_launch("layernorm_bwd_kernel");
}

int slider_geglu_backward(void* dx, const void* gy, const void* x, int64_t rows, int d, int dtype, void* stream) {
  SLIDER_CHECK_ARG(dx && gy && x && rows > 0 && d > 0 && d % 8 == 0, "geglu_backward: bad argument");
  SLIDER_CHECK_ARG((((uintptr_t)dx | (uintptr_t)gy | (uintptr_t)x) & 15) == 0, "geglu_backward: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int V = dtype == SLIDER_BF16 ? 8 : 4;
  const int nv = d / V;
  const int threads = nv >= 256 ? 256 : ((nv + 31) / 32) * 32;
  const unsigned gx = (unsigned)((nv + threads - 1) / threads);
  int64_t gyb = ((int64_t)sm_count() * 16 + gx - 1) / gx;
  if (gyb > rows) gyb = rows;
  if (gyb > 65535) gyb = 65535;
  if (dtype == SLIDER_BF16)
    geglu_bwd_kernel<__nv_bfloat16><<<dim3(gx, (unsigned)gyb), threads, 0, st>>>((__nv_bfloat16*)dx, (const __nv_bfloat16*)gy,
                                                                               (const __nv_bfloat16*)x, rows, d);
  else if (dtype == SLIDER_F32)
    geglu_bwd_kernel<float><<<dim3(gx, (unsigned)gyb), threads, 0, st>>>((float*)dx, (const float*)gy, (const float*)x, rows, d);
  else { set_error("geglu_backward: unsupported dtype %d", dtype); return SLIDER_ERR_UNSUPPORTED; }
  return post_launch("geglu_bwd_kernel");
}

int slider_upsample_nearest2x_nhwc(void* y, const void* x, int batch, int height, int width, int channels, int dtype,
                                   void* stream) {
  SLIDER_CHECK_ARG(y && x && batch > 0 && height > 0 && width > 0 && channels > 0, "upsample2x: bad argument");
  SLIDER_CHECK_ARG(dtype == SLIDER_BF16 || dtype == SLIDER_F32 || dtype == SLIDER_F16, "upsample2x: unsupported dtype %d", dtype);
  const int per = dtype == SLIDER_F32 ? 4 : 8;
  SLIDER_CHECK_ARG(channels % per == 0, "upsample2x: channels must fill 16-byte vectors (got %d)", channels);
  SLIDER_CHECK_ARG((((uintptr_t)y | (uintptr_t)x) & 15) == 0, "upsample2x: pointers must be 16-byte aligned");
  const int nv = channels / per;
  const int wv = width * nv;
  const int threads = wv >= 256 ? 256 : ((wv + 31) / 32) * 32;
  const unsigned gx = (unsigned)((wv + threads - 1) / threads);
  int64_t gy = ((int64_t)sm_count() * 8 * 256 / threads + gx - 1) / gx;
  if (gy > (int64_t)batch * height) gy = (int64_t)batch * height;
  if (gy > 65535) gy = 65535;
  launch_pdl(upsample2x_kernel, dim3(gx, (unsigned)gy), dim3(threads), 0, (cudaStream_t)stream, (uint4*)y, (const uint4*)x,
             (int64_t)batch * height, width, nv);
  return post_launch("upsample2x_kernel");
}

int slider_concat_channels_nhwc(void* y, const void* a, const void* b, int64_t pixels, int channels_a, int channels_b,
                                int dtype, void* stream) {
  SLIDER_CHECK_ARG(y && a && b && pixels > 0 && channels_a > 0 && channels_b > 0, "concat: bad argument");
  SLIDER_CHECK_ARG(dtype == SLIDER_BF16 || dtype == SLIDER_F32 || dtype == SLIDER_F16, "concat: unsupported dtype %d", dtype);
  const int per = dtype == SLIDER_F32 ? 4 : 8;
  SLIDER_CHECK_ARG(channels_a % per == 0 && channels_b % per == 0,
                   "concat: channel counts must fill 16-byte vectors (got %d, %d)", channels_a, channels_b);
  SLIDER_CHECK_ARG((((uintptr_t)y | (uintptr_t)a | (uintptr_t)b) & 15) == 0, "concat: pointers must be 16-byte aligned");
  const int nva = channels_a / per, nvb = channels_b / per;
  const int nvo = nva + nvb;
  const int threads = nvo >= 256 ? 256 : ((nvo + 31) / 32) * 32;
  const unsigned gx = (unsigned)((nvo + threads - 1) / threads);
  int64_t gy = ((int64_t)sm_count() * 8 * 256 / threads + gx - 1) / gx;      // ~8 CTAs of 256 threads per SM
  if (gy > pixels) gy = pixels;
  if (gy > 65535) gy = 65535;
  launch_pdl(concat2_kernel, dim3(gx, (unsigned)gy), dim3(threads), 0, (cudaStream_t)stream, (uint4*)y, (const uint4*)a,
             (const uint4*)b, pixels, nva, nvb);
  return post_launch("concat2_kernel");
}

}  // extern "C"
