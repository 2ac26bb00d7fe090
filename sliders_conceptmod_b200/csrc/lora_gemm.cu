// Fused base + low-rank implicit GEMM: host side (tile / split-K planning, TMA descriptors, C-ABI entry points) and
// the small-M weight-streaming kernels.  The tcgen05 kernel itself is in lora_gemm_kernel.cuh.
#include <stdlib.h>
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.cuh"
#include "lora_gemm_kernel.cuh"

namespace slider {

extern template int launch_gemm_tile<160>(int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                          const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                          const GemmParams&, cudaStream_t);
extern template int launch_gemm_tile<80>(int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                         const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                         const GemmParams&, cudaStream_t);

// ---------------------------------------------------------------------------------------------
// Small-M path (M <= 16 rows, 1x1): time_emb_proj (M = 2).  HBM-bound on the weight matrix:
// one warp per output channel streams W[n,:] with 16-byte loads; the activation rows live in
// shared memory (fp32); h = x.A^T is recomputed per CTA (M*r dot products of length K).
constexpr int GEMV_WARPS = 8;
template <typename T> struct Raw16 {
  __device__ static inline void unpack(const uint4& r, float (&v)[Elem<T>::kVec]);
};
template <> __device__ inline void Raw16<__nv_bfloat16>::unpack(const uint4& t, float (&v)[8]) {
  const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __uint_as_float(w[i] << 16); v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
}
template <> __device__ inline void Raw16<__half>::unpack(const uint4& t, float (&v)[8]) {
  const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) Elem<__half>::unpack2(w[i], v[2 * i], v[2 * i + 1]);
}
template <> __device__ inline void Raw16<float>::unpack(const uint4& t, float (&v)[4]) {
  v[0] = __uint_as_float(t.x); v[1] = __uint_as_float(t.y); v[2] = __uint_as_float(t.z); v[3] = __uint_as_float(t.w);
}

// V consecutive fp32 activations from shared memory as 16-byte loads (the scalar form made the small-M kernels
// shared-memory-issue bound: one LDS.32 per FMA)
template <int V>
__device__ __forceinline__ void lds_vec(const float* p, float (&v)[V]) {
#pragma unroll
  for (int q = 0; q < V / 4; ++q) {
    const float4 t = reinterpret_cast<const float4*>(p)[q];
    v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
  }
}

// KC = 16-byte chunks of a weight row per lane (K <= KC * 32 * kVec).  All global loads of a warp (its W row,
// its lora_down row) are issued before the first barrier, so the kernel pays ONE DRAM latency, not one per chunk.
template <typename T, int MT, int KC>
__device__ __forceinline__ void
lora_gemv_body(T* __restrict__ y, const T* __restrict__ x, const T* __restrict__ W, const T* __restrict__ bias,
               const T* __restrict__ A, const T* __restrict__ Bup, float* __restrict__ h_out, int M, int K, int N,
               int rank, float s, int lora, int blk) {
  constexpr int V = Elem<T>::kVec;
  extern __shared__ float gemv_smem[];
  float* xs = gemv_smem;                  // [MT][K]
  float* hs = gemv_smem + (size_t)MT * K; // [MT][16]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blk * GEMV_WARPS + warp;
  const bool n_ok = n < N;
  uint4 wreg[KC], areg[KC];
  const int pr0 = warp;                   // first (row m, rank j) pair of this warp for h = x . A^T
  const bool pr_ok = lora && pr0 < MT * rank;
  pdl_trigger();
  {
    // the frozen base weight is the one large operand and is never written by a preceding kernel: its loads are
    // issued BEFORE the dependency wait, so under PDL they overlap the previous kernel's tail
    const uint4* wrow = reinterpret_cast<const uint4*>(W + (size_t)(n_ok ? n : 0) * K);
#pragma unroll
    for (int c = 0; c < KC; ++c) {
      const int k = (c * 32 + lane) * V;
      wreg[c] = (n_ok && k < K) ? __ldg(wrow + c * 32 + lane) : make_uint4(0, 0, 0, 0);
    }
    pdl_wait();
    const uint4* arow = reinterpret_cast<const uint4*>(A + (size_t)(pr_ok ? pr0 % rank : 0) * K);
#pragma unroll
    for (int c = 0; c < KC; ++c) {
      const int k = (c * 32 + lane) * V;
      areg[c] = (pr_ok && k < K) ? __ldg(arow + c * 32 + lane) : make_uint4(0, 0, 0, 0);
    }
  }
  for (int i = threadIdx.x * V; i < MT * K; i += blockDim.x * V) {
    const int m = i / K;
    float v[V];
    if (m < M) Elem<T>::load(x + i, v);
#pragma unroll
    for (int k = 0; k < V; ++k) xs[i + k] = m < M ? v[k] : 0.f;
  }
  __syncthreads();
  if (lora) {
    for (int pr = warp; pr < MT * rank; pr += GEMV_WARPS) {
      const int m = pr / rank, j = pr - m * rank;
      float acc = 0.f;
      if (pr == pr0) {
#pragma unroll
        for (int c = 0; c < KC; ++c) {
          const int k = (c * 32 + lane) * V;
          if (k < K) {
            float av[V], xv[V];
            Raw16<T>::unpack(areg[c], av);
            lds_vec<V>(xs + m * K + k, xv);
#pragma unroll
            for (int e = 0; e < V; ++e) acc = fmaf(av[e], xv[e], acc);
          }
        }
      } else {
        for (int k = lane * V; k < K; k += 32 * V) {
          float av[V], xv[V];
          Elem<T>::load(A + (size_t)j * K + k, av);
          lds_vec<V>(xs + m * K + k, xv);
#pragma unroll
          for (int e = 0; e < V; ++e) acc = fmaf(av[e], xv[e], acc);
        }
      }
      acc = warp_sum(acc);
      if (lane == 0) {
        hs[m * 16 + j] = acc;
        if (h_out != nullptr && blk == 0 && m < M) h_out[m * rank + j] = acc;
      }
    }
    __syncthreads();
  }
  if (!n_ok) return;
  float acc[MT];
#pragma unroll
  for (int m = 0; m < MT; ++m) acc[m] = 0.f;
#pragma unroll
  for (int c = 0; c < KC; ++c) {
    const int k = (c * 32 + lane) * V;
    if (k < K) {
      float wv[V];
      Raw16<T>::unpack(wreg[c], wv);
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        float xv[V];
        lds_vec<V>(xs + m * K + k, xv);
#pragma unroll
        for (int e = 0; e < V; ++e) acc[m] = fmaf(wv[e], xv[e], acc[m]);
      }
    }
  }
#pragma unroll
  for (int m = 0; m < MT; ++m) acc[m] = warp_sum(acc[m]);
  if (lane < MT && lane < M) {
    float a = 0.f;
#pragma unroll
    for (int m = 0; m < MT; ++m) a = (lane == m) ? acc[m] : a;
    if (bias != nullptr) a += Elem<T>::to_f(bias[n]);
    if (lora) {
      float l = 0.f;
      for (int j = 0; j < rank; ++j) l = fmaf(hs[lane * 16 + j], Elem<T>::to_f(Bup[(size_t)n * rank + j]), l);
      a = fmaf(s, l, a);
    }
    y[(size_t)lane * N + n] = Elem<T>::from_f(a);
  }
}

template <typename T, int MT, int KC>
__global__ void __launch_bounds__(GEMV_WARPS * 32)
lora_gemv_kernel(T* __restrict__ y, const T* __restrict__ x, const T* __restrict__ W, const T* __restrict__ bias,
                 const T* __restrict__ A, const T* __restrict__ Bup, float* __restrict__ h_out, int M, int K, int N,
                 int rank, float s, int lora) {
  lora_gemv_body<T, MT, KC>(y, x, W, bias, A, Bup, h_out, M, K, N, rank, s, lora, (int)blockIdx.x);
}

// Several small-M adapters on ONE input in one launch (the 16-22 `time_emb_proj` layers of a UNet forward all project
// the same silu(temb) rows): block b serves layer l with blk_end[l-1] <= b < blk_end[l].
constexpr int GEMV_MAX_LAYERS = 32;
struct GemvMultiParams {
  const void* W[GEMV_MAX_LAYERS];
  const void* bias[GEMV_MAX_LAYERS];
  const void* A[GEMV_MAX_LAYERS];
  const void* Bup[GEMV_MAX_LAYERS];
  void* y[GEMV_MAX_LAYERS];
  int N[GEMV_MAX_LAYERS];
  int blk_end[GEMV_MAX_LAYERS];
  float s[GEMV_MAX_LAYERS];
  int layers;
};
// Each warp walks `ch` output channels of its layer (stride GEMV_WARPS) with the NEXT channel's weight row in flight
// while it reduces the current one; the activation rows and h = x . A^T are staged once per block.  Per-channel
// arithmetic is the single-layer kernel's (same chunk order, same shuffles): results are bit-identical to it.
template <typename T, int MT, int KC>
__global__ void __launch_bounds__(GEMV_WARPS * 32)
lora_gemv_multi_kernel(const T* __restrict__ x, int M, int K, int rank, int ch, const __grid_constant__ GemvMultiParams mp) {
  constexpr int V = Elem<T>::kVec;
  extern __shared__ float gemv_smem[];
  float* xs = gemv_smem;                  // [MT][K]
  float* hs = gemv_smem + (size_t)MT * K; // [MT][16]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int l = 0;
  while (l + 1 < mp.layers && (int)blockIdx.x >= mp.blk_end[l]) ++l;
  const int blk = (int)blockIdx.x - (l > 0 ? mp.blk_end[l - 1] : 0);
  const float s = mp.s[l];
  const int N = mp.N[l];
  T* __restrict__ y = reinterpret_cast<T*>(mp.y[l]);
  const T* __restrict__ W = reinterpret_cast<const T*>(mp.W[l]);
  const T* __restrict__ bias = reinterpret_cast<const T*>(mp.bias[l]);
  const T* __restrict__ A = reinterpret_cast<const T*>(mp.A[l]);
  const T* __restrict__ Bup = reinterpret_cast<const T*>(mp.Bup[l]);
  const bool lora = A != nullptr && s != 0.f;
  uint4 w0[KC], w1[KC];
  auto load_row = [&](uint4 (&wr)[KC], int n) {
    const uint4* wrow = reinterpret_cast<const uint4*>(W + (size_t)n * K);
#pragma unroll
    for (int c = 0; c < KC; ++c) {
      const int k = (c * 32 + lane) * V;
      wr[c] = k < K ? __ldg(wrow + c * 32 + lane) : make_uint4(0, 0, 0, 0);
    }
  };
  auto reduce_row = [&](const uint4 (&wr)[KC], int n) {
    float acc[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[m] = 0.f;
#pragma unroll
    for (int c = 0; c < KC; ++c) {
      const int k = (c * 32 + lane) * V;
      if (k < K) {
        float wv[V];
        Raw16<T>::unpack(wr[c], wv);
#pragma unroll
        for (int m = 0; m < MT; ++m) {
          float xv[V];
          lds_vec<V>(xs + m * K + k, xv);
#pragma unroll
          for (int e = 0; e < V; ++e) acc[m] = fmaf(wv[e], xv[e], acc[m]);
        }
      }
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[m] = warp_sum(acc[m]);
    if (lane < MT && lane < M) {
      float a = 0.f;
#pragma unroll
      for (int m = 0; m < MT; ++m) a = (lane == m) ? acc[m] : a;
      if (bias != nullptr) a += Elem<T>::to_f(bias[n]);
      if (lora) {
        float lo = 0.f;
        for (int j = 0; j < rank; ++j) lo = fmaf(hs[lane * 16 + j], Elem<T>::to_f(Bup[(size_t)n * rank + j]), lo);
        a = fmaf(s, lo, a);
      }
      y[(size_t)lane * N + n] = Elem<T>::from_f(a);
    }
  };
  pdl_trigger();
  const int n_first = blk * ch * GEMV_WARPS + warp;
  // frozen weights are never written by a preceding kernel: the first row is requested before the dependency wait
  if (n_first < N) load_row(w0, n_first);
  pdl_wait();
  for (int i = threadIdx.x * V; i < MT * K; i += blockDim.x * V) {
    const int m = i / K;
    float v[V];
    if (m < M) Elem<T>::load(x + i, v);
#pragma unroll
    for (int k = 0; k < V; ++k) xs[i + k] = m < M ? v[k] : 0.f;
  }
  __syncthreads();
  if (lora) {
    for (int pr = warp; pr < MT * rank; pr += GEMV_WARPS) {
      const int m = pr / rank, j = pr - m * rank;
      float acc = 0.f;
      for (int k = lane * V; k < K; k += 32 * V) {
        float av[V], xv[V];
        Elem<T>::load(A + (size_t)j * K + k, av);
        lds_vec<V>(xs + m * K + k, xv);
#pragma unroll
        for (int e = 0; e < V; ++e) acc = fmaf(av[e], xv[e], acc);
      }
      acc = warp_sum(acc);
      if (lane == 0) hs[m * 16 + j] = acc;
    }
  }
  __syncthreads();
  for (int c = 0; c < ch; c += 2) {
    const int n0 = n_first + c * GEMV_WARPS, n1 = n0 + GEMV_WARPS, n2 = n1 + GEMV_WARPS;
    if (n0 >= N) break;
    if (c + 1 < ch && n1 < N) load_row(w1, n1);
    reduce_row(w0, n0);
    if (c + 1 >= ch || n1 >= N) break;
    if (c + 2 < ch && n2 < N) load_row(w0, n2);
    reduce_row(w1, n1);
  }
}

}  // namespace slider

// =================================== host side ===================================
namespace slider {

struct TileGeom { int TW, TH, TB; };

static long long* g_gemm_trace = nullptr;   // set by slider_debug_gemm_trace (profiling aid)
// minimum K blocks per split (slider_debug_splitk_depth / SLIDER_SPLITK_DEPTH)
// 0 = the measured heuristic in run_gemm
static int g_splitk_depth = [] { const char* e = getenv("SLIDER_SPLITK_DEPTH"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 0; }();
// K blocks of L2 look-ahead on the weight stream (SLIDER_W_PREFETCH, 0 = off)
static int g_w_prefetch = [] { const char* e = getenv("SLIDER_W_PREFETCH"); return e ? atoi(e) : 12; }();
// The 128 x 80 tile serves launches whose 128 x 160 tiling would leave at least half of the SMs without a tile
// (tiles * 2 <= this threshold; SLIDER_BN80_TILES / slider_debug_bn80_tiles: 0 = never, -1 = the SM count)
// CTA-pair form (SLIDER_PAIR / slider_debug_pair): 0 = never, 1 = wherever the geometry allows, -1 = the rule in run_gemm
static int g_pair = [] { const char* e = getenv("SLIDER_PAIR"); return e ? atoi(e) : -1; }();
static int g_bn80_tiles = [] { const char* e = getenv("SLIDER_BN80_TILES"); return e ? atoi(e) : -1; }();

static TileGeom choose_tile(int W, int H, int B) {
  TileGeom t;
  t.TW = W < BM ? W : BM;
  int rest = BM / t.TW;
  t.TH = H < rest ? H : rest;
  if (t.TH < 1) t.TH = 1;
  rest = BM / (t.TW * t.TH);
  t.TB = B < rest ? B : rest;
  if (t.TB < 1) t.TB = 1;
  return t;
}

static inline int esize(int dtype) { return dtype == SLIDER_F32 ? 4 : 2; }

template <typename T>
static int launch_gemv(const slider_layer_geom* g, int M, void* y, const void* x, const void* W, const void* bias,
                       const void* A, const void* Bup, float* h_out, float s, bool lora, cudaStream_t st) {
  const int K = g->cin, N = g->cout;
  const int mt = M <= 2 ? 2 : (M <= 4 ? 4 : (M <= 8 ? 8 : 16));
  const size_t smem = ((size_t)mt * K + (size_t)mt * 16) * sizeof(float);
  const int grid = (N + GEMV_WARPS - 1) / GEMV_WARPS;
  const int V = 16 / (int)sizeof(T);
  const int kc = (K + 32 * V - 1) / (32 * V);     // 16-byte chunks per lane (caller guarantees <= 8)
#define GEMV2(MT, KC)                                                                                             \
  launch_pdl(lora_gemv_kernel<T, MT, KC>, dim3(grid), dim3(GEMV_WARPS * 32), smem, st, (T*)y, (const T*)x, (const T*)W, \
             (const T*)bias, (const T*)A, (const T*)Bup, h_out, M, K, N, g->rank, s, lora ? 1 : 0)
#define GEMV(MT) do { if (kc <= 2) GEMV2(MT, 2); else if (kc <= 4) GEMV2(MT, 4); else GEMV2(MT, 8); } while (0)
  if (mt == 2) GEMV(2); else if (mt == 4) GEMV(4); else if (mt == 8) GEMV(8); else GEMV(16);
#undef GEMV
#undef GEMV2
  return post_launch("lora_gemv_kernel");
}

// Shared driver for forward (lora epilogue) and input-gradient (extended-K) modes.
//   act      [B,H,W,kc]      activations (x or gy); g->height / g->width are the INPUT extent
//   wgt      [nn, taps, kc]  weights (W or Wt)
//   lo       [rank, taps, kc] lora_down (forward only)
//   Bup      [nn, bup_ld]    up-projection rows (bup_ld = 0: rank)
//   xe / we  [B,H,W,16] / [nn, taps, 16]  extended-K pair (input gradient only)
static int run_gemm(const slider_layer_geom* g, int kc, int nn, void* y, const void* act, const void* wgt,
                    const void* bias, const void* lo, const void* Bup, float* h_out, float s, const void* xe,
                    const void* we, void* ws, size_t ws_bytes, cudaStream_t st, int groups = 1, int accumulate = 0,
                    int bup_ld = 0, int bup_jd = 1, int geglu = 0) {
  tmap_cache_trim();
  const int es = esize(g->dtype);
  const int taps = g->kh * g->kw;
  const int bk = ROW_BYTES / es;
  const int stride = g->stride == 2 ? 2 : 1;
  SLIDER_CHECK_ARG(g->dtype == SLIDER_BF16 || g->dtype == SLIDER_F32 || g->dtype == SLIDER_F16, "lora gemm: unsupported dtype %d", g->dtype);
  SLIDER_CHECK_ARG(g->dtype != SLIDER_F16 || xe == nullptr, "lora gemm: fp16 storage is forward / inference only");
  SLIDER_CHECK_ARG(g->kh == g->kw && (g->kh == 1 || g->kh == 3) && g->pad == (g->kh - 1) / 2,
                   "lora gemm: only 1x1/pad0 and 3x3/pad1 kernels (got %dx%d pad %d)", g->kh, g->kw, g->pad);
  SLIDER_CHECK_ARG(g->stride >= 0 && g->stride <= 2 && (stride == 1 || g->kh == 3),
                   "lora gemm: stride %d (only 1, and 2 for a 3x3 kernel)", g->stride);
  SLIDER_CHECK_ARG(g->batch > 0 && g->height > 0 && g->width > 0 && kc > 0 && nn > 0, "lora gemm: empty shape");
  SLIDER_CHECK_ARG(kc % 8 == 0 && nn % 8 == 0,
                   "lora gemm: channel counts must be multiples of 8 (cin %d, cout %d)", g->cin, g->cout);
  SLIDER_CHECK_ARG(g->rank >= 0 && g->rank <= 16, "lora gemm: rank %d not in 0..16", g->rank);
  SLIDER_CHECK_ARG(((uintptr_t)act & 15) == 0 && ((uintptr_t)wgt & 15) == 0 && ((uintptr_t)y & 15) == 0,
                   "lora gemm: pointers must be 16-byte aligned");

  const bool lora = lo != nullptr && g->rank > 0 && (s != 0.f || h_out != nullptr);
  const bool ext = xe != nullptr;
  SLIDER_CHECK_ARG(stride == 1 || (!ext && groups == 1), "lora gemm: stride 2 is a forward-only geometry");
  if (groups > 1) {
    SLIDER_CHECK_ARG(lora && !ext && taps == 1 && h_out == nullptr, "grouped lora gemm: needs an active 1x1 adapter, no h_out");
    SLIDER_CHECK_ARG(groups * g->rank <= HCOLS, "grouped lora gemm: groups * rank = %d exceeds %d", groups * g->rank, HCOLS);
    SLIDER_CHECK_ARG(nn % groups == 0 && (nn / groups) % 160 == 0,
                     "grouped lora gemm: columns per group (%d / %d) must be a multiple of 160", nn, groups);
    SLIDER_CHECK_ARG((long long)g->batch * g->height * g->width > 16, "grouped lora gemm: needs more than 16 rows");
  }
  if (lora) SLIDER_CHECK_ARG(Bup != nullptr, "lora gemm: lora_up weight is NULL");
  if (geglu) {
    SLIDER_CHECK_ARG(!lora && !ext && !accumulate && groups == 1 && taps == 1 && nn % 160 == 0,
                     "linear_geglu: plain 1x1 layer whose 2*d output columns are a multiple of 160 (got %d)", nn);
    SLIDER_CHECK_ARG((long long)g->batch * g->height * g->width > 16, "linear_geglu: needs more than 16 rows");
  }
  {
    // small-M path: a 128-row tensor-core tile would be > 87 % padding and the op is weight-bandwidth bound
    const long long M = (long long)g->batch * g->height * g->width;
    const size_t mt = M <= 2 ? 2 : (M <= 4 ? 4 : (M <= 8 ? 8 : 16));
    if (groups == 1 && !accumulate && !ext && taps == 1 && bup_ld == 0 && M <= 16 && mt * kc * sizeof(float) <= 40 * 1024 &&
        !geglu && kc * es <= 8 * 32 * 16 && ((uintptr_t)wgt & 15) == 0 && (lo == nullptr || ((uintptr_t)lo & 15) == 0)) {
      if (g->dtype == SLIDER_BF16)
        return launch_gemv<__nv_bfloat16>(g, (int)M, y, act, wgt, bias, lo, Bup, lora ? h_out : nullptr, s, lora, st);
      if (g->dtype == SLIDER_F16)
        return launch_gemv<__half>(g, (int)M, y, act, wgt, bias, lo, Bup, lora ? h_out : nullptr, s, lora, st);
      return launch_gemv<float>(g, (int)M, y, act, wgt, bias, lo, Bup, lora ? h_out : nullptr, s, lora, st);
    }
  }
  // output extent (the tile grid lives on OUTPUT pixels)
  const int Wo = (g->width + 2 * g->pad - g->kw) / stride + 1;
  const int Ho = (g->height + 2 * g->pad - g->kh) / stride + 1;
  const TileGeom tg = choose_tile(Wo, Ho, g->batch);

  GemmParams p;
  p.TW = tg.TW; p.TH = tg.TH; p.TB = tg.TB; p.rows = tg.TW * tg.TH * tg.TB;
  p.W = Wo; p.H = Ho; p.B = g->batch; p.stride = stride;
  p.mt_w = (Wo + tg.TW - 1) / tg.TW;
  p.mt_h = (Ho + tg.TH - 1) / tg.TH;
  p.mt_b = (g->batch + tg.TB - 1) / tg.TB;
  const int m_tiles = p.mt_w * p.mt_h * p.mt_b;
  // tile width: 160 columns unless that tiling leaves at least half of the SMs without a tile
  int bn = 160;
  {
    // measured on the SD-1.4 layer set (profiles/r02_layerset_bn80.txt): the narrow tile pays where split-K cannot
    // supply the parallelism (shallow K) or there is only a handful of tiles; the deep 3x3 convolutions at 16 x 16 /
    // 32 x 32 are L2-bandwidth bound and lose with it (27 % more operand bytes per FLOP)
    const int thr = g_bn80_tiles < 0 ? sm_count() : g_bn80_tiles;
    const int tiles160 = m_tiles * ((nn + 159) / 160);
    const int kit = taps * ((kc + bk - 1) / bk + (ext ? 1 : 0));
    if (!geglu && (g_bn80_tiles < 0 ? (tiles160 <= 12 || (tiles160 * 2 <= thr && kit <= 40)) : tiles160 * 2 <= thr)) bn = 80;
  }
  p.n_tiles = (nn + bn - 1) / bn; p.N = nn;
  p.taps = taps; p.kw = g->kw; p.pad = g->pad; p.cblocks = (kc + bk - 1) / bk;
  p.rank = g->rank; p.bup_ld = bup_ld > 0 ? bup_ld : g->rank; p.bup_jd = bup_jd; p.s = s; p.lora = lora ? 1 : 0; p.ext = ext ? 1 : 0;
  p.group_n = groups > 1 ? nn / groups : 0;
  p.accumulate = accumulate ? 1 : 0;
  p.geglu = geglu ? 1 : 0;
  p.total_tiles = m_tiles * p.n_tiles;
  p.m_tiles = m_tiles;
  // CTA pairs: bf16, wide tile, forward geometry; a grouped launch's h columns must not straddle the two halves
  bool pair = g_pair != 0 && g->dtype == SLIDER_BF16 && bn == 160 && !ext && m_tiles >= 2 &&
              (groups == 1 || (g->rank <= 8 && 8 % (g->rank > 0 ? g->rank : 1) == 0));
  // measured per signature (profiles/r02_ab_pair.txt): the pair form gains 5-15 % on the 3x3 convolutions with >= 8
  // pixel tiles and 7 % on the SD-XL q/k/v launch (20 K blocks, 384 tiles); its cluster start-up / tear-down costs
  // 1-1.5 us, which loses on launches with a handful of K blocks
  if (pair && g_pair < 0)
    pair = (taps == 9 && m_tiles >= 8) || (taps * p.cblocks >= 18 && p.total_tiles >= 2 * sm_count());
  p.per_tap = p.cblocks + (ext ? 1 : 0);
  p.kiters = taps * p.per_tap;
  // split-K when the tile count leaves more than half of the SMs idle and K is deep enough
  p.ksplit = 1;
  p.partial = reinterpret_cast<float*>(ws);
  const size_t unit_floats = bn == 160 ? TileCfg<160>::UNIT_FLOATS : TileCfg<80>::UNIT_FLOATS;
  const int nchunk = bn == 160 ? TileCfg<160>::NCHUNK : TileCfg<80>::NCHUNK;
  if (!geglu && p.total_tiles * 2 <= sm_count()) {
    int ks = sm_count() / p.total_tiles;
    // A split costs a second launch plus the partial round trip (~6 us at 32+ tiles, ~2.5 us at <= 12): measured on
    // the SD-1.4 layer set (profiles/r01_splitk_depth_sweep.txt) it only pays for deep K -- 64+ K blocks at >= 18 per
    // split, or 40+ at >= 6 per split for the weight-streaming shapes with a handful of tiles.
    int depth = g_splitk_depth, min_k = 0;
    const int tiles_eq = bn == 160 ? p.total_tiles : (p.total_tiles + 1) / 2;     // in 160-column tiles (what the sweep measured)
    if (depth == 0) {
      depth = tiles_eq <= 12 ? 6 : 18;
      min_k = tiles_eq <= 12 ? 40 : 64;
    }
    const int by_depth = p.kiters >= min_k ? p.kiters / depth : 1;
    if (ks > by_depth) ks = by_depth;
    if (ws != nullptr && ((uintptr_t)ws & 15) == 0) {
      const long long cap = (long long)(ws_bytes / (unit_floats * sizeof(float)));
      if ((long long)ks * p.total_tiles > cap) ks = (int)(cap / p.total_tiles);
      if (ks >= 2) p.ksplit = ks;
    }
  }
  p.total_units = p.total_tiles * p.ksplit;
  p.pair_units = p.n_tiles * ((m_tiles + 1) / 2) * p.ksplit;
  p.fd_n_tiles = FastDiv::make((uint32_t)p.n_tiles);
  p.fd_mt_w = FastDiv::make((uint32_t)p.mt_w);
  p.fd_mt_h = FastDiv::make((uint32_t)p.mt_h);
  p.fd_ksplit = FastDiv::make((uint32_t)p.ksplit);
  p.fd_per_tap = FastDiv::make((uint32_t)p.per_tap);
  p.fd_kw = FastDiv::make((uint32_t)p.kw);
  p.trace = g_gemm_trace;
  { static const int skip = [] { const char* e = getenv("SLIDER_DBG_SKIP"); return e ? atoi(e) : 0; }(); p.dbg_skip = skip; }
  // only when several pixel tiles share each weight tile: with ONE pixel tile every CTA streams distinct weights and
  // the pipeline's own requests already fill HBM
  p.w_prefetch = (m_tiles > 1) ? g_w_prefetch : 0;
  p.y = y; p.bias = bias; p.Bup = Bup; p.h_out = lora ? h_out : nullptr;
  if (lora) {
    SLIDER_CHECK_ARG((g->rank * taps * kc * es) % 16 == 0 && ((uintptr_t)lo & 15) == 0, "lora gemm: lora_down misaligned");
  }

  // TMA descriptors come from a per-thread cache keyed by (pointer, shape, box): the eager drop-in trainers issue the
  // same few hundred launches every step and would otherwise re-encode 5-7 descriptors per launch
  const CUtensorMap *mA, *mW, *mL, *mXe, *mWe, *mY, *mP;
  {
    // stride 2: the box spans 2*TW x 2*TH input pixels and the map's element strides pick every second one
    uint64_t dims[4] = {(uint64_t)kc, (uint64_t)g->width, (uint64_t)g->height, (uint64_t)g->batch};
    uint64_t str[3] = {(uint64_t)kc * es, (uint64_t)kc * es * g->width, (uint64_t)kc * es * g->width * g->height};
    uint32_t box[4] = {(uint32_t)bk, (uint32_t)(tg.TW * stride), (uint32_t)(tg.TH * stride), (uint32_t)tg.TB};
    uint32_t est[4] = {1u, (uint32_t)stride, (uint32_t)stride, 1u};
    if (!(mA = cached_tmap(g->dtype, 4, act, dims, str, box, ROW_BYTES, est))) return SLIDER_ERR_CUDA;
  }
  {
    uint64_t dims[3] = {(uint64_t)kc, (uint64_t)taps, (uint64_t)nn};
    uint64_t str[2] = {(uint64_t)kc * es, (uint64_t)kc * es * taps};
    uint32_t box[3] = {(uint32_t)bk, 1u, (uint32_t)(pair ? bn / 2 : bn)};      // pair mode: each CTA stages half of the rows
    if (!(mW = cached_tmap(g->dtype, 3, wgt, dims, str, box, ROW_BYTES))) return SLIDER_ERR_CUDA;
  }
  mL = mW; mXe = mA; mWe = mW;
  if (lora) {
    uint64_t dims[3] = {(uint64_t)kc, (uint64_t)taps, (uint64_t)(g->rank * groups)};
    uint64_t str[2] = {(uint64_t)kc * es, (uint64_t)kc * es * taps};
    uint32_t box[3] = {(uint32_t)bk, 1u, (uint32_t)(pair ? HCOLS / 2 : HCOLS)};
    if (!(mL = cached_tmap(g->dtype, 3, lo, dims, str, box, ROW_BYTES))) return SLIDER_ERR_CUDA;
  }
  if (ext) {
    SLIDER_CHECK_ARG(we != nullptr && ((uintptr_t)xe & 15) == 0 && ((uintptr_t)we & 15) == 0, "lora gemm: bad ext operands");
    const uint32_t eb = EXT_ELEMS * es;
    uint64_t dims[4] = {(uint64_t)EXT_ELEMS, (uint64_t)g->width, (uint64_t)g->height, (uint64_t)g->batch};
    uint64_t str[3] = {(uint64_t)eb, (uint64_t)eb * g->width, (uint64_t)eb * g->width * g->height};
    uint32_t box[4] = {(uint32_t)EXT_ELEMS, (uint32_t)tg.TW, (uint32_t)tg.TH, (uint32_t)tg.TB};
    if (!(mXe = cached_tmap(g->dtype, 4, xe, dims, str, box, eb))) return SLIDER_ERR_CUDA;
    uint64_t wdims[3] = {(uint64_t)EXT_ELEMS, (uint64_t)taps, (uint64_t)nn};
    uint64_t wstr[2] = {(uint64_t)eb, (uint64_t)eb * taps};
    uint32_t wbox[3] = {(uint32_t)EXT_ELEMS, 1u, (uint32_t)bn};
    if (!(mWe = cached_tmap(g->dtype, 3, we, wdims, wstr, wbox, eb))) return SLIDER_ERR_CUDA;
  }
  {
    const uint64_t ny = geglu ? (uint64_t)nn / 2 : (uint64_t)nn;       // GEGLU: y holds d = N / 2 columns
    uint64_t dims[4] = {ny, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)g->batch};
    uint64_t str[3] = {ny * es, ny * es * Wo, ny * es * Wo * Ho};
    uint32_t box[4] = {(uint32_t)(OUT_ROW_BYTES / es), (uint32_t)tg.TW, (uint32_t)tg.TH, (uint32_t)tg.TB};
    if (!(mY = cached_tmap(g->dtype, 4, y, dims, str, box, OUT_ROW_BYTES))) return SLIDER_ERR_CUDA;
  }
  mP = mY;
  if (p.ksplit > 1) {
    uint64_t dims[2] = {(uint64_t)PCW, (uint64_t)p.total_units * nchunk * BM};
    uint64_t str[1] = {(uint64_t)OUT_ROW_BYTES};
    uint32_t box[2] = {(uint32_t)PCW, (uint32_t)BM};
    if (!(mP = cached_tmap(SLIDER_F32, 2, ws, dims, str, box, OUT_ROW_BYTES))) return SLIDER_ERR_CUDA;
  }
  const int r4 = !lora ? 0 : (g->rank <= 4 ? 1 : (g->rank <= 8 ? 2 : 4));
  if (g->dtype == SLIDER_F16) return launch_gemm_f16(bn, r4, *mA, *mW, *mL, *mXe, *mWe, *mY, *mP, p, st);
  if (pair) return launch_gemm_pair(r4, *mA, *mW, *mL, *mXe, *mWe, *mY, *mP, p, st);
  if (bn == 160) return launch_gemm_tile<160>(g->dtype, r4, *mA, *mW, *mL, *mXe, *mWe, *mY, *mP, p, st);
  return launch_gemm_tile<80>(g->dtype, r4, *mA, *mW, *mL, *mXe, *mWe, *mY, *mP, p, st);
}

template <typename T>
static int launch_gemv_multi(const GemvMultiParams& mp, int grid, int ch, const void* x, int M, int K, int rank,
                             cudaStream_t st) {
  const int mt = M <= 2 ? 2 : (M <= 4 ? 4 : (M <= 8 ? 8 : 16));
  const size_t smem = ((size_t)mt * K + (size_t)mt * 16) * sizeof(float);
  const int V = 16 / (int)sizeof(T);
  const int kc = (K + 32 * V - 1) / (32 * V);
#define GEMVM2(MT, KC) launch_pdl(lora_gemv_multi_kernel<T, MT, KC>, dim3(grid), dim3(GEMV_WARPS * 32), smem, st, (const T*)x, M, K, rank, ch, mp)
#define GEMVM(MT) do { if (kc <= 2) GEMVM2(MT, 2); else if (kc <= 5) GEMVM2(MT, 5); else GEMVM2(MT, 8); } while (0)
  if (mt == 2) GEMVM(2); else if (mt == 4) GEMVM(4); else if (mt == 8) GEMVM(8); else GEMVM(16);
#undef GEMVM
#undef GEMVM2
  return post_launch("lora_gemv_multi_kernel");
}

int run_gemm_plain(const slider_layer_geom* g, int kc, int nn, void* y, const void* act, const void* wgt, void* ws,
                   size_t ws_bytes, cudaStream_t st) {
  return run_gemm(g, kc, nn, y, act, wgt, nullptr, nullptr, nullptr, nullptr, 0.f, nullptr, nullptr, ws, ws_bytes, st);
}

// Fused 1x1 backward: the forward kernel on the transposed problem (K = cout, N = cin): activations gy, weights Wt,
// "lora_down" = B^T [rank.., cout], "lora_up" = A read as its own transpose; h output = gy . B.
int run_gemm_fused_bwd(const slider_layer_geom* g, void* dx, float* g_out, const void* gy, const void* Wt,
                       const void* Bt, const void* A, float s, void* ws, size_t ws_bytes, cudaStream_t st) {
  return run_gemm(g, g->cout, g->cin, dx, gy, Wt, nullptr, Bt, A, g_out, s, nullptr, nullptr, ws, ws_bytes, st, 1, 0,
                  /*bup_ld=*/1, /*bup_jd=*/g->cin);
}

}  // namespace slider

using namespace slider;

extern "C" int slider_linear_geglu(const slider_layer_geom* g, void* y, const void* x, const void* W_perm,
                                   const void* bias_perm, void* stream) {
  SLIDER_CHECK_ARG(g && y && x && W_perm, "linear_geglu: null pointer");
  SLIDER_CHECK_ARG(g->kh == 1 && g->kw == 1 && g->stride != 2, "linear_geglu: 1x1 / Linear layers only");
  return run_gemm(g, g->cin, g->cout, y, x, W_perm, bias_perm, nullptr, nullptr, nullptr, 0.f, nullptr, nullptr, nullptr, 0,
                  (cudaStream_t)stream, 1, 0, 0, 1, /*geglu=*/1);
}

extern "C" {

int slider_debug_splitk_depth(int d) { g_splitk_depth = d > 0 ? d : 0; return SLIDER_OK; }
int slider_debug_bn80_tiles(int tiles) { g_bn80_tiles = tiles; return SLIDER_OK; }
int slider_debug_pair(int mode) { g_pair = mode; return SLIDER_OK; }
int slider_debug_gemm_trace(void* buf) { g_gemm_trace = reinterpret_cast<long long*>(buf); return SLIDER_OK; }

size_t slider_gemm_workspace_bytes(void) { return (size_t)160 * TileCfg<160>::UNIT_FLOATS * sizeof(float); }

int slider_lora_forward(const slider_layer_geom* g, void* y, const void* x, const void* W, const void* bias,
                        const void* A, const void* B, float* h_out, float s, void* ws, size_t ws_bytes,
                        void* stream) {
  SLIDER_CHECK_ARG(g && y && x && W, "lora_forward: null pointer");
  return run_gemm(g, g->cin, g->cout, y, x, W, bias, A, B, h_out, s, nullptr, nullptr, ws, ws_bytes,
                  (cudaStream_t)stream);
}

int slider_lora_forward_ex(const slider_layer_geom* g, void* y, const void* x, const void* W, const void* bias,
                           const void* A, const void* B, float* h_out, float s, unsigned flags, void* ws,
                           size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(g && y && x && W, "lora_forward_ex: null pointer");
  SLIDER_CHECK_ARG((flags & ~(unsigned)SLIDER_FWD_ACCUMULATE) == 0, "lora_forward_ex: unknown flags 0x%x", flags);
  return run_gemm(g, g->cin, g->cout, y, x, W, bias, A, B, h_out, s, nullptr, nullptr, ws, ws_bytes,
                  (cudaStream_t)stream, 1, (flags & SLIDER_FWD_ACCUMULATE) ? 1 : 0);
}

int slider_lora_forward_grouped(const slider_layer_geom* g, void* y, const void* x, const void* W, const void* bias,
                                const void* A, const void* B, int groups, float s, void* ws, size_t ws_bytes,
                                void* stream) {
  SLIDER_CHECK_ARG(g && y && x && W && A && B, "lora_forward_grouped: null pointer");
  SLIDER_CHECK_ARG(groups >= 1, "lora_forward_grouped: groups must be >= 1");
  SLIDER_CHECK_ARG(s != 0.f, "lora_forward_grouped: s == 0 (adapter off) is the plain slider_lora_forward call");
  return run_gemm(g, g->cin, g->cout, y, x, W, bias, A, B, nullptr, s, nullptr, nullptr, ws, ws_bytes,
                  (cudaStream_t)stream, groups);
}

int slider_lora_forward_small_multi(int dtype, int layers, int M, int K, int rank, const void* x, void* const* y,
                                    const void* const* W, const void* const* bias, const void* const* A,
                                    const void* const* B, const int* N, const float* s, void* stream) {
  SLIDER_CHECK_ARG(dtype == SLIDER_BF16 || dtype == SLIDER_F32 || dtype == SLIDER_F16, "lora_forward_small_multi: unsupported dtype %d", dtype);
  SLIDER_CHECK_ARG(layers >= 1 && layers <= GEMV_MAX_LAYERS, "lora_forward_small_multi: layers %d not in 1..%d", layers,
                   GEMV_MAX_LAYERS);
  SLIDER_CHECK_ARG(x && y && W && N && s, "lora_forward_small_multi: null pointer");
  SLIDER_CHECK_ARG(M >= 1 && M <= 16, "lora_forward_small_multi: M = %d rows (the weight-streaming path serves 1..16)", M);
  const int es = esize(dtype);
  const size_t mt = M <= 2 ? 2 : (M <= 4 ? 4 : (M <= 8 ? 8 : 16));
  SLIDER_CHECK_ARG(K > 0 && K % 8 == 0 && mt * K * sizeof(float) <= 40 * 1024 && (size_t)K * es <= 8 * 32 * 16,
                   "lora_forward_small_multi: K = %d out of range", K);
  SLIDER_CHECK_ARG(rank >= 0 && rank <= 16, "lora_forward_small_multi: rank %d not in 0..16", rank);
  SLIDER_CHECK_ARG(((uintptr_t)x & 15) == 0, "lora_forward_small_multi: x must be 16-byte aligned");
  GemvMultiParams mp;
  for (int l = 0; l < layers; ++l) {
    SLIDER_CHECK_ARG(y[l] && W[l] && N[l] > 0, "lora_forward_small_multi: layer %d has a null pointer / empty N", l);
    SLIDER_CHECK_ARG(((uintptr_t)W[l] & 15) == 0, "lora_forward_small_multi: W of layer %d must be 16-byte aligned", l);
    const bool lora = A && A[l] && rank > 0 && s[l] != 0.f;
    if (lora) SLIDER_CHECK_ARG(B && B[l] && ((uintptr_t)A[l] & 15) == 0, "lora_forward_small_multi: bad adapter of layer %d", l);
    mp.W[l] = W[l]; mp.bias[l] = bias ? bias[l] : nullptr;
    mp.A[l] = lora ? A[l] : nullptr; mp.Bup[l] = lora ? B[l] : nullptr;
    mp.y[l] = y[l]; mp.N[l] = N[l]; mp.s[l] = s[l];
  }
  mp.layers = layers;
  // channels per warp: the smallest count for which the whole call is ONE wave of 4 resident blocks per SM (64 regs)
  int ch = 1, blocks = 0;
  for (; ch <= 16; ++ch) {
    blocks = 0;
    for (int l = 0; l < layers; ++l) blocks += (N[l] + GEMV_WARPS * ch - 1) / (GEMV_WARPS * ch);
    if (blocks <= 4 * sm_count()) break;
  }
  if (ch > 16) ch = 16;
  blocks = 0;
  for (int l = 0; l < layers; ++l) {
    blocks += (N[l] + GEMV_WARPS * ch - 1) / (GEMV_WARPS * ch);
    mp.blk_end[l] = blocks;
  }
  if (dtype == SLIDER_BF16) return launch_gemv_multi<__nv_bfloat16>(mp, blocks, ch, x, M, K, rank, (cudaStream_t)stream);
  if (dtype == SLIDER_F16) return launch_gemv_multi<__half>(mp, blocks, ch, x, M, K, rank, (cudaStream_t)stream);
  return launch_gemv_multi<float>(mp, blocks, ch, x, M, K, rank, (cudaStream_t)stream);
}

int slider_lora_input_grad(const slider_layer_geom* g, void* dx, const void* gy, const void* Wt, const void* dh,
                           const void* At, void* ws, size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(g && dx && gy && Wt, "lora_input_grad: null pointer");
  SLIDER_CHECK_ARG((dh == nullptr) == (At == nullptr), "lora_input_grad: dh and At must both be set or both NULL");
  return run_gemm(g, g->cout, g->cin, dx, gy, Wt, nullptr, nullptr, nullptr, nullptr, 0.f, dh, At, ws, ws_bytes,
                  (cudaStream_t)stream);
}

}  // extern "C"
