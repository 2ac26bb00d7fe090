// Low-rank side of the backward pass (bandwidth-class work):
//   dh[p, j]  = s * sum_n gy[p, n] * B[n, j]      tensor cores: the forward GEMM kernel on gy x (s B)^T
//   colreduce dB[n, j]     = s * sum_p gy[p, n] * h[p, j]
//             dA[j, t, c]  =     sum_p x[p + d(t), c] * dh[p, j]
//   pack_at / pack_wt      operand re-layouts for the input-gradient GEMM.
// The contraction over pixels is split over CTAs and finished with fp32 atomics.
#include "common.cuh"

namespace slider {

// --------------------------------------------------------------- pack_bt
// Bt[j][n] = s * B[n][j] (j < rank, else 0): the [16, N] K-major weight of the dh = gy . (s B) GEMM.
template <typename T>
__global__ void pack_bt_kernel(T* __restrict__ Bt, const T* __restrict__ Bup, int N, int rank, float s) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 16 * N) return;
  const int j = i / N, n = i - j * N;
  Bt[i] = j < rank ? Elem<T>::from_f(s * Elem<T>::to_f(Bup[(size_t)n * rank + j])) : Elem<T>::from_f(0.f);
}

// --------------------------------------------------------------- colreduce
// out[c, j, t] (generic strides) += scale * sum_p X[p + d(t), c] * V[p, j0 + j],  j = 0..3
// HBM-bound on X (gy for dB, x for dA), which is read exactly once with vector loads:
//   * a thread owns CH adjacent channels (16-byte loads for 1x1, 8/16-byte for 3x3) and keeps
//     TAPS x CH x 4 fp32 accumulators; PL pixel lanes share a CTA;
//   * the V rows (h fp32 / dh) of the CTA's pixel range (+ halo for 3x3) are staged once in shared memory
//     as float4, so the inner loop is one broadcast LDS.128 per tap;
//   * a CTA owns a narrow channel slice (4 vector threads) and 64 pixel lanes; lanes are reduced by warp
//     shuffles, warps through shared memory, then one fp32 atomicAdd per output element and CTA.
struct ColReduceParams {
  int B, H, W, C;        // X is [B,H,W,C]
  int kw, pad;
  int rank, vstride;     // V is [pixels, vstride]
  int NVT;               // channel-vector threads per pixel lane (covering CB = NVT*CH channels per block)
  int PL;                // pixel lanes
  int cpi;               // chunks per image
  int ppc;               // pixels per chunk
  int64_t oc, oj, ot;    // output strides for (channel, rank index, tap)
  float scale;
};

template <typename T, int CH> struct ChanVec;
template <> struct ChanVec<__nv_bfloat16, 8> {
  __device__ static inline void load(const __nv_bfloat16* p, float (&v)[8]) { Elem<__nv_bfloat16>::load(p, v); }
};
template <> struct ChanVec<__nv_bfloat16, 4> {
  __device__ static inline void load(const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 t = *reinterpret_cast<const uint2*>(p);
    v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xffff0000u);
    v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xffff0000u);
  }
};
template <> struct ChanVec<float, 8> {
  __device__ static inline void load(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <> struct ChanVec<float, 4> {
  __device__ static inline void load(const float* p, float (&v)[4]) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  }
};

template <typename T, typename VT, int TAPS, int CH>
__global__ void __launch_bounds__(256)
colreduce_kernel(float* __restrict__ out, const T* __restrict__ X, const VT* __restrict__ V,
                 const ColReduceParams p, int j0) {
  extern __shared__ __align__(16) float cr_smem[];
  const int HW = p.H * p.W;
  const int img = blockIdx.x / p.cpi, chunk = blockIdx.x - img * p.cpi;
  const int q0 = chunk * p.ppc;
  const int q1 = min(HW, q0 + p.ppc);
  const int halo = TAPS > 1 ? p.pad * (p.W + 1) : 0;
  const int v0 = max(0, q0 - halo), v1 = min(HW, q1 + halo);   // V pixels staged: [v0, v1) of image img
  float4* s_v = reinterpret_cast<float4*>(cr_smem);
  for (int i = threadIdx.x; i < v1 - v0; i += blockDim.x) {
    const VT* vp = V + ((size_t)img * HW + v0 + i) * p.vstride + j0;
    float4 f;
    f.x = Elem<VT>::to_f(vp[0]);
    f.y = (j0 + 1 < p.rank) ? Elem<VT>::to_f(vp[1]) : 0.f;
    f.z = (j0 + 2 < p.rank) ? Elem<VT>::to_f(vp[2]) : 0.f;
    f.w = (j0 + 3 < p.rank) ? Elem<VT>::to_f(vp[3]) : 0.f;
    s_v[i] = f;
  }
  __syncthreads();
  const int pl = threadIdx.x / p.NVT, v = threadIdx.x - pl * p.NVT;
  const int c = (blockIdx.y * p.NVT + v) * CH;
  const bool active = pl < p.PL && c < p.C;
  float acc[TAPS][CH][4];
#pragma unroll
  for (int t = 0; t < TAPS; ++t)
#pragma unroll
    for (int k = 0; k < CH; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[t][k][j] = 0.f;
  if (active) {
    const T* xb = X + ((size_t)img * HW) * p.C + c;
    // software pipeline: the X vectors of the next two pixels are in flight while this one is consumed
    float xn0[CH], xn1[CH];
    int q = q0 + pl;
    if (q < q1) ChanVec<T, CH>::load(xb + (size_t)q * p.C, xn0);
    if (q + p.PL < q1) ChanVec<T, CH>::load(xb + (size_t)(q + p.PL) * p.C, xn1);
    for (; q < q1; q += p.PL) {
      float xv[CH];
#pragma unroll
      for (int k = 0; k < CH; ++k) { xv[k] = xn0[k]; xn0[k] = xn1[k]; }
      if (q + 2 * p.PL < q1) ChanVec<T, CH>::load(xb + (size_t)(q + 2 * p.PL) * p.C, xn1);
      if (TAPS == 1) {
        const float4 f = s_v[q - v0];
#pragma unroll
        for (int k = 0; k < CH; ++k) {
          acc[0][k][0] = fmaf(xv[k], f.x, acc[0][k][0]);
          acc[0][k][1] = fmaf(xv[k], f.y, acc[0][k][1]);
          acc[0][k][2] = fmaf(xv[k], f.z, acc[0][k][2]);
          acc[0][k][3] = fmaf(xv[k], f.w, acc[0][k][3]);
        }
      } else {
        const int h = q / p.W, w = q - h * p.W;
#pragma unroll
        for (int t = 0; t < TAPS; ++t) {
          // x[q] meets dh[pv] where q = pv + d(t)  =>  pv = q - d(t)
          const int ph = h - (t / 3 - p.pad), pw = w - (t % 3 - p.pad);
          if (ph >= 0 && ph < p.H && pw >= 0 && pw < p.W) {
            const float4 f = s_v[ph * p.W + pw - v0];
#pragma unroll
            for (int k = 0; k < CH; ++k) {
              acc[t][k][0] = fmaf(xv[k], f.x, acc[t][k][0]);
              acc[t][k][1] = fmaf(xv[k], f.y, acc[t][k][1]);
              acc[t][k][2] = fmaf(xv[k], f.z, acc[t][k][2]);
              acc[t][k][3] = fmaf(xv[k], f.w, acc[t][k][3]);
            }
          }
        }
      }
    }
  }
  // cross-lane reduction: the 8 pixel lanes of a warp by shuffles (lane = pixel-lane * 4 + channel-vector thread,
  // NVT == 4), the warps through shared memory (re-using the V staging area), then one fp32 atomicAdd per output
  // element and CTA.  A CTA covers only NVT * CH channels, so it issues TAPS * NVT * CH * 4 atomics in total.
  constexpr int NACC = TAPS * CH * 4;
#pragma unroll
  for (int t = 0; t < TAPS; ++t)
#pragma unroll
    for (int k = 0; k < CH; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float a = acc[t][k][j];
        a += __shfl_xor_sync(0xffffffffu, a, 4);
        a += __shfl_xor_sync(0xffffffffu, a, 8);
        a += __shfl_xor_sync(0xffffffffu, a, 16);
        acc[t][k][j] = a;
      }
  float* s_red = cr_smem;     // [warps][NVT][NACC]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  __syncthreads();            // every thread is done with the staged V rows
  if (lane < 4) {
    float4* dst = reinterpret_cast<float4*>(s_red + ((size_t)warp * 4 + lane) * NACC);
#pragma unroll
    for (int t = 0; t < TAPS; ++t)
#pragma unroll
      for (int k = 0; k < CH; ++k)
        dst[t * CH + k] = make_float4(acc[t][k][0], acc[t][k][1], acc[t][k][2], acc[t][k][3]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * NACC; i += blockDim.x) {
    const int vv = i / NACC, rem = i - vv * NACC;
    const int t = rem / (CH * 4), k = (rem >> 2) % CH, j = rem & 3;
    const int cc = (blockIdx.y * 4 + vv) * CH + k;
    if (cc < p.C && j0 + j < p.rank) {
      float sum = 0.f;
      for (int w = 0; w < nwarps; ++w) sum += s_red[(size_t)w * 4 * NACC + i];
      atomicAdd(out + (int64_t)cc * p.oc + (int64_t)(j0 + j) * p.oj + (int64_t)t * p.ot, sum * p.scale);
    }
  }
}

template <typename T>
__global__ void pack_at_kernel(T* __restrict__ At, const T* __restrict__ A, int cin, int taps, int rank) {
  const int64_t total = (int64_t)cin * taps * 16;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i % 16);
    const int t = (int)((i / 16) % taps);
    const int c = (int)(i / (16 * taps));
    At[i] = j < rank ? A[((int64_t)j * taps + (taps - 1 - t)) * cin + c] : Elem<T>::from_f(0.f);
  }
}

template <typename T>
__global__ void pack_wt_kernel(T* __restrict__ Wt, const T* __restrict__ W, int cin, int cout, int taps) {
  // Wt[c][t][n] = W[n][taps-1-t][c]; 32x32 shared-memory transpose per (tap, tile)
  __shared__ T tile[32][33];
  const int t = blockIdx.z;
  const int c0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int n = n0 + i, c = c0 + threadIdx.x;
    if (n < cout && c < cin) tile[i][threadIdx.x] = W[((int64_t)n * taps + (taps - 1 - t)) * cin + c];
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, n = n0 + threadIdx.x;
    if (n < cout && c < cin) Wt[((int64_t)c * taps + t) * cout + n] = tile[threadIdx.x][i];
  }
}

template <typename T, typename VT>
static int launch_colreduce(float* out, const T* X, const VT* v, ColReduceParams p, int taps, cudaStream_t st) {
  const int CH = taps == 1 ? 8 : 4;
  const int HW = p.H * p.W;
  // a CTA = 4 channel-vector threads (NVT, fixed: the shuffle reduction relies on it) x 64 pixel lanes: a narrow
  // channel slice keeps the per-CTA output (and its atomics) small, the pixel lanes supply the parallelism
  const int max_threads = 256;
  const int nvt_all = (p.C + CH - 1) / CH;
  const int cblocks = (nvt_all + 3) / 4;
  p.NVT = 4;
  p.PL = max_threads / p.NVT;
  // pixel chunks: ~2 CTAs per SM over (images x chunks x channel blocks); chunks never straddle images
  int want = (sm_count() * 2) / (p.B * cblocks);
  if (want < 1) want = 1;
  int ppc = (HW + want - 1) / want;
  if (ppc < p.PL * 2) ppc = p.PL * 2;
  if (ppc > HW) ppc = HW;
  p.ppc = ppc;
  p.cpi = (HW + ppc - 1) / ppc;
  const int halo = taps > 1 ? p.pad * (p.W + 1) : 0;
  size_t smem_v = (size_t)(ppc + 2 * halo) * sizeof(float4);
  size_t smem_r = (size_t)(max_threads / 32) * p.NVT * taps * CH * 4 * sizeof(float);
  size_t smem = smem_v > smem_r ? smem_v : smem_r;
  if (smem > 200 * 1024) { set_error("colreduce: shared memory %zu too large", smem); return SLIDER_ERR_UNSUPPORTED; }
  const int threads = max_threads;
  dim3 grid(p.B * p.cpi, cblocks);
  for (int j0 = 0; j0 < p.rank; j0 += 4) {
    if (taps == 1) {
      auto k = colreduce_kernel<T, VT, 1, 8>;
      if (smem > 48 * 1024) SLIDER_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, threads, smem, st>>>(out, X, v, p, j0);
    } else {
      auto k = colreduce_kernel<T, VT, 9, 4>;
      if (smem > 48 * 1024) SLIDER_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, threads, smem, st>>>(out, X, v, p, j0);
    }
    int rc = post_launch("colreduce_kernel");
    if (rc) return rc;
  }
  return SLIDER_OK;
}

int run_gemm_plain(const slider_layer_geom* g, int kc, int nn, void* y, const void* act, const void* wgt, void* ws,
                   size_t ws_bytes, cudaStream_t st);   // lora_gemm.cu

template <typename T>
static int lora_grad_impl(const slider_layer_geom* g, float* dA, float* dB, T* dh, const T* gy, const T* x,
                          const float* h, const T* Bup, float s, int accumulate, void* ws, size_t ws_bytes,
                          cudaStream_t st) {
  const int taps = g->kh * g->kw;
  if (!accumulate) {
    SLIDER_CUDA(cudaMemsetAsync(dA, 0, sizeof(float) * (size_t)g->rank * taps * g->cin, st));
    SLIDER_CUDA(cudaMemsetAsync(dB, 0, sizeof(float) * (size_t)g->cout * g->rank, st));
  }
  // dh = gy . (s B): a [pixels, Cout] x [Cout, 16] contraction on the tensor cores (same tcgen05 kernel
  // as the forward, 1x1, no adapter), output already in the zero-padded [pixels, 16] layout.
  const size_t bt_bytes = (((size_t)16 * g->cout * sizeof(T)) + 255) & ~(size_t)255;
  if (ws == nullptr || ws_bytes < bt_bytes) {
    set_error("lora_grad: workspace %zu < %zu bytes", ws_bytes, bt_bytes);
    return SLIDER_ERR_WORKSPACE;
  }
  T* Bt = reinterpret_cast<T*>(ws);
  pack_bt_kernel<T><<<(16 * g->cout + 255) / 256, 256, 0, st>>>(Bt, Bup, g->cout, g->rank, s);
  int rc = post_launch("pack_bt_kernel");
  if (rc) return rc;
  slider_layer_geom g1 = *g;
  g1.kh = g1.kw = 1; g1.pad = 0; g1.rank = 0; g1.cin = g->cout; g1.cout = 16;
  rc = run_gemm_plain(&g1, g->cout, 16, dh, gy, Bt, reinterpret_cast<uint8_t*>(ws) + bt_bytes, ws_bytes - bt_bytes, st);
  if (rc) return rc;
  ColReduceParams pb;
  pb.B = g->batch; pb.H = g->height; pb.W = g->width; pb.C = g->cout; pb.kw = 1; pb.pad = 0;
  pb.rank = g->rank; pb.vstride = g->rank; pb.oc = g->rank; pb.oj = 1; pb.ot = 0; pb.scale = s;
  if ((rc = launch_colreduce<T, float>(dB, gy, h, pb, 1, st))) return rc;
  ColReduceParams pa;
  pa.B = g->batch; pa.H = g->height; pa.W = g->width; pa.C = g->cin; pa.kw = g->kw; pa.pad = g->pad;
  pa.rank = g->rank; pa.vstride = 16; pa.oc = 1; pa.oj = (int64_t)taps * g->cin; pa.ot = g->cin; pa.scale = 1.f;
  return launch_colreduce<T, T>(dA, x, dh, pa, taps, st);
}

}  // namespace slider

using namespace slider;

static int check_geom(const slider_layer_geom* g, const char* who) {
  SLIDER_CHECK_ARG(g != nullptr, "%s: geometry is NULL", who);
  SLIDER_CHECK_ARG(g->dtype == SLIDER_F32 || g->dtype == SLIDER_BF16, "%s: unsupported dtype %d", who, g->dtype);
  SLIDER_CHECK_ARG(g->kh == g->kw && (g->kh == 1 || g->kh == 3) && g->pad == (g->kh - 1) / 2,
                   "%s: only 1x1/pad0 and 3x3/pad1 kernels", who);
  SLIDER_CHECK_ARG(g->batch > 0 && g->height > 0 && g->width > 0 && g->cin > 0 && g->cout > 0, "%s: empty shape", who);
  SLIDER_CHECK_ARG(g->cin % 8 == 0 && g->cout % 8 == 0, "%s: channels must be multiples of 8", who);
  return SLIDER_OK;
}

extern "C" {

size_t slider_lora_grad_workspace_bytes(const slider_layer_geom* g) {
  const size_t bt = g ? ((((size_t)16 * g->cout * 4) + 255) & ~(size_t)255) : 0;
  return bt + slider_gemm_workspace_bytes();
}

int slider_lora_grad(const slider_layer_geom* g, float* dA, float* dB, void* dh_out, const void* gy, const void* x,
                     const float* h, const void* B, float s, int accumulate, void* ws, size_t ws_bytes,
                     void* stream) {
  int rc = check_geom(g, "lora_grad");
  if (rc) return rc;
  SLIDER_CHECK_ARG(dA && dB && dh_out && gy && x && h && B, "lora_grad: null pointer");
  SLIDER_CHECK_ARG(g->rank >= 1 && g->rank <= 16, "lora_grad: rank %d not in 1..16", g->rank);
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    return lora_grad_impl<__nv_bfloat16>(g, dA, dB, (__nv_bfloat16*)dh_out, (const __nv_bfloat16*)gy,
                                         (const __nv_bfloat16*)x, h, (const __nv_bfloat16*)B, s, accumulate, ws, ws_bytes, st);
  return lora_grad_impl<float>(g, dA, dB, (float*)dh_out, (const float*)gy, (const float*)x, h, (const float*)B, s,
                               accumulate, ws, ws_bytes, st);
}

int slider_lora_pack_at(const slider_layer_geom* g, void* At, const void* A, void* stream) {
  int rc = check_geom(g, "lora_pack_at");
  if (rc) return rc;
  SLIDER_CHECK_ARG(At && A && g->rank >= 1 && g->rank <= 16, "lora_pack_at: bad argument");
  const int taps = g->kh * g->kw;
  const int64_t total = (int64_t)g->cin * taps * 16;
  const int grid = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    pack_at_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)At, (const __nv_bfloat16*)A, g->cin, taps, g->rank);
  else
    pack_at_kernel<float><<<grid, 256, 0, st>>>((float*)At, (const float*)A, g->cin, taps, g->rank);
  return post_launch("pack_at_kernel");
}

int slider_pack_wt(const slider_layer_geom* g, void* Wt, const void* W, void* stream) {
  int rc = check_geom(g, "pack_wt");
  if (rc) return rc;
  SLIDER_CHECK_ARG(Wt && W, "pack_wt: null pointer");
  const int taps = g->kh * g->kw;
  dim3 grid((g->cin + 31) / 32, (g->cout + 31) / 32, taps), block(32, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    pack_wt_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((__nv_bfloat16*)Wt, (const __nv_bfloat16*)W, g->cin, g->cout, taps);
  else
    pack_wt_kernel<float><<<grid, block, 0, st>>>((float*)Wt, (const float*)W, g->cin, g->cout, taps);
  return post_launch("pack_wt_kernel");
}

}  // extern "C"
