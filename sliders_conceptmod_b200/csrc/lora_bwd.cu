// Low-rank side of the backward pass (SIMT, bandwidth-class work):
//   rowdot    dh[p, j]     = s * sum_n gy[p, n] * B[n, j]                (warp per pixel)
//   colreduce dB[n, j]     = s * sum_p gy[p, n] * h[p, j]
//             dA[j, t, c]  =     sum_p x[p + d(t), c] * dh[p, j]
//   pack_at / pack_wt      operand re-layouts for the input-gradient GEMM.
// The contraction over pixels is split over CTAs and finished with fp32 atomics.
#include "common.cuh"

namespace slider {

// ------------------------------------------------------------------ rowdot
template <typename T, int R4>
__global__ void __launch_bounds__(256)
rowdot_kernel(T* __restrict__ dh, const T* __restrict__ gy, const T* __restrict__ Bup, int64_t M, int N,
              int rank, float s) {
  constexpr int V = Elem<T>::kVec;
  constexpr int RP = 4 * R4;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  float acc[RP];
#pragma unroll
  for (int j = 0; j < RP; ++j) acc[j] = 0.f;
  const T* grow = gy + row * (int64_t)N;
  for (int n = lane * V; n < N; n += 32 * V) {
    float gv[V];
    Elem<T>::load(grow + n, gv);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const T* b = Bup + (int64_t)(n + k) * rank;
#pragma unroll
      for (int j = 0; j < RP; ++j)
        if (j < rank) acc[j] = fmaf(gv[k], Elem<T>::to_f(b[j]), acc[j]);
    }
  }
  float mine = 0.f;
#pragma unroll
  for (int j = 0; j < RP; ++j) {
    const float t = warp_sum(acc[j]);
    if (lane == j) mine = t * s;
  }
  if (lane < 16) dh[row * 16 + lane] = Elem<T>::from_f(lane < rank ? mine : 0.f);
}

// --------------------------------------------------------------- colreduce
// out[c, j, t] (generic strides) (+)= scale * sum_p X[p + d(t), c] * v[p, j]
// Each thread owns 2 adjacent channels; 64 channel pairs x 4 row lanes per CTA.
struct ColReduceParams {
  int B, H, W, C;        // X is [B,H,W,C]
  int taps, kw, pad;
  int rank, vstride;     // v is [pixels, vstride]
  int rows_per_cta;
  int64_t oc, oj, ot;    // output strides for (channel, rank index, tap)
  float scale;
};

template <typename T, typename VT, int TAPS>
__global__ void __launch_bounds__(256)
colreduce_kernel(float* __restrict__ out, const T* __restrict__ X, const VT* __restrict__ v,
                 const ColReduceParams p, int j0) {
  const int cp = threadIdx.x & 63;   // channel pair within the CTA
  const int rl = threadIdx.x >> 6;   // row lane 0..3
  const int c = (blockIdx.x * 64 + cp) * 2;
  const int64_t P = (int64_t)p.B * p.H * p.W;
  const int64_t r0 = (int64_t)blockIdx.y * p.rows_per_cta;
  int64_t r1 = r0 + p.rows_per_cta;
  if (r1 > P) r1 = P;
  const bool c_ok = c < p.C;
  float acc[TAPS][2][4];
#pragma unroll
  for (int t = 0; t < TAPS; ++t)
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[t][k][j] = 0.f;

  for (int64_t q = r0 + rl; q < r1; q += 4) {
    float x0 = 0.f, x1 = 0.f;
    if (c_ok) {
      x0 = Elem<T>::to_f(X[q * p.C + c]);
      x1 = Elem<T>::to_f(X[q * p.C + c + 1]);
    }
    int w = 0, h = 0, b = 0;
    if (TAPS > 1) {
      w = (int)(q % p.W);
      const int64_t t2 = q / p.W;
      h = (int)(t2 % p.H);
      b = (int)(t2 / p.H);
    }
#pragma unroll
    for (int t = 0; t < TAPS; ++t) {
      int64_t pv = q;
      bool ok = true;
      if (TAPS > 1) {
        // x[q] meets dh[p] where q = p + d(t)  =>  p = q - d(t)
        const int ph = h - (t / p.kw - p.pad), pw = w - (t % p.kw - p.pad);
        ok = ph >= 0 && ph < p.H && pw >= 0 && pw < p.W;
        pv = ((int64_t)b * p.H + ph) * p.W + pw;
      }
      if (ok) {
        const VT* vp = v + pv * p.vstride + j0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (j0 + j < p.rank) {
            const float vv = Elem<VT>::to_f(vp[j]);
            acc[t][0][j] = fmaf(x0, vv, acc[t][0][j]);
            acc[t][1][j] = fmaf(x1, vv, acc[t][1][j]);
          }
        }
      }
    }
  }
  // reduce the 4 row lanes through shared memory, then one atomic per output element
  __shared__ float red[4][64][2 * 4 + 1];
#pragma unroll
  for (int t = 0; t < TAPS; ++t) {
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) red[rl][cp][k * 4 + j] = acc[t][k][j];
    __syncthreads();
    if (rl == 0 && c_ok) {
#pragma unroll
      for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (j0 + j < p.rank) {
            const float sum = red[0][cp][k * 4 + j] + red[1][cp][k * 4 + j] + red[2][cp][k * 4 + j] +
                              red[3][cp][k * 4 + j];
            atomicAdd(out + (int64_t)(c + k) * p.oc + (int64_t)(j0 + j) * p.oj + (int64_t)t * p.ot,
                      sum * p.scale);
          }
        }
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
static int launch_colreduce(float* out, const T* X, const VT* v, ColReduceParams p, cudaStream_t st) {
  const int64_t P = (int64_t)p.B * p.H * p.W;
  const int gx = (p.C + 127) / 128;
  int64_t want = ((int64_t)sm_count() * 4 + gx - 1) / gx;
  int64_t rows = (P + want - 1) / want;
  if (rows < 32) rows = 32;
  rows = (rows + 3) / 4 * 4;
  p.rows_per_cta = (int)rows;
  const int gy = (int)((P + rows - 1) / rows);
  dim3 grid(gx, gy);
  for (int j0 = 0; j0 < p.rank; j0 += 4) {
    if (p.taps == 1) colreduce_kernel<T, VT, 1><<<grid, 256, 0, st>>>(out, X, v, p, j0);
    else colreduce_kernel<T, VT, 9><<<grid, 256, 0, st>>>(out, X, v, p, j0);
    int rc = post_launch("colreduce_kernel");
    if (rc) return rc;
  }
  return SLIDER_OK;
}

template <typename T>
static int lora_grad_impl(const slider_layer_geom* g, float* dA, float* dB, T* dh, const T* gy, const T* x,
                          const float* h, const T* Bup, float s, int accumulate, cudaStream_t st) {
  const int taps = g->kh * g->kw;
  const int64_t M = (int64_t)g->batch * g->height * g->width;
  if (!accumulate) {
    SLIDER_CUDA(cudaMemsetAsync(dA, 0, sizeof(float) * (size_t)g->rank * taps * g->cin, st));
    SLIDER_CUDA(cudaMemsetAsync(dB, 0, sizeof(float) * (size_t)g->cout * g->rank, st));
  }
  const int grid = (int)((M + 7) / 8);
  if (g->rank <= 4) rowdot_kernel<T, 1><<<grid, 256, 0, st>>>(dh, gy, Bup, M, g->cout, g->rank, s);
  else if (g->rank <= 8) rowdot_kernel<T, 2><<<grid, 256, 0, st>>>(dh, gy, Bup, M, g->cout, g->rank, s);
  else rowdot_kernel<T, 4><<<grid, 256, 0, st>>>(dh, gy, Bup, M, g->cout, g->rank, s);
  int rc = post_launch("rowdot_kernel");
  if (rc) return rc;
  ColReduceParams pb;
  pb.B = 1; pb.H = 1; pb.W = (int)M; pb.C = g->cout; pb.taps = 1; pb.kw = 1; pb.pad = 0;
  pb.rank = g->rank; pb.vstride = g->rank; pb.oc = g->rank; pb.oj = 1; pb.ot = 0; pb.scale = s;
  if ((rc = launch_colreduce<T, float>(dB, gy, h, pb, st))) return rc;
  ColReduceParams pa;
  pa.B = g->batch; pa.H = g->height; pa.W = g->width; pa.C = g->cin; pa.taps = taps; pa.kw = g->kw; pa.pad = g->pad;
  pa.rank = g->rank; pa.vstride = 16; pa.oc = 1; pa.oj = (int64_t)taps * g->cin; pa.ot = g->cin; pa.scale = 1.f;
  return launch_colreduce<T, T>(dA, x, dh, pa, st);
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

size_t slider_lora_grad_workspace_bytes(const slider_layer_geom* g) { (void)g; return 0; }

int slider_lora_grad(const slider_layer_geom* g, float* dA, float* dB, void* dh_out, const void* gy, const void* x,
                     const float* h, const void* B, float s, int accumulate, void* ws, size_t ws_bytes,
                     void* stream) {
  (void)ws; (void)ws_bytes;
  int rc = check_geom(g, "lora_grad");
  if (rc) return rc;
  SLIDER_CHECK_ARG(dA && dB && dh_out && gy && x && h && B, "lora_grad: null pointer");
  SLIDER_CHECK_ARG(g->rank >= 1 && g->rank <= 16, "lora_grad: rank %d not in 1..16", g->rank);
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    return lora_grad_impl<__nv_bfloat16>(g, dA, dB, (__nv_bfloat16*)dh_out, (const __nv_bfloat16*)gy,
                                         (const __nv_bfloat16*)x, h, (const __nv_bfloat16*)B, s, accumulate, st);
  return lora_grad_impl<float>(g, dA, dB, (float*)dh_out, (const float*)gy, (const float*)x, h, (const float*)B, s,
                               accumulate, st);
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
