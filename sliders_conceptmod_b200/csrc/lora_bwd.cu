// Backward of the adapter layer (autograd of lora.py:134-138).
//
//   1x1 / Linear layers (100 of the 150 wrapped layers of SD-1.4):  slider_lora_backward_1x1
//     ONE launch of the forward tcgen05 kernel on (x := gy, W := Wt, lora_down := B^T, lora_up := A^T) gives
//         dx = gy . W + s * (gy . B) . A      and      g = gy . B   (fp32 [M, r], the kernel's h output)
//     and ONE reduction launch gives both   dB = s * gy^T . h   and   dA = s * g^T . x.
//   3x3 layers:  slider_lora_grad (dh = s * gy . B on the tensor cores, then the two reductions) followed by
//     slider_lora_input_grad (extended-K GEMM).
//
// The reductions over pixels (colreduce_kernel) are bandwidth-class work and DETERMINISTIC: every CTA writes its
// partial sums to the workspace, and the last CTA of a channel block (ticket) adds them up in chunk order -- no
// floating-point atomics, so two runs give bit-identical LoRA gradients.
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

// Batched form for a whole network: entry e transposes src [rows, cols] into dst [cols, rows] (the B^T operands of
// every 1x1 layer's fused backward launch, rebuilt once per optimizer step from the flat parameter buffer).
template <typename T>
__global__ void transpose_batch_kernel(const slider_transpose_entry* __restrict__ tab, int entries) {
  const int e = blockIdx.y;
  if (e >= entries) return;
  const slider_transpose_entry en = tab[e];
  const T* __restrict__ src = reinterpret_cast<const T*>(en.src);
  T* __restrict__ dst = reinterpret_cast<T*>(en.dst);
  const int total = en.rows * en.cols;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i / en.rows, r = i - c * en.rows;      // dst index i = c * rows + r
    dst[i] = src[(size_t)r * en.cols + c];
  }
}

// --------------------------------------------------------------- colreduce
// out[c, j, t] (generic strides) (+)= scale * sum_p X[stride*p + d(t), c] * V[p, j0 + j],  j = 0..3
// HBM-bound on X (gy for dB, x for dA), which is read exactly once with vector loads:
//   * a thread owns CH adjacent channels (16-byte loads for 1x1, 8/16-byte for 3x3) and keeps
//     TAPS x CH x 4 fp32 accumulators; PL pixel lanes share a CTA;
//   * the V rows (h / dh) the CTA's pixel range meets (+ halo for 3x3) are staged once in shared memory
//     as float4, so the inner loop is one broadcast LDS.128 per tap;
//   * a CTA owns a narrow channel slice (4 vector threads) and 64 pixel lanes; lanes are reduced by warp
//     shuffles, warps through shared memory; the CTA's sums go to the workspace and the LAST CTA of the channel
//     block (ticket) adds all chunks' sums in chunk order and updates `out` (deterministic, no float atomics).
// One launch may carry TWO problems over the same pixels (1x1 layers: dB from (gy, h) and dA from (x, g)):
// blockIdx.y selects the problem and its channel block.
struct ColProblem {
  const void* X;         // [B,H,W,C]
  const void* V;         // [B*VH*VW, vstride]
  float* out;
  int C, cblocks;        // channels; channel blocks (of 4 * CH channels) of this problem
  int vstride;
  int64_t oc, oj, ot;    // output strides for (channel, rank index, tap)
  float scale;
  const float* cscale;   // optional per-channel factor on top of `scale` (DoRA: dA_i *= s * c_i), or NULL
};
struct ColReduceParams {
  int B, H, W;           // X pixels
  int VH, VW, stride;    // V pixels; X pixel = stride * V pixel + d(tap)
  int kw, pad;
  int rank;
  int PL;                // pixel lanes
  int cpi;               // chunks per image
  int ppc;               // X pixels per chunk
  int vmax;              // V rows the staging area holds
  int accumulate;        // out += result instead of out = result
  int nprob;
  ColProblem prob[2];
  float* partial;        // [channel block][chunk][4 * NACC]
  unsigned* tickets;     // [channel block], zero on entry and on exit
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

constexpr int CR_NVT = 4;          // channel-vector threads per pixel lane (the shuffle reduction relies on it)
constexpr int CR_MAX_TICKETS = 1024;

template <typename T, typename VT, int TAPS, int CH>
__global__ void __launch_bounds__(256)
colreduce_kernel(const ColReduceParams p, int j0) {
  extern __shared__ __align__(16) float cr_smem[];
  __shared__ int s_last;
  const int pi = (p.nprob > 1 && (int)blockIdx.y >= p.prob[0].cblocks) ? 1 : 0;
  const ColProblem& pr = p.prob[pi];
  const int yblk = (int)blockIdx.y - (pi ? p.prob[0].cblocks : 0);
  const T* __restrict__ X = reinterpret_cast<const T*>(pr.X);
  const VT* __restrict__ V = reinterpret_cast<const VT*>(pr.V);
  const int HW = p.H * p.W, VHW = p.VH * p.VW;
  const int img = blockIdx.x / p.cpi, chunk = blockIdx.x - img * p.cpi;
  const int q0 = chunk * p.ppc;
  const int q1 = min(HW, q0 + p.ppc);
  // V pixels staged: [v0, v1) of image img -- every V pixel an X pixel of [q0, q1) can meet through a tap
  int v0, v1;
  if (TAPS == 1) { v0 = q0; v1 = q1; }
  else if (p.stride == 1) { const int halo = p.pad * (p.W + 1); v0 = max(0, q0 - halo); v1 = min(HW, q1 + halo); }
  else {
    const int h_lo = q0 / p.W, h_hi = (q1 - 1) / p.W;
    v0 = (max(0, h_lo - p.pad) / p.stride) * p.VW;
    v1 = min(p.VH, (h_hi + p.pad) / p.stride + 1) * p.VW;
  }
  float4* s_v = reinterpret_cast<float4*>(cr_smem);
  for (int i = threadIdx.x; i < v1 - v0; i += blockDim.x) {
    const VT* vp = V + ((size_t)img * VHW + v0 + i) * pr.vstride + j0;
    float4 f;
    f.x = Elem<VT>::to_f(vp[0]);
    f.y = (j0 + 1 < p.rank) ? Elem<VT>::to_f(vp[1]) : 0.f;
    f.z = (j0 + 2 < p.rank) ? Elem<VT>::to_f(vp[2]) : 0.f;
    f.w = (j0 + 3 < p.rank) ? Elem<VT>::to_f(vp[3]) : 0.f;
    s_v[i] = f;
  }
  __syncthreads();
  const int pl = threadIdx.x / CR_NVT, v = threadIdx.x - pl * CR_NVT;
  const int c = (yblk * CR_NVT + v) * CH;
  const bool active = pl < p.PL && c < pr.C;
  float acc[TAPS][CH][4];
#pragma unroll
  for (int t = 0; t < TAPS; ++t)
#pragma unroll
    for (int k = 0; k < CH; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[t][k][j] = 0.f;
  if (active) {
    const T* xb = X + ((size_t)img * HW) * pr.C + c;
    // software pipeline: the X vectors of the next two pixels are in flight while this one is consumed
    float xn0[CH], xn1[CH];
    int q = q0 + pl;
    if (q < q1) ChanVec<T, CH>::load(xb + (size_t)q * pr.C, xn0);
    if (q + p.PL < q1) ChanVec<T, CH>::load(xb + (size_t)(q + p.PL) * pr.C, xn1);
    for (; q < q1; q += p.PL) {
      float xv[CH];
#pragma unroll
      for (int k = 0; k < CH; ++k) { xv[k] = xn0[k]; xn0[k] = xn1[k]; }
      if (q + 2 * p.PL < q1) ChanVec<T, CH>::load(xb + (size_t)(q + 2 * p.PL) * pr.C, xn1);
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
          // x[q] meets V[pv] where q = stride * pv + d(t)  =>  stride * pv = q - d(t)
          int ph = h - (t / 3 - p.pad), pw = w - (t % 3 - p.pad);
          bool ok = ph >= 0 && pw >= 0;
          if (p.stride == 2) { ok = ok && ((ph | pw) & 1) == 0; ph >>= 1; pw >>= 1; }
          if (ok && ph < p.VH && pw < p.VW) {
            const float4 f = s_v[ph * p.VW + pw - v0];
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
  // cross-lane reduction: the 8 pixel lanes of a warp by shuffles (lane = pixel-lane * 4 + channel-vector thread),
  // the warps through shared memory (re-using the V staging area)
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
  // this CTA's sums -> workspace; the last CTA of the channel block finishes in chunk order
  const int nchunks = (int)gridDim.x;
  float* mine = p.partial + ((size_t)blockIdx.y * nchunks + blockIdx.x) * (4 * NACC);
  for (int i = threadIdx.x; i < 4 * NACC; i += blockDim.x) {
    float sum = 0.f;
    for (int w = 0; w < nwarps; ++w) sum += s_red[(size_t)w * 4 * NACC + i];
    __stcg(mine + i, sum);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(p.tickets + blockIdx.y, 1u) == (unsigned)(nchunks - 1)) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const float* all = p.partial + (size_t)blockIdx.y * nchunks * (4 * NACC);
  for (int i = threadIdx.x; i < 4 * NACC; i += blockDim.x) {
    const int vv = i / NACC, rem = i - vv * NACC;
    const int t = rem / (CH * 4), k = (rem >> 2) % CH, j = rem & 3;
    const int cc = (yblk * 4 + vv) * CH + k;
    if (cc < pr.C && j0 + j < p.rank) {
      // chunk order (deterministic); eight loads in flight at a time -- one L2 round trip per chunk made this tail
      // the longest part of the 9-tap launches
      float sum = 0.f;
      for (int ch0 = 0; ch0 < nchunks; ch0 += 8) {
        float t[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) t[u] = ch0 + u < nchunks ? __ldcg(all + (size_t)(ch0 + u) * (4 * NACC) + i) : 0.f;
#pragma unroll
        for (int u = 0; u < 8; ++u) sum += t[u];
      }
      float* o = pr.out + (int64_t)cc * pr.oc + (int64_t)(j0 + j) * pr.oj + (int64_t)t * pr.ot;
      *o = (p.accumulate ? *o : 0.f) + sum * pr.scale * (pr.cscale != nullptr ? pr.cscale[cc] : 1.f);
    }
  }
  if (threadIdx.x == 0) p.tickets[blockIdx.y] = 0u;
}

// --------------------------------------------------------------- colreduce on the tensor cores (bf16 storage)
// The same contraction as a GEMM over the pixels:   out[c, j, t] = scale * sum_q X[q, c] * Vs[(t, j), q],
//   Vs[(t, j), q] = V[p(q, t), j]   (the V pixel the X pixel q meets through tap t; 0 where the tap leaves the image)
// so the taps shift the SMALL operand (rank columns) and X is read exactly once.  The SIMT form above keeps
// TAPS x CH x 4 accumulators per thread and spends most of a 3x3 launch reducing them across 64 pixel lanes (33 us per
// launch in the SD-1.4 train step, r02 launch list); here a warp owns 8 channels over the CTA's whole pixel chunk, so
// nothing is reduced inside the CTA at all.
//   CTA = (chunk of spc x 128 pixels of one image) x (64-channel slice), 8 warps, mma.sync m16n8k16 (bf16, fp32 accum)
//   A (shared): [MT * 16 rows = (tap, j)][128 pixels] bf16, row stride 272 bytes  -> ldmatrix.x4
//   B (shared): [128 pixels][64 channels] bf16, row stride 144 bytes              -> ldmatrix.x2.trans
// The padded strides keep both ldmatrix patterns conflict-free.  V is rounded to bf16 (what the reference's autograd in
// bf16 does with the lora_down output); fp32 storage keeps the SIMT kernel.  Cross-chunk reduction as above: per-CTA
// partials + ticket + the last CTA adds in chunk order (deterministic).
constexpr int TC_PIX = 128;                 // pixels per staged sub-chunk
constexpr int TC_CH = 64;                   // channels per CTA
constexpr int TC_A_STRIDE = TC_PIX * 2 + 16;
constexpr int TC_B_STRIDE = TC_CH * 2 + 16;

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t addr, uint32_t (&r)[2]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <typename VT, int MT, int TAPS>
__global__ void __launch_bounds__(256)
tc_colreduce_kernel(const ColReduceParams p) {
  using T = __nv_bfloat16;
  extern __shared__ __align__(16) unsigned char tc_smem[];
  __shared__ int s_last;
  unsigned char* sA = tc_smem;                                   // [MT * 16][TC_A_STRIDE]
  unsigned char* sB = tc_smem + (size_t)MT * 16 * TC_A_STRIDE;   // [TC_PIX][TC_B_STRIDE]
  pdl_trigger();
  const int pi = (p.nprob > 1 && (int)blockIdx.y >= p.prob[0].cblocks) ? 1 : 0;
  const ColProblem& pr = p.prob[pi];
  const int yblk = (int)blockIdx.y - (pi ? p.prob[0].cblocks : 0);
  const T* __restrict__ X = reinterpret_cast<const T*>(pr.X);
  const VT* __restrict__ V = reinterpret_cast<const VT*>(pr.V);
  const int HW = p.H * p.W, VHW = p.VH * p.VW;
  const int img = blockIdx.x / p.cpi, chunk = blockIdx.x - img * p.cpi;
  const int q0 = chunk * p.ppc, q1 = min(HW, q0 + p.ppc);
  const int c0 = yblk * TC_CH;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int R = p.rank, rows = TAPS * R;                          // live rows of A: m = t * R + j
  float acc[MT][4];
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[m][k] = 0.f;
  // rows >= TAPS * rank never change: zero once
  for (int i = tid; i < (MT * 16 - rows) * (TC_PIX / 2); i += 256) {
    const int m = rows + i / (TC_PIX / 2), w = i - (i / (TC_PIX / 2)) * (TC_PIX / 2);
    *reinterpret_cast<uint32_t*>(sA + (size_t)m * TC_A_STRIDE + w * 4) = 0u;
  }
  pdl_wait();
  const uint32_t sA_u = (uint32_t)__cvta_generic_to_shared(sA), sB_u = (uint32_t)__cvta_generic_to_shared(sB);
  constexpr int B_ITEMS = TC_PIX * (TC_CH / 8) / 256;            // 16-byte X vectors per thread and sub-chunk (4)
  constexpr int A_ITEMS = (TC_PIX * TAPS + 255) / 256;           // (tap, pixel) items per thread and sub-chunk (5 / 1)
  const bool vec4 = R == 4 && (pr.vstride & 3) == 0 && ((uintptr_t)V & 15) == 0;
  for (int s0 = q0; s0 < q1; s0 += TC_PIX) {
    // every global load of the sub-chunk is issued before the first shared-memory store: ONE memory round trip
    // ---- B: X[s0 .. s0+127, c0 .. c0+63] (16-byte vectors; rows past the chunk and channels past C are zero)
    uint4 xb[B_ITEMS];
#pragma unroll
    for (int u = 0; u < B_ITEMS; ++u) {
      const int i = tid + u * 256;
      const int r = i >> 3, v = i & 7;
      const int q = s0 + r, c = c0 + v * 8;
      xb[u] = make_uint4(0u, 0u, 0u, 0u);
      if (q < q1 && c < pr.C) xb[u] = *reinterpret_cast<const uint4*>(X + ((size_t)img * HW + q) * pr.C + c);
    }
    // ---- A: Vs[(t, j), q] for the 128 pixels
    int apv[A_ITEMS];
    float av[A_ITEMS][4];
#pragma unroll
    for (int u = 0; u < A_ITEMS; ++u) {
      const int i = tid + u * 256;
      const int t = i / TC_PIX, r = i - t * TC_PIX;
      const int q = s0 + r;
      int pv = -1;
      if (i < TC_PIX * TAPS && q < q1) {
        if (TAPS == 1) pv = q;
        else {
          const int h = q / p.W, w = q - h * p.W;
          int ph = h - (t / 3 - p.pad), pw = w - (t % 3 - p.pad);
          bool ok = ph >= 0 && pw >= 0;
          if (p.stride == 2) { ok = ok && ((ph | pw) & 1) == 0; ph >>= 1; pw >>= 1; }
          if (ok && ph < p.VH && pw < p.VW) pv = ph * p.VW + pw;
        }
      }
      apv[u] = pv;
      av[u][0] = av[u][1] = av[u][2] = av[u][3] = 0.f;
      if (vec4 && pv >= 0) {
        const VT* vp = V + ((size_t)img * VHW + pv) * pr.vstride;
        if (sizeof(VT) == 4) {
          const float4 f = *reinterpret_cast<const float4*>(vp);
          av[u][0] = f.x; av[u][1] = f.y; av[u][2] = f.z; av[u][3] = f.w;
        } else {
          const uint2 w2 = *reinterpret_cast<const uint2*>(vp);
          av[u][0] = __uint_as_float(w2.x << 16); av[u][1] = __uint_as_float(w2.x & 0xffff0000u);
          av[u][2] = __uint_as_float(w2.y << 16); av[u][3] = __uint_as_float(w2.y & 0xffff0000u);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < B_ITEMS; ++u) {
      const int i = tid + u * 256;
      *reinterpret_cast<uint4*>(sB + (size_t)(i >> 3) * TC_B_STRIDE + (i & 7) * 16) = xb[u];
    }
#pragma unroll
    for (int u = 0; u < A_ITEMS; ++u) {
      const int i = tid + u * 256;
      if (i >= TC_PIX * TAPS) break;
      const int t = i / TC_PIX, r = i - t * TC_PIX;
      if (vec4) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<T*>(sA + (size_t)(t * 4 + j) * TC_A_STRIDE + r * 2) = __float2bfloat16_rn(av[u][j]);
      } else {
        const int pv = apv[u];
        const VT* vp = V + ((size_t)img * VHW + (pv < 0 ? 0 : pv)) * pr.vstride;
        for (int j = 0; j < R; ++j) {
          const float f = pv < 0 ? 0.f : Elem<VT>::to_f(vp[j]);
          *reinterpret_cast<T*>(sA + (size_t)(t * R + j) * TC_A_STRIDE + r * 2) = __float2bfloat16_rn(f);
        }
      }
    }
    __syncthreads();
#pragma unroll 2
    for (int ks = 0; ks < TC_PIX / 16; ++ks) {
      uint32_t b[2];
      ldmatrix_x2_trans(sB_u + (uint32_t)((ks * 16 + (lane & 15)) * TC_B_STRIDE + warp * 16), b);
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        uint32_t a[4];
        ldmatrix_x4(sA_u + (uint32_t)((m * 16 + (lane & 15)) * TC_A_STRIDE + (ks * 16 + (lane >> 4) * 8) * 2), a);
        mma_bf16_16816(acc[m], a, b);
      }
    }
    __syncthreads();
  }
  // ---- this CTA's sums -> workspace [channel block][chunk][MT * 16 rows][64 channels]
  const int rows8 = (rows + 7) & ~7;                 // only the live rows (rounded to the fragment's 8) travel
  const int NOUT = rows8 * TC_CH;
  const int nchunks = (int)gridDim.x;
  float* mine = p.partial + ((size_t)blockIdx.y * nchunks + blockIdx.x) * NOUT;
  {
    const int g = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      const int col = warp * 8 + 2 * tq;
      if (m * 16 + g < rows8)
        __stcg(reinterpret_cast<float2*>(mine + (size_t)(m * 16 + g) * TC_CH + col), make_float2(acc[m][0], acc[m][1]));
      if (m * 16 + g + 8 < rows8)
        __stcg(reinterpret_cast<float2*>(mine + (size_t)(m * 16 + g + 8) * TC_CH + col), make_float2(acc[m][2], acc[m][3]));
    }
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(p.tickets + blockIdx.y, 1u) == (unsigned)(nchunks - 1)) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // Fixed summation order per output (deterministic).  Few outputs (rank rows of a 1x1 problem: 64 pieces): the chunks
  // are dealt to `groups` thread groups (chunk c -> group c % groups, added in increasing c), whose sums meet in shared
  // memory in group order.  Many outputs: a thread owns up to three 16-byte pieces at a time.  Either way eight chunk
  // loads per piece are in flight, so the tail is a handful of L2 round trips.
  const float4* all = reinterpret_cast<const float4*>(p.partial + (size_t)blockIdx.y * nchunks * NOUT);
  const int nq = rows * TC_CH / 4;
  auto emit = [&](int idx, const float4& sm) {
    const int i = idx * 4;
    const int m = i / TC_CH, cl = i - m * TC_CH;
    const int t = m / R, j = m - t * R;
    const float r4[4] = {sm.x, sm.y, sm.z, sm.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int cc = c0 + cl + e;
      if (cc >= pr.C) break;
      float* o = pr.out + (int64_t)cc * pr.oc + (int64_t)j * pr.oj + (int64_t)t * pr.ot;
      *o = (p.accumulate ? *o : 0.f) + r4[e] * pr.scale * (pr.cscale != nullptr ? pr.cscale[cc] : 1.f);
    }
  };
  if (nq <= 128) {
    const int groups = nq <= 64 ? 4 : 2;
    const int idx = tid % nq, cg = tid / nq;
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
    if (cg < groups) {
      for (int ch0 = cg; ch0 < nchunks; ch0 += 8 * groups) {
        float4 v8[8];
#pragma unroll
        for (int w = 0; w < 8; ++w)
          v8[w] = ch0 + w * groups < nchunks ? __ldcg(all + (size_t)(ch0 + w * groups) * (NOUT / 4) + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int w = 0; w < 8; ++w) { sum.x += v8[w].x; sum.y += v8[w].y; sum.z += v8[w].z; sum.w += v8[w].w; }
      }
    }
    float4* s_red = reinterpret_cast<float4*>(tc_smem);      // the operand tiles are dead: [groups][nq]
    if (cg < groups) s_red[cg * nq + idx] = sum;
    __syncthreads();
    if (cg == 0) {
      for (int g = 1; g < groups; ++g) { const float4 o = s_red[g * nq + idx]; sum.x += o.x; sum.y += o.y; sum.z += o.z; sum.w += o.w; }
      emit(idx, sum);
    }
  } else {
    for (int base = tid; base < nq; base += 3 * 256) {
      float4 sum[3];
#pragma unroll
      for (int u = 0; u < 3; ++u) sum[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int ch0 = 0; ch0 < nchunks; ch0 += 8) {
        float4 v8[3][8];
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
          for (int w = 0; w < 8; ++w) {
            const int idx = base + u * 256;
            v8[u][w] = (idx < nq && ch0 + w < nchunks) ? __ldcg(all + (size_t)(ch0 + w) * (NOUT / 4) + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
          for (int w = 0; w < 8; ++w) { sum[u].x += v8[u][w].x; sum[u].y += v8[u][w].y; sum[u].z += v8[u][w].z; sum[u].w += v8[u][w].w; }
      }
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int idx = base + u * 256;
        if (idx >= nq) break;
        emit(idx, sum[u]);
      }
    }
  }
  if (tid == 0) p.tickets[blockIdx.y] = 0u;
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

// Workspace layout of the gradient entry points (caller zero-fills the first CR_TICKET_BYTES once; the kernels
// leave them zero):  [tickets 4 KB][Bt 16 x cout][colreduce partials][split-K scratch of the GEMM]
constexpr size_t CR_TICKET_BYTES = CR_MAX_TICKETS * sizeof(unsigned);
constexpr size_t CR_PARTIAL_BYTES = (size_t)2 << 20;
static inline size_t bt_bytes_of(int cout) { return (((size_t)16 * cout * 4) + 255) & ~(size_t)255; }

// bf16 storage: the tensor-core form (SLIDER_COLREDUCE_TC=0: the SIMT kernel, for A/B measurements)
static int g_colreduce_tc = [] { const char* e = getenv("SLIDER_COLREDUCE_TC"); return e ? atoi(e) : 1; }();

template <typename VT>
static int launch_tc_colreduce(ColReduceParams p, int taps, float* partial, unsigned* tickets, cudaStream_t st) {
  const int HW = p.H * p.W;
  const int rows = taps * p.rank;
  const int mt = (rows + 15) / 16;
  int cblocks = 0;
  for (int i = 0; i < p.nprob; ++i) {
    p.prob[i].cblocks = (p.prob[i].C + TC_CH - 1) / TC_CH;
    cblocks += p.prob[i].cblocks;
  }
  if (cblocks > CR_MAX_TICKETS) { set_error("colreduce: %d channel blocks exceed %d", cblocks, CR_MAX_TICKETS); return SLIDER_ERR_UNSUPPORTED; }
  const int MT = taps == 1 ? 1 : (mt <= 1 ? 1 : (mt <= 2 ? 2 : (mt <= 3 ? 3 : (mt <= 5 ? 5 : 9))));
  const size_t nout_bytes = (size_t)((rows + 7) & ~7) * TC_CH * sizeof(float);      // live rows only (kernel: rows8)
  // pixel chunks of spc x 128 pixels: ~2 CTAs per SM over (images x chunks x channel blocks), bounded by the scratch
  const int nsub = (HW + TC_PIX - 1) / TC_PIX;
  // (3x3: two CTAs per SM; 1x1: four -- one staged sub-chunk per CTA where it fits, the round trips of co-resident CTAs
  // overlap and the finish deals the chunks over four thread groups)
  int want = (sm_count() * (taps == 1 ? 4 : 2)) / (p.B * cblocks);
  if (want < 1) want = 1;
  int spc = (nsub + want - 1) / want;
  if (spc < 1) spc = 1;
  while ((size_t)cblocks * p.B * ((nsub + spc - 1) / spc) * nout_bytes > CR_PARTIAL_BYTES && spc < nsub) ++spc;
  p.ppc = spc * TC_PIX;
  p.cpi = (HW + p.ppc - 1) / p.ppc;
  if ((size_t)cblocks * p.B * p.cpi * nout_bytes > CR_PARTIAL_BYTES) {
    set_error("colreduce: partial-sum scratch too small for %d x %d CTAs", p.B * p.cpi, cblocks);
    return SLIDER_ERR_WORKSPACE;
  }
  p.PL = 0; p.vmax = 0;
  p.partial = partial; p.tickets = tickets;
  const dim3 grid(p.B * p.cpi, cblocks);
  const size_t smem = (size_t)MT * 16 * TC_A_STRIDE + (size_t)TC_PIX * TC_B_STRIDE;
#define SLIDER_TC_LAUNCH(MTV, TAPSV)                                                                              \
  do {                                                                                                            \
    auto k = tc_colreduce_kernel<VT, MTV, TAPSV>;                                                                 \
    if (smem > 48 * 1024) SLIDER_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    launch_pdl(k, grid, dim3(256), smem, st, p);                                                                  \
  } while (0)
  if (taps == 1) SLIDER_TC_LAUNCH(1, 1);
  else if (MT == 1) SLIDER_TC_LAUNCH(1, 9);
  else if (MT == 2) SLIDER_TC_LAUNCH(2, 9);
  else if (MT == 3) SLIDER_TC_LAUNCH(3, 9);
  else if (MT == 5) SLIDER_TC_LAUNCH(5, 9);
  else SLIDER_TC_LAUNCH(9, 9);
#undef SLIDER_TC_LAUNCH
  return post_launch("tc_colreduce_kernel");
}

// taps: 1 or 9.  One or two problems over the same pixel set.
template <typename T, typename VT>
static int launch_colreduce(ColReduceParams p, int taps, float* partial, unsigned* tickets, cudaStream_t st) {
  if constexpr (sizeof(T) == 2) if (g_colreduce_tc && taps * p.rank <= 144 && p.rank <= 16) {
    bool ok = true;
    for (int i = 0; i < p.nprob; ++i) ok = ok && p.prob[i].C % 8 == 0 && ((uintptr_t)p.prob[i].X & 15) == 0;
    if (ok) return launch_tc_colreduce<VT>(p, taps, partial, tickets, st);
  }
  const int CH = taps == 1 ? 8 : 4;
  const int HW = p.H * p.W;
  // a CTA = 4 channel-vector threads x 64 pixel lanes: a narrow channel slice keeps the per-CTA output small, the
  // pixel lanes supply the parallelism
  const int max_threads = 256;
  int cblocks = 0;
  for (int i = 0; i < p.nprob; ++i) {
    const int nvt_all = (p.prob[i].C + CH - 1) / CH;
    p.prob[i].cblocks = (nvt_all + CR_NVT - 1) / CR_NVT;
    cblocks += p.prob[i].cblocks;
  }
  if (cblocks > CR_MAX_TICKETS) { set_error("colreduce: %d channel blocks exceed %d", cblocks, CR_MAX_TICKETS); return SLIDER_ERR_UNSUPPORTED; }
  p.PL = max_threads / CR_NVT;
  const int nacc4 = 4 * taps * CH * 4;
  // pixel chunks: ~2 CTAs per SM over (images x chunks x channel blocks); chunks never straddle images; bounded by the
  // partial-sum scratch
  int want = (sm_count() * 2) / (p.B * cblocks);
  const int cap = (int)(CR_PARTIAL_BYTES / ((size_t)cblocks * nacc4 * sizeof(float))) / p.B;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  int ppc = (HW + want - 1) / want;
  if (ppc < p.PL * 2) ppc = p.PL * 2;
  if (ppc > HW) ppc = HW;
  p.ppc = ppc;
  p.cpi = (HW + ppc - 1) / ppc;
  if ((size_t)p.B * p.cpi * cblocks * nacc4 * sizeof(float) > CR_PARTIAL_BYTES) {
    set_error("colreduce: partial-sum scratch too small for %d x %d CTAs", p.B * p.cpi, cblocks);
    return SLIDER_ERR_WORKSPACE;
  }
  int vmax;
  if (taps == 1) vmax = ppc;
  else if (p.stride == 1) vmax = ppc + 2 * p.pad * (p.W + 1);
  else vmax = (((ppc + p.W - 1) / p.W + 1 + 2 * p.pad) / p.stride + 2) * p.VW;
  p.vmax = vmax;
  size_t smem_v = (size_t)vmax * sizeof(float4);
  size_t smem_r = (size_t)(max_threads / 32) * nacc4 * sizeof(float);
  size_t smem = smem_v > smem_r ? smem_v : smem_r;
  if (smem > 200 * 1024) { set_error("colreduce: shared memory %zu too large", smem); return SLIDER_ERR_UNSUPPORTED; }
  p.partial = partial; p.tickets = tickets;
  dim3 grid(p.B * p.cpi, cblocks);
  for (int j0 = 0; j0 < p.rank; j0 += 4) {
    if (taps == 1) {
      auto k = colreduce_kernel<T, VT, 1, 8>;
      if (smem > 48 * 1024) SLIDER_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, max_threads, smem, st>>>(p, j0);
    } else {
      auto k = colreduce_kernel<T, VT, 9, 4>;
      if (smem > 48 * 1024) SLIDER_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, max_threads, smem, st>>>(p, j0);
    }
    int rc = post_launch("colreduce_kernel");
    if (rc) return rc;
  }
  return SLIDER_OK;
}

int run_gemm_plain(const slider_layer_geom* g, int kc, int nn, void* y, const void* act, const void* wgt, void* ws,
                   size_t ws_bytes, cudaStream_t st);   // lora_gemm.cu
int run_gemm_fused_bwd(const slider_layer_geom* g, void* dx, float* g_out, const void* gy, const void* Wt,
                       const void* Bt, const void* A, float s, void* ws, size_t ws_bytes, cudaStream_t st);   // lora_gemm.cu

static ColProblem make_problem(const void* X, const void* V, float* out, int C, int vstride, int64_t oc, int64_t oj,
                               int64_t ot, float scale) {
  ColProblem q;
  q.X = X; q.V = V; q.out = out; q.C = C; q.cblocks = 0; q.vstride = vstride; q.oc = oc; q.oj = oj; q.ot = ot; q.scale = scale; q.cscale = nullptr;
  return q;
}

struct GradWs {
  unsigned* tickets; void* bt; float* partial; void* gemm; size_t gemm_bytes;
};
static int carve_ws(GradWs* w, void* ws, size_t ws_bytes, int cout, const char* who) {
  const size_t need = CR_TICKET_BYTES + bt_bytes_of(cout) + CR_PARTIAL_BYTES;
  if (ws == nullptr || ws_bytes < need || ((uintptr_t)ws & 255) != 0) {
    set_error("%s: workspace %zu < %zu bytes (or not 256-byte aligned)", who, ws_bytes, need);
    return SLIDER_ERR_WORKSPACE;
  }
  uint8_t* b = reinterpret_cast<uint8_t*>(ws);
  w->tickets = reinterpret_cast<unsigned*>(b);
  w->bt = b + CR_TICKET_BYTES;
  w->partial = reinterpret_cast<float*>(b + CR_TICKET_BYTES + bt_bytes_of(cout));
  w->gemm = b + need;
  w->gemm_bytes = ws_bytes - need;
  return SLIDER_OK;
}

template <typename T>
static int lora_grad_impl(const slider_layer_geom* g, float* dA, float* dB, T* dh, const T* gy, const T* x,
                          const float* h, const T* Bup, float s, int accumulate, const float* dA_chan_scale, void* ws,
                          size_t ws_bytes, cudaStream_t st) {
  const int taps = g->kh * g->kw;
  const int stride = g->stride == 2 ? 2 : 1;
  const int Ho = (g->height + 2 * g->pad - g->kh) / stride + 1, Wo = (g->width + 2 * g->pad - g->kw) / stride + 1;
  GradWs w;
  int rc = carve_ws(&w, ws, ws_bytes, g->cout, "lora_grad");
  if (rc) return rc;
  // dh = gy . (s B): a [pixels, Cout] x [Cout, 16] contraction on the tensor cores (same tcgen05 kernel
  // as the forward, 1x1, no adapter), output already in the zero-padded [pixels, 16] layout.
  T* Bt = reinterpret_cast<T*>(w.bt);
  pack_bt_kernel<T><<<(16 * g->cout + 255) / 256, 256, 0, st>>>(Bt, Bup, g->cout, g->rank, s);
  if ((rc = post_launch("pack_bt_kernel"))) return rc;
  slider_layer_geom g1 = *g;
  g1.kh = g1.kw = 1; g1.pad = 0; g1.rank = 0; g1.cin = g->cout; g1.cout = 16; g1.stride = 1; g1.height = Ho; g1.width = Wo;
  if ((rc = run_gemm_plain(&g1, g->cout, 16, dh, gy, Bt, w.gemm, w.gemm_bytes, st))) return rc;
  if (dA == nullptr && dB == nullptr) return SLIDER_OK;     // frozen adapter: only dh (for the input gradient)
  ColReduceParams pb;
  pb.B = g->batch; pb.H = Ho; pb.W = Wo; pb.VH = Ho; pb.VW = Wo; pb.stride = 1; pb.kw = 1; pb.pad = 0;
  pb.rank = g->rank; pb.accumulate = accumulate; pb.nprob = 1;
  pb.prob[0] = make_problem(gy, h, dB, g->cout, g->rank, g->rank, 1, 0, s);
  if ((rc = launch_colreduce<T, float>(pb, 1, w.partial, w.tickets, st))) return rc;
  ColReduceParams pa;
  pa.B = g->batch; pa.H = g->height; pa.W = g->width; pa.VH = Ho; pa.VW = Wo; pa.stride = stride; pa.kw = g->kw; pa.pad = g->pad;
  pa.rank = g->rank; pa.accumulate = accumulate; pa.nprob = 1;
  pa.prob[0] = make_problem(x, dh, dA, g->cin, 16, 1, (int64_t)taps * g->cin, g->cin, 1.f);
  pa.prob[0].cscale = dA_chan_scale;
  return launch_colreduce<T, T>(pa, taps, w.partial, w.tickets, st);
}

// 1x1 layers: one fused GEMM launch (dx and g = gy . B), one reduction launch (dB and dA).
template <typename T>
static int lora_backward_1x1_impl(const slider_layer_geom* g, T* dx, float* dA, float* dB, float* gbuf, const T* gy,
                                  const T* x, const float* h, const T* Wt, const T* Bt, const T* A, const T* Bup,
                                  float s, int accumulate, void* ws, size_t ws_bytes, cudaStream_t st) {
  GradWs w;
  int rc = carve_ws(&w, ws, ws_bytes, g->cout, "lora_backward_1x1");
  if (rc) return rc;
  if (Bt == nullptr) {
    // B^T [16, cout] (rows >= rank zero) for this layer alone; a network-wide pack (slider_transpose_batch) avoids it
    T* bt = reinterpret_cast<T*>(w.bt);
    pack_bt_kernel<T><<<(16 * g->cout + 255) / 256, 256, 0, st>>>(bt, Bup, g->cout, g->rank, 1.f);
    if ((rc = post_launch("pack_bt_kernel"))) return rc;
    Bt = bt;
  }
  if ((rc = run_gemm_fused_bwd(g, dx, gbuf, gy, Wt, Bt, A, s, w.gemm, w.gemm_bytes, st))) return rc;
  if (dA == nullptr && dB == nullptr) return SLIDER_OK;
  ColReduceParams p;
  p.B = g->batch; p.H = g->height; p.W = g->width; p.VH = g->height; p.VW = g->width; p.stride = 1; p.kw = 1; p.pad = 0;
  p.rank = g->rank; p.accumulate = accumulate; p.nprob = 2;
  p.prob[0] = make_problem(gy, h, dB, g->cout, g->rank, g->rank, 1, 0, s);
  p.prob[1] = make_problem(x, gbuf, dA, g->cin, g->rank, 1, g->cin, 0, s);
  return launch_colreduce<T, float>(p, 1, w.partial, w.tickets, st);
}

}  // namespace slider

using namespace slider;

static int check_geom(const slider_layer_geom* g, const char* who) {
  SLIDER_CHECK_ARG(g != nullptr, "%s: geometry is NULL", who);
  SLIDER_CHECK_ARG(g->dtype == SLIDER_F32 || g->dtype == SLIDER_BF16, "%s: unsupported dtype %d", who, g->dtype);
  SLIDER_CHECK_ARG(g->kh == g->kw && (g->kh == 1 || g->kh == 3) && g->pad == (g->kh - 1) / 2,
                   "%s: only 1x1/pad0 and 3x3/pad1 kernels", who);
  SLIDER_CHECK_ARG(g->stride >= 0 && g->stride <= 2 && (g->stride != 2 || g->kh == 3), "%s: stride %d", who, g->stride);
  SLIDER_CHECK_ARG(g->batch > 0 && g->height > 0 && g->width > 0 && g->cin > 0 && g->cout > 0, "%s: empty shape", who);
  SLIDER_CHECK_ARG(g->cin % 8 == 0 && g->cout % 8 == 0, "%s: channels must be multiples of 8", who);
  return SLIDER_OK;
}

extern "C" {

size_t slider_lora_grad_workspace_bytes(const slider_layer_geom* g) {
  return CR_TICKET_BYTES + (g ? bt_bytes_of(g->cout) : 0) + CR_PARTIAL_BYTES + slider_gemm_workspace_bytes();
}

int slider_lora_grad(const slider_layer_geom* g, float* dA, float* dB, void* dh_out, const void* gy, const void* x,
                     const float* h, const void* B, float s, int accumulate, void* ws, size_t ws_bytes,
                     void* stream) {
  return slider_lora_grad_ex(g, dA, dB, dh_out, gy, x, h, B, s, accumulate, nullptr, ws, ws_bytes, stream);
}

int slider_lora_grad_ex(const slider_layer_geom* g, float* dA, float* dB, void* dh_out, const void* gy, const void* x,
                        const float* h, const void* B, float s, int accumulate, const float* dA_chan_scale, void* ws,
                        size_t ws_bytes, void* stream) {
  int rc = check_geom(g, "lora_grad");
  if (rc) return rc;
  SLIDER_CHECK_ARG((dA == nullptr) == (dB == nullptr), "lora_grad: dA and dB must both be set or both NULL");
  SLIDER_CHECK_ARG(dh_out && gy && B && (dA == nullptr || (x && h)), "lora_grad: null pointer");
  SLIDER_CHECK_ARG(g->rank >= 1 && g->rank <= 16, "lora_grad: rank %d not in 1..16", g->rank);
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    return lora_grad_impl<__nv_bfloat16>(g, dA, dB, (__nv_bfloat16*)dh_out, (const __nv_bfloat16*)gy,
                                         (const __nv_bfloat16*)x, h, (const __nv_bfloat16*)B, s, accumulate,
                                         dA_chan_scale, ws, ws_bytes, st);
  return lora_grad_impl<float>(g, dA, dB, (float*)dh_out, (const float*)gy, (const float*)x, h, (const float*)B, s,
                               accumulate, dA_chan_scale, ws, ws_bytes, st);
}

int slider_lora_backward_1x1(const slider_layer_geom* g, void* dx, float* dA, float* dB, float* g_buf, const void* gy,
                             const void* x, const float* h, const void* Wt, const void* Bt, const void* A,
                             const void* B, float s, int accumulate, void* ws, size_t ws_bytes, void* stream) {
  int rc = check_geom(g, "lora_backward_1x1");
  if (rc) return rc;
  SLIDER_CHECK_ARG(g->kh == 1 && g->stride != 2, "lora_backward_1x1: 1x1 / Linear layers only");
  SLIDER_CHECK_ARG((dA == nullptr) == (dB == nullptr), "lora_backward_1x1: dA and dB must both be set or both NULL");
  SLIDER_CHECK_ARG(dx && g_buf && gy && Wt && A && (Bt || B) && (dA == nullptr || (x && h)), "lora_backward_1x1: null pointer");
  SLIDER_CHECK_ARG(g->rank >= 1 && g->rank <= 16, "lora_backward_1x1: rank %d not in 1..16", g->rank);
  cudaStream_t st = (cudaStream_t)stream;
  if (g->dtype == SLIDER_BF16)
    return lora_backward_1x1_impl<__nv_bfloat16>(g, (__nv_bfloat16*)dx, dA, dB, g_buf, (const __nv_bfloat16*)gy,
                                                 (const __nv_bfloat16*)x, h, (const __nv_bfloat16*)Wt,
                                                 (const __nv_bfloat16*)Bt, (const __nv_bfloat16*)A,
                                                 (const __nv_bfloat16*)B, s, accumulate, ws, ws_bytes, st);
  return lora_backward_1x1_impl<float>(g, (float*)dx, dA, dB, g_buf, (const float*)gy, (const float*)x, h,
                                       (const float*)Wt, (const float*)Bt, (const float*)A, (const float*)B, s,
                                       accumulate, ws, ws_bytes, st);
}

int slider_transpose_batch(const slider_transpose_entry* table_dev, int entries, int max_elems, int dtype, void* stream) {
  SLIDER_CHECK_ARG(table_dev != nullptr && entries > 0 && entries <= 65535 && max_elems > 0, "transpose_batch: bad argument");
  SLIDER_CHECK_ARG(dtype == SLIDER_F32 || dtype == SLIDER_BF16, "transpose_batch: unsupported dtype %d", dtype);
  int bx = (max_elems + 255) / 256;
  if (bx > 32) bx = 32;
  dim3 grid(bx, entries);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLIDER_BF16) transpose_batch_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(table_dev, entries);
  else transpose_batch_kernel<float><<<grid, 256, 0, st>>>(table_dev, entries);
  return post_launch("transpose_batch_kernel");
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
