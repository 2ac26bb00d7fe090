// NHWC GroupNorm (+ per-(batch,channel) pre-add, + SiLU) for the UNet layers that feed the fused
// LoRA convolutions (ResnetBlock2D: norm1 -> SiLU -> conv1, "+ time embedding" -> norm2 -> SiLU -> conv2).
// HBM-bound: x is read twice (statistics, apply) and y written once; channels are innermost so every
// access is a 16-byte vector and the result is already in the layout the implicit-GEMM kernel loads.
//
//   z = x + chan_add[b, c]                      (optional: the time-embedding projection)
//   y = act((z - mean_g) * rstd_g * gamma[c] + beta[c]),   act = SiLU or identity
//
// Statistics are reduced deterministically: fixed per-thread channel ownership, per-chunk partial sums,
// and every block of the apply pass combines the (L2-resident, few KB) partials of its batch item in a
// fixed order — a same-address ticket atomic per chunk cost 11 us of a 24 us GroupNorm on B200.
#include <stdlib.h>
#include "common.cuh"
#include "ptx.cuh"

namespace slider {

struct GnParams {
  int B, HW, C, G, cpg;
  int NV;          // 8-channel vectors per pixel
  int PL;          // pixel lanes per block (threads = PL * NV <= 512)
  int chunks;      // pixel chunks per batch item (stats grid.x)
  int ppc;         // pixels per chunk
  float eps;
  int silu;
  int variant;
};

constexpr int GN_MAX_THREADS = 512;
static int g_gn_variant = 0;         // profiling aid (slider_debug_gn_variant)
constexpr int GN_TICKETS = 64;      // max batch; ticket slots at the head of the workspace

template <typename T> struct Vec8;
template <> struct Vec8<__nv_bfloat16> {
  __device__ static inline void load(const __nv_bfloat16* p, float (&v)[8]) { Elem<__nv_bfloat16>::load(p, v); }
  __device__ static inline void store(__nv_bfloat16* p, const float (&v)[8]) { Elem<__nv_bfloat16>::store(p, v); }
};
template <> struct Vec8<__half> {
  __device__ static inline void load(const __half* p, float (&v)[8]) { Elem<__half>::load(p, v); }
  __device__ static inline void store(__half* p, const float (&v)[8]) { Elem<__half>::store(p, v); }
};
template <> struct Vec8<float> {
  __device__ static inline void load(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  __device__ static inline void store(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};

// ---- pass 1: per-(batch, group) mean / rstd ------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(GN_MAX_THREADS)
gn_stats_kernel(const T* __restrict__ x, const T* __restrict__ chan_add, float* __restrict__ partial,
                const GnParams p) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ __align__(16) float gn_smem[];
  float* s_sum = gn_smem;                       // [PL][C]
  float* s_sq = gn_smem + (size_t)p.PL * p.C;   // [PL][C]
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int tid = threadIdx.x;
  const int pl = tid / p.NV, v = tid - pl * p.NV;
  if (pl < p.PL) {
    float add[8], sum[8], sq[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { add[k] = 0.f; sum[k] = 0.f; sq[k] = 0.f; }
    if (chan_add != nullptr) Vec8<T>::load(chan_add + (size_t)b * p.C + v * 8, add);
    const int p0 = chunk * p.ppc;
    const int p1 = min(p.HW, p0 + p.ppc);
    const T* xb = x + ((size_t)b * p.HW) * p.C + v * 8;
#pragma unroll 4
    for (int pix = p0 + pl; pix < p1; pix += p.PL) {
      float xv[8];
      Vec8<T>::load(xb + (size_t)pix * p.C, xv);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float z = xv[k] + add[k];
        sum[k] += z;
        sq[k] = fmaf(z, z, sq[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      s_sum[pl * p.C + v * 8 + k] = sum[k];
      s_sq[pl * p.C + v * 8 + k] = sq[k];
    }
  }
  __syncthreads();
  // per-group sums of this chunk: one warp per group (fixed lane -> element map, fixed shuffle tree)
  const int warp = tid >> 5, lane = tid & 31, nwarps = blockDim.x >> 5;
  const int per_group = p.PL * p.cpg;
  for (int g = warp; g < p.G; g += nwarps) {
    float a = 0.f, q = 0.f;
    for (int i = lane; i < per_group; i += 32) {
      const int l = i / p.cpg, c = g * p.cpg + (i - l * p.cpg);
      a += s_sum[l * p.C + c];
      q += s_sq[l * p.C + c];
    }
    a = warp_sum(a);
    q = warp_sum(q);
    if (lane == 0) {
      float* dst = partial + (((size_t)b * p.chunks + chunk) * p.G + g) * 2;
      dst[0] = a; dst[1] = q;
    }
  }
}

// ---- pass 2: normalise, affine, activation -----------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
gn_apply_kernel(T* __restrict__ y, const T* __restrict__ x, const T* __restrict__ chan_add,
                const T* __restrict__ gamma, const T* __restrict__ beta, const float* __restrict__ partial,
                const GnParams p, int chunks, int ppc, float* __restrict__ stats_out) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ __align__(16) float gn_smem[];
  float* s_scale = gn_smem;             // [C]
  float* s_shift = gn_smem + p.C;       // [C]
  float* s_stat = gn_smem + 2 * p.C;    // [G][2] mean, rstd
  const int b = blockIdx.y;
  {
    // every block combines the per-chunk partial sums of its batch item itself (fixed order: no ticket,
    // no atomics, no extra launch); the partials are a few KB and L2-resident.  Thread (lane l, group g)
    // sums chunks l, l+nl, ... with independent loads in flight, then the nl lane sums are added in order.
    double* s_part = reinterpret_cast<double*>(gn_smem + 2 * p.C + 2 * p.G);   // [nl][G][2]
    const int g = threadIdx.x % p.G, l = threadIdx.x / p.G, nl = blockDim.x / p.G;
    if (l < nl) {
      double a = 0.0, q = 0.0;
      const float2* src = reinterpret_cast<const float2*>(partial) + ((size_t)b * p.chunks) * p.G + g;
      int ch = l;
      for (; ch + 3 * nl < p.chunks; ch += 4 * nl) {
        const float2 v0 = __ldcg(src + (size_t)ch * p.G), v1 = __ldcg(src + (size_t)(ch + nl) * p.G);
        const float2 v2 = __ldcg(src + (size_t)(ch + 2 * nl) * p.G), v3 = __ldcg(src + (size_t)(ch + 3 * nl) * p.G);
        a += (double)v0.x; q += (double)v0.y; a += (double)v1.x; q += (double)v1.y;
        a += (double)v2.x; q += (double)v2.y; a += (double)v3.x; q += (double)v3.y;
      }
      for (; ch < p.chunks; ch += nl) {
        const float2 v = __ldcg(src + (size_t)ch * p.G);
        a += (double)v.x; q += (double)v.y;
      }
      s_part[(l * p.G + g) * 2] = a;
      s_part[(l * p.G + g) * 2 + 1] = q;
    }
    __syncthreads();
    if (threadIdx.x < p.G) {
      double a = 0.0, q = 0.0;
      for (int i = 0; i < nl; ++i) { a += s_part[(i * p.G + g) * 2]; q += s_part[(i * p.G + g) * 2 + 1]; }
      const double n = (double)p.cpg * (double)p.HW;
      const double mean = a / n;
      double var = q / n - mean * mean;
      if (var < 0.0) var = 0.0;
      s_stat[2 * g] = (float)mean;
      s_stat[2 * g + 1] = rsqrtf((float)var + p.eps);
      if (stats_out != nullptr && blockIdx.x == 0) {      // saved for the backward pass: [B][G]{mean, rstd}
        stats_out[((size_t)b * p.G + g) * 2] = s_stat[2 * g];
        stats_out[((size_t)b * p.G + g) * 2 + 1] = s_stat[2 * g + 1];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < p.C; c += blockDim.x) {
    const int g = c / p.cpg;
    const float mean = s_stat[2 * g], rstd = s_stat[2 * g + 1];
    const float ga = gamma ? Elem<T>::to_f(gamma[c]) : 1.f, be = beta ? Elem<T>::to_f(beta[c]) : 0.f;
    const float add = chan_add ? Elem<T>::to_f(chan_add[(size_t)b * p.C + c]) : 0.f;
    s_scale[c] = rstd * ga;
    s_shift[c] = fmaf(add - mean, rstd * ga, be);
  }
  __syncthreads();
  const int p0 = blockIdx.x * ppc;
  const int p1 = min(p.HW, p0 + ppc);
  if (p0 >= p1) return;
  const long long total = (long long)(p1 - p0) * p.NV;
  const T* xb = x + ((size_t)b * p.HW + p0) * p.C;
  T* yb = y + ((size_t)b * p.HW + p0) * p.C;
  int v = threadIdx.x % p.NV;
  const int dv = blockDim.x % p.NV;
  for (long long idx = threadIdx.x; idx < total; idx += blockDim.x) {
    float xv[8], o[8];
    Vec8<T>::load(xb + idx * 8, xv);
    const float4 sc0 = *reinterpret_cast<const float4*>(s_scale + v * 8), sc1 = *reinterpret_cast<const float4*>(s_scale + v * 8 + 4);
    const float4 sh0 = *reinterpret_cast<const float4*>(s_shift + v * 8), sh1 = *reinterpret_cast<const float4*>(s_shift + v * 8 + 4);
    const float sc[8] = {sc0.x, sc0.y, sc0.z, sc0.w, sc1.x, sc1.y, sc1.z, sc1.w};
    const float sh[8] = {sh0.x, sh0.y, sh0.z, sh0.w, sh1.x, sh1.y, sh1.z, sh1.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float u = fmaf(xv[k], sc[k], sh[k]);
      if (p.silu) u = u / (1.f + __expf(-u));
      o[k] = u;
    }
    Vec8<T>::store(yb + idx * 8, o);
    v += dv;
    if (v >= p.NV) v -= p.NV;
  }
}

// ---- single-launch form: one CTA (or a small cluster) per (image, group set) --------------------------------------------
// The data of GP adjacent groups is a [HW][GP * cpg] slab (rows of GP * cpg * sizeof(T) bytes at a stride of
// C * sizeof(T)); GP = 1, 2, 4 or 8 is the smallest count that makes a row a whole number of 16-byte vectors (SD 1.x:
// 320 channels / 32 groups = 20-byte groups -> GP = 4, 80-byte rows).  k = 1 .. 8 CTAs (one cluster) split the rows of a
// slab so that each holds a few tens of KB of it in shared memory:
//   1. rows -> shared memory, 16-byte vectors (x is read from global memory ONCE);
//   2. statistics from the shared copy by 32-bit words: the block size is a multiple of the words per row, so a thread
//      always sees the same word of a row -> one group, two channels, its {sum, sum of squares} stay in registers;
//      one warp per group then adds the per-thread pairs in a fixed order (fp64);
//   3. k > 1: the per-CTA pairs of the peers over distributed shared memory, in rank order (deterministic);
//   4. scale / shift of the thread's 16-byte column in registers, normalised rows written from the shared copy.
// No partial-sum workspace, no second launch, no second read: the two-kernel form costs ~13.5 us per GroupNorm in the
// denoise forward whatever the size (launch + dependent global-load latency), 17 % of it.
constexpr int GN_GROUP_THREADS = 512;
constexpr int GN_GROUP_MAXGP = 8;
constexpr int GN_GROUP_BATCH = 8;      // unroll of the row loop

// 1 / (1 + exp(-t)) with the two SFU approximations and no range fix-ups (ex2: 2 ulp, rcp: 1 ulp; exp overflow ->
// rcp(inf) = 0, underflow -> 1): 5 instructions per SiLU instead of 10
__device__ __forceinline__ float fast_sigmoid(float t) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t * -1.4426950408889634f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
  return r;
}

template <typename T> struct Word;      // the elements of one 32-bit word
template <> struct Word<__nv_bfloat16> {
  static constexpr int N = 2;
  __device__ static inline void unpack(uint32_t w, float (&f)[2]) {
    f[0] = __uint_as_float(w << 16);
    f[1] = __uint_as_float(w & 0xffff0000u);
  }
  __device__ static inline uint32_t pack(const float (&f)[2]) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(f[0], f[1]);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
};
template <> struct Word<__half> {
  static constexpr int N = 2;
  __device__ static inline void unpack(uint32_t w, float (&f)[2]) { Elem<__half>::unpack2(w, f[0], f[1]); }
  __device__ static inline uint32_t pack(const float (&f)[2]) { return Elem<__half>::pack2(f[0], f[1]); }
};
template <> struct Word<float> {
  static constexpr int N = 1;
  __device__ static inline void unpack(uint32_t w, float (&f)[2]) { f[0] = __uint_as_float(w); f[1] = 0.f; }
  __device__ static inline uint32_t pack(const float (&f)[2]) { return __float_as_uint(f[0]); }
};

struct GnGroupArgs {
  int k;        // CTAs per slab (cluster size)
  int gp;       // groups per slab
  int wr;       // 32-bit words per slab row
  int rows;     // rows (pixels) per CTA
  unsigned long long* trace;   // profiling aid (slider_debug_gn_trace): 8 globaltimer stamps per CTA, or nullptr
};

__device__ __forceinline__ void gn_stamp(unsigned long long* trace, int slot) {
  if (trace != nullptr && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 8 + slot] = t;
  }
}

template <typename T>
__global__ void __launch_bounds__(GN_GROUP_THREADS, 2)       // two CTAs per SM: clusters of 8 must not start in waves
gn_group_kernel(T* __restrict__ y, const T* __restrict__ x, const T* __restrict__ chan_add,
                const T* __restrict__ gamma, const T* __restrict__ beta, const GnParams p, const GnGroupArgs ga,
                float* __restrict__ stats_out) {
  constexpr int EW = Word<T>::N, NE = 4 * EW;                         // elements per word / per 16-byte vector
  extern __shared__ __align__(16) unsigned char gn_gsm[];
  const int tid = threadIdx.x, nt = blockDim.x, b = blockIdx.y;
  const int warp = tid >> 5, lane = tid & 31, nwarps = nt >> 5;
  const int wpr = ga.wr >> 2;                                         // 16-byte vectors per row
  const int nch = ga.gp * p.cpg;                                      // channels of the slab
  const int total = ga.rows * wpr;
  uint4* slab = reinterpret_cast<uint4*>(gn_gsm);
  float2* s_part = reinterpret_cast<float2*>(gn_gsm + (size_t)total * 16);       // [warps][MAXGP] per-warp pairs
  double* s_cta = reinterpret_cast<double*>(s_part + (GN_GROUP_THREADS / 32) * GN_GROUP_MAXGP);   // [gp][2], read by the peers
  float* s_stat = reinterpret_cast<float*>(s_cta + 2 * GN_GROUP_MAXGP);          // [gp][2] mean, rstd
  const int set = blockIdx.x / ga.k;
  const int part = ga.k > 1 ? (int)ptx::cluster_ctarank() : 0;
  const int c0 = set * nch;
  // nt % wpr == 0: a thread stays in one 16-byte column of the slab -> its NE channels, their groups, gamma / beta /
  // pre-add and later scale / shift all live in registers
  const int vpos = tid % wpr, vrow = tid / wpr, vstep = nt / wpr;
  pdl_trigger();
  gn_stamp(ga.trace, 0);
  float add[NE], sum[NE], sq[NE];
  int ge[NE];                                     // group (within the slab) of each of the column's channels
  {
    int c = vpos * NE, gi = c / p.cpg, rem = c - gi * p.cpg;
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      add[e] = 0.f; sum[e] = 0.f; sq[e] = 0.f;
      ge[e] = gi;
      if (++rem == p.cpg) { rem = 0; ++gi; }
    }
  }
  // frozen parameters of the column (not produced by the previous kernel), kept packed until the apply pass
  uint4 graw = make_uint4(0, 0, 0, 0), braw = make_uint4(0, 0, 0, 0);
  if (gamma != nullptr) graw = *reinterpret_cast<const uint4*>(gamma + c0 + vpos * NE);
  if (beta != nullptr) braw = *reinterpret_cast<const uint4*>(beta + c0 + vpos * NE);
  pdl_wait();
  gn_stamp(ga.trace, 1);
  if (chan_add != nullptr) Elem<T>::load(chan_add + (size_t)b * p.C + c0 + vpos * NE, add);
  // 1. rows -> shared memory (x is read from global memory once), statistics of the column on the way
  const size_t rstride = (size_t)p.C * sizeof(T) / 16;                // row stride in vectors
  const size_t base = (((size_t)b * p.HW + (size_t)part * ga.rows) * p.C + c0) * sizeof(T) / 16;
  const uint4* xg = reinterpret_cast<const uint4*>(x) + base + vpos;
  uint4* yg = reinterpret_cast<uint4*>(y) + base + vpos;
  {
    int i = tid, r = vrow;
#pragma unroll GN_GROUP_BATCH
    for (; i < total; i += nt, r += vstep) {
      const uint4 v = xg[(size_t)r * rstride];
      slab[i] = v;
      const uint32_t in[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float f[2];
        Word<T>::unpack(in[j], f);
#pragma unroll
        for (int e = 0; e < EW; ++e) {
          const float z = f[e] + add[j * EW + e];
          sum[j * EW + e] += z;
          sq[j * EW + e] = fmaf(z, z, sq[j * EW + e]);
        }
      }
    }
  }
  gn_stamp(ga.trace, 2);
  // 2. per-warp sums of every group (fixed order over the column's elements, fixed shuffle tree), then one warp per
  //    group over the warps (fp64)
  for (int gi = 0; gi < ga.gp; ++gi) {
    float a = 0.f, q = 0.f;
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      a += ge[e] == gi ? sum[e] : 0.f;
      q += ge[e] == gi ? sq[e] : 0.f;
    }
    a = warp_sum(a);
    q = warp_sum(q);
    if (lane == 0) s_part[warp * GN_GROUP_MAXGP + gi] = make_float2(a, q);
  }
  __syncthreads();
  if (warp < ga.gp) {
    double a = 0.0, q = 0.0;
    if (lane < nwarps) { const float2 v = s_part[lane * GN_GROUP_MAXGP + warp]; a = (double)v.x; q = (double)v.y; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (lane == 0) { s_cta[2 * warp] = a; s_cta[2 * warp + 1] = q; }
  }
  // 3. the peers' pairs: warp gi, lane r reads the pair of peer r; fixed shuffle tree over the ranks
  gn_stamp(ga.trace, 3);
  if (ga.k > 1) { ptx::cluster_arrive(); ptx::cluster_wait(); } else { __syncthreads(); }
  gn_stamp(ga.trace, 4);
  if (warp < ga.gp) {
    double a = 0.0, q = 0.0;
    if (ga.k > 1) {
      if (lane < ga.k) {
        const uint32_t peer = ptx::mapa(ptx::smem_u32(s_cta + 2 * warp), (uint32_t)lane);
        asm volatile("ld.shared::cluster.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(q) : "r"(peer));
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
      }
    } else {
      a = s_cta[2 * warp]; q = s_cta[2 * warp + 1];
    }
    if (lane == 0) {
      const double n = (double)p.cpg * (double)p.HW;
      const double mean = a / n;
      double var = q / n - mean * mean;
      if (var < 0.0) var = 0.0;
      const float mf = (float)mean, rf = rsqrtf((float)var + p.eps);
      s_stat[2 * warp] = mf; s_stat[2 * warp + 1] = rf;
      if (stats_out != nullptr && part == 0) {               // saved for the backward pass: [B][G]{mean, rstd}
        const int g = set * ga.gp + warp;
        stats_out[((size_t)b * p.G + g) * 2] = mf;
        stats_out[((size_t)b * p.G + g) * 2 + 1] = rf;
      }
    }
  }
  if (ga.k > 1) ptx::cluster_arrive();    // the peers may still read s_cta: arrive now, wait before exit
  __syncthreads();
  gn_stamp(ga.trace, 5);
  // 4. normalise from the shared copy
  {
    float gam[NE], bet[NE];
    {
      const uint32_t gw[4] = {graw.x, graw.y, graw.z, graw.w}, bw[4] = {braw.x, braw.y, braw.z, braw.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float f[2], h[2];
        Word<T>::unpack(gw[j], f);
        Word<T>::unpack(bw[j], h);
#pragma unroll
        for (int e = 0; e < EW; ++e) { gam[j * EW + e] = gamma != nullptr ? f[e] : 1.f; bet[j * EW + e] = h[e]; }
      }
    }
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      const float mean = s_stat[2 * ge[e]], rstd = s_stat[2 * ge[e] + 1];
      gam[e] *= rstd;                                    // scale
      bet[e] = fmaf(add[e] - mean, gam[e], bet[e]);      // shift
    }
    int i = tid, r = vrow;
#pragma unroll 2
    for (; i < total; i += nt, r += vstep) {
      const uint4 v = slab[i];
      const uint32_t in[4] = {v.x, v.y, v.z, v.w};
      uint32_t out[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float f[2], u[2] = {0.f, 0.f};
        Word<T>::unpack(in[j], f);
#pragma unroll
        for (int e = 0; e < EW; ++e) {
          float t = fmaf(f[e], gam[j * EW + e], bet[j * EW + e]);
          if (p.silu) t *= fast_sigmoid(t);
          u[e] = t;
        }
        out[j] = Word<T>::pack(u);
      }
      yg[(size_t)r * rstride] = make_uint4(out[0], out[1], out[2], out[3]);
    }
  }
  gn_stamp(ga.trace, 6);
  if (ga.k > 1) ptx::cluster_wait();
  gn_stamp(ga.trace, 7);
}

// per-(image, group set) single-launch form: on by default (SLIDER_GN_GROUP=0 / slider_debug_gn_group(0, 0): two-kernel
// form); SLIDER_GN_GROUP_KB: the slab bytes one CTA may hold before the slab is split over a larger cluster
static int g_gn_group = [] { const char* e = getenv("SLIDER_GN_GROUP"); return e ? atoi(e) : 1; }();
static int g_gn_group_kb = [] { const char* e = getenv("SLIDER_GN_GROUP_KB"); const int v = e ? atoi(e) : 24; return v < 4 ? 4 : (v > 128 ? 128 : v); }();

static unsigned long long* g_gn_trace = nullptr;

struct GnGroupPlan { GnGroupArgs a; int threads; size_t smem; };

template <typename T>
static bool gn_group_plan(const GnParams& p, GnGroupPlan* gp) {
  if (((size_t)p.cpg * sizeof(T)) % 4 != 0) return false;      // a 32-bit word must not straddle two groups
  int g = 1;
  while (g <= GN_GROUP_MAXGP && (((size_t)g * p.cpg * sizeof(T)) % 16 != 0 || p.G % g != 0)) g *= 2;
  if (g > GN_GROUP_MAXGP) return false;
  const size_t seg = (size_t)g * p.cpg * sizeof(T);
  const int wr = (int)(seg / 4);
  int l = wr;                                   // lcm(32, wr): block sizes that keep a thread on one word of a row
  while (l % 32 != 0) l += wr;
  if (l > GN_GROUP_THREADS) return false;
  const size_t slab = (size_t)p.HW * seg;
  int k = 1;
  while (k < 8 && slab / k > (size_t)g_gn_group_kb * 1024) k *= 2;     // clusters of 16 start in waves (measured)
  while (k > 1 && p.HW % k != 0) k /= 2;
  // above ~100 KB per CTA only one CTA fits an SM and clusters of 8 start in waves: measured equal to 14 % slower than
  // the two-kernel form (tools/bench_gn_big.py, profiles/r01_ab_gn_group.txt), which those shapes therefore keep
  if (slab / k > 100 * 1024) return false;
  gp->a.k = k; gp->a.gp = g; gp->a.wr = wr; gp->a.rows = p.HW / k; gp->a.trace = g_gn_trace;
  gp->threads = (GN_GROUP_THREADS / l) * l;
  gp->smem = (size_t)gp->a.rows * seg + (GN_GROUP_THREADS / 32) * GN_GROUP_MAXGP * sizeof(float2) +
             2 * GN_GROUP_MAXGP * sizeof(double) + 2 * GN_GROUP_MAXGP * sizeof(float);
  return true;
}

template <typename T>
static int gn_group_launch(void* y, const void* x, const void* chan_add, const void* gamma, const void* beta,
                           const GnParams& p, const GnGroupPlan& gp, cudaStream_t st, float* stats_out) {
  static thread_local bool configured = false;
  if (!configured) {
    SLIDER_CUDA(cudaFuncSetAttribute(gn_group_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((p.G / gp.a.gp) * gp.a.k, p.B); cfg.blockDim = dim3(gp.threads); cfg.dynamicSmemBytes = gp.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (gp.a.k > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = (unsigned)gp.a.k; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaLaunchKernelEx(&cfg, gn_group_kernel<T>, (T*)y, (const T*)x, (const T*)chan_add, (const T*)gamma, (const T*)beta, p,
                     gp.a, stats_out);
  return post_launch("gn_group_kernel");
}

template <typename T>
static int groupnorm_impl(void* y, const void* x, const void* chan_add, const void* gamma, const void* beta,
                          float* ws, GnParams p, cudaStream_t st, float* stats_out = nullptr) {
  GnGroupPlan gp;
  if (g_gn_group && p.B <= 65535 && gn_group_plan<T>(p, &gp))
    return gn_group_launch<T>(y, x, chan_add, gamma, beta, p, gp, st, stats_out);
  // workspace: per-chunk partial sums [B][chunks][G][2]
  float* partial = ws;
  {
    dim3 grid(p.chunks, p.B);
    const size_t smem = (size_t)p.PL * p.C * 2 * sizeof(float);
    int threads = p.PL * p.NV > p.G ? p.PL * p.NV : p.G;
    threads = (threads + 31) / 32 * 32;
    launch_pdl(gn_stats_kernel<T>, grid, dim3(threads), smem, st, (const T*)x, (const T*)chan_add, partial, p);
    int rc = post_launch("gn_stats_kernel");
    if (rc) return rc;
  }
  {
    int chunks = (sm_count() * 4 + p.B - 1) / p.B;
    int ppc = (p.HW + chunks - 1) / chunks;
    if (ppc < 4) ppc = 4;
    chunks = (p.HW + ppc - 1) / ppc;
    dim3 grid(chunks, p.B);
    launch_pdl(gn_apply_kernel<T>, grid, dim3(256),
               ((size_t)p.C * 2 + (size_t)p.G * 2) * sizeof(float) + 256 * 2 * sizeof(double) + 16, st,
               (T*)y, (const T*)x, (const T*)chan_add, (const T*)gamma, (const T*)beta, (const float*)partial, p, chunks, ppc, stats_out);
  }
  return post_launch("gn_apply_kernel");
}

// ---- backward (frozen gamma / beta: only dx and the gradient of the per-channel pre-add) ---------------------
//   xhat = (x + add - mean) * rstd,  z = xhat * gamma + beta,  d = gy * gamma * act'(z)
//   s1_g = sum d,  s2_g = sum d * xhat  over the group     dx = rstd * (d - (s1 + xhat * s2) / n)
//   dadd[b,c] = sum_pixels dx = rstd * (D_c - HW * s1 / n - (s2 / n) * X_c),  D_c = sum_p d,  X_c = sum_p xhat
__device__ __forceinline__ float silu_grad(float z) {
  const float sg = 1.f / (1.f + __expf(-z));
  return sg * fmaf(z, 1.f - sg, 1.f);
}

// pass 1: per-chunk partial sums {s1, s2} per group (+ {D_c, X_c} per channel when dadd is wanted)
template <typename T>
__global__ void __launch_bounds__(GN_MAX_THREADS)
gn_bwd_stats_kernel(const T* __restrict__ gy, const T* __restrict__ x, const T* __restrict__ chan_add,
                    const T* __restrict__ gamma, const T* __restrict__ beta, const float* __restrict__ stats,
                    float* __restrict__ partial_g, float* __restrict__ partial_c, const GnParams p) {
  extern __shared__ __align__(16) float gn_smem[];
  float* s_a = gn_smem;                       // [PL][C] sum d
  float* s_b = gn_smem + (size_t)p.PL * p.C;  // [PL][C] sum d * xhat
  float* s_x = gn_smem + (size_t)2 * p.PL * p.C;   // [PL][C] sum xhat (only with partial_c)
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int tid = threadIdx.x;
  const int pl = tid / p.NV, v = tid - pl * p.NV;
  if (pl < p.PL) {
    float rs[8], sh[8], ga[8], be[8], A[8], Bq[8], Xs[8];
    float add[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { add[k] = 0.f; A[k] = 0.f; Bq[k] = 0.f; Xs[k] = 0.f; }
    if (chan_add != nullptr) Vec8<T>::load(chan_add + (size_t)b * p.C + v * 8, add);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = v * 8 + k, g = c / p.cpg;
      const float mean = stats[((size_t)b * p.G + g) * 2], rstd = stats[((size_t)b * p.G + g) * 2 + 1];
      rs[k] = rstd;
      sh[k] = (add[k] - mean) * rstd;
      ga[k] = gamma ? Elem<T>::to_f(gamma[c]) : 1.f;
      be[k] = beta ? Elem<T>::to_f(beta[c]) : 0.f;
    }
    const int p0 = chunk * p.ppc;
    const int p1 = min(p.HW, p0 + p.ppc);
    const size_t base = ((size_t)b * p.HW) * p.C + v * 8;
#pragma unroll 2
    for (int pix = p0 + pl; pix < p1; pix += p.PL) {
      float xv[8], gv[8];
      Vec8<T>::load(x + base + (size_t)pix * p.C, xv);
      Vec8<T>::load(gy + base + (size_t)pix * p.C, gv);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float xh = fmaf(xv[k], rs[k], sh[k]);
        float d = gv[k] * ga[k];
        if (p.silu) d *= silu_grad(fmaf(xh, ga[k], be[k]));
        A[k] += d;
        Bq[k] = fmaf(d, xh, Bq[k]);
        Xs[k] += xh;
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      s_a[pl * p.C + v * 8 + k] = A[k];
      s_b[pl * p.C + v * 8 + k] = Bq[k];
      if (partial_c != nullptr) s_x[pl * p.C + v * 8 + k] = Xs[k];
    }
  }
  __syncthreads();
  const int warp = tid >> 5, lane = tid & 31, nwarps = blockDim.x >> 5;
  const int per_group = p.PL * p.cpg;
  for (int g = warp; g < p.G; g += nwarps) {
    float a = 0.f, q = 0.f;
    for (int i = lane; i < per_group; i += 32) {
      const int l = i / p.cpg, c = g * p.cpg + (i - l * p.cpg);
      a += s_a[l * p.C + c];
      q += s_b[l * p.C + c];
    }
    a = warp_sum(a);
    q = warp_sum(q);
    if (lane == 0) {
      float* dst = partial_g + (((size_t)b * p.chunks + chunk) * p.G + g) * 2;
      dst[0] = a; dst[1] = q;
    }
  }
  if (partial_c != nullptr) {       // per-channel sums of this chunk, pixel lanes added in lane order
    for (int c = tid; c < p.C; c += blockDim.x) {
      float a = 0.f, xs = 0.f;
      for (int l = 0; l < p.PL; ++l) { a += s_a[l * p.C + c]; xs += s_x[l * p.C + c]; }
      float* dst = partial_c + (((size_t)b * p.chunks + chunk) * p.C + c) * 2;
      dst[0] = a; dst[1] = xs;
    }
  }
}

// pass 2: dx, and dadd by the first ceil(C / 256) blocks of every batch item
template <typename T>
__global__ void __launch_bounds__(256)
gn_bwd_apply_kernel(T* __restrict__ dx, T* __restrict__ dadd, const T* __restrict__ gy, const T* __restrict__ x,
                    const T* __restrict__ chan_add, const T* __restrict__ gamma, const T* __restrict__ beta,
                    const float* __restrict__ stats, const float* __restrict__ partial_g,
                    const float* __restrict__ partial_c, const GnParams p, int chunks, int ppc) {
  extern __shared__ __align__(16) float gn_smem[];
  float* s_rs = gn_smem;                // [C] rstd
  float* s_sh = gn_smem + p.C;          // [C] (add - mean) * rstd
  float* s_ga = gn_smem + 2 * p.C;      // [C]
  float* s_be = gn_smem + 3 * p.C;      // [C]
  float* s_c1 = gn_smem + 4 * p.C;      // [C] s1_g / n
  float* s_c2 = gn_smem + 5 * p.C;      // [C] s2_g / n
  float* s_grp = gn_smem + 6 * p.C;     // [G][2]
  double* s_part = reinterpret_cast<double*>(gn_smem + 6 * p.C + 2 * p.G);   // [nl][G][2]
  const int b = blockIdx.y;
  {
    const int g = threadIdx.x % p.G, l = threadIdx.x / p.G, nl = blockDim.x / p.G;
    if (l < nl) {
      double a = 0.0, q = 0.0;
      const float2* src = reinterpret_cast<const float2*>(partial_g) + ((size_t)b * p.chunks) * p.G + g;
      int ch = l;
      for (; ch + 3 * nl < p.chunks; ch += 4 * nl) {
        const float2 v0 = __ldcg(src + (size_t)ch * p.G), v1 = __ldcg(src + (size_t)(ch + nl) * p.G);
        const float2 v2 = __ldcg(src + (size_t)(ch + 2 * nl) * p.G), v3 = __ldcg(src + (size_t)(ch + 3 * nl) * p.G);
        a += (double)v0.x; q += (double)v0.y; a += (double)v1.x; q += (double)v1.y;
        a += (double)v2.x; q += (double)v2.y; a += (double)v3.x; q += (double)v3.y;
      }
      for (; ch < p.chunks; ch += nl) {
        const float2 v = __ldcg(src + (size_t)ch * p.G);
        a += (double)v.x; q += (double)v.y;
      }
      s_part[(l * p.G + g) * 2] = a;
      s_part[(l * p.G + g) * 2 + 1] = q;
    }
    __syncthreads();
    if (threadIdx.x < p.G) {
      double a = 0.0, q = 0.0;
      for (int i = 0; i < nl; ++i) { a += s_part[(i * p.G + g) * 2]; q += s_part[(i * p.G + g) * 2 + 1]; }
      const double n = (double)p.cpg * (double)p.HW;
      s_grp[2 * g] = (float)(a / n);
      s_grp[2 * g + 1] = (float)(q / n);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < p.C; c += blockDim.x) {
    const int g = c / p.cpg;
    const float mean = stats[((size_t)b * p.G + g) * 2], rstd = stats[((size_t)b * p.G + g) * 2 + 1];
    const float add = chan_add ? Elem<T>::to_f(chan_add[(size_t)b * p.C + c]) : 0.f;
    s_rs[c] = rstd;
    s_sh[c] = (add - mean) * rstd;
    s_ga[c] = gamma ? Elem<T>::to_f(gamma[c]) : 1.f;
    s_be[c] = beta ? Elem<T>::to_f(beta[c]) : 0.f;
    s_c1[c] = s_grp[2 * g];
    s_c2[c] = s_grp[2 * g + 1];
  }
  __syncthreads();
  if (dadd != nullptr && (int)blockIdx.x * 256 < p.C) {
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c < p.C) {
      double D = 0.0, X = 0.0;
      const float2* src = reinterpret_cast<const float2*>(partial_c) + ((size_t)b * p.chunks) * p.C + c;
      int ch = 0;
      for (; ch + 3 < p.chunks; ch += 4) {
        const float2 v0 = __ldcg(src + (size_t)ch * p.C), v1 = __ldcg(src + (size_t)(ch + 1) * p.C);
        const float2 v2 = __ldcg(src + (size_t)(ch + 2) * p.C), v3 = __ldcg(src + (size_t)(ch + 3) * p.C);
        D += (double)v0.x; X += (double)v0.y; D += (double)v1.x; X += (double)v1.y;
        D += (double)v2.x; X += (double)v2.y; D += (double)v3.x; X += (double)v3.y;
      }
      for (; ch < p.chunks; ++ch) {
        const float2 v = __ldcg(src + (size_t)ch * p.C);
        D += (double)v.x; X += (double)v.y;
      }
      const double r = (double)s_rs[c] * (D - (double)p.HW * (double)s_c1[c] - (double)s_c2[c] * X);
      dadd[(size_t)b * p.C + c] = Elem<T>::from_f((float)r);
    }
  }
  const int p0 = blockIdx.x * ppc;
  const int p1 = min(p.HW, p0 + ppc);
  if (p0 >= p1) return;
  const long long total = (long long)(p1 - p0) * p.NV;
  const size_t base = ((size_t)b * p.HW + p0) * p.C;
  int v = threadIdx.x % p.NV;
  const int dv = blockDim.x % p.NV;
  for (long long idx = threadIdx.x; idx < total; idx += blockDim.x) {
    float xv[8], gv[8], o[8];
    Vec8<T>::load(x + base + idx * 8, xv);
    Vec8<T>::load(gy + base + idx * 8, gv);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = v * 8 + k;
      const float xh = fmaf(xv[k], s_rs[c], s_sh[c]);
      float d = gv[k] * s_ga[c];
      if (p.silu) d *= silu_grad(fmaf(xh, s_ga[c], s_be[c]));
      o[k] = s_rs[c] * (d - s_c1[c] - xh * s_c2[c]);
    }
    Vec8<T>::store(dx + base + idx * 8, o);
    v += dv;
    if (v >= p.NV) v -= p.NV;
  }
}

// Single-launch backward on the forward's slab plan (gn_group_plan): the CTAs of a cluster hold the (image, group set)
// slab of BOTH x and gy in shared memory, so the two tensors are read from global memory once, the group sums
// {s1, s2} travel through the cluster like the forward's statistics, and dx is formed from the shared copies.  The
// two-kernel form above reads x and gy twice and pays two launches with three dependent global round trips each
// (12.6 + 9.8 us per layer inside the train graph, r02 torch profile).
//   dadd[b, c] = sum over the image's pixels of dx[., c]: per-thread column sums -> shared [row lane][channel] ->
//   per-CTA sums (fixed order) -> rank 0 of the cluster adds the peers' arrays in rank order (deterministic).
template <typename T>
__global__ void __launch_bounds__(GN_GROUP_THREADS, 1)
gn_bwd_group_kernel(T* __restrict__ dx, T* __restrict__ dadd, const T* __restrict__ gy, const T* __restrict__ x,
                    const T* __restrict__ chan_add, const T* __restrict__ gamma, const T* __restrict__ beta,
                    const float* __restrict__ stats, const GnParams p, const GnGroupArgs ga) {
  constexpr int EW = Word<T>::N, NE = 4 * EW;
  extern __shared__ __align__(16) unsigned char gn_gsm[];
  const int tid = threadIdx.x, nt = blockDim.x, b = blockIdx.y;
  const int warp = tid >> 5, lane = tid & 31, nwarps = nt >> 5;
  const int wpr = ga.wr >> 2;
  const int nch = ga.gp * p.cpg;
  const int total = ga.rows * wpr;
  uint4* slab_x = reinterpret_cast<uint4*>(gn_gsm);
  uint4* slab_g = slab_x + total;
  float2* s_part = reinterpret_cast<float2*>(gn_gsm + (size_t)2 * total * 16);
  double* s_cta = reinterpret_cast<double*>(s_part + (GN_GROUP_THREADS / 32) * GN_GROUP_MAXGP);
  float* s_c = reinterpret_cast<float*>(s_cta + 2 * GN_GROUP_MAXGP);            // [gp][2] s1 / n, s2 / n
  float* s_colsum = s_c + 2 * GN_GROUP_MAXGP;                                   // [nch] this CTA's column sums of dx
  float* s_col = s_colsum + nch;                                                // [vstep][nch] per row-lane sums
  const int set = blockIdx.x / ga.k;
  const int part = ga.k > 1 ? (int)ptx::cluster_ctarank() : 0;
  const int c0 = set * nch;
  const int vpos = tid % wpr, vrow = tid / wpr, vstep = nt / wpr;
  pdl_trigger();
  float add[NE], rs[NE], sh[NE], gam[NE], bet[NE], sd[NE], sdx[NE];
  int ge[NE];
  {
    int c = vpos * NE, gi = c / p.cpg, rem = c - gi * p.cpg;
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      add[e] = 0.f; sd[e] = 0.f; sdx[e] = 0.f;
      ge[e] = gi;
      if (++rem == p.cpg) { rem = 0; ++gi; }
    }
  }
  {
    uint4 graw = make_uint4(0, 0, 0, 0), braw = make_uint4(0, 0, 0, 0);
    if (gamma != nullptr) graw = *reinterpret_cast<const uint4*>(gamma + c0 + vpos * NE);
    if (beta != nullptr) braw = *reinterpret_cast<const uint4*>(beta + c0 + vpos * NE);
    const uint32_t gw[4] = {graw.x, graw.y, graw.z, graw.w}, bw[4] = {braw.x, braw.y, braw.z, braw.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float f[2], h[2];
      Word<T>::unpack(gw[j], f);
      Word<T>::unpack(bw[j], h);
#pragma unroll
      for (int e = 0; e < EW; ++e) { gam[j * EW + e] = gamma != nullptr ? f[e] : 1.f; bet[j * EW + e] = h[e]; }
    }
  }
  pdl_wait();
  if (chan_add != nullptr) Elem<T>::load(chan_add + (size_t)b * p.C + c0 + vpos * NE, add);
#pragma unroll
  for (int e = 0; e < NE; ++e) {
    const int g = set * ga.gp + ge[e];
    const float mean = stats[((size_t)b * p.G + g) * 2], rstd = stats[((size_t)b * p.G + g) * 2 + 1];
    rs[e] = rstd;
    sh[e] = (add[e] - mean) * rstd;
  }
  const size_t rstride = (size_t)p.C * sizeof(T) / 16;
  const size_t base = (((size_t)b * p.HW + (size_t)part * ga.rows) * p.C + c0) * sizeof(T) / 16;
  const uint4* xg = reinterpret_cast<const uint4*>(x) + base + vpos;
  const uint4* gg = reinterpret_cast<const uint4*>(gy) + base + vpos;
  uint4* dg = reinterpret_cast<uint4*>(dx) + base + vpos;
  // 1. rows of x and gy -> shared memory, {sum d, sum d * xhat} of the column on the way
  {
    int i = tid, r = vrow;
#pragma unroll 4
    for (; i < total; i += nt, r += vstep) {
      const uint4 vx = xg[(size_t)r * rstride], vg = gg[(size_t)r * rstride];
      slab_x[i] = vx;
      slab_g[i] = vg;
      const uint32_t ix[4] = {vx.x, vx.y, vx.z, vx.w}, ig[4] = {vg.x, vg.y, vg.z, vg.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float fx[2], fg[2];
        Word<T>::unpack(ix[j], fx);
        Word<T>::unpack(ig[j], fg);
#pragma unroll
        for (int e = 0; e < EW; ++e) {
          const int k = j * EW + e;
          const float xh = fmaf(fx[e], rs[k], sh[k]);
          float d = fg[e] * gam[k];
          if (p.silu) d *= silu_grad(fmaf(xh, gam[k], bet[k]));
          sd[k] += d;
          sdx[k] = fmaf(d, xh, sdx[k]);
        }
      }
    }
  }
  // 2. group sums: warp -> CTA (fp64) -> cluster, as in the forward
  for (int gi = 0; gi < ga.gp; ++gi) {
    float a = 0.f, q = 0.f;
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      a += ge[e] == gi ? sd[e] : 0.f;
      q += ge[e] == gi ? sdx[e] : 0.f;
    }
    a = warp_sum(a);
    q = warp_sum(q);
    if (lane == 0) s_part[warp * GN_GROUP_MAXGP + gi] = make_float2(a, q);
  }
  __syncthreads();
  if (warp < ga.gp) {
    double a = 0.0, q = 0.0;
    if (lane < nwarps) { const float2 v = s_part[lane * GN_GROUP_MAXGP + warp]; a = (double)v.x; q = (double)v.y; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (lane == 0) { s_cta[2 * warp] = a; s_cta[2 * warp + 1] = q; }
  }
  if (ga.k > 1) { ptx::cluster_arrive(); ptx::cluster_wait(); } else { __syncthreads(); }
  if (warp < ga.gp) {
    double a = 0.0, q = 0.0;
    if (ga.k > 1) {
      if (lane < ga.k) {
        const uint32_t peer = ptx::mapa(ptx::smem_u32(s_cta + 2 * warp), (uint32_t)lane);
        asm volatile("ld.shared::cluster.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(q) : "r"(peer));
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
      }
    } else {
      a = s_cta[2 * warp]; q = s_cta[2 * warp + 1];
    }
    if (lane == 0) {
      const double n = (double)p.cpg * (double)p.HW;
      s_c[2 * warp] = (float)(a / n); s_c[2 * warp + 1] = (float)(q / n);
    }
  }
  __syncthreads();
  // 3. dx from the shared copies; column sums of dx for dadd
  float c1[NE], c2[NE], dcol[NE];
#pragma unroll
  for (int e = 0; e < NE; ++e) { c1[e] = s_c[2 * ge[e]]; c2[e] = s_c[2 * ge[e] + 1]; dcol[e] = 0.f; }
  {
    int i = tid, r = vrow;
#pragma unroll 2
    for (; i < total; i += nt, r += vstep) {
      const uint4 vx = slab_x[i], vg = slab_g[i];
      const uint32_t ix[4] = {vx.x, vx.y, vx.z, vx.w}, ig[4] = {vg.x, vg.y, vg.z, vg.w};
      uint32_t out[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float fx[2], fg[2], u[2] = {0.f, 0.f};
        Word<T>::unpack(ix[j], fx);
        Word<T>::unpack(ig[j], fg);
#pragma unroll
        for (int e = 0; e < EW; ++e) {
          const int k = j * EW + e;
          const float xh = fmaf(fx[e], rs[k], sh[k]);
          float d = fg[e] * gam[k];
          if (p.silu) d *= silu_grad(fmaf(xh, gam[k], bet[k]));
          const float o = rs[k] * (d - c1[k] - xh * c2[k]);
          u[e] = o;
          dcol[k] += o;
        }
        out[j] = Word<T>::pack(u);
      }
      dg[(size_t)r * rstride] = make_uint4(out[0], out[1], out[2], out[3]);
    }
  }
  if (dadd != nullptr) {
#pragma unroll
    for (int e = 0; e < NE; ++e) s_col[(size_t)vrow * nch + vpos * NE + e] = dcol[e];
    __syncthreads();
    for (int c = tid; c < nch; c += nt) {
      float a = 0.f;
      for (int l = 0; l < vstep; ++l) a += s_col[(size_t)l * nch + c];
      s_colsum[c] = a;
    }
    if (ga.k > 1) { ptx::cluster_arrive(); ptx::cluster_wait(); } else { __syncthreads(); }
    if (part == 0) {
      for (int c = tid; c < nch; c += nt) {
        float a = 0.f;
        if (ga.k > 1) {
          for (int r = 0; r < ga.k; ++r) {
            float v;
            const uint32_t peer = ptx::mapa(ptx::smem_u32(s_colsum + c), (uint32_t)r);
            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(peer));
            a += v;
          }
        } else {
          a = s_colsum[c];
        }
        dadd[(size_t)b * p.C + c0 + c] = Elem<T>::from_f(a);
      }
    }
  }
  // the peers may still read this CTA's shared memory
  if (ga.k > 1) { ptx::cluster_arrive(); ptx::cluster_wait(); }
}

static int g_gn_bwd_group = [] { const char* e = getenv("SLIDER_GN_BWD_GROUP"); return e ? atoi(e) : 1; }();

template <typename T>
static bool gn_bwd_group_launch(int* rc, void* dx, void* dadd, const void* gy, const void* x, const void* chan_add,
                                const void* gamma, const void* beta, const float* stats, const GnParams& p,
                                cudaStream_t st) {
  GnGroupPlan gp;
  if (!g_gn_bwd_group || p.B > 65535 || !gn_group_plan<T>(p, &gp)) return false;
  const int nch = gp.a.gp * p.cpg;
  const int wpr = gp.a.wr >> 2;
  const int vstep = gp.threads / wpr;
  const size_t seg = (size_t)gp.a.wr * 4;
  const size_t smem = 2 * (size_t)gp.a.rows * seg + (GN_GROUP_THREADS / 32) * GN_GROUP_MAXGP * sizeof(float2) +
                      2 * GN_GROUP_MAXGP * sizeof(double) + 2 * GN_GROUP_MAXGP * sizeof(float) +
                      (size_t)nch * sizeof(float) + (size_t)vstep * nch * sizeof(float);
  if (smem > 200 * 1024) return false;
  static thread_local size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    if (cudaFuncSetAttribute(gn_bwd_group_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
      cudaGetLastError();
      return false;
    }
    configured = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((p.G / gp.a.gp) * gp.a.k, p.B); cfg.blockDim = dim3(gp.threads); cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (gp.a.k > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = (unsigned)gp.a.k; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaLaunchKernelEx(&cfg, gn_bwd_group_kernel<T>, (T*)dx, (T*)dadd, (const T*)gy, (const T*)x, (const T*)chan_add,
                     (const T*)gamma, (const T*)beta, stats, p, gp.a);
  *rc = post_launch("gn_bwd_group_kernel");
  return true;
}

template <typename T>
static int groupnorm_bwd_impl(void* dx, void* dadd, const void* gy, const void* x, const void* chan_add,
                              const void* gamma, const void* beta, const float* stats, float* ws, GnParams p,
                              cudaStream_t st) {
  {
    int rc = SLIDER_OK;
    if (gn_bwd_group_launch<T>(&rc, dx, dadd, gy, x, chan_add, gamma, beta, stats, p, st)) return rc;
  }
  float* partial_g = ws;                                             // [B][chunks][G][2]
  float* partial_c = ws + (size_t)p.B * p.chunks * p.G * 2;          // [B][chunks][C][2]
  const bool want_add = dadd != nullptr;
  {
    dim3 grid(p.chunks, p.B);
    const size_t smem = (size_t)p.PL * p.C * (want_add ? 3 : 2) * sizeof(float);
    static thread_local size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
      SLIDER_CUDA(cudaFuncSetAttribute(gn_bwd_stats_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      configured = smem;
    }
    int threads = p.PL * p.NV > p.G ? p.PL * p.NV : p.G;
    threads = (threads + 31) / 32 * 32;
    gn_bwd_stats_kernel<T><<<grid, threads, smem, st>>>((const T*)gy, (const T*)x, (const T*)chan_add, (const T*)gamma,
                                                        (const T*)beta, stats, partial_g, want_add ? partial_c : nullptr, p);
    int rc = post_launch("gn_bwd_stats_kernel");
    if (rc) return rc;
  }
  {
    int chunks = (sm_count() * 4 + p.B - 1) / p.B;
    int ppc = (p.HW + chunks - 1) / chunks;
    if (ppc < 4) ppc = 4;
    chunks = (p.HW + ppc - 1) / ppc;
    const int min_blocks = want_add ? (p.C + 255) / 256 : 1;     // the dadd slices need that many blocks per item
    if (chunks < min_blocks) chunks = min_blocks;
    dim3 grid(chunks, p.B);
    const size_t smem = ((size_t)p.C * 6 + (size_t)p.G * 2) * sizeof(float) + 256 * 2 * sizeof(double) + 16;
    static thread_local size_t configured2 = 0;
    if (smem > 48 * 1024 && smem > configured2) {
      SLIDER_CUDA(cudaFuncSetAttribute(gn_bwd_apply_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      configured2 = smem;
    }
    gn_bwd_apply_kernel<T><<<grid, 256, smem, st>>>((T*)dx, (T*)dadd, (const T*)gy, (const T*)x, (const T*)chan_add,
                                                    (const T*)gamma, (const T*)beta, stats, partial_g, partial_c, p,
                                                    chunks, ppc);
  }
  return post_launch("gn_bwd_apply_kernel");
}

static int gn_plan(GnParams* p, int B, int HW, int C, int G, float eps, int silu) {
  SLIDER_CHECK_ARG(B > 0 && HW > 0 && C > 0 && G > 0 && C % G == 0, "groupnorm: bad shape B %d HW %d C %d G %d", B, HW, C, G);
  SLIDER_CHECK_ARG(C % 8 == 0 && C <= 4096, "groupnorm: channels must be a multiple of 8 and <= 4096 (got %d)", C);
  SLIDER_CHECK_ARG(G <= 256, "groupnorm: too many groups (%d)", G);
  p->B = B; p->HW = HW; p->C = C; p->G = G; p->cpg = C / G; p->eps = eps; p->silu = silu;
  p->variant = g_gn_variant;
  p->NV = C / 8;
  p->PL = GN_MAX_THREADS / p->NV;
  if (p->PL < 1) p->PL = 1;
  if (p->PL > HW) p->PL = HW;
  int chunks = g_gn_variant == 3 ? (sm_count() / 2 + B - 1) / B : (g_gn_variant == 4 ? (sm_count() / 4 + B - 1) / B : (sm_count() + B - 1) / B);
  int ppc = (HW + chunks - 1) / chunks;
  if (ppc < p->PL) ppc = p->PL;
  chunks = (HW + ppc - 1) / ppc;
  p->chunks = chunks; p->ppc = ppc;
  return SLIDER_OK;
}

}  // namespace slider

using namespace slider;

extern "C" {

int slider_debug_gn_variant(int v) { g_gn_variant = v; return 0; }
// host-only view of the single-launch plan (no device needed): 1 and out = {k, gp, words per row, rows per CTA,
// threads, dynamic smem bytes} when the shape takes the single launch, 0 when it takes the two-kernel form
int slider_debug_gn_group_plan(int hw, int channels, int groups, int dtype, int* out) {
  if (hw <= 0 || channels <= 0 || groups <= 0 || channels % groups != 0 || out == nullptr) return -1;
  GnParams p = {};
  p.B = 1; p.HW = hw; p.C = channels; p.G = groups; p.cpg = channels / groups;
  GnGroupPlan gp;
  const bool ok = dtype == SLIDER_F32 ? gn_group_plan<float>(p, &gp) : gn_group_plan<__nv_bfloat16>(p, &gp);   // 16-bit types share the plan
  if (!ok) return 0;
  out[0] = gp.a.k; out[1] = gp.a.gp; out[2] = gp.a.wr; out[3] = gp.a.rows; out[4] = gp.threads; out[5] = (int)gp.smem;
  return 1;
}
int slider_debug_gn_trace(void* buf) { g_gn_trace = (unsigned long long*)buf; return 0; }
int slider_debug_gn_group(int on, int slab_kb) {
  g_gn_group = on;
  if (slab_kb >= 4 && slab_kb <= 128) g_gn_group_kb = slab_kb;
  return 0;
}

size_t slider_groupnorm_workspace_bytes(int batch, int groups) {
  // tickets + stats + partials (batch * chunks <= 2 * SMs + 2 * batch)
  return ((size_t)GN_TICKETS + (size_t)batch * groups * 2 +
          ((size_t)sm_count() * 4 + 2 * (size_t)batch) * groups * 2) * sizeof(float);
}

int slider_groupnorm_nhwc(void* y, const void* x, const void* chan_add, const void* gamma, const void* beta,
                          int batch, int hw, int channels, int groups, float eps, int silu, int dtype, void* ws,
                          size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(y && x && ws, "groupnorm: null pointer");
  SLIDER_CHECK_ARG((((uintptr_t)y | (uintptr_t)x | (uintptr_t)chan_add | (uintptr_t)ws) & 15) == 0,
                   "groupnorm: pointers must be 16-byte aligned");
  GnParams p;
  int rc = gn_plan(&p, batch, hw, channels, groups, eps, silu);
  if (rc) return rc;
  SLIDER_CHECK_ARG(p.NV <= GN_MAX_THREADS, "groupnorm: channels %d exceed the kernel's 4096 limit", channels);
  SLIDER_CHECK_ARG(batch <= GN_TICKETS, "groupnorm: batch %d > %d", batch, GN_TICKETS);
  const size_t need = ((size_t)GN_TICKETS + (size_t)batch * groups * 2 + (size_t)batch * p.chunks * groups * 2) * sizeof(float);
  if (ws_bytes < need) {
    set_error("groupnorm: workspace %zu < %zu bytes", ws_bytes, need);
    return SLIDER_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLIDER_BF16) return groupnorm_impl<__nv_bfloat16>(y, x, chan_add, gamma, beta, (float*)ws, p, st);
  if (dtype == SLIDER_F32) return groupnorm_impl<float>(y, x, chan_add, gamma, beta, (float*)ws, p, st);
  if (dtype == SLIDER_F16) return groupnorm_impl<__half>(y, x, chan_add, gamma, beta, (float*)ws, p, st);
  set_error("groupnorm: unsupported dtype %d", dtype);
  return SLIDER_ERR_UNSUPPORTED;
}

int slider_groupnorm_nhwc_train(void* y, float* stats_out, const void* x, const void* chan_add, const void* gamma,
                                const void* beta, int batch, int hw, int channels, int groups, float eps, int silu,
                                int dtype, void* ws, size_t ws_bytes, void* stream) {
  SLIDER_CHECK_ARG(y && x && ws && stats_out, "groupnorm_train: null pointer");
  SLIDER_CHECK_ARG((((uintptr_t)y | (uintptr_t)x | (uintptr_t)chan_add | (uintptr_t)ws) & 15) == 0,
                   "groupnorm_train: pointers must be 16-byte aligned");
  GnParams p;
  int rc = gn_plan(&p, batch, hw, channels, groups, eps, silu);
  if (rc) return rc;
  SLIDER_CHECK_ARG(p.NV <= GN_MAX_THREADS && batch <= GN_TICKETS, "groupnorm_train: shape outside the kernel's limits");
  const size_t need = ((size_t)GN_TICKETS + (size_t)batch * groups * 2 + (size_t)batch * p.chunks * groups * 2) * sizeof(float);
  if (ws_bytes < need) {
    set_error("groupnorm_train: workspace %zu < %zu bytes", ws_bytes, need);
    return SLIDER_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLIDER_BF16) return groupnorm_impl<__nv_bfloat16>(y, x, chan_add, gamma, beta, (float*)ws, p, st, stats_out);
  if (dtype == SLIDER_F32) return groupnorm_impl<float>(y, x, chan_add, gamma, beta, (float*)ws, p, st, stats_out);
  set_error("groupnorm_train: unsupported dtype %d", dtype);
  return SLIDER_ERR_UNSUPPORTED;
}

size_t slider_groupnorm_bwd_workspace_bytes(int batch, int channels, int groups) {
  const size_t chunks = (size_t)sm_count() + 2 * (size_t)batch;       // batch * chunks upper bound (gn_plan)
  return chunks * ((size_t)groups + (size_t)channels) * 2 * sizeof(float);
}

int slider_groupnorm_nhwc_backward(void* dx, void* dadd, const void* gy, const void* x, const void* chan_add,
                                   const void* gamma, const void* beta, const float* stats, int batch, int hw,
                                   int channels, int groups, int silu, int dtype, void* ws, size_t ws_bytes,
                                   void* stream) {
  SLIDER_CHECK_ARG(dx && gy && x && stats && ws, "groupnorm_backward: null pointer");
  SLIDER_CHECK_ARG((dadd == nullptr) || (chan_add != nullptr), "groupnorm_backward: dadd needs chan_add");
  SLIDER_CHECK_ARG((((uintptr_t)dx | (uintptr_t)gy | (uintptr_t)x | (uintptr_t)chan_add | (uintptr_t)dadd | (uintptr_t)ws) & 15) == 0,
                   "groupnorm_backward: pointers must be 16-byte aligned");
  GnParams p;
  int rc = gn_plan(&p, batch, hw, channels, groups, 0.f, silu);
  if (rc) return rc;
  SLIDER_CHECK_ARG(p.NV <= GN_MAX_THREADS, "groupnorm_backward: channels %d exceed the kernel's 4096 limit", channels);
  // three per-channel accumulator arrays in shared memory: fewer pixel lanes than the forward when needed
  while (p.PL > 1 && (size_t)p.PL * p.C * 3 * sizeof(float) > 160 * 1024) p.PL /= 2;
  const size_t need = (size_t)batch * p.chunks * ((size_t)groups + channels) * 2 * sizeof(float);
  if (ws_bytes < need) {
    set_error("groupnorm_backward: workspace %zu < %zu bytes", ws_bytes, need);
    return SLIDER_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLIDER_BF16)
    return groupnorm_bwd_impl<__nv_bfloat16>(dx, dadd, gy, x, chan_add, gamma, beta, stats, (float*)ws, p, st);
  if (dtype == SLIDER_F32) return groupnorm_bwd_impl<float>(dx, dadd, gy, x, chan_add, gamma, beta, stats, (float*)ws, p, st);
  set_error("groupnorm_backward: unsupported dtype %d", dtype);
  return SLIDER_ERR_UNSUPPORTED;
}

}  // extern "C"
