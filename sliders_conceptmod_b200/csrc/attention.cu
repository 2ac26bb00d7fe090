// Cross-attention core for SHORT key sets (the 77 prompt tokens of the UNet's attn2 layers): softmax(Q K^T / sqrt(d)) V
// with every key of a (batch, head) resident in shared memory, so there is no online-softmax loop — one S tile, one
// softmax, one P.V.  Reference call site: the attn2 layers of the UNet evaluated by train_util.predict_noise
// (conceptmod/textsliders/train_util.py:290-294; diffusers Attention with encoder_hidden_states = the 77-token prompt).
//
// Why not the library kernel: at 77 keys the layer is 0.8 GFLOP and 10 MB of traffic, yet the flash / cuDNN /
// mem-efficient SDPA kernels torch picks take 7-18 us per layer (r02 in-graph profile: 253 us per denoise forward for the
// 16 attn2 layers), and their [B, H, L, D] layout costs a transposing copy of the output on top.
//
//   CTA = (batch, head, 64 queries), 4 warps x 16 queries, mma.sync m16n8k16 (16-bit operands, fp32 accumulate)
//   K, V [80 rows (77 keys + zero rows)][DP] and Q [64][DP] in shared memory, DP = head dim padded to a multiple of 16
//   (zero columns), row pitch DP * 2 + 16 bytes (conflict-free ldmatrix)
//   S = Q K^T: 10 key tiles x DP/16 K steps -> 40 fp32 registers per thread; keys >= kv_len masked
//   softmax in registers (a row lives in one quad: two xor-shuffles), P rounded to the storage type and used directly
//   as the A operand of P.V (accumulator layout of two adjacent key tiles == A fragment of one 16-key K step)
//   O = P V (V through ldmatrix.trans), normalised by the row sum, stored straight in the [B, L, H * D] layout the
//   output projection reads — q / k / v are read in that layout as well (no head transposes at all)
#include "common.cuh"

namespace slider {

constexpr int XA_KEYS = 80;       // key rows staged (kv_len <= 80)
// queries per CTA: template parameter QT (64 or 128), one warp per 16 queries

template <typename T> struct XaMma;
template <> struct XaMma<__nv_bfloat16> {
  __device__ static __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
};
template <> struct XaMma<__half> {
  __device__ static __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
};

__device__ __forceinline__ void xa_ldmatrix_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void xa_ldmatrix_x2(uint32_t addr, uint32_t (&r)[2]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void xa_ldmatrix_x2_trans(uint32_t addr, uint32_t (&r)[2]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}

struct XaParams {
  int B, L, S, H, D;      // batch, queries, keys, heads, head dim
  int ldq, ldk, ldv, ldo; // row strides in elements (H * D for dense tensors; 3 * H * D for the views of a fused q/k/v launch)
  float scale_log2;       // log2(e) / sqrt(D)
  int kv_static;          // k / v are not written by any kernel in flight: stage them before the dependent-launch wait
};

// DS = DP / 16 K steps of Q K^T (DP = padded head dim)
template <typename T, int DS, int QT>
__global__ void __launch_bounds__(QT * 2)
cross_attention_kernel(T* __restrict__ o, const T* __restrict__ q, const T* __restrict__ k, const T* __restrict__ v,
                       const XaParams p) {
  constexpr int DP = DS * 16;
  constexpr int PITCH = DP * 2 + 16;              // bytes
  constexpr int VPR = DP / 8;                     // 16-byte vectors per staged row
  constexpr int XA_QT = QT, XA_THREADS = QT * 2;
  extern __shared__ __align__(16) unsigned char xa_smem[];
  unsigned char* sK = xa_smem;
  unsigned char* sV = sK + XA_KEYS * PITCH;
  unsigned char* sQ = sV + XA_KEYS * PITCH;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const int q0 = blockIdx.x * XA_QT;
  pdl_trigger();
  const int dv = p.D >> 3;                        // live vectors per row
  // staging: the global loads of a batch (four rounds of the CTA) are all issued before the first shared-memory store,
  // so a batch costs one memory round trip instead of one per round
  auto stage_kv = [&]() {
    constexpr int N = XA_KEYS * VPR;
    for (int i0 = tid; i0 < N; i0 += 4 * XA_THREADS) {
      uint4 kk[4], vv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        const int r = i / VPR, c = i - r * VPR;
        kk[u] = make_uint4(0u, 0u, 0u, 0u); vv[u] = kk[u];
        if (i < N && r < p.S && c < dv) {
          const size_t row = (size_t)b * p.S + r, col = (size_t)h * p.D + c * 8;
          kk[u] = *reinterpret_cast<const uint4*>(k + row * p.ldk + col);
          vv[u] = *reinterpret_cast<const uint4*>(v + row * p.ldv + col);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        if (i < N) {
          const int r = i / VPR, c = i - r * VPR;
          *reinterpret_cast<uint4*>(sK + r * PITCH + c * 16) = kk[u];
          *reinterpret_cast<uint4*>(sV + r * PITCH + c * 16) = vv[u];
        }
      }
    }
  };
  if (p.kv_static) stage_kv();
  pdl_wait();
  if (!p.kv_static) stage_kv();
  {
    constexpr int N = XA_QT * VPR;
    for (int i0 = tid; i0 < N; i0 += 4 * XA_THREADS) {
      uint4 qq[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        const int r = i / VPR, c = i - r * VPR;
        qq[u] = make_uint4(0u, 0u, 0u, 0u);
        if (i < N && q0 + r < p.L && c < dv)
          qq[u] = *reinterpret_cast<const uint4*>(q + ((size_t)b * p.L + q0 + r) * p.ldq + (size_t)h * p.D + c * 8);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        if (i < N) {
          const int r = i / VPR, c = i - r * VPR;
          *reinterpret_cast<uint4*>(sQ + r * PITCH + c * 16) = qq[u];
        }
      }
    }
  }
  __syncthreads();
  const uint32_t sK_u = (uint32_t)__cvta_generic_to_shared(sK), sV_u = (uint32_t)__cvta_generic_to_shared(sV);
  const uint32_t sQ_u = (uint32_t)__cvta_generic_to_shared(sQ);
  const int g = lane >> 2, t = lane & 3;

  // ---- S = Q K^T: this warp's 16 queries x 80 keys
  uint32_t qa[DS][4];
#pragma unroll
  for (int ks = 0; ks < DS; ++ks)
    xa_ldmatrix_x4(sQ_u + (uint32_t)((warp * 16 + (lane & 15)) * PITCH + (ks * 16 + (lane >> 4) * 8) * 2), qa[ks]);
  float s[XA_KEYS / 8][4];
#pragma unroll
  for (int j = 0; j < XA_KEYS / 8; ++j) {
    s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < DS; ++ks) {
      uint32_t kb[2];
      // B fragment: n = key 8j + (lane & 7), k = d: matrix 0 = d 16ks .. +7, matrix 1 = d 16ks + 8 .. +15
      xa_ldmatrix_x2(sK_u + (uint32_t)((j * 8 + (lane & 7)) * PITCH + (ks * 16 + ((lane >> 3) & 1) * 8) * 2), kb);
      XaMma<T>::mma(s[j], qa[ks], kb);
    }
  }
  // ---- softmax over the live keys; rows g and g + 8 of the warp's 16 queries
  float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
  for (int j = 0; j < XA_KEYS / 8; ++j) {
    const int key = j * 8 + 2 * t;
    if (key >= p.S) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
    if (key + 1 >= p.S) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
    m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
    m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
  }
  m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
  m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
  float l0 = 0.f, l1 = 0.f;
  uint32_t pa[XA_KEYS / 16][4];                   // P as A fragments: K step kk covers keys 16kk .. 16kk + 15
#pragma unroll
  for (int j = 0; j < XA_KEYS / 8; ++j) {
    const float e0 = ex2_approx((s[j][0] - m0) * p.scale_log2), e1 = ex2_approx((s[j][1] - m0) * p.scale_log2);
    const float e2 = ex2_approx((s[j][2] - m1) * p.scale_log2), e3 = ex2_approx((s[j][3] - m1) * p.scale_log2);
    // the row sum uses the ROUNDED probabilities, so that P / l is exactly what multiplies V
    const uint32_t w01 = Elem<T>::pack2(e0, e1), w23 = Elem<T>::pack2(e2, e3);
    float r0, r1, r2, r3;
    Elem<T>::unpack2(w01, r0, r1);
    Elem<T>::unpack2(w23, r2, r3);
    l0 += r0 + r1; l1 += r2 + r3;
    pa[j >> 1][(j & 1) * 2 + 0] = w01;            // (row g,     keys 8j + 2t, +1)
    pa[j >> 1][(j & 1) * 2 + 1] = w23;            // (row g + 8, keys 8j + 2t, +1)
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = 1.f / l0, inv1 = 1.f / l1;

  // ---- O = P V, one 8-wide slice of the head dim at a time
  const int row0 = q0 + warp * 16 + g, row1 = row0 + 8;
  T* o0 = o + ((size_t)b * p.L + row0) * p.ldo + (size_t)h * p.D;
  T* o1 = o + ((size_t)b * p.L + row1) * p.ldo + (size_t)h * p.D;
#pragma unroll
  for (int jd = 0; jd < DP / 8; ++jd) {
    if (jd * 8 >= p.D) break;                     // padded columns (uniform over the CTA)
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int kk = 0; kk < XA_KEYS / 16; ++kk) {
      uint32_t vb[2];
      xa_ldmatrix_x2_trans(sV_u + (uint32_t)((kk * 16 + (lane & 15)) * PITCH + jd * 16), vb);
      XaMma<T>::mma(acc, pa[kk], vb);
    }
    const int d = jd * 8 + 2 * t;
    if (row0 < p.L) *reinterpret_cast<uint32_t*>(o0 + d) = Elem<T>::pack2(acc[0] * inv0, acc[1] * inv0);
    if (row1 < p.L) *reinterpret_cast<uint32_t*>(o1 + d) = Elem<T>::pack2(acc[2] * inv1, acc[3] * inv1);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// Backward for FROZEN keys / values (attn2 under the `noxattn` training methods: to_k / to_v are neither adapted nor
// trainable and the prompt embeddings carry no gradient): only dq is needed, and with every key resident it is the
// forward's structure again, nothing saved but q, k, v:
//   P = softmax(q k^T c) recomputed;  dP = dO V^T;  D_i = sum_j P_ij dP_ij  (= dO_i . O_i, so O is not needed);
//   dS = P o (dP - D) * c;  dq = dS K.
// S and dP are two "Q K^T"-type products (dO staged like Q, V read like K), dq one "P V"-type product with K through
// ldmatrix.trans.  The library's backward for these shapes takes 40-150 us per layer (it also forms dK / dV).
template <typename T, int DS, int QT>
__global__ void __launch_bounds__(QT * 2)
cross_attention_bwd_kernel(T* __restrict__ dq, const T* __restrict__ dout, const T* __restrict__ q,
                           const T* __restrict__ k, const T* __restrict__ v, const XaParams p) {
  constexpr int DP = DS * 16;
  constexpr int PITCH = DP * 2 + 16;
  constexpr int VPR = DP / 8;
  constexpr int XA_THREADS = QT * 2;
  extern __shared__ __align__(16) unsigned char xa_smem[];
  unsigned char* sK = xa_smem;
  unsigned char* sV = sK + XA_KEYS * PITCH;
  unsigned char* sQ = sV + XA_KEYS * PITCH;
  unsigned char* sG = sQ + QT * PITCH;            // dO tile
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const int q0 = blockIdx.x * QT;
  const int dv = p.D >> 3;
  {
    constexpr int N = XA_KEYS * VPR;
    for (int i0 = tid; i0 < N; i0 += 4 * XA_THREADS) {
      uint4 kk[4], vv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        const int r = i / VPR, c = i - r * VPR;
        kk[u] = make_uint4(0u, 0u, 0u, 0u); vv[u] = kk[u];
        if (i < N && r < p.S && c < dv) {
          const size_t row = (size_t)b * p.S + r, col = (size_t)h * p.D + c * 8;
          kk[u] = *reinterpret_cast<const uint4*>(k + row * p.ldk + col);
          vv[u] = *reinterpret_cast<const uint4*>(v + row * p.ldv + col);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        if (i < N) {
          *reinterpret_cast<uint4*>(sK + (i / VPR) * PITCH + (i % VPR) * 16) = kk[u];
          *reinterpret_cast<uint4*>(sV + (i / VPR) * PITCH + (i % VPR) * 16) = vv[u];
        }
      }
    }
  }
  {
    constexpr int N = QT * VPR;
    for (int i0 = tid; i0 < N; i0 += 4 * XA_THREADS) {
      uint4 qq[4], gg[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        const int r = i / VPR, c = i - r * VPR;
        qq[u] = make_uint4(0u, 0u, 0u, 0u); gg[u] = qq[u];
        if (i < N && q0 + r < p.L && c < dv) {
          const size_t row = (size_t)b * p.L + q0 + r, col = (size_t)h * p.D + c * 8;
          qq[u] = *reinterpret_cast<const uint4*>(q + row * p.ldq + col);
          gg[u] = *reinterpret_cast<const uint4*>(dout + row * p.ldo + col);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * XA_THREADS;
        if (i < N) {
          *reinterpret_cast<uint4*>(sQ + (i / VPR) * PITCH + (i % VPR) * 16) = qq[u];
          *reinterpret_cast<uint4*>(sG + (i / VPR) * PITCH + (i % VPR) * 16) = gg[u];
        }
      }
    }
  }
  __syncthreads();
  const uint32_t sK_u = (uint32_t)__cvta_generic_to_shared(sK), sV_u = (uint32_t)__cvta_generic_to_shared(sV);
  const uint32_t sQ_u = (uint32_t)__cvta_generic_to_shared(sQ), sG_u = (uint32_t)__cvta_generic_to_shared(sG);
  const int g = lane >> 2, t = lane & 3;
  float s[XA_KEYS / 8][4], dp[XA_KEYS / 8][4];
  {
    uint32_t qa[DS][4], ga[DS][4];
#pragma unroll
    for (int ks = 0; ks < DS; ++ks) {
      const uint32_t off = (uint32_t)((warp * 16 + (lane & 15)) * PITCH + (ks * 16 + (lane >> 4) * 8) * 2);
      xa_ldmatrix_x4(sQ_u + off, qa[ks]);
      xa_ldmatrix_x4(sG_u + off, ga[ks]);
    }
#pragma unroll
    for (int j = 0; j < XA_KEYS / 8; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      dp[j][0] = dp[j][1] = dp[j][2] = dp[j][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < DS; ++ks) {
        uint32_t kb[2], vb[2];
        const uint32_t off = (uint32_t)((j * 8 + (lane & 7)) * PITCH + (ks * 16 + ((lane >> 3) & 1) * 8) * 2);
        xa_ldmatrix_x2(sK_u + off, kb);
        xa_ldmatrix_x2(sV_u + off, vb);
        XaMma<T>::mma(s[j], qa[ks], kb);
        XaMma<T>::mma(dp[j], ga[ks], vb);
      }
    }
  }
  float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
  for (int j = 0; j < XA_KEYS / 8; ++j) {
    const int key = j * 8 + 2 * t;
    if (key >= p.S) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
    if (key + 1 >= p.S) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
    m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
    m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
  }
  m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
  m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
  float l0 = 0.f, l1 = 0.f, d0 = 0.f, d1 = 0.f;
#pragma unroll
  for (int j = 0; j < XA_KEYS / 8; ++j) {
    s[j][0] = ex2_approx((s[j][0] - m0) * p.scale_log2); s[j][1] = ex2_approx((s[j][1] - m0) * p.scale_log2);
    s[j][2] = ex2_approx((s[j][2] - m1) * p.scale_log2); s[j][3] = ex2_approx((s[j][3] - m1) * p.scale_log2);
    l0 += s[j][0] + s[j][1]; l1 += s[j][2] + s[j][3];
    d0 += s[j][0] * dp[j][0] + s[j][1] * dp[j][1];          // masked keys: p = 0 and dP = dO . 0 = 0
    d1 += s[j][2] * dp[j][2] + s[j][3] * dp[j][3];
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  d0 += __shfl_xor_sync(0xffffffffu, d0, 1); d0 += __shfl_xor_sync(0xffffffffu, d0, 2);
  d1 += __shfl_xor_sync(0xffffffffu, d1, 1); d1 += __shfl_xor_sync(0xffffffffu, d1, 2);
  const float inv0 = 1.f / l0, inv1 = 1.f / l1;
  const float D0 = d0 * inv0, D1 = d1 * inv1;               // sum_j P_ij dP_ij with P = e / l
  const float c = p.scale_log2 * 0.6931471805599453f;       // 1 / sqrt(d)
  uint32_t da[XA_KEYS / 16][4];                             // dS as A fragments
#pragma unroll
  for (int j = 0; j < XA_KEYS / 8; ++j) {
    const float e0 = s[j][0] * inv0 * (dp[j][0] - D0) * c, e1 = s[j][1] * inv0 * (dp[j][1] - D0) * c;
    const float e2 = s[j][2] * inv1 * (dp[j][2] - D1) * c, e3 = s[j][3] * inv1 * (dp[j][3] - D1) * c;
    da[j >> 1][(j & 1) * 2 + 0] = Elem<T>::pack2(e0, e1);
    da[j >> 1][(j & 1) * 2 + 1] = Elem<T>::pack2(e2, e3);
  }
  const int row0 = q0 + warp * 16 + g, row1 = row0 + 8;
  T* o0 = dq + ((size_t)b * p.L + row0) * p.ldq + (size_t)h * p.D;
  T* o1 = dq + ((size_t)b * p.L + row1) * p.ldq + (size_t)h * p.D;
#pragma unroll
  for (int jd = 0; jd < DP / 8; ++jd) {
    if (jd * 8 >= p.D) break;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int kk = 0; kk < XA_KEYS / 16; ++kk) {
      uint32_t kb[2];
      xa_ldmatrix_x2_trans(sK_u + (uint32_t)((kk * 16 + (lane & 15)) * PITCH + jd * 16), kb);
      XaMma<T>::mma(acc, da[kk], kb);
    }
    const int d = jd * 8 + 2 * t;
    if (row0 < p.L) *reinterpret_cast<uint32_t*>(o0 + d) = Elem<T>::pack2(acc[0], acc[1]);
    if (row1 < p.L) *reinterpret_cast<uint32_t*>(o1 + d) = Elem<T>::pack2(acc[2], acc[3]);
  }
}

template <typename T, int DS>
static int launch_xattn_bwd(void* dq, const void* dout, const void* q, const void* k, const void* v, const XaParams& p,
                            cudaStream_t st) {
  constexpr int QT = 64;
  constexpr int PITCH = DS * 32 + 16;
  const size_t smem = (size_t)(2 * XA_KEYS + 2 * QT) * PITCH;
  auto kern = cross_attention_bwd_kernel<T, DS, QT>;
  static thread_local bool configured = false;
  if (!configured && smem > 48 * 1024) {
    SLIDER_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  const dim3 grid((p.L + QT - 1) / QT, p.B * p.H);
  kern<<<grid, QT * 2, smem, st>>>((T*)dq, (const T*)dout, (const T*)q, (const T*)k, (const T*)v, p);
  return post_launch("cross_attention_bwd_kernel");
}

template <typename T>
static int xattn_bwd_dispatch(void* dq, const void* dout, const void* q, const void* k, const void* v, const XaParams& p,
                              cudaStream_t st) {
  const int ds = (p.D + 15) / 16;
  if (ds <= 3) return launch_xattn_bwd<T, 3>(dq, dout, q, k, v, p, st);
  if (ds <= 4) return launch_xattn_bwd<T, 4>(dq, dout, q, k, v, p, st);
  if (ds <= 5) return launch_xattn_bwd<T, 5>(dq, dout, q, k, v, p, st);
  if (ds <= 8) return launch_xattn_bwd<T, 8>(dq, dout, q, k, v, p, st);
  return launch_xattn_bwd<T, 10>(dq, dout, q, k, v, p, st);
}

template <typename T, int DS, int QT>
static int launch_xattn_q(void* o, const void* q, const void* k, const void* v, const XaParams& p, cudaStream_t st) {
  constexpr int PITCH = DS * 32 + 16;
  const size_t smem = (size_t)(2 * XA_KEYS + QT) * PITCH;
  auto kern = cross_attention_kernel<T, DS, QT>;
  static thread_local bool configured = false;
  if (!configured && smem > 48 * 1024) {
    SLIDER_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  const dim3 grid((p.L + QT - 1) / QT, p.B * p.H);
  launch_pdl(kern, grid, dim3(QT * 2), smem, st, (T*)o, (const T*)q, (const T*)k, (const T*)v, p);
  return post_launch("cross_attention_kernel");
}

// 128 queries per CTA halve the K / V staging per query where there are thousands of them (8.6 vs 10.5 us at 2 x 4096
// queries, 21 vs 26 us at batch 6); with a few hundred queries the smaller tile supplies the CTAs (6.0 vs 7.6 us at 256)
template <typename T, int DS>
static int launch_xattn(void* o, const void* q, const void* k, const void* v, const XaParams& p, cudaStream_t st) {
  if (p.L >= 2048) return launch_xattn_q<T, DS, 128>(o, q, k, v, p, st);
  return launch_xattn_q<T, DS, 64>(o, q, k, v, p, st);
}

template <typename T>
static int xattn_dispatch(void* o, const void* q, const void* k, const void* v, const XaParams& p, cudaStream_t st) {
  const int ds = (p.D + 15) / 16;
  if (ds <= 3) return launch_xattn<T, 3>(o, q, k, v, p, st);
  if (ds <= 4) return launch_xattn<T, 4>(o, q, k, v, p, st);
  if (ds <= 5) return launch_xattn<T, 5>(o, q, k, v, p, st);
  if (ds <= 8) return launch_xattn<T, 8>(o, q, k, v, p, st);
  return launch_xattn<T, 10>(o, q, k, v, p, st);
}

}  // namespace slider

using namespace slider;

static int attention_entry(void* o, const void* q, const void* k, const void* v, int batch, int q_len, int kv_len,
                           int heads, int head_dim, int ldq, int ldk, int ldv, int ldo, int dtype, int kv_static,
                           void* stream, const char* who) {
  SLIDER_CHECK_ARG(o && q && k && v, "%s: null pointer", who);
  SLIDER_CHECK_ARG(batch > 0 && q_len > 0 && heads > 0, "%s: empty shape", who);
  SLIDER_CHECK_ARG(kv_len >= 1 && kv_len <= XA_KEYS, "%s: %d keys (this kernel serves 1..%d resident keys; longer key sets are "
                   "the library attention's job)", who, kv_len, XA_KEYS);
  SLIDER_CHECK_ARG(head_dim >= 8 && head_dim <= 160 && head_dim % 8 == 0, "%s: head dim %d not a multiple of 8 in 8..160", who,
                   head_dim);
  SLIDER_CHECK_ARG(batch * heads <= 65535, "%s: batch * heads = %d exceeds the grid limit", who, batch * heads);
  const int hd = heads * head_dim;
  SLIDER_CHECK_ARG(ldq >= hd && ldk >= hd && ldv >= hd && ldo >= hd && ((ldq | ldk | ldv | ldo) & 7) == 0,
                   "%s: row strides must be >= heads * head_dim and multiples of 8 elements", who);
  SLIDER_CHECK_ARG((((uintptr_t)o | (uintptr_t)q | (uintptr_t)k | (uintptr_t)v) & 15) == 0, "%s: pointers must be 16-byte aligned", who);
  XaParams p;
  p.B = batch; p.L = q_len; p.S = kv_len; p.H = heads; p.D = head_dim;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv; p.ldo = ldo;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)head_dim);
  p.kv_static = kv_static ? 1 : 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLIDER_BF16) return xattn_dispatch<__nv_bfloat16>(o, q, k, v, p, st);
  if (dtype == SLIDER_F16) return xattn_dispatch<__half>(o, q, k, v, p, st);
  set_error("%s: unsupported dtype %d (16-bit storage only)", who, dtype);
  return SLIDER_ERR_UNSUPPORTED;
}

extern "C" int slider_cross_attention(void* o, const void* q, const void* k, const void* v, int batch, int q_len,
                                      int kv_len, int heads, int head_dim, int dtype, int kv_static, void* stream) {
  const int hd = heads * head_dim;
  return attention_entry(o, q, k, v, batch, q_len, kv_len, heads, head_dim, hd, hd, hd, hd, dtype, kv_static, stream,
                         "cross_attention");
}

extern "C" int slider_cross_attention_strided(void* o, const void* q, const void* k, const void* v, int batch, int q_len,
                                              int kv_len, int heads, int head_dim, int ldq, int ldk, int ldv, int ldo,
                                              int dtype, int kv_static, void* stream) {
  return attention_entry(o, q, k, v, batch, q_len, kv_len, heads, head_dim, ldq, ldk, ldv, ldo, dtype, kv_static, stream,
                         "cross_attention_strided");
}

extern "C" int slider_cross_attention_backward(void* dq, const void* dout, const void* q, const void* k, const void* v,
                                               int batch, int q_len, int kv_len, int heads, int head_dim, int dtype,
                                               void* stream) {
  SLIDER_CHECK_ARG(dq && dout && q && k && v, "cross_attention_backward: null pointer");
  SLIDER_CHECK_ARG(batch > 0 && q_len > 0 && heads > 0, "cross_attention_backward: empty shape");
  SLIDER_CHECK_ARG(kv_len >= 1 && kv_len <= XA_KEYS, "cross_attention_backward: %d keys (1..%d)", kv_len, XA_KEYS);
  SLIDER_CHECK_ARG(head_dim >= 8 && head_dim <= 160 && head_dim % 8 == 0,
                   "cross_attention_backward: head dim %d not a multiple of 8 in 8..160", head_dim);
  SLIDER_CHECK_ARG(batch * heads <= 65535, "cross_attention_backward: batch * heads = %d exceeds the grid limit", batch * heads);
  SLIDER_CHECK_ARG((((uintptr_t)dq | (uintptr_t)dout | (uintptr_t)q | (uintptr_t)k | (uintptr_t)v) & 15) == 0,
                   "cross_attention_backward: pointers must be 16-byte aligned");
  XaParams p;
  p.B = batch; p.L = q_len; p.S = kv_len; p.H = heads; p.D = head_dim;
  p.ldq = p.ldk = p.ldv = p.ldo = heads * head_dim;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)head_dim);
  p.kv_static = 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLIDER_BF16) return xattn_bwd_dispatch<__nv_bfloat16>(dq, dout, q, k, v, p, st);
  if (dtype == SLIDER_F16) return xattn_bwd_dispatch<__half>(dq, dout, q, k, v, p, st);
  set_error("cross_attention_backward: unsupported dtype %d (16-bit storage only)", dtype);
  return SLIDER_ERR_UNSUPPORTED;
}
