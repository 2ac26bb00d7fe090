// Fused base + low-rank implicit GEMM for sm_100a (tcgen05 + TMEM + TMA): device code.
//
//   y[pix, n] = sum_{tap,c} x[stride*pix + d(tap), c] * W[n, tap, c] + bias[n]
//             + s * sum_j ( sum_{tap,c} x[stride*pix + d(tap), c] * A[j, tap, c] ) * Bup[n, j]
//
// One persistent CTA per SM, 10 warps:
//   warp 0      TMA producer  (4-D box loads of NHWC activations: padding = TMA out-of-bounds zero fill, a stride-2
//                              convolution = the activation map's element strides {1, 2, 2, 1};
//                              3-D box loads of the tap-major weights)
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer
//   warps 2..9  epilogue (two groups of four warps, one column half each):
//               tcgen05.ld -> registers -> (+bias, + s*h.Bup) -> shared staging -> TMA store
// The rank-r intermediate h = x (*) A lives in 16 extra TMEM columns next to the main accumulator: the lora_down
// tile sits behind the W tile in the same shared-memory stage, so ONE tcgen05.mma of N = BN + 16 produces both, and
// h is folded into the output in the epilogue with r FMAs per element.  The accumulators are double buffered; the
// epilogue reads a whole accumulator into registers with ONE tcgen05.wait and hands the TMEM stage back before it
// does any arithmetic, so the main loop of the next tile never waits for stores.
//
// Two tile widths are instantiated (lora_gemm_bn160.cu / lora_gemm_bn80.cu): 128 x 160 for launches with enough
// tiles to fill the GPU, 128 x 80 (six smaller operand stages) for tile-starved ones.
//
// The same kernel computes the input gradient ("extended-K" mode): activations = gy, weights = Wt, and one extra
// 16-wide K block per tap whose operands are dh and At.
#pragma once
#include <stdlib.h>
#include "common.cuh"
#include "ptx.cuh"

namespace slider {

constexpr int BM = 128;          // pixels (GEMM rows) per tile
constexpr int HCOLS = 16;        // TMEM columns of the low-rank accumulator (rank padded to 16)
constexpr int ROW_BYTES = 128;   // K-block row = one 128-byte swizzle row
constexpr int ACC_STAGES = 2;
constexpr int A_BYTES = BM * ROW_BYTES;      // 16 KB
constexpr int L_BYTES = HCOLS * ROW_BYTES;   // 2 KB
constexpr int EXT_ELEMS = 16;    // width of the extended-K operand (dh / At), in elements
constexpr int kGemmThreads = 320;            // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two groups of four)
constexpr int kEpiThreads = 256;
constexpr int OUT_ROW_BYTES = 80;            // staging row of one store chunk: 40 bf16 / 20 fp32 columns; an 80-byte
                                             // pitch makes the 16-byte st.shared of 8 consecutive rows conflict-free
constexpr int OUT_BYTES = BM * OUT_ROW_BYTES;   // 10 KB per chunk buffer
constexpr int PCW = OUT_ROW_BYTES / 4;       // fp32 columns per split-K partial chunk
constexpr int MAX_NSTAGES = 6;

// Everything that depends on the tile width BN (output channels per tile; 160 divides 320/640/960/1280/1920/2560).
template <int BN_> struct TileCfg {
  static constexpr int BN = BN_;
  static constexpr int B_BYTES = BN * ROW_BYTES;                     // 20 KB / 10 KB
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + L_BYTES;    // 38 KB / 28 KB
  static constexpr int NSTAGES = BN_ >= 160 ? 4 : 6;
  static constexpr int HALF_N = BN / 2;                              // output columns owned by one epilogue group
  static constexpr int ACC_COLS = BN + HCOLS;                        // TMEM columns of one accumulator stage
  static constexpr int TMEM_COLS = ACC_STAGES * ACC_COLS <= 256 ? 256 : 512;
  static constexpr int NCHUNK = BN / PCW + 1;                        // partial chunks of a split-K unit (+1 for h)
  static constexpr size_t UNIT_FLOATS = (size_t)NCHUNK * BM * PCW;
  // staging buffers per epilogue group: three when the rank table leaves room, else two
  __host__ __device__ static constexpr int out_bufs(int r4) { return BN_ >= 160 ? (r4 <= 2 ? 3 : 2) : 2; }
  // s*Bup + bias tables (fp32), or -- rank <= 8, bf16 -- the [BN x 16] bf16 operand of the fold MMA (BN * 32 bytes)
  __host__ __device__ static constexpr int epi_floats(int r4) {
    return (BN * 4 * (r4 > 0 ? r4 : 1) + BN) > BN * 8 ? (BN * 4 * (r4 > 0 ? r4 : 1) + BN) : BN * 8;
  }
  __host__ __device__ static constexpr size_t smem_bytes(int r4) {
    return 1024 /*align slack*/ + (size_t)NSTAGES * STAGE_BYTES + (size_t)2 * out_bufs(r4) * OUT_BYTES +
           (size_t)epi_floats(r4) * 4 + 256;
  }
};

// Division by a launch-constant through a host-computed multiplier (n < 2^31): the producer's tile -> coordinate
// chain is on the critical path of every launch and eight hardware divisions cost ~1000 cycles there.
struct FastDiv {
  uint32_t d, mul, shr;
  static FastDiv make(uint32_t d) {
    FastDiv f; f.d = d; f.mul = 0; f.shr = 0;
    if (d <= 1) { f.d = 1; return f; }
    uint32_t l = 0;
    while ((1u << l) < d) ++l;
    f.mul = (uint32_t)(((1ull << (31 + l)) + d - 1) / d);
    f.shr = l - 1;
    return f;
  }
  __device__ __forceinline__ uint32_t div(uint32_t n) const { return d == 1 ? n : (__umulhi(n, mul) >> shr); }
  __device__ __forceinline__ void divmod(uint32_t n, uint32_t& q, uint32_t& r) const { q = div(n); r = n - q * d; }
};

template <typename T> struct GemmTraits;
template <> struct GemmTraits<__nv_bfloat16> {
  static constexpr int kBK = 64;       // elements per 128-byte row
  static constexpr uint32_t kFmt = 1;  // bf16
  static constexpr int kExtRowBytes = EXT_ELEMS * 2;
  static constexpr bool kSplit3 = false;
  __device__ static inline void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    ptx::mma_f16(d, a, b, idesc, acc);
  }
};
template <> struct GemmTraits<__half> {
  static constexpr int kBK = 64;
  static constexpr uint32_t kFmt = 0;  // f16
  static constexpr int kExtRowBytes = EXT_ELEMS * 2;
  static constexpr bool kSplit3 = false;
  __device__ static inline void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    ptx::mma_f16(d, a, b, idesc, acc);
  }
};
template <> struct GemmTraits<float> {
  static constexpr int kBK = 32;
  static constexpr uint32_t kFmt = 2;  // tf32
  static constexpr int kExtRowBytes = EXT_ELEMS * 4;
  // fp32 storage: error-compensated 3xTF32.  Every landed stage is split in shared memory into
  // hi = x with the 13 low mantissa bits cleared (exact in tf32) and lo = x - hi (the stage slot behind it),
  // and each K step issues hi.hi + lo.hi + hi.lo into the same accumulator (~2^-21 rel).  Stages are used in pairs.
  static constexpr bool kSplit3 = true;
  __device__ static inline void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    ptx::mma_tf32(d, a, b, idesc, acc);
  }
};

struct GemmParams {
  int TW, TH, TB, rows;   // tile box in OUTPUT pixels; rows = TW*TH*TB <= 128
  int W, H, B;            // OUTPUT extent
  int stride;             // input pixel step per output pixel (1 or 2)
  int mt_w, mt_h, mt_b;   // number of pixel tiles per dimension
  int n_tiles, N;
  int taps, kw, pad, cblocks;
  int rank;
  int bup_ld, bup_jd;     // Bup[n, j] sits at n * bup_ld + j * bup_jd elements: (rank, 1) for a lora_up weight [N, r];
                          // (1, N) reads a lora_down weight [r, N] as its own transpose (fused 1x1 backward)
  float s;
  int lora;               // accumulate h and apply it in the epilogue
  int ext;                // extended-K operand pair present (input-gradient mode)
  int total_tiles;
  int m_tiles;            // pixel tiles (total_tiles / n_tiles)
  int pair_units;         // CTA-pair mode: work units of a pair = n_tiles * ceil(m_tiles / 2) * ksplit
  int per_tap;            // K iterations per tap = cblocks + ext
  int kiters;             // taps * per_tap
  int ksplit;             // K partitions per tile (1 = no split)
  int total_units;        // total_tiles * ksplit
  float* partial;         // split-K partial tiles, [unit][NCHUNK][128][PCW] fp32
  int accumulate;         // y += result (y holds the residual branch) instead of y = result
  int geglu;              // feed-forward gate fused into the store: the tile's columns are [a(40) | g(40)] x 2 (weight rows
                          // permuted by the caller) and the tile writes 80 columns a * gelu(g) of y [.., N / 2]
  int group_n;            // grouped adapters: output columns per group (0 = one adapter); group g uses h columns [g*rank, (g+1)*rank)
  FastDiv fd_n_tiles, fd_mt_w, fd_mt_h, fd_ksplit, fd_per_tap, fd_kw;
  int w_prefetch;         // K blocks the weight stream is requested into L2 ahead of the shared-memory pipeline
  long long* trace;       // debug: 16 clock64() stamps per CTA (NULL = off)
  int dbg_skip;           // debug (timing experiments only, results are garbage): bit 0 = no A loads, bit 1 = no W loads
  void* y;
  const void* bias;
  const void* Bup;
  float* h_out;
};

struct TileCoord { int w0, h0, b0, n0, n_blk; };

template <int BN>
__device__ __forceinline__ TileCoord tile_coord(const GemmParams& p, int tile) {
  TileCoord t;
  uint32_t mi, nb, wt, ht, bt;
  p.fd_n_tiles.divmod((uint32_t)tile, mi, nb);
  p.fd_mt_w.divmod(mi, mi, wt);
  p.fd_mt_h.divmod(mi, bt, ht);
  t.n_blk = (int)nb;
  t.w0 = (int)wt * p.TW; t.h0 = (int)ht * p.TH; t.b0 = (int)bt * p.TB; t.n0 = t.n_blk * BN;
  return t;
}

// K-iteration range [it0, it1) of split `split` of a tile
__device__ __forceinline__ void split_range(const GemmParams& p, uint32_t split, int& it0, int& it1) {
  it0 = (int)p.fd_ksplit.div(split * (uint32_t)p.kiters);
  it1 = (int)p.fd_ksplit.div((split + 1) * (uint32_t)p.kiters);
}

// N consecutive fp32 accumulator columns of this warp's 32 lanes into registers, NO wait (the caller issues every
// load of a tile and then one tcgen05.wait::ld)
template <int N> __device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t* v);
template <> __device__ __forceinline__ void tmem_ld_cols<40>(uint32_t taddr, uint32_t* v) {
  ptx::tmem_ld_x32(taddr, v);
  ptx::tmem_ld_x8(taddr + 32, v + 32);
}
template <> __device__ __forceinline__ void tmem_ld_cols<20>(uint32_t taddr, uint32_t* v) {
  ptx::tmem_ld_x16(taddr, v);
  ptx::tmem_ld_x4(taddr + 16, v + 16);
}

// TWO = CTA-pair mode (cluster of 2, cta_group::2): the pair owns two M-adjacent tiles of one N tile; each CTA loads its
// own 128 activation rows and HALF of the weight rows (80 W rows + 8 lora_down rows), the leader issues M = 256 MMAs over
// both shared memories, each CTA's 128 accumulator rows land in its own TMEM.  27 KB instead of 38 KB of operands per
// SM and K block: the launches with hundreds of tiles (SD-XL) are bound by L2 -> SM bandwidth, not by the tensor pipe.
// Accumulator columns in this mode: [0,80) W rows of CTA 0 | [80,88) h 0..7 | [88,168) W rows of CTA 1 | [168,176) h 8..15
// (adapter off: [0,80) | [80,160)).
template <typename T, int R4, int BN, bool TWO>
__global__ void __launch_bounds__(kGemmThreads, 1)
lora_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmL, const __grid_constant__ CUtensorMap tmXe,
                 const __grid_constant__ CUtensorMap tmWe, const __grid_constant__ CUtensorMap tmY,
                 const __grid_constant__ CUtensorMap tmP, const GemmParams p) {
  using Tr = GemmTraits<T>;
  using Cfg = TileCfg<BN>;
  constexpr int BK = Tr::kBK;
  constexpr int KSTEPS = ROW_BYTES / 32;   // one tcgen05.mma consumes 32 bytes of K per row
  constexpr int EXT_KSTEPS = Tr::kExtRowBytes / 32;
  constexpr int NSTAGES = Cfg::NSTAGES;
  constexpr int STAGE_BYTES = Cfg::STAGE_BYTES;
  constexpr int B_BYTES = Cfg::B_BYTES;
  constexpr int HALF_N = Cfg::HALF_N;
  constexpr int ACC_COLS = Cfg::ACC_COLS;
  constexpr int NCHUNK = Cfg::NCHUNK;
  constexpr int kStages = Tr::kSplit3 ? NSTAGES / 2 : NSTAGES;            // fp32: hi / lo planes take two slots
  constexpr int kStageStride = Tr::kSplit3 ? 2 * STAGE_BYTES : STAGE_BYTES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int NB = Cfg::out_bufs(R4);                    // staging buffers per epilogue group
  uint8_t* s_out = smem + (size_t)NSTAGES * STAGE_BYTES;   // 2 x NB x 10 KB staging of the TMA-store epilogue
  float* s_epi = reinterpret_cast<float*>(s_out + (size_t)2 * NB * OUT_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_epi + Cfg::epi_floats(R4));
  uint64_t* full_bar = bars;                      // [NSTAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + NSTAGES;           // [NSTAGES]  MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * NSTAGES;       // [ACC_STAGES] MMA -> epilogue
  uint64_t* tempty_bar = tfull_bar + ACC_STAGES;  // [ACC_STAGES] epilogue -> MMA
  uint64_t* xf_bar = tempty_bar + ACC_STAGES;     // [NSTAGES] hi/lo split done (fp32 only) -> MMA
  uint64_t* fold_bar = xf_bar + NSTAGES;          // fold MMA (bias + rank-r update on the tensor core) -> epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(fold_bar + 1);

  // warp index through a shuffle so that the role branches below are provably warp-uniform
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
#define SLIDER_TRACE(slot) do { if (p.trace != nullptr) p.trace[blockIdx.x * 16 + (slot)] = clock64(); } while (0)
  pdl_trigger();
  if (threadIdx.x == 0) SLIDER_TRACE(0);
  const uint32_t smem_base = ptx::smem_u32(smem);
  const uint32_t crank = TWO ? ptx::cluster_ctarank() : 0u;                 // rank in the CTA pair (0 = leader)
  const int first_unit = TWO ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;    // persistent loop over units / pair units
  const int unit_step = TWO ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int unit_end = TWO ? p.pair_units : p.total_units;
  constexpr int W_ROWS = TWO ? BN / 2 : BN;                                 // weight rows this CTA stages per K block
  constexpr int L_OFF = A_BYTES + W_ROWS * ROW_BYTES;                       // lora_down rows sit right behind them
  // unit -> (tile of THIS CTA, K split, index of the (tile, split) unit in the split-K workspace, phantom tile?)
  auto unit_map = [&](int u, uint32_t& tile, uint32_t& split, int& real_unit, bool& phantom) {
    tile = (uint32_t)u; split = 0; phantom = false;
    if (p.ksplit > 1) p.fd_ksplit.divmod((uint32_t)u, tile, split);
    if (TWO) {
      uint32_t mp, nb;
      p.fd_n_tiles.divmod(tile, mp, nb);
      const uint32_t mi = 2 * mp + crank;
      phantom = (int)mi >= p.m_tiles;          // odd tile count: the pair's second CTA computes zeros, stores nothing
      tile = mi * (uint32_t)p.n_tiles + nb;
    }
    real_unit = (int)tile * p.ksplit + (int)split;
  };
  const uint32_t main_tx = TWO ? 2u * (uint32_t)(p.rows * ROW_BYTES + W_ROWS * ROW_BYTES + (p.lora ? L_BYTES / 2 : 0))
                               : (uint32_t)(p.rows * ROW_BYTES + B_BYTES + (p.lora ? L_BYTES : 0));
  const uint32_t ext_tx = (uint32_t)((p.rows + BN) * Tr::kExtRowBytes);

  // ---- TMA producer state (one elected thread).
  int pu = first_unit, pit = 0, pit1 = 0, ptap = 0, pj = 0;
  uint32_t pstage = 0, pphase = 0;
  TileCoord ptc = {0, 0, 0, 0, 0};
  bool phave = false, pfirst_m = false;
  auto prod_begin = [&]() {
    phave = pu < unit_end;
    if (!phave) return;
    uint32_t tile, split;
    int ru;
    bool ph;
    unit_map(pu, tile, split, ru, ph);
    ptc = tile_coord<BN>(p, (int)tile);
    pfirst_m = TWO ? (int)tile < 2 * p.n_tiles : (ptc.w0 == 0 && ptc.h0 == 0 && ptc.b0 == 0);
    pit = 0; pit1 = p.kiters; ptap = 0; pj = 0;
    if (p.ksplit > 1) {
      split_range(p, split, pit, pit1);
      ptap = (int)p.fd_per_tap.div((uint32_t)pit);
      pj = pit - ptap * p.per_tap;
    }
  };
  // Cold weights (in the real forward every layer's weights are cold: 1.7 GB stream through a 126 MB L2): the CTAs
  // that share an N tile ask for the SAME weight rows, so only n_tiles x stages x 20 KB of distinct HBM requests are
  // ever in flight and the first stage lands ~2 us late.  The CTA that owns the first pixel tile of an N tile
  // therefore requests the weight tiles of the next `w_prefetch` K blocks into L2 (TMA tile prefetch, in consumption
  // order), the first burst before the dependency wait -- frozen weights are never written by a preceding kernel.
  auto prod_prefetch_w = [&](int it) {
    const int tap = (int)p.fd_per_tap.div((uint32_t)it);
    const int j = it - tap * p.per_tap;
    if (j < p.cblocks) ptx::tma_prefetch_l2_3d(&tmW, j * BK, tap, ptc.n0 + (int)crank * W_ROWS);
  };
  auto prod_owner = [&]() { return p.w_prefetch > 0 && pfirst_m; };
  auto prod_issue = [&]() {
    const int ty = (int)p.fd_kw.div((uint32_t)ptap);
    const int dy = ty - p.pad, dx = ptap - ty * p.kw - p.pad;
    ptx::mbar_wait(ptx::smem_u32(empty_bar + pstage), pphase ^ 1);
    const uint32_t sa = smem_base + pstage * kStageStride;
    const uint32_t fb = ptx::smem_u32(full_bar + pstage);
    if (TWO) {
      // both CTAs' loads complete on the LEADER's barrier, which expects the bytes of both
      const uint32_t fbl = ptx::mapa(fb, 0u);
      if (crank == 0) ptx::mbar_arrive_expect_tx(fb, main_tx);
      ptx::tma_load_4d_2cta(sa, &tmA, fbl, pj * BK, ptc.w0 * p.stride + dx, ptc.h0 * p.stride + dy, ptc.b0);
      ptx::tma_load_3d_2cta(sa + A_BYTES, &tmW, fbl, pj * BK, ptap, ptc.n0 + (int)crank * W_ROWS);
      if (p.lora) ptx::tma_load_3d_2cta(sa + L_OFF, &tmL, fbl, pj * BK, ptap, (int)crank * (HCOLS / 2));
    } else if (pj < p.cblocks) {
      if (p.dbg_skip) {
        const uint32_t tx = ((p.dbg_skip & 1) ? 0u : (uint32_t)(p.rows * ROW_BYTES)) + ((p.dbg_skip & 2) ? 0u : (uint32_t)B_BYTES) +
                            (p.lora ? (uint32_t)L_BYTES : 0u);
        ptx::mbar_arrive_expect_tx(fb, tx);
        if (!(p.dbg_skip & 1)) ptx::tma_load_4d(sa, &tmA, fb, pj * BK, ptc.w0 * p.stride + dx, ptc.h0 * p.stride + dy, ptc.b0);
        if (!(p.dbg_skip & 2)) ptx::tma_load_3d(sa + A_BYTES, &tmW, fb, pj * BK, ptap, ptc.n0);
        if (p.lora) ptx::tma_load_3d(sa + A_BYTES + B_BYTES, &tmL, fb, pj * BK, ptap, 0);
        if (tx == 0) ptx::mbar_arrive(fb);
      } else {
      ptx::mbar_arrive_expect_tx(fb, main_tx);
      // stride 2: the activation map samples every second pixel (element strides {1, 2, 2, 1}), so the box that
      // starts at input pixel (2*w0 + dx, 2*h0 + dy) lands the TW x TH output pixels' operands densely
      ptx::tma_load_4d(sa, &tmA, fb, pj * BK, ptc.w0 * p.stride + dx, ptc.h0 * p.stride + dy, ptc.b0);
      ptx::tma_load_3d(sa + A_BYTES, &tmW, fb, pj * BK, ptap, ptc.n0);
      if (p.lora) ptx::tma_load_3d(sa + A_BYTES + B_BYTES, &tmL, fb, pj * BK, ptap, 0);
      }
    } else {
      ptx::mbar_arrive_expect_tx(fb, ext_tx);
      ptx::tma_load_4d(sa, &tmXe, fb, 0, ptc.w0 + dx, ptc.h0 + dy, ptc.b0);
      ptx::tma_load_3d(sa + A_BYTES, &tmWe, fb, 0, ptap, ptc.n0);
    }
    if (prod_owner() && pit + p.w_prefetch < pit1) prod_prefetch_w(pit + p.w_prefetch);
    if (++pstage == (uint32_t)kStages) { pstage = 0; pphase ^= 1; }
    if (++pj == p.per_tap) { pj = 0; ++ptap; }
    if (++pit == pit1) {
      pu += unit_step;
      prod_begin();
      if (phave && prod_owner())
        for (int d = 0; d < p.w_prefetch && pit + d < pit1; ++d) prod_prefetch_w(pit + d);
    }
  };

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmW);
    if (p.lora) ptx::prefetch_tmap(&tmL);
    if (p.ext) { ptx::prefetch_tmap(&tmXe); ptx::prefetch_tmap(&tmWe); }
    if (p.ksplit > 1) ptx::prefetch_tmap(&tmP); else ptx::prefetch_tmap(&tmY);
    for (int i = 0; i < NSTAGES; ++i) {
      ptx::mbar_init(ptx::smem_u32(full_bar + i), 1);
      ptx::mbar_init(ptx::smem_u32(empty_bar + i), 1);
      ptx::mbar_init(ptx::smem_u32(xf_bar + i), 8);      // one arrive per epilogue warp
    }
    for (int i = 0; i < ACC_STAGES; ++i) {
      ptx::mbar_init(ptx::smem_u32(tfull_bar + i), 1);
      ptx::mbar_init(ptx::smem_u32(tempty_bar + i), TWO ? 16 : 8);  // one arrive per epilogue warp (of both CTAs)
    }
    ptx::mbar_init(ptx::smem_u32(fold_bar), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    if (TWO) { ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), Cfg::TMEM_COLS); ptx::tmem_relinquish_2cta(); }
    else { ptx::tmem_alloc(ptx::smem_u32(tmem_slot), Cfg::TMEM_COLS); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  // CTA-wide start barrier.  Warp 0 only ARRIVES (its lane 0 initialised the mbarriers above and is the one
  // thread that never needs the TMEM base), so the producer starts fetching operands while the TMEM allocation
  // and the other warps' start-up are still in flight.
  if (TWO) {
    // pair mode: nobody may signal the peer's barriers or write its shared memory before the peer has initialised
    __syncthreads();
    ptx::cluster_arrive();
    ptx::cluster_wait();
  } else if (warp == 0) {
    __syncwarp();
    asm volatile("bar.arrive 1, %0;" ::"n"(kGemmThreads) : "memory");
  } else {
    asm volatile("bar.sync 1, %0;" ::"n"(kGemmThreads) : "memory");
  }
  ptx::tc_fence_after();
  // PDL: nothing of ours reads what the previous kernel wrote before that kernel has completed and flushed (the
  // producer warp waits after its first weight-prefetch burst below)
  if (warp != 0) pdl_wait();
  const uint32_t tmem_base = warp == 0 ? 0u : *tmem_slot;
  if (threadIdx.x == 32) SLIDER_TRACE(1);

  // The two issuing roles run in ONE ELECTED lane (elect.sync, not `lane == 0`): ptxas then knows a single thread
  // is active and feeds UTMALDG / UTCHMMA from uniform registers directly instead of wrapping every issue in an
  // ELECT + R2UR.BROADCAST waterfall loop (which made the producer, ~500 cycles per K block, the main-loop bound).
  if (warp == 0 && ptx::elect_one()) {
    // ===================== TMA producer =====================
    prod_begin();
    SLIDER_TRACE(8);
    if (phave && prod_owner())
      for (int d = 0; d < p.w_prefetch && pit + d < pit1; ++d) prod_prefetch_w(pit + d);
    pdl_wait();
    bool first = true;
    while (phave) {
      prod_issue();
      if (first) { SLIDER_TRACE(2); first = false; }
    }
  } else if (warp == 1 && (!TWO || crank == 0) && ptx::elect_one()) {
    // ===================== MMA issuer (pair mode: the leader only) =====================
    // With the adapter on, W tile (BN rows) and lora_down tile (HCOLS rows) are adjacent in shared
    // memory and their accumulators adjacent in TMEM: ONE tcgen05.mma of N = BN + HCOLS covers both.
    constexpr uint32_t idesc_main = ptx::make_idesc(Tr::kFmt, TWO ? 2 * BM : BM, BN);
    constexpr uint32_t idesc_lora = ptx::make_idesc(Tr::kFmt, TWO ? 2 * BM : BM, BN + HCOLS);
    const uint32_t idesc = p.lora ? idesc_lora : idesc_main;
    uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
    for (int unit = first_unit; unit < unit_end; unit += unit_step) {
      int it0 = 0, it1 = p.kiters, j = 0;
      if (p.ksplit > 1) {
        uint32_t tile, split;
        p.fd_ksplit.divmod((uint32_t)unit, tile, split);
        split_range(p, split, it0, it1);
        j = it0 - (int)p.fd_per_tap.div((uint32_t)it0) * p.per_tap;
      }
      ptx::mbar_wait(ptx::smem_u32(tempty_bar + as), aphase ^ 1);
      ptx::tc_fence_after();
      const uint32_t d_main = tmem_base + as * ACC_COLS;
      uint32_t acc = 0;
      for (int it = it0; it < it1; ++it) {
        ptx::mbar_wait(ptx::smem_u32((Tr::kSplit3 ? xf_bar : full_bar) + stage), phase);
        ptx::tc_fence_after();
        if (unit == first_unit && it == it0) SLIDER_TRACE(3);
        const uint32_t sa = smem_base + stage * kStageStride;
        const bool main_blk = j < p.cblocks;
        const uint32_t rb = main_blk ? (uint32_t)ROW_BYTES : (uint32_t)Tr::kExtRowBytes;
        const uint32_t id = main_blk ? idesc : idesc_main;
        const int ksteps = main_blk ? KSTEPS : EXT_KSTEPS;
        const uint64_t da = ptx::smem_desc_kmajor(sa, rb);
        const uint64_t db = ptx::smem_desc_kmajor(sa + A_BYTES, rb);
        if (Tr::kSplit3) {
          const uint64_t da_lo = ptx::smem_desc_kmajor(sa + STAGE_BYTES, rb);
          const uint64_t db_lo = ptx::smem_desc_kmajor(sa + STAGE_BYTES + A_BYTES, rb);
          for (int k = 0; k < ksteps; ++k) {
            Tr::mma(d_main, da_lo + 2 * k, db + 2 * k, id, acc | (uint32_t)k);
            Tr::mma(d_main, da + 2 * k, db_lo + 2 * k, id, 1u);
            Tr::mma(d_main, da + 2 * k, db + 2 * k, id, 1u);
          }
        } else if (TWO) {
#pragma unroll
          for (int k = 0; k < KSTEPS; ++k) ptx::mma_f16_2cta(d_main, da + 2 * k, db + 2 * k, id, acc | (uint32_t)k);
        } else if (main_blk) {
#pragma unroll
          for (int k = 0; k < KSTEPS; ++k) Tr::mma(d_main, da + 2 * k, db + 2 * k, id, acc | (uint32_t)k);
        } else {
#pragma unroll
          for (int k = 0; k < EXT_KSTEPS; ++k) Tr::mma(d_main, da + 2 * k, db + 2 * k, id, acc | (uint32_t)k);
        }
        acc = 1;
        if (TWO) ptx::mma_commit_2cta(ptx::smem_u32(empty_bar + stage), 3);     // frees the stage in BOTH CTAs
        else ptx::mma_commit(ptx::smem_u32(empty_bar + stage));
        if (++stage == (uint32_t)kStages) { stage = 0; phase ^= 1; }
        if (++j == p.per_tap) j = 0;
      }
      if (TWO) ptx::mma_commit_2cta(ptx::smem_u32(tfull_bar + as), 3);
      else ptx::mma_commit(ptx::smem_u32(tfull_bar + as));
      if (unit == first_unit) SLIDER_TRACE(4);
      if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
    }
  } else if (warp >= 2) {
    // ===================== epilogue =====================
    // Eight warps = two groups of four; a group covers all 128 accumulator rows (one TMEM lane quarter per warp)
    // and owns HALF_N of the tile's columns.  Per tile a thread reads its row of the group's columns (and the h
    // columns) with back-to-back tcgen05.ld and ONE wait, hands the accumulator stage back to the MMA warp, and then,
    // per 80-byte chunk (40 bf16 / 20 fp32 columns): (+bias + s*h.Bup^T, cast) -> shared staging -> one TMA store
    // (box {chunk, TW, TH, TB}: clips out-of-range pixels/channels).  NB staging buffers and one named barrier per
    // group: before the barrier of chunk c the group's issuing thread waits until at most NB-2 stores still read
    // shared memory, so the buffer the group fills next (chunk c+1, last used by store c+1-NB) is free once its
    // threads pass that barrier.
    const int q = warp & 3;               // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;        // tile row == TMEM lane
    const int et = threadIdx.x - 64;      // 0..255 among epilogue threads
    const int grp = et >> 7;              // column half owned by this thread's group
    const int gt = et & 127;
    constexpr int CW = OUT_ROW_BYTES / (int)sizeof(T);    // columns per store chunk
    constexpr int CPG = HALF_N / CW;                      // store chunks per group
    constexpr int PW = 16 / (int)sizeof(T);               // columns per 16-byte piece
    constexpr int PCPG = HALF_N / PCW;                    // split-K partial chunks per group
    float* s_bias = s_epi + BN * 4 * (R4 > 0 ? R4 : 1);
    const uint32_t out_base = ptx::smem_u32(s_out) + (uint32_t)(grp * NB) * OUT_BYTES;
    const uint32_t epi_s = ptx::smem_u32(s_epi), bias_s = ptx::smem_u32(s_bias);
    uint32_t as = 0, aphase = 0, ob = 0, xs = 0, xphase = 0, fphase = 0;
    const T* __restrict__ bias = reinterpret_cast<const T*>(p.bias);
    const T* __restrict__ Bup = reinterpret_cast<const T*>(p.Bup);
#define SLIDER_GROUP_BAR() asm volatile("bar.sync %0, 128;" ::"r"(3 + grp) : "memory")
    // pair mode: accumulator column of this group's first output column / of h column j (see the layout above)
    const uint32_t grp_col = TWO ? (uint32_t)grp * (uint32_t)(HALF_N + (p.lora ? HCOLS / 2 : 0)) : (uint32_t)(grp * HALF_N);
    auto tempty_arrive = [&](uint32_t stage) {      // one arrive per epilogue warp, on the LEADER's barrier in pair mode
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        const uint32_t bar = ptx::smem_u32(tempty_bar + stage);
        if (TWO && crank != 0) ptx::mbar_arrive_cluster(ptx::mapa(bar, 0u));
        else ptx::mbar_arrive(bar);
      }
    };
    for (int unit = first_unit; unit < unit_end; unit += unit_step) {
      uint32_t tile, split;
      int real_unit;
      bool phantom;
      unit_map(unit, tile, split, real_unit, phantom);
      if (Tr::kSplit3) {
        // fp32: split every stage of this unit into hi / lo before the MMA warp may consume it
        int n_it = p.kiters;
        if (p.ksplit > 1) {
          int a0, a1;
          split_range(p, split, a0, a1);
          n_it = a1 - a0;
        }
        for (int it = 0; it < n_it; ++it) {
          ptx::mbar_wait(ptx::smem_u32(full_bar + xs), xphase);
          float4* hi = reinterpret_cast<float4*>(smem + (size_t)xs * kStageStride);
          float4* lo = reinterpret_cast<float4*>(smem + (size_t)xs * kStageStride + STAGE_BYTES);
          for (int i = et; i < STAGE_BYTES / 16; i += kEpiThreads) {
            const float4 v = hi[i];
            float4 h4, l4;
            h4.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); l4.x = v.x - h4.x;
            h4.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); l4.y = v.y - h4.y;
            h4.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); l4.z = v.z - h4.z;
            h4.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); l4.w = v.w - h4.w;
            hi[i] = h4;
            lo[i] = l4;
          }
          ptx::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(xf_bar + xs));
          if (++xs == (uint32_t)kStages) { xs = 0; xphase ^= 1; }
        }
      }
      if (p.ksplit > 1) {
        // split-K: park the fp32 partial tile (main + h columns) in the L2-resident workspace,
        // [unit][chunk][row][PCW]; splitk_finalize_kernel reduces over the splits in a fixed order.
        // one warp polls the accumulator-ready mbarrier, the others sleep on the hardware barrier: eight warps
        // spinning on try_wait steal shared-memory cycles from the operand pipeline for the whole main loop
        if (et < 32) ptx::mbar_wait(ptx::smem_u32(tfull_bar + as), aphase);
        asm volatile("bar.sync 2, 256;" ::: "memory");
        ptx::tc_fence_after();
        if (et == 0 && unit == first_unit) SLIDER_TRACE(5);
        const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + as * ACC_COLS;
        const bool park_h = grp == 0 && R4 > 0;                    // group 0 also parks the h columns
        uint32_t v[PCPG][PCW];
        uint32_t hr[16];
#pragma unroll
        for (int ci = 0; ci < PCPG; ++ci) tmem_ld_cols<PCW>(t_acc + grp_col + ci * PCW, v[ci]);
        if (park_h) {
          if (TWO) { ptx::tmem_ld_x8(t_acc + HALF_N, hr); ptx::tmem_ld_x8(t_acc + BN + HCOLS / 2, hr + 8); }
          else ptx::tmem_ld_x16(t_acc + BN, hr);
        }
        ptx::tmem_ld_wait();
        // the accumulator stage is free again as soon as it sits in registers
        tempty_arrive(as);
        if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
        if (TWO && phantom) continue;                               // uniform over the CTA: nothing to park
#pragma unroll
        for (int ci = 0; ci <= PCPG; ++ci) {
          if (ci == PCPG && !park_h) break;
          const int chunk = ci < PCPG ? grp * PCPG + ci : NCHUNK - 1;
          const uint32_t dst = out_base + ob * OUT_BYTES + row * OUT_ROW_BYTES;
          if (ci < PCPG) {
#pragma unroll
            for (int k = 0; k < OUT_ROW_BYTES / 16; ++k)
              ptx::st_shared_v4(dst + k * 16, v[ci % PCPG][4 * k], v[ci % PCPG][4 * k + 1], v[ci % PCPG][4 * k + 2],
                                v[ci % PCPG][4 * k + 3]);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) ptx::st_shared_v4(dst + k * 16, hr[4 * k], hr[4 * k + 1], hr[4 * k + 2], hr[4 * k + 3]);
            ptx::st_shared_v4(dst + 64, 0u, 0u, 0u, 0u);
          }
          ptx::fence_proxy_async_smem();
          if (gt == 0) ptx::bulk_wait_read<NB - 2>();   // buffer of chunk c+1 (last read by store c+1-NB) is free past the barrier
          SLIDER_GROUP_BAR();
          if (gt == 0) {
            ptx::tma_store_2d(&tmP, out_base + ob * OUT_BYTES, 0, (real_unit * NCHUNK + chunk) * BM);
            ptx::bulk_commit();
          }
          if (++ob == (uint32_t)NB) ob = 0;
        }
        if (et == 0 && unit == first_unit) SLIDER_TRACE(6);
        continue;
      }
      const TileCoord tc = tile_coord<BN>(p, (int)tile);
      // FOLD (bf16, rank <= 8, single-CTA form): bias and the rank-r update go through ONE more tcgen05.mma,
      //   D[128 x BN] += [h_0 .. h_{r-1}, 1, 0 ..] (128 x 16, bf16, written to TMEM by the epilogue)
      //                  . [s*B_n0 .. s*B_n(r-1), bias_n, 0 ..]^T (BN x 16, bf16, staged in shared memory)
      // The per-thread form below reads one broadcast LDS.128 of s*B per output: 8 warps x 50 LDS.128 x 4 cycles of
      // shared-memory return bandwidth = 1600 cycles per 40-column chunk, which WAS the epilogue (traced).
      constexpr bool FOLD = !TWO && sizeof(T) == 2 && (R4 == 1 || R4 == 2);
      const bool fold = FOLD && p.lora && !p.geglu;
      if (fold) {
        asm volatile("bar.sync 2, 256;" ::: "memory");      // the previous tile's fold MMA has been waited for by everyone
        for (int i = et; i < BN; i += kEpiThreads) {
          const int n = tc.n0 + i;
          uint32_t wv[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            float e[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int j = 2 * k + u;
              float f = 0.f;
              if (n < p.N) {
                if (j < p.rank) f = p.s * Elem<T>::to_f(Bup[(size_t)n * p.bup_ld + (size_t)j * p.bup_jd]);
                else if (j == p.rank && bias != nullptr) f = Elem<T>::to_f(bias[n]);
              }
              e[u] = f;
            }
            wv[k] = Elem<T>::pack2(e[0], e[1]);
          }
          // K-major rows of 32 bytes, 32-byte swizzle (16-byte chunk index ^= bit 2 of the row): the layout the
          // extended-K operand is landed in by TMA
          const uint32_t sw = (uint32_t)((i >> 2) & 1);
          const uint32_t rowa = epi_s + (uint32_t)i * 32u;
          ptx::st_shared_v4(rowa + ((0u ^ sw) << 4), wv[0], wv[1], wv[2], wv[3]);
          ptx::st_shared_v4(rowa + ((1u ^ sw) << 4), wv[4], wv[5], wv[6], wv[7]);
        }
        ptx::fence_proxy_async_smem();
        if (et < 32) ptx::mbar_wait(ptx::smem_u32(tfull_bar + as), aphase);
        asm volatile("bar.sync 2, 256;" ::: "memory");
        ptx::tc_fence_after();
        if (et == 0 && unit == first_unit) SLIDER_TRACE(5);
        const uint32_t t_main = tmem_base + ((uint32_t)(q * 32) << 16) + as * ACC_COLS;
        const uint32_t t_fold = tmem_base + ((uint32_t)(q * 32) << 16) + ACC_STAGES * ACC_COLS + as * 8;
        if (grp == 0) {
          uint32_t hr[16];
          const int hoff = p.group_n > 0 ? (tc.n0 / p.group_n) * p.rank : 0;
          ptx::tmem_ld_x16(t_main + BN + hoff, hr);
          ptx::tmem_ld_wait();
          if (p.h_out != nullptr && tc.n_blk == 0) {
            const int lw = row % p.TW, lh = (row / p.TW) % p.TH, lb = row / (p.TW * p.TH);
            const bool row_ok = row < p.rows && tc.w0 + lw < p.W && tc.h0 + lh < p.H && tc.b0 + lb < p.B;
            const size_t pix = ((size_t)(tc.b0 + lb) * p.H + (tc.h0 + lh)) * p.W + (tc.w0 + lw);
            if (row_ok) {
#pragma unroll
              for (int j = 0; j < 4 * R4; ++j)
                if (j < p.rank) p.h_out[pix * p.rank + j] = __uint_as_float(hr[j]);
            }
          }
          uint32_t av[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            float e[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int j = 2 * k + u;
              e[u] = j < p.rank ? __uint_as_float(hr[j]) : (j == p.rank ? 1.f : 0.f);
            }
            av[k] = Elem<T>::pack2(e[0], e[1]);
          }
          ptx::tmem_st_x8(t_fold, av);
          ptx::tmem_st_wait();
        }
        ptx::tc_fence_before();
        asm volatile("bar.sync 2, 256;" ::: "memory");
        if (et == 0) {
          ptx::tc_fence_after();
          constexpr uint32_t idesc_fold = ptx::make_idesc(Tr::kFmt, BM, BN);
          ptx::mma_f16_ts(tmem_base + as * ACC_COLS, tmem_base + ACC_STAGES * ACC_COLS + as * 8,
                          ptx::smem_desc_kmajor(epi_s, 32), idesc_fold, 1u);
          ptx::mma_commit(ptx::smem_u32(fold_bar));
        }
        if (et < 32) ptx::mbar_wait(ptx::smem_u32(fold_bar), fphase);
        fphase ^= 1;
        asm volatile("bar.sync 2, 256;" ::: "memory");
        ptx::tc_fence_after();
        uint32_t v[CPG][CW];
#pragma unroll
        for (int ci = 0; ci < CPG; ++ci) tmem_ld_cols<CW>(t_main + grp_col + ci * CW, v[ci]);
        ptx::tmem_ld_wait();
        tempty_arrive(as);
        if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
        if (et == 0 && unit == first_unit) SLIDER_TRACE(9);
#pragma unroll
        for (int ci = 0; ci < CPG; ++ci) {
          const int col0 = grp * HALF_N + ci * CW;
          if (tc.n0 + col0 >= p.N) break;
          const uint32_t dst = out_base + ob * OUT_BYTES + row * OUT_ROW_BYTES;
#pragma unroll
          for (int pc = 0; pc < OUT_ROW_BYTES / 16; ++pc) {
            uint32_t w[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              w[k] = Elem<T>::pack2(__uint_as_float(v[ci][(pc * PW + 2 * k) % CW]),
                                    __uint_as_float(v[ci][(pc * PW + 2 * k + 1) % CW]));
            }
            ptx::st_shared_v4(dst + pc * 16, w[0], w[1], w[2], w[3]);
          }
          ptx::fence_proxy_async_smem();
          if (gt == 0) ptx::bulk_wait_read<NB - 2>();
          SLIDER_GROUP_BAR();
          if (gt == 0) {
            if (p.accumulate) ptx::tma_reduce_add_4d(&tmY, out_base + ob * OUT_BYTES, tc.n0 + col0, tc.w0, tc.h0, tc.b0);
            else ptx::tma_store_4d(&tmY, out_base + ob * OUT_BYTES, tc.n0 + col0, tc.w0, tc.h0, tc.b0);
            ptx::bulk_commit();
          }
          if (++ob == (uint32_t)NB) ob = 0;
        }
        if (et == 0 && unit == first_unit) SLIDER_TRACE(6);
        continue;
      }
      // stage s*Bup[n0.., :] and bias[n0..] for this tile (fp32) while the MMAs run
      asm volatile("bar.sync 2, 256;" ::: "memory");
      for (int i = et; i < BN; i += kEpiThreads) {
        const int n = tc.n0 + i;
        s_bias[i] = (bias != nullptr && n < p.N) ? Elem<T>::to_f(bias[n]) : 0.f;
        if (R4 > 0) {
#pragma unroll
          for (int j = 0; j < 4 * R4; ++j)
            s_epi[i * (4 * R4) + j] =
                (n < p.N && j < p.rank) ? p.s * Elem<T>::to_f(Bup[(size_t)n * p.bup_ld + (size_t)j * p.bup_jd]) : 0.f;
        }
      }
      // one warp polls the accumulator-ready mbarrier, the others sleep on the hardware barrier (which also
      // publishes the tables staged above): eight warps spinning on try_wait steal shared-memory cycles from the
      // operand pipeline for the whole main loop
      if (et < 32) ptx::mbar_wait(ptx::smem_u32(tfull_bar + as), aphase);
      asm volatile("bar.sync 2, 256;" ::: "memory");
      ptx::tc_fence_after();
      if (et == 0 && unit == first_unit) SLIDER_TRACE(5);
      const uint32_t t_main = tmem_base + ((uint32_t)(q * 32) << 16) + as * ACC_COLS;

      // every TMEM read of this tile, back to back, ONE wait; then the stage goes back to the MMA warp
      uint32_t v[CPG][CW];
      uint32_t hr[16];
#pragma unroll
      for (int ci = 0; ci < CPG; ++ci) tmem_ld_cols<CW>(t_main + grp_col + ci * CW, v[ci]);
      if (R4 > 0) {
        // grouped adapters: this tile's group owns `rank` of the 16 h columns, starting at its offset
        const int hoff = p.group_n > 0 ? (tc.n0 / p.group_n) * p.rank : 0;
        if (!TWO) ptx::tmem_ld_x16(t_main + BN + hoff, hr);
        else if (R4 > 2) { ptx::tmem_ld_x8(t_main + HALF_N, hr); ptx::tmem_ld_x8(t_main + BN + HCOLS / 2, hr + 8); }
        else ptx::tmem_ld_x8(t_main + (hoff < HCOLS / 2 ? HALF_N + hoff : BN + hoff), hr);   // host: a group's h never straddles
      }
      ptx::tmem_ld_wait();
      tempty_arrive(as);
      if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
      if (et == 0 && unit == first_unit) SLIDER_TRACE(9);

      float hv[R4 > 0 ? 4 * R4 : 1];
      if (R4 > 0) {
#pragma unroll
        for (int j = 0; j < 4 * R4; ++j) hv[j] = __uint_as_float(hr[j]);
        if (p.h_out != nullptr && tc.n_blk == 0 && grp == 1) {
          const int lw = row % p.TW, lh = (row / p.TW) % p.TH, lb = row / (p.TW * p.TH);
          const bool row_ok = row < p.rows && tc.w0 + lw < p.W && tc.h0 + lh < p.H && tc.b0 + lb < p.B;
          const size_t pix = ((size_t)(tc.b0 + lb) * p.H + (tc.h0 + lh)) * p.W + (tc.w0 + lw);
          if (row_ok) {
#pragma unroll
            for (int j = 0; j < 4 * R4; ++j)
              if (j < p.rank) p.h_out[pix * p.rank + j] = hv[j];
          }
        }
      }
      if (p.geglu) {
        // GEGLU (diffusers GEGLU.forward: hidden, gate = proj(x).chunk(2); hidden * gelu(gate)): this group's first
        // CPG/2 chunks are the values, the last CPG/2 the matching gates
        if (BN == 160) {
#pragma unroll
          for (int ci = 0; ci < CPG / 2; ++ci) {
            const int ca = grp * HALF_N + ci * CW, cg = ca + HALF_N / 2;
            const uint32_t dst = out_base + ob * OUT_BYTES + row * OUT_ROW_BYTES;
#pragma unroll
            for (int pc = 0; pc < OUT_ROW_BYTES / 16; ++pc) {
              float o[PW], ba[PW], bg[PW];
              // every shared-memory load of the piece first, then the arithmetic (a warp issues in order: a load
              // followed directly by its consumer exposes the full LDS latency per output)
#pragma unroll
              for (int k4 = 0; k4 < PW / 4; ++k4) {
                const float4 t0 = ptx::ld_shared_f4(bias_s + (ca + pc * PW + 4 * k4) * 4);
                const float4 t1 = ptx::ld_shared_f4(bias_s + (cg + pc * PW + 4 * k4) * 4);
                ba[4 * k4] = t0.x; ba[4 * k4 + 1] = t0.y; ba[4 * k4 + 2] = t0.z; ba[4 * k4 + 3] = t0.w;
                bg[4 * k4] = t1.x; bg[4 * k4 + 1] = t1.y; bg[4 * k4 + 2] = t1.z; bg[4 * k4 + 3] = t1.w;
              }
#pragma unroll
              for (int k = 0; k < PW; ++k) {
                const int i = pc * PW + k;
                const float av = __uint_as_float(v[ci][i]) + ba[k];
                const float gv = __uint_as_float(v[ci + CPG / 2][i]) + bg[k];
                o[k] = av * (sizeof(T) == 2 ? gelu_bf16_grade(gv) : 0.5f * gv * (1.f + erff(gv * 0.70710678118654752f)));
              }
              if (sizeof(T) == 2) {
                uint32_t w[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  w[k] = Elem<T>::pack2(o[(2 * k) % PW], o[(2 * k + 1) % PW]);
                }
                ptx::st_shared_v4(dst + pc * 16, w[0], w[1], w[2], w[3]);
              } else {
                ptx::st_shared_v4(dst + pc * 16, __float_as_uint(o[0]), __float_as_uint(o[1]), __float_as_uint(o[2 % PW]),
                                  __float_as_uint(o[3 % PW]));
              }
            }
            ptx::fence_proxy_async_smem();
            if (gt == 0) ptx::bulk_wait_read<NB - 2>();
            SLIDER_GROUP_BAR();
            if (gt == 0) {
              ptx::tma_store_4d(&tmY, out_base + ob * OUT_BYTES, tc.n_blk * (BN / 2) + grp * (HALF_N / 2) + ci * CW, tc.w0,
                                tc.h0, tc.b0);
              ptx::bulk_commit();
            }
            if (++ob == (uint32_t)NB) ob = 0;
          }
        }
        if (et == 0 && unit == first_unit) SLIDER_TRACE(6);
        continue;
      }
#pragma unroll
      for (int ci = 0; ci < CPG; ++ci) {
        const int col0 = grp * HALF_N + ci * CW;
        if (tc.n0 + col0 >= p.N) break;              // ragged last column tile (uniform over the group)
        const uint32_t dst = out_base + ob * OUT_BYTES + row * OUT_ROW_BYTES;
#pragma unroll
        for (int pc = 0; pc < OUT_ROW_BYTES / 16; ++pc) {
          float o[PW], bs[PW];
          // every shared-memory load of a batch first, then the arithmetic: a warp issues in order, so a load followed
          // directly by its consumer exposes the full LDS latency per output (42 cycles per output before this)
#pragma unroll
          for (int k4 = 0; k4 < PW / 4; ++k4) {       // explicit ld.shared: generic pointers off the re-aligned base compile to LD.E
            const float4 t = ptx::ld_shared_f4(bias_s + (col0 + pc * PW + 4 * k4) * 4);
            bs[4 * k4] = t.x; bs[4 * k4 + 1] = t.y; bs[4 * k4 + 2] = t.z; bs[4 * k4 + 3] = t.w;
          }
          constexpr int KB = R4 <= 1 ? PW : (R4 == 2 ? 4 : 2);      // outputs per batch: <= 8 float4 of s*Bup in flight
#pragma unroll
          for (int k0 = 0; k0 < PW; k0 += KB) {
            float4 bb[KB][R4 > 0 ? R4 : 1];
            if (R4 > 0) {
#pragma unroll
              for (int k = 0; k < KB; ++k)
#pragma unroll
                for (int j4 = 0; j4 < R4; ++j4)
                  bb[k][j4] = ptx::ld_shared_f4(epi_s + (col0 + pc * PW + k0 + k) * (16 * R4) + j4 * 16);
            }
#pragma unroll
            for (int k = 0; k < KB; ++k) {
              float a = __uint_as_float(v[ci][pc * PW + k0 + k]) + bs[k0 + k];
              if (R4 > 0) {
#pragma unroll
                for (int j4 = 0; j4 < R4; ++j4) {
                  a = fmaf(hv[4 * j4 + 0], bb[k][j4].x, a);
                  a = fmaf(hv[4 * j4 + 1], bb[k][j4].y, a);
                  a = fmaf(hv[4 * j4 + 2], bb[k][j4].z, a);
                  a = fmaf(hv[4 * j4 + 3], bb[k][j4].w, a);
                }
              }
              o[k0 + k] = a;
            }
          }
          if (sizeof(T) == 2) {
            uint32_t w[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              w[k] = Elem<T>::pack2(o[(2 * k) % PW], o[(2 * k + 1) % PW]);
            }
            ptx::st_shared_v4(dst + pc * 16, w[0], w[1], w[2], w[3]);
          } else {
            ptx::st_shared_v4(dst + pc * 16, __float_as_uint(o[0]), __float_as_uint(o[1]), __float_as_uint(o[2 % PW]),
                              __float_as_uint(o[3 % PW]));
          }
        }
        if (et == 0 && unit == first_unit && ci == 0) SLIDER_TRACE(10);
        ptx::fence_proxy_async_smem();
        if (et == 0 && unit == first_unit && ci == 0) SLIDER_TRACE(11);
        if (gt == 0) ptx::bulk_wait_read<NB - 2>();     // buffer of chunk c+1 (last read by store c+1-NB) is free past the barrier
        SLIDER_GROUP_BAR();
        if (et == 0 && unit == first_unit && ci == 0) SLIDER_TRACE(12);
        if (gt == 0) {
          if (p.accumulate) ptx::tma_reduce_add_4d(&tmY, out_base + ob * OUT_BYTES, tc.n0 + col0, tc.w0, tc.h0, tc.b0);
          else ptx::tma_store_4d(&tmY, out_base + ob * OUT_BYTES, tc.n0 + col0, tc.w0, tc.h0, tc.b0);
          ptx::bulk_commit();
        }
        if (++ob == (uint32_t)NB) ob = 0;
      }
      if (et == 0 && unit == first_unit) SLIDER_TRACE(6);
    }
    // the staging buffers must outlive the stores' READS; the writes themselves complete with the grid (the next
    // kernel's dependency wait covers them), so the CTA does not sit out the store drain (~0.4 us per launch)
    if (gt == 0) ptx::bulk_wait_read<0>();
#undef SLIDER_GROUP_BAR
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (TWO) {      // the peer may still signal this CTA's barriers / the leader's MMAs may still read its operands
    ptx::cluster_arrive();
    ptx::cluster_wait();
  }
  if (warp == 1) {
    ptx::tc_fence_after();
    if (TWO) ptx::tmem_dealloc_2cta(tmem_base, Cfg::TMEM_COLS);
    else ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
  if (threadIdx.x == 0) SLIDER_TRACE(7);
#undef SLIDER_TRACE
}

// ---------------------------------------------------------------------------------------------
// Split-K finalize: y = sum_splits(partial) + bias + s * h . Bup^T.  One thread per (row, 4-column piece, chunk c)
// also owns the same piece of chunk c + 4, so the row's h partials are read once per 8 output columns.
// Thread order (piece, row, chunk) follows the partial layout, so reads are fully coalesced.
template <typename T, int R4, int BN>
__global__ void __launch_bounds__(256)
splitk_finalize_kernel(const GemmParams p) {
  constexpr int NCHUNK = TileCfg<BN>::NCHUNK;
  constexpr size_t UNIT_FLOATS = TileCfg<BN>::UNIT_FLOATS;
  pdl_trigger();
  constexpr int PIECES = PCW / 4;                  // 16-byte pieces per partial row
  constexpr int HC = (NCHUNK - 1) / 2;             // chunk pairs (c, c + HC)
  constexpr int PER_TILE = BM * PIECES * HC;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)p.total_tiles * PER_TILE) return;
  const int tile = (int)(idx / PER_TILE);
  const int rem = (int)(idx - (long long)tile * PER_TILE);
  const int pc = rem % PIECES, row = (rem / PIECES) & (BM - 1), c = rem / (PIECES * BM);
  const TileCoord tc = tile_coord<BN>(p, tile);
  const int lw = row % p.TW, lh = (row / p.TW) % p.TH, lb = row / (p.TW * p.TH);
  const bool row_ok = row < p.rows && tc.w0 + lw < p.W && tc.h0 + lh < p.H && tc.b0 + lb < p.B;
  const int n0 = tc.n0 + c * PCW + pc * 4, n1 = n0 + HC * PCW;
  if (!row_ok || n0 >= p.N) return;
  const bool second = n1 < p.N;
  const size_t pix = ((size_t)(tc.b0 + lb) * p.H + (tc.h0 + lh)) * p.W + (tc.w0 + lw);
  const float* src = p.partial + (size_t)tile * p.ksplit * UNIT_FLOATS + ((size_t)c * BM + row) * PCW + pc * 4;
  const int hoff = p.group_n > 0 ? (tc.n0 / p.group_n) * p.rank : 0;
  const float* hsrc = p.partial + (size_t)tile * p.ksplit * UNIT_FLOATS + ((size_t)(NCHUNK - 1) * BM + row) * PCW + hoff;
  const T* __restrict__ bias = reinterpret_cast<const T*>(p.bias);
  const T* __restrict__ Bup = reinterpret_cast<const T*>(p.Bup);
  T* const dst0 = reinterpret_cast<T*>(p.y) + pix * (size_t)p.N + n0;
  T* const dst1 = reinterpret_cast<T*>(p.y) + pix * (size_t)p.N + n1;
  // Everything the epilogue needs besides the partials is requested BEFORE the split loop, so that its latency hides
  // behind the partial loads instead of following them: the frozen bias even before the dependent-launch wait, the
  // rank-4 bf16 up-projection rows (8 bytes per output column) and the residual right after it.
  float bs[8];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    bs[k] = bias != nullptr ? Elem<T>::to_f(bias[n0 + k]) : 0.f;
    bs[4 + k] = (bias != nullptr && second) ? Elem<T>::to_f(bias[n1 + k]) : 0.f;
  }
  pdl_wait();
  constexpr bool kPackB = R4 == 1 && sizeof(T) == 2;
  const bool packb = kPackB && p.rank == 4 && p.bup_ld == 4 && p.bup_jd == 1 && Bup != nullptr;
  uint2 bq[kPackB ? 8 : 1];
  if (kPackB && packb) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      bq[k] = *reinterpret_cast<const uint2*>(Bup + (size_t)(n0 + k) * 4);
      bq[4 + k] = *reinterpret_cast<const uint2*>(Bup + (size_t)(second ? n1 + k : n0 + k) * 4);
    }
  }
  uint4 resq[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};       // bf16: .x .y used; fp32: all four
  if (p.accumulate) {
    if (sizeof(T) == 2) {
      const uint2 q0 = *reinterpret_cast<const uint2*>(dst0), q1 = *reinterpret_cast<const uint2*>(second ? dst1 : dst0);
      resq[0].x = q0.x; resq[0].y = q0.y; resq[1].x = q1.x; resq[1].y = q1.y;
    } else {
      resq[0] = *reinterpret_cast<const uint4*>(dst0);
      resq[1] = *reinterpret_cast<const uint4*>(second ? dst1 : dst0);
    }
  }
  float o[8];
  float hv[R4 > 0 ? 4 * R4 : 1];
#pragma unroll
  for (int k = 0; k < 8; ++k) o[k] = 0.f;
#pragma unroll
  for (int j = 0; j < (R4 > 0 ? 4 * R4 : 1); ++j) hv[j] = 0.f;
  // branch-free loop bodies (the second chunk of a ragged tile re-reads the first and is discarded), so that the
  // unrolled loads of several splits are in flight together: the kernel is pure L2 latency otherwise
  const float* srcb = second ? src + (size_t)HC * BM * PCW : src;
  if (R4 == 0 || (hoff & 3) == 0) {
#pragma unroll 6
    for (int s = 0; s < p.ksplit; ++s) {
      const float4 a = __ldcg(reinterpret_cast<const float4*>(src + (size_t)s * UNIT_FLOATS));
      const float4 b = __ldcg(reinterpret_cast<const float4*>(srcb + (size_t)s * UNIT_FLOATS));
      o[0] += a.x; o[1] += a.y; o[2] += a.z; o[3] += a.w;
      o[4] += b.x; o[5] += b.y; o[6] += b.z; o[7] += b.w;
      if (R4 > 0) {
#pragma unroll
        for (int j4 = 0; j4 < R4; ++j4) {
          const float4 h4 = __ldcg(reinterpret_cast<const float4*>(hsrc + (size_t)s * UNIT_FLOATS) + j4);
          hv[4 * j4] += h4.x; hv[4 * j4 + 1] += h4.y; hv[4 * j4 + 2] += h4.z; hv[4 * j4 + 3] += h4.w;
        }
      }
    }
  } else {    // grouped adapters whose h offset is not a multiple of four columns: scalar h loads
#pragma unroll 2
    for (int s = 0; s < p.ksplit; ++s) {
      const float4 a = __ldcg(reinterpret_cast<const float4*>(src + (size_t)s * UNIT_FLOATS));
      const float4 b = __ldcg(reinterpret_cast<const float4*>(srcb + (size_t)s * UNIT_FLOATS));
      o[0] += a.x; o[1] += a.y; o[2] += a.z; o[3] += a.w;
      o[4] += b.x; o[5] += b.y; o[6] += b.z; o[7] += b.w;
#pragma unroll
      for (int j = 0; j < (R4 > 0 ? 4 * R4 : 1); ++j)
        if (R4 > 0 && hoff + j < HCOLS) hv[j] += __ldcg(hsrc + (size_t)s * UNIT_FLOATS + j);
    }
  }
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    if (half == 1 && !second) break;
    const int n = half == 0 ? n0 : n1;
    float r[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float a = o[half * 4 + k] + bs[half * 4 + k];
      if (R4 > 0) {
        float l = 0.f;
        if (kPackB && packb) {
          const uint2 q = bq[kPackB ? half * 4 + k : 0];
          float b0, b1, b2, b3;
          Elem<T>::unpack2(q.x, b0, b1);
          Elem<T>::unpack2(q.y, b2, b3);
          l = fmaf(hv[0], b0, l);
          l = fmaf(hv[1 % (R4 > 0 ? 4 * R4 : 1)], b1, l);
          l = fmaf(hv[2 % (R4 > 0 ? 4 * R4 : 1)], b2, l);
          l = fmaf(hv[3 % (R4 > 0 ? 4 * R4 : 1)], b3, l);
        } else {
#pragma unroll
          for (int j = 0; j < 4 * R4; ++j)
            if (j < p.rank) l = fmaf(hv[j], Elem<T>::to_f(Bup[(size_t)(n + k) * p.bup_ld + (size_t)j * p.bup_jd]), l);
        }
        a = fmaf(p.s, l, a);
      }
      r[k] = a;
    }
    T* dst = half == 0 ? dst0 : dst1;
    if (sizeof(T) == 2) {
      uint2* d2 = reinterpret_cast<uint2*>(dst);
      if (p.accumulate) {  // y holds the residual: same two roundings as `residual + module(x)` in element dtype
        const uint4 q = resq[half];
        float res[4];
        Elem<T>::unpack2(q.x, res[0], res[1]);
        Elem<T>::unpack2(q.y, res[2], res[3]);
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = Elem<T>::round(r[k]) + res[k];
      }
      *d2 = make_uint2(Elem<T>::pack2(r[0], r[1]), Elem<T>::pack2(r[2], r[3]));
    } else {
      float4* d4 = reinterpret_cast<float4*>(dst);
      if (p.accumulate) {
        const uint4 q = resq[half];
        r[0] += __uint_as_float(q.x); r[1] += __uint_as_float(q.y); r[2] += __uint_as_float(q.z); r[3] += __uint_as_float(q.w);
      }
      *d4 = make_float4(r[0], r[1], r[2], r[3]);
    }
  }
  if (R4 > 0 && p.h_out != nullptr && tc.n_blk == 0 && c == 0 && pc == 0) {
#pragma unroll
    for (int j = 0; j < 4 * R4; ++j)
      if (j < p.rank) p.h_out[pix * p.rank + j] = hv[j];
  }
}


// ---------------------------------------------------------------------------------------------
// Launch of one tile width (explicitly instantiated in lora_gemm_bn160.cu / lora_gemm_bn80.cu so that the two sets of
// kernels compile in parallel).
template <typename T, int BN, bool TWO>
static int launch_gemm_t(int r4, const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& l,
                         const CUtensorMap& xe, const CUtensorMap& we, const CUtensorMap& my, const CUtensorMap& mp,
                         const GemmParams& p, cudaStream_t st) {
  using Cfg = TileCfg<BN>;
  void (*kern)(CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap,
               GemmParams) = nullptr;
  void (*fin)(GemmParams) = nullptr;
  switch (r4) {
    case 0: kern = lora_gemm_kernel<T, 0, BN, TWO>; fin = splitk_finalize_kernel<T, 0, BN>; break;
    case 1: kern = lora_gemm_kernel<T, 1, BN, TWO>; fin = splitk_finalize_kernel<T, 1, BN>; break;
    case 2: kern = lora_gemm_kernel<T, 2, BN, TWO>; fin = splitk_finalize_kernel<T, 2, BN>; break;
    default: kern = lora_gemm_kernel<T, 4, BN, TWO>; fin = splitk_finalize_kernel<T, 4, BN>; break;
  }
  static thread_local const void* configured[8] = {nullptr};
  bool done = false;
  for (auto c : configured) done |= (c == (const void*)kern);
  if (!done) {
    SLIDER_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes(r4)));
    for (auto& c : configured) if (c == nullptr) { c = (const void*)kern; break; }
  }
  if (TWO) {
    // one cluster of 2 per pair of SMs; a pair walks the pair units
    const int pairs = p.pair_units < sm_count() / 2 ? p.pair_units : sm_count() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = Cfg::smem_bytes(r4); cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaLaunchKernelEx(&cfg, kern, a, w, l, xe, we, my, mp, p);
  } else {
    const int grid = p.total_units < sm_count() ? p.total_units : sm_count();
    launch_pdl(kern, dim3(grid), dim3(kGemmThreads), Cfg::smem_bytes(r4), st, a, w, l, xe, we, my, mp, p);
  }
  int rc = post_launch("lora_gemm_kernel");
  if (rc || p.ksplit == 1) return rc;
  const long long threads = (long long)p.total_tiles * BM * (BN / 8);
  launch_pdl(fin, dim3((unsigned)((threads + 255) / 256)), dim3(256), 0, st, p);
  return post_launch("splitk_finalize_kernel");
}

template <int BN>
int launch_gemm_tile(int dtype, int r4, const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& l,
                     const CUtensorMap& xe, const CUtensorMap& we, const CUtensorMap& my, const CUtensorMap& mp,
                     const GemmParams& p, cudaStream_t st) {
  if (dtype == SLIDER_BF16) return launch_gemm_t<__nv_bfloat16, BN, false>(r4, a, w, l, xe, we, my, mp, p, st);
  return launch_gemm_t<float, BN, false>(r4, a, w, l, xe, we, my, mp, p, st);
}

// fp16 storage (forward / inference): instantiated in lora_gemm_f16.cu
int launch_gemm_f16(int bn, int r4, const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& l, const CUtensorMap& xe,
                    const CUtensorMap& we, const CUtensorMap& my, const CUtensorMap& mp, const GemmParams& p,
                    cudaStream_t st);

// CTA-pair form: bf16 storage, 128 x 160 tiles (instantiated in lora_gemm_pair.cu)
int launch_gemm_pair(int r4, const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& l, const CUtensorMap& xe,
                     const CUtensorMap& we, const CUtensorMap& my, const CUtensorMap& mp, const GemmParams& p,
                     cudaStream_t st);

}  // namespace slider
