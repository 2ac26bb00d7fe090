// Fused base + low-rank implicit GEMM for sm_100a (tcgen05 + TMEM + TMA).
//
//   y[pix, n] = sum_{tap,c} x[pix + d(tap), c] * W[n, tap, c] + bias[n]
//             + s * sum_j ( sum_{tap,c} x[pix + d(tap), c] * A[j, tap, c] ) * Bup[n, j]
//
// One persistent CTA per SM, 6 warps:
//   warp 0      TMA producer  (4-D box loads of NHWC activations: padding = TMA OOB zero fill,
//                              3-D box loads of the tap-major weights)
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer
//   warps 2..5  epilogue: tcgen05.ld -> registers -> (+bias, + s*h.Bup) -> global
// The rank-r intermediate h = x (*) A lives in 16 extra TMEM columns next to the main
// accumulator: it is produced by a second, N=16 MMA that re-uses the activation tile already
// staged in shared memory, and is folded into the output in the epilogue with r FMAs per
// element.  The accumulators are double buffered so the epilogue of tile i overlaps the
// main loop of tile i+1.
//
// The same kernel computes the input gradient ("extended-K" mode): activations = gy,
// weights = Wt, and one extra 16-wide K block per tap whose operands are dh and At.
#include <stdlib.h>
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.cuh"

namespace slider {

constexpr int BM = 128;          // pixels (GEMM rows) per tile
constexpr int BN = 160;          // output channels per tile; divides 320/640/960/1280/1920/2560
constexpr int HCOLS = 16;        // TMEM columns of the low-rank accumulator (rank padded to 16)
constexpr int ROW_BYTES = 128;   // K-block row = one 128-byte swizzle row
constexpr int NSTAGES = 4;
constexpr int ACC_STAGES = 2;
constexpr int A_BYTES = BM * ROW_BYTES;      // 16 KB
constexpr int B_BYTES = BN * ROW_BYTES;      // 20 KB
constexpr int L_BYTES = HCOLS * ROW_BYTES;   // 2 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES + L_BYTES;
constexpr int EXT_ELEMS = 16;    // width of the extended-K operand (dh / At), in elements
constexpr int TMEM_COLS = 512;
constexpr int kGemmThreads = 320;            // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two groups of four)
constexpr int kEpiThreads = 256;
constexpr int HALF_N = BN / 2;               // output columns owned by one epilogue group
constexpr int OUT_ROW_BYTES = 80;            // staging row of one store chunk: 40 bf16 / 20 fp32 columns; an 80-byte
                                             // pitch makes the 16-byte st.shared of 8 consecutive rows conflict-free
constexpr int OUT_BYTES = BM * OUT_ROW_BYTES;   // 10 KB per chunk buffer
// staging buffers per epilogue group: three when the rank table leaves room (rank <= 8), else two
__host__ __device__ constexpr int out_bufs_per_group(int r4) { return r4 <= 2 ? 3 : 2; }
__host__ __device__ constexpr int epi_floats(int r4) { return BN * 4 * (r4 > 0 ? r4 : 1) + BN; }   // s*Bup + bias
__host__ __device__ constexpr size_t gemm_smem_bytes(int r4) {
  return 1024 /*align slack*/ + (size_t)NSTAGES * STAGE_BYTES + (size_t)2 * out_bufs_per_group(r4) * OUT_BYTES +
         (size_t)epi_floats(r4) * 4 + 256;
}
constexpr int PCW = OUT_ROW_BYTES / 4;       // fp32 columns per split-K partial chunk
constexpr int NCHUNK = BN / PCW + 1;         // partial chunks of a split-K unit (+1 for h)
constexpr size_t UNIT_FLOATS = (size_t)NCHUNK * BM * PCW;
constexpr int CPITCH = (BN + HCOLS + 4) * 4;  // bytes per row of a cluster split-K partial tile: 180 words, conflict-free
constexpr int MAX_CLUSTER = 8;               // portable cluster size = largest in-cluster K split
constexpr int MAX_UNITS = 160;               // split-K work units the caller workspace must hold

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
  static constexpr int kStages = NSTAGES;
  static constexpr int kStageStride = STAGE_BYTES;
  __device__ static inline void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    ptx::mma_f16(d, a, b, idesc, acc);
  }
};
template <> struct GemmTraits<float> {
  static constexpr int kBK = 32;
  static constexpr uint32_t kFmt = 2;  // tf32
  static constexpr int kExtRowBytes = EXT_ELEMS * 4;
  // fp32 storage: error-compensated 3xTF32.  Every landed stage is split in shared memory into
  // hi = x with the 13 low mantissa bits cleared (exact in tf32) and lo = x - hi (second half of the
  // stage slot), and each K step issues hi.hi + lo.hi + hi.lo into the same accumulator (~2^-21 rel).
  static constexpr bool kSplit3 = true;
  static constexpr int kStages = NSTAGES / 2;
  static constexpr int kStageStride = 2 * STAGE_BYTES;
  __device__ static inline void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    ptx::mma_tf32(d, a, b, idesc, acc);
  }
};

struct GemmParams {
  int TW, TH, TB, rows;   // tile box in pixels; rows = TW*TH*TB <= 128
  int W, H, B;
  int mt_w, mt_h, mt_b;   // number of pixel tiles per dimension
  int n_tiles, N;
  int taps, kw, pad, cblocks;
  int rank;
  float s;
  int lora;               // accumulate h and apply it in the epilogue
  int ext;                // extended-K operand pair present (input-gradient mode)
  int total_tiles;
  int per_tap;            // K iterations per tap = cblocks + ext
  int kiters;             // taps * per_tap
  int ksplit;             // K partitions per tile (1 = no split)
  int cluster;            // split-K partials are reduced inside a thread-block cluster of ksplit CTAs (DSMEM)
  int total_units;        // total_tiles * ksplit
  float* partial;         // split-K partial tiles, [unit][NCHUNK][128][32] fp32
  int accumulate;         // y += result (y holds the residual branch) instead of y = result
  int group_n;            // grouped adapters: output columns per group (0 = one adapter); group g uses h columns [g*rank, (g+1)*rank)
  FastDiv fd_n_tiles, fd_mt_w, fd_mt_h, fd_ksplit, fd_per_tap, fd_kw;
  int w_prefetch;         // K blocks the weight stream is requested into L2 ahead of the shared-memory pipeline
  long long* trace;       // debug: 16 clock64() stamps per CTA (NULL = off)
  void* y;
  const void* bias;
  const void* Bup;
  float* h_out;
};

struct TileCoord { int w0, h0, b0, n0, n_blk; };

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

template <int CW> struct TmemChunk;
template <> struct TmemChunk<40> {
  __device__ static __forceinline__ void ld(uint32_t taddr, uint32_t (&v)[40]) {
    uint32_t a[32], b[8];
    ptx::tmem_ld_32x32(taddr, a);
    ptx::tmem_ld_32x8(taddr + 32, b);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = a[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[32 + i] = b[i];
  }
};
template <> struct TmemChunk<20> {
  __device__ static __forceinline__ void ld(uint32_t taddr, uint32_t (&v)[20]) {
    uint32_t a[16], b[4];
    ptx::tmem_ld_32x16(taddr, a);
    ptx::tmem_ld_32x4(taddr + 16, b);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = a[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[16 + i] = b[i];
  }
};

template <typename T, int R4>
__global__ void __launch_bounds__(kGemmThreads, 1)
lora_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmL, const __grid_constant__ CUtensorMap tmXe,
                 const __grid_constant__ CUtensorMap tmWe, const __grid_constant__ CUtensorMap tmY,
                 const __grid_constant__ CUtensorMap tmP, const GemmParams p) {
  using Tr = GemmTraits<T>;
  constexpr int BK = Tr::kBK;
  constexpr int KSTEPS = ROW_BYTES / 32;   // one tcgen05.mma consumes 32 bytes of K per row
  constexpr int EXT_KSTEPS = Tr::kExtRowBytes / 32;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int NB = out_bufs_per_group(R4);               // staging buffers per epilogue group
  uint8_t* s_out = smem + (size_t)NSTAGES * STAGE_BYTES;   // 2 x NB x 10 KB staging of the TMA-store epilogue
  float* s_epi = reinterpret_cast<float*>(s_out + (size_t)2 * NB * OUT_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_epi + epi_floats(R4));
  uint64_t* full_bar = bars;                      // [NSTAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + NSTAGES;           // [NSTAGES]  MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * NSTAGES;       // [ACC_STAGES] MMA -> epilogue
  uint64_t* tempty_bar = tfull_bar + ACC_STAGES;  // [ACC_STAGES] epilogue -> MMA
  uint64_t* xf_bar = tempty_bar + ACC_STAGES;     // [NSTAGES] hi/lo split done (fp32 only) -> MMA
  uint64_t* recv_bar = xf_bar + NSTAGES;          // cluster split-K: bytes pushed by the peer CTAs have landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(recv_bar + 1);

  // warp index through a shuffle so that the role branches below are provably warp-uniform
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
#define SLIDER_TRACE(slot) do { if (p.trace != nullptr) p.trace[blockIdx.x * 16 + (slot)] = clock64(); } while (0)
  pdl_trigger();
  if (threadIdx.x == 0) SLIDER_TRACE(0);
  const uint32_t smem_base = ptx::smem_u32(smem);
  const uint32_t main_tx = (uint32_t)(p.rows * ROW_BYTES + B_BYTES + (p.lora ? L_BYTES : 0));
  const uint32_t ext_tx = (uint32_t)((p.rows + BN) * Tr::kExtRowBytes);

  // ---- TMA producer state (thread 0 only).  The producer issues its first kStages K blocks BEFORE the CTA-wide
  // barrier below, so the first operand fetch overlaps the TMEM allocation and the other warps' start-up.
  int pu = (int)blockIdx.x, pit = 0, pit1 = 0, ptap = 0, pj = 0;
  uint32_t pstage = 0, pphase = 0;
  TileCoord ptc = {0, 0, 0, 0, 0};
  bool phave = false;
  auto prod_begin = [&]() {
    phave = pu < p.total_units;
    if (!phave) return;
    uint32_t tile = (uint32_t)pu, split = 0;
    if (p.ksplit > 1) p.fd_ksplit.divmod((uint32_t)pu, tile, split);
    ptc = tile_coord(p, (int)tile);
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
    if (j < p.cblocks) ptx::tma_prefetch_l2_3d(&tmW, j * BK, tap, ptc.n0);
  };
  auto prod_owner = [&]() { return p.w_prefetch > 0 && ptc.w0 == 0 && ptc.h0 == 0 && ptc.b0 == 0; };
  auto prod_issue = [&]() {
    const int ty = (int)p.fd_kw.div((uint32_t)ptap);
    const int dy = ty - p.pad, dx = ptap - ty * p.kw - p.pad;
    ptx::mbar_wait(ptx::smem_u32(empty_bar + pstage), pphase ^ 1);
    const uint32_t sa = smem_base + pstage * Tr::kStageStride;
    const uint32_t fb = ptx::smem_u32(full_bar + pstage);
    if (pj < p.cblocks) {
      ptx::mbar_arrive_expect_tx(fb, main_tx);
      ptx::tma_load_4d(sa, &tmA, fb, pj * BK, ptc.w0 + dx, ptc.h0 + dy, ptc.b0);
      ptx::tma_load_3d(sa + A_BYTES, &tmW, fb, pj * BK, ptap, ptc.n0);
      if (p.lora) ptx::tma_load_3d(sa + A_BYTES + B_BYTES, &tmL, fb, pj * BK, ptap, 0);
    } else {
      ptx::mbar_arrive_expect_tx(fb, ext_tx);
      ptx::tma_load_4d(sa, &tmXe, fb, 0, ptc.w0 + dx, ptc.h0 + dy, ptc.b0);
      ptx::tma_load_3d(sa + A_BYTES, &tmWe, fb, 0, ptap, ptc.n0);
    }
    if (prod_owner() && pit + p.w_prefetch < pit1) prod_prefetch_w(pit + p.w_prefetch);
    if (++pstage == (uint32_t)Tr::kStages) { pstage = 0; pphase ^= 1; }
    if (++pj == p.per_tap) { pj = 0; ++ptap; }
    if (++pit == pit1) {
      pu += (int)gridDim.x;
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
    if (p.ksplit > 1 && !p.cluster) ptx::prefetch_tmap(&tmP); else if (!p.cluster) ptx::prefetch_tmap(&tmY);
    for (int i = 0; i < NSTAGES; ++i) {
      ptx::mbar_init(ptx::smem_u32(full_bar + i), 1);
      ptx::mbar_init(ptx::smem_u32(empty_bar + i), 1);
      ptx::mbar_init(ptx::smem_u32(xf_bar + i), 8);      // one arrive per epilogue warp
    }
    for (int i = 0; i < ACC_STAGES; ++i) {
      ptx::mbar_init(ptx::smem_u32(tfull_bar + i), 1);
      ptx::mbar_init(ptx::smem_u32(tempty_bar + i), 8);  // one arrive per epilogue warp
    }
    ptx::mbar_init(ptx::smem_u32(recv_bar), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  // CTA-wide start barrier.  Warp 0 only ARRIVES (its lane 0 initialised the mbarriers above and is the one
  // thread that never needs the TMEM base), so the producer starts fetching operands while the TMEM allocation
  // and the other warps' start-up are still in flight.
  if (warp == 0) {
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
  } else if (warp == 1 && ptx::elect_one()) {
    // ===================== MMA issuer =====================
    // With the adapter on, W tile (BN rows) and lora_down tile (HCOLS rows) are adjacent in shared
    // memory and their accumulators adjacent in TMEM: ONE tcgen05.mma of N = BN + HCOLS covers both.
    constexpr uint32_t idesc_main = ptx::make_idesc(Tr::kFmt, BM, BN);
    constexpr uint32_t idesc_lora = ptx::make_idesc(Tr::kFmt, BM, BN + HCOLS);
    const uint32_t idesc = p.lora ? idesc_lora : idesc_main;
    uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
    for (int unit = blockIdx.x; unit < p.total_units; unit += gridDim.x) {
      int it0 = 0, it1 = p.kiters, j = 0;
      if (p.ksplit > 1) {
        uint32_t tile, split;
        p.fd_ksplit.divmod((uint32_t)unit, tile, split);
        split_range(p, split, it0, it1);
        j = it0 - (int)p.fd_per_tap.div((uint32_t)it0) * p.per_tap;
      }
      ptx::mbar_wait(ptx::smem_u32(tempty_bar + as), aphase ^ 1);
      ptx::tc_fence_after();
      const uint32_t d_main = tmem_base + as * (BN + HCOLS);
      uint32_t acc = 0;
      for (int it = it0; it < it1; ++it) {
        ptx::mbar_wait(ptx::smem_u32((Tr::kSplit3 ? xf_bar : full_bar) + stage), phase);
        ptx::tc_fence_after();
        if (unit == (int)blockIdx.x && it == it0) SLIDER_TRACE(3);
        const uint32_t sa = smem_base + stage * Tr::kStageStride;
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
        } else if (main_blk) {
#pragma unroll
          for (int k = 0; k < KSTEPS; ++k) Tr::mma(d_main, da + 2 * k, db + 2 * k, id, acc | (uint32_t)k);
        } else {
#pragma unroll
          for (int k = 0; k < EXT_KSTEPS; ++k) Tr::mma(d_main, da + 2 * k, db + 2 * k, id, acc | (uint32_t)k);
        }
        acc = 1;
        ptx::mma_commit(ptx::smem_u32(empty_bar + stage));
        if (++stage == Tr::kStages) { stage = 0; phase ^= 1; }
        if (++j == p.per_tap) j = 0;
      }
      ptx::mma_commit(ptx::smem_u32(tfull_bar + as));
      if (unit == (int)blockIdx.x) SLIDER_TRACE(4);
      if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
    }
  } else if (warp >= 2) {
    // ===================== epilogue =====================
    // Eight warps = two groups of four; a group covers all 128 accumulator rows (one TMEM lane quarter per warp)
    // and owns HALF_N = 80 of the tile's columns.  Per 80-byte chunk (40 bf16 / 20 fp32 columns):
    // tcgen05.ld -> registers -> (+bias + s*h.Bup^T, cast) -> shared staging -> one TMA store (box {chunk, TW, TH,
    // TB}: clips out-of-range pixels/channels).  NB staging buffers and one named barrier per group: before the
    // barrier of chunk c the group's issuing thread waits until at most NB-2 stores still read shared memory, so the
    // buffer the group fills next (chunk c+1, last used by store c+1-NB) is free once its threads pass that barrier.
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
    uint32_t as = 0, aphase = 0, ob = 0, xs = 0, xphase = 0;
    const T* __restrict__ bias = reinterpret_cast<const T*>(p.bias);
    const T* __restrict__ Bup = reinterpret_cast<const T*>(p.Bup);
#define SLIDER_GROUP_BAR() asm volatile("bar.sync %0, 128;" ::"r"(3 + grp) : "memory")
    for (int unit = blockIdx.x; unit < p.total_units; unit += gridDim.x) {
      uint32_t tile = (uint32_t)unit, split = 0;
      if (p.ksplit > 1) p.fd_ksplit.divmod((uint32_t)unit, tile, split);
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
          float4* hi = reinterpret_cast<float4*>(smem + (size_t)xs * Tr::kStageStride);
          float4* lo = reinterpret_cast<float4*>(smem + (size_t)xs * Tr::kStageStride + STAGE_BYTES);
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
          if (++xs == (uint32_t)Tr::kStages) { xs = 0; xphase ^= 1; }
        }
      }
      if (p.cluster) {
        // ---- split-K inside a thread-block cluster (one unit per CTA, cluster rank == split) ----
        // Every CTA parks its fp32 partial tile (main + h columns) in its own shared memory (the operand stages are
        // idle once the accumulator is complete), the cluster synchronises, and CTA r reduces rows
        // [r*R, (r+1)*R) of all partials over distributed shared memory in split order (deterministic), applies the
        // epilogue and writes y directly: no workspace round trip, no second launch.
        const TileCoord tc = tile_coord(p, (int)tile);
        for (int i = et; i < BN; i += kEpiThreads) {
          const int n = tc.n0 + i;
          s_bias[i] = (bias != nullptr && n < p.N) ? Elem<T>::to_f(bias[n]) : 0.f;
          if (R4 > 0) {
#pragma unroll
            for (int j = 0; j < 4 * R4; ++j)
              s_epi[i * (4 * R4) + j] =
                  (n < p.N && j < p.rank) ? p.s * Elem<T>::to_f(Bup[(size_t)n * p.rank + j]) : 0.f;
          }
        }
        if (et < 32) ptx::mbar_wait(ptx::smem_u32(tfull_bar + as), aphase);
        asm volatile("bar.sync 2, 256;" ::: "memory");
        ptx::tc_fence_after();
        if (et == 0) SLIDER_TRACE(5);
        const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + as * (BN + HCOLS);
        const int S = p.ksplit;
        const int R = (BM + S - 1) / S;                                // rows reduced by one CTA
        const int r0 = (int)split * R;
        const int nrows = max(0, min(R, BM - r0));
        const uint32_t part = smem_base;                               // [128][CPITCH] fp32 partial of this CTA
        const uint32_t recv = smem_base + (uint32_t)BM * CPITCH;       // [S-1][R][CPITCH] row slices pushed by the peers
        const uint32_t sums = part + (uint32_t)r0 * CPITCH;            // reduced in place over this CTA's own rows
        const uint32_t rbar = ptx::smem_u32(recv_bar);
        {
          // group 0 parks 16-column pieces 0..5, group 1 pieces 6..10 (the last one is h)
          const int pc0 = grp == 0 ? 0 : 6, pc1 = grp == 0 ? 6 : (BN + HCOLS) / 16;
#pragma unroll 1
          for (int pc = pc0; pc < pc1; ++pc) {
            uint32_t v[16];
            ptx::tmem_ld_32x16(t_acc + pc * 16, v);
            ptx::tmem_ld_wait();
            const uint32_t dst = part + row * CPITCH + pc * 64;
#pragma unroll
            for (int k = 0; k < 4; ++k) ptx::st_shared_v4(dst + k * 16, v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
          }
        }
        ptx::tc_fence_before();
        ptx::fence_proxy_async_smem();          // the partial is read by the bulk-copy (async proxy) below
        if (et == 0) ptx::mbar_arrive_expect_tx(rbar, (uint32_t)((S - 1) * nrows * CPITCH));
        // cluster barrier 1: every CTA's partial is parked, its operand stages are idle (they hold the receive
        // buffer) and its receive barrier is armed
        ptx::cluster_arrive();
        ptx::cluster_wait();
        if (et == 0) {
          SLIDER_TRACE(11);
          // push my rows [r*R, ...) to peer r: one bulk DMA per peer, no thread touches remote shared memory
          for (int r = 0; r < S; ++r) {
            if (r == (int)split) continue;
            const int pr0 = r * R, pn = max(0, min(R, BM - pr0));
            if (pn == 0) continue;
            const int slot = (int)split < r ? (int)split : (int)split - 1;
            ptx::bulk_copy_to_peer(ptx::mapa(recv + (uint32_t)(slot * R) * CPITCH, (uint32_t)r),
                                   part + (uint32_t)pr0 * CPITCH, (uint32_t)(pn * CPITCH), ptx::mapa(rbar, (uint32_t)r));
          }
        }
        if (et < 32) ptx::mbar_wait(rbar, 0);
        asm volatile("bar.sync 2, 256;" ::: "memory");
        if (et == 0) SLIDER_TRACE(12);
        // sum the S row slices in split order (deterministic): mine in place, the peers' from the receive buffer
        constexpr int PIECES = (BN + HCOLS) / 4;                       // 16-byte pieces per row (44)
        for (int i = et; i < nrows * PIECES; i += kEpiThreads) {
          const int rr = i / PIECES, pc = i - rr * PIECES;
          const uint32_t off = (uint32_t)rr * CPITCH + pc * 16;
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int sp = 0; sp < MAX_CLUSTER; ++sp) {
            if (sp < S) {
              const int slot = sp < (int)split ? sp : sp - 1;
              const float4 v = ptx::ld_shared_f4(sp == (int)split ? sums + off : recv + (uint32_t)(slot * R) * CPITCH + off);
              acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
          }
          ptx::st_shared_f4(sums + off, acc);
        }
        asm volatile("bar.sync 2, 256;" ::: "memory");
        const int hoff = p.group_n > 0 ? (tc.n0 / p.group_n) * p.rank : 0;
        for (int i = et; i < nrows * (BN / 8); i += kEpiThreads) {
          const int rr = i / (BN / 8), g8 = i - rr * (BN / 8);
          const int n = tc.n0 + g8 * 8;
          const int trow = r0 + rr;
          const int lw = trow % p.TW, lh = (trow / p.TW) % p.TH, lb = trow / (p.TW * p.TH);
          const bool row_ok = trow < p.rows && tc.w0 + lw < p.W && tc.h0 + lh < p.H && tc.b0 + lb < p.B;
          if (!row_ok || n >= p.N) continue;
          const size_t pix = ((size_t)(tc.b0 + lb) * p.H + (tc.h0 + lh)) * p.W + (tc.w0 + lw);
          const uint32_t srow = sums + rr * CPITCH;
          const float4 a0 = ptx::ld_shared_f4(srow + g8 * 32), a1 = ptx::ld_shared_f4(srow + g8 * 32 + 16);
          float o[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
          float hv[R4 > 0 ? 4 * R4 : 1];
          if (R4 > 0) {
#pragma unroll
            for (int j = 0; j < 4 * R4; ++j) hv[j] = ptx::ld_shared_f32(srow + (BN + min(hoff + j, HCOLS - 1)) * 4);
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int ci = g8 * 8 + k;
            float a = o[k] + ptx::ld_shared_f32(bias_s + ci * 4);
            if (R4 > 0) {
#pragma unroll
              for (int j4 = 0; j4 < R4; ++j4) {
                const float4 b = ptx::ld_shared_f4(epi_s + ci * (16 * R4) + j4 * 16);
                a = fmaf(hv[4 * j4 + 0], b.x, a);
                a = fmaf(hv[4 * j4 + 1], b.y, a);
                a = fmaf(hv[4 * j4 + 2], b.z, a);
                a = fmaf(hv[4 * j4 + 3], b.w, a);
              }
            }
            o[k] = a;
          }
          T* dst = reinterpret_cast<T*>(p.y) + pix * (size_t)p.N + n;
          if (sizeof(T) == 2) {
            __nv_bfloat16* d16 = reinterpret_cast<__nv_bfloat16*>(dst);
            if (p.accumulate) {   // y holds the residual: same two roundings as `residual + module(x)` in element dtype
              float res[8];
              Elem<__nv_bfloat16>::load(d16, res);
#pragma unroll
              for (int k = 0; k < 8; ++k) o[k] = __bfloat162float(__float2bfloat16_rn(o[k])) + res[k];
            }
            Elem<__nv_bfloat16>::store(d16, o);
          } else {
            float4* d4 = reinterpret_cast<float4*>(dst);
            if (p.accumulate) {
              const float4 q0 = d4[0], q1 = d4[1];
              o[0] += q0.x; o[1] += q0.y; o[2] += q0.z; o[3] += q0.w; o[4] += q1.x; o[5] += q1.y; o[6] += q1.z; o[7] += q1.w;
            }
            d4[0] = make_float4(o[0], o[1], o[2], o[3]);
            d4[1] = make_float4(o[4], o[5], o[6], o[7]);
          }
          if (R4 > 0 && p.h_out != nullptr && tc.n_blk == 0 && g8 == 0) {
#pragma unroll
            for (int j = 0; j < 4 * R4; ++j)
              if (j < p.rank) p.h_out[pix * p.rank + j] = hv[j];
          }
        }
        if (et == 0) SLIDER_TRACE(6);
        // the peers may still be reading this CTA's partial: second cluster barrier before anyone exits
        ptx::cluster_arrive();
        ptx::cluster_wait();
        continue;
      }
      if (p.ksplit > 1) {
        // split-K: park the fp32 partial tile (main + h columns) in the L2-resident workspace,
        // [unit][chunk][row][PCW]; splitk_finalize_kernel reduces over the splits in a fixed order.
        // one warp polls the accumulator-ready mbarrier, the others sleep on the hardware barrier: eight warps
        // spinning on try_wait steal shared-memory cycles from the operand pipeline for the whole main loop
        if (et < 32) ptx::mbar_wait(ptx::smem_u32(tfull_bar + as), aphase);
        asm volatile("bar.sync 2, 256;" ::: "memory");
        ptx::tc_fence_after();
        if (et == 0 && unit == (int)blockIdx.x) SLIDER_TRACE(5);
        const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + as * (BN + HCOLS);
        const int nmine = PCPG + ((grp == 0 && R4 > 0) ? 1 : 0);   // group 0 also parks the h columns
#pragma unroll 1
        for (int ci = 0; ci < nmine; ++ci) {
          uint32_t v[PCW];
          int chunk;
          if (ci < PCPG) {
            chunk = grp * PCPG + ci;
            TmemChunk<PCW>::ld(t_acc + chunk * PCW, v);
          } else {
            chunk = NCHUNK - 1;
            uint32_t hr[16];
            ptx::tmem_ld_32x16(t_acc + BN, hr);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = hr[k];
#pragma unroll
            for (int k = 16; k < PCW; ++k) v[k] = 0u;
          }
          const uint32_t dst = out_base + ob * OUT_BYTES + row * OUT_ROW_BYTES;
#pragma unroll
          for (int k = 0; k < OUT_ROW_BYTES / 16; ++k)
            ptx::st_shared_v4(dst + k * 16, v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
          ptx::fence_proxy_async_smem();
          if (gt == 0) ptx::bulk_wait_read<NB - 2>();   // buffer of chunk c+1 (last read by store c+1-NB) is free past the barrier
          SLIDER_GROUP_BAR();
          if (gt == 0) {
            ptx::tma_store_2d(&tmP, out_base + ob * OUT_BYTES, 0, (unit * NCHUNK + chunk) * BM);
            ptx::bulk_commit();
          }
          if (++ob == (uint32_t)NB) ob = 0;
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (et == 0 && unit == (int)blockIdx.x) SLIDER_TRACE(6);
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(tempty_bar + as));
        if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
        continue;
      }
      const TileCoord tc = tile_coord(p, unit);
      // stage s*Bup[n0.., :] and bias[n0..] for this tile (fp32) while the MMAs run
      asm volatile("bar.sync 2, 256;" ::: "memory");
      for (int i = et; i < BN; i += kEpiThreads) {
        const int n = tc.n0 + i;
        s_bias[i] = (bias != nullptr && n < p.N) ? Elem<T>::to_f(bias[n]) : 0.f;
        if (R4 > 0) {
#pragma unroll
          for (int j = 0; j < 4 * R4; ++j)
            s_epi[i * (4 * R4) + j] =
                (n < p.N && j < p.rank) ? p.s * Elem<T>::to_f(Bup[(size_t)n * p.rank + j]) : 0.f;
        }
      }
      // one warp polls the accumulator-ready mbarrier, the others sleep on the hardware barrier (which also
      // publishes the tables staged above): eight warps spinning on try_wait steal shared-memory cycles from the
      // operand pipeline for the whole main loop
      if (et < 32) ptx::mbar_wait(ptx::smem_u32(tfull_bar + as), aphase);
      asm volatile("bar.sync 2, 256;" ::: "memory");
      ptx::tc_fence_after();
      if (et == 0 && unit == (int)blockIdx.x) SLIDER_TRACE(5);
      const uint32_t t_main = tmem_base + ((uint32_t)(q * 32) << 16) + as * (BN + HCOLS);

      float hv[R4 > 0 ? 4 * R4 : 1];
      if (R4 > 0) {
        uint32_t hr[16];
        // grouped adapters: this tile's group owns `rank` of the 16 h columns, starting at its offset
        const int hoff = p.group_n > 0 ? (tc.n0 / p.group_n) * p.rank : 0;
        ptx::tmem_ld_32x16(t_main + BN + hoff, hr);
        ptx::tmem_ld_wait();
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
#pragma unroll 1
      for (int ci = 0; ci < CPG; ++ci) {
        const int col0 = grp * HALF_N + ci * CW;
        if (tc.n0 + col0 >= p.N) break;              // ragged last column tile (uniform over the group)
        uint32_t v[CW];
        TmemChunk<CW>::ld(t_main + col0, v);
        const uint32_t dst = out_base + ob * OUT_BYTES + row * OUT_ROW_BYTES;
#pragma unroll
        for (int pc = 0; pc < OUT_ROW_BYTES / 16; ++pc) {
          float o[PW];
#pragma unroll
          for (int k = 0; k < PW; ++k) {
            const int i = col0 + pc * PW + k;
            // explicit ld.shared: the generic pointers derived from the re-aligned base compile to LD.E
            float a = __uint_as_float(v[pc * PW + k]) + ptx::ld_shared_f32(bias_s + i * 4);
            if (R4 > 0) {
#pragma unroll
              for (int j4 = 0; j4 < R4; ++j4) {
                const float4 b = ptx::ld_shared_f4(epi_s + i * (16 * R4) + j4 * 16);
                a = fmaf(hv[4 * j4 + 0], b.x, a);
                a = fmaf(hv[4 * j4 + 1], b.y, a);
                a = fmaf(hv[4 * j4 + 2], b.z, a);
                a = fmaf(hv[4 * j4 + 3], b.w, a);
              }
            }
            o[k] = a;
          }
          if (sizeof(T) == 2) {
            uint32_t w[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              __nv_bfloat162 h2 = __floats2bfloat162_rn(o[(2 * k) % PW], o[(2 * k + 1) % PW]);
              w[k] = *reinterpret_cast<uint32_t*>(&h2);
            }
            ptx::st_shared_v4(dst + pc * 16, w[0], w[1], w[2], w[3]);
          } else {
            ptx::st_shared_v4(dst + pc * 16, __float_as_uint(o[0]), __float_as_uint(o[1]), __float_as_uint(o[2 % PW]),
                              __float_as_uint(o[3 % PW]));
          }
        }
        ptx::fence_proxy_async_smem();
        if (gt == 0) ptx::bulk_wait_read<NB - 2>();     // buffer of chunk c+1 (last read by store c+1-NB) is free past the barrier
        SLIDER_GROUP_BAR();
        if (gt == 0) {
          if (p.accumulate) ptx::tma_reduce_add_4d(&tmY, out_base + ob * OUT_BYTES, tc.n0 + col0, tc.w0, tc.h0, tc.b0);
          else ptx::tma_store_4d(&tmY, out_base + ob * OUT_BYTES, tc.n0 + col0, tc.w0, tc.h0, tc.b0);
          ptx::bulk_commit();
        }
        if (++ob == (uint32_t)NB) ob = 0;
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (et == 0 && unit == (int)blockIdx.x) SLIDER_TRACE(6);
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(tempty_bar + as));
      if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
    }
    if (gt == 0) ptx::bulk_wait_all();    // all TMA stores of this group have landed before the CTA exits
#undef SLIDER_GROUP_BAR
  }

  if (p.cluster && warp < 2) {
    // cluster split-K: every thread of the cluster takes part in the two cluster barriers of the epilogue
    // (warps 0/1, whose elected lanes have finished their single unit by now, have nothing to publish)
    ptx::cluster_arrive(); ptx::cluster_wait();
    ptx::cluster_arrive(); ptx::cluster_wait();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
  if (threadIdx.x == 0) SLIDER_TRACE(7);
#undef SLIDER_TRACE
}


// ---------------------------------------------------------------------------------------------
// Split-K finalize: y = sum_splits(partial) + bias + s * h . Bup^T.  One thread per (row, 4-column piece, chunk c)
// also owns the same piece of chunk c + 4, so the row's h partials are read once per 8 output columns.
// Thread order (piece, row, chunk) follows the partial layout, so reads are fully coalesced.
template <typename T, int R4>
__global__ void __launch_bounds__(256)
splitk_finalize_kernel(const GemmParams p) {
  pdl_trigger();
  constexpr int PIECES = PCW / 4;                  // 16-byte pieces per partial row
  constexpr int HC = (NCHUNK - 1) / 2;             // chunk pairs (c, c + HC)
  constexpr int PER_TILE = BM * PIECES * HC;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)p.total_tiles * PER_TILE) return;
  const int tile = (int)(idx / PER_TILE);
  const int rem = (int)(idx - (long long)tile * PER_TILE);
  const int pc = rem % PIECES, row = (rem / PIECES) & (BM - 1), c = rem / (PIECES * BM);
  const TileCoord tc = tile_coord(p, tile);
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
  const bool packb = kPackB && p.rank == 4 && Bup != nullptr;
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
          l = fmaf(hv[0], __uint_as_float(q.x << 16), l);
          l = fmaf(hv[1 % (R4 > 0 ? 4 * R4 : 1)], __uint_as_float(q.x & 0xffff0000u), l);
          l = fmaf(hv[2 % (R4 > 0 ? 4 * R4 : 1)], __uint_as_float(q.y << 16), l);
          l = fmaf(hv[3 % (R4 > 0 ? 4 * R4 : 1)], __uint_as_float(q.y & 0xffff0000u), l);
        } else {
#pragma unroll
          for (int j = 0; j < 4 * R4; ++j)
            if (j < p.rank) l = fmaf(hv[j], Elem<T>::to_f(Bup[(size_t)(n + k) * p.rank + j]), l);
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
        const float res[4] = {__uint_as_float(q.x << 16), __uint_as_float(q.x & 0xffff0000u),
                              __uint_as_float(q.y << 16), __uint_as_float(q.y & 0xffff0000u)};
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = __bfloat162float(__float2bfloat16_rn(r[k])) + res[k];
      }
      __nv_bfloat162 lo2 = __floats2bfloat162_rn(r[0], r[1]), hi2 = __floats2bfloat162_rn(r[2], r[3]);
      *d2 = make_uint2(*reinterpret_cast<uint32_t*>(&lo2), *reinterpret_cast<uint32_t*>(&hi2));
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
// 1 = reduce split-K partials inside a thread-block cluster over distributed shared memory (bulk DMA pushes), 0 =
// workspace + finalize kernel.  Opt-in (SLIDER_SPLITK_CLUSTER=1 / slider_debug_splitk_cluster): measured slower on
// B200 -- DSMEM moves ~17 B/clk per SM, the 70-80 KB of partial rows a CTA exchanges cost ~5 K cycles against ~1.5 K
// through L2 -- see profiles/r01_ab_splitk_cluster.txt.
// K blocks of L2 look-ahead on the weight stream (SLIDER_W_PREFETCH, 0 = off)
static int g_w_prefetch = [] { const char* e = getenv("SLIDER_W_PREFETCH"); return e ? atoi(e) : 12; }();
static int g_splitk_cluster = [] { const char* e = getenv("SLIDER_SPLITK_CLUSTER"); return (e && e[0] == '1') ? 1 : 0; }();

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

static inline int esize(int dtype) { return dtype == SLIDER_BF16 ? 2 : 4; }

template <typename T>
static int launch_gemm(int r4, const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& l,
                       const CUtensorMap& xe, const CUtensorMap& we, const CUtensorMap& my, const CUtensorMap& mp,
                       const GemmParams& p, cudaStream_t st) {
  void (*kern)(CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap,
               GemmParams) = nullptr;
  void (*fin)(GemmParams) = nullptr;
  switch (r4) {
    case 0: kern = lora_gemm_kernel<T, 0>; fin = splitk_finalize_kernel<T, 0>; break;
    case 1: kern = lora_gemm_kernel<T, 1>; fin = splitk_finalize_kernel<T, 1>; break;
    case 2: kern = lora_gemm_kernel<T, 2>; fin = splitk_finalize_kernel<T, 2>; break;
    default: kern = lora_gemm_kernel<T, 4>; fin = splitk_finalize_kernel<T, 4>; break;
  }
  static thread_local const void* configured[8] = {nullptr};
  bool done = false;
  for (auto c : configured) done |= (c == (const void*)kern);
  if (!done) {
    SLIDER_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem_bytes(r4)));
    for (auto& c : configured) if (c == nullptr) { c = (const void*)kern; break; }
  }
  int grid = p.total_units < sm_count() ? p.total_units : sm_count();
  if (p.cluster) {
    // one unit per CTA, the ksplit CTAs of a tile form one cluster (cluster rank == split)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.total_units); cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = gemm_smem_bytes(r4); cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.ksplit; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaLaunchKernelEx(&cfg, kern, a, w, l, xe, we, my, mp, p);
    return post_launch("lora_gemm_kernel (cluster split-K)");
  }
  launch_pdl(kern, dim3(grid), dim3(kGemmThreads), gemm_smem_bytes(r4), st, a, w, l, xe, we, my, mp, p);
  int rc = post_launch("lora_gemm_kernel");
  if (rc || p.ksplit == 1) return rc;
  const long long threads = (long long)p.total_tiles * BM * (BN / 8);
  launch_pdl(fin, dim3((unsigned)((threads + 255) / 256)), dim3(256), 0, st, p);
  return post_launch("splitk_finalize_kernel");
}

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
//   act      [B,H,W,kc]      activations (x or gy)
//   wgt      [nn, taps, kc]  weights (W or Wt)
//   lo       [rank, taps, kc] lora_down (forward only)
//   xe / we  [B,H,W,16] / [nn, taps, 16]  extended-K pair (input gradient only)
static int run_gemm(const slider_layer_geom* g, int kc, int nn, void* y, const void* act, const void* wgt,
                    const void* bias, const void* lo, const void* Bup, float* h_out, float s, const void* xe,
                    const void* we, void* ws, size_t ws_bytes, cudaStream_t st, int groups = 1, int accumulate = 0) {
  const int es = esize(g->dtype);
  const int taps = g->kh * g->kw;
  const int bk = ROW_BYTES / es;
  SLIDER_CHECK_ARG(g->dtype == SLIDER_BF16 || g->dtype == SLIDER_F32, "lora gemm: unsupported dtype %d", g->dtype);
  SLIDER_CHECK_ARG(g->kh == g->kw && (g->kh == 1 || g->kh == 3) && g->pad == (g->kh - 1) / 2,
                   "lora gemm: only 1x1/pad0 and 3x3/pad1 stride-1 kernels (got %dx%d pad %d)", g->kh, g->kw, g->pad);
  SLIDER_CHECK_ARG(g->batch > 0 && g->height > 0 && g->width > 0 && kc > 0 && nn > 0, "lora gemm: empty shape");
  SLIDER_CHECK_ARG(kc % 8 == 0 && nn % 8 == 0,
                   "lora gemm: channel counts must be multiples of 8 (cin %d, cout %d)", g->cin, g->cout);
  SLIDER_CHECK_ARG(g->rank >= 0 && g->rank <= 16, "lora gemm: rank %d not in 0..16", g->rank);
  SLIDER_CHECK_ARG(((uintptr_t)act & 15) == 0 && ((uintptr_t)wgt & 15) == 0 && ((uintptr_t)y & 15) == 0,
                   "lora gemm: pointers must be 16-byte aligned");

  const bool lora = lo != nullptr && g->rank > 0 && (s != 0.f || h_out != nullptr);
  const bool ext = xe != nullptr;
  if (groups > 1) {
    SLIDER_CHECK_ARG(lora && !ext && taps == 1 && h_out == nullptr, "grouped lora gemm: needs an active 1x1 adapter, no h_out");
    SLIDER_CHECK_ARG(groups * g->rank <= HCOLS, "grouped lora gemm: groups * rank = %d exceeds %d", groups * g->rank, HCOLS);
    SLIDER_CHECK_ARG(nn % groups == 0 && (nn / groups) % BN == 0,
                     "grouped lora gemm: columns per group (%d / %d) must be a multiple of %d", nn, groups, BN);
    SLIDER_CHECK_ARG((long long)g->batch * g->height * g->width > 16, "grouped lora gemm: needs more than 16 rows");
  }
  if (lora) SLIDER_CHECK_ARG(Bup != nullptr, "lora gemm: lora_up weight is NULL");
  {
    // small-M path: a 128-row tensor-core tile would be > 87 % padding and the op is weight-bandwidth bound
    const long long M = (long long)g->batch * g->height * g->width;
    const size_t mt = M <= 2 ? 2 : (M <= 4 ? 4 : (M <= 8 ? 8 : 16));
    if (groups == 1 && !accumulate && !ext && taps == 1 && M <= 16 && mt * kc * sizeof(float) <= 40 * 1024 && kc * es <= 8 * 32 * 16 &&
        ((uintptr_t)wgt & 15) == 0 && (lo == nullptr || ((uintptr_t)lo & 15) == 0)) {
      if (g->dtype == SLIDER_BF16)
        return launch_gemv<__nv_bfloat16>(g, (int)M, y, act, wgt, bias, lo, Bup, lora ? h_out : nullptr, s, lora, st);
      return launch_gemv<float>(g, (int)M, y, act, wgt, bias, lo, Bup, lora ? h_out : nullptr, s, lora, st);
    }
  }
  const TileGeom tg = choose_tile(g->width, g->height, g->batch);

  GemmParams p;
  p.TW = tg.TW; p.TH = tg.TH; p.TB = tg.TB; p.rows = tg.TW * tg.TH * tg.TB;
  p.W = g->width; p.H = g->height; p.B = g->batch;
  p.mt_w = (g->width + tg.TW - 1) / tg.TW;
  p.mt_h = (g->height + tg.TH - 1) / tg.TH;
  p.mt_b = (g->batch + tg.TB - 1) / tg.TB;
  p.n_tiles = (nn + BN - 1) / BN; p.N = nn;
  p.taps = taps; p.kw = g->kw; p.pad = g->pad; p.cblocks = (kc + bk - 1) / bk;
  p.rank = g->rank; p.s = s; p.lora = lora ? 1 : 0; p.ext = ext ? 1 : 0;
  p.group_n = groups > 1 ? nn / groups : 0;
  p.accumulate = accumulate ? 1 : 0;
  p.total_tiles = p.mt_w * p.mt_h * p.mt_b * p.n_tiles;
  p.per_tap = p.cblocks + (ext ? 1 : 0);
  p.kiters = taps * p.per_tap;
  // split-K when the tile count leaves more than half of the SMs idle and K is deep enough
  p.ksplit = 1;
  p.cluster = 0;
  p.partial = reinterpret_cast<float*>(ws);
  if (p.total_tiles * 2 <= sm_count()) {
    int ks = sm_count() / p.total_tiles;
    // A split costs a second launch plus the partial round trip (~6 us at 32+ tiles, ~2.5 us at <= 12): measured on
    // the SD-1.4 layer set (profiles/r01_splitk_depth_sweep.txt) it only pays for deep K -- 64+ K blocks at >= 18 per
    // split, or 40+ at >= 6 per split for the weight-streaming shapes with a handful of tiles.
    int depth = g_splitk_depth, min_k = 0;
    if (depth == 0) {
      depth = p.total_tiles <= 12 ? 6 : 18;
      min_k = p.total_tiles <= 12 ? 40 : 64;
    }
    const int by_depth = p.kiters >= min_k ? p.kiters / depth : 1;
    if (ks > by_depth) ks = by_depth;
    if (g_splitk_cluster) {
      // in-cluster reduction over distributed shared memory: no workspace, no second launch
      if (ks > MAX_CLUSTER) ks = MAX_CLUSTER;
      if (ks >= 2) { p.ksplit = ks; p.cluster = 1; }
    } else if (ws != nullptr && ((uintptr_t)ws & 15) == 0) {
      const long long cap = (long long)(ws_bytes / (UNIT_FLOATS * sizeof(float)));
      if ((long long)ks * p.total_tiles > cap) ks = (int)(cap / p.total_tiles);
      if (ks >= 2) p.ksplit = ks;
    }
  }
  p.total_units = p.total_tiles * p.ksplit;
  p.fd_n_tiles = FastDiv::make((uint32_t)p.n_tiles);
  p.fd_mt_w = FastDiv::make((uint32_t)p.mt_w);
  p.fd_mt_h = FastDiv::make((uint32_t)p.mt_h);
  p.fd_ksplit = FastDiv::make((uint32_t)p.ksplit);
  p.fd_per_tap = FastDiv::make((uint32_t)p.per_tap);
  p.fd_kw = FastDiv::make((uint32_t)p.kw);
  p.trace = g_gemm_trace;
  // only when several pixel tiles share each weight tile: with ONE pixel tile every CTA streams distinct weights and
  // the pipeline's own requests already fill HBM
  p.w_prefetch = (p.mt_w * p.mt_h * p.mt_b > 1) ? g_w_prefetch : 0;
  p.y = y; p.bias = bias; p.Bup = Bup; p.h_out = lora ? h_out : nullptr;
  if (lora) {
    SLIDER_CHECK_ARG((g->rank * taps * kc * es) % 16 == 0 && ((uintptr_t)lo & 15) == 0, "lora gemm: lora_down misaligned");
  }

  CUtensorMap mA, mW, mL, mXe, mWe;
  int rc;
  {
    uint64_t dims[4] = {(uint64_t)kc, (uint64_t)g->width, (uint64_t)g->height, (uint64_t)g->batch};
    uint64_t str[3] = {(uint64_t)kc * es, (uint64_t)kc * es * g->width, (uint64_t)kc * es * g->width * g->height};
    uint32_t box[4] = {(uint32_t)bk, (uint32_t)tg.TW, (uint32_t)tg.TH, (uint32_t)tg.TB};
    if ((rc = make_tmap(&mA, g->dtype, 4, act, dims, str, box, ROW_BYTES))) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)kc, (uint64_t)taps, (uint64_t)nn};
    uint64_t str[2] = {(uint64_t)kc * es, (uint64_t)kc * es * taps};
    uint32_t box[3] = {(uint32_t)bk, 1u, (uint32_t)BN};
    if ((rc = make_tmap(&mW, g->dtype, 3, wgt, dims, str, box, ROW_BYTES))) return rc;
  }
  mL = mW; mXe = mA; mWe = mW;
  if (lora) {
    uint64_t dims[3] = {(uint64_t)kc, (uint64_t)taps, (uint64_t)(g->rank * groups)};
    uint64_t str[2] = {(uint64_t)kc * es, (uint64_t)kc * es * taps};
    uint32_t box[3] = {(uint32_t)bk, 1u, (uint32_t)HCOLS};
    if ((rc = make_tmap(&mL, g->dtype, 3, lo, dims, str, box, ROW_BYTES))) return rc;
  }
  if (ext) {
    SLIDER_CHECK_ARG(we != nullptr && ((uintptr_t)xe & 15) == 0 && ((uintptr_t)we & 15) == 0, "lora gemm: bad ext operands");
    const uint32_t eb = EXT_ELEMS * es;
    uint64_t dims[4] = {(uint64_t)EXT_ELEMS, (uint64_t)g->width, (uint64_t)g->height, (uint64_t)g->batch};
    uint64_t str[3] = {(uint64_t)eb, (uint64_t)eb * g->width, (uint64_t)eb * g->width * g->height};
    uint32_t box[4] = {(uint32_t)EXT_ELEMS, (uint32_t)tg.TW, (uint32_t)tg.TH, (uint32_t)tg.TB};
    if ((rc = make_tmap(&mXe, g->dtype, 4, xe, dims, str, box, eb))) return rc;
    uint64_t wdims[3] = {(uint64_t)EXT_ELEMS, (uint64_t)taps, (uint64_t)nn};
    uint64_t wstr[2] = {(uint64_t)eb, (uint64_t)eb * taps};
    uint32_t wbox[3] = {(uint32_t)EXT_ELEMS, 1u, (uint32_t)BN};
    if ((rc = make_tmap(&mWe, g->dtype, 3, we, wdims, wstr, wbox, eb))) return rc;
  }
  CUtensorMap mY, mP;
  {
    uint64_t dims[4] = {(uint64_t)nn, (uint64_t)g->width, (uint64_t)g->height, (uint64_t)g->batch};
    uint64_t str[3] = {(uint64_t)nn * es, (uint64_t)nn * es * g->width, (uint64_t)nn * es * g->width * g->height};
    uint32_t box[4] = {(uint32_t)(OUT_ROW_BYTES / es), (uint32_t)tg.TW, (uint32_t)tg.TH, (uint32_t)tg.TB};
    if ((rc = make_tmap(&mY, g->dtype, 4, y, dims, str, box, OUT_ROW_BYTES))) return rc;
  }
  mP = mY;
  if (p.ksplit > 1 && !p.cluster) {
    uint64_t dims[2] = {(uint64_t)PCW, (uint64_t)p.total_units * NCHUNK * BM};
    uint64_t str[1] = {(uint64_t)OUT_ROW_BYTES};
    uint32_t box[2] = {(uint32_t)PCW, (uint32_t)BM};
    if ((rc = make_tmap(&mP, SLIDER_F32, 2, ws, dims, str, box, OUT_ROW_BYTES))) return rc;
  }
  const int r4 = !lora ? 0 : (g->rank <= 4 ? 1 : (g->rank <= 8 ? 2 : 4));
  if (g->dtype == SLIDER_BF16) return launch_gemm<__nv_bfloat16>(r4, mA, mW, mL, mXe, mWe, mY, mP, p, st);
  return launch_gemm<float>(r4, mA, mW, mL, mXe, mWe, mY, mP, p, st);
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

}  // namespace slider

using namespace slider;

extern "C" {

int slider_debug_splitk_depth(int d) { g_splitk_depth = d > 0 ? d : 0; return SLIDER_OK; }
int slider_debug_splitk_cluster(int on) { g_splitk_cluster = on ? 1 : 0; return SLIDER_OK; }
int slider_debug_gemm_trace(void* buf) { g_gemm_trace = reinterpret_cast<long long*>(buf); return SLIDER_OK; }

size_t slider_gemm_workspace_bytes(void) { return (size_t)MAX_UNITS * UNIT_FLOATS * sizeof(float); }

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
  SLIDER_CHECK_ARG(dtype == SLIDER_BF16 || dtype == SLIDER_F32, "lora_forward_small_multi: unsupported dtype %d", dtype);
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
