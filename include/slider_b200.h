/*
 * slider_b200.h — C ABI of libslider_b200.so (sm_100a only).
 *
 * Drop-in boundary for the concept-slider hot path of ntc-ai/sliders-conceptmod.
 * The reference has no native boundary (it is pure PyTorch); each entry point
 * below names the reference Python call site whose arithmetic it replaces.
 * Paths are relative to the reference checkout.
 *
 * Conventions
 *  - every pointer is a CUDA device pointer into caller-owned memory unless
 *    stated otherwise; the library never allocates or frees caller memory;
 *  - `stream` is a cudaStream_t passed as void*; all calls are asynchronous;
 *  - return 0 on success, non-zero error code otherwise; no C++ exception and
 *    no exit() crosses the boundary; slider_last_error() returns thread-local
 *    text describing the last failure on the calling thread;
 *  - activations of conv layers are NHWC ("channels_last") in memory;
 *    linear layers are row-major [M, K].
 */
#ifndef SLIDER_B200_H
#define SLIDER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SLIDER_ABI_VERSION 3

/* SLIDER_F16: forward / inference entry points only (the reference's inference scripts run in fp16,
 * eval-scripts/generate_images_xl.py:443); the training entry points return SLIDER_ERR_UNSUPPORTED for it. */
enum slider_dtype { SLIDER_F32 = 0, SLIDER_BF16 = 1, SLIDER_F16 = 2 };

enum slider_status {
  SLIDER_OK = 0,
  SLIDER_ERR_BAD_ARG = 1,     /* null pointer, bad shape, misalignment        */
  SLIDER_ERR_UNSUPPORTED = 2, /* dtype / rank / geometry not implemented      */
  SLIDER_ERR_CUDA = 3,        /* a CUDA runtime / driver call failed          */
  SLIDER_ERR_WORKSPACE = 4    /* caller workspace too small                   */
};

int slider_version(void);
const char* slider_last_error(void);
/* Number of kernel launches issued by this library since load (all threads). */
int64_t slider_launch_count(void);

/* ---------------------------------------------------------------------------
 * Layer geometry shared by the LoRA entry points.
 * A Linear layer is the degenerate conv: batch=1, height=1, width=M, kh=kw=1.
 * Symmetric padding; stride 1, or stride 2 for a 3x3 kernel on the forward entry points (the three Downsample2D
 * convolutions, lora.py:100-114 builds lora_down with the base conv's stride): height / width are the INPUT extent,
 * the output is [batch, (height-1)/2+1, (width-1)/2+1, cout] and h_out has one row per OUTPUT pixel.
 * ------------------------------------------------------------------------- */
typedef struct slider_layer_geom {
  int32_t dtype;   /* enum slider_dtype of x / W / A / B / y                 */
  int32_t batch;   /* NHWC batch                                             */
  int32_t height;
  int32_t width;
  int32_t cin;     /* K per tap                                              */
  int32_t cout;    /* N                                                      */
  int32_t kh, kw;  /* 1x1 or 3x3                                             */
  int32_t pad;     /* 0 for 1x1, 1 for 3x3                                   */
  int32_t rank;    /* LoRA rank r, 0..16                                     */
  int32_t stride;  /* 0 or 1: stride 1;  2: stride 2 (3x3, forward only)     */
} slider_layer_geom;

/*
 * Fused base + low-rank forward.  Replaces LoRAModule.forward
 * (conceptmod/textsliders/lora.py:134-138, trainscripts/imagesliders/lora.py:107-111):
 *     y = conv(x, W) + bias + s * conv1x1(conv(x, A), B),   s = multiplier*alpha/rank
 *
 *   x     [batch,height,width,cin]                     activations (NHWC)
 *   W     [cout, kh*kw, cin]   tap-major ("KRSC")      frozen base weight
 *   bias  [cout] or NULL
 *   A     [rank, kh*kw, cin]   tap-major               lora_down.weight
 *   B     [cout, rank]                                 lora_up.weight
 *   h_out [batch*height*width, rank] fp32 or NULL      x (*) A, saved for backward
 *   y     [batch,height,width,cout]
 * s == 0 and h_out == NULL selects the plain-GEMM path (adapter switched off,
 * lora.py:299-301).  One tcgen05/TMA kernel launch; M <= 16 rows of a 1x1 layer
 * (time_emb_proj) take a weight-streaming SIMT kernel instead.
 *   ws    optional scratch of slider_gemm_workspace_bytes() bytes (16-byte
 *         aligned, contents don't-care).  When given, tile-starved shapes
 *         (few output tiles, deep K) run split-K: partial tiles are parked in
 *         ws and a second small launch reduces them (deterministic order).
 *         NULL => never split.  ws may be shared by all calls on one stream.
 */
size_t slider_gemm_workspace_bytes(void);
int slider_lora_forward(const slider_layer_geom* g, void* y, const void* x,
                        const void* W, const void* bias, const void* A,
                        const void* B, float* h_out, float s, void* ws,
                        size_t ws_bytes, void* stream);

/*
 * slider_lora_forward with option flags.
 *   SLIDER_FWD_ACCUMULATE: y already holds the residual branch and the layer result is ADDED to it by the store
 *   (TMA reduce-add): `x = x + attn.to_out(...)`, `hidden = conv_shortcut(x) + conv2(h)`, `h + res` around the
 *   wrapped projections of the UNet blocks (diffusers forward reached from train_util.py:290-294) without a
 *   separate elementwise kernel.  Rounding equals `residual + module(x)` evaluated in the element dtype.
 */
#define SLIDER_FWD_ACCUMULATE 1u
int slider_lora_forward_ex(const slider_layer_geom* g, void* y, const void* x,
                           const void* W, const void* bias, const void* A,
                           const void* B, float* h_out, float s, unsigned flags,
                           void* ws, size_t ws_bytes, void* stream);

/*
 * Feed-forward projection with its GEGLU gate fused into the store (diffusers GEGLU.forward reached from the UNet call
 * at train_util.py:290-294: `hidden, gate = proj(x).chunk(2, dim=-1); hidden * gelu(gate)`; frozen, never LoRA-wrapped):
 *     y[p, c] = (x . W[c]^T + b[c]) * gelu(x . W[d + c]^T + b[d + c]),   c < d,   g->cout = 2 d
 * One launch of the tcgen05 kernel; the 2d-wide intermediate never reaches memory.  W_perm / bias_perm are the
 * projection's rows re-ordered ONCE by the caller so that every 160-row block holds [a(40) | g(40) | a(40) | g(40)]:
 *     perm row 160 t + 80 h + 40 s + i  =  original row  s d + 80 t + 40 h + i     (s = 0 value, 1 gate; h = 0, 1; i < 40)
 * d must be a multiple of 80, M > 16 rows; y [M, d].  Exact (erf) GELU in fp32, the bf16-grade erf of slider_geglu
 * for bf16 tensors.
 */
int slider_linear_geglu(const slider_layer_geom* g, void* y, const void* x, const void* W_perm,
                        const void* bias_perm, void* stream);

/*
 * Several adapters on ONE input in one launch: the self-attention to_q / to_k / to_v of a transformer block
 * (three LoRAModule.forward calls on the same x, reference diffusers Attention.forward reached from
 * conceptmod/textsliders/train_util.py:290-294).  Operands are the row-wise concatenations
 *   W [groups*cout_g, cin]   A [groups*rank, cin]   B [groups*cout_g, rank]   y [M, groups*cout_g]
 * with g->cout = groups*cout_g and g->rank = the per-adapter rank; output columns of group i get
 * s * (x . A_i^T) . B_i^T.  Needs groups*rank <= 16, cout_g a multiple of 160, kh = kw = 1, M > 16, s != 0.
 * No h_out: inference / no-grad forwards only (the grad forward uses slider_lora_forward per adapter).
 */
int slider_lora_forward_grouped(const slider_layer_geom* g, void* y, const void* x,
                                const void* W, const void* bias, const void* A,
                                const void* B, int groups, float s, void* ws,
                                size_t ws_bytes, void* stream);

/*
 * Several SMALL-M Linear adapters on ONE input in one launch: the `time_emb_proj` of every ResnetBlock2D projects the
 * same silu(temb) rows (M = 2b <= 16; each one a LoRAModule.forward, conceptmod/textsliders/lora.py:94-98,134-138,
 * reached from the UNet call at train_util.py:290-294).  Host arrays of `layers` (<= 32) device pointers:
 *   x [M, K]   W[l] [N[l], K]   bias[l] [N[l]] or NULL   A[l] [rank, K] or NULL   B[l] [N[l], rank]   y[l] [M, N[l]]
 *   y[l] = x . W[l]^T + bias[l] + s[l] * (x . A[l]^T) . B[l]^T        (s[l] == 0 or A[l] == NULL: plain projection)
 * `bias`, `A`, `B` may be NULL as a whole.  Weight-streaming (HBM-bound) kernel, no tensor cores; no h_out:
 * inference / no-grad forwards only (the grad forward uses slider_lora_forward per adapter).
 */
int slider_lora_forward_small_multi(int dtype, int layers, int M, int K, int rank, const void* x,
                                    void* const* y, const void* const* W, const void* const* bias,
                                    const void* const* A, const void* const* B, const int* N,
                                    const float* s, void* stream);

/*
 * Low-rank part of the backward of the layer above (autograd of lora.py:134-138,
 * triggered by loss.backward() at conceptmod/textsliders/train_lora.py:298):
 *     dh = s * gy . B          [M, rank]  -> dh_out (element dtype, padded to 16)
 *     dB = s * gy^T . h        [cout, rank]
 *     dA = dh^T (*) x          [rank, kh*kw, cin]
 *   gy     [batch, out_h, out_w, cout]  upstream gradient (OUTPUT extent; stride 2 is accepted here)
 *   x      [batch, height, width, cin]
 *   h      [M, rank] fp32 saved by the forward (M = output pixels)
 *   dA,dB  fp32; `accumulate` != 0 adds to their current content (flat gradient buffer; two-sided image-slider
 *          objective, train_lora-scale.py:303-332), 0 overwrites.  Both NULL: only dh is produced (adapter
 *          frozen, the input gradient still needs it).
 *   dh_out [M, 16] element dtype, columns >= rank are zero; consumed by
 *          slider_lora_input_grad
 *   ws     caller workspace of slider_lora_grad_workspace_bytes(g) bytes, 256-byte aligned; its first 4096
 *          bytes are zero-filled ONCE by the caller (ticket words; the kernels leave them zero), the rest is
 *          don't-care.
 * The reductions over pixels are deterministic (per-CTA partial sums, finished in a fixed order by the last
 * CTA of a channel block; no floating-point atomics).
 */
size_t slider_lora_grad_workspace_bytes(const slider_layer_geom* g);
int slider_lora_grad(const slider_layer_geom* g, float* dA, float* dB,
                     void* dh_out, const void* gy, const void* x,
                     const float* h, const void* B, float s, int accumulate,
                     void* ws, size_t ws_bytes, void* stream);

/* slider_lora_grad with a per-input-channel factor on dA: dA[j,t,c] *= dA_chan_scale[c] (fp32 [cin], NULL = none).
 * DoRA (conceptmod/textsliders/dora.py:122-163): the low-rank term of W' is scaled per input channel by s*c_c. */
int slider_lora_grad_ex(const slider_layer_geom* g, float* dA, float* dB,
                        void* dh_out, const void* gy, const void* x,
                        const float* h, const void* B, float s, int accumulate,
                        const float* dA_chan_scale, void* ws, size_t ws_bytes, void* stream);

/*
 * Whole backward of a 1x1 / Linear adapter layer in two launches (same autograd as above):
 *   launch 1: the forward tcgen05 kernel on the transposed problem -- activations gy, weights Wt, "lora_down" = B^T,
 *             "lora_up" = A read as its own transpose -- gives  dx = gy . W + s * (gy . B) . A  and
 *             g_buf = gy . B  (fp32 [M, rank], unscaled);
 *   launch 2: one reduction kernel gives  dB = s * gy^T . h  and  dA = s * g_buf^T . x  (deterministic).
 *   Wt   [cin, 1, cout]   slider_pack_wt of the frozen base weight
 *   Bt   [>= rank, cout]  B^T, or NULL: packed here from B (one more small launch; slider_transpose_batch builds the
 *                         B^T of every layer of a network in one launch)
 *   A    [rank, cin] lora_down.weight;  B [cout, rank] lora_up.weight (only read when Bt == NULL)
 *   dA, dB NULL: only dx (frozen adapter).  ws as for slider_lora_grad.
 */
int slider_lora_backward_1x1(const slider_layer_geom* g, void* dx, float* dA, float* dB, float* g_buf,
                             const void* gy, const void* x, const float* h, const void* Wt, const void* Bt,
                             const void* A, const void* B, float s, int accumulate, void* ws, size_t ws_bytes,
                             void* stream);

/* Batched small transposes, one launch: entry e writes dst[c][r] = src[r][c] for src [rows, cols].  `table_dev` is a
 * DEVICE array of `entries` records; max_elems = the largest rows*cols (sizes the grid). */
typedef struct slider_transpose_entry {
  const void* src;
  void* dst;
  int32_t rows, cols;
} slider_transpose_entry;
int slider_transpose_batch(const slider_transpose_entry* table_dev, int entries, int max_elems, int dtype,
                           void* stream);

/*
 * Input gradient:  dx = convT(gy, W) + convT(dh, A)   (one tcgen05 launch).
 *   Wt  [cin, kh*kw, cout]  frozen, taps flipped: Wt[c][t][n] = W[n][T-1-t][c]
 *   At  [cin, kh*kw, 16]    At[c][t][j] = A[j][T-1-t][c], j >= rank zero;
 *                           produced by slider_lora_pack_at
 *   dh  [M, 16]             from slider_lora_grad (NULL => base term only)
 *   ws  as in slider_lora_forward
 */
int slider_lora_pack_at(const slider_layer_geom* g, void* At, const void* A,
                        void* stream);
int slider_lora_input_grad(const slider_layer_geom* g, void* dx, const void* gy,
                           const void* Wt, const void* dh, const void* At,
                           void* ws, size_t ws_bytes, void* stream);

/*
 * DoRA on the fused kernel.  Replaces DoRAModule.apply_weight_decompose + forward
 * (conceptmod/textsliders/dora.py:122-163; selected by train_lora_xl.py:87-90 --peft_type dora):
 *     W' = (W + B.A) * (m / ||W + B.A||_per input channel, detached),   y = org(x) + s * ops(x, W' - W)
 * which, with c_i = m_i / n_i and v = 1 - s + s*c, equals  x . (W*v)^T + bias + (x . (A * s*c)^T) . B^T :
 * slider_lora_forward on W_eff = W*v and A_s = A*(s c) with adapter scale 1.
 *   slider_dora_prepare: two passes over the frozen weight (column norms; scaling) ->
 *       W_eff [cout, kh*kw, cin], A_s [rank, kh*kw, cin] (element dtype), c, v, ninv = 1/n (fp32 [cin]).
 *       W, A tap-major as in slider_lora_forward; dora_scale m [cin] (element dtype).
 *   slider_dora_finish: tail of the backward.  u = gy.W + (gy.B).A is the input-gradient GEMM of the LoRA path with
 *       s = 1 (slider_lora_backward_1x1 / slider_lora_input_grad on the UNSCALED W and A); then
 *           dx[p,i] = v_i * T1 + s c_i * T2        (T1 = gy.W, T2 = u - T1)
 *           dm[i]   = s/n_i * sum_p x[p,i] * u[p,i]          (gradient of dora_scale; fp32 [cin])
 *       t1 [M, cin] = gy.W from a second slider_lora_input_grad call (3x3), or NULL for a 1x1 layer, where
 *       T2 = g_buf . A is formed here from g_buf = gy.B [M, rank] fp32 (slider_lora_backward_1x1's output).
 *       dx or dm may be NULL.  geometry: batch/height/width = extent of x (the layer INPUT).
 *   The low-rank gradients come from slider_lora_grad_ex(s = 1, h = the forward's h_out, dA_chan_scale = s*c).
 *   ws: slider_lora_grad_workspace_bytes(g) bytes, first 4096 zero-filled once (as for slider_lora_grad).
 * The reference's dense weight gradient (out x in per wrapped layer) is never formed.
 */
int slider_dora_prepare(const slider_layer_geom* g, void* W_eff, void* A_s, float* c, float* v, float* ninv,
                        const void* W, const void* A, const void* B, const void* dora_scale, float s, void* ws,
                        size_t ws_bytes, void* stream);
int slider_dora_finish(const slider_layer_geom* g, void* dx, float* dm, const void* u, const void* t1, const void* x,
                       const float* g_buf, const void* A, const float* c, const float* v, const float* ninv, float s,
                       int accumulate, void* ws, size_t ws_bytes, void* stream);

/* Profiling aid: when buf != NULL, every lora_gemm_kernel CTA writes 8 clock64()
 * stamps (entry, setup done, first TMA issued, first operands landed, last MMA
 * committed, accumulator visible to the epilogue, epilogue done, exit) to
 * buf[cta*8 ..]; buf must hold 148*8 int64.  NULL switches it off. */
int slider_debug_gemm_trace(void* buf);

/* One-time repack of a frozen base weight [cout, kh*kw, cin] into the Wt layout. */
int slider_pack_wt(const slider_layer_geom* g, void* Wt, const void* W,
                   void* stream);

/*
 * Slider guidance objective, forward and gradient in one pass.  Replaces
 * PromptEmbedsPair._enhance/_erase + MSELoss
 * (trainscripts/imagesliders/prompt_util.py:108-148,
 *  conceptmod/textsliders/prompt_util.py:134-174) and its autograd:
 *     tgt  = n + sign*g*(p - q);  loss = mean((t - tgt)^2);
 *     grad = 2*(t - tgt)/n_elem            (grad may be NULL)
 * sign = +1 enhance, -1 erase.  p == NULL selects plain MSE(t, n)
 * (image sliders, trainscripts/imagesliders/train_lora-scale.py:303,327).
 * loss_out: one fp32.  ws: slider_loss_workspace_bytes() bytes, zero-filled by
 * the caller once; the kernel leaves it zeroed.  Deterministic reduction.
 */
size_t slider_loss_workspace_bytes(void);
int slider_guidance_loss(float* loss_out, void* grad, const void* t,
                         const void* p, const void* n, const void* q,
                         int64_t n_elem, float g, int sign, int dtype, void* ws,
                         size_t ws_bytes, void* stream);

/*
 * Allocation-free variant of slider_guidance_loss.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid allocation).  Same arithmetic, same kernel and same results as slider_guidance_loss; the
 * reduction scratch (partial sums + ticket, 4752 bytes) is a file-scope `static __device__` array inside the
 * library instead of a caller workspace, so the call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine) -- it issues one 16-byte cudaMemsetAsync (the ticket word) and one kernel launch on `stream`.
 *   limit   n_elem <= SLIDER_NOALLOC_MAX_ELEMS (2^28 elements, the largest size of the bandwidth sweep); larger
 *           sizes return SLIDER_ERR_UNSUPPORTED with a message naming the limit
 *           (slider_guidance_loss_noalloc_max_elems() returns it);
 *   memory  every GPU that loads the library holds the 4752-byte array until exit;
 *   sharing ALL calls on one device share that array and it keeps the last call's contents (the entry point clears
 *           what must start at zero): calls that overlap on the GPU (different streams, graph branches, host threads
 *           without stream ordering between them) RACE -- serialise them on one stream or use slider_guidance_loss
 *           with separate workspaces.
 * Inputs and outputs are still the caller's (loss_out one fp32, grad like t or NULL).
 */
#define SLIDER_NOALLOC_MAX_ELEMS (1ll << 28)
int64_t slider_guidance_loss_noalloc_max_elems(void);
int slider_guidance_loss_noalloc(float* loss_out, void* grad, const void* t,
                                 const void* p, const void* n, const void* q,
                                 int64_t n_elem, float g, int sign, int dtype,
                                 void* stream);

/*
 * Classifier-free guidance + affine scheduler step.  Replaces
 * train_util.predict_noise tail (conceptmod/textsliders/train_util.py:297-300)
 * and scheduler.step (train_util.py:324; DDIM restated in
 * trainscripts/imagesliders/train_lora-scale.py:31-39):
 *     eps = eps_u + g*(eps_c - eps_u)
 *     out = c_x*x + c_eps*eps + c_noise*noise
 * x == NULL  -> out = eps                (pure CFG combine)
 * eps_u == NULL -> out = c_x*x + c_noise*noise   (scheduler.add_noise)
 * noise may be NULL.  dev_coef, if non-NULL, is a device float[4]
 * {g, c_x, c_eps, c_noise} that overrides the host scalars (CUDA-graph replay).
 */
int slider_cfg_step(void* out, const void* x, const void* eps_u,
                    const void* eps_c, const void* noise, int64_t n_elem,
                    float g, float c_x, float c_eps, float c_noise,
                    const float* dev_coef, int dtype, void* stream);

/*
 * NHWC GroupNorm with optional per-(batch,channel) pre-add and SiLU — the producer of every
 * wrapped convolution's input in the UNet's ResnetBlock2D (the reference
 * delegates to diffusers: norm1 -> SiLU -> conv1, "+ time_emb_proj(temb)" -> norm2 -> SiLU -> conv2):
 *     z = x + chan_add[b, c];  y = act((z - mean_g) * rstd_g * gamma[c] + beta[c])
 *   x, y      [batch, hw, channels]  channels innermost (multiple of 8), element dtype
 *   chan_add  [batch, channels] or NULL;  gamma, beta [channels] or NULL
 *   silu      != 0 applies SiLU
 *   ws        slider_groupnorm_workspace_bytes(batch, groups) bytes, zero-filled ONCE by the caller
 *             (the kernels leave the ticket words zero); may be shared by calls on one stream.
 * One launch per call when a group set fits a thread-block cluster's shared memory (a CTA or a cluster of 2 / 4 / 8
 * per (image, group set): x read once, statistics exchanged over distributed shared memory), otherwise two launches
 * (per-chunk statistics into ws, apply); deterministic reduction order either way.
 * gamma / beta are parameters: with programmatic dependent launch on (default) they are read BEFORE the wait on the
 * preceding kernel, so they must not be the output of a libslider launch issued just before on the same stream.
 */
size_t slider_groupnorm_workspace_bytes(int batch, int groups);
int slider_groupnorm_nhwc(void* y, const void* x, const void* chan_add,
                          const void* gamma, const void* beta, int batch, int hw,
                          int channels, int groups, float eps, int silu, int dtype,
                          void* ws, size_t ws_bytes, void* stream);

/*
 * Training form of the same GroupNorm (the grad-recording forward of train_lora.py:275-289 and its backward at
 * :298).  gamma / beta are frozen (unet.requires_grad_(False), train_lora.py:69), so the backward produces only
 *     dx   [batch, hw, channels]     gradient w.r.t. x
 *     dadd [batch, channels] or NULL gradient w.r.t. chan_add (= sum of dx over pixels; flows into time_emb_proj's
 *                                    LoRA factors)
 * slider_groupnorm_nhwc_train == slider_groupnorm_nhwc + stats_out [batch, groups, 2] fp32 {mean, rstd}, which the
 * backward takes back as `stats`.  Backward = ONE launch on the forward's (image, group set) cluster plan when the
 * slabs of x and gy fit its shared memory (both read once; group sums over distributed shared memory), otherwise two
 * launches (per-chunk partial sums into ws, apply); deterministic order either way;
 * ws of slider_groupnorm_bwd_workspace_bytes(batch, channels, groups) bytes, contents don't-care.
 */
int slider_groupnorm_nhwc_train(void* y, float* stats_out, const void* x, const void* chan_add,
                                const void* gamma, const void* beta, int batch, int hw,
                                int channels, int groups, float eps, int silu, int dtype,
                                void* ws, size_t ws_bytes, void* stream);
size_t slider_groupnorm_bwd_workspace_bytes(int batch, int channels, int groups);
int slider_groupnorm_nhwc_backward(void* dx, void* dadd, const void* gy, const void* x,
                                   const void* chan_add, const void* gamma, const void* beta,
                                   const float* stats, int batch, int hw, int channels, int groups,
                                   int silu, int dtype, void* ws, size_t ws_bytes, void* stream);

/*
 * Row-wise kernels of the UNet's transformer blocks (diffusers BasicTransformerBlock):
 *   slider_layernorm: y = (x - mean) * rstd * gamma + beta over the last dimension of x [rows, channels]
 *                     (the producer of the LoRA-wrapped attn1.to_q/k/v inputs);
 *   slider_geglu:     y[rows, d] = x[:, :d] * gelu(x[:, d:]), x [rows, 2d], exact (erf) GELU.
 * channels and d must be multiples of 8; gamma / beta may be NULL (parameters: read before the dependent-launch wait,
 * see slider_groupnorm_nhwc).
 */
int slider_layernorm(void* y, const void* x, const void* gamma, const void* beta,
                     int64_t rows, int channels, float eps, int dtype, void* stream);
int slider_geglu(void* y, const void* x, int64_t rows, int d, int dtype, void* stream);
/*
 * Training forms (the grad-recording forward of train_lora.py:275-289 and its backward at :298; gamma / beta are frozen,
 * unet.requires_grad_(False) at train_lora.py:69, so only the input gradient exists):
 *   slider_layernorm_train    = slider_layernorm + stats [rows, 2] fp32 {mean, rstd};
 *   slider_layernorm_backward : dx = rstd * (ghat - mean(ghat) - xhat * mean(ghat * xhat)), ghat = gy * gamma;
 *   slider_geglu_backward     : dx[:, :d] = gy * gelu(g), dx[:, d:] = gy * a * gelu'(g) for x = [a | g] (recomputed, no
 *                               saved activation besides x).
 */
int slider_layernorm_train(void* y, float* stats, const void* x, const void* gamma, const void* beta,
                           int64_t rows, int channels, float eps, int dtype, void* stream);
int slider_layernorm_backward(void* dx, const void* gy, const void* x, const void* gamma, const float* stats,
                              int64_t rows, int channels, int dtype, void* stream);
int slider_geglu_backward(void* dx, const void* gy, const void* x, int64_t rows, int d, int dtype, void* stream);

/*
 * Nearest-neighbour 2x upsample of an NHWC tensor: y[b, 2h + i, 2w + j, :] = x[b, h, w, :]  (i, j in {0, 1}).
 * Replaces `F.interpolate(hidden_states, scale_factor=2.0, mode="nearest")` of diffusers' Upsample2D ahead of its
 * LoRA-wrapped 3x3 convolution (UNet calls at train_util.py:290-294).  x [batch, height, width, channels],
 * y [batch, 2*height, 2*width, channels]; channels * sizeof(dtype) must be a multiple of 16.
 */
int slider_upsample_nearest2x_nhwc(void* y, const void* x, int batch, int height, int width, int channels, int dtype,
                                   void* stream);

/*
 * Channel concatenation of two NHWC tensors: y[p, :ca] = a[p, :], y[p, ca:] = b[p, :] for every pixel p.
 * Replaces `torch.cat([hidden_states, res_hidden_states], dim=1)` of diffusers' up blocks (the skip connections; UNet
 * calls at train_util.py:290-294).  channels_a/b * sizeof(dtype) must be multiples of 16.
 */
int slider_concat_channels_nhwc(void* y, const void* a, const void* b, int64_t pixels, int channels_a, int channels_b,
                                int dtype, void* stream);

/*
 * Cross-attention core for short key sets: o = softmax(q k^T / sqrt(head_dim)) v per (batch, head), kv_len <= 80 (the
 * 77 prompt tokens of the UNet's attn2 layers; UNet calls at train_util.py:290-294, diffusers Attention with
 * encoder_hidden_states).  q, o: [batch, q_len, heads * head_dim], k, v: [batch, kv_len, heads * head_dim]
 * — the layout the q / k / v / out projections produce and consume, no head transposes.  16-bit storage (SLIDER_BF16 /
 * SLIDER_F16), fp32 accumulation and softmax; head_dim a multiple of 8 up to 160.  kv_static != 0 promises that k and v
 * are not written by any kernel still in flight on the stream (cached per prompt), so they are staged before the
 * dependent-launch wait.  No-grad forwards only (no backward form).  SLIDER_ERR_BAD_ARG outside these limits.
 */
int slider_cross_attention(void* o, const void* q, const void* k, const void* v, int batch, int q_len, int kv_len,
                           int heads, int head_dim, int dtype, int kv_static, void* stream);

/*
 * The same kernel with explicit row strides (elements): row r of batch item b of a tensor with stride ld starts at
 * (b * len + r) * ld — dense [batch, len, heads * head_dim] tensors (ld = heads * head_dim) or the q / k / v views of one
 * fused [batch, len, 3 * heads * head_dim] projection (ld = 3 * heads * head_dim; the 64-token self-attention of the
 * 8 x 8 level).  Strides are multiples of 8 elements and >= heads * head_dim.
 */
int slider_cross_attention_strided(void* o, const void* q, const void* k, const void* v, int batch, int q_len, int kv_len,
                                   int heads, int head_dim, int ldq, int ldk, int ldv, int ldo, int dtype, int kv_static,
                                   void* stream);

/*
 * Backward of slider_cross_attention for FROZEN keys / values (attn2 under the `noxattn` training methods: to_k / to_v
 * are neither adapted nor trainable and the prompt embeddings carry no gradient — autograd of the attention inside the
 * UNet call at train_lora.py:244-262): dq only, from dout, q, k, v (dense [batch, len, heads * head_dim]); nothing else
 * is saved by the forward.  Same limits as the forward.
 */
int slider_cross_attention_backward(void* dq, const void* dout, const void* q, const void* k, const void* v, int batch,
                                    int q_len, int kv_len, int heads, int head_dim, int dtype, void* stream);

/* x *= *dev_scalar  (chain rule for a non-unit upstream loss gradient). */
int slider_scale_by_device_scalar(void* x, const float* dev_scalar,
                                  int64_t n_elem, int dtype, void* stream);

/*
 * Flat-buffer AdamW over the contiguous LoRA parameter / gradient buffers
 * (torch.optim.AdamW built at conceptmod/textsliders/train_lora.py:80-89 and
 * train_lora_xl.py:104-105; clip_grad_norm_ at train_lora_xl.py:349).
 *   p      element dtype, updated in place;  grad fp32;  m, v fp32 state
 *   grad_scale_dev: optional device float multiplied into grad (1/world_size
 *   and/or the clip coefficient); NULL => 1.
 */
int slider_adamw_flat(void* p, const float* grad, float* m, float* v,
                      int64_t n_elem, float lr, float beta1, float beta2,
                      float eps, float weight_decay, int step,
                      const float* grad_scale_dev, int dtype, void* stream);
/* out[0] = min(1, max_norm/(||grad||_2*pre_scale + 1e-6)) * pre_scale */
int slider_clip_coef(float* out, const float* grad, int64_t n_elem,
                     float max_norm, float pre_scale, void* ws, size_t ws_bytes,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SLIDER_B200_H */
