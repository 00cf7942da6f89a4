/* fit_b200 -- C ABI of the B200-native FiT denoising hot path (libfit_b200.so).
 *
 * The reference (jlnk03/FiT) has no FFI: its hot path sits behind the Python module interface
 * `models.fit.FiT.forward / forward_with_cfg` (FiT/models/fit.py:474-553) and
 * `diffusion.create_diffusion(...).p_sample / ddim_sample / q_sample`
 * (FiT/diffusion/gaussian_diffusion.py:227-572).  fit_b200/models/fit.py and
 * fit_b200/diffusion/ mirror those Python interfaces and call the entry points below through
 * ctypes.  Each entry point cites the reference lines whose arithmetic it replaces.
 *
 * Conventions: plain pointers (device memory unless noted) + explicit sizes; the caller owns and
 * allocates every buffer (no allocation, no synchronisation, no hidden state in the library, so
 * every call is CUDA-graph capturable); all work is enqueued on `stream` (a cudaStream_t);
 * return 0 on success or a negative FITB200_ERR_* code, message in fitb200_last_error().
 * Row-major everywhere; "act dtype" is FITB200_F32 (verification mode) or FITB200_BF16.
 */
#ifndef FIT_B200_H
#define FIT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* fitb200_stream_t; /* cudaStream_t */

enum { FITB200_OK = 0, FITB200_ERR_INVALID_ARG = -1, FITB200_ERR_CUDA = -2, FITB200_ERR_UNSUPPORTED = -3 };
enum { FITB200_F32 = 0, FITB200_BF16 = 1 };

/* GEMM epilogues (all fused into the GEMM kernel) */
enum {
  FITB200_EPI_BIAS = 0,            /* out = acc + bias                      qkv (fit.py:234), adaLN (fit.py:323-324,329) */
  FITB200_EPI_BIAS_SILU = 1,       /* out = silu(acc + bias)                t-embedder first Linear (fit.py:55-57) */
  FITB200_EPI_BIAS_GATE_RESID = 2, /* out = resid + gate[sample] * (acc+bias)   proj / fc2 + adaLN-Zero gate + residual (fit.py:331-337) */
  FITB200_EPI_SWIGLU = 3,          /* out = silu(acc_g + b_g) * (acc_x + b_x)   SwiGLU fc1_g|fc1_x (fit.py:278-280) */
  FITB200_EPI_BIAS_ROPE = 4        /* out = rope(acc + bias): qkv Linear (fit.py:234) fused with apply_rotary_emb (fit.py:160-174) on the
                                      q and k columns [0, rope_cols); bf16 tcgen05 GEMM only */
};

const char* fitb200_last_error(void);
int fitb200_version(void);
/* 1 if the current device is compute capability 10.x (tcgen05 / TMEM available). Host query only. */
int fitb200_device_ok(void);

/* ---- token mask -> per-sample key range ------------------------------------------------------
 * kv_len[r] = 1 + index of the last True in mask[r, :] (0 if none); n_valid[r] = number of True.
 * Replaces the mask broadcast of Attention.forward (fit.py:184-185) by a per-sample bound used
 * for tile skipping; the mask bytes themselves are still honoured inside the attention kernel. */
int fitb200_mask_to_kvlen(const uint8_t* mask, int R, int T, int32_t* kv_len, int32_t* n_valid, fitb200_stream_t stream);

/* ---- patchify + x_embedder -------------------------------------------------------------------
 * FiT.patchify (fit.py:466-472) fused with x_embedder Linear(p*p*C -> D) (fit.py:387-388,499-501).
 * x: fp32; tokens [R_src, T, p*p*C] (x_is_nchw=0) or latents [R_src, C, H, W] (x_is_nchw=1,
 * T must equal (H/p)*(W/p)).  Output row r reads source row r % R_src (forward_with_cfg's
 * cat([half, half]), fit.py:545-546, without the copy).  w: [D, p*p*C] fp32, b: [D] fp32.
 * out: [R, T, D] in out_dtype (the residual stream). */
int fitb200_embed_tokens_fwd(const float* x, int x_is_nchw, int R, int R_src, int C, int H, int W, int p, int T,
                             const float* w, const float* b, int D, void* out, int out_dtype,
                             fitb200_stream_t stream);

/* ---- timestep sinusoid -----------------------------------------------------------------------
 * TimestepEmbedder.timestep_embedding (fit.py:61-80): out[r, :] = [cos(t f) | sin(t f)], dim 256.
 * freqs: optional device fp32 [128] frequency table (the reference builds it on the host, fit.py:73-75;
 * passing that table makes the arguments bit-identical); NULL = computed in-kernel. */
int fitb200_timestep_sincos(const int64_t* t, int R, const float* freqs, float* out /* [R,256] */,
                            fitb200_stream_t stream);

/* ---- c = t_emb + y_emb, silu(c) ---------------------------------------------------------------
 * LabelEmbedder gather (fit.py:114) + `c = t + y` (fit.py:505) + the SiLU that precedes every
 * adaLN Linear (fit.py:323, dit.py:134).  t_emb has t_rows rows (1 = batch-uniform timestep, or R).
 * Labels outside [0, table_rows) are clamped (the reference would raise).  c: fp32 [R,D] (nullable),
 * silu_c: [R,D] in act dtype. */
int fitb200_cond_combine_fwd(const float* t_emb, int t_rows, const int64_t* y, const float* table, int table_rows,
                             int R, int D, float* c, void* silu_c, int act_dtype, fitb200_stream_t stream);

/* ---- GEMM  out[M,N] = epilogue(A[M,K] . B[N,K]^T) --------------------------------------------- */
typedef struct {
  int32_t M, N, K;
  const void* A; int64_t lda;      /* [M,K] activations */
  const void* B; int64_t ldb;      /* [N,K] nn.Linear weight layout (fit.py:208,209,268,269,274,324) */
  const void* B2;                  /* fp32 SWIGLU only: fc1_x weight [N,K] (B = fc1_g) */
  const float* bias;               /* [N] or NULL */
  const float* bias2;              /* fp32 SWIGLU only: fc1_x bias */
  void* out; int64_t ldo; int32_t out_dtype;
  int32_t epilogue;                /* FITB200_EPI_* */
  const float* gate; int64_t ld_gate; int32_t rows_per_sample; /* GATE_RESID: gate[(m / rows_per_sample) * ld_gate + n] */
  const void* resid; int64_t ld_resid; int32_t resid_dtype;    /* GATE_RESID: may alias out */
  /* BIAS_ROPE: output row m is token m % rope_tokens of sample m / rope_tokens; column n < rope_cols belongs to head-dim index
   * n % rope_head_dim; the pair (2j, 2j+1) is rotated by (cos, sin) = rope_freqs[(s or 0), token, 2j .. 2j+1]
   * (fp32 [rope_rows = 1 | samples, rope_tokens, rope_head_dim], the table of pos_embed.py:51-75). */
  const float* rope_freqs; int32_t rope_rows, rope_tokens, rope_head_dim, rope_cols;
  /* Backward GEMMs (autograd of nn.Linear, FiT/train.py:55-66): a_mn = 1: A is stored [K, M] row-major (lda = its row stride);
   * b_mn = 1: B is stored [K, N] row-major.  out[m,n] = sum_k A(m,k) B(n,k) in every case.
   *   dgrad  dX[M,K'] = dY[M,N] W[N,K']   : A = dY,  B = W  with b_mn = 1 (no transposed weight copy)
   *   wgrad  dW[N,K'] = dY[M,N]^T X[M,K'] : A = dY with a_mn = 1, B = X with b_mn = 1 (contraction over the token rows)
   * FITB200_EPI_BIAS (bias may be NULL) only. */
  int32_t a_mn, b_mn;
  int32_t bias_row0;               /* bf16 GEMM: the bias is added to rows m >= bias_row0 only (0 = all rows) */
  /* ---- LayerNorm + adaLN-modulate PROLOGUE (bf16 tcgen05 GEMM; epilogue BIAS, BIAS_ROPE or SWIGLU; M > 128, N > 128) ----
   * ln_g != NULL: A is the raw residual stream x [M,K] and the GEMM computes  epilogue( modulate(LayerNorm_eps(x), shift, scale) . B^T + bias )
   * (fit.py:45-46,303,308,332,336: norm1/norm2 without affine, eps 1e-6) without materialising the modulated rows:
   *     rstd_m * ( sum_k (x[m,k] * g[s,k]) B[n,k]  -  mean_m * c1[s,n] )  +  c2[s,n],      s = m / rows_per_sample
   *   ln_g     bf16 [ln_samples, K] (row stride ld_ln_g):  g = bf16(1 + scale)
   *   ln_c     fp32 [2 * ln_samples, N]:  rows [0, S): c1[s,n] = sum_k g[s,k] B[n,k];  rows [S, 2S): c2[s,n] = bias[n] + sum_k bf16(shift[s,k]) B[n,k]
   *            (fitb200_modulation_pack + one FITB200_EPI_BIAS GEMM of the packed rows with the same B, bias_row0 = S; `bias` is ignored here)
   *   ln_parts fp32 [M, ln_nparts, 2]: partial (sum x, sum x^2) of every row (summed in order; from stats_out of the GEMM that wrote x, or
   *            fitb200_row_stats);  ln_eps: the LayerNorm epsilon. */
  const void* ln_g; int64_t ld_ln_g;
  const float* ln_c; int32_t ln_samples;
  const float* ln_parts; int32_t ln_nparts; float ln_eps;
  /* GATE_RESID (bf16 GEMM): optional fp32 [M, fitb200_gemm_stats_parts(N), 2] partial (sum, sum of squares) of every OUTPUT row -- the
   * statistics of the LayerNorm that follows (fit.py:332,336), so that nothing re-reads the residual stream for them. */
  float* stats_out;
} fitb200_gemm_args;

/* number of partial sums per row written through stats_out for an N-column output */
int fitb200_gemm_stats_parts(int N);

/* fp32 FMA GEMM on CUDA cores: A, B, B2 fp32.  The fp32 verification mode and the small
 * embedder GEMMs.  SWIGLU: N = hidden width, out [M,N]. */
int fitb200_gemm_f32(const fitb200_gemm_args* a, fitb200_stream_t stream);

/* bf16 tcgen05 GEMM: A, B bf16, fp32 accumulation in TMEM, operands staged by TMA.
 * K % 8 == 0, N % 8 == 0, 16-byte aligned rows.  SWIGLU: B is the packed [N = 2*Hp, K] weight whose
 * rows alternate 16 rows of fc1_g with the matching 16 rows of fc1_x (bias packed alike);
 * out is [M, Hp]. */
int fitb200_gemm_bf16_tc(const fitb200_gemm_args* a, fitb200_stream_t stream);

/* ---- inputs of the GEMM LayerNorm + modulate prologue (fitb200_gemm_args.ln_*) -----------------------
 * row_stats: parts[m, 0] = (sum_k x[m,k], sum_k x[m,k]^2), parts[m, 1..nparts) = 0 -- for a residual stream no GATE_RESID GEMM wrote
 * (the embedder's output, fit.py:499-501).  x: [M, D] act dtype; parts: fp32 [M, nparts, 2].
 * modulation_pack: out[set, s, :] = bf16(1 + scale_set[s, :]), out[set, S + s, :] = bf16(shift_set[s, :]) with
 * shift_set = shift + set * set_stride, scale_set = scale + set * set_stride (fp32 rows of stride ld_mod: the adaLN chunks of fit.py:329-330;
 * n_sets = 2 * blocks, set_stride = 3 D packs norm1 and norm2 of every block in one launch).  out: bf16 [n_sets, 2 S, D]. */
int fitb200_row_stats(const void* x, int x_dtype, int M, int D, float* parts, int nparts, fitb200_stream_t stream);
int fitb200_modulation_pack(const float* shift, const float* scale, int64_t ld_mod, int S, int D, int n_sets, int64_t set_stride, void* out,
                            fitb200_stream_t stream);

/* ---- LayerNorm(eps 1e-6, no affine) + adaLN modulate -------------------------------------------
 * modulate(norm(x), shift, scale) (fit.py:45-46,303,308,332,336): out = LN(x) * (1 + scale[s]) + shift[s],
 * s = row / rows_per_sample.  shift/scale: fp32 with row stride ld_mod.  mean/rstd: optional fp32 [M]
 * (saved for backward). */
int fitb200_ln_modulate_fwd(const void* x, int x_dtype, int M, int D, int rows_per_sample, const float* shift,
                            const float* scale, int64_t ld_mod, void* out, int out_dtype, float* mean, float* rstd,
                            fitb200_stream_t stream);

/* ---- fused attention: 2-D RoPE on q,k + key-masked softmax(q k^T / sqrt(d)) v -------------------
 * apply_rotary_emb (fit.py:160-174) + Attention.forward (fit.py:176-188) + _rearange_out (fit.py:220-227).
 * qkv: [R, T, 3, h, d] (the qkv Linear output, column = which*h*d + head*d + i, fit.py:234-235).
 * freqs: fp32 [freqs_rows, T, d] interleaved (cos, sin) pairs (freqs_rows = 1 broadcasts, fit.py:166-167).
 * Key j of sample r takes part iff j < kv_len[r] and (mask == NULL or mask[r*T + j]).
 * Query tiles starting at or beyond kv_len[r] are skipped and their output rows zero-filled (they are
 * padding whose reference output is unspecified).  out: [R, T, h*d].
 * lse: optional fp32 [R, h, T] log-sum-exp (natural log, of the scaled scores) for backward.
 * d in {64, 72}.
 * flags: FITB200_ATTN_PREROTATED = q and k in qkv already carry the RoPE (written by a FITB200_EPI_BIAS_ROPE GEMM); freqs is then ignored
 *        (may be NULL).  REQUIRED for bf16 (un-rotated bf16 operands return FITB200_ERR_UNSUPPORTED); not accepted for fp32.
 * bf16, T <= 256: persistent tcgen05 / TMEM kernel (attn_tc2.cu); bf16, T > 256: KV-blocked tcgen05 kernel (attn_tc3.cu); fp32: CUDA-core
 * kernel with the RoPE applied in the kernel (attn_simt.cu). */
enum { FITB200_ATTN_PREROTATED = 1, FITB200_ATTN_FORCE_SIMT = 2 /* backward only: take the CUDA-core kernel */ };
int fitb200_attn_fwd(const void* qkv, int dtype, const float* freqs, int freqs_rows, const int32_t* kv_len,
                     const uint8_t* mask, int R, int T, int h, int d, int flags, void* out, float* lse,
                     fitb200_stream_t stream);

/* Backward of the above (autograd of fit.py:160-188 as triggered by loss.backward(), FiT/train.py:55-66).
 * qkv / flags / freqs / kv_len / mask exactly as given to the forward; out = the forward's output [R, T, h*d], dout its gradient,
 * lse the forward's log-sum-exp.  dqkv [R, T, 3, h, d] receives the gradient with respect to the UN-rotated q, k (and v): the
 * RoPE's transpose (rotation by -theta) is applied in the kernel, so dqkv feeds the qkv Linear's dgrad / wgrad directly.
 * Rows of padded tokens are written as zeros.  bf16 + PREROTATED + T <= 256: tcgen05 / TMEM kernel (attn_bwd_tc.cu); otherwise the
 * CUDA-core kernel (exact fp32 math; bf16 storage supported). */
int fitb200_attn_bwd(const void* qkv, int dtype, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask,
                     int R, int T, int h, int d, int flags, const void* out, const void* dout, const float* lse, void* dqkv,
                     fitb200_stream_t stream);

/* The same for rows whose partial (sum x, sum x^2) are already known -- written by the GATE_RESID GEMM that produced x (fitb200_gemm_args.stats_out)
 * or by fitb200_row_stats: a streaming pass without the row reduction.  parts: fp32 [M, nparts, 2]; eps: the LayerNorm epsilon (1e-6). */
int fitb200_ln_apply_fwd(const void* x, int x_dtype, int M, int D, int rows_per_sample, const float* parts, int nparts, float eps,
                         const float* shift, const float* scale, int64_t ld_mod, void* out, int out_dtype, fitb200_stream_t stream);

/* fitb200_ln_modulate_fwd without ANY caller- or allocator-provided scratch.  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same
 * result as fitb200_ln_modulate_fwd (eps 1e-6; no mean / rstd outputs): the row sums go through a file-scope `static __device__` array of
 * FITB200_STATIC_LN_MAX_ROWS entries (1 MiB per GPU that loads the library, held until exit) and a streaming modulate pass.  M beyond that
 * limit returns FITB200_ERR_INVALID_ARG.  The entry point calls no cudaMalloc / cudaFree / allocating routine.  All calls on one GPU share
 * the array (it keeps the last call's contents; nothing in it has to start at zero): calls that overlap on the GPU (different streams) RACE. */
#define FITB200_STATIC_LN_MAX_ROWS 131072
int fitb200_ln_modulate_fwd_static(const void* x, int x_dtype, int M, int D, int rows_per_sample, const float* shift, const float* scale,
                                   int64_t ld_mod, void* out, int out_dtype, fitb200_stream_t stream);

/* ---- final layer ---------------------------------------------------------------------------------
 * FinalLayer.forward (dit.py:138-142) after its adaLN Linear: LN + modulate + Linear(D -> P) with
 * P = p*p*C_out, optionally fused with FiT.unpatchify (fit.py:444-456).
 * out: fp32 tokens [R, T, P] (out_nchw = 0) or latents [R, C, H, W] (out_nchw = 1, T = (H/p)*(W/p)). */
int fitb200_final_layer_fwd(const void* x, int x_dtype, int R, int T, int D, const float* shift, const float* scale,
                            int64_t ld_mod, const float* w, const float* b, int P, float* out, int out_nchw, int C,
                            int H, int W, int p, fitb200_stream_t stream);

/* FiT.unpatchify (fit.py:444-456) alone: tokens fp32 [R, T, p*p*C] -> latents fp32 [R, C, H, W], T = (H/p)*(W/p).  (The bf16 engine's final
 * layer is a tensor-core GEMM with the LayerNorm prologue that writes tokens; fitb200_final_layer_fwd fuses the map itself.) */
int fitb200_unpatchify(const float* tokens, int R, int T, int C, int H, int W, int p, float* out, fitb200_stream_t stream);

/* ---- diffusion step ------------------------------------------------------------------------------
 * coef: device fp32 table [S, 8] per (respaced) timestep:
 *   0 sqrt_recip_alphas_cumprod   1 sqrt_recipm1_alphas_cumprod   2 posterior_mean_coef1
 *   3 posterior_mean_coef2        4 model log-variance (FIXED_LARGE / FIXED_SMALL)
 *   5 alphas_cumprod              6 alphas_cumprod_prev           7 (unused)
 * (gaussian_diffusion.py:179-213,307-320), each the float64 table value cast to fp32 exactly as
 * _extract_into_tensor does (:891).
 * cfg_pairs = B > 0: model_out holds the raw network output for rows [0,B) conditional and [B,2B)
 * unconditional and the classifier-free-guidance combine of forward_with_cfg (fit.py:549-552) is
 * applied on the fly: eps[r] = u + s (c - u) with pair index r % B; x / noise / sample then have
 * rows = 2B (the reference's duplicated batch) or rows = B (only the half the reference keeps,
 * sample_lightning.py:69).  cfg_pairs = 0: model_out is eps.
 * ddpm: p_mean_variance EPSILON branch, clip_denoised=False (:332-335,346-351,244-264) + p_sample (:422-428):
 *   x0 = c0 x - c1 eps;  mean = c2 x0 + c3 x;  sample = mean + (t != 0) exp(0.5 logvar) noise.
 * ddim (:553-572): eps' re-derived from x0, sigma = eta sqrt((1-ab_prev)/(1-ab)) sqrt(1-ab/ab_prev),
 *   sample = sqrt(ab_prev) x0 + sqrt(1-ab_prev-sigma^2) eps' + (t != 0) sigma noise.
 * flags: FITB200_STEP_CLIP_DENOISED clamps x0 to [-1,1] (process_xstart, :322-326);
 *        FITB200_STEP_PREDICT_XSTART: the model output is x0 itself (START_X, :328-331).
 * t: device int64 [rows] indices into coef.  noise may be NULL (= 0).  sample / pred_xstart / mean
 * (the posterior mean p_mean_variance returns, :335) are each nullable; sample may alias x. */
enum { FITB200_STEP_CLIP_DENOISED = 1, FITB200_STEP_PREDICT_XSTART = 2 };
int fitb200_cfg_ddpm_step(const float* x, const float* model_out, const float* noise, float* sample,
                          float* pred_xstart, float* mean, int rows, int64_t per_row, int cfg_pairs, float cfg_scale,
                          int flags, const int64_t* t, const float* coef, int S, fitb200_stream_t stream);
int fitb200_cfg_ddim_step(const float* x, const float* model_out, const float* noise, float* sample,
                          float* pred_xstart, int rows, int64_t per_row, int cfg_pairs, float cfg_scale, float eta,
                          int flags, const int64_t* t, const float* coef, int S, fitb200_stream_t stream);

/* forward_with_cfg's combine alone (fit.py:549-553): model_out rows [0,pairs) conditional, [pairs,2 pairs)
 * unconditional; out[r] = out[r + pairs] = u + s (c - u).  out must not alias model_out. */
int fitb200_cfg_combine(const float* model_out, float* out, int pairs, int64_t per_row, float cfg_scale,
                        fitb200_stream_t stream);

/* q_sample (gaussian_diffusion.py:227-242) = DDIMScheduler.add_noise call site (FiT/train.py:53):
 * out = a[t] x0 + b[t] noise with a = sqrt_alphas_cumprod, b = sqrt_one_minus_alphas_cumprod (fp32 [S]). */
int fitb200_q_sample(const float* x0, const float* noise, float* out, int rows, int64_t per_row, const int64_t* t,
                     const float* sqrt_ac, const float* sqrt_1m_ac, int S, fitb200_stream_t stream);

/* Per-step Gaussian noise of p_sample (th.randn_like, gaussian_diffusion.py:422) as a counter-based generator:
 * Philox4x32-10 keyed by (seed, *step), counter = (row_offset + row) * per_row/4 + i, Box-Muller -> N(0,1).
 * Keyed by the GLOBAL image index so that a batch sharded over several GPUs draws, per image, exactly the
 * noise of the unsharded run; `step` and (optionally) `row_offset_dev[0]`, which is added to `row_offset`, are read from device
 * memory, so ONE captured CUDA graph serves every step and every batch position (sample_lightning.py:32-70 loop).  Not the torch
 * generator's stream: parity tests pass the reference's noise explicitly instead. */
int fitb200_philox_normal(float* out, int rows, int64_t per_row, uint64_t seed, const int64_t* step, int64_t row_offset,
                          const int64_t* row_offset_dev, fitb200_stream_t stream);

/* loop plumbing for CUDA-graph replay of the sampling loop (gaussian_diffusion.py:679, respace.py:125-126):
 * out[i] = map[clamp(idx[i])]; and idx[i] += delta. */
int fitb200_gather_i64(const int64_t* map, int map_len, const int64_t* idx, int64_t* out, int n, fitb200_stream_t stream);
int fitb200_add_i64(int64_t* idx, int n, int64_t delta, fitb200_stream_t stream);

/* ==== training step: un-fused forward pieces + backward kernels =====================================================
 * Autograd of FiTBlock.forward (fit.py:326-337) and of the training-step arithmetic (FiT/train.py:45-69).  "act dtype" tensors
 * are bf16 or fp32; reductions, modulation tables and parameter gradients are fp32.  Buffers marked "+=" are accumulated with
 * atomics and must be zero-initialised by the caller. */

/* out = resid + gate[m / rows_per_sample] * y   (x + gate * branch, fit.py:331,337; the training forward keeps y) */
int fitb200_gate_resid_fwd(const void* resid, const void* y, int dtype, const float* gate, int64_t ld_gate, int M, int D,
                           int rows_per_sample, void* out, fitb200_stream_t stream);
/* dy = gate[s] * dx;  dgate[s,:] += sum_t dx * y;  dbias[:] += column sums of dy (nullable) */
int fitb200_gate_bwd(const void* dx, const void* y, int dtype, const float* gate, int64_t ld_gate, int R, int rows_per_sample, int D,
                     void* dy, float* dgate, int64_t ld_dgate, float* dbias, fitb200_stream_t stream);
/* SwiGLU gating (fit.py:278-280) on the packed pre-activations u [M, 2*Hp] (16 gate columns then the matching 16 up columns,
 * the layout of the packed fc1 weight):  g = silu(u_gate) * u_up;  backward: du (same layout), dbias[2*Hp] += column sums */
int fitb200_swiglu_fwd(const void* u, int dtype, int M, int Hp, void* g, fitb200_stream_t stream);
int fitb200_swiglu_bwd(const void* dg, const void* u, int dtype, int M, int Hp, void* du, float* dbias, fitb200_stream_t stream);
/* modulate(LayerNorm(x)) backward (fit.py:45-46,303,332,336): dx = dres (nullable) + LN^T(dh * (1 + scale[s]));
 * dshift[s,:] += sum_t dh;  dscale[s,:] += sum_t dh * xhat  (both nullable together).  mean / rstd from the forward. */
int fitb200_ln_modulate_bwd(const void* dh, const void* x, int dtype, const float* mean, const float* rstd, const float* scale,
                            int64_t ld_mod, int M, int D, int rows_per_sample, const void* dres, void* dx, float* dshift, float* dscale,
                            int64_t ld_dmod, fitb200_stream_t stream);
/* out[n] += sum_m a[m, n]   (bias gradients) */
int fitb200_colsum(const void* a, int dtype, int M, int N, int64_t lda, float* out, fitb200_stream_t stream);
/* y = silu(x);  dx = dy * silu'(x)   (t-embedder MLP fit.py:55-57, adaLN SiLU fit.py:323) */
int fitb200_silu_fwd(const float* x, int64_t n, float* y, fitb200_stream_t stream);
int fitb200_silu_bwd(const float* dy, const float* x, int64_t n, float* dx, fitb200_stream_t stream);
/* dtable[y[r], :] += dc[r, :]   (LabelEmbedder, fit.py:114) */
int fitb200_embedding_bwd(const float* dc, const int64_t* y, int R, int D, int table_rows, float* dtable, fitb200_stream_t stream);
/* dst = (dst dtype) src */
int fitb200_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, fitb200_stream_t stream);
/* F.mse_loss(out[mask], target[mask]) (FiT/train.py:63-66): out / target fp32 [tokens, P], mask bytes [tokens].
 * fwd: acc[0] += sum of squared errors over valid tokens, acc[1] += number of valid tokens (acc zeroed by the caller),
 *      loss[0] = acc[0] / (acc[1] * P).   bwd: dout = gscale[0] (nullable = 1) * 2 (out - target) mask / (acc[1] * P). */
int fitb200_masked_mse_fwd(const float* out, const float* target, const uint8_t* mask, int64_t tokens, int P, float* acc, float* loss,
                           fitb200_stream_t stream);
int fitb200_masked_mse_bwd(const float* out, const float* target, const uint8_t* mask, int64_t tokens, int P, const float* acc,
                           const float* gscale, void* dout, int dout_dtype, fitb200_stream_t stream);


/* Cap on the SMs the persistent kernels (GEMM, attention) occupy; 0 = all.  fit_b200.train.GradSync lowers it while NCCL's all-reduce
 * CTAs run beside the backward pass (DDP, FiT/train.py:287-288 via Lightning) so that both fit in one wave.  Process-wide, host-side. */
int fitb200_set_sm_limit(int n);

/* ---- masked_FiT token sub-sampling: ImageNetLatentIterator.subsample / collate
 * (masked_FiT/preprocess/iterators.py:142-169) for a whole batch in one launch.  Sample s holds n_tokens[s] valid tokens in
 * latent[s, :, :P], pos[s, :, :d] (d = RoPE row width) and mask[s, :] (bytes).  With T' = Tout:
 *   n_tokens[s] >  T': out[s, j] = in[s, perm[s, j]] for j < T'  (random subset in permuted order, iterators.py:143-148; the
 *                      permutation is an INPUT -- torch.randperm stays on the host for RNG parity with the reference);
 *   n_tokens[s] <= T': out[s, j] = in[s, j] for j < n_tokens[s], zeros (mask False) behind (iterators.py:150-154, no permutation).
 * Bit-exact copies.  perm: int64 [N, Tsrc]; latent_out [N, T', P], pos_out [N, T', d] fp32, mask_out [N, T'] bytes. */
int fitb200_subsample_tokens(const float* latent, const float* pos, const uint8_t* mask, const int64_t* perm, const int32_t* n_tokens, int N,
                             int Tsrc, int P, int d, int Tout, float* latent_out, float* pos_out, uint8_t* mask_out, fitb200_stream_t stream);

/* ---- optimizer step: torch.optim.AdamW(lr=1e-4, weight_decay=0) (FiT/train.py:215-217) followed by the
 * EMA update ema = ema * decay + p * (1 - decay) (FiT/ema.py:149-154,256-267) as ONE multi-tensor pass over (p, g, m, v, ema).
 * Per element, in torch's single-tensor AdamW order:  p *= 1 - lr * wd;  m += (g - m) * (1 - beta1);  v = v * beta2 + g * g * (1 - beta2);
 * p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)   with bc1 = 1 - beta1^t, bc2 = 1 - beta2^t computed by the caller in double.
 * tensors: device array of n_tensors descriptors; chunks: device array of n_chunks {tensor index, first element} pairs, every chunk
 * covers FITB200_ADAM_CHUNK elements (the last chunk of a tensor is cut by numel).  ema may be null per tensor (no EMA) and
 * ema_decay < 0 skips the EMA update of this step (every_n_steps).  All buffers fp32, 16-byte aligned.  36 B per element.
 * Hyper-parameters are doubles: torch forms 1 - beta, lr / bc1, 1 - decay in double before the fp32 arithmetic, and so does this. */
#define FITB200_ADAM_CHUNK 16384
typedef struct fitb200_adam_tensor {
  float* p;
  const float* g;
  float* m;
  float* v;
  float* ema;
  int64_t numel;
  /* optional kernel-layout copy of the UPDATED parameter, written in the same pass (NULL = none): the bf16 (or packed fp32) operand the
   * GEMMs read, so that nothing re-casts / re-packs 675 M parameters after every step.  p is viewed as [numel / row_len, row_len];
   * element (row, col) goes to shadow[srow * shadow_ld + col] with srow = row (pack_off < 0) or (row / 16) * 32 + pack_off + row % 16
   * (pack_off = 0 / 16: the fc1_g / fc1_x halves of the packed SwiGLU weight and bias, FITB200_EPI_SWIGLU). */
  void* shadow;
  int64_t shadow_ld;
  int32_t shadow_dtype; /* FITB200_BF16 or FITB200_F32 */
  int32_t row_len;
  int32_t pack_off;
  int32_t reserved_;
} fitb200_adam_tensor;
int fitb200_adamw_ema_step(const fitb200_adam_tensor* tensors, const int32_t* chunks, int n_chunks, double lr, double beta1, double beta2,
                           double eps, double weight_decay, double bias_correction1, double bias_correction2_sqrt, double ema_decay,
                           fitb200_stream_t stream);

/* The same step with every step-dependent scalar read from DEVICE memory, so that a captured CUDA graph of the training step can be replayed
 * for every step (the capturable=True mode of torch.optim.AdamW): hyper = device double[6] {lr, beta1, beta2, eps, weight_decay, ema_decay
 * (< 0: no EMA)}; step = device int64[1], the 1-based number of THIS step (the caller increments it first, e.g. fitb200_add_i64); the bias
 * corrections 1 - beta^step are formed in double in the kernel; the EMA moves when (step - 1) % every_n_steps == 0 (FiT/ema.py:252-257). */
int fitb200_adamw_ema_step_dev(const fitb200_adam_tensor* tensors, const int32_t* chunks, int n_chunks, const double* hyper, const int64_t* step,
                               int every_n_steps, fitb200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* FIT_B200_H */
