// Coalesced, vectorised elementwise / row-reduction kernels of the FiT hot path (family 3 and 4 of
// SURVEY.md section 8): mask scan, patchify+embed, timestep sinusoid, conditioning combine,
// LayerNorm+modulate, final layer (+unpatchify), fused CFG + DDPM/DDIM step, q_sample.
// All HBM-bound: one pass over the data, 16-byte accesses, fp32 math.
#include <cstdarg>

#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace fitb200 {

static thread_local char g_err[512] = "";
int& sm_limit_ref() {
  static int lim = 0;
  return lim;
}
void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// --------------------------------------------------------------------------------------------
// mask -> kv_len / n_valid : one warp per sample
// --------------------------------------------------------------------------------------------
__global__ void mask_to_kvlen_kernel(const uint8_t* __restrict__ mask, int R, int T, int32_t* __restrict__ kv_len,
                                     int32_t* __restrict__ n_valid) {
  int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (r >= R) return;
  int last = 0, cnt = 0;
  for (int j = lane; j < T; j += 32) {
    if (mask[(size_t)r * T + j]) { last = j + 1; ++cnt; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if (lane == 0) {
    kv_len[r] = last;
    if (n_valid) n_valid[r] = cnt;
  }
}

// --------------------------------------------------------------------------------------------
// patchify + x_embedder.  CTA: TOK tokens x 128 output columns.
// --------------------------------------------------------------------------------------------
constexpr int EMB_TOK = 16;
constexpr int EMB_COLS = 4;   // output columns per thread: 8-byte (bf16) / 16-byte (fp32) stores, a warp writes 256 / 512 contiguous bytes per token

template <typename OutT>
__global__ void __launch_bounds__(128) embed_tokens_kernel(const float* __restrict__ x, int nchw, int R, int R_src, int C,
                                                           int H, int W, int p, int T, const float* __restrict__ w,
                                                           const float* __restrict__ b, int D, OutT* __restrict__ out) {
  extern __shared__ float xs[];  // [EMB_TOK][K]
  const int K = p * p * C;
  const long tok0 = (long)blockIdx.x * EMB_TOK;
  const long ntok = (long)R * T;
  const int nw = nchw ? (W / p) : 0;
  for (int i = threadIdx.x; i < EMB_TOK * K; i += blockDim.x) {
    int tl = i / K, k = i - tl * K;
    long tok = tok0 + tl;
    float v = 0.f;
    if (tok < ntok) {
      int r = (int)(tok / T), tk = (int)(tok - (long)r * T);
      int rs = r % R_src;
      if (nchw) {
        int ih = tk / nw, iw = tk - ih * nw;
        int c = k % C, pp = k / C;
        int ph = pp / p, pw = pp - ph * p;
        v = x[(((size_t)rs * C + c) * H + (ih * p + ph)) * W + (iw * p + pw)];
      } else {
        v = x[((size_t)rs * T + tk) * K + k];
      }
    }
    xs[i] = v;
  }
  __syncthreads();
  const int j0 = (blockIdx.y * 128 + threadIdx.x) * EMB_COLS;   // D % 4 == 0 (checked on the host)
  if (j0 >= D) return;
  float acc[EMB_TOK][EMB_COLS];
  const float4 b4 = *reinterpret_cast<const float4*>(b + j0);
#pragma unroll
  for (int t = 0; t < EMB_TOK; ++t) { acc[t][0] = b4.x; acc[t][1] = b4.y; acc[t][2] = b4.z; acc[t][3] = b4.w; }
  for (int k0 = 0; k0 < K; k0 += 4) {
    float4 wv[EMB_COLS];
#pragma unroll
    for (int c = 0; c < EMB_COLS; ++c) wv[c] = *reinterpret_cast<const float4*>(w + (size_t)(j0 + c) * K + k0);
#pragma unroll
    for (int t = 0; t < EMB_TOK; ++t) {
      const float4 xv = *reinterpret_cast<const float4*>(xs + t * K + k0);   // the same address for the whole warp: a broadcast
#pragma unroll
      for (int c = 0; c < EMB_COLS; ++c) {
        float a = acc[t][c];
        a = fmaf(wv[c].x, xv.x, a); a = fmaf(wv[c].y, xv.y, a); a = fmaf(wv[c].z, xv.z, a); a = fmaf(wv[c].w, xv.w, a);
        acc[t][c] = a;
      }
    }
  }
#pragma unroll
  for (int t = 0; t < EMB_TOK; ++t) {
    long tok = tok0 + t;
    if (tok < ntok) {
      if constexpr (sizeof(OutT) == 2) {
        uint2 o; o.x = pack_bf16x2(acc[t][0], acc[t][1]); o.y = pack_bf16x2(acc[t][2], acc[t][3]);
        *reinterpret_cast<uint2*>(out + (size_t)tok * D + j0) = o;
      } else {
        *reinterpret_cast<float4*>(out + (size_t)tok * D + j0) = make_float4(acc[t][0], acc[t][1], acc[t][2], acc[t][3]);
      }
    }
  }
}

// --------------------------------------------------------------------------------------------
// timestep sinusoid
// --------------------------------------------------------------------------------------------
__global__ void timestep_sincos_kernel(const int64_t* __restrict__ t, int R, const float* __restrict__ freqs,
                                       float* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R * 128) return;
  int r = i >> 7, k = i & 127;
  // fit.py:73-76: freqs = exp(-ln(1e4) * arange(half) / half) in fp32 (computed on the HOST by the
  // reference, so the caller may pass that exact table), args = t.float() * freqs
  float f = freqs ? freqs[k] : expf(-9.210340371976184f * (float)k / 128.0f);
  float a = (float)t[r] * f;
  float s, c;
  sincosf(a, &s, &c);
  out[(size_t)r * 256 + k] = c;
  out[(size_t)r * 256 + 128 + k] = s;
}

// --------------------------------------------------------------------------------------------
// c = t_emb + table[y];  silu(c)
// --------------------------------------------------------------------------------------------
template <typename ActT>
__global__ void cond_combine_kernel(const float* __restrict__ t_emb, int t_rows, const int64_t* __restrict__ y,
                                    const float* __restrict__ table, int table_rows, int R, int D, float* __restrict__ c,
                                    ActT* __restrict__ silu_c) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)R * D) return;
  int r = (int)(i / D), j = (int)(i - (long)r * D);
  long lbl = y[r];
  lbl = lbl < 0 ? 0 : (lbl >= table_rows ? table_rows - 1 : lbl);
  float v = t_emb[(size_t)(t_rows == 1 ? 0 : r) * D + j] + table[(size_t)lbl * D + j];
  if (c) c[i] = v;
  silu_c[i] = from_f32<ActT>(silu_acc(v));
}

// --------------------------------------------------------------------------------------------
// LayerNorm + modulate.  One CTA = 8 warps x RPW consecutive rows of ONE sample: (1 + scale) and shift of that sample are staged
// once in shared memory in a lane-major float4 layout (no bank conflicts) and read ONCE per RPW rows; a warp keeps its RPW = 2
// bf16 rows packed in registers (loads of all four in flight together), D <= 2048, D % 8 == 0, 16-byte accesses.
// History on B200 (R=512 XL/2): one row per warp + [D] smem table = 0.158 ms (l1tex 74 %, 10.6 M bank conflicts per launch).
// --------------------------------------------------------------------------------------------
constexpr int LN_MAXV = 8;    // 8 vectors of 8 per lane
constexpr int LN_ITERS = 4;   // row rounds per CTA: the staged modulation is reused by 8 warps x RPW x LN_ITERS rows

// raw 8-element row fragments: bf16 rows stay packed (4 registers) so that one warp can hold several rows in flight
template <typename T> struct Raw8;
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) { a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 4); }
  __device__ __forceinline__ void get(float (&f)[8]) const { f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w; }
};
template <> struct Raw8<__nv_bfloat16> {
  uint4 a;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { a = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void get(float (&f)[8]) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&a);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
};
template <typename InT> struct LnRows { static constexpr int RPW = sizeof(InT) == 2 ? 2 : 1; };  // rows one warp holds in flight

template <typename InT, typename OutT, int NV>
__global__ void __launch_bounds__(256, 2) ln_modulate_kernel(const InT* __restrict__ x, int M, int D, int rows_per_sample,
                                                          int blocks_per_sample, const float* __restrict__ shift,
                                                          const float* __restrict__ scale, long ld_mod, OutT* __restrict__ out,
                                                          float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  constexpr int RPW = LnRows<InT>::RPW;
  // 1+scale and shift of this CTA's sample as float4 [i][half][lane]: lane l's two float4 of vector (l + 32 i) -- conflict-free reads
  extern __shared__ __align__(16) float4 ln_mod4[];
  float4* sc4 = ln_mod4;
  float4* sh4 = ln_mod4 + NV * 64;
  // rows are walked from the END of the tensor: the GEMM that produced x wrote its last rows most recently (still in L2), and the GEMM
  // that consumes `out` starts at row 0, which this order writes last
  const int bid = gridDim.x - 1 - blockIdx.x;
  const int sample = bid / blocks_per_sample, blk = bid - sample * blocks_per_sample;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = D >> 3;
  const size_t mrow = (size_t)sample * ld_mod;
  for (int it = 0; it < LN_ITERS; ++it) {
  const int rl0 = ((blk * LN_ITERS + it) * 8 + warp) * RPW;  // first row (inside the sample) of this warp in this round
  if (it > 0 && rl0 >= rows_per_sample) break;  // (round 0 always reaches the __syncthreads below)
  Raw8<InT> v[RPW][NV];
  bool live[RPW];
#pragma unroll
  for (int j = 0; j < RPW; ++j) {
    const int row = sample * rows_per_sample + rl0 + j;
    live[j] = rl0 + j < rows_per_sample && row < M;
    if (live[j]) {
      const InT* xr = x + (size_t)row * D;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vi = lane + i * 32;
        if (vi < nvec) v[j][i].load(xr + vi * 8);
      }
    }
  }
  // modulation staging after the first row loads are in flight (independent latencies overlap)
  if (it == 0) {
  for (int i = threadIdx.x; i < D / 4; i += blockDim.x) {
    float4 sc = *reinterpret_cast<const float4*>(scale + mrow + i * 4);
    const float4 sh = *reinterpret_cast<const float4*>(shift + mrow + i * 4);
    sc.x += 1.0f; sc.y += 1.0f; sc.z += 1.0f; sc.w += 1.0f;
    const int vi = i >> 1, pos = ((vi >> 5) * 2 + (i & 1)) * 32 + (vi & 31);
    sc4[pos] = sc;
    sh4[pos] = sh;
  }
  __syncthreads();
  }
  float mu[RPW], rstd[RPW];
#pragma unroll
  for (int j = 0; j < RPW; ++j) {
    float s = 0.f;
    if (live[j]) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        if (lane + i * 32 < nvec) {
          float f[8];
          v[j][i].get(f);
#pragma unroll
          for (int e = 0; e < 8; ++e) s += f[e];
        }
      }
    }
    mu[j] = warp_sum(s) / (float)D;
  }
#pragma unroll
  for (int j = 0; j < RPW; ++j) {
    float q = 0.f;
    if (live[j]) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        if (lane + i * 32 < nvec) {
          float f[8];
          v[j][i].get(f);
#pragma unroll
          for (int e = 0; e < 8; ++e) { const float d = f[e] - mu[j]; q = fmaf(d, d, q); }
        }
      }
    }
    rstd[j] = rsqrtf(warp_sum(q) / (float)D + 1e-6f);
    if (live[j] && lane == 0) {
      const int row = sample * rows_per_sample + rl0 + j;
      if (mean_out) mean_out[row] = mu[j];
      if (rstd_out) rstd_out[row] = rstd[j];
    }
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    if (lane + i * 32 < nvec) {
      const float4 c0 = sc4[(i * 2) * 32 + lane], c1 = sc4[(i * 2 + 1) * 32 + lane];
      const float4 h0 = sh4[(i * 2) * 32 + lane], h1 = sh4[(i * 2 + 1) * 32 + lane];
      const float sc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
      const float sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        if (live[j]) {
          float f[8];
          v[j][i].get(f);
          Vec8<OutT> o;
#pragma unroll
          for (int e = 0; e < 8; ++e) o.v[e] = fmaf((f[e] - mu[j]) * rstd[j], sc[e], sh[e]);
          o.store(out + (size_t)(sample * rows_per_sample + rl0 + j) * D + (lane + i * 32) * 8);
        }
      }
    }
  }
  }
}

// --------------------------------------------------------------------------------------------
// final layer: LN + modulate + Linear(D -> P) (+ unpatchify). One warp per token.
// --------------------------------------------------------------------------------------------
template <typename InT>
__global__ void __launch_bounds__(256) final_layer_kernel(const InT* __restrict__ x, int R, int T, int D,
                                                          const float* __restrict__ shift, const float* __restrict__ scale,
                                                          long ld_mod, const float* __restrict__ w, const float* __restrict__ b,
                                                          int P, float* __restrict__ out, int nchw, int C, int H, int W, int p) {
  const long tok = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (tok >= (long)R * T) return;
  const int r = (int)(tok / T), tk = (int)(tok - (long)r * T);
  const int nvec = D >> 3;
  const InT* xr = x + (size_t)tok * D;
  Vec8<InT> v[LN_MAXV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
      v[i] = Vec8<InT>::load(xr + vi * 8);
#pragma unroll
      for (int e = 0; e < 8; ++e) s += v[i].v[e];
    }
  }
  const float mu = warp_sum(s) / (float)D;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
#pragma unroll
      for (int e = 0; e < 8; ++e) { float d = v[i].v[e] - mu; q = fmaf(d, d, q); }
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)D + 1e-6f);
  const size_t mrow = (size_t)r * ld_mod;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
      Vec8<float> sh = Vec8<float>::load(shift + mrow + vi * 8);
      Vec8<float> sc = Vec8<float>::load(scale + mrow + vi * 8);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[i].v[e] = fmaf((v[i].v[e] - mu) * rstd, 1.0f + sc.v[e], sh.v[e]);
    }
  }
  const int nw = nchw ? (W / p) : 0;
  for (int o = 0; o < P; ++o) {
    const float* wr = w + (size_t)o * D;
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i) {
      int vi = lane + i * 32;
      if (vi < nvec) {
        Vec8<float> wv = Vec8<float>::load(wr + vi * 8);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc = fmaf(v[i].v[e], wv.v[e], acc);
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      float val = acc + b[o];
      if (nchw) {
        int ih = tk / nw, iw = tk - ih * nw;
        int c = o % C, pp = o / C;
        int ph = pp / p, pw = pp - ph * p;
        out[(((size_t)r * C + c) * H + (ih * p + ph)) * W + (iw * p + pw)] = val;
      } else {
        out[(size_t)tok * P + o] = val;
      }
    }
  }
}

// The same with the [P, D] weight staged ONCE per CTA in shared memory (lane-contiguous float4 layout) and persistent CTAs that walk the
// tokens two per warp and round: the one-warp-per-token kernel above re-reads the 74 KB weight from L1 / L2 for every token (9.6 GB at the
// headline shape: 1.03 ms for a 0.3 GB stream = 0.05 of the HBM roofline).  Used when P * D * 4 bytes fit (p = 2: P = 16).
template <typename InT, int NV>
__global__ void __launch_bounds__(256, 2) final_layer_smem_kernel(const InT* __restrict__ x, int R, int T, int D,
                                                                 const float* __restrict__ shift, const float* __restrict__ scale,
                                                                 long ld_mod, const float* __restrict__ w, const float* __restrict__ b,
                                                                 int P, float* __restrict__ out, int nchw, int C, int H, int W, int p) {
  extern __shared__ __align__(16) float4 fl_w4[];   // [P][NV][2][32]: lane l's two float4 of vector (l + 32 i) of output row o
  const int nvec = D >> 3;
  for (int idx = threadIdx.x; idx < P * (D >> 2); idx += blockDim.x) {
    const int o = idx / (D >> 2), q4 = idx - o * (D >> 2);      // q4: float4 index inside row o
    const int vi = q4 >> 1, hf = q4 & 1;
    fl_w4[((o * NV + (vi >> 5)) * 2 + hf) * 32 + (vi & 31)] = *reinterpret_cast<const float4*>(w + (size_t)o * D + q4 * 4);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long ntok = (long)R * T;
  const long nwarps = (long)gridDim.x * 8;
  const int nw = nchw ? (W / p) : 0;
  for (long t0 = ((long)blockIdx.x * 8 + warp) * 2; t0 < ntok; t0 += nwarps * 2) {
    float v[2][NV][8];
    bool live[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const long tok = t0 + j;
      live[j] = tok < ntok;
      const InT* xr = x + (size_t)(live[j] ? tok : t0) * D;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vi = lane + i * 32;
        if (vi < nvec) {
          const Vec8<InT> a = Vec8<InT>::load(xr + vi * 8);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[j][i][e] = a.v[e];
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) v[j][i][e] = 0.f;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const long tok = live[j] ? t0 + j : t0;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int e = 0; e < 8; ++e) s += v[j][i][e];
      const float mu = warp_sum(s) / (float)D;
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        if (lane + i * 32 < nvec) {
#pragma unroll
          for (int e = 0; e < 8; ++e) { const float d = v[j][i][e] - mu; q = fmaf(d, d, q); }
        }
      }
      const float rstd = rsqrtf(warp_sum(q) / (float)D + 1e-6f);
      const size_t mrow = (size_t)(tok / T) * ld_mod;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int vi = lane + i * 32;
        if (vi < nvec) {
          const Vec8<float> sh = Vec8<float>::load(shift + mrow + vi * 8);
          const Vec8<float> sc = Vec8<float>::load(scale + mrow + vi * 8);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[j][i][e] = fmaf((v[j][i][e] - mu) * rstd, 1.0f + sc.v[e], sh.v[e]);
        }
      }
    }
    for (int o = 0; o < P; ++o) {
      float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        if (lane + i * 32 < nvec) {
          const float4 wa = fl_w4[((o * NV + i) * 2 + 0) * 32 + lane], wb = fl_w4[((o * NV + i) * 2 + 1) * 32 + lane];
          const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
          for (int e = 0; e < 8; ++e) { acc0 = fmaf(v[0][i][e], wv[e], acc0); acc1 = fmaf(v[1][i][e], wv[e], acc1); }
        }
      }
      acc0 = warp_sum(acc0); acc1 = warp_sum(acc1);
      if (lane < 2 && live[lane]) {
        const long tok = t0 + lane;
        const float val = (lane == 0 ? acc0 : acc1) + b[o];
        if (nchw) {
          const int r = (int)(tok / T), tk = (int)(tok - (long)r * T);
          const int ih = tk / nw, iw = tk - ih * nw;
          const int c = o % C, pp = o / C;
          const int ph = pp / p, pw = pp - ph * p;
          out[(((size_t)r * C + c) * H + (ih * p + ph)) * W + (iw * p + pw)] = val;
        } else {
          out[(size_t)tok * P + o] = val;
        }
      }
    }
  }
}

// --------------------------------------------------------------------------------------------
// fused CFG combine + DDPM / DDIM step
// --------------------------------------------------------------------------------------------
template <bool DDIM>
__global__ void __launch_bounds__(256) cfg_step_kernel(const float* __restrict__ x, const float* __restrict__ mo,
                                                       const float* __restrict__ noise, float* __restrict__ sample,
                                                       float* __restrict__ pred_xstart, float* __restrict__ mean_out,
                                                       int rows, long per_row, int cfg_pairs, float cfg_scale, float eta,
                                                       int flags, const int64_t* __restrict__ t,
                                                       const float* __restrict__ coef, int S) {
  const long nvec = (long)rows * per_row / 4;
  const bool clip = flags & FITB200_STEP_CLIP_DENOISED, start_x = flags & FITB200_STEP_PREDICT_XSTART;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long)gridDim.x * blockDim.x) {
    const long e0 = i * 4;
    const int row = (int)(e0 / per_row);
    const long off = e0 - (long)row * per_row;
    long ti = t[row];
    ti = ti < 0 ? 0 : (ti >= S ? S - 1 : ti);
    const float* cf = coef + ti * 8;
    const float c0 = cf[0], c1 = cf[1], c2 = cf[2], c3 = cf[3], logvar = cf[4];
    const float4 xv = *reinterpret_cast<const float4*>(x + e0);
    float4 nv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (noise) nv = *reinterpret_cast<const float4*>(noise + e0);
    float4 ev;
    if (cfg_pairs > 0) {
      const int pair = row % cfg_pairs;
      const float4 cv = *reinterpret_cast<const float4*>(mo + (size_t)pair * per_row + off);
      const float4 uv = *reinterpret_cast<const float4*>(mo + (size_t)(pair + cfg_pairs) * per_row + off);
      // fit.py:551: half_eps = uncond_eps + cfg_scale * (cond_eps - uncond_eps)
      ev.x = __fadd_rn(uv.x, __fmul_rn(cfg_scale, __fsub_rn(cv.x, uv.x)));
      ev.y = __fadd_rn(uv.y, __fmul_rn(cfg_scale, __fsub_rn(cv.y, uv.y)));
      ev.z = __fadd_rn(uv.z, __fmul_rn(cfg_scale, __fsub_rn(cv.z, uv.z)));
      ev.w = __fadd_rn(uv.w, __fmul_rn(cfg_scale, __fsub_rn(cv.w, uv.w)));
    } else {
      ev = *reinterpret_cast<const float4*>(mo + e0);
    }
    const float nz = (ti != 0) ? 1.f : 0.f;
    float xs[4] = {xv.x, xv.y, xv.z, xv.w}, es[4] = {ev.x, ev.y, ev.z, ev.w}, ns[4] = {nv.x, nv.y, nv.z, nv.w};
    float x0[4], out[4], mean[4];
    // no fma contraction across the reference's separate multiplies/adds: keep its rounding
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      // _predict_xstart_from_eps (gaussian_diffusion.py:346-351) or the START_X branch (:328-331)
      x0[k] = start_x ? es[k] : __fsub_rn(__fmul_rn(c0, xs[k]), __fmul_rn(c1, es[k]));
      if (clip) x0[k] = fminf(fmaxf(x0[k], -1.f), 1.f);  // process_xstart (:322-326)
      mean[k] = __fadd_rn(__fmul_rn(c2, x0[k]), __fmul_rn(c3, xs[k]));  // q_posterior_mean_variance (:253-256)
    }
    if (!DDIM) {
      const float sig = nz * expf(0.5f * logvar);
#pragma unroll
      for (int k = 0; k < 4; ++k) out[k] = __fadd_rn(mean[k], __fmul_rn(sig, ns[k]));
    } else {
      const float ab = cf[5], abp = cf[6];
      const float sigma = eta * sqrtf((1.f - abp) / (1.f - ab)) * sqrtf(1.f - ab / abp);
      const float sa = sqrtf(abp), sb = sqrtf(1.f - abp - sigma * sigma);
      const float sn = nz * sigma;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float e2 = __fdiv_rn(__fsub_rn(__fmul_rn(c0, xs[k]), x0[k]), c1);  // _predict_eps_from_xstart (:353-356)
        float mp = __fadd_rn(__fmul_rn(x0[k], sa), __fmul_rn(sb, e2));
        out[k] = __fadd_rn(mp, __fmul_rn(sn, ns[k]));
      }
    }
    if (sample) *reinterpret_cast<float4*>(sample + e0) = make_float4(out[0], out[1], out[2], out[3]);
    if (pred_xstart) *reinterpret_cast<float4*>(pred_xstart + e0) = make_float4(x0[0], x0[1], x0[2], x0[3]);
    if (mean_out) *reinterpret_cast<float4*>(mean_out + e0) = make_float4(mean[0], mean[1], mean[2], mean[3]);
  }
}

// forward_with_cfg's combine alone (fit.py:549-553): out[r] = u + s (c - u) for both halves
__global__ void __launch_bounds__(256) cfg_combine_kernel(const float* __restrict__ mo, float* __restrict__ out, int pairs,
                                                          long per_row, float s) {
  const long nvec = (long)pairs * per_row / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long)gridDim.x * blockDim.x) {
    const long e0 = i * 4;
    const float4 cv = *reinterpret_cast<const float4*>(mo + e0);
    const float4 uv = *reinterpret_cast<const float4*>(mo + (size_t)pairs * per_row + e0);
    float4 o;
    o.x = __fadd_rn(uv.x, __fmul_rn(s, __fsub_rn(cv.x, uv.x)));
    o.y = __fadd_rn(uv.y, __fmul_rn(s, __fsub_rn(cv.y, uv.y)));
    o.z = __fadd_rn(uv.z, __fmul_rn(s, __fsub_rn(cv.z, uv.z)));
    o.w = __fadd_rn(uv.w, __fmul_rn(s, __fsub_rn(cv.w, uv.w)));
    *reinterpret_cast<float4*>(out + e0) = o;
    *reinterpret_cast<float4*>(out + (size_t)pairs * per_row + e0) = o;
  }
}

__global__ void q_sample_kernel(const float* __restrict__ x0, const float* __restrict__ noise, float* __restrict__ out,
                                int rows, long per_row, const int64_t* __restrict__ t, const float* __restrict__ a,
                                const float* __restrict__ b, int S) {
  const long nvec = (long)rows * per_row / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long)gridDim.x * blockDim.x) {
    const long e0 = i * 4;
    const int row = (int)(e0 / per_row);
    long ti = t[row];
    ti = ti < 0 ? 0 : (ti >= S ? S - 1 : ti);
    const float ca = a[ti], cb = b[ti];
    const float4 xv = *reinterpret_cast<const float4*>(x0 + e0);
    const float4 nv = *reinterpret_cast<const float4*>(noise + e0);
    float4 o;
    o.x = __fadd_rn(__fmul_rn(ca, xv.x), __fmul_rn(cb, nv.x));
    o.y = __fadd_rn(__fmul_rn(ca, xv.y), __fmul_rn(cb, nv.y));
    o.z = __fadd_rn(__fmul_rn(ca, xv.z), __fmul_rn(cb, nv.z));
    o.w = __fadd_rn(__fmul_rn(ca, xv.w), __fmul_rn(cb, nv.w));
    *reinterpret_cast<float4*>(out + e0) = o;
  }
}

// --------------------------------------------------------------------------------------------
// counter-based N(0,1) noise: Philox4x32-10 keyed by (seed, step), counter = global element index / 4,
// Box-Muller.  Keyed by the GLOBAL row index so a batch sharded over GPUs draws the same noise per image.
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

__global__ void __launch_bounds__(256) philox_normal_kernel(float* __restrict__ out, int rows, long per_row, uint64_t seed,
                                                            const int64_t* __restrict__ step_ptr, long row_offset,
                                                            const int64_t* __restrict__ row_offset_ptr) {
  const long nvec = (long)rows * per_row / 4;
  if (row_offset_ptr) row_offset += (long)row_offset_ptr[0];  // device-resident part of the key: a replayed graph serves any batch position
  const uint64_t step = (uint64_t)(step_ptr ? step_ptr[0] : 0);
  const uint32_t k0 = (uint32_t)seed ^ (uint32_t)(step * 0x9E3779B97F4A7C15ull >> 32);
  const uint32_t k1 = (uint32_t)(seed >> 32) ^ (uint32_t)step;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long)gridDim.x * blockDim.x) {
    const uint64_t gi = (uint64_t)(row_offset * (per_row / 4) + i);
    uint32_t c[4] = {(uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)step, 0x46495442u /* "FITB" */};
    philox4x32_10(c, k0, k1);
    float4 o;
    {
      const float u1 = ((float)c[0] + 0.5f) * 2.3283064365386963e-10f, u2 = ((float)c[1] + 0.5f) * 2.3283064365386963e-10f;
      const float r = sqrtf(-2.0f * logf(fminf(u1, 0.99999994f))), a = 6.283185307179586f * u2;
      float sn, cs; sincosf(a, &sn, &cs);
      o.x = r * cs; o.y = r * sn;
    }
    {
      const float u1 = ((float)c[2] + 0.5f) * 2.3283064365386963e-10f, u2 = ((float)c[3] + 0.5f) * 2.3283064365386963e-10f;
      const float r = sqrtf(-2.0f * logf(fminf(u1, 0.99999994f))), a = 6.283185307179586f * u2;
      float sn, cs; sincosf(a, &sn, &cs);
      o.z = r * cs; o.w = r * sn;
    }
    *reinterpret_cast<float4*>(out + i * 4) = o;
  }
}

__global__ void gather_i64_kernel(const int64_t* __restrict__ map, int map_len, const int64_t* __restrict__ idx,
                                  int64_t* __restrict__ out, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  long k = idx[i];
  k = k < 0 ? 0 : (k >= map_len ? map_len - 1 : k);
  out[i] = map[k];
}
__global__ void add_i64_kernel(int64_t* __restrict__ idx, int n, int64_t delta) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) idx[i] += delta;
}

template <typename InT, typename OutT>
static void launch_ln(const void* x, int M, int D, int rows_per_sample, const float* shift, const float* scale, int64_t ld_mod,
                      void* out, float* mean, float* rstd, cudaStream_t s) {
  const int samples = ceil_div(M, rows_per_sample), bps = ceil_div(rows_per_sample, 8 * LnRows<InT>::RPW * LN_ITERS);
  const dim3 grid((unsigned)(samples * bps));
  const int nv = ceil_div(D >> 3, 32);
  const size_t smem = (size_t)nv * 128 * sizeof(float4);
#define FITB200_LN_CASE(NV)                                                                                              \
  case NV:                                                                                                               \
    ln_modulate_kernel<InT, OutT, NV><<<grid, 256, smem, s>>>((const InT*)x, M, D, rows_per_sample, bps, shift, scale, ld_mod, \
                                                              (OutT*)out, mean, rstd);                                   \
    break;
  switch (nv) {
    FITB200_LN_CASE(1) FITB200_LN_CASE(2) FITB200_LN_CASE(3) FITB200_LN_CASE(4)
    FITB200_LN_CASE(5) FITB200_LN_CASE(6) FITB200_LN_CASE(7) FITB200_LN_CASE(8)
  }
#undef FITB200_LN_CASE
}

// ---------------------------------------------------------------------------------------------
// fitb200_ln_modulate_fwd_static: scratch that lives in the library image instead of a caller / allocator buffer
// ---------------------------------------------------------------------------------------------
// Partial row sums (sum x, sum x^2) of up to FITB200_STATIC_LN_MAX_ROWS rows: 1 MiB per GPU that loads the library, held until exit.
// Every call shares it (it keeps the last call's contents; nothing in it needs to start at zero: row_stats overwrites each entry the
// following ln_apply reads).  Calls that overlap on one GPU (two streams) race on it.
static __device__ __align__(256) float2 g_static_ln_parts[FITB200_STATIC_LN_MAX_ROWS];

// The array's name is a host-side stand-in in host code, not a device address: its address is looked up once per device, with that device
// current, and the POINTER is what the kernels receive.
static int static_ln_parts_ptr(float2** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static float2* ptrs[MAX_DEV] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) { set_last_error("ln_modulate_fwd_static: cudaGetDevice failed or device index %d out of range", dev); return FITB200_ERR_CUDA; }
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    float2* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_static_ln_parts);
    if (e != cudaSuccess || ptr == nullptr) { set_last_error("ln_modulate_fwd_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    ptrs[dev] = ptr;
  }
  *out = ptrs[dev];
  return FITB200_OK;
}

}  // namespace fitb200

using namespace fitb200;

// ---------------------------------------------------------------------------------------------
// pieces of the GEMM LayerNorm + modulate prologue (gemm_tc.cu, LNA)
// ---------------------------------------------------------------------------------------------
// parts[m] = (sum_k x[m,k], sum_k x[m,k]^2): one warp per row (the rows the embedder wrote; every later residual stream gets its sums from
// the GATE_RESID GEMM epilogue)
template <typename InT>
__global__ void __launch_bounds__(256) row_stats_kernel(const InT* __restrict__ x, int M, int D, float2* __restrict__ parts, int nparts) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= M) return;
  const InT* xr = x + (size_t)row * D;
  float s1 = 0.f, s2 = 0.f;
  for (int v = lane; v < (D >> 3); v += 32) {
    const Vec8<InT> a = Vec8<InT>::load(xr + v * 8);
#pragma unroll
    for (int e = 0; e < 8; ++e) { s1 += a.v[e]; s2 = fmaf(a.v[e], a.v[e], s2); }
  }
  s1 = warp_sum(s1); s2 = warp_sum(s2);
  if (lane < nparts) parts[(size_t)row * nparts + lane] = lane == 0 ? make_float2(s1, s2) : make_float2(0.f, 0.f);
}

// modulate(LayerNorm(x)) for rows whose (sum, sum of squares) are already known -- fitb200_gemm_args.stats_out of the GATE_RESID GEMM that wrote
// x, or row_stats: a pure streaming pass (16 bytes in, 16 bytes out per thread and step, no row held in registers, no reduction), 32 rows per CTA.
constexpr int LNA_ROWS = 32;
template <typename InT, typename OutT>
__global__ void __launch_bounds__(256) ln_apply_kernel(const InT* __restrict__ x, int M, int D, int rows_per_sample, const float2* __restrict__ parts,
                                                       int nparts, float eps, const float* __restrict__ shift, const float* __restrict__ scale,
                                                       long ld_mod, OutT* __restrict__ out) {
  __shared__ float2 st[LNA_ROWS];   // (rstd, -mean * rstd)
  // rows are walked from the END of the tensor (see ln_modulate_kernel): the producer GEMM wrote its last rows most recently
  const int row0 = (gridDim.x - 1 - blockIdx.x) * LNA_ROWS;
  if (threadIdx.x < LNA_ROWS) {
    const int row = row0 + threadIdx.x;
    if (row < M) {
      float s1 = 0.f, s2 = 0.f;
      for (int i = 0; i < nparts; ++i) { const float2 t = __ldg(parts + (size_t)row * nparts + i); s1 += t.x; s2 += t.y; }
      const float mu = s1 / (float)D;
      const float rstd = 1.0f / sqrtf(fmaxf(s2 / (float)D - mu * mu, 0.f) + eps);
      st[threadIdx.x] = make_float2(rstd, -mu * rstd);
    }
  }
  __syncthreads();
  const int nvec = D >> 3;
  const int total = min(LNA_ROWS, M - row0) * nvec;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int rl = i / nvec, c = i - rl * nvec;
    const int row = row0 + rl;
    const size_t mrow = (size_t)(row / rows_per_sample) * ld_mod + c * 8;
    Vec8<InT> v = Vec8<InT>::load(x + (size_t)row * D + c * 8);
    const Vec8<float> sc = Vec8<float>::load(scale + mrow), sh = Vec8<float>::load(shift + mrow);
    const float2 ab = st[rl];
    Vec8<OutT> o;
#pragma unroll
    for (int e = 0; e < 8; ++e) o.v[e] = fmaf(fmaf(v.v[e], ab.x, ab.y), 1.0f + sc.v[e], sh.v[e]);
    o.store(out + (size_t)row * D + c * 8);
  }
}

// FiT.unpatchify (fit.py:444-456) as an index map: tokens [R, T, P = p*p*C] -> latents [R, C, H, W] (the tensor-core final layer writes tokens)
__global__ void __launch_bounds__(256) unpatchify_kernel(const float* __restrict__ tok, int R, int T, int C, int H, int W, int p, float* __restrict__ out) {
  const long n = (long)R * C * H * W;
  const int nw = W / p, P = p * p * C;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W), y = (int)((i / W) % H), c = (int)((i / ((long)W * H)) % C), r = (int)(i / ((long)W * H * C));
    const int tk = (y / p) * nw + x / p, pp = (y % p) * p + x % p;
    out[i] = tok[((size_t)r * T + tk) * P + pp * C + c];
  }
}

// out[set][s][k] = bf16(1 + scale_set[s,k]),  out[set][S + s][k] = bf16(shift_set[s,k])
__global__ void __launch_bounds__(256) modulation_pack_kernel(const float* __restrict__ shift, const float* __restrict__ scale, long ld_mod, int S, int D,
                                                           int n_sets, long set_stride, __nv_bfloat16* __restrict__ out) {
  const int dv = D >> 2;
  const long total = (long)n_sets * 2 * S * dv;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int v = (int)(i % dv);
    long r = i / dv;
    const int s2 = (int)(r % (2 * S));
    const int set = (int)(r / (2 * S));
    const bool is_shift = s2 >= S;
    const int smp = is_shift ? s2 - S : s2;
    const float* src = (is_shift ? shift : scale) + (size_t)smp * ld_mod + set * set_stride + v * 4;
    float4 f = *reinterpret_cast<const float4*>(src);
    if (!is_shift) { f.x += 1.0f; f.y += 1.0f; f.z += 1.0f; f.w += 1.0f; }
    uint2 o;
    o.x = pack_bf16x2(f.x, f.y); o.y = pack_bf16x2(f.z, f.w);
    *reinterpret_cast<uint2*>(out + ((size_t)set * 2 * S + s2) * D + v * 4) = o;
  }
}

extern "C" {

const char* fitb200_last_error(void) { return g_err; }
int fitb200_version(void) { return 100; }

int fitb200_set_sm_limit(int n) {
  FITB200_CHECK_ARG(n >= 0, "set_sm_limit: n must be >= 0 (0 = every SM)");
  sm_limit_ref() = n & ~1;  // even: the GEMM schedules CTA pairs
  return FITB200_OK;
}

int fitb200_device_ok(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10;
}

int fitb200_mask_to_kvlen(const uint8_t* mask, int R, int T, int32_t* kv_len, int32_t* n_valid, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(mask && kv_len && R > 0 && T > 0, "mask_to_kvlen: bad arguments");
  mask_to_kvlen_kernel<<<ceil_div(R, 4), 128, 0, (cudaStream_t)stream>>>(mask, R, T, kv_len, n_valid);
  FITB200_CHECK_LAUNCH("mask_to_kvlen");
  return FITB200_OK;
}

int fitb200_embed_tokens_fwd(const float* x, int x_is_nchw, int R, int R_src, int C, int H, int W, int p, int T,
                             const float* w, const float* b, int D, void* out, int out_dtype, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && w && b && out && R > 0 && R_src > 0 && T > 0 && D > 0 && p > 0 && C > 0, "embed_tokens: bad arguments");
  const int K = p * p * C;
  FITB200_CHECK_ARG(K % 4 == 0 && K <= 1024, "embed_tokens: p*p*C=%d must be a multiple of 4 and <= 1024", K);
  if (x_is_nchw) FITB200_CHECK_ARG(H % p == 0 && W % p == 0 && (H / p) * (W / p) == T, "embed_tokens: T=%d != (H/p)*(W/p)", T);
  FITB200_CHECK_ARG(D % EMB_COLS == 0, "embed_tokens: D=%d must be a multiple of 4", D);
  dim3 grid((unsigned)ceil_div((long)R * T, (long)EMB_TOK), (unsigned)ceil_div(D, 128 * EMB_COLS));
  size_t smem = (size_t)EMB_TOK * K * sizeof(float);
  FITB200_DISPATCH_ACT(out_dtype, OutT, {
    if (smem > 48 * 1024) cudaFuncSetAttribute(embed_tokens_kernel<OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    embed_tokens_kernel<OutT><<<grid, 128, smem, (cudaStream_t)stream>>>(x, x_is_nchw, R, R_src, C, H, W, p, T, w, b, D, (OutT*)out);
  })
  FITB200_CHECK_LAUNCH("embed_tokens");
  return FITB200_OK;
}

int fitb200_timestep_sincos(const int64_t* t, int R, const float* freqs, float* out, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(t && out && R > 0, "timestep_sincos: bad arguments");
  timestep_sincos_kernel<<<ceil_div(R * 128, 256), 256, 0, (cudaStream_t)stream>>>(t, R, freqs, out);
  FITB200_CHECK_LAUNCH("timestep_sincos");
  return FITB200_OK;
}

int fitb200_cond_combine_fwd(const float* t_emb, int t_rows, const int64_t* y, const float* table, int table_rows, int R,
                             int D, float* c, void* silu_c, int act_dtype, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(t_emb && y && table && silu_c && R > 0 && D > 0 && (t_rows == 1 || t_rows == R), "cond_combine: bad arguments");
  long n = (long)R * D;
  FITB200_DISPATCH_ACT(act_dtype, ActT, {
    cond_combine_kernel<ActT><<<(unsigned)ceil_div(n, 256L), 256, 0, (cudaStream_t)stream>>>(t_emb, t_rows, y, table, table_rows, R, D, c, (ActT*)silu_c);
  })
  FITB200_CHECK_LAUNCH("cond_combine");
  return FITB200_OK;
}

int fitb200_ln_modulate_fwd(const void* x, int x_dtype, int M, int D, int rows_per_sample, const float* shift,
                            const float* scale, int64_t ld_mod, void* out, int out_dtype, float* mean, float* rstd,
                            fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && shift && scale && out && M > 0 && rows_per_sample > 0, "ln_modulate: bad arguments");
  FITB200_CHECK_ARG(D % 8 == 0 && D <= 32 * 8 * LN_MAXV && ld_mod % 4 == 0, "ln_modulate: D=%d must be a multiple of 8 and <= %d", D, 32 * 8 * LN_MAXV);
  FITB200_CHECK_ARG(((uintptr_t)shift & 15) == 0 && ((uintptr_t)scale & 15) == 0, "ln_modulate: shift / scale must be 16-byte aligned");
  FITB200_DISPATCH_ACT(x_dtype, InT, {
    FITB200_DISPATCH_ACT(out_dtype, OutT, {
      launch_ln<InT, OutT>(x, M, D, rows_per_sample, shift, scale, ld_mod, out, mean, rstd, (cudaStream_t)stream);
    })
  })
  FITB200_CHECK_LAUNCH("ln_modulate");
  return FITB200_OK;
}

int fitb200_final_layer_fwd(const void* x, int x_dtype, int R, int T, int D, const float* shift, const float* scale,
                            int64_t ld_mod, const float* w, const float* b, int P, float* out, int out_nchw, int C,
                            int H, int W, int p, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && shift && scale && w && b && out && R > 0 && T > 0 && P > 0, "final_layer: bad arguments");
  FITB200_CHECK_ARG(D % 8 == 0 && D <= 32 * 8 * LN_MAXV && ld_mod % 4 == 0, "final_layer: D=%d must be a multiple of 8 and <= %d", D, 32 * 8 * LN_MAXV);
  if (out_nchw) FITB200_CHECK_ARG(p > 0 && C > 0 && P % C == 0 && P == p * p * C && H % p == 0 && W % p == 0 && (H / p) * (W / p) == T, "final_layer: inconsistent unpatchify shape");
  dim3 grid((unsigned)ceil_div((long)R * T, 8L));
  const int nv = ceil_div(D >> 3, 32);
  const size_t wbytes = (size_t)P * nv * 2 * 32 * sizeof(float4);
  static const bool smem_ok = [] { const char* e = getenv("FITB200_FINAL_SMEM"); return !(e && e[0] == '0'); }();
  if (smem_ok && wbytes <= 100 * 1024 && ((uintptr_t)w & 15) == 0 && (long)R * T >= 4096) {
    const int ctas = (int)std::min<long>(2L * sm_count(), ceil_div((long)R * T, 16L));
#define FITB200_FINAL_CASE(NVV)                                                                                                     \
    case NVV: {                                                                                                                     \
      FITB200_DISPATCH_ACT(x_dtype, InT, {                                                                                          \
        auto kern = final_layer_smem_kernel<InT, NVV>;                                                                              \
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wbytes);                                       \
        kern<<<ctas, 256, wbytes, (cudaStream_t)stream>>>((const InT*)x, R, T, D, shift, scale, ld_mod, w, b, P, out, out_nchw, C, H, W, p); \
      })                                                                                                                            \
    } break;
    switch (nv) {
      FITB200_FINAL_CASE(1) FITB200_FINAL_CASE(2) FITB200_FINAL_CASE(3) FITB200_FINAL_CASE(4) FITB200_FINAL_CASE(5)
      default: set_last_error("final_layer: D=%d too wide", D); return FITB200_ERR_UNSUPPORTED;
    }
#undef FITB200_FINAL_CASE
    FITB200_CHECK_LAUNCH("final_layer");
    return FITB200_OK;
  }
  FITB200_DISPATCH_ACT(x_dtype, InT, {
    final_layer_kernel<InT><<<grid, 256, 0, (cudaStream_t)stream>>>((const InT*)x, R, T, D, shift, scale, ld_mod, w, b, P, out, out_nchw, C, H, W, p);
  })
  FITB200_CHECK_LAUNCH("final_layer");
  return FITB200_OK;
}

static int step_grid(long nvec) {
  long want = ceil_div(nvec, 256L);
  long cap = (long)sm_count() * 8;
  return (int)(want < cap ? want : cap);
}

static int launch_step(bool ddim, const float* x, const float* model_out, const float* noise, float* sample,
                       float* pred_xstart, float* mean, int rows, int64_t per_row, int cfg_pairs, float cfg_scale,
                       float eta, int flags, const int64_t* t, const float* coef, int S, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && model_out && t && coef && rows > 0 && per_row > 0 && S > 0, "diffusion step: bad arguments");
  FITB200_CHECK_ARG(sample || pred_xstart || mean, "diffusion step: no output requested");
  FITB200_CHECK_ARG(per_row % 4 == 0, "diffusion step: per-row element count %lld must be a multiple of 4", (long long)per_row);
  FITB200_CHECK_ARG(cfg_pairs == 0 || rows == 2 * cfg_pairs || rows == cfg_pairs, "diffusion step: cfg_pairs must be 0, rows or rows/2");
  long nvec = (long)rows * per_row / 4;
  if (ddim)
    cfg_step_kernel<true><<<step_grid(nvec), 256, 0, (cudaStream_t)stream>>>(x, model_out, noise, sample, pred_xstart, mean, rows, per_row, cfg_pairs, cfg_scale, eta, flags, t, coef, S);
  else
    cfg_step_kernel<false><<<step_grid(nvec), 256, 0, (cudaStream_t)stream>>>(x, model_out, noise, sample, pred_xstart, mean, rows, per_row, cfg_pairs, cfg_scale, 0.f, flags, t, coef, S);
  FITB200_CHECK_LAUNCH("diffusion step");
  return FITB200_OK;
}

int fitb200_cfg_ddpm_step(const float* x, const float* model_out, const float* noise, float* sample, float* pred_xstart,
                          float* mean, int rows, int64_t per_row, int cfg_pairs, float cfg_scale, int flags,
                          const int64_t* t, const float* coef, int S, fitb200_stream_t stream) {
  return launch_step(false, x, model_out, noise, sample, pred_xstart, mean, rows, per_row, cfg_pairs, cfg_scale, 0.f, flags, t, coef, S, stream);
}

int fitb200_cfg_ddim_step(const float* x, const float* model_out, const float* noise, float* sample, float* pred_xstart,
                          int rows, int64_t per_row, int cfg_pairs, float cfg_scale, float eta, int flags,
                          const int64_t* t, const float* coef, int S, fitb200_stream_t stream) {
  return launch_step(true, x, model_out, noise, sample, pred_xstart, nullptr, rows, per_row, cfg_pairs, cfg_scale, eta, flags, t, coef, S, stream);
}

int fitb200_cfg_combine(const float* model_out, float* out, int pairs, int64_t per_row, float cfg_scale,
                        fitb200_stream_t stream) {
  FITB200_CHECK_ARG(model_out && out && pairs > 0 && per_row > 0 && per_row % 4 == 0, "cfg_combine: bad arguments");
  long nvec = (long)pairs * per_row / 4;
  cfg_combine_kernel<<<step_grid(nvec), 256, 0, (cudaStream_t)stream>>>(model_out, out, pairs, per_row, cfg_scale);
  FITB200_CHECK_LAUNCH("cfg_combine");
  return FITB200_OK;
}

int fitb200_q_sample(const float* x0, const float* noise, float* out, int rows, int64_t per_row, const int64_t* t,
                     const float* sqrt_ac, const float* sqrt_1m_ac, int S, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x0 && noise && out && t && sqrt_ac && sqrt_1m_ac && rows > 0 && per_row > 0 && S > 0, "q_sample: bad arguments");
  FITB200_CHECK_ARG(per_row % 4 == 0, "q_sample: per-row element count must be a multiple of 4");
  long nvec = (long)rows * per_row / 4;
  q_sample_kernel<<<step_grid(nvec), 256, 0, (cudaStream_t)stream>>>(x0, noise, out, rows, per_row, t, sqrt_ac, sqrt_1m_ac, S);
  FITB200_CHECK_LAUNCH("q_sample");
  return FITB200_OK;
}

int fitb200_philox_normal(float* out, int rows, int64_t per_row, uint64_t seed, const int64_t* step, int64_t row_offset, const int64_t* row_offset_dev,
                          fitb200_stream_t stream) {
  FITB200_CHECK_ARG(out && rows > 0 && per_row > 0 && per_row % 4 == 0 && row_offset >= 0, "philox_normal: bad arguments");
  long nvec = (long)rows * per_row / 4;
  philox_normal_kernel<<<step_grid(nvec), 256, 0, (cudaStream_t)stream>>>(out, rows, per_row, seed, step, row_offset, row_offset_dev);
  FITB200_CHECK_LAUNCH("philox_normal");
  return FITB200_OK;
}

int fitb200_gather_i64(const int64_t* map, int map_len, const int64_t* idx, int64_t* out, int n, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(map && idx && out && n > 0 && map_len > 0, "gather_i64: bad arguments");
  gather_i64_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(map, map_len, idx, out, n);
  FITB200_CHECK_LAUNCH("gather_i64");
  return FITB200_OK;
}

int fitb200_add_i64(int64_t* idx, int n, int64_t delta, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(idx && n > 0, "add_i64: bad arguments");
  add_i64_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(idx, n, delta);
  FITB200_CHECK_LAUNCH("add_i64");
  return FITB200_OK;
}

int fitb200_row_stats(const void* x, int x_dtype, int M, int D, float* parts, int nparts, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && parts && M > 0 && D > 0 && D % 8 == 0 && nparts >= 1 && nparts <= 32 && ((uintptr_t)parts & 7) == 0, "row_stats: bad arguments");
  FITB200_DISPATCH_ACT(x_dtype, InT, {
    row_stats_kernel<InT><<<(unsigned)ceil_div(M, 8), 256, 0, (cudaStream_t)stream>>>((const InT*)x, M, D, (float2*)parts, nparts);
  })
  FITB200_CHECK_LAUNCH("row_stats");
  return FITB200_OK;
}

int fitb200_modulation_pack(const float* shift, const float* scale, int64_t ld_mod, int S, int D, int n_sets, int64_t set_stride, void* out,
                            fitb200_stream_t stream) {
  FITB200_CHECK_ARG(shift && scale && out && S > 0 && D > 0 && n_sets > 0, "modulation_pack: bad arguments");
  FITB200_CHECK_ARG(D % 4 == 0 && ld_mod % 4 == 0 && set_stride % 4 == 0 && ((uintptr_t)shift & 15) == 0 && ((uintptr_t)scale & 15) == 0 && ((uintptr_t)out & 7) == 0,
                    "modulation_pack: D, ld_mod, set_stride must be multiples of 4 and the tables 16-byte aligned");
  const long total = (long)n_sets * 2 * S * (D / 4);
  modulation_pack_kernel<<<step_grid(total), 256, 0, (cudaStream_t)stream>>>(shift, scale, ld_mod, S, D, n_sets, set_stride, (__nv_bfloat16*)out);
  FITB200_CHECK_LAUNCH("modulation_pack");
  return FITB200_OK;
}

int fitb200_ln_apply_fwd(const void* x, int x_dtype, int M, int D, int rows_per_sample, const float* parts, int nparts, float eps,
                         const float* shift, const float* scale, int64_t ld_mod, void* out, int out_dtype, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && parts && shift && scale && out && M > 0 && rows_per_sample > 0 && nparts >= 1, "ln_apply: bad arguments");
  FITB200_CHECK_ARG(D % 8 == 0 && ld_mod % 4 == 0 && ((uintptr_t)shift & 15) == 0 && ((uintptr_t)scale & 15) == 0 && ((uintptr_t)parts & 7) == 0,
                    "ln_apply: D=%d must be a multiple of 8, shift / scale 16-byte aligned", D);
  const unsigned grid = (unsigned)ceil_div(M, LNA_ROWS);
  FITB200_DISPATCH_ACT(x_dtype, InT, {
    FITB200_DISPATCH_ACT(out_dtype, OutT, {
      ln_apply_kernel<InT, OutT><<<grid, 256, 0, (cudaStream_t)stream>>>((const InT*)x, M, D, rows_per_sample, (const float2*)parts, nparts, eps, shift,
                                                                        scale, ld_mod, (OutT*)out);
    })
  })
  FITB200_CHECK_LAUNCH("ln_apply");
  return FITB200_OK;
}

int fitb200_unpatchify(const float* tokens, int R, int T, int C, int H, int W, int p, float* out, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(tokens && out && R > 0 && C > 0 && p > 0 && H % p == 0 && W % p == 0 && (H / p) * (W / p) == T, "unpatchify: inconsistent shape");
  unpatchify_kernel<<<step_grid((long)R * C * H * W), 256, 0, (cudaStream_t)stream>>>(tokens, R, T, C, H, W, p, out);
  FITB200_CHECK_LAUNCH("unpatchify");
  return FITB200_OK;
}

int fitb200_ln_modulate_fwd_static(const void* x, int x_dtype, int M, int D, int rows_per_sample, const float* shift, const float* scale,
                                   int64_t ld_mod, void* out, int out_dtype, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && shift && scale && out && M > 0 && rows_per_sample > 0, "ln_modulate_fwd_static: bad arguments");
  FITB200_CHECK_ARG(M <= FITB200_STATIC_LN_MAX_ROWS, "ln_modulate_fwd_static: M=%d rows exceed the static scratch of %d rows (use fitb200_ln_modulate_fwd)",
                    M, FITB200_STATIC_LN_MAX_ROWS);
  FITB200_CHECK_ARG(D % 8 == 0 && ld_mod % 4 == 0 && ((uintptr_t)shift & 15) == 0 && ((uintptr_t)scale & 15) == 0,
                    "ln_modulate_fwd_static: D=%d must be a multiple of 8, shift / scale 16-byte aligned", D);
  float2* parts = nullptr;
  const int rc = static_ln_parts_ptr(&parts);
  if (rc != FITB200_OK) return rc;
  const cudaStream_t s = (cudaStream_t)stream;
  FITB200_DISPATCH_ACT(x_dtype, InT, {
    row_stats_kernel<InT><<<(unsigned)ceil_div(M, 8), 256, 0, s>>>((const InT*)x, M, D, parts, 1);
    FITB200_DISPATCH_ACT(out_dtype, OutT, {
      ln_apply_kernel<InT, OutT><<<(unsigned)ceil_div(M, LNA_ROWS), 256, 0, s>>>((const InT*)x, M, D, rows_per_sample, parts, 1, 1e-6f, shift, scale,
                                                                                ld_mod, (OutT*)out);
    })
  })
  FITB200_CHECK_LAUNCH("ln_modulate_fwd_static");
  return FITB200_OK;
}

}  // extern "C"
