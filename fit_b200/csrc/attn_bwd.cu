// Attention backward (autograd of apply_rotary_emb + Attention.forward, fit.py:160-188), CUDA-core version: exact fp32 math for
// the verification mode, also used for bf16 activations (fp32 accumulation).
//
//   P = exp(q.k^T / sqrt(d) - lse)            (keys masked exactly as in the forward)
//   dV = P^T dO      dP = dO V^T      delta_i = sum_d dO_id O_id      dS = P (dP - delta) / sqrt(d)
//   dQ = dS K        dK = dS^T Q      and the RoPE transposed (rotation by -theta) on dQ, dK   (complex_mult, fit.py:144-155)
//
// Two passes over 32 x 32 (query, key) tiles, each recomputing S and dP, so that no atomics are needed and the result is
// deterministic:  ROLE 0: CTA = 32 keys, loops over query tiles, accumulates dK / dV;  ROLE 1: CTA = 32 queries, loops over key
// tiles, accumulates dQ.  The gradient is written for the UN-rotated q / k (what the qkv Linear's backward consumes), whether or
// not the forward operands in `qkv` were pre-rotated by the GEMM epilogue.
#include "common.cuh"

namespace fitb200 {

int attn_bwd_tc(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d,
                const void* o, const void* dO, const float* lse, void* dqkv, cudaStream_t s);

constexpr int AB_T = 32;         // tile edge
constexpr int AB_THREADS = 256;
constexpr int AB_QSKIP = 128;    // the forward skips (zero-fills) query tiles of 128 starting at or beyond kv_len

template <int HD, typename T, int ROLE>
__global__ void __launch_bounds__(AB_THREADS) attn_bwd_simt_kernel(const T* __restrict__ qkv, const float* __restrict__ freqs, int freqs_rows,
                                                                   const int32_t* __restrict__ kv_len, const uint8_t* __restrict__ mask,
                                                                   int R, int Tn, int h, int prerotated, const T* __restrict__ o,
                                                                   const T* __restrict__ dO, const float* __restrict__ lse,
                                                                   T* __restrict__ dqkv) {
  constexpr int LD = HD + 1;
  __shared__ float Qs[AB_T][LD], Ks[AB_T][LD], Vs[AB_T][LD], Ds[AB_T][LD];  // Ds = dO tile
  __shared__ float Ps[AB_T][AB_T + 1], Ss[AB_T][AB_T + 1];                  // P and dS tiles
  __shared__ float lse_s[AB_T], delta_s[AB_T];
  __shared__ int kval[AB_T];

  const int r = blockIdx.z, head = blockIdx.y, tile0 = blockIdx.x * AB_T;  // tile0: first key (ROLE 0) or first query (ROLE 1)
  const int tid = threadIdx.x;
  const int D = h * HD;
  const size_t rs3 = (size_t)3 * D;
  const T* qbase = qkv + (size_t)r * Tn * rs3 + (size_t)head * HD;
  const T* kbase = qbase + D;
  const T* vbase = qbase + 2 * D;
  const float* fr = freqs + (size_t)(freqs_rows == 1 ? 0 : r) * Tn * HD;
  const int kvl = min(kv_len[r], Tn);
  const int q_end = min(Tn, ceil_div(kvl, AB_QSKIP) * AB_QSKIP);  // queries that took part in the forward
  const float scale = rsqrtf((float)HD);

  // load a [32 x HD] tile of rows t0.. from `src` (row stride rs elements), optionally rotated, zero beyond `tend`
  auto load_tile = [&](float (*dst)[LD], const T* src, size_t rs, int t0, int tend, bool rotate) {
    for (int i = tid; i < AB_T * HD; i += AB_THREADS) {
      const int row = i / HD, dd = i - row * HD, t = t0 + row;
      float v = 0.f;
      if (t < tend) {
        v = to_f32(src[(size_t)t * rs + dd]);
        if (rotate) {
          const float w = to_f32(src[(size_t)t * rs + (dd ^ 1)]);
          const float c = fr[(size_t)t * HD + (dd & ~1)], s = fr[(size_t)t * HD + (dd | 1)];
          v = (dd & 1) ? v * c + w * s : v * c - w * s;
        }
      }
      dst[row][dd] = v;
    }
  };
  // rotate a gradient tile by -theta and store it: out[2j] = g[2j] c + g[2j+1] s, out[2j+1] = g[2j+1] c - g[2j] s
  auto store_unrotated = [&](float (*g)[LD], T* dst, int t0) {
    for (int i = tid; i < AB_T * HD; i += AB_THREADS) {
      const int row = i / HD, dd = i - row * HD, t = t0 + row;
      if (t < Tn) {
        const float c = fr[(size_t)t * HD + (dd & ~1)], s = fr[(size_t)t * HD + (dd | 1)];
        const float a = g[row][dd], b = g[row][dd ^ 1];
        dst[(size_t)t * rs3 + dd] = from_f32<T>((dd & 1) ? a * c - b * s : a * c + b * s);
      }
    }
  };

  const int ty = tid >> 4, tx = tid & 15;  // S / dP micro-tile: queries 2 ty + {0,1}, keys 2 tx + {0,1}
  constexpr int NACC = (AB_T * HD + AB_THREADS - 1) / AB_THREADS;
  float acc0[NACC] = {}, acc1[NACC] = {};  // ROLE 0: dK, dV   ROLE 1: dQ (acc0)

  if (ROLE == 0) {
    load_tile(Ks, kbase, rs3, tile0, kvl, !prerotated);
    load_tile(Vs, vbase, rs3, tile0, kvl, false);
    if (tid < AB_T) { const int k = tile0 + tid; kval[tid] = k < kvl && (mask == nullptr || mask[(size_t)r * Tn + k] != 0); }
  } else {
    load_tile(Qs, qbase, rs3, tile0, q_end, !prerotated);
    load_tile(Ds, (const T*)dO + (size_t)r * Tn * D + head * HD, D, tile0, q_end, false);
    // delta and lse of this query tile
    load_tile(Vs, (const T*)o + (size_t)r * Tn * D + head * HD, D, tile0, q_end, false);  // O tile, temporarily in Vs
    __syncthreads();
    if (tid < AB_T) {
      float dl = 0.f;
      for (int dd = 0; dd < HD; ++dd) dl = fmaf(Ds[tid][dd], Vs[tid][dd], dl);
      delta_s[tid] = dl;
      lse_s[tid] = tile0 + tid < Tn ? lse[((size_t)r * h + head) * Tn + tile0 + tid] : 0.f;
    }
  }
  __syncthreads();

  const bool outer_live = ROLE == 0 ? tile0 < kvl : tile0 < q_end;
  const int inner_end = ROLE == 0 ? q_end : kvl;
  if (outer_live) {
    for (int in0 = 0; in0 < inner_end; in0 += AB_T) {
      __syncthreads();
      if (ROLE == 0) {
        load_tile(Qs, qbase, rs3, in0, q_end, !prerotated);
        load_tile(Ds, (const T*)dO + (size_t)r * Tn * D + head * HD, D, in0, q_end, false);
      } else {
        load_tile(Ks, kbase, rs3, in0, kvl, !prerotated);
        load_tile(Vs, vbase, rs3, in0, kvl, false);
        if (tid < AB_T) { const int k = in0 + tid; kval[tid] = k < kvl && (mask == nullptr || mask[(size_t)r * Tn + k] != 0); }
      }
      __syncthreads();
      if (ROLE == 0) {  // delta / lse of the inner query tile: O rows are read straight from global (one row per thread)
        if (tid < AB_T) {
          const int t = in0 + tid;
          float dl = 0.f;
          if (t < q_end) {
            const T* orow = o + ((size_t)r * Tn + t) * D + head * HD;
            for (int dd = 0; dd < HD; ++dd) dl = fmaf(Ds[tid][dd], to_f32(orow[dd]), dl);
          }
          delta_s[tid] = dl;
          lse_s[tid] = t < Tn ? lse[((size_t)r * h + head) * Tn + t] : 0.f;
        }
        __syncthreads();
      }
      // S and dP micro-tiles
      float s[2][2] = {}, dp[2][2] = {};
      for (int dd = 0; dd < HD; ++dd) {
        const float q0 = Qs[2 * ty][dd], q1 = Qs[2 * ty + 1][dd], d0 = Ds[2 * ty][dd], d1 = Ds[2 * ty + 1][dd];
        const float k0 = Ks[2 * tx][dd], k1 = Ks[2 * tx + 1][dd], v0 = Vs[2 * tx][dd], v1 = Vs[2 * tx + 1][dd];
        s[0][0] = fmaf(q0, k0, s[0][0]); s[0][1] = fmaf(q0, k1, s[0][1]); s[1][0] = fmaf(q1, k0, s[1][0]); s[1][1] = fmaf(q1, k1, s[1][1]);
        dp[0][0] = fmaf(d0, v0, dp[0][0]); dp[0][1] = fmaf(d0, v1, dp[0][1]); dp[1][0] = fmaf(d1, v0, dp[1][0]); dp[1][1] = fmaf(d1, v1, dp[1][1]);
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int ql = 2 * ty + i, kl = 2 * tx + j;
          const int qt = (ROLE == 0 ? in0 : tile0) + ql;
          const bool live = kval[kl] && qt < q_end;
          const float p = live ? expf(s[i][j] * scale - lse_s[ql]) : 0.f;
          Ps[ql][kl] = p;
          Ss[ql][kl] = p * (dp[i][j] - delta_s[ql]) * scale;
        }
      __syncthreads();
      // accumulate
#pragma unroll
      for (int a = 0; a < NACC; ++a) {
        const int e = tid + a * AB_THREADS;
        if (e < AB_T * HD) {
          const int row = e / HD, dd = e - row * HD;
          if (ROLE == 0) {
            float dk = acc0[a], dv = acc1[a];
#pragma unroll 8
            for (int q = 0; q < AB_T; ++q) { dk = fmaf(Ss[q][row], Qs[q][dd], dk); dv = fmaf(Ps[q][row], Ds[q][dd], dv); }
            acc0[a] = dk; acc1[a] = dv;
          } else {
            float dq = acc0[a];
#pragma unroll 8
            for (int k = 0; k < AB_T; ++k) dq = fmaf(Ss[row][k], Ks[k][dd], dq);
            acc0[a] = dq;
          }
        }
      }
    }
  }
  __syncthreads();
  // stage the gradient tiles in shared memory (Qs / Ks are free now) and write them out
#pragma unroll
  for (int a = 0; a < NACC; ++a) {
    const int e = tid + a * AB_THREADS;
    if (e < AB_T * HD) {
      const int row = e / HD, dd = e - row * HD;
      Qs[row][dd] = acc0[a];
      if (ROLE == 0) Ks[row][dd] = acc1[a];
    }
  }
  __syncthreads();
  T* dq_out = dqkv + (size_t)r * Tn * rs3 + (size_t)head * HD;
  if (ROLE == 0) {
    store_unrotated(Qs, dq_out + D, tile0);  // dK
    for (int i = tid; i < AB_T * HD; i += AB_THREADS) {  // dV
      const int row = i / HD, dd = i - row * HD, t = tile0 + row;
      if (t < Tn) dq_out[(size_t)t * rs3 + 2 * D + dd] = from_f32<T>(Ks[row][dd]);
    }
  } else {
    store_unrotated(Qs, dq_out, tile0);      // dQ
  }
}

template <int HD, typename T>
static int launch_attn_bwd(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h,
                           int prerotated, const void* o, const void* dO, const float* lse, void* dqkv, cudaStream_t s) {
  dim3 grid((unsigned)ceil_div(Tn, AB_T), (unsigned)h, (unsigned)R);
  attn_bwd_simt_kernel<HD, T, 0><<<grid, AB_THREADS, 0, s>>>((const T*)qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, prerotated, (const T*)o,
                                                              (const T*)dO, lse, (T*)dqkv);
  attn_bwd_simt_kernel<HD, T, 1><<<grid, AB_THREADS, 0, s>>>((const T*)qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, prerotated, (const T*)o,
                                                              (const T*)dO, lse, (T*)dqkv);
  FITB200_CHECK_LAUNCH("attn_bwd");
  return FITB200_OK;
}

}  // namespace fitb200

using namespace fitb200;

extern "C" int fitb200_attn_bwd(const void* qkv, int dtype, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask,
                                int R, int T, int h, int d, int flags, const void* out, const void* dout, const float* lse, void* dqkv,
                                fitb200_stream_t stream) {
  FITB200_CHECK_ARG(qkv && freqs && kv_len && out && dout && lse && dqkv && R > 0 && T > 0 && h > 0, "attn_bwd: bad arguments");
  FITB200_CHECK_ARG(freqs_rows == 1 || freqs_rows == R, "attn_bwd: freqs_rows must be 1 or R");
  FITB200_CHECK_ARG(h <= 65535 && R <= 65535, "attn_bwd: h and R must be <= 65535");
  const int pre = flags & FITB200_ATTN_PREROTATED;
  cudaStream_t s = (cudaStream_t)stream;
  // bf16 training path (pre-rotated operands, trained sequence lengths): tcgen05 kernel; everything else: CUDA cores
  if (dtype == FITB200_BF16 && pre && T <= 256 && (d == 64 || d == 72) && !(flags & FITB200_ATTN_FORCE_SIMT) &&
      ((uintptr_t)qkv & 15) == 0 && ((uintptr_t)dout & 15) == 0 && ((uintptr_t)out & 15) == 0 && ((uintptr_t)dqkv & 15) == 0)
    return attn_bwd_tc(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, d, out, dout, lse, dqkv, s);
  FITB200_DISPATCH_ACT(dtype, AT, {
    if (d == 64) return launch_attn_bwd<64, AT>(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, pre, out, dout, lse, dqkv, s);
    if (d == 72) return launch_attn_bwd<72, AT>(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, pre, out, dout, lse, dqkv, s);
  })
  set_last_error("attn_bwd: head_dim %d not supported (64 or 72)", d);
  return FITB200_ERR_UNSUPPORTED;
}
