// Elementwise / reduction kernels of the training step (autograd of the FiT block, FiT/train.py:55-66) and the un-fused
// forward pieces the training forward uses so that every tensor the backward needs is materialised once:
//   gate + residual (fit.py:331-337), SwiGLU gating (fit.py:278-280), LayerNorm+modulate backward (fit.py:45-46,332,336),
//   bias-gradient column sums, SiLU, label-embedding scatter (fit.py:114), masked MSE (FiT/train.py:63-66), dtype casts.
// Activations and activation gradients are T (bf16 or fp32); statistics, modulation tables, reductions and parameter gradients
// are fp32.  Reductions over tokens use fp32 atomics into zero-initialised buffers (the caller zeroes them).
#include <cstdlib>

#include "common.cuh"

namespace fitb200 {

// ------------------------------------------------------------------------------------------------ gate + residual forward
// out[m, :] = resid[m, :] + gate[m / rps, :] * y[m, :]
template <typename T>
__global__ void gate_resid_fwd_kernel(const T* __restrict__ resid, const T* __restrict__ y, const float* __restrict__ gate, long ld_gate,
                                      int M, int D, int rps, T* __restrict__ out) {
  const int nv = D >> 3;
  const long total = (long)M * nv;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int m = (int)(i / nv), c = (int)(i - (long)m * nv) * 8;
    const Vec8<T> r = Vec8<T>::load(resid + (size_t)m * D + c), yy = Vec8<T>::load(y + (size_t)m * D + c);
    const Vec8<float> g = Vec8<float>::load(gate + (size_t)(m / rps) * ld_gate + c);
    Vec8<T> o;
#pragma unroll
    for (int e = 0; e < 8; ++e) o.v[e] = fmaf(g.v[e], yy.v[e], r.v[e]);
    o.store(out + (size_t)m * D + c);
  }
}

// ------------------------------------------------------------------------------------------------ gate backward
// dy = gate[s] * dx;  dgate[s, :] += sum_t dx * y;  dbias[:] += sum_m dy.
// One CTA = (256 columns, one sample, ALL its rows): 32 column groups of 8 x 8 row lanes; the row lanes are reduced through shared
// memory, so dgate needs no atomics at all and dbias one per (sample, column).  (The first version split the rows into chunks that
// each ended in atomicAdds on the same addresses: 2.4 ms per XL/2 step, and 4.7 ms with three times the chunks.)
constexpr int CR_LANES = 8;   // row lanes per CTA of the column-reduction kernels
template <typename T>
__global__ void __launch_bounds__(256) gate_bwd_kernel(const T* __restrict__ dx, const T* __restrict__ y, const float* __restrict__ gate,
                                                       long ld_gate, int D, int rps, T* __restrict__ dy, float* __restrict__ dgate,
                                                       long ld_dgate, float* __restrict__ dbias) {
  __shared__ float red[CR_LANES][32][17];
  const int cgl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int cg = blockIdx.x * 32 + cgl, c = cg * 8, s = blockIdx.y;
  const bool live = cg < (D >> 3);
  float ag[8] = {}, ab[8] = {};
  if (live) {
    const Vec8<float> g = Vec8<float>::load(gate + (size_t)s * ld_gate + c);
#pragma unroll 4
    // gridDim.z row chunks per sample (when one CTA per (256 columns, sample) leaves most SMs idle: 160 CTAs at 32 samples x 1152 columns ran at
    // 18 % of the HBM rate); the chunks' partial sums then meet in atomicAdds, gridDim.z per address
    const int t_lo = (int)((long)rps * blockIdx.z / gridDim.z), t_hi = (int)((long)rps * (blockIdx.z + 1) / gridDim.z);
    for (int t = t_lo + rl; t < t_hi; t += CR_LANES) {
      const size_t off = ((size_t)s * rps + t) * D + c;
      const Vec8<T> d = Vec8<T>::load(dx + off), yy = Vec8<T>::load(y + off);
      Vec8<T> o;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        ag[e] = fmaf(d.v[e], yy.v[e], ag[e]);
        o.v[e] = g.v[e] * d.v[e];
        ab[e] += o.v[e];
      }
      o.store(dy + off);
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) { red[rl][cgl][e] = ag[e]; red[rl][cgl][8 + e] = ab[e]; }
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * 16; i += 256) {
    const int g2 = i >> 4, e = i & 15, col = (blockIdx.x * 32 + g2) * 8 + (e & 7);
    if (col < D) {
      float v = 0.f;
#pragma unroll
      for (int r = 0; r < CR_LANES; ++r) v += red[r][g2][e];
      if (e < 8) {
        if (gridDim.z == 1) dgate[(size_t)s * ld_dgate + col] += v;   // this CTA owns (sample, column): plain read-modify-write
        else atomicAdd(dgate + (size_t)s * ld_dgate + col, v);
      } else if (dbias) atomicAdd(dbias + col, v);
    }
  }
}

// sigmoid for the SwiGLU kernels: exact expf / division for fp32 storage (verification mode), the fast intrinsics for bf16 storage (their
// error, ~2 ulp fp32, is far below the bf16 rounding of the result; with expf + IEEE division these kernels were issue-bound: 75 % issue
// slots busy at 41 % of the HBM rate)
template <typename T> __device__ __forceinline__ float sigmoid_t(float x) {
  if constexpr (sizeof(T) == 2) return __fdividef(1.0f, 1.0f + __expf(-x));
  else return 1.0f / (1.0f + expf(-x));
}

// ------------------------------------------------------------------------------------------------ SwiGLU forward / backward
// u [M, 2*Hp]: every 32 columns = 16 gate pre-activations then the matching 16 up pre-activations (the packed fc1 weight).
template <typename T>
__global__ void swiglu_fwd_kernel(const T* __restrict__ u, int M, int Hp, T* __restrict__ g) {
  const int nv = Hp >> 3;
  const long total = (long)M * nv;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int m = (int)(i / nv), j = (int)(i - (long)m * nv);           // j: group of 8 hidden units
    const size_t uo = (size_t)m * 2 * Hp + (j >> 1) * 32 + (j & 1) * 8;  // gate columns; up columns are 16 further
    const Vec8<T> a = Vec8<T>::load(u + uo), b = Vec8<T>::load(u + uo + 16);
    Vec8<T> o;
#pragma unroll
    for (int e = 0; e < 8; ++e) o.v[e] = a.v[e] * sigmoid_t<T>(a.v[e]) * b.v[e];
    o.store(g + (size_t)m * Hp + j * 8);
  }
}

// du_gate = dg * up * silu'(gate), du_up = dg * silu(gate);  dbias[2*Hp] += column sums.  grid = (col groups, row chunks)
template <typename T>
__global__ void swiglu_bwd_kernel(const T* __restrict__ dg, const T* __restrict__ u, int M, int Hp, int rows_per_chunk,
                                  T* __restrict__ du, float* __restrict__ dbias) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= (Hp >> 3)) return;
  const int m0 = blockIdx.y * rows_per_chunk, m1 = min(M, m0 + rows_per_chunk);
  const int uc = (j >> 1) * 32 + (j & 1) * 8;
  float sa[8] = {}, sb[8] = {};
#pragma unroll 4
  for (int m = m0; m < m1; ++m) {
    const size_t uo = (size_t)m * 2 * Hp + uc;
    const Vec8<T> a = Vec8<T>::load(u + uo), b = Vec8<T>::load(u + uo + 16), d = Vec8<T>::load(dg + (size_t)m * Hp + j * 8);
    Vec8<T> oa, ob;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float sg = sigmoid_t<T>(a.v[e]);
      const float si = a.v[e] * sg;
      oa.v[e] = d.v[e] * b.v[e] * (sg * (1.0f + a.v[e] * (1.0f - sg)));
      ob.v[e] = d.v[e] * si;
      sa[e] += oa.v[e]; sb[e] += ob.v[e];
    }
    oa.store(du + uo); ob.store(du + uo + 16);
  }
  if (dbias) {
#pragma unroll
    for (int e = 0; e < 8; ++e) { atomicAdd(dbias + uc + e, sa[e]); atomicAdd(dbias + uc + 16 + e, sb[e]); }
  }
}

// ------------------------------------------------------------------------------------------------ LayerNorm + modulate backward
// rows: dxhat = dh * (1 + scale[s]);  dx = rstd * (dxhat - mean(dxhat) - xhat * mean(dxhat * xhat)) + dres.  One warp per row.
// 8 consecutive elements as loaded (16 bytes of bf16 / 32 bytes of fp32): rows are kept packed between the two passes of ln_bwd_rows
template <typename T> struct Pack8;
template <> struct Pack8<__nv_bfloat16> {
  uint4 r;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { r = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void get(float (&f)[8]) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
};
template <> struct Pack8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) { a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 4); }
  __device__ __forceinline__ void get(float (&f)[8]) const { f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w; }
};

// All three inputs of a row (x, dh, dres) are requested up front (one exposed memory latency per row instead of two) and kept PACKED between
// the statistics pass and the output pass (the fp32 copies cost 98 registers: two CTAs per SM, 22 % of the HBM rate).
template <typename T, int NV>
__global__ void __launch_bounds__(256, 2) ln_bwd_rows_kernel(const T* __restrict__ dh, const T* __restrict__ x, const float* __restrict__ mean,
                                                          const float* __restrict__ rstd, const float* __restrict__ scale, long ld_mod,
                                                          int M, int D, int rps, const T* __restrict__ dres, T* __restrict__ dx) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= M) return;
  const int nvec = D >> 3;
  const float* sc = scale + (size_t)(row / rps) * ld_mod;
  constexpr bool EARLY = sizeof(T) == 2;   // (fp32 rows: three packed copies do not fit the register file; dres is loaded in the output pass)
  Pack8<T> px[NV], pd[NV], pr[EARLY ? NV : 1];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      px[i].load(x + (size_t)row * D + vi * 8);
      pd[i].load(dh + (size_t)row * D + vi * 8);
      if (EARLY && dres) pr[i].load(dres + (size_t)row * D + vi * 8);
    }
  }
  const float mu = mean[row], rs = rstd[row];
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      float a[8], d[8];
      px[i].get(a); pd[i].get(d);
      const Vec8<float> s = Vec8<float>::load(sc + vi * 8);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float dxh = d[e] * (1.0f + s.v[e]);
        s1 += dxh;
        s2 = fmaf(dxh, (a[e] - mu) * rs, s2);
      }
    }
  }
  s1 = warp_sum(s1) / (float)D;
  s2 = warp_sum(s2) / (float)D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      float a[8], d[8], r[8];
      px[i].get(a); pd[i].get(d);
      if (dres) {
        if (!EARLY) pr[0].load(dres + (size_t)row * D + vi * 8);
        pr[EARLY ? i : 0].get(r);
      }
      const Vec8<float> s = Vec8<float>::load(sc + vi * 8);   // an L1 hit: read by the statistics pass just above
      Vec8<T> o;
#pragma unroll
      for (int e = 0; e < 8; ++e) o.v[e] = (dres ? r[e] : 0.f) + rs * (d[e] * (1.0f + s.v[e]) - s1 - (a[e] - mu) * rs * s2);
      o.store(dx + (size_t)row * D + vi * 8);
    }
  }
}

// columns: dshift[s, :] += sum_t dh;  dscale[s, :] += sum_t dh * xhat.   Same CTA shape as gate_bwd_kernel: no atomics.
template <typename T>
__global__ void __launch_bounds__(256) mod_bwd_cols_kernel(const T* __restrict__ dh, const T* __restrict__ x, const float* __restrict__ mean,
                                                           const float* __restrict__ rstd, int D, int rps, float* __restrict__ dshift,
                                                           float* __restrict__ dscale, long ld_mod) {
  __shared__ float red[CR_LANES][32][17];
  const int cgl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int cg = blockIdx.x * 32 + cgl, c = cg * 8, s = blockIdx.y;
  const bool live = cg < (D >> 3);
  float a1[8] = {}, a2[8] = {};
  if (live) {
#pragma unroll 4
    const int t_lo = (int)((long)rps * blockIdx.z / gridDim.z), t_hi = (int)((long)rps * (blockIdx.z + 1) / gridDim.z);   // see gate_bwd_kernel
    for (int t = t_lo + rl; t < t_hi; t += CR_LANES) {
      const size_t row = (size_t)s * rps + t;
      const float mu = mean[row], rs = rstd[row];
      const Vec8<T> d = Vec8<T>::load(dh + row * D + c), a = Vec8<T>::load(x + row * D + c);
#pragma unroll
      for (int e = 0; e < 8; ++e) { a1[e] += d.v[e]; a2[e] = fmaf(d.v[e], (a.v[e] - mu) * rs, a2[e]); }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) { red[rl][cgl][e] = a1[e]; red[rl][cgl][8 + e] = a2[e]; }
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * 16; i += 256) {
    const int g2 = i >> 4, e = i & 15, col = (blockIdx.x * 32 + g2) * 8 + (e & 7);
    if (col < D) {
      float v = 0.f;
#pragma unroll
      for (int r = 0; r < CR_LANES; ++r) v += red[r][g2][e];
      float* dst = (e < 8 ? dshift : dscale) + (size_t)s * ld_mod + col;
      if (gridDim.z == 1) *dst += v; else atomicAdd(dst, v);
    }
  }
}

// ------------------------------------------------------------------------------------------------ column sums (bias gradients)
template <typename T>
__global__ void colsum_kernel(const T* __restrict__ a, int M, int N, long lda, int rows_per_chunk, float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int m0 = blockIdx.y * rows_per_chunk, m1 = min(M, m0 + rows_per_chunk);
  float s = 0.f;
#pragma unroll 8
  for (int m = m0; m < m1; ++m) s += to_f32(a[(size_t)m * lda + n]);
  atomicAdd(out + n, s);
}

// The same with 16-byte loads: CTA = 32 column groups of 8 x 8 row lanes over one row chunk, row lanes reduced through shared memory, one
// atomicAdd per (chunk, column).  (The thread-per-column kernel above reads 2 bytes per thread and row: 18 % of the HBM rate on dqkv.)
template <typename T>
__global__ void __launch_bounds__(256) colsum_vec_kernel(const T* __restrict__ a, int M, int N, long lda, int rows_per_chunk, float* __restrict__ out) {
  __shared__ float red[CR_LANES][32][9];
  const int cgl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + cgl) * 8;
  const int m0 = blockIdx.y * rows_per_chunk, m1 = min(M, m0 + rows_per_chunk);
  float acc[8] = {};
  if (c < N) {
#pragma unroll 4
    for (int m = m0 + rl; m < m1; m += CR_LANES) {
      const Vec8<T> v = Vec8<T>::load(a + (size_t)m * lda + c);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += v.v[e];
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[rl][cgl][e] = acc[e];
  __syncthreads();
  const int g2 = threadIdx.x >> 3, e = threadIdx.x & 7, col = (blockIdx.x * 32 + g2) * 8 + e;
  if (col < N) {
    float v = 0.f;
#pragma unroll
    for (int r = 0; r < CR_LANES; ++r) v += red[r][g2][e];
    atomicAdd(out + col, v);
  }
}

// ------------------------------------------------------------------------------------------------ small pointwise pieces
__global__ void silu_fwd_kernel(const float* __restrict__ x, long n, float* __restrict__ y) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) y[i] = silu_acc(x[i]);
}
// dx = dy * silu'(x)
__global__ void silu_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, long n, float* __restrict__ dx) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const float v = x[i], sg = 1.0f / (1.0f + expf(-v));
    dx[i] = dy[i] * sg * (1.0f + v * (1.0f - sg));
  }
}
// dtable[y[r], :] += dc[r, :]   (LabelEmbedder gather backward, fit.py:114)
__global__ void embedding_bwd_kernel(const float* __restrict__ dc, const int64_t* __restrict__ y, int R, int D, int table_rows,
                                     float* __restrict__ dtable) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < (long)R * D; i += (long)gridDim.x * blockDim.x) {
    const int r = (int)(i / D), c = (int)(i - (long)r * D);
    long lab = y[r];
    lab = lab < 0 ? 0 : (lab >= table_rows ? table_rows - 1 : lab);
    atomicAdd(dtable + lab * D + c, dc[i]);
  }
}
template <typename Src, typename Dst>
__global__ void cast_kernel(const Src* __restrict__ a, long n, Dst* __restrict__ b) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) b[i] = from_f32<Dst>(to_f32(a[i]));
}

// ------------------------------------------------------------------------------------------------ masked MSE (FiT/train.py:63-66)
// acc[0] += sum over valid tokens of (out - target)^2, acc[1] += number of valid tokens
__global__ void masked_mse_fwd_kernel(const float* __restrict__ out, const float* __restrict__ target, const uint8_t* __restrict__ mask,
                                      long tokens, int P, float* __restrict__ acc) {
  float s = 0.f, n = 0.f;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < tokens; i += (long)gridDim.x * blockDim.x) {
    if (mask[i]) {
      n += 1.f;
      for (int p = 0; p < P; ++p) { const float d = out[i * P + p] - target[i * P + p]; s = fmaf(d, d, s); }
    }
  }
  s = warp_sum(s); n = warp_sum(n);
  if ((threadIdx.x & 31) == 0) { atomicAdd(acc, s); atomicAdd(acc + 1, n); }
}
__global__ void masked_mse_finish_kernel(const float* __restrict__ acc, int P, float* __restrict__ loss) {
  loss[0] = acc[0] / (acc[1] * (float)P);   // F.mse_loss over out[mask] (mean over valid tokens x P); 0 valid tokens -> NaN like torch
}
// dout = gscale * 2 (out - target) mask / (count * P)
template <typename T>
__global__ void masked_mse_bwd_kernel(const float* __restrict__ out, const float* __restrict__ target, const uint8_t* __restrict__ mask,
                                      long tokens, int P, const float* __restrict__ acc, const float* __restrict__ gscale, T* __restrict__ dout) {
  const float k = 2.0f * (gscale ? gscale[0] : 1.0f) / (acc[1] * (float)P);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < tokens * P; i += (long)gridDim.x * blockDim.x)
    dout[i] = from_f32<T>(mask[i / P] ? k * (out[i] - target[i]) : 0.f);
}

template <typename T>
static void launch_ln_bwd_rows(int nv, const void* dh, const void* x, const float* mean, const float* rstd, const float* scale, long ld_mod,
                               int M, int D, int rps, const void* dres, void* dx, cudaStream_t s) {
#define FITB200_LNB_CASE(NV) \
  case NV: ln_bwd_rows_kernel<T, NV><<<ceil_div(M, 8), 256, 0, s>>>((const T*)dh, (const T*)x, mean, rstd, scale, ld_mod, M, D, rps, (const T*)dres, (T*)dx); break;
  switch (nv) {
    FITB200_LNB_CASE(1) FITB200_LNB_CASE(2) FITB200_LNB_CASE(3) FITB200_LNB_CASE(4)
    FITB200_LNB_CASE(5) FITB200_LNB_CASE(6) FITB200_LNB_CASE(7) FITB200_LNB_CASE(8)
  }
#undef FITB200_LNB_CASE
}

static inline int grid_for(long n, int block, int cap = 148 * 16) {
  long g = (n + block - 1) / block;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}
// enough CTAs for ~4 waves without tiny chunks.  More, smaller chunks were measured slower (gate_bwd 2.4 -> 4.7 ms per XL/2 step): every
// chunk ends in atomicAdds on the same dgate / dbias addresses, and those, not the row loop, bound these kernels.
// row chunks per sample for the column-reduction kernels: enough CTAs for ~4 per SM, at least 4 rows per row lane
static inline int row_chunks(long ctas, int rows_per_sample) {
  static const int force = [] { const char* e = getenv("FITB200_BWD_ROW_CHUNKS"); return e ? atoi(e) : 0; }();
  if (force > 0) return force;
  long nz = (148L * 4 + ctas - 1) / ctas;
  const long cap = rows_per_sample / (CR_LANES * 4);
  if (nz > cap) nz = cap;
  if (nz > 16) nz = 16;
  return nz < 1 ? 1 : (int)nz;
}

static inline int chunk_rows(int rows, long other_ctas) {
  int chunks = (int)((148L * 4 + other_ctas - 1) / other_ctas);
  if (chunks < 1) chunks = 1;
  int rpc = (rows + chunks - 1) / chunks;
  return rpc < 8 ? (rows < 8 ? rows : 8) : rpc;
}

}  // namespace fitb200

using namespace fitb200;

#define BW_STREAM ((cudaStream_t)stream)

extern "C" {

int fitb200_gate_resid_fwd(const void* resid, const void* y, int dtype, const float* gate, int64_t ld_gate, int M, int D,
                           int rows_per_sample, void* out, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(resid && y && gate && out && M > 0 && D % 8 == 0 && rows_per_sample > 0 && ld_gate % 4 == 0, "gate_resid_fwd: bad arguments");
  FITB200_DISPATCH_ACT(dtype, T, {
    gate_resid_fwd_kernel<T><<<grid_for((long)M * (D / 8), 256), 256, 0, BW_STREAM>>>((const T*)resid, (const T*)y, gate, ld_gate, M, D,
                                                                                     rows_per_sample, (T*)out);
  })
  FITB200_CHECK_LAUNCH("gate_resid_fwd");
  return FITB200_OK;
}

int fitb200_gate_bwd(const void* dx, const void* y, int dtype, const float* gate, int64_t ld_gate, int R, int rows_per_sample, int D,
                     void* dy, float* dgate, int64_t ld_dgate, float* dbias, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(dx && y && gate && dy && dgate && R > 0 && rows_per_sample > 0 && D % 8 == 0 && ld_gate % 4 == 0, "gate_bwd: bad arguments");
  dim3 grid(ceil_div(D / 8, 32), R, row_chunks(ceil_div(D / 8, 32) * R, rows_per_sample));
  FITB200_DISPATCH_ACT(dtype, T, {
    gate_bwd_kernel<T><<<grid, 256, 0, BW_STREAM>>>((const T*)dx, (const T*)y, gate, ld_gate, D, rows_per_sample, (T*)dy, dgate, ld_dgate, dbias);
  })
  FITB200_CHECK_LAUNCH("gate_bwd");
  return FITB200_OK;
}

int fitb200_swiglu_fwd(const void* u, int dtype, int M, int Hp, void* g, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(u && g && M > 0 && Hp > 0 && Hp % 16 == 0, "swiglu_fwd: bad arguments (Hp %% 16)");
  FITB200_DISPATCH_ACT(dtype, T, {
    swiglu_fwd_kernel<T><<<grid_for((long)M * (Hp / 8), 256), 256, 0, BW_STREAM>>>((const T*)u, M, Hp, (T*)g);
  })
  FITB200_CHECK_LAUNCH("swiglu_fwd");
  return FITB200_OK;
}

int fitb200_swiglu_bwd(const void* dg, const void* u, int dtype, int M, int Hp, void* du, float* dbias, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(dg && u && du && M > 0 && Hp > 0 && Hp % 16 == 0, "swiglu_bwd: bad arguments (Hp %% 16)");
  const int bx = ceil_div(Hp / 8, 128);
  const int rpc = chunk_rows(M, bx);
  dim3 grid(bx, ceil_div(M, rpc));
  FITB200_DISPATCH_ACT(dtype, T, {
    swiglu_bwd_kernel<T><<<grid, 128, 0, BW_STREAM>>>((const T*)dg, (const T*)u, M, Hp, rpc, (T*)du, dbias);
  })
  FITB200_CHECK_LAUNCH("swiglu_bwd");
  return FITB200_OK;
}

int fitb200_ln_modulate_bwd(const void* dh, const void* x, int dtype, const float* mean, const float* rstd, const float* scale,
                            int64_t ld_mod, int M, int D, int rows_per_sample, const void* dres, void* dx, float* dshift, float* dscale,
                            int64_t ld_dmod, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(dh && x && mean && rstd && scale && dx && M > 0 && rows_per_sample > 0 && M % rows_per_sample == 0, "ln_modulate_bwd: bad arguments");
  FITB200_CHECK_ARG(D % 8 == 0 && D <= 32 * 8 * 8 && ld_mod % 4 == 0, "ln_modulate_bwd: D=%d must be a multiple of 8 and <= 2048", D);
  const int nv = ceil_div(D / 8, 32);
  FITB200_DISPATCH_ACT(dtype, T, {
    launch_ln_bwd_rows<T>(nv, dh, x, mean, rstd, scale, ld_mod, M, D, rows_per_sample, dres, dx, BW_STREAM);
  })
  FITB200_CHECK_LAUNCH("ln_bwd_rows");
  if (dshift && dscale) {
    const int R = M / rows_per_sample;
    dim3 grid(ceil_div(D / 8, 32), R, row_chunks(ceil_div(D / 8, 32) * R, rows_per_sample));
    FITB200_DISPATCH_ACT(dtype, T, {
      mod_bwd_cols_kernel<T><<<grid, 256, 0, BW_STREAM>>>((const T*)dh, (const T*)x, mean, rstd, D, rows_per_sample, dshift, dscale, ld_dmod);
    })
    FITB200_CHECK_LAUNCH("mod_bwd_cols");
  }
  return FITB200_OK;
}

int fitb200_colsum(const void* a, int dtype, int M, int N, int64_t lda, float* out, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(a && out && M > 0 && N > 0, "colsum: bad arguments");
  if (N % 8 == 0 && lda % 8 == 0 && ((uintptr_t)a & 15) == 0 && M >= 64) {
    const int bx = ceil_div(N / 8, 32);
    int rpc = chunk_rows(M, bx);
    rpc = (rpc + CR_LANES - 1) / CR_LANES * CR_LANES;
    if (rpc < 4 * CR_LANES) rpc = 4 * CR_LANES;
    dim3 grid(bx, ceil_div(M, rpc));
    FITB200_DISPATCH_ACT(dtype, T, { colsum_vec_kernel<T><<<grid, 256, 0, BW_STREAM>>>((const T*)a, M, N, lda, rpc, out); })
    FITB200_CHECK_LAUNCH("colsum");
    return FITB200_OK;
  }
  const int bx = ceil_div(N, 128);
  const int rpc = chunk_rows(M, bx);
  dim3 grid(bx, ceil_div(M, rpc));
  FITB200_DISPATCH_ACT(dtype, T, { colsum_kernel<T><<<grid, 128, 0, BW_STREAM>>>((const T*)a, M, N, lda, rpc, out); })
  FITB200_CHECK_LAUNCH("colsum");
  return FITB200_OK;
}

int fitb200_silu_fwd(const float* x, int64_t n, float* y, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(x && y && n > 0, "silu_fwd: bad arguments");
  silu_fwd_kernel<<<grid_for(n, 256), 256, 0, BW_STREAM>>>(x, n, y);
  FITB200_CHECK_LAUNCH("silu_fwd");
  return FITB200_OK;
}

int fitb200_silu_bwd(const float* dy, const float* x, int64_t n, float* dx, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(dy && x && dx && n > 0, "silu_bwd: bad arguments");
  silu_bwd_kernel<<<grid_for(n, 256), 256, 0, BW_STREAM>>>(dy, x, n, dx);
  FITB200_CHECK_LAUNCH("silu_bwd");
  return FITB200_OK;
}

int fitb200_embedding_bwd(const float* dc, const int64_t* y, int R, int D, int table_rows, float* dtable, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(dc && y && dtable && R > 0 && D > 0 && table_rows > 0, "embedding_bwd: bad arguments");
  embedding_bwd_kernel<<<grid_for((long)R * D, 256), 256, 0, BW_STREAM>>>(dc, y, R, D, table_rows, dtable);
  FITB200_CHECK_LAUNCH("embedding_bwd");
  return FITB200_OK;
}

int fitb200_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(src && dst && n > 0, "cast: bad arguments");
  const int g = grid_for(n, 256);
  if (src_dtype == FITB200_F32 && dst_dtype == FITB200_BF16) cast_kernel<float, __nv_bfloat16><<<g, 256, 0, BW_STREAM>>>((const float*)src, n, (__nv_bfloat16*)dst);
  else if (src_dtype == FITB200_BF16 && dst_dtype == FITB200_F32) cast_kernel<__nv_bfloat16, float><<<g, 256, 0, BW_STREAM>>>((const __nv_bfloat16*)src, n, (float*)dst);
  else if (src_dtype == FITB200_F32 && dst_dtype == FITB200_F32) cast_kernel<float, float><<<g, 256, 0, BW_STREAM>>>((const float*)src, n, (float*)dst);
  else { set_last_error("cast: unsupported dtype pair %d -> %d", src_dtype, dst_dtype); return FITB200_ERR_INVALID_ARG; }
  FITB200_CHECK_LAUNCH("cast");
  return FITB200_OK;
}

int fitb200_masked_mse_fwd(const float* out, const float* target, const uint8_t* mask, int64_t tokens, int P, float* acc /* [2], zeroed */,
                           float* loss, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(out && target && mask && acc && loss && tokens > 0 && P > 0, "masked_mse_fwd: bad arguments");
  masked_mse_fwd_kernel<<<grid_for(tokens, 256, 148 * 4), 256, 0, BW_STREAM>>>(out, target, mask, tokens, P, acc);
  masked_mse_finish_kernel<<<1, 1, 0, BW_STREAM>>>(acc, P, loss);
  FITB200_CHECK_LAUNCH("masked_mse_fwd");
  return FITB200_OK;
}

int fitb200_masked_mse_bwd(const float* out, const float* target, const uint8_t* mask, int64_t tokens, int P, const float* acc,
                           const float* gscale, void* dout, int dout_dtype, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(out && target && mask && acc && dout && tokens > 0 && P > 0, "masked_mse_bwd: bad arguments");
  FITB200_DISPATCH_ACT(dout_dtype, T, {
    masked_mse_bwd_kernel<T><<<grid_for(tokens * P, 256), 256, 0, BW_STREAM>>>(out, target, mask, tokens, P, acc, gscale, (T*)dout);
  })
  FITB200_CHECK_LAUNCH("masked_mse_bwd");
  return FITB200_OK;
}

}  // extern "C"
