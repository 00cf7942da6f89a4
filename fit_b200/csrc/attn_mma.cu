// bf16 fused attention forward: 2-D RoPE applied to q/k in registers on the way from HBM to shared
// memory, flash-style online softmax, per-sample key-range tile skipping (kv_len) + mask bytes.
// Tensor path: mma.sync m16n8k16 (bf16 in, fp32 accumulate).  head_dim 64, or 72 zero-padded to 80
// for the q.k^T contraction only (p.v runs on the exact 72 columns).
#include "common.cuh"

namespace fitb200 {

int attn_fwd_f32(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask,
                 int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s);

constexpr int AM_BQ = 64, AM_BK = 64, AM_THREADS = 128;

__device__ __forceinline__ void ldsm_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(a));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(a));
}
__device__ __forceinline__ void ldsm_x2_t(uint32_t& r0, uint32_t& r1, const void* p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];\n" : "=r"(r0), "=r"(r1) : "r"(a));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// load 8 consecutive bf16 of a q/k row, rotate the 4 (re, im) pairs by the fp32 table, round to bf16
__device__ __forceinline__ uint4 load_rotate8(const __nv_bfloat16* p, const float* f) {
  uint4 raw = *reinterpret_cast<const uint4*>(p);
  float4 f0 = *reinterpret_cast<const float4*>(f), f1 = *reinterpret_cast<const float4*>(f + 4);
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
  const float cs[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
  uint4 o;
  uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 ab = __bfloat1622float2(h2[i]);
    float c = cs[2 * i], s = cs[2 * i + 1];
    ow[i] = pack_bf16x2(ab.x * c - ab.y * s, ab.y * c + ab.x * s);  // complex_mult, fit.py:152-154
  }
  return o;
}

template <int HD>
__global__ void __launch_bounds__(AM_THREADS) attn_mma_kernel(const __nv_bfloat16* __restrict__ qkv,
                                                              const float* __restrict__ freqs, int freqs_rows,
                                                              const int32_t* __restrict__ kv_len,
                                                              const uint8_t* __restrict__ mask, int R, int Tn, int h,
                                                              __nv_bfloat16* __restrict__ out, float* __restrict__ lse) {
  constexpr int HDP = (HD + 15) / 16 * 16;  // contraction width of q.k^T (72 -> 80)
  constexpr int LDS = HDP + 8;              // smem row stride in elements (conflict-free ldmatrix)
  constexpr int KSTEPS = HDP / 16;
  constexpr int NT_O = HD / 8;              // output n-tiles
  constexpr int CH = HD / 8;                // 16-byte chunks per row

  __shared__ __align__(16) __nv_bfloat16 Qs[AM_BQ * LDS];
  __shared__ __align__(16) __nv_bfloat16 Ks[AM_BK * LDS];
  __shared__ __align__(16) __nv_bfloat16 Vs[AM_BK * LDS];
  __shared__ uint8_t kvalid[AM_BK];

  const int r = blockIdx.z, head = blockIdx.y, q0 = blockIdx.x * AM_BQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, tq4 = lane & 3;
  const int kvl = min(kv_len[r], Tn);
  const size_t row_stride = (size_t)3 * h * HD;
  const __nv_bfloat16* base = qkv + (size_t)r * Tn * row_stride + (size_t)head * HD;
  const float* fr = freqs + (size_t)(freqs_rows == 1 ? 0 : r) * Tn * HD;
  const int D = h * HD;

  if (q0 >= kvl) {  // padded query tile: zero output rows, no work
    for (int i = tid; i < AM_BQ * CH; i += AM_THREADS) {
      int row = i / CH, c = i - row * CH;
      if (q0 + row < Tn)
        *reinterpret_cast<uint4*>(out + ((size_t)r * Tn + q0 + row) * D + head * HD + c * 8) = make_uint4(0, 0, 0, 0);
    }
    if (lse) for (int i = tid; i < AM_BQ; i += AM_THREADS) if (q0 + i < Tn) lse[((size_t)r * h + head) * Tn + q0 + i] = 0.f;
    return;
  }

  // ---- Q tile: HBM -> rotate in registers -> smem
  for (int i = tid; i < AM_BQ * (LDS / 8); i += AM_THREADS) {
    int row = i / (LDS / 8), c = i - row * (LDS / 8);
    uint4 v = make_uint4(0, 0, 0, 0);
    int t = q0 + row;
    if (c < CH && t < Tn) v = load_rotate8(base + (size_t)t * row_stride + c * 8, fr + (size_t)t * HD + c * 8);
    *reinterpret_cast<uint4*>(Qs + row * LDS + c * 8) = v;
  }
  __syncthreads();

  // Q A-fragments stay in registers for the whole kernel
  uint32_t qf[KSTEPS][4];
#pragma unroll
  for (int ks = 0; ks < KSTEPS; ++ks) {
    const __nv_bfloat16* p = Qs + (warp * 16 + (lane & 15)) * LDS + ks * 16 + (lane >> 4) * 8;
    ldsm_x4(qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], p);
  }

  float o[NT_O][4];
#pragma unroll
  for (int n = 0; n < NT_O; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const float sl2 = rsqrtf((float)HD) * 1.4426950408889634f;  // softmax scale folded with log2(e)

  for (int k0 = 0; k0 < kvl; k0 += AM_BK) {
    __syncthreads();  // previous tile fully consumed
    for (int i = tid; i < AM_BK * (LDS / 8); i += AM_THREADS) {
      int row = i / (LDS / 8), c = i - row * (LDS / 8);
      uint4 kv4 = make_uint4(0, 0, 0, 0), vv4 = make_uint4(0, 0, 0, 0);
      int t = k0 + row;
      if (c < CH && t < kvl) {
        const __nv_bfloat16* kp = base + (size_t)t * row_stride + (size_t)D + c * 8;
        kv4 = load_rotate8(kp, fr + (size_t)t * HD + c * 8);
        vv4 = *reinterpret_cast<const uint4*>(kp + D);
      }
      *reinterpret_cast<uint4*>(Ks + row * LDS + c * 8) = kv4;
      *reinterpret_cast<uint4*>(Vs + row * LDS + c * 8) = vv4;
    }
    if (tid < AM_BK) {
      int t = k0 + tid;
      kvalid[tid] = (t < kvl) && (mask == nullptr || mask[(size_t)r * Tn + t] != 0);
    }
    __syncthreads();

    // ---- S = Q K^T  (16 x 64 per warp)
    float s[AM_BK / 8][4];
#pragma unroll
    for (int n = 0; n < AM_BK / 8; ++n) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
#pragma unroll
      for (int np = 0; np < AM_BK / 16; ++np) {
        // two key n-tiles (8 keys each) x 16 contraction columns in one ldmatrix.x4
        uint32_t b0, b1, b2, b3;
        const __nv_bfloat16* p = Ks + (np * 16 + (lane & 7) + ((lane >> 4) << 3)) * LDS + ks * 16 + ((lane >> 3) & 1) * 8;
        ldsm_x4(b0, b1, b2, b3, p);
        mma_bf16(s[2 * np], qf[ks], b0, b1);
        mma_bf16(s[2 * np + 1], qf[ks], b2, b3);
      }
    }

    // ---- mask + online softmax (rows g and g+8 of this warp's 16)
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < AM_BK / 8; ++n) {
      const int c = n * 8 + tq4 * 2;
      const bool v0 = kvalid[c], v1 = kvalid[c + 1];
      s[n][0] = v0 ? s[n][0] * sl2 : -INFINITY;
      s[n][1] = v1 ? s[n][1] * sl2 : -INFINITY;
      s[n][2] = v0 ? s[n][2] * sl2 : -INFINITY;
      s[n][3] = v1 ? s[n][3] * sl2 : -INFINITY;
      mx0 = fmaxf(mx0, fmaxf(s[n][0], s[n][1]));
      mx1 = fmaxf(mx1, fmaxf(s[n][2], s[n][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    // all keys of the tile masked -> mn = -inf: keep everything at zero without producing NaN
    const float ms0 = mn0 == -INFINITY ? 0.f : mn0, ms1 = mn1 == -INFINITY ? 0.f : mn1;
    const float c0 = exp2f(m0 - ms0), c1 = exp2f(m1 - ms1);
    m0 = mn0; m1 = mn1;
    float rs0 = 0.f, rs1 = 0.f;
    uint32_t pf[AM_BK / 8][2];
#pragma unroll
    for (int n = 0; n < AM_BK / 8; ++n) {
      float p0 = exp2f(s[n][0] - ms0), p1 = exp2f(s[n][1] - ms0);
      float p2 = exp2f(s[n][2] - ms1), p3 = exp2f(s[n][3] - ms1);
      rs0 += p0 + p1; rs1 += p2 + p3;
      pf[n][0] = pack_bf16x2(p0, p1);
      pf[n][1] = pack_bf16x2(p2, p3);
    }
    l0 = l0 * c0 + rs0; l1 = l1 * c1 + rs1;
#pragma unroll
    for (int n = 0; n < NT_O; ++n) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }

    // ---- O += P V
#pragma unroll
    for (int kk = 0; kk < AM_BK / 16; ++kk) {
      const uint32_t a[4] = {pf[2 * kk][0], pf[2 * kk][1], pf[2 * kk + 1][0], pf[2 * kk + 1][1]};
#pragma unroll
      for (int np = 0; np < NT_O / 2; ++np) {
        uint32_t b0, b1, b2, b3;
        const __nv_bfloat16* p = Vs + (kk * 16 + (lane & 15)) * LDS + np * 16 + (lane >> 4) * 8;
        ldsm_x4_t(b0, b1, b2, b3, p);
        mma_bf16(o[2 * np], a, b0, b1);
        mma_bf16(o[2 * np + 1], a, b2, b3);
      }
      if (NT_O & 1) {
        uint32_t b0, b1;
        const __nv_bfloat16* p = Vs + (kk * 16 + (lane & 15)) * LDS + (NT_O - 1) * 8;
        ldsm_x2_t(b0, b1, p);
        mma_bf16(o[NT_O - 1], a, b0, b1);
      }
    }
  }

  // ---- finalise: row sums across the quad, normalise, stage through this warp's Q rows, 16-byte stores
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = l0 > 0.f ? 1.f / l0 : 0.f, i1 = l1 > 0.f ? 1.f / l1 : 0.f;
  __nv_bfloat16* stage = Qs + warp * 16 * LDS;
#pragma unroll
  for (int n = 0; n < NT_O; ++n) {
    *reinterpret_cast<uint32_t*>(stage + g * LDS + n * 8 + tq4 * 2) = pack_bf16x2(o[n][0] * i0, o[n][1] * i0);
    *reinterpret_cast<uint32_t*>(stage + (g + 8) * LDS + n * 8 + tq4 * 2) = pack_bf16x2(o[n][2] * i1, o[n][3] * i1);
  }
  __syncwarp();
  for (int i = lane; i < 16 * CH; i += 32) {
    int row = i / CH, c = i - row * CH;
    int t = q0 + warp * 16 + row;
    if (t < Tn)
      *reinterpret_cast<uint4*>(out + ((size_t)r * Tn + t) * D + head * HD + c * 8) =
          *reinterpret_cast<const uint4*>(stage + row * LDS + c * 8);
  }
  if (lse && tq4 == 0) {
    int t0 = q0 + warp * 16 + g, t1 = t0 + 8;
    float* lp = lse + ((size_t)r * h + head) * Tn;
    if (t0 < Tn) lp[t0] = l0 > 0.f ? (m0 + log2f(l0)) * 0.6931471805599453f : 0.f;
    if (t1 < Tn) lp[t1] = l1 > 0.f ? (m1 + log2f(l1)) * 0.6931471805599453f : 0.f;
  }
}

template <int HD>
static int launch_attn_mma(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len,
                           const uint8_t* mask, int R, int Tn, int h, void* out, float* lse, cudaStream_t s) {
  dim3 grid((unsigned)ceil_div(Tn, AM_BQ), (unsigned)h, (unsigned)R);
  attn_mma_kernel<HD><<<grid, AM_THREADS, 0, s>>>((const __nv_bfloat16*)qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h,
                                                  (__nv_bfloat16*)out, lse);
  FITB200_CHECK_LAUNCH("attn_mma");
  return FITB200_OK;
}

}  // namespace fitb200

using namespace fitb200;

extern "C" int fitb200_attn_fwd(const void* qkv, int dtype, const float* freqs, int freqs_rows, const int32_t* kv_len,
                                const uint8_t* mask, int R, int T, int h, int d, void* out, float* lse,
                                fitb200_stream_t stream) {
  FITB200_CHECK_ARG(qkv && freqs && kv_len && out && R > 0 && T > 0 && h > 0, "attn_fwd: bad arguments");
  FITB200_CHECK_ARG(freqs_rows == 1 || freqs_rows == R, "attn_fwd: freqs_rows must be 1 or R");
  FITB200_CHECK_ARG(h <= 65535 && R <= 65535, "attn_fwd: h and R must be <= 65535");
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == FITB200_F32) return attn_fwd_f32(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, d, out, lse, s);
  if (dtype == FITB200_BF16) {
    if (d == 64) return launch_attn_mma<64>(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, out, lse, s);
    if (d == 72) return launch_attn_mma<72>(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, out, lse, s);
    set_last_error("attn_fwd: head_dim %d not supported (64 or 72)", d);
    return FITB200_ERR_UNSUPPORTED;
  }
  set_last_error("attn_fwd: unsupported dtype %d", dtype);
  return FITB200_ERR_INVALID_ARG;
}
