// bf16 fused attention forward: 2-D RoPE applied to q/k in registers on the way from HBM to shared
// memory, flash-style online softmax, per-sample key-range tile skipping (kv_len) + mask bytes.
// Tensor path: mma.sync m16n8k16 (bf16 in, fp32 accumulate).  head_dim 64, or 72 zero-padded to 80
// for the q.k^T contraction only (p.v runs on the exact 72 columns).
#include "common.cuh"

namespace fitb200 {

int attn_fwd_f32(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask,
                 int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s);
int attn_fwd_tc3(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s);
int attn_fwd_tc(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask,
                int R, int Tn, int h, int d, int prerotated, void* out, float* lse, cudaStream_t s);

constexpr int AM_BQ = 128;      // queries per CTA: 8 warps x 16 rows
constexpr int AM_BK = 64;       // keys per softmax tile
constexpr int AM_CHUNK = 256;   // keys resident in shared memory at a time (the whole sequence when T <= 256)
constexpr int AM_THREADS = 256;

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
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(a), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory"); }

// rotate the 4 (re, im) pairs of 8 consecutive bf16 by the fp32 (cos, sin) table, round to bf16 (complex_mult, fit.py:152-154)
__device__ __forceinline__ uint4 rotate8(const uint4 raw, const float4 f0, const float4 f1) {
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
  const float cs[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
  uint4 o;
  uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 ab = __bfloat1622float2(h2[i]);
    float c = cs[2 * i], s = cs[2 * i + 1];
    ow[i] = pack_bf16x2(ab.x * c - ab.y * s, ab.y * c + ab.x * s);
  }
  return o;
}

// Load `nrows` rows (HD bf16 each, global row stride `gstride` elements) starting at token t0, rotate them by the RoPE
// table and store them as smem rows of LDS elements; rows in [nvalid, nrows) and the pad columns [HD, LDS) are zeroed.
// Loads are issued UNR at a time so each thread keeps UNR x 48 bytes in flight.
template <int HD, int LDS>
__device__ __forceinline__ void load_rotate_rows(__nv_bfloat16* dst, const __nv_bfloat16* __restrict__ src, size_t gstride,
                                                 const float* __restrict__ fr, int t0, int nrows, int nvalid, int tid) {
  constexpr int CH = HD / 8, CHP = LDS / 8, UNR = 2;
  const int total = nrows * CHP;
  for (int i0 = tid; i0 < total; i0 += AM_THREADS * UNR) {
    uint4 raw[UNR]; float4 f0[UNR], f1[UNR]; bool live[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int i = i0 + u * AM_THREADS;
      const int row = i / CHP, c = i - row * CHP;
      live[u] = i < total && c < CH && row < nvalid;
      if (live[u]) {
        const size_t t = (size_t)(t0 + row);
        raw[u] = *reinterpret_cast<const uint4*>(src + t * gstride + c * 8);
        if (fr) {  // fr == nullptr: q / k were rotated by the qkv GEMM epilogue
          f0[u] = __ldg(reinterpret_cast<const float4*>(fr + t * HD + c * 8));
          f1[u] = __ldg(reinterpret_cast<const float4*>(fr + t * HD + c * 8 + 4));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int i = i0 + u * AM_THREADS;
      if (i < total) {
        const int row = i / CHP, c = i - row * CHP;
        *reinterpret_cast<uint4*>(dst + row * LDS + c * 8) = live[u] ? (fr ? rotate8(raw[u], f0[u], f1[u]) : raw[u]) : make_uint4(0, 0, 0, 0);
      }
    }
  }
}

// One CTA = 128 queries of one (sample, head).  The keys/values of the sample are staged in shared memory in chunks of
// 256 (one chunk when T <= 256): K rotated in registers on the way in, V by cp.async.  Each warp owns 16 query rows
// (Q fragments live in registers) and runs the online softmax over 64-key tiles of the resident chunk.
template <int HD>
__global__ void __launch_bounds__(AM_THREADS, 2) attn_mma_kernel(const __nv_bfloat16* __restrict__ qkv,
                                                                 const float* __restrict__ freqs, int freqs_rows,
                                                                 const int32_t* __restrict__ kv_len,
                                                                 const uint8_t* __restrict__ mask, int R, int Tn, int h,
                                                                 __nv_bfloat16* __restrict__ out, float* __restrict__ lse) {
  constexpr int HDP = (HD + 15) / 16 * 16;  // contraction width of q.k^T (72 -> 80, zero padded)
  constexpr int LDS = HDP + 8;              // smem row stride in elements: odd multiple of 16 bytes => conflict-free ldmatrix
  constexpr int KSTEPS = HDP / 16;
  constexpr int NT_O = HD / 8;              // output n-tiles
  constexpr int CH = HD / 8;                // 16-byte chunks per row

  extern __shared__ __align__(16) uint8_t attn_smem[];
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(attn_smem);  // [AM_CHUNK][LDS]
  __nv_bfloat16* Vs = Ks + AM_CHUNK * LDS;                          // [AM_CHUNK][LDS]; first the Q tile, last the output staging
  uint8_t* kvalid = reinterpret_cast<uint8_t*>(Vs + AM_CHUNK * LDS);  // [AM_CHUNK]

  const int r = blockIdx.z, head = blockIdx.y, q0 = blockIdx.x * AM_BQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, tq4 = lane & 3;
  const int kvl = min(kv_len[r], Tn);
  const size_t row_stride = (size_t)3 * h * HD;
  const __nv_bfloat16* base = qkv + (size_t)r * Tn * row_stride + (size_t)head * HD;
  const float* fr = freqs ? freqs + (size_t)(freqs_rows == 1 ? 0 : r) * Tn * HD : nullptr;
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

  // ---- Q tile: HBM -> rotate in registers -> smem (V region) -> A fragments in registers for the whole kernel
  load_rotate_rows<HD, LDS>(Vs, base, row_stride, fr, q0, AM_BQ, min(AM_BQ, Tn - q0), tid);
  __syncthreads();
  uint32_t qf[KSTEPS][4];
#pragma unroll
  for (int ks = 0; ks < KSTEPS; ++ks) {
    const __nv_bfloat16* p = Vs + (warp * 16 + (lane & 15)) * LDS + ks * 16 + (lane >> 4) * 8;
    ldsm_x4(qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], p);
  }

  float o[NT_O][4];
#pragma unroll
  for (int n = 0; n < NT_O; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const float sl2 = rsqrtf((float)HD) * 1.4426950408889634f;  // softmax scale folded with log2(e)

  for (int c0 = 0; c0 < kvl; c0 += AM_CHUNK) {
    __syncthreads();  // Q fragments read / previous chunk fully consumed: the smem regions may be overwritten
    const int nk = min(AM_CHUNK, kvl - c0);
    const int ntiles = ceil_div(nk, AM_BK);
    const int nrows = ntiles * AM_BK;
    // V: asynchronous 16-byte copies (no transform); rows in [nk, nrows) zeroed so that 0 * garbage cannot make NaN
    for (int i = tid; i < nrows * CH; i += AM_THREADS) {
      const int row = i / CH, c = i - row * CH;
      __nv_bfloat16* dst = Vs + row * LDS + c * 8;
      if (row < nk) cp_async16(dst, base + (size_t)(c0 + row) * row_stride + 2 * (size_t)D + c * 8);
      else *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
    }
    // K: rotate in registers on the way to shared memory
    load_rotate_rows<HD, LDS>(Ks, base + D, row_stride, fr, c0, nrows, nk, tid);
    for (int j = tid; j < nrows; j += AM_THREADS)
      kvalid[j] = (j < nk) && (mask == nullptr || mask[(size_t)r * Tn + c0 + j] != 0);
    cp_async_wait_all();
    __syncthreads();

    for (int kt = 0; kt < ntiles; ++kt) {
      const __nv_bfloat16* Kt = Ks + kt * AM_BK * LDS;
      const __nv_bfloat16* Vt = Vs + kt * AM_BK * LDS;
      const uint8_t* kval = kvalid + kt * AM_BK;
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
          const __nv_bfloat16* p = Kt + (np * 16 + (lane & 7) + ((lane >> 4) << 3)) * LDS + ks * 16 + ((lane >> 3) & 1) * 8;
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
        const bool v0 = kval[c], v1 = kval[c + 1];
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
      // all keys so far masked -> mn = -inf: keep everything at zero without producing NaN
      const float ms0 = mn0 == -INFINITY ? 0.f : mn0, ms1 = mn1 == -INFINITY ? 0.f : mn1;
      const float c0s = exp2f(m0 - ms0), c1s = exp2f(m1 - ms1);
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
      l0 = l0 * c0s + rs0; l1 = l1 * c1s + rs1;
#pragma unroll
      for (int n = 0; n < NT_O; ++n) { o[n][0] *= c0s; o[n][1] *= c0s; o[n][2] *= c1s; o[n][3] *= c1s; }

      // ---- O += P V
#pragma unroll
      for (int kk = 0; kk < AM_BK / 16; ++kk) {
        const uint32_t a[4] = {pf[2 * kk][0], pf[2 * kk][1], pf[2 * kk + 1][0], pf[2 * kk + 1][1]};
#pragma unroll
        for (int np = 0; np < NT_O / 2; ++np) {
          uint32_t b0, b1, b2, b3;
          const __nv_bfloat16* p = Vt + (kk * 16 + (lane & 15)) * LDS + np * 16 + (lane >> 4) * 8;
          ldsm_x4_t(b0, b1, b2, b3, p);
          mma_bf16(o[2 * np], a, b0, b1);
          mma_bf16(o[2 * np + 1], a, b2, b3);
        }
        if (NT_O & 1) {
          uint32_t b0, b1;
          const __nv_bfloat16* p = Vt + (kk * 16 + (lane & 15)) * LDS + (NT_O - 1) * 8;
          ldsm_x2_t(b0, b1, p);
          mma_bf16(o[NT_O - 1], a, b0, b1);
        }
      }
    }
  }

  // ---- finalise: row sums across the quad, normalise, stage through this warp's slice of the V region, 16-byte stores
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = l0 > 0.f ? 1.f / l0 : 0.f, i1 = l1 > 0.f ? 1.f / l1 : 0.f;
  __syncthreads();  // every warp is done reading K / V
  __nv_bfloat16* stage = Vs + warp * 16 * LDS;
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
  constexpr int LDS = (HD + 15) / 16 * 16 + 8;
  constexpr int SMEM = 2 * AM_CHUNK * LDS * 2 + AM_CHUNK;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(attn_mma_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) { set_last_error("attn_mma: cannot set %d bytes of dynamic smem: %s", SMEM, cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    attr_done = true;
  }
  dim3 grid((unsigned)ceil_div(Tn, AM_BQ), (unsigned)h, (unsigned)R);
  attn_mma_kernel<HD><<<grid, AM_THREADS, SMEM, s>>>((const __nv_bfloat16*)qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h,
                                                     (__nv_bfloat16*)out, lse);
  FITB200_CHECK_LAUNCH("attn_mma");
  return FITB200_OK;
}

}  // namespace fitb200

using namespace fitb200;

extern "C" int fitb200_attn_fwd(const void* qkv, int dtype, const float* freqs, int freqs_rows, const int32_t* kv_len,
                                const uint8_t* mask, int R, int T, int h, int d, int flags, void* out, float* lse,
                                fitb200_stream_t stream) {
  const int pre = flags & FITB200_ATTN_PREROTATED;
  FITB200_CHECK_ARG(qkv && (freqs || pre) && kv_len && out && R > 0 && T > 0 && h > 0, "attn_fwd: bad arguments");
  FITB200_CHECK_ARG(pre || freqs_rows == 1 || freqs_rows == R, "attn_fwd: freqs_rows must be 1 or R");
  FITB200_CHECK_ARG(h <= 65535 && R <= 65535, "attn_fwd: h and R must be <= 65535");
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == FITB200_F32 && pre) { set_last_error("attn_fwd: PREROTATED is a bf16-path flag"); return FITB200_ERR_UNSUPPORTED; }
  if (dtype == FITB200_F32) return attn_fwd_f32(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, d, out, lse, s);
  if (dtype == FITB200_BF16) {
    // sequences that fit the resident K/V tile (every trained shape): tcgen05 kernel; longer (extrapolation): chunked mma.sync kernel
    if (T <= 256 && (d == 64 || d == 72)) return attn_fwd_tc(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, d, pre, out, lse, s);
    if (pre && (d == 64 || d == 72) && ((uintptr_t)out & 15) == 0) {
      // KV-blocked tcgen05 kernel (attn_tc3.cu); FITB200_ATTN_MMA=1 keeps the mma.sync kernel below (A/B timing, debugging)
      static const bool use_mma = [] { const char* e = getenv("FITB200_ATTN_MMA"); return e && e[0] == '1'; }();
      if (!use_mma) return attn_fwd_tc3(qkv, kv_len, mask, R, T, h, d, out, lse, s);
    }
    if (pre) freqs = nullptr;
    if (d == 64) return launch_attn_mma<64>(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, out, lse, s);
    if (d == 72) return launch_attn_mma<72>(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, out, lse, s);
    set_last_error("attn_fwd: head_dim %d not supported (64 or 72)", d);
    return FITB200_ERR_UNSUPPORTED;
  }
  set_last_error("attn_fwd: unsupported dtype %d", dtype);
  return FITB200_ERR_INVALID_ARG;
}
