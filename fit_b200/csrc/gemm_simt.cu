// fp32 FMA GEMM on CUDA cores with the fused epilogues of the hot path.
// Used by the fp32 verification mode (north_star: 1e-4 vs the fp32 reference -- tensor-core TF32/BF16
// cannot meet that) and for the tiny embedder GEMMs (M = batch rows).  64x64x16 tiles, 4x4 per thread.
#include "common.cuh"

namespace fitb200 {

constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16, SG_PAD = 4;

template <int EPI, typename OutT, typename ResT>
__global__ void __launch_bounds__(256) gemm_f32_kernel(fitb200_gemm_args a) {
  __shared__ float As[SG_BK][SG_BM + SG_PAD];
  __shared__ float Bs[SG_BK][SG_BN + SG_PAD];
  __shared__ float Bs2[EPI == FITB200_EPI_SWIGLU ? SG_BK : 1][SG_BN + SG_PAD];

  const float* __restrict__ A = (const float*)a.A;
  const float* __restrict__ B = (const float*)a.B;
  const float* __restrict__ B2 = (const float*)a.B2;
  const int M = a.M, N = a.N, K = a.K;
  const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;       // 16 x 16 threads, 4x4 outputs each
  const int lrow = tid >> 2, lk = (tid & 3) * 4;  // loader: 64 rows x 4 float4 along K

  float acc[4][4] = {}, acc2[4][4] = {};

  for (int k0 = 0; k0 < K; k0 += SG_BK) {
    float4 av = make_float4(0.f, 0.f, 0.f, 0.f), bv = av, b2v = av;
    const int kk = k0 + lk;
    // contraction-major operands (backward GEMMs): thread -> k-row tid >> 4, 4 consecutive m / n at (tid & 15) * 4
    const int tk = k0 + (tid >> 4), tmn = (tid & 15) * 4;
    if (a.a_mn) {
      float t[4] = {0.f, 0.f, 0.f, 0.f};
      if (tk < K)
#pragma unroll
        for (int i = 0; i < 4; ++i) if (m0 + tmn + i < M) t[i] = A[(size_t)tk * a.lda + m0 + tmn + i];
      As[tid >> 4][tmn + 0] = t[0]; As[tid >> 4][tmn + 1] = t[1]; As[tid >> 4][tmn + 2] = t[2]; As[tid >> 4][tmn + 3] = t[3];
    } else {
      if (m0 + lrow < M && kk < K) av = *reinterpret_cast<const float4*>(A + (size_t)(m0 + lrow) * a.lda + kk);
      As[lk + 0][lrow] = av.x; As[lk + 1][lrow] = av.y; As[lk + 2][lrow] = av.z; As[lk + 3][lrow] = av.w;
    }
    if (a.b_mn) {
      float t[4] = {0.f, 0.f, 0.f, 0.f};
      if (tk < K)
#pragma unroll
        for (int i = 0; i < 4; ++i) if (n0 + tmn + i < N) t[i] = B[(size_t)tk * a.ldb + n0 + tmn + i];
      Bs[tid >> 4][tmn + 0] = t[0]; Bs[tid >> 4][tmn + 1] = t[1]; Bs[tid >> 4][tmn + 2] = t[2]; Bs[tid >> 4][tmn + 3] = t[3];
    } else {
      if (n0 + lrow < N && kk < K) {
        bv = *reinterpret_cast<const float4*>(B + (size_t)(n0 + lrow) * a.ldb + kk);
        if (EPI == FITB200_EPI_SWIGLU) b2v = *reinterpret_cast<const float4*>(B2 + (size_t)(n0 + lrow) * a.ldb + kk);
      }
      Bs[lk + 0][lrow] = bv.x; Bs[lk + 1][lrow] = bv.y; Bs[lk + 2][lrow] = bv.z; Bs[lk + 3][lrow] = bv.w;
    }
    if (EPI == FITB200_EPI_SWIGLU) {
      Bs2[lk + 0][lrow] = b2v.x; Bs2[lk + 1][lrow] = b2v.y; Bs2[lk + 2][lrow] = b2v.z; Bs2[lk + 3][lrow] = b2v.w;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SG_BK; ++k) {
      const float4 ar = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 br = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float af[4] = {ar.x, ar.y, ar.z, ar.w}, bf[4] = {br.x, br.y, br.z, br.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(af[i], bf[j], acc[i][j]);
      if (EPI == FITB200_EPI_SWIGLU) {
        const float4 b2r = *reinterpret_cast<const float4*>(&Bs2[k][tx * 4]);
        const float b2f[4] = {b2r.x, b2r.y, b2r.z, b2r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc2[i][j] = fmaf(af[i], b2f[j], acc2[i][j]);
      }
    }
    __syncthreads();
  }

  OutT* __restrict__ out = (OutT*)a.out;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j] + (a.bias ? a.bias[n] : 0.f);
      if (EPI == FITB200_EPI_BIAS_SILU) {
        v = silu_acc(v);
      } else if (EPI == FITB200_EPI_BIAS_GATE_RESID) {
        const float g = a.gate[(size_t)(m / a.rows_per_sample) * a.ld_gate + n];
        const float r = to_f32(((const ResT*)a.resid)[(size_t)m * a.ld_resid + n]);
        v = r + g * v;
      } else if (EPI == FITB200_EPI_SWIGLU) {
        const float u = acc2[i][j] + (a.bias2 ? a.bias2[n] : 0.f);
        v = silu_acc(v) * u;
      }
      out[(size_t)m * a.ldo + n] = from_f32<OutT>(v);
    }
  }
}

template <int EPI, typename OutT, typename ResT>
static int launch_f32(const fitb200_gemm_args& a, cudaStream_t s) {
  dim3 grid((unsigned)ceil_div(a.N, SG_BN), (unsigned)ceil_div(a.M, SG_BM));
  gemm_f32_kernel<EPI, OutT, ResT><<<grid, 256, 0, s>>>(a);
  FITB200_CHECK_LAUNCH("gemm_f32");
  return FITB200_OK;
}

template <typename OutT>
static int dispatch_f32(const fitb200_gemm_args& a, cudaStream_t s) {
  switch (a.epilogue) {
    case FITB200_EPI_BIAS: return launch_f32<FITB200_EPI_BIAS, OutT, float>(a, s);
    case FITB200_EPI_BIAS_SILU: return launch_f32<FITB200_EPI_BIAS_SILU, OutT, float>(a, s);
    case FITB200_EPI_SWIGLU: return launch_f32<FITB200_EPI_SWIGLU, OutT, float>(a, s);
    case FITB200_EPI_BIAS_GATE_RESID:
      if (a.resid_dtype == FITB200_F32) return launch_f32<FITB200_EPI_BIAS_GATE_RESID, OutT, float>(a, s);
      if (a.resid_dtype == FITB200_BF16) return launch_f32<FITB200_EPI_BIAS_GATE_RESID, OutT, __nv_bfloat16>(a, s);
      break;
  }
  set_last_error("gemm_f32: unsupported epilogue %d / resid dtype %d", a.epilogue, a.resid_dtype);
  return FITB200_ERR_INVALID_ARG;
}

}  // namespace fitb200

using namespace fitb200;

extern "C" int fitb200_gemm_f32(const fitb200_gemm_args* ap, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(ap, "gemm_f32: null args");
  const fitb200_gemm_args& a = *ap;
  FITB200_CHECK_ARG(a.A && a.B && a.out && a.M > 0 && a.N > 0 && a.K > 0, "gemm_f32: bad arguments");
  FITB200_CHECK_ARG(((a.a_mn && a.b_mn) || a.K % 4 == 0) && (a.a_mn || a.lda % 4 == 0) && (a.b_mn || a.ldb % 4 == 0),
                    "gemm_f32: K, lda, ldb must be multiples of 4 (K=%d)", a.K);
  FITB200_CHECK_ARG(!(a.a_mn || a.b_mn) || a.epilogue == FITB200_EPI_BIAS, "gemm_f32: a_mn / b_mn support FITB200_EPI_BIAS only");
  if (a.epilogue == FITB200_EPI_SWIGLU) FITB200_CHECK_ARG(a.B2, "gemm_f32: SWIGLU needs B2");
  if (a.epilogue == FITB200_EPI_BIAS_GATE_RESID)
    FITB200_CHECK_ARG(a.gate && a.resid && a.rows_per_sample > 0, "gemm_f32: GATE_RESID needs gate, resid, rows_per_sample");
  if (a.out_dtype == FITB200_F32) return dispatch_f32<float>(a, (cudaStream_t)stream);
  if (a.out_dtype == FITB200_BF16) return dispatch_f32<__nv_bfloat16>(a, (cudaStream_t)stream);
  set_last_error("gemm_f32: unsupported out dtype %d", a.out_dtype);
  return FITB200_ERR_INVALID_ARG;
}
