// bf16 GEMM on the 5th-generation tensor cores:  out[M,N] = epilogue(A[M,K] . B[N,K]^T)
//
//   * persistent: one CTA per SM walks output tiles (128 x BN, BN in {128,192,256}), n fastest so the
//     CTAs of a wave share A row-blocks in L2 while the (small) weight stays L2 resident
//   * warp 0 : TMA producer   -- cp.async.bulk.tensor 128B-swizzled A/B k-slabs into a STAGES-deep ring
//   * warp 1 : MMA issuer     -- one thread issues tcgen05.mma (M=128, N=BN, K=16) into TMEM,
//                                tcgen05.commit releases smem slots / publishes the accumulator
//   * warps 2-5 : epilogue    -- tcgen05.ld TMEM -> registers, fused bias / SiLU-gate / adaLN-gate +
//                                residual, 16-byte global stores.  Two TMEM accumulator stages, so the
//                                epilogue of tile i overlaps the mainloop of tile i+1.
// Every mbarrier wait is bounded (ptx::mbar_wait traps) so a pipeline bug cannot hang the device.
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace fitb200 {

constexpr int GT_BM = 128, GT_BK = 64;
constexpr int GT_THREADS = 192;  // producer + mma + 4 epilogue warps
constexpr int GT_EPI_THREADS = 128;

template <int BN> struct GtCfg {
  static constexpr int STAGES = BN == 256 ? 4 : (BN == 192 ? 5 : 6);
  static constexpr int A_BYTES = GT_BM * GT_BK * 2;
  static constexpr int B_BYTES = BN * GT_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int TMEM_COLS = BN == 128 ? 256 : 512;  // 2 accumulator stages, power of two
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct GtParams {
  int M, N, K;
  const float* bias;
  void* out; long ldo;
  const float* gate; long ld_gate; int rows_per_sample;
  const void* resid; long ld_resid;
};

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t lo[16], hi[16];
  ptx::tmem_ld16(taddr, lo);
  ptx::tmem_ld16(taddr + 16, hi);
  ptx::tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 16; ++i) { v[i] = __uint_as_float(lo[i]); v[16 + i] = __uint_as_float(hi[i]); }
}

template <typename T> __device__ __forceinline__ void store8(T* p, const float* v);
template <> __device__ __forceinline__ void store8<float>(float* p, const float* v) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float* v) {
  uint4 o;
  o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]); o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = o;
}
template <typename T> __device__ __forceinline__ void load8(const T* p, float* v);
template <> __device__ __forceinline__ void load8<float>(const float* p, float* v) {
  float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <> __device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float* v) {
  Vec8<__nv_bfloat16> t = Vec8<__nv_bfloat16>::load(p);
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = t.v[i];
}

template <int BN, int EPI, typename OutT, typename ResT>
__global__ void __launch_bounds__(GT_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GtParams p) {
  using Cfg = GtCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);  // SWIZZLE_128B needs 1024-byte alignment
  uint8_t* smemA = smem;
  uint8_t* smemB = smem + STAGES * Cfg::A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* full_bar = bars;                    // [STAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + STAGES;          // [STAGES]  MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * STAGES;      // [2]       MMA -> epilogue
  uint64_t* tempty_bar = bars + 2 * STAGES + 2; // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = ceil_div(p.N, BN), tiles_m = ceil_div(p.M, GT_BM);
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = ceil_div(p.K, GT_BK);

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full_bar[s], 1); ptx::mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(&tfull_bar[s], 1); ptx::mbar_init(&tempty_bar[s], GT_EPI_THREADS); }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {  // one warp owns TMEM allocation and deallocation
    ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / tiles_n) * GT_BM, n0 = (tile % tiles_n) * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          ptx::mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
          ptx::tma_load_2d(smemA + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], kb * GT_BK, m0);
          ptx::tma_load_2d(smemB + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], kb * GT_BK, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::idesc_bf16_f32(GT_BM, BN);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);  // epilogue has drained this accumulator stage
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint64_t da = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(smemA + stage * Cfg::A_BYTES));
          const uint64_t db = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(smemB + stage * Cfg::B_BYTES));
#pragma unroll
          for (int k = 0; k < GT_BK / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle row: +2 in the (addr >> 4) field
            ptx::mma_f16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          }
          ptx::mma_commit(&empty_bar[stage]);  // smem slot reusable once these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        ptx::mma_commit(&tfull_bar[as]);  // accumulator complete
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue warps (2..5)
    const int quarter = warp & 3;            // TMEM lane quarter this warp may read
    const int row_in_tile = quarter * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      const int m0 = (tile / tiles_n) * GT_BM, n0 = (tile % tiles_n) * BN;
      const int m = m0 + row_in_tile;
      const bool row_ok = m < p.M;
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * BN;
      const float* gate_row = nullptr;
      if (EPI == FITB200_EPI_BIAS_GATE_RESID && row_ok) gate_row = p.gate + (size_t)(m / p.rows_per_sample) * p.ld_gate;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int n = n0 + c * 32;
        if (n >= p.N) break;  // warp-uniform
        float acc[32];
        __syncwarp();  // tcgen05.ld is .sync.aligned: reconverge after the row-predicated stores below
        tmem_ld32(taddr + c * 32, acc);
        if (!row_ok) continue;
        if (EPI == FITB200_EPI_SWIGLU) {
          // packed weight: 16 gate rows then the matching 16 up rows -> 16 hidden units per chunk
          const int j = n >> 1;
          float o[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float gv = acc[i] + __ldg(p.bias + n + i);
            const float uv = acc[16 + i] + __ldg(p.bias + n + 16 + i);
            o[i] = silu_f(gv) * uv;
          }
          OutT* op = (OutT*)p.out + (size_t)m * p.ldo + j;
          store8<OutT>(op, o);
          store8<OutT>(op + 8, o + 8);
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int nq = n + q * 8;
            if (nq >= p.N) break;
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = acc[q * 8 + i] + (p.bias ? __ldg(p.bias + nq + i) : 0.f);
            if (EPI == FITB200_EPI_BIAS_SILU) {
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] = silu_f(v[i]);
            } else if (EPI == FITB200_EPI_BIAS_GATE_RESID) {
              float r[8], gt[8];
              load8<ResT>((const ResT*)p.resid + (size_t)m * p.ld_resid + nq, r);
              load8<float>(gate_row + nq, gt);
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] = fmaf(gt[i], v[i], r[i]);
            }
            store8<OutT>((OutT*)p.out + (size_t)m * p.ldo + nq, v);
          }
        }
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(&tempty_bar[as]);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------------------------- host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)sym;
  });
  return fn;
}

// 2-D bf16 row-major [rows, cols] with leading dimension ld (elements); box = box_rows x 64 columns, 128B swizzle
static int make_tmap(CUtensorMap* m, const void* ptr, long rows, long cols, long ld, int box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) { set_last_error("gemm_bf16_tc: cuTensorMapEncodeTiled not available from the driver"); return FITB200_ERR_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)GT_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_error("gemm_bf16_tc: cuTensorMapEncodeTiled failed (%d) rows=%ld cols=%ld ld=%ld", (int)r, rows, cols, ld); return FITB200_ERR_CUDA; }
  return FITB200_OK;
}

template <int BN, int EPI, typename OutT, typename ResT>
static int launch_tc(const fitb200_gemm_args& a, cudaStream_t s) {
  using Cfg = GtCfg<BN>;
  CUtensorMap tmA, tmB;
  int rc = make_tmap(&tmA, a.A, a.M, a.K, a.lda, GT_BM);
  if (rc) return rc;
  rc = make_tmap(&tmB, a.B, a.N, a.K, a.ldb, BN);
  if (rc) return rc;
  GtParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.bias = a.bias; p.out = a.out; p.ldo = a.ldo;
  p.gate = a.gate; p.ld_gate = a.ld_gate; p.rows_per_sample = a.rows_per_sample; p.resid = a.resid; p.ld_resid = a.ld_resid;
  auto kern = gemm_tc_kernel<BN, EPI, OutT, ResT>;
  static bool attr_done = false;  // per template instantiation
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) { set_last_error("gemm_bf16_tc: cannot set %d bytes of dynamic smem: %s", Cfg::SMEM_BYTES, cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    attr_done = true;
  }
  const int tiles = ceil_div(a.M, GT_BM) * ceil_div(a.N, BN);
  const int grid = tiles < sm_count() ? tiles : sm_count();
  kern<<<grid, GT_THREADS, Cfg::SMEM_BYTES, s>>>(tmA, tmB, p);
  FITB200_CHECK_LAUNCH("gemm_bf16_tc");
  return FITB200_OK;
}

template <int BN, typename OutT>
static int dispatch_epi(const fitb200_gemm_args& a, cudaStream_t s) {
  switch (a.epilogue) {
    case FITB200_EPI_BIAS: return launch_tc<BN, FITB200_EPI_BIAS, OutT, float>(a, s);
    case FITB200_EPI_BIAS_SILU: return launch_tc<BN, FITB200_EPI_BIAS_SILU, OutT, float>(a, s);
    case FITB200_EPI_SWIGLU: return launch_tc<BN, FITB200_EPI_SWIGLU, OutT, float>(a, s);
    case FITB200_EPI_BIAS_GATE_RESID:
      if (a.resid_dtype == FITB200_F32) return launch_tc<BN, FITB200_EPI_BIAS_GATE_RESID, OutT, float>(a, s);
      if (a.resid_dtype == FITB200_BF16) return launch_tc<BN, FITB200_EPI_BIAS_GATE_RESID, OutT, __nv_bfloat16>(a, s);
      break;
  }
  set_last_error("gemm_bf16_tc: unsupported epilogue %d / resid dtype %d", a.epilogue, a.resid_dtype);
  return FITB200_ERR_INVALID_ARG;
}

template <typename OutT>
static int dispatch_bn(const fitb200_gemm_args& a, cudaStream_t s) {
  // widest tile with the least padded columns (3456, 1152 -> 192; 6144, 2304, 768 -> 256)
  int best = 256; long best_waste = (long)ceil_div(a.N, 256) * 256 - a.N;
  const int cand[2] = {192, 128};
  for (int c : cand) {
    long w = (long)ceil_div(a.N, c) * c - a.N;
    if (w < best_waste) { best = c; best_waste = w; }
  }
  if (best == 256) return dispatch_epi<256, OutT>(a, s);
  if (best == 192) return dispatch_epi<192, OutT>(a, s);
  return dispatch_epi<128, OutT>(a, s);
}

}  // namespace fitb200

using namespace fitb200;

extern "C" int fitb200_gemm_bf16_tc(const fitb200_gemm_args* ap, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(ap, "gemm_bf16_tc: null args");
  const fitb200_gemm_args& a = *ap;
  FITB200_CHECK_ARG(a.A && a.B && a.out && a.M > 0 && a.N > 0 && a.K > 0, "gemm_bf16_tc: bad arguments");
  FITB200_CHECK_ARG(a.K % 8 == 0 && a.lda % 8 == 0 && a.ldb % 8 == 0, "gemm_bf16_tc: K, lda, ldb must be multiples of 8 (K=%d lda=%lld ldb=%lld)", a.K, (long long)a.lda, (long long)a.ldb);
  FITB200_CHECK_ARG(((uintptr_t)a.A & 15) == 0 && ((uintptr_t)a.B & 15) == 0 && ((uintptr_t)a.out & 15) == 0, "gemm_bf16_tc: A, B, out must be 16-byte aligned");
  FITB200_CHECK_ARG(a.N % 8 == 0 && a.ldo % 8 == 0, "gemm_bf16_tc: N and ldo must be multiples of 8 (N=%d)", a.N);
  if (a.epilogue == FITB200_EPI_SWIGLU) FITB200_CHECK_ARG(a.N % 32 == 0 && a.bias, "gemm_bf16_tc: SWIGLU needs N %% 32 == 0 and a packed bias");
  if (a.epilogue == FITB200_EPI_BIAS_GATE_RESID)
    FITB200_CHECK_ARG(a.gate && a.resid && a.rows_per_sample > 0 && a.ld_gate % 4 == 0 && a.ld_resid % 8 == 0 &&
                          ((uintptr_t)a.gate & 15) == 0 && ((uintptr_t)a.resid & 15) == 0,
                      "gemm_bf16_tc: GATE_RESID needs aligned gate/resid and rows_per_sample");
  if (a.out_dtype == FITB200_BF16) return dispatch_bn<__nv_bfloat16>(a, (cudaStream_t)stream);
  if (a.out_dtype == FITB200_F32) return dispatch_bn<float>(a, (cudaStream_t)stream);
  set_last_error("gemm_bf16_tc: unsupported out dtype %d", a.out_dtype);
  return FITB200_ERR_INVALID_ARG;
}
