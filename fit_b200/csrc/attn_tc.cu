// bf16 fused attention forward on the 5th-generation tensor cores (tcgen05 + TMEM), sequences of up to 256 tokens
// (every trained / headline shape of FiT; longer sequences take the chunked mma.sync kernel in attn_mma.cu).
//
// One CTA = 128 queries of one (sample, head); two CTAs per SM (256 TMEM columns, ~100 KB shared memory each) so that
// the loads / softmax of one overlap the MMAs of the other.
//   load     : q, k rows HBM -> registers -> 2-D RoPE (fp32) -> bf16 -> shared memory in the tcgen05 no-swizzle
//              core-matrix layout  offset(row, chunk) = chunk * ROWS * 16 + row * 16  (chunk = 8 head-dim elements);
//              v rows by cp.async into the same layout.  Only keys < kv_len are loaded: padded tokens cost nothing.
//   S = Q K^T: one tcgen05.mma (M=128, N=round16(kv_len), K=16) per 16 head-dim columns, A and B K-major from smem,
//              accumulator in TMEM columns [0, N)                                              (Attention.forward, fit.py:183)
//   softmax  : one thread per (query row, half of the keys): two passes over its TMEM lane (tcgen05.ld) -- exact row
//              max, then exp2 / sum -- no shuffles; P is written back to TMEM as packed bf16 pairs (tcgen05.st) over
//              columns already consumed; key-padding mask from a 256-bit validity word            (fit.py:184-186)
//   O = P V  : tcgen05.mma with A = P from TMEM and B = V from smem (MN-major descriptor), accumulator in
//              TMEM columns [64, 64 + d)                                                        (fit.py:188)
//   epilogue : tcgen05.ld, 1/rowsum, bf16, staged in smem, 16-byte coalesced stores into [R, T, h*d]   (fit.py:220-227)
// Two operand paths: PRE = q and k already rotated by the qkv GEMM epilogue (FITB200_EPI_BIAS_ROPE): Q / K / V tiles are
// fetched by three-dimensional TMA boxes {8 elements, 128 rows, d/8 chunks} that land directly in the core-matrix layout --
// one asynchronous round, no registers; otherwise the rows go through registers and the RoPE as described above.
// Keys are held as two 128-row halves: [half][chunk][128 rows][16 B].
// The operand conventions used here (no-swizzle LBO/SBO roles, MN-major B, A-from-TMEM packing, accumulator at an
// unaligned column) are pinned by tools/probe/umma_probe.cu.
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "ptx.cuh"
#include "umma_nosw.cuh"

namespace fitb200 {

constexpr int AT_BQ = 128;       // queries per CTA = TMEM lanes
constexpr int AT_KEYS = 256;     // key capacity of the resident K / V tiles
constexpr int AT_THREADS = 256;  // 8 warps: warps w and w+4 share TMEM lane quarter w and split the keys / columns
constexpr int AT_TMEM_COLS = 256;
constexpr int AT_OCOL = 64;      // first TMEM column of the O accumulator

__device__ __forceinline__ void at_cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(ptx::smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void at_cp_async_wait_all() { asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory"); }

// rotate the 4 (re, im) pairs of 8 consecutive bf16 by the fp32 (cos, sin) table, round to bf16 (complex_mult, fit.py:152-154)
__device__ __forceinline__ uint4 at_rotate8(const uint4 raw, const float4 f0, const float4 f1) {
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
  const float cs[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
  uint4 o;
  uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 ab = __bfloat1622float2(h2[i]);
    const float c = cs[2 * i], s = cs[2 * i + 1];
    ow[i] = pack_bf16x2(ab.x * c - ab.y * s, ab.y * c + ab.x * s);
  }
  return o;
}

// Rows [0, nrows) of a [rows][HD] bf16 operand (global row stride `gstride` elements, first token t0) -> RoPE -> smem core-matrix
// layout in blocks of ROWS rows: offset(row, chunk) = (row / ROWS) * BLOCK + chunk * ROWS * 16 + (row % ROWS) * 16.  Rows in [nvalid, nrows) are zero-filled.  Work item = (8-row group, chunk): a quarter
// warp covers 8 consecutive rows of one chunk (one contiguous 128-byte core matrix in smem), a warp 4 consecutive chunks (64
// contiguous bytes per row in HBM).
template <int HD, int ROWS, bool ROTATE>
__device__ __forceinline__ void at_load_rows(uint8_t* dst, const __nv_bfloat16* __restrict__ src, size_t gstride,
                                             const float* __restrict__ fr, int t0, int nrows, int nvalid, int tid) {
  constexpr int CH = HD / 8, UNR = 3;
  constexpr int BLOCK = (HD + 15) / 16 * 2 * ROWS * 16;  // bytes of one ROWS-row block (all chunks incl. contraction padding)
  const int total = (nrows >> 3) * CH * 8;  // nrows is a multiple of 8
  for (int i0 = tid; i0 < total; i0 += AT_THREADS * UNR) {
    uint4 raw[UNR]; float4 f0[UNR], f1[UNR]; bool live[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int i = i0 + u * AT_THREADS;
      const int qd = i >> 3, g = qd / CH, c = qd - g * CH, row = g * 8 + (i & 7);
      live[u] = i < total && row < nvalid;
      if (live[u]) {
        const size_t t = (size_t)(t0 + row);
        if (ROTATE) {
          raw[u] = *reinterpret_cast<const uint4*>(src + t * gstride + c * 8);
          f0[u] = __ldg(reinterpret_cast<const float4*>(fr + t * HD + c * 8));
          f1[u] = __ldg(reinterpret_cast<const float4*>(fr + t * HD + c * 8 + 4));
        } else {
          at_cp_async16(dst + (row / ROWS) * BLOCK + c * (ROWS * 16) + (row % ROWS) * 16, src + t * gstride + c * 8);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int i = i0 + u * AT_THREADS;
      if (i < total) {
        const int qd = i >> 3, g = qd / CH, c = qd - g * CH, row = g * 8 + (i & 7);
        uint8_t* p = dst + (row / ROWS) * BLOCK + c * (ROWS * 16) + (row % ROWS) * 16;
        if (ROTATE) *reinterpret_cast<uint4*>(p) = live[u] ? at_rotate8(raw[u], f0[u], f1[u]) : make_uint4(0, 0, 0, 0);
        else if (!live[u]) *reinterpret_cast<uint4*>(p) = make_uint4(0, 0, 0, 0);
      }
    }
  }
}

template <int HD, bool PRE>
__global__ void __launch_bounds__(AT_THREADS, 2) attn_tc_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                const __nv_bfloat16* __restrict__ qkv, const float* __restrict__ freqs,
                                                                int freqs_rows, const int32_t* __restrict__ kv_len,
                                                                const uint8_t* __restrict__ mask, int R, int Tn, int h,
                                                                __nv_bfloat16* __restrict__ out, float* __restrict__ lse) {
  constexpr int DP = (HD + 15) / 16 * 16;  // contraction width of q.k^T and N of p.v (72 -> 80)
  constexpr int CHP = DP / 8, CH = HD / 8;
  constexpr int HB = 128;                   // rows per operand block
  constexpr int CS = HB * 16;               // chunk-column stride inside a block
  constexpr int BLOCK = CHP * CS;           // bytes of one 128-row block
  constexpr int STG = 144;                  // staging row stride (bytes): conflict-free 16-byte stores

  extern __shared__ __align__(128) uint8_t at_smem[];
  uint8_t* sQ = at_smem;
  uint8_t* sK = sQ + BLOCK;        // [2][CHP][128][16]
  uint8_t* sV = sK + 2 * BLOCK;    // [2][CHP][128][16]
  float* smax = reinterpret_cast<float*>(sV + 2 * BLOCK);  // [2][128]
  float* ssum = smax + 2 * AT_BQ;                           // [2][128]
  uint32_t* vw = reinterpret_cast<uint32_t*>(ssum + 2 * AT_BQ);  // [8] key-validity bits
  uint64_t* bars = reinterpret_cast<uint64_t*>(vw + 8);          // [4]: q/k landed, v landed, S done, O done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int r = blockIdx.z, head = blockIdx.y, q0 = blockIdx.x * AT_BQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kvl = min(kv_len[r], Tn);
  const int D = h * HD;

  if (q0 >= kvl) {  // padded query tile: zero output rows, no work
    for (int i = tid; i < AT_BQ * CH; i += AT_THREADS) {
      const int row = i / CH, c = i - row * CH;
      if (q0 + row < Tn) *reinterpret_cast<uint4*>(out + ((size_t)r * Tn + q0 + row) * D + head * HD + c * 8) = make_uint4(0, 0, 0, 0);
    }
    if (lse) for (int i = tid; i < AT_BQ; i += AT_THREADS) if (q0 + i < Tn) lse[((size_t)r * h + head) * Tn + q0 + i] = 0.f;
    return;
  }
  const int nk = kvl, nk16 = (nk + 15) & ~15;
  const int n0 = min(nk16, HB), n1 = nk16 - n0;  // keys in the two halves (multiples of 16)

  if (tid == 0) {
    for (int i = 0; i < 4; ++i) ptx::mbar_init(&bars[i], 1);
    ptx::fence_mbar_init();
    if (PRE) {
      // ---- one asynchronous round: Q, K (barrier 0) and V (barrier 1) boxes straight into the core-matrix layout
      ptx::prefetch_tmap(&tmQKV);
      const int row0 = r * Tn, cq = head * CH, ck = (D + head * HD) / 8, cv = (2 * D + head * HD) / 8;
      const uint32_t box = CH * CS;
      ptx::mbar_arrive_expect_tx(&bars[0], box * (n1 > 0 ? 3 : 2));
      tma_load_3d(sQ, &tmQKV, &bars[0], 0, row0 + q0, cq);
      tma_load_3d(sK, &tmQKV, &bars[0], 0, row0, ck);
      if (n1 > 0) tma_load_3d(sK + BLOCK, &tmQKV, &bars[0], 0, row0 + HB, ck);
      ptx::mbar_arrive_expect_tx(&bars[1], box * (n1 > 0 ? 2 : 1));
      tma_load_3d(sV, &tmQKV, &bars[1], 0, row0, cv);
      if (n1 > 0) tma_load_3d(sV + BLOCK, &tmQKV, &bars[1], 0, row0 + HB, cv);
    }
  }
  if (!PRE) {
    // ---- register path: V (async copies) first, then Q and K through the RoPE
    const size_t row_stride = (size_t)3 * D;
    const __nv_bfloat16* base = qkv + (size_t)r * Tn * row_stride + (size_t)head * HD;
    const float* fr = freqs + (size_t)(freqs_rows == 1 ? 0 : r) * Tn * HD;
    const int nq = min(AT_BQ, Tn - q0);
    at_load_rows<HD, HB, false>(sV, base + 2 * (size_t)D, row_stride, fr, 0, nk16, nk, tid);
    at_load_rows<HD, HB, true>(sQ, base, row_stride, fr, q0, AT_BQ, nq, tid);
    at_load_rows<HD, HB, true>(sK, base + D, row_stride, fr, 0, nk16, nk, tid);
  }
  if (DP != HD) {  // zero contraction padding (head_dim 72 -> 80) of Q and K
    for (int i = tid; i < AT_BQ + nk16; i += AT_THREADS) {
      const int kr = i - AT_BQ;
      uint8_t* p = i < AT_BQ ? sQ + CH * CS + i * 16 : sK + (kr >> 7) * BLOCK + CH * CS + (kr & 127) * 16;
      *reinterpret_cast<uint4*>(p) = make_uint4(0, 0, 0, 0);
    }
  }
  {
    const int key = warp * 32 + lane;
    const bool v = key < nk && (mask == nullptr || mask[(size_t)r * Tn + key] != 0);
    const uint32_t w = __ballot_sync(0xffffffffu, v);
    if (lane == 0) vw[warp] = w;
  }
  if (warp == 0) { ptx::tmem_alloc(tmem_slot, AT_TMEM_COLS); ptx::tmem_relinquish(); }
  if (!PRE) at_cp_async_wait_all();
  ptx::fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = *tmem_slot;

  // ---- S = Q K^T: keys [0,128) -> TMEM columns [0, n0), keys [128, 256) -> columns [128, 128 + n1)
  if (tid == 0) {
    if (PRE) ptx::mbar_wait(&bars[0], 0);
    ptx::tc_fence_after();
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      const int nn = hf ? n1 : n0;
      if (nn > 0) {
        const uint32_t idesc = ptx::idesc_bf16_f32(AT_BQ, nn);
#pragma unroll
        for (int k = 0; k < DP / 16; ++k) {
          const uint64_t da = desc_nosw(ptx::smem_u32(sQ + 2 * k * CS), CS, 128);
          const uint64_t db = desc_nosw(ptx::smem_u32(sK + hf * BLOCK + 2 * k * CS), CS, 128);
          ptx::mma_f16_ss(tb + hf * 128, da, db, idesc, k != 0);
        }
      }
    }
    ptx::mma_commit(&bars[2]);
  }
  ptx::mbar_wait(&bars[2], 0);
  ptx::tc_fence_after();
  // ---- softmax: thread = (query row, half of the keys)
  const int quarter = warp & 3, half = warp >> 2, row = quarter * 32 + lane;
  const uint32_t lane_base = tb + ((uint32_t)(quarter * 32) << 16);
  const int n_my = max(0, min(nk16 - half * 128, 128));
  const int nchunks = (n_my + 31) >> 5;
  const float sl2 = rsqrtf((float)HD) * 1.4426950408889634f;  // softmax scale folded with log2(e)
  float m = -INFINITY;
  for (int c = 0; c < nchunks; ++c) {
    uint32_t s[32];
    tmem_ld32(lane_base + half * 128 + c * 32, s);
    ptx::tmem_ld_wait();
    const uint32_t w = vw[half * 4 + c];
    if (w == 0xffffffffu) {
#pragma unroll
      for (int i = 0; i < 32; ++i) m = fmaxf(m, __uint_as_float(s[i]));
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) m = (w >> i) & 1u ? fmaxf(m, __uint_as_float(s[i])) : m;
    }
  }
  smax[half * AT_BQ + row] = m;
  __syncthreads();
  m = fmaxf(smax[row], smax[AT_BQ + row]);
  const float msc = (m == -INFINITY ? 0.f : m) * sl2;
  float sum = 0.f;
  for (int cc = 0; cc < nchunks; ++cc) {
    // half 0 writes P over columns [0, 64) ascending; half 1 walks its chunks downwards and writes P over [192, 256):
    // every destination column has already been consumed by the thread that owns it
    const int c = half ? nchunks - 1 - cc : cc;
    uint32_t s[32];
    tmem_ld32(lane_base + half * 128 + c * 32, s);
    ptx::tmem_ld_wait();
    const uint32_t w = vw[half * 4 + c];
    uint32_t pk[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float p0 = ex2f(fmaf(__uint_as_float(s[2 * i]), sl2, -msc));
      float p1 = ex2f(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -msc));
      if (w != 0xffffffffu) {
        p0 = (w >> (2 * i)) & 1u ? p0 : 0.f;
        p1 = (w >> (2 * i + 1)) & 1u ? p1 : 0.f;
      }
      sum += p0 + p1;
      pk[i] = pack_bf16x2(p0, p1);
    }
    tmem_st16(lane_base + (half ? 192 : 0) + c * 16, pk);
  }
  ssum[half * AT_BQ + row] = sum;
  tmem_st_wait();
  ptx::tc_fence_before();
  __syncthreads();

  // ---- O = P V into TMEM columns [AT_OCOL, AT_OCOL + DP)
  if (tid == 0) {
    if (PRE) ptx::mbar_wait(&bars[1], 0);
    ptx::tc_fence_after();
    const uint32_t idesc = ptx::idesc_bf16_f32(AT_BQ, DP) | (1u << 16);  // B (= V) is MN-major
    const int ksteps = nk16 >> 4;
    for (int k = 0; k < ksteps; ++k) {
      const uint32_t a = tb + (k < 8 ? 8 * k : 192 + 8 * (k - 8));
      // LBO: next 8 keys, SBO: next 8 head-dim columns
      const uint64_t db = desc_nosw(ptx::smem_u32(sV + (k >> 3) * BLOCK + (k & 7) * 256), 128, CS);
      mma_f16_ts(tb + AT_OCOL, a, db, idesc, k != 0);
    }
    ptx::mma_commit(&bars[3]);
  }
  ptx::mbar_wait(&bars[3], 0);
  ptx::tc_fence_after();
  // ---- epilogue: normalise, bf16, stage in the (free) K region, coalesced stores
  const float l = ssum[row] + ssum[AT_BQ + row];
  const float inv = l > 0.f ? 1.f / l : 0.f;
  constexpr int C0 = HD == 64 ? 32 : 48;  // columns [0, C0) -> half 0, [C0, HD) -> half 1
  uint8_t* stg = sK + row * STG;
  {
    uint32_t o[32];
    const int col0 = half ? C0 : 0;
    tmem_ld32(lane_base + AT_OCOL + col0, o);
    ptx::tmem_ld_wait();
    const int nvec = half ? (HD - C0) / 8 : 4;  // 16-byte chunks out of these 32 columns
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (j < nvec) {
        uint4 v;
        v.x = pack_bf16x2(__uint_as_float(o[8 * j]) * inv, __uint_as_float(o[8 * j + 1]) * inv);
        v.y = pack_bf16x2(__uint_as_float(o[8 * j + 2]) * inv, __uint_as_float(o[8 * j + 3]) * inv);
        v.z = pack_bf16x2(__uint_as_float(o[8 * j + 4]) * inv, __uint_as_float(o[8 * j + 5]) * inv);
        v.w = pack_bf16x2(__uint_as_float(o[8 * j + 6]) * inv, __uint_as_float(o[8 * j + 7]) * inv);
        *reinterpret_cast<uint4*>(stg + (col0 + 8 * j) * 2) = v;
      }
    }
    if (HD == 72 && half == 0) {  // columns [32, 48)
      uint32_t o2[16];
      ptx::tmem_ld16(lane_base + AT_OCOL + 32, o2);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        uint4 v;
        v.x = pack_bf16x2(__uint_as_float(o2[8 * j]) * inv, __uint_as_float(o2[8 * j + 1]) * inv);
        v.y = pack_bf16x2(__uint_as_float(o2[8 * j + 2]) * inv, __uint_as_float(o2[8 * j + 3]) * inv);
        v.z = pack_bf16x2(__uint_as_float(o2[8 * j + 4]) * inv, __uint_as_float(o2[8 * j + 5]) * inv);
        v.w = pack_bf16x2(__uint_as_float(o2[8 * j + 6]) * inv, __uint_as_float(o2[8 * j + 7]) * inv);
        *reinterpret_cast<uint4*>(stg + (32 + 8 * j) * 2) = v;
      }
    }
  }
  if (lse && half == 0 && q0 + row < Tn)
    lse[((size_t)r * h + head) * Tn + q0 + row] = l > 0.f ? (msc + log2f(l)) * 0.6931471805599453f : 0.f;
  ptx::tc_fence_before();
  __syncthreads();
  for (int i = tid; i < AT_BQ * CH; i += AT_THREADS) {
    const int rw = i / CH, c = i - rw * CH;
    if (q0 + rw < Tn)
      *reinterpret_cast<uint4*>(out + ((size_t)r * Tn + q0 + rw) * D + head * HD + c * 8) = *reinterpret_cast<const uint4*>(sK + rw * STG + c * 16);
  }
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tb, AT_TMEM_COLS);
  }
}

template <int HD, bool PRE>
static int launch_attn_tc(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask, int R, int Tn,
                          int h, void* out, float* lse, cudaStream_t s) {
  constexpr int CHP = (HD + 15) / 16 * 2, CH = HD / 8;
  constexpr int SMEM = 5 * CHP * 128 * 16 + 4 * AT_BQ * 4 + 8 * 4 + 4 * 8 + 16;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(attn_tc_kernel<HD, PRE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) { set_last_error("attn_tc: cannot set %d bytes of dynamic smem: %s", SMEM, cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    attr_done = true;
  }
  CUtensorMap tm;
  memset(&tm, 0, sizeof(tm));
  if (PRE) {
    // qkv [R*T, 3*h*d] bf16 seen as {8 elements, R*T rows, 3*h*d/8 chunks}: a box {8, 128, d/8} is one operand block
    AtEncodeTiledFn enc = at_get_encode();
    if (!enc) { set_last_error("attn_tc: cuTensorMapEncodeTiled not available from the driver"); return FITB200_ERR_CUDA; }
    const int D3 = 3 * h * HD;
    cuuint64_t dims[3] = {8, (cuuint64_t)R * Tn, (cuuint64_t)(D3 / 8)};
    cuuint64_t strides[2] = {(cuuint64_t)D3 * 2, 16};
    cuuint32_t box[3] = {8, 128, (cuuint32_t)CH};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult rc = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(qkv), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) { set_last_error("attn_tc: cuTensorMapEncodeTiled failed (%d)", (int)rc); return FITB200_ERR_CUDA; }
  }
  dim3 grid((unsigned)ceil_div(Tn, AT_BQ), (unsigned)h, (unsigned)R);
  attn_tc_kernel<HD, PRE><<<grid, AT_THREADS, SMEM, s>>>(tm, (const __nv_bfloat16*)qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h,
                                                         (__nv_bfloat16*)out, lse);
  FITB200_CHECK_LAUNCH("attn_tc");
  return FITB200_OK;
}

int attn_fwd_tc2(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s);
int attn_fwd_tc3(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s);

// bf16, T <= 256: the tcgen05 kernel.  Called from fitb200_attn_fwd (attn_mma.cu).  prerotated: q, k already carry the RoPE.
int attn_fwd_tc(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d,
                int prerotated, void* out, float* lse, cudaStream_t s) {
  if (prerotated) {
    if (((uintptr_t)qkv & 15) != 0) { set_last_error("attn_fwd: qkv must be 16-byte aligned"); return FITB200_ERR_INVALID_ARG; }
    // persistent warp-specialised schedule (attn_tc2.cu); FITB200_ATTN_V1=1 keeps the one-tile-per-CTA kernel below
    static const bool v1 = [] { const char* e = getenv("FITB200_ATTN_V1"); return e && e[0] == '1'; }();
    // FITB200_ATTN_TC3=1: the KV-blocked kernel of the long sequences (attn_tc3.cu) on these shapes too (A/B timing)
    static const bool v3 = [] { const char* e = getenv("FITB200_ATTN_TC3"); return e && e[0] == '1'; }();
    if (v3 && (d == 64 || d == 72) && ((uintptr_t)out & 15) == 0) return attn_fwd_tc3(qkv, kv_len, mask, R, Tn, h, d, out, lse, s);
    if (!v1 && (d == 64 || d == 72) && ((uintptr_t)out & 15) == 0) return attn_fwd_tc2(qkv, kv_len, mask, R, Tn, h, d, out, lse, s);
    if (d == 64) return launch_attn_tc<64, true>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, out, lse, s);
    if (d == 72) return launch_attn_tc<72, true>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, out, lse, s);
  } else {
    if (d == 64) return launch_attn_tc<64, false>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, out, lse, s);
    if (d == 72) return launch_attn_tc<72, false>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, out, lse, s);
  }
  set_last_error("attn_fwd: head_dim %d not supported (64 or 72)", d);
  return FITB200_ERR_UNSUPPORTED;
}

}  // namespace fitb200
