// bf16 attention backward on tcgen05 / TMEM for sequences of up to 256 tokens (pre-rotated q, k: the bf16 training path).
//
//   P = exp(q.k^T / sqrt(d) - lse)     dP = dO V^T     delta_i = sum_d dO_id O_id     dS = P (dP - delta) / sqrt(d)
//   dV = P^T dO      dK = dS^T Q      dQ = dS K        then the RoPE's transpose (rotation by -theta) on dQ, dK
//
// Two roles, one CTA = 128 rows of one (sample, head), 512 TMEM columns (one CTA per SM):
//   ROLE 0 (rows = keys):    S^T = Kt Q^T and dP^T = Vt dO^T into TMEM; each thread owns one key row: P^T, dS^T (bf16 pairs) written
//                            back over consumed columns; dV = P^T dO and dK = dS^T Q with the A operand FROM TMEM and dO / Q read
//                            MN-major from the very tiles that fed the first two products.
//   ROLE 1 (rows = queries): S = Qt K^T, dP = dOt V^T; dS in TMEM; dQ = dS K.
// Recomputing S / dP once per role costs 7 instead of 5 small matmuls but needs no atomics and no transposes; operands arrive by
// 3-D TMA boxes straight into the no-swizzle core-matrix layout (see attn_tc.cu).  Only keys < kv_len and the queries that took
// part in the forward are multiplied.  Replaces the CUDA-core kernel of attn_bwd.cu on the bf16 path (63 + 33 ms -> see DESIGN.md).
#include <cstring>

#include "common.cuh"
#include "ptx.cuh"
#include "umma_nosw.cuh"

namespace fitb200 {

constexpr int BT_ROWS = 128;
constexpr int BT_THREADS = 256;
constexpr int BT_QSKIP = 128;  // forward skips query tiles of 128 at or beyond kv_len

template <int HD, int ROLE>
__global__ void __launch_bounds__(BT_THREADS, 1) attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                                                                   const float* __restrict__ freqs, int freqs_rows,
                                                                   const int32_t* __restrict__ kv_len, const uint8_t* __restrict__ mask,
                                                                   int R, int Tn, int h, const __nv_bfloat16* __restrict__ o,
                                                                   const __nv_bfloat16* __restrict__ dO, const float* __restrict__ lse,
                                                                   __nv_bfloat16* __restrict__ dqkv) {
  constexpr int DP = (HD + 15) / 16 * 16, CHP = DP / 8, CH = HD / 8;
  constexpr int CS = BT_ROWS * 16;     // chunk-column stride inside a 128-row block
  constexpr int BLOCK = CHP * CS;      // bytes of one 128-row block
  constexpr int STG = 144;

  extern __shared__ __align__(128) uint8_t bt_smem[];
  uint8_t* sRowA = bt_smem;              // ROLE 0: K tile   ROLE 1: Q tile     (A of product 1)
  uint8_t* sRowB = sRowA + BLOCK;        // ROLE 0: V tile   ROLE 1: dO tile    (A of product 2)
  uint8_t* sColA = sRowB + BLOCK;        // ROLE 0: Q        ROLE 1: K          (B of product 1; two 128-row blocks)
  uint8_t* sColB = sColA + 2 * BLOCK;    // ROLE 0: dO       ROLE 1: V          (B of product 2)
  float* lse2 = reinterpret_cast<float*>(sColB + 2 * BLOCK);  // [256] lse * log2(e) per query
  float* delta = lse2 + 256;                                   // [256]
  uint32_t* vw = reinterpret_cast<uint32_t*>(delta + 256);     // [8] key-validity bits
  uint64_t* bars = reinterpret_cast<uint64_t*>(vw + 8);        // [3]: operands landed, products 1+2 done, products 3 done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const int r = blockIdx.z, head = blockIdx.y, tile0 = blockIdx.x * BT_ROWS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int D = h * HD;
  const size_t rs3 = (size_t)3 * D;
  const int kvl = min(kv_len[r], Tn);
  const int q_end = min(Tn, ceil_div(kvl, BT_QSKIP) * BT_QSKIP);
  const int ncol = ROLE == 0 ? q_end : kvl;            // columns of the score tile: queries (ROLE 0) or keys (ROLE 1)
  const int ncol16 = (ncol + 15) & ~15;
  const int n0 = min(ncol16, 128), n1 = ncol16 - n0;
  const bool tile_live = ROLE == 0 ? tile0 < kvl : tile0 < q_end;
  __nv_bfloat16* gout = dqkv + (size_t)r * Tn * rs3 + (size_t)head * HD;  // q part; k at +D, v at +2D

  if (!tile_live) {  // rows that never took part in the forward: zero gradient
    for (int i = tid; i < BT_ROWS * CH; i += BT_THREADS) {
      const int row = i / CH, c = i - row * CH, t = tile0 + row;
      if (t < Tn) {
        const uint4 z = make_uint4(0, 0, 0, 0);
        if (ROLE == 0) { *reinterpret_cast<uint4*>(gout + (size_t)t * rs3 + D + c * 8) = z; *reinterpret_cast<uint4*>(gout + (size_t)t * rs3 + 2 * D + c * 8) = z; }
        else *reinterpret_cast<uint4*>(gout + (size_t)t * rs3 + c * 8) = z;
      }
    }
    return;
  }

  if (tid == 0) {
    for (int i = 0; i < 3; ++i) ptx::mbar_init(&bars[i], 1);
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&tmQKV);
    ptx::prefetch_tmap(&tmDO);
    const int row0 = r * Tn, cq = head * CH, ck = (D + head * HD) / 8, cv = (2 * D + head * HD) / 8, cdo = head * CH;
    const uint32_t box = CH * CS;
    ptx::mbar_arrive_expect_tx(&bars[0], box * (n1 > 0 ? 6 : 4));
    if (ROLE == 0) {
      tma_load_3d(sRowA, &tmQKV, &bars[0], 0, row0 + tile0, ck);
      tma_load_3d(sRowB, &tmQKV, &bars[0], 0, row0 + tile0, cv);
      tma_load_3d(sColA, &tmQKV, &bars[0], 0, row0, cq);
      tma_load_3d(sColB, &tmDO, &bars[0], 0, row0, cdo);
      if (n1 > 0) { tma_load_3d(sColA + BLOCK, &tmQKV, &bars[0], 0, row0 + 128, cq); tma_load_3d(sColB + BLOCK, &tmDO, &bars[0], 0, row0 + 128, cdo); }
    } else {
      tma_load_3d(sRowA, &tmQKV, &bars[0], 0, row0 + tile0, cq);
      tma_load_3d(sRowB, &tmDO, &bars[0], 0, row0 + tile0, cdo);
      tma_load_3d(sColA, &tmQKV, &bars[0], 0, row0, ck);
      tma_load_3d(sColB, &tmQKV, &bars[0], 0, row0, cv);
      if (n1 > 0) { tma_load_3d(sColA + BLOCK, &tmQKV, &bars[0], 0, row0 + 128, ck); tma_load_3d(sColB + BLOCK, &tmQKV, &bars[0], 0, row0 + 128, cv); }
    }
  }
  if (DP != HD) {  // zero the contraction padding chunk of all six blocks
    for (int i = tid; i < 6 * BT_ROWS; i += BT_THREADS) {
      const int blk = i >> 7, row = i & 127;
      *reinterpret_cast<uint4*>(bt_smem + blk * BLOCK + CH * CS + row * 16) = make_uint4(0, 0, 0, 0);
    }
  }
  {  // per-query statistics: lse (log2 domain) and delta = rowsum(dO * O)
    const int qi = tid;
    float l2 = 0.f, dl = 0.f;
    if (qi < q_end) {
      l2 = lse[((size_t)r * h + head) * Tn + qi] * 1.4426950408889634f;
      const __nv_bfloat16* orow = o + ((size_t)r * Tn + qi) * D + head * HD;
      const __nv_bfloat16* drow = dO + ((size_t)r * Tn + qi) * D + head * HD;
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const Vec8<__nv_bfloat16> a = Vec8<__nv_bfloat16>::load(orow + c * 8), b = Vec8<__nv_bfloat16>::load(drow + c * 8);
#pragma unroll
        for (int e = 0; e < 8; ++e) dl = fmaf(a.v[e], b.v[e], dl);
      }
    }
    lse2[qi] = l2;
    delta[qi] = dl;
    const int key = warp * 32 + lane;
    const bool v = key < kvl && (mask == nullptr || mask[(size_t)r * Tn + key] != 0);
    const uint32_t w = __ballot_sync(0xffffffffu, v);
    if (lane == 0) vw[warp] = w;
  }
  if (warp == 0) { ptx::tmem_alloc(tmem_slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = *tmem_slot;

  // ---- products 1 and 2: X1 = RowA ColA^T -> columns [0, 256), X2 = RowB ColB^T -> columns [256, 512)
  if (tid == 0) {
    ptx::mbar_wait(&bars[0], 0);
    ptx::tc_fence_after();
#pragma unroll
    for (int prod = 0; prod < 2; ++prod) {
      const uint8_t* sa = prod ? sRowB : sRowA;
      const uint8_t* sb = prod ? sColB : sColA;
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const int nn = hf ? n1 : n0;
        if (nn > 0) {
          const uint32_t idesc = ptx::idesc_bf16_f32(BT_ROWS, nn);
#pragma unroll
          for (int k = 0; k < DP / 16; ++k) {
            const uint64_t da = desc_nosw(ptx::smem_u32(sa + 2 * k * CS), CS, 128);
            const uint64_t db = desc_nosw(ptx::smem_u32(sb + hf * BLOCK + 2 * k * CS), CS, 128);
            ptx::mma_f16_ss(tb + prod * 256 + hf * 128, da, db, idesc, k != 0);
          }
        }
      }
    }
    ptx::mma_commit(&bars[1]);
  }
  ptx::mbar_wait(&bars[1], 0);
  ptx::tc_fence_after();

  // ---- P and dS: thread = (row, half of the columns); results go back to TMEM as bf16 pairs over consumed columns
  const int quarter = warp & 3, half = warp >> 2, row = quarter * 32 + lane, trow = tile0 + row;
  const uint32_t lane_base = tb + ((uint32_t)(quarter * 32) << 16);
  const int n_my = max(0, min(ncol16 - half * 128, 128));
  const int nchunks = (n_my + 31) >> 5;
  const float scale = rsqrtf((float)HD), sl2 = scale * 1.4426950408889634f;
  bool row_live;
  float row_l2 = 0.f, row_dl = 0.f;
  if (ROLE == 1) { row_live = trow < q_end; row_l2 = lse2[min(trow, 255)]; row_dl = delta[min(trow, 255)]; }
  else row_live = trow < 256 && ((vw[min(trow, 255) >> 5] >> (trow & 31)) & 1u);
  for (int cc = 0; cc < nchunks; ++cc) {
    const int c = half ? nchunks - 1 - cc : cc;
    const int col0 = half * 128 + c * 32;  // first column (query for ROLE 0, key for ROLE 1) of this chunk
    uint32_t x1[32], x2[32];
    tmem_ld32(lane_base + col0, x1);
    tmem_ld32(lane_base + 256 + col0, x2);
    ptx::tmem_ld_wait();
    const uint32_t kbits = ROLE == 1 ? vw[half * 4 + c] : 0u;
    uint32_t pp[16], ps[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float pv[2], dv[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = 2 * i + e, col = col0 + j;
        const bool live = row_live && (ROLE == 1 ? ((kbits >> j) & 1u) != 0 : col < q_end);
        const float l2 = ROLE == 1 ? row_l2 : lse2[col];
        const float dl = ROLE == 1 ? row_dl : delta[col];
        const float p = live ? ex2f(fmaf(__uint_as_float(x1[j]), sl2, -l2)) : 0.f;
        pv[e] = p;
        dv[e] = live ? p * (__uint_as_float(x2[j]) - dl) * scale : 0.f;
      }
      pp[i] = pack_bf16x2(pv[0], pv[1]);
      ps[i] = pack_bf16x2(dv[0], dv[1]);
    }
    const uint32_t slot = lane_base + (half ? 192 : 0) + c * 16;
    if (ROLE == 0) { tmem_st16(slot, pp); tmem_st16(slot + 256, ps); }
    else tmem_st16(slot, ps);
  }
  tmem_st_wait();
  ptx::tc_fence_before();
  __syncthreads();

  // ---- products 3: A from TMEM, B = column operand read MN-major (LBO: next 8 rows, SBO: next 8 head-dim columns)
  constexpr int OCOL = 64;
  if (tid == 0) {
    ptx::tc_fence_after();
    const uint32_t idesc = ptx::idesc_bf16_f32(BT_ROWS, DP) | (1u << 16);
    const int ksteps = ncol16 >> 4;
    for (int k = 0; k < ksteps; ++k) {
      const uint32_t a = tb + (k < 8 ? 8 * k : 192 + 8 * (k - 8));
      const uint32_t boff = (k >> 3) * BLOCK + (k & 7) * 256;
      if (ROLE == 0) {
        mma_f16_ts(tb + OCOL, a, desc_nosw(ptx::smem_u32(sColB + boff), 128, CS), idesc, k != 0);              // dV = P^T dO
        mma_f16_ts(tb + 256 + OCOL, a + 256, desc_nosw(ptx::smem_u32(sColA + boff), 128, CS), idesc, k != 0);  // dK = dS^T Q
      } else {
        mma_f16_ts(tb + OCOL, a, desc_nosw(ptx::smem_u32(sColA + boff), 128, CS), idesc, k != 0);              // dQ = dS K
      }
    }
    ptx::mma_commit(&bars[2]);
  }
  ptx::mbar_wait(&bars[2], 0);
  ptx::tc_fence_after();

  // ---- epilogue: (RoPE transposed on dQ / dK) -> bf16 -> staging rows -> coalesced stores
  constexpr int C0 = HD == 64 ? 32 : 48;
  const int col_lo = half ? C0 : 0, ncols = half ? HD - C0 : C0;  // this thread's head-dim range: 32 / 32 or 48 / 24 columns
  const float* fr = freqs + ((size_t)(freqs_rows == 1 ? 0 : r) * Tn + min(trow, Tn - 1)) * HD;
  auto drain = [&](uint32_t acc_col, bool rotate, uint8_t* stage) {
#pragma unroll
    for (int part = 0; part < 2; ++part) {  // up to 48 columns: 32 + 16
      const int pc = part * 32;
      if (pc >= ncols) break;
      uint32_t v[32];
      if (part == 0) tmem_ld32(lane_base + acc_col + col_lo, v);
      else { uint32_t v16[16]; ptx::tmem_ld16(lane_base + acc_col + col_lo + 32, v16);
#pragma unroll
             for (int i = 0; i < 16; ++i) v[i] = v16[i]; }
      ptx::tmem_ld_wait();
      const int nhere = min(32, ncols - pc);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j * 8 < nhere) {
          float g[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) g[e] = __uint_as_float(v[j * 8 + e]);
          if (rotate) {
            const float4 f0 = __ldg(reinterpret_cast<const float4*>(fr + col_lo + pc + j * 8));
            const float4 f1 = __ldg(reinterpret_cast<const float4*>(fr + col_lo + pc + j * 8 + 4));
            const float cs[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float a = g[2 * e], b = g[2 * e + 1], c = cs[2 * e], s = cs[2 * e + 1];
              g[2 * e] = a * c + b * s;
              g[2 * e + 1] = b * c - a * s;
            }
          }
          uint4 ov;
          ov.x = pack_bf16x2(g[0], g[1]); ov.y = pack_bf16x2(g[2], g[3]); ov.z = pack_bf16x2(g[4], g[5]); ov.w = pack_bf16x2(g[6], g[7]);
          *reinterpret_cast<uint4*>(stage + row * STG + (col_lo + pc + j * 8) * 2) = ov;
        }
      }
    }
  };
  if (ROLE == 0) { drain(OCOL, false, sColB); drain(256 + OCOL, true, sColA); }
  else drain(OCOL, true, sColA);
  ptx::tc_fence_before();
  __syncthreads();
  for (int i = tid; i < BT_ROWS * CH; i += BT_THREADS) {
    const int rw = i / CH, c = i - rw * CH, t = tile0 + rw;
    if (t < Tn) {
      if (ROLE == 0) {
        *reinterpret_cast<uint4*>(gout + (size_t)t * rs3 + D + c * 8) = *reinterpret_cast<const uint4*>(sColA + rw * STG + c * 16);
        *reinterpret_cast<uint4*>(gout + (size_t)t * rs3 + 2 * D + c * 8) = *reinterpret_cast<const uint4*>(sColB + rw * STG + c * 16);
      } else {
        *reinterpret_cast<uint4*>(gout + (size_t)t * rs3 + c * 8) = *reinterpret_cast<const uint4*>(sColA + rw * STG + c * 16);
      }
    }
  }
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tb, 512);
  }
}

template <int HD>
static int launch_attn_bwd_tc(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h,
                              const void* o, const void* dO, const float* lse, void* dqkv, cudaStream_t s) {
  constexpr int CHP = (HD + 15) / 16 * 2, CH = HD / 8;
  constexpr int SMEM = 6 * CHP * 128 * 16 + 2 * 256 * 4 + 8 * 4 + 3 * 8 + 16;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e0 = cudaFuncSetAttribute(attn_bwd_tc_kernel<HD, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    cudaError_t e1 = cudaFuncSetAttribute(attn_bwd_tc_kernel<HD, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e0 != cudaSuccess || e1 != cudaSuccess) { set_last_error("attn_bwd_tc: cannot set %d bytes of dynamic smem", SMEM); return FITB200_ERR_CUDA; }
    attr_done = true;
  }
  CUtensorMap tmQKV, tmDO;
  memset(&tmQKV, 0, sizeof(tmQKV)); memset(&tmDO, 0, sizeof(tmDO));
  const int D = h * HD;
  if (at_make_tmap3(&tmQKV, qkv, (long)R * Tn, 3L * D, CH) != 0 || at_make_tmap3(&tmDO, dO, (long)R * Tn, D, CH) != 0) {
    set_last_error("attn_bwd_tc: cuTensorMapEncodeTiled failed");
    return FITB200_ERR_CUDA;
  }
  dim3 grid((unsigned)ceil_div(Tn, BT_ROWS), (unsigned)h, (unsigned)R);
  attn_bwd_tc_kernel<HD, 0><<<grid, BT_THREADS, SMEM, s>>>(tmQKV, tmDO, freqs, freqs_rows, kv_len, mask, R, Tn, h, (const __nv_bfloat16*)o,
                                                           (const __nv_bfloat16*)dO, lse, (__nv_bfloat16*)dqkv);
  attn_bwd_tc_kernel<HD, 1><<<grid, BT_THREADS, SMEM, s>>>(tmQKV, tmDO, freqs, freqs_rows, kv_len, mask, R, Tn, h, (const __nv_bfloat16*)o,
                                                           (const __nv_bfloat16*)dO, lse, (__nv_bfloat16*)dqkv);
  FITB200_CHECK_LAUNCH("attn_bwd_tc");
  return FITB200_OK;
}

// bf16, pre-rotated q / k, T <= 256.  Called from fitb200_attn_bwd (attn_bwd.cu).
int attn_bwd_tc(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d,
                const void* o, const void* dO, const float* lse, void* dqkv, cudaStream_t s) {
  if (d == 64) return launch_attn_bwd_tc<64>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, o, dO, lse, dqkv, s);
  if (d == 72) return launch_attn_bwd_tc<72>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, o, dO, lse, dqkv, s);
  set_last_error("attn_bwd: head_dim %d not supported (64 or 72)", d);
  return FITB200_ERR_UNSUPPORTED;
}

}  // namespace fitb200
