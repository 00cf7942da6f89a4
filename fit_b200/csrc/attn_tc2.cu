// Persistent, warp-specialised bf16 attention forward on tcgen05 / TMEM for sequences of up to 256 tokens whose q and k already carry
// the RoPE (qkv GEMM epilogue FITB200_EPI_BIAS_ROPE).  Same arithmetic as attn_tc.cu (Attention.forward, fit.py:176-188; head merge
// fit.py:220-227); what changes is the schedule.  attn_tc.cu runs one 128-query tile per CTA as a serial chain
// load -> S -> softmax -> PV -> store and relies on two resident CTAs for overlap (0.45-0.57 ms at R=512 XL/2, 10 us per tile, K / V
// fetched once per query tile).  Here one CTA per SM walks (sample, head) items:
//   warp 0      TMA producer: per item K (k_full) and V (v_full) once, per 128-query unit Q (q_full); rings: Q x3, K x2, V x2
//   warps 1, 2  MMA issuers, one per softmax group (its units only, blocking waits): S(u) = Q K^T into TMEM region u&1, then P(u) V -> O(u)
//   warps 3-10  softmax / epilogue group 0 (units with even running index), warps 11-18 group 1 (odd): two-pass row-per-thread
//               softmax on TMEM lanes, P back to TMEM as bf16, O normalised and staged in the unit's (already consumed) Q buffer,
//               one TMA store per unit (3-D map, rows clipped at the sample's token count)
// TMEM: region g = columns [256 g, 256 g + 256): S in [0, n), P over [0, 64) and [192, 256), O in [64, 64 + DP)  (as attn_tc.cu).
// Only keys < kv_len are multiplied (MMA N / K = round16(kv_len)); query tiles at or beyond kv_len are skipped (zero rows).
#include <cstring>

#include "common.cuh"
#include "ptx.cuh"
#include "umma_nosw.cuh"

namespace fitb200 {

constexpr int A2_BQ = 128;
constexpr int A2_GROUP = 256;                    // threads of one softmax group (8 warps)
constexpr int A2_THREADS = 96 + 2 * A2_GROUP;    // producer warp + one MMA warp per group + two groups
constexpr int A2_OCOL = 64;
constexpr int A2_NQ = 3;                         // Q ring depth (units)

// 3-D TMA store {cols, rows, samples} <- dense [rows][box cols] staging
__device__ __forceinline__ void a2_tma_store_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];\n" ::"l"(reinterpret_cast<uint64_t>(m)),
               "r"(ptx::smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

template <int HD>
struct A2Cfg {
  static constexpr int DP = (HD + 15) / 16 * 16;      // contraction width of q.k^T and N of p.v (72 -> 80)
  static constexpr int CHP = DP / 8, CH = HD / 8;
  static constexpr int CS = 128 * 16;                  // chunk-column stride inside a 128-row block
  static constexpr int BLOCK = CHP * CS;               // bytes of one 128-row operand block (incl. the contraction pad chunk)
  static constexpr int Q_OFF = 0;
  static constexpr int K_OFF = A2_NQ * BLOCK;
  static constexpr int V_OFF = K_OFF + 2 * 2 * BLOCK;
  static constexpr int SCR_OFF = V_OFF + 2 * 2 * BLOCK;   // per group: smax[2][128], ssum[2][128], vw[8]
  static constexpr int SCR_BYTES = 4 * A2_BQ * 4 + 32;
  static constexpr int BAR_OFF = SCR_OFF + 2 * SCR_BYTES;
  static constexpr int NBARS = 2 * A2_NQ + 6 + 8;         // q_full/q_empty, k_full/v_full/kv_empty x2, s_full/p_ready/o_full/t_empty x2
  static constexpr int SMEM = BAR_OFF + NBARS * 8 + 16 + 1024 /*alignment slack*/;
  static constexpr int OUT_ROW = HD * 2;                  // staging row bytes (dense TMA box)
};

template <int HD>
__global__ void __launch_bounds__(A2_THREADS, 1)
attn_tc2_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmOut, const int32_t* __restrict__ kv_len,
                const uint8_t* __restrict__ mask, int R, int Tn, int h, __nv_bfloat16* __restrict__ out, float* __restrict__ lse) {
  using C = A2Cfg<HD>;
  constexpr int DP = C::DP, CH = C::CH, CS = C::CS, BLOCK = C::BLOCK;
  extern __shared__ uint8_t a2_raw[];
  uint8_t* sm = a2_raw + ((1024u - (ptx::smem_u32(a2_raw) & 1023u)) & 1023u);
  uint8_t* sQ = sm + C::Q_OFF;   // [A2_NQ][CHP][128][16]
  uint8_t* sK = sm + C::K_OFF;   // [2 items][2 halves][CHP][128][16]
  uint8_t* sV = sm + C::V_OFF;   // [2 items][2 halves][CHP][128][16]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + C::BAR_OFF);
  uint64_t* q_full = bars;
  uint64_t* q_empty = bars + A2_NQ;
  uint64_t* k_full = bars + 2 * A2_NQ;
  uint64_t* v_full = k_full + 2;
  uint64_t* kv_empty = k_full + 4;
  uint64_t* s_full = k_full + 6;
  uint64_t* p_ready = s_full + 2;
  uint64_t* o_full = s_full + 4;
  uint64_t* t_empty = s_full + 6;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + C::NBARS);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_items = R * h;
  const int upi = (Tn + A2_BQ - 1) / A2_BQ;  // query units per item (1 or 2)
  const int D = h * HD;

  if (tid == 0) {
    ptx::prefetch_tmap(&tmQKV);
    ptx::prefetch_tmap(&tmOut);
    for (int i = 0; i < A2_NQ; ++i) { ptx::mbar_init(&q_full[i], 1); ptx::mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&k_full[i], 1); ptx::mbar_init(&v_full[i], 1); ptx::mbar_init(&kv_empty[i], 2);  // one arrival per MMA warp
      ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&p_ready[i], A2_GROUP); ptx::mbar_init(&o_full[i], 1); ptx::mbar_init(&t_empty[i], A2_GROUP);
    }
    ptx::fence_mbar_init();
  }
  if (DP != HD) {  // contraction padding (head_dim 72 -> 80) of every Q / K block: zeroed once, TMA boxes never touch it
    for (int i = tid; i < (A2_NQ + 4) * 128; i += A2_THREADS) {
      const int b = i >> 7, rr = i & 127;
      uint8_t* blk = b < A2_NQ ? sQ + b * BLOCK : sK + (b - A2_NQ) * BLOCK;
      *reinterpret_cast<uint4*>(blk + CH * CS + rr * 16) = make_uint4(0, 0, 0, 0);
    }
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = *tmem_slot;
  constexpr uint32_t BOX = CH * CS;  // bytes of one TMA box {8, 128, CH}

  if (warp == 0) {
    // ------------------------------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int uq = 0, ic = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int r = item / h, head = item - r * h;
        const int kvl = min(kv_len[r], Tn);
        const int nlive = min(upi, (kvl + A2_BQ - 1) / A2_BQ);
        if (nlive == 0) continue;
        const int nk16 = (kvl + 15) & ~15, n1 = nk16 - min(nk16, 128);
        const int slot = ic & 1; const uint32_t ph = (ic >> 1) & 1;
        const int row0 = r * Tn, cq = head * CH, ck = (D + head * HD) / 8, cv = (2 * D + head * HD) / 8;
        for (int qt = 0; qt < nlive; ++qt, ++uq) {
          const int qb = uq % A2_NQ; const uint32_t qph = (uq / A2_NQ) & 1;
          ptx::mbar_wait_sleep(&q_empty[qb], qph ^ 1, 64);
          ptx::mbar_arrive_expect_tx(&q_full[qb], BOX);
          tma_load_3d(sQ + qb * BLOCK, &tmQKV, &q_full[qb], 0, row0 + qt * A2_BQ, cq);
          if (qt == 0) {
            ptx::mbar_wait_sleep(&kv_empty[slot], ph ^ 1, 64);
            ptx::mbar_arrive_expect_tx(&k_full[slot], BOX * (n1 > 0 ? 2 : 1));
            tma_load_3d(sK + slot * 2 * BLOCK, &tmQKV, &k_full[slot], 0, row0, ck);
            if (n1 > 0) tma_load_3d(sK + slot * 2 * BLOCK + BLOCK, &tmQKV, &k_full[slot], 0, row0 + 128, ck);
            ptx::mbar_arrive_expect_tx(&v_full[slot], BOX * (n1 > 0 ? 2 : 1));
            tma_load_3d(sV + slot * 2 * BLOCK, &tmQKV, &v_full[slot], 0, row0, cv);
            if (n1 > 0) tma_load_3d(sV + slot * 2 * BLOCK + BLOCK, &tmQKV, &v_full[slot], 0, row0 + 128, cv);
          }
        }
        ++ic;
      }
    }
    __syncwarp();
  } else if (warp <= 2) {
    // ------------------------------------------------------------------------------------------ MMA issuer of group warp - 1
    // One issuing warp PER GROUP: a single thread serving both groups was the bottleneck of the kernel -- its ~500 instructions per
    // unit pair (descriptor arithmetic, polling of six barriers) run at a fifth of an issue port it shares with four saturated
    // softmax warps, so P V and S went out ~2000 clk after their barriers had flipped (clock64 trace, profiles/r1g_attn_tc3.txt).
    // Each warp now walks only its own units with blocking waits, descriptors are advanced by constants, loops are unrolled.
    if (lane == 0) {
      const int g = warp - 1;
      const uint32_t tr = tb + g * 256;
      // cursor over the live units in issue order; the NEXT unit of this group (incl. its kv_len load) is looked up while the softmax of
      // the current one runs, and everything that does not depend on the awaited barrier is computed before the wait: this thread gets
      // an issue slot only every ~20 clk next to the softmax warps of its scheduler, so each instruction after a barrier flip counts
      struct It { int item, qt, nlive, uq, ic, nk16; bool valid; };
      auto step = [&](It& it) {
        if (it.qt + 1 < it.nlive) { ++it.qt; ++it.uq; return; }
        for (;;) {
          it.item += gridDim.x;
          if (it.item >= n_items) { it.valid = false; return; }
          const int kvl = min(kv_len[it.item / h], Tn);
          const int nlive = min(upi, (kvl + A2_BQ - 1) / A2_BQ);
          if (nlive == 0) continue;
          it.nlive = nlive; it.nk16 = (kvl + 15) & ~15; it.qt = 0; ++it.uq; ++it.ic;
          return;
        }
      };
      auto next_mine = [&](It& it) { do step(it); while (it.valid && (it.uq & 1) != g); };
      It cur;
      cur.item = (int)blockIdx.x - (int)gridDim.x; cur.qt = 0; cur.nlive = 0; cur.uq = -1; cur.ic = -1; cur.nk16 = 0; cur.valid = true;
      next_mine(cur);
      constexpr uint64_t KSTEP = (2 * CS) >> 4, HALF = BLOCK >> 4;   // descriptor start-address field is in 16-byte units
      const uint32_t idpv = ptx::idesc_bf16_f32(A2_BQ, DP) | (1u << 16);  // B (= V) is MN-major
      while (cur.valid) {
        const int uq = cur.uq, slot = cur.ic & 1, qb = uq % A2_NQ;
        const uint32_t ph = (cur.ic >> 1) & 1;
        const int n0 = min(cur.nk16, 128), n1 = cur.nk16 - n0, ksteps = cur.nk16 >> 4, nlive = cur.nlive;
        const uint64_t da0 = desc_nosw(ptx::smem_u32(sQ + qb * BLOCK), CS, 128);
        const uint64_t db0 = desc_nosw(ptx::smem_u32(sK + slot * 2 * BLOCK), CS, 128);
        const uint64_t dv0 = desc_nosw(ptx::smem_u32(sV + slot * 2 * BLOCK), 128, CS);
        const uint32_t id0 = ptx::idesc_bf16_f32(A2_BQ, n0), id1 = ptx::idesc_bf16_f32(A2_BQ, n1 > 0 ? n1 : 16);
        ptx::mbar_wait(&q_full[qb], (uq / A2_NQ) & 1);      // operands first (they land early), the TMEM region last
        ptx::mbar_wait(&k_full[slot], ph);
        ptx::mbar_wait(&t_empty[g], ((uq >> 1) & 1) ^ 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < DP / 16; ++k) ptx::mma_f16_ss(tr, da0 + k * KSTEP, db0 + k * KSTEP, id0, k != 0);
        if (n1 > 0) {
#pragma unroll
          for (int k = 0; k < DP / 16; ++k) ptx::mma_f16_ss(tr + 128, da0 + k * KSTEP, db0 + HALF + k * KSTEP, id1, k != 0);
        }
        ptx::mma_commit(&s_full[g]);
        It nxt = cur;
        next_mine(nxt);                                       // hidden under the softmax of this unit
        ptx::mbar_wait(&v_full[slot], ph);
        ptx::mbar_wait(&p_ready[g], (uq >> 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          if (k < ksteps)
            mma_f16_ts(tr + A2_OCOL, tr + (k < 8 ? 8 * k : 192 + 8 * (k - 8)), dv0 + (k >> 3) * HALF + (k & 7) * 16, idpv, k != 0);
        }
        ptx::mma_commit(&o_full[g]);
        // K and V of the item are free once both groups' P V have retired (an item with one live unit arrives for both)
        ptx::mma_commit(&kv_empty[slot]);
        if (nlive == 1) ptx::mma_commit(&kv_empty[slot]);
        cur = nxt;
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------------------------------ softmax / epilogue groups
    const int g = (warp - 3) >> 3;                 // group 0: warps 3-10, group 1: warps 11-18
    const int gt = tid - 96 - g * A2_GROUP;        // thread inside the group
    const int gw = gt >> 5;                        // warp inside the group
    const int quarter = warp & 3, half = gw >> 2;  // TMEM lane quarter this warp may access; which half of the keys / columns
    const int row = quarter * 32 + lane;
    float* smax = reinterpret_cast<float*>(sm + C::SCR_OFF + g * C::SCR_BYTES);  // [2][128]
    float* ssum = smax + 2 * A2_BQ;                                                // [2][128]
    uint32_t* vw = reinterpret_cast<uint32_t*>(ssum + 2 * A2_BQ);                  // [8] key-validity bits
    const uint32_t lane_base = tb + g * 256 + ((uint32_t)(quarter * 32) << 16);
    const float sl2 = rsqrtf((float)HD) * 1.4426950408889634f;  // softmax scale folded with log2(e)
    const int bar_id = 1 + g;
    int uq = 0, pend_qb = -1;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const int r = item / h, head = item - r * h;
      const int kvl = min(kv_len[r], Tn);
      const int nlive = min(upi, (kvl + A2_BQ - 1) / A2_BQ);
      if (g == 0) {  // padded query tiles: zero output rows, no tensor work
        for (int qt = nlive; qt < upi; ++qt) {
          const int q0 = qt * A2_BQ;
          for (int i = gt; i < A2_BQ * CH; i += A2_GROUP) {
            const int rw = i / CH, c = i - rw * CH;
            if (q0 + rw < Tn) *reinterpret_cast<uint4*>(out + ((size_t)r * Tn + q0 + rw) * D + head * HD + c * 8) = make_uint4(0, 0, 0, 0);
          }
          if (lse) for (int i = gt; i < A2_BQ; i += A2_GROUP) if (q0 + i < Tn) lse[((size_t)r * h + head) * Tn + q0 + i] = 0.f;
        }
      }
      if (nlive == 0) continue;
      const int nk = kvl, nk16 = (nk + 15) & ~15;
      for (int qt = 0; qt < nlive; ++qt, ++uq) {
        if ((uq & 1) != g) continue;
        const int q0 = qt * A2_BQ, qb = uq % A2_NQ;
        const uint32_t uph = (uq >> 1) & 1;
        if (gt == 0 && pend_qb >= 0) {
          ptx::tma_store_wait_read<0>();
          ptx::mbar_arrive(&q_empty[pend_qb]);  // the producer may refill that Q buffer
          pend_qb = -1;
        }
        {
          const int key = gw * 32 + lane;
          const bool v = key < nk && (mask == nullptr || mask[(size_t)r * Tn + key] != 0);
          const uint32_t w = __ballot_sync(0xffffffffu, v);
          if (lane == 0) vw[gw] = w;
        }
        ptx::named_bar_sync(bar_id, A2_GROUP);
        ptx::mbar_wait(&s_full[g], uph);
        ptx::tc_fence_after();
        const int n_my = max(0, min(nk16 - half * 128, 128));
        const int nchunks = (n_my + 31) >> 5;
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
        smax[half * A2_BQ + row] = m;
        ptx::named_bar_sync(bar_id, A2_GROUP);
        m = fmaxf(smax[row], smax[A2_BQ + row]);
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
          // two whole-chunk code paths behind ONE warp-uniform branch: a per-element `if (w != ~0)` was if-converted into a select plus
          // two integer ops per element on the unmasked path too (FSEL / LOP3 / ISETP each as frequent as the exp2 in the profile)
          if (w == 0xffffffffu) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float p0 = ex2f(fmaf(__uint_as_float(s[2 * i]), sl2, -msc));
              const float p1 = ex2f(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -msc));
              sum += p0 + p1;
              pk[i] = pack_bf16x2(p0, p1);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              float p0 = ex2f(fmaf(__uint_as_float(s[2 * i]), sl2, -msc));
              float p1 = ex2f(fmaf(__uint_as_float(s[2 * i + 1]), sl2, -msc));
              p0 = (w >> (2 * i)) & 1u ? p0 : 0.f;
              p1 = (w >> (2 * i + 1)) & 1u ? p1 : 0.f;
              sum += p0 + p1;
              pk[i] = pack_bf16x2(p0, p1);
            }
          }
          tmem_st16(lane_base + (half ? 192 : 0) + c * 16, pk);
        }
        ssum[half * A2_BQ + row] = sum;
        tmem_st_wait();
        ptx::tc_fence_before();
        ptx::mbar_arrive(&p_ready[g]);

        // ---- epilogue: O / rowsum -> bf16 -> staging (the unit's Q buffer, consumed by S) -> one TMA store
        ptx::mbar_wait(&o_full[g], uph);
        ptx::tc_fence_after();
        const float l = ssum[row] + ssum[A2_BQ + row];
        const float inv = l > 0.f ? 1.f / l : 0.f;
        constexpr int C0 = HD == 64 ? 32 : 48;  // columns [0, C0) -> half 0, [C0, HD) -> half 1
        uint8_t* stg = sQ + qb * BLOCK + row * C::OUT_ROW;
        const uint32_t swz = HD == 64 ? (uint32_t)(row & 7) : 0u;  // 128-byte rows: 128B-swizzled box (conflict-free 16-byte stores)
        {
          uint32_t o[32];
          const int col0 = half ? C0 : 0;
          tmem_ld32(lane_base + A2_OCOL + col0, o);
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
              *reinterpret_cast<uint4*>(stg + ((((uint32_t)(col0 >> 3) + j) ^ swz) << 4)) = v;
            }
          }
          if (HD == 72 && half == 0) {  // columns [32, 48)
            uint32_t o2[16];
            ptx::tmem_ld16(lane_base + A2_OCOL + 32, o2);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              uint4 v;
              v.x = pack_bf16x2(__uint_as_float(o2[8 * j]) * inv, __uint_as_float(o2[8 * j + 1]) * inv);
              v.y = pack_bf16x2(__uint_as_float(o2[8 * j + 2]) * inv, __uint_as_float(o2[8 * j + 3]) * inv);
              v.z = pack_bf16x2(__uint_as_float(o2[8 * j + 4]) * inv, __uint_as_float(o2[8 * j + 5]) * inv);
              v.w = pack_bf16x2(__uint_as_float(o2[8 * j + 6]) * inv, __uint_as_float(o2[8 * j + 7]) * inv);
              *reinterpret_cast<uint4*>(stg + ((4 + j) << 4)) = v;
            }
          }
        }
        ptx::tc_fence_before();
        ptx::mbar_arrive(&t_empty[g]);  // TMEM region free for S of unit uq + 2
        if (lse && half == 0 && q0 + row < Tn)
          lse[((size_t)r * h + head) * Tn + q0 + row] = l > 0.f ? (msc + log2f(l)) * 0.6931471805599453f : 0.f;
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(bar_id, A2_GROUP);
        if (gt == 0) {
          a2_tma_store_3d(&tmOut, sQ + qb * BLOCK, head * HD, q0, r);
          ptx::tma_store_commit();
          pend_qb = qb;  // released at the start of this group's next unit (the store needs time to read the staging)
        }
      }
    }
    if (gt == 0) ptx::tma_store_wait_all();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tb, 512);
  }
}

template <int HD>
static int launch_attn_tc2(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, void* out, float* lse, cudaStream_t s) {
  using C = A2Cfg<HD>;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(attn_tc2_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) { set_last_error("attn_tc2: cannot set %d bytes of dynamic smem: %s", C::SMEM, cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    attr_done = true;
  }
  CUtensorMap tmQ, tmO;
  const int D = h * HD;
  if (at_make_tmap3(&tmQ, qkv, (long)R * Tn, 3L * D, C::CH) != 0) { set_last_error("attn_tc2: cuTensorMapEncodeTiled (qkv) failed"); return FITB200_ERR_CUDA; }
  {
    // out [R, Tn, D] bf16: box {HD columns, 128 tokens, 1 sample}; rows beyond the sample's Tn tokens are clipped by the store
    AtEncodeTiledFn enc = at_get_encode();
    if (!enc) { set_last_error("attn_tc2: cuTensorMapEncodeTiled not available from the driver"); return FITB200_ERR_CUDA; }
    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)Tn, (cuuint64_t)R};
    cuuint64_t strides[2] = {(cuuint64_t)D * 2, (cuuint64_t)Tn * D * 2};
    cuuint32_t box[3] = {(cuuint32_t)HD, 128, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult rc = enc(&tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      HD == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) { set_last_error("attn_tc2: cuTensorMapEncodeTiled (out) failed (%d)", (int)rc); return FITB200_ERR_CUDA; }
  }
  const int items = R * h;
  const int grid = items < sm_count() ? items : sm_count();
  attn_tc2_kernel<HD><<<grid, A2_THREADS, C::SMEM, s>>>(tmQ, tmO, kv_len, mask, R, Tn, h, (__nv_bfloat16*)out, lse);
  FITB200_CHECK_LAUNCH("attn_tc2");
  return FITB200_OK;
}

// bf16, T <= 256, q / k prerotated.  Returns FITB200_ERR_UNSUPPORTED for other head dims (caller falls back to attn_tc.cu).
int attn_fwd_tc2(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s) {
  if (d == 64) return launch_attn_tc2<64>(qkv, kv_len, mask, R, Tn, h, out, lse, s);
  if (d == 72) return launch_attn_tc2<72>(qkv, kv_len, mask, R, Tn, h, out, lse, s);
  return FITB200_ERR_UNSUPPORTED;
}

}  // namespace fitb200
