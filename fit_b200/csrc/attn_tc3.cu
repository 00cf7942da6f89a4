// KV-blocked bf16 attention forward on tcgen05 / TMEM for sequences LONGER than 256 tokens (resolution extrapolation: BASELINE
// config 5, 320x320 = 400 tokens ... 1024x1024 = 4096 tokens) whose q and k already carry the RoPE (qkv GEMM epilogue
// FITB200_EPI_BIAS_ROPE).  Same arithmetic as Attention.forward (fit.py:176-188; head merge fit.py:220-227) with the softmax
// evaluated block by block over the keys (running row max / row sum, O rescaled when the max moves), the schedule of attn_tc2.cu
// (persistent CTAs, event-driven MMA issuer, row-per-thread softmax groups) extended by a key loop:
//   warp 0      TMA producer: per PAIR of 128-query tiles of one (sample, head) item Q0 / Q1 once, then K_j / V_j blocks of 128 keys
//               through ONE 4-stage ring that both tiles multiply against (half the L2 -> SM traffic of a ring per tile; the first
//               blocks of the next pair are prefetched before its Q tiles, which wait for the previous epilogue)
//   warp 2      MMA issuer for both groups: S_j = Q K_j^T (TMEM region g, columns [0, 128)), then O += P_j V_j (columns [128, 128+DP))
//   warps 4-11  softmax / epilogue group 0 (first tile of the pair), warps 12-19 group 1 (second tile); two threads per query row
//               (TMEM lane quarter = warp & 3, key half = 64 of the block's 128 keys): block max (halves exchanged through shared
//               memory, one named barrier per block), P_j = exp2(S_j - max) as bf16 written over the first 32 of the thread's own 64
//               score columns (keys [0,64) -> columns [0,32), keys [64,128) -> columns [64,96)), partial row sums joined at the end,
//               O row rescaled in TMEM (columns split between the halves) when the reference max moved (skipped per warp when no
//               row's did), at the end O / rowsum -> staging (the unit's Q buffer) -> TMA stores of 32 rows (rows clipped at the
//               sample's token count by the 3-D map).
// The two groups interleave on the tensor pipe: while one runs its softmax the other's S / PV multiply.  tcgen05.mma executes in
// issue order, so S_{j+1} (which overwrites P_j) is simply issued behind P_j V_j; the only waits the issuer has are data (TMA) and
// p_ready.  Only keys < kv_len are multiplied (N / K of a block = round16 of its live keys); query tiles at or beyond kv_len are
// skipped (zero rows).
#include <cstring>

#include "common.cuh"
#include "ptx.cuh"
#include "umma_nosw.cuh"

namespace fitb200 {

constexpr int A3_BQ = 128;                        // queries per unit
constexpr int A3_BK = 128;                        // keys per block
constexpr int A3_GROUP = 256;                     // threads of one softmax group: 8 warps = 4 TMEM lane quarters x 2 key halves
constexpr int A3_THREADS = 128 + 2 * A3_GROUP;    // warps 0-1 producers, 2 MMA, 3 spare, 4-11 group 0, 12-19 group 1
constexpr int A3_OCOL = 128;                      // O accumulator columns inside a group's TMEM region
constexpr float A3_TAU = 8.f;                      // log2 of the growth of the row maximum that triggers a rescale of O
constexpr int A3_NS = 3;                          // K / V ring stages (one ring, read by both groups)
constexpr int A3_NQB = 2;                         // Q buffers per group (the next unit's Q lands while this one computes / stores)

__device__ __forceinline__ void a3_tma_store_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];\n" ::"l"(reinterpret_cast<uint64_t>(m)),
               "r"(ptx::smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

template <int HD>
struct A3Cfg {
  static constexpr int DP = (HD + 15) / 16 * 16;      // contraction width of q.k^T and N of p.v (72 -> 80)
  static constexpr int CHP = DP / 8, CH = HD / 8;
  static constexpr int CS = 128 * 16;                  // chunk-column stride inside a 128-row block
  static constexpr int BLOCK = CHP * CS;               // bytes of one 128-row operand block (incl. the contraction pad chunk)
  static constexpr int Q_OFF = 0;                      // [2 groups][A3_NQB]
  static constexpr int KV_OFF = 2 * A3_NQB * BLOCK;    // [A3_NS][K, V]
  static constexpr int BAR_OFF = KV_OFF + A3_NS * 2 * BLOCK;
  static constexpr int NBARS_G = 4 + 2 * A3_NQB;       // per group: s_full p_ready o_full t_empty | q_full q_empty [NQB]
  static constexpr int NBARS = 2 * NBARS_G + 4 * A3_NS;   // + ring: k_full v_full k_empty v_empty [NS]
  static constexpr int SCR_OFF = BAR_OFF + NBARS * 8 + 16;   // per group: block max [2 parities][2 halves][128] + row sum [2][128]
  static constexpr int SCR_BYTES = 6 * 128 * 4;
  static constexpr int SMEM = SCR_OFF + 2 * SCR_BYTES + 1024 /*alignment slack*/;
  static constexpr int OUT_ROW = HD * 2;               // staging row bytes (dense TMA box)
};

// A pair = two consecutive 128-query tiles of one (sample, head) item: tile 0 goes to softmax group 0, tile 1 to group 1, and both
// multiply against the SAME K_j / V_j blocks in shared memory (half the L2 -> SM traffic of one ring per group).
struct A3Unit { int r, head, q0, kvl, nkb; bool live0, live1; };

// Returns false for pairs whose first tile is padding (no tensor work, zero rows).
__device__ __forceinline__ bool a3_decode(int p, int npair, int h, int Tn, const int32_t* __restrict__ kv_len, A3Unit& o) {
  const int item = p / npair, pi = p - item * npair;
  o.r = item / h; o.head = item - o.r * h;
  o.kvl = min(kv_len[o.r], Tn);
  o.q0 = pi * 2 * A3_BQ;
  o.nkb = (o.kvl + A3_BK - 1) / A3_BK;
  o.live0 = o.q0 < o.kvl;
  o.live1 = o.q0 + A3_BQ < o.kvl;
  return o.live0;
}

template <int HD>
__global__ void __launch_bounds__(A3_THREADS, 1)
attn_tc3_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmOut, const int32_t* __restrict__ kv_len,
                const uint8_t* __restrict__ mask, int R, int Tn, int h, __nv_bfloat16* __restrict__ out, float* __restrict__ lse) {
  using C = A3Cfg<HD>;
  constexpr int DP = C::DP, CH = C::CH, CS = C::CS, BLOCK = C::BLOCK, NB = C::NBARS_G;
  extern __shared__ uint8_t a3_raw[];
  uint8_t* sm = a3_raw + ((1024u - (ptx::smem_u32(a3_raw) & 1023u)) & 1023u);
  uint8_t* sQ = sm + C::Q_OFF;
  uint8_t* sKV = sm + C::KV_OFF;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + C::BAR_OFF);
  auto s_full = [&](int g) { return bars + g * NB; };
  auto p_ready = [&](int g) { return bars + g * NB + 1; };
  auto o_full = [&](int g) { return bars + g * NB + 2; };
  auto t_empty = [&](int g) { return bars + g * NB + 3; };
  auto q_full = [&](int g, int b) { return bars + g * NB + 4 + b; };
  auto q_empty = [&](int g, int b) { return bars + g * NB + 4 + A3_NQB + b; };
  auto qbuf = [&](int g, int b) { return sQ + (g * A3_NQB + b) * BLOCK; };
  auto k_full = [&](int s) { return bars + 2 * NB + s; };
  auto v_full = [&](int s) { return bars + 2 * NB + A3_NS + s; };
  auto k_empty = [&](int s) { return bars + 2 * NB + 2 * A3_NS + s; };
  auto v_empty = [&](int s) { return bars + 2 * NB + 3 * A3_NS + s; };
  auto kbuf = [&](int s) { return sKV + (s * 2) * BLOCK; };
  auto vbuf = [&](int s) { return sKV + (s * 2 + 1) * BLOCK; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + C::NBARS);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int npair = (Tn + 2 * A3_BQ - 1) / (2 * A3_BQ);  // query-tile pairs per item
  const int NP = R * h * npair;
  const int D = h * HD;

  if (tid == 0) {
    ptx::prefetch_tmap(&tmQKV);
    ptx::prefetch_tmap(&tmOut);
    for (int g = 0; g < 2; ++g) {
      for (int b = 0; b < A3_NQB; ++b) { ptx::mbar_init(q_full(g, b), 1); ptx::mbar_init(q_empty(g, b), 1); }
      ptx::mbar_init(s_full(g), 1); ptx::mbar_init(p_ready(g), A3_GROUP); ptx::mbar_init(o_full(g), 1); ptx::mbar_init(t_empty(g), A3_GROUP);
    }
    for (int s = 0; s < A3_NS; ++s) {
      ptx::mbar_init(k_full(s), 1); ptx::mbar_init(v_full(s), 1); ptx::mbar_init(k_empty(s), 1); ptx::mbar_init(v_empty(s), 1);
    }
    ptx::fence_mbar_init();
  }
  if (DP != HD) {  // contraction padding (head_dim 72 -> 80) of every Q / K block: zeroed once, TMA boxes never touch it
    for (int i = tid; i < (2 * A3_NQB + A3_NS) * 128; i += A3_THREADS) {
      const int b = i >> 7, rr = i & 127;
      uint8_t* blk = b < 2 * A3_NQB ? sQ + b * BLOCK : kbuf(b - 2 * A3_NQB);
      *reinterpret_cast<uint4*>(blk + CH * CS + rr * 16) = make_uint4(0, 0, 0, 0);
    }
  }
  if (warp == 2) { ptx::tmem_alloc(tmem_slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = *tmem_slot;
  constexpr uint32_t BOX = CH * CS;  // bytes of one TMA box {8, 128, CH}

  if (warp == 0) {
    // ------------------------------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int uc0 = 0, uc1 = 0, bseq = 0;
      for (int p = blockIdx.x; p < NP; p += gridDim.x) {
        A3Unit pr;
        if (!a3_decode(p, npair, h, Tn, kv_len, pr)) continue;
        const int row0 = pr.r * Tn, cq = pr.head * CH, ck = (D + pr.head * HD) / 8, cv = (2 * D + pr.head * HD) / 8;
        auto load_kv = [&](int j) {
          const int slot = bseq % A3_NS; const uint32_t ph = (bseq / A3_NS) & 1;
          ptx::mbar_wait_sleep(k_empty(slot), ph ^ 1, 64);
          ptx::mbar_arrive_expect_tx(k_full(slot), BOX);
          tma_load_3d(kbuf(slot), &tmQKV, k_full(slot), 0, row0 + j * A3_BK, ck);
          ptx::mbar_wait_sleep(v_empty(slot), ph ^ 1, 64);
          ptx::mbar_arrive_expect_tx(v_full(slot), BOX);
          tma_load_3d(vbuf(slot), &tmQKV, v_full(slot), 0, row0 + j * A3_BK, cv);
          ++bseq;
        };
        // Q tiles first: their buffers were released two units ago (double buffered), so these waits do not hold up the ring
        ptx::mbar_wait_sleep(q_empty(0, uc0 % A3_NQB), ((uc0 / A3_NQB) & 1) ^ 1, 64);
        ptx::mbar_arrive_expect_tx(q_full(0, uc0 % A3_NQB), BOX);
        tma_load_3d(qbuf(0, uc0 % A3_NQB), &tmQKV, q_full(0, uc0 % A3_NQB), 0, row0 + pr.q0, cq);
        ++uc0;
        if (pr.live1) {
          ptx::mbar_wait_sleep(q_empty(1, uc1 % A3_NQB), ((uc1 / A3_NQB) & 1) ^ 1, 64);
          ptx::mbar_arrive_expect_tx(q_full(1, uc1 % A3_NQB), BOX);
          tma_load_3d(qbuf(1, uc1 % A3_NQB), &tmQKV, q_full(1, uc1 % A3_NQB), 0, row0 + pr.q0 + A3_BQ, cq);
          ++uc1;
        }
        for (int j = 0; j < pr.nkb; ++j) load_kv(j);
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // ------------------------------------------------------------------------------------------ MMA issuer (both groups)
    if (lane == 0) {
      // group state: current pair, block j, stage (0 = S_j pending, 1 = P_j V_j pending), ucnt = units done (q / o / t phases),
      // bc = blocks done by THIS group (s_full / p_ready phases), bseq = ring sequence number of its current block
      struct St { int p, nkb, kvl, need, j, stage, ucnt, bc, bseq; bool active; } st[2];
      auto next_unit = [&](St& s, int g, int from) {
        s.active = false;
        for (int p = from; p < NP; p += gridDim.x) {
          A3Unit pr;
          if (!a3_decode(p, npair, h, Tn, kv_len, pr)) continue;
          if (g == 1 && !pr.live1) { s.bseq += pr.nkb; continue; }   // group 0 alone walks this pair's blocks
          s.p = p; s.nkb = pr.nkb; s.kvl = pr.kvl; s.need = pr.live1 ? 2 : 1; s.j = 0; s.stage = 0; s.active = true;
          break;
        }
      };
      int cnt_s[A3_NS], cnt_pv[A3_NS];
#pragma unroll
      for (int i = 0; i < A3_NS; ++i) cnt_s[i] = cnt_pv[i] = 0;
      for (int g = 0; g < 2; ++g) { st[g].ucnt = 0; st[g].bc = 0; st[g].bseq = 0; next_unit(st[g], g, blockIdx.x); }
      uint32_t spins = 0;
      while (st[0].active || st[1].active) {
        bool progressed = false;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          St& s = st[g];
          if (!s.active) continue;
          const uint32_t tr = tb + g * 256;
          if (s.stage == 1) {
            const int slot = s.bseq % A3_NS; const uint32_t ph = (s.bseq / A3_NS) & 1;
            const int nk16 = (min(A3_BK, s.kvl - s.j * A3_BK) + 15) & ~15;
            // P_0 V_0 overwrites O: the previous unit's epilogue must have read it (t_empty)
            if ((s.j > 0 || ptx::mbar_try_wait(t_empty(g), (s.ucnt & 1) ^ 1)) && ptx::mbar_try_wait(p_ready(g), s.bc & 1) &&
                ptx::mbar_try_wait(v_full(slot), ph)) {
              ptx::tc_fence_after();
              const uint32_t idesc = ptx::idesc_bf16_f32(A3_BQ, DP) | (1u << 16);  // B (= V) is MN-major
              const int ksteps = nk16 >> 4;
              const uint8_t* v0 = vbuf(slot);
              for (int k = 0; k < ksteps; ++k) {
                const uint64_t db = desc_nosw(ptx::smem_u32(v0 + k * 256), 128, CS);
                // P of keys [0, 64) sits in columns [0, 32), P of keys [64, 128) in columns [64, 96) (each key half over its own scores)
                mma_f16_ts(tr + A3_OCOL, tr + (k < 4 ? 8 * k : 64 + 8 * (k - 4)), db, idesc, (s.j > 0 || k > 0) ? 1u : 0u);
              }
              if (++cnt_pv[slot] == s.need) { ptx::mma_commit(v_empty(slot)); cnt_pv[slot] = 0; }   // both groups' P_j V_j are out
              ++s.bc; ++s.bseq;
              if (++s.j == s.nkb) {
                ptx::mma_commit(o_full(g));
                ++s.ucnt;
                next_unit(s, g, s.p + gridDim.x);
              } else {
                s.stage = 0;
              }
              progressed = true;
            }
          }
          // S_{j+1} goes out right behind P_j V_j of the SAME group (not after the other group's turn): the group that finished its
          // softmax first gets its next scores first, which keeps the two groups in anti-phase -- one multiplies while the other
          // exponentiates.  Serving the groups strictly in turn interleaved PV(g1) PV(g0) S(g1) S(g0) and locked them in step.
          if (s.active && s.stage == 0) {
            const int slot = s.bseq % A3_NS; const uint32_t ph = (s.bseq / A3_NS) & 1;
            const int nk16 = (min(A3_BK, s.kvl - s.j * A3_BK) + 15) & ~15;
            const int qb = s.ucnt % A3_NQB;
            if ((s.j > 0 || ptx::mbar_try_wait(q_full(g, qb), (s.ucnt / A3_NQB) & 1)) && ptx::mbar_try_wait(k_full(slot), ph)) {
              ptx::tc_fence_after();
              const uint32_t idesc = ptx::idesc_bf16_f32(A3_BQ, nk16);
#pragma unroll
              for (int k = 0; k < DP / 16; ++k) {
                const uint64_t da = desc_nosw(ptx::smem_u32(qbuf(g, qb) + 2 * k * CS), CS, 128);
                const uint64_t db = desc_nosw(ptx::smem_u32(kbuf(slot) + 2 * k * CS), CS, 128);
                ptx::mma_f16_ss(tr, da, db, idesc, k != 0);
              }
              ptx::mma_commit(s_full(g));
              if (++cnt_s[slot] == s.need) { ptx::mma_commit(k_empty(slot)); cnt_s[slot] = 0; }   // both groups' S_j are out
              s.stage = 1;
              progressed = true;
            }
          }
        }
        if (progressed) spins = 0;
        else if (++spins > (1u << 25)) __trap();
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ------------------------------------------------------------------------------------------ softmax / epilogue groups
    const int g = (warp - 4) >> 3;                 // group 0: warps 4-11, group 1: warps 12-19
    const int gt = tid - 128 - g * A3_GROUP;       // thread inside the group
    const int gw = gt >> 5;                        // warp inside the group
    const int quarter = warp & 3, half = gw >> 2;  // TMEM lane quarter this warp may access; which 64 keys of a block / which O columns
    const int row = quarter * 32 + lane;
    float* smax = reinterpret_cast<float*>(sm + C::SCR_OFF + g * C::SCR_BYTES);  // [2][2][128]
    float* ssum = smax + 4 * A3_BQ;                                                // [2][128]
    const uint32_t lane_base = tb + g * 256 + ((uint32_t)(quarter * 32) << 16);
    const uint32_t my_base = lane_base + half * 64;  // this thread's 64 score columns; its P goes over the first 32 of them
    const float sl2 = rsqrtf((float)HD) * 1.4426950408889634f;  // softmax scale folded with log2(e)
    const int bar_id = 1 + g;
    int ucnt = 0, bcnt = 0;
    for (int p = blockIdx.x; p < NP; p += gridDim.x) {
      A3Unit un;
      a3_decode(p, npair, h, Tn, kv_len, un);
      const bool live = g ? un.live1 : un.live0;
      const int r = un.r, head = un.head, q0 = un.q0 + g * A3_BQ;
      if (!live) {  // padded query tile: zero output rows, no tensor work
        for (int i = gt; i < A3_BQ * CH; i += A3_GROUP) {
          const int rw = i / CH, c = i - rw * CH;
          if (q0 + rw < Tn) *reinterpret_cast<uint4*>(out + ((size_t)r * Tn + q0 + rw) * D + head * HD + c * 8) = make_uint4(0, 0, 0, 0);
        }
        if (lse) for (int i = gt; i < A3_BQ; i += A3_GROUP) if (q0 + i < Tn) lse[((size_t)r * h + head) * Tn + q0 + i] = 0.f;
        continue;
      }
      float m_run = -INFINITY, l_run = 0.f;   // l_run: this thread's half of the row sum (both halves use the same reference maximum)
      for (int j = 0; j < un.nkb; ++j, ++bcnt) {
        const int key0 = j * A3_BK + half * 64;
        const int nk = min(A3_BK, un.kvl - j * A3_BK), nk16 = (nk + 15) & ~15;
        const int n_my = max(0, min(nk16 - half * 64, 64));
        const int nch = (n_my + 31) >> 5;
        // key validity words (warp-uniform registers): bit i of vw[c] = key key0 + 32 c + i is < kv_len and not masked out
        uint32_t vw[2];
        {
          bool ok[2];
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const int key = key0 + c * 32 + lane;
            ok[c] = key < un.kvl && (mask == nullptr || mask[(size_t)r * Tn + key] != 0);
          }
#pragma unroll
          for (int c = 0; c < 2; ++c) vw[c] = __ballot_sync(0xffffffffu, ok[c]);
        }
        ptx::mbar_wait(s_full(g), bcnt & 1);
        ptx::tc_fence_after();
        float m = -INFINITY;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c < nch) {
            uint32_t sb[32];
            tmem_ld32(my_base + c * 32, sb);
            ptx::tmem_ld_wait();
            const uint32_t w = vw[c];
            if (w == 0xffffffffu) {
#pragma unroll
              for (int i = 0; i < 32; ++i) m = fmaxf(m, __uint_as_float(sb[i]));
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) m = (w >> i) & 1u ? fmaxf(m, __uint_as_float(sb[i])) : m;
            }
          }
        }
        float* sx = smax + (bcnt & 1) * 2 * A3_BQ;
        sx[half * A3_BQ + row] = m;
        ptx::named_bar_sync(bar_id, A3_GROUP);
        m = fmaxf(m_run, fmaxf(sx[row], sx[A3_BQ + row]));
        // Lazy rescaling: the reference maximum only follows the true one when that grew by more than 2^A3_TAU (exact in the end:
        // the normalisation cancels whatever reference was used; P <= 2^A3_TAU stays far inside bf16 / fp32 range).  With it O is
        // rescaled on the first blocks only; following every new maximum rescaled O on 82 % of the blocks (ncu, T = 4096).
        const bool moved = fmaf(m, sl2, -A3_TAU) > m_run * sl2 || m_run == -INFINITY;   // (-inf > -inf is false: handled by the second test)
        if (!moved) m = m_run;
        const float msc = (m == -INFINITY ? 0.f : m) * sl2;
        const float alpha = moved ? ex2f(fmaf(m_run, sl2, -msc)) : 1.f;  // m_run = -inf (first block / all keys so far masked) -> 0
        float sum = 0.f;
        // P (bf16 pairs) of chunk c goes to columns [16 c, 16 c + 16) of this thread's own score columns: already consumed by it
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c < nch) {
            uint32_t sb[32];
            tmem_ld32(my_base + c * 32, sb);
            ptx::tmem_ld_wait();
            const uint32_t w = vw[c];
            uint32_t pk[16];
            if (w == 0xffffffffu) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float p0 = ex2f(fmaf(__uint_as_float(sb[2 * i]), sl2, -msc));
                const float p1 = ex2f(fmaf(__uint_as_float(sb[2 * i + 1]), sl2, -msc));
                sum += p0 + p1;
                pk[i] = pack_bf16x2(p0, p1);
              }
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                float p0 = ex2f(fmaf(__uint_as_float(sb[2 * i]), sl2, -msc));
                float p1 = ex2f(fmaf(__uint_as_float(sb[2 * i + 1]), sl2, -msc));
                p0 = (w >> (2 * i)) & 1u ? p0 : 0.f;
                p1 = (w >> (2 * i + 1)) & 1u ? p1 : 0.f;
                sum += p0 + p1;
                pk[i] = pack_bf16x2(p0, p1);
              }
            }
            tmem_st16(my_base + c * 16, pk);
          }
        }
        l_run = fmaf(l_run, alpha, sum);
        m_run = m;
        // O row *= alpha (P_{j-1} V_{j-1} has retired: s_full of this block was committed behind it); half 0 takes columns [0, 48),
        // half 1 the rest.  Skipped when no row of the warp moved its reference maximum (the common case after the first blocks).
        if (j > 0 && __any_sync(0xffffffffu, moved)) {
#pragma unroll
          for (int cb = 0; cb < DP / 16; ++cb) {
            if ((cb < 3) == (half == 0)) {
              uint32_t o[16];
              ptx::tmem_ld16(lane_base + A3_OCOL + cb * 16, o);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tmem_st16(lane_base + A3_OCOL + cb * 16, o);
            }
          }
        }
        tmem_st_wait();
        ptx::tc_fence_before();
        ptx::mbar_arrive(p_ready(g));
      }

      // ---- epilogue: O / rowsum -> bf16 -> staging (the unit's Q buffer, consumed by every S_j) -> TMA stores (32 rows each)
      ssum[half * A3_BQ + row] = l_run;
      ptx::mbar_wait(o_full(g), ucnt & 1);
      ptx::tc_fence_after();
      ptx::named_bar_sync(bar_id, A3_GROUP);
      const float l = ssum[row] + ssum[A3_BQ + row];
      const float inv = l > 0.f ? 1.f / l : 0.f;
      constexpr int C0 = HD == 64 ? 32 : 48;  // columns [0, C0) -> half 0, [C0, HD) -> half 1
      uint8_t* stg = qbuf(g, ucnt % A3_NQB) + row * C::OUT_ROW;
      const uint32_t swz = HD == 64 ? (uint32_t)(row & 7) : 0u;  // 128-byte rows: 128B-swizzled box (conflict-free 16-byte stores)
      {
        uint32_t o[32];
        const int col0 = half ? C0 : 0;
        tmem_ld32(lane_base + A3_OCOL + col0, o);
        ptx::tmem_ld_wait();
        const int nvec = half ? (HD - C0) / 8 : 4;  // 16-byte chunks out of these 32 columns
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (q < nvec) {
            uint4 v;
            v.x = pack_bf16x2(__uint_as_float(o[8 * q]) * inv, __uint_as_float(o[8 * q + 1]) * inv);
            v.y = pack_bf16x2(__uint_as_float(o[8 * q + 2]) * inv, __uint_as_float(o[8 * q + 3]) * inv);
            v.z = pack_bf16x2(__uint_as_float(o[8 * q + 4]) * inv, __uint_as_float(o[8 * q + 5]) * inv);
            v.w = pack_bf16x2(__uint_as_float(o[8 * q + 6]) * inv, __uint_as_float(o[8 * q + 7]) * inv);
            *reinterpret_cast<uint4*>(stg + ((((uint32_t)(col0 >> 3) + q) ^ swz) << 4)) = v;
          }
        }
        if (HD == 72 && half == 0) {  // columns [32, 48)
          uint32_t o2[16];
          ptx::tmem_ld16(lane_base + A3_OCOL + 32, o2);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            uint4 v;
            v.x = pack_bf16x2(__uint_as_float(o2[8 * q]) * inv, __uint_as_float(o2[8 * q + 1]) * inv);
            v.y = pack_bf16x2(__uint_as_float(o2[8 * q + 2]) * inv, __uint_as_float(o2[8 * q + 3]) * inv);
            v.z = pack_bf16x2(__uint_as_float(o2[8 * q + 4]) * inv, __uint_as_float(o2[8 * q + 5]) * inv);
            v.w = pack_bf16x2(__uint_as_float(o2[8 * q + 6]) * inv, __uint_as_float(o2[8 * q + 7]) * inv);
            *reinterpret_cast<uint4*>(stg + ((4 + q) << 4)) = v;
          }
        }
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(t_empty(g));  // O free for P_0 V_0 of this group's next unit
      if (lse && half == 0 && q0 + row < Tn)
        lse[((size_t)r * h + head) * Tn + q0 + row] = l > 0.f ? (m_run * sl2 + log2f(l)) * 0.6931471805599453f : 0.f;
      ptx::fence_proxy_async_smem();
      ptx::named_bar_sync(bar_id, A3_GROUP);
      if (gt == 0) {
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) a3_tma_store_3d(&tmOut, qbuf(g, ucnt % A3_NQB) + qq * 32 * C::OUT_ROW, head * HD, q0 + qq * 32, r);
        ptx::tma_store_commit();
        ptx::tma_store_wait_read<0>();       // release the staging (= this unit's Q buffer) as soon as the stores have read it:
        ptx::mbar_arrive(q_empty(g, ucnt % A3_NQB));   // the Q tile of the unit after next lands while the next one computes
      }
      ++ucnt;
    }
    if (gt == 0) ptx::tma_store_wait_all();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tb, 512);
  }
}

template <int HD>
static int launch_attn_tc3(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, void* out, float* lse, cudaStream_t s) {
  using C = A3Cfg<HD>;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(attn_tc3_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) { set_last_error("attn_tc3: cannot set %d bytes of dynamic smem: %s", C::SMEM, cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    attr_done = true;
  }
  CUtensorMap tmQ, tmO;
  const int D = h * HD;
  if (at_make_tmap3(&tmQ, qkv, (long)R * Tn, 3L * D, C::CH) != 0) { set_last_error("attn_tc3: cuTensorMapEncodeTiled (qkv) failed"); return FITB200_ERR_CUDA; }
  {
    // out [R, Tn, D] bf16: box {HD columns, 32 tokens, 1 sample}; rows beyond the sample's Tn tokens are clipped by the store
    AtEncodeTiledFn enc = at_get_encode();
    if (!enc) { set_last_error("attn_tc3: cuTensorMapEncodeTiled not available from the driver"); return FITB200_ERR_CUDA; }
    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)Tn, (cuuint64_t)R};
    cuuint64_t strides[2] = {(cuuint64_t)D * 2, (cuuint64_t)Tn * D * 2};
    cuuint32_t box[3] = {(cuuint32_t)HD, 32, 1};  // one store per warp: its 32 query rows
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult rc = enc(&tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      HD == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) { set_last_error("attn_tc3: cuTensorMapEncodeTiled (out) failed (%d)", (int)rc); return FITB200_ERR_CUDA; }
  }
  const int npair = (Tn + 2 * A3_BQ - 1) / (2 * A3_BQ);
  const long pairs = (long)R * h * npair;
  const int grid = pairs < sm_count() ? (int)pairs : sm_count();
  attn_tc3_kernel<HD><<<grid, A3_THREADS, C::SMEM, s>>>(tmQ, tmO, kv_len, mask, R, Tn, h, (__nv_bfloat16*)out, lse);
  FITB200_CHECK_LAUNCH("attn_tc3");
  return FITB200_OK;
}

// bf16, any T, q / k prerotated.  Returns FITB200_ERR_UNSUPPORTED for other head dims (caller falls back to attn_mma.cu).
int attn_fwd_tc3(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s) {
  if (d == 64) return launch_attn_tc3<64>(qkv, kv_len, mask, R, Tn, h, out, lse, s);
  if (d == 72) return launch_attn_tc3<72>(qkv, kv_len, mask, R, Tn, h, out, lse, s);
  return FITB200_ERR_UNSUPPORTED;
}

}  // namespace fitb200
