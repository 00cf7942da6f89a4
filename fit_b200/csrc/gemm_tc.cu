// bf16 GEMM on the 5th-generation tensor cores:  out[M,N] = epilogue(A[M,K] . B[N,K]^T)
//
//   * persistent: one CTA per SM walks output tiles (128 x BN, BN in {128,192,256}), n fastest so the
//     CTAs of a wave share A row-blocks in L2 while the (small) weight stays L2 resident
//   * warp 0 : TMA producer   -- cp.async.bulk.tensor 128B-swizzled A/B k-slabs into a STAGES-deep ring
//   * warp 1 : MMA issuer     -- one thread issues tcgen05.mma (M=128, N=BN, K=16) into TMEM,
//                                tcgen05.commit releases smem slots / publishes the accumulator
//   * warps 2-9 : epilogue    -- (two warps per TMEM lane quarter, half a slab each) per 128-byte-wide output slab: tcgen05.ld TMEM -> registers, fused
//                                bias / SiLU / SwiGLU / adaLN-gate + residual, 16-byte swizzled st.shared,
//                                TMA store.  The residual slab is TMA-LOADED two slabs ahead, so the
//                                epilogue issues no per-row global access at all (round-1 profile: the
//                                row-per-thread global loads made the epilogue the bottleneck, long_scoreboard).
//     Two TMEM accumulator stages: the epilogue of tile i overlaps the mainloop of tile i+1.
// Every mbarrier wait is bounded (ptx::mbar_wait traps) so a pipeline bug cannot hang the device.
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace fitb200 {

constexpr int GT_BM = 128, GT_BK = 64;
constexpr int GT_THREADS = 320;  // producer + mma + 8 epilogue warps
constexpr int GT_EPI_THREADS = 256;  // two warps per TMEM lane quarter: each takes half of a slab's columns
// LNA kernels (LayerNorm + adaLN modulate prologue): 16 warps in four warpgroups, so that setmaxnreg can move registers between the roles --
//   warps 0..3   TMA producer (raw x rows + B), MMA issuer, relay thread (publishes the rescaled A rows to the tensor core), one idle warp
//   warps 4..11  epilogue
//   warps 12..15 A rescalers: x * bf16(1 + scale) in place in the shared-memory ring slot TMA filled
constexpr int GT_LNA_THREADS = 128;
constexpr int GT_THREADS_LNA = 512;
constexpr int GT_LNA_ROWS = 128 / (GT_LNA_THREADS / 8);   // rows per rescaler thread
constexpr int GT_BAR_BYTES = 512;
constexpr int GT_SLAB_BYTES = GT_BM * 128;  // 128 rows x 128 bytes
#ifndef FITB200_ROPE_SMEM_HD
#define FITB200_ROPE_SMEM_HD 72
#endif
constexpr int GT_ROPE_MAX_HD = FITB200_ROPE_SMEM_HD;  // head dims up to this keep the tile's RoPE rows in shared memory (FiT: 64, 72)
constexpr int GT_SMEM_BUDGET = 232448 - 1024 /*align slack*/ - GT_BAR_BYTES /*barriers*/;

// PAIR: two CTAs of a cluster (one TPC) share one 256 x BN tile through tcgen05 cta_group::2 -- each CTA stages its own 128 rows of A
// and HALF of the B rows, so the shared-memory bytes read per MMA flop drop by a third (a 128 x 192 single-CTA tile needs 106 of the
// SM's 128 B/clk; the pair needs 75) and each CTA's ring holds more stages.
template <int BN, int EPI, bool PAIR = false, bool LNA = false> struct GtCfg {
  static constexpr int A_BYTES = GT_BM * GT_BK * 2;
  static constexpr int BNL = PAIR ? BN / 2 : BN;  // B rows staged by this CTA
  static constexpr int B_BYTES = BNL * GT_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // out x2 (+ resid x2) (+ the RoPE table rows of the tile's 128 tokens, fp32, row stride head_dim + 4 words: conflict-free float4 reads)
  static constexpr int ROPE_BYTES = EPI == FITB200_EPI_BIAS_ROPE ? GT_BM * (GT_ROPE_MAX_HD + 4) * 4 : 0;
  // LNA: the (c1, c2) vectors of the tile's sample and columns, one copy per epilogue lane quarter: [4][2][BN] fp32
  static constexpr int LNC_BYTES = LNA ? 4 * 2 * BN * 4 : 0;
  static constexpr int EPI_BYTES = (EPI == FITB200_EPI_BIAS_GATE_RESID ? 4 : 2) * GT_SLAB_BYTES + ROPE_BYTES + LNC_BYTES;
  static constexpr int STAGES_RAW = (GT_SMEM_BUDGET - EPI_BYTES) / STAGE_BYTES;
#ifdef FITB200_MAX_STAGES   // tuning builds: cap the ring depth (how much of a kernel's rate is its stage count?)
  static constexpr int STAGES_CAP = FITB200_MAX_STAGES;
#else
  static constexpr int STAGES_CAP = PAIR ? 8 : 6;
#endif
  static constexpr int STAGES = STAGES_RAW > STAGES_CAP ? STAGES_CAP : STAGES_RAW;
  static constexpr int TMEM_COLS = BN == 128 ? 256 : 512;  // 2 accumulator stages, power of two
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + GT_BAR_BYTES;
  static_assert(STAGES >= 3, "pipeline too shallow");
  static_assert(!PAIR || (BNL % 8 == 0 && BN % 16 == 0), "pair tile: N a multiple of 16");
};

struct GtParams {
  int M, N, K;
  const float* bias;
  const float* gate; long ld_gate; int rows_per_sample;
  const float* rope_freqs; int rope_rows, rope_tokens, rope_hd, rope_cols;
  int tail_rows;  // K-major B: rows of the narrower TMA box (tmBt) used for the last, cut column tile (0 = always the full box)
  int bias_row0;  // the bias is added to rows >= bias_row0 only (0 = every row)
  int epi_sleep;  // nanoseconds between the epilogue's polls of the accumulator-full barrier (0 = spin)
  // split-K (wgrad GEMMs with few output tiles): work item = (tile, k-range); the partial tiles meet in TMA reduce-add stores on a zeroed output
  int splits, kb_per_split;
  // LNA (LayerNorm + adaLN-modulate prologue, see the A-producer warps below)
  const __nv_bfloat16* lnx; long ld_lnx;   // raw residual-stream rows x [M, K]
  const __nv_bfloat16* lng; long ld_lng;   // bf16(1 + scale) [samples, K]
  const float* lnc; int ln_samples;        // [2 * samples, N]: c1[s, n] = sum_k bf16(1+scale[s,k]) W[n,k], then c2[s, n] = bias[n] + sum_k bf16(shift[s,k]) W[n,k]
  const float2* ln_parts; int ln_nparts;   // [M, nparts] partial (sum x, sum x^2) of every row of x
  float ln_eps;
  // GATE_RESID: partial row sums of the OUTPUT rows (the next LayerNorm's statistics), [M, 2 * column tiles] (nullable)
  float2* stats_out; int stats_nparts;
  int dbg;  // FITB200_GEMM_DBG=1: epilogue releases the accumulator without reading it (mainloop ceiling measurement only)
};

__device__ __forceinline__ void tmem_ld16f(uint32_t taddr, float* v) {
  uint32_t r[16];
  ptx::tmem_ld16(taddr, r);
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// OutT = bf16: a slab is 64 output columns; OutT = float: 32 output columns (both 128 bytes per row).
// AMN / BMN: the operand is stored contraction-major ([K, M] resp. [K, N] row-major: the dgrad / wgrad GEMMs of the backward pass);
// its k-slab is fetched as 64-column TMA boxes and read through MN-major SWIZZLE_128B descriptors (tools/probe/umma_probe_mn.cu).
//
// LNA: the A operand is modulate(LayerNorm(x), shift, scale) (fit.py:45-46,303,308,332,336) of the raw residual stream x, formed inside this kernel
// instead of by a separate pass over HBM.  With g = 1 + scale[s], mu / rstd the row statistics:
//     sum_k ((x - mu) rstd g_k + shift_k) W[n,k]  =  rstd (sum_k (x g_k) W[n,k]  -  mu c1[s,n])  +  c2[s,n]
// so the only per-element work on the operand path is ONE packed bf16 multiply x * bf16(g_k): TMA delivers the raw x rows into the A ring as
// for any GEMM, four extra warps rescale them in place (ld.shared / HMUL2 / st.shared on the 128B-swizzled slot) before the MMA reads them;
// mu, rstd, c1, c2 are applied to the fp32 accumulators in the epilogue.  (First version: the warps loaded x themselves with 16-byte global
// loads -- the same bytes through the LSU / L1 instead of TMA: 65 % tensor pipe against 80 %, profiles/r2b_ln_prologue.txt.)  c1 / c2 are [samples, N] vectors
// from a small GEMM of the modulation rows with the same weight; the row sums come from the GATE_RESID epilogue that wrote x.
template <int BN, int EPI, typename OutT, bool AMN = false, bool BMN = false, bool PAIR = false, bool LNA = false>
__global__ void __launch_bounds__(LNA ? GT_THREADS_LNA : GT_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmBt,
               const GtParams p) {
  using Cfg = GtCfg<BN, EPI, PAIR, LNA>;
  static_assert(!PAIR || !BMN || Cfg::BNL % 64 == 0, "pair tiles with a contraction-major B: whole 64-column boxes per CTA");
  constexpr int STAGES = Cfg::STAGES;
  constexpr int TM = PAIR ? 2 * GT_BM : GT_BM;  // rows of one scheduled tile (both CTAs of a pair)
  constexpr int NGRAN = PAIR ? 32 : 16;         // granularity of the last column tile's MMA N
  constexpr bool SWIGLU = EPI == FITB200_EPI_SWIGLU;
  constexpr bool RESID = EPI == FITB200_EPI_BIAS_GATE_RESID;
  constexpr bool ROPE = EPI == FITB200_EPI_BIAS_ROPE;
  constexpr int OUTC = 128 / (int)sizeof(OutT);       // output columns per slab
  constexpr int ACCC = SWIGLU ? 2 * OUTC : OUTC;      // accumulator columns per slab
  constexpr int NSLAB = BN / ACCC;
  constexpr int CW = 16 / (int)sizeof(OutT);          // output columns per 16-byte chunk: 8 (bf16) or 4 (fp32)
  static_assert(BN % ACCC == 0, "tile width must be a whole number of slabs");
  static_assert(!LNA || (PAIR && !AMN && !BMN && (ROPE || SWIGLU || EPI == FITB200_EPI_BIAS) && (sizeof(OutT) == 2 || EPI == FITB200_EPI_BIAS)),
                "LNA: K-major pair tiles; bf16 output (fp32 with the plain BIAS epilogue: the final layer)");
  constexpr int LNA_WARPS = GT_LNA_THREADS / 32;
  constexpr int EPI_WARP0 = LNA ? 4 : 2;          // first epilogue warp (a multiple of 4 apart from the TMEM lane quarter rule: quarter = warp % 4)

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = ptx::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);  // SWIZZLE_128B needs 1024-byte alignment
  uint8_t* smemA = smem;
  uint8_t* smemB = smem + STAGES * Cfg::A_BYTES;
  uint8_t* smemOut = smem + STAGES * Cfg::STAGE_BYTES;          // [2][128 rows][128 B]
  uint8_t* smemRes = smemOut + 2 * GT_SLAB_BYTES;               // [2][128 rows][128 B]  (RESID only)
  float* smemRope = reinterpret_cast<float*>(smemOut + 2 * GT_SLAB_BYTES);  // [128 rows][head_dim + 4]  (ROPE only)
  float* smemLnc = reinterpret_cast<float*>(smem + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES - Cfg::LNC_BYTES);  // [4][2][BN]  (LNA only)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
  uint64_t* full_bar = bars;                     // [STAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + STAGES;           // [STAGES]  MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * STAGES;       // [2]       MMA -> epilogue
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]       epilogue -> MMA
  uint64_t* res_bar = bars + 2 * STAGES + 4;     // [4][2]    residual slab rows of lane quarter q landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 12);
  uint64_t* abar = bars + 2 * STAGES + 13;       // [STAGES]  LNA: rescaler warps -> relay thread
  uint64_t* rbar = bars + 3 * STAGES + 13;       // [STAGES]  LNA: TMA (raw x rows of THIS CTA landed) -> rescaler warps
  static_assert(4 * STAGES + 13 <= GT_BAR_BYTES / 8, "barrier area");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = ceil_div(p.N, BN), tiles_m = ceil_div(p.M, TM);
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = ceil_div(p.K, GT_BK);
  const int nsplit = LNA ? 1 : p.splits;                       // host: every split is non-empty; 1 for everything but the wgrad GEMMs
  const int num_items = num_tiles * nsplit;                    // item = split * num_tiles + tile
  const int kbs = nsplit > 1 ? p.kb_per_split : num_kb;
  const uint32_t rank = PAIR ? ptx::cluster_ctarank() : 0u;   // 0 = leader: issues the MMAs of the pair
  const int wid = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;      // scheduling unit (CTA or pair)
  const int nwork = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int row_off = (int)rank * GT_BM;                                   // this CTA's rows inside the tile

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    ptx::prefetch_tmap(&tmOut);
    if (RESID) ptx::prefetch_tmap(&tmRes);
    if (p.tail_rows > 0) ptx::prefetch_tmap(&tmBt);
    // LNA: a slot is full once the B bytes have landed AND the relay thread of each CTA of the pair has published the A rows of its producers
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], LNA ? 1 + (PAIR ? 2 : 1) : 1);
      ptx::mbar_init(&empty_bar[s], 1);
      if (LNA) { ptx::mbar_init(&abar[s], LNA_WARPS); ptx::mbar_init(&rbar[s], 1); }
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1);
      ptx::mbar_init(&tempty_bar[s], PAIR ? 2 * (GT_EPI_THREADS / 32) : GT_EPI_THREADS);  // pair: one arrival per epilogue warp of both CTAs
      for (int qd = 0; qd < 4; ++qd) ptx::mbar_init(&res_bar[qd * 2 + s], 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {  // one warp (of each CTA) owns TMEM allocation and deallocation
    if (PAIR) { ptx::tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS); ptx::tmem_relinquish_pair(); }
    else { ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync();  // the peer's barriers must be initialised before anything is signalled across the pair
  else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // LNA: registers follow the work -- every warp of a warpgroup executes the same setmaxnreg, first thing in its role branch
  if (warp == 0) {
    if (LNA) ptx::setmaxnreg_dec<40>();
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int item = wid; item < num_items; item += nwork) {
        const int tile = item % num_tiles, kb0 = (item / num_tiles) * kbs, kb1 = min(num_kb, kb0 + kbs);
        // the last column tile may be narrower (MMA N = nn): in a pair the second CTA's half of B then starts nn / 2 rows in
        const int n0t = (tile % tiles_n) * BN, nn = min(BN, (p.N - n0t + NGRAN - 1) / NGRAN * NGRAN);
        const int m0 = (tile / tiles_n) * TM + row_off, n0 = n0t + (int)rank * (nn / 2);
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          if (PAIR) {
            // both CTAs' bytes are counted on the LEADER's full barrier (the only one the MMA thread waits on)
            // the cut last column tile fetches only the B rows its MMAs read: a full box would make that tile as expensive as a whole one
            const bool tail = !BMN && nn < BN && p.tail_rows > 0;
            const uint32_t stage_bytes = (LNA ? 0 : Cfg::A_BYTES) + (tail ? p.tail_rows * 128 : Cfg::B_BYTES);
            if (rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * stage_bytes);
            const uint32_t fb = ptx::mapa(ptx::smem_u32(&full_bar[stage]), 0);
            if (LNA) {
              // raw x rows of this CTA -> its own ring slot, announced to its own rescaler warps (not to the MMA's full barrier)
              ptx::mbar_arrive_expect_tx(&rbar[stage], Cfg::A_BYTES);
              ptx::tma_load_2d(smemA + stage * Cfg::A_BYTES, &tmA, &rbar[stage], kb * GT_BK, m0);
            } else if (AMN) {
#pragma unroll
              for (int b = 0; b < GT_BM / 64; ++b) ptx::tma_load_2d_pair(smemA + stage * Cfg::A_BYTES + b * 8192, &tmA, fb, m0 + b * 64, kb * GT_BK);
            } else {
              ptx::tma_load_2d_pair(smemA + stage * Cfg::A_BYTES, &tmA, fb, kb * GT_BK, m0);
            }
            if (BMN) {
#pragma unroll
              for (int b = 0; b < Cfg::BNL / 64; ++b) ptx::tma_load_2d_pair(smemB + stage * Cfg::B_BYTES + b * 8192, &tmB, fb, n0 + b * 64, kb * GT_BK);
            } else {
              ptx::tma_load_2d_pair(smemB + stage * Cfg::B_BYTES, tail ? &tmBt : &tmB, fb, kb * GT_BK, n0);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
            continue;
          }
          const bool tail1 = !BMN && nn < BN && p.tail_rows > 0;
          ptx::mbar_arrive_expect_tx(&full_bar[stage], tail1 ? Cfg::A_BYTES + p.tail_rows * 128 : Cfg::STAGE_BYTES);
          if (AMN) {
#pragma unroll
            for (int b = 0; b < GT_BM / 64; ++b) ptx::tma_load_2d(smemA + stage * Cfg::A_BYTES + b * 8192, &tmA, &full_bar[stage], m0 + b * 64, kb * GT_BK);
          } else {
            ptx::tma_load_2d(smemA + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], kb * GT_BK, m0);
          }
          if (BMN) {
#pragma unroll
            for (int b = 0; b < BN / 64; ++b) ptx::tma_load_2d(smemB + stage * Cfg::B_BYTES + b * 8192, &tmB, &full_bar[stage], n0 + b * 64, kb * GT_BK);
          } else {
            ptx::tma_load_2d(smemB + stage * Cfg::B_BYTES, tail1 ? &tmBt : &tmB, &full_bar[stage], kb * GT_BK, n0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (LNA) ptx::setmaxnreg_dec<40>();
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc0 = ptx::idesc_bf16_f32(TM, 0) | (AMN ? 1u << 15 : 0u) | (BMN ? 1u << 16 : 0u);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int item = wid; item < num_items; item += nwork, ++it) {
        const int tile = item % num_tiles, kb0 = (item / num_tiles) * kbs, kb1 = min(num_kb, kb0 + kbs);
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        // epilogue (of both CTAs of a pair) has drained this accumulator stage
        // (plain wait also for the pair: the peer's arrivals order no memory -- the hand-off is "TMEM drained", fenced by tcgen05 -- and a
        // cluster-scope acquire compiles to TRYWAIT + CCTL.IVALL, an L1 invalidate per tile under the epilogue's bias / gate / table reads)
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        // tiles are 256 wide and only the last one of a row is cut to what is left (192- and 128-wide tiles measured 0.73x / 0.50x the
        // throughput of 256-wide ones: the MMA itself runs at full rate for N >= 128 (tools/probe/umma_rate.cu), the operand bytes staged
        // per flop are what grows)
        const int nn = min(BN, (p.N - (tile % tiles_n) * BN + NGRAN - 1) / NGRAN * NGRAN);
        const uint32_t idesc = idesc0 | ((uint32_t)(nn >> 3) << 17);
        for (int kb = kb0; kb < kb1; ++kb) {
          // (LNA: the A rows of both CTAs were published by their relay threads behind a generic -> async proxy fence.  NOT a cluster-scope
          // acquire here: it compiles to CCTL.IVALL -- an L1 invalidate per k-slab, which turned every spill reload and table read of the
          // other warps into an L2 round trip: the fused GEMMs ran 2.5x slower)
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t a_addr = ptx::smem_u32(smemA + stage * Cfg::A_BYTES), b_addr = ptx::smem_u32(smemB + stage * Cfg::B_BYTES);
          const uint64_t da = AMN ? ptx::smem_desc_mnmajor_sw128(a_addr) : ptx::smem_desc_kmajor_sw128(a_addr);
          const uint64_t db = BMN ? ptx::smem_desc_mnmajor_sw128(b_addr) : ptx::smem_desc_kmajor_sw128(b_addr);
#pragma unroll
          for (int k = 0; k < GT_BK / 16; ++k) {
            // 16 elements along K: K-major = 32 bytes inside the 128-byte swizzle row (+2 in the addr >> 4 field);
            // MN-major = 16 k-rows of 128 bytes (+128)
            if (PAIR) ptx::mma_f16_ss_pair(d_tmem, da + (AMN ? 128 : 2) * k, db + (BMN ? 128 : 2) * k, idesc, ((kb - kb0) | k) != 0);
            else ptx::mma_f16_ss(d_tmem, da + (AMN ? 128 : 2) * k, db + (BMN ? 128 : 2) * k, idesc, ((kb - kb0) | k) != 0);
          }
          // smem slot reusable once these MMAs retire (pair: the slot of BOTH CTAs)
          if (PAIR) ptx::mma_commit_pair(&empty_bar[stage], 3); else ptx::mma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (PAIR) ptx::mma_commit_pair(&tfull_bar[as], 3); else ptx::mma_commit(&tfull_bar[as]);  // accumulator complete
      }
    }
    __syncwarp();
  } else if (LNA && warp == 2) {
    ptx::setmaxnreg_dec<40>();
    // ------------------------------------------------------------------ relay thread (LNA): A rows stored -> proxy fence -> MMA's full barrier
    if (lane == 0) {
      const uint32_t full_leader = PAIR ? ptx::mapa(ptx::smem_u32(&full_bar[0]), 0) : ptx::smem_u32(&full_bar[0]);
      const int my_tiles = wid < num_tiles ? (num_tiles - 1 - wid) / nwork + 1 : 0;
      const int total = my_tiles * num_kb;
      int stage = 0; uint32_t phase = 0;
      for (int i = 0; i < total; ++i) {
        ptx::mbar_wait(&abar[stage], phase);    // acquire: the four producer warps' st.shared of this slot
        ptx::fence_proxy_async_smem();          // ... become visible to the async proxy
        if (PAIR) ptx::mbar_arrive_remote(full_leader + stage * 8); else ptx::mbar_arrive(&full_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (LNA && warp == 3) {
    ptx::setmaxnreg_dec<40>();  // idle: fills the first warpgroup
  } else if (LNA && warp >= EPI_WARP0 + GT_EPI_THREADS / 32) {
    // ------------------------------------------------------------------ A-producer warps (LNA): x * bf16(1 + scale) -> swizzled A ring
    // thread = (16-byte chunk c of the 128-byte k-slab row, row group rg): rows rg, rg + 16, ... rg + 112 of this CTA's 128 rows.  One warp
    // instruction reads 4 rows x 128 contiguous bytes and stores 512 contiguous (chunk-permuted) bytes: coalesced and conflict-free.
    ptx::setmaxnreg_dec<72>();
    constexpr int NR = GT_LNA_ROWS, RSTEP = GT_LNA_THREADS / 8;   // rows per thread, row distance between them
    const int pt = (int)threadIdx.x - (EPI_WARP0 * 32 + GT_EPI_THREADS);
    const int c = pt & 7, rg = pt >> 3;
    // thread = (16-byte chunk c of the 128-byte k-slab row, row group rg): rows rg, rg + 16, ... of this CTA's 128 rows; (rg + 16 j) & 7 ==
    // rg & 7, so the 128B-swizzle phase is per thread.  One warp instruction touches 4 rows x 128 contiguous (chunk-permuted) bytes:
    // conflict-free.  32-bit shared-window addresses throughout.
    const uint32_t dst0 = ptx::smem_u32(smemA) + (uint32_t)(rg * 128 + ((c ^ (rg & 7)) << 4));
    const uint32_t rbar0 = ptx::smem_u32(&rbar[0]), abar0 = ptx::smem_u32(&abar[0]);
    const int my_tiles = wid < num_tiles ? (num_tiles - 1 - wid) / nwork + 1 : 0;
    const int total = my_tiles * num_kb;
    // bf16(1 + scale) of the slab's 8 columns of this thread.  rows_per_sample % 128 == 0 (checked on the host): the 128 rows of a CTA tile
    // belong to one sample, so one 16-byte load per thread and k-slab serves all its rows; each chunk is read once per CTA and tile (an L2
    // hit, never an L1 hit), so it is fetched three k-slabs ahead.  K % 64 == 0 (checked on the host): every k-slab is whole.
    int f_it = 0, f_kb = 0;
    const char* gbase = nullptr;
    auto enter_tile = [&]() {
      const int m0 = ((wid + f_it * nwork) / tiles_n) * TM + row_off;
      gbase = reinterpret_cast<const char*>(p.lng + (size_t)(min(m0, p.M - 1) / p.rows_per_sample) * p.ld_lng + c * 8);
    };
    if (total > 0) enter_tile();
    auto fetch = [&](uint4& g) {
      g = __ldg(reinterpret_cast<const uint4*>(gbase + (uint32_t)f_kb * (GT_BK * 2)));
      if (++f_kb == num_kb) { f_kb = 0; ++f_it; if (f_it < my_tiles) enter_tile(); }
    };
    int stage = 0; uint32_t phase = 0;
    auto rescale = [&](const uint4& g) {
      ptx::mbar_wait_addr(rbar0 + stage * 8, phase);   // TMA has written this CTA's raw rows of the slot
      const uint32_t dst = dst0 + stage * Cfg::A_BYTES;
      const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&g);
      uint4 v[NR];
#pragma unroll
      for (int j = 0; j < NR; ++j) v[j] = ptx::ld_shared_v4(dst + j * RSTEP * 128);
#pragma unroll
      for (int j = 0; j < NR; ++j) {
        __nv_bfloat162* v2 = reinterpret_cast<__nv_bfloat162*>(&v[j]);
#pragma unroll
        for (int e = 0; e < 4; ++e) v2[e] = __hmul2(v2[e], g2[e]);
        ptx::st_shared_v4(dst + j * RSTEP * 128, v[j]);
      }
      // hand the rows to the relay thread (CTA-scope arrive); it executes the generic -> async proxy fence the tensor core's operand reads
      // need and arrives on the MMA's full barrier
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_addr(abar0 + stage * 8);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    };
    const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
    uint4 g0 = z4, g1 = z4, g2 = z4;
    if (total > 0) fetch(g0);
    if (total > 1) fetch(g1);
    if (total > 2) fetch(g2);
    for (int i = 0; i < total; i += 3) {
      rescale(g0);
      if (i + 3 < total) fetch(g0);
      if (i + 1 < total) {
        rescale(g1);
        if (i + 4 < total) fetch(g1);
      }
      if (i + 2 < total) {
        rescale(g2);
        if (i + 5 < total) fetch(g2);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (2..9)
    if (LNA) ptx::setmaxnreg_inc<168>();
    const int quarter = warp & 3;            // TMEM lane quarter this warp may read
    const int half = (warp - EPI_WARP0) >> 2;  // which half of each slab's columns this warp handles
    const int row = quarter * 32 + lane;     // row inside the tile == TMEM lane
    // The epilogue is four independent pipelines, one per TMEM lane quarter (two warps = 64 threads, 32 tile rows): each stages, stores
    // and prefetches the residual of ITS rows (32-row TMA boxes) behind its own named barrier, so a slow quarter never stalls the others
    // (one 128-row store per slab behind two CTA-wide barriers kept all eight warps in lock step).
    const bool leader = half == 0 && lane == 0;  // issues this quarter's TMA stores / residual loads
    const int qbar = 1 + quarter;                // named barrier of this quarter's two warps
    constexpr int QROWS = 32, QBYTES = QROWS * 128;
    const uint32_t sw = (uint32_t)(row & 7); // 128B-swizzle phase of this row
    const int my_tiles = wid < num_items ? (num_items - 1 - wid) / nwork + 1 : 0;   // work items of this CTA (pair)
    const uint32_t tempty_leader = PAIR ? ptx::mapa(ptx::smem_u32(&tempty_bar[0]), 0) : 0u;  // the leader's tempty_bar[0]
    const int total_slabs = my_tiles * NSLAB;

    // residual slab q (running index over this CTA's tiles) -> TMA load into resid buffer q & 1
    auto issue_resid = [&](int qq) {
      if (qq >= total_slabs) return;
#ifdef FITB200_TUNING
      if (p.dbg & 4) return;
#endif
      const int tile = (wid + (qq / NSLAB) * nwork) % num_tiles, slab = qq % NSLAB;
      const int rm0 = (tile / tiles_n) * TM + row_off, rn0 = (tile % tiles_n) * BN + slab * OUTC;
      ptx::mbar_arrive_expect_tx(&res_bar[quarter * 2 + (qq & 1)], QBYTES);
      ptx::tma_load_2d(smemRes + (qq & 1) * GT_SLAB_BYTES + quarter * QBYTES, &tmRes, &res_bar[quarter * 2 + (qq & 1)], rn0, rm0 + quarter * QROWS);
    };
    if (RESID && leader) { issue_resid(0); issue_resid(1); }

    int q = 0;  // running slab index
    int rope_m0 = -1;  // key of the row block whose RoPE rows are in smemRope (token offset for a shared table, else the row block itself)
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = (wid + it * nwork) % num_tiles;
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      const int m0 = (tile / tiles_n) * TM + row_off, n0 = (tile % tiles_n) * BN;
      const int m = min(m0 + row, p.M - 1);
      const float* gate_row = nullptr;
      if (RESID) {
        gate_row = p.gate + (size_t)(m / p.rows_per_sample) * p.ld_gate;
        // this tile's segment of the gate row -> L1 now, while the accumulator is still being formed: it changes with the sample (= with every
        // tile at 256 tokens per sample), so each 16-byte read at its point of use below was a cold L2 round trip
        if (half == 0 && lane < BN / 32 && n0 + lane * 32 < p.N) ptx::prefetch_l1(gate_row + n0 + lane * 32);
      }
      float st1 = 0.f, st2 = 0.f;  // RESID: sum / sum of squares of this row's output columns handled by this thread (stats_out)
      // LNA: value = ln_a * acc + (ln_b * c1[s, n] + c2[s, n]) with ln_a = rstd, ln_b = -mean * rstd of this row of x
      float ln_a = 1.f, ln_b = 0.f;
      const float* c1row = nullptr;
      const float* c2row = nullptr;
      if (LNA) {
        float s1 = 0.f, s2 = 0.f;
        const float2* pp = p.ln_parts + (size_t)m * p.ln_nparts;
        for (int i = 0; i < p.ln_nparts; ++i) { const float2 t = __ldg(pp + i); s1 += t.x; s2 += t.y; }
        const float mu = s1 / (float)p.K;
        const float var = fmaxf(s2 / (float)p.K - mu * mu, 0.f);
        ln_a = 1.0f / sqrtf(var + p.ln_eps);
        ln_b = -mu * ln_a;
        // c1 / c2 of this tile's sample (one per CTA tile: rows_per_sample % 128 == 0) and columns -> this quarter's shared-memory copy, loaded
        // BEFORE the accumulator wait: they differ from tile to tile, so reading them from global memory at their point of use was an L2 round
        // trip per 16-byte chunk (the fused GEMMs ran 2.5x slower, epilogue-bound).  The quarter's threads are past the last named barrier of
        // the previous tile (all its reads of the copy), and slab 0's first named barrier below publishes the new one.
        float* lc = smemLnc + quarter * 2 * BN;
        const float* src = p.lnc + (size_t)(min(m0, p.M - 1) / p.rows_per_sample) * p.N + n0;
        const size_t c2off = (size_t)p.ln_samples * p.N;
        for (int i = half * 32 + lane; i < BN; i += 64) {
          const bool in = n0 + i < p.N;
          lc[i] = in ? __ldg(src + i) : 0.f;
          lc[BN + i] = in ? __ldg(src + c2off + i) : 0.f;
        }
        c1row = lc - n0;          // indexed by the global column like the bias
        c2row = lc + BN - n0;
      }
      const float* rope_row = nullptr;  // (cos, sin) pairs of this row's token
      if (ROPE) {
        if (p.rope_hd <= GT_ROPE_MAX_HD) {
          // the table rows of this tile's 128 tokens are staged in shared memory once per row block (column tiles of one row
          // block are consecutive): per-thread global reads of 288-byte-strided rows were the epilogue's critical path
          const int rs = p.rope_hd + 4;
          // a table shared by all samples (rope_rows == 1) depends on the token index only: with 256-token samples every tile of this CTA
          // holds the same 128 tokens, so the rows are staged once per kernel, not once per tile (37 KB from L2 each time)
          const int rope_key = p.rope_rows == 1 && m0 + GT_BM <= p.M ? m0 % p.rope_tokens : -2 - m0;
          if (rope_key != rope_m0) {
            const int q4 = p.rope_hd >> 2, total = QROWS * q4, et = half * 32 + lane;  // this quarter's 32 rows, by its 64 threads
            for (int i = et; i < total; i += 64) {
              const int rr = quarter * QROWS + i / q4, c4 = i - (i / q4) * q4;
              const int mm = min(m0 + rr, p.M - 1), smp = mm / p.rope_tokens, tok = mm - smp * p.rope_tokens;
              const float4 v = __ldg(reinterpret_cast<const float4*>(p.rope_freqs + ((size_t)(p.rope_rows == 1 ? 0 : smp) * p.rope_tokens + tok) * p.rope_hd) + c4);
              *reinterpret_cast<float4*>(smemRope + rr * rs + c4 * 4) = v;
            }
            rope_m0 = rope_key;  // visible to the quarter's threads after the first slab's named barrier
          }
          rope_row = smemRope + row * rs;
        } else {
          const int smp = m / p.rope_tokens, tok = m - smp * p.rope_tokens;
          rope_row = p.rope_freqs + ((size_t)(p.rope_rows == 1 ? 0 : smp) * p.rope_tokens + tok) * p.rope_hd;
        }
      }
      {
        // FITB200_EPI_SLEEP=<ns> (tuning): back off between polls while the accumulator is still being formed (measured: 117.3-118.3 vs
        // 118.0-118.6 ms per step, within box noise; default off)
        if (p.epi_sleep > 0) ptx::mbar_wait_sleep(&tfull_bar[as], aphase, (unsigned)p.epi_sleep); else ptx::mbar_wait(&tfull_bar[as], aphase);
      }
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * BN;
#ifdef FITB200_TUNING  // measurement-only builds (FITB200_TUNING=1 python -m fit_b200.build): mainloop ceiling with the epilogue skipped
      if (p.dbg == 1) {
        ptx::tc_fence_before();
        if (PAIR) { __syncwarp(); if (lane == 0) ptx::mbar_arrive_cluster_relaxed(tempty_leader + as * 8); } else ptx::mbar_arrive(&tempty_bar[as]);
        q += NSLAB;
        continue;
      }
#endif
#pragma unroll 1
      for (int slab = 0; slab < NSLAB; ++slab, ++q) {
        const int buf = q & 1;
        const int nacc = n0 + slab * ACCC;             // first accumulator column of the slab (global n)
        const int nout = SWIGLU ? (nacc >> 1) : nacc;  // first output column
        uint8_t* obuf = smemOut + buf * GT_SLAB_BYTES + row * 128;
        const uint8_t* rbuf = smemRes + buf * GT_SLAB_BYTES + row * 128;
        const bool live = nacc < p.N;  // tile-uniform: slabs beyond N are skipped (their store would be fully clipped)
        constexpr int HACC = ACCC / 2;  // accumulator columns of this warp's half slab
        // the accumulator columns are requested BEFORE the staging-buffer waits and the quarter barrier: the TMEM round trip overlaps them
        uint32_t accr[HACC];
        if (live) {
#pragma unroll
          for (int c = 0; c < HACC / 16; ++c) ptx::tmem_ld16(taddr + slab * ACCC + half * HACC + c * 16, *reinterpret_cast<uint32_t(*)[16]>(accr + c * 16));
        }
        // the TMA store that read this staging buffer two slabs ago must have finished reading it
        if (leader) ptx::tma_store_wait_read<1>();
#ifdef FITB200_TUNING
        if (RESID && !(p.dbg & 4)) ptx::mbar_wait(&res_bar[quarter * 2 + buf], (q >> 1) & 1);
#else
        if (RESID) ptx::mbar_wait(&res_bar[quarter * 2 + buf], (q >> 1) & 1);
#endif
        ptx::named_bar_sync(qbar, 64);
        if (live) {
          float acc[HACC];
          // RoPE table of this thread's 4 chunks, issued ahead of the TMEM loads (L2 latency hidden behind them)
          float4 fq[ROPE ? 4 : 1][CW / 4];
          if (ROPE) {
            int hcol = (nacc + half * 4 * CW) % p.rope_hd;  // head-dim index of the first chunk (chunks never straddle a head: 8 | head_dim)
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
              const bool rot = nacc + (half * 4 + jj) * CW < p.rope_cols;
#pragma unroll
              for (int i4 = 0; i4 < CW / 4; ++i4)
                fq[ROPE ? jj : 0][i4] = rot ? *reinterpret_cast<const float4*>(rope_row + hcol + i4 * 4) : make_float4(1.f, 0.f, 1.f, 0.f);
              hcol += CW;
              if (hcol >= p.rope_hd) hcol -= p.rope_hd;
            }
          }
          ptx::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < HACC; ++i) acc[i] = __uint_as_float(accr[i]);
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {  // this warp's 4 of the 8 x 16-byte chunks per 128-byte row
            const int j = half * 4 + jj;
            float v[8];
            if (SWIGLU) {
              // packed weight: every 32 accumulator columns = 16 gate rows then the matching 16 up rows
              const int a0 = (jj >> 1) * 32 + (jj & 1) * 8;  // inside this warp's half slab
              const int nb = nacc + half * HACC + a0;
              const bool ok = nb < p.N;
              const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
              const float* brow = LNA ? c2row : p.bias;   // LNA: c2 carries the bias (shared-memory copy: plain loads)
              auto ld4 = [&](const float* q) { return LNA ? *reinterpret_cast<const float4*>(q) : __ldg(reinterpret_cast<const float4*>(q)); };
              const float4 g0 = ok ? ld4(brow + nb) : z4;
              const float4 g1 = ok ? ld4(brow + nb + 4) : z4;
              const float4 u0 = ok ? ld4(brow + nb + 16) : z4;
              const float4 u1 = ok ? ld4(brow + nb + 20) : z4;
              float bg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
              float bu[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
              if (LNA) {
                const float4 h0 = ok ? ld4(c1row + nb) : z4;
                const float4 h1 = ok ? ld4(c1row + nb + 4) : z4;
                const float4 w0 = ok ? ld4(c1row + nb + 16) : z4;
                const float4 w1 = ok ? ld4(c1row + nb + 20) : z4;
                const float cg[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
                const float cu[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) { bg[i] = fmaf(ln_b, cg[i], bg[i]); bu[i] = fmaf(ln_b, cu[i], bu[i]); }
              }
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] = silu_f(fmaf(ln_a, acc[a0 + i], bg[i])) * fmaf(ln_a, acc[a0 + 16 + i], bu[i]);
            } else {
              const int nb = nacc + j * CW;
              const bool ok = nb < p.N;
#pragma unroll
              for (int i4 = 0; i4 < CW / 4; ++i4) {
                float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
                if (LNA) {
                  if (ok) {
                    const float4 c1 = *reinterpret_cast<const float4*>(c1row + nb + i4 * 4);
                    b4 = *reinterpret_cast<const float4*>(c2row + nb + i4 * 4);
                    b4.x = fmaf(ln_b, c1.x, b4.x); b4.y = fmaf(ln_b, c1.y, b4.y); b4.z = fmaf(ln_b, c1.z, b4.z); b4.w = fmaf(ln_b, c1.w, b4.w);
                  }
                } else if (p.bias && ok && m >= p.bias_row0) {
                  b4 = __ldg(reinterpret_cast<const float4*>(p.bias + nb + i4 * 4));
                }
                v[i4 * 4 + 0] = fmaf(ln_a, acc[jj * CW + i4 * 4 + 0], b4.x); v[i4 * 4 + 1] = fmaf(ln_a, acc[jj * CW + i4 * 4 + 1], b4.y);
                v[i4 * 4 + 2] = fmaf(ln_a, acc[jj * CW + i4 * 4 + 2], b4.z); v[i4 * 4 + 3] = fmaf(ln_a, acc[jj * CW + i4 * 4 + 3], b4.w);
              }
              if (ROPE) {  // complex_mult (fit.py:152-154): (a + ib)(c + id), pairs (2j, 2j+1); identity (1, 0) on the v columns
#pragma unroll
                for (int i4 = 0; i4 < CW / 4; ++i4) {
                  const float4 f = fq[ROPE ? jj : 0][i4];
                  const float a0 = v[i4 * 4 + 0], b0 = v[i4 * 4 + 1], a1 = v[i4 * 4 + 2], b1 = v[i4 * 4 + 3];
                  v[i4 * 4 + 0] = a0 * f.x - b0 * f.y; v[i4 * 4 + 1] = b0 * f.x + a0 * f.y;
                  v[i4 * 4 + 2] = a1 * f.z - b1 * f.w; v[i4 * 4 + 3] = b1 * f.z + a1 * f.w;
                }
              }
              if (EPI == FITB200_EPI_BIAS_SILU) {
#pragma unroll
                for (int i = 0; i < CW; ++i) v[i] = silu_f(v[i]);
              } else if (RESID) {
                const uint4 rraw = *reinterpret_cast<const uint4*>(rbuf + ((j ^ sw) << 4));
                float r[8];
                if constexpr (sizeof(OutT) == 2) {
                  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&rraw);
#pragma unroll
                  for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h2[i]); r[2 * i] = f.x; r[2 * i + 1] = f.y; }
                } else {
                  r[0] = __uint_as_float(rraw.x); r[1] = __uint_as_float(rraw.y); r[2] = __uint_as_float(rraw.z); r[3] = __uint_as_float(rraw.w);
                }
#pragma unroll
                for (int i4 = 0; i4 < CW / 4; ++i4) {
                  float4 g4 = make_float4(0.f, 0.f, 0.f, 0.f);
#ifdef FITB200_TUNING
                  if (ok && !(p.dbg & 8)) g4 = __ldg(reinterpret_cast<const float4*>(gate_row + nb + i4 * 4));
#else
                  if (ok) g4 = __ldg(reinterpret_cast<const float4*>(gate_row + nb + i4 * 4));
#endif
                  v[i4 * 4 + 0] = fmaf(g4.x, v[i4 * 4 + 0], r[i4 * 4 + 0]); v[i4 * 4 + 1] = fmaf(g4.y, v[i4 * 4 + 1], r[i4 * 4 + 1]);
                  v[i4 * 4 + 2] = fmaf(g4.z, v[i4 * 4 + 2], r[i4 * 4 + 2]); v[i4 * 4 + 3] = fmaf(g4.w, v[i4 * 4 + 3], r[i4 * 4 + 3]);
                }
#ifdef FITB200_TUNING
                if (ok && !(p.dbg & 2)) {
#else
                if (ok) {  // row statistics of the new residual stream: the LayerNorm that follows never re-reads x for them
#endif
#pragma unroll
                  for (int i = 0; i < CW; ++i) { st1 += v[i]; st2 = fmaf(v[i], v[i], st2); }
                }
              }
            }
            uint4 o;
            if constexpr (sizeof(OutT) == 2) {
              o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]); o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
            } else {
              o.x = __float_as_uint(v[0]); o.y = __float_as_uint(v[1]); o.z = __float_as_uint(v[2]); o.w = __float_as_uint(v[3]);
            }
            *reinterpret_cast<uint4*>(obuf + ((j ^ sw) << 4)) = o;
          }
        }
        if (RESID && slab == NSLAB - 1 && p.stats_out && m0 + row < p.M)
          p.stats_out[(size_t)(m0 + row) * p.stats_nparts + (tile % tiles_n) * 2 + half] = make_float2(st1, st2);
        if (slab == NSLAB - 1) {  // accumulator fully read: hand the TMEM stage back to the MMA warp
          ptx::tc_fence_before();
          if (PAIR) {
            // one relaxed arrival per warp: a release.cluster arrive by every thread cost a MEMBAR + ERRBAR each (9 % of the epilogue samples)
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster_relaxed(tempty_leader + as * 8);
          } else {
            ptx::mbar_arrive(&tempty_bar[as]);
          }
        }
        ptx::fence_proxy_async_smem();  // make the st.shared visible to the TMA (async proxy)
        ptx::named_bar_sync(qbar, 64);
        if (leader) {
          if (live) {
            if (nsplit > 1) ptx::tma_reduce_add_2d(&tmOut, smemOut + buf * GT_SLAB_BYTES + quarter * QBYTES, nout, m0 + quarter * QROWS);
            else ptx::tma_store_2d(&tmOut, smemOut + buf * GT_SLAB_BYTES + quarter * QBYTES, nout, m0 + quarter * QROWS);
          }
          ptx::tma_store_commit();
          if (RESID) issue_resid(q + 2);  // this residual buffer has been consumed by all 128 threads
        }
      }
    }
    if (leader) ptx::tma_store_wait_all();
  }

  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync();  // neither CTA may retire (or free TMEM) while the pair's MMAs / barrier traffic are in flight
  else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    if (PAIR) ptx::tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS); else ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
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

// 2-D row-major [rows, cols] tensor with leading dimension ld (elements); box = box_rows x box_cols, 128B swizzle
// (box_cols * element bytes must be 128)
static int make_tmap(CUtensorMap* m, const void* ptr, int dtype, long rows, long cols, long ld, int box_rows, int box_cols) {
  EncodeTiledFn enc = get_encode();
  if (!enc) { set_last_error("gemm_bf16_tc: cuTensorMapEncodeTiled not available from the driver"); return FITB200_ERR_CUDA; }
  const int eb = dtype == FITB200_BF16 ? 2 : 4;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * eb};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, dtype == FITB200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                   const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_error("gemm_bf16_tc: cuTensorMapEncodeTiled failed (%d) rows=%ld cols=%ld ld=%ld", (int)r, rows, cols, ld); return FITB200_ERR_CUDA; }
  return FITB200_OK;
}

template <int BN, int EPI, typename OutT, bool AMN = false, bool BMN = false, bool PAIR = false, bool LNA = false>
static int launch_tc(const fitb200_gemm_args& a, cudaStream_t s) {
  using Cfg = GtCfg<BN, EPI, PAIR, LNA>;
  CUtensorMap tmA, tmB, tmOut, tmRes, tmBt;
  int rc = AMN ? make_tmap(&tmA, a.A, FITB200_BF16, a.K, a.M, a.lda, GT_BK, 64) : make_tmap(&tmA, a.A, FITB200_BF16, a.M, a.K, a.lda, GT_BM, GT_BK);
  if (rc) return rc;
  rc = BMN ? make_tmap(&tmB, a.B, FITB200_BF16, a.K, a.N, a.ldb, GT_BK, 64) : make_tmap(&tmB, a.B, FITB200_BF16, a.N, a.K, a.ldb, Cfg::BNL, GT_BK);
  if (rc) return rc;
  const int outc = 128 / (int)sizeof(OutT);
  const long n_out = EPI == FITB200_EPI_SWIGLU ? a.N / 2 : a.N;
  rc = make_tmap(&tmOut, a.out, a.out_dtype, a.M, n_out, a.ldo, 32, outc);  // one store per lane quarter: 32 rows
  if (rc) return rc;
  tmRes = tmOut;
  tmBt = tmB;
  int tail_rows = 0;
  if (!BMN && a.N % BN != 0 && a.N > BN) {  // the cut last column tile: a box of just the rows each CTA's MMAs read
    const int gran = PAIR ? 32 : 16, nn = (a.N % BN + gran - 1) / gran * gran;
    tail_rows = PAIR ? nn / 2 : nn;
    if (tail_rows < Cfg::BNL) {
      rc = make_tmap(&tmBt, a.B, FITB200_BF16, a.N, a.K, a.ldb, tail_rows, GT_BK);
      if (rc) return rc;
    } else {
      tail_rows = 0;
    }
  }
  if (EPI == FITB200_EPI_BIAS_GATE_RESID) {
    rc = make_tmap(&tmRes, a.resid, a.out_dtype, a.M, a.N, a.ld_resid, 32, outc);
    if (rc) return rc;
  }
  GtParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.bias = a.bias;
  p.gate = a.gate; p.ld_gate = a.ld_gate; p.rows_per_sample = a.rows_per_sample;
  p.rope_freqs = a.rope_freqs; p.rope_rows = a.rope_rows; p.rope_tokens = a.rope_tokens; p.rope_hd = a.rope_head_dim; p.rope_cols = a.rope_cols;
#ifdef FITB200_TUNING
  static const int dbg = [] { const char* e = getenv("FITB200_GEMM_DBG"); return e ? atoi(e) : 0; }();
  p.dbg = dbg;
#else
  p.dbg = 0;
#endif
  p.tail_rows = tail_rows;
  p.bias_row0 = a.bias_row0;
  static const int epi_sleep = [] { const char* e = getenv("FITB200_EPI_SLEEP"); return e ? atoi(e) : 0; }();
  p.epi_sleep = epi_sleep;
  p.lnx = (const __nv_bfloat16*)a.A; p.ld_lnx = a.lda;
  p.lng = (const __nv_bfloat16*)a.ln_g; p.ld_lng = a.ld_ln_g;
  p.lnc = a.ln_c; p.ln_samples = a.ln_samples;
  p.ln_parts = (const float2*)a.ln_parts; p.ln_nparts = a.ln_nparts; p.ln_eps = a.ln_eps;
  p.stats_out = (float2*)a.stats_out; p.stats_nparts = 2 * ceil_div(a.N, BN);
  constexpr int THREADS = LNA ? GT_THREADS_LNA : GT_THREADS;
  auto kern = gemm_tc_kernel<BN, EPI, OutT, AMN, BMN, PAIR, LNA>;
  static bool attr_done = false;  // per template instantiation
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) { set_last_error("gemm_bf16_tc: cannot set %d bytes of dynamic smem: %s", Cfg::SMEM_BYTES, cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    attr_done = true;
  }
  p.splits = 1; p.kb_per_split = 0;
  if (PAIR) {
    int tiles = ceil_div(a.M, 2 * GT_BM) * ceil_div(a.N, BN);
    if constexpr (AMN && BMN && sizeof(OutT) == 4 && EPI == FITB200_EPI_BIAS) {
      // weight-gradient GEMMs (both operands contraction-major, fp32 output, no bias): few output tiles (dW_proj: 25 on 74 CTA pairs) over a
      // long contraction (the token rows).  Split the contraction so that the work items fill whole waves; the partial tiles are ADDED into
      // the zeroed output by TMA reduce stores (fp32 adds in arrival order: the last bits of a gradient may differ from run to run).
      // FITB200_GEMM_SPLITK=0 switches it off.
      // FITB200_GEMM_SPLITK: 0 = off, n >= 2 = force n splits (tuning), unset = cost model
      static const int splitk = [] { const char* e = getenv("FITB200_GEMM_SPLITK"); return e ? atoi(e) : -1; }();
      const int num_kb = ceil_div(a.K, GT_BK), slots = sm_count() / 2;
      if (splitk != 0 && a.bias == nullptr && num_kb >= 32 && slots > 0) {
        // cost of a schedule in k-slab times: waves x (k-slabs per item + the item's epilogue, ~12 k-slabs for a 256 x 256 fp32 tile that is
        // not hidden behind a following mainloop)
        auto cost = [&](int sp) { return (double)ceil_div(tiles * sp, slots) * ((double)ceil_div(num_kb, sp) + 12.0); };
        int best = 1; double best_cost = cost(1);
        for (int sp = 2; sp <= 8 && num_kb / sp >= 16; ++sp)
          if (cost(sp) < 0.9 * best_cost) { best = sp; best_cost = cost(sp); }
        if (splitk >= 2) best = splitk <= num_kb ? splitk : 1;
        if (best > 1) {
          const int kbs = ceil_div(num_kb, best);
          p.splits = ceil_div(num_kb, kbs);          // every split non-empty
          p.kb_per_split = kbs;
          cudaError_t e = cudaMemset2DAsync(a.out, (size_t)a.ldo * sizeof(OutT), 0, (size_t)a.N * sizeof(OutT), (size_t)a.M, s);
          if (e != cudaSuccess) { set_last_error("gemm_bf16_tc: split-K memset failed: %s", cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
          tiles *= p.splits;
        }
      }
    }
    const int pairs = tiles < sm_count() / 2 ? tiles : sm_count() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = Cfg::SMEM_BYTES; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmOut, tmRes, tmBt, p);
    if (e != cudaSuccess) { set_last_error("gemm_bf16_tc: pair launch failed: %s", cudaGetErrorString(e)); return FITB200_ERR_CUDA; }
    return FITB200_OK;
  }
  const int tiles = ceil_div(a.M, GT_BM) * ceil_div(a.N, BN);
  const int grid = tiles < sm_count() ? tiles : sm_count();
  kern<<<grid, THREADS, Cfg::SMEM_BYTES, s>>>(tmA, tmB, tmOut, tmRes, tmBt, p);
  FITB200_CHECK_LAUNCH("gemm_bf16_tc");
  return FITB200_OK;
}

template <int BN, typename OutT>
static int dispatch_epi(const fitb200_gemm_args& a, cudaStream_t s) {
  if (a.a_mn || a.b_mn) {  // backward GEMMs: plain (optional bias) epilogue only
    if (a.epilogue != FITB200_EPI_BIAS) { set_last_error("gemm_bf16_tc: a_mn / b_mn support FITB200_EPI_BIAS only"); return FITB200_ERR_UNSUPPORTED; }
    static const bool pair_mn = [] { const char* e = getenv("FITB200_GEMM_PAIR_MN"); return !(e && e[0] == '0'); }();
    if constexpr (BN != 192) if (pair_mn && a.M > GT_BM) {
      if (a.a_mn && a.b_mn) return launch_tc<BN, FITB200_EPI_BIAS, OutT, true, true, true>(a, s);
      if (a.b_mn) return launch_tc<BN, FITB200_EPI_BIAS, OutT, false, true, true>(a, s);
      return launch_tc<BN, FITB200_EPI_BIAS, OutT, true, false, true>(a, s);
    }
    if (a.a_mn && a.b_mn) return launch_tc<BN, FITB200_EPI_BIAS, OutT, true, true>(a, s);
    if (a.b_mn) return launch_tc<BN, FITB200_EPI_BIAS, OutT, false, true>(a, s);
    return launch_tc<BN, FITB200_EPI_BIAS, OutT, true, false>(a, s);
  }
  // K-major operands: CTA pairs (256 x BN tiles, cta_group::2) whenever there are at least two row blocks; FITB200_GEMM_PAIR=0 forces single-CTA tiles
  static const bool pair_ok = [] { const char* e = getenv("FITB200_GEMM_PAIR"); return !(e && e[0] == '0'); }();
  if (a.ln_g) {  // LayerNorm + modulate prologue: 256-wide pair tiles (any N: a narrow output is one cut tile per row block)
    if constexpr (BN == 256) {
      if (a.M > GT_BM) {
        if (a.epilogue == FITB200_EPI_BIAS) return launch_tc<BN, FITB200_EPI_BIAS, OutT, false, false, true, true>(a, s);
        if constexpr (sizeof(OutT) == 2) {
          if (a.epilogue == FITB200_EPI_BIAS_ROPE) return launch_tc<BN, FITB200_EPI_BIAS_ROPE, OutT, false, false, true, true>(a, s);
          if (a.epilogue == FITB200_EPI_SWIGLU) return launch_tc<BN, FITB200_EPI_SWIGLU, OutT, false, false, true, true>(a, s);
        }
      }
    }
    set_last_error("gemm_bf16_tc: the LN prologue needs M > 128 and a BIAS (bf16 / fp32 output) or BIAS_ROPE / SWIGLU (bf16 output) epilogue");
    return FITB200_ERR_UNSUPPORTED;
  }
  if (pair_ok && a.M > GT_BM) {
    switch (a.epilogue) {
      case FITB200_EPI_BIAS: return launch_tc<BN, FITB200_EPI_BIAS, OutT, false, false, true>(a, s);
      case FITB200_EPI_BIAS_SILU: return launch_tc<BN, FITB200_EPI_BIAS_SILU, OutT, false, false, true>(a, s);
      case FITB200_EPI_BIAS_ROPE: return launch_tc<BN, FITB200_EPI_BIAS_ROPE, OutT, false, false, true>(a, s);
      case FITB200_EPI_SWIGLU:
        if constexpr (BN != 192) return launch_tc<BN, FITB200_EPI_SWIGLU, OutT, false, false, true>(a, s);
        break;
      case FITB200_EPI_BIAS_GATE_RESID: return launch_tc<BN, FITB200_EPI_BIAS_GATE_RESID, OutT, false, false, true>(a, s);
    }
  }
  switch (a.epilogue) {
    case FITB200_EPI_BIAS: return launch_tc<BN, FITB200_EPI_BIAS, OutT>(a, s);
    case FITB200_EPI_BIAS_SILU: return launch_tc<BN, FITB200_EPI_BIAS_SILU, OutT>(a, s);
    case FITB200_EPI_BIAS_ROPE: return launch_tc<BN, FITB200_EPI_BIAS_ROPE, OutT>(a, s);
    case FITB200_EPI_SWIGLU:
      if constexpr (BN != 192) return launch_tc<BN, FITB200_EPI_SWIGLU, OutT>(a, s);  // 128 accumulator columns per slab
      break;
    case FITB200_EPI_BIAS_GATE_RESID: return launch_tc<BN, FITB200_EPI_BIAS_GATE_RESID, OutT>(a, s);
  }
  set_last_error("gemm_bf16_tc: unsupported epilogue %d", a.epilogue);
  return FITB200_ERR_INVALID_ARG;
}

template <typename OutT>
static int dispatch_bn(const fitb200_gemm_args& a, cudaStream_t s) {
  // 256-wide tiles (the last one of a row is cut to the columns that are left): narrower tiles stage more operand bytes per flop.
  // FITB200_GEMM_BN=128|192|256 overrides (tuning / tests)
  static const int force = [] { const char* e = getenv("FITB200_GEMM_BN"); return e ? atoi(e) : 0; }();
  // (tried: 128-wide tiles when they need fewer wave x width units at training sizes -- 47.4 vs 42.5 ms per XL/2 step: N = 128 tiles are
  // shared-memory-bandwidth bound, so 256 stays)
  const int best = a.ln_g ? 256 : force ? force : (a.N <= 128 ? 128 : 256);
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
  FITB200_CHECK_ARG((a.K % 8 == 0 || (a.a_mn && a.b_mn)) && a.lda % 8 == 0 && a.ldb % 8 == 0, "gemm_bf16_tc: K, lda, ldb must be multiples of 8 (K=%d lda=%lld ldb=%lld)", a.K, (long long)a.lda, (long long)a.ldb);
  FITB200_CHECK_ARG(!a.a_mn || a.M % 8 == 0, "gemm_bf16_tc: a_mn needs M %% 8 == 0 (M=%d)", a.M);
  FITB200_CHECK_ARG(((uintptr_t)a.A & 15) == 0 && ((uintptr_t)a.B & 15) == 0 && ((uintptr_t)a.out & 15) == 0, "gemm_bf16_tc: A, B, out must be 16-byte aligned");
  FITB200_CHECK_ARG(a.N % 8 == 0 && a.ldo % 8 == 0, "gemm_bf16_tc: N and ldo must be multiples of 8 (N=%d)", a.N);
  FITB200_CHECK_ARG(a.bias == nullptr || ((uintptr_t)a.bias & 15) == 0, "gemm_bf16_tc: bias must be 16-byte aligned");
  if (a.epilogue == FITB200_EPI_SWIGLU) FITB200_CHECK_ARG(a.N % 32 == 0 && (a.bias || a.ln_g), "gemm_bf16_tc: SWIGLU needs N %% 32 == 0 and a packed bias");
  if (a.epilogue == FITB200_EPI_BIAS_ROPE)
    FITB200_CHECK_ARG(a.rope_freqs && ((uintptr_t)a.rope_freqs & 15) == 0 && a.rope_tokens > 0 && a.rope_head_dim > 0 && a.rope_head_dim % 8 == 0 &&
                          a.rope_cols % a.rope_head_dim == 0 && a.rope_cols <= a.N && (a.rope_rows == 1 || (int64_t)a.rope_rows * a.rope_tokens >= a.M),
                      "gemm_bf16_tc: BIAS_ROPE needs a 16-byte aligned fp32 table [1|samples, tokens, head_dim], head_dim %% 8 == 0, rope_cols a multiple of head_dim");
  if (a.epilogue == FITB200_EPI_BIAS_GATE_RESID)
    FITB200_CHECK_ARG(a.gate && a.resid && a.rows_per_sample > 0 && a.ld_gate % 4 == 0 && a.ld_resid % 8 == 0 &&
                          ((uintptr_t)a.gate & 15) == 0 && ((uintptr_t)a.resid & 15) == 0 && a.resid_dtype == a.out_dtype,
                      "gemm_bf16_tc: GATE_RESID needs 16-byte aligned gate/resid, rows_per_sample, and resid dtype == out dtype");
  if (a.ln_g)
    FITB200_CHECK_ARG(!a.a_mn && !a.b_mn && a.ln_c && a.ln_parts && a.ln_nparts > 0 && a.ln_samples > 0 && a.rows_per_sample > 0 && a.ld_ln_g % 8 == 0 &&
                          ((uintptr_t)a.ln_g & 15) == 0 && ((uintptr_t)a.ln_c & 15) == 0 && ((uintptr_t)a.ln_parts & 7) == 0 &&
                          (int64_t)a.ln_samples * a.rows_per_sample >= a.M,
                      "gemm_bf16_tc: LN prologue needs ln_g [samples, K] bf16, ln_c [2 samples, N] fp32, ln_parts [M, nparts] and rows_per_sample");
  if (a.ln_g) FITB200_CHECK_ARG(a.rows_per_sample % GT_BM == 0 && a.K % GT_BK == 0, "gemm_bf16_tc: LN prologue needs rows_per_sample %% 128 == 0 and K %% 64 == 0 (got %d, %d)", a.rows_per_sample, a.K);
  if (a.stats_out)
    FITB200_CHECK_ARG(a.epilogue == FITB200_EPI_BIAS_GATE_RESID && !a.a_mn && !a.b_mn && ((uintptr_t)a.stats_out & 7) == 0, "gemm_bf16_tc: stats_out goes with GATE_RESID");
  if (a.out_dtype == FITB200_BF16) return dispatch_bn<__nv_bfloat16>(a, (cudaStream_t)stream);
  if (a.out_dtype == FITB200_F32) return dispatch_bn<float>(a, (cudaStream_t)stream);
  set_last_error("gemm_bf16_tc: unsupported out dtype %d", a.out_dtype);
  return FITB200_ERR_INVALID_ARG;
}

extern "C" int fitb200_gemm_stats_parts(int N) {
  static const int force = [] { const char* e = getenv("FITB200_GEMM_BN"); return e ? atoi(e) : 0; }();
  const int bn = force ? force : (N <= 128 ? 128 : 256);
  return 2 * ((N + bn - 1) / bn);
}
