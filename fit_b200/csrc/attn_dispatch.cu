// fitb200_attn_fwd: one entry point, three kernels.
//   fp32 (verification mode)                      -> attn_simt.cu   CUDA cores, RoPE applied in the kernel from the (cos, sin) table
//   bf16, q / k pre-rotated, T <= 256             -> attn_tc2.cu    persistent warp-specialised tcgen05 / TMEM kernel (every trained shape)
//   bf16, q / k pre-rotated, T >  256             -> attn_tc3.cu    KV-blocked tcgen05 kernel (320^2 ... 1024^2 extrapolation)
// bf16 operands always carry the RoPE already: the qkv GEMM's FITB200_EPI_BIAS_ROPE epilogue writes rope(q), rope(k) from its fp32 accumulators
// (apply_rotary_emb, fit.py:160-174), so the attention kernels fetch finished operands with TMA.  (Round 1 also kept a tcgen05 kernel and an
// mma.sync kernel that rotated bf16 q / k in registers; no product path reached them and they were removed.)
#include <cstdlib>

#include "common.cuh"

namespace fitb200 {
int attn_fwd_f32(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask, int R, int T, int h, int d,
                 void* out, float* lse, cudaStream_t s);
int attn_fwd_tc2(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s);
int attn_fwd_tc3(const void* qkv, const int32_t* kv_len, const uint8_t* mask, int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s);
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
  if (dtype == FITB200_F32) {
    if (pre) { set_last_error("attn_fwd: PREROTATED is a bf16-path flag"); return FITB200_ERR_UNSUPPORTED; }
    return attn_fwd_f32(qkv, freqs, freqs_rows, kv_len, mask, R, T, h, d, out, lse, s);
  }
  if (dtype == FITB200_BF16) {
    if (!pre) {
      set_last_error("attn_fwd: bf16 q / k must carry the RoPE already (FITB200_ATTN_PREROTATED: write them with a FITB200_EPI_BIAS_ROPE GEMM); "
                     "the fp32 path rotates in the kernel");
      return FITB200_ERR_UNSUPPORTED;
    }
    if (d != 64 && d != 72) { set_last_error("attn_fwd: head_dim %d not supported (64 or 72)", d); return FITB200_ERR_UNSUPPORTED; }
    FITB200_CHECK_ARG(((uintptr_t)qkv & 15) == 0 && ((uintptr_t)out & 15) == 0, "attn_fwd: qkv and out must be 16-byte aligned");
    // FITB200_ATTN_TC3=1: the KV-blocked kernel on the short shapes too (A/B timing)
    static const bool v3 = [] { const char* e = getenv("FITB200_ATTN_TC3"); return e && e[0] == '1'; }();
    if (T <= 256 && !v3) return attn_fwd_tc2(qkv, kv_len, mask, R, T, h, d, out, lse, s);
    return attn_fwd_tc3(qkv, kv_len, mask, R, T, h, d, out, lse, s);
  }
  set_last_error("attn_fwd: unsupported dtype %d", dtype);
  return FITB200_ERR_INVALID_ARG;
}
