// fp32 CUDA-core attention for the verification mode: thread-per-query online softmax with the 2-D
// RoPE applied while q / k are loaded, key masking from kv_len + mask bytes.
// (The bf16 kernels are attn_tc2.cu / attn_tc3.cu; the shared entry point fitb200_attn_fwd is in attn_dispatch.cu.)
#include "common.cuh"

namespace fitb200 {

constexpr int AS_Q = 128;   // queries per CTA (one per thread)
constexpr int AS_KT = 32;   // keys per shared-memory tile

template <int HD, typename T>
__global__ void __launch_bounds__(AS_Q) attn_simt_kernel(const T* __restrict__ qkv, const float* __restrict__ freqs,
                                                         int freqs_rows, const int32_t* __restrict__ kv_len,
                                                         const uint8_t* __restrict__ mask, int R, int Tn, int h,
                                                         T* __restrict__ out, float* __restrict__ lse) {
  __shared__ float Ks[AS_KT][HD];
  __shared__ float Vs[AS_KT][HD];
  __shared__ int valid[AS_KT];

  const int r = blockIdx.z, head = blockIdx.y, q0 = blockIdx.x * AS_Q;
  const int tq = q0 + threadIdx.x;
  const int kvl = min(kv_len[r], Tn);
  const size_t row_stride = (size_t)3 * h * HD;
  const T* base = qkv + (size_t)r * Tn * row_stride;
  T* orow = out + ((size_t)r * Tn + tq) * (size_t)(h * HD) + head * HD;

  if (q0 >= kvl) {  // padded query tile: defined (zero) output, no work
    if (tq < Tn) {
#pragma unroll
      for (int i = 0; i < HD; ++i) orow[i] = from_f32<T>(0.f);
      if (lse) lse[((size_t)r * h + head) * Tn + tq] = 0.f;
    }
    return;
  }

  const float scale = rsqrtf((float)HD);
  const float* fr = freqs + (size_t)(freqs_rows == 1 ? 0 : r) * Tn * HD;

  float q[HD], acc[HD];
#pragma unroll
  for (int i = 0; i < HD; ++i) acc[i] = 0.f;
  if (tq < Tn) {
    const T* qp = base + (size_t)tq * row_stride + head * HD;
    const float* f = fr + (size_t)tq * HD;
#pragma unroll
    for (int i = 0; i < HD; i += 2) {
      float a = to_f32(qp[i]), b = to_f32(qp[i + 1]);
      float c = f[i], s = f[i + 1];
      // complex_mult (fit.py:152-154), result cast back to the storage dtype (fit.py:169)
      q[i] = to_f32(from_f32<T>(a * c - b * s));
      q[i + 1] = to_f32(from_f32<T>(b * c + a * s));
    }
  } else {
#pragma unroll
    for (int i = 0; i < HD; ++i) q[i] = 0.f;
  }
  float m = -INFINITY, l = 0.f;

  for (int k0 = 0; k0 < kvl; k0 += AS_KT) {
    __syncthreads();
    for (int i = threadIdx.x; i < AS_KT * (HD / 2); i += AS_Q) {
      int j = i / (HD / 2), pr = i - j * (HD / 2);
      int tk = k0 + j;
      float k_re = 0.f, k_im = 0.f, v0 = 0.f, v1 = 0.f;
      if (tk < kvl) {
        const T* kp = base + (size_t)tk * row_stride + (size_t)h * HD + head * HD + 2 * pr;
        const T* vp = kp + (size_t)h * HD;
        float a = to_f32(kp[0]), b = to_f32(kp[1]);
        float c = fr[(size_t)tk * HD + 2 * pr], s = fr[(size_t)tk * HD + 2 * pr + 1];
        k_re = to_f32(from_f32<T>(a * c - b * s));
        k_im = to_f32(from_f32<T>(b * c + a * s));
        v0 = to_f32(vp[0]); v1 = to_f32(vp[1]);
      }
      Ks[j][2 * pr] = k_re; Ks[j][2 * pr + 1] = k_im;
      Vs[j][2 * pr] = v0; Vs[j][2 * pr + 1] = v1;
    }
    if (threadIdx.x < AS_KT) {
      int tk = k0 + threadIdx.x;
      valid[threadIdx.x] = (tk < kvl) && (mask == nullptr || mask[(size_t)r * Tn + tk] != 0);
    }
    __syncthreads();

    float s[AS_KT];
    float tmax = -INFINITY;
#pragma unroll
    for (int j = 0; j < AS_KT; ++j) {
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < HD; ++i) d = fmaf(q[i], Ks[j][i], d);
      s[j] = valid[j] ? d * scale : -INFINITY;
      tmax = fmaxf(tmax, s[j]);
    }
    if (tmax == -INFINITY) continue;  // no valid key in this tile (block-uniform: same for all threads)
    const float m_new = fmaxf(m, tmax);
    const float corr = expf(m - m_new);  // m = -inf -> 0
    l *= corr;
#pragma unroll
    for (int i = 0; i < HD; ++i) acc[i] *= corr;
#pragma unroll
    for (int j = 0; j < AS_KT; ++j) {
      const float p = expf(s[j] - m_new);  // -inf -> 0
      l += p;
#pragma unroll
      for (int i = 0; i < HD; ++i) acc[i] = fmaf(p, Vs[j][i], acc[i]);
    }
    m = m_new;
  }

  if (tq < Tn) {
    const float inv = l > 0.f ? 1.0f / l : 0.f;
#pragma unroll
    for (int i = 0; i < HD; ++i) orow[i] = from_f32<T>(acc[i] * inv);
    if (lse) lse[((size_t)r * h + head) * Tn + tq] = l > 0.f ? m + logf(l) : 0.f;
  }
}

template <int HD, typename T>
static int launch_attn_simt(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len,
                            const uint8_t* mask, int R, int Tn, int h, void* out, float* lse, cudaStream_t s) {
  dim3 grid((unsigned)ceil_div(Tn, AS_Q), (unsigned)h, (unsigned)R);
  attn_simt_kernel<HD, T><<<grid, AS_Q, 0, s>>>((const T*)qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, (T*)out, lse);
  FITB200_CHECK_LAUNCH("attn_simt");
  return FITB200_OK;
}

int attn_fwd_f32(const void* qkv, const float* freqs, int freqs_rows, const int32_t* kv_len, const uint8_t* mask,
                 int R, int Tn, int h, int d, void* out, float* lse, cudaStream_t s) {
  if (d == 64) return launch_attn_simt<64, float>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, out, lse, s);
  if (d == 72) return launch_attn_simt<72, float>(qkv, freqs, freqs_rows, kv_len, mask, R, Tn, h, out, lse, s);
  set_last_error("attn_fwd: head_dim %d not supported (64 or 72)", d);
  return FITB200_ERR_UNSUPPORTED;
}

}  // namespace fitb200
