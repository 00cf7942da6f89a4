// Training-loop neighbours of the hot path.
// (1) Fused multi-tensor AdamW + EMA step (row f3).  Replaces torch.optim.AdamW.step (FiT/train.py:215-217) and the EMA
// wrapper's _foreach_mul_ / _foreach_add_ pair on a side stream (FiT/ema.py:149-154,256-267) by ONE pass: every element of
// (p, g, m, v, ema) is read once and (p, m, v, ema) written once -- 36 B per element, HBM-bound (FiT-XL/2: 674.9 M elements = 24.3 GB);
// optionally the bf16 / packed copy of the parameter that the GEMMs read is written in the same pass (+2 B per element).
// One CTA = one FITB200_ADAM_CHUNK-element chunk of one tensor (chunk table built once by the host), 16-byte accesses, grid = chunks.
#include "common.cuh"

namespace fitb200 {

struct AdamScalars { float step, decay_p, omb1, beta2, omb2, bc2_sqrt, eps, ema_decay, ome; };

// derived scalars in double, as torch forms them (step_size = lr / bc1, value = 1 - beta2, alpha = 1 - decay), then one rounding to fp32
__host__ __device__ inline AdamScalars adam_scalars(double lr, double beta1, double beta2, double eps, double weight_decay, double bc1,
                                                     double bc2_sqrt, double ema_decay) {
  AdamScalars sc;
  sc.step = (float)(lr / bc1); sc.decay_p = (float)(1.0 - lr * weight_decay); sc.omb1 = (float)(1.0 - beta1);
  sc.beta2 = (float)beta2; sc.omb2 = (float)(1.0 - beta2); sc.bc2_sqrt = (float)bc2_sqrt; sc.eps = (float)eps;
  sc.ema_decay = (float)ema_decay; sc.ome = (float)(1.0 - ema_decay);
  return sc;
}

// DEV: the step-dependent scalars come from DEVICE memory (hyper = {lr, beta1, beta2, eps, weight_decay, ema_decay}, step_dev = the number
// of this step, every_n = EMA period), so that one captured CUDA graph serves every step (torch.optim's capturable=True)
template <bool DEV>
__global__ void __launch_bounds__(256) adamw_ema_kernel(const fitb200_adam_tensor* __restrict__ tensors, const int32_t* __restrict__ chunks,
                                                        const AdamScalars sc_host, const double* __restrict__ hyper,
                                                        const int64_t* __restrict__ step_dev, int every_n) {
  __shared__ AdamScalars sc_sh;
  if (DEV) {
    if (threadIdx.x == 0) {
      const double lr = hyper[0], b1 = hyper[1], b2 = hyper[2], t = (double)step_dev[0];
      const bool ema_now = hyper[5] >= 0.0 && (step_dev[0] - 1) % (every_n > 0 ? every_n : 1) == 0;   // FiT/ema.py:252-257
      sc_sh = adam_scalars(lr, b1, b2, hyper[3], hyper[4], 1.0 - pow(b1, t), sqrt(1.0 - pow(b2, t)), ema_now ? hyper[5] : -1.0);
    }
    __syncthreads();
  }
  const AdamScalars sc = DEV ? sc_sh : sc_host;
  const int ti = chunks[2 * blockIdx.x], first = chunks[2 * blockIdx.x + 1];
  const fitb200_adam_tensor t = tensors[ti];
  const int64_t base = (int64_t)first;
  const int n = (int)min((int64_t)FITB200_ADAM_CHUNK, t.numel - base);
  float* __restrict__ p = t.p + base;
  const float* __restrict__ g = t.g + base;
  float* __restrict__ m = t.m + base;
  float* __restrict__ v = t.v + base;
  float* __restrict__ e = t.ema ? t.ema + base : nullptr;
  const float step = sc.step, decay_p = sc.decay_p, omb1 = sc.omb1, beta2 = sc.beta2, omb2 = sc.omb2, bc2_sqrt = sc.bc2_sqrt, eps = sc.eps,
              ema_decay = sc.ema_decay, ome = sc.ome;
  const bool do_ema = e != nullptr && ema_decay >= 0.f;
  // kernel-layout copy of the updated parameter (fitb200_adam_tensor.shadow)
  auto shadow_index = [&](int64_t idx) {
    const int64_t row = idx / t.row_len, col = idx - row * t.row_len;
    const int64_t srow = t.pack_off < 0 ? row : (row >> 4) * 32 + t.pack_off + (row & 15);
    return srow * t.shadow_ld + col;
  };
  auto upd = [&](float& pp, float gg, float& mm, float& vv, float& ee) {
    pp *= decay_p;
    mm = mm + (gg - mm) * omb1;                       // exp_avg.lerp_(grad, 1 - beta1)
    vv = fmaf(gg * gg, omb2, vv * beta2);             // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    const float denom = sqrtf(vv) / bc2_sqrt + eps;   // (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
    pp = pp - step * (mm / denom);                    // param.addcdiv_(exp_avg, denom, value=-step_size)
    if (do_ema) ee = ee * ema_decay + pp * ome;       // _foreach_mul_(ema, decay); _foreach_add_(ema, p, alpha=1 - decay)
  };
  const int n4 = n >> 2;  // chunk starts are multiples of 16384 elements of 16-byte aligned tensors: float4 accesses are aligned
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    float4 p4 = reinterpret_cast<float4*>(p)[i];
    const float4 g4 = reinterpret_cast<const float4*>(g)[i];
    float4 m4 = reinterpret_cast<float4*>(m)[i], v4 = reinterpret_cast<float4*>(v)[i];
    float4 e4 = do_ema ? reinterpret_cast<float4*>(e)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    upd(p4.x, g4.x, m4.x, v4.x, e4.x); upd(p4.y, g4.y, m4.y, v4.y, e4.y);
    upd(p4.z, g4.z, m4.z, v4.z, e4.z); upd(p4.w, g4.w, m4.w, v4.w, e4.w);
    reinterpret_cast<float4*>(p)[i] = p4;
    reinterpret_cast<float4*>(m)[i] = m4;
    reinterpret_cast<float4*>(v)[i] = v4;
    if (do_ema) reinterpret_cast<float4*>(e)[i] = e4;
    if (t.shadow) {
      const int64_t idx = base + 4 * (int64_t)i;
      if ((t.row_len & 3) == 0 && (t.shadow_ld & 3) == 0) {   // the four elements share a row: one 8- or 16-byte store
        const int64_t si = shadow_index(idx);
        if (t.shadow_dtype == FITB200_BF16) {
          uint2 o; o.x = pack_bf16x2(p4.x, p4.y); o.y = pack_bf16x2(p4.z, p4.w);
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(t.shadow) + si) = o;
        } else {
          *reinterpret_cast<float4*>(reinterpret_cast<float*>(t.shadow) + si) = p4;
        }
      } else {
        const float pv[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t si = shadow_index(idx + k);
          if (t.shadow_dtype == FITB200_BF16) reinterpret_cast<__nv_bfloat16*>(t.shadow)[si] = __float2bfloat16_rn(pv[k]);
          else reinterpret_cast<float*>(t.shadow)[si] = pv[k];
        }
      }
    }
  }
  for (int i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) {
    float pp = p[i], mm = m[i], vv = v[i], ee = do_ema ? e[i] : 0.f;
    upd(pp, g[i], mm, vv, ee);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (do_ema) e[i] = ee;
    if (t.shadow) {
      const int64_t si = shadow_index(base + i);
      if (t.shadow_dtype == FITB200_BF16) reinterpret_cast<__nv_bfloat16*>(t.shadow)[si] = __float2bfloat16_rn(pp);
      else reinterpret_cast<float*>(t.shadow)[si] = pp;
    }
  }
}

// masked_FiT token sub-sampling + RoPE row gather (row f4): one warp per output token, lanes copy the P latent and d RoPE floats.
__global__ void __launch_bounds__(256) subsample_tokens_kernel(const float* __restrict__ latent, const float* __restrict__ pos,
                                                               const uint8_t* __restrict__ mask, const int64_t* __restrict__ perm,
                                                               const int32_t* __restrict__ n_tokens, int Tsrc, int P, int d, int Tout,
                                                               float* __restrict__ latent_out, float* __restrict__ pos_out,
                                                               uint8_t* __restrict__ mask_out) {
  const int s = blockIdx.y, j = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (j >= Tout) return;
  const int nt = min(n_tokens[s], Tsrc);
  long src = -1;  // source token, -1 = padding
  if (nt > Tout) {
    src = perm[(size_t)s * Tsrc + j];
    if (src < 0 || src >= nt) src = -1;  // a malformed permutation yields padding instead of an out-of-bounds read
  } else if (j < nt) {
    src = j;
  }
  const size_t so = (size_t)s * Tsrc + (src < 0 ? 0 : src), oo = (size_t)s * Tout + j;
  for (int i = lane; i < P; i += 32) latent_out[oo * P + i] = src < 0 ? 0.f : latent[so * P + i];
  for (int i = lane; i < d; i += 32) pos_out[oo * d + i] = src < 0 ? 0.f : pos[so * d + i];
  if (lane == 0) mask_out[oo] = src < 0 ? (uint8_t)0 : mask[so];
}

}  // namespace fitb200

using namespace fitb200;

extern "C" int fitb200_subsample_tokens(const float* latent, const float* pos, const uint8_t* mask, const int64_t* perm, const int32_t* n_tokens, int N,
                                        int Tsrc, int P, int d, int Tout, float* latent_out, float* pos_out, uint8_t* mask_out,
                                        fitb200_stream_t stream) {
  FITB200_CHECK_ARG(latent && pos && mask && perm && n_tokens && latent_out && pos_out && mask_out, "subsample_tokens: null pointer");
  FITB200_CHECK_ARG(N > 0 && Tsrc > 0 && P > 0 && d > 0 && Tout > 0 && N <= 65535, "subsample_tokens: bad sizes (N=%d Tsrc=%d P=%d d=%d Tout=%d)", N, Tsrc, P, d, Tout);
  dim3 grid((unsigned)ceil_div(Tout, 8), (unsigned)N);
  subsample_tokens_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(latent, pos, mask, perm, n_tokens, Tsrc, P, d, Tout, latent_out, pos_out, mask_out);
  FITB200_CHECK_LAUNCH("subsample_tokens");
  return FITB200_OK;
}

extern "C" int fitb200_adamw_ema_step(const fitb200_adam_tensor* tensors, const int32_t* chunks, int n_chunks, double lr, double beta1, double beta2,
                                      double eps, double weight_decay, double bias_correction1, double bias_correction2_sqrt, double ema_decay,
                                      fitb200_stream_t stream) {
  FITB200_CHECK_ARG(tensors && chunks && n_chunks >= 0, "adamw_ema_step: null table");
  FITB200_CHECK_ARG(bias_correction1 > 0. && bias_correction2_sqrt > 0., "adamw_ema_step: bias corrections must be positive (step >= 1)");
  if (n_chunks == 0) return FITB200_OK;
  const AdamScalars sc = adam_scalars(lr, beta1, beta2, eps, weight_decay, bias_correction1, bias_correction2_sqrt, ema_decay);
  adamw_ema_kernel<false><<<n_chunks, 256, 0, (cudaStream_t)stream>>>(tensors, chunks, sc, nullptr, nullptr, 1);
  FITB200_CHECK_LAUNCH("adamw_ema_step");
  return FITB200_OK;
}

extern "C" int fitb200_adamw_ema_step_dev(const fitb200_adam_tensor* tensors, const int32_t* chunks, int n_chunks, const double* hyper,
                                          const int64_t* step, int every_n_steps, fitb200_stream_t stream) {
  FITB200_CHECK_ARG(tensors && chunks && n_chunks >= 0 && hyper && step && every_n_steps >= 1, "adamw_ema_step_dev: bad arguments");
  if (n_chunks == 0) return FITB200_OK;
  adamw_ema_kernel<true><<<n_chunks, 256, 0, (cudaStream_t)stream>>>(tensors, chunks, AdamScalars{}, hyper, step, every_n_steps);
  FITB200_CHECK_LAUNCH("adamw_ema_step_dev");
  return FITB200_OK;
}
