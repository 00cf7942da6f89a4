// Shared device/host helpers for the fit_b200 kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>

#include "../../include/fit_b200.h"

namespace fitb200 {

// ---------------------------------------------------------------------------------------------
// error plumbing: every C-ABI entry returns 0 or a negative code and records a message
// ---------------------------------------------------------------------------------------------
void set_last_error(const char* fmt, ...);

#define FITB200_CHECK_ARG(cond, ...)                         \
  do {                                                       \
    if (!(cond)) {                                           \
      ::fitb200::set_last_error(__VA_ARGS__);                \
      return FITB200_ERR_INVALID_ARG;                        \
    }                                                        \
  } while (0)

#define FITB200_CHECK_LAUNCH(name)                                                         \
  do {                                                                                     \
    cudaError_t e__ = cudaGetLastError();                                                  \
    if (e__ != cudaSuccess) {                                                              \
      ::fitb200::set_last_error("%s: launch failed: %s", name, cudaGetErrorString(e__));   \
      return FITB200_ERR_CUDA;                                                             \
    }                                                                                      \
  } while (0)

// SMs the persistent kernels (GEMM, attention) may occupy: all of them, or fewer while a collective runs beside the backward pass
// (fitb200_set_sm_limit): NCCL's copy / reduce CTAs need SMs of their own, and a persistent grid that does not fit in one wave next
// to them runs its last CTAs as a second wave (measured: 8 ms of a 42 ms XL/2 step at 2 GPUs).
int& sm_limit_ref();
inline int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  const int lim = sm_limit_ref();
  return lim > 0 && lim < n ? lim : n;
}

template <typename T>
__host__ __device__ constexpr T ceil_div(T a, T b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
// activation-dtype helpers (float or bf16 storage, fp32 math)
// ---------------------------------------------------------------------------------------------
template <typename T> struct Vec8;  // 8 consecutive elements

template <> struct Vec8<float> {
  float v[8];
  __device__ __forceinline__ static Vec8 load(const float* p) {
    Vec8 r;
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
  }
  __device__ __forceinline__ void store(float* p) const {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};

template <> struct Vec8<__nv_bfloat16> {
  float v[8];
  __device__ __forceinline__ static Vec8 load(const __nv_bfloat16* p) {
    Vec8 r;
    uint4 raw = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      r.v[2 * i] = f.x; r.v[2 * i + 1] = f.y;
    }
    return r;
  }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const {
    uint4 raw;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = raw;
  }
};

__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
template <typename T> __device__ __forceinline__ T from_f32(float x);
template <> __device__ __forceinline__ float from_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ float silu_f(float x) { return __fdividef(x, 1.0f + __expf(-x)); }
// accurate variant for the fp32 verification path
__device__ __forceinline__ float silu_acc(float x) { return x / (1.0f + expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// dispatch on the public dtype enum
#define FITB200_DISPATCH_ACT(dt, T, ...)                                    \
  if ((dt) == FITB200_F32) { using T = float; __VA_ARGS__ }                 \
  else if ((dt) == FITB200_BF16) { using T = __nv_bfloat16; __VA_ARGS__ }   \
  else { ::fitb200::set_last_error("unsupported dtype %d", (int)(dt)); return FITB200_ERR_INVALID_ARG; }

}  // namespace fitb200
