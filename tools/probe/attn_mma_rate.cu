// Probe: execution rate of the tcgen05.mma shapes the attention kernels issue (attn_tc2.cu / attn_tc3.cu), one CTA per SM on every SM:
//   S  = Q K^T   SS, M = 128, N = 128, K = 16 per instruction, both operands K-major NO-swizzle core-matrix layout (LBO = 2048, SBO = 128)
//   PV = P V     TS, M = 128, N = 80,  K = 16, A (= P, bf16 pairs) from TMEM, B (= V) MN-major no-swizzle (LBO = 128, SBO = 2048)
// next to the 128B-swizzled K-major SS form the GEMMs use.  Operand values do not matter for timing.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o attn_mma_rate attn_mma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../fit_b200/csrc/ptx.cuh"
#include "../../fit_b200/csrc/umma_nosw.cuh"
using namespace fitb200;

constexpr int CS = 2048, BLOCK = 10 * CS;

__global__ void __launch_bounds__(128, 1) rate_kernel(int mode, int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  uint8_t* base = smem + ((1024u - (ptx::smem_u32(smem) & 1023u)) & 1023u);
  for (int i = threadIdx.x; i < (4 * BLOCK) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(base)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  if (threadIdx.x == 0) {
    const uint32_t tb = slot;
    uint32_t phase = 0;
    long long t0 = 0;
    for (int pass = 0; pass < 2; ++pass) {   // pass 0 = warm-up
      t0 = clock64();
      for (int i = 0; i < reps; ++i) {
        const uint32_t tr = tb + (i & 1) * 256;
        if (mode == 0) {          // S, no swizzle
          const uint32_t idesc = ptx::idesc_bf16_f32(128, 128);
          for (int k = 0; k < 5; ++k)
            ptx::mma_f16_ss(tr, desc_nosw(ptx::smem_u32(base + 2 * k * CS), CS, 128), desc_nosw(ptx::smem_u32(base + BLOCK + 2 * k * CS), CS, 128), idesc, k != 0);
        } else if (mode == 1) {   // PV, A from TMEM, B MN-major no swizzle
          const uint32_t idesc = ptx::idesc_bf16_f32(128, 80) | (1u << 16);
          for (int k = 0; k < 8; ++k)
            mma_f16_ts(tr + 128, tr + 8 * k, desc_nosw(ptx::smem_u32(base + 2 * BLOCK + k * 256), 128, CS), idesc, k != 0);
        } else if (mode == 2) {   // SS, 128B swizzle K-major, N = 128 (4 k-steps of one 64-wide atom + 1)
          const uint32_t idesc = ptx::idesc_bf16_f32(128, 128);
          const uint64_t da = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(base)), db = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(base + 16384));
          for (int k = 0; k < 5; ++k) ptx::mma_f16_ss(tr, da + 2 * (k & 3), db + 2 * (k & 3), idesc, k != 0);
        } else {                  // PV-shaped but SS with 128B-swizzled K-major B (what a swizzled V^T would cost), N = 80
          const uint32_t idesc = ptx::idesc_bf16_f32(128, 80);
          const uint64_t da = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(base)), db = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(base + 16384));
          for (int k = 0; k < 8; ++k) ptx::mma_f16_ss(tr + 128, da + 2 * (k & 3), db + 2 * (k & 3), idesc, k != 0);
        }
      }
      ptx::mma_commit(&bar);
      ptx::mbar_wait(&bar, phase);
      phase ^= 1;
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(slot, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  const int smem = 4 * BLOCK + 1024;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int reps = 2048;
  const char* names[4] = {"S  SS no-swizzle K-major  M128 N128 K16 x5 ", "PV TS A=TMEM, B MN-major no-swizzle N80 K16 x8", "S  SS 128B-swizzle K-major M128 N128 K16 x5 ", "PV-shaped SS 128B-swizzle N80 K16 x8        "};
  const int per[4] = {5, 8, 5, 8};
  const double ideal[4] = {64, 40, 64, 40};
  for (int mode = 0; mode < 4; ++mode) {
    rate_kernel<<<sms, 128, smem>>>(mode, reps, d);
    long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
    cudaError_t e = cudaGetLastError();
    printf("%s: %7.1f clk per group, %6.1f clk per MMA (ideal %.0f)  (%s)\n", names[mode], (double)c / reps, (double)c / reps / per[mode], ideal[mode], cudaGetErrorString(e));
  }
  return 0;
}
