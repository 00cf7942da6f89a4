// Probe: issue rate of tcgen05.mma kind::f16 (bf16 in, fp32 accumulate) as a function of N, for M = 128 (cta_group::1), one CTA per SM on
// every SM.  Operands: garbage shared memory through valid 128B-swizzled K-major descriptors (values do not matter for timing).
// Prints cycles per MMA and the implied flop/clk/SM.   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_rate umma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../fit_b200/csrc/ptx.cuh"
using namespace fitb200;

__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  uint8_t* base = smem + ((1024u - (ptx::smem_u32(smem) & 1023u)) & 1023u);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(base)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  if (threadIdx.x == 0) {
    const uint32_t tb = slot;
    const uint64_t da = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(base)), db = ptx::smem_desc_kmajor_sw128(ptx::smem_u32(base + 16384));
    const uint32_t idesc = ptx::idesc_bf16_f32(128, N);
    // warm-up
    for (int i = 0; i < 64; ++i) ptx::mma_f16_ss(tb, da + 2 * (i & 3), db + 2 * (i & 3), idesc, 1);
    ptx::mma_commit(&bar);
    ptx::mbar_wait(&bar, 0);
    const long long t0 = clock64();
    for (int i = 0; i < reps; ++i) ptx::mma_f16_ss(tb + (i & 1) * 256, da + 2 * (i & 3), db + 2 * (i & 3), idesc, 1);
    ptx::mma_commit(&bar);
    ptx::mbar_wait(&bar, 1);
    const long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(slot, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 + 32768 + 1024);
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int reps = 4096;
  for (int N : {32, 64, 96, 128, 160, 192, 224, 256}) {
    rate_kernel<<<sms, 128, 16384 + 32768 + 1024>>>(N, reps, d);
    long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
    cudaError_t e = cudaGetLastError();
    printf("M=128 N=%3d K=16: %7.1f clk per MMA -> %6.0f flop/clk/SM  (%s)\n", N, (double)c / reps, 2.0 * 128 * N * 16 * reps / (double)c, cudaGetErrorString(e));
  }
  return 0;
}
