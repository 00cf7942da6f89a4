// Probe 2: MN-major SWIZZLE_128B operands for both A and B (the wgrad / dgrad GEMMs):  D[128 x 128] = A^T-stored . B^T-stored
// A is stored [K][M] (M contiguous), B is stored [K][N] (N contiguous), tiles as TMA would write them with 128B swizzle:
// block of 64 MN elements x BK k-rows, row kk at kk*128 bytes, 16-byte chunk c stored at chunk c ^ (kk & 7).
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_bf16.h>
#include "ptx.cuh"
using namespace fitb200;
constexpr int TM = 128, TN = 128, TK = 64;

__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__global__ void __launch_bounds__(128, 1) probe(const __nv_bfloat16* A, const __nv_bfloat16* B, float* Dout, int swap, int kadv) {
  extern __shared__ __align__(1024) uint8_t sm_raw[];
  uint8_t* sm = sm_raw + ((1024u - (ptx::smem_u32(sm_raw) & 1023u)) & 1023u);
  uint8_t* sA = sm; uint8_t* sB = sm + TM * TK * 2;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + TN * TK * 2);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < TK * TM / 8; i += 128) { int k = i / (TM / 8), c = i % (TM / 8); int blk = c / 8, cc = c % 8;
    *reinterpret_cast<uint4*>(sA + blk * (TK * 128) + k * 128 + ((cc ^ (k & 7)) << 4)) = *reinterpret_cast<const uint4*>(A + k * TM + c * 8); }
  for (int i = tid; i < TK * TN / 8; i += 128) { int k = i / (TN / 8), c = i % (TN / 8); int blk = c / 8, cc = c % 8;
    *reinterpret_cast<uint4*>(sB + blk * (TK * 128) + k * 128 + ((cc ^ (k & 7)) << 4)) = *reinterpret_cast<const uint4*>(B + k * TN + c * 8); }
  if (tid == 0) { ptx::mbar_init(bar, 1); ptx::fence_mbar_init(); }
  if (warp == 0) { ptx::tmem_alloc(slot, 128); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tb = *slot;
  if (tid == 0) {
    const uint32_t idesc = ptx::idesc_bf16_f32(TM, TN) | (1u << 15) | (1u << 16);
    for (int k = 0; k < TK / 16; ++k) {
      uint32_t lbo = TK * 128, sbo = 1024;
      if (swap) { uint32_t t = lbo; lbo = sbo; sbo = t; }
      uint64_t da = desc_sw128(ptx::smem_u32(sA + k * kadv), lbo, sbo);
      uint64_t db = desc_sw128(ptx::smem_u32(sB + k * kadv), lbo, sbo);
      ptx::mma_f16_ss(tb, da, db, idesc, k != 0);
    }
    ptx::mma_commit(bar);
  }
  ptx::mbar_wait(bar, 0);
  ptx::tc_fence_after();
  const uint32_t lane_base = tb + ((uint32_t)(warp * 32) << 16);
  for (int c = 0; c < TN / 16; ++c) {
    uint32_t r[16];
    ptx::tmem_ld16(lane_base + c * 16, r); ptx::tmem_ld_wait();
    for (int i = 0; i < 16; ++i) Dout[tid * TN + c * 16 + i] = __uint_as_float(r[i]);
  }
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tb, 128); }
}
int main() {
  std::vector<__nv_bfloat16> a(TK * TM), b(TK * TN); std::vector<float> af(TK * TM), bf(TK * TN), ref(TM * TN);
  srand(2);
  auto rnd = [] { return (rand() % 2001 - 1000) / 1000.0f; };
  for (int i = 0; i < TK * TM; ++i) { a[i] = __float2bfloat16(rnd()); af[i] = __bfloat162float(a[i]); }
  for (int i = 0; i < TK * TN; ++i) { b[i] = __float2bfloat16(rnd()); bf[i] = __bfloat162float(b[i]); }
  for (int m = 0; m < TM; ++m) for (int n = 0; n < TN; ++n) { float s = 0; for (int k = 0; k < TK; ++k) s += af[k * TM + m] * bf[k * TN + n]; ref[m * TN + n] = s; }
  __nv_bfloat16 *da, *db; float* dd;
  cudaMalloc(&da, a.size() * 2); cudaMalloc(&db, b.size() * 2); cudaMalloc(&dd, ref.size() * 4);
  cudaMemcpy(da, a.data(), a.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(db, b.data(), b.size() * 2, cudaMemcpyHostToDevice);
  const int smem = TM * TK * 2 + TN * TK * 2 + 64 + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int sw = 0; sw < 2; ++sw) {
    cudaMemset(dd, 0, ref.size() * 4);
    probe<<<1, 128, smem>>>(da, db, dd, sw, 2048);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("swap=%d: CUDA error %s\n", sw, cudaGetErrorString(e)); return 1; }
    std::vector<float> d(ref.size());
    cudaMemcpy(d.data(), dd, d.size() * 4, cudaMemcpyDeviceToHost);
    double es = 0, ns = 0;
    for (size_t i = 0; i < d.size(); ++i) { es += (d[i] - ref[i]) * (d[i] - ref[i]); ns += ref[i] * ref[i]; }
    printf("MN-major SW128 swap=%d: rel err %.3e\n", sw, sqrt(es / ns));
  }
  return 0;
}
