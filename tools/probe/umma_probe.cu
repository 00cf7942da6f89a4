// Probe of the tcgen05 operand conventions the attention kernel relies on (run on a B200):
//   (1) no-swizzle K-major smem descriptors (LBO / SBO roles)     S = Q K^T
//   (2) no-swizzle MN-major smem descriptor for V                 O = P V
//   (3) A operand from TMEM (bf16 pairs packed per 32-bit column) and an accumulator at an arbitrary column offset
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I fit_b200/csrc -o /tmp/umma_probe tools/probe/umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_bf16.h>
#include "ptx.cuh"
using namespace fitb200;

constexpr int NQ = 128, NK = 256, DP = 80, DV = 80;  // DP: padded contraction width, DV: N of the PV product

__device__ __forceinline__ uint64_t desc_nosw(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d), "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(taddr),
               "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
               "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

// smem layout of a [rows][chunks of 8 bf16] operand: offset(row, chunk) = chunk * rows * 16 + row * 16
__global__ void __launch_bounds__(128, 1) probe(const __nv_bfloat16* Q, const __nv_bfloat16* K, const __nv_bfloat16* V, float* S, float* O,
                                                int swapK, int swapV, int o_col) {
  extern __shared__ __align__(1024) uint8_t sm[];
  uint8_t* sQ = sm; uint8_t* sK = sQ + NQ * DP * 2; uint8_t* sV = sK + NK * DP * 2;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sV + NK * DV * 2);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < NQ * DP / 8; i += 128) { int row = i / (DP / 8), c = i % (DP / 8);
    *reinterpret_cast<uint4*>(sQ + c * NQ * 16 + row * 16) = *reinterpret_cast<const uint4*>(Q + row * DP + c * 8); }
  for (int i = tid; i < NK * DP / 8; i += 128) { int row = i / (DP / 8), c = i % (DP / 8);
    *reinterpret_cast<uint4*>(sK + c * NK * 16 + row * 16) = *reinterpret_cast<const uint4*>(K + row * DP + c * 8); }
  for (int i = tid; i < NK * DV / 8; i += 128) { int row = i / (DV / 8), c = i % (DV / 8);
    *reinterpret_cast<uint4*>(sV + c * NK * 16 + row * 16) = *reinterpret_cast<const uint4*>(V + row * DV + c * 8); }
  if (tid == 0) { ptx::mbar_init(&bar[0], 1); ptx::mbar_init(&bar[1], 1); ptx::fence_mbar_init(); }
  if (warp == 0) { ptx::tmem_alloc(slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tb = *slot;
  if (tid == 0) {
    // S[128 x 256] = Q K^T : A, B K-major no-swizzle; one K step = 2 chunks
    const uint32_t idesc = ptx::idesc_bf16_f32(128, 256);
    for (int k = 0; k < DP / 16; ++k) {
      uint32_t la = NQ * 16, sa = 128, lb = NK * 16, sb = 128;
      if (swapK) { uint32_t t = la; la = sa; sa = t; t = lb; lb = sb; sb = t; }
      uint64_t da = desc_nosw(ptx::smem_u32(sQ + 2 * k * NQ * 16), la, sa);
      uint64_t db = desc_nosw(ptx::smem_u32(sK + 2 * k * NK * 16), lb, sb);
      ptx::mma_f16_ss(tb, da, db, idesc, k != 0);
    }
    ptx::mma_commit(&bar[0]);
  }
  ptx::mbar_wait(&bar[0], 0);
  ptx::tc_fence_after();
  const uint32_t lane_base = tb + ((uint32_t)(warp * 32) << 16);
  // read S, write it out, write P = bf16(S * 0.01) packed into columns [256, 384)
  for (int c = 0; c < NK / 16; ++c) {
    uint32_t r[16];
    ptx::tmem_ld16(lane_base + c * 16, r); ptx::tmem_ld_wait();
    for (int i = 0; i < 16; ++i) S[tid * NK + c * 16 + i] = __uint_as_float(r[i]);
    if (c & 1) continue;
    uint32_t r2[16];
    ptx::tmem_ld16(lane_base + c * 16 + 16, r2); ptx::tmem_ld_wait();
    uint32_t pk[16];
    for (int i = 0; i < 8; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(r[2 * i]) * 0.01f, __uint_as_float(r[2 * i + 1]) * 0.01f);
      pk[i] = *reinterpret_cast<uint32_t*>(&h);
      __nv_bfloat162 g = __floats2bfloat162_rn(__uint_as_float(r2[2 * i]) * 0.01f, __uint_as_float(r2[2 * i + 1]) * 0.01f);
      pk[8 + i] = *reinterpret_cast<uint32_t*>(&g);
    }
    tmem_st16(lane_base + 256 + (c / 2) * 16, pk);
  }
  tmem_st_wait();
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  if (tid == 0) {
    // O[128 x 80] = P[128 x 256] V[256 x 80]: A from TMEM, B = V MN-major no-swizzle; one K step = 16 keys = 2 key groups
    const uint32_t idesc = ptx::idesc_bf16_f32(128, DV) | (1u << 16);
    for (int k = 0; k < NK / 16; ++k) {
      uint32_t lb = 128, sb = NK * 16;   // LBO: next 8 keys (K direction), SBO: next 8 columns of d (N direction)
      if (swapV) { uint32_t t = lb; lb = sb; sb = t; }
      uint64_t db = desc_nosw(ptx::smem_u32(sV + k * 256), lb, sb);
      mma_ts(tb + o_col, tb + 256 + k * 8, db, idesc, k != 0);
    }
    ptx::mma_commit(&bar[1]);
  }
  ptx::mbar_wait(&bar[1], 0);
  ptx::tc_fence_after();
  for (int c = 0; c < DV / 16; ++c) {
    uint32_t r[16];
    ptx::tmem_ld16(lane_base + o_col + c * 16, r); ptx::tmem_ld_wait();
    for (int i = 0; i < 16; ++i) O[tid * DV + c * 16 + i] = __uint_as_float(r[i]);
  }
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tb, 512); }
}

int main() {
  std::vector<__nv_bfloat16> q(NQ * DP), k(NK * DP), v(NK * DV);
  std::vector<float> qf(NQ * DP), kf(NK * DP), vf(NK * DV);
  srand(1);
  auto rnd = [] { return (rand() % 2001 - 1000) / 1000.0f; };
  for (int i = 0; i < NQ * DP; ++i) { q[i] = __float2bfloat16(rnd()); qf[i] = __bfloat162float(q[i]); }
  for (int i = 0; i < NK * DP; ++i) { k[i] = __float2bfloat16(rnd()); kf[i] = __bfloat162float(k[i]); }
  for (int i = 0; i < NK * DV; ++i) { v[i] = __float2bfloat16(rnd()); vf[i] = __bfloat162float(v[i]); }
  std::vector<float> Sref(NQ * NK), Oref(NQ * DV);
  for (int i = 0; i < NQ; ++i) for (int j = 0; j < NK; ++j) { float a = 0; for (int c = 0; c < DP; ++c) a += qf[i * DP + c] * kf[j * DP + c]; Sref[i * NK + j] = a; }
  for (int i = 0; i < NQ; ++i) for (int n = 0; n < DV; ++n) { float a = 0; for (int j = 0; j < NK; ++j) a += __bfloat162float(__float2bfloat16(Sref[i * NK + j] * 0.01f)) * vf[j * DV + n]; Oref[i * DV + n] = a; }
  __nv_bfloat16 *dq, *dk, *dv; float *dS, *dO;
  cudaMalloc(&dq, q.size() * 2); cudaMalloc(&dk, k.size() * 2); cudaMalloc(&dv, v.size() * 2);
  cudaMalloc(&dS, Sref.size() * 4); cudaMalloc(&dO, Oref.size() * 4);
  cudaMemcpy(dq, q.data(), q.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(dk, k.data(), k.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dv, v.data(), v.size() * 2, cudaMemcpyHostToDevice);
  const int smem = NQ * DP * 2 + NK * DP * 2 + NK * DV * 2 + 64;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int ocols[2] = {384, 64 + 256 + 100 - 100 + 8};  // 384 (aligned) and 328 (odd multiple of 8)
  for (int sk = 0; sk < 2; ++sk) for (int sv = 0; sv < 2; ++sv) for (int oc = 0; oc < 2; ++oc) {
    cudaMemset(dS, 0, Sref.size() * 4); cudaMemset(dO, 0, Oref.size() * 4);
    probe<<<1, 128, smem>>>(dq, dk, dv, dS, dO, sk, sv, ocols[oc] == 384 ? 384 : 400);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("swapK=%d swapV=%d ocol=%d: CUDA error %s\n", sk, sv, oc, cudaGetErrorString(e)); return 1; }
    std::vector<float> S(Sref.size()), O(Oref.size());
    cudaMemcpy(S.data(), dS, S.size() * 4, cudaMemcpyDeviceToHost); cudaMemcpy(O.data(), dO, O.size() * 4, cudaMemcpyDeviceToHost);
    double es = 0, ns = 0, eo = 0, no = 0;
    for (size_t i = 0; i < S.size(); ++i) { es += (S[i] - Sref[i]) * (S[i] - Sref[i]); ns += Sref[i] * Sref[i]; }
    for (size_t i = 0; i < O.size(); ++i) { eo += (O[i] - Oref[i]) * (O[i] - Oref[i]); no += Oref[i] * Oref[i]; }
    printf("swapK=%d swapV=%d ocol=%d: S rel err %.3e   O rel err %.3e\n", sk, sv, ocols[oc] == 384 ? 384 : 400, sqrt(es / ns), sqrt(eo / no));
  }
  return 0;
}
