// Probe: tcgen05.mma kind::f16 with MN-major A and B operands (the contraction runs over the ROWS of row-major
// [tokens, channels] planes, as in phi(K)^T V).  Shared-memory image = what a TMA box {64 channels, 64 tokens} with
// SWIZZLE_128B produces: row t (one token) = 128 bytes = 64 channels, 16-byte chunk index XORed with t & 7; the second
// 64 channels of a 128-wide operand follow 8192 bytes later.  Tries both assignments of (LBO, SBO) in the descriptor
// and prints which one reproduces D[m][n] = sum_t A[t][m] * B[t][n].
//   nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../onepose_st_b200/csrc -o mn_major_probe.bin mn_major_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace opb;

__device__ __forceinline__ uint64_t make_desc_mn(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(lbo >> 4) << 16;
  d |= static_cast<uint64_t>(sbo >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;  // SWIZZLE_128B
  return d;
}

__global__ void __launch_bounds__(128, 1) probe(float* out, int variant, int n_cols) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sa = smem;            // [2 boxes][64 tok][128 B]
  uint8_t* sb = smem + 16384;
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int e = tid; e < 64 * 128; e += 128) {
    const int t = e / 128, c = e % 128;
    const int box = c >> 6, cc = c & 63;
    const uint32_t off = box * 8192 + t * 128 + ((((cc >> 3) ^ (t & 7))) << 4) + (cc & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sa + off) = __float2bfloat16(static_cast<float>((t * 7 + c * 3) % 5 - 2));
    *reinterpret_cast<__nv_bfloat16*>(sb + off) = __float2bfloat16(static_cast<float>((t * 5 + c) % 7 - 3));
  }
  if (tid == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (warp == 1) { ptx::tmem_alloc(&slot, 256); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 0 && ptx::elect_one()) {
    const uint32_t idesc = ptx::make_idesc_bf16(128, n_cols) | (1u << 15) | (1u << 16);  // A and B MN-major
    const uint32_t lbo = variant == 0 ? 8192 : 1024, sbo = variant == 0 ? 1024 : 8192;
    const uint64_t ad = make_desc_mn(ptx::smem_u32(sa), lbo, sbo), bd = make_desc_mn(ptx::smem_u32(sb), lbo, sbo);
    for (int k = 0; k < 4; ++k) ptx::mma_bf16(tm, ad + k * 128, bd + k * 128, idesc, k != 0);  // 16 tokens = 2048 bytes
    ptx::mma_commit(&bar);
  }
  ptx::mbar_wait(&bar, 0);
  ptx::tc_fence_after();
  for (int c0 = 0; c0 < n_cols; c0 += 32) {
    uint32_t r[32];
    ptx::tmem_ld32(tm + (static_cast<uint32_t>(warp * 32) << 16) + c0, r);
    ptx::tmem_ld_wait();
    for (int i = 0; i < 32; ++i) out[(warp * 32 + (tid & 31)) * 128 + c0 + i] = __uint_as_float(r[i]);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 256); }
}

int main() {
  float* d;
  cudaMalloc(&d, 128 * 128 * 4);
  const int smem = 34 * 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  static float h[128 * 128];
  for (int n_cols = 128; n_cols >= 64; n_cols -= 64)
    for (int variant = 0; variant < 2; ++variant) {
      cudaMemset(d, 0, sizeof(h));
      probe<<<1, 128, smem>>>(d, variant, n_cols);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("variant %d: %s\n", variant, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < n_cols; ++n) {
          float ref = 0.f;
          for (int t = 0; t < 64; ++t) ref += static_cast<float>((t * 7 + m * 3) % 5 - 2) * static_cast<float>((t * 5 + n) % 7 - 3);
          if (h[m * 128 + n] != ref) {
            if (bad < 4) printf("  N=%d variant %d: D[%d][%d] = %.0f, expected %.0f\n", n_cols, variant, m, n, h[m * 128 + n], ref);
            ++bad;
          }
        }
      printf("N=%d, LBO/SBO = %s: %s (%d mismatches)\n", n_cols, variant == 0 ? "8192/1024" : "1024/8192", bad ? "WRONG" : "correct", bad);
    }
  return 0;
}
