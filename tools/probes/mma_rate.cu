// Microbenchmark: issue rate of tcgen05.mma.kind::f16 (M = 128) from one thread per SM, operands resident in shared
// memory (no TMA traffic).  Variants: N = 64 / 128 / 256, one accumulator vs alternating accumulators, A re-used vs
// walking over a 64 KB block, commit every 4 / 16 MMAs.   nvcc -gencode arch=compute_100a,code=sm_100a -I../../onepose_st_b200/csrc
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace opb;

template <int N, int NACC, int COMMIT_EVERY, bool WALK>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(long long* out, int iters) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar, bar2;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  // fill operands with small finite values (fp16 1.0 = 0x3c00)
  for (int i = threadIdx.x; i < (192 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::mbar_init(&bar2, 1); ptx::fence_barrier_init(); }
  if (warp == 1) { ptx::tmem_alloc(&tmem_slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (warp == 0 && ptx::elect_one()) {
    constexpr uint32_t idesc = ptx::make_idesc_f16(128, N);
    const uint64_t a0 = ptx::make_desc_sw128(base);               // A: 64 KB block [4 kb][128 x 64]
    const uint64_t b0 = ptx::make_desc_sw128(base + 65536);       // B: up to 8 stages of [N x 64]
    uint32_t phase = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int kb = j >> 2, k = j & 3;
        const uint64_t ad = a0 + (WALK ? kb * 1024 + 2 * k : 2 * k);
        const uint64_t bd = b0 + (WALK ? (((it & 1) * 4 + kb) % (N > 128 ? 4 : 8)) * (N * 128 / 16) + 2 * k : 2 * k);
        const uint32_t d = tmem_base + ((NACC > 1 ? (j % NACC) : 0) * N) % 512;
        ptx::mma_bf16(d, ad, bd, idesc, 1);
        if ((j + 1) % COMMIT_EVERY == 0 && j != 15) ptx::mma_commit(&bar2);  // a barrier nobody waits on (as the real kernel's B-stage commits)
      }
      ptx::mma_commit(&bar);
      ptx::mbar_wait(&bar, phase);  // the tile is complete before the next one is issued
      phase ^= 1;
    }
    const long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem_base, 512); }
}

// Same without waiting for completion inside the loop: the thread only issues (commit at the very end)
template <int N, int NACC, int COMMIT_EVERY>
__global__ void __launch_bounds__(128, 1) mma_issue_kernel(long long* out, int iters) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar, bar2;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (192 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::mbar_init(&bar2, 1); ptx::fence_barrier_init(); }
  if (warp == 1) { ptx::tmem_alloc(&tmem_slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (warp == 0 && ptx::elect_one()) {
    constexpr uint32_t idesc = ptx::make_idesc_f16(128, N);
    const uint64_t a0 = ptx::make_desc_sw128(base);
    const uint64_t b0 = ptx::make_desc_sw128(base + 65536);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int kb = j >> 2, k = j & 3;
        const uint32_t d = tmem_base + ((NACC > 1 ? ((it * 16 + j) % NACC) : 0) * N) % 512;
        ptx::mma_bf16(d, a0 + kb * 1024 + 2 * k, b0 + (((it & 1) * 4 + kb) % (N > 128 ? 4 : 8)) * (N * 128 / 16) + 2 * k, idesc, 1);
        if (COMMIT_EVERY > 0 && (j + 1) % COMMIT_EVERY == 0) ptx::mma_commit(&bar2);  // nobody waits on it
      }
    }
    ptx::mma_commit(&bar);
    ptx::mbar_wait(&bar, 0);
    const long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem_base, 512); }
}

template <typename K>
static void run(const char* name, K kern, int n_mma_per_iter_N, int iters) {
  long long* d;
  cudaMalloc(&d, 148 * 8);
  const int smem = 200 * 1024;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    kern<<<148, 128, smem>>>(d, iters);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(err)); return; }
  }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const double clk = double(h[0]) / (double(iters) * 16);
  const double flop = 148.0 * iters * 16 * 2.0 * 128 * n_mma_per_iter_N * 16;
  printf("%-44s %8.1f clk/MMA  (ideal %3d)  %7.1f us  %7.1f TFLOP/s  -> %.0f MHz\n", name, clk, n_mma_per_iter_N / 2, ms * 1e3, flop / ms / 1e9,
         double(h[0]) / (ms * 1e3));
  cudaFree(d);
}

int main() {
  const int iters = 2000;
  run("issue-only N=128 1 acc no commits", mma_issue_kernel<128, 1, 0>, 128, iters);
  run("issue-only N=128 1 acc commit/16", mma_issue_kernel<128, 1, 16>, 128, iters);
  run("issue-only N=128 1 acc commit/4", mma_issue_kernel<128, 1, 4>, 128, iters);
  run("issue-only N=128 1 acc commit/1", mma_issue_kernel<128, 1, 1>, 128, iters);
  run("issue-only N=128 4 acc commit/4", mma_issue_kernel<128, 4, 4>, 128, iters);
  run("issue-only N=256 1 acc no commits", mma_issue_kernel<256, 1, 0>, 256, iters);
  run("issue-only N=256 1 acc commit/4", mma_issue_kernel<256, 1, 4>, 256, iters);
  run("issue-only N=64 1 acc no commits", mma_issue_kernel<64, 1, 0>, 64, iters);
  run("tile-wait N=128 1acc commit/16 fixed", mma_rate_kernel<128, 1, 16, false>, 128, iters);
  run("tile-wait N=128 1acc commit/16 walk", mma_rate_kernel<128, 1, 16, true>, 128, iters);
  run("tile-wait N=128 1acc commit/4 walk", mma_rate_kernel<128, 1, 4, true>, 128, iters);
  run("tile-wait N=256 1acc commit/16 walk", mma_rate_kernel<256, 1, 16, true>, 256, iters);
  return 0;
}
