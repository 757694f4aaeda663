// Microbenchmark 2: the MMA-issuer loop of sim_fast_kernel in isolation (no data movement): per k-block wait on a
// "full" barrier re-armed by a partner thread once the stage's MMAs have completed (8-stage ring), four 128x128x16
// MMAs, commit to the stage's "empty" barrier; per tile a commit to the accumulator's full barrier and (optionally) a
// wait on its empty barrier, which FLAGS & 2 lets 4 x 4 partner warps release right away.
//   FLAGS: 1 = tcgen05.fence::after_thread_sync after every wait, 2 = accumulator handshake with epilogue warps,
//          4 = whole-warp loop with one issuing lane (else the loop runs inside an elected-thread branch)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace opb;

template <int FLAGS>
__global__ void __launch_bounds__(576, 1) pipe_kernel(long long* out, int tiles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t b_full[8], b_empty[8], tfull[4], tempty[4];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (192 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 8; ++s) { ptx::mbar_init(&b_full[s], 1); ptx::mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < 4; ++s) { ptx::mbar_init(&tfull[s], 1); ptx::mbar_init(&tempty[s], 4); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) { ptx::tmem_alloc(&tmem_slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (warp == 0) {  // partner of the B ring: a stage is "full" again as soon as its MMAs have completed
    if (ptx::elect_one()) {
      int stage = 0; uint32_t phase = 0;
      for (int t = 0; t < tiles * 4; ++t) {
        ptx::mbar_wait(&b_empty[stage], phase ^ 1);
        ptx::mbar_arrive(&b_full[stage]);
        if (++stage == 8) stage = 0, phase ^= 1;
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = ptx::make_idesc_f16(128, 128);
    const uint64_t a0 = ptx::make_desc_sw128(base);
    const uint64_t b0 = ptx::make_desc_sw128(base + 65536);
    const bool whole = (FLAGS & 4) != 0;
    const bool leader = ptx::elect_one();
    if (whole || leader) {
      const long long t0 = clock64();
      for (int it0 = 0; it0 < tiles; it0 += 4) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (FLAGS & 2) { ptx::mbar_wait(&tempty[u], ((it0 >> 2) & 1) ^ 1); if (FLAGS & 1) ptx::tc_fence_after(); }
#pragma unroll
          for (int kb = 0; kb < 4; ++kb) {
            const int stage = (u & 1) * 4 + kb;
            ptx::mbar_wait(&b_full[stage], static_cast<uint32_t>(u >> 1));
            if (FLAGS & 1) ptx::tc_fence_after();
            if (leader) {
#pragma unroll
              for (int k = 0; k < 4; ++k) ptx::mma_bf16(tmem_base + u * 128, a0 + kb * 1024 + 2 * k, b0 + stage * 1024 + 2 * k, idesc, (kb | k) != 0);
              ptx::mma_commit(&b_empty[stage]);
            }
          }
          if (leader) ptx::mma_commit(&tfull[u]);
        }
      }
      if (leader) {
        const long long t1 = clock64();
        out[blockIdx.x] = t1 - t0;
      }
    }
  } else if (FLAGS & 2) {  // "epilogue" warps: group g owns accumulator g and hands it back as soon as it is full
    const int g = (warp - 2) >> 2;
    for (int it = g, round = 0; it < tiles; it += 4, ++round) {
      ptx::mbar_wait(&tfull[g], round & 1);
      ptx::tc_fence_after();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty[g]);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem_base, 512); }
}

template <typename K>
static void run(const char* name, K kern, int tiles) {
  long long* d;
  cudaMalloc(&d, 148 * 8);
  const int smem = 200 * 1024;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    kern<<<148, 576, smem>>>(d, tiles);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(err)); return; }
  }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-56s %8.1f clk/tile (ideal 1024)  %7.1f us -> %.0f MHz\n", name, double(h[0]) / tiles, ms * 1e3, double(h[0]) / (ms * 1e3));
  cudaFree(d);
}

int main() {
  const int tiles = 4000;
  run("elected thread, ring only", pipe_kernel<0>, tiles);
  run("elected thread, ring + fences", pipe_kernel<1>, tiles);
  run("elected thread, ring + accumulator handshake", pipe_kernel<2>, tiles);
  run("elected thread, ring + handshake + fences", pipe_kernel<3>, tiles);
  run("whole warp, ring only", pipe_kernel<4>, tiles);
  run("whole warp, ring + handshake + fences", pipe_kernel<7>, tiles);
  return 0;
}
