// Prints which (TMEM lane, column) each register of tcgen05.ld.16x256b.x4 receives: the tile is first written with
// tcgen05.st.32x32b (lane = thread, register = column; value = lane * 1000 + column), then read back in the
// fragment layout from lane offsets 0 and 16.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tmem_probe.bin
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
__global__ void probe(float* out) {
  __shared__ uint32_t slot;
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t taddr = base + ((uint32_t)(warp * 32) << 16);
  for (int c = 0; c < 32; ++c) {
    const uint32_t v = __float_as_uint((float)((warp * 32 + lane) * 1000 + c));
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr + c), "r"(v) : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  for (int half = 0; half < 2; ++half) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr + ((uint32_t)(half * 16) << 16))
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int q = 0; q < 16; ++q) out[((warp * 2 + half) * 32 + lane) * 16 + q] = __uint_as_float(r[q]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(base) : "memory");
}
int main() {
  float* d;
  cudaMalloc(&d, 4 * 2 * 32 * 16 * 4);
  probe<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
  static float h[4 * 2 * 32 * 16];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int w = 0; w < 4; ++w)
    for (int half = 0; half < 2; ++half)
      for (int t = 0; t < 32; ++t)
        for (int q = 0; q < 16; ++q) {
          const int k = q / 4, sub = q % 4;
          const int lane = w * 32 + half * 16 + t / 4 + 8 * (sub / 2), col = 8 * k + 2 * (t % 4) + (sub & 1);
          const float got = h[((w * 2 + half) * 32 + t) * 16 + q];
          if (got != (float)(lane * 1000 + col)) {
            if (bad < 20) printf("warp %d half %d thread %d reg %d: expected lane %d col %d, got %.0f\n", w, half, t, q, lane, col, got);
            ++bad;
          }
        }
  printf("tmem 16x256b.x4 layout: %s (%d mismatches)\n", bad ? "DIFFERENT" : "as assumed", bad);
  return 0;
}
