// Probe: tcgen05.mma kind::f16 with the A operand in TENSOR MEMORY (written by tcgen05.st from registers) and B in shared
// memory (K-major, 64-byte swizzle).  Assumed layout of A: lane = row m, 32-bit column c of the operand holds elements
// (2c, 2c + 1) of that row, a K = 16 step advances the address by 8 columns.  D[m][n] = sum_k A[m][k] * B[n][k].
//   nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../onepose_st_b200/csrc -o tmem_a_probe.bin tmem_a_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace opb;

__device__ __forceinline__ void mma_bf16_ta(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

constexpr int K = 32;
__host__ __device__ inline float aval(int m, int k) { return static_cast<float>((m * 7 + k * 3) % 5 - 2); }
__host__ __device__ inline float bval(int n, int k) { return static_cast<float>((n * 5 + k) % 7 - 3); }

__global__ void __launch_bounds__(128, 1) probe(float* out, int n_cols) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sb = smem;  // [256 rows][64 B], 64-byte swizzle
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int e = tid; e < 256 * K; e += 128) {
    const int n = e / K, k = e % K;
    const uint32_t off = n * 64 + ((((k >> 3) ^ ((n >> 1) & 3))) << 4) + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sb + off) = __float2bfloat16(bval(n, k));
  }
  if (tid == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (warp == 1) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = slot;
  const uint32_t lane_field = static_cast<uint32_t>(warp * 32) << 16;
  {
    uint32_t r[16];
    for (int c = 0; c < 16; ++c) {
      const __nv_bfloat162 v = __floats2bfloat162_rn(aval(tid, 2 * c), aval(tid, 2 * c + 1));
      r[c] = *reinterpret_cast<const uint32_t*>(&v);
    }
    tmem_st16(tm + lane_field + 256, r);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  if (warp == 0 && ptx::elect_one()) {
    const uint32_t idesc = ptx::make_idesc_bf16(128, n_cols);
    const uint64_t bd = ptx::make_desc_sw64(ptx::smem_u32(sb));
    for (int k = 0; k < K / 16; ++k) mma_bf16_ta(tm, tm + 256 + 8 * k, bd + 2 * k, idesc, k != 0);
    ptx::mma_commit(&bar);
  }
  ptx::mbar_wait(&bar, 0);
  ptx::tc_fence_after();
  for (int c0 = 0; c0 < n_cols; c0 += 32) {
    uint32_t r[32];
    ptx::tmem_ld32(tm + lane_field + c0, r);
    ptx::tmem_ld_wait();
    for (int i = 0; i < 32; ++i) out[tid * 256 + c0 + i] = __uint_as_float(r[i]);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 512); }
}

int main() {
  float* d;
  cudaMalloc(&d, 128 * 256 * 4);
  const int smem = 20 * 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  static float h[128 * 256];
  for (int n_cols = 256; n_cols >= 128; n_cols -= 128) {
    cudaMemset(d, 0, sizeof(h));
    probe<<<1, 128, smem>>>(d, n_cols);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N=%d: %s\n", n_cols, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < n_cols; ++n) {
        float ref = 0.f;
        for (int k = 0; k < K; ++k) ref += aval(m, k) * bval(n, k);
        if (h[m * 256 + n] != ref) {
          if (bad < 6) printf("  N=%d: D[%d][%d] = %.0f, expected %.0f\n", n_cols, m, n, h[m * 256 + n], ref);
          ++bad;
        }
      }
    printf("A in TMEM, N=%d: %s (%d mismatches)\n", n_cols, bad ? "WRONG" : "correct", bad);
  }
  return 0;
}
