// Standalone elu+1 linear attention operator (any N, L, S, H, D <= 64) on CUDA cores.
//
// Replaces LinearAttention.forward when it is called on its own:
//   src/models/OnePosePlus/loftr_module/linear_attention.py:29-61
//   submodules/LoFTR/src/loftr/loftr_module/linear_attention.py:20-47
// Inside the encoder stack the same math is fused into the tcgen05 kernels of encoder.cu; this entry
// point exists so that the operator API of the reference is complete.  HBM-bound (8 flop/B).
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

constexpr int LA_CHUNK = 128;

__device__ __forceinline__ float la_elu1(float x) { return x > 0.f ? __fadd_rn(x, 1.0f) : __fadd_rn(expm1f(x), 1.0f); }

// grid (chunks, H, N), block 256.  partial[n][h][chunk][D*D + D] = sum over the chunk's source rows of
// phi(K)[d] * (V[v] / S) and of phi(K)[d]
__global__ void __launch_bounds__(256) la_kv_kernel(const float* __restrict__ k, const float* __restrict__ v,
                                                    const uint8_t* __restrict__ kv_mask, int S, int H, int D, int chunks,
                                                    float* __restrict__ partial) {
  __shared__ float ks[64], vs[64];
  const int chunk = blockIdx.x, h = blockIdx.y, n = blockIdx.z, t = threadIdx.x;
  const int s0 = chunk * LA_CHUNK, s1 = min(S, s0 + LA_CHUNK);
  const int nout = D * D + D;
  float acc[17];  // ceil((64*64+64)/256) = 17
#pragma unroll
  for (int i = 0; i < 17; ++i) acc[i] = 0.f;
  const float inv_s = static_cast<float>(S);
  for (int s = s0; s < s1; ++s) {
    __syncthreads();
    if (t < 2 * D) {
      const float m = (kv_mask == nullptr || kv_mask[static_cast<long long>(n) * S + s]) ? 1.f : 0.f;
      const long long off = ((static_cast<long long>(n) * S + s) * H + h) * D;
      if (t < D) ks[t] = la_elu1(k[off + t]) * m;
      else vs[t - D] = (v[off + t - D] * m) / inv_s;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 17; ++i) {
      const int o = t + 256 * i;
      if (o < D * D) acc[i] = fmaf(ks[o / D], vs[o % D], acc[i]);
      else if (o < nout) acc[i] += ks[o - D * D];
    }
  }
  float* out = partial + ((static_cast<long long>(n) * H + h) * chunks + chunk) * nout;
#pragma unroll
  for (int i = 0; i < 17; ++i) {
    const int o = t + 256 * i;
    if (o < nout) out[o] = acc[i];
  }
}

// grid (H, N), block 256: fixed-order sum over chunks -> kv[n][h][D*D + D]
__global__ void __launch_bounds__(256) la_reduce_kernel(const float* __restrict__ partial, int D, int chunks,
                                                        float* __restrict__ kv) {
  const int nout = D * D + D;
  const long long base = static_cast<long long>(blockIdx.y) * gridDim.x + blockIdx.x;
  for (int o = threadIdx.x; o < nout; o += 256) {
    float s = 0.f;
    for (int c = 0; c < chunks; ++c) s += partial[(base * chunks + c) * nout + o];
    kv[base * nout + o] = s;
  }
}

// grid (ceil(L/128), H, N), block 128: one query row per thread
__global__ void __launch_bounds__(128) la_apply_kernel(const float* __restrict__ q, const uint8_t* __restrict__ q_mask,
                                                       const float* __restrict__ kv, int L, int S, int H, int D, float eps,
                                                       float* __restrict__ out) {
  extern __shared__ float sm[];  // D*D + D
  const int h = blockIdx.y, n = blockIdx.z;
  const int nout = D * D + D;
  const float* src = kv + (static_cast<long long>(n) * H + h) * nout;
  for (int o = threadIdx.x; o < nout; o += blockDim.x) sm[o] = src[o];
  __syncthreads();
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= L) return;
  const long long off = ((static_cast<long long>(n) * L + l) * H + h) * D;
  const float m = (q_mask == nullptr || q_mask[static_cast<long long>(n) * L + l]) ? 1.f : 0.f;
  float Q[64];
  float den = 0.f;
  for (int d = 0; d < D; ++d) {
    Q[d] = la_elu1(q[off + d]) * m;
    den = fmaf(Q[d], sm[D * D + d], den);
  }
  const float z = 1.0f / (den + eps);
  for (int vv = 0; vv < D; ++vv) {
    float a = 0.f;
    for (int d = 0; d < D; ++d) a = fmaf(Q[d], sm[d * D + vv], a);
    out[off + vv] = a * z * static_cast<float>(S);
  }
}

}  // namespace opb

using namespace opb;

extern "C" {

size_t opb200_linear_attention_workspace_bytes(int n_batch, int S, int H, int D) {
  const size_t chunks = (S + LA_CHUNK - 1) / LA_CHUNK;
  return static_cast<size_t>(n_batch) * H * (chunks + 1) * (D * D + D) * 4;
}

int opb200_linear_attention(const float* q, const float* k, const float* v, const uint8_t* q_mask,
                            const uint8_t* kv_mask, int n_batch, int L, int S, int H, int D, float eps, float* out,
                            void* workspace, size_t workspace_bytes, void* stream_) {
  if (int r = check_device()) return r;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  OPB_CHECK_ARG(q && k && v && out && workspace, "null pointer");
  OPB_CHECK_ARG(D >= 1 && D <= 64, "head dim must be in [1,64]");
  OPB_CHECK_ARG(n_batch >= 1 && L >= 1 && S >= 1 && H >= 1, "empty input");
  if (opb200_linear_attention_workspace_bytes(n_batch, S, H, D) > workspace_bytes) {
    set_last_error("linear_attention: workspace too small");
    return OPB_ERR_WORKSPACE;
  }
  const int chunks = (S + LA_CHUNK - 1) / LA_CHUNK;
  float* partial = static_cast<float*>(workspace);
  float* kv = partial + static_cast<size_t>(n_batch) * H * chunks * (D * D + D);
  la_kv_kernel<<<dim3(chunks, H, n_batch), 256, 0, stream>>>(k, v, kv_mask, S, H, D, chunks, partial);
  la_reduce_kernel<<<dim3(H, n_batch), 256, 0, stream>>>(partial, D, chunks, kv);
  la_apply_kernel<<<dim3((L + 127) / 128, H, n_batch), 128, (D * D + D) * 4, stream>>>(q, q_mask, kv, L, S, H, D, eps, out);
  count_launch(3);
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

}  // extern "C"
