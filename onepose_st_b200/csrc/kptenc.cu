// 3D keypoint position encoding of the 2D-3D matcher, fused: normalisation of the keypoints + 4-layer MLP with
// per-point channel normalisation + addition to the coarse descriptors.
//
// Replaces  normalize_3d_keypoints        src/models/OnePosePlus/utils/normalize.py:17-28
//           KeypointEncoding_linear       src/models/OnePosePlus/utils/position_encoding.py:46-79   (norm_method
//           "instancenorm": nn.InstanceNorm1d(c) applied to [B, L, c], i.e. statistics over the CHANNELS of each point,
//           biased variance, eps 1e-5, no affine -- SURVEY.md Appendix B #3)
//   k   = (kpts - mean_L(kpts)) / (0.6 * max_xyz(extent of batch element 0))
//   out = desc + MLP(k)^T,  MLP = Linear(3,32) IN ReLU Linear(32,64) IN ReLU Linear(64,128) IN ReLU Linear(128,256)
// kpt_stats_kernel: per batch element the mean over the points (fixed-order tree: deterministic) and, for element 0,
// the extent.  kpt_encode_kernel: one warp per point, activations of the point in shared memory, lane o computes
// output channels o, o + 32, ... from weights stored input-major (coalesced across the lanes); 32 consecutive points
// per block so that the channels-first result leaves through a shared-memory transpose in 128-byte lines.
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

constexpr int KE_D0 = 3, KE_D1 = 32, KE_D2 = 64, KE_D3 = 128, KE_D4 = 256;
// packed parameter block (floats): W1t [3][32], b1 [32], W2t [32][64], b2 [64], W3t [64][128], b3 [128], W4t [128][256], b4 [256]
constexpr int KE_W1 = 0, KE_B1 = KE_W1 + KE_D0 * KE_D1, KE_W2 = KE_B1 + KE_D1, KE_B2 = KE_W2 + KE_D1 * KE_D2;
constexpr int KE_W3 = KE_B2 + KE_D2, KE_B3 = KE_W3 + KE_D2 * KE_D3, KE_W4 = KE_B3 + KE_D3, KE_B4 = KE_W4 + KE_D3 * KE_D4;
constexpr int KE_PACKED = KE_B4 + KE_D4;

// stats [N][4]: mean x, y, z of element n; stats[n][3] = 1 / (0.6 * max extent of element 0)
__global__ void __launch_bounds__(256) kpt_stats_kernel(const float* __restrict__ kpts, long long sn, long long sl, long long sc,
                                                        int L, int normalize, float* __restrict__ stats) {
  __shared__ float red[8][9];
  const int n = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (!normalize) {  // the caller's keypoints are normalised already (KeypointEncoding_linear.forward on its own)
    if (tid < 3) stats[4 * n + tid] = 0.f;
    if (tid == 3 && n == 0) stats[3] = 1.f;
    return;
  }
  float s[3] = {0.f, 0.f, 0.f}, lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int l = tid; l < L; l += 256) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = kpts[n * sn + l * sl + c * sc];
      s[c] += v;
      lo[c] = fminf(lo[c], v);
      hi[c] = fmaxf(hi[c], v);
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s[c] += __shfl_xor_sync(0xffffffffu, s[c], o);
      lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
  if (lane == 0)
#pragma unroll
    for (int c = 0; c < 3; ++c) red[warp][c] = s[c], red[warp][3 + c] = lo[c], red[warp][6 + c] = hi[c];
  __syncthreads();
  if (tid == 0) {
    float m[3] = {0.f, 0.f, 0.f}, l3[3] = {INFINITY, INFINITY, INFINITY}, h3[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int w = 0; w < 8; ++w)
      for (int c = 0; c < 3; ++c) {
        m[c] += red[w][c];
        l3[c] = fminf(l3[c], red[w][3 + c]);
        h3[c] = fmaxf(h3[c], red[w][6 + c]);
      }
    for (int c = 0; c < 3; ++c) stats[4 * n + c] = m[c] / static_cast<float>(L);
    if (n == 0) stats[3] = 1.f / (fmaxf(fmaxf(h3[0] - l3[0], h3[1] - l3[1]), h3[2] - l3[2]) * 0.6f);
  }
}

// x (DIN values of the warp's point, shared memory) -> y (DOUT values): y[o] = b[o] + sum_i Wt[i][o] x[i], then, unless
// LAST, (y - mean) / sqrt(var + 1e-5) over the DOUT channels and ReLU
template <int DIN, int DOUT, bool LAST>
__device__ __forceinline__ void ke_layer(const float* __restrict__ wt, const float* __restrict__ b, const float* x, float* y, int lane) {
  constexpr int PER = DOUT / 32;
  float acc[PER];
#pragma unroll
  for (int k = 0; k < PER; ++k) acc[k] = __ldg(b + lane + 32 * k);
#pragma unroll 4
  for (int i = 0; i < DIN; ++i) {
    const float xi = x[i];
#pragma unroll
    for (int k = 0; k < PER; ++k) acc[k] = fmaf(__ldg(wt + i * DOUT + lane + 32 * k), xi, acc[k]);
  }
  if (!LAST) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < PER; ++k) s += acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / DOUT;
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < PER; ++k) v = fmaf(acc[k] - mean, acc[k] - mean, v);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const float inv = rsqrtf(v / DOUT + 1e-5f);
#pragma unroll
    for (int k = 0; k < PER; ++k) acc[k] = fmaxf((acc[k] - mean) * inv, 0.f);
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < PER; ++k) y[lane + 32 * k] = acc[k];
  __syncwarp();
}

// block = 8 warps, 32 consecutive points of one batch element (4 per warp)
__global__ void __launch_bounds__(256) kpt_encode_kernel(const float* __restrict__ kpts, long long kn, long long kl, long long kc,
                                                         const float* __restrict__ desc, long long dn, long long dc, long long dl,
                                                         const float* __restrict__ w, const float* __restrict__ stats, int L,
                                                         float* __restrict__ out /* [N, 256, L] */) {
  __shared__ float act[8][2][KE_D3];
  __shared__ float res[KE_D4][33];
  const int n = blockIdx.y, l0 = blockIdx.x * 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float inv = stats[3];
  for (int q = 0; q < 4; ++q) {
    const int pl = warp * 4 + q, l = l0 + pl;
    if (l < L) {
      float* xa = act[warp][0];
      float* xb = act[warp][1];
      if (lane < 3) xa[lane] = (kpts[n * kn + l * kl + lane * kc] - stats[4 * n + lane]) * inv;
      __syncwarp();
      ke_layer<KE_D0, KE_D1, false>(w + KE_W1, w + KE_B1, xa, xb, lane);
      ke_layer<KE_D1, KE_D2, false>(w + KE_W2, w + KE_B2, xb, xa, lane);
      ke_layer<KE_D2, KE_D3, false>(w + KE_W3, w + KE_B3, xa, xb, lane);
      // last layer straight into the transpose buffer
      float acc[KE_D4 / 32];
#pragma unroll
      for (int k = 0; k < KE_D4 / 32; ++k) acc[k] = __ldg(w + KE_B4 + lane + 32 * k);
#pragma unroll 4
      for (int i = 0; i < KE_D3; ++i) {
        const float xi = xb[i];
#pragma unroll
        for (int k = 0; k < KE_D4 / 32; ++k) acc[k] = fmaf(__ldg(w + KE_W4 + i * KE_D4 + lane + 32 * k), xi, acc[k]);
      }
#pragma unroll
      for (int k = 0; k < KE_D4 / 32; ++k) res[lane + 32 * k][pl] = acc[k];
    }
  }
  __syncthreads();
  // out[n, c, l0 + lane] = desc[n, c, l0 + lane] + res[c][lane]: a warp writes one 128-byte line per channel
  const int l = l0 + lane;
  if (l < L)
    for (int c = warp; c < KE_D4; c += 8)
      out[(static_cast<long long>(n) * KE_D4 + c) * L + l] = desc[n * dn + c * dc + l * dl] + res[c][lane];
}

// ---- scratch of the entry point that allocates nothing (opb200_keypoint_encode_static) -------------------------
// The [n_batch][4] statistics block of up to KE_STATIC_MAX_BATCH batch elements, resident in every GPU that loads the
// library until the process exits (16 KB).  In host code the NAME below is a host-side stand-in, not a device address:
// the address a kernel may be given comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int KE_STATIC_MAX_BATCH = 1024;
static __device__ __align__(256) float g_ke_static_stats[4 * KE_STATIC_MAX_BATCH];

constexpr int KE_MAX_DEVICES = 64;
static std::atomic<float*> g_ke_static_ptr[KE_MAX_DEVICES];

// device address of g_ke_static_stats on the CURRENT device (looked up on first use per device, then cached)
static int ke_static_scratch(float** out) {
  int dev = 0;
  OPB_CHECK_CUDA(cudaGetDevice(&dev));
  OPB_CHECK_ARG(dev >= 0 && dev < KE_MAX_DEVICES, "device ordinal beyond the static scratch table (64)");
  float* p = g_ke_static_ptr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    void* sym = nullptr;
    OPB_CHECK_CUDA(cudaGetSymbolAddress(&sym, g_ke_static_stats));
    OPB_CHECK_ARG(sym != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch");
    p = static_cast<float*>(sym);
    g_ke_static_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return OPB_OK;
}

}  // namespace opb

using namespace opb;

extern "C" {

size_t opb200_keypoint_encode_packed_floats(void) { return KE_PACKED; }

int opb200_keypoint_encode(const float* kpts, int64_t kn, int64_t kl, int64_t kc, const float* desc, int64_t dn, int64_t dc,
                           int64_t dl, const float* packed_w, int n_batch, int L, int normalize, float* out, float* stats_ws,
                           void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(kpts && desc && packed_w && out && stats_ws, "null pointer");
  OPB_CHECK_ARG(n_batch >= 1 && L >= 1, "empty input");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  kpt_stats_kernel<<<n_batch, 256, 0, stream>>>(kpts, kn, kl, kc, L, normalize, stats_ws);
  kpt_encode_kernel<<<dim3((L + 31) / 32, n_batch), 256, 0, stream>>>(kpts, kn, kl, kc, desc, dn, dc, dl, packed_w, stats_ws, L, out);
  count_launch(2);
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_keypoint_encode_static_max_batch(void) { return KE_STATIC_MAX_BATCH; }

// Same result as opb200_keypoint_encode, no allocation anywhere and no scratch argument: the statistics block lives in
// g_ke_static_stats.  Every call on a device shares that array (it keeps the last call's contents; nothing in it has to
// start at zero because kpt_stats_kernel writes every word kpt_encode_kernel reads): calls that overlap on the GPU
// (different streams, no ordering between them) race on it.
int opb200_keypoint_encode_static(const float* kpts, int64_t kn, int64_t kl, int64_t kc, const float* desc, int64_t dn,
                                  int64_t dc, int64_t dl, const float* packed_w, int n_batch, int L, int normalize, float* out,
                                  void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(kpts && desc && packed_w && out, "null pointer");
  OPB_CHECK_ARG(n_batch >= 1 && L >= 1, "empty input");
  if (n_batch > KE_STATIC_MAX_BATCH) {
    set_last_error("argument error: opb200_keypoint_encode_static: n_batch " + std::to_string(n_batch) +
                   " exceeds the static scratch (" + std::to_string(KE_STATIC_MAX_BATCH) +
                   " batch elements); use opb200_keypoint_encode with a caller-provided stats_ws");
    return OPB_ERR_ARG;
  }
  float* stats = nullptr;
  if (int r = ke_static_scratch(&stats)) return r;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  kpt_stats_kernel<<<n_batch, 256, 0, stream>>>(kpts, kn, kl, kc, L, normalize, stats);
  kpt_encode_kernel<<<dim3((L + 31) / 32, n_batch), 256, 0, stream>>>(kpts, kn, kl, kc, desc, dn, dc, dl, packed_w, stats, L, out);
  count_launch(2);
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

}  // extern "C"
