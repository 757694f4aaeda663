// Fine level: window gather (FinePreprocess) and correlation + softmax expectation (FineMatching).
//
// Replaces
//   FinePreprocess.forward     src/models/OnePosePlus/loftr_module/fine_preprocess.py:32-55
//                              submodules/LoFTR/src/loftr/loftr_module/fine_preprocess.py:29-59
//   FineMatching.forward       src/models/OnePosePlus/utils/fine_matching.py:28-110
//                              submodules/LoFTR/src/loftr/utils/fine_matching.py:15-74
//
// The reference unfolds ALL H_c*W_c windows of the fine map (52 MB at W=5, 170 MB at W=9 per image) and
// then indexes M of them; here the M windows are gathered directly: algorithmic traffic per match is
// 4*Cf*(WW+1) bytes in + the same out.  Both kernels are HBM-bound (0.5 flop/B).
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

// one block per match, one thread per channel; coalesced when the feature map is channels-last (sc == 1)
__device__ __forceinline__ void store_feat(float* p, float v) { *p = v; }
__device__ __forceinline__ void store_feat(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

template <typename OutT>
__global__ void fine_gather_kernel(const float* __restrict__ feat, long long sn, long long sc, long long sh, long long sw,
                                   int N, int Cf, int Hf, int Wf, const long long* __restrict__ b_ids,
                                   const long long* __restrict__ cell_ids, int M, int W, int stride, int Wc, int n_cells,
                                   OutT* __restrict__ out, int* __restrict__ err_flag) {
  const int m = blockIdx.x;
  if (m >= M) return;
  const long long b = b_ids[m], cell = cell_ids[m];
  const int WW = W * W;
  OutT* o = out + static_cast<long long>(m) * WW * Cf;
  if (b < 0 || b >= N || cell < 0 || cell >= n_cells) {  // the reference raises IndexError here
    if (threadIdx.x == 0) atomicExch(err_flag, 1);
    for (int idx = threadIdx.x; idx < WW * Cf; idx += blockDim.x) store_feat(o + idx, 0.f);
    return;
  }
  const int r = static_cast<int>(cell) / Wc, c = static_cast<int>(cell) % Wc;
  const int y0 = r * stride - W / 2, x0 = c * stride - W / 2;
  const float* fb = feat + b * sn;
  for (int ch = threadIdx.x; ch < Cf; ch += blockDim.x) {
    for (int wy = 0; wy < W; ++wy) {
      const int y = y0 + wy;
      for (int wx = 0; wx < W; ++wx) {
        const int x = x0 + wx;
        float v = 0.f;  // F.unfold zero padding
        if (y >= 0 && y < Hf && x >= 0 && x < Wf) v = __ldg(fb + ch * sc + y * sh + x * sw);
        store_feat(o + (wy * W + wx) * Cf + ch, v);
      }
    }
  }
}

// out[m, c] = src[b_ids[m]*sn + c*sc + i_ids[m]*sl]   (descriptors3d_db.permute(0,2,1)[b_ids, i_ids])
__global__ void gather_rows_kernel(const float* __restrict__ src, long long sn, long long sc, long long sl, int N, int C,
                                   int L, const long long* __restrict__ b_ids, const long long* __restrict__ i_ids, int M,
                                   float* __restrict__ out, int* __restrict__ err_flag) {
  const int m = blockIdx.x;
  if (m >= M) return;
  const long long b = b_ids[m], i = i_ids[m];
  const bool bad = b < 0 || b >= N || i < 0 || i >= L;
  if (bad && threadIdx.x == 0) atomicExch(err_flag, 1);
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    out[static_cast<long long>(m) * C + c] = bad ? 0.f : __ldg(src + b * sn + c * sc + i * sl);
}

// one block (4 warps) per match: warp w correlates window rows w, w+4, ... with the query vector
// (float4 per lane, Cf = 128), then warp 0 does the softmax expectation over the WW scores.
// four consecutive channels of a feature row: fp32 (16 bytes) or bf16 (8 bytes, widened exactly)
__device__ __forceinline__ float4 load4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 load4(const __nv_bfloat16* p) {
  const uint2 w = __ldg(reinterpret_cast<const uint2*>(p));
  return make_float4(__uint_as_float(w.x << 16), __uint_as_float(w.x & 0xffff0000u), __uint_as_float(w.y << 16),
                     __uint_as_float(w.y & 0xffff0000u));
}

template <int CF, typename InT>
__global__ void __launch_bounds__(128) fine_match_kernel(const InT* __restrict__ q, long long q_stride,
                                                         const InT* __restrict__ win, int M, int W,
                                                         float* __restrict__ expec_f) {
  static_assert(CF == 128, "four channels per lane");
  const int m = blockIdx.x;
  if (m >= M) return;
  __shared__ float sim[128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int WW = W * W;
  const float4 qv = load4(q + m * q_stride + 4 * lane);
  const InT* wbase = win + static_cast<long long>(m) * WW * CF;
  for (int r = warp; r < WW; r += 4) {
    const float4 wv = load4(wbase + r * CF + 4 * lane);
    float d = qv.x * wv.x;
    d = fmaf(qv.y, wv.y, d);
    d = fmaf(qv.z, wv.z, d);
    d = fmaf(qv.w, wv.w, d);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) sim[r] = d;
  }
  __syncthreads();
  if (warp != 0) return;
  const float temp = 1.0f / sqrtf(static_cast<float>(CF));  // softmax_temp = 1 / C**.5
  float x[4];
  float mx = -INFINITY;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = lane + 32 * k;
    x[k] = r < WW ? sim[r] * temp : -INFINITY;
    mx = fmaxf(mx, x[k]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float e[4], sum = 0.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    e[k] = (lane + 32 * k) < WW ? expf(x[k] - mx) : 0.f;
    sum += e[k];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  // grid = (linspace(0, W-1, W) / (W-1) - 0.5) * 2, (x, y) order (kornia 0.4.1 create_meshgrid)
  float ex = 0.f, ey = 0.f, exx = 0.f, eyy = 0.f;
  const float wm1 = static_cast<float>(W - 1);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = lane + 32 * k;
    if (r < WW) {
      const float p = e[k] / sum;
      const float gx = (static_cast<float>(r % W) / wm1 - 0.5f) * 2.f;
      const float gy = (static_cast<float>(r / W) / wm1 - 0.5f) * 2.f;
      ex = fmaf(gx, p, ex);
      ey = fmaf(gy, p, ey);
      exx = fmaf(gx * gx, p, exx);
      eyy = fmaf(gy * gy, p, eyy);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ex += __shfl_xor_sync(0xffffffffu, ex, o);
    ey += __shfl_xor_sync(0xffffffffu, ey, o);
    exx += __shfl_xor_sync(0xffffffffu, exx, o);
    eyy += __shfl_xor_sync(0xffffffffu, eyy, o);
  }
  if (lane == 0) {
    const float vx = fmaxf(exx - ex * ex, 1e-10f), vy = fmaxf(eyy - ey * ey, 1e-10f);
    expec_f[3 * m + 0] = ex;
    expec_f[3 * m + 1] = ey;
    expec_f[3 * m + 2] = sqrtf(vx) + sqrtf(vy);
  }
}

// Feature sampling at M sub-pixel keypoints (sample_feature_from_featuremap, point mode: F.grid_sample with
// align_corners=True and zero padding; src/KeypointFreeSfM/loftr_for_sfm/utils/sample_feature_from_featuremap.py:28-81,
// called from loftr.py:130-166 with mode "nearest" for the coarse map and "bilinear" for the fine map).  The coordinate
// arithmetic repeats the reference's fp32 operations one by one (coord_normalization :6-25, then ATen's
// grid_sampler_unnormalize: ((g + 1) / 2) * (size - 1)); one block per keypoint, one thread per channel.
__global__ void sample_features_kernel(const float* __restrict__ feat, long long sc, long long sh, long long sw, int C, int Hf,
                                       int Wf, const float* __restrict__ kpts, int M, float img_h, float img_w, int bilinear,
                                       int normalize, float* __restrict__ out) {
  const int m = blockIdx.x;
  if (m >= M) return;
  float g[2], px[2];
  const float res[2] = {__fmul_rn(img_w - 1.f, 1.f), __fmul_rn(img_h - 1.f, 1.f)};
  const float size1[2] = {static_cast<float>(Wf - 1), static_cast<float>(Hf - 1)};
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    g[a] = __fadd_rn(__fsub_rn(kpts[2 * m + a], 0.5f), 0.5f);             // keypoints - scale / 2 + 0.5, scale = 1
    g[a] = __fsub_rn(__fmul_rn(__fdiv_rn(g[a], res[a]), 2.f), 1.f);      // / ((w - 1) * scale) * 2 - 1
    px[a] = __fmul_rn(__fdiv_rn(__fadd_rn(g[a], 1.f), 2.f), size1[a]);   // align_corners=True un-normalisation
  }
  __shared__ float red[32];
  float sq = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float* fc = feat + c * sc;
    float v;
    if (!bilinear) {
      const float xr = nearbyintf(px[0]), yr = nearbyintf(px[1]);          // round half to even, as ATen
      const bool in = xr >= 0.f && xr <= size1[0] && yr >= 0.f && yr <= size1[1];
      v = in ? __ldg(fc + static_cast<long long>(yr) * sh + static_cast<long long>(xr) * sw) : 0.f;
    } else {
      const float x0 = floorf(px[0]), y0 = floorf(px[1]);
      const float x1 = x0 + 1.f, y1 = y0 + 1.f;
      const float nw = __fmul_rn(x1 - px[0], y1 - px[1]), ne = __fmul_rn(px[0] - x0, y1 - px[1]);
      const float sw_ = __fmul_rn(x1 - px[0], px[1] - y0), se = __fmul_rn(px[0] - x0, px[1] - y0);
      auto at = [&](float x, float y) -> float {
        return (x >= 0.f && x <= size1[0] && y >= 0.f && y <= size1[1])
                   ? __ldg(fc + static_cast<long long>(y) * sh + static_cast<long long>(x) * sw) : 0.f;
      };
      v = 0.f;
      v = fmaf(at(x0, y0), nw, v);
      v = fmaf(at(x1, y0), ne, v);
      v = fmaf(at(x0, y1), sw_, v);
      v = fmaf(at(x1, y1), se, v);
    }
    out[static_cast<long long>(m) * C + c] = v;
    sq = fmaf(v, v, sq);
  }
  if (!normalize) return;
  // F.normalize(p=2, dim=-1): v / max(|v|_2, 1e-12)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sq;
  __syncthreads();
  float tot = 0.f;
  for (int w = 0; w < static_cast<int>(blockDim.x + 31) / 32; ++w) tot += red[w];
  const float inv = 1.f / fmaxf(sqrtf(tot), 1e-12f);
  for (int c = threadIdx.x; c < C; c += blockDim.x) out[static_cast<long long>(m) * C + c] *= inv;
}

// Track-wise feature aggregation (feature_aggregation_and_update, src/KeypointFreeSfM/post_optimization/
// feature_aggregation.py:140-160, aggregation_method "avg"): out[t] = mean of the rows of track t, added one by one in fp32
// in row order and divided by the count -- the arithmetic of np.mean(stack, axis=0) on float32 rows.
__global__ void track_mean_kernel(const float* __restrict__ rows, const long long* __restrict__ start, int T, int C,
                                  float* __restrict__ out) {
  const int t = blockIdx.x;
  if (t >= T) return;
  const long long r0 = start[t], r1 = start[t + 1];
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (long long r = r0; r < r1; ++r) s = __fadd_rn(s, rows[r * C + c]);
    out[static_cast<long long>(t) * C + c] = r1 > r0 ? __fdiv_rn(s, static_cast<float>(r1 - r0)) : 0.f;
  }
}

}  // namespace opb

using namespace opb;

extern "C" {

int opb200_sample_features(const float* feat, int64_t sc, int64_t sh, int64_t sw, int C, int Hf, int Wf, const float* kpts, int M,
                           float img_h, float img_w, int bilinear, int normalize, float* out, void* stream_) {
  if (int r = check_device()) return r;
  if (M <= 0) return OPB_OK;
  OPB_CHECK_ARG(feat && kpts && out, "null pointer");
  OPB_CHECK_ARG(C >= 1 && Hf >= 1 && Wf >= 1 && img_h > 1.f && img_w > 1.f, "bad shape");
  const int threads = C >= 256 ? 256 : (C >= 128 ? 128 : 64);
  sample_features_kernel<<<M, threads, 0, static_cast<cudaStream_t>(stream_)>>>(feat, sc, sh, sw, C, Hf, Wf, kpts, M, img_h, img_w,
                                                                                bilinear, normalize, out);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_track_feature_mean(const float* rows, const int64_t* track_start, int n_tracks, int C, float* out, void* stream_) {
  if (int r = check_device()) return r;
  if (n_tracks <= 0) return OPB_OK;
  OPB_CHECK_ARG(rows && track_start && out && C >= 1, "null pointer / bad shape");
  track_mean_kernel<<<n_tracks, C >= 128 ? 128 : 64, 0, static_cast<cudaStream_t>(stream_)>>>(
      rows, reinterpret_cast<const long long*>(track_start), n_tracks, C, out);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_fine_gather(const float* feat, int64_t sn, int64_t sc, int64_t sh, int64_t sw, int n_batch, int Cf, int Hf,
                       int Wf, const int64_t* b_ids, const int64_t* cell_ids, int M, int W, int stride, int Hc, int Wc,
                       float* out, int* err_flag, void* stream_) {
  if (int r = check_device()) return r;
  if (M <= 0) return OPB_OK;
  OPB_CHECK_ARG(feat && b_ids && cell_ids && out && err_flag, "null pointer");
  OPB_CHECK_ARG(W >= 1 && (W & 1) == 1 && W <= 11, "window size must be odd and <= 11");
  OPB_CHECK_ARG(stride >= 1 && Hc >= 1 && Wc >= 1, "bad grid");
  fine_gather_kernel<float><<<M, 128, 0, static_cast<cudaStream_t>(stream_)>>>(
      feat, sn, sc, sh, sw, n_batch, Cf, Hf, Wf, reinterpret_cast<const long long*>(b_ids),
      reinterpret_cast<const long long*>(cell_ids), M, W, stride, Wc, Hc * Wc, out, err_flag);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_fine_gather_bf16(const float* feat, int64_t sn, int64_t sc, int64_t sh, int64_t sw, int n_batch, int Cf, int Hf,
                            int Wf, const int64_t* b_ids, const int64_t* cell_ids, int M, int W, int stride, int Hc, int Wc,
                            void* out_bf16, int* err_flag, void* stream_) {
  if (int r = check_device()) return r;
  if (M <= 0) return OPB_OK;
  OPB_CHECK_ARG(feat && b_ids && cell_ids && out_bf16 && err_flag, "null pointer");
  OPB_CHECK_ARG(W >= 1 && (W & 1) == 1 && W <= 11, "window size must be odd and <= 11");
  OPB_CHECK_ARG(stride >= 1 && Hc >= 1 && Wc >= 1, "bad grid");
  fine_gather_kernel<__nv_bfloat16><<<M, 128, 0, static_cast<cudaStream_t>(stream_)>>>(
      feat, sn, sc, sh, sw, n_batch, Cf, Hf, Wf, reinterpret_cast<const long long*>(b_ids),
      reinterpret_cast<const long long*>(cell_ids), M, W, stride, Wc, Hc * Wc, static_cast<__nv_bfloat16*>(out_bf16), err_flag);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_gather_rows(const float* src, int64_t sn, int64_t sc, int64_t sl, int n_batch, int C, int L,
                       const int64_t* b_ids, const int64_t* i_ids, int M, float* out, int* err_flag, void* stream_) {
  if (int r = check_device()) return r;
  if (M <= 0) return OPB_OK;
  OPB_CHECK_ARG(src && b_ids && i_ids && out && err_flag, "null pointer");
  gather_rows_kernel<<<M, 128, 0, static_cast<cudaStream_t>(stream_)>>>(
      src, sn, sc, sl, n_batch, C, L, reinterpret_cast<const long long*>(b_ids), reinterpret_cast<const long long*>(i_ids), M,
      out, err_flag);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_fine_match(const float* q, int64_t q_stride, const float* win, int M, int W, int Cf, float* expec_f,
                      void* stream_) {
  if (int r = check_device()) return r;
  if (M <= 0) return OPB_OK;
  OPB_CHECK_ARG(q && win && expec_f, "null pointer");
  OPB_CHECK_ARG(Cf == 128, "fine feature dim must be 128");
  OPB_CHECK_ARG(W >= 3 && (W & 1) == 1 && W * W <= 128, "window size must be odd with W*W <= 128");
  OPB_CHECK_ARG(q_stride % 4 == 0 && (reinterpret_cast<uintptr_t>(q) & 15) == 0 && (reinterpret_cast<uintptr_t>(win) & 15) == 0,
                "q / win must be 16-byte aligned");
  fine_match_kernel<128, float><<<M, 128, 0, static_cast<cudaStream_t>(stream_)>>>(q, q_stride, win, M, W, expec_f);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_fine_match_bf16(const void* q, int64_t q_stride, const void* win, int M, int W, int Cf, float* expec_f,
                           void* stream_) {
  if (int r = check_device()) return r;
  if (M <= 0) return OPB_OK;
  OPB_CHECK_ARG(q && win && expec_f, "null pointer");
  OPB_CHECK_ARG(Cf == 128, "fine feature dim must be 128");
  OPB_CHECK_ARG(W >= 3 && (W & 1) == 1 && W * W <= 128, "window size must be odd with W*W <= 128");
  OPB_CHECK_ARG(q_stride % 4 == 0 && (reinterpret_cast<uintptr_t>(q) & 7) == 0 && (reinterpret_cast<uintptr_t>(win) & 7) == 0,
                "q / win must be 8-byte aligned");
  fine_match_kernel<128, __nv_bfloat16><<<M, 128, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(q), q_stride, static_cast<const __nv_bfloat16*>(win), M, W, expec_f);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

}  // extern "C"
