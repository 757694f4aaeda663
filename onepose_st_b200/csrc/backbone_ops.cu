// Memory-bound glue of the ResNet-FPN backbone (channels-last fp32), fused so each FPN level makes one pass:
//   opb200_upsample2x_add_nhwc   dst += bilinear_x2(src), align_corners=True
//                                (F.interpolate(scale_factor=2, mode='bilinear', align_corners=True) + add;
//                                 reference: src/models/OnePosePlus/backbone/resnet.py:141-164 FPN top-down path,
//                                 submodules/LoFTR/src/loftr/backbone/resnet_fpn.py:100-118)
//   opb200_bias_leaky_relu_nhwc  x = leaky_relu(x + bias[c])   (folded BatchNorm shift + nn.LeakyReLU of the
//                                 layer{1,2}_outconv2 heads)
// The convolutions themselves stay cuDNN; these two kernels only remove extra full
// passes over the 1/2- and 1/4-resolution maps.  HBM-bound: 8 B (read+write dst) + ~1 B (src) per element.
#include <cuda_fp16.h>
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

__global__ void __launch_bounds__(256) upsample2x_add_kernel(const float4* __restrict__ src, int N, int Hs, int Ws, int C4,
                                                             float4* __restrict__ dst, float ry, float rx) {
  const int Ho = 2 * Hs, Wo = 2 * Ws;
  const long long total = static_cast<long long>(N) * Ho * Wo * C4;
  for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % C4);
    long long p = t / C4;
    const int x = static_cast<int>(p % Wo);
    p /= Wo;
    const int y = static_cast<int>(p % Ho);
    const int n = static_cast<int>(p / Ho);
    // torch's area_pixel_compute_source_index(align_corners=True): src = scale * dst
    const float sy = ry * y, sx = rx * x;
    const int y0 = static_cast<int>(sy), x0 = static_cast<int>(sx);
    const int y1 = y0 + (y0 < Hs - 1), x1 = x0 + (x0 < Ws - 1);
    const float ly1 = sy - y0, lx1 = sx - x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
    const float4* s = src + static_cast<long long>(n) * Hs * Ws * C4;
    const float4 v00 = __ldg(s + (static_cast<long long>(y0) * Ws + x0) * C4 + c);
    const float4 v01 = __ldg(s + (static_cast<long long>(y0) * Ws + x1) * C4 + c);
    const float4 v10 = __ldg(s + (static_cast<long long>(y1) * Ws + x0) * C4 + c);
    const float4 v11 = __ldg(s + (static_cast<long long>(y1) * Ws + x1) * C4 + c);
    float4 d = dst[t];
    d.x += ly0 * (lx0 * v00.x + lx1 * v01.x) + ly1 * (lx0 * v10.x + lx1 * v11.x);
    d.y += ly0 * (lx0 * v00.y + lx1 * v01.y) + ly1 * (lx0 * v10.y + lx1 * v11.y);
    d.z += ly0 * (lx0 * v00.z + lx1 * v01.z) + ly1 * (lx0 * v10.z + lx1 * v11.z);
    d.w += ly0 * (lx0 * v00.w + lx1 * v01.w) + ly1 * (lx0 * v10.w + lx1 * v11.w);
    dst[t] = d;
  }
}

__global__ void __launch_bounds__(256) bias_leaky_relu_kernel(float4* __restrict__ x, const float4* __restrict__ bias,
                                                              long long total4, int C4, float slope) {
  for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total4;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 b = __ldg(bias + t % C4);
    float4 v = x[t];
    v.x += b.x, v.y += b.y, v.z += b.z, v.w += b.w;
    v.x = v.x > 0.f ? v.x : v.x * slope;
    v.y = v.y > 0.f ? v.y : v.y * slope;
    v.z = v.z > 0.f ? v.z : v.z * slope;
    v.w = v.w > 0.f ? v.w : v.w * slope;
    x[t] = v;
  }
}


// fp16 NHWC variant: dst (fp16) += bilinear_x2(src (fp16)), interpolation in fp32, 8 channels per thread.
__global__ void __launch_bounds__(256) upsample2x_add_f16_kernel(const uint4* __restrict__ src, int N, int Hs, int Ws, int C8,
                                                                 uint4* __restrict__ dst, float ry, float rx) {
  const int Ho = 2 * Hs, Wo = 2 * Ws;
  const long long total = static_cast<long long>(N) * Ho * Wo * C8;
  for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % C8);
    long long p = t / C8;
    const int x = static_cast<int>(p % Wo);
    p /= Wo;
    const int y = static_cast<int>(p % Ho);
    const int n = static_cast<int>(p / Ho);
    const float sy = ry * y, sx = rx * x;
    const int y0 = static_cast<int>(sy), x0 = static_cast<int>(sx);
    const int y1 = y0 + (y0 < Hs - 1), x1 = x0 + (x0 < Ws - 1);
    const float ly1 = sy - y0, lx1 = sx - x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
    const uint4* s = src + static_cast<long long>(n) * Hs * Ws * C8;
    const uint4 q00 = __ldg(s + (static_cast<long long>(y0) * Ws + x0) * C8 + c);
    const uint4 q01 = __ldg(s + (static_cast<long long>(y0) * Ws + x1) * C8 + c);
    const uint4 q10 = __ldg(s + (static_cast<long long>(y1) * Ws + x0) * C8 + c);
    const uint4 q11 = __ldg(s + (static_cast<long long>(y1) * Ws + x1) * C8 + c);
    uint4 d = dst[t];
    const uint32_t* a00 = reinterpret_cast<const uint32_t*>(&q00);
    const uint32_t* a01 = reinterpret_cast<const uint32_t*>(&q01);
    const uint32_t* a10 = reinterpret_cast<const uint32_t*>(&q10);
    const uint32_t* a11 = reinterpret_cast<const uint32_t*>(&q11);
    uint32_t* dd = reinterpret_cast<uint32_t*>(&d);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 v00 = __half22float2(*reinterpret_cast<const __half2*>(&a00[i]));
      const float2 v01 = __half22float2(*reinterpret_cast<const __half2*>(&a01[i]));
      const float2 v10 = __half22float2(*reinterpret_cast<const __half2*>(&a10[i]));
      const float2 v11 = __half22float2(*reinterpret_cast<const __half2*>(&a11[i]));
      float2 o = __half22float2(*reinterpret_cast<const __half2*>(&dd[i]));
      o.x += ly0 * (lx0 * v00.x + lx1 * v01.x) + ly1 * (lx0 * v10.x + lx1 * v11.x);
      o.y += ly0 * (lx0 * v00.y + lx1 * v01.y) + ly1 * (lx0 * v10.y + lx1 * v11.y);
      const __half2 h = __floats2half2_rn(o.x, o.y);
      dd[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    dst[t] = d;
  }
}

// Stem of ResNetFPN_8_2: 7x7 stride-2 pad-3 convolution of a 1-channel image + folded-BatchNorm bias + ReLU,
// written channels-last (reference: backbone/resnet.py:141 `relu(bn1(conv1(x)))`, conv1 = Conv2d(1, 128, 7, 2, 3)).
// One block = 16 x 16 output pixels of one image; the 37 x 37 input patch and the [49][COUT] weights sit in shared
// memory; a thread owns two output pixels (rows ty and ty + 8) and walks the output channels 32 at a time, so
// every broadcast weight float4 feeds two pixels.  fp32 FMA (exact w.r.t. the fp32 reference up to summation
// order).  Output traffic dominates: 4 * COUT bytes per output pixel.
__device__ __forceinline__ void stem_store32(const float (&v)[32], float* dst) { store_f32x32_v8(v, dst); }
__device__ __forceinline__ void stem_store32(const float (&v)[32], __half* dst) {
  uint32_t h[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const __half2 p = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    h[i] = *reinterpret_cast<const uint32_t*>(&p);
  }
  ptx::st_global_v8(dst, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
  ptx::st_global_v8(dst + 16, h[8], h[9], h[10], h[11], h[12], h[13], h[14], h[15]);
}

template <int COUT, typename OutT>
__global__ void __launch_bounds__(128) stem_conv7x7_s2_kernel(const float* __restrict__ img, const float* __restrict__ w,
                                                              const float* __restrict__ bias, int H, int W, int Ho, int Wo,
                                                              OutT* __restrict__ out) {
  __shared__ float patch[37][38];
  __shared__ __align__(16) float ws[49][COUT];
  const int n = blockIdx.z, oy0 = blockIdx.y * 16, ox0 = blockIdx.x * 16;
  const int t = threadIdx.x;
  for (int i = t; i < 49 * COUT; i += 128) ws[i / COUT][i % COUT] = w[i];  // w is [49][COUT] (tap-major)
  const float* im = img + static_cast<long long>(n) * H * W;
  const int iy0 = oy0 * 2 - 3, ix0 = ox0 * 2 - 3;
  for (int i = t; i < 37 * 37; i += 128) {
    const int py = i / 37, px = i - py * 37;
    const int y = iy0 + py, x = ix0 + px;
    patch[py][px] = (y >= 0 && y < H && x >= 0 && x < W) ? __ldg(im + static_cast<long long>(y) * W + x) : 0.f;
  }
  __syncthreads();
  const int tx = t & 15, ty = t >> 4;  // ty in 0..7: pixels (ty, tx) and (ty + 8, tx)
#pragma unroll 1
  for (int c0 = 0; c0 < COUT; c0 += 32) {
    float a0[32], a1[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) a0[k] = a1[k] = __ldg(bias + c0 + k);
#pragma unroll 1
    for (int ky = 0; ky < 7; ++ky) {
#pragma unroll
      for (int kx = 0; kx < 7; ++kx) {
        const float p0 = patch[2 * ty + ky][2 * tx + kx];
        const float p1 = patch[2 * (ty + 8) + ky][2 * tx + kx];
        const float4* wv = reinterpret_cast<const float4*>(&ws[ky * 7 + kx][c0]);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 f = wv[q];
          a0[4 * q + 0] = fmaf(p0, f.x, a0[4 * q + 0]);
          a0[4 * q + 1] = fmaf(p0, f.y, a0[4 * q + 1]);
          a0[4 * q + 2] = fmaf(p0, f.z, a0[4 * q + 2]);
          a0[4 * q + 3] = fmaf(p0, f.w, a0[4 * q + 3]);
          a1[4 * q + 0] = fmaf(p1, f.x, a1[4 * q + 0]);
          a1[4 * q + 1] = fmaf(p1, f.y, a1[4 * q + 1]);
          a1[4 * q + 2] = fmaf(p1, f.z, a1[4 * q + 2]);
          a1[4 * q + 3] = fmaf(p1, f.w, a1[4 * q + 3]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      a0[k] = fmaxf(a0[k], 0.f);
      a1[k] = fmaxf(a1[k], 0.f);
    }
    const int ox = ox0 + tx;
    if (ox < Wo) {
      const int oya = oy0 + ty, oyb = oy0 + ty + 8;
      if (oya < Ho) stem_store32(a0, out + ((static_cast<long long>(n) * Ho + oya) * Wo + ox) * COUT + c0);
      if (oyb < Ho) stem_store32(a1, out + ((static_cast<long long>(n) * Ho + oyb) * Wo + ox) * COUT + c0);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// The same stem on the tensor cores (fp16 operands, fp32 accumulation: the TF32-grade precision of the other
// backbone convolutions).  GEMM view: 128 output pixels (8 x 16) x 128 channels x K = 49 taps padded to 64.
// The A tile is an im2col of the shared-memory image patch that the threads write straight into the 128-byte-swizzled
// K-major layout tcgen05.mma reads (generic-proxy stores + fence.proxy.async); B = the weights, built once per CTA.
// Persistent CTAs (4 per SM: TMEM 128 columns each); per tile: patch -> A -> 4 MMAs -> epilogue (bias, ReLU, fp16,
// quad-transposed 128-byte-line stores).  Write-bound: 256 bytes per output pixel.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) stem_tc_kernel(const float* __restrict__ img, const float* __restrict__ w,
                                                      const float* __restrict__ bias, int H, int W, int Ho, int Wo,
                                                      int tiles_x, int tiles_y, int num_tiles, __half* __restrict__ out,
                                                      int* __restrict__ range_flag) {
  __shared__ __align__(1024) uint8_t a_tile[128 * 128];
  __shared__ __align__(1024) uint8_t b_tile[128 * 128];
  __shared__ float patch[21][40];
  __shared__ __align__(8) uint64_t mma_bar;
  __shared__ uint32_t tmem_slot;
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  if (t == 0) {
    ptx::mbar_init(&mma_bar, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(&tmem_slot, 128);
    ptx::tmem_relinquish();
  }
  {  // B row n = output channel t: 64 K values (49 taps + zeros), 16-byte chunk c stored at chunk position c ^ (n % 8)
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      uint32_t p[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k0 = 8 * c + 2 * j, k1 = k0 + 1;
        const float f0 = k0 < 49 ? __ldg(w + k0 * 128 + t) : 0.f;
        const float f1 = k1 < 49 ? __ldg(w + k1 * 128 + t) : 0.f;
        const __half2 hh = __floats2half2_rn(f0, f1);
        p[j] = *reinterpret_cast<const uint32_t*>(&hh);
      }
      *reinterpret_cast<uint4*>(b_tile + t * 128 + ((c ^ (t & 7)) << 4)) = make_uint4(p[0], p[1], p[2], p[3]);
    }
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  constexpr uint32_t idesc = ptx::make_idesc_f16(128, 128);
  const int py = t >> 4, px = t & 15;  // this thread's output pixel inside the 8 x 16 tile (GEMM row t)
  float bv[4];  // bias is re-read per 4 columns below; keep nothing resident
  (void)bv;
  uint32_t phase = 0;
  float amax = 0.f;  // largest magnitude read from the image or stored as fp16 by this thread (range_flag)
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int n = tile / (tiles_x * tiles_y);
    const int r = tile - n * tiles_x * tiles_y;
    const int ty = r / tiles_x, tx = r - ty * tiles_x;
    const int oy0 = ty * 8, ox0 = tx * 16;
    const int iy0 = oy0 * 2 - 3, ix0 = ox0 * 2 - 3;
    const float* im = img + static_cast<long long>(n) * H * W;
    for (int i = t; i < 21 * 37; i += 128) {
      const int yy = i / 37, xx = i - yy * 37;
      const int y = iy0 + yy, x = ix0 + xx;
      const float pv = (y >= 0 && y < H && x >= 0 && x < W) ? __ldg(im + static_cast<long long>(y) * W + x) : 0.f;
      amax = pv == pv ? fmaxf(amax, fabsf(pv)) : INFINITY;
      patch[yy][xx] = pv;
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      uint32_t p[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k0 = 8 * c + 2 * j, k1 = k0 + 1;
        const float f0 = k0 < 49 ? patch[2 * py + k0 / 7][2 * px + k0 % 7] : 0.f;
        const float f1 = k1 < 49 ? patch[2 * py + k1 / 7][2 * px + k1 % 7] : 0.f;
        const __half2 hh = __floats2half2_rn(f0, f1);
        p[j] = *reinterpret_cast<const uint32_t*>(&hh);
      }
      *reinterpret_cast<uint4*>(a_tile + t * 128 + ((c ^ (t & 7)) << 4)) = make_uint4(p[0], p[1], p[2], p[3]);
    }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    if (t == 0) {
      ptx::tc_fence_after();
      const uint64_t a_desc = ptx::make_desc_sw128(ptx::smem_u32(a_tile));
      const uint64_t b_desc = ptx::make_desc_sw128(ptx::smem_u32(b_tile));
#pragma unroll
      for (int k = 0; k < 4; ++k) ptx::mma_bf16(tmem_base, a_desc + 2 * k, b_desc + 2 * k, idesc, k != 0);
      ptx::mma_commit(&mma_bar);
    }
    ptx::mbar_wait(&mma_bar, phase);
    phase ^= 1;
    ptx::tc_fence_after();
    // epilogue: lane owns GEMM row 32 * warp + lane; stores go through the quad transpose (one 128-byte line per quad)
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      uint32_t r0[32], r1[32];
      ptx::tmem_ld32(taddr + half * 64, r0);
      ptx::tmem_ld32(taddr + half * 64 + 32, r1);
      ptx::tmem_ld_wait();
      uint32_t it[4][8];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int c0 = half * 64 + 2 * i;
        const float2 b2 = __ldg(reinterpret_cast<const float2*>(bias + c0));
        const float v0 = fmaxf(__uint_as_float(i < 16 ? r0[2 * i] : r1[2 * i - 32]) + b2.x, 0.f);
        const float v1 = fmaxf(__uint_as_float(i < 16 ? r0[2 * i + 1] : r1[2 * i - 31]) + b2.y, 0.f);
        amax = fmaxf(fmaxf(amax, v0), v1);
        const __half2 hh = __floats2half2_rn(v0, v1);
        it[i >> 3][i & 7] = *reinterpret_cast<const uint32_t*>(&hh);
      }
      quad_transpose(it, lane);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int rr = warp * 32 + (lane & ~3) + k;  // GEMM row whose sector (lane % 4) this lane now holds
        const int oy = oy0 + (rr >> 4), ox = ox0 + (rr & 15);
        if (oy < Ho && ox < Wo) {
          __half* dst = out + ((static_cast<long long>(n) * Ho + oy) * Wo + ox) * 128 + half * 64 + (lane & 3) * 16;
          ptx::st_global_v8(dst, it[k][0], it[k][1], it[k][2], it[k][3], it[k][4], it[k][5], it[k][6], it[k][7]);
        }
      }
    }
    ptx::tc_fence_before();
    __syncthreads();  // TMEM accumulator, A tile and patch are free for the next tile
  }
  if (range_flag != nullptr && !(amax <= 65504.f)) atomicOr(range_flag, 1);
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 128);
  }
}

// fp32 NHWC [P, C] -> fp16 [P, 2C] = [hi | lo], hi = rn(x), lo = rn(x - hi): the activation format of the split-precision
// convolutions (conv.cu, terms = 3); 8 channels per thread
__global__ void __launch_bounds__(256) split_f16_planes_kernel(const float4* __restrict__ x, long long n_vec, int C8,
                                                               uint4* __restrict__ out, int* __restrict__ range_flag) {
  float amax = 0.f;
  for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n_vec; t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long p = t / C8;
    const int c8 = static_cast<int>(t - p * C8);
    const float4 a = x[2 * t], b = x[2 * t + 1];
    const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      amax = fmaxf(fmaxf(amax, fabsf(v[2 * q])), fabsf(v[2 * q + 1]));
      const __half2 h = __floats2half2_rn(v[2 * q], v[2 * q + 1]);
      const float2 hf = __half22float2(h);
      const __half2 l = __floats2half2_rn(v[2 * q] - hf.x, v[2 * q + 1] - hf.y);
      hi[q] = *reinterpret_cast<const uint32_t*>(&h);
      lo[q] = *reinterpret_cast<const uint32_t*>(&l);
    }
    out[p * 2 * C8 + c8] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    out[p * 2 * C8 + C8 + c8] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
  if (range_flag != nullptr && !(amax <= 65504.f)) atomicOr(range_flag, 1);
}

}  // namespace opb

using namespace opb;

extern "C" {

int opb200_split_f16_planes(const float* x, long long n_pixels, int C, void* out, int* range_flag, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(x && out, "null pointer");
  OPB_CHECK_ARG(n_pixels >= 1 && C >= 8 && C % 8 == 0, "bad shape (C must be a multiple of 8)");
  OPB_CHECK_ARG((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0, "16-byte alignment");
  const long long n_vec = n_pixels * (C / 8);
  const long long blocks = (n_vec + 255) / 256;
  const int grid = static_cast<int>(blocks < num_sms() * 16LL ? blocks : num_sms() * 16LL);
  split_f16_planes_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream_)>>>(reinterpret_cast<const float4*>(x), n_vec, C / 8,
                                                                                reinterpret_cast<uint4*>(out), range_flag);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_upsample2x_add_nhwc(const float* src, int n_batch, int Hs, int Ws, int C, float* dst, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(src && dst, "null pointer");
  OPB_CHECK_ARG(n_batch >= 1 && Hs >= 1 && Ws >= 1 && C >= 4 && C % 4 == 0, "bad shape (C must be a multiple of 4)");
  OPB_CHECK_ARG((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0, "16-byte alignment");
  const long long total = static_cast<long long>(n_batch) * 4 * Hs * Ws * (C / 4);
  const float ry = Hs > 1 ? static_cast<float>(Hs - 1) / static_cast<float>(2 * Hs - 1) : 0.f;
  const float rx = Ws > 1 ? static_cast<float>(Ws - 1) / static_cast<float>(2 * Ws - 1) : 0.f;
  const long long blocks = (total + 255) / 256;
  const int grid = static_cast<int>(blocks < num_sms() * 16LL ? blocks : num_sms() * 16LL);
  upsample2x_add_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<const float4*>(src), n_batch, Hs, Ws, C / 4, reinterpret_cast<float4*>(dst), ry, rx);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_bias_leaky_relu_nhwc(float* x, const float* bias, long long n_pixels, int C, float slope, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(x && bias, "null pointer");
  OPB_CHECK_ARG(n_pixels >= 1 && C >= 4 && C % 4 == 0, "bad shape (C must be a multiple of 4)");
  OPB_CHECK_ARG((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(bias) & 15) == 0, "16-byte alignment");
  const long long total4 = n_pixels * (C / 4);
  const long long blocks = (total4 + 255) / 256;
  const int grid = static_cast<int>(blocks < num_sms() * 16LL ? blocks : num_sms() * 16LL);
  bias_leaky_relu_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream_)>>>(reinterpret_cast<float4*>(x),
                                                                                reinterpret_cast<const float4*>(bias), total4,
                                                                                C / 4, slope);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_stem_conv7x7_s2_relu_nhwc(const float* image, const float* weight_tap_major, const float* bias, int n_batch, int H,
                                     int W, int Cout, float* out, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(image && weight_tap_major && bias && out, "null pointer");
  OPB_CHECK_ARG(n_batch >= 1 && H >= 1 && W >= 1, "bad shape");
  OPB_CHECK_ARG(Cout == 128, "stem kernel is instantiated for 128 output channels (initial_dim of ResNetFPN_8_2)");
  OPB_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 31) == 0, "output must be 32-byte aligned");
  const int Ho = (H + 2 * 3 - 7) / 2 + 1, Wo = (W + 2 * 3 - 7) / 2 + 1;
  dim3 grid((Wo + 15) / 16, (Ho + 15) / 16, n_batch);
  stem_conv7x7_s2_kernel<128, float><<<grid, 128, 0, static_cast<cudaStream_t>(stream_)>>>(image, weight_tap_major, bias, H, W, Ho, Wo, out);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_stem_conv7x7_s2_relu_nhwc_f16(const float* image, const float* weight_tap_major, const float* bias, int n_batch,
                                         int H, int W, int Cout, void* out_f16, int* range_flag, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(image && weight_tap_major && bias && out_f16, "null pointer");
  OPB_CHECK_ARG(n_batch >= 1 && H >= 1 && W >= 1, "bad shape");
  OPB_CHECK_ARG(Cout == 128, "stem kernel is instantiated for 128 output channels (initial_dim of ResNetFPN_8_2)");
  OPB_CHECK_ARG((reinterpret_cast<uintptr_t>(out_f16) & 31) == 0, "output must be 32-byte aligned");
  const int Ho = (H + 2 * 3 - 7) / 2 + 1, Wo = (W + 2 * 3 - 7) / 2 + 1;
  dim3 grid((Wo + 15) / 16, (Ho + 15) / 16, n_batch);
  (void)grid;
  const int tiles_x = (Wo + 15) / 16, tiles_y = (Ho + 7) / 8;
  const int num_tiles = n_batch * tiles_x * tiles_y;
  const int ctas = num_tiles < 4 * num_sms() ? num_tiles : 4 * num_sms();
  stem_tc_kernel<<<ctas, 128, 12 * 1024 /* unused: caps residency at 4 CTAs per SM = 4 x 128 TMEM columns */, static_cast<cudaStream_t>(stream_)>>>(image, weight_tap_major, bias, H, W, Ho, Wo, tiles_x,
                                                                      tiles_y, num_tiles, static_cast<__half*>(out_f16), range_flag);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_upsample2x_add_nhwc_f16(const void* src, int n_batch, int Hs, int Ws, int C, void* dst, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(src && dst, "null pointer");
  OPB_CHECK_ARG(n_batch >= 1 && Hs >= 1 && Ws >= 1 && C >= 8 && C % 8 == 0, "bad shape (C must be a multiple of 8)");
  OPB_CHECK_ARG((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0, "16-byte alignment");
  const long long total = static_cast<long long>(n_batch) * 4 * Hs * Ws * (C / 8);
  const float ry = Hs > 1 ? static_cast<float>(Hs - 1) / static_cast<float>(2 * Hs - 1) : 0.f;
  const float rx = Ws > 1 ? static_cast<float>(Ws - 1) / static_cast<float>(2 * Ws - 1) : 0.f;
  const long long blocks = (total + 255) / 256;
  const int grid = static_cast<int>(blocks < num_sms() * 16LL ? blocks : num_sms() * 16LL);
  upsample2x_add_f16_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const uint4*>(src), n_batch, Hs, Ws, C / 8, static_cast<uint4*>(dst), ry, rx);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

}  // extern "C"
