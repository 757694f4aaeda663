// Implicit-GEMM convolution (1x1 / 3x3, stride 1 / 2, "same" padding) on tcgen05 tensor cores for the ResNet-FPN
// backbone of the matchers (reference: src/models/OnePosePlus/backbone/resnet.py:11-164 BasicBlock / ResNetFPN_8_2,
// submodules/LoFTR/src/loftr/backbone/resnet_fpn.py).  Activations are NHWC fp16, weights fp16 [tap][Cout][Cin]
// with BatchNorm folded in; accumulation is fp32 in TMEM.  fp16 carries the same 11-bit significand as the TF32
// operands the reference's cuDNN convolutions use on a GPU (torch.backends.cudnn.allow_tf32 defaults to True), at
// twice the tensor-core rate and half the activation traffic.
//
//   D[pixels x Cout] = sum over taps (ky,kx) and 64-channel chunks  X[pixel + (ky,kx) - pad][chunk] * W[tap][Cout][chunk]^T
//
// One output tile = MSUB sub-tiles of 8 x 16 output pixels (128 GEMM rows each) x all BN = Cout channels.  The A operand
// of one (tap, chunk) k-block is ONE 4-D TMA box {64 ch, 16 (x stride), 8*MSUB (x stride), 1} of the NHWC tensor placed at
// the shifted pixel origin: TMA's out-of-bounds zero fill IS the convolution's zero padding, and its element strides
// give the stride-2 sampling, so no im2col buffer ever exists.  The B operand is the [BN x 64] weight slab of that tap.
// Both sub-tiles consume the same weight slab (MSUB = 2 halves the weight refetch per flop).
// Warp roles as in gemm.cuh: warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2..17 epilogue
// (bias + residual + ReLU / LeakyReLU, fp16 and/or fp32 NHWC stores, 32 bytes per lane per store).
#include <cstdlib>
#include <cuda_fp16.h>
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

constexpr int CONV_TW = 16;  // output pixels per tile row
constexpr int CONV_TH = 8;   // output rows per sub-tile (8 x 16 = 128 GEMM rows)

struct ConvPlan {
  CUtensorMap x_map;  // 4-D (C, W, H, N) fp16
  CUtensorMap w_map;  // 2-D [taps * BN, Cin_s] fp16
  int n_img, Ho, Wo;
  int tiles_x, tiles_y, num_tiles;
  int ksize, pad, stride;
  int nchunks;  // 64-channel chunks per tap
  int nk_last;  // K=16 MMAs in the last chunk (1..4)
  // Split precision ("f16x3"): activations and weights are stored as two fp16 planes side by side along the channel axis,
  // x = [hi | lo] with hi = rn(v), lo = rn(v - hi) (22 significant bits together), and every (tap, chunk) k-block is
  // issued for the three products hi*hi, lo*hi, hi*lo -- fp32-grade convolutions on the fp16 tensor pipe (terms = 3,
  // plane = channels of one plane); plain fp16 mode: terms = 1.
  int terms, plane;
  CUtensorMap xh_map;  // HALO kernels: the same tensor with boxes two rows taller (8 * MSUB + 2): one box serves the taps ky = 0..2
  long long* trace;  // debug (opb200_debug_trace): per CTA and tile {first TMA, MMA start, accumulator full, epilogue done}
};

struct ConvEpi {
  const float* bias;  // [BN] or null
  const __half* res;  // NHWC [n, Ho, Wo, BN] residual added before the activation, or null
  __half* y16;        // NHWC fp16 output or null
  float* y32;         // NHWC fp32 output or null
  int act;            // 0 none, 1 ReLU, 2 LeakyReLU(slope)
  float slope;
  // FPN top-down path fused in: y += bilinear x2 (align_corners = True) of up [n, Ho/2, Wo/2, BN] fp16 (null = off)
  const __half* up;
  float up_ry, up_rx;
  // nullable: bit 0 is set when a value stored as fp16 leaves the fp16 range (|v| > 65504): the caller must not trust
  // this executor's result for these inputs (the host side repeats the backbone in fp32)
  int* range_flag;
  // split precision: y16 / res / up hold [hi | lo] planes of BN channels each (pixel stride 2 * BN); acc_scale undoes the
  // power-of-two scaling that keeps the lo plane of the weights in fp16's normal range
  int split;
  float acc_scale;
};

struct UpTaps {  // the four source pixels (element offsets) and weights of one output pixel
  long long o00, o01, o10, o11;
  float w00, w01, w10, w11;
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(ptx::smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(ptx::smem_u32(bar)), "r"(c0), "r"(c1),
      "r"(c2), "r"(c3)
      : "memory");
}

template <int BN, int MSUB>
struct ConvCfg {
  static constexpr int A_BYTES = MSUB * BLOCK_M * BLOCK_K * 2;
  static constexpr int B_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int ACC_COLS = MSUB * BN;
  static constexpr int ACC_STAGES = (2 * ACC_COLS <= 512) ? 2 : 1;
  static constexpr int TMEM_COLS = (ACC_STAGES * ACC_COLS <= 128) ? 128 : (ACC_STAGES * ACC_COLS <= 256) ? 256 : 512;
  static constexpr int STAGES = (200 * 1024) / STAGE_BYTES > 6 ? 6 : (200 * 1024) / STAGE_BYTES;
  static constexpr int SMEM = 1024 + STAGES * STAGE_BYTES + 256;
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N");
  static_assert(STAGE_BYTES % 1024 == 0, "stage alignment");
  static_assert(ACC_COLS <= 512, "TMEM");
  static_assert(STAGES >= 3, "ring too shallow");
};

// HALO (3x3, stride 1): the three taps of one kernel column read the same pixels one row apart, so ONE box of
// 8 * MSUB + 2 rows at the column's x shift serves ky = 0, 1, 2 -- the A operand of tap ky starts ky rows (2048 bytes: a
// whole number of swizzle atoms) into it.  The activations are written into shared memory three times per 64 channels
// instead of nine.  What that buys is bandwidth of the shared-memory port (128 B/clk per SM), which the operand reads of
// 128-wide MMAs (8 KB per 64 clk) need all of: with f bytes of TMA fill per clk of tensor work the pipe is busy at most
// 128 / (128 + f) of the time -- 57 % with one box per tap (f = 96), 70 % with the halo boxes (f = 56); measured 56 % and
// 63-64 %.  Shared memory: three halo buffers + a ring of weight slabs (one per tap), each with its own
// full / empty barriers.
template <int BN, int MSUB>
struct HaloCfg {
  static constexpr int A_BYTES = (CONV_TH * MSUB + 2) * CONV_TW * BLOCK_K * 2;
  static constexpr int B_BYTES = BN * BLOCK_K * 2;
  static constexpr int NA = 3;
  static constexpr int NB = (222 * 1024 - NA * A_BYTES) / B_BYTES;
  static constexpr int SMEM = 1024 + NA * A_BYTES + NB * B_BYTES + 256;
  static_assert(A_BYTES % 1024 == 0 && B_BYTES % 1024 == 0, "swizzle atoms");
  static_assert(NB >= 3 && NB <= 10, "weight ring");
  static_assert((2 * NA + 2 * NB + 4) * 8 + 4 <= 256, "barrier block");
  static_assert(SMEM <= 227 * 1024, "shared memory budget exceeded");
};

// 16 accumulator columns of one pixel: bias, residual, activation, stores.
// 16 fp16 values at p (+ the lo plane BN further on when split) as floats
template <int BN>
__device__ __forceinline__ void load16_planes(const __half* p, bool split, float (&f)[16]) {
  uint32_t h[8];
  ptx::ld_global_v8(p, h);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&h[i]));
    f[2 * i] = t.x, f[2 * i + 1] = t.y;
  }
  if (split) {
    ptx::ld_global_v8(p + BN, h);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&h[i]));
      f[2 * i] += t.x, f[2 * i + 1] += t.y;
    }
  }
}

template <int BN, bool SPLIT>
__device__ __forceinline__ void conv_epi_chunk(const uint32_t (&r)[16], const ConvEpi& ep, long long off16, long long off32, int c0,
                                               const UpTaps& ut, float& amax) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = SPLIT ? __uint_as_float(r[i]) * ep.acc_scale : __uint_as_float(r[i]);
  if (ep.up != nullptr) {
    if constexpr (SPLIT) {  // one source pixel at a time (hi + lo planes): registers
      float t[16];
      load16_planes<BN>(ep.up + ut.o00 + c0, true, t);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaf(ut.w00, t[i], v[i]);
      load16_planes<BN>(ep.up + ut.o01 + c0, true, t);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaf(ut.w01, t[i], v[i]);
      load16_planes<BN>(ep.up + ut.o10 + c0, true, t);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaf(ut.w10, t[i], v[i]);
      load16_planes<BN>(ep.up + ut.o11 + c0, true, t);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaf(ut.w11, t[i], v[i]);
    } else {
      uint32_t a[8], b[8], c[8], d[8];
      ptx::ld_global_v8(ep.up + ut.o00 + c0, a);
      ptx::ld_global_v8(ep.up + ut.o01 + c0, b);
      ptx::ld_global_v8(ep.up + ut.o10 + c0, c);
      ptx::ld_global_v8(ep.up + ut.o11 + c0, d);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 fa = __half22float2(*reinterpret_cast<const __half2*>(&a[i]));
        const float2 fb = __half22float2(*reinterpret_cast<const __half2*>(&b[i]));
        const float2 fc = __half22float2(*reinterpret_cast<const __half2*>(&c[i]));
        const float2 fd = __half22float2(*reinterpret_cast<const __half2*>(&d[i]));
        v[2 * i] += ut.w00 * fa.x + ut.w01 * fb.x + ut.w10 * fc.x + ut.w11 * fd.x;
        v[2 * i + 1] += ut.w00 * fa.y + ut.w01 * fb.y + ut.w10 * fc.y + ut.w11 * fd.y;
      }
    }
  }
  if (ep.bias != nullptr) {
    const float4* b = reinterpret_cast<const float4*>(ep.bias + c0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 f = __ldg(b + q);
      v[4 * q] += f.x, v[4 * q + 1] += f.y, v[4 * q + 2] += f.z, v[4 * q + 3] += f.w;
    }
  }
  if (ep.res != nullptr) {
    if constexpr (SPLIT) {
      float f[16];
      load16_planes<BN>(ep.res + off16 + c0, true, f);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] += f[i];
    } else {
      uint32_t h[8];
      ptx::ld_global_v8(ep.res + off16 + c0, h);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&h[i]));
        v[2 * i] += f.x, v[2 * i + 1] += f.y;
      }
    }
  }
  if (ep.act == 1) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (ep.act == 2) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * ep.slope;
  }
  if (ep.y16 != nullptr) {
    uint32_t h[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      amax = fmaxf(fmaxf(amax, fabsf(v[2 * i])), fabsf(v[2 * i + 1]));  // (all inputs were finite: see ConvEpi::range_flag)
      const __half2 p = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
      h[i] = *reinterpret_cast<const uint32_t*>(&p);
    }
    ptx::st_global_v8(ep.y16 + off16 + c0, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
    if constexpr (SPLIT) {  // lo plane: what the hi plane's rounding left over
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&h[i]));
        const __half2 p = __floats2half2_rn(v[2 * i] - hi.x, v[2 * i + 1] - hi.y);
        h[i] = *reinterpret_cast<const uint32_t*>(&p);
      }
      ptx::st_global_v8(ep.y16 + off16 + BN + c0, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
    }
  }
  if (ep.y32 != nullptr) {
    float* d = ep.y32 + off32 + c0;
#pragma unroll
    for (int q = 0; q < 2; ++q)
      ptx::st_global_v8(d + 8 * q, __float_as_uint(v[8 * q]), __float_as_uint(v[8 * q + 1]), __float_as_uint(v[8 * q + 2]),
                        __float_as_uint(v[8 * q + 3]), __float_as_uint(v[8 * q + 4]), __float_as_uint(v[8 * q + 5]),
                        __float_as_uint(v[8 * q + 6]), __float_as_uint(v[8 * q + 7]));
  }
}

// CL2: the kernel runs as clusters of two CTAs that walk their tiles in lockstep; every k-block each CTA fetches HALF of
// the weight slab and multicasts it into both CTAs' shared memory, so the L2 -> SM weight traffic per flop halves
// (the BN = 208 layers are bound by that fill, not by the tensor pipe).  A stage is refilled only after the MMAs of
// both CTAs have read it (empty barriers count two multicast commits).
template <int BN, int MSUB, bool CL2, bool SPLIT, bool HALO>
__global__ void __launch_bounds__(64 + 512, 1) conv_kernel(const __grid_constant__ ConvPlan plan, const ConvEpi ep) {
  using Cfg = ConvCfg<BN, MSUB>;
  using HCfg = HaloCfg<BN, MSUB>;
  static_assert(!(CL2 && HALO), "the halo kernels do not multicast");
  constexpr int STAGES = HALO ? HCfg::NB : Cfg::STAGES;  // HALO: full_bar / empty_bar belong to the weight ring
  constexpr int NA = HALO ? HCfg::NA : 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_hb = smem + HCfg::NA * HCfg::A_BYTES;  // HALO: weight ring behind the halo buffers
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(HALO ? smem_hb + HCfg::NB * HCfg::B_BYTES : smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* afull_bar = tempty_bar + 2;   // HALO only
  uint64_t* aempty_bar = afull_bar + NA;  // HALO only
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aempty_bar + NA);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&plan.x_map);
    ptx::prefetch_tmap(&plan.w_map);
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], CL2 ? 2 : 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1);
      ptx::mbar_init(&tempty_bar[s], 16);
    }
    if constexpr (HALO) {
      ptx::prefetch_tmap(&plan.xh_map);
      for (int s = 0; s < NA; ++s) {
        ptx::mbar_init(&afull_bar[s], 1);
        ptx::mbar_init(&aempty_bar[s], 1);
      }
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  if constexpr (CL2) ptx::cluster_sync();  // the peer's barriers are initialised before anything is multicast to them
  else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int taps = plan.ksize * plan.ksize;
  const int tiles_per_img = plan.tiles_x * plan.tiles_y;
  // tile walk: plain = blockIdx.x, += gridDim.x; CL2 = the pair takes tiles (2q, 2q + 1), q = pair, += number of pairs;
  // both CTAs of a pair run the same number of iterations (a tile index past the end is a ghost: loads and MMAs on the
  // last tile's data, no stores)
  const int crank = CL2 ? static_cast<int>(ptx::cluster_ctarank()) : 0;
  const int t_first = CL2 ? 2 * static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int t_step = CL2 ? static_cast<int>(gridDim.x) : static_cast<int>(gridDim.x);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (ptx::elect_one()) {
      int stage = 0, astage = 0;
      uint32_t phase = 0, aphase = 0;
      for (int tq = t_first; tq < plan.num_tiles; tq += t_step) {
        const int t = min(tq + crank, plan.num_tiles - 1);
        const int n = t / tiles_per_img;
        const int r = t - n * tiles_per_img;
        const int ty = r / plan.tiles_x, tx = r - ty * plan.tiles_x;
        const int ix0 = tx * CONV_TW * plan.stride - plan.pad;
        const int iy0 = ty * (CONV_TH * MSUB) * plan.stride - plan.pad;
        if (plan.trace != nullptr) {
          const int it = (tq - t_first) / t_step;
          if (it < kTraceTilesPerCta) plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 0] = clock64();
        }
        if constexpr (HALO) {
          for (int kx = 0; kx < 3; ++kx) {
            for (int term = 0; term < plan.terms; ++term) {
              const int a0 = term == 1 ? plan.plane : 0, b0 = term == 2 ? plan.plane : 0;
              for (int ch = 0; ch < plan.nchunks; ++ch) {
                ptx::mbar_wait(&aempty_bar[astage], aphase ^ 1);
                ptx::mbar_arrive_expect_tx(&afull_bar[astage], HCfg::A_BYTES);
                tma_load_4d(smem + astage * HCfg::A_BYTES, &plan.xh_map, &afull_bar[astage], a0 + ch * BLOCK_K, ix0 + kx, iy0, n);
                if (++astage == NA) astage = 0, aphase ^= 1;
                for (int ky = 0; ky < 3; ++ky) {
                  ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                  ptx::mbar_arrive_expect_tx(&full_bar[stage], HCfg::B_BYTES);
                  ptx::tma_load_2d(smem_hb + stage * HCfg::B_BYTES, &plan.w_map, &full_bar[stage], b0 + ch * BLOCK_K, (ky * 3 + kx) * BN);
                  if (++stage == STAGES) stage = 0, phase ^= 1;
                }
              }
            }
          }
          continue;
        }
        for (int tap = 0; tap < taps; ++tap) {
          const int ky = tap / plan.ksize, kx = tap - ky * plan.ksize;
          for (int term = 0; term < plan.terms; ++term) {  // hi*hi | lo*hi | hi*lo (plane offsets along the channel axis)
            const int a0 = term == 1 ? plan.plane : 0, b0 = term == 2 ? plan.plane : 0;
            for (int ch = 0; ch < plan.nchunks; ++ch) {
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              ptx::mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
              uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
              tma_load_4d(sa, &plan.x_map, &full_bar[stage], a0 + ch * BLOCK_K, ix0 + kx, iy0 + ky, n);
              if constexpr (CL2) {
                ptx::tma_load_2d_multicast(sa + Cfg::A_BYTES + crank * (Cfg::B_BYTES / 2), &plan.w_map, &full_bar[stage],
                                           b0 + ch * BLOCK_K, tap * BN + crank * (BN / 2), 3);
              } else {
                ptx::tma_load_2d(sa + Cfg::A_BYTES, &plan.w_map, &full_bar[stage], b0 + ch * BLOCK_K, tap * BN);
              }
              if (++stage == STAGES) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = ptx::make_idesc_f16(BLOCK_M, BN);
    int stage = 0, astage = 0;
    uint32_t phase = 0, aphase = 0;
    int it = 0;
    const int total_kb = taps * plan.terms * plan.nchunks;
    for (int tq = t_first; tq < plan.num_tiles; tq += t_step, ++it) {
      const int acc = it % Cfg::ACC_STAGES;
      const uint32_t acc_phase = (it / Cfg::ACC_STAGES) & 1;
      ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * Cfg::ACC_COLS;
      if (plan.trace != nullptr && it < kTraceTilesPerCta && lane == 0)
        plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 1] = clock64();
      if constexpr (HALO) {
        // one elected thread issues the whole tile: the per-k-block warp-wide election and the run-time k loops of the
        // per-tap path cost ~150 clk per sub-tile, which the 64-clk MMAs of the 128-channel layers do not hide
        if (ptx::elect_one()) {
          const int groups = 3 * plan.terms * plan.nchunks;  // (kx, term, chunk): one halo box, three weight slabs
          int ch = 0;
          for (int g = 0; g < groups; ++g) {
            ptx::mbar_wait(&afull_bar[astage], aphase);
            const bool last_chunk = ch == plan.nchunks - 1;
            const uint32_t a_addr = ptx::smem_u32(smem + astage * HCfg::A_BYTES);
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
              ptx::mbar_wait(&full_bar[stage], phase);
              ptx::tc_fence_after();
              const uint64_t b_desc = ptx::make_desc_sw128(ptx::smem_u32(smem_hb + stage * HCfg::B_BYTES));
              // sub-tile s, tap row ky: the operand starts (8 s + ky) pixel rows into the halo box
              if (!last_chunk || plan.nk_last == BLOCK_K / 16) {
#pragma unroll
                for (int s = 0; s < MSUB; ++s) {
                  const uint64_t a_desc = ptx::make_desc_sw128(a_addr + (s * CONV_TH + ky) * (CONV_TW * BLOCK_K * 2));
#pragma unroll
                  for (int k = 0; k < BLOCK_K / 16; ++k)
                    ptx::mma_bf16(d_tmem + s * BN, a_desc + 2 * k, b_desc + 2 * k, idesc, (g | ky | k) != 0);
                }
              } else {
#pragma unroll
                for (int s = 0; s < MSUB; ++s) {
                  const uint64_t a_desc = ptx::make_desc_sw128(a_addr + (s * CONV_TH + ky) * (CONV_TW * BLOCK_K * 2));
                  for (int k = 0; k < plan.nk_last; ++k)
                    ptx::mma_bf16(d_tmem + s * BN, a_desc + 2 * k, b_desc + 2 * k, idesc, (g | ky | k) != 0);
                }
              }
              ptx::mma_commit(&empty_bar[stage]);
              if (ky == 2) ptx::mma_commit(&aempty_bar[astage]);
              if (g == groups - 1 && ky == 2) ptx::mma_commit(&tfull_bar[acc]);
              if (++stage == STAGES) stage = 0, phase ^= 1;
            }
            if (++astage == NA) astage = 0, aphase ^= 1;
            if (++ch == plan.nchunks) ch = 0;
          }
        }
        __syncwarp();
        continue;
      }
      int ch = 0;
      for (int kb = 0; kb < total_kb; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        const int nk = (ch == plan.nchunks - 1) ? plan.nk_last : BLOCK_K / 16;
        if (ptx::elect_one()) {
          const uint32_t a_addr = ptx::smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint64_t b_desc = ptx::make_desc_sw128(a_addr + Cfg::A_BYTES);
#pragma unroll
          for (int s = 0; s < MSUB; ++s) {
            const uint64_t a_desc = ptx::make_desc_sw128(a_addr + s * (BLOCK_M * BLOCK_K * 2));
            for (int k = 0; k < nk; ++k)
              ptx::mma_bf16(d_tmem + s * BN, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
          }
          if constexpr (CL2) ptx::mma_commit_multicast(&empty_bar[stage], 3);
          else ptx::mma_commit(&empty_bar[stage]);
          if (kb == total_kb - 1) ptx::mma_commit(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++ch == plan.nchunks) ch = 0;
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue: 16 warps, four per TMEM lane quarter
    // (two per SM sub-partition is too few to hide the TMEM / global latencies of this loop)
    const int quarter = warp & 3;
    const int part = (warp - 2) >> 2;  // 0..3: MSUB == 2 -> (sub-tile, column half); MSUB == 1 -> column quarter
    constexpr int NCH = BN / 16;
    constexpr int NSPLIT = MSUB == 2 ? 2 : 4;
    const int sub = MSUB == 2 ? (part & 1) : 0;
    const int cpart = MSUB == 2 ? (part >> 1) : part;
    const int ch_begin = (NCH * cpart) / NSPLIT;
    const int ch_end = (NCH * (cpart + 1)) / NSPLIT;
    const int row = quarter * 32 + lane;  // GEMM row inside the sub-tile
    const int hy = row / CONV_TW, wx = row - hy * CONV_TW;
    int it = 0;
    float amax = 0.f;  // largest magnitude this thread stored as fp16
    for (int tq = t_first; tq < plan.num_tiles; tq += t_step, ++it) {
      const int acc = it % Cfg::ACC_STAGES;
      const uint32_t acc_phase = (it / Cfg::ACC_STAGES) & 1;
      const bool ghost = tq + crank >= plan.num_tiles;
      const int t = min(tq + crank, plan.num_tiles - 1);
      const int n = t / tiles_per_img;
      const int r = t - n * tiles_per_img;
      const int ty = r / plan.tiles_x, tx = r - ty * plan.tiles_x;
      const int oy = ty * (CONV_TH * MSUB) + sub * CONV_TH + hy, ox = tx * CONV_TW + wx;
      const bool valid = oy < plan.Ho && ox < plan.Wo && !ghost;
      const long long pix = (static_cast<long long>(n) * plan.Ho + oy) * plan.Wo + ox;
      constexpr int st16 = SPLIT ? 2 * BN : BN;  // channels per pixel of the fp16 tensors
      const long long off = pix * BN, off16 = pix * st16;
      if (valid && ep.res != nullptr) ptx::prefetch_l2(ep.res + off16 + ch_begin * 16);
      UpTaps ut = {};
      if (ep.up != nullptr && valid) {  // torch's align_corners=True source index: src = dst * (in - 1) / (out - 1)
        const int Hs = plan.Ho >> 1, Ws = plan.Wo >> 1;
        const float sy = ep.up_ry * oy, sx = ep.up_rx * ox;
        const int y0 = static_cast<int>(sy), x0 = static_cast<int>(sx);
        const int y1 = y0 + (y0 < Hs - 1), x1 = x0 + (x0 < Ws - 1);
        const float ly1 = sy - y0, lx1 = sx - x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
        const long long base = static_cast<long long>(n) * Hs * Ws;
        ut.o00 = (base + static_cast<long long>(y0) * Ws + x0) * st16;
        ut.o01 = (base + static_cast<long long>(y0) * Ws + x1) * st16;
        ut.o10 = (base + static_cast<long long>(y1) * Ws + x0) * st16;
        ut.o11 = (base + static_cast<long long>(y1) * Ws + x1) * st16;
        ut.w00 = ly0 * lx0, ut.w01 = ly0 * lx1, ut.w10 = ly1 * lx0, ut.w11 = ly1 * lx1;
      }
      ptx::mbar_wait(&tfull_bar[acc], acc_phase);
      ptx::tc_fence_after();
      const bool tr = plan.trace != nullptr && it < kTraceTilesPerCta && warp == 2 && lane == 0;
      if (tr) plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 2] = clock64();
      const uint32_t taddr = tmem_base + acc * Cfg::ACC_COLS + sub * BN + (static_cast<uint32_t>(quarter * 32) << 16);
      int c = ch_begin;
      for (; c + 1 < ch_end; c += 2) {
        uint32_t r0[16], r1[16];
        ptx::tmem_ld16(taddr + c * 16, r0);
        ptx::tmem_ld16(taddr + c * 16 + 16, r1);
        ptx::tmem_ld_wait();
        if (valid) {
          conv_epi_chunk<BN, SPLIT>(r0, ep, off16, off, c * 16, ut, amax);
          conv_epi_chunk<BN, SPLIT>(r1, ep, off16, off, c * 16 + 16, ut, amax);
        }
      }
      if (c < ch_end) {
        uint32_t r0[16];
        ptx::tmem_ld16(taddr + c * 16, r0);
        ptx::tmem_ld_wait();
        if (valid) conv_epi_chunk<BN, SPLIT>(r0, ep, off16, off, c * 16, ut, amax);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[acc]);
      if (tr) plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 3] = clock64();
    }
    if (ep.range_flag != nullptr && !(amax <= 65504.f)) atomicOr(ep.range_flag, 1);
  }

  ptx::tc_fence_before();
  if constexpr (CL2) ptx::cluster_sync();  // no CTA leaves while its peer may still multicast into it or arrive on its barriers
  else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

template <int BN, int MSUB, bool CL2, bool SPLIT = false, bool HALO = false>
static int launch_conv(const ConvPlan& plan_in, const ConvEpi& ep, cudaStream_t stream) {
  using Cfg = ConvCfg<BN, MSUB>;
  auto kern = conv_kernel<BN, MSUB, CL2, SPLIT, HALO>;
  static_assert(Cfg::SMEM <= 227 * 1024, "shared memory budget exceeded");
  constexpr int kSmem = HALO ? HaloCfg<BN, MSUB>::SMEM : Cfg::SMEM;
  static std::atomic<unsigned long long> configured_dev{0};
  if (int r = opt_in_dyn_smem(kern, kSmem, configured_dev)) return r;
  static std::atomic<int> max_clusters_a{0};  // (all devices of a box are the same part)
  int max_clusters = max_clusters_a.load();
  if (max_clusters == 0) {
    max_clusters = 1;
    if (CL2) {
      cudaLaunchConfig_t q = {};
      q.gridDim = dim3(2 * num_sms());
      q.blockDim = dim3(64 + 512);
      q.dynamicSmemBytes = kSmem;
      cudaLaunchAttribute a[1];
      a[0].id = cudaLaunchAttributeClusterDimension;
      a[0].val.clusterDim.x = 2;
      a[0].val.clusterDim.y = 1;
      a[0].val.clusterDim.z = 1;
      q.attrs = a;
      q.numAttrs = 1;
      OPB_CHECK_CUDA(cudaOccupancyMaxActiveClusters(&max_clusters, kern, &q));
      if (max_clusters < 1) max_clusters = 1;
    }
    max_clusters_a.store(max_clusters);
  }
  ConvPlan plan = plan_in;
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  int grid;
  if (CL2) {
    const int pairs_needed = (plan.num_tiles + 1) / 2;
    int pairs = pairs_needed < max_clusters ? pairs_needed : max_clusters;
    const int waves = (pairs_needed + pairs - 1) / pairs;
    pairs = (pairs_needed + waves - 1) / waves;
    grid = 2 * pairs;
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  } else {
    grid = num_sms();
    const int waves = (plan.num_tiles + grid - 1) / grid;  // fewest CTAs for the same number of tile waves
    grid = (plan.num_tiles + waves - 1) / waves;
  }
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(64 + 512);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = stream;
  if (grid <= 148) plan.trace = next_trace_slot(grid);
  OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, plan, ep));
  count_launch();
  return OPB_OK;
}

// Off by default: measured on B200 the convolution mainloops already run at the tensor pipe's practical rate
// (10.5 / 12.2 / 13 TFLOP/s per SM for BN = 128 / 208 / 256, i.e. at or above the cuBLAS-measured peak), so halving
// the weight fill changed nothing (2.145 vs 2.153 ms for the 23 convolutions of 8 frames).  Kept as a tested option
// (opb200_conv_cluster / OPB200_CONV_CLUSTER=1) for parts where the L2 -> SM fill is the bound.
static int g_conv_cluster = -1;
static bool conv_use_cluster() {
  if (g_conv_cluster < 0) {
    const char* e = std::getenv("OPB200_CONV_CLUSTER");
    g_conv_cluster = e ? (e[0] != '0') : 0;
  }
  return g_conv_cluster != 0;
}

template <int BN>
static int conv_dispatch(ConvPlan& plan, const ConvEpi& ep, const void* x, int Hi, int Wi, int cin_s, const void* w,
                         cudaStream_t stream) {
  // BN = 128: two sub-tiles still double-buffer in TMEM (4 x 128 columns).  BN = 208: two sub-tiles fill TMEM with one
  // accumulator set, so the epilogue runs behind the mainloop, but the halved weight refetch still wins (measured
  // 19.9 vs 24.2 us per 256 pixels).  BN = 256: one sub-tile, double-buffered (23.1 vs 25.1 us).
  const int tx = (plan.Wo + CONV_TW - 1) / CONV_TW;
  const int ty2 = (plan.Ho + 2 * CONV_TH - 1) / (2 * CONV_TH);
  const bool two = BN <= 208 && static_cast<long long>(plan.n_img) * tx * ty2 >= 2LL * num_sms();
  const int msub = two ? 2 : 1;
  plan.tiles_x = tx;
  plan.tiles_y = (plan.Ho + msub * CONV_TH - 1) / (msub * CONV_TH);
  plan.num_tiles = plan.n_img * plan.tiles_x * plan.tiles_y;
  // clusters of two CTAs sharing the weight slab pay off once every SM has tiles to walk in lockstep
  const bool cl2 = !ep.split && conv_use_cluster() && plan.num_tiles >= 2 * num_sms();
  const uint32_t s = static_cast<uint32_t>(plan.stride);
  const uint64_t cst = static_cast<uint64_t>(cin_s) * (plan.terms == 3 ? 2 : 1);  // channels per pixel in memory
  const uint64_t gdim[4] = {cst, static_cast<uint64_t>(Wi), static_cast<uint64_t>(Hi), static_cast<uint64_t>(plan.n_img)};
  const uint64_t gstr[3] = {cst * 2, cst * 2 * Wi, cst * 2 * Wi * Hi};
  const uint32_t box[4] = {BLOCK_K, CONV_TW * s, static_cast<uint32_t>(CONV_TH * msub) * s, 1};
  const uint32_t estr[4] = {1, s, s, 1};
  if (int r = encode_tmap_16bit(&plan.x_map, true, 4, x, gdim, gstr, box, estr)) return r;
  {
    const uint64_t wdim[2] = {cst, static_cast<uint64_t>(plan.ksize) * plan.ksize * BN};
    const uint64_t wstr[1] = {cst * 2};
    const uint32_t wbox[2] = {BLOCK_K, static_cast<uint32_t>(cl2 ? BN / 2 : BN)};
    const uint32_t westr[2] = {1, 1};
    if (int r = encode_tmap_16bit(&plan.w_map, true, 2, w, wdim, wstr, wbox, westr)) return r;
  }
  // 3x3, stride 1: one halo box per kernel column instead of one box per tap (OPB200_CONV_HALO=0: the per-tap kernels)
  static const bool halo_env = std::getenv("OPB200_CONV_HALO") ? std::getenv("OPB200_CONV_HALO")[0] != '0' : true;
  if (halo_env && plan.ksize == 3 && plan.stride == 1 && !cl2) {
    const uint32_t hbox[4] = {BLOCK_K, CONV_TW, static_cast<uint32_t>(CONV_TH * msub + 2), 1};
    if (int r = encode_tmap_16bit(&plan.xh_map, true, 4, x, gdim, gstr, hbox, estr)) return r;
    if (ep.split) return two ? launch_conv<BN, 2, false, true, true>(plan, ep, stream) : launch_conv<BN, 1, false, true, true>(plan, ep, stream);
    return two ? launch_conv<BN, 2, false, false, true>(plan, ep, stream) : launch_conv<BN, 1, false, false, true>(plan, ep, stream);
  }
  if (ep.split) return two ? launch_conv<BN, 2, false, true>(plan, ep, stream) : launch_conv<BN, 1, false, true>(plan, ep, stream);
  if (cl2) return two ? launch_conv<BN, 2, true>(plan, ep, stream) : launch_conv<BN, 1, true>(plan, ep, stream);
  return two ? launch_conv<BN, 2, false>(plan, ep, stream) : launch_conv<BN, 1, false>(plan, ep, stream);
}

}  // namespace opb

using namespace opb;

extern "C" {

int opb200_conv_cluster(int on) {
  const int prev = conv_use_cluster() ? 1 : 0;
  if (on >= 0) g_conv_cluster = on ? 1 : 0;
  return prev;
}

static int conv2d_impl(const void* x, int n_img, int Hi, int Wi, int cin_s, const void* w, int ksize, int stride, int cout_s,
                       const float* bias, const void* res, const void* up, void* y16, float* y32, int act, float slope,
                       int* range_flag, int split, float acc_scale, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(x && w && (y16 || y32), "null pointer");
  OPB_CHECK_ARG(n_img >= 1 && Hi >= 1 && Wi >= 1, "bad shape");
  OPB_CHECK_ARG(ksize == 1 || ksize == 3, "kernel size must be 1 or 3");
  OPB_CHECK_ARG(stride == 1 || stride == 2, "stride must be 1 or 2");
  OPB_CHECK_ARG(cin_s >= 16 && cin_s % 16 == 0, "input channel stride must be a multiple of 16");
  OPB_CHECK_ARG(cout_s == 128 || cout_s == 208 || cout_s == 256, "output channel stride must be 128, 208 or 256");
  OPB_CHECK_ARG(act >= 0 && act <= 2, "act must be 0 (none), 1 (ReLU) or 2 (LeakyReLU)");
  OPB_CHECK_ARG(((reinterpret_cast<uintptr_t>(y16) | reinterpret_cast<uintptr_t>(y32) | reinterpret_cast<uintptr_t>(res) |
                  reinterpret_cast<uintptr_t>(bias)) & 31) == 0, "bias / residual / outputs must be 32-byte aligned");
  ConvPlan plan;
  std::memset(&plan, 0, sizeof(plan));
  plan.n_img = n_img;
  plan.ksize = ksize;
  plan.pad = ksize / 2;
  plan.stride = stride;
  plan.Ho = (Hi + 2 * plan.pad - ksize) / stride + 1;
  plan.Wo = (Wi + 2 * plan.pad - ksize) / stride + 1;
  plan.nchunks = (cin_s + BLOCK_K - 1) / BLOCK_K;
  plan.nk_last = (cin_s - (plan.nchunks - 1) * BLOCK_K) / 16;
  plan.terms = split ? 3 : 1;
  plan.plane = cin_s;
  ConvEpi ep;
  ep.split = split ? 1 : 0;
  ep.acc_scale = acc_scale;
  ep.bias = bias;
  ep.res = static_cast<const __half*>(res);
  ep.y16 = static_cast<__half*>(y16);
  ep.y32 = y32;
  ep.act = act;
  ep.slope = slope;
  ep.up = static_cast<const __half*>(up);
  ep.up_ry = ep.up_rx = 0.f;
  ep.range_flag = range_flag;
  if (up != nullptr) {
    OPB_CHECK_ARG(plan.Ho % 2 == 0 && plan.Wo % 2 == 0, "fused x2 upsampling needs even output sides");
    OPB_CHECK_ARG((reinterpret_cast<uintptr_t>(up) & 31) == 0, "up must be 32-byte aligned");
    const int Hs = plan.Ho / 2, Ws = plan.Wo / 2;
    ep.up_ry = Hs > 1 ? static_cast<float>(Hs - 1) / static_cast<float>(plan.Ho - 1) : 0.f;
    ep.up_rx = Ws > 1 ? static_cast<float>(Ws - 1) / static_cast<float>(plan.Wo - 1) : 0.f;
  }
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  switch (cout_s) {
    case 128: return conv_dispatch<128>(plan, ep, x, Hi, Wi, cin_s, w, stream);
    case 208: return conv_dispatch<208>(plan, ep, x, Hi, Wi, cin_s, w, stream);
    default: return conv_dispatch<256>(plan, ep, x, Hi, Wi, cin_s, w, stream);
  }
}

int opb200_conv2d_nhwc_f16(const void* x, int n_img, int Hi, int Wi, int cin_s, const void* w, int ksize, int stride, int cout_s,
                           const float* bias, const void* res, const void* up, void* y16, float* y32, int act, float slope,
                           int* range_flag, void* stream) {
  return conv2d_impl(x, n_img, Hi, Wi, cin_s, w, ksize, stride, cout_s, bias, res, up, y16, y32, act, slope, range_flag, 0, 1.f, stream);
}

int opb200_conv2d_nhwc_f16x3(const void* x, int n_img, int Hi, int Wi, int cin_s, const void* w, int ksize, int stride, int cout_s,
                             const float* bias, const void* res, const void* up, void* y16, float* y32, int act, float slope,
                             float acc_scale, int* range_flag, void* stream) {
  OPB_CHECK_ARG(acc_scale > 0.f, "acc_scale must be positive");
  return conv2d_impl(x, n_img, Hi, Wi, cin_s, w, ksize, stride, cout_s, bias, res, up, y16, y32, act, slope, range_flag, 1, acc_scale,
                     stream);
}

}  // extern "C"
