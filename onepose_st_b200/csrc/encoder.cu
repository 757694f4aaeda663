// LoFTR encoder layer stack (elu+1 linear attention) on tcgen05 tensor cores.
//
// Replaces, for inference:
//   LoFTREncoderLayer.forward       src/models/OnePosePlus/loftr_module/transformer.py:65-94
//                                   submodules/LoFTR/src/loftr/loftr_module/transformer.py:35-58
//   LinearAttention.forward         src/models/OnePosePlus/loftr_module/linear_attention.py:29-61
//   LocalFeatureTransformer.forward transformer.py:133-171 (OnePosePlus: both streams read pre-update
//                                   tensors) and LoFTR transformer.py:80-101 (sequential cross update)
//
// Data layout ("row space"): every token of both streams of every batch element is one row of a
// [R, C] matrix.  Stream 0 occupies rows [0, N*P0), stream 1 rows [N*P0, N*(P0+P1)); with big groups
// (coarse level) each (stream, batch) group is padded to a multiple of 128 rows so a 128-row MMA tile
// never straddles two groups (the attention B operand differs per group).  For every activation that
// feeds a tensor-core product the row space holds two bf16 planes (hi = rn(x), lo = rn(x - hi)); the
// product A*B is evaluated as A_hi*B_lo + A_lo*B_hi + A_hi*B_hi with fp32 accumulation in TMEM, which
// keeps ~2^-17 relative operand precision (fp32-grade results from the bf16 tensor pipe).
//
// One layer step = 7 launches, all persistent tcgen05 GEMMs except the two small KV kernels:
//   1 QKV      [R,C]x[3C,C]^T   epilogue: phi=elu+1 on q,k; row mask; q -> planes, k,v -> fp32
//   2 KV       per group/head   sum_s phi(K)^T V and sum_s phi(K)            (CUDA cores, fp32)
//   3 KVpack   block-diagonal B operand [KV | Ksum] as planes
//   4 ATTN     phi(Q) x blockdiag(KV)  epilogue: divide by (phi(Q).Ksum + eps)  -> planes
//   5 MERGE    x Wm^T           epilogue: LayerNorm1 -> planes
//   6 MLP1     [x|msg] x W1^T   epilogue: ReLU -> planes
//   7 MLP2     x W2^T           epilogue: LayerNorm2, + residual -> planes of x (in place; x = hi + lo is the master)
#include <cmath>
#include <cuda_pipeline.h>
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

int split_planes(const float* x, long long stride_r, long long stride_c, const float* add, long long add_period_rows,
                 int rows, int cols, __nv_bfloat16* hi, __nv_bfloat16* lo, float* f32out, long long ld_out,
                 cudaStream_t stream);

constexpr int EPI_SPLIT = 4;  // 16 epilogue warps: every thread owns one row and BN/4 of its columns

__device__ __forceinline__ float elu1(float x) {
  // elu(x) + 1 = x + 1 (x > 0) | exp(x) (x <= 0); ex2.approx keeps ~2^-22 relative error
  return x > 0.f ? x + 1.0f : ptx::ex2(x * 1.4426950408889634f);
}

// ------------------------------------------------------------------------------------------------
// Epilogues.  Thread = (row of the tile, column slice `part`); values come out of TMEM 32 columns at a
// time; every global store is a full 32-byte sector per lane (STG.256).
// ------------------------------------------------------------------------------------------------
template <int C>
struct EpiQKV {
  struct Params {
    __nv_bfloat16 *q_hi, *q_lo;  // phi(Q) planes [R,C] (big groups)
    float* q_f32;                // phi(Q) fp32 [R,C]   (small groups)
    float *k_f32, *v_f32;        // phi(K), V fp32 [R,C]
    const uint8_t* row_mask;     // [R], 1 = valid
    __nv_bfloat16 *k_hi, *k_lo, *v_hi, *v_lo;  // big groups, tensor-core KV: phi(K), V as bf16 planes [R,C] instead of fp32
  };
  static constexpr int kSmemBytes = 0;
  static constexpr int CPT = C / EPI_SPLIT;
  static __device__ __forceinline__ void prefetch(const Params&, const EpiCtx&) {}
  static __device__ __forceinline__ void run(const Params& p, const EpiCtx& ctx) {
    const long long row = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32 + ctx.lane;
    const float m = p.row_mask[row] ? 1.f : 0.f;
#pragma unroll 1
    for (int c0 = ctx.part * CPT; c0 < (ctx.part + 1) * CPT; c0 += 32) {
      float v[32];
      load_acc32(ctx.tmem_acc + c0, v);
      if (ctx.n_tile == 2) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] *= m;
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = elu1(v[i]) * m;
      }
      if (ctx.n_tile == 0 && p.q_f32 == nullptr) {
        store_planes32_v8(v, p.q_hi + row * C + c0, p.q_lo ? p.q_lo + row * C + c0 : nullptr);
      } else if (ctx.n_tile != 0 && p.k_hi != nullptr) {
        if (ctx.n_tile == 1) store_planes32_v8(v, p.k_hi + row * C + c0, p.k_lo ? p.k_lo + row * C + c0 : nullptr);
        else store_planes32_v8(v, p.v_hi + row * C + c0, p.v_lo ? p.v_lo + row * C + c0 : nullptr);
      } else {
        store_f32x32_v8(v, (ctx.n_tile == 0 ? p.q_f32 : ctx.n_tile == 1 ? p.k_f32 : p.v_f32) + row * C + c0);
      }
    }
  }
};

// ATTN: tile = 128 rows x [128 value columns of 4 heads | 16 denominator columns], D = 32; part = head
struct EpiAttn {
  struct Params {
    __nv_bfloat16 *o_hi, *o_lo;  // [R, 256]
    float eps;
  };
  static constexpr int kSmemBytes = 0;
  static __device__ __forceinline__ void prefetch(const Params&, const EpiCtx&) {}
  static __device__ __forceinline__ void run(const Params& p, const EpiCtx& ctx) {
    const long long row = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32 + ctx.lane;
    uint32_t dr[16];
    ptx::tmem_ld16(ctx.tmem_acc + 128, dr);
    ptx::tmem_ld_wait();
    const int h = ctx.part;
    float den = __uint_as_float(dr[0]);
    den = h == 1 ? __uint_as_float(dr[1]) : den;
    den = h == 2 ? __uint_as_float(dr[2]) : den;
    den = h == 3 ? __uint_as_float(dr[3]) : den;
    const float z = 1.0f / (den + p.eps);
    float v[32];
    load_acc32(ctx.tmem_acc + h * 32, v);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= z;
    const long long off = row * 256 + ctx.n_tile * 128 + h * 32;
    store_planes32_v8(v, p.o_hi + off, p.o_lo ? p.o_lo + off : nullptr);
  }
};

// Row LayerNorm statistics with the row's columns spread over EPI_SPLIT warps: each thread keeps its CPT
// values in registers (one TMEM read), partial sums meet in shared memory behind a per-quarter named
// barrier; two-pass (mean, then centred squares) like torch.
template <int CPT>
__device__ __forceinline__ void row_mean_rstd_split(const EpiCtx& ctx, const float (&x)[CPT], int C, float& mean, float& rstd) {
  float* red = reinterpret_cast<float*>(ctx.scratch);  // [2][4 quarters][EPI_SPLIT][32]
  const int bar_id = 1 + ctx.group * 4 + ctx.quarter;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
  for (int i = 0; i < CPT; i += 4) {
    s0 += x[i];
    s1 += x[i + 1];
    s2 += x[i + 2];
    s3 += x[i + 3];
  }
  float* slot = red + (ctx.quarter * EPI_SPLIT) * 32 + ctx.lane;
  if (ctx.scratch_small) ptx::named_bar_sync(bar_id, 32 * EPI_SPLIT);  // the previous call's squares have been read
  slot[ctx.part * 32] = (s0 + s1) + (s2 + s3);
  ptx::named_bar_sync(bar_id, 32 * EPI_SPLIT);
  float tot = 0.f;
#pragma unroll
  for (int q = 0; q < EPI_SPLIT; ++q) tot += slot[q * 32];
  mean = tot / static_cast<float>(C);
  float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
  for (int i = 0; i < CPT; i += 4) {
    const float d0 = x[i] - mean, d1 = x[i + 1] - mean, d2 = x[i + 2] - mean, d3 = x[i + 3] - mean;
    q0 = fmaf(d0, d0, q0);
    q1 = fmaf(d1, d1, q1);
    q2 = fmaf(d2, d2, q2);
    q3 = fmaf(d3, d3, q3);
  }
  float* slot2 = ctx.scratch_small ? slot : slot + 4 * EPI_SPLIT * 32;
  if (ctx.scratch_small) ptx::named_bar_sync(bar_id, 32 * EPI_SPLIT);  // the sums have been read: the slots take the squares
  slot2[ctx.part * 32] = (q0 + q1) + (q2 + q3);
  ptx::named_bar_sync(bar_id, 32 * EPI_SPLIT);
  float var = 0.f;
#pragma unroll
  for (int q = 0; q < EPI_SPLIT; ++q) var += slot2[q * 32];
  rstd = rsqrtf(var / static_cast<float>(C) + 1e-5f);
}

constexpr int kLnScratch = 2 * 4 * EPI_SPLIT * 32 * 4;

template <int CPT>
__device__ __forceinline__ void load_row_slice(const EpiCtx& ctx, float (&x)[CPT]) {
#pragma unroll
  for (int c = 0; c < CPT; c += 32) {
    uint32_t r[32];
    ptx::tmem_ld32(ctx.tmem_acc + ctx.part * CPT + c, r);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) x[c + i] = __uint_as_float(r[i]);
  }
}

template <int C>
struct EpiLN {  // MERGE epilogue: LayerNorm1 -> planes
  struct Params {
    const float *gamma, *beta;
    __nv_bfloat16 *o_hi, *o_lo;
  };
  static constexpr int kSmemBytes = kLnScratch;
  static constexpr int CPT = C / EPI_SPLIT;
  static __device__ __forceinline__ void prefetch(const Params&, const EpiCtx&) {}
  static __device__ __forceinline__ void run(const Params& p, const EpiCtx& ctx) {
    const long long row = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32 + ctx.lane;
    float x[CPT];
    load_row_slice<CPT>(ctx, x);
    float mean, rstd;
    row_mean_rstd_split<CPT>(ctx, x, C, mean, rstd);
    const int cb = ctx.part * CPT;
#pragma unroll
    for (int i = 0; i < CPT; i += 4) {  // warp-uniform addresses: one broadcast 16-byte load per 4 columns
      const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.gamma + cb + i));
      const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.beta + cb + i));
      x[i + 0] = (x[i + 0] - mean) * rstd * g4.x + b4.x;
      x[i + 1] = (x[i + 1] - mean) * rstd * g4.y + b4.y;
      x[i + 2] = (x[i + 2] - mean) * rstd * g4.z + b4.z;
      x[i + 3] = (x[i + 3] - mean) * rstd * g4.w + b4.w;
    }
    const long long wrow0 = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32;
    if constexpr (CPT == 64) {
      uint32_t h[4][8], l[4][8];
      pack_planes64(x, h, l);
      store_row64_bf16_quad(h, p.o_hi + wrow0 * C + cb, C, ctx.lane);
      if (p.o_lo) store_row64_bf16_quad(l, p.o_lo + wrow0 * C + cb, C, ctx.lane);
    } else {
#pragma unroll
      for (int c = 0; c < CPT; c += 32) {
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = x[c + i];
        store_planes32_v8(v, p.o_hi + row * C + cb + c, p.o_lo ? p.o_lo + row * C + cb + c : nullptr);
      }
    }
  }
};

template <int BN>
struct EpiRelu {  // MLP1 epilogue: ReLU -> planes [R, ld]
  struct Params {
    __nv_bfloat16 *o_hi, *o_lo;
    int ld;
  };
  static constexpr int kSmemBytes = 0;
  static constexpr int CPT = BN / EPI_SPLIT;
  static __device__ __forceinline__ void prefetch(const Params&, const EpiCtx&) {}
  static __device__ __forceinline__ void run(const Params& p, const EpiCtx& ctx) {
    const long long row = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32 + ctx.lane;
#pragma unroll 1
    for (int c0 = ctx.part * CPT; c0 < (ctx.part + 1) * CPT; c0 += 32) {
      float v[32];
      load_acc32(ctx.tmem_acc + c0, v);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
      const long long off = row * p.ld + ctx.n_tile * BN + c0;
      store_planes32_v8(v, p.o_hi + off, p.o_lo ? p.o_lo + off : nullptr);
    }
  }
};

template <int C>
struct EpiLNRes {  // MLP2 epilogue: x <- x + LayerNorm2(acc).  The residual stream lives in its two bf16 planes
  struct Params {   // (x = hi + lo, ~2^-17 relative): no separate fp32 master copy is read or written.
    const float *gamma, *beta;
    __nv_bfloat16 *x_hi, *x_lo;
  };
  static constexpr int kSmemBytes = kLnScratch;
  static constexpr int CPT = C / EPI_SPLIT;
  static __device__ __forceinline__ void prefetch(const Params& p, const EpiCtx& ctx) {
    // the residual planes of this thread's row slice (2 x CPT bf16 = 2 x 128 B at C = 256) come from HBM: start early
    const long long off = (static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32 + ctx.lane) * C + ctx.part * (C / EPI_SPLIT);
    ptx::prefetch_l2(p.x_hi + off);
    ptx::prefetch_l2(p.x_lo + off);
  }
  static __device__ __forceinline__ void run(const Params& p, const EpiCtx& ctx) {
    const long long row = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32 + ctx.lane;
    float x[CPT];
    load_row_slice<CPT>(ctx, x);
    float mean, rstd;
    row_mean_rstd_split<CPT>(ctx, x, C, mean, rstd);
    const int cb = ctx.part * CPT;
#pragma unroll
    for (int i = 0; i < CPT; i += 4) {
      const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.gamma + cb + i));
      const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.beta + cb + i));
      x[i + 0] = (x[i + 0] - mean) * rstd * g4.x + b4.x;
      x[i + 1] = (x[i + 1] - mean) * rstd * g4.y + b4.y;
      x[i + 2] = (x[i + 2] - mean) * rstd * g4.z + b4.z;
      x[i + 3] = (x[i + 3] - mean) * rstd * g4.w + b4.w;
    }
    if constexpr (CPT == 64) {  // residual planes in and out through the quad transpose (one line per quad and instruction)
      const long long wrow0 = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32;
      __nv_bfloat16* ph = p.x_hi + wrow0 * C + cb;
      __nv_bfloat16* pl = p.x_lo + wrow0 * C + cb;
      uint32_t h[4][8], l[4][8];
      // one plane at a time (x[64] + two planes in flight would not fit the 112 registers of a 576-thread CTA); the small
      // lo plane is added first so that the sum is (lo + LN) + hi
      load_row64_bf16_quad(l, pl, C, ctx.lane);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float2 lf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&l[i >> 3][i & 7]));
        x[2 * i] += lf.x;
        x[2 * i + 1] += lf.y;
      }
      load_row64_bf16_quad(h, ph, C, ctx.lane);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float2 hf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h[i >> 3][i & 7]));
        x[2 * i] += hf.x;
        x[2 * i + 1] += hf.y;
      }
      pack_planes64(x, h, l);
      store_row64_bf16_quad(h, ph, C, ctx.lane);
      store_row64_bf16_quad(l, pl, C, ctx.lane);
    } else {
#pragma unroll
      for (int c = 0; c < CPT; c += 32) {
        __nv_bfloat16* ph = p.x_hi + row * C + cb + c;
        __nv_bfloat16* pl = p.x_lo + row * C + cb + c;
        float v[32];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          uint32_t h[8], l[8];
          ptx::ld_global_v8(ph + 16 * i, h);
          ptx::ld_global_v8(pl + 16 * i, l);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float2 hf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h[k]));
            const float2 lf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&l[k]));
            v[16 * i + 2 * k] = (hf.x + lf.x) + x[c + 16 * i + 2 * k];
            v[16 * i + 2 * k + 1] = (hf.y + lf.y) + x[c + 16 * i + 2 * k + 1];
          }
        }
        store_planes32_v8(v, ph, pl);
      }
    }
  }
};

// ------------------------------------------------------------------------------------------------
// KV reduction for big groups (coarse level: C = 256, H = 8, D = 32)
// ------------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------------
// Fused MLP of a coarse encoder layer (C = 256, split precision):
//   x <- x + LayerNorm2( relu([x | msg] W1^T) W2^T )
// in ONE launch: the [128 x 512] hidden tile of an m tile never leaves the SM (as separate GEMMs its bf16 hi / lo planes make
// a 1.6 GB round trip through HBM per layer step at 32 frames).
//   * Per m tile and 256-wide half of the hidden tile: phase 1 = [x | msg] x W1 rows of the half into D1 (three split
//     products); the 16 epilogue warps read D1, apply the ReLU, split to bf16 hi / lo and write the planes back INTO THE SAME
//     TENSOR-MEMORY COLUMNS, one 128-column quarter at a time; phase 2 = a tcgen05.mma whose A operand is read from tensor
//     memory, x W2 columns of the half, accumulated into D2.  After the last half the warps run the LayerNorm2 + residual
//     epilogue of the unfused path (EpiLNRes) on D2.  (Details at the kernel.)
//   * One TMA thread streams {A planes, W1 rows} / {W2 columns} through four 48 KB stages in exactly the order the MMA
//     thread consumes them.
// History (32 frames per forward, ncu, per layer step; the two GEMM launches: 415 + 318 us): hidden tile in shared memory as
// 128 KB halves, two 48 KB stages beside it, every stage waiting for the latency of its own TMA load: 995 us.  Operand in
// separate tensor-memory columns per 128-wide quarter, seven 32 KB stages, [x | msg] read four times per tile: 823 us.
// In place per 256-wide half (this version): 753-771 us, tensor pipe 65 % active.  In the bench, where the clocks are set
// by the power cap and the 1.6 GB that no longer move count: transformer256 0.314-0.317 ms per frame against 0.338-0.340;
// 8 / 16 frames per forward (second version): 0.363 / 0.342 against 0.376 / 0.351.
constexpr int kMlpFusedMinRounds = 4;  // tiles per CTA from which the fused kernels are the default (OPB200_MLP_FUSED / OPB200_ATTN_MERGE_FUSED override)
constexpr int MF_STAGES = 4;
constexpr int MF_APL = BLOCK_M * 32 * 2;     // one A plane of a 32-wide k-block: 8 KB
constexpr int MF_BPL = 256 * 32 * 2;         // one plane of a 256-row weight slab: 16 KB (phase 2 at C = 128: 128 rows)
constexpr int MF_STAGE_BYTES = 2 * MF_APL + 2 * MF_BPL;  // 48 KB (phase 2 uses the weight part only)
constexpr int MF_LN_SCRATCH = kLnScratch / 2;  // (EpiCtx::scratch_small)
// 64-byte swizzle: 512-byte alignment of the operand tiles is enough
constexpr int MF_SMEM = 512 + MF_STAGES * MF_STAGE_BYTES + 256 + MF_LN_SCRATCH;
static_assert(MF_SMEM <= 232448, "shared memory of a CTA");

template <int C>
struct MlpFusedParams {
  CUtensorMap x_hi, x_lo, m_hi, m_lo;  // [R, C] planes, box {32, 128}, 64-byte swizzle
  CUtensorMap w1_hi, w1_lo;            // [2C, 2C], box {32, 256}
  CUtensorMap w2_hi, w2_lo;            // [C, 2C], box {32, C}
  typename EpiLNRes<C>::Params ln;
  int m_tile_begin, num_tiles;
  // phase 0 (C = 128 only, 128 columns of tensor memory are free there): msg = LayerNorm1(attn_out Wm^T) is computed in
  // the kernel and handed to phase 1 IN TENSOR MEMORY as the A operand of the msg half of K; m_hi / m_lo are not read
  int phase0;
  CUtensorMap ao_hi, ao_lo;  // [R, C] attention output planes, box {32, 128}
  CUtensorMap wm_hi, wm_lo;  // [C, C], box {32, C}
  const float *ln1_g, *ln1_b;
};

// Tensor memory: D1 = columns 0..255 (one 256-wide half of the hidden tile, fp32), D2 = columns 256..511 (the output).
// The epilogue warps turn D1 IN PLACE into the A operand of phase 2, one 128-column quarter at a time: quarter a
// (columns 0..127) becomes bf16 hi (64 columns) | lo (64 columns) of its 128 values, quarter b (128..255) likewise, and
// phase 2 is a tcgen05.mma that reads A from tensor memory -- all of shared memory is left to the operand ring.
//   per half u:  G1(u): 16 k-blocks [x | msg] x W1[256 u ..] -> D1 (N = 256)
//                e1(a) -> G2(a): 4 k-blocks A2a x W2[:, 256 u ..] -> D2;   e1(b) runs under G2(a);   G2(b)
//   G1(u + 1) overwrites D1 after G2(b) -- the pipe executes a thread's MMAs in order -- so the only time the tensor pipe
//   waits for the warps is e1(a) (~600 clk per ~18 500 clk half); LayerNorm2 of a tile runs under the next tile's G1.
// L2 -> SM traffic per tile (C = 256): [x | msg] twice (0.5 MB) + W1 (1 MB) + W2 (0.5 MB).  C = 128 (fine level): the hidden
// tile is one 256-wide "half", D2 has 128 columns.
template <int C>
__global__ void __launch_bounds__(64 + 128 * EPI_SPLIT, 1) mlp_fused_kernel(const __grid_constant__ MlpFusedParams<C> p) {
  constexpr int HALVES = 2 * C / 256;  // 256-wide parts of the hidden tile
  constexpr int KB1 = 2 * C / 32;      // k-blocks of phase 1; the first C / 32 come from x, the rest from msg
  constexpr int B2PL = C * 32 * 2;     // one W2 plane of a k-block
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 511) & ~uintptr_t(511));
  uint8_t* stages = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + MF_STAGES * MF_STAGE_BYTES);
  uint64_t* empty = full + MF_STAGES;
  uint64_t* d1_full = empty + MF_STAGES;
  uint64_t* a2_full = d1_full + 1;  // [2]: quarters a, b
  uint64_t* d2_full = a2_full + 2;
  uint64_t* d2_empty = d2_full + 1;
  uint64_t* d0_full = d2_empty + 1;
  uint64_t* m_full = d0_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(m_full + 1);
  uint8_t* ln_scratch = smem + MF_STAGES * MF_STAGE_BYTES + 256;
  static_assert((2 * MF_STAGES + 7) * 8 + 4 <= 256, "barrier block");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kEpiWarps = 4 * EPI_SPLIT;
  constexpr uint32_t kD1 = 0, kD2 = 256, kD0 = 384;  // tensor-memory columns (D0 / msg operand: phase 0, C = 128)
  constexpr int B0PL = C * 32 * 2;  // one Wm plane of a k-block (phase 0)
  const bool phase0 = C == 128 && p.phase0 != 0;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.x_hi);
    ptx::prefetch_tmap(&p.w1_hi);
    ptx::prefetch_tmap(&p.w2_hi);
    for (int s = 0; s < MF_STAGES; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    ptx::mbar_init(d1_full, 1);
    ptx::mbar_init(&a2_full[0], kEpiWarps);
    ptx::mbar_init(&a2_full[1], kEpiWarps);
    ptx::mbar_init(d2_full, 1);
    ptx::mbar_init(d2_empty, kEpiWarps);
    ptx::mbar_init(d0_full, 1);
    ptx::mbar_init(m_full, kEpiWarps);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_it = (p.num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int U = HALVES * n_it;  // halves of this CTA

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (same order as the MMA thread)
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      auto next = [&]() {
        if (++stage == MF_STAGES) stage = 0, phase ^= 1;
      };
      for (int u = 0; u < U; ++u) {
        const int row = (p.m_tile_begin + static_cast<int>(blockIdx.x) + (u / HALVES) * static_cast<int>(gridDim.x)) * BLOCK_M;
        const int half = u % HALVES;
        if (phase0) {
          for (int kb = 0; kb < C / 32; ++kb) {  // phase 0: the attention output and Wm
            ptx::mbar_wait(&empty[stage], phase ^ 1);
            ptx::mbar_arrive_expect_tx(&full[stage], 2 * MF_APL + 2 * B0PL);
            uint8_t* st = stages + stage * MF_STAGE_BYTES;
            ptx::tma_load_2d(st, &p.ao_hi, &full[stage], kb * 32, row);
            ptx::tma_load_2d(st + MF_APL, &p.ao_lo, &full[stage], kb * 32, row);
            ptx::tma_load_2d(st + 2 * MF_APL, &p.wm_hi, &full[stage], kb * 32, 0);
            ptx::tma_load_2d(st + 2 * MF_APL + B0PL, &p.wm_lo, &full[stage], kb * 32, 0);
            next();
          }
        }
        for (int kb = 0; kb < KB1; ++kb) {  // phase 1: [x | msg] and the W1 slab of this half
          ptx::mbar_wait(&empty[stage], phase ^ 1);
          const bool a_tm = phase0 && kb >= C / 32;  // the msg half of K comes from tensor memory
          ptx::mbar_arrive_expect_tx(&full[stage], a_tm ? 2 * MF_BPL : MF_STAGE_BYTES);
          uint8_t* st = stages + stage * MF_STAGE_BYTES;
          const int ka = (kb % (C / 32)) * 32;
          if (!a_tm) {
            ptx::tma_load_2d(st, kb < C / 32 ? &p.x_hi : &p.m_hi, &full[stage], ka, row);
            ptx::tma_load_2d(st + MF_APL, kb < C / 32 ? &p.x_lo : &p.m_lo, &full[stage], ka, row);
          }
          ptx::tma_load_2d(st + 2 * MF_APL, &p.w1_hi, &full[stage], kb * 32, half * 256);
          ptx::tma_load_2d(st + 2 * MF_APL + MF_BPL, &p.w1_lo, &full[stage], kb * 32, half * 256);
          next();
        }
        for (int kb = 0; kb < 8; ++kb) {  // phase 2: the W2 columns that meet this half (quarter a: kb 0..3, b: 4..7)
          ptx::mbar_wait(&empty[stage], phase ^ 1);
          ptx::mbar_arrive_expect_tx(&full[stage], 2 * B2PL);
          uint8_t* st = stages + stage * MF_STAGE_BYTES;
          ptx::tma_load_2d(st + 2 * MF_APL, &p.w2_hi, &full[stage], half * 256 + kb * 32, 0);
          ptx::tma_load_2d(st + 2 * MF_APL + B2PL, &p.w2_lo, &full[stage], half * 256 + kb * 32, 0);
          next();
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(BLOCK_M, 256), idesc2 = ptx::make_idesc_bf16(BLOCK_M, C);
      const uint32_t d1 = tmem_base + kD1, d2 = tmem_base + kD2;
      int stage = 0;
      uint32_t phase = 0;
      auto next = [&]() {
        if (++stage == MF_STAGES) stage = 0, phase ^= 1;
      };
      for (int u = 0; u < U; ++u) {
        const int half = u % HALVES;
        const uint32_t u1 = static_cast<uint32_t>(u);
        if (phase0) {
          // phase 0: D0 = attn_out Wm^T (N = C); the warps normalise it under the x half of phase 1.  (D0 is free: the MMAs
          // of the previous tile that read the msg operand from these columns were issued before these.)
          constexpr uint32_t idesc0 = ptx::make_idesc_bf16(BLOCK_M, C);
          const uint32_t d0 = tmem_base + kD0;
          for (int kb = 0; kb < C / 32; ++kb) {
            ptx::mbar_wait(&full[stage], phase);
            ptx::tc_fence_after();
            const uint32_t sa = ptx::smem_u32(stages + stage * MF_STAGE_BYTES);
            const uint64_t a_hi = ptx::make_desc_sw64(sa), a_lo = ptx::make_desc_sw64(sa + MF_APL);
            const uint64_t b_hi = ptx::make_desc_sw64(sa + 2 * MF_APL), b_lo = ptx::make_desc_sw64(sa + 2 * MF_APL + B0PL);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d0, a_hi + 2 * k, b_hi + 2 * k, idesc0, !(kb == 0 && k == 0));
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d0, a_lo + 2 * k, b_hi + 2 * k, idesc0, 1);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d0, a_hi + 2 * k, b_lo + 2 * k, idesc0, 1);
            ptx::mma_commit(&empty[stage]);
            next();
          }
          ptx::mma_commit(d0_full);
        }
        // (D1 is free: G2 of the previous half was issued after the warps had read and rewritten it, and runs before this)
        for (int kb = 0; kb < KB1; ++kb) {
          const bool a_tm = phase0 && kb >= C / 32;
          if (a_tm && kb == C / 32) {
            ptx::mbar_wait(m_full, u1 & 1);  // LayerNorm1(D0) is in tensor memory as bf16 planes
            ptx::tc_fence_after();
          }
          ptx::mbar_wait(&full[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(stages + stage * MF_STAGE_BYTES);
          const uint64_t b_hi = ptx::make_desc_sw64(sa + 2 * MF_APL), b_lo = ptx::make_desc_sw64(sa + 2 * MF_APL + MF_BPL);
          if (a_tm) {
            const uint32_t a_hi = tmem_base + kD0 + 16 * (kb - C / 32), a_lo = a_hi + 64;
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(d1, a_hi + 8 * k, b_hi + 2 * k, idesc, 1);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(d1, a_lo + 8 * k, b_hi + 2 * k, idesc, 1);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(d1, a_hi + 8 * k, b_lo + 2 * k, idesc, 1);
          } else {
            const uint64_t a_hi = ptx::make_desc_sw64(sa), a_lo = ptx::make_desc_sw64(sa + MF_APL);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d1, a_hi + 2 * k, b_hi + 2 * k, idesc, !(kb == 0 && k == 0));
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d1, a_lo + 2 * k, b_hi + 2 * k, idesc, 1);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d1, a_hi + 2 * k, b_lo + 2 * k, idesc, 1);
          }
          ptx::mma_commit(&empty[stage]);
          next();
        }
        ptx::mma_commit(d1_full);
        if (half == 0) ptx::mbar_wait(d2_empty, (static_cast<uint32_t>(u / HALVES) & 1) ^ 1);  // the previous tile's output has been read
        for (int kb = 0; kb < 8; ++kb) {
          if ((kb & 3) == 0) {
            ptx::mbar_wait(&a2_full[kb >> 2], u1 & 1);  // this quarter of the hidden tile is in tensor memory as bf16 planes
            ptx::tc_fence_after();
          }
          ptx::mbar_wait(&full[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(stages + stage * MF_STAGE_BYTES);
          const uint64_t b_hi = ptx::make_desc_sw64(sa + 2 * MF_APL), b_lo = ptx::make_desc_sw64(sa + 2 * MF_APL + B2PL);
          const uint32_t a_hi = tmem_base + kD1 + 128 * (kb >> 2) + 16 * (kb & 3), a_lo = a_hi + 64;
#pragma unroll
          for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(d2, a_hi + 8 * k, b_hi + 2 * k, idesc2, !(half == 0 && kb == 0 && k == 0));
#pragma unroll
          for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(d2, a_lo + 8 * k, b_hi + 2 * k, idesc2, 1);
#pragma unroll
          for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(d2, a_hi + 8 * k, b_lo + 2 * k, idesc2, 1);
          ptx::mma_commit(&empty[stage]);
          next();
        }
        if (half == HALVES - 1) ptx::mma_commit(d2_full);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps: (column part, TMEM lane quarter)
    EpiCtx ctx;
    ctx.quarter = warp & 3;
    ctx.lane = lane;
    ctx.epi_warp = (warp - 2) & 3;
    ctx.group = 0;
    ctx.part = (warp - 2) >> 2;
    ctx.n_tile = 0;
    ctx.scratch = ln_scratch;
    ctx.scratch_small = 1;
    const uint32_t lane_field = static_cast<uint32_t>(ctx.quarter * 32) << 16;
    const int bar_id = 1 + ctx.quarter;  // the four warps (column parts) of a TMEM lane quarter
    for (int u = 0; u < U; ++u) {
      const uint32_t u1 = static_cast<uint32_t>(u);
      if constexpr (C == 128) {
        if (phase0) {  // msg = LayerNorm1(D0): this warp's 32 columns of its rows, back into D0's columns as bf16 hi | lo planes
          ptx::mbar_wait(d0_full, u1 & 1);
          ptx::tc_fence_after();
          const uint32_t base0 = tmem_base + kD0 + lane_field;
          float x[32];
          load_acc32(base0 + ctx.part * 32, x);
          float mean, rstd;
          row_mean_rstd_split<32>(ctx, x, C, mean, rstd);  // (its barriers: every column of these lanes has been read)
          uint32_t h[16], l[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int c0 = ctx.part * 32 + 2 * i;
            const float y0 = (x[2 * i] - mean) * rstd * __ldg(p.ln1_g + c0) + __ldg(p.ln1_b + c0);
            const float y1 = (x[2 * i + 1] - mean) * rstd * __ldg(p.ln1_g + c0 + 1) + __ldg(p.ln1_b + c0 + 1);
            __nv_bfloat16 h0, l0, h1, l1;
            split_bf16(y0, h0, l0);
            split_bf16(y1, h1, l1);
            h[i] = pack_bf16(h0, h1);
            l[i] = pack_bf16(l0, l1);
          }
          ptx::tmem_st16(base0 + ctx.part * 16, h);
          ptx::tmem_st16(base0 + 64 + ctx.part * 16, l);
          ptx::tmem_st_wait();
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(m_full);
        }
      }
      ptx::mbar_wait(d1_full, u1 & 1);
      ptx::tc_fence_after();
#pragma unroll 1
      for (int qq = 0; qq < 2; ++qq) {
        // this warp's 32 columns of the quarter -> ReLU -> bf16 hi / lo; the planes go back into the quarter's own columns
        // once the four warps that share these lanes have all read theirs
        const uint32_t qbase = tmem_base + kD1 + lane_field + 128 * qq;
        uint32_t h[16], l[16];
        {
          float x[32];
          load_acc32(qbase + ctx.part * 32, x);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            __nv_bfloat16 h0, l0, h1, l1;
            split_bf16(fmaxf(x[2 * i], 0.f), h0, l0);
            split_bf16(fmaxf(x[2 * i + 1], 0.f), h1, l1);
            h[i] = pack_bf16(h0, h1);
            l[i] = pack_bf16(l0, l1);
          }
        }
        ptx::named_bar_sync(bar_id, 32 * EPI_SPLIT);
        ptx::tmem_st16(qbase + ctx.part * 16, h);
        ptx::tmem_st16(qbase + 64 + ctx.part * 16, l);
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&a2_full[qq]);
      }
      if (u % HALVES == HALVES - 1) {  // LayerNorm2 + residual of this tile (its last phase 2 runs now; the next tile's phase 1 follows it)
        const int it = u / HALVES;
        ctx.m_tile = p.m_tile_begin + static_cast<int>(blockIdx.x) + it * static_cast<int>(gridDim.x);
        EpiLNRes<C>::prefetch(p.ln, ctx);
        ptx::mbar_wait(d2_full, static_cast<uint32_t>(it) & 1);
        ptx::tc_fence_after();
        ctx.tmem_acc = tmem_base + kD2 + lane_field;
        EpiLNRes<C>::run(p.ln, ctx);
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(d2_empty);
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------
// ATTN + MERGE + LayerNorm1 of a coarse encoder layer (C = 256, split precision) in ONE launch:
//   m <- LayerNorm1( (phi(Q) blockdiag(KV) / (phi(Q) Ksum + eps)) Wm^T )
// The attention output of a 128-row tile never leaves the SM: per half of the heads (4 heads = 128 value columns + their
// denominators) the product phi(Q) KV is accumulated into D_attn (tensor-memory columns 0..143), the epilogue warps scale
// it by 1 / (denominator + eps) and rewrite it IN PLACE as bf16 hi | lo planes, and the merge projection reads it from
// tensor memory as the A operand of its K half (same technique as mlp_fused_kernel); D_merge = columns 256..511.
// Per tile: 2 x (4 + 4) k-blocks of 32 through the 48 KB stages of the MLP kernel; the message planes (0.4 GB written and
// read back per layer step at 32 frames) and one launch go away.
struct AttnMergeParams {
  CUtensorMap q_hi, q_lo;      // [R, 256] planes, box {32, 128}, 64-byte swizzle
  CUtensorMap kv_hi, kv_lo;    // [G * 2 * 144, 128] block-diagonal KV records, box {32, 144}
  CUtensorMap wm_hi, wm_lo;    // [256, 256], box {32, 256}
  const int* b_row_base;       // per m tile: first row of its group's KV records
  EpiLN<256>::Params ln;
  float eps;
  int m_tile_begin, num_tiles;
};
constexpr int AM_KVPL = 144 * 32 * 2;  // one KV plane of a k-block: 9216 B

__global__ void __launch_bounds__(64 + 128 * EPI_SPLIT, 1) attn_merge_kernel(const __grid_constant__ AttnMergeParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 511) & ~uintptr_t(511));
  uint8_t* stages = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + MF_STAGES * MF_STAGE_BYTES);
  uint64_t* empty = full + MF_STAGES;
  uint64_t* da_full = empty + MF_STAGES;
  uint64_t* a2_full = da_full + 1;
  uint64_t* dm_full = a2_full + 1;
  uint64_t* dm_empty = dm_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dm_empty + 1);
  uint8_t* ln_scratch = smem + MF_STAGES * MF_STAGE_BYTES + 256;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kEpiWarps = 4 * EPI_SPLIT;
  constexpr uint32_t kDA = 0, kDM = 256;  // tensor-memory columns
  static_assert(EPI_SPLIT == 4, "one epilogue warp per head of a half");

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.q_hi);
    ptx::prefetch_tmap(&p.kv_hi);
    ptx::prefetch_tmap(&p.wm_hi);
    for (int s = 0; s < MF_STAGES; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    ptx::mbar_init(da_full, 1);
    ptx::mbar_init(a2_full, kEpiWarps);
    ptx::mbar_init(dm_full, 1);
    ptx::mbar_init(dm_empty, kEpiWarps);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_it = (p.num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (same order as the MMA thread)
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      auto next = [&]() {
        if (++stage == MF_STAGES) stage = 0, phase ^= 1;
      };
      for (int it = 0; it < n_it; ++it) {
        const int m_tile = p.m_tile_begin + static_cast<int>(blockIdx.x) + it * static_cast<int>(gridDim.x);
        const int row = m_tile * BLOCK_M;
        const int kv_row = p.b_row_base[m_tile];
        for (int h = 0; h < 2; ++h) {
          for (int kb = 0; kb < 4; ++kb) {  // phi(Q)[:, 128 h ..] and the KV records of these four heads
            ptx::mbar_wait(&empty[stage], phase ^ 1);
            ptx::mbar_arrive_expect_tx(&full[stage], 2 * MF_APL + 2 * AM_KVPL);
            uint8_t* st = stages + stage * MF_STAGE_BYTES;
            ptx::tma_load_2d(st, &p.q_hi, &full[stage], h * 128 + kb * 32, row);
            ptx::tma_load_2d(st + MF_APL, &p.q_lo, &full[stage], h * 128 + kb * 32, row);
            ptx::tma_load_2d(st + 2 * MF_APL, &p.kv_hi, &full[stage], kb * 32, kv_row + h * 144);
            ptx::tma_load_2d(st + 2 * MF_APL + AM_KVPL, &p.kv_lo, &full[stage], kb * 32, kv_row + h * 144);
            next();
          }
          for (int kb = 0; kb < 4; ++kb) {  // the Wm columns that meet this half of the message
            ptx::mbar_wait(&empty[stage], phase ^ 1);
            ptx::mbar_arrive_expect_tx(&full[stage], 2 * MF_BPL);
            uint8_t* st = stages + stage * MF_STAGE_BYTES;
            ptx::tma_load_2d(st + 2 * MF_APL, &p.wm_hi, &full[stage], h * 128 + kb * 32, 0);
            ptx::tma_load_2d(st + 2 * MF_APL + MF_BPL, &p.wm_lo, &full[stage], h * 128 + kb * 32, 0);
            next();
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (ptx::elect_one()) {
      constexpr uint32_t idesc_a = ptx::make_idesc_bf16(BLOCK_M, 144), idesc_m = ptx::make_idesc_bf16(BLOCK_M, 256);
      const uint32_t da = tmem_base + kDA, dm = tmem_base + kDM;
      int stage = 0;
      uint32_t phase = 0;
      auto next = [&]() {
        if (++stage == MF_STAGES) stage = 0, phase ^= 1;
      };
      for (int it = 0; it < n_it; ++it) {
        for (int h = 0; h < 2; ++h) {
          const uint32_t u1 = static_cast<uint32_t>(2 * it + h);
          // (D_attn is free: the merge MMAs that read it as their A operand were issued before these and run before them)
          for (int kb = 0; kb < 4; ++kb) {
            ptx::mbar_wait(&full[stage], phase);
            ptx::tc_fence_after();
            const uint32_t sa = ptx::smem_u32(stages + stage * MF_STAGE_BYTES);
            const uint64_t a_hi = ptx::make_desc_sw64(sa), a_lo = ptx::make_desc_sw64(sa + MF_APL);
            const uint64_t b_hi = ptx::make_desc_sw64(sa + 2 * MF_APL), b_lo = ptx::make_desc_sw64(sa + 2 * MF_APL + AM_KVPL);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(da, a_hi + 2 * k, b_hi + 2 * k, idesc_a, !(kb == 0 && k == 0));
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(da, a_lo + 2 * k, b_hi + 2 * k, idesc_a, 1);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(da, a_hi + 2 * k, b_lo + 2 * k, idesc_a, 1);
            ptx::mma_commit(&empty[stage]);
            next();
          }
          ptx::mma_commit(da_full);
          if (h == 0) ptx::mbar_wait(dm_empty, (static_cast<uint32_t>(it) & 1) ^ 1);  // the previous tile's output has been read
          ptx::mbar_wait(a2_full, u1 & 1);  // this half of the message is in tensor memory as bf16 planes
          ptx::tc_fence_after();
          for (int kb = 0; kb < 4; ++kb) {
            ptx::mbar_wait(&full[stage], phase);
            ptx::tc_fence_after();
            const uint32_t sa = ptx::smem_u32(stages + stage * MF_STAGE_BYTES);
            const uint64_t b_hi = ptx::make_desc_sw64(sa + 2 * MF_APL), b_lo = ptx::make_desc_sw64(sa + 2 * MF_APL + MF_BPL);
            const uint32_t a_hi = tmem_base + kDA + 16 * kb, a_lo = a_hi + 64;
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(dm, a_hi + 8 * k, b_hi + 2 * k, idesc_m, !(h == 0 && kb == 0 && k == 0));
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(dm, a_lo + 8 * k, b_hi + 2 * k, idesc_m, 1);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16_ta(dm, a_hi + 8 * k, b_lo + 2 * k, idesc_m, 1);
            ptx::mma_commit(&empty[stage]);
            next();
          }
        }
        ptx::mma_commit(dm_full);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps: (head of the half, TMEM lane quarter)
    EpiCtx ctx;
    ctx.quarter = warp & 3;
    ctx.lane = lane;
    ctx.epi_warp = (warp - 2) & 3;
    ctx.group = 0;
    ctx.part = (warp - 2) >> 2;
    ctx.n_tile = 0;
    ctx.scratch = ln_scratch;
    ctx.scratch_small = 1;
    const uint32_t lane_field = static_cast<uint32_t>(ctx.quarter * 32) << 16;
    const int bar_id = 1 + ctx.quarter;  // the four warps (heads) of a TMEM lane quarter
    auto convert = [&](int it, int h) {  // e(it, h): D_attn -> scaled bf16 planes in place
      const uint32_t u1 = static_cast<uint32_t>(2 * it + h);
      ptx::mbar_wait(da_full, u1 & 1);
      ptx::tc_fence_after();
      const uint32_t base = tmem_base + kDA + lane_field;
      uint32_t hh[16], ll[16];
      {
        uint32_t dr[16];
        ptx::tmem_ld16(base + 128, dr);
        ptx::tmem_ld_wait();
        float den = __uint_as_float(dr[0]);
        den = ctx.part == 1 ? __uint_as_float(dr[1]) : den;
        den = ctx.part == 2 ? __uint_as_float(dr[2]) : den;
        den = ctx.part == 3 ? __uint_as_float(dr[3]) : den;
        const float z = 1.0f / (den + p.eps);
        float v[32];
        load_acc32(base + ctx.part * 32, v);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          __nv_bfloat16 h0, l0, h1, l1;
          split_bf16(v[2 * i] * z, h0, l0);
          split_bf16(v[2 * i + 1] * z, h1, l1);
          hh[i] = pack_bf16(h0, h1);
          ll[i] = pack_bf16(l0, l1);
        }
      }
      ptx::named_bar_sync(bar_id, 32 * EPI_SPLIT);  // the four heads' columns have all been read: the planes may land
      ptx::tmem_st16(base + ctx.part * 16, hh);
      ptx::tmem_st16(base + 64 + ctx.part * 16, ll);
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(a2_full);
    };
    // LayerNorm1 of a tile (EpiLN<256>::run with D_merge handed back as soon as it is in registers)
    auto layer_norm = [&](int it) {
      constexpr int CPT = 64;
      ctx.m_tile = p.m_tile_begin + static_cast<int>(blockIdx.x) + it * static_cast<int>(gridDim.x);
      ptx::mbar_wait(dm_full, static_cast<uint32_t>(it) & 1);
      ptx::tc_fence_after();
      ctx.tmem_acc = tmem_base + kDM + lane_field;
      float x[CPT];
      load_row_slice<CPT>(ctx, x);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(dm_empty);
      float mean, rstd;
      row_mean_rstd_split<CPT>(ctx, x, 256, mean, rstd);
      const int cb = ctx.part * CPT;
#pragma unroll
      for (int i = 0; i < CPT; i += 4) {
        const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.ln.gamma + cb + i));
        const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.ln.beta + cb + i));
        x[i + 0] = (x[i + 0] - mean) * rstd * g4.x + b4.x;
        x[i + 1] = (x[i + 1] - mean) * rstd * g4.y + b4.y;
        x[i + 2] = (x[i + 2] - mean) * rstd * g4.z + b4.z;
        x[i + 3] = (x[i + 3] - mean) * rstd * g4.w + b4.w;
      }
      const long long wrow0 = static_cast<long long>(ctx.m_tile) * BLOCK_M + ctx.quarter * 32;
      uint32_t h[4][8], l[4][8];
      pack_planes64(x, h, l);
      store_row64_bf16_quad(h, p.ln.o_hi + wrow0 * 256 + cb, 256, ctx.lane);
      store_row64_bf16_quad(l, p.ln.o_lo + wrow0 * 256 + cb, 256, ctx.lane);
    };
    // order: e(t, 0) | LN1(t - 1) | e(t, 1): the first conversion of a tile goes before the previous tile's LayerNorm, so
    // the tensor pipe gets its operand (and, right after LN1 has loaded D_merge, its accumulator) back without waiting
    // for the LayerNorm arithmetic and stores
    for (int it = 0; it < n_it; ++it) {
      convert(it, 0);
      if (it > 0) layer_norm(it - 1);
      convert(it, 1);
    }
    if (n_it > 0) layer_norm(n_it - 1);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

constexpr int KV_ROWS_PER_BLOCK = 128;
constexpr int KV_REC = 33;  // per (head, d): 32 KV values + Ksum

struct GroupTable {  // passed by value to small kernels
  int num_groups;
  int row_start[64];
  int len[64];
};

// grid = (max_chunks, G), block = 128: thread t -> head t/16, d pair (2*(t%16), +1); accumulates KV[h][d][0..31] and
// Ksum[h][d] of its two d over the chunk's rows: 64 FMAs per 8 broadcast 16-byte V reads + one 8-byte K read, so the
// FMA pipe is the bound.  Rows are staged 8 at a time through a cp.async double buffer (coalesced 16-byte copies; a third slab
// measured slower: occupancy);
// partials are written j-major ([33][256]) so that both this store and the reduce are coalesced.
__global__ void __launch_bounds__(128) kv_partial_kernel(const float* __restrict__ kphi, const float* __restrict__ v,
                                                         GroupTable gt, unsigned long long group_sel,
                                                         float* __restrict__ partial, int max_chunks) {
  const int g = blockIdx.y;
  if (!((group_sel >> g) & 1ull)) return;
  const int chunk = blockIdx.x;
  const int r0 = chunk * KV_ROWS_PER_BLOCK;
  const int len = gt.len[g];
  if (r0 >= len) return;
  const int nrows = min(KV_ROWS_PER_BLOCK, len - r0);
  __shared__ __align__(16) float vs[2][8][256];
  __shared__ __align__(16) float ks[2][8][256];
  const int t = threadIdx.x;
  const int h = t >> 4, d0 = (t & 15) * 2;
  // accumulators as packed fp32 pairs (v, v + 1): one FFMA2 updates two of them, halving the FMA issue slots this
  // kernel is bound by (ncu: issue slots 74 % busy with scalar FFMA); results are bit-identical to scalar fmaf
  uint64_t acc0[16], acc1[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc0[i] = acc1[i] = 0ull;
  float ksum0 = 0.f, ksum1 = 0.f;
  const float* kbase = kphi + static_cast<long long>(gt.row_start[g] + r0) * 256;
  const float* vbase = v + static_cast<long long>(gt.row_start[g] + r0) * 256;
  const int nb = (nrows + 7) / 8;
  auto issue = [&](int b) {  // 8 rows x 256 floats = 512 float4 per operand, 4 per thread
    const int buf = b & 1;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int f4 = t + 128 * q;           // float4 index within the 8-row batch
      const int row = b * 8 + (f4 >> 6);
      if (row < nrows) {
        const long long off = static_cast<long long>(row) * 256 + (f4 & 63) * 4;
        __pipeline_memcpy_async(&ks[buf][f4 >> 6][(f4 & 63) * 4], kbase + off, 16);
        __pipeline_memcpy_async(&vs[buf][f4 >> 6][(f4 & 63) * 4], vbase + off, 16);
      }
    }
    __pipeline_commit();
  };
  issue(0);
  for (int b = 0; b < nb; ++b) {
    if (b + 1 < nb) issue(b + 1);
    else __pipeline_commit();
    __pipeline_wait_prior(1);
    __syncthreads();
    const int buf = b & 1;
    const int nr = min(8, nrows - b * 8);
    for (int i = 0; i < nr; ++i) {
      const float2 kk = *reinterpret_cast<const float2*>(&ks[buf][i][h * 32 + d0]);
      ksum0 += kk.x;
      ksum1 += kk.y;
      const uint64_t k0 = ptx::pack_f32x2(kk.x, kk.x), k1 = ptx::pack_f32x2(kk.y, kk.y);
      const ulonglong2* v2 = reinterpret_cast<const ulonglong2*>(&vs[buf][i][h * 32]);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const ulonglong2 vv = v2[j];  // (v[4j], v[4j+1]), (v[4j+2], v[4j+3])
        acc0[2 * j + 0] = ptx::fma_f32x2(k0, vv.x, acc0[2 * j + 0]);
        acc0[2 * j + 1] = ptx::fma_f32x2(k0, vv.y, acc0[2 * j + 1]);
        acc1[2 * j + 0] = ptx::fma_f32x2(k1, vv.x, acc1[2 * j + 0]);
        acc1[2 * j + 1] = ptx::fma_f32x2(k1, vv.y, acc1[2 * j + 1]);
      }
    }
    __syncthreads();
  }
  float2* out = reinterpret_cast<float2*>(partial + (static_cast<long long>(g) * max_chunks + chunk) * (KV_REC * 256) + h * 32 + d0);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float a0, a1, b0, b1;
    ptx::unpack_f32x2(acc0[i], a0, a1);
    ptx::unpack_f32x2(acc1[i], b0, b1);
    out[(2 * i) * 128] = make_float2(a0, b0);
    out[(2 * i + 1) * 128] = make_float2(a1, b1);
  }
  out[32 * 128] = make_float2(ksum0, ksum1);
}

// ------------------------------------------------------------------------------------------------
// KV = phi(K)^T V and Ksum on the tensor cores (north_star (a); replaces kv_partial_kernel for the coarse level).
//
// The contraction runs over the TOKENS, i.e. over the rows of the row-major [R, 256] planes, so both operands are
// MN-major: a TMA box {64 channels, 64 tokens} with the 128-byte swizzle is exactly the canonical MN-major tile
// (row = token = 128 bytes, eight tokens per swizzle atom; LBO = 8192 = distance to the next 64 channels, SBO = 1024,
// 16 tokens = +2048 bytes per MMA; tools/probes/mn_major_probe.cu).  One CTA = (split of the token range, group,
// half of the heads): D[128 d-channels x 128 v-channels] += phi(K)^T V over its tokens with the three split products
// (hi*hi + lo*hi + hi*lo), of which the four diagonal 32 x 32 blocks are the KV of heads 4*half .. 4*half+3 (the
// off-diagonal blocks are the price of a 128-row MMA: 4x the flops of an op that is HBM-bound at 8 flop/B anyway), and
// D[:, 128..143] += phi(K)^T 1 against a constant K-major tile of ones = Ksum.  Partials go to the same
// [g][split][33][256] records kv_pack_kernel folds (deterministic: fixed split ranges, fixed order).
//   warp 0: TMA thread (8 boxes = 64 KB per 64-token stage, 3 stages), warp 1: MMA thread, warps 2-5: epilogue.
constexpr int KVT_STAGES = 3;
constexpr int KVT_BOX = 64 * 128;                 // one box: 64 tokens x 64 channels x 2 B
constexpr int KVT_STAGE_BYTES = 8 * KVT_BOX;      // {K hi, K lo, V hi, V lo} x two boxes of 64 channels
constexpr int KVT_ONES_BYTES = 16 * 128;          // K-major [16 x 64] tile of ones
constexpr int KVT_SMEM = 1024 + KVT_STAGES * KVT_STAGE_BYTES + KVT_ONES_BYTES + 256;

struct KvTcParams {
  CUtensorMap k_hi, k_lo, v_hi, v_lo;  // [R, 256] bf16, box {64 channels, 64 tokens}, 128B swizzle
  GroupTable gt;
  unsigned long long group_sel;
  float* partial;
  int max_chunks, n_splits, x3;
};

__device__ __forceinline__ uint64_t make_desc_mn_sw128(uint32_t smem_addr) {  // MN-major, two 64-wide boxes 8192 B apart
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(8192 >> 4) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

__global__ void __launch_bounds__(192, 1) kv_tc_kernel(const __grid_constant__ KvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* ones = smem + KVT_STAGES * KVT_STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(ones + KVT_ONES_BYTES);
  uint64_t* empty = full + KVT_STAGES;
  uint64_t* done = empty + KVT_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int split = blockIdx.x, g = blockIdx.y, half = blockIdx.z;
  if (!((p.group_sel >> g) & 1ull)) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = (p.gt.len[g] + 63) / 64;  // 64-token blocks of the group (its rows are padded to 128 with zeros)
  const int kb0 = static_cast<int>(static_cast<long long>(nkb) * split / p.n_splits);
  const int kb1 = static_cast<int>(static_cast<long long>(nkb) * (split + 1) / p.n_splits);
  float* out = p.partial + (static_cast<long long>(g) * p.max_chunks + split) * (KV_REC * 256) + half * 128;
  if (kb1 <= kb0) {  // nothing to add: the record still has to exist
    for (int i = threadIdx.x; i < KV_REC * 128; i += blockDim.x) out[(i >> 7) * 256 + (i & 127)] = 0.f;
    return;
  }
  for (int i = threadIdx.x; i < KVT_ONES_BYTES / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(ones)[i] = 0x3F803F80u;  // bf16 1.0
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.k_hi);
    ptx::prefetch_tmap(&p.v_hi);
    for (int s = 0; s < KVT_STAGES; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    ptx::mbar_init(done, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 256);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async();  // the ones tile is read by the tensor core (async proxy)
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int row0 = p.gt.row_start[g];
  const int planes = p.x3 ? 4 : 2;

  if (warp == 0) {
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&empty[stage], phase ^ 1);
        ptx::mbar_arrive_expect_tx(&full[stage], planes * 2 * KVT_BOX);
        uint8_t* st = smem + stage * KVT_STAGE_BYTES;
        const int tok = row0 + kb * 64;
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const int ch = half * 128 + b * 64;
          ptx::tma_load_2d(st + (0 + b) * KVT_BOX, &p.k_hi, &full[stage], ch, tok);
          ptx::tma_load_2d(st + (4 + b) * KVT_BOX, &p.v_hi, &full[stage], ch, tok);
          if (p.x3) {
            ptx::tma_load_2d(st + (2 + b) * KVT_BOX, &p.k_lo, &full[stage], ch, tok);
            ptx::tma_load_2d(st + (6 + b) * KVT_BOX, &p.v_lo, &full[stage], ch, tok);
          }
        }
        if (++stage == KVT_STAGES) stage = 0, phase ^= 1;
      }
    }
  } else if (warp == 1) {
    if (ptx::elect_one()) {
      constexpr uint32_t kMN = (1u << 15) | (1u << 16);  // a_major, b_major = MN
      constexpr uint32_t idesc_kv = ptx::make_idesc_bf16(BLOCK_M, 128) | kMN;
      constexpr uint32_t idesc_ks = ptx::make_idesc_bf16(BLOCK_M, 16) | (1u << 15);  // A MN-major, B (ones) K-major
      const uint64_t ones_desc = ptx::make_desc_sw128(ptx::smem_u32(ones));
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&full[stage], phase);
        ptx::tc_fence_after();
        const uint32_t sa = ptx::smem_u32(smem + stage * KVT_STAGE_BYTES);
        const uint64_t kh = make_desc_mn_sw128(sa), kl = make_desc_mn_sw128(sa + 2 * KVT_BOX);
        const uint64_t vh = make_desc_mn_sw128(sa + 4 * KVT_BOX), vl = make_desc_mn_sw128(sa + 6 * KVT_BOX);
#pragma unroll
        for (int k = 0; k < 4; ++k) {  // 16 tokens per MMA: +2048 bytes in the MN-major tiles, +32 bytes in the ones tile
          const bool first = kb == kb0 && k == 0;
          if (p.x3) {
            ptx::mma_bf16(tmem_base, kh + 128 * k, vl + 128 * k, idesc_kv, !first);
            ptx::mma_bf16(tmem_base, kl + 128 * k, vh + 128 * k, idesc_kv, 1);
            ptx::mma_bf16(tmem_base, kh + 128 * k, vh + 128 * k, idesc_kv, 1);
            ptx::mma_bf16(tmem_base + 128, kl + 128 * k, ones_desc + 2 * k, idesc_ks, !first);
            ptx::mma_bf16(tmem_base + 128, kh + 128 * k, ones_desc + 2 * k, idesc_ks, 1);
          } else {
            ptx::mma_bf16(tmem_base, kh + 128 * k, vh + 128 * k, idesc_kv, !first);
            ptx::mma_bf16(tmem_base + 128, kh + 128 * k, ones_desc + 2 * k, idesc_ks, !first);
          }
        }
        ptx::mma_commit(&empty[stage]);
        if (++stage == KVT_STAGES) stage = 0, phase ^= 1;
      }
      ptx::mma_commit(done);
    }
  } else {
    // epilogue: TMEM lanes 32q .. 32q+31 = the d channels of head 4*half + q; its KV block are the columns 32q .. 32q+31
    const int q = warp & 3;
    ptx::mbar_wait(done, 0);
    ptx::tc_fence_after();
    uint32_t kv[32], ks[16];
    ptx::tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + q * 32, kv);
    ptx::tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + 128, ks);
    ptx::tmem_ld_wait();
    float* o = out + q * 32 + lane;  // record j, thread index t = head * 32 + d
#pragma unroll
    for (int j = 0; j < 32; ++j) o[j * 256] = __uint_as_float(kv[j]);
    o[32 * 256] = __uint_as_float(ks[0]);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 256);
  }
}

// grid = (33, G), block = 256 (thread t = head*32 + d).  Sums the chunk partials of record j = blockIdx.x in a
// fixed order (deterministic) and writes the bf16 planes of the block-diagonal B operand of the ATTN GEMM.
// Per (group, half) the operand is 144 rows (n) x 128 (k):
//   n <  128 : n = hh*32 + v, k = hh*32 + d  ->  KV[half*4+hh][d][v]      (zero off the diagonal blocks)
//   n = 128+hh : k = hh*32 + d               ->  Ksum[half*4+hh][d]
// The zero entries are written once per transformer call (memset), never here.
__global__ void __launch_bounds__(256) kv_pack_kernel(const float* __restrict__ partial, GroupTable gt,
                                                      unsigned long long group_sel, int max_chunks,
                                                      __nv_bfloat16* __restrict__ b_hi, __nv_bfloat16* __restrict__ b_lo,
                                                      int n_splits /* > 0: that many partials per group (kv_tc_kernel) */) {
  const int g = blockIdx.y, j = blockIdx.x;
  if (!((group_sel >> g) & 1ull)) return;
  const int nchunks = n_splits > 0 ? n_splits : (gt.len[g] + KV_ROWS_PER_BLOCK - 1) / KV_ROWS_PER_BLOCK;
  const int t = threadIdx.x;
  const float* src = partial + static_cast<long long>(g) * max_chunks * (KV_REC * 256) + j * 256 + t;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  int c = 0;
  for (; c + 4 <= nchunks; c += 4) {
    a0 += src[static_cast<long long>(c + 0) * (KV_REC * 256)];
    a1 += src[static_cast<long long>(c + 1) * (KV_REC * 256)];
    a2 += src[static_cast<long long>(c + 2) * (KV_REC * 256)];
    a3 += src[static_cast<long long>(c + 3) * (KV_REC * 256)];
  }
  for (; c < nchunks; ++c) a0 += src[static_cast<long long>(c) * (KV_REC * 256)];
  const float sum = (a0 + a1) + (a2 + a3);
  const int h = t >> 5, d = t & 31, half = h >> 2, hh = h & 3;
  const int n = (j < 32) ? hh * 32 + j : 128 + hh;
  const long long o = (static_cast<long long>(g) * 2 + half) * 144 * 128 + n * 128 + hh * 32 + d;
  __nv_bfloat16 h_, l_;
  split_bf16(sum, h_, l_);
  b_hi[o] = h_;
  if (b_lo) b_lo[o] = l_;
}

// ------------------------------------------------------------------------------------------------
// Small-group attention (fine level: C = 128, H = 8, D = 16; groups are single matches of <= 128 rows).
// One block per query group: build KV / Ksum of the source group in shared memory, then apply.
// ------------------------------------------------------------------------------------------------
struct SmallAttnJob {  // one (query stream, source stream) pairing of a layer step
  int q_row0, q_len, s_row0, s_len;
};
struct SmallAttnJobs {
  SmallAttnJob job[2];
};

// grid = (groups, jobs), block = 128 threads = channels (head h = t / 16, d or v index = t % 16).  Source / query rows are
// staged 32 at a time through shared memory with coalesced 16-byte loads (three barriers per chunk instead of two
// per row); the 16 x 16 KV block of a head and its Ksum stay in registers for the apply phase.
__global__ void __launch_bounds__(128) attn_small_kernel(const float* __restrict__ qphi, const float* __restrict__ kphi,
                                                         const float* __restrict__ v, SmallAttnJobs jobs, int num_groups,
                                                         float eps, __nv_bfloat16* __restrict__ o_hi,
                                                         __nv_bfloat16* __restrict__ o_lo) {
  constexpr int C = 128, D = 16, CH = 32;
  const int g = blockIdx.x;
  if (g >= num_groups) return;
  const SmallAttnJob jb = jobs.job[blockIdx.y];
  const int t = threadIdx.x;
  const int h = t >> 4, dv = t & 15;
  __shared__ __align__(16) float abuf[CH][C];  // K rows, later Q rows
  __shared__ __align__(16) float bbuf[CH][C];  // V rows
  __shared__ float kv_s[8][D][D + 1];
  __shared__ float ksum_s[8][D];
  const float4* k4 = reinterpret_cast<const float4*>(kphi + static_cast<long long>(jb.s_row0 + g * jb.s_len) * C);
  const float4* v4 = reinterpret_cast<const float4*>(v + static_cast<long long>(jb.s_row0 + g * jb.s_len) * C);
  float acc[D];
#pragma unroll
  for (int i = 0; i < D; ++i) acc[i] = 0.f;
  float ksum = 0.f;
  for (int r0 = 0; r0 < jb.s_len; r0 += CH) {
    const int nr = min(CH, jb.s_len - r0);
    __syncthreads();
    // cp.async: every 16-byte copy of the chunk in flight at once, no staging registers (occupancy)
    for (int i = t; i < nr * (C / 4); i += 128) {
      __pipeline_memcpy_async(reinterpret_cast<float4*>(&abuf[0][0]) + i, k4 + r0 * (C / 4) + i, 16);
      __pipeline_memcpy_async(reinterpret_cast<float4*>(&bbuf[0][0]) + i, v4 + r0 * (C / 4) + i, 16);
    }
    __pipeline_commit();
    __pipeline_wait_prior(0);
    __syncthreads();
    for (int r = 0; r < nr; ++r) {
      const float kk = abuf[r][t];
      ksum += kk;
      const float4* vv = reinterpret_cast<const float4*>(&bbuf[r][h * D]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 f = vv[q];
        acc[4 * q + 0] = fmaf(kk, f.x, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(kk, f.y, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(kk, f.z, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(kk, f.w, acc[4 * q + 3]);
      }
    }
  }
  // thread (h, d) holds KV[h][d][0..15]; the apply phase wants thread (h, i) to hold KV[h][0..15][i]: transpose via smem
#pragma unroll
  for (int i = 0; i < D; ++i) kv_s[h][dv][i] = acc[i];
  ksum_s[h][dv] = ksum;
  __syncthreads();
  float kvr[D], ksr[D];
#pragma unroll
  for (int d = 0; d < D; ++d) {
    kvr[d] = kv_s[h][d][dv];
    ksr[d] = ksum_s[h][d];
  }
  const long long qbase = static_cast<long long>(jb.q_row0 + g * jb.q_len) * C;
  const float4* q4 = reinterpret_cast<const float4*>(qphi + qbase);
  for (int r0 = 0; r0 < jb.q_len; r0 += CH) {
    const int nr = min(CH, jb.q_len - r0);
    __syncthreads();
    for (int i = t; i < nr * (C / 4); i += 128)
      __pipeline_memcpy_async(reinterpret_cast<float4*>(&abuf[0][0]) + i, q4 + r0 * (C / 4) + i, 16);
    __pipeline_commit();
    __pipeline_wait_prior(0);
    __syncthreads();
    for (int r = 0; r < nr; ++r) {
      const float4* qq = reinterpret_cast<const float4*>(&abuf[r][h * D]);
      float num = 0.f, den = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 f = qq[q];
        num = fmaf(f.x, kvr[4 * q + 0], num), den = fmaf(f.x, ksr[4 * q + 0], den);
        num = fmaf(f.y, kvr[4 * q + 1], num), den = fmaf(f.y, ksr[4 * q + 1], den);
        num = fmaf(f.z, kvr[4 * q + 2], num), den = fmaf(f.z, ksr[4 * q + 2], den);
        num = fmaf(f.w, kvr[4 * q + 3], num), den = fmaf(f.w, ksr[4 * q + 3], den);
      }
      const float o = num / (den + eps);
      __nv_bfloat16 hi_, lo_;
      split_bf16(o, hi_, lo_);
      const long long idx = qbase + static_cast<long long>(r0 + r) * C + t;
      o_hi[idx] = hi_;
      if (o_lo) o_lo[idx] = lo_;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Row-space staging
// ------------------------------------------------------------------------------------------------
// x element (n, l, c) at x[n*sn + l*sl + c*sc]; optional additive table add[l*C + c] (position encoding).
// Writes rows [row0 + n*P, +P): valid rows get the data, pad rows zeros; also the row mask.
__global__ void copy_in_kernel(const float* __restrict__ x, long long sn, long long sl, long long sc,
                               const float* __restrict__ add, const uint8_t* __restrict__ mask, int len, int P, int C,
                               int row0, __nv_bfloat16* __restrict__ xh,
                               __nv_bfloat16* __restrict__ xl, uint8_t* __restrict__ row_mask) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int l0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const float* xn = x + n * sn;
  if (sc == 1) {
    for (int i = ty; i < 32; i += 8) {
      const int l = l0 + i, c = c0 + tx;
      tile[i][tx] = (l < len && c < C) ? xn[l * sl + c] : 0.f;
    }
  } else {
    for (int i = ty; i < 32; i += 8) {
      const int c = c0 + i, l = l0 + tx;
      tile[tx][i] = (l < len && c < C) ? xn[l * sl + c * sc] : 0.f;
    }
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int l = l0 + i, c = c0 + tx;
    if (l < P && c < C) {
      float val = tile[i][tx];
      if (add != nullptr && l < len) val += add[static_cast<long long>(l) * C + c];
      const long long o = (static_cast<long long>(row0) + static_cast<long long>(n) * P + l) * C + c;
      __nv_bfloat16 h_, l_;
      split_bf16(val, h_, l_);
      xh[o] = h_;
      xl[o] = l_;
      if (c == 0)
        row_mask[row0 + n * P + l] = (l < len) ? (mask ? (mask[static_cast<long long>(n) * len + l] != 0) : 1) : 0;
    }
  }
}

// Row-major fast path of copy_in_kernel (sc == 1, sl % 4 == 0, 16-byte aligned, C % 128 == 0): one warp per row,
// 16-byte loads, 8-byte plane stores; pad rows (l >= len) are zero-filled.  grid = (ceil(P / 8), 1, n_batch), block 256.
__global__ void __launch_bounds__(256) copy_in_rows_kernel(const float* __restrict__ x, long long sn, long long sl,
                                                           const float* __restrict__ add, const uint8_t* __restrict__ mask,
                                                           int len, int P, int C, int row0, __nv_bfloat16* __restrict__ xh,
                                                           __nv_bfloat16* __restrict__ xl, uint8_t* __restrict__ row_mask) {
  const int n = blockIdx.z;
  const int l = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (l >= P) return;
  const long long orow = static_cast<long long>(row0) + static_cast<long long>(n) * P + l;
  const bool live = l < len;
  const float4* src = reinterpret_cast<const float4*>(x + n * sn + static_cast<long long>(l) * sl);
  const float4* ad = add != nullptr ? reinterpret_cast<const float4*>(add + static_cast<long long>(l) * C) : nullptr;
  for (int c4 = lane; c4 < C / 4; c4 += 32) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) {
      v = __ldg(src + c4);
      if (ad != nullptr) {
        const float4 a = __ldg(ad + c4);
        v.x += a.x, v.y += a.y, v.z += a.z, v.w += a.w;
      }
    }
    const __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
    const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
    const __nv_bfloat162 l0 = __floats2bfloat162_rn(v.x - f0.x, v.y - f0.y), l1 = __floats2bfloat162_rn(v.z - f1.x, v.w - f1.y);
    uint2 ph, pl;
    ph.x = *reinterpret_cast<const uint32_t*>(&h0), ph.y = *reinterpret_cast<const uint32_t*>(&h1);
    pl.x = *reinterpret_cast<const uint32_t*>(&l0), pl.y = *reinterpret_cast<const uint32_t*>(&l1);
    *reinterpret_cast<uint2*>(xh + orow * C + c4 * 4) = ph;
    *reinterpret_cast<uint2*>(xl + orow * C + c4 * 4) = pl;
  }
  if (lane == 0) row_mask[orow] = live ? (mask ? (mask[static_cast<long long>(n) * len + l] != 0) : 1) : 0;
}

__global__ void copy_out_kernel(const __nv_bfloat16* __restrict__ xh, const __nv_bfloat16* __restrict__ xl, int len, int P,
                                int C, int row0, float* __restrict__ out) {
  const int n = blockIdx.z;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;  // group of 8 elements within [len, C]
  const long long total8 = static_cast<long long>(len) * C / 8;
  if (i >= total8) return;
  const long long src = (static_cast<long long>(row0) + static_cast<long long>(n) * P) * C + i * 8;
  const uint4 h = *reinterpret_cast<const uint4*>(xh + src);
  const uint4 l = *reinterpret_cast<const uint4*>(xl + src);
  const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
  float v[8];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 hf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hw[k]));
    const float2 lf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&lw[k]));
    v[2 * k] = hf.x + lf.x;
    v[2 * k + 1] = hf.y + lf.y;
  }
  float4* o = reinterpret_cast<float4*>(out + static_cast<long long>(n) * len * C + i * 8);
  o[0] = make_float4(v[0], v[1], v[2], v[3]);
  o[1] = make_float4(v[4], v[5], v[6], v[7]);
}

__global__ void fill_row_base_kernel(int* base_self, int* base_cross, int tiles0, int tiles1, int N, int rows_per_group) {
  const int T0 = N * tiles0, T1 = N * tiles1;
  for (int t = threadIdx.x + blockIdx.x * blockDim.x; t < T0 + T1; t += blockDim.x * gridDim.x) {
    int g, partner;
    if (t < T0) {
      g = t / tiles0;
      partner = N + g;
    } else {
      g = N + (t - T0) / tiles1;
      partner = g - N;
    }
    base_self[t] = g * rows_per_group;
    base_cross[t] = partner * rows_per_group;
  }
}

// ------------------------------------------------------------------------------------------------
// Host driver
// ------------------------------------------------------------------------------------------------
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct LayerWeights {  // views into one packed buffer (see opb200_pack_layer_weights)
  const __nv_bfloat16 *qkv_hi, *qkv_lo, *m_hi, *m_lo, *w1_hi, *w1_lo, *w2_hi, *w2_lo;
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
};

static size_t packed_layer_bytes(int C) {
  return static_cast<size_t>(10) * C * C * 2 * 2 + static_cast<size_t>(4) * C * 4;
}

static LayerWeights view_layer(const void* buf, int C) {
  const __nv_bfloat16* p = static_cast<const __nv_bfloat16*>(buf);
  const size_t cc = static_cast<size_t>(C) * C;
  LayerWeights w;
  w.qkv_hi = p;
  w.m_hi = w.qkv_hi + 3 * cc;
  w.w1_hi = w.m_hi + cc;
  w.w2_hi = w.w1_hi + 4 * cc;
  const __nv_bfloat16* lo = w.w2_hi + 2 * cc;
  w.qkv_lo = lo;
  w.m_lo = w.qkv_lo + 3 * cc;
  w.w1_lo = w.m_lo + cc;
  w.w2_lo = w.w1_lo + 4 * cc;
  const float* f = reinterpret_cast<const float*>(w.w2_lo + 2 * cc);
  w.ln1_g = f;
  w.ln1_b = f + C;
  w.ln2_g = f + 2 * C;
  w.ln2_b = f + 3 * C;
  return w;
}

struct Workspace {
  __nv_bfloat16 *xh, *xl, *qh, *ql, *aoh, *aol, *mh, *ml, *hh, *hl, *kvb_h, *kvb_l;
  float *qf, *kf, *vf, *kv_partial;
  uint8_t* row_mask;
  int *base_self, *base_cross;
  size_t bytes;
};

struct Geometry {
  int C, N, len0, len1, P0, P1, row1, R, big, G, max_chunks, tiles0, tiles1;
};

static Geometry make_geometry(int C, int N, int len0, int len1, int big) {
  Geometry g;
  g.C = C;
  g.N = N;
  g.len0 = len0;
  g.len1 = len1;
  g.big = big;
  if (big) {  // every (stream, batch) group padded to whole 128-row tiles
    g.P0 = static_cast<int>(align_up(len0, 128));
    g.P1 = static_cast<int>(align_up(len1, 128));
  } else {    // groups packed back to back; only the two stream regions are tile aligned
    g.P0 = len0;
    g.P1 = len1;
  }
  g.row1 = static_cast<int>(align_up(static_cast<size_t>(N) * g.P0, 128));
  g.R = g.row1 + static_cast<int>(align_up(static_cast<size_t>(N) * g.P1, 128));
  g.G = 2 * N;
  const int maxlen = len0 > len1 ? len0 : len1;
  g.max_chunks = (maxlen + KV_ROWS_PER_BLOCK - 1) / KV_ROWS_PER_BLOCK;
  g.tiles0 = g.P0 / 128;
  g.tiles1 = g.P1 / 128;
  return g;
}

static Workspace carve(void* base, const Geometry& g) {
  Workspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? static_cast<uint8_t*>(base) + off : nullptr;
    off += align_up(bytes, 1024);
    return p;
  };
  const size_t RC = static_cast<size_t>(g.R) * g.C;
  w.xh = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.xl = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.qh = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.ql = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.qf = static_cast<float*>(take(g.big ? 0 : RC * 4));
  w.kf = static_cast<float*>(take(RC * 4));
  w.vf = static_cast<float*>(take(RC * 4));
  w.aoh = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.aol = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.mh = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.ml = static_cast<__nv_bfloat16*>(take(RC * 2));
  w.hh = static_cast<__nv_bfloat16*>(take(RC * 2 * 2));
  w.hl = static_cast<__nv_bfloat16*>(take(RC * 2 * 2));
  w.kv_partial = static_cast<float*>(take(g.big ? static_cast<size_t>(g.G) * g.max_chunks * 256 * KV_REC * 4 : 0));
  w.kvb_h = static_cast<__nv_bfloat16*>(take(g.big ? static_cast<size_t>(g.G) * 2 * 144 * 128 * 2 : 0));
  w.kvb_l = static_cast<__nv_bfloat16*>(take(g.big ? static_cast<size_t>(g.G) * 2 * 144 * 128 * 2 : 0));
  w.row_mask = static_cast<uint8_t*>(take(g.R));
  w.base_self = static_cast<int*>(take(static_cast<size_t>(g.R / 128 + 1) * 4));
  w.base_cross = static_cast<int*>(take(static_cast<size_t>(g.R / 128 + 1) * 4));
  w.bytes = off;
  return w;
}

// which streams are updated ("query side") in a launch: bit0 = stream 0, bit1 = stream 1
template <int C>
static int run_step(const Geometry& g, const Workspace& w, const LayerWeights& lw, bool cross, int q_streams,
                    int prec, cudaStream_t stream) {
  const bool x3 = prec == 3;
  const int kbC = C / BLOCK_K;
  const int R = g.R;
  const int T0 = g.row1 / 128;  // m tiles of stream 0 / stream 1
  const int T1 = (R - g.row1) / 128;
  const int Tall = R / 128;
  // m-tile range of the query streams / of the source streams
  auto tile_range = [&](int streams, int& begin, int& count) {
    if (streams == 3) {
      begin = 0;
      count = Tall;
    } else if (streams == 1) {
      begin = 0;
      count = T0;
    } else {
      begin = T0;
      count = T1;
    }
  };
  const int src_streams = cross ? (((q_streams & 1) ? 2 : 0) | ((q_streams & 2) ? 1 : 0)) : q_streams;
  int qb, qc, sb, sc;
  tile_range(q_streams, qb, qc);
  tile_range(src_streams, sb, sc);

  auto add_segs = [&](PlanBuilder& pb, int a_hi, int a_lo, int b_hi, int b_lo, int a_k0, int b_k0, int nkb) {
    if (x3) {
      pb.seg3(a_hi, a_lo, b_hi, b_lo, a_k0, b_k0, 2 * nkb);  // fused split-precision mainloop: 32-wide k-blocks
    } else {
      pb.seg(a_hi, b_hi, a_k0, b_k0, nkb);
    }
  };

  // ---- 1. QKV
  {
    // big groups (coarse level): phi(K), V leave as bf16 planes for the tensor-core KV kernel, in the space of the fp32 buffers
    static const bool kv_tc_env = getenv("OPB200_KV_TC") ? atoi(getenv("OPB200_KV_TC")) != 0 : true;
    const bool kv_tc = kv_tc_env && g.big && C == 256;
    __nv_bfloat16* kh = reinterpret_cast<__nv_bfloat16*>(w.kf);
    __nv_bfloat16* vh = reinterpret_cast<__nv_bfloat16*>(w.vf);
    const size_t plane = static_cast<size_t>(R) * C;
    typename EpiQKV<C>::Params ep{w.qh, x3 ? w.ql : nullptr, g.big ? nullptr : w.qf, w.kf, w.vf, w.row_mask,
                                  kv_tc ? kh : nullptr, kv_tc && x3 ? kh + plane : nullptr,
                                  kv_tc ? vh : nullptr, kv_tc && x3 ? vh + plane : nullptr};
    auto launch = [&](int mb, int mc, int nb, int nc) -> int {
      if (mc <= 0) return OPB_OK;
      PlanBuilder pb(x3);
      pb.add_a(w.xh, R, C, C, 0);
      pb.add_b(lw.qkv_hi, 3 * C, C, C, 0, C);
      if (x3) {
        pb.add_a(w.xl, R, C, C, 1);
        pb.add_b(lw.qkv_lo, 3 * C, C, C, 1, C);
      }
      if (pb.status) return pb.status;
      add_segs(pb, 0, 1, 0, 1, 0, 0, kbC);
      pb.finish();
      pb.plan.m_tile_begin = mb;
      pb.plan.n_tile_begin = nb;
      pb.plan.num_n_tiles = nc;
      pb.plan.num_tiles = mc * nc;
      pb.plan.b_rows_per_ntile = C;
      return x3 ? launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiQKV<C>, EPI_SPLIT, true>(pb.plan, ep, stream)
                : launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiQKV<C>, EPI_SPLIT>(pb.plan, ep, stream);
    };
    if (q_streams == src_streams) {
      if (int r = launch(qb, qc, 0, 3)) return r;
    } else {
      if (int r = launch(qb, qc, 0, 1)) return r;
      if (int r = launch(sb, sc, 1, 2)) return r;
    }
  }
  // ---- 2-4. attention
  bool merge_done = false;  // attn_merge_kernel has run MERGE + LN1 as well
  if (g.big) {
    if constexpr (C == 256) {
      GroupTable gt;
      gt.num_groups = g.G;
      unsigned long long sel = 0;
      for (int i = 0; i < g.G; ++i) {
        const int s = i / g.N, n = i % g.N;
        gt.row_start[i] = s == 0 ? n * g.P0 : g.row1 + n * g.P1;
        gt.len[i] = s == 0 ? g.len0 : g.len1;
        if (src_streams & (1 << s)) sel |= 1ull << i;
      }
      static const bool kv_tc = getenv("OPB200_KV_TC") ? atoi(getenv("OPB200_KV_TC")) != 0 : true;
      if (kv_tc) {
        KvTcParams kp;
        const __nv_bfloat16* kh = reinterpret_cast<const __nv_bfloat16*>(w.kf);
        const __nv_bfloat16* vh = reinterpret_cast<const __nv_bfloat16*>(w.vf);
        const size_t plane = static_cast<size_t>(R) * C;
        int r;
        if ((r = make_tmap_bf16(&kp.k_hi, kh, R, C, C, 64))) return r;
        if ((r = make_tmap_bf16(&kp.v_hi, vh, R, C, C, 64))) return r;
        if ((r = make_tmap_bf16(&kp.k_lo, x3 ? kh + plane : kh, R, C, C, 64))) return r;
        if ((r = make_tmap_bf16(&kp.v_lo, x3 ? vh + plane : vh, R, C, C, 64))) return r;
        kp.gt = gt, kp.group_sel = sel, kp.partial = w.kv_partial, kp.max_chunks = g.max_chunks, kp.x3 = x3 ? 1 : 0;
        // two waves of CTAs over the selected groups x 2 halves, at least 4 token blocks each, no more splits than records
        int n_sel = 0, min_len = 1 << 30;
        for (int i = 0; i < g.G; ++i)
          if ((sel >> i) & 1ull) ++n_sel, min_len = gt.len[i] < min_len ? gt.len[i] : min_len;
        int splits = (2 * num_sms() + 2 * n_sel - 1) / (2 * (n_sel > 0 ? n_sel : 1));
        const int most = (min_len + 255) / 256;
        splits = splits > most ? most : splits;
        splits = splits > g.max_chunks ? g.max_chunks : splits;
        splits = splits < 1 ? 1 : splits;
        kp.n_splits = splits;
        static std::atomic<unsigned long long> configured_kvt{0};
        if (int rr = opt_in_dyn_smem(kv_tc_kernel, KVT_SMEM, configured_kvt)) return rr;
        kv_tc_kernel<<<dim3(splits, g.G, 2), 192, KVT_SMEM, stream>>>(kp);
        kv_pack_kernel<<<dim3(KV_REC, g.G), 256, 0, stream>>>(w.kv_partial, gt, sel, g.max_chunks, w.kvb_h, x3 ? w.kvb_l : nullptr, splits);
      } else {
        kv_partial_kernel<<<dim3(g.max_chunks, g.G), 128, 0, stream>>>(w.kf, w.vf, gt, sel, w.kv_partial, g.max_chunks);
        kv_pack_kernel<<<dim3(KV_REC, g.G), 256, 0, stream>>>(w.kv_partial, gt, sel, g.max_chunks, w.kvb_h, x3 ? w.kvb_l : nullptr, 0);
      }
      count_launch(2);
      OPB_CHECK_CUDA(cudaGetLastError());
      // OPB200_ATTN_MERGE_FUSED = 0 / 1 forces the separate launches / the fused kernel; unset: fused from two tiles per CTA on
      static const int am_env = getenv("OPB200_ATTN_MERGE_FUSED") ? atoi(getenv("OPB200_ATTN_MERGE_FUSED")) : -1;
      if (x3 && (am_env >= 0 ? am_env != 0 : qc >= kMlpFusedMinRounds * num_sms())) {
        AttnMergeParams ap;
        PlanBuilder tb(true);
        int r;
        if ((r = tb.make(&ap.q_hi, w.qh, R, C, C, BLOCK_M))) return r;
        if ((r = tb.make(&ap.q_lo, w.ql, R, C, C, BLOCK_M))) return r;
        if ((r = tb.make(&ap.kv_hi, w.kvb_h, static_cast<uint64_t>(g.G) * 2 * 144, 128, 128, 144))) return r;
        if ((r = tb.make(&ap.kv_lo, w.kvb_l, static_cast<uint64_t>(g.G) * 2 * 144, 128, 128, 144))) return r;
        if ((r = tb.make(&ap.wm_hi, lw.m_hi, C, C, C, 256))) return r;
        if ((r = tb.make(&ap.wm_lo, lw.m_lo, C, C, C, 256))) return r;
        ap.b_row_base = cross ? w.base_cross : w.base_self;
        ap.ln = typename EpiLN<256>::Params{lw.ln1_g, lw.ln1_b, w.mh, w.ml};
        ap.eps = 1e-6f;
        ap.m_tile_begin = qb, ap.num_tiles = qc;
        static std::atomic<unsigned long long> configured_am{0};
        if (int rr = opt_in_dyn_smem(attn_merge_kernel, MF_SMEM, configured_am)) return rr;
        const int grid = qc < num_sms() ? qc : num_sms();
        attn_merge_kernel<<<grid, 64 + 128 * EPI_SPLIT, MF_SMEM, stream>>>(ap);
        count_launch();
        OPB_CHECK_CUDA(cudaGetLastError());
        merge_done = true;
      } else {
      PlanBuilder pb;
      pb.add_a(w.qh, R, C, C, 0);
      pb.add_b(w.kvb_h, static_cast<uint64_t>(g.G) * 2 * 144, 128, 128, 0, 144);
      if (x3) {
        pb.add_a(w.ql, R, C, C, 1);
        pb.add_b(w.kvb_l, static_cast<uint64_t>(g.G) * 2 * 144, 128, 128, 1, 144);
      }
      if (pb.status) return pb.status;
      if (x3) {  // K = 128 per n tile: the three products as three K segments (small; not worth the fused mainloop)
        pb.seg(0, 1, 0, 0, 2);
        pb.seg(1, 0, 0, 0, 2);
      }
      pb.seg(0, 0, 0, 0, 2);
      pb.finish();
      pb.plan.a_k_per_ntile = 128;
      pb.plan.m_tile_begin = qb;
      pb.plan.num_n_tiles = 2;
      pb.plan.num_tiles = qc * 2;
      pb.plan.b_rows_per_ntile = 144;
      pb.plan.b_row_base = cross ? w.base_cross : w.base_self;
      EpiAttn::Params ep{w.aoh, x3 ? w.aol : nullptr, 1e-6f};
      if (int r = launch_gemm<144, 4, 1, EpiAttn, EPI_SPLIT>(pb.plan, ep, stream)) return r;
      }
    } else {
      set_last_error("big-group attention is implemented for d_model 256 only");
      return OPB_ERR_ARG;
    }
  } else {
    if constexpr (C == 128) {
      const int row1 = g.row1;
      SmallAttnJobs jobs;
      int nj = 0;
      for (int s = 0; s < 2; ++s) {
        if (!(q_streams & (1 << s))) continue;
        const int ss = cross ? 1 - s : s;
        jobs.job[nj++] = SmallAttnJob{s == 0 ? 0 : row1, s == 0 ? g.len0 : g.len1, ss == 0 ? 0 : row1,
                                      ss == 0 ? g.len0 : g.len1};
      }
      if (nj > 0) {
        attn_small_kernel<<<dim3(g.N, nj), 128, 0, stream>>>(w.qf, w.kf, w.vf, jobs, g.N, 1e-6f, w.aoh, x3 ? w.aol : nullptr);
        count_launch();
      }
      OPB_CHECK_CUDA(cudaGetLastError());
    } else {
      set_last_error("small-group attention is implemented for d_model 128 only");
      return OPB_ERR_ARG;
    }
  }
  const int ub = qb, uc = qc;  // only the tiles of the streams being updated run the row-wise GEMMs
  // OPB200_MLP_FUSED = 0 / 1 forces the two-launch / the fused path; unset: fused when the CTAs get kMlpFusedMinRounds tiles
  // each or more (a CTA's tiles run back to back, so a short list ends in a round that is mostly empty)
  static const int mlp_fused_env = getenv("OPB200_MLP_FUSED") ? atoi(getenv("OPB200_MLP_FUSED")) : -1;
  const bool mlp_fused = mlp_fused_env >= 0 ? mlp_fused_env != 0 : uc >= kMlpFusedMinRounds * num_sms();
  // fine level: MERGE + LN1 run inside the fused MLP kernel as its phase 0 (OPB200_MLP_PHASE0=0: as a launch of their own)
  static const bool phase0_env = getenv("OPB200_MLP_PHASE0") ? atoi(getenv("OPB200_MLP_PHASE0")) != 0 : true;
  const bool phase0 = C == 128 && mlp_fused && x3 && !merge_done && phase0_env;
  // ---- 5. MERGE + LN1
  if (!merge_done && !phase0) {
    PlanBuilder pb(x3);
    pb.add_a(w.aoh, R, C, C, 0);
    pb.add_b(lw.m_hi, C, C, C, 0, C);
    if (x3) {
      pb.add_a(w.aol, R, C, C, 1);
      pb.add_b(lw.m_lo, C, C, C, 1, C);
    }
    if (pb.status) return pb.status;
    add_segs(pb, 0, 1, 0, 1, 0, 0, kbC);
    pb.finish();
    pb.plan.m_tile_begin = ub;
    pb.plan.num_n_tiles = 1;
    pb.plan.num_tiles = uc;
    pb.plan.b_rows_per_ntile = C;
    typename EpiLN<C>::Params ep{lw.ln1_g, lw.ln1_b, w.mh, x3 ? w.ml : nullptr};
    if (int r = x3 ? launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiLN<C>, EPI_SPLIT, true>(pb.plan, ep, stream)
                   : launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiLN<C>, EPI_SPLIT>(pb.plan, ep, stream))
      return r;
  }
  // ---- 6 + 7 in one launch (split precision): the hidden tile stays on the SM
  if (mlp_fused && x3) {
    MlpFusedParams<C> fp;
    PlanBuilder tb(true);
    int r;
    if ((r = tb.make(&fp.x_hi, w.xh, R, C, C, BLOCK_M))) return r;
    if ((r = tb.make(&fp.x_lo, w.xl, R, C, C, BLOCK_M))) return r;
    if ((r = tb.make(&fp.m_hi, w.mh, R, C, C, BLOCK_M))) return r;
    if ((r = tb.make(&fp.m_lo, w.ml, R, C, C, BLOCK_M))) return r;
    if ((r = tb.make(&fp.w1_hi, lw.w1_hi, 2 * C, 2 * C, 2 * C, 256))) return r;
    if ((r = tb.make(&fp.w1_lo, lw.w1_lo, 2 * C, 2 * C, 2 * C, 256))) return r;
    if ((r = tb.make(&fp.w2_hi, lw.w2_hi, C, 2 * C, 2 * C, C))) return r;
    if ((r = tb.make(&fp.w2_lo, lw.w2_lo, C, 2 * C, 2 * C, C))) return r;
    fp.ln = typename EpiLNRes<C>::Params{lw.ln2_g, lw.ln2_b, w.xh, w.xl};
    fp.m_tile_begin = ub, fp.num_tiles = uc;
    fp.phase0 = phase0 ? 1 : 0;
    fp.ln1_g = lw.ln1_g, fp.ln1_b = lw.ln1_b;
    if (phase0) {
      if ((r = tb.make(&fp.ao_hi, w.aoh, R, C, C, BLOCK_M))) return r;
      if ((r = tb.make(&fp.ao_lo, w.aol, R, C, C, BLOCK_M))) return r;
      if ((r = tb.make(&fp.wm_hi, lw.m_hi, C, C, C, C))) return r;
      if ((r = tb.make(&fp.wm_lo, lw.m_lo, C, C, C, C))) return r;
    } else {
      fp.ao_hi = fp.ao_lo = fp.x_hi;
      fp.wm_hi = fp.wm_lo = fp.w2_hi;
    }
    if (uc > 0) {
      static std::atomic<unsigned long long> configured_mf{0};
      if (int rr = opt_in_dyn_smem(mlp_fused_kernel<C>, MF_SMEM, configured_mf)) return rr;
      const int grid = uc < num_sms() ? uc : num_sms();
      mlp_fused_kernel<C><<<grid, 64 + 128 * EPI_SPLIT, MF_SMEM, stream>>>(fp);
      count_launch();
      OPB_CHECK_CUDA(cudaGetLastError());
    }
    return OPB_OK;
  }
  // ---- 6. MLP1 + ReLU   (K = [x | msg], N = 2C as two tiles of C)
  {
    PlanBuilder pb(x3);
    pb.add_a(w.xh, R, C, C, 0);
    pb.add_a(w.mh, R, C, C, 2);
    pb.add_b(lw.w1_hi, 2 * C, 2 * C, 2 * C, 0, C);
    if (x3) {
      pb.add_a(w.xl, R, C, C, 1);
      pb.add_a(w.ml, R, C, C, 3);
      pb.add_b(lw.w1_lo, 2 * C, 2 * C, 2 * C, 1, C);
    }
    if (pb.status) return pb.status;
    add_segs(pb, 0, 1, 0, 1, 0, 0, kbC);
    add_segs(pb, 2, 3, 0, 1, 0, C, kbC);
    pb.finish();
    pb.plan.m_tile_begin = ub;
    pb.plan.num_n_tiles = 2;
    pb.plan.num_tiles = uc * 2;
    pb.plan.b_rows_per_ntile = C;
    typename EpiRelu<C>::Params ep{w.hh, x3 ? w.hl : nullptr, 2 * C};
    if (int r = x3 ? launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiRelu<C>, EPI_SPLIT, true>(pb.plan, ep, stream)
                   : launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiRelu<C>, EPI_SPLIT>(pb.plan, ep, stream))
      return r;
  }
  // ---- 7. MLP2 + LN2 + residual (in place on the master copy and its planes)
  {
    PlanBuilder pb(x3);
    pb.add_a(w.hh, R, 2 * C, 2 * C, 0);
    pb.add_b(lw.w2_hi, C, 2 * C, 2 * C, 0, C);
    if (x3) {
      pb.add_a(w.hl, R, 2 * C, 2 * C, 1);
      pb.add_b(lw.w2_lo, C, 2 * C, 2 * C, 1, C);
    }
    if (pb.status) return pb.status;
    add_segs(pb, 0, 1, 0, 1, 0, 0, 2 * kbC);
    pb.finish();
    pb.plan.m_tile_begin = ub;
    pb.plan.num_n_tiles = 1;
    pb.plan.num_tiles = uc;
    pb.plan.b_rows_per_ntile = C;
    typename EpiLNRes<C>::Params ep{lw.ln2_g, lw.ln2_b, w.xh, w.xl};
    if (int r = x3 ? launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiLNRes<C>, EPI_SPLIT, true>(pb.plan, ep, stream)
                   : launch_gemm<C, (C == 256 ? 4 : 6), 1, EpiLNRes<C>, EPI_SPLIT>(pb.plan, ep, stream))
      return r;
  }
  return OPB_OK;
}

}  // namespace opb

using namespace opb;

extern "C" {

size_t opb200_packed_layer_bytes(int d_model) { return packed_layer_bytes(d_model); }

int opb200_pack_layer_weights(const float* wq, const float* wk, const float* wv, const float* wm, const float* w1,
                              const float* w2, const float* ln1_w, const float* ln1_b, const float* ln2_w,
                              const float* ln2_b, int C, void* packed, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(C == 256 || C == 128, "d_model must be 256 or 128");
  OPB_CHECK_ARG(wq && wk && wv && wm && w1 && w2 && ln1_w && ln1_b && ln2_w && ln2_b && packed, "null pointer");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LayerWeights v = view_layer(packed, C);
  auto nc = [](const __nv_bfloat16* p) { return const_cast<__nv_bfloat16*>(p); };
  const size_t cc = static_cast<size_t>(C) * C;
  int r;
  if ((r = split_planes(wq, C, 1, nullptr, 0, C, C, nc(v.qkv_hi), nc(v.qkv_lo), nullptr, C, stream))) return r;
  if ((r = split_planes(wk, C, 1, nullptr, 0, C, C, nc(v.qkv_hi) + cc, nc(v.qkv_lo) + cc, nullptr, C, stream))) return r;
  if ((r = split_planes(wv, C, 1, nullptr, 0, C, C, nc(v.qkv_hi) + 2 * cc, nc(v.qkv_lo) + 2 * cc, nullptr, C, stream))) return r;
  if ((r = split_planes(wm, C, 1, nullptr, 0, C, C, nc(v.m_hi), nc(v.m_lo), nullptr, C, stream))) return r;
  if ((r = split_planes(w1, 2 * C, 1, nullptr, 0, 2 * C, 2 * C, nc(v.w1_hi), nc(v.w1_lo), nullptr, 2 * C, stream))) return r;
  if ((r = split_planes(w2, 2 * C, 1, nullptr, 0, C, 2 * C, nc(v.w2_hi), nc(v.w2_lo), nullptr, 2 * C, stream))) return r;
  float* f = const_cast<float*>(v.ln1_g);
  OPB_CHECK_CUDA(cudaMemcpyAsync(f, ln1_w, C * 4, cudaMemcpyDeviceToDevice, stream));
  OPB_CHECK_CUDA(cudaMemcpyAsync(f + C, ln1_b, C * 4, cudaMemcpyDeviceToDevice, stream));
  OPB_CHECK_CUDA(cudaMemcpyAsync(f + 2 * C, ln2_w, C * 4, cudaMemcpyDeviceToDevice, stream));
  OPB_CHECK_CUDA(cudaMemcpyAsync(f + 3 * C, ln2_b, C * 4, cudaMemcpyDeviceToDevice, stream));
  return OPB_OK;
}

size_t opb200_transformer_workspace_bytes(int d_model, int n_batch, int len0, int len1, int big_groups) {
  Geometry g = make_geometry(d_model, n_batch, len0, len1, big_groups);
  return carve(nullptr, g).bytes;
}

int opb200_transformer(const void* const* host_layer_weights, const int* host_layer_kinds, int n_layers, int schedule,
                       int d_model, int nhead, int n_batch, int len0, int len1, int big_groups, const float* x0,
                       int64_t x0_sn, int64_t x0_sl, int64_t x0_sc, const float* add0, const float* x1, int64_t x1_sn,
                       int64_t x1_sl, int64_t x1_sc, const float* add1, const uint8_t* mask0, const uint8_t* mask1,
                       float* out0, float* out1, void* workspace, size_t workspace_bytes, int precision,
                       void* stream_) {
  if (int r = check_device()) return r;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  OPB_CHECK_ARG(d_model == 256 || d_model == 128, "d_model must be 256 (coarse) or 128 (fine)");
  OPB_CHECK_ARG(nhead == 8, "nhead must be 8");
  OPB_CHECK_ARG(precision == 1 || precision == 3, "precision must be 1 or 3");
  OPB_CHECK_ARG(schedule == 0 || schedule == 1, "schedule: 0 = OnePosePlus (simultaneous), 1 = LoFTR (sequential cross)");
  OPB_CHECK_ARG(n_batch >= 1 && len0 >= 1 && len1 >= 1, "empty input");
  OPB_CHECK_ARG(!big_groups || d_model == 256, "big groups need d_model 256");
  OPB_CHECK_ARG(big_groups || d_model == 128, "small groups need d_model 128");
  OPB_CHECK_ARG(!big_groups || 2 * n_batch <= 64, "at most 32 batch elements per call with big groups");
  OPB_CHECK_ARG(big_groups || (len0 <= 128 && len1 <= 128), "small groups hold at most 128 tokens");
  OPB_CHECK_ARG(x0 && x1 && workspace && host_layer_weights && host_layer_kinds, "null pointer");
  OPB_CHECK_ARG((out0 == nullptr) == (out1 == nullptr), "out0 / out1 must both be given or both be null");
  Geometry g = make_geometry(d_model, n_batch, len0, len1, big_groups);
  Workspace w = carve(workspace, g);
  if (w.bytes > workspace_bytes) {
    set_last_error("workspace too small: need " + std::to_string(w.bytes) + " bytes");
    return OPB_ERR_WORKSPACE;
  }
  const int C = d_model;
  // stage in
  const int row1 = g.row1;
  if (!big_groups) {  // zero the pad rows at the end of each stream region (they flow through the GEMMs)
    const size_t pad_lo[2] = {static_cast<size_t>(n_batch) * len0, static_cast<size_t>(row1) + static_cast<size_t>(n_batch) * len1};
    const size_t pad_hi[2] = {static_cast<size_t>(row1), static_cast<size_t>(g.R)};
    for (int s = 0; s < 2; ++s) {
      const size_t n = pad_hi[s] - pad_lo[s];
      if (n == 0) continue;
      OPB_CHECK_CUDA(cudaMemsetAsync(w.xh + pad_lo[s] * C, 0, n * C * 2, stream));
      OPB_CHECK_CUDA(cudaMemsetAsync(w.xl + pad_lo[s] * C, 0, n * C * 2, stream));
      OPB_CHECK_CUDA(cudaMemsetAsync(w.row_mask + pad_lo[s], 0, n, stream));
    }
  }
  auto stage_in = [&](const float* x, int64_t sn, int64_t sl, int64_t sc, const float* add, const uint8_t* mask, int len,
                      int P, int row_begin) {
    const bool rows = sc == 1 && (sl % 4 == 0 || len == 1) && sn % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 &&
                      (add == nullptr || (reinterpret_cast<uintptr_t>(add) & 15) == 0);
    if (rows) {
      copy_in_rows_kernel<<<dim3((P + 7) / 8, 1, n_batch), 256, 0, stream>>>(x, sn, sl, add, mask, len, P, C, row_begin, w.xh,
                                                                               w.xl, w.row_mask);
    } else {  // channels-first / arbitrary strides: 32 x 32 shared-memory transpose
      copy_in_kernel<<<dim3((C + 31) / 32, (P + 31) / 32, n_batch), dim3(32, 8), 0, stream>>>(x, sn, sl, sc, add, mask, len, P, C,
                                                                                             row_begin, w.xh, w.xl, w.row_mask);
    }
  };
  stage_in(x0, x0_sn, x0_sl, x0_sc, add0, mask0, len0, g.P0, 0);
  stage_in(x1, x1_sn, x1_sl, x1_sc, add1, mask1, len1, g.P1, row1);
  count_launch(2 + (big_groups ? 1 : 0));
  if (big_groups) {  // off-diagonal / unused entries of the block-diagonal attention operand stay zero
    OPB_CHECK_CUDA(cudaMemsetAsync(w.kvb_h, 0, static_cast<size_t>(g.G) * 2 * 144 * 128 * 2, stream));
    OPB_CHECK_CUDA(cudaMemsetAsync(w.kvb_l, 0, static_cast<size_t>(g.G) * 2 * 144 * 128 * 2, stream));
  }
  if (big_groups)
    fill_row_base_kernel<<<1, 256, 0, stream>>>(w.base_self, w.base_cross, g.tiles0, g.tiles1, n_batch, 2 * 144);
  OPB_CHECK_CUDA(cudaGetLastError());

  for (int l = 0; l < n_layers; ++l) {
    LayerWeights lw = view_layer(host_layer_weights[l], C);
    const int kind = host_layer_kinds[l];
    const bool cross = kind == 1;
    OPB_CHECK_ARG(kind >= 0 && kind <= 3, "layer kind must be 0 (self), 1 (cross), 2 (self, stream 1 only) or 3 (self, stream 0 only)");
    int r;
    if (kind >= 2) {  // self layer applied to one stream; the other stream's rows are left as they are
      const int qs = kind == 2 ? 2 : 1;
      r = (C == 256) ? run_step<256>(g, w, lw, false, qs, precision, stream) : run_step<128>(g, w, lw, false, qs, precision, stream);
    } else if (cross && schedule == 1) {
      // LoFTR: feat0 = layer(feat0, feat1); then feat1 = layer(feat1, feat0_new)
      r = (C == 256) ? run_step<256>(g, w, lw, true, 1, precision, stream) : run_step<128>(g, w, lw, true, 1, precision, stream);
      if (r) return r;
      r = (C == 256) ? run_step<256>(g, w, lw, true, 2, precision, stream) : run_step<128>(g, w, lw, true, 2, precision, stream);
    } else {
      r = (C == 256) ? run_step<256>(g, w, lw, cross, 3, precision, stream)
                     : run_step<128>(g, w, lw, cross, 3, precision, stream);
    }
    if (r) return r;
  }
  if (out0 == nullptr) return OPB_OK;  // the caller consumes the planes in the workspace (opb200_transformer_planes)
  const int threads = 256;
  copy_out_kernel<<<dim3((static_cast<long long>(len0) * C / 8 + threads - 1) / threads, 1, n_batch), threads, 0, stream>>>(
      w.xh, w.xl, len0, g.P0, C, 0, out0);
  copy_out_kernel<<<dim3((static_cast<long long>(len1) * C / 8 + threads - 1) / threads, 1, n_batch), threads, 0, stream>>>(
      w.xh, w.xl, len1, g.P1, C, row1, out1);
  count_launch(2);
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_transformer_planes(void* workspace, int d_model, int n_batch, int len0, int len1, int big_groups, void** planes,
                              int* padded_len) {
  OPB_CHECK_ARG(workspace && planes && padded_len, "null pointer");
  OPB_CHECK_ARG(d_model == 256 || d_model == 128, "d_model must be 256 or 128");
  Geometry g = make_geometry(d_model, n_batch, len0, len1, big_groups);
  Workspace w = carve(workspace, g);
  planes[0] = w.xh;
  planes[1] = w.xl;
  planes[2] = w.xh + static_cast<size_t>(g.row1) * d_model;
  planes[3] = w.xl + static_cast<size_t>(g.row1) * d_model;
  padded_len[0] = g.P0;
  padded_len[1] = g.P1;
  return OPB_OK;
}

}  // extern "C"
