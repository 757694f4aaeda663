// Warp-specialised persistent GEMM mainloop for sm_100a.
//
//   D[128 x BN] (fp32, TMEM)  =  sum over k-blocks  A[128 x 64] (bf16, smem via TMA)  *  B[BN x 64]^T
//
// One CTA per SM, static round-robin tile schedule.  Warp 0 = TMA producer (one elected lane),
// warp 1 = tcgen05.mma issuer (one elected lane) + TMEM owner, warps 2.. = epilogue (TMEM -> registers
// -> fused math -> global).  Smem ring of STAGES {A,B} tiles (128B-swizzled, K-major) guarded by
// full/empty mbarriers; accumulators double-buffered in TMEM (when 2*BN <= 512 columns) guarded by
// tmem_full/tmem_empty mbarriers, so the MMAs of tile i+1 overlap the epilogue of tile i.
//
// The K loop is described by "segments": each names an A tensor map, a B tensor map and a starting
// K coordinate in each.  This is how the split-precision product (A_hi*B_hi + A_hi*B_lo + A_lo*B_hi,
// fp32-grade accuracy from bf16 tensor cores) and the concat-input MLP (K = [x | msg]) are expressed
// without separate kernels: they are just longer K loops into the same accumulator.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include "ptx.cuh"

namespace opb {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // bf16 elements per k-block = one 128-byte swizzle row
constexpr int kNumMaps = 4;
constexpr int kMaxSegs = 8;

struct KSeg {
  int a_map, b_map;    // indices into GemmPlan::a_maps / b_maps (fused split mode: the hi planes)
  int a_map2, b_map2;  // fused split mode only: the lo planes
  int a_k0, b_k0;      // starting K coordinate (elements) in each
  int nkb;             // number of k-blocks (64 wide; 32 wide in fused split mode)
};

struct GemmPlan {
  CUtensorMap a_maps[kNumMaps];
  CUtensorMap b_maps[kNumMaps];
  KSeg segs[kMaxSegs];
  int nsegs;
  int total_kb;
  int num_tiles;          // = (number of m tiles) * num_n_tiles
  int num_n_tiles;        // tile t -> (m_tile_begin + t / num_n_tiles, n_tile_begin + t % num_n_tiles)
  int m_tile_begin;
  int n_tile_begin;
  int a_k_per_ntile;      // extra A K-offset per n tile (block-diagonal B: each n tile reads its own K slice)
  const int* b_row_base;  // optional per-m-tile base row into B (per-group / per-batch B operand)
  int b_rows_per_ntile;   // row distance between consecutive n tiles of B
  long long* trace;       // debug (opb200_debug_trace): per CTA and tile {first TMA issued, MMA start, accumulator full,
                          // epilogue done} in SM clock cycles; null in normal use
};
constexpr int kTraceTilesPerCta = 32;

// GROUPS = number of epilogue warpgroups (4 warps each).  With one group the accumulator is double-buffered
// (when it fits) so the next tile's MMAs overlap the epilogue; with G > 1 groups there are G accumulator
// stages, group g owns stage g and handles every G-th tile of the CTA, so up to G epilogues run concurrently
// (4*G warps hide each other's TMEM / memory latency) while the MMA warp fills the next free stage.
template <int BN, int GROUPS = 1>
struct GemmCfg {
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
  static constexpr int B_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int ACC_STAGES = GROUPS > 1 ? GROUPS : ((2 * BN <= 512) ? 2 : 1);
  static_assert(ACC_STAGES * BN <= 512, "accumulator stages exceed the 512 TMEM columns");
  static constexpr int TMEM_COLS = (ACC_STAGES * BN <= 32)    ? 32
                                   : (ACC_STAGES * BN <= 64)  ? 64
                                   : (ACC_STAGES * BN <= 128) ? 128
                                   : (ACC_STAGES * BN <= 256) ? 256
                                                              : 512;
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N for M=128 must be a multiple of 16 in [16,256]");
  static_assert(STAGE_BYTES % 1024 == 0, "stage buffers must stay 1024-byte aligned for SWIZZLE_128B");
};

__host__ __device__ constexpr int gemm_smem_bytes(int BN, int stages, int epi_scratch /* all groups */) {
  return 1024 /*alignment slack*/ + stages * (BLOCK_M * BLOCK_K * 2 + BN * BLOCK_K * 2) + 256 /*barriers*/ + epi_scratch;
}

// Context handed to the epilogue functor for one output tile.
struct EpiCtx {
  int m_tile, n_tile;
  uint32_t tmem_acc;  // TMEM address of column 0 of this tile's accumulator, lane field already = 32*quarter
  int quarter;        // TMEM lane quarter of this warp (rows quarter*32 .. +32 of the tile)
  int lane;           // lane id; this thread owns tile row quarter*32 + lane
  int epi_warp;       // 0 .. 3 within the group
  int group;          // epilogue group (named barriers 1 + 4*group .. 4 + 4*group are reserved for it)
  int part;           // with SPLIT > 1: which slice of the tile's columns this warp handles (0 .. SPLIT-1)
  uint8_t* scratch;   // Epi::kSmemBytes of shared memory private to this epilogue group
  int scratch_small = 0;  // 1: only half of kLnScratch is there (mlp_fused_kernel): the LayerNorm statistics take two more barriers
};

// SPLIT = warps per TMEM lane quarter inside one group: the tile's columns are divided among them, so a
// single tile's epilogue runs on 4*SPLIT warps (what matters when a CTA has only one or two tiles).
//
// F3 = fused split-precision mainloop.  The segment form of A_hi*B_lo + A_lo*B_hi + A_hi*B_hi walks K three times
// and so fills shared memory with six operand tiles per k-block for three MMA groups; these GEMMs are bound by the
// L2 -> SM fill rate, not by the tensor pipe.  With F3 one stage holds the four distinct tiles of a 32-wide k-block
// {A_hi, A_lo, B_hi, B_lo} (64-byte swizzle; same stage size as one 64-wide {A, B} pair) and the MMA warp issues
// the three products from it: 2/3 of the shared-memory fill per flop.
template <int BN, int STAGES, int GROUPS, int SPLIT, class Epi, bool F3 = false>
__global__ void __launch_bounds__(64 + 128 * GROUPS * SPLIT, 1)
gemm_kernel(const __grid_constant__ GemmPlan plan, const typename Epi::Params ep) {
  using Cfg = GemmCfg<BN, GROUPS>;
  static_assert(GROUPS >= 1 && GROUPS <= 4, "1..4 epilogue groups");
  static_assert(SPLIT == 1 || SPLIT == 2 || SPLIT == 4, "1, 2 or 4 warps per lane quarter");
  static_assert(GROUPS * SPLIT <= 4, "at most 16 epilogue warps");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 4);
  uint8_t* scratch = smem + STAGES * Cfg::STAGE_BYTES + 256;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < kNumMaps; ++i) {
      ptx::prefetch_tmap(&plan.a_maps[i]);
      ptx::prefetch_tmap(&plan.b_maps[i]);
    }
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 4; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1);
      ptx::mbar_init(&tempty_bar[s], 4 * SPLIT);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto tile_coords = [&](int t, int& m_tile, int& n_tile) {
    const int q = t / plan.num_n_tiles;
    m_tile = plan.m_tile_begin + q;
    n_tile = plan.n_tile_begin + (t - q * plan.num_n_tiles);
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < plan.num_tiles; t += gridDim.x) {
        int m_tile, n_tile;
        tile_coords(t, m_tile, n_tile);
        const int a_row = m_tile * BLOCK_M;
        const int b_row = (plan.b_row_base ? plan.b_row_base[m_tile] : 0) + n_tile * plan.b_rows_per_ntile;
        if (plan.trace != nullptr) {
          const int it = (t - blockIdx.x) / gridDim.x;
          if (it < kTraceTilesPerCta) plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 0] = clock64();
        }
        for (int s = 0; s < plan.nsegs; ++s) {
          const KSeg seg = plan.segs[s];
          for (int kb = 0; kb < seg.nkb; ++kb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            ptx::mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
            uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
            if constexpr (F3) {
              const int ka = seg.a_k0 + n_tile * plan.a_k_per_ntile + kb * 32, kbb = seg.b_k0 + kb * 32;
              ptx::tma_load_2d(sa, &plan.a_maps[seg.a_map], &full_bar[stage], ka, a_row);
              ptx::tma_load_2d(sa + Cfg::A_BYTES / 2, &plan.a_maps[seg.a_map2], &full_bar[stage], ka, a_row);
              ptx::tma_load_2d(sa + Cfg::A_BYTES, &plan.b_maps[seg.b_map], &full_bar[stage], kbb, b_row);
              ptx::tma_load_2d(sa + Cfg::A_BYTES + Cfg::B_BYTES / 2, &plan.b_maps[seg.b_map2], &full_bar[stage], kbb, b_row);
            } else {
              ptx::tma_load_2d(sa, &plan.a_maps[seg.a_map], &full_bar[stage],
                               seg.a_k0 + n_tile * plan.a_k_per_ntile + kb * BLOCK_K, a_row);
              ptx::tma_load_2d(sa + Cfg::A_BYTES, &plan.b_maps[seg.b_map], &full_bar[stage], seg.b_k0 + kb * BLOCK_K,
                               b_row);
            }
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = ptx::make_idesc_bf16(BLOCK_M, BN);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int t = blockIdx.x; t < plan.num_tiles; t += gridDim.x, ++it) {
      const int acc = it % Cfg::ACC_STAGES;
      const uint32_t acc_phase = (it / Cfg::ACC_STAGES) & 1;
      ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
      ptx::tc_fence_after();
      if (plan.trace != nullptr && it < kTraceTilesPerCta && lane == 0)
        plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 1] = clock64();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = 0; kb < plan.total_kb; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t a_addr = ptx::smem_u32(smem + stage * Cfg::STAGE_BYTES);
          if constexpr (F3) {
            const uint64_t a_hi = ptx::make_desc_sw64(a_addr), a_lo = ptx::make_desc_sw64(a_addr + Cfg::A_BYTES / 2);
            const uint64_t b_hi = ptx::make_desc_sw64(a_addr + Cfg::A_BYTES);
            const uint64_t b_lo = ptx::make_desc_sw64(a_addr + Cfg::A_BYTES + Cfg::B_BYTES / 2);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d_tmem, a_hi + 2 * k, b_hi + 2 * k, idesc, (kb | k) != 0);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d_tmem, a_lo + 2 * k, b_hi + 2 * k, idesc, true);
#pragma unroll
            for (int k = 0; k < 2; ++k) ptx::mma_bf16(d_tmem, a_hi + 2 * k, b_lo + 2 * k, idesc, true);
          } else {
            const uint64_t a_desc = ptx::make_desc_sw128(a_addr);
            const uint64_t b_desc = ptx::make_desc_sw128(a_addr + Cfg::A_BYTES);
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) {
              // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in the (addr >> 4) field
              ptx::mma_bf16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
            }
          }
          ptx::mma_commit(&empty_bar[stage]);
          if (kb == plan.total_kb - 1) ptx::mma_commit(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    EpiCtx ctx;
    ctx.quarter = warp & 3;
    ctx.lane = lane;
    ctx.epi_warp = (warp - 2) & 3;
    ctx.group = ((warp - 2) >> 2) / SPLIT;
    ctx.part = ((warp - 2) >> 2) % SPLIT;
    ctx.scratch = scratch + ctx.group * Epi::kSmemBytes;
    int it = 0;
    for (int t = blockIdx.x; t < plan.num_tiles; t += gridDim.x, ++it) {
      if (GROUPS > 1 && it % GROUPS != ctx.group) continue;
      const int acc = it % Cfg::ACC_STAGES;
      const uint32_t acc_phase = (it / Cfg::ACC_STAGES) & 1;
      tile_coords(t, ctx.m_tile, ctx.n_tile);
      Epi::prefetch(ep, ctx);  // global operands of the epilogue that do not depend on the accumulator
      ptx::mbar_wait(&tfull_bar[acc], acc_phase);
      ptx::tc_fence_after();
      const bool tr = plan.trace != nullptr && it < kTraceTilesPerCta && warp == 2 && lane == 0;
      if (tr) plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 2] = clock64();
      ctx.tmem_acc = tmem_base + acc * BN + (static_cast<uint32_t>(ctx.quarter * 32) << 16);
      Epi::run(ep, ctx);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[acc]);
      if (tr) plan.trace[(blockIdx.x * kTraceTilesPerCta + it) * 4 + 3] = clock64();
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// bf16 split helpers -------------------------------------------------------------------------
__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}
__device__ __forceinline__ uint32_t pack_bf16(__nv_bfloat16 a, __nv_bfloat16 b) {
  return static_cast<uint32_t>(__bfloat16_as_ushort(a)) | (static_cast<uint32_t>(__bfloat16_as_ushort(b)) << 16);
}

// Write 32 consecutive fp32 values of one row as bf16 hi (and lo, if lo != nullptr) planes.
__device__ __forceinline__ void store_planes32(const float (&v)[32], __nv_bfloat16* hi, __nv_bfloat16* lo) {
  uint32_t h[16], l[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    __nv_bfloat16 h0, l0, h1, l1;
    split_bf16(v[2 * i], h0, l0);
    split_bf16(v[2 * i + 1], h1, l1);
    h[i] = pack_bf16(h0, h1);
    l[i] = pack_bf16(l0, l1);
  }
  uint4* ph = reinterpret_cast<uint4*>(hi);
#pragma unroll
  for (int i = 0; i < 4; ++i) ph[i] = make_uint4(h[4 * i], h[4 * i + 1], h[4 * i + 2], h[4 * i + 3]);
  if (lo != nullptr) {
    uint4* pl = reinterpret_cast<uint4*>(lo);
#pragma unroll
    for (int i = 0; i < 4; ++i) pl[i] = make_uint4(l[4 * i], l[4 * i + 1], l[4 * i + 2], l[4 * i + 3]);
  }
}

// Same as store_planes32 with 256-bit stores (hi / lo must be 32-byte aligned).
__device__ __forceinline__ void store_planes32_v8(const float (&v)[32], __nv_bfloat16* hi, __nv_bfloat16* lo) {
  uint32_t h[16], l[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const __nv_bfloat162 hh = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    const float2 hf = __bfloat1622float2(hh);
    const __nv_bfloat162 ll = __floats2bfloat162_rn(v[2 * i] - hf.x, v[2 * i + 1] - hf.y);
    h[i] = *reinterpret_cast<const uint32_t*>(&hh);
    l[i] = *reinterpret_cast<const uint32_t*>(&ll);
  }
  ptx::st_global_v8(hi, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
  ptx::st_global_v8(hi + 16, h[8], h[9], h[10], h[11], h[12], h[13], h[14], h[15]);
  if (lo != nullptr) {
    ptx::st_global_v8(lo, l[0], l[1], l[2], l[3], l[4], l[5], l[6], l[7]);
    ptx::st_global_v8(lo + 16, l[8], l[9], l[10], l[11], l[12], l[13], l[14], l[15]);
  }
}
// ---- quad-transposed global access --------------------------------------------------------------------------
// In the TMEM epilogues lane i owns row i, so a 32-byte store per lane touches 32 different 128-byte lines per
// instruction; the LSU serialises on lines, and that -- not bytes -- bounds the LayerNorm epilogues.  A 4 x 4 transpose
// of 32-byte items inside each quad of lanes (two xor-shuffle rounds) lets the four lanes of a quad write the four
// sectors of ONE line per instruction: 8 lines per instruction instead of 32.
// Before: it[j] = sector j of this lane's row.  After: it[k] = sector (lane % 4) of row (lane & ~3) + k.  Involution.
__device__ __forceinline__ void quad_transpose(uint32_t (&it)[4][8], int lane) {
  const bool b0 = lane & 1, b1 = lane & 2;
#pragma unroll
  for (int j = 0; j < 4; j += 2) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const uint32_t x = b0 ? it[j][r] : it[j + 1][r];
      const uint32_t y = __shfl_xor_sync(0xffffffffu, x, 1);
      if (b0) it[j][r] = y; else it[j + 1][r] = y;
    }
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const uint32_t x = b1 ? it[j][r] : it[j + 2][r];
      const uint32_t y = __shfl_xor_sync(0xffffffffu, x, 2);
      if (b1) it[j][r] = y; else it[j + 2][r] = y;
    }
  }
}
// 64 consecutive bf16 of this lane's row (4 sectors, already packed two per register) written through the quad
// transpose.  `tile_row0` points at column c0 of the row of lane 0 of this warp; ld = row pitch in elements.
__device__ __forceinline__ void store_row64_bf16_quad(uint32_t (&it)[4][8], __nv_bfloat16* warp_row0, long long ld, int lane) {
  quad_transpose(it, lane);
  __nv_bfloat16* base = warp_row0 + static_cast<long long>(lane & ~3) * ld + (lane & 3) * 16;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    ptx::st_global_v8(base + k * ld, it[k][0], it[k][1], it[k][2], it[k][3], it[k][4], it[k][5], it[k][6], it[k][7]);
}
__device__ __forceinline__ void load_row64_bf16_quad(uint32_t (&it)[4][8], const __nv_bfloat16* warp_row0, long long ld, int lane) {
  const __nv_bfloat16* base = warp_row0 + static_cast<long long>(lane & ~3) * ld + (lane & 3) * 16;
#pragma unroll
  for (int k = 0; k < 4; ++k) ptx::ld_global_v8(base + k * ld, it[k]);
  quad_transpose(it, lane);
}
// fp32 row slice -> packed bf16 hi / lo sectors (64 columns)
__device__ __forceinline__ void pack_planes64(const float* x, uint32_t (&h)[4][8], uint32_t (&l)[4][8]) {
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const __nv_bfloat162 hh = __floats2bfloat162_rn(x[2 * i], x[2 * i + 1]);
    const float2 hf = __bfloat1622float2(hh);
    const __nv_bfloat162 ll = __floats2bfloat162_rn(x[2 * i] - hf.x, x[2 * i + 1] - hf.y);
    h[i >> 3][i & 7] = *reinterpret_cast<const uint32_t*>(&hh);
    l[i >> 3][i & 7] = *reinterpret_cast<const uint32_t*>(&ll);
  }
}

// 32 consecutive fp32 values of one row with 256-bit stores (dst 32-byte aligned)
__device__ __forceinline__ void store_f32x32_v8(const float (&v)[32], float* dst) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
    ptx::st_global_v8(dst + 8 * i, __float_as_uint(v[8 * i]), __float_as_uint(v[8 * i + 1]), __float_as_uint(v[8 * i + 2]),
                      __float_as_uint(v[8 * i + 3]), __float_as_uint(v[8 * i + 4]), __float_as_uint(v[8 * i + 5]),
                      __float_as_uint(v[8 * i + 6]), __float_as_uint(v[8 * i + 7]));
}

__device__ __forceinline__ void load_acc32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  ptx::tmem_ld32(taddr, r);
  ptx::tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

}  // namespace opb
