// CoarseMatching: dual-softmax confidence + threshold + border + mutual-nearest-neighbour selection,
// without materialising the [L, S] confidence matrix (unless the caller asks for it).
//
// Replaces  CoarseMatching.forward / get_coarse_match
//             src/models/OnePosePlus/utils/coarse_matching.py:76-242   (3D points x query cells)
//             submodules/LoFTR/src/loftr/utils/coarse_matching.py:87-261 (image0 cells x image1 cells)
//
//   sim  = (a . b) * scale                    scale = 1 / (C * (T + eps))
//   conf = softmax_rows-over-L(sim) * softmax_cols-over-S(sim)
//        = exp(2 sim - LSErow_i - LSEcol_j)   (LSE = log-sum-exp of the row / of the column)
//
// Pass 1 (STATS): tcgen05 GEMM tiles of sim; the epilogue keeps an online (max, sum) per row in the
//   owning thread and reduces per-tile column (max, sum) through a 32-column shared-memory transpose.
//   Partials go to small [tiles, rows] / [tiles, cols] arrays; a merge kernel turns them into LSEs.
// Pass 2 (SELECT): the same GEMM tiles again (bit-identical accumulators); the epilogue forms
//   t = 2 sim - LSErow - LSEcol and appends (row, col, conf) for the rare elements with conf > thr.
//   Since sum_j conf_ij <= 1 there are < 1/thr such elements per row, so the candidate list is tiny,
//   and because any element that beats a candidate is itself a candidate, the row/column maxima over
//   candidates equal the maxima over the full matrix wherever a match can exist.
// Resolve: atomicMax of conf bits per row / column, pick the first j per row that equals both maxima
//   and passes the border rule, then an ordered compaction (ascending (b, i), as torch.where yields).
// Optional: pass 2 also streams conf to HBM (conf_matrix of the data-dict contract; training/debug).
#include <cmath>
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

int split_planes(const float* x, long long stride_r, long long stride_c, const float* add, long long add_period_rows,
                 int rows, int cols, __nv_bfloat16* hi, __nv_bfloat16* lo, float* f32out, long long ld_out,
                 cudaStream_t stream);

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kNegInf = -INFINITY;
constexpr int kNoMatch = 0x7f7f7f7f;
constexpr float kTinySum = 1.262177e-29f;  // 2^-96: single-pass sums below this may have lost terms to flush-to-zero
constexpr int SIM_BN = 128;  // rowj sentinel (what cudaMemsetAsync(0x7f) leaves)

struct SimGeom {
  int N, L, S, Lp, Sp;        // valid and padded (per batch element) row / column counts
  float scale;                // 1 / (C * (T + eps))
  const uint8_t* mask_a;      // [N, L] nullable
  const uint8_t* mask_b;      // [N, S] nullable
  int mask_fill;              // 0: sim += -1e9 (OnePosePlus), 1: sim = -1e9 (LoFTR masked_fill_)
};

// named barrier of one epilogue group (4 warps); ids 1.. are reserved for the groups
__device__ __forceinline__ void epi_bar(int group) { ptx::named_bar_sync(1 + group, 128); }

// y = log2(e) * sim with the reference's masking semantics applied to sim first
__device__ __forceinline__ float sim_log2(float acc, float scale, bool masked, int fill) {
  // explicit _rn intrinsics: the compiler must not contract these into FMAs -- both passes (and the
  // subtraction of the row/column maxima that follows) have to see the same ROUNDED value of y
  float sim = __fmul_rn(acc, scale);
  if (masked) sim = fill ? -1e9f : __fadd_rn(sim, -1e9f);
  return __fmul_rn(sim, kLog2e);
}

template <int BN>
struct EpiSimStats {
  struct Params {
    SimGeom g;
    float2* rowpart;  // [Sp/BN][N*Lp]
    float2* colpart;  // [Lp/128][N*Sp]
  };
  static constexpr int kSmemBytes = 2 * 128 * 33 * 4 + 2 * 4 * 32 * 8;
  static __device__ __forceinline__ void prefetch(const Params&, const EpiCtx&) {}
  static __device__ __forceinline__ void run(const Params& p, const EpiCtx& ctx) {
    const SimGeom& g = p.g;
    float(*sc)[128][33] = reinterpret_cast<float(*)[128][33]>(ctx.scratch);
    float2(*part)[4][32] = reinterpret_cast<float2(*)[4][32]>(ctx.scratch + 2 * 128 * 33 * 4);
    const int r = ctx.quarter * 32 + ctx.lane;
    const int grow = ctx.m_tile * BLOCK_M + r;  // padded global row
    const int n = grow / g.Lp;
    const int i = grow - n * g.Lp;
    const bool row_valid = i < g.L;
    const bool row_masked = row_valid && g.mask_a != nullptr && g.mask_a[static_cast<long long>(n) * g.L + i] == 0;
    const int mt_in_batch = ctx.m_tile - n * (g.Lp / BLOCK_M);
    float m = kNegInf, s = 0.f;
    int buf = 0;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32, buf ^= 1) {
      float y[32];
      load_acc32(ctx.tmem_acc + c0, y);
      const int j0 = ctx.n_tile * BN + c0;
      float cmax = kNegInf;
#pragma unroll
      for (int k = 0; k < 32; ++k) {
        const int j = j0 + k;
        const bool col_valid = j < g.S;
        const bool masked = row_masked || (g.mask_b != nullptr && col_valid && g.mask_b[static_cast<long long>(n) * g.S + j] == 0);
        const float v = col_valid ? sim_log2(y[k], g.scale, masked, g.mask_fill) : kNegInf;
        y[k] = v;
        cmax = fmaxf(cmax, v);
      }
      if (cmax > kNegInf) {
        const float mn = fmaxf(m, cmax);
        float acc = s * ptx::ex2(m - mn);
#pragma unroll
        for (int k = 0; k < 32; ++k) acc += ptx::ex2(y[k] - mn);
        s = acc;
        m = mn;
      }
#pragma unroll
      for (int k = 0; k < 32; ++k) sc[buf][r][k] = row_valid ? y[k] : kNegInf;
      epi_bar(ctx.group);
      {
        const int cq = ctx.epi_warp, cc = ctx.lane;  // NB: epi_warp != quarter; any bijection of the 4 warps works here
        float v[32];
        float pm = kNegInf;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          v[k] = sc[buf][cq * 32 + k][cc];
          pm = fmaxf(pm, v[k]);
        }
        float ps = 0.f;
        if (pm > kNegInf) {
#pragma unroll
          for (int k = 0; k < 32; ++k) ps += ptx::ex2(v[k] - pm);
        }
        part[buf][cq][cc] = make_float2(pm, ps);
      }
      epi_bar(ctx.group);
      if (ctx.epi_warp == 0) {
        float M = kNegInf;
#pragma unroll
        for (int q = 0; q < 4; ++q) M = fmaxf(M, part[buf][q][ctx.lane].x);
        float S = 0.f;
        if (M > kNegInf) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 pq = part[buf][q][ctx.lane];
            if (pq.x > kNegInf) S += pq.y * ptx::ex2(pq.x - M);
          }
        }
        const int j = j0 + ctx.lane;
        if (j < g.Sp) p.colpart[static_cast<long long>(mt_in_batch) * g.N * g.Sp + static_cast<long long>(n) * g.Sp + j] = make_float2(M, S);
      }
    }
    p.rowpart[static_cast<long long>(ctx.n_tile) * g.N * g.Lp + grow] = make_float2(m, s);
  }
};

// merge (max, sum) partials -> (max, log2 sum) per row and per column.  Block = 32 outputs x 8 slices of the
// part list: reads stay coalesced along the output index, each thread folds its slice online, the eight
// slices are combined through shared memory in a fixed order (deterministic).
__global__ void __launch_bounds__(256) merge_stats_kernel(const float2* __restrict__ rowpart, int n_row_parts, long long n_rows,
                                                          const float2* __restrict__ colpart, int n_col_parts, long long n_cols,
                                                          float2* __restrict__ lse_row, float2* __restrict__ lse_col,
                                                          int* __restrict__ needs_robust) {
  __shared__ float2 sm[8][33];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const long long t = static_cast<long long>(blockIdx.x) * 32 + lane;
  const long long row_blocks = (n_rows + 31) / 32;  // row outputs and column outputs never share a block
  const float2* part;
  int nparts;
  long long stride, idx;
  float2* out;
  bool valid;
  if (blockIdx.x < row_blocks) {
    part = rowpart, nparts = n_row_parts, stride = n_rows, idx = t, out = lse_row, valid = idx < n_rows;
  } else {
    part = colpart, nparts = n_col_parts, stride = n_cols, idx = t - row_blocks * 32, out = lse_col, valid = idx < n_cols;
  }
  float M = kNegInf, S = 0.f;
  if (valid) {
    for (int k = slice; k < nparts; k += 8) {
      const float2 pk = part[k * stride + idx];
      if (pk.x > kNegInf) {
        const float Mn = fmaxf(M, pk.x);
        S = S * exp2f(M - Mn) + pk.y * exp2f(pk.x - Mn);  // exp2f(-inf) = 0 covers the first term
        M = Mn;
      }
    }
  }
  sm[slice][lane] = make_float2(M, S);
  __syncthreads();
  if (slice == 0 && valid) {
    float Mt = kNegInf;
#pragma unroll
    for (int q = 0; q < 8; ++q) Mt = fmaxf(Mt, sm[q][lane].x);
    float St = 0.f;
    if (Mt > kNegInf) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float2 pq = sm[q][lane];
        if (pq.x > kNegInf) St += pq.y * exp2f(pq.x - Mt);
      }
    }
    // kept as (max, log2 sum) rather than their sum: for a fully masked row/column max = -1.44e9 and
    // adding log2(sum) to it would be lost to rounding, turning 1/(L*S) confidences into 1.0
    out[idx] = make_float2(Mt, (Mt > kNegInf) ? log2f(St) : 0.f);
    // single-pass sweep: a finished sum that is not safely above the flush-to-zero range (or is NaN) cannot be trusted
    if (needs_robust != nullptr && Mt > kNegInf && !(St >= kTinySum)) atomicOr(needs_robust, 1);
  }
}

struct Candidate {
  int row;  // n * L + i
  int col;  // j
  float conf;
  int pad;
};

// The one place conf is computed; used by both the select and the materialise path so that the value a
// match reports, the value compared against thr and the value written to conf_matrix are the same bits.
__device__ __forceinline__ float conf_exponent(float y, float2 lse_r, float2 lse_c) {
  return __fadd_rn(__fsub_rn(__fsub_rn(y, lse_r.x), lse_r.y), __fsub_rn(__fsub_rn(y, lse_c.x), lse_c.y));
}
__device__ __forceinline__ float conf_from(float y, float2 lse_r, float2 lse_c) {
  return exp2f(conf_exponent(y, lse_r, lse_c));
}

template <int BN, bool WRITE_CONF>
struct EpiSimSelect {
  struct Params {
    SimGeom g;
    const float2* lse_row;  // [N*Lp] (max, log2 sum) in log2 units
    const float2* lse_col;  // [N*Sp]
    float thr, log2_thr_lo;
    Candidate* cand;
    int* cand_count;
    int cand_cap;
    float* conf_matrix;  // [N, L, S] (WRITE_CONF only)
  };
  static constexpr int kSmemBytes = 0;
  static __device__ __forceinline__ void prefetch(const Params&, const EpiCtx&) {}
  static __device__ __forceinline__ void run(const Params& p, const EpiCtx& ctx) {
    const SimGeom& g = p.g;
    const int r = ctx.quarter * 32 + ctx.lane;
    const int grow = ctx.m_tile * BLOCK_M + r;
    const int n = grow / g.Lp;
    const int i = grow - n * g.Lp;
    const bool row_valid = i < g.L;
    const bool row_masked = row_valid && g.mask_a != nullptr && g.mask_a[static_cast<long long>(n) * g.L + i] == 0;
    const float2 lr = row_valid ? p.lse_row[grow] : make_float2(0.f, 0.f);
    const float2* lc = p.lse_col + static_cast<long long>(n) * g.Sp;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      float y[32];
      load_acc32(ctx.tmem_acc + c0, y);
      const int j0 = ctx.n_tile * BN + c0;
      if (!row_valid || j0 >= g.S) continue;
      float cf[32];
      float2 lcs[WRITE_CONF ? 1 : 32];
      bool any = false;
#pragma unroll
      for (int k = 0; k < 32; ++k) {
        const int j = j0 + k;
        const bool col_valid = j < g.S;
        const bool masked = row_masked || (g.mask_b != nullptr && col_valid && g.mask_b[static_cast<long long>(n) * g.S + j] == 0);
        const float v = sim_log2(y[k], g.scale, masked, g.mask_fill);
        const float2 lcj = col_valid ? __ldg(lc + j) : make_float2(0.f, 0.f);
        if (WRITE_CONF) {
          cf[k] = col_valid ? conf_from(v, lr, lcj) : 0.f;
          any |= cf[k] > p.thr;
        } else {
          y[k] = v;
          lcs[WRITE_CONF ? 0 : k] = lcj;
          any |= col_valid && (conf_exponent(v, lr, lcj) > p.log2_thr_lo);
        }
      }
      if (WRITE_CONF) {
        float* dst = p.conf_matrix + (static_cast<long long>(n) * g.L + i) * g.S + j0;
        if ((g.S & 7) == 0 && j0 + 32 <= g.S && (reinterpret_cast<uintptr_t>(p.conf_matrix) & 31) == 0) {
          store_f32x32_v8(cf, dst);
        } else {
#pragma unroll
          for (int k = 0; k < 32; ++k)
            if (j0 + k < g.S) dst[k] = cf[k];
        }
      }
      if (any) {
#pragma unroll 1
        for (int k = 0; k < 32; ++k) {
          const int j = j0 + k;
          if (j >= g.S) break;
          const float c = WRITE_CONF ? cf[k] : conf_from(y[k], lr, lcs[WRITE_CONF ? 0 : k]);
          if (c > p.thr) {
            const int slot = atomicAdd(p.cand_count, 1);
            if (slot < p.cand_cap) p.cand[slot] = Candidate{n * g.L + i, j, c, 0};
          }
        }
      }
    }
  }
};


// ================================================================================================
// FAST PATH: one pass over the similarity tiles (no masks, conf matrix not requested, C == 256).
//
//   * The 128-row A block (all of K = 256: 64 KB as one fp16 plane, 128 KB as bf16 hi + lo planes) stays RESIDENT in
//     shared memory while the CTA sweeps its n tiles; only B streams through a TMA ring of 64-wide k-blocks.  CTAs
//     take contiguous, balanced ranges of the m-major tile list (109 tiles each at 8 frames x 8000 x 4096).
//   * One TMA thread, one MMA thread (tcgen05.mma 128x128x16 into four TMEM accumulator stages), four epilogue
//     groups of four warps; group g owns stage g, so up to four tiles are in their epilogue while the MMA thread
//     works on the next one.
//   * Epilogue, per warp = 32 rows x 128 columns, no shared memory, no barrier with other warps (sim_fast_epilogue):
//       e_ij = 2^(y_ij - T),  y = acc * scale * log2(e),  T = one reference exponent for the warp's whole block
//       row sums    -> rowpart[n_tile][row]                 (float; registers until the tile ends)
//       column sums -> colpart[m_tile*4 + quarter][col]     (float; 7-shuffle butterfly per 32-column chunk)
//       T           -> tile_ref[m_tile][quarter][n_tile]
//     merge_fast_kernel folds the records; a finished sum that is not safely above the flush-to-zero range raises
//     *needs_robust (the caller then reruns the two-pass path, which keeps a separate maximum per row / column).
//   * Candidates are captured in the SAME pass: conf_ij > thr requires 2^(y_ij - LSErow_i) > thr, and whatever part
//     of the row the warp has already summed under-estimates the full row sum, so  e_ij > thr * (that sum)  is a
//     necessary condition.  (row, col, y) of the few survivors go to the candidate list; their exact confidence is
//     evaluated once the LSEs are final, with the same conf_from() as the two-pass path.
// ================================================================================================
//
// SCREEN variant (the default single-pass mode): the sweep over all L x S similarities runs ONE fp16 x fp16 product
// (11-bit significands, fp32 accumulation) instead of the three bf16 split products -- a third of the tensor work and
// half of the shared-memory fill.  What the sweep decides is only (a) which elements can reach thr (captured with a
// 10 % margin) and (b) the diffuse part of the row / column sums; every captured element is then re-evaluated from the
// hi + lo planes in fp32 (cand_rescore_kernel) and its term in the two log-sum-exps is REPLACED by the exact one
// (fixed-point integer atomics, so the result does not depend on the order of arrival).  The confidence of a match is
// therefore exact in its own similarity and in every other large term of its row and column; the remaining error is
// the fp16-operand error of the many small terms, which averages out (measured max 1.2e-5 against the two-pass path,
// contract 1e-3; tests/test_gpu_coarse.py).
constexpr int FAST_GROUPS = 4;
constexpr int FAST_C = 256;
constexpr int FAST_KB = FAST_C / BLOCK_K;           // 4 k-blocks
constexpr int FAST_SUB = BLOCK_M * BLOCK_K * 2;     // one [128 x 64] 16-bit sub-tile: 16 KB
template <bool SCREEN>
struct FastCfg {
  static constexpr int PLANES = SCREEN ? 1 : 2;                   // fp16 plane | bf16 hi + lo planes
  // N of one MMA.  At N = 128 a 128 x 128 x 16 MMA reads 8 KB of shared memory in its 64 clocks -- all of the 128 B/clk
  // an SM has, with the TMA fill on top; N = 256 needs 96 B/clk.  The split variant has no room for 256-wide stages.
  static constexpr int BN = 128;
  static constexpr int HALVES = BN / SIM_BN;                      // 128-column halves of an accumulator, one per epilogue group
  static constexpr int TSTAGES = 512 / BN;                        // accumulator stages in TMEM
  static constexpr int BSTAGES = 2;
  static constexpr int A_BYTES = PLANES * FAST_KB * FAST_SUB;     // the resident A block
  // one B stage: SCREEN: a whole tile [kb][BN rows x 64] = 64 KB (one wait and one commit per tile for the MMA thread;
  // four half-tile stages -- a finer-grained refill for two waits and two commits -- measured slower, 185 vs 179 us);
  // split variant: one k-block [plane][BN rows x 64]
  static constexpr int BSTAGE_BYTES = SCREEN ? FAST_KB * FAST_SUB : PLANES * HALVES * FAST_SUB;
  static constexpr int SMEM = 1024 + A_BYTES + BSTAGES * BSTAGE_BYTES + 512;  // + barriers, TMEM slot, tile slots
};
// What the screening sweep does with the bound delta on its similarity error (derived at pack_f16_kernel, below).  Capture: conf_ij > thr needs e_ij / R_i > thr; the screening sweep sees
// y_s >= y - delta and a partial row sum R_s <= 2^delta R, so testing against thr * 2^(-2 delta) (times 0.99 for the
// approximate exponentials) loses no element that can reach thr -- rigorous.  Diffuse part of the sums (everything not
// captured and therefore not re-evaluated exactly): its terms carry independent relative errors of at most
// 2^delta - 1, a match's confidence moves by at most ~0.2 of the relative error of that remainder, and the remainder
// of a row that holds a match consists of at least 1 / thr_lim terms; screening is therefore only used while
// delta <= kScreenMaxDelta (0.25 octave), beyond which the call is repeated on the three split products (algo 2).
constexpr float kScreenMaxDelta = 0.6f;
// ... and the sweep checks its own work: every captured element is evaluated twice (screening value, exact value); the
// largest difference seen is the measured counterpart of the bound, and a call whose measured error exceeds
// kScreenMaxObserved octaves is repeated on the split products as well.
constexpr float kScreenMaxObserved = 0.03f;
constexpr float kScreenMaxNorm = 60000.f;  // |x| <= |row| must stay below the largest fp16 (65504)
struct ScreenPlan {
  bool ok;
  float thr_lim;
  float delta;
};
__device__ __forceinline__ ScreenPlan screen_plan(const unsigned* __restrict__ norm_sq, float c, float thr, float max_delta) {
  const float na = sqrtf(__uint_as_float(norm_sq[0])), nb = sqrtf(__uint_as_float(norm_sq[1]));
  const float delta = c * (9.768e-4f * na * nb + 2.9802322e-8f * 16.f * (na + nb));  // 2^-10 (1 + 2^-12), 2^-25 sqrt(256)
  ScreenPlan sp;
  sp.ok = na <= kScreenMaxNorm && nb <= kScreenMaxNorm && delta <= max_delta;  // (false for NaN / inf)
  sp.thr_lim = 0.99f * thr * exp2f(-2.f * delta);
  sp.delta = delta;
  return sp;
}

constexpr double kDeltaScale = 1099511627776.0;     // 2^40: fixed-point unit of the LSE corrections

struct SimFastParams {
  CUtensorMap a_hi, a_lo, b_hi, b_lo;  // [rows, 256] bf16, box 128 x 64, 128B swizzle (SCREEN: a_hi / b_hi = the fp16 planes)
  int N, L, S, Lp, Sp;
  int n_tiles;      // ceil(Sp / BN) column tiles of the sweep
  int total_tiles;  // N * (Lp/128) * n_tiles, m-major
  float c;          // scale * log2(e)  (> 0)
  float thr_lim;    // 0.99 * thr (slack for the approximate exponentials); the screening sweep takes its own from screen_plan()
  float thr, max_delta, lim_scale;
  const float* row_lb;      // SCREEN: [N*Lp] log2 of a lower bound of each row's full sum (Jensen; pack_f16_kernel), nullable
  const unsigned* norm_sq;  // SCREEN: bits of the largest squared row norm of each side (pack_f16_kernel)
  int* needs_robust;        // SCREEN: bit 1 <- the bound does not allow screening (the caller repeats with algo 2)
  long long* trace; // debug (opb200_debug_trace): per CTA and every fourth tile {MMA issue start, MMA issued, accumulator
                    // seen full by one epilogue warp of its stage, last release among the stage's four warps} SM-clock stamps
  int pair;         // SCREEN: tiles (n even, n + 1) of a row share 128 x 256 x 16 MMAs (see the MMA issuer)
  int debug;        // OPB200_SIM_DEBUG timing experiments (results are garbage): 1 = epilogue warps only hand the accumulators back,
                    // 2 = no MMAs are issued, 3 = epilogue reads TMEM only, 4 = epilogue arithmetic without the TMEM reads,
                    // 6 = mode 1 without the B loads (MMAs on whatever the stages hold), 7 = mode 1 without the MMAs (TMA alone),
                    // 8 = one warp's tile profile instead of the stamps, 9 = accumulator stages taken strictly in turn
  int cand_tag;     // Candidate::pad of captured elements: 1 = y is final, 2 = y is the screening value (rescore)
  float* rowpart;   // [Sp/128][N*Lp]        sum_j e_ij over the 128 columns of a tile
  float* colpart;   // [4*Lp/128][N*Sp]      sum_i e_ij over the 32 rows of a warp's block
  float* tile_ref;  // [N*Lp/128][4][Sp/128] the reference T of each warp block (-inf: nothing there)
  Candidate* cand;
  int* cand_count;
  int cand_cap;
};

// (e0, e1) = 2^(c v - T) of two neighbouring columns: one packed FMA, two MUFU
__device__ __forceinline__ uint64_t exp2_pair(uint32_t v0, uint32_t v1, uint64_t c2, uint64_t t2) {
  float x0, x1;
  ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(v0), __uint_as_float(v1)), c2, t2), x0, x1);
  return ptx::pack_f32x2(ptx::ex2(x0), ptx::ex2(x1));
}

// (Measured and dropped: a share of the exponentials on the FMA pipe -- magic-number floor, cubic for 2^frac, exponent
// added to the bits: 5 packed FMA-pipe instructions + 2 FMNMX + 2 LEA per pair instead of 2 MUFU.  With 1 / 2 / 3 / 4 of
// every four pairs on that route the sweep took 188 / 195 / 206 / 218 us against 182.5: the epilogue warps are bound by
// the latency of their dependent chain at four warps per scheduler, not by the MUFU pipe or by issue slots.)
// order-preserving map of a float's bits to a signed integer (for redux.sync.max)
__device__ __forceinline__ int float_key(float v) {
  const int b = __float_as_int(v);
  return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

constexpr int kFastPipe = 2;  // epilogue schedule of sim_fast_kernel<true> (see sim_fast_epilogue)
constexpr float kRefWindow = 60.f;  // a tile keeps its reference exponent while its maximum stays within +-60 octaves of it

// Epilogue of one warp = 32 rows x 128 columns of a tile, read in the mma C-fragment layout (tmem_ld_16x256b_x4):
// thread t holds rows rsub + 8h (rsub = t/4, h = 0..3) and, per 32-column chunk, the column pairs 8k + 2(t%4) + {0,1}.
// Row sums therefore stay in registers until the tile ends (two xor-shuffles over t%4), and a column sum needs only the
// 8 threads t/4: a 7-shuffle halving butterfly per chunk leaves one finished column in every lane (the lane = row
// layout needs 31).  All adds run as packed f32x2.
//   e_ij = 2^(y_ij - T) with ONE reference T for the warp's whole 32 x 128 block, so a single ex2 serves the row and
// the column sum.  T is a guess -- the block maximum this warp saw on its previous tile -- checked after the fact: the
// raw maximum of the block is tracked alongside (FMNMX3, no dependency on T), and if it left the window
// [T - 60, T + 60] octaves the block is simply swept again from TMEM with T = its own maximum (rare: neighbouring
// tiles of the same rows have similar maxima; always on a warp's first tile).  A warp shares nothing with the other
// warps: its row sums, column sums and T go to global memory as they are and merge_fast_kernel folds them.
// Carried by a warp from tile to tile: the reference guess, and -- while the tiles belong to the same 128-row block -- a
// lower bound of each row's full sum (everything this warp has seen of the row so far), which is what the candidate
// limit is taken from: the more of a row has gone by, the fewer of its elements can still reach thr.  Lane t owns row
// t/4 + 8 (t%4) of the warp's block (its sums arrive there through a 3-shuffle transposing reduction); vlim[] is the
// limit of the thread's own four rows in raw accumulator units, as of the previous tile.
struct FastWarpState {
  float T = INFINITY;     // no guess yet: the first tile is swept twice
  int run_m = -1;         // m tile of the running sums
  float run_T = 0.f;      // owner lane: reference of run_s
  float run_s = 0.f;      // owner lane: running sum of its row
  float vlim[4] = {kNegInf, kNegInf, kNegInf, kNegInf};
  int dbg_flags = 0;      // debug trace: 1 = the tile took the second (candidate) sweep, 2 = it was re-swept for its reference
  float thr_lim = 0.f;    // candidate limit relative to the row sum
  float seed_shift = 0.f; // SCREEN: log2(2^delta * 0.99): what turns row_lb into a bound of the screened sums
  float seed_l2 = kNegInf;  // owner lane: log2 of the a-priori lower bound of its row's sum
  float T_rows = 0.f;     // SCREEN: the reference guess of the current block of rows (a function of the rows alone)
};

// lane t <- reduction over its quad (lanes t ^ 1, t ^ 2, t ^ 3) of v[t % 4]
template <typename Op>
__device__ __forceinline__ float quad_transpose_reduce(const float (&v)[4], int lane, Op op) {
  const bool b0 = (lane & 1) != 0, b1 = (lane & 2) != 0;
  const float a0 = op(b0 ? v[1] : v[0], __shfl_xor_sync(0xffffffffu, b0 ? v[0] : v[1], 1));
  const float a1 = op(b0 ? v[3] : v[2], __shfl_xor_sync(0xffffffffu, b0 ? v[2] : v[3], 1));
  return op(b1 ? a1 : a0, __shfl_xor_sync(0xffffffffu, b1 ? a0 : a1, 2));
}

// PIPE: 0 = the steps of a chunk one after the other, 1 = TMEM read of the next chunk and column butterfly of the
// previous one overlapped with the exponentials (two register sets), 2 = only the butterfly overlapped (one set)
template <bool RAGGED, int NCH, int PIPE = 0>
__device__ __forceinline__ void sim_fast_epilogue(const SimFastParams& p, uint32_t tmem_acc, int m_tile, int n, int mt_in_batch,
                                                  int jt, int quarter, int lane, FastWarpState& st,
                                                  uint64_t* __restrict__ tempty_bar, long long* trace_slot) {
  float T = st.T;
  const bool prof = p.debug == 8 && trace_slot != nullptr;  // debug: where a warp's tile time goes (SM clocks since entry)
  const long long prof_t0 = prof ? clock64() : 0;
  const int cq = (lane & 3) * 2;
  const int i0 = mt_in_batch * BLOCK_M + quarter * 32 + (lane >> 2);  // this thread's rows: i0 + 8h
  const uint64_t c2 = ptx::pack_f32x2(p.c, p.c);
  bool rv[4];
#pragma unroll
  for (int h = 0; h < 4; ++h) rv[h] = !RAGGED || i0 + 8 * h < p.L;
  float* __restrict__ colp = p.colpart + ((static_cast<long long>(mt_in_batch) * 4 + quarter) * p.N + n) * p.Sp + jt;
  if (st.run_m != m_tile) {  // a new block of rows: all that is known about their sums is the a-priori lower bound
    st.run_m = m_tile, st.run_s = 0.f, st.run_T = 0.f;
    const int own0 = lane & 3;
    const bool rv0 = !RAGGED || i0 + 8 * own0 < p.L;
    st.seed_l2 = (p.row_lb != nullptr && rv0) ? p.row_lb[static_cast<long long>(n) * p.Lp + i0 + 8 * own0] + st.seed_shift : kNegInf;
    const float v0 = st.seed_l2 > kNegInf ? __fdividef(__log2f(st.thr_lim) + st.seed_l2, p.c) : kNegInf;
#pragma unroll
    for (int h = 0; h < 4; ++h) st.vlim[h] = __shfl_sync(0xffffffffu, v0, (lane & ~3) + h);
    // Reference guess of the screening sweep: the largest MEAN similarity among the warp's 32 rows (what the Jensen
    // bound is made of) plus 8 octaves.  It depends on the rows alone -- not on which tile this warp happened to
    // process before, which the tile slots leave to chance -- so the sums, and with them every confidence, come out
    // bit-identical from run to run and for any number of GPUs; a block whose maximum lies outside T - 60 .. T + 60 is
    // swept again with T = that maximum (data-determined as well).
    if (p.row_lb != nullptr) {
      const float mean_y = st.seed_l2 > kNegInf ? st.seed_l2 - st.seed_shift - __log2f(static_cast<float>(p.S)) : kNegInf;
      const float top = key_float(__reduce_max_sync(0xffffffffu, float_key(mean_y)));
      st.T_rows = top > kNegInf ? top + 8.f : 0.f;
    }
  }
  if (p.row_lb != nullptr) T = st.T_rows;
  uint64_t racc[4];
  float rmax[4];  // raw accumulator maxima per row
  uint32_t cmask = 0;  // chunks in which this thread saw an element above the limits of the previous tiles
  float ymax = kNegInf;
  bool released = false;
  auto release = [&]() {  // the accumulator is no longer needed: hand the TMEM stage back
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(tempty_bar);
    if (trace_slot != nullptr && !prof) atomicMax(reinterpret_cast<unsigned long long*>(trace_slot), static_cast<unsigned long long>(clock64()));
    released = true;
  };
#pragma unroll 1
  for (int attempt = 0;; ++attempt) {
    const uint64_t t2 = ptx::pack_f32x2(-T, -T);
    bool redo = false;
#pragma unroll
    for (int h = 0; h < 4; ++h) racc[h] = 0ull, rmax[h] = kNegInf;
    cmask = 0;
    // the four steps of a 32-column chunk
    auto load = [&](int ch, uint32_t (&va)[16], uint32_t (&vb)[16]) {
      ptx::tmem_ld_16x256b_x4(tmem_acc + ch * 32, va);
      ptx::tmem_ld_16x256b_x4(tmem_acc + (16u << 16) + ch * 32, vb);
    };
    auto head = [&](int ch, uint32_t (&va)[16], uint32_t (&vb)[16]) {  // masks, row maxima, (last chunk) the tile's verdict
      if (RAGGED) {  // rows / columns that do not exist: -inf, i.e. e = 0 and no say in the maximum
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          const bool cv = jt + ch * 32 + 8 * (q >> 2) + cq + (q & 1) < p.S;
          if (!cv || !rv[(q >> 1) & 1]) va[q] = __float_as_uint(kNegInf);
          if (!cv || !rv[2 + ((q >> 1) & 1)]) vb[q] = __float_as_uint(kNegInf);
        }
      }
      // per-row maxima of the chunk; a chunk with an element above the limits known before this tile is remembered: only
      // those chunks can hold a candidate (the limits only rise), so only they are read again by the candidate sweep
      float cm[4] = {fmaxf(__uint_as_float(va[0]), __uint_as_float(va[1])), fmaxf(__uint_as_float(va[2]), __uint_as_float(va[3])),
                     fmaxf(__uint_as_float(vb[0]), __uint_as_float(vb[1])), fmaxf(__uint_as_float(vb[2]), __uint_as_float(vb[3]))};
#pragma unroll
      for (int k = 1; k < 4; ++k) {
        cm[0] = ptx::max3(cm[0], __uint_as_float(va[4 * k]), __uint_as_float(va[4 * k + 1]));
        cm[1] = ptx::max3(cm[1], __uint_as_float(va[4 * k + 2]), __uint_as_float(va[4 * k + 3]));
        cm[2] = ptx::max3(cm[2], __uint_as_float(vb[4 * k]), __uint_as_float(vb[4 * k + 1]));
        cm[3] = ptx::max3(cm[3], __uint_as_float(vb[4 * k + 2]), __uint_as_float(vb[4 * k + 3]));
      }
#pragma unroll
      for (int h = 0; h < 4; ++h) rmax[h] = fmaxf(rmax[h], cm[h]);
      if (cm[0] > st.vlim[0] || cm[1] > st.vlim[1] || cm[2] > st.vlim[2] || cm[3] > st.vlim[3]) cmask |= 1u << ch;
      if (ch == NCH - 1) {
        // Everything of the tile is in registers or behind us.  Its maximum decides whether T held (else: once more
        // with T = the maximum), and if no element beats even the limits known before this tile -- which can only
        // have risen since -- nothing here can become a candidate: the accumulator goes back right away.
        const float tm = fmaxf(fmaxf(rmax[0], rmax[1]), fmaxf(rmax[2], rmax[3]));
        ymax = key_float(__reduce_max_sync(0xffffffffu, float_key(tm))) * p.c;
        redo = attempt == 0 && ymax > kNegInf && !(fabsf(ymax - T) <= kRefWindow);
        const bool maybe = rmax[0] > st.vlim[0] || rmax[1] > st.vlim[1] || rmax[2] > st.vlim[2] || rmax[3] > st.vlim[3];
        if (!redo && !__any_sync(0xffffffffu, maybe)) release();
      }
    };
    auto exps = [&](const uint32_t (&va)[16], const uint32_t (&vb)[16], float (&f)[4][2]) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t e0 = exp2_pair(va[4 * k], va[4 * k + 1], c2, t2), e1 = exp2_pair(va[4 * k + 2], va[4 * k + 3], c2, t2);
        const uint64_t e2 = exp2_pair(vb[4 * k], vb[4 * k + 1], c2, t2), e3 = exp2_pair(vb[4 * k + 2], vb[4 * k + 3], c2, t2);
        racc[0] = ptx::add_f32x2(racc[0], e0);
        racc[1] = ptx::add_f32x2(racc[1], e1);
        racc[2] = ptx::add_f32x2(racc[2], e2);
        racc[3] = ptx::add_f32x2(racc[3], e3);
        ptx::unpack_f32x2(ptx::add_f32x2(ptx::add_f32x2(e0, e1), ptx::add_f32x2(e2, e3)), f[k][0], f[k][1]);
      }
    };
    auto cols = [&](int ch, const float (&f)[4][2]) {
      // halving butterfly over the row bits of the lane id (16, 8, 4): 4 + 2 + 1 shuffles
      float g[2][2], hh[2];
      {
        const bool up = (lane & 16) != 0;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            const float keep = up ? f[kk + 2][b] : f[kk][b];
            const float send = up ? f[kk][b] : f[kk + 2][b];
            g[kk][b] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
          }
      }
      {
        const bool up = (lane & 8) != 0;
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const float keep = up ? g[1][b] : g[0][b];
          const float send = up ? g[0][b] : g[1][b];
          hh[b] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
      }
      const bool up4 = (lane & 4) != 0;
      const float cs = (up4 ? hh[1] : hh[0]) + __shfl_xor_sync(0xffffffffu, up4 ? hh[0] : hh[1], 4);
      // lane bits 4,3 select k, bit 2 selects the column of the pair
      colp[ch * 32 + ((lane >> 3) & 3) * 8 + cq + ((lane >> 2) & 1)] = cs;  // (a repeated sweep overwrites it)
    };
    if constexpr (PIPE == 1) {
      // Software pipeline over two register sets: the TMEM read of chunk ch + 1 is in flight while chunk ch is
      // exponentiated, and the column butterfly of chunk ch - 1 (shuffle latency) runs between its MUFU instructions.
      // A warp alone needs ~1250 clk for a chunk when the three steps run one after the other -- latency, not
      // throughput -- and twelve to sixteen warps per SM are too few to hide that from each other.
      uint32_t v[2][2][16];
      float f[2][4][2];
      load(0, v[0][0], v[0][1]);
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        ptx::tmem_ld_wait();
        if (prof && ch == 0 && lane == 0) trace_slot[0] = clock64() - prof_t0;
        if (ch + 1 < NCH) load(ch + 1, v[(ch + 1) & 1][0], v[(ch + 1) & 1][1]);
        head(ch, v[ch & 1][0], v[ch & 1][1]);
        exps(v[ch & 1][0], v[ch & 1][1], f[ch & 1]);
        if (ch > 0) cols(ch - 1, f[(ch - 1) & 1]);
      }
      cols(NCH - 1, f[(NCH - 1) & 1]);
    } else if constexpr (PIPE == 2) {
      float f[2][4][2];
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        uint32_t va[16], vb[16];
        load(ch, va, vb);
        ptx::tmem_ld_wait();
        if (prof && ch == 0 && lane == 0) trace_slot[0] = clock64() - prof_t0;
        head(ch, va, vb);
        exps(va, vb, f[ch & 1]);
        if (ch > 0) cols(ch - 1, f[(ch - 1) & 1]);
      }
      cols(NCH - 1, f[(NCH - 1) & 1]);
    } else {
#pragma unroll 1
      for (int ch = 0; ch < NCH; ++ch) {
        uint32_t va[16], vb[16];
        if (p.debug != 4) {
          load(ch, va, vb);
          ptx::tmem_ld_wait();
          if (prof && ch == 0 && lane == 0) trace_slot[0] = clock64() - prof_t0;
        } else {
#pragma unroll
          for (int q = 0; q < 16; ++q) va[q] = vb[q] = __float_as_uint(static_cast<float>(q + ch + lane));
        }
        if (p.debug == 3) continue;
        head(ch, va, vb);
        float f[4][2];
        exps(va, vb, f);
        cols(ch, f);
      }
    }
    if (!redo) break;
    st.dbg_flags |= 2;
    T = ymax;
  }
  if (prof && lane == 0) trace_slot[1] = clock64() - prof_t0;
  const bool empty = ymax == kNegInf;  // ragged edge: no existing element in this warp's block
  const int own = lane & 3;
  float s[4];
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    float lo, hi;
    ptx::unpack_f32x2(racc[h], lo, hi);
    s[h] = lo + hi;
  }
  const float s_own = quad_transpose_reduce(s, lane, [](float x, float y) { return x + y; });
  const float mx_own = quad_transpose_reduce(rmax, lane, [](float x, float y) { return fmaxf(x, y); });
  p.rowpart[(static_cast<long long>(jt / (32 * NCH)) * p.N + n) * p.Lp + i0 + 8 * own] = s_own;  // rows that do not exist hold 0
  if (lane == 0) p.tile_ref[(static_cast<long long>(m_tile) * 4 + quarter) * (p.Sp / (32 * NCH)) + jt / (32 * NCH)] = empty ? kNegInf : T;
  // candidate capture: conf_ij > thr needs e_ij > thr * (row sum), and the part of the row this warp has seen so far
  // bounds that sum from below, i.e. acc_ij > vlim_i in raw accumulator units -- one compare per element on a second
  // read of the flagged chunks, taken only when some row of the warp has its maximum above the new limit
  bool sweep = false;
  if (!empty) {
    // (clamps keep it a finite lower bound when the reference moves far down)
    st.run_s = fminf(st.run_s * ptx::ex2(fminf(st.run_T - T, 60.f)), 1e30f) + s_own;
    st.run_T = T;
    const float lim = st.thr_lim * st.run_s;
    const bool rv_own = !RAGGED || i0 + 8 * own < p.L;
    // log2 of the limit: the part of the row seen so far or the a-priori bound, whichever is larger
    const float vlim_own = (rv_own && lim > 0.f) ? __fdividef(fmaxf(__log2f(lim) + T, __log2f(st.thr_lim) + st.seed_l2), p.c) : INFINITY;
    const bool trig = !released && mx_own > vlim_own;
#pragma unroll
    for (int h = 0; h < 4; ++h) st.vlim[h] = __shfl_sync(0xffffffffu, vlim_own, (lane & ~3) + h);
    if (prof && lane == 0) trace_slot[2] = clock64() - prof_t0;
    sweep = __any_sync(0xffffffffu, trig);
    st.T = ymax;  // the next tile's guess
  }
  if (sweep) {
    st.dbg_flags |= 1;
#pragma unroll 1
    for (uint32_t todo = __reduce_or_sync(0xffffffffu, cmask); todo != 0; todo &= todo - 1) {
      const int ch = __ffs(todo) - 1;
      uint32_t sa[16], sb[16];
      ptx::tmem_ld_16x256b_x4(tmem_acc + ch * 32, sa);
      ptx::tmem_ld_16x256b_x4(tmem_acc + (16u << 16) + ch * 32, sb);
      ptx::tmem_ld_wait();
      if ((todo & (todo - 1)) == 0) release();  // the last flagged chunk is in registers: the accumulator goes back before the walk
      // (the per-row maxima cost 16 + 4 instructions where a compare per element costs 64, and a flagged chunk need not
      // hold an element above the NEW limits)
      float cm[4] = {kNegInf, kNegInf, kNegInf, kNegInf};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        cm[0] = ptx::max3(cm[0], __uint_as_float(sa[4 * k]), __uint_as_float(sa[4 * k + 1]));
        cm[1] = ptx::max3(cm[1], __uint_as_float(sa[4 * k + 2]), __uint_as_float(sa[4 * k + 3]));
        cm[2] = ptx::max3(cm[2], __uint_as_float(sb[4 * k]), __uint_as_float(sb[4 * k + 1]));
        cm[3] = ptx::max3(cm[3], __uint_as_float(sb[4 * k + 2]), __uint_as_float(sb[4 * k + 3]));
      }
      const bool any = cm[0] > st.vlim[0] || cm[1] > st.vlim[1] || cm[2] > st.vlim[2] || cm[3] > st.vlim[3];
      if (!__any_sync(0xffffffffu, any)) continue;
      // row by row, so that the warp only walks the eight elements of rows that do hold a candidate
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        if (cm[h] > st.vlim[h]) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int b = 0; b < 2; ++b) {
              const float val = __uint_as_float(h < 2 ? sa[4 * k + 2 * h + b] : sb[4 * k + 2 * (h - 2) + b]);
              const int j = jt + ch * 32 + 8 * k + cq + b;
              if (val > st.vlim[h] && (!RAGGED || j < p.S)) {
                const int slot = atomicAdd(p.cand_count, 1);
                if (slot < p.cand_cap) p.cand[slot] = Candidate{n * p.L + i0 + 8 * h, j, __fmul_rn(val, p.c), p.cand_tag};
              }
            }
        }
      }
    }
  }
  if (!released) release();
  if (prof && lane == 0) trace_slot[3] = clock64() - prof_t0;
}

template <bool SCREEN>
__global__ void __launch_bounds__(64 + 128 * FAST_GROUPS, 1) sim_fast_kernel(const __grid_constant__ SimFastParams p) {
  constexpr int FAST_A_BYTES = FastCfg<SCREEN>::A_BYTES, FAST_BSTAGES = FastCfg<SCREEN>::BSTAGES;
  constexpr int FAST_BSTAGE_BYTES = FastCfg<SCREEN>::BSTAGE_BYTES;
  constexpr int BN = FastCfg<SCREEN>::BN, HALVES = FastCfg<SCREEN>::HALVES, TSTAGES = FastCfg<SCREEN>::TSTAGES;
  constexpr int PLANES = FastCfg<SCREEN>::PLANES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;                  // [plane 0/1][kb][128 x 64]
  uint8_t* smem_b = smem + FAST_A_BYTES;   // [stage][plane 0/1][BN x 64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + FAST_BSTAGES * FAST_BSTAGE_BYTES);
  uint64_t* a_full = bars;                 // 1
  uint64_t* a_empty = bars + 1;            // 1
  uint64_t* b_full = bars + 2;             // FAST_BSTAGES
  uint64_t* b_empty = b_full + FAST_BSTAGES;
  uint64_t* tfull = b_empty + FAST_BSTAGES;  // TSTAGES
  uint64_t* tempty = tfull + TSTAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + TSTAGES);
  uint64_t* pfull = bars + 16;   // SCREEN, pair mode: one barrier pair per k-block slot of the B region
  uint64_t* pempty = bars + 20;
  // SCREEN: which tile an accumulator stage holds -- (m_tile, n_tile, position in the CTA's list, -) written by the MMA
  // thread before it issues the tile; m_tile < 0 tells the stage's epilogue group that the list is exhausted
  volatile int4* tile_slot = reinterpret_cast<volatile int4*>(reinterpret_cast<uint8_t*>(bars) + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float thr_lim = p.thr_lim, seed_shift = 0.f;
  if (SCREEN) {  // every CTA derives the same plan from the operand norms; no screening -> nothing to do here
    const ScreenPlan plan = screen_plan(p.norm_sq, p.c, p.thr, p.max_delta);
    if (!plan.ok) {
      if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(p.needs_robust, 2);
      return;
    }
    thr_lim = plan.thr_lim * p.lim_scale;
    // e_s > thr 2^-delta R and R >= R_jensen: in units of thr_lim = 0.99 thr 2^(-2 delta) the seed is 2^delta R_jensen
    // (the 0.99 of thr_lim also covers the rounding of the bound itself)
    seed_shift = plan.delta;
  }
  // contiguous, balanced range of the m-major tile list: at most two A blocks per CTA at these sizes
  const int t_begin = static_cast<int>(static_cast<long long>(p.total_tiles) * blockIdx.x / gridDim.x);
  const int t_end = static_cast<int>(static_cast<long long>(p.total_tiles) * (blockIdx.x + 1) / gridDim.x);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_hi);
    ptx::prefetch_tmap(&p.b_hi);
    if (!SCREEN) {
      ptx::prefetch_tmap(&p.a_lo);
      ptx::prefetch_tmap(&p.b_lo);
    }
    ptx::mbar_init(a_full, 1);
    ptx::mbar_init(a_empty, 1);
    for (int s = 0; s < FAST_BSTAGES; ++s) {
      ptx::mbar_init(&b_full[s], 1);
      ptx::mbar_init(&b_empty[s], 1);
    }
    for (int g = 0; g < TSTAGES; ++g) {
      ptx::mbar_init(&tfull[g], 1);
      ptx::mbar_init(&tempty[g], 4 * HALVES);
    }
    if (SCREEN) {
      for (int kb = 0; kb < 4; ++kb) {
        ptx::mbar_init(&pfull[kb], 1);
        ptx::mbar_init(&pempty[kb], 1);
      }
    }
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
  const int tiles_per_batch = p.Lp / BLOCK_M;

  // Every role walks the same tile list t_begin .. t_end with an incrementally kept (m_tile, n_tile): the two
  // single-thread loops below (TMA, MMA issue) are what paces the kernel once the epilogue keeps up, so they carry no
  // divisions, no per-step election and no descriptor construction.
  const int m_first = t_begin / p.n_tiles, n_first = t_begin - m_first * p.n_tiles;
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0, a_phase = 0;
      int m_tile = m_first, n_tile = n_first;
      int b_row0 = (m_tile / tiles_per_batch) * p.Sp, mt_in_batch = m_tile % tiles_per_batch;
      bool new_a = true;
      if (SCREEN && p.pair) {
        // pair mode: the B region is four k-block slots of [256 rows x 128 B]: rows 0..127 = tile n, 128..255 = tile n + 1
        uint32_t unit = 0;
        for (int t = t_begin; t < t_end; ++unit) {
          if (new_a) {
            ptx::mbar_wait(a_empty, a_phase ^ 1);
            ptx::mbar_arrive_expect_tx(a_full, FAST_A_BYTES);
#pragma unroll
            for (int kb = 0; kb < FAST_KB; ++kb) ptx::tma_load_2d(smem_a + kb * FAST_SUB, &p.a_hi, a_full, kb * BLOCK_K, m_tile * BLOCK_M);
            a_phase ^= 1;
            new_a = false;
          }
          const bool pr = (n_tile & 1) == 0 && t + 1 < t_end;
          const int b_row = b_row0 + n_tile * SIM_BN;
#pragma unroll
          for (int kb = 0; kb < FAST_KB; ++kb) {
            ptx::mbar_wait(&pempty[kb], (unit & 1u) ^ 1u);
            ptx::mbar_arrive_expect_tx(&pfull[kb], pr ? 2 * FAST_SUB : FAST_SUB);
            ptx::tma_load_2d(smem_b + kb * (2 * FAST_SUB), &p.b_hi, &pfull[kb], kb * BLOCK_K, b_row);
            if (pr) ptx::tma_load_2d(smem_b + kb * (2 * FAST_SUB) + FAST_SUB, &p.b_hi, &pfull[kb], kb * BLOCK_K, b_row + SIM_BN);
          }
          const int adv = pr ? 2 : 1;
          t += adv;
          n_tile += adv;
          if (n_tile == p.n_tiles) {
            n_tile = 0, ++m_tile, new_a = true;
            if (++mt_in_batch == tiles_per_batch) mt_in_batch = 0, b_row0 += p.Sp;
          }
        }
      } else
      for (int t = t_begin; t < t_end; ++t) {
        if (new_a) {
          ptx::mbar_wait(a_empty, a_phase ^ 1);
          ptx::mbar_arrive_expect_tx(a_full, FAST_A_BYTES);
#pragma unroll
          for (int kb = 0; kb < FAST_KB; ++kb) {
            ptx::tma_load_2d(smem_a + kb * FAST_SUB, &p.a_hi, a_full, kb * BLOCK_K, m_tile * BLOCK_M);
            if (!SCREEN) ptx::tma_load_2d(smem_a + (FAST_KB + kb) * FAST_SUB, &p.a_lo, a_full, kb * BLOCK_K, m_tile * BLOCK_M);
          }
          a_phase ^= 1;
          new_a = false;
        }
        const int b_row = b_row0 + n_tile * BN;
        if constexpr (SCREEN) {  // one stage = the tile's four k-blocks
          ptx::mbar_wait(&b_empty[stage], phase ^ 1);
          if (p.debug == 6) {  // experiment: the B stages are declared full without loading anything
            ptx::mbar_arrive(&b_full[stage]);
          } else {
            ptx::mbar_arrive_expect_tx(&b_full[stage], FAST_BSTAGE_BYTES);
#pragma unroll
            for (int kb = 0; kb < FAST_KB; ++kb)
              ptx::tma_load_2d(smem_b + stage * FAST_BSTAGE_BYTES + kb * FAST_SUB, &p.b_hi, &b_full[stage], kb * BLOCK_K, b_row);
          }
          if (++stage == FAST_BSTAGES) {
            stage = 0;
            phase ^= 1;
          }
        } else {
#pragma unroll
        for (int kb = 0; kb < FAST_KB; ++kb) {
          ptx::mbar_wait(&b_empty[stage], phase ^ 1);
          if (p.debug == 6) {  // experiment: the B stages are declared full without loading anything
            ptx::mbar_arrive(&b_full[stage]);
          } else {
          ptx::mbar_arrive_expect_tx(&b_full[stage], FAST_BSTAGE_BYTES);
          uint8_t* sb = smem_b + stage * FAST_BSTAGE_BYTES;
          // 128-row boxes; rows past the tensor (odd half of the last tile) arrive as zeros and are never read back
#pragma unroll
          for (int hf = 0; hf < HALVES; ++hf) {
            ptx::tma_load_2d(sb + hf * FAST_SUB, &p.b_hi, &b_full[stage], kb * BLOCK_K, b_row + hf * SIM_BN);
            if (!SCREEN) ptx::tma_load_2d(sb + (HALVES + hf) * FAST_SUB, &p.b_lo, &b_full[stage], kb * BLOCK_K, b_row + hf * SIM_BN);
          }
          }
          if (++stage == FAST_BSTAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        }
        if (++n_tile == p.n_tiles) {
          n_tile = 0, ++m_tile, new_a = true;
          if (++mt_in_batch == tiles_per_batch) mt_in_batch = 0, b_row0 += p.Sp;
        }
      }
    }
  } else if (warp == 1 && SCREEN) {
    // ------------------------------------------------------------------ MMA issuer of the screening sweep (one thread)
    // * B ring: the list is walked two tiles at a time so that the B stage ((it % 2) * 4 + kb of the 8-stage ring) and
    //   the parity of its full barrier ((it / 2) % 2) are compile-time constants: a k-block is one wait, four MMAs
    //   whose descriptors are base + constant, one commit (tools/probes/mma_pipe.cu: this loop alone issues a tile in
    //   1080 clk against 1024 clk of tensor work; the rolled loop with run-time stage / phase arithmetic needed 1700, a
    //   whole-warp loop with one issuing lane 1530).
    // * Accumulators: a tile goes to WHICHEVER of the four stages is free (probed round-robin), not to stage it % 4: an
    //   epilogue group that is late (a second sweep for candidates, a re-swept reference) then only skips a turn
    //   instead of holding up the three tiles queued behind it.  The stage's tile slot tells its group what it got.
    //   (Measured and dropped: two groups of eight warps alternating between two stages each, so that a group never
    //   waits for a refill -- 190 us against 181: the two warps of a group that share a scheduler move in step and
    //   their MUFU bursts collide, 1300 clk per 32 x 32 chunk against 950 with four groups on four different tiles.)
    if constexpr (SCREEN) {
      static_assert(FAST_BSTAGES == 2 && TSTAGES == 4, "the unrolled schedule assumes two B stages and 4 accumulators");
      if (ptx::elect_one()) {
        constexpr uint32_t idesc = ptx::make_idesc_f16(BLOCK_M, BN);
        constexpr int kDescPerStage = FAST_BSTAGE_BYTES >> 4;
        const uint64_t a_desc0 = ptx::make_desc_sw128(ptx::smem_u32(smem_a));
        const uint64_t b_desc0 = ptx::make_desc_sw128(ptx::smem_u32(smem_b));
        const uint32_t slot_a = ptx::smem_u32(const_cast<int4*>(tile_slot));
        const bool trace_on = p.debug != 5 && p.debug != 8 && p.trace != nullptr;
        const int n_it = t_end - t_begin;
        uint32_t a_phase = 0;
        uint32_t te_par = 0xfu;  // per stage: the parity of its empty barrier that means "free" (a fresh barrier passes parity 1)
        int g = 3;               // last stage taken
        int m_tile = m_first, n_tile = n_first;
        bool new_a = true;
        auto take_stage = [&]() {
          if (p.debug == 9) {  // experiment: stages strictly in turn (what the kernel did before the tile slots)
            const int c = (g + 1) & (TSTAGES - 1);
            ptx::mbar_wait(&tempty[c], (te_par >> c) & 1u);
            g = c;
            te_par ^= 1u << c;
            return;
          }
          uint32_t spins = 0;
          for (;;) {
#pragma unroll
            for (int q = 1; q <= TSTAGES; ++q) {
              const int c = (g + q) & (TSTAGES - 1);
              if (ptx::mbar_test_wait(&tempty[c], (te_par >> c) & 1u)) {
                g = c;
                te_par ^= 1u << c;
                return;
              }
            }
            if (++spins > (1u << 24)) __trap();
            __nanosleep(32);  // every stage is with its epilogue group: the operand reads of the MMAs in flight want the shared-memory port
          }
        };
        if (p.pair) {
          // Pair mode: tiles (n even, n + 1) of a row go to an adjacent pair of accumulator stages (0, 1) or (2, 3) with
          // 128 x 256 x 16 MMAs -- 4 KB of A + 8 KB of B per 128 clk of tensor work (96 B/clk) where two 128-wide MMAs read
          // 8 KB each per 64 clk (128 B/clk, all an SM's shared memory delivers while the TMA refills the ring through the
          // same port).  A tile without a partner (odd start or end of the CTA's list) takes any free stage with 128-wide MMAs
          // on the first half of the slots.
          constexpr uint32_t idesc2 = ptx::make_idesc_f16(BLOCK_M, 2 * BN);
          uint32_t unit = 0;
          int lastp = 1;
          for (int it = 0; it < n_it; ++unit) {
            if (new_a) {
              ptx::mbar_wait(a_full, a_phase);
              a_phase ^= 1;
              new_a = false;
            }
            const bool pr = (n_tile & 1) == 0 && it + 1 < n_it;
            if (pr) {
              uint32_t spins = 0;
              for (;;) {
                bool got = false;
#pragma unroll
                for (int q = 1; q <= 2; ++q) {
                  const int pp = (lastp + q) & 1;
                  if (ptx::mbar_test_wait(&tempty[2 * pp], (te_par >> (2 * pp)) & 1u) &&
                      ptx::mbar_test_wait(&tempty[2 * pp + 1], (te_par >> (2 * pp + 1)) & 1u)) {
                    lastp = pp;
                    te_par ^= 3u << (2 * pp);
                    got = true;
                    break;
                  }
                }
                if (got) break;
                if (++spins > (1u << 24)) __trap();
                __nanosleep(32);
              }
              g = 2 * lastp + 1;  // (the round-robin of the single tiles continues behind the pair)
            } else {
              take_stage();
            }
            ptx::tc_fence_after();
            const int g0 = pr ? 2 * lastp : g;
            ptx::st_shared_v4(slot_a + 16 * g0, m_tile, n_tile, it, 0);
            if (pr) ptx::st_shared_v4(slot_a + 16 * (g0 + 1), m_tile, n_tile + 1, it + 1, 0);
            const bool tr = trace_on && (it & 3) == 0 && (it >> 2) < kTraceTilesPerCta;
            if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 0] = clock64();
            const uint32_t d_tmem = tmem_base + g0 * BN;
#pragma unroll
            for (int kb = 0; kb < FAST_KB; ++kb) {
              ptx::mbar_wait(&pfull[kb], unit & 1u);
              ptx::tc_fence_after();
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k)
                ptx::mma_bf16(d_tmem, a_desc0 + static_cast<uint64_t>(kb * (FAST_SUB >> 4) + 2 * k),
                              b_desc0 + static_cast<uint64_t>(kb * ((2 * FAST_SUB) >> 4) + 2 * k), pr ? idesc2 : idesc, (kb | k) != 0);
              ptx::mma_commit(&pempty[kb]);
            }
            ptx::mma_commit(&tfull[g0]);
            if (pr) ptx::mma_commit(&tfull[g0 + 1]);
            if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 1] = clock64();
            const int adv = pr ? 2 : 1;
            it += adv;
            n_tile += adv;
            const bool last_n = n_tile == p.n_tiles;
            if (last_n || it >= n_it) ptx::mma_commit(a_empty);
            if (last_n) n_tile = 0, ++m_tile, new_a = true;
          }
        } else
        for (int it0 = 0; it0 < n_it; it0 += 2) {
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int it = it0 + u;
            if (it >= n_it) break;
            if (new_a) {
              ptx::mbar_wait(a_full, a_phase);
              a_phase ^= 1;
              new_a = false;
            }
            take_stage();
            ptx::tc_fence_after();
            // (plain store: the group reads the slot after the stage's full barrier, which the commit below arrives on
            // once the MMAs issued after this store have completed)
            ptx::st_shared_v4(slot_a + 16 * g, m_tile, n_tile, it, 0);
            const bool tr = trace_on && (it & 3) == 0 && (it >> 2) < kTraceTilesPerCta;
            if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 0] = clock64();
            const uint32_t d_tmem = tmem_base + g * BN;
            ptx::mbar_wait(&b_full[u], static_cast<uint32_t>((it0 >> 1) & 1));
            ptx::tc_fence_after();
#pragma unroll
            for (int kb = 0; kb < FAST_KB; ++kb)
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k)
                ptx::mma_bf16(d_tmem, a_desc0 + static_cast<uint64_t>(kb * (FAST_SUB >> 4) + 2 * k),
                              b_desc0 + static_cast<uint64_t>(u * kDescPerStage + kb * (FAST_SUB >> 4) + 2 * k), idesc, (kb | k) != 0);
            ptx::mma_commit(&b_empty[u]);
            ptx::mma_commit(&tfull[g]);
            if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 1] = clock64();
            const bool last_n = ++n_tile == p.n_tiles;
            if (last_n || it + 1 == n_it) ptx::mma_commit(a_empty);
            if (last_n) n_tile = 0, ++m_tile, new_a = true;
          }
        }
        // the list is exhausted: every group gets the end marker once its stage is free again
        for (int q = 0; q < TSTAGES; ++q) {
          ptx::mbar_wait(&tempty[q], (te_par >> q) & 1u);
          ptx::st_shared_v4(slot_a + 16 * q, -1, 0, 0, 0);
          __threadfence_block();
          ptx::mbar_arrive(&tfull[q]);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = SCREEN ? ptx::make_idesc_f16(BLOCK_M, BN) : ptx::make_idesc_bf16(BLOCK_M, BN);
      uint64_t a_desc[PLANES][FAST_KB];
#pragma unroll
      for (int pl = 0; pl < PLANES; ++pl)
#pragma unroll
        for (int kb = 0; kb < FAST_KB; ++kb) a_desc[pl][kb] = ptx::make_desc_sw128(ptx::smem_u32(smem_a + (pl * FAST_KB + kb) * FAST_SUB));
      const uint64_t b_desc0 = ptx::make_desc_sw128(ptx::smem_u32(smem_b));  // + (byte offset >> 4): the start-address field
      int stage = 0;
      uint32_t phase = 0, a_phase = 0;
      int n_tile = n_first;
      bool new_a = true;
      for (int t = t_begin, it = 0; t < t_end; ++t, ++it) {
        if (new_a) {
          ptx::mbar_wait(a_full, a_phase);
          a_phase ^= 1;
          new_a = false;
        }
        const int acc = it % TSTAGES;
        ptx::mbar_wait(&tempty[acc], ((it / TSTAGES) & 1) ^ 1);
        ptx::tc_fence_after();
        const bool tr = p.debug != 5 && p.trace != nullptr && (it & 3) == 0 && (it >> 2) < kTraceTilesPerCta;
        if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 0] = clock64();
        const uint32_t d_tmem = tmem_base + acc * BN;
#pragma unroll
        for (int kb = 0; kb < FAST_KB; ++kb) {
          ptx::mbar_wait(&b_full[stage], phase);
          const uint64_t bh = b_desc0 + static_cast<uint64_t>((stage * FAST_BSTAGE_BYTES) >> 4);
          if (p.debug == 2) {
          } else if (SCREEN) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) ptx::mma_bf16(d_tmem, a_desc[0][kb] + 2 * k, bh + 2 * k, idesc, (kb | k) != 0);
          } else {
            const uint64_t bl = bh + static_cast<uint64_t>((HALVES * FAST_SUB) >> 4);
            // small cross terms first, the dominant hi*hi term last
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) ptx::mma_bf16(d_tmem, a_desc[0][kb] + 2 * k, bl + 2 * k, idesc, (kb | k) != 0);
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) ptx::mma_bf16(d_tmem, a_desc[PLANES - 1][kb] + 2 * k, bh + 2 * k, idesc, 1);
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) ptx::mma_bf16(d_tmem, a_desc[0][kb] + 2 * k, bh + 2 * k, idesc, 1);
          }
          ptx::mma_commit(&b_empty[stage]);
          if (++stage == FAST_BSTAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        ptx::mma_commit(&tfull[acc]);
        if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 1] = clock64();
        const bool last_n = ++n_tile == p.n_tiles;
        if (last_n || t + 1 == t_end) ptx::mma_commit(a_empty);
        if (last_n) n_tile = 0, new_a = true;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue groups
    const int ew = warp - 2;
    const int quarter = warp & 3;
    FastWarpState st;
    st.thr_lim = thr_lim;
    st.seed_shift = seed_shift;
    if constexpr (SCREEN) {
      // group g = the four warps that own accumulator stage g; the stage's tile slot says which tile the MMA thread put
      // there (tiles reach a group in list order, but not every fourth one)
      const int tstage = ew >> 2;
      int m_tile = -1, n_batch = 0, mt_in_batch = 0;
      for (int cnt = 0;; ++cnt) {
        ptx::mbar_wait(&tfull[tstage], cnt & 1);
        ptx::tc_fence_after();
        const int4 slot = ptx::ld_shared_v4(ptx::smem_u32(const_cast<int4*>(tile_slot)) + 16 * tstage);
        const int mt = slot.x;
        if (mt < 0) break;
        const int n_tile = slot.y, it_e = slot.z;
        if (mt != m_tile) m_tile = mt, n_batch = mt / tiles_per_batch, mt_in_batch = mt - n_batch * tiles_per_batch;
        const bool tr = p.debug != 5 && p.trace != nullptr && (it_e & 3) == 0 && (it_e >> 2) < kTraceTilesPerCta &&
                        (p.debug != 8 || quarter == 0);  // all four warps: first / last (mode 8: one warp's profile)
        if (tr && quarter == 2 && lane == 0 && p.debug != 8) p.trace[(blockIdx.x * kTraceTilesPerCta + (it_e >> 2)) * 4 + 2] = clock64();
        const int jt = n_tile * BN;
        if (p.debug != 1 && (p.debug < 6 || p.debug >= 8)) {
          const uint32_t tacc = tmem_base + tstage * BN + (static_cast<uint32_t>(quarter * 32) << 16);
          const bool ragged = ((mt_in_batch + 1) * BLOCK_M > p.L) || (jt + SIM_BN > p.S);
          const bool tr5 = p.debug == 5 && p.trace != nullptr && (it_e & 3) == 0 && (it_e >> 2) < kTraceTilesPerCta;
          const long long t5 = tr5 ? clock64() : 0;
          st.dbg_flags = 0;
          long long* tslot = tr ? &p.trace[(blockIdx.x * kTraceTilesPerCta + (it_e >> 2)) * 4 + (p.debug == 8 ? 0 : 3)] : nullptr;
          if (ragged) sim_fast_epilogue<true, 4>(p, tacc, m_tile, n_batch, mt_in_batch, jt, quarter, lane, st, &tempty[tstage], tslot);
          else sim_fast_epilogue<false, 4, kFastPipe>(p, tacc, m_tile, n_batch, mt_in_batch, jt, quarter, lane, st, &tempty[tstage], tslot);
          if (tr5 && lane == 0)  // mode 5: the slot holds, per quarter, this warp's epilogue time and flags instead of the four stamps
            p.trace[(blockIdx.x * kTraceTilesPerCta + (it_e >> 2)) * 4 + quarter] = (clock64() - t5) | (static_cast<long long>(st.dbg_flags) << 48);
        } else {
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tempty[tstage]);
        }
      }
    } else {
      const int group = ew >> 2;                          // owns half (group % HALVES) of accumulator stage (group / HALVES)
      const int tstage = group / HALVES, half = group % HALVES;
      int m_tile = m_first, n_tile = n_first + tstage;
      int n_batch = m_tile / tiles_per_batch, mt_in_batch = m_tile % tiles_per_batch;
      for (int t = t_begin + tstage, round = 0; t < t_end; t += TSTAGES, n_tile += TSTAGES, ++round) {
        while (n_tile >= p.n_tiles) {
          n_tile -= p.n_tiles, ++m_tile;
          if (++mt_in_batch == tiles_per_batch) mt_in_batch = 0, ++n_batch;
        }
        ptx::mbar_wait(&tfull[tstage], round & 1);
        ptx::tc_fence_after();
        const int it_e = t - t_begin;
        const bool tr = p.debug != 5 && p.trace != nullptr && (it_e & 3) == 0 && (it_e >> 2) < kTraceTilesPerCta && half == 0 && lane == 0;  // all four warps: first / last
        if (tr && quarter == 2) p.trace[(blockIdx.x * kTraceTilesPerCta + (it_e >> 2)) * 4 + 2] = clock64();
        const int jt = n_tile * BN + half * SIM_BN;
        if (jt < p.Sp && p.debug != 1 && p.debug < 6) {  // (the odd half of the last tile lies outside the padded width)
          const uint32_t tacc = tmem_base + tstage * BN + half * SIM_BN + (static_cast<uint32_t>(quarter * 32) << 16);
          const bool ragged = ((mt_in_batch + 1) * BLOCK_M > p.L) || (jt + SIM_BN > p.S);
          st.dbg_flags = 0;
          long long* tslot = tr ? &p.trace[(blockIdx.x * kTraceTilesPerCta + (it_e >> 2)) * 4 + 3] : nullptr;
          if (ragged) sim_fast_epilogue<true, 4>(p, tacc, m_tile, n_batch, mt_in_batch, jt, quarter, lane, st, &tempty[tstage], tslot);
          else sim_fast_epilogue<false, 4>(p, tacc, m_tile, n_batch, mt_in_batch, jt, quarter, lane, st, &tempty[tstage], tslot);
        } else {
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tempty[tstage]);
        }
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

// (Measured and dropped once more in round 2, with the Jensen seed in place and only the 8000 genuine candidates left to
// capture: taking the candidates from the registers of the main pass, against the limits known before the tile, so that
// no accumulator is ever read twice -- 256 us against 181, same box: the walk inside the chunk loop costs the hot path
// its registers (96 bytes of spills) although it runs for about one chunk in thirty.)
// (Measured and dropped -- "mailbox" variant: THREE epilogue groups on the four accumulator stages, tiles addressed to a
// group through a two-deep mailbox so that the fourth stage refills while three are in work, 512 threads = 128 registers,
// the epilogue fully software-pipelined over two register sets (TMEM read of chunk ch + 1 in flight, butterfly of chunk
// ch - 1 between the MUFU instructions).  A warp's tile fell from ~5000 to ~3000 clk, the kernel went from 184 to
// 217-233 us: with epilogue warps that issue densely the single MMA-issuing thread gets its instructions in later
// (1600-1800 clk per tile instead of 1200) and becomes the limit.)

// Partial sums of the single-pass sweep -> (reference, log2 sum) per row and per column.  The reference of a part is
// that of the warp block it came from (tile_ref) and is the same for the 32 consecutive rows / columns a warp owns, so a
// warp first takes the maximum M of its references (lane k holds T_k), turns them into weights 2^(T_k - M) once, and
// then adds the parts with independent FMAs (no rescaling chain; loads coalesced along the output index, many in
// flight).  Rows: one warp per 32 rows over all n tiles.  Columns: 8 warps per 32 columns, each over a slice of the
// (m tile, quarter) list, combined through shared memory in a fixed order (deterministic).  A finished sum of an
// existing row / column that is not safely above the flush-to-zero range (or is NaN) raises bit 0 of *needs_robust.
// CV = columns per lane of the column blocks: 4 (16-byte loads, 128 columns per block) when that still fills the GPU,
// else 1.  (Round 2: the 32-frame launch ran at 1.7 TB/s with 4-byte loads and accumulators indexed by a run-time
// counter, which the compiler kept in local memory.)
template <int CV>
__global__ void __launch_bounds__(256) merge_fast_kernel(const float* __restrict__ rowpart, const float* __restrict__ colpart,
                                                         const float* __restrict__ tile_ref, int N, int L, int S, int Lp, int Sp,
                                                         float2* __restrict__ lse_row, float2* __restrict__ lse_col,
                                                         int* __restrict__ needs_robust, int part_w /* columns per row part: 128 | 64 */) {
  __shared__ float2 sm[8][32 * CV + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long n_rows = static_cast<long long>(N) * Lp;
  const int row_blocks = static_cast<int>(n_rows / 256 + ((n_rows % 256) != 0));  // Lp is a multiple of 128
  const int mtb = Lp / BLOCK_M, ntb = Sp / part_w;
  if (static_cast<int>(blockIdx.x) < row_blocks) {
    const long long idx0 = (static_cast<long long>(blockIdx.x) * 8 + warp) * 32;  // this warp's 32 rows
    if (idx0 >= n_rows) return;
    const long long idx = idx0 + lane;
    const int n = static_cast<int>(idx0 / Lp), i0 = static_cast<int>(idx0 - static_cast<long long>(n) * Lp);
    const float* tr = tile_ref + ((static_cast<long long>(n) * mtb + i0 / BLOCK_M) * 4 + (i0 % BLOCK_M) / 32) * ntb;
    float M = kNegInf;
    for (int k0 = 0; k0 < ntb; k0 += 32) {
      const float T = k0 + lane < ntb ? tr[k0 + lane] : kNegInf;
      M = fmaxf(M, key_float(__reduce_max_sync(0xffffffffu, float_key(T))));
    }
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (M > kNegInf) {
      for (int k0 = 0; k0 < ntb; k0 += 32) {
        const float T = k0 + lane < ntb ? tr[k0 + lane] : kNegInf;
        const float w = T > kNegInf ? exp2f(T - M) : 0.f;
        const int kn = min(32, ntb - k0);
        const float* src = rowpart + static_cast<long long>(k0) * n_rows + idx;
        int k = 0;
        for (; k + 8 <= kn; k += 8) {
          float v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = __ldcs(src + static_cast<long long>(k + u) * n_rows);
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const float wk = __shfl_sync(0xffffffffu, w, k + u);
            acc[u] = fmaf(wk > 0.f ? v[u] : 0.f, wk, acc[u]);  // (a part without a reference holds nothing valid)
          }
        }
        for (; k < kn; ++k) {
          const float wk = __shfl_sync(0xffffffffu, w, k);
          const float v = __ldcs(src + static_cast<long long>(k) * n_rows);
          acc[0] = fmaf(wk > 0.f ? v : 0.f, wk, acc[0]);
        }
      }
    }
    const float St = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
    const bool ok = (i0 + lane) < L && M > kNegInf;
    lse_row[idx] = make_float2(ok ? M : kNegInf, ok ? log2f(St) : 0.f);
    if (ok && !(St >= kTinySum)) atomicOr(needs_robust, 1);
    return;
  }
  // ---- columns: 32 * CV per block (a lane owns CV consecutive ones), 8 slices of the (m tile, quarter) list; the
  // block's columns lie in one tile of part_w columns, or in two (part_w = 64, CV = 4: lanes 16.. are in the second)
  const long long idx0 = (static_cast<long long>(blockIdx.x) - row_blocks) * (32 * CV);
  const int n = static_cast<int>(idx0 / Sp), j0 = static_cast<int>(idx0 - static_cast<long long>(n) * Sp);
  const int jl = j0 + CV * lane;
  const int tile0 = j0 / part_w;
  const bool second = jl / part_w != tile0;
  const bool two = (j0 + 32 * CV - 1) / part_w != tile0;
  const int nparts = 4 * mtb;
  const int per = (nparts + 7) / 8;
  const int kb = warp * per, ke = min(nparts, kb + per);
  const float* tr = tile_ref + static_cast<long long>(n) * nparts * ntb + tile0;
  float M0 = kNegInf, M1 = kNegInf;
  for (int k0 = kb; k0 < ke; k0 += 32) {
    const bool in = k0 + lane < ke;
    const float T0 = in ? tr[static_cast<long long>(k0 + lane) * ntb] : kNegInf;
    const float T1 = (in && two) ? tr[static_cast<long long>(k0 + lane) * ntb + 1] : kNegInf;
    M0 = fmaxf(M0, key_float(__reduce_max_sync(0xffffffffu, float_key(T0))));
    M1 = fmaxf(M1, key_float(__reduce_max_sync(0xffffffffu, float_key(T1))));
  }
  const float M = second ? M1 : M0;
  float acc[2][CV];
#pragma unroll
  for (int c = 0; c < CV; ++c) acc[0][c] = acc[1][c] = 0.f;
  if (M0 > kNegInf || M1 > kNegInf) {  // warp-uniform
    for (int k0 = kb; k0 < ke; k0 += 32) {
      const bool in = k0 + lane < ke;
      const float T0 = in ? tr[static_cast<long long>(k0 + lane) * ntb] : kNegInf;
      const float T1 = (in && two) ? tr[static_cast<long long>(k0 + lane) * ntb + 1] : kNegInf;
      const float w0 = T0 > kNegInf ? exp2f(T0 - M0) : 0.f;
      const float w1 = T1 > kNegInf ? exp2f(T1 - M1) : 0.f;
      const int kn = min(32, ke - k0);
      const float* src = colpart + (static_cast<long long>(k0) * N + n) * Sp + jl;
      const long long pstride = static_cast<long long>(N) * Sp;
      int k = 0;
      for (; k + 8 <= kn; k += 8) {
        float v[8][CV];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if constexpr (CV == 4) {
            const float4 t = __ldcs(reinterpret_cast<const float4*>(src + (k + u) * pstride));
            v[u][0] = t.x, v[u][1] = t.y, v[u][2] = t.z, v[u][3] = t.w;
          } else {
            v[u][0] = __ldcs(src + (k + u) * pstride);
          }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float wa = __shfl_sync(0xffffffffu, w0, k + u), wb = __shfl_sync(0xffffffffu, w1, k + u);
          const float wk = second ? wb : wa;
#pragma unroll
          for (int c = 0; c < CV; ++c) acc[u & 1][c] = fmaf(wk > 0.f ? v[u][c] : 0.f, wk, acc[u & 1][c]);
        }
      }
      for (; k < kn; ++k) {
        const float wa = __shfl_sync(0xffffffffu, w0, k), wb = __shfl_sync(0xffffffffu, w1, k);
        const float wk = second ? wb : wa;
        float v[CV];
        if constexpr (CV == 4) {
          const float4 t = __ldcs(reinterpret_cast<const float4*>(src + k * pstride));
          v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
        } else {
          v[0] = __ldcs(src + k * pstride);
        }
#pragma unroll
        for (int c = 0; c < CV; ++c) acc[0][c] = fmaf(wk > 0.f ? v[c] : 0.f, wk, acc[0][c]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CV; ++c) sm[warp][CV * lane + c] = make_float2(M, acc[0][c] + acc[1][c]);
  __syncthreads();
  if (threadIdx.x < 32 * CV) {
    const int c = threadIdx.x;
    float Mt = kNegInf;
#pragma unroll
    for (int q = 0; q < 8; ++q) Mt = fmaxf(Mt, sm[q][c].x);
    float St = 0.f;
    if (Mt > kNegInf) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float2 pq = sm[q][c];
        if (pq.x > kNegInf) St += pq.y * exp2f(pq.x - Mt);
      }
    }
    const bool ok = (j0 + c) < S && Mt > kNegInf;
    lse_col[idx0 + c] = make_float2(ok ? Mt : kNegInf, ok ? log2f(St) : 0.f);
    if (ok && !(St >= kTinySum)) atomicOr(needs_robust, 1);
  }
}

// ================================================================================================
// SWEEP KERNEL (algo 3): the same single pass as sim_fast_kernel<true>, re-cut so that the tensor pipe and the epilogue
// no longer take turns on an accumulator stage.
//
//   * Tiles are 128 x 128 (MMAs 128 x 128 x 16) in four accumulator stages; the MMA thread runs the statically
//     scheduled loop of sim_fast_kernel (8-stage B ring, stage / parity compile-time constants).
//   * ALL sixteen epilogue warps work on every tile: warp (quarter q, column block c) owns rows 32q .. 32q+31 and
//     columns 32c .. 32c+31.  It reads its 32 x 32 chunk into registers and arrives on the stage's barrier as soon as
//     the read has landed: a stage is held for one TMEM read, not for a tile's epilogue, so the MMA thread is never
//     held up by a warp that is late, and the warps drift apart over the four stages instead of moving in step.
//   * Everything a warp keeps is per (32 rows, its 32 columns of every tile): the reference exponent T (guess = the
//     last chunk's maximum, verified on the registers; when it moves by more than the window the running row sums are
//     written out as one record and a new run starts -- no rescaling, no second TMEM pass), the row sums of the
//     current run, and a lower bound of each row's full sum (the a-priori Jensen bound of pack_f16_kernel or all of the
//     row this warp has seen, whichever is larger) from which the candidate limit is taken.  The limit includes the
//     chunk itself and the candidates are captured from the registers.
//   * Records: column sums per (warp block, chunk) with their reference in col_ref; row sums per run in one of 128
//     slots (n tile and column block of the run's first chunk), references in row_ref (NaN = slot unused, cleared by a
//     memset per call).  merge_sweep_kernel folds them.
// ================================================================================================
constexpr int SW_BN = 128;
constexpr int SW_BSTAGES = 2 * FAST_KB;              // two tiles' worth of 16 KB k-blocks
constexpr int SW_TSTAGES = 4;
constexpr int SW_A_BYTES = FAST_KB * FAST_SUB;       // 64 KB: the resident 128-row A block (fp16)
constexpr int SW_BSTAGE_BYTES = FAST_SUB;            // 16 KB: 128 rows x 64 of B
constexpr int SW_SMEM = 1024 + SW_A_BYTES + SW_BSTAGES * SW_BSTAGE_BYTES + 256;
constexpr int SW_ROW_SLOTS = 128;                    // (n tile of 128: up to 32) x (column block c)
constexpr int SW_MAX_NTILES = SW_ROW_SLOTS / 4;

struct SimSweepParams {
  CUtensorMap a, b;   // fp16 operand planes [rows, 256], box 128 x 64, 128B swizzle
  int N, L, S, Lp, Sp;
  int n_tiles;        // Sp / 128
  int total_tiles;    // N * (Lp/128) * n_tiles, m-major
  float c;            // scale * log2(e)  (> 0)
  float thr, max_delta;
  const float* row_lb;      // [N*Lp] log2 of the a-priori (Jensen) lower bound of each row's full sum (pack_f16_kernel)
  const unsigned* norm_sq;  // bits of the largest squared row norm of each side (pack_f16_kernel)
  int* needs_robust;        // bit 1 <- the bound does not allow screening (the caller repeats with algo 2)
  float* colpart;     // [4*Lp/128][N*Sp]          sum_i e_ij over the 32 rows of a warp block
  float* col_ref;     // [N*Lp/128][4][Sp/32]      reference of that sum (-inf: nothing there)
  float* rowpart;     // [SW_ROW_SLOTS][N*Lp]      sum_j e_ij over one run of a warp
  float* row_ref;     // [SW_ROW_SLOTS][N*Lp/32]   reference of the run (NaN: slot unused)
  Candidate* cand;
  int* cand_count;
  int cand_cap;
  long long* trace;   // debug: per CTA and every fourth tile {MMA issue start, MMA issued, warp (2,0) sees full, releases}
  int debug;          // OPB200_SIM_DEBUG: 1 = epilogue warps only hand the accumulators back, 2 = no MMAs are issued
};

__global__ void __launch_bounds__(64 + 128 * FAST_GROUPS, 1) sim_sweep_kernel(const __grid_constant__ SimSweepParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;                // [kb][128 x 64]
  uint8_t* smem_b = smem + SW_A_BYTES;   // [stage][128 x 64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + SW_BSTAGES * SW_BSTAGE_BYTES);
  uint64_t* a_full = bars;
  uint64_t* a_empty = bars + 1;
  uint64_t* b_full = bars + 2;
  uint64_t* b_empty = b_full + SW_BSTAGES;
  uint64_t* tfull = b_empty + SW_BSTAGES;
  uint64_t* tempty = tfull + SW_TSTAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + SW_TSTAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t_begin = static_cast<int>(static_cast<long long>(p.total_tiles) * blockIdx.x / gridDim.x);
  const int t_end = static_cast<int>(static_cast<long long>(p.total_tiles) * (blockIdx.x + 1) / gridDim.x);
  const ScreenPlan plan = screen_plan(p.norm_sq, p.c, p.thr, p.max_delta);
  if (!plan.ok) {
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(p.needs_robust, 2);
    return;
  }
  const float thr_lim = plan.thr_lim;
  const float seed_shift = plan.delta;  // (see sim_fast_kernel)

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a);
    ptx::prefetch_tmap(&p.b);
    ptx::mbar_init(a_full, 1);
    ptx::mbar_init(a_empty, 1);
    for (int s = 0; s < SW_BSTAGES; ++s) {
      ptx::mbar_init(&b_full[s], 1);
      ptx::mbar_init(&b_empty[s], 1);
    }
    for (int g = 0; g < SW_TSTAGES; ++g) {
      ptx::mbar_init(&tfull[g], 1);
      ptx::mbar_init(&tempty[g], 4 * FAST_GROUPS);  // every epilogue warp
    }
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
  const int tiles_per_batch = p.Lp / BLOCK_M;
  const int m_first = t_begin / p.n_tiles, n_first = t_begin - m_first * p.n_tiles;
  const int n_it = t_end - t_begin;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (one thread)
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0, a_phase = 0;
      int m_tile = m_first, n_tile = n_first;
      int b_row0 = (m_tile / tiles_per_batch) * p.Sp, mt_in_batch = m_tile % tiles_per_batch;
      bool new_a = true;
      for (int t = t_begin; t < t_end; ++t) {
        if (new_a) {
          ptx::mbar_wait(a_empty, a_phase ^ 1);
          ptx::mbar_arrive_expect_tx(a_full, SW_A_BYTES);
#pragma unroll
          for (int kb = 0; kb < FAST_KB; ++kb) ptx::tma_load_2d(smem_a + kb * FAST_SUB, &p.a, a_full, kb * BLOCK_K, m_tile * BLOCK_M);
          a_phase ^= 1;
          new_a = false;
        }
        const int b_row = b_row0 + n_tile * SW_BN;
#pragma unroll
        for (int kb = 0; kb < FAST_KB; ++kb) {
          ptx::mbar_wait(&b_empty[stage], phase ^ 1);
          ptx::mbar_arrive_expect_tx(&b_full[stage], SW_BSTAGE_BYTES);
          ptx::tma_load_2d(smem_b + stage * SW_BSTAGE_BYTES, &p.b, &b_full[stage], kb * BLOCK_K, b_row);
          if (++stage == SW_BSTAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++n_tile == p.n_tiles) {
          n_tile = 0, ++m_tile, new_a = true;
          if (++mt_in_batch == tiles_per_batch) mt_in_batch = 0, b_row0 += p.Sp;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread; see sim_fast_kernel)
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::make_idesc_f16(BLOCK_M, SW_BN);
      constexpr int kDescPerStage = SW_BSTAGE_BYTES >> 4;
      const uint64_t a_desc0 = ptx::make_desc_sw128(ptx::smem_u32(smem_a));
      const uint64_t b_desc0 = ptx::make_desc_sw128(ptx::smem_u32(smem_b));
      const bool no_mma = p.debug == 2;
      uint32_t a_phase = 0;
      int n_left = p.n_tiles - n_first;  // tiles left in the current block of rows
      bool new_a = true;
      for (int it0 = 0; it0 < n_it; it0 += SW_TSTAGES) {
        const uint32_t te_parity = ((it0 >> 2) & 1) ^ 1;
#pragma unroll
        for (int u = 0; u < SW_TSTAGES; ++u) {
          if (it0 + u >= n_it) break;
          if (new_a) {
            ptx::mbar_wait(a_full, a_phase);
            a_phase ^= 1;
            new_a = false;
          }
          ptx::mbar_wait(&tempty[u], te_parity);
          ptx::tc_fence_after();
          const bool tr = p.trace != nullptr && u == 0 && (it0 >> 2) < kTraceTilesPerCta;
          if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it0 >> 2)) * 4 + 0] = clock64();
          const uint32_t d_tmem = tmem_base + u * SW_BN;
#pragma unroll
          for (int kb = 0; kb < FAST_KB; ++kb) {
            const int stage = (u & 1) * FAST_KB + kb;
            ptx::mbar_wait(&b_full[stage], static_cast<uint32_t>(u >> 1));
            ptx::tc_fence_after();
            if (!no_mma) {
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k)
                ptx::mma_bf16(d_tmem, a_desc0 + static_cast<uint64_t>(kb * (FAST_SUB >> 4) + 2 * k),
                              b_desc0 + static_cast<uint64_t>(stage * kDescPerStage + 2 * k), idesc, (kb | k) != 0);
            }
            ptx::mma_commit(&b_empty[stage]);
          }
          ptx::mma_commit(&tfull[u]);
          if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it0 >> 2)) * 4 + 1] = clock64();
          const bool last_n = --n_left == 0;
          if (last_n || it0 + u + 1 == n_it) ptx::mma_commit(a_empty);
          if (last_n) n_left = p.n_tiles, new_a = true;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps: (quarter, column block)
    const int cb = (warp - 2) >> 2;   // columns 32 cb .. 32 cb + 31 of every tile
    const int quarter = warp & 3;     // TMEM lanes 32 quarter .. + 31
    const int cq = (lane & 3) * 2;
    const int own = lane & 3;
    const uint64_t c2 = ptx::pack_f32x2(p.c, p.c);
    // the run of chunks that share the reference T: row sums of the thread's four rows, where the run began
    float T = INFINITY;
    uint64_t racc[4] = {0ull, 0ull, 0ull, 0ull};
    int run_slot = -1;                 // -1: no run open
    // lower bound of the full row sums (owner lane: row lane/4 + 8 (lane%4)) and the limits taken from it
    float run_s = 0.f, run_T = 0.f;
    float vlim[4] = {kNegInf, kNegInf, kNegInf, kNegInf};
    int m_tile = m_first, n_tile = n_first;
    int n_batch = m_tile / tiles_per_batch, mt_in_batch = m_tile % tiles_per_batch;
    // per block of rows (recomputed when m_tile changes): the thread's first row, raggedness, record bases
    int i0;
    bool rows_ragged;
    float* col_base;
    float* cref_base;
    float seed_l2 = kNegInf;  // owner lane: log2 of the a-priori lower bound of its row's sum (in units of the screened sums)
    auto enter_rows = [&]() {
      i0 = mt_in_batch * BLOCK_M + quarter * 32 + (lane >> 2);  // this thread's rows: i0 + 8h
      rows_ragged = (mt_in_batch + 1) * BLOCK_M > p.L;
      col_base = p.colpart + ((static_cast<long long>(mt_in_batch) * 4 + quarter) * p.N + n_batch) * p.Sp;
      cref_base = p.col_ref + (static_cast<long long>(m_tile) * 4 + quarter) * (p.Sp / 32);
      seed_l2 = (i0 + 8 * own < p.L && n_batch < p.N) ? p.row_lb[static_cast<long long>(n_batch) * p.Lp + i0 + 8 * own] + seed_shift : kNegInf;
    };
    enter_rows();

    auto flush_run = [&](int n, int mtb) {  // the open run -> one record per row
      if (run_slot < 0) return;
      float s4[4];
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        float lo, hi;
        ptx::unpack_f32x2(racc[h], lo, hi);
        s4[h] = lo + hi;
        racc[h] = 0ull;
      }
      const float s_own = quad_transpose_reduce(s4, lane, [](float x, float y) { return x + y; });
      const long long row0 = static_cast<long long>(n) * p.Lp + mtb * BLOCK_M + quarter * 32;
      p.rowpart[static_cast<long long>(run_slot) * p.N * p.Lp + row0 + (lane >> 2) + 8 * own] = s_own;  // rows that do not exist hold 0
      if (lane == 0) p.row_ref[static_cast<long long>(run_slot) * (p.N * p.Lp / 32) + row0 / 32] = T;
      run_slot = -1;
    };

    for (int it = 0; it < n_it; ++it) {
      const int stage = it & (SW_TSTAGES - 1);
      ptx::mbar_wait(&tfull[stage], (it >> 2) & 1);
      ptx::tc_fence_after();
      const bool tr = p.trace != nullptr && (it & 3) == 0 && (it >> 2) < kTraceTilesPerCta && warp == 4 && lane == 0;
      if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 2] = clock64();
      const int j0 = n_tile * SW_BN + cb * 32;
      const uint32_t tacc = tmem_base + stage * SW_BN + cb * 32 + (static_cast<uint32_t>(quarter * 32) << 16);
      uint32_t va[16], vb[16];
      if (p.debug != 1) {
        ptx::tmem_ld_16x256b_x4(tacc, va);
        ptx::tmem_ld_16x256b_x4(tacc + (16u << 16), vb);
        ptx::tmem_ld_wait();
      }
      {  // the chunk is in registers: the stage goes back to the MMA thread
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&tempty[stage]);
        if (tr) p.trace[(blockIdx.x * kTraceTilesPerCta + (it >> 2)) * 4 + 3] = clock64();
      }
      if (p.debug != 1) {
        if (rows_ragged || j0 + 32 > p.S) {  // rows / columns that do not exist: -inf, i.e. e = 0 and no say in the maximum
#pragma unroll
          for (int q = 0; q < 16; ++q) {
            const bool cv = j0 + 8 * (q >> 2) + cq + (q & 1) < p.S;
            if (!cv || i0 + 8 * ((q >> 1) & 1) >= p.L) va[q] = __float_as_uint(kNegInf);
            if (!cv || i0 + 8 * (2 + ((q >> 1) & 1)) >= p.L) vb[q] = __float_as_uint(kNegInf);
          }
        }
        float cmax[4] = {kNegInf, kNegInf, kNegInf, kNegInf};  // raw maxima of the chunk per row
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          cmax[0] = ptx::max3(cmax[0], __uint_as_float(va[4 * k]), __uint_as_float(va[4 * k + 1]));
          cmax[1] = ptx::max3(cmax[1], __uint_as_float(va[4 * k + 2]), __uint_as_float(va[4 * k + 3]));
          cmax[2] = ptx::max3(cmax[2], __uint_as_float(vb[4 * k]), __uint_as_float(vb[4 * k + 1]));
          cmax[3] = ptx::max3(cmax[3], __uint_as_float(vb[4 * k + 2]), __uint_as_float(vb[4 * k + 3]));
        }
        const float tm = fmaxf(fmaxf(cmax[0], cmax[1]), fmaxf(cmax[2], cmax[3]));
        const float ymax = key_float(__reduce_max_sync(0xffffffffu, float_key(tm))) * p.c;
        float* __restrict__ colp = col_base + j0;
        float* __restrict__ cref = cref_base + (j0 >> 5);
        if (ymax == kNegInf) {  // ragged edge: nothing exists in this chunk
          colp[lane] = 0.f;
          if (lane == 0) *cref = kNegInf;
        } else {
          if (!(fabsf(ymax - T) <= kRefWindow)) {  // the guess does not hold: close the run, start one at this chunk
            flush_run(n_batch, mt_in_batch);
            T = ymax;
          }
          if (run_slot < 0) run_slot = n_tile * 4 + cb;
          const uint64_t t2 = ptx::pack_f32x2(-T, -T);
          uint64_t rc[4] = {0ull, 0ull, 0ull, 0ull};  // this chunk's row sums
          float f[4][2];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t e0 = exp2_pair(va[4 * k], va[4 * k + 1], c2, t2), e1 = exp2_pair(va[4 * k + 2], va[4 * k + 3], c2, t2);
            const uint64_t e2 = exp2_pair(vb[4 * k], vb[4 * k + 1], c2, t2), e3 = exp2_pair(vb[4 * k + 2], vb[4 * k + 3], c2, t2);
            rc[0] = ptx::add_f32x2(rc[0], e0);
            rc[1] = ptx::add_f32x2(rc[1], e1);
            rc[2] = ptx::add_f32x2(rc[2], e2);
            rc[3] = ptx::add_f32x2(rc[3], e3);
            ptx::unpack_f32x2(ptx::add_f32x2(ptx::add_f32x2(e0, e1), ptx::add_f32x2(e2, e3)), f[k][0], f[k][1]);
          }
          // column sums: halving butterfly over the row bits of the lane id (16, 8, 4): 4 + 2 + 1 shuffles
          float g[2][2], hh[2];
          {
            const bool up = (lane & 16) != 0;
#pragma unroll
            for (int kk = 0; kk < 2; ++kk)
#pragma unroll
              for (int b = 0; b < 2; ++b) {
                const float keep = up ? f[kk + 2][b] : f[kk][b];
                const float send = up ? f[kk][b] : f[kk + 2][b];
                g[kk][b] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
              }
          }
          {
            const bool up = (lane & 8) != 0;
#pragma unroll
            for (int b = 0; b < 2; ++b) {
              const float keep = up ? g[1][b] : g[0][b];
              const float send = up ? g[0][b] : g[1][b];
              hh[b] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
            }
          }
          const bool up4 = (lane & 4) != 0;
          const float cs = (up4 ? hh[1] : hh[0]) + __shfl_xor_sync(0xffffffffu, up4 ? hh[0] : hh[1], 4);
          colp[((lane >> 3) & 3) * 8 + cq + ((lane >> 2) & 1)] = cs;  // lane bits 4,3 select k, bit 2 the column of the pair
          if (lane == 0) *cref = T;
          // row sums: into the run, and (owner lanes) into the lower bound of the full row sum
          float s4[4];
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            racc[h] = ptx::add_f32x2(racc[h], rc[h]);
            float lo, hi;
            ptx::unpack_f32x2(rc[h], lo, hi);
            s4[h] = lo + hi;
          }
          const float s_own = quad_transpose_reduce(s4, lane, [](float x, float y) { return x + y; });
          // candidate capture: conf_ij > thr needs e_ij > thr * (row sum) and the part of the row this warp has seen,
          // this chunk included, bounds that sum from below: acc_ij > vlim_i in raw accumulator units
          run_s = fminf(run_s * ptx::ex2(fminf(run_T - T, 60.f)), 1e30f) + s_own;  // (clamps keep it a finite lower bound)
          run_T = T;
          const float lim = thr_lim * run_s;
          const float vlim_own = (i0 + 8 * own < p.L && lim > 0.f)
                                     ? __fdividef(fmaxf(__log2f(lim) + T, __log2f(thr_lim) + seed_l2), p.c) : INFINITY;
#pragma unroll
          for (int h = 0; h < 4; ++h) vlim[h] = __shfl_sync(0xffffffffu, vlim_own, (lane & ~3) + h);
          if (cmax[0] > vlim[0] || cmax[1] > vlim[1] || cmax[2] > vlim[2] || cmax[3] > vlim[3]) {  // only the lanes that hold one
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const int h = q >> 1, b = q & 1;
                const float val = __uint_as_float(h < 2 ? va[4 * k + 2 * h + b] : vb[4 * k + 2 * (h - 2) + b]);
                if (val > vlim[h]) {  // (-inf where nothing exists)
                  const int slot = atomicAdd(p.cand_count, 1);
                  if (slot < p.cand_cap)
                    p.cand[slot] = Candidate{n_batch * p.L + i0 + 8 * h, j0 + 8 * k + cq + b, __fmul_rn(val, p.c), 2};
                }
              }
          }
        }
      }
      if (++n_tile == p.n_tiles) {  // the next tile belongs to another block of rows
        flush_run(n_batch, mt_in_batch);
        n_tile = 0, ++m_tile;
        if (++mt_in_batch == tiles_per_batch) mt_in_batch = 0, ++n_batch;
        enter_rows();
        run_s = 0.f, run_T = 0.f;
#pragma unroll
        for (int h = 0; h < 4; ++h) vlim[h] = kNegInf;
      }
    }
    flush_run(n_batch, mt_in_batch);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// Records of sim_sweep_kernel -> (reference, log2 sum) per row and per column.  Block = 32 outputs x 8 slices of the
// part list, slices combined through shared memory in a fixed order.  A finished sum of an existing row / column that
// is not safely above the flush-to-zero range (or is NaN / inf) raises bit 0 of *needs_robust.
__global__ void __launch_bounds__(256) merge_sweep_kernel(const float* __restrict__ rowpart, const float* __restrict__ row_ref,
                                                          const float* __restrict__ colpart, const float* __restrict__ col_ref,
                                                          int N, int L, int S, int Lp, int Sp, float2* __restrict__ lse_row,
                                                          float2* __restrict__ lse_col, int* __restrict__ needs_robust) {
  __shared__ float2 sm[8][33];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const long long n_rows = static_cast<long long>(N) * Lp, n_cols = static_cast<long long>(N) * Sp;
  const long long row_blocks = n_rows / 32;  // Lp and Sp are multiples of 128
  const bool is_row = blockIdx.x < row_blocks;
  const long long idx = (is_row ? blockIdx.x : blockIdx.x - row_blocks) * 32ll + lane;
  const int mtb = Lp / BLOCK_M;
  float M = kNegInf, Sm = 0.f;
  bool exists;
  if (is_row) {
    exists = static_cast<int>(idx % Lp) < L;
    const long long blk = idx / 32;  // = blockIdx.x: one 32-row warp block per thread block
#pragma unroll 4
    for (int k = slice; k < SW_ROW_SLOTS; k += 8) {
      const float T = row_ref[k * (n_rows / 32) + blk];
      if (T > -3e38f) {  // (false for the NaN of an unused slot)
        const float v = rowpart[k * n_rows + idx];
        const float Mn = fmaxf(M, T);
        Sm = Sm * exp2f(M - Mn) + v * exp2f(T - Mn);  // exp2f(-inf) = 0 covers the first term
        M = Mn;
      }
    }
  } else {
    const int n = static_cast<int>(idx / Sp), j = static_cast<int>(idx - static_cast<long long>(n) * Sp);
    exists = j < S;
    const float* tr = col_ref + static_cast<long long>(n) * mtb * 4 * (Sp / 32) + j / 32;
#pragma unroll 4
    for (int k = slice; k < 4 * mtb; k += 8) {  // k = (m tile, quarter)
      const float T = tr[static_cast<long long>(k) * (Sp / 32)];
      if (T > -3e38f) {
        const float v = colpart[(static_cast<long long>(k) * N + n) * Sp + j];
        const float Mn = fmaxf(M, T);
        Sm = Sm * exp2f(M - Mn) + v * exp2f(T - Mn);
        M = Mn;
      }
    }
  }
  sm[slice][lane] = make_float2(M, Sm);
  __syncthreads();
  if (slice == 0) {
    float Mt = kNegInf;
#pragma unroll
    for (int q = 0; q < 8; ++q) Mt = fmaxf(Mt, sm[q][lane].x);
    float St = 0.f;
    if (Mt > kNegInf) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float2 pq = sm[q][lane];
        if (pq.x > kNegInf) St += pq.y * exp2f(pq.x - Mt);
      }
    }
    const bool ok = exists && Mt > kNegInf;
    (is_row ? lse_row : lse_col)[idx] = make_float2(ok ? Mt : kNegInf, ok ? log2f(St) : 0.f);
    if (ok && !(St >= kTinySum && St < 3e38f)) atomicOr(needs_robust, 1);
  }
}

struct BorderRule {
  int mode;    // 0: OnePosePlus (query grid, top/left only); 1: LoFTR (both grids, four sides)
  int b;       // border_rm
  int h0, w0;  // grid of the rows (LoFTR image0); unused for mode 0
  int h1, w1;  // grid of the columns (query image / image1)
  const int* valid_hw;  // nullable [N,4] = (h0, w0, h1, w1) valid extents from the padding masks (LoFTR)
};

__device__ __forceinline__ bool border_keep(const BorderRule& br, int n, int i, int j) {
  if (br.b <= 0) return true;
  const int r1 = j / br.w1, c1 = j - r1 * br.w1;
  if (br.mode == 0) return r1 >= br.b && c1 >= br.b;
  const int r0 = i / br.w0, c0 = i - r0 * br.w0;
  int h0 = br.h0, w0 = br.w0, h1 = br.h1, w1 = br.w1;
  if (br.valid_hw != nullptr) {
    h0 = br.valid_hw[4 * n + 0], w0 = br.valid_hw[4 * n + 1], h1 = br.valid_hw[4 * n + 2], w1 = br.valid_hw[4 * n + 3];
  }
  return r0 >= br.b && c0 >= br.b && r1 >= br.b && c1 >= br.b && r0 < h0 - br.b && c0 < w0 - br.b && r1 < h1 - br.b &&
         c1 < w1 - br.b;
}

// fp16 operand planes of the screening sweep from the bf16 hi + lo planes, both sides in one launch (8 elements per
// thread, one warp per 256-wide row).  The same pass reduces what the soundness of the screening rests on: the largest
// L2 norm of an existing row of each side (as the bits of its square, atomicMax into norm_sq[0 / 1]).  With operands
// rounded to fp16 (relative error <= 2^-11, absolute error <= 2^-25 in the sub-normal range) the screening similarity
// y_s = c * sum_k a_k b_k differs from the exact one by at most
//     delta = c * (2^-10 (1 + 2^-12) |a| |b|  +  2^-25 sqrt(C) (|a| + |b|))           (Cauchy-Schwarz, in octaves)
// screen_plan() turns that bound into the candidate limit of the sweep, or into the decision not to screen at all.
struct PackSide {
  const uint4* hi;
  const uint4* lo;
  uint4* out;
  long long n_vec;
  int rows_valid, rows_padded;  // per batch element: rows beyond rows_valid are padding and have no say in the norm
};

// Column mean of side b per batch element (fp32, deterministic: fixed chunks, the block that finishes last adds the
// chunk sums in chunk order).  It gives every row of side a an a-priori lower bound of its softmax sum before the sweep
// has seen a single column:  sum_j 2^(y_ij) >= S 2^(mean_j y_ij) = S 2^(c a_i . mean_j b_j)  (Jensen).  For the rows
// that hold no match -- nearly all of them -- this bound is within a small factor of the sum, which keeps the first
// column tiles of a row block from flooding the candidate list (their own partial sum is a poor bound).
constexpr int MEAN_CHUNKS = 64;
// thread = (8 channels, row lane): 16-byte loads of the hi and lo planes, 8 rows of a chunk in flight per block step
// The same pass writes side b's fp16 operand plane of the screening sweep and takes its largest row norm (out / norm_sq
// non-null): side b is read once, not once for the mean and once for the pack.
__global__ void __launch_bounds__(256) colmean_kernel(const __nv_bfloat16* __restrict__ hi, const __nv_bfloat16* __restrict__ lo,
                                                      int S, int Sp, float* __restrict__ part /* [N][MEAN_CHUNKS][256] */,
                                                      float* __restrict__ mean /* [N][256] */, int* __restrict__ done /* [N] */,
                                                      __half* __restrict__ out /* [N][Sp][256] or null */,
                                                      unsigned* __restrict__ norm_sq /* bits of the largest squared row norm */) {
  __shared__ float red[8][FAST_C];
  __shared__ int last;
  const int n = blockIdx.x / MEAN_CHUNKS, ch = blockIdx.x % MEAN_CHUNKS;
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int per = (S + MEAN_CHUNKS - 1) / MEAN_CHUNKS;
  const int j0 = ch * per, j1 = min(S, j0 + per);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float worst = 0.f;
  const long long base = static_cast<long long>(n) * Sp * FAST_C + cg * 8;
  auto one_row = [&](int j, bool valid) {
    const uint4 h = *reinterpret_cast<const uint4*>(hi + base + static_cast<long long>(j) * FAST_C);
    const uint4 l = *reinterpret_cast<const uint4*>(lo + base + static_cast<long long>(j) * FAST_C);
    const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
    uint32_t ow[4];
    float sq = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float x0 = __uint_as_float(hw[q] << 16) + __uint_as_float(lw[q] << 16);
      const float x1 = __uint_as_float(hw[q] & 0xffff0000u) + __uint_as_float(lw[q] & 0xffff0000u);
      if (valid) acc[2 * q] += x0, acc[2 * q + 1] += x1;
      sq = fmaf(x0, x0, fmaf(x1, x1, sq));
      const __half2 v = __floats2half2_rn(x0, x1);
      ow[q] = *reinterpret_cast<const uint32_t*>(&v);
    }
    if (out != nullptr) {
      *reinterpret_cast<uint4*>(out + base + static_cast<long long>(j) * FAST_C) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);  // (a warp = one row)
      if (valid) worst = fmaxf(worst, sq <= 3e38f ? sq : INFINITY);  // (a NaN norm counts as +inf, as in pack_f16_kernel)
    }
  };
#pragma unroll 4
  for (int j = j0 + rl; j < j1; j += 8) one_row(j, true);  // fixed assignment of rows to row lanes: deterministic
  // the padding rows S .. Sp - 1 are packed by the last chunk; they have no say in the mean or the norm
  if (out != nullptr && ch == MEAN_CHUNKS - 1)
    for (int j = S + rl; j < Sp; j += 8) one_row(j, false);
  if (norm_sq != nullptr && cg == 0 && worst > 0.f) atomicMax(norm_sq, __float_as_uint(worst));
#pragma unroll
  for (int q = 0; q < 8; ++q) red[rl][cg * 8 + q] = acc[q];
  __syncthreads();
  const int c = threadIdx.x;
  float t = 0.f;
#pragma unroll
  for (int r = 0; r < 8; ++r) t += red[r][c];
  part[(static_cast<long long>(n) * MEAN_CHUNKS + ch) * FAST_C + c] = t;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(&done[n], 1) == MEAN_CHUNKS - 1;
  __syncthreads();
  if (last) {
    __threadfence();
    float t4[4] = {0.f, 0.f, 0.f, 0.f};  // fixed order: chunk k goes to slot k % 4
#pragma unroll 8
    for (int k = 0; k < MEAN_CHUNKS; ++k) t4[k & 3] += __ldcg(&part[(static_cast<long long>(n) * MEAN_CHUNKS + k) * FAST_C + c]);
    mean[n * FAST_C + c] = ((t4[0] + t4[1]) + (t4[2] + t4[3])) / static_cast<float>(S);
  }
}

__global__ void __launch_bounds__(256) pack_f16_kernel(PackSide sa, PackSide sb, int blocks_a, unsigned* __restrict__ norm_sq,
                                                       const float* __restrict__ mean_b, float c, int S,
                                                       float* __restrict__ row_lb) {
  const bool second = static_cast<int>(blockIdx.x) >= blocks_a;
  const PackSide& sd = second ? sb : sa;
  const long long bid = second ? blockIdx.x - blocks_a : blockIdx.x, nblk = second ? gridDim.x - blocks_a : blocks_a;
  float worst = 0.f;
  for (long long t = bid * 256 + threadIdx.x; t < sd.n_vec; t += nblk * 256) {  // (n_vec is a multiple of 32: whole warps)
    const uint4 h = sd.hi[t], l = sd.lo[t];
    const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
    uint32_t ow[4];
    float sq = 0.f, dot = 0.f;
    const long long row = t / (FAST_C / 8);
    const int nb = static_cast<int>(row / sd.rows_padded);
    const bool valid = static_cast<int>(row - static_cast<long long>(nb) * sd.rows_padded) < sd.rows_valid;
    float mu[8];
    if (!second && row_lb != nullptr) {
      const float4* m4 = reinterpret_cast<const float4*>(mean_b + nb * FAST_C + (threadIdx.x & 31) * 8);
      const float4 m0 = m4[0], m1 = m4[1];
      mu[0] = m0.x, mu[1] = m0.y, mu[2] = m0.z, mu[3] = m0.w, mu[4] = m1.x, mu[5] = m1.y, mu[6] = m1.z, mu[7] = m1.w;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float x0 = __uint_as_float(hw[q] << 16) + __uint_as_float(lw[q] << 16);
      const float x1 = __uint_as_float(hw[q] & 0xffff0000u) + __uint_as_float(lw[q] & 0xffff0000u);
      sq = fmaf(x0, x0, fmaf(x1, x1, sq));
      if (!second && row_lb != nullptr) dot = fmaf(x0, mu[2 * q], fmaf(x1, mu[2 * q + 1], dot));
      const __half2 v = __floats2half2_rn(x0, x1);
      ow[q] = *reinterpret_cast<const uint32_t*>(&v);
    }
    sd.out[t] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sq += __shfl_xor_sync(0xffffffffu, sq, o);
      dot += __shfl_xor_sync(0xffffffffu, dot, o);
    }
    if (valid) worst = fmaxf(worst, sq <= 3e38f ? sq : INFINITY);  // (a NaN norm must not vanish in the maximum: it counts as +inf)
    if (!second && row_lb != nullptr && (threadIdx.x & 31) == 0)  // log2(S 2^(c a . mean b)); a non-finite one is no bound
      row_lb[row] = (valid && fabsf(dot) <= 3e38f) ? fmaf(c, dot, __log2f(static_cast<float>(S))) : kNegInf;
  }
  if ((threadIdx.x & 31) == 0 && worst > 0.f) atomicMax(&norm_sq[second ? 1 : 0], __float_as_uint(worst));
}

// One warp per captured element of the screening sweep: the similarity again from the hi + lo planes in fp32, and the
// exchange of its term in the row / column sums:  delta = 2^(y_exact - LSE) - 2^(y_screen - LSE)  in units of the
// screening sum, accumulated as 2^-40 fixed point (integer atomics: independent of the order of arrival).
__global__ void __launch_bounds__(256) cand_rescore_kernel(Candidate* __restrict__ cand, const int* __restrict__ count, int cap,
                                                           int L, int Lp, int Sp, const __nv_bfloat16* __restrict__ a_hi,
                                                           const __nv_bfloat16* __restrict__ a_lo,
                                                           const __nv_bfloat16* __restrict__ b_hi,
                                                           const __nv_bfloat16* __restrict__ b_lo, float c,
                                                           const float2* __restrict__ lse_row, const float2* __restrict__ lse_col,
                                                           unsigned long long* __restrict__ rowdelta,
                                                           unsigned long long* __restrict__ coldelta,
                                                           unsigned* __restrict__ observed /* bits of max |y - y_screen| */,
                                                           int* __restrict__ needs_robust, const unsigned* __restrict__ norm_sq,
                                                           float thr, float max_delta) {
  const int n_c = min(*count, cap);
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float log2_thr_lim = log2f(screen_plan(norm_sq, c, thr, max_delta).thr_lim);
  for (int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < n_c; t += warps) {
    const Candidate cd = cand[t];
    if (cd.pad != 2) continue;
    const int n = cd.row / L, i = cd.row - n * L;
    const long long ra = static_cast<long long>(n) * Lp + i, rb = static_cast<long long>(n) * Sp + cd.col;
    const uint4 ah = *reinterpret_cast<const uint4*>(a_hi + ra * FAST_C + lane * 8);
    const uint4 al = *reinterpret_cast<const uint4*>(a_lo + ra * FAST_C + lane * 8);
    const uint4 bh = *reinterpret_cast<const uint4*>(b_hi + rb * FAST_C + lane * 8);
    const uint4 bl = *reinterpret_cast<const uint4*>(b_lo + rb * FAST_C + lane * 8);
    const uint32_t ahw[4] = {ah.x, ah.y, ah.z, ah.w}, alw[4] = {al.x, al.y, al.z, al.w};
    const uint32_t bhw[4] = {bh.x, bh.y, bh.z, bh.w}, blw[4] = {bl.x, bl.y, bl.z, bl.w};
    float acc = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float a0 = __uint_as_float(ahw[q] << 16) + __uint_as_float(alw[q] << 16);
      const float a1 = __uint_as_float(ahw[q] & 0xffff0000u) + __uint_as_float(alw[q] & 0xffff0000u);
      const float b0 = __uint_as_float(bhw[q] << 16) + __uint_as_float(blw[q] << 16);
      const float b1 = __uint_as_float(bhw[q] & 0xffff0000u) + __uint_as_float(blw[q] & 0xffff0000u);
      acc = fmaf(a0, b0, acc);
      acc = fmaf(a1, b1, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      const float y = __fmul_rn(acc, c);
      const float2 lr = lse_row[ra], lc = lse_col[rb];
      // Which near misses the sweep listed depends on the order in which its warps met the tiles; what is exchanged
      // must not.  An element that can reach thr has its screening row-softmax above thr_lim (the sweep's own capture
      // condition, here against the FINAL row sum): everything else is dropped, so the exchanged set -- and every sum
      // and confidence after it -- is a function of the data alone.
      if (!((cd.conf - lr.x) - lr.y > log2_thr_lim)) {
        cand[t].conf = kNegInf;
        continue;
      }
      const float dr = exp2f((y - lr.x) - lr.y) - exp2f((cd.conf - lr.x) - lr.y);
      const float dc = exp2f((y - lc.x) - lc.y) - exp2f((cd.conf - lc.x) - lc.y);
      atomicAdd(&rowdelta[ra], static_cast<unsigned long long>(__double2ll_rn(static_cast<double>(dr) * kDeltaScale)));
      atomicAdd(&coldelta[rb], static_cast<unsigned long long>(__double2ll_rn(static_cast<double>(dc) * kDeltaScale)));
      const float obs = fabsf(y - cd.conf);
      if (obs > __uint_as_float(*observed)) atomicMax(observed, __float_as_uint(obs));  // (non-negative floats order like their bits)
      if (!(obs <= kScreenMaxObserved)) atomicOr(needs_robust, 2);
      cand[t].conf = y;
    }
  }
}

// LSE of a row / column whose captured terms were exchanged for their exact values: log2(sum * (1 + delta))
__device__ __forceinline__ float2 corrected_lse(float2 lse, unsigned long long delta_fx) {
  const float rel = 1.f + static_cast<float>(static_cast<double>(static_cast<long long>(delta_fx)) * (1.0 / kDeltaScale));
  lse.y += log2f(fmaxf(rel, 1e-30f));
  return lse;
}

__global__ void cand_max_kernel(Candidate* __restrict__ cand, const int* __restrict__ count, int cap, int L, int S,
                                const float2* __restrict__ lse_row, const float2* __restrict__ lse_col, int Lp, int Sp, float thr,
                                const unsigned long long* __restrict__ rowdelta, const unsigned long long* __restrict__ coldelta,
                                unsigned* __restrict__ rowmax, unsigned* __restrict__ colmax) {
  const int n_c = min(*count, cap);
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n_c; t += gridDim.x * blockDim.x) {
    Candidate c = cand[t];
    if (c.pad == 2) {  // screening sweep: y is the rescored value, the LSEs take their corrections
      const int n = c.row / L, i = c.row - n * L;
      const long long ra = static_cast<long long>(n) * Lp + i, rb = static_cast<long long>(n) * Sp + c.col;
      const float cf = conf_from(c.conf, corrected_lse(lse_row[ra], rowdelta[ra]), corrected_lse(lse_col[rb], coldelta[rb]));
      c.conf = cf > thr ? cf : 0.f;
      c.pad = 0;
      cand[t] = c;
    }
    if (c.pad == 1) {  // fast path: the entry holds y; its confidence can be formed now that the LSEs are final
      const int n = c.row / L, i = c.row - n * L;
      const float cf = conf_from(c.conf, lse_row[static_cast<long long>(n) * Lp + i], lse_col[static_cast<long long>(n) * Sp + c.col]);
      c.conf = cf > thr ? cf : 0.f;
      c.pad = 0;
      cand[t] = c;
    }
    if (c.conf == 0.f) continue;
    const unsigned bits = __float_as_uint(c.conf);  // conf > thr > 0: the bit pattern orders like the value
    atomicMax(&rowmax[c.row], bits);
    atomicMax(&colmax[(c.row / L) * static_cast<long long>(S) + c.col], bits);
  }
}

__global__ void cand_pick_kernel(const Candidate* __restrict__ cand, const int* __restrict__ count, int cap, int L, int S,
                                 const unsigned* __restrict__ rowmax, const unsigned* __restrict__ colmax, BorderRule br,
                                 int* __restrict__ rowj) {
  const int n_c = min(*count, cap);
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n_c; t += gridDim.x * blockDim.x) {
    const Candidate c = cand[t];
    if (c.conf == 0.f) continue;  // fast-path candidate that did not reach thr
    const unsigned bits = __float_as_uint(c.conf);
    const int n = c.row / L, i = c.row - n * L;
    if (bits == rowmax[c.row] && bits == colmax[n * static_cast<long long>(S) + c.col] && border_keep(br, n, i, c.col))
      atomicMin(&rowj[c.row], c.col);  // ties inside a row: the first j, as mask.max(dim=2) returns
  }
}

// Ordered compaction over the N*L rows (ascending (b, i) like torch.where) on up to COMPACT_BLOCKS blocks.  A block
// takes the next row segment by ticket (so every segment before it is owned by a block that is already running),
// counts its matches, publishes the count as one 64-bit word {1, count}, and sums the words of the segments before it
// (decoupled look-back: all blocks count in parallel, the wait is one L2 round trip).  Inside a block, warp w owns a
// contiguous slice and walks it 32 rows at a time (coalesced loads, ballot + popc for the in-warp order).
constexpr int COMPACT_BLOCKS = 64;
__global__ void __launch_bounds__(256) compact_kernel(const int* __restrict__ rowj, const unsigned* __restrict__ rowmax,
                                                      int n_rows, int L, long long* __restrict__ b_ids,
                                                      long long* __restrict__ i_ids, long long* __restrict__ j_ids,
                                                      float* __restrict__ mconf, int capacity, int* __restrict__ out_count,
                                                      const int* __restrict__ cand_count, int cand_cap, int* __restrict__ overflow,
                                                      int fast, unsigned long long* __restrict__ agg, int* __restrict__ ticket) {
  __shared__ int seg_s, base_s;
  __shared__ int warp_cnt[8], warp_look[8];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (t == 0) seg_s = atomicAdd(ticket, 1);
  __syncthreads();
  const int seg = seg_s;
  const int per_seg = ((n_rows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x) + 255) / 256 * 256;
  const int per_warp = per_seg / 8;  // a multiple of 32
  const int r0 = min(seg * per_seg + warp * per_warp, n_rows), r1 = min(r0 + per_warp, n_rows);
  int cnt = 0;
  for (int base = r0; base < r1; base += 32 * 8) {  // 8 independent loads in flight per lane
    int jv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int r = base + u * 32 + lane;
      jv[u] = r < r1 ? rowj[r] : kNoMatch;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) cnt += __popc(__ballot_sync(0xffffffffu, jv[u] != kNoMatch));
  }
  if (lane == 0) warp_cnt[warp] = cnt;
  __syncthreads();
  int block_total = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) block_total += warp_cnt[q];
  if (t == 0) atomicExch(&agg[seg], (1ull << 32) | static_cast<unsigned>(block_total));
  int look = 0;
  for (int k = t; k < seg; k += 256) {
    unsigned long long v;
    do {
      v = *reinterpret_cast<const volatile unsigned long long*>(&agg[k]);
    } while ((v >> 32) == 0ull);
    look += static_cast<int>(v & 0xffffffffull);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) look += __shfl_xor_sync(0xffffffffu, look, o);
  if (lane == 0) warp_look[warp] = look;
  __syncthreads();
  if (t == 0) {
    int base = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) base += warp_look[q];
    base_s = base;
    if (seg == static_cast<int>(gridDim.x) - 1) {
      const int total = base + block_total;
      *out_count = total;
      const bool cand_over = *cand_count > cand_cap;
      // fast path: a full candidate list or an FTZ-endangered sum asks for the two-pass rerun instead
      overflow[0] = (!fast && cand_over) || (total > capacity);
      // rerun code: 2 = the screening sweep declined (operand norms): repeat single-pass on the three split products -- it
      // wrote no statistics, so bit 0 means nothing then; 1 = take the two-pass path (dynamic range, candidate list full)
      const int flag = cand_count[2];
      overflow[1] = !fast ? 0 : ((flag & 2) ? 2 : ((cand_over || (flag & 1)) ? 1 : 0));
    }
  }
  __syncthreads();
  int pos = base_s;
  for (int q = 0; q < warp; ++q) pos += warp_cnt[q];
  for (int base = r0; base < r1; base += 32 * 8) {
    int jv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int r = base + u * 32 + lane;
      jv[u] = r < r1 ? rowj[r] : kNoMatch;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int r = base + u * 32 + lane;
      const int j = jv[u];
      const unsigned ball = __ballot_sync(0xffffffffu, j != kNoMatch);
      if (j != kNoMatch) {
        const int o = pos + __popc(ball & ((1u << lane) - 1));
        if (o < capacity) {
          b_ids[o] = r / L;
          i_ids[o] = r % L;
          j_ids[o] = j;
          mconf[o] = __uint_as_float(rowmax[r]);
        }
      }
      pos += __popc(ball);
    }
  }
}

__global__ void fill_sim_row_base(int* base, int tiles_per_batch, int N, int Sp) {
  for (int t = threadIdx.x + blockIdx.x * blockDim.x; t < tiles_per_batch * N; t += blockDim.x * gridDim.x)
    base[t] = (t / tiles_per_batch) * Sp;
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct CoarseWs {
  __nv_bfloat16 *a_hi, *a_lo, *b_hi, *b_lo;
  __half *a_f16, *b_f16;                       // operand planes of the screening sweep
  unsigned long long *rowdelta, *coldelta;     // [N*Lp], [N*Sp] fixed-point LSE corrections (cleared per call)
  float2 *rowpart, *colpart;                   // two-pass: (max, sum) records; single pass: plain float sums
  float2 *lse_row, *lse_col;
  float* tile_ref;                             // single pass: reference exponent per (m tile, quarter, n tile | 32-column chunk)
  float* row_ref;                              // sweep kernel: reference of each row-sum record (0xff-filled per call)
  float *row_lb, *mean_part, *mean_b;          // screening: Jensen lower bound per row; column mean of side b
  int* mean_done;                              // [N] chunk counters of colmean_kernel (cleared per call)
  unsigned long long* agg;                     // [COMPACT_BLOCKS] look-back words of the compaction (cleared per call)
  int* ticket;
  Candidate* cand;
  unsigned *rowmax, *colmax;
  int *rowj, *row_base, *counters;  // counters: [0] cand_count, [2] rerun flags, [4], [5] largest squared row norms (bits),
                                    // [6] largest |exact - screening| similarity seen by the rescore pass (bits, octaves)
  int cand_cap;
  int Lp, Sp;
  size_t zero_begin, zero_bytes;  // region that must be cleared per call (rowmax, colmax, counters)
  size_t bytes;
};

static CoarseWs carve_coarse(void* base, int N, int L, int S, int C, float thr) {
  CoarseWs w;
  w.Lp = static_cast<int>(align_up(L, BLOCK_M));
  w.Sp = static_cast<int>(align_up(S, SIM_BN));
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? static_cast<uint8_t*>(base) + off : nullptr;
    off += align_up(bytes, 1024);
    return p;
  };
  const size_t ra = static_cast<size_t>(N) * w.Lp, rb = static_cast<size_t>(N) * w.Sp;
  w.a_hi = static_cast<__nv_bfloat16*>(take(ra * C * 2));
  w.a_lo = static_cast<__nv_bfloat16*>(take(ra * C * 2));
  w.b_hi = static_cast<__nv_bfloat16*>(take(rb * C * 2));
  w.b_lo = static_cast<__nv_bfloat16*>(take(rb * C * 2));
  w.a_f16 = static_cast<__half*>(take(ra * C * 2));
  w.b_f16 = static_cast<__half*>(take(rb * C * 2));
  {
    const size_t two_pass = static_cast<size_t>(w.Sp / SIM_BN) * ra * 8, sweep = static_cast<size_t>(SW_ROW_SLOTS) * ra * 4;
    w.rowpart = static_cast<float2*>(take(two_pass > sweep ? two_pass : sweep));
  }
  w.colpart = static_cast<float2*>(take(static_cast<size_t>(w.Lp / BLOCK_M) * rb * 16));  // single pass: 4 float rows per m tile
  w.tile_ref = static_cast<float*>(take(ra / BLOCK_M * 4 * (w.Sp / 32) * 4));  // (one per 32 columns for the sweep kernel)
  w.row_ref = static_cast<float*>(take(static_cast<size_t>(SW_ROW_SLOTS) * (ra / 32) * 4));
  w.row_lb = static_cast<float*>(take(ra * 4));
  w.mean_part = static_cast<float*>(take(static_cast<size_t>(N) * MEAN_CHUNKS * FAST_C * 4));
  w.mean_b = static_cast<float*>(take(static_cast<size_t>(N) * FAST_C * 4));
  w.lse_row = static_cast<float2*>(take(ra * 8));
  w.lse_col = static_cast<float2*>(take(rb * 8));
  const double bound = std::ceil(static_cast<double>(N) * (L < S ? L : S) / (thr > 1e-6f ? thr : 1e-6f)) + 64;
  const double all = static_cast<double>(N) * L * S;
  const double roomy = bound + 8.0 * N * L + 65536;  // the single-pass capture also lists near misses
  w.cand_cap = static_cast<int>(roomy < all ? roomy : all);
  w.cand = static_cast<Candidate*>(take(static_cast<size_t>(w.cand_cap) * sizeof(Candidate)));
  w.rowj = static_cast<int*>(take(static_cast<size_t>(N) * L * 4));
  w.row_base = static_cast<int*>(take(ra / BLOCK_M * 4 + 4));
  w.zero_begin = off;
  w.rowmax = static_cast<unsigned*>(take(static_cast<size_t>(N) * L * 4));
  w.colmax = static_cast<unsigned*>(take(static_cast<size_t>(N) * S * 4));
  w.counters = static_cast<int*>(take(32));
  w.mean_done = static_cast<int*>(take(static_cast<size_t>(N) * 4));
  w.agg = static_cast<unsigned long long*>(take(COMPACT_BLOCKS * 8));
  w.ticket = static_cast<int*>(take(4));
  w.rowdelta = static_cast<unsigned long long*>(take(ra * 8));
  w.coldelta = static_cast<unsigned long long*>(take(rb * 8));
  w.zero_bytes = off - w.zero_begin;
  w.bytes = off;
  return w;
}

}  // namespace opb

using namespace opb;

#include <mutex>
#include <vector>
namespace {
std::mutex g_kt_mutex;
bool g_kt_armed = false;
float g_screen_max_delta = opb::kScreenMaxDelta;
float g_screen_lim_scale = getenv("OPB200_SCREEN_LIM_SCALE") ? static_cast<float>(atof(getenv("OPB200_SCREEN_LIM_SCALE"))) : 1.f;
std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_kt_events;
}  // namespace

extern "C" {

size_t opb200_coarse_match_workspace_bytes(int n_batch, int L, int S, int C, float thr) {
  return carve_coarse(nullptr, n_batch, L, S, C, thr).bytes;
}

}  // extern "C"

// ext[4] (nullable) = caller-provided operand planes {a_hi, a_lo, b_hi, b_lo} laid out [n_batch*Lp, C] / [n_batch*Sp, C]
static int coarse_match_impl(const float* feat_a, const float* feat_b, const void* const* ext, const uint8_t* mask_a,
                             const uint8_t* mask_b, int mask_fill, int n_batch, int L, int S, int C, float scale, float thr,
                             int border_rm, int border_mode, int h0, int w0, int h1, int w1, const int* valid_hw,
                             int64_t* b_ids, int64_t* i_ids, int64_t* j_ids, float* mconf, int capacity,
                             int* count_and_overflow, float* conf_matrix, void* workspace, size_t workspace_bytes,
                             int precision, int algo, void* stream_) {
  if (int r = check_device()) return r;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  OPB_CHECK_ARG(algo >= 0 && algo <= 3,
                "algo must be 0 (auto: single-pass with the fp16 screening sweep where legal), 1 (two-pass), 2 (single-pass, "
                "three split products) or 3 (screening sweep, register-buffered epilogue variant)");
  OPB_CHECK_ARG(((feat_a && feat_b) || ext) && b_ids && i_ids && j_ids && mconf && count_and_overflow && workspace, "null pointer");
  OPB_CHECK_ARG(n_batch >= 1 && L >= 1 && S >= 1, "empty input");
  OPB_CHECK_ARG(C % BLOCK_K == 0 && C >= BLOCK_K, "feature dim must be a positive multiple of 64");
  OPB_CHECK_ARG(scale > 0.f, "scale must be positive");
  OPB_CHECK_ARG(thr > 0.f && thr < 1.f, "thr must be in (0,1): the candidate-list selection relies on conf > thr > 0");
  OPB_CHECK_ARG(precision == 1 || precision == 3, "precision must be 1 or 3");
  OPB_CHECK_ARG(border_mode == 0 || border_mode == 1, "border_mode must be 0 (OnePosePlus) or 1 (LoFTR)");
  OPB_CHECK_ARG(h1 * w1 == S && (border_mode == 0 || h0 * w0 == L), "grid shape does not match the token count");
  OPB_CHECK_ARG(static_cast<long long>(n_batch) * (L > S ? L : S) < (1ll << 30), "problem too large for 32-bit row ids");
  CoarseWs w = carve_coarse(workspace, n_batch, L, S, C, thr);
  if (w.bytes > workspace_bytes) {
    set_last_error("workspace too small: need " + std::to_string(w.bytes) + " bytes");
    return OPB_ERR_WORKSPACE;
  }
  const bool x3 = precision == 3;
  const int N = n_batch, Lp = w.Lp, Sp = w.Sp;
  // ---- operand planes (pad rows zero so they add nothing and never produce NaN)
  if (ext != nullptr) {  // planes written by the encoder stack; its pad rows are finite and masked by row/col validity
    OPB_CHECK_ARG(ext[0] && ext[2] && (!x3 || (ext[1] && ext[3])), "null operand plane");
    w.a_hi = static_cast<__nv_bfloat16*>(const_cast<void*>(ext[0]));
    w.a_lo = static_cast<__nv_bfloat16*>(const_cast<void*>(ext[1]));
    w.b_hi = static_cast<__nv_bfloat16*>(const_cast<void*>(ext[2]));
    w.b_lo = static_cast<__nv_bfloat16*>(const_cast<void*>(ext[3]));
  }
  if (ext == nullptr && Lp != L) {
    OPB_CHECK_CUDA(cudaMemsetAsync(w.a_hi, 0, static_cast<size_t>(N) * Lp * C * 2, stream));
    if (x3) OPB_CHECK_CUDA(cudaMemsetAsync(w.a_lo, 0, static_cast<size_t>(N) * Lp * C * 2, stream));
  }
  if (ext == nullptr && Sp != S) {
    OPB_CHECK_CUDA(cudaMemsetAsync(w.b_hi, 0, static_cast<size_t>(N) * Sp * C * 2, stream));
    if (x3) OPB_CHECK_CUDA(cudaMemsetAsync(w.b_lo, 0, static_cast<size_t>(N) * Sp * C * 2, stream));
  }
  for (int n = 0; ext == nullptr && n < N; ++n) {
    if (int r = split_planes(feat_a + static_cast<size_t>(n) * L * C, C, 1, nullptr, 0, L, C, w.a_hi + static_cast<size_t>(n) * Lp * C,
                             x3 ? w.a_lo + static_cast<size_t>(n) * Lp * C : nullptr, nullptr, C, stream))
      return r;
    if (int r = split_planes(feat_b + static_cast<size_t>(n) * S * C, C, 1, nullptr, 0, S, C, w.b_hi + static_cast<size_t>(n) * Sp * C,
                             x3 ? w.b_lo + static_cast<size_t>(n) * Sp * C : nullptr, nullptr, C, stream))
      return r;
  }
  OPB_CHECK_CUDA(cudaMemsetAsync(static_cast<uint8_t*>(workspace) + w.zero_begin, 0, w.zero_bytes, stream));
  OPB_CHECK_CUDA(cudaMemsetAsync(w.rowj, 0x7f, static_cast<size_t>(N) * L * 4, stream));  // kNoMatch
  const int m_tiles = N * (Lp / BLOCK_M), n_tiles = Sp / SIM_BN;

  const bool fast = algo != 1 && x3 && C == FAST_C && mask_a == nullptr && mask_b == nullptr && conf_matrix == nullptr;
  const bool screen = fast && (algo == 0 || algo == 3);
  if (fast) {
    SimFastParams fp;
    int r;
    if (screen) {  // fp16 operand planes of the sweep
      const long long va = static_cast<long long>(N) * Lp * C / 8, vb = static_cast<long long>(N) * Sp * C / 8;
      PackSide sa{reinterpret_cast<const uint4*>(w.a_hi), reinterpret_cast<const uint4*>(w.a_lo), reinterpret_cast<uint4*>(w.a_f16), va, L, Lp};
      PackSide sb{reinterpret_cast<const uint4*>(w.b_hi), reinterpret_cast<const uint4*>(w.b_lo), reinterpret_cast<uint4*>(w.b_f16), vb, S, Sp};
      const int ba = static_cast<int>((va + 255) / 256 < 8 * num_sms() ? (va + 255) / 256 : 8 * num_sms());
      // side b: column mean + fp16 plane + largest row norm in one pass; side a: fp16 plane + norm + Jensen bound per row
      colmean_kernel<<<N * MEAN_CHUNKS, 256, 0, stream>>>(w.b_hi, w.b_lo, S, Sp, w.mean_part, w.mean_b, w.mean_done,
                                                          reinterpret_cast<__half*>(w.b_f16), reinterpret_cast<unsigned*>(w.counters + 4) + 1);
      sb.n_vec = 0;
      pack_f16_kernel<<<ba, 256, 0, stream>>>(sa, sb, ba, reinterpret_cast<unsigned*>(w.counters + 4), w.mean_b,
                                              scale * kLog2e, S, w.row_lb);
      count_launch(2);
    }
    if (screen) {  // the data type of a tensor map only matters for its fill value
      if ((r = make_tmap_bf16(&fp.a_hi, w.a_f16, static_cast<uint64_t>(N) * Lp, C, C, BLOCK_M))) return r;
      if ((r = make_tmap_bf16(&fp.b_hi, w.b_f16, static_cast<uint64_t>(N) * Sp, C, C, SIM_BN))) return r;
      fp.a_lo = fp.a_hi, fp.b_lo = fp.b_hi;
    } else {
      if ((r = make_tmap_bf16(&fp.a_hi, w.a_hi, static_cast<uint64_t>(N) * Lp, C, C, BLOCK_M))) return r;
      if ((r = make_tmap_bf16(&fp.a_lo, w.a_lo, static_cast<uint64_t>(N) * Lp, C, C, BLOCK_M))) return r;
      if ((r = make_tmap_bf16(&fp.b_hi, w.b_hi, static_cast<uint64_t>(N) * Sp, C, C, SIM_BN))) return r;
      if ((r = make_tmap_bf16(&fp.b_lo, w.b_lo, static_cast<uint64_t>(N) * Sp, C, C, SIM_BN))) return r;
    }
    // algo 3: the register-buffered variant (sim_sweep_kernel: every epilogue warp on every tile, stage handed back right
    // after the TMEM read); measured slower than sim_fast_kernel<true> on the bench workload (196 vs 181 us at
    // 8 x 8000 x 4096: a warp's 32 x 32 chunk with its per-chunk bookkeeping takes ~2400 clk), kept selectable and
    // parity-tested.  Wider grids run out of row-record slots.
    const bool sweep = screen && algo == 3 && (Sp + SW_BN - 1) / SW_BN <= SW_MAX_NTILES;
    static const int dbg = getenv("OPB200_SIM_DEBUG") ? atoi(getenv("OPB200_SIM_DEBUG")) : 0;
    cudaEvent_t kt0 = nullptr, kt1 = nullptr;
    {
      std::lock_guard<std::mutex> lk(g_kt_mutex);
      if (g_kt_armed && cudaEventCreate(&kt0) == cudaSuccess && cudaEventCreate(&kt1) == cudaSuccess) cudaEventRecord(kt0, stream);
    }
    if (sweep) {
      SimSweepParams sp;
      sp.a = fp.a_hi, sp.b = fp.b_hi;
      sp.N = N, sp.L = L, sp.S = S, sp.Lp = Lp, sp.Sp = Sp;
      sp.n_tiles = (Sp + SW_BN - 1) / SW_BN;
      sp.total_tiles = m_tiles * sp.n_tiles;
      sp.c = scale * kLog2e, sp.thr = thr, sp.max_delta = g_screen_max_delta, sp.row_lb = w.row_lb, sp.norm_sq = reinterpret_cast<const unsigned*>(w.counters + 4), sp.needs_robust = w.counters + 2;
      sp.colpart = reinterpret_cast<float*>(w.colpart), sp.col_ref = w.tile_ref;
      sp.rowpart = reinterpret_cast<float*>(w.rowpart), sp.row_ref = w.row_ref;
      sp.cand = w.cand, sp.cand_count = w.counters, sp.cand_cap = w.cand_cap;
      sp.trace = next_trace_slot(num_sms());
      sp.debug = dbg;
      static std::atomic<unsigned long long> configured_sweep{0};
      if (int rr = opt_in_dyn_smem(sim_sweep_kernel, SW_SMEM, configured_sweep)) return rr;
      OPB_CHECK_CUDA(cudaMemsetAsync(w.row_ref, 0xff, static_cast<size_t>(SW_ROW_SLOTS) * (static_cast<size_t>(N) * Lp / 32) * 4, stream));
      const int grid = sp.total_tiles < num_sms() ? sp.total_tiles : num_sms();
      sim_sweep_kernel<<<grid, 64 + 128 * FAST_GROUPS, SW_SMEM, stream>>>(sp);
      count_launch();
      OPB_CHECK_CUDA(cudaGetLastError());
      if (kt1 != nullptr) {
        cudaEventRecord(kt1, stream);
        std::lock_guard<std::mutex> lk(g_kt_mutex);
        g_kt_events.emplace_back(kt0, kt1);
      }
      const long long nr = static_cast<long long>(N) * Lp, nc = static_cast<long long>(N) * Sp;
      merge_sweep_kernel<<<static_cast<unsigned>(nr / 32 + nc / 32), 256, 0, stream>>>(sp.rowpart, sp.row_ref, sp.colpart, sp.col_ref, N, L, S,
                                                                                       Lp, Sp, w.lse_row, w.lse_col, w.counters + 2);
      count_launch();
    } else {
    fp.N = N, fp.L = L, fp.S = S, fp.Lp = Lp, fp.Sp = Sp;
    fp.n_tiles = screen ? (Sp + FastCfg<true>::BN - 1) / FastCfg<true>::BN : (Sp + FastCfg<false>::BN - 1) / FastCfg<false>::BN;
    fp.total_tiles = m_tiles * fp.n_tiles;
    fp.c = scale * kLog2e;
    fp.thr_lim = 0.99f * thr, fp.thr = thr, fp.max_delta = g_screen_max_delta;
    fp.norm_sq = reinterpret_cast<const unsigned*>(w.counters + 4), fp.needs_robust = w.counters + 2;
    fp.row_lb = screen ? w.row_lb : nullptr;
    fp.lim_scale = g_screen_lim_scale;
    fp.cand_tag = screen ? 2 : 1;
    fp.trace = next_trace_slot(num_sms());
    static const int pair_env = getenv("OPB200_SIM_PAIR") ? atoi(getenv("OPB200_SIM_PAIR")) : 0;
    fp.pair = screen && pair_env != 0 && fp.n_tiles % 2 == 0 ? 1 : 0;
    fp.debug = dbg;
    fp.rowpart = reinterpret_cast<float*>(w.rowpart), fp.colpart = reinterpret_cast<float*>(w.colpart);
    fp.tile_ref = w.tile_ref;
    fp.cand = w.cand, fp.cand_count = w.counters, fp.cand_cap = w.cand_cap;
    static std::atomic<unsigned long long> configured_split{0}, configured_screen{0};
    if (int rr = opt_in_dyn_smem(sim_fast_kernel<false>, FastCfg<false>::SMEM, configured_split)) return rr;
    if (int rr = opt_in_dyn_smem(sim_fast_kernel<true>, FastCfg<true>::SMEM, configured_screen)) return rr;
    const int grid = fp.total_tiles < num_sms() ? fp.total_tiles : num_sms();
    if (screen) sim_fast_kernel<true><<<grid, 64 + 128 * FAST_GROUPS, FastCfg<true>::SMEM, stream>>>(fp);
    else sim_fast_kernel<false><<<grid, 64 + 128 * FAST_GROUPS, FastCfg<false>::SMEM, stream>>>(fp);
    count_launch();
    OPB_CHECK_CUDA(cudaGetLastError());
    if (kt1 != nullptr) {
      cudaEventRecord(kt1, stream);
      std::lock_guard<std::mutex> lk(g_kt_mutex);
      g_kt_events.emplace_back(kt0, kt1);
    }
    const long long nr = static_cast<long long>(N) * Lp, nc = static_cast<long long>(N) * Sp;
    if (nc / 128 >= num_sms())  // 16-byte loads when 128-column blocks still fill the GPU
      merge_fast_kernel<4><<<static_cast<unsigned>((nr + 255) / 256 + nc / 128), 256, 0, stream>>>(fp.rowpart, fp.colpart, w.tile_ref, N, L, S, Lp, Sp,
                                                                                           w.lse_row, w.lse_col, w.counters + 2, SIM_BN);
    else
      merge_fast_kernel<1><<<static_cast<unsigned>((nr + 255) / 256 + nc / 32), 256, 0, stream>>>(fp.rowpart, fp.colpart, w.tile_ref, N, L, S, Lp, Sp,
                                                                                          w.lse_row, w.lse_col, w.counters + 2, SIM_BN);
    count_launch();
    }
  } else {
    fill_sim_row_base<<<(m_tiles + 255) / 256, 256, 0, stream>>>(w.row_base, Lp / BLOCK_M, N, Sp);
    count_launch();
    PlanBuilder pb;
    pb.add_a(w.a_hi, static_cast<uint64_t>(N) * Lp, C, C, 0);
    pb.add_b(w.b_hi, static_cast<uint64_t>(N) * Sp, C, C, 0, SIM_BN);
    if (x3) {
      pb.add_a(w.a_lo, static_cast<uint64_t>(N) * Lp, C, C, 1);
      pb.add_b(w.b_lo, static_cast<uint64_t>(N) * Sp, C, C, 1, SIM_BN);
    }
    if (pb.status) return pb.status;
    const int nkb = C / BLOCK_K;
    if (x3) {
      pb.seg(0, 1, 0, 0, nkb);
      pb.seg(1, 0, 0, 0, nkb);
    }
    pb.seg(0, 0, 0, 0, nkb);
    pb.finish();
    pb.plan.num_n_tiles = n_tiles;
    pb.plan.num_tiles = m_tiles * n_tiles;
    pb.plan.b_rows_per_ntile = SIM_BN;
    pb.plan.b_row_base = w.row_base;

    SimGeom g{N, L, S, Lp, Sp, scale, mask_a, mask_b, mask_fill};
    {
      EpiSimStats<SIM_BN>::Params ep{g, w.rowpart, w.colpart};
      if (int r = launch_gemm<SIM_BN, 4, 2, EpiSimStats<SIM_BN>>(pb.plan, ep, stream)) return r;
    }
    {
      const long long nr = static_cast<long long>(N) * Lp, nc = static_cast<long long>(N) * Sp;
      merge_stats_kernel<<<static_cast<unsigned>((nr + 31) / 32 + (nc + 31) / 32), 256, 0, stream>>>(
          w.rowpart, n_tiles, nr, w.colpart, Lp / BLOCK_M, nc, w.lse_row, w.lse_col, nullptr);
      count_launch();
    }
    {
      const float log2_thr_lo = log2f(thr) - 1e-3f;
      if (conf_matrix != nullptr) {
        EpiSimSelect<SIM_BN, true>::Params ep{g, w.lse_row, w.lse_col, thr, log2_thr_lo, w.cand, w.counters, w.cand_cap, conf_matrix};
        if (int r = launch_gemm<SIM_BN, 5, 4, EpiSimSelect<SIM_BN, true>>(pb.plan, ep, stream)) return r;
      } else {
        EpiSimSelect<SIM_BN, false>::Params ep{g, w.lse_row, w.lse_col, thr, log2_thr_lo, w.cand, w.counters, w.cand_cap, nullptr};
        if (int r = launch_gemm<SIM_BN, 5, 4, EpiSimSelect<SIM_BN, false>>(pb.plan, ep, stream)) return r;
      }
    }
  }
  BorderRule br{border_mode, border_rm, h0, w0, h1, w1, valid_hw};
  const int cb = (w.cand_cap + 255) / 256 < 592 ? (w.cand_cap + 255) / 256 : 592;
  if (screen) {
    cand_rescore_kernel<<<cb, 256, 0, stream>>>(w.cand, w.counters, w.cand_cap, L, Lp, Sp, w.a_hi, w.a_lo, w.b_hi, w.b_lo,
                                                scale * kLog2e, w.lse_row, w.lse_col, w.rowdelta, w.coldelta,
                                                reinterpret_cast<unsigned*>(w.counters + 6), w.counters + 2,
                                                reinterpret_cast<const unsigned*>(w.counters + 4), thr, g_screen_max_delta);
    count_launch();
  }
  cand_max_kernel<<<cb, 256, 0, stream>>>(w.cand, w.counters, w.cand_cap, L, S, w.lse_row, w.lse_col, Lp, Sp, thr, w.rowdelta,
                                          w.coldelta, w.rowmax, w.colmax);
  cand_pick_kernel<<<cb, 256, 0, stream>>>(w.cand, w.counters, w.cand_cap, L, S, w.rowmax, w.colmax, br, w.rowj);
  const int compact_blocks = (N * L + 2047) / 2048 < COMPACT_BLOCKS ? (N * L + 2047) / 2048 : COMPACT_BLOCKS;
  compact_kernel<<<compact_blocks, 256, 0, stream>>>(w.rowj, w.rowmax, N * L, L, reinterpret_cast<long long*>(b_ids),
                                                     reinterpret_cast<long long*>(i_ids), reinterpret_cast<long long*>(j_ids), mconf,
                                                     capacity, count_and_overflow, w.counters, w.cand_cap, count_and_overflow + 1,
                                                     fast ? 1 : 0, w.agg, w.ticket);
  count_launch(3);
  OPB_CHECK_CUDA(cudaGetLastError());
  if (fast && getenv("OPB200_SIM_DEBUG_COUNT") != nullptr) {  // debug: how many candidates the sweep captured
    int hc = 0;
    cudaStreamSynchronize(stream);
    cudaMemcpy(&hc, w.counters, sizeof(hc), cudaMemcpyDeviceToHost);
    fprintf(stderr, "opb200_coarse_match: %d candidates captured (%d slots)\n", hc, w.cand_cap);
  }
  return OPB_OK;
}

extern "C" {

int opb200_coarse_match(const float* feat_a, const float* feat_b, const uint8_t* mask_a, const uint8_t* mask_b,
                        int mask_fill, int n_batch, int L, int S, int C, float scale, float thr, int border_rm,
                        int border_mode, int h0, int w0, int h1, int w1, const int* valid_hw, int64_t* b_ids,
                        int64_t* i_ids, int64_t* j_ids, float* mconf, int capacity, int* count_and_overflow,
                        float* conf_matrix, void* workspace, size_t workspace_bytes, int precision, int algo,
                        void* stream) {
  return coarse_match_impl(feat_a, feat_b, nullptr, mask_a, mask_b, mask_fill, n_batch, L, S, C, scale, thr, border_rm,
                           border_mode, h0, w0, h1, w1, valid_hw, b_ids, i_ids, j_ids, mconf, capacity, count_and_overflow,
                           conf_matrix, workspace, workspace_bytes, precision, algo, stream);
}

int opb200_coarse_match_planes(const void* a_hi, const void* a_lo, const void* b_hi, const void* b_lo, const uint8_t* mask_a,
                               const uint8_t* mask_b, int mask_fill, int n_batch, int L, int S, int C, float scale, float thr,
                               int border_rm, int border_mode, int h0, int w0, int h1, int w1, const int* valid_hw,
                               int64_t* b_ids, int64_t* i_ids, int64_t* j_ids, float* mconf, int capacity,
                               int* count_and_overflow, float* conf_matrix, void* workspace, size_t workspace_bytes,
                               int precision, int algo, void* stream) {
  const void* ext[4] = {a_hi, a_lo, b_hi, b_lo};
  return coarse_match_impl(nullptr, nullptr, ext, mask_a, mask_b, mask_fill, n_batch, L, S, C, scale, thr, border_rm,
                           border_mode, h0, w0, h1, w1, valid_hw, b_ids, i_ids, j_ids, mconf, capacity, count_and_overflow,
                           conf_matrix, workspace, workspace_bytes, precision, algo, stream);
}

float opb200_coarse_screen_max_delta(float max_delta) {
  const float prev = g_screen_max_delta;
  if (max_delta >= 0.f) g_screen_max_delta = max_delta;
  return prev;
}

int opb200_coarse_norms(const void* workspace, int n_batch, int L, int S, int C, float thr, float* norm_a, float* norm_b,
                        int* candidates, float* observed, void* stream) {
  OPB_CHECK_ARG(workspace && norm_a && norm_b && candidates && observed, "null pointer");
  CoarseWs w = carve_coarse(const_cast<void*>(workspace), n_batch, L, S, C, thr);
  int words[8];
  OPB_CHECK_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  OPB_CHECK_CUDA(cudaMemcpy(words, w.counters, sizeof(words), cudaMemcpyDeviceToHost));
  unsigned ua, ub;
  std::memcpy(&ua, &words[4], 4);
  std::memcpy(&ub, &words[5], 4);
  float fa, fb;
  std::memcpy(&fa, &ua, 4);
  std::memcpy(&fb, &ub, 4);
  *norm_a = std::sqrt(fa), *norm_b = std::sqrt(fb), *candidates = words[0];
  std::memcpy(observed, &words[6], 4);
  return OPB_OK;
}

int opb200_coarse_kernel_timing(int enable) {
  std::lock_guard<std::mutex> lk(g_kt_mutex);
  g_kt_armed = enable != 0;
  return OPB_OK;
}

int opb200_coarse_kernel_ms(double* total_ms, long long* launches) {
  OPB_CHECK_ARG(total_ms && launches, "null pointer");
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> evs;
  {
    std::lock_guard<std::mutex> lk(g_kt_mutex);
    evs.swap(g_kt_events);
  }
  double tot = 0.0;
  for (auto& e : evs) {
    float ms = 0.f;
    OPB_CHECK_CUDA(cudaEventSynchronize(e.second));
    OPB_CHECK_CUDA(cudaEventElapsedTime(&ms, e.first, e.second));
    tot += ms;
    cudaEventDestroy(e.first);
    cudaEventDestroy(e.second);
  }
  *total_ms = tot;
  *launches = static_cast<long long>(evs.size());
  return OPB_OK;
}

}  // extern "C"
