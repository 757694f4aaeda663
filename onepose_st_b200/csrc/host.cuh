// Host-side helpers shared by the C-ABI translation units: error reporting, TMA descriptor encoding,
// GEMM plan construction and launch.
#pragma once
#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>
#include <cuda.h>
#include <cuda_runtime.h>
#include "gemm.cuh"

namespace opb {

// ---- status codes returned by every C-ABI entry point
enum : int {
  OPB_OK = 0,
  OPB_ERR_ARG = -1,        // bad argument (null pointer, unsupported size, misalignment)
  OPB_ERR_CUDA = -2,       // a CUDA runtime/driver call failed (see opb200_last_error)
  OPB_ERR_WORKSPACE = -3,  // caller-provided workspace too small
  OPB_ERR_DEVICE = -4,     // not an sm_100 device
};

void set_last_error(const std::string& msg);

#define OPB_CHECK_ARG(cond, msg)                                             \
  do {                                                                       \
    if (!(cond)) {                                                           \
      ::opb::set_last_error(std::string("argument error: ") + (msg));        \
      return ::opb::OPB_ERR_ARG;                                             \
    }                                                                        \
  } while (0)

#define OPB_CHECK_CUDA(expr)                                                                  \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      ::opb::set_last_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e));       \
      return ::opb::OPB_ERR_CUDA;                                                             \
    }                                                                                         \
  } while (0)

void count_launch(int n = 1);  // bumps the counter opb200_launch_count() reports
int check_device();  // OPB_OK iff the current device is compute capability 10.x
int num_sms();
long long* next_trace_slot(int grid);  // debug tracing of GEMM launches (null unless opb200_debug_trace armed it)

// Encode a 2-D bf16 row-major [rows, cols] tensor (row pitch in elements) as a TMA descriptor whose box
// is [box_rows x 64] with 128-byte swizzle.  Out-of-bounds rows/cols are zero-filled on load.
int make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t pitch_elems,
                   uint32_t box_rows);

// General 16-bit (bf16 / fp16) tiled descriptor of rank 2..5 with 128- or 64-byte swizzle (innermost box extent 64 or
// 32 elements = one swizzle row); gdim / box / elem_strides innermost first, gstride_bytes for dims 1..rank-1.
int encode_tmap_16bit(CUtensorMap* out, bool fp16, int rank, const void* base, const uint64_t* gdim,
                      const uint64_t* gstride_bytes, const uint32_t* box, const uint32_t* elem_strides,
                      int swizzle_bytes = 128);

struct PlanBuilder {
  GemmPlan plan;
  int status = OPB_OK;
  bool fused3;  // descriptors for the fused split-precision mainloop: [rows x 32] boxes, 64-byte swizzle
  explicit PlanBuilder(bool fused3_ = false) : fused3(fused3_) { std::memset(&plan, 0, sizeof(plan)); }
  int make(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t pitch, uint32_t box_rows) {
    if (!fused3) return make_tmap_bf16(out, base, rows, cols, pitch, box_rows);
    const uint64_t gdim[2] = {cols, rows};
    const uint64_t gstr[1] = {pitch * 2};
    const uint32_t box[2] = {32, box_rows};
    const uint32_t estr[2] = {1, 1};
    return encode_tmap_16bit(out, false, 2, base, gdim, gstr, box, estr, 64);
  }
  int add_a(const void* base, uint64_t rows, uint64_t cols, uint64_t pitch, int idx) {
    int r = make(&plan.a_maps[idx], base, rows, cols, pitch, BLOCK_M);
    if (r != OPB_OK) status = r;
    return idx;
  }
  int add_b(const void* base, uint64_t rows, uint64_t cols, uint64_t pitch, int idx, int bn) {
    int r = make(&plan.b_maps[idx], base, rows, cols, pitch, bn);
    if (r != OPB_OK) status = r;
    return idx;
  }
  // fused split-precision segment: one pass over K computing A_hi*B_hi + A_lo*B_hi + A_hi*B_lo per 32-wide k-block
  void seg3(int a_hi, int a_lo, int b_hi, int b_lo, int a_k0, int b_k0, int nkb32) {
    KSeg& s = plan.segs[plan.nsegs++];
    s.a_map = a_hi;
    s.a_map2 = a_lo;
    s.b_map = b_hi;
    s.b_map2 = b_lo;
    s.a_k0 = a_k0;
    s.b_k0 = b_k0;
    s.nkb = nkb32;
    plan.total_kb += nkb32;
  }
  void seg(int a_map, int b_map, int a_k0, int b_k0, int nkb) {
    KSeg& s = plan.segs[plan.nsegs++];
    s.a_map = a_map;
    s.b_map = b_map;
    s.a_k0 = a_k0;
    s.b_k0 = b_k0;
    s.nkb = nkb;
    plan.total_kb += nkb;
  }
  // fill unused descriptor slots with a valid one so the kernel's prefetch never sees garbage
  void finish() {
    for (int i = 1; i < kNumMaps; ++i) {
      bool a_used = false, b_used = false;
      for (int s = 0; s < plan.nsegs; ++s) {
        a_used |= plan.segs[s].a_map == i || (fused3 && plan.segs[s].a_map2 == i);
        b_used |= plan.segs[s].b_map == i || (fused3 && plan.segs[s].b_map2 == i);
      }
      if (!a_used) plan.a_maps[i] = plan.a_maps[0];
      if (!b_used) plan.b_maps[i] = plan.b_maps[0];
    }
  }
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute of a kernel: opt in once per (kernel, device).
// ``done`` is the call site's bitmask of devices already configured (one static per kernel instantiation).
template <typename Kern>
inline int opt_in_dyn_smem(Kern kern, int bytes, std::atomic<unsigned long long>& done) {
  int dev = 0;
  OPB_CHECK_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ull << (dev & 63);
  if (!(done.load(std::memory_order_acquire) & bit)) {
    OPB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    done.fetch_or(bit, std::memory_order_release);
  }
  return OPB_OK;
}

template <int BN, int STAGES, int GROUPS, class Epi, int SPLIT = 1, bool F3 = false>
int launch_gemm(const GemmPlan& plan, const typename Epi::Params& ep, cudaStream_t stream, int max_ctas = 0) {
  if (plan.num_tiles <= 0) return OPB_OK;
  auto kern = gemm_kernel<BN, STAGES, GROUPS, SPLIT, Epi, F3>;
  constexpr int smem = gemm_smem_bytes(BN, STAGES, GROUPS * Epi::kSmemBytes);
  static_assert(smem <= 227 * 1024, "shared memory budget exceeded");
  static std::atomic<unsigned long long> configured{0};  // per instantiation
  if (int r = opt_in_dyn_smem(kern, smem, configured)) return r;
  int grid = num_sms();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  // the fewest CTAs that finish in the same number of tile waves: SMs a launch cannot use anyway stay free for
  // kernels of the other streams in flight (pipeline.StreamPipeline)
  const int waves = (plan.num_tiles + grid - 1) / grid;
  grid = (plan.num_tiles + waves - 1) / waves;
  if (long long* tr = next_trace_slot(grid)) {
    GemmPlan traced = plan;
    traced.trace = tr;
    kern<<<grid, 64 + 128 * GROUPS * SPLIT, smem, stream>>>(traced, ep);
  } else {
    kern<<<grid, 64 + 128 * GROUPS * SPLIT, smem, stream>>>(plan, ep);
  }
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

}  // namespace opb
