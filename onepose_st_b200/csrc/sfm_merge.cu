// Keypoint-Free SfM match post-processing on the device: per-image merge of the LoFTR match rows of all image pairs.
//
// Replaces (src/KeypointFreeSfM/coarse_match)
//   Match2Pts2D.__getitem__ + agg_groupby_2d      utils.py:5-60
//   points2D_worker / update_matches / transform_points2D   coarse_match_worker.py:87-184
// The reference walks Python dicts of tuples per image and per pair (the serial tail of the 2k-pair configuration).  Here
// the gathered [sum M, 5] = (u0, v0, u1, v1, conf) rows are consumed as they lie in HBM:
//   sfm_keys_kernel        entry e = 2 * row + side  ->  64-bit key (image, x, y) with x, y = int(u), int(v) (truncation,
//                          as numpy's astype(int)); entry order = (pair, row) order, which is the order in which the
//                          reference concatenates an image's keypoints
//   [stable sort of the keys, segment heads, prefix sum: library primitives driven by sfm_merge.py]
//   sfm_group_sum_kernel   one thread per group of equal keys: the confidences added in fp64 in entry order -- exactly
//                          what np.bincount(weights=) does, so that the ranking by score (and its ties) is bit-identical
// HBM-bound integer work: 8 + 8 bytes per entry for the keys, one pass each.
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

__global__ void __launch_bounds__(256) sfm_keys_kernel(const float* __restrict__ rows, long long n_rows,
                                                       const int* __restrict__ pair_of_row, const int* __restrict__ pair_images,
                                                       long long* __restrict__ keys, int* __restrict__ bad) {
  const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= 2 * n_rows) return;
  const long long r = e >> 1;
  const int side = static_cast<int>(e & 1);
  const float* row = rows + r * 5;
  const int img = pair_images[2 * pair_of_row[r] + side];
  const float u = row[2 * side], v = row[2 * side + 1];
  const long long x = static_cast<long long>(u), y = static_cast<long long>(v);  // truncation toward zero
  if (!(u == u) || !(v == v) || x < 0 || y < 0 || x >= (1 << 20) || y >= (1 << 20) || img < 0 || img >= (1 << 23)) {
    atomicExch(bad, 1);
    keys[e] = 0x7fffffffffffffffLL;
    return;
  }
  keys[e] = (static_cast<long long>(img) << 40) | (x << 20) | y;
}

// group g owns the sorted entries start[g] .. start[g + 1] - 1 (order[] = entry ids in sorted order; the sort is stable,
// so inside a group they ascend); sums[g] = sum of their confidences, added one by one in fp64
__global__ void __launch_bounds__(256) sfm_group_sum_kernel(const float* __restrict__ rows, const long long* __restrict__ order,
                                                            const long long* __restrict__ start, long long n_groups,
                                                            double* __restrict__ sums) {
  const long long g = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (g >= n_groups) return;
  double s = 0.0;
  for (long long k = start[g]; k < start[g + 1]; ++k) s += static_cast<double>(rows[(order[k] >> 1) * 5 + 4]);
  sums[g] = s;
}

}  // namespace opb

using namespace opb;

extern "C" {

int opb200_sfm_keys(const float* rows, long long n_rows, const int* pair_of_row, const int* pair_images, long long* keys,
                    int* bad_flag, void* stream) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(rows && pair_of_row && pair_images && keys && bad_flag, "null pointer");
  OPB_CHECK_ARG(n_rows >= 1, "no rows");
  const long long n = 2 * n_rows;
  sfm_keys_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(rows, n_rows, pair_of_row,
                                                                                                        pair_images, keys, bad_flag);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

int opb200_sfm_group_sums(const float* rows, const long long* order, const long long* group_start, long long n_groups,
                          double* sums, void* stream) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(rows && order && group_start && sums, "null pointer");
  OPB_CHECK_ARG(n_groups >= 1, "no groups");
  sfm_group_sum_kernel<<<static_cast<unsigned>((n_groups + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      rows, order, group_start, n_groups, sums);
  count_launch();
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

}  // extern "C"
