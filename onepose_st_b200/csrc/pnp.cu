// RANSAC-PnP on the device: the pose step that follows the 2D-3D matcher, consuming its match rows where they lie.
//
// Replaces  ransac_PnP   src/utils/metric_utils.py:121-209   (call sites inference.py:181-189, 328-336)
//   reference: OpenCV solvePnPRansac (EPnP model, reprojectionError 5 px, iterationsCount 10000) or pycolmap's
//   absolute_pose_estimation (P3P RANSAC + non-linear refinement) on the host, per frame, after a device -> host copy
//   of the matches.  Here every frame of a batch is solved in three launches, the matches never leave HBM:
//     pnp_hypotheses_kernel   one thread per hypothesis: 4 indices from a counter-based hash, Grunert's P3P in fp64 on
//                             three of them (quartic by Ferrari + Newton), the fourth point picks among <= 4 poses
//     pnp_score_kernel        one warp per hypothesis over the frame's points (staged in shared memory, shared by the
//                             block's hypotheses): number of points with squared reprojection error < thr^2
//     pnp_refine_kernel       one block per frame: arg-max score (ties: lowest index), then twice { inlier set,
//                             10 Gauss-Newton steps on the inliers' reprojection error in fp64 } (left-multiplied
//                             so(3) x R^3 increments; 27 sums per step reduced over the block, 6x6 Cholesky)
//   oracle/pnp_ref.py restates the same scheme in numpy; OpenCV is the end-to-end oracle (tests/test_gpu_pnp.py).
#include <cmath>
#include "host.cuh"
#include "../../include/opb200.h"

namespace opb {

__device__ __forceinline__ unsigned hash32(unsigned x) {  // lowbias32
  x ^= x >> 16;
  x *= 0x7FEB352Du;
  x ^= x >> 15;
  x *= 0x846CA68Bu;
  x ^= x >> 16;
  return x;
}

// Real roots of a4 x^4 + a3 x^3 + a2 x^2 + a1 x + a0 (Ferrari through one positive root of the resolvent cubic; every
// root polished by two Newton steps on the original polynomial).  Returns the number of roots written.
__device__ int solve_quartic(double a4, double a3, double a2, double a1, double a0, double* x) {
  const double big = fmax(fmax(fabs(a3), fabs(a2)), fmax(fabs(a1), fmax(fabs(a0), 1e-300)));
  if (fabs(a4) < 1e-14 * big) return 0;
  const double b = a3 / a4, c = a2 / a4, d = a1 / a4, e = a0 / a4;
  const double p = c - 3.0 * b * b / 8.0;
  const double q = d - b * c / 2.0 + b * b * b / 8.0;
  const double r = e - b * d / 4.0 + b * b * c / 16.0 - 3.0 * b * b * b * b / 256.0;
  double y[4];
  int n = 0;
  if (fabs(q) < 1e-14) {
    const double disc = p * p - 4.0 * r;
    if (disc >= 0.0) {
      const double sd = sqrt(disc);
      const double z0 = (-p + sd) / 2.0, z1 = (-p - sd) / 2.0;
      if (z0 >= 0.0) y[n++] = sqrt(z0), y[n++] = -sqrt(z0);
      if (z1 >= 0.0) y[n++] = sqrt(z1), y[n++] = -sqrt(z1);
    }
  } else {
    const double A = p, B = p * p / 4.0 - r, C = -q * q / 8.0;
    const double P = B - A * A / 3.0;
    const double Q = 2.0 * A * A * A / 27.0 - A * B / 3.0 + C;
    const double D = Q * Q / 4.0 + P * P * P / 27.0;
    double t;
    if (D >= 0.0) {
      const double sD = sqrt(D);
      t = cbrt(-Q / 2.0 + sD) + cbrt(-Q / 2.0 - sD);
    } else {
      const double rho = sqrt(-P / 3.0);
      const double th = acos(fmin(1.0, fmax(-1.0, -Q / (2.0 * rho * rho * rho))));
      t = 2.0 * rho * cos(th / 3.0);
    }
    const double m = t - A / 3.0;
    if (!(m > 0.0)) return 0;
    const double s = sqrt(2.0 * m);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const double sg = k == 0 ? 1.0 : -1.0;
      const double disc = -(2.0 * p + 2.0 * m + sg * 2.0 * q / s);
      if (disc >= 0.0) {
        const double sd = sqrt(disc);
        y[n++] = (sg * s + sd) / 2.0;
        y[n++] = (sg * s - sd) / 2.0;
      }
    }
  }
  for (int k = 0; k < n; ++k) {
    double v = y[k] - b / 4.0;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const double f = (((a4 * v + a3) * v + a2) * v + a1) * v + a0;
      const double df = ((4.0 * a4 * v + 3.0 * a3) * v + 2.0 * a2) * v + a1;
      if (df != 0.0) v -= f / df;
    }
    x[k] = v;
  }
  return n;
}

struct Vec3 {
  double x, y, z;
};
__device__ __forceinline__ Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ Vec3 operator*(double s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ double dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ Vec3 cross(Vec3 a, Vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ Vec3 unit(Vec3 a) { return (1.0 / sqrt(dot(a, a))) * a; }

// orthonormal frame of a triangle: columns e1 = (X1 - X0)^, e3 = (e1 x (X2 - X0))^, e2 = e3 x e1
__device__ __forceinline__ void tri_frame(const Vec3* X, Vec3* e) {
  e[0] = unit(X[1] - X[0]);
  e[2] = unit(cross(e[0], X[2] - X[0]));
  e[1] = cross(e[2], e[0]);
}

struct Intr {
  double fx, fy, cx, cy;
};

// squared reprojection error of one point under (R row-major, t); +inf behind the camera
__device__ __forceinline__ float reproj_sq_f(const float* R, const float* t, float fx, float fy, float cx, float cy, float X,
                                             float Y, float Z, float u, float v) {
  const float xc = R[0] * X + R[1] * Y + R[2] * Z + t[0];
  const float yc = R[3] * X + R[4] * Y + R[5] * Z + t[1];
  const float zc = R[6] * X + R[7] * Y + R[8] * Z + t[2];
  if (!(zc > 1e-9f)) return INFINITY;
  const float iz = 1.f / zc;
  const float du = fx * xc * iz + cx - u, dv = fy * yc * iz + cy - v;
  return du * du + dv * dv;
}

// hyp [n_frames][H][12] float (R row-major, t); an invalid hypothesis has t.z = NaN
__global__ void __launch_bounds__(128) pnp_hypotheses_kernel(const float* __restrict__ p3, int s3, const float* __restrict__ p2, int s2,
                                                             const int* __restrict__ offsets, const double* __restrict__ intr,
                                                             int H, unsigned seed, float* __restrict__ hyp) {
  const int frame = blockIdx.y;
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= H) return;
  float* out = hyp + (static_cast<long long>(frame) * H + h) * 12;
  const int o0 = offsets[frame], m = offsets[frame + 1] - o0;
  out[11] = NAN;
  if (m < 4) return;
  const Intr K = {intr[4 * frame], intr[4 * frame + 1], intr[4 * frame + 2], intr[4 * frame + 3]};
  int idx[4];
  const unsigned ctr = seed + static_cast<unsigned>(h) * 32u;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int pick = -1;
    for (int a = 0; a < 8 && pick < 0; ++a) {
      const int c = static_cast<int>(hash32(ctr + k * 8 + a) % static_cast<unsigned>(m));
      bool dup = false;
      for (int j = 0; j < k; ++j) dup |= idx[j] == c;
      if (!dup) pick = c;
    }
    for (int c = 0; pick < 0 && c < m; ++c) {
      bool dup = false;
      for (int j = 0; j < k; ++j) dup |= idx[j] == c;
      if (!dup) pick = c;
    }
    idx[k] = pick;
  }
  Vec3 P[3], f[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float* a = p3 + static_cast<long long>(o0 + idx[k]) * s3;
    const float* b = p2 + static_cast<long long>(o0 + idx[k]) * s2;
    P[k] = {static_cast<double>(a[0]), static_cast<double>(a[1]), static_cast<double>(a[2])};
    f[k] = unit(Vec3{(static_cast<double>(b[0]) - K.cx) / K.fx, (static_cast<double>(b[1]) - K.cy) / K.fy, 1.0});
  }
  const double a2 = dot(P[1] - P[2], P[1] - P[2]), b2 = dot(P[0] - P[2], P[0] - P[2]), c2 = dot(P[0] - P[1], P[0] - P[1]);
  if (fmin(a2, fmin(b2, c2)) < 1e-18) return;
  const double ca = dot(f[1], f[2]), cb = dot(f[0], f[2]), cg = dot(f[0], f[1]);
  const double k1 = (a2 - c2) / b2, k2 = (a2 + c2) / b2;
  const double A4 = (k1 - 1.0) * (k1 - 1.0) - 4.0 * c2 / b2 * ca * ca;
  const double A3 = 4.0 * (k1 * (1.0 - k1) * cb - (1.0 - k2) * ca * cg + 2.0 * c2 / b2 * ca * ca * cb);
  const double A2 = 2.0 * (k1 * k1 - 1.0 + 2.0 * k1 * k1 * cb * cb + 2.0 * (b2 - c2) / b2 * ca * ca - 4.0 * k2 * ca * cb * cg +
                           2.0 * (b2 - a2) / b2 * cg * cg);
  const double A1 = 4.0 * (-k1 * (1.0 + k1) * cb + 2.0 * a2 / b2 * cg * cg * cb - (1.0 - k2) * ca * cg);
  const double A0 = (1.0 + k1) * (1.0 + k1) - 4.0 * a2 / b2 * cg * cg;
  double roots[4];
  const int nr = solve_quartic(A4, A3, A2, A1, A0, roots);
  const float* a4p = p3 + static_cast<long long>(o0 + idx[3]) * s3;
  const float* b4p = p2 + static_cast<long long>(o0 + idx[3]) * s2;
  const Vec3 P4 = {static_cast<double>(a4p[0]), static_cast<double>(a4p[1]), static_cast<double>(a4p[2])};
  const double u4 = b4p[0], v4 = b4p[1];
  Vec3 eP[3];
  tri_frame(P, eP);
  double best_e = INFINITY;
  double bR[9], bt[3];
  for (int r = 0; r < nr; ++r) {
    const double v = roots[r];
    if (!(v > 0.0) || !isfinite(v)) continue;
    const double den = 2.0 * (cg - v * ca);
    if (fabs(den) < 1e-12) continue;
    const double u = ((k1 - 1.0) * v * v - 2.0 * k1 * cb * v + 1.0 + k1) / den;
    if (!(u > 0.0)) continue;
    const double dd = 1.0 + v * v - 2.0 * v * cb;
    if (!(dd > 0.0)) continue;
    const double s1 = sqrt(b2 / dd);
    Vec3 Q[3] = {s1 * f[0], (u * s1) * f[1], (v * s1) * f[2]};
    if (fabs(dot(Q[1] - Q[2], Q[1] - Q[2]) - a2) > 1e-6 * a2 || fabs(dot(Q[0] - Q[1], Q[0] - Q[1]) - c2) > 1e-6 * c2) continue;
    Vec3 eQ[3];
    tri_frame(Q, eQ);
    // R = [eQ] [eP]^T
    double R[9];
    const double q[3][3] = {{eQ[0].x, eQ[1].x, eQ[2].x}, {eQ[0].y, eQ[1].y, eQ[2].y}, {eQ[0].z, eQ[1].z, eQ[2].z}};
    const double pm[3][3] = {{eP[0].x, eP[1].x, eP[2].x}, {eP[0].y, eP[1].y, eP[2].y}, {eP[0].z, eP[1].z, eP[2].z}};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) R[3 * i + j] = q[i][0] * pm[j][0] + q[i][1] * pm[j][1] + q[i][2] * pm[j][2];
    const double t0 = Q[0].x - (R[0] * P[0].x + R[1] * P[0].y + R[2] * P[0].z);
    const double t1 = Q[0].y - (R[3] * P[0].x + R[4] * P[0].y + R[5] * P[0].z);
    const double t2 = Q[0].z - (R[6] * P[0].x + R[7] * P[0].y + R[8] * P[0].z);
    const double xc = R[0] * P4.x + R[1] * P4.y + R[2] * P4.z + t0;
    const double yc = R[3] * P4.x + R[4] * P4.y + R[5] * P4.z + t1;
    const double zc = R[6] * P4.x + R[7] * P4.y + R[8] * P4.z + t2;
    if (!(zc > 1e-9)) continue;
    const double du = K.fx * xc / zc + K.cx - u4, dv = K.fy * yc / zc + K.cy - v4;
    const double e = du * du + dv * dv;
    if (e < best_e) {
      best_e = e;
#pragma unroll
      for (int i = 0; i < 9; ++i) bR[i] = R[i];
      bt[0] = t0, bt[1] = t1, bt[2] = t2;
    }
  }
  if (best_e < INFINITY) {
#pragma unroll
    for (int i = 0; i < 9; ++i) out[i] = static_cast<float>(bR[i]);
    out[9] = static_cast<float>(bt[0]), out[10] = static_cast<float>(bt[1]), out[11] = static_cast<float>(bt[2]);
  }
}

// score[frame][h] = number of the frame's points within thr of hypothesis h.  Block = 8 warps x PNP_HPW hypotheses each; the
// frame's points pass through shared memory in chunks (structure of arrays), read by every hypothesis of the block.
constexpr int PNP_CHUNK = 2048;
constexpr int PNP_HPW = 4;
__global__ void __launch_bounds__(256) pnp_score_kernel(const float* __restrict__ p3, int s3, const float* __restrict__ p2, int s2,
                                                        const int* __restrict__ offsets, const double* __restrict__ intr, int H,
                                                        float thr2, const float* __restrict__ hyp, int* __restrict__ score) {
  __shared__ float sx[PNP_CHUNK], sy[PNP_CHUNK], sz[PNP_CHUNK], su[PNP_CHUNK], sv[PNP_CHUNK];
  const int frame = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int o0 = offsets[frame], m = offsets[frame + 1] - o0;
  const float fx = static_cast<float>(intr[4 * frame]), fy = static_cast<float>(intr[4 * frame + 1]);
  const float cx = static_cast<float>(intr[4 * frame + 2]), cy = static_cast<float>(intr[4 * frame + 3]);
  const int h0 = (blockIdx.x * 8 + warp) * PNP_HPW;
  float R[PNP_HPW][9], t[PNP_HPW][3];
  int cnt[PNP_HPW];
#pragma unroll
  for (int k = 0; k < PNP_HPW; ++k) {
    cnt[k] = 0;
    const int h = h0 + k;
    const float* src = hyp + (static_cast<long long>(frame) * H + (h < H ? h : 0)) * 12;
#pragma unroll
    for (int i = 0; i < 9; ++i) R[k][i] = src[i];
    t[k][0] = src[9], t[k][1] = src[10], t[k][2] = (h < H) ? src[11] : NAN;
  }
  for (int c0 = 0; c0 < m; c0 += PNP_CHUNK) {
    const int cn = min(PNP_CHUNK, m - c0);
    __syncthreads();
    for (int i = threadIdx.x; i < cn; i += blockDim.x) {
      const float* a = p3 + static_cast<long long>(o0 + c0 + i) * s3;
      const float* b = p2 + static_cast<long long>(o0 + c0 + i) * s2;
      sx[i] = a[0], sy[i] = a[1], sz[i] = a[2], su[i] = b[0], sv[i] = b[1];
    }
    __syncthreads();
    for (int i = lane; i < cn; i += 32) {
      const float X = sx[i], Y = sy[i], Z = sz[i], u = su[i], v = sv[i];
#pragma unroll
      for (int k = 0; k < PNP_HPW; ++k) cnt[k] += reproj_sq_f(R[k], t[k], fx, fy, cx, cy, X, Y, Z, u, v) < thr2 ? 1 : 0;
    }
  }
#pragma unroll
  for (int k = 0; k < PNP_HPW; ++k) {
    const int tot = __reduce_add_sync(0xffffffffu, cnt[k]);
    if (lane == 0 && h0 + k < H) score[static_cast<long long>(frame) * H + h0 + k] = (t[k][2] == t[k][2]) ? tot : -1;
  }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One block per frame: best hypothesis, then twice { inliers, Gauss-Newton on them }.  Outputs: pose [n_frames][12]
// double (rows of [R | t]), inlier mask [sum M] uint8, inlier count / best RANSAC score [n_frames][2] int.
constexpr int PNP_GN_ITERS = 10;
__global__ void __launch_bounds__(256) pnp_refine_kernel(const float* __restrict__ p3, int s3, const float* __restrict__ p2, int s2,
                                                         const int* __restrict__ offsets, const double* __restrict__ intr, int H,
                                                         float thr2, const float* __restrict__ hyp, const int* __restrict__ score,
                                                         double* __restrict__ pose, unsigned char* __restrict__ inlier,
                                                         int* __restrict__ counts) {
  __shared__ unsigned long long red_key[8];
  __shared__ double red[8][27];
  __shared__ double sR[9], st[3], upd[6];
  __shared__ int s_cnt[8], s_ok;
  const int frame = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int o0 = offsets[frame], m = offsets[frame + 1] - o0;
  const double fx = intr[4 * frame], fy = intr[4 * frame + 1], cx = intr[4 * frame + 2], cy = intr[4 * frame + 3];
  // ---- arg-max of the scores; ties go to the lowest hypothesis index
  unsigned long long key = 0ull;
  for (int h = tid; h < H; h += blockDim.x) {
    const int sc = score[static_cast<long long>(frame) * H + h];
    if (sc >= 0) {
      const unsigned long long k = (static_cast<unsigned long long>(sc + 1) << 32) | static_cast<unsigned>(0xFFFFFFFFu - h);
      key = k > key ? k : key;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
    key = other > key ? other : key;
  }
  if (lane == 0) red_key[warp] = key;
  __syncthreads();
  if (tid == 0) {
    unsigned long long k = 0ull;
    for (int w = 0; w < 8; ++w) k = red_key[w] > k ? red_key[w] : k;
    const int best_score = static_cast<int>(k >> 32) - 1;
    s_ok = (m >= 4 && best_score >= 4) ? 1 : 0;
    counts[2 * frame + 1] = best_score < 0 ? 0 : best_score;
    if (s_ok) {
      const int h = static_cast<int>(0xFFFFFFFFu - static_cast<unsigned>(k & 0xFFFFFFFFull));
      const float* src = hyp + (static_cast<long long>(frame) * H + h) * 12;
      for (int i = 0; i < 9; ++i) sR[i] = src[i];
      for (int i = 0; i < 3; ++i) st[i] = src[9 + i];
    }
  }
  __syncthreads();
  if (!s_ok) {  // the reference returns the identity pose and no inliers when the solver fails
    for (int i = tid; i < m; i += blockDim.x) inlier[o0 + i] = 0;
    if (tid < 12) pose[12 * frame + tid] = (tid == 0 || tid == 5 || tid == 10) ? 1.0 : 0.0;
    if (tid == 0) counts[2 * frame] = 0;
    return;
  }
  auto mark_inliers = [&]() -> int {  // inlier[] <- within thr of (sR, st); returns the count (to every thread)
    int c = 0;
    for (int i = tid; i < m; i += blockDim.x) {
      const float* a = p3 + static_cast<long long>(o0 + i) * s3;
      const float* b = p2 + static_cast<long long>(o0 + i) * s2;
      const double X = a[0], Y = a[1], Z = a[2];
      const double xc = sR[0] * X + sR[1] * Y + sR[2] * Z + st[0];
      const double yc = sR[3] * X + sR[4] * Y + sR[5] * Z + st[1];
      const double zc = sR[6] * X + sR[7] * Y + sR[8] * Z + st[2];
      bool in = false;
      if (zc > 1e-9) {
        const double du = fx * xc / zc + cx - b[0], dv = fy * yc / zc + cy - b[1];
        in = du * du + dv * dv < static_cast<double>(thr2);
      }
      inlier[o0 + i] = in ? 1 : 0;
      c += in ? 1 : 0;
    }
    c = __reduce_add_sync(0xffffffffu, c);
    __syncthreads();
    if (lane == 0) s_cnt[warp] = c;
    __syncthreads();
    int tot = 0;
    for (int w = 0; w < 8; ++w) tot += s_cnt[w];
    return tot;
  };
  int n_in = mark_inliers();
  for (int round = 0; round < 2 && n_in >= 4; ++round) {
    for (int it = 0; it < PNP_GN_ITERS; ++it) {
      double acc[27];
#pragma unroll
      for (int k = 0; k < 27; ++k) acc[k] = 0.0;
      for (int i = tid; i < m; i += blockDim.x) {
        if (!inlier[o0 + i]) continue;
        const float* a = p3 + static_cast<long long>(o0 + i) * s3;
        const float* b = p2 + static_cast<long long>(o0 + i) * s2;
        const double X0 = a[0], Y0 = a[1], Z0 = a[2];
        const double X = sR[0] * X0 + sR[1] * Y0 + sR[2] * Z0 + st[0];
        const double Y = sR[3] * X0 + sR[4] * Y0 + sR[5] * Z0 + st[1];
        const double Z = sR[6] * X0 + sR[7] * Y0 + sR[8] * Z0 + st[2];
        const double iz = 1.0 / Z;
        const double ru = b[0] - (fx * X * iz + cx), rv = b[1] - (fy * Y * iz + cy);
        // rows of the 2 x 6 Jacobian: d proj / d (w, dt), X' = Exp(w) X + dt
        const double ju[3] = {fx * iz, 0.0, -fx * X * iz * iz}, jv[3] = {0.0, fy * iz, -fy * Y * iz * iz};
        double Ju[6], Jv[6];
        Ju[0] = ju[2] * Y - ju[1] * Z, Ju[1] = ju[0] * Z - ju[2] * X, Ju[2] = ju[1] * X - ju[0] * Y;
        Jv[0] = jv[2] * Y - jv[1] * Z, Jv[1] = jv[0] * Z - jv[2] * X, Jv[2] = jv[1] * X - jv[0] * Y;
#pragma unroll
        for (int k = 0; k < 3; ++k) Ju[3 + k] = ju[k], Jv[3 + k] = jv[k];
        int q = 0;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
#pragma unroll
          for (int c = r; c < 6; ++c) acc[q++] += Ju[r] * Ju[c] + Jv[r] * Jv[c];
          acc[21 + r] += Ju[r] * ru + Jv[r] * rv;
        }
      }
#pragma unroll
      for (int k = 0; k < 27; ++k) acc[k] = warp_sum(acc[k]);
      __syncthreads();
      if (lane == 0)
#pragma unroll
        for (int k = 0; k < 27; ++k) red[warp][k] = acc[k];
      __syncthreads();
      if (tid == 0) {
        double A[6][6], g[6];
        int q = 0;
        for (int r = 0; r < 6; ++r)
          for (int c = r; c < 6; ++c) {
            double s = 0.0;
            for (int w = 0; w < 8; ++w) s += red[w][q];
            A[r][c] = A[c][r] = s;
            ++q;
          }
        for (int r = 0; r < 6; ++r) {
          double s = 0.0;
          for (int w = 0; w < 8; ++w) s += red[w][21 + r];
          g[r] = s;
          A[r][r] += 1e-9;
        }
        // Cholesky A = L L^T, solve
        bool ok = true;
        for (int j = 0; j < 6 && ok; ++j) {
          double s = A[j][j];
          for (int k = 0; k < j; ++k) s -= A[j][k] * A[j][k];
          if (!(s > 0.0)) { ok = false; break; }
          A[j][j] = sqrt(s);
          for (int i = j + 1; i < 6; ++i) {
            double v = A[i][j];
            for (int k = 0; k < j; ++k) v -= A[i][k] * A[j][k];
            A[i][j] = v / A[j][j];
          }
        }
        double d[6] = {0, 0, 0, 0, 0, 0};
        if (ok) {
          for (int i = 0; i < 6; ++i) {
            double v = g[i];
            for (int k = 0; k < i; ++k) v -= A[i][k] * d[k];
            d[i] = v / A[i][i];
          }
          for (int i = 5; i >= 0; --i) {
            double v = d[i];
            for (int k = i + 1; k < 6; ++k) v -= A[k][i] * d[k];
            d[i] = v / A[i][i];
          }
        }
        // R <- Exp(w) R, t <- Exp(w) t + dt
        const double wx = d[0], wy = d[1], wz = d[2];
        const double th2 = wx * wx + wy * wy + wz * wz, th = sqrt(th2);
        double sa, sb;  // Exp(w) = I + sa [w]x + sb [w]x^2
        if (th < 1e-12) sa = 1.0, sb = 0.5;
        else sa = sin(th) / th, sb = (1.0 - cos(th)) / th2;
        const double Kx[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
        double K2[9], E[9];
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j) K2[3 * i + j] = Kx[3 * i] * Kx[j] + Kx[3 * i + 1] * Kx[3 + j] + Kx[3 * i + 2] * Kx[6 + j];
        for (int i = 0; i < 9; ++i) E[i] = ((i % 4 == 0) ? 1.0 : 0.0) + sa * Kx[i] + sb * K2[i];
        double Rn[9], tn[3];
        for (int i = 0; i < 3; ++i) {
          for (int j = 0; j < 3; ++j) Rn[3 * i + j] = E[3 * i] * sR[j] + E[3 * i + 1] * sR[3 + j] + E[3 * i + 2] * sR[6 + j];
          tn[i] = E[3 * i] * st[0] + E[3 * i + 1] * st[1] + E[3 * i + 2] * st[2] + d[3 + i];
        }
        for (int i = 0; i < 9; ++i) sR[i] = Rn[i];
        for (int i = 0; i < 3; ++i) st[i] = tn[i];
      }
      __syncthreads();
    }
    const int n2 = mark_inliers();
    if (n2 < 4) break;
    n_in = n2;
  }
  if (tid < 12) pose[12 * frame + tid] = (tid % 4 == 3) ? st[tid / 4] : sR[3 * (tid / 4) + tid % 4];
  if (tid == 0) counts[2 * frame] = n_in;
}

}  // namespace opb

using namespace opb;

extern "C" {

size_t opb200_ransac_pnp_workspace_bytes(int n_frames, int iterations) {
  if (n_frames < 1 || iterations < 1) return 0;
  return static_cast<size_t>(n_frames) * iterations * (12 * 4 + 4) + 256;
}

int opb200_ransac_pnp(const float* pts3d, int stride3, const float* pts2d, int stride2, const int* offsets, int n_frames,
                      const double* intrinsics, float reproj_error_px, int iterations, unsigned seed, double* pose,
                      unsigned char* inlier_mask, int* counts, void* workspace, size_t workspace_bytes, void* stream_) {
  if (int r = check_device()) return r;
  OPB_CHECK_ARG(pts3d && pts2d && offsets && intrinsics && pose && inlier_mask && counts && workspace, "null pointer");
  OPB_CHECK_ARG(n_frames >= 1 && iterations >= 1 && stride3 >= 3 && stride2 >= 2, "bad shape");
  OPB_CHECK_ARG(reproj_error_px > 0.f, "reprojection error must be positive");
  if (opb200_ransac_pnp_workspace_bytes(n_frames, iterations) > workspace_bytes) {
    set_last_error("workspace too small");
    return OPB_ERR_WORKSPACE;
  }
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  float* hyp = static_cast<float*>(workspace);
  int* score = reinterpret_cast<int*>(hyp + static_cast<size_t>(n_frames) * iterations * 12);
  const float thr2 = reproj_error_px * reproj_error_px;
  pnp_hypotheses_kernel<<<dim3((iterations + 127) / 128, n_frames), 128, 0, stream>>>(pts3d, stride3, pts2d, stride2, offsets,
                                                                                      intrinsics, iterations, seed, hyp);
  pnp_score_kernel<<<dim3((iterations + 8 * PNP_HPW - 1) / (8 * PNP_HPW), n_frames), 256, 0, stream>>>(
      pts3d, stride3, pts2d, stride2, offsets, intrinsics, iterations, thr2, hyp, score);
  pnp_refine_kernel<<<n_frames, 256, 0, stream>>>(pts3d, stride3, pts2d, stride2, offsets, intrinsics, iterations, thr2, hyp, score,
                                                  pose, inlier_mask, counts);
  count_launch(3);
  OPB_CHECK_CUDA(cudaGetLastError());
  return OPB_OK;
}

}  // extern "C"
