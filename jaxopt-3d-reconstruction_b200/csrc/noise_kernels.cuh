// f3 — `Dataset.with_noise` (src/dataset/dataset.py:43-63) on the device: the step that turns a ground-truth scene into the
// noisy start every benchmark of the reference optimises from.  Every perturbation is `magnitude x (random unit direction in
// the POSITIVE orthant)` (`_random_direction*`, dataset.py:32-41: `rand`, not `randn`); intrinsics get fx, fy, cx, cy, skew
// += U[0, magnitude) each (`CameraIntrinsics.apply_noise`, camera.py:30-47, masked); a pose is `apply_move(d_t)` then
// `apply_transform_3d(Rotation.from_rotvec(d_r))` (camera_pose.py:93-106).
// Counter-based generator (one stream per element and quantity, SplitMix64 of (seed, stream, element, k)): the result does
// not depend on the launch geometry.  Explicit perturbation arrays may replace the generator (parity tests against the
// reference's own operations, tests/golden/noise_ops.npz).
#pragma once
#include "common.cuh"

__device__ __forceinline__ uint64_t nz_mix(uint64_t z) {
  z += 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
// U[0,1) with 53 random bits
__device__ __forceinline__ double nz_uniform(uint64_t seed, uint32_t stream, uint64_t elem, uint32_t k) {
  const uint64_t h = nz_mix(nz_mix(seed ^ ((uint64_t)stream << 56)) ^ nz_mix(elem * 8 + k + 0x632be59bd9b4e019ULL));
  return (double)(h >> 11) * (1.0 / 9007199254740992.0);
}
template <int DIM>
__device__ __forceinline__ void nz_direction(uint64_t seed, uint32_t stream, uint64_t elem, double* d) {
  double n2 = 0.0;
#pragma unroll
  for (int k = 0; k < DIM; ++k) { d[k] = nz_uniform(seed, stream, elem, k); n2 += d[k] * d[k]; }
  const double inv = n2 > 0.0 ? 1.0 / sqrt(n2) : 0.0;
#pragma unroll
  for (int k = 0; k < DIM; ++k) d[k] *= inv;
}

// points (stream 0) and observations (stream 4): x += magnitude * direction   (or the explicit vectors)
template <int DIM>
__global__ void k_noise_vec(int64_t n, double* __restrict__ x, uint64_t seed, uint32_t stream, double mag,
                            const double* __restrict__ vec) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double d[DIM];
  if (vec) {
#pragma unroll
    for (int k = 0; k < DIM; ++k) d[k] = vec[DIM * i + k];
  } else {
    nz_direction<DIM>(seed, stream, (uint64_t)i, d);
#pragma unroll
    for (int k = 0; k < DIM; ++k) d[k] *= mag;
  }
#pragma unroll
  for (int k = 0; k < DIM; ++k) x[DIM * i + k] += d[k];
}

// cameras: pose (3x4 W2C [R|t], row-major) and intrinsics [fx, fy, cx, cy, skew]
__global__ void k_noise_cams(int C, double* __restrict__ poses, double* __restrict__ intr, uint64_t seed, double rot_mag,
                             double trans_mag, double intr_mag, int move_cameras, const double* __restrict__ move_vec,
                             const double* __restrict__ rot_vec, const double* __restrict__ intr_add) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double R[9], t[3], mv[3], w[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int k = 0; k < 3; ++k) R[3 * r + k] = poses[12 * (size_t)c + 4 * r + k];
    t[r] = poses[12 * (size_t)c + 4 * r + 3];
  }
  if (move_vec) { mv[0] = move_vec[3 * c]; mv[1] = move_vec[3 * c + 1]; mv[2] = move_vec[3 * c + 2]; }
  else { nz_direction<3>(seed, 1, (uint64_t)c, mv); mv[0] *= trans_mag; mv[1] *= trans_mag; mv[2] *= trans_mag; }
  if (rot_vec) { w[0] = rot_vec[3 * c]; w[1] = rot_vec[3 * c + 1]; w[2] = rot_vec[3 * c + 2]; }
  else { nz_direction<3>(seed, 2, (uint64_t)c, w); w[0] *= rot_mag; w[1] *= rot_mag; w[2] *= rot_mag; }
  // apply_move: on W2C poses the reference's implementation restores the translation it changed (a measured no-op,
  // tests/golden/make_noise_golden.py); move_cameras = 1 applies what the code intends: camera centre += mv, t -= R mv
  if (move_cameras) {
#pragma unroll
    for (int r = 0; r < 3; ++r) t[r] -= R[3 * r] * mv[0] + R[3 * r + 1] * mv[1] + R[3 * r + 2] * mv[2];
  }
  // Rotation.from_rotvec(w): Rodrigues with the small-angle limit
  const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
  double a, b;  // Rn = I + a [w]x + b [w]x^2
  if (th > 1e-8) { a = sin(th) / th; b = (1.0 - cos(th)) / th2; } else { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; }
  const double K[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
  double Rn[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double k2 = 0.0;
#pragma unroll
      for (int m = 0; m < 3; ++m) k2 += K[3 * i + m] * K[3 * m + j];
      Rn[3 * i + j] = (i == j ? 1.0 : 0.0) + a * K[3 * i + j] + b * k2;
    }
  // apply_transform_3d: [R|t] <- Rn [R|t]
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j)
      poses[12 * (size_t)c + 4 * i + j] = Rn[3 * i] * R[j] + Rn[3 * i + 1] * R[3 + j] + Rn[3 * i + 2] * R[6 + j];
    poses[12 * (size_t)c + 4 * i + 3] = Rn[3 * i] * t[0] + Rn[3 * i + 1] * t[1] + Rn[3 * i + 2] * t[2];
  }
#pragma unroll
  for (int k = 0; k < 5; ++k)
    intr[5 * (size_t)c + k] += intr_add ? intr_add[5 * (size_t)c + k] : intr_mag * nz_uniform(seed, 3, (uint64_t)c, k);
}
