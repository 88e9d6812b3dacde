// kernels_keyconf.cuh -- K9: key-configuration candidates from surface-point pairs.
//
// Reference: UnitWorld::enumFreeConfiguration (/root/reference/lib/osr/unit_world.cc:1340-1389; the sampling
// variant :1295-1337) -- for a robot surface point + normal and an environment surface point + normal:
//   rot_1 aligns the robot normal with the NEGATED environment normal (Eigen setFromTwoVectors),
//   rot_2 = AngleAxis(i * 2 pi / den, env normal), i = 0 .. den-1 (axis not renormalised),
//   q_i   = compose(env_o - (rot_2 rot_1) * rob_o, rot_2 rot_1) with both points pushed out by `margin`,
// every q_i is validity-checked and, with only_median, only the middle element of every run of consecutive valid
// rotations that is closed by an invalid one is kept.  The pipeline calls it for NumberOfRotations x
// SurfacePairsToSample candidates per puzzle (src/GP/pipeline/keyconf.py:117-131).
//
// Here the candidates are built on the device (one thread per (pair, rotation); the den half-angle sines and
// cosines come from the host's libm so that every state is bit-identical to the host construction,
// osr::keyconf_state in se3.h), K1 labels them where they lie, and one thread per pair applies the run filter:
// 96 bytes per pair go up; states, labels and counts come back.
#pragma once
#include "kernels.cuh"

namespace osr {

__global__ void __launch_bounds__(256)
    k_keyconf_states(const double* __restrict__ pts, long long b, int den, double margin,
                     const double* __restrict__ half_cs, double* __restrict__ out_q) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b * den) return;
  const long long r = t / den;
  const int i = (int)(t - r * den);
  double p[12];
#pragma unroll
  for (int k = 0; k < 12; k++) p[k] = __ldg(pts + 12 * r + k);
  double q[7];
  keyconf_state(p, margin, __ldg(half_cs + 2 * i), __ldg(half_cs + 2 * i + 1), q);
#pragma unroll
  for (int k = 0; k < 7; k++) out_q[7 * t + k] = q[k];
}

// in place: valid[r][den] -> kept flags; counts[r] = number kept (unit_world.cc:1375-1386)
__global__ void __launch_bounds__(128)
    k_keyconf_filter(uint8_t* __restrict__ valid, long long b, int den, int only_median, int* __restrict__ counts) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= b) return;
  uint8_t* v = valid + r * den;
  int c = 0;
  if (only_median) {
    int i = 0;
    while (i < den) {
      if (!v[i]) { i++; continue; }
      int j = i;
      while (j + 1 < den && v[j + 1]) j++;
      // a run is only flushed by a following invalid sample; mid = begin + size / 2
      const int mid = (j + 1 < den) ? i + (j - i + 1) / 2 : -1;
      for (int k = i; k <= j; k++) v[k] = (k == mid) ? 1 : 0;
      c += mid >= 0 ? 1 : 0;
      i = j + 1;
    }
  } else {
    for (int i = 0; i < den; i++) c += v[i] ? 1 : 0;
  }
  if (counts) counts[r] = c;
}

}  // namespace osr
