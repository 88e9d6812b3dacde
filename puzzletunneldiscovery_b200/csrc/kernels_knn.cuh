// kernels_knn.cuh -- K6: k nearest milestones under the SE(3) state metric inside windows of consecutive indices
// (config 3 of BASELINE.json / SURVEY.md 8(d) D4: the k-NN edges whose validity K2 then checks; N1 of 8(f)).
//
// Metric: osr::distance (/root/reference/lib/osr/osr_state.cc:130-154) = |dt| + angularDistance(q0, q1) with both
// quaternions normalised, the metric the reference's planners use for their nearest-neighbour structures.
// One thread per query milestone; the candidates of its window stream through shared memory in tiles of
// KNN_TILE states (quaternions normalised once per tile); every thread keeps its k best (distance, index)
// pairs in registers, ordered by (distance, index).  fp64 throughout, so that the neighbour sets equal those of
// the host restatement (workloads.knn_edges) except on exact distance ties.
#pragma once
#include "kernels.cuh"

namespace osr {

constexpr int KNN_TILE = 256;
constexpr int KNN_KMAX = 16;

__global__ void __launch_bounds__(KNN_TILE)
    k_knn_window(const double* __restrict__ q, long long n, int k, long long window, int* __restrict__ out_idx,
                 double* __restrict__ out_dist) {
  __shared__ double cand[7][KNN_TILE];
  const long long i = (long long)blockIdx.x * KNN_TILE + threadIdx.x;   // KNN_TILE divides window: one window per block
  const long long w0 = ((long long)blockIdx.x * KNN_TILE) / window * window;
  const long long w1 = min(n, w0 + window);
  double me[7] = {0, 0, 0, 1, 0, 0, 0};
  const bool have = i < n;
  if (have) {
#pragma unroll
    for (int c = 0; c < 7; c++) me[c] = __ldg(q + 7 * i + c);
    quat_normalize(me[3], me[4], me[5], me[6]);
  }
  double bd[KNN_KMAX];
  int bi[KNN_KMAX];
#pragma unroll
  for (int s = 0; s < KNN_KMAX; s++) { bd[s] = 1e300; bi[s] = -1; }
  for (long long t0 = w0; t0 < w1; t0 += KNN_TILE) {
    __syncthreads();
    {
      long long j = t0 + threadIdx.x;
      double s[7] = {0, 0, 0, 1, 0, 0, 0};
      if (j < w1) {
#pragma unroll
        for (int c = 0; c < 7; c++) s[c] = __ldg(q + 7 * j + c);
        quat_normalize(s[3], s[4], s[5], s[6]);
      }
#pragma unroll
      for (int c = 0; c < 7; c++) cand[c][threadIdx.x] = s[c];
    }
    __syncthreads();
    if (!have) continue;
    const int cnt = (int)min((long long)KNN_TILE, w1 - t0);
    for (int c = 0; c < cnt; c++) {
      const long long j = t0 + c;
      if (j == i) continue;
      double dx = me[0] - cand[0][c], dy = me[1] - cand[1][c], dz = me[2] - cand[2][c];
      double tr = sqrt(dx * dx + dy * dy + dz * dz);
      // d = me * conj(other)
      double bw = cand[3][c], bx = -cand[4][c], by = -cand[5][c], bz = -cand[6][c];
      double dw = me[3] * bw - me[4] * bx - me[5] * by - me[6] * bz;
      double vx = me[3] * bx + me[4] * bw + me[5] * bz - me[6] * by;
      double vy = me[3] * by + me[5] * bw + me[6] * bx - me[4] * bz;
      double vz = me[3] * bz + me[6] * bw + me[4] * by - me[5] * bx;
      double d = tr + 2.0 * atan2(sqrt(vx * vx + vy * vy + vz * vz), fabs(dw));
      if (d < bd[KNN_KMAX - 1] || (d == bd[KNN_KMAX - 1] && (int)j < bi[KNN_KMAX - 1])) {
        bd[KNN_KMAX - 1] = d; bi[KNN_KMAX - 1] = (int)j;
#pragma unroll
        for (int s = KNN_KMAX - 1; s > 0; s--) {
          bool sw = bd[s] < bd[s - 1] || (bd[s] == bd[s - 1] && bi[s] < bi[s - 1]);
          if (sw) { double td = bd[s]; bd[s] = bd[s - 1]; bd[s - 1] = td; int ti = bi[s]; bi[s] = bi[s - 1]; bi[s - 1] = ti; }
        }
      }
    }
  }
  if (have) {
#pragma unroll
    for (int s = 0; s < KNN_KMAX; s++)
      if (s < k) {
        out_idx[(size_t)i * k + s] = bi[s];
        if (out_dist) out_dist[(size_t)i * k + s] = bi[s] >= 0 ? bd[s] : 0.0;
      }
  }
}

}  // namespace osr
