// kernels_knn.cuh -- K6: k nearest milestones under the SE(3) state metric inside windows of consecutive indices
// (config 3 of BASELINE.json / SURVEY.md 8(d) D4: the k-NN edges whose validity K2 then checks; N1 of 8(f)).
//
// Metric: osr::distance (/root/reference/lib/osr/osr_state.cc:130-154) = |dt| + angularDistance(q0, q1) with both
// quaternions normalised, the metric the reference's planners use for their nearest-neighbour structures.
// One thread per query milestone; the candidates of its window stream through shared memory in tiles of
// KNN_TILE states (quaternions normalised once per tile); every thread keeps its k best (distance, index)
// pairs in registers, ordered by (distance, index).  The distance is fp64 throughout, so that the neighbour sets
// equal those of the host restatement (workloads.knn_edges) except on exact distance ties.
// A candidate reaches the fp64 distance (two square roots and an atan2) only if a LOWER BOUND of it, computed in
// fp32 with its rounding subtracted, does not exceed the thread's current k-th best distance:
//     |dt| + 2 atan2(|v|, |w|)  >=  |dt| + 2 * 2 sin(phi / 2),   cos(phi) = |qa . qb|,  (2 sin(phi/2))^2 = 2 (1 - |qa . qb|)
// (arc >= chord).  The k-th best only decreases, so a candidate skipped once could never have entered the list: the
// output is the same, and ~99 % of the candidates of a 4 096-state window stop at the bound.
#pragma once
#include "kernels.cuh"

namespace osr {

constexpr int KNN_TILE = 256;
constexpr int KNN_KMAX = 16;
constexpr int KNN_PBUF = 16;

__global__ void __launch_bounds__(KNN_TILE)
    k_knn_window(const double* __restrict__ q, long long n, int k, long long window, int* __restrict__ out_idx,
                 double* __restrict__ out_dist) {
  __shared__ double cand[7][KNN_TILE];
  __shared__ float candq[4][KNN_TILE];   // the normalised quaternions again, rounded to fp32 (lower bound)
  __shared__ unsigned char pbuf[KNN_PBUF][KNN_TILE];   // per thread: tile offsets of candidates that passed the bound
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
  const float mq0 = (float)me[3], mq1 = (float)me[4], mq2 = (float)me[5], mq3 = (float)me[6];
  double bd[KNN_KMAX];
  int bi[KNN_KMAX];
#pragma unroll
  for (int s = 0; s < KNN_KMAX; s++) { bd[s] = 1e300; bi[s] = -1; }
  double thr = 1e300;        // bd[k - 1]: the k-th best so far
  int thr_i = -1;            // bi[k - 1]
  float thr_up = __int_as_float(0x7f800000);   // thr rounded up to fp32 (+inf while the list is not full)
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
#pragma unroll
      for (int c = 0; c < 4; c++) candq[c][threadIdx.x] = (float)s[3 + c];
    }
    __syncthreads();
    const int cnt = (int)min((long long)KNN_TILE, w1 - t0);
    int np = 0;   // candidates of this tile that passed the bound and wait for their fp64 distance
    for (int c0 = 0; c0 <= cnt; c0++) {
      if (c0 < cnt) {
        const int c = c0;
        const double dx = me[0] - cand[0][c], dy = me[1] - cand[1][c], dz = me[2] - cand[2][c];
        // lower bound, fp32: the differences are rounded once (relative 6e-8), the rest carries < 1e-6 relative;
        // the quaternion dot product is within 4e-7 of the fp64 one, (chord)^2 = 8 (1 - |dot|) within 4e-6 (taken off
        // before the root; it also covers quaternions that are unit only to rounding)
        const float fx = (float)dx, fy = (float)dy, fz = (float)dz;
        const float trf = sqrtf(fmaf(fx, fx, fmaf(fy, fy, fz * fz))) * (1.0f - 2e-6f);
        const float dotf = fabsf(fmaf(mq0, candq[0][c], fmaf(mq1, candq[1][c], fmaf(mq2, candq[2][c], mq3 * candq[3][c]))));
        const float ch2 = fmaf(-8.0f, dotf, 8.0f) - 4e-6f;
        const float chf = ch2 > 0.f ? sqrtf(ch2) * (1.0f - 2e-6f) : 0.f;
        const bool skip = (trf + chf) * (1.0f - 2e-6f) > thr_up || t0 + c == i || !have;
        if (!skip) { pbuf[np][threadIdx.x] = (unsigned char)c; np++; }
      }
      // Passing candidates are evaluated together: in a loop over the candidates of a tile some lane passes in every
      // second iteration, and the fp64 distance would run with one or two lanes active.  Drained when a lane's
      // buffer is full and at the end of the tile (the stale k-th best only lets a few more candidates through).
      if (!__any_sync(0xffffffffu, np == KNN_PBUF || (c0 == cnt && np > 0))) continue;
      const int rounds = __reduce_max_sync(0xffffffffu, np);
      for (int r = 0; r < rounds; r++) {
        if (r >= np) continue;
        const int c = pbuf[r][threadIdx.x];
        const long long j = t0 + c;
        const double dx = me[0] - cand[0][c], dy = me[1] - cand[1][c], dz = me[2] - cand[2][c];
        double tr = sqrt(dx * dx + dy * dy + dz * dz);
        // d = me * conj(other)
        double bw = cand[3][c], bx = -cand[4][c], by = -cand[5][c], bz = -cand[6][c];
        double dw = me[3] * bw - me[4] * bx - me[5] * by - me[6] * bz;
        double vx = me[3] * bx + me[4] * bw + me[5] * bz - me[6] * by;
        double vy = me[3] * by + me[5] * bw + me[6] * bx - me[4] * bz;
        double vz = me[3] * bz + me[6] * bw + me[4] * by - me[5] * bx;
        double d = tr + 2.0 * atan2(sqrt(vx * vx + vy * vy + vz * vz), fabs(dw));
        if (d < thr || (d == thr && (int)j < thr_i)) {
          // the newcomer replaces the k-th best and bubbles up (slots >= k stay at +inf / -1)
#pragma unroll
          for (int s = 0; s < KNN_KMAX; s++)
            if (s == k - 1) { bd[s] = d; bi[s] = (int)j; }
#pragma unroll
          for (int s = KNN_KMAX - 1; s > 0; s--) {
            bool sw = bd[s] < bd[s - 1] || (bd[s] == bd[s - 1] && bi[s] < bi[s - 1]);
            if (sw) { double td = bd[s]; bd[s] = bd[s - 1]; bd[s - 1] = td; int ti = bi[s]; bi[s] = bi[s - 1]; bi[s - 1] = ti; }
          }
#pragma unroll
          for (int s = 0; s < KNN_KMAX; s++)
            if (s == k - 1) { thr = bd[s]; thr_i = bi[s]; }
          thr_up = __double2float_ru(thr);
        }
      }
      np = 0;
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
