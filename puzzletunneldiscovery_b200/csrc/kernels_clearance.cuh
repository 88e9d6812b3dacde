// kernels_clearance.cuh -- K8: clearance statistics of key-configuration candidates.
//
// Reference: for every candidate q0 the pipeline shoots `npoint` touch segments (touchq_util.calc_touch,
// /root/reference/src/GP/pipeline/touchq_util.py:8-27 -> UnitWorld::transitStateTo with contact), ships all
// npoint x 15 doubles back, and reduces them to five numbers on the host:
//   distances = pyosr.multi_distance(q0, FREE_V)              lib/osr/osr_state.cc:157-167
//   [median, max, min, mean, std](distances)                  src/GP/pipeline/preprocess_key.py:275-281
// (estimate_keyconf_clearance keeps the distances themselves: src/GP/pipeline/keyconf.py:312-315.)
//
// Here K2 leaves the last free state of every segment in HBM and this kernel reduces them in place: one block
// per candidate, the distances of its segments in shared memory, a bitonic sort (padding +inf) for the order
// statistics, fixed-shape tree reductions for mean and variance.  min / max / median are exact (numpy's median of
// an even count is the mean of the two middle elements); mean and std agree with numpy's pairwise summation to
// rounding (1e-13 relative in the tests).
#pragma once
#include "kernels.cuh"

namespace osr {

constexpr int CLR_THREADS = 256;
constexpr int CLR_MAX_POINTS = 8192;   // 64 KB of dynamic shared memory

// from[c * npoint + i] = q0[c]: the `from` array of the touch segments of every candidate
__global__ void __launch_bounds__(256)
    k_repeat_states(const double* __restrict__ q0, long long m, int npoint, double* __restrict__ from) {
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // one double per thread
  long long total = m * (long long)npoint * 7;
  if (t >= total) return;
  long long seg = t / 7;
  int c = (int)(t - seg * 7);
  from[t] = __ldg(q0 + 7 * (seg / npoint) + c);
}

__device__ __forceinline__ double clr_block_sum(double v, double* red) {
  // fixed shape: xor-shuffle inside the warp, then the warp sums in order
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < CLR_THREADS / 32; k++) s += red[k];
  return s;
}

// stats[c] = {median, max, min, mean, std}; distances (nullable) receives the unsorted distances
__global__ void __launch_bounds__(CLR_THREADS)
    k_clearance_stats(const double* __restrict__ q0, const double* __restrict__ last_free, int npoint, int padded,
                      double* __restrict__ stats, double* __restrict__ distances) {
  extern __shared__ double sd[];
  __shared__ double red[CLR_THREADS / 32];
  const long long c = blockIdx.x;
  double a[7];
#pragma unroll
  for (int k = 0; k < 7; k++) a[k] = __ldg(q0 + 7 * c + k);
  double part = 0.0;
  for (int i = threadIdx.x; i < padded; i += CLR_THREADS) {
    double d = __longlong_as_double(0x7ff0000000000000LL);
    if (i < npoint) {
      double b[7];
      const double* p = last_free + 7 * (c * (long long)npoint + i);
#pragma unroll
      for (int k = 0; k < 7; k++) b[k] = __ldg(p + k);
      d = distance(a, b);
      if (distances) distances[c * (long long)npoint + i] = d;
      part += d;
    }
    sd[i] = d;
  }
  const double mean = clr_block_sum(part, red) / (double)npoint;
  part = 0.0;
  for (int i = threadIdx.x; i < npoint; i += CLR_THREADS) {
    double e = sd[i] - mean;
    part += e * e;
  }
  const double var = clr_block_sum(part, red) / (double)npoint;
  // bitonic sort, ascending
  for (int k = 2; k <= padded; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < padded; i += CLR_THREADS) {
        int l = i ^ j;
        if (l > i) {
          double x = sd[i], y = sd[l];
          bool up = (i & k) == 0;
          if ((x > y) == up) { sd[i] = y; sd[l] = x; }
        }
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double med = (npoint & 1) ? sd[npoint >> 1] : (sd[(npoint >> 1) - 1] + sd[npoint >> 1]) * 0.5;
    double* o = stats + 5 * c;
    o[0] = med;
    o[1] = sd[npoint - 1];
    o[2] = sd[0];
    o[3] = mean;
    o[4] = sqrt(var);
  }
}

}  // namespace osr
