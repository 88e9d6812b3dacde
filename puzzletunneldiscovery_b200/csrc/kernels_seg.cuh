// kernels_seg.cuh -- K2, second generation: segment walk / first contact on the lane-owned depth-first
// traversal of kernels_dfs.cuh (sm_100a).
//
// Reference semantics (UnitWorld::transitStateToWithContact, /root/reference/lib/osr/unit_world.cc:405-474;
// isValidTransition :477-486): walk the segment at
//     delta_1 = vd;  delta_{k+1} = delta_k < dist ? min(dist, delta_k + vd) : delta_k + vd;  tau_k = delta_k / dist
// and return the FIRST step whose interpolated state collides.  Validity along a segment is not monotone, so
// every step before the first hit has to be evaluated; steps are independent of each other though, which is
// what this kernel uses:
//   * a WARP owns one segment at a time, a LANE owns one step of it and runs the same private depth-first
//     (environment node, robot node) traversal as K1 -- near-contact steps (hundreds of box tests) and free
//     steps (a handful) no longer wait for each other as they did in the 32-step tiles of the first
//     generation, and a colliding step stops at its first intersecting triangle pair;
//   * steps are prepared 32 at a time by the whole warp (the running fp64 delta sum reproduced add by add,
//     Eigen's slerp, the fp32 box pose) and handed to idle lanes IN ORDER; the first hit found cancels every
//     later step in flight, the segment ends when all earlier steps have finished;
//   * triangle pairs go through the same two queues (fp32-filtered SAT-17, then exact fp64 SAT-17) as in K1,
//     with the exact pose taken from the lane's interpolated state kept in shared memory.
#pragma once
#include "kernels_dfs.cuh"

namespace osr {

template <typename E, int DEPTH>
struct __align__(16) SegWarpMem {
  E stack[DEPTH][32];
  unsigned long long leafq[LEAFQ_CAP];
  unsigned long long req[REQ_CAP];
  float pend[17][32];        // BoxPose of the prepared steps: q[4] R[9] t[3] eta
  double pend_state[7][32];  // their interpolated states
  double cur_state[7][32];   // state of the step every lane works on (exact pose of the triangle tests)
  double pend_tau[33];       // [0] tau of the step before the prepared batch, [1 + j] tau of prepared step j
  double lane_tau[2][32];    // tau of each lane's step and of the step before it
  double ab[14];             // from, to
  double hit_tau, hit_ptau;  // taus recorded with the earliest hit so far
  double last_tau;           // tau of the last step of the walk that was prepared
};

struct SegArgs {
  const double* from;
  const double* to;
  long long n;
  const double* vd_arr;
  long long vd_count;
  double vd_scalar;
  int transition_mode;  // vd := min(dist/2, vd) (isValidTransition)
  SegOut O;
};

// Results of one segment (all lanes compute the same values; lane 0 / lanes 0..6 store them).
__device__ __noinline__ void seg_store(const SegArgs& A, long long seg, const double* ab, bool free_flag,
                                       double free_tau, double touch_tau, long long steps, int mode, int lane) {
  // mode 0: lf = from, fh = to (vd >= dist);  1: hit at the first step (lf = from);  2: hit later;  3: all free
  double lf[7], fh[7];
  if (mode == 0) {
#pragma unroll
    for (int k = 0; k < 7; k++) { lf[k] = ab[k]; fh[k] = ab[7 + k]; }
  } else if (mode == 3) {
    interpolate(ab, ab + 7, free_tau, lf);
#pragma unroll
    for (int k = 0; k < 7; k++) fh[k] = lf[k];
  } else {
    interpolate(ab, ab + 7, touch_tau, fh);
    if (mode == 2) interpolate(ab, ab + 7, free_tau, lf);
    else {
#pragma unroll
      for (int k = 0; k < 7; k++) lf[k] = ab[k];
    }
  }
  if (lane == 0) {
    if (A.O.is_free) A.O.is_free[seg] = free_flag ? 1 : 0;
    if (A.O.free_tau) A.O.free_tau[seg] = free_tau;
    if (A.O.touch_tau) A.O.touch_tau[seg] = touch_tau;
    if (A.O.steps) A.O.steps[seg] = steps;
  }
  if (lane < 7) {
    if (A.O.last_free) A.O.last_free[7 * seg + lane] = lf[lane];
    if (A.O.first_hit) A.O.first_hit[7 * seg + lane] = fh[lane];
  }
}

}  // namespace osr
