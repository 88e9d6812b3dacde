// kernels.cuh -- sm_100a device code of the batched rigid-body collision engine.
//
// One warp owns a TILE of up to 32 poses (32 independent configurations for validity, 32
// consecutive interpolation steps of one segment for the segment walk).  The warp keeps ONE
// work stack of (pose, env node, robot node) entries in shared memory; every round the 32 lanes
// pop 32 entries, run 32 OBB-OBB tests (fp32, conservative), and push the surviving child pairs
// with ballot/popc compaction.  Leaf-leaf survivors go to a second queue and are tested 32 at a
// time with the exact-input triangle SAT.  A pose that already has a hit is dropped lazily at
// pop time.  This replaces FCL's recursive per-state DFS
// (/root/reference/third-party/fcl/include/fcl/narrowphase/detail/traversal/traversal_recurse-inl.h:83-128).
//
// Label exactness: the leaf predicate is FCL's 17-axis SAT
// (.../traversal/collision/intersect-inl.h:728-799, project6 :1086-1107).  Triangle vertices are
// transformed in fp64 with exactly the reference's operation order, rounded to fp32 and run
// through a filtered SAT with a proven forward error bound; any pair whose decision is not
// certain is re-evaluated by the fp64 SAT-17, so the returned boolean always equals the fp64
// result.  Compile with -fmad=false: fp64 code must not be contracted; fp32 box tests use
// explicit fmaf().
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "se3.h"

namespace osr {

struct DevModel {
  const float4* nodes;  // 4 x float4 per node (see geometry.h)
  const double* tris;   // 9 doubles per triangle
  int num_nodes;
  int num_tris;
};

struct DevCounters {
  unsigned long long bv, leaf, fp64, poses, spills;
};

struct KParams {
  DevModel env, rob;
  Rigid env_tf;                    // environment transform (perturbation), fp64
  DevCounters* counters;
  unsigned long long* work_counter;  // dynamic tile queue
  unsigned long long* spill;         // [total_warps][spill_cap]
  int spill_cap;
  int* error_flag;                   // set to 1 on spill exhaustion
};

constexpr int STACK_CAP = 512;   // shared-memory stack entries per warp
constexpr int SPILL_CHUNK = 256; // entries moved per spill / refill
constexpr unsigned FULL = 0xffffffffu;

struct __align__(16) WarpShared {
  unsigned long long stack[STACK_CAP];
  unsigned long long leafq[64];
  double Rd[12][32];  // per pose: R (row-major 9) + t (3), fp64, robot in env frame
  float Rf[12][32];   // same rounded to fp32
  float eta[32];      // per-pose absolute inflation of box extents
  double tau[32];     // segment walk: tau of each step in the tile
};

__device__ __forceinline__ unsigned long long make_entry(unsigned pose, unsigned n1, unsigned n2) {
  return ((unsigned long long)pose << 59) | ((unsigned long long)n2 << 32) | n1;
}
__device__ __forceinline__ unsigned entry_pose(unsigned long long e) { return (unsigned)(e >> 59); }
__device__ __forceinline__ unsigned entry_n1(unsigned long long e) { return (unsigned)e; }
__device__ __forceinline__ unsigned entry_n2(unsigned long long e) { return (unsigned)(e >> 32) & 0x07ffffffu; }

// ---------------------------------------------------------------------------------------------
// fp32 OBB-OBB separating-axis test (15 axes).  Conservative: extents get +eta, |B| gets +reps,
// both far above the fp32 rounding of the products below ("box test slack").
// Same axis set as FCL's obbDisjoint (third-party/fcl/include/fcl/math/bv/OBB-inl.h:399-525).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool obb_overlap_f32(const float4 a0, const float4 a1, const float4 a2, const float4 a3,
                                                const float4 b0, const float4 b1, const float4 b2, const float4 b3,
                                                const float* __restrict__ Rf, float eta) {
  const float reps = 2e-5f;
  float r0 = Rf[0 * 32], r1 = Rf[1 * 32], r2 = Rf[2 * 32], r3 = Rf[3 * 32], r4 = Rf[4 * 32], r5 = Rf[5 * 32],
        r6 = Rf[6 * 32], r7 = Rf[7 * 32], r8 = Rf[8 * 32];
  // v_j = R0 * (axis j of box b)
  float v0x = fmaf(r0, b0.x, fmaf(r1, b0.y, r2 * b0.z)), v0y = fmaf(r3, b0.x, fmaf(r4, b0.y, r5 * b0.z)),
        v0z = fmaf(r6, b0.x, fmaf(r7, b0.y, r8 * b0.z));
  float v1x = fmaf(r0, b1.x, fmaf(r1, b1.y, r2 * b1.z)), v1y = fmaf(r3, b1.x, fmaf(r4, b1.y, r5 * b1.z)),
        v1z = fmaf(r6, b1.x, fmaf(r7, b1.y, r8 * b1.z));
  float v2x = fmaf(r0, b2.x, fmaf(r1, b2.y, r2 * b2.z)), v2y = fmaf(r3, b2.x, fmaf(r4, b2.y, r5 * b2.z)),
        v2z = fmaf(r6, b2.x, fmaf(r7, b2.y, r8 * b2.z));
  // centre of b in a's frame
  float tx = fmaf(r0, b0.w, fmaf(r1, b1.w, fmaf(r2, b2.w, Rf[9 * 32]))) - a0.w;
  float ty = fmaf(r3, b0.w, fmaf(r4, b1.w, fmaf(r5, b2.w, Rf[10 * 32]))) - a1.w;
  float tz = fmaf(r6, b0.w, fmaf(r7, b1.w, fmaf(r8, b2.w, Rf[11 * 32]))) - a2.w;
  float T0 = fmaf(a0.x, tx, fmaf(a0.y, ty, a0.z * tz));
  float T1 = fmaf(a1.x, tx, fmaf(a1.y, ty, a1.z * tz));
  float T2 = fmaf(a2.x, tx, fmaf(a2.y, ty, a2.z * tz));
  // B(i,j) = axis_i(a) . v_j
  float B00 = fmaf(a0.x, v0x, fmaf(a0.y, v0y, a0.z * v0z)), B01 = fmaf(a0.x, v1x, fmaf(a0.y, v1y, a0.z * v1z)),
        B02 = fmaf(a0.x, v2x, fmaf(a0.y, v2y, a0.z * v2z));
  float B10 = fmaf(a1.x, v0x, fmaf(a1.y, v0y, a1.z * v0z)), B11 = fmaf(a1.x, v1x, fmaf(a1.y, v1y, a1.z * v1z)),
        B12 = fmaf(a1.x, v2x, fmaf(a1.y, v2y, a1.z * v2z));
  float B20 = fmaf(a2.x, v0x, fmaf(a2.y, v0y, a2.z * v0z)), B21 = fmaf(a2.x, v1x, fmaf(a2.y, v1y, a2.z * v1z)),
        B22 = fmaf(a2.x, v2x, fmaf(a2.y, v2y, a2.z * v2z));
  float F00 = fabsf(B00) + reps, F01 = fabsf(B01) + reps, F02 = fabsf(B02) + reps;
  float F10 = fabsf(B10) + reps, F11 = fabsf(B11) + reps, F12 = fabsf(B12) + reps;
  float F20 = fabsf(B20) + reps, F21 = fabsf(B21) + reps, F22 = fabsf(B22) + reps;
  float ea0 = a3.x + eta, ea1 = a3.y + eta, ea2 = a3.z + eta;
  float eb0 = b3.x + eta, eb1 = b3.y + eta, eb2 = b3.z + eta;
  bool sep = false;
  sep |= fabsf(T0) > ea0 + fmaf(F00, eb0, fmaf(F01, eb1, F02 * eb2));
  sep |= fabsf(T1) > ea1 + fmaf(F10, eb0, fmaf(F11, eb1, F12 * eb2));
  sep |= fabsf(T2) > ea2 + fmaf(F20, eb0, fmaf(F21, eb1, F22 * eb2));
  sep |= fabsf(fmaf(B00, T0, fmaf(B10, T1, B20 * T2))) > eb0 + fmaf(F00, ea0, fmaf(F10, ea1, F20 * ea2));
  sep |= fabsf(fmaf(B01, T0, fmaf(B11, T1, B21 * T2))) > eb1 + fmaf(F01, ea0, fmaf(F11, ea1, F21 * ea2));
  sep |= fabsf(fmaf(B02, T0, fmaf(B12, T1, B22 * T2))) > eb2 + fmaf(F02, ea0, fmaf(F12, ea1, F22 * ea2));
  if (!sep) {
    // A0 x Bj
    sep |= fabsf(fmaf(T2, B10, -T1 * B20)) > fmaf(ea1, F20, fmaf(ea2, F10, fmaf(eb1, F02, eb2 * F01)));
    sep |= fabsf(fmaf(T2, B11, -T1 * B21)) > fmaf(ea1, F21, fmaf(ea2, F11, fmaf(eb0, F02, eb2 * F00)));
    sep |= fabsf(fmaf(T2, B12, -T1 * B22)) > fmaf(ea1, F22, fmaf(ea2, F12, fmaf(eb0, F01, eb1 * F00)));
    // A1 x Bj
    sep |= fabsf(fmaf(T0, B20, -T2 * B00)) > fmaf(ea0, F20, fmaf(ea2, F00, fmaf(eb1, F12, eb2 * F11)));
    sep |= fabsf(fmaf(T0, B21, -T2 * B01)) > fmaf(ea0, F21, fmaf(ea2, F01, fmaf(eb0, F12, eb2 * F10)));
    sep |= fabsf(fmaf(T0, B22, -T2 * B02)) > fmaf(ea0, F22, fmaf(ea2, F02, fmaf(eb0, F11, eb1 * F10)));
    // A2 x Bj
    sep |= fabsf(fmaf(T1, B00, -T0 * B10)) > fmaf(ea0, F10, fmaf(ea1, F00, fmaf(eb1, F22, eb2 * F21)));
    sep |= fabsf(fmaf(T1, B01, -T0 * B11)) > fmaf(ea0, F11, fmaf(ea1, F01, fmaf(eb0, F22, eb2 * F20)));
    sep |= fabsf(fmaf(T1, B02, -T0 * B12)) > fmaf(ea0, F12, fmaf(ea1, F02, fmaf(eb0, F21, eb1 * F20)));
  }
  return !sep;
}

// ---------------------------------------------------------------------------------------------
// Triangle-triangle SAT-17.  Points are relative to P1 (p1 == 0), q already in the env frame.
// ---------------------------------------------------------------------------------------------
template <typename T>
struct Tri6 {  // p2, p3, q1, q2, q3
  T p2[3], p3[3], q1[3], q2[3], q3[3];
};

template <typename T>
__device__ __forceinline__ void crossT(const T* a, const T* b, T* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}
template <typename T>
__device__ __forceinline__ T dotT(const T* a, const T* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
template <typename T>
__device__ __forceinline__ T min3(T a, T b, T c) { return a < (b < c ? b : c) ? a : (b < c ? b : c); }
template <typename T>
__device__ __forceinline__ T max3(T a, T b, T c) { return a > (b > c ? b : c) ? a : (b > c ? b : c); }

// project6 (intersect-inl.h:1086-1107): gap1 = min(P) - max(Q), gap2 = min(Q) - max(P);
// the axis separates iff gap1 > 0 or gap2 > 0.  dot(ax, p1) is exactly 0.
template <typename T>
__device__ __forceinline__ void axis_gaps(const T* ax, const Tri6<T>& t, T& g1, T& g2) {
  T P2 = dotT(ax, t.p2), P3 = dotT(ax, t.p3);
  T Q1 = dotT(ax, t.q1), Q2 = dotT(ax, t.q2), Q3 = dotT(ax, t.q3);
  T mnP = min3<T>(T(0), P2, P3), mxP = max3<T>(T(0), P2, P3);
  T mnQ = min3<T>(Q1, Q2, Q3), mxQ = max3<T>(Q1, Q2, Q3);
  g1 = mnP - mxQ;
  g2 = mnQ - mxP;
}

// Exact restatement (fp64, reference operation order) -> true iff no axis separates.
__device__ __noinline__ bool sat17_exact(const Tri6<double>& t) {
  double e1[3] = {t.p2[0], t.p2[1], t.p2[2]};                                  // p2 - p1
  double e2[3] = {t.p3[0] - t.p2[0], t.p3[1] - t.p2[1], t.p3[2] - t.p2[2]};   // p3 - p2
  double e3[3] = {0.0 - t.p3[0], 0.0 - t.p3[1], 0.0 - t.p3[2]};               // p1 - p3
  double f1[3] = {t.q2[0] - t.q1[0], t.q2[1] - t.q1[1], t.q2[2] - t.q1[2]};
  double f2[3] = {t.q3[0] - t.q2[0], t.q3[1] - t.q2[1], t.q3[2] - t.q2[2]};
  double f3[3] = {t.q1[0] - t.q3[0], t.q1[1] - t.q3[1], t.q1[2] - t.q3[2]};
  double n1[3], m1[3], ax[3], g1, g2;
  crossT(e1, e2, n1);
  axis_gaps(n1, t, g1, g2); if (g1 > 0 || g2 > 0) return false;
  crossT(f1, f2, m1);
  axis_gaps(m1, t, g1, g2); if (g1 > 0 || g2 > 0) return false;
  const double* E[3] = {e1, e2, e3};
  const double* F[3] = {f1, f2, f3};
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      crossT(E[i], F[j], ax);
      axis_gaps(ax, t, g1, g2); if (g1 > 0 || g2 > 0) return false;
    }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    crossT(E[i], n1, ax);
    axis_gaps(ax, t, g1, g2); if (g1 > 0 || g2 > 0) return false;
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    crossT(F[i], m1, ax);
    axis_gaps(ax, t, g1, g2); if (g1 > 0 || g2 > 0) return false;
  }
  return true;
}

// fp32 filter.  Inputs are the fp64 values rounded to nearest; with u = 2^-24 and
// L = max |coordinate|, every gap of the 11 "quadratic" axes (n1, m1, e x f) is within
// 600 u L^3 of its exact value and every gap of the 6 "cubic" axes (e x n1, f x m1) within
// 3000 u L^4 (derivation in DESIGN.md).  Returns 0 = certainly separated, 1 = certainly
// intersecting (no axis separates), 2 = undecided.
__device__ __forceinline__ int sat17_filter(const Tri6<float>& t) {
  float L = 0.f;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    L = fmaxf(L, fmaxf(fabsf(t.p2[k]), fabsf(t.p3[k])));
    L = fmaxf(L, fmaxf(fabsf(t.q1[k]), fmaxf(fabsf(t.q2[k]), fabsf(t.q3[k]))));
  }
  const float u = 5.9604645e-08f;
  float L3 = L * L * L;
  float Eq = 640.f * u * L3 + 1e-37f;
  float Ec = 3200.f * u * L3 * L + 1e-37f;
  float e1[3] = {t.p2[0], t.p2[1], t.p2[2]};
  float e2[3] = {t.p3[0] - t.p2[0], t.p3[1] - t.p2[1], t.p3[2] - t.p2[2]};
  float e3[3] = {-t.p3[0], -t.p3[1], -t.p3[2]};
  float f1[3] = {t.q2[0] - t.q1[0], t.q2[1] - t.q1[1], t.q2[2] - t.q1[2]};
  float f2[3] = {t.q3[0] - t.q2[0], t.q3[1] - t.q2[1], t.q3[2] - t.q2[2]};
  float f3[3] = {t.q1[0] - t.q3[0], t.q1[1] - t.q3[1], t.q1[2] - t.q3[2]};
  float n1[3], m1[3], ax[3], g1, g2;
  bool sep = false, unsure = false;
  crossT(e1, e2, n1);
  axis_gaps(n1, t, g1, g2); sep |= (g1 > Eq) | (g2 > Eq); unsure |= !((g1 < -Eq) & (g2 < -Eq));
  crossT(f1, f2, m1);
  axis_gaps(m1, t, g1, g2); sep |= (g1 > Eq) | (g2 > Eq); unsure |= !((g1 < -Eq) & (g2 < -Eq));
  if (sep) return 0;
  const float* E[3] = {e1, e2, e3};
  const float* F[3] = {f1, f2, f3};
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      crossT(E[i], F[j], ax);
      axis_gaps(ax, t, g1, g2); sep |= (g1 > Eq) | (g2 > Eq); unsure |= !((g1 < -Eq) & (g2 < -Eq));
    }
  if (sep) return 0;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    crossT(E[i], n1, ax);
    axis_gaps(ax, t, g1, g2); sep |= (g1 > Ec) | (g2 > Ec); unsure |= !((g1 < -Ec) & (g2 < -Ec));
    crossT(F[i], m1, ax);
    axis_gaps(ax, t, g1, g2); sep |= (g1 > Ec) | (g2 > Ec); unsure |= !((g1 < -Ec) & (g2 < -Ec));
  }
  if (sep) return 0;
  return unsure ? 2 : 1;
}

// Robot triangle into the env frame with the reference's operation order
// (intersect-inl.h:598-617: Q' = R*Q + T; then :738-743 everything minus P1).
__device__ __forceinline__ void load_pair_exact(const double* __restrict__ P, const double* __restrict__ Q,
                                                const double* __restrict__ Rd, Tri6<double>& t) {
  double R[12];
#pragma unroll
  for (int k = 0; k < 12; k++) R[k] = Rd[k * 32];
  double p1[3] = {__ldg(P), __ldg(P + 1), __ldg(P + 2)};
#pragma unroll
  for (int k = 0; k < 3; k++) {
    t.p2[k] = __ldg(P + 3 + k) - p1[k];
    t.p3[k] = __ldg(P + 6 + k) - p1[k];
  }
  double* qo[3] = {t.q1, t.q2, t.q3};
#pragma unroll
  for (int v = 0; v < 3; v++) {
    double x = __ldg(Q + 3 * v), y = __ldg(Q + 3 * v + 1), z = __ldg(Q + 3 * v + 2);
#pragma unroll
    for (int k = 0; k < 3; k++) {
      double w = (R[3 * k] * x + R[3 * k + 1] * y + R[3 * k + 2] * z) + R[9 + k];
      qo[v][k] = w - p1[k];
    }
  }
}

// returns true iff the fp64 SAT-17 reports an intersection
__device__ __forceinline__ bool tri_pair_intersect(const Tri6<double>& td, unsigned& n_recheck) {
  Tri6<float> tf;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
    tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
  }
  int r = sat17_filter(tf);
  if (r == 2) {
    n_recheck++;
    return sat17_exact(td);
  }
  return r == 1;
}

