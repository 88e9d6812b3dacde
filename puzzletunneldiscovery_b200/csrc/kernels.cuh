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
#include "sat17.h"

namespace osr {

struct DevModel {
  const float4* nodes;  // 4 x float4 per node (see geometry.h)
  const double* tris;   // 9 doubles per triangle
  int num_nodes;
  int num_tris;
};

struct DevCounters {
  unsigned long long bv, leaf, fp64, poses, spills;
  // lane occupancy of the depth-first kernels, in lane-iterations (one iteration = one box test slot)
  unsigned long long warp_iters, lanes_active, lanes_waiting, lanes_idle;
};

struct KParams {
  DevModel env, rob;
  Rigid env_tf;                    // environment transform (perturbation), fp64
  DevCounters* counters;
  unsigned long long* work_counter;  // dynamic tile queue
  unsigned long long* spill;         // [total_warps][spill_cap]
  int spill_cap;
  int* error_flag;                   // set to 1 on spill exhaustion
  double rob_radius;                 // max |p| over the robot's unit-mesh vertices (swept windows of K2)
  int sweep_tiles;                   // K2: largest swept window in tiles of 32 steps (0 = every step traversed)
  double sweep_rmax;                 // K2: windows whose swept radius exceeds this are not attempted
  double sweep_rstart;               // K2: swept radius of the first windows of a walk
  int sweep_short_tile;              // K2: steps of the exact tile that follows a blocked 4-step window (1..32)
  int sweep_nb0;                     // K2: windows in the first swept traversal of a walk (1..32)
  int sweep_resolve;                 // K2: most exact (triangle pair, step) tests spent on one window whose swept
                                     //     traversal left pairs that are not provably clear (0 = such windows are blocked)
  double sweep_keep_rmax;            // K2: ... and only windows whose swept radius stays below this keep pairs at all
};

#ifndef OSR_STACK_CAP
#define OSR_STACK_CAP 512
#endif
constexpr int STACK_CAP = OSR_STACK_CAP;   // shared-memory stack entries per warp
constexpr int SPILL_CHUNK = 256; // entries moved per spill / refill
constexpr unsigned FULL = 0xffffffffu;
// K2: most triangle pairs a swept traversal keeps for exact per-step tests of one window.  Kept small: shared memory
// per block decides how much of the SM's 256 KB is left as L1 for the node and triangle reads.
constexpr int SURV_CAP = 128;

struct __align__(16) WarpShared {
  unsigned long long stack[STACK_CAP];
  unsigned long long leafq[64];
  unsigned long long surv[SURV_CAP];  // (window, env triangle, robot triangle) not separated by more than the window's radius
  double Rd[12][32];  // per pose: R (row-major 9) + t (3), fp64, robot in env frame
  float Rf[12][32];   // same rounded to fp32
  float eta[32];      // per-pose absolute inflation of box extents
  double tau[32];     // segment walk: tau of each step in the tile
  double seg[14];     // K2: end states of the segment the warp walks (from, to) -- warp-uniform, kept out of registers
  SlerpPlan plan;     // K2: what osr::interpolate needs of them (normalised quaternions, arc angle, its sine)
  double envtf[12];   // K2: environment transform (R row-major, t)
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

// Edge vectors of both triangles and the a-th of the 11 "quadratic" axes (0: n1 = e1 x e2, 1: m1 = f1 x f2,
// 2 + 3 i + j: e_i x f_j -- the order of intersect-inl.h:745-777), operands picked with selects.
struct TriEdgesF {
  float e[3][3], f[3][3];
  __device__ __forceinline__ void set(const Tri6<float>& t) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
      e[0][k] = t.p2[k]; e[1][k] = t.p3[k] - t.p2[k]; e[2][k] = -t.p3[k];
      f[0][k] = t.q2[k] - t.q1[k]; f[1][k] = t.q3[k] - t.q2[k]; f[2][k] = t.q1[k] - t.q3[k];
    }
  }
  __device__ __forceinline__ void axis(unsigned a, float* ax) const {
    const unsigned b = a >= 2u ? a - 2u : 0u;
    const unsigned i = b / 3u, j = b - 3u * i;
    float u[3], w[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float ei = i == 0 ? e[0][k] : (i == 1 ? e[1][k] : e[2][k]);
      const float fj = j == 0 ? f[0][k] : (j == 1 ? f[1][k] : f[2][k]);
      u[k] = a == 0 ? e[0][k] : (a == 1 ? f[0][k] : ei);
      w[k] = a == 0 ? e[1][k] : (a == 1 ? f[1][k] : fj);
    }
    crossT(u, w, ax);
  }
};

// Swept windows of K2, fp32: true when one of the 11 quadratic axes separates the two triangles by MORE than r.
// The triangles are the fp64 values rounded to fp32; whatever direction the rounded cross product has, it is a
// valid direction to project on, so only the projections carry error: each within 7 u L |axis| of the exact one
// (input rounding + three products and two sums), the gap within 18 u L |axis| -- r is raised by 32 u L.  A pair
// that is not certainly clear is simply not clear (conservative; nothing is re-tested in fp64).  On `false`, best =
// the axis whose gap at this pose is largest relative to its length (15: no axis separates here) -- the first axis
// worth testing at the neighbouring poses (test_kept_pairs).
__device__ __noinline__ bool tri_gap_clear_f(const Tri6<float>& t, float r, unsigned& best) {
  float L = 0.f;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    L = fmaxf(L, fmaxf(fabsf(t.p2[k]), fabsf(t.p3[k])));
    L = fmaxf(L, fmaxf(fabsf(t.q1[k]), fmaxf(fabsf(t.q2[k]), fabsf(t.q3[k]))));
  }
  const float rp = r + 32.f * 5.9604645e-08f * L;
  const float rr = rp * rp * 1.00001f;
  TriEdgesF E;
  E.set(t);
  float best_g2 = 0.f, best_n2 = 1.f;
  best = 15u;
#pragma unroll 1
  for (unsigned a = 0; a < 11u; a++) {
    float ax[3], g1, g2;
    E.axis(a, ax);
    axis_gaps(ax, t, g1, g2);
    const float g = fmaxf(g1, g2), n2 = dotT(ax, ax);
    if (g > 0.f && n2 > 1e-30f) {
      const float gg = g * g;
      if (gg > rr * n2) return true;
      if (gg * best_n2 > best_g2 * n2) { best_g2 = gg; best_n2 = n2; best = a; }
    }
  }
  return false;
}

// One axis of the fp32 filter: true when quadratic axis a certainly separates (same arithmetic and bound as
// sat17_filter_c for that axis, so the exact fp64 SAT-17 finds the same axis separating and reports "disjoint").
__device__ __forceinline__ bool sat17_axis_separates(const Tri6<float>& t, unsigned a) {
  float L = 0.f;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    L = fmaxf(L, fmaxf(fabsf(t.p2[k]), fabsf(t.p3[k])));
    L = fmaxf(L, fmaxf(fabsf(t.q1[k]), fmaxf(fabsf(t.q2[k]), fabsf(t.q3[k]))));
  }
  TriEdgesF E;
  E.set(t);
  // the two edge vectors the axis is the cross product of (0: e1 x e2, 1: f1 x f2, 2 + 3 i + j: e_i x f_j)
  const unsigned b = a >= 2u ? a - 2u : 0u;
  const unsigned i = b / 3u, j = b - 3u * i;
  const float me0 = maxabs3(E.e[0]), me1 = maxabs3(E.e[1]), me2 = maxabs3(E.e[2]);
  const float mf0 = maxabs3(E.f[0]), mf1 = maxabs3(E.f[1]), mf2 = maxabs3(E.f[2]);
  const float mu = a == 0 ? me0 : (a == 1 ? mf0 : (i == 0 ? me0 : (i == 1 ? me1 : me2)));
  const float mw = a == 0 ? me1 : (a == 1 ? mf1 : (j == 0 ? mf0 : (j == 1 ? mf1 : mf2)));
  const float Eq = sat17_bound_q(1.1f * 5.9604645e-08f * L, L, mu, mw);
  float ax[3], g1, g2;
  E.axis(a, ax);
  axis_gaps(ax, t, g1, g2);
  return (g1 > Eq) | (g2 > Eq);
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

// bit 0: the fp64 SAT-17 reports an intersection; bit 1: the fp32 filter left the pair open and fp64 decided.
// One copy per kernel (noinline): the predicate is ~700 instructions and has several callers.
__device__ __noinline__ int tri_pair_intersect(const Tri6<double>& td) {
  // Tri6 and TriPair (sat17.h) have the same layout; the predicates are the host-tested ones of sat17.h
  const TriPair<double>& pd = reinterpret_cast<const TriPair<double>&>(td);
  TriPair<float> tf;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
    tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
  }
  const int r = sat17_filter_c(tf);
  if (r == 2) return 2 | (sat17_exact_c(pd) ? 1 : 0);
  return r;
}

// ---------------------------------------------------------------------------------------------
// Warp-cooperative traversal of one tile of poses.
// ---------------------------------------------------------------------------------------------
struct WarpCtx {
  unsigned long long* spill;  // this warp's global spill area
  int spill_cap;
  int nspill;
  unsigned n_bv, n_leaf, n_fp64, n_spill;
  bool overflow;
  int nsurv;       // MODE_SWEPT_KEEP: pairs kept for window keep_win (the first surv_limit of them are in WarpShared::surv)
  int surv_limit;  // ... pairs the window may keep before it counts as blocked (<= SURV_CAP)
  int keep_win;    // ... the earliest window that kept a pair (32: none)
};

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// Contact record emitted by the all-contacts mode.
struct ContactRec {
  int state;     // index of the state in the batch
  int env_tri;
  int rob_tri;
  int mult;      // number of rows FCL emits for this pair (0..2)
};

__device__ __forceinline__ unsigned contact_multiplicity_dev(const double* P, const double* Q, const double* Rd);

// MODE_RT: MODE_SEGMENT or MODE_SWEPT_KEEP, chosen at run time (`rmode`), so that k_transit holds ONE copy of the
// traversal loop (its instruction working set has to stay inside the instruction cache shared by the SM's warps).
enum { MODE_VALID = 0, MODE_SEGMENT = 1, MODE_CONTACTS = 2, MODE_SWEPT = 3, MODE_SWEPT_KEEP = 4, MODE_RT = 5 };

// alive: bit p set = pose p takes part.  Returns the mask of poses with at least one
// intersecting triangle pair.  MODE_SEGMENT: once pose p hits, poses >= p are abandoned (only
// the FIRST colliding step matters, unit_world.cc:452-458).  MODE_CONTACTS: no early out, every
// intersecting pair is appended to `recs`.  MODE_SWEPT: the poses are the mid poses of consecutive windows of a
// segment walk, S.eta carries each window's swept radius (boxes) and S.tau the radius itself (triangles); a pose
// "hits" as soon as one triangle pair is NOT separated by more than the radius along one of its 11 face / edge-edge
// axes, and then it and all later windows are abandoned.  MODE_SWEPT_KEEP: the EARLIEST window with such pairs is
// not blocked by them: they are KEPT (WarpShared::surv) so that the caller can run the exact predicate on them at
// every step of that window.  Every triangle pair that intersects at some step of a window is among the pairs kept
// for it (it lies within the window's radius of intersecting at the mid pose), so those tests decide each step of
// the window without another traversal.  The windows behind the keeping one are blocked (the walk will be at a
// contact or needs a fresh traversal there anyway); the keeping window itself blocks only when it keeps more than
// C.surv_limit pairs.
// BOTH (small validity calls only): a pair of two inner nodes is replaced by its four child-by-child pairs, so one
// round descends a level in BOTH trees.  More box tests, about half as many rounds: what the latency of one pose
// is made of.  Pruning stays conservative either way, so the labels do not depend on it.
template <int MODE, bool BOTH = false>
__device__ __forceinline__ unsigned traverse_tile(const KParams& P, WarpShared& S, WarpCtx& C, unsigned alive,
                                                  int lane, int state_base = 0, ContactRec* recs = nullptr,
                                                  unsigned long long* rec_count = nullptr,
                                                  unsigned long long rec_cap = 0, int rmode = MODE_SEGMENT) {
  const bool m_keep = MODE == MODE_RT ? rmode == MODE_SWEPT_KEEP : MODE == MODE_SWEPT_KEEP;
  const bool m_swept = m_keep || MODE == MODE_SWEPT;
  const bool m_segment = MODE == MODE_RT ? rmode == MODE_SEGMENT : MODE == MODE_SEGMENT;
  const unsigned ltmask = (1u << lane) - 1u;
  unsigned dead = ~alive;
  unsigned hits = 0;
  int top = __popc(alive);
  int nleaf = 0;
  if (alive >> lane & 1u) S.stack[__popc(alive & ltmask)] = make_entry(lane, 0, 0);
  __syncwarp();
  while (true) {
    if (top == 0 && C.nspill > 0) {  // refill from the global spill area
      int take = min(C.nspill, SPILL_CHUNK);
      for (int i = lane; i < take; i += 32) S.stack[i] = C.spill[C.nspill - take + i];
      C.nspill -= take;
      top = take;
      __syncwarp();
    }
    if (top == 0 || nleaf >= 32) {
      if (nleaf == 0) break;
      // ---- leaf phase: up to 32 triangle pairs
      int cnt = min(32, nleaf);
      bool act = lane < cnt;
      unsigned long long e = act ? S.leafq[nleaf - 1 - lane] : 0ull;
      nleaf -= cnt;
      unsigned pose = entry_pose(e);
      act = act && !((dead >> pose) & 1u);
      bool hit = false;
      if (act) {
        C.n_leaf++;
        const double* tp = P.env.tris + (size_t)9 * entry_n1(e);
        const double* tq = P.rob.tris + (size_t)9 * entry_n2(e);
        Tri6<double> td;
        load_pair_exact(tp, tq, &S.Rd[0][pose], td);
        if (m_swept) {
          Tri6<float> tf;
#pragma unroll
          for (int k = 0; k < 3; k++) {
            tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
            tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
          }
          unsigned best;
          hit = !tri_gap_clear_f(tf, __double2float_ru(S.tau[pose]), best);
          if (m_keep) e |= (unsigned long long)best << 28;   // triangle ids stay below 2^27
        } else {
          const int r = tri_pair_intersect(td);
          hit = (r & 1) != 0;
          C.n_fp64 += (unsigned)r >> 1;
        }
        if (MODE == MODE_CONTACTS && hit) {
          unsigned long long slot = atomicAdd(rec_count, 1ull);
          if (slot < rec_cap) {
            ContactRec r;
            r.state = state_base + (int)pose;
            r.env_tri = (int)entry_n1(e);
            r.rob_tri = (int)entry_n2(e);
            r.mult = (int)contact_multiplicity_dev(tp, tq, &S.Rd[0][pose]);
            recs[slot] = r;
          }
        }
      }
      unsigned hm;
      if (m_keep) {
        hm = 0;
        if (__any_sync(FULL, hit)) {
          const int pmin = __reduce_min_sync(FULL, hit ? (int)pose : 32);
          if (pmin < C.keep_win) { C.keep_win = pmin; C.nsurv = 0; }   // an earlier window takes over the list
          const bool mine = hit && (int)pose == C.keep_win;
          const unsigned mm = __ballot_sync(FULL, mine);
          if (mine) {
            const int pos = C.nsurv + __popc(mm & ltmask);
            if (pos < C.surv_limit) S.surv[pos] = e;
          }
          C.nsurv += __popc(mm);
          if (C.nsurv > C.surv_limit) hm = 1u << C.keep_win;       // too many pairs to test one by one
          else if (C.keep_win < 31) hm = 2u << C.keep_win;
        }
      } else {
        hm = __reduce_or_sync(FULL, hit ? (1u << pose) : 0u);
      }
      hits |= hm;
      if (MODE == MODE_VALID) dead |= hm;
      else if (m_segment || m_swept) { if (hm) dead |= FULL << (__ffs(hm) - 1); }
      __syncwarp();
      if (m_swept && (dead & alive) == alive) { C.nspill = 0; break; }   // every window abandoned
      continue;
    }
    if (top > STACK_CAP - (BOTH ? 96 : 32)) {  // make room: move the oldest entries to global memory
      if (C.nspill + SPILL_CHUNK > C.spill_cap) {
        if (m_swept) { hits = alive; C.nspill = 0; break; }   // no room: nothing proven, every window blocked
        C.overflow = true;
        break;
      }
      for (int i = lane; i < SPILL_CHUNK; i += 32) C.spill[C.nspill + i] = S.stack[i];
      C.nspill += SPILL_CHUNK;
      C.n_spill++;
      int rem = top - SPILL_CHUNK;
      for (int i0 = 0; i0 < rem; i0 += 32) {
        unsigned long long v = (i0 + lane < rem) ? S.stack[SPILL_CHUNK + i0 + lane] : 0ull;
        __syncwarp();
        if (i0 + lane < rem) S.stack[i0 + lane] = v;
      }
      top = rem;
      __syncwarp();
    }
    // ---- node phase: up to 32 box pairs
    int cnt = min(32, top);
    bool act = lane < cnt;
    unsigned long long e = act ? S.stack[top - 1 - lane] : 0ull;
    top -= cnt;
    __syncwarp();
    unsigned pose = entry_pose(e), n1 = entry_n1(e), n2 = entry_n2(e);
    act = act && !((dead >> pose) & 1u);
    bool leafpair = false, expand = false, four = false;
    unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0, lf = 0;
    if (act) {
      C.n_bv++;
      const float4* na = P.env.nodes + (size_t)4 * n1;
      const float4* nb = P.rob.nodes + (size_t)4 * n2;
      float4 a0 = ldg4(na), a1 = ldg4(na + 1), a2 = ldg4(na + 2), a3 = ldg4(na + 3);
      float4 b0 = ldg4(nb), b1 = ldg4(nb + 1), b2 = ldg4(nb + 2), b3 = ldg4(nb + 3);
      if (obb_overlap_f32(a0, a1, a2, a3, b0, b1, b2, b3, &S.Rf[0][pose], S.eta[pose])) {
        int l1 = __float_as_int(a3.w), l2 = __float_as_int(b3.w);
        if (l1 < 0 && l2 < 0) {
          leafpair = true;
          lf = make_entry(pose, (unsigned)(-(l1 + 1)), (unsigned)(-(l2 + 1)));
        } else {
          expand = true;
          // descent rule: split the box of larger volume (FCL compares extent^2; labels do not depend on it)
          float s1 = a3.x * a3.y * a3.z;
          float s2 = b3.x * b3.y * b3.z;
          if (BOTH && l1 >= 0 && l2 >= 0) {
            four = true;
            c0 = make_entry(pose, (unsigned)l1, (unsigned)l2);
            c1 = make_entry(pose, (unsigned)l1 + 1u, (unsigned)l2);
            c2 = make_entry(pose, (unsigned)l1, (unsigned)l2 + 1u);
            c3 = make_entry(pose, (unsigned)l1 + 1u, (unsigned)l2 + 1u);
          } else if (l2 < 0 || (l1 >= 0 && s1 > s2)) {  // descend the env node
            c0 = make_entry(pose, (unsigned)l1, n2);
            c1 = make_entry(pose, (unsigned)l1 + 1u, n2);
          } else {
            c0 = make_entry(pose, n1, (unsigned)l2);
            c1 = make_entry(pose, n1, (unsigned)l2 + 1u);
          }
        }
      }
    }
    unsigned bl = __ballot_sync(FULL, leafpair);
    if (leafpair) S.leafq[nleaf + __popc(bl & ltmask)] = lf;
    nleaf += __popc(bl);
    unsigned be = __ballot_sync(FULL, expand);
    unsigned b4 = BOTH ? __ballot_sync(FULL, four) : 0u;
    if (expand) {
      int off = top + 2 * (__popc(be & ltmask) + __popc(b4 & ltmask));
      S.stack[off] = c0;
      S.stack[off + 1] = c1;
      if (BOTH && four) {
        S.stack[off + 2] = c2;
        S.stack[off + 3] = c3;
      }
    }
    top += 2 * (__popc(be) + __popc(b4));
    __syncwarp();
  }
  return hits;
}

// Fill pose slot `lane` from a unit state.
__device__ __forceinline__ void set_pose(const KParams& P, WarpShared& S, int lane, const double* q) {
  Rigid rob, rel;
  state_to_rigid(q, rob);
  relative_rigid(P.env_tf, rob, rel);
#pragma unroll
  for (int k = 0; k < 9; k++) { S.Rd[k][lane] = rel.R[k]; S.Rf[k][lane] = (float)rel.R[k]; }
#pragma unroll
  for (int k = 0; k < 3; k++) { S.Rd[9 + k][lane] = rel.t[k]; S.Rf[9 + k][lane] = (float)rel.t[k]; }
  S.eta[lane] = 2e-5f * (1.0f + fabsf((float)rel.t[0]) + fabsf((float)rel.t[1]) + fabsf((float)rel.t[2]));
  // The reference does not renormalise the quaternion (lib/osr/osr_state.cc:16-25): for a non-unit one the map M
  // above is 1 + |q|^2 (Rot - 1), not a rotation, and the 15-axis box test assumes one.  Boxes are then placed with
  // the rotation of the NORMALISED quaternion and grown by |M - Rot|_F x (largest robot vertex radius), which bounds
  // how far any robot point lands from where the boxes put it; the triangle tests keep M (S.Rd).  Same rule as
  // make_box_pose of the lane-owned kernel (qnode.h).
  const double n2 = q[3] * q[3] + q[4] * q[4] + q[5] * q[5] + q[6] * q[6];
  if (fabs(n2 - 1.0) > 1e-9) {
    double w = q[3], x = q[4], y = q[5], z = q[6];
    quat_normalize(w, x, y, z);
    Rigid rn, reln;
    quat_to_rot(w, x, y, z, rn.R);
    rn.t[0] = q[0]; rn.t[1] = q[1]; rn.t[2] = q[2];
    relative_rigid(P.env_tf, rn, reln);
    double dev = 0.0;
#pragma unroll
    for (int k = 0; k < 9; k++) {
      const double d = rel.R[k] - reln.R[k];
      dev += d * d;
      S.Rf[k][lane] = (float)reln.R[k];
    }
    S.eta[lane] = __fadd_ru(S.eta[lane], __double2float_ru(1.0001 * sqrt(dev) * P.rob_radius));
  }
}

__device__ __forceinline__ void ctx_init(const KParams& P, WarpCtx& C) {
  int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  C.spill = P.spill + (size_t)gw * P.spill_cap;
  C.spill_cap = P.spill_cap;
  C.nspill = 0;
  C.n_bv = C.n_leaf = C.n_fp64 = C.n_spill = 0;
  C.overflow = false;
  C.nsurv = 0;
  C.surv_limit = 0;
  C.keep_win = 32;
}
__device__ __forceinline__ void ctx_flush(const KParams& P, WarpCtx& C, unsigned long long poses, int lane) {
  unsigned bv = __reduce_add_sync(FULL, C.n_bv), lf = __reduce_add_sync(FULL, C.n_leaf),
           f64 = __reduce_add_sync(FULL, C.n_fp64);
  if (lane == 0) {
    atomicAdd(&P.counters->bv, (unsigned long long)bv);
    atomicAdd(&P.counters->leaf, (unsigned long long)lf);
    atomicAdd(&P.counters->fp64, (unsigned long long)f64);
    atomicAdd(&P.counters->poses, poses);
    if (C.n_spill) atomicAdd(&P.counters->spills, (unsigned long long)C.n_spill);
    if (C.overflow) atomicExch(P.error_flag, 1);
  }
}

constexpr int WARPS_PER_BLOCK = 4;

// ---------------------------------------------------------------------------------------------
// K1: batched validity.  out[i] = 1 iff state i is collision free (UnitWorld::isValid,
// /root/reference/lib/osr/unit_world.cc:215-247).
// ---------------------------------------------------------------------------------------------
// ppt = poses per tile (32; 1 for small host calls: every warp then puts its 32 lanes on ONE pose's node pairs,
// which is what the round trip of a handful of states is made of)
template <bool BOTH>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32) k_is_valid(KParams P, const double* __restrict__ q,
                                                                    long long n, uint8_t* __restrict__ out, int ppt) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpShared& S = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  WarpCtx C;
  ctx_init(P, C);
  const long long ntiles = (n + ppt - 1) / ppt;
  const bool fixed = P.work_counter == nullptr;
  unsigned long long nposes = 0;
  while (true) {
    unsigned long long tile = 0;
    if (fixed) {   // small host calls: a warp per tile, no counter to reset before the launch
      tile = (unsigned long long)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
    } else {
      if (lane == 0) tile = atomicAdd(P.work_counter, 1ull);
      tile = __shfl_sync(FULL, tile, 0);
    }
    if ((long long)tile >= ntiles) break;
    long long i = (long long)tile * ppt + lane;
    bool have = lane < ppt && i < n;
    if (have) {
      double s[7];
#pragma unroll
      for (int k = 0; k < 7; k++) s[k] = __ldg(q + 7 * i + k);
      set_pose(P, S, lane, s);
    }
    unsigned alive = __ballot_sync(FULL, have);
    nposes += __popc(alive);
    __syncwarp();
    unsigned hits = traverse_tile<MODE_VALID, BOTH>(P, S, C, alive, lane);
    if (have) out[i] = (hits >> lane & 1u) ? 0 : 1;
    __syncwarp();
    if (C.overflow || fixed) break;
  }
  ctx_flush(P, C, nposes, lane);
}

// ---------------------------------------------------------------------------------------------
// K2: segment walk / first contact (UnitWorld::transitStateToWithContact,
// /root/reference/lib/osr/unit_world.cc:405-474; isValidTransition :477-486).
// One warp per segment; each tile is 32 consecutive steps of the reference's linear walk
//   delta_1 = vd;  delta_{k+1} = delta_k < dist ? min(dist, delta_k + vd) : delta_k + vd
// (the running fp64 sum is reproduced add by add so every tau is bit-identical), and the FIRST
// colliding step is returned -- validity along a segment is not monotone, so no bisection.
// ---------------------------------------------------------------------------------------------
struct SegOut {
  double* last_free;   // [n][7] or null
  double* first_hit;   // [n][7] or null
  uint8_t* is_free;    // [n] or null
  double* free_tau;    // [n] or null
  double* touch_tau;   // [n] or null
  long long* steps;    // [n] or null
  // indexed segments (k-NN edges validated where they are, osr_knn_validate_edges): segment e runs from state
  // e / idx_k to state to_idx[e] of ONE state array (`from` == `to`); to_idx[e] < 0 = no such edge, label 0
  const int* to_idx;   // [n] or null
  int idx_k;
};

// ---------------------------------------------------------------------------------------------
// Swept windows that keep pairs (MODE_SWEPT_KEEP): helpers of k_transit.
// ---------------------------------------------------------------------------------------------
// Exact SAT-17 of the ncur kept pairs (S.surv) at each of the 2^cs_log2 step poses of the current chunk (pose slot p =
// step p; `alive` = slots that hold a step).  Work item i = pair (i >> cs_log2) at pose (i & (cs - 1)), 32 items per
// pass, so a 4-step window still fills the warp.  Returns the mask of steps with an intersecting pair; steps behind
// a known hit are not tested (only the first colliding step matters, unit_world.cc:452-458).
__device__ __forceinline__ unsigned test_kept_pairs(const KParams& P, WarpShared& S, WarpCtx& C, int ncur, int cs_log2,
                                                    unsigned alive, int lane) {
  unsigned hits = 0;
  const int total = ncur << cs_log2;
  const unsigned cmask = (1u << cs_log2) - 1u;
  for (int base = 0; base < total; base += 32) {
    const int i = base + lane;
    const unsigned pose = (unsigned)i & cmask;
    const bool act = i < total && (alive >> pose & 1u) && !(hits & (0xffffffffu >> (31u - pose)));
    bool hit = false;
    if (act) {
      C.n_leaf++;
      const unsigned long long e = S.surv[i >> cs_log2];
      const unsigned axis = ((unsigned)e >> 28) & 15u;
      Tri6<double> td;
      load_pair_exact(P.env.tris + (size_t)9 * (entry_n1(e) & 0x0fffffffu), P.rob.tris + (size_t)9 * entry_n2(e),
                      &S.Rd[0][pose], td);
      // the axis that separated the pair best at the window's mid pose usually still separates it at this step
      Tri6<float> tf;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
        tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
      }
      if (!(axis < 11u && sat17_axis_separates(tf, axis))) {
        const int r = tri_pair_intersect(td);
        hit = (r & 1) != 0;
        C.n_fp64 += (unsigned)r >> 1;
      }
    }
    hits |= __reduce_or_sync(FULL, hit ? (1u << pose) : 0u);
  }
  return hits;
}

// Pose slot `slot` := the walk's state at tau (osr::interpolate of the segment, prepared in S.plan) as the traversal
// wants it (robot in the environment frame, fp64 + fp32 + box slack).  One copy per kernel.
__device__ __noinline__ void seg_pose_at(int slot, double tau) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpShared& S = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
  double s[7];
  slerp_at(S.plan, tau, s);
  Rigid env, rob, rel;
#pragma unroll
  for (int k = 0; k < 9; k++) env.R[k] = S.envtf[k];
#pragma unroll
  for (int k = 0; k < 3; k++) env.t[k] = S.envtf[9 + k];
  state_to_rigid(s, rob);
  relative_rigid(env, rob, rel);
#pragma unroll
  for (int k = 0; k < 9; k++) { S.Rd[k][slot] = rel.R[k]; S.Rf[k][slot] = (float)rel.R[k]; }
#pragma unroll
  for (int k = 0; k < 3; k++) { S.Rd[9 + k][slot] = rel.t[k]; S.Rf[9 + k][slot] = (float)rel.t[k]; }
  S.eta[slot] = 2e-5f * (1.0f + fabsf((float)rel.t[0]) + fabsf((float)rel.t[1]) + fabsf((float)rel.t[2]));
}

// One output state of a segment, written by lanes 0..6: which = 0 the `from` state, 1 the `to` state, 2 the walk's
// state at tau.
__device__ __noinline__ void seg_write_state(double* __restrict__ dst, int which, double tau, int lane) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpShared& S = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
  double s[7];
#pragma unroll
  for (int k = 0; k < 7; k++) s[k] = 0.0;
  if (which == 2) slerp_at(S.plan, tau, s);
  double v = 0.0;
#pragma unroll
  for (int k = 0; k < 7; k++) {
    const double x = which == 0 ? S.seg[k] : (which == 1 ? S.seg[7 + k] : s[k]);
    if (k == lane) v = x;
  }
  if (lane < 7) dst[lane] = v;
}

// transition_mode: vd := min(dist/2, vd) (isValidTransition)
template <int MINB>  // resident blocks per SM the register allocation is capped for
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, MINB)
    k_transit(KParams P, const double* __restrict__ from, const double* __restrict__ to, long long n,
              const double* __restrict__ vd_arr, long long vd_count, double vd_scalar, int transition_mode,
              SegOut O) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpShared& S = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  WarpCtx C;
  ctx_init(P, C);
  if (lane < 9) S.envtf[lane] = P.env_tf.R[lane];
  else if (lane < 12) S.envtf[lane] = P.env_tf.t[lane - 9];
  unsigned long long nposes = 0;
  enum { ST_FROM = 0, ST_TO = 1, ST_TAU = 2 };
  enum { ACT_SWEEP = 0, ACT_EXACT = 1, ACT_LISTED = 2 };
  while (true) {
    unsigned long long seg = 0;
    if (lane == 0) seg = atomicAdd(P.work_counter, 1ull);
    seg = __shfl_sync(FULL, seg, 0);
    if ((long long)seg >= n) break;
    __syncwarp();
    long long fi = (long long)seg, ti = (long long)seg;
    if (O.to_idx) {
      fi = div53((long long)seg, (long long)O.idx_k);
      ti = __ldg(O.to_idx + seg);
      if (ti < 0) {
        if (lane == 0) {
          if (O.is_free) O.is_free[seg] = 0;
          if (O.free_tau) O.free_tau[seg] = 0.0;
          if (O.touch_tau) O.touch_tau[seg] = 0.0;
          if (O.steps) O.steps[seg] = 0;
        }
        continue;
      }
    }
    if (lane < 7) { S.seg[lane] = __ldg(from + 7 * fi + lane); S.seg[7 + lane] = __ldg(to + 7 * ti + lane); }
    __syncwarp();
    double vd = vd_arr ? __ldg(vd_arr + (vd_count > 1 ? seg : 0)) : vd_scalar;
    double dist, trans;
    {
      double a[7], b[7];
#pragma unroll
      for (int k = 0; k < 7; k++) { a[k] = S.seg[k]; b[k] = S.seg[7 + k]; }
      {   // once per segment; every pose of the walk reads it.  Computed by every lane (no per-lane branch around
          // the libm calls), stored by one.
        SlerpPlan pl;
        slerp_prepare(a, b, pl);
        if (lane == 0) S.plan = pl;
        __syncwarp();
      }
      dist = distance(a, b);
      const double tdx = a[0] - b[0], tdy = a[1] - b[1], tdz = a[2] - b[2];
      trans = sqrt(tdx * tdx + tdy * tdy + tdz * tdz);
    }
    if (transition_mode) vd = fmin(dist / 2.0, vd);
    // results (warp-uniform).  The two output states are produced at the end from (which, tau): last_free is `from`
    // until a step has been proven free, first_hit is `to` unless the walk says otherwise.
    bool free_flag = false;
    double free_tau = 0.0, touch_tau = 1.0;
    long long steps = 0;
    int lf_which = ST_FROM, fh_which = ST_TO;
    double lf_tau = 0.0, fh_tau = 0.0;
    // A NaN distance or verify_delta fails both `verify_delta >= dist` and `delta <= dist` of the reference
    // (unit_world.cc:443-447): its loop body never runs and it returns (from, from, false, 0, 0).  With
    // verify_delta <= 0 the reference's walk never ends; here that is an argument error (error flag 3, same tuple).
    const bool nan_walk = !(dist == dist) || !(vd == vd);
    const bool bad_vd = !nan_walk && !(vd >= dist) && !(vd > 0.0);
    if (nan_walk || bad_vd) {
      fh_which = ST_FROM;
      touch_tau = 0.0;
      if (bad_vd && lane == 0) atomicExch(P.error_flag, 3);
    } else if (!(vd >= dist)) {
      const double inv_dist = 1.0 / dist;
      double carry = 0.0;       // delta of the last step of the previous tile (0 before the first)
      double prev_tau = 0.0;    // tau of that step
      long long k0 = 0;         // steps completed before this tile
      bool done = false;
      // ---- swept windows -------------------------------------------------------------------------------------
      // A window is a run of consecutive steps.  Between two values of tau a robot point moves at most
      // |tau - tau'| * rate, rate = |dt| + (rotation angle) x (largest vertex radius): translation is linear in
      // tau, the slerp arc has constant angular velocity, and dist = |dt| + angle.  If NO pair of leaf boxes
      // overlaps at the window's mid pose with the boxes grown by the window's radius, no triangle pair can
      // intersect at any step of the window (triangles lie inside their leaf boxes), so the reference's per-step
      // checks (unit_world.cc:445-470) all say "free" and the walk advances over the window without traversing
      // its steps.  A window that does not clear is split, decided from the pairs it kept, or walked exactly.
      // One traversal tests up to 32 consecutive windows (one per pose slot, lane-parallel over box pairs).
      const double rate = (trans + (fmax(dist - trans, 0.0) + 1e-12) * P.rob_radius) * (1.0 + 1e-9);
      // steps per window (a power of two >= 4): never wider than the window whose radius reaches sweep_rmax (a
      // robot grown by more is not expected to clear anything), starts where the radius is sweep_rstart,
      // doubles while everything clears and halves on a blocked window; below 4 steps a tile is walked exactly
      const double step_radius = 0.5 * vd * inv_dist * rate;
      const int ws_cap = (int)fmin(32.0 * (double)P.sweep_tiles, P.sweep_rmax / step_radius);
      int ws = 0;
      if (ws_cap >= 4) {
        ws = 4;
        while (2 * ws <= ws_cap && (double)(2 * ws) * step_radius <= P.sweep_rstart) ws *= 2;
      }
      int nb = P.sweep_nb0;        // windows per traversal
      int fails = 0;
      bool exact_next = false;     // the next tile is known not to clear
      // What the last swept traversal left to consume, in walk order: pend_clear windows of pws steps that kept no
      // pair (free: stepped over), then, if pend_keep, one window whose steps are decided by the exact predicate on
      // exactly the pairs it kept (res_left = its steps still to decide, ncur = its pairs in S.surv).
      int pend_clear = 0, pws = 0, res_left = 0, ncur = 0;
      bool pend_keep = false;
      while (!done) {
        int act, exact_n = 0;
        if (res_left > 0) {
          act = ACT_LISTED;
          exact_n = min(32, res_left);
        } else if (pend_clear > 0) {
          const int ns = pend_clear * pws;
          pend_clear = 0;
          double d = carry, last = carry;
          const int valid = advance_chain(d, vd, dist, ns, last);   // the reference's running sum, rounding included
          carry = d;
          nposes += valid;
          k0 += valid;
          if (valid > 0) prev_tau = last * inv_dist;
          if (valid < ns) {  // walked past dist: the whole segment is free
            free_tau = prev_tau;
            touch_tau = 0.0;  // the reference returns its shadowed outer `tau` (unit_world.cc:452-455,473)
            lf_which = fh_which = ST_TAU;
            lf_tau = fh_tau = prev_tau;
            steps = k0;
            free_flag = true;
            done = true;
          }
          continue;
        } else if (pend_keep) {
          pend_keep = false;
          ncur = C.nsurv;
          res_left = pws;
          continue;
        } else if (ws > 0 && !exact_next && (fails < 1 || ((k0 >> 5) & 7) == 0)) {
          act = ACT_SWEEP;
        } else {
          // a tile entered because four steps did not clear is short: the contact is expected within them, and every
          // pose traversed beyond the first hit is wasted
          act = ACT_EXACT;
          exact_n = exact_next ? P.sweep_short_tile : 32;
          exact_next = false;
        }
        // ---- the poses of this pass: mid poses of consecutive windows, or consecutive steps
        bool have;
        double my_tau, my_radius = 0.0;
        if (act == ACT_SWEEP) {
          // window `lane` holds the steps whose delta lies in (carry + lane w, carry + (lane + 1) w], w = ws vd,
          // up to the rounding of the running sum (< 1e-6 relative to w for any vd/dist >= 1e-9)
          const double w = (double)ws * vd;
          const double lo = carry + (double)lane * w * (1.0 - 1e-6);
          const double hi = carry + (double)(lane + 1) * w * (1.0 + 1e-6);
          have = lane < nb && lo <= dist;
          const double tau_lo = lo * inv_dist, tau_hi = fmin(1.0, hi * inv_dist * (1.0 + 1e-9));
          my_radius = 0.5 * (tau_hi - tau_lo) * rate * (1.0 + 1e-9) + 1e-6;
          my_tau = 0.5 * (tau_lo + tau_hi);
        } else {
          // step deltas of this tile: lane j takes the (j+1)-th value after `carry`.  The chain of fp64 adds is
          // inherently serial (the reference's running sum, rounding included), so all lanes run it; it is kept to
          // add / compare / select per step: with carry = 0 the first value is 0 + vd = vd exactly, and since the
          // sequence increases strictly, "every delta up to mine is <= dist" is just mine <= dist.
          double d = carry, mine = 0.0;
#pragma unroll 1   // (rolled: the chain is latency bound either way, and unrolled it was 525 instructions of cold code)
          for (int j = 0; j < 32; j++) {
            const double t = d + vd;
            d = (d < dist) ? fmin(dist, t) : t;
            if (j == lane) mine = d;
          }
          have = mine <= dist && lane < exact_n;
          carry = __shfl_sync(FULL, mine, exact_n - 1);  // delta of the last slot (only used when all slots were valid)
          my_tau = mine * inv_dist;
        }
        // Every lane fills its own slot, also the lanes without a step / window (their slot is not part of `alive`):
        // a per-lane branch around the out-of-line call here costs the traversal below its convergence -- the two
        // groups of lanes then run every round of it separately (1.75x the warp instructions, measured).
        seg_pose_at(lane, have ? my_tau : 0.0);
        if (act == ACT_SWEEP) {
          // half of the radius on each box: every separating axis gains at least radius x |axis|
          S.eta[lane] = __fadd_ru(S.eta[lane], __double2float_ru(0.5 * my_radius));
          S.tau[lane] = my_radius;   // triangle-level test of the window
        } else {
          S.tau[lane] = my_tau;
        }
        const unsigned alive = __ballot_sync(FULL, have);   // a prefix of the lanes
        __syncwarp();
        if (act == ACT_SWEEP) {
          if (alive == 0) { exact_next = true; continue; }   // cannot happen for finite inputs; never spin
          C.nsurv = 0;
          C.keep_win = 32;
          // pairs are kept only by windows of small radius: a wide window near a contact keeps so many pairs that
          // walking its steps exactly is cheaper than listing and testing them
          C.surv_limit = (double)ws * step_radius <= P.sweep_keep_rmax ? min(SURV_CAP, P.sweep_resolve / ws) : 0;
        }
        unsigned hits;
        if (act == ACT_LISTED) {
          hits = test_kept_pairs(P, S, C, ncur, 31 - __clz(exact_n), alive, lane);
          res_left -= exact_n;
        } else {
          hits = traverse_tile<MODE_RT>(P, S, C, alive, lane, 0, nullptr, nullptr, 0,
                                        act == ACT_SWEEP ? MODE_SWEPT_KEEP : MODE_SEGMENT);
        }
        if (C.overflow) { done = true; break; }
        __syncwarp();
        if (act == ACT_SWEEP) {
          const int ntried = __popc(alive);
          const int first_blocked = hits ? min(ntried, __ffs(hits) - 1) : ntried;
          pend_keep = C.keep_win < first_blocked;      // that window kept at most surv_limit pairs, all of them listed
          pend_clear = pend_keep ? C.keep_win : first_blocked;
          pws = ws;
          if (pend_keep) {           // near a contact: one window per traversal until windows clear again
            fails = 0;
            nb = 1;
          } else if (first_blocked == ntried) {   // everything cleared: widen
            fails = 0;
            if (nb < 32) nb *= 2;
            else if (2 * ws <= ws_cap) ws *= 2;
          } else if (ws > 4) {       // split the blocked window
            ws >>= 1;
            nb = 2;
          } else {                   // four steps with too many pairs to list: walk the tile step by step
            exact_next = true;
            nb = 2;
            if (first_blocked == 0) fails++;
          }
          continue;
        }
        const int nalive = __popc(alive);
        if (hits) {
          const int j = __ffs(hits) - 1;
          nposes += j + 1;
          touch_tau = S.tau[j];
          free_tau = j > 0 ? S.tau[j - 1] : prev_tau;
          fh_which = ST_TAU; fh_tau = touch_tau;
          if (j > 0 || k0 > 0) { lf_which = ST_TAU; lf_tau = free_tau; }
          steps = k0 + j + 1;
          free_flag = false;
          done = true;
        } else {
          nposes += nalive;
          k0 += nalive;
          if (nalive > 0) prev_tau = S.tau[nalive - 1];
          if (nalive < exact_n) {  // walked past dist: the whole segment is free
            free_tau = prev_tau;
            touch_tau = 0.0;  // the reference returns its shadowed outer `tau` (unit_world.cc:452-455,473)
            lf_which = fh_which = ST_TAU;
            lf_tau = fh_tau = prev_tau;
            steps = k0;
            free_flag = true;
            done = true;
          }
        }
        __syncwarp();
      }
    }
    if (lane == 0) {
      if (O.is_free) O.is_free[seg] = free_flag ? 1 : 0;
      if (O.free_tau) O.free_tau[seg] = free_tau;
      if (O.touch_tau) O.touch_tau[seg] = touch_tau;
      if (O.steps) O.steps[seg] = steps;
    }
    if (O.last_free) seg_write_state(O.last_free + 7 * seg, lf_which, lf_tau, lane);
    if (O.first_hit) seg_write_state(O.first_hit + 7 * seg, fh_which, fh_tau, lane);
    if (C.overflow) break;
  }
  ctx_flush(P, C, nposes, lane);
}

// ---------------------------------------------------------------------------------------------
// K2b: OMPL motion validation on OMPL-frame states: what OmplDriver::validateMotionPairs
// (/root/reference/lib/pyse3ompl/ompldriver.cc:592-621) asks of ompl::base::DiscreteMotionValidator::checkMotion
// (third-party/ompl.app/ompl/src/ompl/base/src/DiscreteMotionValidator.cpp:96-150): s2 is valid and so is every
// interior point interpolate(s1, s2, m/nd), m = 1..nd-1 (checkMotion visits them in bisection order; the boolean
// does not depend on the order), with
//   nd = max over the two sub-spaces of ceil(distance / (resolution * maximum extent))
//        (StateSpace::validSegmentCount, CompoundStateSpace takes the maximum; SO(3) extent pi/2, distance
//         acos|q1.q2| with the 1e-9 clamp of SO3StateSpace.cpp:252-260),
//   interpolate = RealVectorStateSpace (from + (to - from) t) x SO3StateSpace::interpolate (:287-316).
// States are rows x y z qx qy qz qw (StateSpace::copyFromEigen3).  Each interpolated OMPL state is mapped to the
// unit frame like UnitWorld::translateOMPLStateToUnitState (unit_world.cc:689-711) with identity perturbation:
// t_unit = scale * t - Rot(q) * dc,  dc = (ompl centre - robot centre) * scale.
// One warp per pair, 32 consecutive sample points per tile, the same warp-shared traversal as K2.
// ---------------------------------------------------------------------------------------------
struct OmplFrame {
  double scale;       // scene_scale (float widened)
  double dc[3];       // (ompl centre - centre) * scale
  double seg_trans;   // resolution * maximum extent of the translation bounds
  double seg_rot;     // resolution * pi/2
  double lo[3], hi[3];  // RealVectorBounds of the translation (OmplDriver::setBounds, ompldriver.cc:680-685)
};

// SpaceInformation::satisfiesBounds of an OMPL-frame SE(3) row x y z qx qy qz qw, the first half of
// FCLStateValidityChecker::isValid (third-party/ompl.app/src/omplapp/geometry/detail/FCLStateValidityChecker.h:103):
// RealVectorStateSpace::satisfiesBounds (spaces/src/RealVectorStateSpace.cpp:200-208, machine-epsilon slack) and
// SO3StateSpace::satisfiesBounds (spaces/src/SO3StateSpace.cpp:175-179, 211-214: |norm - 1| < 1e-9).
OSR_HD bool ompl_satisfies_bounds(const double* lo, const double* hi, const double* o) {
  const double eps = 2.220446049250313e-16;
  for (int k = 0; k < 3; k++)
    if (o[k] - eps > hi[k] || o[k] + eps < lo[k]) return false;
  const double n2 = o[3] * o[3] + o[4] * o[4] + o[5] * o[5] + o[6] * o[6];
  const double nrm = (fabs(n2 - 1.0) > eps) ? sqrt(n2) : 1.0;
  return fabs(nrm - 1.0) < 1e-9;
}
// Both end points so far inside the bounds (and so close to unit norm) that no interpolated point can fail
// satisfiesBounds whatever the rounding of the lerp / slerp: such motions skip the per-point bounds test.
OSR_HD bool ompl_safely_inside(const double* lo, const double* hi, const double* o) {
  for (int k = 0; k < 3; k++) {
    const double m = 1e-12 * (1.0 + fabs(lo[k]) + fabs(hi[k]));
    if (!(o[k] >= lo[k] + m && o[k] <= hi[k] - m)) return false;
  }
  const double n2 = o[3] * o[3] + o[4] * o[4] + o[5] * o[5] + o[6] * o[6];
  return fabs(n2 - 1.0) < 1e-10;
}

__device__ __forceinline__ double ompl_arc_length(const double* a, const double* b) {
  double dq = fabs(a[3] * b[3] + a[4] * b[4] + a[5] * b[5] + a[6] * b[6]);
  if (dq > 1.0 - 1e-9) return 0.0;
  return acos(dq);
}

// OMPL-frame interpolation at t (RealVectorStateSpace lerp x SO3StateSpace::interpolate): row x y z qx qy qz qw
__device__ __forceinline__ void ompl_point(const double* a, const double* b, double theta, double t, double* o) {
#pragma unroll
  for (int k = 0; k < 3; k++) o[k] = a[k] + (b[k] - a[k]) * t;
  if (theta > 2.220446049250313e-16) {
    double d = 1.0 / sin(theta), s0 = sin((1.0 - t) * theta), s1 = sin(t * theta);
    double dq = a[3] * b[3] + a[4] * b[4] + a[5] * b[5] + a[6] * b[6];
    if (dq < 0) s1 = -s1;
#pragma unroll
    for (int k = 0; k < 4; k++) o[3 + k] = (a[3 + k] * s0 + b[3 + k] * s1) * d;
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++) o[3 + k] = a[3 + k];
  }
}
// the unit-frame state (w-first) of an OMPL-frame row (translateOMPLStateToUnitState, identity perturbation)
__device__ __forceinline__ void ompl_row_to_unit(const OmplFrame& F, const double* o, double* s) {
  double w = o[6], x = o[3], y = o[4], z = o[5];
  quat_normalize(w, x, y, z);
  // Eigen quaternion * vector: v + w (2 u x v) + u x (2 u x v)
  double uvx = 2.0 * (y * F.dc[2] - z * F.dc[1]), uvy = 2.0 * (z * F.dc[0] - x * F.dc[2]),
         uvz = 2.0 * (x * F.dc[1] - y * F.dc[0]);
  double rx = F.dc[0] + w * uvx + (y * uvz - z * uvy), ry = F.dc[1] + w * uvy + (z * uvx - x * uvz),
         rz = F.dc[2] + w * uvz + (x * uvy - y * uvx);
  s[0] = F.scale * o[0] - rx; s[1] = F.scale * o[1] - ry; s[2] = F.scale * o[2] - rz;
  s[3] = w; s[4] = x; s[5] = y; s[6] = z;
}
__device__ __forceinline__ void ompl_point_to_unit(const OmplFrame& F, const double* a, const double* b, double theta,
                                                   double t, double* s) {
  double o[7];
  ompl_point(a, b, theta, t, o);
  ompl_row_to_unit(F, o, s);
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, 3)
    k_ompl_motion(KParams P, OmplFrame F, const double* __restrict__ q0, const double* __restrict__ q1, long long n,
                  uint8_t* __restrict__ out, long long* __restrict__ nd_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpShared& S = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  WarpCtx C;
  ctx_init(P, C);
  unsigned long long nposes = 0;
  while (true) {
    unsigned long long pr = 0;
    if (lane == 0) pr = atomicAdd(P.work_counter, 1ull);
    pr = __shfl_sync(FULL, pr, 0);
    if ((long long)pr >= n) break;
    double a[7], b[7];
#pragma unroll
    for (int k = 0; k < 7; k++) { a[k] = __ldg(q0 + 7 * pr + k); b[k] = __ldg(q1 + 7 * pr + k); }
    const double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    const double dtr = sqrt(dx * dx + dy * dy + dz * dz);
    const double theta = ompl_arc_length(a, b);
    const long long nt = (long long)ceil(dtr / F.seg_trans), nr = (long long)ceil(theta / F.seg_rot);
    const long long nd = nt > nr ? nt : nr;
    // sample points: index 0 = s2 itself, index m = interior point m/nd (m = 1..nd-1)
    const long long total = nd >= 2 ? nd : 1;
    bool valid = true;
    // Swept windows as in k_transit (the boolean needs no first-contact step, but the prefix order keeps the early
    // exit): a robot point is R(q(t)) (v - dc) + scale p(t), so between t and t' it moves at most
    // |t - t'| (scale |dp| + 2 theta (R_rob + |dc|)); the slerp of two UNIT quaternions turns at constant rate.
    const double na2 = a[3] * a[3] + a[4] * a[4] + a[5] * a[5] + a[6] * a[6];
    const double nb2 = b[3] * b[3] + b[4] * b[4] + b[5] * b[5] + b[6] * b[6];
    const double dcn = sqrt(F.dc[0] * F.dc[0] + F.dc[1] * F.dc[1] + F.dc[2] * F.dc[2]);
    const double rate = (F.scale * dtr + 2.0 * theta * (P.rob_radius + dcn)) * (1.0 + 1e-9);
    const double step_radius = 0.5 * rate / (double)nd;
    // FCLStateValidityChecker::isValid = satisfiesBounds && collision free, for s2 and every interior point.
    // Motions whose end points are safely inside need no per-point bounds test (the lerp stays inside, the
    // slerp of two unit quaternions stays unit); the others are walked point by point with the literal test.
    const bool safe = ompl_safely_inside(F.lo, F.hi, a) && ompl_safely_inside(F.lo, F.hi, b);
    if (!ompl_satisfies_bounds(F.lo, F.hi, b)) {   // checkMotion tests s2 first (DiscreteMotionValidator.cpp:101)
      if (lane == 0) {
        out[pr] = 0;
        if (nd_out) nd_out[pr] = nd;
      }
      continue;
    }
    int ws = 0;
    if (safe && fabs(na2 - 1.0) < 1e-6 && fabs(nb2 - 1.0) < 1e-6 && step_radius > 0.0) {
      const double cap = fmin(32.0 * (double)P.sweep_tiles, P.sweep_rmax / step_radius);
      if (cap >= 4.0) {
        ws = 4;
        while (2.0 * ws <= cap && (double)(2 * ws) * step_radius <= P.sweep_rstart) ws *= 2;
      }
    }
    const int ws_cap = ws > 0 ? (int)fmin(32.0 * (double)P.sweep_tiles, P.sweep_rmax / step_radius) : 0;
    int nwin = 4;
    bool exact_next = true;   // the first tile (s2 and the first interior points) is always walked
    long long base = 0;
    while (base < total && valid) {
      if (ws > 0 && !exact_next) {
        const long long m_lo = base + (long long)lane * ws;
        const long long m_hi = min(total - 1, m_lo + ws - 1);
        const bool mine_w = lane < nwin && m_lo < total;
        if (mine_w) {
          const double t_lo = (double)m_lo / (double)nd, t_hi = (double)m_hi / (double)nd;
          const double radius = 0.5 * (t_hi - t_lo) * rate * (1.0 + 1e-9) + 1e-6;
          double s[7];
          ompl_point_to_unit(F, a, b, theta, 0.5 * (t_lo + t_hi), s);
          set_pose(P, S, lane, s);
          S.eta[lane] = __fadd_ru(S.eta[lane], __double2float_ru(0.5 * radius));
          S.tau[lane] = radius;
        }
        const unsigned tried = __ballot_sync(FULL, mine_w);
        __syncwarp();
        const unsigned blocked = traverse_tile<MODE_SWEPT>(P, S, C, tried, lane);
        if (C.overflow) { valid = false; break; }
        __syncwarp();
        const int cleared = blocked ? __ffs(blocked) - 1 : __popc(tried);
        const long long adv = min((long long)cleared * ws, total - base);
        base += adv;
        nposes += (unsigned long long)adv;
        if (!blocked) {
          if (nwin < 32) nwin *= 2;
          else if (2 * ws <= ws_cap) ws *= 2;
        } else if (ws > 4) {
          ws >>= 1;
          nwin = 2;
        } else {
          exact_next = true;
          nwin = 2;
        }
        continue;
      }
      const int exact_n = (base > 0 && ws > 0) ? P.sweep_short_tile : 32;
      exact_next = false;
      const long long m = base + lane;
      bool have = m < total && lane < exact_n;
      bool oob = false;
      if (have) {
        double o[7], s[7];
        if (m == 0) {   // s2 exactly, not its slerp image
#pragma unroll
          for (int k = 0; k < 7; k++) o[k] = b[k];
        } else {
          ompl_point(a, b, theta, (double)m / (double)nd, o);
          if (!safe) oob = !ompl_satisfies_bounds(F.lo, F.hi, o);
        }
        if (!oob) {
          ompl_row_to_unit(F, o, s);
          set_pose(P, S, lane, s);
        }
      }
      const unsigned oobm = __ballot_sync(FULL, have && oob);
      if (oobm) have = have && lane < __ffs(oobm) - 1;   // nothing behind an out-of-bounds point matters
      unsigned alive = __ballot_sync(FULL, have);
      __syncwarp();
      unsigned hits = traverse_tile<MODE_SEGMENT>(P, S, C, alive, lane);
      if (C.overflow) { valid = false; break; }
      nposes += hits ? (unsigned)__ffs(hits) : __popc(alive);
      if (hits | oobm) valid = false;
      base += exact_n;
      __syncwarp();
    }
    if (lane == 0) {
      out[pr] = valid ? 1 : 0;
      if (nd_out) nd_out[pr] = nd;
    }
    if (C.overflow) break;
  }
  ctx_flush(P, C, nposes, lane);
}

// ---------------------------------------------------------------------------------------------
// K3: all contacts of a state + Moeller intersection segment per contact
// (CDModel::collideForDetails, /root/reference/lib/osr/cdmodel.cc:200-229;
//  UnitWorld::intersectingSegments, lib/osr/unit_world.cc:959-1012).
// ---------------------------------------------------------------------------------------------

// FCL emits one contact per "deepest point" (0..2) of an intersecting pair in enable_contact mode
// (third-party/fcl/.../intersect-inl.h:801-849 with computeDeepestPoints :853-893 and
// buildTrianglePlane :1040-1053); each contact becomes one facePairs row (cdmodel.cc:213-227).
__device__ __forceinline__ bool tri_plane_dev(const double* v1, const double* v2, const double* v3, double* n,
                                              double& t) {
  double a[3] = {v2[0] - v1[0], v2[1] - v1[1], v2[2] - v1[2]};
  double b[3] = {v3[0] - v1[0], v3[1] - v1[1], v3[2] - v1[2]};
  double c[3];
  crossT(a, b, c);
  double sq = dotT(c, c);
  if (sq > 0) {
    double l = sqrt(sq);
    n[0] = c[0] / l; n[1] = c[1] / l; n[2] = c[2] / l;
    t = dotT(n, v1);
    return true;
  }
  return false;
}
__device__ __forceinline__ void deepest_dev(const double (*pts)[3], const double* n, double t, double& depth,
                                            unsigned& num) {
  const double eps = 1e-5;
  double max_depth = -1.7976931348623157e308;
  unsigned cnt = 0, nneg = 0, npos = 0, nzero = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    double dist = -(dotT(n, pts[i]) - t);
    if (dist > eps) npos++; else if (dist < -eps) nneg++; else nzero++;
    if (dist > max_depth) { max_depth = dist; cnt = 1; }
    else if (dist + 1e-6 >= max_depth) cnt++;
  }
  if (max_depth < -eps) cnt = 0;
  if (nzero == 0 && (nneg == 0 || npos == 0)) cnt = 0;
  depth = max_depth;
  num = cnt;
}
__device__ __forceinline__ unsigned contact_multiplicity_dev(const double* Pt, const double* Qt, const double* Rd) {
  double R[12];
#pragma unroll
  for (int k = 0; k < 12; k++) R[k] = Rd[k * 32];
  double P[3][3], Q[3][3];
#pragma unroll
  for (int v = 0; v < 3; v++) {
    double x = __ldg(Qt + 3 * v), y = __ldg(Qt + 3 * v + 1), z = __ldg(Qt + 3 * v + 2);
#pragma unroll
    for (int k = 0; k < 3; k++) {
      P[v][k] = __ldg(Pt + 3 * v + k);
      Q[v][k] = (R[3 * k] * x + R[3 * k + 1] * y + R[3 * k + 2] * z) + R[9 + k];
    }
  }
  double n1[3] = {0, 0, 0}, n2[3] = {0, 0, 0}, t1 = 0, t2 = 0, d1, d2;
  unsigned k1, k2;
  tri_plane_dev(P[0], P[1], P[2], n1, t1);
  tri_plane_dev(Q[0], Q[1], Q[2], n2, t2);
  deepest_dev(Q, n1, t1, d2, k2);
  deepest_dev(P, n2, t2, d1, k1);
  if (d1 > d2) return min(k2, 2u);
  return min(k1, 2u);
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
    k_contacts(KParams P, const double* __restrict__ q, long long n, ContactRec* recs,
               unsigned long long* rec_count, unsigned long long rec_cap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpShared& S = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  WarpCtx C;
  ctx_init(P, C);
  // tiles of 8 states: contact states run deep traversals, a few poses already fill the warp
  const int TP = 8;
  const long long ntiles = (n + TP - 1) / TP;
  unsigned long long nposes = 0;
  while (true) {
    unsigned long long tile = 0;
    if (lane == 0) tile = atomicAdd(P.work_counter, 1ull);
    tile = __shfl_sync(FULL, tile, 0);
    if ((long long)tile >= ntiles) break;
    long long i = (long long)tile * TP + lane;
    bool have = lane < TP && i < n;
    if (have) {
      double s[7];
#pragma unroll
      for (int k = 0; k < 7; k++) s[k] = __ldg(q + 7 * i + k);
      set_pose(P, S, lane, s);
    }
    unsigned alive = __ballot_sync(FULL, have);
    nposes += __popc(alive);
    __syncwarp();
    traverse_tile<MODE_CONTACTS>(P, S, C, alive, lane, (int)(tile * TP), recs, rec_count, rec_cap);
    __syncwarp();
    if (C.overflow) break;
  }
  ctx_flush(P, C, nposes, lane);
}

// --- Moeller 1997 interval-overlap test with intersection segment, fp64 -----------------------
// Follows the published algorithm as used by the reference (lib/tritri/Moller97d.cc:579-717,
// helpers :474-543, coplanar case :169-217) including its EPSILON=1e-6 snapping of the plane
// distances.  Output: returns 1 if the triangles intersect; *coplanar set when both lie in one
// plane (then no segment is produced).
struct MollerIv {
  double s0, s1;     // interval parameters
  double p0[3], p1[3];
};
__device__ __forceinline__ void moller_isect2(const double* V0, const double* V1, const double* V2, double vv0,
                                              double vv1, double vv2, double d0, double d1, double d2,
                                              MollerIv& iv) {
  double tmp = d0 / (d0 - d1);
  iv.s0 = vv0 + (vv1 - vv0) * tmp;
#pragma unroll
  for (int k = 0; k < 3; k++) iv.p0[k] = tmp * (V1[k] - V0[k]) + V0[k];
  tmp = d0 / (d0 - d2);
  iv.s1 = vv0 + (vv2 - vv0) * tmp;
#pragma unroll
  for (int k = 0; k < 3; k++) iv.p1[k] = V0[k] + tmp * (V2[k] - V0[k]);
}
// returns true when the triangle lies in the other's plane
__device__ __forceinline__ bool moller_interval(const double* V0, const double* V1, const double* V2, double vv0,
                                                double vv1, double vv2, double d0, double d1, double d2,
                                                double d0d1, double d0d2, MollerIv& iv) {
  if (d0d1 > 0.0) moller_isect2(V2, V0, V1, vv2, vv0, vv1, d2, d0, d1, iv);
  else if (d0d2 > 0.0) moller_isect2(V1, V0, V2, vv1, vv0, vv2, d1, d0, d2, iv);
  else if (d1 * d2 > 0.0 || d0 != 0.0) moller_isect2(V0, V1, V2, vv0, vv1, vv2, d0, d1, d2, iv);
  else if (d1 != 0.0) moller_isect2(V1, V0, V2, vv1, vv0, vv2, d1, d0, d2, iv);
  else if (d2 != 0.0) moller_isect2(V2, V0, V1, vv2, vv0, vv1, d2, d0, d1, iv);
  else return true;
  return false;
}
__device__ __forceinline__ bool moller_edge_edge(const double* V0, const double* U0, const double* U1, int i0,
                                                 int i1, double Ax, double Ay) {
  double Bx = U0[i0] - U1[i0], By = U0[i1] - U1[i1];
  double Cx = V0[i0] - U0[i0], Cy = V0[i1] - U0[i1];
  double f = Ay * Bx - Ax * By;
  double d = By * Cx - Bx * Cy;
  if ((f > 0 && d >= 0 && d <= f) || (f < 0 && d <= 0 && d >= f)) {
    double e = Ax * Cy - Ay * Cx;
    if (f > 0) { if (e >= 0 && e <= f) return true; }
    else { if (e <= 0 && e >= f) return true; }
  }
  return false;
}
__device__ __forceinline__ bool moller_edge_tri(const double* V0, const double* V1, const double* U0,
                                                const double* U1, const double* U2, int i0, int i1) {
  double Ax = V1[i0] - V0[i0], Ay = V1[i1] - V0[i1];
  return moller_edge_edge(V0, U0, U1, i0, i1, Ax, Ay) || moller_edge_edge(V0, U1, U2, i0, i1, Ax, Ay) ||
         moller_edge_edge(V0, U2, U0, i0, i1, Ax, Ay);
}
__device__ __forceinline__ bool moller_point_in_tri(const double* V0, const double* U0, const double* U1,
                                                    const double* U2, int i0, int i1) {
  double a = U1[i1] - U0[i1], b = -(U1[i0] - U0[i0]);
  double c = -a * U0[i0] - b * U0[i1];
  double d0 = a * V0[i0] + b * V0[i1] + c;
  a = U2[i1] - U1[i1]; b = -(U2[i0] - U1[i0]);
  c = -a * U1[i0] - b * U1[i1];
  double d1 = a * V0[i0] + b * V0[i1] + c;
  a = U0[i1] - U2[i1]; b = -(U0[i0] - U2[i0]);
  c = -a * U2[i0] - b * U2[i1];
  double d2 = a * V0[i0] + b * V0[i1] + c;
  return d0 * d1 > 0.0 && d0 * d2 > 0.0;
}
__device__ __forceinline__ int moller_coplanar(const double* N, const double* V0, const double* V1,
                                               const double* V2, const double* U0, const double* U1,
                                               const double* U2) {
  double A[3] = {fabs(N[0]), fabs(N[1]), fabs(N[2])};
  int i0, i1;
  if (A[0] > A[1]) {
    if (A[0] > A[2]) { i0 = 1; i1 = 2; } else { i0 = 0; i1 = 1; }
  } else {
    if (A[2] > A[1]) { i0 = 0; i1 = 1; } else { i0 = 0; i1 = 2; }
  }
  if (moller_edge_tri(V0, V1, U0, U1, U2, i0, i1)) return 1;
  if (moller_edge_tri(V1, V2, U0, U1, U2, i0, i1)) return 1;
  if (moller_edge_tri(V2, V0, U0, U1, U2, i0, i1)) return 1;
  if (moller_point_in_tri(V0, U0, U1, U2, i0, i1)) return 1;
  if (moller_point_in_tri(U0, V0, V1, V2, i0, i1)) return 1;
  return 0;
}
__device__ __forceinline__ int moller_isectline(const double (*V)[3], const double (*U)[3], int* coplanar,
                                                double* pt1, double* pt2) {
  const double EPS = 0.000001;
  double E1[3], E2[3], N1[3], N2[3];
#pragma unroll
  for (int k = 0; k < 3; k++) { E1[k] = V[1][k] - V[0][k]; E2[k] = V[2][k] - V[0][k]; }
  crossT(E1, E2, N1);
  double d1 = -dotT(N1, V[0]);
  double du0 = dotT(N1, U[0]) + d1, du1 = dotT(N1, U[1]) + d1, du2 = dotT(N1, U[2]) + d1;
  if (fabs(du0) < EPS) du0 = 0.0;
  if (fabs(du1) < EPS) du1 = 0.0;
  if (fabs(du2) < EPS) du2 = 0.0;
  double du0du1 = du0 * du1, du0du2 = du0 * du2;
  if (du0du1 > 0.0 && du0du2 > 0.0) return 0;
#pragma unroll
  for (int k = 0; k < 3; k++) { E1[k] = U[1][k] - U[0][k]; E2[k] = U[2][k] - U[0][k]; }
  crossT(E1, E2, N2);
  double d2 = -dotT(N2, U[0]);
  double dv0 = dotT(N2, V[0]) + d2, dv1 = dotT(N2, V[1]) + d2, dv2 = dotT(N2, V[2]) + d2;
  if (fabs(dv0) < EPS) dv0 = 0.0;
  if (fabs(dv1) < EPS) dv1 = 0.0;
  if (fabs(dv2) < EPS) dv2 = 0.0;
  double dv0dv1 = dv0 * dv1, dv0dv2 = dv0 * dv2;
  if (dv0dv1 > 0.0 && dv0dv2 > 0.0) return 0;
  double D[3];
  crossT(N1, N2, D);
  double mx = fabs(D[0]);
  int idx = 0;
  if (fabs(D[1]) > mx) { mx = fabs(D[1]); idx = 1; }
  if (fabs(D[2]) > mx) { mx = fabs(D[2]); idx = 2; }
  MollerIv A, B;
  *coplanar = moller_interval(V[0], V[1], V[2], V[0][idx], V[1][idx], V[2][idx], dv0, dv1, dv2, dv0dv1, dv0dv2, A)
                  ? 1 : 0;
  if (*coplanar) return moller_coplanar(N1, V[0], V[1], V[2], U[0], U[1], U[2]);
  if (moller_interval(U[0], U[1], U[2], U[0][idx], U[1][idx], U[2][idx], du0, du1, du2, du0du1, du0du2, B)) {
    // U snapped into V's plane while V is not in U's: the reference reads uninitialised memory
    // here (Moller97d.cc:671-672 ignores this return value); defined as the zero segment.
    return 1;
  }
  bool swapA = A.s0 > A.s1, swapB = B.s0 > B.s1;
  double a0 = swapA ? A.s1 : A.s0, a1 = swapA ? A.s0 : A.s1;
  double b0 = swapB ? B.s1 : B.s0, b1 = swapB ? B.s0 : B.s1;
  if (a1 < b0 || b1 < a0) return 0;
  const double* Alo = swapA ? A.p1 : A.p0; const double* Ahi = swapA ? A.p0 : A.p1;
  const double* Blo = swapB ? B.p1 : B.p0; const double* Bhi = swapB ? B.p0 : B.p1;
  const double* r1; const double* r2;
  if (b0 < a0) { r1 = Alo; r2 = (b1 < a1) ? Bhi : Ahi; }
  else { r1 = Blo; r2 = (b1 > a1) ? Ahi : Bhi; }
#pragma unroll
  for (int k = 0; k < 3; k++) { pt1[k] = r1[k]; pt2[k] = r2[k]; }
  return 1;
}

// barycentric coordinates of p w.r.t. (a,b,c), areas ratio form
__device__ __forceinline__ void barycentric_dev(const double* p, const double* a, const double* b, const double* c,
                                                double* out) {
  double v0[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, v1[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]},
         v2[3] = {p[0] - a[0], p[1] - a[1], p[2] - a[2]};
  double d00 = dotT(v0, v0), d01 = dotT(v0, v1), d11 = dotT(v1, v1), d20 = dotT(v2, v0), d21 = dotT(v2, v1);
  double den = d00 * d11 - d01 * d01;
  double v = (d11 * d20 - d01 * d21) / den, w = (d00 * d21 - d01 * d20) / den;
  out[0] = 1.0 - v - w; out[1] = v; out[2] = w;
}

// One thread per output row.  row_rec[r] = index of the contact record the row belongs to.
__global__ void k_contact_rows(KParams P, const double* __restrict__ q, const ContactRec* __restrict__ recs,
                               const int* __restrict__ row_rec, long long nrows, double* __restrict__ seg,
                               int* __restrict__ faces, uint8_t* __restrict__ coplanar, double* __restrict__ bary) {
  long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrows) return;
  ContactRec rec = recs[row_rec[r]];
  double s[7];
#pragma unroll
  for (int k = 0; k < 7; k++) s[k] = __ldg(q + 7 * (long long)rec.state + k);
  Rigid rob;
  state_to_rigid(s, rob);
  const double* Pt = P.env.tris + (size_t)9 * rec.env_tri;
  const double* Qt = P.rob.tris + (size_t)9 * rec.rob_tri;
  double V[3][3], U[3][3];
#pragma unroll
  for (int v = 0; v < 3; v++) {
    double px = Pt[3 * v], py = Pt[3 * v + 1], pz = Pt[3 * v + 2];
    double qx = Qt[3 * v], qy = Qt[3 * v + 1], qz = Qt[3 * v + 2];
#pragma unroll
    for (int k = 0; k < 3; k++) {   // world space: env_tf * v, rob_tf * v (unit_world.cc:981-982)
      V[v][k] = (P.env_tf.R[3 * k] * px + P.env_tf.R[3 * k + 1] * py + P.env_tf.R[3 * k + 2] * pz) + P.env_tf.t[k];
      U[v][k] = (rob.R[3 * k] * qx + rob.R[3 * k + 1] * qy + rob.R[3 * k + 2] * qz) + rob.t[k];
    }
  }
  int cop = 0;
  double p1[3] = {0, 0, 0}, p2[3] = {0, 0, 0};
  moller_isectline(V, U, &cop, p1, p2);
#pragma unroll
  for (int k = 0; k < 3; k++) { seg[6 * r + k] = p1[k]; seg[6 * r + 3 + k] = p2[k]; }
  faces[2 * r] = rec.env_tri;
  faces[2 * r + 1] = rec.rob_tri;
  coplanar[r] = (uint8_t)(cop != 0);
  if (bary) {
    double mid[3] = {0.5 * (p1[0] + p2[0]), 0.5 * (p1[1] + p2[1]), 0.5 * (p1[2] + p2[2])};
    barycentric_dev(mid, V[0], V[1], V[2], bary + 6 * r);
    barycentric_dev(mid, U[0], U[1], U[2], bary + 6 * r + 3);
  }
}

}  // namespace osr

namespace osr {
// ---------------------------------------------------------------------------------------------
// K4: bounding-box "disentangled" filter (UnitWorld::isDisentangled,
// /root/reference/lib/osr/unit_world.cc:249-264 -> CDModel::collideBB, lib/osr/cdmodel.cc:180-197):
// fcl::collide of the two models' local AABBs as fcl::Box shapes, robot first.  FCL routes
// Box-Box to boxBoxIntersect (third-party/fcl/.../primitive_shape_algorithm/box_box-inl.h:827-1117,
// 1395-1412): a 15-axis SAT, fp64, +1e-6 on |R| for the 9 edge axes only.  out = 1 iff separated.
// ---------------------------------------------------------------------------------------------
struct BoxPair {
  double rob_c[3], rob_side[3];
  double env_c[3], env_side[3];
};
__global__ void k_disentangled(Rigid env_tf, BoxPair bp, const double* __restrict__ q, long long n,
                               uint8_t* __restrict__ out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s[7];
#pragma unroll
  for (int k = 0; k < 7; k++) s[k] = __ldg(q + 7 * i + k);
  Rigid rob;
  state_to_rigid(s, rob);
  // tf1 = robTf * translate(rob centre), tf2 = envTf * translate(env centre)
  double t1[3], t2[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    t1[k] = (rob.R[3 * k] * bp.rob_c[0] + rob.R[3 * k + 1] * bp.rob_c[1] + rob.R[3 * k + 2] * bp.rob_c[2]) + rob.t[k];
    t2[k] = (env_tf.R[3 * k] * bp.env_c[0] + env_tf.R[3 * k + 1] * bp.env_c[1] + env_tf.R[3 * k + 2] * bp.env_c[2]) +
            env_tf.t[k];
  }
  const double* R1 = rob.R;
  const double* R2 = env_tf.R;
  double p[3] = {t2[0] - t1[0], t2[1] - t1[1], t2[2] - t1[2]};
  double pp[3], A[3], B[3], R[3][3], Q[3][3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    pp[k] = R1[k] * p[0] + R1[3 + k] * p[1] + R1[6 + k] * p[2];
    A[k] = bp.rob_side[k] * 0.5;
    B[k] = bp.env_side[k] * 0.5;
  }
#pragma unroll
  for (int a = 0; a < 3; a++)
#pragma unroll
    for (int b = 0; b < 3; b++) {
      R[a][b] = R1[a] * R2[b] + R1[3 + a] * R2[3 + b] + R1[6 + a] * R2[6 + b];
      Q[a][b] = fabs(R[a][b]);
    }
  bool sep = false;
#pragma unroll
  for (int a = 0; a < 3; a++)
    sep |= fabs(pp[a]) - ((Q[a][0] * B[0] + Q[a][1] * B[1] + Q[a][2] * B[2]) + A[a]) > 0;
#pragma unroll
  for (int b = 0; b < 3; b++) {
    double tmp = R2[b] * p[0] + R2[3 + b] * p[1] + R2[6 + b] * p[2];
    sep |= fabs(tmp) - ((Q[0][b] * A[0] + Q[1][b] * A[1] + Q[2][b] * A[2]) + B[b]) > 0;
  }
#pragma unroll
  for (int a = 0; a < 3; a++)
#pragma unroll
    for (int b = 0; b < 3; b++) Q[a][b] += 1.0e-6;
#define OSR_EDGE(t, rhs) sep |= fabs(t) - (rhs) > 0
  OSR_EDGE(pp[2] * R[1][0] - pp[1] * R[2][0], A[1] * Q[2][0] + A[2] * Q[1][0] + B[1] * Q[0][2] + B[2] * Q[0][1]);
  OSR_EDGE(pp[2] * R[1][1] - pp[1] * R[2][1], A[1] * Q[2][1] + A[2] * Q[1][1] + B[0] * Q[0][2] + B[2] * Q[0][0]);
  OSR_EDGE(pp[2] * R[1][2] - pp[1] * R[2][2], A[1] * Q[2][2] + A[2] * Q[1][2] + B[0] * Q[0][1] + B[1] * Q[0][0]);
  OSR_EDGE(pp[0] * R[2][0] - pp[2] * R[0][0], A[0] * Q[2][0] + A[2] * Q[0][0] + B[1] * Q[1][2] + B[2] * Q[1][1]);
  OSR_EDGE(pp[0] * R[2][1] - pp[2] * R[0][1], A[0] * Q[2][1] + A[2] * Q[0][1] + B[0] * Q[1][2] + B[2] * Q[1][0]);
  OSR_EDGE(pp[0] * R[2][2] - pp[2] * R[0][2], A[0] * Q[2][2] + A[2] * Q[0][2] + B[0] * Q[1][1] + B[1] * Q[1][0]);
  OSR_EDGE(pp[1] * R[0][0] - pp[0] * R[1][0], A[0] * Q[1][0] + A[1] * Q[0][0] + B[1] * Q[2][2] + B[2] * Q[2][1]);
  OSR_EDGE(pp[1] * R[0][1] - pp[0] * R[1][1], A[0] * Q[1][1] + A[1] * Q[0][1] + B[0] * Q[2][2] + B[2] * Q[2][0]);
  OSR_EDGE(pp[1] * R[0][2] - pp[0] * R[1][2], A[0] * Q[1][2] + A[1] * Q[0][2] + B[0] * Q[2][1] + B[1] * Q[2][0]);
#undef OSR_EDGE
  out[i] = sep ? 1 : 0;
}
}  // namespace osr
