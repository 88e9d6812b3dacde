// sat17.h -- compact SAT-17 triangle-triangle predicates (host + device).
//
// FCL's Intersect::intersect_Triangle + project6
// (/root/reference/third-party/fcl/include/fcl/narrowphase/detail/traversal/collision/intersect-inl.h:728-799,
// 1086-1107): 17 candidate axes n1, m1, e_i x f_j (9), e_i x n1 (3), f_j x m1 (3); an axis separates iff
// min(P) > max(Q) or min(Q) > max(P) (strict: touching counts as intersecting); no epsilon.
//
// Same arithmetic as the unrolled versions in kernels.cuh, but the 9 + 6 cross-product axes run in rolled loops
// whose operands are picked with selects: ~4x less code.  Code size matters here: the traversal kernel's
// instruction working set has to stay inside the 32 KB L1.5 instruction cache.
//   sat17_exact_c : fp64, the reference's operation order (compile without FMA contraction) -> the label
//   sat17_filter_c: fp32 on the rounded fp64 inputs with a forward error bound -> 0 certainly separated,
//                   1 certainly intersecting, 2 undecided (then the caller runs sat17_exact_c)
#pragma once
#include "se3.h"

namespace osr {

template <typename T>
struct TriPair {  // everything relative to P1 (p1 == 0); q in the environment frame
  T p2[3], p3[3], q1[3], q2[3], q3[3];
};

template <typename T>
OSR_HD void cross3(const T* a, const T* b, T* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}
template <typename T>
OSR_HD T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T>
OSR_HD T tmin3(T a, T b, T c) { T m = b < c ? b : c; return a < m ? a : m; }
template <typename T>
OSR_HD T tmax3(T a, T b, T c) { T m = b > c ? b : c; return a > m ? a : m; }

// project6: g1 = min(P) - max(Q), g2 = min(Q) - max(P); dot(ax, p1) is exactly 0
template <typename T>
OSR_HD void gaps17(const T* ax, const TriPair<T>& t, T& g1, T& g2) {
  T P2 = dot3(ax, t.p2), P3 = dot3(ax, t.p3);
  T Q1 = dot3(ax, t.q1), Q2 = dot3(ax, t.q2), Q3 = dot3(ax, t.q3);
  T mnP = tmin3<T>(T(0), P2, P3), mxP = tmax3<T>(T(0), P2, P3);
  T mnQ = tmin3<T>(Q1, Q2, Q3), mxQ = tmax3<T>(Q1, Q2, Q3);
  g1 = mnP - mxQ;
  g2 = mnQ - mxP;
}

template <typename T>
struct Edges {  // e1 = p2-p1, e2 = p3-p2, e3 = p1-p3; f likewise
  T e[3][3], f[3][3];
  OSR_HD void set(const TriPair<T>& t) {
    for (int k = 0; k < 3; k++) {
      e[0][k] = t.p2[k]; e[1][k] = t.p3[k] - t.p2[k]; e[2][k] = T(0) - t.p3[k];
      f[0][k] = t.q2[k] - t.q1[k]; f[1][k] = t.q3[k] - t.q2[k]; f[2][k] = t.q1[k] - t.q3[k];
    }
  }
  // operand selection without dynamic register indexing (which would go through local memory)
  OSR_HD void pick_e(int i, T* o) const {
    for (int k = 0; k < 3; k++) o[k] = i == 0 ? e[0][k] : (i == 1 ? e[1][k] : e[2][k]);
  }
  OSR_HD void pick_f(int i, T* o) const {
    for (int k = 0; k < 3; k++) o[k] = i == 0 ? f[0][k] : (i == 1 ? f[1][k] : f[2][k]);
  }
};

// true iff no axis separates (fp64, reference operation order)
OSR_HD bool sat17_exact_c(const TriPair<double>& t) {
  Edges<double> E;
  E.set(t);
  double n1[3], m1[3], ax[3], u[3], w[3], g1, g2;
  cross3(E.e[0], E.e[1], n1);
  gaps17(n1, t, g1, g2);
  if (g1 > 0 || g2 > 0) return false;
  cross3(E.f[0], E.f[1], m1);
  gaps17(m1, t, g1, g2);
  if (g1 > 0 || g2 > 0) return false;
#pragma unroll 1
  for (int a = 0; a < 9; a++) {  // e_i x f_j, i major (intersect-inl.h:757-777)
    int i = a / 3, j = a - 3 * i;
    E.pick_e(i, u);
    E.pick_f(j, w);
    cross3(u, w, ax);
    gaps17(ax, t, g1, g2);
    if (g1 > 0 || g2 > 0) return false;
  }
#pragma unroll 1
  for (int a = 0; a < 6; a++) {  // e_i x n1 (i = 0..2), then f_i x m1 (intersect-inl.h:779-797)
    int i = a < 3 ? a : a - 3;
    if (a < 3) E.pick_e(i, u); else E.pick_f(i, u);
    for (int k = 0; k < 3; k++) w[k] = a < 3 ? n1[k] : m1[k];
    cross3(u, w, ax);
    gaps17(ax, t, g1, g2);
    if (g1 > 0 || g2 > 0) return false;
  }
  return true;
}

// fp32 filter.  Inputs are the fp64 values rounded to nearest.  With u = 2^-24, L = max |coordinate| and, for an edge
// vector v, |v| = its largest |component| (as computed), the computed gap of an axis differs from the fp64 value by
// at most
//     u L (48 L (|a| + |b|) + 84 |a| |b|)                           axis a x b of two edges (n1, m1, e_i x f_j)
//     u L (96 L (|a||b| + |v| (|a| + |b|)) + 216 |v| |a| |b|)        axis v x (a x b)  (e_i x n1, f_i x m1)
// (edge components carry 4 u L of input rounding, a cross product of vectors with errors ea, eb gains
// 2 (ea |b| + eb |a|) + 4 u |a| |b| per component, a projection on a point of magnitude <= L gains 3 L times that
// plus 12 u L |axis|, a gap is the difference of two projections plus its own rounding); the factor 1.1 below
// covers the second-order terms and the rounding of the fp64 evaluation itself.  The bounds scale with the edges
// actually involved: with one L^3 / L^4 bound for all axes, the long thin triangles of the alpha puzzle left half
// of all intersecting pairs undecided (their in-plane axes e x n have tiny projections next to L^4).
// Returns 0 = certainly separated, 1 = certainly intersecting (no axis separates), 2 = undecided (the caller runs
// sat17_exact_c).
template <typename T>
OSR_HD T maxabs3(const T* v) {
  T a = v[0] < 0 ? -v[0] : v[0], b = v[1] < 0 ? -v[1] : v[1], c = v[2] < 0 ? -v[2] : v[2];
  T m = a > b ? a : b;
  return m > c ? m : c;
}
OSR_HD float sat17_bound_q(float uL, float L, float a, float b) {
  return uL * (48.f * L * (a + b) + 84.f * a * b) + 1e-37f;
}
OSR_HD float sat17_bound_c(float uL, float L, float v, float a, float b) {
  return uL * (96.f * L * (a * b + v * (a + b)) + 216.f * v * a * b) + 1e-37f;
}
OSR_HD int sat17_filter_c(const TriPair<float>& t) {
  float L = 0.f;
  for (int k = 0; k < 3; k++) {
    L = fmaxf(L, fmaxf(fabsf(t.p2[k]), fabsf(t.p3[k])));
    L = fmaxf(L, fmaxf(fabsf(t.q1[k]), fmaxf(fabsf(t.q2[k]), fabsf(t.q3[k]))));
  }
  const float uL = 1.1f * 5.9604645e-08f * L;
  Edges<float> E;
  E.set(t);
  const float ue0 = maxabs3(E.e[0]), ue1 = maxabs3(E.e[1]), ue2 = maxabs3(E.e[2]);
  const float uf0 = maxabs3(E.f[0]), uf1 = maxabs3(E.f[1]), uf2 = maxabs3(E.f[2]);
  float n1[3], m1[3], ax[3], u[3], w[3], g1, g2;
  bool sep = false, unsure = false;
  cross3(E.e[0], E.e[1], n1);
  gaps17(n1, t, g1, g2);
  float B = sat17_bound_q(uL, L, ue0, ue1);
  sep |= (g1 > B) | (g2 > B); unsure |= !((g1 < -B) & (g2 < -B));
  cross3(E.f[0], E.f[1], m1);
  gaps17(m1, t, g1, g2);
  B = sat17_bound_q(uL, L, uf0, uf1);
  sep |= (g1 > B) | (g2 > B); unsure |= !((g1 < -B) & (g2 < -B));
  if (sep) return 0;
#pragma unroll 1
  for (int a = 0; a < 9; a++) {
    int i = a / 3, j = a - 3 * i;
    E.pick_e(i, u);
    E.pick_f(j, w);
    cross3(u, w, ax);
    gaps17(ax, t, g1, g2);
    B = sat17_bound_q(uL, L, i == 0 ? ue0 : (i == 1 ? ue1 : ue2), j == 0 ? uf0 : (j == 1 ? uf1 : uf2));
    sep |= (g1 > B) | (g2 > B); unsure |= !((g1 < -B) & (g2 < -B));
  }
  if (sep) return 0;
#pragma unroll 1
  for (int a = 0; a < 6; a++) {
    int i = a < 3 ? a : a - 3;
    if (a < 3) E.pick_e(i, u); else E.pick_f(i, u);
    for (int k = 0; k < 3; k++) w[k] = a < 3 ? n1[k] : m1[k];
    cross3(u, w, ax);
    gaps17(ax, t, g1, g2);
    const float v = a < 3 ? (i == 0 ? ue0 : (i == 1 ? ue1 : ue2)) : (i == 0 ? uf0 : (i == 1 ? uf1 : uf2));
    B = sat17_bound_c(uL, L, v, a < 3 ? ue0 : uf0, a < 3 ? ue1 : uf1);
    sep |= (g1 > B) | (g2 > B); unsure |= !((g1 < -B) & (g2 < -B));
  }
  if (sep) return 0;
  return unsure ? 2 : 1;
}

}  // namespace osr
