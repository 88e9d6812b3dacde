// qnode.h -- the 32-byte OBB node of the traversal kernels and its box-box test (host + device).
//
// FCL keeps a 3x3 fp64 axis matrix, a centre and extents per BV (third-party/fcl/include/fcl/math/bv/OBB.h)
// and tests pairs with obbDisjoint (OBB-inl.h:399-525).  Here a node is
//     word 0-2  centre (fp32, model frame)
//     word 3    link   (int32: >=0 index of the left child, right = left+1; <0 leaf, triangle = -(link+1))
//     word 4-5  box orientation as a quaternion (w,x,y,z), 4 x fp16
//     word 6-7  half extents, 3 x fp16 rounded UP (+ 16 spare bits)
// so that both trees of the alpha puzzle (2 x 2015 nodes = 129 KB) are resident in one SM's shared memory and a
// node is two 128-bit loads.  The "true" box of a node is defined by the EXACT rotation of the normalised
// stored quaternion; the builder (geometry.cc) fits the extents to that frame in fp64 and rounds up, so the
// box contains its triangles by construction.  The test composes orientations as quaternions
// (qA* (x) qPose (x) qB -> one 3x3), which needs fewer instructions than decoding two matrices.
#pragma once
#include <stdint.h>
#include <string.h>

#include "se3.h"

#if defined(__CUDACC__)
#include <cuda_fp16.h>
#endif

namespace osr {

struct QNode {  // 32 bytes
  float c[3];
  int32_t link;
  uint32_t q01, q23;  // fp16 pairs (low half first): w x | y z
  uint32_t e01, e2x;  // fp16 pairs: ex ey | ez spare
};

// ---- fp16 <-> fp32 (host: bit manipulation; device: hardware converts) ------------------------------
OSR_HD float half_bits_to_float(uint16_t h) {
#if defined(__CUDA_ARCH__)
  return __half2float(__ushort_as_half(h));
#else
  uint32_t s = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1fu, m = h & 0x3ffu, out;
  if (e == 0) {
    if (m == 0) out = s;
    else {  // subnormal
      int sh = 0;
      while (!(m & 0x400u)) { m <<= 1; sh++; }
      out = s | ((uint32_t)(113 - sh) << 23) | ((m & 0x3ffu) << 13);
    }
  } else if (e == 31) out = s | 0x7f800000u | (m << 13);
  else out = s | ((e + 112) << 23) | (m << 13);
  float f;
  memcpy(&f, &out, 4);
  return f;
#endif
}

#if !defined(__CUDA_ARCH__)
// host only: nearest fp16 (ties to even) and the smallest fp16 >= x (x >= 0)
inline uint16_t float_to_half_bits_rn(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  uint32_t s = (x >> 16) & 0x8000u;
  int32_t e = (int32_t)((x >> 23) & 0xff) - 127 + 15;
  uint32_t m = x & 0x7fffffu;
  if (e >= 31) return (uint16_t)(s | 0x7bffu);  // clamp to the largest finite value
  if (e <= 0) {
    if (e < -10) return (uint16_t)s;
    m |= 0x800000u;
    int sh = 14 - e;
    uint32_t r = m >> sh, rem = m & ((1u << sh) - 1), half = 1u << (sh - 1);
    if (rem > half || (rem == half && (r & 1))) r++;
    return (uint16_t)(s | r);
  }
  uint32_t r = (uint32_t)(e << 10) | (m >> 13), rem = m & 0x1fffu;
  if (rem > 0x1000u || (rem == 0x1000u && (r & 1))) r++;
  return (uint16_t)(s | r);
}
inline uint16_t float_to_half_bits_up(double x) {  // x >= 0
  uint16_t h = float_to_half_bits_rn((float)x);
  while ((double)half_bits_to_float(h) < x) h++;
  return h;
}
#endif

// ---- pose as the box test sees it -----------------------------------------------------------------
// Robot frame -> environment frame: x -> Rot(q) x + t with q a UNIT quaternion (to fp32 rounding) and R its
// matrix; `eta` is the absolute slack added to every extent.  It covers fp32 rounding of the test itself and
// the difference between the reference's transform (quaternions are not renormalised there,
// lib/osr/osr_state.cc:16-25) and this rotation: |M p - Rot p| <= ||M - Rot||_F * |p| for every robot point p.
struct BoxPose {
  float q[4];
  float R[9];
  float t[3];
  float eta;
};

OSR_HD void quat_mul_f(const float* a, const float* b, float* o) {
  o[0] = fmaf(a[0], b[0], fmaf(-a[1], b[1], fmaf(-a[2], b[2], -a[3] * b[3])));
  o[1] = fmaf(a[0], b[1], fmaf(a[1], b[0], fmaf(a[2], b[3], -a[3] * b[2])));
  o[2] = fmaf(a[0], b[2], fmaf(a[2], b[0], fmaf(a[3], b[1], -a[1] * b[3])));
  o[3] = fmaf(a[0], b[3], fmaf(a[3], b[0], fmaf(a[1], b[2], -a[2] * b[1])));
}
OSR_HD void quat_to_rot_f(const float* q, float* R) {
  float tx = q[1] + q[1], ty = q[2] + q[2], tz = q[3] + q[3];
  float twx = tx * q[0], twy = ty * q[0], twz = tz * q[0];
  float txx = tx * q[1], txy = ty * q[1], txz = tz * q[1];
  float tyy = ty * q[2], tyz = tz * q[2], tzz = tz * q[3];
  R[0] = 1.f - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
  R[3] = txy + twz;         R[4] = 1.f - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.f - (txx + tyy);
}
OSR_HD void unpack_half2(uint32_t w, float& lo, float& hi) {
#if defined(__CUDA_ARCH__)
  float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
  lo = f.x; hi = f.y;
#else
  lo = half_bits_to_float((uint16_t)(w & 0xffffu));
  hi = half_bits_to_float((uint16_t)(w >> 16));
#endif
}
// first-order renormalisation of a quaternion whose fp16 components put |q|^2 within ~1e-3 of 1
OSR_HD void quat_decode(uint32_t q01, uint32_t q23, float* q) {
  float w, x, y, z;
  unpack_half2(q01, w, x);
  unpack_half2(q23, y, z);
  float n2 = fmaf(w, w, fmaf(x, x, fmaf(y, y, z * z)));
  float s = fmaf(-0.5f, n2, 1.5f);
  q[0] = w * s; q[1] = x * s; q[2] = y * s; q[3] = z * s;
}

// Box A (environment node) against box B (robot node) under pose P.  Returns true when no axis of the 15
// separates them after the slack (conservative: a `false` proves the exact boxes are disjoint).  szA / szB
// receive the squared size used by the descent rule (bvh_collision_traversal_node-inl.h:78-89).
OSR_HD bool obb_overlap_q(const QNode& A, const QNode& B, const BoxPose& P, float& szA, float& szB) {
  const float reps = 2e-5f;
  float qa[4], qb[4], q1[4], q2[4], Bm[9];
  quat_decode(A.q01, A.q23, qa);
  quat_decode(B.q01, B.q23, qb);
  quat_mul_f(P.q, qb, q1);
  float qac[4] = {qa[0], -qa[1], -qa[2], -qa[3]};
  quat_mul_f(qac, q1, q2);
  quat_to_rot_f(q2, Bm);  // B(i,j) = axis_i(A) . (R axis_j(B))
  // centre of B in A's frame: T = A^T (R cB + t - cA), rotating with the conjugate of qa
  float dx = fmaf(P.R[0], B.c[0], fmaf(P.R[1], B.c[1], fmaf(P.R[2], B.c[2], P.t[0]))) - A.c[0];
  float dy = fmaf(P.R[3], B.c[0], fmaf(P.R[4], B.c[1], fmaf(P.R[5], B.c[2], P.t[1]))) - A.c[1];
  float dz = fmaf(P.R[6], B.c[0], fmaf(P.R[7], B.c[1], fmaf(P.R[8], B.c[2], P.t[2]))) - A.c[2];
  // v' = v + w' (2u x v) + u x (2u x v) with u = -qa.xyz (conjugate), w' = qa.w
  float ux = -qa[1], uy = -qa[2], uz = -qa[3];
  float cx = fmaf(uy, dz, -uz * dy), cy = fmaf(uz, dx, -ux * dz), cz = fmaf(ux, dy, -uy * dx);
  cx += cx; cy += cy; cz += cz;
  float T0 = fmaf(qa[0], cx, dx) + fmaf(uy, cz, -uz * cy);
  float T1 = fmaf(qa[0], cy, dy) + fmaf(uz, cx, -ux * cz);
  float T2 = fmaf(qa[0], cz, dz) + fmaf(ux, cy, -uy * cx);
  float ea0, ea1, ea2, eb0, eb1, eb2, spare;
  unpack_half2(A.e01, ea0, ea1); unpack_half2(A.e2x, ea2, spare);
  unpack_half2(B.e01, eb0, eb1); unpack_half2(B.e2x, eb2, spare);
  szA = fmaf(ea0, ea0, fmaf(ea1, ea1, ea2 * ea2));
  szB = fmaf(eb0, eb0, fmaf(eb1, eb1, eb2 * eb2));
  ea0 += P.eta; ea1 += P.eta; ea2 += P.eta;
  eb0 += P.eta; eb1 += P.eta; eb2 += P.eta;
  const float B00 = Bm[0], B01 = Bm[1], B02 = Bm[2], B10 = Bm[3], B11 = Bm[4], B12 = Bm[5], B20 = Bm[6], B21 = Bm[7],
              B22 = Bm[8];
  float F00 = fabsf(B00) + reps, F01 = fabsf(B01) + reps, F02 = fabsf(B02) + reps;
  float F10 = fabsf(B10) + reps, F11 = fabsf(B11) + reps, F12 = fabsf(B12) + reps;
  float F20 = fabsf(B20) + reps, F21 = fabsf(B21) + reps, F22 = fabsf(B22) + reps;
  bool sep = false;
  sep |= fabsf(T0) > ea0 + fmaf(F00, eb0, fmaf(F01, eb1, F02 * eb2));
  sep |= fabsf(T1) > ea1 + fmaf(F10, eb0, fmaf(F11, eb1, F12 * eb2));
  sep |= fabsf(T2) > ea2 + fmaf(F20, eb0, fmaf(F21, eb1, F22 * eb2));
  sep |= fabsf(fmaf(B00, T0, fmaf(B10, T1, B20 * T2))) > eb0 + fmaf(F00, ea0, fmaf(F10, ea1, F20 * ea2));
  sep |= fabsf(fmaf(B01, T0, fmaf(B11, T1, B21 * T2))) > eb1 + fmaf(F01, ea0, fmaf(F11, ea1, F21 * ea2));
  sep |= fabsf(fmaf(B02, T0, fmaf(B12, T1, B22 * T2))) > eb2 + fmaf(F02, ea0, fmaf(F12, ea1, F22 * ea2));
  sep |= fabsf(fmaf(T2, B10, -T1 * B20)) > fmaf(ea1, F20, fmaf(ea2, F10, fmaf(eb1, F02, eb2 * F01)));
  sep |= fabsf(fmaf(T2, B11, -T1 * B21)) > fmaf(ea1, F21, fmaf(ea2, F11, fmaf(eb0, F02, eb2 * F00)));
  sep |= fabsf(fmaf(T2, B12, -T1 * B22)) > fmaf(ea1, F22, fmaf(ea2, F12, fmaf(eb0, F01, eb1 * F00)));
  sep |= fabsf(fmaf(T0, B20, -T2 * B00)) > fmaf(ea0, F20, fmaf(ea2, F00, fmaf(eb1, F12, eb2 * F11)));
  sep |= fabsf(fmaf(T0, B21, -T2 * B01)) > fmaf(ea0, F21, fmaf(ea2, F01, fmaf(eb0, F12, eb2 * F10)));
  sep |= fabsf(fmaf(T0, B22, -T2 * B02)) > fmaf(ea0, F22, fmaf(ea2, F02, fmaf(eb0, F11, eb1 * F10)));
  sep |= fabsf(fmaf(T1, B00, -T0 * B10)) > fmaf(ea0, F10, fmaf(ea1, F00, fmaf(eb1, F22, eb2 * F21)));
  sep |= fabsf(fmaf(T1, B01, -T0 * B11)) > fmaf(ea0, F11, fmaf(ea1, F01, fmaf(eb0, F22, eb2 * F20)));
  sep |= fabsf(fmaf(T1, B02, -T0 * B12)) > fmaf(ea0, F12, fmaf(ea1, F02, fmaf(eb0, F21, eb1 * F20)));
  return !sep;
}

// The same test with the two by-products the sibling-pair traversal wants (kernels_dfs.cuh):
//   slack  = min over the six FACE axes of (sum of radii - |centre distance|): how deep the boxes overlap; of two
//            overlapping sibling pairs the deeper one is expanded first, which finds the first intersecting triangle
//            pair of a colliding pose in ~40 % of the box tests of a fixed left-to-right order;
//   volA/B = volume of each box: the descent rule expands the box of larger volume (FCL's rule,
//            bvh_collision_traversal_node-inl.h:78-89, compares extent^2; the label does not depend on it).
OSR_HD bool obb_overlap_q2(const QNode& A, const QNode& B, const BoxPose& P, float& slack, float& volA, float& volB) {
  const float reps = 2e-5f;
  float qa[4], qb[4], q1[4], q2[4], Bm[9];
  quat_decode(A.q01, A.q23, qa);
  quat_decode(B.q01, B.q23, qb);
  quat_mul_f(P.q, qb, q1);
  float qac[4] = {qa[0], -qa[1], -qa[2], -qa[3]};
  quat_mul_f(qac, q1, q2);
  quat_to_rot_f(q2, Bm);
  float dx = fmaf(P.R[0], B.c[0], fmaf(P.R[1], B.c[1], fmaf(P.R[2], B.c[2], P.t[0]))) - A.c[0];
  float dy = fmaf(P.R[3], B.c[0], fmaf(P.R[4], B.c[1], fmaf(P.R[5], B.c[2], P.t[1]))) - A.c[1];
  float dz = fmaf(P.R[6], B.c[0], fmaf(P.R[7], B.c[1], fmaf(P.R[8], B.c[2], P.t[2]))) - A.c[2];
  float ux = -qa[1], uy = -qa[2], uz = -qa[3];
  float cx = fmaf(uy, dz, -uz * dy), cy = fmaf(uz, dx, -ux * dz), cz = fmaf(ux, dy, -uy * dx);
  cx += cx; cy += cy; cz += cz;
  float T0 = fmaf(qa[0], cx, dx) + fmaf(uy, cz, -uz * cy);
  float T1 = fmaf(qa[0], cy, dy) + fmaf(uz, cx, -ux * cz);
  float T2 = fmaf(qa[0], cz, dz) + fmaf(ux, cy, -uy * cx);
  float ea0, ea1, ea2, eb0, eb1, eb2, spare;
  unpack_half2(A.e01, ea0, ea1); unpack_half2(A.e2x, ea2, spare);
  unpack_half2(B.e01, eb0, eb1); unpack_half2(B.e2x, eb2, spare);
  volA = ea0 * ea1 * ea2;
  volB = eb0 * eb1 * eb2;
  ea0 += P.eta; ea1 += P.eta; ea2 += P.eta;
  eb0 += P.eta; eb1 += P.eta; eb2 += P.eta;
  const float B00 = Bm[0], B01 = Bm[1], B02 = Bm[2], B10 = Bm[3], B11 = Bm[4], B12 = Bm[5], B20 = Bm[6], B21 = Bm[7],
              B22 = Bm[8];
  float F00 = fabsf(B00) + reps, F01 = fabsf(B01) + reps, F02 = fabsf(B02) + reps;
  float F10 = fabsf(B10) + reps, F11 = fabsf(B11) + reps, F12 = fabsf(B12) + reps;
  float F20 = fabsf(B20) + reps, F21 = fabsf(B21) + reps, F22 = fabsf(B22) + reps;
  // face axes: the slack of each is kept
  float s0 = (ea0 + fmaf(F00, eb0, fmaf(F01, eb1, F02 * eb2))) - fabsf(T0);
  float s1 = (ea1 + fmaf(F10, eb0, fmaf(F11, eb1, F12 * eb2))) - fabsf(T1);
  float s2 = (ea2 + fmaf(F20, eb0, fmaf(F21, eb1, F22 * eb2))) - fabsf(T2);
  float s3 = (eb0 + fmaf(F00, ea0, fmaf(F10, ea1, F20 * ea2))) - fabsf(fmaf(B00, T0, fmaf(B10, T1, B20 * T2)));
  float s4 = (eb1 + fmaf(F01, ea0, fmaf(F11, ea1, F21 * ea2))) - fabsf(fmaf(B01, T0, fmaf(B11, T1, B21 * T2)));
  float s5 = (eb2 + fmaf(F02, ea0, fmaf(F12, ea1, F22 * ea2))) - fabsf(fmaf(B02, T0, fmaf(B12, T1, B22 * T2)));
  slack = fminf(fminf(fminf(s0, s1), fminf(s2, s3)), fminf(s4, s5));
  bool sep = slack < 0.f;
  // (the nine edge x edge axes are not optional: without them K1 runs 30 % more box tests and 4.4x the triangle tests)
  sep |= fabsf(fmaf(T2, B10, -T1 * B20)) > fmaf(ea1, F20, fmaf(ea2, F10, fmaf(eb1, F02, eb2 * F01)));
  sep |= fabsf(fmaf(T2, B11, -T1 * B21)) > fmaf(ea1, F21, fmaf(ea2, F11, fmaf(eb0, F02, eb2 * F00)));
  sep |= fabsf(fmaf(T2, B12, -T1 * B22)) > fmaf(ea1, F22, fmaf(ea2, F12, fmaf(eb0, F01, eb1 * F00)));
  sep |= fabsf(fmaf(T0, B20, -T2 * B00)) > fmaf(ea0, F20, fmaf(ea2, F00, fmaf(eb1, F12, eb2 * F11)));
  sep |= fabsf(fmaf(T0, B21, -T2 * B01)) > fmaf(ea0, F21, fmaf(ea2, F01, fmaf(eb0, F12, eb2 * F10)));
  sep |= fabsf(fmaf(T0, B22, -T2 * B02)) > fmaf(ea0, F22, fmaf(ea2, F02, fmaf(eb0, F11, eb1 * F10)));
  sep |= fabsf(fmaf(T1, B00, -T0 * B10)) > fmaf(ea0, F10, fmaf(ea1, F00, fmaf(eb1, F22, eb2 * F21)));
  sep |= fabsf(fmaf(T1, B01, -T0 * B11)) > fmaf(ea0, F11, fmaf(ea1, F01, fmaf(eb0, F22, eb2 * F20)));
  sep |= fabsf(fmaf(T1, B02, -T0 * B12)) > fmaf(ea0, F12, fmaf(ea1, F02, fmaf(eb0, F21, eb1 * F20)));
  return !sep;
}

// fp32 box pose from the fp64 relative transform `rel` (reference semantics) and the two quaternions it
// was built from (environment perturbation, robot state; neither normalised).  r_rob = largest |vertex| of
// the robot in its own frame.
OSR_HD void quat_normalize_f(const double* q, float* o) {
  float w = (float)q[0], x = (float)q[1], y = (float)q[2], z = (float)q[3];
  float n2 = fmaf(w, w, fmaf(x, x, fmaf(y, y, z * z)));
  if (!(n2 > 0.f)) { o[0] = 1.f; o[1] = 0.f; o[2] = 0.f; o[3] = 0.f; return; }
#if defined(__CUDA_ARCH__)
  float s = rsqrtf(n2);
#else
  float s = 1.0f / sqrtf(n2);
#endif
  o[0] = w * s; o[1] = x * s; o[2] = y * s; o[3] = z * s;
}
OSR_HD void make_box_pose(const Rigid& rel, const double* q_env, const double* q_rob, float r_rob, BoxPose& P) {
  // fp32 throughout: whatever rotation comes out, its distance to the reference's matrix goes into eta
  float e[4], r[4];
  quat_normalize_f(q_env, e);
  quat_normalize_f(q_rob, r);
  float ec[4] = {e[0], -e[1], -e[2], -e[3]};
  quat_mul_f(ec, r, P.q);
  quat_to_rot_f(P.q, P.R);
  float dev = 0.f;
  for (int k = 0; k < 9; k++) {
    float d = (float)rel.R[k] - P.R[k];
    dev = fmaf(d, d, dev);
  }
  for (int k = 0; k < 3; k++) P.t[k] = (float)rel.t[k];
  P.eta = 2e-5f * (1.0f + fabsf(P.t[0]) + fabsf(P.t[1]) + fabsf(P.t[2])) + 1.0001f * sqrtf(dev) * r_rob;
}

}  // namespace osr
