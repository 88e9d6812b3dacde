// se3.h -- SE(3) unit-state arithmetic shared by host code and CUDA kernels.
//
// Same semantics as the reference's lib/osr/osr_state.cc (paths relative to /root/reference):
//   state layout [tx ty tz qw qx qy qz]                     lib/osr/osr_state.h:20-33
//   state -> rigid transform (quaternion NOT renormalised)   lib/osr/osr_state.cc:16-25
//   interpolate = lerp + Eigen-3.3 slerp                     lib/osr/osr_state.cc:28-44
//   distance = |dt| + angularDistance                        lib/osr/osr_state.cc:130-154
//   apply(from, tr, angle-axis)                              lib/osr/osr_state.cc:244-259
// All arithmetic is plain fp64 multiply/add in a fixed order; this file must be compiled
// without FMA contraction (nvcc -fmad=false, g++ -ffp-contract=off) so that host, device and
// the test oracle round identically.
#pragma once
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define OSR_HD __host__ __device__ __forceinline__
#else
#define OSR_HD inline
#endif

namespace osr {

struct Rigid {   // x -> R x + t, R row-major
  double R[9];
  double t[3];
};

OSR_HD void quat_normalize(double& w, double& x, double& y, double& z) {
  double n2 = x * x + y * y + z * z + w * w;
  if (n2 > 0) {
    double n = sqrt(n2);
    w /= n; x /= n; y /= n; z /= n;
  }
}

// The same, out of line and by value, for per-segment code of the segment kernels: four fp64 divisions are ~100
// instructions inline, a segment normalises four quaternions, and code size matters there.
struct Quat4 { double w, x, y, z; };
#if defined(__CUDACC__)
static __host__ __device__ __noinline__
#else
inline
#endif
Quat4 quat_normalized(double w, double x, double y, double z) {
  quat_normalize(w, x, y, z);
  Quat4 q = {w, x, y, z};
  return q;
}

OSR_HD void quat_mul(const double* a, const double* b, double* o) {  // (w,x,y,z)
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] + a[2] * b[0] + a[3] * b[1] - a[1] * b[3];
  double z = a[0] * b[3] + a[3] * b[0] + a[1] * b[2] - a[2] * b[1];
  o[0] = w; o[1] = x; o[2] = y; o[3] = z;
}

OSR_HD void quat_to_rot(double w, double x, double y, double z, double* R) {
  double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  double twx = tx * w, twy = ty * w, twz = tz * w;
  double txx = tx * x, txy = ty * x, txz = tz * x;
  double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz;       R[2] = txz + twy;
  R[3] = txy + twz;       R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;       R[7] = tyz + twx;       R[8] = 1 - (txx + tyy);
}

OSR_HD void state_to_rigid(const double* q, Rigid& r) {
  quat_to_rot(q[3], q[4], q[5], q[6], r.R);
  r.t[0] = q[0]; r.t[1] = q[1]; r.t[2] = q[2];
}

// robot pose expressed in the environment's frame: R = R1^T R2, t = R1^T (t2 - t1)
// (what FCL's relativeTransform hands to the traversal, fcl math/geometry-inl.h:758-784).
OSR_HD void relative_rigid(const Rigid& env, const Rigid& rob, Rigid& o) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      o.R[3 * i + j] = env.R[i] * rob.R[j] + env.R[3 + i] * rob.R[3 + j] + env.R[6 + i] * rob.R[6 + j];
  double d0 = rob.t[0] - env.t[0], d1 = rob.t[1] - env.t[1], d2 = rob.t[2] - env.t[2];
  for (int i = 0; i < 3; i++) o.t[i] = env.R[i] * d0 + env.R[3 + i] * d1 + env.R[6 + i] * d2;
}

// osr::interpolate (lib/osr/osr_state.cc:28-44: lerp of the translation, both quaternions normalised, Eigen 3.3 slerp)
// in two parts: what depends on the two end states only (SlerpPlan) and what depends on tau.  interpolate() is their
// composition, operation for operation, so a walk that prepares the plan once per segment gets the same bits as one
// that calls interpolate() per step -- without two normalisations (2 sqrt, 8 divisions), acos and one sin per step.
struct SlerpPlan {
  double ta[3], tb[3];   // translations
  double qa[4], qb[4];   // normalised quaternions (w, x, y, z)
  double d;              // qa . qb
  double th, st;         // acos|d| and its sine (unused when linear)
  int linear;            // |d| >= 1 - eps: Eigen's slerp lerps the coefficients
};

OSR_HD void slerp_prepare(const double* a, const double* b, SlerpPlan& p) {
  const Quat4 qa = quat_normalized(a[3], a[4], a[5], a[6]), qb = quat_normalized(b[3], b[4], b[5], b[6]);
  const double aw = qa.w, ax = qa.x, ay = qa.y, az = qa.z, bw = qb.w, bx = qb.x, by = qb.y, bz = qb.z;
  const double one = 1.0 - 2.220446049250313e-16;
  p.d = ax * bx + ay * by + az * bz + aw * bw;
  const double ad = fabs(p.d);
  p.linear = ad >= one ? 1 : 0;
  p.th = 0.0; p.st = 1.0;
  if (!p.linear) {
    p.th = acos(ad);
    p.st = sin(p.th);
  }
  for (int k = 0; k < 3; k++) { p.ta[k] = a[k]; p.tb[k] = b[k]; }
  p.qa[0] = aw; p.qa[1] = ax; p.qa[2] = ay; p.qa[3] = az;
  p.qb[0] = bw; p.qb[1] = bx; p.qb[2] = by; p.qb[3] = bz;
}

OSR_HD void slerp_at(const SlerpPlan& p, double tau, double* o) {
  double s0, s1;
  if (p.linear) {
    s0 = 1.0 - tau; s1 = tau;
  } else {
    s0 = sin((1.0 - tau) * p.th) / p.st;
    s1 = sin(tau * p.th) / p.st;
  }
  if (p.d < 0) s1 = -s1;
  o[0] = p.ta[0] * (1 - tau) + p.tb[0] * tau;
  o[1] = p.ta[1] * (1 - tau) + p.tb[1] * tau;
  o[2] = p.ta[2] * (1 - tau) + p.tb[2] * tau;
  o[3] = s0 * p.qa[0] + s1 * p.qb[0]; o[4] = s0 * p.qa[1] + s1 * p.qb[1];
  o[5] = s0 * p.qa[2] + s1 * p.qb[2]; o[6] = s0 * p.qa[3] + s1 * p.qb[3];
}

OSR_HD void interpolate(const double* a, const double* b, double tau, double* o) {
  SlerpPlan p;
  slerp_prepare(a, b, p);
  slerp_at(p, tau, o);
}

OSR_HD double distance(const double* a, const double* b) {
  const Quat4 na = quat_normalized(a[3], a[4], a[5], a[6]), nb = quat_normalized(b[3], b[4], b[5], b[6]);
  double qa[4] = {na.w, na.x, na.y, na.z}, qb[4] = {nb.w, nb.x, nb.y, nb.z};
  double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
  double tr = sqrt(dx * dx + dy * dy + dz * dz);
  double cb[4] = {qb[0], -qb[1], -qb[2], -qb[3]}, d[4];
  quat_mul(qa, cb, d);
  double vn = sqrt(d[1] * d[1] + d[2] * d[2] + d[3] * d[3]);
  return tr + 2.0 * atan2(vn, fabs(d[0]));
}

OSR_HD void apply(const double* from, const double* tr, const double* aa, double* o) {
  double q0[4] = {from[3], from[4], from[5], from[6]};
  quat_normalize(q0[0], q0[1], q0[2], q0[3]);
  double ang = sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]);
  double ax, ay, az;
  if (ang == 0.0) { ax = 1.0; ay = 0.0; az = 0.0; }
  else { ax = aa[0] / ang; ay = aa[1] / ang; az = aa[2] / ang; }
  double s = sin(0.5 * ang);
  double qa[4] = {cos(0.5 * ang), s * ax, s * ay, s * az}, q[4];
  quat_mul(qa, q0, q);
  quat_normalize(q[0], q[1], q[2], q[3]);
  o[0] = from[0] + tr[0]; o[1] = from[1] + tr[1]; o[2] = from[2] + tr[2];
  o[3] = q[0]; o[4] = q[1]; o[5] = q[2]; o[6] = q[3];
}

// v rotated by the (not necessarily unit) quaternion q, Eigen's Quaternion * vector: v + w t + u x t, t = 2 u x v
OSR_HD void quat_rotate(const double* q, const double* v, double* o) {
  double u[3] = {q[1], q[2], q[3]};
  double uv[3] = {u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]};
  for (int k = 0; k < 3; k++) uv[k] += uv[k];
  double c[3] = {u[1] * uv[2] - u[2] * uv[1], u[2] * uv[0] - u[0] * uv[2], u[0] * uv[1] - u[1] * uv[0]};
  for (int k = 0; k < 3; k++) o[k] = v[k] + q[0] * uv[k] + c[k];
}

// Eigen Quaternion::setFromTwoVectors, the generic branch; in the antiparallel branch Eigen takes the axis from an
// SVD, here any perpendicular vector (the reference only gets there when rob_n == env_n).
OSR_HD void quat_from_two_vectors(const double* a, const double* b, double* q) {
  double la = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]), lb = sqrt(b[0] * b[0] + b[1] * b[1] + b[2] * b[2]);
  double v0[3] = {a[0] / la, a[1] / la, a[2] / la}, v1[3] = {b[0] / lb, b[1] / lb, b[2] / lb};
  double c = v1[0] * v0[0] + v1[1] * v0[1] + v1[2] * v0[2];
  if (c < -1.0 + 1e-12) {
    double h[3] = {1, 0, 0};
    if (fabs(v0[0]) > 0.9) { h[0] = 0; h[1] = 1; }
    double ax[3] = {v0[1] * h[2] - v0[2] * h[1], v0[2] * h[0] - v0[0] * h[2], v0[0] * h[1] - v0[1] * h[0]};
    double l = sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]);
    q[0] = 0; q[1] = ax[0] / l; q[2] = ax[1] / l; q[3] = ax[2] / l;
    return;
  }
  double ax[3] = {v0[1] * v1[2] - v0[2] * v1[1], v0[2] * v1[0] - v0[0] * v1[2], v0[0] * v1[1] - v0[1] * v1[0]};
  double s = sqrt((1.0 + c) * 2.0), inv = 1.0 / s;
  q[0] = s * 0.5; q[1] = ax[0] * inv; q[2] = ax[1] * inv; q[3] = ax[2] * inv;
}

// One key-configuration candidate (unit_world.cc:1340-1373): pts = rob point, rob normal, env point, env normal;
// ch / sh = cos / sin of half the spin angle about the environment normal.
OSR_HD void keyconf_state(const double* pts, double margin, double ch, double sh, double* o) {
  const double* ro = pts; const double* rn = pts + 3; const double* eo = pts + 6; const double* en = pts + 9;
  double men[3] = {-en[0], -en[1], -en[2]}, rot1[4];
  quat_from_two_vectors(rn, men, rot1);
  double eo_m[3] = {eo[0] + en[0] * margin, eo[1] + en[1] * margin, eo[2] + en[2] * margin};
  double ro_m[3] = {ro[0] + rn[0] * margin, ro[1] + rn[1] * margin, ro[2] + rn[2] * margin};
  double rot2[4] = {ch, sh * en[0], sh * en[1], sh * en[2]}, rot[4], p[3];   // AngleAxis -> Quaternion keeps the axis as is
  quat_mul(rot2, rot1, rot);
  quat_rotate(rot, ro_m, p);
  quat_normalize(rot[0], rot[1], rot[2], rot[3]);
  o[0] = eo_m[0] - p[0]; o[1] = eo_m[1] - p[1]; o[2] = eo_m[2] - p[2];
  o[3] = rot[0]; o[4] = rot[1]; o[5] = rot[2]; o[6] = rot[3];
}

OSR_HD long long dbits(double x) {
#if defined(__CUDA_ARCH__)
  return __double_as_longlong(x);
#else
  long long b;
  memcpy(&b, &x, 8);
  return b;
#endif
}
OSR_HD double bits_to_double(long long b) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double(b);
#else
  double x;
  memcpy(&x, &b, 8);
  return x;
#endif
}
OSR_HD long long imin64(long long a, long long b) { return a < b ? a : b; }
// floor(a / b) for 0 <= a < 2^53, 0 < b < 2^53: both are exact doubles, the rounded quotient is off by at most one,
// and two multiplications put it right -- the GPU has no 64-bit integer divide (a ~70-instruction subroutine).
OSR_HD long long div53(long long a, long long b) {
  long long q = (long long)((double)a / (double)b);
  if (q * b > a) q--;
  else if ((q + 1) * b <= a) q++;
  return q;
}

// The reference's running sum over `ns` steps, d <- (d < dist) ? min(dist, d + vd) : d + vd, stopping at the first
// value beyond dist -- with the result of every fp64 add reproduced exactly but not one add per step.  While d stays
// inside one binade [2^e, 2^(e+1)) it is a multiple of u = ulp(d), the exact sum d + vd rounds to the multiple of u
// nearest to it, and unless vd / u ends in exactly one half (a tie, whose rounding depends on the parity of d) that
// is d + c with the same c for every d of the binade.  So inside a binade k steps are d + k c, computed in integers;
// binade crossings, the step that reaches dist, ties and d = 0 take ordinary steps.  Returns the number of valid
// steps (delta <= dist); `last` = delta of the last valid one; `d` = the value after the last executed step.
OSR_HD int advance_chain(double& d, const double vd, const double dist, const int ns, double& last) {
  int valid = 0, left = ns;
  while (left > 0) {
    int k = 0;
    long long D = 0, Ci = 0;
    int e = 0;
    if (d > 0.0 && d < dist) {
      const long long bits = dbits(d);
      e = (int)(bits >> 52);                                      // biased exponent (d > 0: no sign bit)
      const double t = d + vd;
      const long long tb = dbits(t);
      if (e > 0 && e < 2046 && (int)(tb >> 52) == e && t <= dist) {   // normal, same binade, not past dist
        // 2 vd / u = vd 2^(53 - (e - 1023)), an exact scaling (x < 2^54 because t stays in d's binade); the power
        // of two is built from its exponent bits (e >= 53 always holds for distances a walk can have)
        const double x = e >= 53 ? vd * bits_to_double((long long)(2099 - e) << 52) : scalbn(vd, 53 - (e - 1023));
        const long long xi = (long long)x;
        const bool tie = x < 4.0e18 && (double)xi == x && (xi & 1);
        if (!tie) {
          D = (bits & 0xfffffffffffffLL) | (1LL << 52);
          Ci = ((tb & 0xfffffffffffffLL) | (1LL << 52)) - D;      // c / u  (> 0: t > d)
          if (Ci > 0) {                                           // (vd below half an ulp of d: the sum stalls)
            long long kk = div53((1LL << 53) - 1 - D, Ci);        // stay inside the binade
            const long long db = dbits(dist);
            if ((int)(db >> 52) == e) kk = imin64(kk, div53(((db & 0xfffffffffffffLL) | (1LL << 52)) - D, Ci));   // <= dist
            k = (int)imin64(kk, (long long)left);
          }
        }
      }
    }
    if (k > 0) {
      const long long M = D + (long long)k * Ci;                  // < 2^53, same exponent
      d = bits_to_double(((long long)e << 52) | (M & 0xfffffffffffffLL));
      valid += k;
      left -= k;
      last = d;
    } else {
      const double t = d + vd;
      d = (d < dist) ? fmin(dist, t) : t;
      left--;
      if (!(d <= dist)) break;
      valid++;
      last = d;
    }
  }
  return valid;
}

}  // namespace osr
