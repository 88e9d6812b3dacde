// osr_oracle.cc -- CPU ORACLE for the pyosr/UnitWorld collision hot path.
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, the smoke check in
// __graft_entry__.py and bench.py's cpu_baseline / --impl reference legs may load it.
// The product (puzzletunneldiscovery_b200/) never links, imports or calls it.
//
// It restates, in plain fp64 C++ (compile with -O3 -ffp-contract=off: no FMA, like
// an SSE2 build of FCL), the algorithm the reference runs on the CPU.  All paths are
// relative to /root/reference:
//   * unit-world geometry        lib/osr/unit_world.cc:123-161, lib/osr/bbox.h:21-46,
//                                lib/osr/cdmodel.cc:120-160, lib/osr/scene.cc:211-229
//   * OBB-BVH build              third-party/fcl/include/fcl/geometry/bvh/BVH_model-inl.h:855-925,
//                                .../geometry/bvh/detail/BV_fitter-inl.h:449-478,
//                                .../math/geometry-inl.h:293-361,501-585,590-630,1431-1515
//   * OBB-OBB SAT                .../math/bv/OBB-inl.h:384-525
//   * DFS traversal              .../narrowphase/detail/traversal/traversal_recurse-inl.h:83-128,
//                                .../traversal/collision/bvh_collision_traversal_node-inl.h:78-89
//   * triangle SAT-17            .../traversal/collision/intersect-inl.h:598-617,728-849,1086-1107
//   * contact multiplicity       .../intersect-inl.h:801-849,853-893,1033-1053
//   * SE(3) state math           lib/osr/osr_state.cc:16-44,86-154,244-259
//   * segment walk               lib/osr/unit_world.cc:395-506
//   * contact segments           lib/osr/unit_world.cc:959-1012, lib/tritri/tritri_igl.hh:15-67
//                                (Moeller97d itself is NOT restated: the reference file is
//                                compiled into oracle/_ref/libmoller97d.so and called via the
//                                function pointer set with orc_set_moller()).
//
// PARITY STATUS: the reference binary (pyosr + FCL + Eigen + assimp) cannot be built in this
// image (Eigen/assimp/libccd/glm absent), and the reference ships no golden labels for bulk
// samples.  The oracle is pinned against every known answer the reference holds
// (tests/test_oracle_known_answers.py): sancheck/fclcontacts triangle pair, the two
// "Expecting True" states of src/RL/test_collide.py, validity + connectivity of
// res/alpha/alpha-1.2.org.path, validity of TUNNEL_V on the .wt2 meshes.  Bulk labels on
// random samples are "parity unpinned" by the reference; the BVH mode is cross-checked against
// the brute-force all-pairs SAT-17 mode of this same file.  Eigen quaternion semantics
// (slerp, angularDistance, toRotationMatrix, normalize) are restated from Eigen 3.3 and are
// likewise unpinned by any reference test.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <string>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {

struct V3 {
  double x, y, z;
  double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
static inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// 3x3 row-major
struct M3 {
  double m[9];
  double operator()(int r, int c) const { return m[3 * r + c]; }
  double& operator()(int r, int c) { return m[3 * r + c]; }
};
static inline V3 mul(const M3& A, V3 v) {
  return {A(0, 0) * v.x + A(0, 1) * v.y + A(0, 2) * v.z, A(1, 0) * v.x + A(1, 1) * v.y + A(1, 2) * v.z,
          A(2, 0) * v.x + A(2, 1) * v.y + A(2, 2) * v.z};
}
static inline V3 mulT(const M3& A, V3 v) {  // A^T v
  return {A(0, 0) * v.x + A(1, 0) * v.y + A(2, 0) * v.z, A(0, 1) * v.x + A(1, 1) * v.y + A(2, 1) * v.z,
          A(0, 2) * v.x + A(1, 2) * v.y + A(2, 2) * v.z};
}
static inline M3 matmul(const M3& A, const M3& B) {
  M3 C;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) C(i, j) = A(i, 0) * B(0, j) + A(i, 1) * B(1, j) + A(i, 2) * B(2, j);
  return C;
}
static inline M3 matmulTA(const M3& A, const M3& B) {  // A^T B
  M3 C;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) C(i, j) = A(0, i) * B(0, j) + A(1, i) * B(1, j) + A(2, i) * B(2, j);
  return C;
}

// ---------------------------------------------------------------------------------------------
// SE(3) state math (lib/osr/osr_state.cc).  State = [tx ty tz qw qx qy qz].
// ---------------------------------------------------------------------------------------------
struct Quat {
  double w, x, y, z;
};
static inline Quat qnormalized(Quat q) {  // Eigen normalize(): z=squaredNorm; if z>0 coeffs/=sqrt(z)
  double z = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
  if (z > 0) {
    double n = std::sqrt(z);
    q.w /= n; q.x /= n; q.y /= n; q.z /= n;
  }
  return q;
}
static inline Quat qmul(Quat a, Quat b) {
  return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x};
}
// Eigen 3.3 QuaternionBase::toRotationMatrix -- the quaternion is NOT renormalised
// (osr_state.cc:16-25 passes the raw state quaternion).
static inline M3 qtorot(Quat q) {
  double tx = 2 * q.x, ty = 2 * q.y, tz = 2 * q.z;
  double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  M3 R;
  R(0, 0) = 1 - (tyy + tzz); R(0, 1) = txy - twz;       R(0, 2) = txz + twy;
  R(1, 0) = txy + twz;       R(1, 1) = 1 - (txx + tzz); R(1, 2) = tyz - twx;
  R(2, 0) = txz - twy;       R(2, 1) = tyz + twx;       R(2, 2) = 1 - (txx + tyy);
  return R;
}
static inline Quat qfromaa(double angle, V3 axis) {  // Eigen AngleAxis -> Quaternion
  double ha = 0.5 * angle;
  double s = std::sin(ha);
  return {std::cos(ha), s * axis.x, s * axis.y, s * axis.z};
}
// Eigen 3.3 slerp: threshold 1-eps, no renormalisation of the result.
static inline Quat qslerp(Quat a, Quat b, double t) {
  const double one = 1.0 - std::numeric_limits<double>::epsilon();
  double d = a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
  double absD = std::fabs(d);
  double s0, s1;
  if (absD >= one) {
    s0 = 1.0 - t; s1 = t;
  } else {
    double theta = std::acos(absD);
    double sinTheta = std::sin(theta);
    s0 = std::sin((1.0 - t) * theta) / sinTheta;
    s1 = std::sin(t * theta) / sinTheta;
  }
  if (d < 0) s1 = -s1;
  return {s0 * a.w + s1 * b.w, s0 * a.x + s1 * b.x, s0 * a.y + s1 * b.y, s0 * a.z + s1 * b.z};
}
// Eigen 3.3 angularDistance: d = a * conj(b); 2*atan2(|vec(d)|, |w(d)|)
static inline double qangdist(Quat a, Quat b) {
  Quat c = {b.w, -b.x, -b.y, -b.z};
  Quat d = qmul(a, c);
  double vn = std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
  return 2.0 * std::atan2(vn, std::fabs(d.w));
}

void interpolate(const double* a, const double* b, double tau, double* out) {  // osr_state.cc:28-44
  Quat qa = qnormalized({a[3], a[4], a[5], a[6]});
  Quat qb = qnormalized({b[3], b[4], b[5], b[6]});
  Quat q = qslerp(qa, qb, tau);
  for (int i = 0; i < 3; i++) out[i] = a[i] * (1 - tau) + b[i] * tau;
  out[3] = q.w; out[4] = q.x; out[5] = q.y; out[6] = q.z;
}
double distance(const double* a, const double* b) {  // osr_state.cc:130-154
  Quat qa = qnormalized({a[3], a[4], a[5], a[6]});
  Quat qb = qnormalized({b[3], b[4], b[5], b[6]});
  double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
  double tr = std::sqrt(dx * dx + dy * dy + dz * dz);
  return tr + qangdist(qa, qb);
}
void apply(const double* from, const double* tr, const double* aa, double* out) {  // osr_state.cc:244-259
  Quat q0 = qnormalized({from[3], from[4], from[5], from[6]});
  double angle = std::sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]);
  V3 axis;
  if (angle == 0.0) axis = {1.0, 0.0, 0.0};
  else axis = {aa[0] / angle, aa[1] / angle, aa[2] / angle};
  Quat q = qnormalized(qmul(qfromaa(angle, axis), q0));
  for (int i = 0; i < 3; i++) out[i] = from[i] + tr[i];
  out[3] = q.w; out[4] = q.x; out[5] = q.y; out[6] = q.z;
}

// ---------------------------------------------------------------------------------------------
// Geometry: OBJ (v/f only), the reference's unit-world transform, FCL-like OBB BVH.
// ---------------------------------------------------------------------------------------------
struct RawMesh {
  std::vector<float> V;  // referenced vertices only, xyz
  std::vector<int> F;    // triangles
};

static bool load_obj(const char* path, RawMesh& out) {
  std::ifstream in(path);
  if (!in.good()) return false;
  std::vector<float> allv;
  std::vector<int> faces;
  std::string line;
  while (std::getline(in, line)) {
    if (line.size() < 2) continue;
    if (line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) {
      const char* p = line.c_str() + 1;
      char* e;
      for (int k = 0; k < 3; k++) { allv.push_back(strtof(p, &e)); p = e; }
    } else if (line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) {
      std::istringstream ss(line.substr(1));
      std::string tok;
      std::vector<int> idx;
      while (ss >> tok) {
        int v = atoi(tok.c_str());
        int nv = (int)(allv.size() / 3);
        idx.push_back(v > 0 ? v - 1 : nv + v);
      }
      for (size_t k = 2; k < idx.size(); k++) {  // fan triangulation
        faces.push_back(idx[0]); faces.push_back(idx[k - 1]); faces.push_back(idx[k]);
      }
    }
  }
  // keep referenced vertices only, first-use order preserved by original index
  std::vector<int> remap(allv.size() / 3, -1);
  std::vector<char> used(allv.size() / 3, 0);
  for (int f : faces) used[f] = 1;
  for (size_t i = 0; i < used.size(); i++)
    if (used[i]) {
      remap[i] = (int)(out.V.size() / 3);
      out.V.push_back(allv[3 * i]); out.V.push_back(allv[3 * i + 1]); out.V.push_back(allv[3 * i + 2]);
    }
  for (int f : faces) out.F.push_back(remap[f]);
  return !out.F.empty();
}

// lib/osr/bbox.h:21-46: box is seeded at the origin, span = max extent (float arithmetic).
static float bbox_span(const std::vector<float>& V) {
  float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
  for (size_t i = 0; i < V.size(); i += 3)
    for (int k = 0; k < 3; k++) {
      hi[k] = std::max(V[i + k], hi[k]);
      lo[k] = std::min(V[i + k], lo[k]);
    }
  return std::max(std::max(hi[0] - lo[0], hi[1] - lo[1]), hi[2] - lo[2]);
}
static void mean_center(const std::vector<float>& V, float c[3]) {  // scene.cc:40-50 (float sum)
  float s[3] = {0, 0, 0};
  size_t n = V.size() / 3;
  for (size_t i = 0; i < n; i++)
    for (int k = 0; k < 3; k++) s[k] += V[3 * i + k];
  for (int k = 0; k < 3; k++) c[k] = s[k] / (float)n;
}

struct Node {       // one OBB node; children at first_child, first_child+1
  M3 axis;          // columns are the box axes (axis(r,c): component r of axis c)
  V3 To;            // centre
  V3 ext;           // half extents
  int first_child;  // >=0: index of left child ; <0 : leaf, primitive = -(first_child+1)
  bool leaf() const { return first_child < 0; }
  int prim() const { return -(first_child + 1); }
  double size() const { return ext.x * ext.x + ext.y * ext.y + ext.z * ext.z; }
};

struct Model {
  std::vector<V3> V;     // unit-space vertices, fp64
  std::vector<int> F;    // 3 per triangle
  std::vector<Node> bvh; // 2F-1 nodes, root 0
  int ntri() const { return (int)(F.size() / 3); }
};

// Cyclic Jacobi eigen-solver for a symmetric 3x3 (same family as FCL's eigen_old: rows of
// `vec` are eigenvectors after transposition below).  Any orthonormal eigenbasis gives a valid
// (conservative) OBB, labels do not depend on it.
static void sym_eigen3(const M3& A, double eval[3], M3& evec) {
  double a[3][3], v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) a[i][j] = A(i, j);
  for (int sweep = 0; sweep < 50; sweep++) {
    double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
    if (off == 0.0) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        if (a[p][q] == 0.0) continue;
        double theta = 0.5 * (a[q][q] - a[p][p]) / a[p][q];
        double t = 1.0 / (std::fabs(theta) + std::sqrt(1.0 + theta * theta));
        if (theta < 0) t = -t;
        double c = 1.0 / std::sqrt(1 + t * t), s = t * c;
        for (int k = 0; k < 3; k++) {  // A <- A J
          double akp = a[k][p], akq = a[k][q];
          a[k][p] = c * akp - s * akq; a[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; k++) {  // A <- J^T A
          double apk = a[p][k], aqk = a[q][k];
          a[p][k] = c * apk - s * aqk; a[q][k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 3; k++) {
          double vkp = v[k][p], vkq = v[k][q];
          v[k][p] = c * vkp - s * vkq; v[k][q] = s * vkp + c * vkq;
        }
      }
  }
  for (int i = 0; i < 3; i++) {
    eval[i] = a[i][i];
    for (int j = 0; j < 3; j++) evec(i, j) = v[i][j];  // column i = eigenvector i
  }
}

static void fit_obb(const Model& M, const int* prims, int n, Node& nd) {
  // covariance over the 3n triangle vertices (geometry-inl.h:1431-1515)
  double S1[3] = {0, 0, 0}, S2[3][3] = {{0}};
  for (int i = 0; i < n; i++)
    for (int j = 0; j < 3; j++) {
      V3 p = M.V[M.F[3 * prims[i] + j]];
      double pp[3] = {p.x, p.y, p.z};
      for (int a = 0; a < 3; a++) {
        S1[a] += pp[a];
        for (int b = a; b < 3; b++) S2[a][b] += pp[a] * pp[b];
      }
    }
  M3 C;
  double np = 3.0 * n;
  for (int a = 0; a < 3; a++)
    for (int b = a; b < 3; b++) C(a, b) = C(b, a) = S2[a][b] - S1[a] * S1[b] / np;
  double ev[3];
  M3 E;
  sym_eigen3(C, ev, E);
  // axisFromEigen (geometry-inl.h:590-630): axis0 = largest, axis1 = middle, axis2 = axis0 x axis1
  int mn, md, mx;
  if (ev[0] > ev[1]) { mx = 0; mn = 1; } else { mn = 0; mx = 1; }
  if (ev[2] < ev[mn]) { md = mn; mn = 2; } else if (ev[2] > ev[mx]) { md = mx; mx = 2; } else { md = 2; }
  V3 a0 = {E(0, mx), E(1, mx), E(2, mx)}, a1 = {E(0, md), E(1, md), E(2, md)};
  V3 a2 = cross(a0, a1);
  V3 ax[3] = {a0, a1, a2};
  for (int c = 0; c < 3; c++) { nd.axis(0, c) = ax[c].x; nd.axis(1, c) = ax[c].y; nd.axis(2, c) = ax[c].z; }
  // getExtentAndCenter_mesh (geometry-inl.h:293-361)
  double lo[3], hi[3];
  for (int k = 0; k < 3; k++) { lo[k] = std::numeric_limits<double>::max(); hi[k] = -lo[k]; }
  for (int i = 0; i < n; i++)
    for (int j = 0; j < 3; j++) {
      V3 p = M.V[M.F[3 * prims[i] + j]];
      for (int k = 0; k < 3; k++) {
        double pr = dot(ax[k], p);
        if (pr > hi[k]) hi[k] = pr;
        if (pr < lo[k]) lo[k] = pr;
      }
    }
  V3 o = {(hi[0] + lo[0]) / 2, (hi[1] + lo[1]) / 2, (hi[2] + lo[2]) / 2};
  nd.To = mul(nd.axis, o);
  nd.ext = {(hi[0] - lo[0]) / 2, (hi[1] - lo[1]) / 2, (hi[2] - lo[2]) / 2};
}

// BVH_model-inl.h:855-925 with SPLIT_METHOD_MEAN (BV_splitter-inl.h:422-435,586-620):
// split vector = OBB axis 0, split value = mean projected centroid, ">" goes right.
static void build_rec(Model& M, std::vector<int>& prims, int bv_id, int first, int n) {
  Node nd;
  fit_obb(M, prims.data() + first, n, nd);
  if (n == 1) {
    nd.first_child = -(prims[first] + 1);
    M.bvh[bv_id] = nd;
    return;
  }
  int fc = (int)M.bvh.size();
  nd.first_child = fc;
  M.bvh[bv_id] = nd;
  M.bvh.resize(M.bvh.size() + 2);
  V3 sv = {nd.axis(0, 0), nd.axis(1, 0), nd.axis(2, 0)};
  double sum = 0;
  std::vector<V3> cent(n);
  for (int i = 0; i < n; i++) {
    int t = prims[first + i];
    V3 p1 = M.V[M.F[3 * t]], p2 = M.V[M.F[3 * t + 1]], p3 = M.V[M.F[3 * t + 2]];
    cent[i] = {(p1.x + p2.x + p3.x) / 3.0, (p1.y + p2.y + p3.y) / 3.0, (p1.z + p2.z + p3.z) / 3.0};
    sum += dot(V3{p1.x + p2.x + p3.x, p1.y + p2.y + p3.y, p1.z + p2.z + p3.z}, sv);
  }
  double split = sum / (3.0 * n);
  int c1 = 0;
  for (int i = 0; i < n; i++) {
    if (dot(sv, cent[i]) > split) {
      // right side
    } else {
      std::swap(prims[first + i], prims[first + c1]);
      std::swap(cent[i], cent[c1]);
      c1++;
    }
  }
  if (c1 == 0 || c1 == n) c1 = n / 2;
  build_rec(M, prims, fc, first, c1);
  build_rec(M, prims, fc + 1, first + c1, n - c1);
}
static void build_bvh(Model& M) {
  int n = M.ntri();
  std::vector<int> prims(n);
  for (int i = 0; i < n; i++) prims[i] = i;
  M.bvh.clear();
  M.bvh.reserve(2 * n);
  M.bvh.resize(1);
  build_rec(M, prims, 0, 0, n);
}

// ---------------------------------------------------------------------------------------------
// Collision: obbDisjoint, SAT-17, DFS.
// ---------------------------------------------------------------------------------------------
// OBB-inl.h:399-525.  B = rotation of box2 in box1's frame, T = centre of box2 in box1's frame.
static bool obb_disjoint(const M3& B, V3 Tv, V3 av, V3 bv) {
  const double reps = 1e-6;
  const double T[3] = {Tv.x, Tv.y, Tv.z}, a[3] = {av.x, av.y, av.z}, b[3] = {bv.x, bv.y, bv.z};
  double Bf[3][3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) Bf[i][j] = std::fabs(B(i, j)) + reps;
  double t, s;
  // A0
  t = std::fabs(T[0]);
  if (t > (a[0] + (Bf[0][0] * b[0] + Bf[0][1] * b[1] + Bf[0][2] * b[2]))) return true;
  // B0
  s = B(0, 0) * T[0] + B(1, 0) * T[1] + B(2, 0) * T[2];
  t = std::fabs(s);
  if (t > (b[0] + (Bf[0][0] * a[0] + Bf[1][0] * a[1] + Bf[2][0] * a[2]))) return true;
  // A1
  t = std::fabs(T[1]);
  if (t > (a[1] + (Bf[1][0] * b[0] + Bf[1][1] * b[1] + Bf[1][2] * b[2]))) return true;
  // A2
  t = std::fabs(T[2]);
  if (t > (a[2] + (Bf[2][0] * b[0] + Bf[2][1] * b[1] + Bf[2][2] * b[2]))) return true;
  // B1
  s = B(0, 1) * T[0] + B(1, 1) * T[1] + B(2, 1) * T[2];
  t = std::fabs(s);
  if (t > (b[1] + (Bf[0][1] * a[0] + Bf[1][1] * a[1] + Bf[2][1] * a[2]))) return true;
  // B2
  s = B(0, 2) * T[0] + B(1, 2) * T[1] + B(2, 2) * T[2];
  t = std::fabs(s);
  if (t > (b[2] + (Bf[0][2] * a[0] + Bf[1][2] * a[1] + Bf[2][2] * a[2]))) return true;
  // A0 x B0
  s = T[2] * B(1, 0) - T[1] * B(2, 0); t = std::fabs(s);
  if (t > (a[1] * Bf[2][0] + a[2] * Bf[1][0] + b[1] * Bf[0][2] + b[2] * Bf[0][1])) return true;
  // A0 x B1
  s = T[2] * B(1, 1) - T[1] * B(2, 1); t = std::fabs(s);
  if (t > (a[1] * Bf[2][1] + a[2] * Bf[1][1] + b[0] * Bf[0][2] + b[2] * Bf[0][0])) return true;
  // A0 x B2
  s = T[2] * B(1, 2) - T[1] * B(2, 2); t = std::fabs(s);
  if (t > (a[1] * Bf[2][2] + a[2] * Bf[1][2] + b[0] * Bf[0][1] + b[1] * Bf[0][0])) return true;
  // A1 x B0
  s = T[0] * B(2, 0) - T[2] * B(0, 0); t = std::fabs(s);
  if (t > (a[0] * Bf[2][0] + a[2] * Bf[0][0] + b[1] * Bf[1][2] + b[2] * Bf[1][1])) return true;
  // A1 x B1
  s = T[0] * B(2, 1) - T[2] * B(0, 1); t = std::fabs(s);
  if (t > (a[0] * Bf[2][1] + a[2] * Bf[0][1] + b[0] * Bf[1][2] + b[2] * Bf[1][0])) return true;
  // A1 x B2
  s = T[0] * B(2, 2) - T[2] * B(0, 2); t = std::fabs(s);
  if (t > (a[0] * Bf[2][2] + a[2] * Bf[0][2] + b[0] * Bf[1][1] + b[1] * Bf[1][0])) return true;
  // A2 x B0
  s = T[1] * B(0, 0) - T[0] * B(1, 0); t = std::fabs(s);
  if (t > (a[0] * Bf[1][0] + a[1] * Bf[0][0] + b[1] * Bf[2][2] + b[2] * Bf[2][1])) return true;
  // A2 x B1
  s = T[1] * B(0, 1) - T[0] * B(1, 1); t = std::fabs(s);
  if (t > (a[0] * Bf[1][1] + a[1] * Bf[0][1] + b[0] * Bf[2][2] + b[2] * Bf[2][0])) return true;
  // A2 x B2
  s = T[1] * B(0, 2) - T[0] * B(1, 2); t = std::fabs(s);
  if (t > (a[0] * Bf[1][2] + a[1] * Bf[0][2] + b[0] * Bf[2][1] + b[1] * Bf[2][0])) return true;
  return false;
}

// OBB-inl.h:384-395
static bool obb_overlap(const M3& R0, V3 T0, const Node& b1, const Node& b2) {
  M3 R0b2 = matmul(R0, b2.axis);
  M3 R = matmulTA(b1.axis, R0b2);
  V3 Ttemp = (mul(R0, b2.To) + T0) - b1.To;
  V3 T = mulT(b1.axis, Ttemp);
  return !obb_disjoint(R, T, b1.ext, b2.ext);
}

// intersect-inl.h:1086-1107
static inline int project6(V3 ax, V3 p1, V3 p2, V3 p3, V3 q1, V3 q2, V3 q3) {
  double P1 = dot(ax, p1), P2 = dot(ax, p2), P3 = dot(ax, p3);
  double Q1 = dot(ax, q1), Q2 = dot(ax, q2), Q3 = dot(ax, q3);
  double mn1 = std::min(P1, std::min(P2, P3));
  double mx2 = std::max(Q1, std::max(Q2, Q3));
  if (mn1 > mx2) return 0;
  double mx1 = std::max(P1, std::max(P2, P3));
  double mn2 = std::min(Q1, std::min(Q2, Q3));
  if (mn2 > mx1) return 0;
  return 1;
}

// intersect-inl.h:728-799: Q already in model-1 frame.
bool sat17(V3 P1, V3 P2, V3 P3, V3 Q1, V3 Q2, V3 Q3) {
  V3 p1 = P1 - P1, p2 = P2 - P1, p3 = P3 - P1;
  V3 q1 = Q1 - P1, q2 = Q2 - P1, q3 = Q3 - P1;
  V3 e1 = p2 - p1, e2 = p3 - p2;
  V3 n1 = cross(e1, e2);
  if (!project6(n1, p1, p2, p3, q1, q2, q3)) return false;
  V3 f1 = q2 - q1, f2 = q3 - q2;
  V3 m1 = cross(f1, f2);
  if (!project6(m1, p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e1, f1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e1, f2), p1, p2, p3, q1, q2, q3)) return false;
  V3 f3 = q1 - q3;
  if (!project6(cross(e1, f3), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e2, f1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e2, f2), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e2, f3), p1, p2, p3, q1, q2, q3)) return false;
  V3 e3 = p1 - p3;
  if (!project6(cross(e3, f1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e3, f2), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e3, f3), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e1, n1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e2, n1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(e3, n1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(f1, m1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(f2, m1), p1, p2, p3, q1, q2, q3)) return false;
  if (!project6(cross(f3, m1), p1, p2, p3, q1, q2, q3)) return false;
  return true;
}

// Contact multiplicity of an intersecting pair in enable_contact mode
// (intersect-inl.h:801-849, computeDeepestPoints :853-893, buildTrianglePlane :1040-1053).
// FCL emits one Contact (hence one facePairs row, cdmodel.cc:213-227) per contact point.
static bool tri_plane(V3 v1, V3 v2, V3 v3, V3& n, double& t) {
  V3 n_ = cross(v2 - v1, v3 - v1);
  double sq = dot(n_, n_);
  if (sq > 0) {
    double l = std::sqrt(sq);
    n = {n_.x / l, n_.y / l, n_.z / l};
    t = dot(n, v1);
    return true;
  }
  return false;
}
static void deepest_points(const V3* pts, V3 n, double t, double* depth, unsigned* num) {
  const double eps = 1e-5;
  double max_depth = -std::numeric_limits<double>::max();
  unsigned cnt = 0, nneg = 0, npos = 0, nzero = 0;
  for (int i = 0; i < 3; i++) {
    double dist = -(dot(n, pts[i]) - t);
    if (dist > eps) npos++; else if (dist < -eps) nneg++; else nzero++;
    if (dist > max_depth) { max_depth = dist; cnt = 1; }
    else if (dist + 1e-6 >= max_depth) cnt++;
  }
  if (max_depth < -eps) cnt = 0;
  if (nzero == 0 && (nneg == 0 || npos == 0)) cnt = 0;
  *depth = max_depth;
  *num = cnt;
}
unsigned contact_multiplicity(V3 P1, V3 P2, V3 P3, V3 Q1, V3 Q2, V3 Q3) {
  V3 n1 = {0, 0, 0}, n2 = {0, 0, 0};
  double t1 = 0, t2 = 0;
  // the reference ignores buildTrianglePlane's return value; n/t stay uninitialised (here: 0)
  tri_plane(P1, P2, P3, n1, t1);
  tri_plane(Q1, Q2, Q3, n2, t2);
  V3 P[3] = {P1, P2, P3}, Q[3] = {Q1, Q2, Q3};
  double d1, d2;
  unsigned k1, k2;
  deepest_points(Q, n1, t1, &d2, &k2);
  deepest_points(P, n2, t2, &d1, &k1);
  if (d1 > d2) return std::min(k2, 2u);
  return std::min(k1, 2u);
}

struct Counters {
  long long bv = 0, leaf = 0;
};

struct Collider {
  const Model* m1;  // env
  const Model* m2;  // robot
  M3 R;
  V3 T;
  Counters cnt;
  bool want_all = false;
  bool hit = false;
  std::vector<int>* pairs = nullptr;  // (b1,b2) rows, one per contact point (FCL semantics)

  void leaf_test(int b1, int b2) {
    cnt.leaf++;
    int t1 = m1->bvh[b1].prim(), t2 = m2->bvh[b2].prim();
    V3 p1 = m1->V[m1->F[3 * t1]], p2 = m1->V[m1->F[3 * t1 + 1]], p3 = m1->V[m1->F[3 * t1 + 2]];
    V3 q1 = mul(R, m2->V[m2->F[3 * t2]]) + T;
    V3 q2 = mul(R, m2->V[m2->F[3 * t2 + 1]]) + T;
    V3 q3 = mul(R, m2->V[m2->F[3 * t2 + 2]]) + T;
    if (sat17(p1, p2, p3, q1, q2, q3)) {
      hit = true;
      if (pairs) {
        unsigned k = contact_multiplicity(p1, p2, p3, q1, q2, q3);
        for (unsigned i = 0; i < k; i++) { pairs->push_back(t1); pairs->push_back(t2); }
      }
    }
  }
  bool can_stop() const { return !want_all && hit; }
  // traversal_recurse-inl.h:83-128
  void recurse(int b1, int b2) {
    const Node& n1 = m1->bvh[b1];
    const Node& n2 = m2->bvh[b2];
    bool l1 = n1.leaf(), l2 = n2.leaf();
    cnt.bv++;
    bool disjoint = !obb_overlap(R, T, n1, n2);
    if (l1 && l2) {
      if (disjoint) return;
      leaf_test(b1, b2);
      return;
    }
    if (disjoint) return;
    // firstOverSecond: bvh_collision_traversal_node-inl.h:78-89
    if (l2 || (!l1 && (n1.size() > n2.size()))) {
      recurse(n1.first_child, b2);
      if (can_stop()) return;
      recurse(n1.first_child + 1, b2);
    } else {
      recurse(b1, n2.first_child);
      if (can_stop()) return;
      recurse(b1, n2.first_child + 1);
    }
  }
};

// ---------------------------------------------------------------------------------------------
// UnitWorld restatement.
// ---------------------------------------------------------------------------------------------
typedef int (*moller_fn)(const double*, const double*, const double*, const double*, const double*,
                         const double*, int*, double*, double*);
static moller_fn g_moller = nullptr;

struct World {
  RawMesh raw_env, raw_rob;
  bool has_env = false, has_rob = false;
  bool center_enforced = false;
  float center[3] = {0, 0, 0};
  float scene_scale = 1.0f;
  double perturb[7] = {0, 0, 0, 1, 0, 0, 0};
  Model env, rob;
  bool finalized = false;

  // unit_world.cc:123-161 + cdmodel.cc:120-160 (float glm matrix widened to double)
  void finalize() {
    float es = bbox_span(raw_env.V);
    float rs = has_rob ? bbox_span(raw_rob.V) : 1.0f;
    scene_scale = (float)(1.0 / std::max(es, rs));
    if (!center_enforced) mean_center(raw_rob.V, center);
    double s = (double)scene_scale;
    env.V.clear(); rob.V.clear();
    for (size_t i = 0; i < raw_env.V.size(); i += 3)
      env.V.push_back({s * (double)raw_env.V[i] + 0.0, s * (double)raw_env.V[i + 1] + 0.0,
                       s * (double)raw_env.V[i + 2] + 0.0});
    env.F = raw_env.F;
    float tr[3];
    for (int k = 0; k < 3; k++) tr[k] = scene_scale * (-center[k]);  // float product (glm mat4*mat4)
    for (size_t i = 0; i < raw_rob.V.size(); i += 3)
      rob.V.push_back({s * (double)raw_rob.V[i] + (double)tr[0], s * (double)raw_rob.V[i + 1] + (double)tr[1],
                       s * (double)raw_rob.V[i + 2] + (double)tr[2]});
    rob.F = raw_rob.F;
    build_bvh(env);
    build_bvh(rob);
    finalized = true;
  }


  // unit_world.cc:249-264 -> cdmodel.cc:180-197 -> FCL boxBoxIntersect
  // (third-party/fcl/include/fcl/narrowphase/detail/primitive_shape_algorithm/box_box-inl.h:827-1117,1395-1412):
  // the two local AABBs as boxes, robot first; 15-axis SAT, +1e-6 on |R| for the edge axes only.
  bool is_disentangled(const double* q) const {
    double lo1[3], hi1[3], lo2[3], hi2[3];
    for (int k = 0; k < 3; k++) { lo1[k] = lo2[k] = std::numeric_limits<double>::max(); hi1[k] = hi2[k] = -lo1[k]; }
    for (const V3& v : rob.V) for (int k = 0; k < 3; k++) { lo1[k] = std::min(lo1[k], v[k]); hi1[k] = std::max(hi1[k], v[k]); }
    for (const V3& v : env.V) for (int k = 0; k < 3; k++) { lo2[k] = std::min(lo2[k], v[k]); hi2[k] = std::max(hi2[k], v[k]); }
    V3 c1 = {(lo1[0] + hi1[0]) * 0.5, (lo1[1] + hi1[1]) * 0.5, (lo1[2] + hi1[2]) * 0.5};
    V3 c2 = {(lo2[0] + hi2[0]) * 0.5, (lo2[1] + hi2[1]) * 0.5, (lo2[2] + hi2[2]) * 0.5};
    double A[3], B[3];
    for (int k = 0; k < 3; k++) { A[k] = (hi1[k] - lo1[k]) * 0.5; B[k] = (hi2[k] - lo2[k]) * 0.5; }
    M3 R1 = qtorot({q[3], q[4], q[5], q[6]});
    M3 R2 = qtorot({perturb[3], perturb[4], perturb[5], perturb[6]});
    V3 t1 = mul(R1, c1) + V3{q[0], q[1], q[2]};
    V3 t2 = mul(R2, c2) + V3{perturb[0], perturb[1], perturb[2]};
    V3 p = t2 - t1;
    V3 ppv = mulT(R1, p);
    double pp[3] = {ppv.x, ppv.y, ppv.z};
    M3 R = matmulTA(R1, R2);
    double Q[3][3];
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) Q[a][b] = std::fabs(R(a, b));
    for (int a = 0; a < 3; a++)
      if (std::fabs(pp[a]) - ((Q[a][0] * B[0] + Q[a][1] * B[1] + Q[a][2] * B[2]) + A[a]) > 0) return true;
    for (int b = 0; b < 3; b++) {
      double tmp = R2(0, b) * p.x + R2(1, b) * p.y + R2(2, b) * p.z;
      if (std::fabs(tmp) - ((Q[0][b] * A[0] + Q[1][b] * A[1] + Q[2][b] * A[2]) + B[b]) > 0) return true;
    }
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) Q[a][b] += 1.0e-6;
#define ORC_EDGE(t, rhs) if (std::fabs(t) - (rhs) > 0) return true
    ORC_EDGE(pp[2] * R(1, 0) - pp[1] * R(2, 0), A[1] * Q[2][0] + A[2] * Q[1][0] + B[1] * Q[0][2] + B[2] * Q[0][1]);
    ORC_EDGE(pp[2] * R(1, 1) - pp[1] * R(2, 1), A[1] * Q[2][1] + A[2] * Q[1][1] + B[0] * Q[0][2] + B[2] * Q[0][0]);
    ORC_EDGE(pp[2] * R(1, 2) - pp[1] * R(2, 2), A[1] * Q[2][2] + A[2] * Q[1][2] + B[0] * Q[0][1] + B[1] * Q[0][0]);
    ORC_EDGE(pp[0] * R(2, 0) - pp[2] * R(0, 0), A[0] * Q[2][0] + A[2] * Q[0][0] + B[1] * Q[1][2] + B[2] * Q[1][1]);
    ORC_EDGE(pp[0] * R(2, 1) - pp[2] * R(0, 1), A[0] * Q[2][1] + A[2] * Q[0][1] + B[0] * Q[1][2] + B[2] * Q[1][0]);
    ORC_EDGE(pp[0] * R(2, 2) - pp[2] * R(0, 2), A[0] * Q[2][2] + A[2] * Q[0][2] + B[0] * Q[1][1] + B[1] * Q[1][0]);
    ORC_EDGE(pp[1] * R(0, 0) - pp[0] * R(1, 0), A[0] * Q[1][0] + A[1] * Q[0][0] + B[1] * Q[2][2] + B[2] * Q[2][1]);
    ORC_EDGE(pp[1] * R(0, 1) - pp[0] * R(1, 1), A[0] * Q[1][1] + A[1] * Q[0][1] + B[0] * Q[2][2] + B[2] * Q[2][0]);
    ORC_EDGE(pp[1] * R(0, 2) - pp[0] * R(1, 2), A[0] * Q[1][2] + A[1] * Q[0][2] + B[0] * Q[2][1] + B[1] * Q[2][0]);
#undef ORC_EDGE
    return false;
  }

  // unit_world.cc:197-212 + geometry-inl.h:758-784
  void rel_tf(const double* q, M3& R, V3& T) const {
    M3 R1 = qtorot({perturb[3], perturb[4], perturb[5], perturb[6]});
    M3 R2 = qtorot({q[3], q[4], q[5], q[6]});
    V3 t1 = {perturb[0], perturb[1], perturb[2]}, t2 = {q[0], q[1], q[2]};
    R = matmulTA(R1, R2);
    T = mulT(R1, t2 - t1);
  }

  bool is_valid(const double* q, Counters* c = nullptr) const {  // unit_world.cc:215-247
    Collider col;
    col.m1 = &env; col.m2 = &rob;
    rel_tf(q, col.R, col.T);
    col.recurse(0, 0);
    if (c) { c->bv += col.cnt.bv; c->leaf += col.cnt.leaf; }
    return !col.hit;
  }
  bool is_valid_brute(const double* q) const {  // ground truth: all pairs SAT-17
    M3 R; V3 T;
    rel_tf(q, R, T);
    int n2 = rob.ntri(), n1 = env.ntri();
    std::vector<V3> Q(3 * (size_t)n2);
    for (int j = 0; j < 3 * n2; j++) Q[j] = mul(R, rob.V[rob.F[j]]) + T;
    for (int i = 0; i < n1; i++) {
      V3 p1 = env.V[env.F[3 * i]], p2 = env.V[env.F[3 * i + 1]], p3 = env.V[env.F[3 * i + 2]];
      for (int j = 0; j < n2; j++)
        if (sat17(p1, p2, p3, Q[3 * j], Q[3 * j + 1], Q[3 * j + 2])) return false;
    }
    return true;
  }

  // unit_world.cc:405-474.  out: last_free[7], first_hit[7], is_free, free_tau, touch_tau
  // steps_out = number of isValid calls made.
  void transit(const double* from, const double* to, double vd, double* last_free, double* state_out,
               uint8_t* is_free, double* free_tau, double* touch_tau, long long* steps_out,
               Counters* c = nullptr) const {
    double dist = distance(from, to);
    long long steps = 0;
    if (vd >= dist) {
      memcpy(last_free, from, 56); memcpy(state_out, to, 56);
      *is_free = 0; *free_tau = 0.0; *touch_tau = 1.0;
      if (steps_out) *steps_out = 0;
      return;
    }
    double lf[7], st[7];
    memcpy(lf, from, 56); memcpy(st, from, 56);
    double delta = vd, inv_dist = 1.0 / dist, last_tau = 0.0;
    while (delta <= dist) {
      double tau = delta * inv_dist;
      interpolate(from, to, tau, st);
      steps++;
      if (!is_valid(st, c)) {
        memcpy(last_free, lf, 56); memcpy(state_out, st, 56);
        *is_free = 0; *free_tau = last_tau; *touch_tau = tau;
        if (steps_out) *steps_out = steps;
        return;
      }
      last_tau = tau;
      memcpy(lf, st, 56);
      if (delta < dist) { delta += vd; delta = std::min(dist, delta); }
      else delta += vd;
    }
    memcpy(last_free, lf, 56); memcpy(state_out, st, 56);
    *is_free = (delta > dist) ? 1 : 0; *free_tau = last_tau;
    *touch_tau = 0.0;  // the reference returns the shadowed outer `tau` (unit_world.cc:452-455,473)
    if (steps_out) *steps_out = steps;
  }
  bool is_valid_transition(const double* from, const double* to, double ivd, long long* steps,
                           Counters* c = nullptr) const {  // unit_world.cc:477-486
    double d = distance(from, to);
    double vd = std::min(d / 2.0, ivd);
    double lf[7], st[7], ft, tt; uint8_t fr;
    transit(from, to, vd, lf, st, &fr, &ft, &tt, steps, c);
    return fr != 0;
  }

  // cdmodel.cc:200-229 + unit_world.cc:959-1012.  Returns rows; fills pairs (env_tri, rob_tri),
  // segs (6 per row), coplanar flags.
  int intersecting_segments(const double* q, std::vector<int>& pairs, std::vector<double>& segs,
                            std::vector<uint8_t>& cop, Counters* c = nullptr) const {
    Collider col;
    col.m1 = &env; col.m2 = &rob;
    rel_tf(q, col.R, col.T);
    col.want_all = true;
    col.pairs = &pairs;
    size_t first = pairs.size() / 2;
    col.recurse(0, 0);
    if (c) { c->bv += col.cnt.bv; c->leaf += col.cnt.leaf; }
    size_t rows = pairs.size() / 2 - first;
    // world-space vertices: env_tf * V, rob_tf * V (Eigen affine: R*v + t)
    M3 R1 = qtorot({perturb[3], perturb[4], perturb[5], perturb[6]});
    M3 R2 = qtorot({q[3], q[4], q[5], q[6]});
    V3 t1 = {perturb[0], perturb[1], perturb[2]}, t2 = {q[0], q[1], q[2]};
    for (size_t r = first; r < first + rows; r++) {
      int te = pairs[2 * r], tr = pairs[2 * r + 1];
      double v[3][3], u[3][3];
      for (int k = 0; k < 3; k++) {
        V3 a = mul(R1, env.V[env.F[3 * te + k]]) + t1;
        V3 b = mul(R2, rob.V[rob.F[3 * tr + k]]) + t2;
        v[k][0] = a.x; v[k][1] = a.y; v[k][2] = a.z;
        u[k][0] = b.x; u[k][1] = b.y; u[k][2] = b.z;
      }
      int coplanar = 0;
      double p0[3] = {0, 0, 0}, p1[3] = {0, 0, 0};
      if (g_moller) g_moller(v[0], v[1], v[2], u[0], u[1], u[2], &coplanar, p0, p1);
      // Moller97d.cc:671-672 ignores the "coplanar" return of the second interval computation: when
      // all three snapped distances of U to V's plane are zero (but V is not in U's plane) the
      // reference copies uninitialised stack memory into the end points.  That output is
      // unspecified; oracle and engine both define it as the zero segment.
      if (!coplanar) {
        double e1[3], e2[3], n1[3];
        for (int k = 0; k < 3; k++) { e1[k] = v[1][k] - v[0][k]; e2[k] = v[2][k] - v[0][k]; }
        n1[0] = e1[1] * e2[2] - e1[2] * e2[1]; n1[1] = e1[2] * e2[0] - e1[0] * e2[2]; n1[2] = e1[0] * e2[1] - e1[1] * e2[0];
        double d1 = -(n1[0] * v[0][0] + n1[1] * v[0][1] + n1[2] * v[0][2]);
        bool allzero = true;
        for (int k = 0; k < 3; k++) {
          double du = (n1[0] * u[k][0] + n1[1] * u[k][1] + n1[2] * u[k][2]) + d1;
          if (!((du >= 0 ? du : -du) < 0.000001)) allzero = false;
        }
        if (allzero) for (int k = 0; k < 3; k++) p0[k] = p1[k] = 0.0;
      }
      for (int k = 0; k < 3; k++) segs.push_back(p0[k]);
      for (int k = 0; k < 3; k++) segs.push_back(p1[k]);
      cop.push_back((uint8_t)(coplanar != 0));
    }
    return (int)rows;
  }
};

}  // namespace orc

// ---------------------------------------------------------------------------------------------
// C API for ctypes (oracle/oracle.py).
// ---------------------------------------------------------------------------------------------
using orc::World;
extern "C" {
void* orc_create() { return new World(); }
void orc_destroy(void* h) { delete (World*)h; }
int orc_load_env_obj(void* h, const char* p) { World* w = (World*)h; w->raw_env = {}; w->has_env = orc::load_obj(p, w->raw_env); return w->has_env ? 0 : -1; }
int orc_load_rob_obj(void* h, const char* p) { World* w = (World*)h; w->raw_rob = {}; w->has_rob = orc::load_obj(p, w->raw_rob); return w->has_rob ? 0 : -1; }
int orc_set_meshes(void* h, const float* Ve, int nVe, const int* Fe, int nFe, const float* Vr, int nVr,
                   const int* Fr, int nFr) {
  World* w = (World*)h;
  w->raw_env.V.assign(Ve, Ve + 3 * (size_t)nVe); w->raw_env.F.assign(Fe, Fe + 3 * (size_t)nFe);
  w->raw_rob.V.assign(Vr, Vr + 3 * (size_t)nVr); w->raw_rob.F.assign(Fr, Fr + 3 * (size_t)nFr);
  w->has_env = w->has_rob = true;
  return 0;
}
void orc_enforce_robot_center(void* h, const double* c) {
  World* w = (World*)h;
  for (int k = 0; k < 3; k++) w->center[k] = (float)c[k];
  w->center_enforced = true;
}
void orc_set_perturbation(void* h, const double* q) { memcpy(((World*)h)->perturb, q, 56); }
void orc_finalize(void* h) { ((World*)h)->finalize(); }
double orc_scene_scale(void* h) { return (double)((World*)h)->scene_scale; }
void orc_get_center(void* h, double* c) { for (int k = 0; k < 3; k++) c[k] = ((World*)h)->center[k]; }
int orc_num_tris(void* h, int which) { World* w = (World*)h; return which ? w->rob.ntri() : w->env.ntri(); }
int orc_num_verts(void* h, int which) { World* w = (World*)h; return (int)(which ? w->rob.V.size() : w->env.V.size()); }
void orc_get_unit_vertices(void* h, int which, double* out) {
  World* w = (World*)h;
  const auto& V = which ? w->rob.V : w->env.V;
  for (size_t i = 0; i < V.size(); i++) { out[3 * i] = V[i].x; out[3 * i + 1] = V[i].y; out[3 * i + 2] = V[i].z; }
}
void orc_get_faces(void* h, int which, int* out) {
  World* w = (World*)h;
  const auto& F = which ? w->rob.F : w->env.F;
  memcpy(out, F.data(), F.size() * sizeof(int));
}
void orc_set_moller(void* fn) { orc::g_moller = (orc::moller_fn)fn; }
int orc_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// mode 0: BVH DFS (FCL order); mode 1: brute-force all pairs.  threads<=0: all cores.
// counters (optional): per-sample N_bv, N_leaf (int32 each, [N][2]).
void orc_is_valid_states(void* h, const double* q, long long N, uint8_t* out, int mode, int threads,
                         int* counters) {
  World* w = (World*)h;
#ifdef _OPENMP
  int nt = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 64) num_threads(nt)
#endif
  for (long long i = 0; i < N; i++) {
    if (mode == 1) { out[i] = w->is_valid_brute(q + 7 * i); continue; }
    orc::Counters c;
    out[i] = w->is_valid(q + 7 * i, &c);
    if (counters) { counters[2 * i] = (int)c.bv; counters[2 * i + 1] = (int)c.leaf; }
  }
}
// vd: 1 value (vd_stride 0) or N values (vd_stride 1). counters: [N][3] = steps, bv, leaf (int64).
void orc_transit_states(void* h, const double* from, const double* to, long long N, const double* vd,
                        int vd_stride, double* last_free, double* first_hit, uint8_t* is_free,
                        double* free_tau, double* touch_tau, long long* counters, int threads) {
  World* w = (World*)h;
#ifdef _OPENMP
  int nt = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nt)
#endif
  for (long long i = 0; i < N; i++) {
    orc::Counters c; long long steps = 0;
    w->transit(from + 7 * i, to + 7 * i, vd[vd_stride ? i : 0], last_free + 7 * i, first_hit + 7 * i,
               is_free + i, free_tau + i, touch_tau + i, &steps, &c);
    if (counters) { counters[3 * i] = steps; counters[3 * i + 1] = c.bv; counters[3 * i + 2] = c.leaf; }
  }
}
void orc_is_valid_transitions(void* h, const double* from, const double* to, long long N, double ivd,
                              uint8_t* out, long long* counters, int threads) {
  World* w = (World*)h;
#ifdef _OPENMP
  int nt = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nt)
#endif
  for (long long i = 0; i < N; i++) {
    orc::Counters c; long long steps = 0;
    out[i] = w->is_valid_transition(from + 7 * i, to + 7 * i, ivd, &steps, &c);
    if (counters) { counters[3 * i] = steps; counters[3 * i + 1] = c.bv; counters[3 * i + 2] = c.leaf; }
  }
}
// Contacts for N states. offsets[N+1]; rows beyond cap are dropped (returns total rows found).
long long orc_intersecting_segments(void* h, const double* q, long long N, long long cap, long long* offsets,
                                    double* segs, int* faces, uint8_t* coplanar, long long* counters) {
  World* w = (World*)h;
  long long total = 0;
  offsets[0] = 0;
  for (long long i = 0; i < N; i++) {
    std::vector<int> pairs; std::vector<double> sg; std::vector<uint8_t> cp;
    orc::Counters c;
    int rows = w->intersecting_segments(q + 7 * i, pairs, sg, cp, &c);
    for (int r = 0; r < rows; r++) {
      long long o = total + r;
      if (o >= cap) break;
      faces[2 * o] = pairs[2 * r]; faces[2 * o + 1] = pairs[2 * r + 1];
      memcpy(segs + 6 * o, sg.data() + 6 * r, 48);
      coplanar[o] = cp[r];
    }
    total += rows;
    offsets[i + 1] = std::min(total, cap);
    if (counters) { counters[2 * i] = c.bv; counters[2 * i + 1] = c.leaf; }
  }
  return total;
}
void orc_is_disentangled_states(void* h, const double* q, long long N, uint8_t* out) {
  World* w = (World*)h;
  for (long long i = 0; i < N; i++) out[i] = w->is_disentangled(q + 7 * i);
}
// state math
void orc_interpolate(const double* a, const double* b, double tau, double* out) { orc::interpolate(a, b, tau, out); }
double orc_distance(const double* a, const double* b) { return orc::distance(a, b); }
void orc_apply(const double* from, const double* tr, const double* aa, double* out) { orc::apply(from, tr, aa, out); }
// raw triangle predicates (known-answer tests): 6 points x 3 doubles, Q already in P's frame
int orc_sat17(const double* t) {
  orc::V3 v[6];
  for (int i = 0; i < 6; i++) v[i] = {t[3 * i], t[3 * i + 1], t[3 * i + 2]};
  return orc::sat17(v[0], v[1], v[2], v[3], v[4], v[5]);
}
int orc_contact_multiplicity(const double* t) {
  orc::V3 v[6];
  for (int i = 0; i < 6; i++) v[i] = {t[3 * i], t[3 * i + 1], t[3 * i + 2]};
  return (int)orc::contact_multiplicity(v[0], v[1], v[2], v[3], v[4], v[5]);
}
// UnitWorld::intersectingToSurface (unit_world.cc:1092-1176): TriTriCopIsect (lib/tritri/tritri_cop.hh:115-151)
// over every (target face i, mesh face j) pair with the REFERENCE's Moeller routine (oracle/_ref), kept when
// it intersects and is coplanar; rows in Eigen's column-major SparseMatrix order (j outer, i inner); libigl
// barycentric_coordinates (third-party/libigl/include/igl/barycentric_coordinates.cpp) per vertex of face j.
// TV = world-space target vertices.  Returns the number of rows found; writes at most `cap`.
long long orc_surface_map(const double* TV, const int* TF, long long nTF, const double* V, const int* F,
                          long long nF, long long cap, int* outF, double* outL) {
  if (!orc::g_moller) return -1;
  long long rows = 0;
  for (long long j = 0; j < nF; j++) {
    const double* u[3] = {V + 3 * F[3 * j], V + 3 * F[3 * j + 1], V + 3 * F[3 * j + 2]};
    for (long long i = 0; i < nTF; i++) {
      const double* v[3] = {TV + 3 * TF[3 * i], TV + 3 * TF[3 * i + 1], TV + 3 * TF[3 * i + 2]};
      int cop = 0;
      double p0[3], p1[3];
      int isect = orc::g_moller(v[0], v[1], v[2], u[0], u[1], u[2], &cop, p0, p1);
      if (!(isect && cop)) continue;
      if (rows < cap) {
        for (int k = 0; k < 3; k++) outF[3 * rows + k] = TF[3 * i + k];
        for (int p = 0; p < 3; p++) {
          double v0[3], v1[3], v2[3];
          for (int k = 0; k < 3; k++) { v0[k] = v[1][k] - v[0][k]; v1[k] = v[2][k] - v[0][k]; v2[k] = u[p][k] - v[0][k]; }
          double d00 = v0[0] * v0[0] + v0[1] * v0[1] + v0[2] * v0[2], d01 = v0[0] * v1[0] + v0[1] * v1[1] + v0[2] * v1[2];
          double d11 = v1[0] * v1[0] + v1[1] * v1[1] + v1[2] * v1[2], d20 = v2[0] * v0[0] + v2[1] * v0[1] + v2[2] * v0[2];
          double d21 = v2[0] * v1[0] + v2[1] * v1[1] + v2[2] * v1[2];
          double den = d00 * d11 - d01 * d01;
          double* L = outL + 9 * rows + 3 * p;
          L[1] = (d11 * d20 - d01 * d21) / den;
          L[2] = (d00 * d21 - d01 * d20) / den;
          L[0] = 1.0 - (L[1] + L[2]);
        }
      }
      rows++;
    }
  }
  return rows;
}

// OmplDriver::validateMotionPairs (lib/pyse3ompl/ompldriver.cc:592-621) -> DiscreteMotionValidator::checkMotion
// (third-party/ompl.app/ompl/src/ompl/base/src/DiscreteMotionValidator.cpp:96-150), literally: s2 first, then
// the interior points in bisection order from a FIFO queue.  OMPL pieces restated from the vendored sources:
// validSegmentCount (StateSpace.cpp, CompoundStateSpace: maximum over sub-spaces), SO3StateSpace arcLength /
// interpolate (spaces/src/SO3StateSpace.cpp:252-316), RealVectorStateSpace::interpolate.  Each OMPL-frame state
// is mapped to the unit frame as UnitWorld::translateOMPLStateToUnitState (unit_world.cc:689-711, identity
// perturbation): t = scale*t - Rot(q)*dc.  PARITY UNPINNED: OMPL itself cannot be built here (boost absent).
static void ompl_to_unit(const double* o, double scale, const double* dc, double* u) {
  double w = o[6], x = o[3], y = o[4], z = o[5];
  double n2 = x * x + y * y + z * z + w * w;
  if (n2 > 0) { double n = std::sqrt(n2); w /= n; x /= n; y /= n; z /= n; }
  double uvx = 2.0 * (y * dc[2] - z * dc[1]), uvy = 2.0 * (z * dc[0] - x * dc[2]), uvz = 2.0 * (x * dc[1] - y * dc[0]);
  double rx = dc[0] + w * uvx + (y * uvz - z * uvy), ry = dc[1] + w * uvy + (z * uvx - x * uvz),
         rz = dc[2] + w * uvz + (x * uvy - y * uvx);
  u[0] = scale * o[0] - rx; u[1] = scale * o[1] - ry; u[2] = scale * o[2] - rz;
  u[3] = w; u[4] = x; u[5] = y; u[6] = z;
}
// SpaceInformation::satisfiesBounds for SE(3): RealVectorStateSpace (spaces/src/RealVectorStateSpace.cpp:200-208)
// and SO3StateSpace (spaces/src/SO3StateSpace.cpp:175-179, 211-214).  bounds = lo[3] hi[3]; null = no bounds test.
static bool ompl_in_bounds(const double* bounds, const double* o) {
  if (!bounds) return true;
  const double eps = std::numeric_limits<double>::epsilon();
  for (int k = 0; k < 3; k++)
    if (o[k] - eps > bounds[3 + k] || o[k] + eps < bounds[k]) return false;
  double n2 = o[3] * o[3] + o[4] * o[4] + o[5] * o[5] + o[6] * o[6];
  double nrm = (std::fabs(n2 - 1.0) > eps) ? std::sqrt(n2) : 1.0;
  return std::fabs(nrm - 1.0) < 1e-9;
}
// FCLStateValidityChecker::isValid (omplapp/geometry/detail/FCLStateValidityChecker.h:103)
static bool ompl_state_valid(World* w, const double* bounds, const double* o, double scale, const double* dc) {
  if (!ompl_in_bounds(bounds, o)) return false;
  double u[7];
  ompl_to_unit(o, scale, dc, u);
  return w->is_valid(u);
}
// OmplDriver::validateStates (lib/pyse3ompl/ompldriver.cc:566-589): si->isValid per row
void orc_ompl_validate_states(void* h, const double* q, long long N, double scale, const double* dc,
                              const double* bounds, uint8_t* out) {
  World* w = (World*)h;
  for (long long i = 0; i < N; i++) out[i] = ompl_state_valid(w, bounds, q + 7 * i, scale, dc) ? 1 : 0;
}
void orc_ompl_check_motions(void* h, const double* q0, const double* q1, long long N, double scale, const double* dc,
                            const double* bounds, double seg_trans, double seg_rot, uint8_t* out, long long* nd_out,
                            int threads) {
  World* w = (World*)h;
#ifdef _OPENMP
  int nt = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nt)
#endif
  for (long long i = 0; i < N; i++) {
    const double* a = q0 + 7 * i; const double* b = q1 + 7 * i;
    double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    double dtr = std::sqrt(dx * dx + dy * dy + dz * dz);
    double dq = std::fabs(a[3] * b[3] + a[4] * b[4] + a[5] * b[5] + a[6] * b[6]);
    double theta = dq > 1.0 - 1e-9 ? 0.0 : std::acos(dq);
    long long n1 = (long long)std::ceil(dtr / seg_trans), n2 = (long long)std::ceil(theta / seg_rot);
    long long nd = n1 > n2 ? n1 : n2;
    if (nd_out) nd_out[i] = nd;
    bool result = ompl_state_valid(w, bounds, b, scale, dc);
    if (result && nd >= 2) {
      std::vector<std::pair<long long, long long>> pos;
      size_t head = 0;
      pos.emplace_back(1, nd - 1);
      while (head < pos.size()) {
        std::pair<long long, long long> x = pos[head];
        long long mid = (x.first + x.second) / 2;
        double t = (double)mid / (double)nd, o[7];
        for (int k = 0; k < 3; k++) o[k] = a[k] + (b[k] - a[k]) * t;
        if (theta > 2.220446049250313e-16) {
          double d = 1.0 / std::sin(theta), s0 = std::sin((1.0 - t) * theta), s1 = std::sin(t * theta);
          double dot = a[3] * b[3] + a[4] * b[4] + a[5] * b[5] + a[6] * b[6];
          if (dot < 0) s1 = -s1;
          for (int k = 0; k < 4; k++) o[3 + k] = (a[3 + k] * s0 + b[3 + k] * s1) * d;
        } else {
          for (int k = 0; k < 4; k++) o[3 + k] = a[3 + k];
        }
        if (!ompl_state_valid(w, bounds, o, scale, dc)) { result = false; break; }
        head++;
        if (x.first < mid) pos.emplace_back(x.first, mid - 1);
        if (x.second > mid) pos.emplace_back(mid + 1, x.second);
      }
    }
    out[i] = result ? 1 : 0;
  }
}

}  // extern "C"
