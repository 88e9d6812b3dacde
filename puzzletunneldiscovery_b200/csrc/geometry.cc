// geometry.cc -- OBJ reader, unit-world transform and the GPU-oriented OBB-BVH builder.
// See geometry.h for the reference functions this replaces.
#include "geometry.h"
#include "qnode.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <fstream>
#include <future>
#include <limits>
#include <queue>
#include <sstream>

namespace osr {

bool load_obj(const std::string& path, TriMesh& out, std::string& err) {
  std::ifstream in(path.c_str());
  if (!in.good()) { err = "cannot open " + path; return false; }
  std::vector<float> pos;
  std::vector<int> tri;
  std::string line;
  while (std::getline(in, line)) {
    if (line.size() < 2 || (line[1] != ' ' && line[1] != '\t')) continue;
    if (line[0] == 'v') {
      const char* p = line.c_str() + 1;
      char* e = nullptr;
      for (int k = 0; k < 3; k++) { pos.push_back(strtof(p, &e)); p = e; }
    } else if (line[0] == 'f') {
      std::istringstream ss(line.substr(1));
      std::string tok;
      std::vector<int> poly;
      int nv = (int)(pos.size() / 3);
      while (ss >> tok) {
        int v = atoi(tok.c_str());
        int idx = v > 0 ? v - 1 : nv + v;
        if (idx < 0 || idx >= nv) { err = "face index out of range in " + path; return false; }
        poly.push_back(idx);
      }
      for (size_t k = 2; k < poly.size(); k++) {
        tri.push_back(poly[0]); tri.push_back(poly[k - 1]); tri.push_back(poly[k]);
      }
    }
  }
  if (tri.empty()) { err = "no faces in " + path; return false; }
  // drop unreferenced vertices (a mesh importer only keeps vertices used by faces)
  std::vector<int> remap(pos.size() / 3, -1);
  std::vector<char> used(pos.size() / 3, 0);
  for (int t : tri) used[t] = 1;
  out.V.clear(); out.F.clear();
  for (size_t i = 0; i < used.size(); i++)
    if (used[i]) {
      remap[i] = out.nv();
      out.V.insert(out.V.end(), pos.begin() + 3 * i, pos.begin() + 3 * i + 3);
    }
  for (int t : tri) out.F.push_back(remap[t]);
  return true;
}

float bbox_span(const TriMesh& m) {
  float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
  for (size_t i = 0; i < m.V.size(); i += 3)
    for (int k = 0; k < 3; k++) {
      hi[k] = std::max(m.V[i + k], hi[k]);
      lo[k] = std::min(m.V[i + k], lo[k]);
    }
  return std::max(std::max(hi[0] - lo[0], hi[1] - lo[1]), hi[2] - lo[2]);
}

void mean_of_vertices(const TriMesh& m, float c[3]) {
  float s[3] = {0, 0, 0};
  for (int i = 0; i < m.nv(); i++)
    for (int k = 0; k < 3; k++) s[k] += m.V[3 * i + k];
  for (int k = 0; k < 3; k++) c[k] = s[k] / (float)m.nv();
}

void make_unit_model(const TriMesh& m, float scale, const float* centre, UnitModel& out) {
  double s = (double)scale;
  double tr[3] = {0, 0, 0};
  if (centre)
    for (int k = 0; k < 3; k++) tr[k] = (double)(scale * (-centre[k]));  // float product, then widened
  out.V.resize(m.V.size());
  for (size_t i = 0; i < m.V.size(); i++) out.V[i] = s * (double)m.V[i] + tr[i % 3];
  out.F = m.F;
  out.tris.resize((size_t)9 * out.nf());
  for (int f = 0; f < out.nf(); f++)
    for (int k = 0; k < 3; k++)
      for (int d = 0; d < 3; d++) out.tris[(size_t)9 * f + 3 * k + d] = out.V[3 * (size_t)out.F[3 * f + k] + d];
  for (int d = 0; d < 3; d++) {
    out.aabb_min[d] = std::numeric_limits<double>::max();
    out.aabb_max[d] = -std::numeric_limits<double>::max();
  }
  // FCL computeLocalAABB runs over all model vertices
  for (size_t i = 0; i < out.V.size(); i++) {
    out.aabb_min[i % 3] = std::min(out.aabb_min[i % 3], out.V[i]);
    out.aabb_max[i % 3] = std::max(out.aabb_max[i % 3], out.V[i]);
  }
}

// ------------------------------------------------------------------------------------------
// BVH builder
// ------------------------------------------------------------------------------------------
namespace {

struct Vec3 { double x, y, z; };
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline double dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

// Jacobi rotations on a symmetric 3x3; columns of v become eigenvectors.
void jacobi3(double a[3][3], double v[3][3]) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) v[i][j] = (i == j);
  for (int sweep = 0; sweep < 64; sweep++) {
    double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
    double diag = std::fabs(a[0][0]) + std::fabs(a[1][1]) + std::fabs(a[2][2]);
    if (off <= 1e-18 * diag || off == 0.0) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        if (a[p][q] == 0.0) continue;
        double th = 0.5 * (a[q][q] - a[p][p]) / a[p][q];
        double t = (th >= 0 ? 1.0 : -1.0) / (std::fabs(th) + std::sqrt(1.0 + th * th));
        double c = 1.0 / std::sqrt(1.0 + t * t), s = t * c;
        for (int k = 0; k < 3; k++) { double x = a[k][p], y = a[k][q]; a[k][p] = c * x - s * y; a[k][q] = s * x + c * y; }
        for (int k = 0; k < 3; k++) { double x = a[p][k], y = a[q][k]; a[p][k] = c * x - s * y; a[q][k] = s * x + c * y; }
        for (int k = 0; k < 3; k++) { double x = v[k][p], y = v[k][q]; v[k][p] = c * x - s * y; v[k][q] = s * x + c * y; }
      }
  }
}

struct BuildNode {
  float axis[9];   // axis k = axis[3k..3k+2]  (fp32-rounded, what the GPU will use)
  float centre[3];
  float ext[3];
  int left = -1;   // child build-node ids, or -1
  int right = -1;
  int tri = -1;
  uint16_t qh[4];  // QNode: orientation quaternion (fp16 bits) ...
  uint16_t eh[3];  // ... and extents fitted to the exact rotation of that quaternion, rounded up to fp16
};

// Builder options (development switches; the defaults are the measured best, see DESIGN.md "BVH quality").
//   OSR_BVH_REFINE  0: covariance (PCA) frame as FCL's BVFitter   1: + local search minimising the box AREA   2: ... VOLUME
//   OSR_BVH_SPLIT   0: mean of centroids on the longest axis (FCL SPLIT_METHOD_MEAN)   1: surface-area sweep in the box frame
static std::atomic<long long> g_t_refine{0}, g_t_sweep{0}, g_t_fit{0};
struct ScopedNs {
  std::atomic<long long>& acc;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  explicit ScopedNs(std::atomic<long long>& a) : acc(a) {}
  ~ScopedNs() { acc += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count(); }
};
static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

struct Builder {
  const UnitModel& m;
  std::vector<BuildNode> nodes;
  std::vector<int> prims;
  std::vector<Vec3> cent;
  int depth = 0;
  int max_depth = 64;
  int par_levels = 5;   // levels whose left subtree is built by another thread (<= 2^5 threads in flight)
  const int opt_refine = env_int("OSR_BVH_REFINE", 1);
  const int opt_split = env_int("OSR_BVH_SPLIT", 1);
  explicit Builder(const UnitModel& mm) : m(mm) {}

  // half extents of the points `pts` in the orthonormal frame ax -> cost (area or volume of the box)
  static double frame_cost(const std::vector<Vec3>& pts, const Vec3 ax[3], int kind) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (const Vec3& p : pts)
      for (int a = 0; a < 3; a++) {
        double pr = dot(ax[a], p);
        lo[a] = std::min(lo[a], pr); hi[a] = std::max(hi[a], pr);
      }
    double e0 = hi[0] - lo[0], e1 = hi[1] - lo[1], e2 = hi[2] - lo[2];
    return kind == 2 ? e0 * e1 * e2 : e0 * e1 + e1 * e2 + e0 * e2;
  }
  static void rotate_frame(const Vec3 in[3], int about, double ang, Vec3 out[3]) {
    const int i = (about + 1) % 3, j = (about + 2) % 3;
    const double c = std::cos(ang), s = std::sin(ang);
    out[about] = in[about];
    out[i] = {c * in[i].x + s * in[j].x, c * in[i].y + s * in[j].y, c * in[i].z + s * in[j].z};
    out[j] = {c * in[j].x - s * in[i].x, c * in[j].y - s * in[i].y, c * in[j].z - s * in[i].z};
  }
  // Local search over orientations from the covariance frame: rotate about each box axis by +-step, keep what
  // shrinks the box, halve the step.  Large nodes are searched on their extreme points along 96 fixed directions
  // (the final extents are always taken over every vertex, so containment never depends on the search).
  void refine_frame(int first, int n, Vec3 ax[3]) const {
    ScopedNs tm(g_t_refine);
    std::vector<Vec3> pts;
    pts.reserve((size_t)3 * n);
    for (int i = 0; i < n; i++)
      for (int k = 0; k < 3; k++) pts.push_back(vert(prims[first + i], k));
    if (pts.size() > 768) {
      std::vector<Vec3> ex;
      for (int d = 0; d < 96; d++) {
        // Fibonacci directions on the half sphere (both extremes are taken per direction)
        double z = (d + 0.5) / 96.0, r = std::sqrt(std::max(0.0, 1.0 - z * z)), ph = 2.399963229728653 * d;
        Vec3 dir = {r * std::cos(ph), r * std::sin(ph), z};
        size_t imin = 0, imax = 0;
        double vmin = 1e300, vmax = -1e300;
        for (size_t i = 0; i < pts.size(); i++) {
          double pr = dot(dir, pts[i]);
          if (pr < vmin) { vmin = pr; imin = i; }
          if (pr > vmax) { vmax = pr; imax = i; }
        }
        ex.push_back(pts[imin]); ex.push_back(pts[imax]);
      }
      pts.swap(ex);
    }
    double best = frame_cost(pts, ax, opt_refine);
    for (double step = 0.4; step > 0.01; step *= 0.5) {
      bool improved = true;
      for (int it = 0; improved && it < 12; it++) {
        improved = false;
        for (int about = 0; about < 3; about++)
          for (int sg = -1; sg <= 1; sg += 2) {
            Vec3 cand[3];
            rotate_frame(ax, about, sg * step, cand);
            double c = frame_cost(pts, cand, opt_refine);
            if (c < best * (1.0 - 1e-9)) { best = c; ax[0] = cand[0]; ax[1] = cand[1]; ax[2] = cand[2]; improved = true; }
          }
      }
    }
    // re-orthonormalise (rounding of the accumulated rotations)
    double l0 = std::sqrt(dot(ax[0], ax[0]));
    ax[0] = {ax[0].x / l0, ax[0].y / l0, ax[0].z / l0};
    double d01 = dot(ax[0], ax[1]);
    ax[1] = {ax[1].x - d01 * ax[0].x, ax[1].y - d01 * ax[0].y, ax[1].z - d01 * ax[0].z};
    double l1 = std::sqrt(dot(ax[1], ax[1]));
    ax[1] = {ax[1].x / l1, ax[1].y / l1, ax[1].z / l1};
    ax[2] = cross(ax[0], ax[1]);
  }

  Vec3 vert(int t, int k) const {
    const double* p = &m.tris[(size_t)9 * t + 3 * k];
    return {p[0], p[1], p[2]};
  }

  // PCA box of the triangle vertices; axes rounded to fp32 FIRST, then centre/extents are
  // derived from the exact (fp64) projections on the rounded axes and rounded outwards, so the
  // fp32 box is a superset of the fp64 triangles.
  void fit(int first, int n, BuildNode& nd) {
    ScopedNs tm(g_t_fit);
    double S1[3] = {0, 0, 0}, S2[3][3] = {{0}};
    for (int i = 0; i < n; i++)
      for (int k = 0; k < 3; k++) {
        Vec3 p = vert(prims[first + i], k);
        double pp[3] = {p.x, p.y, p.z};
        for (int a = 0; a < 3; a++) {
          S1[a] += pp[a];
          for (int b = 0; b < 3; b++) S2[a][b] += pp[a] * pp[b];
        }
      }
    double C[3][3], E[3][3];
    double np = 3.0 * n;
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) C[a][b] = S2[a][b] - S1[a] * S1[b] / np;
    jacobi3(C, E);
    int order[3] = {0, 1, 2};
    std::sort(order, order + 3, [&](int a, int b) { return C[a][a] > C[b][b]; });
    Vec3 a0 = {E[0][order[0]], E[1][order[0]], E[2][order[0]]};
    Vec3 a1 = {E[0][order[1]], E[1][order[1]], E[2][order[1]]};
    if (n == 1) {  // a single triangle: longest edge / in-plane normal / face normal
      Vec3 p0 = vert(prims[first], 0), p1 = vert(prims[first], 1), p2 = vert(prims[first], 2);
      Vec3 e[3] = {p1 - p0, p2 - p1, p0 - p2};
      int lg = 0;
      for (int k = 1; k < 3; k++) if (dot(e[k], e[k]) > dot(e[lg], e[lg])) lg = k;
      Vec3 nrm = cross(e[0], e[1]);
      if (dot(nrm, nrm) > 0 && dot(e[lg], e[lg]) > 0) { a0 = e[lg]; a1 = cross(nrm, a0); }
    }
    double l0 = std::sqrt(dot(a0, a0));
    if (!(l0 > 0)) { a0 = {1, 0, 0}; l0 = 1; }
    a0 = {a0.x / l0, a0.y / l0, a0.z / l0};
    double d01 = dot(a0, a1);
    a1 = {a1.x - d01 * a0.x, a1.y - d01 * a0.y, a1.z - d01 * a0.z};
    double l1 = std::sqrt(dot(a1, a1));
    if (!(l1 > 1e-12)) {
      Vec3 h = std::fabs(a0.x) < 0.9 ? Vec3{1, 0, 0} : Vec3{0, 1, 0};
      a1 = cross(a0, h);
      l1 = std::sqrt(dot(a1, a1));
    }
    a1 = {a1.x / l1, a1.y / l1, a1.z / l1};
    Vec3 a2 = cross(a0, a1);
    Vec3 ax[3] = {a0, a1, a2};
    if (opt_refine > 0 && n > 1) refine_frame(first, n, ax);
    for (int k = 0; k < 3; k++) {
      nd.axis[3 * k] = (float)ax[k].x; nd.axis[3 * k + 1] = (float)ax[k].y; nd.axis[3 * k + 2] = (float)ax[k].z;
      ax[k] = {(double)nd.axis[3 * k], (double)nd.axis[3 * k + 1], (double)nd.axis[3 * k + 2]};
    }
    double lo[3], hi[3];
    for (int k = 0; k < 3; k++) { lo[k] = std::numeric_limits<double>::max(); hi[k] = -lo[k]; }
    for (int i = 0; i < n; i++)
      for (int k = 0; k < 3; k++) {
        Vec3 p = vert(prims[first + i], k);
        for (int a = 0; a < 3; a++) {
          double pr = dot(ax[a], p);
          lo[a] = std::min(lo[a], pr); hi[a] = std::max(hi[a], pr);
        }
      }
    // centre in world coordinates, rounded to fp32; extents recomputed about the rounded centre.
    // The rounded axes are orthonormal only to ~1e-7, so measure the extents as the largest
    // |axis . (p - centre)| directly (exact definition of the box the GPU tests against).
    double o[3] = {(lo[0] + hi[0]) / 2, (lo[1] + hi[1]) / 2, (lo[2] + hi[2]) / 2};
    Vec3 c = {ax[0].x * o[0] + ax[1].x * o[1] + ax[2].x * o[2], ax[0].y * o[0] + ax[1].y * o[1] + ax[2].y * o[2],
              ax[0].z * o[0] + ax[1].z * o[1] + ax[2].z * o[2]};
    nd.centre[0] = (float)c.x; nd.centre[1] = (float)c.y; nd.centre[2] = (float)c.z;
    Vec3 cf = {(double)nd.centre[0], (double)nd.centre[1], (double)nd.centre[2]};
    double ext[3] = {0, 0, 0};
    for (int i = 0; i < n; i++)
      for (int k = 0; k < 3; k++) {
        Vec3 d = vert(prims[first + i], k) - cf;
        for (int a = 0; a < 3; a++) ext[a] = std::max(ext[a], std::fabs(dot(ax[a], d)));
      }
    for (int a = 0; a < 3; a++) {
      // slack for the non-orthonormality of the rounded frame (<= ~3e-7 relative) + round up
      double e = ext[a] * (1.0 + 1e-6) + 1e-9;
      nd.ext[a] = std::nextafterf((float)e, std::numeric_limits<float>::infinity());
    }
  }

  // QNode: quaternion of the (near-orthonormal) frame above, rounded to fp16.  The box of the node is DEFINED
  // by the exact rotation of the normalised rounded quaternion; extents are re-fitted to it in fp64 about the
  // same fp32 centre.  The relative slack covers the kernel's first-order renormalisation and fp32 products
  // (frame error < 2e-6 rad); the absolute one the same error times the lever arm to the centre.
  void fit_q(int first, int n, BuildNode& nd) {
    double m[3][3];  // columns = axes
    for (int k = 0; k < 3; k++) { m[0][k] = nd.axis[3 * k]; m[1][k] = nd.axis[3 * k + 1]; m[2][k] = nd.axis[3 * k + 2]; }
    double q[4];
    double tr = m[0][0] + m[1][1] + m[2][2];
    if (tr > 0) {
      double s = std::sqrt(tr + 1.0) * 2;
      q[0] = 0.25 * s; q[1] = (m[2][1] - m[1][2]) / s; q[2] = (m[0][2] - m[2][0]) / s; q[3] = (m[1][0] - m[0][1]) / s;
    } else if (m[0][0] > m[1][1] && m[0][0] > m[2][2]) {
      double s = std::sqrt(1.0 + m[0][0] - m[1][1] - m[2][2]) * 2;
      q[0] = (m[2][1] - m[1][2]) / s; q[1] = 0.25 * s; q[2] = (m[0][1] + m[1][0]) / s; q[3] = (m[0][2] + m[2][0]) / s;
    } else if (m[1][1] > m[2][2]) {
      double s = std::sqrt(1.0 + m[1][1] - m[0][0] - m[2][2]) * 2;
      q[0] = (m[0][2] - m[2][0]) / s; q[1] = (m[0][1] + m[1][0]) / s; q[2] = 0.25 * s; q[3] = (m[1][2] + m[2][1]) / s;
    } else {
      double s = std::sqrt(1.0 + m[2][2] - m[0][0] - m[1][1]) * 2;
      q[0] = (m[1][0] - m[0][1]) / s; q[1] = (m[0][2] + m[2][0]) / s; q[2] = (m[1][2] + m[2][1]) / s; q[3] = 0.25 * s;
    }
    double qn = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    double qd[4];
    for (int k = 0; k < 4; k++) {
      nd.qh[k] = float_to_half_bits_rn((float)(q[k] / qn));
      qd[k] = (double)half_bits_to_float(nd.qh[k]);
    }
    qn = std::sqrt(qd[0] * qd[0] + qd[1] * qd[1] + qd[2] * qd[2] + qd[3] * qd[3]);
    double R[9];
    quat_to_rot(qd[0] / qn, qd[1] / qn, qd[2] / qn, qd[3] / qn, R);  // columns = exact box axes
    Vec3 cf = {(double)nd.centre[0], (double)nd.centre[1], (double)nd.centre[2]};
    double ext[3] = {0, 0, 0}, lever = 0;
    for (int i = 0; i < n; i++)
      for (int k = 0; k < 3; k++) {
        Vec3 d = vert(prims[first + i], k) - cf;
        lever = std::max(lever, std::sqrt(dot(d, d)));
        for (int a = 0; a < 3; a++)
          ext[a] = std::max(ext[a], std::fabs(R[a] * d.x + R[3 + a] * d.y + R[6 + a] * d.z));
      }
    for (int a = 0; a < 3; a++) nd.eh[a] = float_to_half_bits_up(ext[a] * (1.0 + 1e-5) + 4e-6 * lever + 1e-9);
  }

  // Surface-area sweep in the node's own box frame: triangles sorted by centroid along each box axis, cost of a cut
  // = area(box of the left part) * count + area(box of the right part) * count with both boxes axis-aligned in that
  // frame; the cut must leave both sides small enough for the depth budget.  Reorders prims[first, first+n).
  int sweep_split(int first, int n, const BuildNode& nd, int lvl) {
    ScopedNs tm(g_t_sweep);
    Vec3 ax[3];
    for (int k = 0; k < 3; k++) ax[k] = {(double)nd.axis[3 * k], (double)nd.axis[3 * k + 1], (double)nd.axis[3 * k + 2]};
    const int room = max_depth - lvl - 1;                       // levels left below the children
    const long long cap = room >= 30 ? (1LL << 30) : (1LL << std::max(0, room));
    const int kmin = (int)std::max<long long>(1, n - cap), kmax = (int)std::min<long long>(n - 1, cap);
    if (kmin > kmax) return -1;
    struct Box { double lo[3], hi[3]; };
    std::vector<Box> tb(n);
    std::vector<double> key(n);
    std::vector<int> idx(n), order[3];
    for (int i = 0; i < n; i++) {
      Box& b = tb[i];
      for (int a = 0; a < 3; a++) { b.lo[a] = 1e300; b.hi[a] = -1e300; }
      for (int k = 0; k < 3; k++) {
        Vec3 p = vert(prims[first + i], k);
        for (int a = 0; a < 3; a++) {
          double pr = dot(ax[a], p);
          b.lo[a] = std::min(b.lo[a], pr); b.hi[a] = std::max(b.hi[a], pr);
        }
      }
    }
    auto area = [](const Box& b) {
      double e0 = b.hi[0] - b.lo[0], e1 = b.hi[1] - b.lo[1], e2 = b.hi[2] - b.lo[2];
      return e0 * e1 + e1 * e2 + e0 * e2;
    };
    auto grow = [](Box& b, const Box& o) {
      for (int a = 0; a < 3; a++) { b.lo[a] = std::min(b.lo[a], o.lo[a]); b.hi[a] = std::max(b.hi[a], o.hi[a]); }
    };
    double best = 1e300;
    int best_k = -1, best_axis = -1;
    std::vector<double> right(n + 1);
    for (int a = 0; a < 3; a++) {
      for (int i = 0; i < n; i++) { idx[i] = i; key[i] = dot(ax[a], cent[prims[first + i]]); }
      std::sort(idx.begin(), idx.end(), [&](int x, int y) { return key[x] < key[y] || (key[x] == key[y] && x < y); });
      Box acc;
      for (int d = 0; d < 3; d++) { acc.lo[d] = 1e300; acc.hi[d] = -1e300; }
      right[n] = 0;
      for (int i = n - 1; i >= 1; i--) { grow(acc, tb[idx[i]]); right[i] = area(acc); }
      for (int d = 0; d < 3; d++) { acc.lo[d] = 1e300; acc.hi[d] = -1e300; }
      for (int k = 1; k < n; k++) {
        grow(acc, tb[idx[k - 1]]);
        if (k < kmin || k > kmax) continue;
        double c = area(acc) * k + right[k] * (n - k);
        if (c < best) { best = c; best_k = k; best_axis = a; }
      }
      order[a] = idx;
    }
    if (best_k < 0) return -1;
    const std::vector<int>& best_order = order[best_axis];
    std::vector<int> tmp(n);
    for (int i = 0; i < n; i++) tmp[i] = prims[first + best_order[i]];
    std::copy(tmp.begin(), tmp.end(), prims.begin() + first);
    return best_k;
  }

  // Builds the subtree over prims[first, first+n) into `out` (child ids are indices into `out`); returns its root id.
  // Subtrees operate on disjoint ranges of `prims`, so the first levels hand their left child to another thread.
  int build(std::vector<BuildNode>& out, int& max_lvl, int first, int n, int lvl) {
    max_lvl = std::max(max_lvl, lvl);
    int id = (int)out.size();
    out.emplace_back();
    BuildNode nd;
    fit(first, n, nd);
    fit_q(first, n, nd);
    if (n == 1) { nd.tri = prims[first]; out[id] = nd; return id; }
    int c1 = -1;
    if (opt_split == 1 && n > 2) c1 = sweep_split(first, n, nd, lvl);
    if (c1 < 0) {
      // split along the box's longest axis at the mean centroid (FCL SPLIT_METHOD_MEAN); fall back to the median
      // when the mean split is badly unbalanced (keeps the depth, hence the traversal stack, small)
      int la = 0;
      for (int a = 1; a < 3; a++) if (nd.ext[a] > nd.ext[la]) la = a;
      Vec3 sv = {(double)nd.axis[3 * la], (double)nd.axis[3 * la + 1], (double)nd.axis[3 * la + 2]};
      double mean = 0;
      for (int i = 0; i < n; i++) mean += dot(sv, cent[prims[first + i]]);
      mean /= n;
      auto mid = std::partition(prims.begin() + first, prims.begin() + first + n,
                                [&](int t) { return dot(sv, cent[t]) <= mean; });
      c1 = (int)(mid - (prims.begin() + first));
      if (c1 < n / 8 + (n < 8 ? 1 : 0) || n - c1 < n / 8 + (n < 8 ? 1 : 0) || c1 == 0 || c1 == n) {
        c1 = n / 2;
        std::nth_element(prims.begin() + first, prims.begin() + first + c1, prims.begin() + first + n,
                         [&](int a, int b) { return dot(sv, cent[a]) < dot(sv, cent[b]); });
      }
    }
    if (lvl < par_levels && n >= 4096) {
      std::vector<BuildNode> lsub;
      int llvl = 0;
      std::future<int> fut = std::async(std::launch::async, [&, first, c1, lvl] { return build(lsub, llvl, first, c1, lvl + 1); });
      int r = build(out, max_lvl, first + c1, n - c1, lvl + 1);
      int lroot = fut.get();
      const int off = (int)out.size();
      for (BuildNode b : lsub) {
        if (b.left >= 0) { b.left += off; b.right += off; }
        out.push_back(b);
      }
      max_lvl = std::max(max_lvl, llvl);
      nd.left = lroot + off; nd.right = r;
    } else {
      int l = build(out, max_lvl, first, c1, lvl + 1);
      int r = build(out, max_lvl, first + c1, n - c1, lvl + 1);
      nd.left = l; nd.right = r;
    }
    out[id] = nd;
    return id;
  }
};

}  // namespace

void build_gpu_bvh(UnitModel& m) {
  Builder b(m);
  int n = m.nf();
  b.prims.resize(n);
  b.cent.resize(n);
  for (int t = 0; t < n; t++) {
    b.prims[t] = t;
    Vec3 p0 = b.vert(t, 0), p1 = b.vert(t, 1), p2 = b.vert(t, 2);
    b.cent[t] = {(p0.x + p1.x + p2.x) / 3, (p0.y + p1.y + p2.y) / 3, (p0.z + p1.z + p2.z) / 3};
  }
  b.nodes.reserve(2 * (size_t)n);
  {
    int lg = 0;
    while ((1LL << lg) < n) lg++;
    b.max_depth = lg + 3;   // the sweep split may unbalance the tree by at most this much (traversal stack depth)
  }
  auto t_start = std::chrono::steady_clock::now();
  b.build(b.nodes, b.depth, 0, n, 0);
  if (getenv("OSR_BVH_VERBOSE"))
    fprintf(stderr, "[bvh] %d triangles, depth %d: %.2f s wall; thread-seconds fit %.2f (of which refine %.2f), sweep split %.2f\n", n,
            b.depth, std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count(), g_t_fit.load() * 1e-9,
            g_t_refine.load() * 1e-9, g_t_sweep.load() * 1e-9);
  // breadth-first renumbering with sibling pairs adjacent
  int N = (int)b.nodes.size();
  std::vector<int> order;   // new index -> build id
  std::vector<int> newid(N, -1);
  order.reserve(N);
  order.push_back(0); newid[0] = 0;
  for (size_t h = 0; h < order.size(); h++) {
    const BuildNode& nd = b.nodes[order[h]];
    if (nd.left >= 0) {
      newid[nd.left] = (int)order.size(); order.push_back(nd.left);
      newid[nd.right] = (int)order.size(); order.push_back(nd.right);
    }
  }
  m.bvh.num_nodes = N;
  m.bvh.depth = b.depth;
  m.bvh.nodes.assign((size_t)16 * N, 0.f);
  m.bvh.qnodes.assign((size_t)8 * N, 0u);
  for (int i = 0; i < N; i++) {
    const BuildNode& nd = b.nodes[order[i]];
    float* o = &m.bvh.nodes[(size_t)16 * i];
    for (int k = 0; k < 3; k++) {
      o[4 * k] = nd.axis[3 * k]; o[4 * k + 1] = nd.axis[3 * k + 1]; o[4 * k + 2] = nd.axis[3 * k + 2];
      o[4 * k + 3] = nd.centre[k];
    }
    o[12] = nd.ext[0]; o[13] = nd.ext[1]; o[14] = nd.ext[2];
    int32_t link = nd.left >= 0 ? newid[nd.left] : -(nd.tri + 1);
    std::memcpy(&o[15], &link, 4);
    QNode qn;
    for (int k = 0; k < 3; k++) qn.c[k] = nd.centre[k];
    qn.link = link;
    qn.q01 = (uint32_t)nd.qh[0] | ((uint32_t)nd.qh[1] << 16);
    qn.q23 = (uint32_t)nd.qh[2] | ((uint32_t)nd.qh[3] << 16);
    qn.e01 = (uint32_t)nd.eh[0] | ((uint32_t)nd.eh[1] << 16);
    qn.e2x = (uint32_t)nd.eh[2];
    std::memcpy(&m.bvh.qnodes[(size_t)8 * i], &qn, 32);
  }
  double r2 = 0;
  for (size_t i = 0; i < m.V.size(); i += 3) r2 = std::max(r2, m.V[i] * m.V[i] + m.V[i + 1] * m.V[i + 1] + m.V[i + 2] * m.V[i + 2]);
  m.bvh.max_radius = std::nextafterf((float)std::sqrt(r2), std::numeric_limits<float>::infinity());
}

}  // namespace osr
