// trav_sim.cc -- CPU counting experiment for the GPU traversal design (development tool, not product code).
// Builds the engine's own BVHs (csrc/geometry.cc), replays seeded states and counts the box / leaf tests
// of candidate traversal orders and box predicates.  No collision labels are produced for anyone: the
// leaf predicate here is a plain fp64 SAT used only to know where a depth-first walk would stop.
//   g++ -O2 -std=c++17 -o /tmp/trav_sim scripts/trav_sim.cc puzzletunneldiscovery_b200/csrc/geometry.cc
//   /tmp/trav_sim env.obj rob.obj states.bin N
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../puzzletunneldiscovery_b200/csrc/geometry.h"
#include "../puzzletunneldiscovery_b200/csrc/se3.h"
#include "../puzzletunneldiscovery_b200/csrc/qnode.h"

using namespace osr;

struct Pose { float R[9], t[3]; double Rd[9], td[3]; float eta; BoxPose bp; };

static int g_axes = 15;

static const UnitModel *g_env, *g_rob;
static bool box_overlap(const float* a, const float* b, const Pose& P) {
  if (g_axes == 99) {  // QNode test
    size_t ia = (a - g_env->bvh.nodes.data()) / 16, ib = (b - g_rob->bvh.nodes.data()) / 16;
    QNode A, B; float s1, s2;
    memcpy(&A, &g_env->bvh.qnodes[8 * ia], 32); memcpy(&B, &g_rob->bvh.qnodes[8 * ib], 32);
    return obb_overlap_q(A, B, P.bp, s1, s2);
  }
  const float reps = 2e-5f;
  const float* r = P.R;
  float v[3][3];
  for (int j = 0; j < 3; j++)
    for (int i = 0; i < 3; i++) v[j][i] = r[3 * i] * b[4 * j] + r[3 * i + 1] * b[4 * j + 1] + r[3 * i + 2] * b[4 * j + 2];
  float c[3];
  for (int i = 0; i < 3; i++) c[i] = r[3 * i] * b[3] + r[3 * i + 1] * b[7] + r[3 * i + 2] * b[11] + P.t[i];
  float d[3] = {c[0] - a[3], c[1] - a[7], c[2] - a[11]};
  float T[3], B[3][3], F[3][3];
  for (int i = 0; i < 3; i++) {
    T[i] = a[4 * i] * d[0] + a[4 * i + 1] * d[1] + a[4 * i + 2] * d[2];
    for (int j = 0; j < 3; j++) {
      B[i][j] = a[4 * i] * v[j][0] + a[4 * i + 1] * v[j][1] + a[4 * i + 2] * v[j][2];
      F[i][j] = fabsf(B[i][j]) + reps;
    }
  }
  float ea[3] = {a[12] + P.eta, a[13] + P.eta, a[14] + P.eta}, eb[3] = {b[12] + P.eta, b[13] + P.eta, b[14] + P.eta};
  for (int i = 0; i < 3; i++)
    if (fabsf(T[i]) > ea[i] + F[i][0] * eb[0] + F[i][1] * eb[1] + F[i][2] * eb[2]) return false;
  for (int j = 0; j < 3; j++)
    if (fabsf(B[0][j] * T[0] + B[1][j] * T[1] + B[2][j] * T[2]) > eb[j] + F[0][j] * ea[0] + F[1][j] * ea[1] + F[2][j] * ea[2])
      return false;
  if (g_axes == 6) return true;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      float lhs = fabsf(T[i2] * B[i1][j] - T[i1] * B[i2][j]);
      float rhs = ea[i1] * F[i2][j] + ea[i2] * F[i1][j] + eb[j1] * F[i][j2] + eb[j2] * F[i][j1];
      if (lhs > rhs) return false;
    }
  return true;
}

// plain fp64 SAT-17 (touching counts as intersecting)
static bool tri_hit(const double* P, const double* Q, const Pose& po) {
  double p[3][3], q[3][3];
  for (int v = 0; v < 3; v++)
    for (int k = 0; k < 3; k++) {
      p[v][k] = P[3 * v + k] - P[k];
      q[v][k] = (po.Rd[3 * k] * Q[3 * v] + po.Rd[3 * k + 1] * Q[3 * v + 1] + po.Rd[3 * k + 2] * Q[3 * v + 2]) + po.td[k] - P[k];
    }
  double e[3][3], f[3][3];
  for (int k = 0; k < 3; k++) {
    e[0][k] = p[1][k] - p[0][k]; e[1][k] = p[2][k] - p[1][k]; e[2][k] = p[0][k] - p[2][k];
    f[0][k] = q[1][k] - q[0][k]; f[1][k] = q[2][k] - q[1][k]; f[2][k] = q[0][k] - q[2][k];
  }
  auto cross = [](const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
  };
  auto sep = [&](const double* ax) {
    double mnp = 1e300, mxp = -1e300, mnq = 1e300, mxq = -1e300;
    for (int v = 0; v < 3; v++) {
      double a = ax[0] * p[v][0] + ax[1] * p[v][1] + ax[2] * p[v][2];
      double b = ax[0] * q[v][0] + ax[1] * q[v][1] + ax[2] * q[v][2];
      mnp = std::min(mnp, a); mxp = std::max(mxp, a); mnq = std::min(mnq, b); mxq = std::max(mxq, b);
    }
    return mnp > mxq || mnq > mxp;
  };
  double n1[3], m1[3], ax[3];
  cross(e[0], e[1], n1); if (sep(n1)) return false;
  cross(f[0], f[1], m1); if (sep(m1)) return false;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { cross(e[i], f[j], ax); if (sep(ax)) return false; }
  for (int i = 0; i < 3; i++) { cross(e[i], n1, ax); if (sep(ax)) return false; }
  for (int i = 0; i < 3; i++) { cross(f[i], m1, ax); if (sep(ax)) return false; }
  return true;
}

struct Sim {
  const UnitModel *env, *rob;
  long long bv = 0, leaf = 0;
  bool hit = false;
  int mode = 0;  // 0: FCL order (test after descend); 1: test both children eagerly, nearer first
  const float* en(int i) const { return &env->bvh.nodes[16 * (size_t)i]; }
  const float* rn(int i) const { return &rob->bvh.nodes[16 * (size_t)i]; }
  static int link(const float* n) { int l; memcpy(&l, n + 15, 4); return l; }
  void rec(int n1, int n2, const Pose& P) {  // pair already known to overlap
    const float *a = en(n1), *b = rn(n2);
    int l1 = link(a), l2 = link(b);
    if (l1 < 0 && l2 < 0) {
      leaf++;
      if (tri_hit(&env->tris[9 * (size_t)(-(l1 + 1))], &rob->tris[9 * (size_t)(-(l2 + 1))], P)) hit = true;
      return;
    }
    float s1 = a[12] * a[12] + a[13] * a[13] + a[14] * a[14], s2 = b[12] * b[12] + b[13] * b[13] + b[14] * b[14];
    int c[2][2];
    if (l2 < 0 || (l1 >= 0 && s1 > s2)) { c[0][0] = l1; c[0][1] = n2; c[1][0] = l1 + 1; c[1][1] = n2; }
    else { c[0][0] = n1; c[0][1] = l2; c[1][0] = n1; c[1][1] = l2 + 1; }
    if (mode == 0) {
      for (int k = 0; k < 2 && !hit; k++) {
        bv++;
        if (box_overlap(en(c[k][0]), rn(c[k][1]), P)) rec(c[k][0], c[k][1], P);
      }
    } else {
      bool o[2];
      for (int k = 0; k < 2; k++) { bv++; o[k] = box_overlap(en(c[k][0]), rn(c[k][1]), P); }
      for (int k = 0; k < 2 && !hit; k++) if (o[k]) rec(c[k][0], c[k][1], P);
    }
  }
  bool run(const Pose& P) {
    hit = false;
    bv++;
    if (box_overlap(en(0), rn(0), P)) rec(0, 0, P);
    return hit;
  }
};

int main(int argc, char** argv) {
  if (argc < 5) { fprintf(stderr, "usage: trav_sim env.obj rob.obj states.bin N\n"); return 1; }
  TriMesh em, rm; std::string err;
  if (!load_obj(argv[1], em, err) || !load_obj(argv[2], rm, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
  int N = atoi(argv[4]);
  std::vector<double> q(7 * (size_t)N);
  FILE* f = fopen(argv[3], "rb");
  if (!f || fread(q.data(), 56, N, f) != (size_t)N) { fprintf(stderr, "states?\n"); return 1; }
  fclose(f);
  float scale = (float)(1.0 / std::max(bbox_span(em), bbox_span(rm)));
  float c[3]; mean_of_vertices(rm, c);
  UnitModel env, rob;
  make_unit_model(em, scale, nullptr, env); make_unit_model(rm, scale, c, rob);
  build_gpu_bvh(env); build_gpu_bvh(rob);
  printf("env %d tris depth %d | rob %d tris depth %d\n", env.nf(), env.bvh.depth, rob.nf(), rob.bvh.depth);
  std::vector<Pose> poses(N);
  for (int i = 0; i < N; i++) {
    Rigid r; state_to_rigid(&q[7 * (size_t)i], r);
    Pose& P = poses[i];
    for (int k = 0; k < 9; k++) { P.Rd[k] = r.R[k]; P.R[k] = (float)r.R[k]; }
    for (int k = 0; k < 3; k++) { P.td[k] = r.t[k]; P.t[k] = (float)r.t[k]; }
    P.eta = 2e-5f * (1.0f + fabsf(P.t[0]) + fabsf(P.t[1]) + fabsf(P.t[2]));
    double ident[4] = {1, 0, 0, 0};
    make_box_pose(r, ident, &q[7 * (size_t)i + 3], rob.bvh.max_radius, P.bp);
  }
  g_env = &env; g_rob = &rob;
  std::vector<char> ref_hit(N, 0);
  for (int axes : {15, 99, 6})
    for (int mode : {0, 1}) {
      g_axes = axes;
      Sim S; S.env = &env; S.rob = &rob; S.mode = mode;
      long long bvv = 0, bvi = 0, lfv = 0, lfi = 0, nv = 0, ni = 0;
      for (int i = 0; i < N; i++) {
        long long b0 = S.bv, l0 = S.leaf;
        bool h = S.run(poses[i]);
        if (axes == 15 && mode == 0) ref_hit[i] = h;
        else if (ref_hit[i] != (char)h) { printf("LABEL MISMATCH pose %d axes %d mode %d\n", i, axes, mode); return 2; }
        if (h) { ni++; bvi += S.bv - b0; lfi += S.leaf - l0; } else { nv++; bvv += S.bv - b0; lfv += S.leaf - l0; }
      }
      printf("axes %2d mode %d: all bv %.1f leaf %.2f | valid(%.3f) bv %.1f leaf %.2f | invalid bv %.1f leaf %.2f\n", axes, mode,
             (double)S.bv / N, (double)S.leaf / N, (double)nv / N, (double)bvv / nv, (double)lfv / nv, (double)bvi / ni, (double)lfi / ni);
    }
  return 0;
}
