// geometry.h -- host-side geometry front end of the B200 collision engine.
//
// Replaces, for the hot path only, the reference's loader + unit-world set-up
// (/root/reference/lib/osr/scene.cc:78-156, lib/osr/unit_world.cc:113-161,
// lib/osr/cdmodel.cc:85-160) and FCL's BVH construction
// (third-party/fcl/include/fcl/geometry/bvh/BVH_model-inl.h:855-925).  The BVH here is NOT
// FCL's tree: it is built for the GPU traversal (breadth-first node order, fp32 boxes that
// are provably supersets of the fp64 triangles).  Validity labels do not depend on the tree
// as long as every box contains its triangles.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace osr {

struct TriMesh {
  std::vector<float> V;  // xyz, referenced vertices only
  std::vector<int> F;    // 3 indices per triangle
  int nv() const { return (int)(V.size() / 3); }
  int nf() const { return (int)(F.size() / 3); }
};

// Reads "v" and "f" records (fan triangulation; negative indices; v/vt/vn tokens).
bool load_obj(const std::string& path, TriMesh& out, std::string& err);

// Device-facing BVH: 64-byte nodes (16 floats) in breadth-first order.
//   [0..2]  axis0.xyz   [3]  centre.x
//   [4..6]  axis1.xyz   [7]  centre.y
//   [8..10] axis2.xyz   [11] centre.z
//   [12..14] half extents (already inflated to cover fp32 rounding of axes/centre)
//   [15] int32 bit pattern: >=0 index of left child (right = left+1); <0 leaf, triangle = -(v+1)
struct GpuBvh {
  std::vector<float> nodes;       // 16 floats per node (layout above)
  std::vector<uint32_t> qnodes;   // 8 words per node: the 32-byte QNode of qnode.h (same numbering)
  int num_nodes = 0;
  int depth = 0;
  float max_radius = 0.f;         // largest |vertex| of the model (bounds |p| in the pose slack, qnode.h)
};

struct UnitModel {
  std::vector<double> V;     // unit-space vertices (fp64), xyz
  std::vector<int> F;        // triangles
  std::vector<double> tris;  // 9 doubles per triangle (soup, face order)
  GpuBvh bvh;
  double aabb_min[3], aabb_max[3];  // local AABB (cdmodel.cc:91-100)
  int nf() const { return (int)(F.size() / 3); }
};

// bbox.h:21-46 semantics: box seeded at the origin, float arithmetic.
float bbox_span(const TriMesh& m);
void mean_of_vertices(const TriMesh& m, float c[3]);

// unit-space vertices: double(s)*double(v) + double(float(s*-c)) (cdmodel.cc:120-139 with the
// float glm matrix S*T(-c) of unit_world.cc:137-161).  centre==nullptr -> no translation (env).
void make_unit_model(const TriMesh& m, float scale, const float* centre, UnitModel& out);

void build_gpu_bvh(UnitModel& m);

}  // namespace osr
