// kernels_surface.cuh -- K5: map an intersection mesh onto the surface of the robot / environment
// (UnitWorld::intersectingToSurface, /root/reference/lib/osr/unit_world.cc:1092-1176, reached through
// intersectingToRobotSurface / intersectingToModelSurface :928-957).
//
// Reference: tritri::TriTriCopIsect (lib/tritri/tritri_cop.hh:115-151) runs the Moeller interval test
// (lib/tritri/Moller97d.cc:579-717) on every (target face i, intersection face j) pair and keeps the pairs
// that intersect AND are coplanar; the kept pairs are visited column-major (j outer, i inner -- the default
// Eigen::SparseMatrix order) and for each the barycentric coordinates of the three vertices of face j with
// respect to target triangle i are computed with libigl's formula
// (third-party/libigl/include/igl/barycentric_coordinates.cpp: v = (d11 d20 - d01 d21)/den, ..., u = 1-(v+w)).
//
// Here: one thread per pair (p = j * nT + i, so the flat order IS the reference's output order), a flag byte
// and a per-block count in pass 1; the host scans the block counts; pass 2 ranks the flagged pairs inside
// each block with ballots and writes the rows in order.  All arithmetic fp64, -fmad=false.
#pragma once
#include "kernels.cuh"

namespace osr {

constexpr int SURF_BLOCK = 256;

// world-space vertices of target triangle i: Tf * v, same operation order as k_contact_rows
__device__ __forceinline__ void surf_target_tri(const double* __restrict__ tris, const Rigid& tf, long long i,
                                                double (*T)[3]) {
  const double* Pt = tris + (size_t)9 * i;
#pragma unroll
  for (int v = 0; v < 3; v++) {
    double x = __ldg(Pt + 3 * v), y = __ldg(Pt + 3 * v + 1), z = __ldg(Pt + 3 * v + 2);
#pragma unroll
    for (int k = 0; k < 3; k++) T[v][k] = (tf.R[3 * k] * x + tf.R[3 * k + 1] * y + tf.R[3 * k + 2] * z) + tf.t[k];
  }
}

__device__ __forceinline__ void surf_isect_tri(const double* __restrict__ V, const int* __restrict__ F, long long j,
                                               double (*U)[3]) {
#pragma unroll
  for (int v = 0; v < 3; v++) {
    int vi = __ldg(F + 3 * j + v);
#pragma unroll
    for (int k = 0; k < 3; k++) U[v][k] = __ldg(V + 3 * (size_t)vi + k);
  }
}

// pass 1: flags[p] = intersecting && coplanar; block_count[b] = number of flagged pairs of block b
__global__ void __launch_bounds__(SURF_BLOCK)
    k_surface_flags(const double* __restrict__ tgt_tris, Rigid tf, long long nT, const double* __restrict__ V,
                    const int* __restrict__ F, long long nF, uint8_t* __restrict__ flags, int* __restrict__ block_count) {
  __shared__ int warp_cnt[SURF_BLOCK / 32];
  long long p = (long long)blockIdx.x * SURF_BLOCK + threadIdx.x;
  bool keep = false;
  if (p < nT * nF) {
    long long j = p / nT, i = p - j * nT;
    double T[3][3], U[3][3];
    surf_target_tri(tgt_tris, tf, i, T);
    surf_isect_tri(V, F, j, U);
    int cop = 0;
    double p1[3], p2[3];
    int isect = moller_isectline(T, U, &cop, p1, p2);
    keep = isect && cop;
    flags[p] = keep ? 1 : 0;
  }
  unsigned m = __ballot_sync(FULL, keep);
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = __popc(m);
  __syncthreads();
  if (threadIdx.x == 0) {
    int c = 0;
#pragma unroll
    for (int k = 0; k < SURF_BLOCK / 32; k++) c += warp_cnt[k];
    block_count[blockIdx.x] = c;
  }
}

// libigl barycentric_coordinates for one point
__device__ __forceinline__ void barycentric_igl(const double* p, const double* a, const double* b, const double* c,
                                                double* L) {
  double v0[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, v1[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]},
         v2[3] = {p[0] - a[0], p[1] - a[1], p[2] - a[2]};
  double d00 = dotT(v0, v0), d01 = dotT(v0, v1), d11 = dotT(v1, v1), d20 = dotT(v2, v0), d21 = dotT(v2, v1);
  double den = d00 * d11 - d01 * d01;
  L[1] = (d11 * d20 - d01 * d21) / den;
  L[2] = (d00 * d21 - d01 * d20) / den;
  L[0] = 1.0 - (L[1] + L[2]);
}

// pass 2: row = block_offset[b] + rank inside the block; out_face[row] = i, out_L[3 row + v] = barycentrics
__global__ void __launch_bounds__(SURF_BLOCK)
    k_surface_emit(const double* __restrict__ tgt_tris, Rigid tf, long long nT, const double* __restrict__ V,
                   const int* __restrict__ F, long long nF, const uint8_t* __restrict__ flags,
                   const long long* __restrict__ block_offset, long long cap, int* __restrict__ out_face,
                   double* __restrict__ out_L) {
  __shared__ int warp_cnt[SURF_BLOCK / 32];
  long long p = (long long)blockIdx.x * SURF_BLOCK + threadIdx.x;
  bool keep = p < nT * nF && flags[p] != 0;
  unsigned m = __ballot_sync(FULL, keep);
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) warp_cnt[wid] = __popc(m);
  __syncthreads();
  if (!keep) return;
  int before = __popc(m & ((1u << lane) - 1u));
  for (int k = 0; k < wid; k++) before += warp_cnt[k];
  long long row = block_offset[blockIdx.x] + before;
  if (row >= cap) return;
  long long j = p / nT, i = p - j * nT;
  double T[3][3], U[3][3];
  surf_target_tri(tgt_tris, tf, i, T);
  surf_isect_tri(V, F, j, U);
  out_face[row] = (int)i;
#pragma unroll
  for (int v = 0; v < 3; v++) barycentric_igl(U[v], T[0], T[1], T[2], out_L + 9 * row + 3 * v);
}

}  // namespace osr
