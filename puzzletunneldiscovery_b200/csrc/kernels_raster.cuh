// kernels_raster.cuh -- K7: the barycentric UV-chart splat of the reference's renderer as a CUDA rasteriser
// (SURVEY.md 8(f) N4).  Reference: Renderer::addBarycentric / renderBarycentric
// (/root/reference/lib/osr/osr_render.cc:538-569, 590-759) with shaders lib/osr/shader/bary.{vert,geom,frag}:
//   geometry stage : vertex n of a primitive sits at  blend = uv0*b[n].x + uv1*b[n].y + uv2*b[n].z  (the three UV
//                    corners of the surface face blended by that vertex's barycentrics), NDC = 2*blend - 1,
//                    viewport res x res  ->  window position blend * res;
//   fragment stage : the barycentrics interpolated across the primitive; discard if their min < 0 or max > 1
//                    (the part of the primitive outside its surface face), else output the weight;
//   blending       : GL_FUNC_ADD, GL_ONE / GL_ONE into a float render target (additive), no depth, no culling;
//   read-back      : glReadPixels rows bottom-up: out[y][x], x along u, y along v.
// The GL path itself cannot run here (no graphics stack on the box), so coverage follows the usual hardware
// rule stated explicitly: a pixel is covered when its centre (x+0.5, y+0.5) is inside the triangle, edges by
// the top-left rule; attributes are interpolated linearly (w = 1).  Vertex positions are formed in fp32 like
// the shader (mul, add, add); edge functions in fp64.  One warp per primitive, lanes stride over the pixels of
// its bounding box, atomicAdd into the chart.
#pragma once
#include "kernels.cuh"

namespace osr {

__device__ __forceinline__ bool raster_top_left(double ex, double ey) {
  // edge vector (ex, ey) of a counter-clockwise triangle in a y-up window: "left" edges go down, "top" edges are
  // horizontal and go left
  return ey < 0.0 || (ey == 0.0 && ex < 0.0);
}

__global__ void __launch_bounds__(256)
    k_bary_raster(const float* __restrict__ uv, const float* __restrict__ bary, const float* __restrict__ weight,
                  long long nprim, int res_x, int res_y, float* __restrict__ out) {
  const long long p = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= nprim) return;
  float u[3][2], b[3][3], wt[3];
#pragma unroll
  for (int v = 0; v < 3; v++) {
    u[v][0] = __ldg(uv + 6 * p + 2 * v); u[v][1] = __ldg(uv + 6 * p + 2 * v + 1);
#pragma unroll
    for (int k = 0; k < 3; k++) b[v][k] = __ldg(bary + 9 * p + 3 * v + k);
    wt[v] = __ldg(weight + 3 * p + v);
  }
  double X[3], Y[3];
#pragma unroll
  for (int v = 0; v < 3; v++) {
    float bx = u[0][0] * b[v][0]; bx = bx + u[1][0] * b[v][1]; bx = bx + u[2][0] * b[v][2];
    float by = u[0][1] * b[v][0]; by = by + u[1][1] * b[v][1]; by = by + u[2][1] * b[v][2];
    X[v] = (double)bx * (double)res_x;
    Y[v] = (double)by * (double)res_y;
  }
  double area = (X[1] - X[0]) * (Y[2] - Y[0]) - (X[2] - X[0]) * (Y[1] - Y[0]);
  if (!(area != 0.0) || !isfinite(area)) return;
  int i1 = 1, i2 = 2;
  if (area < 0.0) { i1 = 2; i2 = 1; area = -area; }   // make it counter-clockwise
  const double x0 = X[0], y0 = Y[0], x1 = X[i1], y1 = Y[i1], x2 = X[i2], y2 = Y[i2];
  const bool tl0 = raster_top_left(x2 - x1, y2 - y1), tl1 = raster_top_left(x0 - x2, y0 - y2),
             tl2 = raster_top_left(x1 - x0, y1 - y0);
  double fx0 = floor(fmin(x0, fmin(x1, x2))), fx1 = ceil(fmax(x0, fmax(x1, x2)));
  double fy0 = floor(fmin(y0, fmin(y1, y2))), fy1 = ceil(fmax(y0, fmax(y1, y2)));
  int px0 = (int)fmax(fx0, 0.0), px1 = (int)fmin(fx1, (double)res_x), py0 = (int)fmax(fy0, 0.0),
      py1 = (int)fmin(fy1, (double)res_y);
  if (px1 <= px0 || py1 <= py0) return;
  const int bw = px1 - px0;
  const long long npx = (long long)bw * (py1 - py0);
  for (long long t = lane; t < npx; t += 32) {
    const int px = px0 + (int)(t % bw), py = py0 + (int)(t / bw);
    const double cx = px + 0.5, cy = py + 0.5;
    const double e0 = (x2 - x1) * (cy - y1) - (y2 - y1) * (cx - x1);   // weight of vertex 0
    const double e1 = (x0 - x2) * (cy - y2) - (y0 - y2) * (cx - x2);   // weight of vertex i1
    const double e2 = (x1 - x0) * (cy - y0) - (y1 - y0) * (cx - x0);   // weight of vertex i2
    const bool in = (e0 > 0.0 || (e0 == 0.0 && tl0)) && (e1 > 0.0 || (e1 == 0.0 && tl1)) &&
                    (e2 > 0.0 || (e2 == 0.0 && tl2));
    if (!in) continue;
    const double l0 = e0 / area, l1 = e1 / area, l2 = e2 / area;
    double fb[3];
#pragma unroll
    for (int k = 0; k < 3; k++) fb[k] = l0 * (double)b[0][k] + l1 * (double)b[i1][k] + l2 * (double)b[i2][k];
    const double mn = fmin(fb[0], fmin(fb[1], fb[2])), mx = fmax(fb[0], fmax(fb[1], fb[2]));
    if (mn < 0.0 || mx > 1.0) continue;   // bary.frag: discard
    const double fw = l0 * (double)wt[0] + l1 * (double)wt[i1] + l2 * (double)wt[i2];
    atomicAdd(out + (size_t)py * res_x + px, (float)fw);
  }
}

}  // namespace osr
