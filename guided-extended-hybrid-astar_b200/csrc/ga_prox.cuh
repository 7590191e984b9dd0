// ga_prox.cuh -- device side of the proximity scoring (HybridAStar::getProxOfCorners and the bilinear
// interpolation under it), shared by ga_prox.cu and the close-pose search in ga_collision.cu.
#pragma once

#include "ga_internal.cuh"

struct prox_args
{
  const double* grid;  // AStar::h_prox_arr_, dim x dim row-major
  int dim;
  ga_prox_params p;
};

// libstdc++'s std::lerp (C++20, <cmath> __lerp): exact at t = 0 / 1, monotone
__device__ __forceinline__ double lerp_std(const double a, const double b, const double t)
{
  if ((a <= 0 && b >= 0) || (a >= 0 && b <= 0)) return t * b + (1 - t) * a;
  if (t == 1) return b;
  const double x = a + t * (b - a);
  return (t > 1) == (b > a) ? (b < x ? x : b) : (b > x ? x : b);
}

// util::bilinearInterpolation (util1.cpp:152-200)
__device__ __forceinline__ double bilinear(const double q11, const double q12, const double q21, const double q22, const int x1,
                                           const int x2, const int y1, const int y2, const double x, const double y)
{
  const int x2x1 = x2 - x1, y2y1 = y2 - y1;
  const double yy1 = y - y1, xx1 = x - x1;
  double result;
  if (x2x1 == 0 && y2y1 == 0) result = q11;
  else if (x2x1 == 0) result = lerp_std(q11, q12, yy1 / y2y1);
  else if (y2y1 == 0) result = lerp_std(q11, q21, xx1 / x2x1);
  else
  {
    const double x2x = static_cast<double>(x2) - x, y2y = y2 - y;
    result = 1.0 / (x2x1 * y2y1) * (q11 * x2x * y2y + q21 * xx1 * y2y + q12 * x2x * yy1 + q22 * xx1 * yy1);
  }
  return isnan(result) ? 0.0 : result;
}

// util::getBilinInterp (util2.cpp:41-52).  The reference clamps the cell indices at 0 only and reads
// past the array for a point beyond the last row / column; here they are clamped into the grid.
__device__ __forceinline__ double bilin_at(const double x, const double y, const double* __restrict__ grid, const int dim)
{
  const int hi = dim - 1;
  const int x1 = min(max(static_cast<int>(x), 0), hi), x2 = min(max(static_cast<int>(ceil(x)), 0), hi);
  const int y1 = min(max(static_cast<int>(y), 0), hi), y2 = min(max(static_cast<int>(ceil(y)), 0), hi);
  const double q11 = __ldg(grid + (size_t)y1 * dim + x1), q12 = __ldg(grid + (size_t)y2 * dim + x1);
  const double q21 = __ldg(grid + (size_t)y1 * dim + x2), q22 = __ldg(grid + (size_t)y2 * dim + x2);
  return bilinear(q11, q12, q21, q22, x1, x2, y1, y2, x, y);
}

// getProxOfCorners(pose * con2star_): the corner offsets stay in METRES while the pose is in astar
// cells -- that is what the reference computes (hybrid_a_star_lib.cpp:270-274), kept as it is
__device__ __forceinline__ double prox_of_corners(const double px, const double py, const double yaw, const prox_args& a)
{
  double s, c;
  sincos(yaw, &s, &c);
  const double ox = px * a.p.con2star, oy = py * a.p.con2star;
  const double hw = a.p.width / 2;
  const double cx[4] = { a.p.lf, a.p.lf, -a.p.lb, -a.p.lb }, cy[4] = { -hw, hw, -hw, hw };  // f_r, f_l, r_r, r_l (vehicle.cpp:24-27)
  double best = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k)
  {
    const double x = (cx[k] * c - cy[k] * s) + ox, y = (cx[k] * s + cy[k] * c) + oy;  // rotatePreCalc + offset
    const double v = bilin_at(x, y, a.grid, a.dim);
    best = k == 0 ? v : (best < v ? v : best);  // std::max({v_1 .. v_4}); bilinear never returns NaN
  }
  return best;
}

