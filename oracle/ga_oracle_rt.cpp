// TEST INFRASTRUCTURE ONLY -- oracle of the simulated lidar ray-cast (GridMapSim,
// LIB/src/gridmap_sim_lib/gridmap_sim.cpp).  C++ and not C for one reason: the reference orders its
// precast lists with std::sort (gridmap_sim.cpp:36-39), whose treatment of equal distances decides
// which of two equidistant cells a ray meets first; the restatement has to use the same algorithm
// on the same input order.  Part of libga_oracle.so; used by tests/ and smoke() only.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

namespace
{
constexpr int kDim = 801;         // GM_DIM, gridmap_sim.hpp:17
constexpr int kLists = 180;       // NB_PRECASTS, :21
constexpr double kPi = 3.14159265358979323846;
constexpr double kYawRes = 2 * (kPi / 180);  // YAW_RES, :20

struct Cast  // PrecastDB, gridmap_sim.hpp:74-90
{
  double d;
  int ix, iy;
  bool operator<(const Cast& o) const { return d < o.d; }
};
}  // namespace

extern "C" {

// gridmap_sim.cpp:14-40: every cell goes into the list of floor(angle / 2deg) and of ceil(angle / 2deg)
// (twice into the same list when both agree), lists sorted by distance from the centre cell.
// entries = ix * 801 + iy (the flat index both grid_map_data(ix, iy) and pmap_(ix, iy) address).
int gao_precast_lists(int* entries, int cap, int* offsets)
{
  const int middle = static_cast<int>(static_cast<double>(kDim - 1) / 2);  // gridmap_sim.hpp:96
  std::vector<std::vector<Cast>> lists(kLists);
  for (int ix = 0; ix < kDim; ++ix)
  {
    for (int iy = 0; iy < kDim; ++iy)
    {
      const double dist = hypot(ix - middle, iy - middle);
      double angle = atan2(iy - middle, ix - middle);  // atan2Pi, :82-90
      if (angle < 0) angle += 2 * kPi;
      const int a1 = std::min(static_cast<int>(floor(angle / kYawRes)), kLists - 1);
      const int a2 = std::min(static_cast<int>(ceil(angle / kYawRes)), kLists - 1);
      const Cast c{ dist, ix, iy };
      lists[a1].push_back(c);
      lists[a2].push_back(c);
    }
  }
  int n = 0;
  for (int k = 0; k < kLists; ++k)
  {
    std::sort(lists[k].begin(), lists[k].end());
    offsets[k] = n;
    for (const Cast& c : lists[k])
    {
      if (n < cap) entries[n] = c.ix * kDim + c.iy;
      ++n;
    }
  }
  offsets[kLists] = n;
  return n;
}

// gridmap_sim.cpp:42-80: per list, the first OCCUPIED (255) cell of the ground truth is marked
// occupied and everything before it free; a list without a hit is free throughout.  Lists run in
// order, later ones overwrite earlier ones.
void gao_raytrace(const uint8_t* grid_map_gt, const int* entries, const int* offsets, uint8_t* pmap)
{
  for (int i = 0; i < kDim * kDim; ++i) pmap[i] = 127;
  for (int k = 0; k < kLists; ++k)
  {
    const int b = offsets[k], e = offsets[k + 1];
    int hit = e;
    for (int i = b; i < e; ++i)
    {
      if (grid_map_gt[entries[i]] == 255)
      {
        hit = i;
        break;
      }
    }
    if (hit < e) pmap[entries[hit]] = 255;
    for (int i = b; i < hit; ++i) pmap[entries[i]] = 0;
  }
}
}  // extern "C"
