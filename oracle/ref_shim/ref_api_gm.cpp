// TEST INFRASTRUCTURE ONLY -- C entry points over the reference's GridMapSim
// (LIB/src/gridmap_sim_lib/gridmap_sim.cpp, compiled unmodified into oracle/_ref/libga_ref.so).
// A translation unit of its own: gridmap_sim.hpp's global enum (FREE / UNKNOWN / OCCUPIED) collides
// with data_structures2.hpp's PATH_TYPE when both are seen by one TU.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <cstdio>
#include <cstring>
#include <exception>
#define private public  // pmap_, precast_list_, is_initialized_ are private statics the checker must read
#include "gridmap_sim_lib/gridmap_sim.hpp"
#undef private

static char g_err[512];
#define GAR_TRY(body)                                   \
  try                                                   \
  {                                                     \
    body;                                               \
    return 0;                                           \
  }                                                     \
  catch (const std::exception& e)                       \
  {                                                     \
    std::snprintf(g_err, sizeof g_err, "%s", e.what()); \
    return -1;                                          \
  }

extern "C" {
const char* gar_gm_last_error() { return g_err; }

// ---- GridMapSim (LIB/src/gridmap_sim_lib/gridmap_sim.cpp): the simulated lidar ray-cast that
//      produces each cycle's sensor window.  raytracing takes a py::array_t: built here under the GIL
//      (the caller is a Python process; ctypes released the GIL for the call).
int gar_raytrace(const uint8_t* grid_map_gt, uint8_t* pmap_out)
{
  GAR_TRY({
    py::gil_scoped_acquire gil;
    py::array_t<uint8_t> gt({ GM_DIM, GM_DIM }, grid_map_gt);
    if (!GridMapSim::is_initialized_) GridMapSim::initialize();
    GridMapSim::raytracing(gt);
    std::memcpy(pmap_out, GridMapSim::pmap_.getPtr(), GM_DIM * GM_DIM);
  })
}
// the reference's precast lists (flattened) -- to compare list ORDER, which std::sort's handling of
// equal distances decides
int gar_precast_lists(int* entries, int cap, int* offsets)
{
  if (!GridMapSim::is_initialized_) GridMapSim::initialize();
  int n = 0;
  for (int k = 0; k < NB_PRECASTS; ++k)
  {
    offsets[k] = n;
    for (const PrecastDB& e : GridMapSim::precast_list_[k])
    {
      if (n < cap) entries[n] = e.ix_ * (int)GM_DIM + e.iy_;
      ++n;
    }
  }
  offsets[NB_PRECASTS] = n;
  return n;
}
}  // extern "C"
