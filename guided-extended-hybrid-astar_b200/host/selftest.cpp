// selftest.cpp -- exercises the C++ host mirror end to end on cuda:0 (run by tests/test_host_cpp.py
// on the GPU box).  Mirrors the reference's gtest (LIB/test/cpp/collision_checker_lib_test/
// test_collision_checking.cpp:9-30) and checks size-independent properties of the grid builders.
#include <cstdio>
#include <cstdlib>
#include <random>

#include "gridaccel_host.hpp"

using namespace gridaccel;

#define EXPECT(c)                                                     \
  do                                                                  \
  {                                                                   \
    if (!(c))                                                         \
    {                                                                 \
      std::fprintf(stderr, "FAIL %s:%d %s\n", __FILE__, __LINE__, #c); \
      return 1;                                                       \
    }                                                                 \
  } while (0)

static void write_config(const char* path)
{
  std::FILE* f = std::fopen(path, "w");
  std::fputs("YAW_RES_COLL: 3 # must be a divisor of 15\nDOUBLE_DISK_ROWS: False\nLEN_PER_DISK: 2.0\n"
             "SAFETY_DISTANCE_M: 0.1\nSEARCH_DIST: 5.0\nMIN_THRESH: 100\nMAX_THRESH: 130\nMAX_PATCH_INS_DIST: 40\n"
             "ASTAR_UNKNOWN_COST: 0\nMAX_EXTRA_NODES_ASTAR: 10000\nHEURISTIC_EARLY_EXIT: True\n"
             "MOTION_RES_MIN: 0.1\nMOTION_RES_MAX: 1.0\nMIN_DIST_VAL: 3\nMAX_DIST_VAL: 10\n",
             f);
  std::fclose(f);
}

int main()
{
  const char* cfg = "/tmp/gridaccel_selftest_config.yml";
  write_config(cfg);
  Context ctx(0);

  // ---- the reference's known-answer test: vehicle (0.6, 2.7, 3.0, 1.5, 2.0), patch 1000, all UNKNOWN
  {
    Vehicle veh;
    veh.initialize(0.6, 2.7, 3.0, 1.5, 2.0, false);
    CollisionChecker cc(ctx, veh);
    cc.initialize(1000, cfg);
    EXPECT(cc.checkGrid(Pose<int>(500, 500, 0)) == true);
    EXPECT(cc.checkGrid(Pose<int>(0, 0, 0)) == false);
    EXPECT(cc.checkGrid(Pose<int>(999, 999, 0)) == false);
    EXPECT(cc.checkGridVariant(Pose<int>(0, 0, 0)) == true);
  }

  // ---- default vehicle: 3 disks, 21x21 dilation kernel, 120 yaw bins
  Vehicle veh;
  CollisionChecker cc(ctx, veh);
  AStar astar(ctx);
  const int dim = 1000;
  astar.initialize(dim, Point<double>(0, 0), cfg);
  cc.initialize(dim, cfg);
  astar.reinit(Point<double>(0, 0), dim);
  EXPECT(cc.nb_disks() == 3 && cc.disk_r_c_ == 10 && astar.astar_dim_ == 250);
  EXPECT(cc.getYawIdx(0.0) == 0 && cc.getYawIdx(-1e-9) == 119 && cc.getYawIdx(3.0 * TO_RAD + 1e-9) == 1);

  // free patch with one obstacle block and a 2-cell wall on the border
  Vec2DFlat<uint8_t> win;
  win.resize_and_reset(dim, dim, 0);
  for (int y = 0; y < dim; ++y)
    for (int x = 0; x < dim; ++x)
      if (x < 2 || y < 2 || x >= dim - 2 || y >= dim - 2 || (x >= 400 && x < 440 && y >= 300 && y < 360)) win(y, x) = 255;
  cc.passLocalMap(win, Point<int>(0, 0), dim);
  cc.processSafetyPatch();
  cc.fetchSafetyPatch();
  astar.fetchGrids();
  // dilation by the radius-10 disk: block grows by exactly 10 cells along the axes
  EXPECT(cc.patch_safety_arr_(330, 390) == CollisionChecker::OCC && cc.patch_safety_arr_(330, 389) == CollisionChecker::FREE);
  EXPECT(cc.patch_safety_arr_(290, 420) == CollisionChecker::OCC && cc.patch_safety_arr_(289, 420) == CollisionChecker::FREE);
  // pooled grid == 4x4 block max of the safety grid
  for (int Y = 0; Y < 250; Y += 7)
    for (int X = 0; X < 250; X += 5)
    {
      uint8_t m = 0;
      for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i) m = std::max(m, cc.patch_safety_arr_(4 * Y + j, 4 * X + i));
      EXPECT(astar.astar_grid_(Y, X) == m);
    }

  // pose checks: free space valid, inside the dilated block invalid, path index = first hit
  EXPECT(cc.checkPose(Pose<double>(30.0, 100.0, 0.3)) == true);
  EXPECT(cc.checkPose(Pose<double>(420 * GM_RES, 330 * GM_RES, 0.0)) == false);
  {
    std::vector<double> xs, ys, ws;
    for (int i = 0; i < 60; ++i)
    {
      xs.push_back((300 + 2 * i) * GM_RES);
      ys.push_back(330 * GM_RES);
      ws.push_back(0.0);
    }
    const int hit = cc.getPathCollisionIndex(xs, ys, ws);
    EXPECT(hit > 0 && hit < 60);
    for (int i = 0; i < hit; ++i) EXPECT(cc.checkPose(Pose<double>(xs[i], ys[i], ws[i])));
    EXPECT(!cc.checkPose(Pose<double>(xs[hit], ys[hit], ws[hit])));
    EXPECT(cc.checkPathCollision(xs, ys, ws) == false);
    EXPECT(cc.getPathCollisionIndex({}, {}, {}) == -1);
  }

  // heuristic: goal cost 0, every free cell closed, parent walk reaches the goal with cost
  // decreasing, Bellman consistency of the cost plane along the path
  const Point<int> goal(20, 20), start(230, 225);
  astar.reference_order_ = false;  // the GPU fixed point: every reachable cell closed
  astar.calcDistanceHeuristic(goal, start, false);
  const ClosedSet& cs = astar.closed_set_guidance_;
  EXPECT(cs.contains(astar.calcIndex(goal.x, goal.y)) && cs.at(astar.calcIndex(goal.x, goal.y)).cost_ == 0.0);
  size_t n_free = 0;
  for (size_t i = 0; i < astar.astar_grid_.size(); ++i) n_free += astar.astar_grid_.getPtr()[i] != CollisionChecker::OCC;
  EXPECT(cs.size() == n_free);
  size_t n_iter = 0;
  for (auto it = cs.begin(); it != cs.end(); ++it) ++n_iter;
  EXPECT(n_iter == cs.size());
  const auto path = astar.getAstarPath(start.x, start.y);
  EXPECT(!path.first.empty() && path.first.back() == goal.x && path.second.back() == goal.y);
  for (size_t i = 0; i + 1 < path.first.size(); ++i)
  {
    const NodeDisc a = cs.at(astar.calcIndex(path.first[i], path.second[i]));
    const NodeDisc b = cs.at(astar.calcIndex(path.first[i + 1], path.second[i + 1]));
    const bool diag = path.first[i] != path.first[i + 1] && path.second[i] != path.second[i + 1];
    const double step = diag ? ASTAR_RES * std::sqrt(2.0) : ASTAR_RES;
    EXPECT(a.parent_index_ == (int)astar.calcIndex(path.first[i + 1], path.second[i + 1]));
    EXPECT(a.cost_ == b.cost_ + step * 5.0);           // uniform movement cost 5, no prox yet
    EXPECT(std::fabs(a.cost_dist_ - (b.cost_dist_ + step)) < 1e-9);
  }
  EXPECT(cs.find(astar.calcIndex(105, 82)) == cs.end());  // inside the dilated block: not in the set

  // potential field with an empty end-point list: obstacle distances only
  const bool had = astar.calcVoronoiPotentialField(Point<int>(110, 100), [](const std::vector<int>&) { return std::vector<int>(); });
  EXPECT(had);
  astar.fetchGrids();
  EXPECT(astar.obstacle_dist_(0, 0) >= 0.0f);

  // expansion batch incl. rear-axis primitives
  cc.setExpandParams(ASTAR_RES * 1.5, 10.0, 90.0);
  std::vector<ga_node> nodes(64);
  std::mt19937 rng(1);
  std::uniform_real_distribution<double> u(10.0, 140.0), a(-PI, PI);
  for (auto& n : nodes) n = ga_node{ u(rng), u(rng), a(rng), 0.5, 1, 0 };
  std::vector<ga_succ> succ;
  cc.expandNodes(nodes, succ);
  EXPECT(cc.expandStride() == 32 && succ.size() == 64u * 32u);
  int n_valid = 0;
  for (const auto& s : succ) n_valid += s.valid;
  EXPECT(n_valid > 0);

  // proximity scoring: the potential field above left h_prox_arr_ on the device
  HybridAStar hy(ctx, veh);
  const double v0 = hy.getProxOfCorners(Pose<double>(70.0, 60.0, 0.3));
  EXPECT(v0 >= 0.0 && v0 <= 1.0);
  const std::vector<double> path_poses = { 70.0, 60.0, 0.3, 70.1, 60.0, 0.3, 70.2, 60.1, 0.31 };
  const std::vector<double> costs = hy.getPathProxCosts(path_poses, { 0, 1, 1, 3 });
  EXPECT(costs.size() == 3 && costs[1] == 0.0);
  EXPECT(costs[0] == v0 * 5.0 * 0.1);

  // close pose: the neighbourhood of a free cell; no cell list -> the ego pose comes back
  std::vector<int> near_cells;
  for (int dy = -6; dy <= 6; ++dy)
    for (int dx = -6; dx <= 6; ++dx) near_cells.push_back((int)astar.calcIndex(112 + dx, 96 + dy));
  const Pose<double> ego(1.0, 2.0, 0.5);
  const Pose<double> cp = hy.getValidClosePose(Pose<double>(70.0, 60.0, 0.3), near_cells, ego);
  EXPECT(std::fabs(cp.x - 70.0) <= 5.0 && std::fabs(cp.y - 60.0) <= 5.0);
  EXPECT(hy.getValidClosePose(Pose<double>(70.0, 60.0, 0.3), {}, ego).x == 1.0);

  std::printf("host selftest ok: %zu closed cells, path %zu cells, %d valid successors, %lld kernel launches\n", cs.size(),
              path.first.size(), n_valid, ga_launch_count(ctx.handle()));
  return 0;
}
