/*
 * ga_oracle.c -- CPU restatement of the reference's grid hot path (see ga_oracle.h).
 * TEST INFRASTRUCTURE ONLY: never linked into, loaded by or called from libgridaccel.so.
 * Parity: pinned (cv2 4.13.0 golden vectors + the reference TUs built into oracle/_ref).
 *
 * Build: gcc -O2 -std=c11 -fPIC -shared -ffp-contract=off -o libga_oracle.so ga_oracle.c -lm
 * (-ffp-contract=off: the reference is built for baseline x86-64, i.e. without FMA.)
 */
#include "ga_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ========================================================================================
 * occupancy: threshold, dilation, pooling
 * ====================================================================================== */

/* OpenCV saturate_cast<int>(double) == cvRound == round-half-to-even (lrint in the default
 * rounding mode). */
static int cv_round(double v) { return (int)lrint(v); }

int gao_ellipse_rows(int r, int* half_w, uint8_t* se)
{
  /* OpenCV 4.8.0 modules/imgproc/src/morph.dispatch.cpp getStructuringElement, MORPH_ELLIPSE:
   * r = ksize.height/2, c = ksize.width/2, inv_r2 = 1/r^2,
   * row i: dy = i-r, dx = saturate_cast<int>(c*sqrt((r*r-dy*dy)*inv_r2)), cols [c-dx, c+dx] */
  const int d = 2 * r + 1;
  const int c = r;
  const double inv_r2 = r ? 1.0 / ((double)r * r) : 0.0;
  int taps = 0;
  for (int i = 0; i < d; ++i)
  {
    const int dy = i - r;
    int dx = cv_round(c * sqrt((double)(r * r - dy * dy) * inv_r2));
    int j1 = c - dx;
    int j2 = c + dx + 1;
    if (j1 < 0) j1 = 0;
    if (j2 > d) j2 = d;
    half_w[i] = dx;
    taps += j2 - j1;
    if (se)
    {
      for (int j = 0; j < d; ++j) se[i * d + j] = (uint8_t)(j >= j1 && j < j2);
    }
  }
  return taps;
}

void gao_threshold(const uint8_t* patch, uint8_t* out, size_t n, int min_thresh, int max_thresh)
{
  /* collision_checking.cpp:213-221; min_thresh_/max_thresh_ are uint8_t (hpp:51-52) */
  const uint8_t lo = (uint8_t)min_thresh;
  const uint8_t hi = (uint8_t)max_thresh;
  for (size_t i = 0; i < n; ++i)
  {
    const uint8_t v = patch[i];
    /* setTo(FREE), then setTo(OCC), then setTo(UNKNOWN where neither): OCC wins over FREE */
    out[i] = (v > hi) ? GAO_OCC : ((v < lo) ? GAO_FREE : GAO_UNKNOWN);
  }
}

void gao_dilate_disk(const uint8_t* in, uint8_t* out, int dim, int r)
{
  const int d = 2 * r + 1;
  int* hw = (int*)malloc(sizeof(int) * (size_t)d);
  gao_ellipse_rows(r, hw, NULL);
  /* values are {0,1,2}: a max filter.  Row-span decomposition keeps the brute force fast
   * enough for 4096^2 inside the bounded cpu_baseline sample; it is still a plain max over
   * the structuring element's in-image support. */
  for (int y = 0; y < dim; ++y)
  {
    for (int x = 0; x < dim; ++x)
    {
      uint8_t m = 0;
      for (int i = 0; i < d && m < 2; ++i)
      {
        const int yy = y + i - r;
        if (yy < 0 || yy >= dim) continue;
        int x0 = x - hw[i];
        int x1 = x + hw[i];
        if (x0 < 0) x0 = 0;
        if (x1 > dim - 1) x1 = dim - 1;
        const uint8_t* row = in + (size_t)yy * dim;
        for (int xx = x0; xx <= x1; ++xx)
        {
          if (row[xx] > m) m = row[xx];
        }
      }
      out[(size_t)y * dim + x] = m;
    }
  }
  free(hw);
}

void gao_process_safety_patch(const uint8_t* patch, uint8_t* safety, int dim, int min_thresh,
                              int max_thresh, int disk_r_c)
{
  const size_t n = (size_t)dim * dim;
  uint8_t* tmp = (uint8_t*)malloc(n);
  gao_threshold(patch, tmp, n, min_thresh, max_thresh);
  gao_dilate_disk(tmp, safety, dim, disk_r_c);
  free(tmp);
}

void gao_pool_max(const uint8_t* safety, uint8_t* astar, int patch_dim, int astar_dim, int win)
{
  for (int Y = 0; Y < astar_dim; ++Y)
  {
    for (int X = 0; X < astar_dim; ++X)
    {
      uint8_t m = 0;
      for (int j = 0; j < win; ++j)
      {
        const uint8_t* row = safety + (size_t)(Y * win + j) * patch_dim + (size_t)X * win;
        for (int i = 0; i < win; ++i)
        {
          if (row[i] > m) m = row[i];
        }
      }
      astar[(size_t)Y * astar_dim + X] = m;
    }
  }
}

void gao_pass_local_map(uint8_t* patch, int patch_dim, const uint8_t* local_map, int ox, int oy,
                        int dim)
{
  /* collision_checking.cpp:287-312: drop the whole window when it overhangs in x, skip rows
   * that overhang in y */
  if (ox < 0 || ox + dim > patch_dim) return;
  for (int i = 0; i < dim; ++i)
  {
    const int row = oy + i;
    if (row > patch_dim - 1 || row < 0) continue;
    memcpy(patch + (size_t)row * patch_dim + ox, local_map + (size_t)i * dim, (size_t)dim);
  }
}

/* cartographing.cpp:17-96 (both overloads are the same loop) */
void gao_cartograph(uint8_t* cart, int patch_dim, const uint8_t* local_map, int ox, int oy, int dim)
{
  for (int x = 0; x < dim; ++x)
  {
    for (int y = 0; y < dim; ++y)
    {
      const int px = ox + x, py = oy + y;
      if (px >= patch_dim || px < 0 || py >= patch_dim || py < 0) continue;
      uint8_t* cell = cart + (size_t)py * patch_dim + px;
      if (*cell != 127) continue; /* only overwrite unknown parts */
      const uint8_t v = local_map[(size_t)y * dim + x];
      if (v == 255 || v == 0) *cell = v; /* SENSOR_FREE = 255, SENSOR_OCC = 0 (collision_checking.hpp:44-46) */
    }
  }
}

/* util2.hpp:32-67 */
void gao_copy_patch_to_patch(const uint8_t* prev, int prev_dim, int prev_ox, int prev_oy, uint8_t* next,
                             int next_dim, int ox, int oy)
{
  const int opp_x = prev_ox + prev_dim, opp_y = prev_oy + prev_dim;
  const int sp_x = ox - prev_ox > 0 ? ox - prev_ox : 0, sp_y = oy - prev_oy > 0 ? oy - prev_oy : 0;
  const int sn_x = prev_ox - ox > 0 ? prev_ox - ox : 0, sn_y = prev_oy - oy > 0 ? prev_oy - oy : 0;
  const int en_x = opp_x - ox < next_dim - 1 ? opp_x - ox : next_dim - 1;
  const int en_y = opp_y - oy < next_dim - 1 ? opp_y - oy : next_dim - 1;
  const int nx = en_x - sn_x, ny = en_y - sn_y;
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y)
      next[(size_t)(sn_y + y) * next_dim + sn_x + x] = prev[(size_t)(sp_y + y) * prev_dim + sp_x + x];
}

/* ========================================================================================
 * footprint disks + collision lookups
 * ====================================================================================== */

static const double TO_RAD = M_PI / 180.0;
static const double CON2GM = 1.0 / GAO_GM_RES;     /* transforms.hpp:73-74 */
static const double CON2STAR = 1.0 / GAO_ASTAR_RES; /* transforms.hpp:76-77 */

int gao_calc_disks(double veh_length, double veh_width, double veh_lb, double len_per_disk,
                   int double_disk_rows, double safety_distance_m, int yaw_res_deg,
                   gao_disks_info* info, int* centers, int centers_cap)
{
  const unsigned int eff_nb = (unsigned int)(int)ceil(veh_length / len_per_disk);
  const int mult = double_disk_rows ? 2 : 1;
  const int nb_disks = mult * (int)eff_nb;
  const double eff_width = veh_width / mult;
  const int nb_yaws = 360 / yaw_res_deg;
  const double yaw_res_rad = yaw_res_deg * TO_RAD;

  /* getDiskRadius :44-47 */
  const double radius =
      sqrt(((veh_length * veh_length) / (eff_nb * eff_nb) + (eff_width * eff_width)) / 4);
  /* getDiskPositions :57-70 */
  const double dist = 2 * sqrt((radius * radius) - (eff_width * eff_width) / 4);
  const double first_pos = dist / 2 - veh_lb;

  const double safety_r = radius + safety_distance_m;
  info->nb_disks = nb_disks;
  info->nb_yaws = nb_yaws;
  info->yaw_res_deg = yaw_res_deg;
  info->disk_r = safety_r;
  info->disk_r_c = (int)ceil(safety_r * CON2GM);
  if (nb_disks * nb_yaws * 2 > centers_cap) return -1;

  for (int yi = 0; yi < nb_yaws; ++yi)
  {
    const double yaw = yi * yaw_res_rad;
    int k = 0;
    for (unsigned int i = 0; i < eff_nb; ++i)
    {
      const double px = first_pos + (int)i * dist;
      for (int rrow = 0; rrow < mult; ++rrow)
      {
        const double py = double_disk_rows ? (rrow == 0 ? eff_width / 2.0 : -eff_width / 2.0) : 0.0;
        /* Point::rotate data_structures1.hpp:62-65, operator* :73-76, toInt :57-60 */
        const double rx = px * cos(yaw) - py * sin(yaw);
        const double ry = px * sin(yaw) + py * cos(yaw);
        centers[(yi * nb_disks + k) * 2 + 0] = (int)round(rx * CON2GM);
        centers[(yi * nb_disks + k) * 2 + 1] = (int)round(ry * CON2GM);
        ++k;
      }
    }
  }
  return 0;
}

double gao_angle_0_2pi(double a)
{
  while (a < 0) a += 2 * M_PI;
  return fmod(a, 2 * M_PI);
}

double gao_angle_mpi_pi(double a)
{
  while (a < 0) a += 2 * M_PI;
  return fmod(a + M_PI, 2.0 * M_PI) - M_PI;
}

double gao_angle_diff(double a1, double a2)
{
  a2 = gao_angle_mpi_pi(a2);
  a1 = gao_angle_mpi_pi(a1);
  return gao_angle_mpi_pi(a2 - a1);
}

int gao_yaw_idx(double yaw, int yaw_res_deg)
{
  const double res_rad = (unsigned int)yaw_res_deg * TO_RAD;
  const int nb = 360 / yaw_res_deg;
  const int idx = (int)(gao_angle_0_2pi(yaw) / res_rad);
  return idx < nb - 1 ? idx : nb - 1;
}

int gao_check_grid(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks, int x,
                   int y, int yaw_idx, int variant)
{
  for (int k = 0; k < nb_disks; ++k)
  {
    const int px = centers[(yaw_idx * nb_disks + k) * 2 + 0] + x;
    const int py = centers[(yaw_idx * nb_disks + k) * 2 + 1] + y;
    if (px >= patch_dim || py >= patch_dim || px < 0 || py < 0)
    {
      if (variant) continue;
      return 0;
    }
    if (safety[(size_t)py * patch_dim + px] == GAO_OCC) return 0;
  }
  return 1;
}

int gao_check_pose(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks,
                   int yaw_res_deg, double x, double y, double yaw, int variant)
{
  const int gx = (int)(x * CON2GM);
  const int gy = (int)(y * CON2GM);
  return gao_check_grid(safety, patch_dim, centers, nb_disks, gx, gy,
                        gao_yaw_idx(yaw, yaw_res_deg), variant);
}

int gao_path_collision_index(const uint8_t* safety, int patch_dim, const int* centers,
                             int nb_disks, int yaw_res_deg, const double* x, const double* y,
                             const double* yaw, int n, int variant)
{
  for (int i = 0; i < n; ++i)
  {
    if (!gao_check_pose(safety, patch_dim, centers, nb_disks, yaw_res_deg, x[i], y[i], yaw[i],
                        variant))
      return i;
  }
  return -1;
}

void gao_check_segments(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks,
                        int yaw_res_deg, const double* x, const double* y, const double* yaw,
                        const int* seg_off, int nseg, int variant, int* first_hit)
{
  for (int s = 0; s < nseg; ++s)
  {
    const int b = seg_off[s];
    first_hit[s] = gao_path_collision_index(safety, patch_dim, centers, nb_disks, yaw_res_deg,
                                            x + b, y + b, yaw + b, seg_off[s + 1] - b, variant);
  }
}

/* ========================================================================================
 * motion primitives
 * ====================================================================================== */

void gao_vehicle_move(double* x, double* y, double* yaw, double distance, double steer,
                      double wheel_base)
{
  *x += distance * cos(*yaw);
  *y += distance * sin(*yaw);
  *yaw += distance * tan(steer) / wheel_base;
  *yaw = gao_angle_mpi_pi(*yaw);
}

int gao_move_car_some_steps(double x, double y, double yaw, double arc_l, double motion_res,
                            int direction, double steer, double wheel_base, double* xs,
                            double* ys, double* yaws, int cap)
{
  const size_t nb = (size_t)round(arc_l / motion_res);
  int n = 0;
  for (size_t i = 0; i < nb && n < cap; ++i)
  {
    gao_vehicle_move(&x, &y, &yaw, motion_res * direction, steer, wheel_base);
    xs[n] = x;
    ys[n] = y;
    yaws[n] = yaw;
    ++n;
  }
  return n;
}

static int sgn_d(double v) { return (0.0 < v) - (v < 0.0); }

int gao_turn_on_rear_axis(double yaw, double delta_angle, double yaw_res_deg, double* yaws,
                          int cap)
{
  const double goal_yaw = gao_angle_mpi_pi(yaw + delta_angle);
  const double res = yaw_res_deg * TO_RAD;
  double margin;
  int n = 0;
  do
  {
    yaw = gao_angle_mpi_pi(yaw + sgn_d(delta_angle) * res);
    if (n < cap) yaws[n] = yaw;
    ++n;
    margin = fabs(gao_angle_diff(yaw, goal_yaw));
  } while (margin >= res && n < cap);
  if (yaws[n - 1] != goal_yaw) yaws[n - 1] = goal_yaw;
  return n;
}

void gao_steering_inputs(double max_steer, int nb_steer, double* out)
{
  const double delta = (2 * max_steer) / (double)(nb_steer - 1);
  const int side = nb_steer / 2;
  for (int j = 0, i = -side; i <= side; ++i, ++j) out[j] = i * delta;
}

int gao_expand_node(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks,
                    int yaw_res_coll_deg, const gao_expand_params* p, double x, double y,
                    double yaw, double motion_res, uint8_t* valid, int* disc, double* end_pose,
                    int* n_poses, int cap)
{
  double steers[64];
  double xs[512], ys[512], yaws[512];
  int k = 0;
  /* a_star_yaw_res_inv_ = 1 / (YAW_RES * TO_RAD), hybrid_a_star_lib.cpp:38-40 */
  const double yaw_res_inv = 1.0 / (p->yaw_res_astar_deg * TO_RAD);
  gao_steering_inputs(p->max_steer, p->nb_steer, steers);
  for (int s = 0; s < p->nb_steer; ++s)
  {
    for (int d = 0; d < 2 && k < cap; ++d)
    {
      const int dir = d == 0 ? 1 : -1; /* direction_inputs_ = {1,-1}, hybrid_a_star_lib.hpp:50 */
      const int n = gao_move_car_some_steps(x, y, yaw, p->arc_l, motion_res, dir, steers[s],
                                            p->wheel_base, xs, ys, yaws, 512);
      n_poses[k] = n;
      valid[k] = 0;
      if (n > 0)
      {
        const int xi = (int)(xs[n - 1] * CON2STAR);
        const int yi = (int)(ys[n - 1] * CON2STAR);
        const int yawi = (int)round(yaws[n - 1] * yaw_res_inv);
        disc[k * 3 + 0] = xi;
        disc[k * 3 + 1] = yi;
        disc[k * 3 + 2] = yawi;
        end_pose[k * 3 + 0] = xs[n - 1];
        end_pose[k * 3 + 1] = ys[n - 1];
        end_pose[k * 3 + 2] = yaws[n - 1];
        const int on_grid = (-1 < xi && xi < p->astar_dim) && (-1 < yi && yi < p->astar_dim);
        if (on_grid)
        {
          valid[k] = (uint8_t)(gao_path_collision_index(safety, patch_dim, centers, nb_disks,
                                                        yaw_res_coll_deg, xs, ys, yaws, n, 0) < 0);
        }
      }
      ++k;
    }
  }
  if (p->with_rear_axis)
  {
    const double step = p->turn_on_point_angle_deg * TO_RAD;
    /* same accumulating loop as hybrid_a_star_lib.cpp:519-531 */
    for (double da = -2 * M_PI + step; da <= 2 * M_PI - step && k < cap; da += step)
    {
      if (da == 0) continue;
      const int n = gao_turn_on_rear_axis(yaw, da, (double)yaw_res_coll_deg, yaws, 512);
      for (int i = 0; i < n; ++i)
      {
        xs[i] = x;
        ys[i] = y;
      }
      n_poses[k] = n;
      disc[k * 3 + 0] = (int)(x * CON2STAR);
      disc[k * 3 + 1] = (int)(y * CON2STAR);
      disc[k * 3 + 2] = (int)round(yaws[n - 1] * yaw_res_inv);
      end_pose[k * 3 + 0] = x;
      end_pose[k * 3 + 1] = y;
      end_pose[k * 3 + 2] = yaws[n - 1];
      valid[k] = (uint8_t)(gao_path_collision_index(safety, patch_dim, centers, nb_disks,
                                                    yaw_res_coll_deg, xs, ys, yaws, n, 0) < 0);
      ++k;
    }
  }
  return k;
}

/* ========================================================================================
 * heuristic grid
 * ====================================================================================== */

/* a_star.hpp:50-57 */
static const int MOT_DX[8] = { 1, 0, -1, 0, -1, -1, 1, 1 };
static const int MOT_DY[8] = { 0, 1, 0, -1, -1, 1, -1, 1 };

typedef struct
{
  double prio;
  size_t id;
} heap_el;

typedef struct
{
  heap_el* a;
  size_t n, cap;
} min_heap;

/* (priority, id) lexicographic, as std::greater<std::pair<double,size_t>>
 * (data_structures2.hpp:236-237) */
static int el_less(const heap_el* u, const heap_el* v)
{
  if (u->prio < v->prio) return 1;
  if (v->prio < u->prio) return 0;
  return u->id < v->id;
}

static void heap_push(min_heap* h, double prio, size_t id)
{
  if (h->n == h->cap)
  {
    h->cap = h->cap ? h->cap * 2 : 1024;
    h->a = (heap_el*)realloc(h->a, h->cap * sizeof(heap_el));
  }
  size_t i = h->n++;
  heap_el e = { prio, id };
  while (i > 0)
  {
    const size_t parent = (i - 1) / 2;
    if (!el_less(&e, &h->a[parent])) break;
    h->a[i] = h->a[parent];
    i = parent;
  }
  h->a[i] = e;
}

static heap_el heap_pop(min_heap* h)
{
  const heap_el top = h->a[0];
  const heap_el last = h->a[--h->n];
  size_t i = 0;
  for (;;)
  {
    size_t c = 2 * i + 1;
    if (c >= h->n) break;
    if (c + 1 < h->n && el_less(&h->a[c + 1], &h->a[c])) ++c;
    if (!el_less(&h->a[c], &last)) break;
    h->a[i] = h->a[c];
    i = c;
  }
  if (h->n) h->a[i] = last;
  return top;
}

static int verify_node(int dim, const uint8_t* grid, const uint8_t* fence, int x, int y)
{
  /* a_star.cpp:325-343 */
  if (x < 0 || y < 0 || x >= dim || y >= dim) return 0;
  if (grid[(size_t)y * dim + x] == GAO_OCC) return 0;
  if (fence && fence[(size_t)y * dim + x] == 1) return 0;
  return 1;
}

int gao_heuristic_reforder(const gao_heur_params* p, const uint8_t* astar_grid,
                           const uint8_t* geofence, const double* movement_cost,
                           const double* h_prox, int goal_x, int goal_y, int start_x,
                           int start_y, double* cost, double* cost_dist, int32_t* parent,
                           uint8_t* flags, int32_t* closed_order, int32_t* near_ids,
                           double* near_dist, int* n_near)
{
  const int dim = p->astar_dim;
  const size_t ncell = (size_t)dim * dim;
  const double len[8] = { GAO_ASTAR_RES,           GAO_ASTAR_RES,           GAO_ASTAR_RES,
                          GAO_ASTAR_RES,           GAO_ASTAR_RES * sqrt(2), GAO_ASTAR_RES * sqrt(2),
                          GAO_ASTAR_RES * sqrt(2), GAO_ASTAR_RES * sqrt(2) };
  /* open-set node payloads (dense instead of unordered_map<size_t, NodeDisc>) */
  uint8_t* state = (uint8_t*)calloc(ncell, 1); /* 0 unseen, 1 open, 2 closed */
  double* o_cost = (double*)malloc(ncell * sizeof(double));
  double* o_dist = (double*)malloc(ncell * sizeof(double));
  int32_t* o_par = (int32_t*)malloc(ncell * sizeof(int32_t));
  uint8_t* o_unk = (uint8_t*)malloc(ncell);
  min_heap heap = { NULL, 0, 0 };

  for (size_t i = 0; i < ncell; ++i)
  {
    cost[i] = INFINITY;
    cost_dist[i] = INFINITY;
    parent[i] = -2;
    flags[i] = 0;
  }
  if (n_near) *n_near = 0;

  const size_t goal_id = (size_t)goal_y * dim + goal_x;
  const size_t start_id = (size_t)start_y * dim + start_x;
  state[goal_id] = 1;
  o_cost[goal_id] = 0.0;
  o_dist[goal_id] = 0.0;
  o_par[goal_id] = -1;
  o_unk[goal_id] = (uint8_t)(astar_grid[goal_id] == GAO_UNKNOWN);
  heap_push(&heap, 0.0, goal_id);

  int start_found = (start_id == goal_id);
  unsigned int nr_extra = 0;
  int is_near = 0;
  size_t n_closed = 0;
  const size_t limit = (size_t)round(0.95 * ((double)dim * dim));

  for (;;)
  {
    if (n_closed > limit) break;
    if (heap.n == 0) break;
    const size_t c_id = heap_pop(&heap).id;
    if (state[c_id] != 1) continue;

    state[c_id] = 2;
    const double c_cost = o_cost[c_id];
    const double c_dist = o_dist[c_id];
    cost[c_id] = c_cost;
    cost_dist[c_id] = c_dist;
    parent[c_id] = o_par[c_id];
    flags[c_id] = (uint8_t)(GAO_H_CLOSED | (o_unk[c_id] ? GAO_H_UNKNOWN : 0));
    if (closed_order) closed_order[n_closed] = (int32_t)c_id;
    ++n_closed;

    const int cx = (int)(c_id % (size_t)dim);
    const int cy = (int)(c_id / (size_t)dim);

    if (p->get_only_near)
    {
      const double xd = cx - start_x;
      const double yd = cy - start_y;
      const double d = sqrt(xd * xd + yd * yd);
      if (d < 10)
      {
        if (near_ids && n_near)
        {
          near_ids[*n_near] = (int32_t)c_id;
          near_dist[*n_near] = d;
          ++*n_near;
        }
        is_near = 1;
      }
    }
    if (p->heuristic_early_exit)
    {
      const int diff = abs((int)c_id - (int)start_id);
      if (diff == 0)
        start_found = 1;
      else if (diff == 1 || diff == dim)
        is_near = 1;
    }
    if (p->heuristic_early_exit || p->get_only_near)
    {
      if (start_found || is_near)
      {
        if (nr_extra > p->max_extra_nodes) continue;
        ++nr_extra;
      }
    }

    /* per-current terms, a_star.cpp:251-254,260 */
    const double prox_cost = h_prox[c_id] * p->astar_prox_cost;
    const int is_unknown = (astar_grid[c_id] == GAO_UNKNOWN);
    const double unknown_cost = is_unknown ? p->unknown_cost_w : 0;
    const double xd = cx - start_x;
    const double yd = cy - start_y;
    const double h_euclid = p->astar_lane_movement_cost * fabs(sqrt(xd * xd + yd * yd)) * GAO_ASTAR_RES;

    for (int i = 0; i < 8; ++i)
    {
      const int nx = cx + MOT_DX[i];
      const int ny = cy + MOT_DY[i];
      if (!verify_node(dim, astar_grid, geofence, nx, ny)) continue;
      const size_t n_id = (size_t)ny * dim + nx;
      if (state[n_id] == 2) continue;

      const double movement_costs = len[i] * movement_cost[n_id];
      const double nd = c_dist + len[i];
      const double ncost = c_cost + movement_costs + prox_cost + unknown_cost;
      const double est = h_euclid + ncost;

      if (state[n_id] == 0 || o_cost[n_id] > ncost)
      {
        state[n_id] = 1;
        o_cost[n_id] = ncost;
        o_dist[n_id] = nd;
        o_par[n_id] = (int32_t)c_id;
        o_unk[n_id] = (uint8_t)is_unknown;
        heap_push(&heap, est, n_id);
      }
    }
  }

  free(heap.a);
  free(state);
  free(o_cost);
  free(o_dist);
  free(o_par);
  free(o_unk);
  return (int)n_closed;
}

int gao_heuristic_fixedpoint(const gao_heur_params* p, const uint8_t* astar_grid,
                             const uint8_t* geofence, const double* movement_cost,
                             const double* h_prox, int goal_x, int goal_y, double* cost,
                             double* cost_dist, int32_t* parent, uint8_t* flags)
{
  const int dim = p->astar_dim;
  const size_t ncell = (size_t)dim * dim;
  const double L1 = GAO_ASTAR_RES;
  const double L2 = GAO_ASTAR_RES * sqrt(2);
  const double len[8] = { L1, L1, L1, L1, L2, L2, L2, L2 };
  uint8_t* done = (uint8_t*)calloc(ncell, 1);
  min_heap heap = { NULL, 0, 0 };

  for (size_t i = 0; i < ncell; ++i)
  {
    cost[i] = INFINITY;
    cost_dist[i] = INFINITY;
    parent[i] = -2;
    flags[i] = 0;
  }
  const size_t goal_id = (size_t)goal_y * dim + goal_x;
  cost[goal_id] = 0.0;
  heap_push(&heap, 0.0, goal_id);

  /* Dijkstra: all edge terms are >= 0 and fl(+) is monotone, so the settled values are the
   * least fixed point of the relaxation operator. */
  int n_reached = 0;
  while (heap.n)
  {
    const heap_el e = heap_pop(&heap);
    const size_t c = e.id;
    if (done[c] || e.prio > cost[c]) continue;
    done[c] = 1;
    ++n_reached;
    const int cx = (int)(c % (size_t)dim);
    const int cy = (int)(c / (size_t)dim);
    const double prox_cost = h_prox[c] * p->astar_prox_cost;
    const double unknown_cost = (astar_grid[c] == GAO_UNKNOWN) ? p->unknown_cost_w : 0;
    for (int i = 0; i < 8; ++i)
    {
      const int nx = cx + MOT_DX[i];
      const int ny = cy + MOT_DY[i];
      if (!verify_node(dim, astar_grid, geofence, nx, ny)) continue;
      const size_t n = (size_t)ny * dim + nx;
      if (n == goal_id) continue; /* the root keeps cost 0 / parent -1 */
      const double nc = cost[c] + len[i] * movement_cost[n] + prox_cost + unknown_cost;
      if (nc < cost[n])
      {
        cost[n] = nc;
        heap_push(&heap, nc, n);
      }
    }
  }

  /* canonical parent: smallest motion index i with n = p + motion[i] whose candidate
   * reproduces cost[n] bit for bit */
  for (size_t n = 0; n < ncell; ++n)
  {
    if (!done[n]) continue;
    flags[n] = GAO_H_CLOSED;
    if (n == goal_id)
    {
      parent[n] = -1;
      if (astar_grid[n] == GAO_UNKNOWN) flags[n] |= GAO_H_UNKNOWN;
      continue;
    }
    const int nx = (int)(n % (size_t)dim);
    const int ny = (int)(n / (size_t)dim);
    for (int i = 0; i < 8; ++i)
    {
      const int px = nx - MOT_DX[i];
      const int py = ny - MOT_DY[i];
      if (px < 0 || py < 0 || px >= dim || py >= dim) continue;
      const size_t q = (size_t)py * dim + px;
      if (!done[q]) continue;
      const double prox_cost = h_prox[q] * p->astar_prox_cost;
      const double unknown_cost = (astar_grid[q] == GAO_UNKNOWN) ? p->unknown_cost_w : 0;
      const double nc = cost[q] + len[i] * movement_cost[n] + prox_cost + unknown_cost;
      if (nc == cost[n])
      {
        parent[n] = (int32_t)q;
        if (astar_grid[q] == GAO_UNKNOWN) flags[n] |= GAO_H_UNKNOWN;
        break;
      }
    }
  }

  /* cost_dist from step counts along the chain (order-free definition) */
  int32_t* stack = (int32_t*)malloc(ncell * sizeof(int32_t));
  int32_t* ns = (int32_t*)calloc(ncell, sizeof(int32_t));
  int32_t* nd = (int32_t*)calloc(ncell, sizeof(int32_t));
  uint8_t* have = (uint8_t*)calloc(ncell, 1);
  have[goal_id] = 1;
  for (size_t n = 0; n < ncell; ++n)
  {
    if (!done[n] || have[n]) continue;
    size_t sp = 0;
    size_t q = n;
    while (!have[q])
    {
      stack[sp++] = (int32_t)q;
      q = (size_t)parent[q];
    }
    while (sp)
    {
      const size_t c = (size_t)stack[--sp];
      const size_t par = (size_t)parent[c];
      const int diag = (c % (size_t)dim != par % (size_t)dim) && (c / (size_t)dim != par / (size_t)dim);
      ns[c] = ns[par] + (diag ? 0 : 1);
      nd[c] = nd[par] + (diag ? 1 : 0);
      have[c] = 1;
    }
  }
  for (size_t n = 0; n < ncell; ++n)
  {
    if (done[n]) cost_dist[n] = (double)ns[n] * L1 + (double)nd[n] * L2;
  }
  free(stack);
  free(ns);
  free(nd);
  free(have);
  free(done);
  free(heap.a);
  return n_reached;
}

int gao_astar_path(int astar_dim, const int32_t* parent, const uint8_t* flags, int x, int y,
                   int* xs, int* ys, int cap)
{
  const long ncell = (long)astar_dim * astar_dim;
  long idx = (long)y * astar_dim + x;
#define IN_SET(i) ((i) >= 0 && (i) < ncell && (flags[(i)] & GAO_H_CLOSED))
  if (!IN_SET(idx))
  {
    /* findValidNeighborIndex: right, left, above (+dim), below (-dim) */
    long b = idx + 1;
    if (!IN_SET(b))
    {
      b = idx - 1;
      if (!IN_SET(b))
      {
        b = idx + astar_dim;
        if (!IN_SET(b))
        {
          b = idx - astar_dim;
          if (!IN_SET(b)) return 0;
        }
      }
    }
    idx = b;
  }
#undef IN_SET
  int n = 0;
  while (idx != -1 && n < cap)
  {
    xs[n] = (int)(idx % astar_dim);
    ys[n] = (int)(idx / astar_dim);
    ++n;
    idx = parent[idx];
  }
  return n;
}

/* ========================================================================================
 * distance transform + potential field
 * ====================================================================================== */

static void dt_fixedpoint(const uint8_t* src, int stride, int w, int h, float* dist)
{
  /* OpenCV 4.8.0 modules/imgproc/src/distransform.cpp distanceTransform_5x5 with the DIST_L2
   * 5x5 metrics {1, 1.4, 2.1969} in 16.16 fixed point */
  const int B = 2;
  const unsigned int HV = 65536u, DG = 91750u, LG = 143976u; /* cvRound(m * 65536) */
  const unsigned int INIT = 0x7fffffffu >> 2;               /* INIT_DIST0 == DIST_MAX */
  const float scale = 1.f / 65536.f;
  const int step = w + 2 * B;
  unsigned int* tmp = (unsigned int*)malloc(sizeof(unsigned int) * (size_t)step * (h + 2 * B));
  for (size_t i = 0; i < (size_t)step * (h + 2 * B); ++i) tmp[i] = INIT;
#define T(y, x) tmp[(size_t)((y) + B) * step + (x) + B]
  for (int i = 0; i < h; ++i)
  {
    const uint8_t* s = src + (size_t)i * stride;
    for (int j = 0; j < w; ++j)
    {
      if (!s[j])
      {
        T(i, j) = 0;
      }
      else
      {
        unsigned int t0 = T(i - 2, j - 1) + LG, t;
        t = T(i - 2, j + 1) + LG; if (t0 > t) t0 = t;
        t = T(i - 1, j - 2) + LG; if (t0 > t) t0 = t;
        t = T(i - 1, j - 1) + DG; if (t0 > t) t0 = t;
        t = T(i - 1, j) + HV;     if (t0 > t) t0 = t;
        t = T(i - 1, j + 1) + DG; if (t0 > t) t0 = t;
        t = T(i - 1, j + 2) + LG; if (t0 > t) t0 = t;
        t = T(i, j - 1) + HV;     if (t0 > t) t0 = t;
        T(i, j) = t0;
      }
    }
  }
  for (int i = h - 1; i >= 0; --i)
  {
    for (int j = w - 1; j >= 0; --j)
    {
      unsigned int t0 = T(i, j), t;
      if (t0 > HV)
      {
        t = T(i + 2, j + 1) + LG; if (t0 > t) t0 = t;
        t = T(i + 2, j - 1) + LG; if (t0 > t) t0 = t;
        t = T(i + 1, j + 2) + LG; if (t0 > t) t0 = t;
        t = T(i + 1, j + 1) + DG; if (t0 > t) t0 = t;
        t = T(i + 1, j) + HV;     if (t0 > t) t0 = t;
        t = T(i + 1, j - 1) + DG; if (t0 > t) t0 = t;
        t = T(i + 1, j - 2) + LG; if (t0 > t) t0 = t;
        t = T(i, j + 1) + HV;     if (t0 > t) t0 = t;
        T(i, j) = t0;
      }
      t0 = (t0 > INIT) ? INIT : t0;
      dist[(size_t)i * w + j] = (float)(t0 * scale);
    }
  }
#undef T
  free(tmp);
}

static void dt_ippfloat(const uint8_t* src, int stride, int w, int h, float* dist)
{
  /* ippiDistanceTransform_5x5_8u32f_C1R as dispatched by cv::distanceTransform when OpenCV
   * is built WITH_IPP: float32 two-pass raster chamfer, same neighbourhoods as above, with
   * out-of-image neighbours ignored.  (Closed-source; behaviour pinned against cv2 4.13.0
   * goldens, tests/golden/dt_*.npz.) */
  const float HV = 1.0f, DG = 1.4f, LG = 2.1969f;
  const float BIG = 3.4e38f;
#define D(y, x) (((y) < 0 || (y) >= h || (x) < 0 || (x) >= w) ? BIG : dist[(size_t)(y) * w + (x)])
  for (int i = 0; i < h; ++i)
  {
    const uint8_t* s = src + (size_t)i * stride;
    for (int j = 0; j < w; ++j)
    {
      if (!s[j])
      {
        dist[(size_t)i * w + j] = 0.f;
      }
      else
      {
        float t0 = BIG, t;
        t = D(i - 2, j - 1) + LG; if (t < t0) t0 = t;
        t = D(i - 2, j + 1) + LG; if (t < t0) t0 = t;
        t = D(i - 1, j - 2) + LG; if (t < t0) t0 = t;
        t = D(i - 1, j - 1) + DG; if (t < t0) t0 = t;
        t = D(i - 1, j) + HV;     if (t < t0) t0 = t;
        t = D(i - 1, j + 1) + DG; if (t < t0) t0 = t;
        t = D(i - 1, j + 2) + LG; if (t < t0) t0 = t;
        t = D(i, j - 1) + HV;     if (t < t0) t0 = t;
        dist[(size_t)i * w + j] = t0;
      }
    }
  }
  for (int i = h - 1; i >= 0; --i)
  {
    for (int j = w - 1; j >= 0; --j)
    {
      float t0 = dist[(size_t)i * w + j], t;
      t = D(i + 2, j + 1) + LG; if (t < t0) t0 = t;
      t = D(i + 2, j - 1) + LG; if (t < t0) t0 = t;
      t = D(i + 1, j + 2) + LG; if (t < t0) t0 = t;
      t = D(i + 1, j + 1) + DG; if (t < t0) t0 = t;
      t = D(i + 1, j) + HV;     if (t < t0) t0 = t;
      t = D(i + 1, j - 1) + DG; if (t < t0) t0 = t;
      t = D(i + 1, j - 2) + LG; if (t < t0) t0 = t;
      t = D(i, j + 1) + HV;     if (t < t0) t0 = t;
      dist[(size_t)i * w + j] = t0;
    }
  }
#undef D
}

void gao_distance_transform(const uint8_t* src, int stride, int w, int h, float* dist,
                            int variant)
{
  if (variant == GAO_DT_IPPFLOAT)
    dt_ippfloat(src, stride, w, h, dist);
  else
    dt_fixedpoint(src, stride, w, h, dist);
}

void gao_costmap_value(const gao_pot_params* p, float d_obs, float d_vor, double* h_prox,
                       double* motion_res)
{
  /* a_star.cpp:492-507.  float operands are promoted to double by the mixed expressions;
   * (obstacle_dist + voronoi_dist) is a float+float sum evaluated in float. */
  if (d_obs < p->do_max)
  {
    *h_prox = (p->alpha / (p->alpha + d_obs)) * (d_vor / (d_obs + d_vor)) *
              pow((d_obs - p->do_max), 2) / pow(p->do_max, 2);
  }
  else
  {
    *h_prox = 0.0;
  }
  const float lo = (float)p->dist_val_min;
  const float hi = (float)p->dist_val_max;
  const float cl = d_obs < lo ? lo : (hi < d_obs ? hi : d_obs); /* std::clamp */
  const double input = cl;
  const double mr =
      p->motion_res_min + ((p->motion_res_max - p->motion_res_min) / (p->dist_val_max - p->dist_val_min)) *
                              (input - p->dist_val_min);
  *motion_res = mr > p->motion_res_min ? mr : p->motion_res_min; /* std::max(mr, min) */
}

int gao_voronoi_potential_field(const gao_pot_params* p, const uint8_t* astar_grid,
                                int astar_dim, int ego_x, int ego_y, const int* vor_xy,
                                int n_vor, float* obstacle_dist, float* voronoi_dist,
                                double* h_prox, double* motion_res)
{
  const int V = p->vor_dim;
  const int off = V / 2;
  const int ox = ego_x - off > 0 ? ego_x - off : 0; /* getCurrentMapOrigin :527-531 */
  const int oy = ego_y - off > 0 ? ego_y - off : 0;
  uint8_t* mask = (uint8_t*)malloc((size_t)V * V);
  const float scale = (float)GAO_ASTAR_RES;

  /* DT#1: 1 - (OCC == window) */
  int n_occ = 0;
  for (int y = 0; y < V; ++y)
  {
    for (int x = 0; x < V; ++x)
    {
      const int occ = astar_grid[(size_t)(y + oy) * astar_dim + (x + ox)] == GAO_OCC;
      mask[(size_t)y * V + x] = (uint8_t)(occ ? 0 : 1);
      n_occ += occ;
    }
  }
  gao_distance_transform(mask, V, V, V, obstacle_dist, p->dt_variant);
  /* dist_map *= star2con_: float Mat times double scalar, per-element float multiply */
  for (int i = 0; i < V * V; ++i) obstacle_dist[i] = obstacle_dist[i] * scale;

  if (n_occ == 0)
  {
    free(mask);
    return 0;
  }

  /* endpoints marked with OCC (=2) into the persistent float window, then DT in place */
  for (int i = 0; i < n_vor; ++i)
  {
    const int x = vor_xy[2 * i], y = vor_xy[2 * i + 1];
    if (x >= 0 && y >= 0 && x < V && y < V) voronoi_dist[(size_t)y * V + x] = (float)GAO_OCC;
  }
  for (int i = 0; i < V * V; ++i) mask[i] = (uint8_t)(voronoi_dist[i] == 2.0f ? 0 : 1);
  gao_distance_transform(mask, V, V, V, voronoi_dist, p->dt_variant);
  for (int i = 0; i < V * V; ++i) voronoi_dist[i] = voronoi_dist[i] * scale;

  for (int y = 0; y < V; ++y)
  {
    for (int x = 0; x < V; ++x)
    {
      const size_t g = (size_t)(y + oy) * astar_dim + (x + ox);
      gao_costmap_value(p, obstacle_dist[(size_t)y * V + x], voronoi_dist[(size_t)y * V + x],
                        &h_prox[g], &motion_res[g]);
    }
  }
  free(mask);
  return 1;
}

/* ---------------------------------------------------------------------------------------------
 * proximity scoring of sampled poses
 * ------------------------------------------------------------------------------------------- */

/* libstdc++ 13 <cmath> __lerp (what std::lerp in util1.cpp:179,185 resolves to) */
static double gao_lerp(double a, double b, double t)
{
  if ((a <= 0 && b >= 0) || (a >= 0 && b <= 0)) return t * b + (1 - t) * a;
  if (t == 1) return b;
  {
    const double x = a + t * (b - a);
    return (t > 1) == (b > a) ? (b < x ? x : b) : (b > x ? x : b);
  }
}

/* util::bilinearInterpolation, util1.cpp:152-200 */
double gao_bilinear(double q11, double q12, double q21, double q22, int x1, int x2, int y1, int y2, double x, double y)
{
  const int x2x1 = x2 - x1, y2y1 = y2 - y1;
  const double yy1 = y - y1, xx1 = x - x1;
  double result;
  if (abs(x2x1) == 0 && abs(y2y1) < 0.0001) result = q11;
  else if (abs(x2x1) == 0) result = gao_lerp(q11, q12, yy1 / y2y1);
  else if (abs(y2y1) < 0.0001) result = gao_lerp(q11, q21, xx1 / x2x1);
  else
  {
    const double x2x = (double)x2 - x, y2y = y2 - y;
    result = 1.0 / (x2x1 * y2y1) * (q11 * x2x * y2y + q21 * xx1 * y2y + q12 * x2x * yy1 + q22 * xx1 * yy1);
  }
  if (isnan(result)) result = 0.0;
  return result;
}

/* util::getBilinInterp, util2.cpp:41-52.  The reference clamps the indices at 0 only; a point
 * beyond the last row / column reads past its array -- here (as in the CUDA path) the indices
 * are clamped into the grid, which is the same for every point the reference may legally ask. */
double gao_bilin_at(double x, double y, const double* grid, int dim)
{
  const int hi = dim - 1;
  int x1 = (int)x, x2 = (int)ceil(x), y1 = (int)y, y2 = (int)ceil(y);
  x1 = x1 < 0 ? 0 : x1 > hi ? hi : x1;
  x2 = x2 < 0 ? 0 : x2 > hi ? hi : x2;
  y1 = y1 < 0 ? 0 : y1 > hi ? hi : y1;
  y2 = y2 < 0 ? 0 : y2 > hi ? hi : y2;
  return gao_bilinear(grid[(size_t)y1 * dim + x1], grid[(size_t)y2 * dim + x1], grid[(size_t)y1 * dim + x2],
                      grid[(size_t)y2 * dim + x2], x1, x2, y1, y2, x, y);
}

/* HybridAStar::getProxOfCorners(pose * con2star_), hybrid_a_star_lib.cpp:265-289: the corners
 * (vehicle.cpp:24-27, METRES) are rotated and added to the pose in ASTAR CELLS, as the reference does */
double gao_prox_of_corners(const double* grid, int dim, const gao_prox_params* p, double x_m, double y_m, double yaw)
{
  const double cos_yaw = cos(yaw), sin_yaw = sin(yaw);
  const double ox = x_m * p->con2star, oy = y_m * p->con2star;
  const double cx[4] = { p->lf, p->lf, -p->lb, -p->lb };
  const double cy[4] = { -p->width / 2, p->width / 2, -p->width / 2, p->width / 2 };
  double best = 0;
  int k;
  for (k = 0; k < 4; ++k)
  {
    const double x = (cx[k] * cos_yaw - cy[k] * sin_yaw) + ox, y = (cx[k] * sin_yaw + cy[k] * cos_yaw) + oy;
    const double v = gao_bilin_at(x, y, grid, dim);
    if (k == 0 || best < v) best = v;
  }
  return isnan(best) ? 0.0 : best;
}

/* the prox_cost loops of getRAPathCosts / getRSPathCosts, hybrid_a_star_lib.cpp:588-594, 670-676 */
void gao_path_prox_costs(const double* grid, int dim, const gao_prox_params* p, const double* poses, const int* offsets,
                         int n_paths, double* out)
{
  int i, j;
  for (i = 0; i < n_paths; ++i)
  {
    double prox_cost = 0;
    for (j = offsets[i]; j < offsets[i + 1]; ++j)
      prox_cost += gao_prox_of_corners(grid, dim, p, poses[3 * j], poses[3 * j + 1], poses[3 * j + 2]) * p->h_prox_cost * p->interp_res;
    out[i] = prox_cost;
  }
}

/* step 3 of HybridAStar::getValidClosePose, hybrid_a_star_lib.cpp:1604-1668.  member = astar_dim^2
 * bytes, 1 where the cell is one of AStar::nodes_near_goal_.  Returns 1 and the pose, or 0 (the
 * reference then keeps the ego pose). */
int gao_valid_close_pose(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks, int yaw_res_deg,
                         const double* prox, int astar_dim, const gao_prox_params* p, const uint8_t* member,
                         const double* center, double* out_pose)
{
  const double angle_weight = 20.0, dist_weight = 1.0, prox_weight = 1.0;
  const double dxy_max = 5.0, dphi_max = 3.14159265358979323846;
  const double gm_res = 0.15625;                                      /* gm_res_ = GM_RES, :34 */
  const double phi_res = yaw_res_deg * (3.14159265358979323846 / 180); /* yaw_res_coll_ * util::TO_RAD */
  double min_cost = 1.7976931348623157e308;
  int found_any = 0, dxi, dyi, dphii;
  const int dxy_max_idx = (int)(floor(dxy_max) / gm_res);
  const int dphi_max_idx = (int)(floor(dphi_max) / phi_res);
  for (dxi = -dxy_max_idx; dxi < dxy_max_idx; ++dxi)
    for (dyi = -dxy_max_idx; dyi < dxy_max_idx; ++dyi)
      for (dphii = -dphi_max_idx; dphii < dphi_max_idx; ++dphii)
      {
        const double x_diff = dxi * gm_res, y_diff = dyi * gm_res, phi_diff = dphii * phi_res;
        const double dist2 = x_diff * x_diff + y_diff * y_diff;
        const double px = center[0] + x_diff, py = center[1] + y_diff, pyaw = center[2] + phi_diff;
        const double prox_cost = gao_prox_of_corners(prox, astar_dim, p, px, py, pyaw);
        const double cost = dist2 * dist_weight + fabs(phi_diff) * angle_weight + prox_cost * prox_weight;
        long long ix, iy, index;
        if (cost > min_cost) continue;
        ix = (long long)round(px * p->con2star); /* Point<double>::toInt = std::round */
        iy = (long long)round(py * p->con2star);
        index = iy * astar_dim + ix;               /* calcIndex(size_t, size_t): wraps like this sum */
        if (index < 0 || index >= (long long)astar_dim * astar_dim || !member[index]) continue;
        if (gao_check_pose(safety, patch_dim, centers, nb_disks, yaw_res_deg, px, py, pyaw, 0))
        {
          out_pose[0] = px;
          out_pose[1] = py;
          out_pose[2] = pyaw;
          min_cost = cost;
          found_any = 1;
        }
      }
  return found_any;
}
