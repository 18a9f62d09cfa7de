// class_capi.cpp -- plain-C facade over the amcl:: C++ classes and the few
// caller-side helpers amcl_node.cpp implements itself (map conversion, the
// uniform free-space pose generator).  FFI callers (ctypes in tests/bench)
// cannot construct C++ objects, so every method gets one extern "C" twin.
//
// This file is written ONLY against the public headers "amcl/..." and is
// compiled twice: into libamcl_b200.so against include/ of this repo, and
// (by oracle/Makefile, when /root/reference is mounted) into
// oracle/_ref/libamcl_ref.so against the reference's own headers and
// objects.  That it compiles unchanged against both header sets is the
// source-compatibility check of the drop-in boundary.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include <utility>

#include "amcl/map/map.h"
#include "amcl/pf/pf.h"
#include "amcl/sensors/amcl_odom.h"
#include "amcl/sensors/amcl_laser.h"

using namespace amcl;

extern "C" {

// ---- maps -----------------------------------------------------------------

// What AmclNode::convertMap does (amcl_node.cpp:987-1012), minus the ROS
// message: occ[] holds -1 / 0 / +1 per cell, row-major with x fastest;
// (origin_x, origin_y) is the world position of the map CENTRE.
map_t *amcl_c_map_from_occupancy(const signed char *occ, int size_x, int size_y,
                                 double scale, double origin_x, double origin_y)
{
  map_t *map = map_alloc();
  map->size_x = size_x;
  map->size_y = size_y;
  map->scale = scale;
  map->origin_x = origin_x;
  map->origin_y = origin_y;
  size_t n = (size_t)size_x * (size_t)size_y;
  map->cells = (map_cell_t *)malloc(sizeof(map_cell_t) * n);
  for (size_t k = 0; k < n; k++)
  {
    map->cells[k].occ_state = occ[k];
    map->cells[k].occ_dist = 0.0;
  }
  return map;
}

void amcl_c_map_get_occ_dist(map_t *map, double *out)
{
  size_t n = (size_t)map->size_x * (size_t)map->size_y;
  for (size_t k = 0; k < n; k++)
    out[k] = map->cells[k].occ_dist;
}

void amcl_c_map_set_occ_dist(map_t *map, const double *in, double max_occ_dist)
{
  size_t n = (size_t)map->size_x * (size_t)map->size_y;
  for (size_t k = 0; k < n; k++)
    map->cells[k].occ_dist = in[k];
  map->max_occ_dist = max_occ_dist;
}

// ---- uniform free-space pose generator --------------------------------------
// Restates AmclNode::uniformPoseGenerator, NEW_UNIFORM_SAMPLING branch
// (amcl_node.cpp:1050-1060) and its free-space index (amcl_node.cpp:901-908).
// The index is a process-wide static there as well.

static std::vector<std::pair<int, int> > g_free_cells;

int amcl_c_freespace_build(map_t *map)
{
  g_free_cells.clear();
  for (int i = 0; i < map->size_x; i++)
    for (int j = 0; j < map->size_y; j++)
      if (map->cells[MAP_INDEX(map, i, j)].occ_state == -1)
        g_free_cells.push_back(std::make_pair(i, j));
  return (int)g_free_cells.size();
}

pf_vector_t amcl_c_uniform_pose(void *arg)
{
  map_t *map = (map_t *)arg;
  unsigned int pick = drand48() * g_free_cells.size();
  std::pair<int, int> cell = g_free_cells[pick];
  pf_vector_t p;
  p.v[0] = MAP_WXGX(map, cell.first);
  p.v[1] = MAP_WYGY(map, cell.second);
  p.v[2] = drand48() * 2 * M_PI - M_PI;
  return p;
}

// Address of the generator, for pf_alloc / pf_init_model through an FFI.
pf_init_model_fn_t amcl_c_uniform_pose_fn(void) { return amcl_c_uniform_pose; }

// ---- AMCLOdom ---------------------------------------------------------------

void *amcl_c_odom_new(void) { return new AMCLOdom(); }
void amcl_c_odom_delete(void *o) { delete (AMCLOdom *)o; }

void amcl_c_odom_set_model(void *o, int type, double a1, double a2, double a3,
                           double a4, double a5)
{
  ((AMCLOdom *)o)->SetModel((odom_model_t)type, a1, a2, a3, a4, a5);
}

// AMCLOdom::UpdateAction with an AMCLOdomData built from plain arrays.
int amcl_c_odom_update_action(void *o, pf_t *pf, const double pose[3],
                              const double delta[3])
{
  AMCLOdomData od;
  od.sensor = (AMCLOdom *)o;
  od.time = 0.0;
  for (int k = 0; k < 3; k++)
  {
    od.pose.v[k] = pose[k];
    od.delta.v[k] = delta[k];
  }
  return ((AMCLOdom *)o)->UpdateAction(pf, &od) ? 1 : 0;
}

// ---- AMCLLaser ----------------------------------------------------------------

void *amcl_c_laser_new(int max_beams, map_t *map)
{
  return new AMCLLaser((size_t)max_beams, map);
}
void amcl_c_laser_delete(void *l) { delete (AMCLLaser *)l; }

// Mirrors `new AMCLLaser(*laser_)` in amcl_node.cpp:1143.
void *amcl_c_laser_clone(void *l) { return new AMCLLaser(*(AMCLLaser *)l); }

void amcl_c_laser_set_pose(void *l, double x, double y, double a)
{
  pf_vector_t p;
  p.v[0] = x; p.v[1] = y; p.v[2] = a;
  ((AMCLLaser *)l)->SetLaserPose(p);
}

void amcl_c_laser_set_model_beam(void *l, double z_hit, double z_short,
                                 double z_max, double z_rand, double sigma_hit,
                                 double lambda_short, double chi_outlier)
{
  ((AMCLLaser *)l)->SetModelBeam(z_hit, z_short, z_max, z_rand, sigma_hit,
                                 lambda_short, chi_outlier);
}

void amcl_c_laser_set_model_lf(void *l, double z_hit, double z_rand,
                               double sigma_hit, double max_occ_dist)
{
  ((AMCLLaser *)l)->SetModelLikelihoodField(z_hit, z_rand, sigma_hit, max_occ_dist);
}

void amcl_c_laser_set_model_lfp(void *l, double z_hit, double z_rand,
                                double sigma_hit, double max_occ_dist,
                                int do_beamskip, double beam_skip_distance,
                                double beam_skip_threshold,
                                double beam_skip_error_threshold)
{
  ((AMCLLaser *)l)->SetModelLikelihoodFieldProb(
      z_hit, z_rand, sigma_hit, max_occ_dist, do_beamskip != 0,
      beam_skip_distance, beam_skip_threshold, beam_skip_error_threshold);
}

// AMCLLaser::UpdateSensor with an AMCLLaserData built the way the node does
// (amcl_node.cpp:1296-1327): ranges_rb holds range_count (range, bearing) pairs.
int amcl_c_laser_update_sensor(void *l, pf_t *pf, const double *ranges_rb,
                               int range_count, double range_max)
{
  AMCLLaserData ld;
  ld.sensor = (AMCLLaser *)l;
  ld.time = 0.0;
  ld.range_count = range_count;
  ld.range_max = range_max;
  ld.ranges = new double[range_count > 0 ? range_count : 1][2];
  memcpy(ld.ranges, ranges_rb, sizeof(double) * 2 * (size_t)range_count);
  bool ok = ((AMCLLaser *)l)->UpdateSensor(pf, &ld);
  return ok ? 1 : 0;  // ld's destructor releases ranges, as in the node
}

// ---- struct access helpers (so FFI callers need not mirror layouts) ---------

pf_sample_set_t *amcl_c_pf_current_set(pf_t *pf) { return pf->sets + pf->current_set; }
int amcl_c_sizeof_pf(void) { return (int)sizeof(pf_t); }
int amcl_c_sizeof_laser(void) { return (int)sizeof(AMCLLaser); }
int amcl_c_sizeof_odom(void) { return (int)sizeof(AMCLOdom); }

}  // extern "C"
