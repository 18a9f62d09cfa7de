// amcl_node_probe.cpp -- a separately compiled C++ caller that uses the AMCL
// libraries the way amcl_node.cpp does (amcl_node.cpp:896-958 map + filter +
// sensors, :1143 copy-constructed laser, :1257-1329 odometry and laser updates,
// :1342-1361 reading set->samples[] through the struct) and prints what it sees.
//
// The same source is built twice: here, on the test box, against include/ and
// libamcl_b200.so; and in the build container against the reference's own
// headers and the compiled reference (oracle/Makefile -> oracle/_ref/).  It uses
// nothing outside the reference's API and no random numbers: poses come from a
// callback, the odometry noise parameters are zero.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "amcl/map/map.h"
#include "amcl/pf/pf.h"
#include "amcl/sensors/amcl_laser.h"
#include "amcl/sensors/amcl_odom.h"

using namespace amcl;

struct PoseSource
{
  unsigned long long state;
  double half_x, half_y;
};

// pf_init_model_fn_t: deterministic poses inside the walls
static pf_vector_t next_pose(void *arg)
{
  PoseSource *s = (PoseSource *)arg;
  double u[3];
  for (int k = 0; k < 3; k++)
  {
    s->state = s->state * 6364136223846793005ull + 1442695040888963407ull;
    u[k] = (double)(s->state >> 11) / 9007199254740992.0;
  }
  pf_vector_t p;
  p.v[0] = (2.0 * u[0] - 1.0) * (s->half_x - 0.3);
  p.v[1] = (2.0 * u[1] - 1.0) * (s->half_y - 0.3);
  p.v[2] = (2.0 * u[2] - 1.0) * M_PI;
  return p;
}

static void dump(const char *tag, pf_t *pf)
{
  pf_sample_set_t *set = pf->sets + pf->current_set;
  printf("%s %d\n", tag, set->sample_count);
  for (int i = 0; i < set->sample_count; i++)
    printf("%.17g %.17g %.17g %.17g\n", set->samples[i].pose.v[0], set->samples[i].pose.v[1],
           set->samples[i].pose.v[2], set->samples[i].weight);
}

int main()
{
  // AmclNode::convertMap (amcl_node.cpp:987-1012): a walled box with a cross inside
  const int sx = 240, sy = 200;
  map_t *map = map_alloc();
  map->size_x = sx;
  map->size_y = sy;
  map->scale = 0.05;
  map->origin_x = 0.0 + (sx / 2) * map->scale;
  map->origin_y = 0.0 + (sy / 2) * map->scale;
  map->cells = (map_cell_t *)malloc(sizeof(map_cell_t) * sx * sy);
  for (int j = 0; j < sy; j++)
    for (int i = 0; i < sx; i++)
    {
      bool wall = i == 0 || j == 0 || i == sx - 1 || j == sy - 1 || (j == 100 && i >= 60 && i < 180) ||
                  (i == 120 && j >= 60 && j < 140);
      map->cells[MAP_INDEX(map, i, j)].occ_state = wall ? +1 : -1;
      map->cells[MAP_INDEX(map, i, j)].occ_dist = 0.0;
    }

  PoseSource src;
  src.state = 12345;
  src.half_x = sx * map->scale / 2;
  src.half_y = sy * map->scale / 2;
  // the callback's poses are relative to the map centre
  struct Shifted
  {
    static pf_vector_t fn(void *arg)
    {
      PoseSource *s = (PoseSource *)arg;
      pf_vector_t p = next_pose(arg);
      p.v[0] += s->half_x;
      p.v[1] += s->half_y;
      return p;
    }
  };
  pf_t *pf = pf_alloc(100, 2000, 0.0, 0.0, (pf_init_model_fn_t)Shifted::fn, (void *)&src);
  pf->pop_err = 0.01;
  pf->pop_z = 0.99;
  pf_init_model(pf, (pf_init_model_fn_t)Shifted::fn, (void *)&src);
  dump("init", pf);
  {
    double w;
    pf_vector_t mean;
    pf_matrix_t cov;
    int k = 0;
    while (pf_get_cluster_stats(pf, k, &w, &mean, &cov)) k++;
    printf("clusters %d\n", k);
  }

  AMCLOdom *odom = new AMCLOdom();
  odom->SetModel(ODOM_MODEL_DIFF, 0.0, 0.0, 0.0, 0.0, 0.0);
  AMCLLaser *laser = new AMCLLaser(60, map);
  laser->SetModelLikelihoodField(0.95, 0.05, 0.2, 2.0);
  AMCLLaser *copy = new AMCLLaser(*laser);      // amcl_node.cpp:1143
  pf_vector_t lp;
  lp.v[0] = 0.1; lp.v[1] = 0.0; lp.v[2] = 0.0;
  copy->SetLaserPose(lp);

  // a 60-beam scan from a pose in the lower-left room
  const int nb = 60;
  const double range_max = 8.0;
  AMCLLaserData ldata;
  ldata.sensor = copy;
  ldata.range_count = nb;
  ldata.range_max = range_max;
  ldata.ranges = new double[nb][2];
  for (int i = 0; i < nb; i++)
  {
    const double bearing = -2.0 + 4.0 * i / (nb - 1);
    double r = map_calc_range(map, 3.0, 2.5, 0.4 + bearing, range_max);
    if (r > range_max) r = range_max;
    ldata.ranges[i][0] = r;
    ldata.ranges[i][1] = bearing;
  }

  AMCLOdomData odata;
  odata.pose.v[0] = 0.1; odata.pose.v[1] = 0.05; odata.pose.v[2] = 0.02;
  odata.delta = odata.pose;
  bool ok = odom->UpdateAction(pf, (AMCLSensorData *)&odata);
  dump("action", pf);
  ok = copy->UpdateSensor(pf, (AMCLSensorData *)&ldata) && ok;
  dump("lf", pf);
  laser->SetModelBeam(0.95, 0.1, 0.05, 0.05, 0.2, 0.1, 0.0);
  laser->SetLaserPose(lp);                      // the constructor leaves laser_pose unset
  ldata.sensor = laser;
  ok = laser->UpdateSensor(pf, (AMCLSensorData *)&ldata) && ok;
  dump("beam", pf);
  printf("ok %d occ_dist %.17g %.17g\n", ok ? 1 : 0, map->cells[MAP_INDEX(map, 30, 40)].occ_dist,
         map->cells[MAP_INDEX(map, 119, 99)].occ_dist);

  delete copy;
  delete laser;
  delete odom;
  pf_free(pf);
  map_free(map);
  return ok ? 0 : 1;
}
