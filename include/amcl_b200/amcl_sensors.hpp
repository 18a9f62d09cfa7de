/*
 * amcl_sensors.hpp -- the C++ half of the drop-in boundary (namespace amcl).
 *
 * Class layouts and the virtual-function order are ABI: amcl_node.cpp does
 * `new AMCLLaser(max_beams, map)`, copy-constructs lasers and reads/writes
 * data objects in its own translation unit, so no member may be added,
 * removed or reordered.  Written from scratch; each class names the
 * reference declaration it stands in for (path:line under
 * /root/reference/amcl/include/amcl/sensors/).
 *
 * GPU state is NOT held in these objects.  It sits in registries keyed by
 * pf_t* / map_t* inside the library (see navigation_b200/csrc/context.hpp).
 */
#ifndef AMCL_B200_AMCL_SENSORS_HPP
#define AMCL_B200_AMCL_SENSORS_HPP

#include <stddef.h>
#include "amcl_abi.h"

namespace amcl
{

class AMCLSensorData;

/* amcl_sensor.h:41-80.  sizeof == 40: vptr, is_action, pose. */
class AMCLSensor
{
public:
  AMCLSensor();
  virtual ~AMCLSensor();

  /* Each returns true when the filter was touched.  Base versions: false. */
  virtual bool UpdateAction(pf_t *pf, AMCLSensorData *data);
  virtual bool InitSensor(pf_t *pf, AMCLSensorData *data);
  virtual bool UpdateSensor(pf_t *pf, AMCLSensorData *data);

  bool is_action;
  pf_vector_t pose;
};

/* amcl_sensor.h:85-94.  sizeof == 24: vptr, sensor, time. */
class AMCLSensorData
{
public:
  AMCLSensor *sensor;          /* the model instance this reading belongs to */
  virtual ~AMCLSensorData() {}
  double time;
};

/* ----------------------------------------------------------------------- */
/* Odometry                                           amcl_odom.h:36-93    */
/* ----------------------------------------------------------------------- */

typedef enum                                        /* amcl_odom.h:36-42 */
{
  ODOM_MODEL_DIFF,
  ODOM_MODEL_OMNI,
  ODOM_MODEL_DIFF_CORRECTED,
  ODOM_MODEL_OMNI_CORRECTED
} odom_model_t;

/* amcl_odom.h:45-53.  sizeof == 72. */
class AMCLOdomData : public AMCLSensorData
{
public:
  pf_vector_t pose;            /* odometric pose at this reading */
  pf_vector_t delta;           /* change since the previous filter update */
};

/* amcl_odom.h:57-93.  sizeof == 96. */
class AMCLOdom : public AMCLSensor
{
public:
  AMCLOdom();

  void SetModelDiff(double alpha1, double alpha2, double alpha3, double alpha4);
  void SetModelOmni(double alpha1, double alpha2, double alpha3, double alpha4,
                    double alpha5);
  void SetModel(odom_model_t type, double alpha1, double alpha2, double alpha3,
                double alpha4, double alpha5 = 0);

  /* sample_motion_model on every particle (device kernel). */
  virtual bool UpdateAction(pf_t *pf, AMCLSensorData *data);

private:
  double time;
  odom_model_t model_type;
  double alpha1, alpha2, alpha3, alpha4, alpha5;
};

/* ----------------------------------------------------------------------- */
/* Planar laser                                      amcl_laser.h:37-151   */
/* ----------------------------------------------------------------------- */

typedef enum                                        /* amcl_laser.h:37-42 */
{
  LASER_MODEL_BEAM,
  LASER_MODEL_LIKELIHOOD_FIELD,
  LASER_MODEL_LIKELIHOOD_FIELD_PROB
} laser_model_t;

/* amcl_laser.h:45-55.  sizeof == 48.  ranges is new[]'d by the caller. */
class AMCLLaserData : public AMCLSensorData
{
public:
  AMCLLaserData() { ranges = NULL; }
  virtual ~AMCLLaserData() { delete[] ranges; }
  int range_count;
  double range_max;
  double (*ranges)[2];         /* (range, bearing) pairs */
};

/* amcl_laser.h:59-151.  sizeof == 192. */
class AMCLLaser : public AMCLSensor
{
public:
  AMCLLaser(size_t max_beams, map_t *map);
  virtual ~AMCLLaser();

  void SetModelBeam(double z_hit, double z_short, double z_max, double z_rand,
                    double sigma_hit, double lambda_short, double chi_outlier);

  /* Both likelihood-field setters rebuild the distance map
   * (map_update_cspace) exactly as the reference does. */
  void SetModelLikelihoodField(double z_hit, double z_rand, double sigma_hit,
                               double max_occ_dist);
  void SetModelLikelihoodFieldProb(double z_hit, double z_rand, double sigma_hit,
                                   double max_occ_dist, bool do_beamskip,
                                   double beam_skip_distance,
                                   double beam_skip_threshold,
                                   double beam_skip_error_threshold);

  /* false when max_beams < 2, otherwise weights every particle on device. */
  virtual bool UpdateSensor(pf_t *pf, AMCLSensorData *data);

  void SetLaserPose(pf_vector_t &laser_pose) { this->laser_pose = laser_pose; }

private:
  /* pf_sensor_model_fn_t-shaped entry points (amcl_laser.h:98-107). */
  static double BeamModel(AMCLLaserData *data, pf_sample_set_t *set);
  static double LikelihoodFieldModel(AMCLLaserData *data, pf_sample_set_t *set);
  static double LikelihoodFieldModelProb(AMCLLaserData *data, pf_sample_set_t *set);
  void reallocTempData(int max_samples, int max_obs);

  laser_model_t model_type;
  double time;
  map_t *map;                  /* not owned */
  pf_vector_t laser_pose;      /* sensor offset in the robot frame */
  int max_beams;

  bool do_beamskip;
  double beam_skip_distance;
  double beam_skip_threshold;
  double beam_skip_error_threshold;

  /* The reference keeps an N x B host scratch here.  This library never
   * allocates it (the beam-skip pass runs as two device passes), so
   * temp_obs stays NULL and shallow copies of the object are harmless. */
  int max_samples;
  int max_obs;
  double **temp_obs;

  double z_hit, z_short, z_max, z_rand;
  double sigma_hit;
  double lambda_short;
  double chi_outlier;

  friend struct LaserAccess;   /* library-internal view of the private block */
};

#if defined(__x86_64__)
static_assert(sizeof(AMCLSensor) == 40, "AMCLSensor");
static_assert(sizeof(AMCLSensorData) == 24, "AMCLSensorData");
static_assert(sizeof(AMCLOdomData) == 72, "AMCLOdomData");
static_assert(sizeof(AMCLOdom) == 96, "AMCLOdom");
static_assert(sizeof(AMCLLaserData) == 48, "AMCLLaserData");
static_assert(sizeof(AMCLLaser) == 192, "AMCLLaser");
#endif

}  /* namespace amcl */

#endif /* AMCL_B200_AMCL_SENSORS_HPP */
