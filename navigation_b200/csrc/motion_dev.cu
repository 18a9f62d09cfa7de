// motion_dev.cu -- odometry sample_motion_model on the device
// (/root/reference/amcl/src/amcl/sensors/amcl_odom.cpp:113-310) and the
// Gaussian sampler it draws from (/root/reference/amcl/src/amcl/pf/pf_pdf.c:133-147).
//
// Everything that does not depend on the particle (delta_rot1/2, delta_trans,
// the noise terms and the three sigmas) is computed on the host with libm in
// the reference's expression order and handed to the kernel; one thread then
// moves one particle.
//
// Randomness.  Production mode draws from a per-particle Philox4x32-10 stream
// (key = seed ^ update counter, counter = particle index) and runs the same
// polar rejection sampler as pf_ran_gaussian.  Parity mode (RNG_DRAND48) takes
// the uniforms from libc drand48() on the host IN THE REFERENCE'S ORDER --
// particle by particle, three Gaussians each, rejections included -- and ships
// the accepted (x2, sqrt(-2 ln w / w)) pairs, so both implementations consume
// the same stream and produce the same poses up to device-libm rounding.
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "amcl/sensors/amcl_odom.h"
#include "context.hpp"
#include "philox.cuh"

namespace amclb200
{

struct MotionArgs
{
  Particle *set;
  int n;
  int model;                 // odom_model_t
  // DIFF: rot1_n, trans, rot2_n, sd1, sd2, sd3   (rot*_n already wrapped)
  // OMNI: bearing0, trans, rot, sd_trans, sd_rot, sd_strafe
  double k[6];
  const double *draws;       // parity: 6 doubles per particle (global index), else null
  uint64_t key;              // Philox key
  int offset;                // global index of set[0] (sharded runs own a slice)
};

// normalize() of amcl_odom.cpp:38-42, atan2(sin z, cos z).  For |z| <= 3 (< pi)
// the mathematical result is z itself and the reference's libm chain returns it
// to within an ulp or two, the same size as the difference between host and
// device libm; those angles (every noise sample, in practice) skip the three
// transcendental calls.
__device__ __forceinline__ double wrap_angle(double z)
{
  if (fabs(z) <= 3.0) return z;
  return atan2(sin(z), cos(z));
}

// amcl_odom.cpp:44-58 with the first argument already wrapped on the host.
__device__ __forceinline__ double angle_diff_wrapped_a(double a_wrapped, double b)
{
  const double bw = wrap_angle(b);
  const double d1 = a_wrapped - bw;
  double d2 = 2 * M_PI - fabs(d1);
  if (d1 > 0) d2 *= -1.0;
  return fabs(d1) < fabs(d2) ? d1 : d2;
}

__global__ void motion_kernel(const MotionArgs a)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  Particle s = load_particle(a.set + i);

  double g[3];
  if (a.draws)
  {
    const double *d = a.draws + 6 * (size_t)(a.offset + i);
#pragma unroll
    for (int q = 0; q < 3; q++) g[q] = a.k[3 + q] * d[2 * q] * d[2 * q + 1];  // sigma*x2*s
  }
  else
  {
    PhiloxStream rng(a.key, (uint64_t)(a.offset + i));
    rng.gaussian3(a.k[3], a.k[4], a.k[5], g);
  }

  if (a.model == amcl::ODOM_MODEL_DIFF || a.model == amcl::ODOM_MODEL_DIFF_CORRECTED)
  {
    // amcl_odom.cpp:193-213 / :283-303
    const double rot1_hat = angle_diff_wrapped_a(a.k[0], g[0]);
    const double trans_hat = a.k[1] - g[1];
    const double rot2_hat = angle_diff_wrapped_a(a.k[2], g[2]);
    double sn, cs;
    sincos(s.th + rot1_hat, &sn, &cs);
    s.x += trans_hat * cs;
    s.y += trans_hat * sn;
    s.th += rot1_hat + rot2_hat;
  }
  else
  {
    // amcl_odom.cpp:143-163 / :236-256
    const double bearing = a.k[0] + s.th;
    double sn, cs;
    sincos(bearing, &sn, &cs);
    const double trans_hat = a.k[1] + g[0];
    const double rot_hat = a.k[2] + g[1];
    const double strafe_hat = 0 + g[2];
    s.x += (trans_hat * cs + strafe_hat * sn);
    s.y += (trans_hat * sn - strafe_hat * cs);
    s.th += rot_hat;
  }
  store_particle(a.set + i, s);
}

// pf.c:136-174 in production mode: x = mean + cr * (cd .* N(0,1)^3), w = 1/n.
__global__ void gaussian_init_kernel(Particle *set, int n, double m0, double m1, double m2,
                                     const double *cr_cd /* 9 + 3 */, uint64_t key)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  PhiloxStream rng(key, (uint64_t)i);
  double r[3];
  rng.gaussian3(cr_cd[9], cr_cd[10], cr_cd[11], r);
  Particle s;
  double mean[3] = {m0, m1, m2};
  double out[3];
#pragma unroll
  for (int q = 0; q < 3; q++)
  {
    double x = mean[q];
#pragma unroll
    for (int j = 0; j < 3; j++) x += cr_cd[3 * q + j] * r[j];
    out[q] = x;
  }
  s.x = out[0]; s.y = out[1]; s.th = out[2];
  s.w = 1.0 / n;
  store_particle(set + i, s);
}

void run_motion(PfCtx *c, int which, int model, const double consts[8],
                const double *host_draws_or_null)
{
  cudaStream_t st = globals().stream;
  const int n = c->host->sets[which].sample_count;
  const int lo = c->shard.lo(n), cnt = c->shard.count(n);   // whole set when not sharded
  MotionArgs a;
  a.set = c->d_set[which] + lo;
  a.n = cnt;
  a.offset = lo;
  a.model = model;
  for (int q = 0; q < 6; q++) a.k[q] = consts[q];
  a.draws = nullptr;
  if (host_draws_or_null)
  {
    if (n > c->motion_draws_cap)
    {
      if (c->d_motion_draws) cudaFree(c->d_motion_draws);
      c->motion_draws_cap = n + 1024;
      AMCL_CUDA(cudaMalloc(&c->d_motion_draws, sizeof(double) * 6 * (size_t)c->motion_draws_cap));
    }
    AMCL_CUDA(cudaMemcpyAsync(c->d_motion_draws, host_draws_or_null, sizeof(double) * 6 * (size_t)n,
                              cudaMemcpyHostToDevice, st));
    a.draws = c->d_motion_draws;
  }
  a.key = globals().philox_seed ^ (0x9E3779B97F4A7C15ull * (++c->step));
  if (cnt > 0)
  {
    motion_kernel<<<(cnt + 127) / 128, 128, 0, st>>>(a);
    AMCL_LAUNCH_CHECK();
  }
  if (host_draws_or_null) AMCL_CUDA(cudaStreamSynchronize(st));  // caller's buffer lifetime
  c->host_fresh[which] = false;
  if (c->shard.on()) c->replicated[which] = false;
}

void run_philox_gaussian_init(PfCtx *c, int which, const double mean[3], const double cr[9],
                              const double cd[3])
{
  cudaStream_t st = globals().stream;
  const int n = c->host->sets[which].sample_count;
  double pack[12];
  for (int q = 0; q < 9; q++) pack[q] = cr[q];
  for (int q = 0; q < 3; q++) pack[9 + q] = cd[q];
  AMCL_CUDA(cudaMemcpyAsync(c->d_scalars + 16, pack, sizeof(pack), cudaMemcpyHostToDevice, st));
  const uint64_t key = globals().philox_seed ^ (0xD1B54A32D192ED03ull * (++c->step));
  gaussian_init_kernel<<<(n + 127) / 128, 128, 0, st>>>(c->d_set[which], n, mean[0], mean[1],
                                                        mean[2], c->d_scalars + 16, key);
  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaStreamSynchronize(st));
  c->host_fresh[which] = false;
  c->replicated[which] = true;   // every rank generates the whole (identical) cloud
}

}  // namespace amclb200

// ======================================================= amcl::AMCLOdom (C++) ==

using namespace amclb200;

namespace amcl
{

static double normalize(double z) { return atan2(sin(z), cos(z)); }

// amcl_odom.cpp:44-58
static double angle_diff(double a, double b)
{
  a = normalize(a);
  b = normalize(b);
  const double d1 = a - b;
  double d2 = 2 * M_PI - fabs(d1);
  if (d1 > 0) d2 *= -1.0;
  return fabs(d1) < fabs(d2) ? d1 : d2;
}

AMCLOdom::AMCLOdom() : AMCLSensor()
{
  this->time = 0.0;
  this->model_type = ODOM_MODEL_DIFF;
  alpha1 = alpha2 = alpha3 = alpha4 = alpha5 = 0.0;
}

void AMCLOdom::SetModelDiff(double a1, double a2, double a3, double a4)
{
  this->model_type = ODOM_MODEL_DIFF;
  alpha1 = a1; alpha2 = a2; alpha3 = a3; alpha4 = a4;
}

void AMCLOdom::SetModelOmni(double a1, double a2, double a3, double a4, double a5)
{
  this->model_type = ODOM_MODEL_OMNI;
  alpha1 = a1; alpha2 = a2; alpha3 = a3; alpha4 = a4; alpha5 = a5;
}

void AMCLOdom::SetModel(odom_model_t type, double a1, double a2, double a3, double a4, double a5)
{
  this->model_type = type;
  alpha1 = a1; alpha2 = a2; alpha3 = a3; alpha4 = a4; alpha5 = a5;
}

// One accepted polar pair from libc drand48(), consumed exactly as
// pf_ran_gaussian does (pf_pdf.c:137-144).
static inline void drand48_polar(double *x2_out, double *s_out)
{
  double x1, x2, w, r;
  do
  {
    do { r = drand48(); } while (r == 0.0);
    x1 = 2.0 * r - 1.0;
    do { r = drand48(); } while (r == 0.0);
    x2 = 2.0 * r - 1.0;
    w = x1 * x1 + x2 * x2;
  } while (w > 1.0 || w == 0.0);
  *x2_out = x2;
  *s_out = sqrt(-2.0 * log(w) / w);
}

// amcl_odom.cpp:113-310
bool AMCLOdom::UpdateAction(pf_t *pf, AMCLSensorData *data)
{
  AMCLOdomData *ndata = (AMCLOdomData *)data;
  PfCtx *c = pf_ctx(pf);
  const int which = pf->current_set;
  const pf_vector_t old_pose = pf_vector_sub(ndata->pose, ndata->delta);
  const double *dl = ndata->delta.v;
  double k[8] = {0, 0, 0, 0, 0, 0, 0, 0};

  switch (this->model_type)
  {
    case ODOM_MODEL_OMNI:
    case ODOM_MODEL_OMNI_CORRECTED:
    {
      const double delta_trans = sqrt(dl[0] * dl[0] + dl[1] * dl[1]);
      const double delta_rot = dl[2];
      double sd_trans, sd_rot, sd_strafe;
      if (this->model_type == ODOM_MODEL_OMNI)
      {
        // :131-136 -- the uncorrected model passes these variances as sigma
        sd_trans = (alpha3 * (delta_trans * delta_trans) + alpha1 * (delta_rot * delta_rot));
        sd_rot = (alpha4 * (delta_rot * delta_rot) + alpha2 * (delta_trans * delta_trans));
        sd_strafe = (alpha1 * (delta_rot * delta_rot) + alpha5 * (delta_trans * delta_trans));
      }
      else
      {
        // :227-232
        sd_trans = sqrt(alpha3 * (delta_trans * delta_trans) + alpha4 * (delta_rot * delta_rot));
        sd_rot = sqrt(alpha1 * (delta_rot * delta_rot) + alpha2 * (delta_trans * delta_trans));
        sd_strafe = sqrt(alpha4 * (delta_rot * delta_rot) + alpha5 * (delta_trans * delta_trans));
      }
      k[0] = angle_diff(atan2(dl[1], dl[0]), old_pose.v[2]);   // :142-143
      k[1] = delta_trans;
      k[2] = delta_rot;
      k[3] = sd_trans;
      k[4] = sd_rot;
      k[5] = sd_strafe;
      break;
    }
    default:
    {
      double delta_rot1;
      // :174-180 -- in-place rotation guard
      if (sqrt(dl[1] * dl[1] + dl[0] * dl[0]) < 0.01)
        delta_rot1 = 0.0;
      else
        delta_rot1 = angle_diff(atan2(dl[1], dl[0]), old_pose.v[2]);
      const double delta_trans = sqrt(dl[0] * dl[0] + dl[1] * dl[1]);
      const double delta_rot2 = angle_diff(dl[2], delta_rot1);
      // :187-190
      const double n1a = fabs(angle_diff(delta_rot1, 0.0)), n1b = fabs(angle_diff(delta_rot1, M_PI));
      const double n2a = fabs(angle_diff(delta_rot2, 0.0)), n2b = fabs(angle_diff(delta_rot2, M_PI));
      const double rot1_noise = n1b < n1a ? n1b : n1a;
      const double rot2_noise = n2b < n2a ? n2b : n2a;
      // :198-206 (variance passed as sigma) / :288-296 (sqrt, corrected)
      double s1 = alpha1 * rot1_noise * rot1_noise + alpha2 * delta_trans * delta_trans;
      double s2 = alpha3 * delta_trans * delta_trans + alpha4 * rot1_noise * rot1_noise +
                  alpha4 * rot2_noise * rot2_noise;
      double s3 = alpha1 * rot2_noise * rot2_noise + alpha2 * delta_trans * delta_trans;
      if (this->model_type == ODOM_MODEL_DIFF_CORRECTED)
      {
        s1 = sqrt(s1); s2 = sqrt(s2); s3 = sqrt(s3);
      }
      k[0] = normalize(delta_rot1);
      k[1] = delta_trans;
      k[2] = normalize(delta_rot2);
      k[3] = s1;
      k[4] = s2;
      k[5] = s3;
      break;
    }
  }

  ApiRange range("amcl_b200:AMCLOdom::UpdateAction");
  StageTimer timer(c, 0);
  const int n = pf->sets[which].sample_count;
  if (globals().rng == RNG_DRAND48)
  {
    std::vector<double> draws((size_t)6 * (n > 0 ? n : 1));
    for (int i = 0; i < n; i++)
      for (int q = 0; q < 3; q++) drand48_polar(&draws[6 * (size_t)i + 2 * q], &draws[6 * (size_t)i + 2 * q + 1]);
    run_motion(c, which, (int)this->model_type, k, draws.data());
  }
  else
    run_motion(c, which, (int)this->model_type, k, nullptr);
  pf_after_update(c, which);
  return true;
}

}  // namespace amcl
