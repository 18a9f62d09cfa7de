// ext_api.cu -- Part 3 of include/amcl_b200.h: the amclb200_* entry points,
// which have no counterpart in the reference.  They expose device residency,
// the host-mirror policy, the random-stream choice and integer by-products of
// resampling (source indices, labels) that the reference only holds implicitly.
#include <string.h>

#include "amcl_b200.h"
#include "context.hpp"

using namespace amclb200;

extern "C" {

const char *amclb200_version(void) { return "amcl_b200 0.1 sm_100a"; }

void amclb200_set_stream(void *cuda_stream) { globals().stream = (cudaStream_t)cuda_stream; }

void amclb200_set_rng(int mode, uint64_t seed)
{
  globals().rng = mode == 1 ? RNG_DRAND48 : RNG_PHILOX;
  if (seed) globals().philox_seed = seed;
}

void amclb200_set_mirror(pf_t *pf, int eager) { pf_ctx(pf)->mirror_eager = eager != 0; }

void amclb200_pf_pull_samples(pf_t *pf)
{
  PfCtx *c = pf_ctx(pf);
  c->host_fresh[pf->current_set] = false;
  pf_pull(c, pf->current_set);
}

void amclb200_pf_push_samples(pf_t *pf) { pf_push(pf_ctx(pf), pf->current_set); }

void *amclb200_pf_device_samples(pf_t *pf, int set_index)
{
  return pf_ctx(pf)->d_set[set_index & 1];
}

void amclb200_map_push_distances(map_t *map)
{
  MapCtx *m = map_ctx(map, true);
  if (!m->occ_uploaded) map_upload_occupancy(m);
  map_codes_from_host_cells(m);
}

static int copy_ints(const int *d_src, int n, int *out, int capacity)
{
  if (n > capacity) n = capacity;
  if (n <= 0) return 0;
  AMCL_CUDA(cudaMemcpyAsync(out, d_src, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost,
                            globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
  return n;
}

int amclb200_pf_resample_sources(pf_t *pf, int *out, int capacity)
{
  PfCtx *c = pf_ctx(pf);
  return copy_ints(c->d_src, pf->sets[pf->current_set].sample_count, out, capacity);
}

int amclb200_pf_cluster_labels(pf_t *pf, int *out, int capacity)
{
  PfCtx *c = pf_ctx(pf);
  return copy_ints(c->d_label, pf->sets[pf->current_set].sample_count, out, capacity);
}

int amclb200_pf_cumulative_table(pf_t *pf, double *out, int capacity)
{
  PfCtx *c = pf_ctx(pf);
  const int n = pf->sets[pf->current_set].sample_count;
  if (capacity < n + 1) return 0;
  exact_cumsum(c, c->d_set[pf->current_set], n, c->d_cum);
  AMCL_CUDA(cudaMemcpyAsync(out, c->d_cum, sizeof(double) * (size_t)(n + 1),
                            cudaMemcpyDeviceToHost, globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
  return n + 1;
}

void amclb200_pf_last_timings(pf_t *pf, float out_ms[3])
{
  PfCtx *c = pf_ctx(pf);
  out_ms[0] = c->ms_action;
  out_ms[1] = c->ms_sensor;
  out_ms[2] = c->ms_resample;
}

void amclb200_map_calc_ranges(map_t *map, const double *rays_xya, int n, double max_range,
                              double *out_ranges)
{
  MapCtx *m = map_ctx(map, true);
  if (!m->occ_uploaded) map_upload_occupancy(m);
  run_calc_ranges(m, rays_xya, n, max_range, out_ranges);
}

}  // extern "C"
