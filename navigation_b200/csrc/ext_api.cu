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

void amclb200_map_set_mirror(map_t *map, int eager) { map_ctx(map, true)->mirror_eager = eager != 0; }

void amclb200_map_pull_distances(map_t *map)
{
  MapCtx *m = map_ctx(map, true);
  if (!m->dist_valid) AMCL_FATAL("amclb200_map_pull_distances: no distance map has been computed");
  map_codes_to_host_cells(m);
}

int amclb200_map_push_distances(map_t *map)
{
  MapCtx *m = map_ctx(map, true);
  if (!m->occ_uploaded) map_upload_occupancy(m);
  return map_codes_from_host_cells(m) ? 0 : -1;
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
  stage_times_resolve(c);
  out_ms[0] = c->ms_action;
  out_ms[1] = c->ms_sensor;
  out_ms[2] = c->ms_resample;
}

// Device milliseconds of the distance-transform kernels of the last map_update_cspace.
double amclb200_map_last_edt_ms(map_t *map)
{
  MapCtx *m = map_ctx(map, false);
  return m ? (double)m->ms_edt : -1.0;
}

void amclb200_map_calc_ranges(map_t *map, const double *rays_xya, int n, double max_range,
                              double *out_ranges)
{
  MapCtx *m = map_ctx(map, true);
  if (!m->occ_uploaded) map_upload_occupancy(m);
  run_calc_ranges(m, rays_xya, n, max_range, out_ranges);
}

}  // extern "C"

// ---- measurement helpers --------------------------------------------------------

// Uncorrelated 2-byte gathers over a table of `table_bytes`: every lane of every
// warp reads an independent pseudo-random element, so each gather costs one
// 32-byte L2 sector and nothing coalesces.  This is the roofline the laser
// kernels are held against ("measured random-gather peak").
__global__ void gather_probe_kernel(const uint16_t *__restrict__ table, unsigned int elems,
                                    int per_thread, unsigned int *sink)
{
  unsigned int x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  unsigned int acc = 0;
  for (int k = 0; k < per_thread; k += 8)
  {
    unsigned int idx[8];
#pragma unroll
    for (int u = 0; u < 8; u++)
    {
      x ^= x << 13; x ^= x >> 17; x ^= x << 5;   // xorshift32
      idx[u] = __umulhi(x, elems);                // uniform over the whole table, any size
    }
#pragma unroll
    for (int u = 0; u < 8; u++) acc += __ldg(table + idx[u]);
  }
  if (acc == 0xFFFFFFFFu) *sink = acc;   // keeps the loads alive
}

extern "C" {

// Copies n particles from a DEVICE array (32-byte records) into the current set
// and makes n its sample_count: lets a benchmark restart every step from the
// same cloud without touching the host.
void amclb200_pf_load_device(pf_t *pf, const void *d_particles, int n)
{
  PfCtx *c = pf_ctx(pf);
  if (n < 0 || n > c->max_samples) AMCL_FATAL("amclb200_pf_load_device: n = %d", n);
  const int which = pf->current_set;
  AMCL_CUDA(cudaMemcpyAsync(c->d_set[which], d_particles, sizeof(Particle) * (size_t)n,
                            cudaMemcpyDeviceToDevice, globals().stream));
  pf->sets[which].sample_count = n;
  c->host_fresh[which] = false;
}

// The node's hypothesis selection (amcl_node.cpp:1369-1394): the cluster of
// the current set with the largest weight.  Returns its label, or -1.
int amclb200_pf_best_cluster(pf_t *pf, double *weight, pf_vector_t *mean, pf_matrix_t *cov)
{
  pf_sample_set_t *set = pf->sets + pf->current_set;
  double max_weight = 0.0;
  int best = -1;
  for (int k = 0; k < set->cluster_count; k++)
  {
    double w;
    pf_vector_t m;
    pf_matrix_t c;
    if (!pf_get_cluster_stats(pf, k, &w, &m, &c)) break;
    if (w > max_weight)
    {
      max_weight = w;
      best = k;
      *weight = w;
      *mean = m;
      *cov = c;
    }
  }
  return best;
}

// ---- sharding over GPUs (one process per GPU) --------------------------------------

// Declares this filter to be rank `rank` of `world` cooperating processes that
// all hold a filter of the same max_samples and issue the same call sequence.
// From then on UpdateAction / UpdateSensor / pf_update_resample touch only the
// rank's slice [rank*ceil(n/world), ...) of the particle arrays and call the
// two host-supplied collectives (in place, on device buffers, ordered with the
// library's stream) where the reference's algorithm needs global data.
void amclb200_pf_set_shard(pf_t *pf, int rank, int world, shard_allreduce_fn allreduce,
                           shard_allgather_fn allgather, void *user)
{
  PfCtx *c = pf_ctx(pf);
  if (world < 1 || rank < 0 || rank >= world || world > 1024)
    AMCL_FATAL("amclb200_pf_set_shard: rank %d of %d", rank, world);
  if (world > 1 && (!allreduce || !allgather))
    AMCL_FATAL("amclb200_pf_set_shard: collectives are required when world > 1");
  c->shard.rank = rank;
  c->shard.world = world;
  c->shard.align = world > 1 && world <= kShardMaxAligned ? kShardTile : 1;
  c->shard.cb_allreduce = allreduce;
  c->shard.cb_allgather = allgather;
  c->shard.user = user;
}

// Collective (every rank calls it once, after amclb200_pf_set_shard): maps every
// rank's particle arrays into this process with CUDA IPC so that resampling can
// read the poses it picks straight from their owners over NVLink instead of
// all-gathering the whole set.
void amclb200_pf_enable_peer_access(pf_t *pf) { pf_enable_peers(pf_ctx(pf)); }

// Loads THIS RANK'S slice of a cloud of n_global particles from a device array
// holding just that slice (ceil(n_global/world) records or fewer for the last rank).
void amclb200_pf_load_device_shard(pf_t *pf, const void *d_slice, int n_global)
{
  PfCtx *c = pf_ctx(pf);
  if (n_global < 0 || n_global > c->max_samples) AMCL_FATAL("load_device_shard: n = %d", n_global);
  const int which = pf->current_set;
  const int lo = c->shard.lo(n_global), cnt = c->shard.count(n_global);
  if (cnt > 0)
    AMCL_CUDA(cudaMemcpyAsync(c->d_set[which] + lo, d_slice, sizeof(Particle) * (size_t)cnt,
                              cudaMemcpyDeviceToDevice, globals().stream));
  pf->sets[which].sample_count = n_global;
  c->host_fresh[which] = false;
  c->replicated[which] = !c->shard.on();
}

// [lo, hi) of this rank's slice for a set of n particles.
void amclb200_pf_shard_range(pf_t *pf, int n, int out[2])
{
  PfCtx *c = pf_ctx(pf);
  out[0] = c->shard.lo(n);
  out[1] = c->shard.hi(n);
}

unsigned long long amclb200_launch_count(void) { return amcl_launch_counter(); }

// How often the exact cumulative table had to fall back to the single-thread walk.
unsigned long long amclb200_cumsum_fallbacks(void) { return exact_cumsum_fallback_count(); }

// out = {tiles walked sequentially, tiles} of the last tile-parallel table
void amclb200_cumsum_last_tiles(int out[2]) { exact_cumsum_last_tiles(out); }

// 1: tile-parallel exact table for large sets (default); 0: always the single-thread walk.
void amclb200_set_parallel_cumsum(int on) { globals().exact_cumsum_parallel = on != 0; }

// out[0] = laser-model kernel ms of the last UpdateSensor, out[1] = its
// particle-beam evaluation count
void amclb200_pf_sensor_kernel_stats(pf_t *pf, double out[2])
{
  PfCtx *c = pf_ctx(pf);
  if (c->sensor_kernel_pending && c->ev_b)
  {
    AMCL_CUDA(cudaEventSynchronize(c->ev_b));
    AMCL_CUDA(cudaEventElapsedTime(&c->ms_sensor_kernel, c->ev_a, c->ev_b));
    c->sensor_kernel_pending = false;
  }
  out[0] = (double)c->ms_sensor_kernel;
  out[1] = (double)c->sensor_evals;
}

// Random 2-byte gathers per second over a table of the given footprint.
double amclb200_measure_gather_rate(unsigned long long table_bytes, int reps)
{
  ensure_device();
  if (table_bytes < 64) table_bytes = 64;
  if (table_bytes > (8ull << 30)) table_bytes = 8ull << 30;
  const unsigned int elems = (unsigned int)(table_bytes / 2);   // the footprint asked for, not a power of two
  uint16_t *d_table = nullptr;
  unsigned int *d_sink = nullptr;
  AMCL_CUDA(cudaMalloc(&d_table, (size_t)elems * 2));
  AMCL_CUDA(cudaMalloc(&d_sink, 4));
  AMCL_CUDA(cudaMemset(d_table, 1, (size_t)elems * 2));
  const int threads = 256, blocks = globals().sm_count * 8 * 4, per_thread = 256;
  cudaEvent_t a, b;
  AMCL_CUDA(cudaEventCreate(&a));
  AMCL_CUDA(cudaEventCreate(&b));
  cudaStream_t st = globals().stream;
  gather_probe_kernel<<<blocks, threads, 0, st>>>(d_table, elems, per_thread, d_sink);  // warm
  float best = 1e30f;
  for (int r = 0; r < reps; r++)
  {
    AMCL_CUDA(cudaEventRecord(a, st));
    gather_probe_kernel<<<blocks, threads, 0, st>>>(d_table, elems, per_thread, d_sink);
    AMCL_CUDA(cudaEventRecord(b, st));
    AMCL_CUDA(cudaEventSynchronize(b));
    float ms = 0;
    AMCL_CUDA(cudaEventElapsedTime(&ms, a, b));
    if (ms < best) best = ms;
  }
  AMCL_CUDA(cudaGetLastError());
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(d_table);
  cudaFree(d_sink);
  return (double)blocks * threads * per_thread / (best * 1e-3);
}

}  // extern "C"
