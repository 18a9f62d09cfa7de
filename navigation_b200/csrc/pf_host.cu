// pf_host.cu -- the pf_* entry points of /root/reference/amcl/include/amcl/pf/pf.h:146-203
// re-hosted on the device context: allocation, initialisation, the action /
// sensor / resample plumbing and the statistics getters.  The O(N) work of
// every call runs in the kernels of motion_dev.cu / sensor_dev.cu /
// resample_dev.cu; what stays on the host is the scalar bookkeeping the
// reference keeps in pf_t and the callbacks that are host functions by
// contract (pf_init_model_fn_t, foreign action/sensor models).
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <unordered_map>
#include <vector>

#include "amcl/pf/pf.h"
#include "amcl/pf/pf_pdf.h"
#include "context.hpp"

extern "C" int amclb200_get_reseed(void);   // host_math.cpp

namespace amcl
{
bool amclb200_is_device_laser_model(pf_sensor_model_fn_t fn);  // sensor_dev.cu
}

namespace amclb200
{

static std::unordered_map<pf_t *, PfCtx *> &pf_registry()
{
  static std::unordered_map<pf_t *, PfCtx *> r;
  return r;
}

PfCtx *pf_ctx(pf_t *pf)
{
  auto &r = pf_registry();
  auto it = r.find(pf);
  if (it == r.end()) AMCL_FATAL("pf_t %p was not created by this library's pf_alloc", (void *)pf);
  return it->second;
}

PfCtx *pf_ctx_of_set(pf_sample_set_t *set, int *which)
{
  for (auto &kv : pf_registry())
  {
    pf_t *pf = kv.first;
    if (set == pf->sets + 0 || set == pf->sets + 1)
    {
      *which = set == pf->sets + 0 ? 0 : 1;
      return kv.second;
    }
  }
  AMCL_FATAL("sample set %p does not belong to a filter of this library", (void *)set);
  return nullptr;
}

static unsigned int next_pow2(unsigned int v)
{
  unsigned int p = 1024;
  while (p < v) p <<= 1;
  return p;
}

PfCtx *pf_ctx_create(pf_t *pf)
{
  ensure_device();
  PfCtx *c = new PfCtx();
  c->host = pf;
  // arrays are padded so that world * chunk(n) slices always fit (Shard::chunk)
  const size_t maxs = (size_t)pf->max_samples + kShardPad;
  c->max_samples = pf->max_samples;
  c->alloc_samples = (int)maxs;
  c->hash_cap = next_pow2((unsigned int)(2 * (size_t)pf->max_samples + 2));   // bins <= samples: load factor <= 1/2
  for (int j = 0; j < 2; j++) AMCL_CUDA(cudaMalloc(&c->d_set[j], sizeof(Particle) * maxs));
  // entry 0 of the table sits 8 bytes in FRONT of a 256-byte boundary, so that
  // entries 1.. -- the part a sharded filter all-gathers slice by slice -- are aligned
  AMCL_CUDA(cudaMalloc(&c->d_cum_raw, sizeof(double) * (maxs + 1 + 32)));
  c->d_cum = c->d_cum_raw + 31;
  AMCL_CUDA(cudaMalloc(&c->d_draws, sizeof(double) * 2 * maxs));
  AMCL_CUDA(cudaMalloc(&c->d_src, sizeof(int) * maxs));
  AMCL_CUDA(cudaMalloc(&c->d_label, sizeof(int) * maxs));
  AMCL_CUDA(cudaMalloc(&c->d_keys, sizeof(unsigned long long) * maxs));
  AMCL_CUDA(cudaMalloc(&c->d_slot, sizeof(unsigned int) * maxs));
  AMCL_CUDA(cudaMalloc(&c->d_flag, sizeof(unsigned int) * maxs));
  AMCL_CUDA(cudaMalloc(&c->d_scan, sizeof(unsigned int) * (maxs + c->hash_cap)));
  c->scan_tiles_cap = (int)(maxs / 4096 + 4);
  AMCL_CUDA(cudaMalloc(&c->d_scan_tiles, sizeof(unsigned int) * (size_t)c->scan_tiles_cap));
  AMCL_CUDA(cudaMalloc(&c->d_tab_key, sizeof(unsigned long long) * (size_t)c->hash_cap));
  AMCL_CUDA(cudaMalloc(&c->d_tab_first, sizeof(unsigned int) * (size_t)c->hash_cap));
  AMCL_CUDA(cudaMalloc(&c->d_tab_aux, sizeof(unsigned int) * (size_t)c->hash_cap));
  AMCL_CUDA(cudaMalloc(&c->d_scalars, sizeof(double) * 64));
  AMCL_CUDA(cudaMalloc(&c->d_ints, sizeof(int) * 64));
  AMCL_CUDA(cudaMallocHost(&c->h_scalars, sizeof(double) * 64));
  AMCL_CUDA(cudaMallocHost(&c->h_ints, sizeof(int) * 64));
  pf_registry()[pf] = c;
  return c;
}

void pf_ctx_release(pf_t *pf)
{
  auto &r = pf_registry();
  auto it = r.find(pf);
  if (it == r.end()) return;
  PfCtx *c = it->second;
  for (int j = 0; j < 2; j++) cudaFree(c->d_set[j]);
  if (c->d_motion_draws) cudaFree(c->d_motion_draws);
  for (void *p : c->peer_opened) cudaIpcCloseMemHandle(p);
  for (int j = 0; j < 2; j++) if (c->d_peer_tab[j]) cudaFree((void *)c->d_peer_tab[j]);
  if (c->d_wpack) cudaFree(c->d_wpack);
  if (c->shard.nccl_comm) nccl_release(c->shard.nccl_comm);
  for (int j = 0; j < 2; j++)
    if (c->small_graph[j].exec) cudaGraphExecDestroy(c->small_graph[j].exec);
  if (c->capture_stream) cudaStreamDestroy(c->capture_stream);
  for (int k = 0; k < 3; k++)
    for (int j = 0; j < 2; j++)
      if (c->ev_stage[k][j]) cudaEventDestroy(c->ev_stage[k][j]);
  cudaFree(c->d_cum_raw); cudaFree(c->d_draws); cudaFree(c->d_src); cudaFree(c->d_label);
  cudaFree(c->d_keys); cudaFree(c->d_slot); cudaFree(c->d_flag); cudaFree(c->d_scan);
  cudaFree(c->d_scan_tiles); cudaFree(c->d_tab_key); cudaFree(c->d_tab_first); cudaFree(c->d_tab_aux); cudaFree(c->d_scalars); cudaFree(c->d_ints);
  if (c->d_cluster_acc) cudaFree(c->d_cluster_acc);
  if (c->d_partials) cudaFree(c->d_partials);
  if (c->d_scan_tab) cudaFree(c->d_scan_tab);
  if (c->d_lut) cudaFree(c->d_lut);
  if (c->d_skip) cudaFree(c->d_skip);
  if (c->d_cloud) cudaFree(c->d_cloud);
  if (c->d_logsum) cudaFree(c->d_logsum);
  if (c->d_bucket_index) cudaFree(c->d_bucket_index);
  if (c->d_rootlist) cudaFree(c->d_rootlist);
  if (c->d_tile_sum) cudaFree(c->d_tile_sum);
  if (c->d_tile_carry) cudaFree(c->d_tile_carry);
  if (c->d_tile_fn) cudaFree(c->d_tile_fn);
  if (c->d_tile_class) cudaFree(c->d_tile_class);
  if (c->d_dirty_w) cudaFree(c->d_dirty_w);
  if (c->d_dirty_row) cudaFree(c->d_dirty_row);
  cudaFreeHost(c->h_scalars);
  cudaFreeHost(c->h_ints);
  delete c;
  r.erase(it);
}

void pf_replicate(PfCtx *c, int which)
{
  if (!c->shard.on() || c->replicated[which]) return;
  const int n = c->host->sets[which].sample_count;
  if (n > 0)
    c->shard.allgather(c->d_set[which],
                       (unsigned long long)c->shard.chunk(n) * sizeof(Particle), c->shard.user);
  c->replicated[which] = true;
}

// Exchange CUDA IPC handles of the two particle arrays through the host-supplied
// all-gather and open every peer's arrays, so that kernels can read remote
// particles directly over NVLink (resample_dev.cu, pick_and_copy).
void pf_enable_peers(PfCtx *c)
{
  const Shard &sh = c->shard;
  if (!sh.on() || c->peers_ready) return;
  cudaStream_t st = globals().stream;
  const size_t per_rank = 2 * sizeof(cudaIpcMemHandle_t);   // 128 bytes
  unsigned char *d_stage = nullptr;
  AMCL_CUDA(cudaMalloc(&d_stage, per_rank * (size_t)sh.world));
  cudaIpcMemHandle_t mine[2];
  for (int j = 0; j < 2; j++) AMCL_CUDA(cudaIpcGetMemHandle(&mine[j], c->d_set[j]));
  AMCL_CUDA(cudaMemcpyAsync(d_stage + per_rank * (size_t)sh.rank, mine, per_rank,
                            cudaMemcpyHostToDevice, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  sh.allgather(d_stage, per_rank, sh.user);
  std::vector<cudaIpcMemHandle_t> all((size_t)2 * sh.world);
  AMCL_CUDA(cudaMemcpyAsync(all.data(), d_stage, per_rank * (size_t)sh.world, cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  cudaFree(d_stage);
  for (int j = 0; j < 2; j++)
  {
    std::vector<const Particle *> tab((size_t)sh.world);
    for (int r = 0; r < sh.world; r++)
    {
      if (r == sh.rank)
      {
        tab[r] = c->d_set[j];
        continue;
      }
      void *ptr = nullptr;
      AMCL_CUDA(cudaIpcOpenMemHandle(&ptr, all[(size_t)2 * r + j], cudaIpcMemLazyEnablePeerAccess));
      c->peer_opened.push_back(ptr);
      tab[r] = reinterpret_cast<const Particle *>(ptr);
    }
    AMCL_CUDA(cudaMalloc((void **)&c->d_peer_tab[j], sizeof(Particle *) * (size_t)sh.world));
    AMCL_CUDA(cudaMemcpy((void *)c->d_peer_tab[j], tab.data(), sizeof(Particle *) * (size_t)sh.world,
                         cudaMemcpyHostToDevice));
  }
  if (!c->d_wpack) AMCL_CUDA(cudaMalloc(&c->d_wpack, sizeof(double) * (size_t)c->alloc_samples));
  c->peers_ready = true;
}

void pf_pull(PfCtx *c, int which)
{
  if (c->host_fresh[which]) return;
  pf_replicate(c, which);   // a collective when sharded: every rank pulls together
  pf_sample_set_t *set = c->host->sets + which;
  if (set->sample_count > 0)
    AMCL_CUDA(cudaMemcpyAsync(set->samples, c->d_set[which],
                              sizeof(Particle) * (size_t)set->sample_count,
                              cudaMemcpyDeviceToHost, globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
  c->host_fresh[which] = true;
}

void pf_push(PfCtx *c, int which)
{
  pf_sample_set_t *set = c->host->sets + which;
  if (set->sample_count > c->max_samples)
    AMCL_FATAL("sample_count %d exceeds max_samples %d", set->sample_count, c->max_samples);
  if (set->sample_count > 0)
    AMCL_CUDA(cudaMemcpyAsync(c->d_set[which], set->samples,
                              sizeof(Particle) * (size_t)set->sample_count,
                              cudaMemcpyHostToDevice, globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
  c->host_fresh[which] = true;
  c->replicated[which] = true;
}

void pf_after_update(PfCtx *c, int which)
{
  if (c->mirror_eager) pf_pull(c, which);
}

}  // namespace amclb200

using namespace amclb200;

extern "C" {

// ---------------------------------------------------------------- alloc / free --

// pf.c:46-115
pf_t *pf_alloc(int min_samples, int max_samples, double alpha_slow, double alpha_fast,
               pf_init_model_fn_t random_pose_fn, void *random_pose_data)
{
  if (max_samples < 1) AMCL_FATAL("pf_alloc: max_samples = %d", max_samples);
  // The reference seeds the process-wide drand48 stream here (pf.c:55); the
  // caller's own pose generator draws from that stream too, so keep doing it.
  if (amclb200_get_reseed()) srand48(time(NULL));

  pf_t *pf = (pf_t *)calloc(1, sizeof(pf_t));
  pf->random_pose_fn = random_pose_fn;
  pf->random_pose_data = random_pose_data;
  pf->min_samples = min_samples;
  pf->max_samples = max_samples;
  pf->pop_err = 0.01;          // pf.c:70-72
  pf->pop_z = 3;
  pf->dist_threshold = 0.5;
  pf->current_set = 0;

  PfCtx *c = pf_ctx_create(pf);
  pf->limit_cache = reinterpret_cast<int *>(c);   // opaque handle (see amcl_abi.h)

  const double w0 = 1.0 / max_samples;
  for (int j = 0; j < 2; j++)
  {
    pf_sample_set_t *set = pf->sets + j;
    set->sample_count = max_samples;
    // pinned, so the mirror copy runs at full PCIe rate
    AMCL_CUDA(cudaMallocHost((void **)&set->samples, sizeof(pf_sample_t) * (size_t)max_samples));
    for (int i = 0; i < max_samples; i++)
    {
      set->samples[i].pose.v[0] = set->samples[i].pose.v[1] = set->samples[i].pose.v[2] = 0.0;
      set->samples[i].weight = w0;
    }
    set->kdtree = pf_kdtree_alloc(3 * max_samples);
    set->cluster_count = 0;
    set->cluster_max_count = max_samples;
    set->clusters = (pf_cluster_t *)calloc((size_t)max_samples, sizeof(pf_cluster_t));
    set->mean = pf_vector_zero();
    set->cov = pf_matrix_zero();
    pf_push(c, j);
  }
  pf->w_slow = 0.0;
  pf->w_fast = 0.0;
  pf->alpha_slow = alpha_slow;
  pf->alpha_fast = alpha_fast;
  pf_init_converged(pf);
  return pf;
}

// pf.c:118-133
void pf_free(pf_t *pf)
{
  if (!pf) return;
  pf_ctx_release(pf);
  for (int j = 0; j < 2; j++)
  {
    free(pf->sets[j].clusters);
    pf_kdtree_free(pf->sets[j].kdtree);
    cudaFreeHost(pf->sets[j].samples);
  }
  free(pf);
}

// ----------------------------------------------------------------------- init --

// pf.c:136-174
void pf_init(pf_t *pf, pf_vector_t mean, pf_matrix_t cov)
{
  ApiRange range("amcl_b200:pf_init");
  PfCtx *c = pf_ctx(pf);
  const int which = pf->current_set;
  pf_sample_set_t *set = pf->sets + which;
  pf_kdtree_clear(set->kdtree);
  set->sample_count = pf->max_samples;
  // eigen-decomposition and (reference-compatible) re-seeding, pf_pdf.c:47-71
  pf_pdf_gaussian_t *pdf = pf_pdf_gaussian_alloc(mean, cov);
  if (globals().rng == RNG_DRAND48)
  {
    for (int i = 0; i < set->sample_count; i++)
    {
      set->samples[i].weight = 1.0 / pf->max_samples;
      set->samples[i].pose = pf_pdf_gaussian_sample(pdf);
    }
    pf_push(c, which);
  }
  else
  {
    double cr[9];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) cr[3 * i + j] = pdf->cr.m[i][j];
    run_philox_gaussian_init(c, which, mean.v, cr, pdf->cd.v);
  }
  pf->w_slow = pf->w_fast = 0.0;
  pf_pdf_gaussian_free(pdf);
  run_cluster_stats(c, which, false);
  pf_init_converged(pf);
  pf_after_update(c, which);
}

// pf.c:178-211 -- the generator is a host callback by contract.
void pf_init_model(pf_t *pf, pf_init_model_fn_t init_fn, void *init_data)
{
  PfCtx *c = pf_ctx(pf);
  const int which = pf->current_set;
  pf_sample_set_t *set = pf->sets + which;
  pf_kdtree_clear(set->kdtree);
  set->sample_count = pf->max_samples;
  for (int i = 0; i < set->sample_count; i++)
  {
    set->samples[i].weight = 1.0 / pf->max_samples;
    set->samples[i].pose = (*init_fn)(init_data);
  }
  pf_push(c, which);
  pf->w_slow = pf->w_fast = 0.0;
  run_cluster_stats(c, which, false);
  pf_init_converged(pf);
}

// pf.c:213-218
void pf_init_converged(pf_t *pf)
{
  pf->sets[pf->current_set].converged = 0;
  pf->converged = 0;
}

// pf.c:220-251
int pf_update_converged(pf_t *pf)
{
  return run_update_converged(pf_ctx(pf), pf->current_set);
}

// -------------------------------------------------------------------- updates --

// pf.c:254-263.  A foreign action model is a host function that edits the
// host sample array, so the set is mirrored down and pushed back around it.
// AMCLOdom::UpdateAction does not come through here (amcl_node calls it
// directly, amcl_node.cpp:1247) and runs on the device.
void pf_update_action(pf_t *pf, pf_action_model_fn_t action_fn, void *action_data)
{
  PfCtx *c = pf_ctx(pf);
  const int which = pf->current_set;
  pf_pull(c, which);
  (*action_fn)(action_data, pf->sets + which);
  pf_push(c, which);
}

// pf.c:268-318
void pf_update_sensor(pf_t *pf, pf_sensor_model_fn_t sensor_fn, void *sensor_data)
{
  ApiRange range("amcl_b200:pf_update_sensor");
  PfCtx *c = pf_ctx(pf);
  const int which = pf->current_set;
  pf_sample_set_t *set = pf->sets + which;
  StageTimer timer(c, 1);
  double total;
  if (amcl::amclb200_is_device_laser_model(sensor_fn))
  {
    c->defer_total = true;                       // the total is fetched once, with n_effective
    total = (*sensor_fn)(sensor_data, set);      // device kernel, see sensor_dev.cu
  }
  else
  {
    pf_pull(c, which);
    total = (*sensor_fn)(sensor_data, set);      // foreign host model
    pf_push(c, which);
  }
  run_sensor_normalise(c, which, total);
  c->defer_total = false;
  pf_after_update(c, which);
}

// pf.c:363-510
void pf_update_resample(pf_t *pf)
{
  ApiRange range("amcl_b200:pf_update_resample");
  PfCtx *c = pf_ctx(pf);
  StageTimer timer(c, 2);
  run_resample(c);
  pf_after_update(c, pf->current_set);
}

// pf.c:695-698
void pf_set_selective_resampling(pf_t *pf, int selective_resampling)
{
  pf->selective_resampling = selective_resampling;
}

// pf.c:321-360 (the reference defines copy_set and declares pf_copy_set)
void copy_set(pf_sample_set_t *set_a, pf_sample_set_t *set_b)
{
  int wa = 0, wb = 0;
  PfCtx *c = pf_ctx_of_set(set_a, &wa);
  PfCtx *cb = pf_ctx_of_set(set_b, &wb);
  if (c != cb) AMCL_FATAL("copy_set across two filters");
  pf_kdtree_clear(set_b->kdtree);
  run_copy_set(c, wa, wb);
  if (c->mirror_eager) pf_pull(c, wb);
}
void pf_copy_set(pf_sample_set_t *set_a, pf_sample_set_t *set_b) { copy_set(set_a, set_b); }

// ----------------------------------------------------------------- statistics --

// pf.c:559-693
void pf_cluster_stats(pf_t *pf, pf_sample_set_t *set)
{
  int which = 0;
  PfCtx *c = pf_ctx_of_set(set, &which);
  (void)pf;
  run_cluster_stats(c, which, false);
}

// pf.c:701-740
void pf_get_cep_stats(pf_t *pf, pf_vector_t *mean, double *var)
{
  PfCtx *c = pf_ctx(pf);
  double s[4];
  run_cep_stats(c, pf->current_set, s);
  const double mn = s[0], mx = s[1], my = s[2], mrr = s[3];
  if (fabs(mn) < 2.220446049250313e-16)
  {
    printf("ERROR : divide-by-zero exception : mn is zero\n");
    return;
  }
  mean->v[0] = mx / mn;
  mean->v[1] = my / mn;
  mean->v[2] = 0.0;
  *var = mrr / mn - (mx * mx / (mn * mn) + my * my / (mn * mn));
}

// pf.c:744-761
int pf_get_cluster_stats(pf_t *pf, int clabel, double *weight, pf_vector_t *mean, pf_matrix_t *cov)
{
  pf_sample_set_t *set = pf->sets + pf->current_set;
  if (clabel >= set->cluster_count) return 0;
  pf_cluster_t *cluster = set->clusters + clabel;
  *weight = cluster->weight;
  *mean = cluster->mean;
  *cov = cluster->cov;
  return 1;
}

}  // extern "C"
