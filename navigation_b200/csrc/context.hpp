// context.hpp -- device-side state behind the reference's host structs.
//
// The reference structs (map_t, pf_t, AMCLLaser) cannot grow (they are ABI,
// SURVEY.md 8b), so everything that lives in HBM hangs off two registries:
//   MapCtx  keyed by map_t*   -- occupancy bytes, distance-code map, ray bitmap
//   PfCtx   keyed by pf_t*    -- the two particle arrays and all resampling
//                                scratch (also reachable through
//                                pf_t::limit_cache, which only this library
//                                interprets)
#pragma once

#include <stdint.h>
#include <functional>
#include <utility>
#include <vector>

#include "amcl/map/map.h"
#include "amcl/pf/pf.h"
#include "common.cuh"

namespace amclb200
{

// ---- library-wide settings ----------------------------------------------------

enum RngMode
{
  RNG_PHILOX = 0,   // production: counter-based streams generated on device
  RNG_DRAND48 = 1   // parity: uniforms drawn from libc drand48() on the host in
                    // the reference's exact order, then shipped to the device
};

struct Globals
{
  cudaStream_t stream = 0;     // all launches; 0 = legacy default stream
  RngMode rng = RNG_PHILOX;
  uint64_t philox_seed = 0x243F6A8885A308D3ull;
  int sm_count = 0;
  int device = -1;
  bool exact_cumsum_parallel = true;  // tile-classified exact scan vs 1-thread walk
};
Globals &globals();
void ensure_device();           // fills sm_count/device once; aborts without a GPU

// ---- distance-code map ----------------------------------------------------------
//
// Every value occ_dist can take is sqrt(d2)*scale with integer d2 <= R^2, or
// max_occ_dist (map_cspace.cpp:49,102-105,148).  The device keeps a uint16
// CODE per cell: d2 itself (0..R^2), R^2+1 for "farther than R".  Cells are
// stored in 4x4 blocks of 16 codes = 32 B = one L2 sector, blocks row-major,
// so neighbouring beam end points (in any direction) share sectors.
constexpr int kTile = 4;

struct MapCtx
{
  map_t *host = nullptr;
  int sx = 0, sy = 0;
  int tiles_x = 0, tiles_y = 0;
  double scale = 0, origin_x = 0, origin_y = 0;
  int R = 0;                      // cell radius of the current distance map
  double max_occ_dist = 0;
  bool occ_uploaded = false;
  bool dist_valid = false;
  bool mirror_eager = true;       // map_update_cspace re-reads cells[].occ_state and writes cells[].occ_dist
  uint64_t generation = 0;        // bumps whenever codes / R / scale change

  int8_t *d_occ = nullptr;        // [sy][sx] occupancy states
  uint16_t *d_code = nullptr;     // tiled distance codes
  uint64_t *d_blocked = nullptr;  // 8x8-cell bit tiles: bit set = ray stops here
  int btiles_x = 0, btiles_y = 0;
  uint8_t *d_rowdist = nullptr;   // EDT scratch (search kernels): distance along the row, one byte per cell
  uint16_t *d_rowsq = nullptr;    // EDT scratch: its square, rows of (sx rounded up to 8) cells
  size_t rowsq_cap = 0;
  int8_t *h_occ_pin = nullptr;    // page-locked staging of the occupancy states / the codes
  uint16_t *h_code_pin = nullptr; // (eager mirror: map_update_cspace through the unchanged API)
  size_t h_occ_cap = 0, h_code_cap = 0;
  unsigned long long *d_rowbits = nullptr;  // EDT scratch: occupancy bit strings per row
  size_t rowbits_cap = 0;
  float ms_edt = 0;               // device time of the last distance transform (3 kernels)
  std::vector<double> dist_of_code;  // host: code -> metres (incl. sentinel)

  // Laser-model view of the distance map (sensor_dev.cu).  Per cell the RANK of
  // its code among the codes that can occur at all -- d2 = i^2 + j^2 <= R^2 (505
  // values for R = 40, not 1601), then "far", then "off the map" -- so that the
  // rank -> likelihood table is small enough to sit in shared memory once per
  // bank pair.  4x4 tiles as in d_code but stored column-first (the offset of a
  // cell is three integer operations), and an apron of off-map cells on every
  // side: a particle whose beams all end inside the array needs no bounds clamp.
  uint16_t *d_lf = nullptr;
  size_t lf_cap = 0;              // elements allocated
  int lf_apron = 0;               // cells on every side, a multiple of 4
  int lf_w = 0, lf_h = 0;         // cells including the apron, multiples of 4
  cudaTextureObject_t lf_tex = 0; // d_lf as a linear texture of uint16 (0: not created)
  cudaArray_t lf_array = nullptr; // the laser view row by row in a 2-D array (texture-unit layout)
  cudaTextureObject_t lf_tex2d = 0;
  cudaSurfaceObject_t lf_surf = 0;
  int lf_array_w = 0, lf_array_h = 0;
  uint64_t lf_generation = 0;     // generation of d_code it was derived from (0: never)
  uint16_t *d_rank_of_code = nullptr;
  int rank_cap = 0;
  std::vector<uint16_t> rank_of_code;  // host: code -> rank (0xffff: cannot occur)
  std::vector<int> code_of_rank;       // host: rank -> code; the last two are far, off the map

  int code_far() const { return R * R + 1; }
};

MapCtx *map_ctx(map_t *map, bool create);
void map_ctx_release(map_t *map);
void map_upload_occupancy(MapCtx *m);            // host cells -> d_occ, d_blocked
void map_finish_occupancy_on_device(MapCtx *m);  // d_occ already written on the device -> d_blocked
void parallel_rows(int rows, const std::function<void(int, int)> &body);  // host row loops on several cores
void map_build_distance_codes(MapCtx *m, double max_occ_dist);  // exact EDT
void map_codes_to_host_cells(MapCtx *m);         // fills cells[].occ_dist
bool map_codes_from_host_cells(MapCtx *m);       // encodes cells[].occ_dist; false: a value has no code
void map_prepare_laser_view(MapCtx *m, int reach_cells);  // (re)builds d_lf when stale or its apron is too small

// ---- particle filter -------------------------------------------------------------

constexpr unsigned long long kEmptyKey = ~0ull;

// ---- sharding over GPUs -------------------------------------------------------------
//
// One process per GPU.  Every rank holds a filter sized for the WHOLE particle
// set (arrays are indexed by the global particle index), but owns -- computes
// and keeps current -- only the contiguous slice [rank*chunk, (rank+1)*chunk)
// with chunk = ceil(n / world).  Steps that need another rank's data
// (resampling draws over the global table) first all-gather the slices in
// place.  The collectives themselves are supplied by the host program as
// callbacks (torch.distributed over NCCL in bench.py): they act IN PLACE on
// device buffers and are stream-ordered with the library's stream.
typedef void (*shard_allreduce_fn)(void *d_buf, int count, int dtype, int op, void *user);
typedef void (*shard_allgather_fn)(void *d_buf, unsigned long long bytes_per_rank, void *user);
enum { SHARD_F64 = 0, SHARD_I32 = 1, SHARD_I64 = 2 };
enum { SHARD_SUM = 0, SHARD_MIN = 1, SHARD_MAX = 2 };

constexpr int kShardTile = 1024;      // = the tile length of the exact cumulative table
constexpr int kShardMaxAligned = 32;  // largest world with tile-aligned slices
constexpr int kShardPad = kShardTile * (kShardMaxAligned + 1);   // extra records every particle array carries

struct Shard
{
  int rank = 0, world = 1;
  // Either the host program's callbacks (amclb200_pf_set_shard: torch.distributed
  // in Python programs, gloo on host buffers in the CPU tests) ...
  shard_allreduce_fn cb_allreduce = nullptr;
  shard_allgather_fn cb_allgather = nullptr;
  void *user = nullptr;
  // ... or an NCCL communicator the library owns (amclb200_pf_set_shard_nccl,
  // shard_nccl.cu): collectives are enqueued on the library's stream straight
  // from C++, no host round trip through the caller.
  void *nccl_comm = nullptr;
  // Slices are whole tiles of the cumulative table (cumsum_dev.cu), so that every
  // rank can build its own part of the table; 1 when the world is too large for
  // the padding the arrays carry (kShardPad).
  int align = 1;
  bool on() const { return world > 1; }
  int chunk(int n) const
  {
    const int per = (n + world - 1) / world;
    return (per + align - 1) / align * align;
  }
  int lo(int n) const { int v = rank * chunk(n); return v < n ? v : n; }
  int hi(int n) const { int v = (rank + 1) * chunk(n); return v < n ? v : n; }
  int count(int n) const { return hi(n) - lo(n); }
  // same call shape as the callbacks (the last argument is ignored)
  void allreduce(void *d_buf, int count, int dtype, int op, void * = nullptr) const;
  void allgather(void *d_buf, unsigned long long bytes_per_rank, void * = nullptr) const;
};

// shard_nccl.cu
void nccl_allreduce(void *comm, void *d_buf, int count, int dtype, int op);
void nccl_allgather(void *comm, int rank, void *d_buf, unsigned long long bytes_per_rank);
void nccl_release(void *comm);

// The captured launch chain of the small-filter resampling tail (resample_dev.cu).
struct SmallGraph
{
  cudaGraphExec_t exec = nullptr;
  double key[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // what the captured arguments depend on
  int launches = 0;                           // kernels inside, for amclb200_launch_count
};

struct PfCtx
{
  pf_t *host = nullptr;
  int max_samples = 0;
  int alloc_samples = 0;          // max_samples padded so world*chunk always fits
  Shard shard;
  bool replicated[2] = {true, true};  // whole set valid on this rank (not only the owned slice)
  // peer-memory path: every rank's particle arrays mapped into this process (CUDA IPC
  // over NVLink); d_peer_tab[j][r] = rank r's array for set j, as seen from this GPU
  bool peers_ready = false;
  const Particle **d_peer_tab[2] = {nullptr, nullptr};
  std::vector<void *> peer_opened;
  double *d_wpack = nullptr;      // packed weights of the whole set (sharded resample)
  Particle *d_set[2] = {nullptr, nullptr};
  bool mirror_eager = true;       // D2H the sample array after every update
  bool host_fresh[2] = {false, false};

  // resampling scratch
  double *d_cum = nullptr;        // max+1 cumulative table (d_cum + 1 is 256-byte aligned)
  double *d_cum_raw = nullptr;    // the allocation behind it
  double *d_draws = nullptr;      // 2*max uniforms (u_test, r) of a resample
  double *d_motion_draws = nullptr;  // parity mode: 6 doubles per particle, grown on demand
  int motion_draws_cap = 0;
  int *d_src = nullptr;           // source index per output sample
  unsigned long long *d_keys = nullptr;
  unsigned int *d_slot = nullptr; // hash slot per sample
  unsigned int *d_flag = nullptr; // scan input
  unsigned int *d_scan = nullptr; // scan output [max] + per-slot scratch [hash_cap]
  unsigned int *d_scan_tiles = nullptr;  // per-tile sums of the u32 scan
  int scan_tiles_cap = 0;
  int *d_label = nullptr;         // cluster label per sample (canonical numbering)
  // bin table (see resample_dev.cu): parallel arrays of hash_cap slots
  unsigned long long *d_tab_key = nullptr;
  unsigned int *d_tab_first = nullptr;
  unsigned int *d_tab_aux = nullptr;
  unsigned int hash_cap = 0;      // power of two
  int tab_dense = 0;              // last histogram: dense grid (1) or hash (0)
  int tab_lo[3] = {0, 0, 0}, tab_ext[3] = {0, 0, 0};
  double *d_cluster_acc = nullptr;  // [cluster][10]
  int cluster_acc_cap = 0;
  double *d_partials = nullptr;   // per-block partial sums (several lanes)
  int partials_cap = 0;
  double *d_scalars = nullptr;    // small device scratch (64 doubles)
  double *h_scalars = nullptr;    // pinned mirror of d_scalars
  int *d_ints = nullptr;          // small device scratch (64 ints + beam counts)
  int *h_ints = nullptr;          // pinned
  int ints_cap = 0;
  // exact cumulative-sum scratch
  double *d_tile_sum = nullptr;
  double *d_tile_carry = nullptr;
  unsigned long long *d_tile_fn = nullptr;  // 2 words per tile
  int *d_tile_class = nullptr;
  int tile_cap = 0;
  double *d_dirty_w = nullptr;    // sharded table: weights of the tiles that need the element-wise walk
  int *d_dirty_row = nullptr;     // tile -> row of d_dirty_w
  int dirty_row_cap = 0;

  // laser staging: beam table, sized on demand; rank -> likelihood table, each
  // entry repeated lut_copies times, re-uploaded only when its version changes
  double *d_scan_tab = nullptr;
  int scan_tab_cap = 0;
  std::vector<double> h_beam_tab, h_lut;
  double *d_lut = nullptr;
  size_t lut_cap = 0;
  uint64_t lut_version = 0;
  int lut_copies = 0;
  double *d_skip = nullptr;       // beam-skip pass: per-beam agreement counts + code->agree bytes
  unsigned int *d_rootlist = nullptr;  // cluster roots + their earliest samples (short-list ranking)
  int *d_bucket_index = nullptr;  // bucket index of the cumulative table (resample_dev.cu)
  size_t bucket_cap = 0;
  double *d_logsum = nullptr;     // beam skipping: per-particle log-likelihood over all beams (first pass)
  double *d_cloud = nullptr;      // packed PoseArray for the caller (amclb200_pf_pack_cloud)
  size_t cloud_cap = 0;
  int skip_cap = 0;

  uint64_t step = 0;              // counter feeding the Philox key
  unsigned long long draws_used = 0;

  // last-call timings (ms, CUDA events), for bench/diagnostics
  float ms_action = 0, ms_sensor = 0, ms_resample = 0;
  cudaEvent_t ev_stage[3][2] = {{nullptr, nullptr}, {nullptr, nullptr}, {nullptr, nullptr}};
  bool stage_pending[3] = {false, false, false};
  float ms_sensor_kernel = 0;     // the laser-model kernel alone
  bool sensor_kernel_pending = false;  // ev_a / ev_b recorded, not read yet
  SmallGraph small_graph[2];      // one per destination set
  cudaStream_t capture_stream = nullptr;
  bool defer_total = false;       // pf_update_sensor is calling one of the library's own models: the
                                  // total weight stays on the device until the normalisation reads it
  long long sensor_evals = 0;     // particle-beam evaluations of that launch
  cudaEvent_t ev_a = nullptr, ev_b = nullptr;
};

PfCtx *pf_ctx(pf_t *pf);                       // aborts if unknown
void pf_replicate(PfCtx *c, int which);        // all-gather the owned slices if needed
void pf_enable_peers(PfCtx *c);                // collective: map every rank's arrays (CUDA IPC)
PfCtx *pf_ctx_of_set(pf_sample_set_t *set, int *which);
PfCtx *pf_ctx_create(pf_t *pf);
void pf_ctx_release(pf_t *pf);
void pf_pull(PfCtx *c, int which);             // device -> host mirror
void pf_push(PfCtx *c, int which);             // host mirror -> device
void pf_after_update(PfCtx *c, int which);     // honours mirror_eager

// ---- stage launchers (device pointers, run on globals().stream) -----------------

struct LaserScanHost
{
  // beams that survive the reference's skip rules, in scan order
  std::vector<double> a, b;      // r*cos(bearing)/scale, r*sin(bearing)/scale
  std::vector<double> range;     // r (beam model)
  std::vector<double> cb, sb;    // cos/sin(bearing) (beam model)
  std::vector<int> beam_ind;     // position in the reference's beam_ind numbering
};

// likelihood_field (model 1) / likelihood_field_prob (model 2); returns the
// unnormalised total weight (pf_sensor_model_fn_t contract).
double run_likelihood_field(PfCtx *c, int which, MapCtx *m, int model,
                            const double laser_pose[3], const LaserScanHost &scan,
                            const std::vector<double> &lut, uint64_t lut_version,
                            const int *mask_or_null, int *agree_counts_or_null,
                            const std::vector<unsigned char> *agree_lut, bool subtract_from_logsum = false);
double run_beam_model(PfCtx *c, int which, MapCtx *m, const double laser_pose[3],
                      const LaserScanHost &scan, double range_max, const double zp[6]);
// pf.c:282-316 on device; returns n_effective, updates pf w_slow/w_fast
void run_sensor_normalise(PfCtx *c, int which, double total);
void run_calc_ranges(MapCtx *m, const double *rays_xya, int n, double max_range, double *out);

void run_motion(PfCtx *c, int which, int model, const double consts[8],
                const double *host_draws_or_null);
void run_resample(PfCtx *c);
void run_cluster_stats(PfCtx *c, int which, bool histogram_ready);
int run_update_converged(PfCtx *c, int which);
void run_cep_stats(PfCtx *c, int which, double out[4]);   // mn, mx, my, mrr
void run_copy_set(PfCtx *c, int from, int to);
void run_philox_gaussian_init(PfCtx *c, int which, const double mean[3],
                              const double cr[9], const double cd[3]);

// A strided view of weights: the w field of an AoS particle array (stride 4
// doubles) or a packed array (stride 1).
struct Weights
{
  const double *p;
  int stride;
#ifdef __CUDACC__
  __device__ __forceinline__ double operator[](int i) const { return p[(size_t)i * stride]; }
#endif
};
inline Weights weights_of(const Particle *set)
{
  Weights w;
  w.p = reinterpret_cast<const double *>(set) + 3;
  w.stride = 4;
  return w;
}
void exact_cumsum_w(PfCtx *c, Weights w, int n, double *d_cum);
// Sharded form: every rank builds the table of its own (tile-aligned) slice from
// the weights it owns, the ranks exchange only per-tile summaries, and the table
// itself is all-gathered.  False: not applicable here (the caller uses the
// replicated form); the result is the same table either way.
bool exact_cumsum_sharded(PfCtx *c, const Particle *set, int n, double *d_cum);

// exact sequential-order cumulative table (pf.c:399-402) of the weights of a set
void exact_cumsum(PfCtx *c, const Particle *set, int n, double *d_cum);
unsigned long long exact_cumsum_fallback_count();
void exact_cumsum_last_tiles(int out[2]);   // {dirty tiles, tiles} of the last parallel table

// Optional stage trace (AMCLB200_TRACE=1): CUDA-event marks inside a call, printed as
// a one-line breakdown when the call ends.  Off by default; costs nothing then.
struct Trace
{
  std::vector<std::pair<const char *, cudaEvent_t> > marks;
  bool on;
  Trace();
  void mark(const char *name);
  void report(const char *what);
};

// NVTX range around one API call (tracing).  Header-only NVTX3: a
// no-op unless a tool (nsys, ncu --nvtx) has injected its library.
struct ApiRange
{
  explicit ApiRange(const char *name);
  ~ApiRange();
};

// Device time of one API call (action / sensor / resample): two events on the
// library's stream, kept in the filter's context.  Nothing waits for them here;
// amclb200_pf_last_timings() resolves them when somebody asks.
struct StageTimer
{
  PfCtx *c;
  int stage;   // 0 action, 1 sensor, 2 resample
  StageTimer(PfCtx *ctx, int which_stage);
  ~StageTimer();
};
void stage_times_resolve(PfCtx *c);   // fills ms_action / ms_sensor / ms_resample

}  // namespace amclb200
