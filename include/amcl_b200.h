/*
 * amcl_b200.h -- everything libamcl_b200.so exports, in one place.
 *
 * Part 1 is the drop-in boundary: the reference's own C API, re-declared in
 * amcl_b200/amcl_abi.h with the reference file:line each entry point replaces
 * (pf.h:147-203, map.h:88-111, pf_vector.h, pf_pdf.h, pf_kdtree.h, eig3.h), and
 * the C++ classes in amcl_b200/amcl_sensors.hpp (amcl_sensor.h, amcl_odom.h,
 * amcl_laser.h).  amcl_node links against those unchanged.
 *
 * Part 2 (amcl_c_*) is a plain-C facade over the C++ classes for FFI callers
 * (ctypes, cgo, JNI ...) that cannot construct C++ objects.  Each function
 * forwards to exactly one reference method; see navigation_b200/csrc/class_capi.cpp.
 *
 * Part 3 (amclb200_*) has no counterpart in the reference.  It exposes what a
 * host-only API cannot express: where the particles live (HBM), when the host
 * mirror of pf_sample_set_t::samples is refreshed, which random stream is used,
 * and integer results (source indices, labels) that the reference only holds
 * implicitly.  Nothing in Part 3 has to be called by a drop-in user.
 *
 * All pointers are HOST pointers unless a name says "device".  No function
 * here takes or returns a torch / C++ library type.
 */
#ifndef AMCL_B200_H
#define AMCL_B200_H

#include <stdint.h>
#include "amcl_b200/amcl_abi.h"   /* Part 1 (C half) */

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- Part 2 -- */
/* maps: what AmclNode::convertMap builds (amcl_node.cpp:987-1012) */
map_t *amcl_c_map_from_occupancy(const signed char *occ, int size_x, int size_y,
                                 double scale, double origin_x, double origin_y);
void amcl_c_map_get_occ_dist(map_t *map, double *out);
void amcl_c_map_set_occ_dist(map_t *map, const double *in, double max_occ_dist);
/* AmclNode::uniformPoseGenerator + its free-space index (amcl_node.cpp:901-908,1050-1060) */
int amcl_c_freespace_build(map_t *map);
pf_vector_t amcl_c_uniform_pose(void *map);
pf_init_model_fn_t amcl_c_uniform_pose_fn(void);
/* AMCLOdom (amcl_odom.h:57-93) */
void *amcl_c_odom_new(void);
void amcl_c_odom_delete(void *odom);
void amcl_c_odom_set_model(void *odom, int type, double a1, double a2, double a3, double a4, double a5);
int amcl_c_odom_update_action(void *odom, pf_t *pf, const double pose[3], const double delta[3]);
/* AMCLLaser (amcl_laser.h:59-151) */
void *amcl_c_laser_new(int max_beams, map_t *map);
void amcl_c_laser_delete(void *laser);
void *amcl_c_laser_clone(void *laser);
void amcl_c_laser_set_pose(void *laser, double x, double y, double a);
void amcl_c_laser_set_model_beam(void *laser, double z_hit, double z_short, double z_max,
                                 double z_rand, double sigma_hit, double lambda_short,
                                 double chi_outlier);
void amcl_c_laser_set_model_lf(void *laser, double z_hit, double z_rand, double sigma_hit,
                               double max_occ_dist);
void amcl_c_laser_set_model_lfp(void *laser, double z_hit, double z_rand, double sigma_hit,
                                double max_occ_dist, int do_beamskip, double beam_skip_distance,
                                double beam_skip_threshold, double beam_skip_error_threshold);
int amcl_c_laser_update_sensor(void *laser, pf_t *pf, const double *ranges_rb, int range_count,
                               double range_max);
pf_sample_set_t *amcl_c_pf_current_set(pf_t *pf);
int amcl_c_sizeof_pf(void);
int amcl_c_sizeof_laser(void);
int amcl_c_sizeof_odom(void);

/* ---------------------------------------------------------------- Part 3 -- */

/* "amcl_b200 <major>.<minor> sm_100a" */
const char *amclb200_version(void);

/* CUDA stream (a cudaStream_t passed as void*) all later launches go to.
 * Default: the legacy default stream. */
void amclb200_set_stream(void *cuda_stream);

/* Random streams.  mode 0 = Philox4x32-10 on the device, keyed by `seed`
 * (production default).  mode 1 = libc drand48() consumed on the host in the
 * reference's exact order (parity with the reference for a given srand48). */
void amclb200_set_rng(int mode, uint64_t seed);

/* on = 1 (default): pf_alloc seeds drand48 from the clock and pf_init re-seeds it
 * from a per-process counter, exactly as the reference does (pf.c:55, pf_pdf.c:68).
 * on = 0: the library never seeds; a caller that seeds the stream itself gets
 * reproducible runs. */
void amclb200_set_reseed(int on);
int amclb200_get_reseed(void);

/* Host mirror policy for pf_sample_set_t::samples.  eager = 1 (default): every
 * pf_update_* / UpdateAction / UpdateSensor call ends with a device-to-host
 * copy of the current set, so callers that read samples[] directly
 * (amcl_node.cpp:1342-1361) keep working.  eager = 0: the mirror is refreshed
 * only by amclb200_pf_pull_samples. */
void amclb200_set_mirror(pf_t *pf, int eager);
void amclb200_pf_pull_samples(pf_t *pf);   /* device -> samples[] of the current set */
void amclb200_pf_push_samples(pf_t *pf);   /* samples[] (+ sample_count) -> device */

/* Device address of set `set_index`'s particle array: max_samples records of
 * 32 bytes (x, y, theta, weight as fp64), the layout of pf_sample_t. */
void *amclb200_pf_device_samples(pf_t *pf, int set_index);

/* Re-encode the distances currently held in map->cells[].occ_dist (and
 * map->max_occ_dist) into the device distance map, e.g. to evaluate the laser
 * models on the reference wavefront's output.  Returns 0; or -1, with one line
 * on stderr and the device map unchanged, when a value is not
 * sqrt(di^2 + dj^2) * scale within the radius, or the radius exceeds 254 cells. */
int amclb200_map_push_distances(map_t *map);

/* Source index of every sample of the current set as chosen by the last
 * pf_update_resample (-1 = injected random pose); returns the count written. */
int amclb200_pf_resample_sources(pf_t *pf, int *out, int capacity);

/* Cluster label of every sample of the current set (this library's canonical
 * numbering: clusters in order of first appearance); returns the count. */
int amclb200_pf_cluster_labels(pf_t *pf, int *out, int capacity);

/* The sequential-order cumulative weight table pf_update_resample searches
 * (pf.c:399-402), recomputed for the current set: n+1 doubles. */
int amclb200_pf_cumulative_table(pf_t *pf, double *out, int capacity);

/* Device milliseconds (CUDA events) of the last action / sensor / resample call. */
void amclb200_pf_last_timings(pf_t *pf, float out_ms[3]);

/* Device milliseconds the distance-transform kernels of the last map_update_cspace took. */
double amclb200_map_last_edt_ms(map_t *map);

/* Batched map_calc_range on the device: n rays (ox, oy, oa) -> ranges. */
void amclb200_map_calc_ranges(map_t *map, const double *rays_xya, int n, double max_range,
                              double *out_ranges);

/* Device-to-device load of n particles (32-byte records) into the current set;
 * sample_count becomes n.  For benchmarks that restart each step from one cloud. */
void amclb200_pf_load_device(pf_t *pf, const void *d_particles, int n);

/* The hypothesis amcl_node publishes (amcl_node.cpp:1369-1394): label, weight,
 * mean and covariance of the heaviest cluster of the current set; -1 if none. */
int amclb200_pf_best_cluster(pf_t *pf, double *weight, pf_vector_t *mean, pf_matrix_t *cov);

/* Host mirror of the map.  eager (default 1): map_update_cspace re-reads
 * cells[].occ_state and writes cells[].occ_dist on every call, as the reference
 * does.  eager 0: the device copy is authoritative -- the cells are read once
 * (not at all for a map made by amclb200_map_from_image) and cells[].occ_dist
 * is written only by amclb200_map_pull_distances.  The laser models never read
 * the host cells either way. */
void amclb200_map_set_mirror(map_t *map, int eager);
void amclb200_map_pull_distances(map_t *map);

/* ---- the data formats either side of the filter (N2, N4) -----------
 *
 * Map ingest: map_server's image -> occupancy rule
 * (map_server/src/image_loader.cpp:57-165: channel mean, negate, occupied / free
 * thresholds, modes 0 = TRINARY, 1 = SCALE, 2 = RAW, rows flipped so that cell
 * (0,0) is the lower-left pixel) fused with AmclNode::convertMap
 * (amcl/src/amcl_node.cpp:987-1012: 0 -> free, 100 -> occupied, anything else
 * unknown; origin moved to the map centre).  `pixels` are DECODED image bytes
 * as SDL hands them over (rowstride = pitch, n_channels = BytesPerPixel,
 * has_alpha_mask = format->Amask != 0); origin is the yaml's [x, y, yaw].
 * Returns a map_t (free with map_free) whose occupancy is already resident on
 * the device; grid_out, if not NULL, receives the width*height OccupancyGrid
 * bytes map_server would have published. */
map_t *amclb200_map_from_image(const unsigned char *pixels, int width, int height, int rowstride,
                               int n_channels, int has_alpha_mask, double resolution,
                               const double origin[3], int negate, double occ_th, double free_th,
                               int mode, signed char *grid_out);

/* The particle cloud AmclNode publishes (amcl_node.cpp:1349-1362): every
 * `stride`-th sample of the current set as 7 doubles -- position x, y, 0 and the
 * quaternion x, y, z, w of tf2::Quaternion::setRPY(0, 0, yaw) -- packed on the
 * device; only these bytes are copied to `out` (host).  Returns the number of
 * poses written (<= max_poses). */
int amclb200_pf_pack_cloud(pf_t *pf, int stride, int max_poses, double *out);

/* Costmap inflation (N3): costmap_2d::InflationLayer::updateCosts
 * (costmap_2d/plugins/inflation_layer.cpp:176-263) on a caller-owned cost grid
 * (row-major unsigned char, x fastest, cost_values.h encoding), in place.  The
 * window [min_i, max_i) x [min_j, max_j) is the layer's update window BEFORE it
 * is grown by the inflation radius; resolution, inscribed_radius,
 * inflation_radius and cost_scaling_factor are the layer's resolution_,
 * inscribed_radius_, inflation_radius_ and weight_.  Every cell within the cell
 * radius of a LETHAL_OBSTACLE cell of the grown window gets
 * max(old, computeCost(distance to the nearest such cell)), with the
 * NO_INFORMATION / inflate_unknown rule of :253-256.  `costs` is a host pointer
 * (only the rectangle the update can touch is copied each way) or, with on_device
 * != 0, a device pointer.  Returns the number of cells visited (0 when the cell
 * radius is 0, as the reference returns early). */
int amclb200_inflate_costs(unsigned char *costs, int on_device, int size_x, int size_y, int min_i,
                           int min_j, int max_i, int max_j, double resolution,
                           double inscribed_radius, double inflation_radius,
                           double cost_scaling_factor, int inflate_unknown);

/* Device time (ms) of the three kernels of the last amclb200_inflate_costs call. */
float amclb200_inflate_last_ms(void);

/* ---- several GPUs, one process each -------------------------------------------
 * Collectives are supplied by the host program (torch.distributed / NCCL in
 * bench.py).  Both act IN PLACE on DEVICE buffers and must be ordered with the
 * library's stream.  dtype: 0 = fp64, 1 = int32, 2 = int64; op: 0 = sum, 1 = min, 2 = max.
 * all-gather: rank r's block of bytes_per_rank bytes sits at offset
 * r * bytes_per_rank of d_buf on entry; all blocks are present on return. */
typedef void (*amclb200_allreduce_fn)(void *d_buf, int count, int dtype, int op, void *user);
typedef void (*amclb200_allgather_fn)(void *d_buf, unsigned long long bytes_per_rank, void *user);

/* Makes this filter rank `rank` of `world`.  Every rank allocates the filter
 * with the same (global) max_samples and issues the same calls; each keeps only
 * its contiguous slice of the particle arrays current and the library inserts
 * the exchanges the algorithm needs (weight totals, the source set and the bin
 * keys before resampling, per-cluster sums).  Results equal the one-GPU run. */
void amclb200_pf_set_shard(pf_t *pf, int rank, int world, amclb200_allreduce_fn allreduce,
                           amclb200_allgather_fn allgather, void *user);
/* The same, with the exchanges on an NCCL communicator the LIBRARY owns: no
 * callbacks, the collectives are enqueued on the library's stream from C++ (this
 * is the path bench.py uses on GPUs; the callback form stays for programs that
 * bring their own transport, e.g. gloo on host buffers).  Rank 0 obtains an id
 * with amclb200_nccl_unique_id (128 bytes, an ncclUniqueId) and distributes it
 * by any host-side means; then every rank calls amclb200_pf_set_shard_nccl
 * (collective).  libnccl.so.2 is opened on first use only.  Returns 0 / -1. */
int amclb200_nccl_unique_id(unsigned char out[128]);
void amclb200_pf_set_shard_nccl(pf_t *pf, int rank, int world, const unsigned char id[128]);
/* Collective, optional, once after amclb200_pf_set_shard: map every rank's particle
 * arrays into this process (CUDA IPC).  pf_update_resample then all-gathers only
 * the 8-byte weights and reads the poses it picks from their owners' memory over
 * NVLink inside the pick kernel, instead of all-gathering the whole 32-byte set. */
void amclb200_pf_enable_peer_access(pf_t *pf);
/* Device-to-device load of this rank's slice of an n_global-particle cloud. */
void amclb200_pf_load_device_shard(pf_t *pf, const void *d_slice, int n_global);
/* out = [lo, hi): the slice of a set of n particles this rank owns. */
void amclb200_pf_shard_range(pf_t *pf, int n, int out[2]);

/* Kernels launched by this library so far (all filters). */
unsigned long long amclb200_launch_count(void);

/* Exact cumulative table: number of times the tile-parallel form failed its
 * recurrence check and the single-thread walk was used instead (expected 0);
 * and a switch to force the single-thread walk (on = 0). */
unsigned long long amclb200_cumsum_fallbacks(void);
void amclb200_set_parallel_cumsum(int on);
/* out = {tiles that had to be walked element by element, tiles} of the last table. */
void amclb200_cumsum_last_tiles(int out[2]);

/* out[0] = milliseconds of the laser-model kernel of the last UpdateSensor,
 * out[1] = particle-beam evaluations it performed. */
void amclb200_pf_sensor_kernel_stats(pf_t *pf, double out[2]);

/* Measured rate (gathers/s) of uncorrelated 2-byte gathers over a table of
 * `table_bytes`: the L2-gather roofline of the laser kernels. */
double amclb200_measure_gather_rate(unsigned long long table_bytes, int reps);

/* Distance transform for callers that must not allocate device memory (a
 * real-time loop or certified runtime with a no-allocation-after-start-up rule,
 * or a process whose device memory belongs entirely to another allocator).
 * It has no counterpart in the reference.  The caller owns both buffers:
 * d_occ holds size_x*size_y occupancy states (-1 / 0 / +1, x fastest) and
 * d_codes receives, row-major, the squared distance in cells to the nearest
 * occupied cell, or cell_radius^2 + 1 where that exceeds cell_radius^2
 * (occ_dist = sqrt(code)*scale, resp. max_occ_dist).  The only scratch is a
 * fixed 4 MiB array inside the library, held by every GPU that loads it, so:
 * at most amclb200_edt_static_max_cells() = 4 Mi cells and radius <= 254, else
 * -2 is returned and nothing is written; calls that overlap on one GPU race on
 * that array (serialise them); the first call on each GPU resolves the array's
 * address and should happen at start-up, before any stream capture.  No
 * cudaMalloc / cudaFree or other allocating call is made.  Returns 0, or
 * -1 bad argument, -2 over the limit, -3 CUDA error. */
int amclb200_edt_codes_static(const signed char *d_occ, int size_x, int size_y, int cell_radius,
                              unsigned short *d_codes, void *cuda_stream);
unsigned long long amclb200_edt_static_max_cells(void);

#ifdef __cplusplus
}
#endif

#endif /* AMCL_B200_H */
