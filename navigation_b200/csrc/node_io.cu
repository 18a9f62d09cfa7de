// node_io.cu -- the data formats either side of the filter.
//
// N4, map ingest: what map_server does to an image
// (/root/reference/map_server/src/image_loader.cpp:57-165) followed by what
// AmclNode::convertMap does to the resulting occupancy grid
// (/root/reference/amcl/src/amcl_node.cpp:987-1012), fused into one kernel that
// leaves the occupancy states where the distance transform and the ray caster
// read them.  The caller hands in decoded pixels (decoding PNG/BMP files is
// SDL_image's job in the reference and stays outside this library).
//
// N2, particle cloud for the caller: the PoseArray AmclNode publishes after
// every update (amcl_node.cpp:1349-1362) packed on the device, optionally
// decimated, so only the packed poses cross PCIe instead of the whole sample
// array.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "amcl_b200.h"
#include "context.hpp"

namespace amclb200
{

enum { MAP_MODE_TRINARY = 0, MAP_MODE_SCALE = 1, MAP_MODE_RAW = 2 };  // map_server/image_loader.h:46

struct IngestArgs
{
  const unsigned char *pixels;
  int width, height, rowstride, n_channels, avg_channels;
  int negate, mode;
  double occ_th, free_th;
  signed char *grid;   // nav_msgs/OccupancyGrid data, row 0 = bottom image row
  int8_t *occ;         // occ_state: -1 free, 0 unknown, +1 occupied
};

// One thread per cell of the OUTPUT grid (x fastest), reading image row
// height-1-j: image_loader.cpp:139,163 flips the rows.
__global__ void ingest_image_kernel(const IngestArgs a)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;
  if (i >= a.width) return;
  const unsigned char *p = a.pixels + (size_t)(a.height - 1 - j) * a.rowstride + (size_t)i * a.n_channels;
  int color_sum = 0;
  for (int k = 0; k < a.avg_channels; k++) color_sum += p[k];
  double color_avg = __ddiv_rn((double)color_sum, (double)a.avg_channels);   // :119-122
  const int alpha = a.n_channels == 1 ? 1 : p[a.n_channels - 1];            // :124-127
  if (a.negate) color_avg = 255 - color_avg;                                 // :129-130
  unsigned char value;
  if (a.mode == MAP_MODE_RAW)
    value = (unsigned char)color_avg;                                        // :132-136
  else
  {
    const double occ = __ddiv_rn(255 - color_avg, 255.0);                    // :141
    if (occ > a.occ_th)
      value = 100;
    else if (occ < a.free_th)
      value = 0;
    else if (a.mode == MAP_MODE_TRINARY || alpha < 1.0)
      value = (unsigned char)-1;
    else
    {
      const double ratio = __ddiv_rn(occ - a.free_th, a.occ_th - a.free_th);
      value = (unsigned char)(1 + 98 * ratio);                               // :155-158
    }
  }
  const signed char cell = (signed char)value;
  const size_t idx = (size_t)j * a.width + i;
  a.grid[idx] = cell;
  a.occ[idx] = cell == 0 ? -1 : (cell == 100 ? 1 : 0);                       // amcl_node.cpp:1002-1009
}

// PoseArray entry for a sample: position (x, y, 0), orientation
// tf2::Quaternion::setRPY(0, 0, yaw) (geometry2 tf2/LinearMath/Quaternion.h,
// not vendored in the reference tree) written out term by term so that the
// signs of the zero components come out as tf2's do.
// own_lo / own_hi: the samples this rank holds (a sharded filter packs only those; the
// entries of the others stay zero and a SUM all-reduce completes the cloud).
__global__ void pack_cloud_kernel(const Particle *__restrict__ set, int n, int stride, int count,
                                  double *__restrict__ out, int own_lo, int own_hi)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= count) return;
  const long long idx = (long long)k * stride;
  if (idx < own_lo || idx >= own_hi) return;
  const Particle s = load_particle(set + (size_t)idx);
  const double half_yaw = s.th * 0.5, half_pitch = 0.0 * 0.5, half_roll = 0.0 * 0.5;
  double sy, cy;
  sincos(half_yaw, &sy, &cy);
  const double cp = cos(half_pitch), sp = sin(half_pitch), cr = cos(half_roll), sr = sin(half_roll);
  double *o = out + (size_t)k * 7;
  o[0] = s.x;
  o[1] = s.y;
  o[2] = 0.0;
  o[3] = sr * cp * cy - cr * sp * sy;
  o[4] = cr * sp * cy + sr * cp * sy;
  o[5] = cr * cp * sy - sr * sp * cy;
  o[6] = cr * cp * cy + sr * sp * sy;
}

}  // namespace amclb200

using namespace amclb200;

extern "C" {

map_t *amclb200_map_from_image(const unsigned char *pixels, int width, int height, int rowstride,
                               int n_channels, int has_alpha_mask, double resolution,
                               const double origin[3], int negate, double occ_th, double free_th,
                               int mode, signed char *grid_out)
{
  if (!pixels || width <= 0 || height <= 0 || n_channels < 1 || rowstride < width * n_channels)
    AMCL_FATAL("amclb200_map_from_image: bad image (%d x %d, %d channels, stride %d)", width, height,
               n_channels, rowstride);
  if (height > 65535) AMCL_FATAL("maps taller than 65535 rows are not supported (%d)", height);
  // convertMap: the map_t origin is the world position of the map centre
  map_t *map = map_alloc();
  map->size_x = width;
  map->size_y = height;
  map->scale = resolution;
  map->origin_x = origin[0] + (width / 2) * resolution;
  map->origin_y = origin[1] + (height / 2) * resolution;
  const size_t cells = (size_t)width * height;
  map->cells = (map_cell_t *)malloc(sizeof(map_cell_t) * cells);
  if (!map->cells) AMCL_FATAL("out of host memory for %zu map cells", cells);
  MapCtx *m = map_ctx(map, true);

  cudaStream_t st = globals().stream;
  unsigned char *d_pix = nullptr;
  signed char *d_grid = nullptr;
  const size_t pix_bytes = (size_t)rowstride * height;
  AMCL_CUDA(cudaMalloc(&d_pix, pix_bytes));
  AMCL_CUDA(cudaMalloc(&d_grid, cells));
  AMCL_CUDA(cudaMemcpyAsync(d_pix, pixels, pix_bytes, cudaMemcpyHostToDevice, st));
  IngestArgs a;
  a.pixels = d_pix;
  a.width = width;
  a.height = height;
  a.rowstride = rowstride;
  a.n_channels = n_channels;
  // image_loader.cpp:106-109: trinary mode averages the alpha channel in as well
  a.avg_channels = (mode == MAP_MODE_TRINARY || !has_alpha_mask) ? n_channels : n_channels - 1;
  if (a.avg_channels < 1) a.avg_channels = 1;
  a.negate = negate;
  a.mode = mode;
  a.occ_th = occ_th;
  a.free_th = free_th;
  a.grid = d_grid;
  a.occ = m->d_occ;
  dim3 grid((width + 255) / 256, height);
  ingest_image_kernel<<<grid, 256, 0, st>>>(a);
  AMCL_LAUNCH_CHECK();
  map_finish_occupancy_on_device(m);   // ray-casting bit tiles; marks the occupancy as uploaded

  std::vector<signed char> host_grid;
  signed char *g = grid_out;
  if (!g)
  {
    host_grid.resize(cells);
    g = host_grid.data();
  }
  AMCL_CUDA(cudaMemcpyAsync(g, d_grid, cells, cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  parallel_rows(height, [&](int j0, int j1) {
    for (int j = j0; j < j1; j++)
    {
      const signed char *row = g + (size_t)j * width;
      map_cell_t *crow = map->cells + (size_t)j * width;
      for (int i = 0; i < width; i++)
      {
        crow[i].occ_state = row[i] == 0 ? -1 : (row[i] == 100 ? +1 : 0);
        crow[i].occ_dist = 0.0;
      }
    }
  });
  cudaFree(d_pix);
  cudaFree(d_grid);
  return map;
}

int amclb200_pf_pack_cloud(pf_t *pf, int stride, int max_poses, double *out)
{
  PfCtx *c = pf_ctx(pf);
  const int which = pf->current_set;
  const int n = pf->sets[which].sample_count;
  if (stride < 1) stride = 1;
  int count = (n + stride - 1) / stride;
  if (count > max_poses) count = max_poses;
  if (count <= 0 || !out) return 0;
  // A sharded filter does not gather the whole set for a decimated cloud: every rank
  // packs the poses it holds, the rest of its buffer is zero, one SUM all-reduce of
  // the packed cloud (56 B per pose) gives everybody all of it.
  const bool sliced = c->shard.on() && !c->replicated[which];
  cudaStream_t st = globals().stream;
  const size_t bytes = sizeof(double) * 7 * (size_t)count;
  if (bytes > c->cloud_cap)
  {
    if (c->d_cloud) cudaFree(c->d_cloud);
    c->cloud_cap = bytes + 7 * sizeof(double) * 1024;
    AMCL_CUDA(cudaMalloc(&c->d_cloud, c->cloud_cap));
  }
  if (sliced) AMCL_CUDA(cudaMemsetAsync(c->d_cloud, 0, bytes, st));
  pack_cloud_kernel<<<(count + 255) / 256, 256, 0, st>>>(c->d_set[which], n, stride, count, c->d_cloud,
                                                         sliced ? c->shard.lo(n) : 0, sliced ? c->shard.hi(n) : n);
  AMCL_LAUNCH_CHECK();
  if (sliced)
    for (size_t off = 0; off < 7 * (size_t)count; off += (size_t)1 << 28)
    {
      const size_t len = 7 * (size_t)count - off < ((size_t)1 << 28) ? 7 * (size_t)count - off : (size_t)1 << 28;
      c->shard.allreduce(c->d_cloud + off, (int)len, SHARD_F64, SHARD_SUM);
    }
  AMCL_CUDA(cudaMemcpyAsync(out, c->d_cloud, bytes, cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  return count;
}

}  // extern "C"
