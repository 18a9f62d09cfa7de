/*
 * node_io_oracle.c -- TEST INFRASTRUCTURE (see amcl_oracle.c for the rules).
 * CPU restatement of the steps either side of the AMCL filter that SURVEY.md
 * 8(f) ranks next: map ingest (N4), the published particle cloud (N2) and
 * costmap inflation (N3).  Paths are under /root/reference.
 *
 * Parity status
 *   - orc_image_to_grid: PINNED by the golden vectors of the reference's own
 *     unit tests (map_server/test/test_constants.cpp: testmap.png / .bmp /
 *     testmap2.png contents; map_server/test/utest.cpp:121-152: value counts
 *     of spectrum.png in the three modes) -- tests/golden/map_server_golden.npz,
 *     written by tests/golden/make_map_golden.py.  The loader itself cannot be
 *     compiled here (SDL_image, nav_msgs).
 *   - orc_convert_map: trivially restated from amcl_node.cpp:1002-1009 (the node
 *     cannot be compiled here: roscpp).  Parity unpinned by execution.
 *   - orc_pack_cloud: amcl_node.cpp:1349-1362 with tf2::Quaternion::setRPY
 *     restated from its published formula (ros/geometry2, tf2/LinearMath/
 *     Quaternion.h, the 0.6/0.7 series ROS melodic/noetic ship; tf2 is a
 *     package dependency of amcl, not vendored in /root/reference).  Parity
 *     unpinned by execution.
 *   - orc_inflate: costmap_2d/plugins/inflation_layer.cpp:176-306; see there.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { ORC_TRINARY = 0, ORC_SCALE = 1, ORC_RAW = 2 }; /* map_server/image_loader.h:46 */

/* map_server/src/image_loader.cpp:104-164.  pixels: rowstride bytes per image
 * row, n_channels bytes per pixel, top row first; grid: width*height values,
 * bottom row first. */
void orc_image_to_grid(const unsigned char *pixels, int width, int height, int rowstride,
                       int n_channels, int has_alpha_mask, int negate, double occ_th,
                       double free_th, int mode, signed char *grid)
{
  int avg_channels;
  if (mode == ORC_TRINARY || !has_alpha_mask) /* :106-109 */
    avg_channels = n_channels;
  else
    avg_channels = n_channels - 1;
  for (int j = 0; j < height; j++)
    for (int i = 0; i < width; i++)
    {
      const unsigned char *p = pixels + (size_t)j * rowstride + (size_t)i * n_channels;
      unsigned char value;
      int color_sum = 0;
      for (int k = 0; k < avg_channels; k++) color_sum += p[k];
      double color_avg = color_sum / (double)avg_channels;
      int alpha = n_channels == 1 ? 1 : p[n_channels - 1];
      if (negate) color_avg = 255 - color_avg;
      if (mode == ORC_RAW)
      {
        value = color_avg;
        grid[(size_t)width * (height - j - 1) + i] = value;
        continue;
      }
      double occ = (255 - color_avg) / 255.0;
      if (occ > occ_th)
        value = +100;
      else if (occ < free_th)
        value = 0;
      else if (mode == ORC_TRINARY || alpha < 1.0)
        value = -1;
      else
      {
        double ratio = (occ - free_th) / (occ_th - free_th);
        value = 1 + 98 * ratio;
      }
      grid[(size_t)width * (height - j - 1) + i] = value;
    }
}

/* amcl/src/amcl_node.cpp:1002-1009 */
void orc_convert_map(const signed char *grid, size_t n, int8_t *occ_state)
{
  for (size_t i = 0; i < n; i++)
  {
    if (grid[i] == 0)
      occ_state[i] = -1;
    else if (grid[i] == 100)
      occ_state[i] = +1;
    else
      occ_state[i] = 0;
  }
}

/* amcl/src/amcl_node.cpp:1349-1362; samples are rows (x, y, theta, w).
 * out: 7 doubles per pose (position xyz, orientation xyzw). */
int orc_pack_cloud(const double *samples, int n, int stride, int max_poses, double *out)
{
  int count = 0;
  for (int i = 0; i < n && count < max_poses; i += stride, count++)
  {
    const double roll = 0, pitch = 0, yaw = samples[4 * (size_t)i + 2];
    /* tf2::Quaternion::setRPY */
    double halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
    double cosYaw = cos(halfYaw), sinYaw = sin(halfYaw);
    double cosPitch = cos(halfPitch), sinPitch = sin(halfPitch);
    double cosRoll = cos(halfRoll), sinRoll = sin(halfRoll);
    double *o = out + 7 * (size_t)count;
    o[0] = samples[4 * (size_t)i + 0];
    o[1] = samples[4 * (size_t)i + 1];
    o[2] = 0;
    o[3] = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
    o[4] = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
    o[5] = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
    o[6] = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  }
  return count;
}

/* ======================================================================= */
/* N3: costmap_2d::InflationLayer                                           */
/* ======================================================================= */
/*
 * Parity status: restated from costmap_2d/plugins/inflation_layer.cpp and
 * include/costmap_2d/inflation_layer.h; the layer cannot be compiled here
 * (roscpp, pluginlib, dynamic_reconfigure, boost).  PINNED by the known answers
 * of the reference's own tests, restated on bare grids in
 * tests/test_inflation_oracle.py (costmap_2d/test/inflation_tests.cpp:137-410,
 * with the rostest's cost_scaling_factor = 1: testAdjacentToObstacleCanStillMove,
 * testInflationShouldNotCreateUnknowns, testCostFunctionCorrectness,
 * testInflationOrderCorrectness with its validatePointInflation property,
 * testInflation3's cell counts, and -- chained with orc_image_to_grid on the
 * reference's TenByTen.pgm -- testInflation's 20 lethal / 28 inscribed cells and
 * testInflation2's L-shape).
 */
enum { ORC_NO_INFORMATION = 255, ORC_LETHAL = 254, ORC_INSCRIBED = 253, ORC_FREE = 0 }; /* cost_values.h:42-45 */

/* inflation_layer.h:104-119 */
unsigned char orc_inflation_cost(double distance, double resolution, double inscribed_radius, double weight)
{
  unsigned char cost = 0;
  if (distance == 0)
    cost = ORC_LETHAL;
  else if (distance * resolution <= inscribed_radius)
    cost = ORC_INSCRIBED;
  else
  {
    double euclidean_distance = distance * resolution;
    double factor = exp(-1.0 * weight * (euclidean_distance - inscribed_radius));
    cost = (unsigned char)((ORC_INSCRIBED - 1) * factor);
  }
  return cost;
}

/* costmap_2d/src/costmap_2d.cpp:181-185 */
unsigned int orc_cell_distance(double world_dist, double resolution)
{
  double cells_dist = fmax(0.0, ceil(world_dist / resolution));
  return (unsigned int)cells_dist;
}

typedef struct { unsigned int index, x, y, sx, sy; } orc_celldata; /* inflation_layer.h:56-79 */
typedef struct { orc_celldata *v; size_t n, cap; } orc_bin;

static void orc_bin_push(orc_bin *b, orc_celldata c)
{
  if (b->n == b->cap)
  {
    b->cap = b->cap ? 2 * b->cap : 64;
    b->v = (orc_celldata *)realloc(b->v, b->cap * sizeof(orc_celldata));
  }
  b->v[b->n++] = c;
}

static int orc_cmp_double(const void *a, const void *b)
{
  double x = *(const double *)a, y = *(const double *)b;
  return x < y ? -1 : (x > y ? 1 : 0);
}

/* InflationLayer::updateCosts (inflation_layer.cpp:176-263) with enqueue
 * (:275-291) and computeCaches (:293-340), on a bare row-major cost grid.
 * The std::map<double, vector<CellData>> is an array of bins over the sorted
 * distinct values of the distance cache; a push into a bin the sweep has
 * already left is kept but never processed, exactly like an insertion behind
 * a std::map iterator. */
void orc_inflate(unsigned char *master, int size_x, int size_y, int min_i, int min_j, int max_i,
                 int max_j, double resolution, double inscribed_radius, double inflation_radius,
                 double weight, int inflate_unknown)
{
  const unsigned int R = orc_cell_distance(inflation_radius, resolution); /* :167, :371 */
  if (R == 0) return;                                                       /* :179-180 */
  const unsigned int W = R + 2;
  double *dist = (double *)malloc(sizeof(double) * W * W);
  unsigned char *cost = (unsigned char *)malloc(W * W);
  double *keys = (double *)malloc(sizeof(double) * W * W);
  for (unsigned int i = 0; i < W; i++)
    for (unsigned int j = 0; j < W; j++)
    {
      dist[i * W + j] = hypot(i, j);                                        /* :313 */
      cost[i * W + j] = orc_inflation_cost(dist[i * W + j], resolution, inscribed_radius, weight);
      keys[i * W + j] = dist[i * W + j];
    }
  qsort(keys, W * W, sizeof(double), orc_cmp_double);
  size_t nkeys = 0;
  for (size_t k = 0; k < (size_t)W * W; k++)
    if (nkeys == 0 || keys[k] != keys[nkeys - 1]) keys[nkeys++] = keys[k];
  int *bin_of = (int *)malloc(sizeof(int) * W * W);
  for (size_t k = 0; k < (size_t)W * W; k++)
  {
    size_t lo = 0, hi = nkeys;
    while (lo + 1 < hi)
    {
      size_t mid = (lo + hi) / 2;
      if (keys[mid] <= dist[k]) lo = mid; else hi = mid;
    }
    bin_of[k] = (int)lo;
  }
  orc_bin *bins = (orc_bin *)calloc(nkeys, sizeof(orc_bin));
  unsigned char *seen = (unsigned char *)calloc((size_t)size_x * size_y, 1);  /* :200 */

  min_i -= (int)R; min_j -= (int)R; max_i += (int)R; max_j += (int)R;         /* :206-209 */
  if (min_i < 0) min_i = 0;
  if (min_j < 0) min_j = 0;
  if (max_i > size_x) max_i = size_x;
  if (max_j > size_y) max_j = size_y;

  for (int j = min_j; j < max_j; j++)                                         /* :220-232 */
    for (int i = min_i; i < max_i; i++)
    {
      unsigned int index = (unsigned int)(j * size_x + i);
      if (master[index] == ORC_LETHAL)
      {
        orc_celldata c = {index, (unsigned)i, (unsigned)j, (unsigned)i, (unsigned)j};
        orc_bin_push(&bins[0], c);
      }
    }

#define ORC_ENQUEUE(IDX, MX, MY)                                                          \
  do                                                                                      \
  {                                                                                       \
    if (!seen[IDX])                                                                       \
    {                                                                                     \
      unsigned int dx_ = (unsigned)abs((int)(MX) - (int)sx), dy_ = (unsigned)abs((int)(MY) - (int)sy); \
      double distance = dist[dx_ * W + dy_];                                              \
      if (!(distance > R))                                                                \
      {                                                                                   \
        orc_celldata c_ = {(IDX), (MX), (MY), sx, sy};                                    \
        orc_bin_push(&bins[bin_of[dx_ * W + dy_]], c_);                                   \
      }                                                                                   \
    }                                                                                     \
  } while (0)

  for (size_t b = 0; b < nkeys; b++)                                          /* :237-272 */
    for (size_t q = 0; q < bins[b].n; q++)
    {
      const orc_celldata cell = bins[b].v[q];
      const unsigned int index = cell.index;
      if (seen[index]) continue;
      seen[index] = 1;
      const unsigned int mx = cell.x, my = cell.y, sx = cell.sx, sy = cell.sy;
      const unsigned int dx = (unsigned)abs((int)mx - (int)sx), dy = (unsigned)abs((int)my - (int)sy);
      const unsigned char c = cost[dx * W + dy];
      const unsigned char old_cost = master[index];
      if (old_cost == ORC_NO_INFORMATION && (inflate_unknown ? (c > ORC_FREE) : (c >= ORC_INSCRIBED)))
        master[index] = c;
      else
        master[index] = old_cost > c ? old_cost : c;
      if (mx > 0) ORC_ENQUEUE(index - 1, mx - 1, my);
      if (my > 0) ORC_ENQUEUE(index - size_x, mx, my - 1);
      if (mx < (unsigned)size_x - 1) ORC_ENQUEUE(index + 1, mx + 1, my);
      if (my < (unsigned)size_y - 1) ORC_ENQUEUE(index + size_x, mx, my + 1);
    }
#undef ORC_ENQUEUE

  for (size_t b = 0; b < nkeys; b++) free(bins[b].v);
  free(bins); free(seen); free(bin_of); free(keys); free(cost); free(dist);
}

/* The same update with the distance taken to the NEAREST lethal cell of the
 * window by exhaustive search: what the sweep computes whenever its
 * source-carrying propagation is exact (the property the reference's
 * validatePointInflation asserts, inflation_tests.cpp:73-135). */
void orc_inflate_exact(unsigned char *master, int size_x, int size_y, int min_i, int min_j,
                       int max_i, int max_j, double resolution, double inscribed_radius,
                       double inflation_radius, double weight, int inflate_unknown)
{
  const int R = (int)orc_cell_distance(inflation_radius, resolution);
  if (R == 0) return;
  min_i -= R; min_j -= R; max_i += R; max_j += R;
  if (min_i < 0) min_i = 0;
  if (min_j < 0) min_j = 0;
  if (max_i > size_x) max_i = size_x;
  if (max_j > size_y) max_j = size_y;
  unsigned char *in = (unsigned char *)malloc((size_t)size_x * size_y);
  memcpy(in, master, (size_t)size_x * size_y);
  for (int y = 0; y < size_y; y++)
    for (int x = 0; x < size_x; x++)
    {
      int best = -1, bdx = 0, bdy = 0;
      for (int dy = -R; dy <= R; dy++)
      {
        const int j = y + dy;
        if (j < min_j || j >= max_j) continue;
        for (int dx = -R; dx <= R; dx++)
        {
          const int i = x + dx;
          if (i < min_i || i >= max_i) continue;
          if (in[(size_t)j * size_x + i] != ORC_LETHAL) continue;
          const int d2 = dx * dx + dy * dy;
          if (best < 0 || d2 < best) { best = d2; bdx = abs(dx); bdy = abs(dy); }
        }
      }
      if (best < 0) continue;
      const double distance = hypot(bdx, bdy);
      if (distance > R) continue;
      const unsigned char c = orc_inflation_cost(distance, resolution, inscribed_radius, weight);
      const unsigned char old_cost = in[(size_t)y * size_x + x];
      unsigned char *out = &master[(size_t)y * size_x + x];
      if (old_cost == ORC_NO_INFORMATION && (inflate_unknown ? (c > ORC_FREE) : (c >= ORC_INSCRIBED)))
        *out = c;
      else
        *out = old_cost > c ? old_cost : c;
    }
  free(in);
}
