// inflation_dev.cu -- costmap_2d::InflationLayer::updateCosts on the device
// (/root/reference/costmap_2d/plugins/inflation_layer.cpp:176-340,
// include/costmap_2d/inflation_layer.h:104-167).
//
// The reference sweeps outwards from the lethal cells of the (enlarged) update
// window through 4-neighbours, carrying each cell's source obstacle, in bins of
// increasing distance, and writes cost(distance to the carried source).  Its
// own tests demand that this equals the cost of the NEAREST lethal cell
// (validatePointInflation, costmap_2d/test/inflation_tests.cpp:73-135); the
// sweep misses that in a few cells per 10^5 (the carried source is not always
// the nearest one).  Here the nearest lethal cell is found exactly, with the
// separable bounded distance transform of map_dev.cu restricted to the window:
//   (1) infl_pack_rows : lethal cells of the seed window as one bit per cell
//   (2) infl_row_dist  : distance along the row to the nearest set bit (clz/ffs)
//   (3) infl_col_apply : d2 = min_dj dj^2 + g^2 over the column, then the
//                        reference's update rule with cost looked up by d2.
// Only the rows and columns the update can touch are visited.
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "amcl_b200.h"
#include "context.hpp"
#include "edt_rows.cuh"
#include "inflation_kernels.cuh"

namespace amclb200
{

// One warp per 64-cell word of a seed row.
__global__ void infl_pack_rows(const unsigned char *__restrict__ master, int pitch, InflBox b,
                               int words, int pad, unsigned long long *__restrict__ bits)
{
  const int lane = threadIdx.x & 31;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int stride = words + 2 * pad;
  const int rows = b.e_j1 - b.e_j0;
  const int r = (int)(warp / stride), w = (int)(warp % stride) - pad;
  if (r >= rows) return;
  unsigned long long word = 0ull;
  if (w >= 0 && w < words)
  {
    const unsigned char *row = master + (size_t)(b.e_j0 + r) * pitch;
    const int i0 = b.a_i0 + w * 64 + lane, i1 = i0 + 32;
    const unsigned lo = __ballot_sync(kFullMask, i0 >= b.e_i0 && i0 < b.e_i1 && row[i0] == COST_LETHAL);
    const unsigned hi = __ballot_sync(kFullMask, i1 >= b.e_i0 && i1 < b.e_i1 && row[i1] == COST_LETHAL);
    word = (unsigned long long)lo | ((unsigned long long)hi << 32);
  }
  if (lane == 0) bits[(size_t)r * stride + (w + pad)] = word;
}

// g(c, r) = distance along seed row r from column a_i0 + c to the nearest lethal cell, capped at R + 1.
__global__ void infl_row_dist(const unsigned long long *__restrict__ bits, int aw, int words, int pad,
                              int R, unsigned char *__restrict__ g)
{
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= aw) return;
  const int stride = words + 2 * pad;
  const unsigned long long *A = bits + (size_t)r * stride + pad;
  const int best = row_distance(A, c, R);
  g[(size_t)r * aw + c] = (unsigned char)best;
}

// One thread per reachable cell.
__global__ void infl_col_apply(const unsigned char *__restrict__ g, int aw, InflBox b, int R,
                               const unsigned char *__restrict__ cost_of_d2, int inflate_unknown,
                               int pitch, unsigned char *master)
{
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = b.a_j0 + blockIdx.y;
  if (c >= aw) return;
  const int R2 = R * R;
  const int rows = b.e_j1 - b.e_j0;
  int best = R2 + 1;
  for (int d = 0; d <= R; d++)
  {
    const int dd = d * d;
    if (dd >= best) break;
    const int ra = j - d - b.e_j0, rb = j + d - b.e_j0;
    if ((unsigned)ra < (unsigned)rows)
    {
      const int gg = g[(size_t)ra * aw + c];
      best = min(best, dd + gg * gg);
    }
    if ((unsigned)rb < (unsigned)rows)
    {
      const int gg = g[(size_t)rb * aw + c];
      best = min(best, dd + gg * gg);
    }
  }
  if (best > R2) return;                       // enqueue: distance > cell_inflation_radius_ (:285-286)
  const unsigned char cost = cost_of_d2[best];
  unsigned char *cell = master + (size_t)j * pitch + (b.a_i0 + c);
  const unsigned char old_cost = *cell;
  // inflation_layer.cpp:253-256
  if (old_cost == COST_NO_INFORMATION &&
      (inflate_unknown ? (cost > COST_FREE) : (cost >= COST_INSCRIBED)))
    *cell = cost;
  else
    *cell = old_cost > cost ? old_cost : cost;
}

// inflation_layer.h:104-119
static unsigned char compute_cost(double distance, double resolution, double inscribed_radius, double weight)
{
  unsigned char cost = 0;
  if (distance == 0)
    cost = COST_LETHAL;
  else if (distance * resolution <= inscribed_radius)
    cost = COST_INSCRIBED;
  else
  {
    double euclidean_distance = distance * resolution;
    double factor = exp(-1.0 * weight * (euclidean_distance - inscribed_radius));
    cost = (unsigned char)((COST_INSCRIBED - 1) * factor);
  }
  return cost;
}

struct InflScratch
{
  unsigned long long *d_bits = nullptr;
  size_t bits_cap = 0;
  unsigned char *d_g = nullptr;      // row distances, one byte per cell (search kernels)
  size_t g_cap = 0;
  uint16_t *d_g2 = nullptr;          // their squares, rows of (width rounded up to 8) cells
  size_t g2_cap = 0;
  unsigned char *d_lut = nullptr;
  size_t lut_cap = 0;
  unsigned char *d_rows = nullptr;   // staging for host-resident grids
  size_t rows_cap = 0;
  float ms_kernels = 0;
};

static InflScratch &scratch()
{
  static InflScratch s;
  return s;
}

template <typename T>
static void grow(T *&p, size_t &cap, size_t need)
{
  if (need <= cap) return;
  if (p) cudaFree(p);
  cap = need + need / 4;
  AMCL_CUDA(cudaMalloc(&p, cap * sizeof(T)));
}

}  // namespace amclb200

using namespace amclb200;

extern "C" {

int amclb200_inflate_costs(unsigned char *costs, int on_device, int size_x, int size_y, int min_i,
                           int min_j, int max_i, int max_j, double resolution,
                           double inscribed_radius, double inflation_radius,
                           double cost_scaling_factor, int inflate_unknown)
{
  if (!costs || size_x <= 0 || size_y <= 0) AMCL_FATAL("amclb200_inflate_costs: empty grid");
  // Costmap2D::cellDistance (costmap_2d.cpp:181-185)
  const double cells_dist = fmax(0.0, ceil(inflation_radius / resolution));
  const unsigned int Ru = (unsigned int)cells_dist;
  if (Ru == 0) return 0;                                         // inflation_layer.cpp:179-180
  if (Ru > 254)
    AMCL_FATAL("inflation radius of %u cells exceeds the 254 the byte row distances are sized for", Ru);
  const int R = (int)Ru;
  ensure_device();
  InflBox b;
  if (!infl_box(size_x, size_y, min_i, min_j, max_i, max_j, R, &b)) return 0;
  const int aw = b.a_i1 - b.a_i0, ah = b.a_j1 - b.a_j0, rows = b.e_j1 - b.e_j0;

  // cost by squared cell distance: cached_costs_[dx][dy] = computeCost(hypot(dx, dy)) (:313, :336)
  const int R2 = R * R;
  std::vector<unsigned char> lut((size_t)R2 + 1, 0);
  std::vector<unsigned char> have((size_t)R2 + 1, 0);
  for (int dx = 0; dx <= R; dx++)
    for (int dy = dx; dy <= R && dx * dx + dy * dy <= R2; dy++)
    {
      const int d2 = dx * dx + dy * dy;
      const unsigned char cst = compute_cost(hypot((double)dx, (double)dy), resolution, inscribed_radius,
                                             cost_scaling_factor);
      if (have[d2] && lut[d2] != cst)
        AMCL_FATAL("hypot() gives cell offsets of equal length %d different costs (%d vs %d)", d2,
                   (int)lut[d2], (int)cst);
      lut[d2] = cst;
      have[d2] = 1;
    }

  InflScratch &s = scratch();
  cudaStream_t st = globals().stream;
  const int words = (aw + 63) / 64;
  const int pad = (R + 1 + 63) / 64 + 1;
  grow(s.d_bits, s.bits_cap, (size_t)rows * (words + 2 * pad));
  // AMCLB200_EDT=search keeps the per-cell search kernels for any radius (A/B timing)
  static const bool force_search = [] {
    const char *e = getenv("AMCLB200_EDT");
    return e && strcmp(e, "search") == 0;
  }();
  const bool fast = R <= kFastRadius && !force_search;
  const int g2_pitch = (aw + 7) & ~7;
  if (fast)
    grow(s.d_g2, s.g2_cap, (size_t)rows * g2_pitch);
  else
    grow(s.d_g, s.g_cap, (size_t)rows * aw);
  grow(s.d_lut, s.lut_cap, (size_t)R2 + 1);
  AMCL_CUDA(cudaMemcpyAsync(s.d_lut, lut.data(), (size_t)R2 + 1, cudaMemcpyHostToDevice, st));

  unsigned char *d_master = costs;
  int row0 = 0;                       // first grid row held at d_master
  int pitch = size_x;                 // bytes between rows at d_master
  if (!on_device)
  {
    // only the rectangle the update can touch crosses PCIe; the staging buffer
    // holds columns [a_i0, a_i1) with pitch aw, addressed through a base pointer
    // moved back by a_i0 so that the kernels keep using grid column indices
    grow(s.d_rows, s.rows_cap, (size_t)ah * aw);
    AMCL_CUDA(cudaMemcpy2DAsync(s.d_rows, (size_t)aw, costs + (size_t)b.a_j0 * size_x + b.a_i0, (size_t)size_x,
                                (size_t)aw, (size_t)ah, cudaMemcpyHostToDevice, st));
    d_master = s.d_rows - b.a_i0;
    row0 = b.a_j0;
    pitch = aw;
  }
  InflBox k = b;                      // the kernels index rows relative to d_master
  k.e_j0 -= row0; k.e_j1 -= row0; k.a_j0 -= row0; k.a_j1 -= row0;

  cudaEvent_t ev0, ev1;
  AMCL_CUDA(cudaEventCreate(&ev0));
  AMCL_CUDA(cudaEventCreate(&ev1));
  AMCL_CUDA(cudaEventRecord(ev0, st));
  {
    const long long warps = (long long)rows * (words + 2 * pad);
    infl_pack_rows<<<(unsigned)((warps + 7) / 8), 256, 0, st>>>(d_master, pitch, k, words, pad, s.d_bits);
    AMCL_LAUNCH_CHECK();
  }
  if (rows > 65535 || ah > 65535) AMCL_FATAL("update windows taller than 65535 rows are not supported");
  if (fast)
  {
    edt_row_squares<<<dim3((g2_pitch / 8 + 255) / 256, rows), 256, 0, st>>>(s.d_bits, g2_pitch, words, pad, R,
                                                                           s.d_g2);
    AMCL_LAUNCH_CHECK();
    const int chunks = (2 * R + 8 + 7) / 8;
    const size_t smem = (size_t)(kColY - kColV + chunks * 8) * kColX * sizeof(uint16_t);
    if (smem > 48 * 1024)
      AMCL_CUDA(cudaFuncSetAttribute(infl_col_squares, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    infl_col_squares<<<dim3((aw + kColX - 1) / kColX, (ah + kColY - 1) / kColY), 256, smem, st>>>(
        s.d_g2, g2_pitch, k, R, chunks, s.d_lut, inflate_unknown, pitch, d_master);
    AMCL_LAUNCH_CHECK();
  }
  else
  {
    infl_row_dist<<<dim3((aw + 255) / 256, rows), 256, 0, st>>>(s.d_bits, aw, words, pad, R, s.d_g);
    AMCL_LAUNCH_CHECK();
    infl_col_apply<<<dim3((aw + 255) / 256, ah), 256, 0, st>>>(s.d_g, aw, k, R, s.d_lut, inflate_unknown,
                                                               pitch, d_master);
    AMCL_LAUNCH_CHECK();
  }
  AMCL_CUDA(cudaEventRecord(ev1, st));
  if (!on_device)
    AMCL_CUDA(cudaMemcpy2DAsync(costs + (size_t)b.a_j0 * size_x + b.a_i0, (size_t)size_x, s.d_rows, (size_t)aw,
                                (size_t)aw, (size_t)ah, cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  AMCL_CUDA(cudaEventElapsedTime(&s.ms_kernels, ev0, ev1));
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  return aw * ah;
}

float amclb200_inflate_last_ms(void) { return scratch().ms_kernels; }

}  // extern "C"
