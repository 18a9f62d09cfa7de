// inflation_kernels.cuh -- the column step of inflation_dev.cu with O(1) work per
// cell (radius <= kFastRadius): the register-block column minimum of
// edt_squares.cuh, finished with InflationLayer's update rule
// (/root/reference/costmap_2d/plugins/inflation_layer.cpp:245-256, 285-286).
// Kernel and window arithmetic only, so that tests/native/edt_kernels_host.cpp
// can run them on host threads against the oracle.
#pragma once

#include <stdint.h>

#include "edt_squares.cuh"

namespace amclb200
{

enum { COST_NO_INFORMATION = 255, COST_LETHAL = 254, COST_INSCRIBED = 253, COST_FREE = 0 };  // cost_values.h:42-45

struct InflBox
{
  int e_i0, e_j0, e_i1, e_j1;   // seed window: update window grown by R, clipped (inflation_layer.cpp:206-214)
  int a_i0, a_j0, a_i1, a_j1;   // cells the update can reach: seed window grown by R again, clipped
};

// false: the seed window is empty, nothing to do
inline bool infl_box(int size_x, int size_y, int min_i, int min_j, int max_i, int max_j, int R, InflBox *b)
{
  b->e_i0 = min_i - R > 0 ? min_i - R : 0;
  b->e_j0 = min_j - R > 0 ? min_j - R : 0;
  b->e_i1 = max_i + R < size_x ? max_i + R : size_x;
  b->e_j1 = max_j + R < size_y ? max_j + R : size_y;
  if (b->e_i0 >= b->e_i1 || b->e_j0 >= b->e_j1) return false;
  b->a_i0 = b->e_i0 - R > 0 ? b->e_i0 - R : 0;
  b->a_j0 = b->e_j0 - R > 0 ? b->e_j0 - R : 0;
  b->a_i1 = b->e_i1 + R < size_x ? b->e_i1 + R : size_x;
  b->a_j1 = b->e_j1 + R < size_y ? b->e_j1 + R : size_y;
  return true;
}

// g2: squares of the row distances of the seed rows e_j0 .. e_j1, columns counted
// from a_i0, `pitch` cells per row.  Block (bx, by) owns reachable cells
// [128 bx, 128 bx + 128) x [64 by, 64 by + 64) counted from (a_i0, a_j0); b's rows
// are relative to the first row held at `master`.
static __global__ void __launch_bounds__(256)
infl_col_squares(const uint16_t *__restrict__ g2, int pitch, InflBox b, int R, int chunks,
                 const unsigned char *__restrict__ cost_of_d2, int inflate_unknown, int master_pitch,
                 unsigned char *master)
{
  extern __shared__ __align__(16) unsigned char s_raw[];
  const int i0 = blockIdx.x * kColX, j0 = blockIdx.y * kColY;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned lo[kColV], hi[kColV];
  edt_col_block(g2, pitch, b.e_j1 - b.e_j0, b.a_j0 - b.e_j0, R, chunks, i0, j0, s_raw, lo, hi);
  const int aw = b.a_i1 - b.a_i0, ah = b.a_j1 - b.a_j0;
  const unsigned R2 = (unsigned)(R * R);
#pragma unroll
  for (int v = 0; v < kColV; v++)
  {
    const int jj = j0 + warp * kColV + v;
    if (jj >= ah) continue;
    const unsigned d2s[4] = {lo[v] & 0xffffu, lo[v] >> 16, hi[v] & 0xffffu, hi[v] >> 16};
#pragma unroll
    for (int q = 0; q < 4; q++)
    {
      const int c = i0 + lane * 4 + q;
      if (c >= aw || d2s[q] > R2) continue;        // enqueue: distance > cell_inflation_radius_ (:285-286)
      const unsigned char cost = cost_of_d2[d2s[q]];
      unsigned char *cell = master + (size_t)(b.a_j0 + jj) * master_pitch + (b.a_i0 + c);
      const unsigned char old_cost = *cell;
      // inflation_layer.cpp:253-256
      if (old_cost == COST_NO_INFORMATION &&
          (inflate_unknown ? (cost > COST_FREE) : (cost >= COST_INSCRIBED)))
        *cell = cost;
      else
        *cell = old_cost > cost ? old_cost : cost;
    }
  }
}

}  // namespace amclb200
