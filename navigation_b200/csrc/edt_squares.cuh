// edt_squares.cuh -- the exact bounded distance transform of map_dev.cu with O(1)
// work per cell (map_update_cspace, /root/reference/amcl/src/amcl/map/map_cspace.cpp:120-176,
// restated as an exact separable transform).  Kernels only, no host code and no
// dependence on the library's context: tests/native/edt_kernels_host.cpp compiles
// this file for the host with a thread-per-CUDA-thread shim and checks the three
// kernels against the oracle without a GPU.
#pragma once

#include <stdint.h>

#include "edt_rows.cuh"

namespace amclb200
{

// ---- the same transform with O(1) work per cell (radius <= kFastRadius) ----------
//
// (1') edt_pack_rows16: 16 cells per thread (one 16-byte load), four threads make
//      a word.  Needs 16-byte aligned rows (sx % 16 == 0); other widths keep (1).
// (2') edt_row_squares: a thread resolves the 8 cells of one byte of the bit
//      string with ONE search to the left of the byte and ONE to the right, then
//      a scan over its 8 bits; it stores g^2 (uint16, (R+1)^2 = no obstacle
//      within R) with one 16-byte store into rows of their own pitch.
// (3') edt_col_squares: d2(i, j) = min_dj dj^2 + g^2(i, j + dj).  A thread owns 4
//      columns x 8 rows of outputs as 16 registers of two uint16 each; every row
//      of g^2 it reads from shared memory (8 bytes) feeds all 8 output rows with
//      one add-min instruction per register (VIADDMNMX.U16x2), so a cell costs
//      about (8 + 2R)/8 * 2.8 instructions whatever its distance.  The sums stay
//      below 2^16 for R <= kFastRadius: (R+14)^2 + (R+1)^2 <= 65535.  A thread's
//      outputs are two whole 4x4 tiles: two 32-byte stores, no staging.
constexpr int kFastRadius = 160;
constexpr int kColX = 128, kColY = 64, kColV = 8;

// Launch geometry of the three kernels for a map of sx x sy cells and radius R
// (shared by map_dev.cu and the host-side check of the kernels).
struct EdtPlan
{
  int words, pad, stride;      // bit strings: words per row, zero words either side
  int pitch;                   // cells per row of g^2
  int chunks;                  // groups of 8 source rows a warp of the column kernel walks
  long long pack_threads;      // edt_pack_rows16, 256 per block
  int row_blocks_x;            // edt_row_squares: grid (row_blocks_x, sy), 256 per block
  int col_blocks_x, col_blocks_y;
  size_t col_smem;
};

inline EdtPlan edt_plan(int sx, int sy, int R)
{
  EdtPlan p;
  p.words = (sx + 63) / 64;
  p.pad = (R + 1 + 63) / 64 + 1;
  p.stride = p.words + 2 * p.pad;
  p.pitch = (sx + 7) & ~7;
  p.chunks = (2 * R + 8 + 7) / 8;
  p.pack_threads = (long long)sy * p.stride * 4;
  p.row_blocks_x = (p.pitch / 8 + 255) / 256;
  p.col_blocks_x = (sx + 1 + kColX - 1) / kColX;     // + the padding column / row of tiles
  p.col_blocks_y = (sy + 1 + kColY - 1) / kColY;
  p.col_smem = (size_t)(kColY - kColV + p.chunks * 8) * kColX * sizeof(uint16_t);
  return p;
}

static __global__ void edt_pack_rows16(const int8_t *__restrict__ occ, int sx, int sy, int words, int pad,
                                unsigned long long *__restrict__ bits)
{
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int stride = words + 2 * pad;
  const long long q = t >> 2;                       // word
  const int part = (int)(t & 3);                    // which 16 cells of it
  const int j = (int)(q / stride), w = (int)(q % stride) - pad;
  unsigned mask = 0;
  if (j < sy && w >= 0 && w < words)
  {
    const int i = w * 64 + part * 16;
    if (i < sx)                                     // sx % 16 == 0: all 16 cells are on the map
    {
      const uint4 v = __ldg(reinterpret_cast<const uint4 *>(occ + (size_t)j * sx + i));
      const unsigned c[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; k++)
      {
        const unsigned eq = __vcmpeq4(c[k], 0x01010101u);                   // 0xff where occ == +1
        mask |= (((eq & 0x08040201u) * 0x01010101u) >> 24) << (4 * k);      // one bit per byte
      }
    }
  }
  // the four parts of a word sit in four neighbouring lanes
  const unsigned m1 = __shfl_down_sync(0xffffffffu, mask, 1);
  const unsigned m2 = __shfl_down_sync(0xffffffffu, mask, 2);
  const unsigned m3 = __shfl_down_sync(0xffffffffu, mask, 3);
  if (part == 0 && j < sy)
    bits[(size_t)j * stride + (w + pad)] = (unsigned long long)(mask | (m1 << 16)) |
                                           ((unsigned long long)(m2 | (m3 << 16)) << 32);
}

static __global__ void __launch_bounds__(256)
edt_row_squares(const unsigned long long *__restrict__ bits, int pitch, int words, int pad, int R,
                uint16_t *__restrict__ g2)
{
  const int j = blockIdx.y;
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i >= pitch) return;
  const int stride = words + 2 * pad;
  const unsigned long long *A = bits + (size_t)j * stride + pad;   // A[-pad .. words+pad)
  const unsigned own = (unsigned)(A[i >> 6] >> (i & 63)) & 255u;
  int run = row_distance_left(A, i - 1, R);          // from the cell left of the byte
  run = run > R + 1 ? R + 1 : run;
  int left[8];
#pragma unroll
  for (int b = 0; b < 8; b++)
  {
    run = (own >> b) & 1u ? 0 : run + 1;
    left[b] = run;
  }
  run = row_distance_right(A, i + 8, R);             // from the cell right of the byte
  run = run > R + 1 ? R + 1 : run;
  unsigned sq[8];
#pragma unroll
  for (int b = 7; b >= 0; b--)
  {
    run = (own >> b) & 1u ? 0 : run + 1;
    int d = left[b] < run ? left[b] : run;
    d = d > R + 1 ? R + 1 : d;
    sq[b] = (unsigned)(d * d);
  }
  uint4 out;
  out.x = sq[0] | (sq[1] << 16);
  out.y = sq[2] | (sq[3] << 16);
  out.z = sq[4] | (sq[5] << 16);
  out.w = sq[6] | (sq[7] << 16);
  *reinterpret_cast<uint4 *>(g2 + (size_t)j * pitch + i) = out;
}

// The column minimum for the block's 128 x 64 outputs.  Output row j (block row
// j0 + warp * 8 + v) reads source rows j + src_shift - R .. j + src_shift + R of g2,
// which holds src_rows rows; rows outside it count as "no obstacle".  Leaves
// min_dj dj^2 + g^2 of columns i0 + 4 * lane .. + 3 in lo[v] (two columns) and hi[v].
// Every thread of the block must call it (it holds the block's barrier).
__device__ __forceinline__ void edt_col_block(const uint16_t *__restrict__ g2, int pitch, int src_rows,
                                              int src_shift, int R, int chunks, int i0, int j0,
                                              unsigned char *s_raw, unsigned (&lo)[kColV],
                                              unsigned (&hi)[kColV])
{
  // rows j0 - R .. of the patch, 128 cells (256 bytes) each; the rows past the
  // halo that the last chunk of 8 touches hold "no obstacle"
  const int rows = kColY - kColV + chunks * 8;
  const unsigned far2 = (unsigned)((R + 1) * (R + 1)) * 0x10001u;
  {
    uint4 *s4 = reinterpret_cast<uint4 *>(s_raw);
    for (int k = threadIdx.x; k < rows * (kColX / 8); k += blockDim.x)
    {
      const int lj = k / (kColX / 8), c = (k % (kColX / 8)) * 8;
      const int r = j0 - R + lj + src_shift, i = i0 + c;
      uint4 v = make_uint4(far2, far2, far2, far2);
      if (lj < kColY + 2 * R && r >= 0 && r < src_rows && i < pitch)
        v = __ldg(reinterpret_cast<const uint4 *>(g2 + (size_t)r * pitch + i));
      s4[k] = v;
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint2 *col = reinterpret_cast<const uint2 *>(s_raw) + (size_t)warp * kColV * (kColX / 4) + lane;
#pragma unroll
  for (int v = 0; v < kColV; v++) lo[v] = hi[v] = 0xffffffffu;
  // D[k] = dj^2 (in both halves) for dj = t0 + k - 7 - R: source row t0 + u of
  // the warp's window meets output row v at dj = t0 + u - v - R = D[7 + u - v]
  unsigned D[15];
#pragma unroll
  for (int k = 0; k < 7; k++) D[k + 8] = (unsigned)((k - 7 - R) * (k - 7 - R)) * 0x10001u;
  for (int t0 = 0; t0 < chunks * 8; t0 += 8)
  {
#pragma unroll
    for (int k = 0; k < 7; k++) D[k] = D[k + 8];
#pragma unroll
    for (int k = 7; k < 15; k++)
    {
      const int dj = t0 + k - 7 - R;
      D[k] = (unsigned)(dj * dj) * 0x10001u;
    }
#pragma unroll
    for (int u = 0; u < 8; u++)
    {
      const uint2 row = col[(size_t)(t0 + u) * (kColX / 4)];
      if (row.x == far2 && row.y == far2) continue;      // nothing within R along this row
#pragma unroll
      for (int v = 0; v < kColV; v++)
      {
        lo[v] = __viaddmin_u16x2(row.x, D[7 + u - v], lo[v]);
        hi[v] = __viaddmin_u16x2(row.y, D[7 + u - v], hi[v]);
      }
    }
  }
}

static __global__ void __launch_bounds__(256)
edt_col_squares(const uint16_t *__restrict__ g2, int pitch, int sx, int sy, int R, int chunks,
                int tiles_x, int tiles_y, uint16_t *__restrict__ code)
{
  extern __shared__ __align__(16) unsigned char s_raw[];
  const int i0 = blockIdx.x * kColX, j0 = blockIdx.y * kColY;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned lo[kColV], hi[kColV];
  edt_col_block(g2, pitch, sy, 0, R, chunks, i0, j0, s_raw, lo, hi);
  // codes: d2 <= R^2 as they are, R^2 + 1 beyond, R^2 + 2 off the map
  const unsigned R2 = (unsigned)(R * R);
  const unsigned beyond2 = (R2 + 1) * 0x10001u, off2 = (R2 + 2) * 0x10001u;
  const int ci = i0 + lane * 4;
  const unsigned keep_lo = (ci < sx ? 0x0000ffffu : 0u) | (ci + 1 < sx ? 0xffff0000u : 0u);
  const unsigned keep_hi = (ci + 2 < sx ? 0x0000ffffu : 0u) | (ci + 3 < sx ? 0xffff0000u : 0u);
#pragma unroll
  for (int v = 0; v < kColV; v++)
  {
    const bool row_in = j0 + warp * kColV + v < sy;
    const unsigned a = __vminu2(lo[v], beyond2), b = __vminu2(hi[v], beyond2);
    lo[v] = row_in ? (a & keep_lo) | (off2 & ~keep_lo) : off2;
    hi[v] = row_in ? (b & keep_hi) | (off2 & ~keep_hi) : off2;
  }
  const int gi = (i0 >> 2) + lane;
  if (gi >= tiles_x) return;
#pragma unroll
  for (int half = 0; half < 2; half++)
  {
    const int gj = (j0 >> 2) + warp * 2 + half;
    if (gj >= tiles_y) continue;
    uint4 *dst = reinterpret_cast<uint4 *>(code + ((size_t)gj * tiles_x + gi) * 16);
    dst[0] = make_uint4(lo[half * 4 + 0], hi[half * 4 + 0], lo[half * 4 + 1], hi[half * 4 + 1]);
    dst[1] = make_uint4(lo[half * 4 + 2], hi[half * 4 + 2], lo[half * 4 + 3], hi[half * 4 + 3]);
  }
}

}  // namespace amclb200
