// edt_rows.cuh -- the row step shared by the distance transforms of map_dev.cu
// (likelihood-field map) and inflation_dev.cu (costmap inflation): obstacle
// cells of a row are a bit string, and the distance from a cell to the nearest
// set bit is read off with count-leading / count-trailing-zero instructions.
#pragma once

namespace amclb200
{

// distance from cell i down the row (towards lower x, cell i itself included) to
// the nearest set bit of A (bit strings with `pad` zero words on both sides);
// R + 1 or more when there is none within R
__device__ __forceinline__ int row_distance_left(const unsigned long long *__restrict__ A, int i, int R)
{
  const int k = i >> 6, sft = i & 63;
  unsigned long long w = A[k] << (63 - sft);          // bit i at the top
  int base = 0;
  int kk = k;
  int have = sft + 1;                                  // valid bits in w
  for (;;)
  {
    if (w) return base + __clzll((long long)w);
    base += have;
    if (base > R) return R + 1;
    kk--;
    w = A[kk];
    have = 64;
  }
}

// the same towards higher x
__device__ __forceinline__ int row_distance_right(const unsigned long long *__restrict__ A, int i, int R)
{
  const int k = i >> 6, sft = i & 63;
  unsigned long long w = A[k] >> sft;                  // bit i at the bottom
  int base = 0;
  int kk = k;
  int have = 64 - sft;
  for (;;)
  {
    if (w) return base + __ffsll((long long)w) - 1;
    base += have;
    if (base > R) return R + 1;
    kk++;
    w = A[kk];
    have = 64;
  }
}

// distance along the row from cell i to the nearest set bit of A, capped at R + 1
__device__ __forceinline__ int row_distance(const unsigned long long *__restrict__ A, int i, int R)
{
  const int left = row_distance_left(A, i, R), right = row_distance_right(A, i, R);
  const int best = left < right ? left : right;
  return best > R + 1 ? R + 1 : best;
}

}  // namespace amclb200
