// edt_rows.cuh -- the row step shared by the distance transforms of map_dev.cu
// (likelihood-field map) and inflation_dev.cu (costmap inflation): obstacle
// cells of a row are a bit string, and the distance from a cell to the nearest
// set bit is read off with count-leading / count-trailing-zero instructions.
#pragma once

namespace amclb200
{

// distance along the row from cell i to the nearest set bit of A (bit strings
// with `pad` zero words on both sides), capped at R + 1
__device__ __forceinline__ int row_distance(const unsigned long long *__restrict__ A, int i, int R)
{
  const int k = i >> 6, sft = i & 63;
  // towards lower x: bits i, i-1, ... ; towards higher x: bits i, i+1, ...
  int left = R + 1, right = R + 1;
  {
    unsigned long long w = A[k] << (63 - sft);          // bit i at the top
    int base = 0;
    int kk = k;
    int have = sft + 1;                                  // valid bits in w
    for (;;)
    {
      if (w) { left = base + __clzll((long long)w); break; }
      base += have;
      if (base > R) break;
      kk--;
      w = A[kk];
      have = 64;
    }
  }
  {
    unsigned long long w = A[k] >> sft;                  // bit i at the bottom
    int base = 0;
    int kk = k;
    int have = 64 - sft;
    for (;;)
    {
      if (w) { right = base + __ffsll((long long)w) - 1; break; }
      base += have;
      if (base > R) break;
      kk++;
      w = A[kk];
      have = 64;
    }
  }
  int best = left < right ? left : right;
  return best > R + 1 ? R + 1 : best;
}

}  // namespace amclb200
