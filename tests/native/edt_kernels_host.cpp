// edt_kernels_host.cpp -- the three kernels of navigation_b200/csrc/edt_squares.cuh
// run on host threads (cuda_thread_shim.hpp) with the launch geometry the library
// uses (edt_plan).  Built and called by tests/test_edt_kernels_host.py, which
// compares the tiled codes with the oracle's exact transform.
#include "cuda_thread_shim.hpp"

namespace amclb200
{
alignas(16) unsigned char s_raw[232448];   // the dynamic shared memory of the block that is running
}

#include "edt_squares.cuh"

using namespace amclb200;

extern "C" int edt_kernels_host_max_radius(void) { return kFastRadius; }

// occ: sx * sy states (-1, 0, +1), x fastest.  code: tiles_x * tiles_y 4x4 tiles
// with tiles_x = (sx + 1 + 3) / 4, tiles_y = (sy + 1 + 3) / 4 (map_dev.cu:160).
// use_pack16 = 0 packs the bit strings on the host (what edt_pack_rows does on
// the device for widths that are not a multiple of 16).
extern "C" int edt_kernels_host(const int8_t *occ, int sx, int sy, int R, int use_pack16, uint16_t *code)
{
  if (R > kFastRadius) return -1;
  const EdtPlan p = edt_plan(sx, sy, R);
  if (p.col_smem > sizeof(s_raw)) return -2;
  const int tiles_x = (sx + 1 + 3) / 4, tiles_y = (sy + 1 + 3) / 4;
  std::vector<unsigned long long> bits((size_t)sy * p.stride, 0xdeadbeefdeadbeefull);
  if (use_pack16)
  {
    if (sx & 15) return -3;
    launch(dim3((unsigned)((p.pack_threads + 255) / 256)), dim3(256),
           [&] { edt_pack_rows16(occ, sx, sy, p.words, p.pad, bits.data()); });
  }
  else
  {
    for (int j = 0; j < sy; j++)
      for (int w = -p.pad; w < p.words + p.pad; w++)
      {
        unsigned long long word = 0;
        for (int b = 0; b < 64; b++)
        {
          const int i = w * 64 + b;
          if (w >= 0 && w < p.words && i < sx && occ[(size_t)j * sx + i] == 1) word |= 1ull << b;
        }
        bits[(size_t)j * p.stride + (w + p.pad)] = word;
      }
  }
  std::vector<uint16_t> g2((size_t)p.pitch * sy, 0xabcd);
  launch(dim3(p.row_blocks_x, sy), dim3(256),
         [&] { edt_row_squares(bits.data(), p.pitch, p.words, p.pad, R, g2.data()); });
  launch(dim3(p.col_blocks_x, p.col_blocks_y), dim3(256), [&] {
    edt_col_squares(g2.data(), p.pitch, sx, sy, R, p.chunks, tiles_x, tiles_y, code);
  });
  return 0;
}

// ---- costmap inflation: edt_row_squares + infl_col_squares (inflation_kernels.cuh) ----
#include "inflation_kernels.cuh"

// costs: size_x * size_y cost values, updated in place as amclb200_inflate_costs
// does on a device-resident grid.  cost_of_d2: R * R + 1 entries.  The lethal cells
// of the seed window are packed on the host (infl_pack_rows does it on the device).
extern "C" int infl_kernels_host(unsigned char *costs, int size_x, int size_y, int min_i, int min_j, int max_i,
                                 int max_j, int R, const unsigned char *cost_of_d2, int inflate_unknown)
{
  if (R < 1 || R > kFastRadius) return -1;
  InflBox b;
  if (!infl_box(size_x, size_y, min_i, min_j, max_i, max_j, R, &b)) return 0;
  const int aw = b.a_i1 - b.a_i0, ah = b.a_j1 - b.a_j0, rows = b.e_j1 - b.e_j0;
  const int words = (aw + 63) / 64, pad = (R + 1 + 63) / 64 + 1, stride = words + 2 * pad;
  std::vector<unsigned long long> bits((size_t)rows * stride, 0ull);
  for (int r = 0; r < rows; r++)
    for (int c = 0; c < aw; c++)
    {
      const int i = b.a_i0 + c;
      if (i >= b.e_i0 && i < b.e_i1 && costs[(size_t)(b.e_j0 + r) * size_x + i] == COST_LETHAL)
        bits[(size_t)r * stride + pad + (c >> 6)] |= 1ull << (c & 63);
    }
  const int pitch = (aw + 7) & ~7;
  std::vector<uint16_t> g2((size_t)rows * pitch, 0xabcd);
  launch(dim3((pitch / 8 + 255) / 256, rows), dim3(256),
         [&] { edt_row_squares(bits.data(), pitch, words, pad, R, g2.data()); });
  const int chunks = (2 * R + 8 + 7) / 8;
  if ((size_t)(kColY - kColV + chunks * 8) * kColX * sizeof(uint16_t) > sizeof(s_raw)) return -2;
  launch(dim3((aw + kColX - 1) / kColX, (ah + kColY - 1) / kColY), dim3(256), [&] {
    infl_col_squares(g2.data(), pitch, b, R, chunks, cost_of_d2, inflate_unknown, size_x, costs);
  });
  return aw * ah;
}
