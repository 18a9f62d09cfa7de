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
