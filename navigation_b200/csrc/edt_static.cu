// edt_static.cu -- amclb200_edt_codes_static(): the bounded exact distance
// transform for callers that are not allowed to allocate device memory.
//
// Who needs this.  Some users of this library run under a "no allocation after
// start-up" rule: a hard real-time localisation loop where a cudaMalloc (which
// may take a driver lock and stall for milliseconds) is forbidden once the
// robot is moving, a safety-certified runtime that bans dynamic allocation
// outright, or a process whose device memory has been handed in full to
// another allocator so that a stray cudaMalloc would simply fail.  The regular
// entry point, map_update_cspace(), allocates its device buffers (occupancy
// bytes, codes, row-distance scratch) when it first sees a map.  This variant
// allocates nothing at all: the caller passes the input and the output as
// device pointers it already owns, and the one piece of scratch the transform
// needs -- the per-cell row distances handed from the row pass to the column
// pass -- lives in a fixed array inside the library image.
//
// The price.
//   * A hard size limit: at most AMCLB200_EDT_STATIC_MAX_CELLS cells (4 Mi,
//     e.g. 2048 x 2048) and a radius of at most 254 cells; larger requests are
//     refused with an error, never truncated.
//   * The array (4 MiB) is part of the loaded module, so every GPU that loads
//     libamcl_b200.so carries it until the process exits, used or not.
//   * There is ONE array per GPU, shared by every call and every stream.  It
//     keeps whatever the previous call left in it.  That is harmless for a
//     single call -- the row pass writes every byte the column pass will read
//     before the column pass starts, so nothing has to begin at zero and
//     nothing is cleared -- but two calls whose kernels overlap on the same GPU
//     (different streams, or different host threads) race on it and both
//     results are garbage.  Serialise calls per GPU.
//   * The first call on each GPU looks up the array's device address
//     (cudaGetSymbolAddress, with that GPU current) and remembers it.  Make
//     that first call during start-up, outside any CUDA stream capture.
//
// The arithmetic is the same two-pass transform as map_dev.cu (row pass:
// distance to the nearest occupied cell of the row, capped at R+1; column
// pass: min over |dj| <= R of dj^2 + g^2, "far" above R^2).  This file has no
// counterpart in the reference and touches none of the existing code paths.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#define AMCLB200_EDT_STATIC_MAX_CELLS (4u * 1024u * 1024u)
#define AMCLB200_EDT_STATIC_MAX_RADIUS 254
#define AMCLB200_EDT_STATIC_MAX_DEVICES 64

// Row distances are at most R+1 <= 255, so one byte per cell is enough.
// In HOST code this name is only a handle for cudaGetSymbolAddress; it is not
// a device address and must never be passed to a kernel as one.
static __device__ __align__(256) unsigned char g_edt_row_scratch[AMCLB200_EDT_STATIC_MAX_CELLS];

namespace
{

constexpr int kSeg = 512;    // cells of one row handled by a block in the row pass
constexpr int kPatch = 32;   // the column pass works on 32 x 32 cell patches

// Row pass.  scratch[j*sx + i] = min(distance along row j from cell i to the
// nearest cell with occ == +1, R+1).
__global__ void static_row_pass(const signed char *__restrict__ occ, int sx, int sy, int R,
                                unsigned char *__restrict__ scratch)
{
  extern __shared__ unsigned char s_hit[];   // kSeg + 2R flags
  const int j = blockIdx.y;
  const int i0 = blockIdx.x * kSeg;
  const int span = kSeg + 2 * R;
  const signed char *row = occ + (size_t)j * sx;
  for (int k = threadIdx.x; k < span; k += blockDim.x)
  {
    const int i = i0 - R + k;
    s_hit[k] = (i >= 0 && i < sx && row[i] == 1) ? 1 : 0;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < kSeg; k += blockDim.x)
  {
    const int i = i0 + k;
    if (i >= sx) break;
    const int c = k + R;
    int best = R + 1;
    for (int d = 0; d <= R; d++)
      if (s_hit[c - d] | s_hit[c + d])
      {
        best = d;
        break;
      }
    scratch[(size_t)j * sx + i] = (unsigned char)best;
  }
}

// Column pass.  codes[j*sx + i] = squared distance in cells to the nearest
// occupied cell if that is <= R^2, otherwise R^2 + 1.  Row-major output.
__global__ void static_col_pass(const unsigned char *__restrict__ scratch, int sx, int sy, int R,
                                unsigned short *__restrict__ codes)
{
  extern __shared__ unsigned char s_g[];     // (kPatch + 2R) x kPatch row distances
  const int i0 = blockIdx.x * kPatch, j0 = blockIdx.y * kPatch;
  const int rows = kPatch + 2 * R;
  for (int k = threadIdx.x; k < rows * kPatch; k += blockDim.x)
  {
    const int lj = k / kPatch, li = k % kPatch;
    const int j = j0 - R + lj, i = i0 + li;
    int g = R + 1;                            // rows above / below the map: nothing there
    if (j >= 0 && j < sy && i < sx) g = scratch[(size_t)j * sx + i];
    s_g[k] = (unsigned char)g;
  }
  __syncthreads();
  const int R2 = R * R;
  for (int k = threadIdx.x; k < kPatch * kPatch; k += blockDim.x)
  {
    const int lj = k / kPatch, li = k % kPatch;
    const int i = i0 + li, j = j0 + lj;
    if (i >= sx || j >= sy) continue;
    int best = R2 + 1;
    for (int d = 0; d <= R; d++)
    {
      const int dd = d * d;
      if (dd >= best) break;                  // farther rows cannot improve any more
      const int ga = s_g[(lj + R - d) * kPatch + li];
      const int gb = s_g[(lj + R + d) * kPatch + li];
      const int g = ga < gb ? ga : gb;
      const int v = dd + g * g;
      if (v < best) best = v;
    }
    codes[(size_t)j * sx + i] = (unsigned short)(best > R2 ? R2 + 1 : best);
  }
}

// Device address of the scratch array, looked up once per GPU.  Atomic: two host
// threads may make the first call on a GPU at the same time; both then resolve
// the same address and store it, which is harmless.
std::atomic<unsigned char *> g_scratch_addr[AMCLB200_EDT_STATIC_MAX_DEVICES];

}  // namespace

extern "C" {

unsigned long long amclb200_edt_static_max_cells(void) { return AMCLB200_EDT_STATIC_MAX_CELLS; }

// See the file header and include/amcl_b200.h.  Returns 0 on success; on any
// failure prints one line to stderr, writes nothing to d_codes and returns
//   -1 bad argument, -2 map or radius beyond the fixed limits, -3 CUDA error.
int amclb200_edt_codes_static(const signed char *d_occ, int size_x, int size_y, int cell_radius,
                              unsigned short *d_codes, void *cuda_stream)
{
  if (d_occ == NULL || d_codes == NULL || size_x <= 0 || size_y <= 0 || cell_radius < 0)
  {
    fprintf(stderr, "[amcl_b200] amclb200_edt_codes_static: bad argument (size %d x %d, radius %d)\n",
            size_x, size_y, cell_radius);
    return -1;
  }
  const unsigned long long cells = (unsigned long long)size_x * (unsigned long long)size_y;
  if (cells > AMCLB200_EDT_STATIC_MAX_CELLS || cell_radius > AMCLB200_EDT_STATIC_MAX_RADIUS ||
      size_y > 65535)
  {
    fprintf(stderr,
            "[amcl_b200] amclb200_edt_codes_static: %d x %d cells with radius %d exceeds the fixed "
            "scratch (at most %u cells, radius %d, 65535 rows); use map_update_cspace() instead\n",
            size_x, size_y, cell_radius, AMCLB200_EDT_STATIC_MAX_CELLS,
            AMCLB200_EDT_STATIC_MAX_RADIUS);
    return -2;
  }

  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess || dev < 0 || dev >= AMCLB200_EDT_STATIC_MAX_DEVICES)
  {
    fprintf(stderr, "[amcl_b200] amclb200_edt_codes_static: no usable current device (%s)\n",
            err == cudaSuccess ? "device index out of range" : cudaGetErrorString(err));
    return -3;
  }
  unsigned char *scratch = g_scratch_addr[dev].load(std::memory_order_acquire);
  if (scratch == NULL)
  {
    // One-time set-up for this GPU (it is the current device): resolve the
    // array's address in THIS device's copy of the module.
    unsigned char *addr = NULL;
    err = cudaGetSymbolAddress((void **)&addr, g_edt_row_scratch);
    if (err != cudaSuccess || addr == NULL)
    {
      fprintf(stderr, "[amcl_b200] amclb200_edt_codes_static: cudaGetSymbolAddress failed on "
                      "device %d (%s)\n", dev, cudaGetErrorString(err));
      return -3;
    }
    g_scratch_addr[dev].store(addr, std::memory_order_release);
    scratch = addr;
  }
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const int R = cell_radius;

  // An error left pending by earlier, unrelated work would otherwise be taken
  // for a failure of these launches: report it as such and launch nothing.
  err = cudaPeekAtLastError();
  if (err != cudaSuccess)
  {
    fprintf(stderr, "[amcl_b200] amclb200_edt_codes_static: an earlier CUDA error is pending (%s); "
                    "nothing launched\n", cudaGetErrorString(err));
    return -3;
  }
  dim3 g1((size_x + kSeg - 1) / kSeg, size_y);
  static_row_pass<<<g1, 256, kSeg + 2 * R, st>>>(d_occ, size_x, size_y, R, scratch);
  err = cudaGetLastError();
  if (err != cudaSuccess)
  {
    fprintf(stderr, "[amcl_b200] amclb200_edt_codes_static: row pass not launched (%s)\n",
            cudaGetErrorString(err));
    return -3;     // d_codes untouched: only the second kernel writes it
  }
  dim3 g2((size_x + kPatch - 1) / kPatch, (size_y + kPatch - 1) / kPatch);
  const size_t smem = (size_t)(kPatch + 2 * R) * kPatch;   // <= 17.3 KB at R = 254
  static_col_pass<<<g2, 256, smem, st>>>(scratch, size_x, size_y, R, d_codes);
  err = cudaGetLastError();
  if (err != cudaSuccess)
  {
    fprintf(stderr, "[amcl_b200] amclb200_edt_codes_static: column pass not launched (%s)\n",
            cudaGetErrorString(err));
    return -3;
  }
  return 0;
}

}  // extern "C"
