// common.cuh -- error handling and small device helpers shared by every .cu
// file of libamcl_b200.  sm_100a only; there is no host fallback anywhere: a
// CUDA failure prints the failing call and aborts, because the reference API
// (void functions, no status codes -- SURVEY.md 8b "Error conventions") has no
// channel to report it through.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define AMCL_CUDA(call)                                                        \
  do                                                                           \
  {                                                                            \
    cudaError_t err__ = (call);                                                \
    if (err__ != cudaSuccess)                                                  \
    {                                                                          \
      fprintf(stderr, "[amcl_b200] CUDA failure %s (%d) at %s:%d in `%s`\n",   \
              cudaGetErrorString(err__), (int)err__, __FILE__, __LINE__, #call); \
      abort();                                                                 \
    }                                                                          \
  } while (0)

#define AMCL_FATAL(...)                                                        \
  do                                                                           \
  {                                                                            \
    fprintf(stderr, "[amcl_b200] fatal: ");                                    \
    fprintf(stderr, __VA_ARGS__);                                              \
    fprintf(stderr, " (%s:%d)\n", __FILE__, __LINE__);                         \
    abort();                                                                   \
  } while (0)

// After every kernel launch.  With AMCLB200_DEBUG_SYNC=1 in the environment the
// launch is also waited for, so a device fault is reported at the kernel that
// raised it rather than at the next blocking call.
inline unsigned long long &amcl_launch_counter()
{
  static unsigned long long n = 0;
  return n;
}
inline bool amcl_debug_sync()
{
  static int flag = -1;
  if (flag < 0)
  {
    const char *e = getenv("AMCLB200_DEBUG_SYNC");
    flag = (e && e[0] == '1') ? 1 : 0;
  }
  return flag == 1;
}
#define AMCL_LAUNCH_CHECK()                                \
  do                                                       \
  {                                                        \
    AMCL_CUDA(cudaGetLastError());                         \
    amcl_launch_counter()++;                               \
    if (amcl_debug_sync()) AMCL_CUDA(cudaDeviceSynchronize()); \
  } while (0)

namespace amclb200
{

constexpr int kWarp = 32;
constexpr unsigned kFullMask = 0xffffffffu;

// One particle exactly as pf_sample_t lays it out (pf.h:59-67): 32 bytes,
// 32-byte aligned, so a particle is one L2 sector and two 16-byte loads.
struct __align__(32) Particle
{
  double x, y, th, w;
};
static_assert(sizeof(Particle) == 32, "Particle must match pf_sample_t");

#ifdef __CUDACC__

__device__ __forceinline__ Particle load_particle(const Particle *p)
{
  // Two LDG.128: both halves of the same sector.
  const double2 *q = reinterpret_cast<const double2 *>(p);
  double2 a = q[0], b = q[1];
  Particle r;
  r.x = a.x; r.y = a.y; r.th = b.x; r.w = b.y;
  return r;
}

__device__ __forceinline__ void store_particle(Particle *p, const Particle &v)
{
  double2 *q = reinterpret_cast<double2 *>(p);
  q[0] = make_double2(v.x, v.y);
  q[1] = make_double2(v.th, v.w);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
  return v;
}

__device__ __forceinline__ int warp_sum_int(int v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
  return v;
}

// Block-wide sum in a fixed order (lane tree, then warp 0 over warp totals):
// the result depends only on blockDim, never on scheduling.  `scratch` needs
// one double per warp.  Valid in thread 0.
__device__ __forceinline__ double block_sum(double v, double *scratch)
{
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  double t = 0.0;
  if (wid == 0)
  {
    t = lane < nw ? scratch[lane] : 0.0;
    t = warp_sum(t);
  }
  return t;
}

// floor(x) for |x| < 2^31 without the slow F2I.F64 path: adding 1.5*2^52 with
// round-toward-minus-infinity leaves floor(x) in the low mantissa word.
__device__ __forceinline__ int floor_to_int(double x)
{
  return __double2loint(__dadd_rd(x, 6755399441055744.0));
}

#endif  // __CUDACC__

}  // namespace amclb200
