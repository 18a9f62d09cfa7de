// cuda_thread_shim.hpp -- just enough of the CUDA execution model to run a kernel
// header on host cores, one OS thread per CUDA thread of a block, blocks one
// after another.  Test infrastructure only (tests/test_edt_kernels_host.py): it
// lets the integer kernels of navigation_b200/csrc/edt_squares.cuh be checked
// against the oracle where there is no GPU.  Nothing in the library uses it.
#pragma once

#include <pthread.h>
#include <stdint.h>
#include <string.h>

#include <functional>
#include <thread>
#include <vector>

#define __global__
#define __device__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__
#define __align__(n)

struct dim3
{
  unsigned x, y, z;
  dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};
struct alignas(16) uint4 { unsigned x, y, z, w; };
struct alignas(8) uint2 { unsigned x, y; };
static inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }

static thread_local dim3 threadIdx, blockIdx;
static dim3 blockDim, gridDim;

namespace shim
{
static pthread_barrier_t block_barrier;
static pthread_barrier_t warp_barrier[32];
static unsigned warp_slot[1024];
}  // namespace shim

template <class T> static inline T __ldg(const T *p) { return *p; }
static inline int __clzll(long long v) { return v ? __builtin_clzll((unsigned long long)v) : 64; }
static inline int __ffsll(long long v) { return __builtin_ffsll(v); }

static inline unsigned __vcmpeq4(unsigned a, unsigned b)
{
  unsigned r = 0;
  for (int k = 0; k < 4; k++)
    if (((a >> (8 * k)) & 255u) == ((b >> (8 * k)) & 255u)) r |= 255u << (8 * k);
  return r;
}
static inline unsigned __vminu2(unsigned a, unsigned b)
{
  const unsigned lo = (a & 0xffffu) < (b & 0xffffu) ? (a & 0xffffu) : (b & 0xffffu);
  const unsigned hi = (a >> 16) < (b >> 16) ? (a >> 16) : (b >> 16);
  return lo | (hi << 16);
}
// per halfword min(a + b, c), the sum taken modulo 2^16 as the hardware does
static inline unsigned __viaddmin_u16x2(unsigned a, unsigned b, unsigned c)
{
  const unsigned lo = ((a & 0xffffu) + (b & 0xffffu)) & 0xffffu;
  const unsigned hi = ((a >> 16) + (b >> 16)) & 0xffffu;
  return __vminu2(lo | (hi << 16), c);
}

static inline void __syncthreads() { pthread_barrier_wait(&shim::block_barrier); }

static inline unsigned __shfl_down_sync(unsigned, unsigned v, int delta)
{
  const int t = (int)threadIdx.x, lane = t & 31;
  shim::warp_slot[t] = v;
  pthread_barrier_wait(&shim::warp_barrier[t >> 5]);
  const unsigned r = lane + delta < 32 ? shim::warp_slot[t + delta] : v;
  pthread_barrier_wait(&shim::warp_barrier[t >> 5]);
  return r;
}

// kernel<<<grid, block>>>(args...) as launch(grid, block, [&] { kernel(args...); })
static inline void launch(dim3 grid, dim3 block, const std::function<void()> &body)
{
  gridDim = grid;
  blockDim = block;
  const unsigned nt = block.x;
  pthread_barrier_init(&shim::block_barrier, nullptr, nt);
  for (unsigned w = 0; w < (nt + 31) / 32; w++)
    pthread_barrier_init(&shim::warp_barrier[w], nullptr, nt - w * 32 < 32 ? nt - w * 32 : 32);
  for (unsigned by = 0; by < grid.y; by++)
    for (unsigned bx = 0; bx < grid.x; bx++)
    {
      std::vector<std::thread> pool;
      pool.reserve(nt);
      for (unsigned t = 0; t < nt; t++)
        pool.emplace_back([&, t] {
          threadIdx = dim3(t, 0, 0);
          blockIdx = dim3(bx, by, 0);
          body();
        });
      for (auto &th : pool) th.join();
    }
  pthread_barrier_destroy(&shim::block_barrier);
  for (unsigned w = 0; w < (nt + 31) / 32; w++) pthread_barrier_destroy(&shim::warp_barrier[w]);
}
