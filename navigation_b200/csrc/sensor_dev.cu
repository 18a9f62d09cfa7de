// sensor_dev.cu -- the three laser measurement models of
// /root/reference/amcl/src/amcl/sensors/amcl_laser.cpp on the device, and the
// weight normalisation of pf_update_sensor (/root/reference/amcl/src/amcl/pf/pf.c:268-318).
//
// Work decomposition (all three models): ONE WARP PER PARTICLE, LANES OVER
// BEAMS.  Neighbouring beams of one scan end in neighbouring cells, so the 32
// gathers of one warp instruction fall into a handful of 32-byte sectors of the
// 4x4-tiled distance-code map (context.hpp) instead of 32 unrelated ones; the
// per-particle sum is a warp-shuffle reduction; the scan table and the
// code -> likelihood table are staged once per CTA in shared memory by a bulk
// (TMA) copy.
//
// Arithmetic: fp64 throughout (H6 in SURVEY.md: fp32 end points flip cells).
// cos/sin(theta + bearing) come from one sincos per particle and the angle
// addition formulas with host-computed cos/sin(bearing); cell coordinates are
// formed directly in cell units.  Both change end points by a few ulp only
// (cell flips ~1e-12 per beam), inside the 1e-5 weight tolerance.
#include <math.h>
#include <float.h>
#include <string.h>

#include <vector>

#include "amcl/sensors/amcl_laser.h"
#include "context.hpp"

namespace amclb200
{

// ------------------------------------------------------------ bulk staging --

__device__ __forceinline__ uint32_t smem_addr(const void *p)
{
  return (uint32_t)__cvta_generic_to_shared(p);
}

// Thread 0 arms an mbarrier with the byte count and issues 1-D bulk copies
// (cp.async.bulk, the TMA engine); every thread then waits on the barrier.
__device__ __forceinline__ void bulk_stage_begin(uint64_t *bar, uint32_t total_bytes)
{
  if (threadIdx.x == 0)
  {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)),
                 "r"(total_bytes)
                 : "memory");
  }
}

__device__ __forceinline__ void bulk_stage_copy(void *dst, const void *src, uint32_t bytes,
                                                uint64_t *bar)
{
  // bytes: multiple of 16; dst/src 16-byte aligned.  Called by thread 0 only.
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_addr(dst)),
      "l"(src), "r"(bytes), "r"(smem_addr(bar))
      : "memory");
}

__device__ __forceinline__ void bulk_stage_wait(uint64_t *bar)
{
  uint32_t done = 0;
  const uint32_t a = smem_addr(bar);
  while (!done)
  {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(a)
        : "memory");
  }
}

// --------------------------------------------------- likelihood-field kernels --

struct LfArgs
{
  Particle *set;
  int n;
  const double2 *beams;     // (r cos b / scale, r sin b / scale) per used beam
  int nb;                   // number of used beams
  int nb_pad;               // rounded up to an even count (16-byte granules)
  const double *lut;        // rank -> pz^3 (LF) | log pz (LFP), each entry repeated 2^LOGC times
  const unsigned char *agree;  // rank -> z < beam_skip_distance (COUNT mode)
  int nranks_pad;           // even
  const uint16_t *lf;       // laser view of the distance map (MapCtx::d_lf)
  cudaTextureObject_t lf_tex;  // the same array as a linear texture (TEXQ > 0 only)
  cudaTextureObject_t lf_tex2d;  // the laser view as a 2-D array texture, clamp addressing (TEXQ == kTex2D)
  unsigned max_i, max_j;    // lf_w - 1, lf_h - 1: both lie in the apron
  unsigned k_row;           // lf_w - 1 (see lf_index)
  double lx, ly, la;        // laser pose in the robot frame
  double ox, oy, scale;     // map origin (centre) and resolution
  double half_x, half_y;    // 0.5 + size/2 + apron, folded into the cell coordinate
  double in_lo, in_hi_x, in_hi_y;  // a laser at in_lo <= X0 <= in_hi_x (Y alike) keeps every beam inside the view
  double *block_totals;     // one partial per CTA
  int *beam_counts;         // COUNT mode: agreements per used beam
  double *logsum;           // COUNT mode: out, sum of log pz over ALL used beams per particle;
                            // LFP mode, if not null: in, weight *= exp(logsum - sum over the beams given)
};

enum { MODE_LF = 0, MODE_LFP = 1, MODE_COUNT = 2 };

constexpr int kSensorThreads = 256;     // beam model
#ifndef AMCL_LF_THREADS
#define AMCL_LF_THREADS 512
#endif
#ifndef AMCL_LF_MINBLOCKS
#define AMCL_LF_MINBLOCKS 1
#endif
constexpr int kLfThreads = AMCL_LF_THREADS;   // likelihood-field kernels: one CTA of 16 warps per SM (~126 registers)
constexpr double kCellGuard = 536870912.0;  // 2^29: beyond this nothing is on a map

// Particles handled together by one warp: every beam-table entry fetched from
// shared memory is applied to kBlockP particles, which divides the shared-
// memory traffic of the table by kBlockP and keeps kBlockP independent map
// gathers in flight per lane.  Measured at 16 Mi x 1059 (profiles/r2_lf_kernel.md):
// 8 particles x 16 warps per SM 15.5 ms, 4 x 24 warps 16.1 ms, 16 x 8 warps 17.9 ms.
#ifndef AMCL_LF_BLOCKP
#define AMCL_LF_BLOCKP 8
#endif
constexpr int kBlockP = AMCL_LF_BLOCKP;
#ifndef AMCL_LF_UNROLL
#define AMCL_LF_UNROLL 3
#endif
constexpr int kLfUnroll = AMCL_LF_UNROLL;   // beams per lane in flight, linear texture
#ifndef AMCL_LF_UNROLL_2D
#define AMCL_LF_UNROLL_2D 2
#endif
constexpr int kLfUnroll2D = AMCL_LF_UNROLL_2D;   // ... 2-D array texture
constexpr int kTex2D = 100;                 // TEXQ value: every gather is a 2-D texel fetch

// Texel (x, y) of a 2-D uint16 texture, integer coordinates, which must lie inside
// the array.
__device__ __forceinline__ unsigned texel_2d(cudaTextureObject_t tex, int x, int y)
{
  unsigned r0, r1, r2, r3;
  asm("tex.2d.v4.u32.s32 {%0, %1, %2, %3}, [%4, {%5, %6}];"
      : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
      : "l"(tex), "r"(x), "r"(y));
  return r0;
}

// Offset of cell (mi, mj) in the laser view: 4x4 tiles, tiles row-major, cells
// of a tile column-first,
//   ((mj >> 2) * (w / 4) + (mi >> 2)) * 16 + (mi & 3) * 4 + (mj & 3)
//     = (mj & ~3) * (w - 1) + mj + 4 mi
// -- one AND, one multiply-add, one shift-add.
__device__ __forceinline__ unsigned lf_index(unsigned mi, unsigned mj, unsigned k_row)
{

  return (mj & ~3u) * k_row + mj + (mi << 2);
}

// The beam loop of one group of kBlockP particles.  CLAMP = false is the path of
// particles whose beams all end inside the view (apron included); CLAMP = true
// moves end points beyond it (negative indices wrap to huge unsigned values) onto
// the last column / row, which lie in the apron and hold the off-map rank
// (!MAP_VALID end points, amcl_laser.cpp:283-286): no branch either way.
//
// LOGC: every table entry is stored 2^LOGC times, copy l in bank pair l; lane
// l & (2^LOGC - 1) reads its own copy.  With 16 copies the 8-byte reads of a
// half-warp never meet in a bank (ncu on the single-copy table: 4.4 wavefronts
// per read, the largest share of the LSU pipe that bounds this kernel).
template <int MODE, int LOGC, bool CLAMP, int TEXQ>
__device__ __forceinline__ void lf_beam_loop(const LfArgs &a, const double2 *s_beams,
                                             const double *lut_lane, const unsigned char *s_agree,
                                             int *s_count_row, int lane, int live,
                                             const double (&pX)[kBlockP], const double (&pY)[kBlockP],
                                             const double (&pc)[kBlockP], const double (&ps)[kBlockP],
                                             double (&acc)[kBlockP])
{
  const uint16_t *__restrict__ lf = a.lf;
  const unsigned k_row = a.k_row, max_i = a.max_i, max_j = a.max_j;
  constexpr int kShift = LOGC < 0 ? 0 : LOGC;
  auto load_cell = [&](int q, unsigned idx) -> unsigned {
    // the first TEXQ of the kBlockP gathers are texture fetches (TLD: the index is
    // the operand, no 64-bit address is formed), the rest global loads
    if (q < TEXQ) return tex1Dfetch<unsigned short>(a.lf_tex, (int)idx);
    return __ldg(lf + idx);
  };
  auto gather = [&](int i, unsigned (&r)[kBlockP]) {
    const double2 b = s_beams[i];
#pragma unroll
    for (int q = 0; q < kBlockP; q++)
    {
      unsigned mi = (unsigned)floor_to_int(fma(pc[q], b.x, fma(-ps[q], b.y, pX[q])));
      unsigned mj = (unsigned)floor_to_int(fma(ps[q], b.x, fma(pc[q], b.y, pY[q])));
      if (CLAMP)
      {
        mi = min(mi, max_i);
        mj = min(mj, max_j);
      }
      // 2-D array: the texture unit does the addressing (a texel fetch outside the
      // array returns 0, not the edge, so the clamp above stays)
      if (TEXQ == kTex2D)
        r[q] = texel_2d(a.lf_tex2d, (int)mi, (int)mj);
      else
        r[q] = load_cell(q, lf_index(mi, mj, k_row));
    }
  };
  auto consume = [&](int i, const unsigned (&r)[kBlockP]) {
    if (MODE == MODE_COUNT)
    {
      int agreeing = 0;
#pragma unroll
      for (int q = 0; q < kBlockP; q++)
        if (q < live) agreeing += s_agree[r[q]];
      s_count_row[i] += agreeing;
    }
#pragma unroll
    for (int q = 0; q < kBlockP; q++) acc[q] += lut_lane[r[q] << kShift];
  };
  // (Issuing the gathers of beam i + 32 before the table reads of beam i was
  // measured twice and loses: the L1TEX pipe, not the gather latency, is what
  // this loop waits for -- profiles/r2_lf_kernel.md.)
  // With texel fetches several beams per lane are in flight together.  16 Mi x 1059,
  // linear texture: one beam 14.66 ms, two 13.97, three 13.25, four 13.82 (spills);
  // 2-D array: two 12.27, three 12.75, four 12.75, five 12.40.  With loads, unrolling
  // gains nothing (15.15).
#pragma unroll (TEXQ == kTex2D ? kLfUnroll2D : (TEXQ > 0 ? kLfUnroll : 1))
  for (int i = lane; i < a.nb; i += 32)
  {
    unsigned r[kBlockP];
    gather(i, r);
    consume(i, r);
  }
}

template <int MODE, int LOGC, int TEXQ = 0>
__global__ void __launch_bounds__(kLfThreads, AMCL_LF_MINBLOCKS)
likelihood_field_kernel(const LfArgs a)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2 *s_beams = reinterpret_cast<double2 *>(smem_raw);
  double *s_lut = reinterpret_cast<double *>(smem_raw + (size_t)a.nb_pad * sizeof(double2));
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ double s_red[kLfThreads / 32];
  int *s_counts = nullptr;

  constexpr bool kLutInSmem = LOGC >= 0;
  constexpr int kCopies = kLutInSmem ? (1 << LOGC) : 1;
  const uint32_t beam_bytes = (uint32_t)a.nb_pad * (uint32_t)sizeof(double2);
  const uint32_t lut_bytes = kLutInSmem ? (uint32_t)a.nranks_pad * 8u * (uint32_t)kCopies : 0u;
  bulk_stage_begin(&s_bar, beam_bytes + lut_bytes);
  if (threadIdx.x == 0)
  {
    bulk_stage_copy(s_beams, a.beams, beam_bytes, &s_bar);
    if (kLutInSmem) bulk_stage_copy(s_lut, a.lut, lut_bytes, &s_bar);
  }
  unsigned char *s_agree = nullptr;
  if (MODE == MODE_COUNT)
  {
    // COUNT mode (first pass of beam skipping): besides the log-likelihood sum
    // over every beam it counts, per beam, the particles whose end point lies
    // within beam_skip_distance of an obstacle.  After the beams and the table
    // come one private row of per-beam counters per warp (beam i is only ever
    // touched by lane i % 32 of that warp: plain read-modify-write, no atomics)
    // and the rank -> "agrees" bytes.
    s_counts = reinterpret_cast<int *>(smem_raw + (size_t)a.nb_pad * sizeof(double2) + lut_bytes);
    const int rows = blockDim.x >> 5;
    for (int k = threadIdx.x; k < rows * a.nb; k += blockDim.x) s_counts[k] = 0;
    s_agree = reinterpret_cast<unsigned char *>(s_counts + rows * a.nb);
    for (int k = threadIdx.x; k < a.nranks_pad; k += blockDim.x) s_agree[k] = a.agree[k];
  }
  __syncthreads();  // barrier init visible to all before anyone polls it
  bulk_stage_wait(&s_bar);

  const int lane = threadIdx.x & 31;
  const double *lut_lane = (kLutInSmem ? s_lut : a.lut) + (lane & (kCopies - 1));
  const int warps_per_block = blockDim.x >> 5;
  const int warp_global = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  const int warp_stride = gridDim.x * warps_per_block;
  int *s_count_row = MODE == MODE_COUNT ? s_counts + (threadIdx.x >> 5) * a.nb : nullptr;
  double warp_total = 0.0;  // lanes 0 .. kBlockP-1: total weight of the particles they finished

  for (int p0 = warp_global * kBlockP; p0 < a.n; p0 += warp_stride * kBlockP)
  {
    // Lane q (< kBlockP) prepares particle p0+q: laser pose by
    // pf_vector_coord_add(laser_pose, pose) (pf_vector.c:106-116) -- only the
    // sine / cosine of the summed heading are needed -- and the cell-unit
    // origin MAP_GXWX / MAP_GYWY (map.h:137-138) with size/2 and the apron folded in.
    const int mine = min(p0 + (lane & (kBlockP - 1)), a.n - 1);
    const Particle s = load_particle(a.set + mine);
    double sn, cs;
    sincos(s.th, &sn, &cs);
    const double px = s.x + a.lx * cs - a.ly * sn;
    const double py = s.y + a.lx * sn + a.ly * cs;
    double st = sn, ct = cs;
    if (a.la != 0.0) sincos(s.th + a.la, &st, &ct);
    double X0 = (px - a.ox) / a.scale + a.half_x;
    double Y0 = (py - a.oy) / a.scale + a.half_y;
    if (!(fabs(X0) < kCellGuard && fabs(Y0) < kCellGuard))
    {
      X0 = -kCellGuard; Y0 = -kCellGuard; st = 0.0; ct = 0.0;   // nothing of it is on the map
    }
    const bool inside = X0 >= a.in_lo && X0 <= a.in_hi_x && Y0 >= a.in_lo && Y0 <= a.in_hi_y;
#ifdef AMCL_LF_FORCE_CLAMP   /* kernel-variant experiments only */
    const bool all_inside = false;
#else
    const bool all_inside = __all_sync(kFullMask, inside);
#endif
    double pX[kBlockP], pY[kBlockP], pc[kBlockP], ps[kBlockP], acc[kBlockP];
#pragma unroll
    for (int q = 0; q < kBlockP; q++)
    {
      pX[q] = __shfl_sync(kFullMask, X0, q);
      pY[q] = __shfl_sync(kFullMask, Y0, q);
      pc[q] = __shfl_sync(kFullMask, ct, q);
      ps[q] = __shfl_sync(kFullMask, st, q);
      acc[q] = 0.0;
    }
    const int live = min(kBlockP, a.n - p0);
    if (all_inside)
      lf_beam_loop<MODE, LOGC, false, TEXQ>(a, s_beams, lut_lane, s_agree, s_count_row, lane, live, pX, pY, pc, ps, acc);
    else
      lf_beam_loop<MODE, LOGC, true, TEXQ>(a, s_beams, lut_lane, s_agree, s_count_row, lane, live, pX, pY, pc, ps, acc);

    // Every lane receives the kBlockP totals; lane q (< kBlockP) finishes
    // particle p0 + q, the one it loaded above, so that the exp() of the LFP
    // model runs once per warp instead of once per particle.
    double mine_sum = 0.0;
#pragma unroll
    for (int q = 0; q < kBlockP; q++)
    {
      const double sum = warp_sum(acc[q]);
      if ((lane & (kBlockP - 1)) == q) mine_sum = sum;
    }
    if (lane < kBlockP && p0 + lane < a.n)
    {
      if (MODE == MODE_COUNT)
        a.logsum[p0 + lane] = mine_sum;   // the weight waits for the beam mask
      else
      {
        // LF : p = 1 + sum pz^3          (amcl_laser.cpp:241,295,298)
        // LFP: weight *= exp(sum log pz)  (amcl_laser.cpp:437,444); after a COUNT
        //      pass the beams given here are the ones to LEAVE OUT of its sum
        double factor;
        if (MODE == MODE_LF) factor = 1.0 + mine_sum;
        else factor = a.logsum ? exp(a.logsum[p0 + lane] - mine_sum) : exp(mine_sum);
        const double w = s.w * factor;
        a.set[p0 + lane].w = w;
        warp_total += w;
      }
    }
  }

  if (MODE == MODE_COUNT)
  {
    __syncthreads();
    const int rows = blockDim.x >> 5;
    for (int k = threadIdx.x; k < a.nb; k += blockDim.x)
    {
      int total = 0;
      for (int r = 0; r < rows; r++) total += s_counts[r * a.nb + k];
      if (total) atomicAdd(&a.beam_counts[k], total);
    }
  }
  else
  {
    const double t = block_sum(lane < kBlockP ? warp_total : 0.0, s_red);
    if (threadIdx.x == 0) a.block_totals[blockIdx.x] = t;
  }
}

// Sums `count` partials in index order with one block; out[slot] = sum.
__global__ void finish_partials(const double *__restrict__ partials, int count,
                                double *__restrict__ out, int slot)
{
  __shared__ double s_red[32];
  double v = 0.0;
  for (int k = threadIdx.x; k < count; k += blockDim.x) v += partials[k];
  const double t = block_sum(v, s_red);
  if (threadIdx.x == 0) out[slot] = t;
}

// ----------------------------------------------------------- beam model kernel --

struct BeamArgs
{
  Particle *set;
  int n;
  const double *tab;        // 5 doubles per beam: range, cos b, sin b, short term, tail term
  int nb;
  const uint64_t *blocked;  // 8x8 bit tiles
  int btx, sx, sy;
  double lx, ly, la;
  double ox, oy, scale;
  double range_max;
  double z_hit, hit_denom;  // 2 sigma^2
  double *block_totals;
};

// floor(t / d) for 0 <= t < 2^24, 0 < d, with d's reciprocal at hand: the float
// product is within 1e-6 of the quotient, one correction step makes it exact.
__device__ __forceinline__ int floor_div_small(int t, int d, float inv_d)
{
  int q = (int)((float)t * inv_d);
  if ((q + 1) * d <= t) q++;
  else if (q * d > t) q--;
  return q;
}

// map_calc_range (map_range.c:38-120) on the bit-tiled occupancy: returns the
// range in metres.  The current 8x8 tile word is cached in a register.
//
// The reference steps one cell at a time.  The error term of its line stepping
// has a closed form: with the invariant -dx <= 2 err < dx, after k more steps the
// minor coordinate has advanced m(k) = floor((2 (err + k dy) + dx) / (2 dx)) times
// and err has become err + k dy - m(k) dx.  When the tile under the ray is
// entirely free (word == 0; cells outside the map count as blocked in edge
// tiles) the walk therefore jumps straight to the first cell outside the tile --
// the same cells are visited afterwards, in the same state.
//
// The walk is a resumable state: begin() sets a ray up, visit() performs ONE
// iteration of the reference's loop (test a cell, then jump or step) and returns
// true with `range` set when the ray is finished.
struct RayWalk
{
  int x0, y0, x1, x, y, err, dx, dy, xs, ys, cur_tile;
  bool steep, can_jump;
  float inv_2dx, inv_2dy;
  uint64_t word;
  double range;

  __device__ __forceinline__ void begin(int ax0, int ay0, int ax1, int ay1)
  {
    steep = abs(ay1 - ay0) > abs(ax1 - ax0);
    if (steep)
    {
      int t = ax0; ax0 = ay0; ay0 = t;
      t = ax1; ax1 = ay1; ay1 = t;
    }
    x0 = ax0; y0 = ay0; x1 = ax1;
    dx = abs(ax1 - ax0);
    dy = abs(ay1 - ay0);
    xs = ax0 < ax1 ? 1 : -1;
    ys = ay0 < ay1 ? 1 : -1;
    x = ax0; y = ay0; err = 0;
    cur_tile = -1;
    word = 0;
    can_jump = dx > 0 && dx < (1 << 18);
    inv_2dx = can_jump ? 1.0f / (float)(2 * dx) : 0.0f;
    inv_2dy = (can_jump && dy > 0) ? 1.0f / (float)(2 * dy) : 0.0f;
  }

  __device__ __forceinline__ bool visit(const BeamArgs &a)
  {
    const int ci = steep ? y : x, cj = steep ? x : y;
    bool stop = !((unsigned)ci < (unsigned)a.sx && (unsigned)cj < (unsigned)a.sy);
    if (!stop)
    {
      const int tile = (cj >> 3) * a.btx + (ci >> 3);
      if (tile != cur_tile)
      {
        cur_tile = tile;
        word = __ldg(a.blocked + tile);
      }
      if (word == 0ull && can_jump)
      {
        const int rem = (x1 + xs - x) * xs;
        int k = xs > 0 ? 8 - (x & 7) : (x & 7) + 1;
        if (dy > 0)
        {
          const int my = ys > 0 ? 8 - (y & 7) : (y & 7) + 1;
          const int num = my * 2 * dx - dx - 2 * err;
          k = min(k, floor_div_small(num + 2 * dy - 1, 2 * dy, inv_2dy));
        }
        if (k > rem)
        {
          range = a.range_max;
          return true;
        }
        const int m = floor_div_small(2 * (err + k * dy) + dx, 2 * dx, inv_2dx);
        x += k * xs;
        y += m * ys;
        err += k * dy - m * dx;
        return false;
      }
      stop = (word >> (((cj & 7) << 3) | (ci & 7))) & 1ull;
    }
    if (stop)
    {
      const int ddx = x - x0, ddy = y - y0;
      range = sqrt((double)((long long)ddx * ddx + (long long)ddy * ddy)) * a.scale;
      return true;
    }
    if (x == x1 + xs)
    {
      range = a.range_max;
      return true;
    }
    x += xs;
    err += dy;
    if (2 * err >= dx)
    {
      y += ys;
      err -= dx;
    }
    return false;
  }
};

__device__ __forceinline__ double cast_ray(const BeamArgs &a, int x0, int y0, int x1, int y1)
{
  RayWalk ray;
  ray.begin(x0, y0, x1, y1);
  while (!ray.visit(a)) {}
  return ray.range;
}

// Warp per particle.  The rays of a warp end at different cells; with one ray
// per lane per round half of the lanes idle on average (ncu).  Here the warp
// hands out beams in order: whenever kRefillIdle lanes have finished their
// rays, the idle lanes take the next beams (neighbouring beams still go to the
// same warp at the same time, so their tile loads stay coherent), walk for a few
// visits, and so on.  Ranges land in a per-warp shared array indexed by beam;
// the mixture likelihood is then evaluated as uniform work in beam order.
constexpr int kRefillIdle = 16;
constexpr int kVisitsPerRound = 8;

__global__ void __launch_bounds__(kSensorThreads) beam_model_kernel(const BeamArgs a)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *s_tab = reinterpret_cast<double *>(smem_raw);
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ double s_red[kSensorThreads / 32];
  const uint32_t bytes = (uint32_t)(((size_t)a.nb * 5 * 8 + 15) & ~(size_t)15);
  bulk_stage_begin(&s_bar, bytes);
  if (threadIdx.x == 0) bulk_stage_copy(s_tab, a.tab, bytes, &s_bar);
  __syncthreads();
  bulk_stage_wait(&s_bar);

  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int warp_global = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  const int warp_stride = gridDim.x * warps_per_block;
  double *s_rng = reinterpret_cast<double *>(smem_raw + bytes) + (size_t)(threadIdx.x >> 5) * a.nb;
  double warp_total = 0.0;

  for (int p = warp_global; p < a.n; p += warp_stride)
  {
    const Particle s = load_particle(a.set + p);
    double sn, cs;
    sincos(s.th, &sn, &cs);
    const double px = s.x + a.lx * cs - a.ly * sn;
    const double py = s.y + a.lx * sn + a.ly * cs;
    double st = sn, ct = cs;
    if (a.la != 0.0) sincos(s.th + a.la, &st, &ct);
    const double fx0 = floor((px - a.ox) / a.scale + 0.5), fy0 = floor((py - a.oy) / a.scale + 0.5);
    const bool sane = fabs(fx0) < kCellGuard && fabs(fy0) < kCellGuard;
    const int x0 = sane ? (int)fx0 + a.sx / 2 : -1;
    const int y0 = sane ? (int)fy0 + a.sy / 2 : -1;

    RayWalk ray;
    int next = 0;          // next beam to hand out (warp-uniform)
    int my_beam = 0;
    bool active = false;
    for (;;)
    {
      const unsigned idle_mask = __ballot_sync(kFullMask, !active);
      const int idle = __popc(idle_mask);
      if (next < a.nb && (idle >= kRefillIdle || idle == 32))
      {
        const int rank = __popc(idle_mask & ((1u << lane) - 1u));
        if (!active && next + rank < a.nb)
        {
          my_beam = next + rank;
          const double *t = s_tab + 5 * my_beam;
          const double cdir = ct * t[1] - st * t[2];   // cos(theta + bearing)
          const double sdir = st * t[1] + ct * t[2];
          const double ex = floor((px + a.range_max * cdir - a.ox) / a.scale + 0.5);
          const double ey = floor((py + a.range_max * sdir - a.oy) / a.scale + 0.5);
          ray.begin(x0, y0, (int)fmin(fmax(ex, -kCellGuard), kCellGuard) + a.sx / 2,
                    (int)fmin(fmax(ey, -kCellGuard), kCellGuard) + a.sy / 2);
          active = true;
        }
        next = min(a.nb, next + idle);
      }
      else if (idle == 32)
        break;               // nothing in flight and nothing left to hand out
#pragma unroll 1
      for (int v = 0; v < kVisitsPerRound; v++)
        if (active && ray.visit(a))
        {
          s_rng[my_beam] = ray.range;
          active = false;
        }
    }
    __syncwarp();

    double acc = 0.0;
    for (int i = lane; i < a.nb; i += 32)
    {
      const double *t = s_tab + 5 * i;
      // amcl_laser.cpp:179-196
      const double z = t[0] - s_rng[i];
      double pz = a.z_hit * exp(-(z * z) / a.hit_denom);
      if (z < 0) pz += t[3];
      pz += t[4];
      acc += pz * pz * pz;
    }
    __syncwarp();
    const double sum = warp_sum(acc);
    if (lane == 0)
    {
      const double w = s.w * (1.0 + sum);
      a.set[p].w = w;
      warp_total += w;
    }
  }
  const double tt = block_sum(lane == 0 ? warp_total : 0.0, s_red);
  if (threadIdx.x == 0) a.block_totals[blockIdx.x] = tt;
}

// One thread per ray: the batched form of map_calc_range (map_range.c:38-120).
__global__ void calc_ranges_kernel(const BeamArgs a, const double *__restrict__ rays, int n,
                                   double *__restrict__ out)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double ox = rays[3 * (size_t)i], oy = rays[3 * (size_t)i + 1], oa = rays[3 * (size_t)i + 2];
  const double fx0 = floor((ox - a.ox) / a.scale + 0.5), fy0 = floor((oy - a.oy) / a.scale + 0.5);
  const double fx1 = floor((ox + a.range_max * cos(oa) - a.ox) / a.scale + 0.5);
  const double fy1 = floor((oy + a.range_max * sin(oa) - a.oy) / a.scale + 0.5);
  const int x0 = (int)fmin(fmax(fx0, -kCellGuard), kCellGuard) + a.sx / 2;
  const int y0 = (int)fmin(fmax(fy0, -kCellGuard), kCellGuard) + a.sy / 2;
  const int x1 = (int)fmin(fmax(fx1, -kCellGuard), kCellGuard) + a.sx / 2;
  const int y1 = (int)fmin(fmax(fy1, -kCellGuard), kCellGuard) + a.sy / 2;
  out[i] = cast_ray(a, x0, y0, x1, y1);
}

// ------------------------------------------------ normalisation (pf.c:282-316) --

// total > 0: w /= total, accumulates sum w^2 per block.  total <= 0: w = 1/n.
// total_dev: the total still sits on the device (the library's own laser models
// leave it there, so that a sensor update costs one trip to the host, not two).
__global__ void normalise_weights(Particle *set, int n, double total, double uniform,
                                  double *block_sq, const double *total_dev)
{
  __shared__ double s_red[32];
  if (total_dev) total = *total_dev;
  double sq = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    double w;
    if (total > 0.0)
    {
      w = set[i].w / total;
      sq += w * w;
    }
    else
      w = uniform;
    set[i].w = w;
  }
  const double t = block_sum(sq, s_red);
  if (threadIdx.x == 0) block_sq[blockIdx.x] = t;
}

// ------------------------------------------------------------------ launchers --

static void ensure_partials(PfCtx *c, int count)
{
  if (count <= c->partials_cap) return;
  if (c->d_partials) cudaFree(c->d_partials);
  c->partials_cap = count;
  AMCL_CUDA(cudaMalloc(&c->d_partials, sizeof(double) * (size_t)count));
}

static double read_scalar(PfCtx *c, int slot)
{
  AMCL_CUDA(cudaMemcpyAsync(c->h_scalars + slot, c->d_scalars + slot, sizeof(double),
                            cudaMemcpyDeviceToHost, globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
  return c->h_scalars[slot];
}

template <typename K>
static int persistent_grid(K kernel, int threads, size_t smem)
{
  int per_sm = 0;
  if (smem > 48 * 1024)
    AMCL_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  AMCL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
  if (per_sm < 1) AMCL_FATAL("kernel does not fit an SM with %zu bytes of shared memory", smem);
  return per_sm * globals().sm_count;
}

// `lut` is indexed by code RANK (MapCtx::code_of_rank); `lut_version` changes
// whenever its contents do, so the replicated device copy is rebuilt only then.
double run_likelihood_field(PfCtx *c, int which, MapCtx *m, int model,
                            const double laser_pose[3], const LaserScanHost &scan,
                            const std::vector<double> &lut, uint64_t lut_version,
                            const int *mask_or_null, int *agree_counts_or_null,
                            const std::vector<unsigned char> *agree_lut, bool subtract_from_logsum)
{
  cudaStream_t st = globals().stream;
  const int n_all = c->host->sets[which].sample_count;
  const int lo = c->shard.lo(n_all);
  const int n = c->shard.count(n_all);     // this rank's slice (everything when not sharded)
  // beam table: only beams that are integrated (mask applied on the host)
  std::vector<double> &tab = c->h_beam_tab;
  std::vector<int> used;
  tab.clear();
  double reach = 0.0;
  for (size_t k = 0; k < scan.a.size(); k++)
  {
    if (mask_or_null && !mask_or_null[scan.beam_ind[k]]) continue;
    tab.push_back(scan.a[k]);
    tab.push_back(scan.b[k]);
    const double len = sqrt(scan.a[k] * scan.a[k] + scan.b[k] * scan.b[k]) + 1.0;   // cells
    if (len > reach) reach = len;
    used.push_back((int)k);
  }
  const int nb = (int)used.size();
  const int nb_pad = (nb + 1) & ~1;
  tab.resize((size_t)(nb_pad > 0 ? nb_pad : 2) * 2, 0.0);
  const int nranks = (int)lut.size();
  const int nranks_pad = (nranks + 1) & ~1;
  const bool count_mode = agree_counts_or_null != nullptr;
  if (nranks != (int)m->code_of_rank.size())
    AMCL_FATAL("likelihood table has %d entries, the map's codes have %zu ranks", nranks,
               m->code_of_rank.size());
  // the laser view of the distance map, with an apron as wide as the beams reach
  if (!(reach < 1e8)) reach = 1e8;
  map_prepare_laser_view(m, (int)reach + 2);

  if ((int)tab.size() > c->scan_tab_cap)
  {
    if (c->d_scan_tab) cudaFree(c->d_scan_tab);
    c->scan_tab_cap = (int)tab.size() + 1024;
    AMCL_CUDA(cudaMalloc(&c->d_scan_tab, sizeof(double) * (size_t)c->scan_tab_cap));
  }
  AMCL_CUDA(cudaMemcpyAsync(c->d_scan_tab, tab.data(), tab.size() * sizeof(double),
                            cudaMemcpyHostToDevice, st));

  // shared memory: [beams | table x copies | COUNT: counters per warp, agree bytes]
  const size_t beam_bytes = (size_t)(nb_pad > 0 ? nb_pad : 2) * sizeof(double2);
  const size_t count_bytes =
      count_mode ? sizeof(int) * (size_t)(kLfThreads / 32) * (size_t)nb + (size_t)nranks_pad + 16 : 0;
  const size_t per_cta = (size_t)(220 / AMCL_LF_MINBLOCKS) * 1024;   // AMCL_LF_MINBLOCKS CTAs per SM
  int logc = -1;
  for (int cand = 4; cand >= 0; cand -= 2)
    if (beam_bytes + count_bytes + ((size_t)nranks_pad * 8 << cand) <= per_cta)
    {
      logc = cand;
      break;
    }
  if (logc < 0 && beam_bytes + count_bytes + (size_t)nranks_pad * 8 <= 200 * 1024) logc = 0;
  if (beam_bytes + count_bytes > 200 * 1024)
    AMCL_FATAL("a scan of %d beams needs %zu bytes of shared memory per block", nb,
               beam_bytes + count_bytes);
#ifdef AMCL_LF_FORCE_LOGC   /* kernel-variant experiments only */
  if (logc > AMCL_LF_FORCE_LOGC) logc = AMCL_LF_FORCE_LOGC;
#endif
  const int copies = logc < 0 ? 1 : (1 << logc);

  // device copy of the table, each entry repeated `copies` times
  if (c->lut_version != lut_version || c->lut_copies != copies)
  {
    const size_t need = (size_t)nranks_pad * copies;
    if (need > c->lut_cap)
    {
      if (c->d_lut) cudaFree(c->d_lut);
      c->lut_cap = need + 1024;
      AMCL_CUDA(cudaMalloc(&c->d_lut, sizeof(double) * c->lut_cap));
    }
    c->h_lut.assign(need, 0.0);
    for (int r = 0; r < nranks; r++)
      for (int k = 0; k < copies; k++) c->h_lut[(size_t)r * copies + k] = lut[r];
    AMCL_CUDA(cudaMemcpyAsync(c->d_lut, c->h_lut.data(), need * sizeof(double),
                              cudaMemcpyHostToDevice, st));
    c->lut_version = lut_version;
    c->lut_copies = copies;
  }

  unsigned char *d_agree = nullptr;
  int *d_counts = nullptr;
  if (count_mode)
  {
    size_t bytes = (size_t)nranks_pad + sizeof(int) * (size_t)(nb + 1);
    if ((int)bytes > c->skip_cap)
    {
      if (c->d_skip) cudaFree(c->d_skip);
      c->skip_cap = (int)bytes + 4096;
      AMCL_CUDA(cudaMalloc(&c->d_skip, (size_t)c->skip_cap));
    }
    d_counts = reinterpret_cast<int *>(c->d_skip);
    d_agree = reinterpret_cast<unsigned char *>(c->d_skip) + sizeof(int) * (size_t)(nb + 1);
    AMCL_CUDA(cudaMemsetAsync(d_counts, 0, sizeof(int) * (size_t)(nb + 1), st));
    AMCL_CUDA(cudaMemcpyAsync(d_agree, agree_lut->data(), agree_lut->size(),
                              cudaMemcpyHostToDevice, st));
  }

  LfArgs a;
  a.set = c->d_set[which] + lo;
  a.n = n;
  a.beams = reinterpret_cast<const double2 *>(c->d_scan_tab);
  a.nb = nb;
  a.nb_pad = nb_pad > 0 ? nb_pad : 2;
  a.lut = c->d_lut;
  a.agree = d_agree;
  a.nranks_pad = nranks_pad;
  a.lf = m->d_lf;
  a.lf_tex = m->lf_tex;
  a.lf_tex2d = m->lf_tex2d;
  a.max_i = (unsigned)m->lf_w - 1u;
  a.max_j = (unsigned)m->lf_h - 1u;
  a.k_row = (unsigned)m->lf_w - 1u;
  a.lx = laser_pose[0];
  a.ly = laser_pose[1];
  a.la = laser_pose[2];
  a.ox = m->origin_x;
  a.oy = m->origin_y;
  a.scale = m->scale;
  a.half_x = 0.5 + (double)(m->sx / 2) + (double)m->lf_apron;
  a.half_y = 0.5 + (double)(m->sy / 2) + (double)m->lf_apron;
  a.in_lo = reach + 1.0;
  a.in_hi_x = (double)m->lf_w - reach - 1.0;
  a.in_hi_y = (double)m->lf_h - reach - 1.0;
  a.beam_counts = d_counts;

  const size_t smem = beam_bytes + (logc >= 0 ? (size_t)nranks_pad * 8 * copies : 0) + count_bytes;
  a.logsum = nullptr;
  if (count_mode || subtract_from_logsum)
  {
    if (!c->d_logsum) AMCL_CUDA(cudaMalloc(&c->d_logsum, sizeof(double) * (size_t)c->alloc_samples));
    a.logsum = c->d_logsum + lo;
  }

  int grid = 1;
  if (!c->ev_a)
  {
    AMCL_CUDA(cudaEventCreate(&c->ev_a));
    AMCL_CUDA(cudaEventCreate(&c->ev_b));
  }
  AMCL_CUDA(cudaEventRecord(c->ev_a, st));
  // The map gathers are texel fetches with three beams per lane in flight.  Measured
  // at 16 Mi x 1059 (profiles/r2_lf_kernel.md), likelihood_field / likelihood_field_prob:
  // loads 15.52 / 15.10 ms, fetches from the linear texture 13.26 / 14.06 ms, fetches
  // from the 2-D array 12.78 / 12.93 ms.  The counting pass of beam skipping keeps the
  // loads; a view too large for an array falls back to the linear texture, then to loads.
  // AMCLB200_LF_TEXQ=0|8|100 overrides (A/B timing).
  static const int texq_env = [] {
    const char *e = getenv("AMCLB200_LF_TEXQ");
    return e ? atoi(e) : -1;
  }();
  int texq = texq_env >= 0 ? texq_env : kTex2D;
  if (texq == kTex2D && !m->lf_tex2d) texq = kBlockP;
  if ((texq != kTex2D && !m->lf_tex) || logc != 4 || count_mode) texq = 0;
#define AMCL_LF_LAUNCH(MODE, LOGC, TEXQ)                                                 \
  do                                                                                     \
  {                                                                                      \
    grid = persistent_grid(likelihood_field_kernel<MODE, LOGC, TEXQ>, kLfThreads, smem); \
    int need_blocks = (n + (kLfThreads / 32) * kBlockP - 1) / ((kLfThreads / 32) * kBlockP); \
    if (grid > need_blocks) grid = need_blocks > 0 ? need_blocks : 1;                    \
    ensure_partials(c, grid);                                                            \
    a.block_totals = c->d_partials;                                                      \
    likelihood_field_kernel<MODE, LOGC, TEXQ><<<grid, kLfThreads, smem, st>>>(a);        \
  } while (0)
#define AMCL_LF_DISPATCH(MODE)                       \
  do                                                 \
  {                                                  \
    if (logc == 4 && texq == kTex2D) AMCL_LF_LAUNCH(MODE, 4, kTex2D);   \
    else if (logc == 4 && texq == kBlockP) AMCL_LF_LAUNCH(MODE, 4, kBlockP); \
    else if (logc == 4) AMCL_LF_LAUNCH(MODE, 4, 0);  \
    else if (logc == 2) AMCL_LF_LAUNCH(MODE, 2, 0);  \
    else if (logc == 0) AMCL_LF_LAUNCH(MODE, 0, 0);  \
    else AMCL_LF_LAUNCH(MODE, -1, 0);                \
  } while (0)

  if (count_mode) AMCL_LF_DISPATCH(MODE_COUNT);
  else if (model == 1) AMCL_LF_DISPATCH(MODE_LF);
  else AMCL_LF_DISPATCH(MODE_LFP);
#undef AMCL_LF_DISPATCH
#undef AMCL_LF_LAUNCH

  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaEventRecord(c->ev_b, st));

  if (count_mode)
  {
    if (c->shard.on() && nb > 0) c->shard.allreduce(d_counts, nb, SHARD_I32, SHARD_SUM, c->shard.user);
    std::vector<int> counts((size_t)nb + 1);
    AMCL_CUDA(cudaMemcpyAsync(counts.data(), d_counts, sizeof(int) * (size_t)nb,
                              cudaMemcpyDeviceToHost, st));
    AMCL_CUDA(cudaStreamSynchronize(st));
    for (int k = 0; k < nb; k++) agree_counts_or_null[scan.beam_ind[used[k]]] = counts[k];
    return 0.0;
  }
  finish_partials<<<1, 256, 0, st>>>(c->d_partials, grid, c->d_scalars, 0);
  AMCL_LAUNCH_CHECK();
  if (c->shard.on())
  {
    c->shard.allreduce(c->d_scalars, 1, SHARD_F64, SHARD_SUM, c->shard.user);
    c->replicated[which] = false;
  }
  c->sensor_evals = (long long)n * nb;
  c->host_fresh[which] = false;
  c->sensor_kernel_pending = true;         // ev_a / ev_b are read when somebody asks for the time
  if (c->defer_total) return 0.0;          // pf_update_sensor picks the total up together with n_effective
  return read_scalar(c, 0);
}

double run_beam_model(PfCtx *c, int which, MapCtx *m, const double laser_pose[3],
                      const LaserScanHost &scan, double range_max, const double zp[6])
{
  cudaStream_t st = globals().stream;
  const int n_all = c->host->sets[which].sample_count;
  const int lo = c->shard.lo(n_all);
  const int n = c->shard.count(n_all);
  const int nb = (int)scan.range.size();
  // zp: z_hit, z_short, z_max, z_rand, sigma_hit, lambda_short
  std::vector<double> tab((size_t)(nb > 0 ? nb : 1) * 5 + 2, 0.0);
  for (int k = 0; k < nb; k++)
  {
    const double r = scan.range[k];
    tab[5 * k + 0] = r;
    tab[5 * k + 1] = scan.cb[k];
    tab[5 * k + 2] = scan.sb[k];
    tab[5 * k + 3] = zp[1] * zp[5] * exp(-zp[5] * r);          // amcl_laser.cpp:185
    double tail = 0.0;
    if (r == range_max) tail += zp[2] * 1.0;                     // :189
    if (r < range_max) tail += zp[3] * 1.0 / range_max;          // :193
    tab[5 * k + 4] = tail;
  }
  if ((int)tab.size() > c->scan_tab_cap)
  {
    if (c->d_scan_tab) cudaFree(c->d_scan_tab);
    c->scan_tab_cap = (int)tab.size() + 1024;
    AMCL_CUDA(cudaMalloc(&c->d_scan_tab, sizeof(double) * (size_t)c->scan_tab_cap));
  }
  AMCL_CUDA(cudaMemcpyAsync(c->d_scan_tab, tab.data(), tab.size() * sizeof(double),
                            cudaMemcpyHostToDevice, st));
  BeamArgs a;
  a.set = c->d_set[which] + lo;
  a.n = n;
  a.tab = c->d_scan_tab;
  a.nb = nb;
  a.blocked = m->d_blocked;
  a.btx = m->btiles_x;
  a.sx = m->sx;
  a.sy = m->sy;
  a.lx = laser_pose[0];
  a.ly = laser_pose[1];
  a.la = laser_pose[2];
  a.ox = m->origin_x;
  a.oy = m->origin_y;
  a.scale = m->scale;
  a.range_max = range_max;
  a.z_hit = zp[0];
  a.hit_denom = 2 * zp[4] * zp[4];
  // beam table, then one row of ranges per warp
  const size_t smem = (((size_t)nb * 5 * 8 + 15) & ~(size_t)15) +
                      (size_t)(kSensorThreads / 32) * (size_t)(nb > 0 ? nb : 1) * sizeof(double);
  int grid = persistent_grid(beam_model_kernel, kSensorThreads, smem);
  int need_blocks = (n + (kSensorThreads / 32) - 1) / (kSensorThreads / 32);
  if (grid > need_blocks) grid = need_blocks > 0 ? need_blocks : 1;
  ensure_partials(c, grid);
  a.block_totals = c->d_partials;
  beam_model_kernel<<<grid, kSensorThreads, smem, st>>>(a);
  AMCL_LAUNCH_CHECK();
  finish_partials<<<1, 256, 0, st>>>(c->d_partials, grid, c->d_scalars, 0);
  AMCL_LAUNCH_CHECK();
  if (c->shard.on())
  {
    c->shard.allreduce(c->d_scalars, 1, SHARD_F64, SHARD_SUM, c->shard.user);
    c->replicated[which] = false;
  }
  c->host_fresh[which] = false;
  if (c->defer_total) return 0.0;
  return read_scalar(c, 0);
}

void run_sensor_normalise(PfCtx *c, int which, double total)
{
  cudaStream_t st = globals().stream;
  pf_t *pf = c->host;
  pf_sample_set_t *set = pf->sets + which;
  const int n = set->sample_count;                 // the global count
  // a foreign host model rewrote the whole (replicated) set: normalise all of it
  const bool slice = c->shard.on() && !c->replicated[which];
  const int lo = slice ? c->shard.lo(n) : 0;
  const int cnt = slice ? c->shard.count(n) : n;
  int grid = globals().sm_count * 4;
  int need_blocks = (cnt + 255) / 256;
  if (grid > need_blocks) grid = need_blocks > 0 ? need_blocks : 1;
  ensure_partials(c, grid);
  const bool deferred = c->defer_total;
  normalise_weights<<<grid, 256, 0, st>>>(c->d_set[which] + lo, cnt, total, 1.0 / n, c->d_partials,
                                          deferred ? c->d_scalars : nullptr);
  AMCL_LAUNCH_CHECK();
  double neff = 0.0;
  if (deferred)
  {
    // total (slot 0, left there by the laser model) and the sum of squares (slot 1)
    // come back together: the one trip to the host of this update
    finish_partials<<<1, 256, 0, st>>>(c->d_partials, grid, c->d_scalars, 1);
    AMCL_LAUNCH_CHECK();
    if (slice) c->shard.allreduce(c->d_scalars + 1, 1, SHARD_F64, SHARD_SUM, c->shard.user);
    AMCL_CUDA(cudaMemcpyAsync(c->h_scalars, c->d_scalars, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    AMCL_CUDA(cudaStreamSynchronize(st));
    total = c->h_scalars[0];
  }
  if (total > 0.0)
  {
    if (deferred)
      neff = c->h_scalars[1];
    else
    {
      finish_partials<<<1, 256, 0, st>>>(c->d_partials, grid, c->d_scalars, 1);
      AMCL_LAUNCH_CHECK();
      if (slice) c->shard.allreduce(c->d_scalars + 1, 1, SHARD_F64, SHARD_SUM, c->shard.user);
      neff = read_scalar(c, 1);
    }
    // pf.c:293-302.  w_avg accumulates the very numbers the model summed into
    // `total`, in the same order, so it equals total / n.
    double w_avg = total / n;
    if (pf->w_slow == 0.0) pf->w_slow = w_avg;
    else pf->w_slow += pf->alpha_slow * (w_avg - pf->w_slow);
    if (pf->w_fast == 0.0) pf->w_fast = w_avg;
    else pf->w_fast += pf->alpha_fast * (w_avg - pf->w_fast);
  }
  set->n_effective = 1.0 / neff;  // pf.c:316 (1/0 -> inf on the zero-total path)
  c->host_fresh[which] = false;
}

void run_calc_ranges(MapCtx *m, const double *rays_xya, int n, double max_range, double *out)
{
  cudaStream_t st = globals().stream;
  if (n <= 0) return;
  double *d_rays = nullptr, *d_out = nullptr;
  AMCL_CUDA(cudaMalloc(&d_rays, sizeof(double) * 3 * (size_t)n));
  AMCL_CUDA(cudaMalloc(&d_out, sizeof(double) * (size_t)n));
  AMCL_CUDA(cudaMemcpyAsync(d_rays, rays_xya, sizeof(double) * 3 * (size_t)n,
                            cudaMemcpyHostToDevice, st));
  BeamArgs a;
  memset(&a, 0, sizeof(a));
  a.blocked = m->d_blocked;
  a.btx = m->btiles_x;
  a.sx = m->sx;
  a.sy = m->sy;
  a.ox = m->origin_x;
  a.oy = m->origin_y;
  a.scale = m->scale;
  a.range_max = max_range;
  calc_ranges_kernel<<<(n + 127) / 128, 128, 0, st>>>(a, d_rays, n, d_out);
  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaMemcpyAsync(out, d_out, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  cudaFree(d_rays);
  cudaFree(d_out);
}

}  // namespace amclb200

// ====================================================== amcl::AMCLLaser (C++) ==

using namespace amclb200;

namespace amcl
{

// The private block of AMCLLaser, reachable through the friend declaration in
// the header.
struct LaserAccess
{
  static int model(const AMCLLaser *l) { return (int)l->model_type; }
  static map_t *map(const AMCLLaser *l) { return l->map; }
  static int max_beams(const AMCLLaser *l) { return l->max_beams; }
  static const pf_vector_t &pose(const AMCLLaser *l) { return l->laser_pose; }
  static void params(const AMCLLaser *l, double zp[6])
  {
    zp[0] = l->z_hit; zp[1] = l->z_short; zp[2] = l->z_max;
    zp[3] = l->z_rand; zp[4] = l->sigma_hit; zp[5] = l->lambda_short;
  }
  static bool is_ours(pf_sensor_model_fn_t fn)
  {
    return fn == (pf_sensor_model_fn_t)AMCLLaser::BeamModel ||
           fn == (pf_sensor_model_fn_t)AMCLLaser::LikelihoodFieldModel ||
           fn == (pf_sensor_model_fn_t)AMCLLaser::LikelihoodFieldModelProb;
  }
  static void skip(const AMCLLaser *l, bool *on, double v[3])
  {
    *on = l->do_beamskip;
    v[0] = l->beam_skip_distance;
    v[1] = l->beam_skip_threshold;
    v[2] = l->beam_skip_error_threshold;
  }
};

// pf_update_sensor must know whether a model function is one of ours (runs on
// the device-resident set) or a foreign host callback (needs the host mirror).
bool amclb200_is_device_laser_model(pf_sensor_model_fn_t fn)
{
  return LaserAccess::is_ours(fn);
}

// amcl_sensor.cpp:37-95
AMCLSensor::AMCLSensor() {}
AMCLSensor::~AMCLSensor() {}
bool AMCLSensor::UpdateAction(pf_t *, AMCLSensorData *) { return false; }
bool AMCLSensor::InitSensor(pf_t *, AMCLSensorData *) { return false; }
bool AMCLSensor::UpdateSensor(pf_t *, AMCLSensorData *) { return false; }

// amcl_laser.cpp:44-55
AMCLLaser::AMCLLaser(size_t max_beams, map_t *map)
    : AMCLSensor(), max_samples(0), max_obs(0), temp_obs(NULL)
{
  this->time = 0.0;
  this->max_beams = max_beams;
  this->map = map;
  this->model_type = LASER_MODEL_BEAM;
  this->do_beamskip = false;
}

// amcl_laser.cpp:56-64: nothing to release, temp_obs is never allocated here.
AMCLLaser::~AMCLLaser() {}

// amcl_laser.cpp:66-83
void AMCLLaser::SetModelBeam(double z_hit, double z_short, double z_max, double z_rand,
                             double sigma_hit, double lambda_short, double chi_outlier)
{
  this->model_type = LASER_MODEL_BEAM;
  this->z_hit = z_hit;
  this->z_short = z_short;
  this->z_max = z_max;
  this->z_rand = z_rand;
  this->sigma_hit = sigma_hit;
  this->lambda_short = lambda_short;
  this->chi_outlier = chi_outlier;
}

// amcl_laser.cpp:85-97
void AMCLLaser::SetModelLikelihoodField(double z_hit, double z_rand, double sigma_hit,
                                        double max_occ_dist)
{
  this->model_type = LASER_MODEL_LIKELIHOOD_FIELD;
  this->z_hit = z_hit;
  this->z_rand = z_rand;
  this->sigma_hit = sigma_hit;
  map_update_cspace(this->map, max_occ_dist);
}

// amcl_laser.cpp:99-118
void AMCLLaser::SetModelLikelihoodFieldProb(double z_hit, double z_rand, double sigma_hit,
                                            double max_occ_dist, bool do_beamskip,
                                            double beam_skip_distance,
                                            double beam_skip_threshold,
                                            double beam_skip_error_threshold)
{
  this->model_type = LASER_MODEL_LIKELIHOOD_FIELD_PROB;
  this->z_hit = z_hit;
  this->z_rand = z_rand;
  this->sigma_hit = sigma_hit;
  this->do_beamskip = do_beamskip;
  this->beam_skip_distance = beam_skip_distance;
  this->beam_skip_threshold = beam_skip_threshold;
  this->beam_skip_error_threshold = beam_skip_error_threshold;
  map_update_cspace(this->map, max_occ_dist);
}

// amcl_laser.cpp:496-510.  Kept for link compatibility; the device path needs
// no per-(particle, beam) scratch, so this only records the sizes.
void AMCLLaser::reallocTempData(int new_max_samples, int new_max_obs)
{
  max_obs = new_max_obs;
  max_samples = new_max_samples > max_samples ? new_max_samples : max_samples;
}

// amcl_laser.cpp:123-139
bool AMCLLaser::UpdateSensor(pf_t *pf, AMCLSensorData *data)
{
  if (this->max_beams < 2) return false;
  if (this->model_type == LASER_MODEL_LIKELIHOOD_FIELD)
    pf_update_sensor(pf, (pf_sensor_model_fn_t)LikelihoodFieldModel, data);
  else if (this->model_type == LASER_MODEL_LIKELIHOOD_FIELD_PROB)
    pf_update_sensor(pf, (pf_sensor_model_fn_t)LikelihoodFieldModelProb, data);
  else
    pf_update_sensor(pf, (pf_sensor_model_fn_t)BeamModel, data);
  return true;
}

// Host-side preparation shared by the three models: walk the scan with the
// reference's stride and skip rules and tabulate the per-beam constants.
static void tabulate_scan(AMCLLaserData *data, int step, bool skip_invalid, double scale,
                          LaserScanHost *out)
{
  int beam_ind = 0;
  for (int i = 0; i < data->range_count; i += step, beam_ind++)
  {
    const double r = data->ranges[i][0], b = data->ranges[i][1];
    if (skip_invalid)
    {
      if (r >= data->range_max) continue;   // amcl_laser.cpp:258-259, 383-385
      if (r != r) continue;                 // :262-263, 388-390
    }
    const double cb = cos(b), sb = sin(b);
    out->a.push_back(r * cb / scale);
    out->b.push_back(r * sb / scale);
    out->range.push_back(r);
    out->cb.push_back(cb);
    out->sb.push_back(sb);
    out->beam_ind.push_back(beam_ind);
  }
}

// Versions of the rank -> likelihood tables: a fresh number whenever one is rebuilt.
static uint64_t next_lut_version()
{
  static uint64_t v = 0;
  return ++v;
}

static MapCtx *ready_map(map_t *map, bool need_dist)
{
  MapCtx *m = map_ctx(map, true);
  if (!m->occ_uploaded) map_upload_occupancy(m);
  if (need_dist && !m->dist_valid)
    AMCL_FATAL("likelihood-field model used before map_update_cspace()");
  return m;
}

// amcl_laser.cpp:215-303
double AMCLLaser::LikelihoodFieldModel(AMCLLaserData *data, pf_sample_set_t *set)
{
  AMCLLaser *self = (AMCLLaser *)data->sensor;
  int which = 0;
  PfCtx *c = pf_ctx_of_set(set, &which);
  MapCtx *m = ready_map(self->map, true);
  int step = (data->range_count - 1) / (self->max_beams - 1);
  if (step < 1) step = 1;
  LaserScanHost scan;
  tabulate_scan(data, step, true, m->scale, &scan);
  // code -> pz^3 with the host libm in the reference's expression order; the
  // table only depends on the model parameters, range_max and the map's code
  // table, so it is rebuilt only when one of those changes
  const double z_hit_denom = 2 * self->sigma_hit * self->sigma_hit;
  const double z_rand_mult = 1.0 / data->range_max;
  static std::vector<double> lut;
  static uint64_t lut_version = 0;
  static double lut_key[7] = {0, 0, 0, 0, -1, 0, 0};
  // one entry per code rank; dist_of_code is sqrt(code) * scale for codes 0..R^2,
  // then max_occ_dist, which is also what an off-map end point gets (:278-281)
  const double key[7] = {self->z_hit, self->z_rand, self->sigma_hit, data->range_max,
                         (double)m->R, m->scale, m->max_occ_dist};
  if (memcmp(key, lut_key, sizeof(key)) != 0 || lut.size() != m->code_of_rank.size())
  {
    lut.assign(m->code_of_rank.size(), 0.0);
    for (size_t r = 0; r < lut.size(); r++)
    {
      const size_t k = (size_t)m->code_of_rank[r];
      const double z = k < m->dist_of_code.size() ? m->dist_of_code[k] : m->max_occ_dist;
      double pz = 0.0;
      pz += self->z_hit * exp(-(z * z) / z_hit_denom);
      pz += self->z_rand * z_rand_mult;
      lut[r] = pz * pz * pz;
    }
    memcpy(lut_key, key, sizeof(key));
    lut_version = next_lut_version();
  }
  return run_likelihood_field(c, which, m, 1, self->laser_pose.v, scan, lut, lut_version, nullptr,
                              nullptr, nullptr);
}

// amcl_laser.cpp:305-494
double AMCLLaser::LikelihoodFieldModelProb(AMCLLaserData *data, pf_sample_set_t *set)
{
  AMCLLaser *self = (AMCLLaser *)data->sensor;
  int which = 0;
  PfCtx *c = pf_ctx_of_set(set, &which);
  MapCtx *m = ready_map(self->map, true);
  int step = ceil((data->range_count) / static_cast<double>(self->max_beams));
  if (step < 1) step = 1;
  LaserScanHost scan;
  tabulate_scan(data, step, true, m->scale, &scan);
  const double z_hit_denom = 2 * self->sigma_hit * self->sigma_hit;
  const double z_rand_mult = 1.0 / data->range_max;
  const double max_dist_prob = exp(-(m->max_occ_dist * m->max_occ_dist) / z_hit_denom);
  // code -> log(pz).  Last entry = off-map (:414-416).  Cached like the
  // likelihood_field table.
  static std::vector<double> lut;
  static std::vector<unsigned char> agree;
  static uint64_t lut_version = 0;
  static double lut_key[8] = {0, 0, 0, 0, -1, 0, 0, 0};
  const double key[8] = {self->z_hit, self->z_rand, self->sigma_hit, data->range_max,
                         (double)m->R, m->scale, m->max_occ_dist, self->beam_skip_distance};
  if (memcmp(key, lut_key, sizeof(key)) != 0 || lut.size() != m->code_of_rank.size())
  {
    lut.assign(m->code_of_rank.size(), 0.0);
    agree.assign(((lut.size() + 1) & ~(size_t)1), 0);
    for (size_t r = 0; r < lut.size(); r++)
    {
      const size_t k = (size_t)m->code_of_rank[r];
      double pz = 0.0;
      if (k < m->dist_of_code.size())
      {
        const double z = m->dist_of_code[k];
        agree[r] = z < self->beam_skip_distance ? 1 : 0;   // :419-421
        pz += self->z_hit * exp(-(z * z) / z_hit_denom);
      }
      else
        pz += self->z_hit * max_dist_prob;
      pz += self->z_rand * z_rand_mult;
      lut[r] = log(pz);
    }
    memcpy(lut_key, key, sizeof(key));
    lut_version = next_lut_version();
  }
  const bool do_beamskip = self->do_beamskip && set->converged;   // :342-344
  if (!do_beamskip)
    return run_likelihood_field(c, which, m, 2, self->laser_pose.v, scan, lut, lut_version, nullptr,
                                nullptr, nullptr);

  // Beam skipping (:446-489) without the reference's N x B host scratch.  Pass 1
  // goes over every beam once: per beam it counts the particles whose end point
  // lies within beam_skip_distance of an obstacle, per particle it keeps the sum
  // of log pz over ALL beams.  The host picks the beams as the reference does;
  // pass 2 only visits the beams that are to be left out and applies
  // weight *= exp(sum_all - sum_left_out) -- nothing but that N-sized finish when
  // every beam is kept.
  const int slots = self->max_beams;
  std::vector<int> obs_count(slots, 0), obs_mask(slots, 0);
  for (size_t k = 0; k < scan.beam_ind.size(); k++)
    if (scan.beam_ind[k] >= slots) AMCL_FATAL("beam index beyond max_beams");
  run_likelihood_field(c, which, m, 2, self->laser_pose.v, scan, lut, lut_version, nullptr,
                       obs_count.data(), &agree);
  int skipped = 0;
  for (int b = 0; b < slots; b++)
  {
    if ((obs_count[b] / static_cast<double>(set->sample_count)) > self->beam_skip_threshold)
      obs_mask[b] = 1;
    else
      skipped++;
  }
  bool error = false;
  if (skipped >= (slots * self->beam_skip_error_threshold))
  {
    fprintf(stderr, "Over %f%% of the observations were not in the map - pf may have converged "
                    "to wrong pose - integrating all observations\n",
            (100 * self->beam_skip_error_threshold));
    error = true;
  }
  // Readings the scan loop skipped (max range / NaN) are not integrated in
  // either case: the reference would read stale scratch for them when `error`
  // is set (:479-482), which has no defined value.  See DESIGN.md.
  std::vector<int> leave_out(slots, 0);
  if (!error)
    for (int b = 0; b < slots; b++) leave_out[b] = obs_mask[b] ? 0 : 1;
  return run_likelihood_field(c, which, m, 2, self->laser_pose.v, scan, lut, lut_version,
                              leave_out.data(), nullptr, nullptr, true);
}

// amcl_laser.cpp:144-213
double AMCLLaser::BeamModel(AMCLLaserData *data, pf_sample_set_t *set)
{
  AMCLLaser *self = (AMCLLaser *)data->sensor;
  int which = 0;
  PfCtx *c = pf_ctx_of_set(set, &which);
  MapCtx *m = ready_map(self->map, false);
  int step = (data->range_count - 1) / (self->max_beams - 1);
  if (step < 1) step = 1;   // the reference would loop forever here (:171)
  LaserScanHost scan;
  tabulate_scan(data, step, false, m->scale, &scan);
  double zp[6];
  LaserAccess::params(self, zp);
  return run_beam_model(c, which, m, self->laser_pose.v, scan, data->range_max, zp);
}

}  // namespace amcl
