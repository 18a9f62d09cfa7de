// cumsum_dev.cu -- the cumulative weight table of pf_update_resample
// (/root/reference/amcl/src/amcl/pf/pf.c:399-402), computed in parallel but
// BIT-IDENTICAL to the reference's sequential  c[i+1] = fl(c[i] + w[i]).
//
// Why it matters: a draw r picks the index i with
// c[i] <= r < c[i+1].  A parallel fp64 scan rounds differently from the
// sequential sum, so at 16 M particles ~100 draws per resample would land on a
// neighbouring particle.  "Resampled indices bit-exact" needs the exact table.
//
// How.  While the running sum stays inside one binade [2^e, 2^(e+1)) it is an
// integer multiple C of u = 2^(e-52), and adding w rounds to
//     C' = C + W + (f > 1/2) ,   w/u = W + f ,
// with the tie f == 1/2 going to the even neighbour, i.e. depending on the
// parity of C + W.  So every element is a function "C -> C + A[parity of C]"
// described by a pair (A[0], A[1]), and such functions compose associatively:
//     (g then f)[p] = g[p] + f[(p + g[p]) & 1].
// Integer pair functions can be scanned in parallel with no rounding at all.
//
// The only places where this breaks are the crossings from one binade into the
// next (a few dozen per table).  The table is cut into tiles of 1024 weights;
// a rough fp64 prefix of the tile sums, widened by a rigorous error margin,
// tells which tiles certainly lie inside one binade ("clean") and which might
// contain a crossing ("dirty", typically ~20 of 16 K tiles, always including
// tile 0).  Clean tiles go through the integer scan; dirty tiles are walked
// element by element with real fp64 additions by one thread.  Finally EVERY
// entry is checked against the defining recurrence c[i+1] == fl(c[i] + w[i]);
// by induction that proves the whole table equals the sequential one.  Should
// the check ever fail (it cannot, unless the margin reasoning is violated by
// non-finite or negative weights) the single-thread walk is used instead.
#include <math.h>

#include "context.hpp"

namespace amclb200
{

void launch_cumsum_sequential(Weights W, int n, double *d_cum);   // resample_dev.cu

namespace
{

constexpr int kTileLen = 1024;
constexpr int kTileThreads = 256;
constexpr int kPerThread = kTileLen / kTileThreads;   // 4 consecutive weights per thread
constexpr int kDirty = -100000;
constexpr int kFlat = -200000;    // every weight of the tile is exactly 0: the table does not move

struct PairFn
{
  unsigned long long a0, a1;   // increment when the incoming C is even / odd
};

__device__ __forceinline__ PairFn fn_identity()
{
  PairFn r;
  r.a0 = 0ull;
  r.a1 = 0ull;
  return r;
}

// g is applied first, f second.
__device__ __forceinline__ PairFn fn_compose(const PairFn &g, const PairFn &f)
{
  PairFn r;
  r.a0 = g.a0 + ((g.a0 & 1ull) ? f.a1 : f.a0);
  r.a1 = g.a1 + (((1ull + g.a1) & 1ull) ? f.a1 : f.a0);
  return r;
}

__device__ __forceinline__ unsigned long long fn_apply(const PairFn &f, unsigned long long C)
{
  return C + ((C & 1ull) ? f.a1 : f.a0);
}

// The function of one weight inside binade e (shift = 52 - e).
__device__ __forceinline__ PairFn fn_of_weight(double w, int shift)
{
  const double q = scalbn(w, shift);                 // exact: power-of-two scaling
  const unsigned long long W = (unsigned long long)q;   // q >= 0: truncation == floor
  const double frac = q - (double)W;                 // exact
  PairFn r;
  if (frac > 0.5) { r.a0 = W + 1ull; r.a1 = W + 1ull; }
  else if (frac < 0.5) { r.a0 = W; r.a1 = W; }
  else { r.a0 = W + (W & 1ull); r.a1 = W + ((W + 1ull) & 1ull); }   // tie -> even
  return r;
}

__device__ __forceinline__ PairFn fn_shfl_up(const PairFn &v, int delta)
{
  PairFn r;
  r.a0 = __shfl_up_sync(kFullMask, v.a0, delta);
  r.a1 = __shfl_up_sync(kFullMask, v.a1, delta);
  return r;
}

// Block-wide exclusive scan (in thread order) under fn_compose.  s_warp needs
// 33 entries.  Returns the exclusive prefix; *total = composition of all.
__device__ PairFn block_scan_fn(PairFn v, PairFn *total, PairFn *s_warp)
{
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int nw = blockDim.x >> 5;
  PairFn inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    PairFn t = fn_shfl_up(inc, o);
    if (lane >= o) inc = fn_compose(t, inc);
  }
  __syncthreads();            // protects s_warp against a previous use
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  if (wid == 0)
  {
    PairFn w = lane < nw ? s_warp[lane] : fn_identity();
    PairFn winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
      PairFn t = fn_shfl_up(winc, o);
      if (lane >= o) winc = fn_compose(t, winc);
    }
    PairFn wex = fn_shfl_up(winc, 1);
    if (lane == 0) wex = fn_identity();
    s_warp[lane] = wex;                       // exclusive warp prefixes
    if (lane == 31) s_warp[32] = winc;        // everything
  }
  __syncthreads();
  *total = s_warp[32];
  PairFn ex = fn_shfl_up(inc, 1);
  if (lane == 0) ex = fn_identity();
  return fn_compose(s_warp[wid], ex);
}

// ---- A: fp64 sum of each tile (any fixed order; only used for classification)
// When the weights sit inside 32-byte particle records (stride 4) this pass also
// writes them out contiguously: the four later passes then read 8 B per weight
// instead of dragging the whole record through the memory system each time.
__global__ void cs_tile_sums(const Weights W, int n, double *__restrict__ tile_sum,
                             double *__restrict__ tile_max, double *__restrict__ packed, int tile0,
                             int ts /* element stride of tile_sum / tile_max: 2 when they interleave */)
{
  __shared__ double s_red[32];
  const int tile = (int)blockIdx.x + tile0;
  const int base = tile * kTileLen;
  double v = 0.0, mx = 0.0;
#pragma unroll
  for (int k = 0; k < kPerThread; k++)
  {
    const int i = base + k * kTileThreads + threadIdx.x;
    if (i < n)
    {
      const double w = W[i];
      if (packed) packed[i] = w;
      v += w;
      mx = fmax(mx, w);
    }
  }
  const double t = block_sum(v, s_red);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(kFullMask, mx, o));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    for (int k = 1; k < (int)(blockDim.x >> 5); k++) mx = fmax(mx, s_red[k]);
    tile_sum[(size_t)tile * ts] = t;
    tile_max[(size_t)tile * ts] = mx;
  }
}

// ---- B: rough prefix of the tile sums and clean/dirty classification ---------
// dirty_row (sharded table only): tile -> row of the buffer in which the owners of
// the dirty tiles publish their weights; a count beyond max_rows raises *overflow.
__global__ void cs_classify(const double *__restrict__ tile_sum, const double *__restrict__ tile_max,
                            int tiles, int n, int *__restrict__ tile_class, int *dirty_count,
                            int *__restrict__ dirty_row, int max_rows, int *overflow, int ts)
{
  __shared__ double s_part[1024];
  // pass 1: grand total (each thread a contiguous slab, then a serial combine)
  const int per = (tiles + blockDim.x - 1) / blockDim.x;
  const int lo = threadIdx.x * per, hi = min(tiles, lo + per);
  double acc = 0.0;
  for (int t = lo; t < hi; t++) acc += tile_sum[(size_t)t * ts];
  s_part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    double run = 0.0;
    for (int k = 0; k < (int)blockDim.x; k++)
    {
      const double v = s_part[k];
      s_part[k] = run;            // exclusive prefix of the slabs
      run += v;
    }
  }
  __syncthreads();
  // Weights are non-negative, so every partial sum -- the sequential one and this
  // rough prefix alike -- carries a RELATIVE error of at most (#additions) * 2^-53
  // <= n * 2^-53.  The margin below is four times that, relative to the local
  // magnitude, so tiles deep in the tiny-weight region (likelihood_field_prob
  // leaves weights 300 orders of magnitude apart) are still classified exactly.
  const double eps = 4.0 * (double)n * 1.1102230246251565e-16;
  double run = s_part[threadIdx.x];
  int my_dirty = 0;
  for (int t = lo; t < hi; t++)
  {
    const double before = run;
    const double start = run * (1.0 - eps);
    run += tile_sum[(size_t)t * ts];
    const double end = run * (1.0 + eps);
    int cls = kDirty;
    (void)before;
    // Flat: no addition in the tile can move the table.  True when every weight
    // is 0, and also when the largest weight is below a quarter ulp of the
    // smallest value the running sum can have here (c + w rounds back to c,
    // whatever the binade -- the tail of a likelihood_field_prob posterior).
    if (tile_sum[(size_t)t * ts] == 0.0 ||
        (start > 0.0 && tile_max[(size_t)t * ts] < start * 2.7755575615628914e-17))
      cls = kFlat;
    else if (start > 0.0 && isfinite(end))
    {
      const int e0 = ilogb(start), e1 = ilogb(end);
      if (e0 == e1 && e0 > -1020 && e0 < 1020) cls = e0;
    }
    tile_class[t] = cls;
    if (cls == kDirty) my_dirty++;
  }
  // Rows of the dirty tiles in TILE ORDER: every rank of a sharded filter runs this
  // kernel on the same sums and must give the same tile the same row.
  __shared__ int s_cnt[1024];
  __syncthreads();
  s_cnt[threadIdx.x] = my_dirty;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    int acc = 0;
    for (int k = 0; k < (int)blockDim.x; k++)
    {
      const int v = s_cnt[k];
      s_cnt[k] = acc;
      acc += v;
    }
    *dirty_count = acc;
    if (overflow && acc > max_rows) *overflow = 1;
  }
  __syncthreads();
  if (dirty_row)
  {
    int row = s_cnt[threadIdx.x];
    for (int t = lo; t < hi; t++)
      if (tile_class[t] == kDirty)
      {
        dirty_row[t] = row < max_rows ? row : -1;
        row++;
      }
  }
}

// Sharded table: the owner of a dirty tile copies its weights into the tile's row.
__global__ void cs_publish_dirty(const Weights W, int n, const int *__restrict__ tile_class,
                                 const int *__restrict__ dirty_row, int tile0, double *__restrict__ rows)
{
  const int tile = (int)blockIdx.x + tile0;
  if (tile_class[tile] != kDirty) return;
  const int row = dirty_row[tile];
  if (row < 0) return;
  for (int k = threadIdx.x; k < kTileLen; k += blockDim.x)
  {
    const int i = tile * kTileLen + k;
    rows[(size_t)row * kTileLen + k] = i < n ? W[i] : 0.0;
  }
}

// ---- C: composed function of every clean tile ----------------------------------
__global__ void cs_tile_fns(const Weights W, int n, const int *__restrict__ tile_class,
                            PairFn *__restrict__ tile_fn, int tile0)
{
  __shared__ PairFn s_warp[33];
  const int tile = (int)blockIdx.x + tile0;
  const int cls = tile_class[tile];
  if (cls == kDirty || cls == kFlat) return;     // uniform across the block
  const int shift = 52 - cls;
  const int first = tile * kTileLen + threadIdx.x * kPerThread;
  PairFn f = fn_identity();
#pragma unroll
  for (int k = 0; k < kPerThread; k++)
  {
    const int i = first + k;
    if (i < n) f = fn_compose(f, fn_of_weight(W[i], shift));
  }
  PairFn total;
  block_scan_fn(f, &total, s_warp);
  if (threadIdx.x == 0) tile_fn[tile] = total;
}

// ---- the same algorithm inside ONE block, on weights held in shared memory ------------
//
// Used twice: for the dirty 1024-weight tiles of a large table (cs_carries), so
// that only the 32-weight stretch around a binade crossing is walked element by
// element instead of the whole tile, and for the whole table of a node-sized
// filter (cs_small: amcl_node runs 500 .. 10 000 particles, amcl/cfg/AMCL.cfg,
// where a single-thread walk is the longest kernel of the update).  Sub-tiles of
// 32 weights: warp sums classify them, a warp scan of pair functions gives each
// clean sub-tile its function, one thread chains the sub-tiles in integer
// arithmetic (converting only where the binade changes) and walks the few dirty
// ones, then the warps write the entries.
constexpr int kSubPer = 2;                     // weights per lane
constexpr int kSubTile = 32 * kSubPer;         // weights per sub-tile (one warp pass)

struct SubScratch           // shared memory for up to MAXT sub-tiles
{
  double *tsum, *tmax, *cin;
  unsigned long long *cin_int;
  int *cls;
  PairFn *fn;
  double *carry;            // one double: carry out
};

__device__ __forceinline__ int exponent_of(double v)   // ilogb for normal, positive v
{
  return (int)((__double_as_longlong(v) >> 52) & 0x7ff) - 1023;
}
__device__ __forceinline__ double pow2(int e)          // 2^e, |e| < 1022
{
  return __longlong_as_double((long long)(1023 + e) << 52);
}

// Inclusive warp scan of a lane's composed function (lane order = element order).
__device__ __forceinline__ PairFn warp_scan_fn(PairFn inc, int lane)
{
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    const PairFn up = fn_shfl_up(inc, o);
    if (lane >= o) inc = fn_compose(up, inc);
  }
  return inc;
}

// s_w[0..m): the weights; carry0: the table entry in front of them; n_total: the
// length of the whole table (error margin of the rough prefix); out[k + 1] receives
// the entry behind weight k.  All threads of the block must call (at most
// blockDim.x sub-tiles); the carry out is returned to all of them.
__device__ double block_exact_table(const double *s_w, int m, double carry0, int n_total,
                                    double *__restrict__ out, const SubScratch S)
{
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int tiles = (m + kSubTile - 1) / kSubTile;
  // A: sum and maximum of every sub-tile
  for (int t = warp; t < tiles; t += nwarps)
  {
    double v = 0.0, mx = 0.0;
#pragma unroll
    for (int q = 0; q < kSubPer; q++)
    {
      const int i = t * kSubTile + lane * kSubPer + q;
      const double w = i < m ? s_w[i] : 0.0;
      v += w;
      mx = fmax(mx, w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
      v += __shfl_xor_sync(kFullMask, v, o);
      mx = fmax(mx, __shfl_xor_sync(kFullMask, mx, o));
    }
    if (lane == 0) { S.tsum[t] = v; S.tmax[t] = mx; }
  }
  __syncthreads();
  // B: rough prefix (block scan, one sub-tile per thread) and classification with
  // the margins of cs_classify
  {
    const double v = tid < tiles ? S.tsum[tid] : 0.0;
    double inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
      const double up = __shfl_up_sync(kFullMask, inc, o);
      if (lane >= o) inc += up;
    }
    if (lane == 31 && warp * 32 < tiles) S.cin[warp] = inc;   // warp totals (cin is free until step D)
    __syncthreads();
    double base = carry0;
    if (tid < tiles)
      for (int w = 0; w < warp; w++) base += S.cin[w];
    __syncthreads();
    if (tid < tiles)
    {
      const double eps = 4.0 * (double)n_total * 1.1102230246251565e-16;
      const double start = (base + (inc - v)) * (1.0 - eps);
      const double end = (base + inc) * (1.0 + eps);
      int cls = kDirty;
      if (v == 0.0 || (start > 0.0 && S.tmax[tid] < start * 2.7755575615628914e-17))
        cls = kFlat;
      else if (start > 0.0 && end < 1.0e300)
      {
        const int e0 = exponent_of(start), e1 = exponent_of(end);
        if (e0 == e1 && e0 > -1020 && e0 < 1020) cls = e0;
      }
      S.cls[tid] = cls;
    }
  }
  __syncthreads();
  // C: composed function of every clean sub-tile
  for (int t = warp; t < tiles; t += nwarps)
  {
    const int cls = S.cls[t];
    if (cls == kDirty || cls == kFlat) continue;            // uniform across the warp
    PairFn f = fn_identity();
#pragma unroll
    for (int q = 0; q < kSubPer; q++)
    {
      const int i = t * kSubTile + lane * kSubPer + q;
      if (i < m) f = fn_compose(f, fn_of_weight(s_w[i], 52 - cls));
    }
    f = warp_scan_fn(f, lane);
    if (lane == 31) S.fn[t] = f;
  }
  __syncthreads();
  // D: the exact carry into every sub-tile; dirty ones are walked here
  if (tid == 0)
  {
    double carry = carry0;
    unsigned long long C = 0ull;
    int e = 0;
    bool as_int = false;
    int cls_next = tiles > 0 ? S.cls[0] : kFlat;
    PairFn fn_next = tiles > 0 ? S.fn[0] : fn_identity();
    for (int t = 0; t < tiles; t++)
    {
      const int cls = cls_next;
      const PairFn fn = fn_next;
      if (t + 1 < tiles)                     // fetched a round early: this loop is all latency
      {
        cls_next = S.cls[t + 1];
        fn_next = S.fn[t + 1];
      }
      if (cls != kDirty && cls != kFlat)
      {
        if (!as_int || e != cls)
        {
          if (as_int) carry = (double)C * pow2(e - 52);
          C = (unsigned long long)(carry * pow2(52 - cls));    // exact: carry is a multiple of 2^(cls-52)
          e = cls;
          as_int = true;
        }
        S.cin_int[t] = C;
        C = fn_apply(fn, C);
        continue;
      }
      if (as_int)
      {
        carry = (double)C * pow2(e - 52);
        as_int = false;
      }
      S.cin[t] = carry;
      if (cls == kDirty)
      {
        const int a = t * kSubTile, cnt = min(kSubTile, m - a);
        for (int k = 0; k < cnt; k++)
        {
          carry = __dadd_rn(carry, s_w[a + k]);
          out[a + k + 1] = carry;
        }
      }
    }
    if (as_int) carry = (double)C * pow2(e - 52);
    S.carry[0] = carry;
  }
  __syncthreads();
  // E: the entries of the clean and flat sub-tiles
  for (int t = warp; t < tiles; t += nwarps)
  {
    const int cls = S.cls[t];
    if (cls == kDirty) continue;
    const int i0 = t * kSubTile + lane * kSubPer;
    if (cls == kFlat)
    {
#pragma unroll
      for (int q = 0; q < kSubPer; q++)
        if (i0 + q < m) out[i0 + q + 1] = S.cin[t];
      continue;
    }
    const int shift = 52 - cls;
    PairFn mine[kSubPer];
    PairFn f = fn_identity();
#pragma unroll
    for (int q = 0; q < kSubPer; q++)
    {
      mine[q] = i0 + q < m ? fn_of_weight(s_w[i0 + q], shift) : fn_identity();
      f = fn_compose(f, mine[q]);
    }
    const PairFn inc = warp_scan_fn(f, lane);
    PairFn ex = fn_shfl_up(inc, 1);
    if (lane == 0) ex = fn_identity();
    unsigned long long C = fn_apply(ex, S.cin_int[t]);
    const double unit = pow2(-shift);
#pragma unroll
    for (int q = 0; q < kSubPer; q++)
    {
      C = fn_apply(mine[q], C);
      if (i0 + q < m) out[i0 + q + 1] = (double)C * unit;
    }
  }
  const double result = S.carry[0];
  __syncthreads();
  return result;
}

// ---- D: exact carry into every tile; dirty tiles are walked here ---------------
// dirty_rows / dirty_row: where the weights of the dirty tiles are read from in
// the sharded table (this rank owns only some of them); null: from W itself.
__global__ void cs_carries(const Weights W, int n, int tiles,
                           const int *__restrict__ tile_class, const PairFn *__restrict__ tile_fn,
                           double *__restrict__ carry_in, double *__restrict__ cum,
                           const double *__restrict__ dirty_rows, const int *__restrict__ dirty_row)
{
  __shared__ PairFn s_warp[33];
  __shared__ double s_w[kTileLen];
  __shared__ double s_carry;
  __shared__ int s_min[32];
  __shared__ int s_first_dirty, s_run_class;
  constexpr int kSub = kTileLen / kSubTile;
  __shared__ double s_sub_d[3 * kSub + 1];
  __shared__ unsigned long long s_sub_c[kSub];
  __shared__ int s_sub_cls[kSub];
  __shared__ PairFn s_sub_fn[kSub];
  SubScratch sub;
  sub.tsum = s_sub_d; sub.tmax = s_sub_d + kSub; sub.cin = s_sub_d + 2 * kSub; sub.carry = s_sub_d + 3 * kSub;
  sub.cin_int = s_sub_c; sub.cls = s_sub_cls; sub.fn = s_sub_fn;
  const int tid = threadIdx.x;
  if (tid == 0)
  {
    s_carry = 0.0;
    cum[0] = 0.0;
  }
  __syncthreads();
  for (int base = 0; base < tiles; base += blockDim.x)
  {
    const int t = base + tid;
    const bool live = t < tiles;
    const int cls = live ? tile_class[t] : 0;
    const bool dirty = live && cls == kDirty;
    int cursor = 0;
    for (;;)
    {
      // first dirty tile at or after the cursor (block-wide minimum)
      int cand = (dirty && tid >= cursor) ? tid : (int)blockDim.x;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cand = min(cand, __shfl_xor_sync(kFullMask, cand, o));
      if ((tid & 31) == 0) s_min[tid >> 5] = cand;
      __syncthreads();
      if (tid < 32)
      {
        int v = tid < (int)(blockDim.x >> 5) ? s_min[tid] : (int)blockDim.x;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(kFullMask, v, o));
        if (tid == 0) s_first_dirty = v;
      }
      __syncthreads();
      const int d = s_first_dirty;
      const int run_end = min(d, tiles - base);          // non-dirty tiles [cursor, run_end)
      if (run_end > cursor)
      {
        const bool in_run = tid >= cursor && tid < run_end;
        // The binade of the run is that of its first tile that has weight in it:
        // flat tiles fit any binade, and a change of binade needs a crossing,
        // which makes a tile dirty and ends the run.
        int pick = (in_run && cls != kFlat) ? tid : (int)blockDim.x;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pick = min(pick, __shfl_xor_sync(kFullMask, pick, o));
        if ((tid & 31) == 0) s_min[tid >> 5] = pick;
        __syncthreads();
        if (tid < 32)
        {
          int v = tid < (int)(blockDim.x >> 5) ? s_min[tid] : (int)blockDim.x;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(kFullMask, v, o));
          if (tid == 0) s_first_dirty = v;                // reused as "first non-flat"
        }
        __syncthreads();
        const int first_solid = s_first_dirty;
        if (tid == first_solid) s_run_class = cls;
        __syncthreads();
        if (first_solid >= (int)blockDim.x)
        {
          // nothing but flat tiles: the carry passes through untouched
          if (in_run) carry_in[t] = s_carry;
          __syncthreads();
        }
        else
        {
          const int e = s_run_class;
          PairFn f = (in_run && cls != kFlat) ? tile_fn[t] : fn_identity();
          PairFn total;
          const PairFn ex = block_scan_fn(f, &total, s_warp);
          const unsigned long long C0 = (unsigned long long)scalbn(s_carry, 52 - e);
          if (in_run) carry_in[t] = scalbn((double)fn_apply(ex, C0), e - 52);
          __syncthreads();                                  // everyone has read s_carry
          if (tid == 0) s_carry = scalbn((double)fn_apply(total, C0), e - 52);
          __syncthreads();
        }
      }
      if (d >= (int)blockDim.x || base + d >= tiles) break;
      // walk the dirty tile with real fp64 additions
      const int tt = base + d;
      const int a = tt * kTileLen;
      const int m = min(kTileLen, n - a);
      if (dirty_rows)
      {
        const int row = dirty_row[tt];
        for (int k = tid; k < m; k += blockDim.x) s_w[k] = row >= 0 ? dirty_rows[(size_t)row * kTileLen + k] : 0.0;
      }
      else
        for (int k = tid; k < m; k += blockDim.x) s_w[k] = W[a + k];
      __syncthreads();
      {
        // only the 32-weight stretches that may hold a crossing are walked
        if (tid == 0) cum[a] = s_carry;      // (already there unless the tile opens a rank's slice)
        const double carry_out = block_exact_table(s_w, m, s_carry, n, cum + a, sub);
        if (tid == 0) s_carry = carry_out;
      }
      cursor = d + 1;
      __syncthreads();
    }
    __syncthreads();
  }
}

// ---- E: per-element table of the clean tiles ----------------------------------------
__global__ void cs_apply(const Weights W, int n, const int *__restrict__ tile_class,
                         const double *__restrict__ carry_in, double *__restrict__ cum, int tile0)
{
  __shared__ PairFn s_warp[33];
  const int tile = (int)blockIdx.x + tile0;
  const int cls = tile_class[tile];
  if (cls == kDirty) return;
  // the entry in front of the tile: written by the previous tile as well, except
  // where that tile belongs to another rank
  if (threadIdx.x == 0) cum[(size_t)tile * kTileLen] = carry_in[tile];
  if (cls == kFlat)
  {
    const double v = carry_in[tile];
    for (int k = threadIdx.x; k < kTileLen; k += blockDim.x)
    {
      const int i = tile * kTileLen + k;
      if (i < n) cum[i + 1] = v;
    }
    return;
  }
  const int shift = 52 - cls;
  const int first = tile * kTileLen + threadIdx.x * kPerThread;
  double w[kPerThread];
  PairFn f = fn_identity();
#pragma unroll
  for (int k = 0; k < kPerThread; k++)
  {
    const int i = first + k;
    w[k] = i < n ? W[i] : 0.0;
    f = fn_compose(f, fn_of_weight(w[k], shift));
  }
  PairFn total;
  const PairFn ex = block_scan_fn(f, &total, s_warp);
  unsigned long long C = (unsigned long long)scalbn(carry_in[tile], shift);
  C = fn_apply(ex, C);
#pragma unroll
  for (int k = 0; k < kPerThread; k++)
  {
    const int i = first + k;
    C = fn_apply(fn_of_weight(w[k], shift), C);
    if (i < n) cum[i + 1] = scalbn((double)C, -shift);
  }
}

// ---- F: the defining recurrence, checked everywhere -------------------------------------
__global__ void cs_verify(const Weights W, int n, const double *__restrict__ cum,
                          int *bad, int first, int count)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  const int i = first + t;
  if (i >= n) return;
  const double want = __dadd_rn(cum[i], W[i]);
  if (!(cum[i + 1] == want) || (i == 0 && cum[0] != 0.0)) *bad = 1;
}

// ---- node-sized filters: the whole table in one launch, nothing returns to the host ---
constexpr int kSmallMaxTiles = 16384 / kSubTile;          // n <= 16384
__device__ unsigned long long g_small_fallbacks_dev = 0;

__global__ void __launch_bounds__(1024)
cs_small(const Weights W, int n, double *__restrict__ cum)
{
  extern __shared__ double s_w[];                          // n weights
  __shared__ double s_d[3 * kSmallMaxTiles + 1];
  __shared__ unsigned long long s_c[kSmallMaxTiles];
  __shared__ int s_cls[kSmallMaxTiles];
  __shared__ PairFn s_fn[kSmallMaxTiles];
  __shared__ int s_bad;
  SubScratch sub;
  sub.tsum = s_d; sub.tmax = s_d + kSmallMaxTiles; sub.cin = s_d + 2 * kSmallMaxTiles;
  sub.carry = s_d + 3 * kSmallMaxTiles;
  sub.cin_int = s_c; sub.cls = s_cls; sub.fn = s_fn;
  const int tid = threadIdx.x;
  if (tid == 0)
  {
    s_bad = 0;
    cum[0] = 0.0;
  }
  for (int i = tid; i < n; i += blockDim.x) s_w[i] = W[i];
  __syncthreads();
  block_exact_table(s_w, n, 0.0, n, cum, sub);
  // the defining recurrence, everywhere (this block's own global writes)
  for (int i = tid; i < n; i += blockDim.x)
    if (!(cum[i + 1] == __dadd_rn(cum[i], s_w[i]))) s_bad = 1;
  __syncthreads();
  if (s_bad && tid == 0)
  {
    atomicAdd(&g_small_fallbacks_dev, 1ull);
    double run = 0.0;
    for (int i = 0; i < n; i++)
    {
      run = __dadd_rn(run, s_w[i]);
      cum[i + 1] = run;
    }
  }
}

unsigned long long g_fallbacks = 0;
int g_last_dirty = 0, g_last_tiles = 0;

}  // namespace

unsigned long long exact_cumsum_fallback_count()
{
  unsigned long long small = 0;
  if (globals().device >= 0)
    AMCL_CUDA(cudaMemcpyFromSymbol(&small, g_small_fallbacks_dev, sizeof(small)));   // waits for the device
  return g_fallbacks + small;
}

// n <= 16384: one launch, no trip to the host.
void exact_cumsum_small(Weights W, int n, double *d_cum)
{
  static bool attr_set = false;
  if (!attr_set)
  {
    AMCL_CUDA(cudaFuncSetAttribute(cs_small, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(sizeof(double) * kSubTile * kSmallMaxTiles)));
    attr_set = true;
  }
  cs_small<<<1, 1024, sizeof(double) * (size_t)(n > 0 ? n : 1), globals().stream>>>(W, n, d_cum);
  AMCL_LAUNCH_CHECK();
}
void exact_cumsum_last_tiles(int out[2]) { out[0] = g_last_dirty; out[1] = g_last_tiles; }

static void ensure_tile_scratch(PfCtx *c, int tiles)
{
  if (tiles <= c->tile_cap) return;
  if (c->d_tile_sum) { cudaFree(c->d_tile_sum); cudaFree(c->d_tile_carry); cudaFree(c->d_tile_fn); cudaFree(c->d_tile_class); }
  c->tile_cap = tiles + 64;
  AMCL_CUDA(cudaMalloc(&c->d_tile_sum, sizeof(double) * 2 * (size_t)c->tile_cap));   // room for (sum, max) pairs
  AMCL_CUDA(cudaMalloc(&c->d_tile_carry, sizeof(double) * (size_t)c->tile_cap));
  AMCL_CUDA(cudaMalloc(&c->d_tile_fn, sizeof(PairFn) * (size_t)c->tile_cap));
  AMCL_CUDA(cudaMalloc(&c->d_tile_class, sizeof(int) * (size_t)c->tile_cap));
}

// ---- the table of a sharded filter ---------------------------------------------------
//
// Every rank owns a run of whole tiles (Shard::chunk is a multiple of the tile
// length) and only its own weights.  The O(n) passes -- tile sums, tile
// functions, the per-element table, the recurrence check -- run on the own tiles
// only; what the ranks exchange is one record per TILE (sum + maximum for the
// classification, the composed pair function of the clean tiles: 32 B per 1024
// weights) and the weights of the few dirty tiles; the single-block pass that
// turns those into the exact carry of every tile is repeated by everybody
// (16 Ki tiles at 16 Mi particles).  The finished slices are then all-gathered:
// 8 B per particle, like the weights before, but without the 7/8 of the table
// work that every rank used to redo.  Same bits as the replicated form: it IS the
// same algorithm, tile by tile.
constexpr int kMaxDirtyRows = 96;

bool exact_cumsum_sharded(PfCtx *c, const Particle *set, int n, double *d_cum)
{
  const Shard &sh = c->shard;
  if (!sh.on() || sh.align != kTileLen || !globals().exact_cumsum_parallel || n < (1 << 16)) return false;
  cudaStream_t st = globals().stream;
  const int chunk = sh.chunk(n);                    // multiple of kTileLen
  const int chunk_tiles = chunk / kTileLen;
  const int tiles = (n + kTileLen - 1) / kTileLen;
  const int tile0 = sh.rank * chunk_tiles;
  int own = tiles - tile0;
  if (own > chunk_tiles) own = chunk_tiles;
  if (own < 0) own = 0;
  ensure_tile_scratch(c, chunk_tiles * sh.world);
  if (!c->d_dirty_w)
    AMCL_CUDA(cudaMalloc(&c->d_dirty_w, sizeof(double) * (size_t)kMaxDirtyRows * kTileLen));
  if (c->dirty_row_cap < c->tile_cap)
  {
    if (c->d_dirty_row) cudaFree(c->d_dirty_row);
    c->dirty_row_cap = c->tile_cap;
    AMCL_CUDA(cudaMalloc(&c->d_dirty_row, sizeof(int) * (size_t)c->dirty_row_cap));
  }
  PairFn *tile_fn = reinterpret_cast<PairFn *>(c->d_tile_fn);
  if (!c->d_wpack) AMCL_CUDA(cudaMalloc(&c->d_wpack, sizeof(double) * (size_t)c->alloc_samples));
  Weights W = weights_of(set);
  Trace tr;
  tr.mark("start");
  // tiles beyond this rank's (and beyond the table's end) must not carry stale sums
  // (sum, maximum) of a tile side by side: one all-gather moves both
  AMCL_CUDA(cudaMemsetAsync(c->d_tile_sum + 2 * (size_t)tile0, 0, sizeof(double) * 2 * (size_t)chunk_tiles, st));
  if (own > 0)
  {
    cs_tile_sums<<<own, kTileThreads, 0, st>>>(W, n, c->d_tile_sum, c->d_tile_sum + 1, c->d_wpack, tile0, 2);
    AMCL_LAUNCH_CHECK();
  }
  W.p = c->d_wpack;
  W.stride = 1;
  sh.allgather(c->d_tile_sum, (unsigned long long)chunk_tiles * 2 * sizeof(double));
  tr.mark("sums+ag");
  const int bad_slot = 8, dirty_slot = 9, overflow_slot = 10;
  AMCL_CUDA(cudaMemsetAsync(c->d_ints + bad_slot, 0, 3 * sizeof(int), st));
  cs_classify<<<1, 1024, 0, st>>>(c->d_tile_sum, c->d_tile_sum + 1, tiles, n, c->d_tile_class,
                                 c->d_ints + dirty_slot, c->d_dirty_row, kMaxDirtyRows,
                                 c->d_ints + overflow_slot, 2);
  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaMemsetAsync(c->d_dirty_w, 0, sizeof(double) * (size_t)kMaxDirtyRows * kTileLen, st));
  if (own > 0)
  {
    cs_tile_fns<<<own, kTileThreads, 0, st>>>(W, n, c->d_tile_class, tile_fn, tile0);
    AMCL_LAUNCH_CHECK();
    cs_publish_dirty<<<own, kTileThreads, 0, st>>>(W, n, c->d_tile_class, c->d_dirty_row, tile0, c->d_dirty_w);
    AMCL_LAUNCH_CHECK();
  }
  sh.allgather(tile_fn, (unsigned long long)chunk_tiles * sizeof(PairFn));
  // rows of tiles this rank does not own are zero here: the sum is a gather
  sh.allreduce(c->d_dirty_w, kMaxDirtyRows * kTileLen, SHARD_F64, SHARD_SUM);
  tr.mark("classify+fns+ag+ar");
  cs_carries<<<1, 1024, 0, st>>>(W, n, tiles, c->d_tile_class, tile_fn, c->d_tile_carry, d_cum,
                                c->d_dirty_w, c->d_dirty_row);
  AMCL_LAUNCH_CHECK();
  if (own > 0)
  {
    cs_apply<<<own, kTileThreads, 0, st>>>(W, n, c->d_tile_class, c->d_tile_carry, d_cum, tile0);
    AMCL_LAUNCH_CHECK();
    const int first = tile0 * kTileLen;
    const int count = (n - first < chunk ? n - first : chunk);
    cs_verify<<<(count + 255) / 256, 256, 0, st>>>(W, n, d_cum, c->d_ints + bad_slot, first, count);
    AMCL_LAUNCH_CHECK();
  }
  tr.mark("carries+apply");
  // entry i + 1 of the table belongs to the owner of weight i
  sh.allgather(d_cum + 1, (unsigned long long)chunk * sizeof(double));
  tr.mark("ag table");
  // one verdict for everybody: any rank's failed check, or too many dirty tiles
  sh.allreduce(c->d_ints + bad_slot, 3, SHARD_I32, SHARD_MAX);
  AMCL_CUDA(cudaMemcpyAsync(c->h_ints + bad_slot, c->d_ints + bad_slot, 3 * sizeof(int),
                            cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  tr.mark("verdict");
  tr.report("exact_cumsum_sharded");
  g_last_dirty = c->h_ints[dirty_slot];
  g_last_tiles = tiles;
  if (c->h_ints[bad_slot] || c->h_ints[overflow_slot])
  {
    if (c->h_ints[bad_slot])
    {
      g_fallbacks++;
      fprintf(stderr, "[amcl_b200] exact_cumsum (sharded): recurrence check failed; using the replicated table\n");
    }
    return false;
  }
  return true;
}

void exact_cumsum_parallel(PfCtx *c, Weights W, int n, double *d_cum)
{
  cudaStream_t st = globals().stream;
  const int tiles = (n + kTileLen - 1) / kTileLen;
  ensure_tile_scratch(c, tiles);
  PairFn *tile_fn = reinterpret_cast<PairFn *>(c->d_tile_fn);
  double *packed = nullptr;
  if (W.stride != 1)
  {
    if (!c->d_wpack) AMCL_CUDA(cudaMalloc(&c->d_wpack, sizeof(double) * (size_t)c->alloc_samples));
    packed = c->d_wpack;
  }
  cs_tile_sums<<<tiles, kTileThreads, 0, st>>>(W, n, c->d_tile_sum, c->d_tile_carry, packed, 0, 1);
  AMCL_LAUNCH_CHECK();
  if (packed)
  {
    W.p = packed;
    W.stride = 1;
  }
  const int bad_slot = 8, dirty_slot = 9;
  AMCL_CUDA(cudaMemsetAsync(c->d_ints + bad_slot, 0, 2 * sizeof(int), st));
  cs_classify<<<1, 1024, 0, st>>>(c->d_tile_sum, c->d_tile_carry, tiles, n, c->d_tile_class,
                                 c->d_ints + dirty_slot, nullptr, 0, nullptr, 1);
  AMCL_LAUNCH_CHECK();
  cs_tile_fns<<<tiles, kTileThreads, 0, st>>>(W, n, c->d_tile_class, tile_fn, 0);
  AMCL_LAUNCH_CHECK();
  cs_carries<<<1, 1024, 0, st>>>(W, n, tiles, c->d_tile_class, tile_fn, c->d_tile_carry, d_cum, nullptr,
                                nullptr);
  AMCL_LAUNCH_CHECK();
  cs_apply<<<tiles, kTileThreads, 0, st>>>(W, n, c->d_tile_class, c->d_tile_carry, d_cum, 0);
  AMCL_LAUNCH_CHECK();
  cs_verify<<<(n + 255) / 256, 256, 0, st>>>(W, n, d_cum, c->d_ints + bad_slot, 0, n);
  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaMemcpyAsync(c->h_ints + bad_slot, c->d_ints + bad_slot, 2 * sizeof(int),
                            cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  g_last_dirty = c->h_ints[dirty_slot];
  g_last_tiles = tiles;
  if (c->h_ints[bad_slot])
  {
    g_fallbacks++;
    fprintf(stderr, "[amcl_b200] exact_cumsum: recurrence check failed (non-finite or negative "
                    "weights?); using the single-thread walk\n");
    launch_cumsum_sequential(W, n, d_cum);
  }
}

}  // namespace amclb200
