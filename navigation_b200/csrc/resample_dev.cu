// resample_dev.cu -- KLD-adaptive multinomial resampling, pose histogram,
// cluster labelling and cluster / set statistics on the device.  Stands in for
//   pf_update_resample, copy_set, pf_resample_limit   pf.c:321-555
//   pf_kdtree_insert / _cluster / _get_cluster         pf_kdtree.c:112-439
//   pf_cluster_stats, pf_update_converged, cep stats   pf.c:220-251, 559-740
// (paths under /root/reference/amcl/src/amcl/pf/).
//
// How the sequential reference loop becomes data-parallel without changing its
// results:
//  * The live sampler is MULTINOMIAL (pf.c:458-466; the low-variance code is
//    commented out): sample m takes the source i with c[i] <= r_m < c[i+1].
//    That is "how many of c[1..N] are <= r_m" -- a binary search per draw.
//  * c[] must be the SEQUENTIAL fp64 running sum (pf.c:399-402) bit for bit,
//    or draws that land within rounding of a boundary pick another index:
//    exact_cumsum() below reproduces it exactly (see that function).
//  * leaf_count after m insertions == number of distinct bin keys among
//    samples 0..m: a hash insert that keeps the minimum sample index per key
//    marks "first occurrences"; an inclusive scan of those marks is k_m.
//  * The loop stops at the first m with (m+1) > limit(k_m) (pf.c:483): a min-
//    reduction.  limit() is evaluated with IEEE fp64 ops and no contraction,
//    so it returns the integers the host formula returns.
//  * Clusters are the 26-connected components of occupied bins
//    (pf_kdtree.c:407-438): lock-free union-find over the hash slots.  Cluster
//    NUMBERS here are canonical -- clusters are numbered in the order their
//    first sample appears -- because the reference's numbering is an artefact
//    of kd-tree node allocation order (C4).  Statistics are per-cluster sums.
#include <math.h>
#include <float.h>
#include <string.h>

#include <vector>

#include "context.hpp"
#include "philox.cuh"

namespace amclb200
{

// ------------------------------------------------------------------ bin keys --

// pf_kdtree.c:116-118: key = floor(pose / size) per axis, IEEE division.
// Packed as 21 | 21 | 22 bits with a bias so that neighbours differ by fixed
// increments.  Poses beyond +-2^20 bins (524 km) are clamped.
constexpr int kBiasXY = 1 << 20;
constexpr int kBiasTh = 1 << 21;

__host__ __device__ inline unsigned long long pack_key(int kx, int ky, int kt)
{
  kx = kx < -kBiasXY ? -kBiasXY : (kx >= kBiasXY ? kBiasXY - 1 : kx);
  ky = ky < -kBiasXY ? -kBiasXY : (ky >= kBiasXY ? kBiasXY - 1 : ky);
  kt = kt < -kBiasTh ? -kBiasTh : (kt >= kBiasTh ? kBiasTh - 1 : kt);
  return ((unsigned long long)(kx + kBiasXY) << 43) | ((unsigned long long)(ky + kBiasXY) << 22) |
         (unsigned long long)(kt + kBiasTh);
}

__device__ __forceinline__ int floor_div_to_int(double v, double size)
{
  const double q = floor(__ddiv_rn(v, size));
  // (int) of an out-of-range double is what x86's cvttsd2si makes of it
  if (!(q > -2147483649.0 && q < 2147483648.0)) return (int)0x80000000;
  return (int)q;
}

__device__ __forceinline__ unsigned long long pose_key(const Particle &s, double sx, double sy,
                                                       double st)
{
  return pack_key(floor_div_to_int(s.x, sx), floor_div_to_int(s.y, sy),
                  floor_div_to_int(s.th, st));
}

__host__ __device__ inline unsigned int hash_key(unsigned long long k)
{
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (unsigned int)k;
}

// --------------------------------------------------------------- u32 scan ----

constexpr int kScanThreads = 256;
constexpr int kScanItems = 16;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ unsigned int block_exclusive_scan_u32(unsigned int v, unsigned int *total,
                                                                 unsigned int *s_warp)
{
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  unsigned int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    unsigned int t = __shfl_up_sync(kFullMask, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  if (wid == 0)
  {
    unsigned int w = lane < (int)(blockDim.x >> 5) ? s_warp[lane] : 0u;
    unsigned int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
      unsigned int t = __shfl_up_sync(kFullMask, winc, o);
      if (lane >= o) winc += t;
    }
    s_warp[lane] = winc - w;          // exclusive warp offsets
    if (lane == 31) s_warp[32] = winc;  // block total
  }
  __syncthreads();
  *total = s_warp[32];
  return inc - v + s_warp[wid];
}

__global__ void scan_tile_sums(const unsigned int *__restrict__ in, int n,
                               unsigned int *__restrict__ tile_sums)
{
  __shared__ unsigned int s_warp[33];
  const int base = blockIdx.x * kScanTile;
  unsigned int v = 0;
  for (int k = 0; k < kScanItems; k++)
  {
    int i = base + k * kScanThreads + threadIdx.x;
    if (i < n) v += in[i];
  }
  unsigned int total;
  block_exclusive_scan_u32(v, &total, s_warp);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// single block: exclusive scan of the tile sums in place; total -> *grand
__global__ void scan_tile_offsets(unsigned int *tile_sums, int tiles, unsigned int *grand)
{
  __shared__ unsigned int s_warp[33];
  __shared__ unsigned int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < tiles; base += blockDim.x)
  {
    int i = base + threadIdx.x;
    unsigned int v = i < tiles ? tile_sums[i] : 0u;
    unsigned int total;
    unsigned int ex = block_exclusive_scan_u32(v, &total, s_warp);
    if (i < tiles) tile_sums[i] = ex + s_carry;
    __syncthreads();
    if (threadIdx.x == 0) s_carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *grand = s_carry;
}

__global__ void scan_apply(const unsigned int *__restrict__ in, int n,
                           const unsigned int *__restrict__ tile_offsets,
                           unsigned int *__restrict__ out_inclusive)
{
  __shared__ unsigned int s_warp[33];
  const int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
  unsigned int item[kScanItems];
  unsigned int v = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; k++)
  {
    int i = base + k;
    item[k] = i < n ? in[i] : 0u;
    v += item[k];
  }
  unsigned int total;
  unsigned int run = block_exclusive_scan_u32(v, &total, s_warp) + tile_offsets[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kScanItems; k++)
  {
    run += item[k];
    int i = base + k;
    if (i < n) out_inclusive[i] = run;
  }
}

// inclusive scan of n flags; returns nothing, grand total lands in d_ints[slot]
static void scan_u32(PfCtx *c, const unsigned int *in, unsigned int *out, int n, int total_slot)
{
  cudaStream_t st = globals().stream;
  const int tiles = (n + kScanTile - 1) / kScanTile;
  unsigned int *tile_sums = c->d_scan_tiles;
  if (tiles > c->scan_tiles_cap) AMCL_FATAL("scan tile scratch too small (%d > %d)", tiles, c->scan_tiles_cap);
  scan_tile_sums<<<tiles, kScanThreads, 0, st>>>(in, n, tile_sums);
  AMCL_LAUNCH_CHECK();
  scan_tile_offsets<<<1, 1024, 0, st>>>(tile_sums, tiles,
                                        reinterpret_cast<unsigned int *>(c->d_ints + total_slot));
  AMCL_LAUNCH_CHECK();
  scan_apply<<<tiles, kScanThreads, 0, st>>>(in, n, tile_sums, out);
  AMCL_LAUNCH_CHECK();
}

// ------------------------------------------------- exact cumulative table ----
//
// pf.c:399-402 computes c[i+1] = fl(c[i] + w[i]) one element after another.
// Version 1 of the device path keeps that order literally: a single thread
// performs the dependent additions while the block streams weights in and
// sums out through shared memory.  (The tile-parallel exact form replaces it
// for large N; see exact_cumsum_parallel.cu when present.)
constexpr int kSeqChunk = 4096;

__global__ void cumsum_sequential(const Weights W, int n, double *__restrict__ c)
{
  __shared__ double s_w[kSeqChunk];
  __shared__ double s_carry;
  if (threadIdx.x == 0)
  {
    s_carry = 0.0;
    c[0] = 0.0;
  }
  __syncthreads();
  for (int base = 0; base < n; base += kSeqChunk)
  {
    const int m = min(kSeqChunk, n - base);
    for (int k = threadIdx.x; k < m; k += blockDim.x) s_w[k] = W[base + k];
    __syncthreads();
    if (threadIdx.x == 0)
    {
      double run = s_carry;
      for (int k = 0; k < m; k++)
      {
        run = __dadd_rn(run, s_w[k]);
        s_w[k] = run;
      }
      s_carry = run;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < m; k += blockDim.x) c[base + k + 1] = s_w[k];
    __syncthreads();
  }
}

void exact_cumsum_parallel(PfCtx *c, Weights W, int n, double *d_cum);  // cumsum_dev.cu

void launch_cumsum_sequential(Weights W, int n, double *d_cum)
{
  cumsum_sequential<<<1, 256, 0, globals().stream>>>(W, n, d_cum);
  AMCL_LAUNCH_CHECK();
}

void exact_cumsum_small(Weights W, int n, double *d_cum);   // cumsum_dev.cu

void exact_cumsum_w(PfCtx *c, Weights W, int n, double *d_cum)
{
  if (globals().exact_cumsum_parallel && n > 16384)
    exact_cumsum_parallel(c, W, n, d_cum);
  else if (globals().exact_cumsum_parallel && n >= 64)
    exact_cumsum_small(W, n, d_cum);
  else
    launch_cumsum_sequential(W, n, d_cum);
}

void exact_cumsum(PfCtx *c, const Particle *set, int n, double *d_cum)
{
  exact_cumsum_w(c, weights_of(set), n, d_cum);
}

// ------------------------------------------------------- draw + search + copy --

__global__ void philox_resample_draws(double *draws, int count, uint64_t key)
{
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= count) return;
  PhiloxStream rng(key, (uint64_t)m);
  draws[2 * (size_t)m] = rng.uniform();      // compared with w_diff (pf.c:430)
  draws[2 * (size_t)m + 1] = rng.uniform();  // the multinomial draw (pf.c:460)
}

constexpr int kCoarse = 256;

// coarse[j] = cum[(j + 1) * kCoarse]
__global__ void sample_table(const double *__restrict__ cum, int n_coarse, double *__restrict__ coarse)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n_coarse) coarse[j] = cum[(size_t)(j + 1) * kCoarse];
}

// #{ j in 1..n : cum[j] <= r }, searched inside [lo, hi] (the caller knows that
// cum[j] <= r for j <= lo and cum[j] > r for j > hi).
__device__ __forceinline__ int table_rank_between(const double *__restrict__ cum, double r, int lo, int hi)
{
  while (lo < hi)
  {
    const int mid = lo + ((hi - lo) >> 1);
    if (cum[mid + 1] <= r) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// The same count over the whole table, through the coarse level when present:
// every kCoarse-th entry lives in a small array that stays in L1/L2, so only
// the last ~8 probes touch the big table.
__device__ __forceinline__ int table_rank(const double *__restrict__ coarse, int n_coarse,
                                          const double *__restrict__ cum, int n, double r)
{
  int lo = 0, hi = n;
  if (coarse)
  {
    int clo = 0, chi = n_coarse;   // #{ j in 1..n_coarse : cum[j*kCoarse] <= r }
    while (clo < chi)
    {
      const int mid = clo + ((chi - clo) >> 1);
      if (__ldg(coarse + mid) <= r) clo = mid + 1;
      else chi = mid;
    }
    lo = clo * kCoarse;
    hi = min(n, lo + kCoarse - 1);
    // entries lo+1 .. lo+kCoarse-1 are undecided; cum[lo+kCoarse] > r is known
  }
  return table_rank_between(cum, r, lo, hi);
}

// Bucket index of the table: index[b] = #{ j : cum[j] <= b / buckets } for
// b = 0..buckets (buckets is a power of two, so b / buckets and r * buckets are
// exact).  A draw r in [b/buckets, (b+1)/buckets) then has its answer inside
// [index[b], index[b+1]] -- a handful of neighbouring entries instead of a
// 24-level search through 128 MB.  Built with one full search per bucket,
// 1/16 of the searches the draws themselves would need.
__global__ void build_bucket_index(const double *__restrict__ coarse, int n_coarse,
                                   const double *__restrict__ cum, int n, int buckets,
                                   int *__restrict__ index)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > buckets) return;
  index[b] = table_rank(coarse, n_coarse, cum, n, (double)b / (double)buckets);
}

struct PickArgs
{
  const int *bucket_index;  // see build_bucket_index, or null
  int buckets;
  const double *coarse;     // every kCoarse-th table entry, or null
  int n_coarse;
  const Particle *src;
  int n_src;
  const double *cum;        // n_src + 1 entries
  const double *draws;      // (u_test, r) per candidate; null: drawn here from the Philox stream
  unsigned long long philox_key;
  const Particle *injected; // random poses for candidates flagged on the host, or null
  const int *inject_slot;   // candidate -> 1 + index into injected, 0 otherwise; or null
  Particle *dst;
  int *src_index;
  unsigned long long *keys;
  const Particle *const *peers;  // sharded + peer memory: rank r's array of the source set
  int peer_chunk;           // particles per rank slice (owner of source i = i / peer_chunk)
  int first;                // first candidate index handled by this launch
  int count;                // number of candidates handled
  int write_keys;           // 0: poses only (re-pick after the cut)
  double weight;            // written to every picked sample
  double bin_x, bin_y, bin_t;
};

// One thread per candidate sample m: binary search of r_m in the cumulative
// table, pose copy, bin key.  The reference writes weight 1.0 (pf.c:475) and
// rescales to 1/n once n is known (pf.c:493-498); here the weight written is
// 1/max_samples, final whenever the KLD bound does not cut the set (only a cut
// set is rewritten).
__global__ void pick_and_copy(const PickArgs a)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= a.count) return;
  const int m = a.first + t;
  Particle s;
  int from = -1;
  if (a.inject_slot && a.inject_slot[m] > 0)      // slot + 1; 0 = not injected
    s = load_particle(a.injected + (a.inject_slot[m] - 1));
  else
  {
    double r;
    if (a.draws)
      r = a.draws[2 * (size_t)m + 1];
    else
    {
      // the same two uniforms philox_resample_draws would have stored: the first is
      // the recovery test (pf.c:430), unused here because w_diff == 0
      PhiloxStream rng(a.philox_key, (uint64_t)m);
      (void)rng.uniform();
      r = rng.uniform();
    }
    int lo;
    if (a.bucket_index)
    {
      const int b = min((int)(r * (double)a.buckets), a.buckets - 1);   // r < 1
      const int2 range = make_int2(__ldg(a.bucket_index + b), __ldg(a.bucket_index + b + 1));
      lo = table_rank_between(a.cum, r, range.x, min(range.y, a.n_src));
    }
    else
      lo = table_rank(a.coarse, a.n_coarse, a.cum, a.n_src, r);
    if (lo >= a.n_src) lo = a.n_src - 1;  // r beyond the table's last entry
    from = lo;
    // the owner's copy is the current one; a remote owner is read over NVLink
    const Particle *base = a.peers ? a.peers[lo / a.peer_chunk] : a.src;
    s = load_particle(base + lo);
  }
  s.w = a.weight;
  store_particle(a.dst + m, s);
  a.src_index[m] = from;
  if (a.write_keys) a.keys[m] = pose_key(s, a.bin_x, a.bin_y, a.bin_t);
}

__global__ void negate_ints(int *v, int n)
{
  if ((int)threadIdx.x < n) v[threadIdx.x] = -v[threadIdx.x];
}

__global__ void pack_weights(const Particle *__restrict__ set, int n, double *__restrict__ out)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = set[i].w;
}

__global__ void keys_of_set(const Particle *__restrict__ set, int n, unsigned long long *keys,
                            double bin_x, double bin_y, double bin_t)
{
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= n) return;
  keys[m] = pose_key(load_particle(set + m), bin_x, bin_y, bin_t);
}

// ------------------------------------------------------------ pose histogram --
//
// The histogram of occupied bins (the reference's kd-tree leaves) lives in three
// parallel arrays indexed by SLOT: first[] = smallest sample index that fell
// into the bin, aux[] = union-find parent / root, key[] = packed bin key.
//  * DENSE mode: when the bounding box of the keys is small enough (it always
//    is for poses on a map: 400 x 400 x ~40 bins for a 200 m map), a slot IS a
//    cell of a 3-D grid over that box.  Insertion is one atomicMin, neighbour
//    lookup is index arithmetic, and the 4-byte arrays stay L2-resident.
//  * HASH mode (any other case): open addressing on the 64-bit key.
struct BinTable
{
  unsigned long long *key;   // hash mode only
  unsigned int *first;
  unsigned int *aux;
  unsigned int mask;         // hash mode: capacity - 1
  int dense;
  int x0, y0, t0;            // dense: smallest key per axis
  int ex, ey, et;            // dense: extent per axis
  // Number of samples the table is valid for when the host does not know it yet
  // (the single-trip path of small filters): high word of the (n, k) pair
  // find_kld_cut left on the device.  Null: the kernels' n_valid argument counts.
  const unsigned long long *n_dev;
};

__device__ __forceinline__ unsigned int table_valid(const BinTable &T, int n_valid)
{
  return T.n_dev ? (unsigned int)(__ldg(T.n_dev) >> 32) : (unsigned int)n_valid;
}

__host__ __device__ inline void unpack_key(unsigned long long k, int *kx, int *ky, int *kt)
{
  *kx = (int)(k >> 43) - kBiasXY;
  *ky = (int)((k >> 22) & 0x1FFFFFull) - kBiasXY;
  *kt = (int)(k & 0x3FFFFFull) - kBiasTh;
}

// min / max of each key component: out = {minx, miny, mint, maxx, maxy, maxt}
__global__ void key_bounds(const unsigned long long *__restrict__ keys, int n, int *out)
{
  int lo[3] = {0x7FFFFFFF, 0x7FFFFFFF, 0x7FFFFFFF}, hi[3] = {-0x7FFFFFFF, -0x7FFFFFFF, -0x7FFFFFFF};
  for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < n; m += gridDim.x * blockDim.x)
  {
    int k[3];
    unpack_key(keys[m], &k[0], &k[1], &k[2]);
#pragma unroll
    for (int q = 0; q < 3; q++)
    {
      lo[q] = min(lo[q], k[q]);
      hi[q] = max(hi[q], k[q]);
    }
  }
#pragma unroll
  for (int q = 0; q < 3; q++)
  {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
      lo[q] = min(lo[q], __shfl_xor_sync(kFullMask, lo[q], o));
      hi[q] = max(hi[q], __shfl_xor_sync(kFullMask, hi[q], o));
    }
    if ((threadIdx.x & 31) == 0)
    {
      atomicMin(&out[q], lo[q]);
      atomicMax(&out[3 + q], hi[q]);
    }
  }
}

// Samples [base, base + n): keys / slot_of are indexed by the GLOBAL sample index.
__global__ void bin_insert(const unsigned long long *__restrict__ keys, int base, int n,
                           const BinTable T, unsigned int *slot_of)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int m = base + t;
  const unsigned long long key = keys[m];
  if (T.dense)
  {
    int kx, ky, kt;
    unpack_key(key, &kx, &ky, &kt);
    const unsigned int cell =
        (unsigned int)(((kt - T.t0) * T.ey + (ky - T.y0)) * T.ex + (kx - T.x0));
    // converged clouds put millions of samples into a handful of cells: look
    // first, and only issue the atomic when this sample can lower the value
    if ((unsigned int)m < __ldcg(&T.first[cell])) atomicMin(&T.first[cell], (unsigned int)m);
    slot_of[m] = cell;
    return;
  }
  unsigned int h = hash_key(key) & T.mask;
  for (;;)
  {
    const unsigned long long prev = atomicCAS(&T.key[h], kEmptyKey, key);
    if (prev == kEmptyKey || prev == key)
    {
      if ((unsigned int)m < __ldcg(&T.first[h])) atomicMin(&T.first[h], (unsigned int)m);
      slot_of[m] = h;
      return;
    }
    h = (h + 1) & T.mask;
  }
}

// Slot of samples [base, base + n) of `set` in an already built table (read-only).
__global__ void bin_lookup(const Particle *__restrict__ set, int base, int n, const BinTable T,
                           double bin_x, double bin_y, double bin_t, unsigned int *slot_of)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int m = base + t;
  const unsigned long long key = pose_key(load_particle(set + m), bin_x, bin_y, bin_t);
  if (T.dense)
  {
    int kx, ky, kt;
    unpack_key(key, &kx, &ky, &kt);
    const int dx = kx - T.x0, dy = ky - T.y0, dt = kt - T.t0;
    // a pose outside the grid cannot have been entered into it: no slot
    const bool inside = dx >= 0 && dx < T.ex && dy >= 0 && dy < T.ey && dt >= 0 && dt < T.et;
    slot_of[m] = inside ? (unsigned int)((dt * T.ey + dy) * T.ex + dx) : 0xFFFFFFFFu;
    return;
  }
  unsigned int h = hash_key(key) & T.mask;
  while (T.key[h] != key && T.key[h] != kEmptyKey) h = (h + 1) & T.mask;
  slot_of[m] = T.key[h] == key ? h : 0xFFFFFFFFu;
}

__global__ void mark_first(const unsigned int *__restrict__ first, const unsigned int *__restrict__ slot_of,
                           int base, int n, unsigned int *flag)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int m = base + t;
  flag[m] = first[slot_of[m]] == (unsigned int)m ? 1u : 0u;
}

// pf.c:515-555 with IEEE ops in source order and no contraction.  The final
// (int) conversion follows x86 (cvttsd2si): out-of-range -> INT_MIN.
__device__ __forceinline__ int resample_limit(int k, int min_samples, int max_samples,
                                              double pop_err, double pop_z)
{
  if (k < 1 || k > max_samples) return max_samples;
  if (k == 1) return max_samples;
  const double km1 = (double)k - 1.0;
  const double b = __ddiv_rn(2.0, __dmul_rn(9.0, km1));
  const double cc = __dmul_rn(__dsqrt_rn(__ddiv_rn(2.0, __dmul_rn(9.0, km1))), pop_z);
  const double x = __dadd_rn(__dsub_rn(1.0, b), cc);
  double v = __ddiv_rn((double)(k - 1), __dmul_rn(2.0, pop_err));
  v = __dmul_rn(v, x);
  v = __dmul_rn(v, x);
  v = __dmul_rn(v, x);
  v = ceil(v);
  int n;
  if (!(v > -2147483649.0 && v < 2147483648.0)) n = (int)0x80000000;
  else n = (int)v;
  if (n < min_samples) return min_samples;
  if (n > max_samples) return max_samples;
  return n;
}

// n = min over m of (m+1) with (m+1) > limit(k_m)   (pf.c:483), and with it
// leaf_count = k_{n-1}: one 64-bit minimum over (m + 1) << 32 | k_m.  The last
// candidate always offers (max_samples, k_last), the answer when nothing cuts.
// Samples [base, base + count); k_m = k_offset + k_inclusive[m], the offset being
// the number of distinct bins first seen on lower ranks of a sharded run (summed
// here from the all-reduced per-rank counts: no trip to the host).
__global__ void find_kld_cut(const unsigned int *__restrict__ k_inclusive, int base, int count,
                             const int *__restrict__ rank_counts, int rank, int min_samples,
                             int max_samples, double pop_err, double pop_z, unsigned long long *cutk)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  int k_offset = 0;
  for (int r = 0; r < rank; r++) k_offset += rank_counts[r];
  const int m = base + t;
  const int k = k_offset + (int)k_inclusive[m];
  const unsigned long long mine = ((unsigned long long)(unsigned int)(m + 1) << 32) | (unsigned int)k;
  if (m + 1 == max_samples)
  {
    atomicMin(cutk, mine);
    return;
  }
  // limit() only changes where k changes, but evaluating it everywhere keeps
  // the kernel branch-free; min_samples gives a cheap early reject.
  if (m + 1 <= min_samples) return;
  if (mine >= __ldcg(cutk)) return;   // cannot lower the cut: skip the formula and the atomic
  if (m + 1 > resample_limit(k, min_samples, max_samples, pop_err, pop_z)) atomicMin(cutk, mine);
}

__global__ void set_uniform_weight(Particle *set, int n, double w)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) set[i].w = w;
}

// ------------------------------------------------------------ cluster labels --

// Parent pointers are read with L2 (.cg) loads: L1 is not coherent across SMs
// and a stale "I am a root" would make the CAS in uf_union spin.
__device__ __forceinline__ unsigned int uf_find(unsigned int *aux, unsigned int x)
{
  unsigned int p = __ldcg(&aux[x]);
  while (p != x)
  {
    const unsigned int gp = __ldcg(&aux[p]);
    if (gp != p) aux[x] = gp;  // path halving (benign race: gp is an ancestor)
    x = p;
    p = gp;
  }
  return x;
}

__device__ __forceinline__ void uf_union(unsigned int *aux, unsigned int a, unsigned int b)
{
  for (;;)
  {
    a = uf_find(aux, a);
    b = uf_find(aux, b);
    if (a == b) return;
    if (a > b) { unsigned int s = a; a = b; b = s; }
    // hang the larger root under the smaller one
    if (atomicCAS(&aux[b], b, a) == b) return;
  }
}

// The labelling kernels below run ONE THREAD PER TABLE SLOT (grid cell or hash
// slot), not per sample: neighbouring threads then work on neighbouring bins,
// the arrays are read in order, and the unions arrive in a scan-line-like
// pattern that keeps the trees shallow.  A slot is occupied when a sample below
// n_valid fell into it (empty slots hold 0xFFFFFFFF).

// every occupied bin starts as its own component
__global__ void cc_init(const BinTable T, unsigned int nslots, int n_valid, unsigned int *cmin)
{
  const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots || T.first[s] >= table_valid(T, n_valid)) return;
  T.aux[s] = s;
  cmin[s] = 0xFFFFFFFFu;
}

// pf_kdtree.c:407-438: bins are neighbours when every key component differs
// by at most one (26-neighbourhood, heading NOT wrapped).  Only the 13
// "forward" neighbours are visited: the relation is symmetric.
__global__ void cc_link(const BinTable T, unsigned int nslots, int n_valid)
{
  const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots || T.first[s] >= table_valid(T, n_valid)) return;
  if (T.dense)
  {
    const int ix = (int)(s % (unsigned)T.ex);
    const int iy = (int)((s / (unsigned)T.ex) % (unsigned)T.ey);
    const int it = (int)(s / ((unsigned)T.ex * (unsigned)T.ey));
    for (int q = 14; q < 27; q++)
    {
      const int dt = q / 9 - 1, dy = (q % 9) / 3 - 1, dx = (q % 9) % 3 - 1;   // x fastest
      const int nx = ix + dx, ny = iy + dy, nt = it + dt;
      if ((unsigned)nx >= (unsigned)T.ex || (unsigned)ny >= (unsigned)T.ey ||
          (unsigned)nt >= (unsigned)T.et)
        continue;
      const unsigned int nb = (unsigned int)((nt * T.ey + ny) * T.ex + nx);
      if (T.first[nb] < table_valid(T, n_valid)) uf_union(T.aux, s, nb);
    }
    return;
  }
  const unsigned long long key = T.key[s];
  for (int q = 14; q < 27; q++)
  {
    const long long dx = q / 9 - 1, dy = (q % 9) / 3 - 1, dt = (q % 9) % 3 - 1;
    const unsigned long long nk = key + (unsigned long long)(dx * (1ll << 43) + dy * (1ll << 22) + dt);
    unsigned int h = hash_key(nk) & T.mask;
    for (;;)
    {
      const unsigned long long kk = T.key[h];
      if (kk == kEmptyKey) break;
      if (kk == nk)
      {
        if (T.first[h] < table_valid(T, n_valid)) uf_union(T.aux, s, h);
        break;
      }
      h = (h + 1) & T.mask;
    }
  }
}

// ---- dense grid: tile-local labelling in shared memory, then the tile borders -----
//
// A CTA owns a 32 x 8 x 4 brick of the grid (one thread per cell).  The 13
// forward neighbours that lie inside the brick are united in a shared-memory
// union-find (tens of cycles per hop instead of an L2 round trip); each cell
// then points at its brick-local root in the global array.  A second kernel
// unites across brick faces only -- about a quarter of all edges.
constexpr int kBrickX = 32, kBrickY = 8, kBrickT = 4;
constexpr int kBrickCells = kBrickX * kBrickY * kBrickT;

__device__ __forceinline__ unsigned int suf_find(unsigned int *lab, unsigned int x)
{
  unsigned int p = lab[x];
  while (p != x)
  {
    const unsigned int gp = lab[p];
    if (gp != p) lab[x] = gp;
    x = p;
    p = gp;
  }
  return x;
}

__device__ __forceinline__ void suf_union(unsigned int *lab, unsigned int a, unsigned int b)
{
  for (;;)
  {
    a = suf_find(lab, a);
    b = suf_find(lab, b);
    if (a == b) return;
    if (a > b) { unsigned int t = a; a = b; b = t; }
    if (atomicCAS(&lab[b], b, a) == b) return;
  }
}

__global__ void __launch_bounds__(kBrickCells)
cc_brick_local(const BinTable T, int bricks_x, int bricks_y, int n_valid, unsigned int *cmin, int band,
               int owner)
{
  // band > 0 (sharded labelling): rank `owner` labels the bricks of its band of
  // `band` brick rows along y, every block of another band leaves at once
  const int brick = (int)blockIdx.x;
  if (band > 0 && ((brick / bricks_x) % bricks_y) / band != owner) return;
  __shared__ unsigned int s_lab[kBrickCells];
  __shared__ unsigned char s_occ[kBrickCells];
  const int lx = threadIdx.x & (kBrickX - 1);
  const int ly = (threadIdx.x / kBrickX) & (kBrickY - 1);
  const int lt = threadIdx.x / (kBrickX * kBrickY);
  const int bx = brick % bricks_x;
  const int by = (brick / bricks_x) % bricks_y;
  const int bt = brick / (bricks_x * bricks_y);
  const int ix = bx * kBrickX + lx, iy = by * kBrickY + ly, it = bt * kBrickT + lt;
  const bool inside = ix < T.ex && iy < T.ey && it < T.et;
  const unsigned int s = inside ? (unsigned int)((it * T.ey + iy) * T.ex + ix) : 0u;
  const bool occ = inside && T.first[s] < table_valid(T, n_valid);
  s_lab[threadIdx.x] = threadIdx.x;
  s_occ[threadIdx.x] = occ ? 1 : 0;
  __syncthreads();
  if (occ)
  {
    for (int q = 14; q < 27; q++)
    {
      const int dt = q / 9 - 1, dy = (q % 9) / 3 - 1, dx = (q % 9) % 3 - 1;
      const int nx = lx + dx, ny = ly + dy, nt = lt + dt;
      if ((unsigned)nx >= (unsigned)kBrickX || (unsigned)ny >= (unsigned)kBrickY ||
          (unsigned)nt >= (unsigned)kBrickT)
        continue;
      const unsigned int nb = (unsigned int)((nt * kBrickY + ny) * kBrickX + nx);
      if (!s_occ[nb]) continue;
      // A diagonal link is implied when a cell "between" the two -- the
      // neighbour with one of the offset's components set back to zero -- is
      // occupied: that cell is a face neighbour of nb and a less diagonal
      // neighbour of this cell, so by induction on the number of non-zero
      // components the face links alone already connect them.  (Only cells of
      // this brick are consulted here; cc_brick_borders applies the same rule
      // with the global occupancy.)
      bool implied = false;
      if ((dx != 0) + (dy != 0) + (dt != 0) >= 2)
      {
        if (dx != 0 && s_occ[(nt * kBrickY + ny) * kBrickX + lx]) implied = true;
        if (dy != 0 && s_occ[(nt * kBrickY + ly) * kBrickX + nx]) implied = true;
        if (dt != 0 && s_occ[(lt * kBrickY + ny) * kBrickX + nx]) implied = true;
      }
      if (!implied) suf_union(s_lab, threadIdx.x, nb);
    }
  }
  __syncthreads();
  if (occ)
  {
    const unsigned int r = suf_find(s_lab, threadIdx.x);
    const int rx = r & (kBrickX - 1), ry = (r / kBrickX) & (kBrickY - 1), rt = r / (kBrickX * kBrickY);
    T.aux[s] = (unsigned int)(((bt * kBrickT + rt) * T.ey + (by * kBrickY + ry)) * T.ex +
                              (bx * kBrickX + rx));
    cmin[s] = 0xFFFFFFFFu;
  }
}

__global__ void __launch_bounds__(kBrickCells)
cc_brick_borders(const BinTable T, int bricks_x, int bricks_y, int n_valid, int band, int owner, int phase)
{
  // Sharded labelling gives every rank a band of `band` brick rows along y, all
  // headings and all x (band > 0): bands only touch on y faces.  Phase 1 (the
  // rank's own bricks): pairs inside the band.  Phase 2 (after the ranks' forests
  // have been merged; only the brick rows at a band edge stay): the pairs that
  // straddle two bands.  Unsharded (band == 0): one pass, every pair.
  //
  // A warp is 32 x-neighbours of one (y, heading) row of the brick; on a brick
  // face most of them want to unite the SAME two brick-local trees (after
  // cc_brick_local, aux[] of a cell is its brick-local root).  Lanes that hold the
  // same (tree of this cell, tree of the neighbour) pair elect one to do the
  // union: every pair of trees is united once per warp instead of once per cell.
  const int brick = (int)blockIdx.x;
  const int lx = threadIdx.x & (kBrickX - 1);
  const int ly = (threadIdx.x / kBrickX) & (kBrickY - 1);
  const int lt = threadIdx.x / (kBrickX * kBrickY);
  const int bx = brick % bricks_x;
  const int by = (brick / bricks_x) % bricks_y;
  const int bt = brick / (bricks_x * bricks_y);
  if (band > 0)
  {
    if (phase == 1 && by / band != owner) return;                     // block-uniform
    if (phase == 2 && !((by % band == band - 1 && by + 1 < bricks_y) || (by % band == 0 && by > 0))) return;
  }
  const int ix = bx * kBrickX + lx, iy = by * kBrickY + ly, it = bt * kBrickT + lt;
  const unsigned int nv = table_valid(T, n_valid);
  // occupancy of the brick and of the one-cell shell its forward neighbours and
  // the "cells between" can lie in (x, y: -1 .. size; heading: 0 .. size), read
  // once, in rows, instead of up to four scattered loads per neighbour and cell
  constexpr int kHX = kBrickX + 2, kHY = kBrickY + 2, kHT = kBrickT + 1;
  __shared__ unsigned char s_occ[kHX * kHY * kHT];
  for (int k = threadIdx.x; k < kHX * kHY * kHT; k += blockDim.x)
  {
    const int hx = k % kHX - 1, hy = (k / kHX) % kHY - 1, ht = k / (kHX * kHY);
    const int gx = bx * kBrickX + hx, gy = by * kBrickY + hy, gt = bt * kBrickT + ht;
    bool o = false;
    if ((unsigned)gx < (unsigned)T.ex && (unsigned)gy < (unsigned)T.ey && (unsigned)gt < (unsigned)T.et)
      o = T.first[(gt * T.ey + gy) * T.ex + gx] < nv;
    s_occ[k] = o ? 1 : 0;
  }
  __syncthreads();
#define AMCL_OCC(x, y, t) s_occ[((t) * kHY + ((y) + 1)) * kHX + ((x) + 1)]
  bool active = ix < T.ex && iy < T.ey && it < T.et;
  // only cells on a forward face (or next to the low-x / low-y face, whose
  // forward neighbours with dx or dy = -1 leave the brick) have work to do
  if (lx != 0 && lx != kBrickX - 1 && ly != 0 && ly != kBrickY - 1 && lt != kBrickT - 1) active = false;
  if (active && !AMCL_OCC(lx, ly, lt)) active = false;
  // Pairs of trees to unite are first collected per BLOCK (a brick has a handful
  // of local trees and so have its neighbours: ~10 distinct pairs, against ~150
  // warp-level candidates), then united once each: the unions are chains of
  // dependent L2 round trips on a few hot roots, their count is what costs.
  constexpr int kPairSlots = 256;
  __shared__ unsigned long long s_pair[kPairSlots];
  __shared__ uint2 s_pair_cells[kPairSlots];
  for (int k = threadIdx.x; k < kPairSlots; k += blockDim.x) s_pair[k] = ~0ull;
  __syncthreads();
  const unsigned int s = active ? (unsigned int)((it * T.ey + iy) * T.ex + ix) : 0u;
  const unsigned int mine = active ? __ldcg(&T.aux[s]) : 0u;   // this cell's tree
  const bool warp_has_work = __ballot_sync(kFullMask, active) != 0u;
  for (int q = 14; warp_has_work && q < 27; q++)
  {
    const int dt = q / 9 - 1, dy = (q % 9) / 3 - 1, dx = (q % 9) % 3 - 1;
    const int mx = lx + dx, my = ly + dy, mt = lt + dt;
    bool want = active;
    if ((unsigned)mx < (unsigned)kBrickX && (unsigned)my < (unsigned)kBrickY && (unsigned)mt < (unsigned)kBrickT)
      want = false;                                 // inside the brick: done locally
    if (want && !AMCL_OCC(mx, my, mt)) want = false;   // (cells beyond the grid read as empty)
    if (want && band > 0)
    {
      const bool same_owner = ((iy + dy) / kBrickY) / band == by / band;
      if (same_owner != (phase == 1)) want = false;
    }
    if (want && (dx != 0) + (dy != 0) + (dt != 0) >= 2)
    {
      // same pruning as in cc_brick_local, now with any cell of the grid in between
      if ((dx != 0 && AMCL_OCC(lx, my, mt)) || (dy != 0 && AMCL_OCC(mx, ly, mt)) ||
          (dt != 0 && AMCL_OCC(mx, my, lt)))
        want = false;
    }
    const unsigned int voters = __ballot_sync(kFullMask, want);
    if (voters == 0u) continue;
    if (want)
    {
      const unsigned int nb = (unsigned int)(((it + dt) * T.ey + (iy + dy)) * T.ex + (ix + dx));
      const unsigned long long pair = ((unsigned long long)mine << 32) | __ldcg(&T.aux[nb]);
      const unsigned int peers = __match_any_sync(voters, pair);
      if ((int)(threadIdx.x & 31) == __ffs(peers) - 1)
      {
        // enter the pair into the block's set; a full set falls back to uniting at once
        unsigned int h = (unsigned int)((pair * 0x9E3779B97F4A7C15ull) >> 56);   // 0..255
        bool placed = false;
        for (int probe = 0; probe < 16 && !placed; probe++, h = (h + 1) & (kPairSlots - 1))
        {
          const unsigned long long prev = atomicCAS(&s_pair[h], ~0ull, pair);
          if (prev == ~0ull)
          {
            s_pair_cells[h] = make_uint2(s, nb);
            placed = true;
          }
          else if (prev == pair)
            placed = true;
        }
        if (!placed) uf_union(T.aux, s, nb);
      }
    }
  }
#undef AMCL_OCC
  __syncthreads();
  for (int k = threadIdx.x; k < kPairSlots; k += blockDim.x)
    if (s_pair[k] != ~0ull) uf_union(T.aux, s_pair_cells[k].x, s_pair_cells[k].y);
}

// Every bin follows its parents to the root and points at it directly.  Other
// threads only ever overwrite a parent pointer with that bin's root, so a chain
// read while it is being shortened still ends at the same root.
__global__ void cc_flatten(const BinTable T, unsigned int nslots, int n_valid)
{
  const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots || T.first[s] >= table_valid(T, n_valid)) return;
  unsigned int p = __ldcg(&T.aux[s]);
  for (;;)
  {
    const unsigned int gp = __ldcg(&T.aux[p]);
    if (gp == p) break;
    p = gp;
  }
  T.aux[s] = p;
}

// Few clusters (the usual case): the roots are collected in a short list and
// ranked by their earliest sample directly, instead of flagging and scanning
// all n sample indices.
constexpr int kRootListCap = 4096;

__global__ void cc_collect_roots(const BinTable T, unsigned int nslots, int n_valid,
                                 const unsigned int *__restrict__ cmin, unsigned int *list,
                                 unsigned int *keys, int *count)
{
  const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots || T.first[s] >= table_valid(T, n_valid)) return;
  if (T.aux[s] != s) return;
  const int k = atomicAdd(count, 1);
  if (k < kRootListCap)
  {
    list[k] = s;
    keys[k] = cmin[s];
  }
}

// id of a root = number of roots whose earliest sample comes before its own
__global__ void cc_rank_roots(const unsigned int *__restrict__ list, const unsigned int *__restrict__ keys,
                              int count, unsigned int *cmin_to_id, const int *count_dev = nullptr)
{
  __shared__ unsigned int s_keys[256];
  if (count_dev) count = min(*count_dev, kRootListCap);
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned int mine = t < count ? keys[t] : 0u;
  unsigned int rank = 0;
  for (int base = 0; base < count; base += 256)
  {
    __syncthreads();
    if (base + (int)threadIdx.x < count) s_keys[threadIdx.x] = keys[base + threadIdx.x];
    __syncthreads();
    const int m = min(256, count - base);
    for (int k = 0; k < m; k++) rank += s_keys[k] < mine;
  }
  if (t < count) cmin_to_id[list[t]] = rank;
}

// smallest first-sample index of every component, kept at its root
__global__ void cc_component_min(const BinTable T, unsigned int nslots, int n_valid,
                                 unsigned int *cmin)
{
  const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots) return;
  const unsigned int f = T.first[s];
  if (f >= table_valid(T, n_valid)) return;
  // A giant component would otherwise serialise millions of atomics on one
  // address: lanes of a warp that share a root combine first, and the group's
  // leader looks before it competes.
  const unsigned int root = T.aux[s];
  const unsigned int peers = __match_any_sync(__activemask(), root);
  const unsigned int best = __reduce_min_sync(peers, f);
  if ((int)(threadIdx.x & 31) == __ffs(peers) - 1)
  {
    unsigned int *slot = &cmin[root];
    if (best < __ldcg(slot)) atomicMin(slot, best);
  }
}

// Canonical cluster numbers: clusters are numbered in the order in which their
// earliest sample appears.  Every root bin flags that sample's index; an
// inclusive scan over sample indices ranks the flags; the root reads its rank
// back.  Works on the table alone, so a sharded rank needs no other rank's
// per-sample data.
__global__ void cc_flag_roots(const BinTable T, unsigned int nslots, int n_valid,
                              const unsigned int *__restrict__ cmin, unsigned int *cflag)
{
  const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots || T.first[s] >= table_valid(T, n_valid)) return;
  if (T.aux[s] == s) cflag[cmin[s]] = 1u;
}

__global__ void cc_assign_ids(const BinTable T, unsigned int nslots, int n_valid,
                              const unsigned int *__restrict__ cscan, unsigned int *cmin_to_id)
{
  const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots || T.first[s] >= table_valid(T, n_valid)) return;
  if (T.aux[s] == s) cmin_to_id[s] = cscan[cmin_to_id[s]] - 1u;   // cmin[root] -> cluster id
}

// ----------------------------------------------------------------- statistics --

// pf.c:603-643: per sample, ten weighted sums into its cluster's accumulators
// [count, w, w x, w y, w cos, w sin, w xx, w xy, w yx, w yy].
//
// Persistent warps walk the samples 32 at a time.  Each warp keeps a "hot"
// label and every lane keeps private partial sums for it, so the common cases
// -- a converged cloud (one cluster) or a global cloud (one giant free-space
// component plus stragglers) -- cost ten atomics per WARP LIFETIME instead of
// ten per sample.  Lanes on other labels are combined with shuffles (up to
// three further labels per batch) and only then fall back to single atomics.
__device__ __forceinline__ void flush_hot(double *acc, unsigned int hot, double h[10], int lane)
{
  if (hot == 0xFFFFFFFFu) return;   // warp-uniform
#pragma unroll
  for (int q = 0; q < 10; q++)
  {
    const double t = warp_sum(h[q]);
    if (lane == 0 && t != 0.0) atomicAdd(&acc[(size_t)hot * 10 + q], t);
    h[q] = 0.0;
  }
}

__global__ void cluster_accumulate(const Particle *__restrict__ set, int n,
                                   const unsigned int *__restrict__ root_of_slot,
                                   const unsigned int *__restrict__ slot_of,
                                   const unsigned int *__restrict__ cid_of_root, double *acc,
                                   int *label_out, int first, const unsigned long long *n_dev = nullptr)
{
  // samples [first, first + n) of the set (a sharded rank sums its own slice);
  // n_dev: the count is the high word of the (n, k) pair still on the device
  if (n_dev) n = (int)(__ldg(n_dev) >> 32);
  set += first;
  slot_of += first;
  if (label_out) label_out += first;
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int warp_global = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  const int warp_stride = gridDim.x * warps_per_block;
  unsigned int hot = 0xFFFFFFFFu;
  double h[10];
#pragma unroll
  for (int q = 0; q < 10; q++) h[q] = 0.0;

  for (long long base = (long long)warp_global * 32; base < n; base += (long long)warp_stride * 32)
  {
    const int i = (int)base + lane;
    const bool live = i < n;
    unsigned int label = 0xFFFFFFFFu;
    double v[10];
#pragma unroll
    for (int q = 0; q < 10; q++) v[q] = 0.0;
    if (live)
    {
      const Particle s = load_particle(set + i);
      // (a sample without a slot or in a bin nobody entered gets no label and is
      // left out of the sums; cannot happen while every rank sees the same poses)
      const unsigned int slot = slot_of[i];
      const unsigned int root = slot != 0xFFFFFFFFu ? root_of_slot[slot] : 0xFFFFFFFFu;
      label = root != 0xFFFFFFFFu ? cid_of_root[root] : 0xFFFFFFFFu;
      double sn, cs;
      sincos(s.th, &sn, &cs);
      v[0] = 1.0;
      v[1] = s.w;
      v[2] = s.w * s.x;
      v[3] = s.w * s.y;
      v[4] = s.w * cs;
      v[5] = s.w * sn;
      v[6] = s.w * s.x * s.x;
      v[7] = s.w * s.x * s.y;
      v[8] = s.w * s.y * s.x;
      v[9] = s.w * s.y * s.y;
      if (label_out) label_out[i] = (int)label;
    }
    const bool valid = live && label != 0xFFFFFFFFu;
    unsigned remaining = __ballot_sync(kFullMask, valid);
    if (remaining == 0u) continue;
    unsigned hot_group = __ballot_sync(kFullMask, valid && label == hot);
    if (hot_group == 0u)
    {
      // nobody here belongs to the hot cluster any more: retire it, adopt the
      // label of the first valid lane
      flush_hot(acc, hot, h, lane);
      hot = __shfl_sync(kFullMask, label, __ffs(remaining) - 1);
      hot_group = __ballot_sync(kFullMask, valid && label == hot);
    }
    if (valid && label == hot)
    {
#pragma unroll
      for (int q = 0; q < 10; q++) h[q] += v[q];
    }
    remaining &= ~hot_group;
    int rounds = 0;
    while (remaining != 0u && rounds < 3)
    {
      const int leader = __ffs(remaining) - 1;
      const unsigned int lab = __shfl_sync(kFullMask, label, leader);
      const bool mine = valid && label == lab;
      const unsigned group = __ballot_sync(kFullMask, mine);
#pragma unroll
      for (int q = 0; q < 10; q++)
      {
        const double t = warp_sum(mine ? v[q] : 0.0);
        if (lane == leader) atomicAdd(&acc[(size_t)lab * 10 + q], t);
      }
      remaining &= ~group;
      rounds++;
    }
    if (valid && ((remaining >> lane) & 1u))
    {
#pragma unroll
      for (int q = 0; q < 10; q++) atomicAdd(&acc[(size_t)label * 10 + q], v[q]);
    }
  }
  flush_hot(acc, hot, h, lane);
}

// sum x, sum y (pf.c:231-236) -> block partials (two lanes of the partial array)
__global__ void sum_xy(const Particle *__restrict__ set, int n, double *px, double *py)
{
  __shared__ double s_red[32];
  double ax = 0.0, ay = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    const double2 xy = *reinterpret_cast<const double2 *>(set + i);
    ax += xy.x;
    ay += xy.y;
  }
  const double tx = block_sum(ax, s_red);
  const double ty = block_sum(ay, s_red);
  if (threadIdx.x == 0)
  {
    px[blockIdx.x] = tx;
    py[blockIdx.x] = ty;
  }
}

__global__ void finish_xy(const double *px, const double *py, int count, int n, double *out)
{
  __shared__ double s_red[32];
  double ax = 0.0, ay = 0.0;
  for (int k = threadIdx.x; k < count; k += blockDim.x)
  {
    ax += px[k];
    ay += py[k];
  }
  const double tx = block_sum(ax, s_red);
  const double ty = block_sum(ay, s_red);
  if (threadIdx.x == 0)
  {
    out[0] = tx / n;   // pf.c:237-238
    out[1] = ty / n;
  }
}

// pf.c:240-247: any sample farther than dist_threshold from the mean on either axis
__global__ void any_outside(const Particle *__restrict__ set, int n, const double *mean,
                            double thr, int *flag)
{
  const double mx = mean[0], my = mean[1];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    const double2 xy = *reinterpret_cast<const double2 *>(set + i);
    if (fabs(xy.x - mx) > thr || fabs(xy.y - my) > thr)
    {
      *flag = 1;
      return;
    }
  }
}

// pf.c:701-740 raw sums: mn, mx, my, mrr
__global__ void cep_sums(const Particle *__restrict__ set, int n, double *partials, int stride)
{
  __shared__ double s_red[32];
  double a[4] = {0, 0, 0, 0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    const Particle s = load_particle(set + i);
    a[0] += s.w;
    a[1] += s.w * s.x;
    a[2] += s.w * s.y;
    a[3] += s.w * s.x * s.x;
    a[3] += s.w * s.y * s.y;
  }
  for (int q = 0; q < 4; q++)
  {
    const double t = block_sum(a[q], s_red);
    if (threadIdx.x == 0) partials[q * stride + blockIdx.x] = t;
  }
}

__global__ void finish_multi(const double *partials, int stride, int count, int lanes, double *out)
{
  __shared__ double s_red[32];
  for (int q = 0; q < lanes; q++)
  {
    double v = 0.0;
    for (int k = threadIdx.x; k < count; k += blockDim.x) v += partials[q * stride + k];
    const double t = block_sum(v, s_red);
    if (threadIdx.x == 0) out[q] = t;
    __syncthreads();
  }
}

__global__ void sum_weights(const Particle *__restrict__ set, int n, double *partials)
{
  __shared__ double s_red[32];
  double a = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    a += set[i].w;
  const double t = block_sum(a, s_red);
  if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

__global__ void copy_scaled(const Particle *__restrict__ src, Particle *dst, int n,
                            const double *total)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Particle s = load_particle(src + i);
  s.w = s.w / total[0];   // pf.c:353-357
  store_particle(dst + i, s);
}

// ---- single-trip path of small filters: kernels that take n from the device ----------

// pf.c:493-498 for a set whose size the host has not seen yet
__global__ void set_uniform_weight_dev(Particle *set, int max_n, const unsigned long long *cutk)
{
  const int n = (int)(__ldg(cutk) >> 32);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && i < max_n) set[i].w = 1.0 / (double)n;
}

// pf_update_converged (pf.c:220-251) in one block: mean x / y in a fixed order,
// then "is any sample farther than the threshold from it".  out[0] = 1 if so.
__global__ void __launch_bounds__(1024)
converged_small(const Particle *__restrict__ set, const unsigned long long *cutk, double thr, int *out)
{
  __shared__ double s_red[32];
  __shared__ double s_mean[2];
  __shared__ int s_flag;
  const int n = (int)(__ldg(cutk) >> 32);
  double ax = 0.0, ay = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x)
  {
    const double2 xy = *reinterpret_cast<const double2 *>(set + i);
    ax += xy.x;
    ay += xy.y;
  }
  const double tx = block_sum(ax, s_red);
  const double ty = block_sum(ay, s_red);
  if (threadIdx.x == 0)
  {
    s_mean[0] = tx / n;
    s_mean[1] = ty / n;
    s_flag = 0;
  }
  __syncthreads();
  const double mx = s_mean[0], my = s_mean[1];
  for (int i = threadIdx.x; i < n; i += blockDim.x)
  {
    const double2 xy = *reinterpret_cast<const double2 *>(set + i);
    if (fabs(xy.x - mx) > thr || fabs(xy.y - my) > thr) s_flag = 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) out[0] = s_flag;
}

// Filters up to this many samples take the single-trip path of run_resample.
constexpr int kSmallFilter = 16384;
constexpr int kSmallHeadPinned = 4;    // clusters whose sums ride along in the pinned scalar block

// ------------------------------------------------------------------- launchers --

static inline int blocks_for(int n, int threads) { return n > 0 ? (n + threads - 1) / threads : 1; }

static int reduce_grid(int n)
{
  int grid = globals().sm_count * 4;
  int need = blocks_for(n, 256);
  return grid < need ? grid : need;
}

static void ensure_partials(PfCtx *c, int count)
{
  if (count <= c->partials_cap) return;
  if (c->d_partials) cudaFree(c->d_partials);
  c->partials_cap = count;
  AMCL_CUDA(cudaMalloc(&c->d_partials, sizeof(double) * (size_t)count));
}

static void sync_ints(PfCtx *c, int first, int count)
{
  AMCL_CUDA(cudaMemcpyAsync(c->h_ints + first, c->d_ints + first, sizeof(int) * (size_t)count,
                            cudaMemcpyDeviceToHost, globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
}

static void sync_scalars(PfCtx *c, int first, int count)
{
  AMCL_CUDA(cudaMemcpyAsync(c->h_scalars + first, c->d_scalars + first, sizeof(double) * (size_t)count,
                            cudaMemcpyDeviceToHost, globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
}

// d_ints slots
enum { I_DISTINCT = 0, I_CUT = 1, I_CLUSTERS = 2, I_OUTSIDE = 3, I_MOVED = 4, I_KCUT = 5,
       I_CUTK = 6 /* and 7: one 64-bit word */, I_RANKCNT = 24 };

static void bin_sizes(const pf_sample_set_t *set, double *bx, double *by, double *bt)
{
  *bx = set->kdtree->size[0];
  *by = set->kdtree->size[1];
  *bt = set->kdtree->size[2];
}

// Histogram over samples [0, count) of `which` from precomputed keys.
// d_ints slots 16..21 hold the key bounds
enum { I_BOUNDS = 16 };

static BinTable table_of(const PfCtx *c)
{
  BinTable T;
  T.key = c->d_tab_key;
  T.first = c->d_tab_first;
  T.aux = c->d_tab_aux;
  T.mask = c->hash_cap - 1;
  T.dense = c->tab_dense;
  T.x0 = c->tab_lo[0]; T.y0 = c->tab_lo[1]; T.t0 = c->tab_lo[2];
  T.ex = c->tab_ext[0]; T.ey = c->tab_ext[1]; T.et = c->tab_ext[2];
  T.n_dev = nullptr;
  return T;
}

// Histogram over the first `count` keys of d_keys: slot per sample, first-
// occurrence flags.  Chooses the dense grid when the keys' bounding box fits
// the table, the hash otherwise.
static void build_histogram(PfCtx *c, int count)
{
  cudaStream_t st = globals().stream;
  int *h = c->h_ints + I_BOUNDS;
  for (int q = 0; q < 3; q++) { h[q] = 0x7FFFFFFF; h[3 + q] = -0x7FFFFFFF; }
  AMCL_CUDA(cudaMemcpyAsync(c->d_ints + I_BOUNDS, h, sizeof(int) * 6, cudaMemcpyHostToDevice, st));
  int grid = globals().sm_count * 8;
  if (grid > blocks_for(count, 256)) grid = blocks_for(count, 256);
  key_bounds<<<grid, 256, 0, st>>>(c->d_keys, count, c->d_ints + I_BOUNDS);
  AMCL_LAUNCH_CHECK();
  sync_ints(c, I_BOUNDS, 6);
  double cells = 1.0;
  for (int q = 0; q < 3; q++)
  {
    c->tab_lo[q] = h[q];
    c->tab_ext[q] = count > 0 ? h[3 + q] - h[q] + 1 : 1;
    cells *= (double)c->tab_ext[q];
  }
  c->tab_dense = (count > 0 && cells <= (double)c->hash_cap) ? 1 : 0;
  if (c->tab_dense)
    AMCL_CUDA(cudaMemsetAsync(c->d_tab_first, 0x7F, sizeof(unsigned int) * (size_t)cells, st));
  else
  {
    AMCL_CUDA(cudaMemsetAsync(c->d_tab_key, 0xFF, sizeof(unsigned long long) * (size_t)c->hash_cap, st));
    AMCL_CUDA(cudaMemsetAsync(c->d_tab_first, 0xFF, sizeof(unsigned int) * (size_t)c->hash_cap, st));
  }
  const BinTable T = table_of(c);
  bin_insert<<<blocks_for(count, 256), 256, 0, st>>>(c->d_keys, 0, count, T, c->d_slot);
  AMCL_LAUNCH_CHECK();
  mark_first<<<blocks_for(count, 256), 256, 0, st>>>(c->d_tab_first, c->d_slot, 0, count, c->d_flag);
  AMCL_LAUNCH_CHECK();
}

static void publish_clusters(pf_sample_set_t *set, const double *acc, int ncl);

// Labels + statistics for samples [0, n) of set `which`; the histogram
// (d_slot and the bin table) must describe at least those samples.
static void cluster_and_stats(PfCtx *c, int which, int n)
{
  // Labelling runs on the whole (replicated) histogram on every rank and is
  // deterministic, so all ranks agree on the labels; the per-sample sums are
  // taken over the owned slice only and all-reduced.
  const bool sliced = c->shard.on() && !c->replicated[which];
  const int acc_lo = sliced ? c->shard.lo(n) : 0;
  const int acc_cnt = sliced ? c->shard.count(n) : n;
  cudaStream_t st = globals().stream;
  pf_t *pf = c->host;
  pf_sample_set_t *set = pf->sets + which;
  unsigned int *cmin = c->d_scan + c->alloc_samples;  // second half: per-slot scratch (hash_cap)
  const int nb = blocks_for(n, 256);

  const BinTable T = table_of(c);
  const unsigned int nslots =
      T.dense ? (unsigned int)T.ex * (unsigned int)T.ey * (unsigned int)T.et : c->hash_cap;
  const int sb = (int)((nslots + 255u) / 256u);
  if (T.dense)
  {
    const int bxn = (T.ex + kBrickX - 1) / kBrickX, byn = (T.ey + kBrickY - 1) / kBrickY,
              btn = (T.et + kBrickT - 1) / kBrickT;
    const int bricks = bxn * byn * btn;
    const Shard &sh = c->shard;
    if (sh.on() && bricks >= 64 * sh.world)
    {
      // Every rank labels a band of brick rows along y (all headings, all x);
      // the forests (disjoint sets of cells, everything else 0xFFFFFFFF = -1) are
      // merged with one MAX all-reduce; the links between bands -- y faces of the
      // brick rows at a band edge only -- are then added by everybody.  Roots are
      // always the smallest cell index of their tree, so all ranks end up with
      // identical forests.
      const int band = (byn + sh.world - 1) / sh.world;
      AMCL_CUDA(cudaMemsetAsync(T.aux, 0xFF, sizeof(unsigned int) * (size_t)nslots, st));
      AMCL_CUDA(cudaMemsetAsync(cmin, 0xFF, sizeof(unsigned int) * (size_t)nslots, st));
      cc_brick_local<<<bricks, kBrickCells, 0, st>>>(T, bxn, byn, n, cmin, band, sh.rank);
      AMCL_LAUNCH_CHECK();
      cc_brick_borders<<<bricks, kBrickCells, 0, st>>>(T, bxn, byn, n, band, sh.rank, 1);
      AMCL_LAUNCH_CHECK();
      sh.allreduce(T.aux, (int)nslots, SHARD_I32, SHARD_MAX, sh.user);
      cc_brick_borders<<<bricks, kBrickCells, 0, st>>>(T, bxn, byn, n, band, sh.rank, 2);
      AMCL_LAUNCH_CHECK();
    }
    else
    {
      cc_brick_local<<<bricks, kBrickCells, 0, st>>>(T, bxn, byn, n, cmin, 0, 0);
      AMCL_LAUNCH_CHECK();
      cc_brick_borders<<<bricks, kBrickCells, 0, st>>>(T, bxn, byn, n, 0, 0, 0);
      AMCL_LAUNCH_CHECK();
    }
  }
  else
  {
    cc_init<<<sb, 256, 0, st>>>(T, nslots, n, cmin);
    AMCL_LAUNCH_CHECK();
    cc_link<<<sb, 256, 0, st>>>(T, nslots, n);
    AMCL_LAUNCH_CHECK();
  }
  cc_flatten<<<sb, 256, 0, st>>>(T, nslots, n);
  AMCL_LAUNCH_CHECK();
  cc_component_min<<<sb, 256, 0, st>>>(T, nslots, n, cmin);
  AMCL_LAUNCH_CHECK();
  // canonical ids: short root list when there are few clusters ...
  if (!c->d_rootlist) AMCL_CUDA(cudaMalloc(&c->d_rootlist, sizeof(unsigned int) * 2 * kRootListCap));
  AMCL_CUDA(cudaMemsetAsync(c->d_ints + I_CLUSTERS, 0, sizeof(int), st));
  cc_collect_roots<<<sb, 256, 0, st>>>(T, nslots, n, cmin, c->d_rootlist, c->d_rootlist + kRootListCap,
                                       c->d_ints + I_CLUSTERS);
  AMCL_LAUNCH_CHECK();
  sync_ints(c, I_CLUSTERS, 1);
  int ncl = c->h_ints[I_CLUSTERS];
  if (ncl <= kRootListCap)
  {
    if (ncl > 0)
    {
      cc_rank_roots<<<blocks_for(ncl, 256), 256, 0, st>>>(c->d_rootlist, c->d_rootlist + kRootListCap, ncl, cmin);
      AMCL_LAUNCH_CHECK();
    }
  }
  else
  {
    // ... otherwise flag every cluster's earliest sample and scan the flags.
    // The flags reuse d_keys' storage as u32 scratch (keys are spent).
    unsigned int *cflag = reinterpret_cast<unsigned int *>(c->d_keys);
    unsigned int *cscan = cflag + c->alloc_samples;
    AMCL_CUDA(cudaMemsetAsync(cflag, 0, sizeof(unsigned int) * (size_t)(n > 0 ? n : 1), st));
    cc_flag_roots<<<sb, 256, 0, st>>>(T, nslots, n, cmin, cflag);
    AMCL_LAUNCH_CHECK();
    scan_u32(c, cflag, cscan, n, I_CLUSTERS);
    cc_assign_ids<<<sb, 256, 0, st>>>(T, nslots, n, cscan, cmin);
    AMCL_LAUNCH_CHECK();
    sync_ints(c, I_CLUSTERS, 1);
    ncl = c->h_ints[I_CLUSTERS];
  }

  if (ncl > c->cluster_acc_cap)
  {
    if (c->d_cluster_acc) cudaFree(c->d_cluster_acc);
    c->cluster_acc_cap = ncl + ncl / 2 + 64;
    AMCL_CUDA(cudaMalloc(&c->d_cluster_acc, sizeof(double) * 10 * (size_t)c->cluster_acc_cap));
  }
  AMCL_CUDA(cudaMemsetAsync(c->d_cluster_acc, 0, sizeof(double) * 10 * (size_t)(ncl > 0 ? ncl : 1), st));
  {
    int grid = globals().sm_count * 8;
    const int need = blocks_for(acc_cnt, 256);
    if (grid > need) grid = need;
    cluster_accumulate<<<grid, 256, 0, st>>>(c->d_set[which], acc_cnt, c->d_tab_aux, c->d_slot, cmin,
                                             c->d_cluster_acc, c->d_label, acc_lo);
  }
  AMCL_LAUNCH_CHECK();
  if (sliced && ncl > 0)
    c->shard.allreduce(c->d_cluster_acc, 10 * ncl, SHARD_F64, SHARD_SUM, c->shard.user);
  std::vector<double> acc((size_t)10 * (ncl > 0 ? ncl : 1));
  AMCL_CUDA(cudaMemcpyAsync(acc.data(), c->d_cluster_acc, sizeof(double) * 10 * (size_t)ncl,
                            cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaStreamSynchronize(st));
  publish_clusters(set, acc.data(), ncl);
}

// pf.c:576-693 on the host: the per-cluster sums [count, w, w x, w y, w cos, w sin,
// w xx, w xy, w yx, w yy] become pf_cluster_t entries and the set's mean / covariance.
static void publish_clusters(pf_sample_set_t *set, const double *acc, int ncl)
{
  // pf.c:576-589 zeroes every slot; here only the ones that were in use
  const int prev = set->cluster_count;
  int limit = ncl < set->cluster_max_count ? ncl : set->cluster_max_count;
  for (int k = limit; k < prev && k < set->cluster_max_count; k++)
    memset(set->clusters + k, 0, sizeof(pf_cluster_t));
  set->cluster_count = limit;

  double m[4] = {0, 0, 0, 0}, cc[2][2] = {{0, 0}, {0, 0}}, weight = 0.0;
  for (int k = 0; k < ncl; k++)
  {
    const double *a = acc + (size_t)10 * k;
    weight += a[1];
    for (int q = 0; q < 4; q++) m[q] += a[2 + q];
    cc[0][0] += a[6]; cc[0][1] += a[7]; cc[1][0] += a[8]; cc[1][1] += a[9];
    if (k >= limit) continue;   // pf.c:609-610
    pf_cluster_t *cl = set->clusters + k;
    cl->count = (int)a[0];
    cl->weight = a[1];
    for (int q = 0; q < 4; q++) cl->m[q] = a[2 + q];
    cl->c[0][0] = a[6]; cl->c[0][1] = a[7]; cl->c[1][0] = a[8]; cl->c[1][1] = a[9];
    // pf.c:646-666
    cl->mean.v[0] = cl->m[0] / cl->weight;
    cl->mean.v[1] = cl->m[1] / cl->weight;
    cl->mean.v[2] = atan2(cl->m[3], cl->m[2]);
    cl->cov = pf_matrix_zero();
    for (int j = 0; j < 2; j++)
      for (int q = 0; q < 2; q++)
        cl->cov.m[j][q] = cl->c[j][q] / cl->weight - cl->mean.v[j] * cl->mean.v[q];
    cl->cov.m[2][2] = -2 * log(sqrt(cl->m[2] * cl->m[2] + cl->m[3] * cl->m[3]) / cl->weight);
  }
  set->mean = pf_vector_zero();
  set->cov = pf_matrix_zero();
  if (fabs(weight) < DBL_EPSILON)
  {
    printf("ERROR : divide-by-zero exception : weight is zero\n");   // pf.c:672-677
    return;
  }
  // pf.c:678-690
  set->mean.v[0] = m[0] / weight;
  set->mean.v[1] = m[1] / weight;
  set->mean.v[2] = atan2(m[3], m[2]);
  for (int j = 0; j < 2; j++)
    for (int q = 0; q < 2; q++)
      set->cov.m[j][q] = cc[j][q] / weight - set->mean.v[j] * set->mean.v[q];
  set->cov.m[2][2] = -2 * log(sqrt(m[2] * m[2] + m[3] * m[3]));
}

void run_cluster_stats(PfCtx *c, int which, bool histogram_ready)
{
  if (!histogram_ready) pf_replicate(c, which);   // keys are needed for every sample
  cudaStream_t st = globals().stream;
  pf_sample_set_t *set = c->host->sets + which;
  const int n = set->sample_count;
  if (!histogram_ready)
  {
    double bx, by, bt;
    bin_sizes(set, &bx, &by, &bt);
    keys_of_set<<<blocks_for(n, 256), 256, 0, st>>>(c->d_set[which], n, c->d_keys, bx, by, bt);
    AMCL_LAUNCH_CHECK();
    build_histogram(c, n);
    scan_u32(c, c->d_flag, c->d_scan, n, I_DISTINCT);
    sync_ints(c, I_DISTINCT, 1);
    set->kdtree->leaf_count = c->h_ints[I_DISTINCT];
  }
  cluster_and_stats(c, which, n);
}

__global__ void divide_pair(double *v, int n)
{
  v[0] = v[0] / n;
  v[1] = v[1] / n;
}

int run_update_converged(PfCtx *c, int which)
{
  cudaStream_t st = globals().stream;
  pf_t *pf = c->host;
  pf_sample_set_t *set = pf->sets + which;
  const int n = set->sample_count;
  const bool sliced = c->shard.on() && !c->replicated[which];
  const int lo = sliced ? c->shard.lo(n) : 0;
  const int cnt = sliced ? c->shard.count(n) : n;
  const Particle *base = c->d_set[which] + lo;
  const int grid = reduce_grid(cnt);
  ensure_partials(c, 2 * grid);
  sum_xy<<<grid, 256, 0, st>>>(base, cnt, c->d_partials, c->d_partials + grid);
  AMCL_LAUNCH_CHECK();
  if (!sliced)
    finish_xy<<<1, 256, 0, st>>>(c->d_partials, c->d_partials + grid, grid, n, c->d_scalars + 4);
  else
  {
    finish_xy<<<1, 256, 0, st>>>(c->d_partials, c->d_partials + grid, grid, 1, c->d_scalars + 4);
    AMCL_LAUNCH_CHECK();
    c->shard.allreduce(c->d_scalars + 4, 2, SHARD_F64, SHARD_SUM, c->shard.user);
    divide_pair<<<1, 1, 0, st>>>(c->d_scalars + 4, n);
  }
  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaMemsetAsync(c->d_ints + I_OUTSIDE, 0, sizeof(int), st));
  any_outside<<<grid, 256, 0, st>>>(base, cnt, c->d_scalars + 4, pf->dist_threshold,
                                    c->d_ints + I_OUTSIDE);
  AMCL_LAUNCH_CHECK();
  if (sliced) c->shard.allreduce(c->d_ints + I_OUTSIDE, 1, SHARD_I32, SHARD_MAX, c->shard.user);
  sync_ints(c, I_OUTSIDE, 1);
  const int conv = c->h_ints[I_OUTSIDE] ? 0 : 1;
  set->converged = conv;
  pf->converged = conv;
  return conv;
}

void run_cep_stats(PfCtx *c, int which, double out[4])
{
  pf_replicate(c, which);
  cudaStream_t st = globals().stream;
  pf_sample_set_t *set = c->host->sets + which;
  const int n = set->sample_count;
  const int grid = reduce_grid(n);
  ensure_partials(c, 4 * grid);
  cep_sums<<<grid, 256, 0, st>>>(c->d_set[which], n, c->d_partials, grid);
  AMCL_LAUNCH_CHECK();
  finish_multi<<<1, 256, 0, st>>>(c->d_partials, grid, grid, 4, c->d_scalars + 8);
  AMCL_LAUNCH_CHECK();
  sync_scalars(c, 8, 4);
  for (int q = 0; q < 4; q++) out[q] = c->h_scalars[8 + q];   // mn, mx, my, mrr
}

// copy_set (pf.c:321-360): b = a with weights renormalised by their sum.
void run_copy_set(PfCtx *c, int from, int to)
{
  pf_replicate(c, from);
  c->replicated[to] = true;
  cudaStream_t st = globals().stream;
  pf_t *pf = c->host;
  const int n = pf->sets[from].sample_count;
  const int grid = reduce_grid(n);
  ensure_partials(c, grid);
  sum_weights<<<grid, 256, 0, st>>>(c->d_set[from], n, c->d_partials);
  AMCL_LAUNCH_CHECK();
  finish_multi<<<1, 256, 0, st>>>(c->d_partials, grid, grid, 1, c->d_scalars + 12);
  AMCL_LAUNCH_CHECK();
  copy_scaled<<<blocks_for(n, 256), 256, 0, st>>>(c->d_set[from], c->d_set[to], n,
                                                  c->d_scalars + 12);
  AMCL_LAUNCH_CHECK();
  pf->sets[to].sample_count = n;
  pf->sets[to].converged = pf->sets[from].converged;
  c->host_fresh[to] = false;
}

// Everything the small tail puts on the stream, and nothing else (no allocation,
// no wait): this is what is captured into a CUDA graph.
static void small_tail_enqueue(PfCtx *c, int ib, int maxs)
{
  cudaStream_t st = globals().stream;
  pf_t *pf = c->host;
  AMCL_CUDA(cudaMemsetAsync(c->d_tab_key, 0xFF, sizeof(unsigned long long) * (size_t)c->hash_cap, st));
  AMCL_CUDA(cudaMemsetAsync(c->d_tab_first, 0xFF, sizeof(unsigned int) * (size_t)c->hash_cap, st));
  BinTable T = table_of(c);
  bin_insert<<<blocks_for(maxs, 256), 256, 0, st>>>(c->d_keys, 0, maxs, T, c->d_slot);
  AMCL_LAUNCH_CHECK();
  mark_first<<<blocks_for(maxs, 256), 256, 0, st>>>(c->d_tab_first, c->d_slot, 0, maxs, c->d_flag);
  AMCL_LAUNCH_CHECK();
  scan_u32(c, c->d_flag, c->d_scan, maxs, I_DISTINCT);
  unsigned long long *d_cutk = reinterpret_cast<unsigned long long *>(c->d_ints + I_CUTK);
  AMCL_CUDA(cudaMemsetAsync(d_cutk, 0x7F, sizeof(unsigned long long), st));
  find_kld_cut<<<blocks_for(maxs, 256), 256, 0, st>>>(c->d_scan, 0, maxs, nullptr, 0, pf->min_samples, maxs,
                                                      pf->pop_err, pf->pop_z, d_cutk);
  AMCL_LAUNCH_CHECK();
  set_uniform_weight_dev<<<blocks_for(maxs, 256), 256, 0, st>>>(c->d_set[ib], maxs, d_cutk);
  AMCL_LAUNCH_CHECK();
  // labels of the occupied bins among the first n samples
  T.n_dev = d_cutk;
  unsigned int *cmin = c->d_scan + c->alloc_samples;
  const unsigned int nslots = c->hash_cap;
  const int sb = (int)((nslots + 255u) / 256u);
  cc_init<<<sb, 256, 0, st>>>(T, nslots, 0, cmin);
  AMCL_LAUNCH_CHECK();
  cc_link<<<sb, 256, 0, st>>>(T, nslots, 0);
  AMCL_LAUNCH_CHECK();
  cc_flatten<<<sb, 256, 0, st>>>(T, nslots, 0);
  AMCL_LAUNCH_CHECK();
  cc_component_min<<<sb, 256, 0, st>>>(T, nslots, 0, cmin);
  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaMemsetAsync(c->d_ints + I_CLUSTERS, 0, sizeof(int), st));
  cc_collect_roots<<<sb, 256, 0, st>>>(T, nslots, 0, cmin, c->d_rootlist, c->d_rootlist + kRootListCap,
                                       c->d_ints + I_CLUSTERS);
  AMCL_LAUNCH_CHECK();
  cc_rank_roots<<<kRootListCap / 256, 256, 0, st>>>(c->d_rootlist, c->d_rootlist + kRootListCap, 0, cmin,
                                                    c->d_ints + I_CLUSTERS);
  AMCL_LAUNCH_CHECK();
  // Should there be more clusters than the short list ranks, a root keeps its
  // earliest sample index (< maxs) as label: room for that many rows keeps the
  // sums in bounds; they are thrown away and redone the long way below.
  const int acc_rows = maxs + 1;
  const int zero_rows = acc_rows < kRootListCap ? acc_rows : kRootListCap;
  AMCL_CUDA(cudaMemsetAsync(c->d_cluster_acc, 0, sizeof(double) * 10 * (size_t)zero_rows, st));
  {
    int grid = globals().sm_count * 8;
    if (grid > blocks_for(maxs, 256)) grid = blocks_for(maxs, 256);
    cluster_accumulate<<<grid, 256, 0, st>>>(c->d_set[ib], maxs, c->d_tab_aux, c->d_slot, cmin,
                                             c->d_cluster_acc, c->d_label, 0, d_cutk);
    AMCL_LAUNCH_CHECK();
  }
  converged_small<<<1, 1024, 0, st>>>(c->d_set[ib], d_cutk, pf->dist_threshold, c->d_ints + I_OUTSIDE);
  AMCL_LAUNCH_CHECK();
  // the one trip: ints 0..7 (cluster count, outside flag, the (n, k) pair) and
  // the sums of the first clusters
  const int head_rows = kSmallHeadPinned < zero_rows ? kSmallHeadPinned : zero_rows;
  AMCL_CUDA(cudaMemcpyAsync(c->h_ints, c->d_ints, sizeof(int) * 8, cudaMemcpyDeviceToHost, st));
  AMCL_CUDA(cudaMemcpyAsync(c->h_scalars + 16, c->d_cluster_acc, sizeof(double) * 10 * (size_t)head_rows,
                            cudaMemcpyDeviceToHost, st));
}

// The tail of pf_update_resample for a small, unsharded filter: histogram, cut,
// weights, labels, statistics and convergence without the host in between.
static void resample_small_tail(PfCtx *c, int ia, int ib, int maxs, bool host_stream, double w_diff,
                                unsigned short saved_state[3], Particle *d_injected, int *d_inject_slot,
                                Trace &tr)
{
  (void)ia;
  cudaStream_t st = globals().stream;
  pf_t *pf = c->host;
  pf_sample_set_t *set_b = pf->sets + ib;
  c->tab_dense = 0;        // the histogram of this path is always the hash
  // buffers the chain needs, made before anything is captured
  if (!c->d_rootlist) AMCL_CUDA(cudaMalloc(&c->d_rootlist, sizeof(unsigned int) * 2 * kRootListCap));
  const int acc_rows = maxs + 1;
  if (acc_rows > c->cluster_acc_cap)
  {
    if (c->d_cluster_acc) cudaFree(c->d_cluster_acc);
    c->cluster_acc_cap = acc_rows + 64;
    AMCL_CUDA(cudaMalloc(&c->d_cluster_acc, sizeof(double) * 10 * (size_t)c->cluster_acc_cap));
  }
  const int zero_rows = acc_rows < kRootListCap ? acc_rows : kRootListCap;
  const int head_rows = kSmallHeadPinned < zero_rows ? kSmallHeadPinned : zero_rows;
  // The chain is ~20 tiny launches with fixed arguments: replayed as ONE graph
  // launch.  It is captured again whenever something it has baked in changes
  // (the caller may poke pop_err / pop_z / dist_threshold between updates).
  SmallGraph &g = c->small_graph[ib];
  const double key[8] = {(double)maxs, (double)pf->min_samples, pf->pop_err, pf->pop_z, pf->dist_threshold,
                         (double)c->cluster_acc_cap, (double)(size_t)c->d_cluster_acc,
                         (double)set_b->kdtree->size[0] + 1e3 * set_b->kdtree->size[1] + 1e6 * set_b->kdtree->size[2]};
  static const bool use_graph = !(getenv("AMCLB200_NO_GRAPH") && getenv("AMCLB200_NO_GRAPH")[0] == '1');
  if (use_graph && !amcl_debug_sync())
  {
    if (!g.exec || memcmp(key, g.key, sizeof(key)) != 0)
    {
      if (g.exec) cudaGraphExecDestroy(g.exec);
      g.exec = nullptr;
      if (!c->capture_stream) AMCL_CUDA(cudaStreamCreateWithFlags(&c->capture_stream, cudaStreamNonBlocking));
      cudaStream_t user_stream = globals().stream;
      const unsigned long long launches0 = amcl_launch_counter();
      globals().stream = c->capture_stream;
      AMCL_CUDA(cudaStreamBeginCapture(c->capture_stream, cudaStreamCaptureModeThreadLocal));
      small_tail_enqueue(c, ib, maxs);
      cudaGraph_t graph = nullptr;
      AMCL_CUDA(cudaStreamEndCapture(c->capture_stream, &graph));
      globals().stream = user_stream;
      AMCL_CUDA(cudaGraphInstantiate(&g.exec, graph, 0));
      cudaGraphDestroy(graph);
      g.launches = (int)(amcl_launch_counter() - launches0);
      amcl_launch_counter() = launches0;
      memcpy(g.key, key, sizeof(key));
    }
    AMCL_CUDA(cudaGraphLaunch(g.exec, st));
    amcl_launch_counter() += (unsigned long long)g.launches;
  }
  else
    small_tail_enqueue(c, ib, maxs);
  AMCL_CUDA(cudaStreamSynchronize(st));
  const unsigned long long cutk = *reinterpret_cast<unsigned long long *>(c->h_ints + I_CUTK);
  const int n = (int)(cutk >> 32);
  set_b->sample_count = n;                                      // pf.c:483
  set_b->kdtree->leaf_count = (int)(cutk & 0xFFFFFFFFull);      // k_{n-1}
  tr.mark("one trip");
  if (host_stream)
  {
    // leave libc's stream where the reference would have left it
    seed48(saved_state);
    for (int m = 0; m < n; m++)
    {
      if (drand48() < w_diff) (void)(pf->random_pose_fn)(pf->random_pose_data);
      else (void)drand48();
    }
  }
  c->draws_used = 2ull * (unsigned long long)n;
  if (w_diff > 0.0) pf->w_slow = pf->w_fast = 0.0;     // pf.c:487-489
  c->host_fresh[ib] = false;
  c->replicated[ib] = true;
  if (d_injected)
  {
    cudaFree(d_injected);
    cudaFree(d_inject_slot);
  }
  const int ncl = c->h_ints[I_CLUSTERS];
  if (ncl > kRootListCap)
    cluster_and_stats(c, ib, n);                       // rare: the long way (flag + scan ranking)
  else if (ncl <= head_rows)
    publish_clusters(set_b, c->h_scalars + 16, ncl);
  else
  {
    static std::vector<double> all;
    all.resize((size_t)10 * ncl);
    AMCL_CUDA(cudaMemcpyAsync(all.data(), c->d_cluster_acc, sizeof(double) * 10 * (size_t)ncl,
                              cudaMemcpyDeviceToHost, st));
    AMCL_CUDA(cudaStreamSynchronize(st));
    publish_clusters(set_b, all.data(), ncl);
  }
  pf->current_set = ib;
  const int conv = c->h_ints[I_OUTSIDE] ? 0 : 1;
  set_b->converged = conv;
  pf->converged = conv;
  tr.mark("clusters");
  tr.report("pf_update_resample (single trip)");
}

// pf_update_resample (pf.c:363-510)
void run_resample(PfCtx *c)
{
  cudaStream_t st = globals().stream;
  pf_t *pf = c->host;
  const int ia = pf->current_set, ib = (pf->current_set + 1) % 2;
  pf_sample_set_t *set_a = pf->sets + ia, *set_b = pf->sets + ib;
  const int n_a = set_a->sample_count;
  const int maxs = pf->max_samples;
  const Shard &sh = c->shard;

  // Draws range over the GLOBAL table.  A sharded run
  // needs every weight for the table and any particle's pose for the copy:
  //  * peer-memory path: all-gather the 8-byte weights only, and let the pick
  //    kernel read the chosen poses from their owners' memory over NVLink
  //    (2 x 32 B per output sample instead of shipping the whole set);
  //  * otherwise: all-gather the whole source set in place.
  Trace tr;
  tr.mark("start");
  const bool selective_copy =
      pf->selective_resampling != 0 && set_a->n_effective > 0.5 * (set_a->sample_count);
  const bool p2p = sh.on() && c->peers_ready && !c->replicated[ia] && !selective_copy;
  if (!p2p) pf_replicate(c, ia);

  // pf.c:378-394
  if (pf->selective_resampling != 0 && set_a->n_effective > 0.5 * (set_a->sample_count))
  {
    run_copy_set(c, ia, ib);
    run_cluster_stats(c, ib, false);
    pf->current_set = ib;
    return;
  }

  if (p2p && exact_cumsum_sharded(c, c->d_set[ia], n_a, c->d_cum))
  {
    // every rank built its slice of the table from its own weights; the all-gather
    // of the slices is also the rendezvous that makes the source set final
  }
  else if (p2p)
  {
    const int lo = sh.lo(n_a), cnt = sh.count(n_a);
    if (cnt > 0)
    {
      pack_weights<<<blocks_for(cnt, 256), 256, 0, st>>>(c->d_set[ia] + lo, cnt, c->d_wpack + lo);
      AMCL_LAUNCH_CHECK();
    }
    // also the rendezvous that makes every rank's slice final before anyone reads it
    sh.allgather(c->d_wpack, (unsigned long long)sh.chunk(n_a) * sizeof(double), sh.user);
    Weights W;
    W.p = c->d_wpack;
    W.stride = 1;
    exact_cumsum_w(c, W, n_a, c->d_cum);
  }
  else
    exact_cumsum(c, c->d_set[ia], n_a, c->d_cum);

  tr.mark("table");
  // pf.c:411-413
  double w_diff = 1.0 - pf->w_fast / pf->w_slow;
  if (w_diff < 0.0) w_diff = 0.0;

  // ---- draws for every candidate sample -----------------------------------
  std::vector<int> inject_slot;
  std::vector<Particle> injected;
  unsigned short saved_state[3] = {0, 0, 0};
  const bool host_stream = globals().rng == RNG_DRAND48;
  uint64_t philox_key = 0;
  bool draws_in_kernel = false;
  if (host_stream)
  {
    // Draw exactly as pf.c:426-460 would for a full run to max_samples, then
    // rewind and re-consume only what the first n samples used (below).
    unsigned short probe[3] = {0, 0, 0};
    unsigned short *old = seed48(probe);
    memcpy(saved_state, old, sizeof(saved_state));
    seed48(saved_state);
    std::vector<double> draws((size_t)2 * maxs);
    for (int m = 0; m < maxs; m++)
    {
      const double u = drand48();
      draws[2 * (size_t)m] = u;
      if (u < w_diff)
      {
        if (inject_slot.empty()) inject_slot.assign(maxs, -1);
        const pf_vector_t p = (pf->random_pose_fn)(pf->random_pose_data);
        inject_slot[m] = (int)injected.size();
        Particle q; q.x = p.v[0]; q.y = p.v[1]; q.th = p.v[2]; q.w = 1.0;
        injected.push_back(q);
        draws[2 * (size_t)m + 1] = 0.0;
      }
      else
        draws[2 * (size_t)m + 1] = drand48();
    }
    AMCL_CUDA(cudaMemcpyAsync(c->d_draws, draws.data(), sizeof(double) * 2 * (size_t)maxs,
                              cudaMemcpyHostToDevice, st));
    AMCL_CUDA(cudaStreamSynchronize(st));
  }
  else
  {
    philox_key = globals().philox_seed ^ (0xA0761D6478BD642Full * (++c->step));
    // without recovery injection nobody but the pick kernel needs the draws: it
    // regenerates them from (key, sample index) instead of reading 16 B per sample
    draws_in_kernel = !(w_diff > 0.0);
    if (w_diff > 0.0)
    {
      philox_resample_draws<<<blocks_for(maxs, 256), 256, 0, st>>>(c->d_draws, maxs, philox_key);
      AMCL_LAUNCH_CHECK();
      // recovery path: candidates with u < w_diff take a host-generated pose
      std::vector<double> draws((size_t)2 * maxs);
      AMCL_CUDA(cudaMemcpyAsync(draws.data(), c->d_draws, sizeof(double) * 2 * (size_t)maxs,
                                cudaMemcpyDeviceToHost, st));
      AMCL_CUDA(cudaStreamSynchronize(st));
      for (int m = 0; m < maxs; m++)
        if (draws[2 * (size_t)m] < w_diff)
        {
          if (inject_slot.empty()) inject_slot.assign(maxs, -1);
          const pf_vector_t p = (pf->random_pose_fn)(pf->random_pose_data);
          inject_slot[m] = (int)injected.size();
          Particle q; q.x = p.v[0]; q.y = p.v[1]; q.th = p.v[2]; q.w = 1.0;
          injected.push_back(q);
        }
    }
  }

  // Sharded: what the HOST contributed must be the same on every rank -- the
  // caller's random_pose_fn draws from each process's own libc stream, and in
  // drand48 mode so do the uniforms.  Rank 0's version wins: the others zero
  // their copy and a SUM all-reduce hands it round (x + 0 is exact).
  int n_injected = (int)injected.size();
  if (sh.on() && (host_stream || w_diff > 0.0))
  {
    c->h_ints[I_MOVED] = sh.rank == 0 ? n_injected : 0;
    AMCL_CUDA(cudaMemcpyAsync(c->d_ints + I_MOVED, c->h_ints + I_MOVED, sizeof(int), cudaMemcpyHostToDevice, st));
    sh.allreduce(c->d_ints + I_MOVED, 1, SHARD_I32, SHARD_SUM, sh.user);
    sync_ints(c, I_MOVED, 1);
    n_injected = c->h_ints[I_MOVED];
    if (host_stream)
    {
      if (sh.rank != 0) AMCL_CUDA(cudaMemsetAsync(c->d_draws, 0, sizeof(double) * 2 * (size_t)maxs, st));
      for (size_t off = 0; off < 2 * (size_t)maxs; off += (size_t)1 << 28)   // int-sized pieces
      {
        const size_t len = 2 * (size_t)maxs - off < ((size_t)1 << 28) ? 2 * (size_t)maxs - off : (size_t)1 << 28;
        sh.allreduce(c->d_draws + off, (int)len, SHARD_F64, SHARD_SUM, sh.user);
      }
    }
  }
  Particle *d_injected = nullptr;
  int *d_inject_slot = nullptr;
  if (n_injected > 0)
  {
    const bool mine = !sh.on() || sh.rank == 0;
    if (inject_slot.empty()) inject_slot.assign(maxs, -1);
    injected.resize((size_t)n_injected);
    AMCL_CUDA(cudaMalloc(&d_injected, sizeof(Particle) * (size_t)n_injected));
    AMCL_CUDA(cudaMalloc(&d_inject_slot, sizeof(int) * (size_t)maxs));
    if (mine)
    {
      // slots travel as slot + 1 so that "none" is 0, the neutral element of the sum
      for (int m = 0; m < maxs; m++) inject_slot[m] += 1;
      AMCL_CUDA(cudaMemcpyAsync(d_injected, injected.data(), sizeof(Particle) * (size_t)n_injected,
                                cudaMemcpyHostToDevice, st));
      AMCL_CUDA(cudaMemcpyAsync(d_inject_slot, inject_slot.data(), sizeof(int) * (size_t)maxs,
                                cudaMemcpyHostToDevice, st));
    }
    else
    {
      AMCL_CUDA(cudaMemsetAsync(d_injected, 0, sizeof(Particle) * (size_t)n_injected, st));
      AMCL_CUDA(cudaMemsetAsync(d_inject_slot, 0, sizeof(int) * (size_t)maxs, st));
    }
    if (sh.on())
    {
      sh.allreduce(d_injected, 4 * n_injected, SHARD_F64, SHARD_SUM, sh.user);
      sh.allreduce(d_inject_slot, maxs, SHARD_I32, SHARD_SUM, sh.user);
    }
  }

  // ---- pick sources, copy poses, bin keys -----------------------------------
  PickArgs pa;
  pa.src = c->d_set[ia];
  pa.peers = p2p ? c->d_peer_tab[ia] : nullptr;
  pa.peer_chunk = sh.chunk(n_a) > 0 ? sh.chunk(n_a) : 1;
  pa.n_src = n_a;
  pa.cum = c->d_cum;
  pa.coarse = nullptr;
  pa.n_coarse = n_a / kCoarse;
  if (pa.n_coarse >= 64)
  {
    // the draws array has room behind the 2*maxs uniforms? no: use the partial-sum scratch
    ensure_partials(c, pa.n_coarse);
    sample_table<<<blocks_for(pa.n_coarse, 256), 256, 0, st>>>(c->d_cum, pa.n_coarse, c->d_partials);
    AMCL_LAUNCH_CHECK();
    pa.coarse = c->d_partials;
  }
  pa.bucket_index = nullptr;
  pa.buckets = 0;
  if (n_a >= (1 << 16))
  {
    int buckets = 1;
    while (buckets * 16 < n_a) buckets <<= 1;     // about 16 table entries per bucket
    if ((size_t)buckets + 1 > c->bucket_cap)
    {
      if (c->d_bucket_index) cudaFree(c->d_bucket_index);
      c->bucket_cap = (size_t)buckets + 1;
      AMCL_CUDA(cudaMalloc(&c->d_bucket_index, sizeof(int) * c->bucket_cap));
    }
    build_bucket_index<<<blocks_for(buckets + 1, 256), 256, 0, st>>>(pa.coarse, pa.n_coarse, c->d_cum, n_a,
                                                                     buckets, c->d_bucket_index);
    AMCL_LAUNCH_CHECK();
    pa.bucket_index = c->d_bucket_index;
    pa.buckets = buckets;
  }
  pa.draws = draws_in_kernel ? nullptr : c->d_draws;
  pa.philox_key = philox_key;
  pa.injected = d_injected;
  pa.inject_slot = d_inject_slot;
  pa.dst = c->d_set[ib];
  pa.src_index = c->d_src;
  pa.keys = c->d_keys;
  pa.write_keys = 1;
  pa.weight = 1.0 / (double)maxs;
  // a sharded rank picks only its slice of the candidate samples ...
  pa.first = sh.lo(maxs);
  pa.count = sh.count(maxs);
  bin_sizes(set_b, &pa.bin_x, &pa.bin_y, &pa.bin_t);
  pick_and_copy<<<blocks_for(pa.count, 256), 256, 0, st>>>(pa);
  AMCL_LAUNCH_CHECK();
  tr.mark("pick+keys");

  // ---- node-sized filters: everything below in ONE trip to the host -----------
  // (amcl_node runs 500 .. 10 000 particles; at that size the update is launch
  // and synchronisation latency, not work.)  The histogram is always the hash --
  // no bounding box to fetch --, the cut (n, k) stays on the device and the
  // kernels after it read n from there; the host gets n, k, the cluster count,
  // the convergence flag and the first clusters' sums in a single copy.
  if (!sh.on() && maxs <= kSmallFilter)
  {
    resample_small_tail(c, ia, ib, maxs, host_stream, w_diff, saved_state, d_injected, d_inject_slot, tr);
    return;
  }

  // ---- histogram, k_m, KLD cut ---------------------------------------------
  // The KLD count is a property of the global sample ORDER: k_m = number of
  // distinct bins among samples 0..m.
  int n = maxs;
  unsigned int k_at_cut = 0;
  bool shared_grid = false;
  if (sh.on() && sh.world <= 32)
  {
    // Sharded, preferred way: agree on the key bounding box; if the dense grid
    // fits, every rank enters only ITS samples and the grids are merged with one
    // MIN all-reduce (23 MB for a 200 m map) instead of shipping 8 B per sample.
    const int lo_c = sh.lo(maxs), cnt_c = sh.count(maxs);
    int *h = c->h_ints + I_BOUNDS;
    for (int q = 0; q < 3; q++) { h[q] = 0x7FFFFFFF; h[3 + q] = -0x7FFFFFFF; }
    AMCL_CUDA(cudaMemcpyAsync(c->d_ints + I_BOUNDS, h, sizeof(int) * 6, cudaMemcpyHostToDevice, st));
    if (cnt_c > 0)
    {
      int grid = globals().sm_count * 8;
      if (grid > blocks_for(cnt_c, 256)) grid = blocks_for(cnt_c, 256);
      key_bounds<<<grid, 256, 0, st>>>(c->d_keys + lo_c, cnt_c, c->d_ints + I_BOUNDS);
      AMCL_LAUNCH_CHECK();
    }
    negate_ints<<<1, 32, 0, st>>>(c->d_ints + I_BOUNDS + 3, 3);      // max -> -max: one MIN serves all six
    AMCL_LAUNCH_CHECK();
    sh.allreduce(c->d_ints + I_BOUNDS, 6, SHARD_I32, SHARD_MIN, sh.user);
    sync_ints(c, I_BOUNDS, 6);
    double cells = 1.0;
    for (int q = 0; q < 3; q++)
    {
      c->tab_lo[q] = h[q];
      c->tab_ext[q] = -h[3 + q] - h[q] + 1;
      cells *= (double)c->tab_ext[q];
    }
    shared_grid = cells <= (double)c->hash_cap;
    if (shared_grid)
    {
      c->tab_dense = 1;
      const BinTable T = table_of(c);
      // 0x7F7F7F7F = empty: larger than any sample index, and positive as the
      // int32 the all-reduce works on
      AMCL_CUDA(cudaMemsetAsync(c->d_tab_first, 0x7F, sizeof(unsigned int) * (size_t)cells, st));
      if (cnt_c > 0)
      {
        bin_insert<<<blocks_for(cnt_c, 256), 256, 0, st>>>(c->d_keys, lo_c, cnt_c, T, c->d_slot);
        AMCL_LAUNCH_CHECK();
      }
      sh.allreduce(c->d_tab_first, (int)cells, SHARD_I32, SHARD_MIN, sh.user);
      tr.mark("histogram");
      // own flags, own inclusive scan, then the counts of the lower ranks
      int *h_cnt = c->h_ints + I_RANKCNT;
      for (int r = 0; r < sh.world; r++) h_cnt[r] = 0;
      AMCL_CUDA(cudaMemcpyAsync(c->d_ints + I_RANKCNT, h_cnt, sizeof(int) * sh.world,
                                cudaMemcpyHostToDevice, st));
      if (cnt_c > 0)
      {
        mark_first<<<blocks_for(cnt_c, 256), 256, 0, st>>>(c->d_tab_first, c->d_slot, lo_c, cnt_c,
                                                           c->d_flag);
        AMCL_LAUNCH_CHECK();
        scan_u32(c, c->d_flag + lo_c, c->d_scan + lo_c, cnt_c, I_RANKCNT + sh.rank);
      }
      sh.allreduce(c->d_ints + I_RANKCNT, sh.world, SHARD_I32, SHARD_SUM, sh.user);
      // (n, leaf_count) in one 64-bit minimum over all ranks, one trip to the host
      unsigned long long *d_cutk = reinterpret_cast<unsigned long long *>(c->d_ints + I_CUTK);
      AMCL_CUDA(cudaMemsetAsync(d_cutk, 0x7F, sizeof(unsigned long long), st));   // positive as an int64
      if (cnt_c > 0)
      {
        find_kld_cut<<<blocks_for(cnt_c, 256), 256, 0, st>>>(c->d_scan, lo_c, cnt_c, c->d_ints + I_RANKCNT,
                                                             sh.rank, pf->min_samples, maxs, pf->pop_err,
                                                             pf->pop_z, d_cutk);
        AMCL_LAUNCH_CHECK();
      }
      sh.allreduce(d_cutk, 1, SHARD_I64, SHARD_MIN, sh.user);
      sync_ints(c, I_CUTK, 2);
      const unsigned long long cutk = *reinterpret_cast<unsigned long long *>(c->h_ints + I_CUTK);
      n = (int)(cutk >> 32);
      k_at_cut = (unsigned int)(cutk & 0xFFFFFFFFull);
    }
  }
  if (!shared_grid)
  {
    // One GPU, or the keys spread too far for a dense grid: every rank looks at
    // the full key list (all-gathered when sharded) and reaches the same n and k.
    if (sh.on())
      sh.allgather(c->d_keys, (unsigned long long)sh.chunk(maxs) * sizeof(unsigned long long), sh.user);
    build_histogram(c, maxs);
    tr.mark("histogram");
    scan_u32(c, c->d_flag, c->d_scan, maxs, I_DISTINCT);
    unsigned long long *d_cutk = reinterpret_cast<unsigned long long *>(c->d_ints + I_CUTK);
    AMCL_CUDA(cudaMemsetAsync(d_cutk, 0x7F, sizeof(unsigned long long), st));
    find_kld_cut<<<blocks_for(maxs, 256), 256, 0, st>>>(c->d_scan, 0, maxs, nullptr, 0, pf->min_samples, maxs,
                                                        pf->pop_err, pf->pop_z, d_cutk);
    AMCL_LAUNCH_CHECK();
    sync_ints(c, I_CUTK, 2);
    const unsigned long long cutk = *reinterpret_cast<unsigned long long *>(c->h_ints + I_CUTK);
    n = (int)(cutk >> 32);                               // pf.c:483
    k_at_cut = (unsigned int)(cutk & 0xFFFFFFFFull);     // leaf_count = k_{n-1}
  }
  set_b->sample_count = n;
  set_b->kdtree->leaf_count = (int)k_at_cut;
  tr.mark("scan+cut");

  if (host_stream)
  {
    // leave libc's stream where the reference would have left it: replay the
    // first n samples' draws from the saved state
    seed48(saved_state);
    for (int m = 0; m < n; m++)
    {
      if (drand48() < w_diff) (void)(pf->random_pose_fn)(pf->random_pose_data);
      else (void)drand48();
    }
  }
  c->draws_used = 2ull * (unsigned long long)n;

  // pf.c:487-489
  if (w_diff > 0.0) pf->w_slow = pf->w_fast = 0.0;

  const int own_lo = sh.lo(n), own_cnt = sh.count(n);
  if (sh.on() && n != maxs && own_cnt > 0)
  {
    // The cut changed the slicing (slices of n, not of max_samples).  Picking
    // is a pure function of (table, draw m), so each rank simply re-picks the
    // poses of its NEW slice instead of fetching them from their old owners.
    pa.first = own_lo;
    pa.count = own_cnt;
    pa.write_keys = 0;
    pa.weight = 1.0 / (double)n;
    pick_and_copy<<<blocks_for(own_cnt, 256), 256, 0, st>>>(pa);
    AMCL_LAUNCH_CHECK();
    if (shared_grid)
    {
      // this rank entered only its CANDIDATE slice into the grid; the samples of
      // its new slice find their bins in the merged grid
      bin_lookup<<<blocks_for(own_cnt, 256), 256, 0, st>>>(c->d_set[ib], own_lo, own_cnt, table_of(c),
                                                          pa.bin_x, pa.bin_y, pa.bin_t, c->d_slot);
      AMCL_LAUNCH_CHECK();
    }
  }
  // pf.c:493-498: every weight is 1/n (already so unless the set was cut on one GPU)
  if (own_cnt > 0 && n != maxs && !sh.on())
  {
    set_uniform_weight<<<blocks_for(own_cnt, 256), 256, 0, st>>>(c->d_set[ib] + own_lo, own_cnt,
                                                                 1.0 / (double)n);
    AMCL_LAUNCH_CHECK();
  }
  c->host_fresh[ib] = false;
  c->replicated[ib] = !sh.on();
  if (d_injected)
  {
    cudaFree(d_injected);     // synchronises; the picks above are done with it
    cudaFree(d_inject_slot);
  }

  tr.mark("weights");
  // pf.c:501-506
  cluster_and_stats(c, ib, n);
  tr.mark("clusters");
  pf->current_set = ib;
  run_update_converged(c, ib);
  tr.mark("converged");
  tr.report("pf_update_resample");
}

}  // namespace amclb200
