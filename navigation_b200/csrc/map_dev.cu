// map_dev.cu -- occupancy grid on the device: exact bounded Euclidean distance
// transform (stands in for the priority-queue wavefront of
// /root/reference/amcl/src/amcl/map/map_cspace.cpp:120-176), the ray bitmap for
// the beam model, and the map_* entry points of map.h:88-111.
//
// Distance semantics.  The reference's wavefront is NOT an exact transform: a
// cell keeps whatever source reached it first (map_cspace.cpp:95-116).
// north_star asks for an exact EDT, so this file computes, per cell,
//     d2 = min over occupied cells of (di^2 + dj^2)
// and stores d2 (or "far" when d2 > R^2, R = (int)(max_occ_dist/scale),
// map_cspace.cpp:42,102).  occ_dist on the host is then sqrt(d2)*scale with one
// IEEE sqrt and one multiply, the reference's own arithmetic (:49,:105), so it
// equals the reference bit-for-bit wherever the reference itself is exact, and
// is never larger.  See DESIGN.md "C1".
#include <errno.h>
#include <math.h>
#include <string.h>

#include <functional>
#include <mutex>
#include <thread>
#include <unordered_map>

#include <nvtx3/nvToolsExt.h>

#include "context.hpp"
#include "edt_rows.cuh"
#include "edt_squares.cuh"

namespace amclb200
{

// ------------------------------------------------------------------ globals --

Globals &globals()
{
  static Globals g;
  return g;
}

void ensure_device()
{
  Globals &g = globals();
  if (g.device >= 0) return;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    AMCL_FATAL("no CUDA device: libamcl_b200 has no CPU path (%s)",
               e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  int dev = 0;
  AMCL_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  AMCL_CUDA(cudaGetDeviceProperties(&prop, dev));
  g.device = dev;
  g.sm_count = prop.multiProcessorCount;
}

Trace::Trace()
{
  const char *e = getenv("AMCLB200_TRACE");
  on = e && e[0] == '1';
}

void Trace::mark(const char *name)
{
  if (!on) return;
  cudaEvent_t ev;
  AMCL_CUDA(cudaEventCreate(&ev));
  AMCL_CUDA(cudaEventRecord(ev, globals().stream));
  marks.push_back(std::make_pair(name, ev));
}

void Trace::report(const char *what)
{
  if (!on || marks.size() < 2) return;
  AMCL_CUDA(cudaEventSynchronize(marks.back().second));
  fprintf(stderr, "[amcl_b200 trace] %s:", what);
  for (size_t k = 1; k < marks.size(); k++)
  {
    float ms = 0;
    cudaEventElapsedTime(&ms, marks[k - 1].second, marks[k].second);
    fprintf(stderr, " %s=%.3f", marks[k].first, ms);
  }
  float tot = 0;
  cudaEventElapsedTime(&tot, marks.front().second, marks.back().second);
  fprintf(stderr, " total=%.3f ms\n", tot);
  for (auto &m : marks) cudaEventDestroy(m.second);
  marks.clear();
}

ApiRange::ApiRange(const char *name) { nvtxRangePushA(name); }
ApiRange::~ApiRange() { nvtxRangePop(); }

StageTimer::StageTimer(PfCtx *ctx, int which_stage) : c(ctx), stage(which_stage)
{
  if (!c->ev_stage[stage][0])
  {
    AMCL_CUDA(cudaEventCreate(&c->ev_stage[stage][0]));
    AMCL_CUDA(cudaEventCreate(&c->ev_stage[stage][1]));
  }
  AMCL_CUDA(cudaEventRecord(c->ev_stage[stage][0], globals().stream));
}

StageTimer::~StageTimer()
{
  AMCL_CUDA(cudaEventRecord(c->ev_stage[stage][1], globals().stream));
  c->stage_pending[stage] = true;
}

void stage_times_resolve(PfCtx *c)
{
  float *dst[3] = {&c->ms_action, &c->ms_sensor, &c->ms_resample};
  for (int k = 0; k < 3; k++)
  {
    if (!c->stage_pending[k]) continue;
    AMCL_CUDA(cudaEventSynchronize(c->ev_stage[k][1]));
    AMCL_CUDA(cudaEventElapsedTime(dst[k], c->ev_stage[k][0], c->ev_stage[k][1]));
    c->stage_pending[k] = false;
  }
}

// ----------------------------------------------------------------- registry --

static std::unordered_map<map_t *, MapCtx *> &map_registry()
{
  static std::unordered_map<map_t *, MapCtx *> r;
  return r;
}

MapCtx *map_ctx(map_t *map, bool create)
{
  auto &r = map_registry();
  auto it = r.find(map);
  if (it != r.end())
  {
    MapCtx *m = it->second;
    // the caller may have re-filled the struct (reconfigure path)
    if (m->sx != map->size_x || m->sy != map->size_y)
    {
      map_ctx_release(map);
      return map_ctx(map, create);
    }
    m->scale = map->scale;
    m->origin_x = map->origin_x;
    m->origin_y = map->origin_y;
    return m;
  }
  if (!create) return nullptr;
  ensure_device();
  if (map->size_x <= 0 || map->size_y <= 0 || map->cells == nullptr)
    AMCL_FATAL("map_t has no cells (size %d x %d)", map->size_x, map->size_y);
  MapCtx *m = new MapCtx();
  m->host = map;
  m->sx = map->size_x;
  m->sy = map->size_y;
  m->scale = map->scale;
  m->origin_x = map->origin_x;
  m->origin_y = map->origin_y;
  // one padding column and row (index sx / sy) hold the "off the map" code, so
  // the sensor kernels clamp instead of branching
  m->tiles_x = (m->sx + 1 + kTile - 1) / kTile;
  m->tiles_y = (m->sy + 1 + kTile - 1) / kTile;
  m->btiles_x = (m->sx + 7) / 8;
  m->btiles_y = (m->sy + 7) / 8;
  size_t cells = (size_t)m->sx * m->sy;
  AMCL_CUDA(cudaMalloc(&m->d_occ, cells));
  AMCL_CUDA(cudaMalloc(&m->d_code, (size_t)m->tiles_x * m->tiles_y * 16 * sizeof(uint16_t)));
  AMCL_CUDA(cudaMalloc(&m->d_blocked, (size_t)m->btiles_x * m->btiles_y * sizeof(uint64_t)));
  r[map] = m;
  return m;
}

void map_ctx_release(map_t *map)
{
  auto &r = map_registry();
  auto it = r.find(map);
  if (it == r.end()) return;
  MapCtx *m = it->second;
  cudaFree(m->d_occ);
  cudaFree(m->d_code);
  cudaFree(m->d_blocked);
  if (m->d_rowdist) cudaFree(m->d_rowdist);
  if (m->d_rowsq) cudaFree(m->d_rowsq);
  if (m->h_occ_pin) cudaFreeHost(m->h_occ_pin);
  if (m->h_code_pin) cudaFreeHost(m->h_code_pin);
  if (m->d_rowbits) cudaFree(m->d_rowbits);
  if (m->lf_tex) cudaDestroyTextureObject(m->lf_tex);
  if (m->d_lf) cudaFree(m->d_lf);
  if (m->lf_tex2d) cudaDestroyTextureObject(m->lf_tex2d);
  if (m->lf_surf) cudaDestroySurfaceObject(m->lf_surf);
  if (m->lf_array) cudaFreeArray(m->lf_array);
  if (m->d_rank_of_code) cudaFree(m->d_rank_of_code);
  delete m;
  r.erase(it);
}

// ------------------------------------------------------------------ kernels --

// One thread per 8x8 bit tile: bit (lj*8+li) is set when a ray must stop in
// that cell, i.e. occ_state > -1 (map_range.c:89,110).  Cells outside the map
// inside a partial edge tile are marked blocked too.
__global__ void build_blocked_bits(const int8_t *__restrict__ occ, int sx, int sy,
                                   int btx, int bty, uint64_t *__restrict__ bits)
{
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= btx * bty) return;
  int bi = t % btx, bj = t / btx;
  uint64_t word = 0;
  for (int lj = 0; lj < 8; lj++)
  {
    int j = bj * 8 + lj;
    for (int li = 0; li < 8; li++)
    {
      int i = bi * 8 + li;
      bool blocked = true;
      if (i < sx && j < sy) blocked = occ[(size_t)j * sx + i] > -1;
      if (blocked) word |= 1ull << (lj * 8 + li);
    }
  }
  bits[t] = word;
}

// ---- exact bounded EDT, three kernels --------------------------------------------
//
// (1) edt_pack_rows: occupancy (== +1) of every row as a bit string, 64 cells per
//     word, zero words of padding on both sides of each row.
// (2) edt_row_bits:  g(i, j) = distance along row j from cell i to the nearest
//     occupied cell, capped at R+1, read off the bit string with count-leading /
//     trailing-zero instructions (no neighbour scan).  One byte per cell.
// (3) edt_col_pass:  d2(i, j) = min over |dj| <= R of dj^2 + g(i, j+dj)^2,
//     thresholded at R^2; each thread resolves four neighbouring cells from
//     packed bytes; results leave through shared memory so that every 4x4 tile
//     of codes is one contiguous 32-byte sector.

__global__ void edt_pack_rows(const int8_t *__restrict__ occ, int sx, int sy, int words, int pad,
                              unsigned long long *__restrict__ bits)
{
  const int lane = threadIdx.x & 31;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int stride = words + 2 * pad;
  const int j = (int)(warp / stride), w = (int)(warp % stride) - pad;
  if (j >= sy) return;
  unsigned long long word = 0ull;
  if (w >= 0 && w < words)
  {
    const int8_t *row = occ + (size_t)j * sx;
    const int i0 = w * 64 + lane, i1 = i0 + 32;
    const unsigned lo = __ballot_sync(kFullMask, i0 < sx && row[i0] == 1);
    const unsigned hi = __ballot_sync(kFullMask, i1 < sx && row[i1] == 1);
    word = (unsigned long long)lo | ((unsigned long long)hi << 32);
  }
  if (lane == 0) bits[(size_t)j * stride + (w + pad)] = word;
}

// One thread per four neighbouring cells (one 4-byte store) when the row pitch
// allows it, else per cell.
__global__ void edt_row_bits(const unsigned long long *__restrict__ bits, int sx, int sy, int words,
                             int pad, int R, unsigned char *__restrict__ g)
{
  const int j = blockIdx.y;
  const int stride = words + 2 * pad;
  const unsigned long long *A = bits + (size_t)j * stride + pad;   // A[-pad .. words+pad)
  if ((sx & 3) == 0)
  {
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= sx) return;
    unsigned v = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) v |= (unsigned)row_distance(A, i + b, R) << (8 * b);
    *reinterpret_cast<unsigned *>(g + (size_t)j * sx + i) = v;
    return;
  }
  for (int b = 0; b < 4; b++)
  {
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4 + b;
    if (i < sx) g[(size_t)j * sx + i] = (unsigned char)row_distance(A, i, R);
  }
}

constexpr int kPatchX = 64, kPatchY = 32;

__global__ void __launch_bounds__(256)
edt_col_pass(const unsigned char *__restrict__ g, int sx, int sy, int R, int tiles_x,
             uint16_t *__restrict__ code)
{
  extern __shared__ __align__(16) unsigned char s_raw[];
  unsigned char *s_g = s_raw;                                        // (kPatchY + 2R) x kPatchX
  const int rows = kPatchY + 2 * R;
  uint16_t *s_out = reinterpret_cast<uint16_t *>(s_raw + (((size_t)rows * kPatchX + 15) & ~(size_t)15));
  const int i0 = blockIdx.x * kPatchX, j0 = blockIdx.y * kPatchY;
  const unsigned char far = (unsigned char)(R + 1);
  if ((sx & 3) == 0)
  {
    // rows of g are 4-byte aligned: move the patch as words (16 per row)
    const unsigned far4 = far * 0x01010101u;
    unsigned *s_g4 = reinterpret_cast<unsigned *>(s_g);
    for (int k = threadIdx.x; k < rows * (kPatchX / 4); k += blockDim.x)
    {
      const int lj = k / (kPatchX / 4), li = (k % (kPatchX / 4)) * 4;
      const int j = j0 - R + lj, i = i0 + li;
      s_g4[k] = (j >= 0 && j < sy && i < sx)
                    ? __ldg(reinterpret_cast<const unsigned *>(g + (size_t)j * sx + i)) : far4;
    }
  }
  else
  {
    for (int k = threadIdx.x; k < rows * kPatchX; k += blockDim.x)
    {
      const int lj = k / kPatchX, li = k % kPatchX;
      const int j = j0 - R + lj, i = i0 + li;
      s_g[k] = (j >= 0 && j < sy && i < sx) ? g[(size_t)j * sx + i] : far;
    }
  }
  __syncthreads();
  const int R2 = R * R;
  for (int quad = threadIdx.x; quad < kPatchX * kPatchY / 4; quad += blockDim.x)
  {
    const int lj = quad / (kPatchX / 4), li = (quad % (kPatchX / 4)) * 4;
    int best[4] = {R2 + 1, R2 + 1, R2 + 1, R2 + 1};
    for (int d = 0; d <= R; d++)
    {
      const int dd = d * d;
      const int worst = max(max(best[0], best[1]), max(best[2], best[3]));
      if (dd >= worst) break;                       // farther rows cannot improve any of the four
      const unsigned ga = *reinterpret_cast<const unsigned *>(s_g + (lj + R - d) * kPatchX + li);
      const unsigned gb = *reinterpret_cast<const unsigned *>(s_g + (lj + R + d) * kPatchX + li);
      const unsigned m = __vminu4(ga, gb);
#pragma unroll
      for (int b = 0; b < 4; b++)
      {
        const int gg = (int)((m >> (8 * b)) & 255u);
        const int v = dd + gg * gg;
        best[b] = v < best[b] ? v : best[b];
      }
    }
    const int tj = lj >> 2, ti = li >> 2;           // the quad is one row of one 4x4 tile
    uint16_t *o = s_out + (tj * (kPatchX / kTile) + ti) * 16 + (lj & 3) * 4;
#pragma unroll
    for (int b = 0; b < 4; b++)
    {
      const bool inside = i0 + li + b < sx && j0 + lj < sy;
      o[b] = (uint16_t)(!inside ? R2 + 2 : (best[b] > R2 ? R2 + 1 : best[b]));   // R2 + 2: off the map
    }
  }
  __syncthreads();
  // (kPatchX/4) x (kPatchY/4) tiles of 32 bytes; a thread moves one 16-byte half
  const int halves = (kPatchX / kTile) * (kPatchY / kTile) * 2;
  for (int q = threadIdx.x; q < halves; q += blockDim.x)
  {
    const int tile = q >> 1, part = q & 1;
    const int tj = tile / (kPatchX / kTile), ti = tile % (kPatchX / kTile);
    const int gi = (i0 >> 2) + ti, gj = (j0 >> 2) + tj;
    if (gi >= tiles_x || gj * kTile > sy) continue;
    const uint4 *src = reinterpret_cast<const uint4 *>(s_out + tile * 16 + part * 8);
    uint4 *dst = reinterpret_cast<uint4 *>(code + ((size_t)gj * tiles_x + gi) * 16 + part * 8);
    *dst = *src;
  }
}

// ------------------------------------------------------------ host-side glue --

// The caller's grid is an array of 16-byte structs; converting 16 M of them to
// and from the packed device formats is host-memory bound, so the row loops are
// spread over the host cores.
void parallel_rows(int rows, const std::function<void(int, int)> &body)
{
  unsigned hw = std::thread::hardware_concurrency();
  int nt = (int)(hw ? (hw > 16 ? 16 : hw) : 4);
  if (rows < 256 || nt < 2)
  {
    body(0, rows);
    return;
  }
  std::vector<std::thread> pool;
  const int per = (rows + nt - 1) / nt;
  for (int t = 0; t < nt; t++)
  {
    const int lo = t * per, hi = lo + per < rows ? lo + per : rows;
    if (lo >= hi) break;
    pool.emplace_back(body, lo, hi);
  }
  for (auto &th : pool) th.join();
}


// Page-locked staging that lives as long as the map: a pageable std::vector per
// call costs a page fault per 4 KB and a bounce copy inside the driver.
template <typename T>
static T *pinned_staging(T **slot, size_t *cap, size_t count)
{
  if (count > *cap)
  {
    if (*slot) cudaFreeHost(*slot);
    *slot = nullptr;
    AMCL_CUDA(cudaHostAlloc((void **)slot, count * sizeof(T), cudaHostAllocDefault));
    *cap = count;
  }
  return *slot;
}

// Bands of rows the host <-> device conversions are pipelined in: the copy of
// one band runs while the host cores convert the next.
static int staging_bands(size_t cells) { return cells >= ((size_t)1 << 20) ? 4 : 1; }

void map_upload_occupancy(MapCtx *m)
{
  const map_t *map = m->host;
  size_t cells = (size_t)m->sx * m->sy;
  int8_t *occ = pinned_staging(&m->h_occ_pin, &m->h_occ_cap, cells);
  const int sx = m->sx;
  cudaStream_t st = globals().stream;
  const int bands = staging_bands(cells);
  for (int b = 0; b < bands; b++)
  {
    const int r0 = (int)((long long)m->sy * b / bands), r1 = (int)((long long)m->sy * (b + 1) / bands);
    if (r1 <= r0) continue;
    parallel_rows(r1 - r0, [&](int j0, int j1) {
      for (int j = r0 + j0; j < r0 + j1; j++)
      {
        const map_cell_t *row = map->cells + (size_t)j * sx;
        int8_t *o = occ + (size_t)j * sx;
        for (int i = 0; i < sx; i++)
        {
          const int state = row[i].occ_state;
          o[i] = (int8_t)(state > 0 ? 1 : (state < 0 ? -1 : 0));
        }
      }
    });
    AMCL_CUDA(cudaMemcpyAsync(m->d_occ + (size_t)r0 * sx, occ + (size_t)r0 * sx, (size_t)(r1 - r0) * sx,
                              cudaMemcpyHostToDevice, st));
  }
  int nt = m->btiles_x * m->btiles_y;
  build_blocked_bits<<<(nt + 255) / 256, 256, 0, st>>>(m->d_occ, m->sx, m->sy, m->btiles_x,
                                                         m->btiles_y, m->d_blocked);
  AMCL_LAUNCH_CHECK();
  AMCL_CUDA(cudaStreamSynchronize(st));  // the staging buffer is free for the next call
  m->occ_uploaded = true;
}

// d_occ was written by a kernel (amclb200_map_from_image): derive the
// ray-casting bit tiles from it and mark the occupancy as resident.
void map_finish_occupancy_on_device(MapCtx *m)
{
  cudaStream_t st = globals().stream;
  int nt = m->btiles_x * m->btiles_y;
  build_blocked_bits<<<(nt + 255) / 256, 256, 0, st>>>(m->d_occ, m->sx, m->sy, m->btiles_x,
                                                         m->btiles_y, m->d_blocked);
  AMCL_LAUNCH_CHECK();
  m->occ_uploaded = true;
}

// The largest radius the uint16 codes (d2 <= 254^2, two sentinels) can hold.
constexpr int kMaxCellRadius = 254;

static int cell_radius_of(const MapCtx *m, double max_occ_dist)
{
  // map_cspace.cpp:42 -- truncation of an fp64 quotient, e.g. 0.3/0.1 -> 2
  int R = (int)(max_occ_dist / m->scale);
  return R < 0 ? 0 : R;
}

static void set_code_table(MapCtx *m, double max_occ_dist)
{
  const int R = cell_radius_of(m, max_occ_dist);
  if (R > kMaxCellRadius)
    AMCL_FATAL("max_occ_dist/scale = %d cells exceeds the %d-cell radius the uint16 "
               "distance codes are sized for", R, kMaxCellRadius);
  m->R = R;
  m->max_occ_dist = max_occ_dist;
  m->dist_of_code.assign((size_t)R * R + 2, 0.0);
  for (int d2 = 0; d2 <= R * R; d2++)
    m->dist_of_code[d2] = sqrt((double)d2) * m->scale;   // map_cspace.cpp:49,105
  m->dist_of_code[(size_t)R * R + 1] = max_occ_dist;      // map_cspace.cpp:148
  // ranks of the codes that can occur: squared distances between grid cells,
  // then "far" (R^2 + 1) and "off the map" (R^2 + 2)
  const int R2 = R * R;
  std::vector<char> occurs((size_t)R2 + 1, 0);
  for (int i = 0; i <= R; i++)
    for (int j = i; i * i + j * j <= R2; j++) occurs[(size_t)(i * i + j * j)] = 1;
  m->rank_of_code.assign((size_t)R2 + 3, (uint16_t)0xffff);
  m->code_of_rank.clear();
  for (int d2 = 0; d2 <= R2; d2++)
    if (occurs[d2])
    {
      m->rank_of_code[d2] = (uint16_t)m->code_of_rank.size();
      m->code_of_rank.push_back(d2);
    }
  for (int extra = 1; extra <= 2; extra++)
  {
    m->rank_of_code[(size_t)R2 + extra] = (uint16_t)m->code_of_rank.size();
    m->code_of_rank.push_back(R2 + extra);
  }
  if ((int)m->rank_of_code.size() > m->rank_cap)
  {
    if (m->d_rank_of_code) cudaFree(m->d_rank_of_code);
    m->rank_cap = (int)m->rank_of_code.size();
    AMCL_CUDA(cudaMalloc(&m->d_rank_of_code, sizeof(uint16_t) * (size_t)m->rank_cap));
  }
  AMCL_CUDA(cudaMemcpyAsync(m->d_rank_of_code, m->rank_of_code.data(),
                            sizeof(uint16_t) * m->rank_of_code.size(), cudaMemcpyHostToDevice,
                            globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
}

// One thread per 4x4 tile of the laser view: fetch the tile of codes `apron`
// cells up and to the left (32 bytes), replace every code by its rank, and store
// the tile column-first.  Tiles outside the map hold the off-map rank.
__global__ void build_laser_view(const uint16_t *__restrict__ code, int tiles_x, int sx, int sy,
                                 const uint16_t *__restrict__ rank_of_code, int off_rank,
                                 int apron, int w, int h, uint16_t *__restrict__ out)
{
  const int tw = w >> 2;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)tw * (h >> 2)) return;
  const int ti = (int)(t % tw), tj = (int)(t / tw);
  const int si = ti - (apron >> 2), sj = tj - (apron >> 2);
  __align__(16) uint16_t v[16];
  if (si >= 0 && sj >= 0 && si * 4 < sx && sj * 4 < sy)
  {
    __align__(16) uint16_t c[16];
    const uint4 *src = reinterpret_cast<const uint4 *>(code + ((size_t)sj * tiles_x + si) * 16);
    *reinterpret_cast<uint4 *>(c) = __ldg(src);
    *reinterpret_cast<uint4 *>(c + 8) = __ldg(src + 1);
#pragma unroll
    for (int lj = 0; lj < 4; lj++)
#pragma unroll
      for (int li = 0; li < 4; li++) v[li * 4 + lj] = __ldg(rank_of_code + c[lj * 4 + li]);
  }
  else
  {
#pragma unroll
    for (int k = 0; k < 16; k++) v[k] = (uint16_t)off_rank;
  }
  uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)t * 16);
  dst[0] = *reinterpret_cast<const uint4 *>(v);
  dst[1] = *reinterpret_cast<const uint4 *>(v + 8);
}

// `reach_cells`: how far from its laser a particle's beams can end.  The apron is
// that wide when memory allows (at most the map's own size again plus 64 MB), so
// that every particle on the map takes the clamp-free path; a narrower apron only
// sends the particles near the border through the clamped path.
// The same ranks into a 2-D array through a surface: one thread per 4x4 tile of the
// view, four 8-byte stores.
__global__ void build_laser_array(const uint16_t *__restrict__ code, int tiles_x, int sx, int sy,
                                  const uint16_t *__restrict__ rank_of_code, int off_rank,
                                  int apron, int w, int h, cudaSurfaceObject_t surf)
{
  const int tw = w >> 2;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)tw * (h >> 2)) return;
  const int ti = (int)(t % tw), tj = (int)(t / tw);
  const int si = ti - (apron >> 2), sj = tj - (apron >> 2);
  const bool on_map = si >= 0 && sj >= 0 && si * 4 < sx && sj * 4 < sy;
  __align__(16) uint16_t c[16];
  if (on_map)
  {
    const uint4 *src = reinterpret_cast<const uint4 *>(code + ((size_t)sj * tiles_x + si) * 16);
    *reinterpret_cast<uint4 *>(c) = __ldg(src);
    *reinterpret_cast<uint4 *>(c + 8) = __ldg(src + 1);
  }
#pragma unroll
  for (int lj = 0; lj < 4; lj++)
  {
    ushort4 v;
    v.x = on_map ? __ldg(rank_of_code + c[lj * 4 + 0]) : (uint16_t)off_rank;
    v.y = on_map ? __ldg(rank_of_code + c[lj * 4 + 1]) : (uint16_t)off_rank;
    v.z = on_map ? __ldg(rank_of_code + c[lj * 4 + 2]) : (uint16_t)off_rank;
    v.w = on_map ? __ldg(rank_of_code + c[lj * 4 + 3]) : (uint16_t)off_rank;
    surf2Dwrite(v, surf, ti * 4 * (int)sizeof(uint16_t), tj * 4 + lj);
  }
}

static void release_laser_array(MapCtx *m)
{
  if (m->lf_tex2d) cudaDestroyTextureObject(m->lf_tex2d);
  if (m->lf_surf) cudaDestroySurfaceObject(m->lf_surf);
  if (m->lf_array) cudaFreeArray(m->lf_array);
  m->lf_tex2d = 0;
  m->lf_surf = 0;
  m->lf_array = nullptr;
  m->lf_array_w = m->lf_array_h = 0;
}

void map_prepare_laser_view(MapCtx *m, int reach_cells)
{
  if (!m->dist_valid) AMCL_FATAL("laser view requested before map_update_cspace()");
  int want = ((reach_cells > 4 ? reach_cells : 4) + 3) & ~3;
  const size_t budget = 2 * (size_t)m->sx * m->sy + ((size_t)32 << 20);   // elements
  while (want > 4 && (size_t)(m->sx + 2 * want) * (size_t)(m->sy + 2 * want) > budget) want -= 4;
  if (m->lf_generation == m->generation && m->lf_apron >= want) return;
  const int apron = want > m->lf_apron ? want : m->lf_apron;
  const int w = (m->sx + 2 * apron + 7) & ~7, h = (m->sy + 2 * apron + 7) & ~7;
  const size_t elems = (size_t)w * h;
  if (elems > m->lf_cap)
  {
    if (m->lf_tex) cudaDestroyTextureObject(m->lf_tex);
    m->lf_tex = 0;
    if (m->d_lf) cudaFree(m->d_lf);
    m->lf_cap = elems;
    AMCL_CUDA(cudaMalloc(&m->d_lf, elems * sizeof(uint16_t)));
    // the same memory as a linear texture, for kernels that fetch through the texture pipe
    int max_linear = 0;
    cudaDeviceGetAttribute(&max_linear, cudaDevAttrMaxTexture1DLinearWidth, globals().device);
    if (elems <= (size_t)max_linear)
    {
      cudaResourceDesc res;
      memset(&res, 0, sizeof(res));
      res.resType = cudaResourceTypeLinear;
      res.res.linear.devPtr = m->d_lf;
      res.res.linear.desc = cudaCreateChannelDesc<unsigned short>();
      res.res.linear.sizeInBytes = elems * sizeof(uint16_t);
      cudaTextureDesc td;
      memset(&td, 0, sizeof(td));
      td.readMode = cudaReadModeElementType;
      AMCL_CUDA(cudaCreateTextureObject(&m->lf_tex, &res, &td, nullptr));
    }
  }
  const long long tiles = (long long)(w >> 2) * (h >> 2);
  build_laser_view<<<(unsigned)((tiles + 255) / 256), 256, 0, globals().stream>>>(
      m->d_code, m->tiles_x, m->sx, m->sy, m->d_rank_of_code, (int)m->code_of_rank.size() - 1,
      apron, w, h, m->d_lf);
  AMCL_LAUNCH_CHECK();
  // The view row by row in a 2-D array as well: the likelihood-field kernels fetch
  // texels from it (the texture unit does the addressing).  AMCLB200_LF_TEX2D=0 leaves
  // it out (A/B timing; the kernels then fetch from the linear texture).
  static const bool want_array = [] {
    const char *e = getenv("AMCLB200_LF_TEX2D");
    return !e || atoi(e) != 0;
  }();
  if (want_array)
  {
    if (m->lf_array_w != w || m->lf_array_h != h)
    {
      release_laser_array(m);
      const cudaChannelFormatDesc desc = cudaCreateChannelDesc<unsigned short>();
      if (cudaMallocArray(&m->lf_array, &desc, (size_t)w, (size_t)h, cudaArraySurfaceLoadStore) == cudaSuccess)
      {
        cudaResourceDesc res;
        memset(&res, 0, sizeof(res));
        res.resType = cudaResourceTypeArray;
        res.res.array.array = m->lf_array;
        cudaTextureDesc td;
        memset(&td, 0, sizeof(td));
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint;
        td.readMode = cudaReadModeElementType;
        td.normalizedCoords = 0;
        AMCL_CUDA(cudaCreateTextureObject(&m->lf_tex2d, &res, &td, nullptr));
        AMCL_CUDA(cudaCreateSurfaceObject(&m->lf_surf, &res));
        m->lf_array_w = w;
        m->lf_array_h = h;
      }
      else
      {
        cudaGetLastError();          // too large for an array: the linear view serves
        m->lf_array = nullptr;
      }
    }
    if (m->lf_array)
    {
      build_laser_array<<<(unsigned)((tiles + 255) / 256), 256, 0, globals().stream>>>(
          m->d_code, m->tiles_x, m->sx, m->sy, m->d_rank_of_code, (int)m->code_of_rank.size() - 1,
          apron, w, h, m->lf_surf);
      AMCL_LAUNCH_CHECK();
    }
  }
  m->lf_apron = apron;
  m->lf_w = w;
  m->lf_h = h;
  m->lf_generation = m->generation;
}

void map_build_distance_codes(MapCtx *m, double max_occ_dist)
{
  set_code_table(m, max_occ_dist);
  cudaStream_t st = globals().stream;
  const int R = m->R;
  const EdtPlan plan = edt_plan(m->sx, m->sy, R);
  const int words = plan.words, pad = plan.pad;
  const size_t nbits = (size_t)m->sy * plan.stride;
  if (nbits > m->rowbits_cap)
  {
    if (m->d_rowbits) cudaFree(m->d_rowbits);
    m->rowbits_cap = nbits;
    AMCL_CUDA(cudaMalloc(&m->d_rowbits, nbits * sizeof(unsigned long long)));
  }
  if (m->sy > 65535) AMCL_FATAL("maps taller than 65535 rows are not supported (%d)", m->sy);
  // AMCLB200_EDT=search keeps the per-cell search kernels for any radius (A/B timing)
  static const bool force_search = [] {
    const char *e = getenv("AMCLB200_EDT");
    return e && strcmp(e, "search") == 0;
  }();
  const bool fast = R <= kFastRadius && !force_search;
  const int pitch = plan.pitch;
  if (fast && (size_t)pitch * m->sy > m->rowsq_cap)
  {
    if (m->d_rowsq) cudaFree(m->d_rowsq);
    m->rowsq_cap = (size_t)pitch * m->sy;
    AMCL_CUDA(cudaMalloc(&m->d_rowsq, m->rowsq_cap * sizeof(uint16_t)));
  }
  if (!fast && !m->d_rowdist) AMCL_CUDA(cudaMalloc(&m->d_rowdist, (size_t)m->sx * m->sy));
  cudaEvent_t ev0, ev1;
  AMCL_CUDA(cudaEventCreate(&ev0));
  AMCL_CUDA(cudaEventCreate(&ev1));
  AMCL_CUDA(cudaEventRecord(ev0, st));
  if (fast && (m->sx & 15) == 0)
  {
    edt_pack_rows16<<<(unsigned)((plan.pack_threads + 255) / 256), 256, 0, st>>>(m->d_occ, m->sx, m->sy, words, pad,
                                                                       m->d_rowbits);
    AMCL_LAUNCH_CHECK();
  }
  else
  {
    const long long warps = (long long)m->sy * (words + 2 * pad);
    edt_pack_rows<<<(unsigned)((warps + 7) / 8), 256, 0, st>>>(m->d_occ, m->sx, m->sy, words, pad,
                                                               m->d_rowbits);
    AMCL_LAUNCH_CHECK();
  }
  if (fast)
  {
    dim3 g1(plan.row_blocks_x, m->sy);          // eight cells per thread
    edt_row_squares<<<g1, 256, 0, st>>>(m->d_rowbits, pitch, words, pad, R, m->d_rowsq);
    AMCL_LAUNCH_CHECK();
    const int chunks = plan.chunks;
    dim3 g2(plan.col_blocks_x, plan.col_blocks_y);
    const size_t smem = plan.col_smem;
    if (smem > 48 * 1024)
      AMCL_CUDA(cudaFuncSetAttribute(edt_col_squares, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)smem));
    edt_col_squares<<<g2, 256, smem, st>>>(m->d_rowsq, pitch, m->sx, m->sy, R, chunks, m->tiles_x,
                                           m->tiles_y, m->d_code);
    AMCL_LAUNCH_CHECK();
  }
  else
  {
    dim3 g1((m->sx + 1023) / 1024, m->sy);     // four cells per thread
    edt_row_bits<<<g1, 256, 0, st>>>(m->d_rowbits, m->sx, m->sy, words, pad, R, m->d_rowdist);
    AMCL_LAUNCH_CHECK();
    dim3 g2((m->sx + 1 + kPatchX - 1) / kPatchX, (m->sy + 1 + kPatchY - 1) / kPatchY);  // + padding
    size_t smem = (((size_t)(kPatchY + 2 * R) * kPatchX + 15) & ~(size_t)15) +
                  (size_t)kPatchX * kPatchY * sizeof(uint16_t);
    if (smem > 48 * 1024)
      AMCL_CUDA(cudaFuncSetAttribute(edt_col_pass, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)smem));
    edt_col_pass<<<g2, 256, smem, st>>>(m->d_rowdist, m->sx, m->sy, R, m->tiles_x, m->d_code);
    AMCL_LAUNCH_CHECK();
  }
  AMCL_CUDA(cudaEventRecord(ev1, st));
  AMCL_CUDA(cudaEventSynchronize(ev1));
  AMCL_CUDA(cudaEventElapsedTime(&m->ms_edt, ev0, ev1));
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  m->dist_valid = true;
  m->generation++;
}

void map_codes_to_host_cells(MapCtx *m)
{
  size_t ncodes = (size_t)m->tiles_x * m->tiles_y * 16;
  uint16_t *codes = pinned_staging(&m->h_code_pin, &m->h_code_cap, ncodes);
  cudaStream_t st = globals().stream;
  // bands of tile rows (a tile row is contiguous): band b + 1 crosses PCIe while the
  // host cores spread band b over the caller's 16-byte cells
  const int bands = staging_bands((size_t)m->sx * m->sy);
  cudaEvent_t arrived[4];
  int t_lo[4], t_hi[4];
  for (int b = 0; b < bands; b++)
  {
    t_lo[b] = (int)((long long)m->tiles_y * b / bands);
    t_hi[b] = (int)((long long)m->tiles_y * (b + 1) / bands);
    AMCL_CUDA(cudaEventCreateWithFlags(&arrived[b], cudaEventDisableTiming));
    const size_t off = (size_t)t_lo[b] * m->tiles_x * 16, len = (size_t)(t_hi[b] - t_lo[b]) * m->tiles_x * 16;
    if (len)
      AMCL_CUDA(cudaMemcpyAsync(codes + off, m->d_code + off, len * sizeof(uint16_t), cudaMemcpyDeviceToHost, st));
    AMCL_CUDA(cudaEventRecord(arrived[b], st));
  }
  map_t *map = m->host;
  const double *lut = m->dist_of_code.data();
  for (int b = 0; b < bands; b++)
  {
    AMCL_CUDA(cudaEventSynchronize(arrived[b]));
    cudaEventDestroy(arrived[b]);
    const int r0 = t_lo[b] * kTile, r1 = t_hi[b] * kTile < m->sy ? t_hi[b] * kTile : m->sy;
    if (r1 <= r0) continue;
    parallel_rows(r1 - r0, [&](int j0, int j1) {
      for (int j = r0 + j0; j < r0 + j1; j++)
      {
        const uint16_t *trow = codes + (size_t)(j >> 2) * m->tiles_x * 16 + (j & 3) * 4;
        map_cell_t *crow = map->cells + (size_t)j * m->sx;
        for (int i = 0; i < m->sx; i++)
          crow[i].occ_dist = lut[trow[(size_t)(i >> 2) * 16 + (i & 3)]];
      }
    });
  }
  map->max_occ_dist = m->max_occ_dist;
}

// Inverse of the above: take whatever distances the caller wrote into
// cells[].occ_dist (e.g. the reference wavefront's output, for like-for-like
// sensor parity) and encode them.  Every value must be one the code table can
// express exactly; anything else is a caller error: one line on stderr, the
// device map is left as it was, false is returned.
bool map_codes_from_host_cells(MapCtx *m)
{
  map_t *map = m->host;
  if (cell_radius_of(m, map->max_occ_dist) > kMaxCellRadius)
  {
    fprintf(stderr, "[amcl_b200] max_occ_dist/scale = %d cells exceeds the %d-cell radius the distance "
                    "codes are sized for: cannot encode, device map unchanged\n",
            cell_radius_of(m, map->max_occ_dist), kMaxCellRadius);
    return false;
  }
  const double old_max = m->max_occ_dist;
  const bool had_table = !m->dist_of_code.empty();
  set_code_table(m, map->max_occ_dist);
  const int R2 = m->R * m->R;
  size_t ncodes = (size_t)m->tiles_x * m->tiles_y * 16;
  std::vector<uint16_t> codes(ncodes, (uint16_t)(R2 + 2));   // padding: off the map
  for (int j = 0; j < m->sy; j++)
    for (int i = 0; i < m->sx; i++)
    {
      double d = map->cells[(size_t)j * m->sx + i].occ_dist;
      int code;
      if (d == m->max_occ_dist)
        code = R2 + 1;
      else
      {
        double q = d / m->scale;
        long d2 = lround(q * q);
        if (d2 < 0 || d2 > R2 || m->dist_of_code[d2] != d || m->rank_of_code[d2] == 0xffff)
        {
          fprintf(stderr, "[amcl_b200] occ_dist %.17g at cell (%d,%d) is not sqrt(di^2 + dj^2)*scale "
                          "with di^2 + dj^2 <= %d: cannot encode, device map unchanged\n", d, i, j, R2);
          if (had_table) set_code_table(m, old_max);   // the codes on the device still mean what they meant
          return false;
        }
        code = (int)d2;
      }
      codes[((size_t)(j >> 2) * m->tiles_x + (i >> 2)) * 16 + (j & 3) * 4 + (i & 3)] =
          (uint16_t)code;
    }
  AMCL_CUDA(cudaMemcpyAsync(m->d_code, codes.data(), ncodes * sizeof(uint16_t),
                            cudaMemcpyHostToDevice, globals().stream));
  AMCL_CUDA(cudaStreamSynchronize(globals().stream));
  m->dist_valid = true;
  m->generation++;
  return true;
}

}  // namespace amclb200

// ===================================================================== C API ==

using namespace amclb200;

extern "C" {

// map.c:38-57
map_t *map_alloc(void)
{
  map_t *map = (map_t *)malloc(sizeof(map_t));
  map->origin_x = 0;
  map->origin_y = 0;
  map->size_x = 0;
  map->size_y = 0;
  map->scale = 0;
  map->cells = (map_cell_t *)NULL;
  map->max_occ_dist = 0;
  return map;
}

// map.c:61-66 -- cells were malloc'd by the caller and are freed here.
void map_free(map_t *map)
{
  if (!map) return;
  map_ctx_release(map);
  free(map->cells);
  free(map);
}

// map.c:70-84 -- a single host-side lookup into the host mirror of the grid.
map_cell_t *map_get_cell(map_t *map, double ox, double oy, double oa)
{
  (void)oa;
  int i = MAP_GXWX(map, ox);
  int j = MAP_GYWY(map, oy);
  if (!MAP_VALID(map, i, j)) return NULL;
  return map->cells + MAP_INDEX(map, i, j);
}

// map_store.c:39-136: fills the grid from a binary PGM ("P5") file.  Host code;
// the occupancy reaches the device with the next map_update_cspace().  Rules of
// the reference: '#' comment lines may follow the magic number; the first image
// row is the TOP row of the map (j = height - 1); a pixel below depth/4 is
// occupied, above 3*depth/4 free, anything between unknown (the other way round
// with `negate`); a map that already has cells must have the file's dimensions;
// a short file leaves fgetc()'s EOF (-1) as the pixel value, as there.
int map_load_occ(map_t *map, const char *filename, double scale, int negate)
{
  FILE *file = fopen(filename, "r");
  if (file == NULL)
  {
    fprintf(stderr, "%s: %s\n", strerror(errno), filename);
    return -1;
  }
  char magic[3] = {0, 0, 0};
  if (fscanf(file, "%2s \n", magic) != 1 || strcmp(magic, "P5") != 0)
  {
    fprintf(stderr, "incorrect image format; must be PGM/binary");
    fclose(file);
    return -1;
  }
  int ch;
  while ((ch = fgetc(file)) == '#')
    while ((ch = fgetc(file)) != '\n' && ch != EOF) {}
  ungetc(ch, file);
  int width = 0, height = 0, depth = 0;
  if (fscanf(file, " %d %d \n %d \n", &width, &height, &depth) != 3 || width <= 0 || height <= 0)
  {
    fprintf(stderr, "Failed ot read image dimensions");
    fclose(file);
    return -1;
  }
  if (map->cells == NULL)
  {
    map->scale = scale;
    map->size_x = width;
    map->size_y = height;
    map->cells = (map_cell_t *)calloc((size_t)width * height, sizeof(map->cells[0]));
  }
  else if (width != map->size_x || height != map->size_y)
  {
    fclose(file);
    return -1;
  }
  const int lo = depth / 4, hi = 3 * depth / 4;
  for (int j = height - 1; j >= 0; j--)
    for (int i = 0; i < width; i++)
    {
      ch = fgetc(file);
      const int dark = ch < lo, light = ch > hi;
      const int occ = dark ? (negate ? -1 : +1) : (light ? (negate ? +1 : -1) : 0);
      if (!MAP_VALID(map, i, j)) continue;
      map->cells[MAP_INDEX(map, i, j)].occ_state = occ;
    }
  fclose(file);
  // a device copy made from earlier contents of these cells is stale now
  MapCtx *m = map_ctx(map, false);
  if (m)
  {
    m->occ_uploaded = false;
    m->dist_valid = false;
  }
  return 0;
}

// map_cspace.cpp:120-176.  Uploads the occupancy states, runs the exact EDT on
// the device and writes occ_dist back into the caller's cells.
void map_update_cspace(map_t *map, double max_occ_dist)
{
  ApiRange range("amcl_b200:map_update_cspace");
  MapCtx *m = map_ctx(map, true);
  // amclb200_map_set_mirror(map, 0): the device copy is authoritative -- the
  // caller's cells are read once (or never, after amclb200_map_from_image) and
  // cells[].occ_dist is filled only by amclb200_map_pull_distances
  if (m->mirror_eager || !m->occ_uploaded) map_upload_occupancy(m);
  map_build_distance_codes(m, max_occ_dist);
  if (m->mirror_eager)
    map_codes_to_host_cells(m);
  else
    map->max_occ_dist = max_occ_dist;
}

// map_range.c:38-120.  A single ray is a host call in the reference API (the
// node never issues it; the beam model's rays run in the device kernel, see
// sensor_dev.cu).  This walks the caller's own cells with the same integer
// line stepping so isolated queries need no device round trip.
double map_calc_range(map_t *map, double ox, double oy, double oa, double max_range)
{
  int x0 = MAP_GXWX(map, ox), y0 = MAP_GYWY(map, oy);
  int x1 = MAP_GXWX(map, ox + max_range * cos(oa));
  int y1 = MAP_GYWY(map, oy + max_range * sin(oa));
  const bool steep = abs(y1 - y0) > abs(x1 - x0);
  if (steep)
  {
    int t = x0; x0 = y0; y0 = t;
    t = x1; x1 = y1; y1 = t;
  }
  const int dx = abs(x1 - x0), dy = abs(y1 - y0);
  const int xs = x0 < x1 ? 1 : -1, ys = y0 < y1 ? 1 : -1;
  int x = x0, y = y0, err = 0;
  // The reference tests the start cell, then keeps stepping until x has gone
  // one cell PAST x1 (its loop condition is x != x1 + xstep), testing after
  // every step.
  for (;;)
  {
    int ci = steep ? y : x, cj = steep ? x : y;
    if (!MAP_VALID(map, ci, cj) || map->cells[MAP_INDEX(map, ci, cj)].occ_state > -1)
      return sqrt((double)((x - x0) * (x - x0) + (y - y0) * (y - y0))) * map->scale;
    if (x == x1 + xs) break;
    x += xs;
    err += dy;
    if (2 * err >= dx)
    {
      y += ys;
      err -= dx;
    }
  }
  return max_range;
}

}  // extern "C"
