/*
 * amcl_oracle.c -- TEST INFRASTRUCTURE.  CPU restatement ("port") of the AMCL
 * particle-filter hot path of ros-planning/navigation 1.17.3 (package amcl).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this file; the product library never does.
 *
 * Parity status: the reference's own tests hold NO function-level golden
 * vectors for this path, so this restatement is pinned against
 * the compiled, unmodified reference (oracle/_ref/libamcl_ref.so) by
 * tests/test_oracle_vs_reference.py and by the fixtures in tests/golden/
 * (generated from that reference by tests/golden/make_golden.py).
 *
 * Conventions: a sample set is a flat double array of rows
 * (x, y, theta, weight) -- the memory layout of pf_sample_t (pf.h:59-67).
 * A map is (occ int8[size_y*size_x], x fastest) plus origin/scale, cell index
 * i + j*size_x as MAP_INDEX (map.h:144).  Every function names the reference
 * lines it follows (paths under /root/reference/amcl).
 *
 * Two places deliberately restate a *faster, proven-equal* form of a
 * reference loop, because the as-is loop is infeasible at 16 M particles:
 *   - the multinomial draw is a binary search instead of the linear scan
 *     (pf.c:461-465) -- same index for every r;
 *   - KLD bin counting uses a hash set instead of the kd-tree -- same
 *     leaf_count; a kd-tree replay is kept for the cluster label ORDER.
 * And one place restates what north_star asks for instead of what the
 * reference does: orc_edt_exact is an exact Euclidean distance transform,
 * while map_cspace.cpp's wavefront is approximate.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

/* ======================================================================= */
/* drand48-compatible stream (glibc LCG); pf.c:55, pf_pdf.c:68,139-141      */
/* ======================================================================= */

static uint64_t orc_state = 0x1234ABCD330EULL;
static uint64_t orc_draws = 0;

void orc_srand48(long seed)
{
  orc_state = ((((uint64_t)seed) & 0xFFFFFFFFULL) << 16) | 0x330EULL;
}
uint64_t orc_get_state(void) { return orc_state; }
void orc_set_state(uint64_t s) { orc_state = s & 0xFFFFFFFFFFFFULL; }
uint64_t orc_draw_count(void) { return orc_draws; }
void orc_reset_count(void) { orc_draws = 0; }

double orc_drand48(void)
{
  orc_state = (0x5DEECE66DULL * orc_state + 0xBULL) & 0xFFFFFFFFFFFFULL;
  orc_draws++;
  return (double)orc_state * (1.0 / 281474976710656.0);
}

/* pf_pdf.c:133-147 -- polar Box-Muller with rejection. */
double orc_ran_gaussian(double sigma)
{
  double x1, x2, w, r;
  do
  {
    do { r = orc_drand48(); } while (r == 0.0);
    x1 = 2.0 * r - 1.0;
    do { r = orc_drand48(); } while (r == 0.0);
    x2 = 2.0 * r - 1.0;
    w = x1 * x1 + x2 * x2;
  } while (w > 1.0 || w == 0.0);
  return sigma * x2 * sqrt(-2.0 * log(w) / w);
}

/* Same draw, but hands back the accepted (x2, sqrt(-2 ln w / w)) pair so a
 * device kernel can form sigma * x2 * s in the reference's order. */
void orc_ran_gaussian_parts(double *x2_out, double *s_out)
{
  double x1, x2, w, r;
  do
  {
    do { r = orc_drand48(); } while (r == 0.0);
    x1 = 2.0 * r - 1.0;
    do { r = orc_drand48(); } while (r == 0.0);
    x2 = 2.0 * r - 1.0;
    w = x1 * x1 + x2 * x2;
  } while (w > 1.0 || w == 0.0);
  *x2_out = x2;
  *s_out = sqrt(-2.0 * log(w) / w);
}

/* ======================================================================= */
/* small algebra: pf_vector.c:106-116, eig3.c (JAMA tred2 + tql2, n = 3)     */
/* ======================================================================= */

void orc_coord_add(const double a[3], const double b[3], double out[3])
{
  double c2;
  out[0] = b[0] + a[0] * cos(b[2]) - a[1] * sin(b[2]);
  out[1] = b[1] + a[0] * sin(b[2]) + a[1] * cos(b[2]);
  c2 = b[2] + a[2];
  out[2] = atan2(sin(c2), cos(c2));
}

/* Householder tridiagonalisation of a symmetric 3x3 (EISPACK tred2 as
 * published in JAMA); eig3.c:20-130.  V in: matrix, out: accumulated Q. */
static void orc_tred2(double V[3][3], double d[3], double e[3])
{
  const int n = 3;
  int i, j, k;
  for (j = 0; j < n; j++) d[j] = V[n - 1][j];
  for (i = n - 1; i > 0; i--)
  {
    double scale = 0.0, h = 0.0, f, g, hh;
    for (k = 0; k < i; k++) scale = scale + fabs(d[k]);
    if (scale == 0.0)
    {
      e[i] = d[i - 1];
      for (j = 0; j < i; j++)
      {
        d[j] = V[i - 1][j];
        V[i][j] = 0.0;
        V[j][i] = 0.0;
      }
    }
    else
    {
      for (k = 0; k < i; k++)
      {
        d[k] /= scale;
        h += d[k] * d[k];
      }
      f = d[i - 1];
      g = sqrt(h);
      if (f > 0) g = -g;
      e[i] = scale * g;
      h = h - f * g;
      d[i - 1] = f - g;
      for (j = 0; j < i; j++) e[j] = 0.0;
      for (j = 0; j < i; j++)
      {
        f = d[j];
        V[j][i] = f;
        g = e[j] + V[j][j] * f;
        for (k = j + 1; k <= i - 1; k++)
        {
          g += V[k][j] * d[k];
          e[k] += V[k][j] * f;
        }
        e[j] = g;
      }
      f = 0.0;
      for (j = 0; j < i; j++)
      {
        e[j] /= h;
        f += e[j] * d[j];
      }
      hh = f / (h + h);
      for (j = 0; j < i; j++) e[j] -= hh * d[j];
      for (j = 0; j < i; j++)
      {
        f = d[j];
        g = e[j];
        for (k = j; k <= i - 1; k++) V[k][j] -= (f * e[k] + g * d[k]);
        d[j] = V[i - 1][j];
        V[i][j] = 0.0;
      }
    }
    d[i] = h;
  }
  for (i = 0; i < n - 1; i++)
  {
    double h, g;
    V[n - 1][i] = V[i][i];
    V[i][i] = 1.0;
    h = d[i + 1];
    if (h != 0.0)
    {
      for (k = 0; k <= i; k++) d[k] = V[k][i + 1] / h;
      for (j = 0; j <= i; j++)
      {
        g = 0.0;
        for (k = 0; k <= i; k++) g += V[k][i + 1] * V[k][j];
        for (k = 0; k <= i; k++) V[k][j] -= g * d[k];
      }
    }
    for (k = 0; k <= i; k++) V[k][i + 1] = 0.0;
  }
  for (j = 0; j < n; j++)
  {
    d[j] = V[n - 1][j];
    V[n - 1][j] = 0.0;
  }
  V[n - 1][n - 1] = 1.0;
  e[0] = 0.0;
}

static double orc_hypot2(double x, double y) { return sqrt(x * x + y * y); }

/* Implicit-shift QL on the tridiagonal form (EISPACK tql2 / JAMA);
 * eig3.c:134-255, including the final ascending sort. */
static void orc_tql2(double V[3][3], double d[3], double e[3])
{
  const int n = 3;
  int i, j, k, l, m;
  double f = 0.0, tst1 = 0.0;
  const double eps = pow(2.0, -52.0);
  for (i = 1; i < n; i++) e[i - 1] = e[i];
  e[n - 1] = 0.0;
  for (l = 0; l < n; l++)
  {
    double t = fabs(d[l]) + fabs(e[l]);
    if (t > tst1) tst1 = t;
    m = l;
    while (m < n)
    {
      if (fabs(e[m]) <= eps * tst1) break;
      m++;
    }
    if (m > l)
    {
      do
      {
        double g = d[l];
        double p = (d[l + 1] - g) / (2.0 * e[l]);
        double r = orc_hypot2(p, 1.0);
        double dl1, h, c, c2, c3, el1, s, s2;
        if (p < 0) r = -r;
        d[l] = e[l] / (p + r);
        d[l + 1] = e[l] * (p + r);
        dl1 = d[l + 1];
        h = g - d[l];
        for (i = l + 2; i < n; i++) d[i] -= h;
        f = f + h;
        p = d[m];
        c = 1.0; c2 = c; c3 = c;
        el1 = e[l + 1];
        s = 0.0; s2 = 0.0;
        for (i = m - 1; i >= l; i--)
        {
          c3 = c2;
          c2 = c;
          s2 = s;
          g = c * e[i];
          h = c * p;
          r = orc_hypot2(p, e[i]);
          e[i + 1] = s * r;
          s = e[i] / r;
          c = p / r;
          p = c * d[i] - s * g;
          d[i + 1] = h + s * (c * g + s * d[i]);
          for (k = 0; k < n; k++)
          {
            h = V[k][i + 1];
            V[k][i + 1] = s * V[k][i] + c * h;
            V[k][i] = c * V[k][i] - s * h;
          }
        }
        p = -s * s2 * c3 * el1 * e[l] / dl1;
        e[l] = s * p;
        d[l] = c * p;
      } while (fabs(e[l]) > eps * tst1);
    }
    d[l] = d[l] + f;
    e[l] = 0.0;
  }
  for (i = 0; i < n - 1; i++)
  {
    double p = d[i];
    k = i;
    for (j = i + 1; j < n; j++)
      if (d[j] < p) { k = j; p = d[j]; }
    if (k != i)
    {
      d[k] = d[i];
      d[i] = p;
      for (j = 0; j < n; j++)
      {
        p = V[j][i];
        V[j][i] = V[j][k];
        V[j][k] = p;
      }
    }
  }
}

/* eig3.c:257-268 */
void orc_eigen3(const double A[9], double V[9], double d[3])
{
  double Vm[3][3], e[3];
  int i, j;
  for (i = 0; i < 3; i++)
    for (j = 0; j < 3; j++) Vm[i][j] = A[3 * i + j];
  orc_tred2(Vm, d, e);
  orc_tql2(Vm, d, e);
  for (i = 0; i < 3; i++)
    for (j = 0; j < 3; j++) V[3 * i + j] = Vm[i][j];
}

/* pf.c:136-174 + pf_pdf.c:47-127: Gaussian initial cloud.  seed_counter is the
 * value the reference's static pf_pdf_seed takes for this call (it re-seeds
 * with srand48(++pf_pdf_seed), pf_pdf.c:68); pass <= 0 to keep the stream. */
void orc_pf_init(double *samples, int n, const double mean[3], const double cov[9],
                 long seed_counter)
{
  double cr[9], cd[3], ev[3];
  int i, j, k;
  orc_eigen3(cov, cr, ev);
  for (i = 0; i < 3; i++) cd[i] = sqrt(ev[i]);
  if (seed_counter > 0) orc_srand48(seed_counter);
  for (k = 0; k < n; k++)
  {
    double r[3];
    for (i = 0; i < 3; i++) r[i] = orc_ran_gaussian(cd[i]);
    for (i = 0; i < 3; i++)
    {
      double x = mean[i];
      for (j = 0; j < 3; j++) x += cr[3 * i + j] * r[j];
      samples[4 * k + i] = x;
    }
    samples[4 * k + 3] = 1.0 / n;
  }
}

/* ======================================================================= */
/* map: distance transform and ray casting                                   */
/* ======================================================================= */

/* map_cspace.cpp:42 -- int cell_radius_ = max_dist / scale (fp64, truncated). */
int orc_cell_radius(double max_dist, double scale)
{
  int r = max_dist / scale;
  return r;
}

/* Exact squared Euclidean distance (in cells) to the nearest cell with
 * occ_state == +1, for every cell; -1 where it exceeds R^2 (those cells keep
 * max_occ_dist, map_cspace.cpp:102,148).  Separable two-pass form. */
void orc_edt_exact(const int8_t *occ, int sx, int sy, int R, int32_t *d2)
{
  const int INF = R + 1;
  int *g = (int *)malloc(sizeof(int) * (size_t)sx * sy);
  int i, j, dj;
  for (j = 0; j < sy; j++)
  {
    const int8_t *row = occ + (size_t)j * sx;
    int *grow = g + (size_t)j * sx;
    int run = INF;
    for (i = 0; i < sx; i++)
    {
      if (row[i] == 1) run = 0;
      else if (run < INF) run++;
      grow[i] = run;
    }
    run = INF;
    for (i = sx - 1; i >= 0; i--)
    {
      if (row[i] == 1) run = 0;
      else if (run < INF) run++;
      if (run < grow[i]) grow[i] = run;
    }
  }
  for (j = 0; j < sy; j++)
    for (i = 0; i < sx; i++)
    {
      int best = R * R + 1;
      int lo = j - R < 0 ? 0 : j - R, hi = j + R >= sy ? sy - 1 : j + R;
      for (dj = lo; dj <= hi; dj++)
      {
        int gg = g[(size_t)dj * sx + i];
        int v = (dj - j) * (dj - j) + gg * gg;
        if (v < best) best = v;
      }
      d2[(size_t)j * sx + i] = best <= R * R ? best : -1;
    }
  free(g);
}

/* O(cells * obstacles-in-window) check of the same definition; small maps. */
void orc_edt_bruteforce(const int8_t *occ, int sx, int sy, int R, int32_t *d2)
{
  int i, j, a, b;
  for (j = 0; j < sy; j++)
    for (i = 0; i < sx; i++)
    {
      int best = R * R + 1;
      for (b = j - R; b <= j + R; b++)
      {
        if (b < 0 || b >= sy) continue;
        for (a = i - R; a <= i + R; a++)
        {
          int v;
          if (a < 0 || a >= sx) continue;
          if (occ[(size_t)b * sx + a] != 1) continue;
          v = (a - i) * (a - i) + (b - j) * (b - j);
          if (v < best) best = v;
        }
      }
      d2[(size_t)j * sx + i] = best <= R * R ? best : -1;
    }
}

/* occ_dist = sqrt(di^2 + dj^2) * scale (map_cspace.cpp:49,105), else
 * max_occ_dist (:148). */
void orc_d2_to_occ_dist(const int32_t *d2, size_t n, double scale,
                        double max_occ_dist, double *out)
{
  size_t k;
  for (k = 0; k < n; k++)
    out[k] = d2[k] < 0 ? max_occ_dist : sqrt((double)d2[k]) * scale;
}

typedef struct
{
  const int8_t *occ;
  const double *occ_dist; /* may be NULL for the beam model */
  int sx, sy;
  double origin_x, origin_y, scale, max_occ_dist;
} orc_map;

/* map.h:137-138 */
static int orc_gx(const orc_map *m, double x)
{
  return (int)(floor((x - m->origin_x) / m->scale + 0.5) + m->sx / 2);
}
static int orc_gy(const orc_map *m, double y)
{
  return (int)(floor((y - m->origin_y) / m->scale + 0.5) + m->sy / 2);
}
static int orc_valid(const orc_map *m, int i, int j)
{
  return i >= 0 && i < m->sx && j >= 0 && j < m->sy;
}

/* map_range.c:38-120: integer line walk; unknown and off-map cells stop the
 * ray; returns sqrt(dx^2+dy^2)*scale of the stopping cell or max_range. */
static double orc_calc_range_m(const orc_map *m, double ox, double oy, double oa,
                               double max_range)
{
  int x0 = orc_gx(m, ox), y0 = orc_gy(m, oy);
  int x1 = orc_gx(m, ox + max_range * cos(oa));
  int y1 = orc_gy(m, oy + max_range * sin(oa));
  int steep = abs(y1 - y0) > abs(x1 - x0);
  int x, y, xstep, ystep, deltax, deltay, error, deltaerr, tmp;
  if (steep)
  {
    tmp = x0; x0 = y0; y0 = tmp;
    tmp = x1; x1 = y1; y1 = tmp;
  }
  deltax = abs(x1 - x0);
  deltay = abs(y1 - y0);
  error = 0;
  deltaerr = deltay;
  x = x0;
  y = y0;
  xstep = x0 < x1 ? 1 : -1;
  ystep = y0 < y1 ? 1 : -1;
  {
    int ci = steep ? y : x, cj = steep ? x : y;
    if (!orc_valid(m, ci, cj) || m->occ[(size_t)cj * m->sx + ci] > -1)
      return sqrt((double)((x - x0) * (x - x0) + (y - y0) * (y - y0))) * m->scale;
  }
  while (x != (x1 + xstep * 1))
  {
    int ci, cj;
    x += xstep;
    error += deltaerr;
    if (2 * error >= deltax)
    {
      y += ystep;
      error -= deltax;
    }
    ci = steep ? y : x;
    cj = steep ? x : y;
    if (!orc_valid(m, ci, cj) || m->occ[(size_t)cj * m->sx + ci] > -1)
      return sqrt((double)((x - x0) * (x - x0) + (y - y0) * (y - y0))) * m->scale;
  }
  return max_range;
}

double orc_calc_range(const int8_t *occ, int sx, int sy, double origin_x,
                      double origin_y, double scale, double ox, double oy,
                      double oa, double max_range)
{
  orc_map m = {occ, NULL, sx, sy, origin_x, origin_y, scale, 0.0};
  return orc_calc_range_m(&m, ox, oy, oa, max_range);
}

/* ======================================================================= */
/* laser models: amcl_laser.cpp                                              */
/* ======================================================================= */

typedef struct
{
  double laser_pose[3];
  int max_beams;
  double z_hit, z_short, z_max, z_rand, sigma_hit, lambda_short;
  int do_beamskip;
  double beam_skip_distance, beam_skip_threshold, beam_skip_error_threshold;
} orc_laser;

/* amcl_laser.cpp:215-303.  Returns the total weight; weights updated in place. */
double orc_lf_model(double *samples, int n, const orc_laser *L, const orc_map *m,
                    const double *ranges, int range_count, double range_max)
{
  double total = 0.0;
  int i, j;
  for (j = 0; j < n; j++)
  {
    double *s = samples + 4 * j;
    double pose[3], p = 1.0;
    double z_hit_denom = 2 * L->sigma_hit * L->sigma_hit;
    double z_rand_mult = 1.0 / range_max;
    int step = (range_count - 1) / (L->max_beams - 1);
    orc_coord_add(L->laser_pose, s, pose);
    if (step < 1) step = 1;
    for (i = 0; i < range_count; i += step)
    {
      double obs_range = ranges[2 * i], obs_bearing = ranges[2 * i + 1];
      double pz = 0.0, z, hx, hy;
      int mi, mj;
      if (obs_range >= range_max) continue;
      if (obs_range != obs_range) continue;
      hx = pose[0] + obs_range * cos(pose[2] + obs_bearing);
      hy = pose[1] + obs_range * sin(pose[2] + obs_bearing);
      mi = orc_gx(m, hx);
      mj = orc_gy(m, hy);
      if (!orc_valid(m, mi, mj)) z = m->max_occ_dist;
      else z = m->occ_dist[(size_t)mj * m->sx + mi];
      pz += L->z_hit * exp(-(z * z) / z_hit_denom);
      pz += L->z_rand * z_rand_mult;
      p += pz * pz * pz;
    }
    s[3] *= p;
    total += s[3];
  }
  return total;
}

/* amcl_laser.cpp:305-494.  temp_obs is the caller-held n x max_beams scratch
 * (the reference keeps it across calls inside the AMCLLaser object and never
 * clears it, :440,:496-510); it is only touched when beam skipping is active
 * (do_beamskip && converged, :342). */
double orc_lfp_model(double *samples, int n, const orc_laser *L, const orc_map *m,
                     const double *ranges, int range_count, double range_max,
                     int set_converged, double *temp_obs)
{
  double total = 0.0;
  int step = (int)ceil(range_count / (double)L->max_beams);
  double z_hit_denom = 2 * L->sigma_hit * L->sigma_hit;
  double z_rand_mult = 1.0 / range_max;
  double max_dist_prob = exp(-(m->max_occ_dist * m->max_occ_dist) / z_hit_denom);
  int do_beamskip = L->do_beamskip && set_converged;
  int *obs_count = (int *)calloc(L->max_beams, sizeof(int));
  char *obs_mask = (char *)calloc(L->max_beams, 1);
  int i, j, beam_ind = 0;
  if (step < 1) step = 1;

  for (j = 0; j < n; j++)
  {
    double *s = samples + 4 * j;
    double pose[3], log_p = 0;
    orc_coord_add(L->laser_pose, s, pose);
    beam_ind = 0;
    for (i = 0; i < range_count; i += step, beam_ind++)
    {
      double obs_range = ranges[2 * i], obs_bearing = ranges[2 * i + 1];
      double pz = 0.0, z, hx, hy;
      int mi, mj;
      if (obs_range >= range_max) continue;
      if (obs_range != obs_range) continue;
      hx = pose[0] + obs_range * cos(pose[2] + obs_bearing);
      hy = pose[1] + obs_range * sin(pose[2] + obs_bearing);
      mi = orc_gx(m, hx);
      mj = orc_gy(m, hy);
      if (!orc_valid(m, mi, mj))
        pz += L->z_hit * max_dist_prob;
      else
      {
        z = m->occ_dist[(size_t)mj * m->sx + mi];
        if (z < L->beam_skip_distance) obs_count[beam_ind] += 1;
        pz += L->z_hit * exp(-(z * z) / z_hit_denom);
      }
      pz += L->z_rand * z_rand_mult;
      if (!do_beamskip) log_p += log(pz);
      else temp_obs[(size_t)j * L->max_beams + beam_ind] = pz;
    }
    if (!do_beamskip)
    {
      s[3] *= exp(log_p);
      total += s[3];
    }
  }

  if (do_beamskip)
  {
    int skipped = 0, error = 0;
    for (beam_ind = 0; beam_ind < L->max_beams; beam_ind++)
    {
      if ((obs_count[beam_ind] / (double)n) > L->beam_skip_threshold)
        obs_mask[beam_ind] = 1;
      else
      {
        obs_mask[beam_ind] = 0;
        skipped++;
      }
    }
    if (skipped >= (beam_ind * L->beam_skip_error_threshold)) error = 1;
    for (j = 0; j < n; j++)
    {
      double *s = samples + 4 * j;
      double log_p = 0;
      for (beam_ind = 0; beam_ind < L->max_beams; beam_ind++)
        if (error || obs_mask[beam_ind])
          log_p += log(temp_obs[(size_t)j * L->max_beams + beam_ind]);
      s[3] *= exp(log_p);
      total += s[3];
    }
  }
  free(obs_count);
  free(obs_mask);
  return total;
}

/* amcl_laser.cpp:144-213 (no step<1 guard there: the caller must keep
 * range_count-1 >= max_beams-1). */
double orc_beam_model(double *samples, int n, const orc_laser *L, const orc_map *m,
                      const double *ranges, int range_count, double range_max)
{
  double total = 0.0;
  int i, j;
  for (j = 0; j < n; j++)
  {
    double *s = samples + 4 * j;
    double pose[3], p = 1.0;
    int step = (range_count - 1) / (L->max_beams - 1);
    orc_coord_add(L->laser_pose, s, pose);
    for (i = 0; i < range_count; i += step)
    {
      double obs_range = ranges[2 * i], obs_bearing = ranges[2 * i + 1];
      double map_range = orc_calc_range_m(m, pose[0], pose[1], pose[2] + obs_bearing, range_max);
      double pz = 0.0, z = obs_range - map_range;
      pz += L->z_hit * exp(-(z * z) / (2 * L->sigma_hit * L->sigma_hit));
      if (z < 0) pz += L->z_short * L->lambda_short * exp(-L->lambda_short * obs_range);
      if (obs_range == range_max) pz += L->z_max * 1.0;
      if (obs_range < range_max) pz += L->z_rand * 1.0 / range_max;
      p += pz * pz * pz;
    }
    s[3] *= p;
    total += s[3];
  }
  return total;
}

/* pf.c:268-318: normalise, n_effective, slow/fast likelihood averages.
 * scalars = {w_slow, w_fast} in/out; returns n_effective. */
double orc_sensor_finish(double *samples, int n, double total, double alpha_slow,
                         double alpha_fast, double *w_slow, double *w_fast)
{
  double neff = 0.0;
  int i;
  if (total > 0.0)
  {
    double w_avg = 0.0;
    for (i = 0; i < n; i++)
    {
      double *s = samples + 4 * i;
      w_avg += s[3];
      s[3] /= total;
      neff += s[3] * s[3];
    }
    w_avg /= n;
    if (*w_slow == 0.0) *w_slow = w_avg;
    else *w_slow += alpha_slow * (w_avg - *w_slow);
    if (*w_fast == 0.0) *w_fast = w_avg;
    else *w_fast += alpha_fast * (w_avg - *w_fast);
  }
  else
  {
    for (i = 0; i < n; i++) samples[4 * i + 3] = 1.0 / n;
  }
  return 1.0 / neff;
}

/* Flat entry points for FFI (ctypes cannot pass the structs above by value). */
static orc_map orc_mk_map(const int8_t *occ, const double *occ_dist, int sx, int sy,
                          double ox, double oy, double scale, double max_occ_dist)
{
  orc_map m = {occ, occ_dist, sx, sy, ox, oy, scale, max_occ_dist};
  return m;
}

/* laser_params: [lx, ly, la, z_hit, z_short, z_max, z_rand, sigma_hit,
 *                lambda_short, beam_skip_distance, beam_skip_threshold,
 *                beam_skip_error_threshold] */
static orc_laser orc_mk_laser(const double *lp, int max_beams, int do_beamskip)
{
  orc_laser L;
  L.laser_pose[0] = lp[0]; L.laser_pose[1] = lp[1]; L.laser_pose[2] = lp[2];
  L.max_beams = max_beams;
  L.z_hit = lp[3]; L.z_short = lp[4]; L.z_max = lp[5]; L.z_rand = lp[6];
  L.sigma_hit = lp[7]; L.lambda_short = lp[8];
  L.do_beamskip = do_beamskip;
  L.beam_skip_distance = lp[9];
  L.beam_skip_threshold = lp[10];
  L.beam_skip_error_threshold = lp[11];
  return L;
}

/* model: 0 beam, 1 likelihood_field, 2 likelihood_field_prob (laser_model_t,
 * amcl_laser.h:37-42). */
double orc_laser_model(int model, double *samples, int n, const double *laser_params,
                       int max_beams, int do_beamskip, int set_converged,
                       double *temp_obs, const int8_t *occ, const double *occ_dist,
                       int sx, int sy, double origin_x, double origin_y, double scale,
                       double max_occ_dist, const double *ranges, int range_count,
                       double range_max)
{
  orc_map m = orc_mk_map(occ, occ_dist, sx, sy, origin_x, origin_y, scale, max_occ_dist);
  orc_laser L = orc_mk_laser(laser_params, max_beams, do_beamskip);
  if (model == 1) return orc_lf_model(samples, n, &L, &m, ranges, range_count, range_max);
  if (model == 2)
    return orc_lfp_model(samples, n, &L, &m, ranges, range_count, range_max,
                         set_converged, temp_obs);
  return orc_beam_model(samples, n, &L, &m, ranges, range_count, range_max);
}

/* ======================================================================= */
/* odometry motion models: amcl_odom.cpp                                     */
/* ======================================================================= */

static double orc_normalize(double z) { return atan2(sin(z), cos(z)); }

/* amcl_odom.cpp:44-58 */
static double orc_angle_diff(double a, double b)
{
  double d1, d2;
  a = orc_normalize(a);
  b = orc_normalize(b);
  d1 = a - b;
  d2 = 2 * M_PI - fabs(d1);
  if (d1 > 0) d2 *= -1.0;
  if (fabs(d1) < fabs(d2)) return d1;
  return d2;
}

/* amcl_odom.cpp:113-310.  model: 0 DIFF, 1 OMNI, 2 DIFF_CORRECTED,
 * 3 OMNI_CORRECTED (odom_model_t, amcl_odom.h:36-42).  Draws from the orc
 * stream in the reference's order (three Gaussians per particle). */
void orc_motion(double *samples, int n, int model, const double alpha[5],
                const double pose[3], const double delta[3])
{
  double a1 = alpha[0], a2 = alpha[1], a3 = alpha[2], a4 = alpha[3], a5 = alpha[4];
  double old_pose2 = pose[2] - delta[2];
  int i;
  if (model == 1 || model == 3)
  {
    double delta_trans = sqrt(delta[0] * delta[0] + delta[1] * delta[1]);
    double delta_rot = delta[2];
    double trans_sd, rot_sd, strafe_sd;
    if (model == 1)
    {
      trans_sd = (a3 * (delta_trans * delta_trans) + a1 * (delta_rot * delta_rot));
      rot_sd = (a4 * (delta_rot * delta_rot) + a2 * (delta_trans * delta_trans));
      strafe_sd = (a1 * (delta_rot * delta_rot) + a5 * (delta_trans * delta_trans));
    }
    else
    {
      trans_sd = sqrt(a3 * (delta_trans * delta_trans) + a4 * (delta_rot * delta_rot));
      rot_sd = sqrt(a1 * (delta_rot * delta_rot) + a2 * (delta_trans * delta_trans));
      strafe_sd = sqrt(a4 * (delta_rot * delta_rot) + a5 * (delta_trans * delta_trans));
    }
    for (i = 0; i < n; i++)
    {
      double *s = samples + 4 * i;
      double bearing = orc_angle_diff(atan2(delta[1], delta[0]), old_pose2) + s[2];
      double cs = cos(bearing), sn = sin(bearing);
      double trans_hat = delta_trans + orc_ran_gaussian(trans_sd);
      double rot_hat = delta_rot + orc_ran_gaussian(rot_sd);
      double strafe_hat = 0 + orc_ran_gaussian(strafe_sd);
      s[0] += (trans_hat * cs + strafe_hat * sn);
      s[1] += (trans_hat * sn - strafe_hat * cs);
      s[2] += rot_hat;
    }
  }
  else
  {
    double delta_rot1, delta_trans, delta_rot2, rot1_noise, rot2_noise;
    double n1a, n1b, n2a, n2b;
    if (sqrt(delta[1] * delta[1] + delta[0] * delta[0]) < 0.01)
      delta_rot1 = 0.0;
    else
      delta_rot1 = orc_angle_diff(atan2(delta[1], delta[0]), old_pose2);
    delta_trans = sqrt(delta[0] * delta[0] + delta[1] * delta[1]);
    delta_rot2 = orc_angle_diff(delta[2], delta_rot1);
    n1a = fabs(orc_angle_diff(delta_rot1, 0.0));
    n1b = fabs(orc_angle_diff(delta_rot1, M_PI));
    rot1_noise = n1b < n1a ? n1b : n1a;   /* std::min(a, b) */
    n2a = fabs(orc_angle_diff(delta_rot2, 0.0));
    n2b = fabs(orc_angle_diff(delta_rot2, M_PI));
    rot2_noise = n2b < n2a ? n2b : n2a;
    for (i = 0; i < n; i++)
    {
      double *s = samples + 4 * i;
      double v1 = a1 * rot1_noise * rot1_noise + a2 * delta_trans * delta_trans;
      double v2 = a3 * delta_trans * delta_trans + a4 * rot1_noise * rot1_noise +
                  a4 * rot2_noise * rot2_noise;
      double v3 = a1 * rot2_noise * rot2_noise + a2 * delta_trans * delta_trans;
      double rot1_hat, trans_hat, rot2_hat;
      if (model == 2) { v1 = sqrt(v1); v2 = sqrt(v2); v3 = sqrt(v3); }
      rot1_hat = orc_angle_diff(delta_rot1, orc_ran_gaussian(v1));
      trans_hat = delta_trans - orc_ran_gaussian(v2);
      rot2_hat = orc_angle_diff(delta_rot2, orc_ran_gaussian(v3));
      s[0] += trans_hat * cos(s[2] + rot1_hat);
      s[1] += trans_hat * sin(s[2] + rot1_hat);
      s[2] += rot1_hat + rot2_hat;
    }
  }
}

/* ======================================================================= */
/* KLD resampling: pf.c:363-555                                              */
/* ======================================================================= */

/* pf.c:515-555 without the memo table (the cached value is the same number). */
int orc_resample_limit(int k, int min_samples, int max_samples, double pop_err,
                       double pop_z)
{
  double a, b, c, x;
  int n;
  if (k < 1 || k > max_samples) return max_samples;
  if (k == 1) return max_samples;
  a = 1;
  b = 2 / (9 * ((double)k - 1));
  c = sqrt(2 / (9 * ((double)k - 1))) * pop_z;
  x = a - b + c;
  n = (int)ceil((k - 1) / (2 * pop_err) * x * x * x);
  if (n < min_samples) return min_samples;
  if (n > max_samples) return max_samples;
  return n;
}

/* pf_kdtree.c:72-74,116-118 */
static const double ORC_BIN[3] = {0.50, 0.50, (10 * M_PI / 180)};

void orc_bin_key(const double pose[3], int key[3])
{
  key[0] = (int)floor(pose[0] / ORC_BIN[0]);
  key[1] = (int)floor(pose[1] / ORC_BIN[1]);
  key[2] = (int)floor(pose[2] / ORC_BIN[2]);
}

/* open-addressing set of 3-int keys */
typedef struct
{
  int *keys;       /* 3 ints per slot */
  int *val;        /* payload per slot (first index / bin id) */
  char *used;
  size_t cap;      /* power of two */
  size_t count;
} orc_keyset;

static void orc_keyset_init(orc_keyset *s, size_t want)
{
  size_t cap = 64;
  while (cap < 2 * want + 16) cap <<= 1;
  s->cap = cap;
  s->count = 0;
  s->keys = (int *)malloc(sizeof(int) * 3 * cap);
  s->val = (int *)malloc(sizeof(int) * cap);
  s->used = (char *)calloc(cap, 1);
}
static void orc_keyset_free(orc_keyset *s)
{
  free(s->keys); free(s->val); free(s->used);
}
static size_t orc_keyhash(const int k[3])
{
  uint64_t h = (uint64_t)(uint32_t)k[0] * 0x9E3779B97F4A7C15ULL;
  h ^= ((uint64_t)(uint32_t)k[1] + 0x7F4A7C15ULL) * 0xC2B2AE3D27D4EB4FULL;
  h ^= ((uint64_t)(uint32_t)k[2] + 0x165667B1ULL) * 0x94D049BB133111EBULL;
  h ^= h >> 29;
  return (size_t)h;
}
/* returns slot; *is_new set when the key was absent (and is inserted) */
static size_t orc_keyset_insert(orc_keyset *s, const int k[3], int *is_new)
{
  size_t p = orc_keyhash(k) & (s->cap - 1);
  for (;;)
  {
    if (!s->used[p])
    {
      s->used[p] = 1;
      s->keys[3 * p] = k[0]; s->keys[3 * p + 1] = k[1]; s->keys[3 * p + 2] = k[2];
      s->count++;
      *is_new = 1;
      return p;
    }
    if (s->keys[3 * p] == k[0] && s->keys[3 * p + 1] == k[1] && s->keys[3 * p + 2] == k[2])
    {
      *is_new = 0;
      return p;
    }
    p = (p + 1) & (s->cap - 1);
  }
}
static long orc_keyset_find(const orc_keyset *s, const int k[3])
{
  size_t p = orc_keyhash(k) & (s->cap - 1);
  for (;;)
  {
    if (!s->used[p]) return -1;
    if (s->keys[3 * p] == k[0] && s->keys[3 * p + 1] == k[1] && s->keys[3 * p + 2] == k[2])
      return (long)p;
    p = (p + 1) & (s->cap - 1);
  }
}

typedef void (*orc_pose_fn)(void *arg, double out[3]);

/* pf.c:396-498, non-selective branch.  samples_a: n_a rows; samples_b gets the
 * new set (weights 1/n); src_index[m] = source row of output m, or -1 when it
 * was a random pose (w_diff branch, pf.c:430-431).  The multinomial search is
 * restated as "number of table entries c[1..n_a] that are <= r", which is the
 * index the linear scan at pf.c:461-465 stops at.  A draw at or beyond the
 * last table entry (possible only through rounding of the total) is clamped
 * to the last row; the reference asserts there.
 * Returns n; *leaf_count_out = occupied bins. */
int orc_resample(const double *samples_a, int n_a, double *samples_b, int *src_index,
                 int min_samples, int max_samples, double pop_err, double pop_z,
                 double w_diff, orc_pose_fn random_pose, void *random_pose_arg,
                 int *leaf_count_out)
{
  double *c = (double *)malloc(sizeof(double) * ((size_t)n_a + 1));
  orc_keyset bins;
  int count = 0, i;
  c[0] = 0.0;
  for (i = 0; i < n_a; i++) c[i + 1] = c[i] + samples_a[4 * i + 3];
  orc_keyset_init(&bins, (size_t)max_samples);
  while (count < max_samples)
  {
    double *sb = samples_b + 4 * (size_t)count;
    int key[3], is_new;
    if (orc_drand48() < w_diff)
    {
      random_pose(random_pose_arg, sb);
      src_index[count] = -1;
    }
    else
    {
      double r = orc_drand48();
      int lo = 0, hi = n_a; /* count of c[1..n_a] <= r */
      while (lo < hi)
      {
        int mid = lo + (hi - lo) / 2;
        if (c[mid + 1] <= r) lo = mid + 1;
        else hi = mid;
      }
      if (lo >= n_a) lo = n_a - 1;
      sb[0] = samples_a[4 * lo]; sb[1] = samples_a[4 * lo + 1]; sb[2] = samples_a[4 * lo + 2];
      src_index[count] = lo;
    }
    sb[3] = 1.0;
    count++;
    orc_bin_key(sb, key);
    orc_keyset_insert(&bins, key, &is_new);
    if (count > orc_resample_limit((int)bins.count, min_samples, max_samples, pop_err, pop_z))
      break;
  }
  for (i = 0; i < count; i++) samples_b[4 * (size_t)i + 3] /= (double)count;
  *leaf_count_out = (int)bins.count;
  orc_keyset_free(&bins);
  free(c);
  return count;
}

/* Sequential fp64 cumulative table alone (pf.c:399-402), for the
 * "binade-integer scan" parity gate. */
void orc_cumsum(const double *samples, int n, double *c)
{
  int i;
  c[0] = 0.0;
  for (i = 0; i < n; i++) c[i + 1] = c[i] + samples[4 * i + 3];
}

/* ======================================================================= */
/* clustering and statistics: pf_kdtree.c, pf.c:559-693                      */
/* ======================================================================= */

/* Index-based replay of the reference's tree (pf_kdtree.c:215-300), kept only
 * because the cluster label ORDER is defined by node allocation order
 * (pf_kdtree.c:368-387). */
typedef struct
{
  int leaf, pivot_dim;
  double pivot_value;
  int key[3];
  int cluster;
  int child[2];
} orc_node;

typedef struct
{
  orc_node *nodes;
  int count, cap, root, leaf_count;
} orc_tree;

static int orc_tree_new_leaf(orc_tree *t, const int key[3])
{
  orc_node *nd;
  if (t->count == t->cap)
  {
    t->cap = t->cap ? 2 * t->cap : 1024;
    t->nodes = (orc_node *)realloc(t->nodes, sizeof(orc_node) * (size_t)t->cap);
  }
  nd = t->nodes + t->count;
  memset(nd, 0, sizeof(*nd));
  nd->leaf = 1;
  nd->key[0] = key[0]; nd->key[1] = key[1]; nd->key[2] = key[2];
  nd->child[0] = nd->child[1] = -1;
  nd->cluster = -1;
  t->leaf_count++;
  return t->count++;
}

static void orc_tree_insert(orc_tree *t, const int key[3])
{
  int cur, i;
  if (t->root < 0)
  {
    t->root = orc_tree_new_leaf(t, key);
    return;
  }
  cur = t->root;
  for (;;)
  {
    orc_node *nd = t->nodes + cur;
    if (nd->leaf)
    {
      int max_split = 0, split, c0, c1, okey[3];
      if (nd->key[0] == key[0] && nd->key[1] == key[1] && nd->key[2] == key[2]) return;
      nd->pivot_dim = -1;
      for (i = 0; i < 3; i++)
      {
        split = abs(key[i] - nd->key[i]);
        if (split > max_split) { max_split = split; nd->pivot_dim = i; }
      }
      nd->pivot_value = (key[nd->pivot_dim] + nd->key[nd->pivot_dim]) / 2.0;
      okey[0] = nd->key[0]; okey[1] = nd->key[1]; okey[2] = nd->key[2];
      /* child[0] is allocated before child[1], whichever holds the new key */
      if (key[nd->pivot_dim] < nd->pivot_value)
      {
        c0 = orc_tree_new_leaf(t, key);
        c1 = orc_tree_new_leaf(t, okey);
      }
      else
      {
        c0 = orc_tree_new_leaf(t, okey);
        c1 = orc_tree_new_leaf(t, key);
      }
      nd = t->nodes + cur; /* realloc may have moved the array */
      nd->child[0] = c0;
      nd->child[1] = c1;
      nd->leaf = 0;
      t->leaf_count--;
      return;
    }
    cur = key[nd->pivot_dim] < nd->pivot_value ? nd->child[0] : nd->child[1];
  }
}

static int orc_tree_find(const orc_tree *t, const int key[3])
{
  int cur = t->root;
  if (cur < 0) return -1;
  for (;;)
  {
    const orc_node *nd = t->nodes + cur;
    if (nd->leaf)
      return (nd->key[0] == key[0] && nd->key[1] == key[1] && nd->key[2] == key[2]) ? cur : -1;
    cur = key[nd->pivot_dim] < nd->pivot_value ? nd->child[0] : nd->child[1];
  }
}

/* pf_kdtree.c:358-439 with an explicit stack instead of recursion (labels do
 * not depend on the visiting order inside one component). */
static int orc_tree_cluster(orc_tree *t)
{
  int *queue = (int *)malloc(sizeof(int) * (size_t)(t->count ? t->count : 1));
  int *stack = (int *)malloc(sizeof(int) * (size_t)(t->count ? t->count : 1));
  int qn = 0, clusters = 0, i;
  for (i = 0; i < t->count; i++)
    if (t->nodes[i].leaf)
    {
      t->nodes[i].cluster = -1;
      queue[qn++] = i;
    }
  while (qn > 0)
  {
    int seed = queue[--qn], sp = 0;
    if (t->nodes[seed].cluster >= 0) continue;
    t->nodes[seed].cluster = clusters++;
    stack[sp++] = seed;
    while (sp > 0)
    {
      int cur = stack[--sp];
      for (i = 0; i < 27; i++)
      {
        int nkey[3], nn;
        nkey[0] = t->nodes[cur].key[0] + (i / 9) - 1;
        nkey[1] = t->nodes[cur].key[1] + ((i % 9) / 3) - 1;
        nkey[2] = t->nodes[cur].key[2] + ((i % 9) % 3) - 1;
        nn = orc_tree_find(t, nkey);
        if (nn < 0 || t->nodes[nn].cluster >= 0) continue;
        t->nodes[nn].cluster = t->nodes[cur].cluster;
        stack[sp++] = nn;
      }
    }
  }
  free(queue);
  free(stack);
  return clusters;
}

/* Labels every sample with the reference's cluster number.  labels[n] out;
 * returns the cluster count; *leaf_count_out = occupied bins. */
int orc_cluster_labels(const double *samples, int n, int *labels, int *leaf_count_out)
{
  orc_tree t;
  int i, clusters;
  memset(&t, 0, sizeof(t));
  t.root = -1;
  for (i = 0; i < n; i++)
  {
    int key[3];
    orc_bin_key(samples + 4 * (size_t)i, key);
    orc_tree_insert(&t, key);
  }
  clusters = orc_tree_cluster(&t);
  for (i = 0; i < n; i++)
  {
    int key[3], nd;
    orc_bin_key(samples + 4 * (size_t)i, key);
    nd = orc_tree_find(&t, key);
    labels[i] = nd < 0 ? -1 : t.nodes[nd].cluster;
  }
  if (leaf_count_out) *leaf_count_out = t.leaf_count;
  free(t.nodes);
  return clusters;
}

/* pf.c:559-693.  cluster_out: per cluster 15 doubles
 * [count, weight, mean(3), cov 00 01 10 11 22, m(4)... see below];
 * layout: count, weight, mean0, mean1, mean2, cov00, cov01, cov10, cov11, cov22,
 *         m0, m1, m2, m3, c00, c01, c10, c11  (18 doubles).
 * set_out: mean(3), cov00, cov01, cov10, cov11, cov22 (8 doubles). */
void orc_cluster_stats(const double *samples, int n, const int *labels,
                       int cluster_count, double *cluster_out, double *set_out)
{
  double m[4] = {0, 0, 0, 0}, c[2][2] = {{0, 0}, {0, 0}}, weight = 0.0;
  int i, j, k;
  memset(cluster_out, 0, sizeof(double) * 18 * (size_t)(cluster_count > 0 ? cluster_count : 1));
  for (i = 0; i < n; i++)
  {
    const double *s = samples + 4 * (size_t)i;
    double *co;
    if (labels[i] < 0 || labels[i] >= cluster_count) continue;
    co = cluster_out + 18 * (size_t)labels[i];
    co[0] += 1;
    co[1] += s[3];
    weight += s[3];
    co[10] += s[3] * s[0];
    co[11] += s[3] * s[1];
    co[12] += s[3] * cos(s[2]);
    co[13] += s[3] * sin(s[2]);
    m[0] += s[3] * s[0];
    m[1] += s[3] * s[1];
    m[2] += s[3] * cos(s[2]);
    m[3] += s[3] * sin(s[2]);
    for (j = 0; j < 2; j++)
      for (k = 0; k < 2; k++)
      {
        co[14 + 2 * j + k] += s[3] * s[j] * s[k];
        c[j][k] += s[3] * s[j] * s[k];
      }
  }
  for (i = 0; i < cluster_count; i++)
  {
    double *co = cluster_out + 18 * (size_t)i;
    co[2] = co[10] / co[1];
    co[3] = co[11] / co[1];
    co[4] = atan2(co[13], co[12]);
    for (j = 0; j < 2; j++)
      for (k = 0; k < 2; k++)
        co[5 + 2 * j + k] = co[14 + 2 * j + k] / co[1] - co[2 + j] * co[2 + k];
    co[9] = -2 * log(sqrt(co[12] * co[12] + co[13] * co[13]) / co[1]);
  }
  set_out[0] = m[0] / weight;
  set_out[1] = m[1] / weight;
  set_out[2] = atan2(m[3], m[2]);
  for (j = 0; j < 2; j++)
    for (k = 0; k < 2; k++)
      set_out[3 + 2 * j + k] = c[j][k] / weight - set_out[j] * set_out[k];
  set_out[7] = -2 * log(sqrt(m[2] * m[2] + m[3] * m[3]));
}

/* pf.c:220-251 */
int orc_update_converged(const double *samples, int n, double dist_threshold)
{
  double mean_x = 0, mean_y = 0;
  int i;
  for (i = 0; i < n; i++)
  {
    mean_x += samples[4 * (size_t)i];
    mean_y += samples[4 * (size_t)i + 1];
  }
  mean_x /= n;
  mean_y /= n;
  for (i = 0; i < n; i++)
    if (fabs(samples[4 * (size_t)i] - mean_x) > dist_threshold ||
        fabs(samples[4 * (size_t)i + 1] - mean_y) > dist_threshold)
      return 0;
  return 1;
}

/* pf.c:701-740.  out = mean_x, mean_y, var */
void orc_cep_stats(const double *samples, int n, double out[3])
{
  double mn = 0, mx = 0, my = 0, mrr = 0;
  int i;
  for (i = 0; i < n; i++)
  {
    const double *s = samples + 4 * (size_t)i;
    mn += s[3];
    mx += s[3] * s[0];
    my += s[3] * s[1];
    mrr += s[3] * s[0] * s[0];
    mrr += s[3] * s[1] * s[1];
  }
  out[0] = mx / mn;
  out[1] = my / mn;
  out[2] = mrr / mn - (mx * mx / (mn * mn) + my * my / (mn * mn));
}

/* Partition check helper: number of occupied bins and, per sample, a dense id
 * of its bin in first-occurrence order (the GPU's canonical numbering). */
int orc_bin_first_occurrence(const double *samples, int n, int *bin_id)
{
  orc_keyset bins;
  int i;
  orc_keyset_init(&bins, (size_t)n);
  for (i = 0; i < n; i++)
  {
    int key[3], is_new;
    size_t slot;
    orc_bin_key(samples + 4 * (size_t)i, key);
    slot = orc_keyset_insert(&bins, key, &is_new);
    if (is_new) bins.val[slot] = (int)bins.count - 1;
    bin_id[i] = bins.val[slot];
  }
  i = (int)bins.count;
  orc_keyset_free(&bins);
  return i;
}

/* silence -Wunused for the lookup helper kept for tests */
long orc_keyset_probe_selftest(void)
{
  orc_keyset s;
  int k[3] = {1, 2, 3}, is_new;
  long r;
  orc_keyset_init(&s, 4);
  orc_keyset_insert(&s, k, &is_new);
  r = orc_keyset_find(&s, k);
  orc_keyset_free(&s);
  return r;
}
