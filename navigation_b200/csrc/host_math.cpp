// host_math.cpp -- the small host-side helpers the reference exports next to the
// filter: 3-vector / 3x3 algebra (pf_vector.c), the symmetric 3x3 eigen-solver
// (eig3.c), the Gaussian pdf object (pf_pdf.c) and the pf_kdtree_* histogram
// API (pf_kdtree.c).  None of it is per-particle work: it runs a handful of
// times per filter initialisation.  Paths are under /root/reference/amcl/src/amcl/pf/.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <unordered_map>
#include <vector>

#include "amcl/pf/pf.h"
#include "amcl/pf/pf_pdf.h"
#include "amcl/pf/pf_kdtree.h"

extern "C" {

// ------------------------------------------------------- pf_vector.c:36-276 --

pf_vector_t pf_vector_zero(void)
{
  pf_vector_t c = {{0.0, 0.0, 0.0}};
  return c;
}

int pf_vector_finite(pf_vector_t a)
{
  return isfinite(a.v[0]) && isfinite(a.v[1]) && isfinite(a.v[2]) ? 1 : 0;
}

void pf_vector_fprintf(pf_vector_t a, FILE *file, const char *fmt)
{
  for (int i = 0; i < 3; i++)
  {
    fprintf(file, fmt, a.v[i]);
    fprintf(file, " ");
  }
  fprintf(file, "\n");
}

pf_vector_t pf_vector_add(pf_vector_t a, pf_vector_t b)
{
  pf_vector_t c = {{a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2]}};
  return c;
}

pf_vector_t pf_vector_sub(pf_vector_t a, pf_vector_t b)
{
  pf_vector_t c = {{a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2]}};
  return c;
}

// pf_vector.c:106-116: rotate a by b's heading, translate, wrap the heading
pf_vector_t pf_vector_coord_add(pf_vector_t a, pf_vector_t b)
{
  pf_vector_t c;
  c.v[0] = b.v[0] + a.v[0] * cos(b.v[2]) - a.v[1] * sin(b.v[2]);
  c.v[1] = b.v[1] + a.v[0] * sin(b.v[2]) + a.v[1] * cos(b.v[2]);
  c.v[2] = b.v[2] + a.v[2];
  c.v[2] = atan2(sin(c.v[2]), cos(c.v[2]));
  return c;
}

// pf_vector.c:120-130
pf_vector_t pf_vector_coord_sub(pf_vector_t a, pf_vector_t b)
{
  pf_vector_t c;
  c.v[0] = +(a.v[0] - b.v[0]) * cos(b.v[2]) + (a.v[1] - b.v[1]) * sin(b.v[2]);
  c.v[1] = -(a.v[0] - b.v[0]) * sin(b.v[2]) + (a.v[1] - b.v[1]) * cos(b.v[2]);
  c.v[2] = a.v[2] - b.v[2];
  c.v[2] = atan2(sin(c.v[2]), cos(c.v[2]));
  return c;
}

pf_matrix_t pf_matrix_zero(void)
{
  pf_matrix_t c;
  memset(&c, 0, sizeof(c));
  return c;
}

int pf_matrix_finite(pf_matrix_t a)
{
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      if (!isfinite(a.m[i][j])) return 0;
  return 1;
}

void pf_matrix_fprintf(pf_matrix_t a, FILE *file, const char *fmt)
{
  for (int i = 0; i < 3; i++)
  {
    for (int j = 0; j < 3; j++)
    {
      fprintf(file, fmt, a.m[i][j]);
      fprintf(file, " ");
    }
    fprintf(file, "\n");
  }
}

// pf_vector.c:222-275
void pf_matrix_unitary(pf_matrix_t *r, pf_matrix_t *d, pf_matrix_t a)
{
  double aa[3][3], eval[3], evec[3][3];
  memcpy(aa, a.m, sizeof(aa));
  eigen_decomposition(aa, evec, eval);
  *d = pf_matrix_zero();
  for (int i = 0; i < 3; i++)
  {
    d->m[i][i] = eval[i];
    for (int j = 0; j < 3; j++) r->m[i][j] = evec[i][j];
  }
}

// ------------------------------------------------------------------- eig3.c --
//
// Symmetric 3x3 eigen-decomposition by Householder tridiagonalisation followed
// by implicit-shift QL -- the EISPACK tred2/tql2 pair as published in JAMA,
// which is what eig3.c:20-268 uses.  The operation ORDER is kept, because
// pf_init multiplies Gaussian draws by these eigenvectors: a different (equally
// valid) eigenbasis would give different initial particles for the same seed.

namespace
{

struct Sym3
{
  double V[3][3];
  double d[3];
  double e[3];
};

void tridiagonalise(Sym3 &s)
{
  double(&V)[3][3] = s.V;
  double *d = s.d, *e = s.e;
  for (int j = 0; j < 3; j++) d[j] = V[2][j];
  for (int i = 2; i > 0; i--)
  {
    double scale = 0.0, h = 0.0;
    for (int k = 0; k < i; k++) scale = scale + fabs(d[k]);
    if (scale == 0.0)
    {
      e[i] = d[i - 1];
      for (int j = 0; j < i; j++)
      {
        d[j] = V[i - 1][j];
        V[i][j] = 0.0;
        V[j][i] = 0.0;
      }
    }
    else
    {
      for (int k = 0; k < i; k++)
      {
        d[k] /= scale;
        h += d[k] * d[k];
      }
      double f = d[i - 1];
      double g = sqrt(h);
      if (f > 0) g = -g;
      e[i] = scale * g;
      h = h - f * g;
      d[i - 1] = f - g;
      for (int j = 0; j < i; j++) e[j] = 0.0;
      for (int j = 0; j < i; j++)
      {
        f = d[j];
        V[j][i] = f;
        g = e[j] + V[j][j] * f;
        for (int k = j + 1; k <= i - 1; k++)
        {
          g += V[k][j] * d[k];
          e[k] += V[k][j] * f;
        }
        e[j] = g;
      }
      f = 0.0;
      for (int j = 0; j < i; j++)
      {
        e[j] /= h;
        f += e[j] * d[j];
      }
      const double hh = f / (h + h);
      for (int j = 0; j < i; j++) e[j] -= hh * d[j];
      for (int j = 0; j < i; j++)
      {
        f = d[j];
        g = e[j];
        for (int k = j; k <= i - 1; k++) V[k][j] -= (f * e[k] + g * d[k]);
        d[j] = V[i - 1][j];
        V[i][j] = 0.0;
      }
    }
    d[i] = h;
  }
  // accumulate the Householder reflections into V
  for (int i = 0; i < 2; i++)
  {
    V[2][i] = V[i][i];
    V[i][i] = 1.0;
    const double h = d[i + 1];
    if (h != 0.0)
    {
      for (int k = 0; k <= i; k++) d[k] = V[k][i + 1] / h;
      for (int j = 0; j <= i; j++)
      {
        double g = 0.0;
        for (int k = 0; k <= i; k++) g += V[k][i + 1] * V[k][j];
        for (int k = 0; k <= i; k++) V[k][j] -= g * d[k];
      }
    }
    for (int k = 0; k <= i; k++) V[k][i + 1] = 0.0;
  }
  for (int j = 0; j < 3; j++)
  {
    d[j] = V[2][j];
    V[2][j] = 0.0;
  }
  V[2][2] = 1.0;
  e[0] = 0.0;
}

inline double pythag(double x, double y) { return sqrt(x * x + y * y); }

void ql_implicit(Sym3 &s)
{
  double(&V)[3][3] = s.V;
  double *d = s.d, *e = s.e;
  for (int i = 1; i < 3; i++) e[i - 1] = e[i];
  e[2] = 0.0;
  double f = 0.0, tst1 = 0.0;
  const double eps = pow(2.0, -52.0);
  for (int l = 0; l < 3; l++)
  {
    const double mag = fabs(d[l]) + fabs(e[l]);
    if (mag > tst1) tst1 = mag;
    int m = l;
    while (m < 3)
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
        double r = pythag(p, 1.0);
        if (p < 0) r = -r;
        d[l] = e[l] / (p + r);
        d[l + 1] = e[l] * (p + r);
        const double dl1 = d[l + 1];
        double h = g - d[l];
        for (int i = l + 2; i < 3; i++) d[i] -= h;
        f = f + h;
        p = d[m];
        double c = 1.0, c2 = c, c3 = c;
        const double el1 = e[l + 1];
        double sn = 0.0, sn2 = 0.0;
        for (int i = m - 1; i >= l; i--)
        {
          c3 = c2;
          c2 = c;
          sn2 = sn;
          g = c * e[i];
          h = c * p;
          r = pythag(p, e[i]);
          e[i + 1] = sn * r;
          sn = e[i] / r;
          c = p / r;
          p = c * d[i] - sn * g;
          d[i + 1] = h + sn * (c * g + sn * d[i]);
          for (int k = 0; k < 3; k++)
          {
            h = V[k][i + 1];
            V[k][i + 1] = sn * V[k][i] + c * h;
            V[k][i] = c * V[k][i] - sn * h;
          }
        }
        p = -sn * sn2 * c3 * el1 * e[l] / dl1;
        e[l] = sn * p;
        d[l] = c * p;
      } while (fabs(e[l]) > eps * tst1);
    }
    d[l] = d[l] + f;
    e[l] = 0.0;
  }
  // ascending eigenvalues, columns follow
  for (int i = 0; i < 2; i++)
  {
    int k = i;
    double p = d[i];
    for (int j = i + 1; j < 3; j++)
      if (d[j] < p)
      {
        k = j;
        p = d[j];
      }
    if (k != i)
    {
      d[k] = d[i];
      d[i] = p;
      for (int j = 0; j < 3; j++)
      {
        const double t = V[j][i];
        V[j][i] = V[j][k];
        V[j][k] = t;
      }
    }
  }
}

}  // namespace

// eig3.c:257-268
void eigen_decomposition(double A[3][3], double V[3][3], double d[3])
{
  Sym3 s;
  memcpy(s.V, A, sizeof(s.V));
  tridiagonalise(s);
  ql_implicit(s);
  memcpy(V, s.V, sizeof(s.V));
  memcpy(d, s.d, sizeof(s.d));
}

// ----------------------------------------------------------- pf_pdf.c:40-147 --

static unsigned int g_pdf_seed = 0;   // pf_pdf.c:38
static int g_reseed = 1;

// amclb200_set_reseed(0): the library stops re-seeding the process-wide drand48
// stream (pf_alloc's srand48(time), pf_pdf_gaussian_alloc's srand48(++counter)),
// so a caller that seeds the stream itself gets reproducible runs.
void amclb200_set_reseed(int on) { g_reseed = on != 0; }
int amclb200_get_reseed(void) { return g_reseed; }

pf_pdf_gaussian_t *pf_pdf_gaussian_alloc(pf_vector_t x, pf_matrix_t cx)
{
  pf_pdf_gaussian_t *pdf = (pf_pdf_gaussian_t *)calloc(1, sizeof(pf_pdf_gaussian_t));
  pf_matrix_t cd;
  pdf->x = x;
  pdf->cx = cx;
  pf_matrix_unitary(&pdf->cr, &cd, pdf->cx);
  pdf->cd.v[0] = sqrt(cd.m[0][0]);
  pdf->cd.v[1] = sqrt(cd.m[1][1]);
  pdf->cd.v[2] = sqrt(cd.m[2][2]);
  // The reference re-seeds the global stream with a per-process counter here
  // (pf_pdf.c:68), which makes pf_init deterministic; kept for that reason.
  ++g_pdf_seed;
  if (g_reseed) srand48(g_pdf_seed);
  return pdf;
}

void pf_pdf_gaussian_free(pf_pdf_gaussian_t *pdf) { free(pdf); }

// pf_pdf.c:133-147
double pf_ran_gaussian(double sigma)
{
  double x1, x2, w, r;
  do
  {
    do { r = drand48(); } while (r == 0.0);
    x1 = 2.0 * r - 1.0;
    do { r = drand48(); } while (r == 0.0);
    x2 = 2.0 * r - 1.0;
    w = x1 * x1 + x2 * x2;
  } while (w > 1.0 || w == 0.0);
  return (sigma * x2 * sqrt(-2.0 * log(w) / w));
}

// pf_pdf.c:105-127
pf_vector_t pf_pdf_gaussian_sample(pf_pdf_gaussian_t *pdf)
{
  pf_vector_t r, x;
  for (int i = 0; i < 3; i++) r.v[i] = pf_ran_gaussian(pdf->cd.v[i]);
  for (int i = 0; i < 3; i++)
  {
    x.v[i] = pdf->x.v[i];
    for (int j = 0; j < 3; j++) x.v[i] += pdf->cr.m[i][j] * r.v[j];
  }
  return x;
}

// -------------------------------------------------- pf_kdtree.c (host facade) --
//
// Inside the filter the histogram is a device hash table (resample_dev.cu).
// These entry points keep the reference's standalone API alive for C callers:
// a host-side bin map hung off pf_kdtree_t::nodes, with the same bin keys
// (pf_kdtree.c:116-118), the same 26-neighbour clustering (:407-438) and
// canonical cluster numbers (order of first insertion).

namespace
{

struct BinKey
{
  int k[3];
  bool operator==(const BinKey &o) const { return k[0] == o.k[0] && k[1] == o.k[1] && k[2] == o.k[2]; }
};
struct BinHash
{
  size_t operator()(const BinKey &b) const
  {
    size_t h = (size_t)(unsigned)b.k[0] * 0x9E3779B97F4A7C15ull;
    h ^= ((size_t)(unsigned)b.k[1] + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
    h ^= ((size_t)(unsigned)b.k[2] + 0x165667B1ull) * 0x94D049BB133111EBull;
    return h ^ (h >> 29);
  }
};
struct Bin
{
  double value;
  int cluster;
  int order;
};
typedef std::unordered_map<BinKey, Bin, BinHash> BinMap;

BinMap *bins_of(pf_kdtree_t *self) { return reinterpret_cast<BinMap *>(self->nodes); }

BinKey key_of(pf_kdtree_t *self, pf_vector_t pose)
{
  BinKey b;
  for (int i = 0; i < 3; i++) b.k[i] = (int)floor(pose.v[i] / self->size[i]);
  return b;
}

}  // namespace

// pf_kdtree.c:66-86
pf_kdtree_t *pf_kdtree_alloc(int max_size)
{
  pf_kdtree_t *self = (pf_kdtree_t *)calloc(1, sizeof(pf_kdtree_t));
  self->size[0] = 0.50;
  self->size[1] = 0.50;
  self->size[2] = (10 * M_PI / 180);
  self->root = NULL;
  self->node_count = 0;
  self->node_max_count = max_size;
  self->nodes = reinterpret_cast<pf_kdtree_node_t *>(new BinMap());
  self->leaf_count = 0;
  return self;
}

void pf_kdtree_free(pf_kdtree_t *self)
{
  if (!self) return;
  delete bins_of(self);
  free(self);
}

void pf_kdtree_clear(pf_kdtree_t *self)
{
  bins_of(self)->clear();
  self->root = NULL;
  self->leaf_count = 0;
  self->node_count = 0;
}

void pf_kdtree_insert(pf_kdtree_t *self, pf_vector_t pose, double value)
{
  BinMap *m = bins_of(self);
  BinKey k = key_of(self, pose);
  BinMap::iterator it = m->find(k);
  if (it == m->end())
  {
    Bin b = {value, -1, (int)m->size()};
    (*m)[k] = b;
    self->leaf_count = (int)m->size();
    self->node_count = (int)m->size();
  }
  else
    it->second.value += value;
}

double pf_kdtree_get_prob(pf_kdtree_t *self, pf_vector_t pose)
{
  BinMap *m = bins_of(self);
  BinMap::iterator it = m->find(key_of(self, pose));
  return it == m->end() ? 0.0 : it->second.value;
}

int pf_kdtree_get_cluster(pf_kdtree_t *self, pf_vector_t pose)
{
  BinMap *m = bins_of(self);
  BinMap::iterator it = m->find(key_of(self, pose));
  return it == m->end() ? -1 : it->second.cluster;
}

void pf_kdtree_cluster(pf_kdtree_t *self)
{
  BinMap *m = bins_of(self);
  std::vector<BinMap::iterator> by_order(m->size());
  for (BinMap::iterator it = m->begin(); it != m->end(); ++it)
  {
    it->second.cluster = -1;
    by_order[it->second.order] = it;
  }
  int clusters = 0;
  std::vector<BinMap::iterator> stack;
  for (size_t s = 0; s < by_order.size(); s++)
  {
    if (by_order[s]->second.cluster >= 0) continue;
    by_order[s]->second.cluster = clusters;
    stack.push_back(by_order[s]);
    while (!stack.empty())
    {
      BinMap::iterator cur = stack.back();
      stack.pop_back();
      for (int q = 0; q < 27; q++)
      {
        BinKey nk = cur->first;
        nk.k[0] += q / 9 - 1;
        nk.k[1] += (q % 9) / 3 - 1;
        nk.k[2] += (q % 9) % 3 - 1;
        BinMap::iterator nb = m->find(nk);
        if (nb == m->end() || nb->second.cluster >= 0) continue;
        nb->second.cluster = clusters;
        stack.push_back(nb);
      }
    }
    clusters++;
  }
}

}  // extern "C"
