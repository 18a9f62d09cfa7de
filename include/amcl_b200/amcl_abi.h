/*
 * amcl_abi.h -- the C half of the drop-in boundary.
 *
 * Every type and prototype in this file is an interface fact of the reference
 * (ros-planning/navigation 1.17.3, package amcl): amcl_node.cpp is compiled
 * against these layouts and pokes the fields directly, so sizes and offsets
 * are ABI.  The text was written from scratch for this library; each block
 * names the reference declaration it stands in for (path:line under
 * /root/reference/amcl).  The layout is pinned by static asserts at the end
 * (x86-64 SysV).
 *
 * The forwarding headers under include/amcl/{pf,map}/ include this file, so a
 * caller that says  #include "amcl/pf/pf.h"  keeps compiling unchanged.
 *
 * What is different from the reference implementation behind these symbols:
 *   - the sample arrays, the distance map and all O(N) state live in B200 HBM;
 *     pf_sample_set_t::samples is a host *mirror* (see amclb200_set_mirror in
 *     amcl_b200.h),
 *   - pf_t::limit_cache, pf_sample_set_t::kdtree and AMCLLaser::temp_obs are
 *     pointers only this library interprets; they hold device-context handles.
 */
#ifndef AMCL_B200_AMCL_ABI_H
#define AMCL_B200_AMCL_ABI_H

#include <stdio.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------ */
/* 3-vectors and 3x3 matrices          (include/amcl/pf/pf_vector.h:36-88)  */
/* ------------------------------------------------------------------------ */

typedef struct { double v[3]; } pf_vector_t;      /* pf_vector.h:39-42 */
typedef struct { double m[3][3]; } pf_matrix_t;   /* pf_vector.h:46-49 */

pf_vector_t pf_vector_zero(void);                                  /* :53 */
int         pf_vector_finite(pf_vector_t a);                       /* :56 */
void        pf_vector_fprintf(pf_vector_t s, FILE *file, const char *fmt); /* :59 */
pf_vector_t pf_vector_add(pf_vector_t a, pf_vector_t b);           /* :62 */
pf_vector_t pf_vector_sub(pf_vector_t a, pf_vector_t b);           /* :65 */
/* local -> global: b (+) a, heading re-wrapped with atan2(sin,cos)   :68 */
pf_vector_t pf_vector_coord_add(pf_vector_t a, pf_vector_t b);
/* global -> local: a (-) b                                           :71 */
pf_vector_t pf_vector_coord_sub(pf_vector_t a, pf_vector_t b);
pf_matrix_t pf_matrix_zero(void);                                  /* :75 */
int         pf_matrix_finite(pf_matrix_t a);                       /* :78 */
void        pf_matrix_fprintf(pf_matrix_t s, FILE *file, const char *fmt); /* :81 */
/* a = r d r^T for symmetric a (eigenvalues ascending)                :88 */
void        pf_matrix_unitary(pf_matrix_t *r, pf_matrix_t *d, pf_matrix_t a);

/* include/amcl/pf/eig3.h:9 -- symmetric 3x3 eigen-solver (Householder + QL) */
void eigen_decomposition(double A[3][3], double V[3][3], double d[3]);

/* ------------------------------------------------------------------------ */
/* Gaussian pdf                          (include/amcl/pf/pf_pdf.h:45-79)   */
/* ------------------------------------------------------------------------ */

typedef struct                                     /* pf_pdf.h:45-60 */
{
  pf_vector_t x;       /* mean */
  pf_matrix_t cx;      /* covariance */
  double      cxdet;   /* unused by the reference, kept for layout */
  pf_matrix_t cr;      /* rotation part of cx = cr diag(cd^2) cr^T */
  pf_vector_t cd;      /* per-axis standard deviations */
} pf_pdf_gaussian_t;

pf_pdf_gaussian_t *pf_pdf_gaussian_alloc(pf_vector_t x, pf_matrix_t cx); /* :64 */
void        pf_pdf_gaussian_free(pf_pdf_gaussian_t *pdf);                /* :67 */
/* polar Box-Muller on drand48(), zero mean, std-dev sigma               :76 */
double      pf_ran_gaussian(double sigma);
pf_vector_t pf_pdf_gaussian_sample(pf_pdf_gaussian_t *pdf);              /* :79 */

/* ------------------------------------------------------------------------ */
/* Pose histogram ("kd-tree")            (include/amcl/pf/pf_kdtree.h:37-99) */
/* ------------------------------------------------------------------------ */

/* The reference keeps a binary tree over integer bin keys.  Here the
 * histogram is a device hash table; the struct is kept so that sizeof and the
 * size[] / leaf_count fields stay meaningful to a reader of the struct.  */
typedef struct pf_kdtree_node                      /* pf_kdtree.h:37-58 */
{
  int    leaf, depth;
  int    pivot_dim;
  double pivot_value;
  int    key[3];
  double value;
  int    cluster;
  struct pf_kdtree_node *children[2];
} pf_kdtree_node_t;

typedef struct                                     /* pf_kdtree.h:62-77 */
{
  double size[3];             /* bin edge: 0.5 m, 0.5 m, 10 deg */
  pf_kdtree_node_t *root;     /* unused here (always NULL) */
  int node_count, node_max_count;
  pf_kdtree_node_t *nodes;    /* opaque: device histogram handle */
  int leaf_count;             /* number of occupied bins */
} pf_kdtree_t;

pf_kdtree_t *pf_kdtree_alloc(int max_size);                        /* :81 */
void   pf_kdtree_free(pf_kdtree_t *self);                          /* :84 */
void   pf_kdtree_clear(pf_kdtree_t *self);                         /* :87 */
void   pf_kdtree_insert(pf_kdtree_t *self, pf_vector_t pose, double value); /* :90 */
void   pf_kdtree_cluster(pf_kdtree_t *self);                       /* :93 */
double pf_kdtree_get_prob(pf_kdtree_t *self, pf_vector_t pose);    /* :96 */
int    pf_kdtree_get_cluster(pf_kdtree_t *self, pf_vector_t pose); /* :99 */

/* ------------------------------------------------------------------------ */
/* Particle filter                          (include/amcl/pf/pf.h:38-203)   */
/* ------------------------------------------------------------------------ */

struct _pf_t;
struct _rtk_fig_t;
struct _pf_sample_set_t;

typedef pf_vector_t (*pf_init_model_fn_t)(void *init_data);        /* pf.h:45 */
typedef void   (*pf_action_model_fn_t)(void *action_data,          /* pf.h:49 */
                                       struct _pf_sample_set_t *set);
typedef double (*pf_sensor_model_fn_t)(void *sensor_data,          /* pf.h:54 */
                                       struct _pf_sample_set_t *set);

typedef struct                                     /* pf.h:59-67, 32 B */
{
  pf_vector_t pose;
  double      weight;
} pf_sample_t;

typedef struct                                     /* pf.h:71-87, 176 B */
{
  int         count;
  double      weight;
  pf_vector_t mean;
  pf_matrix_t cov;
  double      m[4], c[2][2];   /* raw weighted sums the mean/cov come from */
} pf_cluster_t;

typedef struct _pf_sample_set_t                    /* pf.h:91-108, 152 B */
{
  int          sample_count;
  pf_sample_t *samples;        /* host mirror of the device AoS array */
  pf_kdtree_t *kdtree;
  int          cluster_count, cluster_max_count;
  pf_cluster_t *clusters;
  pf_vector_t  mean;
  pf_matrix_t  cov;
  int          converged;
  double       n_effective;
} pf_sample_set_t;

typedef struct _pf_t                               /* pf.h:112-143, 408 B */
{
  int    min_samples, max_samples;
  double pop_err, pop_z;       /* the caller writes these after pf_alloc */
  int   *limit_cache;          /* opaque here: device-context handle */
  int    current_set;
  pf_sample_set_t sets[2];
  double w_slow, w_fast;
  double alpha_slow, alpha_fast;
  pf_init_model_fn_t random_pose_fn;
  void  *random_pose_data;
  double dist_threshold;
  int    converged;
  int    selective_resampling;
} pf_t;

pf_t *pf_alloc(int min_samples, int max_samples,                   /* pf.h:147 */
               double alpha_slow, double alpha_fast,
               pf_init_model_fn_t random_pose_fn, void *random_pose_data);
void  pf_free(pf_t *pf);                                           /* :152 */
void  pf_init(pf_t *pf, pf_vector_t mean, pf_matrix_t cov);        /* :155 */
void  pf_init_model(pf_t *pf, pf_init_model_fn_t init_fn, void *init_data); /* :158 */
void  pf_update_action(pf_t *pf, pf_action_model_fn_t action_fn, void *action_data); /* :161 */
void  pf_update_sensor(pf_t *pf, pf_sensor_model_fn_t sensor_fn, void *sensor_data); /* :164 */
void  pf_update_resample(pf_t *pf);                                /* :167 */
void  pf_set_selective_resampling(pf_t *pf, int selective_resampling); /* :170 */
void  pf_get_cep_stats(pf_t *pf, pf_vector_t *mean, double *var);  /* :173 */
int   pf_get_cluster_stats(pf_t *pf, int cluster, double *weight,  /* :177 */
                           pf_vector_t *mean, pf_matrix_t *cov);
void  pf_cluster_stats(pf_t *pf, pf_sample_set_t *set);            /* :181 */
int   pf_update_converged(pf_t *pf);                               /* :198 */
void  pf_init_converged(pf_t *pf);                                 /* :201 */
/* Declared by the reference (:203) but only defined there as copy_set();
 * both spellings are exported here.  */
void  pf_copy_set(pf_sample_set_t *set_a, pf_sample_set_t *set_b);
void  copy_set(pf_sample_set_t *set_a, pf_sample_set_t *set_b);

/* ------------------------------------------------------------------------ */
/* Occupancy grid + likelihood field        (include/amcl/map/map.h:38-144) */
/* ------------------------------------------------------------------------ */

#define MAP_WIFI_MAX_LEVELS 8                      /* map.h:42 */

typedef struct                                     /* map.h:46-58, 16 B */
{
  int    occ_state;   /* -1 free, 0 unknown, +1 occupied */
  double occ_dist;    /* metres to the nearest occupied cell, capped */
} map_cell_t;

typedef struct                                     /* map.h:62-79, 48 B */
{
  double origin_x, origin_y;   /* world coordinates of the map *centre* */
  double scale;                /* metres per cell */
  int    size_x, size_y;
  map_cell_t *cells;           /* malloc'd by the caller, freed by map_free */
  double max_occ_dist;
} map_t;

map_t      *map_alloc(void);                                       /* map.h:88 */
void        map_free(map_t *map);                                  /* :91 */
map_cell_t *map_get_cell(map_t *map, double ox, double oy, double oa); /* :94 */
int         map_load_occ(map_t *map, const char *filename, double scale, int negate); /* :97 */
/* Rebuilds occ_dist (and the device distance map the laser models read). :103 */
void        map_update_cspace(map_t *map, double max_occ_dist);
/* One simulated range reading by integer line walking.                  :111 */
double      map_calc_range(map_t *map, double ox, double oy, double oa, double max_range);

/* map.h:133-144 -- grid <-> world conversions (the map origin is its centre) */
#define MAP_WXGX(map, i) (map->origin_x + ((i) - map->size_x / 2) * map->scale)
#define MAP_WYGY(map, j) (map->origin_y + ((j) - map->size_y / 2) * map->scale)
#define MAP_GXWX(map, x) (floor((x - map->origin_x) / map->scale + 0.5) + map->size_x / 2)
#define MAP_GYWY(map, y) (floor((y - map->origin_y) / map->scale + 0.5) + map->size_y / 2)
#define MAP_VALID(map, i, j) ((i >= 0) && (i < map->size_x) && (j >= 0) && (j < map->size_y))
#define MAP_INDEX(map, i, j) ((i) + (j) * map->size_x)

#ifdef __cplusplus
}
#endif

/* Layout pins (probed from the compiled reference). */
#if defined(__cplusplus) && defined(__x86_64__)
#include <cstddef>
static_assert(sizeof(pf_sample_t) == 32, "pf_sample_t");
static_assert(sizeof(pf_cluster_t) == 176, "pf_cluster_t");
static_assert(sizeof(pf_kdtree_node_t) == 72, "pf_kdtree_node_t");
static_assert(sizeof(pf_sample_set_t) == 152, "pf_sample_set_t");
static_assert(offsetof(pf_sample_set_t, samples) == 8, "set.samples");
static_assert(offsetof(pf_sample_set_t, cluster_count) == 24, "set.cluster_count");
static_assert(offsetof(pf_sample_set_t, clusters) == 32, "set.clusters");
static_assert(offsetof(pf_sample_set_t, mean) == 40, "set.mean");
static_assert(offsetof(pf_sample_set_t, cov) == 64, "set.cov");
static_assert(offsetof(pf_sample_set_t, converged) == 136, "set.converged");
static_assert(offsetof(pf_sample_set_t, n_effective) == 144, "set.n_effective");
static_assert(sizeof(pf_t) == 408, "pf_t");
static_assert(offsetof(pf_t, pop_err) == 8, "pf.pop_err");
static_assert(offsetof(pf_t, limit_cache) == 24, "pf.limit_cache");
static_assert(offsetof(pf_t, current_set) == 32, "pf.current_set");
static_assert(offsetof(pf_t, sets) == 40, "pf.sets");
static_assert(offsetof(pf_t, w_slow) == 344, "pf.w_slow");
static_assert(offsetof(pf_t, random_pose_fn) == 376, "pf.random_pose_fn");
static_assert(offsetof(pf_t, dist_threshold) == 392, "pf.dist_threshold");
static_assert(offsetof(pf_t, converged) == 400, "pf.converged");
static_assert(offsetof(pf_t, selective_resampling) == 404, "pf.selective_resampling");
static_assert(sizeof(map_cell_t) == 16 && offsetof(map_cell_t, occ_dist) == 8, "map_cell_t");
static_assert(sizeof(map_t) == 48, "map_t");
#endif

#endif /* AMCL_B200_AMCL_ABI_H */
