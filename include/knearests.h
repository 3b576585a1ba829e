/* knearests.h -- drop-in for the reference's knearests.h:1-29 (ssloy/cuda_knearests).
 *
 * Exact k-nearest neighbours of every point of a 3-D cloud among all the other points, for points in
 * [0,1000]^3, on one NVIDIA B200 (sm_100a).  The struct below has the reference's field names, order
 * and types, and the seven functions have the reference's C++ signatures, so a TU written against the
 * reference (test_knearests.cu:143-171) compiles and LINKS unchanged: like the reference, these are
 * C++-linkage symbols (_Z10kn_prepareP6float3i, _Z8kn_solveP10kn_problem, ...).  A true
 * C ABI with status codes, run-time k and device-resident inputs is in knearests_c.h -- the kn_*
 * functions here are thin wrappers over it that keep the reference's behaviour on errors: a message on
 * stderr and exit(EXIT_FAILURE) (knearests.cu:163-167,206-210).
 *
 * Differences a caller can observe (all within what the reference leaves undefined):
 *  - the layout is deterministic: cells in ascending cell id, points inside a cell in ascending
 *    original index; rows of kn_get_knearests are ascending in (squared distance, sorted index);
 *  - the grid is not a cube unless asked for: dimx, dimy, dimz differ (narrow cells in x, cells as wide as the search
 *    radius in y and z: knc_fast_dims in knearests_c.h).  The environment variable KN_GRID_DENSITY=3.1 gives the
 *    reference's cube of 3.1 points per cell (knearests.cu:249).  The grid decides the cell order only;
 *  - d_ptrs has dimx*dimy*dimz + 1 entries (exclusive prefix sum; last = number of points);
 *  - there is no ring offset table of cells: num_cell_offsets is the size of the (dj, dk) row table the query
 *    kernels use in its place (121), d_cell_offsets/d_cell_offset_dists are NULL;
 *  - d_cell_max holds ONE float: the largest search radius (squared) any query needed;
 *  - any number of points >= 1 is accepted (the reference exits below 11 545, knearests.cu:255-258),
 *    and the search is exact for any distribution (the reference silently stops after ring 15).
 */
#ifndef KNEARESTS_B200_H
#define KNEARESTS_B200_H

#if !defined(__VECTOR_TYPES_H__) && !defined(__CUDACC__)
#include <vector_types.h> /* float3; the reference header relies on nvcc having included it */
#endif

typedef struct {
    int allocated_points;        /* number of input points                                  (knearests.h:4)  */
    int dimx, dimy, dimz;        /* grid resolution, cubic: round(cbrt(n / 3.1))            (knearests.h:5)  */
    int num_cell_offsets;        /* rows of the (dj, dk) table: 121                         (knearests.h:6)  */
    int *d_cell_offsets;         /* always NULL here                                        (knearests.h:7)  */
    float *d_cell_offset_dists;  /* always NULL here                                        (knearests.h:8)  */
    float *d_cell_max;           /* 1 float: max squared search radius used                 (knearests.h:9)  */
    unsigned int *d_permutation; /* [n] permutation[sorted position] = original index       (knearests.h:10) */
    int *d_counters;             /* [dimx*dimy*dimz] points per cell                        (knearests.h:11) */
    int *d_ptrs;                 /* [dimx*dimy*dimz + 1] first sorted position of each cell (knearests.h:12) */
    int *d_globcounter;          /* 1 int: number of points stored                          (knearests.h:13) */
    float3 *d_stored_points;     /* [n] the points in cell order                            (knearests.h:14) */
    unsigned int *d_knearests;   /* [n * k] neighbour rows, sorted-position space           (knearests.h:15) */
} kn_problem;

/* Copies `numpoints` HOST points (AoS xyz, f32, expected in [0,1000]^3) to the device and builds the
 * uniform grid (count, exclusive scan, stable counting sort + permutation).  k = DEFAULT_NB_PLANES of
 * the library build, or the environment variable KN_NB_PLANES when set.   (knearests.h:21, knearests.cu:235) */
kn_problem *kn_prepare(float3 *points, int numpoints);

/* Runs the all-points kNN query; blocking; results stay on the device.   (knearests.h:22, knearests.cu:348) */
void kn_solve(kn_problem *kn);

/* Frees device and host state and sets *kn = NULL.                        (knearests.h:23, knearests.cu:396) */
void kn_free(kn_problem **kn);

/* Each returns a fresh malloc()'d host buffer that the caller free()s.    (knearests.h:25-27, knearests.cu:410-438)
 *   kn_get_points       float3[n]        points in sorted (cell) order
 *   kn_get_knearests    unsigned[n * k]  row = sorted position, entries = sorted positions, nearest
 *                                        first, UINT_MAX = no such neighbour
 *   kn_get_permutation  unsigned[n]      sorted position -> original index
 * Original-index result, as test_knearests.cu:158: neighbors[perm[i]*k + j] = perm[knn[i*k + j]]. */
float3 *kn_get_points(kn_problem *kn);
unsigned int *kn_get_knearests(kn_problem *kn);
unsigned int *kn_get_permutation(kn_problem *kn);

/* Grid occupancy statistics on stderr.                                    (knearests.h:29, knearests.cu:440) */
void kn_print_stats(kn_problem *kn);

#endif /* KNEARESTS_B200_H */
