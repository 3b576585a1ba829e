// kn_internal.h -- private state and launcher prototypes of libknearests_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "knearests.h"
#include "knearests_c.h"

struct knc_multi;  // multi.cu: the slabs of a handle that spans several GPUs

// The handle.  Its first member is the reference-compatible public struct (knearests.h), so a
// kn_problem* handed to reference-style callers and a knc_problem* are the same address.
struct knc_problem {
    kn_problem pub;
    knc_multi *multi;            // non-NULL: a multi-GPU handle (knc_prepare_multi / KN_NUM_GPUS); the work is in its slabs
    int k;
    int device;
    float density;
    float search_scale;
    int algorithm;               // knc_options.algorithm
    unsigned int index_offset;   // knc_set_index_offset: added to every stored neighbour index
    int z_begin, z_end;          // cell planes held (z_end - z_begin == pub.dimz planes LOCALLY)
    int own_z_begin, own_z_end;
    int dims[3];                 // GLOBAL grid resolution (x, y, z) = pub.dimx / dimy / dimz
    long long cells;             // dimx * dimy * (z_end - z_begin)
    int query_begin, query_count;
    int solved;
    int owns_stream;             // 0: the caller's stream (knc_options.stream): never synchronised by prepare, never recycled
    int timings_pending;         // bit 0: prepare's events, bit 1: solve's events not yet read (read needs them complete)
    int *d_retry;                // scratch of an asynchronous solve in flight (freed by knc_solve_finish)
    float4 *d_points4;           // [n] x, y, z, bits(own sorted position): what the query kernels read
    unsigned int *d_tags;        // [n] 4th word of every input record in sorted order (point_stride 4 only, else NULL)
    unsigned int *d_diag;        // [KN_DIAG_WORDS] restarts, max radius^2 bits, halo misses, retry count, block fallbacks
    cudaStream_t stream;         // where the handle's work runs: its own, or the caller's (knc_options.stream)
    cudaStream_t kit_stream;     // the library-owned stream that came with the events (recycled at knc_free)
    cudaStream_t copy_stream;    // device->host row copies of knc_solve_to_host (created on first use)
    cudaEvent_t ev[5];             // h2d start, build start, build end, solve start, solve end
    knc_timings t;
};

#define KN_DIAG_WORDS 8
#define KN_ROW_TABLE_SIZE 121  // (2 * 5 + 1)^2 rows (dj, dk) around a query's own row: knn_query.cu's table

// Grid geometry handed to kernels.  The reference's grid is a cube of D^3 cells (knearests.cu:249); here the three
// resolutions are independent (kn_problem has dimx / dimy / dimz, knearests.h:5-7): the default grid has cells that
// are as wide as the first search radius in y and z and narrow in x (knc_fast_dims).
struct GridGeom {
    int Dx, Dy, Dz;       // GLOBAL resolution per axis
    int zb, ze;           // cell planes held locally: [zb, ze)
    float Dfx, Dfy, Dfz;  // (float)D per axis -- cellFromPoint multiplies by it (knearests.cu:23-25)
    float wx, wy, wz;     // cell widths 1000 / D
    float invwx;          // Dx / 1000
};

inline GridGeom kn_make_geom(const int dims[3], int zb, int ze) {
    GridGeom g;
    g.Dx = dims[0]; g.Dy = dims[1]; g.Dz = dims[2];
    g.zb = zb; g.ze = ze;
    g.Dfx = (float)g.Dx; g.Dfy = (float)g.Dy; g.Dfz = (float)g.Dz;
    g.wx = 1000.f / g.Dfx; g.wy = 1000.f / g.Dfy; g.wz = 1000.f / g.Dfz;
    g.invwx = g.Dfx / 1000.f;
    return g;
}

// ---- grid_build.cu -----------------------------------------------------------------------------
struct BuildScratch {
    float4 *d_recs;              // [n] records in cell order, arrival order inside a cell
    unsigned int *d_rec_tags;    // [n] their 4th input words (point_stride 4, else NULL)
    unsigned long long *d_tile;  // scan tile states
    int *d_ticket;               // scan ticket + big-cell worklist counter
    int *d_biglist;              // [n / (KN_BIG_CELL + 1) + 1] cells too large for the per-thread ordering
    size_t tile_count;
    float4 *d_bin_recs;          // [n] z-banded copy of the input (large clouds only, else NULL)
    int *d_bin_counts;           // [kn_bin_counts_ints(n)] band totals + band cursors
    unsigned int *d_bin_tags;    // [n] 4th input words in banded order (large clouds with point_stride 4, else NULL)
};
size_t kn_bin_counts_ints(int n);
#define KN_BIN_MIN_POINTS 4000000  // below this the records being placed stay in L2 anyway (build, ms, with / without the
                                   // split: 2 M points 0.13 / 0.10, 5 M 0.26 / 0.28, 10 M 0.48 / 0.67, 14 M 0.65 / 1.07, 19 M 0.85 / 1.67)
size_t kn_build_scratch_tiles(long long cells);
cudaError_t kn_launch_build(const float *d_in_xyz, int stride, int n, const GridGeom &g, long long cells, int *d_counters,
                            int *d_ptrs, int *d_globcounter, unsigned int *d_permutation, float *d_points3,
                            float4 *d_points4, unsigned int *d_tags, const BuildScratch &s, cudaStream_t st);
cudaError_t kn_launch_generate(float *d_xyz, long long first, long long count, uint64_t seed, cudaStream_t st);
cudaError_t kn_launch_remap_original(const unsigned int *d_knn, const unsigned int *d_perm, int n, int k,
                                     unsigned int *d_out, cudaStream_t st);

// ---- knn_query.cu ------------------------------------------------------------------------------
// d_retry_list: scratch of kn_retry_ints(query_count) ints (queries handed on to the retry kernel + blocks handed
// on to the previous-generation select kernel), or NULL for the general kernel alone
size_t kn_retry_ints(int query_count);
cudaError_t kn_launch_query(const float4 *d_points4, const int *d_ptrs, const GridGeom &g, int n, int k,
                            int query_begin, int query_count, float radius0_sq, unsigned int *d_knn,
                            unsigned int *d_diag, int *d_retry_list, int force_general, unsigned int index_offset,
                            const float *d_ext_queries, cudaStream_t st);
cudaError_t kn_launch_rows_dists(const float4 *d_points4, const unsigned int *d_rows, const float *d_ext_queries,
                                 int query_begin, unsigned int index_offset, long long rows, int k, float *d_out,
                                 cudaStream_t st);

// ---- multi.cu ----------------------------------------------------------------------------------
int kn_multi_solve(knc_problem *parent);
int kn_multi_get(knc_problem *parent, int what, void *dst);  // 0 sorted points, 1 permutation, 2 neighbour rows
int kn_multi_counters(knc_problem *parent, int *host_counters);
void kn_multi_free(knc_problem *parent);

// ---- api.cu ------------------------------------------------------------------------------------
int kn_set_error(int code, const char *fmt, ...);  // records the message knc_last_error() returns
cudaError_t kn_dev_alloc(void **p, size_t bytes, cudaStream_t st);  // stream-ordered, from the library's private pool

#define KN_BIG_CELL 32  // a cell whose records see more than this many of their own kind to one side is ordered by a whole block
