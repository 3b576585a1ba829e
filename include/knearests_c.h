/* knearests_c.h -- the C ABI of the B200 kNN engine (libknearests_b200.so).
 *
 * Plain pointers and sizes, status codes instead of exit(), no C++ or torch types: this is what an FFI
 * (ctypes, cgo, JNI, dlopen) binds.  Every entry point names the reference interface it replaces
 * (file:line in ssloy/cuda_knearests).  The reference's own header has no extern "C" (SURVEY F4); its
 * seven C++-linkage functions are kept, as wrappers over these, in knearests.h.
 *
 * Model: one handle = one point cloud on one GPU (the current device at prepare time).  A handle is
 * not thread-safe; distinct handles are independent.  All calls are synchronous: when they return,
 * the device work they describe is complete (as in the reference, knearests.cu:196,373).
 *
 * Index spaces (knearests.cu:57, test_knearests.cu:158):
 *   original index  i : position of a point in the caller's input array
 *   sorted position s : position in cell order; permutation[s] = i
 * Cell order is deterministic: ascending cell id (x fastest, z slowest, knearests.cu:29), and inside a
 * cell ascending original index.  A neighbour row is ascending in the total order
 * (squared distance, sorted position), where the squared distance is the reference's f32 expression
 * fma(dz,dz, fma(dx,dx, dy*dy)) (knearests.cu:125 as nvcc contracts it).  The point itself is excluded
 * by position, coincident points are not (knearests.cu:123).
 */
#ifndef KNEARESTS_B200_C_H
#define KNEARESTS_B200_C_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes ------------------------------------------------------------------------------- */
#define KNC_OK 0
#define KNC_ERR_INVALID 1     /* bad argument (null pointer, n < 1, k out of range, bad slab)       */
#define KNC_ERR_CUDA 2        /* a CUDA runtime call failed; knc_last_error() has the CUDA string   */
#define KNC_ERR_NOMEM 3       /* host or device allocation failed                                   */
#define KNC_ERR_STATE 4       /* call order: results requested before knc_solve                     */
#define KNC_ERR_HALO 5        /* slab mode: some query needed points beyond the halo it was given   */

#define KNC_EMPTY 0xFFFFFFFFu /* row entry meaning "no such neighbour" (UINT_MAX, knearests.cu:108) */
#define KNC_MAX_K 256         /* largest supported k                                                */

typedef struct knc_problem knc_problem; /* opaque; its first bytes are a kn_problem (knearests.h) */

/* Options for knc_prepare_ex.  Zero-initialise, set struct_size = sizeof(knc_options), then set
 * what you need; a zero field means "default". */
typedef struct {
    size_t struct_size;
    int k;                  /* neighbours per point; default = library's DEFAULT_NB_PLANES (params.h:4) */
    float density;          /* > 0: a CUBIC grid with this many points per cell, D = round(cbrt(n / density))
                               -- 3.1f gives the reference's grid (knearests.cu:249).  0: see grid_dims     */
    int points_on_device;   /* 1: `xyz` is a DEVICE pointer on the current device (no H2D copy)         */
    int grid_dim;           /* force the cubic grid resolution D                                        */
    /* z-slab mode (multi-GPU): the handle holds only cell planes [z_begin, z_end) of a
     * global grid (grid_dim or grid_dims must be set) and answers queries only for the points in planes
     * [own_z_begin, own_z_end); the other planes are a read-only halo.  All zero = whole grid. */
    int z_begin, z_end;
    int own_z_begin, own_z_end;
    float search_scale;     /* tuning: multiplies the first-guess search radius; default 1             */
    int algorithm;          /* 0 = automatic (tile / select / insertion kernel + retry), 1 = general kernel only,
                               2 = insertion kernel for every k, 3 = grid build without the band
                               split, 4 = no tile kernel, 5 = no second-chance pass over a long retry
                               list, 6 = second-chance pass over any retry list (testing, A/B)          */
    int point_stride;       /* floats per input point: 3 (xyz, default) or 4 (xyz + one word that is carried
                               along, see knc_get_tags; e.g. the records knc_slab_route writes)         */
    int grid_dims[3];       /* force the grid resolution per axis (x, y, z; all three > 0).  The reference's grid
                               is a cube (knearests.cu:249) but its handle has dimx / dimy / dimz
                               (knearests.h:5-7) and nothing in its API depends on the cube.  Precedence:
                               grid_dims, grid_dim, density, environment KN_GRID_DENSITY, else knc_fast_dims(n, k).
                               The grid decides the cell order (hence the permutation and the row numbering),
                               never which neighbours a point gets.                                       */
    /* --- for callers that drive several GPUs and cannot afford a host synchronisation per call --- */
    void *stream;           /* a cudaStream_t of the caller on the current device: the handle does ALL its work there
                               (instead of on a stream of its own) and knc_prepare_ex returns without waiting for
                               it -- which requires query_hint below in slab mode.  NULL: own stream, blocking. */
    int query_hint;         /* 1: query_begin_hint / query_count_hint are the sorted positions of the owned planes'
                               points (known to the caller from the routing counts): the prepare then needs no
                               device->host read to learn them                                                    */
    int query_begin_hint, query_count_hint;
} knc_options;

/* What the handle holds; device pointers are valid until knc_free. */
typedef struct {
    int n;                  /* points held (slab mode: owned + halo)                                    */
    int k;
    int dimx, dimy, dimz;   /* GLOBAL grid resolution                                                   */
    int z_begin, z_end;     /* cell planes held                                                         */
    int query_begin;        /* first sorted position that is a query                                    */
    int query_count;        /* number of queries = rows of d_knearests                                  */
    int solved;             /* 1 after knc_solve                                                        */
    const float *d_points;          /* [n*3]  sorted points, AoS xyz (kn_problem.d_stored_points)       */
    const float *d_points4;         /* [n*4]  sorted points as x,y,z,bits(own sorted position)          */
    const unsigned int *d_permutation; /* [n]                                                           */
    const int *d_counters;          /* [cells]                                                          */
    const int *d_ptrs;              /* [cells + 1]                                                      */
    const unsigned int *d_knearests;   /* [query_count * k], row r = query sorted position query_begin+r */
} knc_info;

/* Grid occupancy, the numbers kn_print_stats prints (knearests.cu:440-466). */
typedef struct {
    long long cells;
    int min_per_cell, max_per_cell;
    double avg_per_cell;
    long long total;            /* sum of the counters = n                                              */
    long long histogram[64];    /* histogram[c] = cells holding c points; the last bin also counts more */
} knc_stats;

/* Device time of the last prepare / solve on this handle, in milliseconds (CUDA events; what the
 * reference prints as "kn_firstbuild: X msec" / "kn_solve: X msec", knearests.cu:200,376). */
typedef struct {
    float build_ms;         /* grid build kernels only (band split, count, scan, scatter, order)         */
    float solve_ms;         /* query kernel only                                                        */
    float h2d_ms;           /* host->device copy of the input inside knc_prepare (0 for device input)   */
    unsigned int restarts;  /* queries whose first-guess radius was too small (diagnostic)              */
    float max_radius2;      /* largest squared search radius any query ended with                       */
    unsigned int handed_over_blocks; /* blocks of 128 queries the tile kernel left to the select kernel
                                        (their candidates did not fit its shared-memory tile; diagnostic)  */
    unsigned int second_retries;     /* queries still unfinished after the second, larger sphere of the select
                                        kernel (0 when the retry list was short: that pass is skipped; diagnostic) */
} knc_timings;

/* ---- the hot path -------------------------------------------------------------------------------- */

/* Replaces kn_prepare (knearests.h:21, knearests.cu:235-344): copy n HOST points (AoS xyz f32) to the
 * device, build the grid.  k <= 0 selects the default.  On success *out is a new handle. */
int knc_prepare(const float *xyz, int n, int k, knc_problem **out);

/* Same with options: device-resident input, grid density, forced resolution, z-slab mode. */
int knc_prepare_ex(const float *xyz, int n, const knc_options *opt, knc_problem **out);

/* Replaces kn_solve (knearests.h:22, knearests.cu:348-392): the all-points kNN query.  Blocking.
 * May be called again; it recomputes the same rows. */
int knc_solve(knc_problem *p);

/* knc_solve in two halves: knc_solve_async enqueues the query on the handle's stream and returns;
 * knc_solve_finish waits for it, reads the diagnostics and reports like knc_solve (KNC_ERR_HALO included).  Between
 * the two the caller may enqueue more work on the same stream (the multi-GPU driver: the ranks' vote on
 * knc_halo_flag_ptr).  knc_halo_flag_ptr: DEVICE address of the word that counts this slab's queries that reached
 * beyond their halo planes (valid after knc_solve_async, in stream order; 0 = the halo was wide enough). */
int knc_solve_async(knc_problem *p);
int knc_solve_finish(knc_problem *p);
int knc_halo_flag_ptr(knc_problem *p, const unsigned int **d_flag);

/* knc_solve and kn_get_knearests in one call, overlapped: the query runs in `chunks` pieces (<= 0: 8) and
 * the rows of each finished piece are copied to out_knn (HOST, query_count*k unsigned; page-locked
 * memory is needed for the copy to overlap) while the next piece is computed.  The rows also stay on the
 * device.  Replaces the pair kn_solve + kn_get_knearests (knearests.cu:348,430) for callers whose result
 * must end up on the host: the 4*k*n-byte read-back otherwise costs more than the query itself. */
int knc_solve_to_host(knc_problem *p, unsigned int *out_knn, int chunks);

/* Replaces kn_free (knearests.h:23, knearests.cu:396-408).  NULL is accepted. */
int knc_free(knc_problem *p);

/* Device memory the library frees is kept in a PRIVATE stream-ordered pool per device (cudaMemPoolCreate; the
 * device's default pool, which other libraries of the process share, is never touched), so that preparing
 * cloud after cloud does not pay the driver's page mapping again (the reference cudaMallocs and cudaFrees
 * every buffer per problem, knearests.cu:205-231,396-408).  After knc_free the memory therefore still counts
 * as used for everyone else: knc_trim gives all of it back to the driver (current device; blocking). */
int knc_trim(void);

/* Replace kn_get_points / kn_get_permutation / kn_get_knearests (knearests.h:25-27,
 * knearests.cu:410-438), writing into CALLER-provided host buffers instead of malloc()ing:
 *   out_xyz   3*n floats            out_perm  n unsigned          out_knn  query_count*k unsigned  */
int knc_get_points(knc_problem *p, float *out_xyz);
int knc_get_permutation(knc_problem *p, unsigned int *out_perm);
int knc_get_knearests(knc_problem *p, unsigned int *out_knn);

/* The same rows in ORIGINAL index space -- the remap every caller of the reference does on the CPU
 * (test_knearests.cu:155-160), done on the device: out[perm[s]*k + j] = perm[knn[s*k + j]].
 * Whole-grid handles only (not slab mode).  out: n*k unsigned, host. */
int knc_get_knearests_original(knc_problem *p, unsigned int *out_neighbors);

/* Squared distances of the stored rows (HOST, query_count*k floats; FLT_MAX where a row entry is KNC_EMPTY),
 * recomputed with the reference's f32 expression fma(dz,dz, fma(dx,dx, dy*dy)) (knearests.cu:125 as compiled).
 * The reference API never returns distances (knearests.h:21-29); its kd-tree does
 * (kd_tree.cpp:168, neighbors_sq_dist). */
int knc_get_dists(knc_problem *p, float *out_d2);

/* k nearest neighbours of m query points that are NOT members of the cloud (HOST xyz, m*3 floats), as the
 * reference's kd-tree offers (kd_tree.cpp:168-192: query by coordinates); the reference's GPU path has no such
 * entry point.  Rows (HOST, m*k unsigned) hold sorted positions, ascending (d2, sorted position), nothing is
 * excluded; out_d2 (HOST, m*k floats) may be NULL.  Works on any prepared handle, solved or not. */
int knc_query_points(knc_problem *p, const float *xyz, int m, unsigned int *out_knn, float *out_d2);

/* The 4th word of every input record, in SORTED order (DEVICE pointer to allocated_points words, owned by the
 * handle, valid until knc_free): tags[s] belongs to the point at sorted position s, the queries are
 * tags[query_begin .. query_begin + query_count).  Only for handles prepared with point_stride = 4
 * (KNC_ERR_STATE otherwise).  No counterpart in the reference (its input is float3, knearests.h:21): the
 * multi-GPU caller keeps each point's GLOBAL id there, so this run IS its slice of the global permutation --
 * gathered during the build, where the record is being read anyway, instead of by a random gather afterwards
 * (knc_get_permutation_global). */
int knc_get_tags(knc_problem *p, const unsigned int **d_tags);

/* Replaces kn_print_stats (knearests.h:29, knearests.cu:440-466), as data. */
int knc_get_stats(knc_problem *p, knc_stats *out);

/* ---- introspection ------------------------------------------------------------------------------- */
int knc_get_info(knc_problem *p, knc_info *out);
int knc_get_timings(knc_problem *p, knc_timings *out);

/* Resolution of the cubic grid for a number of points per cell: max(1, round(cbrt(n / density)))
 * (knearests.cu:249; density <= 0 means the reference's 3.1). */
int knc_grid_dim(int n, float density);

/* The grid the library chooses when the caller names none: dims[0..2] = cells in x, y, z.  Cells are 2 % wider in
 * y and z than the first search radius (the sphere expected to hold k + 2.5 sqrt(k) + 2 points of a uniform cloud)
 * and narrow in x, about one point per cell: a query then needs the 3 x 3 x-rows of cells around its own and
 * nothing else, which is the tile the query kernel stages in shared memory.  k <= 0: the default k.  The reference
 * sizes a cube for 3.1 points per cell (knearests.cu:249); KN_GRID_DENSITY=3.1 or density = 3.1f restores it. */
int knc_fast_dims(int n, int k, int *dims);

/* Message of the last failure on the calling thread ("" if none). */
const char *knc_last_error(void);

/* Library build information: "knearests_b200 <version> sm_100a k_default=<K>". */
const char *knc_version(void);

/* ---- synthetic input (bench / tests) --------------------------------------------------
 * Fills d_xyz (DEVICE, 3*count floats) with points [first, first+count) of the counter-based uniform
 * cloud: u = splitmix64(seed ^ (3*i + c)), coord = 1000 * (u >> 40) * 2^-24.  No reference counterpart
 * (the reference has no generator); bit-identical to oracle/knn_oracle.c:kno_generate_uniform. */
int knc_generate_uniform_device(float *d_xyz, long long first, long long count, uint64_t seed);

/* ---- multi-GPU, one process per GPU ---------------------------------------------------
 * No reference counterpart: the reference is single-GPU.  The cloud is cut into z-slabs of
 * cell planes of one GLOBAL grid: rank d owns planes [zsplit[d], zsplit[d+1]) and additionally needs the
 * `halo` planes on either side.  The caller exchanges the buffers (NCCL all-to-all); these two calls are
 * the device work before and after.
 *
 * knc_slab_route: stable multi-split of n local points (DEVICE xyz, global ids first_id .. first_id+n-1)
 * into per-destination groups: d_out_rec (DEVICE, room for `capacity` 16-byte records x, y, z, id -- one
 * buffer, so that one all-to-all moves everything) receives, for d = 0..nranks-1 in turn, every point whose
 * plane lies in [zsplit[d]-halo, zsplit[d+1]+halo), in input order; out_counts[d] (HOST) is the size of
 * group d.  A received buffer goes to knc_prepare_ex with point_stride = 4; its 4th words are the id table
 * for knc_get_permutation_global / knc_get_knearests_global (id_stride = 4).  zsplit is a HOST array of nranks+1 planes rising
 * from 0 to grid_dim.  KNC_ERR_NOMEM if the groups do not fit `capacity` (out_counts is still filled). */
int knc_slab_route(const float *d_xyz, int n, unsigned int first_id, int grid_dim, int nranks, const int *zsplit,
                   int halo, float *d_out_rec, long long capacity, long long *out_counts);

/* The same routing in two steps, for an exchange WITHOUT a staging copy: knc_route_begin counts the groups
 * (out_counts, HOST) and keeps its scratch; the ranks agree where each group lands (an all-gather of the
 * counts); knc_route_scatter then writes group d straight to dest_bufs[d] + dest_first[d] records -- a buffer in
 * this GPU's memory or in a PEER GPU's memory mapped into this process (knc_peer_open), in which case the
 * 16-byte stores travel over NVLink while the kernel runs: routing and exchange are one kernel, no NCCL
 * all-to-all, no send buffer.  dest_bufs / dest_first: HOST arrays of nranks entries.  Blocking; after it returns
 * and the ranks have met at a barrier, every receive buffer is complete.  knc_route_end releases the scratch. */
typedef struct knc_route knc_route;
int knc_route_begin(const float *d_xyz, int n, unsigned int first_id, int grid_dim, int nranks, const int *zsplit,
                    int halo, knc_route **out, long long *out_counts);
int knc_route_scatter(knc_route *r, void *const *dest_bufs, const long long *dest_first);
int knc_route_end(knc_route *r);

/* The same two steps WITHOUT host synchronisation, on a stream of the caller (a cudaStream_t, NULL = the default
 * stream).  knc_route_begin_async leaves, in stream order, 3 * nranks counts in d_counts3 (DEVICE, long long): for every
 * destination d the records it will receive from this rank, how many of them lie in d's OWN planes (they become d's
 * queries), and how many lie in d's halo planes BELOW its own (they come before d's queries in sorted order) -- so that
 * ONE all-gather of these rows tells every rank where its groups land, how many points and queries every rank holds
 * and at which global sorted position each rank's queries begin.  knc_route_scatter_async enqueues the scatter. */
int knc_route_begin_async(const float *d_xyz, int n, unsigned int first_id, int grid_dim, int nranks, const int *zsplit,
                          int halo, void *stream, long long *d_counts3, knc_route **out);
int knc_route_scatter_async(knc_route *r, void *const *dest_bufs, const long long *dest_first, void *stream);

/* Receive buffers other processes of the node can write into (CUDA IPC): knc_peer_alloc = cudaMalloc + a 64-byte
 * handle to hand to the other ranks; knc_peer_open maps another rank's buffer into this process (peer access is
 * enabled on first use); knc_peer_close / knc_peer_free undo them. */
int knc_peer_alloc(size_t bytes, void **d_ptr, unsigned char *handle64);
int knc_peer_open(const unsigned char *handle64, void **d_ptr);
int knc_peer_close(void *d_ptr);
int knc_peer_free(void *d_ptr);

/* ---- multi-GPU, ONE process (SURVEY 8e "API") -------------------------------------------------------
 * No reference counterpart (single-GPU).  knc_prepare_multi is kn_prepare (knearests.cu:235-344) over
 * `ngpus` ranks (<= 0: every device; more ranks than devices share devices): the host points are cut into ngpus
 * chunks, routed into z-slabs + halo of the library's own grid (knc_fast_dims) with the scatter kernel storing
 * straight into the other GPUs' memory (peer access), and every slab is built.  The handle then answers knc_solve,
 * knc_get_points / knc_get_permutation / knc_get_knearests (the single-GPU arrays, gathered from the slabs),
 * knc_get_stats, knc_get_info (no device pointers), knc_get_timings (slowest rank) and knc_free; a solve whose
 * halo proves too thin rebuilds with a wider one.  The reference's seven entry points do the same when the
 * environment variable KN_NUM_GPUS is set (> 1).  knc_solve_multi: prepare + solve + rows (+ permutation) in one
 * call; out_rows: HOST n*k, out_perm: HOST n or NULL. */
int knc_prepare_multi(const float *xyz, int n, int k, int ngpus, knc_problem **out);
int knc_solve_multi(const float *xyz, int n, int k, int ngpus, unsigned int *out_rows, unsigned int *out_perm);

/* Slab handles keep the single-GPU contract globally.  Because every rank holds its planes in global
 * order, local sorted position + a constant = GLOBAL sorted position (the constant: points in all lower
 * planes minus this handle's query_begin; the caller gets it from an all-gather of the ranks' query
 * counts).  knc_set_index_offset makes the next knc_solve store rows in global sorted positions
 * directly (no extra pass); knc_get_permutation_global writes this rank's slice of the global
 * permutation, d_out[i] = d_ids[id_stride * perm[query_begin + i]] (DEVICE, query_count entries).  Concatenated over
 * the ranks, rows and permutation are exactly what one GPU would return (knearests.cu:57,144-147). */
int knc_set_index_offset(knc_problem *p, long long offset);
int knc_get_permutation_global(knc_problem *p, const unsigned int *d_ids, int id_stride, unsigned int *d_out);

/* Rows of a solved (slab) handle in GLOBAL ids: d_ids[id_stride * i] (DEVICE) is the global id of the handle's
 * i-th input point; d_out_rows (DEVICE, query_count*k) gets ids[perm[knn]], d_out_query_ids (DEVICE,
 * query_count, may be NULL) the global id of each row's query.  (The reference's index remap,
 * test_knearests.cu:155-160, extended by the global id table.) */
int knc_get_knearests_global(knc_problem *p, const unsigned int *d_ids, int id_stride, unsigned int *d_out_rows,
                             unsigned int *d_out_query_ids);

/* ---- for callers that must not allocate --------------------------------------------------------
 * No counterpart in the reference (its every entry point cudaMallocs, knearests.cu:205-231).
 * The grid build's exclusive scan on its own: d_ptrs[c] = sum of d_counters[0..c), d_ptrs[cells] =
 * total; both arrays are DEVICE memory the caller owns (cells and cells + 1 ints), `stream` is a
 * cudaStream_t or NULL.  Meant for programs that have handed the whole GPU to their own memory pool
 * (frameworks, servers with a fixed arena) or sit in a latency-critical loop, where a cudaMalloc made
 * by a library is unwelcome: this call allocates nothing, on host or device.  Its scratch (one status
 * word per 4096 cells) is a static device array inside the library, ~33 KB held on every GPU that
 * loads the library until exit, which is why at most KNC_NOALLOC_MAX_CELLS cells are accepted (a
 * 256^3 grid, ~52 M points at the default density); larger sizes return KNC_ERR_INVALID.
 * The first call on a device resolves the scratch address: make it at start-up, before any stream
 * capture.  The address is remembered per device for the life of the process; it does not survive a
 * cudaDeviceReset (do not reset a device this entry point has been used on).  All calls on a device share the scratch; each clears what it uses first, so calls whose GPU
 * work could overlap (other streams, other threads) race and must be serialised by the caller.
 * Blocking: returns when `stream` has drained. */
#define KNC_NOALLOC_MAX_CELLS 16777216LL
int knc_scan_cells_noalloc(const int *d_counters, long long cells, int *d_ptrs, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* KNEARESTS_B200_C_H */
