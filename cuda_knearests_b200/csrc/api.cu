// api.cu -- host side of libknearests_b200.so: the C ABI (knearests_c.h) and, on top of it, the seven
// reference-compatible C++-linkage entry points (knearests.h).  Replaces the host half of the
// reference's knearests.cu: kn_prepare (:235-344), kn_firstbuild (:152-201), kn_solve (:348-392),
// kn_free (:396-408), kn_get_* (:410-438), kn_print_stats (:440-466), gpuMalloc* (:205-231).
//
// There is no CPU fallback anywhere in this file: every data-path step is a CUDA kernel launched from
// grid_build.cu / knn_query.cu, and a machine without a usable GPU gets KNC_ERR_CUDA (kn_*: exit).
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <limits.h>

#include <mutex>
#include <vector>

#include "kn_internal.h"
#include "params.h"

#define KN_VERSION "0.1.0"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

bool verbose() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("KN_VERBOSE");
        v = (e && atoi(e) != 0) ? 1 : 0;
    }
    return v == 1;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            rc = fail(e__ == cudaErrorMemoryAllocation ? KNC_ERR_NOMEM : KNC_ERR_CUDA, "%s: %s [%s:%d]", #call, \
                      cudaGetErrorString(e__), __FILE__, __LINE__);                                \
            goto out;                                                                              \
        }                                                                                          \
    } while (0)

// Stream-ordered allocations from a PRIVATE memory pool per device, which keeps freed memory (release threshold
// = never): a caller that prepares cloud after cloud (bench, services) pays the driver's page mapping once.  The
// device's default pool -- which other libraries in the process share -- is left untouched; knc_trim() hands the
// cached memory back to the driver.
cudaMemPool_t g_pool[64] = {nullptr};
std::mutex g_pool_lock;

cudaError_t device_pool(int device, cudaMemPool_t *out) {
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> hold(g_pool_lock);
    if (!g_pool[device]) {
        cudaMemPoolProps props;
        memset(&props, 0, sizeof(props));
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        cudaMemPool_t pool;
        cudaError_t e = cudaMemPoolCreate(&pool, &props);
        if (e != cudaSuccess) return e;
        unsigned long long keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        g_pool[device] = pool;
    }
    *out = g_pool[device];
    return cudaSuccess;
}

cudaError_t dev_alloc(void **p, size_t bytes, cudaStream_t st) { return kn_dev_alloc(p, bytes, st); }

// Streams and events are recycled: creating and destroying them are kernel-mode driver calls that cost
// a noticeable fraction of a millisecond each, more when several processes share the host (one per GPU).
struct StreamKit { cudaStream_t stream; cudaEvent_t ev[5]; };
std::mutex g_kit_lock;
std::vector<StreamKit> g_kits[64];

cudaError_t take_kit(int device, StreamKit *k) {
    {
        std::lock_guard<std::mutex> hold(g_kit_lock);
        if (device >= 0 && device < 64 && !g_kits[device].empty()) {
            *k = g_kits[device].back();
            g_kits[device].pop_back();
            return cudaSuccess;
        }
    }
    cudaError_t e = cudaStreamCreateWithFlags(&k->stream, cudaStreamNonBlocking);
    for (int i = 0; i < 5 && e == cudaSuccess; ++i) e = cudaEventCreate(&k->ev[i]);
    return e;
}

void give_kit(int device, const StreamKit &k) {
    std::lock_guard<std::mutex> hold(g_kit_lock);
    if (device >= 0 && device < 64 && g_kits[device].size() < 16) { g_kits[device].push_back(k); return; }
    cudaStreamDestroy(k.stream);
    for (int i = 0; i < 5; ++i) cudaEventDestroy(k.ev[i]);
}

int default_k() {
    const char *e = getenv("KN_NB_PLANES");
    if (e && atoi(e) > 0) return atoi(e);
    return DEFAULT_NB_PLANES;
}

void free_device_state(knc_problem *p) {
    cudaStream_t st = p->stream;
    if (p->d_retry) { cudaFreeAsync(p->d_retry, st); p->d_retry = NULL; }
    void *ptrs[] = {p->pub.d_cell_max, p->pub.d_permutation, p->pub.d_counters, p->pub.d_ptrs,
                    p->pub.d_globcounter, p->pub.d_stored_points, p->pub.d_knearests, p->d_points4, p->d_diag, p->d_tags};
    for (void *q : ptrs)
        if (q) cudaFreeAsync(q, st);
    if (st) cudaStreamSynchronize(st);
    if (p->kit_stream) {
        StreamKit k;
        k.stream = p->kit_stream;
        for (int i = 0; i < 5; ++i) k.ev[i] = p->ev[i];
        give_kit(p->device, k);
    }
    if (p->copy_stream) cudaStreamDestroy(p->copy_stream);
}

}  // namespace

// Allocation for every translation unit of the library: stream-ordered, from the library's own pool on the
// current device.  Freed with cudaFreeAsync.
cudaError_t kn_dev_alloc(void **p, size_t bytes, cudaStream_t st) {
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    cudaMemPool_t pool;
    if (e == cudaSuccess) e = device_pool(device, &pool);
    if (e != cudaSuccess) return e;
    return cudaMallocFromPoolAsync(p, bytes ? bytes : 1, pool, st);
}

// Lets the other translation units of the library report through knc_last_error().
int kn_set_error(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

extern "C" {

const char *knc_last_error(void) { return g_err; }

const char *knc_version(void) {
    static char v[96];
    snprintf(v, sizeof(v), "knearests_b200 %s sm_100a k_default=%d", KN_VERSION, DEFAULT_NB_PLANES);
    return v;
}

int knc_trim(void) {
    int rc = KNC_OK, device = 0;
    CU(cudaGetDevice(&device));
    CU(cudaDeviceSynchronize());
    {
        std::lock_guard<std::mutex> hold(g_pool_lock);
        if (device >= 0 && device < 64 && g_pool[device]) CU(cudaMemPoolTrimTo(g_pool[device], 0));
    }
out:
    return rc;
}

int knc_grid_dim(int n, float density) {
    if (!(density > 0.f)) density = 3.1f;
    // knearests.cu:249 -- f32 divide, double pow, round
    int sz = (int)round(pow((double)((float)n / density), (double)1.0f / 3.0));
    return sz < 1 ? 1 : sz;
}

// Event times of the last prepare / solve, once their events have completed (bit 0: prepare, bit 1: solve).
static void read_timings(knc_problem *p) {
    if (p->timings_pending & 1) {
        if (cudaEventSynchronize(p->ev[2]) == cudaSuccess) {
            cudaEventElapsedTime(&p->t.h2d_ms, p->ev[0], p->ev[1]);
            cudaEventElapsedTime(&p->t.build_ms, p->ev[1], p->ev[2]);
        }
        p->timings_pending &= ~1;
    }
}

// Radius of the sphere expected to hold k + 2.5 sqrt(k) + 2 points at `rho` points per unit volume: the first search
// radius of every query kernel (knn_query.cu).
static double radius_for(int k, double rho) {
    const double lambda = k + 2.5 * sqrt((double)k) + 2.0;
    return cbrt(3.0 * lambda / (4.0 * M_PI * rho));
}

int knc_fast_dims(int n, int k, int *dims) {
    if (!dims || n < 1) return fail(KNC_ERR_INVALID, "knc_fast_dims: bad argument");
    if (k <= 0) k = default_k();
    // y and z: cells 2 % wider than the first search radius, so that a query's sphere never reaches past the 3 x 3
    // x-rows of cells around its own (what the tile kernel stages in shared memory);  x: as many cells as points
    const double r0 = radius_for(k, (double)n / 1e9);
    int dyz = (int)floor(1000.0 / (1.02 * r0));
    if (dyz < 1) dyz = 1;
    if (dyz > 1024) dyz = 1024;
    long long dx = llround((double)n / ((double)dyz * dyz));
    if (dx < 1) dx = 1;
    if (dx > 1048576) dx = 1048576;
    dims[0] = (int)dx; dims[1] = dyz; dims[2] = dyz;
    return KNC_OK;
}

int knc_prepare_ex(const float *xyz, int n, const knc_options *opt, knc_problem **out) {
    int rc = KNC_OK;
    knc_options o;
    memset(&o, 0, sizeof(o));
    if (opt) memcpy(&o, opt, opt->struct_size && opt->struct_size < sizeof(o) ? opt->struct_size : sizeof(o));
    if (!out) return fail(KNC_ERR_INVALID, "knc_prepare: out is NULL");
    *out = NULL;
    if (!xyz) return fail(KNC_ERR_INVALID, "knc_prepare: points pointer is NULL");
    if (n < 1) return fail(KNC_ERR_INVALID, "knc_prepare: n = %d, need at least 1 point", n);
    const int k = o.k > 0 ? o.k : default_k();
    if (k > KNC_MAX_K) return fail(KNC_ERR_INVALID, "knc_prepare: k = %d exceeds KNC_MAX_K = %d", k, KNC_MAX_K);
    // The grid, in order of precedence: explicit resolutions (grid_dims, or the cube grid_dim), a cube from a number
    // of points per cell (the caller's `density`, else the environment's KN_GRID_DENSITY -- drop-in callers cannot
    // pass options; 3.1 gives the reference's grid, knearests.cu:249), else the grid the B200 kernels are built for
    // (knc_fast_dims: as fast as any cube and it changes no result, only dimx/dimy/dimz and the cell order).
    float density = 0.f;
    if (o.density > 0.f) density = o.density;
    else if (const char *e = getenv("KN_GRID_DENSITY")) { if (atof(e) > 0.) density = (float)atof(e); }
    const int stride = o.point_stride > 0 ? o.point_stride : 3;
    if (stride != 3 && stride != 4) return fail(KNC_ERR_INVALID, "knc_prepare: point_stride = %d, supported 3 or 4", stride);
    int dims[3];
    const bool dims_given = o.grid_dims[0] > 0 && o.grid_dims[1] > 0 && o.grid_dims[2] > 0;
    if (dims_given) { dims[0] = o.grid_dims[0]; dims[1] = o.grid_dims[1]; dims[2] = o.grid_dims[2]; }
    else if (o.grid_dim > 0) dims[0] = dims[1] = dims[2] = o.grid_dim;
    else if (density > 0.f) dims[0] = dims[1] = dims[2] = knc_grid_dim(n, density);
    else knc_fast_dims(n, k, dims);
    if ((long long)dims[0] * dims[1] * dims[2] > 2000000000ll)
        return fail(KNC_ERR_INVALID, "knc_prepare: grid %d x %d x %d too large", dims[0], dims[1], dims[2]);
    const int D = dims[2];  // planes: slabs cut the z axis
    int zb = 0, ze = D, ozb = 0, oze = D;
    if (o.z_begin || o.z_end || o.own_z_begin || o.own_z_end) {
        zb = o.z_begin; ze = o.z_end; ozb = o.own_z_begin; oze = o.own_z_end;
        if ((o.grid_dim <= 0 && !dims_given) || zb < 0 || ze > D || zb >= ze || ozb < zb || oze > ze || ozb > oze)
            return fail(KNC_ERR_INVALID, "knc_prepare: bad slab planes [%d,%d) own [%d,%d) of %d (grid_dim / grid_dims must be set)",
                        zb, ze, ozb, oze, D);
    }

    knc_problem *p = (knc_problem *)calloc(1, sizeof(knc_problem));
    if (!p) return fail(KNC_ERR_NOMEM, "knc_prepare: out of host memory");
    const long long cells = (long long)dims[0] * dims[1] * (ze - zb);
    const long long plane = (long long)dims[0] * dims[1];
    float *d_in = NULL;
    BuildScratch s;
    memset(&s, 0, sizeof(s));
    cudaStream_t st = NULL;
    GridGeom g;

    CU(cudaGetDevice(&p->device));
    {
        StreamKit kit;
        CU(take_kit(p->device, &kit));
        p->kit_stream = kit.stream;
        for (int i = 0; i < 5; ++i) p->ev[i] = kit.ev[i];
    }
    p->owns_stream = o.stream ? 0 : 1;
    p->stream = o.stream ? (cudaStream_t)o.stream : p->kit_stream;
    st = p->stream;
    p->k = k; p->density = density; p->search_scale = o.search_scale > 0.f ? o.search_scale : 1.f;
    p->algorithm = o.algorithm;
    p->z_begin = zb; p->z_end = ze; p->own_z_begin = ozb; p->own_z_end = oze;
    p->dims[0] = dims[0]; p->dims[1] = dims[1]; p->dims[2] = dims[2]; p->cells = cells;
    p->pub.allocated_points = n;
    p->pub.dimx = dims[0]; p->pub.dimy = dims[1]; p->pub.dimz = dims[2];
    p->pub.num_cell_offsets = KN_ROW_TABLE_SIZE;  // rows in the query kernels' (dj, dk) table; the reference: 29 791 cell offsets

    // persistent buffers (knearests.cu:313-335, without the `bufsize +=` over-allocation)
    CU(dev_alloc((void **)&p->pub.d_counters, sizeof(int) * (size_t)cells, st));
    CU(dev_alloc((void **)&p->pub.d_ptrs, sizeof(int) * (size_t)(cells + 1), st));
    CU(dev_alloc((void **)&p->pub.d_globcounter, sizeof(int), st));
    CU(dev_alloc((void **)&p->pub.d_cell_max, sizeof(float), st));
    CU(dev_alloc((void **)&p->d_diag, sizeof(unsigned) * KN_DIAG_WORDS, st));
    CU(dev_alloc((void **)&p->pub.d_permutation, sizeof(unsigned) * (size_t)n, st));
    CU(dev_alloc((void **)&p->pub.d_stored_points, sizeof(float) * 3 * (size_t)n, st));
    CU(dev_alloc((void **)&p->d_points4, sizeof(float4) * (size_t)n, st));
    if (stride == 4) CU(dev_alloc((void **)&p->d_tags, sizeof(unsigned) * (size_t)n, st));
    // temporaries of the build
    s.tile_count = kn_build_scratch_tiles(cells);
    CU(dev_alloc((void **)&s.d_recs, sizeof(float4) * (size_t)n, st));
    if (stride == 4) CU(dev_alloc((void **)&s.d_rec_tags, sizeof(unsigned) * (size_t)n, st));
    CU(dev_alloc((void **)&s.d_tile, sizeof(unsigned long long) * s.tile_count, st));
    CU(dev_alloc((void **)&s.d_ticket, sizeof(int) * 2, st));
    CU(dev_alloc((void **)&s.d_biglist, sizeof(int) * ((size_t)n / (KN_BIG_CELL + 1) + 1), st));
    if (n >= KN_BIN_MIN_POINTS && o.algorithm != 3) {  // algorithm 3: build without the z-band pre-pass (A/B, tests)
        CU(dev_alloc((void **)&s.d_bin_recs, sizeof(float4) * (size_t)n, st));
        CU(dev_alloc((void **)&s.d_bin_counts, sizeof(int) * kn_bin_counts_ints(n), st));
        if (stride == 4) CU(dev_alloc((void **)&s.d_bin_tags, sizeof(unsigned) * (size_t)n, st));
    }

    CU(cudaEventRecord(p->ev[0], st));
    if (o.points_on_device) {
        d_in = const_cast<float *>(xyz);
    } else {  // knearests.cu:308-311
        CU(dev_alloc((void **)&d_in, sizeof(float) * stride * (size_t)n, st));
        CU(cudaMemcpyAsync(d_in, xyz, sizeof(float) * stride * (size_t)n, cudaMemcpyHostToDevice, st));
    }
    g = kn_make_geom(dims, zb, ze);
    CU(cudaEventRecord(p->ev[1], st));
    CU(kn_launch_build(d_in, stride, n, g, cells, p->pub.d_counters, p->pub.d_ptrs, p->pub.d_globcounter,
                       p->pub.d_permutation, (float *)p->pub.d_stored_points, p->d_points4, p->d_tags, s, st));
    CU(cudaEventRecord(p->ev[2], st));
    CU(cudaMemsetAsync(p->pub.d_cell_max, 0, sizeof(float), st));
    {   // the queries are the points of the owned planes: one contiguous run of sorted positions
        int range[2] = {0, n};
        const bool slab = ozb != zb || oze != ze;
        if (slab && o.query_hint) {  // the caller knows (from the routing counts): no device->host read, no wait
            range[0] = o.query_begin_hint;
            range[1] = o.query_begin_hint + o.query_count_hint;
            if (range[0] < 0 || o.query_count_hint < 0 || range[1] > n) {
                rc = fail(KNC_ERR_INVALID, "knc_prepare: query hint [%d, +%d) does not fit %d points", o.query_begin_hint,
                          o.query_count_hint, n);
                goto out;
            }
        } else if (slab) {
            if (!p->owns_stream) {
                rc = fail(KNC_ERR_INVALID, "knc_prepare: a slab on the caller's stream needs query_hint (no host wait there)");
                goto out;
            }
            CU(cudaMemcpyAsync(&range[0], p->pub.d_ptrs + (size_t)(ozb - zb) * plane, sizeof(int),
                               cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(&range[1], p->pub.d_ptrs + (size_t)(oze - zb) * plane, sizeof(int),
                               cudaMemcpyDeviceToHost, st));
        }
        if (p->owns_stream) CU(cudaStreamSynchronize(st));  // kn_prepare returns with the grid built (knearests.cu:196)
        p->query_begin = range[0];
        p->query_count = range[1] - range[0];
    }
    CU(dev_alloc((void **)&p->pub.d_knearests, sizeof(unsigned) * (size_t)(p->query_count > 0 ? p->query_count : 1) * k, st));
    p->timings_pending |= 1;  // ev[0..2] are read when somebody asks (knc_get_timings), not here: that would wait
    if (p->owns_stream) read_timings(p);
    if (verbose() && p->owns_stream) {
        fprintf(stderr, "kn_firstbuild: %g msec (n=%d, grid %dx%dx%d planes [%d,%d), k=%d, h2d %g msec)\n",
                p->t.build_ms, n, dims[0], dims[1], dims[2], zb, ze, k, p->t.h2d_ms);
        // what the handle keeps on the device (knearests.cu:342 prints its own running total the same way)
        const size_t held = sizeof(int) * (size_t)(2 * cells + 1) + sizeof(int) + sizeof(float) + sizeof(unsigned) * KN_DIAG_WORDS +
                            (sizeof(unsigned) + sizeof(float) * 3 + sizeof(float4)) * (size_t)n +
                            (stride == 4 ? sizeof(unsigned) * (size_t)n : 0) +
                            sizeof(unsigned) * (size_t)(p->query_count > 0 ? p->query_count : 1) * k;
        fprintf(stderr, "GPU memory used: %zu Mb\n", held / 1048576);
    }
out:
    if (!o.points_on_device && d_in) cudaFreeAsync(d_in, st);
    {
        void *tmp[] = {s.d_recs, s.d_rec_tags, s.d_tile, s.d_ticket, s.d_biglist, s.d_bin_recs, s.d_bin_counts, s.d_bin_tags};
        for (void *q : tmp)
            if (q) cudaFreeAsync(q, st);
    }
    if (rc != KNC_OK) {
        free_device_state(p);
        free(p);
        return rc;
    }
    *out = p;
    return KNC_OK;
}

int knc_prepare(const float *xyz, int n, int k, knc_problem **out) {
    knc_options o;
    memset(&o, 0, sizeof(o));
    o.struct_size = sizeof(o);
    o.k = k;
    return knc_prepare_ex(xyz, n, &o, out);
}

static double first_radius(const knc_problem *p) {
    const double volume = 1000.0 * 1000.0 * (1000.0 * (p->z_end - p->z_begin) / p->dims[2]);
    return radius_for(p->k, (double)p->pub.allocated_points / volume) * p->search_scale;
}

// Second half of a solve: wait for the stream, read the diagnostics, report.
static int solve_finish_impl(knc_problem *p, const char *who) {
    int rc = KNC_OK;
    unsigned diag[KN_DIAG_WORDS] = {0};
    cudaStream_t st = p->stream;
    if (!(p->timings_pending & 2)) return fail(KNC_ERR_STATE, "%s: no solve in flight", who);
    CU(cudaMemcpyAsync(diag, p->d_diag, sizeof(diag), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));  // blocking, like knearests.cu:373
    if (p->copy_stream) CU(cudaStreamSynchronize(p->copy_stream));
    CU(cudaEventElapsedTime(&p->t.solve_ms, p->ev[3], p->ev[4]));
    p->timings_pending &= ~2;
    read_timings(p);
    p->t.restarts = diag[0];  // restarts inside the general kernel + queries it took over from the fast one
    p->t.handed_over_blocks = diag[4];  // last chunk's (the counter restarts with every chunk)
    p->t.second_retries = diag[5];      // likewise
    memcpy(&p->t.max_radius2, &diag[1], sizeof(float));
    p->solved = 1;
    if (verbose()) {
        const GridGeom g = kn_make_geom(p->dims, p->z_begin, p->z_end);
        fprintf(stderr, "kn_solve: %g msec\n", p->t.solve_ms);
        fprintf(stderr, "Max search radius: %g cells (first guess %g cells, %u of %d queries restarted)\n",
                sqrt((double)p->t.max_radius2) / g.wy, first_radius(p) / g.wy, diag[0], p->query_count);
    }
    if (diag[2])
        rc = fail(KNC_ERR_HALO, "%s: %u queries reach beyond the halo planes [%d,%d) of this slab", who, diag[2],
                  p->z_begin, p->z_end);
out:
    if (p->d_retry) { cudaFreeAsync(p->d_retry, st); p->d_retry = NULL; }
    return rc;
}

// The query, optionally streamed to the host: the query range is cut into `chunks` pieces; the rows of
// piece c travel device->host on a second stream while piece c+1 is being computed.  `finish`: wait and report
// (knc_solve); otherwise only enqueue (knc_solve_async; knc_solve_finish does the rest).
static int solve_impl(knc_problem *p, unsigned int *out_knn_host, int chunks, const char *who, bool finish) {
    int rc = KNC_OK;
    if (!p) return fail(KNC_ERR_INVALID, "%s: NULL handle", who);
    if (p->multi) {  // a handle that spans several GPUs: every slab solves, blocking
        if (!finish) return fail(KNC_ERR_STATE, "%s: not available on a multi-GPU handle", who);
        rc = kn_multi_solve(p);
        if (rc == KNC_OK && out_knn_host) rc = kn_multi_get(p, 2, out_knn_host);
        return rc;
    }
    if (p->timings_pending & 2) return fail(KNC_ERR_STATE, "%s: the previous knc_solve_async was not finished", who);
    cudaStream_t st = p->stream;
    const GridGeom g = kn_make_geom(p->dims, p->z_begin, p->z_end);
    // first-guess radius: sphere expected to hold k + 2.5 sqrt(k) + 2 points at the cloud's mean density
    const double r0 = first_radius(p);
    const int force_general = p->algorithm;  // 0 auto, 1 general kernel only, 2 insertion kernel for every k
    if (chunks < 1 || !out_knn_host) chunks = 1;
    if (chunks > 64) chunks = 64;
    const int per = (p->query_count + chunks - 1) / chunks;
    cudaEvent_t done[64];
    int made = 0;
    // scratch of the solve: queries the fast kernels hand to the general kernel, blocks the tile kernel hands on
    CU(dev_alloc((void **)&p->d_retry, sizeof(int) * kn_retry_ints(per), st));
    if (out_knn_host && !p->copy_stream) CU(cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking));
    CU(cudaMemsetAsync(p->d_diag, 0, sizeof(unsigned) * KN_DIAG_WORDS, st));
    CU(cudaEventRecord(p->ev[3], st));
    for (int c = 0; c < chunks; ++c) {
        const int q0 = c * per, qn = (p->query_count - q0) < per ? (p->query_count - q0) : per;
        if (qn <= 0) break;
        unsigned int *rows = p->pub.d_knearests + (size_t)q0 * p->k;
        if (c) CU(cudaMemsetAsync(p->d_diag + 3, 0, sizeof(unsigned) * 3, st));  // the retry and block lists start empty
        CU(kn_launch_query(p->d_points4, p->pub.d_ptrs, g, p->pub.allocated_points, p->k, p->query_begin + q0, qn,
                           (float)(r0 * r0), rows, p->d_diag, p->d_retry, force_general, p->index_offset, NULL, st));
        if (out_knn_host) {
            CU(cudaEventCreateWithFlags(&done[made], cudaEventDisableTiming));
            ++made;
            CU(cudaEventRecord(done[made - 1], st));
            CU(cudaStreamWaitEvent(p->copy_stream, done[made - 1], 0));
            CU(cudaMemcpyAsync(out_knn_host + (size_t)q0 * p->k, rows, sizeof(unsigned) * (size_t)qn * p->k,
                               cudaMemcpyDeviceToHost, p->copy_stream));
        }
    }
    CU(cudaEventRecord(p->ev[4], st));
    CU(cudaMemcpyAsync(p->pub.d_cell_max, p->d_diag + 1, sizeof(float), cudaMemcpyDeviceToDevice, st));
    p->timings_pending |= 2;
out:
    if (rc == KNC_OK && finish) rc = solve_finish_impl(p, who);
    else if (rc != KNC_OK && p->d_retry) { cudaFreeAsync(p->d_retry, st); p->d_retry = NULL; }
    for (int i = 0; i < made; ++i) cudaEventDestroy(done[i]);  // (the events' work is enqueued or done: safe to drop)
    return rc;
}

int knc_solve_async(knc_problem *p) { return solve_impl(p, NULL, 1, "knc_solve_async", false); }

int knc_solve_finish(knc_problem *p) {
    if (!p) return fail(KNC_ERR_INVALID, "knc_solve_finish: NULL handle");
    return solve_finish_impl(p, "knc_solve_finish");
}

int knc_halo_flag_ptr(knc_problem *p, const unsigned int **d_flag) {
    if (!p || !d_flag) return fail(KNC_ERR_INVALID, "knc_halo_flag_ptr: NULL argument");
    *d_flag = p->d_diag + 2;
    return KNC_OK;
}

int knc_solve(knc_problem *p) { return solve_impl(p, NULL, 1, "knc_solve", true); }

int knc_solve_to_host(knc_problem *p, unsigned int *out_knn, int chunks) {
    if (!out_knn) return fail(KNC_ERR_INVALID, "knc_solve_to_host: out_knn is NULL");
    return solve_impl(p, out_knn, chunks > 0 ? chunks : 8, "knc_solve_to_host", true);
}


int knc_get_dists(knc_problem *p, float *out_d2) {
    int rc = KNC_OK;
    float *d_out = NULL;
    if (!p || !out_d2) return fail(KNC_ERR_INVALID, "knc_get_dists: NULL argument");
    if (p->multi) return fail(KNC_ERR_STATE, "knc_get_dists: not available on a multi-GPU handle (its slabs live on several devices)");
    if (!p->solved) return fail(KNC_ERR_STATE, "knc_get_dists: call knc_solve first");
    {
        const size_t bytes = sizeof(float) * (size_t)p->query_count * p->k;
        CU(dev_alloc((void **)&d_out, bytes, p->stream));
        CU(kn_launch_rows_dists(p->d_points4, p->pub.d_knearests, NULL, p->query_begin, p->index_offset, p->query_count,
                                p->k, d_out, p->stream));
        CU(cudaMemcpyAsync(out_d2, d_out, bytes, cudaMemcpyDeviceToHost, p->stream));
        CU(cudaStreamSynchronize(p->stream));
    }
out:
    if (d_out) cudaFreeAsync(d_out, p->stream);
    return rc;
}

int knc_query_points(knc_problem *p, const float *xyz, int m, unsigned int *out_knn, float *out_d2) {
    int rc = KNC_OK;
    float *d_q = NULL, *d_d2 = NULL;
    unsigned *d_rows = NULL, *d_diag = NULL;
    unsigned diag[4] = {0, 0, 0, 0};
    if (!p || !xyz || !out_knn || m < 1) return fail(KNC_ERR_INVALID, "knc_query_points: bad argument");
    if (p->multi) return fail(KNC_ERR_STATE, "knc_query_points: not available on a multi-GPU handle (its slabs live on several devices)");
    {
        cudaStream_t st = p->stream;
        const GridGeom g = kn_make_geom(p->dims, p->z_begin, p->z_end);
        const double r0 = first_radius(p);
        const size_t rows_bytes = sizeof(unsigned) * (size_t)m * p->k;
        CU(dev_alloc((void **)&d_q, sizeof(float) * 3 * (size_t)m, st));
        CU(dev_alloc((void **)&d_rows, rows_bytes, st));
        CU(dev_alloc((void **)&d_diag, sizeof(unsigned) * KN_DIAG_WORDS, st));
        CU(cudaMemcpyAsync(d_q, xyz, sizeof(float) * 3 * (size_t)m, cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(d_diag, 0, sizeof(unsigned) * KN_DIAG_WORDS, st));
        CU(kn_launch_query(p->d_points4, p->pub.d_ptrs, g, p->pub.allocated_points, p->k, 0, m, (float)(r0 * r0), d_rows,
                           d_diag, NULL, 1, p->index_offset, d_q, st));
        CU(cudaMemcpyAsync(out_knn, d_rows, rows_bytes, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(diag, d_diag, sizeof(diag), cudaMemcpyDeviceToHost, st));
        if (out_d2) {
            CU(dev_alloc((void **)&d_d2, sizeof(float) * (size_t)m * p->k, st));
            CU(kn_launch_rows_dists(p->d_points4, d_rows, d_q, 0, p->index_offset, m, p->k, d_d2, st));
            CU(cudaMemcpyAsync(out_d2, d_d2, sizeof(float) * (size_t)m * p->k, cudaMemcpyDeviceToHost, st));
        }
        CU(cudaStreamSynchronize(st));
        if (diag[2])
            rc = fail(KNC_ERR_HALO, "knc_query_points: %u queries reach beyond the planes [%d,%d) of this slab", diag[2],
                      p->z_begin, p->z_end);
    }
out:
    if (d_q) cudaFreeAsync(d_q, p->stream);
    if (d_rows) cudaFreeAsync(d_rows, p->stream);
    if (d_d2) cudaFreeAsync(d_d2, p->stream);
    if (d_diag) cudaFreeAsync(d_diag, p->stream);
    return rc;
}

int knc_set_index_offset(knc_problem *p, long long offset) {
    if (!p) return fail(KNC_ERR_INVALID, "knc_set_index_offset: NULL handle");
    if (p->multi) return fail(KNC_ERR_STATE, "knc_set_index_offset: not available on a multi-GPU handle (its slabs live on several devices)");
    if (offset < 0 || offset + p->pub.allocated_points > 0xFFFFFFFEll)
        return fail(KNC_ERR_INVALID, "knc_set_index_offset: %lld + %d points does not fit 32-bit indices", offset,
                    p->pub.allocated_points);
    p->index_offset = (unsigned)offset;
    return KNC_OK;
}

int knc_free(knc_problem *p) {
    if (!p) return KNC_OK;
    if (p->multi) {
        kn_multi_free(p);
        free(p);
        return KNC_OK;
    }
    free_device_state(p);
    free(p);
    return KNC_OK;
}

static int copy_out(knc_problem *p, void *dst, const void *src, size_t bytes, const char *what) {
    int rc = KNC_OK;
    if (!p || !dst) return fail(KNC_ERR_INVALID, "%s: NULL argument", what);
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
out:
    return rc;
}

int knc_get_points(knc_problem *p, float *out_xyz) {
    if (p && p->multi) return out_xyz ? kn_multi_get(p, 0, out_xyz) : fail(KNC_ERR_INVALID, "knc_get_points: NULL argument");
    return copy_out(p, out_xyz, p ? p->pub.d_stored_points : NULL, p ? sizeof(float) * 3 * (size_t)p->pub.allocated_points : 0,
                    "knc_get_points");
}

int knc_get_permutation(knc_problem *p, unsigned int *out_perm) {
    if (p && p->multi) return out_perm ? kn_multi_get(p, 1, out_perm) : fail(KNC_ERR_INVALID, "knc_get_permutation: NULL argument");
    return copy_out(p, out_perm, p ? p->pub.d_permutation : NULL, p ? sizeof(unsigned) * (size_t)p->pub.allocated_points : 0,
                    "knc_get_permutation");
}

int knc_get_knearests(knc_problem *p, unsigned int *out_knn) {
    if (p && !p->solved) return fail(KNC_ERR_STATE, "knc_get_knearests: call knc_solve first");
    if (p && p->multi) return out_knn ? kn_multi_get(p, 2, out_knn) : fail(KNC_ERR_INVALID, "knc_get_knearests: NULL argument");
    return copy_out(p, out_knn, p ? p->pub.d_knearests : NULL, p ? sizeof(unsigned) * (size_t)p->query_count * p->k : 0,
                    "knc_get_knearests");
}

int knc_get_knearests_original(knc_problem *p, unsigned int *out_neighbors) {
    int rc = KNC_OK;
    unsigned *d_tmp = NULL;
    if (!p || !out_neighbors) return fail(KNC_ERR_INVALID, "knc_get_knearests_original: NULL argument");
    if (p->multi) return fail(KNC_ERR_STATE, "knc_get_knearests_original: not available on a multi-GPU handle (its slabs live on several devices)");
    if (!p->solved) return fail(KNC_ERR_STATE, "knc_get_knearests_original: call knc_solve first");
    if (p->query_count != p->pub.allocated_points)
        return fail(KNC_ERR_INVALID, "knc_get_knearests_original: not available in slab mode");
    {
        const size_t bytes = sizeof(unsigned) * (size_t)p->pub.allocated_points * p->k;
        CU(dev_alloc((void **)&d_tmp, bytes, p->stream));
        CU(kn_launch_remap_original(p->pub.d_knearests, p->pub.d_permutation, p->pub.allocated_points, p->k, d_tmp,
                                    p->stream));
        CU(cudaMemcpyAsync(out_neighbors, d_tmp, bytes, cudaMemcpyDeviceToHost, p->stream));
        CU(cudaStreamSynchronize(p->stream));
    }
out:
    if (d_tmp) cudaFreeAsync(d_tmp, p->stream);
    return rc;
}

int knc_get_tags(knc_problem *p, const unsigned int **d_tags) {
    if (!p || !d_tags) return fail(KNC_ERR_INVALID, "knc_get_tags: NULL argument");
    if (p->multi) return fail(KNC_ERR_STATE, "knc_get_tags: not available on a multi-GPU handle (its slabs live on several devices)");
    if (!p->d_tags) return fail(KNC_ERR_STATE, "knc_get_tags: the handle was prepared with point_stride 3 (no 4th word)");
    *d_tags = p->d_tags;
    return KNC_OK;
}

int knc_get_stats(knc_problem *p, knc_stats *out) {
    int rc = KNC_OK;
    int *h = NULL;
    if (!p || !out) return fail(KNC_ERR_INVALID, "knc_get_stats: NULL argument");
    memset(out, 0, sizeof(*out));
    h = (int *)malloc(sizeof(int) * (size_t)p->cells);
    if (!h) return fail(KNC_ERR_NOMEM, "knc_get_stats: out of host memory");
    if (p->multi) {
        rc = kn_multi_counters(p, h);
        if (rc != KNC_OK) goto out;
    } else {
        CU(cudaMemcpyAsync(h, p->pub.d_counters, sizeof(int) * (size_t)p->cells, cudaMemcpyDeviceToHost, p->stream));
        CU(cudaStreamSynchronize(p->stream));
    }
    out->cells = p->cells;
    out->min_per_cell = INT_MAX;
    for (long long c = 0; c < p->cells; c++) {  // knearests.cu:455-460
        const int v = h[c];
        out->histogram[v < 63 ? v : 63]++;
        if (v < out->min_per_cell) out->min_per_cell = v;
        if (v > out->max_per_cell) out->max_per_cell = v;
        out->total += v;
    }
    out->avg_per_cell = (double)p->pub.allocated_points / (double)p->cells;
out:
    free(h);
    return rc;
}

int knc_get_info(knc_problem *p, knc_info *out) {
    if (!p || !out) return fail(KNC_ERR_INVALID, "knc_get_info: NULL argument");
    memset(out, 0, sizeof(*out));
    out->n = p->pub.allocated_points; out->k = p->k;
    out->dimx = p->pub.dimx; out->dimy = p->pub.dimy; out->dimz = p->pub.dimz;
    out->z_begin = p->z_begin; out->z_end = p->z_end;
    out->query_begin = p->query_begin; out->query_count = p->query_count; out->solved = p->solved;
    out->d_points = (const float *)p->pub.d_stored_points;
    out->d_points4 = (const float *)p->d_points4;
    out->d_permutation = p->pub.d_permutation;
    out->d_counters = p->pub.d_counters;
    out->d_ptrs = p->pub.d_ptrs;
    out->d_knearests = p->pub.d_knearests;
    return KNC_OK;
}

int knc_get_timings(knc_problem *p, knc_timings *out) {
    if (!p || !out) return fail(KNC_ERR_INVALID, "knc_get_timings: NULL argument");
    if (!p->multi) read_timings(p);  // a prepare on the caller's stream leaves its event times unread until somebody asks
    *out = p->t;
    return KNC_OK;
}

int knc_generate_uniform_device(float *d_xyz, long long first, long long count, uint64_t seed) {
    int rc = KNC_OK;
    if (!d_xyz || count < 0) return fail(KNC_ERR_INVALID, "knc_generate_uniform_device: bad argument");
    CU(kn_launch_generate(d_xyz, first, count, seed, 0));
    CU(cudaStreamSynchronize(0));
out:
    return rc;
}

}  // extern "C"

// =====================================================================================================
// The reference's seven entry points (knearests.h), C++ linkage so that the mangled names match the
// reference's objects.  Error behaviour as in the reference: message on stderr, exit.
// =====================================================================================================
static void die_on(int rc, const char *where) {
    if (rc == KNC_OK) return;
    fprintf(stderr, "[%s] Failed (%s)!\n", where, knc_last_error());
    exit(EXIT_FAILURE);
}

kn_problem *kn_prepare(float3 *points, int numpoints) {
    knc_problem *p = NULL;
    const char *g = getenv("KN_NUM_GPUS");  // several GPUs for callers of the reference's API: the same seven calls
    if (g && atoi(g) > 1) die_on(knc_prepare_multi((const float *)points, numpoints, 0, atoi(g), &p), "kn_prepare");
    else die_on(knc_prepare((const float *)points, numpoints, 0, &p), "kn_prepare");
    // printed unconditionally by the reference (knearests.cu:288); here: the rows of the (dj, dk) table that takes
    // the place of its list of cell offsets
    fprintf(stderr, "num_cell_offsets = %d\n", p->pub.num_cell_offsets);
    return &p->pub;
}

void kn_solve(kn_problem *kn) {
    knc_problem *p = (knc_problem *)kn;
    die_on(knc_solve(p), "kn_solve");
    // knearests.cu:389, unconditional there: the farthest ring any query reached / the last ring of the table.  Here a
    // "ring" is a row of cells in y / z: 1 + the largest final search radius in y/z cell widths, over the table's 5.
    const double cells_reached = sqrt((double)p->t.max_radius2) / 1000. * (double)(p->dims[1] < p->dims[2] ? p->dims[1] : p->dims[2]) + 1.;
    fprintf(stderr, "Max visited ring: %g / %g\n", cells_reached, (sqrt((double)p->pub.num_cell_offsets) - 1.) / 2.);
}

void kn_free(kn_problem **kn) {
    if (!kn || !*kn) return;
    knc_free((knc_problem *)*kn);
    *kn = NULL;
}

float3 *kn_get_points(kn_problem *kn) {
    float3 *h = (float3 *)malloc(sizeof(float3) * (size_t)kn->allocated_points);
    die_on(knc_get_points((knc_problem *)kn, (float *)h), "kn_get_points");
    return h;
}

unsigned int *kn_get_permutation(kn_problem *kn) {
    unsigned int *h = (unsigned int *)malloc(sizeof(unsigned) * (size_t)kn->allocated_points);
    die_on(knc_get_permutation((knc_problem *)kn, h), "kn_get_permutation");
    return h;
}

unsigned int *kn_get_knearests(kn_problem *kn) {
    knc_problem *p = (knc_problem *)kn;
    unsigned int *h = (unsigned int *)malloc(sizeof(unsigned) * (size_t)(p->query_count > 0 ? p->query_count : 1) * p->k);
    die_on(knc_get_knearests(p, h), "kn_get_knearests");
    return h;
}

void kn_print_stats(kn_problem *kn) {
    knc_stats s;
    die_on(knc_get_stats((knc_problem *)kn, &s), "kn_print_stats");
    // same two kinds of lines as knearests.cu:461-464
    fprintf(stderr, "Grid:  points per cell: %d (min), %d (max), %g avg, total %lld\n", s.min_per_cell,
            s.max_per_cell, s.avg_per_cell, s.total);
    for (int c = 0; c < 64; c++)
        if (s.histogram[c]) fprintf(stderr, "[%d] => %lld\n", c, s.histogram[c]);
}
