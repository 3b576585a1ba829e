// multi.cu -- one kNN problem on several GPUs of a node from ONE process, behind the C ABI ("API":
// additive entry point + the environment variable KN_NUM_GPUS for callers of knearests.h).
//
// No reference counterpart: the reference is single-GPU.  Same decomposition as the multi-process
// driver (cuda_knearests_b200/slabs.py): z-slabs of cell planes of one global grid, every rank (one host thread
// per rank, device = rank modulo the number of devices, so that more ranks than GPUs also work) routes its
// index-contiguous chunk of the cloud with knc_route_begin_async / knc_route_scatter_async -- the scatter kernel
// stores each group straight into the destination GPU's receive buffer (peer access), no staging copy, no NCCL --
// then builds and queries its slab + halo.  What is a collective in the multi-process driver is a barrier between
// host threads and a few numbers in shared host memory here.  Rows come out in GLOBAL sorted positions; the ranks'
// rows, permutation slices and points, concatenated, are exactly what one GPU returns (tests/test_multi_gpu.py).
#include <cuda_runtime.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "kn_internal.h"
#include "params.h"

namespace {

struct Barrier {  // std::barrier is C++20
    std::mutex m;
    std::condition_variable cv;
    int waiting = 0, generation = 0, parties;
    explicit Barrier(int n) : parties(n) {}
    void arrive_and_wait() {
        std::unique_lock<std::mutex> lk(m);
        const int gen = generation;
        if (++waiting == parties) { waiting = 0; ++generation; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != generation; });
    }
};

}  // namespace

// The state of a multi-GPU handle (knc_problem.multi).
struct knc_multi {
    int world = 0, ndev = 0, n = 0, k = 0, halo = 0;
    int dims[3] = {0, 0, 0};
    std::vector<int> zsplit, device;
    std::vector<float *> d_chunk;        // every rank's index-contiguous part of the input, kept for a wider halo
    std::vector<long long> chunk_lo;     // first global index of every chunk (+ n at the end)
    std::vector<void *> recv;            // receive buffers (16-byte records), written by all ranks
    long long capacity = 0;              // records per receive buffer
    std::vector<knc_problem *> slab;     // one slab handle per rank (NULL: the rank owns no plane or no point)
    std::vector<long long> base, qcount; // global sorted position of every rank's first query; its queries
};

namespace {

int first_error(const std::vector<int> &rc, const std::vector<std::string> &msg, const char *who) {
    for (size_t r = 0; r < rc.size(); ++r)
        if (rc[r] != KNC_OK) return kn_set_error(rc[r], "%s: rank %zu: %s", who, r, msg[r].c_str());
    return KNC_OK;
}

#define RANK_CU(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess && rc[r] == KNC_OK) {                                                     \
            rc[r] = e__ == cudaErrorMemoryAllocation ? KNC_ERR_NOMEM : KNC_ERR_CUDA;                     \
            msg[r] = std::string(#call) + ": " + cudaGetErrorString(e__);                                \
        }                                                                                                \
    } while (0)
#define RANK_KN(call)                                            \
    do {                                                         \
        const int c__ = (call);                                  \
        if (c__ != KNC_OK && rc[r] == KNC_OK) { rc[r] = c__; msg[r] = knc_last_error(); } \
    } while (0)

// Route every chunk, build every slab.  All ranks run the same sequence; a failed rank keeps walking through the
// barriers (doing nothing) so that nobody is left waiting.
int build_slabs(knc_multi *M, const char *who) {
    const int W = M->world;
    std::vector<int> rc(W, KNC_OK);
    std::vector<std::string> msg(W);
    std::vector<long long> counts(3 * (size_t)W * W, 0);  // counts[s][0..3W): what rank s sends (records, own, low halo)
    Barrier bar(W);
    const int D = M->dims[2];
    auto rank_main = [&](int r) {
        RANK_CU(cudaSetDevice(M->device[r]));
        cudaStream_t st = nullptr;
        RANK_CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        long long *d_counts3 = nullptr;
        knc_route *route = nullptr;
        const int n_r = (int)(M->chunk_lo[r + 1] - M->chunk_lo[r]);
        RANK_CU(kn_dev_alloc((void **)&d_counts3, sizeof(long long) * 3 * W, st));
        if (rc[r] == KNC_OK)
            RANK_KN(knc_route_begin_async(M->d_chunk[r], n_r, (unsigned)M->chunk_lo[r], D, W, M->zsplit.data(), M->halo, st,
                                          d_counts3, &route));
        if (rc[r] == KNC_OK) {
            RANK_CU(cudaMemcpyAsync(&counts[3 * (size_t)W * r], d_counts3, sizeof(long long) * 3 * W, cudaMemcpyDeviceToHost, st));
            RANK_CU(cudaStreamSynchronize(st));
        }
        bar.arrive_and_wait();  // ---- every rank's counts are in `counts` ("the all-gather")
        bool all_ok = true;
        for (int s = 0; s < W; ++s) all_ok = all_ok && rc[s] == KNC_OK;
        long long incoming = 0, qbegin = 0, qcount = 0, base = 0;
        std::vector<long long> first(W, 0);
        if (all_ok) {
            for (int s = 0; s < W; ++s) {
                const long long *row = &counts[3 * (size_t)W * s];
                incoming += row[r];
                qcount += row[W + r];
                qbegin += row[2 * W + r];
                for (int d = 0; d < r; ++d) base += row[W + d];
                if (s < r)
                    for (int d = 0; d < W; ++d) first[d] += row[d];
            }
            if (incoming > M->capacity && rc[r] == KNC_OK) {
                rc[r] = KNC_ERR_NOMEM;
                msg[r] = "receive buffer holds " + std::to_string(M->capacity) + " records, " + std::to_string(incoming) +
                         " are on their way (a very uneven cloud: use fewer GPUs)";
            }
        }
        bar.arrive_and_wait();  // ---- capacity verdicts are in
        for (int s = 0; s < W; ++s) all_ok = all_ok && rc[s] == KNC_OK;
        if (all_ok && route) {
            RANK_KN(knc_route_scatter_async(route, M->recv.data(), first.data(), st));
            RANK_CU(cudaStreamSynchronize(st));  // this rank's stores (also those into peers' memory) have landed
        }
        if (route) knc_route_end(route);
        if (d_counts3) cudaFreeAsync(d_counts3, st);
        bar.arrive_and_wait();  // ---- every receive buffer is complete
        for (int s = 0; s < W; ++s) all_ok = all_ok && rc[s] == KNC_OK;
        M->base[r] = base;
        M->qcount[r] = qcount;
        const int zlo = M->zsplit[r], zhi = M->zsplit[r + 1];
        if (all_ok && zhi > zlo && incoming > 0) {
            knc_options o;
            memset(&o, 0, sizeof(o));
            o.struct_size = sizeof(o);
            o.k = M->k;
            o.points_on_device = 1;
            o.point_stride = 4;
            o.grid_dims[0] = M->dims[0]; o.grid_dims[1] = M->dims[1]; o.grid_dims[2] = M->dims[2];
            o.z_begin = zlo - M->halo > 0 ? zlo - M->halo : 0;
            o.z_end = zhi + M->halo < D ? zhi + M->halo : D;
            o.own_z_begin = zlo; o.own_z_end = zhi;
            o.query_hint = 1; o.query_begin_hint = (int)qbegin; o.query_count_hint = (int)qcount;
            knc_problem *p = nullptr;
            RANK_KN(knc_prepare_ex((const float *)M->recv[r], (int)incoming, &o, &p));
            if (p) RANK_KN(knc_set_index_offset(p, base - qbegin));
            M->slab[r] = p;
        }
        if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
    };
    std::vector<std::thread> th;
    for (int r = 0; r < W; ++r) th.emplace_back(rank_main, r);
    for (auto &t : th) t.join();
    return first_error(rc, msg, who);
}

void free_slabs(knc_multi *M) {
    for (int r = 0; r < M->world; ++r)
        if (M->slab[r]) {
            cudaSetDevice(M->device[r]);
            knc_free(M->slab[r]);
            M->slab[r] = nullptr;
        }
}

}  // namespace

void kn_multi_free(knc_problem *parent) {
    knc_multi *M = parent->multi;
    if (!M) return;
    int keep = 0;
    cudaGetDevice(&keep);
    free_slabs(M);
    for (int r = 0; r < M->world; ++r) {
        cudaSetDevice(M->device[r]);
        if (M->d_chunk[r]) cudaFree(M->d_chunk[r]);
        if (M->recv[r]) cudaFree(M->recv[r]);
    }
    cudaSetDevice(keep);
    delete M;
    parent->multi = nullptr;
}

extern "C" int knc_prepare_multi(const float *xyz, int n, int k, int ngpus, knc_problem **out) {
    if (!out) return kn_set_error(KNC_ERR_INVALID, "knc_prepare_multi: out is NULL");
    *out = nullptr;
    if (!xyz || n < 1) return kn_set_error(KNC_ERR_INVALID, "knc_prepare_multi: bad points / n");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return kn_set_error(KNC_ERR_CUDA, "knc_prepare_multi: no CUDA device");
    if (ngpus <= 0) ngpus = ndev;
    if (ngpus > 32) return kn_set_error(KNC_ERR_INVALID, "knc_prepare_multi: at most 32 ranks");
    knc_problem *parent = (knc_problem *)calloc(1, sizeof(knc_problem));
    if (!parent) return kn_set_error(KNC_ERR_NOMEM, "knc_prepare_multi: out of host memory");
    knc_multi *M = new knc_multi;
    parent->multi = M;
    int keep = 0;
    cudaGetDevice(&keep);
    parent->device = keep;
    if (k <= 0) {
        const char *e = getenv("KN_NB_PLANES");
        k = (e && atoi(e) > 0) ? atoi(e) : DEFAULT_NB_PLANES;
    }
    M->world = ngpus; M->ndev = ndev; M->n = n; M->k = k;
    knc_fast_dims(n, k, M->dims);
    const int D = M->dims[2], W = ngpus;
    // 1.75 x the first search radius, in planes: 2 on this grid; verified after the solve and widened if a query reached past it
    const double lambda = k + 2.5 * sqrt((double)k) + 2.0;
    const double r0 = cbrt(3.0 * lambda / (4.0 * M_PI * ((double)n / 1e9)));
    M->halo = (int)ceil(1.75 * r0 / (1000.0 / D));
    if (M->halo < 2) M->halo = 2;
    M->zsplit.resize(W + 1);
    for (int r = 0; r <= W; ++r) M->zsplit[r] = (int)((long long)r * D / W);
    M->device.resize(W); M->d_chunk.assign(W, nullptr); M->recv.assign(W, nullptr); M->slab.assign(W, nullptr);
    M->chunk_lo.resize(W + 1); M->base.assign(W, 0); M->qcount.assign(W, 0);
    for (int r = 0; r <= W; ++r) M->chunk_lo[r] = (long long)r * n / W;
    M->capacity = (long long)(1.35 * n / W) + (1 << 16) + 8ll * M->halo * ((long long)n / D + 1);
    parent->k = k; parent->dims[0] = M->dims[0]; parent->dims[1] = M->dims[1]; parent->dims[2] = M->dims[2];
    parent->z_end = parent->own_z_end = D;
    parent->pub.allocated_points = n;
    parent->pub.dimx = M->dims[0]; parent->pub.dimy = M->dims[1]; parent->pub.dimz = M->dims[2];
    parent->pub.num_cell_offsets = KN_ROW_TABLE_SIZE;
    parent->query_count = n;
    parent->cells = (long long)M->dims[0] * M->dims[1] * D;

    int rc = KNC_OK;
    for (int r = 0; r < W && rc == KNC_OK; ++r) {  // devices, peer access, buffers, the chunks (knearests.cu:308-311)
        M->device[r] = r % ndev;
        cudaError_t e = cudaSetDevice(M->device[r]);
        for (int j = 0; j < ndev && j < W && e == cudaSuccess; ++j)
            if (j != M->device[r]) {
                const cudaError_t pe = cudaDeviceEnablePeerAccess(j, 0);
                if (pe == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (pe != cudaSuccess) e = pe;
            }
        const size_t n_r = (size_t)(M->chunk_lo[r + 1] - M->chunk_lo[r]);
        if (e == cudaSuccess) e = cudaMalloc((void **)&M->d_chunk[r], sizeof(float) * 3 * (n_r ? n_r : 1));
        if (e == cudaSuccess) e = cudaMalloc(&M->recv[r], 16 * (size_t)M->capacity);
        if (e == cudaSuccess && n_r)
            e = cudaMemcpy(M->d_chunk[r], xyz + 3 * M->chunk_lo[r], sizeof(float) * 3 * n_r, cudaMemcpyHostToDevice);
        if (e != cudaSuccess)
            rc = kn_set_error(e == cudaErrorMemoryAllocation ? KNC_ERR_NOMEM : KNC_ERR_CUDA, "knc_prepare_multi: rank %d on device %d: %s",
                              r, M->device[r], cudaGetErrorString(e));
    }
    if (rc == KNC_OK) rc = build_slabs(M, "knc_prepare_multi");
    cudaSetDevice(keep);
    if (rc != KNC_OK) {
        kn_multi_free(parent);
        free(parent);
        return rc;
    }
    *out = parent;
    return KNC_OK;
}

// kn_solve on a multi-GPU handle: every rank's query at once; if some query reached beyond its halo planes, every
// slab is rebuilt with a wider halo and the solve repeated (never silently wrong).
int kn_multi_solve(knc_problem *parent) {
    knc_multi *M = parent->multi;
    const int W = M->world;
    int keep = 0;
    cudaGetDevice(&keep);
    int result = KNC_OK;
    for (int attempt = 0; attempt < 6; ++attempt) {
        std::vector<int> rc(W, KNC_OK);
        std::vector<std::string> msg(W);
        std::vector<std::thread> th;
        for (int r = 0; r < W; ++r)
            th.emplace_back([&, r] {
                if (!M->slab[r]) return;
                RANK_CU(cudaSetDevice(M->device[r]));
                if (rc[r] == KNC_OK) RANK_KN(knc_solve(M->slab[r]));
            });
        for (auto &t : th) t.join();
        bool halo_miss = false;
        for (int r = 0; r < W; ++r)
            if (rc[r] == KNC_ERR_HALO) { halo_miss = true; rc[r] = KNC_OK; }
        result = first_error(rc, msg, "knc_solve (multi-GPU)");
        if (result != KNC_OK || !halo_miss) break;
        free_slabs(M);
        M->halo += 2;
        result = build_slabs(M, "knc_solve (multi-GPU, wider halo)");
        if (result != KNC_OK) break;
        if (attempt == 5) result = kn_set_error(KNC_ERR_HALO, "knc_solve (multi-GPU): halo of %d planes still too thin", M->halo);
    }
    float build_ms = 0.f, solve_ms = 0.f;
    unsigned restarts = 0;
    for (int r = 0; r < W; ++r)
        if (M->slab[r]) {
            knc_timings t;
            cudaSetDevice(M->device[r]);
            knc_get_timings(M->slab[r], &t);
            build_ms = t.build_ms > build_ms ? t.build_ms : build_ms;  // the ranks work side by side: the slowest counts
            solve_ms = t.solve_ms > solve_ms ? t.solve_ms : solve_ms;
            restarts += t.restarts;
            if (t.max_radius2 > parent->t.max_radius2) parent->t.max_radius2 = t.max_radius2;
        }
    parent->t.build_ms = build_ms; parent->t.solve_ms = solve_ms; parent->t.restarts = restarts;
    cudaSetDevice(keep);
    if (result == KNC_OK) parent->solved = 1;
    return result;
}

// what: 0 = sorted points (3 floats), 1 = permutation, 2 = neighbour rows (k words) -- of the owned queries of every
// rank, one after the other: the single-GPU arrays (knearests.cu:410-438).
int kn_multi_get(knc_problem *parent, int what, void *dst) {
    knc_multi *M = parent->multi;
    int keep = 0, rc = KNC_OK;
    cudaGetDevice(&keep);
    for (int r = 0; r < M->world && rc == KNC_OK; ++r) {
        knc_problem *p = M->slab[r];
        if (!p || p->query_count == 0) continue;
        cudaSetDevice(M->device[r]);
        const size_t q0 = (size_t)p->query_begin, qn = (size_t)p->query_count, b = (size_t)M->base[r];
        cudaError_t e = cudaSuccess;
        if (what == 0)
            e = cudaMemcpy((float *)dst + 3 * b, (const float *)p->pub.d_stored_points + 3 * q0, sizeof(float) * 3 * qn, cudaMemcpyDeviceToHost);
        else if (what == 1)  // the records' 4th words -- the global ids -- in sorted order: the global permutation
            e = cudaMemcpy((unsigned *)dst + b, p->d_tags + q0, sizeof(unsigned) * qn, cudaMemcpyDeviceToHost);
        else
            e = cudaMemcpy((unsigned *)dst + b * (size_t)M->k, p->pub.d_knearests, sizeof(unsigned) * qn * M->k, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = kn_set_error(KNC_ERR_CUDA, "multi-GPU read-back, rank %d: %s", r, cudaGetErrorString(e));
    }
    cudaSetDevice(keep);
    return rc;
}

// counters of every rank's OWNED planes, one rank after the other = the counters of the whole grid
int kn_multi_counters(knc_problem *parent, int *h) {
    knc_multi *M = parent->multi;
    int keep = 0, rc = KNC_OK;
    cudaGetDevice(&keep);
    const size_t plane = (size_t)M->dims[0] * M->dims[1];
    memset(h, 0, sizeof(int) * plane * M->dims[2]);
    for (int r = 0; r < M->world && rc == KNC_OK; ++r) {
        knc_problem *p = M->slab[r];
        if (!p) continue;
        cudaSetDevice(M->device[r]);
        const size_t planes = (size_t)(p->own_z_end - p->own_z_begin);
        const cudaError_t e = cudaMemcpy(h + plane * p->own_z_begin, p->pub.d_counters + plane * (p->own_z_begin - p->z_begin),
                                         sizeof(int) * plane * planes, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = kn_set_error(KNC_ERR_CUDA, "multi-GPU statistics, rank %d: %s", r, cudaGetErrorString(e));
    }
    cudaSetDevice(keep);
    return rc;
}

extern "C" int knc_solve_multi(const float *xyz, int n, int k, int ngpus, unsigned int *out_rows, unsigned int *out_perm) {
    if (!out_rows) return kn_set_error(KNC_ERR_INVALID, "knc_solve_multi: out_rows is NULL");
    knc_problem *p = nullptr;
    int rc = knc_prepare_multi(xyz, n, k, ngpus, &p);
    if (rc != KNC_OK) return rc;
    rc = knc_solve(p);
    if (rc == KNC_OK) rc = knc_get_knearests(p, out_rows);
    if (rc == KNC_OK && out_perm) rc = knc_get_permutation(p, out_perm);
    knc_free(p);
    return rc;
}
