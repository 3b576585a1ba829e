// slab_route.cu -- multi-GPU plumbing on the device: route points to the z-slab owners that need them
// (owner + halo), and translate finished neighbour rows to global point ids.
//
// No reference counterpart: the reference is single-GPU.  This is the device half of the
// one exchange step of SURVEY 8e; the collective itself (NCCL all-to-all) is issued by the caller
// (cuda_knearests_b200/slabs.py) on the buffers written here.
//
// kn_route_*: STABLE multi-split.  Rank d must receive every local point whose cell plane lies in
// [zsplit[d] - halo, zsplit[d+1] + halo); a point near a slab border goes to two (rarely three)
// ranks.  Records leave grouped by destination and, inside a group, in input order -- so that a
// receiver that concatenates the groups of ranks 0..W-1 (each holding an index-contiguous chunk of
// the cloud) sees its slab in global index order, and the engine's stable cell sort then yields
// exactly the single-GPU order (and therefore bit-identical rows).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "kn_internal.h"

namespace {

constexpr int kRouteThreads = 256;
constexpr int kMaxRanks = 32;

struct RoutePlan {
    int nranks;
    int lo[kMaxRanks];   // first plane rank d wants
    int hi[kMaxRanks];   // one past the last plane rank d wants
    int own[kMaxRanks];  // first plane rank d OWNS (its queries live in [own[d], own[d + 1]))
    int own_end;         // one past the last owned plane of the last rank
};

__device__ __forceinline__ int plane_of(float z, float Df, int D) {
    // cellFromPoint, z axis (knearests.cu:25,28) -- same rounding as the grid build
    const int c = (int)floorf(__fdiv_rn(__fmul_rn(z, Df), 1000.f));
    return max(0, min(c, D - 1));
}

__device__ __forceinline__ unsigned dest_mask(int plane, const RoutePlan &pl) {
    unsigned m = 0;
    for (int d = 0; d < pl.nranks; ++d)
        if (plane >= pl.lo[d] && plane < pl.hi[d]) m |= 1u << d;
    return m;
}

// A block takes a TILE of kRouteTile consecutive points, kRouteItems per thread, all loaded before the first
// barrier (one point per thread left the kernels waiting on a single load per barrier, 5-8x off the HBM time).
// Item r of thread t is point  tile*kRouteTile + r*kRouteThreads + t,  so the warp-sized run  v = r*8 + warp
// precedes run v + 1 in input order.
constexpr int kRouteItems = 4;
constexpr int kRouteTile = kRouteThreads * kRouteItems;
constexpr int kRouteRuns = kRouteItems * (kRouteThreads / 32);

// per tile and destination: how many of the tile's points go there.  counts[d * nblocks + tile]
// own_low (optional, [2][nranks] long long, zeroed by the caller): totals of the records that lie in the destination's
// OWN planes / in its halo planes BELOW them (what the driver needs to know every rank's query range without a read-back)
__global__ void __launch_bounds__(kRouteThreads)
kn_route_count_kernel(const float *__restrict__ xyz, int n, float Df, int D, const __grid_constant__ RoutePlan pl,
                      int nblocks, int *__restrict__ counts, unsigned long long *__restrict__ own_low) {
    __shared__ int s_cnt[kMaxRanks];
    __shared__ int s_own[kMaxRanks], s_low[kMaxRanks];
    const int tid = threadIdx.x;
    if (tid < kMaxRanks) { s_cnt[tid] = 0; s_own[tid] = 0; s_low[tid] = 0; }
    const long long base = (long long)blockIdx.x * kRouteTile + tid;
    float z[kRouteItems];
#pragma unroll
    for (int r = 0; r < kRouteItems; ++r) {
        const long long i = base + (long long)r * kRouteThreads;
        z[r] = i < n ? __ldg(xyz + 3 * i + 2) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kRouteItems; ++r) {
        const bool in = base + (long long)r * kRouteThreads < n;
        const int plane = plane_of(z[r], Df, D);
        const unsigned m = in ? dest_mask(plane, pl) : 0u;
        for (int d = 0; d < pl.nranks; ++d) {
            const bool goes = (m >> d) & 1u;
            const unsigned b = __ballot_sync(0xffffffffu, goes);
            if ((tid & 31) == 0 && b) atomicAdd(&s_cnt[d], __popc(b));
            if (own_low) {  // uniform
                const int o0 = pl.own[d], o1 = d + 1 < pl.nranks ? pl.own[d + 1] : pl.own_end;
                const unsigned bo = __ballot_sync(0xffffffffu, goes && plane >= o0 && plane < o1);
                const unsigned bl = __ballot_sync(0xffffffffu, goes && plane < o0);
                if ((tid & 31) == 0) {
                    if (bo) atomicAdd(&s_own[d], __popc(bo));
                    if (bl) atomicAdd(&s_low[d], __popc(bl));
                }
            }
        }
    }
    __syncthreads();
    if (tid < pl.nranks) {
        counts[(size_t)tid * nblocks + blockIdx.x] = s_cnt[tid];
        if (own_low) {
            if (s_own[tid]) atomicAdd(own_low + tid, (unsigned long long)s_own[tid]);
            if (s_low[tid]) atomicAdd(own_low + pl.nranks + tid, (unsigned long long)s_low[tid]);
        }
    }
}

// one block per destination: exclusive scan of its row of per-block counts, total to totals[d]
__global__ void __launch_bounds__(1024) kn_route_scan_kernel(int *counts, int nblocks, long long *totals) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    int *row = counts + (size_t)blockIdx.x * nblocks;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 1024) {
        const int idx = base + tid;
        const long long v = idx < nblocks ? row[idx] : 0;
        long long incl = v;
        for (int d = 1; d < 32; d <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        long long off = s_carry;
        for (int wq = 0; wq < warp; ++wq) off += s_warp[wq];
        if (idx < nblocks) row[idx] = (int)(off + incl - v);  // fits: at most n < 2^31 records per destination
        __syncthreads();
        if (tid == 1023) s_carry = off + incl;
        __syncthreads();
    }
    if (tid == 0) totals[blockIdx.x] = s_carry;
}

// Where the group of each destination goes: a buffer (this GPU's memory, or a PEER GPU's memory mapped into this
// process -- the stores then travel over NVLink while the kernel runs) and the first record to write there.
struct RouteDest {
    float4 *buf[kMaxRanks];
    long long first[kMaxRanks];
};

// stable scatter: position = first[d] + tile offset + rank among the tile's earlier points
__global__ void __launch_bounds__(kRouteThreads)
kn_route_scatter_kernel(const float *__restrict__ xyz, int n, unsigned first_id, float Df, int D,
                        const __grid_constant__ RoutePlan pl, int nblocks, const int *__restrict__ offsets,
                        const __grid_constant__ RouteDest dst) {
    __shared__ int s_run[kMaxRanks][kRouteRuns];      // per destination: count of each run, then its exclusive prefix
    __shared__ int s_part[kMaxRanks][kRouteRuns / 4];
    static_assert(kMaxRanks * (kRouteRuns / 4) == kRouteThreads, "one thread per (destination, 4 runs)");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long base = (long long)blockIdx.x * kRouteTile + tid;
    float x[kRouteItems], y[kRouteItems], z[kRouteItems];
    unsigned m[kRouteItems];
#pragma unroll
    for (int r = 0; r < kRouteItems; ++r) {
        const long long i = base + (long long)r * kRouteThreads;
        const long long j = i < n ? i : 0;
        x[r] = __ldg(xyz + 3 * j); y[r] = __ldg(xyz + 3 * j + 1); z[r] = __ldg(xyz + 3 * j + 2);
        m[r] = i < n ? dest_mask(plane_of(z[r], Df, D), pl) : 0u;
    }
#pragma unroll
    for (int r = 0; r < kRouteItems; ++r)
        for (int d = 0; d < pl.nranks; ++d) {
            const unsigned b = __ballot_sync(0xffffffffu, (m[r] >> d) & 1u);
            if (lane == 0) s_run[d][r * (kRouteThreads / 32) + warp] = __popc(b);
        }
    __syncthreads();
    {   // exclusive prefix over the runs of each destination: thread (d, 4 runs) scans its part, then adds the parts below
        const int d = tid / (kRouteRuns / 4), part = tid % (kRouteRuns / 4);
        int run = 0;
        if (d < pl.nranks) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int c = s_run[d][part * 4 + j];
                s_run[d][part * 4 + j] = run;
                run += c;
            }
        }
        s_part[d][part] = run;
        __syncthreads();
        if (d < pl.nranks) {
            int below = 0;
            for (int q = 0; q < part; ++q) below += s_part[d][q];
#pragma unroll
            for (int j = 0; j < 4; ++j) s_run[d][part * 4 + j] += below;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kRouteItems; ++r)
        for (int d = 0; d < pl.nranks; ++d) {
            const unsigned b = __ballot_sync(0xffffffffu, (m[r] >> d) & 1u);
            if ((m[r] >> d) & 1u) {
                const int before = __popc(b & ((1u << lane) - 1u)) + s_run[d][r * (kRouteThreads / 32) + warp];
                const long long pos = dst.first[d] + offsets[(size_t)d * nblocks + blockIdx.x] + before;
                dst.buf[d][pos] = make_float4(x[r], y[r], z[r],  // one 16-byte store
                                              __uint_as_float(first_id + (unsigned)(base + (long long)r * kRouteThreads)));
            }
        }
}

// rows[r*k + j] = ids[perm[knn[r*k + j]]], query_ids[r] = ids[perm[query_begin + r]]
// (the stored rows hold local sorted position + index_offset, knc_set_index_offset)
__global__ void kn_rows_to_global_kernel(const unsigned *__restrict__ knn, const unsigned *__restrict__ perm,
                                         const unsigned *__restrict__ ids, int id_stride, int query_begin,
                                         unsigned index_offset, long long total, int k,
                                         unsigned *__restrict__ out_rows, unsigned *__restrict__ out_query_ids) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const unsigned v = __ldg(knn + e);
    out_rows[e] = v == KNC_EMPTY ? KNC_EMPTY : __ldg(ids + (size_t)id_stride * __ldg(perm + (v - index_offset)));
    const long long r = e / k;
    if (out_query_ids && e - r * k == 0)
        out_query_ids[r] = __ldg(ids + (size_t)id_stride * __ldg(perm + query_begin + r));
}

// out[i] = ids[perm[query_begin + i]]: this rank's slice of the global permutation
__global__ void kn_perm_to_global_kernel(const unsigned *__restrict__ perm, const unsigned *__restrict__ ids,
                                         int id_stride, int query_begin, int count, unsigned *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = __ldg(ids + (size_t)id_stride * __ldg(perm + query_begin + i));
}

}  // namespace

extern "C" int knc_get_permutation_global(knc_problem *p, const unsigned int *d_ids, int id_stride,
                                          unsigned int *d_out) {
    if (!p || !d_ids || !d_out || id_stride < 1)
        return kn_set_error(KNC_ERR_INVALID, "knc_get_permutation_global: bad argument");
    if (p->query_count > 0) {
        kn_perm_to_global_kernel<<<(unsigned)((p->query_count + 255) / 256), 256, 0, p->stream>>>(
            p->pub.d_permutation, d_ids, id_stride, p->query_begin, p->query_count, d_out);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
        if (e != cudaSuccess) return kn_set_error(KNC_ERR_CUDA, "knc_get_permutation_global: %s", cudaGetErrorString(e));
    }
    return KNC_OK;
}

// A routing pass in two steps, so that the caller can learn the group sizes (and agree with the other ranks
// on where each group lands) before any record is written.
struct knc_route {
    const float *d_xyz;
    int n, grid_dim, nblocks;
    unsigned first_id;
    RoutePlan pl;
    int *d_counts;       // [nranks][nblocks] per-block group sizes, then their exclusive scan
    long long *d_tot;    // [nranks] group totals
    long long counts[kMaxRanks];
    int counts_known;    // knc_route_begin read the group sizes back; the asynchronous variant leaves them on the device
    cudaStream_t stream; // where the last scatter ran (the scratch is freed in its order)
};

static int route_begin_impl(const float *d_xyz, int n, unsigned int first_id, int grid_dim, int nranks, const int *zsplit,
                            int halo, cudaStream_t st, long long *d_counts3, knc_route **out, long long *out_counts,
                            const char *who) {
    if (!d_xyz || !zsplit || !out || n < 0 || grid_dim < 1 || halo < 0 || (!out_counts && !d_counts3))
        return kn_set_error(KNC_ERR_INVALID, "%s: bad argument", who);
    if (nranks < 1 || nranks > kMaxRanks)
        return kn_set_error(KNC_ERR_INVALID, "%s: nranks = %d, supported 1..%d", who, nranks, kMaxRanks);
    *out = nullptr;
    knc_route *r = (knc_route *)calloc(1, sizeof(knc_route));
    if (!r) return kn_set_error(KNC_ERR_NOMEM, "%s: out of host memory", who);
    r->d_xyz = d_xyz; r->n = n; r->grid_dim = grid_dim; r->first_id = first_id;
    r->pl.nranks = nranks;
    r->pl.own_end = zsplit[nranks];
    for (int d = 0; d < nranks; ++d) {
        if (zsplit[d] > zsplit[d + 1] || zsplit[0] != 0 || zsplit[nranks] != grid_dim) {
            free(r);
            return kn_set_error(KNC_ERR_INVALID, "%s: zsplit must rise from 0 to grid_dim", who);
        }
        r->pl.lo[d] = zsplit[d] - halo;
        r->pl.hi[d] = zsplit[d + 1] + halo;
        r->pl.own[d] = zsplit[d];
        if (zsplit[d] == zsplit[d + 1]) r->pl.lo[d] = r->pl.hi[d] = 0;  // an empty slab wants nothing
        if (out_counts) out_counts[d] = 0;
    }
    r->nblocks = (n + kRouteTile - 1) / kRouteTile;
    cudaError_t e = cudaSuccess;
    if (d_counts3) e = cudaMemsetAsync(d_counts3, 0, sizeof(long long) * 3 * nranks, st);
    if (n > 0 && e == cudaSuccess) {
        e = kn_dev_alloc((void **)&r->d_counts, sizeof(int) * (size_t)nranks * r->nblocks, st);
        if (e == cudaSuccess && !d_counts3) e = kn_dev_alloc((void **)&r->d_tot, sizeof(long long) * kMaxRanks, st);
        if (e == cudaSuccess) {
            long long *tot = d_counts3 ? d_counts3 : r->d_tot;  // [0, nranks): records per destination
            kn_route_count_kernel<<<r->nblocks, kRouteThreads, 0, st>>>(
                d_xyz, n, (float)grid_dim, grid_dim, r->pl, r->nblocks, r->d_counts,
                d_counts3 ? reinterpret_cast<unsigned long long *>(d_counts3) + nranks : nullptr);
            kn_route_scan_kernel<<<nranks, 1024, 0, st>>>(r->d_counts, r->nblocks, tot);
            e = cudaGetLastError();
            if (e == cudaSuccess && out_counts)
                e = cudaMemcpyAsync(r->counts, tot, sizeof(long long) * nranks, cudaMemcpyDeviceToHost, st);
        }
        if (e == cudaSuccess && out_counts) e = cudaStreamSynchronize(st);
    }
    if (e != cudaSuccess) {
        if (r->d_counts) cudaFreeAsync(r->d_counts, st);
        if (r->d_tot) cudaFreeAsync(r->d_tot, st);
        free(r);
        return kn_set_error(KNC_ERR_CUDA, "%s: %s", who, cudaGetErrorString(e));
    }
    if (out_counts) for (int d = 0; d < nranks; ++d) out_counts[d] = r->counts[d];
    r->counts_known = out_counts != nullptr;
    *out = r;
    return KNC_OK;
}

extern "C" int knc_route_begin(const float *d_xyz, int n, unsigned int first_id, int grid_dim, int nranks,
                               const int *zsplit, int halo, knc_route **out, long long *out_counts) {
    if (!out_counts) return kn_set_error(KNC_ERR_INVALID, "knc_route_begin: bad argument");
    return route_begin_impl(d_xyz, n, first_id, grid_dim, nranks, zsplit, halo, 0, nullptr, out, out_counts, "knc_route_begin");
}

extern "C" int knc_route_begin_async(const float *d_xyz, int n, unsigned int first_id, int grid_dim, int nranks,
                                     const int *zsplit, int halo, void *stream, long long *d_counts3, knc_route **out) {
    if (!d_counts3) return kn_set_error(KNC_ERR_INVALID, "knc_route_begin_async: bad argument");
    return route_begin_impl(d_xyz, n, first_id, grid_dim, nranks, zsplit, halo, (cudaStream_t)stream, d_counts3, out, nullptr,
                            "knc_route_begin_async");
}

static int route_scatter_impl(knc_route *r, void *const *dest_bufs, const long long *dest_first, cudaStream_t st, bool wait,
                              const char *who) {
    if (!r || !dest_bufs || !dest_first) return kn_set_error(KNC_ERR_INVALID, "%s: NULL argument", who);
    if (r->n == 0) return KNC_OK;
    RouteDest dst;
    for (int d = 0; d < r->pl.nranks; ++d) {
        const bool used = r->counts_known ? r->counts[d] > 0 : true;  // without host counts every destination must be usable
        if (used && (!dest_bufs[d] || ((uintptr_t)dest_bufs[d] & 15u) || dest_first[d] < 0))
            return kn_set_error(KNC_ERR_INVALID, "%s: destination %d has no usable buffer", who, d);
        dst.buf[d] = reinterpret_cast<float4 *>(dest_bufs[d]);
        dst.first[d] = dest_first[d];
    }
    kn_route_scatter_kernel<<<r->nblocks, kRouteThreads, 0, st>>>(r->d_xyz, r->n, r->first_id, (float)r->grid_dim, r->grid_dim,
                                                                 r->pl, r->nblocks, r->d_counts, dst);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && wait) e = cudaStreamSynchronize(st);  // the records (also those sent to peers) have landed
    if (e != cudaSuccess) return kn_set_error(KNC_ERR_CUDA, "%s: %s", who, cudaGetErrorString(e));
    r->stream = st;
    return KNC_OK;
}

extern "C" int knc_route_scatter(knc_route *r, void *const *dest_bufs, const long long *dest_first) {
    return route_scatter_impl(r, dest_bufs, dest_first, 0, true, "knc_route_scatter");
}

extern "C" int knc_route_scatter_async(knc_route *r, void *const *dest_bufs, const long long *dest_first, void *stream) {
    return route_scatter_impl(r, dest_bufs, dest_first, (cudaStream_t)stream, false, "knc_route_scatter_async");
}

extern "C" int knc_route_end(knc_route *r) {
    if (!r) return KNC_OK;
    if (r->d_counts) cudaFreeAsync(r->d_counts, r->stream);  // stream-ordered: after the scatter that reads it
    if (r->d_tot) cudaFreeAsync(r->d_tot, r->stream);
    free(r);
    return KNC_OK;
}

extern "C" int knc_slab_route(const float *d_xyz, int n, unsigned int first_id, int grid_dim, int nranks,
                              const int *zsplit, int halo, float *d_out_rec, long long capacity,
                              long long *out_counts) {
    if (!d_out_rec || ((uintptr_t)d_out_rec & 15u)) return kn_set_error(KNC_ERR_INVALID, "knc_slab_route: bad output buffer");
    knc_route *r = nullptr;
    int rc = knc_route_begin(d_xyz, n, first_id, grid_dim, nranks, zsplit, halo, &r, out_counts);
    if (rc != KNC_OK) return rc;
    void *bufs[kMaxRanks];
    long long first[kMaxRanks], run = 0;
    for (int d = 0; d < nranks; ++d) { bufs[d] = d_out_rec; first[d] = run; run += out_counts[d]; }
    if (run > capacity)
        rc = kn_set_error(KNC_ERR_NOMEM, "knc_slab_route: %lld records to send, output capacity %lld", run, capacity);
    else
        rc = knc_route_scatter(r, bufs, first);
    knc_route_end(r);
    return rc;
}

// ---- peer buffers: device memory that other ranks (processes) of the node can write into -------------------------
extern "C" int knc_peer_alloc(size_t bytes, void **d_ptr, unsigned char *handle64) {
    if (!d_ptr || !handle64 || bytes == 0) return kn_set_error(KNC_ERR_INVALID, "knc_peer_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaError_t e = cudaMalloc(d_ptr, bytes);  // plain cudaMalloc: pool memory cannot be exported
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) {
        e = cudaIpcGetMemHandle(&h, *d_ptr);
        if (e != cudaSuccess) { cudaFree(*d_ptr); *d_ptr = nullptr; }  // not exportable: do not leak the buffer
    }
    if (e != cudaSuccess) return kn_set_error(e == cudaErrorMemoryAllocation ? KNC_ERR_NOMEM : KNC_ERR_CUDA, "knc_peer_alloc: %s",
                                              cudaGetErrorString(e));
    memcpy(handle64, &h, 64);
    return KNC_OK;
}

extern "C" int knc_peer_open(const unsigned char *handle64, void **d_ptr) {
    if (!handle64 || !d_ptr) return kn_set_error(KNC_ERR_INVALID, "knc_peer_open: NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    cudaError_t e = cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return kn_set_error(KNC_ERR_CUDA, "knc_peer_open: %s", cudaGetErrorString(e));
    return KNC_OK;
}

extern "C" int knc_peer_close(void *d_ptr) {
    if (d_ptr && cudaIpcCloseMemHandle(d_ptr) != cudaSuccess) return kn_set_error(KNC_ERR_CUDA, "knc_peer_close failed");
    return KNC_OK;
}

extern "C" int knc_peer_free(void *d_ptr) {
    if (d_ptr && cudaFree(d_ptr) != cudaSuccess) return kn_set_error(KNC_ERR_CUDA, "knc_peer_free failed");
    return KNC_OK;
}

extern "C" int knc_get_knearests_global(knc_problem *p, const unsigned int *d_ids, int id_stride,
                                        unsigned int *d_out_rows, unsigned int *d_out_query_ids) {
    if (!p || !d_ids || !d_out_rows || id_stride < 1)
        return kn_set_error(KNC_ERR_INVALID, "knc_get_knearests_global: bad argument");
    if (!p->solved) return kn_set_error(KNC_ERR_STATE, "knc_get_knearests_global: call knc_solve first");
    const long long total = (long long)p->query_count * p->k;
    if (total > 0) {
        kn_rows_to_global_kernel<<<(unsigned)((total + 255) / 256), 256, 0, p->stream>>>(
            p->pub.d_knearests, p->pub.d_permutation, d_ids, id_stride, p->query_begin, p->index_offset, total, p->k,
            d_out_rows, d_out_query_ids);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
        if (e != cudaSuccess) return kn_set_error(KNC_ERR_CUDA, "knc_get_knearests_global: %s", cudaGetErrorString(e));
    }
    return KNC_OK;
}
