// knn_query.cu -- the all-points exact kNN query for sm_100a.  Replaces the reference's `knearest`
// kernel with its heap helpers (knearests.cu:62-148) and the ring offset table it walks
// (knearests.cu:254-300).
//
// Same contract, different algorithm.  The reference walks a precomputed list of 29 791 cell offsets
// ring by ring (Chebyshev rings of whole cells, ~125 cells / ~378 candidates per query at k=50) and
// keeps a binary max-heap per thread in shared memory.  Here:
//
//  * Candidates come from X-ROWS of cells.  Cell ids are x-fastest (knearests.cu:29) and the build
//    lays cells out in ascending id, so the cells [ilo..ihi] of one (j,k) row are ONE contiguous span
//    ptrs[row+ilo] .. ptrs[row+ihi+1] of 16-byte point records: two pointer loads per row instead of
//    two per cell, and a straight streaming loop over the span.
//  * The search region is a SPHERE, not a cube of rings.  Each query starts with a radius guess
//    r0 (from the cloud's density, so that ~k + 2.5 sqrt(k) points are expected inside) and the
//    bound shrinks to the current k-th distance as soon as k candidates are held.  Rows are visited
//    nearest-first (table sorted by the distance from the own cell); a row is skipped when its yz
//    distance exceeds the bound, and inside a row only the cells within sqrt(bound^2 - dyz^2) in x
//    are touched.  ~90-110 candidates per query at k=20 and 3.1 points/cell instead of ~364.
//  * Exact for ANY input: if the guess was too small (fewer than k points inside, or the k-th found
//    lies outside the sphere that was fully searched) the query restarts with radius = that k-th
//    distance (guaranteed sufficient) or 1.5x the radius; it stops growing only when the whole grid
//    has been searched.  The reference gives up silently after ring 15 (knearests.cu:137-139).
//  * All pruning tests are conservative against f32 rounding (MARGIN / SHRINK below): a candidate is
//    discarded only when its squared distance is provably greater than the current k-th.
//  * Candidates are ordered by the 64-bit key bits(d2) << 32 | sorted position: one unsigned compare gives the
//    total order (d2, position), so the result does not depend on visiting order -- deterministic, and
//    identical across the three kernels below, across grid densities and across GPUs.
//  * Five kernels (details at each):
//      kn_query_tile_kernel     k = 8 / 20 / 32 / 50 on grids whose y/z cells are at least as wide as the first radius
//                               (the default grid, knc_fast_dims).  A block of 128 consecutive queries stages the
//                               3 x 3 x-rows of cells around them ONCE in shared memory with bulk asynchronous copies
//                               (cp.async.bulk + mbarrier, one copy per row segment), every lane then streams its own
//                               windows of that tile with 128-bit shared-memory loads, remembers the candidates inside
//                               its sphere as 16-bit tile slots, selects like the select kernel below and the block's
//                               rows leave as one bulk copy shared -> global.
//      kn_query_select_kernel   k = 8 / 20 / 32 / 50, any grid: the previous generation, still the path for blocks
//                               whose tile does not fit (clustered clouds) and for cubic grids, and -- over the retry
//                               list, with a larger sphere and buffer -- the second chance of the queries the first pass
//                               could not finish.  Warp-uniform control flow: span lists, lock-step
//                               streaming of the records through two register buffers, candidates inside the
//                               sphere are only remembered; one sort-select at the end on 32-bit truncated keys
//                               (Batcher network in registers) with an exactness guard.  Sizes its sphere by the local
//                               density where the query's own row is far off the cloud's mean.
//      kn_query_fast_kernel     other k: same streaming, candidates are merged into a sorted list of 64-bit keys
//                               in shared memory in warp-wide bursts.
//      kn_query_general_kernel  per-lane nested loops, restarts, rings beyond the row table: exact for any input;
//                               finishes the few queries the others hand over (retry list), densely packed,
//                               serves clouds whose first radius exceeds the table, and external query points.
//      kn_query_retry_warp_kernel  the same search for a SHORT retry list, one warp per query (small clouds).
//    (v1 walked rows and candidates in per-lane nested loops for every query: 9 of 32 lanes active on average,
//    profiles/r1_query_v1_*; the select kernel runs at 31.6.)
//  * Rows are written through shared memory as one contiguous, coalesced run per block
//    (the reference writes stride-k, knearests.cu:144-147).
//
// Distance arithmetic is the reference's, bit for bit: dx = c.x - p.x, ..,
// d2 = fma(dz,dz, fma(dx,dx, dy*dy)) (knearests.cu:125 under nvcc's default contraction).
//
// Roofline: HBM.  Algorithmic bytes per query = 12 + 4k (sorted point read once, k indices written).
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include "kn_internal.h"
#include "sort_network.h"

#ifndef KN_QSLOTS
#define KN_QSLOTS 8
#endif

namespace {

constexpr int kTableRing = 5;                                             // table covers |dj|,|dk| <= 5
constexpr int kTableSize = (2 * kTableRing + 1) * (2 * kTableRing + 1);  // 121 rows
static_assert(kTableSize == KN_ROW_TABLE_SIZE, "kn_problem.num_cell_offsets reports the table size");
constexpr float kMargin = 4e-3f;               // absolute slack on a cell-border distance, coordinate units
constexpr float kShrink = 1.0f - 1.0f / 262144.0f;  // relative slack on a squared lower bound (1 - 2^-18)
constexpr float kGrow = 1.0f + 1.0f / 262144.0f;

struct RowTable {
    int n;
    signed char dj[kTableSize];
    signed char dk[kTableSize];
    float lb2[kTableSize];  // conservative squared distance from anywhere in the own cell to that row
};

struct QueryParams {
    const float4 *pts;
    const int *ptrs;
    unsigned int *knn;
    unsigned int *diag;
    GridGeom g;
    int n, k, query_begin, query_count;
    float radius0_sq;
    int use_select;        // small k: remember-then-select kernel instead of the insertion kernel
    unsigned index_offset; // added to every stored neighbour index (slab mode: local -> global sorted position)
    const float *ext_queries;  // EXTERNAL mode of the general kernel: query points that are not in the cloud (xyz)
    int *retry_list;       // sorted positions of the queries the fast kernel could not finish
    unsigned int *retry_count;
    int *block_list;       // blocks (of QB queries) the tile kernel hands to the select kernel; NULL: every block
    unsigned int *block_count;
    int halo_cells;        // tile kernel: cells staged beyond the block's own x-range on either side
    int bulk_rows_ok;      // tile kernel: the row array is 16-byte aligned, rows may leave as one bulk copy
    float retry_grow;      // retry kernels: factor on the squared radius when a sphere held fewer than k points
    int *retry_list2;      // queries the second-chance pass of the select kernel could not finish either
    unsigned int *retry_count2;
    float retry_radius_sq; // squared radius of the second-chance pass (0: there is none) ...
    float retry_radius_max_sq;  // ... and the most it may be scaled to for a query at the domain's border (the table's reach)
    float points_per_cell; // the cloud's mean: what the select kernel measures the local density against
    unsigned retry_short;  // a retry list of at most this many queries goes to the warp-per-query kernel
    RowTable tab;
};

__device__ __forceinline__ float sqrt_approx(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Per-query geometry shared by both kernels.
struct QueryGeom {
    float px, py, pz;   // the query point
    float tx, fy, fz;   // x in cell units; position inside the own cell in y and z (cell units)
    float tz;
    int cj, ck;         // own cell row (global plane index for ck)
};

__device__ __forceinline__ QueryGeom query_geom(const float4 me, const GridGeom &g) {
    QueryGeom s;
    s.px = me.x; s.py = me.y; s.pz = me.z;
    // same rounding as the build (knearests.cu:23-28): mul.f32, div.rn.f32, floor
    s.tx = __fdiv_rn(__fmul_rn(me.x, g.Dfx), 1000.f);
    const float ty = __fdiv_rn(__fmul_rn(me.y, g.Dfy), 1000.f);
    s.tz = __fdiv_rn(__fmul_rn(me.z, g.Dfz), 1000.f);
    s.cj = max(0, min((int)floorf(ty), g.Dy - 1));
    s.ck = max(g.zb, min((int)floorf(s.tz), g.ze - 1));
    s.fy = ty - (float)s.cj;
    s.fz = s.tz - (float)s.ck;
    return s;
}

// Conservative squared distance from the query to the x-row (j, k) = (cj+dj, ck+dk) of cells.
// A border cell also holds the points that lie OUTSIDE the domain on its side (cellFromPoint clamps,
// knearests.cu:26-28), so towards that side the cell is unbounded: no gap may be claimed there.  (It
// matters when the query itself is outside the domain: its own border row then holds points that are
// closer to it than the nominal cell box is.)
__device__ __forceinline__ float row_lower_bound2(const QueryGeom &s, int dj, int dk, const GridGeom &g) {
    const int j = s.cj + dj, k = s.ck + dk;
    const float below_y = j > 0 ? (float)dj - s.fy : 0.f;            // query below the row (row's low side bounded)
    const float above_y = j < g.Dy - 1 ? s.fy - (float)dj - 1.f : 0.f;  // query above the row
    const float below_z = k > g.zb ? (float)dk - s.fz : 0.f;
    const float above_z = k < g.ze - 1 ? s.fz - (float)dk - 1.f : 0.f;
    const float gy = fmaxf(0.f, fmaxf(below_y, above_y));
    const float gz = fmaxf(0.f, fmaxf(below_z, above_z));
    const float ly = fmaxf(0.f, fmaf(gy, g.wy, -kMargin)), lz = fmaxf(0.f, fmaf(gz, g.wz, -kMargin));
    return (ly * ly + lz * lz) * kShrink;
}

// Cells [ilo, ihi] of a row that can hold a point within sqrt(bound2 - lbyz2) of the query in x.
__device__ __forceinline__ void row_x_range(const QueryGeom &s, float bound2, float lbyz2, float invw, int D,
                                            int &ilo, int &ihi) {
    const float rxc = fmaf(sqrt_approx(bound2 - lbyz2) * kGrow, invw, kMargin * invw + 1e-3f);
    // both ends clamped like cellFromPoint clamps: a query outside the domain still reaches its border cell
    ilo = max(0, min(D - 1, (int)floorf(s.tx - rxc)));
    ihi = max(0, min(D - 1, (int)floorf(s.tx + rxc)));
}

__device__ __forceinline__ unsigned long long make_key(const float4 c, const QueryGeom &s, int p) {
    const float dx = c.x - s.px, dy = c.y - s.py, dz = c.z - s.pz;
    const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));  // knearests.cu:125 as compiled
    return ((unsigned long long)__float_as_uint(d2) << 32) | (unsigned)p;
}

// Sorted insertion into the slot-major list L (stride QB); the worst entry drops off the end.
template <int QB>
__device__ __forceinline__ void list_insert(unsigned long long *L, int K, int &cnt, unsigned long long key) {
    int s;
    if (cnt < K) s = cnt++; else s = K - 1;
    while (s > 0) {
        const unsigned long long e = L[(s - 1) * QB];
        if (e < key) break;
        L[s * QB] = e;
        --s;
    }
    L[s * QB] = key;
}

// Restart counter, largest final radius, slab halo misses (planes that exist globally but are not held).
__device__ __forceinline__ void report_diag(const QueryParams &P, const QueryGeom &s, bool finished, float kd2,
                                            unsigned restarts, int tid) {
    unsigned rad_bits = 0, halo_miss = 0;
    if (finished) {
        rad_bits = __float_as_uint(kd2);
        if (P.g.zb > 0) {
            const float gap = fmaxf(0.f, (s.tz - (float)P.g.zb) * P.g.wz - kMargin);
            if (gap * gap * kShrink <= kd2) halo_miss = 1;
        }
        if (P.g.ze < P.g.Dz) {
            const float gap = fmaxf(0.f, ((float)P.g.ze - s.tz) * P.g.wz - kMargin);
            if (gap * gap * kShrink <= kd2) halo_miss = 1;
        }
    }
    const unsigned wr = __reduce_add_sync(0xffffffffu, restarts);
    const unsigned wm = __reduce_max_sync(0xffffffffu, rad_bits);
    const unsigned wh = __reduce_add_sync(0xffffffffu, halo_miss);
    if ((tid & 31) == 0) {
        if (wr) atomicAdd(P.diag + 0, wr);
        atomicMax(P.diag + 1, wm);
        if (wh) atomicAdd(P.diag + 2, wh);
    }
}

// ======================================================================================== fast path
#ifndef KN_FIRST_BATCH
#define KN_FIRST_BATCH 9
#define KN_NEXT_BATCH 8
#endif
constexpr int kFirstBatch = KN_FIRST_BATCH;  // rows of the first batch: the own row and its 8 neighbours (lower bound 0)
constexpr int kNextBatch = KN_NEXT_BATCH;    // rows per later batch
constexpr int kSpanSlots = KN_FIRST_BATCH > KN_NEXT_BATCH ? KN_FIRST_BATCH : KN_NEXT_BATCH;  // spans held per batch
constexpr int kQueueSlots = KN_QSLOTS;   // pending candidates per lane between insertion bursts
constexpr int kChunk = 4;        // candidates a lane takes per lock-step iteration

__device__ __forceinline__ void cswap(unsigned long long &a, unsigned long long &b) {
    const unsigned long long lo = a < b ? a : b, hi = a < b ? b : a;
    a = lo; b = hi;
}

template <int K_T, int QB>
__global__ void __launch_bounds__(QB) kn_query_fast_kernel(const __grid_constant__ QueryParams P) {
    extern __shared__ unsigned long long smem[];
    const int K = K_T ? K_T : P.k;
    const int tid = threadIdx.x;
    unsigned long long *L = smem + tid;                                        // [K][QB] sorted k best
    int2 *SP = reinterpret_cast<int2 *>(smem + (size_t)K * QB) + tid;          // [kSpanSlots][QB] spans
    unsigned long long *QU = smem + (size_t)(K + kSpanSlots) * QB + tid;       // [kQueueSlots][QB] pending
    const int Dx = P.g.Dx, Dy = P.g.Dy, zb = P.g.zb, ze = P.g.ze;
    const float invw = P.g.invwx;
    const long long row0 = (long long)blockIdx.x * QB;
    const bool valid = row0 + tid < P.query_count;
    const int q = P.query_begin + (int)min(row0 + tid, (long long)P.query_count - 1);
    const QueryGeom s = query_geom(__ldg(P.pts + q), P.g);

    const float r2 = P.radius0_sq;
    // Only candidates inside the first-guess sphere are ever held: a full list is then final.
    unsigned long long kth = ((unsigned long long)__float_as_uint(r2) << 32) | 0xFFFFFFFFu;
    float bound2 = r2;
    int cnt = 0, qlen = 0;

    // Insertion burst for the whole warp.  Small bursts insert one by one; larger ones sort the
    // pending keys (19-comparator network in registers) and merge them into the list in ONE backward
    // pass, whose length is about the same in every lane.
    auto drain = [&]() {
        const int m = __reduce_max_sync(0xffffffffu, qlen);
        if (m == 0) return;
        if (m <= 2) {
            for (int i = 0; i < m; ++i) {
                if (i < qlen) {
                    const unsigned long long key = QU[i * QB];
                    if (key < kth) {  // the bound may have tightened since the candidate was queued
                        list_insert<QB>(L, K, cnt, key);
                        if (cnt == K) kth = L[(K - 1) * QB];
                    }
                }
            }
        } else {
            unsigned long long b[kQueueSlots];
            int nb = 0;
#pragma unroll
            for (int u = 0; u < kQueueSlots; ++u) {
                b[u] = ~0ull;
                if (u < qlen) {
                    const unsigned long long key = QU[u * QB];
                    if (key < kth) { b[u] = key; ++nb; }
                }
            }
#define KN_CE(i, j) cswap(b[i], b[j]);
#if KN_QSLOTS == 16
            KN_SORT_NETWORK_16(KN_CE)
#else
            KN_SORT_NETWORK_8(KN_CE)
#endif
#undef KN_CE
#pragma unroll
            for (int u = 0; u < kQueueSlots; ++u) QU[u * QB] = b[u];
            // Merge from the back; entries that fall beyond slot K-1 are dropped.  Branch-free body: the
            // side that was consumed is reloaded through ONE predicated load from a selected address,
            // so the lanes of a burst never split (0 = "exhausted" sentinel, smaller than any live key
            // that could still have to move).
            int i = cnt - 1, j = nb - 1, o = cnt + nb - 1;
            unsigned long long e = i >= 0 ? L[i * QB] : 0ull;
            unsigned long long bj = j >= 0 ? QU[j * QB] : 0ull;
            while (j >= 0) {
                const bool from_list = e > bj;
                if (o < K) L[o * QB] = from_list ? e : bj;
                i -= from_list ? 1 : 0;
                j -= from_list ? 0 : 1;
                --o;
                const int idx = from_list ? i : j;
                const unsigned long long *src = from_list ? L + i * QB : QU + j * QB;
                unsigned long long val = 0ull;
                if (idx >= 0) val = *src;
                e = from_list ? val : e;
                bj = from_list ? bj : val;
            }
            cnt = min(K, cnt + nb);
            if (cnt == K) kth = L[(K - 1) * QB];
        }
        if (cnt == K) bound2 = __uint_as_float((unsigned)(kth >> 32));
        qlen = 0;
    };

    int e0 = 0, e1 = min(kFirstBatch, P.tab.n);
    while (true) {
        const bool want = valid && P.tab.lb2[e0] <= bound2;  // table is sorted by lower bound
        if (!__any_sync(0xffffffffu, want)) break;

        // ---- 1. rows e0..e1 -> spans of candidate positions (uniform loop, 4 rows' pointers in flight)
        int nsp = 0;
        for (int e = e0; e < e1; e += 4) {
            int alo[4], ahi[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                ok[u] = false;
                const int ee = e + u;
                if (ee < e1 && want && P.tab.lb2[ee] <= bound2) {
                    const int dj = P.tab.dj[ee], dk = P.tab.dk[ee];
                    const int j = s.cj + dj, k = s.ck + dk;
                    if (j >= 0 && j < Dy && k >= zb && k < ze) {
                        const float lbyz2 = row_lower_bound2(s, dj, dk, P.g);
                        if (lbyz2 <= bound2) {
                            int ilo, ihi;
                            row_x_range(s, bound2, lbyz2, invw, Dx, ilo, ihi);
                            const int row = ((k - zb) * Dy + j) * Dx;
                            alo[u] = row + ilo;
                            ahi[u] = row + ihi + 1;
                            ok[u] = ilo <= ihi;
                        }
                    }
                }
            }
            int b4[4], e4[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) { b4[u] = __ldg(P.ptrs + alo[u]); e4[u] = __ldg(P.ptrs + ahi[u]); }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u] && b4[u] < e4[u]) { SP[nsp * QB] = make_int2(b4[u], e4[u]); ++nsp; }
        }

        // ---- 2. every lane streams its own spans, kChunk candidates per iteration, in lock step
        int sp = 0, p = 0, end = 0;
        if (nsp > 0) { const int2 t = SP[0]; p = t.x; end = t.y; sp = 1; }
        while (__any_sync(0xffffffffu, p < end)) {
            float4 c[kChunk];
#pragma unroll
            for (int u = 0; u < kChunk; ++u) {
                c[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (p + u < end) c[u] = __ldg(P.pts + p + u);
            }
            // branch-free body: the arithmetic runs for all four slots, only the append is predicated
#pragma unroll
            for (int u = 0; u < kChunk; ++u) {
                const unsigned long long key = make_key(c[u], s, p + u);
                const bool pass = (p + u < end) & (key < kth) & (p + u != q);  // self excluded, knearests.cu:123
                if (pass) QU[qlen * QB] = key;
                qlen += pass ? 1 : 0;
            }
            // the unused .w lanes stay allocated until here: otherwise they are recycled at once and the
            // next instruction that writes them waits for the loads (seen in profiles/r1_query_v2*)
            asm volatile("" ::"f"(c[0].w), "f"(c[1].w), "f"(c[2].w), "f"(c[3].w));
            p += kChunk;
            if (p >= end && sp < nsp) { const int2 t = SP[sp * QB]; p = t.x; end = t.y; ++sp; }
            if (__any_sync(0xffffffffu, qlen > kQueueSlots - kChunk)) drain();
        }
        drain();

        e0 = e1;
        if (e0 >= P.tab.n) break;  // the launcher only takes this path when the table covers the sphere
        e1 = min(e0 + kNextBatch, P.tab.n);
    }

    // fewer than k points inside the first-guess sphere: hand the query to the general kernel
    const bool finished = valid && cnt == K;
    if (valid && !finished) P.retry_list[atomicAdd(P.retry_count, 1u)] = q | (int)0x80000000;  // needs a larger sphere
    report_diag(P, s, finished, finished ? __uint_as_float((unsigned)(kth >> 32)) : 0.f, 0u, tid);

    // ---- coalesced row output --------------------------------------------------------------------
    for (int t = cnt; t < K; ++t) L[t * QB] = ~0ull;  // low word = UINT_MAX = "no such neighbour"
    __syncthreads();
    const int rows = (int)min((long long)QB, (long long)P.query_count - row0);
    unsigned int *out = P.knn + row0 * K;
    for (int idx = tid; idx < rows * K; idx += QB) {
        const int t = idx / K, sl = idx - t * K;
        const unsigned v = (unsigned)smem[sl * QB + t];
        out[idx] = v == KNC_EMPTY ? v : v + P.index_offset;
    }
}

// ===================================================================================== select path
// For small compile-time k (8, 20).  Nothing is kept sorted while scanning: every candidate inside the
// first-guess sphere is only REMEMBERED (its 32-bit position, appended to a per-lane buffer in shared
// memory; no k-th distance, no insertion bursts, fully warp-uniform row loop because the bound never
// changes).  When the sphere has been searched, the scan state is dead and each lane rebuilds the keys of
// its buffered candidates (the points are still in L1/L2), sorts 32 of them at a time with a fixed
// Batcher network in registers -- branch-free, all lanes busy -- and keeps the k smallest.
// k <= count <= capacity  =>  the row is exact (the whole sphere was searched and holds >= k points);
// anything else goes to the retry list of the general kernel.
// NET keys are sorted per round, CAP candidates can be buffered per lane (32-bit positions):
//   k <= 20 : NET 32, CAP 56      k = 32 : NET 64, CAP 92      k = 50 : NET 64, CAP 100   (the last kSelChunk slots are overrun room,
//   and the query itself takes one)
#ifndef KN_SEL_CHUNK
#define KN_SEL_CHUNK 4
#endif
constexpr int kSelChunk = KN_SEL_CHUNK;  // candidates a lane takes per lock-step iteration

// c[U] = src[U] for every U < kSelChunk with at + U < stop, as predicated 16-byte loads at immediate offsets
// from one base address.  A record that is not loaded is left UNDEFINED (whatever its registers held): the
// operands are declared write-only so that the buffers are not live -- 32 registers -- outside the scan loop.
template <int U>
__device__ __forceinline__ void load_records_if(float4 (&c)[kSelChunk], int at, int stop, const float4 *src) {
    if constexpr (U < kSelChunk) {
        asm("{\n\t.reg .pred p;\n\tsetp.lt.s32 p, %4, %5;\n\t"
            "@p ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%6+%7];\n\t}"
            : "=f"(c[U].x), "=f"(c[U].y), "=f"(c[U].z), "=f"(c[U].w)
            : "r"(at + U), "r"(stop), "l"(src), "n"(16 * U));
        load_records_if<U + 1>(c, at, stop, src);
    }
}

// MODE 0: block b works on queries [b * QB, (b + 1) * QB).
// MODE 1: the blocks to work on are named by P.block_list (what the tile kernel handed over), each resident block takes
//         several in turn.
// MODE 2: the SECOND CHANCE of the queries on the retry list, when that list is long (100 M points: 1 % of the queries,
//         half of them because their first sphere happened to hold fewer than k points, a third because a face of the
//         domain cut it): thread t of a block takes entry b * QB + t of the list and searches the sphere of
//         P.retry_radius_sq -- large enough for nearly all of them -- with a buffer that holds it.  Lock step and a sorting
//         network instead of the general kernel's per-lane loops and insertions, which ran with 8 of 32 lanes active
//         (profiles/r2_retry_general_100M_k20_*).  What still fails goes to P.retry_list2 and the general kernel.
template <int K_T, int QB, int NET, int CAP, int MODE>
__global__ void __launch_bounds__(QB) kn_query_select_kernel(const __grid_constant__ QueryParams P) {
    constexpr bool LIST = MODE == 1, QLIST = MODE == 2;
    static_assert(K_T > 0 && K_T + 1 <= NET - 8, "select path: k must leave room for new keys in a round");
    static_assert(NET == 32 || NET == 64, "select path: networks exist for 32 and 64 keys");
    constexpr int kSelectCap = CAP, kSelectNet = NET;
    constexpr unsigned kSlotMask = CAP <= 64 ? 63u : 127u;  // low key bits that hold the buffer slot
    constexpr int kSlotBits = CAP <= 64 ? 6 : 7;
    extern __shared__ unsigned long long smem[];
    const int tid = threadIdx.x;
    unsigned int *BUF = reinterpret_cast<unsigned int *>(smem) + tid;                                   // [kSelectCap][QB]
    int2 *SP = reinterpret_cast<int2 *>(reinterpret_cast<unsigned int *>(smem) + kSelectCap * QB) + tid;  // [kSpanSlots][QB]
    const int Dx = P.g.Dx, Dy = P.g.Dy, zb = P.g.zb, ze = P.g.ze;
    const float invw = P.g.invwx;
    const unsigned qtotal = QLIST ? *P.retry_count : 0u;
    if (QLIST && qtotal <= P.retry_short) return;  // a short list: the warp-per-query kernel takes it as it is
    if (QLIST && blockIdx.x == 0 && tid == 0) atomicAdd(P.diag + 0, qtotal);  // counted as restarts
    const unsigned nlist = QLIST ? (qtotal + QB - 1) / QB : LIST ? *P.block_count : gridDim.x;
  for (unsigned bi = blockIdx.x; bi < nlist; bi += gridDim.x) {
    const long long row0 = (long long)(LIST ? (unsigned)P.block_list[bi] : bi) * QB;
    const bool valid = QLIST ? row0 + tid < (long long)qtotal : row0 + tid < P.query_count;
    const int entry = QLIST ? P.retry_list[min(row0 + tid, (long long)qtotal - 1)] : 0;
    const int q = QLIST ? (entry & 0x7FFFFFFF) : P.query_begin + (int)min(row0 + tid, (long long)P.query_count - 1);
    const QueryGeom s = query_geom(__ldg(P.pts + q), P.g);
    // The second chance scales its sphere by the part of it that lies inside the domain: a query on a face of the cube
    // sees half a sphere and needs twice the volume for the same number of points (an estimate; what makes a row exact
    // is checked below, as ever).  Rows are walked as far as the widest sphere of the warp needs.
    float r2 = P.radius0_sq, r2max = P.radius0_sq;
    if constexpr (QLIST) {
        // per axis: 1 - the two spherical caps beyond the domain's faces, cap(t) = (1 - t)^2 (2 + t) / 4 at distance t r
        const float r1 = sqrtf(P.retry_radius_sq), invr = 1.f / r1;
        auto caps = [&](float p) {
            const float a = fminf(fmaxf(p, 0.f) * invr, 1.f), b = fminf(fmaxf(1000.f - p, 0.f) * invr, 1.f);
            return 1.f - 0.25f * ((1.f - a) * (1.f - a) * (2.f + a) + (1.f - b) * (1.f - b) * (2.f + b));
        };
        const float vx = caps(s.px), vy = caps(s.py), vz = caps(s.pz);
        const float visible = fmaxf(vx * vy * vz, 1.f / 64.f);
        // (an entry without bit 31 had its k points in the first sphere, but too many for the first kernel's buffer,
        // or a tie of truncated keys: the same sphere again, with this pass's larger buffer and its tie handling)
        r2 = entry < 0 ? fminf(P.retry_radius_sq * __powf(visible, -2.f / 3.f), P.retry_radius_max_sq) : P.radius0_sq;
    }
    // Local density.  The first radius comes from the cloud's MEAN density; inside a cluster its sphere holds many
    // times k points and overflows every buffer, in a void it holds none.  The points of the query's own row within the
    // first sphere's reach in x (two pointer loads) say how dense the cloud is HERE against the mean: a row holding
    // more than 2.5 x its share shrinks the sphere to the volume expected to hold the usual number (+ 20 %), in every
    // pass; the second chance also grows it for a query whose first sphere came up short in a row holding less than a
    // quarter of its share, and shrinks
    // once more for one that overflowed in a dense row.  (On a uniform cloud one row in 10^7 is 2.5 x its mean, one in 10^4
    // a quarter of it.)
    // An estimate only: what makes a row exact is checked below, as ever.
    {
        const float rx = sqrt_approx(P.radius0_sq) * invw;
        const int xlo = max(0, min(Dx - 1, (int)floorf(s.tx - rx))), xhi = max(0, min(Dx - 1, (int)floorf(s.tx + rx)));
        const int own = ((s.ck - zb) * Dy + s.cj) * Dx;
        const float have_row = (float)(__ldg(P.ptrs + own + xhi + 1) - __ldg(P.ptrs + own + xlo));
        const float mean_row = (float)(xhi - xlo + 1) * P.points_per_cell;
        if (have_row > 2.5f * mean_row) {
            r2 *= __powf((QLIST && entry >= 0 ? 0.7f : 1.2f) * mean_row / have_row, 2.f / 3.f);
        } else if (QLIST && entry < 0 && have_row < 0.25f * mean_row) {
            r2 = fminf(r2 * __powf(mean_row / fmaxf(have_row, 1.f), 2.f / 3.f), P.retry_radius_max_sq);
        }
    }
    r2max = r2;  // rows are walked as far as the widest sphere of the warp needs
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) r2max = fmaxf(r2max, __shfl_xor_sync(0xffffffffu, r2max, d));
    // Append pointer (shared-memory address of the lane's next free buffer slot).  It is clamped once per
    // chunk, so a chunk may run kSelChunk slots past `wlim`: reaching wlim at all counts as an overflow.
    const unsigned wp0 = (unsigned)__cvta_generic_to_shared(BUF);
    const unsigned wlim = wp0 + (unsigned)(kSelectCap - kSelChunk) * QB * 4u;
    unsigned wp = wp0;

    // ---- search the sphere: rows in batches, spans first, then a lock-step stream over them -----------------
    for (int e0 = 0, e1 = min(kFirstBatch, P.tab.n); e0 < P.tab.n && P.tab.lb2[e0] <= r2max;
         e0 = e1, e1 = min(e0 + kNextBatch, P.tab.n)) {
        int nsp = 0;
        for (int e = e0; e < e1; e += 4) {
            int alo[4], ahi[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                ok[u] = false;
                const int ee = e + u;
                if (ee < e1 && valid && P.tab.lb2[ee] <= r2) {
                    const int dj = P.tab.dj[ee], dk = P.tab.dk[ee];
                    const int j = s.cj + dj, k = s.ck + dk;
                    if (j >= 0 && j < Dy && k >= zb && k < ze) {
                        const float lbyz2 = row_lower_bound2(s, dj, dk, P.g);
                        if (lbyz2 <= r2) {
                            int ilo, ihi;
                            row_x_range(s, r2, lbyz2, invw, Dx, ilo, ihi);
                            const int row = ((k - zb) * Dy + j) * Dx;
                            alo[u] = row + ilo;
                            ahi[u] = row + ihi + 1;
                            ok[u] = ilo <= ihi;
                        }
                    }
                }
            }
            int b4[4], e4[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) { b4[u] = __ldg(P.ptrs + alo[u]); e4[u] = __ldg(P.ptrs + ahi[u]); }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u] && b4[u] < e4[u]) { SP[nsp * QB] = make_int2(b4[u], e4[u]); ++nsp; }
        }
        // Lock-step stream over the lane's spans, kSelChunk candidates per iteration, software-pipelined:
        // the records of the NEXT chunk are requested before the current one is evaluated (two register
        // buffers used in turn; the scan is bound by load latency at this occupancy).
        int sp = 0, p = 0, end = 0;
        if (nsp > 0) { const int2 t = SP[0]; p = t.x; end = t.y; sp = 1; }
        // Loads and the append are predicated instructions written out by hand.  The compiler's version
        // zeroed the four destination registers before every predicated load and spent seven instructions
        // on "if (pass) BUF[min(cnt, cap)] = i; cnt += pass" (select 0/1, add, clamp, multiply-add, store):
        // 23 instructions per candidate in all, in a kernel that is bound by instruction issue.  A record that
        // is not loaded holds stale values, which is harmless: the append tests the same index bound again.
        auto fetch = [&](float4(&c)[kSelChunk], int at, int stop) {
            load_records_if<0>(c, at, stop, P.pts + at);
        };
        // The query itself is appended too (it is inside every sphere) and dropped when the keys are rebuilt:
        // one test per buffered candidate instead of one per scanned candidate (self excluded, knearests.cu:123).
        auto evaluate = [&](float4(&c)[kSelChunk], int at, int stop) {
#pragma unroll
            for (int u = 0; u < kSelChunk; ++u) {
                const float dx = c[u].x - s.px, dy = c[u].y - s.py, dz = c[u].z - s.pz;
                const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
                asm volatile("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\tsetp.lt.and.s32 p, %3, %4, p;\n\t"
                             "@p st.shared.u32 [%0], %3;\n\t@p add.u32 %0, %0, %5;\n\t}"
                             : "+r"(wp)
                             : "f"(d2), "f"(r2), "r"(at + u), "r"(stop), "n"(QB * 4));
            }
            wp = min(wp, wlim);
#pragma unroll
            for (int u = 0; u < kSelChunk; ++u) asm volatile("" ::"f"(c[u].w));  // keep .w allocated (see above)
        };
        auto next_chunk = [&](int &pn, int &endn, int &spn) {
            pn = p + kSelChunk; endn = end; spn = sp;
            if (pn >= end && sp < nsp) { const int2 t = SP[sp * QB]; pn = t.x; endn = t.y; spn = sp + 1; }
        };
        float4 ca[kSelChunk], cb[kSelChunk];
        fetch(ca, p, end);
        while (__any_sync(0xffffffffu, p < end)) {
            int pn, endn, spn;
            next_chunk(pn, endn, spn);
            fetch(cb, pn, endn);
            evaluate(ca, p, end);
            p = pn; end = endn; sp = spn;
            if (!__any_sync(0xffffffffu, p < end)) break;
            next_chunk(pn, endn, spn);
            fetch(ca, pn, endn);
            evaluate(cb, p, end);
            p = pn; end = endn; sp = spn;
        }
    }

    // ---- select ------------------------------------------------------------------------------------------
    // Rebuild each buffered candidate's d2 (its record is still in L1/L2) and sort 32-BIT keys
    //     (bits(d2) with the low 6 or 7 bits replaced by the buffer slot)
    // 32 at a time with the Batcher network: a compare-exchange is one integer min and one max, the keys
    // take 32 registers (64-bit keys took 64 and cost the scan its occupancy).  The 6 dropped mantissa bits
    // can only matter when two of the k+1 smallest truncated distances coincide; that is checked (k adjacent
    // compares) and such a query is retried exactly by the general kernel -- about 0.5 % on uniform clouds,
    // every query on a lattice.  Otherwise ascending truncated keys ARE ascending (d2, position).
    static_assert(kSelectCap <= 128, "slot must fit the low key bits");
    asm volatile("" ::: "memory");  // the appends above were written behind the compiler's back
    const int cnt = (int)((wp - wp0) / (QB * 4u)) - 1;  // candidates inside the sphere, the query itself not counted
    const bool searched = valid && cnt >= K_T && wp < wlim;
    const int have = searched ? cnt + 1 : 0;
    unsigned a[kSelectNet];
    for (int base = 0, round = 0; round == 0 || __any_sync(0xffffffffu, have > base); ++round) {
        // Later rounds keep the K_T + 1 smallest so far: the extra one guards the border -- whatever is
        // dropped has a key >= a[K_T], so if a[K_T-1] and a[K_T] differ in their truncated distance at
        // the end, nothing that was dropped could have belonged to the row.
        const int lo = round == 0 ? 0 : K_T + 1;
#pragma unroll
        for (int u = 0; u < kSelectNet; ++u) {
            if (u >= lo) {
                const int slot = base + (u - lo);
                a[u] = 0xFFFFFFFFu;
                if (slot < have) {
                    const int pos = (int)BUF[slot * QB];
                    const unsigned key = ((unsigned)(make_key(__ldg(P.pts + pos), s, pos) >> 32) & ~kSlotMask) | (unsigned)slot;
                    a[u] = pos == q ? 0xFFFFFFFFu : key;
                }
            }
        }
        base += kSelectNet - lo;
#define KN_CE(i, j)                                     \
    {                                                   \
        const unsigned lo_ = min(a[i], a[j]);           \
        a[j] = max(a[i], a[j]);                         \
        a[i] = lo_;                                     \
    }
        if constexpr (NET == 32) { KN_SORT_NETWORK_32(KN_CE) } else { KN_SORT_NETWORK_64(KN_CE) }
#undef KN_CE
    }
    // Exact if the truncated distances decide the order.  The second chance (MODE 2) also accepts EQUAL truncated
    // distances next to each other inside the row and settles each such pair from the exact keys once the row is
    // written, like the tile kernel; a tie at the k / k+1 border, or three in a row, still go on to the general kernel.
    bool exact = searched;
    unsigned long long ties = 0;  // bit u: entries u and u + 1 have the same truncated distance
#pragma unroll
    for (int u = 0; u < K_T; ++u) {
        if constexpr (QLIST) {
            if ((a[u] >> kSlotBits) == (a[u + 1] >> kSlotBits)) ties |= 1ull << u;  // a[K_T] exists: K_T < NET
        } else {
            exact = exact && (a[u] >> kSlotBits) < (a[u + 1] >> kSlotBits);
        }
    }
    if (QLIST) exact = exact && !(ties >> (K_T - 1)) && !(ties & (ties << 1));
    // bit 31 of a retry entry: the sphere held fewer than k points, so the retry must search a larger one
    // (an overflowing buffer or a truncated-key tie is retried with the same radius)
    if (valid && !exact) {
        if (QLIST) P.retry_list2[atomicAdd(P.retry_count2, 1u)] = q | (cnt < K_T ? (int)0x80000000 : 0);
        else P.retry_list[atomicAdd(P.retry_count, 1u)] = q | (cnt < K_T ? (int)0x80000000 : 0);
    }
    report_diag(P, s, exact, exact ? __uint_as_float((a[K_T - 1] | kSlotMask)) : 0.f, 0u, tid);

    // ---- coalesced row output (through the buffer area: read all k positions, then overwrite) -----------------
    unsigned res[K_T];
#pragma unroll
    for (int u = 0; u < K_T; ++u) res[u] = exact ? BUF[(a[u] & kSlotMask) * QB] + P.index_offset : KNC_EMPTY;
    if constexpr (QLIST) {  // scattered rows: each thread writes its own (and only a finished one)
        if (exact) {
            unsigned int *out = P.knn + (long long)(q - P.query_begin) * K_T;
#pragma unroll
            for (int u = 0; u < K_T; ++u) out[u] = res[u];
            while (ties) {  // rare, divergent: order each tied pair by its exact (d2, position)
                const int u = __ffsll((long long)ties) - 1;
                ties &= ties - 1;
                const unsigned pa = out[u], pb = out[u + 1];
                const unsigned long long ka = make_key(__ldg(P.pts + (pa - P.index_offset)), s, (int)pa);
                const unsigned long long kb = make_key(__ldg(P.pts + (pb - P.index_offset)), s, (int)pb);
                if (kb < ka) { out[u] = pb; out[u + 1] = pa; }
            }
        }
        continue;
    }
    // Row stride QB + 1: the transposed read below (consecutive threads walk along a row's k entries, i.e.
    // down a column of OUT) would otherwise hit one bank 16 times over (370 of 2600 L1 wavefronts per warp,
    // profiles/r1_notes.md).
    __syncthreads();
    unsigned int *OUT = reinterpret_cast<unsigned int *>(smem);
#pragma unroll
    for (int u = 0; u < K_T; ++u) OUT[u * (QB + 1) + tid] = res[u];
    __syncthreads();
    const int rows = (int)min((long long)QB, (long long)P.query_count - row0);
    unsigned int *out = P.knn + row0 * K_T;
    for (int idx = tid; idx < rows * K_T; idx += QB) {
        const int t = idx / K_T, sl = idx - t * K_T;
        out[idx] = OUT[sl * (QB + 1) + t];
    }
    if (LIST) __syncthreads();  // the staging area is the next block's candidate buffer
  }
}

// ======================================================================================== tile path
// The candidates of a block of QB consecutive queries staged ONCE in shared memory.
//
// Precondition (checked by the launcher): the cells are at least as wide in y and z as the first search radius,
// so a query's sphere cannot reach past the 3 x 3 x-rows of cells around its own row.  QB consecutive queries lie in
// 1 (sometimes 2 or 3) x-rows of cells; for each such "query row" the block stages the 9 surrounding rows, each
// cut to the x-range of its queries plus the reach of the sphere: <= 27 SEGMENTS, every one a contiguous run of
// 16-byte records in global memory, hence one cp.async.bulk each, all completing on one mbarrier.  Lane L of warp 0
// sets up segment L (two pointer loads, a warp prefix sum for the place in the tile) and issues its copy.
//
// While the copies fly every lane works out its own windows ("spans") of those rows as before, in global
// positions, then rebases them to tile slots.  The scan needs NO bounds test and no predicated load: a span
// lies inside a segment, between records of the same row that are beyond the sphere's reach in x (they fail the
// distance test like any other); a segment lies between runs of kTilePad records of +inf; and a lane that has run
// out of spans parks on +inf records at the head of the tile -- so the body is 4 x LDS.128, the arithmetic, and a
// predicated 16-bit append of the tile slot.
//
// Bank conflicts: 32 lanes reading 32 unrelated 16-byte records cost ~8.3 shared-memory wavefronts per LDS.128
// where 4 is the minimum (profiles/r2_notes.md), and the shared-memory pipe is what bounds this kernel.  So every
// lane starts a span up to 7 records EARLY, on a slot that is congruent to (lane + 4 x iteration) modulo 8: the
// eight lanes of a quarter warp then always touch eight different 16-byte bank groups -- 4 wavefronts, always --
// at the price of 3.5 extra records per span (records before a span fail the distance test like those after it).  Selection as in the select kernel, keys rebuilt from the tile; the neighbour's
// sorted position is the 4th word of its record.  The block's rows are laid out in shared memory exactly as in
// global memory and leave as ONE bulk copy.
//
// A block whose queries span more than kTileQueryRows rows, or whose tile would not fit (clustered clouds), names
// itself in P.block_list and is served by the select kernel afterwards.
constexpr int kTileQueryRows = 3;
constexpr int kTileSegs = 9 * kTileQueryRows;  // <= 32: one lane of warp 0 each
constexpr int kTilePad = 8;                    // +inf records after every segment: 3 read past a span's end, 7 before the next
constexpr int kTileHead = 16;                  // +inf records at the head of the tile: where lanes without work park

__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");  // visible to the async proxy
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned bar, unsigned parity) {
    unsigned done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done != 0;
}
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completion counted on the mbarrier
__device__ __forceinline__ void bulk_load(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// shared -> global; the caller waits (bulk_store_wait) before the shared memory may go away
__device__ __forceinline__ void bulk_store(void *dst, unsigned src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_wait() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// {c.x - px, c.y - py} in ONE instruction (FADD2 on sm_100a): the record's x, y and the query's x, y each sit in an
// aligned register pair.  IEEE per component, so the differences are the reference's (knearests.cu:125).
__device__ __forceinline__ float2 sub2(float ax, float ay, unsigned long long pxy) {
    unsigned long long a, r;
    float2 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(ax), "f"(ay));
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(pxy));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(r));
    return d;
}
__device__ __forceinline__ float2 sq2(float2 v) {  // {v.x * v.x, v.y * v.y}, each rounded (FMUL2)
    unsigned long long a, r;
    float2 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(v.x), "f"(v.y));
    asm("mul.rn.f32x2 %0, %1, %1;" : "=l"(r) : "l"(a));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(r));
    return d;
}

template <int K_T, int QB, int NET, int CAP, int TCAP>
__global__ void __launch_bounds__(QB) kn_query_tile_kernel(const __grid_constant__ QueryParams P) {
    static_assert(K_T > 0 && K_T + 1 <= NET - 8, "tile path: k must leave room for new keys in a round");
    static_assert(NET == 32 || NET == 64, "tile path: networks exist for 32 and 64 keys");
    static_assert(TCAP < 65536 && TCAP % 8 == 0, "tile slots are stored in 16 bits");
    constexpr unsigned kSlotMask = CAP <= 64 ? 63u : 127u;  // low key bits that hold the buffer slot
    constexpr int kSlotBits = CAP <= 64 ? 6 : 7;
    static_assert(CAP <= 128, "slot must fit the low key bits");
    // Shared memory: the tile, then one PRIVATE region per warp (private, so that a warp that is done can stage its
    // rows there without a block barrier): CAP + 2 rows of 32 16-bit words, one column per lane.  A lane's column holds
    // its candidate buffer growing UP from row 0 and its span list -- two rows per span, begin and end slot -- in the
    // rows just below CAP, first span lowest; rows CAP and CAP + 1 hold the terminator.  Spans are consumed as the
    // buffer grows, so the two share the column: the buffer may use every row below the next unread span (a lane that
    // would need more is an overflow and goes to the retry kernel).  That sharing is what lets 5 blocks fit an SM.
    // (Sharing is used where it buys a fifth block, the 32-key variants; the 64-key variants keep their 9 spans in rows
    // of their own above the buffer: at k = 50 sharing was 8 % slower at the same 3 blocks per SM, and no faster when a
    // tile of 1920 records made it a fourth block -- 59.9 against 59.4 ms at 100 M points.)
    constexpr bool kShareColumn = NET == 32;
    constexpr int kTermRow = kShareColumn ? CAP : CAP + 18;  // the terminator's rows: kTermRow, kTermRow + 1
    constexpr int kWarpBytes = (kTermRow + 2) * 64;
    static_assert(kWarpBytes >= 32 * K_T * 4, "a warp's rows are staged in its own region");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ int s_adj[32];      // per segment: tile slot = s_adj + global sorted position
    __shared__ int s_ends[4];      // [0]: query row of the block's first query
    __shared__ int s_fallback;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float4 *tile = reinterpret_cast<float4 *>(smem_raw);                                  // [TCAP] records
    unsigned char *mine = smem_raw + (size_t)TCAP * 16 + (size_t)warp * kWarpBytes;          // this warp's region
    unsigned short *BUF = reinterpret_cast<unsigned short *>(mine) + lane;                   // [CAP + 2][32], this lane's column
    const int Dx = P.g.Dx, Dy = P.g.Dy, zb = P.g.zb, ze = P.g.ze;
    const long long row0 = (long long)blockIdx.x * QB;
    const int rows = (int)min((long long)QB, (long long)P.query_count - row0);
    const bool valid = tid < rows;
    const int q = P.query_begin + (int)row0 + min(tid, rows - 1);
    const QueryGeom s = query_geom(__ldg(P.pts + q), P.g);
    const int qrow = (s.ck - zb) * Dy + s.cj;
    const unsigned bar = smem_addr(&s_bar);
    const float4 inf4 = make_float4(__int_as_float(0x7f800000), __int_as_float(0x7f800000), __int_as_float(0x7f800000), 0.f);

    // ---- warp 0: the segments.  Lane = (query row qi, candidate row dj, dk).  It looks at the block's first and last
    // query itself (two broadcast loads), so nobody waits for anybody before the copies are on their way. ----------
    if (tid < 32) {
        if (tid == 0) mbar_init(bar, 1);
        const QueryGeom sf = query_geom(__ldg(P.pts + P.query_begin + (int)row0), P.g);
        const QueryGeom sl = query_geom(__ldg(P.pts + P.query_begin + (int)row0 + rows - 1), P.g);
        const int cif = max(0, min(Dx - 1, (int)floorf(sf.tx)));  // the build's cells (same arithmetic, knearests.cu:23-28)
        const int cil = max(0, min(Dx - 1, (int)floorf(sl.tx)));
        const int qrow0 = (sf.ck - zb) * Dy + sf.cj, nqr = (sl.ck - zb) * Dy + sl.cj - qrow0 + 1;
        __syncwarp();  // the barrier is initialised before any lane uses it
        int len = 0, g0 = 0;
        const bool rows_ok = nqr <= kTileQueryRows;
        if (rows_ok && lane < nqr * 9) {
            const int qi = lane / 9, rr = lane - 9 * qi, dj = rr % 3 - 1, dk = rr / 3 - 1;
            const int R = qrow0 + qi, j = R % Dy + dj, k = R / Dy + zb + dk;
            if (j >= 0 && j < Dy && k >= zb && k < ze) {
                const int xlo = qi == 0 ? cif : 0, xhi = qi == nqr - 1 ? cil : Dx - 1;
                const int xa = max(0, xlo - P.halo_cells), xb = min(Dx - 1, xhi + P.halo_cells);
                const int base = ((k - zb) * Dy + j) * Dx;
                g0 = __ldg(P.ptrs + base + xa);
                len = __ldg(P.ptrs + base + xb + 1) - g0;
            }
        }
        if (lane == 0) s_ends[0] = qrow0;
        const int need = len > 0 ? len + kTilePad : 0;
        int incl = need, bytes = len * 16;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) bytes += __shfl_xor_sync(0xffffffffu, bytes, d);
        const int total = __shfl_sync(0xffffffffu, incl, 31) + kTileHead;
        const int off = kTileHead + incl - need;
        const bool ok = rows_ok && total <= TCAP;
        if (ok) {
            s_adj[lane] = off - g0;
            if (lane == 0) mbar_arrive_expect_tx(bar, (unsigned)bytes);  // the one arrival of the phase, with the byte
            __syncwarp();                                                // count of all copies, before any of them
            if (lane < kTileHead) tile[lane] = inf4;
            if (len > 0) {
                bulk_load(smem_addr(tile + off), P.pts + g0, (unsigned)len * 16u, bar);
#pragma unroll
                for (int u = 0; u < kTilePad; ++u) tile[off + len + u] = inf4;
            }
        }
        if (lane == 0) s_fallback = ok ? 0 : 1;
    }

    // ---- every lane: its windows of the 3 x 3 rows, in global sorted positions -----------------------------------
    // row_lower_bound2 specialised for |dj|, |dk| <= 1 (same values): the gap to the row below is the query's offset
    // in its cell, to the row above one cell less that, to its own row none -- a query outside the domain sits in a
    // border cell and the row beyond that border does not exist.
    const float r2 = P.radius0_sq;
    int gb[9], ge[9];
    {
        const float lym = fmaxf(0.f, fmaf(s.fy, P.g.wy, -kMargin)), lyp = fmaxf(0.f, fmaf(1.f - s.fy, P.g.wy, -kMargin));
        const float lzm = fmaxf(0.f, fmaf(s.fz, P.g.wz, -kMargin)), lzp = fmaxf(0.f, fmaf(1.f - s.fz, P.g.wz, -kMargin));
        const float ly2[3] = {lym * lym, 0.f, lyp * lyp}, lz2[3] = {lzm * lzm, 0.f, lzp * lzp};
        const bool ey[3] = {s.cj > 0, true, s.cj < Dy - 1}, ez[3] = {s.ck > zb, true, s.ck < ze - 1};
        const int base0 = ((s.ck - zb) * Dy + s.cj) * Dx;
        const float reach = kMargin * P.g.invwx + 1e-3f;
#pragma unroll
        for (int rr = 0; rr < 9; ++rr) {
            const int dj = rr % 3 - 1, dk = rr / 3 - 1;
            gb[rr] = 0; ge[rr] = 0;
            const float lbyz2 = (ly2[dj + 1] + lz2[dk + 1]) * kShrink;
            if (valid && ey[dj + 1] && ez[dk + 1] && lbyz2 <= r2) {
                const float rxc = fmaf(sqrt_approx(r2 - lbyz2) * kGrow, P.g.invwx, reach);  // = row_x_range
                const int ilo = max(0, min(Dx - 1, (int)floorf(s.tx - rxc))), ihi = max(0, min(Dx - 1, (int)floorf(s.tx + rxc)));
                const int base = base0 + (dk * Dy + dj) * Dx;
                gb[rr] = __ldg(P.ptrs + base + ilo);
                ge[rr] = __ldg(P.ptrs + base + ihi + 1);
            }
        }
    }
    __syncthreads();  // the only block-wide barrier: the segment table (and the mbarrier) are ready
    if (s_fallback) {  // block-uniform
        if (tid == 0) P.block_list[atomicAdd(P.block_count, 1u)] = (int)blockIdx.x;
        return;
    }
    const int qi = qrow - s_ends[0];
    int nsp = 0;
#pragma unroll
    for (int rr = 0; rr < 9; ++rr) nsp += ge[rr] > gb[rr] ? 1 : 0;
    {
        // Scan order: corner rows, face rows, the own row LAST -- the rows with the fewest points inside the sphere come
        // first, while the buffer's limit (below the unread spans) is lowest.
        constexpr int kOrder[9] = {0, 2, 6, 8, 1, 3, 5, 7, 4};
        unsigned short *sp = BUF + (kTermRow - 2 * nsp) * 32;  // first span lowest, the last one just below the terminator
#pragma unroll
        for (int o = 0; o < 9; ++o) {
            const int rr = kOrder[o];
            if (ge[rr] > gb[rr]) {
                const int adj = s_adj[qi * 9 + rr];
                sp[0] = (unsigned short)(gb[rr] + adj);
                sp[32] = (unsigned short)(ge[rr] + adj);
                sp += 64;
            }
        }
        BUF[kTermRow * 32] = 8;       // terminator: begin 8 (parks inside the head of the tile) ...
        BUF[(kTermRow + 1) * 32] = 0; // ... end 0 -- recognised by end == 0
    }
    const int my_slot = q + s_adj[qi * 9 + 4];

    // Append pointer (shared-memory address of the lane's next free buffer entry), clamped once per chunk to `wlim`,
    // kSelChunk rows below the next unread span (a chunk may run kSelChunk entries past it): reaching the limit at any
    // time counts as an overflow -- `slack` keeps the smallest distance seen when the limit moved.
    const unsigned wp0 = smem_addr(BUF);
    const unsigned spterm = wp0 + (unsigned)kTermRow * 64u;  // the terminator entry
    unsigned wp = wp0;
    unsigned wlim = wp0 + (unsigned)(CAP - kSelChunk) * 64u;  // (own span rows: the buffer's limit never moves)
    if (kShareColumn) wlim = spterm - (unsigned)max(nsp - 1, 0) * 128u - 256u;  // below the SECOND span: the first is taken at once
    int slack = 1 << 20;

    while (!mbar_try_wait(bar, 0)) {}  // the tile has landed

    // ---- scan: lock step over the lanes' own span lists, kSelChunk records per iteration, software-pipelined: the
    // records of the NEXT chunk are requested before the current one is evaluated.  The test here is a PREFILTER
    // (packed arithmetic: {dx, dy} and their squares in one instruction each; the sum is within 2 ulp of the
    // reference's expression, the radius is widened by 2^-20 to cover it): whatever passes gets its exact distance
    // when the keys are built, and the row is accepted only if its k-th exact distance lies inside the true sphere.
    {
        static_assert(kSelChunk == 4, "the conflict-free start below assumes chunks of 4 records");
        const float r2a = r2 * (1.0f + 1.0f / 1048576.0f);
        unsigned long long pxy;
        asm("mov.b64 %0, {%1, %2};" : "=l"(pxy) : "f"(s.px), "f"(s.py));
        int ph = lane;  // slot residue (mod 8) this lane must read at in the coming iteration: lane + 4 x (iteration & 1)
        unsigned spaddr = spterm - (unsigned)nsp * 128u;          // the entry the next switch reads: the first span ...
        int ts = (int)BUF[(kTermRow - 2 * nsp) * 32], end = (int)BUF[(kTermRow - 2 * nsp + 1) * 32];  // ... which is taken here
        spaddr = min(spaddr + 128u, spterm);
        ts -= (ts - ph) & 7;
        auto fetch = [&](float4(&c)[kSelChunk], int at) {
#pragma unroll
            for (int u = 0; u < kSelChunk; ++u) c[u] = tile[at + u];
        };
        auto evaluate = [&](const float4(&c)[kSelChunk], int at) {
#pragma unroll
            for (int u = 0; u < kSelChunk; ++u) {
                const float2 q2 = sq2(sub2(c[u].x, c[u].y, pxy));
                const float dz = c[u].z - s.pz;
                const float d2 = __fmaf_rn(dz, dz, __fadd_rn(q2.x, q2.y));
                asm volatile("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t"
                             "@p st.shared.u16 [%0], %3;\n\t@p add.u32 %0, %0, 64;\n\t}"
                             : "+r"(wp)
                             : "f"(d2), "f"(r2a), "r"(at + u));
            }
            wp = min(wp, wlim);
        };
        // One step along the lane's span list, straight-line predicated code (the compiler made a divergent branch
        // of it): the next chunk of the span, or -- past its end -- the first chunk of the next span, started on the
        // slot residue of the coming iteration; the buffer's limit moves up with the span list.
        auto advance = [&]() {
            if constexpr (kShareColumn) {
                asm volatile("{\n\t.reg .pred p;\n\t.reg .u32 t, b;\n\t"
                             "add.s32 %0, %0, 4;\n\t"
                             "xor.b32 %3, %3, 4;\n\t"
                             "setp.ge.s32 p, %0, %1;\n\t"
                             "@p ld.shared.u16 b, [%2];\n\t"
                             "@p ld.shared.u16 %1, [%2+64];\n\t"
                             "@p sub.s32 t, %4, %7;\n\t"         // distance to the limit that is about to move
                             "@p min.s32 %5, %5, t;\n\t"
                             "@p sub.s32 t, b, %3;\n\t"
                             "@p and.b32 t, t, 7;\n\t"
                             "@p sub.s32 %0, b, t;\n\t"
                             "@p add.u32 %2, %2, 128;\n\t"
                             "@p min.u32 %2, %2, %6;\n\t"
                             "@p sub.u32 %4, %2, 256;\n\t}"       // kSelChunk rows below the next unread span
                             : "+r"(ts), "+r"(end), "+r"(spaddr), "+r"(ph), "+r"(wlim), "+r"(slack)
                             : "r"(spterm), "r"(wp)
                             : "memory");
            } else {
                asm volatile("{\n\t.reg .pred p;\n\t.reg .u32 t, b;\n\t"
                             "add.s32 %0, %0, 4;\n\t"
                             "xor.b32 %3, %3, 4;\n\t"
                             "setp.ge.s32 p, %0, %1;\n\t"
                             "@p ld.shared.u16 b, [%2];\n\t"
                             "@p ld.shared.u16 %1, [%2+64];\n\t"
                             "@p sub.s32 t, b, %3;\n\t"
                             "@p and.b32 t, t, 7;\n\t"
                             "@p sub.s32 %0, b, t;\n\t"
                             "@p add.u32 %2, %2, 128;\n\t"
                             "@p min.u32 %2, %2, %4;\n\t}"
                             : "+r"(ts), "+r"(end), "+r"(spaddr), "+r"(ph)
                             : "r"(spterm)
                             : "memory");
            }
        };
        float4 ca[kSelChunk], cb[kSelChunk];
        fetch(ca, ts);
        while (true) {
            int tc = ts;
            advance();
            fetch(cb, ts);
            evaluate(ca, tc);
            if (!__any_sync(0xffffffffu, end != 0)) break;
            tc = ts;
            advance();
            fetch(ca, ts);
            evaluate(cb, tc);
            if (!__any_sync(0xffffffffu, end != 0)) break;
        }
    }

    // ---- select (see kn_query_select_kernel): 32-bit keys = bits(d2) with the low bits replaced by the buffer slot ----
    asm volatile("" ::: "memory");  // the appends above were written behind the compiler's back
    const int cnt = (int)((wp - wp0) / 64u) - 1;  // candidates that passed the prefilter, the query itself not counted
    const bool searched = valid && cnt >= K_T && wp < wlim && slack > 0;  // wlim ends kSelChunk rows below the terminator
    const int have = searched ? cnt + 1 : 0;
    unsigned a[NET];
    for (int base = 0, round = 0; round == 0 || __any_sync(0xffffffffu, have > base); ++round) {
        // Keys of the next NET - lo buffered candidates, eight at a time and without branches (all slot reads, then all
        // record reads, then the arithmetic: the loads of a group overlap).  A slot beyond `have` reads the +inf record
        // at the head of the tile and becomes the largest key.
        const int lo = round == 0 ? 0 : K_T + 1;
#pragma unroll
        for (int g = 0; g < NET; g += 8) {
            if (g + 8 > lo) {
                int ti[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int slot = base + (g + j - lo);
                    ti[j] = (g + j >= lo && slot < have) ? (int)BUF[min(slot, kTermRow + 1) * 32] : 0;  // (the row index stays inside the warp's region even if the load is hoisted)
                }
                float4 rec[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) rec[j] = tile[ti[j]];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (g + j >= lo) {
                        const int slot = base + (g + j - lo);
                        const unsigned key = ((unsigned)(make_key(rec[j], s, 0) >> 32) & ~kSlotMask) | (unsigned)slot;
                        a[g + j] = (slot < have && ti[j] != my_slot) ? key : 0xFFFFFFFFu;  // self excluded, knearests.cu:123
                    }
                }
            }
        }
        base += NET - lo;
        // Round 0 sorts all NET keys.  Later rounds hold K_T + 1 sorted keys and NET - K_T - 1 new ones: the block of the
        // last kTail keys (it contains every new one) is sorted DESCENDING, which leaves the whole array bitonic, and
        // one bitonic merge -- log2(NET) stages -- sorts it: 140 compare-exchanges instead of 191 at k = 20, 252
        // instead of 543 at k = 50.
        // (Used where it pays: k = 50, 75.6 -> 70.2 ms at 100 M points.  At k = 20 and 32 the second network made the
        // kernel larger and it ran 1 % slower -- instruction fetch already stalls this kernel -- so they sort again.)
        constexpr int kTail = (NET == 64 && K_T + 1 >= NET - 16) ? 16 : NET;
        if (round == 0 || kTail == NET) {
#define KN_CE(i, j)                                     \
    {                                                   \
        const unsigned lo_ = min(a[i], a[j]);           \
        a[j] = max(a[i], a[j]);                         \
        a[i] = lo_;                                     \
    }
            if constexpr (NET == 32) { KN_SORT_NETWORK_32(KN_CE) } else { KN_SORT_NETWORK_64(KN_CE) }
#undef KN_CE
        } else {
#define KN_CE(i, j)                                                      \
    {                                                                    \
        const unsigned hi_ = max(a[NET - kTail + i], a[NET - kTail + j]); \
        a[NET - kTail + j] = min(a[NET - kTail + i], a[NET - kTail + j]); \
        a[NET - kTail + i] = hi_;                                        \
    }
            if constexpr (kTail == 16) { KN_SORT_NETWORK_16(KN_CE) } else if constexpr (kTail == 32) { KN_SORT_NETWORK_32(KN_CE) }
#undef KN_CE
#pragma unroll
            for (int stride = NET / 2; stride >= 1; stride /= 2)
#pragma unroll
                for (int i = 0; i < NET; ++i)
                    if ((i & stride) == 0) {
                        const unsigned lo_ = min(a[i], a[i + stride]);
                        a[i + stride] = max(a[i], a[i + stride]);
                        a[i] = lo_;
                    }
        }
    }
    // Exact if the k-th lies inside the TRUE sphere (the prefilter's is wider) and the truncated distances decide the
    // order.  Two EQUAL truncated distances next to each other inside the row (0.5 % of the queries have such a pair)
    // only leave the order of those two open: it is settled below from their exact distances, once the row is staged.
    // A tie at the k / k+1 border, or three in a row, still go to the retry kernel.
    const bool inside = (a[K_T - 1] | kSlotMask) <= __float_as_uint(r2);
    // (Worth it for the 64-key variants only -- 7 bits truncated, more neighbours: k = 50 loses a third of its retries,
    // 83.3 -> 77.6 ms at 100 M points; at k = 20 such pairs are 0.08 % of the queries and the bookkeeping cost more
    // than the retries it saved.)
    constexpr bool kSettleTies = NET == 64;
    unsigned long long ties = 0;  // bit u: entries u and u + 1 have the same truncated distance
    bool strictly = true;
#pragma unroll
    for (int u = 0; u < K_T; ++u) {
        if constexpr (kSettleTies) {
            if ((a[u] >> kSlotBits) == (a[u + 1] >> kSlotBits)) ties |= 1ull << u;
        } else {
            strictly = strictly && (a[u] >> kSlotBits) < (a[u + 1] >> kSlotBits);
        }
    }
    const bool exact = searched && inside && strictly && !(ties >> (K_T - 1)) && !(ties & (ties << 1));
    if (!exact) ties = 0;
    if (valid && !exact)
        P.retry_list[atomicAdd(P.retry_count, 1u)] = q | ((cnt < K_T || (searched && !inside)) ? (int)0x80000000 : 0);
    report_diag(P, s, exact, exact ? __uint_as_float((a[K_T - 1] | kSlotMask)) : 0.f, 0u, tid);

    unsigned res[K_T];
#pragma unroll
    for (int u = 0; u < K_T; ++u)
        res[u] = exact ? __float_as_uint(tile[BUF[(a[u] & kSlotMask) * 32]].w) + P.index_offset : KNC_EMPTY;

    // ---- rows: every warp stages its 32 rows in its own region, laid out exactly as in global memory, and sends them
    // off as one bulk copy; no block-wide barrier, a warp that is done is done ------------------------------------------
    __syncwarp();  // every lane of the warp is done reading the warp's buffer
    unsigned *OUT = reinterpret_cast<unsigned *>(mine);
    if constexpr (K_T % 4 == 0) {
        uint4 *dst = reinterpret_cast<uint4 *>(OUT + lane * K_T);
#pragma unroll
        for (int v = 0; v < K_T / 4; ++v) dst[v] = make_uint4(res[4 * v], res[4 * v + 1], res[4 * v + 2], res[4 * v + 3]);
    } else {
        static_assert(K_T % 2 == 0, "row staging writes pairs");
        uint2 *dst = reinterpret_cast<uint2 *>(OUT + lane * K_T);
#pragma unroll
        for (int v = 0; v < K_T / 2; ++v) dst[v] = make_uint2(res[2 * v], res[2 * v + 1]);
    }
    if (ties) {  // rare, divergent: order each tied pair by its exact (d2, position) -- the records come from global memory
        unsigned *row = OUT + lane * K_T;
        do {
            const int u = __ffsll((long long)ties) - 1;
            ties &= ties - 1;
            const unsigned pa = row[u], pb = row[u + 1];
            const unsigned long long ka = make_key(__ldg(P.pts + (pa - P.index_offset)), s, (int)pa);
            const unsigned long long kb = make_key(__ldg(P.pts + (pb - P.index_offset)), s, (int)pb);
            if (kb < ka) { row[u] = pb; row[u + 1] = pa; }
        } while (ties);
    }
    const int wrows = min(32, rows - 32 * warp);  // rows of this warp (<= 0: none)
    unsigned int *out = P.knn + (row0 + 32 * warp) * K_T;
    const int words = wrows * K_T;
    if (P.bulk_rows_ok && (words & 3) == 0) {  // warp-uniform
        fence_async_smem();  // the staged rows must be visible to the copy engine
        __syncwarp();
        if (lane == 0 && words > 0) {
            bulk_store(out, smem_addr(OUT), (unsigned)words * 4u);
            bulk_store_wait();
        }
    } else {
        __syncwarp();
        for (int idx = lane; idx < words; idx += 32) out[idx] = OUT[idx];
    }
}

constexpr int kRetryWarpFactor = 16;  // see kn_query_retry_warp_kernel: up to ~76 k queries (measured: 50 k retries 0.35 ms here against ~0.9 ms in the general kernel, 161 k retries 1.1 ms against ~0.4 ms)

// ===================================================================================== general path
// Per-lane nested loops over rows and candidates, restarts with a growing radius, plain square rings
// of rows beyond the table: exact for any input.  MODE 0: one thread per query of the range (used when
// the fast path does not apply); MODE 1 (RETRY): the queries listed by the fast kernel; MODE 2
// (EXTERNAL): query points that are not members of the cloud (P.ext_queries), nothing excluded.
template <int K_T, int QB, int MODE>
__global__ void __launch_bounds__(QB) kn_query_general_kernel(const __grid_constant__ QueryParams P) {
    constexpr bool RETRY = MODE == 1, EXTERNAL = MODE == 2;
    extern __shared__ unsigned long long smem[];  // [K][QB]: slot-major, thread-minor
    const int K = K_T ? K_T : P.k;
    const int tid = threadIdx.x;
    const int Dx = P.g.Dx, Dy = P.g.Dy, zb = P.g.zb, ze = P.g.ze;
    const float invw = P.g.invwx, wmin = fminf(P.g.wy, P.g.wz);
    unsigned long long *L = smem + tid;
    // which list: the second-chance pass of the select kernel, if it ran, left its own
    const bool second = RETRY && P.retry_radius_sq > 0.f && *P.retry_count > P.retry_short;
    const int *list = second ? P.retry_list2 : P.retry_list;
    const long long total = RETRY ? (long long)(second ? *P.retry_count2 : *P.retry_count) : (long long)P.query_count;
    const long long stride = (long long)gridDim.x * QB;
    if (RETRY && total <= (long long)P.retry_short) return;  // a short list: the warp-per-query kernel took it
    if (RETRY && blockIdx.x == 0 && tid == 0 && total && !second) atomicAdd(P.diag + 0, (unsigned)total);  // counted as restarts
    const float retry_r2 = second ? P.retry_radius_sq : P.radius0_sq;  // the sphere these queries were last searched in

    for (long long row0 = (long long)blockIdx.x * QB; row0 < total; row0 += stride) {
        const bool valid = row0 + tid < total;
        const long long slot = min(row0 + tid, total - 1);
        const int entry = RETRY ? list[slot] : P.query_begin + (int)slot;
        const int q = EXTERNAL ? -1 : (entry & 0x7FFFFFFF);  // -1: no candidate position equals it, nothing is excluded
        const QueryGeom s = EXTERNAL ? query_geom(make_float4(__ldg(P.ext_queries + 3 * slot), __ldg(P.ext_queries + 3 * slot + 1),
                                                              __ldg(P.ext_queries + 3 * slot + 2), 0.f), P.g)
                                     : query_geom(__ldg(P.pts + q), P.g);

        int cnt = 0;
        unsigned long long kth = ~0ull;
        float r2 = RETRY ? (entry < 0 ? retry_r2 * P.retry_grow : retry_r2) : P.radius0_sq;
        float bound2 = r2;
        bool pruned = false;
        unsigned restarts = 0;

        auto scan_row = [&](int dj, int dk) {
            const int j = s.cj + dj, k = s.ck + dk;
            if (j < 0 || j >= Dy || k < zb || k >= ze) return;
            const float lbyz2 = row_lower_bound2(s, dj, dk, P.g);
            if (lbyz2 > bound2) { pruned = true; return; }
            int ilo, ihi;
            row_x_range(s, bound2, lbyz2, invw, Dx, ilo, ihi);
            if (ilo > 0 || ihi < Dx - 1) pruned = true;
            if (ilo > ihi) return;
            const int row = ((k - zb) * Dy + j) * Dx;
            int p = __ldg(P.ptrs + row + ilo);
            const int end = __ldg(P.ptrs + row + ihi + 1);
            if (p >= end) return;
            float4 c = __ldg(P.pts + p);
            while (true) {
                unsigned long long key;
                bool found;
                while (true) {  // run to the next candidate that beats the current k-th
                    const float4 nx = __ldg(P.pts + min(p + 1, end - 1));
                    key = make_key(c, s, p);
                    found = key < kth && p != q;  // self excluded by position, knearests.cu:123
                    c = nx;
                    ++p;
                    if (found || p >= end) break;
                }
                if (found) {
                    list_insert<QB>(L, K, cnt, key);
                    if (cnt == K) {
                        kth = L[(K - 1) * QB];
                        bound2 = fminf(r2, __uint_as_float((unsigned)(kth >> 32)));
                    }
                }
                if (p >= end) break;
            }
        };

        bool done = !valid;
        while (!done) {
            cnt = 0; kth = ~0ull; bound2 = r2; pruned = false;
            for (int e = 0; e < P.tab.n; ++e) {  // rows nearest-first
                if (P.tab.lb2[e] > bound2) { pruned = true; break; }
                scan_row(P.tab.dj[e], P.tab.dk[e]);
            }
            // Beyond the table: plain square rings of rows.  NOT conditional on having walked the whole table:
            // the table ends with its corner rows (lower bound 4 sqrt(2) cells), which lie farther than the face
            // rows of the next ring (5 cells), so leaving the table early says nothing about ring 6.
            {
                const int rmax = max(max(s.cj, Dy - 1 - s.cj), max(s.ck - zb, ze - 1 - s.ck));
                for (int r = kTableRing + 1; r <= rmax; ++r) {
                    const float lb = fmaxf(0.f, (float)(r - 1) * wmin - kMargin);
                    if (lb * lb * kShrink > bound2) { pruned = true; break; }
                    for (int dk = -r; dk <= r; ++dk)
                        for (int dj = -r; dj <= r; dj += (abs(dk) == r ? 1 : 2 * r)) scan_row(dj, dk);
                }
            }
            const float kd2 = __uint_as_float((unsigned)(kth >> 32));
            if ((cnt == K && kd2 <= r2) || !pruned) {
                done = true;  // everything within the k-th distance was searched, or the whole grid was
            } else {
                r2 = (cnt == K) ? kd2 : r2 * (RETRY ? P.retry_grow : 2.25f);  // the k-th found is a proven sufficient radius
                ++restarts;
            }
        }
        report_diag(P, s, valid, (cnt == K) ? __uint_as_float((unsigned)(kth >> 32)) : FLT_MAX, restarts, tid);

        for (int t = cnt; t < K; ++t) L[t * QB] = ~0ull;  // low word = UINT_MAX = "no such neighbour"
        if (RETRY) {  // scattered rows: each thread writes its own
            if (valid) {
                unsigned int *out = P.knn + (long long)(q - P.query_begin) * K;
                for (int t = 0; t < K; ++t) {
                    const unsigned v = (unsigned)L[t * QB];
                    out[t] = v == KNC_EMPTY ? v : v + P.index_offset;
                }
            }
        } else {      // contiguous rows: one coalesced run per block
            __syncthreads();
            const int rows = (int)min((long long)QB, total - row0);
            unsigned int *out = P.knn + row0 * K;
            for (int idx = tid; idx < rows * K; idx += QB) {
                const int t = idx / K, sl = idx - t * K;
                const unsigned v = (unsigned)smem[sl * QB + t];
                out[idx] = v == KNC_EMPTY ? v : v + P.index_offset;
            }
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------ retry path, few queries
// The same search as MODE 1 above for a SHORT retry list, one WARP per query.  The general kernel gives every query a
// lane, and lanes of a warp that walk different rows and restart at different times run one after the other: fine
// when a million retries keep the machine full (100 M points), ruinous for a small cloud, whose few thousand
// retries -- a third of the reference's 20 k-point fixture lies at the domain's faces -- leave most of the GPU idle
// while each warp serialises 32 searches (0.46 ms of the fixture's 0.49 ms).  Here the 32 lanes read 32 candidates of
// a row at once and every candidate inside the sphere goes, by ballot, into the warp's buffer in shared memory; when
// the sphere has been searched the warp sorts the buffer (bitonic, in place) and the first k keys are the row.  Too
// few: a larger sphere; more than the buffer holds: a smaller one, and if that does not help (a heap of coincident
// points) the warp falls back to keeping a sorted list of k by insertion, which is exact for anything.
// The list counts as short when it holds at most kRetryWarpFactor queries per resident warp.
__device__ __forceinline__ int retry_warp_cap(int k) {  // buffer entries per warp: a power of two, >= 8 k and >= 256
    int c = 256;
    while (c < 8 * k) c <<= 1;
    return c;
}

template <int K_T, int QB>
__global__ void __launch_bounds__(QB) kn_query_retry_warp_kernel(const __grid_constant__ QueryParams P) {
    extern __shared__ unsigned long long smem[];  // [QB / 32][cap]: one buffer per warp
    const int K = K_T ? K_T : P.k;
    const int cap = retry_warp_cap(K);
    const int tid = threadIdx.x, lane = tid & 31;
    const int Dx = P.g.Dx, Dy = P.g.Dy, zb = P.g.zb, ze = P.g.ze;
    const float invw = P.g.invwx, wmin = fminf(P.g.wy, P.g.wz);
    unsigned long long *B = smem + (size_t)(tid >> 5) * cap;
    const bool second = P.retry_radius_sq > 0.f && *P.retry_count > P.retry_short;  // see the general kernel
    const int *list = second ? P.retry_list2 : P.retry_list;
    const long long total = (long long)(second ? *P.retry_count2 : *P.retry_count);
    const long long warps = (long long)gridDim.x * (QB / 32);
    if (total > (long long)P.retry_short) return;  // a long list: the general kernel's turn
    if (blockIdx.x == 0 && tid == 0 && total && !second) atomicAdd(P.diag + 0, (unsigned)total);  // counted as restarts
    const float retry_r2 = second ? P.retry_radius_sq : P.radius0_sq;
    for (long long slot = (long long)blockIdx.x * (QB / 32) + (tid >> 5); slot < total; slot += warps) {
        const int entry = list[slot];
        const int q = entry & 0x7FFFFFFF;
        const QueryGeom s = query_geom(__ldg(P.pts + q), P.g);
        float r2 = entry < 0 ? retry_r2 * 2.25f : retry_r2;
        unsigned restarts = 0;
        int cnt = 0;        // keys held (collect mode: inside the sphere; insertion mode: the sorted list)
        bool pruned = false;
        bool collect = true;
        unsigned long long kth = ~0ull;
        float bound2 = r2;

        // insertion mode: sorted insertion into B[0..K) by the whole warp (every lane looks at the entries lane,
        // lane + 32, ..; the number of entries below the new key is its place; the entries from there on move up by one)
        auto insert = [&](unsigned long long cand) {
            int pos = 0;
            for (int i0 = 0; i0 < cnt; i0 += 32) {
                const int i = i0 + lane;
                const unsigned long long e = i < cnt ? B[i] : ~0ull;
                pos += __popc(__ballot_sync(0xffffffffu, e < cand));
            }
            const int last = min(cnt, K - 1);  // highest index written
            for (int i0 = ((last - pos) / 32) * 32; i0 >= 0; i0 -= 32) {  // from the top down, a chunk of 32 at a time
                const int i = pos + i0 + lane;
                const bool mv = i < last;
                const unsigned long long e = mv ? B[i] : 0ull;
                __syncwarp();
                if (mv) B[i + 1] = e;
                __syncwarp();
            }
            if (lane == 0) B[pos] = cand;
            __syncwarp();
            cnt = min(cnt + 1, K);
        };

        auto scan_row = [&](int dj, int dk) {  // warp-uniform
            const int j = s.cj + dj, k = s.ck + dk;
            if (j < 0 || j >= Dy || k < zb || k >= ze) return;
            const float lbyz2 = row_lower_bound2(s, dj, dk, P.g);
            if (lbyz2 > bound2) { pruned = true; return; }
            int ilo, ihi;
            row_x_range(s, bound2, lbyz2, invw, Dx, ilo, ihi);
            if (ilo > 0 || ihi < Dx - 1) pruned = true;
            if (ilo > ihi) return;
            const int row = ((k - zb) * Dy + j) * Dx;
            const int p0 = __ldg(P.ptrs + row + ilo), end = __ldg(P.ptrs + row + ihi + 1);
            for (int base = p0; base < end; base += 32) {
                const int p = base + lane;
                unsigned long long key = ~0ull;
                if (p < end && p != q) key = make_key(__ldg(P.pts + p), s, p);  // self excluded by position, knearests.cu:123
                if (collect) {
                    const bool in = (unsigned)(key >> 32) <= __float_as_uint(r2);  // d2 <= r2 (non-negative floats order like their bits)
                    const unsigned m = __ballot_sync(0xffffffffu, in);
                    const int at = cnt + __popc(m & ((1u << lane) - 1u));
                    if (in && at < cap) B[at] = key;
                    cnt += __popc(m);
                } else {
                    unsigned m = __ballot_sync(0xffffffffu, key < kth);
                    while (m) {
                        const int src = __ffs(m) - 1;
                        m &= m - 1;
                        const unsigned long long cand = __shfl_sync(0xffffffffu, key, src);
                        if (cand < kth) {  // the k-th may have tightened since the ballot
                            insert(cand);
                            if (cnt == K) {
                                kth = B[K - 1];
                                bound2 = fminf(r2, __uint_as_float((unsigned)(kth >> 32)));
                            }
                        }
                    }
                }
            }
        };

        int shrinks = 0;
        while (true) {
            cnt = 0; kth = ~0ull; bound2 = r2; pruned = false;
            for (int e = 0; e < P.tab.n; ++e) {  // rows nearest-first
                if (P.tab.lb2[e] > bound2) { pruned = true; break; }
                scan_row(P.tab.dj[e], P.tab.dk[e]);
            }
            {   // beyond the table: plain square rings of rows (see the general kernel)
                const int rmax = max(max(s.cj, Dy - 1 - s.cj), max(s.ck - zb, ze - 1 - s.ck));
                for (int r = kTableRing + 1; r <= rmax; ++r) {
                    const float lb = fmaxf(0.f, (float)(r - 1) * wmin - kMargin);
                    if (lb * lb * kShrink > bound2) { pruned = true; break; }
                    for (int dk = -r; dk <= r; ++dk)
                        for (int dj = -r; dj <= r; dj += (abs(dk) == r ? 1 : 2 * r)) scan_row(dj, dk);
                }
            }
            __syncwarp();
            if (collect) {
                if (cnt > cap) {                       // more than the buffer holds: a smaller sphere, or -- a heap of
                    if (++shrinks > 3) collect = false;  // coincident points -- the insertion mode, exact for anything
                    else r2 *= 0.5f;
                    continue;
                }
                if (cnt >= K) break;                   // the sphere holds k points
                if (!pruned && r2 == FLT_MAX) break;   // the whole grid was searched and every point taken: fewer than k exist
                r2 = pruned ? r2 * 2.25f : FLT_MAX;    // every row was walked in full, but points beyond r2 were left out
                ++restarts;
            } else {
                const float kd2 = __uint_as_float((unsigned)(kth >> 32));
                if ((cnt == K && kd2 <= r2) || !pruned) break;
                r2 = (cnt == K) ? kd2 : r2 * 2.25f;
                ++restarts;
            }
        }
        if (collect) {  // sort what the sphere held: bitonic network over the next power of two, ~0 padding sorts last
            int n2 = 32;
            while (n2 < cnt) n2 <<= 1;
            for (int i = cnt + lane; i < n2; i += 32) B[i] = ~0ull;
            __syncwarp();
            for (int k2 = 2; k2 <= n2; k2 <<= 1)
                for (int j = k2 >> 1; j > 0; j >>= 1) {
                    for (int t = lane; t < n2 / 2; t += 32) {
                        const int i = 2 * t - (t & (j - 1)), partner = i + j;
                        const unsigned long long x = B[i], y = B[partner];
                        if ((x > y) == ((i & k2) == 0)) { B[i] = y; B[partner] = x; }
                    }
                    __syncwarp();
                }
            cnt = min(cnt, K);
            kth = cnt ? B[cnt - 1] : ~0ull;
        }
        report_diag(P, s, lane == 0, (cnt == K) ? __uint_as_float((unsigned)(kth >> 32)) : FLT_MAX, lane == 0 ? restarts : 0u, tid);
        __syncwarp();
        unsigned int *out = P.knn + (long long)(q - P.query_begin) * K;
        for (int t = lane; t < K; t += 32) {
            const unsigned v = t < cnt ? (unsigned)B[t] : KNC_EMPTY;
            out[t] = v == KNC_EMPTY ? v : v + P.index_offset;
        }
        __syncwarp();
    }
}

template <int K_T, int QB>
cudaError_t launch_retry_warp(const QueryParams &P, unsigned blocks, cudaStream_t st) {
    int cap = 256;
    while (cap < 8 * P.k) cap <<= 1;  // = retry_warp_cap
    const size_t bytes = sizeof(unsigned long long) * (size_t)cap * (QB / 32);
    cudaError_t e = cudaFuncSetAttribute(kn_query_retry_warp_kernel<K_T, QB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return e;
    kn_query_retry_warp_kernel<K_T, QB><<<blocks, QB, bytes, st>>>(P);
    return cudaGetLastError();
}

// Second chance (select kernel, MODE 2): the sphere is `second_chance_volume` times the first one and the buffer holds
// its expected count (k + 2.5 sqrt(k) + 2 times that) + 3 sigma + the chunk of overrun room.
constexpr float second_chance_volume(int k) { return k <= 32 ? 1.6f : 1.3f; }
constexpr int second_chance_cap(int k) { return k <= 8 ? 48 : k <= 20 ? 80 : k <= 32 ? 108 : 116; }

template <int K_T, int QB>
cudaError_t launch(const QueryParams &P, bool fast, bool tile_ok, cudaStream_t st) {
    const size_t list_bytes = sizeof(unsigned long long) * (size_t)P.k * QB;
    const unsigned blocks = (unsigned)(((long long)P.query_count + QB - 1) / QB);
    cudaError_t e;
    if (P.ext_queries) {
        e = cudaFuncSetAttribute(kn_query_general_kernel<K_T, QB, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)list_bytes);
        if (e != cudaSuccess) return e;
        kn_query_general_kernel<K_T, QB, 2><<<blocks, QB, list_bytes, st>>>(P);
        return cudaGetLastError();
    }
    constexpr bool kHasSelect = K_T == 8 || K_T == 20 || K_T == 32 || K_T == 50;
    if (fast && kHasSelect && P.use_select) {
        if constexpr (kHasSelect) {
            constexpr int NET = K_T <= 20 ? 32 : 64, CAP = K_T <= 20 ? 56 : K_T <= 32 ? 92 : 100;  // 56 and 92: the most that still lets 6 resp. 4 blocks share an SM (228 KB)
            const size_t smem = (sizeof(unsigned) * CAP + sizeof(int2) * kSpanSlots) * QB;
            if (tile_ok) {
                // the tile kernel's buffer for k <= 20: 60 rows (16-bit entries in a column shared with the span list)
                // k = 32: 84 rows (49 points expected in the sphere, + 4.5 sigma) and a tile of 1880 records let a FOURTH block
                // share an SM (61 KB -> 56.7 KB per block; 127 registers allow four)
                constexpr int TCAPB = K_T <= 8 ? 40 : K_T <= 20 ? 60 : K_T <= 32 ? 84 : CAP;
                // Tile capacity in records: 9 rows x (the x-extent of QB queries + 2 x reach) at ~1 point per cell is
                // ~1340 at k = 20, + head and padding (~110), + 4 sigma: the x-extent of 128 consecutive queries varies
                // by sqrt(128) cells and all 9 rows follow it, sigma ~ 107 records (profiles/r2_notes.md).
                constexpr int TCAP = K_T <= 8 ? 1672 : K_T <= 20 ? 1760 : K_T <= 32 ? 1880 : 2048;  // 1760 (mean + 3.5 sigma) with 60 buffer rows: 5 blocks share an SM at k <= 20
                const size_t tsmem = (size_t)TCAP * 16 + (size_t)(QB / 32) * ((size_t)(TCAPB + (NET == 32 ? 2 : 20)) * 64);
                e = cudaFuncSetAttribute(kn_query_tile_kernel<K_T, QB, NET, TCAPB, TCAP>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem);
                if (e != cudaSuccess) return e;
                kn_query_tile_kernel<K_T, QB, NET, TCAPB, TCAP><<<blocks, QB, tsmem, st>>>(P);
                e = cudaFuncSetAttribute(kn_query_select_kernel<K_T, QB, NET, CAP, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                const unsigned list_blocks = blocks < 148u * 4u ? blocks : 148u * 4u;  // loops over the block list
                kn_query_select_kernel<K_T, QB, NET, CAP, 1><<<list_blocks, QB, smem, st>>>(P);
            } else {
                e = cudaFuncSetAttribute(kn_query_select_kernel<K_T, QB, NET, CAP, 0>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                kn_query_select_kernel<K_T, QB, NET, CAP, 0><<<blocks, QB, smem, st>>>(P);
            }
            if (P.retry_radius_sq > 0.f) {  // second chance of a long retry list, see the kernel (MODE 2)
                constexpr int NET2 = K_T <= 8 ? 32 : 64, CAP2 = second_chance_cap(K_T);  // (k = 20 with 32 keys a round: a fourth block per SM, yet 1.65 ms instead of 1.38 at 100 M points)
                const size_t smem2 = (sizeof(unsigned) * CAP2 + sizeof(int2) * kSpanSlots) * QB;
                e = cudaFuncSetAttribute(kn_query_select_kernel<K_T, QB, NET2, CAP2, 2>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
                if (e != cudaSuccess) return e;
                const unsigned want = 148u * (unsigned)(227 * 1024 / (smem2 + 1024));  // loops over the list
                kn_query_select_kernel<K_T, QB, NET2, CAP2, 2><<<blocks < want ? blocks : want, QB, smem2, st>>>(P);
            }
        }
        e = cudaFuncSetAttribute(kn_query_general_kernel<K_T, QB, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)list_bytes);
        if (e != cudaSuccess) return e;
        const unsigned retry_blocks = 148u * 8u;  // both retry kernels loop over the list (which of the two: P.retry_short)
        kn_query_general_kernel<K_T, QB, 1><<<retry_blocks, QB, list_bytes, st>>>(P);
        e = launch_retry_warp<K_T, QB>(P, retry_blocks, st);
        if (e != cudaSuccess) return e;
    } else if (fast) {
        const size_t smem = list_bytes + sizeof(unsigned long long) * (size_t)(kSpanSlots + kQueueSlots) * QB;
        e = cudaFuncSetAttribute(kn_query_fast_kernel<K_T, QB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        kn_query_fast_kernel<K_T, QB><<<blocks, QB, smem, st>>>(P);
        e = cudaFuncSetAttribute(kn_query_general_kernel<K_T, QB, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)list_bytes);
        if (e != cudaSuccess) return e;
        const unsigned retry_blocks = 148u * 8u;  // fixed: both retry kernels loop over the list
        kn_query_general_kernel<K_T, QB, 1><<<retry_blocks, QB, list_bytes, st>>>(P);
        e = launch_retry_warp<K_T, QB>(P, retry_blocks, st);
        if (e != cudaSuccess) return e;
    } else {
        e = cudaFuncSetAttribute(kn_query_general_kernel<K_T, QB, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)list_bytes);
        if (e != cudaSuccess) return e;
        kn_query_general_kernel<K_T, QB, 0><<<blocks, QB, list_bytes, st>>>(P);
    }
    return cudaGetLastError();
}

void fill_table(RowTable &t, float wy, float wz) {
    struct Ent { int dj, dk; float lbc2, ly, lz; };
    Ent ent[kTableSize];
    int n = 0;
    for (int dk = -kTableRing; dk <= kTableRing; dk++)
        for (int dj = -kTableRing; dj <= kTableRing; dj++) {
            const int a = abs(dj) > 0 ? abs(dj) - 1 : 0, b = abs(dk) > 0 ? abs(dk) - 1 : 0;
            const float ly = fmaxf(0.f, (float)a * wy - kMargin), lz = fmaxf(0.f, (float)b * wz - kMargin);
            ent[n++] = {dj, dk, ly * ly + lz * lz, ly, lz};
        }
    // stable insertion sort by (lower bound, Manhattan distance): nearest rows first
    for (int i = 1; i < n; i++) {
        Ent e = ent[i];
        int s = i;
        auto less = [](const Ent &x, const Ent &y) {
            if (x.lbc2 != y.lbc2) return x.lbc2 < y.lbc2;
            return abs(x.dj) + abs(x.dk) < abs(y.dj) + abs(y.dk);
        };
        while (s > 0 && less(e, ent[s - 1])) { ent[s] = ent[s - 1]; s--; }
        ent[s] = e;
    }
    t.n = n;
    for (int i = 0; i < n; i++) {
        t.dj[i] = (signed char)ent[i].dj;
        t.dk[i] = (signed char)ent[i].dk;
        t.lb2[i] = ent[i].lbc2 * kShrink;  // same form as row_lower_bound2 at the nearest corner of the own cell
    }
}

}  // namespace

cudaError_t kn_launch_query(const float4 *d_points4, const int *d_ptrs, const GridGeom &g, int n, int k,
                            int query_begin, int query_count, float radius0_sq, unsigned int *d_knn,
                            unsigned int *d_diag, int *d_retry_list, int force_general, unsigned int index_offset,
                            const float *d_ext_queries, cudaStream_t st) {
    if (query_count <= 0) return cudaSuccess;
    QueryParams P;
    P.pts = d_points4; P.ptrs = d_ptrs; P.knn = d_knn; P.diag = d_diag; P.g = g;
    P.n = n; P.k = k; P.query_begin = query_begin; P.query_count = query_count;
    P.radius0_sq = radius0_sq;
    P.retry_list = d_retry_list;
    P.index_offset = index_offset;
    P.ext_queries = d_ext_queries;
    P.use_select = force_general != 2;  // algorithm 2 = insertion kernel even for small k (A/B, tests)
    P.retry_count = d_diag + 3;
    P.block_list = d_retry_list ? d_retry_list + query_count : nullptr;
    P.block_count = d_diag + 4;
    P.points_per_cell = (float)((double)n / ((double)g.Dx * g.Dy * (g.ze - g.zb)));
    P.retry_list2 = d_retry_list ? d_retry_list + query_count + (query_count + 31) / 32 + 1 : nullptr;
    P.retry_count2 = d_diag + 5;
    P.retry_short = force_general == 6 ? 0u : 148u * 8u * 4u * (unsigned)kRetryWarpFactor;  // algorithm 6: every list counts as long (tests)
    fill_table(P.tab, g.wy, g.wz);
    const float wmin = fminf(g.wy, g.wz), r0 = sqrtf(radius0_sq);
    // The fast kernels search table rows only: usable when every row outside the table is provably
    // farther than the first-guess radius.
    const float beyond = fmaxf(0.f, (float)kTableRing * wmin - kMargin);
    const bool fast = force_general != 1 && d_retry_list && radius0_sq < beyond * beyond * kShrink;
    // The tile kernel stages the 3 x 3 rows around a query's own: usable when rows two cells away are provably out
    // of reach (their lower bound is one cell width less the margin), i.e. on grids like knc_fast_dims'.
    // algorithm 4 = without it (A/B, tests).
    const float reach = fmaxf(0.f, wmin - kMargin);
    const bool tile_ok = force_general != 4 && radius0_sq < reach * reach * kShrink;
    P.halo_cells = (int)ceilf(r0 * kGrow * g.invwx + kMargin * g.invwx + 1e-3f) + 1;  // >= row_x_range's reach, in cells
    P.bulk_rows_ok = ((uintptr_t)d_knn & 15u) == 0;
    P.retry_grow = 2.25f;
    {   // The second chance exists where the select kernel does (launch<> below) and searches table rows only, like the
        // first pass: its sphere must stay inside the table too.  algorithm 5 = without it (A/B, tests).
        const bool has_select = k == 8 || k == 20 || k == 32 || k == 50;
        const float r1sq = radius0_sq * powf(second_chance_volume(k), 2.f / 3.f);
        P.retry_radius_sq = (fast && has_select && P.use_select && force_general != 5 && r1sq < beyond * beyond * kShrink) ? r1sq : 0.f;
        P.retry_radius_max_sq = beyond * beyond * kShrink * 0.999f;
    }
    switch (k) {
        case 8: return launch<8, 128>(P, fast, tile_ok, st);
        case 20: return launch<20, 128>(P, fast, tile_ok, st);
        case 32: return launch<32, 128>(P, fast, tile_ok, st);
        case 50: return launch<50, 128>(P, fast, tile_ok, st);
        default: break;
    }
    if (k <= 96) return launch<0, 128>(P, fast, false, st);
    if (k <= 192) return launch<0, 64>(P, fast, false, st);
    return launch<0, 32>(P, fast, false, st);
}

size_t kn_retry_ints(int query_count) {
    const size_t q = query_count > 0 ? (size_t)query_count : 1;
    // retried queries + blocks handed from the tile kernel to the select kernel + queries retried a second time
    return q + ((size_t)(query_count + 31) / 32 + 1) + q;
}

// d2 of every stored row entry, with the reference's expression (knearests.cu:125 as compiled) -- the reference
// never returns distances; this recomputes them from the rows.
namespace {
__global__ void kn_rows_dists_kernel(const float4 *__restrict__ pts, const unsigned *__restrict__ rows,
                                     const float *__restrict__ ext_queries, int query_begin, unsigned index_offset,
                                     long long total, int k, float *__restrict__ out) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const long long r = e / k;
    const unsigned v = __ldg(rows + e);
    float d2 = FLT_MAX;
    if (v != KNC_EMPTY) {
        float qx, qy, qz;
        if (ext_queries) { qx = __ldg(ext_queries + 3 * r); qy = __ldg(ext_queries + 3 * r + 1); qz = __ldg(ext_queries + 3 * r + 2); }
        else { const float4 q = __ldg(pts + query_begin + r); qx = q.x; qy = q.y; qz = q.z; }
        const float4 c = __ldg(pts + (v - index_offset));
        const float dx = c.x - qx, dy = c.y - qy, dz = c.z - qz;
        d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    }
    out[e] = d2;
}
}  // namespace

cudaError_t kn_launch_rows_dists(const float4 *d_points4, const unsigned int *d_rows, const float *d_ext_queries,
                                 int query_begin, unsigned int index_offset, long long rows, int k, float *d_out,
                                 cudaStream_t st) {
    const long long total = rows * k;
    if (total > 0)
        kn_rows_dists_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_points4, d_rows, d_ext_queries, query_begin,
                                                                             index_offset, total, k, d_out);
    return cudaGetLastError();
}
