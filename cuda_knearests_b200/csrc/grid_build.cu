// grid_build.cu -- uniform-grid construction for sm_100a: per-cell count, exclusive scan, counting-sort scatter of
// the records, per-cell ordering.  Replaces the reference's count / reserve / store kernels (knearests.cu:32-60) and
// their launcher kn_firstbuild (knearests.cu:152-201).
//
// What is different from the reference, and why:
//  * count   : one returnless atomic per point (warp-aggregated with match.any when the input arrives spatially
//              coherent); nothing else is written -- no cell id, no rank.
//  * reserve : a real exclusive prefix sum (single-pass, decoupled look-back) instead of atomicAdd(globalcounter, cnt)
//              in arrival order (:45) => cells are laid out in ascending cell id, deterministic, and an x-row of cells
//              is one contiguous span of points.  The scan writes the start of cell c to ptrs[c + 1]:
//  * store   : the scatter takes its slot with atomicAdd(ptrs + cell + 1, 1) -- the cursor of cell c ends at the
//              start of cell c + 1, so after the scatter ptrs IS the pointer array, without a second counter array --
//              and writes the whole 16-byte record (x, y, z, original index) there, not an index to chase later.
//  * order   : slots come from atomics, so the order inside a cell is arrival order; the last pass reads the records
//              once, coalesced, puts every cell's few records into ascending original index (each thread looks at
//              its neighbours of the same cell in shared memory) and writes the three arrays the API returns =>
//              the layout is the stable counting sort, bit-identical from run to run (the reference's is not,
//              SURVEY F6).  Nothing before that pass has to be stable or deterministic.
//  * the sorted points are written twice: float3 (the API's d_stored_points) and float4
//              (x,y,z,bits(own sorted position)) -- 16-byte records the query kernels load with one 128-bit
//              load or copy into shared memory in bulk; the 4th word lets a staged copy name its source.
//
// HBM traffic per point at one point per cell: count 16 r + 8 (counters cleared, written back); scan 4 r + 4 w; scatter
// 16 r + 8 (cursors) + 16 w; order 16 r + 32 w (float4, float3, permutation); z-band split (large clouds) 12 r, then
// 12 r + 16 w ~= 160 B/point, 44 of them algorithmic (12 r + 32 w).
#include <cuda_runtime.h>
#include <stdint.h>

#include "kn_internal.h"

namespace {

constexpr int kBuildThreads = 256;

// cellFromPoint, one axis -- knearests.cu:23-28.  Must round exactly like the reference's PTX
// (mul.f32, div.rn.f32, cvt.rmi, cvt.rzi): do NOT replace the division by a reciprocal multiply.
__device__ __forceinline__ int axis_cell(float p, float Df, int lo, int hi) {
    float t = floorf(__fdiv_rn(__fmul_rn(p, Df), 1000.f));
    int c = (int)t;  // saturating, NaN -> 0
    return max(lo, min(c, hi));
}

__device__ __forceinline__ int cell_of(float x, float y, float z, const GridGeom &g) {
    const int i = axis_cell(x, g.Dfx, 0, g.Dx - 1);
    const int j = axis_cell(y, g.Dfy, 0, g.Dy - 1);
    const int k = axis_cell(z, g.Dfz, g.zb, g.ze - 1) - g.zb;
    return i + (j + k * g.Dy) * g.Dx;  // z-major, knearests.cu:29
}

// one input record: a single 16-byte load for aligned 4-float records, else scalars (4th word = 0 for xyz input)
__device__ __forceinline__ float4 load_record(const float *rec, int stride, bool vec4) {
    if (vec4) return __ldg(reinterpret_cast<const float4 *>(rec));
    return make_float4(__ldg(rec), __ldg(rec + 1), __ldg(rec + 2), stride == 4 ? __ldg(rec + 3) : 0.f);
}

// The kPointItems points of a thread: point r of thread t is input point  first + r * kBuildThreads + t  (4th word: the
// record's, 0 for xyz input).  All loads are issued before anything is used.  Packed xyz input goes through shared
// memory with float4 loads, so the global reads are 16-byte vectorised and fully coalesced although the input is
// AoS.  All threads of the block must call it.
// (One point per thread: with 2 or 4 the scatter ran slower -- 1.25 -> 1.4 / 1.55 ms for 100 M points, the count the same.)
constexpr int kPointItems = 1;
constexpr int kPointTile = kBuildThreads * kPointItems;

__device__ __forceinline__ void block_points(const float *__restrict__ xyz, int stride, long long first, int here,
                                             int aligned16, float *tile, float4 (&v)[kPointItems]) {
    const int tid = threadIdx.x;
    if (stride == 4) {
#pragma unroll
        for (int r = 0; r < kPointItems; ++r) {
            const int at = r * kBuildThreads + tid;
            v[r] = at < here ? load_record(xyz + (first + at) * 4, 4, aligned16) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    const float *src = xyz + first * 3;
    if (aligned16 && here == kBuildThreads * kPointItems) {
#pragma unroll
        for (int r = 0; r < (3 * kPointItems + 3) / 4; ++r)
            if (r * kBuildThreads + tid < kBuildThreads * kPointItems * 3 / 4)
                reinterpret_cast<float4 *>(tile)[r * kBuildThreads + tid] = __ldg(reinterpret_cast<const float4 *>(src) + r * kBuildThreads + tid);
    } else {
        for (int i = tid; i < here * 3; i += kBuildThreads) tile[i] = __ldg(src + i);
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kPointItems; ++r) {
        const int at = r * kBuildThreads + tid;
        v[r] = at < here ? make_float4(tile[3 * at], tile[3 * at + 1], tile[3 * at + 2], 0.f) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// ---------------------------------------------------------------------------------------- count
// One returnless atomic per point, kPointItems points per thread.  Lanes of a warp in the same cell elect a leader
// that adds the group size once: match.any costs one round per distinct value, so it is only used when the warp looks
// coherent (some lane shares its cell with its neighbour: sorted or clustered input); a warp of a randomly ordered
// cloud -- 32 different cells -- goes straight to 32 plain atomics.
__global__ void __launch_bounds__(kBuildThreads)
kn_count_kernel(const float *__restrict__ xyz, int stride, int n, GridGeom g, int *__restrict__ counters, int aligned16) {
    __shared__ __align__(16) float tile[kPointTile * 3];
    const int tid = threadIdx.x, lane = tid & 31;
    const long long first = (long long)blockIdx.x * kPointTile;
    const int here = (int)min((long long)kPointTile, (long long)n - first);
    float4 v[kPointItems];
    block_points(xyz, stride, first, here, aligned16, tile, v);
#pragma unroll
    for (int r = 0; r < kPointItems; ++r) {
        const bool valid = r * kBuildThreads + tid < here;
        const int cell = valid ? cell_of(v[r].x, v[r].y, v[r].z, g) : -1 - lane;  // invalid lanes: unique keys, never grouped
        const int next = __shfl_down_sync(0xffffffffu, cell, 1);
        const bool coherent = __any_sync(0xffffffffu, lane < 31 && next == cell);
        if (coherent) {
            const unsigned peers = __match_any_sync(0xffffffffu, cell);
            if (valid && lane == __ffs(peers) - 1) atomicAdd(counters + cell, __popc(peers));
        } else if (valid) {
            atomicAdd(counters + cell, 1);
        }
    }
}

// ----------------------------------------------------------------------------------------- scan
// Single-pass exclusive prefix sum with decoupled look-back.  Tile = 256 threads x 32 entries.
// Tile state word: status in the top 2 bits (0 = not ready, 1 = tile aggregate, 2 = inclusive prefix),
// value in the low 32 bits -- one 64-bit store publishes both, so no fence is needed between them.
//
// The scan runs over cells + 1 entries  v[j] = counters[j - 1]  (v[0] = 0) and writes ptrs[j] = sum of v below j:
// ptrs[0] = 0 and ptrs[c + 1] = START of cell c, the cursor the scatter advances to the start of cell c + 1.
constexpr int kScanItems = 32;
constexpr int kScanTile = kBuildThreads * kScanItems;

__global__ void __launch_bounds__(kBuildThreads)
kn_scan_kernel(const int *__restrict__ counters, int *__restrict__ ptrs, long long cells,
               unsigned long long *tile_state, int *ticket, int *globcounter) {
    __shared__ int s_tile;
    __shared__ int s_warp[kBuildThreads / 32];
    __shared__ int s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1);  // tiles are numbered in the order blocks START
    __syncthreads();
    const int tile = s_tile;
    const long long base = (long long)tile * kScanTile + (long long)tid * kScanItems;

    int v[kScanItems];
    if (base + kScanItems <= cells) {  // aligned vector loads of counters[base ..], shifted by one entry
        const int4 *src = reinterpret_cast<const int4 *>(counters + base);
        v[0] = base > 0 ? __ldg(counters + base - 1) : 0;
#pragma unroll
        for (int q = 0; q < kScanItems / 4; q++) {
            int4 t = __ldg(src + q);
            v[4 * q + 1] = t.x; v[4 * q + 2] = t.y; v[4 * q + 3] = t.z;
            if (q < kScanItems / 4 - 1) v[4 * q + 4] = t.w;
        }
    } else {
#pragma unroll
        for (int q = 0; q < kScanItems; q++) v[q] = (base + q >= 1 && base + q <= cells) ? __ldg(counters + base + q - 1) : 0;
    }
    int sum = 0;
#pragma unroll
    for (int q = 0; q < kScanItems; q++) sum += v[q];
    int incl = sum;  // inclusive scan of the per-thread sums across the warp
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int warp_off = 0, tile_sum = 0;
#pragma unroll
    for (int wq = 0; wq < kBuildThreads / 32; wq++) {
        int t = s_warp[wq];
        if (wq < warp) warp_off += t;
        tile_sum += t;
    }
    const int thread_excl = warp_off + incl - sum;

    if (warp == 0) {
        volatile unsigned long long *st = tile_state;
        if (lane == 0)
            st[tile] = ((unsigned long long)(tile == 0 ? 2 : 1) << 62) | (unsigned)tile_sum;
        int exclusive = 0;
        if (tile > 0) {
            int look = tile - 1;  // lane L inspects tile look - L
            while (true) {
                unsigned long long wrd = 0;
                const int idx = look - lane;
                if (idx >= 0) {
                    do { wrd = st[idx]; } while ((wrd >> 62) == 0);
                } else {
                    wrd = 2ull << 62;  // before tile 0: inclusive prefix 0
                }
                const unsigned full = __ballot_sync(0xffffffffu, (wrd >> 62) == 2);
                const int stop = full ? __ffs(full) - 1 : 32;  // nearest tile with an inclusive prefix
                int contrib = (lane <= stop) ? (int)(unsigned)wrd : 0;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
                exclusive += contrib;
                if (full) break;
                look -= 32;
            }
            if (lane == 0) st[tile] = (2ull << 62) | (unsigned)(exclusive + tile_sum);
        }
        if (lane == 0) s_prefix = exclusive;
    }
    __syncthreads();
    int run = s_prefix + thread_excl;
    if (base + kScanItems <= cells) {
        int4 *dst = reinterpret_cast<int4 *>(ptrs + base);
#pragma unroll
        for (int q = 0; q < kScanItems / 4; q++) {
            int4 t;
            t.x = run; run += v[4 * q];
            t.y = run; run += v[4 * q + 1];
            t.z = run; run += v[4 * q + 2];
            t.w = run; run += v[4 * q + 3];
            dst[q] = t;
        }
    } else {
#pragma unroll
        for (int q = 0; q < kScanItems; q++) {
            if (base + q <= cells) ptrs[base + q] = run;
            run += v[q];
        }
    }
    // the tile that covers the last entry also writes the public counter (the number of points)
    if ((long long)(tile + 1) * kScanTile > cells && tid == kBuildThreads - 1) *globcounter = s_prefix + tile_sum;
}

// -------------------------------------------------------------------------------------- scatter
// Every point takes the next free slot of its cell (the cursor in ptrs[cell + 1], see the scan) and its record goes
// there: (x, y, z, original index).  `binned`: the input is the z-banded copy, whose 4th word is the original index
// already and whose caller-side 4th words (tags) sit in bin_tags; otherwise the original index is the input position
// and a tag, if asked for, the record's own 4th word.
__global__ void __launch_bounds__(kBuildThreads)
kn_scatter_kernel(const float *__restrict__ xyz, int stride, int n, GridGeom g, int *__restrict__ ptrs,
                  float4 *__restrict__ recs, int binned, const unsigned int *__restrict__ bin_tags,
                  unsigned int *__restrict__ rec_tags, int aligned16) {
    __shared__ __align__(16) float tile[kPointTile * 3];
    const int tid = threadIdx.x, lane = tid & 31;
    const long long first = (long long)blockIdx.x * kPointTile;
    const int here = (int)min((long long)kPointTile, (long long)n - first);
    float4 v[kPointItems];
    block_points(xyz, stride, first, here, aligned16, tile, v);
    int slot[kPointItems];
#pragma unroll
    for (int r = 0; r < kPointItems; ++r) {  // all the atomics first: their round trips overlap
        const bool valid = r * kBuildThreads + tid < here;
        const int cell = valid ? cell_of(v[r].x, v[r].y, v[r].z, g) : -1 - lane;
        const int next = __shfl_down_sync(0xffffffffu, cell, 1);
        const bool coherent = __any_sync(0xffffffffu, lane < 31 && next == cell);
        if (coherent) {
            const unsigned peers = __match_any_sync(0xffffffffu, cell);
            const int leader = __ffs(peers) - 1;
            int base = 0;
            if (valid && lane == leader) base = atomicAdd(ptrs + cell + 1, __popc(peers));
            slot[r] = __shfl_sync(0xffffffffu, base, leader) + __popc(peers & ((1u << lane) - 1u));
        } else {
            slot[r] = valid ? atomicAdd(ptrs + cell + 1, 1) : 0;
        }
    }
#pragma unroll
    for (int r = 0; r < kPointItems; ++r) {
        const int at = r * kBuildThreads + tid;
        if (at < here) {
            const unsigned idx = binned ? __float_as_uint(v[r].w) : (unsigned)(first + at);
            recs[slot[r]] = make_float4(v[r].x, v[r].y, v[r].z, __uint_as_float(idx));
            if (rec_tags) rec_tags[slot[r]] = binned ? __ldg(bin_tags + first + at) : __float_as_uint(v[r].w);
        }
    }
}

// ---------------------------------------------------------------------------------------- order
// The records are in cell order, arrival order inside a cell.  One thread per record: with the cells and original
// indices of the block's records (+ KN_BIG_CELL on either side) in shared memory it walks left and right over the
// records of its own cell, counting them and those with a smaller original index -- that is its place in the cell.
// It then writes its point to that place in the three output arrays.  A thread that cannot see an end of its cell
// (more than KN_BIG_CELL records that way) writes nothing; the first record of such a cell queues the cell for
// kn_order_big_kernel.  (Threads of a queued cell that do see both ends write the same values the big kernel will.)
constexpr int kHalo = KN_BIG_CELL;
constexpr int kOrderItems = 2;  // records per thread
constexpr int kOrderTile = kBuildThreads * kOrderItems;

__global__ void __launch_bounds__(kBuildThreads)
kn_order_kernel(const float4 *__restrict__ recs, const unsigned int *__restrict__ rec_tags, int n, GridGeom g,
                unsigned int *__restrict__ permutation, float *__restrict__ points3, float4 *__restrict__ points4,
                unsigned int *__restrict__ tags, int *big_count, int *__restrict__ big_list) {
    __shared__ int s_cell[kOrderTile + 2 * kHalo];
    __shared__ unsigned s_idx[kOrderTile + 2 * kHalo];
    __shared__ float s_p3[kOrderTile * 3];
    __shared__ unsigned char s_filled[kOrderTile];
    const int tid = threadIdx.x;
    const long long first = (long long)blockIdx.x * kOrderTile;
    const int here = (int)min((long long)kOrderTile, (long long)n - first);
    float4 r[kOrderItems];
    unsigned tag[kOrderItems];
#pragma unroll
    for (int u = 0; u < kOrderItems; ++u) {
        const int at = u * kBuildThreads + tid;
        r[u] = at < here ? __ldg(recs + first + at) : make_float4(0.f, 0.f, 0.f, 0.f);
        tag[u] = (tags && at < here) ? __ldg(rec_tags + first + at) : 0u;
    }
    float4 h = make_float4(0.f, 0.f, 0.f, 0.f);
    bool hvalid = false;
    if (tid < 2 * kHalo) {
        const long long pos = tid < kHalo ? first - kHalo + tid : first + kOrderTile + (tid - kHalo);
        hvalid = pos >= 0 && pos < n;
        if (hvalid) h = __ldg(recs + pos);
    }
#pragma unroll
    for (int u = 0; u < kOrderItems; ++u) {
        const int at = u * kBuildThreads + tid;
        s_cell[kHalo + at] = at < here ? cell_of(r[u].x, r[u].y, r[u].z, g) : -1;  // beyond the array: matches no cell
        s_idx[kHalo + at] = __float_as_uint(r[u].w);
        s_filled[at] = 0;
    }
    if (tid < 2 * kHalo) {
        const int at = tid < kHalo ? tid : kOrderTile + tid;
        s_cell[at] = hvalid ? cell_of(h.x, h.y, h.z, g) : -1;
        s_idx[at] = __float_as_uint(h.w);
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < kOrderItems; ++u) {
        const int at = u * kBuildThreads + tid;
        if (at < here) {
            const long long s = first + at;
            const int p = kHalo + at, mycell = s_cell[p];
            const unsigned myidx = __float_as_uint(r[u].w);
            int left = 0, right = 0, smaller = 0;
            while (left < kHalo && s_cell[p - 1 - left] == mycell) { smaller += s_idx[p - 1 - left] < myidx; ++left; }
            while (right < kHalo && s_cell[p + 1 + right] == mycell) { smaller += s_idx[p + 1 + right] < myidx; ++right; }
            if (left < kHalo && right < kHalo) {  // both ends of the cell seen
                const long long to = s - left + smaller;
                permutation[to] = myidx;
                points4[to] = make_float4(r[u].x, r[u].y, r[u].z, __uint_as_float((unsigned)to));
                if (tags) tags[to] = tag[u];
                const long long rel = to - first;
                if (rel >= 0 && rel < kOrderTile) {
                    s_p3[3 * rel] = r[u].x; s_p3[3 * rel + 1] = r[u].y; s_p3[3 * rel + 2] = r[u].z;
                    s_filled[rel] = 1;
                } else {  // a cell across the block's border
                    points3[3 * to] = r[u].x; points3[3 * to + 1] = r[u].y; points3[3 * to + 2] = r[u].z;
                }
            } else if (left == 0) {
                big_list[atomicAdd(big_count, 1)] = mycell;
            }
        }
    }
    __syncthreads();
    // AoS float3 output written as one contiguous, coalesced run per block (places filled from elsewhere skipped)
    float *dst = points3 + first * 3;
    for (int q = tid; q < here * 3; q += kBuildThreads)
        if (s_filled[q / 3]) dst[q] = s_p3[q];
}

// One block per oversized cell: bitonic network over the cell's records in place, keyed by the original index, every
// compare-exchange ascending (partner = mirror in the first step of a merge, i ^ j afterwards), so the virtual +inf
// padding beyond `cnt` never moves; then the cell's part of the output arrays.
__global__ void __launch_bounds__(1024)
kn_order_big_kernel(const int *big_count, const int *__restrict__ big_list, const int *__restrict__ ptrs,
                    float4 *__restrict__ recs, unsigned int *__restrict__ rec_tags,
                    unsigned int *__restrict__ permutation, float *__restrict__ points3, float4 *__restrict__ points4,
                    unsigned int *__restrict__ tags) {
    const int nbig = *big_count;
    for (int w = blockIdx.x; w < nbig; w += gridDim.x) {
        const int c = big_list[w];
        const long long b = ptrs[c];
        const unsigned cnt = (unsigned)(ptrs[c + 1] - ptrs[c]);
        float4 *seg = recs + b;
        for (unsigned long long k = 2; (k >> 1) < cnt; k <<= 1) {
            for (unsigned long long j = k >> 1; j > 0; j >>= 1) {
                for (unsigned long long t = threadIdx.x; t < cnt; t += blockDim.x) {
                    const unsigned long long partner = (j == (k >> 1)) ? (t ^ (k - 1)) : (t ^ j);
                    if (partner > t && partner < cnt) {
                        const float4 a = seg[t], d = seg[partner];
                        if (__float_as_uint(a.w) > __float_as_uint(d.w)) {
                            seg[t] = d; seg[partner] = a;
                            if (rec_tags) {
                                const unsigned ta = rec_tags[b + t];
                                rec_tags[b + t] = rec_tags[b + partner];
                                rec_tags[b + partner] = ta;
                            }
                        }
                    }
                }
                __syncthreads();
            }
        }
        for (unsigned long long t = threadIdx.x; t < cnt; t += blockDim.x) {
            const long long to = b + (long long)t;
            const float4 a = seg[t];
            permutation[to] = __float_as_uint(a.w);
            points4[to] = make_float4(a.x, a.y, a.z, __uint_as_float((unsigned)to));
            points3[3 * to] = a.x; points3[3 * to + 1] = a.y; points3[3 * to + 2] = a.z;
            if (tags) tags[to] = rec_tags[to];
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------ bins
// Large clouds arrive in random order, so the histogram atomics and the cursor atomics and record writes of the
// scatter would each touch a random DRAM sector per point (13.4 ms for 100 M points).  One coalesced pre-pass
// removes that: a split of the input into 256 bands of consecutive x-rows of cells, written as 16-byte records
// (x, y, z, input index).  All later random accesses of a band then stay inside ~8 MB (6 MB of records being
// placed, 1.5 MB of cursors).  The scatter fills 32-byte sectors half at a time: with 64 bands (31 MB each, next to
// the records streaming through the same L2) a third of them left the L2 half-filled and came back for the other
// half -- 2.95 GB read and 2.87 GB written where 2.0 + 2.0 are needed, 2.05 ms; with 256 bands 2.02 + 1.96 GB, 1.25 ms.
// The split need not be stable (the order inside a cell is settled from the input indices at the end): a tile of
// kBinTile points ranks its points per band with shared-memory atomics and reserves its run of each band with one
// global atomic per band.
constexpr int kBinThreads = 256;
constexpr int kBinItems = 8;    // points per thread, all loaded before the first barrier (memory-level parallelism)
constexpr int kBinTile = kBinThreads * kBinItems;
constexpr int kMaxBins = 256;

// band of a point: its x-row of cells (plane-major) divided by the rows per band
__device__ __forceinline__ int bin_of(float y, float z, const GridGeom &g, int rows_per_bin) {
    const int j = axis_cell(y, g.Dfy, 0, g.Dy - 1);
    const int k = axis_cell(z, g.Dfz, g.zb, g.ze - 1) - g.zb;
    return (k * g.Dy + j) / rows_per_bin;
}

// Band totals: totals[b] (cleared by the launcher).  A block walks over many tiles and adds its histogram to the
// global one once, at the end: with one tile per block -- 12.5 M atomics on 1 KB of counters -- the kernel took 0.3 or
// 0.5 ms from run to run.  (Spreading the counters and the cursors below over 128-byte lines did not help: the
// scatter ran 0.77 instead of 0.69 ms.)
constexpr int kBinPad = 1;

template <int BINS>
__global__ void __launch_bounds__(kBinThreads)
kn_bin_count_kernel(const float *__restrict__ xyz, int stride, int n, GridGeom g, int rows_per_bin, int *__restrict__ totals) {
    __shared__ int s_hist[kBinThreads / 32][BINS];  // one histogram per warp: the atomics of a warp never meet another's
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (kBinThreads / 32) * BINS; i += kBinThreads) (&s_hist[0][0])[i] = 0;
    __syncthreads();
    for (long long tile = blockIdx.x; tile * kBinTile < n; tile += gridDim.x) {
        const long long base = tile * kBinTile + tid;
        float y[kBinItems], z[kBinItems];
#pragma unroll
        for (int r = 0; r < kBinItems; ++r) {
            const long long i = base + (long long)r * kBinThreads;
            y[r] = i < n ? __ldg(xyz + i * stride + 1) : 0.f;
            z[r] = i < n ? __ldg(xyz + i * stride + 2) : 0.f;
        }
#pragma unroll
        for (int r = 0; r < kBinItems; ++r)
            if (base + (long long)r * kBinThreads < n) atomicAdd(&s_hist[warp][bin_of(y[r], z[r], g, rows_per_bin)], 1);
    }
    __syncthreads();
    for (int b = tid; b < BINS; b += kBinThreads) {
        int c = 0;
#pragma unroll
        for (int w = 0; w < kBinThreads / 32; ++w) c += s_hist[w][b];
        if (c) atomicAdd(totals + b * kBinPad, c);
    }
}

// one block: starts[b] = sum of totals below b
__global__ void __launch_bounds__(kMaxBins) kn_bin_starts_kernel(const int *__restrict__ totals, int bins, int *__restrict__ starts) {
    __shared__ int s_warp[kMaxBins / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int v = tid < bins ? totals[tid * kBinPad] : 0;
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int off = 0;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    if (tid < bins) starts[tid * kBinPad] = off + incl - v;
}

// cursors[BINS]: starts[] on entry (kn_bin_starts_kernel), advanced by every tile
template <int BINS>
__global__ void __launch_bounds__(kBinThreads)
kn_bin_scatter_kernel(const float *__restrict__ xyz, int stride, int n, GridGeom g, int rows_per_bin,
                      int *__restrict__ cursors, float4 *__restrict__ recs,
                      unsigned int *__restrict__ tags, int vec4) {  // tags: 4th input words in banded order (stride 4) or NULL
    static_assert(BINS <= kBinThreads && BINS <= 256, "one thread per band; the band of a staged slot is a byte");
    __shared__ float4 s_stage[kBinTile];        // the tile's records, band by band
    __shared__ unsigned char s_band[kBinTile];  // band of each staged slot
    __shared__ int s_cnt[BINS];                 // the tile's count per band
    __shared__ int s_first[BINS];               // first staged slot of each band
    __shared__ int s_base[BINS];                // where the tile's run of each band starts in recs
    __shared__ int s_warp[kBinThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long base = (long long)blockIdx.x * kBinTile + tid;
    float4 v[kBinItems];
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const long long i = base + (long long)r * kBinThreads;
        v[r] = load_record(xyz + (i < n ? i : 0) * stride, stride, vec4);
    }
    if (tid < BINS) s_cnt[tid] = 0;
    __syncthreads();
    int bin[kBinItems], rank[kBinItems];
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        bin[r] = bin_of(v[r].y, v[r].z, g, rows_per_bin);
        rank[r] = base + (long long)r * kBinThreads < n ? atomicAdd(&s_cnt[bin[r]], 1) : 0;
    }
    __syncthreads();
    {   // exclusive scan of the tile's band counts (one per thread) + the tile's run of every band
        const int c = tid < BINS ? s_cnt[tid] : 0;
        int incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        if (tid < BINS) s_base[tid] = c ? atomicAdd(cursors + tid * kBinPad, c) : 0;
        __syncthreads();
        int off = 0;
        for (int w = 0; w < warp; ++w) off += s_warp[w];
        if (tid < BINS) s_first[tid] = off + incl - c;
    }
    __syncthreads();
    // The records leave through shared memory in tile-sorted order (band by band): a warp then writes contiguous
    // runs instead of 32 half-sectors in as many different bands.
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const long long i = base + (long long)r * kBinThreads;
        if (i < n) {
            const int slot = s_first[bin[r]] + rank[r];
            s_stage[slot] = make_float4(v[r].x, v[r].y, v[r].z, __uint_as_float((unsigned)i));
            s_band[slot] = (unsigned char)bin[r];
            if (tags) tags[s_base[bin[r]] + rank[r]] = __float_as_uint(v[r].w);
        }
    }
    __syncthreads();
    const int here = (int)min((long long)kBinTile, (long long)n - (long long)blockIdx.x * kBinTile);
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const int slot = r * kBinThreads + tid;
        if (slot < here) {
            const int b = s_band[slot];
            recs[s_base[b] + slot - s_first[b]] = s_stage[slot];
        }
    }
}

// ------------------------------------------------------------------------------ synthetic input
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__global__ void kn_generate_kernel(float *xyz, long long first, long long count3, unsigned long long seed) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count3) return;
    const unsigned long long u = splitmix64(seed ^ (unsigned long long)(3 * first + t));
    xyz[t] = __fmul_rn(1000.0f, __fmul_rn((float)(u >> 40), 1.0f / 16777216.0f));
}

// out[perm[s]*k + j] = perm[knn[s*k + j]]  (test_knearests.cu:155-160 on the device)
__global__ void kn_remap_original_kernel(const unsigned int *__restrict__ knn,
                                         const unsigned int *__restrict__ perm, long long total, int k,
                                         unsigned int *__restrict__ out) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const long long s = e / k;
    const int j = (int)(e - s * k);
    const unsigned v = __ldg(knn + e);
    out[(long long)__ldg(perm + s) * k + j] = (v == KNC_EMPTY) ? KNC_EMPTY : __ldg(perm + v);
}

inline unsigned blocks_for(long long items, int per_block) {
    return (unsigned)((items + per_block - 1) / per_block);
}

}  // namespace

size_t kn_build_scratch_tiles(long long cells) { return (size_t)((cells + 1 + kScanTile - 1) / kScanTile); }

cudaError_t kn_launch_build(const float *d_in_xyz, int stride, int n, const GridGeom &g, long long cells, int *d_counters,
                            int *d_ptrs, int *d_globcounter, unsigned int *d_permutation, float *d_points3,
                            float4 *d_points4, unsigned int *d_tags, const BuildScratch &s, cudaStream_t st) {
    cudaError_t e;
    if (d_tags && stride != 4) return cudaErrorInvalidValue;  // tags are the 4th word of 16-byte input records
    if (d_tags && !s.d_rec_tags) return cudaErrorInvalidValue;
    if ((e = cudaMemsetAsync(d_counters, 0, sizeof(int) * (size_t)cells, st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(s.d_tile, 0, sizeof(unsigned long long) * s.tile_count, st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(s.d_ticket, 0, sizeof(int) * 2, st)) != cudaSuccess) return e;
    const float *src = d_in_xyz;
    int binned = 0;
    if (s.d_bin_recs) {  // large cloud: split into bands of x-rows first (see "bins" above)
        constexpr int bins = 256;  // measured at 100 M points: 64 bands 5.10 ms, 128 bands 4.54, 256 bands 4.34
        const int rows = (g.ze - g.zb) * g.Dy;
        const int rows_per_bin = (rows + bins - 1) / bins;
        const unsigned nblocks = blocks_for(n, kBinTile);
        const int vec4 = stride == 4 && ((uintptr_t)d_in_xyz & 15u) == 0;
        int *totals = s.d_bin_counts, *cursors = s.d_bin_counts + kMaxBins * kBinPad;
        unsigned int *btags = d_tags ? s.d_bin_tags : nullptr;
        if ((e = cudaMemsetAsync(totals, 0, sizeof(int) * kMaxBins * kBinPad, st)) != cudaSuccess) return e;
#define KN_BIN_LAUNCH(B)                                                                                                 \
    kn_bin_count_kernel<B><<<nblocks < 148u * 8u ? nblocks : 148u * 8u, kBinThreads, 0, st>>>(d_in_xyz, stride, n, g, rows_per_bin, totals);               \
    kn_bin_starts_kernel<<<1, kMaxBins, 0, st>>>(totals, B, cursors);                                                    \
    kn_bin_scatter_kernel<B><<<nblocks, kBinThreads, 0, st>>>(d_in_xyz, stride, n, g, rows_per_bin, cursors, s.d_bin_recs, btags, vec4);
        KN_BIN_LAUNCH(bins)
#undef KN_BIN_LAUNCH
        src = reinterpret_cast<const float *>(s.d_bin_recs);
        stride = 4;
        binned = 1;
    }
    const int aligned16 = ((uintptr_t)src & 15u) == 0;
    const unsigned pblocks = blocks_for(n, kPointTile);
    kn_count_kernel<<<pblocks, kBuildThreads, 0, st>>>(src, stride, n, g, d_counters, aligned16);
    kn_scan_kernel<<<(unsigned)s.tile_count, kBuildThreads, 0, st>>>(d_counters, d_ptrs, cells, s.d_tile,
                                                                    s.d_ticket, d_globcounter);
    kn_scatter_kernel<<<pblocks, kBuildThreads, 0, st>>>(src, stride, n, g, d_ptrs, s.d_recs, binned,
                                                         binned ? s.d_bin_tags : nullptr, d_tags ? s.d_rec_tags : nullptr,
                                                         aligned16);
    kn_order_kernel<<<blocks_for(n, kOrderTile), kBuildThreads, 0, st>>>(s.d_recs, d_tags ? s.d_rec_tags : nullptr, n, g, d_permutation,
                                                       d_points3, d_points4, d_tags, s.d_ticket + 1, s.d_biglist);
    kn_order_big_kernel<<<148, 1024, 0, st>>>(s.d_ticket + 1, s.d_biglist, d_ptrs, s.d_recs,
                                              d_tags ? s.d_rec_tags : nullptr, d_permutation, d_points3, d_points4, d_tags);
    return cudaGetLastError();
}

size_t kn_bin_counts_ints(int) { return (size_t)2 * kMaxBins * kBinPad; }

cudaError_t kn_launch_generate(float *d_xyz, long long first, long long count, uint64_t seed, cudaStream_t st) {
    const long long count3 = count * 3;
    if (count3 > 0)
        kn_generate_kernel<<<blocks_for(count3, 256), 256, 0, st>>>(d_xyz, first, count3, seed);
    return cudaGetLastError();
}

cudaError_t kn_launch_remap_original(const unsigned int *d_knn, const unsigned int *d_perm, int n, int k,
                                     unsigned int *d_out, cudaStream_t st) {
    const long long total = (long long)n * k;
    if (total > 0)
        kn_remap_original_kernel<<<blocks_for(total, 256), 256, 0, st>>>(d_knn, d_perm, total, k, d_out);
    return cudaGetLastError();
}
