// grid_build.cu -- uniform-grid construction for sm_100a: per-cell count, exclusive scan, stable
// counting-sort scatter with the permutation.  Replaces the reference's count / reserve / store kernels
// (knearests.cu:32-60) and their launcher kn_firstbuild (knearests.cu:152-201).
//
// What is different from the reference, and why:
//  * count   : the histogram atomic also hands every point its arrival rank inside its cell, so the
//              scatter needs no second round of atomics (the reference does 2 atomics per point,
//              :36 and :56); atomics are warp-aggregated with match.any, which matters when the
//              input arrives spatially coherent (many lanes of a warp in one cell).
//  * reserve : a real exclusive prefix sum (single-pass, decoupled look-back) instead of
//              atomicAdd(globalcounter, cnt) in arrival order (:45) => cells are laid out in ascending
//              cell id, deterministic, and an x-row of cells is one contiguous span of points.
//  * store   : ranks come from atomics, so the order inside a cell is arrival order; one extra pass
//              sorts each cell's few entries by original index => the whole layout is the stable
//              counting sort, bit-identical from run to run (the reference's is not).
//  * the sorted points are written twice: float3 (the API's d_stored_points) and float4
//              (x,y,z,bits(own sorted position)) -- 16-byte records the query kernels load with one 128-bit
//              load or copy into shared memory in bulk; the 4th word lets a staged copy name its source.
//
// HBM traffic per point (algorithmic): read 12 (count) + write 8 (cell id, rank) + read 8 / write 4
// (scatter) + read 4+12 / write 12+16 (gather) + 8 per cell for the scan ~= 79 B/point.
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

// one input record: a single 16-byte load for aligned 4-float records, else scalars (4th word = 0 for xyz input)
__device__ __forceinline__ float4 load_record(const float *rec, int stride, bool vec4) {
    if (vec4) return __ldg(reinterpret_cast<const float4 *>(rec));
    return make_float4(__ldg(rec), __ldg(rec + 1), __ldg(rec + 2), stride == 4 ? __ldg(rec + 3) : 0.f);
}

// ---------------------------------------------------------------------------------------- count
// One point per thread; a block stages its 256 points (768 floats) through shared memory with
// float4 loads, so global reads are 16-byte vectorised and fully coalesced although the input is AoS.
__global__ void __launch_bounds__(kBuildThreads)
kn_count_kernel(const float *__restrict__ xyz, int stride, int n, GridGeom g, int *__restrict__ counters,
                int *__restrict__ cellid, int *__restrict__ rank, int aligned16) {
    __shared__ __align__(16) float tile[kBuildThreads * 3];
    const int tid = threadIdx.x;
    const long long first = (long long)blockIdx.x * kBuildThreads;
    const int here = (int)min((long long)kBuildThreads, (long long)n - first);
    if (stride == 4) {  // 16-byte records (x, y, z, payload): one aligned float4 load per point, payload ignored
        if (tid < here) {
            const float *rec = xyz + (first + tid) * 4;
            float4 v;
            if (aligned16) v = __ldg(reinterpret_cast<const float4 *>(rec));
            else v = make_float4(__ldg(rec), __ldg(rec + 1), __ldg(rec + 2), 0.f);
            tile[3 * tid] = v.x; tile[3 * tid + 1] = v.y; tile[3 * tid + 2] = v.z;
        }
    } else {
        const float *src = xyz + first * 3;
        if (aligned16 && here == kBuildThreads) {
            if (tid < kBuildThreads * 3 / 4)
                reinterpret_cast<float4 *>(tile)[tid] = __ldg(reinterpret_cast<const float4 *>(src) + tid);
        } else {
            for (int i = tid; i < here * 3; i += kBuildThreads) tile[i] = __ldg(src + i);
        }
    }
    __syncthreads();
    const bool valid = tid < here;
    int cell = -1 - (tid & 31);  // invalid lanes: unique keys, never grouped
    if (valid) {
        int i = axis_cell(tile[3 * tid], g.Dfx, 0, g.Dx - 1);
        int j = axis_cell(tile[3 * tid + 1], g.Dfy, 0, g.Dy - 1);
        int k = axis_cell(tile[3 * tid + 2], g.Dfz, g.zb, g.ze - 1) - g.zb;
        cell = i + (j + k * g.Dy) * g.Dx;  // z-major, knearests.cu:29
    }
    // Warp-aggregated atomic: lanes in the same cell elect a leader that adds the group size once.
    // match.any costs one round per distinct value, so it is only used when the warp looks coherent
    // (some lane shares its cell with its neighbour: sorted or clustered input); a warp of a randomly
    // ordered cloud -- 32 different cells -- goes straight to 32 plain atomics.
    const int lane = tid & 31;
    const int next = __shfl_down_sync(0xffffffffu, cell, 1);
    const bool coherent = __any_sync(0xffffffffu, lane < 31 && next == cell);
    int r;
    if (coherent) {
        const unsigned peers = __match_any_sync(0xffffffffu, cell);
        const int leader = __ffs(peers) - 1;
        const int within = __popc(peers & ((1u << lane) - 1u));
        int base = 0;
        if (valid && lane == leader) base = atomicAdd(counters + cell, __popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        r = base + within;
    } else {
        r = valid ? atomicAdd(counters + cell, 1) : 0;
    }
    if (valid) {
        cellid[first + tid] = cell;
        rank[first + tid] = r;
    }
}

// ----------------------------------------------------------------------------------------- scan
// Single-pass exclusive prefix sum with decoupled look-back.  Tile = 256 threads x 16 cells.
// Tile state word: status in the top 2 bits (0 = not ready, 1 = tile aggregate, 2 = inclusive prefix),
// value in the low 32 bits -- one 64-bit store publishes both, so no fence is needed between them.
constexpr int kScanItems = 16;
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
    if (base + kScanItems <= cells) {
        const int4 *src = reinterpret_cast<const int4 *>(counters + base);
#pragma unroll
        for (int q = 0; q < kScanItems / 4; q++) {
            int4 t = __ldg(src + q);
            v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int q = 0; q < kScanItems; q++) v[q] = (base + q < cells) ? __ldg(counters + base + q) : 0;
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
            if (base + q < cells) ptrs[base + q] = run;
            run += v[q];
        }
    }
    // the tile that covers the last cell also writes the total (ptrs[cells]) and the public counter
    if ((long long)(tile + 1) * kScanTile >= cells && tid == kBuildThreads - 1) {
        ptrs[cells] = s_prefix + tile_sum;
        *globcounter = s_prefix + tile_sum;
    }
}

// -------------------------------------------------------------------------------------- scatter
__global__ void __launch_bounds__(kBuildThreads)
kn_scatter_kernel(int n, const int *__restrict__ cellid, const int *__restrict__ rank,
                  const int *__restrict__ ptrs, unsigned int *__restrict__ permutation) {
    const long long i = (long long)blockIdx.x * kBuildThreads + threadIdx.x;
    if (i < n) permutation[__ldg(ptrs + __ldg(cellid + i)) + __ldg(rank + i)] = (unsigned)i;
}

// ---------------------------------------------------------------------------------------- order
// Puts each cell's slice of the permutation into ascending original index.  One thread per cell:
// up to 8 entries go through a fixed 19-comparator network in registers, up to KN_BIG_CELL through an
// in-place insertion sort, larger cells are queued for kn_order_big_kernel.
__device__ __forceinline__ void cmpswap(unsigned &a, unsigned &b) {
    const unsigned lo = min(a, b), hi = max(a, b);
    a = lo; b = hi;
}

__global__ void __launch_bounds__(kBuildThreads)
kn_order_kernel(long long cells, const int *__restrict__ ptrs, unsigned int *__restrict__ permutation,
                int *big_count, int *__restrict__ big_list) {
    const long long c = (long long)blockIdx.x * kBuildThreads + threadIdx.x;
    if (c >= cells) return;
    const int b = __ldg(ptrs + c), cnt = __ldg(ptrs + c + 1) - b;
    if (cnt < 2) return;
    unsigned *seg = permutation + b;
    if (cnt <= 8) {
        unsigned v[8];
#pragma unroll
        for (int q = 0; q < 8; q++) v[q] = q < cnt ? seg[q] : 0xFFFFFFFFu;
        cmpswap(v[0], v[1]); cmpswap(v[2], v[3]); cmpswap(v[4], v[5]); cmpswap(v[6], v[7]);
        cmpswap(v[0], v[2]); cmpswap(v[1], v[3]); cmpswap(v[4], v[6]); cmpswap(v[5], v[7]);
        cmpswap(v[1], v[2]); cmpswap(v[5], v[6]); cmpswap(v[0], v[4]); cmpswap(v[3], v[7]);
        cmpswap(v[1], v[5]); cmpswap(v[2], v[6]);
        cmpswap(v[1], v[4]); cmpswap(v[3], v[6]);
        cmpswap(v[2], v[4]); cmpswap(v[3], v[5]);
        cmpswap(v[3], v[4]);
#pragma unroll
        for (int q = 0; q < 8; q++)
            if (q < cnt) seg[q] = v[q];
    } else if (cnt <= KN_BIG_CELL) {
        for (int a = 1; a < cnt; a++) {
            const unsigned key = seg[a];
            int s = a;
            while (s > 0 && seg[s - 1] > key) { seg[s] = seg[s - 1]; s--; }
            seg[s] = key;
        }
    } else {
        big_list[atomicAdd(big_count, 1)] = (int)c;
    }
}

// One block per oversized cell: bitonic network, every compare-exchange ascending (partner = mirror in
// the first step of a merge, i ^ j afterwards), so the virtual +inf padding beyond `cnt` never moves.
__global__ void __launch_bounds__(1024)
kn_order_big_kernel(const int *big_count, const int *__restrict__ big_list, const int *__restrict__ ptrs,
                    unsigned int *__restrict__ permutation) {
    const int nbig = *big_count;
    for (int w = blockIdx.x; w < nbig; w += gridDim.x) {
        const int c = big_list[w];
        const int b = ptrs[c];
        const unsigned cnt = (unsigned)(ptrs[c + 1] - b);
        unsigned *seg = permutation + b;
        for (unsigned long long k = 2; (k >> 1) < cnt; k <<= 1) {
            for (unsigned long long j = k >> 1; j > 0; j >>= 1) {
                for (unsigned long long t = threadIdx.x; t < cnt; t += blockDim.x) {
                    const unsigned long long partner = (j == (k >> 1)) ? (t ^ (k - 1)) : (t ^ j);
                    if (partner > t && partner < cnt) {
                        const unsigned a = seg[t], d = seg[partner];
                        if (a > d) { seg[t] = d; seg[partner] = a; }
                    }
                }
                __syncthreads();
            }
        }
    }
}

// ------------------------------------------------------------------------------------------ bins
// Large clouds arrive in random order, so the histogram atomics, the pointer reads of the scatter and
// the record reads of the gather each touch a random DRAM sector per point (13.4 ms for 100 M points).
// One coalesced pre-pass removes that: a STABLE split of the input into kBins z-bands of planes, written
// as 16-byte records (x, y, z, input index).  All later random accesses of a band then stay inside a few
// tens of MB -- L2-resident on B200 (126 MB) -- and the build drops to ~1/3 (profiles/r1_notes.md).
// Stable (input order kept inside a band), so "ascending record position" inside a cell is still
// "ascending original index" and the layout stays the deterministic stable counting sort.
constexpr int kBins = 64;       // 6 bits: the warp-level grouping below votes once per bit
constexpr int kBinThreads = 1024;
constexpr int kBinItems = 4;    // points per thread, all loaded before the first barrier (memory-level parallelism)
constexpr int kBinTile = kBinThreads * kBinItems;
constexpr int kBinGroups = kBinItems * (kBinThreads / 32);  // warp-sized runs of consecutive points in a tile

__device__ __forceinline__ int bin_of_z(float z, const GridGeom &g, int planes_per_bin) {
    return (axis_cell(z, g.Dfz, g.zb, g.ze - 1) - g.zb) / planes_per_bin;
}

// lanes of the warp holding the same band (6 votes; only meaningful on valid lanes)
__device__ __forceinline__ unsigned same_bin_lanes(int bin, bool valid) {
    unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < 6; ++b) {
        const bool bit = (bin >> b) & 1;
        const unsigned m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

// A tile is kBinTile consecutive points; item r of thread t is point  tile*kBinTile + r*kBinThreads + t,  so the
// warp-sized run  v = r*32 + warp  comes before run v+1 in input order.  On return before[r] is the number of
// tile points of the same band that precede item r, and s_cnt[kBinGroups][b] the tile's count of band b.
__device__ __forceinline__ void bin_ranks_in_tile(const int (&bin)[kBinItems], const bool (&valid)[kBinItems],
                                                  int (&before)[kBinItems], int (*s_cnt)[kBins]) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < (kBinGroups + 1) * kBins; i += kBinThreads) (&s_cnt[0][0])[i] = 0;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const unsigned peers = same_bin_lanes(bin[r], valid[r]);
        if (valid[r] && lane == __ffs(peers) - 1) s_cnt[r * 32 + warp][bin[r]] = __popc(peers);
        before[r] = __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    // exclusive prefix over the runs, per band: thread (band, chunk of 8 runs) scans its chunk, then adds the chunks below
    constexpr int kChunk = 8, kChunks = kBinGroups / kChunk;
    static_assert(kChunks * kBins == kBinThreads, "one thread per (band, chunk)");
    __shared__ int s_chunk[kChunks][kBins];
    const int band = tid & (kBins - 1), chunk = tid / kBins;
    int run = 0;
#pragma unroll
    for (int j = 0; j < kChunk; ++j) {
        const int c = s_cnt[chunk * kChunk + j][band];
        s_cnt[chunk * kChunk + j][band] = run;
        run += c;
    }
    s_chunk[chunk][band] = run;
    __syncthreads();
    int below = 0;
    for (int q = 0; q < chunk; ++q) below += s_chunk[q][band];
#pragma unroll
    for (int j = 0; j < kChunk; ++j) s_cnt[chunk * kChunk + j][band] += below;
    if (chunk == kChunks - 1) s_cnt[kBinGroups][band] = below + run;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kBinItems; ++r)
        if (valid[r]) before[r] += s_cnt[r * 32 + warp][bin[r]];
}

__global__ void __launch_bounds__(kBinThreads)
kn_bin_count_kernel(const float *__restrict__ xyz, int stride, int n, GridGeom g, int planes_per_bin, int nblocks,
                    int *__restrict__ blockcounts) {
    __shared__ int s_hist[kBins];
    if (threadIdx.x < kBins) s_hist[threadIdx.x] = 0;
    const long long base = (long long)blockIdx.x * kBinTile + threadIdx.x;
    float z[kBinItems];
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const long long i = base + (long long)r * kBinThreads;
        z[r] = i < n ? __ldg(xyz + i * stride + 2) : 0.f;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const bool valid = base + (long long)r * kBinThreads < n;
        const int bin = bin_of_z(z[r], g, planes_per_bin);
        const unsigned peers = same_bin_lanes(bin, valid);
        if (valid && lane == __ffs(peers) - 1) atomicAdd(&s_hist[bin], __popc(peers));
    }
    __syncthreads();
    if (threadIdx.x < kBins) blockcounts[(size_t)threadIdx.x * nblocks + blockIdx.x] = s_hist[threadIdx.x];
}

// one block per bin: exclusive scan of its row of per-block counts (in place), row total to totals[bin]
__global__ void __launch_bounds__(1024) kn_bin_scan_kernel(int *blockcounts, int nblocks, int *totals) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    int *row = blockcounts + (size_t)blockIdx.x * nblocks;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 1024) {
        const int idx = base + tid;
        const int v = idx < nblocks ? row[idx] : 0;
        int incl = v;
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        int off = s_carry;
        for (int wq = 0; wq < warp; ++wq) off += s_warp[wq];
        if (idx < nblocks) row[idx] = off + incl - v;
        __syncthreads();
        if (tid == 1023) s_carry = off + incl;
        __syncthreads();
    }
    if (tid == 0) totals[blockIdx.x] = s_carry;
}

__global__ void __launch_bounds__(kBinThreads)
kn_bin_scatter_kernel(const float *__restrict__ xyz, int stride, int n, GridGeom g, int planes_per_bin, int nblocks,
                      const int *__restrict__ offsets, const int *__restrict__ totals, float4 *__restrict__ recs,
                      unsigned int *__restrict__ tags, int vec4) {  // tags: 4th input words in banded order (stride 4) or NULL
    __shared__ int s_cnt[kBinGroups + 1][kBins];
    __shared__ int s_base[kBins];
    const long long base = (long long)blockIdx.x * kBinTile + threadIdx.x;
    float x[kBinItems], y[kBinItems], z[kBinItems];
    unsigned tag[kBinItems];
    bool valid[kBinItems];
    int bin[kBinItems], before[kBinItems];
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const long long i = base + (long long)r * kBinThreads;
        valid[r] = i < n;
        const float4 v = load_record(xyz + (valid[r] ? i : 0) * stride, stride, vec4);
        x[r] = v.x; y[r] = v.y; z[r] = v.z;
        tag[r] = __float_as_uint(v.w);
    }
    if (threadIdx.x < kBins) {  // first record of each band (the totals of the lower bands) + this tile's start inside it
        int b = 0;
        for (int q = 0; q < (int)threadIdx.x; ++q) b += __ldg(totals + q);
        s_base[threadIdx.x] = b + __ldg(offsets + (size_t)threadIdx.x * nblocks + blockIdx.x);
    }
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) bin[r] = bin_of_z(z[r], g, planes_per_bin);
    bin_ranks_in_tile(bin, valid, before, s_cnt);  // contains the barriers that publish s_base
    // The records leave through shared memory in tile-sorted order (band by band): a warp then writes one
    // contiguous run of 512 bytes instead of 32 half-sectors in ~25 different bands.
    extern __shared__ float4 s_stage[];          // [kBinTile]
    __shared__ unsigned char s_band[kBinTile];   // band of each staged slot
    __shared__ int s_first[kBins];               // first staged slot of each band
    if (threadIdx.x < 32) {                      // exclusive scan of the tile's 64 band counts, two per lane
        const int lane = threadIdx.x;
        const int c0 = s_cnt[kBinGroups][2 * lane], c1 = s_cnt[kBinGroups][2 * lane + 1];
        int incl = c0 + c1;
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        s_first[2 * lane] = incl - c0 - c1;
        s_first[2 * lane + 1] = incl - c1;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kBinItems; ++r)
        if (valid[r]) {
            const int slot = s_first[bin[r]] + before[r];
            s_stage[slot] = make_float4(x[r], y[r], z[r], __uint_as_float((unsigned)(base + (long long)r * kBinThreads)));
            s_band[slot] = (unsigned char)bin[r];
            if (tags) tags[s_base[bin[r]] + before[r]] = tag[r];
        }
    __syncthreads();
    const int here = (int)min((long long)kBinTile, (long long)n - (long long)blockIdx.x * kBinTile);
#pragma unroll
    for (int r = 0; r < kBinItems; ++r) {
        const int slot = r * kBinThreads + threadIdx.x;
        if (slot < here) {
            const int b = s_band[slot];
            recs[s_base[b] + slot - s_first[b]] = s_stage[slot];
        }
    }
}

// --------------------------------------------------------------------------------------- gather
__global__ void __launch_bounds__(kBuildThreads)
kn_gather_kernel(const float *__restrict__ xyz, int stride, int n, unsigned int *__restrict__ permutation,
                 float *__restrict__ points3, float4 *__restrict__ points4, int binned,
                 const unsigned int *__restrict__ bin_tags, unsigned int *__restrict__ tags, int vec4) {
    __shared__ float tile[kBuildThreads * 3];
    const int tid = threadIdx.x;
    const long long first = (long long)blockIdx.x * kBuildThreads;
    const long long s = first + tid;
    const int here = (int)min((long long)kBuildThreads, (long long)n - first);
    if (s < n) {
        unsigned i = permutation[s];
        const float *rec = xyz + (long long)stride * i;
        const float4 v = load_record(rec, stride, vec4);
        const float x = v.x, y = v.y, z = v.z;
        // the caller's 4th word travels with the point (stride-4 input): a multi-GPU caller keeps global ids there
        if (tags) tags[s] = binned ? __ldg(bin_tags + i) : __float_as_uint(v.w);
        if (binned) {  // `i` was a position in the binned records; their 4th word is the original index
            i = __float_as_uint(v.w);
            permutation[s] = i;
        }
        points4[s] = make_float4(x, y, z, __uint_as_float((unsigned)s));
        tile[3 * tid] = x; tile[3 * tid + 1] = y; tile[3 * tid + 2] = z;
    }
    __syncthreads();
    // AoS float3 output written as one contiguous, coalesced run per block
    float *dst = points3 + first * 3;
    for (int q = tid; q < here * 3; q += kBuildThreads) dst[q] = tile[q];
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

size_t kn_build_scratch_tiles(long long cells) { return (size_t)((cells + kScanTile - 1) / kScanTile); }

cudaError_t kn_launch_build(const float *d_in_xyz, int stride, int n, const GridGeom &g, long long cells, int *d_counters,
                            int *d_ptrs, int *d_globcounter, unsigned int *d_permutation, float *d_points3,
                            float4 *d_points4, unsigned int *d_tags, const BuildScratch &s, cudaStream_t st) {
    cudaError_t e;
    if (d_tags && stride != 4) return cudaErrorInvalidValue;  // tags are the 4th word of 16-byte input records
    if ((e = cudaMemsetAsync(d_counters, 0, sizeof(int) * (size_t)cells, st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(s.d_tile, 0, sizeof(unsigned long long) * s.tile_count, st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(s.d_ticket, 0, sizeof(int) * 2, st)) != cudaSuccess) return e;
    const float *src = d_in_xyz;
    int binned = 0;
    if (s.d_bin_recs) {  // large cloud: stable split into z-bands first (see "bins" above)
        const int planes = g.ze - g.zb;
        const int planes_per_bin = (planes + kBins - 1) / kBins;
        const int nblocks = (int)blocks_for(n, kBinTile);
        kn_bin_count_kernel<<<nblocks, kBinThreads, 0, st>>>(d_in_xyz, stride, n, g, planes_per_bin, nblocks, s.d_bin_counts);
        kn_bin_scan_kernel<<<kBins, 1024, 0, st>>>(s.d_bin_counts, nblocks, s.d_bin_counts + (size_t)kBins * nblocks);
        if ((e = cudaFuncSetAttribute(kn_bin_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(sizeof(float4) * kBinTile))) != cudaSuccess) return e;
        kn_bin_scatter_kernel<<<nblocks, kBinThreads, sizeof(float4) * kBinTile, st>>>(d_in_xyz, stride, n, g, planes_per_bin, nblocks, s.d_bin_counts,
                                                              s.d_bin_counts + (size_t)kBins * nblocks, s.d_bin_recs,
                                                              d_tags ? s.d_bin_tags : nullptr,
                                                              stride == 4 && ((uintptr_t)d_in_xyz & 15u) == 0);
        src = reinterpret_cast<const float *>(s.d_bin_recs);
        stride = 4;
        binned = 1;
    }
    const int aligned16 = ((uintptr_t)src & 15u) == 0;
    kn_count_kernel<<<blocks_for(n, kBuildThreads), kBuildThreads, 0, st>>>(src, stride, n, g, d_counters,
                                                                           s.d_cellid, s.d_rank, aligned16);
    kn_scan_kernel<<<(unsigned)s.tile_count, kBuildThreads, 0, st>>>(d_counters, d_ptrs, cells, s.d_tile,
                                                                    s.d_ticket, d_globcounter);
    kn_scatter_kernel<<<blocks_for(n, kBuildThreads), kBuildThreads, 0, st>>>(n, s.d_cellid, s.d_rank, d_ptrs,
                                                                             d_permutation);
    kn_order_kernel<<<blocks_for(cells, kBuildThreads), kBuildThreads, 0, st>>>(cells, d_ptrs, d_permutation,
                                                                               s.d_ticket + 1, s.d_biglist);
    kn_order_big_kernel<<<148, 1024, 0, st>>>(s.d_ticket + 1, s.d_biglist, d_ptrs, d_permutation);
    kn_gather_kernel<<<blocks_for(n, kBuildThreads), kBuildThreads, 0, st>>>(src, stride, n, d_permutation,
                                                                            d_points3, d_points4, binned,
                                                                            binned ? s.d_bin_tags : nullptr, d_tags,
                                                                            stride == 4 && aligned16);
    return cudaGetLastError();
}

size_t kn_bin_counts_ints(int n) { return (size_t)kBins * blocks_for(n, kBinTile) + kBins; }

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
