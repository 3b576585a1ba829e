// noalloc.cu -- knc_scan_cells_noalloc: the grid build's exclusive scan for callers that may not
// allocate device memory.
//
// Who that is: programs that gave the whole GPU to their own memory pool at start-up (training and
// inference frameworks, simulation codes with a fixed arena), so that a cudaMalloc made behind their
// back can fail or eat the head-room they planned for; and latency-sensitive loops, where cudaMalloc /
// cudaFree stall the device.  Such a caller builds the per-cell counters itself (or keeps the ones of
// a knc_problem) in memory it owns, owns the output array too, and only needs the scan done.
//
// The single-pass scan needs a little scratch that is neither input nor output: one status word per
// tile of 4096 cells plus a ticket counter.  The regular path (grid_build.cu, behind knc_prepare)
// takes it from the stream-ordered allocator.  Here it lives in a file-scope __device__ array of
// fixed size instead -- KNC_NOALLOC_MAX_CELLS cells at most, about 33 KB, resident on every GPU that
// loads this library until the process exits.
//
// Nothing in this file calls cudaMalloc, cudaFree, their async/pool forms, or any library routine
// that allocates.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "kn_internal.h"

namespace {

constexpr int kThreads = 256;
constexpr int kItems = 16;
constexpr int kTile = kThreads * kItems;  // 4096 cells per tile, as in grid_build.cu
constexpr long long kMaxTiles = (KNC_NOALLOC_MAX_CELLS + kTile - 1) / kTile;
constexpr int kStateOffset = 32;  // words; word 0 is the ticket, tile states start 256 bytes in

// The scratch itself.  In HOST code the name `g_scan_scratch` is only a handle for the runtime, not
// an address the GPU understands: handing it to a kernel would fault (or, where the GPU can read host
// memory, quietly scribble on the wrong bytes).  Host code must ask cudaGetSymbolAddress for the
// device address -- see scratch_for_current_device() -- and pass THAT to the kernel.
static __device__ __align__(256) unsigned long long g_scan_scratch[kStateOffset + kMaxTiles];

std::mutex g_setup_lock;
unsigned long long *g_scratch_addr[64] = {nullptr};  // per device, filled on first use

// Once per device, with that device current: resolve the device address of the scratch array.
cudaError_t scratch_for_current_device(unsigned long long **out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> hold(g_setup_lock);
    if (!g_scratch_addr[dev]) {
        unsigned long long *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_scan_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[dev] = ptr;
    }
    *out = g_scratch_addr[dev];
    return cudaSuccess;
}

// Variant of kn_scan_kernel (grid_build.cu) whose tile states and ticket come from `scratch`.
// Status in the top 2 bits of a state word (0 not ready, 1 tile aggregate, 2 inclusive prefix), value
// in the low 32 bits; tiles are numbered by a ticket in the order blocks start, so a block only ever
// waits for blocks that are already running.
__global__ void __launch_bounds__(kThreads)
kn_scan_noalloc_kernel(const int *__restrict__ counters, int *__restrict__ ptrs, long long cells,
                       unsigned long long *scratch) {
    __shared__ int s_tile, s_prefix;
    __shared__ int s_warp[kThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    volatile unsigned long long *state = scratch + kStateOffset;
    if (tid == 0) s_tile = (int)atomicAdd(reinterpret_cast<unsigned int *>(scratch), 1u);
    __syncthreads();
    const int tile = s_tile;
    const long long base = (long long)tile * kTile + (long long)tid * kItems;

    int v[kItems];
    int sum = 0;
#pragma unroll
    for (int q = 0; q < kItems; q++) {
        v[q] = (base + q < cells) ? __ldg(counters + base + q) : 0;
        sum += v[q];
    }
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int warp_off = 0, tile_sum = 0;
#pragma unroll
    for (int wq = 0; wq < kThreads / 32; wq++) {
        const int t = s_warp[wq];
        if (wq < warp) warp_off += t;
        tile_sum += t;
    }
    if (warp == 0) {
        if (lane == 0) state[tile] = ((unsigned long long)(tile == 0 ? 2 : 1) << 62) | (unsigned)tile_sum;
        int exclusive = 0;
        if (tile > 0) {
            int look = tile - 1;
            while (true) {
                unsigned long long wrd = 2ull << 62;  // before tile 0: inclusive prefix 0
                const int idx = look - lane;
                if (idx >= 0) {
                    do { wrd = state[idx]; } while ((wrd >> 62) == 0);
                }
                const unsigned full = __ballot_sync(0xffffffffu, (wrd >> 62) == 2);
                const int stop = full ? __ffs(full) - 1 : 32;
                int contrib = (lane <= stop) ? (int)(unsigned)wrd : 0;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, d);
                exclusive += contrib;
                if (full) break;
                look -= 32;
            }
            if (lane == 0) state[tile] = (2ull << 62) | (unsigned)(exclusive + tile_sum);
        }
        if (lane == 0) s_prefix = exclusive;
    }
    __syncthreads();
    int run = s_prefix + warp_off + incl - sum;
#pragma unroll
    for (int q = 0; q < kItems; q++) {
        if (base + q < cells) ptrs[base + q] = run;
        run += v[q];
    }
    if ((long long)(tile + 1) * kTile >= cells && tid == kThreads - 1) ptrs[cells] = s_prefix + tile_sum;
}

}  // namespace

// Exclusive prefix sum of d_counters[0..cells) into d_ptrs[0..cells], d_ptrs[cells] = total.
// Caller owns both arrays (device memory on the current device); `stream` is a cudaStream_t or NULL.
//  * Allocates nothing, on the host or the device; scratch is the static array above, so at most
//    KNC_NOALLOC_MAX_CELLS cells -- more is refused with KNC_ERR_INVALID before anything is launched.
//  * The first call on each device looks up the scratch address (cudaGetSymbolAddress).  Make that
//    first call during start-up, before any stream capture is begun.
//  * All calls on a device share the one scratch array and it still holds the previous call's tile
//    states, so every call first clears the part it will use (stream-ordered memset), then scans.
//    Two calls whose GPU work can overlap -- different streams, or different host threads -- would
//    clear and write each other's states: serialise them.  The call itself returns only after its
//    stream has drained, like the rest of this ABI.
extern "C" int knc_scan_cells_noalloc(const int *d_counters, long long cells, int *d_ptrs, void *stream) {
    if (!d_counters || !d_ptrs) return kn_set_error(KNC_ERR_INVALID, "knc_scan_cells_noalloc: NULL array");
    if (cells < 1 || cells > KNC_NOALLOC_MAX_CELLS)
        return kn_set_error(KNC_ERR_INVALID,
                            "knc_scan_cells_noalloc: %lld cells is outside 1..%lld (KNC_NOALLOC_MAX_CELLS, the "
                            "size of the built-in scratch); use knc_prepare for larger grids",
                            cells, (long long)KNC_NOALLOC_MAX_CELLS);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long *scratch = nullptr;
    cudaError_t e = scratch_for_current_device(&scratch);
    const long long tiles = (cells + kTile - 1) / kTile;
    if (e == cudaSuccess) e = cudaMemsetAsync(scratch, 0, sizeof(unsigned long long) * (size_t)(kStateOffset + tiles), st);
    if (e == cudaSuccess) {
        kn_scan_noalloc_kernel<<<(unsigned)tiles, kThreads, 0, st>>>(d_counters, d_ptrs, cells, scratch);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return kn_set_error(KNC_ERR_CUDA, "knc_scan_cells_noalloc: %s", cudaGetErrorString(e));
    return KNC_OK;
}
