// ref_shims.cpp -- extern "C" doors into the UNMODIFIED reference, for ctypes.
//
// TEST INFRASTRUCTURE ONLY (see oracle/knn_oracle.c header).  This file contains no algorithm:
// it only calls the reference's own classes/functions, which oracle/Makefile compiles from
// /root/reference into oracle/_ref/ (git-ignored).  Nothing of the reference is copied here.
//
//   libref_kdtree.so   = /root/reference/kd_tree.cpp  + this file (REF_SHIM_KDTREE)
//   libref_loader.so   = /root/reference/test_knearests.cu (main renamed) + knearests.cu
//                        + kd_tree.cpp + this file (REF_SHIM_LOADER)
#include <chrono>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef REF_SHIM_KDTREE
#include "kd_tree.h"  // the reference header, found through -I/root/reference

extern "C" {

// Mirrors the CPU arm of the reference test, test_knearests.cu:197-212: build KdTree(3) on the
// points (single-threaded), then for every query v ask for K+1 neighbours and drop slot 0 (the
// point itself at distance 0), OpenMP over queries.  Queries are the points [q0,q1).
// out_idx: (q1-q0)*K ints in ORIGINAL index space, nearest first.  out_d2 may be null.
// times[0] = tree build seconds, times[1] = query seconds; returns the thread count used.
int ref_kdtree_knn(const float *xyz, int n, int q0, int q1, int K, int *out_idx, float *out_d2,
                   double *times, int num_threads) {
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    KdTree KD(3);
    KD.set_points(n, xyz);
    auto t1 = clk::now();
    int used = 1;
#ifdef _OPENMP
    if (num_threads > 0) omp_set_num_threads(num_threads);
    used = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 256)
    for (int v = q0; v < q1; ++v) {
        std::vector<int> neigh(K + 1);
        std::vector<float> sq(K + 1);
        KD.get_nearest_neighbors(K + 1, v, neigh.data(), sq.data());
        for (int j = 0; j < K; ++j) {
            out_idx[(size_t)(v - q0) * K + j] = neigh[j + 1];
            if (out_d2) out_d2[(size_t)(v - q0) * K + j] = sq[j + 1];
        }
    }
    auto t2 = clk::now();
    if (times) {
        times[0] = std::chrono::duration<double>(t1 - t0).count();
        times[1] = std::chrono::duration<double>(t2 - t1).count();
    }
    return used;
}

// Same, with the query pass repeated `passes` times over the ONE tree (the tree build of a 100 M-point cloud takes
// minutes, a query pass over a sample seconds): times[0] = build seconds, times[1 + i] = seconds of pass i.
// Only the last pass's rows are kept.  Used by bench.py's CPU arm for a warm-up pass and timed passes.
int ref_kdtree_knn_passes(const float *xyz, int n, int q0, int q1, int K, int *out_idx, double *times,
                          int num_threads, int passes) {
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    KdTree KD(3);
    KD.set_points(n, xyz);
    times[0] = std::chrono::duration<double>(clk::now() - t0).count();
    int used = 1;
#ifdef _OPENMP
    if (num_threads > 0) omp_set_num_threads(num_threads);
    used = omp_get_max_threads();
#endif
    for (int pass = 0; pass < passes; ++pass) {
        auto t1 = clk::now();
#pragma omp parallel for schedule(dynamic, 256)
        for (int v = q0; v < q1; ++v) {  // test_knearests.cu:203-212
            std::vector<int> neigh(K + 1);
            std::vector<float> sq(K + 1);
            KD.get_nearest_neighbors(K + 1, v, neigh.data(), sq.data());
            for (int j = 0; j < K; ++j) out_idx[(size_t)(v - q0) * K + j] = neigh[j + 1];
        }
        times[1 + pass] = std::chrono::duration<double>(clk::now() - t1).count();
    }
    return used;
}

int ref_kdtree_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"
#endif  // REF_SHIM_KDTREE

#ifdef REF_SHIM_LOADER
// The reference's loader lives in test_knearests.cu:40-80 next to main(); the Makefile renames
// main and this declaration binds to the reference's definition.
bool load_file(const char *filename, std::vector<float> &xyz, bool normalize);

extern "C" int ref_load_xyz(const char *path, float *xyz, int capacity) {
    std::vector<float> v;
    if (!load_file(path, v, true)) return -1;
    int n = (int)(v.size() / 3);
    if (n <= capacity) std::memcpy(xyz, v.data(), v.size() * sizeof(float));
    return n;
}
#endif  // REF_SHIM_LOADER
