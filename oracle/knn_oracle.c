/*
 * knn_oracle.c -- CPU restatement of the ssloy/cuda_knearests hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the checker (or as the timed CPU arm).  The shipped engine
 * (cuda_knearests_b200/csrc) never links, loads or calls it and has no CPU fallback.
 *
 * Parity status: PINNED.  oracle/Makefile compiles the reference's own kd_tree.cpp (CPU) and
 * knearests.cu (GPU, run on the B200 box) from /root/reference into oracle/_ref/, and
 * tests/test_oracle_pinned.py + tests/test_parity_reference_gpu.py check this restatement
 * against both, plus the committed golden vectors in tests/golden/ (made by
 * tests/golden/make_golden.py from the reference's kd-tree on the reference's pts20K fixture).
 *
 * All citations are file:line in the reference repository.
 *
 * Build: gcc -O2 -std=c11 -ffp-contract=off -fopenmp -shared -fPIC   (see oracle/Makefile)
 *   -ffp-contract=off is REQUIRED: every fused multiply-add below is an explicit fmaf(), every
 *   other operation must round separately, exactly as the PTX of the reference does.
 */
#define _GNU_SOURCE
#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define KNO_EMPTY 0xFFFFFFFFu /* knearests.cu:108 -- unfilled slot sentinel (UINT_MAX) */

/* ------------------------------------------------------------------------------------------
 * Grid sizing -- knearests.cu:249:  max(1,(int)round(pow(numpoints / 3.1f, 1.0f / 3.0)))
 * f32 divide, then double pow with exponent (double)1.0f/3.0, then round-half-away.
 * `density` generalises the literal 3.1f (points per cell); pass 3.1f for the reference value.
 * ------------------------------------------------------------------------------------------ */
int kno_grid_dim(int numpoints, float density) {
    float per = (float)numpoints / density;
    int sz = (int)round(pow((double)per, (double)1.0f / 3.0));
    return sz < 1 ? 1 : sz;
}

/* ------------------------------------------------------------------------------------------
 * cellFromPoint -- knearests.cu:22-30.  (int)floor(p * (float)dim / 1000.f), clamped.
 * PTX: mul.f32, div.rn.f32, cvt.rmi, cvt.rzi.s32 (saturating; NaN -> 0).
 * ------------------------------------------------------------------------------------------ */
static inline int kno_axis_cell(float p, int dim) {
    float f = floorf((p * (float)dim) / 1000.f);
    if (!(f == f)) return 0;          /* NaN: cvt.rzi.s32.f32 gives 0 */
    if (f < 0.f) return 0;            /* max(0, .) -- also covers the saturated negative side */
    if (f > (float)(dim - 1)) return dim - 1;
    return (int)f;
}

int kno_cell_from_point(int xdim, int ydim, int zdim, float x, float y, float z) {
    int i = kno_axis_cell(x, xdim);
    int j = kno_axis_cell(y, ydim);
    int k = kno_axis_cell(z, zdim);
    return i + j * xdim + k * xdim * ydim; /* z-major: knearests.cu:29 */
}

/* ------------------------------------------------------------------------------------------
 * Grid build -- knearests.cu:32-60 (count / reserve / store), made deterministic.
 * The reference reserves cell blocks in atomic arrival order (:45) and fills each cell in
 * atomic arrival order (:56); both are non-deterministic.  The restatement -- and the engine --
 * fix the one canonical instance of that family: cells laid out in ascending cell id
 * (exclusive prefix sum of the counters) and points inside a cell in ascending original index
 * (stable counting sort).  Any reference run is this layout up to a permutation of cell blocks
 * and of points inside a cell, which the canonical comparison (kno_canonical_rows) removes.
 *
 * Outputs: counters[ncell], ptrs[ncell+1] (ptrs[ncell] == n), sorted xyz [3n],
 *          permutation[n] with permutation[sorted_pos] = original index (:57).
 * ------------------------------------------------------------------------------------------ */
void kno_build(const float *xyz, int n, int xdim, int ydim, int zdim,
               int *counters, int *ptrs, float *sorted_xyz, unsigned *permutation) {
    long ncell = (long)xdim * ydim * zdim;
    int *cell = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    memset(counters, 0, sizeof(int) * (size_t)ncell);
    for (int i = 0; i < n; i++) { /* count, :32-38 */
        cell[i] = kno_cell_from_point(xdim, ydim, zdim, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
        counters[cell[i]]++;
    }
    int run = 0; /* reserve, :40-48, as an exclusive scan */
    for (long c = 0; c < ncell; c++) { ptrs[c] = run; run += counters[c]; }
    ptrs[ncell] = run;
    int *fill = (int *)calloc((size_t)ncell, sizeof(int));
    for (int i = 0; i < n; i++) { /* store, :51-60, visiting points in index order => stable */
        int pos = ptrs[cell[i]] + fill[cell[i]]++;
        permutation[pos] = (unsigned)i;
        sorted_xyz[3 * pos] = xyz[3 * i];
        sorted_xyz[3 * pos + 1] = xyz[3 * i + 1];
        sorted_xyz[3 * pos + 2] = xyz[3 * i + 2];
    }
    free(fill);
    free(cell);
}

/* ------------------------------------------------------------------------------------------
 * Ring offset table -- knearests.cu:254-288.  Nmax = 16, rings 1..15, Chebyshev norm == ring,
 * k-major then j then i inside a ring; bound d = 1000.*(float)(ring-1)/(float)D evaluated in
 * double, stored to float, squared in float (:278-279).  Returns the entry count (31^3).
 * offsets/dists must hold 16^4 entries (:260).
 * ------------------------------------------------------------------------------------------ */
int kno_ring_table(int xdim, int ydim, int zdim, int *offsets, float *dists) {
    const int Nmax = 16;
    int n = 1;
    offsets[0] = 0;
    dists[0] = 0.0f;
    int dmax = xdim > ydim ? xdim : ydim;
    if (zdim > dmax) dmax = zdim;
    for (int ring = 1; ring < Nmax; ring++)
        for (int k = -Nmax; k <= Nmax; k++)
            for (int j = -Nmax; j <= Nmax; j++)
                for (int i = -Nmax; i <= Nmax; i++) {
                    int ai = abs(i), aj = abs(j), ak = abs(k);
                    int cheb = ai > aj ? ai : aj;
                    if (ak > cheb) cheb = ak;
                    if (cheb != ring) continue;
                    offsets[n] = i + j * xdim + k * xdim * ydim;
                    float d = (float)(1000. * (double)(float)(ring - 1) / (double)(float)dmax);
                    dists[n] = d * d;
                    n++;
                }
    return n;
}

/* Distance as the reference's GPU code computes it once nvcc contracts :125 (default -fmad=true):
 *   dx = c.x - p.x ...;  d = fma(dz,dz, fma(dx,dx, rn(dy*dy)))       (PTX probe) */
static inline float kno_dist_gpu(float cx, float cy, float cz, float px, float py, float pz) {
    float dx = cx - px, dy = cy - py, dz = cz - pz;
    return fmaf(dz, dz, fmaf(dx, dx, dy * dy));
}
float kno_distance_gpu(float cx, float cy, float cz, float px, float py, float pz) {
    return kno_dist_gpu(cx, cy, cz, px, py, pz);
}

/* Distance as the CPU kd-tree computes it -- kd_tree.cpp:13-19: ((0+dx*dx)+dy*dy)+dz*dz, d = p2-p1,
 * a rounding after every operation (the reference is built without FMA contraction). */
float kno_distance_kdtree(float qx, float qy, float qz, float px, float py, float pz) {
    float r = 0.0f, t;
    t = px - qx; r += t * t;
    t = py - qy; r += t * t;
    t = pz - qz; r += t * t;
    return r;
}

/* heapify / heapsort -- knearests.cu:66-91 (binary max-heap on vals, keys follow). */
static void kno_heapify(unsigned *keys, float *vals, int node, int size) {
    int j = node;
    for (;;) {
        int left = 2 * j + 1, right = 2 * j + 2, largest = j;
        if (left < size && vals[left] > vals[largest]) largest = left;
        if (right < size && vals[right] > vals[largest]) largest = right;
        if (largest == j) return;
        float tv = vals[j]; vals[j] = vals[largest]; vals[largest] = tv;
        unsigned tk = keys[j]; keys[j] = keys[largest]; keys[largest] = tk;
        j = largest;
    }
}
static void kno_heapsort(unsigned *keys, float *vals, int size) {
    while (size) {
        float tv = vals[0]; vals[0] = vals[size - 1]; vals[size - 1] = tv;
        unsigned tk = keys[0]; keys[0] = keys[size - 1]; keys[size - 1] = tk;
        kno_heapify(keys, vals, 0, --size);
    }
}

/* ------------------------------------------------------------------------------------------
 * knearest -- knearests.cu:93-148, one loop iteration per GPU thread.  Faithful restatement:
 * heap initialised to (UINT_MAX, FLT_MAX) (:107-110); do/while over the ring table with the
 * reference's off-by-one bound (:113,136; the extra entry is never reached on valid input and is
 * guarded here instead of read out of bounds); strict early exit (:116); linear cell range check
 * only (:119); self excluded by sorted position (:123); strict insert (:127); heapsort (:141);
 * row of sorted positions, ascending distance (:144-147).
 * Candidate visit order inside a cell follows this oracle's deterministic layout.
 * Returns the number of queries that exhausted the table (the reference's "no guarantee" case).
 * ------------------------------------------------------------------------------------------ */
int kno_knearest_refwalk(int xdim, int ydim, int zdim, int n, const int *ptrs, const int *counters,
                         const float *sorted_xyz, int num_offsets, const int *offsets,
                         const float *offset_dists, int K, unsigned *knn) {
    int exhausted = 0;
    long ncell = (long)xdim * ydim * zdim;
#pragma omp parallel for schedule(dynamic, 1024) reduction(+ : exhausted)
    for (int q = 0; q < n; q++) {
        unsigned keys[K];
        float vals[K];
        float px = sorted_xyz[3 * q], py = sorted_xyz[3 * q + 1], pz = sorted_xyz[3 * q + 2];
        int cell_in = kno_cell_from_point(xdim, ydim, zdim, px, py, pz);
        for (int i = 0; i < K; i++) { keys[i] = KNO_EMPTY; vals[i] = FLT_MAX; }
        int o = 0;
        do {
            if (o >= num_offsets) break; /* reference reads one past the table here */
            float min_dist = offset_dists[o];
            if (vals[0] < min_dist) break;
            long cell = (long)cell_in + offsets[o];
            if (cell >= 0 && cell < ncell) {
                int base = ptrs[cell], num = counters[cell];
                for (int p = base; p < base + num; p++) {
                    if (p == q) continue;
                    float d = kno_dist_gpu(sorted_xyz[3 * p], sorted_xyz[3 * p + 1], sorted_xyz[3 * p + 2],
                                           px, py, pz);
                    if (d < vals[0]) {
                        keys[0] = (unsigned)p;
                        vals[0] = d;
                        kno_heapify(keys, vals, 0, K);
                    }
                }
            }
        } while (o++ < num_offsets);
        if (o >= num_offsets) exhausted++;
        kno_heapsort(keys, vals, K);
        for (int i = 0; i < K; i++) knn[(size_t)q * K + i] = keys[i];
    }
    return exhausted;
}

/* ------------------------------------------------------------------------------------------
 * The engine's stated contract, computed exactly (not a restatement of a reference function: the
 * reference leaves ties to thread arrival order).  For every sorted position q:
 * the K smallest candidates p != q under the TOTAL order (d2_gpu(p,q), p) -- p the sorted
 * position -- listed ascending in that order; rows shorter than K are padded with UINT_MAX.
 * Exact for any input: rings are walked per axis with bounds checks until the ring bound
 * (knearests.cu:278-279 formula generalised past Nmax) strictly exceeds the k-th distance (:116).
 * ------------------------------------------------------------------------------------------ */
typedef struct { float d; unsigned p; } kno_ent;

static inline int kno_ent_less(float d, unsigned p, const kno_ent *e) {
    return d < e->d || (d == e->d && p < e->p);
}

void kno_knn_exact(int xdim, int ydim, int zdim, int n, const int *ptrs,
                   const float *sorted_xyz, int K, unsigned *knn, float *knn_d2 /* may be NULL */) {
    /* The reference's grid is a cube and its rings are Chebyshev shells of whole cells (:262-277).  For a grid with
     * different resolutions per axis the shells are kept equally thick in SPACE: shell t is the box of cells within
     * ext_a(t) = ceil(t * wmax / w_a) cells of the query's cell along axis a, minus the box of shell t-1 (for a
     * cube ext = t: the reference's rings).  A cell outside box t is more than ext_a(t) cells away along some axis
     * a, so every point in it is at least ext_a(t) * w_a >= t * wmax away. */
    const float wx = 1000.f / (float)xdim, wy = 1000.f / (float)ydim, wz = 1000.f / (float)zdim;
    float wmax = wx > wy ? wx : wy;
    if (wz > wmax) wmax = wz;
    int dmin = xdim < ydim ? xdim : ydim;
    if (zdim < dmin) dmin = zdim;
    int dmax = xdim > ydim ? xdim : ydim;
    if (zdim > dmax) dmax = zdim;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int q = 0; q < n; q++) {
        kno_ent best[K];
        int cnt = 0;
        float px = sorted_xyz[3 * q], py = sorted_xyz[3 * q + 1], pz = sorted_xyz[3 * q + 2];
        int ci = kno_axis_cell(px, xdim), cj = kno_axis_cell(py, ydim), ck = kno_axis_cell(pz, zdim);
        int pex = -1, pey = -1, pez = -1; /* extents of the previous shell's box */
        for (int ring = 0; ring <= dmax; ring++) {
            if (cnt == K && ring >= 1) {
                /* :278-279 bound (distance to the shell = (ring-1) shell widths), shrunk by 0.1% of a cell so that
                 * f32 rounding at cell borders can never end the walk early; visiting extra shells cannot change
                 * a total-order result */
                float b = (float)((double)(float)(ring - 1) * (double)wmax) - 1e-3f * wmax;
                if (b > 0.f && best[K - 1].d < b * b) break;
            }
            int ex = (int)ceil((double)ring * wmax / wx - 1e-9), ey = (int)ceil((double)ring * wmax / wy - 1e-9),
                ez = (int)ceil((double)ring * wmax / wz - 1e-9);
            if (pex >= xdim && pey >= ydim && pez >= zdim) break; /* the whole grid has been visited */
            for (int k = ck - ez; k <= ck + ez; k++) {
                if (k < 0 || k >= zdim) continue;
                for (int j = cj - ey; j <= cj + ey; j++) {
                    if (j < 0 || j >= ydim) continue;
                    int inner_yz = abs(k - ck) <= pez && abs(j - cj) <= pey; /* inside the previous box in y and z */
                    for (int i = ci - ex; i <= ci + ex; i++) {
                        if (inner_yz && abs(i - ci) <= pex) { i = ci + pex; continue; } /* skip the previous box */
                        if (i < 0 || i >= xdim) continue;
                        long cell = i + (long)j * xdim + (long)k * xdim * ydim;
                        for (int p = ptrs[cell]; p < ptrs[cell + 1]; p++) {
                            if (p == q) continue;
                            float d = kno_dist_gpu(sorted_xyz[3 * p], sorted_xyz[3 * p + 1],
                                                   sorted_xyz[3 * p + 2], px, py, pz);
                            if (cnt == K && !kno_ent_less(d, (unsigned)p, &best[K - 1])) continue;
                            int s = cnt < K ? cnt++ : K - 1;
                            while (s > 0 && kno_ent_less(d, (unsigned)p, &best[s - 1])) {
                                best[s] = best[s - 1];
                                s--;
                            }
                            best[s].d = d;
                            best[s].p = (unsigned)p;
                        }
                    }
                }
            }
            pex = ex; pey = ey; pez = ez;
        }
        for (int i = 0; i < K; i++) {
            knn[(size_t)q * K + i] = i < cnt ? best[i].p : KNO_EMPTY;
            if (knn_d2) knn_d2[(size_t)q * K + i] = i < cnt ? best[i].d : FLT_MAX;
        }
    }
    (void)dmin;
}

/* The engine's default grid (no reference counterpart: the reference sizes a cube for 3.1 points per cell,
 * knearests.cu:249).  Restated from the rule in include/knearests_c.h (knc_fast_dims): y and z cells 2 % wider
 * than the radius r0 of the sphere expected to hold k + 2.5 sqrt(k) + 2 points of a uniform cloud of n points in
 * [0,1000]^3, and one cell per point overall. */
void kno_fast_dims(int n, int k, int *dims) {
    double lambda = k + 2.5 * sqrt((double)k) + 2.0;
    double r0 = cbrt(3.0 * lambda / (4.0 * 3.14159265358979323846 * ((double)n / 1e9)));
    int dyz = (int)floor(1000.0 / (1.02 * r0));
    if (dyz < 1) dyz = 1;
    if (dyz > 1024) dyz = 1024;
    long long dx = llround((double)n / ((double)dyz * dyz));
    if (dx < 1) dx = 1;
    if (dx > 1048576) dx = 1048576;
    dims[0] = (int)dx; dims[1] = dyz; dims[2] = dyz;
}

/* Brute force O(n*m) in an arbitrary index space (no grid at all): for queries q in [q0,q1) the K
 * best of all points p != q under (dist(p,q), p).  formula 0 = GPU (F11), 1 = kd-tree (:13-19).
 * Used to check kno_knn_exact itself and to emulate the kd-tree's answer on tie-free inputs. */
void kno_brute(const float *xyz, int n, int q0, int q1, int K, int formula, unsigned *knn, float *knn_d2) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int q = q0; q < q1; q++) {
        kno_ent best[K];
        int cnt = 0;
        float px = xyz[3 * q], py = xyz[3 * q + 1], pz = xyz[3 * q + 2];
        for (int p = 0; p < n; p++) {
            if (p == q) continue;
            float d = formula == 0 ? kno_dist_gpu(xyz[3 * p], xyz[3 * p + 1], xyz[3 * p + 2], px, py, pz)
                                   : kno_distance_kdtree(px, py, pz, xyz[3 * p], xyz[3 * p + 1], xyz[3 * p + 2]);
            if (cnt == K && !kno_ent_less(d, (unsigned)p, &best[K - 1])) continue;
            int s = cnt < K ? cnt++ : K - 1;
            while (s > 0 && kno_ent_less(d, (unsigned)p, &best[s - 1])) { best[s] = best[s - 1]; s--; }
            best[s].d = d;
            best[s].p = (unsigned)p;
        }
        size_t row = (size_t)(q - q0) * K;
        for (int i = 0; i < K; i++) {
            knn[row + i] = i < cnt ? best[i].p : KNO_EMPTY;
            if (knn_d2) knn_d2[row + i] = i < cnt ? best[i].d : FLT_MAX;
        }
    }
}

/* Brute force for query points that are NOT members of the cloud (the reference's kd-tree takes an
 * arbitrary query point, kd_tree.cpp:168-192; its GPU path does not).  K best of all n points under
 * (d2_gpu, index); nothing is excluded. */
void kno_brute_external(const float *xyz, int n, const float *queries, int m, int K, unsigned *knn, float *knn_d2) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int q = 0; q < m; q++) {
        kno_ent best[K];
        int cnt = 0;
        float px = queries[3 * q], py = queries[3 * q + 1], pz = queries[3 * q + 2];
        for (int p = 0; p < n; p++) {
            float d = kno_dist_gpu(xyz[3 * p], xyz[3 * p + 1], xyz[3 * p + 2], px, py, pz);
            if (cnt == K && !kno_ent_less(d, (unsigned)p, &best[K - 1])) continue;
            int s = cnt < K ? cnt++ : K - 1;
            while (s > 0 && kno_ent_less(d, (unsigned)p, &best[s - 1])) { best[s] = best[s - 1]; s--; }
            best[s].d = d;
            best[s].p = (unsigned)p;
        }
        for (int i = 0; i < K; i++) {
            knn[(size_t)q * K + i] = i < cnt ? best[i].p : KNO_EMPTY;
            if (knn_d2) knn_d2[(size_t)q * K + i] = i < cnt ? best[i].d : FLT_MAX;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Canonical comparison form (the remap is test_knearests.cu:155-160).
 * in : knn[n*K] rows indexed by sorted position holding sorted positions, permutation[n],
 *      xyz_orig[3n] original-order coordinates.
 * out: rows indexed by ORIGINAL index holding ORIGINAL indices, each row sorted ascending by
 *      (d2_gpu, original index); d2 rows alongside.  UINT_MAX entries stay UINT_MAX, sorted last.
 * ------------------------------------------------------------------------------------------ */
static int kno_ent_cmp(const void *a, const void *b) {
    const kno_ent *x = (const kno_ent *)a, *y = (const kno_ent *)b;
    if (x->d < y->d) return -1;
    if (x->d > y->d) return 1;
    return (x->p > y->p) - (x->p < y->p);
}

void kno_canonical_rows(const unsigned *knn, const unsigned *permutation, const float *xyz_orig,
                        int n, int K, unsigned *out_idx, float *out_d2) {
#pragma omp parallel for schedule(static)
    for (int s = 0; s < n; s++) {
        kno_ent row[K];
        unsigned q = permutation ? permutation[s] : (unsigned)s;
        for (int i = 0; i < K; i++) {
            unsigned v = knn[(size_t)s * K + i];
            if (v == KNO_EMPTY) { row[i].p = KNO_EMPTY; row[i].d = FLT_MAX; continue; }
            unsigned p = permutation ? permutation[v] : v;
            row[i].p = p;
            row[i].d = kno_dist_gpu(xyz_orig[3 * p], xyz_orig[3 * p + 1], xyz_orig[3 * p + 2],
                                    xyz_orig[3 * q], xyz_orig[3 * q + 1], xyz_orig[3 * q + 2]);
        }
        qsort(row, (size_t)K, sizeof(kno_ent), kno_ent_cmp);
        for (int i = 0; i < K; i++) {
            out_idx[(size_t)q * K + i] = row[i].p;
            out_d2[(size_t)q * K + i] = row[i].d;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * .xyz loader + normaliser -- test_knearests.cu:15-80.  First non-empty line = count, then
 * "x y z" per line (parsed as float, :47,56).  bbox (:15-38) grown by 0.001f*max_extent on both
 * sides; coordinates mapped with 1000.*(v-min)/maxside evaluated in double, stored float (:72-74).
 * Returns the number of points read (caller passes capacity), or -1 if the file cannot be opened.
 * ------------------------------------------------------------------------------------------ */
static void kno_bbox(const float *xyz, int n, float *mn, float *mx) {
    for (int c = 0; c < 3; c++) mn[c] = mx[c] = xyz[c];
    for (int i = 1; i < n; i++)
        for (int c = 0; c < 3; c++) {
            float v = xyz[3 * i + c];
            if (v < mn[c]) mn[c] = v;
            if (v > mx[c]) mx[c] = v;
        }
    float d = mx[0] - mn[0];
    if (mx[1] - mn[1] > d) d = mx[1] - mn[1];
    if (mx[2] - mn[2] > d) d = mx[2] - mn[2];
    d = 0.001f * d;
    for (int c = 0; c < 3; c++) { mn[c] -= d; mx[c] += d; }
}

void kno_normalize(float *xyz, int n) {
    float mn[3], mx[3];
    kno_bbox(xyz, n, mn, mx);
    float maxside = mx[0] - mn[0];
    if (mx[1] - mn[1] > maxside) maxside = mx[1] - mn[1];
    if (mx[2] - mn[2] > maxside) maxside = mx[2] - mn[2];
    for (int i = 0; i < n; i++)
        for (int c = 0; c < 3; c++)
            xyz[3 * i + c] = (float)(1000. * (double)(xyz[3 * i + c] - mn[c]) / (double)maxside);
}

int kno_load_xyz(const char *path, float *xyz, int capacity, int normalize) {
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    char *line = NULL;
    size_t cap = 0;
    int first = 1, n = 0;
    while (getline(&line, &cap, f) >= 0) {
        size_t len = strlen(line);
        while (len && (line[len - 1] == '\n' || line[len - 1] == '\r')) line[--len] = 0;
        if (!len) continue;
        if (first) { first = 0; continue; } /* declared count; :62 only asserts it */
        float x = 0, y = 0, z = 0;
        sscanf(line, "%f %f %f", &x, &y, &z);
        if (n < capacity) { xyz[3 * n] = x; xyz[3 * n + 1] = y; xyz[3 * n + 2] = z; }
        n++;
    }
    free(line);
    fclose(f);
    if (normalize && n > 0 && n <= capacity) kno_normalize(xyz, n);
    return n;
}

/* ------------------------------------------------------------------------------------------
 * Synthetic clouds: counter-based, so any slice can be generated independently on
 * the host or on the device: u = splitmix64(seed ^ (3*i + c)); coord = 1000 * (u >> 40) * 2^-24.
 * Not a reference function -- the reference has no generator; listed here so that the CPU side
 * and the engine's device generator can be compared bit for bit.
 * ------------------------------------------------------------------------------------------ */
static inline uint64_t kno_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
void kno_generate_uniform(uint64_t seed, long first, long count, float *xyz) {
#pragma omp parallel for schedule(static)
    for (long t = 0; t < count; t++)
        for (int c = 0; c < 3; c++) {
            uint64_t u = kno_splitmix64(seed ^ (uint64_t)(3 * (first + t) + c));
            xyz[3 * t + c] = 1000.0f * ((float)(u >> 40) * (1.0f / 16777216.0f));
        }
}
