// dropin_main.cu -- a caller written against the REFERENCE interface only (params.h + knearests.h,
// C++ linkage, k from DEFAULT_NB_PLANES), linked against libknearests_b200.so.  It follows the call
// sequence of the reference's driver (test_knearests.cu:134-172): kn_prepare, kn_solve, kn_print_stats,
// kn_get_knearests, kn_get_permutation, the index remap of :155-160, the bijection check of :162-168,
// kn_free.  Input: a raw f32 xyz file; output: the remapped neighbour table as raw u32.
//   usage: dropin_test points.f32 neighbors.u32
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "params.h"
#include "knearests.h"

int main(int argc, char **argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: %s points.f32 neighbors.u32\n", argv[0]); return 2; }
    std::FILE *f = std::fopen(argv[1], "rb");
    if (!f) { std::perror(argv[1]); return 2; }
    std::fseek(f, 0, SEEK_END);
    const long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<float> points(bytes / sizeof(float));
    if (std::fread(points.data(), 1, bytes, f) != (size_t)bytes) return 2;
    std::fclose(f);
    const int nb_points = (int)(points.size() / 3);

    kn_problem *kn = kn_prepare((float3 *)points.data(), nb_points);
    kn_solve(kn);
    kn_print_stats(kn);
    std::fprintf(stderr, "grid %d x %d x %d, %d points, k = %d\n", kn->dimx, kn->dimy, kn->dimz, kn->allocated_points,
                 DEFAULT_NB_PLANES);
    unsigned int *knearests = kn_get_knearests(kn);
    unsigned int *permutation = kn_get_permutation(kn);
    float3 *sorted = kn_get_points(kn);
    std::vector<unsigned int> neighbors((size_t)nb_points * DEFAULT_NB_PLANES);
    for (int i = 0; i < nb_points; i++)
        for (int j = 0; j < DEFAULT_NB_PLANES; j++)
            neighbors[(size_t)permutation[i] * DEFAULT_NB_PLANES + j] = permutation[knearests[(size_t)i * DEFAULT_NB_PLANES + j]];
    int bad = 0;
    for (int i = 0; i < nb_points && !bad; i++)  // sorted points really are the permuted input
        bad = sorted[i].x != points[3 * permutation[i]] || sorted[i].z != points[3 * permutation[i] + 2];
    std::vector<unsigned int> p(permutation, permutation + nb_points);
    std::sort(p.begin(), p.end());
    for (int i = 0; i < nb_points && !bad; i++) bad = p[i] != (unsigned)i;  // bijection, :162-168
    std::free(permutation);
    std::free(knearests);
    std::free(sorted);
    kn_free(&kn);
    if (bad || kn != NULL) { std::fprintf(stderr, "permutation / kn_free check failed\n"); return 1; }
    f = std::fopen(argv[2], "wb");
    std::fwrite(neighbors.data(), sizeof(unsigned), neighbors.size(), f);
    std::fclose(f);
    return 0;
}
