// kn_cli -- command-line driver in the shape of the reference's test_knearests (test_knearests.cu:117-172):
//   kn_cli points.xyz [neighbors.u32]
// Loads a .xyz cloud (first line = count, then "x y z" per line), rescales it into [0,1000]^3 exactly as the
// reference's loader does (bbox grown by 0.1 % of the largest extent, 1000.*(v-min)/maxside evaluated in double,
// test_knearests.cu:15-80), runs kn_prepare + kn_solve through the reference-compatible API and prints the same
// kind of lines the reference prints (bbox, kn_firstbuild: X msec, kn_solve: X msec, grid statistics), so that
// scripts that parse the reference's output keep working.  Optionally writes the neighbour table in ORIGINAL
// index space (the remap of :155-160) as raw u32.  The CPU kd-tree cross-check of the reference driver is not
// part of the product (the kd-tree is the reference's; see oracle/ and tests/).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "params.h"
#include "knearests.h"
#include "knearests_c.h"

static void bbox(const std::vector<float> &xyz, float mn[3], float mx[3]) {
    for (int c = 0; c < 3; c++) mn[c] = mx[c] = xyz[c];
    for (size_t i = 1; i < xyz.size() / 3; i++)
        for (int c = 0; c < 3; c++) {
            mn[c] = std::min(mn[c], xyz[3 * i + c]);
            mx[c] = std::max(mx[c], xyz[3 * i + c]);
        }
    float d = std::max(mx[0] - mn[0], std::max(mx[1] - mn[1], mx[2] - mn[2]));
    d = 0.001f * d;
    for (int c = 0; c < 3; c++) { mn[c] -= d; mx[c] += d; }
}

static bool load_xyz(const char *path, std::vector<float> &xyz) {
    std::ifstream in(path);
    if (!in) return false;
    std::string line;
    bool first = true;
    long declared = 0;
    while (std::getline(in, line)) {
        if (line.empty()) continue;
        std::istringstream iss(line);
        if (first) { iss >> declared; first = false; continue; }
        float x = 0, y = 0, z = 0;
        iss >> x >> y >> z;
        xyz.push_back(x); xyz.push_back(y); xyz.push_back(z);
    }
    if ((long)xyz.size() != declared * 3)
        std::cerr << "warning: header says " << declared << " points, file holds " << xyz.size() / 3 << std::endl;
    if (xyz.empty()) return false;
    float mn[3], mx[3];
    bbox(xyz, mn, mx);
    const float maxside = std::max(mx[0] - mn[0], std::max(mx[1] - mn[1], mx[2] - mn[2]));
    for (size_t i = 0; i < xyz.size() / 3; i++)
        for (int c = 0; c < 3; c++) xyz[3 * i + c] = (float)(1000. * (double)(xyz[3 * i + c] - mn[c]) / (double)maxside);
    bbox(xyz, mn, mx);
    std::cerr << "bbox [" << mn[0] << ":" << mx[0] << "], [" << mn[1] << ":" << mx[1] << ", [" << mn[2] << ":" << mx[2] << "]"
              << std::endl;
    return true;
}

int main(int argc, char **argv) {
    if (argc < 2) { std::cerr << "Usage: " << argv[0] << " points.xyz [neighbors.u32] [normalized_points.f32]" << std::endl; return 1; }
    setenv("KN_VERBOSE", "1", 0);  // kn_firstbuild / kn_solve lines, as the reference prints with IF_VERBOSE
    std::vector<float> points;
    if (!load_xyz(argv[1], points)) { std::cerr << argv[1] << ": could not load file" << std::endl; return 1; }
    const int nb_points = (int)(points.size() / 3);
    std::cerr << knc_version() << std::endl;

    kn_problem *kn = kn_prepare((float3 *)points.data(), nb_points);
    kn_solve(kn);
    kn_print_stats(kn);
    unsigned int *knearests = kn_get_knearests(kn);
    unsigned int *permutation = kn_get_permutation(kn);
    std::vector<unsigned int> neighbors((size_t)nb_points * DEFAULT_NB_PLANES);
    for (int i = 0; i < nb_points; i++)
        for (int j = 0; j < DEFAULT_NB_PLANES; j++) {
            const unsigned v = knearests[(size_t)i * DEFAULT_NB_PLANES + j];
            neighbors[(size_t)permutation[i] * DEFAULT_NB_PLANES + j] = v == KNC_EMPTY ? KNC_EMPTY : permutation[v];
        }
    std::free(permutation);
    std::free(knearests);
    kn_free(&kn);
    if (argc > 2) {
        FILE *f = std::fopen(argv[2], "wb");
        if (!f) { std::perror(argv[2]); return 1; }
        std::fwrite(neighbors.data(), sizeof(unsigned), neighbors.size(), f);
        std::fclose(f);
    }
    if (argc > 3) {
        FILE *f = std::fopen(argv[3], "wb");
        if (!f) { std::perror(argv[3]); return 1; }
        std::fwrite(points.data(), sizeof(float), points.size(), f);
        std::fclose(f);
    }
    return 0;
}
