// kn_cli -- command-line driver in the shape of the reference's test_knearests (test_knearests.cu:117-172):
//   kn_cli points.xyz [neighbors.u32]
// Loads a .xyz cloud (first line = count, then "x y z" per line), rescales it into [0,1000]^3 exactly as the
// reference's loader does (bbox grown by 0.1 % of the largest extent, 1000.*(v-min)/maxside evaluated in double,
// test_knearests.cu:15-80), runs kn_prepare + kn_solve through the reference-compatible API and prints the same
// kind of lines the reference prints (bbox, kn_firstbuild: X msec, kn_solve: X msec, grid statistics), so that
// scripts that parse the reference's output keep working.  Optionally writes the neighbour table in ORIGINAL
// index space (the remap of :155-160) as raw u32.  The CPU kd-tree cross-check of the reference driver is not
// part of the product (the kd-tree is the reference's; see oracle/ and tests/).
//
// Parsing a 100 M-point text file takes minutes, so the loaded and rescaled cloud is kept beside the source as
// <points.xyz>.f32 (raw little-endian floats x y z, preceded by a 16-byte header: magic "KNF32\0\0\0" and the point
// count as u64) and read from there the next time, as long as it is not older than the .xyz file.  KN_NO_CACHE=1
// switches that off.
#include <sys/stat.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
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

static const char kCacheMagic[8] = {'K', 'N', 'F', '3', '2', 0, 0, 0};

static bool load_cache(const std::string &xyz_path, std::vector<float> &xyz) {
    const std::string path = xyz_path + ".f32";
    struct stat src, bin;
    if (stat(xyz_path.c_str(), &src) != 0 || stat(path.c_str(), &bin) != 0 || bin.st_mtime < src.st_mtime) return false;
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    char magic[8];
    uint64_t n = 0;
    bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, kCacheMagic, 8) == 0 && std::fread(&n, 8, 1, f) == 1 &&
              n > 0 && (uint64_t)bin.st_size == 16 + 12 * n;
    if (ok) {
        xyz.resize(3 * n);
        ok = std::fread(xyz.data(), sizeof(float), 3 * n, f) == 3 * n;
    }
    std::fclose(f);
    if (ok) {
        float mn[3], mx[3];
        bbox(xyz, mn, mx);  // same line as after a fresh load (test_knearests.cu:77)
        std::cerr << "bbox [" << mn[0] << ":" << mx[0] << "], [" << mn[1] << ":" << mx[1] << ", [" << mn[2] << ":" << mx[2] << "]"
                  << std::endl;
        std::cerr << "(points from the binary cache " << path << ")" << std::endl;
    } else {
        xyz.clear();
    }
    return ok;
}

static void save_cache(const std::string &xyz_path, const std::vector<float> &xyz) {
    const std::string path = xyz_path + ".f32", tmp = path + ".tmp";
    FILE *f = std::fopen(tmp.c_str(), "wb");
    if (!f) return;  // read-only directory: no cache, no error
    const uint64_t n = xyz.size() / 3;
    const bool ok = std::fwrite(kCacheMagic, 1, 8, f) == 8 && std::fwrite(&n, 8, 1, f) == 1 &&
                    std::fwrite(xyz.data(), sizeof(float), xyz.size(), f) == xyz.size();
    std::fclose(f);
    if (ok) std::rename(tmp.c_str(), path.c_str());
    else std::remove(tmp.c_str());
}

int main(int argc, char **argv) {
    if (argc < 2) { std::cerr << "Usage: " << argv[0] << " points.xyz [neighbors.u32] [normalized_points.f32]" << std::endl; return 1; }
    setenv("KN_VERBOSE", "1", 0);  // kn_firstbuild / kn_solve lines, as the reference prints with IF_VERBOSE
    std::vector<float> points;
    const char *nc = std::getenv("KN_NO_CACHE");
    const bool use_cache = !(nc && std::atoi(nc) != 0);
    if (!(use_cache && load_cache(argv[1], points))) {
        if (!load_xyz(argv[1], points)) { std::cerr << argv[1] << ": could not load file" << std::endl; return 1; }
        if (use_cache) save_cache(argv[1], points);
    }
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
