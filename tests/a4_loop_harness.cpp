// a4_loop_harness.cpp -- times the reference's in-process interface for the projection,
//     VectorXi transform_set_into_vector(const std::unordered_set<unsigned long>&, int d)   (include/random_projection.h:11),
// in the shape of the sample loop of sketch() (src/project_everything.cpp:287-298): an OpenMP
// `parallel for schedule(dynamic)` over in-memory hash sets.  Test infrastructure, not product code:
// oracle/Makefile links it twice, once against the reference's own src/random_projection.cpp and once
// against our link-time replacement (csrc/shim/random_projection_b200.cpp); tests/perf_shim.py runs both
// and compares the checksums.
//
//   a4_loop_<impl> <n_samples> <hashes_per_sample> <d> <passes>
#include "random_projection.h"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <omp.h>
#include <unordered_set>
#include <vector>

static inline uint64_t mix(uint64_t x)          // any bijection will do: the sets only have to be distinct
{
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}

int main(int argc, char** argv)
{
    if (argc < 5) {
        std::fprintf(stderr, "usage: %s <n_samples> <hashes_per_sample> <d> <passes>\n", argv[0]);
        return 1;
    }
    const long n = std::atol(argv[1]), h = std::atol(argv[2]);
    const int d = std::atoi(argv[3]), passes = std::atoi(argv[4]);
    std::vector<std::unordered_set<unsigned long>> sets(n);
#pragma omp parallel for schedule(dynamic)
    for (long s = 0; s < n; ++s) {
        sets[s].reserve(h);
        for (long k = 0; k < h; ++k) sets[s].insert(mix((static_cast<uint64_t>(s) << 32) | static_cast<uint64_t>(k)) >> 10);
    }
    std::vector<VectorXi> out(n);
    out[0] = transform_set_into_vector(sets[0], d);          // first call: device context, not timed
    double best = 1e300;
    for (int p = 0; p < passes; ++p) {
        const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic)
        for (long s = 0; s < n; ++s) out[s] = transform_set_into_vector(sets[s], d);
        const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (dt < best) best = dt;
    }
    long long checksum = 0, total = 0;
    for (long s = 0; s < n; ++s) {
        total += static_cast<long long>(sets[s].size());
        for (int j = 0; j < d; ++j) checksum += static_cast<long long>(out[s][j]) * ((j % 7) + 1 + (s % 5));
    }
    std::printf("{\"samples\": %ld, \"hashes\": %lld, \"d\": %d, \"threads\": %d, \"seconds\": %.6f, "
                "\"samples_per_s\": %.1f, \"hashes_per_s\": %.1f, \"checksum\": %lld}\n",
                n, total, d, omp_get_max_threads(), best, n / best, total / best, checksum);
    return 0;
}
