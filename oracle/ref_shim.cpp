// ref_shim.cpp -- C wrapper around the UNMODIFIED reference kernel.
//
// TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/mvs_oracle.c header).
// Compiled by oracle/Makefile together with
//   /root/reference/src/random_projection.cpp   (compiled where it lies)
// into oracle/_ref/libref_projection.so.  Nothing from the reference is copied
// into this repository: this file only *calls*
//   VectorXi transform_set_into_vector(const std::unordered_set<unsigned long>&, int)
//   (/root/reference/include/random_projection.h:11).
#include <cstdint>
#include <cstring>
#include <unordered_set>
#include <vector>

#include "random_projection.h"   // from /root/reference/include via -I

extern "C" {

// One set -> one vector through the reference function.
void ref_transform_set(const uint64_t* hashes, int64_t n, int d, int32_t* out)
{
    std::unordered_set<unsigned long int> s;
    s.reserve(static_cast<size_t>(n) * 2 + 1);
    for (int64_t k = 0; k < n; ++k) s.insert(hashes[k]);
    VectorXi v = transform_set_into_vector(s, d);
    std::memcpy(out, v.data(), sizeof(int32_t) * static_cast<size_t>(d));
}

// The sample loop of sketch() (/root/reference/src/project_everything.cpp:289-298):
// in-memory unordered_sets, `#pragma omp parallel for schedule(dynamic)`, one
// call of the reference function per sample.  Set construction happens before
// the loop, as in the reference (its parse phase), and is excluded from
// *seconds_kernel, which times the loop only.
void ref_project_csr(const uint64_t* hashes, const uint64_t* offsets, int64_t n_samples,
                     int d, int32_t* out, double* seconds_kernel);

}  // extern "C"

#include <chrono>

extern "C" void ref_project_csr(const uint64_t* hashes, const uint64_t* offsets, int64_t n_samples,
                                int d, int32_t* out, double* seconds_kernel)
{
    std::vector<std::unordered_set<unsigned long int>> sets(static_cast<size_t>(n_samples));
    for (int64_t s = 0; s < n_samples; ++s) {
        auto& st = sets[static_cast<size_t>(s)];
        for (uint64_t k = offsets[s]; k < offsets[s + 1]; ++k) st.insert(hashes[k]);
    }
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic)
    for (int64_t s = 0; s < n_samples; ++s) {
        VectorXi v = transform_set_into_vector(sets[static_cast<size_t>(s)], d);
        std::memcpy(out + static_cast<size_t>(s) * static_cast<size_t>(d), v.data(),
                    sizeof(int32_t) * static_cast<size_t>(d));
    }
    auto t1 = std::chrono::steady_clock::now();
    if (seconds_kernel) *seconds_kernel = std::chrono::duration<double>(t1 - t0).count();
}

// The norm column of vector_norms.txt: the two statements of
// /root/reference/src/project_everything.cpp:328-329 on an arbitrary int32 row, compiled with the
// reference's Eigen and flags.  (They live inside sketch(), which only the binary can execute; the
// binary's own files are the primary pin, this is the per-row live check for any d and any values.)
#include <cmath>
extern "C" void ref_norms(const int32_t* vec, int64_t n, int d, double* out)
{
    for (int64_t s = 0; s < n; ++s) {
        VectorXi v = Eigen::Map<const VectorXi>(vec + static_cast<size_t>(s) * static_cast<size_t>(d), d);
        Eigen::VectorXf vec_f = v.cast<float>() / std::sqrt(static_cast<float>(d));
        double norm = vec_f.norm();
        out[s] = norm;
    }
}
