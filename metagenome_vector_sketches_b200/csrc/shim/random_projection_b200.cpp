// random_projection_b200.cpp -- link-time replacement for the reference's src/random_projection.cpp.
//
// Same signature as reference include/random_projection.h:11, so the reference's own
// project_everything.cpp / standalone_projection.cpp compile and link against it UNMODIFIED
// (INTEGRATION.md option 2).  Needs the reference's include dirs for <Eigen/Dense>; it is therefore
// not part of libmvs_b200.so and is built only where the reference tree is present.
#include "random_projection.h"

#include "mvs_b200.h"

#include <mutex>
#include <stdexcept>
#include <vector>

VectorXi transform_set_into_vector(const std::unordered_set<unsigned long int>& hashes, int d)
{
    static mvs_ctx* ctx = nullptr;
    static std::mutex mu;                       // a context is single-threaded; sketch() calls from OpenMP threads
    std::vector<uint64_t> h(hashes.begin(), hashes.end());   // a set: already unique
    VectorXi vec(d);
    std::lock_guard<std::mutex> lock(mu);
    if (!ctx) {
        const int rc = mvs_create(&ctx, 0);
        if (rc != MVS_OK) throw std::runtime_error(mvs_strerror(rc));
    }
    const int rc = mvs_transform_set(ctx, h.data(), static_cast<int64_t>(h.size()), d, 0, vec.data());
    if (rc != MVS_OK) throw std::runtime_error(mvs_strerror(rc));
    return vec;
}
