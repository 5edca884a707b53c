// random_projection_b200.cpp -- link-time replacement for the reference's src/random_projection.cpp.
//
// Same signature as reference include/random_projection.h:11, so the reference's own
// project_everything.cpp / standalone_projection.cpp compile and link against it UNMODIFIED
// (INTEGRATION.md option 2).  Needs the reference's include dirs for <Eigen/Dense>; it is therefore
// not part of libmvs_b200.so and is built only where the reference tree is present.
//
// The reference calls this function from the threads of an OpenMP loop (project_everything.cpp:289-292)
// and the function is re-entrant there.  A context is single-threaded, so every calling thread gets its
// own (created on first use, kept for the life of the thread; contexts are cheap: three streams and
// grow-only buffers).  Calls of different threads then overlap on the device: one thread's upload runs
// under another one's kernel.  Threads are dealt round-robin over the visible B200s, so the unmodified
// reference driver uses every GPU of the box; MVS_SHIM_DEVICES=n limits that to the first n.
// MVS_SHIM_SERIAL=1 takes one lock around the device call (the behaviour of the first version of this
// file, kept for the A/B measurement in tests/perf_shim.py).
#include "random_projection.h"

#include "mvs_b200.h"

#include <atomic>
#include <cstdlib>
#include <mutex>
#include <stdexcept>
#include <vector>

namespace {

struct ThreadCtx {
    mvs_ctx* ctx = nullptr;
    std::vector<uint64_t> staging;              // reused between calls of this thread
    ~ThreadCtx() { mvs_destroy(ctx); }
};

int usable_devices()
{
    static const int n = [] {
        int found = 0;
        for (int dev = 0; dev < 64; ++dev) {    // mvs_create refuses anything that is not an sm_100 device
            mvs_ctx* probe = nullptr;
            if (mvs_create(&probe, dev) != MVS_OK) break;
            mvs_destroy(probe);
            ++found;
        }
        if (const char* cap = std::getenv("MVS_SHIM_DEVICES")) {
            const int c = std::atoi(cap);
            if (c > 0 && c < found) found = c;
        }
        return found;
    }();
    return n;
}

mvs_ctx* context_of_this_thread(ThreadCtx& t)
{
    if (t.ctx) return t.ctx;
    static std::atomic<unsigned> next{0};
    const int ndev = usable_devices();
    if (ndev <= 0) throw std::runtime_error(mvs_strerror(MVS_ERR_NO_DEVICE));
    const int rc = mvs_create(&t.ctx, static_cast<int>(next.fetch_add(1) % static_cast<unsigned>(ndev)));
    if (rc != MVS_OK) throw std::runtime_error(mvs_strerror(rc));
    return t.ctx;
}

}  // namespace

VectorXi transform_set_into_vector(const std::unordered_set<unsigned long int>& hashes, int d)
{
    thread_local ThreadCtx t;
    mvs_ctx* ctx = context_of_this_thread(t);
    t.staging.assign(hashes.begin(), hashes.end());          // a set: already unique
    VectorXi vec(d);
    static const bool serial = std::getenv("MVS_SHIM_SERIAL") != nullptr;
    static std::mutex mu;
    std::unique_lock<std::mutex> lock(mu, std::defer_lock);
    if (serial) lock.lock();
    const int rc = mvs_transform_set(ctx, t.staging.data(), static_cast<int64_t>(t.staging.size()), d, 0, vec.data());
    if (rc != MVS_OK) throw std::runtime_error(mvs_strerror(rc));
    return vec;
}
