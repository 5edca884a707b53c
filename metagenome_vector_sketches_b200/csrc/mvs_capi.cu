// mvs_capi.cu -- host side of the C ABI declared in include/mvs_b200.h.
//
// Replaces, for the GPU, the sample loop of sketch()
// (reference src/project_everything.cpp:287-298) and transform_set_into_vector
// (reference src/random_projection.cpp:9-26).  Planning (splitting samples into warp
// tasks), slice streaming (H2D / kernels / D2H on several streams) and the partial
// reduce live here; the device code is in mvs_kernels.cuh.
//
// There is deliberately no CPU implementation of the projection in this file: if the
// device is missing or a launch fails the call returns an error.
#include "../../include/mvs_b200.h"
#include "mvs_kernels.cuh"
#include "mvs_internal.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include <sched.h>
#include <sys/syscall.h>
#include <unistd.h>

namespace {

using mvs::ReduceItem;
using mvs::WarpTask;

constexpr int kSlots = 3;                        // slices in flight in mvs_project_batch
constexpr int kAbiVersion = 2;                   // 2: reference-order float32 norms (ref_norms) in the epilogue entry points
constexpr int kTimingRing = 64;

constexpr int kPairwiseSlots = 16;
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;                  // everything queued on this slot has finished
    cudaEvent_t uploaded = nullptr;              // the pinned plan buffers have been read by the copy engine
    bool busy = false;
    bool upload_busy = false;
    DevBuf hashes, out, partials, tasks, reduce, sumsq, out16, refnorm, packed;
    PinBuf h_tasks, h_reduce;
};

// Upload relay (option "relay_device"): the hash slices of this context travel host -> PEER device -> this device
// (cudaMemcpyAsync into a staging buffer on the peer, then cudaMemcpyPeerAsync over NVLink) instead of over this
// device's own PCIe link.  For hosts whose PCIe roots have unequal ceilings (profiles/r02l_h2d_matrix_8gpu.json: the
// four GPUs of the first root share 116 GB/s, the other four get 54 GB/s each and NVLink idles): ranks on the starved
// root route through a partner on the other one.  One staging buffer, stream and event per slice slot.
struct Relay {
    int device = -1;
    void* buf[kSlots] = {nullptr, nullptr, nullptr};
    size_t cap[kSlots] = {0, 0, 0};
    cudaStream_t stream[kSlots] = {nullptr, nullptr, nullptr};
    cudaEvent_t landed[kSlots] = {nullptr, nullptr, nullptr};
    uint64_t bytes = 0;                          // relayed by the last batch call
    int percent = 100;                           // share of the slice bytes that takes the relay (the rest goes direct)
    int64_t credit = 0;                          // error-diffusion accumulator of that share
};

struct Geometry {
    int d = 0;
    int nb = 0;                                  // 64-column blocks
    int lph_log2 = 5;
    int S = 1;                                   // hash slots per warp
    int passes = 1;                              // 32-block column passes
    uint64_t cap = 0;                            // max hashes per warp task
    uint32_t quantum = 64;                       // hashes of one loop trip of the kernel: chunk sizes are multiples of it
};

}  // namespace

struct mvs_ctx {
    int device = 0;
    int sm_count = 0;
    int last_cuda = 0;
    Slot slots[kSlots];
    // ring of CUDA-event triples (start, after projection, after reduce), one per
    // mvs_project_device call; mvs_get_timing averages the calls since the last query
    cudaEvent_t ev[kTimingRing][3] = {};
    unsigned device_turn = 0;                    // mvs_project_device calls so far (slot rotation)
    Relay relay;
    uint64_t timing_calls = 0;                   // device-path calls recorded so far
    uint64_t timing_read = 0;                    // calls already consumed by mvs_get_timing
    mvs_timing timing{};
    uint64_t launches = 0;
    uint64_t slice_hashes = 16ull << 20;         // hashes per streamed slice (128 MB)
    uint64_t slice_rows_bytes = 256ull << 20;    // output bytes per streamed slice
    uint64_t task_cap = 0;                       // experiment knob: rounds per warp task (0 = planner's choice)
    int pairwise_path = 0;                       // 0 = tcgen05 limb GEMM, 1 = IMAD GEMM, 2 = mma.sync limb GEMM
    int pairwise_pair = 0;                       // tcgen05 path: 0 = default, 1 = CTA pairs (cta_group::2, 256 x 128 tiles), 2 = single CTAs
    int pairwise_fuse = 0;                       // tcgen05 path: 0 = keep-test fused into the GEMM epilogue (default), 2 = separate count pass
    int pairwise_cluster = 0;                    // tcgen05 path: CTAs per cluster: 0 / 1 = 1 x 1 (default), 2 = 1 x 2, 4 = 2 x 2 (TMA multicast)
    int pairwise_stream = 0;                     // 1 = tcgen05 path streams limb pairs even when the planes would fit one stage
    int pairwise_limbs = 0;                      // 0 = from the operands' min/max, 1..4 = force at least that many limbs
    uint32_t* d_sink = nullptr;
    DevBuf pw_ws[kPairwiseSlots];                // grow-only device workspace of the pairwise entry points (freed by mvs_destroy
                                                 // or option "release_workspace"): no cudaMalloc / cudaFree in the steady state
    void* pw_pinned = nullptr;                   // 256 bytes of pinned host memory for small read-backs
    void* pw_stage = nullptr;                    // pinned staging ring for uploads from pageable caller memory
    size_t pw_stage_bytes = 0;
    cudaEvent_t stage_ev[2] = {nullptr, nullptr};    // "the DMA out of ring slot k has finished"
    bool stage_busy[2] = {false, false};
    unsigned stage_turn = 0;
    int stage_copy = 1;                          // 0 = memcpy into the ring, 1 = streaming stores (mvs_internal_copy_streaming)
    int stage_threads = 8;                       // host threads filling one ring slot
    int stage_uploads = 0;                       // 0 = pageable sources go through the pinned ring (default), 2 = plain cudaMemcpyAsync
};

namespace {

#define MVS_CUDA(ctx, call)                                   \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) {                              \
            (ctx)->last_cuda = static_cast<int>(e_);          \
            cudaGetLastError();                               \
            return e_ == cudaErrorMemoryAllocation ? MVS_ERR_OOM : MVS_ERR_CUDA; \
        }                                                     \
    } while (0)

int ensure_dev(mvs_ctx* ctx, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return MVS_OK;
    if (b.p) {
        MVS_CUDA(ctx, cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        want = bytes;
        MVS_CUDA(ctx, cudaMalloc(&b.p, want));
    }
    b.cap = want;
    return MVS_OK;
}

int ensure_pin(mvs_ctx* ctx, PinBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return MVS_OK;
    if (b.p) {
        MVS_CUDA(ctx, cudaFreeHost(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    const size_t want = bytes + bytes / 4 + 4096;
    MVS_CUDA(ctx, cudaMallocHost(&b.p, want));
    b.cap = want;
    return MVS_OK;
}

Geometry make_geometry(int d)
{
    Geometry g;
    g.d = d;
    g.nb = (d + 63) / 64;
    if (g.nb >= 32) {
        g.lph_log2 = 5;
        g.passes = (g.nb + 31) / 32;
    } else {
        int l = 0;
        while ((1 << l) < g.nb) ++l;
        g.lph_log2 = l;
        g.passes = 1;
    }
    g.S = 32 >> g.lph_log2;
    // chunk sizes are multiples of 64 rounds (a multiple of the kernel's loop trip of one or two Harley-Seal groups of 16
    // rounds for all S hash slots), never more than a staging tile
    g.quantum = std::min<uint32_t>(64u * static_cast<uint32_t>(g.S), static_cast<uint32_t>(mvs::tile_hashes(g.lph_log2)));
    g.cap = static_cast<uint64_t>(mvs::max_rounds(g.lph_log2)) * static_cast<uint64_t>(g.S);
    return g;
}

struct PlanCounts {
    uint64_t tasks = 0, reduce = 0, prows = 0;
};

// How a batch is cut into warp tasks.  A sample larger than the cap is split into equal chunks (multiples of the
// kernel's trip quantum); the cap is chosen per launch by choose_cap below.
struct Chunker {
    uint64_t cap, quantum;

    // chunks of a sample of n hashes: count and size
    void split(uint64_t n, uint64_t& nchunks, uint64_t& per) const
    {
        nchunks = 1;
        per = n;
        if (n > cap) {
            const uint64_t k = (n + cap - 1) / cap;
            per = (n + k - 1) / k;
            per = (per + quantum - 1) / quantum * quantum;
            nchunks = (n + per - 1) / per;
        }
    }
};

uint64_t count_tasks(const uint64_t* offsets, int64_t s0, int64_t s1, const Chunker& ch)
{
    uint64_t tasks = 0;
    for (int64_t s = s0; s < s1; ++s) {
        const uint64_t n = offsets[s + 1] - offsets[s];
        tasks += n > ch.cap ? (n + ch.cap - 1) / ch.cap : 1;             // == nchunks of split() up to quantum rounding
    }
    return tasks;
}

// The per-launch task cap (hashes).  Large launches take the largest task the bit planes allow (fewest flushes and
// partial rows).  Small launches are scored with a three-term model of the machine and the best cap on a geometric grid
// between 512 rounds and the maximum wins (launches with >= 64 CTAs per SM skip the search):
//   * SM quantisation: CTAs of kWarpsPerCta tasks are spread over the SMs and an SM shares its issue slots among its
//     resident warps, so with equal tasks the launch lasts as long as the SM that received the most CTAs:
//     ceil(CTAs / SMs) / (CTAs / SMs).  Config 3 at d = 256 (1024 x 100 000 hashes): 7168 tasks = 12.1 CTAs per SM
//     run 0.507 ms, 8192 tasks = 13.8 per SM 0.479 ms.
//   * fixed cost per task (first tile's latency, flush): ~110 rounds.
//   * tail: the last tasks of a launch run on nearly empty SMs, where a warp is no longer bound by issue slots but by
//     the latency of its two-stage hash pipeline -- about 2.5 ns per hash of the task whatever d, i.e. 0.145 of the
//     17.3 ns a warp round takes on a full sub-partition.  This is what makes 8192 tasks beat 4096 at d <= 1024
//     (same quantisation, twice the fixed cost, half the tail).
// Fitted on profiles/r02z_ksweep_c3.jsonl (config 3, k = 1..16 chunks per sample, d = 256..4096): the model's choice is the
// measured optimum at d = 256 / 512 / 1024 and within 0.1 % of it at 2048 / 4096.
uint64_t choose_cap(const uint64_t* offsets, int64_t s0, int64_t s1, const Geometry& g, int sm_count)
{
    const uint64_t cap_max = g.cap;
    const uint64_t cap_min = std::min(cap_max, std::max<uint64_t>(512ull * g.S / g.quantum * g.quantum, g.quantum));
    const uint64_t total = offsets[s1] - offsets[s0];
    if (total == 0) return cap_max;
    const double fixed = 110.0 * g.S;                                   // flush + start-up of a task, in hashes (~110 rounds)
    const double sms = static_cast<double>(sm_count);
    const double warp_rounds = static_cast<double>(total) * g.passes / g.S;          // of the whole launch
    auto score = [&](uint64_t cap, uint64_t tasks) {
        const double ctas = std::ceil(static_cast<double>(tasks) * g.passes / mvs::kWarpsPerCta);
        const double per_sm = ctas / sms;
        const double avg = static_cast<double>(total) / static_cast<double>(tasks);
        const double tail = 0.145 * avg * (4.0 * sms) / warp_rounds;    // one task at 2.5 ns per hash over the launch's time
        return std::ceil(per_sm) / per_sm * (1.0 + fixed / avg) + tail; // relative to perfectly divisible work
    };
    Chunker ch{cap_max, g.quantum};
    const uint64_t tasks_max = count_tasks(offsets, s0, s1, ch);
    if (static_cast<double>(tasks_max) * g.passes / mvs::kWarpsPerCta >= 64.0 * sms) return cap_max;
    uint64_t best_cap = cap_max;
    double best = score(cap_max, tasks_max);
    constexpr int kGrid = 32;
    const double ratio = static_cast<double>(cap_max) / static_cast<double>(cap_min);
    uint64_t prev = cap_max;
    for (int i = 1; i <= kGrid; ++i) {
        uint64_t cap = static_cast<uint64_t>(static_cast<double>(cap_max) / std::pow(ratio, static_cast<double>(i) / kGrid));
        cap = std::max(cap_min, cap / g.quantum * g.quantum);
        if (cap == prev) continue;
        prev = cap;
        ch.cap = cap;
        const double sc = score(cap, count_tasks(offsets, s0, s1, ch));
        if (sc < best * 0.995) {                                        // prefer the larger cap on ties
            best = sc;
            best_cap = cap;
        }
    }
    return best_cap;
}

// First pass: how many tasks / reduce items / partial rows does [s0, s1) need?
int count_plan(const uint64_t* offsets, int64_t s0, int64_t s1, const Geometry& g, const Chunker& ch, PlanCounts& pc)
{
    pc = PlanCounts{};
    for (int64_t s = s0; s < s1; ++s) {
        if (offsets[s + 1] < offsets[s]) return MVS_ERR_INVALID_ARG;
        const uint64_t n = offsets[s + 1] - offsets[s];
        if (n >= (1ull << 31)) return MVS_ERR_TOO_LARGE;
        uint64_t nchunks, per;
        ch.split(n, nchunks, per);
        pc.tasks += nchunks * static_cast<uint64_t>(g.passes);
        if (nchunks > 1) {
            pc.reduce += 1;
            pc.prows += nchunks;
        }
    }
    if (pc.tasks >= 0xFFFFFFF0ull || pc.prows >= 0xFFFFFFF0ull) return MVS_ERR_TOO_LARGE;
    return MVS_OK;
}

// Second pass: emit.  `hash_base` is subtracted from the CSR
// offsets (index of the first hash present in the device array), `row_base` from the sample index (first output row).
void emit_plan(const uint64_t* offsets, int64_t s0, int64_t s1, uint64_t hash_base, int64_t row_base,
               const Geometry& g, const Chunker& ch, WarpTask* tasks, ReduceItem* reduce)
{
    uint64_t ti = 0, ri = 0, prow = 0;
    for (int64_t s = s0; s < s1; ++s) {
        const uint64_t n = offsets[s + 1] - offsets[s];
        const uint64_t b0 = offsets[s] - hash_base;
        uint64_t nchunks, per;
        ch.split(n, nchunks, per);
        const uint32_t row = static_cast<uint32_t>(s - row_base);
        for (uint64_t c = 0; c < nchunks; ++c) {
            const uint64_t cb = c * per;
            const uint64_t cn = std::min<uint64_t>(per, n - cb);
            for (int p = 0; p < g.passes; ++p) {
                WarpTask& t = tasks[ti++];
                t.begin = b0 + cb;
                t.count = static_cast<uint32_t>(cn);
                t.sample = row;
                t.pass = static_cast<uint32_t>(p);
                t.prow = nchunks > 1 ? static_cast<uint32_t>(prow + c) : mvs::kDirect;
                t.pad0 = 0;
                t.pad1 = 0;
            }
        }
        if (nchunks > 1) {
            ReduceItem& r = reduce[ri++];
            r.sample = row;
            r.prow = static_cast<uint32_t>(prow);
            r.nchunks = static_cast<uint32_t>(nchunks);
            r.n = static_cast<uint32_t>(n);
            prow += nchunks;
        }
    }
}

template <int L>
cudaError_t launch_project_t(cudaStream_t st, uint32_t ntasks, const uint64_t* d_hashes, const WarpTask* d_tasks,
                             int d, uint64_t seed, int32_t* d_out, int32_t* d_partials, int vec4)
{
    const uint32_t grid = (ntasks + mvs::kWarpsPerCta - 1) / mvs::kWarpsPerCta;
    // per warp: kStages raw tiles (8 B per hash) + one cooked tile (16 B per hash) = 8 KB; 32 KB per CTA
    const size_t smem = static_cast<size_t>(mvs::kWarpsPerCta) * mvs::tile_hashes(L) *
                        (mvs::kStages * sizeof(uint64_t) + sizeof(ulonglong2));
    if (smem > 48 * 1024) {                      // only L = 0 (d <= 64): 64 KB
        cudaError_t e = cudaFuncSetAttribute(mvs::project_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return e;
    }
    mvs::project_kernel<L><<<grid, mvs::kWarpsPerCta * 32, smem, st>>>(d_hashes, d_tasks, ntasks, d,
                                                                     seed + mvs::kGolden, d_out, d_partials, vec4,
                                                                     mvs::kFold);
    return cudaGetLastError();
}

cudaError_t launch_project(const Geometry& g, cudaStream_t st, uint32_t ntasks, const uint64_t* d_hashes,
                           const WarpTask* d_tasks, uint64_t seed, int32_t* d_out, int32_t* d_partials)
{
    const int vec4 = ((g.d & 3) == 0) && ((reinterpret_cast<uintptr_t>(d_out) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(d_partials) & 15) == 0);
    switch (g.lph_log2) {
        case 0: return launch_project_t<0>(st, ntasks, d_hashes, d_tasks, g.d, seed, d_out, d_partials, vec4);
        case 1: return launch_project_t<1>(st, ntasks, d_hashes, d_tasks, g.d, seed, d_out, d_partials, vec4);
        case 2: return launch_project_t<2>(st, ntasks, d_hashes, d_tasks, g.d, seed, d_out, d_partials, vec4);
        case 3: return launch_project_t<3>(st, ntasks, d_hashes, d_tasks, g.d, seed, d_out, d_partials, vec4);
        case 4: return launch_project_t<4>(st, ntasks, d_hashes, d_tasks, g.d, seed, d_out, d_partials, vec4);
        default: return launch_project_t<5>(st, ntasks, d_hashes, d_tasks, g.d, seed, d_out, d_partials, vec4);
    }
}

// Plan [s0, s1) into the slot's pinned buffers, upload, launch projection (+ reduce).
// d_hashes holds the hashes from index `hash_base` on; d_out row 0 = sample `s0`
// when rows_local, else sample 0.
// Plan [s0, s1) into the slot's pinned buffers and copy the plan to the device on stream `up`.  The copy needs nothing
// from the device, so callers enqueue it BEFORE the (much larger) hash upload of the same slice: the host-to-device
// copy queue then never holds a small copy that waits for a kernel while the next slice's hashes queue up behind it.
int plan_range(mvs_ctx* ctx, Slot& sl, const uint64_t* offsets, int64_t s0, int64_t s1, uint64_t hash_base, int64_t row_base,
               const Geometry& g, cudaStream_t up, PlanCounts& pc)
{
    Chunker ch{choose_cap(offsets, s0, s1, g, ctx->sm_count), g.quantum};
#ifdef MVS_EXPERIMENTS
    if (ctx->task_cap) ch.cap = std::min<uint64_t>(g.cap, (ctx->task_cap + 63) / 64 * 64 * static_cast<uint64_t>(g.S));   // fixed caps for the sweep tools
#endif
    int rc = count_plan(offsets, s0, s1, g, ch, pc);
    if (rc != MVS_OK) return rc;
    if (pc.tasks == 0) return MVS_OK;
    if ((rc = ensure_pin(ctx, sl.h_tasks, pc.tasks * sizeof(WarpTask))) != MVS_OK) return rc;
    if ((rc = ensure_pin(ctx, sl.h_reduce, (pc.reduce + 1) * sizeof(ReduceItem))) != MVS_OK) return rc;
    if ((rc = ensure_dev(ctx, sl.tasks, pc.tasks * sizeof(WarpTask))) != MVS_OK) return rc;
    if ((rc = ensure_dev(ctx, sl.reduce, (pc.reduce + 1) * sizeof(ReduceItem))) != MVS_OK) return rc;
    if ((rc = ensure_dev(ctx, sl.partials, (pc.prows + 1) * static_cast<size_t>(g.d) * sizeof(int32_t))) != MVS_OK)
        return rc;
    emit_plan(offsets, s0, s1, hash_base, row_base, g, ch, static_cast<WarpTask*>(sl.h_tasks.p),
              static_cast<ReduceItem*>(sl.h_reduce.p));
    MVS_CUDA(ctx, cudaMemcpyAsync(sl.tasks.p, sl.h_tasks.p, pc.tasks * sizeof(WarpTask), cudaMemcpyHostToDevice, up));
    if (pc.reduce)
        MVS_CUDA(ctx, cudaMemcpyAsync(sl.reduce.p, sl.h_reduce.p, pc.reduce * sizeof(ReduceItem),
                                      cudaMemcpyHostToDevice, up));
    MVS_CUDA(ctx, cudaEventRecord(sl.uploaded, up));
    sl.upload_busy = true;
    ctx->timing.warp_tasks = pc.tasks;
    ctx->timing.partial_rows = pc.prows;
    return MVS_OK;
}

// Launch projection (+ reduce) of a planned range on `st`.  d_hashes holds the hashes from index `hash_base` (of the
// plan) on; d_out row 0 = sample `row_base`.  `up`: the stream the plan was copied on.
int launch_range(mvs_ctx* ctx, Slot& sl, const PlanCounts& pc, const uint64_t* d_hashes, const Geometry& g, uint64_t seed,
                 int32_t* d_out, cudaStream_t st, bool record_timing, cudaStream_t up)
{
    if (pc.tasks == 0) return MVS_OK;
    if (up != st) MVS_CUDA(ctx, cudaStreamWaitEvent(st, sl.uploaded, 0));
    cudaEvent_t* ev = ctx->ev[ctx->timing_calls % kTimingRing];
    if (record_timing) MVS_CUDA(ctx, cudaEventRecord(ev[0], st));
    MVS_CUDA(ctx, launch_project(g, st, static_cast<uint32_t>(pc.tasks), d_hashes,
                                 static_cast<const WarpTask*>(sl.tasks.p), seed, d_out,
                                 static_cast<int32_t*>(sl.partials.p)));
    ctx->launches += 1;
    if (record_timing) MVS_CUDA(ctx, cudaEventRecord(ev[1], st));
    if (pc.reduce) {
        const unsigned gy = static_cast<unsigned>(std::min<int64_t>((g.d + 255) / 256, 65535));
        dim3 grid(static_cast<unsigned>(pc.reduce), gy);
        mvs::reduce_partials_kernel<<<grid, 256, 0, st>>>(static_cast<const ReduceItem*>(sl.reduce.p),
                                                        static_cast<uint32_t>(pc.reduce), g.d,
                                                        static_cast<const int32_t*>(sl.partials.p), d_out);
        MVS_CUDA(ctx, cudaGetLastError());
        ctx->launches += 1;
    }
    if (record_timing) {
        MVS_CUDA(ctx, cudaEventRecord(ev[2], st));
        ctx->timing_calls += 1;
    }
    return MVS_OK;
}

void relay_release(mvs_ctx* ctx)
{
    Relay& r = ctx->relay;
    if (r.device >= 0) {
        cudaSetDevice(r.device);
        for (int k = 0; k < kSlots; ++k) {
            if (r.stream[k]) { cudaStreamSynchronize(r.stream[k]); cudaStreamDestroy(r.stream[k]); }
            if (r.landed[k]) cudaEventDestroy(r.landed[k]);
            if (r.buf[k]) cudaFree(r.buf[k]);
            r.stream[k] = nullptr; r.landed[k] = nullptr; r.buf[k] = nullptr; r.cap[k] = 0;
        }
        cudaSetDevice(ctx->device);
        cudaGetLastError();
    }
    r.device = -1;
}

int relay_setup(mvs_ctx* ctx, int peer)
{
    relay_release(ctx);
    if (peer < 0) return MVS_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || peer >= ndev || peer == ctx->device) { cudaGetLastError(); return MVS_ERR_INVALID_ARG; }
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, ctx->device, peer) != cudaSuccess || !can) { cudaGetLastError(); return MVS_ERR_INVALID_ARG; }
    auto enable = [](int from, int to) {
        cudaSetDevice(from);
        const cudaError_t e = cudaDeviceEnablePeerAccess(to, 0);
        if (e != cudaSuccess) cudaGetLastError();
        return e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled;
    };
    const bool ok = enable(ctx->device, peer) && enable(peer, ctx->device);
    cudaError_t e = cudaSetDevice(peer);
    for (int k = 0; k < kSlots && e == cudaSuccess; ++k) {
        e = cudaStreamCreateWithFlags(&ctx->relay.stream[k], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->relay.landed[k], cudaEventDisableTiming);
    }
    cudaSetDevice(ctx->device);
    ctx->relay.device = peer;
    if (!ok || e != cudaSuccess) {
        ctx->last_cuda = static_cast<int>(e);
        relay_release(ctx);
        return MVS_ERR_CUDA;
    }
    return MVS_OK;
}

// One slice of hashes host -> device on the slot's stream: over the relay when it is on and the source is pinned
// (the DMA engine of the peer reads it), else the direct path (which also stages pageable sources).
int upload_slice(mvs_ctx* ctx, Slot& sl, int k, void* dst, const void* src, size_t bytes)
{
    Relay& r = ctx->relay;
    bool pinned = false;
    if (r.device >= 0 && bytes >= (1u << 20)) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, src) == cudaSuccess) pinned = at.type == cudaMemoryTypeHost;
        else cudaGetLastError();
    }
    if (pinned && r.percent < 100) {             // split by bytes: this slice takes the relay when its share is due
        r.credit += static_cast<int64_t>(bytes / 1024) * r.percent;
        if (r.credit >= static_cast<int64_t>(bytes / 1024) * 50) r.credit -= static_cast<int64_t>(bytes / 1024) * 100;
        else pinned = false;
    }
    if (!pinned) return mvs_internal_upload(ctx, sl.stream, dst, src, bytes);
    MVS_CUDA(ctx, cudaSetDevice(r.device));
    cudaError_t e = cudaSuccess;
    if (r.cap[k] < bytes) {
        if (r.buf[k]) e = cudaFree(r.buf[k]);
        r.buf[k] = nullptr;
        r.cap[k] = 0;
        if (e == cudaSuccess) e = cudaMalloc(&r.buf[k], bytes + bytes / 8 + 256);
        if (e == cudaSuccess) r.cap[k] = bytes + bytes / 8 + 256;
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(r.buf[k], src, bytes, cudaMemcpyHostToDevice, r.stream[k]);
    if (e == cudaSuccess) e = cudaEventRecord(r.landed[k], r.stream[k]);
    const cudaError_t back = cudaSetDevice(ctx->device);
    if (e == cudaSuccess) e = back;
    MVS_CUDA(ctx, e);
    MVS_CUDA(ctx, cudaStreamWaitEvent(sl.stream, r.landed[k], 0));
    MVS_CUDA(ctx, cudaMemcpyPeerAsync(dst, ctx->device, r.buf[k], r.device, bytes, sl.stream));
    r.bytes += bytes;
    return MVS_OK;
}

int wait_slot(mvs_ctx* ctx, Slot& sl)
{
    if (sl.busy) {
        MVS_CUDA(ctx, cudaEventSynchronize(sl.done));
        sl.busy = false;
        sl.upload_busy = false;
    }
    return MVS_OK;
}

int wait_upload(mvs_ctx* ctx, Slot& sl)
{
    if (sl.upload_busy) {
        MVS_CUDA(ctx, cudaEventSynchronize(sl.uploaded));
        sl.upload_busy = false;
    }
    return MVS_OK;
}

}  // namespace

// ======================================================================================
extern "C" {

int mvs_abi_version(void) { return kAbiVersion; }

const char* mvs_strerror(int status)
{
    switch (status) {
        case MVS_OK: return "ok";
        case MVS_ERR_INVALID_ARG: return "invalid argument";
        case MVS_ERR_NO_DEVICE: return "no usable sm_100 CUDA device (there is no CPU fallback)";
        case MVS_ERR_CUDA: return "CUDA call failed";
        case MVS_ERR_OOM: return "out of memory";
        case MVS_ERR_TOO_LARGE: return "sample or batch too large (>= 2^31 hashes in one sample, or >= 2^32 tasks)";
        case MVS_ERR_IO: return "I/O error";
        case MVS_ERR_SCRATCH_LIMIT:
            return "the call needs more scratch than MVS_STATIC_SCRATCH_BYTES (mvs_project_device_static): use mvs_project_device";
        default: return "unknown mvs status";
    }
}

int mvs_last_cuda_error(const mvs_ctx* ctx) { return ctx ? ctx->last_cuda : 0; }

}  // extern "C"  (the internal accessors of mvs_internal.h have C++ linkage)
int mvs_internal_device(const mvs_ctx* ctx) { return ctx->device; }
cudaStream_t mvs_internal_stream(mvs_ctx* ctx) { return ctx->slots[0].stream; }
void mvs_internal_set_cuda_error(mvs_ctx* ctx, cudaError_t e) { ctx->last_cuda = static_cast<int>(e); }
void mvs_internal_count_launches(mvs_ctx* ctx, uint64_t n) { ctx->launches += n; }
int mvs_internal_pairwise_path(const mvs_ctx* ctx) { return ctx->pairwise_path; }
int mvs_internal_pairwise_limbs(const mvs_ctx* ctx) { return ctx->pairwise_limbs; }
int mvs_internal_pairwise_stream(const mvs_ctx* ctx) { return ctx->pairwise_stream; }
int mvs_internal_pairwise_cluster(const mvs_ctx* ctx) { return ctx->pairwise_cluster; }
int mvs_internal_pairwise_pair(const mvs_ctx* ctx) { return ctx->pairwise_pair; }
int mvs_internal_pairwise_fuse(const mvs_ctx* ctx) { return ctx->pairwise_fuse; }
int mvs_internal_workspace(mvs_ctx* ctx, int slot, size_t bytes, void** out)
{
    if (slot < 0 || slot >= kPairwiseSlots) return MVS_ERR_INVALID_ARG;
    const int rc = ensure_dev(ctx, ctx->pw_ws[slot], bytes ? bytes : 16);
    *out = rc == MVS_OK ? ctx->pw_ws[slot].p : nullptr;
    return rc;
}
void* mvs_internal_pinned_scratch(mvs_ctx* ctx)
{
    if (!ctx->pw_pinned && cudaHostAlloc(&ctx->pw_pinned, 256, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        ctx->pw_pinned = nullptr;
    }
    return ctx->pw_pinned;
}
void* mvs_internal_pinned_stage(mvs_ctx* ctx, size_t bytes)
{
    if (ctx->pw_stage_bytes < bytes) {
        if (ctx->pw_stage) cudaFreeHost(ctx->pw_stage);
        ctx->pw_stage = nullptr;
        ctx->pw_stage_bytes = 0;
        if (cudaHostAlloc(&ctx->pw_stage, bytes, cudaHostAllocDefault) != cudaSuccess) {
            cudaGetLastError();
            ctx->pw_stage = nullptr;
            return nullptr;
        }
        ctx->pw_stage_bytes = bytes;
    }
    return ctx->pw_stage;
}
// Host -> device copy of a caller buffer.  cudaMemcpyAsync from PAGEABLE memory goes through the driver's
// single-threaded staging copy (~8 GB/s on this box) and blocks the host; here such a buffer is cut into 8 MB chunks
// that several host threads copy into a two-slot pinned ring while the previous chunk's DMA runs (PCIe rate).
// Pinned / registered sources are copied directly.  Asynchronous on `st` like cudaMemcpyAsync.
int mvs_internal_upload(mvs_ctx* ctx, cudaStream_t st, void* dst, const void* src, size_t bytes)
{
    constexpr size_t kChunk = 8u << 20;
    bool pageable = false;
#ifdef MVS_EXPERIMENTS                           // A/B of round 1: the driver's own staging copy, memcpy instead of streaming stores
    const bool ring_allowed = ctx->stage_uploads != 2, streaming = ctx->stage_copy == 1;
#else
    constexpr bool ring_allowed = true, streaming = true;
#endif
    if (bytes >= (2u << 20) && ring_allowed) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, src) == cudaSuccess) pageable = at.type == cudaMemoryTypeUnregistered;
        else cudaGetLastError();
    }
    char* stage = pageable ? static_cast<char*>(mvs_internal_pinned_stage(ctx, 2 * kChunk)) : nullptr;
    if (stage)
        for (auto& e : ctx->stage_ev)
            if (!e && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
                cudaGetLastError();
                e = nullptr;
                stage = nullptr;
            }
    if (!stage) {
        MVS_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
        return MVS_OK;
    }
    for (size_t off = 0; off < bytes; off += kChunk) {
        const size_t n = std::min(kChunk, bytes - off);
        const unsigned k = ctx->stage_turn++ & 1u;
        char* slot = stage + k * kChunk;
        if (ctx->stage_busy[k]) MVS_CUDA(ctx, cudaEventSynchronize(ctx->stage_ev[k]));
        const char* from = static_cast<const char*>(src) + off;
        const int kParts = ctx->stage_threads;
#pragma omp parallel for num_threads(kParts) schedule(static)
        for (int p = 0; p < kParts; ++p) {
            const size_t a = (n * p / kParts) & ~size_t(63), b = p + 1 == kParts ? n : (n * (p + 1) / kParts) & ~size_t(63);
            if (streaming) mvs_internal_copy_streaming(slot + a, from + a, b - a);
            else std::memcpy(slot + a, from + a, b - a);
        }
        MVS_CUDA(ctx, cudaMemcpyAsync(static_cast<char*>(dst) + off, slot, n, cudaMemcpyHostToDevice, st));
        MVS_CUDA(ctx, cudaEventRecord(ctx->stage_ev[k], st));
        ctx->stage_busy[k] = true;
    }
    return MVS_OK;
}
static void release_workspace(mvs_ctx* ctx)
{
    if (ctx->pw_stage) cudaFreeHost(ctx->pw_stage);
    ctx->pw_stage = nullptr;
    ctx->pw_stage_bytes = 0;
    ctx->stage_busy[0] = ctx->stage_busy[1] = false;
    for (auto& b : ctx->pw_ws) {
        if (b.p) cudaFree(b.p);
        b.p = nullptr;
        b.cap = 0;
    }
}
extern "C" {

int mvs_create(mvs_ctx** out, int device)
{
    if (!out) return MVS_ERR_INVALID_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
        cudaGetLastError();
        return MVS_ERR_NO_DEVICE;
    }
    cudaDeviceProp prop{};
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        cudaGetLastError();
        return MVS_ERR_NO_DEVICE;
    }
    if (prop.major != 10) return MVS_ERR_NO_DEVICE;          // sm_100a cubin only
    mvs_ctx* ctx = new (std::nothrow) mvs_ctx();
    if (!ctx) return MVS_ERR_OOM;
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    int rc = MVS_OK;
    auto chk = [&](cudaError_t e) {
        if (e != cudaSuccess && rc == MVS_OK) {
            ctx->last_cuda = static_cast<int>(e);
            rc = MVS_ERR_CUDA;
        }
    };
    chk(cudaSetDevice(device));
    for (auto& sl : ctx->slots) {
        chk(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
        chk(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
        chk(cudaEventCreateWithFlags(&sl.uploaded, cudaEventDisableTiming));
    }
    for (auto& e3 : ctx->ev)
        for (auto& e : e3) chk(cudaEventCreate(&e));
    chk(cudaMalloc(reinterpret_cast<void**>(&ctx->d_sink), 256));
    if (rc != MVS_OK) {
        mvs_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return MVS_OK;
}

void mvs_destroy(mvs_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    for (auto& sl : ctx->slots) {
        if (sl.stream) cudaStreamSynchronize(sl.stream);
        for (DevBuf* b : {&sl.hashes, &sl.out, &sl.partials, &sl.tasks, &sl.reduce, &sl.sumsq, &sl.out16, &sl.refnorm, &sl.packed})
            if (b->p) cudaFree(b->p);
        for (PinBuf* b : {&sl.h_tasks, &sl.h_reduce})
            if (b->p) cudaFreeHost(b->p);
        if (sl.done) cudaEventDestroy(sl.done);
        if (sl.uploaded) cudaEventDestroy(sl.uploaded);
        if (sl.stream) cudaStreamDestroy(sl.stream);
    }
    for (auto& e3 : ctx->ev)
        for (auto& e : e3)
            if (e) cudaEventDestroy(e);
    if (ctx->d_sink) cudaFree(ctx->d_sink);
    relay_release(ctx);
    release_workspace(ctx);
    for (auto& e : ctx->stage_ev)
        if (e) cudaEventDestroy(e);
    if (ctx->pw_pinned) cudaFreeHost(ctx->pw_pinned);
    cudaGetLastError();
    delete ctx;
}

int mvs_project_device(mvs_ctx* ctx, const uint64_t* d_hashes, const uint64_t* h_offsets, int64_t n_samples, int d,
                       uint64_t seed, int32_t* d_out, void* stream)
{
    if (!ctx || !h_offsets || n_samples < 0 || d <= 0 || (!d_out && n_samples > 0)) return MVS_ERR_INVALID_ARG;
    if (n_samples == 0) return MVS_OK;
    if (n_samples >= 0xFFFFFFF0ll) return MVS_ERR_TOO_LARGE;
    if (!d_hashes && h_offsets[n_samples] > h_offsets[0]) return MVS_ERR_INVALID_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    // Calls rotate over the slots' plan buffers (pinned + device): the plan of call i is copied on the slot's own stream
    // while the kernel of call i-1 still runs on the caller's stream, so back-to-back calls leave no copy-sized gap
    // between kernels.  A slot is reused every kSlots calls: its previous plan copy must have left the pinned buffers
    // (host wait), and its previous kernels must have finished with the device copy (stream wait on `done`).
    Slot& sl = ctx->slots[ctx->device_turn++ % kSlots];
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->slots[0].stream;
    int rc = stream ? wait_upload(ctx, sl) : wait_slot(ctx, sl);
    if (rc != MVS_OK) return rc;
    if (sl.busy && sl.stream != st) MVS_CUDA(ctx, cudaStreamWaitEvent(sl.stream, sl.done, 0));
    const Geometry g = make_geometry(d);
    PlanCounts pc;
    if ((rc = plan_range(ctx, sl, h_offsets, 0, n_samples, 0, 0, g, sl.stream, pc)) != MVS_OK) return rc;
    if ((rc = launch_range(ctx, sl, pc, d_hashes, g, seed, d_out, st, true, sl.stream)) != MVS_OK) return rc;
    MVS_CUDA(ctx, cudaEventRecord(sl.done, st));
    sl.busy = true;
    if (!stream) MVS_CUDA(ctx, cudaStreamSynchronize(st));
    return MVS_OK;
}

}  // extern "C"

// ---- mvs_project_device_static: the device path without any allocation -------------------------------------------
// No counterpart in the reference; added for callers that may not allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The three scratch areas of a projection
// launch -- warp tasks, reduce items, partial rows (what Slot::tasks / reduce / partials hold on the allocating paths)
// -- live in ONE file-scope device array of MVS_STATIC_SCRATCH_BYTES that every GPU which loads this library holds
// until the process exits.  In host code `g_static_scratch` is only the host-side stand-in of the symbol, never a device
// address: static_scratch_base() resolves it with cudaGetSymbolAddress once per device, with that device current, and
// only that pointer reaches cudaMemcpyAsync and the kernels.
static __device__ __align__(256) unsigned char g_static_scratch[MVS_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;                                   // guards the two tables below (host side only)
void* g_static_scratch_base[kStaticMaxDevices] = {};         // device address of g_static_scratch, per device
// host image of the plan (tasks + reduce items) of the call being issued: a pageable cudaMemcpyAsync has left it when
// it returns (the driver stages it), so one image serves all calls; .bss, pages are touched only as far as plans reach
alignas(256) unsigned char g_static_plan_image[MVS_STATIC_SCRATCH_BYTES];

constexpr size_t align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

// device must be current; caller holds g_static_mutex
int static_scratch_base(mvs_ctx* ctx, unsigned char** base)
{
    if (ctx->device < 0 || ctx->device >= kStaticMaxDevices) return MVS_ERR_INVALID_ARG;
    void*& slot = g_static_scratch_base[ctx->device];
    if (!slot) {
        void* p = nullptr;
        MVS_CUDA(ctx, cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255) != 0) return MVS_ERR_CUDA;
        slot = p;
    }
    *base = static_cast<unsigned char*>(slot);
    return MVS_OK;
}

struct StaticLayout {
    size_t reduce_off = 0, partials_off = 0, total = 0;
};

StaticLayout static_layout(const PlanCounts& pc, int d)
{
    StaticLayout l;
    l.reduce_off = align256(pc.tasks * sizeof(WarpTask));
    l.partials_off = align256(l.reduce_off + pc.reduce * sizeof(ReduceItem));
    l.total = l.partials_off + pc.prows * static_cast<size_t>(d) * sizeof(int32_t);
    return l;
}

}  // namespace

extern "C" {

int mvs_project_device_static(mvs_ctx* ctx, const uint64_t* d_hashes, const uint64_t* h_offsets, int64_t n_samples,
                              int d, uint64_t seed, int32_t* d_out, void* stream, uint64_t* scratch_needed)
{
    if (scratch_needed) *scratch_needed = 0;
    if (!ctx || !h_offsets || n_samples < 0 || d <= 0 || (!d_out && n_samples > 0)) return MVS_ERR_INVALID_ARG;
    if (n_samples == 0) return MVS_OK;
    if (n_samples >= 0xFFFFFFF0ll) return MVS_ERR_TOO_LARGE;
    if (!d_hashes && h_offsets[n_samples] > h_offsets[0]) return MVS_ERR_INVALID_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->slots[0].stream;
    const Geometry g = make_geometry(d);

    // the planner's cap first (same tasks as mvs_project_device); if its partial rows do not fit, the largest cap
    Chunker ch{choose_cap(h_offsets, 0, n_samples, g, ctx->sm_count), g.quantum};
    PlanCounts pc;
    int rc = count_plan(h_offsets, 0, n_samples, g, ch, pc);
    if (rc != MVS_OK) return rc;
    StaticLayout lay = static_layout(pc, d);
    if (lay.total > MVS_STATIC_SCRATCH_BYTES && ch.cap < g.cap) {
        ch.cap = g.cap;
        if ((rc = count_plan(h_offsets, 0, n_samples, g, ch, pc)) != MVS_OK) return rc;
        lay = static_layout(pc, d);
    }
    if (scratch_needed) *scratch_needed = lay.total;
    if (lay.total > MVS_STATIC_SCRATCH_BYTES) return MVS_ERR_SCRATCH_LIMIT;
    if (pc.tasks == 0) return MVS_OK;

    std::lock_guard<std::mutex> lock(g_static_mutex);
    unsigned char* base = nullptr;
    if ((rc = static_scratch_base(ctx, &base)) != MVS_OK) return rc;
    emit_plan(h_offsets, 0, n_samples, 0, 0, g, ch, reinterpret_cast<WarpTask*>(g_static_plan_image),
              reinterpret_cast<ReduceItem*>(g_static_plan_image + lay.reduce_off));
    // one copy: tasks, padding, reduce items.  Stream order puts it after the previous call's kernels on `st`.
    const size_t plan_bytes = lay.reduce_off + pc.reduce * sizeof(ReduceItem);
    MVS_CUDA(ctx, cudaMemcpyAsync(base, g_static_plan_image, plan_bytes, cudaMemcpyHostToDevice, st));
    // Nothing in the scratch has to start at zero: the projection kernel STORES every column of every partial row of
    // this plan before the reduce kernel reads it, and rows of earlier calls are never read.
    Slot view;                                               // pointers into the static array, nothing owned
    view.tasks.p = base;
    view.reduce.p = base + lay.reduce_off;
    view.partials.p = base + lay.partials_off;
    if ((rc = launch_range(ctx, view, pc, d_hashes, g, seed, d_out, st, false, st)) != MVS_OK) return rc;
    if (!stream) MVS_CUDA(ctx, cudaStreamSynchronize(st));
    return MVS_OK;
}

// hashes: the CSR payload as u64 (packed56 == false) or as 7 little-endian bytes per hash (packed56 == true)
static int sketch_batch_impl(mvs_ctx* ctx, const void* hashes_any, bool packed56, const uint64_t* offsets, int64_t n_samples,
                             int d, uint64_t seed, int32_t* out, int16_t* out16, uint64_t* sumsq, float* ref_norms)
{
    const uint64_t* hashes = packed56 ? nullptr : static_cast<const uint64_t*>(hashes_any);
    const unsigned char* packed = packed56 ? static_cast<const unsigned char*>(hashes_any) : nullptr;
    if (!ctx || !offsets || n_samples < 0 || d <= 0 || (!out && !out16 && !sumsq && !ref_norms && n_samples > 0))
        return MVS_ERR_INVALID_ARG;
    if (n_samples == 0) return MVS_OK;
    if (n_samples >= 0xFFFFFFF0ll) return MVS_ERR_TOO_LARGE;
    if (!hashes_any && offsets[n_samples] > offsets[0]) return MVS_ERR_INVALID_ARG;
    for (int64_t s = 0; s < n_samples; ++s)
        if (offsets[s + 1] < offsets[s]) return MVS_ERR_INVALID_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    const Geometry g = make_geometry(d);
    const uint64_t row_bytes = static_cast<uint64_t>(d) * sizeof(int32_t);
    const int64_t max_rows = std::max<int64_t>(1, static_cast<int64_t>(ctx->slice_rows_bytes / row_bytes));
    ctx->timing.h2d_bytes = 0;
    ctx->timing.d2h_bytes = 0;
    ctx->relay.bytes = 0;

    int rc = MVS_OK;
    int64_t s0 = 0;
    int turn = 0;
    while (s0 < n_samples && rc == MVS_OK) {
        int64_t s1 = s0 + 1;
        while (s1 < n_samples && s1 - s0 < max_rows && offsets[s1 + 1] - offsets[s0] <= ctx->slice_hashes) ++s1;
        const uint64_t h0 = offsets[s0];
        const uint64_t nh = offsets[s1] - h0;
        const int64_t rows = s1 - s0;
        Slot& sl = ctx->slots[turn % kSlots];
        ++turn;
        if ((rc = wait_slot(ctx, sl)) != MVS_OK) break;
        if ((rc = ensure_dev(ctx, sl.hashes, (nh + 2) * sizeof(uint64_t))) != MVS_OK) break;
        if ((rc = ensure_dev(ctx, sl.out, static_cast<size_t>(rows) * row_bytes)) != MVS_OK) break;
        if (sumsq && (rc = ensure_dev(ctx, sl.sumsq, static_cast<size_t>(rows) * sizeof(uint64_t))) != MVS_OK) break;
        if (out16 && (rc = ensure_dev(ctx, sl.out16, static_cast<size_t>(rows) * row_bytes / 2)) != MVS_OK) break;
        if (ref_norms && (rc = ensure_dev(ctx, sl.refnorm, static_cast<size_t>(rows) * sizeof(float))) != MVS_OK) break;
        PlanCounts pc;                           // plan first: its small copy must not queue behind a kernel (see plan_range)
        if ((rc = plan_range(ctx, sl, offsets, s0, s1, h0, s0, g, sl.stream, pc)) != MVS_OK) break;
        if (nh && !packed56) {
            if ((rc = upload_slice(ctx, sl, (turn - 1) % kSlots, sl.hashes.p, hashes + h0, nh * sizeof(uint64_t))) != MVS_OK) break;
            ctx->timing.h2d_bytes += nh * sizeof(uint64_t);
        } else if (nh) {                         // 7 bytes per hash over the link, widened on the device
            if ((rc = ensure_dev(ctx, sl.packed, nh * 7 + 16)) != MVS_OK) break;
            if ((rc = upload_slice(ctx, sl, (turn - 1) % kSlots, sl.packed.p, packed + 7 * h0, nh * 7)) != MVS_OK) break;
            ctx->timing.h2d_bytes += nh * 7;
            const uint64_t grid = (nh + 255) / 256;
            mvs::unpack56_kernel<<<static_cast<unsigned>(grid), 256, 0, sl.stream>>>(
                static_cast<const uint64_t*>(sl.packed.p), nh, static_cast<uint64_t*>(sl.hashes.p));
            cudaError_t ue = cudaGetLastError();
            if (ue != cudaSuccess) { ctx->last_cuda = ue; rc = MVS_ERR_CUDA; break; }
            ctx->launches += 1;
        }
        rc = launch_range(ctx, sl, pc, static_cast<const uint64_t*>(sl.hashes.p), g, seed, static_cast<int32_t*>(sl.out.p),
                          sl.stream, false, sl.stream);
        if (rc != MVS_OK) break;
        ctx->timing.h2d_bytes += ctx->timing.warp_tasks * sizeof(WarpTask);
        cudaError_t e = cudaSuccess;
        if (out16 || sumsq || ref_norms) {       // fused epilogue of sketch(): norms + int16 saturation
            rc = mvs_finalize_device(ctx, static_cast<const int32_t*>(sl.out.p), rows, d,
                                     sumsq ? static_cast<uint64_t*>(sl.sumsq.p) : nullptr,
                                     out16 ? static_cast<int16_t*>(sl.out16.p) : nullptr,
                                     ref_norms ? static_cast<float*>(sl.refnorm.p) : nullptr, sl.stream);
            if (rc != MVS_OK) break;
        }
        if (out) {
            e = cudaMemcpyAsync(out + static_cast<size_t>(s0) * static_cast<size_t>(d), sl.out.p,
                                static_cast<size_t>(rows) * row_bytes, cudaMemcpyDeviceToHost, sl.stream);
            ctx->timing.d2h_bytes += static_cast<uint64_t>(rows) * row_bytes;
        }
        if (out16 && e == cudaSuccess) {
            e = cudaMemcpyAsync(out16 + static_cast<size_t>(s0) * static_cast<size_t>(d), sl.out16.p,
                                static_cast<size_t>(rows) * row_bytes / 2, cudaMemcpyDeviceToHost, sl.stream);
            ctx->timing.d2h_bytes += static_cast<uint64_t>(rows) * row_bytes / 2;
        }
        if (sumsq && e == cudaSuccess) {
            e = cudaMemcpyAsync(sumsq + s0, sl.sumsq.p, static_cast<size_t>(rows) * sizeof(uint64_t),
                                cudaMemcpyDeviceToHost, sl.stream);
            ctx->timing.d2h_bytes += static_cast<uint64_t>(rows) * sizeof(uint64_t);
        }
        if (ref_norms && e == cudaSuccess) {
            e = cudaMemcpyAsync(ref_norms + s0, sl.refnorm.p, static_cast<size_t>(rows) * sizeof(float),
                                cudaMemcpyDeviceToHost, sl.stream);
            ctx->timing.d2h_bytes += static_cast<uint64_t>(rows) * sizeof(float);
        }
        if (e == cudaSuccess) e = cudaEventRecord(sl.done, sl.stream);
        if (e != cudaSuccess) { ctx->last_cuda = e; cudaGetLastError(); rc = MVS_ERR_CUDA; break; }
        sl.busy = true;
        s0 = s1;
    }
    for (auto& sl : ctx->slots) {
        const int r2 = wait_slot(ctx, sl);
        if (rc == MVS_OK) rc = r2;
    }
    return rc;
}

int mvs_sketch_batch(mvs_ctx* ctx, const uint64_t* hashes, const uint64_t* offsets, int64_t n_samples, int d,
                     uint64_t seed, int32_t* out, int16_t* out16, uint64_t* sumsq, float* ref_norms)
{
    return sketch_batch_impl(ctx, hashes, false, offsets, n_samples, d, seed, out, out16, sumsq, ref_norms);
}

int mvs_sketch_batch_packed56(mvs_ctx* ctx, const unsigned char* packed, const uint64_t* offsets, int64_t n_samples, int d,
                              uint64_t seed, int32_t* out, int16_t* out16, uint64_t* sumsq, float* ref_norms)
{
    return sketch_batch_impl(ctx, packed, true, offsets, n_samples, d, seed, out, out16, sumsq, ref_norms);
}

int mvs_pack56(const uint64_t* hashes, uint64_t n, unsigned char* packed)
{
    if ((!hashes || !packed) && n) return MVS_ERR_INVALID_ARG;
    uint64_t bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (int64_t i = 0; i < static_cast<int64_t>(n); ++i) {
        const uint64_t v = hashes[i];
        bad |= v >> 56;
        for (int b = 0; b < 7; ++b) packed[7 * static_cast<uint64_t>(i) + b] = static_cast<unsigned char>(v >> (8 * b));
    }
    return bad ? MVS_ERR_INVALID_ARG : MVS_OK;   // a hash >= 2^56 does not fit: use the u64 entry point
}

int mvs_project_batch(mvs_ctx* ctx, const uint64_t* hashes, const uint64_t* offsets, int64_t n_samples, int d,
                      uint64_t seed, int32_t* out)
{
    if (!out && n_samples > 0) return MVS_ERR_INVALID_ARG;
    return mvs_sketch_batch(ctx, hashes, offsets, n_samples, d, seed, out, nullptr, nullptr, nullptr);
}

int mvs_transform_set(mvs_ctx* ctx, const uint64_t* hashes, int64_t n_hashes, int d, uint64_t seed, int32_t* out)
{
    if (n_hashes < 0) return MVS_ERR_INVALID_ARG;
    const uint64_t offs[2] = {0, static_cast<uint64_t>(n_hashes)};
    return mvs_project_batch(ctx, hashes, offs, 1, d, seed, out);
}

int mvs_finalize_device(mvs_ctx* ctx, const int32_t* d_vec, int64_t n_samples, int d, uint64_t* d_sumsq,
                        int16_t* d_out16, float* d_ref_norms, void* stream)
{
    if (!ctx || n_samples < 0 || d <= 0 || (!d_vec && n_samples > 0)) return MVS_ERR_INVALID_ARG;
    if (n_samples == 0) return MVS_OK;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->slots[0].stream;
    const int64_t rows_per_cta = 8;
    const int64_t grid = (n_samples + rows_per_cta - 1) / rows_per_cta;
    if (grid > 0x7FFFFFFFll) return MVS_ERR_TOO_LARGE;
    mvs::finalize_kernel<<<static_cast<unsigned>(grid), 256, 0, st>>>(
        d_vec, n_samples, d, reinterpret_cast<unsigned long long*>(d_sumsq), d_out16,
        d_ref_norms ? mvs_reference_inv_sqrt_d(d) : 0.f, d_ref_norms);
    MVS_CUDA(ctx, cudaGetLastError());
    ctx->launches += 1;
    if (!stream) MVS_CUDA(ctx, cudaStreamSynchronize(st));
    return MVS_OK;
}

int mvs_finalize(mvs_ctx* ctx, const int32_t* vec, int64_t n_samples, int d, uint64_t* sumsq, int16_t* out16,
                 float* ref_norms)
{
    if (!ctx || n_samples < 0 || d <= 0 || (!vec && n_samples > 0)) return MVS_ERR_INVALID_ARG;
    if (n_samples == 0) return MVS_OK;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    Slot& sl = ctx->slots[0];
    int rc = wait_slot(ctx, sl);
    if (rc != MVS_OK) return rc;
    const size_t n = static_cast<size_t>(n_samples) * static_cast<size_t>(d);
    // layout in sl.out: [vec int32 n][sumsq u64 n_samples][ref norms f32 n_samples][out16 int16 n]
    const size_t off_ss = (n * sizeof(int32_t) + 255) / 256 * 256;
    const size_t off_rn = off_ss + (static_cast<size_t>(n_samples) * sizeof(uint64_t) + 255) / 256 * 256;
    const size_t off_16 = off_rn + (static_cast<size_t>(n_samples) * sizeof(float) + 255) / 256 * 256;
    const size_t total = off_16 + n * sizeof(int16_t);
    if ((rc = ensure_dev(ctx, sl.out, total)) != MVS_OK) return rc;
    char* base = static_cast<char*>(sl.out.p);
    MVS_CUDA(ctx, cudaMemcpyAsync(base, vec, n * sizeof(int32_t), cudaMemcpyHostToDevice, sl.stream));
    rc = mvs_finalize_device(ctx, reinterpret_cast<int32_t*>(base), n_samples, d,
                             reinterpret_cast<uint64_t*>(base + off_ss),
                             out16 ? reinterpret_cast<int16_t*>(base + off_16) : nullptr,
                             ref_norms ? reinterpret_cast<float*>(base + off_rn) : nullptr, sl.stream);
    if (rc != MVS_OK) return rc;
    if (ref_norms)
        MVS_CUDA(ctx, cudaMemcpyAsync(ref_norms, base + off_rn, static_cast<size_t>(n_samples) * sizeof(float),
                                      cudaMemcpyDeviceToHost, sl.stream));
    if (sumsq)
        MVS_CUDA(ctx, cudaMemcpyAsync(sumsq, base + off_ss, static_cast<size_t>(n_samples) * sizeof(uint64_t),
                                      cudaMemcpyDeviceToHost, sl.stream));
    if (out16)
        MVS_CUDA(ctx, cudaMemcpyAsync(out16, base + off_16, n * sizeof(int16_t), cudaMemcpyDeviceToHost, sl.stream));
    MVS_CUDA(ctx, cudaStreamSynchronize(sl.stream));
    return MVS_OK;
}

int64_t mvs_sort_unique(uint64_t* hashes, int64_t n)
{
    if (!hashes || n <= 0) return 0;
    std::sort(hashes, hashes + n);
    return static_cast<int64_t>(std::unique(hashes, hashes + n) - hashes);
}

void* mvs_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void* mvs_host_alloc_write_combined(size_t bytes)
{
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocWriteCombined) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void mvs_host_free(void* p)
{
    if (p) cudaFreeHost(p);
}
// ---- host locality ---------------------------------------------------------------------------------------
namespace {

bool read_small_file(const char* path, char* buf, size_t cap)
{
    FILE* f = std::fopen(path, "r");
    if (!f) return false;
    const size_t n = std::fread(buf, 1, cap - 1, f);
    std::fclose(f);
    buf[n] = 0;
    return n > 0;
}

// "0-15,64-79" -> cpu_set_t
bool parse_cpulist(const char* s, cpu_set_t* set)
{
    CPU_ZERO(set);
    bool any = false;
    while (*s) {
        char* end = nullptr;
        const long a = std::strtol(s, &end, 10);
        if (end == s) break;
        long b = a;
        s = end;
        if (*s == '-') {
            b = std::strtol(s + 1, &end, 10);
            s = end;
        }
        for (long c = a; c <= b && c < CPU_SETSIZE; ++c) {
            if (c >= 0) {
                CPU_SET(static_cast<int>(c), set);
                any = true;
            }
        }
        while (*s == ',' || *s == ' ' || *s == '\n') ++s;
    }
    return any;
}

}  // namespace

int mvs_bind_host_to_device(int device, int* numa_node, int* applied)
{
    if (numa_node) *numa_node = -1;
    if (applied) *applied = 0;
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) {
        cudaGetLastError();
        return MVS_ERR_NO_DEVICE;
    }
    for (char* c = bus; *c; ++c)
        if (*c >= 'A' && *c <= 'F') *c = static_cast<char>(*c - 'A' + 'a');       // sysfs spells the bus id in lower case
    char path[128], buf[4096];
    int node = -1, did = 0;
    std::snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
    if (read_small_file(path, buf, sizeof buf)) node = std::atoi(buf);
    std::snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/local_cpulist", bus);
    cpu_set_t local, allowed, both;
    if (read_small_file(path, buf, sizeof buf) && parse_cpulist(buf, &local) &&
        sched_getaffinity(0, sizeof allowed, &allowed) == 0) {
        CPU_AND(&both, &local, &allowed);
        // fewer than 4 local CPUs would starve the library's own copy threads: leave the affinity alone then
        if (CPU_COUNT(&both) >= 4 && sched_setaffinity(0, sizeof both, &both) == 0) did |= 1;
    }
    if (node >= 0 && node < 1024) {
        unsigned long mask[16] = {0};                                             // 1024 nodes
        mask[node / (8 * sizeof(unsigned long))] |= 1ul << (node % (8 * sizeof(unsigned long)));
        constexpr int kMpolPreferred = 1;                                         // <linux/mempolicy.h>; no libnuma here
        if (syscall(SYS_set_mempolicy, kMpolPreferred, mask, 1025ul) == 0) did |= 2;
    }
    if (numa_node) *numa_node = node;
    if (applied) *applied = did;
    return MVS_OK;
}

int mvs_host_register(void* p, size_t bytes)
{
    if (!p) return MVS_ERR_INVALID_ARG;
    if (cudaHostRegister(p, bytes, cudaHostRegisterDefault) != cudaSuccess) {
        cudaGetLastError();
        return MVS_ERR_CUDA;
    }
    return MVS_OK;
}
int mvs_host_unregister(void* p)
{
    if (!p) return MVS_ERR_INVALID_ARG;
    if (cudaHostUnregister(p) != cudaSuccess) {
        cudaGetLastError();
        return MVS_ERR_CUDA;
    }
    return MVS_OK;
}

int mvs_get_timing(mvs_ctx* ctx, mvs_timing* t)
{
    if (!ctx || !t) return MVS_ERR_INVALID_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    // mean over the device-path calls since the previous query (at most the ring size)
    uint64_t first = ctx->timing_read;
    if (ctx->timing_calls - first > static_cast<uint64_t>(kTimingRing)) first = ctx->timing_calls - kTimingRing;
    double sum_a = 0.0, sum_b = 0.0;
    uint32_t n = 0;
    for (uint64_t c = first; c < ctx->timing_calls; ++c, ++n) {
        cudaEvent_t* ev = ctx->ev[c % kTimingRing];
        MVS_CUDA(ctx, cudaEventSynchronize(ev[2]));
        float a = 0.f, b = 0.f;
        MVS_CUDA(ctx, cudaEventElapsedTime(&a, ev[0], ev[2]));
        MVS_CUDA(ctx, cudaEventElapsedTime(&b, ev[1], ev[2]));
        sum_a += a;
        sum_b += b;
    }
    ctx->timing_read = ctx->timing_calls;
    ctx->timing.kernel_ms = n ? static_cast<float>(sum_a / n) : 0.f;
    ctx->timing.reduce_ms = n ? static_cast<float>(sum_b / n) : 0.f;
    ctx->timing.calls = n;
    ctx->timing.launches = ctx->launches;
    *t = ctx->timing;
    return MVS_OK;
}

int mvs_int_issue_peak(mvs_ctx* ctx, int which, double* ops_per_second)
{
    if (!ctx || !ops_per_second || which < 0 || which > 9) return MVS_ERR_INVALID_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->slots[0].stream;
    const int iters = 4096;
    const unsigned grid = static_cast<unsigned>(ctx->sm_count) * 8u;
    cudaEvent_t e0, e1;
    MVS_CUDA(ctx, cudaEventCreate(&e0));
    MVS_CUDA(ctx, cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0, st);
#define MVS_BENCH_CASE(M) \
    case M: mvs::issue_bench_kernel<M><<<grid, 256, 0, st>>>(ctx->d_sink, iters, 0x9e3779b9u, 0x85ebca6bu); break;
        switch (which) {
            MVS_BENCH_CASE(0) MVS_BENCH_CASE(1) MVS_BENCH_CASE(2) MVS_BENCH_CASE(3) MVS_BENCH_CASE(4)
            MVS_BENCH_CASE(5) MVS_BENCH_CASE(6) MVS_BENCH_CASE(7) MVS_BENCH_CASE(8) MVS_BENCH_CASE(9)
            default: break;
        }
#undef MVS_BENCH_CASE
        cudaEventRecord(e1, st);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) {
            ctx->last_cuda = e;
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            return MVS_ERR_CUDA;
        }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0) best = std::min(best, ms);
        ctx->launches += 1;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double ops = static_cast<double>(grid) * 256.0 * static_cast<double>(iters) * 32.0;
    *ops_per_second = ops / (static_cast<double>(best) * 1e-3);
    return MVS_OK;
}

// The planner without a device: how a batch would be cut into warp tasks on a GPU with `sm_count` SMs.  Host only.
int mvs_plan_tasks(const uint64_t* offsets, int64_t n_samples, int d, int sm_count, uint64_t* cap_hashes, uint64_t* n_tasks,
                   uint64_t* n_partial_rows, uint64_t* task_begin, uint32_t* task_count, uint32_t* task_sample,
                   uint32_t* task_pass, uint64_t max_tasks)
{
    if (!offsets || n_samples < 0 || d <= 0 || sm_count <= 0) return MVS_ERR_INVALID_ARG;
    const Geometry g = make_geometry(d);
    const Chunker ch{choose_cap(offsets, 0, n_samples, g, sm_count), g.quantum};
    PlanCounts pc;
    const int rc = count_plan(offsets, 0, n_samples, g, ch, pc);
    if (rc != MVS_OK) return rc;
    if (cap_hashes) *cap_hashes = ch.cap;
    if (n_tasks) *n_tasks = pc.tasks;
    if (n_partial_rows) *n_partial_rows = pc.prows;
    if (task_begin && task_count && task_sample && task_pass) {
        if (pc.tasks > max_tasks) return MVS_ERR_TOO_LARGE;
        std::vector<WarpTask> tasks(static_cast<size_t>(pc.tasks) + 1);
        std::vector<ReduceItem> reduce(static_cast<size_t>(pc.reduce) + 1);
        emit_plan(offsets, 0, n_samples, offsets[0], 0, g, ch, tasks.data(), reduce.data());
        for (uint64_t i = 0; i < pc.tasks; ++i) {
            task_begin[i] = tasks[i].begin;
            task_count[i] = tasks[i].count;
            task_sample[i] = tasks[i].sample;
            task_pass[i] = tasks[i].pass;
        }
    }
    return MVS_OK;
}

int mvs_loop_mix_peak(mvs_ctx* ctx, double* rounds_per_second)
{
    if (!ctx || !rounds_per_second) return MVS_ERR_INVALID_ARG;
    MVS_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->slots[0].stream;
    const int iters = 4096;
    const unsigned grid = static_cast<unsigned>(ctx->sm_count) * 6u;     // the projection's residency: 6 CTAs of 4 warps per SM
    cudaEvent_t e0, e1;
    MVS_CUDA(ctx, cudaEventCreate(&e0));
    MVS_CUDA(ctx, cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0, st);
        mvs::loop_mix_kernel<<<grid, 128, 0, st>>>(ctx->d_sink, iters, 0x9e3779b9u, 0x85ebca6bu, 1u);
        cudaEventRecord(e1, st);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) {
            ctx->last_cuda = e;
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            return MVS_ERR_CUDA;
        }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0) best = std::min(best, ms);
        ctx->launches += 1;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double rounds = static_cast<double>(grid) * 4.0 /*warps*/ * static_cast<double>(iters) * 4.0;   // warp rounds
    *rounds_per_second = rounds / (static_cast<double>(best) * 1e-3);
    return MVS_OK;
}

int mvs_set_option(mvs_ctx* ctx, const char* key, int64_t value)
{
    if (!ctx || !key) return MVS_ERR_INVALID_ARG;
    // a key is accepted only with a value in its range; anything else is an error, never silently ignored
    auto set = [&](const char* name, int64_t lo, int64_t hi, auto apply) -> int {
        if (std::strcmp(key, name) != 0) return 1;                       // not this key
        if (value < lo || value > hi) return MVS_ERR_INVALID_ARG;
        apply();
        return MVS_OK;
    };
    int r;
    if ((r = set("slice_hashes", 1, INT64_MAX, [&] { ctx->slice_hashes = static_cast<uint64_t>(value); })) != 1) return r;
    if ((r = set("slice_rows_bytes", 1, INT64_MAX, [&] { ctx->slice_rows_bytes = static_cast<uint64_t>(value); })) != 1) return r;
    if ((r = set("stage_threads", 1, 64, [&] { ctx->stage_threads = static_cast<int>(value); })) != 1) return r;
    if (!std::strcmp(key, "relay_device")) {     // -1 = off; a peer device index = route hash uploads through it
        if (value < -1 || value > 1023) return MVS_ERR_INVALID_ARG;
        cudaSetDevice(ctx->device);
        for (auto& sl : ctx->slots)
            if (sl.stream) cudaStreamSynchronize(sl.stream);
        return relay_setup(ctx, static_cast<int>(value));
    }
    if ((r = set("relay_percent", 0, 100, [&] { ctx->relay.percent = static_cast<int>(value); ctx->relay.credit = 0; })) != 1) return r;
    if ((r = set("release_workspace", INT64_MIN, INT64_MAX, [&] { cudaSetDevice(ctx->device); cudaDeviceSynchronize(); release_workspace(ctx); })) != 1) return r;
#ifdef MVS_EXPERIMENTS                           // A/B knobs of the experiment build (libmvs_b200_exp.so); the shipped library rejects them
    if ((r = set("task_cap", 0, INT64_MAX, [&] { ctx->task_cap = static_cast<uint64_t>(value); })) != 1) return r;
    if ((r = set("pairwise_path", 0, 2, [&] { ctx->pairwise_path = static_cast<int>(value); })) != 1) return r;
    if ((r = set("stage_copy", 0, 1, [&] { ctx->stage_copy = static_cast<int>(value); })) != 1) return r;
    if (!std::strcmp(key, "stage_uploads")) {
        if (value != 0 && value != 2) return MVS_ERR_INVALID_ARG;
        ctx->stage_uploads = static_cast<int>(value);
        return MVS_OK;
    }
    if (!std::strcmp(key, "pairwise_fuse")) {
        if (value != 0 && value != 2) return MVS_ERR_INVALID_ARG;
        ctx->pairwise_fuse = static_cast<int>(value);
        return MVS_OK;
    }
    if ((r = set("pairwise_pair", 0, 2, [&] { ctx->pairwise_pair = static_cast<int>(value); })) != 1) return r;
    if (!std::strcmp(key, "pairwise_cluster")) {
        if (value != 0 && value != 1 && value != 2 && value != 4) return MVS_ERR_INVALID_ARG;
        ctx->pairwise_cluster = static_cast<int>(value);
        return MVS_OK;
    }
    if ((r = set("pairwise_stream", 0, 1, [&] { ctx->pairwise_stream = static_cast<int>(value); })) != 1) return r;
    if ((r = set("pairwise_limbs", 0, 4, [&] { ctx->pairwise_limbs = static_cast<int>(value); })) != 1) return r;
#endif
    return MVS_ERR_INVALID_ARG;                  // unknown key
}

int mvs_has_experiments(void)
{
#ifdef MVS_EXPERIMENTS
    return 1;
#else
    return 0;
#endif
}

}  // extern "C"
