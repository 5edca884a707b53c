// mvs_pairwise.cu -- host side of the all-pairs comparison entry points of include/mvs_b200.h
// (SURVEY 8f rank 4).  Replaces, for the GPU, compute_sparse_dot_products_optimized
// (reference src/pairwise_comp_optimized.cpp:57-160) and the block loops that drive it (:949-983).
// Device code: mvs_pairwise.cuh.  No CPU implementation exists here: without a device the calls fail.
#include "../../include/mvs_b200.h"
#include "mvs_internal.h"
#include "mvs_pairwise.cuh"
#include "mvs_pairwise_umma.cuh"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace {

// Device buffers live in the context's grow-only workspace (mvs_internal_workspace): repeated calls -- the reference
// calls its function once per block pair -- allocate nothing.  Slots:
enum Slot { kWsA = 0, kWsB, kWsNi, kWsNj, kWsC, kWsSeg, kWsRowoff, kWsMinmax, kWsPlanesA, kWsPlanesB, kWsRows, kWsCols, kWsVals };

struct Dev {                                     // a view of one workspace slot
    void* p = nullptr;
    int alloc(mvs_ctx* ctx, int slot, size_t bytes) { return mvs_internal_workspace(ctx, slot, bytes, &p); }
    template <typename T> T* as() const { return static_cast<T*>(p); }
};
#define PW_RC(call) do { const int rc_ = (call); if (rc_ != MVS_OK) return rc_; } while (0)

struct OutBuf {                                  // device triples of one row block, reused across blocks
    Dev rows, cols, vals;
    unsigned long long cap = 0;
    int reserve(mvs_ctx* ctx, unsigned long long n)
    {
        PW_RC(rows.alloc(ctx, kWsRows, n * sizeof(int32_t)));
        PW_RC(cols.alloc(ctx, kWsCols, n * sizeof(int32_t)));
        PW_RC(vals.alloc(ctx, kWsVals, n * sizeof(int64_t)));
        cap = n;
        return MVS_OK;
    }
};

struct Triples {                                 // the arrays handed to the caller in mvs_sparse_result (malloc / realloc / free)
    int32_t* rows = nullptr;
    int32_t* cols = nullptr;
    int64_t* values = nullptr;
    size_t n = 0, cap = 0;
    ~Triples() { std::free(rows); std::free(cols); std::free(values); }
    bool reserve(size_t more)                    // room for `more` further triples; no zero fill, no second copy later
    {
        if (n + more <= cap) return true;
        const size_t want = std::max(n + more, cap + cap / 2);
        void* r = std::realloc(rows, want * sizeof(int32_t));
        if (r) rows = static_cast<int32_t*>(r);
        void* c = std::realloc(cols, want * sizeof(int32_t));
        if (c) cols = static_cast<int32_t*>(c);
        void* v = std::realloc(values, want * sizeof(int64_t));
        if (v) values = static_cast<int64_t*>(v);
        if (!r || !c || !v) return false;
        cap = want;
        return true;
    }
};

#define PW_CUDA(ctx, call)                                                    \
    do {                                                                      \
        cudaError_t e_ = (call);                                              \
        if (e_ != cudaSuccess) {                                              \
            mvs_internal_set_cuda_error((ctx), e_);                           \
            cudaGetLastError();                                               \
            return e_ == cudaErrorMemoryAllocation ? MVS_ERR_OOM : MVS_ERR_CUDA; \
        }                                                                     \
    } while (0)

// An int32 operand matrix resident in HBM, plus (tensor-core path) its int8 limb planes.
struct Operand {
    const int32_t* v = nullptr;                  // [n x d] int32
    int64_t n = 0;
    Dev planes;                                  // [P][n_pad x k_pad] int8
    int P = 0;
    int64_t n_pad = 0;
    int k_pad = 0;
};

// cuTensorMapEncodeTiled through the runtime's driver entry point (libcuda is not linked directly)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled()
{
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess) p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// 3-D map {k, row, plane} over the limb planes; box = 128 bytes of k x box_rows rows x 1 plane, 128-byte swizzle
bool make_tensor_map(const Operand& op, int box_rows, CUtensorMap* tm)
{
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return false;
    const cuuint64_t dims[3] = {static_cast<cuuint64_t>(op.k_pad), static_cast<cuuint64_t>(op.n_pad), static_cast<cuuint64_t>(op.P)};
    const cuuint64_t strides[2] = {static_cast<cuuint64_t>(op.k_pad), static_cast<cuuint64_t>(op.k_pad) * static_cast<cuuint64_t>(op.n_pad)};
    const cuuint32_t box[3] = {mvs::kUmBKB, static_cast<cuuint32_t>(box_rows), 1}, estr[3] = {1, 1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, op.planes.p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int CM, int CN, bool F22>
cudaError_t launch_umma(cudaStream_t st, int sms, size_t smem, const CUtensorMap& tmA, const CUtensorMap& tmB, const mvs::UmmaPlan& plan,
                        int k_pad, int a_row0, int64_t mi, int64_t mj, int32_t* d_C, int64_t ldc, int tiles_m, int tiles_n,
                        const mvs::UmFilter& filt)
{
    auto kernel = mvs::pairwise_dot_umma_kernel<CM, CN, F22>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    constexpr int CS = CM * CN;
    const int super_m = (tiles_m + CM - 1) / CM, super_n = (tiles_n + CN - 1) / CN;
    const int64_t nsuper = static_cast<int64_t>(super_m) * super_n;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(mvs::kUmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int max_clusters = sms / CS;
    if (CS > 1) {                                // clusters that can be co-resident (GPC boundaries)
        cfg.gridDim = dim3(static_cast<unsigned>(sms / CS * CS));
        int q = 0;
        if (cudaOccupancyMaxActiveClusters(&q, kernel, &cfg) == cudaSuccess && q > 0) max_clusters = q;
        else cudaGetLastError();
    }
    const int nclusters = static_cast<int>(std::min<int64_t>(nsuper, max_clusters));
    if (CS == 1) {                               // plain launch, no cluster attribute
        kernel<<<static_cast<unsigned>(nclusters), mvs::kUmThreads, smem, st>>>(tmA, tmB, plan, k_pad, a_row0, mi, mj, d_C, ldc, super_n,
                                                                              static_cast<int>(nsuper), filt);
        return cudaGetLastError();
    }
    cfg.gridDim = dim3(static_cast<unsigned>(nclusters * CS));
    return cudaLaunchKernelEx(&cfg, kernel, tmA, tmB, plan, k_pad, a_row0, mi, mj, d_C, ldc, super_n, static_cast<int>(nsuper), filt);
}

// CTA pairs (tcgen05 cta_group::2): one cluster of two CTAs per 256 x 128 super-tile
cudaError_t launch_umma_pair(cudaStream_t st, int sms, const CUtensorMap& tmA, const CUtensorMap& tmB, mvs::UmmaPlan plan, int k_pad,
                             int a_row0, int a_row_last, int64_t mi, int64_t mj, int32_t* d_C, int64_t ldc, int tiles_m, int tiles_n,
                             const mvs::UmFilter& filt)
{
    const bool f22 = plan.ngroups == 1 && plan.a_count == 2 && plan.b_count == 2;
    auto kernel = f22 ? mvs::pairwise_dot_umma_pair_kernel<true> : mvs::pairwise_dot_umma_pair_kernel<false>;
    const int stage_bytes = plan.a_count * mvs::kUmTile + plan.b_count * mvs::kUmPairBTile;
    plan.stages = std::max(2, std::min(8, mvs::kUmSmemBudget / stage_bytes));
    const size_t smem = static_cast<size_t>(plan.stages) * stage_bytes + 1024 + 256;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    const int64_t nsuper = static_cast<int64_t>((tiles_m + 1) / 2) * tiles_n;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(mvs::kUmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3(static_cast<unsigned>(sms / 2 * 2));
    int max_pairs = sms / 2, q = 0;
    if (cudaOccupancyMaxActiveClusters(&q, kernel, &cfg) == cudaSuccess && q > 0) max_pairs = std::min(max_pairs, q);
    else cudaGetLastError();
    const int npairs = static_cast<int>(std::min<int64_t>(nsuper, max_pairs));
    cfg.gridDim = dim3(static_cast<unsigned>(npairs * 2));
    return cudaLaunchKernelEx(&cfg, kernel, tmA, tmB, plan, k_pad, a_row0, a_row_last, mi, mj, d_C, ldc, tiles_n, static_cast<int>(nsuper), filt);
}

mvs::UmmaPlan make_umma_plan(int PA, int PB, bool allow_fused)
{
    mvs::UmmaPlan pl{};
    pl.smax = std::min(PA + PB - 2, 3);
    if (allow_fused && PA + PB <= 4) {           // FUSED: all planes of a k-slice in one stage
        pl.ngroups = 1;
        pl.a_count = PA;
        pl.b_count = PB;
        pl.a_first[0] = 0;
        pl.b_first[0] = 0;
    } else {                                     // STREAM: one limb pair per group
        pl.a_count = pl.b_count = 1;
        for (int s = pl.smax; s >= 0; --s)
            for (int p = 0; p < PA; ++p) {
                const int q = s - p;
                if (q < 0 || q >= PB) continue;
                pl.a_first[pl.ngroups] = static_cast<unsigned char>(p);
                pl.b_first[pl.ngroups] = static_cast<unsigned char>(q);
                ++pl.ngroups;
            }
    }
    const int stage_bytes = (pl.a_count + pl.b_count) * mvs::kUmTile;
    pl.stages = std::max(2, std::min(6, mvs::kUmSmemBudget / stage_bytes));
    return pl;
}

// limbs needed so that every value in [lo, hi] is a sum of P balanced base-256 digits (P = 4 covers int32 mod 2^32)
int limbs_for_range(int lo, int hi)
{
    long long top = 0;                           // (256^P - 1) / 255
    for (int P = 1; P <= 3; ++P) {
        top = top * 256 + 1;
        if (static_cast<long long>(lo) >= -128 * top && static_cast<long long>(hi) <= 127 * top) return P;
    }
    return 4;
}

int prepare_operand(mvs_ctx* ctx, cudaStream_t st, Operand& op, int planes_slot, int d, int* d_mm, int* h_pinned)
{
    // one extra zero tile of rows: a row block may start at any row (shards), its last tile then reads past n
    op.n_pad = (op.n + mvs::kImBM - 1) / mvs::kImBM * mvs::kImBM + mvs::kImBM;
    op.k_pad = (d + mvs::kUmBKB - 1) / mvs::kUmBKB * mvs::kUmBKB;      // 128: a multiple of both paths' k-slices
    const int forced = mvs_internal_pairwise_limbs(ctx);
    int sms = 148;
    PW_CUDA(ctx, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mvs_internal_device(ctx)));
    const int64_t quads = op.n_pad * (op.k_pad / 4);
    const unsigned grid = static_cast<unsigned>(std::min<int64_t>((quads + 255) / 256, static_cast<int64_t>(sms) * 16));
    auto split = [&](int P, int* mm) {
        const int rc = op.planes.alloc(ctx, planes_slot, static_cast<size_t>(P) * op.n_pad * op.k_pad);
        if (rc != MVS_OK) return rc;
        mvs::pairwise_split_kernel<<<grid, 256, 0, st>>>(op.v, op.n, d, P, op.planes.as<int8_t>(), op.n_pad, op.k_pad, mm);
        mvs_internal_count_launches(ctx, 1);
        if (cudaGetLastError() != cudaSuccess) return static_cast<int>(MVS_ERR_CUDA);
        return static_cast<int>(MVS_OK);
    };
    // sketch entries are ~sqrt(#hashes): two limbs is the normal case.  Split for the guess and measure the range in the
    // same pass; only if the range asks for another limb count is V read a second time.
    const int guess = std::max(2, forced);
    int* mm = h_pinned;                          // pinned: the two small copies are truly asynchronous
    mm[0] = 2147483647;
    mm[1] = -2147483647 - 1;
    PW_CUDA(ctx, cudaMemcpyAsync(d_mm, mm, 2 * sizeof(int), cudaMemcpyHostToDevice, st));
    PW_RC(split(guess, d_mm));
    PW_CUDA(ctx, cudaMemcpyAsync(mm, d_mm, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    PW_CUDA(ctx, cudaStreamSynchronize(st));
    op.P = std::max(limbs_for_range(mm[0], mm[1]), forced);
    if (op.P != guess) PW_RC(split(op.P, nullptr));
    return MVS_OK;
}

#ifdef MVS_EXPERIMENTS
mvs::ImmaPlan make_plan(int PA, int PB)
{
    mvs::ImmaPlan plan{};
    const int smax = std::min(PA + PB - 2, 3);   // weights 256^s with s > 3 vanish modulo 2^32
    for (int s = smax; s >= 0; --s) {
        bool first = true;
        for (int p = 0; p < PA; ++p) {
            const int q = s - p;
            if (q < 0 || q >= PB) continue;
            plan.p[plan.npairs] = static_cast<unsigned char>(p);
            plan.q[plan.npairs] = static_cast<unsigned char>(q);
            plan.shift_before[plan.npairs] = (first && plan.npairs > 0) ? 1 : 0;
            first = false;
            ++plan.npairs;
        }
    }
    return plan;
}
#endif

// Rows [r0, r0 + mi) of A against all of B, operands resident in HBM.  Appends the kept triples
// (row_base + i, col_base + j, dot) in (i, j) order to `out`; adds the CUDA-event times.
int pairwise_core(mvs_ctx* ctx, cudaStream_t st, const Operand& A, int64_t r0, int64_t mi, const Operand& B, int d,
                  const double* d_ni, const double* d_nj, int32_t row_base, int32_t col_base, int32_t* d_C,
                  uint32_t* d_segcount, unsigned long long* d_rowoff, OutBuf& ob, void* h_pinned, Triples& out, double& gemm_ms,
                  double& filter_ms, uint64_t& d2h_bytes)
{
    const int64_t mj = B.n;
    if (mi == 0 || mj == 0) return MVS_OK;
    cudaEvent_t e0, e1, e2;
    PW_CUDA(ctx, cudaEventCreate(&e0));
    PW_CUDA(ctx, cudaEventCreate(&e1));
    PW_CUDA(ctx, cudaEventCreate(&e2));
    struct EvGuard { cudaEvent_t a, b, c; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); cudaEventDestroy(c); } } guard{e0, e1, e2};

    const int64_t ldc = (mj + 3) & ~int64_t{3};
    [[maybe_unused]] const int32_t* d_A = A.v + r0 * d;   // int32 rows: read by the experiment GEMMs only
    const int nseg = static_cast<int>((mj + mvs::kPwSeg - 1) / mvs::kPwSeg);
    const unsigned seg_ctas = static_cast<unsigned>(((mi + mvs::kPwRows - 1) / mvs::kPwRows * nseg + 7) / 8);   // 8 warps per CTA
    bool fused = false;
    PW_CUDA(ctx, cudaEventRecord(e0, st));
#ifdef MVS_EXPERIMENTS
    const int path = mvs_internal_pairwise_path(ctx);
#else
    constexpr int path = 0;                      // the shipped library has ONE GEMM: the tcgen05 limb GEMM
#endif
    if (A.P > 0 && path == 0) {                  // tcgen05 limb GEMM (TMA + TMEM)
        mvs::UmmaPlan plan = make_umma_plan(A.P, B.P, mvs_internal_pairwise_stream(ctx) == 0);
#ifdef MVS_UMMA_PROBE                            // diagnostic build only: switch kernel roles off (wrong results), see UmmaPlan::probe
        if (const char* e = std::getenv("MVS_UMMA_PROBE")) plan.probe = std::atoi(e);
#endif
        const size_t smem = static_cast<size_t>(plan.stages) * (plan.a_count + plan.b_count) * mvs::kUmTile + 1024 + 256;
        const int tiles_n = nseg, tiles_m = static_cast<int>((mi + mvs::kUmBM - 1) / mvs::kUmBM);
        if (static_cast<int64_t>(tiles_n) * tiles_m > 0x3fffffffLL) return MVS_ERR_TOO_LARGE;
        int sms = 0;
        PW_CUDA(ctx, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mvs_internal_device(ctx)));
        // pre-filter fused into the epilogue (option "pairwise_fuse" = 2 switches it off): only rows x 128-column segments
        // that may keep something are stored and marked; the recount kernel below visits those instead of all of C
        fused = mvs_internal_pairwise_fuse(ctx) != 2;
        const mvs::UmFilter filt{fused ? d_ni : nullptr, d_nj, d_segcount, nseg, d};
        // cluster shape: 1 x 1 by default.  Clusters of 1 x 2 / 2 x 2 CTAs sharing operand tiles by TMA multicast (option
        // "pairwise_cluster" = 2 / 4) halve the L2 reads but are NOT faster: the main loop is bound by the bytes an SM
        // can take in (~32 B/clk, the same with multicast), see DESIGN.md section 9
#ifdef MVS_EXPERIMENTS
        int cshape = mvs_internal_pairwise_cluster(ctx);
        if (cshape == 0) cshape = 1;
#else
        constexpr int cshape = 1;
#endif
        const int cm = cshape == 4 ? 2 : 1, cn = cshape >= 2 ? 2 : 1;
        // default: CTA pairs (tcgen05 cta_group::2, 256 x 128 super-tiles) whenever there are two tile rows to pair
        const int pair_opt = mvs_internal_pairwise_pair(ctx);
        if (cshape == 1 && (pair_opt == 1 || (pair_opt == 0 && tiles_m >= 2))) {
            CUtensorMap tmA, tmB;
            if (!make_tensor_map(A, mvs::kUmBM, &tmA) || !make_tensor_map(B, mvs::kUmBN / 2, &tmB)) return MVS_ERR_CUDA;
            PW_CUDA(ctx, launch_umma_pair(st, sms, tmA, tmB, plan, A.k_pad, static_cast<int>(r0), static_cast<int>(A.n_pad - mvs::kUmBM), mi,
                                          mj, d_C, ldc, tiles_m, tiles_n, filt));
        } else {
        CUtensorMap tmA, tmB;
        if (!make_tensor_map(A, mvs::kUmBM / cn, &tmA) || !make_tensor_map(B, mvs::kUmBN / cm, &tmB)) return MVS_ERR_CUDA;
        cudaError_t le;
        const bool f22 = plan.ngroups == 1 && plan.a_count == 2 && plan.b_count == 2;
#ifdef MVS_EXPERIMENTS
        if (cshape == 4) le = launch_umma<2, 2, false>(st, sms, smem, tmA, tmB, plan, A.k_pad, static_cast<int>(r0), mi, mj, d_C, ldc, tiles_m, tiles_n, filt);
        else if (cshape == 2) le = launch_umma<1, 2, false>(st, sms, smem, tmA, tmB, plan, A.k_pad, static_cast<int>(r0), mi, mj, d_C, ldc, tiles_m, tiles_n, filt);
        else
#endif
        if (f22) le = launch_umma<1, 1, true>(st, sms, smem, tmA, tmB, plan, A.k_pad, static_cast<int>(r0), mi, mj, d_C, ldc, tiles_m, tiles_n, filt);
        else le = launch_umma<1, 1, false>(st, sms, smem, tmA, tmB, plan, A.k_pad, static_cast<int>(r0), mi, mj, d_C, ldc, tiles_m, tiles_n, filt);
        PW_CUDA(ctx, le);
#ifdef MVS_UMMA_PROBE
        if (cshape == 1 && std::getenv("MVS_UMMA_PROBE_PRINT")) {
            static unsigned long long h[1024][8];
            PW_CUDA(ctx, cudaStreamSynchronize(st));
            PW_CUDA(ctx, cudaMemcpyFromSymbol(h, mvs::um_probe_cycles, sizeof h));
            double a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const int n = static_cast<int>(std::min<int64_t>(static_cast<int64_t>(tiles_m) * tiles_n, sms));
            for (int b = 0; b < n; ++b) for (int k = 0; k < 8; ++k) a[k] += static_cast<double>(h[b][k]) / n;
            std::fprintf(stderr, "[umma probe] per CTA: MMA warp: wait tmem_empty %.0f clk, wait full %.0f clk, total %.0f clk, tiles %.1f; epilogue warp: wait tmem_full %.0f, drain %.0f, test+store %.0f clk\n", a[0], a[1], a[2], a[3], a[4], a[5], a[6]);
        }
#endif
        }
    }
#ifdef MVS_EXPERIMENTS
    else if (A.P > 0) {                          // warp-level mma.sync limb GEMM
        const mvs::ImmaPlan plan = make_plan(A.P, B.P);
        const dim3 grid(static_cast<unsigned>((mj + mvs::kImBN - 1) / mvs::kImBN), static_cast<unsigned>((mi + mvs::kImBM - 1) / mvs::kImBM));
        mvs::pairwise_dot_imma_kernel<<<grid, mvs::kImThreads, mvs::kImSmemBytes, st>>>(
            A.planes.as<int8_t>() + r0 * A.k_pad, A.n_pad * A.k_pad, B.planes.as<int8_t>(), B.n_pad * B.k_pad, A.k_pad, plan,
            mi, mj, d_C, ldc);
    } else {                                     // IMAD GEMM
        const dim3 grid(static_cast<unsigned>((mj + mvs::kPwBN - 1) / mvs::kPwBN), static_cast<unsigned>((mi + mvs::kPwBM - 1) / mvs::kPwBM));
        const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_A) & 15) == 0) && ((reinterpret_cast<uintptr_t>(B.v) & 15) == 0);
        if (vec) mvs::pairwise_dot_kernel<true><<<grid, mvs::kPwThreads, 0, st>>>(d_A, mi, B.v, mj, d, d_C, ldc);
        else mvs::pairwise_dot_kernel<false><<<grid, mvs::kPwThreads, 0, st>>>(d_A, mi, B.v, mj, d, d_C, ldc);
    }
#else
    else return MVS_ERR_INVALID_ARG;             // operands are always split in the shipped build
#endif
    PW_CUDA(ctx, cudaGetLastError());
    PW_CUDA(ctx, cudaEventRecord(e1, st));
#ifdef MVS_EXPERIMENTS
    if (!fused) {
        mvs::pairwise_count_kernel<<<seg_ctas, 256, 0, st>>>(d_C, mi, mj, ldc, d, d_ni, d_nj, d_segcount, nseg);
        PW_CUDA(ctx, cudaGetLastError());
        mvs_internal_count_launches(ctx, 1);
    }
#else
    (void)seg_ctas;
    if (!fused) return MVS_ERR_INVALID_ARG;
#endif
    const unsigned row_ctas = static_cast<unsigned>((mi + 7) / 8);                       // one warp per row
    mvs::pairwise_rowsum_kernel<<<row_ctas, 256, 0, st>>>(d_C, mi, mj, ldc, d, d_ni, d_nj, d_segcount, nseg, mvs::kUmMaybe, d_rowoff);
    PW_CUDA(ctx, cudaGetLastError());
    mvs::pairwise_scan_kernel<<<1, 1024, 0, st>>>(d_rowoff, mi);
    PW_CUDA(ctx, cudaGetLastError());
    mvs_internal_count_launches(ctx, 1);
    // emit is launched without knowing the total: it writes nothing if the triples do not fit the output buffers, the
    // host then grows them and launches it again (rare: the buffers start at 4 Mi triples)
    unsigned long long* h_total = static_cast<unsigned long long*>(h_pinned);
    auto emit = [&]() {
        mvs::pairwise_emit_kernel<<<row_ctas, 256, 0, st>>>(d_C, mi, mj, ldc, d, d_ni, d_nj, d_segcount, nseg, d_rowoff, row_base, col_base,
                                                           ob.cap, ob.rows.as<int32_t>(), ob.cols.as<int32_t>(), ob.vals.as<long long>());
        mvs_internal_count_launches(ctx, 1);
        return cudaGetLastError();
    };
    PW_CUDA(ctx, emit());
    PW_CUDA(ctx, cudaEventRecord(e2, st));
    PW_CUDA(ctx, cudaMemcpyAsync(h_total, d_rowoff + mi, sizeof *h_total, cudaMemcpyDeviceToHost, st));
    PW_CUDA(ctx, cudaStreamSynchronize(st));
    mvs_internal_count_launches(ctx, 2);
    const unsigned long long total = *h_total;
    d2h_bytes += sizeof total;
    if (total > ob.cap) {
        PW_RC(ob.reserve(ctx, total + total / 4));
        PW_CUDA(ctx, emit());
    }
    if (total) {
        if (!out.reserve(total)) return MVS_ERR_OOM;
        const size_t at = out.n;
        PW_CUDA(ctx, cudaMemcpyAsync(out.rows + at, ob.rows.p, total * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PW_CUDA(ctx, cudaMemcpyAsync(out.cols + at, ob.cols.p, total * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PW_CUDA(ctx, cudaMemcpyAsync(out.values + at, ob.vals.p, total * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        PW_CUDA(ctx, cudaStreamSynchronize(st));
        out.n += total;
        d2h_bytes += total * 16;
    }
    float a = 0.f, b = 0.f;
    PW_CUDA(ctx, cudaEventElapsedTime(&a, e0, e1));
    PW_CUDA(ctx, cudaEventElapsedTime(&b, e1, e2));
    gemm_ms += a;
    filter_ms += b;
    return MVS_OK;
}

// All row blocks of A[begin, end) against B; scratch sized for `rb` rows.
int run_blocks(mvs_ctx* ctx, cudaStream_t st, Operand& A, int64_t begin, int64_t end, Operand& B, bool same, int d,
               const double* d_ni, const double* d_nj, Triples& t, double& gemm_ms, double& filter_ms, double& prep_ms,
               uint64_t& d2h)
{
    const int64_t mj = B.n;
    const int64_t ldc = (mj + 3) & ~int64_t{3};
    const int64_t rb = std::min<int64_t>(std::max<int64_t>(128, std::min<int64_t>(16384, (int64_t{1} << 30) / (4 * std::max<int64_t>(ldc, 1)) / 256 * 256)),
                                         (end - begin + 127) / 128 * 128);
    const int nseg = static_cast<int>((mj + mvs::kPwSeg - 1) / mvs::kPwSeg);
    Dev C, segcount, rowoff, mm;
    OutBuf ob;
    void* h_pinned = mvs_internal_pinned_scratch(ctx);
    if (!h_pinned) return MVS_ERR_OOM;
    PW_RC(ob.reserve(ctx, 4ull << 20));
    PW_RC(C.alloc(ctx, kWsC, static_cast<size_t>(rb) * ldc * 4));
    PW_RC(segcount.alloc(ctx, kWsSeg, static_cast<size_t>(rb) * nseg * 4));
    PW_RC(rowoff.alloc(ctx, kWsRowoff, static_cast<size_t>(rb + 1) * 8));
    if (mvs_internal_pairwise_path(ctx) != 1) {
        PW_CUDA(ctx, cudaStreamSynchronize(st));                      // the asynchronous upload is not part of prep_ms
        const auto t0 = std::chrono::steady_clock::now();
        PW_RC(mm.alloc(ctx, kWsMinmax, 8));
#ifdef MVS_EXPERIMENTS
        PW_CUDA(ctx, cudaFuncSetAttribute(mvs::pairwise_dot_imma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, mvs::kImSmemBytes));
#endif
        PW_RC(prepare_operand(ctx, st, B, kWsPlanesB, d, mm.as<int>(), static_cast<int*>(h_pinned)));
        if (!same) PW_RC(prepare_operand(ctx, st, A, kWsPlanesA, d, mm.as<int>(), static_cast<int*>(h_pinned)));
        PW_CUDA(ctx, cudaStreamSynchronize(st));
        prep_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    const Operand& Aref = same ? B : A;
    for (int64_t r0 = begin; r0 < end; r0 += rb) {
        const int64_t m = std::min(rb, end - r0);
        const int rc = pairwise_core(ctx, st, Aref, r0, m, B, d, d_ni + r0, d_nj, static_cast<int32_t>(r0), 0, C.as<int32_t>(),
                                     segcount.as<uint32_t>(), rowoff.as<unsigned long long>(), ob, h_pinned, t, gemm_ms, filter_ms, d2h);
        if (rc != MVS_OK) return rc;
    }
    return MVS_OK;
}

int finish(Triples& t, double gemm_ms, double filter_ms, double prep_ms, double wall_ms, uint64_t h2d, uint64_t d2h, mvs_sparse_result** out)
{
    mvs_sparse_result* r = static_cast<mvs_sparse_result*>(std::calloc(1, sizeof(mvs_sparse_result)));
    if (!r || !t.reserve(t.n ? 0 : 1)) {         // an empty result still carries valid (1-element) arrays
        std::free(r);
        return MVS_ERR_OOM;
    }
    r->n = static_cast<int64_t>(t.n);
    r->rows = t.rows;                            // ownership moves to the result (mvs_free_sparse_result frees them)
    r->cols = t.cols;
    r->values = t.values;
    t.rows = t.cols = nullptr;
    t.values = nullptr;
    r->gemm_ms = gemm_ms;
    r->filter_ms = filter_ms;
    r->prep_ms = prep_ms;
    r->wall_ms = wall_ms;
    r->h2d_bytes = h2d;
    r->d2h_bytes = d2h;
    *out = r;
    return MVS_OK;
}

}  // namespace

extern "C" {

void mvs_free_sparse_result(mvs_sparse_result* r)
{
    if (!r) return;
    std::free(r->rows);
    std::free(r->cols);
    std::free(r->values);
    std::free(r);
}

int mvs_pairwise_block(mvs_ctx* ctx, const int32_t* block_i, int64_t mi, const int32_t* block_j, int64_t mj, int d,
                       const double* norms_i, const double* norms_j, mvs_sparse_result** out)
{
    if (!ctx || !out || mi < 0 || mj < 0 || d <= 0 || (mi > 0 && (!block_i || !norms_i)) ||
        (mj > 0 && (!block_j || !norms_j)))
        return MVS_ERR_INVALID_ARG;
    *out = nullptr;
    if (mi > 0x7fffffffLL || mj > 0x7fffffffLL) return MVS_ERR_TOO_LARGE;
    const auto t0 = std::chrono::steady_clock::now();
    PW_CUDA(ctx, cudaSetDevice(mvs_internal_device(ctx)));
    cudaStream_t st = mvs_internal_stream(ctx);
    Triples t;
    double gemm_ms = 0, filter_ms = 0, prep_ms = 0;
    uint64_t h2d = 0, d2h = 0;
    if (mi > 0 && mj > 0) {
        Dev A, B, ni, nj;
        const size_t ab = static_cast<size_t>(mi) * d * 4, bb = static_cast<size_t>(mj) * d * 4;
        PW_RC(A.alloc(ctx, kWsA, ab));
        PW_RC(B.alloc(ctx, kWsB, bb));
        PW_RC(ni.alloc(ctx, kWsNi, static_cast<size_t>(mi) * 8));
        PW_RC(nj.alloc(ctx, kWsNj, static_cast<size_t>(mj) * 8));
        PW_RC(mvs_internal_upload(ctx, st, A.p, block_i, ab));
        PW_RC(mvs_internal_upload(ctx, st, B.p, block_j, bb));
        PW_CUDA(ctx, cudaMemcpyAsync(ni.p, norms_i, static_cast<size_t>(mi) * 8, cudaMemcpyHostToDevice, st));
        PW_CUDA(ctx, cudaMemcpyAsync(nj.p, norms_j, static_cast<size_t>(mj) * 8, cudaMemcpyHostToDevice, st));
        h2d = ab + bb + static_cast<uint64_t>(mi + mj) * 8;
        Operand opA, opB;
        opA.v = A.as<int32_t>();
        opA.n = mi;
        opB.v = B.as<int32_t>();
        opB.n = mj;
        const int rc = run_blocks(ctx, st, opA, 0, mi, opB, false, d, ni.as<double>(), nj.as<double>(), t, gemm_ms, filter_ms,
                                  prep_ms, d2h);
        if (rc != MVS_OK) return rc;
    }
    const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return finish(t, gemm_ms, filter_ms, prep_ms, wall, h2d, d2h, out);
}

int mvs_pairwise_shard(mvs_ctx* ctx, const int32_t* vectors, int64_t n_total, int d, const double* norms_sq,
                       int64_t begin_row, int64_t end_row, mvs_sparse_result** out)
{
    if (!ctx || !out || n_total < 0 || d <= 0 || begin_row < 0 || end_row < begin_row || end_row > n_total ||
        (n_total > 0 && (!vectors || !norms_sq)))
        return MVS_ERR_INVALID_ARG;
    *out = nullptr;
    if (n_total > 0x7fffffffLL) return MVS_ERR_TOO_LARGE;
    const auto t0 = std::chrono::steady_clock::now();
    PW_CUDA(ctx, cudaSetDevice(mvs_internal_device(ctx)));
    cudaStream_t st = mvs_internal_stream(ctx);
    Triples t;
    double gemm_ms = 0, filter_ms = 0, prep_ms = 0;
    uint64_t h2d = 0, d2h = 0;
    if (end_row > begin_row) {
        // the whole db stays resident (n_total x d x 4 bytes: 8 GB per million vectors at d = 2048)
        Dev V, nrm;
        const size_t vb = static_cast<size_t>(n_total) * d * 4;
        PW_RC(V.alloc(ctx, kWsB, vb));
        PW_RC(nrm.alloc(ctx, kWsNj, static_cast<size_t>(n_total) * 8));
        PW_RC(mvs_internal_upload(ctx, st, V.p, vectors, vb));
        PW_CUDA(ctx, cudaMemcpyAsync(nrm.p, norms_sq, static_cast<size_t>(n_total) * 8, cudaMemcpyHostToDevice, st));
        h2d = vb + static_cast<uint64_t>(n_total) * 8;
        Operand op, unused;
        op.v = V.as<int32_t>();
        op.n = n_total;
        const int rc = run_blocks(ctx, st, unused, begin_row, end_row, op, true, d, nrm.as<double>(), nrm.as<double>(), t, gemm_ms,
                                  filter_ms, prep_ms, d2h);
        if (rc != MVS_OK) return rc;
    }
    const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return finish(t, gemm_ms, filter_ms, prep_ms, wall, h2d, d2h, out);
}

}  // extern "C"
