// mvs_pairwise.cu -- host side of the all-pairs comparison entry points of include/mvs_b200.h
// (SURVEY 8f rank 4).  Replaces, for the GPU, compute_sparse_dot_products_optimized
// (reference src/pairwise_comp_optimized.cpp:57-160) and the block loops that drive it (:949-983).
// Device code: mvs_pairwise.cuh.  No CPU implementation exists here: without a device the calls fail.
#include "../../include/mvs_b200.h"
#include "mvs_internal.h"
#include "mvs_pairwise.cuh"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace {

struct Dev {                                     // device allocation freed on scope exit
    void* p = nullptr;
    ~Dev() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16); }
    template <typename T> T* as() const { return static_cast<T*>(p); }
};

struct Triples {
    std::vector<int32_t> rows, cols;
    std::vector<int64_t> values;
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

// One block of rows against one block of columns, all operands resident in HBM.  Appends the kept
// triples (row_base + i, col_base + j, dot) in (i, j) order to `out`; adds the CUDA-event times.
int pairwise_core(mvs_ctx* ctx, cudaStream_t st, const int32_t* d_A, int64_t mi, const int32_t* d_B, int64_t mj, int d,
                  const double* d_ni, const double* d_nj, int32_t row_base, int32_t col_base, int32_t* d_C,
                  unsigned long long* d_counts, Triples& out, double& gemm_ms, double& filter_ms, uint64_t& d2h_bytes)
{
    if (mi == 0 || mj == 0) return MVS_OK;
    cudaEvent_t e0, e1, e2;
    PW_CUDA(ctx, cudaEventCreate(&e0));
    PW_CUDA(ctx, cudaEventCreate(&e1));
    PW_CUDA(ctx, cudaEventCreate(&e2));
    struct EvGuard { cudaEvent_t a, b, c; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); cudaEventDestroy(c); } } guard{e0, e1, e2};

    const int64_t ldc = (mj + 3) & ~int64_t{3};
    const dim3 grid(static_cast<unsigned>((mj + mvs::kPwBN - 1) / mvs::kPwBN), static_cast<unsigned>((mi + mvs::kPwBM - 1) / mvs::kPwBM));
    const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_A) & 15) == 0) && ((reinterpret_cast<uintptr_t>(d_B) & 15) == 0);
    PW_CUDA(ctx, cudaEventRecord(e0, st));
    if (vec) mvs::pairwise_dot_kernel<true><<<grid, mvs::kPwThreads, 0, st>>>(d_A, mi, d_B, mj, d, d_C, ldc);
    else mvs::pairwise_dot_kernel<false><<<grid, mvs::kPwThreads, 0, st>>>(d_A, mi, d_B, mj, d, d_C, ldc);
    PW_CUDA(ctx, cudaGetLastError());
    PW_CUDA(ctx, cudaEventRecord(e1, st));
    const unsigned row_ctas = static_cast<unsigned>((mi + 7) / 8);
    mvs::pairwise_count_kernel<<<row_ctas, 256, 0, st>>>(d_C, mi, mj, ldc, d, d_ni, d_nj, d_counts);
    PW_CUDA(ctx, cudaGetLastError());
    mvs::pairwise_scan_kernel<<<1, 1024, 0, st>>>(d_counts, mi);
    PW_CUDA(ctx, cudaGetLastError());
    unsigned long long total = 0;
    PW_CUDA(ctx, cudaMemcpyAsync(&total, d_counts + mi, sizeof total, cudaMemcpyDeviceToHost, st));
    PW_CUDA(ctx, cudaStreamSynchronize(st));
    mvs_internal_count_launches(ctx, 3);
    d2h_bytes += sizeof total;

    if (total) {
        Dev rows, cols, vals;
        PW_CUDA(ctx, rows.alloc(total * sizeof(int32_t)));
        PW_CUDA(ctx, cols.alloc(total * sizeof(int32_t)));
        PW_CUDA(ctx, vals.alloc(total * sizeof(int64_t)));
        mvs::pairwise_emit_kernel<<<row_ctas, 256, 0, st>>>(d_C, mi, mj, ldc, d, d_ni, d_nj, d_counts, row_base, col_base,
                                                           rows.as<int32_t>(), cols.as<int32_t>(), vals.as<long long>());
        PW_CUDA(ctx, cudaGetLastError());
        mvs_internal_count_launches(ctx, 1);
        PW_CUDA(ctx, cudaEventRecord(e2, st));
        const size_t at = out.rows.size();
        out.rows.resize(at + total);
        out.cols.resize(at + total);
        out.values.resize(at + total);
        PW_CUDA(ctx, cudaMemcpyAsync(out.rows.data() + at, rows.p, total * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PW_CUDA(ctx, cudaMemcpyAsync(out.cols.data() + at, cols.p, total * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        PW_CUDA(ctx, cudaMemcpyAsync(out.values.data() + at, vals.p, total * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        PW_CUDA(ctx, cudaStreamSynchronize(st));
        d2h_bytes += total * 16;
    } else {
        PW_CUDA(ctx, cudaEventRecord(e2, st));
        PW_CUDA(ctx, cudaStreamSynchronize(st));
    }
    float a = 0.f, b = 0.f;
    PW_CUDA(ctx, cudaEventElapsedTime(&a, e0, e1));
    PW_CUDA(ctx, cudaEventElapsedTime(&b, e1, e2));
    gemm_ms += a;
    filter_ms += b;
    return MVS_OK;
}

int finish(Triples& t, double gemm_ms, double filter_ms, double wall_ms, uint64_t h2d, uint64_t d2h, mvs_sparse_result** out)
{
    mvs_sparse_result* r = static_cast<mvs_sparse_result*>(std::calloc(1, sizeof(mvs_sparse_result)));
    if (!r) return MVS_ERR_OOM;
    const size_t n = t.rows.size();
    r->n = static_cast<int64_t>(n);
    r->rows = static_cast<int32_t*>(std::malloc(sizeof(int32_t) * (n ? n : 1)));
    r->cols = static_cast<int32_t*>(std::malloc(sizeof(int32_t) * (n ? n : 1)));
    r->values = static_cast<int64_t*>(std::malloc(sizeof(int64_t) * (n ? n : 1)));
    if (!r->rows || !r->cols || !r->values) {
        mvs_free_sparse_result(r);
        return MVS_ERR_OOM;
    }
    if (n) {
        std::memcpy(r->rows, t.rows.data(), n * sizeof(int32_t));
        std::memcpy(r->cols, t.cols.data(), n * sizeof(int32_t));
        std::memcpy(r->values, t.values.data(), n * sizeof(int64_t));
    }
    r->gemm_ms = gemm_ms;
    r->filter_ms = filter_ms;
    r->wall_ms = wall_ms;
    r->h2d_bytes = h2d;
    r->d2h_bytes = d2h;
    *out = r;
    return MVS_OK;
}

// rows per GEMM call: the dense int32 C block stays <= 1 GiB
int64_t rows_per_call(int64_t mj)
{
    const int64_t ldc = (mj + 3) & ~int64_t{3};
    int64_t rb = (int64_t{1} << 30) / (4 * std::max<int64_t>(ldc, 1));
    rb = std::max<int64_t>(128, std::min<int64_t>(8192, rb / 128 * 128));
    return rb;
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
    double gemm_ms = 0, filter_ms = 0;
    uint64_t h2d = 0, d2h = 0;
    if (mi > 0 && mj > 0) {
        Dev A, B, ni, nj, C, counts;
        const size_t ab = static_cast<size_t>(mi) * d * 4, bb = static_cast<size_t>(mj) * d * 4;
        const int64_t rb = std::min<int64_t>(rows_per_call(mj), mi);
        const int64_t ldc = (mj + 3) & ~int64_t{3};
        PW_CUDA(ctx, A.alloc(ab));
        PW_CUDA(ctx, B.alloc(bb));
        PW_CUDA(ctx, ni.alloc(static_cast<size_t>(mi) * 8));
        PW_CUDA(ctx, nj.alloc(static_cast<size_t>(mj) * 8));
        PW_CUDA(ctx, C.alloc(static_cast<size_t>(rb) * ldc * 4));
        PW_CUDA(ctx, counts.alloc(static_cast<size_t>(rb + 1) * 8));
        PW_CUDA(ctx, cudaMemcpyAsync(A.p, block_i, ab, cudaMemcpyHostToDevice, st));
        PW_CUDA(ctx, cudaMemcpyAsync(B.p, block_j, bb, cudaMemcpyHostToDevice, st));
        PW_CUDA(ctx, cudaMemcpyAsync(ni.p, norms_i, static_cast<size_t>(mi) * 8, cudaMemcpyHostToDevice, st));
        PW_CUDA(ctx, cudaMemcpyAsync(nj.p, norms_j, static_cast<size_t>(mj) * 8, cudaMemcpyHostToDevice, st));
        h2d = ab + bb + static_cast<uint64_t>(mi + mj) * 8;
        for (int64_t r0 = 0; r0 < mi; r0 += rb) {
            const int64_t m = std::min(rb, mi - r0);
            const int rc = pairwise_core(ctx, st, A.as<int32_t>() + r0 * d, m, B.as<int32_t>(), mj, d, ni.as<double>() + r0,
                                         nj.as<double>(), static_cast<int32_t>(r0), 0, C.as<int32_t>(),
                                         counts.as<unsigned long long>(), t, gemm_ms, filter_ms, d2h);
            if (rc != MVS_OK) return rc;
        }
    }
    const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return finish(t, gemm_ms, filter_ms, wall, h2d, d2h, out);
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
    double gemm_ms = 0, filter_ms = 0;
    uint64_t h2d = 0, d2h = 0;
    const int64_t mi = end_row - begin_row;
    if (mi > 0) {
        // the whole db stays resident (n_total x d x 4 bytes: 8 GB per million vectors at d = 2048)
        Dev V, nrm, C, counts;
        const size_t vb = static_cast<size_t>(n_total) * d * 4;
        const int64_t rb = std::min<int64_t>(rows_per_call(n_total), mi);
        const int64_t ldc = (n_total + 3) & ~int64_t{3};
        PW_CUDA(ctx, V.alloc(vb));
        PW_CUDA(ctx, nrm.alloc(static_cast<size_t>(n_total) * 8));
        PW_CUDA(ctx, C.alloc(static_cast<size_t>(rb) * ldc * 4));
        PW_CUDA(ctx, counts.alloc(static_cast<size_t>(rb + 1) * 8));
        PW_CUDA(ctx, cudaMemcpyAsync(V.p, vectors, vb, cudaMemcpyHostToDevice, st));
        PW_CUDA(ctx, cudaMemcpyAsync(nrm.p, norms_sq, static_cast<size_t>(n_total) * 8, cudaMemcpyHostToDevice, st));
        h2d = vb + static_cast<uint64_t>(n_total) * 8;
        for (int64_t r0 = begin_row; r0 < end_row; r0 += rb) {
            const int64_t m = std::min(rb, end_row - r0);
            const int rc = pairwise_core(ctx, st, V.as<int32_t>() + r0 * d, m, V.as<int32_t>(), n_total, d,
                                         nrm.as<double>() + r0, nrm.as<double>(), static_cast<int32_t>(r0), 0,
                                         C.as<int32_t>(), counts.as<unsigned long long>(), t, gemm_ms, filter_ms, d2h);
            if (rc != MVS_OK) return rc;
        }
    }
    const double wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return finish(t, gemm_ms, filter_ms, wall, h2d, d2h, out);
}

}  // extern "C"
