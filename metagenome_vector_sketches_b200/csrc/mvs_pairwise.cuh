// mvs_pairwise.cuh -- sm_100a device code of the all-pairs comparison of projected vectors
// (reference src/pairwise_comp_optimized.cpp:57-160, live code :135-147).
//
// What is computed, for a block of rows A [mi x d] and a block of columns B [mj x d] (int32, row-major):
//     dot(i, j) = sum_k A[i][k] * B[j][k]            in int32 arithmetic, i.e. modulo 2^32 (Eigen MatrixXi product)
//     keep (i, j, dot) iff (int64)dot / d > 0.05 * (norms_i[i] + norms_j[j])      (integer division, double compare)
// and the kept triples are returned in row-major (i, then j) order, like the reference's double loop.
//
// How:
//   1. pairwise_dot_kernel: register-tiled int32 GEMM on the FMA-heavy pipe (IMAD is exact modulo 2^32 by
//      construction, for ANY int32 input; a tensor-core path would need 10 u8 x u8 limb products for that and is
//      left to a later round).  128 x 128 x 16 CTA tiles, 256 threads, 8 x 8 outputs per thread,
//      operands staged k-major in shared memory (double buffered, register prefetch), dense C written once.
//   2. pairwise_count_kernel: one warp per row applies the threshold and counts the kept columns.
//   3. pairwise_scan_kernel: exclusive scan of the row counts (single CTA; rows per call <= a few thousand).
//   4. pairwise_emit_kernel: one warp per row writes its kept triples in column order at the scanned offset
//      (ballot + popc prefix inside the warp): deterministic order, no atomics.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mvs {

constexpr int kPwBM = 128, kPwBN = 128, kPwBK = 16;
constexpr int kPwThreads = 256;

// keep-test of the reference (:140-143); norms are squared norms.  `(int64)dot / d > threshold` with the
// truncating quotient q.  For threshold >= 0 (always, with real norms) q > threshold <=> q >= floor(threshold) + 1
// <=> dot >= (floor(threshold) + 1) * d, which needs no division; a negative threshold takes the division.
__device__ __forceinline__ bool pw_keep(int32_t dot, int d, double ni, double nj)
{
    const double threshold = 0.05 * (ni + nj);
    if (threshold >= 0.0) {
        if (threshold >= 2147483647.0) return false;                  // |q| <= 2^31
        const long long need = (static_cast<long long>(threshold) + 1) * d;   // floor == trunc for >= 0
        return static_cast<long long>(dot) >= need;
    }
    const int q = dot / d;                                            // same truncation as int64 / int
    return static_cast<double>(q) > threshold;                        // NaN thresholds keep nothing, as in C++
}

// One 16-byte (4 x int32) slice of a row starting at column k (p = row + k); zero past the row end / for rows
// outside the matrix.  VEC: d % 4 == 0 and 16-byte aligned base, so k < d implies k + 3 < d.
template <bool VEC>
__device__ __forceinline__ int4 pw_load4(const int32_t* __restrict__ p, bool row_ok, int k, int d)
{
    int4 v = make_int4(0, 0, 0, 0);
    if (row_ok) {
        if (VEC) {
            if (k < d) v = *reinterpret_cast<const int4*>(p);
        } else {
            if (k + 0 < d) v.x = p[0];
            if (k + 1 < d) v.y = p[1];
            if (k + 2 < d) v.z = p[2];
            if (k + 3 < d) v.w = p[3];
        }
    }
    return v;
}

// C[i][j] = sum_k A[i][k] * B[j][k]  (mod 2^32).  grid = (ceil(mj/128), ceil(mi/128)).
template <bool VEC>
__global__ void __launch_bounds__(kPwThreads, 2)
pairwise_dot_kernel(const int32_t* __restrict__ A, int64_t mi, const int32_t* __restrict__ B, int64_t mj, int d,
                    int32_t* __restrict__ C, int64_t ldc)
{
    __shared__ __align__(16) int32_t As[2][kPwBK][kPwBM];
    __shared__ __align__(16) int32_t Bs[2][kPwBK][kPwBN];

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int64_t row0 = static_cast<int64_t>(blockIdx.y) * kPwBM;
    const int64_t col0 = static_cast<int64_t>(blockIdx.x) * kPwBN;

    // loader mapping: warp w stages k-quad (w & 3) of rows (w >> 2) * 64 + {lane, lane + 32}: a warp's 32
    // lanes hit 32 consecutive shared-memory words per store (conflict free); the 64-byte row slices
    // it reads from global are re-used from L2 by the other three warps of the quad.
    const int lk = (warp & 3) * 4;
    const int lr = (warp >> 2) * 64 + lane;

    const bool okA0 = row0 + lr < mi, okA1 = row0 + lr + 32 < mi, okB0 = col0 + lr < mj, okB1 = col0 + lr + 32 < mj;
    const int32_t* gA0 = A + (okA0 ? row0 + lr : 0) * d + lk;
    const int32_t* gA1 = A + (okA1 ? row0 + lr + 32 : 0) * d + lk;
    const int32_t* gB0 = B + (okB0 ? col0 + lr : 0) * d + lk;
    const int32_t* gB1 = B + (okB1 ? col0 + lr + 32 : 0) * d + lk;
    int4 pa0, pa1, pb0, pb1;
    auto fetch = [&](int k0) {
        pa0 = pw_load4<VEC>(gA0 + k0, okA0, k0 + lk, d);
        pa1 = pw_load4<VEC>(gA1 + k0, okA1, k0 + lk, d);
        pb0 = pw_load4<VEC>(gB0 + k0, okB0, k0 + lk, d);
        pb1 = pw_load4<VEC>(gB1 + k0, okB1, k0 + lk, d);
    };
    auto stage = [&](int buf) {
        As[buf][lk + 0][lr] = pa0.x; As[buf][lk + 1][lr] = pa0.y; As[buf][lk + 2][lr] = pa0.z; As[buf][lk + 3][lr] = pa0.w;
        As[buf][lk + 0][lr + 32] = pa1.x; As[buf][lk + 1][lr + 32] = pa1.y; As[buf][lk + 2][lr + 32] = pa1.z; As[buf][lk + 3][lr + 32] = pa1.w;
        Bs[buf][lk + 0][lr] = pb0.x; Bs[buf][lk + 1][lr] = pb0.y; Bs[buf][lk + 2][lr] = pb0.z; Bs[buf][lk + 3][lr] = pb0.w;
        Bs[buf][lk + 0][lr + 32] = pb1.x; Bs[buf][lk + 1][lr + 32] = pb1.y; Bs[buf][lk + 2][lr + 32] = pb1.z; Bs[buf][lk + 3][lr + 32] = pb1.w;
    };

    // compute mapping: thread (tx, ty) owns rows ty*4 + {0..3} and 64 + ty*4 + {0..3}, columns likewise with tx
    const int tx = tid & 15, ty = tid >> 4;
    int32_t acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0;

    const int ksteps = (d + kPwBK - 1) / kPwBK;
    fetch(0);
    stage(0);
    __syncthreads();
    for (int ks = 0; ks < ksteps; ++ks) {
        const int buf = ks & 1;
        if (ks + 1 < ksteps) fetch((ks + 1) * kPwBK);
#pragma unroll
        for (int k = 0; k < kPwBK; ++k) {
            const int4 a0 = *reinterpret_cast<const int4*>(&As[buf][k][ty * 4]);
            const int4 a1 = *reinterpret_cast<const int4*>(&As[buf][k][64 + ty * 4]);
            const int4 b0 = *reinterpret_cast<const int4*>(&Bs[buf][k][tx * 4]);
            const int4 b1 = *reinterpret_cast<const int4*>(&Bs[buf][k][64 + tx * 4]);
            const int32_t a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const int32_t b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    acc[i][j] = static_cast<int32_t>(static_cast<uint32_t>(acc[i][j]) +
                                                     static_cast<uint32_t>(a[i]) * static_cast<uint32_t>(b[j]));
        }
        if (ks + 1 < ksteps) {
            stage(buf ^ 1);                      // the other buffer was last read before the previous barrier
            __syncthreads();
        }
    }

    const bool vec_store = ((ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int64_t r = row0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (r >= mi) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t c = col0 + h * 64 + tx * 4;
            int32_t* dst = C + r * ldc + c;
            if (vec_store && c + 3 < mj) {
                *reinterpret_cast<int4*>(dst) = make_int4(acc[i][h * 4], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (c + j < mj) dst[j] = acc[i][h * 4 + j];
            }
        }
    }
}

// counts[i] = number of columns kept in row i.  One warp per row.
__global__ void __launch_bounds__(256)
pairwise_count_kernel(const int32_t* __restrict__ C, int64_t mi, int64_t mj, int64_t ldc, int d,
                      const double* __restrict__ norms_i, const double* __restrict__ norms_j,
                      unsigned long long* __restrict__ counts)
{
    const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= mi) return;
    const int lane = threadIdx.x & 31;
    const double ni = norms_i[row];
    const int32_t* c = C + row * ldc;
    unsigned int n = 0;
    for (int64_t j = lane; j < mj; j += 32) n += pw_keep(c[j], d, ni, norms_j[j]) ? 1u : 0u;
#pragma unroll
    for (int m = 16; m; m >>= 1) n += __shfl_xor_sync(0xffffffffu, n, m);
    if (lane == 0) counts[row] = n;
}

// In-place exclusive scan of counts[0..n), total in counts[n].  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024)
pairwise_scan_kernel(unsigned long long* __restrict__ counts, int64_t n)
{
    __shared__ unsigned long long warp_sums[32];
    __shared__ unsigned long long carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const unsigned long long v = i < n ? counts[i] : 0ull;
        unsigned long long x = v;
#pragma unroll
        for (int m = 1; m < 32; m <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, m);
            if (lane >= m) x += y;
        }
        if (lane == 31) warp_sums[warp] = x;
        __syncthreads();
        if (warp == 0) {
            unsigned long long s = warp_sums[lane];
#pragma unroll
            for (int m = 1; m < 32; m <<= 1) {
                const unsigned long long y = __shfl_up_sync(0xffffffffu, s, m);
                if (lane >= m) s += y;
            }
            warp_sums[lane] = s;                 // inclusive over warps
        }
        __syncthreads();
        const unsigned long long before = carry + (warp ? warp_sums[warp - 1] : 0ull) + (x - v);
        if (i < n) counts[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) counts[n] = carry;
}

// Row i writes its kept (row, col, dot) triples in column order starting at offsets[i].
__global__ void __launch_bounds__(256)
pairwise_emit_kernel(const int32_t* __restrict__ C, int64_t mi, int64_t mj, int64_t ldc, int d,
                     const double* __restrict__ norms_i, const double* __restrict__ norms_j,
                     const unsigned long long* __restrict__ offsets, int32_t row_base, int32_t col_base,
                     int32_t* __restrict__ rows, int32_t* __restrict__ cols, long long* __restrict__ values)
{
    const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= mi) return;
    const int lane = threadIdx.x & 31;
    const double ni = norms_i[row];
    const int32_t* c = C + row * ldc;
    unsigned long long pos = offsets[row];
    for (int64_t j0 = 0; j0 < mj; j0 += 32) {
        const int64_t j = j0 + lane;
        int32_t dot = 0;
        bool keep = false;
        if (j < mj) {
            dot = c[j];
            keep = pw_keep(dot, d, ni, norms_j[j]);
        }
        const unsigned int m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const unsigned long long at = pos + __popc(m & ((1u << lane) - 1u));
            rows[at] = row_base + static_cast<int32_t>(row);
            cols[at] = col_base + static_cast<int32_t>(j);
            values[at] = static_cast<long long>(dot);
        }
        pos += __popc(m);
    }
}

}  // namespace mvs
