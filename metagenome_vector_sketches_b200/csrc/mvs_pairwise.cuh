// mvs_pairwise.cuh -- sm_100a device code of the all-pairs comparison of projected vectors
// (reference src/pairwise_comp_optimized.cpp:57-160, live code :135-147).
//
// What is computed, for a block of rows A [mi x d] and a block of columns B [mj x d] (int32, row-major):
//     dot(i, j) = sum_k A[i][k] * B[j][k]            in int32 arithmetic, i.e. modulo 2^32 (Eigen MatrixXi product)
//     keep (i, j, dot) iff (int64)dot / d > 0.05 * (norms_i[i] + norms_j[j])      (integer division, double compare)
// and the kept triples are returned in row-major (i, then j) order, like the reference's double loop.
//
// How:
//   1a. the int32 operands are split into balanced base-256 limbs (pairwise_split_kernel) and the limb products run
//       on the tensor cores, combined modulo 2^32 by Horner's rule -- exact for any int32 input.  Default: the
//       tcgen05 kernels of mvs_pairwise_umma.cuh; here: pairwise_dot_imma_kernel, the same GEMM with warp-level
//       mma.sync (option "pairwise_path" = 2), see the section header below.
//   1b. pairwise_dot_kernel (cross-check path, option "pairwise_path" = 1): register-tiled int32 GEMM on the
//       FMA-heavy pipe (IMAD is exact modulo 2^32 by construction).  128 x 128 x 16 CTA tiles, 256 threads,
//       8 x 8 outputs per thread, operands staged k-major in shared memory (double buffered).
//   2.  pairwise_count_kernel / pairwise_rowsum_kernel / pairwise_scan_kernel / pairwise_emit_kernel: threshold +
//       ordered compaction of the C block into (row, col, dot) triples in the reference's loop order, no atomics.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mvs {

constexpr int kPwBM = 128, kPwBN = 128, kPwBK = 16;
constexpr int kPwThreads = 256;

// keep-test of the reference (:140-143); norms are squared norms.  `(int64)dot / d > threshold` with the
// truncating quotient q.  For 0 <= threshold < 2^31 (always, with real norms) q > threshold <=> q >= floor(threshold) + 1
// <=> dot >= (floor(threshold) + 1) * d: no division, and floor() without the conversion pipe -- adding 2^52 with
// round-down leaves floor(threshold) in the low mantissa word.  Per element that is 2 FP64 ops for the threshold
// (exactly the reference's expression), one round-down add, one wide multiply-add and a 64-bit compare.
// Anything else (negative, -0, NaN, huge thresholds) takes the literal path.
__device__ __noinline__ bool pw_keep_literal(int32_t dot, int d, double threshold)
{
    if (threshold >= 2147483648.0) return false;                      // |q| <= 2^31
    const int q = dot / d;                                            // same truncation as int64 / int
    return static_cast<double>(q) > threshold;                        // NaN thresholds keep nothing, as in C++
}
__device__ __forceinline__ bool pw_keep(int32_t dot, int d, double ni, double nj)
{
    const double threshold = 0.05 * (ni + nj);
    if (static_cast<unsigned>(__double2hiint(threshold)) < 0x41E00000u) {       // +0 <= threshold < 2^31
        const unsigned fl = static_cast<unsigned>(__double2loint(__dadd_rd(threshold, 4503599627370496.0)));
        const unsigned long long need = static_cast<unsigned long long>(fl) * static_cast<unsigned>(d) + static_cast<unsigned>(d);
        return dot >= 0 && static_cast<unsigned long long>(static_cast<unsigned>(dot)) >= need;
    }
    return pw_keep_literal(dot, d, threshold);
}

#ifdef MVS_EXPERIMENTS                           // cross-check GEMM of the experiment build only (option "pairwise_path" = 1)
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

#endif  // MVS_EXPERIMENTS

// ------------------------------------------------------------------------------------------------------
// Tensor-core path: int32 = up to 4 balanced base-256 limbs, v = sum_p 256^p * s_p with s_p in [-128, 127]
// (exact modulo 2^32), so   dot mod 2^32 = sum_{p + q <= 3} 256^(p+q) * (A_p . B_q)   with int8 x int8 -> int32
// limb products on the tensor cores (IMMA m16n8k32 via mma.sync: 572 TMAC/s measured on B200, 31x the IMAD
// pipe; tools/kexp/mma_probe.cu).  Sketch entries are small (|v| ~ sqrt(#hashes)), so P = 2 limbs (4 limb
// products) is the common case; P is chosen from the min/max of the operands and P = 4 (10 products) is exact
// for ANY int32 input.  One accumulator set serves all products by Horner's rule: the limb pairs are run
// in order of decreasing p + q and the accumulators are shifted left by 8 between weights; a limb product
// accumulates at most d * 2^14 per weight, and int32 wrap-around is exactly what the reference computes.
// ------------------------------------------------------------------------------------------------------
constexpr int kImBM = 128, kImBN = 128, kImBK = 64;       // k in BYTES (= int8 elements) per stage
constexpr int kImPitch = kImBK + 16;                      // 80-byte rows: ldmatrix reads are bank-conflict free
constexpr int kImStages = 4;
constexpr int kImThreads = 256;
constexpr int kImStageBytes = (kImBM + kImBN) * kImPitch; // 20480
constexpr int kImSmemBytes = kImStages * kImStageBytes;   // 81920

struct ImmaPlan {                                         // limb pairs in order of decreasing weight
    int npairs;
    unsigned char p[10], q[10], shift_before[10];         // shift_before: accumulators <<= 8 before this pair
};

// V [n x d] int32 -> P planes of int8 [n_pad x k_pad] (row-major, zero padded).  One thread = 4 consecutive k per
// step of a grid-stride loop.  mm != nullptr: also the min / max of V (two words initialised to INT_MAX / INT_MIN),
// so that the usual case -- the guessed P turns out right -- needs one pass over V instead of two.
__global__ void __launch_bounds__(256)
pairwise_split_kernel(const int32_t* __restrict__ V, int64_t n, int d, int P, int8_t* __restrict__ planes, int64_t n_pad, int k_pad,
                      int* __restrict__ mm)
{
    const int64_t quads = static_cast<int64_t>(k_pad) / 4;
    int lo = 2147483647, hi = -2147483647 - 1;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < n_pad * quads;
         t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t row = t / quads;
        const int k = static_cast<int>(t - row * quads) * 4;
        long long v[4] = {0, 0, 0, 0};
        if (row < n)
            for (int e = 0; e < 4; ++e)
                if (k + e < d) {
                    const int x = V[row * d + k + e];
                    v[e] = x;
                    lo = min(lo, x);
                    hi = max(hi, x);
                }
        for (int p = 0; p < P; ++p) {
            uint32_t word = 0;
            for (int e = 0; e < 4; ++e) {
                const int s = static_cast<int>(static_cast<signed char>(v[e] & 0xFF));
                word |= static_cast<uint32_t>(s & 0xFF) << (8 * e);
                v[e] = (v[e] - s) >> 8;                              // exact: v - s is a multiple of 256
            }
            *reinterpret_cast<uint32_t*>(planes + (static_cast<int64_t>(p) * n_pad + row) * k_pad + k) = word;
        }
    }
    if (mm) {
        __shared__ int s_lo[8], s_hi[8];
        lo = __reduce_min_sync(0xffffffffu, lo);
        hi = __reduce_max_sync(0xffffffffu, hi);
        if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < 8; ++w) { lo = min(lo, s_lo[w]); hi = max(hi, s_hi[w]); }
            atomicMin(mm, lo);
            atomicMax(mm + 1, hi);
        }
    }
}

#ifdef MVS_EXPERIMENTS                           // warp-level mma.sync GEMM of the experiment build only (option "pairwise_path" = 2)
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void imma16832(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// C[i][j] = sum_k A[i][k] * B[j][k] (mod 2^32) from limb planes.  grid = (nj_pad/128, mi_pad/128); planes are
// padded to multiples of 128 rows and 64 k, so the main loop has no edge handling; only the C store is guarded.
// 8 warps = 2 (m) x 4 (n); a warp owns 64 x 32 outputs = 4 x 4 IMMA tiles (64 accumulator registers).
__global__ void __launch_bounds__(kImThreads, 2)
pairwise_dot_imma_kernel(const int8_t* __restrict__ Ap, int64_t a_plane_stride, const int8_t* __restrict__ Bp,
                         int64_t b_plane_stride, int k_pad, ImmaPlan plan, int64_t mi, int64_t mj,
                         int32_t* __restrict__ C, int64_t ldc)
{
    extern __shared__ __align__(128) unsigned char im_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 1, wn = warp >> 1;
    const int64_t row0 = static_cast<int64_t>(blockIdx.y) * kImBM;
    const int64_t col0 = static_cast<int64_t>(blockIdx.x) * kImBN;
    const uint32_t smem_base = static_cast<uint32_t>(__cvta_generic_to_shared(im_smem));

    const int nk = k_pad / kImBK;
    const int steps = plan.npairs * nk;

    // the plan in shared memory (dynamic indexing of kernel parameters would go through local memory)
    __shared__ int s_p[10], s_q[10], s_shift[10];
    if (tid < 10) { s_p[tid] = plan.p[tid]; s_q[tid] = plan.q[tid]; s_shift[tid] = plan.shift_before[tid]; }
    __syncthreads();

    // loader: 1024 16-byte chunks per stage (512 of A, 512 of B), 4 per thread: thread -> rows tid / 4 and tid / 4 + 64,
    // k-chunk tid % 4.  Producer state (pair, k-step, stage) advances by one per issue: no division in the loop.
    const int lrow = tid >> 2, lkc = (tid & 3) * 16;
    const int64_t a_thread = (row0 + lrow) * k_pad + lkc, b_thread = (col0 + lrow) * k_pad + lkc;
    const int64_t half = static_cast<int64_t>(64) * k_pad;
    const uint32_t s_thread = static_cast<uint32_t>(lrow * kImPitch + lkc);
    int p_pair = 0, p_kk = 0, p_stage = 0;
    const int8_t* a_src = Ap + s_p[0] * a_plane_stride + a_thread;
    const int8_t* b_src = Bp + s_q[0] * b_plane_stride + b_thread;
    auto issue = [&]() {
        const uint32_t st = smem_base + p_stage * kImStageBytes + s_thread;
        cp_async16(st, a_src);
        cp_async16(st + 64 * kImPitch, a_src + half);
        cp_async16(st + kImBM * kImPitch, b_src);
        cp_async16(st + (kImBM + 64) * kImPitch, b_src + half);
        a_src += kImBK;
        b_src += kImBK;
        if (++p_kk == nk) {
            p_kk = 0;
            if (++p_pair < plan.npairs) {
                a_src = Ap + s_p[p_pair] * a_plane_stride + a_thread;
                b_src = Bp + s_q[p_pair] * b_plane_stride + b_thread;
            }
        }
        if (++p_stage == kImStages) p_stage = 0;
    };

    int acc[4][4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[i][j][e] = 0;

    // ldmatrix lane addressing (see header of this section): A: row = (l & 7) + 8 * ((l >> 3) & 1), k += 16 * (l >> 4)
    //                                                         B: n   = (l & 7) + 8 * (l >> 4),       k += 16 * ((l >> 3) & 1)
    const uint32_t a_lane = static_cast<uint32_t>((wm * 64 + (lane & 7) + ((lane >> 3) & 1) * 8) * kImPitch + (lane >> 4) * 16);
    const uint32_t b_lane = static_cast<uint32_t>((kImBM + wn * 32 + (lane & 7) + (lane >> 4) * 8) * kImPitch + ((lane >> 3) & 1) * 16);

#pragma unroll
    for (int t = 0; t < kImStages - 1; ++t) {
        if (t < steps) issue();
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    int c_pair = 0, c_kk = 0, c_stage = 0;
    for (int t = 0; t < steps; ++t) {
        asm volatile("cp.async.wait_group %0;" ::"n"(kImStages - 2) : "memory");
        __syncthreads();
        if (t + kImStages - 1 < steps) issue();
        asm volatile("cp.async.commit_group;" ::: "memory");

        if (c_kk == 0 && s_shift[c_pair]) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int e = 0; e < 4; ++e) acc[i][j][e] = static_cast<int>(static_cast<uint32_t>(acc[i][j][e]) << 8);
        }
        const uint32_t st = smem_base + c_stage * kImStageBytes;
#pragma unroll
        for (int ks = 0; ks < kImBK / 32; ++ks) {
            uint32_t a[4][4], b[4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i) ldmatrix_x4(st + a_lane + i * 16 * kImPitch + ks * 32, a[i][0], a[i][1], a[i][2], a[i][3]);
#pragma unroll
            for (int j = 0; j < 2; ++j)
                ldmatrix_x4(st + b_lane + j * 16 * kImPitch + ks * 32, b[2 * j][0], b[2 * j][1], b[2 * j + 1][0], b[2 * j + 1][1]);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) imma16832(acc[i][j], a[i], b[j][0], b[j][1]);
        }
        if (++c_kk == nk) { c_kk = 0; ++c_pair; }
        if (++c_stage == kImStages) c_stage = 0;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");

    // accumulator layout of m16n8: c0,c1 -> (row g, cols 2t, 2t+1), c2,c3 -> (row g + 8, same cols); g = lane / 4, t = lane % 4
    const int g = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t r = row0 + wm * 64 + i * 16 + g + h * 8;
            if (r >= mi) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t c = col0 + wn * 32 + j * 8 + tq * 2;
                int32_t* dst = C + r * ldc + c;
                if (c + 1 < mj) *reinterpret_cast<int2*>(dst) = make_int2(acc[i][j][2 * h], acc[i][j][2 * h + 1]);   // ldc even, c even
                else if (c < mj) dst[0] = acc[i][j][2 * h];
            }
        }
}

#endif  // MVS_EXPERIMENTS

// ------------------------------------------------------------------------------------------------------
// Threshold + ordered compaction.  The columns of a row are cut into segments of 128; one warp owns one
// (row, segment): lane L holds columns 4L..4L+3 of the segment (one 16-byte load).
//   pairwise_count_kernel : segcount[row][seg] = kept columns of the segment
//   pairwise_rowsum_kernel / pairwise_scan_kernel : rowoff[row] = exclusive scan of the row totals, total in rowoff[mi]
//   pairwise_emit_kernel  : non-empty segments write their triples at rowoff[row] + sum of the earlier
//                           segments of the row + rank inside the segment: (row, col) order, no atomics;
//                           empty segments (almost all of them) do not read C again.
// ------------------------------------------------------------------------------------------------------
constexpr int kPwSeg = 128;

__device__ __forceinline__ unsigned pw_keep_mask4(const int32_t* __restrict__ crow, int64_t j, int64_t mj, int d, double ni,
                                                 const double* __restrict__ norms_j, int4& v)
{
    v = make_int4(0, 0, 0, 0);
    unsigned m = 0;
    if (j + 3 < mj) {                                                // ldc % 4 == 0 and C 16-byte aligned (host)
        v = *reinterpret_cast<const int4*>(crow + j);
        m |= pw_keep(v.x, d, ni, norms_j[j]) ? 1u : 0u;
        m |= pw_keep(v.y, d, ni, norms_j[j + 1]) ? 2u : 0u;
        m |= pw_keep(v.z, d, ni, norms_j[j + 2]) ? 4u : 0u;
        m |= pw_keep(v.w, d, ni, norms_j[j + 3]) ? 8u : 0u;
    } else {
        int32_t* e = &v.x;
        for (int t = 0; t < 4; ++t)
            if (j + t < mj) {
                e[t] = crow[j + t];
                m |= pw_keep(e[t], d, ni, norms_j[j + t]) ? (1u << t) : 0u;
            }
    }
    return m;
}

// keep-mask of four columns whose squared norms are already in registers (`valid` = columns that exist)
__device__ __forceinline__ unsigned pw_keep_mask4r(const int4& v, int d, double ni, const double (&nj)[4], unsigned valid)
{
    unsigned m = 0;
    m |= pw_keep(v.x, d, ni, nj[0]) ? 1u : 0u;
    m |= pw_keep(v.y, d, ni, nj[1]) ? 2u : 0u;
    m |= pw_keep(v.z, d, ni, nj[2]) ? 4u : 0u;
    m |= pw_keep(v.w, d, ni, nj[3]) ? 8u : 0u;
    return m & valid;
}

// four columns of one row of C starting at column j (zeros past the row end); ldc % 4 == 0, C 16-byte aligned (host)
__device__ __forceinline__ int4 pw_load_c4(const int32_t* __restrict__ crow, int64_t j, int64_t mj)
{
    if (j + 3 < mj) return *reinterpret_cast<const int4*>(crow + j);
    int4 v = make_int4(0, 0, 0, 0);
    if (j + 0 < mj) v.x = crow[j + 0];
    if (j + 1 < mj) v.y = crow[j + 1];
    if (j + 2 < mj) v.z = crow[j + 2];
    return v;
}

constexpr int kPwRows = 8;
#ifdef MVS_EXPERIMENTS                           // separate count pass (option "pairwise_fuse" = 2); the shipped path filters in the GEMM epilogue
// One warp = one 128-column segment of kPwRows consecutive rows: the lane keeps its four column norms in registers,
// has all kPwRows 16-byte loads of C in flight at once (the pass is HBM-bound: C is read exactly once), and the
// per-row counts are reduced with one redux each.

__global__ void __launch_bounds__(256)
pairwise_count_kernel(const int32_t* __restrict__ C, int64_t mi, int64_t mj, int64_t ldc, int d,
                      const double* __restrict__ norms_i, const double* __restrict__ norms_j,
                      uint32_t* __restrict__ segcount, int nseg)
{
    const int64_t w = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t groups = (mi + kPwRows - 1) / kPwRows;
    if (w >= groups * nseg) return;
    const int64_t row0 = (w / nseg) * kPwRows;
    const int seg = static_cast<int>(w % nseg);
    const int lane = threadIdx.x & 31;
    const int64_t j = static_cast<int64_t>(seg) * kPwSeg + lane * 4;
    double nj[4];
    unsigned valid = 0;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        nj[t] = j + t < mj ? norms_j[j + t] : 0.0;
        valid |= j + t < mj ? (1u << t) : 0u;
    }
    int4 v[kPwRows];
    const int32_t* p = C + row0 * ldc + j;
    if (row0 + kPwRows <= mi && j + 3 < mj) {                        // everywhere but the last row group / last columns
#pragma unroll
        for (int r = 0; r < kPwRows; ++r) v[r] = *reinterpret_cast<const int4*>(p + r * ldc);
    } else {
#pragma unroll
        for (int r = 0; r < kPwRows; ++r) v[r] = row0 + r < mi ? pw_load_c4(C + (row0 + r) * ldc, j, mj) : make_int4(0, 0, 0, 0);
    }
#pragma unroll
    for (int r = 0; r < kPwRows; ++r) {
        if (row0 + r >= mi) break;                                   // warp-uniform
        const unsigned n = __reduce_add_sync(0xffffffffu, static_cast<unsigned>(__popc(pw_keep_mask4r(v[r], d, norms_i[row0 + r], nj, valid))));
        if (lane == 0) segcount[(row0 + r) * nseg + seg] = n;
    }
}
#endif  // MVS_EXPERIMENTS

// One warp per row: reads the row's segment counts (coalesced), replaces every count marked `maybe` -- a segment the
// GEMM's fused pre-filter (mvs_pairwise_umma.cuh) stored without deciding it -- by the exact count, and writes the
// row total to rowoff[row].  After the separate count pass nothing is marked and this is just the row sum.
__global__ void __launch_bounds__(256)
pairwise_rowsum_kernel(const int32_t* __restrict__ C, int64_t mi, int64_t mj, int64_t ldc, int d,
                       const double* __restrict__ norms_i, const double* __restrict__ norms_j,
                       uint32_t* __restrict__ segcount, int nseg, uint32_t maybe, unsigned long long* __restrict__ rowoff)
{
    const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= mi) return;
    const int lane = threadIdx.x & 31;
    unsigned long long total = 0;
    for (int base = 0; base < nseg; base += 32) {
        const int seg = base + lane;
        unsigned c = seg < nseg ? segcount[row * nseg + seg] : 0u;
        unsigned todo = __ballot_sync(0xffffffffu, c == maybe);
        while (todo) {                                               // warp-uniform
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            int4 v;
            const int64_t j = static_cast<int64_t>(base + src) * kPwSeg + lane * 4;
            const unsigned n = __reduce_add_sync(0xffffffffu, static_cast<unsigned>(__popc(pw_keep_mask4(C + row * ldc, j, mj, d, norms_i[row], norms_j, v))));
            if (lane == src) {
                c = n;
                segcount[row * nseg + seg] = n;
            }
        }
        total += c;
    }
#pragma unroll
    for (int m = 16; m; m >>= 1) total += __shfl_xor_sync(0xffffffffu, total, m);
    if (lane == 0) rowoff[row] = total;
}

// In-place exclusive scan of rowoff[0..n), total in rowoff[n].  One CTA of 1024 threads (n <= 16384 rows per call: a few sequential chunks).
__global__ void __launch_bounds__(1024)
pairwise_scan_kernel(unsigned long long* __restrict__ rowoff, int64_t n)
{
    __shared__ unsigned long long warp_sums[32];
    __shared__ unsigned long long carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const unsigned long long v = i < n ? rowoff[i] : 0ull;
        unsigned long long x = v;
#pragma unroll
        for (int m = 1; m < 32; m <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, m);
            if (lane >= m) x += y;
        }
        if (lane == 31) warp_sums[warp] = x;
        __syncthreads();
        if (warp == 0) {
            unsigned long long t = warp_sums[lane];
#pragma unroll
            for (int m = 1; m < 32; m <<= 1) {
                const unsigned long long y = __shfl_up_sync(0xffffffffu, t, m);
                if (lane >= m) t += y;
            }
            warp_sums[lane] = t;                 // inclusive over warps
        }
        __syncthreads();
        const unsigned long long before = carry + (warp ? warp_sums[warp - 1] : 0ull) + (x - v);
        if (i < n) rowoff[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) rowoff[n] = carry;
}

// One warp per row: walks the row's segment counts and writes the triples of the non-empty segments (about 1 % of
// them) at rowoff[row] + kept entries to the left: (row, col) order, no atomics.
__global__ void __launch_bounds__(256)
pairwise_emit_kernel(const int32_t* __restrict__ C, int64_t mi, int64_t mj, int64_t ldc, int d,
                     const double* __restrict__ norms_i, const double* __restrict__ norms_j,
                     const uint32_t* __restrict__ segcount, int nseg, const unsigned long long* __restrict__ rowoff,
                     int32_t row_base, int32_t col_base, unsigned long long cap,
                     int32_t* __restrict__ rows, int32_t* __restrict__ cols, long long* __restrict__ values)
{
    const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= mi) return;
    if (rowoff[mi] > cap) return;                                    // output buffers too small: the host grows them and re-launches
    const int lane = threadIdx.x & 31;
    unsigned long long run = rowoff[row];                             // position of the next triple of this row
    for (int base = 0; base < nseg; base += 32) {
        const int seg = base + lane;
        const unsigned c = seg < nseg ? segcount[row * nseg + seg] : 0u;
        unsigned todo = __ballot_sync(0xffffffffu, c != 0);
        if (todo == 0) continue;
        unsigned incl = c;                                            // inclusive scan of the 32 counts
#pragma unroll
        for (int m = 1; m < 32; m <<= 1) {
            const unsigned y = __shfl_up_sync(0xffffffffu, incl, m);
            if (lane >= m) incl += y;
        }
        while (todo) {                                               // warp-uniform
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const unsigned before = __shfl_sync(0xffffffffu, incl - c, src);
            int4 v;
            const int64_t j = static_cast<int64_t>(base + src) * kPwSeg + lane * 4;
            const unsigned m4 = pw_keep_mask4(C + row * ldc, j, mj, d, norms_i[row], norms_j, v);
            unsigned mine = __popc(m4), inc2 = mine;
#pragma unroll
            for (int m = 1; m < 32; m <<= 1) {
                const unsigned y = __shfl_up_sync(0xffffffffu, inc2, m);
                if (lane >= m) inc2 += y;
            }
            unsigned long long at = run + before + (inc2 - mine);
            const int32_t e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int t = 0; t < 4; ++t)
                if (m4 & (1u << t)) {
                    rows[at] = row_base + static_cast<int32_t>(row);
                    cols[at] = col_base + static_cast<int32_t>(j + t);
                    values[at] = static_cast<long long>(e[t]);
                    ++at;
                }
        }
        run += __shfl_sync(0xffffffffu, incl, 31);
    }
}

}  // namespace mvs
