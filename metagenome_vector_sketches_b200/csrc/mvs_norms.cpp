// mvs_norms.cpp -- the norm column of vector_norms.txt, bit for bit as the reference computes it.
//
// reference src/project_everything.cpp:327-330
//     VectorXf vec_f = pair.second.cast<float>() / std::sqrt(static_cast<float>(dimension));
//     double norm = vec_f.norm();      norm_out << base_name << " " << norm << "\n";
// built with -O3 -march=native -ffast-math (CMakeLists.txt:8) against the vendored Eigen 3.2.10
// (include/Eigen, SSE packets of 4 floats only -- AVX came with Eigen 3.3).  What the compiled
// reference executes (read off the disassembly of the reference binary, g++ 13, both its
// -march=native and x86-64-v3 builds, which print identical files):
//     inv = rsqrtss(float(d)), one Newton-Raphson step:  t = d*r;  t = fma(t, r, -3);  inv = t * (r * -0.5)
//     x_j = float(v_j) * inv
//     Eigen redux, LinearVectorizedTraversal, two packet accumulators (see mvs_kernels.cuh
//     reference_norm_row for the order), multiply-adds contracted to FMA, hadd, scalar FMA tail,
//     sqrtss, widened to double and printed with the default ostream precision (%g).
// The downstream tools square the PRINTED value (src/pairwise_comp_optimized.cpp:899-900) and use it
// in the keep-test (:139-141) and the 8-bit Jaccard (:660-665), so the 6th digit matters.
//
// rsqrtss is an approximation whose bits are the CPU's own (Intel and AMD differ): "identical to the
// reference on the same host" means executing the same instruction here.  mvs_reference_inv_sqrt_d
// is therefore host code; the device epilogue (finalize_kernel) receives the value as an argument
// and does everything else.  mvs_reference_norms is the same computation on the host for callers
// that hold int32 rows and no GPU (and the independent check of the device code in the tests).
#include "../../include/mvs_b200.h"

#include <cmath>
#include <cstdint>

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
#define MVS_X86 1
#else
#define MVS_X86 0
#endif

namespace {

#if MVS_X86
__attribute__((target("fma"))) float inv_sqrt_fma(float x)
{
    const float r = _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(x)));
    float t = x * r;
    t = __builtin_fmaf(t, r, -3.0f);
    return t * (r * -0.5f);
}
#endif

// One row.  The body exists twice: compiled for the FMA instruction (every x86-64-v3 host) and with
// libm's fmaf (single rounding as well) for CPUs without it.
#define MVS_NORM_ROW_BODY(FMA)                                                                   \
    const int a2 = (d / 8) * 8, a1 = (d / 4) * 4;                                                \
    float res;                                                                                   \
    if (a1 == 0) {                                                                               \
        const float x0 = static_cast<float>(v[0]) * inv;                                         \
        res = x0 * x0;                                                                           \
        for (int j = 1; j < d; ++j) { const float t = static_cast<float>(v[j]) * inv; res = FMA(t, t, res); } \
        return std::sqrt(res);                                                                   \
    }                                                                                            \
    float acc[8];                                                                                \
    for (int c = 0; c < 4; ++c) { const float t = static_cast<float>(v[c]) * inv; acc[c] = t * t; } \
    if (a1 > 4) {                                                                                \
        for (int c = 4; c < 8; ++c) { const float t = static_cast<float>(v[c]) * inv; acc[c] = t * t; } \
        for (int j = 8; j < a2; j += 8)                                                          \
            for (int c = 0; c < 8; ++c) { const float t = static_cast<float>(v[j + c]) * inv; acc[c] = FMA(t, t, acc[c]); } \
        for (int c = 0; c < 4; ++c) acc[c] = acc[c] + acc[c + 4];                                \
        if (a1 > a2)                                                                             \
            for (int c = 0; c < 4; ++c) { const float t = static_cast<float>(v[a2 + c]) * inv; acc[c] = FMA(t, t, acc[c]); } \
    }                                                                                            \
    res = (acc[0] + acc[1]) + (acc[2] + acc[3]);                                                 \
    for (int j = a1; j < d; ++j) { const float t = static_cast<float>(v[j]) * inv; res = FMA(t, t, res); } \
    return std::sqrt(res);

#if MVS_X86
__attribute__((target("fma"))) float norm_row_fma(const int32_t* v, int d, float inv) { MVS_NORM_ROW_BODY(__builtin_fmaf) }
__attribute__((target("fma"))) void norms_fma(const int32_t* vec, int64_t n, int d, float inv, float* out)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) out[i] = norm_row_fma(vec + static_cast<size_t>(i) * d, d, inv);
}
#endif

float norm_row_libm(const int32_t* v, int d, float inv) { MVS_NORM_ROW_BODY(std::fmaf) }
void norms_libm(const int32_t* vec, int64_t n, int d, float inv, float* out)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) out[i] = norm_row_libm(vec + static_cast<size_t>(i) * d, d, inv);
}

}  // namespace

extern "C" {

float mvs_reference_inv_sqrt_d(int d)
{
    const float x = static_cast<float>(d);
#if MVS_X86
    if (__builtin_cpu_supports("fma")) return inv_sqrt_fma(x);
    // a reference built for a CPU without FMA: same refinement, separate multiply and add
    const float r = _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(x)));
    const float t = (x * r) * r + -3.0f;
    return t * (r * -0.5f);
#else
    return 1.0f / std::sqrt(x);                  // no x86 reference build exists on this host
#endif
}

int mvs_reference_norms(const int32_t* vec, int64_t n_samples, int d, float* norms)
{
    if (n_samples < 0 || d <= 0 || ((!vec || !norms) && n_samples > 0)) return MVS_ERR_INVALID_ARG;
    const float inv = mvs_reference_inv_sqrt_d(d);
#if MVS_X86
    if (__builtin_cpu_supports("fma")) {
        norms_fma(vec, n_samples, d, inv, norms);
        return MVS_OK;
    }
#endif
    norms_libm(vec, n_samples, d, inv, norms);
    return MVS_OK;
}

}  // extern "C"
