/*
 * mvs_oracle.c -- CPU ORACLE for the random-projection sketching path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, bench.py's
 * cpu_baseline / --impl reference legs and __graft_entry__.smoke() may load
 * it.  The shipped path (metagenome_vector_sketches_b200/csrc) never links
 * or calls anything in this directory.
 *
 * It restates, loop for loop, the reference algorithm
 *   transform_set_into_vector   /root/reference/src/random_projection.cpp:9-26
 * and the two host-side post-passes of the batch driver
 *   int16 saturation            /root/reference/src/project_everything.cpp:332-347
 *   norm = ||v / sqrt(d)||      /root/reference/src/project_everything.cpp:327-330
 *
 * Parity pin: oracle/_ref (the unmodified reference sources compiled where
 * they lie, see oracle/Makefile) is compared with this file on the golden
 * vectors in tests/golden (KAT-1..3 of SURVEY.md section 8c) and on seeded
 * random sets by tests/test_oracle.py.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <math.h>

/* random_projection.cpp:13-17 -- the splitmix64 finaliser applied to
 * (hash + i), i = first column of the 64-wide block.  `seed` is the additive
 * extension of SURVEY.md D2: seed 0 is the reference. */
static inline uint64_t oracle_mix(uint64_t hash, uint64_t i, uint64_t seed)
{
    uint64_t x = hash + i + seed;
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    x = x ^ (x >> 31);
    return x;
}

uint64_t oracle_splitmix64(uint64_t hash, uint64_t i, uint64_t seed)
{
    return oracle_mix(hash, i, seed);
}

/* random_projection.cpp:9-26.  `hashes` must already be a set (unique values),
 * as the reference receives an unordered_set.  out[d] is overwritten. */
void oracle_transform_set(const uint64_t *hashes, int64_t n, int d,
                          uint64_t seed, int32_t *out)
{
    if (d <= 0) return;
    memset(out, 0, (size_t)d * sizeof(int32_t));
    for (int64_t k = 0; k < n; ++k) {                       /* :11 */
        for (int i = 0; i < d; i += 64) {                   /* :12 */
            uint64_t x = oracle_mix(hashes[k], (uint64_t)i, seed);
            for (int b = 0; b < 64 && (i + b) < d; ++b) {   /* :19 */
                int projected = 1 - 2 * (int)((x >> b) & 1);/* :20 */
                out[i + b] += projected;                    /* :21 */
            }
        }
    }
}

/* The sample loop of sketch(), project_everything.cpp:289-298, over a CSR
 * batch: sample s owns hashes[offsets[s] .. offsets[s+1]).  OpenMP dynamic
 * schedule as in the reference when compiled with -fopenmp. */
void oracle_project_csr(const uint64_t *hashes, const uint64_t *offsets,
                        int64_t n_samples, int d, uint64_t seed, int32_t *out)
{
#pragma omp parallel for schedule(dynamic)
    for (int64_t s = 0; s < n_samples; ++s)
        oracle_transform_set(hashes + offsets[s],
                             (int64_t)(offsets[s + 1] - offsets[s]), d, seed,
                             out + (size_t)s * (size_t)d);
}

/* project_everything.cpp:332-347 */
void oracle_saturate_int16(const int32_t *v, int64_t n, int16_t *out)
{
    for (int64_t i = 0; i < n; ++i) {
        int32_t x = v[i];
        out[i] = (int16_t)(x > 32767 ? 32767 : (x < -32768 ? -32768 : x));
    }
}

/* Exact value the reference's float expression approximates:
 * sqrt(sum v^2 / d).  (project_everything.cpp:327-329 evaluates it in float32
 * through Eigen; the printed text has 6 significant digits.) */
double oracle_norm_exact(const int32_t *v, int d)
{
    uint64_t ss = 0;
    for (int i = 0; i < d; ++i) ss += (uint64_t)((int64_t)v[i] * (int64_t)v[i]);
    return sqrt((double)ss / (double)d);
}

/* The value the reference actually prints (project_everything.cpp:328-330), restated operation by
 * operation from its compiled code (g++ 13, -O3 -ffast-math, vendored Eigen 3.2.10 = SSE packets of 4;
 * the -march=native and x86-64-v3 builds agree):
 *   inv = 1/sqrt(float(d)) as  r = rsqrtss(d);  inv = ((d*r)*r - 3 [one fma]) * (r * -0.5)
 *   x_j = float(v_j) * inv
 *   Eigen redux with two packet accumulators: acc0 = x[0..3]^2, acc1 = x[4..7]^2 (products), then
 *   fma per following packet pair, acc0 += acc1, fma for an odd last packet, hadd (a0+a1)+(a2+a3),
 *   scalar fma tail, sqrtss; d < 4: x0^2 then scalar fma.
 * `inv` is an argument: rsqrtss is an approximation defined by the executing CPU, so the caller asks
 * the CPU for it (oracle_inv_sqrt_d below). Compile with -ffp-contract=off: every rounding is explicit. */
float oracle_norm_reference(const int32_t *v, int d, float inv)
{
    const int ps = 4;
    const int aligned2 = (d / (2 * ps)) * (2 * ps), aligned = (d / ps) * ps;
    float res;
    if (aligned == 0) {
        float x = (float)v[0] * inv;
        res = x * x;
        for (int j = 1; j < d; ++j) { x = (float)v[j] * inv; res = fmaf(x, x, res); }
        return sqrtf(res);
    }
    float p0[4], p1[4];
    for (int l = 0; l < ps; ++l) { float x = (float)v[l] * inv; p0[l] = x * x; }
    if (aligned > ps) {
        for (int l = 0; l < ps; ++l) { float x = (float)v[ps + l] * inv; p1[l] = x * x; }
        for (int idx = 2 * ps; idx < aligned2; idx += 2 * ps)
            for (int l = 0; l < ps; ++l) {
                float x = (float)v[idx + l] * inv, y = (float)v[idx + ps + l] * inv;
                p0[l] = fmaf(x, x, p0[l]);
                p1[l] = fmaf(y, y, p1[l]);
            }
        for (int l = 0; l < ps; ++l) p0[l] = p0[l] + p1[l];
        if (aligned > aligned2)
            for (int l = 0; l < ps; ++l) { float x = (float)v[aligned2 + l] * inv; p0[l] = fmaf(x, x, p0[l]); }
    }
    res = (p0[0] + p0[1]) + (p0[2] + p0[3]);
    for (int j = aligned; j < d; ++j) { float x = (float)v[j] * inv; res = fmaf(x, x, res); }
    return sqrtf(res);
}

#if defined(__x86_64__)
#include <xmmintrin.h>
float oracle_inv_sqrt_d(int d)
{
    const float x = (float)d;
    const float r = _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(x)));
    return fmaf(x * r, r, -3.0f) * (r * -0.5f);
}
#else
float oracle_inv_sqrt_d(int d) { return 1.0f / sqrtf((float)d); }
#endif

uint64_t oracle_sumsq(const int32_t *v, int d)
{
    uint64_t ss = 0;
    for (int i = 0; i < d; ++i) ss += (uint64_t)((int64_t)v[i] * (int64_t)v[i]);
    return ss;
}

/* ------------------------------------------------------------------------------------------
 * Pairwise comparison of projected vectors:
 *   compute_sparse_dot_products_optimized   /root/reference/src/pairwise_comp_optimized.cpp:57-160
 * The live code is :135-147: an Eigen MatrixXi product (int32 multiply-add, i.e. arithmetic
 * modulo 2^32 reinterpreted as int32), then for every (i, j) in row-major order
 *   threshold = 0.05 * (norms_i(i) + norms_j(j))          (norms are squared norms, double)
 *   keep (i, j, dot) iff  (int64)dot / dimension > threshold   (integer division, truncating)
 * Blocks are row-major [m x d] (the memory layout of the reference's column-major d x m block).
 * Returns the number of kept pairs; writes at most cap of them.
 * ------------------------------------------------------------------------------------------ */
int64_t oracle_pairwise_block(const int32_t *block_i, int64_t mi, const int32_t *block_j, int64_t mj, int d,
                              const double *norms_i, const double *norms_j,
                              int32_t *rows, int32_t *cols, int64_t *values, int64_t cap)
{
    int64_t n = 0;
    for (int64_t i = 0; i < mi; ++i) {
        for (int64_t j = 0; j < mj; ++j) {
            uint32_t acc = 0;                                  /* int32 wraparound, well defined */
            const int32_t *a = block_i + i * (int64_t)d, *b = block_j + j * (int64_t)d;
            for (int k = 0; k < d; ++k) acc += (uint32_t)a[k] * (uint32_t)b[k];
            const int64_t dot_product = (int64_t)(int32_t)acc;
            const double threshold = 0.05 * (norms_i[i] + norms_j[j]);
            if (dot_product / d > threshold) {
                if (n < cap) { rows[n] = (int32_t)i; cols[n] = (int32_t)j; values[n] = dot_product; }
                ++n;
            }
        }
    }
    return n;
}

/* Jaccard estimate and its 8-bit quantisation as stored in shard_k/matrix.bin
 *   /root/reference/src/pairwise_comp_optimized.cpp:654-665 */
uint16_t oracle_quantize_jaccard(int64_t value, int d, double norm_row, double norm_col)
{
    const double MULT_CONST = 255.0;
    double inter = (double)value / d;
    double jaccard = inter / (norm_row + norm_col - inter);
    if (jaccard > 1) jaccard = 1;
    return (uint16_t)round(jaccard * MULT_CONST);
}
