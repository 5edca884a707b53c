// umma_i8.cu -- experiment: the limb GEMM of the pairwise path on the 5th-generation tensor cores
// (tcgen05.mma kind::i8, accumulators in TMEM, operands staged by TMA with the 128-byte swizzle).  NOT product
// code until it is parity-green; the product path is the mma.sync kernel in csrc/mvs_pairwise.cuh.
//
//   C[i][j] = sum over limb pairs (p, q) of 256^(p+q) * sum_k A_p[i][k] * B_q[j][k]      (mod 2^32)
//
// One CTA = one 128 x 128 tile of C.  Warp 0 (one lane) = TMA producer, warp 1 (one lane) = MMA issuer,
// warps 2..5 = epilogue.  Every weight s = p + q owns its own 128-column accumulator in TMEM (<= 4 weights =
// 512 columns), so the limb pairs need no ordering and the Horner combine happens once, in the epilogue.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o umma_i8 umma_i8.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at line %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

constexpr int BM = 128, BN = 128, BKB = 128, STAGES = 5, UMMA_K = 32;
constexpr int STAGE_BYTES = (BM + BN) * BKB;     // 32768
constexpr int TMEM_COLS = 512;

struct Plan { int npairs; unsigned char p[10], q[10]; int smax; };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok = 0;
    unsigned long long spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && ++spins > (1ull << 26)) { printf("mbarrier timeout block (%d,%d) thread %d bar %u parity %u\n", blockIdx.x, blockIdx.y, threadIdx.x, bar, parity); __trap(); }
    } while (!ok);
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr)
{
    return static_cast<uint64_t>((addr & 0x3FFFFu) >> 4) | (static_cast<uint64_t>(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
                   "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
                   "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(192, 1)
umma_limb_gemm(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Plan plan, int k_pad, int a_row0,
               long long mi, long long mj, int32_t* __restrict__ C, long long ldc)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;          // SW128 tiles need 1024-byte alignment
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t bars = base + STAGES * STAGE_BYTES;                    // full[STAGES], empty[STAGES], tmem_full
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen + STAGES * STAGE_BYTES + (2 * STAGES + 1) * 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(bars + s * 8, 1); mbar_init(bars + (STAGES + s) * 8, 1); }
        mbar_init(bars + 2 * STAGES * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    const int nkb = k_pad / BKB;
    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int pr = 0; pr < plan.npairs; ++pr)
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(bars + (STAGES + stage) * 8, phase ^ 1);
                    mbar_expect_tx(bars + stage * 8, STAGE_BYTES);
                    tma_load_3d(base + stage * STAGE_BYTES, &tmA, bars + stage * 8, kb * BKB, a_row0 + blockIdx.y * BM, plan.p[pr]);
                    tma_load_3d(base + stage * STAGE_BYTES + BM * BKB, &tmB, bars + stage * 8, kb * BKB, blockIdx.x * BN, plan.q[pr]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D = S32, A = B = signed 8 bit, both K-major, N = 128, M = 128
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(BN >> 3) << 17) | (static_cast<uint32_t>(BM >> 4) << 24);
            int stage = 0;
            uint32_t phase = 0, started = 0;
            for (int pr = 0; pr < plan.npairs; ++pr) {
                const int s = plan.p[pr] + plan.q[pr];
                const uint32_t tacc = tmem_base + static_cast<uint32_t>(s * BN);
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(bars + stage * 8, phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a0 = base + stage * STAGE_BYTES, b0 = a0 + BM * BKB;
#pragma unroll
                    for (int k = 0; k < BKB / UMMA_K; ++k) {
                        umma_i8(tacc, smem_desc_sw128(a0 + k * UMMA_K), smem_desc_sw128(b0 + k * UMMA_K), idesc, (started >> s) & 1u);
                        started |= 1u << s;
                    }
                    umma_commit(bars + (STAGES + stage) * 8);             // frees the smem slot when these MMAs are done
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
            umma_commit(bars + 2 * STAGES * 8);                           // all accumulators complete
        }
    } else {
        mbar_wait(bars + 2 * STAGES * 8, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int q = warp & 3;                                           // TMEM lane quarter this warp may read
        const long long row = static_cast<long long>(blockIdx.y) * BM + q * 32 + lane;
        for (int c0 = 0; c0 < BN; c0 += 32) {
            uint32_t v[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = 0;
            for (int s = plan.smax; s >= 0; --s) {                        // Horner over the weights
                uint32_t r[32];
                tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(s * BN + c0), r);
#pragma unroll
                for (int e = 0; e < 32; ++e) v[e] = (v[e] << 8) + r[e];
            }
            const long long col = static_cast<long long>(blockIdx.x) * BN + c0;
            if (row < mi) {
                int32_t* dst = C + row * ldc + col;
                if (col + 31 < mj && (ldc & 3) == 0) {
#pragma unroll
                    for (int e = 0; e < 32; e += 4) *reinterpret_cast<int4*>(dst + e) = make_int4(v[e], v[e + 1], v[e + 2], v[e + 3]);
                } else {
                    for (int e = 0; e < 32; ++e)
                        if (col + e < mj) dst[e] = static_cast<int32_t>(v[e]);
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

// MMA-only issue-rate probe: no TMA, the operands are whatever is in shared memory; one CTA per SM issues
// iters x 4 MMAs of 128 x N x 32 (N = 64 / 128 / 256) into TMEM and waits for the commit.
template <int N>
__global__ void __launch_bounds__(128, 1) umma_rate_kernel(int iters)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(BM >> 4) << 24);
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_i8(tmem_base + (it & 1) * N, smem_desc_sw128(base + k * UMMA_K), smem_desc_sw128(base + 32768 + k * UMMA_K), idesc, 1u);
        umma_commit(smem_u32(&bar));
        mbar_wait(smem_u32(&bar), 0);
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

template <int N>
static void run_rate()
{
    const size_t smem = 32768 + 256 * 128 + 2048;
    CK(cudaFuncSetAttribute(umma_rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int iters = 20000;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        umma_rate_kernel<N><<<148, 128, smem>>>(iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep && ms < best) best = ms;
    }
    const double macs = 148.0 * iters * 4 * 128.0 * N * 32;
    printf("tcgen05.mma kind::i8 128 x %3d x 32, issue only: %8.3f ms  %7.1f T MAC/s  (%.1f clk per MMA)\n", N, best, macs / (best * 1e-3) / 1e12,
           best * 1e-3 * 1.965e9 / (iters * 4.0));
}

// ---------------------------------------------------------------- host ----------------------------------------------------------------
typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(EncodeTiled enc, const int8_t* planes, int P, long long n_pad, int k_pad)
{
    CUtensorMap tm;
    const cuuint64_t dims[3] = {static_cast<cuuint64_t>(k_pad), static_cast<cuuint64_t>(n_pad), static_cast<cuuint64_t>(P)};
    const cuuint64_t strides[2] = {static_cast<cuuint64_t>(k_pad), static_cast<cuuint64_t>(k_pad) * static_cast<cuuint64_t>(n_pad)};
    const cuuint32_t box[3] = {BKB, BM, 1}, estr[3] = {1, 1, 1};
    const CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<int8_t*>(planes), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d\n", static_cast<int>(r)); exit(1); }
    return tm;
}

static Plan make_plan(int PA, int PB)
{
    Plan pl{};
    pl.smax = PA + PB - 2 < 3 ? PA + PB - 2 : 3;
    for (int s = pl.smax; s >= 0; --s)
        for (int p = 0; p < PA; ++p) {
            const int q = s - p;
            if (q < 0 || q >= PB) continue;
            pl.p[pl.npairs] = static_cast<unsigned char>(p);
            pl.q[pl.npairs] = static_cast<unsigned char>(q);
            ++pl.npairs;
        }
    return pl;
}

int main(int argc, char** argv)
{
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    EncodeTiled enc = reinterpret_cast<EncodeTiled>(fn);
    const size_t smem = STAGES * STAGE_BYTES + 1024 + 256;
    CK(cudaFuncSetAttribute(umma_limb_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));

    // ---- correctness: every limb count, ragged mi / mj, row offset ----
    int bad = 0;
    for (int P = 1; P <= 4; ++P) {
        const long long mi = 200, mj = 300, n_pad_a = 512, n_pad_b = 512;
        const int k_pad = 256, a_row0 = (P == 3) ? 128 : 0;
        std::vector<int8_t> hA(static_cast<size_t>(P) * n_pad_a * k_pad, 0), hB(static_cast<size_t>(P) * n_pad_b * k_pad, 0);
        uint32_t rng = 12345u + P;
        auto next = [&] { rng = rng * 1664525u + 1013904223u; return static_cast<int8_t>(rng >> 24); };
        for (int p = 0; p < P; ++p) {
            for (long long r = 0; r < a_row0 + mi; ++r) for (int k = 0; k < k_pad; ++k) hA[(p * n_pad_a + r) * k_pad + k] = next();
            for (long long r = 0; r < mj; ++r) for (int k = 0; k < k_pad; ++k) hB[(p * n_pad_b + r) * k_pad + k] = next();
        }
        int8_t *dA, *dB;
        int32_t* dC;
        const long long ldc = 304;
        CK(cudaMalloc(&dA, hA.size()));
        CK(cudaMalloc(&dB, hB.size()));
        CK(cudaMalloc(&dC, mi * ldc * 4));
        CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
        CK(cudaMemset(dC, 0x55, mi * ldc * 4));
        const CUtensorMap tmA = make_map(enc, dA, P, n_pad_a, k_pad), tmB = make_map(enc, dB, P, n_pad_b, k_pad);
        const Plan pl = make_plan(P, P);
        umma_limb_gemm<<<dim3((mj + BN - 1) / BN, (mi + BM - 1) / BM), 192, smem>>>(tmA, tmB, pl, k_pad, a_row0, mi, mj, dC, ldc);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        std::vector<int32_t> hC(mi * ldc);
        CK(cudaMemcpy(hC.data(), dC, hC.size() * 4, cudaMemcpyDeviceToHost));
        long long wrong = 0;
        for (long long i = 0; i < mi; ++i)
            for (long long j = 0; j < mj; ++j) {
                uint32_t want = 0;
                for (int pr = 0; pr < pl.npairs; ++pr) {
                    int32_t acc = 0;
                    const int8_t* a = &hA[(pl.p[pr] * n_pad_a + a_row0 + i) * k_pad];
                    const int8_t* b = &hB[(pl.q[pr] * n_pad_b + j) * k_pad];
                    for (int k = 0; k < k_pad; ++k) acc += static_cast<int32_t>(a[k]) * b[k];
                    want += static_cast<uint32_t>(acc) << (8 * (pl.p[pr] + pl.q[pr]));
                }
                if (static_cast<uint32_t>(hC[i * ldc + j]) != want) { if (wrong < 3) printf("  P=%d mismatch (%lld,%lld): got %d want %d\n", P, i, j, hC[i * ldc + j], static_cast<int32_t>(want)); ++wrong; }
            }
        printf("P=%d pairs=%d: %lld wrong of %lld\n", P, pl.npairs, wrong, mi * mj);
        bad += wrong != 0;
        cudaFree(dA); cudaFree(dB); cudaFree(dC);
    }
    if (bad || (argc > 1 && !strcmp(argv[1], "check"))) return bad;

    run_rate<64>();
    run_rate<128>();
    run_rate<256>();

    // ---- timing: one 8192-row block of a 10000-vector db, d = 2048 ----
    for (int P = 1; P <= 4; P += (P == 2 ? 2 : 1)) {
        const long long n = 10000, n_pad = 10112 + 128, mi = 8192;
        const int k_pad = 2048;
        int8_t* dV;
        int32_t* dC;
        const long long ldc = 10000;
        CK(cudaMalloc(&dV, static_cast<size_t>(P) * n_pad * k_pad));
        CK(cudaMemset(dV, 3, static_cast<size_t>(P) * n_pad * k_pad));
        CK(cudaMalloc(&dC, mi * ldc * 4));
        const CUtensorMap tm = make_map(enc, dV, P, n_pad, k_pad);
        const Plan pl = make_plan(P, P);
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep) {
            CK(cudaEventRecord(e0));
            umma_limb_gemm<<<dim3((n + BN - 1) / BN, (mi + BM - 1) / BM), 192, smem>>>(tm, tm, pl, k_pad, 0, mi, n, dC, ldc);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaGetLastError());
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep && ms < best) best = ms;
        }
        const double macs = static_cast<double>(mi) * n * k_pad * pl.npairs;
        printf("P=%d (%d limb products): %8.3f ms  %7.1f T int8 MAC/s  = %6.1f T int32-MAC-equivalents/s\n", P, pl.npairs, best, macs / (best * 1e-3) / 1e12,
               static_cast<double>(mi) * n * k_pad / (best * 1e-3) / 1e12);
        cudaFree(dV); cudaFree(dC);
    }
    return 0;
}
