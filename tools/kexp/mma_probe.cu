// mma_probe.cu -- issue-rate probe of the legacy warp-level tensor instructions on sm_100a (NOT product code):
//   mma.sync.m16n8k32.s32.s8.s8.s32, mma.sync.m16n8k32.s32.u8.u8.s32, mma.sync.m16n8k16.f32.bf16.bf16.f32,
//   mma.sync.m16n8k16.f32.f16.f16.f32
// 8 independent accumulators per warp, 8 warps per CTA, 2 CTAs per SM.  Prints T MAC/s.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_probe mma_probe.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) probe(int* sink, int iters, uint32_t seed)
{
    uint32_t a[4] = {seed + threadIdx.x, seed * 3 + threadIdx.x, seed * 5, seed * 7 + blockIdx.x};
    uint32_t b[2] = {seed * 11 + threadIdx.x, seed * 13};
    int c[8][4];
    float f[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { c[i][j] = 0; f[i][j] = 0.f; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
            else if (MODE == 1)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
            else if (MODE == 2)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(f[i][0]), "+f"(f[i][1]), "+f"(f[i][2]), "+f"(f[i][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
            else
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(f[i][0]), "+f"(f[i][1]), "+f"(f[i][2]), "+f"(f[i][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
        }
    }
    int x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) x += c[i][j] + static_cast<int>(f[i][j]);
    if (x == 0x7fffffff) sink[0] = x;
}

template <int MODE>
void run(const char* name, int kdim)
{
    int* sink;
    cudaMalloc(&sink, 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 20000;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        probe<MODE><<<148 * 2, 256>>>(sink, iters, 12345u);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double macs = 148.0 * 2 * 8 * iters * 8.0 * 16 * 8 * kdim;
    printf("%-46s %8.3f ms  %8.1f TMAC/s  (%s)\n", name, best, macs / (best * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    run<0>("mma.sync m16n8k32 s8.s8 -> s32", 32);
    run<1>("mma.sync m16n8k32 u8.u8 -> s32", 32);
    run<2>("mma.sync m16n8k16 bf16 -> f32", 16);
    run<3>("mma.sync m16n8k16 f16 -> f32", 16);
    return 0;
}
