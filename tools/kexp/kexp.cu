// kexp.cu -- experiment bench for the inner loop of project_kernel (NOT product code).
//
// Builds several variants of the mixer / carry-save loop for the d = 2048 geometry
// (one hash broadcast to 32 lanes, lane = 64-column block) and times them on the same
// device-generated input; every variant must produce the same checksum.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o kexp kexp.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

constexpr uint64_t kGolden = 0x9e3779b97f4a7c15ULL;
#ifndef KPLANES
#define KPLANES 14
#endif
constexpr int kPlanes = KPLANES;          // product: 14 (csrc/mvs_kernels.cuh); the round-1 tables were taken with 12

struct Consts { uint32_t m4, m32, m2; };   // 2^(32-30), 2^(32-27), 2^(32-31): opaque to ptxas

__device__ __forceinline__ void csa(uint32_t& carry, uint32_t& sum, uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t s, m;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(s) : "r"(a), "r"(b), "r"(c));
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(m) : "r"(a), "r"(b), "r"(c));
    sum = s; carry = m;
}

template <int L>
__device__ __forceinline__ void ripple(uint32_t (&P)[kPlanes], uint32_t c)
{
#pragma unroll
    for (int p = L; p < kPlanes; ++p) { uint32_t t = P[p] & c; P[p] ^= c; c = t; }
}

// ---- mixer variants ------------------------------------------------------------------
// V bit 0..1: number of xorshifts (0..3) whose high-word shift runs as IMAD.HI on the FMA pipe
// V bit 2   : multiply as  mad.wide(ylo, Mlo, {0, ylo*Mhi + yhi*Mlo})  (4 FMA slots instead of 5)
// V bit 3   : 64-bit add as mad.wide(1, clo, h) + add hi (all FMA pipe)
template <int V>
__device__ __forceinline__ uint32_t shr_hi(uint32_t hi, int k, uint32_t mulc, int idx)
{
    if (idx < (V & 3)) {
        uint32_t r;
        asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(hi), "r"(mulc));
        return r;
    }
    return hi >> k;
}

template <int V>
__device__ __forceinline__ void mul64(uint32_t ylo, uint32_t yhi, uint32_t mlo, uint32_t mhi, uint32_t& lo, uint32_t& hi)
{
    if constexpr (V & 4) {
        uint64_t r;
        asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(ylo), "r"(mlo));
        lo = static_cast<uint32_t>(r);
        hi = static_cast<uint32_t>(r >> 32);
        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(ylo), "r"(mhi));
        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(yhi), "r"(mlo));
    } else {
        const uint64_t y = (static_cast<uint64_t>(yhi) << 32) | ylo;
        const uint64_t r = y * ((static_cast<uint64_t>(mhi) << 32) | mlo);
        lo = static_cast<uint32_t>(r);
        hi = static_cast<uint32_t>(r >> 32);
    }
}

template <int V>
__device__ __forceinline__ void mix64(uint64_t h, uint64_t cval, const Consts& kc, uint32_t& olo, uint32_t& ohi, uint32_t& eacc)
{
    uint32_t lo, hi;
    if constexpr (V & 16) {
        // h already holds hash + base constant; cval's low word is 64*block (no carry possible: host/tile-checked)
        lo = static_cast<uint32_t>(h) + static_cast<uint32_t>(cval);
        hi = static_cast<uint32_t>(h >> 32);
    } else if constexpr (V & 8) {
        uint64_t x;
        asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(x) : "r"(static_cast<uint32_t>(cval)), "r"(kc.m2 >> 1), "l"(h));
        lo = static_cast<uint32_t>(x);
        hi = static_cast<uint32_t>(x >> 32) + static_cast<uint32_t>(cval >> 32);
    } else {
        const uint64_t x = h + cval;
        lo = static_cast<uint32_t>(x);
        hi = static_cast<uint32_t>(x >> 32);
    }
    uint32_t ylo = lo ^ __funnelshift_r(lo, hi, 30);
    uint32_t yhi = hi ^ shr_hi<V>(hi, 30, kc.m4, 0);
    mul64<V>(ylo, yhi, 0x1ce4e5b9u, 0xbf58476du, lo, hi);
    ylo = lo ^ __funnelshift_r(lo, hi, 27);
    yhi = hi ^ shr_hi<V>(hi, 27, kc.m32, 1);
    mul64<V>(ylo, yhi, 0x133111ebu, 0x94d049bbu, lo, hi);
    olo = lo ^ __funnelshift_r(lo, hi, 31);
    if constexpr (V & 64) {
        // column 32 of the block is bit0 ^ bit31 of hi: t = hi * 0x80000001 has that xor in bit 31 and
        // hi's bits 0..30 below it, so popc(t) = popc(hi & 0x7fffffff) + (bit0 ^ bit31); the first term
        // summed over the hashes is known from the plane counts at the flush.  IMAD + POPC + IMAD.IADD:
        // no ALU-pipe instruction, instead of SHF + LOP3.
        uint32_t t;
        asm("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(hi), "r"(kc.m2 * 0x40000000u + 1u));
        eacc += __popc(t);
        ohi = hi;
    } else {
        ohi = hi ^ shr_hi<V>(hi, 31, kc.m2, 2);
    }
}

__device__ __forceinline__ uint32_t hs4(uint32_t (&P)[kPlanes], uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    uint32_t tA, tB, f;
    csa(tA, P[0], P[0], a, b);
    csa(tB, P[0], P[0], c, d);
    csa(f, P[1], P[1], tA, tB);
    return f;
}

template <int V>
__device__ __forceinline__ void group16(const uint64_t* __restrict__ sp, uint64_t cval, const Consts& kc,
                                        uint32_t (&Plo)[kPlanes], uint32_t (&Phi)[kPlanes],
                                        uint32_t& c16lo, uint32_t& c16hi, uint32_t& eacc)
{
    uint32_t lo[16], hi[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const ulonglong2 h2 = *reinterpret_cast<const ulonglong2*>(sp + 2 * i);
        mix64<V>(h2.x, cval, kc, lo[2 * i], hi[2 * i], eacc);
        mix64<V>(h2.y, cval, kc, lo[2 * i + 1], hi[2 * i + 1], eacc);
    }
    uint32_t fl[4], fh[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        fl[q] = hs4(Plo, lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
        fh[q] = hs4(Phi, hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
    }
    uint32_t eA, eB;
    csa(eA, Plo[2], Plo[2], fl[0], fl[1]);
    csa(eB, Plo[2], Plo[2], fl[2], fl[3]);
    csa(c16lo, Plo[3], Plo[3], eA, eB);
    csa(eA, Phi[2], Phi[2], fh[0], fh[1]);
    csa(eB, Phi[2], Phi[2], fh[2], fh[3]);
    csa(c16hi, Phi[3], Phi[3], eA, eB);
}

// ---- V & 128: "cooked" tiles -----------------------------------------------------------
// The tile pass stores per hash 16 bytes {lo32(hb), hi32(hb), 0, M}, M = (xh ^ (xh >> 30)) * K1lo with
// xh = hi32(hb): when lo32(hb) + 64*block cannot carry, the high word of x ^ (x >> 30) and its product
// with the low word of the first multiplier are the same for all 32 lanes, so the lane skips one
// SHF + one LOP3 (ALU pipe) and one IMAD, and {0, M} is the 64-bit addend of the IMAD.WIDE.
// V & 8 here: lane add as IMAD (FMA pipe).  V & 4: second multiply as wide + 2 mad.lo.
// V & 512: record {lo32(hb), F, 0, M}: the second half is a ready-made 64-bit addend (M << 32) for the IMAD.WIDE, which
//          then yields (lo, hi + M) at once; hi += l1 * K1hi follows as one IMAD.  One add less per (hash, lane).
//          The exact path takes hb from the raw tile (rawv = raw hash + golden).
template <int V, bool EXACT>
__device__ __forceinline__ void mix_cooked(ulonglong2 c, uint64_t rawv, uint32_t off, const Consts& kc, uint32_t& olo, uint32_t& ohi,
                                           uint32_t& eacc)
{
    uint32_t lo, hi;
    if constexpr (EXACT) {
        const uint64_t x = ((V & 512) ? rawv : c.x) + off;
        const uint64_t y = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
        lo = static_cast<uint32_t>(y);
        hi = static_cast<uint32_t>(y >> 32);
    } else if constexpr (V & 512) {
        const uint32_t xl = static_cast<uint32_t>(c.x) + off;
        const uint32_t l1 = xl ^ static_cast<uint32_t>(c.x >> 32);
        uint64_t y;
        asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(y) : "r"(l1), "r"(0x1ce4e5b9u), "l"(c.y));
        lo = static_cast<uint32_t>(y);
        hi = static_cast<uint32_t>(y >> 32);
        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(l1), "r"(0xbf58476du));
    } else {
        uint32_t xl;
        if constexpr (V & 8) asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(xl) : "r"(off), "r"(kc.m2 >> 1), "r"(static_cast<uint32_t>(c.x)));
        else xl = static_cast<uint32_t>(c.x) + off;
        // V & 256: the funnel word (xl >> 30) | (xh << 2) is per hash too unless + off reaches bit 30
        const uint32_t l1 = (V & 256) ? (xl ^ static_cast<uint32_t>(c.y))
                                      : (xl ^ __funnelshift_r(xl, static_cast<uint32_t>(c.x >> 32), 30));
        uint32_t t = static_cast<uint32_t>(c.y >> 32);                  // M
        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(t) : "r"(l1), "r"(0xbf58476du));
        uint64_t y;
        asm("mul.wide.u32 %0, %1, %2;" : "=l"(y) : "r"(l1), "r"(0x1ce4e5b9u));
        lo = static_cast<uint32_t>(y);
        hi = static_cast<uint32_t>(y >> 32) + t;
    }
    const uint32_t ylo = lo ^ __funnelshift_r(lo, hi, 27);
    const uint32_t yhi = hi ^ (hi >> 27);
    mul64<V>(ylo, yhi, 0x133111ebu, 0x94d049bbu, lo, hi);
    olo = lo ^ __funnelshift_r(lo, hi, 31);
    if constexpr (V & 64) {
        uint32_t t;
        asm("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(hi), "r"(kc.m2 * 0x40000000u + 1u));
        eacc += __popc(t);
        ohi = hi;
    } else {
        ohi = hi ^ (hi >> 31);                   // the plain last xor-shift: SHF + LOP3 on the ALU pipe
    }
}

template <int V, bool EXACT>
__device__ __forceinline__ void group16c(const ulonglong2* __restrict__ sp, const uint64_t* __restrict__ raw, uint32_t off, const Consts& kc,
                                         uint32_t (&Plo)[kPlanes], uint32_t (&Phi)[kPlanes],
                                         uint32_t& c16lo, uint32_t& c16hi, uint32_t& eacc)
{
    uint32_t fl[4], fh[4];
    if constexpr (V & 1024) {                    // source order: mix 4, compress 4, mix 4, ... (both pipes fed in every window)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t lo[4], hi[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                mix_cooked<V, EXACT>(sp[4 * q + i], (EXACT && (V & 512)) ? raw[4 * q + i] + kGolden : 0ull, off, kc, lo[i], hi[i], eacc);
            fl[q] = hs4(Plo, lo[0], lo[1], lo[2], lo[3]);
            fh[q] = hs4(Phi, hi[0], hi[1], hi[2], hi[3]);
        }
    } else {
        uint32_t lo[16], hi[16];
#pragma unroll
        for (int i = 0; i < 16; ++i)
            mix_cooked<V, EXACT>(sp[i], (EXACT && (V & 512)) ? raw[i] + kGolden : 0ull, off, kc, lo[i], hi[i], eacc);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            fl[q] = hs4(Plo, lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
            fh[q] = hs4(Phi, hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
        }
    }
    uint32_t eA, eB;
    csa(eA, Plo[2], Plo[2], fl[0], fl[1]);
    csa(eB, Plo[2], Plo[2], fl[2], fl[3]);
    csa(c16lo, Plo[3], Plo[3], eA, eB);
    csa(eA, Phi[2], Phi[2], fh[0], fh[1]);
    csa(eB, Phi[2], Phi[2], fh[2], fh[3]);
    csa(c16hi, Phi[3], Phi[3], eA, eB);
}

// V & 4096: 64 hashes per loop trip: four groups with the carries merged directly (no pending-carry state, no branches)
template <int V, bool EXACT>
__device__ __forceinline__ void group64c(const ulonglong2* __restrict__ sp, const uint64_t* __restrict__ raw, uint32_t off, const Consts& kc,
                                         uint32_t (&Plo)[kPlanes], uint32_t (&Phi)[kPlanes], uint32_t& eacc)
{
    uint32_t a_lo, a_hi, b_lo, b_hi, c32lo[2], c32hi[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        group16c<V, EXACT>(sp + 32 * h, raw + 32 * h, off, kc, Plo, Phi, a_lo, a_hi, eacc);
        group16c<V, EXACT>(sp + 32 * h + 16, raw + 32 * h + 16, off, kc, Plo, Phi, b_lo, b_hi, eacc);
        csa(c32lo[h], Plo[4], Plo[4], a_lo, b_lo);
        csa(c32hi[h], Phi[4], Phi[4], a_hi, b_hi);
    }
    uint32_t c64lo, c64hi;
    csa(c64lo, Plo[5], Plo[5], c32lo[0], c32lo[1]);
    csa(c64hi, Phi[5], Phi[5], c32hi[0], c32hi[1]);
    ripple<6>(Plo, c64lo);
    ripple<6>(Phi, c64hi);
}

__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc, uint32_t src_bytes)
{
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc)
{
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One warp = `per` hashes (multiple of 64) of one "sample" -> out row of 2048 int32.
template <int V, int WARPS, int TILE, int MINB = 1>
__global__ void __launch_bounds__(WARPS * 32, MINB)
kexp_kernel(const uint64_t* __restrict__ hashes, uint32_t ntasks, uint32_t per, Consts kc, int32_t* __restrict__ out)
{
    extern __shared__ __align__(16) uint64_t smem[];
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t wid = blockIdx.x * WARPS + warp;
    if (wid >= ntasks) return;
    uint64_t* ring = smem + warp * (2 * TILE);
    const uint64_t* src = hashes + static_cast<uint64_t>(wid) * per;
    const uint64_t cfull = kGolden + (static_cast<uint64_t>(lane) << 6);
    const uint64_t cval = (V & 16) ? (static_cast<uint64_t>(lane) << 6) : cfull;

    uint32_t Plo[kPlanes], Phi[kPlanes];
#pragma unroll
    for (int p = 0; p < kPlanes; ++p) { Plo[p] = 0u; Phi[p] = 0u; }
    uint32_t pend16lo = 0, pend16hi = 0, pend32lo = 0, pend32hi = 0, pend64lo = 0, pend64hi = 0;
    uint32_t g = 0, eacc = 0;
    const uint32_t ntiles = (per + TILE - 1) / TILE;

    auto issue_tile = [&](uint32_t t) {
        uint64_t* st = ring + (t & 1) * TILE;
        const uint32_t base = t * TILE;
#pragma unroll
        for (int i = 0; i < TILE / 32; ++i) {
            const uint32_t pos = base + i * 32 + lane;
            const bool ok = pos < per;
            cp_async8(st + i * 32 + lane, src + (ok ? pos : 0u), ok ? 8u : 0u);
        }
    };
    if (ntiles) issue_tile(0);
    cp_async_commit();
    for (uint32_t t = 0; t < ntiles; ++t) {
        if (t + 1 < ntiles) issue_tile(t + 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncwarp();
        uint64_t* st = ring + (t & 1) * TILE;
        bool tile_risky = false;
        ulonglong2* cooked = reinterpret_cast<ulonglong2*>(smem + WARPS * 2 * TILE) + warp * TILE;
        if constexpr (V & 128) {
            bool risky = false;
#pragma unroll
            for (int i = 0; i < TILE / 32; ++i) {
                const uint64_t v = st[i * 32 + lane] + kGolden;
                const uint32_t xl0 = static_cast<uint32_t>(v);
                risky |= (V & 256) ? ((xl0 | 0xC0000000u) >= 0xFFFFF800u) : (xl0 >= 0xFFFFF840u);
                const uint32_t xh = static_cast<uint32_t>(v >> 32);
                const uint32_t m = (xh ^ (xh >> 30)) * 0x1ce4e5b9u;
                const uint32_t f = (xl0 >> 30) | (xh << 2);
                if constexpr (V & 512) cooked[i * 32 + lane] = make_ulonglong2((static_cast<uint64_t>(f) << 32) | xl0, static_cast<uint64_t>(m) << 32);
                else cooked[i * 32 + lane] = make_ulonglong2(v, (static_cast<uint64_t>(m) << 32) | ((V & 256) ? f : 0u));
            }
            tile_risky = __any_sync(0xffffffffu, risky);
            __syncwarp();
        } else if constexpr (V & 16) {
            bool risky = false;
#pragma unroll
            for (int i = 0; i < TILE / 32; ++i) {
                const uint64_t v = st[i * 32 + lane] + kGolden;
                risky |= static_cast<uint32_t>(v) >= 0xFFFFF840u;
                st[i * 32 + lane] = v;
            }
            tile_risky = __any_sync(0xffffffffu, risky);
            __syncwarp();
        }
        const uint32_t valid = min(per - t * TILE, static_cast<uint32_t>(TILE));
        const uint32_t ngroups = (valid + 15) >> 4;
        if constexpr ((V & 4096) != 0) {         // per is a multiple of 64 in this bench
            for (uint32_t gi = 0; gi < ngroups; gi += 4) {
                if (tile_risky) group64c<V, true>(cooked + gi * 16, st + gi * 16, static_cast<uint32_t>(lane) << 6, kc, Plo, Phi, eacc);
                else group64c<V, false>(cooked + gi * 16, st + gi * 16, static_cast<uint32_t>(lane) << 6, kc, Plo, Phi, eacc);
            }
        } else
        for (uint32_t gi = 0; gi < ngroups; ++gi, ++g) {
            uint32_t c16lo, c16hi;
            if constexpr (V & 128) {
                if (tile_risky) group16c<V, true>(cooked + gi * 16, st + gi * 16, static_cast<uint32_t>(lane) << 6, kc, Plo, Phi, c16lo, c16hi, eacc);
                else group16c<V, false>(cooked + gi * 16, st + gi * 16, static_cast<uint32_t>(lane) << 6, kc, Plo, Phi, c16lo, c16hi, eacc);
            } else
            if ((V & 16) && tile_risky) group16<(V & ~16)>(st + gi * 16, cval, kc, Plo, Phi, c16lo, c16hi, eacc);
            else group16<V>(st + gi * 16, cval, kc, Plo, Phi, c16lo, c16hi, eacc);
            if ((g & 1u) == 0u) {
                pend16lo = c16lo; pend16hi = c16hi;
            } else {
                uint32_t c32lo, c32hi;
                csa(c32lo, Plo[4], Plo[4], pend16lo, c16lo);
                csa(c32hi, Phi[4], Phi[4], pend16hi, c16hi);
                if ((g & 2u) == 0u) {
                    pend32lo = c32lo; pend32hi = c32hi;
                } else {
                    uint32_t c64lo, c64hi;
                    csa(c64lo, Plo[5], Plo[5], pend32lo, c32lo);
                    csa(c64hi, Phi[5], Phi[5], pend32hi, c32hi);
                    if constexpr (V & 32) {
                        if ((g & 4u) == 0u) {
                            pend64lo = c64lo; pend64hi = c64hi;
                        } else {
                            uint32_t c128lo, c128hi;
                            csa(c128lo, Plo[6], Plo[6], pend64lo, c64lo);
                            csa(c128hi, Phi[6], Phi[6], pend64hi, c64hi);
                            ripple<7>(Plo, c128lo);
                            ripple<7>(Phi, c128hi);
                        }
                    } else {
                        ripple<6>(Plo, c64lo);
                        ripple<6>(Phi, c64hi);
                    }
                }
            }
        }
        __syncwarp();
    }
    if (g & 1u) { ripple<4>(Plo, pend16lo); ripple<4>(Phi, pend16hi); }
    if (g & 2u) { ripple<5>(Plo, pend32lo); ripple<5>(Phi, pend32hi); }
    if ((V & 32) && (g & 4u)) { ripple<6>(Plo, pend64lo); ripple<6>(Phi, pend64hi); }

    int32_t* row = out + static_cast<size_t>(wid) * 2048;
    uint32_t ecount = 0;
    if constexpr (V & 64) {
        uint32_t low31 = 0;                      // sum over hashes of popc(hi & 0x7fffffff) = sum_p 2^p * popc(plane_p & 0x7fffffff)
#pragma unroll
        for (int p = 0; p < kPlanes; ++p) low31 += static_cast<uint32_t>(__popc(Phi[p] & 0x7FFFFFFFu)) << p;
        ecount = eacc - low31;
    }
#pragma unroll
    for (int w = 0; w < 2; ++w) {
#pragma unroll
        for (int jq = 0; jq < 8; ++jq) {
            int32_t v[4];
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                const int j = jq * 4 + jj;
                uint32_t val = 0;
#pragma unroll
                for (int p = 0; p < kPlanes; ++p) {
                    const uint32_t plane = w ? Phi[p] : Plo[p];
                    val |= ((plane >> j) & 1u) << p;
                }
                if ((V & 64) && w == 1 && j == 0) val = ecount;
                v[jj] = static_cast<int32_t>(per) - 2 * static_cast<int32_t>(val);
            }
            *reinterpret_cast<int4*>(row + lane * 64 + w * 32 + jq * 4) = make_int4(v[0], v[1], v[2], v[3]);
        }
    }
}

__global__ void gen_kernel(uint64_t* h, uint64_t n)
{
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        uint64_t x = i * 0x9e3779b97f4a7c15ULL + 12345;
        x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
        x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
        h[i] = (x ^ (x >> 31)) >> 10;
    }
}

__global__ void checksum_kernel(const int32_t* v, uint64_t n, unsigned long long* acc)
{
    unsigned long long s = 0;
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
        s += static_cast<unsigned long long>(static_cast<long long>(v[i]) * static_cast<long long>((i % 1021) + 1));
    atomicAdd(acc, s);
}

static uint64_t* d_h;
static unsigned long long* d_acc;

// ---- issue-mix microbenchmarks: 8 independent chains per thread ----------------------
template <int MODE>
__global__ void __launch_bounds__(256) mix_kernel(uint32_t* sink, int iters, uint32_t k0, uint32_t k1, uint32_t k2)
{
    uint32_t a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x * 2654435761u + i * k0 + blockIdx.x; b[i] = a[i] ^ k1; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) {            // LOP3 three distinct registers
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(a[(i + 1) & 7]));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(b[i]) : "r"(a[i]), "r"(b[(i + 3) & 7]));
            } else if (MODE == 1) {     // IMAD.HI : LOP3 = 1 : 1
                asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(k2));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(k0), "r"(k1));
            } else if (MODE == 2) {     // IMAD.HI : LOP3 = 1 : 2
                asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(k2));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(k0), "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(b[i]) : "r"(k0), "r"(k1));
            } else if (MODE == 3) {     // IMAD : LOP3 = 1 : 2
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(k0), "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(b[i]) : "r"(k0), "r"(k1));
            } else if (MODE == 4) {     // IMAD : LOP3 : SHF = 1 : 1 : 1
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(k0), "r"(k1));
                asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(b[i]) : "r"(k1));
            } else if (MODE == 5) {     // IMAD.WIDE : LOP3 = 1 : 2 (wide result both halves used)
                unsigned long long w;
                asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[i]), "r"(k0));
                asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a[i]) : "r"(static_cast<uint32_t>(w)), "r"(static_cast<uint32_t>(w >> 32)), "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(b[i]) : "r"(k0), "r"(k1));
            } else if (MODE == 6) {     // FFMA : LOP3 = 1 : 1  (does fp32 co-issue with ALU?)
                float f = __uint_as_float(a[i]);
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(__uint_as_float(k0)), "f"(__uint_as_float(k1)));
                a[i] = __float_as_uint(f);
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(k0), "r"(k1));
            } else if (MODE == 7) {     // FFMA : IMAD : LOP3 = 1 : 1 : 1
                float f = __uint_as_float(a[i]);
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(__uint_as_float(k0)), "f"(__uint_as_float(k1)));
                a[i] = __float_as_uint(f);
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(k0), "r"(k1));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(b[i]) : "r"(k0), "r"(k1));
            } else if (MODE == 8) {     // LOP3 : IMAD = 3 : 2 (close to the loop's mix)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(b[i]) : "r"(k0), "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(b[i]) : "r"(k0), "r"(k1));
                asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(k1));
            }
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x ^= a[i] ^ b[i];
    if (x == 0x12345u) sink[0] = x;
}

// ---- the projection loop's own instruction mix without its data dependencies -------------------------------------
// Per (hash, lane) round the product kernel issues 9 LOP3, 3 SHF, 5 IMAD, 3 IMAD.IADD, 2 IMAD.WIDE, 1 POPC, 1 LDS.128
// (+ ~1 of loop overhead): profiles/r02a_project_kernel_ncu_full.txt.  loopmix_kernel issues exactly that mix on 4
// independent register sets per thread, so the time per round is what the issue ports / pipes allow for this mix when
// no instruction ever waits for an operand.  DROP bits remove one ingredient (1: WIDE -> IMAD, 2: no POPC, 4: no LDS,
// 8: IMAD.IADD -> nothing, 16: 3 fewer LOP3) to price it.
template <int DROP>
__global__ void __launch_bounds__(128) loopmix_kernel(uint32_t* sink, int iters, uint32_t k0, uint32_t k1, uint32_t one)
{
    __shared__ uint4 tab[64];
    if (threadIdx.x < 64) tab[threadIdx.x] = make_uint4(threadIdx.x * k0, k1, k0 ^ threadIdx.x, k1 + threadIdx.x);
    __syncthreads();
    uint32_t a[4], b[4], c[4], e[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[i] = threadIdx.x * 2654435761u + i * k0 + blockIdx.x; b[i] = a[i] ^ k1; c[i] = a[i] + k0; e[i] = i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            uint4 t = make_uint4(k0, k1, k0, k1);
            if (!(DROP & 4)) t = tab[(it + i) & 63];                                      // LDS.128, warp-uniform address
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(a[i]) : "r"(one), "r"(a[i]), "r"(t.x));   // IMAD.IADD (lane add)
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(t.y), "r"(k0));          // ^ F
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b[i]) : "r"(a[i]), "r"(k1), "r"(t.w));    // l1 * K1hi + M
            if (DROP & 1) {
                asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(a[i]), "r"(k0), "r"(c[i]));
            } else {
                unsigned long long w;
                asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[i]), "r"(k0));   // IMAD.WIDE
                a[i] = static_cast<uint32_t>(w);
                c[i] = static_cast<uint32_t>(w >> 32);
            }
            if (!(DROP & 8)) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(one), "r"(c[i]), "r"(b[i]));  // hi + t
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 27;" : "=r"(b[i]) : "r"(a[i]), "r"(c[i]));       // funnel 27
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(k1));
            asm volatile("shf.r.clamp.b32 %0, %1, %2, 27;" : "=r"(b[i]) : "r"(c[i]), "r"(k0));        // hi >> 27
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(c[i]) : "r"(b[i]), "r"(k0));
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b[i]) : "r"(c[i]), "r"(k0), "r"(k1));     // yh * K2lo
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b[i]) : "r"(a[i]), "r"(k1), "r"(b[i]));   // + yl * K2hi
            if (DROP & 1) {
                asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(a[i]), "r"(k1), "r"(c[i]));
            } else {
                unsigned long long w;
                asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[i]), "r"(k1));
                a[i] = static_cast<uint32_t>(w);
                c[i] = static_cast<uint32_t>(w >> 32);
            }
            if (!(DROP & 8)) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(one), "r"(c[i]), "r"(b[i]));
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 31;" : "=r"(b[i]) : "r"(a[i]), "r"(c[i]));       // funnel 31
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(k1));
            if (!(DROP & 2)) {
                uint32_t p;
                asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(p) : "r"(c[i]), "r"(k0), "r"(k1));    // hi * fold
                asm volatile("popc.b32 %0, %0;" : "+r"(p));
                if (!(DROP & 8)) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(e[i]) : "r"(one), "r"(e[i]), "r"(p));
                else e[i] ^= p;
            }
            // carry-save tree: ~4.5 LOP3 per round on two words
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(a[i]), "r"(c[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(e[i]) : "r"(a[i]), "r"(c[i]));
            if (!(DROP & 16)) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(e[i]), "r"(k0));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(e[i]) : "r"(b[i]), "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(c[i]) : "r"(b[i]), "r"(e[i]));
            }
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) x ^= a[i] ^ b[i] ^ c[i] ^ e[i];
    if (x == 0x12345u) sink[0] = x;
}

template <int DROP>
void run_loopmix(const char* name)
{
    uint32_t* sink = reinterpret_cast<uint32_t*>(d_acc);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int iters = 4096;
    for (int ctas = 4; ctas <= 8; ctas += 2) {   // resident CTAs of 128 threads per SM: 4, 6 (the product's), 8 warps per scheduler
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            CK(cudaEventRecord(e0));
            loopmix_kernel<DROP><<<148 * ctas, 128>>>(sink, iters, 0x9e3779b9u, 0x85ebca6bu, 1u);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0 && ms < best) best = ms;
        }
        const double rounds = 148.0 * ctas * 4 /*warps*/ * iters * 4.0;        // (hash, lane) rounds per launch, all SMs
        const double clk = best * 1e-3 * 1.965e9 * 592.0;
        printf("loopmix %-34s drop=%2d  %d warps/scheduler  %7.3f ms  %6.2f clk per round\n", name, DROP, ctas, best, clk / rounds);
    }
}

template <int MODE>
void run_mix(const char* name, int per_iter)
{
    uint32_t* sink = reinterpret_cast<uint32_t*>(d_acc);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int iters = 4096;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        mix_kernel<MODE><<<148 * 8, 256>>>(sink, iters, 0x9e3779b9u, 0x85ebca6bu, 32u);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    const double instr = 148.0 * 8 * 8 /*warps*/ * iters * 8.0 * per_iter;   // warp instructions
    const double clk = best * 1e-3 * 1.965e9 * 592.0;                      // SMSP-clocks
    printf("mix %-40s %7.3f ms  %.3f warp-instr/clk/SMSP  (%.2f clk per group of %d)\n", name, best, instr / clk,
           clk / (instr / per_iter), per_iter);
}

static int32_t* d_out;
static uint32_t g_ntasks, g_per;
static int g_reps = 6;

template <int V, int WARPS, int TILE, int MINB = 1>
void run(const char* name)
{
    const Consts kc{4u, 32u, 2u};
    const uint32_t grid = (g_ntasks + WARPS - 1) / WARPS;
    const size_t smem = static_cast<size_t>(WARPS) * 2 * TILE * 8 + ((V & 128) ? static_cast<size_t>(WARPS) * TILE * 16 : 0);
    CK(cudaFuncSetAttribute(kexp_kernel<V, WARPS, TILE, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    CK(cudaMemset(d_out, 0, static_cast<size_t>(g_ntasks) * 2048 * 4));
    for (int rep = 0; rep < g_reps; ++rep) {
        CK(cudaEventRecord(e0));
        kexp_kernel<V, WARPS, TILE, MINB><<<grid, WARPS * 32, smem>>>(d_h, g_ntasks, g_per, kc, d_out);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if ((rep >= 2 || g_reps < 3) && ms < best) best = ms;
    }
    CK(cudaMemset(d_acc, 0, 8));
    checksum_kernel<<<1184, 256>>>(d_out, static_cast<uint64_t>(g_ntasks) * 2048, d_acc);
    unsigned long long acc;
    CK(cudaMemcpy(&acc, d_acc, 8, cudaMemcpyDeviceToHost));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, kexp_kernel<V, WARPS, TILE, MINB>));
    const double hashes = static_cast<double>(g_ntasks) * g_per;
    printf("%-34s V=%4d W=%d T=%4d minb=%d regs=%3d  %8.3f ms  %7.3f Ghash/s  clk/hash/warp=%6.2f  sum=%016llx\n", name, V, WARPS,
           TILE, MINB, fa.numRegs, best, hashes / best / 1e6, 1.965e9 * 592.0 / (hashes / (best * 1e-3)), acc);
    fflush(stdout);
}

int main(int argc, char** argv)
{
    g_ntasks = argc > 1 ? atoi(argv[1]) : 26000;
    g_per = argc > 2 ? atoi(argv[2]) : 3840;
    g_reps = argc > 3 ? atoi(argv[3]) : 6;
    const uint64_t n = static_cast<uint64_t>(g_ntasks) * g_per;
    CK(cudaMalloc(&d_h, n * 8));
    CK(cudaMalloc(&d_out, static_cast<size_t>(g_ntasks) * 2048 * 4));
    CK(cudaMalloc(&d_acc, 8));
    gen_kernel<<<1184, 256>>>(d_h, n);
    CK(cudaDeviceSynchronize());
    if (argc > 4) {
        run_mix<0>("LOP3 x2, three distinct regs", 2);
        run_mix<1>("IMAD.HI : LOP3 1:1", 2);
        run_mix<2>("IMAD.HI : LOP3 1:2", 3);
        run_mix<3>("IMAD : LOP3 1:2", 3);
        run_mix<4>("IMAD : LOP3 : SHF 1:1:1", 3);
        run_mix<5>("IMAD.WIDE : LOP3 1:2", 3);
        run_mix<6>("FFMA : LOP3 1:1", 2);
        run_mix<7>("FFMA : LOP3 : IMAD 1:1:1", 3);
        run_mix<8>("LOP3 IMAD LOP3 IMAD SHF", 5);
    }
    if (argc > 4) {
        run_loopmix<0>("full mix");
        run_loopmix<1>("WIDE -> IMAD");
        run_loopmix<2>("no POPC (+IMAD)");
        run_loopmix<4>("no LDS");
        run_loopmix<8>("no IMAD.IADD");
        run_loopmix<16>("3 LOP3 fewer");
        run_loopmix<1 + 2 + 4>("no WIDE, POPC, LDS");
    }
#include "variants.inc"
    return 0;
}
