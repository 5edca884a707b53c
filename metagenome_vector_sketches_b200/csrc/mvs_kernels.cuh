// mvs_kernels.cuh -- sm_100a device code of the random-projection sketching path.
//
// What is computed (reference src/random_projection.cpp:9-26):
//     vec[j] = sum over hashes h of ( 1 - 2 * bit_{j%64}( splitmix64(h + 64*(j/64)) ) )
//            = |H| - 2*C[j],   C[j] = number of hashes whose bit is 1.
//
// How: a *bit-sliced positional popcount*.  A lane owns one 64-column block b of one
// hash stream.  For every hash it evaluates the 64-bit mixer once (two 32-bit words =
// 64 columns) and feeds the two words into a Harley-Seal carry-save-adder tree whose
// bit planes (weights 1,2,4,...,2^13: kPlanes = 14) live in registers; one CSA is two LOP3.  No
// per-column add is ever issued in the loop; the planes are expanded into int32
// columns once per warp task (<= kMaxRounds = 16320 hashes per lane).  No matrix is materialised,
// no atomics are used, integer adds make every partition order bit-exact.
//
// Work decomposition: a *warp task* = (sample, contiguous hash range, 32-block column
// pass).  Lanes of a warp = LPH blocks x S hash slots (LPH*S = 32): d = 2048 gives
// LPH = 32, S = 1 (one hash broadcast to 32 blocks); d = 256 gives LPH = 4, S = 8.
// Hashes are staged global -> shared with cp.async (LDGSTS, zero-filled past the end
// of the range) in a two-stage ring private to the warp and read back as broadcast
// LDS.  Padding zeros are processed like real hashes and their known contribution
// (splitmix64(0 + c) per lane) is subtracted at the flush, so the inner loop has no
// masking at all.
//
// Everything of the mixer that does not depend on the lane is hoisted out of the (hash, lane) loop
// into a per-tile pass ("cooking"): once a tile of raw hashes has landed, the warp computes for every
// HASH, not per hash and lane,
//     hb = hash + base,  base = golden + seed + 2048*pass            (the 64-bit add)
//     F  = (lo32(hb) >> 30) | (hi32(hb) << 2)                         (low word of hb >> 30)
//     M  = (hi32(hb) ^ (hi32(hb) >> 30)) * lo32(K1)                   (high-word term of the 1st multiply)
// and stores the 16-byte record {hb, F, M}.  A lane then only adds its own  off = 64*block < 2048  to
// the low word:  xl = lo32(hb) + off,  l1 = xl ^ F,  and the first multiply is
// (hi32(l1*lo32(K1)) + l1*hi32(K1) + M, lo32(l1*lo32(K1))).  This is exact as long as the add neither
// carries out of the low word nor changes its bits 30..31, i.e. unless
// (lo32(hb) mod 2^30) >= 2^30 - 2048, about one hash in 5e5; the tile pass votes on it and a flagged
// tile (5e-4 of them) runs the same loop with the plain 64-bit arithmetic on hb.  Per (hash, lane)
// this removes 3 SHF/LOP3/IADD from the ALU pipe and 1 IMAD (tools/kexp: 38.5 -> 34.2 clocks).
//
// The last mixer step  x ^= x >> 31  changes only bit 0 of the high word (bit0 ^= bit31), i.e.
// one column (32) of the 64-column block, for two ALU-pipe instructions per (hash, lane).
// The loop therefore feeds the carry-save tree with the high word BEFORE that step (all its
// other 31 columns are already final) and counts column 32 on the side without the ALU pipe:
// t = hi * 0x80000001 (IMAD) carries bit0 ^ bit31 in bit 31 above hi's bits 0..30, so
// popc(t) = popc(hi & 0x7fffffff) + (bit0 ^ bit31)  (POPC runs on the XU pipe, the running sum
// on the FMA pipe), and the first term summed over the hashes equals the sum of the plane
// counts of slots 0..30, known exactly at the flush.  Measured on B200 (tools/kexp): the ALU
// pipe is the binding unit (IMAD.WIDE / IMAD.HI also take an ALU issue slot), 42.7 -> 38.6
// clocks per (hash, warp) with both hoists.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mvs {

constexpr int kWarpsPerCta = 4;
// hashes per staging tile per warp: 256 (2 KB raw x 2 stages + 4 KB cooked), but at least one
// Harley-Seal group of 16 rounds for all S hash slots (S = 32, i.e. d <= 64, needs 512)
__host__ __device__ constexpr int tile_hashes(int lph_log2) { return (32 >> lph_log2) * 16 > 256 ? (32 >> lph_log2) * 16 : 256; }
constexpr int kStages = 2;
constexpr int kPlanes = 14;                      // bit planes: counts up to 16383 per lane
constexpr uint32_t kMaxRounds = 16320;           // hashes per lane per task (255 trips of 64; 14 planes count to 16383)
// The flush adds the S hash slots of a warp as PACKED pairs of 16-bit column counts, so a task may hold at most
// 65535 hashes per 64-column block: rounds per lane <= 65535 / S, a multiple of 64 (S = 32: of 16, one group per trip).
__host__ __device__ constexpr uint32_t max_rounds(int lph_log2)
{
    const uint32_t S = 32u >> lph_log2;
    const uint32_t r = (65535u / S) / 64u * 64u;
    return r < kMaxRounds ? r : kMaxRounds;
}
constexpr uint32_t kDirect = 0xFFFFFFFFu;
constexpr uint64_t kGolden = 0x9e3779b97f4a7c15ULL;
constexpr uint32_t kFold = 0x80000001u;          // passed as a kernel argument so that the multiply stays an IMAD
constexpr uint32_t kCarryRisk = 0xFFFFF800u;     // (lo32(hb) | 0xC0000000) >= this: + 64*block may reach bit 30
constexpr uint32_t kK1lo = 0x1ce4e5b9u, kK1hi = 0xbf58476du;   // 0xbf58476d1ce4e5b9

struct __align__(16) WarpTask {                  // 32 bytes
    uint64_t begin;                              // first hash (index into the batch array)
    uint32_t count;                              // hashes of this chunk
    uint32_t sample;                             // output row
    uint32_t pass;                               // column pass: blocks [32*pass, 32*pass+32)
    uint32_t prow;                               // partial row, or kDirect
    uint32_t pad0, pad1;
};

struct __align__(16) ReduceItem {                // 16 bytes
    uint32_t sample;
    uint32_t prow;                               // first partial row
    uint32_t nchunks;
    uint32_t n;                                  // |H_s|
};

// ---- splitmix64 finaliser (reference src/random_projection.cpp:14-17) --------------
__device__ __forceinline__ void mix64(uint64_t x, uint32_t& lo, uint32_t& hi)
{
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    x = x ^ (x >> 31);
    lo = static_cast<uint32_t>(x);
    hi = static_cast<uint32_t>(x >> 32);
}

// c = cooked record of one hash: c.x = hb = hash + base, c.y = {F, M} (see header); off = 64*block of
// this lane (< 2048).  EXACT = false: the tile pass has proved that lo32(hb) + off stays below bit 30.
// Outputs: lo = final low word; hi = high word BEFORE the last xor-shift (columns 33..63 are
// final, slot 0 holds the raw bit 0); eacc += popc(hi * fold), fold = 0x80000001 (see header).
template <bool EXACT>
__device__ __forceinline__ void mix_staged(ulonglong2 c, uint32_t off, uint32_t fold, uint32_t& lo, uint32_t& hi,
                                           uint32_t& eacc)
{
    uint64_t x;
    if constexpr (EXACT) {
        x = c.x + off;
        x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    } else {
        const uint32_t xl = static_cast<uint32_t>(c.x) + off;
        const uint32_t l1 = xl ^ static_cast<uint32_t>(c.y);
        uint32_t t = static_cast<uint32_t>(c.y >> 32);
        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(t) : "r"(l1), "r"(kK1hi));
        uint64_t w;
        asm("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(l1), "r"(kK1lo));
        x = (static_cast<uint64_t>(static_cast<uint32_t>(w >> 32) + t) << 32) | static_cast<uint32_t>(w);
    }
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    const uint32_t xl = static_cast<uint32_t>(x);
    const uint32_t xh = static_cast<uint32_t>(x >> 32);
    lo = xl ^ __funnelshift_r(xl, xh, 31);
    hi = xh;
    uint32_t t;
    asm("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(xh), "r"(fold));
    eacc += static_cast<uint32_t>(__popc(t));
}

// tile pass: raw hash -> cooked record {hb, F, M} (see the header), stored as the two 8-byte halves of the 16-byte
// record at shared address rec + OFF: {hb} and {F, M} are each produced in a register pair, so no moves into a register
// quad are needed (a 16-byte store cost four IMAD.MOV per hash).  worst = max over the hashes of lo32(hb) << 2.
template <int OFF>
__device__ __forceinline__ void cook_store(uint32_t rec, uint64_t raw, uint64_t base, uint32_t& worst)
{
    const uint64_t hb = raw + base;
    const uint32_t xl = static_cast<uint32_t>(hb);
    const uint32_t xh = static_cast<uint32_t>(hb >> 32);
    worst = max(worst, xl << 2);                 // (xl mod 2^30) >= 2^30 - 2048  <=>  (xl << 2) >= 0xFFFFE000
    const uint32_t f = (xl >> 30) | (xh << 2);
    const uint32_t m = (xh ^ (xh >> 30)) * kK1lo;
    asm volatile("st.shared.v2.u32 [%0 + %3], {%1, %2};" ::"r"(rec), "r"(xl), "r"(xh), "n"(OFF) : "memory");
    asm volatile("st.shared.v2.u32 [%0 + %3], {%1, %2};" ::"r"(rec), "r"(f), "r"(m), "n"(OFF + 8) : "memory");
}
template <int I, int N>
__device__ __forceinline__ void cook_rows(uint32_t rec, const uint64_t* __restrict__ st, uint64_t base, uint32_t& worst)
{
    if constexpr (I < N) {
        cook_store<I * 512>(rec, st[I * 32], base, worst);               // rows of 32 lanes x 16 bytes
        cook_rows<I + 1, N>(rec, st, base, worst);
    }
}

// carry-save adder on 32 bit columns at once: s = a^b^c, c = maj(a,b,c)
__device__ __forceinline__ void csa(uint32_t& carry, uint32_t& sum, uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t s, m;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(s) : "r"(a), "r"(b), "r"(c));
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(m) : "r"(a), "r"(b), "r"(c));
    sum = s;
    carry = m;
}

// add the one-bit-per-column word c (weight 2^L) into planes L..kPlanes-1
template <int L>
__device__ __forceinline__ void ripple(uint32_t (&P)[kPlanes], uint32_t c)
{
#pragma unroll
    for (int p = L; p < kPlanes; ++p) {
        uint32_t t = P[p] & c;
        P[p] ^= c;
        c = t;
    }
}

// 4 weight-1 words into (ones, twos) -> one weight-4 carry
__device__ __forceinline__ uint32_t hs4(uint32_t (&P)[kPlanes], uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    uint32_t tA, tB, f;
    csa(tA, P[0], P[0], a, b);
    csa(tB, P[0], P[0], c, d);
    csa(f, P[1], P[1], tA, tB);
    return f;
}

// One Harley-Seal group: 16 hashes of this lane's stream -> planes 0..3 + a weight-16
// carry per word.  sp points at the lane's first cooked record of the group in shared
// memory; consecutive hashes of the lane are S records apart.
template <int S, bool EXACT>
__device__ __forceinline__ void group16(const ulonglong2* __restrict__ sp, uint32_t off, uint32_t fold,
                                        uint32_t (&Plo)[kPlanes], uint32_t (&Phi)[kPlanes],
                                        uint32_t& c16lo, uint32_t& c16hi, uint32_t& eacc)
{
    uint32_t lo[16], hi[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) mix_staged<EXACT>(sp[i * S], off, fold, lo[i], hi[i], eacc);   // LDS.128 (broadcast if S == 1)
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

// One loop trip = TRIP consecutive Harley-Seal groups of this lane's stream (TRIP = 4, 2 or 1: as many as a tile holds,
// at most 4).  With four groups per trip (d >= 1024) the carries of the groups meet inside the trip -- weight 16 pairwise
// into weight 32, those into weight 64, rippled into planes 6.. -- so the loop carries no pending-carry state and has no
// branch; shorter trips keep the pending words across trips (g = groups done so far, a multiple of TRIP).
struct Pending { uint32_t c16lo, c16hi, c32lo, c32hi, c64lo, c64hi, c128lo, c128hi; };

// A weight-64 carry (one per 64 rounds of the lane) joins the planes.  Rippling it through planes 6..13 at once costs
// 32 LOP3 per 64 rounds (0.5 per round, 3.5 % of the loop's ALU-slot instructions); instead it is paired with the next
// one by a carry-save adder on plane 6, those pairs on plane 7, and only the weight-256 carry is rippled (planes
// 8..13): 0.14 LOP3 per round.  g = groups done before this trip (warp-uniform), so the branches do not diverge.
// DEEP = false ripples at once: the one-group-per-trip instantiations (d <= 256) lose 1.3 % to the deeper branch nest
// (config 3 at d = 256: 0.470 -> 0.477 ms), the others gain 1.1 % (config 2: 14.48 -> 14.32 ms).  All paths of one
// instantiation (packed and plain 64-bit arithmetic) use the same scheme, because they share the pending words.
template <bool DEEP>
__device__ __forceinline__ void carry64(uint32_t (&Plo)[kPlanes], uint32_t (&Phi)[kPlanes], Pending& pend, uint32_t g,
                                        uint32_t c64lo, uint32_t c64hi)
{
    if constexpr (!DEEP) {
        ripple<6>(Plo, c64lo);
        ripple<6>(Phi, c64hi);
        return;
    }
    if ((g & 4u) == 0u) {
        pend.c64lo = c64lo; pend.c64hi = c64hi;
        return;
    }
    uint32_t c128lo, c128hi;
    csa(c128lo, Plo[6], Plo[6], pend.c64lo, c64lo);
    csa(c128hi, Phi[6], Phi[6], pend.c64hi, c64hi);
    if ((g & 8u) == 0u) {
        pend.c128lo = c128lo; pend.c128hi = c128hi;
        return;
    }
    uint32_t c256lo, c256hi;
    csa(c256lo, Plo[7], Plo[7], pend.c128lo, c128lo);
    csa(c256hi, Phi[7], Phi[7], pend.c128hi, c128hi);
    ripple<8>(Plo, c256lo);
    ripple<8>(Phi, c256hi);
}

template <int S, bool EXACT, int TRIP, bool DEEP>
__device__ __forceinline__ void trip_groups(const ulonglong2* __restrict__ sp, uint32_t off, uint32_t fold,
                                            uint32_t (&Plo)[kPlanes], uint32_t (&Phi)[kPlanes], Pending& pend, uint32_t g,
                                            uint32_t& eacc)
{
    if constexpr (TRIP == 4) {
        uint32_t c32lo[2], c32hi[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            uint32_t alo, ahi, blo, bhi;
            group16<S, EXACT>(sp + (2 * h) * (16 * S), off, fold, Plo, Phi, alo, ahi, eacc);
            group16<S, EXACT>(sp + (2 * h + 1) * (16 * S), off, fold, Plo, Phi, blo, bhi, eacc);
            csa(c32lo[h], Plo[4], Plo[4], alo, blo);
            csa(c32hi[h], Phi[4], Phi[4], ahi, bhi);
        }
        uint32_t c64lo, c64hi;
        csa(c64lo, Plo[5], Plo[5], c32lo[0], c32lo[1]);
        csa(c64hi, Phi[5], Phi[5], c32hi[0], c32hi[1]);
        carry64<DEEP>(Plo, Phi, pend, g, c64lo, c64hi);
    } else if constexpr (TRIP == 2) {
        uint32_t alo, ahi, blo, bhi, c32lo, c32hi;
        group16<S, EXACT>(sp, off, fold, Plo, Phi, alo, ahi, eacc);
        group16<S, EXACT>(sp + 16 * S, off, fold, Plo, Phi, blo, bhi, eacc);
        csa(c32lo, Plo[4], Plo[4], alo, blo);
        csa(c32hi, Phi[4], Phi[4], ahi, bhi);
        if ((g & 2u) == 0u) {
            pend.c32lo = c32lo; pend.c32hi = c32hi;
        } else {
            uint32_t c64lo, c64hi;
            csa(c64lo, Plo[5], Plo[5], pend.c32lo, c32lo);
            csa(c64hi, Phi[5], Phi[5], pend.c32hi, c32hi);
            carry64<DEEP>(Plo, Phi, pend, g, c64lo, c64hi);
        }
    } else {
        uint32_t c16lo, c16hi;
        group16<S, EXACT>(sp, off, fold, Plo, Phi, c16lo, c16hi, eacc);
        if ((g & 1u) == 0u) {
            pend.c16lo = c16lo; pend.c16hi = c16hi;
        } else {
            uint32_t c32lo, c32hi;
            csa(c32lo, Plo[4], Plo[4], pend.c16lo, c16lo);
            csa(c32hi, Phi[4], Phi[4], pend.c16hi, c16hi);
            if ((g & 2u) == 0u) {
                pend.c32lo = c32lo; pend.c32hi = c32hi;
            } else {
                uint32_t c64lo, c64hi;
                csa(c64lo, Plo[5], Plo[5], pend.c32lo, c32lo);
                csa(c64hi, Phi[5], Phi[5], pend.c32hi, c32hi);
                carry64<DEEP>(Plo, Phi, pend, g, c64lo, c64hi);
            }
        }
    }
}

// Bit planes -> column counts.  A[p] = plane p (weight 2^p) of 32 columns, A[14] = A[15] = 0.  A 16 x 16 bit-matrix
// transpose carried out in both 16-bit halves of the words at once (four butterfly stages of 8 row pairs; one stage of
// a pair = 2 shifts + 2 bit-selects) leaves  A[c] = count(column c) | count(column 16 + c) << 16 : 128 instructions for
// 32 columns instead of 3 per (column, plane).
__device__ __forceinline__ void planes_to_packed_counts(uint32_t (&A)[16])
{
#pragma unroll
    for (int stage = 0; stage < 4; ++stage) {
        const int j = 8 >> stage;
        const uint32_t m = stage == 0 ? 0x00FF00FFu : stage == 1 ? 0x0F0F0F0Fu : stage == 2 ? 0x33333333u : 0x55555555u;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k & j) continue;
            const uint32_t lo = A[k], hi = A[k + j];
            A[k] = (lo & m) | ((hi << j) & ~m);
            A[k + j] = (hi & ~m) | ((lo >> j) & m);
        }
    }
}

__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc, uint32_t src_bytes)
{
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
// full copy with the same compile-time byte offset on both sides: one LDGSTS with immediate offsets, no address arithmetic
template <int OFF>
__device__ __forceinline__ void cp_async8_at(uint32_t smem_dst, const void* gsrc)
{
    asm volatile("cp.async.ca.shared.global [%0 + %2], [%1 + %2], 8;" ::"r"(smem_dst), "l"(gsrc), "n"(OFF) : "memory");
}
template <int I, int N>
__device__ __forceinline__ void cp_async8_rows(uint32_t smem_dst, const void* gsrc)   // rows of 32 lanes x 8 bytes
{
    if constexpr (I < N) {
        cp_async8_at<I * 256>(smem_dst, gsrc);
        cp_async8_rows<I + 1, N>(smem_dst, gsrc);
    }
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ------------------------------------------------------------------------------------
// The projection kernel.  One warp = one WarpTask.  LPH_LOG2 = log2(lanes per hash).
// ------------------------------------------------------------------------------------
template <int LPH_LOG2>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
project_kernel(const uint64_t* __restrict__ hashes, const WarpTask* __restrict__ tasks, uint32_t ntasks,
               int d, uint64_t seed_golden, int32_t* __restrict__ out, int32_t* __restrict__ partials,
               int vec4_ok, uint32_t fold)
{
    constexpr int LPH = 1 << LPH_LOG2;
    constexpr int S = 32 >> LPH_LOG2;
    constexpr int kTileHashes = tile_hashes(LPH_LOG2);
    constexpr int kGroupsPerTile = kTileHashes / (16 * S);
    static_assert(kGroupsPerTile >= 1, "a tile holds at least one group");
    // groups per loop trip: 2 when a tile holds at least 4 groups (d >= 512), else 1.  Measured on the full kernel
    // (profiles/r02e_ab_kernel.jsonl, same task caps): config 2 14.61 ms with 2, 14.66 with 1, 16.68 with 4 -- the four-group
    // body that wins by 2 % in the isolated loop (tools/kexp) loses 13 % here; d = 256 prefers 1 (0.517 vs 0.528 ms).
#ifndef MVS_TRIP_MAX
#define MVS_TRIP_MAX 2
#endif
#ifndef MVS_TRIP_MIN_GROUPS
#define MVS_TRIP_MIN_GROUPS 4
#endif
    constexpr int kTrip = kGroupsPerTile >= MVS_TRIP_MIN_GROUPS ? (MVS_TRIP_MAX < 4 ? MVS_TRIP_MAX : 4) : 1;
    static_assert(kGroupsPerTile % kTrip == 0, "a tile is a whole number of trips");
    constexpr bool kDeep = kTrip >= 2;           // weight-64 carries deferred through two more adder levels (carry64)

    extern __shared__ __align__(16) uint64_t smem[];
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t wid = blockIdx.x * kWarpsPerCta + warp;
    if (wid >= ntasks) return;                   // warps are independent: no CTA-wide barrier below

    const WarpTask task = tasks[wid];
    uint64_t* ring = smem + warp * (kStages * kTileHashes);                                  // raw hashes, 2 stages
    ulonglong2* cooked = reinterpret_cast<ulonglong2*>(smem + kWarpsPerCta * kStages * kTileHashes) + warp * kTileHashes;
    const uint64_t* src = hashes + task.begin;
    const uint32_t count = task.count;

    const int slot = lane >> LPH_LOG2;
    const int blk = static_cast<int>(task.pass) * 32 + (lane & (LPH - 1));
    const uint64_t base = seed_golden + (static_cast<uint64_t>(task.pass) << 11);   // added per hash in the tile pass
    const uint32_t off = static_cast<uint32_t>(lane & (LPH - 1)) << 6;              // added per lane, < 2048
    const uint64_t cval = base + off;                                               // == seed_golden + 64*blk

    uint32_t Plo[kPlanes], Phi[kPlanes];
#pragma unroll
    for (int p = 0; p < kPlanes; ++p) { Plo[p] = 0u; Phi[p] = 0u; }
    Pending pend{0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    uint32_t g = 0;                              // groups done (warp-uniform, a multiple of kTrip)
    uint32_t eacc = 0;                           // sum of popc(hi * fold) over this lane's hashes

    const uint32_t ntiles = (count + kTileHashes - 1) / kTileHashes;

    auto issue_tile = [&](uint32_t t) {
        uint64_t* st = ring + (t & (kStages - 1)) * kTileHashes;
        const uint32_t base = t * kTileHashes;
        if (base + kTileHashes <= count) {       // a full tile (all but the last of a task): no bounds, no zero-fill
            cp_async8_rows<0, kTileHashes / 32>(static_cast<uint32_t>(__cvta_generic_to_shared(st + lane)), src + base + lane);
            return;
        }
#pragma unroll
        for (int i = 0; i < kTileHashes / 32; ++i) {
            const uint32_t pos = base + i * 32 + lane;
            const bool ok = pos < count;
            cp_async8(st + i * 32 + lane, src + (ok ? pos : 0u), ok ? 8u : 0u);   // zero-fill past the end
        }
    };

    if (ntiles) issue_tile(0);
    cp_async_commit();
    for (uint32_t t = 0; t < ntiles; ++t) {
        if (t + 1 < ntiles) issue_tile(t + 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncwarp();
        const uint64_t* st = ring + (t & (kStages - 1)) * kTileHashes;
        // tile pass: raw -> cooked records; vote whether any low word could reach bit 30 under + off
        uint32_t worst = 0;
        {
            // four rows (of 32 hashes) at a time: enough independent work to hide the LDS latency, few enough live
            // register quads to stay at 6 CTAs per SM
            uint32_t rec = static_cast<uint32_t>(__cvta_generic_to_shared(cooked + lane));
            const uint64_t* raw = st + lane;
#pragma unroll 1
            for (int c = 0; c < kTileHashes / 128; ++c, rec += 4 * 512, raw += 4 * 32) cook_rows<0, 4>(rec, raw, base, worst);
        }
        const bool exact = __any_sync(0xffffffffu, worst >= (kCarryRisk << 2));
        __syncwarp();
        const uint32_t valid = min(count - t * kTileHashes, static_cast<uint32_t>(kTileHashes));
        const uint32_t rounds = (valid + S - 1) / S;
        const uint32_t ntrips = (rounds + 16 * kTrip - 1) / (16 * kTrip);   // the tile's tail is cooked zero padding
        if (!exact) {
            for (uint32_t ti = 0; ti < ntrips; ++ti, g += kTrip)
                trip_groups<S, false, kTrip, kDeep>(cooked + ti * (kTrip * 16 * S) + slot, off, fold, Plo, Phi, pend, g, eacc);
        } else {                                 // 5e-4 of the tiles: plain 64-bit arithmetic, one group at a time (code size)
#pragma unroll 1
            for (uint32_t gi = 0; gi < ntrips * kTrip; ++gi, ++g)
                trip_groups<S, true, 1, kDeep>(cooked + gi * (16 * S) + slot, off, fold, Plo, Phi, pend, g, eacc);
        }
        __syncwarp();                            // stage and cooked tile may be overwritten next
    }
    // fold the pending carries (g = number of groups done)
    if (g & 1u) { ripple<4>(Plo, pend.c16lo); ripple<4>(Phi, pend.c16hi); }
    if (g & 2u) { ripple<5>(Plo, pend.c32lo); ripple<5>(Phi, pend.c32hi); }
    if constexpr (kDeep) {
        if (g & 4u) { ripple<6>(Plo, pend.c64lo); ripple<6>(Phi, pend.c64hi); }
        if (g & 8u) { ripple<7>(Plo, pend.c128lo); ripple<7>(Phi, pend.c128hi); }
    }

    // padding correction: this lane consumed 16*g hashes of which `mine` were real; the
    // others were zeros whose mixer output is the per-lane constant z.
    const uint32_t mine = (count > static_cast<uint32_t>(slot)) ? (count - slot + S - 1) / S : 0u;
    const uint32_t npad = 16u * g - mine;
    uint32_t zlo, zhi;
    mix64(cval, zlo, zhi);

    // ---- flush: expand planes into int32 columns, fold hash slots, store -------------
    const int nb = (d + 63) >> 6;
    const bool writer = (slot == 0) && (blk < nb);
    const bool direct = (task.prow == kDirect);
    int32_t* row = direct ? out + static_cast<size_t>(task.sample) * static_cast<size_t>(d)
                          : partials + static_cast<size_t>(task.prow) * static_cast<size_t>(d);
    const int32_t n_total = static_cast<int32_t>(count);
    const bool vec4 = vec4_ok != 0;             // d % 4 == 0 and 16-byte aligned rows (host-checked)

    // column 32 (bit0 ^ bit31 of the high word): eacc minus the hashes' popc(hi & 0x7fffffff),
    // which is sum_p 2^p * popc(plane_p & 0x7fffffff).  Padding zeros are included on both sides
    // and removed below like in every other column (zhi is the true final word of the padding).
    uint32_t col32 = eacc;
#pragma unroll
    for (int p = 0; p < kPlanes; ++p) col32 -= static_cast<uint32_t>(__popc(Phi[p] & 0x7FFFFFFFu)) << p;

    // planes -> packed counts: A[c] = columns c | 16 + c of the low word, B[c] = columns 32 + c | 48 + c
    uint32_t A[16], B[16];
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        A[p] = p < kPlanes ? Plo[p] : 0u;
        B[p] = p < kPlanes ? Phi[p] : 0u;
    }
    planes_to_packed_counts(A);
    planes_to_packed_counts(B);
    B[0] = (B[0] & 0xFFFF0000u) | col32;
    if (__any_sync(0xffffffffu, npad != 0u)) {   // only the last chunk of a sample is padded
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            A[c] -= npad * ((zlo >> c) & 0x00010001u);
            B[c] -= npad * ((zhi >> c) & 0x00010001u);
        }
    }
    // fold the S hash slots (packed adds: a task holds at most 65535 hashes per block, see max_rounds)
#pragma unroll
    for (int m = LPH; m < 32; m <<= 1) {
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            A[c] += __shfl_xor_sync(0xffffffffu, A[c], m);
            B[c] += __shfl_xor_sync(0xffffffffu, B[c], m);
        }
    }
    if (writer) {
#pragma unroll
        for (int part = 0; part < 4; ++part) {   // columns [16 * part, 16 * part + 16) of the block
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                int32_t v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t w = (part & 2) ? B[4 * q + i] : A[4 * q + i];
                    const uint32_t val = (part & 1) ? (w >> 16) : (w & 0xFFFFu);
                    v[i] = direct ? (n_total - 2 * static_cast<int32_t>(val)) : static_cast<int32_t>(val);
                }
                const int col = blk * 64 + part * 16 + q * 4;
                if (vec4) {
                    if (col < d) *reinterpret_cast<int4*>(row + col) = make_int4(v[0], v[1], v[2], v[3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (col + i < d) row[col + i] = v[i];
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------
// Partial reduce: samples split over several warp tasks.  Fixed chunk order, no atomics.
//   out[s][j] = |H_s| - 2 * sum_c partial[prow + c][j]
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
reduce_partials_kernel(const ReduceItem* __restrict__ items, uint32_t nitems, int d,
                       const int32_t* __restrict__ partials, int32_t* __restrict__ out)
{
    const uint32_t it = blockIdx.x;              // grid = (items, column blocks)
    if (it >= nitems) return;
    const ReduceItem r = items[it];
    const int32_t* p = partials + static_cast<size_t>(r.prow) * static_cast<size_t>(d);
    for (int j = blockIdx.y * blockDim.x + threadIdx.x; j < d; j += gridDim.y * blockDim.x) {
        int32_t acc = 0;
        for (uint32_t c = 0; c < r.nchunks; ++c) acc += p[static_cast<size_t>(c) * d + j];
        out[static_cast<size_t>(r.sample) * d + j] = static_cast<int32_t>(r.n) - 2 * acc;
    }
}

// ------------------------------------------------------------------------------------
// Epilogue of sketch() (reference src/project_everything.cpp:327-347): per row the exact
// sum of squares, the int16-saturated copy and the norm the reference prints.  One warp per row.
//
// The reference norm (:328-329) is  (vec.cast<float>() / std::sqrt(float(d))).norm()  under
// -O3 -ffast-math with the vendored Eigen 3.2.10, i.e. (read off the compiled reference, both its
// -march=native and x86-64-v3 builds):
//     inv  = rsqrtss(float(d)) refined by one Newton-Raphson step   (host: the instruction is the CPU's)
//     x_j  = float(v_j) * inv                                       (one rounding each)
//     SSE packets of 4, two packet accumulators: acc0 = x[0:4]^2, acc1 = x[4:8]^2 (plain products),
//     then acc_k = fma(x, x, acc_k) over the following packet pairs, acc0 += acc1, one more fma for an
//     odd last packet, horizontal add (a0 + a1) + (a2 + a3), scalar fma for the <= 3 tail columns,
//     sqrtss.  Fewer than 4 columns: x_0^2 then scalar fma.
// Every step is a correctly rounded IEEE binary32 operation, so the same sequence of __fmul_rn /
// __fmaf_rn / __fadd_rn / __fsqrt_rn gives the same bits; lanes 0..7 of the warp own the 8 chains.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ float reference_norm_row(const int32_t* __restrict__ v, int d, float inv, int lane)
{
    const int a2 = (d / 8) * 8, a1 = (d / 4) * 4;
    float acc = 0.f;
    if (lane < 8 && lane < a1) {                 // chain `lane`: columns lane, lane + 8, ... below a2 (or the only packet)
        const float x0 = __fmul_rn(__int2float_rn(v[lane]), inv);
        acc = __fmul_rn(x0, x0);
        if (a1 > 4)
            for (int j = lane + 8; j < a2; j += 8) {
                const float x = __fmul_rn(__int2float_rn(v[j]), inv);
                acc = __fmaf_rn(x, x, acc);
            }
    }
    float res = 0.f;
    if (a1 >= 4) {
        if (a1 > 4) {
            const float other = __shfl_down_sync(0xffffffffu, acc, 4);
            acc = __fadd_rn(acc, other);         // lanes 0..3: acc0 + acc1
            if (a1 > a2 && lane < 4) {           // odd last packet
                const float x = __fmul_rn(__int2float_rn(v[a2 + lane]), inv);
                acc = __fmaf_rn(x, x, acc);
            }
        }
        const float nb = __shfl_xor_sync(0xffffffffu, acc, 1);
        const float pair = __fadd_rn(acc, nb);   // lane 0: a0 + a1, lane 2: a2 + a3
        const float hi = __shfl_down_sync(0xffffffffu, pair, 2);
        res = __fadd_rn(pair, hi);               // lane 0: (a0 + a1) + (a2 + a3)
        if (lane == 0)
            for (int j = a1; j < d; ++j) {
                const float x = __fmul_rn(__int2float_rn(v[j]), inv);
                res = __fmaf_rn(x, x, res);
            }
    } else if (lane == 0) {
        const float x0 = __fmul_rn(__int2float_rn(v[0]), inv);
        res = __fmul_rn(x0, x0);
        for (int j = 1; j < d; ++j) {
            const float x = __fmul_rn(__int2float_rn(v[j]), inv);
            res = __fmaf_rn(x, x, res);
        }
    }
    return __fsqrt_rn(res);                      // valid in lane 0
}

__global__ void __launch_bounds__(256)
finalize_kernel(const int32_t* __restrict__ vec, int64_t n_rows, int d,
                unsigned long long* __restrict__ sumsq, int16_t* __restrict__ out16,
                float inv_sqrt_d, float* __restrict__ ref_norms)
{
    const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n_rows) return;
    const int lane = threadIdx.x & 31;
    const int32_t* v = vec + row * d;
    if (sumsq || out16) {
        unsigned long long ss = 0;
        for (int j = lane; j < d; j += 32) {
            const int32_t x = v[j];
            ss += static_cast<unsigned long long>(static_cast<long long>(x) * static_cast<long long>(x));
            if (out16) out16[row * d + j] = static_cast<int16_t>(max(-32768, min(32767, x)));
        }
#pragma unroll
        for (int m = 16; m; m >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, m);
        if (lane == 0 && sumsq) sumsq[row] = ss;
    }
    if (ref_norms) {
        const float nrm = reference_norm_row(v, d, inv_sqrt_d, lane);
        if (lane == 0) ref_norms[row] = nrm;
    }
}

// ------------------------------------------------------------------------------------
// Packed input: hashes below 2^56 travel as 7 little-endian bytes each (12.5 % fewer bytes on the host link, which
// bounds the end-to-end rate) and are widened to the u64 array the projection reads.  Thread t reads the two aligned
// 8-byte words that hold bytes [7t, 7t + 7) of the packed slice (the slice starts at an 8-byte aligned device address
// and is over-allocated by 16 bytes) and writes hashes[t]: 7 B read + 8 B written per hash, HBM-bound, ~0.3 % of the
// projection's time.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
unpack56_kernel(const uint64_t* __restrict__ packed_words, uint64_t n, uint64_t* __restrict__ hashes)
{
    const uint64_t t = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const uint64_t byte = 7ull * t;
    const uint64_t w = byte >> 3;
    const unsigned sh = static_cast<unsigned>(byte & 7ull) * 8u;
    const uint64_t lo = packed_words[w], hi = packed_words[w + 1];
    const uint64_t v = sh ? ((lo >> sh) | (hi << (64u - sh))) : lo;
    hashes[t] = v & 0x00FFFFFFFFFFFFFFull;
}

// ------------------------------------------------------------------------------------
// Integer issue-rate microbenchmarks (roofline denominator).
// ------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256)
issue_bench_kernel(uint32_t* sink, int iters, uint32_t k0, uint32_t k1)
{
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 2654435761u + i * k0 + blockIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (MODE == 0) {
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                } else if (MODE == 1) {
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                } else if (MODE == 2) {
                    if (i & 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                    else       asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                } else if (MODE == 3) {
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(a[(i + 3) & 7]));   // not foldable
                } else if (MODE == 4) {
                    asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(k1));
                } else if (MODE == 5) {          // IMAD.WIDE.U32 (64-bit result, low word fed back)
                    unsigned long long w;
                    asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[i]), "r"(k0));
                    a[i] = static_cast<uint32_t>(w) ^ static_cast<uint32_t>(w >> 32);
                } else if (MODE == 6) {          // IMAD.HI.U32
                    asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                } else if (MODE == 7) {          // 64-bit funnel shift SHF.R.U64 + LOP3 pairs
                    asm volatile("shf.r.clamp.b32 %0, %0, %1, 27;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
                } else if (MODE == 8) {          // LOP3 : IMAD = 2 : 1
                    if ((i % 3) == 2) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                    else              asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                } else {                         // LOP3 : SHF = 1 : 1 (both ALU pipe?)
                    if (i & 1) asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(k1));
                    else       asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(k0), "r"(k1));
                }
            }
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x ^= a[i];
    if (x == 0x12345u) sink[0] = x;              // practically never; keeps the chains alive
}

// ------------------------------------------------------------------------------------
// The projection loop's own instruction mix WITHOUT its data dependencies: per (hash, lane) round project_kernel issues
// ~8.3 LOP3, 3 SHF, 4 IMAD, 3 IMAD.IADD, 2 IMAD.WIDE, 1 POPC, 1 LDS.128 and ~0.5 IADD3 (SASS of the hot loop; ncu
// per-opcode table in profiles/).  This kernel issues that mix (8 LOP3 + 1 every other round, the adds as multiply-adds
// by an opaque 1 so that they stay on the FMA-heavy pipe like ptxas places them) on four independent register sets per
// thread, 6 CTAs of 128 threads per SM like the projection.  Its time per round is what the SM's issue ports and integer
// pipes allow for this mix when no instruction ever waits for an operand: the denominator of roofline.mix_limit.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
loop_mix_kernel(uint32_t* sink, int iters, uint32_t k0, uint32_t k1, uint32_t one)
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
            const uint4 t = tab[(it + i) & 63];                                                       // LDS.128
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(a[i]) : "r"(one), "r"(a[i]), "r"(t.x));   // lane add
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(t.y), "r"(k0));          // ^ F
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b[i]) : "r"(a[i]), "r"(k1), "r"(t.w));    // l1 * K1hi + M
            unsigned long long w;
            asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[i]), "r"(k0));                  // IMAD.WIDE
            a[i] = static_cast<uint32_t>(w);
            c[i] = static_cast<uint32_t>(w >> 32);
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(one), "r"(c[i]), "r"(b[i]));  // hi + t
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 27;" : "=r"(b[i]) : "r"(a[i]), "r"(c[i]));       // xor-shift 27
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(k1));
            asm volatile("shf.r.clamp.b32 %0, %1, %2, 27;" : "=r"(b[i]) : "r"(c[i]), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(c[i]) : "r"(b[i]), "r"(k0));
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b[i]) : "r"(c[i]), "r"(k0), "r"(k1));     // yh * K2lo
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b[i]) : "r"(a[i]), "r"(k1), "r"(b[i]));   // + yl * K2hi
            asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[i]), "r"(k1));                  // IMAD.WIDE
            a[i] = static_cast<uint32_t>(w);
            c[i] = static_cast<uint32_t>(w >> 32);
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(one), "r"(c[i]), "r"(b[i]));  // hi add
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 31;" : "=r"(b[i]) : "r"(a[i]), "r"(c[i]));       // xor-shift 31, low word
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(k1));
            uint32_t p;
            asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(p) : "r"(c[i]), "r"(k0), "r"(k1));        // hi * fold
            asm volatile("popc.b32 %0, %0;" : "+r"(p));
            e[i] += p;                                                                                // running popcount sum
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(a[i]), "r"(c[i]));       // carry-save tree
            asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(e[i]) : "r"(a[i]), "r"(c[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(e[i]), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(c[i]) : "r"(b[i]), "r"(k1));
            if (it & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(c[i]) : "r"(b[i]), "r"(e[i]));
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) x ^= a[i] ^ b[i] ^ c[i] ^ e[i];
    if (x == 0x12345u) sink[0] = x;
}

}  // namespace mvs
