// mvs_pairwise_umma.cuh -- the limb GEMM of the all-pairs comparison on the 5th-generation tensor cores
// (tcgen05.mma kind::i8, accumulators in TMEM, operands staged by TMA with the 128-byte swizzle); default path of
// mvs_pairwise_block / mvs_pairwise_shard.  Math and limb decomposition: see mvs_pairwise.cuh.
//
//   C[i][j] = sum over limb pairs (p, q), p + q <= 3, of 256^(p+q) * sum_k A_p[i][k] * B_q[j][k]      (mod 2^32)
//
// Persistent CTAs (one per SM) loop over 128 x 128 tiles of C (CTA pairs with tcgen05 cta_group::2: over 256 x 128
// tiles, see pairwise_dot_umma_pair_kernel); 192 threads: warp 0 = TMA producer, warp 1 = MMA issuer (all lanes run
// the loops with uniform state, one elected lane issues), warps 2..5 = epilogue.  Every weight s = p + q owns its own
// 128-column int32 accumulator in TMEM (<= 4 weights = the SM's 512 columns), so limb pairs may run in any order and
// the Horner combine v = (v << 8) + acc_s happens once, in the epilogue, on the way from TMEM to registers.  Because
// of that freedom a pipeline stage holds ALL limb planes of one 128-byte k-slice (FUSED, when they fit: PA + PB <= 4,
// i.e. the normal 2-limb case) and every staged byte feeds up to four MMAs; with more limbs the planes are streamed
// pair by pair (STREAM, twice the operand bytes per MMA: that is L2-ingress bound, FUSED is tensor-pipe bound).
// The epilogue frees TMEM as soon as its row is in registers, then applies a lower-bound pre-filter of the keep-test
// (um_store_row) and stores only rows that may keep a pair; the exact test runs later on those segments only.
#pragma once
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

#include "mvs_pairwise.cuh"

namespace mvs {

constexpr int kUmBM = 128, kUmBN = 128, kUmBKB = 128, kUmK = 32;     // k in bytes (= int8 elements)
constexpr int kUmTile = kUmBM * kUmBKB;                              // 16 KB: one plane tile of A or of B
constexpr int kUmThreads = 192;
constexpr int kUmTmemCols = 512;
constexpr int kUmSmemBudget = 200 * 1024;

struct UmmaPlan {
    int ngroups;                                 // FUSED: 1, STREAM: number of limb pairs
    int a_count, b_count;                        // planes of A / of B per stage
    int stages;
    int smax;                                    // highest weight p + q in use
    unsigned char a_first[10], b_first[10];      // first plane of each group
    int probe;                                   // 0 in the product.  Diagnostic builds (-DMVS_UMMA_PROBE, tools/kexp) set bits to
                                                 // switch one role off and time the others: 1 = no TMA after the first ring fill,
                                                 // 2 = no TMEM loads, 4 = no stores of C, 8 = no MMAs (results are then wrong)
};

__device__ __forceinline__ uint32_t um_smem(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void um_mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void um_mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void um_mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok = 0;
    unsigned long long spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && ++spins > (1ull << 28)) __trap();                  // a lost arrival must not hang the GPU
    } while (!ok);
}
__device__ __forceinline__ void um_tma_load_3d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// shared-memory matrix descriptor: K-major, 128-byte swizzle, 8-row groups 1024 bytes apart, sm_100 descriptor version
__device__ __forceinline__ uint64_t um_desc_sw128(uint32_t addr)
{
    return static_cast<uint64_t>((addr & 0x3FFFFu) >> 4) | (static_cast<uint64_t>(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void um_mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void um_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void um_tmem_ld32(uint32_t taddr, uint32_t (&r)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
                   "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
                   "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void um_tmem_ld32_nowait(uint32_t taddr, uint32_t (&r)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
                   "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
                   "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
}
__device__ __forceinline__ void um_tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// TMEM -> registers for one thread's row of a tile: v[c] = sum_s 256^s * acc_s[c] (mod 2^32).  With three weights (the
// two-limb case) the three loads of a 32-column chunk are in flight together and share one wait.
__device__ __forceinline__ void um_drain_row(uint32_t tmem_row, int smax, bool skip, uint32_t (&v)[kUmBN])
{
    if (smax == 2 && !skip) {
#pragma unroll
        for (int c0 = 0; c0 < kUmBN; c0 += 32) {
            uint32_t r0[32], r1[32], r2[32];
            um_tmem_ld32_nowait(tmem_row + static_cast<uint32_t>(0 * kUmBN + c0), r0);
            um_tmem_ld32_nowait(tmem_row + static_cast<uint32_t>(1 * kUmBN + c0), r1);
            um_tmem_ld32_nowait(tmem_row + static_cast<uint32_t>(2 * kUmBN + c0), r2);
            um_tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e) v[c0 + e] = (((r2[e] << 8) + r1[e]) << 8) + r0[e];
        }
        return;
    }
#pragma unroll
    for (int c0 = 0; c0 < kUmBN; c0 += 32) {
#pragma unroll
        for (int e = 0; e < 32; ++e) v[c0 + e] = 0;
        for (int s = skip ? -1 : smax; s >= 0; --s) {
            uint32_t r[32];
            um_tmem_ld32(tmem_row + static_cast<uint32_t>(s * kUmBN + c0), r);
#pragma unroll
            for (int e = 0; e < 32; ++e) v[c0 + e] = (v[c0 + e] << 8) + r[e];
        }
    }
}

__device__ __forceinline__ void um_mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void um_tma_load_3d_mc(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2, uint16_t mask)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4, %5}], [%2], %6;"
                 ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "h"(mask) : "memory");
}
__device__ __forceinline__ void um_commit_mc(uint32_t bar, uint16_t mask)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask) : "memory");
}
// one lane of the (converged) warp; ptxas recognises the elect.sync predicate as "exactly one thread"
__device__ __forceinline__ bool um_elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint32_t um_cluster_rank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void um_cluster_sync()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}


// Pre-filter fused into the epilogue.  The thread holds one row x 128 columns = one segment of the compaction pass in
// registers.  An entry is kept iff dot >= need(i, j) = (floor(0.05 * (n_i + n_j)) + 1) * d (see pw_keep), and
//     need(i, j) > 0.05 * (n_i + n_j) * d * (1 - 2^-51) >= lo_i + lo_j,   lo_x = floor(0.05 * n_x * d * (1 - 2^-40)),
// so "dot < lo_i + lo_j" (one add, one compare) proves that an entry is dropped.  A row none of whose 128 entries
// passes that test is not stored at all and gets segcount 0; the others (about 1 % of the segments for sketches)
// are stored and marked kUmMaybe, and pairwise_rowsum_kernel applies the exact test to those segments only.  That
// removes the dense C round trip through HBM (4 B written + 4 B read per pair) and the full count pass.  Rows or
// columns whose norm is negative or NaN (no valid bound) are always marked.  norms_i == nullptr: plain GEMM.
constexpr uint32_t kUmMaybe = 0xFFFFFFFFu;
constexpr int kUmLoMax = (1 << 30) - 1;          // bounds are clamped (still lower bounds), their sum fits int32

struct UmFilter {
    const double* norms_i;                       // squared norms of the block's rows (block-local index)
    const double* norms_j;
    uint32_t* segcount;                          // [mi][nseg], nseg = tiles of 128 columns
    int nseg;
    int d;
};

// lower bound lo_x of the contribution of one squared norm to need(i, j); false if there is none (negative / NaN norm)
__device__ __forceinline__ bool um_lower_bound(double norm_sq, int d, int& lo)
{
    const double x = 0.05 * norm_sq * static_cast<double>(d) * (1.0 - 0x1p-40);
    lo = __double2int_rd(fmin(x, static_cast<double>(kUmLoMax)));
    return x >= 0.0;
}

// Called by all 128 epilogue threads of the CTA (warps 2..5) once per tile; lo_tile = 128 ints + 1 flag of shared memory.
__device__ __forceinline__ void um_store_row(const uint32_t (&v)[kUmBN], int64_t row, int64_t col, int tile_n, int64_t mi, int64_t mj,
                                             int32_t* __restrict__ C, int64_t ldc, const UmFilter& f, int* lo_tile)
{
    const int ncols = static_cast<int>(max(int64_t{0}, min(static_cast<int64_t>(kUmBN), mj - col)));
    if (f.norms_i) {                             // stage the tile's column bounds: every thread needs all 128 of them
        const int t = static_cast<int>(threadIdx.x) - 64;
        asm volatile("bar.sync 1, 128;" ::: "memory");                 // the previous tile's readers are done
        if (t == 0) lo_tile[kUmBN] = 0;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        int lo = kUmLoMax;                                           // columns past the end: never "maybe" (their dot is 0)
        if (t < ncols && !um_lower_bound(f.norms_j[col + t], f.d, lo)) lo_tile[kUmBN] = 1;
        lo_tile[t] = lo;
        asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    if (row >= mi || ncols == 0) return;
    if (f.norms_i) {
        int lo_i;
        bool maybe = !um_lower_bound(f.norms_i[row], f.d, lo_i) || lo_tile[kUmBN] != 0;
#pragma unroll
        for (int e = 0; e < kUmBN; ++e) maybe |= static_cast<int32_t>(v[e]) >= lo_i + lo_tile[e];
        f.segcount[row * f.nseg + tile_n] = maybe ? kUmMaybe : 0u;
        if (!maybe) return;
    }
    int32_t* dst = C + row * ldc + col;
    if (ncols == kUmBN) {                        // ldc % 4 == 0, C 16-byte aligned (host)
#pragma unroll
        for (int e = 0; e < kUmBN; e += 4) *reinterpret_cast<int4*>(dst + e) = make_int4(v[e], v[e + 1], v[e + 2], v[e + 3]);
    } else {
#pragma unroll
        for (int e = 0; e < kUmBN; ++e)
            if (e < ncols) dst[e] = static_cast<int32_t>(v[e]);
    }
}

#ifdef MVS_UMMA_PROBE
__device__ unsigned long long um_probe_cycles[1024][8];                // per CTA: MMA thread waits (tmem_empty, full), total, tiles; epilogue wait, drain, rest
#define UM_PROBE_T(var) const long long var = clock64()
#else
#define UM_PROBE_T(var)
#endif

// Persistent clusters of CM x CN CTAs (1-D grid, cluster = CM * CN consecutive CTAs); a cluster owns a super-tile of
// (CM * 128) x (CN * 128) outputs, CTA (rm, rn) of it the 128 x 128 tile (rm, rn).  The kernel is bound by L2 -> SM
// operand traffic, so the CTAs of a cluster row share their A tile and those of a cluster column their B tile: each
// CTA fetches 1/CN of the A rows and 1/CM of the B rows of a stage and TMA-multicasts them to its peers (2 x 2:
// half the traffic).  A stage may be refilled once every CTA that receives data from this producer has consumed it,
// so the MMA commit of a CTA arrives (multicast) on the empty barriers of its row and column peers.
// The three roles run decoupled across tiles: the producer keeps the stage ring full while the epilogue of the
// previous tile drains TMEM, the MMA issuer restarts as soon as the epilogue has COPIED the accumulators to registers
// (tmem_empty), and the global stores of a tile overlap the main loop of the next one.
// A rows start at plane row a_row0 (any row); tiles past the edge read zeros (TMA out-of-bounds fill) and store nothing.
// F22 = the normal case (two limbs each side, FUSED: a stage = planes A0 A1 B0 B1 of a k-slice) with the MMA loop fully
// unrolled: 16 MMAs per k-slice whose descriptors differ from the stage base by constants.
template <int CM, int CN, bool F22>
__global__ void __launch_bounds__(kUmThreads, 1)
pairwise_dot_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, UmmaPlan plan, int k_pad,
                         int a_row0, int64_t mi, int64_t mj, int32_t* __restrict__ C, int64_t ldc, int super_n, int nsuper, UmFilter filt)
{
    constexpr int CS = CM * CN;
    constexpr int kARows = kUmBM / CN, kBRows = kUmBN / CM;            // rows of a tile this CTA fetches
    extern __shared__ uint8_t um_smem_raw[];
    __shared__ int um_lo_tile[kUmBN + 1];                              // column bounds of the epilogue's current tile + "no bound" flag
    const uint32_t base = (um_smem(um_smem_raw) + 1023u) & ~1023u;     // SW128 tiles need 1024-byte alignment
    uint8_t* gen = um_smem_raw + (base - um_smem(um_smem_raw));
    const int stage_bytes = (plan.a_count + plan.b_count) * kUmTile;
    const int stages = plan.stages;
    const uint32_t bars = base + stages * stage_bytes;                 // full[stages], empty[stages], tmem_full, tmem_empty
    const uint32_t bar_tmem_full = bars + 2 * stages * 8, bar_tmem_empty = bar_tmem_full + 8;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen + stages * stage_bytes + (2 * stages + 2) * 8);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp index provably uniform
    const int rank = CS > 1 ? static_cast<int>(um_cluster_rank()) : 0;
    const int rm = rank / CN, rn = rank % CN;
    uint16_t row_mask = 0, col_mask = 0;                               // peers sharing my A tile / my B tile (self included)
    for (int j = 0; j < CN; ++j) row_mask |= static_cast<uint16_t>(1u << (rm * CN + j));
    for (int i = 0; i < CM; ++i) col_mask |= static_cast<uint16_t>(1u << (i * CN + rn));
    const int cluster_id = blockIdx.x / CS, nclusters = gridDim.x / CS;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) { um_mbar_init(bars + s * 8, 1); um_mbar_init(bars + (stages + s) * 8, CM + CN - 1); }
        um_mbar_init(bar_tmem_full, 1);
        um_mbar_init(bar_tmem_empty, 4);                               // one arrival per epilogue warp
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(um_smem(tmem_slot)), "r"(kUmTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (CS > 1) um_cluster_sync();                                     // peers' barriers are initialised before any remote arrival
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    const int nkb = k_pad / kUmBKB;
    // Producer and MMA warps run their loops with ALL lanes and warp-uniform state; only the TMA / MMA / commit
    // instructions themselves are predicated on lane 0.  Their operands then live in uniform registers: with the loops
    // inside "if (lane == 0)" the compiler has to move every descriptor into uniform registers through an
    // ELECT / R2UR loop per instruction, and the single issuing thread (131 clk per MMA) -- not the tensor pipe
    // (67 clk) -- bounds the kernel.
    if (warp == 0) {                                                   // ---- TMA producer ----
        int stage = 0;
        uint32_t phase = 0;
        for (int st_id = cluster_id; st_id < nsuper; st_id += nclusters) {
            const int sm_ = st_id / super_n, sn_ = st_id - sm_ * super_n;
            const int tile_m = sm_ * CM + rm, tile_n = sn_ * CN + rn;
            const int ra = a_row0 + tile_m * kUmBM + rn * kARows, rb = tile_n * kUmBN + rm * kBRows;
            for (int g = 0; g < plan.ngroups; ++g) {
                const int af = plan.a_first[g], bf = plan.b_first[g];
                for (int kb = 0; kb < nkb; ++kb) {
                    um_mbar_wait(bars + (stages + stage) * 8, phase ^ 1);
                    const uint32_t full = bars + stage * 8;
                    const uint32_t st = base + stage * stage_bytes;
                    if ((plan.probe & 1) && (phase || st_id != cluster_id)) {
                        if (lane == 0) um_mbar_arrive(full);
                    } else {
                        if (lane == 0) um_mbar_expect_tx(full, static_cast<uint32_t>(stage_bytes));
                        for (int i = 0; i < plan.a_count; ++i) {
                            const uint32_t dst = st + i * kUmTile + rn * kARows * kUmBKB;
                            if (lane == 0) {
                                if (CN > 1) um_tma_load_3d_mc(dst, &tmA, full, kb * kUmBKB, ra, af + i, row_mask);
                                else um_tma_load_3d(dst, &tmA, full, kb * kUmBKB, ra, af + i);
                            }
                        }
                        for (int j = 0; j < plan.b_count; ++j) {
                            const uint32_t dst = st + (plan.a_count + j) * kUmTile + rm * kBRows * kUmBKB;
                            if (lane == 0) {
                                if (CM > 1) um_tma_load_3d_mc(dst, &tmB, full, kb * kUmBKB, rb, bf + j, col_mask);
                                else um_tma_load_3d(dst, &tmB, full, kb * kUmBKB, rb, bf + j);
                            }
                        }
                    }
                    __syncwarp();
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {                                            // ---- MMA issuer ----
        // instruction descriptor: D = S32 (2 << 4), A and B signed 8 bit (1 << 7, 1 << 10), both K-major, N >> 3, M >> 4
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kUmBN >> 3) << 17) |
                               (static_cast<uint32_t>(kUmBM >> 4) << 24);
        const uint16_t peers = row_mask | col_mask;
        const bool leader = um_elect_one();                            // the one lane that issues every MMA and commit of this CTA
        int stage = 0;
        uint32_t phase = 0, tile_phase = 0;
#ifdef MVS_UMMA_PROBE
        long long w_tmem = 0, w_full = 0, ntiles = 0;
#endif
        UM_PROBE_T(t_begin);
        for (int st_id = cluster_id; st_id < nsuper; st_id += nclusters, tile_phase ^= 1) {
            UM_PROBE_T(t0);
            um_mbar_wait(bar_tmem_empty, tile_phase ^ 1);              // the epilogue has copied the previous tile out of TMEM
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#ifdef MVS_UMMA_PROBE
            w_tmem += clock64() - t0;
            ++ntiles;
#endif
            if constexpr (F22) {
                for (int kb = 0; kb < nkb; ++kb) {
                    UM_PROBE_T(t1);
                    um_mbar_wait(bars + stage * 8, phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#ifdef MVS_UMMA_PROBE
                    w_full += clock64() - t1;
#endif
                    const uint64_t d0 = um_desc_sw128(base + stage * (4 * kUmTile));
                    const uint32_t acc0 = kb > 0 ? 1u : 0u;
                    if (leader) {
                        if (!(plan.probe & 8)) {
#pragma unroll
                            for (int i = 0; i < 2; ++i)
#pragma unroll
                                for (int j = 0; j < 2; ++j)
#pragma unroll
                                    for (int k = 0; k < kUmBKB / kUmK; ++k)    // weight 1 = (0,1) then (1,0): the second always accumulates
                                        um_mma_i8(tmem_base + static_cast<uint32_t>((i + j) * kUmBN),
                                                  d0 + static_cast<uint64_t>(i * (kUmTile >> 4) + k * (kUmK >> 4)),
                                                  d0 + static_cast<uint64_t>((2 + j) * (kUmTile >> 4) + k * (kUmK >> 4)), idesc,
                                                  (k || (i == 1 && j == 0)) ? 1u : acc0);
                        }
                        if (CS > 1) um_commit_mc(bars + (stages + stage) * 8, peers);
                        else um_commit(bars + (stages + stage) * 8);
                    }
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            } else {
                uint32_t started = 0;                                      // weights whose accumulator holds data (uniform)
                for (int g = 0; g < plan.ngroups; ++g) {
                    const int sf = plan.a_first[g] + plan.b_first[g];
                    for (int kb = 0; kb < nkb; ++kb) {
                        UM_PROBE_T(t1);
                        um_mbar_wait(bars + stage * 8, phase);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#ifdef MVS_UMMA_PROBE
                        w_full += clock64() - t1;
#endif
                        const uint32_t st = base + stage * stage_bytes;
                        for (int i = 0; i < plan.a_count; ++i)
                            for (int j = 0; j < plan.b_count; ++j) {
                                const int s = sf + i + j;
                                if (s > 3) continue;                       // 256^s vanishes modulo 2^32
                                const uint32_t tacc = tmem_base + static_cast<uint32_t>(s * kUmBN);
                                const uint64_t da = um_desc_sw128(st + i * kUmTile), db = um_desc_sw128(st + (plan.a_count + j) * kUmTile);
                                const uint32_t acc0 = (started >> s) & 1u;
                                started |= 1u << s;
                                if (leader && !(plan.probe & 8)) {
#pragma unroll
                                    for (int k = 0; k < kUmBKB / kUmK; ++k)    // + 32 bytes inside the swizzle atom = + 2 in the address field
                                        um_mma_i8(tacc, da + static_cast<uint64_t>(k * (kUmK >> 4)), db + static_cast<uint64_t>(k * (kUmK >> 4)), idesc, k ? 1u : acc0);
                                }
                            }
                        // frees the stage, here and in the peers whose producers write into this CTA
                        if (leader) {
                            if (CS > 1) um_commit_mc(bars + (stages + stage) * 8, peers);
                            else um_commit(bars + (stages + stage) * 8);
                        }
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
            if (leader) um_commit(bar_tmem_full);                      // all accumulators of this tile complete
        }
#ifdef MVS_UMMA_PROBE
        if (lane == 0) {
            um_probe_cycles[blockIdx.x][0] = w_tmem;
            um_probe_cycles[blockIdx.x][1] = w_full;
            um_probe_cycles[blockIdx.x][2] = clock64() - t_begin;
            um_probe_cycles[blockIdx.x][3] = ntiles;
        }
#endif
    } else {                                                           // ---- epilogue: TMEM -> Horner -> registers -> C ----
        const int q = warp & 3;                                        // TMEM lane quarter this warp may read
        uint32_t tile_phase = 0;
#ifdef MVS_UMMA_PROBE
        long long e_wait = 0, e_drain = 0, e_rest = 0;
#endif
        for (int st_id = cluster_id; st_id < nsuper; st_id += nclusters, tile_phase ^= 1) {
            const int sm_ = st_id / super_n, sn_ = st_id - sm_ * super_n;
            const int tile_m = sm_ * CM + rm, tile_n = sn_ * CN + rn;
            UM_PROBE_T(te0);
            um_mbar_wait(bar_tmem_full, tile_phase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            UM_PROBE_T(te1);
            uint32_t v[kUmBN];                                         // this thread's row of the tile: 128 registers
            um_drain_row(tmem_base + (static_cast<uint32_t>(q * 32) << 16), plan.smax, (plan.probe & 2) != 0, v);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) um_mbar_arrive(bar_tmem_empty);             // TMEM may be overwritten; the stores below overlap the next tile
            UM_PROBE_T(te2);
            const int64_t row = static_cast<int64_t>(tile_m) * kUmBM + q * 32 + lane;
            const int64_t col = static_cast<int64_t>(tile_n) * kUmBN;
            if (!(plan.probe & 4)) um_store_row(v, row, col, tile_n, mi, mj, C, ldc, filt, um_lo_tile);
#ifdef MVS_UMMA_PROBE
            e_wait += te1 - te0;
            e_drain += te2 - te1;
            e_rest += clock64() - te2;
#endif
        }
#ifdef MVS_UMMA_PROBE
        if (threadIdx.x == 64) {
            um_probe_cycles[blockIdx.x][4] = e_wait;
            um_probe_cycles[blockIdx.x][5] = e_drain;
            um_probe_cycles[blockIdx.x][6] = e_rest;
        }
#endif
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (CS > 1) um_cluster_sync();                                     // no CTA leaves while a peer may still multicast into it
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kUmTmemCols) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------------------------
// CTA-pair variant: tcgen05.mma.cta_group::2.  The two CTAs of a cluster (two SMs of one TPC) own a 256 x 128 tile:
// CTA r stages its own 128 A rows and HALF of the B rows (64) of every k-slice, the leader (rank 0) issues one
// M = 256 MMA that runs on both SMs' tensor cores, each reading its A from its own shared memory and the B halves
// from both.  Per SM that is 48 KB instead of 64 KB of operands per fused k-slice (and 6 KB instead of 8 KB of
// shared-memory reads per MMA) for the same 128 x 128 x 128 MACs per limb pair -- the resource the 1-CTA kernel is
// bound by.  Barrier protocol (same ring as above):
//   full[s]   lives in the leader; its producer arms it with the bytes of BOTH CTAs, both producers' TMA loads
//             (cp.async.bulk.tensor .cta_group::2) complete on it;
//   empty[s]  one per CTA; the leader's commit arrives on both (multicast), each producer waits on its own;
//   tmem_full one per CTA, multicast commit;  tmem_empty in the leader, 8 arrivals (4 epilogue warps x 2 CTAs).
__device__ __forceinline__ uint32_t um_mapa(uint32_t addr, uint32_t rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void um_mbar_arrive_remote(uint32_t cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void um_mbar_wait_cluster(uint32_t bar, uint32_t parity)
{
    uint32_t ok = 0;
    unsigned long long spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && ++spins > (1ull << 28)) __trap();
    } while (!ok);
}
__device__ __forceinline__ void um_tma_load_3d_pair(uint32_t dst, const CUtensorMap* tm, uint32_t leader_bar, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(tm), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void um_mma_i8_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void um_commit_pair(uint32_t bar, uint16_t mask)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask) : "memory");
}

constexpr int kUmPairBTile = (kUmBN / 2) * kUmBKB;                   // 8 KB: this CTA's half of a B plane tile

// 1-D grid of CTA pairs (cluster = 2 consecutive CTAs); pair c loops over super-tiles (256 rows x 128 columns) c,
// c + npairs, ...; super-tile t = (t / tiles_n, t % tiles_n).  plan.stages is sized by the host for the 48 KB stage.
template <bool F22>
__global__ void __launch_bounds__(kUmThreads, 1)
pairwise_dot_umma_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, UmmaPlan plan, int k_pad,
                              int a_row0, int a_row_last, int64_t mi, int64_t mj, int32_t* __restrict__ C, int64_t ldc, int tiles_n,
                              int nsuper, UmFilter filt)
{
    extern __shared__ uint8_t um_smem_raw[];
    __shared__ int um_lo_tile[kUmBN + 1];                              // column bounds of the epilogue's current tile + "no bound" flag
    const uint32_t base = (um_smem(um_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = um_smem_raw + (base - um_smem(um_smem_raw));
    const int stage_bytes = plan.a_count * kUmTile + plan.b_count * kUmPairBTile;
    const int stages = plan.stages;
    const uint32_t bars = base + stages * stage_bytes;                 // full[stages], empty[stages], tmem_full, tmem_empty
    const uint32_t bar_tmem_full = bars + 2 * stages * 8, bar_tmem_empty = bar_tmem_full + 8;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(gen + stages * stage_bytes + (2 * stages + 2) * 8);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int rank = static_cast<int>(um_cluster_rank());              // 0 = leader
    const int pair_id = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) { um_mbar_init(bars + s * 8, 1); um_mbar_init(bars + (stages + s) * 8, 1); }
        um_mbar_init(bar_tmem_full, 1);
        um_mbar_init(bar_tmem_empty, 8);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {                                                   // the same warp of both CTAs allocates / frees
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(um_smem(tmem_slot)), "r"(kUmTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    um_cluster_sync();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    const int nkb = k_pad / kUmBKB;
    if (warp == 0) {                                                   // ---- TMA producer (both CTAs; uniform loops, lane 0 issues) ----
        int stage = 0;
        uint32_t phase = 0;
        for (int st_id = pair_id; st_id < nsuper; st_id += npairs) {
            const int sm_ = st_id / tiles_n, tile_n = st_id - sm_ * tiles_n;
            // my 128 A rows (clamped: a tile wholly past the last row loads rows that are never stored)
            const int ar = min(a_row0 + (sm_ * 2 + rank) * kUmBM, a_row_last);
            const int br = tile_n * kUmBN + rank * (kUmBN / 2);
            for (int g = 0; g < plan.ngroups; ++g) {
                const int af = plan.a_first[g], bf = plan.b_first[g];
                for (int kb = 0; kb < nkb; ++kb) {
                    um_mbar_wait(bars + (stages + stage) * 8, phase ^ 1);
                    const uint32_t st = base + stage * stage_bytes;
                    if ((plan.probe & 1) && (phase || st_id != pair_id)) {
                        if (lane == 0 && rank == 0) um_mbar_arrive(bars + stage * 8);
                    } else {
                        const uint32_t full = um_mapa(bars + stage * 8, 0);
                        if (lane == 0 && rank == 0) um_mbar_expect_tx(bars + stage * 8, static_cast<uint32_t>(2 * stage_bytes));
                        for (int i = 0; i < plan.a_count; ++i)
                            if (lane == 0) um_tma_load_3d_pair(st + i * kUmTile, &tmA, full, kb * kUmBKB, ar, af + i);
                        for (int j = 0; j < plan.b_count; ++j)
                            if (lane == 0) um_tma_load_3d_pair(st + plan.a_count * kUmTile + j * kUmPairBTile, &tmB, full, kb * kUmBKB, br, bf + j);
                    }
                    __syncwarp();
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {                                               // ---- MMA issuer (leader CTA; uniform loops, lane 0 issues) ----
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kUmBN >> 3) << 17) |
                                   (static_cast<uint32_t>((2 * kUmBM) >> 4) << 24);
            const bool leader = um_elect_one();                        // the one lane that issues every MMA and commit of the pair
            int stage = 0;
            uint32_t phase = 0, tile_phase = 0;
            for (int st_id = pair_id; st_id < nsuper; st_id += npairs, tile_phase ^= 1) {
                um_mbar_wait_cluster(bar_tmem_empty, tile_phase ^ 1);  // both CTAs' epilogues have copied the previous tile
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if constexpr (F22) {                                   // stage = A0 A1 (16 KB each) B0 B1 (8 KB each), 16 MMAs unrolled
                    for (int kb = 0; kb < nkb; ++kb) {
                        um_mbar_wait(bars + stage * 8, phase);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint64_t d0 = um_desc_sw128(base + stage * (2 * kUmTile + 2 * kUmPairBTile));
                        const uint32_t acc0 = kb > 0 ? 1u : 0u;
                        if (leader) {
                            if (!(plan.probe & 8)) {
#pragma unroll
                                for (int i = 0; i < 2; ++i)
#pragma unroll
                                    for (int j = 0; j < 2; ++j)
#pragma unroll
                                        for (int k = 0; k < kUmBKB / kUmK; ++k)
                                            um_mma_i8_pair(tmem_base + static_cast<uint32_t>((i + j) * kUmBN),
                                                           d0 + static_cast<uint64_t>(i * (kUmTile >> 4) + k * (kUmK >> 4)),
                                                           d0 + static_cast<uint64_t>(2 * (kUmTile >> 4) + j * (kUmPairBTile >> 4) + k * (kUmK >> 4)), idesc,
                                                           (k || (i == 1 && j == 0)) ? 1u : acc0);
                            }
                            um_commit_pair(bars + (stages + stage) * 8, 3);
                        }
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                } else {
                    uint32_t started = 0;
                    for (int g = 0; g < plan.ngroups; ++g) {
                        const int sf = plan.a_first[g] + plan.b_first[g];
                        for (int kb = 0; kb < nkb; ++kb) {
                            um_mbar_wait(bars + stage * 8, phase);
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            const uint32_t st = base + stage * stage_bytes;
                            for (int i = 0; i < plan.a_count; ++i)
                                for (int j = 0; j < plan.b_count; ++j) {
                                    const int s = sf + i + j;
                                    if (s > 3) continue;
                                    const uint32_t tacc = tmem_base + static_cast<uint32_t>(s * kUmBN);
                                    const uint64_t da = um_desc_sw128(st + i * kUmTile);
                                    const uint64_t db = um_desc_sw128(st + plan.a_count * kUmTile + j * kUmPairBTile);
                                    const uint32_t acc0 = (started >> s) & 1u;
                                    started |= 1u << s;
                                    if (leader && !(plan.probe & 8)) {
#pragma unroll
                                        for (int k = 0; k < kUmBKB / kUmK; ++k)
                                            um_mma_i8_pair(tacc, da + static_cast<uint64_t>(k * (kUmK >> 4)), db + static_cast<uint64_t>(k * (kUmK >> 4)), idesc,
                                                           k ? 1u : acc0);
                                    }
                                }
                            if (leader) um_commit_pair(bars + (stages + stage) * 8, 3);   // frees the stage in both CTAs
                            if (++stage == stages) { stage = 0; phase ^= 1; }
                        }
                    }
                }
                if (leader) um_commit_pair(bar_tmem_full, 3);
            }
        }
    } else {                                                           // ---- epilogue (both CTAs, own 128 rows) ----
        const int q = warp & 3;
        const uint32_t leader_tmem_empty = um_mapa(bar_tmem_empty, 0);
        uint32_t tile_phase = 0;
        for (int st_id = pair_id; st_id < nsuper; st_id += npairs, tile_phase ^= 1) {
            const int sm_ = st_id / tiles_n, tile_n = st_id - sm_ * tiles_n;
            const int tile_m = sm_ * 2 + rank;
            um_mbar_wait(bar_tmem_full, tile_phase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            uint32_t v[kUmBN];
            um_drain_row(tmem_base + (static_cast<uint32_t>(q * 32) << 16), plan.smax, (plan.probe & 2) != 0, v);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) um_mbar_arrive_remote(leader_tmem_empty);
            const int64_t row = static_cast<int64_t>(tile_m) * kUmBM + q * 32 + lane;
            const int64_t col = static_cast<int64_t>(tile_n) * kUmBN;
            if (!(plan.probe & 4)) um_store_row(v, row, col, tile_n, mi, mj, C, ldc, filt, um_lo_tile);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    um_cluster_sync();                                                 // nobody leaves while the peer may still signal into it
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kUmTmemCols) : "memory");
    }
}

}  // namespace mvs
