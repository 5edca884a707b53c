/*
 * mvs_b200.h -- C ABI of the B200-native random-projection sketching path.
 *
 * This is the drop-in boundary for the hot path of
 * RolandFaure/metagenome_vector_sketches:
 *
 *   VectorXi transform_set_into_vector(const std::unordered_set<unsigned long>&, int d)
 *       reference: include/random_projection.h:11, src/random_projection.cpp:9-26
 *   the sample loop, int16 saturation and norm pass of sketch()
 *       reference: src/project_everything.cpp:287-298, 327-353
 *
 * The reference has no FFI layer for this path (its pybind module, src/bindings.cpp,
 * binds the query side only), so the entry points below are what a cgo / ctypes /
 * C++ caller of that path binds instead; INTEGRATION.md shows the C++ shim with the
 * reference's exact signature and the ctypes stub.
 *
 * Conventions
 *   - plain pointers and sizes only; the caller owns every buffer;
 *   - every function returns 0 (MVS_OK) or a negative mvs_status; mvs_strerror()
 *     gives the text, mvs_last_cuda_error() the underlying cudaError_t;
 *   - a batch is CSR: sample s owns hashes[offsets[s] .. offsets[s+1]); offsets has
 *     n_samples+1 entries, offsets[0] is the index of the first hash of sample 0;
 *   - hashes inside one sample must be UNIQUE (the reference receives a set; a
 *     duplicate would be counted twice).  mvs_sort_unique() is the helper the
 *     drivers use;
 *   - `seed` is additive: entry(h, j) uses splitmix64(h + 64*(j/64) + seed);
 *     seed 0 is bit-identical to the reference (which has no seed);
 *   - a context is bound to ONE GPU (one process / host thread per GPU); it is not
 *     thread-safe, use one context per host thread;
 *   - there is NO CPU fallback: without a usable sm_100 device mvs_create() fails.
 */
#ifndef MVS_B200_H
#define MVS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mvs_ctx mvs_ctx;

typedef enum mvs_status {
    MVS_OK = 0,
    MVS_ERR_INVALID_ARG = -1,   /* null pointer, d <= 0, negative sizes, offsets not monotone */
    MVS_ERR_NO_DEVICE = -2,     /* no CUDA device / device is not sm_100 */
    MVS_ERR_CUDA = -3,          /* a CUDA call failed: see mvs_last_cuda_error() */
    MVS_ERR_OOM = -4,           /* host or device allocation failed */
    MVS_ERR_TOO_LARGE = -5,     /* a sample has >= 2^31 hashes (int32 accumulator, as the reference) */
    MVS_ERR_IO = -6,            /* file could not be opened / written (driver helpers) */
    MVS_ERR_SCRATCH_LIMIT = -7  /* mvs_project_device_static: the call needs more than MVS_STATIC_SCRATCH_BYTES */
} mvs_status;

/* ---- lifetime --------------------------------------------------------------------- */
int mvs_create(mvs_ctx** ctx, int device);          /* binds to cuda:<device> */
void mvs_destroy(mvs_ctx* ctx);
const char* mvs_strerror(int status);
int mvs_last_cuda_error(const mvs_ctx* ctx);        /* cudaError_t of the last failure, 0 if none */
int mvs_abi_version(void);                          /* bumps when a signature changes */

/* ---- the hot path ----------------------------------------------------------------- */

/* One set -> one vector.  Replaces transform_set_into_vector
 * (reference src/random_projection.cpp:9).  hashes/out are HOST pointers. */
int mvs_transform_set(mvs_ctx* ctx, const uint64_t* hashes, int64_t n_hashes, int d,
                      uint64_t seed, int32_t* out);

/* A CSR batch of sets -> n_samples x d int32 row-major.  Replaces the OpenMP sample
 * loop of sketch() (reference src/project_everything.cpp:289-298).  All pointers are
 * HOST memory (pinned memory from mvs_host_alloc / mvs_host_register makes the copies
 * asynchronous); the call streams slices of the batch through the GPU on several CUDA
 * streams (H2D, kernel, D2H overlapped) and returns when `out` is complete. */
int mvs_project_batch(mvs_ctx* ctx, const uint64_t* hashes, const uint64_t* offsets,
                      int64_t n_samples, int d, uint64_t seed, int32_t* out);

/* mvs_project_batch fused with the epilogue of sketch() (reference
 * src/project_everything.cpp:327-347): any of the four HOST outputs may be NULL.
 *   out        n_samples*d int32   (vectors.bin, int32 mode)
 *   out16      n_samples*d int16   (vectors.bin, --int16 mode: saturated to [-32768, 32767])
 *   sumsq      n_samples uint64    (exact sum of squares; sqrt(sumsq / d) is the mathematically exact norm)
 *   ref_norms  n_samples float     (the norm AS THE REFERENCE COMPUTES IT, :328-329: float32, Eigen 3.2 SSE
 *                                   reduction order, FMA, 1/sqrt(d) from the host CPU's rsqrtss -- widened to
 *                                   double and printed with %g it is the reference's vector_norms.txt column
 *                                   on the same host; computed from the int32 values before saturation) */
int mvs_sketch_batch(mvs_ctx* ctx, const uint64_t* hashes, const uint64_t* offsets,
                     int64_t n_samples, int d, uint64_t seed, int32_t* out, int16_t* out16,
                     uint64_t* sumsq, float* ref_norms);

/* The same with the CSR payload PACKED: hash k occupies bytes [7k, 7k+7) of `packed`, little endian (offsets still count
 * hashes).  Every hash must be below 2^56 -- FracMinHash hashes are below 2^64 / scaled, i.e. below 2^56 for any
 * scaled >= 256 (sourmash's default 1000 gives 2^54.03).  The host link carries 7 instead of 8 bytes per hash and the
 * device widens them (unpack56_kernel): the end-to-end rate of a batch call is bounded by that link, so it rises by
 * the same 12.5 %.  mvs_pack56 is the reference packer (MVS_ERR_INVALID_ARG if a hash does not fit; `packed` holds 7*n
 * bytes).  The u64 entry points are untouched. */
int mvs_sketch_batch_packed56(mvs_ctx* ctx, const unsigned char* packed, const uint64_t* offsets,
                              int64_t n_samples, int d, uint64_t seed, int32_t* out, int16_t* out16,
                              uint64_t* sumsq, float* ref_norms);
int mvs_pack56(const uint64_t* hashes, uint64_t n, unsigned char* packed);

/* The same over several GPUs in ONE call (SURVEY 8b "one call may use all GPUs"; the reference's analogue is
 * the dynamic OpenMP schedule of src/project_everything.cpp:289): the batch is cut into n_ctx contiguous
 * size-balanced shards (mvs_partition_samples), each context projects its shard on its own host thread and
 * writes its own row range of the caller's buffers.  No inter-GPU exchange.  Contexts may share a device. */
int mvs_sketch_batch_multi(mvs_ctx* const* ctxs, int n_ctx, const uint64_t* hashes, const uint64_t* offsets,
                           int64_t n_samples, int d, uint64_t seed, int32_t* out, int16_t* out16,
                           uint64_t* sumsq, float* ref_norms);

/* Same computation with the hash array and the output already resident in HBM.
 * d_hashes / d_out are DEVICE pointers on the context's GPU, h_offsets is a HOST array
 * (n_samples+1 entries indexing d_hashes).  `stream` is a cudaStream_t (NULL = the
 * context's own stream).  Asynchronous w.r.t. the host unless `stream` is NULL.      */
int mvs_project_device(mvs_ctx* ctx, const uint64_t* d_hashes, const uint64_t* h_offsets,
                       int64_t n_samples, int d, uint64_t seed, int32_t* d_out, void* stream);

/* mvs_project_device WITHOUT ANY ALLOCATION.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments, same result (bit-identical rows) as mvs_project_device, but the
 * launch's scratch -- warp tasks, reduce items and the partial rows of samples that are split over several warps --
 * lives in a file-scope `static __device__` array of MVS_STATIC_SCRATCH_BYTES inside the library, which every GPU
 * that loads the library holds until the process exits.  The call performs no cudaMalloc / cudaFree / cudaMallocHost
 * and calls no allocating library routine; the plan is built in a static host image and copied with one
 * cudaMemcpyAsync from pageable memory (the host therefore waits for earlier work on `stream` before the call returns).
 *   limit    the plan must fit the array: 32 B per warp task (>= one per sample and 32-block column pass),
 *            16 B per split sample, d * 4 B per partial row.  `*scratch_needed` (may be NULL) receives the bytes
 *            this call needs; beyond MVS_STATIC_SCRATCH_BYTES the call does nothing and returns
 *            MVS_ERR_SCRATCH_LIMIT -- use mvs_project_device for such batches.
 *   sharing  ALL calls on one GPU (any context, any stream) use the same array.  Calls on one stream are ordered by
 *            the stream; calls whose kernels may overlap on the GPU (different streams, different contexts or host
 *            threads on one device) RACE on it and produce garbage: serialise them.  The array keeps the last
 *            call's contents; nothing in it has to start at zero (every partial row is stored in full before it is
 *            read), so nothing is cleared.
 *   first call per device resolves the array's address (cudaGetSymbolAddress) and may load the CUDA module: issue
 *   one call before entering a region that forbids it. */
#define MVS_STATIC_SCRATCH_BYTES (16u << 20)
int mvs_project_device_static(mvs_ctx* ctx, const uint64_t* d_hashes, const uint64_t* h_offsets,
                              int64_t n_samples, int d, uint64_t seed, int32_t* d_out, void* stream,
                              uint64_t* scratch_needed);

/* Epilogue of sketch() on device-resident vectors (reference
 * src/project_everything.cpp:327-347): per row the exact integer sum of squares, the reference's
 * float32 norm (see mvs_sketch_batch) and the int16-saturated copy; each output may be NULL.
 * d_vec, d_sumsq (n_samples uint64), d_out16 (n_samples*d int16) and d_ref_norms (n_samples float)
 * are DEVICE pointers. */
int mvs_finalize_device(mvs_ctx* ctx, const int32_t* d_vec, int64_t n_samples, int d,
                        uint64_t* d_sumsq, int16_t* d_out16, float* d_ref_norms, void* stream);

/* Host-pointer convenience wrapper of the above (copies in and out). */
int mvs_finalize(mvs_ctx* ctx, const int32_t* vec, int64_t n_samples, int d,
                 uint64_t* sumsq, int16_t* out16, float* ref_norms);

/* The reference's norm on the HOST (needs no GPU): the same float32 sequence as the device epilogue, for
 * callers that hold int32 rows only.  mvs_reference_inv_sqrt_d is the 1/sqrt(d) factor the reference binary
 * uses (rsqrtss + one Newton-Raphson step; the approximation instruction is the executing CPU's). */
int mvs_reference_norms(const int32_t* vec, int64_t n_samples, int d, float* norms);
float mvs_reference_inv_sqrt_d(int d);

/* ---- drivers: the reference's two CLIs as library calls ------------------------------ */
typedef struct mvs_hash_batch {
    int64_t n_samples;
    uint64_t n_hashes;
    uint64_t* hashes;        /* CSR payload, sorted + unique inside each sample (pinned if possible) */
    uint64_t* offsets;       /* n_samples + 1 */
    char** names;            /* n_samples C strings (text before the first ':'); NULL if !named */
    int hashes_pinned;
} mvs_hash_batch;

/* Parser of sketch() (named = 1: "<name>: h1 h2 ...", lines without ':' skipped; reference
 * src/project_everything.cpp:264-282) and of standalone_projection (named = 0: one set per
 * line, empty line = empty set; reference src/standalone_projection.cpp:27-34).  Token
 * semantics are those of `istringstream >> unsigned long`.  Needs no GPU. */
int mvs_parse_hash_file(const char* path, int named, mvs_hash_batch** out);
void mvs_free_hash_batch(mvs_hash_batch* b);

/* Binary form of a parsed hash file (the optional ".u64 sidecar" of SURVEY 8f rank 1; the reference has only the text
 * form, src/project_everything.cpp:257-284): 64-byte header (magic "MVSU64\0\1", n_samples, n_hashes, names_bytes,
 * payload_offset), u64 offsets[n_samples + 1], the names as NUL-terminated strings, zero padding to a multiple of
 * 4096, then the CSR payload as u64 -- the batch exactly as mvs_parse_hash_file returned it.  mvs_parse_hash_file
 * (and with it mvs_sketch_file, mvs_standalone_file and the CLIs) recognises the magic and loads such a file instead
 * of parsing text: the same batch, hence the same db, without the parse.  A file whose sizes or offsets are
 * inconsistent is refused with MVS_ERR_IO; a file without names where names are needed with MVS_ERR_INVALID_ARG.
 * CLI: `project_everything pack <hash_file> <binary_file>`.  Needs no GPU. */
int mvs_write_hash_batch(const mvs_hash_batch* b, const char* path);

/* db/ writer of sketch() (reference src/project_everything.cpp:306-361): vectors.bin (int32
 * rows, or int16 if vec16 != NULL), vector_norms.txt ("<name> <norm>", %g), dimension.txt, dtype.txt.
 * The norm column is ref_norms (the reference's float32 value, file identical to the reference's) when
 * given, else sqrt(sumsq/d) when sumsq is given, else mvs_reference_norms(vec32).  Needs no GPU. */
int mvs_write_db(const char* index_folder, int64_t n_samples, int d, const char* const* names,
                 const float* ref_norms, const uint64_t* sumsq, const int32_t* vec32, const int16_t* vec16);

/* Which norm mvs_sketch_file / the CLI print (process-wide): 0 = the reference's float32 norm (default;
 * vector_norms.txt byte-identical to the reference's on the same host), 1 = sqrt(exact sum of squares / d). */
int mvs_set_norms_mode(int mode);
int mvs_get_norms_mode(void);

/* Contiguous size-balanced static partition of a CSR batch over n_parts GPUs; writes
 * n_parts + 1 sample bounds.  No collective is needed afterwards: samples are independent. */
void mvs_partition_samples(const uint64_t* offsets, int64_t n_samples, int n_parts, int64_t* bounds);

/* sketch() end to end (reference src/project_everything.cpp:238-362): wipes/creates
 * index_folder, parses, projects on n_ctx GPUs (one context + host thread each), writes db/. */
int mvs_sketch_file(mvs_ctx* const* ctxs, int n_ctx, const char* hash_file, const char* index_folder,
                    int d, int use_int16, uint64_t seed, int verbose);

/* The CLI's entry: mvs_sketch_file with the contexts of devices[0..n_dev) created on a second host
 * thread while the hash file is parsed (CUDA start-up overlaps the parse); the contexts are
 * destroyed before returning.  Same results, same printed lines. */
int mvs_sketch_file_on_devices(const int* devices, int n_dev, const char* hash_file, const char* index_folder,
                               int d, int use_int16, uint64_t seed, int verbose);

/* standalone_projection main (reference src/standalone_projection.cpp:11-46): one vector per
 * input line printed as floats separated by ' '.  out_path NULL = stdout. */
int mvs_standalone_file(mvs_ctx* ctx, const char* hashes_file, int d, uint64_t seed, const char* out_path);

/* `project_everything convert` (reference src/project_everything.cpp:181-235) in process:
 * every *.sig.zip of the folder -> one "<name>: h1 h2 ..." line (k=31 `mins`), files in sorted
 * order, zip/gzip read with zlib instead of system("unzip"/"gunzip").  Needs no GPU. */
int mvs_convert_signatures(const char* signature_folder, const char* hash_file, int threads, int verbose);

/* ---- helpers ---------------------------------------------------------------------- */
/* In-place sort + unique of one sample's hashes (set semantics of the reference's
 * unordered_set).  Returns the new count (>= 0). */
int64_t mvs_sort_unique(uint64_t* hashes, int64_t n);

void* mvs_host_alloc(size_t bytes);                 /* pinned host memory, NULL on failure */
/* The same, write-combined (cudaHostAllocWriteCombined): not snooped by the CPU caches, so the device's DMA reads need
 * no cache probes on the host side.  For INPUT buffers the host only ever writes (sequentially); host reads of
 * write-combined memory are very slow.  Freed with mvs_host_free. */
void* mvs_host_alloc_write_combined(size_t bytes);
void mvs_host_free(void* p);
int mvs_host_register(void* p, size_t bytes);       /* pin caller-owned memory */
int mvs_host_unregister(void* p);

/* Host locality (no counterpart in the reference, which has no device): binds the CALLING host thread to the
 * NUMA node the device hangs off -- CPU affinity = the device's local CPUs that the thread may already run on
 * (/sys/bus/pci/devices/<bus id>/local_cpulist; skipped when fewer than 4 remain), memory policy = prefer that node (/sys/.../numa_node) -- so that
 * pinned buffers the thread allocates afterwards (mvs_host_alloc, the library's own staging) are read by the
 * device's DMA engine without crossing the socket interconnect.  On a multi-GPU box every rank / per-GPU thread
 * calls it once before it allocates.  *numa_node (may be NULL) receives the node or -1 when the platform reports
 * none; *applied (may be NULL) receives a bit mask: 1 = CPU affinity set, 2 = memory policy set.  A platform
 * without the sysfs entries, or a container that forbids the node, is not an error (MVS_OK, *applied = 0). */
int mvs_bind_host_to_device(int device, int* numa_node, int* applied);

/* ---- all-pairs comparison of projected vectors --------------------
 * Replaces compute_sparse_dot_products_optimized (reference src/pairwise_comp_optimized.cpp:57-160)
 * and the block loops of its main (:949-983).  dot(i, j) = sum_k a[i][k] * b[j][k] in int32
 * arithmetic (modulo 2^32, as the reference's Eigen MatrixXi product); a pair is kept iff
 * (int64)dot / d > 0.05 * (norms_i[i] + norms_j[j]) with truncating integer division (:140-143);
 * norms are SQUARED norms as the reference holds them (:895-901).  Kept pairs are returned sorted
 * by (row, col) -- the order of the reference's double loop -- in a library-allocated result. */
typedef struct mvs_sparse_result {
    int64_t n;             /* kept pairs */
    int32_t* rows;         /* n row indices (block-relative for _block, global for _shard) */
    int32_t* cols;         /* n column indices */
    int64_t* values;       /* n dot products (the wrapped int32 value, widened) */
    double gemm_ms;        /* CUDA-event time of the dot-product kernel(s) */
    double filter_ms;      /* CUDA-event time of the exact count of the pre-filtered segments + scan + emit
                              (the pre-filter itself runs in the GEMM epilogue) */
    double prep_ms;        /* host wall time of the operand preparation (min/max + int8 limb planes) */
    double wall_ms;        /* host wall clock of the call, copies included */
    uint64_t h2d_bytes, d2h_bytes;
} mvs_sparse_result;

/* One block pair, host buffers in row-major [m x d] (the memory image of the reference's d x m
 * column-major MatrixXi blocks).  Same argument meaning as the reference function. */
int mvs_pairwise_block(mvs_ctx* ctx, const int32_t* block_i, int64_t mi, const int32_t* block_j, int64_t mj,
                       int d, const double* norms_i, const double* norms_j, mvs_sparse_result** out);

/* Rows [begin_row, end_row) of a db against all n_total vectors (one shard of the reference's main,
 * :936-983): vectors = the int32 image of vectors.bin, norms_sq[n_total].  The db is uploaded once
 * and stays resident in HBM; indices in the result are global. */
int mvs_pairwise_shard(mvs_ctx* ctx, const int32_t* vectors, int64_t n_total, int d, const double* norms_sq,
                       int64_t begin_row, int64_t end_row, mvs_sparse_result** out);

void mvs_free_sparse_result(mvs_sparse_result* r);

/* ---- measurement ------------------------------------------------------------------ */
typedef struct mvs_timing {
    float kernel_ms;       /* mean CUDA-event time (on the launching stream) of the projection +
                              partial-reduce kernels over the mvs_project_device calls made since
                              the previous mvs_get_timing (at most the last 64) */
    float reduce_ms;       /* the partial-reduce kernel alone, same averaging */
    uint32_t calls;        /* number of calls averaged */
    uint32_t reserved;
    uint64_t launches;     /* kernels launched by this context since creation */
    uint64_t warp_tasks;   /* warp tasks of the last projection call */
    uint64_t partial_rows; /* partial rows of the last projection call */
    uint64_t h2d_bytes;    /* bytes copied host->device by the last mvs_project_batch */
    uint64_t d2h_bytes;    /* bytes copied device->host by the last mvs_project_batch */
} mvs_timing;
int mvs_get_timing(mvs_ctx* ctx, mvs_timing* t);    /* synchronises the context's streams */

/* Integer issue-rate microbenchmark used as the roofline denominator.
 * which: 0 = LOP3 (ALU pipe), 1 = IMAD (FMA pipe), 2 = LOP3+IMAD interleaved 1:1,
 *        3 = IADD3, 4 = SHF, 5 = IMAD.WIDE, 6 = IMAD.HI, 7 = SHF funnel, 8 = LOP3:IMAD 2:1,
 *        9 = LOP3:SHF 1:1.  Writes 32-bit lane-instructions per second (mode 5 issues one
 *        IMAD.WIDE + one LOP3 per counted op). */
int mvs_int_issue_peak(mvs_ctx* ctx, int which, double* ops_per_second);

/* The planner of the projection without a device (needs no GPU): how a CSR batch would be cut into warp tasks --
 * (sample, hash range, 32-block column pass) -- on a GPU with `sm_count` SMs at dimension d.  Writes the chosen cap
 * (hashes per task), the number of tasks and of partial rows; if the four task arrays are given (room for max_tasks
 * entries) also the task list (begin = index into the batch's hash array).  For tests and capacity planning. */
int mvs_plan_tasks(const uint64_t* offsets, int64_t n_samples, int d, int sm_count, uint64_t* cap_hashes,
                   uint64_t* n_tasks, uint64_t* n_partial_rows, uint64_t* task_begin, uint32_t* task_count,
                   uint32_t* task_sample, uint32_t* task_pass, uint64_t max_tasks);

/* Second roofline denominator: warp rounds per second (one round = one hash x 32 lanes x 64 columns) that the GPU
 * sustains for the projection loop's own instruction mix when no instruction waits for an operand (loop_mix_kernel:
 * the same opcodes in the same proportions on independent register sets, same residency).  The projection kernel
 * cannot run faster than this without issuing fewer instructions. */
int mvs_loop_mix_peak(mvs_ctx* ctx, double* rounds_per_second);

/* Tuning knobs.  Returns MVS_ERR_INVALID_ARG for an unknown key or a value outside the key's range (nothing is
 * ever ignored silently).  The shipped library knows:
 *   "slice_hashes" (> 0): hashes per streamed slice of mvs_project_batch / mvs_sketch_batch (default 16 Mi = 128 MB);
 *   "slice_rows_bytes" (> 0): output bytes per streamed slice (default 256 MB);
 *   "stage_threads" (1..64): host threads that copy a pageable caller buffer into the pinned staging ring (default 8);
 *   "relay_device" (-1 = off, default; else a peer device index with P2P access): the hash slices of the batch entry
 *       points travel host -> peer device -> this device (cudaMemcpyAsync into a staging buffer on the peer, then
 *       cudaMemcpyPeerAsync over NVLink) instead of over this device's own PCIe link.  For hosts whose PCIe roots have
 *       unequal host-memory ceilings: ranks on the starved root route through a partner on the other one (bench.py
 *       probes the ranks' concurrent H2D rates and pairs them).  Pinned sources only; pageable
 *       ones keep the direct path;
 *   "relay_percent" (0..100, default 100): share of the slice bytes that takes the relay, the rest keeps this device's
 *       own link (slices alternate by error diffusion) -- when both PCIe roots still add bandwidth to each other the
 *       balanced split is 1 - 2s/(f+s) of a slow rank's bytes for concurrent rates s (slow) and f (fast);
 *   "release_workspace" (any value): frees the device and pinned buffers the pairwise entry points keep between calls
 *       (they are grow-only and otherwise live until mvs_destroy).
 * The A/B knobs of round 1 ("task_cap", "pairwise_path", "pairwise_limbs", "pairwise_pair", "pairwise_cluster",
 * "pairwise_stream", "pairwise_fuse", "stage_copy", "stage_uploads") select code that only exists in the experiment
 * build (make -C csrc experiments -> libmvs_b200_exp.so, compiled with -DMVS_EXPERIMENTS; used by tools/ and
 * tests/test_gpu_experiments.py); the shipped library has one GEMM per shape class (tcgen05 CTA pair / single CTA),
 * one upload path, and rejects those keys.  mvs_has_experiments() tells which build is loaded. */
int mvs_set_option(mvs_ctx* ctx, const char* key, int64_t value);
int mvs_has_experiments(void);

#ifdef __cplusplus
}
#endif
#endif /* MVS_B200_H */
