// mvs_internal.h -- what the other translation units of libmvs_b200.so may use of a context
// (the struct itself stays private to mvs_capi.cu).  Not part of the C ABI.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

struct mvs_ctx;

int mvs_internal_device(const mvs_ctx* ctx);
cudaStream_t mvs_internal_stream(mvs_ctx* ctx);                  // the stream of slot 0
void mvs_internal_set_cuda_error(mvs_ctx* ctx, cudaError_t e);
void mvs_internal_count_launches(mvs_ctx* ctx, uint64_t n);
int mvs_internal_pairwise_path(const mvs_ctx* ctx);               // option "pairwise_path"
int mvs_internal_pairwise_limbs(const mvs_ctx* ctx);              // option "pairwise_limbs"
int mvs_internal_pairwise_stream(const mvs_ctx* ctx);             // option "pairwise_stream"
int mvs_internal_pairwise_cluster(const mvs_ctx* ctx);            // option "pairwise_cluster"
int mvs_internal_pairwise_pair(const mvs_ctx* ctx);               // option "pairwise_pair"
int mvs_internal_pairwise_fuse(const mvs_ctx* ctx);               // option "pairwise_fuse"
// grow-only device buffer number `slot` (0..15) of at least `bytes` bytes, kept until mvs_destroy / option "release_workspace"
int mvs_internal_workspace(mvs_ctx* ctx, int slot, size_t bytes, void** out);
void* mvs_internal_pinned_scratch(mvs_ctx* ctx);                 // 256 bytes of pinned host memory (nullptr on failure)
void* mvs_internal_pinned_stage(mvs_ctx* ctx, size_t bytes);      // persistent pinned staging buffer (nullptr on failure)
// cudaMemcpyAsync(host -> device) that takes pageable sources through a pinned ring filled by several host threads
int mvs_internal_upload(mvs_ctx* ctx, cudaStream_t st, void* dst, const void* src, size_t bytes);
// memcpy whose stores bypass the caches (AVX2 streaming stores; plain memcpy on hosts without AVX2): for filling the
// pinned staging ring, whose lines the host never reads again.  mvs_driver.cpp.
void mvs_internal_copy_streaming(void* dst, const void* src, size_t bytes);
