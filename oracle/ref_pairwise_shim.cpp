// ref_pairwise_shim.cpp -- C wrapper around the UNMODIFIED reference function
//   SparseResult compute_sparse_dot_products_optimized(const MatrixXi&, const MatrixXi&, const VectorXd&,
//                                                      const VectorXd&, int)
//   (/root/reference/src/pairwise_comp_optimized.cpp:57-160).
//
// TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/mvs_oracle.c header).  The reference keeps that function
// in the translation unit of its CLI, so oracle/Makefile compiles that file WHERE IT LIES (passed as the
// macro REF_PAIRWISE_CPP) with its `main` renamed; nothing of it is copied into this repository.
#define main ref_pairwise_cli_main
#include REF_PAIRWISE_CPP
#undef main

#include <chrono>

extern "C" {

// Blocks are row-major [m x d] = the memory image of the reference's column-major (d x m) MatrixXi.
// Returns the number of pairs kept by the reference; copies at most cap of them out.
int64_t ref_pairwise_block(const int32_t* block_i, int64_t mi, const int32_t* block_j, int64_t mj, int d,
                           const double* norms_i, const double* norms_j, int32_t* rows, int32_t* cols,
                           int64_t* values, int64_t cap, double* seconds)
{
    Eigen::Map<const MatrixXi> mi_map(block_i, d, mi), mj_map(block_j, d, mj);
    const MatrixXi bi = mi_map, bj = mj_map;                  // the reference passes owning matrices
    const VectorXd ni = Eigen::Map<const VectorXd>(norms_i, mi), nj = Eigen::Map<const VectorXd>(norms_j, mj);
    const auto t0 = std::chrono::steady_clock::now();
    SparseResult r = compute_sparse_dot_products_optimized(bi, bj, ni, nj, d);
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const int64_t n = static_cast<int64_t>(r.values.size());
    for (int64_t k = 0; k < n && k < cap; ++k) {
        rows[k] = r.rows[static_cast<size_t>(k)];
        cols[k] = r.cols[static_cast<size_t>(k)];
        values[k] = r.values[static_cast<size_t>(k)];
    }
    return n;
}

}  // extern "C"
