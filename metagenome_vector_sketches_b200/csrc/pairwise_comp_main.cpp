// pairwise_comp_b200 -- drop-in for the reference CLI `pairwise_comp_optimized` (int32 dbs), SURVEY 8f rank 4.
//
//   pairwise_comp_b200 --db <db>/ --max_memory_gb G --num_threads T --output_folder <out>/ --num_shards S --shard_idx K
//                      [--start_shard a] [--end_shard b] [--gpu N]
//   (same required options as the reference, src/pairwise_comp_optimized.cpp:833-849; --gpu is additive)
//
// The N^2 * d part -- the block products and the keep-test (reference :57-160, :949-983) -- runs on the GPU through
// mvs_pairwise_shard.  The succinct shard files (reference :645-817) are written here on the host from the kept
// pairs: std::unordered_map for the per-row grouping like the reference (its iteration order decides the row order
// inside matrix.bin) and the in-tree writers of mvs_succinct.h for the two containers of the third-party `bits`
// library the reference vendors (compact_vector, rice_sequence; pinned byte for byte against those headers by
// tests/test_succinct.py).  Feeding the map in the reference's order -- block pair by block pair, rows then columns
// inside a block, with the reference's chunk size max_bytes / (4d)^2 -- makes shard_K/{matrix.bin,row_index.bin,
// neighbor_start.bin} byte-identical to the reference's, which tests/test_gpu_cli.py checks.
// --max_memory_gb only fixes that order; the GPU tiling is independent of it.  --num_threads is accepted and unused.
#include "../../include/mvs_b200.h"

#include "mvs_succinct.h"

#include <algorithm>
#include <chrono>
#include <thread>
#include <new>
#include <memory>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <numeric>
#include <string>
#include <unordered_map>
#include <vector>

namespace fs = std::filesystem;

namespace {

struct Args {
    std::string db, out;
    double max_memory_gb = 0.0;
    int num_threads = 1, num_shards = 1, shard_idx = 0, gpu = 0;
    bool have[6] = {false, false, false, false, false, false};
    bool help = false;
};

void usage(const char* argv0)
{
    std::cout << "Usage:\n" << argv0
              << " --db <folder> --max_memory_gb <float> --num_threads <int> --output_folder <folder> --num_shards <int>"
                 " --shard_idx <int> [--start_shard <int>] [--end_shard <int>] [--gpu <int>] [--help]" << std::endl;
}

bool parse(int argc, char** argv, Args& a)
{
    for (int i = 1; i < argc; ++i) {
        const std::string k = argv[i];
        if (k == "--help") { a.help = true; continue; }
        if (i + 1 >= argc) return false;
        const char* v = argv[++i];
        char* end = nullptr;
        if (k == "--db") { a.db = v; a.have[0] = true; }
        else if (k == "--max_memory_gb") { a.max_memory_gb = std::strtod(v, &end); a.have[1] = end && !*end; }
        else if (k == "--num_threads") { a.num_threads = static_cast<int>(std::strtol(v, &end, 10)); a.have[2] = end && !*end; }
        else if (k == "--output_folder") { a.out = v; a.have[3] = true; }
        else if (k == "--num_shards") { a.num_shards = static_cast<int>(std::strtol(v, &end, 10)); a.have[4] = end && !*end; }
        else if (k == "--shard_idx") { a.shard_idx = static_cast<int>(std::strtol(v, &end, 10)); a.have[5] = end && !*end; }
        else if (k == "--start_shard" || k == "--end_shard") { std::strtol(v, &end, 10); if (!end || *end) return false; }   // parsed, unused (as in the reference)
        else if (k == "--gpu") { a.gpu = static_cast<int>(std::strtol(v, &end, 10)); if (!end || *end) return false; }
        else return false;
    }
    for (bool h : a.have)
        if (!h) return false;
    return a.num_shards > 0 && a.shard_idx >= 0;
}

struct Pair {
    int32_t row, col;
    int64_t value;
};

// Jaccard estimate of a kept pair, quantised to 8 bits (reference :654-665)
uint16_t quantize(const Pair& p, const std::vector<double>& norms_sq, int dimension)
{
    const double inter = static_cast<double>(p.value) / dimension;
    double jaccard = inter / (norms_sq[static_cast<size_t>(p.row)] + norms_sq[static_cast<size_t>(p.col)] - inter);
    if (jaccard > 1) jaccard = 1;
    return static_cast<uint16_t>(std::round(jaccard * 255.0));
}

// shard_K/ writer: same file formats as reference :645-817 (see the header of this file)
void write_shard(const std::string& folder, const std::vector<Pair>& ordered, const std::vector<double>& norms_sq, int dimension)
{
    if (!fs::exists(folder)) fs::create_directories(folder);
    std::unordered_map<uint32_t, std::vector<std::pair<uint32_t, uint16_t>>> by_row;
    for (const Pair& p : ordered)
        by_row[static_cast<uint32_t>(p.row)].push_back({static_cast<uint32_t>(p.col), quantize(p, norms_sq, dimension)});

    std::ofstream matrix(folder + "matrix.bin", std::ios::binary);
    std::ofstream index(folder + "row_index.bin", std::ios::binary);
    const size_t nrows = by_row.size();
    std::vector<uint32_t> row_ids(nrows), first_neighbor(nrows);
    std::vector<uint64_t> row_pos(nrows);
    uint64_t jaccard_bytes = 0, neighbor_bytes = 0;
    size_t at = 0;
    for (auto& [row, neighbors] : by_row) {                            // unordered_map order = the reference's row order
        std::vector<uint16_t> jac(neighbors.size());
        std::vector<uint64_t> gaps(neighbors.size() - 1);
        for (size_t k = 0; k < neighbors.size(); ++k) {
            jac[k] = neighbors[k].second;
            if (k) gaps[k - 1] = neighbors[k].first - neighbors[k - 1].first;   // columns ascend within a row
        }
        row_ids[at] = row;
        row_pos[at] = static_cast<uint64_t>(matrix.tellp());
        first_neighbor[at] = neighbors[0].first;
        ++at;
        mvs_succinct::CompactVector cv;
        cv.build(jac.begin(), jac.size());
        cv.save(matrix);
        jaccard_bytes += cv.num_bytes();
        if (neighbors.size() == 1) continue;                           // a lone neighbour has no gap sequence
        mvs_succinct::RiceSequence rs;
        rs.encode(gaps.begin(), gaps.size());
        rs.save(matrix);
        neighbor_bytes += rs.num_bytes();
    }
    matrix.flush();
    matrix.close();

    mvs_succinct::CompactVector cv_rows;
    cv_rows.build(row_ids.begin(), row_ids.size());
    cv_rows.save(index);
    std::vector<uint64_t> pos_gaps(nrows ? nrows - 1 : 0);
    for (size_t i = 1; i < nrows; ++i) pos_gaps[i - 1] = row_pos[i] - row_pos[i - 1];
    mvs_succinct::CompactVector cv_pos;
    cv_pos.build(pos_gaps.begin(), pos_gaps.size());
    cv_pos.save(index);
    index.close();

    std::ofstream starts(folder + "neighbor_start.bin", std::ios::binary);
    mvs_succinct::RiceSequence rs_start;
    rs_start.encode(first_neighbor.begin(), first_neighbor.size());
    rs_start.save(starts);
    starts.close();
    neighbor_bytes += rs_start.num_bytes();
    std::cout << "Jac space: " << jaccard_bytes << " ngh space: " << neighbor_bytes << std::endl;
}

}  // namespace

int main(int argc, char** argv)
{
    Args a;
    if (!parse(argc, argv, a) || a.help) {
        usage(argv[0]);
        return a.help ? 0 : 1;
    }
    // string-concatenated paths like the reference (:852-873): --db must end in '/'
    const std::string norms_file = a.db + "vector_norms.txt";
    if (!fs::exists(norms_file)) {
        std::cerr << "Error: Required file 'vector_norms.txt' not found in output folder: " << a.db << std::endl;
        return 1;
    }
    std::string dtype = "int32";
    {
        std::ifstream in(a.db + "dtype.txt");
        if (in) std::getline(in, dtype);
    }
    int dimension = 0;
    {
        std::ifstream in(a.db + "dimension.txt");
        if (in) in >> dimension;
    }
    std::cout << "dtypeqs: " << dtype << std::endl;
    // int16 dbs (`project_everything sketch --int16`): the reference hands them to pairwise_comp_optimized_16bits
    // (:876-879), a different tool with an older shard format (out of scope).  Here the int16 rows are
    // widened to int32 on load and take the SAME path and writer as an int32 db: the shard files are the int32 tool's
    // format, which query_pc_mat reads; the pairs are those the int32 tool finds on the saturated vectors.
    const bool db_is_int16 = dtype == "int16";
    if (db_is_int16) std::cerr << "pairwise_comp_b200: int16 db, rows widened to int32 (int32 shard format)" << std::endl;
    if (dimension <= 0) {
        std::cerr << "pairwise_comp_b200: bad dimension.txt" << std::endl;
        return 1;
    }
    std::string out = a.out;
    if (!out.empty() && out.back() != '/' && out.back() != '\\') out += '/';

    std::vector<double> norms_sq;                                      // :889-901
    {
        std::ifstream in(norms_file);
        std::string line;
        while (std::getline(in, line)) {
            const size_t pos = line.find(' ');
            if (pos == std::string::npos) continue;
            const double n = std::stod(line.substr(pos + 1));
            norms_sq.push_back(n * n);
        }
    }
    const int bytes_per_vector = dimension * static_cast<int>(sizeof(int32_t));
    const int64_t max_bytes = static_cast<int64_t>(a.max_memory_gb * 1024 * 1024 * 1024);
    std::cout << "max bytes " << max_bytes << " " << a.max_memory_gb << std::endl;
    const int chunk = static_cast<int>(max_bytes / (static_cast<int64_t>(bytes_per_vector) * bytes_per_vector));   // :905
    std::cout << "Using chunks of size " << chunk << std::endl;
    if (chunk <= 0) {
        std::cerr << "pairwise_comp_b200: --max_memory_gb gives an empty chunk (the reference would not terminate)" << std::endl;
        return 1;
    }
    const std::string matrix_file = a.db + "vectors.bin";
    std::ifstream vf(matrix_file, std::ios::ate | std::ios::binary);
    const int64_t file_size = vf ? static_cast<int64_t>(vf.tellg()) : 0;
    const int total = static_cast<int>(file_size / (db_is_int16 ? bytes_per_vector / 2 : bytes_per_vector));
    std::cout << "Total vectors: " << total << std::endl;
    if (static_cast<size_t>(total) > norms_sq.size()) {
        std::cerr << "pairwise_comp_b200: vector_norms.txt has fewer lines than vectors.bin has vectors" << std::endl;
        return 1;
    }
    const auto t0 = std::chrono::high_resolution_clock::now();
    auto since = [&](std::chrono::high_resolution_clock::time_point t) {
        return std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t).count();
    };
    // CUDA context creation (0.5 - 1.3 s) runs on a second thread while vectors.bin is read
    mvs_ctx* ctx = nullptr;
    int create_rc = MVS_OK;
    std::thread creator([&] { create_rc = mvs_create(&ctx, a.gpu); });
    const size_t db_count = static_cast<size_t>(total) * dimension;
    std::unique_ptr<int32_t[]> db(new (std::nothrow) int32_t[db_count ? db_count : 1]);   // not zero-filled: read() fills it
    if (!db) {
        creator.join();
        if (ctx) mvs_destroy(ctx);
        std::cerr << "pairwise_comp_b200: out of memory for " << matrix_file << std::endl;
        return 1;
    }
    vf.seekg(0);
    if (db_is_int16) {                                                 // read into the upper half, widen front to back
        char* half = reinterpret_cast<char*>(db.get()) + db_count * sizeof(int16_t);
        vf.read(half, static_cast<std::streamsize>(db_count * sizeof(int16_t)));
        for (size_t k = 0; k < db_count; ++k) {
            int16_t v;
            std::memcpy(&v, half + k * sizeof(int16_t), sizeof v);
            db[k] = v;
        }
    } else {
        vf.read(reinterpret_cast<char*>(db.get()), static_cast<std::streamsize>(db_count * sizeof(int32_t)));
    }
    vf.close();
    const double read_s = since(t0);

    const std::string shard_folder = out + "shard_" + std::to_string(a.shard_idx) + "/";
    const int rows_per_shard = (total + a.num_shards - 1) / a.num_shards;
    const int begin_row = std::min(a.shard_idx * rows_per_shard, total);
    const int end_row = std::min(begin_row + rows_per_shard, total);
    std::cout << "Shard " << a.shard_idx << " processing rows " << begin_row << " to " << end_row << std::endl;

    const auto t_wait = std::chrono::high_resolution_clock::now();
    creator.join();
    const double wait_ctx_s = since(t_wait);
    int rc = create_rc;
    if (rc != MVS_OK) {
        std::cerr << "pairwise_comp_b200: cuda:" << a.gpu << ": " << mvs_strerror(rc) << std::endl;
        return 2;
    }
    if (!fs::exists(shard_folder)) fs::create_directories(shard_folder);   // only once a device exists
    const auto t_shard = std::chrono::high_resolution_clock::now();
    mvs_sparse_result* r = nullptr;
    rc = mvs_pairwise_shard(ctx, db.get(), total, dimension, norms_sq.data(), begin_row, end_row, &r);
    if (rc != MVS_OK) {
        std::cerr << "pairwise_comp_b200: " << mvs_strerror(rc) << std::endl;
        mvs_destroy(ctx);
        return 2;
    }
    const double shard_s = since(t_shard);
    db.reset();
    const auto t_write = std::chrono::high_resolution_clock::now();
    // (row, col)-sorted GPU output -> the order in which the reference appends to all_results (:949-983):
    // row chunks outermost, then column chunks, then rows, then columns (a stable sort on the two chunk ids)
    std::vector<Pair> ordered(static_cast<size_t>(r->n));
    for (int64_t k = 0; k < r->n; ++k) ordered[static_cast<size_t>(k)] = {r->rows[k], r->cols[k], r->values[k]};
    const double gemm_ms = r->gemm_ms, filter_ms = r->filter_ms, prep_ms = r->prep_ms;
    mvs_free_sparse_result(r);
    std::thread destroyer([&] { mvs_destroy(ctx); });                  // context teardown overlaps the writer
    std::stable_sort(ordered.begin(), ordered.end(), [&](const Pair& x, const Pair& y) {
        const int xi = (x.row - begin_row) / chunk, yi = (y.row - begin_row) / chunk;
        if (xi != yi) return xi < yi;
        return x.col / chunk < y.col / chunk;
    });
    write_shard(shard_folder, ordered, norms_sq, dimension);
    const double write_s = since(t_write);
    destroyer.join();
    const auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - t0).count();
    std::cout << "Total computation time: " << ms << " ms" << std::endl;
    if (std::getenv("MVS_TIMING"))
        std::fprintf(stderr, "mvs_timing pairs_kept=%zu gemm_ms=%.3f filter_ms=%.3f prep_ms=%.3f read_s=%.3f wait_ctx_s=%.3f shard_s=%.3f write_s=%.3f\n",
                     ordered.size(), gemm_ms, filter_ms, prep_ms, read_s, wait_ctx_s, shard_s, write_s);
    return 0;
}
