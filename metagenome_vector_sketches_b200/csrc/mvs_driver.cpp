// mvs_driver.cpp -- the reference's two projection drivers re-built over the C ABI.
//
//   mvs_parse_hash_file   hash-file parser of sketch()            reference src/project_everything.cpp:257-284
//                         and the per-line parser of standalone   reference src/standalone_projection.cpp:27-34
//   mvs_write_db          db/ writer of sketch()                  reference src/project_everything.cpp:306-361
//   mvs_sketch_file       sketch() end to end                     reference src/project_everything.cpp:238-362
//   mvs_standalone_file   standalone_projection main              reference src/standalone_projection.cpp:11-46
//
// Host-only C++ (no CUDA here): all projection work goes through mvs_sketch_batch /
// mvs_project_batch.  Parsing mirrors `istringstream >> unsigned long` exactly (see
// parse_tokens) because the reference's set contents, and therefore vectors.bin, depend on it.
#include "../../include/mvs_b200.h"

#include <sys/mman.h>
#include <sys/stat.h>
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <string>
#include <thread>
#include <vector>

namespace fs = std::filesystem;

namespace {

inline bool is_space(unsigned char c)            // isspace() in the "C" locale, as used by operator>>
{
    return c == ' ' || (c >= '\t' && c <= '\r');
}

// `while (iss >> hash) set.insert(hash)` for unsigned long (libstdc++ num_get, base 10):
//   * leading whitespace is skipped; an optional '+' or '-' sign is accepted; "-N" wraps mod 2^64;
//   * at least one digit is required, otherwise the extraction fails and the loop ends;
//   * a magnitude above 2^64-1 sets failbit: nothing is inserted and the loop ends;
//   * extraction stops at the first non-digit; the NEXT extraction starts right there, so "5x"
//     yields 5 and then ends the line, while "5-3" yields 5 and 2^64-3.
void parse_tokens(const char* p, const char* e, std::vector<uint64_t>& out)
{
    for (;;) {
        while (p < e && is_space(static_cast<unsigned char>(*p))) ++p;
        if (p >= e) return;
        bool neg = false;
        if (*p == '+' || *p == '-') {
            neg = (*p == '-');
            ++p;
        }
        if (p >= e || *p < '0' || *p > '9') return;
        uint64_t v = 0;
        bool overflow = false;
        while (p < e && *p >= '0' && *p <= '9') {
            const uint64_t dgt = static_cast<uint64_t>(*p - '0');
            if (v > (UINT64_MAX - dgt) / 10) overflow = true;
            v = v * 10 + dgt;
            ++p;
        }
        if (overflow) return;
        out.push_back(neg ? (0 - v) : v);
    }
}

struct Mapped {
    const char* data = nullptr;
    size_t size = 0;
    int fd = -1;
    bool ok = false;
    explicit Mapped(const char* path)
    {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) return;
        struct stat st;
        if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) return;
        size = static_cast<size_t>(st.st_size);
        if (size == 0) {
            ok = true;
            return;
        }
        void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) return;
        madvise(m, size, MADV_SEQUENTIAL);
        data = static_cast<const char*>(m);
        ok = true;
    }
    ~Mapped()
    {
        if (data) munmap(const_cast<char*>(data), size);
        if (fd >= 0) ::close(fd);
    }
};

void* host_alloc_any(size_t bytes, int* pinned)
{
    void* p = mvs_host_alloc(bytes);             // pinned when a CUDA device is present
    if (p) {
        *pinned = 1;
        return p;
    }
    *pinned = 0;
    return std::malloc(bytes ? bytes : 1);
}
void host_free_any(void* p, int pinned)
{
    if (!p) return;
    if (pinned) mvs_host_free(p);
    else std::free(p);
}

std::string format_g(double x)                   // default ostream formatting of a double (precision 6)
{
    char buf[64];
    std::snprintf(buf, sizeof buf, "%g", x);
    return buf;
}

}  // namespace

extern "C" {

void mvs_free_hash_batch(mvs_hash_batch* b)
{
    if (!b) return;
    host_free_any(b->hashes, b->hashes_pinned);
    std::free(b->offsets);
    if (b->names) {
        for (int64_t i = 0; i < b->n_samples; ++i) std::free(b->names[i]);
        std::free(b->names);
    }
    std::free(b);
}

int mvs_parse_hash_file(const char* path, int named, mvs_hash_batch** out)
{
    if (!path || !out) return MVS_ERR_INVALID_ARG;
    *out = nullptr;
    Mapped mf(path);
    if (!mf.ok) return MVS_ERR_IO;

    // 1. line boundaries (std::getline semantics: '\n' separated, last line kept if non-empty)
    struct Line { const char* name_b; const char* name_e; const char* b; const char* e; };
    std::vector<Line> lines;
    const char* p = mf.data;
    const char* end = mf.data + mf.size;
    while (p < end) {
        const char* nl = static_cast<const char*>(std::memchr(p, '\n', static_cast<size_t>(end - p)));
        const char* le = nl ? nl : end;
        if (named) {
            const char* colon = static_cast<const char*>(std::memchr(p, ':', static_cast<size_t>(le - p)));
            if (colon) lines.push_back({p, colon, colon + 1, le});     // lines without ':' are skipped (:267)
        } else {
            lines.push_back({nullptr, nullptr, p, le});
        }
        if (!nl) break;
        p = nl + 1;
    }

    // 2. parse + dedupe every line (parallel over lines)
    const int64_t n = static_cast<int64_t>(lines.size());
    std::vector<std::vector<uint64_t>> sets(static_cast<size_t>(n));
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t i = 0; i < n; ++i) {
        std::vector<uint64_t>& v = sets[static_cast<size_t>(i)];
        const Line& ln = lines[static_cast<size_t>(i)];
        v.reserve(static_cast<size_t>(ln.e - ln.b) / 16 + 4);
        parse_tokens(ln.b, ln.e, v);
        std::sort(v.begin(), v.end());                                  // unordered_set semantics
        v.erase(std::unique(v.begin(), v.end()), v.end());
    }

    // 3. CSR in (pinned) host memory
    mvs_hash_batch* b = static_cast<mvs_hash_batch*>(std::calloc(1, sizeof(mvs_hash_batch)));
    if (!b) return MVS_ERR_OOM;
    b->n_samples = n;
    b->offsets = static_cast<uint64_t*>(std::malloc(sizeof(uint64_t) * static_cast<size_t>(n + 1)));
    if (!b->offsets) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
    uint64_t total = 0;
    for (int64_t i = 0; i < n; ++i) {
        b->offsets[i] = total;
        total += sets[static_cast<size_t>(i)].size();
    }
    b->offsets[n] = total;
    b->n_hashes = total;
    b->hashes = static_cast<uint64_t*>(host_alloc_any(sizeof(uint64_t) * (total + 2), &b->hashes_pinned));
    if (!b->hashes) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t i = 0; i < n; ++i) {
        const auto& v = sets[static_cast<size_t>(i)];
        if (!v.empty()) std::memcpy(b->hashes + b->offsets[i], v.data(), v.size() * sizeof(uint64_t));
    }
    if (named) {
        b->names = static_cast<char**>(std::calloc(static_cast<size_t>(n ? n : 1), sizeof(char*)));
        if (!b->names) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
        for (int64_t i = 0; i < n; ++i) {
            const Line& ln = lines[static_cast<size_t>(i)];
            const size_t len = static_cast<size_t>(ln.name_e - ln.name_b);
            b->names[i] = static_cast<char*>(std::malloc(len + 1));
            if (!b->names[i]) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
            std::memcpy(b->names[i], ln.name_b, len);
            b->names[i][len] = '\0';
        }
    }
    *out = b;
    return MVS_OK;
}

}  // extern "C"

// ======================================================================================
// db/ writer, sketch(), standalone_projection
// ======================================================================================
extern "C" {

int mvs_write_db(const char* index_folder, int64_t n_samples, int d, const char* const* names,
                 const uint64_t* sumsq, const int32_t* vec32, const int16_t* vec16)
{
    if (!index_folder || n_samples < 0 || d <= 0 || (!vec32 && !vec16 && n_samples > 0) ||
        (n_samples > 0 && (!names || !sumsq)))
        return MVS_ERR_INVALID_ARG;
    std::string dir(index_folder);
    if (dir.empty() || dir.back() != '/') dir += '/';
    FILE* fn = std::fopen((dir + "vector_norms.txt").c_str(), "w");
    FILE* fd = std::fopen((dir + "dimension.txt").c_str(), "w");
    FILE* ft = std::fopen((dir + "dtype.txt").c_str(), "w");
    FILE* fb = std::fopen((dir + "vectors.bin").c_str(), "wb");
    int rc = MVS_OK;
    if (!fn) std::fprintf(stderr, "Error opening vector_norms.txt for writing.\n");
    if (!fb) std::fprintf(stderr, "Error opening vectors.bin for writing.\n");
    if (fn && fd && ft && fb) {
        std::fprintf(fd, "%d\n", d);                                    // :319
        std::fprintf(ft, "%s\n", vec16 ? "int16" : "int32");            // :320
        std::string text;
        text.reserve(static_cast<size_t>(n_samples) * 24);
        for (int64_t i = 0; i < n_samples; ++i) {
            // norm = || v / sqrt(d) ||  (:327-329), from the exact integer sum of squares
            const double norm = std::sqrt(static_cast<double>(sumsq[i]) / static_cast<double>(d));
            text += names[i];
            text += ' ';
            text += format_g(norm);                                     // default ostream precision (:330)
            text += '\n';
        }
        if (std::fwrite(text.data(), 1, text.size(), fn) != text.size()) rc = MVS_ERR_IO;
        const size_t count = static_cast<size_t>(n_samples) * static_cast<size_t>(d);
        const size_t wrote = vec16 ? std::fwrite(vec16, sizeof(int16_t), count, fb)
                                   : std::fwrite(vec32, sizeof(int32_t), count, fb);
        if (wrote != count) rc = MVS_ERR_IO;
    } else {
        rc = MVS_ERR_IO;
    }
    for (FILE* f : {fn, fd, ft, fb})
        if (f && std::fclose(f) != 0) rc = MVS_ERR_IO;
    return rc;
}

// Contiguous size-balanced static partition of samples over n_parts GPUs:
// bounds[k] = first sample of part k, bounds[n_parts] = n_samples; part k ends at the first
// sample boundary where the running hash count reaches k/n_parts of the total.
void mvs_partition_samples(const uint64_t* offsets, int64_t n_samples, int n_parts, int64_t* bounds)
{
    const uint64_t base = offsets[0];
    const uint64_t total = offsets[n_samples] - base;
    bounds[0] = 0;
    int64_t s = 0;
    for (int k = 1; k < n_parts; ++k) {
        // weight = hashes + a small per-sample cost so that empty samples are spread too
        const long double target = static_cast<long double>(total + static_cast<uint64_t>(n_samples)) * k / n_parts;
        while (s < n_samples &&
               static_cast<long double>(offsets[s + 1] - base + static_cast<uint64_t>(s + 1)) <= target)
            ++s;
        bounds[k] = s;
    }
    bounds[n_parts] = n_samples;
}

int mvs_sketch_file(mvs_ctx* const* ctxs, int n_ctx, const char* hash_file, const char* index_folder, int d,
                    int use_int16, uint64_t seed, int verbose)
{
    if (!ctxs || n_ctx <= 0 || !hash_file || !index_folder || d <= 0) return MVS_ERR_INVALID_ARG;
    std::string dir(index_folder);
    if (dir.empty() || dir.back() != '/') dir += '/';                   // :239-241
    std::error_code ec;
    if (fs::exists(dir, ec)) {                                          // :244-252 (wipes the folder)
        for (const auto& entry : fs::directory_iterator(dir, ec)) fs::remove_all(entry.path(), ec);
    } else {
        fs::create_directories(dir, ec);
    }
    const auto t0 = std::chrono::steady_clock::now();                   // :255 timer includes the parse

    mvs_hash_batch* hb = nullptr;
    int rc = mvs_parse_hash_file(hash_file, 1, &hb);
    if (rc != MVS_OK) {
        std::fprintf(stderr, "Error opening %s for reading.\n", hash_file);   // :260
        return rc;
    }
    if (verbose) std::printf("Loaded %lld hash sets from %s\n", static_cast<long long>(hb->n_samples), hash_file);

    const int64_t n = hb->n_samples;
    const size_t count = static_cast<size_t>(n) * static_cast<size_t>(d);
    int pin32 = 0, pin16 = 0;
    int32_t* v32 = use_int16 ? nullptr : static_cast<int32_t*>(host_alloc_any(count * sizeof(int32_t), &pin32));
    int16_t* v16 = use_int16 ? static_cast<int16_t*>(host_alloc_any(count * sizeof(int16_t), &pin16)) : nullptr;
    std::vector<uint64_t> sumsq(static_cast<size_t>(n ? n : 1));
    if ((!use_int16 && !v32) || (use_int16 && !v16)) rc = MVS_ERR_OOM;

    if (rc == MVS_OK && n > 0) {
        const int parts = static_cast<int>(std::min<int64_t>(n_ctx, n));
        std::vector<int64_t> bounds(static_cast<size_t>(parts) + 1);
        mvs_partition_samples(hb->offsets, n, parts, bounds.data());
        std::vector<int> rcs(static_cast<size_t>(parts), MVS_OK);
        auto work = [&](int k) {
            const int64_t a = bounds[static_cast<size_t>(k)], b = bounds[static_cast<size_t>(k) + 1];
            if (b <= a) return;
            rcs[static_cast<size_t>(k)] = mvs_sketch_batch(
                ctxs[k], hb->hashes, hb->offsets + a, b - a, d, seed,
                v32 ? v32 + static_cast<size_t>(a) * d : nullptr, v16 ? v16 + static_cast<size_t>(a) * d : nullptr,
                sumsq.data() + a);
        };
        std::vector<std::thread> th;
        for (int k = 1; k < parts; ++k) th.emplace_back(work, k);      // one host thread per GPU
        work(0);
        for (auto& t : th) t.join();
        for (int r : rcs)
            if (r != MVS_OK && rc == MVS_OK) rc = r;
    }
    if (rc == MVS_OK) {
        if (verbose)
            for (int64_t i = 0; i < n; ++i)                             // :294-297 (order is arbitrary there)
                std::printf("Projected %s, vector dimension %d, index %lld\n", hb->names[i], d,
                            static_cast<long long>(i));
        const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::printf("Time to compute all projected vectors: %g seconds\n", sec);   // :300-303
        rc = mvs_write_db(dir.c_str(), n, d, hb->names, sumsq.data(), v32, v16);
    } else {
        std::fprintf(stderr, "mvs: projection failed: %s\n", mvs_strerror(rc));
    }
    host_free_any(v32, pin32);
    host_free_any(v16, pin16);
    mvs_free_hash_batch(hb);
    return rc;
}

int mvs_standalone_file(mvs_ctx* ctx, const char* hashes_file, int d, uint64_t seed, const char* out_path)
{
    if (!ctx || !hashes_file || d <= 0) return MVS_ERR_INVALID_ARG;
    mvs_hash_batch* hb = nullptr;
    int rc = mvs_parse_hash_file(hashes_file, 0, &hb);
    if (rc != MVS_OK) {
        std::fprintf(stderr, "Error opening file: %s\n", hashes_file);  // standalone_projection.cpp:23
        return rc;
    }
    const int64_t n = hb->n_samples;
    std::vector<int32_t> vec(static_cast<size_t>(n) * static_cast<size_t>(d) + 1);
    rc = mvs_project_batch(ctx, hb->hashes, hb->offsets, n, d, seed, vec.data());
    mvs_free_hash_batch(hb);
    if (rc != MVS_OK) return rc;
    FILE* fo = out_path ? std::fopen(out_path, "w") : stdout;
    if (!fo) return MVS_ERR_IO;
    std::string line;
    for (int64_t i = 0; i < n; ++i) {                                   // :38-42: static_cast<float>, ' ' separated
        line.clear();
        const int32_t* v = vec.data() + static_cast<size_t>(i) * d;
        char buf[32];
        for (int j = 0; j < d; ++j) {
            std::snprintf(buf, sizeof buf, "%g", static_cast<double>(static_cast<float>(v[j])));
            line += buf;
            if (j != d - 1) line += ' ';
        }
        line += '\n';
        std::fwrite(line.data(), 1, line.size(), fo);
    }
    if (out_path) std::fclose(fo);
    else std::fflush(fo);
    return MVS_OK;
}

}  // extern "C"
