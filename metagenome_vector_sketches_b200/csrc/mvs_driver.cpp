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

#include <omp.h>

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

#include <atomic>
static std::atomic<int> g_norms_mode{0};         // mvs_set_norms_mode: 0 = reference float32 (default), 1 = exact

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
// Every token owns one maximal run of digits, so the number of digit runs of a line is an upper
// bound of its token count (exact for well-formed lines): count_digit_runs() sizes the output.

inline bool is_digit(unsigned char c) { return static_cast<unsigned>(c - '0') < 10u; }

// eight ASCII digits at once (little endian: first character in the lowest byte)
inline bool all_digits8(uint64_t w)
{
    return ((w & 0xF0F0F0F0F0F0F0F0ull) | (((w + 0x0606060606060606ull) & 0xF0F0F0F0F0F0F0F0ull) >> 4)) ==
           0x3333333333333333ull;
}
inline uint64_t parse8(uint64_t w)
{
    w -= 0x3030303030303030ull;
    w = (w * 10) + (w >> 8);
    return (((w & 0x000000FF000000FFull) * (100ull + (1000000ull << 32))) +
            (((w >> 16) & 0x000000FF000000FFull) * (1ull + (10000ull << 32)))) >> 32;
}

// high bit of every byte that is an ASCII digit
inline uint64_t digit_mask8(uint64_t w)
{
    const uint64_t x = w & 0x7F7F7F7F7F7F7F7Full;
    const uint64_t ge30 = x + 0x5050505050505050ull;                    // bit 7 set iff low 7 bits >= 0x30
    const uint64_t ge3a = x + 0x4646464646464646ull;                    // bit 7 set iff low 7 bits >= 0x3a
    return ge30 & ~ge3a & ~w & 0x8080808080808080ull;
}

uint64_t count_digit_runs(const char* b, const char* e)
{
    uint64_t n = 0;
    uint64_t prev = 0;                                                  // was the byte before p a digit (bit 7)
    const char* p = b;
    for (; p + 8 <= e; p += 8) {
        uint64_t w;
        std::memcpy(&w, p, 8);
        const uint64_t dm = digit_mask8(w);
        n += static_cast<uint64_t>(__builtin_popcountll(dm & ~((dm << 8) | prev)));   // digit whose predecessor is none
        prev = dm >> 56;
    }
    bool pd = prev != 0;
    for (; p < e; ++p) {
        const bool dg = is_digit(static_cast<unsigned char>(*p));
        n += static_cast<uint64_t>(dg & !pd);
        pd = dg;
    }
    return n;
}

// Parses [p, e) into dst (room for count_digit_runs(p, e) values).  `safe_end` is the end of the
// mapping: 8-byte loads are only issued while p + 8 <= safe_end.  Returns the token count;
// *strictly_increasing tells whether the values came sorted and unique (then no dedupe is needed).
size_t parse_tokens(const char* p, const char* e, const char* safe_end, uint64_t* dst, bool* strictly_increasing)
{
    size_t n = 0;
    bool inc = true;
    uint64_t prev = 0;
    for (;;) {
        while (p < e && is_space(static_cast<unsigned char>(*p))) ++p;
        if (p >= e) break;
        bool neg = false;
        if (*p == '+' || *p == '-') {
            neg = (*p == '-');
            ++p;
        }
        if (p >= e || !is_digit(static_cast<unsigned char>(*p))) break;
        uint64_t v = 0;
        int nd = 0;
        // up to 16 digits in two 8-digit steps: no overflow possible below 19 digits
        while (nd <= 8 && p + 8 <= e && p + 8 <= safe_end) {
            uint64_t w;
            std::memcpy(&w, p, 8);
            if (!all_digits8(w)) break;
            v = v * 100000000ull + parse8(w);
            p += 8;
            nd += 8;
        }
        bool overflow = false;
        while (p < e && is_digit(static_cast<unsigned char>(*p))) {
            const uint64_t dgt = static_cast<uint64_t>(*p - '0');
            if (nd >= 19 && v > (UINT64_MAX - dgt) / 10) overflow = true;
            v = v * 10 + dgt;
            ++p;
            ++nd;
        }
        if (overflow) break;
        if (neg) v = 0 - v;
        inc &= (n == 0) | (v > prev);
        prev = v;
        dst[n++] = v;
    }
    *strictly_increasing = inc;
    return n;
}

// Drops repeated values in place keeping the first occurrence of each (unordered_set semantics; the
// order inside a set does not matter to the projection, a sum).  Open addressing in a per-thread
// table of >= 2n slots, ~10x cheaper than sort + unique for the 5e4-hash lines of config 2.
size_t dedupe_first_occurrence(uint64_t* v, size_t n)
{
    static thread_local std::vector<uint64_t> table;
    int lg = 4;
    while ((size_t{1} << lg) < 2 * n) ++lg;
    const size_t cap = size_t{1} << lg;
    table.assign(cap, 0);                                               // 0 = empty; the value 0 is tracked apart
    bool seen_zero = false;
    size_t m = 0;
    for (size_t i = 0; i < n; ++i) {
        const uint64_t x = v[i];
        if (x == 0) {
            if (!seen_zero) { seen_zero = true; v[m++] = 0; }
            continue;
        }
        size_t h = static_cast<size_t>((x * 0x9E3779B97F4A7C15ull) >> (64 - lg));
        while (table[h] != 0 && table[h] != x) h = (h + 1) & (cap - 1);
        if (table[h] == 0) { table[h] = x; v[m++] = x; }
    }
    return m;
}

// One line -> a set: already one if the tokens came strictly increasing (sourmash writes its mins
// sorted), otherwise deduplicated.
size_t parse_line_set(const char* b, const char* e, const char* safe_end, uint64_t* dst)
{
    bool inc = true;
    size_t n = parse_tokens(b, e, safe_end, dst, &inc);
    if (!inc) n = dedupe_first_occurrence(dst, n);
    return n;
}

struct LineRef {
    const char* name_b;
    const char* name_e;
    const char* b;                               // token text
    const char* e;
    uint64_t bound;                              // upper bound of the token count
};

// Line index of a mapped hash file: std::getline semantics ('\n' separated, a last line without
// '\n' is kept if non-empty), lines without ':' skipped in named mode (reference :267).  The
// newline scan and the per-line digit-run count both run in parallel.
void index_lines(const char* data, size_t size, int named, std::vector<LineRef>& lines)
{
    lines.clear();
    if (size == 0) return;
    int T = 1;
#pragma omp parallel
    {
#pragma omp single
        T = omp_get_num_threads();
    }
    const size_t piece = std::max<size_t>(1 << 20, (size + static_cast<size_t>(T) - 1) / static_cast<size_t>(T));
    const size_t npieces = (size + piece - 1) / piece;
    std::vector<std::vector<size_t>> nl(npieces);
#pragma omp parallel for schedule(dynamic, 1)
    for (size_t k = 0; k < npieces; ++k) {
        const char* p = data + k * piece;
        const char* pe = data + std::min(size, (k + 1) * piece);
        while (p < pe) {
            const char* q = static_cast<const char*>(std::memchr(p, '\n', static_cast<size_t>(pe - p)));
            if (!q) break;
            nl[k].push_back(static_cast<size_t>(q - data));
            p = q + 1;
        }
    }
    size_t pos = 0;
    auto add_line = [&](const char* b, const char* e) {
        if (named) {
            const char* colon = static_cast<const char*>(std::memchr(b, ':', static_cast<size_t>(e - b)));
            if (colon) lines.push_back({b, colon, colon + 1, e, 0});
        } else {
            lines.push_back({nullptr, nullptr, b, e, 0});
        }
    };
    for (size_t k = 0; k < npieces; ++k)
        for (size_t q : nl[k]) {
            add_line(data + pos, data + q);
            pos = q + 1;
        }
    if (pos < size) add_line(data + pos, data + size);
    const int64_t n = static_cast<int64_t>(lines.size());
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t i = 0; i < n; ++i) lines[static_cast<size_t>(i)].bound = count_digit_runs(lines[i].b, lines[i].e);
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

// ---- binary hash file (".u64 sidecar") ---------------------------------------------------------------
// The parsed batch as it sits in memory, so that a second run over the same samples skips the text parse:
//   header (64 B)   magic "MVSU64\0\1" | u64 n_samples | u64 n_hashes | u64 names_bytes | u64 payload_offset | zeros
//   offsets         u64[n_samples + 1]
//   names           n_samples NUL-terminated strings, names_bytes in total (0 = a file without names)
//   padding         zeros up to payload_offset (a multiple of 4096)
//   payload         u64[n_hashes], the CSR payload exactly as mvs_parse_hash_file returns it
// Little endian like everything else here.  mvs_parse_hash_file recognises the magic (no text hash file starts with
// it) and reads the payload straight into the pinned CSR buffer with parallel pread calls.
constexpr char kBinMagic[8] = {'M', 'V', 'S', 'U', '6', '4', '\0', '\1'};
constexpr uint64_t kBinHeader = 64, kBinAlign = 4096;

struct BinHeader {
    char magic[8];
    uint64_t n_samples, n_hashes, names_bytes, payload_offset;
    uint64_t reserved[3];
};
static_assert(sizeof(BinHeader) == kBinHeader, "binary hash file header is 64 bytes");

bool write_all(int fd, const void* p, size_t n)
{
    const char* c = static_cast<const char*>(p);
    while (n) {
        const ssize_t w = ::write(fd, c, std::min<size_t>(n, size_t(1) << 30));
        if (w <= 0) return false;
        c += w;
        n -= static_cast<size_t>(w);
    }
    return true;
}

bool pread_all(int fd, void* p, size_t n, uint64_t at)
{
    char* c = static_cast<char*>(p);
    while (n) {
        const ssize_t r = ::pread(fd, c, std::min<size_t>(n, size_t(1) << 30), static_cast<off_t>(at));
        if (r <= 0) return false;
        c += r;
        n -= static_cast<size_t>(r);
        at += static_cast<uint64_t>(r);
    }
    return true;
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

// Binary hash file -> batch.  Everything a hostile or truncated file could break is checked before a byte of the
// payload is trusted: sizes against the file size, monotone offsets ending at n_hashes, one NUL per name.
static int load_binary_batch(int fd, const char* data, size_t size, int named, bool want_pinned, mvs_hash_batch** out)
{
    BinHeader h;
    std::memcpy(&h, data, sizeof h);
    const uint64_t n = h.n_samples;
    if (n >= (uint64_t(1) << 40) || h.n_hashes >= (uint64_t(1) << 48) || h.names_bytes >= (uint64_t(1) << 40)) return MVS_ERR_IO;
    const uint64_t off_bytes = (n + 1) * sizeof(uint64_t);
    const uint64_t names_at = kBinHeader + off_bytes;
    if (h.payload_offset % kBinAlign || h.payload_offset < names_at + h.names_bytes ||
        h.payload_offset + h.n_hashes * sizeof(uint64_t) != size)
        return MVS_ERR_IO;
    if (named && n > 0 && h.names_bytes == 0) return MVS_ERR_INVALID_ARG;       // sketch() needs the names
    const uint64_t* off = reinterpret_cast<const uint64_t*>(data + kBinHeader);   // 8-byte aligned: mmap base + 64
    if (off[0] != 0 || off[n] != h.n_hashes) return MVS_ERR_IO;
    for (uint64_t i = 0; i < n; ++i)
        if (off[i + 1] < off[i]) return MVS_ERR_IO;

    mvs_hash_batch* b = static_cast<mvs_hash_batch*>(std::calloc(1, sizeof(mvs_hash_batch)));
    if (!b) return MVS_ERR_OOM;
    b->n_samples = static_cast<int64_t>(n);
    b->n_hashes = h.n_hashes;
    b->offsets = static_cast<uint64_t*>(std::malloc(off_bytes));
    if (!b->offsets) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
    std::memcpy(b->offsets, off, off_bytes);
    if (named) {
        b->names = static_cast<char**>(std::calloc(static_cast<size_t>(n ? n : 1), sizeof(char*)));
        if (!b->names) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
        const char* p = data + names_at;
        const char* e = p + h.names_bytes;
        for (uint64_t i = 0; i < n; ++i) {
            const char* z = static_cast<const char*>(std::memchr(p, 0, static_cast<size_t>(e - p)));
            if (!z) { mvs_free_hash_batch(b); return MVS_ERR_IO; }
            b->names[i] = static_cast<char*>(std::malloc(static_cast<size_t>(z - p) + 1));
            if (!b->names[i]) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
            std::memcpy(b->names[i], p, static_cast<size_t>(z - p) + 1);
            p = z + 1;
        }
    }
    const size_t bytes = sizeof(uint64_t) * (h.n_hashes + 2);
    if (want_pinned) {
        b->hashes = static_cast<uint64_t*>(host_alloc_any(bytes, &b->hashes_pinned));
    } else {
        b->hashes_pinned = 0;
        b->hashes = static_cast<uint64_t*>(std::malloc(bytes));
    }
    if (!b->hashes) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
    // payload: parallel pread into the (pinned) buffer, 64 MB pieces
    const uint64_t total = h.n_hashes * sizeof(uint64_t), piece = uint64_t(64) << 20;
    const int64_t pieces = static_cast<int64_t>((total + piece - 1) / piece);
    int failed = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(std::min(8, omp_get_max_threads()))
    for (int64_t k = 0; k < pieces; ++k) {
        const uint64_t at = static_cast<uint64_t>(k) * piece, len = std::min(piece, total - at);
        if (!pread_all(fd, reinterpret_cast<char*>(b->hashes) + at, len, h.payload_offset + at)) {
#pragma omp atomic write
            failed = 1;
        }
    }
    if (failed) { mvs_free_hash_batch(b); return MVS_ERR_IO; }
    *out = b;
    return MVS_OK;
}

int mvs_write_hash_batch(const mvs_hash_batch* b, const char* path)
{
    if (!b || !path || b->n_samples < 0 || !b->offsets || (!b->hashes && b->n_hashes > 0)) return MVS_ERR_INVALID_ARG;
    const uint64_t n = static_cast<uint64_t>(b->n_samples);
    if (b->offsets[0] != 0 || b->offsets[n] != b->n_hashes) return MVS_ERR_INVALID_ARG;
    BinHeader h{};
    std::memcpy(h.magic, kBinMagic, sizeof kBinMagic);
    h.n_samples = n;
    h.n_hashes = b->n_hashes;
    std::string names;
    if (b->names)
        for (uint64_t i = 0; i < n; ++i) names.append(b->names[i] ? b->names[i] : "").push_back('\0');
    h.names_bytes = names.size();
    const uint64_t head = kBinHeader + (n + 1) * sizeof(uint64_t) + names.size();
    h.payload_offset = (head + kBinAlign - 1) / kBinAlign * kBinAlign;
    const int fd = ::open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return MVS_ERR_IO;
    const std::string pad(static_cast<size_t>(h.payload_offset - head), '\0');
    const bool ok = write_all(fd, &h, sizeof h) && write_all(fd, b->offsets, (n + 1) * sizeof(uint64_t)) &&
                    write_all(fd, names.data(), names.size()) && write_all(fd, pad.data(), pad.size()) &&
                    write_all(fd, b->hashes, b->n_hashes * sizeof(uint64_t));
    const bool closed = ::close(fd) == 0;
    return ok && closed ? MVS_OK : MVS_ERR_IO;
}

static int parse_hash_file_impl(const char* path, int named, bool want_pinned, mvs_hash_batch** out);

int mvs_parse_hash_file(const char* path, int named, mvs_hash_batch** out)
{
    return parse_hash_file_impl(path, named, true, out);
}

static int parse_hash_file_impl(const char* path, int named, bool want_pinned, mvs_hash_batch** out)
{
    if (!path || !out) return MVS_ERR_INVALID_ARG;
    *out = nullptr;
    Mapped mf(path);
    if (!mf.ok) return MVS_ERR_IO;
    if (mf.size >= kBinHeader && std::memcmp(mf.data, kBinMagic, sizeof kBinMagic) == 0)
        return load_binary_batch(mf.fd, mf.data, mf.size, named, want_pinned, out);

    // 1. line index + an upper bound of every line's token count (both parallel)
    const bool timing = std::getenv("MVS_TIMING") != nullptr;
    const auto tp0 = std::chrono::steady_clock::now();
    std::vector<LineRef> lines;
    index_lines(mf.data, mf.size, named, lines);
    const auto tp1 = std::chrono::steady_clock::now();
    const int64_t n = static_cast<int64_t>(lines.size());

    mvs_hash_batch* b = static_cast<mvs_hash_batch*>(std::calloc(1, sizeof(mvs_hash_batch)));
    if (!b) return MVS_ERR_OOM;
    b->n_samples = n;
    b->offsets = static_cast<uint64_t*>(std::malloc(sizeof(uint64_t) * static_cast<size_t>(n + 1)));
    if (!b->offsets) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
    uint64_t bound_total = 0;
    for (int64_t i = 0; i < n; ++i) {
        b->offsets[i] = bound_total;                                    // tentative: start of the line's slot
        bound_total += lines[static_cast<size_t>(i)].bound;
    }
    b->offsets[n] = bound_total;
    if (want_pinned) {
        b->hashes = static_cast<uint64_t*>(host_alloc_any(sizeof(uint64_t) * (bound_total + 2), &b->hashes_pinned));
    } else {                                     // pageable: no CUDA call, so a context can be created meanwhile
        b->hashes_pinned = 0;
        b->hashes = static_cast<uint64_t*>(std::malloc(sizeof(uint64_t) * (bound_total + 2)));
    }
    if (!b->hashes) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }

    const auto tp2 = std::chrono::steady_clock::now();
    // 2. parse + dedupe every line straight into its slot of the (pinned) CSR payload
    std::vector<uint64_t> count(static_cast<size_t>(n ? n : 1));
    const char* safe_end = mf.data + mf.size;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t i = 0; i < n; ++i) {
        const LineRef& ln = lines[static_cast<size_t>(i)];
        count[static_cast<size_t>(i)] = parse_line_set(ln.b, ln.e, safe_end, b->hashes + b->offsets[i]);
    }

    const auto tp3 = std::chrono::steady_clock::now();
    if (timing)
        std::fprintf(stderr, "mvs_timing parse: index_s=%.4f alloc_s=%.4f tokens_s=%.4f threads=%d\n",
                     std::chrono::duration<double>(tp1 - tp0).count(), std::chrono::duration<double>(tp2 - tp1).count(),
                     std::chrono::duration<double>(tp3 - tp2).count(), omp_get_max_threads());
    // 3. close the gaps left by malformed / duplicated tokens (none for well-formed files)
    uint64_t total = 0;
    for (int64_t i = 0; i < n; ++i) {
        const uint64_t from = b->offsets[i], c = count[static_cast<size_t>(i)];
        if (from != total && c) std::memmove(b->hashes + total, b->hashes + from, c * sizeof(uint64_t));
        b->offsets[i] = total;
        total += c;
    }
    b->offsets[n] = total;
    b->n_hashes = total;
    if (named) {
        b->names = static_cast<char**>(std::calloc(static_cast<size_t>(n ? n : 1), sizeof(char*)));
        if (!b->names) { mvs_free_hash_batch(b); return MVS_ERR_OOM; }
        for (int64_t i = 0; i < n; ++i) {
            const LineRef& ln = lines[static_cast<size_t>(i)];
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
                 const float* ref_norms, const uint64_t* sumsq, const int32_t* vec32, const int16_t* vec16)
{
    if (!index_folder || n_samples < 0 || d <= 0 || (!vec32 && !vec16 && n_samples > 0) ||
        (n_samples > 0 && (!names || (!ref_norms && !sumsq && !vec32))))
        return MVS_ERR_INVALID_ARG;
    std::vector<float> host_norms;
    if (!ref_norms && !sumsq && n_samples > 0) {                        // int32 rows only: the reference's expression on the host
        host_norms.resize(static_cast<size_t>(n_samples));
        const int nrc = mvs_reference_norms(vec32, n_samples, d, host_norms.data());
        if (nrc != MVS_OK) return nrc;
        ref_norms = host_norms.data();
    }
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
            // norm = || v / sqrt(d) ||  (:327-329): the reference's own float32 evaluation widened to double
            // (mvs_norms.cpp), or -- the caller's choice -- from the exact integer sum of squares
            const double norm = ref_norms ? static_cast<double>(ref_norms[i])
                                          : std::sqrt(static_cast<double>(sumsq[i]) / static_cast<double>(d));
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

int mvs_set_norms_mode(int mode)
{
    if (mode != 0 && mode != 1) return MVS_ERR_INVALID_ARG;
    g_norms_mode.store(mode);
    return MVS_OK;
}
int mvs_get_norms_mode(void) { return g_norms_mode.load(); }

// The sample loop of sketch() (:289-298) over several GPUs: contiguous size-balanced shards
// (mvs_partition_samples), one host thread per context, every shard written straight into its own
// row range of the caller's buffers.  No exchange between the GPUs: samples are independent.
int mvs_sketch_batch_multi(mvs_ctx* const* ctxs, int n_ctx, const uint64_t* hashes, const uint64_t* offsets,
                           int64_t n_samples, int d, uint64_t seed, int32_t* out, int16_t* out16, uint64_t* sumsq,
                           float* ref_norms)
{
    if (!ctxs || n_ctx <= 0 || !offsets || n_samples < 0 || d <= 0) return MVS_ERR_INVALID_ARG;
    for (int k = 0; k < n_ctx; ++k)
        if (!ctxs[k]) return MVS_ERR_INVALID_ARG;
    if (n_samples == 0) return MVS_OK;
    if (!out && !out16 && !sumsq && !ref_norms) return MVS_ERR_INVALID_ARG;
    for (int64_t s = 0; s < n_samples; ++s)
        if (offsets[s + 1] < offsets[s]) return MVS_ERR_INVALID_ARG;
    const int parts = static_cast<int>(std::min<int64_t>(n_ctx, n_samples));
    std::vector<int64_t> bounds(static_cast<size_t>(parts) + 1);
    mvs_partition_samples(offsets, n_samples, parts, bounds.data());
    std::vector<int> rcs(static_cast<size_t>(parts), MVS_OK);
    auto work = [&](int k) {
        const int64_t a = bounds[static_cast<size_t>(k)], b = bounds[static_cast<size_t>(k) + 1];
        if (b <= a) return;
        const size_t row0 = static_cast<size_t>(a) * static_cast<size_t>(d);
        rcs[static_cast<size_t>(k)] =
            mvs_sketch_batch(ctxs[k], hashes, offsets + a, b - a, d, seed, out ? out + row0 : nullptr,
                             out16 ? out16 + row0 : nullptr, sumsq ? sumsq + a : nullptr,
                             ref_norms ? ref_norms + a : nullptr);
    };
    std::vector<std::thread> th;
    for (int k = 1; k < parts; ++k) th.emplace_back(work, k);          // one host thread per GPU
    work(0);
    for (auto& t : th) t.join();
    for (int r : rcs)
        if (r != MVS_OK) return r;
    return MVS_OK;
}

// sketch() after the parse: project hb on the given contexts, print the reference's lines, write db/.
// t0 = start of the reference's timer (:255, includes the parse).  norms_mode 0 = the reference's float32
// norm (vector_norms.txt identical to the reference's on this host), 1 = sqrt(exact sum of squares / d).
static int sketch_parsed(mvs_ctx* const* ctxs, int n_ctx, mvs_hash_batch* hb, const std::string& dir, int d,
                         int use_int16, uint64_t seed, int verbose, std::chrono::steady_clock::time_point t0,
                         double parse_s, double wait_ctx_s)
{
    const auto t_start = std::chrono::steady_clock::now();
    int rc = MVS_OK;
    const int64_t n = hb->n_samples;
    const size_t count = static_cast<size_t>(n) * static_cast<size_t>(d);
    int pin32 = 0, pin16 = 0;
    int32_t* v32 = use_int16 ? nullptr : static_cast<int32_t*>(host_alloc_any(count * sizeof(int32_t), &pin32));
    int16_t* v16 = use_int16 ? static_cast<int16_t*>(host_alloc_any(count * sizeof(int16_t), &pin16)) : nullptr;
    const bool exact_norms = g_norms_mode.load() == 1;
    std::vector<uint64_t> sumsq(exact_norms ? static_cast<size_t>(n ? n : 1) : 0);
    std::vector<float> ref_norms(exact_norms ? 0 : static_cast<size_t>(n ? n : 1));
    if ((!use_int16 && !v32) || (use_int16 && !v16)) rc = MVS_ERR_OOM;

    if (rc == MVS_OK && n > 0)
        rc = mvs_sketch_batch_multi(ctxs, n_ctx, hb->hashes, hb->offsets, n, d, seed, v32, v16,
                                    exact_norms ? sumsq.data() : nullptr, exact_norms ? nullptr : ref_norms.data());
    if (rc == MVS_OK) {
        if (verbose)
            for (int64_t i = 0; i < n; ++i)                             // :294-297 (order is arbitrary there)
                std::printf("Projected %s, vector dimension %d, index %lld\n", hb->names[i], d,
                            static_cast<long long>(i));
        const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::printf("Time to compute all projected vectors: %g seconds\n", sec);   // :300-303
        const auto t_projected = std::chrono::steady_clock::now();
        rc = mvs_write_db(dir.c_str(), n, d, hb->names, exact_norms ? nullptr : ref_norms.data(),
                          exact_norms ? sumsq.data() : nullptr, v32, v16);
        if (rc != MVS_OK) std::fprintf(stderr, "mvs: writing %s failed: %s\n", dir.c_str(), mvs_strerror(rc));
        if (std::getenv("MVS_TIMING")) {                                // stage split for tests/perf_cli.py
            const auto t_written = std::chrono::steady_clock::now();
            std::fprintf(stderr,
                         "mvs_timing parse_s=%.6f wait_ctx_s=%.6f project_s=%.6f write_s=%.6f samples=%lld hashes=%llu\n",
                         parse_s, wait_ctx_s, std::chrono::duration<double>(t_projected - t_start).count(),
                         std::chrono::duration<double>(t_written - t_projected).count(), static_cast<long long>(n),
                         static_cast<unsigned long long>(hb->n_hashes));
        }
    } else {
        std::fprintf(stderr, "mvs: projection failed: %s\n", mvs_strerror(rc));
    }
    host_free_any(v32, pin32);
    host_free_any(v16, pin16);
    return rc;
}

static std::string prepare_index_folder(const char* index_folder)
{
    std::string dir(index_folder);
    if (dir.empty() || dir.back() != '/') dir += '/';                   // :239-241
    std::error_code ec;
    if (fs::exists(dir, ec)) {                                          // :244-252 (wipes the folder)
        for (const auto& entry : fs::directory_iterator(dir, ec)) fs::remove_all(entry.path(), ec);
    } else {
        fs::create_directories(dir, ec);
    }
    return dir;
}

int mvs_sketch_file(mvs_ctx* const* ctxs, int n_ctx, const char* hash_file, const char* index_folder, int d,
                    int use_int16, uint64_t seed, int verbose)
{
    if (!ctxs || n_ctx <= 0 || !hash_file || !index_folder || d <= 0) return MVS_ERR_INVALID_ARG;
    const std::string dir = prepare_index_folder(index_folder);
    const auto t0 = std::chrono::steady_clock::now();                   // :255 timer includes the parse
    mvs_hash_batch* hb = nullptr;
    int rc = mvs_parse_hash_file(hash_file, 1, &hb);
    if (rc != MVS_OK) {
        std::fprintf(stderr, "Error opening %s for reading.\n", hash_file);   // :260
        return rc;
    }
    if (verbose) std::printf("Loaded %lld hash sets from %s\n", static_cast<long long>(hb->n_samples), hash_file);
    const double parse_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    rc = sketch_parsed(ctxs, n_ctx, hb, dir, d, use_int16, seed, verbose, t0, parse_s, 0.0);
    mvs_free_hash_batch(hb);
    return rc;
}

// The CLI's entry: same as mvs_sketch_file, but the contexts of `devices` are created on a second
// host thread WHILE the hash file is parsed (CUDA start-up costs about a second on an 8-GPU host,
// the parse of config 2's text about as much).  The parse then cannot use pinned memory (that needs
// a context), so the payload is page-locked in place (cudaHostRegister) once the context is up.
int mvs_sketch_file_on_devices(const int* devices, int n_dev, const char* hash_file, const char* index_folder, int d,
                               int use_int16, uint64_t seed, int verbose)
{
    if (!devices || n_dev <= 0 || !hash_file || !index_folder || d <= 0) return MVS_ERR_INVALID_ARG;
    std::vector<mvs_ctx*> ctxs(static_cast<size_t>(n_dev), nullptr);
    int create_rc = MVS_OK;
    std::vector<int> create_rcs(static_cast<size_t>(n_dev), MVS_OK);
    std::thread creator([&] {                                          // one creating thread per device
        std::vector<std::thread> th;
        for (int k = 1; k < n_dev; ++k)
            th.emplace_back([&, k] { create_rcs[static_cast<size_t>(k)] = mvs_create(&ctxs[static_cast<size_t>(k)], devices[k]); });
        create_rcs[0] = mvs_create(&ctxs[0], devices[0]);
        for (auto& t : th) t.join();
        for (int r : create_rcs)
            if (r != MVS_OK && create_rc == MVS_OK) create_rc = r;
    });
    const auto t0 = std::chrono::steady_clock::now();
    mvs_hash_batch* hb = nullptr;
    int rc = parse_hash_file_impl(hash_file, 1, false, &hb);
    const auto t1 = std::chrono::steady_clock::now();
    creator.join();
    const auto t2 = std::chrono::steady_clock::now();
    if (rc != MVS_OK && create_rc == MVS_OK) {
        std::fprintf(stderr, "Error opening %s for reading.\n", hash_file);   // :260
        prepare_index_folder(index_folder);                            // the reference has wiped it by then
    }
    if (create_rc != MVS_OK) rc = create_rc;                            // no usable GPU: index_folder stays untouched
    if (rc == MVS_OK) {
        const std::string dir = prepare_index_folder(index_folder);    // the reference wipes it before the parse (:244)
        if (verbose) std::printf("Loaded %lld hash sets from %s\n", static_cast<long long>(hb->n_samples), hash_file);
        const auto t3 = std::chrono::steady_clock::now();
        const int registered = hb->n_hashes ? mvs_host_register(hb->hashes, sizeof(uint64_t) * hb->n_hashes) : -1;
        if (std::getenv("MVS_TIMING"))
            std::fprintf(stderr, "mvs_timing host_register_s=%.6f\n",
                         std::chrono::duration<double>(std::chrono::steady_clock::now() - t3).count());
        rc = sketch_parsed(ctxs.data(), n_dev, hb, dir, d, use_int16, seed, verbose, t0,
                           std::chrono::duration<double>(t1 - t0).count(), std::chrono::duration<double>(t2 - t1).count());
        if (registered == MVS_OK) mvs_host_unregister(hb->hashes);
    }
    mvs_free_hash_batch(hb);
    for (mvs_ctx* c : ctxs) mvs_destroy(c);
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

// ---- streaming copy for the pinned staging ring (mvs_internal.h) -------------------------------------------
#include <immintrin.h>

__attribute__((target("avx2"))) static void copy_streaming_avx2(char* dst, const char* src, size_t n)
{
    size_t head = (32 - (reinterpret_cast<uintptr_t>(dst) & 31)) & 31;
    if (head > n) head = n;
    std::memcpy(dst, src, head);
    size_t i = head;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i));
        const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 32));
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 64));
        const __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 96));
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i), a);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 32), b);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 64), c);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 96), d);
    }
    std::memcpy(dst + i, src + i, n - i);
    _mm_sfence();
}

void mvs_internal_copy_streaming(void* dst, const void* src, size_t bytes)
{
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) copy_streaming_avx2(static_cast<char*>(dst), static_cast<const char*>(src), bytes);
    else std::memcpy(dst, src, bytes);
}
