// mvs_convert.cpp -- `project_everything convert` without shelling out.
//
// Reference: src/project_everything.cpp:73-235 (convert / load_signatures / read_gzipped_file), which
// runs `unzip`, `gunzip` and `rm -rf` through system() and scans the JSON text by hand.  Here the zip
// central directory is walked in memory, members are inflated with zlib, and the same text scan is
// applied: first "ksize" must be 31, then the first "mins":[...] array, split at ',' and parsed like
// std::stoull (errors ignored).  Differences, all deliberate: input files are processed in sorted
// file-name order (the reference uses directory order, which is file-system dependent), hashes are
// written sorted instead of in unordered_set order (the set is the same), and the side-effect file
// ./all_hashes.txt (reference :160-176) is not written.
#include "../../include/mvs_b200.h"

#include <zlib.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <string>
#include <vector>

namespace fs = std::filesystem;

namespace {

inline uint32_t rd16(const unsigned char* p) { return p[0] | (p[1] << 8); }
inline uint32_t rd32(const unsigned char* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | (static_cast<uint32_t>(p[3]) << 24); }

bool inflate_buf(const unsigned char* src, size_t n, int window_bits, std::string& out)
{
    z_stream zs;
    std::memset(&zs, 0, sizeof zs);
    if (inflateInit2(&zs, window_bits) != Z_OK) return false;
    zs.next_in = const_cast<unsigned char*>(src);
    zs.avail_in = static_cast<uInt>(n);
    out.clear();
    char buf[1 << 16];
    int rc = Z_OK;
    while (rc != Z_STREAM_END) {
        zs.next_out = reinterpret_cast<unsigned char*>(buf);
        zs.avail_out = sizeof buf;
        rc = inflate(&zs, Z_NO_FLUSH);
        if (rc != Z_OK && rc != Z_STREAM_END) {
            inflateEnd(&zs);
            return false;
        }
        out.append(buf, sizeof buf - zs.avail_out);
        if (rc == Z_OK && zs.avail_in == 0 && zs.avail_out != 0) break;     // truncated input
    }
    inflateEnd(&zs);
    return rc == Z_STREAM_END;
}

struct ZipMember {
    std::string name;
    std::string data;                            // uncompressed member bytes
};

// Members of a zip archive held in memory (no zip64, no encryption: sourmash writes plain archives).
bool read_zip(const std::string& path, std::vector<ZipMember>& members)
{
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    std::string z((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    const unsigned char* b = reinterpret_cast<const unsigned char*>(z.data());
    const size_t n = z.size();
    if (n < 22) return false;
    size_t eocd = std::string::npos;
    for (size_t i = n - 22;; --i) {              // end-of-central-directory signature PK\5\6
        if (rd32(b + i) == 0x06054b50u) { eocd = i; break; }
        if (i == 0 || n - i > 22 + 65535) break;
    }
    if (eocd == std::string::npos) return false;
    const uint32_t count = rd16(b + eocd + 10);
    size_t p = rd32(b + eocd + 16);
    for (uint32_t k = 0; k < count; ++k) {
        if (p + 46 > n || rd32(b + p) != 0x02014b50u) return false;
        const uint32_t method = rd16(b + p + 10);
        const uint32_t csize = rd32(b + p + 20);
        const uint32_t nlen = rd16(b + p + 28), xlen = rd16(b + p + 30), clen = rd16(b + p + 32);
        const size_t lho = rd32(b + p + 42);
        if (p + 46 + nlen > n) return false;
        ZipMember m;
        m.name.assign(z, p + 46, nlen);
        p += 46 + nlen + xlen + clen;
        if (lho + 30 > n || rd32(b + lho) != 0x04034b50u) return false;
        const size_t data = lho + 30 + rd16(b + lho + 26) + rd16(b + lho + 28);
        if (data + csize > n) return false;
        if (method == 0) m.data.assign(z, data, csize);
        else if (method == 8) {
            if (!inflate_buf(b + data, csize, -15, m.data)) return false;
        } else return false;
        members.push_back(std::move(m));
    }
    return true;
}

// std::stoull(num_str) after whitespace removal (reference :141-148): optional sign, digits, trailing
// garbage ignored; no digits or out of range -> exception -> value ignored.
bool stoull_like(const std::string& s, uint64_t& v)
{
    size_t i = 0;
    bool neg = false;
    if (i < s.size() && (s[i] == '+' || s[i] == '-')) neg = s[i++] == '-';
    if (i >= s.size() || s[i] < '0' || s[i] > '9') return false;
    uint64_t x = 0;
    for (; i < s.size() && s[i] >= '0' && s[i] <= '9'; ++i) {
        const uint64_t dgt = static_cast<uint64_t>(s[i] - '0');
        if (x > (UINT64_MAX - dgt) / 10) return false;
        x = x * 10 + dgt;
    }
    v = neg ? (0 - x) : x;
    return true;
}

inline bool c_isspace(char c) { return c == ' ' || (c >= '\t' && c <= '\r'); }

// The text scan of load_signatures (reference :109-150) on one decompressed .sig JSON string.
void scan_signature_json(const std::string& js, std::vector<uint64_t>& hashes)
{
    const size_t ksize_pos = js.find("\"ksize\"");
    if (ksize_pos == std::string::npos) return;
    const size_t colon = js.find(':', ksize_pos);
    if (colon == std::string::npos) return;
    const size_t kend = js.find_first_of(",}", colon);
    std::string ks = js.substr(colon + 1, kend == std::string::npos ? std::string::npos : kend - colon - 1);
    ks.erase(std::remove_if(ks.begin(), ks.end(), c_isspace), ks.end());
    if (ks != "31") return;
    const size_t mins = js.find("\"mins\"");
    if (mins == std::string::npos) return;
    const size_t a0 = js.find('[', mins);
    if (a0 == std::string::npos) return;
    const size_t a1 = js.find(']', a0);
    if (a1 == std::string::npos) return;
    size_t pos = a0 + 1;
    std::string tok;
    while (pos < a1) {
        size_t comma = js.find(',', pos);
        if (comma == std::string::npos || comma > a1) comma = a1;
        tok.assign(js, pos, comma - pos);
        tok.erase(std::remove_if(tok.begin(), tok.end(), c_isspace), tok.end());
        uint64_t v;
        if (!tok.empty() && stoull_like(tok, v)) hashes.push_back(v);
        pos = comma + 1;
    }
}

}  // namespace

extern "C" int mvs_convert_signatures(const char* signature_folder, const char* hash_file, int threads, int verbose)
{
    if (!signature_folder || !hash_file) return MVS_ERR_INVALID_ARG;
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::string> files;
    std::error_code ec;
    for (const auto& entry : fs::directory_iterator(signature_folder, ec)) files.push_back(entry.path().string());
    if (ec) {
        std::fprintf(stderr, "Error listing %s\n", signature_folder);
        return MVS_ERR_IO;
    }
    std::sort(files.begin(), files.end());
    FILE* fo = std::fopen(hash_file, "w");
    if (!fo) {
        std::fprintf(stderr, "Error opening %s for writing.\n", hash_file);      // reference :194
        return MVS_ERR_IO;
    }
    const int64_t n = static_cast<int64_t>(files.size());
    std::vector<std::vector<uint64_t>> sets(static_cast<size_t>(n));
    std::vector<std::string> names(static_cast<size_t>(n));
    (void)threads;
#pragma omp parallel for schedule(dynamic) num_threads(threads > 0 ? threads : 1)
    for (int64_t i = 0; i < n; ++i) {
        std::vector<uint64_t>& hs = sets[static_cast<size_t>(i)];
        std::vector<ZipMember> members;
        if (!read_zip(files[static_cast<size_t>(i)], members)) {
#pragma omp critical
            std::fprintf(stderr, "Failed to unzip: %s\n", files[static_cast<size_t>(i)].c_str());   // reference :100
        }
        for (const ZipMember& m : members) {
            // reference :103-105: every *.gz directly inside signatures/
            if (m.name.rfind("signatures/", 0) != 0 || m.name.size() < 3 ||
                m.name.compare(m.name.size() - 3, 3, ".gz") != 0)
                continue;
            std::string js;
            if (!inflate_buf(reinterpret_cast<const unsigned char*>(m.data.data()), m.data.size(), 16 + 15, js)) {
#pragma omp critical
                std::fprintf(stderr, "Error running gunzip on: %s\n", m.name.c_str());                // reference :80
                continue;
            }
            scan_signature_json(js, hs);
        }
        std::sort(hs.begin(), hs.end());
        hs.erase(std::unique(hs.begin(), hs.end()), hs.end());
        const std::string stem = fs::path(files[static_cast<size_t>(i)]).stem().string();
        names[static_cast<size_t>(i)] = stem.substr(0, stem.find('.'));                               // reference :210-211
        if (verbose) {
#pragma omp critical
            std::printf("Processed %s, hashes size %zu, file number %lld\n", files[static_cast<size_t>(i)].c_str(),
                        hs.size(), static_cast<long long>(i));
        }
    }
    int rc = MVS_OK;
    std::string line;
    for (int64_t i = 0; i < n; ++i) {                                                                 // reference :222-229
        line = names[static_cast<size_t>(i)];
        line += ':';
        char buf[24];
        for (uint64_t h : sets[static_cast<size_t>(i)]) {
            const int len = std::snprintf(buf, sizeof buf, " %llu", static_cast<unsigned long long>(h));
            line.append(buf, static_cast<size_t>(len));
        }
        line += '\n';
        if (std::fwrite(line.data(), 1, line.size(), fo) != line.size()) rc = MVS_ERR_IO;
    }
    if (std::fclose(fo) != 0) rc = MVS_ERR_IO;
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("Time to convert all signatures: %g seconds\n", sec);                                 // reference :231-234
    return rc;
}
