// project_everything -- drop-in for the reference CLI (reference src/project_everything.cpp:364-417).
//
//   project_everything sketch  <hash_file> <index_folder> [-t N] [-d D] [--int16]     (reference grammar)
//   project_everything convert <signature_folder> <hash_file> [-t N]
//   project_everything pack    <hash_file> <binary_file>          (additive: the parsed batch in binary form, see
//                                                                   mvs_write_hash_batch; sketch reads either form)
// additive, all optional:      [-s|--seed S] [--gpus G | --devices a,b,..] [--norms reference|exact] [--quiet]
// README / BASELINE spelling:  project_everything <input> <db>/ [-t N] [-d D] [-s S]
//   (<input> a hash file -> sketch; a folder of *.sig.zip -> convert to <db>/hashes.txt.tmp, then sketch)
//
// Exit codes follow the reference: 1 + usage on a parse error, otherwise 0 (the reference returns 0
// even when the hash file cannot be opened); 2 only when no usable GPU exists, which the reference
// cannot encounter.  `-t` is accepted and, as in the reference's sketch mode, unused there.
#include "../../include/mvs_b200.h"

#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

void usage(const char* argv0)
{
    std::fprintf(stderr,
                 "Usage:\n"
                 "  Convert mode:\n"
                 "    %s convert <signature_folder> <hash_file> [-t threads]\n"
                 "      signature_folder : Path to folder containing signature files\n"
                 "      hash_file        : Output hash file path\n"
                 "      -t, --threads    : Number of threads (default: 1)\n\n"
                 "  Pack mode (optional: a binary hash file that sketch reads without parsing text):\n"
                 "    %s pack <hash_file> <binary_file>\n\n"
                 "  Sketch mode:\n"
                 "    %s sketch <hash_file> <index_folder> [-t threads] [-d dimension] [--int16]\n"
                 "      hash_file        : Input hash file path\n"
                 "      index_folder     : Output folder for index files\n"
                 "      -t, --threads    : Number of threads (default: 1)\n"
                 "      -d, --dimension  : Vector dimension (default: 2048)\n"
                 "      --int16          : Use int16 instead of int32 for vector storage\n"
                 "      -s, --seed       : additive generator seed (default 0 = reference)\n"
                 "      --gpus           : number of GPUs to shard samples over (default 1)\n"
                 "      --devices        : explicit CUDA device list, e.g. 0,1,2,3 (one shard per entry)\n"
                 "      --norms          : reference (default: vector_norms.txt as the reference binary prints it\n"
                 "                         on this host) | exact (sqrt of the exact integer sum of squares / d)\n"
                 "      --quiet          : no per-sample progress lines\n",
                 argv0, argv0, argv0);
}

bool parse_int(const char* s, long long& v)
{
    if (!s || !*s) return false;
    char* end = nullptr;
    v = std::strtoll(s, &end, 10);
    return end && *end == '\0';
}

bool is_dir(const std::string& p)
{
    struct stat st;
    return stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}

}  // namespace

int main(int argc, char** argv)
{
    enum { NONE, CONVERT, SKETCH, PACK, ONESHOT } mode = NONE;
    std::vector<std::string> pos;
    long long t = 1, d = 2048, gpus = 1;
    unsigned long long seed = 0;
    bool int16 = false, quiet = false, bad = false;
    std::vector<int> devices;
    int norms_mode = 0;

    int i = 1;
    if (argc > 1 && !std::strcmp(argv[1], "convert")) { mode = CONVERT; i = 2; }
    else if (argc > 1 && !std::strcmp(argv[1], "sketch")) { mode = SKETCH; i = 2; }
    else if (argc > 1 && !std::strcmp(argv[1], "pack")) { mode = PACK; i = 2; }
    else if (argc > 1) mode = ONESHOT;
    for (; i < argc && !bad; ++i) {
        const std::string a = argv[i];
        auto need = [&](long long& dst) {
            if (i + 1 >= argc || !parse_int(argv[i + 1], dst)) bad = true;
            else ++i;
        };
        if (a == "-t" || a == "--threads") need(t);
        else if ((a == "-d" || a == "--dimension") && mode != CONVERT) need(d);
        else if (a == "--int16" && mode != CONVERT) int16 = true;
        else if ((a == "-s" || a == "--seed") && mode != CONVERT) {
            if (i + 1 >= argc) bad = true;
            else {
                char* end = nullptr;
                seed = std::strtoull(argv[++i], &end, 10);
                if (!end || *end) bad = true;
            }
        } else if (a == "--gpus" && mode != CONVERT) need(gpus);
        else if (a == "--devices" && mode != CONVERT) {
            if (i + 1 >= argc) bad = true;
            else {
                const char* p = argv[++i];
                while (*p && !bad) {
                    char* end = nullptr;
                    const long v = std::strtol(p, &end, 10);
                    if (end == p || v < 0 || v > 1023) bad = true;
                    else devices.push_back(static_cast<int>(v));
                    p = (*end == ',') ? end + 1 : end;
                    if (*end && *end != ',') bad = true;
                }
                if (devices.empty()) bad = true;
            }
        } else if (a == "--norms" && mode != CONVERT) {
            if (i + 1 >= argc) bad = true;
            else if (!std::strcmp(argv[i + 1], "reference")) { norms_mode = 0; ++i; }
            else if (!std::strcmp(argv[i + 1], "exact")) { norms_mode = 1; ++i; }
            else bad = true;
        }
        else if (a == "--quiet") quiet = true;
        else if (!a.empty() && a[0] == '-' && a.size() > 1) bad = true;      // unknown option
        else pos.push_back(a);
    }
    if (mode == NONE || bad || pos.size() != 2 || d <= 0 || gpus <= 0 || d > 0x7fffffffLL) {
        usage(argv[0]);
        return 1;
    }

    std::string input = pos[0], output = pos[1];
    if (mode == CONVERT) {
        const int rc = mvs_convert_signatures(input.c_str(), output.c_str(), static_cast<int>(t), quiet ? 0 : 1);
        if (rc != MVS_OK) std::fprintf(stderr, "convert: %s\n", mvs_strerror(rc));
        return 0;
    }
    if (mode == PACK) {                                                         // host only: parse, write the binary form
        mvs_hash_batch* hb = nullptr;
        int rc = mvs_parse_hash_file(input.c_str(), 1, &hb);
        if (rc == MVS_OK) rc = mvs_write_hash_batch(hb, output.c_str());
        if (rc == MVS_OK && !quiet)
            std::printf("Packed %lld samples, %llu hashes into %s\n", static_cast<long long>(hb->n_samples),
                        static_cast<unsigned long long>(hb->n_hashes), output.c_str());
        mvs_free_hash_batch(hb);
        if (rc != MVS_OK) {
            std::fprintf(stderr, "pack: %s\n", mvs_strerror(rc));
            return 2;
        }
        return 0;
    }
    std::string tmp_hash_file;
    if (mode == ONESHOT && is_dir(input)) {                                    // README one-shot spelling
        tmp_hash_file = output + (output.empty() || output.back() == '/' ? "" : "/");
        tmp_hash_file = tmp_hash_file.substr(0, tmp_hash_file.size() - 1) + ".hashes.txt";
        const int rc = mvs_convert_signatures(input.c_str(), tmp_hash_file.c_str(), static_cast<int>(t), quiet ? 0 : 1);
        if (rc != MVS_OK) {
            std::fprintf(stderr, "convert: %s\n", mvs_strerror(rc));
            return 0;
        }
        input = tmp_hash_file;
    }

    if (devices.empty())
        for (long long g = 0; g < gpus; ++g) devices.push_back(static_cast<int>(g));
    mvs_set_norms_mode(norms_mode);
    // contexts are created while the hash file is parsed (CUDA start-up overlaps the parse)
    const int rc = mvs_sketch_file_on_devices(devices.data(), static_cast<int>(devices.size()), input.c_str(),
                                              output.c_str(), static_cast<int>(d), int16 ? 1 : 0, seed, quiet ? 0 : 1);
    if (rc == MVS_ERR_NO_DEVICE) {                                              // index_folder was not touched
        std::fprintf(stderr, "project_everything: %s\n", mvs_strerror(rc));
        return 2;
    }
    if (rc != MVS_OK && rc != MVS_ERR_IO) {
        std::fprintf(stderr, "project_everything: %s\n", mvs_strerror(rc));
        return 2;
    }
    return 0;                                                                   // reference: 0 even on I/O errors
}
