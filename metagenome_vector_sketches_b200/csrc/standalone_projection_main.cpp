// standalone_projection <hashes_file> <dimension>  -- drop-in for the reference tool
// (reference src/standalone_projection.cpp:11-46; called by jaccard.py:98-103, which expects
// the binary next to jaccard.py).  Same argv, same stdout format, same exit codes.
#include "../../include/mvs_b200.h"

#include <cstdio>
#include <cstdlib>
#include <string>

int main(int argc, char** argv)
{
    if (argc < 3) {
        std::fprintf(stderr, "Usage: %s <hashes_file> <dimension>\n", argv[0]);
        return 1;
    }
    int d = 0;
    try {
        d = std::stoi(argv[2]);                  // the reference throws (abort) on a bad number too
    } catch (...) {
        std::fprintf(stderr, "terminate: invalid dimension '%s'\n", argv[2]);
        return 134;
    }
    unsigned long long seed = 0;                 // additive extension: optional 3rd argument
    if (argc >= 4) seed = std::strtoull(argv[3], nullptr, 10);
    mvs_ctx* ctx = nullptr;
    int rc = mvs_create(&ctx, 0);
    if (rc != MVS_OK) {
        std::fprintf(stderr, "standalone_projection: %s\n", mvs_strerror(rc));
        return 2;
    }
    rc = mvs_standalone_file(ctx, argv[1], d, seed, nullptr);
    mvs_destroy(ctx);
    if (rc == MVS_ERR_IO) return 1;              // "Error opening file" -> exit 1 like the reference
    if (rc != MVS_OK) {
        std::fprintf(stderr, "standalone_projection: %s\n", mvs_strerror(rc));
        return 2;
    }
    return 0;
}
