// succinct_harness.cpp -- the in-tree shard-file writers (csrc/mvs_succinct.h) behind the same C interface as
// oracle/ref_bits_shim.cpp; compiled by tests/test_succinct.py at test time (host compiler only, no GPU).
#include "../metagenome_vector_sketches_b200/csrc/mvs_succinct.h"

#include <cstdint>
#include <cstring>
#include <sstream>
#include <string>

namespace {
uint64_t emit(const std::string& s, unsigned char* out, uint64_t cap)
{
    if (out && cap) std::memcpy(out, s.data(), s.size() < cap ? s.size() : cap);
    return s.size();
}
}  // namespace

extern "C" {

uint64_t mvs_compact_vector(const uint64_t* v, uint64_t n, unsigned char* out, uint64_t cap, uint64_t* num_bytes)
{
    mvs_succinct::CompactVector cv;
    cv.build(v, n);
    std::ostringstream os(std::ios::binary);
    cv.save(os);
    if (num_bytes) *num_bytes = cv.num_bytes();
    return emit(os.str(), out, cap);
}

// the element types the shard writer feeds: u16 Jaccard codes and u32 row ids go through the same template
uint64_t mvs_compact_vector_u16(const uint16_t* v, uint64_t n, unsigned char* out, uint64_t cap, uint64_t* num_bytes)
{
    mvs_succinct::CompactVector cv;
    cv.build(v, n);
    std::ostringstream os(std::ios::binary);
    cv.save(os);
    if (num_bytes) *num_bytes = cv.num_bytes();
    return emit(os.str(), out, cap);
}

uint64_t mvs_rice_sequence(const uint64_t* v, uint64_t n, unsigned char* out, uint64_t cap, uint64_t* num_bytes)
{
    mvs_succinct::RiceSequence rs;
    rs.encode(v, n);
    std::ostringstream os(std::ios::binary);
    rs.save(os);
    if (num_bytes) *num_bytes = rs.num_bytes();
    return emit(os.str(), out, cap);
}

}  // extern "C"
