// ref_bits_shim.cpp -- TEST INFRASTRUCTURE ONLY (see oracle/Makefile): the reference's vendored `bits` containers
// behind a C interface, compiled from the headers where they lie under $(REF)/bits (nothing is copied).  Used by
// tests/test_succinct.py and tests/golden/make_golden_succinct.py to pin the in-tree writers of
// metagenome_vector_sketches_b200/csrc/mvs_succinct.h byte for byte.
//
//   ref_compact_vector / ref_rice_sequence: build the container from n values the way the reference's shard writer
//   does (src/pairwise_comp_optimized.cpp:700-716, :775-807: compact_vector::build(begin, n), rice_sequence<>::encode
//   (begin, n)), serialise it with its own save(), copy at most `cap` bytes to `out`, return the serialised length and
//   the container's num_bytes().
#include "compact_vector.hpp"
#include "rice_sequence.hpp"

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

uint64_t ref_compact_vector(const uint64_t* v, uint64_t n, unsigned char* out, uint64_t cap, uint64_t* num_bytes)
{
    bits::compact_vector cv;
    cv.build(v, n);
    std::ostringstream os(std::ios::binary);
    cv.save(os);
    if (num_bytes) *num_bytes = cv.num_bytes();
    return emit(os.str(), out, cap);
}

uint64_t ref_rice_sequence(const uint64_t* v, uint64_t n, unsigned char* out, uint64_t cap, uint64_t* num_bytes)
{
    bits::rice_sequence<> rs;
    rs.encode(v, n);
    std::ostringstream os(std::ios::binary);
    rs.save(os);
    if (num_bytes) *num_bytes = rs.num_bytes();
    return emit(os.str(), out, cap);
}

}  // extern "C"
