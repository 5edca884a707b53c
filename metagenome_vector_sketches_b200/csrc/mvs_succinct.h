// mvs_succinct.h -- the two succinct containers of the all-pairs shard files, as WRITERS, in-tree.
//
// The reference writes shard_K/{matrix.bin, row_index.bin, neighbor_start.bin} (src/pairwise_comp_optimized.cpp:645-817)
// with its vendored third-party library `bits` (jermp/bits, bits/include/compact_vector.hpp and rice_sequence.hpp, which
// pull in bit_vector.hpp, darray.hpp and essentials.hpp).  pairwise_comp_b200 used to compile against those headers,
// so it could only be built where the reference tree is present.  This header restates the on-disk format of the two
// containers -- nothing else of the library: no access, no select, no loading -- so that the tool builds from this
// repository alone.  tests/test_succinct.py pins every byte against the reference's headers (golden digests and, where
// the reference tree is present, a live comparison) and tests/test_gpu_cli.py compares whole shard files with the
// reference binary's.
//
// Format (all integers little endian, native widths; a "vec" is a u64 element count followed by the raw elements):
//   compact vector   u64 size | u64 width | u64 mask | vec<u64> words            (compact_vector.hpp: save)
//                    value i occupies bits [i*width, (i+1)*width) of the word array, which has
//                    ceil(size*width / 64) + 1 words (one spare word for unaligned reads)
//   bit vector       u64 num_bits | vec<u64> words                                (bit_vector.hpp: save)
//   darray1          u64 ones | vec<i64> block_inventory | vec<u16> subblock_inventory | vec<u64> overflow_positions
//                    (darray.hpp: the select index over the 1 bits, blocks of 1024 ones, sub-blocks of 32)
//   rice sequence    bit vector (unary high parts, each terminated by a 1) | darray1 over it | compact vector of the
//                    low parts, width l = the Rice parameter                       (rice_sequence.hpp: encode, save)
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <ostream>
#include <vector>

namespace mvs_succinct {

template <typename T>
inline void put_pod(std::ostream& os, const T& v)
{
    os.write(reinterpret_cast<const char*>(&v), sizeof(T));
}

template <typename T>
inline void put_vec(std::ostream& os, const std::vector<T>& v)
{
    const uint64_t n = v.size();                 // size_t on the LP64 hosts the reference runs on
    put_pod(os, n);
    os.write(reinterpret_cast<const char*>(v.data()), static_cast<std::streamsize>(sizeof(T) * n));
}

template <typename T>
inline uint64_t vec_bytes(const std::vector<T>& v)
{
    return v.size() * sizeof(T) + sizeof(uint64_t);
}

// ---- fixed-width packed integers (bits/include/compact_vector.hpp) ---------------------------------------------------
class CompactVector {
public:
    void resize(uint64_t n, uint64_t width)
    {
        size_ = n;
        width_ = width;
        mask_ = width == 64 ? ~uint64_t(0) : (uint64_t(1) << width) - 1;
        words_.assign((n * width + 63) / 64 + 1, 0);
    }

    void set(uint64_t i, uint64_t v)             // width > 0, v < 2^width
    {
        const uint64_t pos = i * width_, block = pos >> 6, shift = pos & 63;
        words_[block] &= ~(mask_ << shift);
        words_[block] |= v << shift;
        const uint64_t room = 64 - shift;
        if (room < width_) {
            words_[block + 1] &= ~(mask_ >> room);
            words_[block + 1] |= v >> room;
        }
    }

    // width from the largest value, as compact_vector::build(begin, n) chooses it: ceil(log2(max + 1)) evaluated in
    // double precision, 1 for an all-zero sequence.  (An empty sequence is undefined in the library; here: width 1.)
    template <typename It>
    void build(It begin, uint64_t n)
    {
        uint64_t max = 0;
        It it = begin;
        for (uint64_t i = 0; i < n; ++i, ++it) max = std::max<uint64_t>(max, static_cast<uint64_t>(*it));
        const uint64_t width = max == 0 ? 1 : static_cast<uint64_t>(std::ceil(std::log2(static_cast<double>(max + 1))));
        resize(n, width);
        it = begin;
        for (uint64_t i = 0; i < n; ++i, ++it) set(i, static_cast<uint64_t>(*it));
    }

    uint64_t size() const { return size_; }
    uint64_t width() const { return width_; }
    uint64_t num_bytes() const { return 3 * sizeof(uint64_t) + vec_bytes(words_); }

    void save(std::ostream& os) const
    {
        put_pod(os, size_);
        put_pod(os, width_);
        put_pod(os, mask_);
        put_vec(os, words_);
    }

private:
    uint64_t size_ = 0, width_ = 0, mask_ = 0;
    std::vector<uint64_t> words_;
};

// ---- Rice-coded sequence (bits/include/rice_sequence.hpp with darray1 over the unary parts) --------------------------
class RiceSequence {
public:
    // Rice parameter by Kiely's rule (rice_sequence.hpp: optimal_parameter_kiely), same double arithmetic:
    //   p = n / (sum + n),   l = max(0, 1 + floor(log2(log(phi - 1) / log(1 - p))))
    template <typename It>
    static uint64_t parameter(It begin, uint64_t n)
    {
        uint64_t sum = 0;
        It it = begin;
        for (uint64_t i = 0; i < n; ++i, ++it) sum += static_cast<uint64_t>(*it);
        const double p = static_cast<double>(n) / (static_cast<double>(sum) + static_cast<double>(n));
        const double gold = (std::sqrt(5.0) + 1.0) / 2.0;
        const double l = 1.0 + std::floor(std::log2(std::log(gold - 1.0) / std::log(1.0 - p)));
        // an all-zero sequence gives p = 1 and l = -inf, which the library converts to int64 (INT64_MIN on x86) before
        // clamping at 0; the comparison below is the same result without the undefined conversion
        return l > 0.0 ? static_cast<uint64_t>(l) : 0;
    }

    template <typename It>
    void encode(It begin, uint64_t n)
    {
        if (n == 0) return;                      // the library leaves all three members default-constructed
        const uint64_t l = parameter(begin, n);
        low_.resize(n, l);
        const uint64_t low_mask = (uint64_t(1) << l) - 1;
        It it = begin;
        for (uint64_t i = 0; i < n; ++i, ++it) {
            const uint64_t v = static_cast<uint64_t>(*it);
            if (l > 0) low_.set(i, v & low_mask);
            num_bits_ += v >> l;                 // the zeros of the unary part
            push_one();
        }
        index_ones();
    }

    uint64_t size() const { return low_.size(); }

    uint64_t num_bytes() const
    {
        return sizeof(uint64_t) + vec_bytes(high_) +                                                       // bit vector
               sizeof(uint64_t) + vec_bytes(block_inventory_) + vec_bytes(subblock_inventory_) + vec_bytes(overflow_) +
               low_.num_bytes();
    }

    void save(std::ostream& os) const
    {
        put_pod(os, num_bits_);
        put_vec(os, high_);
        put_pod(os, ones_);
        put_vec(os, block_inventory_);
        put_vec(os, subblock_inventory_);
        put_vec(os, overflow_);
        low_.save(os);
    }

private:
    static constexpr uint64_t kBlock = 1024, kSubblock = 32;

    void push_one()                              // zeros are appended by advancing num_bits_; words appear on demand
    {
        const uint64_t word = num_bits_ >> 6;
        if (high_.size() <= word) high_.resize(word + 1, 0);
        high_[word] |= uint64_t(1) << (num_bits_ & 63);
        ++num_bits_;
    }

    // darray1::build: positions of the ones in blocks of 1024; a block spanning < 2^16 bits is "dense" (its start in
    // block_inventory, every 32nd position relative to it in subblock_inventory), otherwise "sparse" (all positions
    // verbatim in overflow_positions, block_inventory = -(first overflow index) - 1, subblock entries 0xFFFF)
    void index_ones()
    {
        std::vector<uint64_t> cur;
        cur.reserve(kBlock);
        auto flush = [&]() {
            if (cur.back() - cur.front() < (uint64_t(1) << 16)) {
                block_inventory_.push_back(static_cast<int64_t>(cur.front()));
                for (size_t i = 0; i < cur.size(); i += kSubblock)
                    subblock_inventory_.push_back(static_cast<uint16_t>(cur[i] - cur.front()));
            } else {
                block_inventory_.push_back(-static_cast<int64_t>(overflow_.size()) - 1);
                overflow_.insert(overflow_.end(), cur.begin(), cur.end());
                for (size_t i = 0; i < cur.size(); i += kSubblock) subblock_inventory_.push_back(static_cast<uint16_t>(-1));
            }
            cur.clear();
        };
        for (uint64_t w = 0; w < high_.size(); ++w) {
            uint64_t word = high_[w];
            while (word) {
                cur.push_back((w << 6) + static_cast<uint64_t>(__builtin_ctzll(word)));
                word &= word - 1;
                ++ones_;
                if (cur.size() == kBlock) flush();
            }
        }
        if (!cur.empty()) flush();
    }

    uint64_t num_bits_ = 0, ones_ = 0;
    std::vector<uint64_t> high_;
    std::vector<int64_t> block_inventory_;
    std::vector<uint16_t> subblock_inventory_;
    std::vector<uint64_t> overflow_;
    CompactVector low_;
};

}  // namespace mvs_succinct
