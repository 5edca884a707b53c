// gen_hash_file.cpp -- writes a synthetic hash file ("<name>: h1 h2 ...\n" per sample, the input format of
// `project_everything sketch`, reference src/project_everything.cpp:264-282) with the closed-form generator
// mvs-synth-v1 of metagenome_vector_sketches_b200/synthetic.py.  Test/bench tooling, not product code.
//   g++ -O2 -fopenmp -o gen_hash_file gen_hash_file.cpp ;  ./gen_hash_file out.txt n_samples hashes_per_sample [seed]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

static const uint64_t MASK54 = (1ull << 54) - 1;
static inline uint64_t synth(uint64_t s, uint64_t k, uint64_t seed)
{
    uint64_t x = (s << 24) | k;
    x ^= (seed * 0x9E3779B97F4A7C15ull) & MASK54;
    x = (x * 0x9E3779B97F4A7C15ull) & MASK54;
    x ^= x >> 29;
    x = (x * 0xBF58476D1CE4E5B9ull) & MASK54;
    x ^= x >> 27;
    x = (x * 0x94D049BB133111EBull) & MASK54;
    x ^= x >> 31;
    return x;
}

int main(int argc, char** argv)
{
    if (argc < 4) { std::fprintf(stderr, "usage: %s out.txt n_samples hashes_per_sample [seed]\n", argv[0]); return 1; }
    const long n = std::atol(argv[2]), per = std::atol(argv[3]);
    const uint64_t seed = argc > 4 ? std::strtoull(argv[4], nullptr, 10) : 0;
    FILE* f = std::fopen(argv[1], "wb");
    if (!f) return 1;
    const long chunk = 64;
    std::vector<std::string> lines(chunk);
    for (long s0 = 0; s0 < n; s0 += chunk) {
        const long s1 = s0 + chunk < n ? s0 + chunk : n;
#pragma omp parallel for schedule(dynamic, 1)
        for (long s = s0; s < s1; ++s) {
            std::string& ln = lines[s - s0];
            ln.clear();
            ln.reserve(static_cast<size_t>(per) * 18 + 32);
            ln += "sample_" + std::to_string(s) + ":";
            char buf[24];
            for (long k = 0; k < per; ++k) {
                uint64_t v = synth(static_cast<uint64_t>(s), static_cast<uint64_t>(k), seed);
                int i = 24;
                do { buf[--i] = static_cast<char>('0' + v % 10); v /= 10; } while (v);
                ln += ' ';
                ln.append(buf + i, static_cast<size_t>(24 - i));
            }
            ln += '\n';
        }
        for (long s = s0; s < s1; ++s) std::fwrite(lines[s - s0].data(), 1, lines[s - s0].size(), f);
    }
    std::fclose(f);
    return 0;
}
