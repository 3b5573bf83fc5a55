// device_units.cpp — TEST harness (CPU only): the pure-arithmetic device functions of the CUDA path compiled for the
// HOST and checked against naive restatements / the oracle by tests/test_device_units.py:
//   * text -> masks (frame.cu: eq7, gather2, masks_word, chunk_masks, chunk_newlines, count_nl, before)
//   * 32-mer codes from bit-planes (eref_internal.cuh: window32, window_bits, stream_code, valid_windows)
//   * which table part a code falls into, without forming it (scan.cu: part_mask)
//   * the associative FASTA framing summary (eref_internal.cuh: TileSum / TileOp)
//   * the cut points of a sharded table build (api.cu: part_first_slice, part_weight; host arithmetic)
// The function bodies are NOT copied here: eref_internal.cuh is included as it is, and the test cuts the two source
// regions out of frame.cu / scan.cu / api.cu into tests/_build/device_snippets.inc at build time, so what runs is the product's
// own source text with the handful of CUDA integer intrinsics below supplied for the host.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

using std::max;
using std::min;
static inline uint32_t __brev(uint32_t x)
{
    uint32_t r = 0;
    for (int i = 0; i < 32; i++) r |= ((x >> i) & 1u) << (31 - i);
    return r;
}
static inline int __popc(uint32_t x) { return __builtin_popcount(x); }
static inline int __clz(uint32_t x) { return x ? __builtin_clz(x) : 32; }
// upper 32 bits of (hi:lo) << (s & 31)
static inline uint32_t __funnelshift_l(uint32_t lo, uint32_t hi, uint32_t s)
{
    return (uint32_t)(((((uint64_t)hi << 32) | lo) << (s & 31)) >> 32);
}
// lower 32 bits of (hi:lo) >> min(s, 32)
static inline uint32_t __funnelshift_rc(uint32_t lo, uint32_t hi, uint32_t s)
{
    const uint64_t v = ((uint64_t)hi << 32) | lo;
    return (uint32_t)(s >= 32 ? v >> 32 : v >> s);
}

#include "../palace_b200/csrc/eref_internal.cuh"
#include "_build/device_snippets.inc"

namespace {

struct Rng {                       // xorshift64*: the same stream on every platform
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 1) {}
    uint64_t next() { s ^= s >> 12; s ^= s << 25; s ^= s >> 27; return s * 0x2545F4914F6CDD1Dull; }
    uint32_t u32() { return (uint32_t)(next() >> 32); }
};

// ---- naive restatements ----
bool naive_ok(unsigned c) { return c == 'A' || c == 'C' || c == 'G' || c == 'T' || c == 'a' || c == 'c' || c == 'g' || c == 't'; }
int naive_code(unsigned c)         // A=00 C=01 G=10 T=11 (eref_internal.cuh: Planes)
{
    switch (c & 0xDFu) { case 'A': return 0; case 'C': return 1; case 'G': return 2; default: return 3; }
}

int check_masks(uint64_t seed, int rounds)
{
    Rng rng(seed);
    long bad = 0;
    uint8_t text[32];
    auto one = [&] {
        uint32_t w[8];
        memcpy(w, text, 32);                                   // little-endian words, as the LDG.128 pair delivers them
        const Masks m = chunk_masks(w);
        uint32_t nl = 0, ok = 0, hi = 0, lo = 0, cnt = 0;
        for (int k = 0; k < 32; k++) {
            const uint32_t bit = 0x80000000u >> k;
            if (text[k] == '\n') { nl |= bit; cnt++; }
            if (naive_ok(text[k])) { ok |= bit; if (naive_code(text[k]) & 2) hi |= bit; if (naive_code(text[k]) & 1) lo |= bit; }
        }
        bad += m.nl != nl || m.ok != ok || m.hi != hi || m.lo != lo || chunk_newlines(w) != nl || count_nl(w) != cnt;
    };
    // every byte value at every position, among random neighbours of every kind
    for (int r = 0; r < rounds; r++)
        for (int k = 0; k < 32; k++)
            for (int b = 0; b < 256; b++) {
                for (int j = 0; j < 32; j++) {
                    const uint32_t x = rng.u32();
                    text[j] = (x & 3) == 0 ? (uint8_t)"ACGTacgt\nN>E"[(x >> 8) % 12] : (uint8_t)(x >> 16);
                }
                text[k] = (uint8_t)b;
                one();
            }
    for (int b = 0; b < 256; b++) { memset(text, b, 32); one(); }
    for (int k = 0; k <= 40; k++) bad += before(k) != (k >= 32 ? 0xFFFFFFFFu : k == 0 ? 0u : 0xFFFFFFFFu << (32 - k));
    printf("masks mismatches %ld\n", bad);
    return bad != 0;
}

CoderMasks masks_of(const int16_t coder[96])                   // the loop of eref_create (api.cu)
{
    CoderMasks m{};
    for (int z = 0; z < 32; z++)
        for (int i = 0; i < 3; i++) {
            if (coder[z * 3 + i] == 0) m.sel0[i] |= 1u << (31 - z);
            if (coder[z * 3 + i] == 2) m.sel2[i] |= 1u << (31 - z);
        }
    return m;
}

void planes_of(const std::string& seq, uint32_t hi[2], uint32_t lo[2], uint32_t ok[2])
{
    hi[0] = hi[1] = lo[0] = lo[1] = ok[0] = ok[1] = 0;
    for (size_t p = 0; p < 64 && p < seq.size(); p++) {
        const unsigned c = (unsigned char)seq[p];
        if (!naive_ok(c)) continue;
        const uint32_t bit = 0x80000000u >> (p & 31);
        ok[p >> 5] |= bit;
        if (naive_code(c) & 2) hi[p >> 5] |= bit;
        if (naive_code(c) & 1) lo[p >> 5] |= bit;
    }
}

bool read_coder(int16_t coder[96])
{
    for (int j = 0; j < 96; j++) { int v; if (!(std::cin >> v)) return false; coder[j] = (int16_t)v; }
    std::string rest;
    std::getline(std::cin, rest);
    return true;
}

// stdin: 96 coder entries, then 64-character sequences; per sequence and window start s = 0..31:
// "<valid> <code0> <code1> <code2>" (window_bits + stream_code on the planes; valid from valid_windows)
int print_codes()
{
    int16_t coder[96];
    if (!read_coder(coder)) return 2;
    const CoderMasks m = masks_of(coder);
    std::string seq;
    while (std::getline(std::cin, seq)) {
        uint32_t hi[2], lo[2], ok[2];
        planes_of(seq, hi, lo, ok);
        const uint32_t vm = valid_windows(ok[0], ok[1]);
        for (int s = 0; s < 32; s++) {
            const WindowBits b = window_bits(window32(hi[0], hi[1], s), window32(lo[0], lo[1], s));
            printf("%u %u %u %u\n", (vm >> (31 - s)) & 1u, stream_code(m, b, 0), stream_code(m, b, 1), stream_code(m, b, 2));
        }
    }
    return 0;
}

template <int LOG2P>
long parts_bad(const CoderMasks& m, const uint32_t hi[2], const uint32_t lo[2])
{
    long bad = 0;
    for (int i = 0; i < 3; i++) {
        uint32_t seen = 0;
        for (uint32_t part = 0; part < (1u << LOG2P); part++) {
            const uint32_t pm = part_mask<LOG2P>(m, i, hi[0], hi[1], lo[0], lo[1], part);
            uint32_t want = 0;
            for (int s = 0; s < 32; s++) {
                const WindowBits b = window_bits(window32(hi[0], hi[1], s), window32(lo[0], lo[1], s));
                if (stream_code(m, b, i) >> (32 - LOG2P) == part) want |= 0x80000000u >> s;
            }
            bad += pm != want;
            bad += (seen & pm) != 0;                            // the parts partition the 32 starts
            seen |= pm;
        }
        bad += seen != 0xFFFFFFFFu;
    }
    return bad;
}

int check_parts(uint64_t seed, int rounds)
{
    Rng rng(seed);
    long bad = 0;
    for (int r = 0; r < rounds; r++) {
        int16_t coder[96];
        static const int16_t orderings[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 2, 0}, {1, 0, 2}, {2, 0, 1}, {2, 1, 0}};
        for (int z = 0; z < 32; z++) { const int o = rng.u32() % 6; for (int i = 0; i < 3; i++) coder[z * 3 + i] = orderings[o][i]; }
        const CoderMasks m = masks_of(coder);
        for (int t = 0; t < 64; t++) {
            uint32_t hi[2] = {rng.u32(), rng.u32()}, lo[2] = {rng.u32(), rng.u32()};
            if (t % 8 == 0) { hi[0] = hi[1] = t & 8 ? 0xFFFFFFFFu : 0u; }          // homopolymer-like planes
            if (t % 16 == 1) { hi[1] = __brev(~hi[0]); lo[1] = __brev(~lo[0]); }   // a reverse-complement palindrome (fwd == rc)
            bad += parts_bad<1>(m, hi, lo) + parts_bad<2>(m, hi, lo) + parts_bad<3>(m, hi, lo) + parts_bad<4>(m, hi, lo) +
                   parts_bad<5>(m, hi, lo);
        }
    }
    printf("parts mismatches %ld\n", bad);
    return bad != 0;
}

// ---- TileSum / TileOp against the sequential framing rule (:686-756): a line is a header iff its first byte is '>',
// every other non-newline byte counts as sequence.  summary of text[lo, hi): a line starts at p iff p == 0 or
// text[p-1] == '\n'.
TileSum summarize(const std::string& t, size_t lo, size_t hi)
{
    TileSum s{0, 0, 0, 0};
    bool header = false;
    for (size_t p = lo; p < hi; p++) {
        const bool start = p == 0 || t[p - 1] == '\n';
        if (start) {
            s.flags |= 1u;
            header = t[p] == '>';
            s.flags = header ? (s.flags | 2u) : (s.flags & ~2u);
            if (header) s.hdrs++;
        }
        if (t[p] == '\n') continue;
        if (!(s.flags & 1u)) s.seqA++;                          // belongs to a line that started in an earlier tile
        else if (!header) s.seqB++;
    }
    return s;
}

bool same(const TileSum& a, const TileSum& b) { return a.flags == b.flags && a.hdrs == b.hdrs && a.seqA == b.seqA && a.seqB == b.seqB; }

int check_tileop(uint64_t seed, int rounds)
{
    Rng rng(seed);
    long bad = 0;
    for (int r = 0; r < rounds; r++) {
        const size_t n = 1 + rng.u32() % 300;
        std::string t(n, 'A');
        for (size_t p = 0; p < n; p++) { const uint32_t x = rng.u32() % 16; t[p] = x < 3 ? '\n' : x < 5 ? '>' : "ACGTN"[x % 5]; }
        // cut into random tiles, fold left to right, and in a different association
        std::vector<size_t> cut(1, 0);
        while (cut.back() < n) cut.push_back(std::min(n, cut.back() + 1 + rng.u32() % 40));
        std::vector<TileSum> tiles;
        for (size_t k = 0; k + 1 < cut.size(); k++) tiles.push_back(summarize(t, cut[k], cut[k + 1]));
        TileOp op;
        TileSum left = tiles[0];
        for (size_t k = 1; k < tiles.size(); k++) left = op(left, tiles[k]);
        TileSum right = tiles.back();
        for (size_t k = tiles.size() - 1; k-- > 0;) right = op(tiles[k], right);
        const TileSum whole = summarize(t, 0, n);
        bad += !same(left, whole) || !same(right, whole);
        // prefixes: fold of the first k tiles = summary of text[0, cut[k])
        TileSum pre = tiles[0];
        for (size_t k = 1; k < tiles.size(); k++) { bad += !same(pre, summarize(t, 0, cut[k])); pre = op(pre, tiles[k]); }
    }
    printf("tileop mismatches %ld\n", bad);
    return bad != 0;
}

// ---- table parts of a sharded build (api.cu: part_first_slice, part_weight): the cut points of `parts` GPUs at weight
// w partition the NSLICES slices, every part gets at least one, and they equalise w * code mass + (1 - w) * width, the
// mass of canonical codes min(fwd, rc) falling as 2(1 - x) over the code space
int check_cuts()
{
    long bad = 0;
    for (int parts = 1; parts <= 16; parts++)
        for (int wi = 0; wi <= 19; wi++) {
            const double w = wi * 0.05;
            bad += part_first_slice(0, parts, w) != 0 || part_first_slice(parts, parts, w) != (uint32_t)NSLICES;
            double cost_lo = 1e9, cost_hi = 0;
            for (int p = 0; p < parts; p++) {
                const uint32_t a = part_first_slice(p, parts, w), b = part_first_slice(p + 1, parts, w);
                bad += b <= a;
                const double x0 = (double)a / NSLICES, x1 = (double)b / NSLICES;
                const double mass = (1 - x0) * (1 - x0) - (1 - x1) * (1 - x1), cost = w * mass + (1 - w) * (x1 - x0);
                cost_lo = std::min(cost_lo, cost); cost_hi = std::max(cost_hi, cost);
            }
            bad += cost_hi - cost_lo > 4.0 / NSLICES;          // equal up to the rounding to whole slices
        }
    printf("cuts mismatches %ld\n", bad);
    for (int parts : {2, 8}) {                                 // the numbers DESIGN.md section 6 quotes (w = 0.7)
        for (int p = 0; p < parts; p++) printf("%u ", part_first_slice(p, parts, 0.7));
        printf("\n");
    }
    // the weight rises with the query and falls with the number of GPUs, within [0, 0.95]
    long wbad = 0;
    for (int parts = 2; parts <= 16; parts++) {
        double last = -1;
        for (uint64_t q = 1u << 20; q < (1ull << 36); q <<= 1) {
            const double w = part_weight(parts, q);
            wbad += w < last || w < 0 || w > 0.95 || (parts > 2 && w > part_weight(parts - 1, q));
            last = w;
        }
    }
    printf("weight mismatches %ld\n", wbad);
    return bad != 0 || wbad != 0;
}

} // namespace

int main(int argc, char** argv)
{
    const std::string cmd = argc > 1 ? argv[1] : "";
    const uint64_t seed = argc > 2 ? strtoull(argv[2], nullptr, 10) : 1;
    const int rounds = argc > 3 ? atoi(argv[3]) : 1;
    if (cmd == "masks") return check_masks(seed, rounds);
    if (cmd == "codes") return print_codes();
    if (cmd == "parts") return check_parts(seed, rounds);
    if (cmd == "tileop") return check_tileop(seed, rounds);
    if (cmd == "cuts") return check_cuts();
    fprintf(stderr, "usage: device_units masks|parts|tileop <seed> <rounds>  |  codes < coder+sequences\n");
    return 2;
}
