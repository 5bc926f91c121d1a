#include "fast_deflate.h"

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>


namespace sens_loc::io {
namespace {

constexpr std::size_t kBlock   = std::size_t(1) << 18;  // input bytes per deflate block
constexpr int         kLitLen  = 286;
constexpr int         kMaxBits = 15;

// RFC 1951 3.2.5: length 3..258 -> (symbol 257.., extra bits, extra value)
struct len_code {
    std::uint16_t sym;
    std::uint8_t  ebits;
    std::uint8_t  evalue;
};
struct len_table {
    len_code t[259];
    len_table() {
        static const int base[29]  = {3,  4,  5,  6,  7,  8,  9,  10, 11,  13,  15,  17,  19,  23, 27,
                                      31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
        static const int extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2,
                                      2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
        for (int len = 3; len <= 258; ++len) {
            int c = 28;
            while (base[c] > len)
                --c;
            if (len == 258)
                c = 28;
            t[len] = {std::uint16_t(257 + c), std::uint8_t(extra[c]), std::uint8_t(len - base[c])};
        }
    }
};
const len_table kLen;

// 64-bit accumulator, flushed without a branch (the eight bytes at p are always
// writable: rle_deflate_bound leaves slack): after flush() fewer than 8 bits stay
// pending, so up to 56 bits can be added before the next flush.
struct bit_writer {
    unsigned char* p;
    std::uint64_t  acc = 0;
    int            n   = 0;
    void add(std::uint64_t v, int bits) {
        acc |= v << n;
        n += bits;
    }
    void flush() {
        std::memcpy(p, &acc, 8);  // little endian host (x86-64)
        p += n >> 3;
        acc >>= (n & ~7);
        n &= 7;
    }
    void put(std::uint32_t v, int bits) {
        add(v, bits);
        flush();
    }
    void align_byte() {
        flush();
        if (n > 0) {
            *p++ = (unsigned char)acc;
            acc = 0, n = 0;
        }
    }
};

std::uint32_t reverse_bits(std::uint32_t v, int n) {
    std::uint32_t r = 0;
    for (int i = 0; i < n; ++i)
        r |= ((v >> i) & 1u) << (n - 1 - i);
    return r;
}

// Length-limited Huffman code lengths for freq[0..n): Moffat-Katajainen in-place
// minimum-redundancy lengths, then the usual Kraft repair for lengths > kMaxBits.
void code_lengths(const std::uint32_t* freq, int n, std::uint8_t* len) {
    struct sf {
        std::uint32_t f;
        std::uint16_t s;
    };
    sf  used[kLitLen];
    int m = 0;
    for (int i = 0; i < n; ++i) {
        len[i] = 0;
        if (freq[i] != 0)
            used[m++] = {freq[i], std::uint16_t(i)};
    }
    if (m == 1) {  // a complete code needs two symbols: add a dummy
        const int other = used[0].s == 0 ? 1 : 0;
        len[used[0].s] = 1, len[other] = 1;
        return;
    }
    std::sort(used, used + m, [](const sf& a, const sf& b) { return a.f < b.f || (a.f == b.f && a.s < b.s); });
    std::uint32_t A[kLitLen] = {0};
    for (int i = 0; i < m; ++i)
        A[i] = used[i].f;
    {  // Moffat & Katajainen, "In-place calculation of minimum-redundancy codes"
        A[0] += A[1];
        int root = 0, leaf = 2;
        for (int next = 1; next < m - 1; ++next) {
            if (leaf >= m || A[root] < A[leaf]) {
                A[next]   = A[root];
                A[root++] = std::uint32_t(next);
            } else
                A[next] = A[leaf++];
            if (leaf >= m || (root < next && A[root] < A[leaf])) {
                A[next] += A[root];
                A[root++] = std::uint32_t(next);
            } else
                A[next] += A[leaf++];
        }
        A[m - 2] = 0;
        for (int next = m - 3; next >= 0; --next)
            A[next] = A[A[next]] + 1;
        int avbl = 1, usedn = 0, dpth = 0;
        root      = m - 2;
        int next  = m - 1;
        while (avbl > 0) {
            while (root >= 0 && int(A[root]) == dpth) {
                ++usedn;
                --root;
            }
            while (avbl > usedn) {
                A[next--] = std::uint32_t(dpth);
                --avbl;
            }
            avbl  = 2 * usedn;
            ++dpth;
            usedn = 0;
        }
    }
    // A[i] = length of the i-th least frequent symbol (non-increasing in i)
    int count[64] = {0};
    for (int i = 0; i < m; ++i)
        count[std::min<std::uint32_t>(A[i], 63)]++;
    for (int l = kMaxBits + 1; l < 64; ++l) {
        count[kMaxBits] += count[l];
        count[l] = 0;
    }
    std::uint32_t total = 0;
    for (int l = kMaxBits; l >= 1; --l)
        total += std::uint32_t(count[l]) << (kMaxBits - l);
    while (total != (1u << kMaxBits)) {
        count[kMaxBits]--;
        for (int l = kMaxBits - 1; l >= 1; --l)
            if (count[l] != 0) {
                count[l]--;
                count[l + 1] += 2;
                break;
            }
        total--;
    }
    // hand the lengths out again: longest codes to the least frequent symbols
    int i = 0;
    for (int l = kMaxBits; l >= 1; --l)
        for (int c = 0; c < count[l]; ++c)
            len[used[i++].s] = std::uint8_t(l);
}

// canonical codes (RFC 1951 3.2.2), stored bit-reversed for the LSB-first stream
void assign_codes(const std::uint8_t* len, int n, std::uint16_t* code) {
    int bl_count[kMaxBits + 1] = {0};
    for (int i = 0; i < n; ++i)
        bl_count[len[i]]++;
    bl_count[0] = 0;
    std::uint32_t next[kMaxBits + 2];
    std::uint32_t c = 0;
    for (int b = 1; b <= kMaxBits; ++b) {
        c       = (c + std::uint32_t(bl_count[b - 1])) << 1;
        next[b] = c;
    }
    for (int i = 0; i < n; ++i)
        code[i] = len[i] != 0 ? std::uint16_t(reverse_bits(next[len[i]]++, len[i])) : 0;
}

void stored_blocks(bit_writer& bw, const unsigned char* in, std::size_t n, bool final_block) {
    std::size_t off = 0;
    do {
        const std::size_t k    = std::min<std::size_t>(n - off, 65535);
        const bool        last = final_block && off + k == n;
        bw.put(last ? 1u : 0u, 1);
        bw.put(0u, 2);
        bw.align_byte();
        const std::uint16_t len = std::uint16_t(k), nlen = std::uint16_t(~len);
        bw.p[0] = (unsigned char)len, bw.p[1] = (unsigned char)(len >> 8);
        bw.p[2] = (unsigned char)nlen, bw.p[3] = (unsigned char)(nlen >> 8);
        std::memcpy(bw.p + 4, in + off, k);
        bw.p += 4 + k;
        off += k;
    } while (off < n);
}

}  // namespace

std::size_t rle_deflate_bound(std::size_t n) {
    return n + (n / 65535 + 2) * 5 + (n / kBlock + 2) * 8 + 64 + 16;
}

namespace {

struct run {
    std::uint32_t start;  // offset in the block
    std::uint32_t len;    // 3..258, every byte equal to the byte before `start`
};

inline std::uint64_t load64(const unsigned char* p) {
    std::uint64_t v;
    std::memcpy(&v, p, 8);
    return v;
}

// Finds the distance-1 matches of d[0..blk). `has_prev`: d[-1] is readable (a
// byte of the previous block: the deflate window spans blocks).
void find_runs(const unsigned char* d, std::size_t blk, bool has_prev, std::vector<run>& runs) {
    runs.clear();
    std::size_t i = has_prev ? 0 : 1;
    while (i < blk) {
        if (i + 9 <= blk) {
            // byte j of e is zero iff d[i-1+j] == d[i+j]; a match can start at i+j
            // iff bytes j, j+1, j+2 are zero
            const std::uint64_t e = load64(d + i - 1) ^ load64(d + i);
            const std::uint64_t k = 0x7f7f7f7f7f7f7f7fULL;
            const std::uint64_t z = ~(((e & k) + k) | e | k);  // 0x80 in every zero byte
            const std::uint64_t t = z & (z >> 8) & (z >> 16);
            if (t == 0) {
                i += 6;
                continue;
            }
            i += std::size_t(__builtin_ctzll(t)) >> 3;
        } else if (!(i + 2 < blk && d[i] == d[i - 1] && d[i + 1] == d[i - 1] && d[i + 2] == d[i - 1])) {
            ++i;
            continue;
        }
        const unsigned char prev = d[i - 1];
        const std::size_t   cap  = std::min<std::size_t>(258, blk - i);
        std::size_t         len  = 3;
        while (len < cap && d[i + len] == prev)
            ++len;
        runs.push_back({std::uint32_t(i), std::uint32_t(len)});
        i += len;
    }
}

void histogram(const unsigned char* p, std::size_t n, std::uint32_t (&h)[4][256]) {
    std::size_t i = 0;
    for (; i + 4 <= n; i += 4) {
        h[0][p[i]]++;
        h[1][p[i + 1]]++;
        h[2][p[i + 2]]++;
        h[3][p[i + 3]]++;
    }
    for (; i < n; ++i)
        h[0][p[i]]++;
}

}  // namespace

std::size_t rle_deflate(const unsigned char* in, std::size_t n, unsigned char* out) {
    bit_writer bw{out + 2};
    out[0] = 0x78, out[1] = 0x01;  // zlib header: deflate, 32 KB window, fastest
    std::vector<run> runs;
    std::size_t      pos = 0;
    do {
        const std::size_t    blk   = std::min(kBlock, n - pos);
        const bool           final = pos + blk == n;
        const unsigned char* d     = in + pos;
        find_runs(d, blk, pos > 0, runs);
        // symbol frequencies: literals between the runs, one length symbol per run
        std::uint32_t h4[4][256];
        std::memset(h4, 0, sizeof(h4));
        std::uint32_t freq[kLitLen] = {0};
        std::size_t   extra_bits = 0, at = 0;
        for (const run& r : runs) {
            histogram(d + at, r.start - at, h4);
            const len_code& lc = kLen.t[r.len];
            freq[lc.sym]++;
            extra_bits += lc.ebits;
            at = std::size_t(r.start) + r.len;
        }
        histogram(d + at, blk - at, h4);
        for (int b = 0; b < 256; ++b)
            freq[b] = h4[0][b] + h4[1][b] + h4[2][b] + h4[3][b];
        freq[256] = 1;  // end of block
        std::uint8_t  len[kLitLen];
        std::uint16_t code[kLitLen];
        code_lengths(freq, kLitLen, len);
        std::size_t bits = 3 + 5 + 5 + 4 + 19 * 3 + (kLitLen + 1) * 4 + extra_bits + runs.size();
        for (int s = 0; s < kLitLen; ++s)
            bits += std::size_t(freq[s]) * len[s];
        if ((bits + 7) / 8 >= blk + 5 * (blk / 65535 + 1)) {
            stored_blocks(bw, d, blk, final);
        } else {
            assign_codes(len, kLitLen, code);
            bw.put(final ? 1u : 0u, 1);
            bw.put(2u, 2);             // dynamic Huffman
            bw.put(kLitLen - 257, 5);  // HLIT
            bw.put(0u, 5);             // HDIST: one distance code
            bw.put(19 - 4, 4);         // HCLEN: all 19 code length codes
            // code length alphabet: symbols 0..15 with 4 bits each (complete), 16-18 unused
            static const int order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            for (int k = 0; k < 19; ++k)
                bw.put(order[k] < 16 ? 4u : 0u, 3);
            for (int s = 0; s < kLitLen; ++s)
                bw.put(reverse_bits(len[s], 4), 4);
            bw.put(reverse_bits(1, 4), 4);  // distance code 0: length 1
            // one table entry per literal: code | length << 16
            std::uint32_t lit[256];
            for (int b = 0; b < 256; ++b)
                lit[b] = std::uint32_t(code[b]) | (std::uint32_t(len[b]) << 16);
            auto literals = [&](const unsigned char* p, std::size_t m) {
                std::size_t i = 0;
                for (; i + 3 <= m; i += 3) {  // three codes (<= 45 bits) per flush
                    const std::uint32_t e0 = lit[p[i]], e1 = lit[p[i + 1]], e2 = lit[p[i + 2]];
                    bw.add(e0 & 0xffffu, int(e0 >> 16));
                    bw.add(e1 & 0xffffu, int(e1 >> 16));
                    bw.add(e2 & 0xffffu, int(e2 >> 16));
                    bw.flush();
                }
                for (; i < m; ++i) {
                    const std::uint32_t e = lit[p[i]];
                    bw.put(e & 0xffffu, int(e >> 16));
                }
            };
            at = 0;
            for (const run& r : runs) {
                literals(d + at, r.start - at);
                const len_code& lc = kLen.t[r.len];
                // length code, its extra bits, then the 1-bit distance code '0'
                bw.put(code[lc.sym] | (std::uint32_t(lc.evalue) << len[lc.sym]), len[lc.sym] + lc.ebits + 1);
                at = std::size_t(r.start) + r.len;
            }
            literals(d + at, blk - at);
            bw.put(code[256], len[256]);
        }
        pos += blk;
    } while (pos < n);
    bw.align_byte();
    const std::uint32_t a = adler32_fast(1, in, n);
    bw.p[0] = (unsigned char)(a >> 24), bw.p[1] = (unsigned char)(a >> 16);
    bw.p[2] = (unsigned char)(a >> 8), bw.p[3] = (unsigned char)a;
    return std::size_t(bw.p + 4 - out);
}

}  // namespace sens_loc::io
