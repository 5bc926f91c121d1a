// A whole-buffer inflate (RFC 1950/1951) for the PNG reader: 64-bit bit buffer
// refilled a word at a time, 11-bit (literal/length) and 8-bit (distance) primary
// decode tables with sub-tables for longer codes, word-wise match copies. It
// verifies the Adler-32 trailer itself; the PNG reader falls back to zlib on ANY
// failure reported here, so this decoder only has to be fast and never touch
// memory out of bounds; it does not have to diagnose corrupt streams precisely.
#include "fast_deflate.h"

#include <cstdint>
#include <cstdlib>
#include <cstring>
#if defined(__x86_64__)
#include <immintrin.h>
#define DCB200_INFLATE_X86 1
#endif

namespace sens_loc::io {
namespace {

constexpr int kLitBits = 11, kDistBits = 8;
constexpr int kLitSize = 1 << kLitBits, kDistSize = 1 << kDistBits;
constexpr int kLitCap = 4096, kDistCap = 1024;  // primary + sub-tables

// entry: bits 0-7 code bits consumed at this level, bits 8-11 extra bit count
// (or sub-table index bits), bits 12-15 kind, bits 16-31 value / base / sub-table start
// kinds: bit 2 = one or two literals (value = first | second << 8), bit 3 = two
enum : std::uint32_t { K_BAD = 0, K_BASE = 1, K_EOB = 2, K_SUB = 3, K_LIT = 4, K_LIT2 = 12 };
constexpr std::uint32_t kLitFlag = 4u << 12, kLit2Flag = 8u << 12;
inline std::uint32_t entry(std::uint32_t kind, std::uint32_t value, std::uint32_t extra, std::uint32_t bits) {
    return (value << 16) | (kind << 12) | (extra << 8) | bits;
}

const std::uint16_t kLenBase[29]  = {3,  4,  5,  6,  7,  8,  9,  10, 11,  13,  15,  17,  19,  23, 27,
                                     31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const std::uint8_t  kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2,
                                     2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const std::uint16_t kDistBase[30] = {1,   2,   3,   4,   5,   7,    9,    13,   17,   25,
                                     33,  49,  65,  97,  129, 193,  257,  385,  513,  769,
                                     1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const std::uint8_t  kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8,
                                      9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

// what a symbol decodes to, without the code length (bits 0-7)
struct symbol_table {
    std::uint32_t litlen[320], dist[32];
    symbol_table() {
        for (int s = 0; s < 320; ++s)
            litlen[s] = s < 256    ? entry(K_LIT, std::uint32_t(s), 0, 0)
                        : s == 256 ? entry(K_EOB, 0, 0, 0)
                        : s <= 285 ? entry(K_BASE, kLenBase[s - 257], kLenExtra[s - 257], 0)
                                   : entry(K_BAD, 0, 0, 0);
        for (int s = 0; s < 32; ++s)
            dist[s] = s < 30 ? entry(K_BASE, kDistBase[s], kDistExtra[s], 0) : entry(K_BAD, 0, 0, 0);
    }
};
const symbol_table kSymbols;
inline std::uint32_t symbol_entry(bool litlen, int sym, std::uint32_t bits) {
    return (litlen ? kSymbols.litlen[sym] : kSymbols.dist[sym]) | bits;
}

struct rev8_table {
    std::uint8_t v[256];
    rev8_table() {
        for (int i = 0; i < 256; ++i) {
            int r = 0;
            for (int k = 0; k < 8; ++k)
                r |= ((i >> k) & 1) << (7 - k);
            v[i] = std::uint8_t(r);
        }
    }
};
const rev8_table kRev8;
inline std::uint32_t reverse(std::uint32_t v, int n) {  // n <= 15
    return ((std::uint32_t(kRev8.v[v & 0xff]) << 8) | kRev8.v[(v >> 8) & 0xff]) >> (16 - n);
}

// Canonical Huffman decode table from code lengths. Returns false for an
// over-subscribed code or when the sub-tables do not fit (caller falls back).
bool build(const std::uint8_t* len, int n, bool litlen, int tbits, std::uint32_t* table, int cap,
           bool code_lengths = false, bool fixed = false) {
    int count[16] = {0};
    for (int i = 0; i < n; ++i)
        count[len[i]]++;
    count[0] = 0;
    std::uint32_t next[16], code = 0;
    std::int64_t  left = 1;
    for (int b = 1; b <= 15; ++b) {
        left = (left << 1) - count[b];
        if (left < 0)
            return false;
        code    = (code + std::uint32_t(count[b - 1])) << 1;
        next[b] = code;
    }
    // like zlib: an incomplete code is only legal as a single code of one bit (or, for
    // distances, no code at all); never for the code-length code. (The fixed distance
    // code has 30 of 32 codes.)
    int max_len = 15;
    while (max_len > 0 && count[max_len] == 0)
        --max_len;
    if (left != 0 && !fixed && (code_lengths || max_len > 1))
        return false;
    const int           tsize = 1 << tbits;
    const std::uint32_t bad   = entry(K_BAD, 0, 0, 1);
    if (left != 0)  // some entries stay unassigned
        for (int i = 0; i < tsize; ++i)
            table[i] = bad;
    // symbols in canonical order (by length, then value): the fill loops below then run
    // the same number of times for a whole length class, which the branch predictor likes
    std::uint16_t sorted[320];
    {
        int offs[16], at = 0;
        for (int b = 1; b <= 15; ++b) {
            offs[b] = at;
            at += count[b];
        }
        for (int s = 0; s < n; ++s)
            if (len[s] != 0)
                sorted[offs[len[s]]++] = std::uint16_t(s);
    }
    // short codes fill the primary table; long ones are collected for the sub-tables
    std::uint16_t long_sym[320];
    std::uint32_t long_code[320];
    int           n_long = 0, k = 0;
    for (int l = 1; l <= max_len && l <= tbits; ++l) {
        const std::uint32_t step = 1u << l;
        for (int q = 0; q < count[l]; ++q, ++k) {
            const int           sym = sorted[k];
            const std::uint32_t c   = reverse(next[l]++, l);
            const std::uint32_t e   = symbol_entry(litlen, sym, std::uint32_t(l));
            for (std::uint32_t i = c; i < std::uint32_t(tsize); i += step)
                table[i] = e;
        }
    }
    for (int l = tbits + 1; l <= max_len; ++l)
        for (int q = 0; q < count[l]; ++q, ++k) {
            const std::uint32_t c = reverse(next[l]++, l);
            long_sym[n_long]      = sorted[k];
            long_code[n_long++]   = c;
            table[c & std::uint32_t(tsize - 1)] = 0;  // sub-table bits so far
        }
    if (n_long == 0)
        return true;
    for (int k = 0; k < n_long; ++k) {  // longest code behind each prefix
        std::uint32_t&      t    = table[long_code[k] & std::uint32_t(tsize - 1)];
        const std::uint32_t need = std::uint32_t(len[long_sym[k]] - tbits);
        if (need > t)
            t = need;
    }
    int used = tsize;
    for (int k = 0; k < n_long; ++k) {
        std::uint32_t& t = table[long_code[k] & std::uint32_t(tsize - 1)];
        if (t >= 16)  // already a pointer (sub-table bits are 1..7)
            continue;
        const int sb = int(t), size = 1 << sb;
        if (used + size > cap)
            return false;
        t = entry(K_SUB, std::uint32_t(used), std::uint32_t(sb), std::uint32_t(tbits));
        for (int i = 0; i < size; ++i)
            table[used + i] = bad;
        used += size;
    }
    for (int k = 0; k < n_long; ++k) {
        const std::uint32_t t  = table[long_code[k] & std::uint32_t(tsize - 1)];
        const std::uint32_t st = t >> 16, sb = (t >> 8) & 0xf;
        const int           l  = len[long_sym[k]] - tbits;
        const std::uint32_t e  = symbol_entry(litlen, long_sym[k], std::uint32_t(l));
        for (std::uint32_t i = long_code[k] >> tbits; i < (1u << sb); i += 1u << l)
            table[st + i] = e;
    }
    return true;
}

struct bits {
    const unsigned char *in, *end;  // `end` = last byte + 1; 8 readable bytes follow it
    std::uint64_t        buf  = 0;
    int                  left = 0;
    void refill() {
        std::uint64_t w;
        std::memcpy(&w, in, 8);
        buf |= w << left;
        in += (63 - left) >> 3;
        left |= 56;
    }
    std::uint32_t peek(int n) const { return std::uint32_t(buf & ((std::uint64_t(1) << n) - 1)); }
    void          drop(int n) {
        buf >>= n;
        left -= n;
    }
    std::uint32_t take(int n) {
        const std::uint32_t v = peek(n);
        drop(n);
        return v;
    }
};

// Two literals whose codes fit the primary index together become one entry: depth
// PNGs are mostly literals (sensor noise in the low byte). `prim` is scratch.
void pair_literals_scalar(std::uint32_t* lit, std::uint32_t* prim) {
    std::memcpy(prim, lit, sizeof(std::uint32_t) * kLitSize);
    for (int i = 0; i < kLitSize; ++i) {  // (branch-free)
        const std::uint32_t e1 = prim[i], l1 = e1 & 0xff;
        const std::uint32_t e2 = prim[i >> l1], l12 = l1 + (e2 & 0xff);
        const std::uint32_t ok  = ((e1 & e2) >> 14) & ((l12 - std::uint32_t(kLitBits + 1)) >> 31);
        const std::uint32_t two = ((e2 >> 16) << 24) | (e1 & 0xff0000u) | (K_LIT2 << 12) | l12;
        const std::uint32_t m   = 0u - (ok & 1u);
        lit[i]                  = (two & m) | (e1 & ~m);
    }
}

#ifdef DCB200_INFLATE_X86
__attribute__((target("avx2"))) void pair_literals_avx2(std::uint32_t* lit, std::uint32_t* prim) {
    std::memcpy(prim, lit, sizeof(std::uint32_t) * kLitSize);
    const __m256i byte = _mm256_set1_epi32(0xff), flag = _mm256_set1_epi32(int(kLitFlag));
    const __m256i room = _mm256_set1_epi32(kLitBits + 1), first = _mm256_set1_epi32(0xff0000);
    const __m256i kind = _mm256_set1_epi32(int(K_LIT2 << 12));
    __m256i       idx  = _mm256_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7);
    for (int i = 0; i < kLitSize; i += 8) {
        const __m256i e1  = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(prim + i));
        const __m256i l1  = _mm256_and_si256(e1, byte);
        const __m256i e2  = _mm256_i32gather_epi32(reinterpret_cast<const int*>(prim), _mm256_srlv_epi32(idx, l1), 4);
        const __m256i l12 = _mm256_add_epi32(l1, _mm256_and_si256(e2, byte));
        const __m256i ok  = _mm256_and_si256(
            _mm256_cmpeq_epi32(_mm256_and_si256(_mm256_and_si256(e1, e2), flag), flag), _mm256_cmpgt_epi32(room, l12));
        const __m256i two = _mm256_or_si256(_mm256_or_si256(_mm256_slli_epi32(_mm256_srli_epi32(e2, 16), 24),
                                                            _mm256_and_si256(e1, first)),
                                            _mm256_or_si256(kind, l12));
        _mm256_storeu_si256(reinterpret_cast<__m256i*>(lit + i), _mm256_blendv_epi8(e1, two, ok));
        idx = _mm256_add_epi32(idx, _mm256_set1_epi32(8));
    }
}
#endif

void pair_literals(std::uint32_t* lit, std::uint32_t* prim) {
#ifdef DCB200_INFLATE_X86
    // (DC_INFLATE_SCALAR=1: the portable loop, for the tests)
    static const bool has = __builtin_cpu_supports("avx2") && std::getenv("DC_INFLATE_SCALAR") == nullptr;
    if (has)
        return pair_literals_avx2(lit, prim);
#endif
    pair_literals_scalar(lit, prim);
}

}  // namespace

bool fast_inflate(const unsigned char* in, std::size_t n_in, unsigned char* out, std::size_t n_out) {
    if (n_in < 6)
        return false;
    // zlib header: deflate, window <= 32 KB, no preset dictionary, check bits
    if ((in[0] & 0x0f) != 8 || (in[0] >> 4) > 7 || (in[1] & 0x20) != 0 || ((in[0] << 8) | in[1]) % 31 != 0)
        return false;
    bits b;
    b.in  = in + 2;
    b.end = in + n_in;
    unsigned char *      op = out, *const oend = out + n_out;
    static thread_local std::uint32_t lit[kLitCap], dist[kDistCap], prim[kLitSize];
    bool final_block = false;
    while (!final_block) {
        if (b.in > b.end + 8)
            return false;
        b.refill();
        final_block        = b.take(1) != 0;
        const unsigned typ = b.take(2);
        if (typ == 0) {  // stored
            b.drop(b.left & 7);
            // bytes still in the buffer belong to the stream: step back
            b.in -= b.left >> 3;
            b.buf = 0, b.left = 0;
            if (b.in + 4 > b.end)
                return false;
            const unsigned len = b.in[0] | (b.in[1] << 8), nlen = b.in[2] | (b.in[3] << 8);
            if ((len ^ 0xffffu) != nlen || b.in + 4 + len > b.end || op + len > oend)
                return false;
            std::memcpy(op, b.in + 4, len);
            op += len;
            b.in += 4 + len;
            continue;
        }
        if (typ == 3)
            return false;
        std::uint8_t lens[320];
        int          nlit = 288, ndist = 30;
        if (typ == 1) {  // fixed codes, RFC 1951 3.2.6
            for (int i = 0; i < 144; ++i) lens[i] = 8;
            for (int i = 144; i < 256; ++i) lens[i] = 9;
            for (int i = 256; i < 280; ++i) lens[i] = 7;
            for (int i = 280; i < 288; ++i) lens[i] = 8;
            for (int i = 0; i < 30; ++i) lens[288 + i] = 5;
        } else {
            nlit              = int(b.take(5)) + 257;
            ndist             = int(b.take(5)) + 1;
            const int ncl     = int(b.take(4)) + 4;
            if (nlit > 286 || ndist > 30)
                return false;
            static const int order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            std::uint8_t     cl[19]    = {0};
            b.refill();
            for (int i = 0; i < ncl; ++i) {
                if (b.left < 3)
                    b.refill();
                cl[order[i]] = std::uint8_t(b.take(3));
            }
            std::uint32_t clt[128];
            if (!build(cl, 19, true, 7, clt, 128, true))  // (symbols 0..18 come out as "literals")
                return false;
            int i = 0;
            while (i < nlit + ndist) {
                if (b.in > b.end + 8)
                    return false;
                b.refill();
                const std::uint32_t e = clt[b.peek(7)];
                if (((e >> 12) & 0xf) != K_LIT)
                    return false;
                b.drop(int(e & 0xff));
                const unsigned sym = e >> 16;
                if (sym < 16) {
                    lens[i++] = std::uint8_t(sym);
                    continue;
                }
                unsigned rep, val = 0;
                if (sym == 16) {
                    if (i == 0)
                        return false;
                    val = lens[i - 1];
                    rep = 3 + b.take(2);
                } else if (sym == 17) {
                    rep = 3 + b.take(3);
                } else {
                    rep = 11 + b.take(7);
                }
                if (i + int(rep) > nlit + ndist)
                    return false;
                while (rep-- != 0)
                    lens[i++] = std::uint8_t(val);
            }
            if (lens[256] == 0)
                return false;
            // distance lengths follow the literal/length lengths: move them to a fixed place
            std::uint8_t tmp[30];
            std::memcpy(tmp, lens + nlit, std::size_t(ndist));
            std::memset(lens + nlit, 0, std::size_t(320 - nlit));
            std::memcpy(lens + 288, tmp, std::size_t(ndist));
        }
        if (!build(lens, nlit, true, kLitBits, lit, kLitCap))
            return false;
        // (a single distance code of one bit, or none at all, is legal: the unused
        // entries decode to K_BAD)
        if (!build(lens + 288, ndist, false, kDistBits, dist, kDistCap, false, typ == 1))
            return false;

        pair_literals(lit, prim);

        b.refill();
        std::uint32_t e = lit[b.peek(kLitBits)];
        for (;;) {  // invariant: e is the entry for the next bits and those bits are in the buffer
            if (b.in > b.end + 8)
                return false;
            b.refill();
            if (op + 8 <= oend && (e & kLitFlag) != 0) {
                // up to four entries of one or two literals from one refill (<= 44 of
                // >= 56 bits), without a branch on one-or-two: always store both bytes
                int n = 0;
                do {
                    b.drop(int(e & 0xff));
                    const std::uint16_t two = std::uint16_t(e >> 16);
                    std::memcpy(op, &two, 2);
                    op += 1 + ((e >> 15) & 1);
                    e = lit[b.peek(kLitBits)];
                } while (++n < 4 && (e & kLitFlag) != 0);
                if (n == 4)
                    continue;
                b.refill();
            }
            if (((e >> 12) & 0xf) == K_SUB) {
                b.drop(kLitBits);
                e = lit[(e >> 16) + b.peek(int((e >> 8) & 0xf))];
            }
            b.drop(int(e & 0xff));
            const std::uint32_t kind = (e >> 12) & 0xf;
            if (kind == K_LIT) {
                if (op >= oend)
                    return false;
                *op++ = (unsigned char)(e >> 16);
                e     = lit[b.peek(kLitBits)];  // >= 41 bits left
                continue;
            }
            if (kind == K_LIT2) {  // (single literals and pairs: only within 8 bytes of the end)
                if (op + 2 > oend)
                    return false;
                *op++ = (unsigned char)(e >> 16);
                *op++ = (unsigned char)(e >> 24);
                e     = lit[b.peek(kLitBits)];
                continue;
            }
            if (kind == K_EOB)
                break;
            if (kind != K_BASE)
                return false;
            std::size_t len = (e >> 16) + b.take(int((e >> 8) & 0xf));
            // at most 15 + 5 bits used: >= 36 left, a distance needs <= 15 + 13
            std::uint32_t d = dist[b.peek(kDistBits)];
            if (((d >> 12) & 0xf) == K_SUB) {
                b.drop(kDistBits);
                d = dist[(d >> 16) + b.peek(int((d >> 8) & 0xf))];
            }
            b.drop(int(d & 0xff));
            if (((d >> 12) & 0xf) != K_BASE)
                return false;
            const std::size_t off = (d >> 16) + b.take(int((d >> 8) & 0xf));
            if (off > std::size_t(op - out) || len > std::size_t(oend - op))
                return false;
            const unsigned char* src = op - off;
            if (off == 1) {
                std::memset(op, *src, len);
                op += len;
            } else if (off >= 8 && len + 8 <= std::size_t(oend - op)) {
                unsigned char* const stop = op + len;
                do {  // may write up to 7 bytes past `stop`, inside the buffer
                    std::memcpy(op, src, 8);
                    op += 8, src += 8;
                } while (op < stop);
                op = stop;
            } else {
                for (std::size_t i = 0; i < len; ++i)
                    op[i] = src[i];
                op += len;
            }
            b.refill();
            e = lit[b.peek(kLitBits)];
        }
    }
    if (op != oend)
        return false;
    // Adler-32 trailer, big endian, at the next byte boundary
    b.drop(b.left & 7);
    const unsigned char* t = b.in - (b.left >> 3);
    if (t + 4 > in + n_in)
        return false;
    const std::uint32_t want = (std::uint32_t(t[0]) << 24) | (std::uint32_t(t[1]) << 16) |
                               (std::uint32_t(t[2]) << 8) | std::uint32_t(t[3]);
    return adler32_fast(1, out, n_out) == want;
}

}  // namespace sens_loc::io
