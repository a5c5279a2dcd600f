// rt_png_host.cpp -- the deflate stream of save_image's PNG (base.py:143-151), host code only.
//
// The reference hands its 8-bit array to Pillow, which runs zlib level 6 (0.45 s per 1080p
// frame; level 1 still takes 42 ms on one core).  For animation batches the encoder, not the
// GPU, bounds the frame rate, so the rows are compressed here
// with what rendered frames need and nothing else, in the manner of fpng:
//   * PNG filter 1 (Sub) on every row: flat areas become zeros, gradients small residuals;
//   * matches only at distance 1 (a run of one byte) and distance 3 (a run of one RGB
//     residual): found by comparing with the bytes just behind, no hash chains;
//   * one dynamic-Huffman block per call, code lengths from the band's own histogram
//     (length-limited by flattening the histogram when the tree is too deep);
//   * a stored block instead when that would be smaller.
// What a consumer compares is the decoded image, which does not depend on the encoder.
// A call compresses one band of rows into raw deflate; every band but the last ends on an
// empty stored block (byte-aligned, BFINAL = 0), so bands compressed concurrently concatenate
// into one stream (the pigz construction infrastructure/png.py already uses with zlib).
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/rt_abi.h"

extern "C" void rt_internal_set_error(const char* message);

namespace {

constexpr int kMaxLitCodes = 286, kDistCodes = 3, kClCodes = 19;
constexpr uint32_t kMatchFlag = 0x80000000u;

struct LengthTables {
    uint16_t code[259];
    uint8_t extra_bits[259];
    uint16_t extra_val[259];
    LengthTables() {
        static const int base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
        static const int bits[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
        for (int len = 3; len <= 258; ++len) {
            int k = 28;
            while (base[k] > len) --k;
            if (len != 258 && k == 28) k = 27;  // 258 has its own code; 227..257 belong to code 284
            code[len] = (uint16_t)(257 + k);
            extra_bits[len] = (uint8_t)bits[k];
            extra_val[len] = (uint16_t)(len - base[k]);
        }
    }
};
const LengthTables kLen;

// LSB-first bit writer without a flush branch: every put stores the 8-byte accumulator at the
// write position and advances by the whole bytes it now holds (the buffer has 8 bytes of slack).
struct BitWriter {
    uint8_t* out;
    size_t cap, pos = 0;
    uint64_t acc = 0;
    int nbits = 0;  // < 8 between puts
    bool overflow = false;
    BitWriter(uint8_t* o, size_t c) : out(o), cap(c) {}
    inline void put(uint32_t bits, int n) {  // n <= 32
        acc |= (uint64_t)bits << nbits;
        nbits += n;
        if (__builtin_expect(pos + 8 > cap, 0)) {
            overflow = true;
            pos = 0;  // keep going harmlessly; the caller reports the overflow
        }
        memcpy(out + pos, &acc, 8);  // little-endian hosts only (x86-64, aarch64)
        pos += (size_t)(nbits >> 3);
        acc >>= (nbits & ~7);
        nbits &= 7;
    }
    inline void align() {
        if (nbits > 0) put(0, 8 - nbits);
    }
    inline void bytes(const uint8_t* src, size_t n) {  // after align()
        if (pos + n + 8 > cap) {
            overflow = true;
            pos = 0;
            return;
        }
        memcpy(out + pos, src, n);
        pos += n;
    }
};

// Huffman code lengths (<= limit) of the symbols with freq > 0.  Plain two-queue Huffman on the
// sorted frequencies; a tree deeper than `limit` is rebuilt from a flattened histogram
// (f -> (f + 1) / 2), which terminates at all-ones (depth ceil(log2 n) <= 9 for 286 symbols).
void huffman_lengths(const uint32_t* freq_in, int n, int limit, uint8_t* len) {
    std::vector<uint32_t> freq(freq_in, freq_in + n);
    for (;;) {
        struct Node {
            uint64_t weight;
            int left, right;
        };
        std::vector<std::pair<uint64_t, int>> leaves;
        for (int s = 0; s < n; ++s)
            if (freq[s] > 0) leaves.push_back({freq[s], s});
        for (int s = 0; s < n; ++s) len[s] = 0;
        if (leaves.empty()) return;
        if (leaves.size() == 1) {
            len[leaves[0].second] = 1;
            return;
        }
        std::sort(leaves.begin(), leaves.end());
        const int m = (int)leaves.size();
        std::vector<Node> nodes;
        nodes.reserve(2 * m);
        for (int i = 0; i < m; ++i) nodes.push_back({leaves[i].first, -1, -1});
        int leaf = 0, inner = m;  // next unused leaf / inner node (inner nodes come out sorted)
        auto take = [&]() {
            if (leaf < m && (inner >= (int)nodes.size() || nodes[leaf].weight <= nodes[inner].weight)) return leaf++;
            return inner++;
        };
        for (int i = 0; i < m - 1; ++i) {
            const int a = take(), b = take();
            nodes.push_back({nodes[a].weight + nodes[b].weight, a, b});
        }
        // depths, root = last node
        std::vector<int> depth(nodes.size(), 0);
        int deepest = 0;
        for (int i = (int)nodes.size() - 1; i >= m; --i) {
            depth[nodes[i].left] = depth[i] + 1;
            depth[nodes[i].right] = depth[i] + 1;
        }
        for (int i = 0; i < m; ++i) deepest = std::max(deepest, depth[i]);
        if (deepest <= limit) {
            for (int i = 0; i < m; ++i) len[leaves[i].second] = (uint8_t)depth[i];
            return;
        }
        for (int s = 0; s < n; ++s)
            if (freq[s] > 0) freq[s] = (freq[s] + 1) / 2;
    }
}

// Canonical codes of RFC 1951 section 3.2.2, bit-reversed for the LSB-first writer.
void canonical_codes(const uint8_t* len, int n, uint16_t* code) {
    int count[16] = {0};
    for (int s = 0; s < n; ++s) count[len[s]]++;
    count[0] = 0;
    int next[16] = {0};
    int c = 0;
    for (int b = 1; b <= 15; ++b) {
        c = (c + count[b - 1]) << 1;
        next[b] = c;
    }
    for (int s = 0; s < n; ++s) {
        const int l = len[s];
        if (l == 0) {
            code[s] = 0;
            continue;
        }
        int v = next[l]++, r = 0;
        for (int i = 0; i < l; ++i) {
            r = (r << 1) | (v & 1);
            v >>= 1;
        }
        code[s] = (uint16_t)r;
    }
}

// Over a run of n <= 5552 bytes:  a' = a + S1,  b' = b + n a + n S1 - S2  with S1 = sum p[i] and
// S2 = sum i p[i] (< 2^32 for n <= 5552): two plain reductions the compiler vectorises.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
__attribute__((target_clones("avx2", "default")))  // resolved once at load time
#endif
uint32_t adler32_update(uint32_t adler, const uint8_t* p, size_t n) {
    uint32_t a = adler & 0xffffu, b = adler >> 16;
    while (n > 0) {
        const uint32_t chunk = (uint32_t)(n < 5552 ? n : 5552);
        uint32_t s1 = 0, s2 = 0;
        for (uint32_t i = 0; i < chunk; ++i) {
            s1 += p[i];
            s2 += i * (uint32_t)p[i];
        }
        const uint64_t bb = (uint64_t)b + (uint64_t)chunk * a + (uint64_t)chunk * s1 - s2;
        a = (a + s1) % 65521u;
        b = (uint32_t)(bb % 65521u);
        p += chunk;
        n -= chunk;
    }
    return (b << 16) | a;
}

// number of bytes from p[0] on that equal the byte `dist` behind them, at most `cap`
inline int run_length(const uint8_t* p, int dist, int cap) {
    int k = 0;
    while (k + 8 <= cap) {
        uint64_t x, y;
        memcpy(&x, p + k, 8);
        memcpy(&y, p + k - dist, 8);
        const uint64_t diff = x ^ y;
        if (diff != 0) return k + (__builtin_ctzll(diff) >> 3);
        k += 8;
    }
    while (k < cap && p[k] == p[k - dist]) ++k;
    return k;
}

struct Scratch {
    uint8_t* filtered = nullptr;
    uint32_t* tokens = nullptr;
    size_t capacity = 0;
    bool reserve(size_t raw) {
        if (raw <= capacity) return true;
        free(filtered);
        free(tokens);
        capacity = 0;
        filtered = static_cast<uint8_t*>(malloc(raw + 16));  // slack for the 4- and 8-byte compares
        tokens = static_cast<uint32_t*>(malloc((raw + 16) * sizeof(uint32_t)));
        if (filtered == nullptr || tokens == nullptr) {
            free(filtered);
            free(tokens);
            filtered = nullptr;
            tokens = nullptr;
            return false;
        }
        capacity = raw;
        return true;
    }
    ~Scratch() {
        free(filtered);
        free(tokens);
    }
};

void write_stored(BitWriter& bw, const uint8_t* data, size_t n, bool last) {
    size_t off = 0;
    do {
        const size_t chunk = std::min<size_t>(n - off, 65535);
        const bool final_block = last && off + chunk == n;
        bw.put(final_block ? 1u : 0u, 1);
        bw.put(0u, 2);
        bw.align();
        const uint8_t head[4] = {(uint8_t)(chunk & 255), (uint8_t)(chunk >> 8), (uint8_t)(~chunk & 255), (uint8_t)((~chunk >> 8) & 255)};
        bw.bytes(head, 4);
        bw.bytes(data + off, chunk);
        off += chunk;
    } while (off < n);
}

}  // namespace

extern "C" {

int64_t rt_png_deflate_bound(int32_t width, int32_t rows) {
    if (width <= 0 || rows <= 0) return -1;
    const int64_t raw = (int64_t)rows * (1 + 3 * (int64_t)width);
    return raw + (raw / 65535 + 2) * 5 + 64;  // all stored blocks, the sync block, slack for the 32-bit writer
}

int rt_png_deflate_rows(const uint8_t* pixels, int32_t width, int32_t rows, int32_t pixel_bytes, int64_t row_stride,
                        int32_t last, uint8_t* out, int64_t capacity, int64_t* bytes_out, uint32_t* adler_out,
                        int64_t* raw_bytes_out) {
    if (pixels == nullptr || out == nullptr || bytes_out == nullptr || adler_out == nullptr || width <= 0 || rows <= 0 ||
        (pixel_bytes != 3 && pixel_bytes != 4) || row_stride < (int64_t)width * pixel_bytes ||
        capacity < rt_png_deflate_bound(width, rows)) {
        rt_internal_set_error("rt_png_deflate_rows: NULL pointer, bad size, pixel_bytes not 3 or 4, or capacity below rt_png_deflate_bound");
        return RT_ERR_BAD_ARG;
    }
    const size_t line = 1 + 3 * (size_t)width, raw = line * (size_t)rows;
    // per-thread scratch, kept between calls (an encoder pool compresses band after band of the
    // same size): the filtered rows and at most one token per byte
    thread_local Scratch scratch;
    if (!scratch.reserve(raw)) {
        rt_internal_set_error("rt_png_deflate_rows: out of host memory");
        return RT_ERR_UNSUPPORTED;
    }
    uint8_t* const filt = scratch.filtered;
    uint32_t* const tokens = scratch.tokens;
    size_t n_tokens = 0;
    // ---- filter 1 (Sub): byte - byte of the pixel to the left, first pixel as is
    for (int32_t r = 0; r < rows; ++r) {
        const uint8_t* src = pixels + (size_t)r * (size_t)row_stride;
        uint8_t* dst = filt + (size_t)r * line;
        *dst++ = 1;
        uint8_t pr = 0, pg = 0, pb = 0;
        for (int32_t x = 0; x < width; ++x) {
            const uint8_t cr = src[0], cg = src[1], cb = src[2];
            dst[0] = (uint8_t)(cr - pr);
            dst[1] = (uint8_t)(cg - pg);
            dst[2] = (uint8_t)(cb - pb);
            pr = cr;
            pg = cg;
            pb = cb;
            src += pixel_bytes;
            dst += 3;
        }
    }
    const uint8_t* data = filt;
    *adler_out = adler32_update(1u, data, raw);
    if (raw_bytes_out != nullptr) *raw_bytes_out = (int64_t)raw;

    // ---- tokens + histograms
    uint32_t lit_freq[kMaxLitCodes] = {0};
    uint32_t lit_hist[4][256] = {{0}};  // literals counted in four tables: no chain through one counter
    uint32_t dist_freq[kDistCodes] = {0};
    uint64_t extra_bits_total = 0;
    // A position starts a match when the four bytes there equal the four bytes 1 (or 3) behind:
    // one unaligned 32-bit compare each, runs shorter than four stay literals.
    size_t i = 0;
    for (; i < 3 && i < raw; ++i) {
        tokens[n_tokens++] = data[i];
        lit_hist[0][data[i]]++;
    }
    while (i + 5 <= raw) {
        // bytes i-3 .. i+4 in one load (little-endian): the three windows are shifts of it
        uint64_t window;
        memcpy(&window, data + i - 3, 8);
        const uint32_t here = (uint32_t)(window >> 24), back1 = (uint32_t)(window >> 16), back3 = (uint32_t)window;
        if (here != back1 && here != back3) {
            tokens[n_tokens++] = data[i];
            lit_hist[i & 3][data[i]]++;
            ++i;
            continue;
        }
        // a run of one byte is also a run at distance 3 from its third byte on: the period-3 run is
        // only looked for where the period-1 run does not start
        const int cap = (int)std::min<size_t>(258, raw - i);
        const int sel = here == back1 ? 0 : 1;
        const int best = run_length(data + i, sel ? 3 : 1, cap);
        tokens[n_tokens++] = kMatchFlag | ((uint32_t)sel << 16) | (uint32_t)best;
        lit_freq[kLen.code[best]]++;
        dist_freq[sel ? 2 : 0]++;
        extra_bits_total += kLen.extra_bits[best];
        i += (size_t)best;
    }
    for (; i < raw; ++i) {
        tokens[n_tokens++] = data[i];
        lit_hist[0][data[i]]++;
    }
    for (int v = 0; v < 256; ++v) lit_freq[v] = lit_hist[0][v] + lit_hist[1][v] + lit_hist[2][v] + lit_hist[3][v];
    lit_freq[256] = 1;  // end of block

    // ---- code lengths
    uint8_t lit_len[kMaxLitCodes], dist_len[kDistCodes] = {1, 0, 1};
    uint16_t lit_code[kMaxLitCodes], dist_code[kDistCodes];
    huffman_lengths(lit_freq, kMaxLitCodes, 15, lit_len);
    canonical_codes(lit_len, kMaxLitCodes, lit_code);
    canonical_codes(dist_len, kDistCodes, dist_code);
    int n_lit = kMaxLitCodes;
    while (n_lit > 257 && lit_len[n_lit - 1] == 0) --n_lit;
    // the code-length code: every length sent as itself (no repeat symbols 16-18)
    uint32_t cl_freq[kClCodes] = {0};
    for (int s = 0; s < n_lit; ++s) cl_freq[lit_len[s]]++;
    for (int s = 0; s < kDistCodes; ++s) cl_freq[dist_len[s]]++;
    uint8_t cl_len[kClCodes];
    uint16_t cl_code[kClCodes];
    huffman_lengths(cl_freq, kClCodes, 7, cl_len);  // lengths 0 and 1 always occur (the distance part): a complete code
    canonical_codes(cl_len, kClCodes, cl_code);

    uint64_t bits = 3 + 5 + 5 + 4 + 3 * kClCodes + extra_bits_total;
    for (int s = 0; s < n_lit; ++s) bits += (uint64_t)cl_len[lit_len[s]] + (uint64_t)lit_freq[s] * lit_len[s];
    for (int s = 0; s < kDistCodes; ++s) bits += (uint64_t)cl_len[dist_len[s]] + (uint64_t)dist_freq[s] * dist_len[s];
    const uint64_t stored_bytes = raw + (raw / 65535 + 1) * 5;

    BitWriter bw(out, (size_t)capacity);
    if ((bits + 7) / 8 >= stored_bytes) {
        write_stored(bw, data, raw, last != 0);
    } else {
        bw.put(last ? 1u : 0u, 1);
        bw.put(2u, 2);
        bw.put((uint32_t)(n_lit - 257), 5);
        bw.put((uint32_t)(kDistCodes - 1), 5);
        bw.put((uint32_t)(kClCodes - 4), 4);
        static const int order[kClCodes] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
        for (int k = 0; k < kClCodes; ++k) bw.put(cl_len[order[k]], 3);
        for (int s = 0; s < n_lit; ++s) bw.put(cl_code[lit_len[s]], cl_len[lit_len[s]]);
        for (int s = 0; s < kDistCodes; ++s) bw.put(cl_code[dist_len[s]], cl_len[dist_len[s]]);
        uint32_t lit_entry[kMaxLitCodes];  // code | length << 16: one load per symbol
        for (int s = 0; s < kMaxLitCodes; ++s) lit_entry[s] = (uint32_t)lit_code[s] | ((uint32_t)lit_len[s] << 16);
        for (size_t k = 0; k < n_tokens; ++k) {
            const uint32_t t = tokens[k];
            if (t & kMatchFlag) {
                const int len = (int)(t & 0xffffu);
                const uint32_t e = lit_entry[kLen.code[len]];
                // length code (<= 15 bits), its extra bits (<= 5), distance code (1 bit): one put
                const int dc = (t >> 16) & 1u ? 2 : 0;
                uint32_t bits = e & 0xffffu;
                int nb = (int)(e >> 16);
                bits |= (uint32_t)kLen.extra_val[len] << nb;
                nb += kLen.extra_bits[len];
                bits |= (uint32_t)dist_code[dc] << nb;
                nb += dist_len[dc];
                bw.put(bits, nb);
            } else {
                const uint32_t e = lit_entry[t];
                bw.put(e & 0xffffu, (int)(e >> 16));
            }
        }
        bw.put(lit_code[256], lit_len[256]);
    }
    if (!last) {  // empty stored block: byte-aligns the band so the next one can follow
        bw.put(0u, 1);
        bw.put(0u, 2);
        bw.align();
        const uint8_t sync[4] = {0, 0, 0xff, 0xff};
        bw.bytes(sync, 4);
    } else {
        bw.align();
    }
    if (bw.overflow) {
        rt_internal_set_error("rt_png_deflate_rows: output buffer too small");
        return RT_ERR_BAD_ARG;
    }
    *bytes_out = (int64_t)bw.pos;
    return RT_OK;
}

}  // extern "C"
