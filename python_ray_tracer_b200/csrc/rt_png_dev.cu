// rt_png_dev.cu -- the PNG save_image writes (base.py:143-151), encoded ON THE DEVICE.
//
// The reference hands its 8-bit array to Pillow (zlib level 6: 0.45 s per 1080p frame on one core).  Once the
// frame is traced in ~0.06 ms the 8-bit image itself is the bottleneck wherever it has to travel: 6.2 MB over
// PCIe to the host (0.11 ms), 87 MB over NVLink into rank 0 when eight GPUs share an 8K frame (0.11 ms at best),
// 2.5 ms of host deflate on eight cores.  So the file is made where the pixels are.  Rendered frames compress
// ~10:1 with very little work (rt_png_host.cpp does the same on the host):
//   * PNG filter 1 (Sub) on every row: flat areas become zeros, gradients small residuals;
//   * matches only at distance 3 (a run of one RGB residual; a run of one byte is one too from its fourth
//     byte on), at least 3 long, cut at 258: found by comparing with the byte three behind, no hashing;
//   * ONE fixed Huffman code (rt_png_table.h, made from the token statistics of rendered frames by
//     tools/png_static_table.py) sent in every band's dynamic-block header: no histogram, no tree building;
//   * a band that would grow is stored instead.
// The image is cut into bands of a few rows, one thread block each.  A band is one byte stream.  Whether a byte
// continues a run is one bit per byte, formed four bytes per compare; the bytes that bear tokens (literals, and the
// first byte of every run, which owns the run's tokens) are counted per 32-byte word, prefix-summed, and dealt to the
// threads in equal numbers; a run's length is bit arithmetic on its words; bit positions come from a second scan.  Every
// band becomes one IDAT chunk (PNG allows any number): its deflate block ends byte-aligned on an empty stored
// block (BFINAL = 0), its CRC-32 is computed by the block that made it, so nothing serial is left but the
// Adler-32 of the whole image (combined from the bands' sums) in a 9-byte tail chunk.  A second small kernel packs
// the chunks, header and tail into one contiguous file -- in device memory or straight into mapped host memory.
// What a consumer compares is the decoded image, which does not depend on the encoder (tests decode with zlib).
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../../include/rt_abi.h"
#include "rt_png_table.h"

extern "C" void rt_internal_set_error(const char* message);

namespace {

#ifndef RT_PNG_THREADS
#define RT_PNG_THREADS 256
#endif
#ifndef RT_PNG_BANDS_PER_SM
#define RT_PNG_BANDS_PER_SM 2  // bands per SM the image is cut into at least (rows_per_band)
#endif
constexpr int kThreads = RT_PNG_THREADS;  // per band (measured on B200, 1080p / 8K band kernel: 128: 65 / 300 us, 256: 44 / 240, 512: 43 / 269;
                                          // 4 bands per SM instead of 2: 42 / 240 with twice the chunks, 1: 68 / 240)
constexpr int kHeaderWords = 40;       // the dynamic-block header, LSB-first, at most 1280 bits
constexpr uint32_t kCrcPoly = 0xedb88320u;
constexpr uint32_t kAdlerMod = 65521u;
constexpr int kChunkOverhead = 12;     // length + type + CRC
constexpr int kBandSlack = 64;         // stored-block header, zlib header, sync block, rounding

// a(x) * b(x) mod the CRC-32 polynomial, reflected bit order (zlib crc32.c: multmodp)
__host__ __device__ inline uint32_t multmodp(uint32_t a, uint32_t b) {
    uint32_t m = 1u << 31, p = 0;
    for (;;) {
        if (a & m) {
            p ^= b;
            if ((a & (m - 1)) == 0) break;
        }
        m >>= 1;
        b = b & 1u ? (b >> 1) ^ kCrcPoly : b >> 1;
    }
    return p;
}
// x^(n * 2^k) mod the polynomial
__host__ __device__ inline uint32_t x2nmodp(const uint32_t* x2n, uint64_t n, unsigned k) {
    uint32_t p = 1u << 31;  // x^0
    while (n) {
        if (n & 1) p = multmodp(x2n[k & 31], p);
        n >>= 1;
        k++;
    }
    return p;
}

// Everything of the code a block needs, as a kernel parameter (constant bank), copied to shared memory for
// indexed access.  Entries: bits | count << 24, bits already reversed for the LSB-first stream.
struct PngCode {
    uint32_t lit[256];        // literal byte
    uint32_t match[256];      // match of length 3 + index: length symbol, its extra bits, the 1-bit distance code
    uint32_t eob;             // end of block
    uint32_t header[kHeaderWords];
    uint32_t header_bits;
    uint32_t x2n[32];         // x^(2^k) mod the CRC-32 polynomial (zlib's crc32_combine arithmetic)
};

struct BandJob {
    const uint8_t* pixels;    // first row of the launch's pixel buffer
    long long row_stride;     // bytes between rows of that buffer
    int pixel_bytes;          // 3 or 4 (alpha skipped)
    int width, height;        // of the IMAGE
    int rows_per_band;
    int n_rows;               // rows in the pixel buffer
    // local row r of the buffer is image row row0 + r + (r / strip_rows) * (strip_step - 1) * strip_rows
    // (a contiguous band: strip_step = 1; an interleaved shard of a frame rendered by several GPUs otherwise);
    // strip_rows is a multiple of rows_per_band, so a band never straddles two strips
    int row0, strip_rows, strip_step;
    uint8_t* slots;           // band b of the IMAGE writes its chunk at slots + b * slot_bytes (may be peer memory)
    long long slot_bytes;
    uint32_t* sizes;          // [n_bands_of_image] bytes of band b's chunk
    uint32_t* adlers;         // [n_bands_of_image] Adler-32 of band b's filtered bytes (as if it started a stream)
    // CRC-32 of a chunk by all threads: thread t takes `crc_slice` bytes (slices aligned to the chunk's END, so every
    // slice but the first is full), then a tree of crc32_combine steps  crc(A B) = crc(A) x^(8 |B|) + crc(B)  whose
    // right-hand sides are crc_slice * 2^level bytes long.  The multipliers are fixed per launch, so the products are
    // sums over the set bits of crc(A) of crc_mul[level][k] = x^k * x^(8 crc_slice 2^level) mod the polynomial (made on
    // the host): 32 independent AND / XORs against constant-bank operands instead of a 32-step shift-and-add loop.
    int crc_slice;
    uint32_t crc_mul[10][32];  // level l: right-hand sides of crc_slice * 2^l bytes
};

// ---------------------------------------------------------------------------------------------------------------
// host: the fixed code
// ---------------------------------------------------------------------------------------------------------------
void canonical(const uint8_t* len, int n, uint16_t* code) {  // RFC 1951 3.2.2, bit-reversed
    int count[16] = {0}, next[16] = {0};
    for (int s = 0; s < n; ++s) count[len[s]]++;
    count[0] = 0;
    int c = 0;
    for (int b = 1; b <= 15; ++b) {
        c = (c + count[b - 1]) << 1;
        next[b] = c;
    }
    for (int s = 0; s < n; ++s) {
        const int l = len[s];
        int v = l ? next[l]++ : 0, r = 0;
        for (int i = 0; i < l; ++i) {
            r = (r << 1) | (v & 1);
            v >>= 1;
        }
        code[s] = (uint16_t)r;
    }
}

struct HostBits {
    uint32_t words[kHeaderWords] = {0};
    uint32_t n = 0;
    bool overflow = false;
    void put(uint32_t bits, int count) {
        for (int i = 0; i < count; ++i, ++n) {
            if (n >= 32u * kHeaderWords) {
                overflow = true;
                return;
            }
            if ((bits >> i) & 1u) words[n >> 5] |= 1u << (n & 31);
        }
    }
};

const PngCode& png_code() {
    static PngCode code;
    static bool made = false;
    if (made) return code;
    static const int base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const int bits[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    uint16_t lit_code[286];
    canonical(rt::kPngLitLen, 286, lit_code);
    // distances: codes 0 and 2 with one bit each (a complete code); only code 2 (distance 3, no extra bits) is used
    const uint8_t dist_len[3] = {1, 0, 1};
    uint16_t dist_code[3];
    canonical(dist_len, 3, dist_code);
    for (int v = 0; v < 256; ++v) code.lit[v] = (uint32_t)lit_code[v] | ((uint32_t)rt::kPngLitLen[v] << 24);
    for (int len = 3; len <= 258; ++len) {
        int k = 28;
        while (base[k] > len) --k;
        if (len != 258 && k == 28) k = 27;  // 258 has its own symbol; 227..257 belong to symbol 284
        const int sym = 257 + k;
        uint32_t v = lit_code[sym];
        int n = rt::kPngLitLen[sym];
        v |= (uint32_t)(len - base[k]) << n;
        n += bits[k];
        v |= (uint32_t)dist_code[2] << n;
        n += dist_len[2];
        code.match[len - 3] = v | ((uint32_t)n << 24);
    }
    code.eob = (uint32_t)lit_code[256] | ((uint32_t)rt::kPngLitLen[256] << 24);
    // ---- the block header: BFINAL = 0, BTYPE = 2, the code lengths run-length coded (symbol 16: repeat the
    // previous length 3..6 times), themselves Huffman coded by a code built here
    struct Sym {
        int sym, extra, extra_bits;
    };
    static Sym seq[400];
    int n_seq = 0;
    uint8_t all[286 + 3];
    memcpy(all, rt::kPngLitLen, 286);
    memcpy(all + 286, dist_len, 3);
    for (int i = 0; i < 289;) {
        int run = 1;
        while (i + run < 289 && all[i + run] == all[i]) ++run;
        seq[n_seq++] = {all[i], 0, 0};
        int left = run - 1;
        if (all[i] == 0) {  // (only the unused distance code: a single zero)
            i += 1;
            continue;
        }
        while (left >= 3) {
            const int r = left > 6 ? 6 : left;
            seq[n_seq++] = {16, r - 3, 2};
            left -= r;
        }
        for (; left > 0; --left) seq[n_seq++] = {all[i], 0, 0};
        i += run;
    }
    uint32_t freq[19] = {0};
    for (int i = 0; i < n_seq; ++i) freq[seq[i].sym]++;
    // code-length code (<= 7 bits): 19 symbols, plain Huffman by repeated merging of the two rarest; a tree
    // deeper than 7 is rebuilt from a flattened histogram
    uint8_t cl_len[19] = {0};
    for (bool fits = false; !fits;) {
        memset(cl_len, 0, sizeof(cl_len));
        struct Node {
            uint32_t w;
            int left, right;
        } nodes[40];
        int alive[40], n_alive = 0, n_nodes = 0;
        for (int s = 0; s < 19; ++s)
            if (freq[s]) {
                nodes[n_nodes] = {freq[s], -1 - s, 0};
                alive[n_alive++] = n_nodes++;
            }
        while (n_alive > 1) {
            int a = 0, b = 1;
            if (nodes[alive[b]].w < nodes[alive[a]].w) a = 1, b = 0;
            for (int i = 2; i < n_alive; ++i) {
                if (nodes[alive[i]].w < nodes[alive[a]].w) b = a, a = i;
                else if (nodes[alive[i]].w < nodes[alive[b]].w) b = i;
            }
            nodes[n_nodes] = {nodes[alive[a]].w + nodes[alive[b]].w, alive[a], alive[b]};
            const int hi = a > b ? a : b, lo = a > b ? b : a;
            alive[hi] = alive[--n_alive];
            alive[lo] = n_nodes++;
        }
        int depth[40] = {0};
        for (int i = n_nodes - 1; i >= 0; --i)
            if (nodes[i].left >= 0) depth[nodes[i].left] = depth[nodes[i].right] = depth[i] + 1;
            else cl_len[-1 - nodes[i].left] = (uint8_t)(depth[i] ? depth[i] : 1);
        fits = true;
        for (int k = 0; k < 19; ++k) fits = fits && cl_len[k] <= 7;
        if (!fits)
            for (int k = 0; k < 19; ++k) freq[k] = freq[k] ? (freq[k] + 1) / 2 : 0;
    }
    uint16_t cl_code[19];
    canonical(cl_len, 19, cl_code);
    HostBits hb;
    hb.put(0u, 1);
    hb.put(2u, 2);
    hb.put(286 - 257, 5);
    hb.put(3 - 1, 5);
    hb.put(19 - 4, 4);
    static const int order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    for (int k = 0; k < 19; ++k) hb.put(cl_len[order[k]], 3);
    for (int i = 0; i < n_seq; ++i) {
        hb.put(cl_code[seq[i].sym], cl_len[seq[i].sym]);
        if (seq[i].extra_bits) hb.put((uint32_t)seq[i].extra, seq[i].extra_bits);
    }
    memcpy(code.header, hb.words, sizeof(code.header));
    code.header_bits = hb.overflow ? 0u : hb.n;  // 0 = the header did not fit: every band is stored
    uint32_t p = 1u << 30;  // x^1
    code.x2n[0] = p;
    for (int k = 1; k < 32; ++k) code.x2n[k] = p = multmodp(p, p);
    made = true;
    return code;
}

// ---------------------------------------------------------------------------------------------------------------
// device: one band -> one IDAT chunk
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int job_image_row(const BandJob& j, int local_row) {
    return j.strip_step == 1 ? j.row0 + local_row
                             : j.row0 + local_row + (local_row / j.strip_rows) * (j.strip_step - 1) * j.strip_rows;
}

// exclusive prefix sum over the block's kThreads values; *total = their sum.  `scratch`: kThreads / 32 + 1 words.
__device__ __forceinline__ uint32_t block_exclusive_sum(uint32_t v, uint32_t* scratch, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += up;
    }
    __syncthreads();  // scratch may still be read from the previous use
    if (lane == 31) scratch[warp] = inc;
    __syncthreads();
    uint32_t before = 0, all = 0;
    for (int w = 0; w < kThreads / 32; ++w) {
        const uint32_t s = scratch[w];
        if (w < warp) before += s;
        all += s;
    }
    *total = all;
    return before + inc - v;
}
__device__ __forceinline__ unsigned long long block_sum64(unsigned long long v, unsigned long long* scratch) {
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long all = 0;
    for (int w = 0; w < kThreads / 32; ++w) all += scratch[w];
    return all;
}

// ---------------------------------------------------------------------------------------------------------------
// device: the chunks of all bands -> one contiguous PNG file
// ---------------------------------------------------------------------------------------------------------------
struct PackJob {
    const uint8_t* slots;
    long long slot_bytes;
    const uint32_t* sizes;
    const uint32_t* adlers;
    int n_bands;
    int band_bytes;        // filtered bytes of a full band (rows_per_band * (1 + 3 width))
    long long raw_bytes;   // filtered bytes of the whole image
    uint8_t head[33];      // signature + IHDR chunk
    uint8_t* out;          // the file: device memory or mapped host memory
    long long capacity;
    unsigned long long* bytes_out;  // its size (0 if it did not fit)
};

// One launch for a whole image (png_band_kernel<true>): no pack kernel, no slots.  Blocks take their band from a
// ticket, so a block only ever waits for blocks that started before it; a band's chunk size is known half-way through
// its block (once its bits are counted) and published at once (status 1 << 32 | size), and the position of the chunk in
// the file is the sum of the sizes before it, which the block adds up when its chunk is finished.  The block that
// finishes last writes the tail.
struct FusedCtl {
    unsigned int* ticket;        // zeroed before the launch
    unsigned int* done;
    unsigned long long* state;   // [n_bands] status << 32 | value, zeroed before the launch
};

// `size` bytes from src (16-byte aligned, readable one word past the end) to dst (any alignment): 16-byte stores
// wherever the DESTINATION is aligned, whatever that makes of the source (a chunk lands at any byte offset, and three in
// four of them were copied byte by byte -- over PCIe when the file is mapped host memory): bytes up to the first aligned
// address, then units built from five source words by funnel shifts, then the rest
__device__ __forceinline__ void copy_chunk(uint8_t* dst, const uint8_t* src, int size, int t) {
    const int head = min(size, (int)((16u - (unsigned)((uintptr_t)dst & 15u)) & 15u));
    for (int i = t; i < head; i += kThreads) dst[i] = src[i];
    const int units = (size - head) >> 4;
    const unsigned int shift = 8u * ((unsigned)head & 3u);
    const uint32_t* s4 = reinterpret_cast<const uint32_t*>(src) + (head >> 2);
    uint4* d16 = reinterpret_cast<uint4*>(dst + head);
    for (int u = t; u < units; u += kThreads) {
        const uint32_t* w = s4 + 4 * u;
        const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = w[4];
        d16[u] = make_uint4(__funnelshift_r(w0, w1, shift), __funnelshift_r(w1, w2, shift), __funnelshift_r(w2, w3, shift),
                            __funnelshift_r(w3, w4, shift));
    }
    for (int i = head + 16 * units + t; i < size; i += kThreads) dst[i] = src[i];
}

// The end of the file, by one whole block: Adler-32 of the image from the bands' sums, the tail chunk, IEND, the size.
// A band that starts at offset O in a stream of N bytes, with sums (a, b) taken as if it started a stream
// and n bytes, adds a - 1 to A and (b - n) + (N - O - n) (a - 1) to B;  A = 1 + sum, B = N + sum.
// `all` = bytes of all the bands' chunks.
__device__ __forceinline__ void write_tail(const PackJob& job, unsigned long long all, unsigned long long* s_sum) {
    const int t = threadIdx.x;
    const unsigned long long total = 33ull + all + 21ull + 12ull;  // head, bands, tail chunk (9 data bytes), IEND
    const bool fits = total <= (unsigned long long)job.capacity;
    unsigned long long A = 0, B = 0;
    for (int j = t; j < job.n_bands; j += kThreads) {
        const uint32_t ad = job.adlers[j];
        const unsigned long long a1 = ((ad & 0xffffu) + kAdlerMod - 1u) % kAdlerMod, bj = ad >> 16;
        const long long start = (long long)j * job.band_bytes;
        const long long n = (j == job.n_bands - 1) ? job.raw_bytes - start : job.band_bytes;
        const unsigned long long after = (unsigned long long)(job.raw_bytes - start - n) % kAdlerMod;
        A += a1;
        B += (bj + kAdlerMod - (unsigned long long)(n % kAdlerMod)) % kAdlerMod + (after * a1) % kAdlerMod;
    }
    A = block_sum64(A, s_sum);
    B = block_sum64(B, s_sum);
    if (t != 0) return;
    if (!fits) {
        *job.bytes_out = 0ull;
        return;
    }
    const uint32_t adler_a = (uint32_t)((1ull + A) % kAdlerMod);
    const uint32_t adler_b = (uint32_t)(((unsigned long long)(job.raw_bytes % kAdlerMod) + B) % kAdlerMod);
    uint8_t tail[21 + 12] = {0, 0, 0, 9, 'I', 'D', 'A', 'T',
                             0x01, 0x00, 0x00, 0xff, 0xff,  // the final block: stored, empty, BFINAL = 1
                             (uint8_t)(adler_b >> 8), (uint8_t)adler_b, (uint8_t)(adler_a >> 8), (uint8_t)adler_a,
                             0, 0, 0, 0,
                             0, 0, 0, 0, 'I', 'E', 'N', 'D', 0xae, 0x42, 0x60, 0x82};
    uint32_t c = 0xffffffffu;
    for (int i = 4; i < 17; ++i) {
        c ^= tail[i];
        for (int k = 0; k < 8; ++k) c = c & 1u ? (c >> 1) ^ kCrcPoly : c >> 1;
    }
    c ^= 0xffffffffu;
    tail[17] = (uint8_t)(c >> 24), tail[18] = (uint8_t)(c >> 16), tail[19] = (uint8_t)(c >> 8), tail[20] = (uint8_t)c;
    uint8_t* dst = job.out + 33 + all;
    for (int i = 0; i < 33; ++i) dst[i] = tail[i];
    __threadfence_system();
    *job.bytes_out = total;
}

// ---- tokens dealt evenly to the threads ---------------------------------------------------------------------------
// Bytes are cheap, tokens are not, and tokens are unevenly spread: a thread that owns 46 contiguous bytes of a
// gradient emits 46 literals while its neighbour's 46 bytes are the inside of one run -- and a warp takes as long as
// its slowest thread.  So the band's run bits m(i) are formed word-parallel (32 bytes per thread and step, four bytes
// per compare), the token-bearing bytes -- the literals, m = 0, and the first byte of every run, which owns all the
// tokens of its run -- are counted per 32-byte word and prefix-summed, and every thread takes the same NUMBER of them.
struct TokenCursor {  // where a thread's tokens begin: a word of the run-bit array and what is left of its candidates
    int word;
    uint32_t left;
};
// candidates of word w: bytes that do not continue a run, and bytes that start one
__device__ __forceinline__ uint32_t token_candidates(const uint32_t* m, int w, int n) {
    const uint32_t M = m[w];
    const uint32_t carry = w > 0 ? m[w - 1] >> 31 : 0u;
    uint32_t c = ~M | (M & ~((M << 1) | carry));
    const int valid = n - 32 * w;
    if (valid < 32) c &= (1u << valid) - 1u;
    return c;
}
// The tokens of `count` candidates from `cur` on, in stream order: lit(entry) for one token, rep(entry, k) for k equal
// ones (the 258-byte pieces of a long run).  m[] ends in a zero word, so every run ends; next[w] = the first word at or
// after w that is not all ones, so neither a run's length nor the way to the next candidate is walked word by word.
template <class Lit, class Rep>
__device__ __forceinline__ void deal_tokens(const uint8_t* S, const uint32_t* m, const uint16_t* next, int n, TokenCursor cur, int count,
                                            const uint32_t* lit_table, const uint32_t* match_table, Lit&& lit, Rep&& rep) {
    int w = cur.word;
    uint32_t c = cur.left;
    for (int j = 0; j < count; ++j) {
        if (c == 0u) {
            // the next word with a candidate: the one after this, unless that lies wholly inside a run (all ones and
            // entered from a one) -- then the first word that is not all ones, however far (next[])
            ++w;
            if (m[w] == 0xffffffffu && (m[w - 1] >> 31) != 0u) w = next[w];
            c = token_candidates(m, w, n);
        }
        const int b = __ffs(c) - 1;
        c &= c - 1u;
        const int i = 32 * w + b;
        const uint32_t M = m[w];
        if (((M >> b) & 1u) == 0u) {
            lit(lit_table[S[i]]);
            continue;
        }
        // a run starts here: its length, however many words it spans
        const uint32_t zeros = ~(M >> b);  // (the shift brings in zeros: the run is at most 32 - b long here; a full word at b = 0 has none)
        int len = zeros != 0u ? __ffs(zeros) - 1 : 32;
        if (len == 32 - b) {
            const int w2 = next[w + 1];
            len += 32 * (w2 - (w + 1)) + __ffs(~m[w2]) - 1;
        }
        const int whole = len / 258, head = len - 258 * whole;
        if (head >= 3) {
            lit(match_table[head - 3]);
        } else if (head >= 1) {  // too short for a match: one or two literals
            lit(lit_table[S[i]]);
            if (head == 2) lit(lit_table[S[i + 1]]);
        }
        if (whole > 0) rep(match_table[255], whole);
    }
}

// c(x) * g(x) mod the CRC-32 polynomial for a multiplier fixed per launch: T[k] = x^k g(x), bit 31 - k of c selects it
__device__ __forceinline__ uint32_t crc_mul_fixed(const uint32_t (&T)[32], uint32_t c) {
    uint32_t r = 0u;
#pragma unroll
    for (int k = 0; k < 32; ++k) r ^= (uint32_t)((int32_t)(c << k) >> 31) & T[k];
    return r;
}

template <bool FUSED>
__global__ void __launch_bounds__(kThreads) png_band_kernel(const __grid_constant__ PngCode code, const __grid_constant__ BandJob job,
                                                            const __grid_constant__ PackJob file, const __grid_constant__ FusedCtl ctl) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint32_t s_lit[256], s_match[256], s_crc[256], s_wcrc[1];
    __shared__ uint32_t s_scan[kThreads / 32 + 1];
    __shared__ unsigned long long s_sum[kThreads / 32];
    __shared__ uint32_t s_wsum[kThreads / 32];
    __shared__ uint32_t s_coff[kThreads + 1];
    __shared__ unsigned int s_ticket, s_last;

    const int t = threadIdx.x;
    const int line = 1 + 3 * job.width;
    unsigned int block_index = blockIdx.x;
    if (FUSED) {  // bands in the order the blocks start: a block only ever waits for blocks that are already running
        if (t == 0) s_ticket = atomicAdd(ctl.ticket, 1u);
        __syncthreads();
        block_index = s_ticket;
    }
    const int first_local = (int)block_index * job.rows_per_band;
    const int first_image = job_image_row(job, first_local);
    const int band = first_image / job.rows_per_band;
    const int rows = min(job.rows_per_band, min(job.n_rows - first_local, job.height - first_image));
    const int n = rows * line;                               // bytes of the band's stream
    const int cap = job.rows_per_band * line;
    uint8_t* filt = smem;                                    // [cap] the filtered rows
    uint8_t* out = smem + ((cap + 15) & ~15);                // the chunk: length, type, data, CRC
    uint32_t* out_words = reinterpret_cast<uint32_t*>(out);
    const int out_words_n = (cap + kBandSlack + kChunkOverhead + 3) / 4;

    for (int i = t; i < 256; i += kThreads) {
        s_lit[i] = code.lit[i];
        s_match[i] = code.match[i];
        uint32_t c = (uint32_t)i;
        for (int k = 0; k < 8; ++k) c = c & 1u ? (c >> 1) ^ kCrcPoly : c >> 1;
        s_crc[i] = c;
    }

    // ---- filter 1 (Sub): byte minus the byte of the pixel to the left; Adler-32 sums on the way.
    // Packed RGB rows are first brought into shared memory (the `out` area is free until the tokens are
    // written) with whole-word loads that are all in flight at once; a filter loop that reads its two bytes from
    // global memory waits a memory latency per byte.
    unsigned long long s1 = 0, s2 = 0;
    const int row_bytes = 3 * job.width;
    const bool staged = job.pixel_bytes == 3 && job.row_stride == (long long)row_bytes;
    if (staged) {
        const uint8_t* src = job.pixels + (long long)first_local * job.row_stride;
        const int total = rows * row_bytes;
        if ((((uintptr_t)src) & 3) == 0) {
            const uint32_t* src4 = reinterpret_cast<const uint32_t*>(src);
            for (int i = t; i < total / 4; i += kThreads) out_words[i] = __ldg(src4 + i);
            for (int i = (total & ~3) + t; i < total; i += kThreads) out[i] = __ldg(src + i);
        } else {
            for (int i = t; i < total; i += kThreads) out[i] = __ldg(src + i);
        }
        __syncthreads();
    }
    for (int r = 0; r < rows; ++r) {
        const uint8_t* src = job.pixels + (long long)(first_local + r) * job.row_stride;
        const uint8_t* raw = out + r * row_bytes;
        uint8_t* dst = filt + r * line;
        if (t == 0) {
            dst[0] = 1;
            s1 += 1u;
            s2 += (unsigned long long)(r * line);
        }
        if (staged && (row_bytes & 3) == 0) {
            // four bytes a step: the word against the funnel of its left neighbour and itself (the bytes three back), a
            // per-byte subtract, the two Adler sums by dot products with (1,1,1,1) and (0,1,2,3)
            const uint32_t* raw4 = out_words + r * (row_bytes >> 2);
            for (int k = t; k < (row_bytes >> 2); k += kThreads) {
                const uint32_t cur4 = raw4[k], prev4 = k > 0 ? raw4[k - 1] : 0u;
                const uint32_t v4 = __vsub4(cur4, __funnelshift_r(prev4, cur4, 8));
                uint8_t* d = dst + 1 + 4 * k;
                d[0] = (uint8_t)v4, d[1] = (uint8_t)(v4 >> 8), d[2] = (uint8_t)(v4 >> 16), d[3] = (uint8_t)(v4 >> 24);
                const uint32_t sum4 = __dp4a(v4, 0x01010101u, 0u);
                s1 += sum4;
                s2 += (unsigned long long)(r * line + 1 + 4 * k) * sum4 + __dp4a(v4, 0x03020100u, 0u);
            }
            continue;
        }
        for (int x = t; x < row_bytes; x += kThreads) {
            uint8_t cur, left;
            if (staged) {
                cur = raw[x];
                left = x >= 3 ? raw[x - 3] : (uint8_t)0;
            } else {
                const int px = x / 3, ch = x - 3 * px;
                cur = __ldg(src + (long long)px * job.pixel_bytes + ch);
                left = px > 0 ? __ldg(src + (long long)(px - 1) * job.pixel_bytes + ch) : (uint8_t)0;
            }
            const uint8_t v = (uint8_t)(cur - left);
            dst[1 + x] = v;
            s1 += v;
            s2 += (unsigned long long)(r * line + 1 + x) * v;
        }
    }
    s1 = block_sum64(s1, s_sum);
    s2 = block_sum64(s2, s_sum);  // (block_sum64 synchronises: filt is complete from here on, the staged rows are done with)
    for (int i = t; i < out_words_n; i += kThreads) out_words[i] = 0u;  // (visible to all before the first token: syncs below)
    const uint32_t adler_a = (uint32_t)((1ull + s1) % kAdlerMod);
    const uint32_t adler_b = (uint32_t)((((unsigned long long)n % kAdlerMod) + ((unsigned long long)n % kAdlerMod) * (s1 % kAdlerMod) +
                                         (unsigned long long)kAdlerMod - (s2 % kAdlerMod)) % kAdlerMod);

    // ---- run bits, 32 bytes per thread and step: byte i against byte i - 3 is word k against the funnel of words
    // k - 1 and k, four bytes per compare (filt is 16-byte aligned; the bytes past n are masked off)
    uint32_t* s_m = reinterpret_cast<uint32_t*>(out + (((cap + kBandSlack + kChunkOverhead + 3) / 4) * 4 + 16));
    const int n_words = (n + 31) >> 5;
    {
        const uint32_t* F = reinterpret_cast<const uint32_t*>(filt);
        const int n_f = (n + 3) >> 2;
        for (int w = t; w < n_words; w += kThreads) {
            uint32_t bits = 0u, prev = w > 0 ? F[8 * w - 1] : 0u;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const uint32_t curw = 8 * w + k < n_f ? F[8 * w + k] : 0u;
                const uint32_t eq = __vcmpeq4(curw, __funnelshift_r(prev, curw, 8));
                bits |= (((eq & 0x08040201u) * 0x01010101u) >> 24) << (4 * k);
                prev = curw;
            }
            if (w == 0) bits &= ~7u;  // (the first three bytes have nothing three behind them)
            const int valid = n - 32 * w;
            if (valid < 32) bits &= (1u << valid) - 1u;
            s_m[w] = bits;
        }
        if (t == 0) s_m[n_words] = 0u;
    }
    uint16_t* s_next = reinterpret_cast<uint16_t*>(s_m + (cap + 31) / 32 + 2);  // [n_words + 1]
    __syncthreads();
    // ---- candidates per thread's words, their prefix sum, and every thread's equal share of them
    const int words_each = (n_words + kThreads - 1) / kThreads;
    const int w_first = min(t * words_each, n_words), w_last = min(w_first + words_each, n_words);
    {
        // next[w] (see deal_tokens): the first word of the thread's own range that is not all ones, then the minimum
        // over the ranges to the right (a suffix scan: shuffles inside the warps, a walk over the warps to the right)
        int own = n_words;  // (the end word is zero: it is "not all ones" for every run that reaches it)
        for (int w = w_last - 1; w >= w_first; --w)
            if (s_m[w] != 0xffffffffu) own = w;
        int suffix = own;
        for (int d = 1; d < 32; d <<= 1) {
            const int other = __shfl_down_sync(0xffffffffu, suffix, d);
            if ((t & 31) + d < 32) suffix = min(suffix, other);
        }
        if ((t & 31) == 0) s_wsum[t >> 5] = (uint32_t)suffix;
        int beyond = __shfl_down_sync(0xffffffffu, suffix, 1);  // the ranges to the right of this thread's, inside its warp
        if ((t & 31) == 31) beyond = n_words;
        __syncthreads();
        for (int wp = (t >> 5) + 1; wp < kThreads / 32; ++wp) beyond = min(beyond, (int)s_wsum[wp]);
        int running = beyond;
        for (int w = w_last - 1; w >= w_first; --w) {
            if (s_m[w] != 0xffffffffu) running = w;
            s_next[w] = (uint16_t)running;
        }
        if (t == 0) s_next[n_words] = (uint16_t)n_words;
    }
    uint32_t mine = 0u;
    for (int w = w_first; w < w_last; ++w) mine += (uint32_t)__popc(token_candidates(s_m, w, n));
    uint32_t n_candidates = 0u;
    const uint32_t before_mine = block_exclusive_sum(mine, s_scan, &n_candidates);
    s_coff[t] = before_mine;
    if (t == 0) s_coff[kThreads] = n_candidates;
    __syncthreads();
    const int share = ((int)n_candidates + kThreads - 1) / kThreads;
    const int c_first = min(t * share, (int)n_candidates), c_count = min(share, (int)n_candidates - c_first);
    TokenCursor cursor = {0, 0u};
    if (c_count > 0) {
        int lo = 0, hi = kThreads;  // the thread u whose words hold candidate c_first: s_coff[u] <= c_first < s_coff[u + 1]
        while (hi - lo > 1) {
            const int midp = (lo + hi) >> 1;
            if (s_coff[midp] <= (uint32_t)c_first) lo = midp; else hi = midp;
        }
        int skip = c_first - (int)s_coff[lo];
        int w = min(lo * words_each, n_words);
        uint32_t c = token_candidates(s_m, w, n);
        for (int pc = __popc(c); skip >= pc; pc = __popc(c)) {
            skip -= pc;
            c = token_candidates(s_m, ++w, n);
        }
        for (; skip > 0; --skip) c &= c - 1u;
        cursor.word = w, cursor.left = c;
    }
    // ---- bit positions
    uint32_t my_bits = 0;
    deal_tokens(filt, s_m, s_next, n, cursor, c_count, s_lit, s_match, [&](uint32_t entry) { my_bits += entry >> 24; },
                [&](uint32_t entry, int k) { my_bits += (uint32_t)k * (entry >> 24); });
    uint32_t token_bits = 0;
    const uint32_t my_bit = block_exclusive_sum(my_bits, s_scan, &token_bits);
    const int prefix = band == 0 ? 2 : 0;  // the zlib header travels in the first band's chunk
    const int data0 = 8 + prefix;
    const uint32_t block_bits = code.header_bits + token_bits + (code.eob >> 24) + 3u;  // + the sync block's header
    const uint32_t deflated = (block_bits + 7u) / 8u + 4u;  // ... its LEN / NLEN
    const bool stored = code.header_bits == 0u || deflated >= (uint32_t)n + 5u;
    int data_len;
    auto put_bits = [&](uint32_t bit, uint32_t entry) {  // a token of at most 24 bits at an absolute bit position
        const unsigned long long v = (unsigned long long)(entry & 0xffffffu) << (bit & 31u);
        atomicOr(out_words + (bit >> 5), (uint32_t)v);
        if ((uint32_t)(v >> 32) != 0u) atomicOr(out_words + (bit >> 5) + 1, (uint32_t)(v >> 32));
    };
    if (!stored) {
        const uint32_t bit0 = 8u * (uint32_t)data0;
        if (t < kHeaderWords && code.header[t] != 0u) {
            // the header starts byte-aligned at bit0, which need not be word-aligned
            const unsigned long long v = (unsigned long long)code.header[t] << (bit0 & 31u);
            const uint32_t w = (bit0 >> 5) + (uint32_t)t;
            atomicOr(out_words + w, (uint32_t)v);
            if ((uint32_t)(v >> 32) != 0u) atomicOr(out_words + w + 1, (uint32_t)(v >> 32));
        }
        // the thread's tokens in stream order, each at its own bit position
        uint32_t bit = bit0 + code.header_bits + my_bit;
        deal_tokens(filt, s_m, s_next, n, cursor, c_count, s_lit, s_match,
                    [&](uint32_t entry) {
                        put_bits(bit, entry);
                        bit += entry >> 24;
                    },
                    [&](uint32_t entry, int k) {
                        for (; k > 0; --k) {
                            put_bits(bit, entry);
                            bit += entry >> 24;
                        }
                    });
        if (t == 0) put_bits(bit0 + code.header_bits + token_bits, code.eob);
        data_len = prefix + (int)deflated;
    } else {
        for (int i = t; i < n; i += kThreads) out[data0 + 5 + i] = filt[i];
        data_len = prefix + 5 + n;
    }
    __syncthreads();
    if (t == 0) {
        if (prefix) out[8] = 0x78, out[9] = 0x01;  // zlib header: deflate, 32 KB window, no preset dictionary, fastest
        if (stored) {
            out[data0] = 0;  // BFINAL = 0, BTYPE = 00
            out[data0 + 1] = (uint8_t)(n & 255), out[data0 + 2] = (uint8_t)(n >> 8);
            out[data0 + 3] = (uint8_t)(~n & 255), out[data0 + 4] = (uint8_t)((~n >> 8) & 255);
        } else {
            // the empty stored block that byte-aligns the band: its three header bits are zero already
            uint8_t* sync = out + 8 + data_len - 4;
            sync[0] = 0, sync[1] = 0, sync[2] = 0xff, sync[3] = 0xff;
        }
        out[0] = (uint8_t)(data_len >> 24), out[1] = (uint8_t)(data_len >> 16), out[2] = (uint8_t)(data_len >> 8), out[3] = (uint8_t)data_len;
        out[4] = 'I', out[5] = 'D', out[6] = 'A', out[7] = 'T';
        if (FUSED)  // the chunk's size is known: the bands after this one can place themselves
            *reinterpret_cast<volatile unsigned long long*>(ctl.state + band) = (1ull << 32) | (unsigned long long)(kChunkOverhead + data_len);
    }
    __syncthreads();

    // ---- CRC-32 of type + data by every thread (BandJob::crc_slice): a band of noise or fine gradients is a chunk of
    // several KB, and one warp walking it byte by byte (two dependent shared-memory loads per byte) followed by bit-serial
    // combine steps was the longest serial stretch of such a block.  The slices' CRCs are then combined by ONE warp --
    // a lane folds kThreads / 32 neighbouring slices (Horner, multiplier of one slice), a shuffle tree the 32 results:
    // a combine step is ~100 instructions whoever needs its result, so eight warps climbing the tree side by side cost
    // four times the instructions for the same depth
    const int n_crc = 4 + data_len;
    {
        const int end = n_crc - (kThreads - 1 - t) * job.crc_slice, begin = max(end - job.crc_slice, 0);
        uint32_t c = 0u;
        if (end > begin) {
            c = 0xffffffffu;
            for (int i = begin; i < end; ++i) c = s_crc[(c ^ out[4 + i]) & 255u] ^ (c >> 8);
            c ^= 0xffffffffu;
        }
        s_coff[t] = c;  // (the candidates' offsets are done with)
    }
    __syncthreads();
    if (t < 32) {
        constexpr int kFold = kThreads / 32;  // 4, 8, 16 or 32 slices per lane: levels log2(kFold) .. log2(kFold) + 4 of the tree follow
        constexpr int kFirstLevel = kFold == 4 ? 2 : kFold == 8 ? 3 : kFold == 16 ? 4 : 5;
        uint32_t c = s_coff[kFold * t];
#pragma unroll
        for (int j = 1; j < kFold; ++j) c = crc_mul_fixed(job.crc_mul[0], c) ^ s_coff[kFold * t + j];
#pragma unroll
        for (int l = 0; l < 5; ++l) {
            const uint32_t right = __shfl_down_sync(0xffffffffu, c, 1 << l);
            const uint32_t left = crc_mul_fixed(job.crc_mul[kFirstLevel + l], c);
            if ((t & ((2 << l) - 1)) == 0) c = left ^ right;
        }
        if (t == 0) s_wcrc[0] = c;
    }
    __syncthreads();
    if (t == 0) {
        const uint32_t crc = s_wcrc[0];
        uint8_t* p = out + 8 + data_len;
        p[0] = (uint8_t)(crc >> 24), p[1] = (uint8_t)(crc >> 16), p[2] = (uint8_t)(crc >> 8), p[3] = (uint8_t)crc;
        job.sizes[band] = (uint32_t)(kChunkOverhead + data_len);
        job.adlers[band] = (adler_b << 16) | adler_a;
    }
    __syncthreads();
    if (!FUSED) {
        // ---- the finished chunk to its slot (whole words; the slot is word-aligned and has the slack)
        uint32_t* slot = reinterpret_cast<uint32_t*>(job.slots + (long long)band * job.slot_bytes);
        const int chunk_words = (kChunkOverhead + data_len + 3) / 4;
        for (int i = t; i < chunk_words; i += kThreads) slot[i] = out_words[i];
        return;
    }
    // ---- the finished chunk to its place in the file (device memory or mapped host memory), the head with band 0.
    // Its place: behind the chunks of all bands before it, whose sizes were published when their bits were counted --
    // half a block ago for bands that started with this one -- so the whole block just adds them up, waiting for the
    // few that are not there yet (they belong to blocks that took their ticket earlier and are therefore running).
    unsigned long long before = 0ull;
    for (int j = t; j < band; j += kThreads) {
        unsigned long long w;
        while (((w = *reinterpret_cast<volatile unsigned long long*>(ctl.state + j)) >> 32) == 0ull) __nanosleep(64);
        before += w & 0xffffffffull;
    }
    before = block_sum64(before, s_sum);
    // a chunk that would not leave room for the tail is not written (the last block then reports 0 bytes)
    const int size = kChunkOverhead + data_len;
    if (33ull + before + (unsigned long long)size + 33ull <= (unsigned long long)file.capacity) {
        copy_chunk(file.out + 33 + before, out, size, t);
        if (band == 0)
            for (int i = t; i < 33; i += kThreads) file.out[i] = file.head[i];
    }
    __syncthreads();
    if (t == 0) {
        __threadfence_system();  // (the file may be host memory: the block's bytes, and its sums above, before the count)
        s_last = atomicAdd(ctl.done, 1u) == (unsigned int)(file.n_bands - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (s_last == 0u) return;
    __threadfence();
    unsigned long long all = 0ull;
    for (int j = t; j < file.n_bands; j += kThreads) all += *reinterpret_cast<volatile unsigned long long*>(ctl.state + j) & 0xffffffffull;
    all = block_sum64(all, s_sum);
    write_tail(file, all, s_sum);
}

__global__ void __launch_bounds__(kThreads) png_pack_kernel(const __grid_constant__ PackJob job) {
    __shared__ unsigned long long s_sum[kThreads / 32];
    const int t = threadIdx.x, b = blockIdx.x;
    // where this band's chunk goes: after the head and the chunks of the bands before it
    // (both sums in one reduction: a PNG is far below 4 GB, so each fits 32 bits of the 64-bit word)
    unsigned long long both = 0;
    for (int j = t; j < job.n_bands; j += kThreads) {
        const unsigned long long s = job.sizes[j];
        both += (s << 32) | (j < b ? s : 0ull);
    }
    both = block_sum64(both, s_sum);
    const unsigned long long before = both & 0xffffffffull, all = both >> 32;
    const bool fits = 33ull + all + 21ull + 12ull <= (unsigned long long)job.capacity;
    if (fits) copy_chunk(job.out + 33 + before, job.slots + (long long)b * job.slot_bytes, (int)job.sizes[b], t);
    if (b == 0 && fits)
        for (int i = t; i < 33; i += kThreads) job.out[i] = job.head[i];
    if (b != job.n_bands - 1) return;
    write_tail(job, all, s_sum);
}

uint32_t host_crc32(const uint8_t* p, size_t n) {
    uint32_t c = 0xffffffffu;
    for (size_t i = 0; i < n; ++i) {
        c ^= p[i];
        for (int k = 0; k < 8; ++k) c = c & 1u ? (c >> 1) ^ kCrcPoly : c >> 1;
    }
    return c ^ 0xffffffffu;
}

int png_fail(int status, const char* message) {
    rt_internal_set_error(message);
    return status;
}

// Rows per band: about two bands per SM (one wave of blocks), a divisor of 16 (the strip height of the multi-GPU
// partition: a band never straddles two strips), and small enough for the band to fit shared memory twice over
// and a stored block (65 535 bytes).
int rows_per_band(int width, int height) {
    const long long line = 1 + 3ll * width;
    int r = 16;
    while (r > 1 && ((long long)height + r - 1) / r < RT_PNG_BANDS_PER_SM * 148) r >>= 1;
    if (((long long)height + r - 1) / r < 148 && r > 1) r >>= 1;
    while (r > 1 && (r * line > 40000)) r >>= 1;
    return r;
}
size_t band_smem(int width, int rows) {
    const size_t cap = (size_t)rows * (1 + 3 * (size_t)width);
    // the filtered rows, the chunk (+ 16: a word past its end is read when it is copied out), the run bits (+ an end word) and their next[] (16 bits per word)
    return ((cap + 15) & ~(size_t)15) + ((cap + kBandSlack + kChunkOverhead + 3) / 4) * 4 + 16 + ((cap + 31) / 32 + 2) * 6 + 4;
}
long long slot_bytes_of(int width, int rows) {
    const long long cap = (long long)rows * (1 + 3ll * width);
    return (cap + kBandSlack + kChunkOverhead + 15) & ~15ll;
}

// what the band kernel needs to know about the image and the scratch area (everything but the pixels)
void fill_band_job(BandJob& job, int width, int height, int r, void* scratch_dev) {
    const long long bands = (height + r - 1) / r;
    job.width = width, job.height = height;
    job.rows_per_band = r;
    job.slots = static_cast<uint8_t*>(scratch_dev);
    job.slot_bytes = slot_bytes_of(width, r);
    job.sizes = reinterpret_cast<uint32_t*>(job.slots + bands * job.slot_bytes);
    job.adlers = job.sizes + bands;
    // the chunk CRC's per-launch multipliers (see BandJob); kept for the next call of the same band size
    static thread_local int made_for = 0;
    static thread_local uint32_t made[10][32];
    const int slice = (int)((r * (1 + 3ll * width) + kBandSlack + kChunkOverhead + kThreads - 1) / kThreads);
    if (made_for != slice) {
        for (int l = 0; l < 10; ++l) {
            uint32_t g = x2nmodp(png_code().x2n, (uint64_t)slice, 3u + (unsigned)l);  // x^(8 slice 2^l)
            for (int k = 0; k < 32; ++k) {
                made[l][k] = g;
                g = g & 1u ? (g >> 1) ^ kCrcPoly : g >> 1;  // times x
            }
        }
        made_for = slice;
    }
    job.crc_slice = slice;
    memcpy(job.crc_mul, made, sizeof(made));
}

// ... and the file: header bytes, where the bands' sums are, where the file and its size go
void fill_pack_job(PackJob& job, int width, int height, int r, const void* scratch_dev, uint8_t* png_out, long long capacity,
                   uint64_t* png_bytes_out) {
    const long long bands = (height + r - 1) / r;
    memset(&job, 0, sizeof(job));
    job.slots = static_cast<const uint8_t*>(scratch_dev);
    job.slot_bytes = slot_bytes_of(width, r);
    job.sizes = reinterpret_cast<const uint32_t*>(job.slots + bands * job.slot_bytes);
    job.adlers = job.sizes + bands;
    job.n_bands = (int)bands;
    job.band_bytes = r * (1 + 3 * width);
    job.raw_bytes = (long long)height * (1 + 3ll * width);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    memcpy(job.head, sig, 8);
    uint8_t* ih = job.head + 8;
    ih[0] = 0, ih[1] = 0, ih[2] = 0, ih[3] = 13;
    memcpy(ih + 4, "IHDR", 4);
    for (int i = 0; i < 4; ++i) ih[8 + i] = (uint8_t)((uint32_t)width >> (24 - 8 * i)), ih[12 + i] = (uint8_t)((uint32_t)height >> (24 - 8 * i));
    ih[16] = 8, ih[17] = 2, ih[18] = 0, ih[19] = 0, ih[20] = 0;  // 8 bits, RGB, deflate, adaptive filters, not interlaced
    const uint32_t crc = host_crc32(ih + 4, 17);
    for (int i = 0; i < 4; ++i) ih[21 + i] = (uint8_t)(crc >> (24 - 8 * i));
    job.out = png_out;
    job.capacity = capacity;
    job.bytes_out = reinterpret_cast<unsigned long long*>(png_bytes_out);
}

// the band kernel's dynamic shared memory beyond the default 48 KB (static shared memory counts towards it), once per
// device, thread and instantiation
template <bool FUSED>
cudaError_t opt_in_smem(int device, size_t smem) {
    static thread_local size_t opted[64] = {0};
    if (smem + 8 * 1024 <= 48 * 1024 || device < 0 || device >= 64 || opted[device] >= smem) return cudaSuccess;
    const cudaError_t err = cudaFuncSetAttribute(png_band_kernel<FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err == cudaSuccess) opted[device] = smem;
    return err;
}

}  // namespace

extern "C" {

int32_t rt_png_rows_per_band(int32_t width, int32_t height) {
    if (width <= 0 || height <= 0 || 1 + 3ll * width > 65000) return -1;
    return rows_per_band(width, height);
}

int64_t rt_png_scratch_bytes(int32_t width, int32_t height) {
    if (rt_png_rows_per_band(width, height) < 0) return -1;
    const int r = rows_per_band(width, height);
    const long long bands = (height + r - 1) / r;
    // slots, sizes, Adler sums; ticket, finished-count and one status word per band of the single-launch encoder
    return bands * slot_bytes_of(width, r) + bands * 8 + 16 + bands * 8 + 64;
}

int64_t rt_png_file_bound(int32_t width, int32_t height) {
    if (rt_png_rows_per_band(width, height) < 0) return -1;
    const int r = rows_per_band(width, height);
    const long long bands = (height + r - 1) / r;
    return 33 + bands * (slot_bytes_of(width, r)) + 33;
}

int rt_png_encode_bands_dev(int device, const uint8_t* pixels_dev, int32_t pixel_bytes, int64_t row_stride, int32_t n_rows,
                            int32_t width, int32_t height, int32_t row0, int32_t strip_rows, int32_t strip_step,
                            void* scratch_dev, int64_t scratch_bytes, void* stream) {
    if (pixels_dev == nullptr || scratch_dev == nullptr || (pixel_bytes != 3 && pixel_bytes != 4) || width <= 0 || height <= 0 ||
        n_rows < 0 || row_stride < (int64_t)width * pixel_bytes || row0 < 0 || strip_rows <= 0 || strip_step <= 0)
        return png_fail(RT_ERR_BAD_ARG, "rt_png_encode_bands_dev: NULL pointer or bad size");
    if (rt_png_rows_per_band(width, height) < 0)
        return png_fail(RT_ERR_UNSUPPORTED, "rt_png_encode_bands_dev: rows wider than 21 666 pixels do not fit a stored block");
    if (scratch_bytes < rt_png_scratch_bytes(width, height)) return png_fail(RT_ERR_BAD_ARG, "rt_png_encode_bands_dev: scratch smaller than rt_png_scratch_bytes");
    const int r = rows_per_band(width, height);
    if (strip_step > 1 && (strip_rows % r != 0 || row0 % r != 0))
        return png_fail(RT_ERR_BAD_ARG, "rt_png_encode_bands_dev: strips must be whole bands (rt_png_rows_per_band)");
    if (strip_step == 1 && row0 % r != 0) return png_fail(RT_ERR_BAD_ARG, "rt_png_encode_bands_dev: row0 must be a multiple of rt_png_rows_per_band");
    if (n_rows == 0) return RT_OK;
    int previous = -1;
    cudaGetDevice(&previous);
    if (previous != device && cudaSetDevice(device) != cudaSuccess) return png_fail(RT_ERR_NO_DEVICE, "rt_png_encode_bands_dev: cudaSetDevice failed");
    BandJob job;
    job.pixels = pixels_dev;
    job.row_stride = row_stride;
    job.pixel_bytes = pixel_bytes;
    job.n_rows = n_rows;
    job.row0 = row0, job.strip_rows = strip_step == 1 ? (n_rows > 0 ? n_rows : 1) : strip_rows, job.strip_step = strip_step;
    fill_band_job(job, width, height, r, scratch_dev);
    const size_t smem = band_smem(width, r);
    cudaError_t err = opt_in_smem<false>(device, smem);
    if (err == cudaSuccess) {
        png_band_kernel<false><<<(unsigned)((n_rows + r - 1) / r), kThreads, smem, static_cast<cudaStream_t>(stream)>>>(png_code(), job, PackJob(), FusedCtl());
        err = cudaGetLastError();
    }
    if (previous >= 0 && previous != device) cudaSetDevice(previous);
    if (err != cudaSuccess) return png_fail(RT_ERR_CUDA, cudaGetErrorString(err));
    return RT_OK;
}

int rt_png_pack_dev(int device, int32_t width, int32_t height, const void* scratch_dev, uint8_t* png_out, int64_t capacity,
                    uint64_t* png_bytes_out, void* stream) {
    if (scratch_dev == nullptr || png_out == nullptr || png_bytes_out == nullptr || rt_png_rows_per_band(width, height) < 0 || capacity < 128)
        return png_fail(RT_ERR_BAD_ARG, "rt_png_pack_dev: NULL pointer or bad size");
    const int r = rows_per_band(width, height);
    const long long bands = (height + r - 1) / r;
    PackJob job;
    fill_pack_job(job, width, height, r, scratch_dev, png_out, capacity, png_bytes_out);
    int previous = -1;
    cudaGetDevice(&previous);
    if (previous != device && cudaSetDevice(device) != cudaSuccess) return png_fail(RT_ERR_NO_DEVICE, "rt_png_pack_dev: cudaSetDevice failed");
    png_pack_kernel<<<(unsigned)bands, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(job);
    const cudaError_t err = cudaGetLastError();
    if (previous >= 0 && previous != device) cudaSetDevice(previous);
    if (err != cudaSuccess) return png_fail(RT_ERR_CUDA, cudaGetErrorString(err));
    return RT_OK;
}

int rt_png_encode_file_dev(int device, const uint8_t* pixels_dev, int32_t pixel_bytes, int64_t row_stride, int32_t width, int32_t height,
                           void* scratch_dev, int64_t scratch_bytes, uint8_t* png_out, int64_t capacity, uint64_t* png_bytes_out,
                           void* stream) {
    if (pixels_dev == nullptr || scratch_dev == nullptr || png_out == nullptr || png_bytes_out == nullptr || (pixel_bytes != 3 && pixel_bytes != 4) ||
        width <= 0 || height <= 0 || row_stride < (int64_t)width * pixel_bytes || capacity < 128)
        return png_fail(RT_ERR_BAD_ARG, "rt_png_encode_file_dev: NULL pointer or bad size");
    if (rt_png_rows_per_band(width, height) < 0)
        return png_fail(RT_ERR_UNSUPPORTED, "rt_png_encode_file_dev: rows wider than 21 666 pixels do not fit a stored block");
    if (scratch_bytes < rt_png_scratch_bytes(width, height)) return png_fail(RT_ERR_BAD_ARG, "rt_png_encode_file_dev: scratch smaller than rt_png_scratch_bytes");
    const int r = rows_per_band(width, height);
    const long long bands = (height + r - 1) / r;
    BandJob job;
    job.pixels = pixels_dev;
    job.row_stride = row_stride;
    job.pixel_bytes = pixel_bytes;
    job.n_rows = height;
    job.row0 = 0, job.strip_rows = height, job.strip_step = 1;
    fill_band_job(job, width, height, r, scratch_dev);
    PackJob file;
    fill_pack_job(file, width, height, r, scratch_dev, png_out, capacity, png_bytes_out);
    FusedCtl ctl;
    unsigned char* control = reinterpret_cast<unsigned char*>(job.adlers + bands);  // (8-byte aligned: slots are multiples of 16, 8 bytes per band before it)
    ctl.ticket = reinterpret_cast<unsigned int*>(control);
    ctl.done = ctl.ticket + 1;
    ctl.state = reinterpret_cast<unsigned long long*>(control + 8);
    int previous = -1;
    cudaGetDevice(&previous);
    if (previous != device && cudaSetDevice(device) != cudaSuccess) return png_fail(RT_ERR_NO_DEVICE, "rt_png_encode_file_dev: cudaSetDevice failed");
    const size_t smem = band_smem(width, r);
    cudaError_t err = opt_in_smem<true>(device, smem);
    if (err == cudaSuccess) err = cudaMemsetAsync(control, 0, 8 + 8 * (size_t)bands, static_cast<cudaStream_t>(stream));
    if (err == cudaSuccess) {
        png_band_kernel<true><<<(unsigned)bands, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(png_code(), job, file, ctl);
        err = cudaGetLastError();
    }
    if (previous >= 0 && previous != device) cudaSetDevice(previous);
    if (err != cudaSuccess) return png_fail(RT_ERR_CUDA, cudaGetErrorString(err));
    return RT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// rt_png_encode_static_dev: the two calls above for a whole image, with the band scratch in a file-scope device
// array instead of memory the caller allocated -- for callers that cannot call cudaMalloc (graph capture, serving
// loops, harnesses that forbid it).  Every GPU that loads the library holds the array until the process exits.
// In HOST code `g_png_static_scratch` is only a stand-in for the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_png_static_scratch[RT_PNG_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticDevices = 64;
void* g_static_scratch_ptr[kStaticDevices] = {nullptr};  // device address of the array on device d, once asked for
}  // namespace

extern "C" int rt_png_encode_static_dev(int device, const uint8_t* pixels_dev, int32_t pixel_bytes, int64_t row_stride,
                                        int32_t width, int32_t height, uint8_t* png_out, int64_t capacity,
                                        uint64_t* png_bytes_out, void* stream) {
    if (pixels_dev == nullptr || png_out == nullptr || png_bytes_out == nullptr || (pixel_bytes != 3 && pixel_bytes != 4) ||
        width <= 0 || height <= 0 || row_stride < (int64_t)width * pixel_bytes || capacity < 128)
        return png_fail(RT_ERR_BAD_ARG, "rt_png_encode_static_dev: NULL pointer or bad size");
    if (device < 0 || device >= kStaticDevices) return png_fail(RT_ERR_NO_DEVICE, "rt_png_encode_static_dev: device index out of range");
    const int64_t need = rt_png_scratch_bytes(width, height);
    if (need < 0 || need > (int64_t)RT_PNG_STATIC_SCRATCH_BYTES) {
        char message[320];
        snprintf(message, sizeof(message),
                 "rt_png_encode_static_dev: a %d x %d image needs %lld bytes of band scratch, the static array holds %lld "
                 "(RT_PNG_STATIC_SCRATCH_BYTES); use rt_png_encode_bands_dev + rt_png_pack_dev with your own scratch",
                 (int)width, (int)height, (long long)need, (long long)RT_PNG_STATIC_SCRATCH_BYTES);
        return png_fail(RT_ERR_UNSUPPORTED, message);
    }
    void* scratch = __atomic_load_n(&g_static_scratch_ptr[device], __ATOMIC_ACQUIRE);
    if (scratch == nullptr) {  // set-up, once per device, with that device current (two threads racing here ask twice and get the same answer)
        int previous = -1;
        cudaGetDevice(&previous);
        if (previous != device && cudaSetDevice(device) != cudaSuccess) return png_fail(RT_ERR_NO_DEVICE, "rt_png_encode_static_dev: cudaSetDevice failed");
        const cudaError_t err = cudaGetSymbolAddress(&scratch, g_png_static_scratch);
        if (previous >= 0 && previous != device) cudaSetDevice(previous);
        if (err != cudaSuccess || scratch == nullptr) {
            char message[160];
            snprintf(message, sizeof(message), "rt_png_encode_static_dev: cudaGetSymbolAddress failed: %s", cudaGetErrorString(err));
            return png_fail(RT_ERR_CUDA, message);
        }
        __atomic_store_n(&g_static_scratch_ptr[device], scratch, __ATOMIC_RELEASE);
    }
    // Nothing in the array has to start at zero: the band kernel writes every byte the pack kernel reads (each
    // band's chunk, size and Adler sums, for every band of the image), so what the last call left is never read.
    const int status = rt_png_encode_bands_dev(device, pixels_dev, pixel_bytes, row_stride, height, width, height, 0, height, 1,
                                               scratch, (int64_t)RT_PNG_STATIC_SCRATCH_BYTES, stream);
    if (status != RT_OK) return status;
    return rt_png_pack_dev(device, width, height, scratch, png_out, capacity, png_bytes_out, stream);
}
