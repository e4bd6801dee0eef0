// the baseline JPEG encoder behind the reference's
// `Image.fromarray(combined_image).save(...jpg)` (Style_3DGS/localized_style_transfer.py:241-243), on the
// device, so that a transferred frame leaves the GPU as a ~10x smaller byte stream instead of H*W*3 bytes.
// Output = the bytes Pillow writes with its defaults (libjpeg baseline, quality 75, 4:2:0, Annex-K Huffman
// tables, islow DCT).  All integer work, bit-exact by construction (tests: Pillow itself and the numpy restatement).
//
// Pipeline per frame (blocks in scan order: MCU-major, Y00 Y01 Y10 Y11 Cb Cr inside a 16x16 MCU):
//   k_jpeg_dct      thread per block: RGB -> component samples (edge rules of jcprepct.c/jcsample.c),
//                   islow DCT, quantise -> int16[64] in zigzag order (3 B/px written), plus per block
//                   (bits of its AC codes, DC) for the bit count
//   k_jpeg_bits     thread per block: length in bits of the block's Huffman code; sum per group of 256 blocks
//   k_jpeg_scan     one CTA per frame: exclusive scan over the groups -> bit offset of each group, total bits
//   k_jpeg_emit     thread per block: offset = group offset + scan inside the CTA; codes OR-ed into the
//                   (zeroed) unstuffed bit stream
//   k_jpeg_ff_count / k_jpeg_ff_scan / k_jpeg_ff_write   byte stuffing (0xFF -> 0xFF 0x00) as a
//                   count / scan / scatter over 4 KiB tiles, header in front, EOI behind, byte count out
// The per-block bodies are __host__ __device__ so that tools/jpeg_selftest.cu can run the same arithmetic
// serially on the host (tests/test_jpeg_device_core.py pins it to Pillow without a GPU); the emulation lives there.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>

#ifdef __CUDACC__
#define JHD __host__ __device__ __forceinline__
#else
#define JHD inline
#endif

namespace jpeg {

struct Geom {
    int H, W;
    int mx, my;        // MCUs across / down
    int bx, by;        // real luma blocks across / down; luma blocks beyond are libjpeg's dummy blocks
    int ch;            // rows of a downsampled chroma plane holding image data = ceil(H/2)
    unsigned nmcu, nblk;
    size_t frame_stride;   // bytes between frames of the input
};

inline Geom make_geom(int H, int W, size_t frame_stride)
{
    Geom g;
    g.H = H, g.W = W;
    g.mx = (W + 15) / 16, g.my = (H + 15) / 16;
    g.bx = (W + 7) / 8, g.by = (H + 7) / 8;
    g.ch = (H + 1) / 2;
    g.nmcu = (unsigned)g.mx * (unsigned)g.my;
    g.nblk = g.nmcu * 6u;
    g.frame_stride = frame_stride ? frame_stride : (size_t)H * (size_t)W * 3u;
    return g;
}

// derived tables, passed to the kernels by value (2.4 KB of kernel parameters)
struct Tables {
    unsigned short q8[2][64];   // 8 * quantiser, indexed by ZIGZAG POSITION (the order the blocks are stored in)
    unsigned rq[2][64];         // ceil(2^32 / q8): floor(n / q8) == mulhi(n, rq) for every n < 2^20 (q8 <= 2040)
    unsigned dc[2][12];         // (code << 5) | length
    unsigned ac[2][256];
};

// natural index of zigzag position k (a constexpr local: folded away where k is a compile-time constant)
JHD int zigzag_natural(int k)
{
    constexpr unsigned char t[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                     41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                     30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    return t[k];
}

// ---- sample preparation ---------------------------------------------------------------------
// jccolor.c rgb_ycc_convert in 16-bit fixed point
JHD int ycc_y(int r, int g, int b) { return (19595 * r + 38470 * g + 7471 * b + 32768) >> 16; }
JHD int ycc_cb(int r, int g, int b) { return (-11059 * r - 21709 * g + 32768 * b + (128 << 16) + 32767) >> 16; }
JHD int ycc_cr(int r, int g, int b) { return (32768 * r - 27439 * g - 5329 * b + (128 << 16) + 32767) >> 16; }

JHD bool is_dummy(const Geom &g, unsigned mcu, int b)
{
    if (b >= 4) return false;
    const int mxi = (int)(mcu % (unsigned)g.mx), myi = (int)(mcu / (unsigned)g.mx);
    return (2 * myi + (b >> 1)) >= g.by || (2 * mxi + (b & 1)) >= g.bx;
}

// NPX pixels of one image row starting at x0, as packed little-endian words; pixels right of the image repeat the
// last column.  Spans inside the image whose first byte is 4-byte aligned are read as words.
template <int NPX>
JHD void load_span(const uint8_t *row, int x0, int W, unsigned *wv)
{
    const uint8_t *p = row + (size_t)x0 * 3u;
    if (x0 + NPX <= W && ((uintptr_t)p & 3u) == 0) {
        const unsigned *pw = reinterpret_cast<const unsigned *>(p);
#pragma unroll
        for (int k = 0; k < NPX * 3 / 4; ++k) wv[k] = pw[k];
        return;
    }
#pragma unroll
    for (int k = 0; k < NPX * 3 / 4; ++k) wv[k] = 0u;
#pragma unroll
    for (int j = 0; j < NPX * 3; ++j) {
        const int x = x0 + j / 3 < W ? x0 + j / 3 : W - 1;
        wv[j >> 2] |= (unsigned)row[(size_t)x * 3u + (unsigned)(j % 3)] << (8 * (j & 3));
    }
}
JHD int span_byte(const unsigned *wv, int j) { return (int)((wv[j >> 2] >> (8 * (j & 3))) & 255u); }

// level-shifted samples of block b of one MCU.  Luma: edge pixels replicated.  Chroma: the right edge is
// replicated at full resolution before the 2x2 box filter (bias 1, 2, 1, 2 ... along a row), an odd last row
// is doubled, and rows below the image repeat the last DOWNSAMPLED row (jcprepct.c pre_process_data).
JHD void block_samples(const uint8_t *rgb, const Geom &g, unsigned mcu, int b, int *blk)
{
    const int mxi = (int)(mcu % (unsigned)g.mx), myi = (int)(mcu / (unsigned)g.mx);
    const size_t rowb = (size_t)g.W * 3u;
    if (b < 4) {
        const int y0 = myi * 16 + (b >> 1) * 8, x0 = mxi * 16 + (b & 1) * 8;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int y = y0 + r < g.H ? y0 + r : g.H - 1;
            unsigned wv[6];
            load_span<8>(rgb + (size_t)y * rowb, x0 < g.W ? x0 : g.W - 1, g.W, wv);
#pragma unroll
            for (int c = 0; c < 8; ++c)
                blk[r * 8 + c] = ycc_y(span_byte(wv, 3 * c), span_byte(wv, 3 * c + 1), span_byte(wv, 3 * c + 2)) - 128;
        }
        return;
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int cr = myi * 8 + r;
        if (cr > g.ch - 1) cr = g.ch - 1;
        const int ya = 2 * cr, yb = 2 * cr + 1 < g.H ? 2 * cr + 1 : g.H - 1;
        unsigned wa[12], wb[12];
        load_span<16>(rgb + (size_t)ya * rowb, mxi * 16, g.W, wa);
        load_span<16>(rgb + (size_t)yb * rowb, mxi * 16, g.W, wb);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            int s = 0;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int j = 3 * (2 * c + k);
                const int r0 = span_byte(wa, j), g0 = span_byte(wa, j + 1), b0 = span_byte(wa, j + 2);
                const int r1 = span_byte(wb, j), g1 = span_byte(wb, j + 1), b1 = span_byte(wb, j + 2);
                s += b == 4 ? ycc_cb(r0, g0, b0) + ycc_cb(r1, g1, b1) : ycc_cr(r0, g0, b0) + ycc_cr(r1, g1, b1);
            }
            blk[r * 8 + c] = ((s + 1 + (c & 1)) >> 2) - 128;
        }
    }
}

// ---- DCT + quantiser ------------------------------------------------------------------------
JHD int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// one 8-point pass of jfdctint.c jpeg_fdct_islow (CONST_BITS 13, PASS1_BITS 2) over d[0], d[s], ..., d[7s]
template <bool FIRST>
JHD void fdct8(int *d, int s)
{
    const int d0 = d[0], d1 = d[s], d2 = d[2 * s], d3 = d[3 * s], d4 = d[4 * s], d5 = d[5 * s], d6 = d[6 * s], d7 = d[7 * s];
    int t0 = d0 + d7, t7 = d0 - d7, t1 = d1 + d6, t6 = d1 - d6, t2 = d2 + d5, t5 = d2 - d5, t3 = d3 + d4, t4 = d3 - d4;
    const int t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
    constexpr int sh = FIRST ? 13 - 2 : 13 + 2;
    if (FIRST) {
        d[0] = (t10 + t11) * 4;
        d[4 * s] = (t10 - t11) * 4;
    } else {
        d[0] = descale(t10 + t11, 2);
        d[4 * s] = descale(t10 - t11, 2);
    }
    int z1 = (t12 + t13) * 4433;
    d[2 * s] = descale(z1 + t13 * 6270, sh);
    d[6 * s] = descale(z1 - t12 * 15137, sh);
    z1 = t4 + t7;
    int z2 = t5 + t6, z3 = t4 + t6, z4 = t5 + t7;
    const int z5 = (z3 + z4) * 9633;
    t4 *= 2446, t5 *= 16819, t6 *= 25172, t7 *= 12299;
    z1 *= -7373, z2 *= -20995;
    z3 = z3 * -16069 + z5;
    z4 = z4 * -3196 + z5;
    d[7 * s] = descale(t4 + z1 + z3, sh);
    d[5 * s] = descale(t5 + z2 + z4, sh);
    d[3 * s] = descale(t6 + z2 + z3, sh);
    d[s] = descale(t7 + z1 + z4, sh);
}

JHD unsigned mulhi32(unsigned a, unsigned b)
{
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return (unsigned)(((unsigned long long)a * (unsigned long long)b) >> 32);
#endif
}

// jcdctmgr.c forward_DCT: rows, columns, then (|c| + q8/2) / q8 with the sign put back; zigzag order out.
// The division is a multiplication by ceil(2^32 / q8): exact while n * q8 < 2^32, and |c| < 2^16 here.
JHD void fdct_quantise(int *blk, const unsigned short *q8zz, const unsigned *rqzz, short *out)
{
#pragma unroll
    for (int r = 0; r < 8; ++r) fdct8<true>(blk + 8 * r, 1);
#pragma unroll
    for (int c = 0; c < 8; ++c) fdct8<false>(blk + c, 8);
#pragma unroll
    for (int k = 0; k < 64; ++k) {
        const int v = blk[zigzag_natural(k)], q = (int)q8zz[k];
        const int m = (int)mulhi32((unsigned)((v < 0 ? -v : v) + (q >> 1)), rqzz[k]);
        out[k] = (short)(v < 0 ? -m : m);
    }
}

// ---- entropy coder --------------------------------------------------------------------------
JHD int bit_length(unsigned v)
{
#ifdef __CUDA_ARCH__
    return 32 - __clz((int)v);
#else
    int n = 0;
    while (v) ++n, v >>= 1;
    return n;
#endif
}

// DC of block b of an MCU as the entropy coder sees it: a dummy block repeats the DC of the block before it
JHD int resolved_dc(const short *coef, const Geom &g, unsigned mcu, int b)
{
    while (b > 0 && is_dummy(g, mcu, b)) --b;
    return coef[((size_t)mcu * 6u + (unsigned)b) * 64u];
}
// the predictor: DC of the previous block of the same component in scan order
JHD int dc_predictor(const short *coef, const Geom &g, unsigned mcu, int b)
{
    if (b >= 1 && b <= 3) return resolved_dc(coef, g, mcu, b - 1);
    if (mcu == 0) return 0;
    return b == 0 ? resolved_dc(coef, g, mcu - 1, 3) : (int)coef[((size_t)(mcu - 1) * 6u + (unsigned)b) * 64u];
}

// jchuff.c encode_one_block; SINK::put(code, length).  zz = the block's 64 coefficients in zigzag order
// (zz[0] is ignored: the DC difference is passed in), dummy = AC all zero.
template <class SINK, bool UNROLL = true>
JHD void encode_block(SINK &sink, const short *zz, bool dummy, int dc_diff, const unsigned *dc_tab, const unsigned *ac_tab)
{
    {
        const int mag = dc_diff < 0 ? -dc_diff : dc_diff, low = dc_diff < 0 ? dc_diff - 1 : dc_diff;
        const int nb = bit_length((unsigned)mag);
        const unsigned e = dc_tab[nb];
        sink.put(e >> 5, (int)(e & 31u));
        if (nb) sink.put((unsigned)low & ((1u << nb) - 1u), nb);
    }
    int run = 0;
    if (!dummy) {
#pragma unroll(UNROLL ? 63 : 1)
        for (int k = 1; k < 64; ++k) {
            const int v = zz[k];
            if (v == 0) {
                ++run;
                continue;
            }
            while (run > 15) {
                const unsigned z = ac_tab[0xF0];
                sink.put(z >> 5, (int)(z & 31u));
                run -= 16;
            }
            const int mag = v < 0 ? -v : v, low = v < 0 ? v - 1 : v;
            const int nb = bit_length((unsigned)mag);
            const unsigned e = ac_tab[(run << 4) + nb];
            sink.put(e >> 5, (int)(e & 31u));
            sink.put((unsigned)low & ((1u << nb) - 1u), nb);
            run = 0;
        }
    } else {
        run = 63;
    }
    if (run) {
        const unsigned e = ac_tab[0];
        sink.put(e >> 5, (int)(e & 31u));
    }
}

struct CountSink {
    unsigned bits = 0;
    JHD void put(unsigned, int len) { bits += (unsigned)len; }
};

// what k_jpeg_dct leaves per block for the bit count: (bits of the AC codes << 16) | (DC & 0xffff)
JHD unsigned ac_code_bits(const short *zz, bool dummy, const unsigned *dc_tab, const unsigned *ac_tab)
{
    CountSink sink;
    encode_block(sink, zz, dummy, 0, dc_tab, ac_tab);
    return sink.bits - (dc_tab[0] & 31u);
}
JHD unsigned dc_code_bits(int diff, const unsigned *dc_tab)
{
    const int nb = bit_length((unsigned)(diff < 0 ? -diff : diff));
    return (dc_tab[nb] & 31u) + (unsigned)nb;
}
JHD int meta_dc(unsigned m) { return (int)(short)(unsigned short)(m & 0xffffu); }
JHD int resolved_dc_meta(const unsigned *meta, const Geom &g, unsigned mcu, int b)
{
    while (b > 0 && is_dummy(g, mcu, b)) --b;
    return meta_dc(meta[(size_t)mcu * 6u + (unsigned)b]);
}
JHD int dc_predictor_meta(const unsigned *meta, const Geom &g, unsigned mcu, int b)
{
    if (b >= 1 && b <= 3) return resolved_dc_meta(meta, g, mcu, b - 1);
    if (mcu == 0) return 0;
    return b == 0 ? resolved_dc_meta(meta, g, mcu - 1, 3) : meta_dc(meta[(size_t)(mcu - 1) * 6u + (unsigned)b]);
}

JHD unsigned bswap32(unsigned v)
{
#ifdef __CUDA_ARCH__
    return __byte_perm(v, 0, 0x0123);
#else
    return (v >> 24) | ((v >> 8) & 0xff00u) | ((v << 8) & 0xff0000u) | (v << 24);
#endif
}

// writes a run of bits into a zero-initialised big-endian bit stream, starting at any bit offset; words shared
// with the neighbouring blocks are OR-ed atomically (every word is, it keeps the sink small)
struct StreamSink {
    unsigned *words;
    unsigned long long acc;
    unsigned long long w;
    int n;
    JHD StreamSink(unsigned *stream, unsigned long long bit_offset)
        : words(stream), acc(0ull), w(bit_offset >> 5), n((int)(bit_offset & 31ull)) {}
    JHD void or_word(unsigned v)
    {
        if (v == 0u) return;
#ifdef __CUDA_ARCH__
        atomicOr(words + w, bswap32(v));
#else
        words[w] |= bswap32(v);
#endif
    }
    JHD void put(unsigned code, int len)
    {
        acc = (acc << len) | (unsigned long long)code;
        n += len;
        if (n >= 32) {
            or_word((unsigned)(acc >> (n - 32)));
            ++w;
            n -= 32;
            acc &= (1ull << n) - 1ull;
        }
    }
    JHD unsigned long long position() const { return w * 32ull + (unsigned long long)n; }
    JHD void finish()
    {
        if (n > 0) or_word((unsigned)(acc << (32 - n)));
    }
};

// ---- host side: tables, header, workspace layout --------------------------------------------
constexpr unsigned kTile = 4096;         // bytes of unstuffed stream per stuffing tile
#ifndef RCT_JENT
#define RCT_JENT 256
#endif
#ifndef RCT_JEMIT_UNROLL
#define RCT_JEMIT_UNROLL 1               // 0: k_jpeg_emit walks the 63 AC coefficients in a rolled loop (A/B knob)
#endif
constexpr int kEntThreads = RCT_JENT;    // blocks per CTA of the entropy kernels = one group of the offset scan
constexpr unsigned kBlockBitsMax = 1664; // >= 20 (DC) + 63 * 26 (AC) bits: the worst block
constexpr int kHeaderMax = 640;

struct Header {
    unsigned char b[kHeaderMax];
    int len;
};

struct Std {  // ITU T.81 Annex K
    static const unsigned char *luma_q50()
    {
        static const unsigned char t[64] = {16, 11, 12, 14, 12, 10, 16, 14, 13, 14, 18, 17, 16, 19, 24, 40,
                                            26, 24, 22, 22, 24, 49, 35, 37, 29, 40, 58, 51, 61, 60, 57, 51,
                                            56, 55, 64, 72, 92, 78, 64, 68, 87, 69, 55, 56, 80, 109, 81, 87,
                                            95, 98, 103, 104, 103, 62, 77, 113, 121, 112, 100, 120, 92, 101, 103, 99};
        return t;  // zigzag order
    }
    static const unsigned char *chroma_q50()
    {
        static unsigned char t[64];
        static const unsigned char head[16] = {17, 18, 18, 24, 21, 24, 47, 26, 26, 47, 99, 66, 56, 66, 99, 99};
        for (int i = 0; i < 64; ++i) t[i] = i < 16 ? head[i] : 99;
        return t;
    }
    static const unsigned char *bits(int tc_th)
    {
        static const unsigned char dc0[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
        static const unsigned char dc1[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
        static const unsigned char ac0[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 125};
        static const unsigned char ac1[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 119};
        return tc_th == 0x00 ? dc0 : tc_th == 0x01 ? dc1 : tc_th == 0x10 ? ac0 : ac1;
    }
    static const unsigned char *vals(int tc_th, int *n)
    {
        static const unsigned char dc[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
        static const unsigned char ac0[162] = {
            1,   2,   3,   0,   4,   17,  5,   18,  33,  49,  65,  6,   19,  81,  97,  7,   34,  113, 20,  50,  129, 145, 161,
            8,   35,  66,  177, 193, 21,  82,  209, 240, 36,  51,  98,  114, 130, 9,   10,  22,  23,  24,  25,  26,  37,  38,
            39,  40,  41,  42,  52,  53,  54,  55,  56,  57,  58,  67,  68,  69,  70,  71,  72,  73,  74,  83,  84,  85,  86,
            87,  88,  89,  90,  99,  100, 101, 102, 103, 104, 105, 106, 115, 116, 117, 118, 119, 120, 121, 122, 131, 132, 133,
            134, 135, 136, 137, 138, 146, 147, 148, 149, 150, 151, 152, 153, 154, 162, 163, 164, 165, 166, 167, 168, 169, 170,
            178, 179, 180, 181, 182, 183, 184, 185, 186, 194, 195, 196, 197, 198, 199, 200, 201, 202, 210, 211, 212, 213, 214,
            215, 216, 217, 218, 225, 226, 227, 228, 229, 230, 231, 232, 233, 234, 241, 242, 243, 244, 245, 246, 247, 248, 249,
            250};
        static const unsigned char ac1[162] = {
            0,   1,   2,   3,   17,  4,   5,   33,  49,  6,   18,  65,  81,  7,   97,  113, 19,  34,  50,  129, 8,   20,  66,
            145, 161, 177, 193, 9,   35,  51,  82,  240, 21,  98,  114, 209, 10,  22,  36,  52,  225, 37,  241, 23,  24,  25,
            26,  38,  39,  40,  41,  42,  53,  54,  55,  56,  57,  58,  67,  68,  69,  70,  71,  72,  73,  74,  83,  84,  85,
            86,  87,  88,  89,  90,  99,  100, 101, 102, 103, 104, 105, 106, 115, 116, 117, 118, 119, 120, 121, 122, 130, 131,
            132, 133, 134, 135, 136, 137, 138, 146, 147, 148, 149, 150, 151, 152, 153, 154, 162, 163, 164, 165, 166, 167, 168,
            169, 170, 178, 179, 180, 181, 182, 183, 184, 185, 186, 194, 195, 196, 197, 198, 199, 200, 201, 202, 210, 211, 212,
            213, 214, 215, 216, 217, 218, 226, 227, 228, 229, 230, 231, 232, 233, 234, 242, 243, 244, 245, 246, 247, 248, 249,
            250};
        if (tc_th < 0x10) {
            *n = 12;
            return dc;
        }
        *n = 162;
        return tc_th == 0x10 ? ac0 : ac1;
    }
};

// jcparam.c jpeg_quality_scaling + jpeg_add_quant_table(force_baseline); zigzag order
inline void quant_tables(int quality, unsigned char *luma, unsigned char *chroma)
{
    quality = quality < 1 ? 1 : quality > 100 ? 100 : quality;
    const int scale = quality < 50 ? 5000 / quality : 200 - 2 * quality;
    const unsigned char *base[2] = {Std::luma_q50(), Std::chroma_q50()};
    unsigned char *out[2] = {luma, chroma};
    for (int t = 0; t < 2; ++t)
        for (int i = 0; i < 64; ++i) {
            int v = ((int)base[t][i] * scale + 50) / 100;
            out[t][i] = (unsigned char)(v < 1 ? 1 : v > 255 ? 255 : v);
        }
}

inline void make_tables(int quality, Tables *t, Header *h, int H, int W)
{
    unsigned char q[2][64];
    quant_tables(quality, q[0], q[1]);
    for (int c = 0; c < 2; ++c)
        for (int i = 0; i < 64; ++i) {
            t->q8[c][i] = (unsigned short)(q[c][i] * 8);
            t->rq[c][i] = (unsigned)(((1ull << 32) + t->q8[c][i] - 1u) / t->q8[c][i]);
        }
    memset(t->dc, 0, sizeof t->dc);
    memset(t->ac, 0, sizeof t->ac);
    const int ids[4] = {0x00, 0x10, 0x01, 0x11};
    for (int id : ids) {  // jchuff.c jpeg_make_c_derived_tbl: canonical codes
        int n;
        const unsigned char *bits = Std::bits(id), *vals = Std::vals(id, &n);
        unsigned code = 0;
        int k = 0;
        for (int len = 1; len <= 16; ++len) {
            for (int i = 0; i < bits[len - 1]; ++i, ++k, ++code) {
                unsigned *dst = (id & 0x10) ? t->ac[id & 1] : t->dc[id & 1];
                dst[vals[k]] = (code << 5) | (unsigned)len;
            }
            code <<= 1;
        }
    }
    // the header Pillow's defaults produce: SOI, JFIF APP0 (density 1:1, no units), DQT x2, SOF0, DHT x4, SOS
    unsigned char *p = h->b;
    auto seg = [&](int marker, const unsigned char *payload, int n) {
        *p++ = 0xFF, *p++ = (unsigned char)marker;
        *p++ = (unsigned char)((n + 2) >> 8), *p++ = (unsigned char)((n + 2) & 255);
        memcpy(p, payload, (size_t)n);
        p += n;
    };
    *p++ = 0xFF, *p++ = 0xD8;
    const unsigned char app0[14] = {'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
    seg(0xE0, app0, 14);
    for (int c = 0; c < 2; ++c) {
        unsigned char dqt[65];
        dqt[0] = (unsigned char)c;
        memcpy(dqt + 1, q[c], 64);
        seg(0xDB, dqt, 65);
    }
    const unsigned char sof[15] = {8, (unsigned char)(H >> 8), (unsigned char)(H & 255), (unsigned char)(W >> 8),
                                   (unsigned char)(W & 255), 3, 1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1};
    seg(0xC0, sof, 15);
    for (int id : ids) {
        int n;
        const unsigned char *bits = Std::bits(id), *vals = Std::vals(id, &n);
        unsigned char dht[1 + 16 + 162];
        dht[0] = (unsigned char)id;
        memcpy(dht + 1, bits, 16);
        memcpy(dht + 17, vals, (size_t)n);
        seg(0xC4, dht, 17 + n);
    }
    const unsigned char sos[10] = {3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0};
    seg(0xDA, sos, 10);
    h->len = (int)(p - h->b);
}

// device scratch of one call (n frames).  Every array is 256-byte aligned.
struct Workspace {
    short *coef;                  // [n][nblk][64] zigzag order
    unsigned *meta;               // [n][nblk] (AC code bits << 16) | DC
    unsigned *bits;               // [n][nblk]
    unsigned *part;               // [n][nparts] bits of each group of kEntThreads consecutive blocks
    unsigned long long *pbase;    // [n][nparts] bit offset of each group in its frame's stream
    unsigned long long *total;    // [n] bits of entropy-coded data before padding
    unsigned *stream;             // [n][stream_words] unstuffed, zeroed by the caller of the kernels
    unsigned *tile_ff;            // [n][ntiles] 0xFF bytes per tile, then their exclusive scan
    size_t stream_words, ntiles, nparts, bytes;
};
inline size_t align256(size_t v) { return (v + 255u) & ~(size_t)255u; }
inline Workspace carve(const Geom &g, int n, void *base)
{
    Workspace w;
    const size_t stream_bytes = align256(((size_t)g.nblk * kBlockBitsMax + 7u) / 8u + 8u);
    w.stream_words = stream_bytes / 4u;
    w.ntiles = (stream_bytes + kTile - 1u) / kTile;
    char *p = (char *)base;
    size_t o = 0;
    auto take = [&](size_t bytes) {
        char *r = p ? p + o : nullptr;
        o += align256(bytes);
        return (void *)r;
    };
    w.coef = (short *)take((size_t)n * g.nblk * 64u * sizeof(short));
    w.meta = (unsigned *)take((size_t)n * g.nblk * sizeof(unsigned));
    w.bits = (unsigned *)take((size_t)n * g.nblk * sizeof(unsigned));
    w.nparts = ((size_t)g.nblk + kEntThreads - 1u) / kEntThreads;
    w.part = (unsigned *)take((size_t)n * w.nparts * sizeof(unsigned));
    w.pbase = (unsigned long long *)take((size_t)n * w.nparts * sizeof(unsigned long long));
    w.total = (unsigned long long *)take((size_t)n * sizeof(unsigned long long));
    w.stream = (unsigned *)take((size_t)n * stream_bytes);
    w.tile_ff = (unsigned *)take((size_t)n * w.ntiles * sizeof(unsigned));
    w.bytes = o;
    return w;
}

// ---- kernels --------------------------------------------------------------------------------
#ifdef __CUDACC__
template <class T>
__device__ __forceinline__ T jwarp_excl_scan(T v, T &total)
{
    const unsigned lane = threadIdx.x & 31u;
    T inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const T t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}
// exclusive scan of one value per thread across the CTA (blockDim.x a multiple of 32, <= 1024); s_warp[33]
template <class T>
__device__ __forceinline__ T jblock_excl_scan(T v, T *s_warp, T &block_total)
{
    T wtotal;
    const T ex = jwarp_excl_scan(v, wtotal);
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nwarps = blockDim.x >> 5;
    __syncthreads();  // s_warp may still be read from the previous call
    if (lane == 31) s_warp[warp] = wtotal;
    __syncthreads();
    if (warp == 0) {
        T t = lane < nwarps ? s_warp[lane] : (T)0, tt;
        const T e = jwarp_excl_scan(t, tt);
        if (lane < nwarps) s_warp[lane] = e;
        if (lane == 0) s_warp[32] = tt;
    }
    __syncthreads();
    block_total = s_warp[32];
    return ex + s_warp[warp];
}

constexpr int kDctThreads = 192, kFfThreads = kTile / 16;

__device__ __forceinline__ void load_tables(const Tables &tab, unsigned (*s_dc)[12], unsigned (*s_ac)[256])
{
    for (int k = threadIdx.x; k < 512; k += blockDim.x) s_ac[k >> 8][k & 255] = tab.ac[k >> 8][k & 255];
    for (int k = threadIdx.x; k < 24; k += blockDim.x) s_dc[k / 12][k % 12] = tab.dc[k / 12][k % 12];
    __syncthreads();
}
__device__ __forceinline__ void load_block(const short *coef, size_t block, short *zz)
{
    const uint4 *src = reinterpret_cast<const uint4 *>(coef + block * 64u);
    uint4 *dst = reinterpret_cast<uint4 *>(zz);
#pragma unroll
    for (int k = 0; k < 8; ++k) dst[k] = __ldg(src + k);
}

// 192-thread CTAs over 32 consecutive MCUs: warps 0-3 take the 128 luma blocks, warp 4 the Cb and warp 5 the Cr
// blocks (a chroma block reads four times the pixels of a luma block: warps of one kind finish together)
#ifndef RCT_JDB
#define RCT_JDB 3
#endif
__global__ void __launch_bounds__(kDctThreads, RCT_JDB) k_jpeg_dct(const uint8_t *rgb, Geom g, Tables tab, short *coef, unsigned *meta)
{
    __shared__ unsigned s_dc[2][12], s_ac[2][256];
    load_tables(tab, s_dc, s_ac);
    const unsigned frame = blockIdx.y, t = threadIdx.x;
    const unsigned mcu = blockIdx.x * 32u + (t < 128u ? t >> 2 : (t - 128u) & 31u);
    const int b = t < 128u ? (int)(t & 3u) : 4 + (int)((t - 128u) >> 5);
    if (mcu >= g.nmcu) return;
    __align__(16) short out[64];
    if (is_dummy(g, mcu, b)) {
#pragma unroll
        for (int k = 0; k < 64; ++k) out[k] = 0;
    } else {
        int blk[64];
        block_samples(rgb + (size_t)frame * g.frame_stride, g, mcu, b, blk);
        fdct_quantise(blk, tab.q8[b < 4 ? 0 : 1], tab.rq[b < 4 ? 0 : 1], out);
    }
    const size_t blk_index = (size_t)frame * g.nblk + (size_t)mcu * 6u + (unsigned)b;
    const int c = b < 4 ? 0 : 1;
    meta[blk_index] = (ac_code_bits(out, is_dummy(g, mcu, b), s_dc[c], s_ac[c]) << 16) | (unsigned)(unsigned short)out[0];
    uint4 *dst = reinterpret_cast<uint4 *>(coef + blk_index * 64u);
    const uint4 *src = reinterpret_cast<const uint4 *>(out);
#pragma unroll
    for (int k = 0; k < 8; ++k) dst[k] = src[k];
}

// grid (nparts, n): thread per block -> bits[] = AC bits + bits of the DC difference, and the sum over the CTA's
// blocks -> part[]
__global__ void __launch_bounds__(kEntThreads) k_jpeg_bits(const unsigned *meta, Geom g, Tables tab, unsigned *bits,
                                                           unsigned *part, size_t nparts)
{
    __shared__ unsigned s_dc[2][12], s_warp[33];
    if (threadIdx.x < 24) s_dc[threadIdx.x / 12][threadIdx.x % 12] = tab.dc[threadIdx.x / 12][threadIdx.x % 12];
    __syncthreads();
    const unsigned frame = blockIdx.y, r = blockIdx.x * kEntThreads + threadIdx.x;
    unsigned mine = 0;
    if (r < g.nblk) {
        const unsigned mcu = r / 6u;
        const int b = (int)(r % 6u);
        const unsigned *fm = meta + (size_t)frame * g.nblk;
        const int diff = resolved_dc_meta(fm, g, mcu, b) - dc_predictor_meta(fm, g, mcu, b);
        mine = (fm[r] >> 16) + dc_code_bits(diff, s_dc[b < 4 ? 0 : 1]);
        bits[(size_t)frame * g.nblk + r] = mine;
    }
    unsigned all;
    jblock_excl_scan(mine, s_warp, all);
    if (threadIdx.x == 0) part[(size_t)frame * nparts + blockIdx.x] = all;
}

// one CTA of 1024 threads per frame: bit offset of every group of blocks; total[frame]
__global__ void __launch_bounds__(1024) k_jpeg_scan(const unsigned *part, size_t nparts, unsigned long long *pbase,
                                                    unsigned long long *total)
{
    __shared__ unsigned s_warp[33];
    const unsigned frame = blockIdx.x;
    const unsigned *fp = part + (size_t)frame * nparts;
    unsigned long long *fb = pbase + (size_t)frame * nparts;
    unsigned long long carry = 0;
    for (size_t base = 0; base < nparts; base += 1024u) {   // 1024 groups hold < 2^32 bits
        const size_t i = base + threadIdx.x;
        const unsigned v = i < nparts ? fp[i] : 0u;
        unsigned all;
        const unsigned ex = jblock_excl_scan(v, s_warp, all);
        if (i < nparts) fb[i] = carry + ex;
        carry += all;
    }
    if (threadIdx.x == 0) total[frame] = carry;
}

// zero the words of each frame's stream that the entropy coder will OR into
__global__ void __launch_bounds__(256) k_jpeg_zero(unsigned *stream, size_t stream_words, const unsigned long long *total)
{
    const unsigned frame = blockIdx.y;
    const size_t words = min((size_t)((total[frame] + 31ull) / 32ull + 2ull), stream_words);
    unsigned *s = stream + (size_t)frame * stream_words;
    for (size_t k = (size_t)blockIdx.x * 256u + threadIdx.x; k < words; k += (size_t)gridDim.x * 256u) s[k] = 0u;
}

// grid (nparts, n): offset of a block = offset of its group + exclusive scan of bits[] inside the CTA
__global__ void __launch_bounds__(kEntThreads) k_jpeg_emit(const short *coef, Geom g, Tables tab, const unsigned *bits,
                                                           const unsigned long long *pbase, size_t nparts,
                                                           unsigned *stream, size_t stream_words)
{
    __shared__ unsigned s_dc[2][12], s_ac[2][256], s_warp[33];
    load_tables(tab, s_dc, s_ac);
    const unsigned frame = blockIdx.y, r = blockIdx.x * kEntThreads + threadIdx.x;
    const bool live = r < g.nblk;
    unsigned all;
    const unsigned ex = jblock_excl_scan(live ? bits[(size_t)frame * g.nblk + r] : 0u, s_warp, all);
    if (!live) return;
    const unsigned mcu = r / 6u;
    const int b = (int)(r % 6u), c = b < 4 ? 0 : 1;
    const short *fc = coef + (size_t)frame * g.nblk * 64u;
    __align__(16) short zz[64];
    load_block(fc, r, zz);
    const int diff = resolved_dc(fc, g, mcu, b) - dc_predictor(fc, g, mcu, b);
    StreamSink sink(stream + (size_t)frame * stream_words, pbase[(size_t)frame * nparts + blockIdx.x] + ex);
    encode_block<StreamSink, (RCT_JEMIT_UNROLL != 0)>(sink, zz, is_dummy(g, mcu, b), diff, s_dc[c], s_ac[c]);
    if (r == g.nblk - 1u) {  // jchuff.c flush_bits: pad the last byte with 1-bits
        const int k = (int)((8ull - (sink.position() & 7ull)) & 7ull);
        if (k) sink.put((1u << k) - 1u, k);
    }
    sink.finish();
}

__device__ __forceinline__ unsigned ff_bytes(const uint4 &v, size_t first, size_t nbytes)
{
    const unsigned w[4] = {v.x, v.y, v.z, v.w};
    unsigned c = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j)
        c += (first + (size_t)j < nbytes && ((w[j >> 2] >> (8 * (j & 3))) & 255u) == 255u) ? 1u : 0u;
    return c;
}
__device__ __forceinline__ uint4 stream_chunk(const unsigned *s, size_t first, size_t stream_words)
{
    return first + 16u <= stream_words * 4u ? *reinterpret_cast<const uint4 *>(reinterpret_cast<const char *>(s) + first)
                                            : make_uint4(0u, 0u, 0u, 0u);
}

__global__ void __launch_bounds__(kFfThreads) k_jpeg_ff_count(const unsigned *stream, size_t stream_words, size_t ntiles,
                                                              const unsigned long long *total, unsigned *tile_ff)
{
    __shared__ unsigned s_warp[33];
    const unsigned frame = blockIdx.y;
    const size_t nbytes = (size_t)((total[frame] + 7ull) / 8ull), used = (nbytes + kTile - 1u) / kTile;
    const unsigned *s = stream + (size_t)frame * stream_words;
    for (size_t tile = blockIdx.x; tile < used; tile += gridDim.x) {
        const size_t first = tile * kTile + (size_t)threadIdx.x * 16u;
        const unsigned c = ff_bytes(stream_chunk(s, first, stream_words), first, nbytes);
        unsigned all;
        jblock_excl_scan(c, s_warp, all);
        if (threadIdx.x == 0) tile_ff[(size_t)frame * ntiles + tile] = all;
    }
}

// one CTA of 1024 threads per frame: scan of the tile counts; header, EOI and the byte count
__global__ void __launch_bounds__(1024) k_jpeg_ff_scan(unsigned *tile_ff, size_t ntiles, const unsigned long long *total,
                                                       Header hdr, uint8_t *out, size_t out_stride, size_t out_cap,
                                                       unsigned long long *nbytes_out)
{
    __shared__ unsigned s_warp[33];
    const unsigned frame = blockIdx.x;
    const size_t nbytes = (size_t)((total[frame] + 7ull) / 8ull), used = (nbytes + kTile - 1u) / kTile;
    unsigned *t = tile_ff + (size_t)frame * ntiles;
    unsigned carry = 0;
    for (size_t base = 0; base < used; base += 1024u) {
        const size_t i = base + threadIdx.x;
        const unsigned v = i < used ? t[i] : 0u;
        unsigned all;
        const unsigned ex = jblock_excl_scan(v, s_warp, all);
        if (i < used) t[i] = carry + ex;
        carry += all;
    }
    uint8_t *o = out + (size_t)frame * out_stride;
    if ((int)threadIdx.x < hdr.len && threadIdx.x < out_cap) o[threadIdx.x] = hdr.b[threadIdx.x];
    if (threadIdx.x == 0) {
        const size_t end = (size_t)hdr.len + nbytes + carry;
        if (end < out_cap) o[end] = 0xFF;
        if (end + 1u < out_cap) o[end + 1u] = 0xD9;
        nbytes_out[frame] = end + 2u;   // > out_cap: the stream did not fit and is truncated
    }
}

__global__ void __launch_bounds__(kFfThreads) k_jpeg_ff_write(const unsigned *stream, size_t stream_words, size_t ntiles,
                                                              const unsigned long long *total, const unsigned *tile_ff,
                                                              int hdr_len, uint8_t *out, size_t out_stride, size_t out_cap)
{
    __shared__ unsigned s_warp[33];
    const unsigned frame = blockIdx.y;
    const size_t nbytes = (size_t)((total[frame] + 7ull) / 8ull), used = (nbytes + kTile - 1u) / kTile;
    const unsigned *s = stream + (size_t)frame * stream_words;
    uint8_t *o = out + (size_t)frame * out_stride;
    for (size_t tile = blockIdx.x; tile < used; tile += gridDim.x) {
        const size_t first = tile * kTile + (size_t)threadIdx.x * 16u;
        const uint4 v = stream_chunk(s, first, stream_words);
        const unsigned c = ff_bytes(v, first, nbytes);
        unsigned all;
        const unsigned ex = jblock_excl_scan(c, s_warp, all);
        size_t pos = (size_t)hdr_len + first + tile_ff[(size_t)frame * ntiles + tile] + ex;
        const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (first + (size_t)j >= nbytes) break;
            const unsigned byte = (w[j >> 2] >> (8 * (j & 3))) & 255u;
            if (pos < out_cap) o[pos] = (uint8_t)byte;
            ++pos;
            if (byte == 255u) {
                if (pos < out_cap) o[pos] = 0;
                ++pos;
            }
        }
    }
}

// enqueue the encoder for n frames on `st`.  rgb [n] frames of H x W x 3 at g.frame_stride; out [n][out_stride] bytes,
// nbytes[n] device counters (bytes needed; a value above out_cap marks a truncated stream).
inline cudaError_t encode_launch(const uint8_t *rgb, const Geom &g, int n, int quality, const Workspace &ws, uint8_t *out,
                                 size_t out_stride, size_t out_cap, unsigned long long *nbytes, int sm_count,
                                 cudaStream_t st, int *launches)
{
    Tables tab;
    Header hdr;
    make_tables(quality, &tab, &hdr, g.H, g.W);
    const unsigned wide = (unsigned)(sm_count > 0 ? sm_count : 148) * 8u;
    const unsigned gx = (unsigned)(ws.ntiles < wide ? ws.ntiles : wide);
    k_jpeg_dct<<<dim3((g.nmcu + 31u) / 32u, (unsigned)n), kDctThreads, 0, st>>>(rgb, g, tab, ws.coef, ws.meta);
    const dim3 ent((unsigned)ws.nparts, (unsigned)n);
    k_jpeg_bits<<<ent, kEntThreads, 0, st>>>(ws.meta, g, tab, ws.bits, ws.part, ws.nparts);
    k_jpeg_scan<<<n, 1024, 0, st>>>(ws.part, ws.nparts, ws.pbase, ws.total);
    k_jpeg_zero<<<dim3(wide, n), 256, 0, st>>>(ws.stream, ws.stream_words, ws.total);
    k_jpeg_emit<<<ent, kEntThreads, 0, st>>>(ws.coef, g, tab, ws.bits, ws.pbase, ws.nparts, ws.stream, ws.stream_words);
    k_jpeg_ff_count<<<dim3(gx, n), kFfThreads, 0, st>>>(ws.stream, ws.stream_words, ws.ntiles, ws.total, ws.tile_ff);
    k_jpeg_ff_scan<<<n, 1024, 0, st>>>(ws.tile_ff, ws.ntiles, ws.total, hdr, out, out_stride, out_cap, nbytes);
    k_jpeg_ff_write<<<dim3(gx, n), kFfThreads, 0, st>>>(ws.stream, ws.stream_words, ws.ntiles, ws.total, ws.tile_ff,
                                                       hdr.len, out, out_stride, out_cap);
    if (launches) *launches += 8;
    return cudaGetLastError();
}
#endif  // __CUDACC__

}  // namespace jpeg
