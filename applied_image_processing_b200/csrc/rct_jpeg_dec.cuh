// SURVEY.md 8f row f-3, decode half: `Image.open(path)` of the content / stylised files
// (Style_3DGS/localized_style_transfer.py:200, 219) -- baseline sequential Huffman JPEG, 8-bit, grayscale or
// YCbCr with 4:4:4 / 4:2:2 / 4:2:0 sampling -> uint8 RGB (or gray) on the device, bit for bit what Pillow decodes
// (libjpeg defaults: islow IDCT, fancy chroma upsampling, table-equivalent YCbCr -> RGB).
//   host       parses the header (tables, frame, scan, restart interval) and the restart-marker positions
//   k_jdec_huff  thread per restart interval (one thread for a file without restart markers: correct, serial --
//                the self-synchronising parallel decode is the planned replacement) -> int16 coefficients
//   k_jdec_idct  thread per 8x8 block: dequantise, jidctint islow, range limit -> uint8 component planes
//   k_jdec_rgb   thread per output pixel: jdsample fancy upsampling (replication for chroma widths <= 2),
//                jdcolor arithmetic -> interleaved RGB
// The bodies are __host__ __device__; tools/jpeg_selftest --host-decode runs them serially on the CPU.
#pragma once
#include "rct_jpeg.cuh"

namespace jpeg {

struct DecFrame {
    int H, W, ncomp;
    int hs[3], vs[3], tq[3], td[3], ta[3];   // sampling factors, quantiser / DC / AC table of each component
    int hmax, vmax, mx, my;                   // MCUs across / down
    int bw[3], bh[3];                         // blocks across / down of each component (whole MCUs)
    int cw[3], ch[3];                         // real (downsampled) size of each component
    size_t coef_off[3];                       // first block of each component in the coefficient array
    size_t plane_off[3];                      // first sample of each component plane
    size_t nblocks, nsamples;
    int restart_interval;                     // MCUs per restart interval, 0 = none
    unsigned nintervals;
    size_t scan_bytes;
};

struct DecTables {
    unsigned short q[4][64];                  // natural order
    // canonical Huffman decoding (jdhuff.c slow path): for length l, codes <= maxcode[l] have that length
    int maxcode[4][18];                       // [table: 0,1 = DC 0/1; 2,3 = AC 0/1][length], -1 = none
    int valptr[4][17];                        // index into vals of the first code of each length, minus its code
    unsigned char vals[4][256];
};

// ---- bit reader over the entropy-coded segment (0xFF 0x00 unstuffed; a marker feeds zero bits) ----------
struct BitReader {
    const unsigned char *d;
    size_t p, end;
    unsigned long long acc;
    int n;
    JHD BitReader(const unsigned char *data, size_t begin, size_t end_) : d(data), p(begin), end(end_), acc(0ull), n(0) {}
    JHD void fill()
    {
        unsigned b = 0;
        if (p < end) {
            b = d[p];
            if (b == 0xFFu) {
                const unsigned nx = p + 1 < end ? d[p + 1] : 0xD9u;
                if (nx == 0u)
                    p += 2;
                else
                    b = 0u;  // marker: stay in front of it
            } else {
                ++p;
            }
        }
        acc = (acc << 8) | b;
        n += 8;
    }
    JHD unsigned bits(int k)
    {
        if (k == 0) return 0u;
        while (n < k) fill();
        n -= k;
        const unsigned v = (unsigned)(acc >> n) & ((1u << k) - 1u);
        acc &= (1ull << n) - 1ull;
        return v;
    }
};

JHD int huff_symbol(BitReader &br, const DecTables &t, int table)
{
    int code = (int)br.bits(1), l = 1;
    while (l <= 16 && code > t.maxcode[table][l]) {
        code = (code << 1) | (int)br.bits(1);
        ++l;
    }
    if (l > 16) return 0;  // corrupt data: libjpeg warns and returns 0
    return t.vals[table][(t.valptr[table][l] + code) & 255];
}
JHD int huff_extend(unsigned v, int s) { return s == 0 || (int)v >= (1 << (s - 1)) ? (int)v : (int)v - (1 << s) + 1; }

// decode MCUs [m0, m1) of one restart interval; coef must be zero beforehand (natural order, int16)
JHD void decode_interval(const unsigned char *scan, size_t begin, size_t end, const DecFrame &f, const DecTables &t,
                         unsigned m0, unsigned m1, short *coef)
{
    BitReader br(scan, begin, end);
    int pred[3] = {0, 0, 0};
    for (unsigned m = m0; m < m1; ++m) {
        const int j = (int)(m / (unsigned)f.mx), i = (int)(m % (unsigned)f.mx);
        for (int c = 0; c < f.ncomp; ++c)
            for (int dy = 0; dy < f.vs[c]; ++dy)
                for (int dx = 0; dx < f.hs[c]; ++dx) {
                    short *blk = coef + (f.coef_off[c] + (size_t)(j * f.vs[c] + dy) * (size_t)f.bw[c] + (size_t)(i * f.hs[c] + dx)) * 64u;
                    int s = huff_symbol(br, t, f.td[c]);
                    pred[c] += huff_extend(br.bits(s), s);
                    blk[0] = (short)pred[c];
                    for (int k = 1; k < 64;) {
                        const int rs = huff_symbol(br, t, 2 + f.ta[c]), r = rs >> 4;
                        s = rs & 15;
                        if (s == 0) {
                            if (r != 15) break;
                            k += 16;
                            continue;
                        }
                        k += r;
                        if (k > 63) break;  // corrupt data
                        blk[zigzag_natural(k)] = (short)huff_extend(br.bits(s), s);
                        ++k;
                    }
                }
    }
}

// ---- self-synchronising parallel Huffman decode (files without restart markers) -----------------------
// The scan is cut into subsequences of kSubBytes stuffed bytes.  A decoder state is (p, n, acc, bi, k): next byte to
// load, buffered bits, their value, block inside the MCU, next zigzag index (0 = a DC code comes next); its position
// label is 8p - n.  Subsequence i owns the code units that START at a label in [8 s_i, 8 s_{i+1}).
//   speculative pass  thread i decodes subsequence i from (s_i, 0, 0, 0, 0) -- wrong in general -- and keeps the state
//                     in which it leaves the subsequence, E[i].  Thread 0 starts from the true state.
//   rounds            E'[i] = state after decoding subsequence i from E[i-1].  Huffman codes resynchronise within a
//                     few dozen code units, so after a few rounds E' == E (measured on Pillow files at 256 bytes per
//                     subsequence: 1-2 rounds).  E' == E for every i is a proof, not a heuristic: E[0] is exact and every
//                     E[i] follows from E[i-1] by a real decode.  Without a fixed point after kSyncRounds the serial
//                     kernel takes over.
//   count / scan      blocks that start in each subsequence -> ordinal of its first block
//   write             decode once more from the exact entry states, coefficients to their blocks, DC as differences
//   DC prefix         running sum of the DC differences per component
#ifndef RCT_JSUB
#define RCT_JSUB 256
#endif
constexpr unsigned kSubBytes = RCT_JSUB;
#ifndef RCT_JSYNC
#define RCT_JSYNC 32
#endif
constexpr int kSyncRounds = RCT_JSYNC;

struct HState {
    unsigned long long p;
    unsigned n, acc, bi, k;
};
JHD bool same_state(const HState &a, const HState &b) { return a.p == b.p && a.n == b.n && a.bi == b.bi && a.k == b.k; }

struct SpanReader {
    const unsigned char *d;
    size_t p, end;
    unsigned long long acc;
    int n;
    bool over;
    JHD SpanReader(const unsigned char *data, size_t end_, const HState &s)
        : d(data), p((size_t)s.p), end(end_), acc(s.acc), n((int)s.n), over(false) {}
    JHD void fill()
    {
        unsigned b = 0;
        if (p < end) {
            b = d[p];
            if (b == 0xFFu) {
                if (p + 1 < end && d[p + 1] == 0u) {
                    p += 2;
                } else {
                    b = 0u;
                    p = end;
                }
            } else {
                ++p;
            }
        } else {
            over = true;
        }
        acc = (acc << 8) | b;
        n += 8;
    }
    JHD unsigned bits(int k)
    {
        if (k == 0) return 0u;
        while (n < k) fill();
        n -= k;
        const unsigned v = (unsigned)(acc >> n) & ((1u << k) - 1u);
        acc &= (1ull << n) - 1ull;
        return v;
    }
    JHD long long label() const { return 8ll * (long long)p - (long long)n; }
};

template <class R>
JHD int huff_symbol_r(R &br, const DecTables &t, int table)
{
    int code = (int)br.bits(1), l = 1;
    while (l <= 16 && code > t.maxcode[table][l]) {
        code = (code << 1) | (int)br.bits(1);
        ++l;
    }
    if (l > 16) return 0;
    return t.vals[table][(t.valptr[table][l] + code) & 255];
}

// component of block `bi` of an MCU
JHD int mcu_component(const DecFrame &f, unsigned bi)
{
    const unsigned n0 = (unsigned)(f.hs[0] * f.vs[0]);
    return f.ncomp == 1 || bi < n0 ? 0 : (int)(bi - n0) + 1;
}
JHD unsigned blocks_per_mcu(const DecFrame &f) { return f.ncomp == 1 ? 1u : (unsigned)(f.hs[0] * f.vs[0]) + 2u; }

// first byte of subsequence i (a stuffed 0x00 belongs to the 0xFF in front of it)
JHD size_t sub_start(const unsigned char *scan, size_t nbytes, unsigned i)
{
    size_t p = (size_t)i * kSubBytes;
    if (i && p < nbytes && scan[p] == 0u && scan[p - 1] == 0xFFu) ++p;
    return p < nbytes ? p : nbytes;
}

struct NoSink {
    JHD void dc(unsigned long long, int) {}
    JHD void ac(unsigned long long, int, int) {}
};
// coefficient block of the ord-th block of the scan
JHD short *scan_block(short *coef, const DecFrame &f, unsigned long long ord)
{
    const unsigned bpm = blocks_per_mcu(f);
    const unsigned m = (unsigned)(ord / bpm), bi = (unsigned)(ord % bpm);
    const int c = mcu_component(f, bi);
    const unsigned within = c == 0 ? bi : 0u;
    const int j = (int)(m / (unsigned)f.mx), i = (int)(m % (unsigned)f.mx);
    const int dy = (int)within / f.hs[c], dx = (int)within % f.hs[c];
    return coef + (f.coef_off[c] + (size_t)(j * f.vs[c] + dy) * (size_t)f.bw[c] + (size_t)(i * f.hs[c] + dx)) * 64u;
}
struct CoefSink {
    short *coef;
    const DecFrame *f;
    unsigned long long total;
    JHD void dc(unsigned long long ord, int diff)
    {
        if (ord < total) scan_block(coef, *f, ord)[0] = (short)diff;
    }
    JHD void ac(unsigned long long ord, int k, int v)
    {
        if (ord < total) scan_block(coef, *f, ord)[zigzag_natural(k)] = (short)v;
    }
};

// decode the code units that start before `limit` (a label), from state `s`; `first_ord` = ordinal of the first block
// that STARTS here (a unit continuing an earlier block belongs to first_ord - 1).  Returns the blocks started.
template <class SINK>
JHD unsigned decode_span(const unsigned char *scan, size_t nbytes, HState &s, long long limit, const DecFrame &f,
                         const DecTables &t, unsigned long long first_ord, SINK &sink)
{
    SpanReader br(scan, nbytes, s);
    const unsigned bpm = blocks_per_mcu(f);
    unsigned bi = s.bi, k = s.k, started = 0;
    while (br.label() < limit && !br.over) {
        const int c = mcu_component(f, bi);
        if (k == 0) {
            const int sz = huff_symbol_r(br, t, f.td[c]);
            const int diff = huff_extend(br.bits(sz), sz);
            sink.dc(first_ord + started, diff);
            ++started;
            k = 1;
        } else {
            const int rs = huff_symbol_r(br, t, 2 + f.ta[c]), r = rs >> 4, sz = rs & 15;
            if (sz == 0) {
                k = r == 15 ? k + 16 : 64;
            } else {
                k += (unsigned)r;
                const int v = huff_extend(br.bits(sz), sz);
                if (k < 64) sink.ac(first_ord + started - 1ull, (int)k, v);
                ++k;
            }
        }
        if (k >= 64) {
            k = 0;
            bi = bi + 1 == bpm ? 0 : bi + 1;
        }
    }
    s.p = br.p, s.n = (unsigned)br.n, s.acc = (unsigned)br.acc, s.bi = bi, s.k = k;
    return started;
}
JHD long long sub_limit(const unsigned char *scan, size_t nbytes, unsigned i, unsigned nsub)
{
    return i + 1 < nsub ? 8ll * (long long)sub_start(scan, nbytes, i + 1) : 0x7fffffffffffffffll;
}
JHD HState sub_entry(const unsigned char *scan, size_t nbytes, unsigned i, const HState *E)
{
    if (i) return E[i - 1];
    HState s;
    s.p = 0, s.n = 0, s.acc = 0, s.bi = 0, s.k = 0;
    return s;
}

// ---- IDCT (jidctint.c jpeg_idct_islow) ---------------------------------------------------------
template <bool FIRST>
JHD void idct8(int *d, int s)
{
    const int i0 = d[0], i1 = d[s], i2 = d[2 * s], i3 = d[3 * s], i4 = d[4 * s], i5 = d[5 * s], i6 = d[6 * s], i7 = d[7 * s];
    int z1 = (i2 + i6) * 4433;
    const int t2 = z1 - i6 * 15137, t3 = z1 + i2 * 6270;
    const int t0 = (i0 + i4) * 8192, t1 = (i0 - i4) * 8192;
    const int t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
    int o0 = i7, o1 = i5, o2 = i3, o3 = i1;
    z1 = o0 + o3;
    int z2 = o1 + o2, z3 = o0 + o2, z4 = o1 + o3;
    const int z5 = (z3 + z4) * 9633;
    o0 *= 2446, o1 *= 16819, o2 *= 25172, o3 *= 12299;
    z1 *= -7373, z2 *= -20995;
    z3 = z3 * -16069 + z5;
    z4 = z4 * -3196 + z5;
    o0 += z1 + z3, o1 += z2 + z4, o2 += z2 + z3, o3 += z1 + z4;
    constexpr int sh = FIRST ? 13 - 2 : 13 + 2 + 3;
    d[0] = descale(t10 + o3, sh), d[7 * s] = descale(t10 - o3, sh);
    d[s] = descale(t11 + o2, sh), d[6 * s] = descale(t11 - o2, sh);
    d[2 * s] = descale(t12 + o1, sh), d[5 * s] = descale(t12 - o1, sh);
    d[3 * s] = descale(t13 + o0, sh), d[4 * s] = descale(t13 - o0, sh);
}

// one block: coefficients (natural order) -> 64 samples 0..255 written into the component plane (row stride `stride`)
JHD void idct_block(const short *coef, const unsigned short *q, unsigned char *dst, size_t stride)
{
    int blk[64];
#pragma unroll
    for (int k = 0; k < 64; ++k) blk[k] = (int)coef[k] * (int)q[k];
#pragma unroll
    for (int c = 0; c < 8; ++c) idct8<true>(blk + c, 8);
#pragma unroll
    for (int r = 0; r < 8; ++r) idct8<false>(blk + 8 * r, 1);
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int v = blk[r * 8 + c] + 128;
            dst[(size_t)r * stride + (size_t)c] = (unsigned char)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
}

// ---- upsampling + colour ---------------------------------------------------------------------
// sample of a chroma plane at full-resolution position (y, x); fh, fv = hmax / hs, vmax / vs of the component
JHD int upsampled(const unsigned char *P, size_t stride, int cw, int ch, int fh, int fv, int y, int x)
{
    if (fh == 1 && fv == 1) return P[(size_t)y * stride + (size_t)x];
    const int c = fh == 2 ? x >> 1 : x, r = fv == 2 ? y >> 1 : y;
    if (fh == 2 && cw <= 2) return P[(size_t)r * stride + (size_t)c];   // jinit_upsampler: fancy needs downsampled_width > 2
    if (fv == 1) {                                                       // h2v1_fancy_upsample
        const int v = P[(size_t)r * stride + (size_t)c];
        if ((x & 1) == 0) return c == 0 ? v : (3 * v + P[(size_t)r * stride + (size_t)(c - 1)] + 1) >> 2;
        return c == cw - 1 ? v : (3 * v + P[(size_t)r * stride + (size_t)(c + 1)] + 2) >> 2;
    }
    const int rn = (y & 1) == 0 ? (r > 0 ? r - 1 : 0) : (r < ch - 1 ? r + 1 : ch - 1);
    const unsigned char *a = P + (size_t)r * stride, *b = P + (size_t)rn * stride;
    const int cur = 3 * a[c] + b[c];
    if ((x & 1) == 0) return c == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + 3 * a[c - 1] + b[c - 1] + 8) >> 4;
    return c == cw - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + 3 * a[c + 1] + b[c + 1] + 7) >> 4;
}

JHD int clamp255(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }
JHD void ycc_rgb(int y, int cb, int cr, unsigned char *rgb)
{
    const int xb = cb - 128, xr = cr - 128;
    rgb[0] = (unsigned char)clamp255(y + ((91881 * xr + 32768) >> 16));
    rgb[1] = (unsigned char)clamp255(y + ((-22554 * xb + 32768 - 46802 * xr) >> 16));
    rgb[2] = (unsigned char)clamp255(y + ((116130 * xb + 32768) >> 16));
}

JHD void output_pixel(const unsigned char *planes, const DecFrame &f, int y, int x, unsigned char *out)
{
    const int Y = planes[f.plane_off[0] + (size_t)y * (size_t)(f.bw[0] * 8) + (size_t)x];
    if (f.ncomp == 1) {
        out[(size_t)y * (size_t)f.W + (size_t)x] = (unsigned char)Y;
        return;
    }
    int cc[2];
#pragma unroll
    for (int c = 1; c < 3; ++c)
        cc[c - 1] = upsampled(planes + f.plane_off[c], (size_t)(f.bw[c] * 8), f.cw[c], f.ch[c], f.hmax / f.hs[c],
                              f.vmax / f.vs[c], y, x);
    ycc_rgb(Y, cc[0], cc[1], out + ((size_t)y * (size_t)f.W + (size_t)x) * 3u);
}

// ---- host: header parse -------------------------------------------------------------------------
struct DecPlan {
    DecFrame f;
    DecTables t;
    size_t scan_begin, scan_end;                         // entropy-coded bytes of the file: [scan_begin, scan_end)
    unsigned long long *intervals;                       // malloc'ed [nintervals][2]: byte range of each restart interval,
                                                         // relative to scan_begin
    const char *error;
};
inline void free_plan(DecPlan *p)
{
    free(p->intervals);
    p->intervals = nullptr;
}

// returns false with p->error set for anything that is not a baseline file of the supported layouts
inline bool parse_jpeg(const unsigned char *d, size_t n, DecPlan *p)
{
    memset(p, 0, sizeof *p);
    auto bad = [&](const char *why) {
        p->error = why;
        return false;
    };
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) return bad("not a JPEG file");
    size_t pos = 2;
    bool have_frame = false, have_q[4] = {false, false, false, false}, have_h[4] = {false, false, false, false};
    DecFrame &f = p->f;
    for (;;) {
        if (pos + 4 > n) return bad("truncated file");
        if (d[pos] != 0xFF) return bad("marker expected");
        const unsigned marker = d[pos + 1];
        if (marker == 0xFF) {
            ++pos;
            continue;
        }
        pos += 2;
        const size_t len = ((size_t)d[pos] << 8) | d[pos + 1];
        if (len < 2 || pos + len > n) return bad("truncated segment");
        const unsigned char *seg = d + pos + 2;
        const size_t sl = len - 2;
        if (marker == 0xDB) {
            for (size_t k = 0; k + 65 <= sl; k += 65) {
                if (seg[k] >> 4) return bad("16-bit quantisation table");
                const int id = seg[k] & 15;
                if (id > 3) return bad("quantisation table id");
                for (int i = 0; i < 64; ++i) p->t.q[id][zigzag_natural(i)] = seg[k + 1 + i];
                have_q[id] = true;
            }
        } else if (marker == 0xC4) {
            size_t k = 0;
            while (k + 17 <= sl) {
                const int tc = seg[k] >> 4, th = seg[k] & 15;
                if (tc > 1 || th > 1) return bad("Huffman table id (baseline allows 0 and 1)");
                const int tab = tc * 2 + th;
                int total = 0;
                for (int l = 1; l <= 16; ++l) total += seg[k + l];
                if (total > 256 || k + 17 + (size_t)total > sl) return bad("bad Huffman table");
                int code = 0, q = 0;
                for (int l = 1; l <= 16; ++l) {  // jdhuff.c jpeg_make_d_derived_tbl
                    const int cnt = seg[k + l];
                    if (cnt) {
                        p->t.valptr[tab][l] = q - code;
                        q += cnt;
                        code += cnt;
                        p->t.maxcode[tab][l] = code - 1;
                    } else {
                        p->t.valptr[tab][l] = 0;
                        p->t.maxcode[tab][l] = -1;
                    }
                    code <<= 1;
                }
                p->t.maxcode[tab][17] = 0x7fffffff;
                // symbol ranges of 8-bit baseline data (jdhuff.c): DC size categories 0..11; AC run/size bytes with
                // size 0..10 -- anything else would feed shifts by up to 255 bits into the decoder
                for (int i = 0; i < total; ++i) {
                    const int v = seg[k + 17 + i];
                    if (tc == 0 ? v > 11 : (v & 15) > 10) return bad("Huffman table holds symbols outside the 8-bit baseline range");
                }
                memcpy(p->t.vals[tab], seg + k + 17, (size_t)total);
                have_h[tab] = true;
                k += 17 + (size_t)total;
            }
        } else if (marker == 0xC0 || marker == 0xC1) {
            if (sl < 6 || seg[0] != 8) return bad("8-bit samples only");
            f.H = (seg[1] << 8) | seg[2], f.W = (seg[3] << 8) | seg[4], f.ncomp = seg[5];
            if (f.H <= 0 || f.W <= 0) return bad("empty frame");
            if (f.ncomp != 1 && f.ncomp != 3) return bad("1 or 3 components only");
            if (sl < 6 + 3 * (size_t)f.ncomp) return bad("truncated frame header");
            for (int c = 0; c < f.ncomp; ++c) {
                f.hs[c] = seg[7 + 3 * c] >> 4, f.vs[c] = seg[7 + 3 * c] & 15, f.tq[c] = seg[8 + 3 * c];
                if (f.tq[c] > 3) return bad("quantisation table id");
            }
            if (f.ncomp == 1) f.hs[0] = f.vs[0] = 1;  // a single component is never interleaved
            have_frame = true;
        } else if (marker >= 0xC2 && marker <= 0xCF && marker != 0xC4 && marker != 0xC8 && marker != 0xCC) {
            return bad("only baseline sequential Huffman JPEG (SOF0/SOF1)");
        } else if (marker == 0xDD) {
            if (sl < 2) return bad("truncated DRI");
            f.restart_interval = (seg[0] << 8) | seg[1];
        } else if (marker == 0xDA) {
            if (!have_frame) return bad("scan before frame header");
            if (sl < 1 || seg[0] != f.ncomp || sl < 1 + 2 * (size_t)f.ncomp + 3) return bad("one interleaved scan of all components expected");
            for (int c = 0; c < f.ncomp; ++c) {
                f.td[c] = seg[2 + 2 * c] >> 4, f.ta[c] = seg[2 + 2 * c] & 15;
                if (f.td[c] > 1 || f.ta[c] > 1 || !have_h[f.td[c]] || !have_h[2 + f.ta[c]]) return bad("missing Huffman table");
                if (!have_q[f.tq[c]]) return bad("missing quantisation table");
            }
            pos += len;
            break;
        }
        pos += len;
    }
    if (f.ncomp == 3) {
        if (f.hs[1] != 1 || f.vs[1] != 1 || f.hs[2] != 1 || f.vs[2] != 1) return bad("chroma sampling factors other than 1x1");
        if (!((f.hs[0] == 1 || f.hs[0] == 2) && (f.vs[0] == 1 || f.vs[0] == 2)) || (f.hs[0] == 1 && f.vs[0] == 2))
            return bad("luma sampling other than 1x1, 2x1, 2x2");
    }
    f.hmax = f.hs[0], f.vmax = f.vs[0];
    f.mx = (f.W + 8 * f.hmax - 1) / (8 * f.hmax), f.my = (f.H + 8 * f.vmax - 1) / (8 * f.vmax);
    size_t blocks = 0, samples = 0;
    for (int c = 0; c < f.ncomp; ++c) {
        f.bw[c] = f.mx * f.hs[c], f.bh[c] = f.my * f.vs[c];
        f.cw[c] = (f.W * f.hs[c] + f.hmax - 1) / f.hmax, f.ch[c] = (f.H * f.vs[c] + f.vmax - 1) / f.vmax;
        f.coef_off[c] = blocks, f.plane_off[c] = samples;
        blocks += (size_t)f.bw[c] * (size_t)f.bh[c];
        samples += (size_t)f.bw[c] * (size_t)f.bh[c] * 64u;
    }
    f.nblocks = blocks, f.nsamples = samples;
    // the entropy-coded segment and its restart intervals
    p->scan_begin = pos;
    const unsigned nmcu = (unsigned)f.mx * (unsigned)f.my;
    const unsigned want = f.restart_interval ? (nmcu + (unsigned)f.restart_interval - 1u) / (unsigned)f.restart_interval : 1u;
    p->intervals = (unsigned long long *)malloc(sizeof(unsigned long long) * 2u * want);
    if (!p->intervals) return bad("out of host memory");
    unsigned found = 0;
    size_t start = pos, q = pos;
    while (q < n) {
        const unsigned char *hit = (const unsigned char *)memchr(d + q, 0xFF, n - q);
        if (!hit) {
            q = n;
            break;
        }
        q = (size_t)(hit - d);
        const unsigned nx = q + 1 < n ? d[q + 1] : 0xD9u;
        if (nx == 0x00 || nx == 0xFF) {
            q += nx == 0x00 ? 2 : 1;
            continue;
        }
        if (nx >= 0xD0 && nx <= 0xD7 && f.restart_interval && found + 1 < want) {
            p->intervals[2 * found] = start - pos, p->intervals[2 * found + 1] = q - pos;
            ++found;
            q += 2;
            start = q;
            continue;
        }
        break;  // EOI or another marker: end of the scan
    }
    p->intervals[2 * found] = start - pos, p->intervals[2 * found + 1] = q - pos;
    ++found;
    if (found != want) {
        free_plan(p);
        return bad("restart markers do not match the restart interval");
    }
    f.nintervals = found;
    p->scan_end = q;
    f.scan_bytes = q - pos;
    return true;
}

constexpr unsigned kDcChunk = 256;  // blocks per CTA of the DC prefix kernels
struct DecWorkspace {
    unsigned char *scan;
    unsigned long long *intervals;
    short *coef;
    unsigned char *planes;
    // self-synchronising decode (files without restart markers)
    HState *E0, *E1;
    unsigned *flags;              // [kSyncRounds] 1 = that round still changed a state
    unsigned *nb;                 // [nsub] blocks started per subsequence
    unsigned long long *base;     // [nsub] ordinal of the first block started in the subsequence
    int *dcsum;                   // [3][dc_chunks] sums of the DC differences per chunk of kDcChunk blocks
    unsigned dc_chunks;
    unsigned nsub;
    size_t bytes;
};
inline DecWorkspace carve_dec(const DecFrame &f, void *base)
{
    DecWorkspace w;
    char *p = (char *)base;
    size_t o = 0;
    auto take = [&](size_t bytes) {
        char *r = p ? p + o : nullptr;
        o += align256(bytes);
        return (void *)r;
    };
    w.scan = (unsigned char *)take(f.scan_bytes + 16u);
    w.intervals = (unsigned long long *)take((size_t)f.nintervals * 16u);
    w.coef = (short *)take(f.nblocks * 64u * sizeof(short));
    w.planes = (unsigned char *)take(f.nsamples);
    w.nsub = (unsigned)((f.scan_bytes + kSubBytes - 1u) / kSubBytes);
    w.E0 = (HState *)take((size_t)w.nsub * sizeof(HState));
    w.E1 = (HState *)take((size_t)w.nsub * sizeof(HState));
    w.flags = (unsigned *)take((size_t)kSyncRounds * sizeof(unsigned));
    w.nb = (unsigned *)take((size_t)w.nsub * sizeof(unsigned));
    w.base = (unsigned long long *)take((size_t)w.nsub * sizeof(unsigned long long));
    w.dc_chunks = (unsigned)(f.nblocks / kDcChunk + 1u);
    w.dcsum = (int *)take((size_t)3 * w.dc_chunks * sizeof(int));
    w.bytes = o;
    return w;
}

#ifdef __CUDACC__
__global__ void __launch_bounds__(64) k_jdec_huff(const unsigned char *scan, const unsigned long long *intervals, DecFrame f,
                                                  DecTables t, short *coef)
{
    const unsigned k = blockIdx.x * 64u + threadIdx.x;
    if (k >= f.nintervals) return;
    const unsigned nmcu = (unsigned)f.mx * (unsigned)f.my;
    const unsigned ri = f.restart_interval ? (unsigned)f.restart_interval : nmcu;
    const unsigned m0 = k * ri, m1 = min(m0 + ri, nmcu);
    decode_interval(scan, (size_t)intervals[2 * k], (size_t)intervals[2 * k + 1], f, t, m0, m1, coef);
}

__global__ void __launch_bounds__(128) k_jdec_idct(const short *coef, DecFrame f, DecTables t, unsigned char *planes)
{
    const size_t i = (size_t)blockIdx.x * 128u + threadIdx.x;
    if (i >= f.nblocks) return;
    const int c = (f.ncomp == 3 && i >= f.coef_off[2]) ? 2 : (f.ncomp == 3 && i >= f.coef_off[1]) ? 1 : 0;
    const size_t local = i - f.coef_off[c], stride = (size_t)f.bw[c] * 8u;
    const size_t br = local / (size_t)f.bw[c], bc = local % (size_t)f.bw[c];
    idct_block(coef + i * 64u, t.q[f.tq[c]], planes + f.plane_off[c] + br * 8u * stride + bc * 8u, stride);
}

__global__ void __launch_bounds__(256) k_jdec_rgb(const unsigned char *planes, DecFrame f, unsigned char *out)
{
    const size_t i = (size_t)blockIdx.x * 256u + threadIdx.x;
    if (i >= (size_t)f.H * (size_t)f.W) return;
    output_pixel(planes, f, (int)(i / (size_t)f.W), (int)(i % (size_t)f.W), out);
}

// ---- self-synchronising decode: kernels (thread per subsequence) ------------------------------------
constexpr int kSyncThreads = 64;
__global__ void __launch_bounds__(kSyncThreads) k_jsync_spec(const unsigned char *scan, size_t nbytes, unsigned nsub, DecFrame f,
                                                             DecTables t, HState *E)
{
    const unsigned i = blockIdx.x * kSyncThreads + threadIdx.x;
    if (i >= nsub) return;
    HState s;
    s.p = sub_start(scan, nbytes, i), s.n = 0, s.acc = 0, s.bi = 0, s.k = 0;
    NoSink none;
    decode_span(scan, nbytes, s, sub_limit(scan, nbytes, i, nsub), f, t, 0ull, none);
    E[i] = s;
}

// round r: Eout[i] = state after subsequence i entered in Ein[i-1]; flags[r] = 1 if any state but the last changed.
// Once a round changed nothing the states are exact and the remaining rounds only copy.
__global__ void __launch_bounds__(kSyncThreads) k_jsync_round(const unsigned char *scan, size_t nbytes, unsigned nsub, DecFrame f,
                                                              DecTables t, const HState *Ein, HState *Eout, unsigned *flags, int r)
{
    const unsigned i = blockIdx.x * kSyncThreads + threadIdx.x;
    if (i >= nsub) return;
    if (r > 0 && flags[r - 1] == 0u) {
        Eout[i] = Ein[i];
        return;
    }
    HState s = sub_entry(scan, nbytes, i, Ein);
    NoSink none;
    decode_span(scan, nbytes, s, sub_limit(scan, nbytes, i, nsub), f, t, 0ull, none);
    if (i + 1u < nsub && !same_state(s, Ein[i])) flags[r] = 1u;
    Eout[i] = s;
}

__global__ void __launch_bounds__(kSyncThreads) k_jsync_count(const unsigned char *scan, size_t nbytes, unsigned nsub, DecFrame f,
                                                              DecTables t, const HState *E, const unsigned *flags, unsigned *nb)
{
    const unsigned i = blockIdx.x * kSyncThreads + threadIdx.x;
    if (i >= nsub || flags[kSyncRounds - 1] != 0u) return;
    HState s = sub_entry(scan, nbytes, i, E);
    NoSink none;
    nb[i] = decode_span(scan, nbytes, s, sub_limit(scan, nbytes, i, nsub), f, t, 0ull, none);
}

// one CTA of 1024 threads: exclusive scan of nb[] -> base[]
__global__ void __launch_bounds__(1024) k_jsync_scan(const unsigned *nb, unsigned nsub, const unsigned *flags, unsigned long long *base)
{
    __shared__ unsigned long long s_warp[33];
    if (flags[kSyncRounds - 1] != 0u) return;
    const unsigned chunk = (nsub + 1023u) / 1024u;
    const unsigned lo = min(threadIdx.x * chunk, nsub), hi = min(lo + chunk, nsub);
    unsigned long long sum = 0;
    for (unsigned j = lo; j < hi; ++j) sum += nb[j];
    unsigned long long all;
    unsigned long long run = jblock_excl_scan(sum, s_warp, all);
    for (unsigned j = lo; j < hi; ++j) {
        base[j] = run;
        run += nb[j];
    }
}

__global__ void __launch_bounds__(kSyncThreads) k_jsync_write(const unsigned char *scan, size_t nbytes, unsigned nsub, DecFrame f,
                                                              DecTables t, const HState *E, const unsigned *flags,
                                                              const unsigned long long *base, short *coef)
{
    const unsigned i = blockIdx.x * kSyncThreads + threadIdx.x;
    if (i >= nsub || flags[kSyncRounds - 1] != 0u) return;
    HState s = sub_entry(scan, nbytes, i, E);
    CoefSink sink;
    sink.coef = coef, sink.f = &f, sink.total = (unsigned long long)((unsigned)f.mx * (unsigned)f.my) * blocks_per_mcu(f);
    decode_span(scan, nbytes, s, sub_limit(scan, nbytes, i, nsub), f, t, base[i], sink);
}

// running sum of the DC differences of every component in scan order, in two launches: chunk sums (grid = chunks of
// kDcChunk blocks x components), then every chunk adds the sums of the chunks before it to its own inclusive scan.
// (One CTA per component walking the whole plane took 0.37 ms on a 4K file: 48 dependent strided loads per thread.)
__device__ __forceinline__ bool dc_geom(const DecFrame &f, int c, unsigned &n, unsigned &per, unsigned &first, unsigned &bpm)
{
    const unsigned nmcu = (unsigned)f.mx * (unsigned)f.my;
    bpm = blocks_per_mcu(f);
    per = c == 0 ? (unsigned)(f.hs[0] * f.vs[0]) : 1u;
    first = c == 0 ? 0u : (unsigned)(f.hs[0] * f.vs[0]) + (unsigned)c - 1u;
    n = nmcu * per;
    return true;
}
__global__ void __launch_bounds__(kDcChunk) k_jsync_dc_sum(const short *coef, DecFrame f, const unsigned *flags, int *dcsum, unsigned dc_chunks)
{
    __shared__ int s_warp[33];
    if (flags[kSyncRounds - 1] != 0u) return;
    const int c = blockIdx.y;
    unsigned n, per, first, bpm;
    dc_geom(f, c, n, per, first, bpm);
    const unsigned j = blockIdx.x * kDcChunk + threadIdx.x;
    if (blockIdx.x * kDcChunk >= n) return;
    int v = j < n ? (int)scan_block(const_cast<short *>(coef), f, (unsigned long long)(j / per) * bpm + first + j % per)[0] : 0;
    int all;
    jblock_excl_scan(v, s_warp, all);
    if (threadIdx.x == 0) dcsum[(size_t)c * dc_chunks + blockIdx.x] = all;
}
__global__ void __launch_bounds__(kDcChunk) k_jsync_dc_apply(short *coef, DecFrame f, const unsigned *flags, const int *dcsum, unsigned dc_chunks)
{
    __shared__ int s_warp[33];
    if (flags[kSyncRounds - 1] != 0u) return;
    const int c = blockIdx.y;
    unsigned n, per, first, bpm;
    dc_geom(f, c, n, per, first, bpm);
    if (blockIdx.x * kDcChunk >= n) return;
    int before = 0;  // sum of the chunks before this one
    for (unsigned q = threadIdx.x; q < blockIdx.x; q += kDcChunk) before += dcsum[(size_t)c * dc_chunks + q];
    int all;
    jblock_excl_scan(before, s_warp, all);
    const int prefix = all;
    __syncthreads();
    const unsigned j = blockIdx.x * kDcChunk + threadIdx.x;
    short *b = j < n ? scan_block(coef, f, (unsigned long long)(j / per) * bpm + first + j % per) : nullptr;
    const int v = b ? (int)b[0] : 0;
    int tot;
    const int ex = jblock_excl_scan(v, s_warp, tot);
    if (b) b[0] = (short)(prefix + ex + v);
}

// the serial decoder as the fallback of the parallel one: runs only if the rounds did not reach a fixed point
__global__ void k_jdec_huff_fallback(const unsigned char *scan, size_t nbytes, DecFrame f, DecTables t, const unsigned *flags, short *coef)
{
    if (flags[kSyncRounds - 1] == 0u) return;
    decode_interval(scan, 0, nbytes, f, t, 0u, (unsigned)f.mx * (unsigned)f.my, coef);
}

// file bytes on the host -> pixels on the device (H x W x 3, or H x W for a one-component file), enqueued on `st`
inline cudaError_t decode_launch(const unsigned char *file_host, const DecPlan &pl, const DecWorkspace &ws, unsigned char *out,
                                 cudaStream_t st, int *launches, bool parallel = false)
{
    const DecFrame &f = pl.f;
    cudaError_t e = cudaMemcpyAsync(ws.scan, file_host + pl.scan_begin, f.scan_bytes, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(ws.intervals, pl.intervals, (size_t)f.nintervals * 16u, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(ws.coef, 0, f.nblocks * 64u * sizeof(short), st);
    if (e != cudaSuccess) return e;
    if (parallel && f.nintervals == 1 && ws.nsub >= 4u) {
        e = cudaMemsetAsync(ws.flags, 0, (size_t)kSyncRounds * sizeof(unsigned), st);
        if (e != cudaSuccess) return e;
        const unsigned g = (ws.nsub + kSyncThreads - 1u) / kSyncThreads;
        k_jsync_spec<<<g, kSyncThreads, 0, st>>>(ws.scan, f.scan_bytes, ws.nsub, f, pl.t, ws.E0);
        for (int r = 0; r < kSyncRounds; ++r)
            k_jsync_round<<<g, kSyncThreads, 0, st>>>(ws.scan, f.scan_bytes, ws.nsub, f, pl.t, (r & 1) ? ws.E1 : ws.E0,
                                                     (r & 1) ? ws.E0 : ws.E1, ws.flags, r);
        const HState *E = (kSyncRounds & 1) ? ws.E1 : ws.E0;
        k_jsync_count<<<g, kSyncThreads, 0, st>>>(ws.scan, f.scan_bytes, ws.nsub, f, pl.t, E, ws.flags, ws.nb);
        k_jsync_scan<<<1, 1024, 0, st>>>(ws.nb, ws.nsub, ws.flags, ws.base);
        k_jsync_write<<<g, kSyncThreads, 0, st>>>(ws.scan, f.scan_bytes, ws.nsub, f, pl.t, E, ws.flags, ws.base, ws.coef);
        const unsigned dcg = (unsigned)(((unsigned long long)f.mx * f.my * (unsigned)(f.hs[0] * f.vs[0]) + kDcChunk - 1u) / kDcChunk);
        k_jsync_dc_sum<<<dim3(dcg, f.ncomp), kDcChunk, 0, st>>>(ws.coef, f, ws.flags, ws.dcsum, ws.dc_chunks);
        k_jsync_dc_apply<<<dim3(dcg, f.ncomp), kDcChunk, 0, st>>>(ws.coef, f, ws.flags, ws.dcsum, ws.dc_chunks);
        k_jdec_huff_fallback<<<1, 1, 0, st>>>(ws.scan, f.scan_bytes, f, pl.t, ws.flags, ws.coef);
        if (launches) *launches += 7 + kSyncRounds;
    } else {
        k_jdec_huff<<<(f.nintervals + 63u) / 64u, 64, 0, st>>>(ws.scan, ws.intervals, f, pl.t, ws.coef);
        if (launches) *launches += 1;
    }
    k_jdec_idct<<<(unsigned)((f.nblocks + 127u) / 128u), 128, 0, st>>>(ws.coef, f, pl.t, ws.planes);
    const size_t px = (size_t)f.H * (size_t)f.W;
    k_jdec_rgb<<<(unsigned)((px + 255u) / 256u), 256, 0, st>>>(ws.planes, f, out);
    if (launches) *launches += 2;
    return cudaGetLastError();
}
#endif

}  // namespace jpeg
