// rct_png.h -- PNG -> RGB8, HOST side (`Image.open(path).convert("RGB")`, LST:200, 219, accepts
// any format Pillow reads; the repository's inputs include PNG files).
//
// Deliberately on the host: a PNG is one zlib stream (inflate is a serial bit-level state machine with a 32 KB
// back-reference window) followed by per-scanline filters whose Average / Paeth predictors chain every byte to its
// left and upper neighbours -- neither stage has data parallelism a GPU could use short of a wavefront over
// anti-diagonals (H + W dependent steps).  So the file is inflated with zlib and unfiltered here, and the RGB pixels
// are uploaded once; everything after that (resample, transfer, JPEG encode) stays on the device.
// Scope: non-interlaced, bit depth 8 for grey / grey+alpha / RGB / RGBA, 1/2/4/8 for palette and grey; alpha is dropped
// exactly as Pillow's convert("RGB") does (no compositing), 1/2/4-bit grey is scaled to 0..255 as Pillow does.
// Interlaced and 16-bit files are refused with a reason, never decoded approximately.
// Damaged files follow Pillow 's verdicts (probed against PIL.Image.open().convert("RGB"), tests/test_png.py): a wrong
// CRC on any chunk in front of the first IDAT is an error, the CRCs of IDAT chunks and of what follows them are not
// looked at; a zlib stream that delivers every row is accepted even when its Adler-32 trailer or the IEND chunk is cut
// off or more data follows, a WRONG Adler-32 is an error, and so is a stream that ends before the last row.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <zlib.h>

#ifndef RCT_PNG_ZCHUNK
#define RCT_PNG_ZCHUNK (1u << 30)  // bytes handed to zlib per step: its counters are 32-bit, a PNG can hold > 4 GiB
#endif

namespace rct_png {

struct Info {
    int W = 0, H = 0, depth = 0, ctype = 0, interlace = 0;
    std::string error;
};

inline uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline bool parse_header(const uint8_t *f, size_t n, Info *info)
{
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (n < 33 || memcmp(f, sig, 8) != 0) { info->error = "not a PNG file"; return false; }
    if (be32(f + 8) != 13 || memcmp(f + 12, "IHDR", 4) != 0) { info->error = "PNG: IHDR missing"; return false; }
    const uint8_t *h = f + 16;
    info->W = (int)be32(h); info->H = (int)be32(h + 4);
    info->depth = h[8]; info->ctype = h[9]; info->interlace = h[12];
    if (info->W <= 0 || info->H <= 0 || info->W > 65535 || info->H > 65535) { info->error = "PNG: bad dimensions"; return false; }
    if (h[10] != 0 || h[11] != 0) { info->error = "PNG: unknown compression / filter method"; return false; }
    if (info->interlace) { info->error = "PNG: interlaced (Adam7) files are not supported"; return false; }
    const int d = info->depth, c = info->ctype;
    const bool ok = (c == 0 && (d == 1 || d == 2 || d == 4 || d == 8)) || (c == 3 && (d == 1 || d == 2 || d == 4 || d == 8)) ||
                    ((c == 2 || c == 4 || c == 6) && d == 8);
    if (!ok) {
        info->error = d == 16 ? "PNG: 16-bit samples are not supported" : "PNG: unsupported colour type / bit depth";
        return false;
    }
    return true;
}

inline int paeth(int a, int b, int c)
{
    const int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// file bytes -> rgb [H*W*3]; returns false with info->error set
inline bool decode(const uint8_t *f, size_t n, Info *info, uint8_t *rgb)
{
    if (!parse_header(f, n, info)) return false;
    std::vector<uint8_t> idat;
    uint8_t pal[256][3];
    memset(pal, 0, sizeof pal);
    int npal = 0;
    size_t pos = 8;
    bool end = false, seen_idat = false;
    while (pos + 12 <= n && !end) {
        const uint32_t len = be32(f + pos);
        const uint8_t *type = f + pos + 4, *data = f + pos + 8;
        if ((size_t)len > n - pos - 12) { info->error = "PNG: truncated chunk"; return false; }
        if (!memcmp(type, "IDAT", 4)) seen_idat = true;
        if (!seen_idat) {  // header, palette and everything else in front of the pixels: checksum enforced (as Pillow)
            uLong crc = crc32(0L, Z_NULL, 0);
            for (size_t o = 0, left = (size_t)len + 4; left;) {
                const uInt step = left > RCT_PNG_ZCHUNK ? (uInt)RCT_PNG_ZCHUNK : (uInt)left;
                crc = crc32(crc, type + o, step);
                o += step; left -= step;
            }
            if ((uint32_t)crc != be32(data + len)) {
                info->error = std::string("PNG: bad checksum in chunk ") + std::string((const char *)type, 4);
                return false;
            }
        }
        if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
        else if (!memcmp(type, "PLTE", 4)) {
            npal = (int)(len / 3) > 256 ? 256 : (int)(len / 3);
            for (int i = 0; i < npal; ++i) memcpy(pal[i], data + 3 * i, 3);
        } else if (!memcmp(type, "IEND", 4)) end = true;
        pos += 12 + (size_t)len;
    }
    if (idat.empty()) { info->error = "PNG: no image data"; return false; }
    const int W = info->W, H = info->H, d = info->depth, c = info->ctype;
    const int channels = c == 0 ? 1 : c == 2 ? 3 : c == 3 ? 1 : c == 4 ? 2 : 4;
    const int bpp = (channels * d + 7) / 8;                       // filter unit in bytes
    const size_t rowbytes = ((size_t)W * channels * d + 7) / 8;
    std::vector<uint8_t> raw((rowbytes + 1) * (size_t)H);
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    if (inflateInit(&zs) != Z_OK) { info->error = "PNG: zlib init failed"; return false; }
    // zlib counts avail_in / avail_out in 32 bits: feed both sides in steps so that rows x bytes >= 4 GiB inflate fully
    size_t in_left = idat.size(), out_left = raw.size();
    zs.next_in = idat.data(); zs.next_out = raw.data();
    int zr = Z_OK;
    do {  // Z_OK = progress was made; Z_BUF_ERROR = none possible: the input is used up or every row is there
        if (zs.avail_in == 0 && in_left) { zs.avail_in = in_left > RCT_PNG_ZCHUNK ? (uInt)RCT_PNG_ZCHUNK : (uInt)in_left; in_left -= zs.avail_in; }
        if (zs.avail_out == 0 && out_left) { zs.avail_out = out_left > RCT_PNG_ZCHUNK ? (uInt)RCT_PNG_ZCHUNK : (uInt)out_left; out_left -= zs.avail_out; }
        zr = inflate(&zs, Z_NO_FLUSH);
    } while (zr == Z_OK);
    const size_t got = raw.size() - out_left - zs.avail_out;
    inflateEnd(&zs);
    if ((zr != Z_STREAM_END && zr != Z_OK && zr != Z_BUF_ERROR) || got != raw.size()) {
        info->error = "PNG: corrupt or truncated zlib stream";
        return false;
    }
    std::vector<uint8_t> prev(rowbytes, 0), cur(rowbytes);
    for (int y = 0; y < H; ++y) {
        const uint8_t *src = raw.data() + (rowbytes + 1) * (size_t)y;
        const int ft = src[0];
        ++src;
        for (size_t i = 0; i < rowbytes; ++i) {
            const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = prev[i], cc = i >= (size_t)bpp ? prev[i - bpp] : 0;
            int v = src[i];
            switch (ft) {
            case 0: break;
            case 1: v += a; break;
            case 2: v += b; break;
            case 3: v += (a + b) >> 1; break;
            case 4: v += paeth(a, b, cc); break;
            default: info->error = "PNG: bad filter type"; return false;
            }
            cur[i] = (uint8_t)v;
        }
        uint8_t *o = rgb + (size_t)y * W * 3;
        for (int x = 0; x < W; ++x) {
            uint8_t r, g, bl;
            if (c == 2 || c == 6) { const uint8_t *p = cur.data() + (size_t)x * channels; r = p[0]; g = p[1]; bl = p[2]; }
            else if (c == 4) { r = g = bl = cur[(size_t)x * 2]; }
            else {
                unsigned s;
                if (d == 8) s = cur[x];
                else {
                    const int per = 8 / d, sh = (per - 1 - x % per) * d;
                    s = (cur[x / per] >> sh) & ((1u << d) - 1u);
                }
                if (c == 3) { const uint8_t *p = pal[s & 255u]; r = p[0]; g = p[1]; bl = p[2]; }
                else { r = g = bl = (uint8_t)(d == 8 ? s : s * 255u / ((1u << d) - 1u)); }
            }
            o[3 * x] = r; o[3 * x + 1] = g; o[3 * x + 2] = bl;
        }
        prev.swap(cur);
    }
    return true;
}

}  // namespace rct_png
