"""TEST INFRASTRUCTURE -- numpy restatement of the baseline JPEG encoder behind the reference's
`Image.fromarray(combined_image).save(".../localized_style_transfer_result.jpg")` (LST:241-243;
SURVEY.md 8f row f-3): Pillow defaults = libjpeg(-turbo) baseline, quality 75, 4:2:0, standard
Huffman tables, JDCT_ISLOW -- and, further down, of the baseline decoder behind `Image.open` (LST:200, 219).
The device encoder is csrc/rct_jpeg.cuh; the decoder half is the oracle only so far.  Only tests/ may
import this; the product never does.

The arithmetic is the published libjpeg algorithm (jccolor.c, jcsample.c, jfdctint.c,
jcdctmgr.c, jccoefct.c, jchuff.c; ITU T.81 Annex K tables), restated -- the dependency is not
vendored under /root/reference.  Pinned in tests/test_jpeg_oracle.py: the byte stream equals
Pillow's own output (Pillow 12.2.0 / libjpeg-turbo of this image) for every size tried.
"""
import numpy as np

# ITU T.81 Annex K.1 quantisation tables at quality 50, in zigzag order (as they sit in a DQT segment)
ZZ_LUMA_Q50 = [16, 11, 12, 14, 12, 10, 16, 14, 13, 14, 18, 17, 16, 19, 24, 40, 26, 24, 22, 22, 24, 49, 35, 37, 29, 40,
               58, 51, 61, 60, 57, 51, 56, 55, 64, 72, 92, 78, 64, 68, 87, 69, 55, 56, 80, 109, 81, 87, 95, 98, 103,
               104, 103, 62, 77, 113, 121, 112, 100, 120, 92, 101, 103, 99]
ZZ_CHROMA_Q50 = [17, 18, 18, 24, 21, 24, 47, 26, 26, 47, 99, 66, 56, 66, 99, 99] + [99] * 48

# Annex K.3 Huffman tables: (bits[1..16], values)
DHT = {
    0x00: ([0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0], list(range(12))),
    0x01: ([0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0], list(range(12))),
    0x10: ([0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 125],
           [1, 2, 3, 0, 4, 17, 5, 18, 33, 49, 65, 6, 19, 81, 97, 7, 34, 113, 20, 50, 129, 145, 161, 8, 35, 66, 177, 193,
            21, 82, 209, 240, 36, 51, 98, 114, 130, 9, 10, 22, 23, 24, 25, 26, 37, 38, 39, 40, 41, 42, 52, 53, 54, 55,
            56, 57, 58, 67, 68, 69, 70, 71, 72, 73, 74, 83, 84, 85, 86, 87, 88, 89, 90, 99, 100, 101, 102, 103, 104,
            105, 106, 115, 116, 117, 118, 119, 120, 121, 122, 131, 132, 133, 134, 135, 136, 137, 138, 146, 147, 148,
            149, 150, 151, 152, 153, 154, 162, 163, 164, 165, 166, 167, 168, 169, 170, 178, 179, 180, 181, 182, 183,
            184, 185, 186, 194, 195, 196, 197, 198, 199, 200, 201, 202, 210, 211, 212, 213, 214, 215, 216, 217, 218,
            225, 226, 227, 228, 229, 230, 231, 232, 233, 234, 241, 242, 243, 244, 245, 246, 247, 248, 249, 250]),
    0x11: ([0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 119],
           [0, 1, 2, 3, 17, 4, 5, 33, 49, 6, 18, 65, 81, 7, 97, 113, 19, 34, 50, 129, 8, 20, 66, 145, 161, 177, 193, 9,
            35, 51, 82, 240, 21, 98, 114, 209, 10, 22, 36, 52, 225, 37, 241, 23, 24, 25, 26, 38, 39, 40, 41, 42, 53,
            54, 55, 56, 57, 58, 67, 68, 69, 70, 71, 72, 73, 74, 83, 84, 85, 86, 87, 88, 89, 90, 99, 100, 101, 102, 103,
            104, 105, 106, 115, 116, 117, 118, 119, 120, 121, 122, 130, 131, 132, 133, 134, 135, 136, 137, 138, 146,
            147, 148, 149, 150, 151, 152, 153, 154, 162, 163, 164, 165, 166, 167, 168, 169, 170, 178, 179, 180, 181,
            182, 183, 184, 185, 186, 194, 195, 196, 197, 198, 199, 200, 201, 202, 210, 211, 212, 213, 214, 215, 216,
            217, 218, 226, 227, 228, 229, 230, 231, 232, 233, 234, 242, 243, 244, 245, 246, 247, 248, 249, 250]),
}

# zigzag position k -> natural (row-major) index inside the 8x8 block
ZIGZAG = [0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14,
          21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53,
          60, 61, 54, 47, 55, 62, 63]


def quant_tables(quality=75):
    """jcparam.c: jpeg_quality_scaling + jpeg_add_quant_table(force_baseline) -> two zigzag-ordered lists."""
    quality = min(max(int(quality), 1), 100)
    scale = 5000 // quality if quality < 50 else 200 - 2 * quality
    out = []
    for base in (ZZ_LUMA_Q50, ZZ_CHROMA_Q50):
        out.append([min(max((b * scale + 50) // 100, 1), 255) for b in base])
    return out


def rgb_to_ycc(rgb):
    """jccolor.c rgb_ycc_convert: 16-bit fixed point, Y rounds at ONE_HALF, Cb/Cr at CBCR_OFFSET + ONE_HALF - 1."""
    r, g, b = (rgb[..., i].astype(np.int64) for i in range(3))
    half, off = 1 << 15, 128 << 16
    y = (19595 * r + 38470 * g + 7471 * b + half) >> 16
    cb = (-11059 * r - 21709 * g + 32768 * b + off + half - 1) >> 16
    cr = (32768 * r - 27439 * g - 5329 * b + off + half - 1) >> 16
    return y, cb, cr


def h2v2_downsample(plane):
    """jcsample.c h2v2_downsample: 2x2 box sum, bias alternating 1, 2 along each output row."""
    s = plane[0::2, 0::2] + plane[0::2, 1::2] + plane[1::2, 0::2] + plane[1::2, 1::2]
    bias = 1 + (np.arange(s.shape[1]) & 1)
    return (s + bias[None, :]) >> 2


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def _fdct_pass(d, first):
    """One pass of jfdctint.c jpeg_fdct_islow along the last axis (CONST_BITS 13, PASS1_BITS 2)."""
    d0, d1, d2, d3, d4, d5, d6, d7 = (d[..., i] for i in range(8))
    t0, t7, t1, t6 = d0 + d7, d0 - d7, d1 + d6, d1 - d6
    t2, t5, t3, t4 = d2 + d5, d2 - d5, d3 + d4, d3 - d4
    t10, t13, t11, t12 = t0 + t3, t0 - t3, t1 + t2, t1 - t2
    sh = 13 - 2 if first else 13 + 2
    out = [None] * 8
    if first:
        out[0], out[4] = (t10 + t11) << 2, (t10 - t11) << 2
    else:
        out[0], out[4] = _descale(t10 + t11, 2), _descale(t10 - t11, 2)
    z1 = (t12 + t13) * 4433
    out[2] = _descale(z1 + t13 * 6270, sh)
    out[6] = _descale(z1 - t12 * 15137, sh)
    z1, z2, z3, z4 = t4 + t7, t5 + t6, t4 + t6, t5 + t7
    z5 = (z3 + z4) * 9633
    t4, t5, t6, t7 = t4 * 2446, t5 * 16819, t6 * 25172, t7 * 12299
    z1, z2, z3, z4 = z1 * -7373, z2 * -20995, z3 * -16069 + z5, z4 * -3196 + z5
    out[7] = _descale(t4 + z1 + z3, sh)
    out[5] = _descale(t5 + z2 + z4, sh)
    out[3] = _descale(t6 + z2 + z3, sh)
    out[1] = _descale(t7 + z1 + z4, sh)
    return np.stack(out, axis=-1)


def fdct_quantise(plane, zz_table):
    """jcdctmgr.c forward_DCT: level shift, islow DCT (rows then columns), quantise against 8*q with
    round-half-away.  plane (8*by, 8*bx) -> int64 (by, bx, 64) coefficients in natural order."""
    by, bx = plane.shape[0] // 8, plane.shape[1] // 8
    blk = plane.reshape(by, 8, bx, 8).transpose(0, 2, 1, 3).astype(np.int64) - 128
    blk = _fdct_pass(blk, True)                                     # rows
    blk = _fdct_pass(blk.swapaxes(-1, -2), False).swapaxes(-1, -2)  # columns
    q = np.empty(64, np.int64)
    q[ZIGZAG] = zz_table
    q8 = (q * 8).reshape(8, 8)
    mag = (np.abs(blk) + (q8 >> 1)) // q8
    return (np.sign(blk) * mag).reshape(by, bx, 64)


def component_blocks(rgb, quality=75):
    """Quantised blocks of the three components exactly as jccoefct.c compress_data hands them to the
    entropy coder: Y (2*my, 2*mx, 64) with libjpeg's dummy blocks, Cb and Cr (my, mx, 64)."""
    h, w = rgb.shape[:2]
    mx, my = -(-w // 16), -(-h // 16)
    by, bx = -(-h // 8), -(-w // 8)
    qy, qc = quant_tables(quality)
    y, cb, cr = rgb_to_ycc(rgb)
    # jcprepct.c / jcsample.c: right edge replicated at full resolution, bottom edge of each component
    ypad = np.pad(y, ((0, my * 16 - h), (0, mx * 16 - w)), mode="edge")
    yq = fdct_quantise(ypad, qy)
    # dummy blocks: AC = 0, DC = DC of the previous block of the MCU (order Y00 Y01 Y10 Y11)
    for j in range(my):
        for i in range(mx):
            prev = None
            for dy in range(2):
                for dx in range(2):
                    r, c = 2 * j + dy, 2 * i + dx
                    if r >= by or c >= bx:
                        yq[r, c, :] = 0
                        yq[r, c, 0] = prev
                    prev = yq[r, c, 0]
    out = [yq]
    for plane in (cb, cr):
        full = np.pad(plane, ((0, h & 1), (0, mx * 16 - w)), mode="edge")
        half = h2v2_downsample(full)
        half = np.pad(half, ((0, my * 8 - half.shape[0]), (0, 0)), mode="edge")
        out.append(fdct_quantise(half, qc))
    return out


def huffman_codes(bits, vals):
    """jchuff.c jpeg_make_c_derived_tbl: canonical codes -> {symbol: (code, length)}."""
    table, code, k = {}, 0, 0
    for length in range(1, 17):
        for _ in range(bits[length - 1]):
            table[vals[k]] = (code, length)
            code += 1
            k += 1
        code <<= 1
    return table


class _BitWriter:
    def __init__(self):
        self.out, self.acc, self.n = bytearray(), 0, 0

    def put(self, code, length):
        self.acc = (self.acc << length) | (code & ((1 << length) - 1))
        self.n += length
        while self.n >= 8:
            byte = (self.acc >> (self.n - 8)) & 0xFF
            self.out.append(byte)
            if byte == 0xFF:              # byte stuffing
                self.out.append(0)
            self.n -= 8
        self.acc &= (1 << self.n) - 1

    def flush(self):
        if self.n:
            self.put(0x7F, 8 - self.n)     # pad with 1-bits


def _encode_block(bw, block, last_dc, dc_tab, ac_tab):
    """jchuff.c encode_one_block."""
    diff = int(block[0]) - last_dc
    mag, low = (-diff, diff - 1) if diff < 0 else (diff, diff)
    nbits = mag.bit_length()
    bw.put(*dc_tab[nbits])
    if nbits:
        bw.put(low, nbits)
    run = 0
    for k in range(1, 64):
        v = int(block[ZIGZAG[k]])
        if v == 0:
            run += 1
            continue
        while run > 15:
            bw.put(*ac_tab[0xF0])
            run -= 16
        mag, low = (-v, v - 1) if v < 0 else (v, v)
        nbits = mag.bit_length()
        bw.put(*ac_tab[(run << 4) + nbits])
        bw.put(low, nbits)
        run = 0
    if run:
        bw.put(*ac_tab[0x00])
    return int(block[0])


def entropy_scan(yq, cbq, crq):
    """One interleaved scan, MCU = Y00 Y01 Y10 Y11 Cb Cr."""
    tabs = {k: huffman_codes(*v) for k, v in DHT.items()}
    bw, last = _BitWriter(), [0, 0, 0]
    my, mx = cbq.shape[:2]
    for j in range(my):
        for i in range(mx):
            for dy in range(2):
                for dx in range(2):
                    last[0] = _encode_block(bw, yq[2 * j + dy, 2 * i + dx], last[0], tabs[0x00], tabs[0x10])
            last[1] = _encode_block(bw, cbq[j, i], last[1], tabs[0x01], tabs[0x11])
            last[2] = _encode_block(bw, crq[j, i], last[2], tabs[0x01], tabs[0x11])
    bw.flush()
    return bytes(bw.out)


def _segment(marker, payload):
    return bytes([0xFF, marker]) + (len(payload) + 2).to_bytes(2, "big") + bytes(payload)


def encode_jpeg(rgb, quality=75):
    """uint8 (H, W, 3) -> the bytes Pillow's `Image.fromarray(rgb).save(f, "JPEG")` writes (LST:243)."""
    rgb = np.asarray(rgb)
    assert rgb.dtype == np.uint8 and rgb.ndim == 3 and rgb.shape[2] == 3
    h, w = rgb.shape[:2]
    qy, qc = quant_tables(quality)
    out = bytearray(b"\xff\xd8")
    out += _segment(0xE0, b"JFIF\x00\x01\x01\x00\x00\x01\x00\x01\x00\x00")
    out += _segment(0xDB, bytes([0] + qy))
    out += _segment(0xDB, bytes([1] + qc))
    out += _segment(0xC0, bytes([8]) + h.to_bytes(2, "big") + w.to_bytes(2, "big") +
                    bytes([3, 1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1]))
    for tc_th in (0x00, 0x10, 0x01, 0x11):
        bits, vals = DHT[tc_th]
        out += _segment(0xC4, bytes([tc_th] + bits + vals))
    out += _segment(0xDA, bytes([3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0]))
    out += entropy_scan(*component_blocks(rgb, quality))
    out += b"\xff\xd9"
    return bytes(out)


# ---------------------------------------------------------------------------------------------
# Decoder: the other half of row f-3, `Image.open(path)` of the content / stylised images
# (localized_style_transfer.py:200, 219).  Baseline sequential JPEG, 8-bit, Huffman, 1 or 3 components with
# luma sampling 1x1, 2x1, 1x2 or 2x2; libjpeg's defaults as Pillow uses them: islow IDCT (jidctint.c), fancy
# (triangle) chroma upsampling (jdsample.c), table-driven YCbCr -> RGB (jdcolor.c).  Pinned to Pillow's decode
# in tests/test_jpeg_oracle.py.
# ---------------------------------------------------------------------------------------------
def _idct_pass(d, first):
    """One pass of jidctint.c jpeg_idct_islow along the last axis."""
    i0, i1, i2, i3, i4, i5, i6, i7 = (d[..., k] for k in range(8))
    z1 = (i2 + i6) * 4433
    t2 = z1 - i6 * 15137
    t3 = z1 + i2 * 6270
    t0, t1 = (i0 + i4) << 13, (i0 - i4) << 13
    t10, t13, t11, t12 = t0 + t3, t0 - t3, t1 + t2, t1 - t2
    o0, o1, o2, o3 = i7, i5, i3, i1
    z1, z2, z3, z4 = o0 + o3, o1 + o2, o0 + o2, o1 + o3
    z5 = (z3 + z4) * 9633
    o0, o1, o2, o3 = o0 * 2446, o1 * 16819, o2 * 25172, o3 * 12299
    z1, z2, z3, z4 = z1 * -7373, z2 * -20995, z3 * -16069 + z5, z4 * -3196 + z5
    o0, o1, o2, o3 = o0 + z1 + z3, o1 + z2 + z4, o2 + z2 + z3, o3 + z1 + z4
    sh = 13 - 2 if first else 13 + 2 + 3
    out = [t10 + o3, t11 + o2, t12 + o1, t13 + o0, t13 - o0, t12 - o1, t11 - o2, t10 - o3]
    return np.stack([_descale(v, sh) for v in out], axis=-1)


def idct_blocks(coef, zz_table):
    """coef (..., 64) quantised, natural order -> samples (..., 8, 8) in 0..255 (dequantise, columns, rows, range limit)."""
    q = np.empty(64, np.int64)
    q[ZIGZAG] = zz_table
    blk = (coef.astype(np.int64) * q).reshape(coef.shape[:-1] + (8, 8))
    blk = _idct_pass(blk.swapaxes(-1, -2), True).swapaxes(-1, -2)   # columns
    blk = _idct_pass(blk, False)                                     # rows
    return np.clip(blk + 128, 0, 255)


class _BitReader:
    def __init__(self, data, pos):
        self.d, self.p, self.acc, self.n = data, pos, 0, 0

    def _fill(self):
        b = self.d[self.p]
        if b == 0xFF:
            nxt = self.d[self.p + 1]
            if nxt == 0:
                self.p += 2
            else:                       # a marker: feed zeros (libjpeg's behaviour at the end of the scan)
                b = 0
        else:
            self.p += 1
        self.acc = (self.acc << 8) | b
        self.n += 8

    def bits(self, k):
        if k == 0:
            return 0
        while self.n < k:
            self._fill()
        self.n -= k
        v = (self.acc >> self.n) & ((1 << k) - 1)
        self.acc &= (1 << self.n) - 1
        return v

    def symbol(self, lookup):
        code, length = 0, 0
        while True:
            code = (code << 1) | self.bits(1)
            length += 1
            s = lookup.get((length, code))
            if s is not None:
                return s
            if length > 16:
                raise ValueError("bad Huffman code")

    def restart(self):
        self.acc = self.n = 0
        assert self.d[self.p] == 0xFF and 0xD0 <= self.d[self.p + 1] <= 0xD7, "restart marker expected"
        self.p += 2


def _extend(v, nbits):
    return v if nbits == 0 or v >= (1 << (nbits - 1)) else v - (1 << nbits) + 1


def _fancy_h2(row_sums, bias_a, bias_b, shift):
    """jdsample.c triangle filter along a row: 3/4 nearer + 1/4 further sample; edges use the sample alone."""
    this = row_sums
    last = np.concatenate([this[:, :1], this[:, :-1]], axis=1)
    nxt = np.concatenate([this[:, 1:], this[:, -1:]], axis=1)
    out = np.empty((this.shape[0], this.shape[1] * 2), np.int64)
    out[:, 0::2] = (this * 3 + last + bias_a) >> shift
    out[:, 1::2] = (this * 3 + nxt + bias_b) >> shift
    return out


def upsample(plane, hs, vs):
    """Chroma plane (real downsampled size) -> (rows*vs, cols*hs), jdsample.c with do_fancy_upsampling."""
    p = plane.astype(np.int64)
    if hs == 1 and vs == 1:
        return p
    if hs == 2 and p.shape[1] <= 2:                            # jinit_upsampler: fancy only if downsampled_width > 2
        return np.repeat(np.repeat(p, 2, axis=1), vs, axis=0)  # h2v1_upsample / h2v2_upsample: replication
    if hs == 2 and vs == 1:                                    # h2v1_fancy_upsample
        out = _fancy_h2(p, 1, 2, 2)
        out[:, 0], out[:, -1] = p[:, 0], p[:, -1]
        return out
    if hs == 2 and vs == 2:                                    # h2v2_fancy_upsample
        above = np.concatenate([p[:1], p[:-1]], axis=0)
        below = np.concatenate([p[1:], p[-1:]], axis=0)
        rows = np.empty((p.shape[0] * 2, p.shape[1]), np.int64)
        rows[0::2], rows[1::2] = p * 3 + above, p * 3 + below
        out = _fancy_h2(rows, 8, 7, 4)
        out[:, 0], out[:, -1] = (rows[:, 0] * 4 + 8) >> 4, (rows[:, -1] * 4 + 7) >> 4
        return out
    if hs == 1 and vs == 2:                                    # h1v2_fancy_upsample (libjpeg-turbo)
        above = np.concatenate([p[:1], p[:-1]], axis=0)
        below = np.concatenate([p[1:], p[-1:]], axis=0)
        out = np.empty((p.shape[0] * 2, p.shape[1]), np.int64)
        out[0::2], out[1::2] = (p * 3 + above + 1) >> 2, (p * 3 + below + 2) >> 2
        return out
    raise NotImplementedError((hs, vs))


def ycc_to_rgb(y, cb, cr):
    """jdcolor.c build_ycc_rgb_table + ycc_rgb_convert."""
    xb, xr = cb - 128, cr - 128
    r = y + ((91881 * xr + 32768) >> 16)
    g = y + ((-22554 * xb + 32768 - 46802 * xr) >> 16)
    b = y + ((116130 * xb + 32768) >> 16)
    return np.clip(np.stack([r, g, b], -1), 0, 255).astype(np.uint8)


def decode_jpeg(data):
    """bytes of a baseline JFIF/JPEG file -> uint8 (H, W, 3) (or (H, W) for one component), as Pillow decodes it."""
    data = bytes(data)
    assert data[:2] == b"\xff\xd8", "not a JPEG"
    pos, qt, huff, restart_interval = 2, {}, {}, 0
    frame = None
    while True:
        assert data[pos] == 0xFF, "marker expected"
        marker = data[pos + 1]
        pos += 2
        if marker == 0xFF:
            pos -= 1
            continue
        length = int.from_bytes(data[pos:pos + 2], "big")
        seg = data[pos + 2:pos + length]
        if marker == 0xDB:
            k = 0
            while k < len(seg):
                assert seg[k] >> 4 == 0, "16-bit quantisation tables are not baseline"
                qt[seg[k] & 15] = list(seg[k + 1:k + 65])
                k += 65
        elif marker == 0xC4:
            k = 0
            while k < len(seg):
                bits = list(seg[k + 1:k + 17])
                n = sum(bits)
                vals = list(seg[k + 17:k + 17 + n])
                huff[seg[k]] = {(ln, code): sym for sym, (code, ln) in huffman_codes(bits, vals).items()}
                k += 17 + n
        elif marker == 0xC0 or marker == 0xC1:
            assert seg[0] == 8, "8-bit samples only"
            h, w = int.from_bytes(seg[1:3], "big"), int.from_bytes(seg[3:5], "big")
            comps = [(seg[6 + 3 * i], seg[7 + 3 * i] >> 4, seg[7 + 3 * i] & 15, seg[8 + 3 * i]) for i in range(seg[5])]
            frame = (h, w, comps)
        elif marker in (0xC2, 0xC3, 0xC5, 0xC6, 0xC7, 0xC9, 0xCA, 0xCB, 0xCD, 0xCE, 0xCF):
            raise NotImplementedError("only baseline sequential Huffman JPEG (SOF0/SOF1)")
        elif marker == 0xDD:
            restart_interval = int.from_bytes(seg[:2], "big")
        elif marker == 0xDA:
            ns = seg[0]
            scan = {seg[1 + 2 * i]: (seg[2 + 2 * i] >> 4, seg[2 + 2 * i] & 15) for i in range(ns)}
            pos += length
            break
        pos += length
    h, w, comps = frame
    assert len(scan) == len(comps), "one interleaved scan expected"
    hmax, vmax = max(c[1] for c in comps), max(c[2] for c in comps)
    mx, my = -(-w // (8 * hmax)), -(-h // (8 * vmax))
    coef = [np.zeros((my * c[2], mx * c[1], 64), np.int64) for c in comps]
    br, pred = _BitReader(data, pos), [0] * len(comps)
    for m in range(mx * my):
        if restart_interval and m and m % restart_interval == 0:
            br.restart()
            pred = [0] * len(comps)
        j, i = divmod(m, mx)
        for ci, (cid, hs, vs, _tq) in enumerate(comps):
            dc_tab, ac_tab = huff[scan[cid][0]], huff[0x10 | scan[cid][1]]
            for dy in range(vs):
                for dx in range(hs):
                    blk = coef[ci][j * vs + dy, i * hs + dx]
                    s = br.symbol(dc_tab)
                    pred[ci] += _extend(br.bits(s), s)
                    blk[0] = pred[ci]
                    k = 1
                    while k < 64:
                        rs = br.symbol(ac_tab)
                        r, s = rs >> 4, rs & 15
                        if s == 0:
                            if r != 15:
                                break
                            k += 16
                            continue
                        k += r
                        blk[ZIGZAG[k]] = _extend(br.bits(s), s)
                        k += 1
    planes = []
    for ci, (cid, hs, vs, tq) in enumerate(comps):
        px = idct_blocks(coef[ci], qt[tq])                                   # (by, bx, 8, 8)
        full = px.transpose(0, 2, 1, 3).reshape(px.shape[0] * 8, px.shape[1] * 8)
        ch, cw = -(-h * vs // vmax), -(-w * hs // hmax)                       # downsampled_height / _width
        fh, fv = hmax // hs, vmax // vs
        planes.append(upsample(full[:ch, :cw], fh, fv)[:h, :w])
    if len(planes) == 1:
        return planes[0].astype(np.uint8)
    return ycc_to_rgb(*planes)
