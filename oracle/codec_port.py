"""CPU restatement of the two image decoders on the reference's wire (test infrastructure only - product code never
imports this; the GPU kernels in csrc/codec.cu are checked against it and against OpenCV itself).

The reference server receives every frame as a JPEG colour image and a 16-bit PNG depth image and decodes both with
OpenCV (`services/vision_analysis/server.py:117-118`); the client produces them with `cv2.imencode('.jpg', ...)` and
`cv2.imencode('.png', ...)` (`services/vision_analysis/client.py:65-67`).  The arithmetic lives in third-party
libraries that are not under /root/reference: libjpeg-turbo (3.1.2 in this image's OpenCV 4.13 build) and zlib/libpng
(1.2.11 / 1.6.53).  What is restated here is their published algorithm for exactly the streams OpenCV writes:

* baseline sequential JPEG (SOF0), 8 bit, 1 or 3 components, Huffman coded, chroma 1x1 / 2x1 / 2x2 subsampled:
  Huffman decoding (ITU T.81 F.2.2), the accurate integer inverse DCT (`jidctint.c: jpeg_idct_islow`), "fancy"
  triangle-filter chroma up-sampling (`jdsample.c: h2v1_fancy_upsample, h2v2_fancy_upsample`) and the fixed-point
  YCbCr -> RGB conversion (`jdcolor.c`), written out as BGR like `cv2.imdecode(..., IMREAD_COLOR)`;
* PNG: chunk walk, zlib/deflate inflate (RFC 1950/1951: stored, fixed and dynamic Huffman blocks), the five scan-line
  filters (PNG spec 9), 8/16-bit grey, big-endian samples, as `cv2.imdecode(..., IMREAD_UNCHANGED)` returns them.

Pin: every function is compared with OpenCV run live on the same bytes (tests/test_oracle_codec.py), bit for bit.
"""
from __future__ import annotations

import struct

import numpy as np

# ------------------------------------------------------------------------------------------------------------
# JPEG
# ------------------------------------------------------------------------------------------------------------
ZIGZAG = np.array([0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14,
                   21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53,
                   60, 61, 54, 47, 55, 62, 63], dtype=np.int64)


class JpegHeader:
    def __init__(self):
        self.width = self.height = 0
        self.comps = []          # dicts: id, h, v, tq, td, ta
        self.qt = {}             # table id -> (64,) int32 in natural order
        self.huff = {}           # (class, id) -> (counts[16], symbols)
        self.restart_interval = 0
        self.scan_offset = 0     # first byte of entropy-coded data


def parse_jpeg(data: bytes) -> JpegHeader:
    """Marker segments up to the first scan (ITU T.81 B.2).  Anything but baseline SOF0 raises ValueError."""
    h = JpegHeader()
    if data[:2] != b"\xff\xd8":
        raise ValueError("not a JPEG (no SOI)")
    i = 2
    while i + 4 <= len(data):
        if data[i] != 0xFF:
            raise ValueError(f"marker expected at byte {i}")
        m = data[i + 1]
        if m == 0xFF:           # fill byte
            i += 1
            continue
        L = (data[i + 2] << 8) | data[i + 3]
        seg = data[i + 4:i + 2 + L]
        if m == 0xDB:
            j = 0
            while j < len(seg):
                pq, tq = seg[j] >> 4, seg[j] & 15
                if pq:
                    raise ValueError("16-bit quantisation tables are not baseline")
                t = np.zeros(64, dtype=np.int32)
                t[ZIGZAG] = np.frombuffer(seg[j + 1:j + 65], dtype=np.uint8)
                h.qt[tq] = t
                j += 65
        elif m == 0xC0:
            if seg[0] != 8:
                raise ValueError("only 8-bit samples")
            h.height, h.width = (seg[1] << 8) | seg[2], (seg[3] << 8) | seg[4]
            for c in range(seg[5]):
                cid, hv, tq = seg[6 + 3 * c:9 + 3 * c]
                h.comps.append({"id": cid, "h": hv >> 4, "v": hv & 15, "tq": tq})
        elif m in (0xC1, 0xC2, 0xC3, 0xC5, 0xC6, 0xC7, 0xC9, 0xCA, 0xCB, 0xCD, 0xCE, 0xCF):
            raise ValueError(f"unsupported JPEG process (SOF marker {m:#x}); only baseline SOF0")
        elif m == 0xC4:
            j = 0
            while j < len(seg):
                tc, th = seg[j] >> 4, seg[j] & 15
                counts = list(seg[j + 1:j + 17])
                n = sum(counts)
                h.huff[(tc, th)] = (counts, list(seg[j + 17:j + 17 + n]))
                j += 17 + n
        elif m == 0xDD:
            h.restart_interval = (seg[0] << 8) | seg[1]
        elif m == 0xDA:
            ns = seg[0]
            if ns != len(h.comps):
                raise ValueError("only one interleaved scan is supported")
            for c in range(ns):
                cs, tt = seg[1 + 2 * c], seg[2 + 2 * c]
                comp = next(k for k in h.comps if k["id"] == cs)
                comp["td"], comp["ta"] = tt >> 4, tt & 15
            h.scan_offset = i + 2 + L
            return h
        i += 2 + L
    raise ValueError("no scan found")


def _huff_tables(counts, symbols):
    """code -> symbol maps per length (T.81 C.2): returns (mincode[17], maxcode[17], valptr[17], symbols)."""
    mincode, maxcode, valptr = [0] * 17, [-1] * 17, [0] * 17
    code, k = 0, 0
    for length in range(1, 17):
        valptr[length] = k
        mincode[length] = code
        code += counts[length - 1]
        k += counts[length - 1]
        maxcode[length] = code - 1 if counts[length - 1] else -1
        code <<= 1
    return mincode, maxcode, valptr, symbols


class _Bits:
    def __init__(self, data: bytes, pos: int):
        self.d, self.p, self.acc, self.n = data, pos, 0, 0

    def _fill(self):
        while self.n <= 24:
            if self.p >= len(self.d):
                b = 0
            else:
                b = self.d[self.p]
                if b == 0xFF:
                    nxt = self.d[self.p + 1] if self.p + 1 < len(self.d) else 0
                    if nxt == 0:
                        self.p += 2
                    else:       # a marker: feed zeros, stay on it (T.81 F.2.2.5)
                        b = 0
                else:
                    self.p += 1
            self.acc = ((self.acc << 8) | b) & 0xFFFFFFFFFF
            self.n += 8

    def get(self, k: int) -> int:
        if k == 0:
            return 0
        if self.n < k:
            self._fill()
        self.n -= k
        return (self.acc >> self.n) & ((1 << k) - 1)

    def restart(self):
        """Skip to just behind the next RSTn marker and drop buffered bits."""
        self.acc = self.n = 0
        while self.p + 1 < len(self.d) and not (self.d[self.p] == 0xFF and 0xD0 <= self.d[self.p + 1] <= 0xD7):
            self.p += 1
        self.p += 2


def _decode_symbol(bits: _Bits, tab) -> int:
    mincode, maxcode, valptr, symbols = tab
    code = 0
    for length in range(1, 17):
        code = (code << 1) | bits.get(1)
        if maxcode[length] >= 0 and code <= maxcode[length] and code >= mincode[length]:
            return symbols[valptr[length] + code - mincode[length]]
    raise ValueError("bad Huffman code")


def _extend(v: int, s: int) -> int:
    return v if v >= (1 << (s - 1)) else v - (1 << s) + 1


def decode_coefficients(data: bytes, h: JpegHeader):
    """Entropy decoding: per component an int16 array (blocks_y, blocks_x, 64) of quantised coefficients in natural order
    (padded to whole MCUs), T.81 F.2.2."""
    hmax, vmax = max(c["h"] for c in h.comps), max(c["v"] for c in h.comps)
    mcux, mcuy = -(-h.width // (8 * hmax)), -(-h.height // (8 * vmax))
    coefs = [np.zeros((mcuy * c["v"], mcux * c["h"], 64), dtype=np.int16) for c in h.comps]
    dct = {k: _huff_tables(*v) for k, v in h.huff.items()}
    bits = _Bits(data, h.scan_offset)
    pred = [0] * len(h.comps)
    count = 0
    for my in range(mcuy):
        for mx in range(mcux):
            if h.restart_interval and count and count % h.restart_interval == 0:
                bits.restart()
                pred = [0] * len(h.comps)
            count += 1
            for ci, c in enumerate(h.comps):
                dc_tab, ac_tab = dct[(0, c["td"])], dct[(1, c["ta"])]
                for by in range(c["v"]):
                    for bx in range(c["h"]):
                        blk = coefs[ci][my * c["v"] + by, mx * c["h"] + bx]
                        s = _decode_symbol(bits, dc_tab)
                        pred[ci] += _extend(bits.get(s), s) if s else 0
                        blk[0] = pred[ci]
                        k = 1
                        while k < 64:
                            rs = _decode_symbol(bits, ac_tab)
                            r, s = rs >> 4, rs & 15
                            if s == 0:
                                if r != 15:
                                    break
                                k += 16
                                continue
                            k += r
                            blk[ZIGZAG[k]] = _extend(bits.get(s), s)
                            k += 1
    return coefs


# jidctint.c (CONST_BITS = 13, PASS1_BITS = 2)
_F = dict(f0_298=2446, f0_390=3196, f0_541=4433, f0_765=6270, f0_899=7373, f1_175=9633, f1_501=12299, f1_847=15137,
          f1_961=16069, f2_053=16819, f2_562=20995, f3_072=25172)


def _idct_1d(x, shift_out, first_pass):
    """One pass of jpeg_idct_islow over the LAST axis of x (.., 8) int64; returns the descaled result."""
    x0, x1, x2, x3, x4, x5, x6, x7 = (x[..., i] for i in range(8))
    z1 = (x2 + x6) * _F["f0_541"]
    tmp2 = z1 + x6 * (-_F["f1_847"])
    tmp3 = z1 + x2 * _F["f0_765"]
    tmp0 = (x0 + x4) << 13
    tmp1 = (x0 - x4) << 13
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = x7, x5, x3, x1
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * _F["f1_175"]
    t0 = t0 * _F["f0_298"]
    t1 = t1 * _F["f2_053"]
    t2 = t2 * _F["f3_072"]
    t3 = t3 * _F["f1_501"]
    z1 = z1 * (-_F["f0_899"])
    z2 = z2 * (-_F["f2_562"])
    z3 = z3 * (-_F["f1_961"]) + z5
    z4 = z4 * (-_F["f0_390"]) + z5
    t0 = t0 + z1 + z3
    t1 = t1 + z2 + z4
    t2 = t2 + z2 + z3
    t3 = t3 + z1 + z4
    out = np.stack([tmp10 + t3, tmp11 + t2, tmp12 + t1, tmp13 + t0, tmp13 - t0, tmp12 - t1, tmp11 - t2, tmp10 - t3], axis=-1)
    return (out + (1 << (shift_out - 1))) >> shift_out


def _range_limit_idct(v):
    """`range_limit[v & RANGE_MASK]` of jdmaster.c's table: clamp(v + 128) with the table's own wrap-around."""
    i = v & 1023
    out = np.where(i < 128, i + 128, np.where(i < 512, 255, np.where(i < 896, 0, i - 896)))
    return out.astype(np.uint8)


def idct_islow(coefs: np.ndarray, qt: np.ndarray) -> np.ndarray:
    """(by, bx, 64) quantised coefficients -> (by*8, bx*8) uint8 samples (dequantisation included)."""
    by, bx, _ = coefs.shape
    blk = (coefs.astype(np.int64) * qt.astype(np.int64)).reshape(by, bx, 8, 8)     # [row][col]
    ws = _idct_1d(blk.transpose(0, 1, 3, 2), 13 - 2, True).transpose(0, 1, 3, 2)   # pass 1 works down the columns
    out = _idct_1d(ws, 13 + 2 + 3, False)                                           # pass 2 along the rows
    px = _range_limit_idct(out)
    return px.transpose(0, 2, 1, 3).reshape(by * 8, bx * 8)


def upsample_h2v1_fancy(c: np.ndarray) -> np.ndarray:
    """jdsample.c h2v1_fancy_upsample: (rows, w) -> (rows, 2w); 3/4 nearer + 1/4 further, rounding 1 / 2 alternately."""
    c = c.astype(np.int32)
    rows, w = c.shape
    out = np.empty((rows, 2 * w), dtype=np.int32)
    if w == 1:
        out[:, 0] = out[:, 1] = c[:, 0]
        return out.astype(np.uint8)
    left = np.concatenate([c[:, :1], c[:, :-1]], axis=1)
    right = np.concatenate([c[:, 1:], c[:, -1:]], axis=1)
    out[:, 0::2] = (c * 3 + left + 1) >> 2
    out[:, 1::2] = (c * 3 + right + 2) >> 2
    out[:, 0] = c[:, 0]
    out[:, -1] = c[:, -1]
    return out.astype(np.uint8)


def upsample_h2v2_fancy(c: np.ndarray) -> np.ndarray:
    """jdsample.c h2v2_fancy_upsample on the real (rows, w) samples of a component -> (2 rows, 2 w).  Rows above the first
    and below the last are the first / last row again (jdmainct.c context rows)."""
    c = c.astype(np.int32)
    rows, w = c.shape
    above = np.concatenate([c[:1], c[:-1]], axis=0)
    below = np.concatenate([c[1:], c[-1:]], axis=0)
    out = np.empty((2 * rows, 2 * w), dtype=np.int32)
    for v, other in ((0, above), (1, below)):
        colsum = c * 3 + other                                         # this column's vertical blend (x4)
        last = np.concatenate([colsum[:, :1], colsum[:, :-1]], axis=1)
        nxt = np.concatenate([colsum[:, 1:], colsum[:, -1:]], axis=1)
        even = (colsum * 3 + last + 8) >> 4
        odd = (colsum * 3 + nxt + 7) >> 4
        even[:, 0] = (colsum[:, 0] * 4 + 8) >> 4
        odd[:, -1] = (colsum[:, -1] * 4 + 7) >> 4
        out[v::2, 0::2] = even
        out[v::2, 1::2] = odd
    return out.astype(np.uint8)


def ycc_to_bgr(y: np.ndarray, cb: np.ndarray, cr: np.ndarray) -> np.ndarray:
    """jdcolor.c build_ycc_rgb_table + ycc_rgb_convert (SCALEBITS = 16), written out in B, G, R order."""
    i = np.arange(256, dtype=np.int64) - 128
    fix = lambda f: int(f * 65536 + 0.5)
    cr_r = (fix(1.40200) * i + 32768) >> 16
    cb_b = (fix(1.77200) * i + 32768) >> 16
    cr_g = -fix(0.71414) * i
    cb_g = -fix(0.34414) * i + 32768
    yy = y.astype(np.int64)
    r = np.clip(yy + cr_r[cr], 0, 255)
    g = np.clip(yy + ((cb_g[cb] + cr_g[cr]) >> 16), 0, 255)
    b = np.clip(yy + cb_b[cb], 0, 255)
    return np.stack([b, g, r], axis=-1).astype(np.uint8)


def decode_jpeg_bgr(data: bytes) -> np.ndarray:
    """What `cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)` returns for a baseline JPEG (`server.py:117`)."""
    h = parse_jpeg(data)
    coefs = decode_coefficients(data, h)
    hmax, vmax = max(c["h"] for c in h.comps), max(c["v"] for c in h.comps)
    planes = []
    for c, cf in zip(h.comps, coefs):
        full = idct_islow(cf, h.qt[c["tq"]])
        dw, dh = -(-h.width * c["h"] // hmax), -(-h.height * c["v"] // vmax)     # downsampled_width / _height
        real = full[:dh, :dw]
        if (c["h"], c["v"]) == (hmax, vmax):
            up = real
        elif (c["h"] * 2, c["v"] * 2) == (hmax, vmax):
            up = upsample_h2v2_fancy(real)
        elif (c["h"] * 2, c["v"]) == (hmax, vmax):
            up = upsample_h2v1_fancy(real)
        else:
            raise ValueError(f"unsupported sampling factors {c['h']}x{c['v']} of {hmax}x{vmax}")
        planes.append(up[:h.height, :h.width])
    if len(planes) == 1:
        return np.repeat(planes[0][:, :, None], 3, axis=2)
    return ycc_to_bgr(*planes)


# ------------------------------------------------------------------------------------------------------------
# PNG
# ------------------------------------------------------------------------------------------------------------
_LEN_BASE = [3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258]
_LEN_EXTRA = [0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0]
_DIST_BASE = [1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145,
              8193, 12289, 16385, 24577]
_DIST_EXTRA = [0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13]
_CL_ORDER = [16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15]


class _LsbBits:
    def __init__(self, data: bytes):
        self.d, self.p, self.acc, self.n = data, 0, 0, 0

    def get(self, k: int) -> int:
        while self.n < k:
            self.acc |= (self.d[self.p] if self.p < len(self.d) else 0) << self.n
            self.p += 1
            self.n += 8
        v = self.acc & ((1 << k) - 1)
        self.acc >>= k
        self.n -= k
        return v

    def align(self):
        self.acc >>= self.n & 7
        self.n -= self.n & 7


def _canonical(lengths):
    """RFC 1951 3.2.2: (length, code) -> symbol for a list of code lengths."""
    maxlen = max(lengths) if lengths else 0
    bl_count = [0] * (maxlen + 2)
    for L in lengths:
        bl_count[L] += 1
    bl_count[0] = 0
    code, next_code = 0, [0] * (maxlen + 2)
    for bits in range(1, maxlen + 1):
        code = (code + bl_count[bits - 1]) << 1
        next_code[bits] = code
    table = {}
    for sym, L in enumerate(lengths):
        if L:
            table[(L, next_code[L])] = sym
            next_code[L] += 1
    return table


def _read_sym(bits: _LsbBits, table) -> int:
    code = 0
    for L in range(1, 16):
        code = (code << 1) | bits.get(1)          # Huffman codes are packed most significant bit first
        s = table.get((L, code))
        if s is not None:
            return s
    raise ValueError("bad deflate code")


def inflate(z: bytes) -> bytes:
    """zlib stream (RFC 1950) -> bytes; the Adler-32 trailer is checked."""
    if (z[0] & 15) != 8 or ((z[0] << 8) | z[1]) % 31 or (z[1] & 0x20):
        raise ValueError("not a zlib/deflate stream")
    bits = _LsbBits(z[2:])
    out = bytearray()
    fixed_lit = _canonical([8] * 144 + [9] * 112 + [7] * 24 + [8] * 8)
    fixed_dist = _canonical([5] * 30)
    while True:
        final, btype = bits.get(1), bits.get(2)
        if btype == 0:
            bits.align()
            n, nn = bits.get(16), bits.get(16)
            if n ^ nn != 0xFFFF:
                raise ValueError("stored block length mismatch")
            for _ in range(n):
                out.append(bits.get(8))
        elif btype in (1, 2):
            if btype == 1:
                lit, dist = fixed_lit, fixed_dist
            else:
                hlit, hdist, hclen = bits.get(5) + 257, bits.get(5) + 1, bits.get(4) + 4
                cl = [0] * 19
                for i in range(hclen):
                    cl[_CL_ORDER[i]] = bits.get(3)
                clt = _canonical(cl)
                lens = []
                while len(lens) < hlit + hdist:
                    s = _read_sym(bits, clt)
                    if s < 16:
                        lens.append(s)
                    elif s == 16:
                        lens += [lens[-1]] * (3 + bits.get(2))
                    elif s == 17:
                        lens += [0] * (3 + bits.get(3))
                    else:
                        lens += [0] * (11 + bits.get(7))
                lit, dist = _canonical(lens[:hlit]), _canonical(lens[hlit:hlit + hdist])
            while True:
                s = _read_sym(bits, lit)
                if s < 256:
                    out.append(s)
                elif s == 256:
                    break
                else:
                    length = _LEN_BASE[s - 257] + bits.get(_LEN_EXTRA[s - 257])
                    d = _read_sym(bits, dist)
                    distance = _DIST_BASE[d] + bits.get(_DIST_EXTRA[d])
                    for _ in range(length):
                        out.append(out[-distance])
        else:
            raise ValueError("reserved deflate block type")
        if final:
            break
    bits.align()
    a, b = 1, 0
    for chunk in (out[i:i + 5552] for i in range(0, len(out), 5552)):
        for v in chunk:
            a += v
            b += a
        a %= 65521
        b %= 65521
    tail = bits.d[bits.p - bits.n // 8:bits.p - bits.n // 8 + 4]
    if len(tail) == 4 and struct.unpack(">I", bytes(tail))[0] != ((b << 16) | a):
        raise ValueError("Adler-32 mismatch")
    return bytes(out)


def png_chunks(data: bytes):
    if data[:8] != b"\x89PNG\r\n\x1a\n":
        raise ValueError("not a PNG")
    i = 8
    while i + 12 <= len(data):
        L, = struct.unpack(">I", data[i:i + 4])
        yield data[i + 4:i + 8], data[i + 8:i + 8 + L]
        i += 12 + L


def unfilter(raw: bytes, rows: int, stride: int, bpp: int) -> np.ndarray:
    """PNG filters 0-4 (None, Sub, Up, Average, Paeth) over `rows` scan lines of `stride` bytes, `bpp` bytes per pixel."""
    out = np.zeros((rows, stride), dtype=np.uint8)
    prev = np.zeros(stride, dtype=np.int32)
    pos = 0
    for r in range(rows):
        ft = raw[pos]
        line = np.frombuffer(raw[pos + 1:pos + 1 + stride], dtype=np.uint8).astype(np.int32)
        pos += 1 + stride
        cur = np.zeros(stride, dtype=np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        elif ft == 1:                                   # Sub: a running sum per byte lane, modulo 256
            cur = line.copy()
            for lane in range(bpp):
                cur[lane::bpp] = np.cumsum(line[lane::bpp]) & 255
        else:
            for i in range(stride):
                a = cur[i - bpp] if i >= bpp else 0
                b = prev[i]
                c = prev[i - bpp] if i >= bpp else 0
                if ft == 3:
                    cur[i] = (line[i] + ((a + b) >> 1)) & 255
                elif ft == 4:
                    p = a + b - c
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
                    pr = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                    cur[i] = (line[i] + pr) & 255
                else:
                    raise ValueError(f"bad PNG filter type {ft}")
        out[r] = cur
        prev = cur
    return out


def decode_png_gray(data: bytes) -> np.ndarray:
    """What `cv2.imdecode(..., cv2.IMREAD_UNCHANGED)` returns for an 8- or 16-bit greyscale PNG (`server.py:118`)."""
    ihdr, idat = None, b""
    for t, body in png_chunks(data):
        if t == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", body)
        elif t == b"IDAT":
            idat += body
    w, h, depth, ctype, _, _, interlace = ihdr
    if ctype != 0 or depth not in (8, 16) or interlace:
        raise ValueError("only non-interlaced 8/16-bit greyscale PNGs (the depth images of the reference client)")
    bpp = depth // 8
    px = unfilter(inflate(idat), h, w * bpp, bpp)
    if depth == 8:
        return px
    return (px[:, 0::2].astype(np.uint16) << 8) | px[:, 1::2]
