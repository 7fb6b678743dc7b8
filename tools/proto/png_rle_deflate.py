"""Prototype (pure Python) of the GPU PNG mask encoder's bit logic: fixed-Huffman deflate with distance-1 matches,
row-wise Adler-32 partials, chunked CRC-32 with x^n shifts.  Mirrors csrc/service.cu so the formulas can be checked
against zlib and OpenCV before any GPU time is spent."""
import struct
import zlib

import numpy as np

POLY = 0xEDB88320


def brev(v, n):
    return int(format(v, f"0{n}b")[::-1], 2)


class Bits:
    def __init__(self):
        self.acc, self.n = 0, 0

    def put(self, v, n):
        self.acc |= v << self.n
        self.n += n


def put_literal(bw, v):
    if v < 144:
        bw.put(brev(0x30 + v, 8), 8)
    else:
        bw.put(brev(0x190 + (v - 144), 9), 9)


def put_run_match(bw, length):
    d = length - 3
    extra_bits = extra = 0
    if length == 258:
        idx = 28
    elif d < 8:
        idx = d
    else:
        e = d.bit_length() - 1 - 2
        idx = 4 * (e + 1) + ((d - (4 << e)) >> e)
        extra_bits, extra = e, (d - (4 << e)) & ((1 << e) - 1)
    code = 257 + idx
    if code < 280:
        bw.put(brev(code - 256, 7), 7)
    else:
        bw.put(brev(0xC0 + (code - 280), 8), 8)
    if extra_bits:
        bw.put(extra, extra_bits)
    bw.put(0, 5)


def flush_run(bw, v, run):
    put_literal(bw, v)
    rem = run - 1
    while rem >= 3:
        n = min(rem, 258)
        put_run_match(bw, n)
        rem -= n
    for _ in range(rem):
        put_literal(bw, v)


def encode_row(row_bytes):
    """row_bytes: packed scan line without the filter byte."""
    bw = Bits()
    cur, run = 0, 1
    L = len(row_bytes) + 1
    a = s = 0
    for j, byte in enumerate(row_bytes):
        a += byte
        s += (L - (j + 1)) * byte
        if byte == cur:
            run += 1
        else:
            flush_run(bw, cur, run)
            cur, run = byte, 1
    flush_run(bw, cur, run)
    return bw.acc, bw.n, a % 65521, s % 65521


def multmodp(a, b):
    m, p = 1 << 31, 0
    while True:
        if a & m:
            p ^= b
            if (a & (m - 1)) == 0:
                break
        m >>= 1
        b = (b >> 1) ^ POLY if b & 1 else b >> 1
    return p


def x_pow_8n(n):
    """x^(8 n) mod P in zlib's reflected representation (x^0 = 1 << 31)."""
    p, sq = 1 << 31, 1 << 23          # sq = x^8
    while n:
        if n & 1:
            p = multmodp(sq, p)
        sq = multmodp(sq, sq)
        n >>= 1
    return p


def crc_table():
    t = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = POLY ^ (c >> 1) if c & 1 else c >> 1
        t.append(c)
    return t


TABLE = crc_table()


def crc_chunked(data, nchunks=7):
    n = len(data)
    per = (n + nchunks - 1) // nchunks
    total = 0
    for t in range(nchunks):
        lo, hi = min(t * per, n), min((t + 1) * per, n)
        r = 0xFFFFFFFF if t == 0 else 0
        for b in data[lo:hi]:
            r = TABLE[(r ^ b) & 0xFF] ^ (r >> 8)
        total ^= multmodp(x_pow_8n(n - hi), r)
    return total ^ 0xFFFFFFFF


def encode_png(mask):
    H, W = mask.shape
    packed = np.packbits(mask != 0, axis=1)
    L = packed.shape[1] + 1
    n_raw = H * L
    stream, nbits = 3, 3                       # BFINAL = 1, BTYPE = 01
    A = S = 0
    for r in range(H):
        acc, n, a, s = encode_row(packed[r].tolist())
        stream |= acc << nbits
        nbits += n
        A += a
        S += (s + ((n_raw - (r + 1) * L) % 65521) * a) % 65521
    nbits += 7                                  # end of block: seven zero bits
    deflate = stream.to_bytes((nbits + 7) // 8, "little")
    adler = (((n_raw + S) % 65521) << 16) | ((1 + A) % 65521)
    idat = b"\x78\x01" + deflate + struct.pack(">I", adler)
    ihdr = struct.pack(">IIBBBBB", W, H, 1, 0, 0, 0, 0)

    def chunk(kind, data):
        return struct.pack(">I", len(data)) + kind + data + struct.pack(">I", crc_chunked(kind + data))

    return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", ihdr) + chunk(b"IDAT", idat) + chunk(b"IEND", b""), idat


def main():
    import cv2
    rng = np.random.default_rng(0)
    cases = []
    yy, xx = np.mgrid[0:480, 0:640]
    cases.append(((np.sin(yy / 37.0) * 80 + 240 - np.abs(xx - 320) * 0.3) > yy).astype(np.uint8))   # band-like
    cases.append(np.zeros((480, 640), np.uint8))
    cases.append(np.ones((480, 640), np.uint8))
    cases.append((rng.random((97, 131)) > 0.5).astype(np.uint8))                                    # noise, odd width
    cases.append((rng.random((33, 2600)) > 0.999).astype(np.uint8))                                 # very long runs
    cases.append(np.ones((3, 5), np.uint8))
    for m in cases:
        png, idat = encode_png(m)
        raw = zlib.decompress(idat)                                      # checks the Adler-32 too
        H, W = m.shape
        assert len(raw) == H * ((W + 7) // 8 + 1)
        assert zlib.crc32(b"IDAT" + idat) == crc_chunked(b"IDAT" + idat)
        dec = cv2.imdecode(np.frombuffer(png, np.uint8), cv2.IMREAD_GRAYSCALE)
        assert dec is not None and np.array_equal(dec, m * 255), m.shape
        ref = cv2.imencode(".png", m * 255)[1]
        print(m.shape, "png bytes", len(png), "cv2", len(ref))


if __name__ == "__main__":
    main()
