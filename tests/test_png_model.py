"""The pure-Python model of the GPU PNG encoder's bit logic (tools/proto/png_rle_deflate.py mirrors csrc/service.cu):
fixed-Huffman run-length deflate, row-wise Adler-32 partials, sliced CRC-32 - against zlib and OpenCV, no GPU needed."""
import os
import struct
import sys
import zlib

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "proto"))
import png_rle_deflate as model  # noqa: E402

from oracle import service_port as sp  # noqa: E402


def _masks():
    rng = np.random.default_rng(0)
    yy, xx = np.mgrid[0:120, 0:160]
    yield ((np.sin(xx / 25.0) * 15 + 50 < yy) & (yy < np.sin(xx / 25.0) * 15 + 85)).astype(np.uint8)
    yield np.zeros((64, 96), np.uint8)
    yield np.ones((64, 96), np.uint8)
    yield (rng.random((97, 131)) > 0.5).astype(np.uint8)
    yield (rng.random((9, 2600)) > 0.999).astype(np.uint8)        # runs longer than one 258-byte match
    yield np.ones((3, 5), np.uint8)


@pytest.mark.parametrize("mask", list(_masks()), ids=lambda m: "x".join(map(str, m.shape)))
def test_model_png_is_valid_and_decodes_to_the_mask(mask):
    png, idat = model.encode_png(mask)
    H, W = mask.shape
    assert len(zlib.decompress(idat)) == H * ((W + 7) // 8 + 1)                 # inflates; zlib checks the Adler-32
    assert zlib.crc32(b"IDAT" + idat) == model.crc_chunked(b"IDAT" + idat, nchunks=256)
    assert struct.unpack(">I", png[-16:-12])[0] == zlib.crc32(b"IDAT" + idat)      # IDAT CRC sits before the 12-byte IEND
    assert np.array_equal(sp.decode_mask_png(png), mask * 255)                  # what client.py:112 would see
    assert np.array_equal(sp.decode_mask_png(sp.encode_mask_png(mask)), mask * 255)


def test_length_codes_cover_every_match_length():
    base = [3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258]
    extra = [0] * 8 + [1] * 4 + [2] * 4 + [3] * 4 + [4] * 4 + [5] * 4 + [0]
    for length in range(3, 259):
        bw = model.Bits()
        model.put_run_match(bw, length)
        idx = max(i for i, b in enumerate(base) if b <= length) if length < 258 else 28
        nbits = (7 if 257 + idx < 280 else 8) + extra[idx] + 5
        assert bw.n == nbits, length
        code_bits = 7 if 257 + idx < 280 else 8
        assert (bw.acc >> code_bits) & ((1 << extra[idx]) - 1) == length - base[idx], length
