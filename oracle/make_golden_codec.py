"""Generates tests/golden/codec_cases.npz: small JPEG / PNG files as the reference client writes them
(`cv2.imencode`, `services/vision_analysis/client.py:65-67`) plus variants that exercise every branch of the decoders
(chroma sampling 4:2:0 / 4:2:2 / 4:4:4, grey, restart intervals, odd sizes; zlib stored / fixed / dynamic blocks, all
five PNG filters, 8-bit), each with what `cv2.imdecode` (`services/vision_analysis/server.py:117-118`) returns for it.

    python oracle/make_golden_codec.py          (needs OpenCV; run in the authoring container)

Test infrastructure only.
"""
import json
import os
import struct
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "robotic-discovery-platform_b200"))
OUT = os.path.join(ROOT, "tests", "golden")


def png_with_filters(img16: np.ndarray, level: int = 6, strategy: int = zlib.Z_DEFAULT_STRATEGY) -> bytes:
    """A 16-bit greyscale PNG whose rows cycle through filter types 0..4 (PNG spec 9), general LZ77 matches allowed."""
    h, w = img16.shape
    rows = img16.astype(">u2").view(np.uint8).reshape(h, 2 * w).astype(np.int32)
    raw = bytearray()
    prev = np.zeros(2 * w, dtype=np.int32)
    for r in range(h):
        ft, cur = r % 5, rows[r]
        a = np.concatenate([np.zeros(2, np.int32), cur[:-2]])
        c = np.concatenate([np.zeros(2, np.int32), prev[:-2]])
        if ft == 0:
            f = cur
        elif ft == 1:
            f = cur - a
        elif ft == 2:
            f = cur - prev
        elif ft == 3:
            f = cur - ((a + prev) >> 1)
        else:
            p = a + prev - c
            pa, pb, pc = np.abs(p - a), np.abs(p - prev), np.abs(p - c)
            f = cur - np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, prev, c))
        raw.append(ft)
        raw += bytes((f & 255).astype(np.uint8))
        prev = cur
    co = zlib.compressobj(level, zlib.DEFLATED, 15, 8, strategy)
    z = co.compress(bytes(raw)) + co.flush()

    def chunk(t, body):
        return struct.pack(">I", len(body)) + t + body + struct.pack(">I", zlib.crc32(t + body))
    # two IDAT chunks, split at an odd place, so that the bit reader has to cross a chunk boundary
    cut = len(z) // 3 + 1
    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 16, 0, 0, 0, 0)) + chunk(b"tEXt", b"k\0v")
            + chunk(b"IDAT", z[:cut]) + chunk(b"IDAT", z[cut:]) + chunk(b"IEND", b""))


def main():
    import cv2
    from rdp_b200 import synth
    rec, manifest = {}, {"opencv": cv2.__version__, "cases": {}}

    def add(name, payload, kind):
        flag = cv2.IMREAD_COLOR if kind == "jpeg" else cv2.IMREAD_UNCHANGED
        dec = cv2.imdecode(np.frombuffer(payload, np.uint8), flag)
        assert dec is not None, name
        rec[name + "/bytes"] = np.frombuffer(payload, np.uint8)
        rec[name + "/decoded"] = dec
        manifest["cases"][name] = dict(kind=kind, shape=list(dec.shape), dtype=str(dec.dtype), nbytes=len(payload))
        print(name, manifest["cases"][name])

    def frame(seed, H, W):
        f = synth.make_frame(seed, H, W, 0.25, "u16")
        return np.ascontiguousarray(f["rgb"][:, :, ::-1]), f["depth"]

    def jpg(img, *params):
        ok, buf = cv2.imencode(".jpg", img, list(params))
        assert ok
        return buf.tobytes()

    def png(img, *params):
        ok, buf = cv2.imencode(".png", img, list(params))
        assert ok
        return buf.tobytes()

    bgr, depth = frame(50, 96, 128)
    add("jpg_client_96x128", jpg(bgr), "jpeg")                       # exactly what client.py:65 writes: q95, 4:2:0
    add("png_client_96x128", png(depth), "png")                      # client.py:66: 16-bit, zlib level 1, Z_RLE, filter Sub
    bgr, depth = frame(51, 37, 53)
    add("jpg_odd_37x53", jpg(bgr), "jpeg")
    add("png_odd_37x53", png(depth), "png")
    bgr, depth = frame(52, 40, 56)
    add("jpg_444_40x56", jpg(bgr, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444), "jpeg")
    add("jpg_422_40x57", jpg(bgr[:, :, :][:, :56], cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422), "jpeg")
    add("jpg_422_odd_41x55", jpg(frame(53, 41, 55)[0], cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422), "jpeg")
    add("jpg_q50_rst_64x80", jpg(frame(54, 64, 80)[0], cv2.IMWRITE_JPEG_QUALITY, 50, cv2.IMWRITE_JPEG_RST_INTERVAL, 3), "jpeg")
    add("jpg_q100_48x64", jpg(frame(55, 48, 64)[0], cv2.IMWRITE_JPEG_QUALITY, 100), "jpeg")
    add("jpg_grey_48x64", jpg(cv2.cvtColor(frame(56, 48, 64)[0], cv2.COLOR_BGR2GRAY)), "jpeg")
    add("jpg_tiny_8x8", jpg(frame(57, 16, 16)[0][:8, :8]), "jpeg")
    add("jpg_1x1", jpg(frame(57, 16, 16)[0][:1, :1]), "jpeg")
    d = frame(58, 48, 64)[1]
    add("png_stored_48x64", png(d, cv2.IMWRITE_PNG_COMPRESSION, 0), "png")
    add("png_level9_48x64", png(d, cv2.IMWRITE_PNG_COMPRESSION, 9), "png")
    add("png_8bit_48x64", png((d >> 2).astype(np.uint8)), "png")
    add("png_fixed_huffman_4x4", png(d[:4, :4]), "png")
    add("png_all_filters_45x61", png_with_filters(frame(59, 45, 61)[1]), "png")
    add("png_all_filters_fixed_45x61", png_with_filters(frame(59, 45, 61)[1], 6, zlib.Z_FIXED), "png")
    add("png_long_matches_64x64", png_with_filters(np.tile(frame(60, 16, 64)[1], (4, 1)), 9), "png")
    np.savez_compressed(os.path.join(OUT, "codec_cases.npz"), **rec)
    with open(os.path.join(OUT, "codec_manifest.json"), "w") as fh:
        json.dump(manifest, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
