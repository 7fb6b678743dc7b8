"""`k_png_inflate` (csrc/codec.cu) compiled as host code - its text taken verbatim from the .cu file, with one-thread
stand-ins for the CUDA built-ins - against zlib on the golden PNG files, so the bit reader, the chunk walk and the
Huffman table construction are covered without a GPU (the same idea as tests/test_cr_host.py)."""
import json
import os
import shutil
import subprocess
import zlib

import numpy as np
import pytest

from oracle import codec_port as cp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

PRE = r'''
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
static inline unsigned __brev(unsigned v) { unsigned r = 0; for (int i = 0; i < 32; i++) { r = (r << 1) | (v & 1); v >>= 1; } return r; }
struct PngParams { const uint8_t* data; const int64_t* offsets; int B, H, W; uint8_t* raw; size_t raw_stride; int* geom; uint16_t* out; int32_t* status; };
'''
POST = r'''
int main(int argc, char** argv) {
  FILE* f = fopen(argv[1], "rb"); std::vector<uint8_t> d; int c; while ((c = fgetc(f)) != EOF) d.push_back((uint8_t)c); fclose(f);
  int H = atoi(argv[2]), W = atoi(argv[3]);
  int64_t offs[2] = {0, (int64_t)d.size()}; d.resize(d.size() + 16);
  size_t stride = (size_t)H * (1 + 2 * (size_t)W) + 256; std::vector<uint8_t> raw(stride); int geom[8] = {0}; int32_t st = -9;
  PngParams p{d.data(), offs, 1, H, W, raw.data(), stride, geom, nullptr, &st};
  k_png_inflate(p, 0);
  printf("%d %d\n", st, geom[0]);
  if (st == 0) { FILE* o = fopen(argv[4], "wb"); fwrite(raw.data(), 1, (size_t)H * (1 + (size_t)W * geom[0]), o); fclose(o); }
  return 0; }
'''


@pytest.fixture(scope="module")
def host_inflate(tmp_path_factory):
    if not shutil.which("g++"):
        pytest.skip("no host compiler")
    src = open(os.path.join(ROOT, "robotic-discovery-platform_b200", "csrc", "codec.cu")).read()
    body = src[src.index("constexpr int PNG_LOOK = 10;"):src.index("__device__ __forceinline__ int paeth(")]
    body = body.replace("__global__ void __launch_bounds__(32) k_png_inflate(const PngParams p) {", "void k_png_inflate(const PngParams p, int bidx) {")
    body = body.replace("__shared__ ", "static ").replace("__device__ ", "").replace("__forceinline__ ", "")
    body = body.replace("const int b = blockIdx.x;", "const int b = bidx;").replace("if (threadIdx.x != 0) return;", "")
    d = tmp_path_factory.mktemp("host_inflate")
    (d / "t.cpp").write_text(PRE + body + POST)
    subprocess.run(["g++", "-O1", "-w", "-o", str(d / "t"), str(d / "t.cpp")], check=True)
    return d


def _run(d, payload, H, W):
    (d / "in.png").write_bytes(payload)
    out = d / "out.bin"
    if out.exists():
        out.unlink()
    r = subprocess.run([str(d / "t"), str(d / "in.png"), str(H), str(W), str(out)], capture_output=True, text=True, check=True)
    status, bps = (int(x) for x in r.stdout.split())
    return status, bps, out.read_bytes() if out.exists() else None


def test_inflate_kernel_text_against_zlib(host_inflate):
    meta = json.load(open(os.path.join(GOLDEN, "codec_manifest.json")))["cases"]
    with np.load(os.path.join(GOLDEN, "codec_cases.npz")) as z:
        data = {k: z[k] for k in z.files}
    n = 0
    for name, v in meta.items():
        if v["kind"] != "png":
            continue
        payload = data[name + "/bytes"].tobytes()
        H, W = v["shape"]
        status, bps, got = _run(host_inflate, payload, H, W)
        want = zlib.decompress(b"".join(b for t, b in cp.png_chunks(payload) if t == b"IDAT"))
        assert status == 0 and bps == (2 if v["dtype"] == "uint16" else 1) and got == want, name
        n += 1
    assert n >= 8
    # error paths: wrong size, a flipped bit in the stream, a truncated file, not a PNG
    good = data["png_client_96x128/bytes"].tobytes()
    assert _run(host_inflate, good, 96, 127)[0] == 4
    broken = bytearray(good)
    broken[len(broken) // 2] ^= 0x55
    assert _run(host_inflate, bytes(broken), 96, 128)[0] != 0
    assert _run(host_inflate, good[:300], 96, 128)[0] != 0
    assert _run(host_inflate, data["jpg_tiny_8x8/bytes"].tobytes(), 8, 8)[0] == 1
