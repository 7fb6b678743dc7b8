// Image decoders of the reference's wire format on sm_100a (the decode half).
//
// The reference server receives every frame as a JPEG colour image and a 16-bit PNG depth image and decodes both on the
// host with OpenCV (services/vision_analysis/server.py:117-118); the client writes them with cv2.imencode
// (services/vision_analysis/client.py:65-67).  Here the encoded bytes go to the GPU as they arrive - a fifth of the
// bytes of the decoded frame, which matters where the host-to-device link is the end-to-end limit - and come out as the
// arrays cv2.imdecode would have produced, bit for bit:
//
//   rdp_pre_decode_jpeg   baseline JPEG (SOF0, 8 bit, YCbCr 4:2:0 / 4:2:2 / 4:4:4 or grey) -> BGR uint8 (B,H,W,3)
//       k_jpeg_blocks   one warp per image: lane 0 walks the marker segments and the entropy-coded scan (ITU T.81
//                       F.2.2; 9-bit look-ahead tables in shared memory); after every MCU the whole warp dequantises and
//                       inverse-transforms its blocks with libjpeg's accurate integer IDCT (jidctint.c) into planar
//                       Y / Cb / Cr.  A Huffman-coded scan is one serial bit stream, so the parallelism is across the
//                       images of a batch (and the 31 other lanes share the arithmetic).
//       k_jpeg_colour   per output pixel: libjpeg's "fancy" triangle-filter chroma up-sampling (jdsample.c) and the
//                       fixed-point YCbCr -> RGB tables (jdcolor.c), stored as B, G, R like IMREAD_COLOR.
//   rdp_pre_decode_png16  non-interlaced 8/16-bit greyscale PNG -> uint16 (B,H,W)
//       k_png_inflate   one warp per image: lane 0 walks the chunks and inflates the zlib stream that the IDAT chunks
//                       carry (RFC 1950/1951: stored, fixed and dynamic Huffman blocks, any match distance).
//       k_png_unfilter  one CTA per image: scan-line filters None/Sub/Up/Average/Paeth in row order (Sub - all OpenCV
//                       writes - as a block-wide scan), big-endian samples to uint16.
// Per-image failures (not a baseline JPEG, wrong size, corrupt stream) are reported in `status`, never as garbage.
#include <stdarg.h>

#include "rdp_common.cuh"

namespace rdp {
namespace codec {

// ---------------------------------------------------------------------------------------------------------
// JPEG
// ---------------------------------------------------------------------------------------------------------
constexpr int JPEG_LOOK = 9;  // look-ahead bits of the fast Huffman path

struct HuffTable {
  uint16_t look[1 << JPEG_LOOK];  // (length << 8) | symbol for codes of up to JPEG_LOOK bits, 0 = longer code
  int32_t maxcode[18];            // largest code of each length (-1: none), [17] = sentinel
  int32_t valoff[17];             // symbol index = code + valoff[length]
  uint8_t symbols[256];
};

struct JpegFrame {  // filled by lane 0 while it parses the header
  int width, height, ncomp;
  int h[3], v[3], tq[3], td[3], ta[3];
  int hmax, vmax, mcux, mcuy;
  int restart;
};

struct JpegParams {
  const uint8_t* data;
  const int64_t* offsets;  // [B+1] byte ranges of the images inside `data`
  int B, H, W;
  uint8_t* planes;         // [B][plane_stride]: Y (padded to whole MCUs), then Cb, then Cr
  size_t plane_stride;
  int* geom;               // [B][8]: per image ncomp, hmax, vmax, padded Y width, padded chroma width, Y bytes, C bytes
  uint8_t* out;            // [B][H][W][3]
  int32_t* status;
};

__device__ __forceinline__ int zigzag_at(int k) {
  static const __device__ uint8_t zz[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                                            41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22,
                                            15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
                                            62, 63};
  return zz[k];
}

// Entropy-coded segment reader (T.81 F.2.2.5): FF00 is a stuffed FF, any other marker ends the data and feeds zeros.
struct JpegBits {
  const uint8_t* p;
  const uint8_t* end;
  uint64_t acc;
  int n;
  int fake;  // zero bits fed behind a marker or the end of the file; the reader looks ahead, so only CONSUMING them is an error
  __device__ __forceinline__ void fill() {
    while (n <= 56) {
      uint32_t b = 0;
      bool real = false;
      if (p < end) {
        b = *p;
        if (b == 0xFF) {
          const uint32_t nx = (p + 1 < end) ? p[1] : 0xD9u;
          if (nx == 0) {
            p += 2;
            real = true;
          } else {
            b = 0;  // a marker: stay on it
          }
        } else {
          ++p;
          real = true;
        }
      }
      if (!real) fake += 8;
      acc = (acc << 8) | b;
      n += 8;
    }
  }
  __device__ __forceinline__ bool overrun() const { return n < fake; }
  __device__ __forceinline__ uint32_t peek(int k) { return (uint32_t)(acc >> (n - k)) & ((1u << k) - 1u); }
  __device__ __forceinline__ void skip(int k) { n -= k; }
  __device__ __forceinline__ int receive_extend(int s) {  // F.2.2.1 RECEIVE + EXTEND
    if (s == 0) return 0;
    const int v = (int)peek(s);
    skip(s);
    return v >= (1 << (s - 1)) ? v : v - (1 << s) + 1;
  }
  __device__ void restart() {  // drop buffered bits, continue behind the next RSTn
    acc = 0;
    n = 0;
    fake = 0;
    while (p + 1 < end && !(p[0] == 0xFF && p[1] >= 0xD0 && p[1] <= 0xD7)) ++p;
    p += 2;
  }
};

__device__ __forceinline__ int huff_decode(JpegBits& bits, const HuffTable& t) {
  const uint32_t look = t.look[bits.peek(JPEG_LOOK)];
  if (look) {
    bits.skip(look >> 8);
    return look & 0xff;
  }
  int len = JPEG_LOOK + 1;
  int code = (int)bits.peek(len);
  while (len <= 16 && code > t.maxcode[len]) {
    ++len;
    code = (int)bits.peek(len);
  }
  if (len > 16) return -1;
  bits.skip(len);
  return t.symbols[(code + t.valoff[len]) & 0xff];
}

// T.81 C.2 / F.2.2.3: canonical codes from the 16 counts; the look-ahead table for the short ones.
__device__ void build_huff(HuffTable& t, const uint8_t* counts, const uint8_t* symbols, int nsym) {
  for (int i = 0; i < (1 << JPEG_LOOK); ++i) t.look[i] = 0;
  for (int i = 0; i < nsym && i < 256; ++i) t.symbols[i] = symbols[i];
  int code = 0, k = 0;
  for (int len = 1; len <= 16; ++len) {
    t.valoff[len] = k - code;
    for (int i = 0; i < counts[len - 1]; ++i, ++k, ++code) {
      if (len <= JPEG_LOOK) {
        const int lo = code << (JPEG_LOOK - len);
        for (int f = 0; f < (1 << (JPEG_LOOK - len)); ++f) t.look[lo + f] = (uint16_t)((len << 8) | symbols[k]);
      }
    }
    t.maxcode[len] = counts[len - 1] ? code - 1 : -1;
    code <<= 1;
  }
  t.maxcode[17] = 0x7fffffff;
}

// jidctint.c: jpeg_idct_islow (CONST_BITS 13, PASS1_BITS 2), one 8-point pass.
__device__ __forceinline__ void idct8(const int (&x)[8], int (&o)[8], int shift) {
  int z1 = (x[2] + x[6]) * 4433;
  const int tmp2 = z1 + x[6] * (-15137);
  const int tmp3 = z1 + x[2] * 6270;
  const int tmp0 = (x[0] + x[4]) << 13;
  const int tmp1 = (x[0] - x[4]) << 13;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  int t0 = x[7], t1 = x[5], t2 = x[3], t3 = x[1];
  z1 = t0 + t3;
  int z2 = t1 + t2, z3 = t0 + t2, z4 = t1 + t3;
  const int z5 = (z3 + z4) * 9633;
  t0 *= 2446;
  t1 *= 16819;
  t2 *= 25172;
  t3 *= 12299;
  z1 *= -7373;
  z2 *= -20995;
  z3 = z3 * (-16069) + z5;
  z4 = z4 * (-3196) + z5;
  t0 += z1 + z3;
  t1 += z2 + z4;
  t2 += z2 + z3;
  t3 += z1 + z4;
  const int r = 1 << (shift - 1);
  o[0] = (tmp10 + t3 + r) >> shift;
  o[7] = (tmp10 - t3 + r) >> shift;
  o[1] = (tmp11 + t2 + r) >> shift;
  o[6] = (tmp11 - t2 + r) >> shift;
  o[2] = (tmp12 + t1 + r) >> shift;
  o[5] = (tmp12 - t1 + r) >> shift;
  o[3] = (tmp13 + t0 + r) >> shift;
  o[4] = (tmp13 - t0 + r) >> shift;
}
// range_limit[v & RANGE_MASK] of jdmaster.c's table: clamp(v + 128), with the table's own wrap-around
__device__ __forceinline__ uint8_t idct_range_limit(int v) {
  const int i = v & 1023;
  return (uint8_t)(i < 128 ? i + 128 : (i < 512 ? 255 : (i < 896 ? 0 : i - 896)));
}

constexpr int JPEG_MAX_BLOCKS = 6;  // blocks per MCU: 4 Y + Cb + Cr at 4:2:0

#define JPEG_FAIL(code)                      \
  do {                                       \
    if (lane == 0) p.status[b] = (code);     \
    return;                                  \
  } while (0)

__global__ void __launch_bounds__(32) k_jpeg_blocks(const JpegParams p) {
  __shared__ HuffTable s_huff[4];          // DC0, DC1, AC0, AC1
  __shared__ uint16_t s_qt[4][64];         // natural order
  __shared__ int16_t s_coef[JPEG_MAX_BLOCKS][64];
  __shared__ int s_ws[JPEG_MAX_BLOCKS][64];
  __shared__ JpegFrame s_f;
  __shared__ int s_scan, s_err, s_more;
  const int b = blockIdx.x, lane = threadIdx.x;
  const uint8_t* d = p.data + p.offsets[b];
  const int len = (int)(p.offsets[b + 1] - p.offsets[b]);

  // ---- marker segments up to the first scan (T.81 B.2), lane 0 ----
  if (lane == 0) {
    int err = 0, scan = 0, have_sof = 0;
    JpegFrame f{};
    unsigned have_q = 0, have_h = 0;
    if (len < 4 || d[0] != 0xFF || d[1] != 0xD8) err = 1;
    int i = 2;
    while (!err && !scan) {
      if (i + 4 > len || d[i] != 0xFF) { err = 2; break; }
      const int m = d[i + 1];
      if (m == 0xFF) { ++i; continue; }
      const int L = (d[i + 2] << 8) | d[i + 3];
      if (L < 2 || i + 2 + L > len) { err = 2; break; }
      const uint8_t* seg = d + i + 4;
      const int sl = L - 2;
      if (m == 0xDB) {
        for (int j = 0; j + 65 <= sl; j += 65) {
          const int pq = seg[j] >> 4, tq = seg[j] & 15;
          if (pq || tq > 3) { err = 3; break; }
          for (int k = 0; k < 64; ++k) s_qt[tq][zigzag_at(k)] = seg[j + 1 + k];
          have_q |= 1u << tq;
        }
      } else if (m == 0xC0) {
        if (sl < 6 || seg[0] != 8) { err = 3; break; }
        f.height = (seg[1] << 8) | seg[2];
        f.width = (seg[3] << 8) | seg[4];
        f.ncomp = seg[5];
        if ((f.ncomp != 1 && f.ncomp != 3) || sl < 6 + 3 * f.ncomp) { err = 3; break; }
        for (int c = 0; c < f.ncomp; ++c) {
          f.td[c] = seg[6 + 3 * c];  // component id, replaced by the table selectors at SOS
          f.h[c] = seg[7 + 3 * c] >> 4;
          f.v[c] = seg[7 + 3 * c] & 15;
          f.tq[c] = seg[8 + 3 * c];
          if (f.tq[c] > 3) err = 3;
        }
        have_sof = 1;
      } else if (m >= 0xC1 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
        err = 3;  // extended, progressive, lossless or arithmetic-coded: not what cv2.imencode writes
      } else if (m == 0xC4) {
        int j = 0;
        while (j + 17 <= sl) {
          const int tc = seg[j] >> 4, th = seg[j] & 15;
          int n = 0;
          for (int k = 0; k < 16; ++k) n += seg[j + 1 + k];
          if (tc > 1 || th > 1 || n > 256 || j + 17 + n > sl) { err = 3; break; }
          build_huff(s_huff[tc * 2 + th], seg + j + 1, seg + j + 17, n);
          have_h |= 1u << (tc * 2 + th);
          j += 17 + n;
        }
      } else if (m == 0xDD) {
        f.restart = (seg[0] << 8) | seg[1];
      } else if (m == 0xDA) {
        if (!have_sof || seg[0] != f.ncomp) { err = 3; break; }
        for (int c = 0; c < f.ncomp; ++c) {
          if (seg[1 + 2 * c] != f.td[c]) err = 3;   // components in frame order, one interleaved scan
          const int tt = seg[2 + 2 * c];
          f.ta[c] = tt & 15;
          f.td[c] = tt >> 4;
          if (f.td[c] > 1 || f.ta[c] > 1 || !(have_h & (1u << f.td[c])) || !(have_h & (1u << (2 + f.ta[c]))) ||
              !(have_q & (1u << f.tq[c])))
            err = 3;
        }
        scan = i + 2 + L;
      }
      i += 2 + L;
    }
    if (!err) {
      f.hmax = f.vmax = 1;
      for (int c = 0; c < f.ncomp; ++c) {
        f.hmax = max(f.hmax, f.h[c]);
        f.vmax = max(f.vmax, f.v[c]);
      }
      // luma carries the sampling factor, chroma is 1x1: 4:4:4, 4:2:2 (2x1) or 4:2:0 (2x2); grey is 1x1
      if (f.h[0] != f.hmax || f.v[0] != f.vmax || f.hmax > 2 || f.vmax > 2 || (f.hmax == 1 && f.vmax == 2)) err = 3;
      for (int c = 1; c < f.ncomp; ++c)
        if (f.h[c] != 1 || f.v[c] != 1) err = 3;
      if (f.ncomp == 1 && (f.hmax != 1 || f.vmax != 1)) err = 3;
      if (f.width != p.W || f.height != p.H) err = 4;
      f.mcux = (f.width + 8 * f.hmax - 1) / (8 * f.hmax);
      f.mcuy = (f.height + 8 * f.vmax - 1) / (8 * f.vmax);
    }
    s_f = f;
    s_scan = scan;
    s_err = err;
  }
  __syncwarp();
  if (s_err) JPEG_FAIL(s_err);
  const JpegFrame f = s_f;
  const int yw = f.mcux * 8 * f.hmax, yh = f.mcuy * 8 * f.vmax, cw = f.mcux * 8, ch = f.mcuy * 8;
  uint8_t* plane_y = p.planes + (size_t)b * p.plane_stride;
  uint8_t* plane_c[2] = {plane_y + (size_t)yw * yh, plane_y + (size_t)yw * yh + (size_t)cw * ch};
  if ((size_t)yw * yh + (f.ncomp == 3 ? 2 * (size_t)cw * ch : 0) > p.plane_stride) JPEG_FAIL(4);
  if (lane == 0) {
    int* g = p.geom + 8 * b;
    g[0] = f.ncomp; g[1] = f.hmax; g[2] = f.vmax; g[3] = yw; g[4] = cw; g[5] = yw * yh; g[6] = cw * ch;
  }
  const int blocks_per_mcu = f.hmax * f.vmax + (f.ncomp == 3 ? 2 : 0);

  JpegBits bits;
  bits.p = d + s_scan;
  bits.end = d + len;
  bits.acc = 0;
  bits.n = 0;
  bits.fake = 0;
  int pred[3] = {0, 0, 0};
  int count = 0;
  for (int my = 0; my < f.mcuy; ++my) {
    for (int mx = 0; mx < f.mcux; ++mx, ++count) {
      // ---- entropy decoding of one MCU into shared memory, lane 0 (F.2.2) ----
      for (int i = lane; i < blocks_per_mcu * 64; i += 32) (&s_coef[0][0])[i] = 0;
      __syncwarp();
      if (lane == 0) {
        int err = 0;
        if (f.restart && count && count % f.restart == 0) {
          bits.restart();
          pred[0] = pred[1] = pred[2] = 0;
        }
        int blk = 0;
        for (int c = 0; c < f.ncomp && !err; ++c) {
          const HuffTable& dc = s_huff[f.td[c]];
          const HuffTable& ac = s_huff[2 + f.ta[c]];
          const int nb = c == 0 ? f.hmax * f.vmax : 1;
          for (int q = 0; q < nb && !err; ++q, ++blk) {
            bits.fill();
            int s = huff_decode(bits, dc);
            if (s < 0 || s > 11) { err = 5; break; }
            pred[c] += bits.receive_extend(s);
            s_coef[blk][0] = (int16_t)pred[c];
            for (int k = 1; k < 64;) {
              bits.fill();
              const int rs = huff_decode(bits, ac);
              if (rs < 0) { err = 5; break; }
              const int r = rs >> 4;
              s = rs & 15;
              if (s == 0) {
                if (r != 15) break;
                k += 16;
                continue;
              }
              k += r;
              if (k > 63) { err = 5; break; }
              s_coef[blk][zigzag_at(k)] = (int16_t)bits.receive_extend(s);
              ++k;
            }
          }
        }
        if (!err && bits.overrun()) err = 2;  // the scan ended (marker or end of file) before this MCU did
        s_more = err;
      }
      __syncwarp();
      if (s_more) JPEG_FAIL(s_more);
      // ---- dequantisation + inverse DCT of the MCU's blocks, whole warp ----
      for (int t = lane; t < blocks_per_mcu * 8; t += 32) {  // pass 1: columns
        const int blk = t >> 3, col = t & 7;
        const int c = blk < f.hmax * f.vmax ? 0 : blk - f.hmax * f.vmax + 1;
        const uint16_t* q = s_qt[f.tq[c]];
        int x[8], o[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) x[r] = (int)s_coef[blk][r * 8 + col] * (int)q[r * 8 + col];
        idct8(x, o, 13 - 2);
#pragma unroll
        for (int r = 0; r < 8; ++r) s_ws[blk][r * 8 + col] = o[r];
      }
      __syncwarp();
      for (int t = lane; t < blocks_per_mcu * 8; t += 32) {  // pass 2: rows, range limit, store
        const int blk = t >> 3, row = t & 7;
        int x[8], o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = s_ws[blk][row * 8 + k];
        idct8(x, o, 13 + 2 + 3);
        uint8_t* dst;
        if (blk < f.hmax * f.vmax) {
          const int by = blk / f.hmax, bx = blk - by * f.hmax;
          dst = plane_y + (size_t)((my * f.vmax + by) * 8 + row) * yw + (mx * f.hmax + bx) * 8;
        } else {
          dst = plane_c[blk - f.hmax * f.vmax] + (size_t)(my * 8 + row) * cw + mx * 8;
        }
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          lo |= (uint32_t)idct_range_limit(o[k]) << (8 * k);
          hi |= (uint32_t)idct_range_limit(o[4 + k]) << (8 * k);
        }
        *reinterpret_cast<uint2*>(dst) = make_uint2(lo, hi);  // rows of a plane are multiples of 8 bytes, base 256-aligned
      }
      __syncwarp();
    }
  }
  if (lane == 0) p.status[b] = 0;
}

// jdsample.c fancy up-sampling + jdcolor.c colour conversion, one thread per output pixel.
__global__ void __launch_bounds__(256) k_jpeg_colour(const JpegParams p) {
  const int b = blockIdx.z, y = blockIdx.y, x = blockIdx.x * 256 + threadIdx.x;
  if (x >= p.W) return;
  uint8_t* out = p.out + (((size_t)b * p.H + y) * p.W + x) * 3;
  if (p.status[b] != 0) {
    out[0] = out[1] = out[2] = 0;
    return;
  }
  const int* g = p.geom + 8 * b;
  const int ncomp = g[0], hmax = g[1], vmax = g[2], yw = g[3], cw = g[4];
  const uint8_t* py = p.planes + (size_t)b * p.plane_stride;
  const int Y = py[(size_t)y * yw + x];
  if (ncomp == 1) {
    out[0] = out[1] = out[2] = (uint8_t)Y;
    return;
  }
  const uint8_t* pc[2] = {py + g[5], py + g[5] + g[6]};
  const int dw = (p.W + hmax - 1) / hmax, dh = (p.H + vmax - 1) / vmax;  // real (not padded) chroma size
  int cbcr[2];
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const uint8_t* pl = pc[c];
    int v;
    if (hmax == 1) {
      v = pl[(size_t)y * cw + x];
    } else if (vmax == 1) {  // h2v1_fancy_upsample
      const int cx = x >> 1;
      const int cur = pl[(size_t)y * cw + cx];
      if (dw == 1) v = cur;
      else if (x & 1) v = cx == dw - 1 ? cur : (3 * cur + pl[(size_t)y * cw + cx + 1] + 2) >> 2;
      else v = cx == 0 ? cur : (3 * cur + pl[(size_t)y * cw + cx - 1] + 1) >> 2;
    } else {                 // h2v2_fancy_upsample: rows above the first / below the last repeat (jdmainct.c)
      const int cy = y >> 1, cx = x >> 1;
      int oy = (y & 1) ? cy + 1 : cy - 1;
      oy = oy < 0 ? 0 : (oy > dh - 1 ? dh - 1 : oy);
      const uint8_t* r0 = pl + (size_t)cy * cw;
      const uint8_t* r1 = pl + (size_t)oy * cw;
      const int cur = 3 * r0[cx] + r1[cx];
      if (x & 1) {
        v = cx == dw - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
      } else {
        v = cx == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
      }
    }
    cbcr[c] = v;
  }
  // jdcolor.c build_ycc_rgb_table (SCALEBITS 16): FIX(1.40200) = 91881, FIX(1.77200) = 116130, FIX(0.71414) = 46802,
  // FIX(0.34414) = 22554; >> is an arithmetic shift
  const int cb = cbcr[0] - 128, cr = cbcr[1] - 128;
  const int r = Y + ((91881 * cr + 32768) >> 16);
  const int gg = Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
  const int bl = Y + ((116130 * cb + 32768) >> 16);
  out[0] = (uint8_t)min(max(bl, 0), 255);
  out[1] = (uint8_t)min(max(gg, 0), 255);
  out[2] = (uint8_t)min(max(r, 0), 255);
}

// ---------------------------------------------------------------------------------------------------------
// PNG
// ---------------------------------------------------------------------------------------------------------
struct PngParams {
  const uint8_t* data;
  const int64_t* offsets;  // [B+1]
  int B, H, W;
  uint8_t* raw;            // [B][raw_stride] inflated, still filtered scan lines
  size_t raw_stride;
  int* geom;               // [B][8]: bytes per sample (1 or 2)
  uint16_t* out;           // [B][H][W]
  int32_t* status;
};

constexpr int PNG_LOOK = 10;
struct InflateTable {
  uint16_t look[1 << PNG_LOOK];  // (length << 12) | symbol for codes of up to PNG_LOOK bits (bit-reversed index), 0 = longer
  uint16_t first[16], offset[16], count[16];
  uint16_t sorted[288];
};

// Byte reader over the IDAT chunks of one PNG: chunk framing (length, type, CRC) is skipped on the fly.
struct PngBytes {
  const uint8_t* d;
  int len, pos, seg_end;
  __device__ bool next_chunk() {  // positions `pos` at the start of the next IDAT payload
    while (pos + 12 <= len) {
      const uint32_t L = ((uint32_t)d[pos] << 24) | (d[pos + 1] << 16) | (d[pos + 2] << 8) | d[pos + 3];
      const bool idat = d[pos + 4] == 'I' && d[pos + 5] == 'D' && d[pos + 6] == 'A' && d[pos + 7] == 'T';
      if ((size_t)pos + 12 + L > (size_t)len) return false;
      if (idat) {
        pos += 8;
        seg_end = pos + (int)L;
        return true;
      }
      if (d[pos + 4] == 'I' && d[pos + 5] == 'E') return false;  // IEND
      pos += 12 + (int)L;
    }
    return false;
  }
  __device__ __forceinline__ int get() {
    while (pos >= seg_end) {
      pos = seg_end + 4;  // CRC of the finished chunk
      if (!next_chunk()) return -1;
    }
    return d[pos++];
  }
};

struct InflateBits {
  PngBytes in;
  uint64_t acc;
  int n;
  int fake;  // zero bits appended behind the end of the data (the reader looks up to 8 bytes ahead of what is consumed)
  __device__ __forceinline__ void fill() {
    while (n <= 56) {
      const int b = in.get();
      if (b < 0) {
        fake += 8;
        n += 8;
      } else {
        acc |= (uint64_t)b << n;
        n += 8;
      }
    }
  }
  __device__ __forceinline__ bool overrun() const { return n < fake; }  // bits that were never in the file were consumed
  __device__ __forceinline__ uint32_t peek(int k) { return (uint32_t)acc & ((1u << k) - 1u); }
  __device__ __forceinline__ void skip(int k) {
    acc >>= k;
    n -= k;
  }
  __device__ __forceinline__ uint32_t get(int k) {
    fill();
    const uint32_t v = peek(k);
    skip(k);
    return v;
  }
};

// RFC 1951 3.2.2 canonical code; symbols sorted by (length, value); short codes also in a bit-reversed look-up table.
__device__ bool build_inflate(InflateTable& t, const uint8_t* lengths, int n) {
  for (int i = 0; i < 16; ++i) t.count[i] = 0;
  for (int i = 0; i < n; ++i) t.count[lengths[i]]++;
  t.count[0] = 0;
  int code = 0, off = 0;
  for (int L = 1; L < 16; ++L) {
    code = (code + t.count[L - 1]) << 1;
    t.first[L] = (uint16_t)code;
    t.offset[L] = (uint16_t)off;
    off += t.count[L];
    if (code + t.count[L] > (1 << L)) return false;  // over-subscribed
  }
  uint16_t next[16];
  for (int L = 1; L < 16; ++L) next[L] = t.offset[L];
  for (int i = 0; i < n; ++i)
    if (lengths[i]) t.sorted[next[lengths[i]]++] = (uint16_t)i;
  for (int i = 0; i < (1 << PNG_LOOK); ++i) t.look[i] = 0;
  for (int L = 1; L <= PNG_LOOK; ++L) {
    for (int j = 0; j < t.count[L]; ++j) {
      const int c = t.first[L] + j, sym = t.sorted[t.offset[L] + j];
      const int rev = (int)(__brev((unsigned)c) >> (32 - L));
      for (int f = rev; f < (1 << PNG_LOOK); f += 1 << L) t.look[f] = (uint16_t)((L << 12) | sym);
    }
  }
  return true;
}

__device__ __forceinline__ int inflate_symbol(InflateBits& bits, const InflateTable& t) {
  const uint32_t e = t.look[bits.peek(PNG_LOOK)];
  if (e) {
    bits.skip(e >> 12);
    return e & 0xfff;
  }
  // longer codes: walk the canonical code bit by bit (most significant bit of the code comes first in the stream)
  int code = (int)(__brev(bits.peek(PNG_LOOK)) >> (32 - PNG_LOOK));
  for (int L = PNG_LOOK + 1; L < 16; ++L) {
    code = (code << 1) | (int)((bits.peek(L) >> (L - 1)) & 1u);
    const int idx = code - t.first[L];
    if (idx >= 0 && idx < t.count[L]) {
      bits.skip(L);
      return t.sorted[t.offset[L] + idx];
    }
  }
  return -1;
}

__global__ void __launch_bounds__(32) k_png_inflate(const PngParams p) {
  __shared__ InflateTable s_lit, s_dist;
  __shared__ uint8_t s_len[320];
  const int b = blockIdx.x;
  if (threadIdx.x != 0) return;  // a deflate stream is one serial bit stream: one lane per image
  const uint8_t* d = p.data + p.offsets[b];
  const int len = (int)(p.offsets[b + 1] - p.offsets[b]);
  auto fail = [&](int code) { p.status[b] = code; };
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (len < 8 + 25) return fail(1);
  for (int i = 0; i < 8; ++i)
    if (d[i] != sig[i]) return fail(1);
  // IHDR must come first (PNG 5.6)
  if (!(d[12] == 'I' && d[13] == 'H' && d[14] == 'D' && d[15] == 'R')) return fail(2);
  const uint32_t w = ((uint32_t)d[16] << 24) | (d[17] << 16) | (d[18] << 8) | d[19];
  const uint32_t h = ((uint32_t)d[20] << 24) | (d[21] << 16) | (d[22] << 8) | d[23];
  const int depth = d[24], ctype = d[25], interlace = d[28];
  if (ctype != 0 || (depth != 8 && depth != 16) || interlace) return fail(3);
  if ((int)w != p.W || (int)h != p.H) return fail(4);
  const int bps = depth / 8;
  p.geom[8 * b] = bps;
  const size_t want = (size_t)p.H * (1 + (size_t)p.W * bps);
  if (want > p.raw_stride) return fail(4);
  uint8_t* out = p.raw + (size_t)b * p.raw_stride;
  size_t op = 0;

  InflateBits bits;
  bits.in.d = d;
  bits.in.len = len;
  bits.in.pos = 8;
  bits.in.seg_end = 8;
  bits.acc = 0;
  bits.n = 0;
  bits.fake = 0;
  if (!bits.in.next_chunk()) return fail(2);
  {
    const uint32_t cmf = bits.get(8), flg = bits.get(8);  // RFC 1950
    if ((cmf & 15) != 8 || ((cmf << 8) | flg) % 31 || (flg & 0x20)) return fail(5);
  }
  static const uint16_t len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
  static const uint8_t len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
  static const uint16_t dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
  static const uint8_t dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
  static const uint8_t cl_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
  for (;;) {
    const uint32_t final = bits.get(1), type = bits.get(2);
    if (type == 0) {
      bits.skip(bits.n & 7);
      const uint32_t n = bits.get(16), nn = bits.get(16);
      if ((n ^ nn) != 0xffffu) return fail(5);
      if (op + n > want) return fail(6);
      for (uint32_t i = 0; i < n; ++i) out[op++] = (uint8_t)bits.get(8);
    } else if (type == 1 || type == 2) {
      if (type == 1) {
        for (int i = 0; i < 288; ++i) s_len[i] = i < 144 ? 8 : (i < 256 ? 9 : (i < 280 ? 7 : 8));
        build_inflate(s_lit, s_len, 288);
        for (int i = 0; i < 30; ++i) s_len[i] = 5;
        build_inflate(s_dist, s_len, 30);
      } else {
        const int hlit = (int)bits.get(5) + 257, hdist = (int)bits.get(5) + 1, hclen = (int)bits.get(4) + 4;
        if (hlit > 286 || hdist > 30) return fail(5);
        uint8_t cl[19];
        for (int i = 0; i < 19; ++i) cl[i] = 0;
        for (int i = 0; i < hclen; ++i) cl[cl_order[i]] = (uint8_t)bits.get(3);
        if (!build_inflate(s_dist, cl, 19)) return fail(5);  // the code-length code borrows the distance table
        int i = 0;
        while (i < hlit + hdist) {
          bits.fill();
          const int s = inflate_symbol(bits, s_dist);
          if (s < 0) return fail(5);
          if (s < 16) {
            s_len[i++] = (uint8_t)s;
          } else {
            int rep, val = 0;
            if (s == 16) {
              if (i == 0) return fail(5);
              val = s_len[i - 1];
              rep = 3 + (int)bits.get(2);
            } else if (s == 17) {
              rep = 3 + (int)bits.get(3);
            } else {
              rep = 11 + (int)bits.get(7);
            }
            if (i + rep > hlit + hdist) return fail(5);
            while (rep--) s_len[i++] = (uint8_t)val;
          }
        }
        uint8_t dl[32];
        for (int k = 0; k < hdist; ++k) dl[k] = s_len[hlit + k];
        if (!build_inflate(s_lit, s_len, hlit)) return fail(5);
        if (!build_inflate(s_dist, dl, hdist)) return fail(5);
      }
      for (;;) {
        bits.fill();
        int s = inflate_symbol(bits, s_lit);
        if (s < 0) return fail(5);
        if (s < 256) {
          if (op >= want) return fail(6);
          out[op++] = (uint8_t)s;
        } else if (s == 256) {
          break;
        } else {
          s -= 257;
          if (s >= 29) return fail(5);
          const int length = len_base[s] + (int)bits.get(len_extra[s]);
          bits.fill();
          const int ds = inflate_symbol(bits, s_dist);
          if (ds < 0 || ds >= 30) return fail(5);
          const size_t dist = dist_base[ds] + bits.get(dist_extra[ds]);
          if (dist > op || op + length > want) return fail(6);
          for (int i = 0; i < length; ++i, ++op) out[op] = out[op - dist];
        }
        if (bits.overrun()) return fail(2);
      }
    } else {
      return fail(5);
    }
    if (bits.overrun()) return fail(2);
    if (final) break;
  }
  if (op != want) return fail(6);
  p.status[b] = 0;
}

__device__ __forceinline__ int paeth(int a, int b, int c) {
  const int pp = a + b - c, pa = abs(pp - a), pb = abs(pp - b), pc = abs(pp - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// PNG filters (spec section 9) in row order; `raw` is overwritten with the reconstructed bytes, then samples are emitted.
constexpr int UNF_THREADS = 256;
__global__ void __launch_bounds__(UNF_THREADS) k_png_unfilter(const PngParams p) {
  __shared__ int s_scan[UNF_THREADS / 32][2];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint16_t* out = p.out + (size_t)b * p.H * p.W;
  if (p.status[b] != 0) {
    for (size_t i = tid; i < (size_t)p.H * p.W; i += UNF_THREADS) out[i] = 0;
    return;
  }
  const int bps = p.geom[8 * b];
  const int stride = p.W * bps;
  uint8_t* raw = p.raw + (size_t)b * p.raw_stride;
  for (int r = 0; r < p.H; ++r) {
    uint8_t* cur = raw + (size_t)r * (stride + 1) + 1;
    const uint8_t* prev = r ? raw + (size_t)(r - 1) * (stride + 1) + 1 : nullptr;
    const int ft = cur[-1];
    if (ft == 2) {
      for (int i = tid; i < stride; i += UNF_THREADS) cur[i] = (uint8_t)(cur[i] + (prev ? prev[i] : 0));
    } else if (ft == 1) {
      // Sub: byte i adds byte i - bps, i.e. a running sum (mod 256) per byte lane; block-wide scan over pixel chunks
      int carry[2] = {0, 0};
      for (int base = 0; base < p.W; base += UNF_THREADS) {
        const int px = base + tid;
        int v[2] = {0, 0};
        if (px < p.W)
          for (int k = 0; k < bps; ++k) v[k] = cur[px * bps + k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int a0 = __shfl_up_sync(0xffffffffu, v[0], o), a1 = __shfl_up_sync(0xffffffffu, v[1], o);
          if (lane >= o) {
            v[0] += a0;
            v[1] += a1;
          }
        }
        if (lane == 31) {
          s_scan[warp][0] = v[0];
          s_scan[warp][1] = v[1];
        }
        __syncthreads();
        int off0 = carry[0], off1 = carry[1];
        for (int w = 0; w < warp; ++w) {
          off0 += s_scan[w][0];
          off1 += s_scan[w][1];
        }
        for (int w = 0; w < UNF_THREADS / 32; ++w) {
          carry[0] += s_scan[w][0];
          carry[1] += s_scan[w][1];
        }
        if (px < p.W) {
          cur[px * bps] = (uint8_t)(v[0] + off0);
          if (bps == 2) cur[px * bps + 1] = (uint8_t)(v[1] + off1);
        }
        __syncthreads();
      }
    } else if (ft == 3 || ft == 4) {
      if (tid == 0) {  // Average / Paeth depend on the byte to the left: one thread walks the row
        for (int i = 0; i < stride; ++i) {
          const int a = i >= bps ? cur[i - bps] : 0, up = prev ? prev[i] : 0, c = (prev && i >= bps) ? prev[i - bps] : 0;
          cur[i] = (uint8_t)(cur[i] + (ft == 3 ? ((a + up) >> 1) : paeth(a, up, c)));
        }
      }
    } else if (ft != 0) {
      if (tid == 0) p.status[b] = 7;
    }
    __syncthreads();
    for (int x = tid; x < p.W; x += UNF_THREADS)
      out[(size_t)r * p.W + x] = bps == 2 ? (uint16_t)((cur[2 * x] << 8) | cur[2 * x + 1]) : (uint16_t)cur[x];
  }
}

}  // namespace codec
}  // namespace rdp

using namespace rdp;
using namespace rdp::codec;

static size_t jpeg_plane_stride(int H, int W) {
  const size_t yw = (size_t)(W + 15) / 16 * 16, yh = (size_t)(H + 15) / 16 * 16;
  return align_up(yw * yh * 3, 256);  // Y + two chroma planes, at most full resolution each (4:4:4)
}

extern "C" int rdp_pre_decode_sizes(int B, int H, int W, size_t* jpeg_workspace_bytes, size_t* png_workspace_bytes) {
  RDP_REQUIRE(B > 0 && H > 0 && W > 0 && H <= 16384 && W <= 16384, RDP_ERR_BAD_ARGUMENT, "rdp_pre_decode_sizes: bad shape %dx%dx%d", B, H, W);
  if (jpeg_workspace_bytes) *jpeg_workspace_bytes = align_up((size_t)B * 8 * sizeof(int), 256) + (size_t)B * jpeg_plane_stride(H, W);
  if (png_workspace_bytes) *png_workspace_bytes = align_up((size_t)B * 8 * sizeof(int), 256) + (size_t)B * align_up((size_t)H * (1 + 2 * (size_t)W), 256);
  return RDP_OK;
}

extern "C" int rdp_pre_decode_jpeg(const uint8_t* data, const int64_t* offsets, int B, int H, int W, uint8_t* bgr, int32_t* status,
                                   void* workspace, size_t workspace_bytes, void* stream_) {
  RDP_REQUIRE(data && offsets && bgr && status && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_pre_decode_jpeg: null pointer");
  RDP_REQUIRE(B > 0 && H > 0 && W > 0 && H <= 16384 && W <= 16384, RDP_ERR_BAD_ARGUMENT, "rdp_pre_decode_jpeg: bad shape %dx%dx%d", B, H, W);
  size_t need = 0;
  rdp_pre_decode_sizes(B, H, W, &need, nullptr);
  RDP_REQUIRE(workspace_bytes >= need, RDP_ERR_WORKSPACE_TOO_SMALL, "rdp_pre_decode_jpeg: workspace %zu < required %zu bytes", workspace_bytes, need);
  RDP_REQUIRE(((uintptr_t)workspace & 255) == 0, RDP_ERR_BAD_ARGUMENT, "rdp_pre_decode_jpeg: workspace must be 256-byte aligned");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  JpegParams p{};
  p.data = data; p.offsets = offsets; p.B = B; p.H = H; p.W = W;
  p.geom = static_cast<int*>(workspace);
  p.planes = static_cast<uint8_t*>(workspace) + align_up((size_t)B * 8 * sizeof(int), 256);
  p.plane_stride = jpeg_plane_stride(H, W);
  p.out = bgr;
  p.status = status;
  k_jpeg_blocks<<<B, 32, 0, stream>>>(p);
  k_jpeg_colour<<<dim3((W + 255) / 256, H, B), 256, 0, stream>>>(p);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

extern "C" int rdp_pre_decode_png16(const uint8_t* data, const int64_t* offsets, int B, int H, int W, uint16_t* depth,
                                    int32_t* status, void* workspace, size_t workspace_bytes, void* stream_) {
  RDP_REQUIRE(data && offsets && depth && status && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_pre_decode_png16: null pointer");
  RDP_REQUIRE(B > 0 && H > 0 && W > 0 && H <= 16384 && W <= 16384, RDP_ERR_BAD_ARGUMENT, "rdp_pre_decode_png16: bad shape %dx%dx%d", B, H, W);
  size_t need = 0;
  rdp_pre_decode_sizes(B, H, W, nullptr, &need);
  RDP_REQUIRE(workspace_bytes >= need, RDP_ERR_WORKSPACE_TOO_SMALL, "rdp_pre_decode_png16: workspace %zu < required %zu bytes", workspace_bytes, need);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  PngParams p{};
  p.data = data; p.offsets = offsets; p.B = B; p.H = H; p.W = W;
  p.geom = static_cast<int*>(workspace);
  p.raw = static_cast<uint8_t*>(workspace) + align_up((size_t)B * 8 * sizeof(int), 256);
  p.raw_stride = align_up((size_t)H * (1 + 2 * (size_t)W), 256);
  p.out = depth;
  p.status = status;
  k_png_inflate<<<B, 32, 0, stream>>>(p);
  k_png_unfilter<<<B, UNF_THREADS, 0, stream>>>(p);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}
