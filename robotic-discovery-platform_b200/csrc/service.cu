// Pre/post-processing the reference server does around the hot path, on the GPU so that a
// frame does not bounce through the host between decode, U-Net and geometry:
//   k_resize_aa     ToTensor + Resize((S,S), antialias=True)          services/vision_analysis/server.py:107-110,121
//   k_mask_nearest  cv2.resize(mask, (W,H), INTER_NEAREST) + coverage  server.py:125,133
#include "rdp_common.cuh"

namespace rdp {
namespace service {

// torchvision's antialiased bilinear resize = ATen _upsample_bilinear2d_aa: a separable triangle filter whose
// support grows with the down-scaling factor; weights are normalised per output index, horizontal pass first.
struct AaAxis {
  float scale, support, invscale;
  int in_size;
};
// Every step is rounded to fp32 as ATen rounds it: with the compiler free to contract, (j + lo) - scale * (i + 0.5)
// becomes one fma on the unrounded centre, which at coordinates of several hundred shifts a weight by 1e-5.
__device__ __forceinline__ void aa_window(const AaAxis& a, int i, int& lo, int& n, float& center) {
  center = __fmul_rn(a.scale, __fadd_rn((float)i, 0.5f));
  lo = max(0, (int)__fadd_rn(__fsub_rn(center, a.support), 0.5f));
  n = min(a.in_size, (int)__fadd_rn(__fadd_rn(center, a.support), 0.5f)) - lo;
}
__device__ __forceinline__ float aa_weight(const AaAxis& a, int j, int lo, float center) {
  const float x = __fmul_rn(__fadd_rn(__fsub_rn((float)(j + lo), center), 0.5f), a.invscale);
  const float w = __fsub_rn(1.0f, fabsf(x));
  return w > 0.f ? w : 0.f;
}

constexpr int MAX_C = 8;
constexpr int AA_TAPS = 32;  // horizontal taps kept per thread (a 15x reduction); wider windows recompute them

__global__ void __launch_bounds__(256) k_resize_aa(const uint8_t* __restrict__ in, int B, int H, int W, int C, int OH, int OW,
                                                   AaAxis ay, AaAxis ax, int swap_rb, float* __restrict__ out) {
  const int ox = blockIdx.x * blockDim.x + threadIdx.x;
  const int oy = blockIdx.y, n = blockIdx.z;
  if (ox >= OW) return;
  int x0, nx, y0, ny;
  float cx, cy;
  aa_window(ax, ox, x0, nx, cx);
  aa_window(ay, oy, y0, ny, cy);
  float sumx = 0.f, sumy = 0.f;
  for (int j = 0; j < nx; ++j) sumx += aa_weight(ax, j, x0, cx);
  for (int j = 0; j < ny; ++j) sumy += aa_weight(ay, j, y0, cy);
  // the normalised horizontal weights are the same for every source row: one division per tap, not one per tap and row
  float wxs[AA_TAPS];
  const bool cached = nx <= AA_TAPS;
  if (cached)
    for (int j = 0; j < nx; ++j) wxs[j] = aa_weight(ax, j, x0, cx) / sumx;
  float acc[MAX_C];
#pragma unroll
  for (int c = 0; c < MAX_C; ++c) acc[c] = 0.f;
  for (int jy = 0; jy < ny; ++jy) {
    const float wy = aa_weight(ay, jy, y0, cy) / sumy;
    const uint8_t* row = in + (((size_t)n * H + (y0 + jy)) * W + x0) * C;
    float h[MAX_C];
#pragma unroll
    for (int c = 0; c < MAX_C; ++c) h[c] = 0.f;
    for (int jx = 0; jx < nx; ++jx) {
      const float wx = cached ? wxs[jx] : aa_weight(ax, jx, x0, cx) / sumx;
#pragma unroll
      for (int c = 0; c < MAX_C; ++c)
        if (c < C) h[c] += wx * (__fdiv_rn((float)row[jx * C + c], 255.0f));  // ToTensor: uint8 / 255 (a shared-memory table of the 256 quotients was slower)
    }
#pragma unroll
    for (int c = 0; c < MAX_C; ++c) acc[c] += wy * h[c];
  }
#pragma unroll
  for (int c = 0; c < MAX_C; ++c)
    if (c < C) out[(((size_t)n * C + ((swap_rb && c < 3) ? 2 - c : c)) * OH + oy) * OW + ox] = acc[c];
}

// OpenCV's INTER_NEAREST: source index = min(floor(dst * (1 / (dst_size / src_size))), src_size - 1), in double.
__global__ void __launch_bounds__(256) k_mask_nearest(const uint8_t* __restrict__ in, int h, int w, int H, int W, double ify,
                                                      double ifx, uint8_t* __restrict__ out, int* __restrict__ nonzero) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y, n = blockIdx.z;
  int v = 0;
  if (x < W) {
    const int sy = min((int)floor((double)y * ify), h - 1), sx = min((int)floor((double)x * ifx), w - 1);
    const uint8_t m = in[((size_t)n * h + sy) * w + sx];
    out[((size_t)n * H + y) * W + x] = m;
    v = m != 0;
  }
  if (nonzero != nullptr) {
    const unsigned ball = __ballot_sync(0xffffffffu, v);
    __shared__ int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && ball) atomicAdd(&s_cnt, __popc(ball));
    __syncthreads();
    if (threadIdx.x == 0 && s_cnt) atomicAdd(&nonzero[n], s_cnt);
  }
}

static AaAxis make_axis(int in_size, int out_size) {
  AaAxis a;
  a.scale = (float)in_size / (float)out_size;
  a.support = a.scale >= 1.0f ? a.scale : 1.0f;
  a.invscale = a.scale >= 1.0f ? 1.0f / a.scale : 1.0f;
  a.in_size = in_size;
  return a;
}

}  // namespace service
}  // namespace rdp

using namespace rdp;
using namespace rdp::service;

extern "C" int rdp_pre_resize(const uint8_t* rgb_nhwc, int B, int H, int W, int C, int out_h, int out_w, int flags,
                              float* out_nchw, void* stream) {
  RDP_REQUIRE(rgb_nhwc && out_nchw, RDP_ERR_BAD_ARGUMENT, "rdp_pre_resize: null pointer");
  RDP_REQUIRE(B > 0 && H > 0 && W > 0 && out_h > 0 && out_w > 0 && B <= 65535 && out_h <= 65535, RDP_ERR_BAD_ARGUMENT,
              "rdp_pre_resize: bad shape %dx%dx%d -> %dx%d", B, H, W, out_h, out_w);
  RDP_REQUIRE(C >= 1 && C <= MAX_C, RDP_ERR_UNSUPPORTED, "rdp_pre_resize: %d channels unsupported (1..%d)", C, MAX_C);
  RDP_REQUIRE((flags & ~RDP_PRE_SWAP_RB) == 0 && (!(flags & RDP_PRE_SWAP_RB) || C >= 3), RDP_ERR_BAD_ARGUMENT,
              "rdp_pre_resize: bad flags 0x%x for %d channels", flags, C);
  k_resize_aa<<<dim3((out_w + 255) / 256, out_h, B), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      rgb_nhwc, B, H, W, C, out_h, out_w, make_axis(H, out_h), make_axis(W, out_w),
      (flags & RDP_PRE_SWAP_RB) != 0, out_nchw);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

extern "C" int rdp_post_mask_nearest(const uint8_t* mask, int B, int h, int w, int H, int W, uint8_t* out, int32_t* nonzero,
                                     void* stream_) {
  RDP_REQUIRE(mask && out, RDP_ERR_BAD_ARGUMENT, "rdp_post_mask_nearest: null pointer");
  RDP_REQUIRE(B > 0 && h > 0 && w > 0 && H > 0 && W > 0 && B <= 65535 && H <= 65535, RDP_ERR_BAD_ARGUMENT,
              "rdp_post_mask_nearest: bad shape %dx%dx%d -> %dx%d", B, h, w, H, W);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (nonzero) RDP_CUDA_TRY(cudaMemsetAsync(nonzero, 0, sizeof(int32_t) * B, stream));
  const double ify = 1.0 / ((double)H / (double)h), ifx = 1.0 / ((double)W / (double)w);
  k_mask_nearest<<<dim3((W + 255) / 256, H, B), 256, 0, stream>>>(mask, h, w, H, W, ify, ifx, out, nonzero);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

// ---------------------------------------------------------------------------------------------------------
// PNG encoding of the final mask on the GPU: replaces cv2.imencode('.png', final_mask * 255) (server.py:142-143)
// ---------------------------------------------------------------------------------------------------------
// The client decodes with cv2.imdecode(.., IMREAD_GRAYSCALE) (client.py:111-112), so any valid grey-scale PNG of the
// same picture is a drop-in.  Written here: 1 bit per pixel (1 = white), filter type 0 on every scan line, one zlib
// stream holding one fixed-Huffman deflate block.  A mask is long runs of 0x00 / 0xFF bytes once packed, so each row
// is tokenised as literal + distance-1 matches (run-length coding in deflate's terms).
//   k_png_rows      one thread per scan line: pack 8 pixels per byte, emit the row's deflate bits into its own
//                   96-byte slot, and the row's two Adler-32 partial sums
//   k_png_assemble  one CTA per frame: prefix sum of the rows' bit counts, rows OR-ed into place at bit granularity,
//                   Adler-32 combined from the partials, CRC-32 of the chunks, PNG framing
namespace rdp {
namespace service {

constexpr int PNG_THREADS = 256;
constexpr unsigned ADLER_MOD = 65521u;

struct BitWriter {
  uint32_t* out;
  unsigned long long acc;
  int nbits, words;
  __device__ __forceinline__ void put(uint32_t v, int n) {  // n <= 24 bits, LSB first
    acc |= (unsigned long long)v << nbits;
    nbits += n;
    if (nbits >= 32) {
      out[words++] = (uint32_t)acc;
      acc >>= 32;
      nbits -= 32;
    }
  }
  __device__ __forceinline__ int finish() {  // returns the number of bits written
    const int total = words * 32 + nbits;
    if (nbits > 0) out[words] = (uint32_t)acc;
    return total;
  }
};
// fixed Huffman codes (RFC 1951, 3.2.6); codes go into the stream starting from their most significant bit
__device__ __forceinline__ void put_literal(BitWriter& bw, uint32_t v) {
  if (v < 144) bw.put(__brev(0x30u + v) >> 24, 8);
  else bw.put(__brev(0x190u + (v - 144u)) >> 23, 9);
}
__device__ __forceinline__ void put_run_match(BitWriter& bw, int len) {  // length 3..258, distance 1
  int idx, extra_bits = 0, extra = 0;
  const int d = len - 3;
  if (len == 258) idx = 28;
  else if (d < 8) idx = d;
  else {
    const int e = (31 - __clz(d)) - 2;
    idx = 4 * (e + 1) + ((d - (4 << e)) >> e);
    extra_bits = e;
    extra = (d - (4 << e)) & ((1 << e) - 1);
  }
  const uint32_t code = 257u + idx;
  if (code < 280u) bw.put(__brev(code - 256u) >> 25, 7);
  else bw.put(__brev(0xC0u + (code - 280u)) >> 24, 8);
  if (extra_bits) bw.put((uint32_t)extra, extra_bits);
  bw.put(0u, 5);  // distance code 0 = distance 1
}
__device__ __forceinline__ void flush_run(BitWriter& bw, uint32_t v, int run) {
  put_literal(bw, v);
  int rem = run - 1;
  while (rem >= 3) {
    const int len = rem < 258 ? rem : 258;
    put_run_match(bw, len);
    rem -= len;
  }
  for (; rem > 0; --rem) put_literal(bw, v);
}

__global__ void __launch_bounds__(PNG_THREADS) k_png_rows(const uint8_t* __restrict__ mask, int B, int H, int W, int row_words,
                                                          uint32_t* __restrict__ row_bits, int* __restrict__ row_nbits,
                                                          uint2* __restrict__ row_adler) {
  const size_t r = (size_t)blockIdx.x * PNG_THREADS + threadIdx.x;
  if (r >= (size_t)B * H) return;
  const uint8_t* src = mask + r * W;
  const int wb = (W + 7) >> 3, L = wb + 1;
  BitWriter bw{row_bits + r * row_words, 0ull, 0, 0};
  uint32_t cur = 0u;  // the filter-type byte (0 = None) starts the scan line
  int run = 1;
  unsigned long long a = 0, s = 0;  // Adler partials of the row: a = sum d_j, s = sum (L - j) d_j  (the filter byte adds 0)
  const bool vec = (W % 8 == 0) && ((reinterpret_cast<uintptr_t>(src) & 7) == 0);
  for (int j = 0; j < wb; ++j) {
    uint32_t byte;
    if (vec) {
      const uint2 px = *reinterpret_cast<const uint2*>(src + 8 * j);
      const unsigned long long ones = (unsigned long long)(__vcmpne4(px.x, 0u) & 0x01010101u) |
                                      ((unsigned long long)(__vcmpne4(px.y, 0u) & 0x01010101u) << 32);
      byte = (uint32_t)((ones * 0x8040201008040201ull) >> 56);  // first pixel -> most significant bit
    } else {
      byte = 0;
      for (int k = 0; k < 8; ++k) {
        const int x = 8 * j + k;
        if (x < W && src[x] != 0) byte |= 0x80u >> k;
      }
    }
    a += byte;
    s += (unsigned long long)(L - (j + 1)) * byte;
    if (byte == cur) {
      ++run;
    } else {
      flush_run(bw, cur, run);
      cur = byte;
      run = 1;
    }
  }
  flush_run(bw, cur, run);
  row_nbits[r] = bw.finish();
  row_adler[r] = make_uint2((uint32_t)(a % ADLER_MOD), (uint32_t)(s % ADLER_MOD));
}

// CRC-32 register update over bytes that other threads of the CTA may have produced with atomics: read past L1.
__device__ __forceinline__ uint32_t crc_bytes(const uint32_t* table, uint32_t crc, const uint8_t* p, size_t n) {
  for (size_t i = 0; i < n; ++i) crc = table[(crc ^ __ldcg(p + i)) & 0xffu] ^ (crc >> 8);
  return crc;
}
// Polynomial arithmetic mod the CRC-32 polynomial in the reflected representation (x^0 = 1 << 31), as zlib's
// crc32_combine does it: the register after A||B is x^(8|B|) * register(A) + register(B from zero), so a chunk's
// register only has to be multiplied by x^(8 * bytes after the chunk) before the chunks are XOR-ed together.
__device__ __forceinline__ uint32_t crc_mulmod(uint32_t a, uint32_t b) {
  uint32_t p = 0;
  for (uint32_t m = 1u << 31; m != 0; m >>= 1) {
    if (a & m) p ^= b;
    b = (b & 1u) ? (b >> 1) ^ 0xEDB88320u : b >> 1;
  }
  return p;
}
__device__ __forceinline__ uint32_t crc_x_pow_8n(size_t n) {
  uint32_t p = 1u << 31, sq = 1u << 23;  // x^0, x^8
  for (; n != 0; n >>= 1) {
    if (n & 1) p = crc_mulmod(sq, p);
    sq = crc_mulmod(sq, sq);
  }
  return p;
}
__device__ __forceinline__ void put_be32(uint8_t* p, uint32_t v) {
  p[0] = (uint8_t)(v >> 24);
  p[1] = (uint8_t)(v >> 16);
  p[2] = (uint8_t)(v >> 8);
  p[3] = (uint8_t)v;
}

constexpr int PNG_DEFLATE_AT = 8 + 25 + 8 + 2;  // signature, IHDR chunk, IDAT length + type, zlib header

__global__ void __launch_bounds__(PNG_THREADS) k_png_assemble(int H, int W, int row_words, const uint32_t* __restrict__ row_bits,
                                                              const int* __restrict__ row_nbits,
                                                              const uint2* __restrict__ row_adler, uint8_t* __restrict__ png,
                                                              size_t stride, int* __restrict__ png_bytes,
                                                              unsigned long long* __restrict__ row_off) {
  __shared__ uint32_t s_crc[256];
  __shared__ unsigned long long s_scan[PNG_THREADS / 32];
  __shared__ unsigned long long s_a[PNG_THREADS / 32], s_s[PNG_THREADS / 32];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint8_t* out = png + (size_t)b * stride;
  uint32_t* out32 = reinterpret_cast<uint32_t*>(out);
  {
    uint32_t c = (uint32_t)tid;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
    s_crc[tid] = c;
  }
  // ---- bit offset of every row (exclusive scan over the frame's rows) and the Adler-32 partials ----
  const int wb = (W + 7) >> 3, L = wb + 1;
  const unsigned long long n_raw = (unsigned long long)H * L;
  unsigned long long carry = 0, asum = 0, ssum = 0;
  for (int r0 = 0; r0 < H; r0 += PNG_THREADS) {
    const int r = r0 + tid;
    const unsigned long long v = r < H ? (unsigned long long)row_nbits[(size_t)b * H + r] : 0ull;
    unsigned long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    unsigned long long woff = 0, tot = 0;
    for (int k = 0; k < PNG_THREADS / 32; ++k) {
      if (k < warp) woff += s_scan[k];
      tot += s_scan[k];
    }
    if (r < H) {
      row_off[(size_t)b * H + r] = carry + woff + incl - v;
      const uint2 ad = row_adler[(size_t)b * H + r];
      // a byte at position j of row r sits at i = r L + j of the raw stream and adds (n - i) d to the second sum
      const unsigned long long after = (n_raw - (unsigned long long)(r + 1) * L) % ADLER_MOD;
      asum += ad.x;
      ssum += (ad.y + after * ad.x) % ADLER_MOD;
    }
    carry += tot;
    __syncthreads();
  }
  const unsigned long long total_bits = 3 + carry + 7;  // block header, rows, end-of-block
  const size_t deflate_bytes = (size_t)((total_bits + 7) >> 3);
  const size_t end_byte = PNG_DEFLATE_AT + deflate_bytes;  // Adler-32 goes here
  const size_t png_total = end_byte + 4 + 4 + 12;          // Adler, IDAT CRC, IEND chunk
  // ---- clear the part of the output the rows are OR-ed into ----
  for (size_t i = tid; i < (end_byte + 3) / 4 + 1; i += PNG_THREADS) out32[i] = 0u;
  __syncthreads();
  // ---- rows into place ----
  const unsigned long long base_bit = (unsigned long long)PNG_DEFLATE_AT * 8 + 3;
  for (int r = tid; r < H; r += PNG_THREADS) {
    const int nb = row_nbits[(size_t)b * H + r];
    const uint32_t* src = row_bits + ((size_t)b * H + r) * row_words;
    const unsigned long long at = base_bit + row_off[(size_t)b * H + r];
    for (int w = 0; w * 32 < nb; ++w) {
      uint32_t v = src[w];
      const int valid = nb - w * 32;
      if (valid < 32) v &= (1u << valid) - 1u;
      const unsigned long long bit = at + 32ull * w;
      const size_t idx = (size_t)(bit >> 5);
      const int sh = (int)(bit & 31);
      atomicOr(&out32[idx], v << sh);
      if (sh) atomicOr(&out32[idx + 1], v >> (32 - sh));
    }
  }
  if (tid == 0) atomicOr(&out32[(PNG_DEFLATE_AT * 8) >> 5], 3u << ((PNG_DEFLATE_AT * 8) & 31));  // BFINAL = 1, BTYPE = 01
  // ---- Adler-32 of the raw scan lines ----
  for (int o = 16; o > 0; o >>= 1) {
    asum += __shfl_xor_sync(0xffffffffu, asum, o);
    ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
  }
  if (lane == 0) {
    s_a[warp] = asum;
    s_s[warp] = ssum;
  }
  __syncthreads();
  if (tid == 0) {
    unsigned long long A = 0, S = 0;
    for (int k = 0; k < PNG_THREADS / 32; ++k) {
      A += s_a[k];
      S += s_s[k];
    }
    const uint32_t s1 = (uint32_t)((1 + A) % ADLER_MOD), s2 = (uint32_t)((n_raw + S) % ADLER_MOD);
    // ---- framing (byte stores; every OR above is complete) ----
    const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    for (int i = 0; i < 8; ++i) out[i] = sig[i];
    put_be32(out + 8, 13u);
    out[12] = 'I'; out[13] = 'H'; out[14] = 'D'; out[15] = 'R';
    put_be32(out + 16, (uint32_t)W);
    put_be32(out + 20, (uint32_t)H);
    out[24] = 1;  // bit depth
    out[25] = 0;  // grey scale
    out[26] = 0; out[27] = 0; out[28] = 0;  // deflate, adaptive filtering (type 0 used), no interlace
    put_be32(out + 29, ~crc_bytes(s_crc, 0xffffffffu, out + 12, 17));
    put_be32(out + 33, (uint32_t)(2 + deflate_bytes + 4));
    out[37] = 'I'; out[38] = 'D'; out[39] = 'A'; out[40] = 'T';
    out[41] = 0x78; out[42] = 0x01;  // zlib: deflate, 32 KiB window, no preset dictionary, check bits
    put_be32(out + end_byte, (s2 << 16) | s1);
    const uint8_t iend[12] = {0, 0, 0, 0, 'I', 'E', 'N', 'D', 0xAE, 0x42, 0x60, 0x82};
    for (int i = 0; i < 12; ++i) out[end_byte + 8 + i] = iend[i];
    png_bytes[b] = (int)png_total;
  }
  __syncthreads();
  // ---- CRC-32 of the IDAT chunk (type + data), one slice per thread ----
  {
    const size_t n = 4 + 2 + deflate_bytes + 4, per = (n + PNG_THREADS - 1) / PNG_THREADS;
    const size_t lo = min((size_t)tid * per, n), hi = min((size_t)(tid + 1) * per, n);
    uint32_t r = crc_bytes(s_crc, tid == 0 ? 0xffffffffu : 0u, out + 37 + lo, hi - lo);
    r = r ? crc_mulmod(crc_x_pow_8n(n - hi), r) : 0u;
    for (int o = 16; o > 0; o >>= 1) r ^= __shfl_xor_sync(0xffffffffu, r, o);
    __shared__ uint32_t s_part[PNG_THREADS / 32];
    if (lane == 0) s_part[warp] = r;
    __syncthreads();
    if (tid == 0) {
      uint32_t c = 0;
      for (int k = 0; k < PNG_THREADS / 32; ++k) c ^= s_part[k];
      put_be32(out + end_byte + 4, ~c);
    }
  }
}

static size_t png_max_bytes(int H, int W, int* row_words_out) {
  const int wb = (W + 7) / 8;
  const int row_words = ((wb + 1) * 9 + 31) / 32 + 1;  // every byte a 9-bit literal
  if (row_words_out) *row_words_out = row_words;
  const size_t bits = 3 + (size_t)H * (size_t)(wb + 1) * 9 + 7;
  return align_up(PNG_DEFLATE_AT + (bits + 7) / 8 + 4 + 4 + 12 + 8, 16);
}
struct PngWs {
  uint32_t* row_bits;
  int* row_nbits;
  uint2* row_adler;
  unsigned long long* row_off;
  size_t total;
};
static PngWs png_carve(void* ws, int B, int H, int row_words) {
  Carver c(ws);
  PngWs p;
  p.row_bits = c.take<uint32_t>((size_t)B * H * row_words);
  p.row_nbits = c.take<int>((size_t)B * H);
  p.row_adler = c.take<uint2>((size_t)B * H);
  p.row_off = c.take<unsigned long long>((size_t)B * H);
  p.total = c.total();
  return p;
}

}  // namespace service
}  // namespace rdp

extern "C" int rdp_post_mask_png_sizes(int B, int H, int W, size_t* max_png_bytes, size_t* workspace_bytes) {
  RDP_REQUIRE(B > 0 && H > 0 && W > 0 && max_png_bytes && workspace_bytes, RDP_ERR_BAD_ARGUMENT,
              "rdp_post_mask_png_sizes: bad arguments (%dx%dx%d)", B, H, W);
  int row_words = 0;
  *max_png_bytes = png_max_bytes(H, W, &row_words);
  *workspace_bytes = png_carve(nullptr, B, H, row_words).total;
  return RDP_OK;
}

extern "C" int rdp_post_mask_png(const uint8_t* mask, int B, int H, int W, uint8_t* png, size_t png_stride, int32_t* png_bytes,
                                 void* workspace, size_t workspace_bytes, void* stream_) {
  RDP_REQUIRE(mask && png && png_bytes && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_post_mask_png: null pointer");
  RDP_REQUIRE(B > 0 && H > 0 && W > 0 && (size_t)B * H < (1ull << 31), RDP_ERR_BAD_ARGUMENT, "rdp_post_mask_png: bad shape %dx%dx%d",
              B, H, W);
  int row_words = 0;
  const size_t need_png = png_max_bytes(H, W, &row_words);
  RDP_REQUIRE(png_stride >= need_png && png_stride % 4 == 0 && (reinterpret_cast<uintptr_t>(png) & 3) == 0, RDP_ERR_BAD_ARGUMENT,
              "rdp_post_mask_png: png_stride %zu too small (need %zu) or buffer not 4-byte aligned", png_stride, need_png);
  const PngWs ws = png_carve(workspace, B, H, row_words);
  RDP_REQUIRE(workspace_bytes >= ws.total, RDP_ERR_WORKSPACE_TOO_SMALL, "rdp_post_mask_png: workspace %zu < %zu", workspace_bytes,
              ws.total);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const size_t rows = (size_t)B * H;
  k_png_rows<<<(unsigned)((rows + PNG_THREADS - 1) / PNG_THREADS), PNG_THREADS, 0, stream>>>(mask, B, H, W, row_words, ws.row_bits,
                                                                                           ws.row_nbits, ws.row_adler);
  k_png_assemble<<<B, PNG_THREADS, 0, stream>>>(H, W, row_words, ws.row_bits, ws.row_nbits, ws.row_adler, png, png_stride, png_bytes,
                                                ws.row_off);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}
