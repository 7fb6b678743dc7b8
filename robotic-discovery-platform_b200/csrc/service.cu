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
__device__ __forceinline__ void aa_window(const AaAxis& a, int i, int& lo, int& n, float& center) {
  center = a.scale * ((float)i + 0.5f);
  lo = max(0, (int)(center - a.support + 0.5f));
  n = min(a.in_size, (int)(center + a.support + 0.5f)) - lo;
}
__device__ __forceinline__ float aa_weight(const AaAxis& a, int j, int lo, float center) {
  const float x = ((float)(j + lo) - center + 0.5f) * a.invscale;
  const float w = 1.0f - fabsf(x);
  return w > 0.f ? w : 0.f;
}

constexpr int MAX_C = 8;

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
      const float wx = aa_weight(ax, jx, x0, cx) / sumx;
#pragma unroll
      for (int c = 0; c < MAX_C; ++c)
        if (c < C) h[c] += wx * (__fdiv_rn((float)row[jx * C + c], 255.0f));  // ToTensor: uint8 / 255
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
