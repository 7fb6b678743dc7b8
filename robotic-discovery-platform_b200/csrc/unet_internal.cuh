// Pieces of csrc/unet.cu that the training step (csrc/train.cu) reuses: same kernels, other epilogue options.
#pragma once

namespace rdp {
namespace unet {

// plain convolution without BatchNorm / ReLU (k_conv3x3<BN, RAW>), bf16 NHWC in and out; see unet.cu
int raw_conv(const void* src0, int c0, const void* src1, int c1, const void* w, int cout, int taps, int B, int H, int W,
             const float* zero_shift, void* out, int num_sms, void* stream);
// 3x3xC neighbourhoods of the network input as 64 bf16 per pixel (k = (3*ky+kx)*C + c), see k_im2col_first
int im2col_first(const void* in, int in_fmt, int B, int H, int W, int C, void* out, void* stream);
// MaxPool2d(2) and the bilinear x2 up-sampling (+ pad to Hs x Ws) on bf16 NHWC
int maxpool2(const void* in, int B, int H, int W, int C, void* out, void* stream);
int upsample2x(const void* in, int B, int h, int w, int C, int Hs, int Ws, void* out, void* stream);

}  // namespace unet
}  // namespace rdp
