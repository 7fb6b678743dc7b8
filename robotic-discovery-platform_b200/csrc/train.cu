// One training step of the U-Net on sm_100a: the forward pass in training mode and the backward pass
// that `loss.backward()` runs in the reference's loop (scripts/train_segmenter.py:151-165; `model.train()`, so every
// BatchNorm2d normalises with the statistics of the batch, pkg/segmentation_model.py:31-36).  The loss
// (BCEWithLogitsLoss) and the optimiser (Adam) stay torch's, exactly as in that loop: this file provides what sits
// between `model(images)` and the parameters' `.grad`.
//
// Data: activations NHWC bf16 as in the inference path, gradients of activations bf16, parameter gradients fp32,
// batch statistics accumulated in fp64.  Per 3x3 conv + BatchNorm + ReLU stage i, forward
//     Z_i = conv(X_i, W_i)            k_conv3x3_halo<.., RAW> (tcgen05 implicit GEMM, csrc/unet.cu)
//     mu, var = stats(Z_i)            k_bn_stats + k_bn_finalize   per-channel sum and sum of squares over B*H*W
//     Y_i = relu(g (Z_i - mu)/sqrt(var + eps) + b)     k_bn_relu
// and backward, given dY_i,
//     dbeta, dgamma                   k_bn_bwd_reduce + k_bn_bwd_finalize (dY masked by Y > 0, re-derived from Z)
//     dZ_i = g/sigma (dY~ - dbeta/n - xhat dgamma/n)   k_bn_bwd_apply
//     dW_i = sum_p dZ_i[p] (x) X_i[p + tap]            k_wgrad: tcgen05 GEMM with K = the pixels, both operands read in place
//                                                      from the NHWC tensors as MN-major operands; k_dw_layout
//     dX_i = conv(dZ_i, W_i flipped and transposed)    k_conv3x3_halo<.., RAW> again
// plus the adjoints of MaxPool2d(2) (first maximum wins, like ATen), of the bilinear x2 up-sampling with its pad, of the
// channel concatenation, and the 1x1 OutConv in both directions.
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdarg.h>
#include <stdlib.h>

#include <algorithm>

#include "rdp_common.cuh"
#include "sm100_ptx.cuh"
#include "unet_internal.cuh"

// Development builds (-DRDP_DEVELOPMENT) with RDP_TRAIN_SYNC set wait after every launch group and name the one that failed.
#ifdef RDP_DEVELOPMENT
#define TRAIN_CHECKPOINT(stream, ...)                                                     \
  do {                                                                                    \
    if (getenv("RDP_TRAIN_SYNC")) {                                                       \
      const cudaError_t _e = cudaStreamSynchronize(stream);                               \
      fprintf(stderr, "train checkpoint: ");                                              \
      fprintf(stderr, __VA_ARGS__);                                                       \
      fprintf(stderr, " -> %s\n", cudaGetErrorString(_e));                                \
      if (_e != cudaSuccess) {                                                            \
        ::rdp::set_error("training step failed at a checkpoint: %s", cudaGetErrorString(_e)); \
        return RDP_ERR_CUDA;                                                              \
      }                                                                                   \
    }                                                                                     \
  } while (0)
#else
#define TRAIN_CHECKPOINT(stream, ...) do { } while (0)
#endif

namespace rdp {
namespace train {

using namespace ptx;

// ---------------------------------------------------------------------------------------------------------
// small helpers on 8 bf16 (16 bytes)
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 v;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return v;
}

// ---------------------------------------------------------------------------------------------------------
// BatchNorm (training mode) forward
// ---------------------------------------------------------------------------------------------------------
// Per-thread partial sums of one 8-channel group -> the block's [2][C] array `s_acc` (zeroed by the caller), then the
// block's column of the partial-sum table `part` ([2C][gridDim.x]).  No atomics anywhere: lanes of a warp that hold the same
// group (c8 < 32: lanes cg, cg + c8, ...) are folded with shuffles, the eight warps then add their values one warp at a
// time (within a warp the active lanes own distinct channels), and the finalize kernel adds the blocks' rows in a fixed
// order.  The statistics are therefore bit-reproducible from run to run.  Called by every thread of the block; needs
// c8 to be a power of two no larger than the block (threads without rows pass zeros).
constexpr int ST_THREADS = 256;
__device__ __forceinline__ void block_channel_sums(double (&a)[8], double (&q)[8], double* s_acc, int C, int c8, int cg, double* part) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  bool active = true;
  if (c8 < 32) {
    for (int off = 16; off >= c8; off >>= 1) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        a[k] += __shfl_xor_sync(0xffffffffu, a[k], off);
        q[k] += __shfl_xor_sync(0xffffffffu, q[k], off);
      }
    }
    active = lane < c8;
  }
  for (int w = 0; w < ST_THREADS / 32; ++w) {
    if (warp == w && active) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        s_acc[cg * 8 + k] += a[k];
        s_acc[C + cg * 8 + k] += q[k];
      }
    }
    __syncthreads();
  }
  // table layout [2C][gridDim.x]: the finalize kernel's warp reads one row of consecutive blocks per channel
  for (int i = tid; i < 2 * C; i += ST_THREADS) part[(size_t)i * gridDim.x + blockIdx.x] = s_acc[i];
}

// the two sums of channel c over the blocks' rows of the table: one warp per channel, fixed order
__device__ __forceinline__ void reduce_partials(const double* __restrict__ part, int nb, int C, int c, int lane, double& s0, double& s1) {
  const double* px = part + (size_t)c * nb;
  const double* py = part + (size_t)(C + c) * nb;
  double x = 0.0, y = 0.0;
  int b = lane;
  for (; b + 96 < nb; b += 128) {  // four coalesced loads per sum in flight per lane, added in a fixed order
    const double x0 = px[b], x1 = px[b + 32], x2 = px[b + 64], x3 = px[b + 96];
    const double y0 = py[b], y1 = py[b + 32], y2 = py[b + 64], y3 = py[b + 96];
    x += (x0 + x1) + (x2 + x3);
    y += (y0 + y1) + (y2 + y3);
  }
  for (; b < nb; b += 32) {
    x += px[b];
    y += py[b];
  }
#pragma unroll
  for (int off = 16; off; off >>= 1) {
    x += __shfl_xor_sync(0xffffffffu, x, off);
    y += __shfl_xor_sync(0xffffffffu, y, off);
  }
  s0 = x;
  s1 = y;
}

// part[block][c] = sum_p z[p][c], part[block][C + c] = sum_p z[p][c]^2 over the block's rows (z is bf16 [P][C], C/8 a
// power of two <= 256).  A thread owns 8 channels and walks rows.
__global__ void __launch_bounds__(ST_THREADS) k_bn_stats(const __nv_bfloat16* __restrict__ z, size_t P, int C, double* part) {
  extern __shared__ double s_acc[];  // [2][C]
  const int c8 = C / 8, tid = threadIdx.x;
  const int lanes = ST_THREADS / c8;  // rows handled side by side
  for (int i = tid; i < 2 * C; i += ST_THREADS) s_acc[i] = 0.0;
  __syncthreads();
  const int cg = tid % c8, rl = tid / c8;
  double a[8], q[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = q[k] = 0.0;
  // four rows in flight; their sum and sum of squares in fp32 (exact to ~2^-22 relative for four bf16 values), then fp64
  const size_t rstride = (size_t)gridDim.x * lanes;
  for (size_t r = (size_t)blockIdx.x * lanes + rl; r < P; r += 4 * rstride) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t ru = r + u * rstride;
      v[u] = ru < P ? *reinterpret_cast<const uint4*>(z + ru * C + cg * 8) : make_uint4(0, 0, 0, 0);
    }
    float sa[8], sq[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) sa[k] = sq[k] = 0.f;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float f[8];
      unpack8(v[u], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        sa[k] += f[k];
        sq[k] = fmaf(f[k], f[k], sq[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a[k] += (double)sa[k];
      q[k] += (double)sq[k];
    }
  }
  block_channel_sums(a, q, s_acc, C, c8, cg, part);
}

// mean, biased variance, and the affine form of the normalisation: y = z * scale + shift.  One warp per channel.
__global__ void __launch_bounds__(256) k_bn_finalize(const double* __restrict__ part, int nb, int C, double n, const float* gamma,
                                                     const float* beta, float eps, float* mean, float* var, float* invstd, float* scale,
                                                     float* shift) {
  const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  double s0, s1;
  reduce_partials(part, nb, C, c, lane, s0, s1);
  if (lane) return;
  const double m = s0 / n;
  double v = s1 / n - m * m;
  v = v > 0.0 ? v : 0.0;
  const double is = 1.0 / sqrt(v + (double)eps);
  mean[c] = (float)m;
  var[c] = (float)v;
  invstd[c] = (float)is;
  scale[c] = (float)((double)gamma[c] * is);
  shift[c] = (float)((double)beta[c] - m * (double)gamma[c] * is);
}

// A thread keeps one 8-channel group for the whole walk (the grid stride is a multiple of c8: c8 divides 256), so the
// affine coefficients sit in registers.
__global__ void __launch_bounds__(256) k_bn_relu(const __nv_bfloat16* __restrict__ z, size_t n8, int c8, const float* __restrict__ scale,
                                                 const float* __restrict__ shift, __nv_bfloat16* __restrict__ y) {
  const int cg = threadIdx.x % c8;
  float sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    sc[k] = scale[cg * 8 + k];
    sh[k] = shift[cg * 8 + k];
  }
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (; i + stride < n8; i += 2 * stride) {
    const uint4 v0 = reinterpret_cast<const uint4*>(z)[i], v1 = reinterpret_cast<const uint4*>(z)[i + stride];
    float f0[8], f1[8];
    unpack8(v0, f0);
    unpack8(v1, f1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      f0[k] = fmaxf(fmaf(f0[k], sc[k], sh[k]), 0.f);
      f1[k] = fmaxf(fmaf(f1[k], sc[k], sh[k]), 0.f);
    }
    reinterpret_cast<uint4*>(y)[i] = pack8(f0);
    reinterpret_cast<uint4*>(y)[i + stride] = pack8(f1);
  }
  if (i < n8) {
    float f[8];
    unpack8(reinterpret_cast<const uint4*>(z)[i], f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = fmaxf(fmaf(f[k], sc[k], sh[k]), 0.f);
    reinterpret_cast<uint4*>(y)[i] = pack8(f);
  }
}

// ---------------------------------------------------------------------------------------------------------
// OutConv (1x1, 64 -> n_classes, bias), both directions
// ---------------------------------------------------------------------------------------------------------
// Eight lanes share a pixel (one 8-channel group each, so a warp reads 512 contiguous bytes per load), partial dot
// products are folded with three shuffles, lane 0 of the eight writes.  Four pixels in flight per lane group.
__global__ void __launch_bounds__(256) k_outconv_fwd(const __nv_bfloat16* __restrict__ y, size_t P, size_t HW, int ncls,
                                                     const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ logits) {
  const int cg = threadIdx.x & 7, rl = threadIdx.x >> 3;
  const size_t pstride = (size_t)gridDim.x * 32;
  for (size_t p0 = (size_t)blockIdx.x * 32 + rl; p0 < P; p0 += 4 * pstride) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t pu = p0 + u * pstride;
      v[u] = pu < P ? reinterpret_cast<const uint4*>(y + pu * 64)[cg] : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t pu = p0 + u * pstride;   // uniform across the eight lanes of a pixel; the shuffles below need every lane
      float f[8];
      unpack8(v[u], f);
      for (int cls = 0; cls < ncls; ++cls) {
        float a = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) a = fmaf(f[k], w[cls * 64 + cg * 8 + k], a);
        a += __shfl_xor_sync(0xffffffffu, a, 1);
        a += __shfl_xor_sync(0xffffffffu, a, 2);
        a += __shfl_xor_sync(0xffffffffu, a, 4);
        if (cg == 0 && pu < P) {
          const size_t n = pu / HW, off = pu % HW;
          logits[(n * ncls + cls) * HW + off] = a + bias[cls];
        }
      }
    }
  }
}

// dY[p][c] = sum_cls dlogit[cls][p] w[cls][c];  dW[cls][c] += sum_p dlogit y;  db[cls] += sum_p dlogit.
// A thread owns one 8-channel group and walks pixels (32 side by side per block); its share of dW stays in registers
// until the end, then one shared-memory and one global atomic per value.
template <int NC>  // classes the registers are sized for (>= ncls)
__global__ void __launch_bounds__(256) k_outconv_bwd(const float* __restrict__ dlogits, const __nv_bfloat16* __restrict__ y, size_t P,
                                                     size_t HW, int ncls, const float* __restrict__ w, __nv_bfloat16* __restrict__ dy,
                                                     float* dw, float* db) {
  __shared__ float s_dw[8 * 64 + 8];
  for (int i = threadIdx.x; i < 8 * 64 + 8; i += blockDim.x) s_dw[i] = 0.f;
  __syncthreads();
  const int cg = threadIdx.x & 7, rl = threadIdx.x >> 3;
  float acc[NC][8], accb[NC], wv[NC][8];
#pragma unroll
  for (int cls = 0; cls < NC; ++cls) {
    accb[cls] = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      acc[cls][k] = 0.f;
      wv[cls][k] = cls < ncls ? w[cls * 64 + cg * 8 + k] : 0.f;
    }
  }
  const size_t pstride = (size_t)gridDim.x * 32;
  for (size_t p0 = (size_t)blockIdx.x * 32 + rl; p0 < P; p0 += 4 * pstride) {  // four pixels in flight
    uint4 yv[4];
    float gv[4][NC];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t pu = p0 + u * pstride;
      const bool in = pu < P;
      yv[u] = in ? reinterpret_cast<const uint4*>(y + pu * 64)[cg] : make_uint4(0, 0, 0, 0);
      if (NC == 1) {
        gv[u][0] = in ? dlogits[pu] : 0.f;  // one class: (n, 0, off) is pixel pu itself
      } else {
        const size_t n = pu / HW, off = pu % HW;
#pragma unroll
        for (int cls = 0; cls < NC; ++cls) gv[u][cls] = (in && cls < ncls) ? dlogits[(n * ncls + cls) * HW + off] : 0.f;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t pu = p0 + u * pstride;
      if (pu >= P) break;
      float f[8], o[8];
      unpack8(yv[u], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = 0.f;
#pragma unroll
      for (int cls = 0; cls < NC; ++cls) {
        if (cls < ncls) {
          const float g = gv[u][cls];
          accb[cls] += g;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            o[k] = fmaf(g, wv[cls][k], o[k]);
            acc[cls][k] = fmaf(g, f[k], acc[cls][k]);
          }
        }
      }
      reinterpret_cast<uint4*>(dy + pu * 64)[cg] = pack8(o);
    }
  }
#pragma unroll
  for (int cls = 0; cls < NC; ++cls) {
    if (cls < ncls) {
#pragma unroll
      for (int k = 0; k < 8; ++k) atomicAdd(&s_dw[cls * 64 + cg * 8 + k], acc[cls][k]);
      if (cg == 0) atomicAdd(&s_dw[8 * 64 + cls], accb[cls]);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < ncls * 64; i += blockDim.x) atomicAdd(&dw[i], s_dw[i]);
  if (threadIdx.x < ncls) atomicAdd(&db[threadIdx.x], s_dw[8 * 64 + threadIdx.x]);
}

// ---------------------------------------------------------------------------------------------------------
// BatchNorm + ReLU backward
// ---------------------------------------------------------------------------------------------------------
// part[block][c] = sum_p dY~, part[block][C + c] = sum_p dY~ xhat over the block's rows,  dY~ = dY [Y > 0],
// xhat = (Z - mean) invstd.  Y > 0 is re-derived from Z with the forward pass's own expression
// (fmaf(z, scale, shift) > 0: bf16 rounding keeps the sign), which saves reading Y.
__global__ void __launch_bounds__(ST_THREADS) k_bn_bwd_reduce(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ scale,
                                                              const float* __restrict__ shift,
                                                              const __nv_bfloat16* __restrict__ z, size_t P, int C,
                                                              const float* __restrict__ mean, const float* __restrict__ invstd, double* part) {
  extern __shared__ double s_acc[];
  const int c8 = C / 8, tid = threadIdx.x;
  const int lanes = ST_THREADS / c8;
  for (int i = tid; i < 2 * C; i += ST_THREADS) s_acc[i] = 0.0;
  __syncthreads();
  const int cg = tid % c8, rl = tid / c8;
  double a[8], q[8];
  float mu[8], is[8], sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = q[k] = 0.0;
    mu[k] = mean[cg * 8 + k];
    is[k] = invstd[cg * 8 + k];
    sc[k] = scale[cg * 8 + k];
    sh[k] = shift[cg * 8 + k];
  }
  const size_t rstride = (size_t)gridDim.x * lanes;
  for (size_t r = (size_t)blockIdx.x * lanes + rl; r < P; r += 4 * rstride) {
    uint4 vg[4], vz[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t ru = r + u * rstride;
      const bool in = ru < P;
      vg[u] = in ? *reinterpret_cast<const uint4*>(dy + ru * C + cg * 8) : make_uint4(0, 0, 0, 0);
      vz[u] = in ? *reinterpret_cast<const uint4*>(z + ru * C + cg * 8) : make_uint4(0, 0, 0, 0);
    }
    float sa[8], sq[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) sa[k] = sq[k] = 0.f;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float g[8], zz[8];
      unpack8(vg[u], g);   // rows beyond P carry dY = 0 and add nothing
      unpack8(vz[u], zz);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float gm = fmaf(zz[k], sc[k], sh[k]) > 0.f ? g[k] : 0.f;
        sa[k] += gm;
        sq[k] = fmaf(gm, (zz[k] - mu[k]) * is[k], sq[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a[k] += (double)sa[k];
      q[k] += (double)sq[k];
    }
  }
  block_channel_sums(a, q, s_acc, C, c8, cg, part);
}

// dbeta, dgamma out (fp32) and the two per-channel terms of dZ.  One warp per channel.
__global__ void __launch_bounds__(256) k_bn_bwd_finalize(const double* __restrict__ part, int nb, int C, double n, const float* gamma,
                                                         const float* invstd, float* dgamma, float* dbeta, float* k_scale,
                                                         float* k_mean_db, float* k_mean_dg) {
  const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  double s0, s1;
  reduce_partials(part, nb, C, c, lane, s0, s1);
  if (lane) return;
  dbeta[c] = (float)s0;
  dgamma[c] = (float)s1;
  k_scale[c] = gamma[c] * invstd[c];
  k_mean_db[c] = (float)(s0 / n);
  k_mean_dg[c] = (float)(s1 / n);
}

__global__ void __launch_bounds__(256) k_bn_bwd_apply(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ scale,
                                                      const float* __restrict__ shift,
                                                      const __nv_bfloat16* __restrict__ z, size_t n8, int c8, const float* __restrict__ mean,
                                                      const float* __restrict__ invstd, const float* __restrict__ k_scale,
                                                      const float* __restrict__ k_mean_db, const float* __restrict__ k_mean_dg,
                                                      __nv_bfloat16* __restrict__ dz) {
  // one 8-channel group per thread for the whole walk (c8 divides the grid stride): the seven per-channel terms in registers
  const int cg = threadIdx.x % c8;
  float sc[8], sh[8], mu[8], is[8], ks[8], mdb[8], mdg[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cg * 8 + k;
    sc[k] = scale[c];
    sh[k] = shift[c];
    mu[k] = mean[c];
    is[k] = invstd[c];
    ks[k] = k_scale[c];
    mdb[k] = k_mean_db[c];
    mdg[k] = k_mean_dg[c];
  }
  constexpr int U = 4;  // elements in flight per thread (two 16-byte loads each)
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i0 < n8; i0 += U * stride) {
    uint4 gv[U], zv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + u * stride;
      if (i < n8) {
        gv[u] = reinterpret_cast<const uint4*>(dy)[i];
        zv[u] = reinterpret_cast<const uint4*>(z)[i];
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + u * stride;
      if (i >= n8) break;
      float g[8], zz[8], o[8];
      unpack8(gv[u], g);
      unpack8(zv[u], zz);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float gm = fmaf(zz[k], sc[k], sh[k]) > 0.f ? g[k] : 0.f;
        const float xh = (zz[k] - mu[k]) * is[k];
        o[k] = ks[k] * (gm - mdb[k] - xh * mdg[k]);
      }
      reinterpret_cast<uint4*>(dz)[i] = pack8(o);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// adjoints of the shape-changing ops
// ---------------------------------------------------------------------------------------------------------
// MaxPool2d(2) backward: the gradient of a pooled pixel goes to the first maximum of its window in row-major order
// (ATen's max_pool2d keeps the first index where `val > max`); rows / columns beyond the last full window get none.
// `accumulate`: add to what dst already holds (the skip connection's other consumer wrote it before).
__global__ void __launch_bounds__(256) k_maxpool_bwd(const __nv_bfloat16* __restrict__ src_y, const __nv_bfloat16* __restrict__ dpool, int B,
                                                     int H, int W, int C8, int accumulate, __nv_bfloat16* __restrict__ dst) {
  // A thread owns one 2x2 window (8 channels): the four inputs, the pooled gradient and the four destinations are each
  // read once.  Rows / columns beyond the last full window (odd H or W) get no gradient: handled by the tail loop below.
  // grid = (ceil(Wo * C8 / 256), Ho, B)
  const int Ho = H / 2, Wo = W / 2;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t < Wo * C8 && (int)blockIdx.y < Ho) {
    const int xo = t / C8, cg = t - xo * C8;
    const int yo = blockIdx.y, n = blockIdx.z;
    const size_t i = (((size_t)n * Ho + yo) * Wo + xo) * C8 + cg;
    size_t at[4];
    uint4 yv[4], dv[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      at[q] = (((size_t)n * H + 2 * yo + (q >> 1)) * W + 2 * xo + (q & 1)) * C8 + cg;
      yv[q] = reinterpret_cast<const uint4*>(src_y)[at[q]];
      dv[q] = accumulate ? reinterpret_cast<const uint4*>(dst)[at[q]] : make_uint4(0, 0, 0, 0);
    }
    float w4[4][8], o[4][8], g[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      unpack8(yv[q], w4[q]);
      unpack8(dv[q], o[q]);
    }
    unpack8(reinterpret_cast<const uint4*>(dpool)[i], g);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      int best = 0;
      float m = w4[0][k];
#pragma unroll
      for (int q = 1; q < 4; ++q)
        if (w4[q][k] > m) {
          m = w4[q][k];
          best = q;
        }
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (best == q) o[q][k] += g[k];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) reinterpret_cast<uint4*>(dst)[at[q]] = pack8(o[q]);
  }
  if (!accumulate && ((H | W) & 1)) {  // the uncovered last row / column of an odd-sized tensor: zero gradient
    const size_t total = (size_t)B * H * W * C8;
    const size_t nthreads = (size_t)gridDim.x * gridDim.y * gridDim.z * 256;
    for (size_t i = (((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 256 + threadIdx.x; i < total; i += nthreads) {
      size_t r = i / C8;
      const int x = (int)(r % W);
      r /= W;
      const int y = (int)(r % H);
      if ((y >> 1) >= Ho || (x >> 1) >= Wo) reinterpret_cast<uint4*>(dst)[i] = make_uint4(0, 0, 0, 0);
    }
  }
}

// Adjoint of `Upsample(x2, bilinear, align_corners=True)` + the zero pad of Up.forward, as a gather: a low-resolution
// pixel collects, with the forward pass's own weights (same fp32 expressions as k_upsample2x), every output pixel that
// read it.  src = gradient of the padded, up-sampled tensor: channels [c_off, c_off + 8*C8) of rows of `src_c` channels.
__global__ void __launch_bounds__(256) k_upsample_bwd(const __nv_bfloat16* __restrict__ src, int src_c, int c_off, int B, int h, int w,
                                                      int C8, int Hs, int Ws, int py, int px, __nv_bfloat16* __restrict__ dst) {
  // grid = (ceil(w * C8 / 256), h, B): no 64-bit divisions on the streaming path
  const int Ho = 2 * h, Wo = 2 * w;
  const float ry = Ho > 1 ? (float)(h - 1) / (float)(Ho - 1) : 0.f, rx = Wo > 1 ? (float)(w - 1) / (float)(Wo - 1) : 0.f;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t < w * C8) {
    const int xl = t / C8, cg = t - xl * C8;
    const int yl = blockIdx.y, n = blockIdx.z;
    const size_t i = (((size_t)n * h + yl) * w + xl) * C8 + cg;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    // output rows whose source interval touches yl: fy = ry * yo in (yl - 1, yl + 1); scan a safe window of eight rows
    // and eight columns, test each exactly with the forward pass's own expressions, ONCE per row and per column
    const int y_lo = 2 * yl - 3, x_lo = 2 * xl - 3;
    float wy[8], wx[8];
#pragma unroll
    for (int a = 0; a < 8; ++a) {
      const int yo = y_lo + a, xo = x_lo + a;
      wy[a] = 0.f;
      wx[a] = 0.f;
      if (yo >= 0 && yo < Ho && yo + py >= 0 && yo + py < Hs) {
        const float fy = ry * (float)yo;
        const int y0 = (int)fy, y1 = y0 + (y0 < h - 1 ? 1 : 0);
        const float ly1 = fy - (float)y0, ly0 = 1.f - ly1;
        wy[a] = (y0 == yl ? ly0 : 0.f) + (y1 == yl ? ly1 : 0.f);
      }
      if (xo >= 0 && xo < Wo && xo + px >= 0 && xo + px < Ws) {
        const float fx = rx * (float)xo;
        const int x0 = (int)fx, x1 = x0 + (x0 < w - 1 ? 1 : 0);
        const float lx1 = fx - (float)x0, lx0 = 1.f - lx1;
        wx[a] = (x0 == xl ? lx0 : 0.f) + (x1 == xl ? lx1 : 0.f);
      }
    }
#pragma unroll
    for (int a = 0; a < 8; ++a) {
      if (wy[a] == 0.f) continue;
      const int ys = y_lo + a + py;
      const __nv_bfloat16* rowp = src + ((size_t)n * Hs + ys) * Ws * src_c + c_off + cg * 8;
      uint4 v[8];
#pragma unroll
      for (int b = 0; b < 8; ++b)   // the row's loads first (up to four of them), then the arithmetic: loads in flight together
        if (wx[b] != 0.f) v[b] = *reinterpret_cast<const uint4*>(rowp + (size_t)(x_lo + b + px) * src_c);
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        if (wx[b] == 0.f) continue;
        float g[8];
        unpack8(v[b], g);
        const float wgt = wy[a] * wx[b];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(wgt, g[k], acc[k]);
      }
    }
    reinterpret_cast<uint4*>(dst)[i] = pack8(acc);
  }
}

// dst[p][0..C) (+)= src[p][c_off .. c_off + C): the skip half of a concatenated gradient
__global__ void __launch_bounds__(256) k_slice_channels(const __nv_bfloat16* __restrict__ src, int src_c8, int off8, size_t P, int C8,
                                                        __nv_bfloat16* __restrict__ dst) {
  const size_t total = P * C8;
  const bool pow2 = (C8 & (C8 - 1)) == 0;
  const int sh = __ffs(C8) - 1;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t p = pow2 ? i >> sh : i / C8;   // (a 64-bit division costs more than the copy itself)
    const int cg = (int)(i - p * C8);
    reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[p * src_c8 + off8 + cg];
  }
}

// ---------------------------------------------------------------------------------------------------------
// weight gradient: tcgen05 GEMM with K = pixels, both operands straight from the NHWC tensors
// ---------------------------------------------------------------------------------------------------------
// dW[co][ci][tap] = sum over pixels p of dZ[p][co] * X[p + tap][ci].  The contraction index is the pixel, and in NHWC
// the pixel is the STRIDED index of both operands: exactly the "MN-major" operand form of tcgen05.mma (instruction
// descriptor bits 15/16; shared-memory descriptor: 64 channels = one 128-byte swizzled row per pixel, eight pixel rows
// per group, SBO = distance between groups along K, LBO = distance between 64-channel blocks along M / N).  So the
// tiles arrive by TMA exactly as the convolution loads them - no transposed copies - and a tap is a shifted view:
//   * one pipeline stage = a patch of 16 x 8 pixels (K = 128): dZ as 64-channel blocks of 128 pixel rows (16 KB each),
//     X as 64-channel blocks of an 18 x 8 halo (row y + dy of the tap row, columns x0 - 1 .. x0 + 16; TMA zero fill
//     outside the image = the convolution's padding);
//   * one MMA (K = 16) = one patch row; its B operand for the horizontal tap dx starts (dx + 1) pixels = 128 bytes
//     further into the halo row;
//   * a CTA owns one vertical tap dy and the three horizontal ones: three accumulators of 128 (co) x N (ci) in
//     tensor memory (N = 128: 384 of the 512 columns), so dZ and the halo are loaded once for three taps;
//   * split-K over the patches; partial sums go to dW with fp32 atomics.
// Work items that share pixels (the three dy, the ci tiles, the co tiles) have neighbouring block indices and run
// together, so HBM sees each tensor once and L2 serves the other readers.
constexpr int WG_PW = 16, WG_PH = 8;                      // patch: 16 wide, 8 tall
constexpr int WG_HALO_W = WG_PW + 2;
constexpr int WG_DZ_BLOCK = WG_PW * WG_PH * 128;          // 16 KB: 128 pixels x 64 output channels
constexpr int WG_X_BLOCK = WG_HALO_W * WG_PH * 128;       // 18 KB: 18 x 8 pixels x 64 input channels
constexpr int WG_MAX_STAGES = 8;
constexpr int WG_THREADS = 192;  // warp 0: TMA, warp 1: MMA, warps 2-5: epilogue
constexpr int WG_SMEM_BUDGET = 225 * 1024;
struct WgradParams {
  int B, H, W;
  int cout, cin_real, c0;       // the input is src0[c0 channels] ++ src1[the rest]
  int co_blocks;                // 64-channel blocks of dZ per M tile: 2, or 1 when cout = 64 (the upper 64 rows repeat the lower)
  int m_tiles, n_tiles, dys, ksplit, stages;
  int px_tiles, py_tiles;       // patches along x / y
  int first_layer;              // the input is the im2col'ed image: column k = tap * cin_real + c  ->  dW[co][c][tap]; one "tap"
  float* dw;                    // fp32 scratch [9 * cin_real][cout] (k = tap * cin_real + ci), see k_dw_layout
};

template <int N>
__global__ void __launch_bounds__(WG_THREADS, 1)
k_wgrad(const __grid_constant__ CUtensorMap map_dz, const __grid_constant__ CUtensorMap map_x0, const __grid_constant__ CUtensorMap map_x1,
        const WgradParams p) {
  constexpr int B_BYTES = (N / 64) * WG_X_BLOCK;
  constexpr uint32_t TMEM_COLS = N == 64 ? 256 : 512;  // three accumulators of N columns
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int a_bytes = p.co_blocks * WG_DZ_BLOCK, stage_bytes = a_bytes + B_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.stages * stage_bytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + WG_MAX_STAGES;
  uint64_t* done = bars + 2 * WG_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int item = blockIdx.x;
  const int dyi = item % p.dys; item /= p.dys;
  const int nt = item % p.n_tiles; item /= p.n_tiles;
  const int mt = item % p.m_tiles;
  const int ks = item / p.m_tiles;
  const int dy = p.dys == 3 ? dyi - 1 : 0;
  const int ntaps = p.dys == 3 ? 3 : 1;
  const long long total = (long long)p.B * p.py_tiles * p.px_tiles;
  const long long s0 = total * ks / p.ksplit, s1 = total * (ks + 1) / p.ksplit;

  if (threadIdx.x == 0) {
    prefetch_tensormap(&map_dz);
    prefetch_tensormap(&map_x0);
    prefetch_tensormap(&map_x1);
    for (int s = 0; s < WG_MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (long long s = s0; s < s1; ++s) {
        const int px = (int)(s % p.px_tiles);
        long long r = s / p.px_tiles;
        const int py = (int)(r % p.py_tiles), n = (int)(r / p.py_tiles);
        const int x0 = px * WG_PW, y0 = py * WG_PH;
        uint8_t* st = smem + stage * stage_bytes;
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full[stage], stage_bytes);
        for (int cb = 0; cb < p.co_blocks; ++cb)
          tma_load_4d(st + cb * WG_DZ_BLOCK, &map_dz, (mt * 2 + cb) * 64, x0, y0, n, &full[stage]);
#pragma unroll
        for (int jb = 0; jb < N / 64; ++jb) {
          const int c = nt * N + jb * 64;
          if (c < p.c0) tma_load_4d(st + a_bytes + jb * WG_X_BLOCK, &map_x0, c, x0 - 1, y0 + dy, n, &full[stage]);
          else tma_load_4d(st + a_bytes + jb * WG_X_BLOCK, &map_x1, c - p.c0, x0 - 1, y0 + dy, n, &full[stage]);
        }
        if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // the whole warp walks the loop converged and one elected lane issues: descriptors stay in uniform registers
    constexpr uint32_t idesc = make_idesc_bf16(128, N) | (1u << 15) | (1u << 16);  // A and B MN-major
    constexpr uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);                // SBO = 1024 bytes, version 1, 128-byte swizzle
    const uint32_t lbo_a = (p.co_blocks == 2 ? (uint32_t)WG_DZ_BLOCK >> 4 : 0u) << 16;
    constexpr uint32_t lbo_b = ((uint32_t)WG_X_BLOCK >> 4) << 16;
    uint32_t stage = 0, phase = 0;
    for (long long s = s0; s < s1; ++s) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      const uint32_t a0 = smem_u32(smem + stage * stage_bytes);
      const uint32_t a_lo = (a0 >> 4) | lbo_a, b_lo = ((a0 + a_bytes) >> 4) | lbo_b;
      const uint32_t first = s != s0;
      if (ntaps == 3) {
#pragma unroll
        for (int r = 0; r < WG_PH; ++r) {
          const uint64_t da = ((uint64_t)hi << 32) | (a_lo + r * (WG_PW * 128 >> 4));
#pragma unroll
          for (int t = 0; t < 3; ++t)
            mma_bf16_ss_elect(tmem_base + t * N, da, ((uint64_t)hi << 32) | (b_lo + (r * WG_HALO_W + t) * 8), idesc, r != 0 ? 1u : first);
        }
      } else {
#pragma unroll
        for (int r = 0; r < WG_PH; ++r)
          mma_bf16_ss_elect(tmem_base, ((uint64_t)hi << 32) | (a_lo + r * (WG_PW * 128 >> 4)),
                            ((uint64_t)hi << 32) | (b_lo + (r * WG_HALO_W + 1) * 8), idesc, r != 0 ? 1u : first);
      }
      tc_commit_elect(&empty[stage]);
      if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
    }
    tc_commit_elect(done);
  } else if (s1 > s0) {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int co = mt * 128 + row;
    mbar_wait(done, 0);
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const bool row_ok = co < p.cout && (p.co_blocks == 2 || row < 64);
    for (int t = 0; t < ntaps; ++t) {
      const int tap = (dy + 1) * 3 + t;
#pragma unroll 1
      for (int c = 0; c < N / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + t * N + c * 32, r);
        tmem_ld_wait();
        if (row_ok) {
          // scratch layout [k][co], k = tap * cin + ci (the first layer's im2col column is that k already): the lanes of
          // a warp are consecutive co, so one warp-level reduction is one 128-byte line
          const int k_lim = p.first_layer ? 9 * p.cin_real : p.cin_real;
          float* dst = p.dw + (size_t)(p.first_layer ? 0 : tap * p.cin_real) * p.cout + co;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int ci = nt * N + c * 32 + j;
            if (ci < k_lim) atomicAdd(dst + (size_t)ci * p.cout, __uint_as_float(r[j]));
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// Conv2d.weight fp32 [cout][cin][3][3] -> the two bf16 operand layouts of RdpTrainLayer: w [cout][tap][cin] (first layer:
// [cout][64] with k = tap * cin + c, zero padded) and w_t [cin][8 - tap][cout] (the flipped, transposed kernel of the
// input gradient).  One thread per weight; a block of the first layer also clears that layer's padding columns.
__global__ void __launch_bounds__(256) k_pack_train_weights(const float* __restrict__ w, int cout, int cin, int first,
                                                            __nv_bfloat16* __restrict__ fwd, __nv_bfloat16* __restrict__ wt) {
  const int total = cout * cin * 9;
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e < total) {
    const int co = e / (cin * 9), rem = e - co * cin * 9, ci = rem / 9, tap = rem - ci * 9;
    const __nv_bfloat16 v = __float2bfloat16_rn(w[e]);
    if (first) fwd[co * 64 + tap * cin + ci] = v;
    else fwd[((size_t)co * 9 + tap) * cin + ci] = v;
    if (wt) wt[((size_t)ci * 9 + (8 - tap)) * cout + co] = v;
  }
  if (first) {
    const int pad = 64 - 9 * cin;
    for (int i = e; i < cout * pad; i += gridDim.x * 256) fwd[(i / pad) * 64 + 9 * cin + i % pad] = __float2bfloat16_rn(0.f);
  }
}

// The same two layouts for every layer but the first, through shared memory so that both sides are contiguous: a block
// moves 32 output channels x 32 input channels x 9 taps (the source rows [co][ci0 .. ci0+31][9] are 288 consecutive floats).
// grid = (cin / 32, cout / 32); cin and cout are multiples of 64 here.
__global__ void __launch_bounds__(256) k_pack_train_weights_tiled(const float* __restrict__ w, int cout, int cin, __nv_bfloat16* __restrict__ fwd,
                                                                  __nv_bfloat16* __restrict__ wt) {
  __shared__ __nv_bfloat16 tile[32][9 * 32 + 2];  // [co][ci * 9 + tap]
  const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 32;
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  for (int r = wp; r < 32; r += 8) {
    const float* src = w + ((size_t)(co0 + r) * cin + ci0) * 9;
#pragma unroll
    for (int e = lane; e < 288; e += 32) tile[r][e] = __float2bfloat16_rn(src[e]);
  }
  __syncthreads();
  // fwd[co][tap][ci]: 32 consecutive input channels per (co, tap)
  for (int r = wp; r < 32 * 9; r += 8) {
    const int co = r / 9, tap = r - co * 9;
    fwd[((size_t)(co0 + co) * 9 + tap) * cin + ci0 + lane] = tile[co][lane * 9 + tap];
  }
  // wt[ci][8 - tap][co]: 32 consecutive output channels per (ci, tap)
  for (int r = wp; r < 32 * 9; r += 8) {
    const int ci = r / 9, tap = r - ci * 9;
    wt[((size_t)(ci0 + ci) * 9 + (8 - tap)) * cout + co0 + lane] = tile[lane][ci * 9 + tap];
  }
}

// scratch [9 * cin][cout] (what the GEMM epilogue can add to with coalesced reductions) -> the parameter's layout
// [cout][cin][3][3].  A block moves 32 output channels x 32 input channels x 9 taps through shared memory so that both
// sides are contiguous.  grid = (ceil(cin / 32), ceil(cout / 32)).
__global__ void __launch_bounds__(256) k_dw_layout(const float* __restrict__ src, int cin, int cout, float* __restrict__ dst) {
  __shared__ float tile[32][9 * 32 + 1];  // [co][ci * 9 + tap]
  const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 32;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int co_l = co0 + lane;
#pragma unroll 1
  for (int r0 = w; r0 < 9 * 32; r0 += 8 * 6) {  // r = tap * 32 + ci; six rows of the scratch in flight per warp
    float v[6];
#pragma unroll
    for (int u = 0; u < 6; ++u) {
      const int r = r0 + 8 * u, tap = r >> 5, ci = ci0 + (r & 31);
      v[u] = (r < 9 * 32 && ci < cin && co_l < cout) ? src[((size_t)tap * cin + ci) * cout + co_l] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 6; ++u) {
      const int r = r0 + 8 * u;
      if (r < 9 * 32) tile[lane][(r & 31) * 9 + (r >> 5)] = v[u];
    }
  }
  __syncthreads();
  const int span = min(32, cin - ci0) * 9;
  for (int r = w; r < 32; r += 8) {
    const int co = co0 + r;
    if (co >= cout) break;
    for (int e = lane; e < span; e += 32) dst[((size_t)co * cin + ci0) * 9 + e] = tile[r][e];
  }
}

// The same GEMM for the layers with 64 output channels (the four full-resolution layers: their pixel count makes them the
// long ones, and in k_wgrad they fill only half of the M = 128 rows), with the operands' roles exchanged and ALL NINE taps
// in one CTA:
//   * M = input channels, N = the 64 output channels; one 64-channel block of X per CTA (a 128-channel input is two
//     independent problems, blockIdx selects the block);
//   * M = 128 holds TWO TAPS of that block: the MN-major descriptor's block stride (LBO) is free, so the second 64 rows
//     are the same halo entered one tap further (128 bytes for the next column, 16 pixels = 2048 bytes from the last
//     column of a halo row to the first of the next).  Taps (0,1) (2,3) (4,5) (6,7) share an MMA each, tap 8 takes a
//     fifth whose upper rows repeat the lower and are dropped: five MMAs per patch row for nine taps;
//   * one stage = dZ of a 16x8 patch (16 KB) + the 18x10 halo of X (22.5 KB): 39 KB from L2 for nine taps, where three
//     CTAs with one tap row each moved 102 KB - these layers were bound by L2 throughput, not by the tensor pipe.
// D rows are (tap, input channel), columns the output channels: a thread adds its 64 contiguous floats of the [k][co]
// scratch with 16-byte vector reductions.  The im2col'ed first layer is the one-tap case (halo row 1, column 1).
constexpr int WG9_HALO_H = WG_PH + 2;
constexpr int WG9_X_BYTES = WG_HALO_W * WG9_HALO_H * 128;                 // 23,040
constexpr int WG9_X_STRIDE = (WG9_X_BYTES + 1023) / 1024 * 1024;          // 23,552: the next stage starts 1024-byte aligned
constexpr int WG9_STAGE = WG_DZ_BLOCK + WG9_X_STRIDE;
__global__ void __launch_bounds__(WG_THREADS, 1)
k_wgrad_co64(const __grid_constant__ CUtensorMap map_dz, const __grid_constant__ CUtensorMap map_x0, const __grid_constant__ CUtensorMap map_x1,
             const WgradParams p) {
  constexpr uint32_t TMEM_COLS = 512;  // five accumulators of 64 columns
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.stages * WG9_STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + WG_MAX_STAGES;
  uint64_t* done = bars + 2 * WG_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const int nt = blockIdx.x % p.n_tiles;   // 64-channel block of the input
  const int ks = blockIdx.x / p.n_tiles;
  const bool one_tap = p.first_layer != 0;   // a 1x1 convolution over the im2col'ed image
  const long long total = (long long)p.B * p.py_tiles * p.px_tiles;
  const long long s0 = total * ks / p.ksplit, s1 = total * (ks + 1) / p.ksplit;

  if (threadIdx.x == 0) {
    prefetch_tensormap(&map_dz);
    prefetch_tensormap(&map_x0);
    prefetch_tensormap(&map_x1);
    for (int s = 0; s < WG_MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      const int c = nt * 64;
      for (long long s = s0; s < s1; ++s) {
        const int px = (int)(s % p.px_tiles);
        long long r = s / p.px_tiles;
        const int py = (int)(r % p.py_tiles), n = (int)(r / p.py_tiles);
        const int x0 = px * WG_PW, y0 = py * WG_PH;
        uint8_t* st = smem + stage * WG9_STAGE;
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full[stage], WG_DZ_BLOCK + WG9_X_BYTES);
        tma_load_4d(st, &map_dz, 0, x0, y0, n, &full[stage]);
        if (c < p.c0) tma_load_4d(st + WG_DZ_BLOCK, &map_x0, c, x0 - 1, y0 - 1, n, &full[stage]);
        else tma_load_4d(st + WG_DZ_BLOCK, &map_x1, c - p.c0, x0 - 1, y0 - 1, n, &full[stage]);
        if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc_bf16(128, 64) | (1u << 15) | (1u << 16);  // A and B MN-major
    constexpr uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);                 // SBO = 1024 bytes, version 1, 128-byte swizzle
    constexpr uint32_t lbo_next_col = (128u >> 4) << 16;                            // second tap = one pixel on
    constexpr uint32_t lbo_next_row = ((uint32_t)(WG_HALO_W - 2) * 128u >> 4) << 16;  // from column 2 of a halo row to column 0 of the next
    uint32_t stage = 0, phase = 0;
    for (long long s = s0; s < s1; ++s) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      const uint32_t z0 = smem_u32(smem + stage * WG9_STAGE);
      const uint32_t z_lo = z0 >> 4, x_lo = (z0 + WG_DZ_BLOCK) >> 4;
      const uint32_t first = s != s0;
#pragma unroll
      for (int r = 0; r < WG_PH; ++r) {
        const uint64_t db = ((uint64_t)hi << 32) | (z_lo + r * (WG_PW * 128 >> 4));   // dZ: N = 64 output channels
        const uint32_t acc = r != 0 ? 1u : first;
        // halo pixel (row, col) in 16-byte units: ((r + dy + 1) * 18 + dx + 1) * 8
        auto view = [&](int tap) { return x_lo + ((r + tap / 3) * WG_HALO_W + tap % 3) * 8; };
        if (one_tap) {
          mma_bf16_ss_elect(tmem_base, ((uint64_t)hi << 32) | view(4), db, idesc, acc);   // the centre tap only; upper rows repeat
        } else {
          mma_bf16_ss_elect(tmem_base + 0 * 64, ((uint64_t)hi << 32) | (view(0) | lbo_next_col), db, idesc, acc);  // taps 0, 1
          mma_bf16_ss_elect(tmem_base + 1 * 64, ((uint64_t)hi << 32) | (view(2) | lbo_next_row), db, idesc, acc);  // taps 2, 3
          mma_bf16_ss_elect(tmem_base + 2 * 64, ((uint64_t)hi << 32) | (view(4) | lbo_next_col), db, idesc, acc);  // taps 4, 5
          mma_bf16_ss_elect(tmem_base + 3 * 64, ((uint64_t)hi << 32) | (view(6) | lbo_next_col), db, idesc, acc);  // taps 6, 7
          mma_bf16_ss_elect(tmem_base + 4 * 64, ((uint64_t)hi << 32) | view(8), db, idesc, acc);                   // tap 8; upper rows repeat
        }
      }
      tc_commit_elect(&empty[stage]);
      if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
    }
    tc_commit_elect(done);
  } else if (s1 > s0) {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int ci = nt * 64 + (row & 63);
    mbar_wait(done, 0);
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const int naccs = one_tap ? 1 : 5;
#pragma unroll 1
    for (int a = 0; a < naccs; ++a) {
      // which row k = tap * cin + ci of the scratch this row of accumulator `a` is; k < 0: the row carries nothing
      int k = -1;
      if (one_tap) {
        if (row < 64 && row < 9 * p.cin_real) k = row;            // the im2col column is k = tap * cin + c already
      } else if (ci < p.cin_real) {
        const int tap = 2 * a + (row >> 6);
        if (tap < 9 && (a < 4 || row < 64)) k = tap * p.cin_real + ci;
      }
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + a * 64 + c * 32, r);
        tmem_ld_wait();
        if (k >= 0) {
          float4* dst = reinterpret_cast<float4*>(p.dw + (size_t)k * p.cout + c * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            atomicAdd(dst + j, make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                           __uint_as_float(r[4 * j + 3])));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}
// map over an NHWC bf16 tensor: dims (C, W, H, B), box (64 channels, box_w pixels, box_h rows, 1), 128-byte swizzle,
// zero fill outside
static int make_nhwc_map(CUtensorMap* m, const void* ptr, int B, int H, int W, int C, int box_w, int box_h = WG_PH) {
  EncodeTiledFn fn = encode_fn();
  RDP_REQUIRE(fn, RDP_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_w, (cuuint32_t)box_h, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RDP_REQUIRE(r == CUDA_SUCCESS, RDP_ERR_CUDA, "cuTensorMapEncodeTiled(NHWC %dx%dx%dx%d) failed: %d", B, H, W, C, (int)r);
  return RDP_OK;
}

template <int N>
static int launch_wgrad_t(const CUtensorMap& mz, const CUtensorMap& mx0, const CUtensorMap& mx1, WgradParams p, cudaStream_t stream) {
  const int stage_bytes = p.co_blocks * WG_DZ_BLOCK + (N / 64) * WG_X_BLOCK;
  p.stages = std::min(WG_MAX_STAGES, (WG_SMEM_BUDGET - 2048) / stage_bytes);
  const int smem_bytes = p.stages * stage_bytes + 1024 + 256;
  static bool set[64] = {};
  int dev = 0;
  RDP_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 64 && !set[dev]) {
    RDP_CUDA_TRY(cudaFuncSetAttribute(k_wgrad<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_BUDGET));
    set[dev] = true;
  }
  k_wgrad<N><<<p.m_tiles * p.n_tiles * p.dys * p.ksplit, WG_THREADS, smem_bytes, stream>>>(mz, mx0, mx1, p);
  return RDP_OK;
}

static int launch_wgrad_co64(const CUtensorMap& mz, const CUtensorMap& mx0, const CUtensorMap& mx1, WgradParams p, cudaStream_t stream) {
  p.stages = std::min(WG_MAX_STAGES, (WG_SMEM_BUDGET - 2048) / WG9_STAGE);
  const int smem_bytes = p.stages * WG9_STAGE + 1024 + 256;
  static bool set[64] = {};
  int dev = 0;
  RDP_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 64 && !set[dev]) {
    RDP_CUDA_TRY(cudaFuncSetAttribute(k_wgrad_co64, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_BUDGET));
    set[dev] = true;
  }
  k_wgrad_co64<<<p.n_tiles * p.ksplit, WG_THREADS, smem_bytes, stream>>>(mz, mx0, mx1, p);
  return RDP_OK;
}

}  // namespace train
}  // namespace rdp

using namespace rdp;
using namespace rdp::train;

namespace {

const int kLevel[18] = {0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 3, 3, 2, 2, 1, 1, 0, 0};

struct Shape { int H, W; };

// Everything one forward/backward pair keeps in the workspace.
constexpr int STAT_MAX_BLOCKS = 1184;  // rows of the partial-sum table (8 CTAs on each of 148 SMs)

struct TrainBuffers {
  Shape lv[5];
  __nv_bfloat16* col;        // im2col of the input (64 channels)
  __nv_bfloat16* z[18];      // conv outputs before BatchNorm
  __nv_bfloat16* y[18];      // after BatchNorm + ReLU
  __nv_bfloat16* pool[4];    // pooled x1..x4
  __nv_bfloat16* up[4];      // up-sampled + padded low branch of up1..up4
  float* zero_shift;         // 1024 zeros
  float *mean[18], *var[18], *invstd[18], *scale[18], *shift[18];
  double* part;              // [2 * C][blocks] per-block partial sums of the BatchNorm reductions (room for STAT_MAX_BLOCKS x 2048)
  // backward scratch
  __nv_bfloat16* dy[18];     // gradient wrt y[i]
  __nv_bfloat16* dz;         // gradient wrt the current layer's z (largest layer)
  __nv_bfloat16* dx;         // gradient wrt the current layer's (concatenated) input
  __nv_bfloat16* dpool;      // gradient wrt a pooled tensor (alias of dx is not possible: both live)
  float *kscale, *kdb, *kdg;  // [1024] each
  float* dws;                 // weight-gradient scratch of the current layer, [9 * cin][cout]
  size_t bytes;
};

TrainBuffers carve_train(const RdpTrainWeights* w, void* ws, int B, int H, int W) {
  TrainBuffers b{};
  b.lv[0] = {H, W};
  for (int i = 1; i < 5; ++i) b.lv[i] = {b.lv[i - 1].H / 2, b.lv[i - 1].W / 2};
  Carver c(ws);
  auto px = [&](int l) { return (size_t)B * b.lv[l].H * b.lv[l].W; };
  b.col = c.take<__nv_bfloat16>(px(0) * 64);
  size_t max_act = 0, max_in = 0, max_w = 0;
  for (int i = 0; i < 18; ++i) {
    const size_t n = px(kLevel[i]) * w->conv[i].cout;
    b.z[i] = c.take<__nv_bfloat16>(n);
    b.y[i] = c.take<__nv_bfloat16>(n);
    b.dy[i] = c.take<__nv_bfloat16>(n);
    max_act = std::max(max_act, n);
    const int cin_pad = i == 0 ? 64 : (w->conv[i].cin + 63) / 64 * 64;
    max_in = std::max(max_in, px(kLevel[i]) * cin_pad);
    max_w = std::max(max_w, (size_t)9 * w->conv[i].cin * w->conv[i].cout);
    b.mean[i] = c.take<float>(w->conv[i].cout);
    b.var[i] = c.take<float>(w->conv[i].cout);
    b.invstd[i] = c.take<float>(w->conv[i].cout);
    b.scale[i] = c.take<float>(w->conv[i].cout);
    b.shift[i] = c.take<float>(w->conv[i].cout);
  }
  for (int i = 0; i < 4; ++i) b.pool[i] = c.take<__nv_bfloat16>(px(i + 1) * w->conv[2 * i + 1].cout);
  for (int j = 0; j < 4; ++j) b.up[j] = c.take<__nv_bfloat16>(px(3 - j) * w->conv[9 + 2 * j].cout);
  b.zero_shift = c.take<float>(1024);
  b.part = c.take<double>((size_t)STAT_MAX_BLOCKS * 2048);
  b.dz = c.take<__nv_bfloat16>(max_act);
  b.dx = c.take<__nv_bfloat16>(max_in);
  b.dpool = c.take<__nv_bfloat16>(max_act);
  b.kscale = c.take<float>(1024);
  b.kdb = c.take<float>(1024);
  b.kdg = c.take<float>(1024);
  b.dws = c.take<float>(max_w);
  b.bytes = c.total();
  return b;
}

// CTAs of a BatchNorm reduction: one resident wave (`per_sm` = what the kernel's registers allow) on the large layers,
// never fewer than 8 rows per thread; every CTA costs the finalize kernel one row of the table
int stat_blocks(size_t P, int C, int num_sms, int per_sm) {
  const int lanes = ST_THREADS / (C / 8);
  const size_t by_rows = (P + (size_t)lanes * 8 - 1) / ((size_t)lanes * 8);
  return (int)std::max<size_t>(1, std::min<size_t>(by_rows, std::min<size_t>((size_t)num_sms * per_sm, STAT_MAX_BLOCKS)));
}

int ew_grid(size_t items) { return (int)std::min<size_t>((items + 255) / 256, (size_t)148 * 16); }

int check_train_args(const RdpTrainWeights* w, int B, int H, int W, const char* who) {
  RDP_REQUIRE(w, RDP_ERR_BAD_ARGUMENT, "%s: null weights", who);
  RDP_REQUIRE(B > 0 && H >= 16 && W >= 16 && H % 16 == 0 && W % 16 == 0, RDP_ERR_UNSUPPORTED,
              "%s: training needs H and W divisible by 16 (got %dx%dx%d; the reference trains at 256x256)", who, B, H, W);
  RDP_REQUIRE(w->n_channels >= 1 && 9 * w->n_channels <= 64 && w->n_classes >= 1 && w->n_classes <= 8, RDP_ERR_UNSUPPORTED,
              "%s: n_channels %d / n_classes %d unsupported", who, w->n_channels, w->n_classes);
  static const int cin[18] = {0, 64, 64, 128, 128, 256, 256, 512, 512, 512, 1024, 512, 512, 256, 256, 128, 128, 64};
  static const int cout[18] = {64, 64, 128, 128, 256, 256, 512, 512, 512, 512, 512, 256, 256, 128, 128, 64, 64, 64};
  for (int i = 0; i < 18; ++i) {
    const RdpTrainLayer& L = w->conv[i];
    RDP_REQUIRE(L.w && L.gamma && L.beta && (i == 0 || L.w_t) && L.cout == cout[i] && L.cin == (i == 0 ? w->n_channels : cin[i]),
                RDP_ERR_BAD_ARGUMENT, "%s: layer %d is incomplete or has the wrong shape", who, i);
  }
  RDP_REQUIRE(w->outc_weight && w->outc_bias, RDP_ERR_BAD_ARGUMENT, "%s: missing OutConv parameters", who);
  return RDP_OK;
}

}  // namespace

#ifdef RDP_DEVELOPMENT
extern "C" int rdp_debug_timeouts(int* out8) {  // development builds only: first mbarrier wait of this module that timed out
  return cudaMemcpyFromSymbol(out8, rdp::ptx::g_rdp_timeout, 8 * sizeof(int)) == cudaSuccess ? 0 : -3;
}
#endif

extern "C" int rdp_unet_train_pack(const float* w_oihw, int cout, int cin, int first_layer, void* w, void* w_t, void* stream_) {
  RDP_REQUIRE(w_oihw && w && cout > 0 && cin > 0, RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_pack: null pointer or empty layer");
  RDP_REQUIRE(!first_layer || 9 * cin <= 64, RDP_ERR_UNSUPPORTED, "rdp_unet_train_pack: first layer with %d channels (9*C must be <= 64)", cin);
  RDP_REQUIRE(first_layer || w_t, RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_pack: w_t is required for every layer but the first");
  const int total = cout * cin * 9;
  if (!first_layer && cin % 32 == 0 && cout % 32 == 0)
    k_pack_train_weights_tiled<<<dim3(cin / 32, cout / 32), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        w_oihw, cout, cin, static_cast<__nv_bfloat16*>(w), static_cast<__nv_bfloat16*>(w_t));
  else
    k_pack_train_weights<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        w_oihw, cout, cin, first_layer ? 1 : 0, static_cast<__nv_bfloat16*>(w), first_layer ? nullptr : static_cast<__nv_bfloat16*>(w_t));
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

extern "C" int rdp_unet_train_workspace_bytes(const RdpTrainWeights* w, int B, int H, int W, size_t* bytes) {
  RDP_REQUIRE(bytes, RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_workspace_bytes: null argument");
  const int rc = check_train_args(w, B, H, W, "rdp_unet_train_workspace_bytes");
  if (rc != RDP_OK) return rc;
  *bytes = carve_train(w, nullptr, B, H, W).bytes;
  return RDP_OK;
}

extern "C" int rdp_unet_train_forward(const RdpTrainWeights* w, const float* x_nchw, int B, int H, int W, float* logits,
                                      const RdpTrainBatchStats* stats, void* workspace, size_t workspace_bytes, void* stream_) {
  int rc = check_train_args(w, B, H, W, "rdp_unet_train_forward");
  if (rc != RDP_OK) return rc;
  RDP_REQUIRE(x_nchw && logits && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_forward: null pointer");
  RDP_REQUIRE(((uintptr_t)workspace & 255) == 0, RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_forward: workspace must be 256-byte aligned");
  const TrainBuffers buf = carve_train(w, workspace, B, H, W);
  RDP_REQUIRE(workspace_bytes >= buf.bytes, RDP_ERR_WORKSPACE_TOO_SMALL, "rdp_unet_train_forward: workspace %zu < required %zu bytes",
              workspace_bytes, buf.bytes);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int dev = 0, num_sms = 148;
  RDP_CUDA_TRY(cudaGetDevice(&dev));
  RDP_CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  RDP_CUDA_TRY(cudaMemsetAsync(buf.zero_shift, 0, 1024 * sizeof(float), stream));
  const float eps = w->bn_eps > 0.f ? w->bn_eps : 1e-5f;

  // conv -> batch statistics -> normalise + ReLU, for stage i with input (src0[c0] ++ src1[c1]) at level `lvl`
  auto stage = [&](int i, const __nv_bfloat16* src0, int c0, const __nv_bfloat16* src1, int c1, int taps) -> int {
    const Shape L = buf.lv[kLevel[i]];
    const int C = w->conv[i].cout;
    const size_t P = (size_t)B * L.H * L.W;
    int r = unet::raw_conv(src0, c0, src1, c1, w->conv[i].w, C, taps, B, L.H, L.W, buf.zero_shift, buf.z[i], num_sms, stream);
    if (r != RDP_OK) return r;
    TRAIN_CHECKPOINT(stream, "forward conv %d (%dx%d, %d+%d -> %d channels)", i, L.H, L.W, c0, c1, C);
    const int blocks = stat_blocks(P, C, num_sms, 3);
    k_bn_stats<<<blocks, ST_THREADS, 2 * C * sizeof(double), stream>>>(buf.z[i], P, C, buf.part);
    k_bn_finalize<<<(C + 7) / 8, 256, 0, stream>>>(buf.part, blocks, C, (double)P, w->conv[i].gamma, w->conv[i].beta, eps, buf.mean[i], buf.var[i],
                                                   buf.invstd[i], buf.scale[i], buf.shift[i]);
    k_bn_relu<<<ew_grid(P * C / 8), 256, 0, stream>>>(buf.z[i], P * C / 8, C / 8, buf.scale[i], buf.shift[i], buf.y[i]);
    TRAIN_CHECKPOINT(stream, "forward batch norm %d", i);
    if (stats && stats->mean[i]) RDP_CUDA_TRY(cudaMemcpyAsync(stats->mean[i], buf.mean[i], C * sizeof(float), cudaMemcpyDeviceToDevice, stream));
    if (stats && stats->var_biased[i]) RDP_CUDA_TRY(cudaMemcpyAsync(stats->var_biased[i], buf.var[i], C * sizeof(float), cudaMemcpyDeviceToDevice, stream));
    return RDP_OK;
  };
#define RDP_TRY(expr)              \
  do {                             \
    const int _rc = (expr);        \
    if (_rc != RDP_OK) return _rc; \
  } while (0)
  RDP_TRY(unet::im2col_first(x_nchw, RDP_IN_F32_NCHW, B, H, W, w->n_channels, buf.col, stream));
  RDP_TRY(stage(0, buf.col, 64, nullptr, 0, 1));
  RDP_TRY(stage(1, buf.y[0], 64, nullptr, 0, 9));
  for (int d = 0; d < 4; ++d) {
    const Shape L = buf.lv[d];
    const int C = w->conv[2 * d + 1].cout;
    RDP_TRY(unet::maxpool2(buf.y[2 * d + 1], B, L.H, L.W, C, buf.pool[d], stream));
    RDP_TRY(stage(2 * d + 2, buf.pool[d], C, nullptr, 0, 9));
    RDP_TRY(stage(2 * d + 3, buf.y[2 * d + 2], w->conv[2 * d + 2].cout, nullptr, 0, 9));
  }
  for (int j = 0; j < 4; ++j) {
    const int low_i = 9 + 2 * j, skip_i = 7 - 2 * j;
    const Shape Ls = buf.lv[3 - j], Ll = buf.lv[4 - j];
    const int Cl = w->conv[low_i].cout, Cs = w->conv[skip_i].cout;
    RDP_TRY(unet::upsample2x(buf.y[low_i], B, Ll.H, Ll.W, Cl, Ls.H, Ls.W, buf.up[j], stream));
    RDP_TRY(stage(10 + 2 * j, buf.y[skip_i], Cs, buf.up[j], Cl, 9));
    RDP_TRY(stage(11 + 2 * j, buf.y[10 + 2 * j], w->conv[10 + 2 * j].cout, nullptr, 0, 9));
  }
  const size_t P0 = (size_t)B * H * W;
  k_outconv_fwd<<<(int)std::min<size_t>((P0 + 127) / 128, (size_t)num_sms * 8), 256, 0, stream>>>(buf.y[17], P0, (size_t)H * W, w->n_classes, w->outc_weight,
                                                                                                   w->outc_bias, logits);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

extern "C" int rdp_unet_train_backward(const RdpTrainWeights* w, const float* dlogits, int B, int H, int W, const RdpTrainGrads* g,
                                       void* workspace, size_t workspace_bytes, void* stream_) {
  int rc = check_train_args(w, B, H, W, "rdp_unet_train_backward");
  if (rc != RDP_OK) return rc;
  RDP_REQUIRE(dlogits && g && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_backward: null pointer");
  const TrainBuffers buf = carve_train(w, workspace, B, H, W);
  RDP_REQUIRE(workspace_bytes >= buf.bytes, RDP_ERR_WORKSPACE_TOO_SMALL, "rdp_unet_train_backward: workspace %zu < required %zu bytes",
              workspace_bytes, buf.bytes);
  for (int i = 0; i < 18; ++i)
    RDP_REQUIRE(g->dw[i] && g->dgamma[i] && g->dbeta[i], RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_backward: missing gradient buffer of layer %d", i);
  RDP_REQUIRE(g->d_outc_weight && g->d_outc_bias, RDP_ERR_BAD_ARGUMENT, "rdp_unet_train_backward: missing OutConv gradient buffers");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int dev = 0, num_sms = 148;
  RDP_CUDA_TRY(cudaGetDevice(&dev));
  RDP_CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));

  const size_t P0 = (size_t)B * H * W;
  RDP_CUDA_TRY(cudaMemsetAsync(g->d_outc_weight, 0, (size_t)w->n_classes * 64 * sizeof(float), stream));
  RDP_CUDA_TRY(cudaMemsetAsync(g->d_outc_bias, 0, (size_t)w->n_classes * sizeof(float), stream));
  {
    const int ob_grid = (int)std::min<size_t>((P0 + 127) / 128, (size_t)num_sms * 8);
    const size_t HW0 = (size_t)H * W;
    if (w->n_classes == 1)
      k_outconv_bwd<1><<<ob_grid, 256, 0, stream>>>(dlogits, buf.y[17], P0, HW0, 1, w->outc_weight, buf.dy[17], g->d_outc_weight, g->d_outc_bias);
    else if (w->n_classes <= 4)
      k_outconv_bwd<4><<<ob_grid, 256, 0, stream>>>(dlogits, buf.y[17], P0, HW0, w->n_classes, w->outc_weight, buf.dy[17], g->d_outc_weight, g->d_outc_bias);
    else
      k_outconv_bwd<8><<<ob_grid, 256, 0, stream>>>(dlogits, buf.y[17], P0, HW0, w->n_classes, w->outc_weight, buf.dy[17], g->d_outc_weight, g->d_outc_bias);
  }

  // the input of stage i as (src0[c0] ++ src1[c1]); what the gradient of that input feeds
  auto input_of = [&](int i, const __nv_bfloat16*& s0, int& c0, const __nv_bfloat16*& s1, int& c1) {
    s1 = nullptr;
    c1 = 0;
    if (i == 0) { s0 = buf.col; c0 = 64; }
    else if (i >= 10 && i % 2 == 0) {
      const int j = (i - 10) / 2;
      s0 = buf.y[7 - 2 * j]; c0 = w->conv[7 - 2 * j].cout;
      s1 = buf.up[j]; c1 = w->conv[9 + 2 * j].cout;
    } else if (i >= 2 && i <= 8 && i % 2 == 0) { s0 = buf.pool[i / 2 - 1]; c0 = w->conv[i - 1].cout; }
    else { s0 = buf.y[i - 1]; c0 = w->conv[i - 1].cout; }
  };

  for (int i = 17; i >= 0; --i) {
    const Shape L = buf.lv[kLevel[i]];
    const int C = w->conv[i].cout;
    const size_t P = (size_t)B * L.H * L.W;
    // ---- BatchNorm + ReLU backward: dgamma, dbeta, dz ----
    const int blocks = stat_blocks(P, C, num_sms, 2);
    k_bn_bwd_reduce<<<blocks, ST_THREADS, 2 * C * sizeof(double), stream>>>(buf.dy[i], buf.scale[i], buf.shift[i], buf.z[i], P, C, buf.mean[i], buf.invstd[i], buf.part);
    k_bn_bwd_finalize<<<(C + 7) / 8, 256, 0, stream>>>(buf.part, blocks, C, (double)P, w->conv[i].gamma, buf.invstd[i], g->dgamma[i], g->dbeta[i],
                                                       buf.kscale, buf.kdb, buf.kdg);
    k_bn_bwd_apply<<<ew_grid(P * C / 8), 256, 0, stream>>>(buf.dy[i], buf.scale[i], buf.shift[i], buf.z[i], P * C / 8, C / 8, buf.mean[i], buf.invstd[i], buf.kscale,
                                                           buf.kdb, buf.kdg, buf.dz);
    TRAIN_CHECKPOINT(stream, "backward batch norm %d", i);
    // ---- weight gradient: tcgen05 GEMM over the pixels, operands read in place ----
    const __nv_bfloat16 *s0, *s1;
    int c0, c1;
    input_of(i, s0, c0, s1, c1);
    const int cin_pad = c0 + c1, cin_real = w->conv[i].cin;
    RDP_CUDA_TRY(cudaMemsetAsync(buf.dws, 0, (size_t)C * cin_real * 9 * sizeof(float), stream));
    {
      WgradParams wp{};
      wp.B = B; wp.H = L.H; wp.W = L.W;
      wp.cout = C; wp.cin_real = cin_real; wp.c0 = c0;
      wp.co_blocks = C >= 128 ? 2 : 1;
      const int n_tile = cin_pad >= 128 ? 128 : 64;
      wp.m_tiles = (C + 127) / 128;
      wp.n_tiles = cin_pad / n_tile;
      wp.dys = i == 0 ? 1 : 3;
      wp.px_tiles = (L.W + WG_PW - 1) / WG_PW;
      wp.py_tiles = (L.H + WG_PH - 1) / WG_PH;
      const long long patches = (long long)B * wp.px_tiles * wp.py_tiles;
      const int items = wp.m_tiles * wp.n_tiles * wp.dys;
      // Split-K factor: every CTA ends with 128 x N x 3 fp32 reductions into the scratch, and L2 retires about one of
      // them per slice and clock (14 M of them = 40 us on the 512 x 512 layers with six CTAs per work item), so more
      // CTAs per work item buy tensor time only while the layer has it to give.  Pick the factor that minimises
      // waves x patches-per-CTA x (cycles per patch) + reductions / (L2 slices), both in cycles; never fewer than
      // four patches per CTA.
      {
        const double patch_cycles = 24.0 * (n_tile == 128 ? 100.0 : 77.0), reds_per_cta = 128.0 * n_tile * 3.0, l2_slices = 184.0;
        const int ks_max = (int)std::max<long long>(1, std::min<long long>(2 * num_sms / items, patches / 4));
        double best = 1e300;
        for (int ks = 1; ks <= ks_max; ++ks) {
          const double waves = (double)((items * ks + num_sms - 1) / num_sms);
          const double cost = waves * (double)((patches + ks - 1) / ks) * patch_cycles + (double)items * ks * reds_per_cta / l2_slices;
          if (cost < best) {
            best = cost;
            wp.ksplit = ks;
          }
        }
      }
      wp.first_layer = i == 0;
      wp.dw = buf.dws;
#ifdef RDP_DEVELOPMENT
      static const bool no_co64 = getenv("RDP_WGRAD_NO_CO64") != nullptr;  // A/B: the general kernel for these layers too
#else
      const bool no_co64 = false;
#endif
      const bool co64 = C == 64 && !no_co64;   // operands exchanged, nine taps per CTA (k_wgrad_co64)
      CUtensorMap mz, mx0, mx1;
      RDP_TRY(make_nhwc_map(&mz, buf.dz, B, L.H, L.W, C, WG_PW));
      RDP_TRY(make_nhwc_map(&mx0, s0, B, L.H, L.W, c0, WG_HALO_W, co64 ? WG9_HALO_H : WG_PH));
      if (c1) RDP_TRY(make_nhwc_map(&mx1, s1, B, L.H, L.W, c1, WG_HALO_W, co64 ? WG9_HALO_H : WG_PH));
      else mx1 = mx0;
      if (co64) {
        wp.n_tiles = cin_pad / 64;
        wp.ksplit = (int)std::max<long long>(1, std::min<long long>(num_sms / wp.n_tiles, patches / 4));
        RDP_TRY(launch_wgrad_co64(mz, mx0, mx1, wp, stream));
      } else if (n_tile == 64) RDP_TRY(launch_wgrad_t<64>(mz, mx0, mx1, wp, stream));
      else RDP_TRY(launch_wgrad_t<128>(mz, mx0, mx1, wp, stream));
      k_dw_layout<<<dim3((cin_real + 31) / 32, (C + 31) / 32), 256, 0, stream>>>(buf.dws, cin_real, C, g->dw[i]);
    }
    TRAIN_CHECKPOINT(stream, "weight gradient %d (%dx%d, %d x %d)", i, L.H, L.W, C, cin_pad);
    if (i == 0) break;  // the image needs no gradient
    // ---- input gradient: the same convolution kernel with the flipped, transposed weights ----
    const bool up_first = i >= 10 && i % 2 == 0, down_first = i >= 2 && i <= 8 && i % 2 == 0;
    // second conv of a block: its input is y[i-1] and nothing else reads that, so dX is dY[i-1] itself
    __nv_bfloat16* dx_out = (up_first || down_first) ? buf.dx : buf.dy[i - 1];
    RDP_TRY(unet::raw_conv(buf.dz, C, nullptr, 0, w->conv[i].w_t, cin_pad, 9, B, L.H, L.W, buf.zero_shift, dx_out, num_sms, stream));
    TRAIN_CHECKPOINT(stream, "input gradient %d", i);
    // ---- route it to the producers of the input ----
    if (up_first) {                                    // Up block: [skip | up-sampled low branch]
      const int j = (i - 10) / 2, skip_i = 7 - 2 * j, low_i = 9 + 2 * j;
      const Shape Ll = buf.lv[4 - j];
      k_slice_channels<<<ew_grid(P * c0 / 8), 256, 0, stream>>>(buf.dx, cin_pad / 8, 0, P, c0 / 8, buf.dy[skip_i]);
      const int dyp = L.H - 2 * Ll.H, dxp = L.W - 2 * Ll.W;
      const size_t Pl = (size_t)B * Ll.H * Ll.W;
      k_upsample_bwd<<<dim3((Ll.W * (c1 / 8) + 255) / 256, Ll.H, B), 256, 0, stream>>>(buf.dx, cin_pad, c0, B, Ll.H, Ll.W, c1 / 8, L.H, L.W, dyp / 2, dxp / 2, buf.dy[low_i]);
    } else if (down_first) {                           // Down block: the input was pooled from y[i-1], which also feeds a skip
      const Shape Lp = buf.lv[kLevel[i] - 1];
      const size_t Pp = (size_t)B * Lp.H * Lp.W;
      // the skip consumer (an Up block) ran earlier in this loop and wrote dy[i-1]: accumulate
      k_maxpool_bwd<<<dim3(((Lp.W / 2) * (c0 / 8) + 255) / 256, std::max(1, Lp.H / 2), B), 256, 0, stream>>>(buf.y[i - 1], buf.dx, B, Lp.H, Lp.W, c0 / 8, 1,
                                                                                                        buf.dy[i - 1]);
    }
    TRAIN_CHECKPOINT(stream, "routing %d", i);
  }
#undef RDP_TRY
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}
