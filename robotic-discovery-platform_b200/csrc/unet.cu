// U-Net forward on sm_100a: replaces UNet.forward of the reference (pkg/segmentation_model.py:109-120) plus the
// server's mask rule (services/vision_analysis/server.py:124).
//
// Activations live in HBM as NHWC bf16.  Every 3x3 convolution (+ folded eval-BatchNorm + ReLU) is one launch of an
// implicit GEMM on the 5th-generation tensor cores (`k_conv3x3_halo`; `k_conv_first` for inc.0; `k_conv3x3` is the
// per-tap fallback):
//   * M = a 128-pixel spatial patch of one frame, N = 64..256 output channels, K = 9 taps x Cin;
//   * TMA loads one (tile + halo) box per 64 input channels from a 4-D NHWC tensor map - out-of-bounds pixels are
//     zero-filled by the TMA unit, which *is* the convolution padding - and the 9 taps address it in place through
//     shifted shared-memory descriptors; weight slices are resident or streamed, all with 128-byte swizzle;
//   * one elected thread issues tcgen05.mma (bf16 x bf16 -> fp32) into double-buffered TMEM accumulators, so the
//     epilogue of tile i overlaps the MMAs of tile i+1 (persistent CTAs, static round-robin over tiles);
//   * epilogue warps read TMEM (tcgen05.ld), add the BN shift, apply ReLU and store bf16 NHWC; MaxPool2d(2) leaves as
//     a second, pooled output of the same epilogue; the last layer applies the 1x1 OutConv there as well and writes
//     fp32 logits and/or the thresholded mask;
//   * skip connections are read in place: the K loop takes its first channel blocks from the skip tensor and the
//     rest from the up-sampled branch, so `torch.cat` never materialises; the bilinear x2 up-sampling
//     (align_corners=True, position-dependent weights) is done by interpolation warps inside the convolution that
//     consumes it, straight into the swizzled operand tile;
//   * inc.0 (3 input channels) gathers its 3x3 neighbourhoods from the uint8 image inside the kernel (no im2col in HBM).
// Development switches (environment variables that change kernel selection or skip work) exist only in builds with
// -DRDP_DEVELOPMENT; the shipped library ignores them.
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <set>
#include <utility>

#include "rdp_common.cuh"
#include "sm100_ptx.cuh"
#include "unet_internal.cuh"

// Development switches are compiled out of the shipped library: an environment variable must never change what a
// server computes.  Build with -DRDP_DEVELOPMENT (RDP_DEVELOPMENT=1 python rdp_b200/build.py --force) to get them.
#ifdef RDP_DEVELOPMENT
#define RDP_DEV(...) __VA_ARGS__
static inline const char* dev_env(const char* name) { return getenv(name); }
#else
#define RDP_DEV(...)
static inline const char* dev_env(const char*) { return nullptr; }
#endif

namespace rdp {
namespace unet {

using namespace ptx;

constexpr int TILE_M = 128;   // pixels per tile = UMMA M
constexpr int BLOCK_K = 64;   // bf16 elements per K-step = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int CONV_THREADS = 192;  // warp 0: TMA producer, warp 1: MMA issuer, warps 2-5: epilogue
constexpr int A_STAGE_BYTES = TILE_M * BLOCK_K * 2;
constexpr int SMEM_BUDGET = 200 * 1024;

struct ConvParams {
  int B, H, W;
  int tile_w, tile_h, tiles_x, tiles_y;
  int n_tiles;     // Cout / BN
  int kb0, kb1;    // 64-channel blocks taken from source 0 (skip / plain input) and source 1 (upsampled)
  int taps;        // 9, or 1 for the im2col'ed first layer
  int cin_pad;     // (kb0 + kb1) * 64: K stride between taps in the packed weights
  int cout;
  int total_work;  // B * tiles_y * tiles_x * n_tiles
  const float* shift;
  __nv_bfloat16* out;  // NHWC, may be null when only the head output is wanted
  // fused OutConv head (only with BN == cout == 64)
  const float* head_w;
  const float* head_b;
  int n_classes;
  float* logits;
  uint8_t* mask;
  // fused bilinear x2 upsample of source 1 (Up blocks): the low-resolution tensor the interpolation reads
  const __nv_bfloat16* low;  // (B, low_h, low_w, kb1*64) NHWC, null when source 1 is a plain tensor map
  int low_h, low_w;
  int pad_y, pad_x;          // offset of the interpolated image inside the skip-sized frame (Up.forward's F.pad)
  float ry, rx;              // align_corners=True scales (h-1)/(2h-1), (w-1)/(2w-1)
  int interp_fast;           // exact doubling without pad (H = 2 low_h, W = 2 low_w): the producers work on 2x2 blocks
  int debug_waits;           // development aid: CTA 0 prints where its MMA and epilogue threads waited
  // fused MaxPool2d(2): second, pooled output of the layers that feed a Down block (8-wide tiles only)
  __nv_bfloat16* pool_out;   // (B, H/2, W/2, cout) NHWC or null
};

constexpr int AUX_SHIFT_FLOATS = 512;   // BN shift of every output channel of the layer
constexpr int AUX_HEAD_FLOATS = 8 * 64;  // OutConv weights
constexpr int AUX_BYTES = (AUX_SHIFT_FLOATS + AUX_HEAD_FLOATS + 8) * 4;

template <int BN>
struct ConvCfg {
  static constexpr int B_STAGE_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = SMEM_BUDGET / STAGE_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/ + AUX_BYTES;
};

// sigmoid(x) > 0.5 evaluated in fp32 as torch does: true from x = 1.5 * 2^-24 upwards (probed on torch 2.11 CPU)
__device__ __forceinline__ bool mask_rule(float logit) { return logit > 8.940697e-8f; }


// All threads: stage the per-channel shift (and the OutConv weights / bias) in shared memory once per CTA.
__device__ __forceinline__ void stage_aux(const ConvParams& p, float* s_aux) {
  for (int i = threadIdx.x; i < p.cout; i += blockDim.x) s_aux[i] = __ldg(p.shift + i);
  if (p.head_w != nullptr) {
    for (int i = threadIdx.x; i < p.n_classes * 64; i += blockDim.x) s_aux[p.cout + i] = __ldg(p.head_w + i);
    if (threadIdx.x < p.n_classes) s_aux[p.cout + AUX_HEAD_FLOATS + threadIdx.x] = __ldg(p.head_b + threadIdx.x);
  }
}

// 16-byte read from shared memory by shared-window address (a generic LD through a pointer the compiler cannot
// prove to be shared is slower than LDS)
__device__ __forceinline__ float4 lds_f4(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}

// two fp32 -> packed bf16x2 (lo = first element), optionally clamped at zero by the conversion itself
template <bool RELU>
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t d;
  if (RELU) asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// Drains one 128 x BN fp32 accumulator tile: BN shift + ReLU -> bf16 NHWC store; for the last layer also the 1x1
// OutConv (+ mask rule).  Thread = one pixel (TMEM lane); 32-column chunks, the next chunk's tcgen05.ld is in
// flight while the current one is processed.
// acc_empty_bar: local barrier to arrive on once the accumulator stage is drained (may be null); remote_empty: the same
// as a shared::cluster address when the MMA issuer lives in the other CTA of a pair (0 = use the local barrier).
template <int BN, bool POOL = false, bool RAW = false>
__device__ __forceinline__ void epilogue_tile(const ConvParams& p, const float* s_aux, uint32_t taddr, int nt, int n, int y,
                                              int x, bool inside, uint64_t* acc_empty_bar, int lane, uint32_t remote_empty = 0) {
  const size_t pixel = ((size_t)n * p.H + y) * p.W + x;
  const bool head_on = BN == 64 && p.head_w != nullptr;
  float head[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) head[c] = 0.f;
  uint32_t r[2][32];
  tmem_ld_32x32(taddr, r[0]);
  if (BN == 64) {
    // 64 columns fit the registers: read the whole accumulator and hand the TMEM stage back before any arithmetic.
    // What matters to the MMA warp is how long a stage is held, and the 64-wide tiles are short.
    tmem_ld_32x32(taddr + 32, r[1]);
    tmem_ld_wait();
    if (acc_empty_bar != nullptr || remote_empty != 0) {
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (remote_empty) mbar_arrive_remote(remote_empty);
        else mbar_arrive(acc_empty_bar);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < BN / 32; ++c) {
    if (BN != 64) {
      tmem_ld_wait();
      if (c + 1 < BN / 32) tmem_ld_32x32(taddr + (c + 1) * 32, r[(c + 1) & 1]);
    }
    const uint32_t sh_addr = smem_u32(s_aux + nt * BN + c * 32);
    float v[32];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 sh = lds_f4(sh_addr + 16 * q);
      v[4 * q + 0] = __uint_as_float(r[c & 1][4 * q + 0]) + sh.x;
      v[4 * q + 1] = __uint_as_float(r[c & 1][4 * q + 1]) + sh.y;
      v[4 * q + 2] = __uint_as_float(r[c & 1][4 * q + 2]) + sh.z;
      v[4 * q + 3] = __uint_as_float(r[c & 1][4 * q + 3]) + sh.w;
    }
    if (BN == 64 && head_on) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
      // One rolled loop over the classes (the reference uses a single class): unrolling it eight times made the
      // kernel outgrow the instruction cache and cost more than the arithmetic itself.
#pragma unroll 1
      for (int cls = 0; cls < p.n_classes; ++cls) {
        const uint32_t hw_addr = smem_u32(s_aux + p.cout + cls * 64 + c * 32);
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;  // four chains instead of one 64-long dependency
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 w = lds_f4(hw_addr + 16 * q);
          a0 = fmaf(v[4 * q + 0], w.x, a0);
          a1 = fmaf(v[4 * q + 1], w.y, a1);
          a2 = fmaf(v[4 * q + 2], w.z, a2);
          a3 = fmaf(v[4 * q + 3], w.w, a3);
        }
        head[cls] += (a0 + a1) + (a2 + a3);
      }
    }
    if (p.out != nullptr) {
      // 32 bytes (one full DRAM sector) per store instruction
      __nv_bfloat16* dst = p.out + pixel * p.cout + nt * BN + c * 32;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        uint32_t o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = pack_bf16x2<!RAW>(v[16 * q + 2 * e], v[16 * q + 2 * e + 1]);
        if (inside)
          asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 16 * q), "r"(o[0]), "r"(o[1]),
                       "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7])
                       : "memory");
        if (POOL && p.pool_out != nullptr) {
          // MaxPool2d(2) on the rounded values (rounding is monotone, so max commutes with it).  In an 8-wide tile
          // the pixel is (lane >> 3, lane & 7) within the warp's four rows: window partners are lanes ^1 and ^8.
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&o[e]);
            uint32_t t1 = __shfl_xor_sync(0xffffffffu, o[e], 1);
            a = __hmax2(a, *reinterpret_cast<__nv_bfloat162*>(&t1));
            uint32_t au = *reinterpret_cast<uint32_t*>(&a);
            uint32_t t8 = __shfl_xor_sync(0xffffffffu, au, 8);
            a = __hmax2(a, *reinterpret_cast<__nv_bfloat162*>(&t8));
            o[e] = *reinterpret_cast<uint32_t*>(&a);
          }
          const int Ho = p.H >> 1, Wo = p.W >> 1, yo = y >> 1, xo = x >> 1;
          if (!(lane & 9) && yo < Ho && xo < Wo) {
            __nv_bfloat16* pd = p.pool_out + (((size_t)n * Ho + yo) * Wo + xo) * p.cout + nt * BN + c * 32 + 16 * q;
            asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(pd), "r"(o[0]), "r"(o[1]), "r"(o[2]),
                         "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7])
                         : "memory");
          }
        }
      }
    }
  }
  // accumulator stage drained: hand it back to the MMA warp
  if (BN != 64 && (acc_empty_bar != nullptr || remote_empty != 0)) {
    tc_fence_before();
    __syncwarp();
    if (lane == 0) {
      if (remote_empty) mbar_arrive_remote(remote_empty);
      else mbar_arrive(acc_empty_bar);
    }
  }
  if (BN == 64 && head_on && inside) {
    const size_t hw_px = (size_t)p.H * p.W, off = (size_t)y * p.W + x;
#pragma unroll 1
    for (int cls = 0; cls < p.n_classes; ++cls) {
      const float logit = head[cls] + s_aux[p.cout + AUX_HEAD_FLOATS + cls];
      const size_t o = ((size_t)n * p.n_classes + cls) * hw_px + off;
      if (p.logits) p.logits[o] = logit;
      if (p.mask) p.mask[o] = mask_rule(logit) ? 1 : 0;
    }
  }
}

// RAW: the plain convolution (no ReLU; the caller passes a zero shift) for the training step, where BatchNorm needs the
// pre-normalisation values and the gradient convolutions have no activation at all.
template <int BN, bool RAW = false>
__global__ void __launch_bounds__(CONV_THREADS, 1)
k_conv3x3(const __grid_constant__ CUtensorMap map_src0, const __grid_constant__ CUtensorMap map_src1,
          const __grid_constant__ CUtensorMap map_w, const ConvParams p) {
  using Cfg = ConvCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + Cfg::STAGES * A_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
  uint64_t* full = bars;                       // [STAGES] TMA -> MMA
  uint64_t* empty = bars + Cfg::STAGES;        // [STAGES] MMA -> TMA
  uint64_t* acc_full = bars + 2 * Cfg::STAGES; // [2] MMA -> epilogue
  uint64_t* acc_empty = acc_full + 2;          // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* s_aux = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_steps = (p.kb0 + p.kb1) * p.taps;
  stage_aux(p, s_aux);

  if (threadIdx.x == 0) {
    prefetch_tensormap(&map_src0);
    prefetch_tensormap(&map_src1);
    prefetch_tensormap(&map_w);
    for (int s = 0; s < Cfg::STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 4);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      uint32_t stage = 0, phase = 0;
      for (int work = blockIdx.x; work < p.total_work; work += gridDim.x) {
        const int nt = work % p.n_tiles;
        int m = work / p.n_tiles;
        const int tx = m % p.tiles_x; m /= p.tiles_x;
        const int ty = m % p.tiles_y;
        const int n = m / p.tiles_y;
        const int x0 = tx * p.tile_w, y0 = ty * p.tile_h;
        for (int cb = 0; cb < p.kb0 + p.kb1; ++cb) {
          const CUtensorMap* src = cb < p.kb0 ? &map_src0 : &map_src1;
          const int c0 = (cb < p.kb0 ? cb : cb - p.kb0) * BLOCK_K;
          for (int tap = 0; tap < p.taps; ++tap) {
            const int dy = p.taps == 9 ? tap / 3 - 1 : 0, dx = p.taps == 9 ? tap % 3 - 1 : 0;
            mbar_wait(&empty[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full[stage], Cfg::STAGE_BYTES);
            tma_load_4d(smem_a + stage * A_STAGE_BYTES, src, c0, x0 + dx, y0 + dy, n, &full[stage]);
            tma_load_2d(smem_b + stage * Cfg::B_STAGE_BYTES, &map_w, tap * p.cin_pad + cb * BLOCK_K, nt * BN, &full[stage]);
            if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    {
      // ================= MMA issuer: the warp walks the loop converged, one elected lane issues =================
      constexpr uint32_t idesc = make_idesc_bf16(TILE_M, BN);
      uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
      for (int work = blockIdx.x; work < p.total_work; work += gridDim.x) {
        mbar_wait(&acc_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_acc = tmem_base + as * BN;
        for (int ks = 0; ks < k_steps; ++ks) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t da = make_sw128_desc(smem_u32(smem_a + stage * A_STAGE_BYTES));
          const uint64_t db = make_sw128_desc(smem_u32(smem_b + stage * Cfg::B_STAGE_BYTES));
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // advancing 16 bf16 = 32 bytes inside the 128-byte swizzle row: +2 in 16-byte descriptor units
            mma_bf16_ss_elect(tmem_acc, da + 2 * k, db + 2 * k, idesc, (ks | k) != 0);
          }
          tc_commit_elect(&empty[stage]);  // frees the smem stage once these MMAs have read it
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        tc_commit_elect(&acc_full[as]);    // accumulator complete
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ================= epilogue (warps 2..5) =================
    const int quarter = warp & 3;  // TMEM lanes this warp may read: 32*quarter .. +31
    const int row = quarter * 32 + lane;
    uint32_t as = 0, aphase = 0;
    for (int work = blockIdx.x; work < p.total_work; work += gridDim.x) {
      const int nt = work % p.n_tiles;
      int m = work / p.n_tiles;
      const int tx = m % p.tiles_x; m /= p.tiles_x;
      const int ty = m % p.tiles_y;
      const int n = m / p.tiles_y;
      const int y = ty * p.tile_h + row / p.tile_w, x = tx * p.tile_w + row % p.tile_w;
      const bool inside = y < p.H && x < p.W;
      mbar_wait(&acc_full[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + as * BN + ((uint32_t)(quarter * 32) << 16);
      epilogue_tile<BN, false, RAW>(p, s_aux, taddr, nt, n, y, x, inside, &acc_empty[as], lane);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}


// ---------------------------------------------------------------------------------------------------------
// k_conv3x3_halo: the 3x3 convolution with the activation patch loaded ONCE per 64-channel block
// ---------------------------------------------------------------------------------------------------------
// The 9 taps of a 3x3 convolution read the same (tile_h+2) x (tile_w+2) halo patch shifted by one pixel.  With an
// 8-wide tile the halo is 18 x 10 pixels x 128 bytes, laid out by one TMA box as 180 swizzled 128-byte rows
// (row = hy*10 + hx).  For tap (dy,dx) the MMA's 8-row core groups are the 8 pixels of tile row ty, which sit at
// byte offset ((ty+dy)*10 + dx) * 128: a descriptor with start = stage + (dy*10+dx)*128 and a stride of
// 10 rows = 1280 bytes between groups addresses them in place.  The 128-byte swizzle is a function of the shared
// memory address bits, so a start address that is not 1024-byte aligned needs no other adjustment.
// L2 -> shared-memory traffic for A drops from 9 x 16 KB to 22.5 KB per channel block (6.4x).
// Weights stay resident in shared memory for the whole kernel when all 9*Cin/64 slices fit (the 64-wide
// full-resolution layers), otherwise they stream through their own ring, one 64-wide slice per tap.
// TILES = 1: one 8x16 tile per work item (halo 18x10).  TILES = 2: a 16x16 super-tile, i.e. two 8x16 tiles that share
// one 18x18 halo box and, above all, every streamed weight slice (half the L2 -> shared-memory weight traffic).
constexpr int HALO_H = 18;
__host__ __device__ constexpr int halo_w(int tiles) { return 8 * tiles + 2; }
__host__ __device__ constexpr int halo_bytes(int tiles) { return halo_w(tiles) * HALO_H * 128; }  // 23040 / 41472
__host__ __device__ constexpr int halo_stride(int tiles) { return (halo_bytes(tiles) + 1023) / 1024 * 1024; }
constexpr int HALO_SMEM_TOTAL = 227 * 1024;
constexpr int HALO_MAX_STAGES = 12;
constexpr int HALO_BAR_BYTES = 512;  // barriers; the epilogue constants follow (aux_bytes())
// shift[cout], then (only for the head layer) OutConv weights and bias
inline int aux_bytes(int cout, bool head) { return (cout + (head ? AUX_HEAD_FLOATS + 8 : 0)) * 4; }

// Accumulator stages in tensor memory: the 64-wide layers finish a tile's MMAs in about a microsecond while their
// epilogue stores a whole tile (and its pooled copy) to HBM, so four stages (256 of the 512 columns) let the MMA warp
// run ahead of a slow store instead of waiting for it; the wider layers keep two.
__host__ __device__ constexpr int acc_stages(int bn, int tiles) { return (bn == 64 && tiles == 1) ? 4 : 2; }
constexpr int MAX_ACC_STAGES = 4;

struct HaloCfg {
  int sa, sb;  // A (halo) stages, B (weight slice) stages
  int sl;      // source-patch stages of the fused upsample (1 or 2)
};

// High word of a SWIZZLE_128B K-major descriptor: stride between 8-row groups, version 1, layout type 2.
__host__ __device__ constexpr uint32_t desc_hi(uint32_t sbo_bytes) { return (sbo_bytes >> 4) | (1u << 14) | (2u << 29); }
// Low word: start address in 16-byte units (smem < 256 KB, so it fits 14 bits) and LBO = 1.
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr) { return (smem_addr >> 4) | (1u << 16); }
__device__ __forceinline__ uint64_t desc_pack(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }


// MMA issue loop of the halo kernel (one thread).  The issue loop is the critical path of the 64-wide layers (a
// 128x64x16 MMA occupies the tensor pipe for only ~32-48 cycles), so descriptors are a constant high word plus
// (base + immediate) low words, taps are unrolled, and when the TMEM allocation starts at column 0 (BASE0, the
// usual case with one CTA per SM) the accumulator address is an immediate as well: the tile loop is unrolled
// over the two accumulator stages.
template <int BN, bool RESIDENT, bool BASE0, int TILES>
__device__ __forceinline__ void halo_mma_loop(const ConvParams& p, const HaloCfg& cfg, uint8_t* smem_a, uint8_t* smem_b,
                                              uint64_t* full_a, uint64_t* empty_a, uint64_t* full_b, uint64_t* empty_b,
                                              uint64_t* w_full, uint64_t* acc_full, uint64_t* acc_empty, uint32_t tmem_base) {
  constexpr int B_BLOCK = BN * BLOCK_K * 2;
  constexpr uint32_t idesc = make_idesc_bf16(TILE_M, BN);
  constexpr int HW = halo_w(TILES);
  constexpr uint32_t hi_a = desc_hi(HW * 128), hi_b = desc_hi(1024);
  const int ncb = p.kb0 + p.kb1;
  uint32_t sa = 0, pa = 0, sb = 0, pb = 0, aphase = 0;
  const uint32_t a_lo0 = desc_lo(smem_u32(smem_a)), b_lo0 = desc_lo(smem_u32(smem_b));
  RDP_DEV(long long w_acc = 0, w_a = 0, w_b = 0, t_begin = clock64(), t0;)
  if (RESIDENT) {
    mbar_wait(w_full, 0);
    tc_fence_after();
  }
  constexpr int NACC = acc_stages(BN, TILES);
  int work = blockIdx.x;
  while (work < p.total_work) {
#pragma unroll
    for (int as = 0; as < NACC; ++as) {
      if (work < p.total_work) {
        RDP_DEV(t0 = clock64();)
        mbar_wait(&acc_empty[as], aphase ^ 1);
        RDP_DEV(w_acc += clock64() - t0;)
        tc_fence_after();
        const uint32_t tmem_acc = BASE0 ? (uint32_t)(as * TILES * BN) : tmem_base + as * TILES * BN;
        for (int cb = 0; cb < ncb; ++cb) {
          RDP_DEV(t0 = clock64();)
          mbar_wait(&full_a[sa], pa);
          RDP_DEV(w_a += clock64() - t0;)
          tc_fence_after();
          const uint32_t a_lo = a_lo0 + sa * (halo_stride(TILES) >> 4);
          const uint32_t bres_lo = b_lo0 + cb * (9 * B_BLOCK >> 4);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            uint32_t b_lo;
            if (RESIDENT) {
              b_lo = bres_lo + tap * (B_BLOCK >> 4);
            } else {
              RDP_DEV(t0 = clock64();)
              mbar_wait(&full_b[sb], pb);
              RDP_DEV(w_b += clock64() - t0;)
              tc_fence_after();
              b_lo = b_lo0 + sb * (B_BLOCK >> 4);
            }
            const uint32_t at_lo = a_lo + ((tap / 3) * HW + tap % 3) * 8;  // 128 bytes per halo row
#pragma unroll
            for (int h = 0; h < TILES; ++h) {  // the tiles of a super-tile reuse the weight slice
#pragma unroll
              for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                mma_bf16_ss_elect(tmem_acc + h * BN, desc_pack(at_lo + h * 64 + 2 * k, hi_a), desc_pack(b_lo + 2 * k, hi_b), idesc,
                            (cb | tap | k) != 0);
            }
            if (!RESIDENT) {
              tc_commit_elect(&empty_b[sb]);
              if (++sb == (uint32_t)cfg.sb) { sb = 0; pb ^= 1; }
            }
          }
          tc_commit_elect(&empty_a[sa]);
          if (++sa == (uint32_t)cfg.sa) { sa = 0; pa ^= 1; }
        }
        tc_commit_elect(&acc_full[as]);
        work += gridDim.x;
      }
    }
    aphase ^= 1;
  }
  RDP_DEV(if (p.debug_waits && blockIdx.x == 0 && (threadIdx.x & 31) == 0)
    printf("  mma thread: total %lld cyc, waiting acc_empty %lld, full_a %lld, full_b %lld (BN %d tiles %d resident %d)\n",
           clock64() - t_begin, w_acc, w_a, w_b, BN, TILES, (int)RESIDENT);)
}

// interpolation warps: 8 where a block's MMA time is short (64-wide layers) or its halo large (super-tiles), else 4
__host__ __device__ constexpr int interp_warps(int bn, int tiles) { return (bn == 64 || tiles == 2) ? 8 : 4; }
// epilogue groups of 4 warps: the 64-wide layers finish a tile's MMAs faster than one group drains it, so two groups
// alternate tiles (each owns one accumulator stage)
__host__ __device__ constexpr int epi_groups(int bn, bool interp = false) { return (bn == 64 && !interp) ? 2 : 1; }
__host__ __device__ constexpr int conv_threads(bool interp, int bn, int tiles) {
  return 64 + 128 * epi_groups(bn, interp) + (interp ? interp_warps(bn, tiles) * 32 : 0);
}

// Low-resolution source patch of one halo block: 18x10 output pixels at x2 span at most 11x7 source pixels.
constexpr int PATCH_H = 11;
__host__ __device__ constexpr int patch_w(int tiles) { return tiles == 1 ? 7 : 11; }  // 10 / 18 output columns at x2
__host__ __device__ constexpr int patch_bytes(int tiles) { return patch_w(tiles) * PATCH_H * 128; }
__host__ __device__ constexpr int patch_stride(int tiles) { return (patch_bytes(tiles) + 1023) / 1024 * 1024; }
constexpr int PATCH_STAGES = 2;  // upper bound; HaloCfg::sl is what a launch uses

// Bilinear x2 (align_corners=True) + zero pad of one 16-byte chunk (8 channels) of one halo pixel, read from the
// source patch in shared memory, in the same fp32 operation order as k_upsample2x, so that fusing the interpolation
// into the convolution changes no bit.
__device__ __forceinline__ uint4 interp_chunk(const ConvParams& p, const uint8_t* patch, int pw, int sy0, int sx0, int ys, int xs, int j) {
  uint4 o = make_uint4(0, 0, 0, 0);
  const int yo = ys - p.pad_y, xo = xs - p.pad_x;
  if (ys < 0 || ys >= p.H || xs < 0 || xs >= p.W || yo < 0 || yo >= 2 * p.low_h || xo < 0 || xo >= 2 * p.low_w) return o;
  const int h = p.low_h, w = p.low_w;
  const float fy = p.ry * (float)yo, fx = p.rx * (float)xo;
  const int y0 = (int)fy, x0 = (int)fx;
  const int y1 = y0 + (y0 < h - 1 ? 1 : 0), x1 = x0 + (x0 < w - 1 ? 1 : 0);
  const float ly1 = fy - (float)y0, ly0 = 1.f - ly1, lx1 = fx - (float)x0, lx0 = 1.f - lx1;
  const uint4* base = reinterpret_cast<const uint4*>(patch) + j;
  const int r0 = (y0 - sy0) * pw - sx0, r1 = (y1 - sy0) * pw - sx0;
  const uint4 v00 = base[(r0 + x0) * 8], v01 = base[(r0 + x1) * 8];
  const uint4 v10 = base[(r1 + x0) * 8], v11 = base[(r1 + x1) * 8];
  const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&v00);
  const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&v01);
  const __nv_bfloat162* c = reinterpret_cast<const __nv_bfloat162*>(&v10);
  const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(&v11);
  __nv_bfloat162* po = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float2 fa = __bfloat1622float2(a[q]), fb = __bfloat1622float2(b[q]);
    const float2 fc = __bfloat1622float2(c[q]), fd = __bfloat1622float2(d[q]);
    const float e0 = ly0 * (lx0 * fa.x + lx1 * fb.x) + ly1 * (lx0 * fc.x + lx1 * fd.x);
    const float e1 = ly0 * (lx0 * fa.y + lx1 * fb.y) + ly1 * (lx0 * fc.y + lx1 * fd.y);
    po[q] = __floats2bfloat162_rn(e0, e1);
  }
  return o;
}

// The same interpolation for the exact doubling without pad, a 2x2 block of halo pixels at a time.  The halo of a tile
// starts at an odd output row and column (16 ty - 1, 8 tx TILES - 1), so it splits into blocks of rows (2k-1, 2k) x
// columns (2j-1, 2j), and under align_corners those four pixels read the SAME four source pixels: four shared-memory
// loads and one conversion of each per block instead of per pixel.  Indices and weights come from the expressions of
// interp_chunk per row and column, a block whose pixels do not share their sources (first / last row or column of
// the image) is done pixel by pixel, and pixels outside the image are the zero padding: no bit changes.
struct InterpAxis { int i0, i1; float l0, l1; };
__device__ __forceinline__ InterpAxis interp_axis(float r, int o, int n) {
  InterpAxis a;
  const float f = r * (float)o;
  a.i0 = (int)f;
  a.i1 = a.i0 + (a.i0 < n - 1 ? 1 : 0);
  a.l1 = f - (float)a.i0;
  a.l0 = 1.f - a.l1;
  return a;
}
__device__ __forceinline__ void interp_unpack(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 interp_emit(const float (&a)[8], const float (&b)[8], const float (&c)[8], const float (&d)[8],
                                             const InterpAxis& r, const InterpAxis& cx) {
  uint4 o;
  __nv_bfloat162* po = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float e0 = r.l0 * (cx.l0 * a[2 * q] + cx.l1 * b[2 * q]) + r.l1 * (cx.l0 * c[2 * q] + cx.l1 * d[2 * q]);
    const float e1 = r.l0 * (cx.l0 * a[2 * q + 1] + cx.l1 * b[2 * q + 1]) + r.l1 * (cx.l0 * c[2 * q + 1] + cx.l1 * d[2 * q + 1]);
    po[q] = __floats2bfloat162_rn(e0, e1);
  }
  return o;
}
// One 64-channel halo block of HW x HALO_H pixels, written with TMA's 128-byte swizzle; IT producer threads, thread t.
template <int HW, int PW, int IT>
__device__ __forceinline__ void interp_halo_fast(const ConvParams& p, const uint8_t* patch, uint8_t* stage, int sy0, int sx0, int y0,
                                                 int x0, int t) {
  constexpr int BX = HW / 2, BY = HALO_H / 2;
  const int j = t & 7;
  const uint4* src = reinterpret_cast<const uint4*>(patch) + j;
  auto put = [&](int hy, int hx, const uint4& v) {
    const int px = hy * HW + hx;
    *reinterpret_cast<uint4*>(stage + px * 128 + ((j ^ (px & 7)) << 4)) = v;
  };
  auto at = [&](int yy, int xx) { return src[((yy - sy0) * PW + (xx - sx0)) * 8]; };
  auto single = [&](const InterpAxis& r, const InterpAxis& c) {
    float a[8], b[8], cc[8], d[8];
    interp_unpack(at(r.i0, c.i0), a);
    interp_unpack(at(r.i0, c.i1), b);
    interp_unpack(at(r.i1, c.i0), cc);
    interp_unpack(at(r.i1, c.i1), d);
    return interp_emit(a, b, cc, d, r, c);
  };
  const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll 1
  for (int b = t >> 3; b < BX * BY; b += IT / 8) {
    const int by = b / BX, bx = b - by * BX;
    const int hy = 2 * by, hx = 2 * bx;
    const int ya = y0 + hy, yb = ya + 1, xa = x0 + hx, xb = xa + 1;
    const bool ua = ya >= 0 && ya < p.H, ub = yb >= 0 && yb < p.H, va = xa >= 0 && xa < p.W, vb = xb >= 0 && xb < p.W;
    if (!(ua || ub) || !(va || vb)) {  // the whole block lies in the padding
      put(hy, hx, zero);
      put(hy, hx + 1, zero);
      put(hy + 1, hx, zero);
      put(hy + 1, hx + 1, zero);
      continue;
    }
    const InterpAxis ra = interp_axis(p.ry, ua ? ya : yb, p.low_h), rb = interp_axis(p.ry, ub ? yb : ya, p.low_h);
    const InterpAxis ca = interp_axis(p.rx, va ? xa : xb, p.low_w), cb = interp_axis(p.rx, vb ? xb : xa, p.low_w);
    if (ra.i0 == rb.i0 && ra.i1 == rb.i1 && ca.i0 == cb.i0 && ca.i1 == cb.i1) {
      float v00[8], v01[8], v10[8], v11[8];
      interp_unpack(at(ra.i0, ca.i0), v00);
      interp_unpack(at(ra.i0, ca.i1), v01);
      interp_unpack(at(ra.i1, ca.i0), v10);
      interp_unpack(at(ra.i1, ca.i1), v11);
      put(hy, hx, (ua && va) ? interp_emit(v00, v01, v10, v11, ra, ca) : zero);
      put(hy, hx + 1, (ua && vb) ? interp_emit(v00, v01, v10, v11, ra, cb) : zero);
      put(hy + 1, hx, (ub && va) ? interp_emit(v00, v01, v10, v11, rb, ca) : zero);
      put(hy + 1, hx + 1, (ub && vb) ? interp_emit(v00, v01, v10, v11, rb, cb) : zero);
    } else {
      put(hy, hx, (ua && va) ? single(ra, ca) : zero);
      put(hy, hx + 1, (ua && vb) ? single(ra, cb) : zero);
      put(hy + 1, hx, (ub && va) ? single(rb, ca) : zero);
      put(hy + 1, hx + 1, (ub && vb) ? single(rb, cb) : zero);
    }
  }
}

// first source row / column a tile's halo needs (monotone in the output coordinate, so it matches interp_chunk)
__device__ __forceinline__ int patch_origin(float r, int first_out, int pad) {
  const int o = first_out - pad;
  return (int)(r * (float)(o > 0 ? o : 0));
}

template <int BN, bool RESIDENT, bool INTERP, int TILES, bool RAW = false>
__global__ void __launch_bounds__(conv_threads(INTERP, BN, TILES), 1)
k_conv3x3_halo(const __grid_constant__ CUtensorMap map_src0, const __grid_constant__ CUtensorMap map_src1,
               const __grid_constant__ CUtensorMap map_w, const ConvParams p, const HaloCfg cfg) {
  constexpr int B_BLOCK = BN * BLOCK_K * 2;  // one 64-wide weight slice
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int ncb = p.kb0 + p.kb1;
  uint8_t* smem_a = smem;
  constexpr int HW = halo_w(TILES), HALO_BYTES = halo_bytes(TILES), HALO_STRIDE = halo_stride(TILES);
  constexpr int PATCH_W = patch_w(TILES), PATCH_BYTES = patch_bytes(TILES), PATCH_STRIDE = patch_stride(TILES);
  static_assert(acc_stages(BN, TILES) * TILES * BN <= 512, "accumulators exceed TMEM");
  uint8_t* smem_b = smem + cfg.sa * HALO_STRIDE;
  const int b_bytes = RESIDENT ? 9 * ncb * B_BLOCK : cfg.sb * B_BLOCK;
  uint8_t* smem_l = smem_b + b_bytes;  // low-resolution source patches (INTERP only)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_l + (INTERP ? cfg.sl * PATCH_STRIDE : 0));
  uint64_t* full_a = bars;                          // [HALO_MAX_STAGES]
  uint64_t* empty_a = bars + HALO_MAX_STAGES;       // [HALO_MAX_STAGES]
  uint64_t* full_b = bars + 2 * HALO_MAX_STAGES;    // [HALO_MAX_STAGES]
  uint64_t* empty_b = bars + 3 * HALO_MAX_STAGES;   // [HALO_MAX_STAGES]
  uint64_t* w_full = bars + 4 * HALO_MAX_STAGES;    // [1]
  uint64_t* acc_full = w_full + 1;                  // [MAX_ACC_STAGES]
  uint64_t* acc_empty = acc_full + MAX_ACC_STAGES;  // [MAX_ACC_STAGES]
  uint64_t* full_l = acc_empty + MAX_ACC_STAGES;    // [PATCH_STAGES]
  uint64_t* empty_l = full_l + PATCH_STAGES;        // [PATCH_STAGES]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(empty_l + PATCH_STAGES);
  float* s_aux = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + HALO_BAR_BYTES);
  constexpr int NACC = acc_stages(BN, TILES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  stage_aux(p, s_aux);

  if (threadIdx.x == 0) {
    prefetch_tensormap(&map_src0);
    prefetch_tensormap(&map_src1);
    prefetch_tensormap(&map_w);
    for (int s = 0; s < HALO_MAX_STAGES; ++s) {
      mbar_init(&full_a[s], 1);
      mbar_init(&empty_a[s], 1);
      mbar_init(&full_b[s], 1);
      mbar_init(&empty_b[s], 1);
    }
    mbar_init(w_full, 1);
    for (int s = 0; s < MAX_ACC_STAGES; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 4);
    }
    for (int s = 0; s < PATCH_STAGES; ++s) {
      mbar_init(&full_l[s], 1);
      mbar_init(&empty_l[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, NACC * TILES * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      if (RESIDENT) {
        mbar_arrive_expect_tx(w_full, 9 * ncb * B_BLOCK);
        for (int cb = 0; cb < ncb; ++cb)
          for (int tap = 0; tap < 9; ++tap)
            tma_load_2d(smem_b + (cb * 9 + tap) * B_BLOCK, &map_w, tap * p.cin_pad + cb * BLOCK_K, 0, w_full);
      }
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, sl = 0, pl = 0;
      for (int work = blockIdx.x; work < p.total_work; work += gridDim.x) {
        const int nt = work % p.n_tiles;
        int m = work / p.n_tiles;
        const int tx = m % p.tiles_x; m /= p.tiles_x;
        const int ty = m % p.tiles_y;
        const int n = m / p.tiles_y;
        const int x0 = tx * 8 * TILES, y0 = ty * 16;
        // INTERP: source patches of the interpolated blocks, PATCH_STAGES ahead of the block being consumed (a patch
        // slot frees when the interpolation warps finish a block, which never depends on loads not yet issued)
        const int sy0 = INTERP ? patch_origin(p.ry, y0 - 1, p.pad_y) : 0, sx0 = INTERP ? patch_origin(p.rx, x0 - 1, p.pad_x) : 0;
        int next_patch = p.kb0;
        auto issue_patch = [&]() {
          mbar_wait(&empty_l[sl], pl ^ 1);
          mbar_arrive_expect_tx(&full_l[sl], PATCH_BYTES);
          tma_load_4d(smem_l + sl * PATCH_STRIDE, &map_src1, (next_patch - p.kb0) * BLOCK_K, sx0, sy0, n, &full_l[sl]);
          if (++sl == (uint32_t)cfg.sl) { sl = 0; pl ^= 1; }
          ++next_patch;
        };
        if (INTERP) {
          for (int i = 0; i < cfg.sl && next_patch < ncb; ++i) issue_patch();
        }
        for (int cb = 0; cb < ncb; ++cb) {
          const CUtensorMap* src = cb < p.kb0 ? &map_src0 : &map_src1;
          const int c0 = (cb < p.kb0 ? cb : cb - p.kb0) * BLOCK_K;
          // Every party that walks the ring waits for every stage, also the ones it does not fill: parity waits only
          // tell adjacent phases apart, so nobody may run more than one round ahead of the consumer.
          mbar_wait(&empty_a[sa], pa ^ 1);
          if (!INTERP || cb < p.kb0) {  // with INTERP the interpolation warps fill the stages of source 1
            mbar_arrive_expect_tx(&full_a[sa], HALO_BYTES);
            tma_load_4d(smem_a + sa * HALO_STRIDE, src, c0, x0 - 1, y0 - 1, n, &full_a[sa]);
          }
          if (++sa == (uint32_t)cfg.sa) { sa = 0; pa ^= 1; }
          if (!RESIDENT) {
#pragma unroll 1
            for (int tap = 0; tap < 9; ++tap) {
              mbar_wait(&empty_b[sb], pb ^ 1);
              mbar_arrive_expect_tx(&full_b[sb], B_BLOCK);
              tma_load_2d(smem_b + sb * B_BLOCK, &map_w, tap * p.cin_pad + cb * BLOCK_K, nt * BN, &full_b[sb]);
              if (++sb == (uint32_t)cfg.sb) { sb = 0; pb ^= 1; }
            }
          }
          if (INTERP && cb >= p.kb0 && next_patch < ncb) issue_patch();
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer: the whole warp walks the loop, one elected lane issues =================
    if (tmem_base == 0)
      halo_mma_loop<BN, RESIDENT, true, TILES>(p, cfg, smem_a, smem_b, full_a, empty_a, full_b, empty_b, w_full, acc_full, acc_empty, 0);
    else
      halo_mma_loop<BN, RESIDENT, false, TILES>(p, cfg, smem_a, smem_b, full_a, empty_a, full_b, empty_b, w_full, acc_full, acc_empty, tmem_base);
  } else if (INTERP && warp >= 2 + 4 * epi_groups(BN, INTERP)) {
    // ================= interpolation producers (last warps): halo blocks of the upsampled branch =================
    constexpr int IT = interp_warps(BN, TILES) * 32;
    const int t = threadIdx.x - (2 + 4 * epi_groups(BN, INTERP)) * 32;   // 0..IT-1
    const int j = t & 7;                  // 16-byte chunk (8 channels) this thread owns
    uint32_t sa = 0, pa = 0, sl = 0, pl = 0;
    for (int work = blockIdx.x; work < p.total_work; work += gridDim.x) {
      int m = work / p.n_tiles;
      const int tx = m % p.tiles_x; m /= p.tiles_x;
      const int ty = m % p.tiles_y;
      const int x0 = tx * 8 * TILES - 1, y0 = ty * 16 - 1;
      const int sy0 = patch_origin(p.ry, y0, p.pad_y), sx0 = patch_origin(p.rx, x0, p.pad_x);
      for (int cb = 0; cb < ncb; ++cb) {
        mbar_wait(&empty_a[sa], pa ^ 1);  // also for stages the TMA thread fills (see the producer's note)
        if (cb >= p.kb0) {
          mbar_wait(&full_l[sl], pl);
          uint8_t* stage = smem_a + sa * HALO_STRIDE;
          const uint8_t* patch = smem_l + sl * PATCH_STRIDE;
          if (p.interp_fast) {
            interp_halo_fast<HW, PATCH_W, IT>(p, patch, stage, sy0, sx0, y0, x0, t);
          } else {
#pragma unroll 4
            for (int px = t >> 3; px < HW * HALO_H; px += IT / 8) {
              const int hy = px / HW, hx = px - hy * HW;
              const uint4 v = interp_chunk(p, patch, PATCH_W, sy0, sx0, y0 + hy, x0 + hx, j);
              *reinterpret_cast<uint4*>(stage + px * 128 + ((j ^ (px & 7)) << 4)) = v;  // TMA's 128-byte swizzle
            }
          }
          fence_proxy_async();
          asm volatile("bar.sync 1, %0;" ::"n"(IT) : "memory");
          if (t == 0) {
            mbar_arrive(&full_a[sa]);
            mbar_arrive(&empty_l[sl]);
          }
          if (++sl == (uint32_t)cfg.sl) { sl = 0; pl ^= 1; }
        }
        if (++sa == (uint32_t)cfg.sa) { sa = 0; pa ^= 1; }
      }
    }
  } else {
    // ================= epilogue (groups of 4 warps; with two groups each owns one accumulator stage) =================
    constexpr int EG = epi_groups(BN, INTERP);
    const int group = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    RDP_DEV(long long epi_wait = 0; const long long epi_begin = clock64();)
    int it = 0;
    for (int work = blockIdx.x; work < p.total_work; work += gridDim.x, ++it) {
      if (EG > 1 && (it % EG) != group) continue;
      const uint32_t as = it % NACC, aphase = (it / NACC) & 1;
      const int nt = work % p.n_tiles;
      int m = work / p.n_tiles;
      const int tx = m % p.tiles_x; m /= p.tiles_x;
      const int ty = m % p.tiles_y;
      const int n = m / p.tiles_y;
      const int y = ty * 16 + (row >> 3);
      RDP_DEV(const long long e0 = clock64();)
      mbar_wait(&acc_full[as], aphase);
      RDP_DEV(epi_wait += clock64() - e0;)
      tc_fence_after();
#pragma unroll
      for (int h = 0; h < TILES; ++h) {
        const int x = tx * 8 * TILES + h * 8 + (row & 7);
        const bool inside = y < p.H && x < p.W;
        const uint32_t taddr = tmem_base + (as * TILES + h) * BN + ((uint32_t)(quarter * 32) << 16);
        epilogue_tile<BN, true, RAW>(p, s_aux, taddr, nt, n, y, x, inside, h == TILES - 1 ? &acc_empty[as] : nullptr, lane);
      }
    }
    RDP_DEV(if (p.debug_waits && blockIdx.x == 0 && threadIdx.x == 64)
      printf("  epilogue warp: total %lld cyc, waiting acc_full %lld\n", clock64() - epi_begin, epi_wait);)
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, NACC * TILES * BN);
  }
}


// ---------------------------------------------------------------------------------------------------------
// k_conv3x3_pair: the halo kernel on CTA pairs (tcgen05 cta_group::2) for the 64- and 128-wide layers
// ---------------------------------------------------------------------------------------------------------
// With both operands in shared memory a 128 x N x 16 MMA reads 4 KB of A and N*32 bytes of B: at 128 bytes per cycle
// that is 48 cycles for 32 cycles of tensor work at N = 64 and 64 for 64 at N = 128 - the full-resolution layers are
// bounded by operand reads, not by the tensor pipe.  A pair of CTAs on the two SMs of a TPC issues one M = 256 MMA:
// each CTA keeps its own 128-pixel tile (its own halo blocks) and only HALF of the weight slice, so the reads per MMA
// drop to 4 KB + N*16 bytes per SM (40 cycles at N = 64, 48 at N = 128), and the halved weights stay resident in shared
// memory for every 64-wide layer and for down1.  Layout, halo addressing, epilogue and the fused up-sampling are those
// of k_conv3x3_halo (TILES = 1); what changes is the synchronisation:
//   * the leader (cluster rank 0) issues every MMA; the barriers it waits on (full_a, full_b, w_full, acc_empty) live
//     in ITS shared memory and expect one arrival per CTA (acc_empty: four epilogue warps per CTA); the peer reaches them
//     through their shared::cluster address, its TMA loads complete on them directly (cp.async.bulk.tensor .cta_group::2);
//   * the barriers the producers and epilogues wait on (empty_a, empty_b, acc_full) are local to each CTA and are
//     signalled in both CTAs at once by tcgen05.commit ... .multicast::cluster;
//   * tensor memory is allocated and freed by one warp in each CTA with the cta_group::2 forms, with cluster-wide
//     barriers after set-up and before tear-down (the peer's shared memory and barriers must outlive the leader's MMAs).
// Work item = two spatial tiles with consecutive indices (the leader takes the even one); an odd tile count leaves the
// last peer a tile beyond the batch: its loads are zero-filled by the TMA unit and its epilogue stores nothing.
template <int BN, bool RESIDENT, bool INTERP>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(conv_threads(INTERP, BN, 1), 1)
k_conv3x3_pair(const __grid_constant__ CUtensorMap map_src0, const __grid_constant__ CUtensorMap map_src1,
               const __grid_constant__ CUtensorMap map_w, const ConvParams p, const HaloCfg cfg) {
  constexpr int B_HALF = (BN / 2) * BLOCK_K * 2;  // this CTA's half of one 64-wide weight slice
  constexpr int HW = halo_w(1), HALO_BYTES = halo_bytes(1), HALO_STRIDE = halo_stride(1);
  constexpr int PATCH_W = patch_w(1), PATCH_BYTES = patch_bytes(1), PATCH_STRIDE = patch_stride(1);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int ncb = p.kb0 + p.kb1;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + cfg.sa * HALO_STRIDE;
  const int b_bytes = RESIDENT ? 9 * ncb * B_HALF : cfg.sb * 9 * B_HALF;  // streamed: a stage = the nine taps of one channel block
  uint8_t* smem_l = smem_b + b_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_l + (INTERP ? cfg.sl * PATCH_STRIDE : 0));
  uint64_t* full_a = bars;                          // [HALO_MAX_STAGES]  leader's is used (2 arrivals)
  uint64_t* empty_a = bars + HALO_MAX_STAGES;       // [HALO_MAX_STAGES]  local, multicast commit
  uint64_t* full_b = bars + 2 * HALO_MAX_STAGES;    // [HALO_MAX_STAGES]  leader's
  uint64_t* empty_b = bars + 3 * HALO_MAX_STAGES;   // [HALO_MAX_STAGES]  local
  uint64_t* w_full = bars + 4 * HALO_MAX_STAGES;    // [1]                leader's
  uint64_t* acc_full = w_full + 1;                  // [MAX_ACC_STAGES]   local
  uint64_t* acc_empty = acc_full + MAX_ACC_STAGES;  // [MAX_ACC_STAGES]   leader's (8 arrivals)
  uint64_t* full_l = acc_empty + MAX_ACC_STAGES;    // [PATCH_STAGES]     local
  uint64_t* empty_l = full_l + PATCH_STAGES;        // [PATCH_STAGES]     local
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(empty_l + PATCH_STAGES);
  float* s_aux = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + HALO_BAR_BYTES);
  constexpr int NACC = acc_stages(BN, 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int m_total = p.B * p.tiles_y * p.tiles_x;       // spatial tiles (n_tiles == 1 for these layers)
  const int pair_work = (m_total + 1) >> 1;
  stage_aux(p, s_aux);

  if (threadIdx.x == 0) {
    prefetch_tensormap(&map_src0);
    prefetch_tensormap(&map_src1);
    prefetch_tensormap(&map_w);
    for (int s = 0; s < HALO_MAX_STAGES; ++s) {
      mbar_init(&full_a[s], 2);
      mbar_init(&empty_a[s], 1);
      mbar_init(&full_b[s], 2);
      mbar_init(&empty_b[s], 1);
    }
    mbar_init(w_full, 2);
    for (int s = 0; s < MAX_ACC_STAGES; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 8);
    }
    for (int s = 0; s < PATCH_STAGES; ++s) {
      mbar_init(&full_l[s], 1);
      mbar_init(&empty_l[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc_pair(tmem_slot, NACC * BN);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // both CTAs' barriers are initialised and both allocations done before anyone touches the peer
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // tile of this CTA for pair-level work item w
  auto decode = [&](int w, int& tx, int& ty, int& n) {
    int m = 2 * w + (int)rank;
    tx = m % p.tiles_x; m /= p.tiles_x;
    ty = m % p.tiles_y;
    n = m / p.tiles_y;   // n >= B: the tile beyond the batch (loads zero-filled, nothing stored)
  };

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer (both CTAs; completion on the leader's barriers) =================
      if (RESIDENT) {
        const uint32_t wbar = mapa_u32(smem_u32(w_full), 0);
        mbar_arrive_expect_tx_cluster(wbar, 9 * ncb * B_HALF);
        for (int cb = 0; cb < ncb; ++cb)
          for (int tap = 0; tap < 9; ++tap)
            tma_load_2d_pair(smem_b + (cb * 9 + tap) * B_HALF, &map_w, tap * p.cin_pad + cb * BLOCK_K, (int)rank * (BN / 2), wbar);
      }
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, sl = 0, pl = 0;
      for (int work = pair; work < pair_work; work += n_pairs) {
        int tx, ty, n;
        decode(work, tx, ty, n);
        const int x0 = tx * 8, y0 = ty * 16;
        const int sy0 = INTERP ? patch_origin(p.ry, y0 - 1, p.pad_y) : 0, sx0 = INTERP ? patch_origin(p.rx, x0 - 1, p.pad_x) : 0;
        int next_patch = p.kb0;
        auto issue_patch = [&]() {
          mbar_wait(&empty_l[sl], pl ^ 1);
          mbar_arrive_expect_tx(&full_l[sl], PATCH_BYTES);
          tma_load_4d(smem_l + sl * PATCH_STRIDE, &map_src1, (next_patch - p.kb0) * BLOCK_K, sx0, sy0, n, &full_l[sl]);
          if (++sl == (uint32_t)cfg.sl) { sl = 0; pl ^= 1; }
          ++next_patch;
        };
        if (INTERP) {
          for (int i = 0; i < cfg.sl && next_patch < ncb; ++i) issue_patch();
        }
        for (int cb = 0; cb < ncb; ++cb) {
          const CUtensorMap* src = cb < p.kb0 ? &map_src0 : &map_src1;
          const int c0 = (cb < p.kb0 ? cb : cb - p.kb0) * BLOCK_K;
          mbar_wait(&empty_a[sa], pa ^ 1);
          if (!INTERP || cb < p.kb0) {
            const uint32_t abar = mapa_u32(smem_u32(&full_a[sa]), 0);
            mbar_arrive_expect_tx_cluster(abar, HALO_BYTES);
            tma_load_4d_pair(smem_a + sa * HALO_STRIDE, src, c0, x0 - 1, y0 - 1, n, abar);
          }
          if (++sa == (uint32_t)cfg.sa) { sa = 0; pa ^= 1; }
          if (!RESIDENT) {  // one hand-shake per channel block, not per tap: nine half slices land on one barrier
            mbar_wait(&empty_b[sb], pb ^ 1);
            const uint32_t bbar = mapa_u32(smem_u32(&full_b[sb]), 0);
            mbar_arrive_expect_tx_cluster(bbar, 9 * B_HALF);
#pragma unroll 1
            for (int tap = 0; tap < 9; ++tap)
              tma_load_2d_pair(smem_b + (sb * 9 + tap) * B_HALF, &map_w, tap * p.cin_pad + cb * BLOCK_K, (int)rank * (BN / 2), bbar);
            if (++sb == (uint32_t)cfg.sb) { sb = 0; pb ^= 1; }
          }
          if (INTERP && cb >= p.kb0 && next_patch < ncb) issue_patch();
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer: the leader's warp walks the loop converged, one elected lane issues =================
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * TILE_M, BN);
      constexpr uint32_t hi_a = desc_hi(HW * 128), hi_b = desc_hi(1024);
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, aphase = 0;
      const uint32_t a_lo0 = desc_lo(smem_u32(smem_a)), b_lo0 = desc_lo(smem_u32(smem_b));
      RDP_DEV(long long w_acc = 0, w_a = 0, t_begin = clock64(), t0;)
      if (RESIDENT) {
        mbar_wait(w_full, 0);
        tc_fence_after();
      }
      int work = pair;
      while (work < pair_work) {
#pragma unroll
        for (int as = 0; as < NACC; ++as) {
          if (work < pair_work) {
            RDP_DEV(t0 = clock64();)
            mbar_wait(&acc_empty[as], aphase ^ 1);   // a drained tensor-memory stage: no memory to acquire
            RDP_DEV(w_acc += clock64() - t0;)
            tc_fence_after();
            const uint32_t tmem_acc = tmem_base + as * BN;
            for (int cb = 0; cb < ncb; ++cb) {
              RDP_DEV(t0 = clock64();)
              if (INTERP) mbar_wait_cluster(&full_a[sa], pa);   // the peer's interpolation warps wrote its block with ordinary stores
              else mbar_wait(&full_a[sa], pa);                   // TMA completions only
              RDP_DEV(w_a += clock64() - t0;)
              tc_fence_after();
              const uint32_t a_lo = a_lo0 + sa * (HALO_STRIDE >> 4);
              uint32_t bres_lo = b_lo0 + cb * (9 * B_HALF >> 4);
              if (!RESIDENT) {
                mbar_wait(&full_b[sb], pb);
                tc_fence_after();
                bres_lo = b_lo0 + sb * (9 * B_HALF >> 4);
              }
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                const uint32_t b_lo = bres_lo + tap * (B_HALF >> 4);
                const uint32_t at_lo = a_lo + ((tap / 3) * HW + tap % 3) * 8;  // 128 bytes per halo row
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                  mma_bf16_ss_pair_elect(tmem_acc, desc_pack(at_lo + 2 * k, hi_a), desc_pack(b_lo + 2 * k, hi_b), idesc, (cb | tap | k) != 0);
              }
              if (!RESIDENT) {
                tc_commit_pair_elect(&empty_b[sb]);
                if (++sb == (uint32_t)cfg.sb) { sb = 0; pb ^= 1; }
              }
              tc_commit_pair_elect(&empty_a[sa]);
              if (++sa == (uint32_t)cfg.sa) { sa = 0; pa ^= 1; }
            }
            tc_commit_pair_elect(&acc_full[as]);
            work += n_pairs;
          }
        }
        aphase ^= 1;
      }
      RDP_DEV(if (p.debug_waits && blockIdx.x == 0 && (threadIdx.x & 31) == 0)
        printf("  pair mma thread: total %lld cyc, waiting acc_empty %lld, full_a %lld (BN %d resident %d)\n", clock64() - t_begin, w_acc, w_a, BN, (int)RESIDENT);)
    }
  } else if (INTERP && warp >= 2 + 4 * epi_groups(BN, INTERP)) {
    // ================= interpolation producers: halo blocks of the up-sampled branch, each CTA its own tile =================
    constexpr int IT = interp_warps(BN, 1) * 32;
    const int t = threadIdx.x - (2 + 4 * epi_groups(BN, INTERP)) * 32;
    const int j = t & 7;
    uint32_t sa = 0, pa = 0, sl = 0, pl = 0;
    for (int work = pair; work < pair_work; work += n_pairs) {
      int tx, ty, n;
      decode(work, tx, ty, n);
      const int x0 = tx * 8 - 1, y0 = ty * 16 - 1;
      const int sy0 = patch_origin(p.ry, y0, p.pad_y), sx0 = patch_origin(p.rx, x0, p.pad_x);
      for (int cb = 0; cb < ncb; ++cb) {
        mbar_wait(&empty_a[sa], pa ^ 1);
        if (cb >= p.kb0) {
          mbar_wait(&full_l[sl], pl);
          uint8_t* stage = smem_a + sa * HALO_STRIDE;
          const uint8_t* patch = smem_l + sl * PATCH_STRIDE;
          if (p.interp_fast) {
            interp_halo_fast<HW, PATCH_W, IT>(p, patch, stage, sy0, sx0, y0, x0, t);
          } else {
#pragma unroll 4
            for (int px = t >> 3; px < HW * HALO_H; px += IT / 8) {
              const int hy = px / HW, hx = px - hy * HW;
              const uint4 v = interp_chunk(p, patch, PATCH_W, sy0, sx0, y0 + hy, x0 + hx, j);
              *reinterpret_cast<uint4*>(stage + px * 128 + ((j ^ (px & 7)) << 4)) = v;
            }
          }
          fence_proxy_async();  // generic-proxy writes -> visible to the tensor core (the leader's MMA reads both CTAs)
          asm volatile("bar.sync 1, %0;" ::"n"(IT) : "memory");
          if (t == 0) {
            mbar_arrive_cluster(mapa_u32(smem_u32(&full_a[sa]), 0));
            mbar_arrive(&empty_l[sl]);
          }
          if (++sl == (uint32_t)cfg.sl) { sl = 0; pl ^= 1; }
        }
        if (++sa == (uint32_t)cfg.sa) { sa = 0; pa ^= 1; }
      }
    }
  } else {
    // ================= epilogue: each CTA drains its own 128 rows =================
    constexpr int EG = epi_groups(BN, INTERP);
    const int group = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    int it = 0;
    RDP_DEV(long long epi_wait = 0; const long long epi_begin = clock64();)
    for (int work = pair; work < pair_work; work += n_pairs, ++it) {
      if (EG > 1 && (it % EG) != group) continue;
      const uint32_t as = it % NACC, aphase = (it / NACC) & 1;
      int tx, ty, n;
      decode(work, tx, ty, n);
      const bool real = n < p.B;
      const int y = real ? ty * 16 + (row >> 3) : p.H + 16, x = tx * 8 + (row & 7);
      const bool inside = y < p.H && x < p.W;
      RDP_DEV(const long long e0 = clock64();)
      mbar_wait(&acc_full[as], aphase);
      RDP_DEV(epi_wait += clock64() - e0;)
      tc_fence_after();
      const uint32_t taddr = tmem_base + as * BN + ((uint32_t)(quarter * 32) << 16);
      epilogue_tile<BN, true>(p, s_aux, taddr, 0, real ? n : 0, y, x, inside, nullptr, lane, mapa_u32(smem_u32(&acc_empty[as]), 0));
    }
    RDP_DEV(if (p.debug_waits && blockIdx.x < 2 && threadIdx.x == 64)
      printf("  pair epilogue warp (rank %u): total %lld cyc, waiting acc_full %lld\n", rank, clock64() - epi_begin, epi_wait);)
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's shared memory, barriers and tensor memory stay alive until the leader is done with them
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, NACC * BN);
  }
}


// ---------------------------------------------------------------------------------------------------------
// k_conv_first: inc.0 (Cin = n_channels, K = 9*Cin <= 64) straight from the camera image
// ---------------------------------------------------------------------------------------------------------
// With 3 input channels the layer is bandwidth-bound (it reads 3 bytes and writes 128 bytes per pixel), so the
// point is to never materialise an im2col matrix in HBM: twelve producer warps gather each pixel's 3x3xC
// neighbourhood (uint8 -> x/255 -> bf16 through a 256-entry table, exactly ToTensor's division) and write it as one
// 128-byte row of the swizzled A tile in shared memory; the tensor core multiplies by the resident 64x64 weight
// block; the usual epilogue adds the BN shift, applies ReLU and stores bf16 NHWC.
constexpr int FIRST_STAGES = 4;
constexpr int FIRST_PRODUCER_WARPS = 12;  // three groups of 128 threads
constexpr int FIRST_MMA_WARP = FIRST_PRODUCER_WARPS;
constexpr int FIRST_EPI_GROUPS = 2;  // each group of 4 warps owns one accumulator stage and one output staging tile
constexpr int FIRST_THREADS = (FIRST_PRODUCER_WARPS + 1 + 4 * FIRST_EPI_GROUPS) * 32;
constexpr int FIRST_OUT_BYTES = TILE_M * 64 * 2;  // bf16 output tile staged for the TMA store
constexpr int FIRST_SMEM = FIRST_STAGES * A_STAGE_BYTES + 64 * BLOCK_K * 2 + 2 * FIRST_OUT_BYTES + 1024 + 256 + AUX_BYTES + 512;

template <int C, bool U8>
__global__ void __launch_bounds__(FIRST_THREADS, 1)
k_conv_first(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_out, const ConvParams p,
             const void* __restrict__ in) {
  constexpr int BN = 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + FIRST_STAGES * A_STAGE_BYTES;
  uint8_t* smem_o = smem_b + BN * BLOCK_K * 2;  // two output tiles (one per accumulator stage), 1024-byte aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_o + 2 * FIRST_OUT_BYTES);
  uint64_t* a_full = bars;                   // [FIRST_STAGES], 128 producer arrivals each
  uint64_t* a_empty = bars + FIRST_STAGES;   // [FIRST_STAGES]
  uint64_t* w_full = bars + 2 * FIRST_STAGES;
  uint64_t* acc_full = w_full + 1;           // [2]
  uint64_t* acc_empty = w_full + 3;          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 5);
  float* s_aux = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);
  uint16_t* s_lut = reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(s_aux) + AUX_BYTES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    prefetch_tensormap(&map_w);
    prefetch_tensormap(&map_out);
    for (int s = 0; s < FIRST_STAGES; ++s) {
      mbar_init(&a_full[s], 128);
      mbar_init(&a_empty[s], 1);
    }
    mbar_init(w_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 4);
    }
    fence_mbar_init();
  }
  if (threadIdx.x < 256) {
    const __nv_bfloat16 v = __float2bfloat16_rn(__fdiv_rn((float)threadIdx.x, 255.0f));  // ToTensor: uint8 / 255
    s_lut[threadIdx.x] = *reinterpret_cast<const uint16_t*>(&v);
  }
  stage_aux(p, s_aux);
  if (warp == FIRST_MMA_WARP) tmem_alloc(tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < FIRST_PRODUCER_WARPS) {
    // ================= im2col producers: thread = one pixel row of the A tile =================
    const int group = warp >> 2;                 // three groups of 128 threads take every third tile
    const int r = (warp & 3) * 32 + lane;
    const int ry = r / p.tile_w, rx = r % p.tile_w;
    int it = 0;
    for (int work = blockIdx.x; work < p.total_work; work += gridDim.x, ++it) {
      if (it % 3 != group) continue;
      const int stage = it % FIRST_STAGES;
      const uint32_t phase = (uint32_t)(it / FIRST_STAGES) & 1u;
      int m = work;
      const int tx = m % p.tiles_x; m /= p.tiles_x;
      const int ty = m % p.tiles_y;
      const int n = m / p.tiles_y;
      const int y = ty * p.tile_h + ry, x = tx * p.tile_w + rx;
      // all 9*C loads are issued before the first table look-up (one dependent chain per tile, not 27)
      uint16_t v[64];
#pragma unroll
      for (int k = 0; k < 64; ++k) v[k] = 0;
      if (y < p.H && x < p.W) {
        if (U8) {
          uint8_t raw[9 * C];
          bool ok[9];
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const int yy = y + tap / 3 - 1, xx = x + tap % 3 - 1;
            ok[tap] = yy >= 0 && yy < p.H && xx >= 0 && xx < p.W;
            const uint8_t* src = (const uint8_t*)in + (((size_t)n * p.H + (ok[tap] ? yy : y)) * p.W + (ok[tap] ? xx : x)) * C;
#pragma unroll
            for (int c = 0; c < C; ++c) raw[tap * C + c] = __ldg(src + c);
          }
#pragma unroll
          for (int tap = 0; tap < 9; ++tap)
#pragma unroll
            for (int c = 0; c < C; ++c) v[tap * C + c] = ok[tap] ? s_lut[raw[tap * C + c]] : (uint16_t)0;
        } else {
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const int yy = y + tap / 3 - 1, xx = x + tap % 3 - 1;
            if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) {
#pragma unroll
              for (int c = 0; c < C; ++c) {
                const __nv_bfloat16 b = __float2bfloat16_rn(__ldg((const float*)in + (((size_t)n * C + c) * p.H + yy) * p.W + xx));
                v[tap * C + c] = *reinterpret_cast<const uint16_t*>(&b);
              }
            }
          }
        }
      }
      mbar_wait(&a_empty[stage], phase ^ 1);
      uint8_t* rowp = smem_a + stage * A_STAGE_BYTES + r * 128;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint4 o;
        o.x = (uint32_t)v[8 * j + 0] | ((uint32_t)v[8 * j + 1] << 16);
        o.y = (uint32_t)v[8 * j + 2] | ((uint32_t)v[8 * j + 3] << 16);
        o.z = (uint32_t)v[8 * j + 4] | ((uint32_t)v[8 * j + 5] << 16);
        o.w = (uint32_t)v[8 * j + 6] | ((uint32_t)v[8 * j + 7] << 16);
        *reinterpret_cast<uint4*>(rowp + ((j ^ (r & 7)) << 4)) = o;  // the 128-byte swizzle TMA would have applied
      }
      fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's async proxy
      mbar_arrive(&a_full[stage]);
    }
  } else if (warp == FIRST_MMA_WARP) {
    {
      // ================= weights + MMA issuer: the warp walks the loop converged, one elected lane issues =================
      // (a single-lane loop makes ptxas wrap every tcgen05.mma in a vector-to-uniform loop: ~180 cycles per 48-cycle MMA)
      if (lane == 0) {
        mbar_arrive_expect_tx(w_full, BN * BLOCK_K * 2);
        tma_load_2d(smem_b, &map_w, 0, 0, w_full);
      }
      __syncwarp();
      mbar_wait(w_full, 0);
      tc_fence_after();
      constexpr uint32_t idesc = make_idesc_bf16(TILE_M, BN);
      const uint64_t db = make_sw128_desc(smem_u32(smem_b));
      uint32_t as = 0, aphase = 0;
      int it = 0;
      RDP_DEV(long long w_acc = 0, w_a = 0, t_begin = clock64(), t0;)
      for (int work = blockIdx.x; work < p.total_work; work += gridDim.x, ++it) {
        const int stage = it % FIRST_STAGES;
        const uint32_t phase = (uint32_t)(it / FIRST_STAGES) & 1u;
        RDP_DEV(t0 = clock64();)
        mbar_wait(&acc_empty[as], aphase ^ 1);
        RDP_DEV(w_acc += clock64() - t0; t0 = clock64();)
        mbar_wait(&a_full[stage], phase);
        RDP_DEV(w_a += clock64() - t0;)
        tc_fence_after();
        const uint64_t da = make_sw128_desc(smem_u32(smem_a + stage * A_STAGE_BYTES));
#pragma unroll
        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) mma_bf16_ss_elect(tmem_base + as * BN, da + 2 * k, db + 2 * k, idesc, k != 0);
        tc_commit_elect(&a_empty[stage]);
        tc_commit_elect(&acc_full[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      RDP_DEV(if (p.debug_waits && blockIdx.x == 0 && lane == 0)
        printf("first layer mma thread: total %lld cyc, waiting acc_empty (epilogue) %lld, a_full (producers) %lld\n",
               clock64() - t_begin, w_acc, w_a);)
    }
  } else {
    // ================= epilogue: TMEM -> shift, ReLU, bf16 -> swizzled smem tile -> one TMA store =================
    // A thread owns a pixel (128 bytes of channels); storing that straight to global memory touches 32 different
    // 128-byte lines per warp instruction.  Staging the tile and letting the TMA unit write it (full lines, image
    // borders clipped by the tensor map) removes that cost from the epilogue, which is this layer's bottleneck.
    const int group = (warp - FIRST_MMA_WARP - 1) >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const bool issuer = ((warp - FIRST_MMA_WARP - 1) & 3) == 0 && lane == 0;
    int it = 0;
    for (int work = blockIdx.x; work < p.total_work; work += gridDim.x, ++it) {
      if ((it % FIRST_EPI_GROUPS) != group) continue;
      const uint32_t as = it & 1, aphase = (it >> 1) & 1;   // with two groups: as == group
      int m = work;
      const int tx = m % p.tiles_x; m /= p.tiles_x;
      const int ty = m % p.tiles_y;
      const int n = m / p.tiles_y;
      mbar_wait(&acc_full[as], aphase);
      tc_fence_after();
      if (issuer) tma_store_wait_read<0>();       // this group's previous store has finished reading the buffer
      asm volatile("bar.sync %0, 128;" ::"r"(2 + group) : "memory");
      const uint32_t taddr = tmem_base + as * BN + ((uint32_t)(quarter * 32) << 16);
      uint8_t* orow = smem_o + as * FIRST_OUT_BYTES + row * 128;
      uint32_t r[2][32];
      tmem_ld_32x32(taddr, r[0]);
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        tmem_ld_wait();
        if (c + 1 < BN / 32) tmem_ld_32x32(taddr + (c + 1) * 32, r[(c + 1) & 1]);
        const uint32_t sh_addr = smem_u32(s_aux + c * 32);
#pragma unroll
        for (int e = 0; e < 4; ++e) {   // four 16-byte pieces of 8 channels
          const float4 s0 = lds_f4(sh_addr + 32 * e), s1 = lds_f4(sh_addr + 32 * e + 16);
          uint4 o;
          o.x = pack_bf16x2<true>(__uint_as_float(r[c & 1][8 * e + 0]) + s0.x, __uint_as_float(r[c & 1][8 * e + 1]) + s0.y);
          o.y = pack_bf16x2<true>(__uint_as_float(r[c & 1][8 * e + 2]) + s0.z, __uint_as_float(r[c & 1][8 * e + 3]) + s0.w);
          o.z = pack_bf16x2<true>(__uint_as_float(r[c & 1][8 * e + 4]) + s1.x, __uint_as_float(r[c & 1][8 * e + 5]) + s1.y);
          o.w = pack_bf16x2<true>(__uint_as_float(r[c & 1][8 * e + 6]) + s1.z, __uint_as_float(r[c & 1][8 * e + 7]) + s1.w);
          *reinterpret_cast<uint4*>(orow + (((c * 4 + e) ^ (row & 7)) << 4)) = o;   // the tensor map's 128-byte swizzle
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);
      fence_proxy_async();                        // smem writes -> visible to the TMA unit
      asm volatile("bar.sync %0, 128;" ::"r"(2 + group) : "memory");
      if (issuer && p.out != nullptr) {
        tma_store_4d(&map_out, smem_o + as * FIRST_OUT_BYTES, 0, tx * p.tile_w, ty * p.tile_h, n);
        tma_store_commit();
      }
    }
    if (issuer) tma_store_wait_read<0>();         // shared memory must outlive the last store's reads
  }

  tc_fence_before();
  __syncthreads();
  if (warp == FIRST_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ---------------------------------------------------------------------------------------------------------
// bandwidth-bound helpers
// ---------------------------------------------------------------------------------------------------------
// First layer: gather the 3x3xC neighbourhood of every pixel into 64 bf16 (k = (3*ky+kx)*C + c, zero padded),
// so that inc.0 runs through the same tensor-core kernel as a 1x1 convolution with K = 64.
// grid = (ceil(W / 32), H, B): a CTA makes 32 pixels of one image row.  The three input rows it needs (34 pixels x C
// channels each, zero outside the image) are staged in shared memory with coalesced loads; a thread then assembles one
// 16-byte chunk (8 of the 64 values) of one pixel.  C is a template parameter so that k / C and k % C are constants.
template <int C>
__global__ void __launch_bounds__(256) k_im2col_first(const void* __restrict__ in, int in_fmt, int H, int W, __nv_bfloat16* __restrict__ out) {
  __shared__ float s_in[3][34][C];
  const int x0 = blockIdx.x * 32, y = blockIdx.y, n = blockIdx.z;
  for (int i = threadIdx.x; i < 3 * 34 * C; i += 256) {
    int r, xx, c;
    if (in_fmt == RDP_IN_RGB_U8_NHWC) {   // (row, pixel, channel): channels contiguous in memory
      c = i % C;
      xx = (i / C) % 34;
      r = i / (C * 34);
    } else {                              // NCHW: pixels contiguous
      xx = i % 34;
      c = (i / 34) % C;
      r = i / (34 * C);
    }
    const int yy = y + r - 1, gx = x0 + xx - 1;
    float v = 0.f;
    if (yy >= 0 && yy < H && gx >= 0 && gx < W) {
      if (in_fmt == RDP_IN_RGB_U8_NHWC)
        v = (float)__ldg((const uint8_t*)in + (((size_t)n * H + yy) * W + gx) * C + c) / 255.0f;  // ToTensor
      else
        v = __ldg((const float*)in + (((size_t)n * C + c) * H + yy) * W + gx);
    }
    s_in[r][xx][c] = v;
  }
  __syncthreads();
  const int chunk = threadIdx.x & 7, px = threadIdx.x >> 3;
  if (x0 + px >= W) return;
  __align__(16) __nv_bfloat16 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int k = chunk * 8 + j;
    float val = 0.f;
    if (k < 9 * C) {
      const int tap = k / C, c = k - tap * C;
      val = s_in[tap / 3][px + tap % 3][c];
    }
    v[j] = __float2bfloat16_rn(val);
  }
  *reinterpret_cast<uint4*>(out + (((size_t)n * H + y) * W + x0 + px) * 64 + chunk * 8) = *reinterpret_cast<const uint4*>(v);
}

static void launch_im2col_first(const void* in, int in_fmt, int B, int H, int W, int C, __nv_bfloat16* out, cudaStream_t stream) {
  const dim3 grid((W + 31) / 32, H, B);
  switch (C) {
    case 1: k_im2col_first<1><<<grid, 256, 0, stream>>>(in, in_fmt, H, W, out); break;
    case 2: k_im2col_first<2><<<grid, 256, 0, stream>>>(in, in_fmt, H, W, out); break;
    case 3: k_im2col_first<3><<<grid, 256, 0, stream>>>(in, in_fmt, H, W, out); break;
    case 4: k_im2col_first<4><<<grid, 256, 0, stream>>>(in, in_fmt, H, W, out); break;
    case 5: k_im2col_first<5><<<grid, 256, 0, stream>>>(in, in_fmt, H, W, out); break;
    case 6: k_im2col_first<6><<<grid, 256, 0, stream>>>(in, in_fmt, H, W, out); break;
    default: k_im2col_first<7><<<grid, 256, 0, stream>>>(in, in_fmt, H, W, out); break;
  }
}

__device__ __forceinline__ uint4 bf16x8_max(uint4 a, uint4 b) {
  uint4 r;
  const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
  __nv_bfloat162* pr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) pr[i] = __hmax2(pa[i], pb[i]);
  return r;
}

// MaxPool2d(2) (segmentation_model.py:47) on NHWC bf16, 8 channels (16 bytes) per thread.
// grid = (ceil(Wo*C8 / 256), Ho, B): no 64-bit divisions on the streaming path.
__global__ void __launch_bounds__(256) k_maxpool2(const __nv_bfloat16* in, int H, int W, int C8, int c8_shift,
                                                  __nv_bfloat16* out) {
  const int Ho = H >> 1, Wo = W >> 1;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= Wo * C8) return;
  const int xo = c8_shift >= 0 ? t >> c8_shift : t / C8, c8 = t - xo * C8;
  const int yo = blockIdx.y, n = blockIdx.z;
  const uint4* src = reinterpret_cast<const uint4*>(in) + (((size_t)n * H + 2 * yo) * W + 2 * xo) * C8 + c8;
  const size_t rowstride = (size_t)W * C8;
  const uint4 a = __ldg(src), b = __ldg(src + C8), c = __ldg(src + rowstride), d = __ldg(src + rowstride + C8);
  reinterpret_cast<uint4*>(out)[(((size_t)n * Ho + yo) * Wo + xo) * C8 + c8] = bf16x8_max(bf16x8_max(a, b), bf16x8_max(c, d));
}

// nn.Upsample(scale_factor=2, bilinear, align_corners=True) followed by the zero pad of Up.forward
// (segmentation_model.py:61,68-74): (B,h,w,C) -> (B,Hs,Ws,C), interpolated values placed at offset (py, px).
// grid = (ceil(Ws*C8 / 256), Hs, B).
__global__ void __launch_bounds__(256) k_upsample2x(const __nv_bfloat16* in, int h, int w, int C8, int c8_shift, int Hs,
                                                    int Ws, int py, int px, __nv_bfloat16* out) {
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= Ws * C8) return;
  const int xs = c8_shift >= 0 ? t >> c8_shift : t / C8, c8 = t - xs * C8;
  const int ys = blockIdx.y, n = blockIdx.z;
  const int Ho = 2 * h, Wo = 2 * w;
  const float ry = Ho > 1 ? (float)(h - 1) / (float)(Ho - 1) : 0.f;
  const float rx = Wo > 1 ? (float)(w - 1) / (float)(Wo - 1) : 0.f;
  const int yo = ys - py, xo = xs - px;
  uint4 o = make_uint4(0, 0, 0, 0);
  if (yo >= 0 && yo < Ho && xo >= 0 && xo < Wo) {
    const float fy = ry * (float)yo, fx = rx * (float)xo;
    const int y0 = (int)fy, x0 = (int)fx;
    const int y1 = y0 + (y0 < h - 1 ? 1 : 0), x1 = x0 + (x0 < w - 1 ? 1 : 0);
    const float ly1 = fy - (float)y0, ly0 = 1.f - ly1, lx1 = fx - (float)x0, lx0 = 1.f - lx1;
    const uint4* base = reinterpret_cast<const uint4*>(in) + (size_t)n * h * w * C8 + c8;
    const uint4 v00 = __ldg(base + (y0 * w + x0) * C8), v01 = __ldg(base + (y0 * w + x1) * C8);
    const uint4 v10 = __ldg(base + (y1 * w + x0) * C8), v11 = __ldg(base + (y1 * w + x1) * C8);
    const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&v00);
    const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&v01);
    const __nv_bfloat162* c = reinterpret_cast<const __nv_bfloat162*>(&v10);
    const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(&v11);
    __nv_bfloat162* po = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 fa = __bfloat1622float2(a[j]), fb = __bfloat1622float2(b[j]);
      const float2 fc = __bfloat1622float2(c[j]), fd = __bfloat1622float2(d[j]);
      const float r0 = ly0 * (lx0 * fa.x + lx1 * fb.x) + ly1 * (lx0 * fc.x + lx1 * fd.x);
      const float r1 = ly0 * (lx0 * fa.y + lx1 * fb.y) + ly1 * (lx0 * fc.y + lx1 * fd.y);
      po[j] = __floats2bfloat162_rn(r0, r1);
    }
  }
  reinterpret_cast<uint4*>(out)[(((size_t)n * Hs + ys) * Ws + xs) * C8 + c8] = o;
}

// The same interpolation with the per-thread overhead taken out (the kernel above is bound by instruction issue, about
// 250 instructions per 16 bytes written): the two scale factors come from the host (the same IEEE fp32 divisions), the
// row-dependent part (source rows, vertical weights) is computed once per thread, and a thread then walks UPS_PX pixels
// of its output row, one 8-channel group each, 256 / C8 pixels apart so that a warp still reads and writes whole
// lines.  The arithmetic per element is the expression above, term for term: results are bit-identical.
constexpr int UPS_PX = 4;
__global__ void __launch_bounds__(256) k_upsample2x_row(const __nv_bfloat16* __restrict__ in, int h, int w, int C8, int c8_shift,
                                                        int Hs, int Ws, int py, int px, float ry, float rx,
                                                        __nv_bfloat16* __restrict__ out) {
  const int per_cta = (256 >> c8_shift) * UPS_PX;  // pixels of one output row a CTA covers
  const int c8 = threadIdx.x & (C8 - 1), lane_px = threadIdx.x >> c8_shift;
  const int ys = blockIdx.y, n = blockIdx.z;
  const int Ho = 2 * h, Wo = 2 * w;
  const int yo = ys - py;
  const bool row_in = yo >= 0 && yo < Ho;
  const float fy = ry * (float)yo;
  const int y0 = row_in ? (int)fy : 0;
  const int y1 = y0 + (y0 < h - 1 ? 1 : 0);
  const float ly1 = fy - (float)y0, ly0 = 1.f - ly1;
  const uint4* row0 = reinterpret_cast<const uint4*>(in) + ((size_t)n * h + y0) * w * C8 + c8;
  const uint4* row1 = reinterpret_cast<const uint4*>(in) + ((size_t)n * h + y1) * w * C8 + c8;
  uint4* orow = reinterpret_cast<uint4*>(out) + ((size_t)n * Hs + ys) * Ws * C8 + c8;
  const int x_first = blockIdx.x * per_cta + lane_px, x_step = 256 >> c8_shift;
#pragma unroll
  for (int k = 0; k < UPS_PX; ++k) {
    const int xs = x_first + k * x_step;
    if (xs >= Ws) break;
    const int xo = xs - px;
    uint4 o = make_uint4(0, 0, 0, 0);
    if (row_in && xo >= 0 && xo < Wo) {
      const float fx = rx * (float)xo;
      const int x0 = (int)fx;
      const int x1 = x0 + (x0 < w - 1 ? 1 : 0);
      const float lx1 = fx - (float)x0, lx0 = 1.f - lx1;
      const uint4 v00 = __ldg(row0 + x0 * C8), v01 = __ldg(row0 + x1 * C8);
      const uint4 v10 = __ldg(row1 + x0 * C8), v11 = __ldg(row1 + x1 * C8);
      const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&v00);
      const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&v01);
      const __nv_bfloat162* c = reinterpret_cast<const __nv_bfloat162*>(&v10);
      const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(&v11);
      __nv_bfloat162* po = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 fa = __bfloat1622float2(a[j]), fb = __bfloat1622float2(b[j]);
        const float2 fc = __bfloat1622float2(c[j]), fd = __bfloat1622float2(d[j]);
        const float r0 = ly0 * (lx0 * fa.x + lx1 * fb.x) + ly1 * (lx0 * fc.x + lx1 * fd.x);
        const float r1 = ly0 * (lx0 * fa.y + lx1 * fb.y) + ly1 * (lx0 * fc.y + lx1 * fd.y);
        po[j] = __floats2bfloat162_rn(r0, r1);
      }
    }
    orow[(size_t)xs * C8] = o;
  }
}

// Fast path of the stand-alone interpolation for the exact doubling without pad (Hs = 2h, Ws = 2w: every level of an
// image whose sides divide by 16).  With align_corners the output rows 2k-1 and 2k read the same two source rows k-1 and
// k, and likewise the columns: a thread produces that 2x2 block of output pixels for one 8-channel group from FOUR
// loads and one conversion of each (the row kernel above loads and converts four source pixels per output pixel, and
// is bound by instruction issue).  Indices and weights still come from the same fp32 expressions per output row and
// column, and a block whose outputs do not share their sources (the first and last row / column) falls back to
// per-output loads, so the results are those of k_upsample2x bit for bit.
// grid = (ceil((w+1)*C8 / 256), ceil((h+1) / UQ_ROWS), B); thread = (column pair j = 0..w, channel group).
constexpr int UQ_ROWS = 4;
struct UpsAxis { int i0, i1; float l0, l1; };
__device__ __forceinline__ UpsAxis ups_axis(float r, int o, int n) {
  UpsAxis a;
  const float f = r * (float)o;
  a.i0 = (int)f;
  a.i1 = a.i0 + (a.i0 < n - 1 ? 1 : 0);
  a.l1 = f - (float)a.i0;
  a.l0 = 1.f - a.l1;
  return a;
}
__device__ __forceinline__ void bf16x8_to_f32(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__global__ void __launch_bounds__(256) k_upsample2x_quad(const __nv_bfloat16* __restrict__ in, int h, int w, int C8, int c8_shift,
                                                         float ry, float rx, __nv_bfloat16* __restrict__ out) {
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= (w + 1) * C8) return;
  const int j = t >> c8_shift, c8 = t & (C8 - 1), n = blockIdx.z;
  const int Ho = 2 * h, Wo = 2 * w;
  const int xa = 2 * j - 1, xb = 2 * j;
  const bool va = xa >= 0, vb = xb < Wo;
  const UpsAxis ca = ups_axis(rx, va ? xa : xb, w), cb = ups_axis(rx, vb ? xb : xa, w);
  const bool same_c = ca.i0 == cb.i0 && ca.i1 == cb.i1;
  const uint4* src = reinterpret_cast<const uint4*>(in) + (size_t)n * h * w * C8 + c8;
  uint4* dst = reinterpret_cast<uint4*>(out) + (size_t)n * Ho * Wo * C8 + c8;
  auto emit = [&](const float (&top0)[8], const float (&top1)[8], const float (&bot0)[8], const float (&bot1)[8], const UpsAxis& r,
                  const UpsAxis& c, int yo, int xo) {
    uint4 o;
    __nv_bfloat162* po = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float r0 = r.l0 * (c.l0 * top0[2 * q] + c.l1 * top1[2 * q]) + r.l1 * (c.l0 * bot0[2 * q] + c.l1 * bot1[2 * q]);
      const float r1 = r.l0 * (c.l0 * top0[2 * q + 1] + c.l1 * top1[2 * q + 1]) + r.l1 * (c.l0 * bot0[2 * q + 1] + c.l1 * bot1[2 * q + 1]);
      po[q] = __floats2bfloat162_rn(r0, r1);
    }
    dst[((size_t)yo * Wo + xo) * C8] = o;
  };
  auto generic = [&](const UpsAxis& r, const UpsAxis& c, int yo, int xo) {
    float a[8], b[8], cc[8], d[8];
    bf16x8_to_f32(__ldg(src + ((size_t)r.i0 * w + c.i0) * C8), a);
    bf16x8_to_f32(__ldg(src + ((size_t)r.i0 * w + c.i1) * C8), b);
    bf16x8_to_f32(__ldg(src + ((size_t)r.i1 * w + c.i0) * C8), cc);
    bf16x8_to_f32(__ldg(src + ((size_t)r.i1 * w + c.i1) * C8), d);
    emit(a, b, cc, d, r, c, yo, xo);
  };
#pragma unroll 1
  for (int k = blockIdx.y * UQ_ROWS; k < min(h + 1, (blockIdx.y + 1) * UQ_ROWS); ++k) {
    const int ya = 2 * k - 1, yb = 2 * k;
    const bool ua = ya >= 0, ub = yb < Ho;
    const UpsAxis ra = ups_axis(ry, ua ? ya : yb, h), rb = ups_axis(ry, ub ? yb : ya, h);
    const bool same_r = ra.i0 == rb.i0 && ra.i1 == rb.i1;
    if (same_r && same_c) {
      float v00[8], v01[8], v10[8], v11[8];
      const uint4 q00 = __ldg(src + ((size_t)ra.i0 * w + ca.i0) * C8), q01 = __ldg(src + ((size_t)ra.i0 * w + ca.i1) * C8);
      const uint4 q10 = __ldg(src + ((size_t)ra.i1 * w + ca.i0) * C8), q11 = __ldg(src + ((size_t)ra.i1 * w + ca.i1) * C8);
      bf16x8_to_f32(q00, v00);
      bf16x8_to_f32(q01, v01);
      bf16x8_to_f32(q10, v10);
      bf16x8_to_f32(q11, v11);
      if (ua && va) emit(v00, v01, v10, v11, ra, ca, ya, xa);
      if (ua && vb) emit(v00, v01, v10, v11, ra, cb, ya, xb);
      if (ub && va) emit(v00, v01, v10, v11, rb, ca, yb, xa);
      if (ub && vb) emit(v00, v01, v10, v11, rb, cb, yb, xb);
    } else {
      if (ua && va) generic(ra, ca, ya, xa);
      if (ua && vb) generic(ra, cb, ya, xb);
      if (ub && va) generic(rb, ca, yb, xa);
      if (ub && vb) generic(rb, cb, yb, xb);
    }
  }
}

// NHWC bf16 -> NCHW fp32 (parity hook, not on the hot path)
__global__ void k_nhwc_to_nchw_f32(const __nv_bfloat16* in, int B, int H, int W, int C, float* out) {
  const size_t total = (size_t)B * C * H * W;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    size_t r = i / W;
    const int y = (int)(r % H); r /= H;
    const int c = (int)(r % C);
    const int n = (int)(r / C);
    out[i] = __bfloat162float(in[(((size_t)n * H + y) * W + x) * C + c]);
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
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

// 4-D map over an NHWC bf16 activation: dims (C, W, H, B), box (64, tile_w, tile_h, 1), 128-byte swizzle.
static int make_act_map(CUtensorMap* m, const void* ptr, int B, int H, int W, int C, int box_w, int box_h,
                        bool swizzle = true) {
  EncodeTiledFn fn = encode_fn();
  RDP_REQUIRE(fn, RDP_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {BLOCK_K, (cuuint32_t)box_w, (cuuint32_t)box_h, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RDP_REQUIRE(r == CUDA_SUCCESS, RDP_ERR_CUDA, "cuTensorMapEncodeTiled(activation %dx%dx%dx%d) failed: %d", B, H, W, C, (int)r);
  return RDP_OK;
}

// 2-D map over packed weights [cout][K]: dims (K, cout), box (64, bn).
static int make_weight_map(CUtensorMap* m, const void* ptr, int K, int cout, int bn) {
  EncodeTiledFn fn = encode_fn();
  RDP_REQUIRE(fn, RDP_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)cout};
  cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  cuuint32_t box[2] = {BLOCK_K, (cuuint32_t)bn};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RDP_REQUIRE(r == CUDA_SUCCESS, RDP_ERR_CUDA, "cuTensorMapEncodeTiled(weights %dx%d) failed: %d", cout, K, (int)r);
  return RDP_OK;
}

static void pick_tile(int H, int W, int& tw, int& th) {
  const int cand[5][2] = {{16, 8}, {8, 16}, {32, 4}, {64, 2}, {128, 1}};
  long best = -1;
  for (auto& c : cand) {
    const long cover = (long)((W + c[0] - 1) / c[0]) * ((H + c[1] - 1) / c[1]);
    if (best < 0 || cover < best) {
      best = cover;
      tw = c[0];
      th = c[1];
    }
  }
}

struct Level { int H, W; };

static int log2_exact(int v) {  // shift for a power of two, -1 otherwise
  int s = 0;
  while ((1 << s) < v) ++s;
  return (1 << s) == v ? s : -1;
}

}  // namespace unet
}  // namespace rdp

using namespace rdp;
using namespace rdp::unet;

// Encoded tensor maps of one (workspace, shape) pair, in the order forward_impl asks for them.
struct MapSet {
  const void* ws = nullptr;
  int B = 0, H = 0, W = 0, keep = 0, n = 0;
  unsigned long long stamp = 0;
  CUtensorMap maps[40];
};
constexpr int MAP_CACHE_ENTRIES = 8;

struct RdpUnet {
  RdpUnetWeights w;
  CUtensorMap wmap[18];
  CUtensorMap wmap_half[18];  // box of bn/2 rows: each CTA of a pair loads its half of a weight slice (bn <= 128 only)
  int bn[18];
  int cin_pad[18];
  int num_sms;
  // The handle's parameters never change after creation; this cache of activation tensor maps (about 40
  // cuTensorMapEncodeTiled calls per forward otherwise) is the only mutable part and has its own lock.
  mutable std::mutex map_mu;
  mutable unsigned long long map_clock = 0;
  mutable MapSet map_cache[MAP_CACHE_ENTRIES];
};

namespace {

// Activation buffers of one forward call, all NHWC bf16, carved from the caller's workspace.
struct Buffers {
  Level lv[5];
  __nv_bfloat16* col;      // im2col of the input, 64 ch @ level 0
  __nv_bfloat16* act[18];  // output of conv i
  __nv_bfloat16* pool[4];  // pooled x1..x4
  __nv_bfloat16* up[4];    // upsampled + padded low-resolution branch for up1..up4
  size_t bytes;
};

const int kConvLevel[18] = {0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 3, 3, 2, 2, 1, 1, 0, 0};

// keep_all: every buffer has its own storage (rdp_unet_read_activation can then read any layer afterwards).
// Otherwise buffers whose lifetimes do not overlap share storage.  Time runs in half steps: conv i launches at 2i, the
// stand-alone up-sampling kernel of Up block j at 2(10+2j)-1, the im2col fallback at -1; a buffer lives from the launch
// that writes it to the last launch that reads it, and two buffers may share bytes only if those closed intervals are
// disjoint.  Largest first, lowest free offset: x1 (live from inc.3 to up4.0) ends up at the bottom, and the peak is
// x1 + up4's two inputs + its output, about 30 % of the keep_all size.
Buffers carve_unet(const RdpUnet* net, void* ws, int B, int H, int W, bool keep_all) {
  Buffers b;
  b.lv[0] = {H, W};
  for (int i = 1; i < 5; ++i) b.lv[i] = {b.lv[i - 1].H / 2, b.lv[i - 1].W / 2};
  auto px = [&](int l) { return (size_t)B * b.lv[l].H * b.lv[l].W; };
  struct Span { size_t bytes, off; int t0, t1; __nv_bfloat16** slot; };
  Span sp[27];
  int n = 0;
  auto add = [&](__nv_bfloat16** slot, size_t elems, int t0, int t1) { sp[n++] = {align_up(elems * 2, 256), 0, t0, t1, slot}; };
  add(&b.col, px(0) * 64, -1, 0);
  static const int last_read[18] = {2, 32, 6, 28, 10, 24, 14, 20, 18, 20, 22, 24, 26, 28, 30, 32, 34, 35};
  for (int i = 0; i < 18; ++i) add(&b.act[i], px(kConvLevel[i]) * net->w.conv[i].cout, 2 * i, last_read[i]);
  for (int i = 0; i < 4; ++i) add(&b.pool[i], px(i + 1) * net->w.conv[2 * i + 1].cout, 2 * (2 * i + 1), 2 * (2 * i + 2));
  // up[j] feeds conv 10+2j: channels = cout of the layer below (conv 9+2j), spatial = skip level 3-j
  for (int j = 0; j < 4; ++j) add(&b.up[j], px(3 - j) * net->w.conv[9 + 2 * j].cout, 2 * (10 + 2 * j) - 1, 2 * (10 + 2 * j));
  size_t total = 0;
  if (keep_all) {
    for (int i = 0; i < n; ++i) {
      sp[i].off = total;
      total += sp[i].bytes;
    }
  } else {
    int order[27];
    for (int i = 0; i < n; ++i) order[i] = i;
    std::stable_sort(order, order + n, [&](int a, int c) { return sp[a].bytes > sp[c].bytes; });
    bool placed[27] = {};
    for (int oi = 0; oi < n; ++oi) {
      Span& cur = sp[order[oi]];
      size_t off = 0;
      for (bool moved = true; moved;) {  // lowest offset not overlapping a placed buffer that is live at the same time
        moved = false;
        for (int k = 0; k < n; ++k) {
          const Span& o = sp[k];
          if (!placed[k] || o.t1 < cur.t0 || cur.t1 < o.t0) continue;
          if (off < o.off + o.bytes && o.off < off + cur.bytes) {
            off = o.off + o.bytes;
            moved = true;
          }
        }
      }
      cur.off = off;
      placed[order[oi]] = true;
      total = std::max(total, off + cur.bytes);
    }
  }
  for (int i = 0; i < n; ++i) *sp[i].slot = ws ? reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(ws) + sp[i].off) : nullptr;
  b.bytes = total;
  return b;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of the (kernel, device) pair: a process that drives several
// GPUs (threads instead of one process per GPU) needs it set on each of them, and two first calls may race.
int ensure_dyn_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static std::set<std::pair<const void*, int>> done;
  int dev = 0;
  RDP_CUDA_TRY(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> guard(mu);
  if (done.count({kernel, dev})) return RDP_OK;
  RDP_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  done.insert({kernel, dev});
  return RDP_OK;
}
#define RDP_TRY_RC(expr)            \
  do {                              \
    const int _rc = (expr);         \
    if (_rc != RDP_OK) return _rc;  \
  } while (0)

template <int BN>
int launch_conv_bn(const CUtensorMap& s0, const CUtensorMap& s1, const CUtensorMap& wm, const ConvParams& p, int num_sms,
                   cudaStream_t stream) {
  RDP_TRY_RC(ensure_dyn_smem(reinterpret_cast<const void*>(&k_conv3x3<BN>), (int)ConvCfg<BN>::SMEM_BYTES));
  const int grid = p.total_work < num_sms ? p.total_work : num_sms;
  k_conv3x3<BN><<<grid, CONV_THREADS, ConvCfg<BN>::SMEM_BYTES, stream>>>(s0, s1, wm, p);
  return RDP_OK;
}

template <int C, bool U8>
int launch_conv_first_t(const CUtensorMap& wm, const CUtensorMap& om, const ConvParams& p, const void* in, int num_sms,
                        cudaStream_t stream) {
  RDP_TRY_RC(ensure_dyn_smem(reinterpret_cast<const void*>(&k_conv_first<C, U8>), FIRST_SMEM));
  const int grid = p.total_work < num_sms ? p.total_work : num_sms;
  k_conv_first<C, U8><<<grid, FIRST_THREADS, FIRST_SMEM, stream>>>(wm, om, p, in);
  return RDP_OK;
}

// returns 1 when a specialised kernel exists for this channel count (else the im2col path is used)
int launch_conv_first(int C, bool u8, const CUtensorMap& wm, const CUtensorMap& om, const ConvParams& p, const void* in,
                      int num_sms, cudaStream_t stream, int* rc) {
  switch (C * 2 + (u8 ? 1 : 0)) {
    case 3: *rc = launch_conv_first_t<1, true>(wm, om, p, in, num_sms, stream); return 1;
    case 2: *rc = launch_conv_first_t<1, false>(wm, om, p, in, num_sms, stream); return 1;
    case 7: *rc = launch_conv_first_t<3, true>(wm, om, p, in, num_sms, stream); return 1;
    case 6: *rc = launch_conv_first_t<3, false>(wm, om, p, in, num_sms, stream); return 1;
    case 9: *rc = launch_conv_first_t<4, true>(wm, om, p, in, num_sms, stream); return 1;
    case 8: *rc = launch_conv_first_t<4, false>(wm, om, p, in, num_sms, stream); return 1;
    default: return 0;
  }
}

template <int BN, bool RESIDENT, bool INTERP, int TILES, bool RAW = false>
int launch_conv_halo_t(const CUtensorMap& s0, const CUtensorMap& s1, const CUtensorMap& wm, const ConvParams& p,
                       const HaloCfg& cfg, int num_sms, cudaStream_t stream) {
  RDP_TRY_RC(ensure_dyn_smem(reinterpret_cast<const void*>(&k_conv3x3_halo<BN, RESIDENT, INTERP, TILES, RAW>), HALO_SMEM_TOTAL));
  const int grid = p.total_work < num_sms ? p.total_work : num_sms;
  k_conv3x3_halo<BN, RESIDENT, INTERP, TILES, RAW><<<grid, conv_threads(INTERP, BN, TILES), HALO_SMEM_TOTAL, stream>>>(s0, s1, wm, p, cfg);
  return RDP_OK;
}

template <int BN, bool RESIDENT, bool INTERP>
int launch_conv_pair_t(const CUtensorMap& s0, const CUtensorMap& s1, const CUtensorMap& wm_half, const ConvParams& p,
                       const HaloCfg& cfg, int num_sms, cudaStream_t stream) {
  RDP_TRY_RC(ensure_dyn_smem(reinterpret_cast<const void*>(&k_conv3x3_pair<BN, RESIDENT, INTERP>), HALO_SMEM_TOTAL));
  const int pair_work = (p.B * p.tiles_y * p.tiles_x + 1) / 2;
  const int pairs = std::min(pair_work, num_sms / 2);
  k_conv3x3_pair<BN, RESIDENT, INTERP><<<2 * pairs, conv_threads(INTERP, BN, 1), HALO_SMEM_TOTAL, stream>>>(s0, s1, wm_half, p, cfg);
  return RDP_OK;
}

// Pairs pay where the halved weight slices stay resident in shared memory; with streamed slices every tap is a
// two-CTA hand-shake and the pair runs several times slower than two independent CTAs (measured on up3.0, up2.3).
inline bool pair_weights_resident(int bn, int ncb, int cout, bool head, bool low) {
  const int fixed = 1024 + HALO_BAR_BYTES + aux_bytes(cout, head) + (low ? patch_stride(1) : 0);
  return 9 * ncb * (bn / 2) * BLOCK_K * 2 + 3 * halo_stride(1) <= HALO_SMEM_TOTAL - fixed;
}
// streamed pair: two stages of nine half slices plus at least two halo stages must fit
inline bool pair_weights_streamable(int bn, int cout, bool head, bool low) {
  const int fixed = 1024 + HALO_BAR_BYTES + aux_bytes(cout, head) + (low ? patch_stride(1) : 0);
  return 2 * 9 * (bn / 2) * BLOCK_K * 2 + 2 * halo_stride(1) <= HALO_SMEM_TOTAL - fixed;
}

// CTA pairs (k_conv3x3_pair): 8x16 tiles, each CTA holds half of every weight slice.
template <int BN>
int launch_conv_pair(const CUtensorMap& s0, const CUtensorMap& s1, const CUtensorMap& wm_half, const ConvParams& p, int num_sms,
                     cudaStream_t stream) {
  const int ncb = p.kb0 + p.kb1, b_half = (BN / 2) * BLOCK_K * 2;
  const int fixed = 1024 + HALO_BAR_BYTES + aux_bytes(p.cout, p.head_w != nullptr) + (p.low ? patch_stride(1) : 0);
  int budget = HALO_SMEM_TOTAL - fixed;
  HaloCfg cfg{};
  cfg.sl = 1;
  const int w_bytes = 9 * ncb * b_half;
  if (w_bytes + 3 * halo_stride(1) <= budget) {  // half slices resident: every 64-wide layer, down1.0 and down1.3
    cfg.sb = 0;
    cfg.sa = std::min(HALO_MAX_STAGES, (budget - w_bytes) / halo_stride(1));
    if (p.low && budget - w_bytes - cfg.sa * halo_stride(1) >= patch_stride(1)) cfg.sl = 2;
    return p.low ? launch_conv_pair_t<BN, true, true>(s0, s1, wm_half, p, cfg, num_sms, stream)
                 : launch_conv_pair_t<BN, true, false>(s0, s1, wm_half, p, cfg, num_sms, stream);
  }
  // streamed: two stages of nine half slices each (a whole channel block per hand-shake), the rest goes to halo stages
  cfg.sb = 2;
  budget -= cfg.sb * 9 * b_half;
  if (p.low && budget - 2 * halo_stride(1) >= patch_stride(1)) {
    cfg.sl = 2;
    budget -= patch_stride(1);
  }
  cfg.sa = std::min(HALO_MAX_STAGES, budget / halo_stride(1));
  RDP_REQUIRE(cfg.sa >= 2, RDP_ERR_UNSUPPORTED, "launch_conv_pair: no room for two halo stages (bn %d, %d channel blocks)", BN, ncb);
  return p.low ? launch_conv_pair_t<BN, false, true>(s0, s1, wm_half, p, cfg, num_sms, stream)
               : launch_conv_pair_t<BN, false, false>(s0, s1, wm_half, p, cfg, num_sms, stream);
}

// Super-tiles (two 8x16 tiles per work item) pay off where weights stream from L2 and two accumulator pairs fit TMEM.
inline int halo_tiles_for(int bn, int n_tiles, int ncb, bool interp) {
  static const int env = dev_env("RDP_HALO_TILES") ? atoi(dev_env("RDP_HALO_TILES")) : 0;
  if (env == 1) return 1;
  const int budget = HALO_SMEM_TOTAL - 1024 - HALO_BAR_BYTES - aux_bytes(bn * n_tiles, false) - (interp ? patch_stride(1) : 0);
  const bool resident = n_tiles == 1 && 9 * ncb * bn * BLOCK_K * 2 + 3 * halo_stride(1) <= budget;
  return (!resident && bn == 128) ? 2 : 1;
}

template <int BN>
int launch_conv_halo(const CUtensorMap& s0, const CUtensorMap& s1, const CUtensorMap& wm, const ConvParams& p, int tiles,
                     int num_sms, cudaStream_t stream) {
  const int ncb = p.kb0 + p.kb1, b_block = BN * BLOCK_K * 2;
  const int HALO_STRIDE = halo_stride(tiles);
  // everything but the rings: alignment slack, barriers, epilogue constants, one source-patch stage when fused
  const int fixed = 1024 + HALO_BAR_BYTES + aux_bytes(p.cout, p.head_w != nullptr) + (p.low ? patch_stride(tiles) : 0);
  int budget = HALO_SMEM_TOTAL - fixed;
  static const int env_sa = dev_env("RDP_HALO_SA") ? atoi(dev_env("RDP_HALO_SA")) : 0;
  HaloCfg cfg{};
  cfg.sl = 1;
  const int w_bytes = 9 * ncb * b_block;
  if (tiles == 1 && p.n_tiles == 1 && w_bytes + 3 * HALO_STRIDE <= budget) {
    cfg.sb = 0;
    cfg.sa = std::min(HALO_MAX_STAGES, (budget - w_bytes) / HALO_STRIDE);
    if (p.low && budget - w_bytes - cfg.sa * HALO_STRIDE >= patch_stride(tiles)) cfg.sl = 2;
    return p.low ? launch_conv_halo_t<BN, true, true, 1>(s0, s1, wm, p, cfg, num_sms, stream)
                 : launch_conv_halo_t<BN, true, false, 1>(s0, s1, wm, p, cfg, num_sms, stream);
  }
  // streamed weights: a halo block lasts 9 taps, so few A stages suffice; the weight ring must cover the
  // bandwidth-latency product of one 64-wide slice per tap
  cfg.sa = env_sa ? env_sa : ((BN == 256 || tiles == 2) ? 2 : 3);
  if (p.low) {  // streamed layers have room for a second patch stage
    cfg.sl = 2;
    budget -= patch_stride(tiles);
  }
  cfg.sb = std::min(HALO_MAX_STAGES, (budget - cfg.sa * HALO_STRIDE) / b_block);
  if constexpr (BN == 128) {
    if (tiles == 2)
      return p.low ? launch_conv_halo_t<BN, false, true, 2>(s0, s1, wm, p, cfg, num_sms, stream)
                   : launch_conv_halo_t<BN, false, false, 2>(s0, s1, wm, p, cfg, num_sms, stream);
  }
  return p.low ? launch_conv_halo_t<BN, false, true, 1>(s0, s1, wm, p, cfg, num_sms, stream)
               : launch_conv_halo_t<BN, false, false, 1>(s0, s1, wm, p, cfg, num_sms, stream);
}

}  // namespace

namespace {
// The halo kernel without ReLU (training step): same stage planning as launch_conv_halo, no fused up-sampling.
template <int BN>
int launch_conv_halo_raw(const CUtensorMap& s0, const CUtensorMap& s1, const CUtensorMap& wm, const ConvParams& p, int tiles,
                         int num_sms, cudaStream_t stream) {
  const int ncb = p.kb0 + p.kb1, b_block = BN * BLOCK_K * 2;
  const int HALO_STRIDE = halo_stride(tiles);
  const int budget = HALO_SMEM_TOTAL - (1024 + HALO_BAR_BYTES + aux_bytes(p.cout, false));
  HaloCfg cfg{};
  cfg.sl = 1;
  const int w_bytes = 9 * ncb * b_block;
  if (tiles == 1 && p.n_tiles == 1 && w_bytes + 3 * HALO_STRIDE <= budget) {
    cfg.sb = 0;
    cfg.sa = std::min(HALO_MAX_STAGES, (budget - w_bytes) / HALO_STRIDE);
    return launch_conv_halo_t<BN, true, false, 1, true>(s0, s1, wm, p, cfg, num_sms, stream);
  }
  cfg.sa = (BN == 256 || tiles == 2) ? 2 : 3;
  cfg.sb = std::min(HALO_MAX_STAGES, (budget - cfg.sa * HALO_STRIDE) / b_block);
  if constexpr (BN == 128) {
    if (tiles == 2) return launch_conv_halo_t<BN, false, false, 2, true>(s0, s1, wm, p, cfg, num_sms, stream);
  }
  return launch_conv_halo_t<BN, false, false, 1, true>(s0, s1, wm, p, cfg, num_sms, stream);
}

template <int BN>
int launch_raw_bn(const CUtensorMap& s0, const CUtensorMap& s1, const CUtensorMap& wm, const ConvParams& p, int num_sms, cudaStream_t stream) {
  RDP_TRY_RC(ensure_dyn_smem(reinterpret_cast<const void*>(&k_conv3x3<BN, true>), (int)ConvCfg<BN>::SMEM_BYTES));
  const int grid = p.total_work < num_sms ? p.total_work : num_sms;
  k_conv3x3<BN, true><<<grid, CONV_THREADS, ConvCfg<BN>::SMEM_BYTES, stream>>>(s0, s1, wm, p);
  return RDP_OK;
}
}  // namespace

// Used by the training step (csrc/train.cu): out[B,H,W,cout] = conv(cat(src0[c0], src1[c1]), w) with `taps` = 9 (3x3, pad 1)
// or 1 (1x1 over an im2col'ed input), bf16 in and out, no BatchNorm, no ReLU.  `w` = bf16 [cout][taps][c0+c1];
// `zero_shift` = cout zeros (fp32).  Channel counts are multiples of 64.
int rdp::unet::raw_conv(const void* src0, int c0, const void* src1, int c1, const void* w, int cout, int taps, int B, int H, int W,
                        const float* zero_shift, void* out, int num_sms, void* stream_) {
  RDP_REQUIRE(c0 % 64 == 0 && c1 % 64 == 0 && cout % 64 == 0 && (cout <= 256 || cout % 256 == 0), RDP_ERR_UNSUPPORTED,
              "raw_conv: channel counts %d+%d -> %d unsupported", c0, c1, cout);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ConvParams p{};
  p.B = B; p.H = H; p.W = W;
  const int bn = cout >= 256 ? 256 : cout;
  const bool halo = taps == 9;   // 3x3: the halo-reuse kernel; the im2col'ed first layer (taps == 1) goes per tap
  const int tiles = halo ? halo_tiles_for(bn, cout / bn, (c0 + c1) / 64, false) : 1;
  if (halo) { p.tile_w = 8 * tiles; p.tile_h = 16; }
  else pick_tile(H, W, p.tile_w, p.tile_h);
  p.tiles_x = (W + p.tile_w - 1) / p.tile_w;
  p.tiles_y = (H + p.tile_h - 1) / p.tile_h;
  p.cout = cout;
  p.n_tiles = cout / bn;
  p.kb0 = c0 / 64; p.kb1 = c1 / 64;
  p.taps = taps;
  p.cin_pad = c0 + c1;
  p.total_work = B * p.tiles_y * p.tiles_x * p.n_tiles;
  p.shift = zero_shift;
  p.out = static_cast<__nv_bfloat16*>(out);
  CUtensorMap m0, m1, wm;
  const int box_w = halo ? halo_w(tiles) : p.tile_w, box_h = halo ? HALO_H : p.tile_h;
  RDP_TRY_RC(make_act_map(&m0, src0, B, H, W, c0, box_w, box_h));
  RDP_TRY_RC(make_act_map(&m1, c1 ? src1 : src0, B, H, W, c1 ? c1 : c0, box_w, box_h));
  RDP_TRY_RC(make_weight_map(&wm, w, taps * (c0 + c1), cout, bn));
  int rc;
  if (halo) {
    switch (bn) {
      case 64: rc = launch_conv_halo_raw<64>(m0, m1, wm, p, tiles, num_sms, stream); break;
      case 128: rc = launch_conv_halo_raw<128>(m0, m1, wm, p, tiles, num_sms, stream); break;
      default: rc = launch_conv_halo_raw<256>(m0, m1, wm, p, tiles, num_sms, stream); break;
    }
    RDP_CUDA_TRY(cudaGetLastError());
    return rc;
  }
  switch (bn) {
    case 64: rc = launch_raw_bn<64>(m0, m1, wm, p, num_sms, stream); break;
    case 128: rc = launch_raw_bn<128>(m0, m1, wm, p, num_sms, stream); break;
    default: rc = launch_raw_bn<256>(m0, m1, wm, p, num_sms, stream); break;
  }
  RDP_CUDA_TRY(cudaGetLastError());
  return rc;
}

int rdp::unet::im2col_first(const void* in, int in_fmt, int B, int H, int W, int C, void* out, void* stream_) {
  launch_im2col_first(in, in_fmt, B, H, W, C, static_cast<__nv_bfloat16*>(out), static_cast<cudaStream_t>(stream_));
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

int rdp::unet::maxpool2(const void* in, int B, int H, int W, int C, void* out, void* stream_) {
  const int C8 = C / 8;
  k_maxpool2<<<dim3(((W / 2) * C8 + 255) / 256, H / 2, B), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(in), H, W, C8, log2_exact(C8), static_cast<__nv_bfloat16*>(out));
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

int rdp::unet::upsample2x(const void* in, int B, int h, int w, int C, int Hs, int Ws, void* out, void* stream_) {
  const int C8 = C / 8, dy = Hs - 2 * h, dx = Ws - 2 * w, sh = log2_exact(C8);
  if (sh >= 0 && dy == 0 && dx == 0 && h > 1 && w > 1) {
    const float ry = (float)(h - 1) / (float)(2 * h - 1), rx = (float)(w - 1) / (float)(2 * w - 1);
    k_upsample2x_quad<<<dim3(((w + 1) * C8 + 255) / 256, (h + 1 + UQ_ROWS - 1) / UQ_ROWS, B), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        static_cast<const __nv_bfloat16*>(in), h, w, C8, sh, ry, rx, static_cast<__nv_bfloat16*>(out));
  } else {
    k_upsample2x<<<dim3((Ws * C8 + 255) / 256, Hs, B), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        static_cast<const __nv_bfloat16*>(in), h, w, C8, sh, Hs, Ws, dy / 2, dx / 2, static_cast<__nv_bfloat16*>(out));
  }
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

extern "C" int rdp_unet_create(const RdpUnetWeights* w, RdpUnet** out) {
  RDP_REQUIRE(w && out, RDP_ERR_BAD_ARGUMENT, "rdp_unet_create: null argument");
  RDP_REQUIRE(w->n_channels >= 1 && 9 * w->n_channels <= 64, RDP_ERR_UNSUPPORTED, "rdp_unet_create: n_channels %d unsupported (9*C must be <= 64)", w->n_channels);
  RDP_REQUIRE(w->n_classes >= 1 && w->n_classes <= 8, RDP_ERR_UNSUPPORTED, "rdp_unet_create: n_classes %d unsupported (1..8)", w->n_classes);
  RDP_REQUIRE(w->outc_weight && w->outc_bias, RDP_ERR_BAD_ARGUMENT, "rdp_unet_create: missing OutConv parameters");
  static const int cin[18] = {0, 64, 64, 128, 128, 256, 256, 512, 512, 512, 1024, 512, 512, 256, 256, 128, 128, 64};
  static const int cout[18] = {64, 64, 128, 128, 256, 256, 512, 512, 512, 512, 512, 256, 256, 128, 128, 64, 64, 64};
  RdpUnet* net = new (std::nothrow) RdpUnet;
  RDP_REQUIRE(net, RDP_ERR_CUDA, "rdp_unet_create: out of host memory");
  net->w = *w;
  int dev = 0;
  cudaDeviceProp prop;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
    delete net;
    set_error("rdp_unet_create: no CUDA device");
    return RDP_ERR_CUDA;
  }
  if (prop.major != 10) {
    delete net;
    set_error("rdp_unet_create: this library is built for sm_100a only (device is sm_%d%d)", prop.major, prop.minor);
    return RDP_ERR_UNSUPPORTED;
  }
  net->num_sms = prop.multiProcessorCount;
  for (int i = 0; i < 18; ++i) {
    const RdpConvLayer& L = w->conv[i];
    const int want_cin = i == 0 ? w->n_channels : cin[i];
    if (!L.weight || !L.shift || L.cin != want_cin || L.cout != cout[i]) {
      delete net;
      set_error("rdp_unet_create: layer %d expects cin=%d cout=%d with non-null weight/shift (got %d, %d)", i, want_cin, cout[i], L.cin, L.cout);
      return RDP_ERR_BAD_ARGUMENT;
    }
    net->cin_pad[i] = i == 0 ? 64 : cin[i];
    net->bn[i] = cout[i] >= 256 ? 256 : cout[i];
    const int K = i == 0 ? 64 : 9 * cin[i];
    int rc = make_weight_map(&net->wmap[i], L.weight, K, cout[i], net->bn[i]);
    if (rc == RDP_OK && i > 0 && net->bn[i] <= 128) rc = make_weight_map(&net->wmap_half[i], L.weight, K, cout[i], net->bn[i] / 2);
    if (rc != RDP_OK) {
      delete net;
      return rc;
    }
  }
  *out = net;
  return RDP_OK;
}

extern "C" void rdp_unet_destroy(RdpUnet* net) { delete net; }

extern "C" int rdp_unet_launch_count(const RdpUnet*) { return 1 + 18 + 4 + 4; }

extern "C" int rdp_unet_workspace_bytes_ex(const RdpUnet* net, int in_fmt_flags, int B, int H, int W, size_t* bytes) {
  RDP_REQUIRE(net && bytes, RDP_ERR_BAD_ARGUMENT, "rdp_unet_workspace_bytes: null argument");
  RDP_REQUIRE(B > 0 && H >= 16 && W >= 16, RDP_ERR_BAD_ARGUMENT, "rdp_unet_workspace_bytes: bad shape %dx%dx%d (H, W >= 16)", B, H, W);
  *bytes = carve_unet(net, nullptr, B, H, W, (in_fmt_flags & RDP_IN_FLAG_KEEP_ACTIVATIONS) != 0).bytes;
  return RDP_OK;
}

extern "C" int rdp_unet_workspace_bytes(const RdpUnet* net, int B, int H, int W, size_t* bytes) {
  return rdp_unet_workspace_bytes_ex(net, 0, B, H, W, bytes);
}

static int forward_impl(const RdpUnet* net, const void* input, int in_fmt, int B, int H, int W, float* logits,
                        uint8_t* mask, void* workspace, size_t workspace_bytes, void* stream_, cudaEvent_t* marks) {
  RDP_REQUIRE(net && input && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_unet_forward: null argument");
  RDP_REQUIRE(logits || mask, RDP_ERR_BAD_ARGUMENT, "rdp_unet_forward: neither logits nor mask requested");
  const bool keep_all = (in_fmt & RDP_IN_FLAG_KEEP_ACTIVATIONS) != 0;
  in_fmt &= ~RDP_IN_FLAG_KEEP_ACTIVATIONS;
  RDP_REQUIRE(in_fmt == RDP_IN_RGB_U8_NHWC || in_fmt == RDP_IN_F32_NCHW, RDP_ERR_BAD_ARGUMENT, "rdp_unet_forward: unknown input format %d", in_fmt);
  RDP_REQUIRE(B > 0 && H >= 16 && W >= 16, RDP_ERR_BAD_ARGUMENT, "rdp_unet_forward: bad shape %dx%dx%d (H, W >= 16)", B, H, W);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const Buffers buf = carve_unet(net, workspace, B, H, W, keep_all);
  RDP_REQUIRE(workspace_bytes >= buf.bytes, RDP_ERR_WORKSPACE_TOO_SMALL, "rdp_unet_forward: workspace %zu < required %zu bytes", workspace_bytes, buf.bytes);
  RDP_REQUIRE(((uintptr_t)workspace & 255) == 0, RDP_ERR_BAD_ARGUMENT, "rdp_unet_forward: workspace must be 256-byte aligned");

  // Activation tensor maps depend on the workspace address and the shape only: encode them once per (workspace, shape)
  // and replay them, in call order, on later forwards (the batch-1 server loop calls with the same pair every frame).
  MapSet maps;
  bool maps_cached = false;
  {
    std::lock_guard<std::mutex> guard(net->map_mu);
    for (MapSet& e : net->map_cache)
      if (e.n > 0 && e.ws == workspace && e.B == B && e.H == H && e.W == W && e.keep == (int)keep_all) {
        e.stamp = ++net->map_clock;
        maps = e;
        maps_cached = true;
        break;
      }
  }
  int map_cursor = 0;
  // returns the next map of the sequence: replayed when cached, otherwise built by `make` and recorded
  auto next_map = [&](const CUtensorMap** out, auto make) -> int {
    if (map_cursor >= (int)(sizeof(maps.maps) / sizeof(maps.maps[0]))) {
      set_error("rdp_unet_forward: tensor-map table overflow");
      return RDP_ERR_CUDA;
    }
    CUtensorMap* m = &maps.maps[map_cursor++];
    *out = m;
    return maps_cached ? RDP_OK : make(m);
  };

  static const bool debug_waits = dev_env("RDP_DEBUG_WAITS") != nullptr;  // development aid, see ConvParams::debug_waits
  static const bool debug_nohead = dev_env("RDP_DEBUG_NOHEAD") != nullptr;    // timing experiments only
  static const bool debug_nostore = dev_env("RDP_DEBUG_NOSTORE") != nullptr;  // timing experiments only: results are garbage
  static const bool no_fuse_pool = dev_env("RDP_NO_FUSE_POOL") != nullptr;  // development switch: separate pooling kernel
  static const bool fuse_up4 = dev_env("RDP_FUSE_UP4") != nullptr;  // development switch: fuse up4's upsample as well
  static const bool no_fuse_up = dev_env("RDP_NO_FUSE_UP") != nullptr;  // development switch: separate upsample kernel
  static const bool force_v1 = dev_env("RDP_CONV_V1") != nullptr;  // development switch: per-tap loads everywhere
  // per-layer kernel choice, 'h' = halo reuse, '1' = per-tap loads (RDP_CONV_SEL overrides, 18 characters)
  static const char* sel_env = dev_env("RDP_CONV_SEL");
  // Streamed weights on a pair (two stages of nine half slices, one hand-shake per channel block) still lose to two
  // independent CTAs with super-tiles: up3.0 3.14 vs 2.67 ms, up2.3 0.69 vs 0.63 (profiles/r02_pair_streamed_layer_times.txt);
  // the path stays behind a development switch.
  static const bool pair_streamed_ok = dev_env("RDP_PAIR_STREAMED") != nullptr;
  // 'p' = halo reuse on CTA pairs (cta_group::2): the 64- and 128-wide layers, whose MMAs are bounded by operand reads
  // (measured, batch 64 at 640x480, pair vs two independent CTAs: down1.3 0.94 vs 1.19 ms, up3.3 0.59 vs 0.69, up4.0 2.34
  // vs 2.80, up4.3 1.40 vs 1.49; inc.3 and down1.0, one 64-channel block per tile, lose to the per-tile hand-shake)
  const char* sel = (sel_env && strlen(sel_env) == 18) ? sel_env : "1hhphhhhhhhhhhhppp";
  int n_marks = 0;
  auto mark = [&]() {
    if (marks) cudaEventRecord(marks[n_marks++], stream);
  };
  mark();

  // first layer: fused gather + tensor-core kernel when the channel count has a specialisation (1, 3, 4);
  // otherwise an explicit im2col buffer feeds the generic kernel.  Both occupy two launch slots.
  bool first_fused = false;
  if (!force_v1) {
    ConvParams p{};
    p.B = B; p.H = H; p.W = W;
    pick_tile(H, W, p.tile_w, p.tile_h);
    p.tiles_x = (W + p.tile_w - 1) / p.tile_w;
    p.tiles_y = (H + p.tile_h - 1) / p.tile_h;
    p.cout = 64; p.n_tiles = 1; p.kb0 = 1; p.taps = 1; p.cin_pad = 64;
    p.total_work = B * p.tiles_y * p.tiles_x;
    p.shift = net->w.conv[0].shift;
    p.debug_waits = debug_waits;
    p.out = debug_nostore ? nullptr : buf.act[0];
    int rc = RDP_OK;
    mark();  // (slot of the im2col launch)
    const CUtensorMap* omap = nullptr;
    rc = next_map(&omap, [&](CUtensorMap* m) { return make_act_map(m, buf.act[0], B, H, W, 64, p.tile_w, p.tile_h); });
    if (rc != RDP_OK) return rc;
    if (launch_conv_first(net->w.n_channels, in_fmt == RDP_IN_RGB_U8_NHWC, net->wmap[0], *omap, p, input, net->num_sms, stream, &rc)) {
      if (rc != RDP_OK) return rc;
      first_fused = true;
      mark();
    } else if (marks) {
      --n_marks;
    }
  }
  if (!first_fused) {
    launch_im2col_first(input, in_fmt, B, H, W, net->w.n_channels, buf.col, stream);
    mark();
  }

  // one conv launch: sources (src0 with c0 channels [+ src1 with c1 channels]) at level `lvl` -> act[i]
  struct LowSrc { const __nv_bfloat16* ptr; int h, w, pad_y, pad_x; };
  auto conv = [&](int i, const __nv_bfloat16* src0, int c0, const __nv_bfloat16* src1, int c1, int taps,
                  const LowSrc* low = nullptr, __nv_bfloat16* pool_out = nullptr) -> int {
    const Level L = buf.lv[kConvLevel[i]];
    ConvParams p{};
    p.B = B; p.H = L.H; p.W = L.W;
    const int cout_i = net->w.conv[i].cout;
    const bool pair = taps == 9 && !force_v1 && sel[i] == 'p' && net->bn[i] <= 128 && cout_i == net->bn[i] && net->num_sms >= 2 &&
                      (pair_weights_resident(net->bn[i], (c0 + c1) / 64, cout_i, i == 17 && !debug_nohead, low != nullptr) ||
                       (pair_streamed_ok && pair_weights_streamable(net->bn[i], cout_i, false, low != nullptr)));
    const bool halo = taps == 9 && !force_v1 && (sel[i] == 'h' || sel[i] == 'p');
    const int tiles = (halo && !pair) ? halo_tiles_for(net->bn[i], cout_i / net->bn[i], (c0 + c1) / 64, low != nullptr) : 1;
    if (halo) { p.tile_w = 8 * tiles; p.tile_h = 16; }
    else pick_tile(L.H, L.W, p.tile_w, p.tile_h);
    p.tiles_x = (L.W + p.tile_w - 1) / p.tile_w;
    p.tiles_y = (L.H + p.tile_h - 1) / p.tile_h;
    p.cout = net->w.conv[i].cout;
    p.n_tiles = p.cout / net->bn[i];
    p.kb0 = c0 / 64; p.kb1 = c1 / 64;
    p.taps = taps;
    p.cin_pad = net->cin_pad[i];
    p.total_work = B * p.tiles_y * p.tiles_x * p.n_tiles;
    p.shift = net->w.conv[i].shift;
    p.debug_waits = debug_waits;
    p.pool_out = pool_out;
    if (debug_waits) { cudaStreamSynchronize(stream); printf("conv %d:\n", i); fflush(stdout); }
    p.out = debug_nostore ? nullptr : buf.act[i];
    if (i == 17 && !debug_nohead) {
      if (!keep_all) p.out = nullptr;  // nobody reads up4.3's activation: the head consumes it in registers
      p.head_w = net->w.outc_weight;
      p.head_b = net->w.outc_bias;
      p.n_classes = net->w.n_classes;
      p.logits = logits;
      p.mask = mask;
    }
    if (low) {
      p.low = low->ptr;
      p.low_h = low->h; p.low_w = low->w;
      p.pad_y = low->pad_y; p.pad_x = low->pad_x;
      p.ry = 2 * low->h > 1 ? (float)(low->h - 1) / (float)(2 * low->h - 1) : 0.f;
      p.rx = 2 * low->w > 1 ? (float)(low->w - 1) / (float)(2 * low->w - 1) : 0.f;
      static const bool interp_generic = dev_env("RDP_INTERP_GENERIC") != nullptr;  // development switch: per-pixel producer everywhere
      p.interp_fast = !interp_generic && low->pad_y == 0 && low->pad_x == 0 && L.H == 2 * low->h && L.W == 2 * low->w && low->h > 1 && low->w > 1;
    }
    const CUtensorMap *pm0 = nullptr, *pm1 = nullptr;
    const int box_w = halo ? halo_w(tiles) : p.tile_w, box_h = halo ? HALO_H : p.tile_h;
    int rc = next_map(&pm0, [&](CUtensorMap* m) { return make_act_map(m, src0, B, L.H, L.W, c0, box_w, box_h); });
    if (rc != RDP_OK) return rc;
    rc = next_map(&pm1, [&](CUtensorMap* m) {
      if (low) return make_act_map(m, low->ptr, B, low->h, low->w, c1, patch_w(tiles), PATCH_H, /*swizzle=*/false);
      return make_act_map(m, src1 ? src1 : src0, B, L.H, L.W, src1 ? c1 : c0, box_w, box_h);
    });
    if (rc != RDP_OK) return rc;
    const CUtensorMap &m0 = *pm0, &m1 = *pm1;
    if (pair) {
      rc = net->bn[i] == 64 ? launch_conv_pair<64>(m0, m1, net->wmap_half[i], p, net->num_sms, stream)
                            : launch_conv_pair<128>(m0, m1, net->wmap_half[i], p, net->num_sms, stream);
      mark();
      return rc;
    }
    if (halo) {
      switch (net->bn[i]) {
        case 64: rc = launch_conv_halo<64>(m0, m1, net->wmap[i], p, tiles, net->num_sms, stream); break;
        case 128: rc = launch_conv_halo<128>(m0, m1, net->wmap[i], p, tiles, net->num_sms, stream); break;
        default: rc = launch_conv_halo<256>(m0, m1, net->wmap[i], p, tiles, net->num_sms, stream); break;
      }
      mark();
      return rc;
    }
    switch (net->bn[i]) {
      case 64: rc = launch_conv_bn<64>(m0, m1, net->wmap[i], p, net->num_sms, stream); break;
      case 128: rc = launch_conv_bn<128>(m0, m1, net->wmap[i], p, net->num_sms, stream); break;
      default: rc = launch_conv_bn<256>(m0, m1, net->wmap[i], p, net->num_sms, stream); break;
    }
    mark();
    return rc;
  };
#define RDP_TRY(expr)          \
  do {                         \
    const int _rc = (expr);    \
    if (_rc != RDP_OK) return _rc; \
  } while (0)

  // encoder: inc, down1..down4 (segmentation_model.py:110-114)
  if (!first_fused) RDP_TRY(conv(0, buf.col, 64, nullptr, 0, 1));
  // MaxPool2d(2) of x1..x4 is a second output of the producing conv's epilogue (halo kernels); the stand-alone
  // pooling kernel remains for the per-tap fallback
  auto fused_pool = [&](int i) { return !force_v1 && !no_fuse_pool && (sel[i] == 'h' || sel[i] == 'p'); };
  RDP_TRY(conv(1, buf.act[0], 64, nullptr, 0, 9, nullptr, fused_pool(1) ? buf.pool[0] : nullptr));
  for (int d = 0; d < 4; ++d) {
    const Level L = buf.lv[d];
    const int C = net->w.conv[2 * d + 1].cout;
    const int C8 = C / 8;
    if (!fused_pool(2 * d + 1))
      k_maxpool2<<<dim3(((L.W / 2) * C8 + 255) / 256, L.H / 2, B), 256, 0, stream>>>(buf.act[2 * d + 1], L.H, L.W, C8, log2_exact(C8), buf.pool[d]);
    mark();
    RDP_TRY(conv(2 * d + 2, buf.pool[d], C, nullptr, 0, 9));
    RDP_TRY(conv(2 * d + 3, buf.act[2 * d + 2], net->w.conv[2 * d + 2].cout, nullptr, 0, 9, nullptr,
                 (d < 3 && fused_pool(2 * d + 3)) ? buf.pool[d + 1] : nullptr));
  }
  // decoder: up1..up4 (segmentation_model.py:115-118); skip channels first, then the upsampled branch
  for (int j = 0; j < 4; ++j) {
    const int low_i = 9 + 2 * j;                 // conv whose output is upsampled
    const int skip_i = 7 - 2 * j;                // x4, x3, x2, x1
    const Level Ls = buf.lv[3 - j], Ll = buf.lv[4 - j];
    const int Cl = net->w.conv[low_i].cout, Cs = net->w.conv[skip_i].cout;
    const int dy = Ls.H - 2 * Ll.H, dx = Ls.W - 2 * Ll.W;
    const int C8 = Cl / 8;
    // up4.0 (64 output channels): interpolation warps slow the MMA issue warp down as much as the copy costs - measured
    // equal, so it keeps the simpler stand-alone upsample unless asked
    const bool fuse_up = !force_v1 && !no_fuse_up && (sel[10 + 2 * j] == 'h' || sel[10 + 2 * j] == 'p') && (net->bn[10 + 2 * j] >= 128 || fuse_up4);
    if (fuse_up) {
      // the convolution's interpolation warps read the low-resolution tensor directly: no upsampled copy in HBM
      mark();
      const LowSrc low{buf.act[low_i], Ll.H, Ll.W, dy / 2, dx / 2};
      RDP_TRY(conv(10 + 2 * j, buf.act[skip_i], Cs, nullptr, Cl, 9, &low));
    } else {
      static const bool ups_v1 = dev_env("RDP_UPSAMPLE_V1") != nullptr;  // the one-output-per-thread kernel (A/B, cross-check)
      const int sh = log2_exact(C8);
      static const bool ups_row = dev_env("RDP_UPSAMPLE_ROW") != nullptr;  // the kernel before the 2x2-block one (A/B, cross-check)
      if (!ups_v1 && !ups_row && sh >= 0 && dy == 0 && dx == 0 && Ls.H == 2 * Ll.H && Ls.W == 2 * Ll.W && Ll.H > 1 && Ll.W > 1) {
        const float ry = (float)(Ll.H - 1) / (float)(2 * Ll.H - 1), rx = (float)(Ll.W - 1) / (float)(2 * Ll.W - 1);
        k_upsample2x_quad<<<dim3(((Ll.W + 1) * C8 + 255) / 256, (Ll.H + 1 + UQ_ROWS - 1) / UQ_ROWS, B), 256, 0, stream>>>(buf.act[low_i], Ll.H, Ll.W, C8, sh,
                                                                                                                       ry, rx, buf.up[j]);
      } else if (ups_v1 || sh < 0 || C8 > 256) {
        k_upsample2x<<<dim3((Ls.W * C8 + 255) / 256, Ls.H, B), 256, 0, stream>>>(buf.act[low_i], Ll.H, Ll.W, C8, sh, Ls.H, Ls.W, dy / 2, dx / 2, buf.up[j]);
      } else {
        const int Ho = 2 * Ll.H, Wo = 2 * Ll.W;
        const float ry = Ho > 1 ? (float)(Ll.H - 1) / (float)(Ho - 1) : 0.f, rx = Wo > 1 ? (float)(Ll.W - 1) / (float)(Wo - 1) : 0.f;
        const int per_cta = (256 >> sh) * UPS_PX;
        k_upsample2x_row<<<dim3((Ls.W + per_cta - 1) / per_cta, Ls.H, B), 256, 0, stream>>>(buf.act[low_i], Ll.H, Ll.W, C8, sh, Ls.H, Ls.W,
                                                                                           dy / 2, dx / 2, ry, rx, buf.up[j]);
      }
      mark();
      RDP_TRY(conv(10 + 2 * j, buf.act[skip_i], Cs, buf.up[j], Cl, 9));
    }
    RDP_TRY(conv(11 + 2 * j, buf.act[10 + 2 * j], net->w.conv[10 + 2 * j].cout, nullptr, 0, 9));
  }
#undef RDP_TRY
  RDP_CUDA_TRY(cudaGetLastError());
  if (!maps_cached) {  // remember this (workspace, shape): least recently used entry goes
    maps.ws = workspace; maps.B = B; maps.H = H; maps.W = W; maps.keep = (int)keep_all; maps.n = map_cursor;
    std::lock_guard<std::mutex> guard(net->map_mu);
    MapSet* victim = nullptr;
    for (MapSet& e : net->map_cache)  // another thread may have stored the same key meanwhile
      if (e.n > 0 && e.ws == workspace && e.B == B && e.H == H && e.W == W && e.keep == (int)keep_all) victim = &e;
    for (MapSet& e : net->map_cache)
      if (!victim && e.n == 0) victim = &e;
    if (!victim) {
      victim = &net->map_cache[0];
      for (MapSet& e : net->map_cache)
        if (e.stamp < victim->stamp) victim = &e;
    }
    maps.stamp = ++net->map_clock;
    *victim = maps;
  }
  return RDP_OK;
}

extern "C" int rdp_unet_forward(const RdpUnet* net, const void* input, int in_fmt, int B, int H, int W, float* logits,
                                uint8_t* mask, void* workspace, size_t workspace_bytes, void* stream) {
  return forward_impl(net, input, in_fmt, B, H, W, logits, mask, workspace, workspace_bytes, stream, nullptr);
}

extern "C" int rdp_unet_forward_profiled(const RdpUnet* net, const void* input, int in_fmt, int B, int H, int W,
                                         float* logits, uint8_t* mask, void* workspace, size_t workspace_bytes,
                                         void* stream, float* launch_ms) {
  RDP_REQUIRE(launch_ms, RDP_ERR_BAD_ARGUMENT, "rdp_unet_forward_profiled: launch_ms is null");
  const int n = rdp_unet_launch_count(net);
  cudaEvent_t ev[64];
  for (int i = 0; i <= n; ++i) RDP_CUDA_TRY(cudaEventCreate(&ev[i]));
  int rc = forward_impl(net, input, in_fmt, B, H, W, logits, mask, workspace, workspace_bytes, stream, ev);
  if (rc == RDP_OK && cudaEventSynchronize(ev[n]) != cudaSuccess) {
    set_error("rdp_unet_forward_profiled: %s", cudaGetErrorString(cudaGetLastError()));
    rc = RDP_ERR_CUDA;
  }
  for (int i = 0; i < n && rc == RDP_OK; ++i) cudaEventElapsedTime(&launch_ms[i], ev[i], ev[i + 1]);
  for (int i = 0; i <= n; ++i) cudaEventDestroy(ev[i]);
  return rc;
}

extern "C" int rdp_unet_read_activation(const RdpUnet* net, int index, int B, int H, int W, const void* workspace,
                                        float* out_nchw, void* stream_) {
  RDP_REQUIRE(net && workspace && out_nchw && index >= 0 && index < 18, RDP_ERR_BAD_ARGUMENT, "rdp_unet_read_activation: bad argument");
  const Buffers buf = carve_unet(net, const_cast<void*>(workspace), B, H, W, /*keep_all=*/true);
  const Level L = buf.lv[kConvLevel[index]];
  const int C = net->w.conv[index].cout;
  const size_t total = (size_t)B * C * L.H * L.W;
  k_nhwc_to_nchw_f32<<<(int)std::min<size_t>((total + 255) / 256, 65535), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      buf.act[index], B, L.H, L.W, C, out_nchw);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}
