// Geometry stage of the vision hot path on sm_100a: masked depth deprojection with ordered stream
// compaction, x-binning, per-bin top-5% selection, x ordering, least-squares cubic fit and curvature.
//
// Replaces, for a batch of frames, compute_curvature_profile and its helpers in the reference
// (pkg/geometry_utils.py:42-162).  All *decisions* (validity, min/max, bin index, selection, ordering) are
// taken in fp64 with the reference's operation order and no FMA contraction, so they are bit-identical to
// NumPy's; the emitted point cloud is fp32.  Ties, which NumPy's argpartition/argsort leave unspecified, are
// broken by (larger y, then smaller pixel index) and (x, then pixel index).
//
// Data layout (HBM): mask u8 [B][P], depth u16|f32 [B][P] (P = H*W), workspace arrays indexed [b][i].
// Kernels (one launch each per batch, all on the caller's stream, no host sync):
//   k_deproject    HBM-bound   reads 3P (u16) bytes/frame once; writes 4N (pixel ids) + 16N (fp64 x and y as ordered keys,
//                              computed while the tile's depth is still in L1) + 12N (fp32 xyz, on request); x range by
//                              one pair of atomics per tile; 16384-pixel tiles, ordered compaction by decoupled look-back
//   k_bin_hist     L2/HBM      reads 8N, writes N (bin ids); the last block of a frame turns the 50 counts into k per bin,
//                              offsets and the frame status (no separate launch)
//   k_scatter      L2/HBM      reads 13N, writes 12N (pixel ids and y keys grouped by bin)
//   k_select_sort  latency     one CTA per (frame, bin): the bin's keys staged in shared memory, radix select of the k
//                              largest y, bitonic sort by x
//   k_fit          latency     one CTA per frame: chord-length scan, 4x4 normal equations in fp64, curvature
#include <stdarg.h>

#include <mutex>
#include <type_traits>

#include "rdp_common.cuh"

// Development switches are compiled out of the shipped library (see unet.cu): build with -DRDP_DEVELOPMENT to get them.
#ifdef RDP_DEVELOPMENT
static inline const char* dev_env(const char* name) { return getenv(name); }
#else
static inline const char* dev_env(const char*) { return nullptr; }
#endif

namespace rdp {
namespace geom {

constexpr int NUM_BINS = 50;      // geometry_utils.py:119
constexpr int MIN_POINTS = 100;   // :64
constexpr int MIN_EDGE = 20;      // :69
constexpr int NUM_SAMPLES = 100;  // :83,144
constexpr double SMOOTH_S = 0.1;  // :78
constexpr double FIT_TOL = 1e-3;  // FITPACK parcur tol

constexpr int DP_THREADS = 256;
constexpr int DP_CHUNKS = 4;                             // 16-pixel chunks per thread, all loads issued before use
constexpr int DP_PX_PER_THREAD = 16 * DP_CHUNKS;
constexpr int TILE_PX = DP_THREADS * DP_PX_PER_THREAD;  // 16384 pixels per deprojection tile
constexpr int PT_THREADS = 256;
constexpr int PT_PER_BLOCK = 2048;  // points per block in the histogram / scatter passes
constexpr int SEL_THREADS = 128;
constexpr int SORT_CAP = 1024;   // selected points sorted in shared memory up to this many (global memory beyond)

struct FrameState {
  unsigned long long xmax_key;      // atomicMax over ordered keys of x
  unsigned long long xmin_inv_key;  // atomicMax over ~key, so that a zeroed struct is the identity
  int n_points;
  int n_edge;
  int status;
  int hist_done;  // blocks of k_bin_hist that have added their counts (the last one finalises the frame)
  int bin_count[NUM_BINS];
  int bin_k[NUM_BINS];
  int bin_off[NUM_BINS + 1];
  int edge_off[NUM_BINS + 1];
  int bin_cursor[NUM_BINS];
  unsigned char bin_done[NUM_BINS + 2];  // set by k_select_sort_warp for the bins it finished (the CTA kernel takes the others)
};

struct Params {
  const uint8_t* mask;
  const void* depth;
  int depth_fmt;
  int B, H, W, P, tiles;
  double fx, fy, cx, cy, scale;
  float scale_f32;
  int ecap;
  // workspace
  FrameState* fs;
  unsigned long long* tile_state;  // [B][tiles] decoupled look-back descriptors
  unsigned int* ticket;
  uint32_t* pix;       // [B][P] valid pixel ids, row-major order
  uint8_t* bins;       // [B][P] bin id per valid point
  uint32_t* binned;    // [B][P] pixel ids grouped by bin
  unsigned long long* xkey_c;  // [B][P] ordered fp64 x key per valid point (compacted order), written by k_deproject
  unsigned long long* ykey_c;  // [B][P] ordered fp64 y key per valid point (compacted order)
  unsigned long long* ykey_b;  // [B][P] the same, grouped by bin like `binned`
  uint32_t* edge_pix;  // [B][ecap] selected pixel ids, ordered by x
  unsigned long long* edge_key;  // [B][ecap] sort keys (only touched when a bin overflows SORT_CAP)
  double* edge_xyz;    // [B][ecap][3]
  double* edge_u;      // [B][ecap] cumulative chord length
  // outputs
  RdpCurvatureOut out;
};

// ---------------------------------------------------------------------------------------------------------
// fp64 arithmetic in NumPy's order, never contracted into FMA
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double depth_to_z(const Params& p, size_t gi) {
  if (p.depth_fmt == RDP_DEPTH_U16) return __dmul_rn((double)((const uint16_t*)p.depth)[gi], p.scale);
  if (p.depth_fmt == RDP_DEPTH_F64) return __dmul_rn(((const double*)p.depth)[gi], p.scale);
  const float d = ((const float*)p.depth)[gi];
  if (p.depth_fmt == RDP_DEPTH_F32) return (double)__fmul_rn(d, p.scale_f32);
  return __dmul_rn((double)d, p.scale);
}
__device__ __forceinline__ double deproject_axis(double pixel, double c, double z, double f) {
  return __ddiv_rn(__dmul_rn(__dsub_rn(pixel, c), z), f);  // (u - cx) * z / fx   geometry_utils.py:114
}
__device__ __forceinline__ unsigned long long ordered_key(double v) {
  const long long b = __double_as_longlong(v);
  return (unsigned long long)(b ^ ((b >> 63) | (long long)0x8000000000000000ull));
}
__device__ __forceinline__ double key_to_double(unsigned long long k) {
  const long long b = (long long)k;
  return __longlong_as_double((b >= 0) ? ~b : (b ^ (long long)0x8000000000000000ull));
}
__device__ __forceinline__ int bin_of(double x, double x_min, double width) {
  const double q = floor(__ddiv_rn(__dsub_rn(x, x_min), width));  // geometry_utils.py:130
  return q < 0.0 ? 0 : (q > (double)(NUM_BINS - 1) ? NUM_BINS - 1 : (int)q);  // clip, :132
}
__device__ __forceinline__ bool frame_bins(const FrameState& f, double& x_min, double& width) {
  if (f.n_points < MIN_POINTS) return false;
  x_min = key_to_double(~f.xmin_inv_key);
  const double x_max = key_to_double(f.xmax_key);
  width = __ddiv_rn(__dsub_rn(x_max, x_min), (double)NUM_BINS);  // :125
  return width > 0.0;                                             // :127
}

__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ unsigned long long warp_max(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  return v;
}

// ---------------------------------------------------------------------------------------------------------
// k_deproject: one pass over mask+depth, ordered compaction across the tiles of a frame by decoupled look-back
// ---------------------------------------------------------------------------------------------------------
#define FLAG_AGG (1ull << 62)
#define FLAG_INC (2ull << 62)
#define FLAG_MASK (3ull << 62)

__device__ __forceinline__ int lookback_exclusive(volatile unsigned long long* st, int t, int total, int lane) {
  if (lane == 0) st[t] = (t == 0 ? FLAG_INC : FLAG_AGG) | (unsigned)total;
  if (t == 0) return 0;
  int excl = 0;
  for (int j = t - 1;; j -= 32) {
    const int idx = j - lane;
    unsigned long long s = idx >= 0 ? st[idx] : FLAG_INC;
    while (__any_sync(0xffffffffu, (s & FLAG_MASK) == 0)) {
      if (idx >= 0 && (s & FLAG_MASK) == 0) s = st[idx];
    }
    const unsigned inc = __ballot_sync(0xffffffffu, (s & FLAG_MASK) == FLAG_INC);
    const int first = inc ? __ffs(inc) - 1 : 32;
    excl += warp_sum(lane <= first ? (int)(unsigned)(s & 0xffffffffull) : 0);
    if (inc) break;
  }
  if (lane == 0) st[t] = FLAG_INC | (unsigned)(excl + total);
  return excl;
}

template <bool VEC>
__global__ void __launch_bounds__(DP_THREADS) k_deproject(const Params p) {
  __shared__ uint16_t s_pix[TILE_PX];  // tile-relative pixel offsets of the valid points
  __shared__ int s_warp[DP_THREADS / 32];
  __shared__ int s_ticket, s_base, s_total;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_ticket = (int)atomicAdd(p.ticket, 1u);  // tiles are claimed in launch order: no look-back deadlock
  __syncthreads();
  const int b = s_ticket / p.tiles, t = s_ticket % p.tiles;
  const size_t fbase = (size_t)b * p.P;
  const int px0 = t * TILE_PX + tid * DP_PX_PER_THREAD;

  // 64 consecutive pixels per thread: 4 x (16 mask bytes + 16 depths), every load in flight before the first use
  unsigned valid[DP_CHUNKS];
#pragma unroll
  for (int c = 0; c < DP_CHUNKS; ++c) valid[c] = 0;
  if (VEC && px0 + DP_PX_PER_THREAD <= p.P) {
    uint4 m4[DP_CHUNKS];
    const uint4* mp = reinterpret_cast<const uint4*>(p.mask + fbase + px0);
#pragma unroll
    for (int c = 0; c < DP_CHUNKS; ++c) m4[c] = __ldg(mp + c);
    if (p.depth_fmt == RDP_DEPTH_U16) {
      uint4 d4[2 * DP_CHUNKS];
      const uint4* dp = reinterpret_cast<const uint4*>((const uint16_t*)p.depth + fbase + px0);
#pragma unroll
      for (int c = 0; c < 2 * DP_CHUNKS; ++c) d4[c] = __ldg(dp + c);
#pragma unroll
      for (int c = 0; c < DP_CHUNKS; ++c) {
        const uint32_t mw[4] = {m4[c].x, m4[c].y, m4[c].z, m4[c].w};
        const uint32_t dw[8] = {d4[2 * c].x, d4[2 * c].y, d4[2 * c].z, d4[2 * c].w,
                                d4[2 * c + 1].x, d4[2 * c + 1].y, d4[2 * c + 1].z, d4[2 * c + 1].w};
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const uint32_t mb = (mw[i >> 2] >> (8 * (i & 3))) & 0xffu;
          const uint32_t dv = (dw[i >> 1] >> (16 * (i & 1))) & 0xffffu;
          if (mb && __dmul_rn((double)dv, p.scale) > 0.0) valid[c] |= 1u << i;
        }
      }
    } else {
      const float4* dp = reinterpret_cast<const float4*>((const float*)p.depth + fbase + px0);
#pragma unroll
      for (int c = 0; c < DP_CHUNKS; ++c) {
        float4 d[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) d[q] = __ldg(dp + 4 * c + q);
        const uint32_t mw[4] = {m4[c].x, m4[c].y, m4[c].z, m4[c].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float dv[4] = {d[q].x, d[q].y, d[q].z, d[q].w};
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const int i = q * 4 + r;
            const uint32_t mb = (mw[i >> 2] >> (8 * (i & 3))) & 0xffu;
            const double z = p.depth_fmt == RDP_DEPTH_F32 ? (double)__fmul_rn(dv[r], p.scale_f32)
                                                          : __dmul_rn((double)dv[r], p.scale);
            if (mb && z > 0.0) valid[c] |= 1u << i;
          }
        }
      }
    }
  } else {
    for (int i = 0; i < DP_PX_PER_THREAD; ++i) {
      const int px = px0 + i;
      if (px < p.P && p.mask[fbase + px] && depth_to_z(p, fbase + px) > 0.0) valid[i >> 4] |= 1u << (i & 15);
    }
  }

  // block-wide exclusive scan of the per-thread counts
  int cnt = 0;
#pragma unroll
  for (int c = 0; c < DP_CHUNKS; ++c) cnt += __popc(valid[c]);
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = lane < DP_THREADS / 32 ? s_warp[lane] : 0;
    int wi = w;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += v;
    }
    if (lane < DP_THREADS / 32) s_warp[lane] = wi - w;
    const int total = __shfl_sync(0xffffffffu, wi, DP_THREADS / 32 - 1);
    const int base = lookback_exclusive(p.tile_state + (size_t)b * p.tiles, t, total, lane);
    if (lane == 0) {
      s_base = base;
      s_total = total;
      if (t == p.tiles - 1) p.fs[b].n_points = base + total;
    }
  }
  __syncthreads();
  int local = s_warp[warp] + incl - cnt;
#pragma unroll
  for (int c = 0; c < DP_CHUNKS; ++c) {
    unsigned v = valid[c];
    while (v) {
      const int i = __ffs(v) - 1;
      v &= v - 1;
      s_pix[local++] = (uint16_t)(tid * DP_PX_PER_THREAD + c * 16 + i);
    }
  }
  __syncthreads();

  // cooperative, coalesced emission of the tile's points: pixel id, the fp64 x and y as ordered integer keys (what the
  // binning, the selection and the ordering decide on), the frame's x range, and the fp32 cloud when the caller wants
  // it.  The depth the keys need was loaded by this CTA a moment ago (L1/L2 hit), so the stage reads HBM once.
  const int total = s_total, base = s_base;
  const bool keys = p.xkey_c != nullptr;
  unsigned long long kmax = 0, kmin_inv = 0;
  for (int j = tid; j < total; j += DP_THREADS) {
    const uint32_t px = (uint32_t)(t * TILE_PX) + s_pix[j];
    const size_t o = fbase + base + j;
    p.pix[o] = px;
    if (keys || p.out.points_xyz) {
      const int v = (int)(px / (uint32_t)p.W), u = (int)(px - (uint32_t)v * (uint32_t)p.W);
      const double z = depth_to_z(p, fbase + px);
      const double x = deproject_axis((double)u, p.cx, z, p.fx), y = deproject_axis((double)v, p.cy, z, p.fy);
      if (keys) {
        const unsigned long long k = ordered_key(x);
        p.xkey_c[o] = k;
        p.ykey_c[o] = ordered_key(y);
        kmax = k > kmax ? k : kmax;
        kmin_inv = ~k > kmin_inv ? ~k : kmin_inv;
      }
      if (p.out.points_xyz) {
        float* dst = p.out.points_xyz + 3 * o;
        dst[0] = (float)x;
        dst[1] = (float)y;
        dst[2] = (float)z;
      }
    }
  }
  if (keys) {  // x range of the frame (geometry_utils.py:124): block reduction, then one atomic pair per tile
    __shared__ unsigned long long s_red[2][DP_THREADS / 32];
    kmax = warp_max(kmax);
    kmin_inv = warp_max(kmin_inv);
    if (lane == 0) {
      s_red[0][warp] = kmax;
      s_red[1][warp] = kmin_inv;
    }
    __syncthreads();
    if (tid == 0 && total > 0) {
      for (int w = 1; w < DP_THREADS / 32; ++w) {
        kmax = s_red[0][w] > kmax ? s_red[0][w] : kmax;
        kmin_inv = s_red[1][w] > kmin_inv ? s_red[1][w] : kmin_inv;
      }
      atomicMax(&p.fs[b].xmax_key, kmax);
      atomicMax(&p.fs[b].xmin_inv_key, kmin_inv);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// k_bin_hist / k_scatter: counting sort of the points by x-bin
// ---------------------------------------------------------------------------------------------------------
// (Points arrive in row-major pixel order, so the 32 points of a warp fall into one or two bins; grouping the lanes with
// match.any before the shared-memory atomic was measured slower than the plain atomic, which the hardware already folds
// for lanes that hit one address.)
// k per bin, bin offsets, edge offsets and the frame status from the 50 counts (geometry_utils.py:121-138)
// Executed by one whole warp: lane l owns bins l and l + 32, the running offsets are two warp scans.
__device__ void finalize_bins(const Params& p, int b, const volatile int* counts, int lane) {
  FrameState& f = p.fs[b];
  double x_min, width;
  int status = RDP_FRAME_OK;
  if (f.n_points < MIN_POINTS) {
    status = RDP_FRAME_TOO_FEW_POINTS;
  } else if (!frame_bins(f, x_min, width)) {
    status = RDP_FRAME_DEGENERATE_BINS;
  }
  int c[2], k[2], coff[2], koff[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int i = lane + 32 * h;
    c[h] = (status == RDP_FRAME_OK && i < NUM_BINS) ? counts[i] : 0;
    k[h] = 0;
    if (c[h] > 0) {
      k[h] = (int)__dmul_rn((double)c[h], 0.05);  // int(cnt * top_k_percent), geometry_utils.py:138
      k[h] = k[h] < 1 ? 1 : k[h];
    }
  }
  int cbase = 0, kbase = 0;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    int ci = c[h], ki = k[h];  // inclusive scans over the lanes
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int a = __shfl_up_sync(0xffffffffu, ci, o), e = __shfl_up_sync(0xffffffffu, ki, o);
      if (lane >= o) {
        ci += a;
        ki += e;
      }
    }
    coff[h] = cbase + ci - c[h];
    koff[h] = kbase + ki - k[h];
    cbase += __shfl_sync(0xffffffffu, ci, 31);
    kbase += __shfl_sync(0xffffffffu, ki, 31);
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int i = lane + 32 * h;
    if (i < NUM_BINS) {
      f.bin_k[i] = k[h];
      f.bin_off[i] = coff[h];
      f.edge_off[i] = koff[h];
      if (p.out.bin_counts) p.out.bin_counts[b * NUM_BINS + i] = c[h];
    }
  }
  if (lane == 0) {
    f.bin_off[NUM_BINS] = cbase;
    f.edge_off[NUM_BINS] = kbase;
    f.n_edge = kbase;
    if (status == RDP_FRAME_OK && kbase < MIN_EDGE) status = RDP_FRAME_TOO_FEW_EDGE;
    f.status = status;
  }
}

// Points arrive in row-major pixel order, so the 32 points a warp handles at once fall into one or two bins.  The lanes of
// a run (consecutive lanes with the same bin) are counted with one ballot and the run's first lane makes ONE atomic for
// all of them; the others take their place behind it.  Returns the lane's offset inside its run, the run's length
// (valid on its first lane) and the first lane's index.  bin < 0 = no point on this lane.
struct BinRun { int head_lane, offset, length; bool is_head; };
__device__ __forceinline__ BinRun bin_run(int bin, int lane) {
  const int prev = __shfl_up_sync(0xffffffffu, bin, 1);
  const bool head = lane == 0 || prev != bin;
  const unsigned heads = __ballot_sync(0xffffffffu, head);
  BinRun r;
  r.head_lane = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));  // the last head at or below this lane
  r.offset = lane - r.head_lane;
  const unsigned above = lane == 31 ? 0u : (heads >> (lane + 1));
  r.length = above ? __ffs(above) : 32 - lane;                      // distance to the next head (meaningful on heads)
  r.is_head = head;
  return r;
}

__global__ void __launch_bounds__(PT_THREADS) k_bin_hist(const Params p) {
  __shared__ int s_hist[NUM_BINS];
  __shared__ int s_last;
  const int b = blockIdx.y;
  FrameState& f = p.fs[b];
  const int n = f.n_points, i0 = blockIdx.x * PT_PER_BLOCK;
  double x_min, width;
  if (!frame_bins(f, x_min, width)) {  // too few points or zero width: nothing to count, block 0 records the status
    if (blockIdx.x == 0 && threadIdx.x < 32) finalize_bins(p, b, f.bin_count, threadIdx.x);
    return;
  }
  if (i0 >= n) return;
  if (threadIdx.x < NUM_BINS) s_hist[threadIdx.x] = 0;
  __syncthreads();
  const size_t fbase = (size_t)b * p.P;
  const int i1 = min(n, i0 + PT_PER_BLOCK);
  for (int i = i0 + threadIdx.x; i < i1; i += PT_THREADS) {
    const int bin = bin_of(key_to_double(p.xkey_c[fbase + i]), x_min, width);
    p.bins[fbase + i] = (uint8_t)bin;
    atomicAdd(&s_hist[bin], 1);
  }
  __syncthreads();
  if (threadIdx.x < NUM_BINS && s_hist[threadIdx.x]) atomicAdd(&f.bin_count[threadIdx.x], s_hist[threadIdx.x]);
  // the block that adds its counts last turns them into offsets (what used to be a launch of its own)
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int blocks = (n + PT_PER_BLOCK - 1) / PT_PER_BLOCK;
    s_last = atomicAdd(&f.hist_done, 1) == blocks - 1;
  }
  __syncthreads();
  if (s_last && threadIdx.x < 32) {
    __threadfence();
    finalize_bins(p, b, f.bin_count, threadIdx.x);
  }
}

__global__ void __launch_bounds__(PT_THREADS) k_scatter(const Params p) {
  __shared__ int s_cnt[NUM_BINS], s_base[NUM_BINS];
  const int b = blockIdx.y;
  FrameState& f = p.fs[b];
  const int n = f.n_points, i0 = blockIdx.x * PT_PER_BLOCK;
  if (i0 >= n || f.status != RDP_FRAME_OK) return;
  if (threadIdx.x < NUM_BINS) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const size_t fbase = (size_t)b * p.P;
  const int i1 = min(n, i0 + PT_PER_BLOCK);
  constexpr int ITERS = PT_PER_BLOCK / PT_THREADS;
  const int lane = threadIdx.x & 31;
  int bins[ITERS];
  uint32_t pix[ITERS];
  unsigned long long yk[ITERS];
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {  // all loads of the block's share in flight
    const int i = i0 + threadIdx.x + it * PT_THREADS;
    const bool in = i < i1;
    bins[it] = in ? (int)p.bins[fbase + i] : -1;
    pix[it] = in ? p.pix[fbase + i] : 0u;
    yk[it] = in ? p.ykey_c[fbase + i] : 0ull;
  }
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {
    const BinRun r = bin_run(bins[it], lane);
    if (r.is_head && bins[it] >= 0) atomicAdd(&s_cnt[bins[it]], r.length);
  }
  __syncthreads();
  if (threadIdx.x < NUM_BINS) {
    const int c = s_cnt[threadIdx.x];
    s_base[threadIdx.x] = f.bin_off[threadIdx.x] + (c ? atomicAdd(&f.bin_cursor[threadIdx.x], c) : 0);
    s_cnt[threadIdx.x] = 0;
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {
    const int bin = bins[it];
    const BinRun r = bin_run(bin, lane);
    int first = 0;
    if (r.is_head && bin >= 0) first = s_base[bin] + atomicAdd(&s_cnt[bin], r.length);
    first = __shfl_sync(0xffffffffu, first, r.head_lane);
    if (bin >= 0) {  // the order inside a bin is free: the selection and the final sort are total orders on (y, x, pixel)
      const int pos = first + r.offset;
      p.binned[fbase + pos] = pix[it];
      p.ykey_b[fbase + pos] = yk[it];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// k_select_sort: per (frame, bin) the k largest y (ties: smaller pixel id), then ordered by (x, pixel id)
// ---------------------------------------------------------------------------------------------------------
// Composite selection key, most significant first: ordered fp64 y (8 digits of 8 bits), then ~pixel (4 digits).
__device__ __forceinline__ int key_digit(unsigned long long yk, uint32_t npx, int d) {
  return d < 8 ? (int)((yk >> (56 - 8 * d)) & 0xffull) : (int)((npx >> (24 - 8 * (d - 8))) & 0xffu);
}
__device__ __forceinline__ bool key_has_prefix(unsigned long long yk, uint32_t npx, int d, unsigned long long ph,
                                               uint32_t pl) {
  if (d == 0) return true;
  if (d < 8) return (yk >> (64 - 8 * d)) == (ph >> (64 - 8 * d));
  if (yk != ph) return false;
  if (d == 8) return true;
  return (npx >> (32 - 8 * (d - 8))) == (pl >> (32 - 8 * (d - 8)));
}

// Shared memory of k_select_sort.  A bin of up to SEL_STAGE_CAP points is copied in once (y key + pixel id, 12 bytes a
// point) and every radix pass, the gather and the sort then stay on chip; its k <= 5 % selected points are sorted in the
// small arrays.  Larger bins (HD frames, thick bands) keep reading global memory and use the staging area as their
// SORT_CAP-sized sort buffer instead.  The kernel is a chain of short phases between barriers, so what it needs most is
// resident CTAs: a 3072-point staging area (39 KB, 5 CTAs per SM) measured 1.7x slower than this one (14 KB, 16 per SM).
constexpr int SEL_STAGE_CAP = 1024;
constexpr int SEL_SMALL_SORT = SEL_STAGE_CAP / 20 + 4;   // k = max(1, int(cnt * 0.05)) for cnt <= SEL_STAGE_CAP
static_assert(SEL_STAGE_CAP * 12 >= SORT_CAP * 12, "the staging area doubles as the large sort buffer");

__global__ void __launch_bounds__(SEL_THREADS) k_select_sort(const Params p) {
  __shared__ unsigned long long s_stage_key[SEL_STAGE_CAP];   // staged y keys | x sort keys of a large bin
  __shared__ uint32_t s_stage_pix[SEL_STAGE_CAP];             // staged pixel ids | pixel ids of a large bin's selection
  __shared__ unsigned long long s_small_key[SEL_SMALL_SORT];  // x sort keys of a staged bin's selection
  __shared__ uint32_t s_small_pix[SEL_SMALL_SORT];
  __shared__ int s_hist[256];
  __shared__ unsigned long long s_ph, s_mm[2][SEL_THREADS / 32];
  __shared__ uint32_t s_pl;
  __shared__ int s_krem, s_done, s_nsel, s_start;

  const int bin = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const FrameState& f = p.fs[b];
  if (f.status != RDP_FRAME_OK) return;
  const int cnt = f.bin_count[bin], k = f.bin_k[bin];
  if (cnt == 0 || f.bin_done[bin]) return;  // done: k_select_sort_warp had it
  const size_t fbase = (size_t)b * p.P;
  const bool staged = cnt <= SEL_STAGE_CAP;
  const uint32_t* seg = p.binned + fbase + f.bin_off[bin];              // pixel ids of the bin
  const unsigned long long* ykeys = p.ykey_b + fbase + f.bin_off[bin];  // their ordered fp64 y keys

  // one read of the bin from global memory: stage it (when it fits) and find the leading digits every key shares
  // (the keys are clustered: a bin spans a few centimetres)
  unsigned long long kmax = 0, kmin_inv = 0;
  for (int i = tid; i < cnt; i += SEL_THREADS) {
    const unsigned long long yk = ykeys[i];
    if (staged) {
      s_stage_key[i] = yk;
      s_stage_pix[i] = seg[i];
    }
    kmax = yk > kmax ? yk : kmax;
    kmin_inv = ~yk > kmin_inv ? ~yk : kmin_inv;
  }
  kmax = warp_max(kmax);
  kmin_inv = warp_max(kmin_inv);
  if (lane == 0) {
    s_mm[0][warp] = kmax;
    s_mm[1][warp] = kmin_inv;
  }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < SEL_THREADS / 32; ++w) {
      kmax = s_mm[0][w] > kmax ? s_mm[0][w] : kmax;
      kmin_inv = s_mm[1][w] > kmin_inv ? s_mm[1][w] : kmin_inv;
    }
    const unsigned long long diff = kmax ^ ~kmin_inv;
    const int d0 = diff ? __clzll((long long)diff) / 8 : 8;  // digits shared by every key in the segment
    s_start = d0;
    s_ph = d0 == 0 ? 0ull : (d0 == 8 ? kmax : (kmax >> (64 - 8 * d0)) << (64 - 8 * d0));
    s_pl = 0;
    s_krem = k;
    s_done = (k >= cnt);  // everything is selected (tiny bins)
    s_nsel = 0;
  }
  __syncthreads();
  auto key_at = [&](int i) { return staged ? s_stage_key[i] : ykeys[i]; };
  auto pix_at = [&](int i) { return staged ? s_stage_pix[i] : seg[i]; };

  if (!s_done) {
    for (int d = s_start; d < 12; ++d) {
      for (int i = tid; i < 256; i += SEL_THREADS) s_hist[i] = 0;
      __syncthreads();
      const unsigned long long ph = s_ph;
      const uint32_t pl = s_pl;
      for (int i = tid; i < cnt; i += SEL_THREADS) {
        const unsigned long long yk = key_at(i);
        const uint32_t npx = d >= 8 ? ~pix_at(i) : 0u;  // the pixel id only matters once the whole y key is tied
        if (key_has_prefix(yk, npx, d, ph, pl)) atomicAdd(&s_hist[key_digit(yk, npx, d)], 1);
      }
      __syncthreads();
      if (warp == 0) {
        // digit of the k-th largest key: lane l owns buckets 8l..8l+7, suffix sums across lanes by shuffles
        int h[8], own = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          h[e] = s_hist[8 * lane + e];
          own += h[e];
        }
        int above = 0;  // keys in buckets of higher lanes
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int v = __shfl_down_sync(0xffffffffu, own + above, o);
          if (lane + o < 32) above += v;
        }
        const int krem = s_krem;
        const bool mine = above < krem && krem <= above + own;
        if (mine) {
          int cum = above, dig = 7;
          for (; dig > 0; --dig) {
            if (cum + h[dig] >= krem) break;
            cum += h[dig];
          }
          const int digit = 8 * lane + dig;
          if (d < 8) s_ph |= (unsigned long long)digit << (56 - 8 * d);
          else s_pl |= (uint32_t)digit << (24 - 8 * (d - 8));
          s_krem = krem - cum;
          if (h[dig] == krem - cum) s_done = 1;  // the whole bucket is taken: threshold = prefix with zeros below
        }
      }
      __syncthreads();
      if (s_done) break;
    }
  }

  // gather the selected points with their x sort key
  const unsigned long long th = s_done && k >= cnt ? 0ull : s_ph;
  const uint32_t tl = s_done && k >= cnt ? 0u : s_pl;
  const bool sort_in_smem = staged || k <= SORT_CAP;
  const size_t ebase = (size_t)b * p.ecap + f.edge_off[bin];
  unsigned long long* sk = staged ? s_small_key : (sort_in_smem ? s_stage_key : p.edge_key + ebase);
  uint32_t* sp = staged ? s_small_pix : (sort_in_smem ? s_stage_pix : p.edge_pix + ebase);
  for (int i = tid; i < cnt; i += SEL_THREADS) {
    const unsigned long long yk = key_at(i);
    if (yk < th) continue;
    const uint32_t px = pix_at(i);
    if (yk > th || ~px >= tl) {
      const int v = (int)(px / (uint32_t)p.W), u = (int)(px - (uint32_t)v * (uint32_t)p.W);
      const int pos = atomicAdd(&s_nsel, 1);
      if (pos < k) {
        sk[pos] = ordered_key(deproject_axis((double)u, p.cx, depth_to_z(p, fbase + px), p.fx));
        sp[pos] = px;
      }
    }
  }
  __syncthreads();

  // bitonic network on k (not necessarily a power of two) items: indices >= k act as +inf and never move
  auto cmpx = [&](int i, int j) {
    const unsigned long long a = sk[i], c = sk[j];
    const uint32_t ap = sp[i], cp = sp[j];
    if (a > c || (a == c && ap > cp)) {
      sk[i] = c; sk[j] = a;
      sp[i] = cp; sp[j] = ap;
    }
  };
  for (int size = 2; (size >> 1) < k; size <<= 1) {
    for (int i = tid; i < k; i += SEL_THREADS) {
      const int j = i ^ (size - 1);
      if (j > i && j < k) cmpx(i, j);
    }
    __syncthreads();
    for (int stride = size >> 2; stride > 0; stride >>= 1) {
      for (int i = tid; i < k; i += SEL_THREADS) {
        const int j = i | stride;
        if (!(i & stride) && j < k) cmpx(i, j);
      }
      __syncthreads();
    }
  }

  for (int i = tid; i < k; i += SEL_THREADS) {
    const uint32_t px = sp[i];
    const int v = (int)(px / (uint32_t)p.W), u = (int)(px - (uint32_t)v * (uint32_t)p.W);
    const double z = depth_to_z(p, fbase + px);
    double* dst = p.edge_xyz + 3 * (ebase + i);
    dst[0] = deproject_axis((double)u, p.cx, z, p.fx);
    dst[1] = deproject_axis((double)v, p.cy, z, p.fy);
    dst[2] = z;
    p.edge_pix[ebase + i] = px;
    if (p.out.edge_pixels) p.out.edge_pixels[ebase + i] = px;
  }
}

// ---------------------------------------------------------------------------------------------------------
// k_select_sort_warp: the same selection and ordering, one WARP per (frame, bin), no block barriers
// ---------------------------------------------------------------------------------------------------------
// k_select_sort is a chain of ~40 short phases separated by block barriers; with 50 bins per frame of ~1000 points each
// the launch is bound by that latency chain, not by memory.  Here a warp owns a bin and needs three passes over its keys:
//   1. min / max of the y keys;
//   2. a 1024-bucket histogram of the keys, bucket = floor((y - ymin) * 1023 / (ymax - ymin)) - monotone in y, so the
//      k largest keys are every key in the buckets above a boundary bucket plus the best `need` keys of the boundary
//      bucket, which are ranked exactly on the composite key (y, then smaller pixel id);
//   3. gather of the selected points (with their fp64 x) into a per-warp list, ordered by (x, pixel id) by counting ranks.
// Bins with more than WS_MAX_CNT points, more than WS_LIST selected points or more than WS_LIST tied candidates in the
// boundary bucket are left to k_select_sort (bin_done stays 0).  Same total order, same result, bit for bit.
constexpr int WS_WARPS = 4;         // warps (bins) per CTA (8 KB of shared memory each)
constexpr int WS_BUCKETS = 1024;
constexpr int WS_LIST = 256;        // selected points / boundary candidates a warp can hold
constexpr int WS_MAX_CNT = 5000;    // k = int(cnt * 0.05) <= 250

__global__ void __launch_bounds__(WS_WARPS * 32) k_select_sort_warp(const Params p) {
  __shared__ int s_hist[WS_WARPS][WS_BUCKETS];
  __shared__ unsigned long long s_key[WS_WARPS][WS_LIST];   // boundary candidates' y keys, then the selected points' x keys
  __shared__ uint32_t s_px[WS_WARPS][WS_LIST];
  __shared__ uint32_t s_cand_px[WS_WARPS][WS_LIST];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int item = blockIdx.x * WS_WARPS + warp;
  if (item >= p.B * NUM_BINS) return;
  const int b = item / NUM_BINS, bin = item - b * NUM_BINS;
  FrameState& f = p.fs[b];
  if (f.status != RDP_FRAME_OK) return;
  const int cnt = f.bin_count[bin], k = f.bin_k[bin];
  if (cnt == 0) {
    if (lane == 0) f.bin_done[bin] = 1;
    return;
  }
  if (cnt > WS_MAX_CNT || k > WS_LIST) return;
  const size_t fbase = (size_t)b * p.P;
  const uint32_t* seg = p.binned + fbase + f.bin_off[bin];
  const unsigned long long* ykeys = p.ykey_b + fbase + f.bin_off[bin];
  int* hist = s_hist[warp];
  unsigned long long* lkey = s_key[warp];
  uint32_t* lpx = s_px[warp];
  uint32_t* cpx = s_cand_px[warp];

  // pass 1: range of the y keys.  (The passes are latency-bound - one warp, L2 round trips - so every loop keeps
  // several independent loads in flight.)
  unsigned long long kmax = 0, kmin_inv = 0;
#pragma unroll 8
  for (int i = lane; i < cnt; i += 32) {
    const unsigned long long yk = ykeys[i];
    kmax = yk > kmax ? yk : kmax;
    kmin_inv = ~yk > kmin_inv ? ~yk : kmin_inv;
  }
  kmax = warp_max(kmax);
  const unsigned long long kmin = ~warp_max(kmin_inv);
  const double ymin = key_to_double(kmin), ymax = key_to_double(kmax);
  // bucket of a key: monotone non-decreasing in y (a product of non-negative doubles, then truncation)
  const double scale = ymax > ymin ? (double)(WS_BUCKETS - 1) / (ymax - ymin) : 0.0;
  auto bucket_of = [&](unsigned long long yk) {
    const int q = (int)((key_to_double(yk) - ymin) * scale);
    return q < 0 ? 0 : (q > WS_BUCKETS - 1 ? WS_BUCKETS - 1 : q);
  };
  // bucket q lives at hist[(q & 31) * 32 + (q >> 5)]: lane l then reads its 32 consecutive buckets from 32 different banks
  auto slot = [](int q) { return ((q & 31) << 5) | (q >> 5); };
  for (int i = lane; i < WS_BUCKETS; i += 32) hist[i] = 0;
  __syncwarp();
  // pass 2: histogram
#pragma unroll 8
  for (int i = lane; i < cnt; i += 32) atomicAdd(&hist[slot(bucket_of(ykeys[i]))], 1);
  __syncwarp();
  // boundary bucket: the highest bucket t with (number of keys in buckets >= t) >= k.  Lane l owns buckets 32l..32l+31.
  int own = 0;
  for (int e = 0; e < 32; ++e) own += hist[slot(32 * lane + e)];
  int above = 0;  // keys in the buckets of higher lanes
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_down_sync(0xffffffffu, own + above, o);
    if (lane + o < 32) above += v;
  }
  const bool mine = above < k && k <= above + own;
  int tb = 0, need = 0;  // boundary bucket, how many of its keys are selected
  if (mine) {
    int cum = above, e = 31;
    for (; e > 0; --e) {
      if (cum + hist[slot(32 * lane + e)] >= k) break;
      cum += hist[slot(32 * lane + e)];
    }
    tb = 32 * lane + e;
    need = k - cum;
  }
  const unsigned who = __ballot_sync(0xffffffffu, mine);
  const int src = __ffs(who) - 1;  // exactly one lane (k >= 1 and k <= cnt)
  tb = __shfl_sync(0xffffffffu, tb, src);
  need = __shfl_sync(0xffffffffu, need, src);
  const int n_tb = hist[slot(tb)];
  if (n_tb > WS_LIST) return;  // heavy ties at the boundary: the CTA kernel resolves them

  // pass 3: everything above the boundary bucket is selected; the boundary bucket's keys become candidates
  int n_sel = 0, n_cand = 0;  // warp-uniform running counts
  for (int i0 = 0; i0 < cnt; i0 += 128) {
    unsigned long long yk4[4];
    uint32_t px4[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {  // all eight loads of the iteration are issued before the first use
      const int i = i0 + 32 * q + lane;
      yk4[q] = i < cnt ? ykeys[i] : 0ull;
      px4[q] = i < cnt ? seg[i] : 0u;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int i = i0 + 32 * q + lane;
      const int bk = i < cnt ? bucket_of(yk4[q]) : -1;
      const bool sel = bk > tb, cand = bk == tb;
      const unsigned ms = __ballot_sync(0xffffffffu, sel), mc = __ballot_sync(0xffffffffu, cand);
      const unsigned lt = (1u << lane) - 1u;
      if (sel) lpx[n_sel + __popc(ms & lt)] = px4[q];
      if (cand) {
        const int pos = n_cand + __popc(mc & lt);
        lkey[WS_LIST - 1 - pos] = yk4[q];  // candidates are parked at the top of the key list (the x keys fill it from the bottom)
        cpx[pos] = px4[q];
      }
      n_sel += __popc(ms);
      n_cand += __popc(mc);
    }
  }
  __syncwarp();
  // the `need` best candidates: rank on (y descending, pixel id ascending); ranks are distinct (pixel ids are)
  for (int c = lane; c < n_cand; c += 32) {
    const unsigned long long yk = lkey[WS_LIST - 1 - c];
    const uint32_t px = cpx[c];
    int rank = 0;
    for (int j = 0; j < n_cand; ++j) {
      const unsigned long long yj = lkey[WS_LIST - 1 - j];
      const uint32_t pj = cpx[j];
      rank += (yj > yk || (yj == yk && pj < px)) ? 1 : 0;
    }
    if (rank < need) lpx[n_sel + rank] = px;  // positions n_sel .. n_sel + need - 1, any order: the sort follows
  }
  __syncwarp();
  // x keys of the k selected points
  for (int i = lane; i < k; i += 32) {
    const uint32_t px = lpx[i];
    const int v = (int)(px / (uint32_t)p.W), u = (int)(px - (uint32_t)v * (uint32_t)p.W);
    lkey[i] = ordered_key(deproject_axis((double)u, p.cx, depth_to_z(p, fbase + px), p.fx));
  }
  __syncwarp();
  // order by (x, pixel id): rank by counting, then emit
  const size_t ebase = (size_t)b * p.ecap + f.edge_off[bin];
  for (int i = lane; i < k; i += 32) {
    const unsigned long long xk = lkey[i];
    const uint32_t px = lpx[i];
    int rank = 0;
    for (int j = 0; j < k; ++j) rank += (lkey[j] < xk || (lkey[j] == xk && lpx[j] < px)) ? 1 : 0;
    const int v = (int)(px / (uint32_t)p.W), u = (int)(px - (uint32_t)v * (uint32_t)p.W);
    const double z = depth_to_z(p, fbase + px);
    double* dst = p.edge_xyz + 3 * (ebase + rank);
    dst[0] = deproject_axis((double)u, p.cx, z, p.fx);
    dst[1] = deproject_axis((double)v, p.cy, z, p.fy);
    dst[2] = z;
    p.edge_pix[ebase + rank] = px;
    if (p.out.edge_pixels) p.out.edge_pixels[ebase + rank] = px;
  }
  if (lane == 0) f.bin_done[bin] = 1;
}

// ---------------------------------------------------------------------------------------------------------
// k_fit: chord-length parameter, least-squares cubic on knots [0,0,0,0,1,1,1,1], curvature at 100 samples
// ---------------------------------------------------------------------------------------------------------
constexpr int FIT_THREADS = 256;

template <int N>
__device__ __forceinline__ void block_sum(double (&v)[N], double* s_part /*[FIT_THREADS/32][N]*/, double* s_out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < N; ++i) v[i] = warp_sum(v[i]);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < N; ++i) s_part[warp * N + i] = v[i];
  }
  __syncthreads();
  if (threadIdx.x < N) {
    double a = 0.0;
    for (int w = 0; w < FIT_THREADS / 32; ++w) a += s_part[w * N + threadIdx.x];
    s_out[threadIdx.x] = a;
  }
  __syncthreads();
}

__device__ __forceinline__ void bernstein3(double u, double (&bb)[4]) {
  const double w = 1.0 - u;
  bb[0] = w * w * w;
  bb[1] = 3.0 * u * w * w;
  bb[2] = 3.0 * u * u * w;
  bb[3] = u * u * u;
}

__global__ void __launch_bounds__(FIT_THREADS) k_fit(const Params p) {
  __shared__ double s_part[(FIT_THREADS / 32) * 22];
  __shared__ double s_sum[22];
  __shared__ double s_c[12];  // control points [4][3]
  __shared__ double s_carry;
  __shared__ int s_bad;
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const FrameState& f = p.fs[b];
  int status = f.status;
  const int E = f.n_edge;
  const double* X = p.edge_xyz + 3 * (size_t)b * p.ecap;
  double* U = p.edge_u + (size_t)b * p.ecap;
  double mean_k = 0.0, max_k = 0.0, fp = 0.0;

  if (status == RDP_FRAME_OK) {
    // cumulative chord length (parcur, ipar=0)
    if (tid == 0) {
      s_carry = 0.0;
      s_bad = 0;
    }
    __syncthreads();
    for (int i0 = 0; i0 < E; i0 += FIT_THREADS) {
      const int i = i0 + tid;
      double d = 0.0;
      if (i > 0 && i < E) {
        const double dx = X[3 * i] - X[3 * i - 3], dy = X[3 * i + 1] - X[3 * i - 2], dz = X[3 * i + 2] - X[3 * i - 1];
        d = sqrt(dx * dx + dy * dy + dz * dz);
      }
      double incl = d;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const double v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) s_part[warp] = incl;
      __syncthreads();
      double woff = 0.0;
      for (int w = 0; w < warp; ++w) woff += s_part[w];
      const double carry = s_carry;
      if (i < E) U[i] = carry + woff + incl;
      __syncthreads();
      if (tid == FIT_THREADS - 1) s_carry = carry + woff + incl;
      __syncthreads();
    }
    const double total = s_carry;
    auto unorm = [&](int i) { return i == 0 ? 0.0 : (i == E - 1 ? 1.0 : U[i] / total); };
    if (!(total > 0.0)) {
      status = RDP_FRAME_FIT_INVALID;
    } else {
      // normal equations of the Bernstein collocation matrix: 10 Gram entries + 12 right-hand sides
      double acc[22];
#pragma unroll
      for (int q = 0; q < 22; ++q) acc[q] = 0.0;
      int bad = 0;
      for (int i = tid; i < E; i += FIT_THREADS) {
        const double u = unorm(i);
        if (i > 0 && !(u > unorm(i - 1))) bad = 1;  // parcur: u must be strictly increasing (ier=10 -> ValueError)
        double bb[4];
        bernstein3(u, bb);
        int q = 0;
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int c = r; c < 4; ++c) acc[q++] += bb[r] * bb[c];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 3; ++c) acc[q++] += bb[r] * X[3 * i + c];
      }
      if (bad) s_bad = 1;
      block_sum<22>(acc, s_part, s_sum);
      if (s_bad) {
        status = RDP_FRAME_FIT_INVALID;
      } else {
        if (tid == 0) {
          // Cholesky of the 4x4 Gram matrix, three right-hand sides
          double G[4][4], L[4][4] = {};
          int q = 0;
          for (int r = 0; r < 4; ++r)
            for (int c = r; c < 4; ++c) G[r][c] = G[c][r] = s_sum[q++];
          for (int r = 0; r < 4; ++r)
            for (int c = 0; c <= r; ++c) {
              double s = G[r][c];
              for (int m = 0; m < c; ++m) s -= L[r][m] * L[c][m];
              L[r][c] = r == c ? sqrt(s) : s / L[c][c];
            }
          for (int c = 0; c < 3; ++c) {
            double y[4], x[4];
            for (int r = 0; r < 4; ++r) {
              double s = s_sum[10 + r * 3 + c];
              for (int m = 0; m < r; ++m) s -= L[r][m] * y[m];
              y[r] = s / L[r][r];
            }
            for (int r = 3; r >= 0; --r) {
              double s = y[r];
              for (int m = r + 1; m < 4; ++m) s -= L[m][r] * x[m];
              x[r] = s / L[r][r];
            }
            for (int r = 0; r < 4; ++r) s_c[r * 3 + c] = x[r];
          }
        }
        __syncthreads();
        // residual of the 8-knot fit decides between FITPACK's two regimes
        double res[1] = {0.0};
        for (int i = tid; i < E; i += FIT_THREADS) {
          double bb[4];
          bernstein3(unorm(i), bb);
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const double e = bb[0] * s_c[c] + bb[1] * s_c[3 + c] + bb[2] * s_c[6 + c] + bb[3] * s_c[9 + c] - X[3 * i + c];
            res[0] += e * e;
          }
        }
        block_sum<1>(res, s_part, s_sum);
        fp = s_sum[0];
        const double fpms = fp - SMOOTH_S;
        if (!(fabs(fpms) < FIT_TOL * SMOOTH_S || fpms < 0.0)) status = RDP_FRAME_NEEDS_KNOTS;
      }
    }
  }

  // curvature and spline samples (geometry_utils.py:144-162, :83-87); zeros for frames without a fit
  double samp[3] = {0.0, 0.0, 0.0};  // sum of kappa, count, (unused)
  unsigned long long kmax_key = 0;
  if (tid < NUM_SAMPLES) {
    double pt[3] = {0.0, 0.0, 0.0};
    if (status == RDP_FRAME_OK) {
      const double u = tid == NUM_SAMPLES - 1 ? 1.0 : (double)tid * (1.0 / (double)(NUM_SAMPLES - 1));
      const double w = 1.0 - u;
      double bb[4];
      bernstein3(u, bb);
      const double q0 = w * w, q1 = 2.0 * u * w, q2 = u * u;
      double d1[3], d2[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const double c0 = s_c[c], c1 = s_c[3 + c], c2 = s_c[6 + c], c3 = s_c[9 + c];
        pt[c] = bb[0] * c0 + bb[1] * c1 + bb[2] * c2 + bb[3] * c3;
        const double e0 = 3.0 * (c1 - c0), e1 = 3.0 * (c2 - c1), e2 = 3.0 * (c3 - c2);
        d1[c] = q0 * e0 + q1 * e1 + q2 * e2;
        d2[c] = w * (2.0 * (e1 - e0)) + u * (2.0 * (e2 - e1));
      }
      const double cx = d1[1] * d2[2] - d1[2] * d2[1], cy = d1[2] * d2[0] - d1[0] * d2[2],
                   cz = d1[0] * d2[1] - d1[1] * d2[0];
      const double n1 = sqrt(d1[0] * d1[0] + d1[1] * d1[1] + d1[2] * d1[2]);
      if (n1 > 1e-6) {
        const double kappa = sqrt(cx * cx + cy * cy + cz * cz) / (n1 * n1 * n1);
        samp[0] = kappa;
        samp[1] = 1.0;
        kmax_key = ordered_key(kappa);
      }
    }
    double* dst = p.out.spline_points + ((size_t)b * NUM_SAMPLES + tid) * 3;
    dst[0] = pt[0];
    dst[1] = pt[1];
    dst[2] = pt[2];
  }
  __syncthreads();
  block_sum<3>(samp, s_part, s_sum);
  kmax_key = warp_max(kmax_key);
  __shared__ unsigned long long s_km[FIT_THREADS / 32];
  if (lane == 0) s_km[warp] = kmax_key;
  __syncthreads();
  if (tid == 0) {
    if (status == RDP_FRAME_OK && s_sum[1] > 0.0) {
      for (int w = 1; w < FIT_THREADS / 32; ++w) kmax_key = s_km[w] > kmax_key ? s_km[w] : kmax_key;
      mean_k = s_sum[0] / s_sum[1];
      max_k = key_to_double(kmax_key);
    }
    p.out.mean_curvature[b] = mean_k;
    p.out.max_curvature[b] = max_k;
    p.out.status[b] = status;
    if (p.out.n_points) p.out.n_points[b] = f.n_points;
    if (p.out.n_edge) p.out.n_edge[b] = E;
    if (p.out.fit_residual) p.out.fit_residual[b] = fp;
    if (p.out.n_knots) p.out.n_knots[b] = 8;
    if (p.out.fit_ier) p.out.fit_ier[b] = status == RDP_FRAME_OK ? -2 : 0;
    if (p.out.fit_flags) p.out.fit_flags[b] = 0;
  }
}


// ---------------------------------------------------------------------------------------------------------
// k_fit_knots: FITPACK's general regime (knot insertion + smoothing) for frames whose 8-knot cubic leaves fp > s
// ---------------------------------------------------------------------------------------------------------
// Restates parcur/fppara (P. Dierckx; reached through scipy.interpolate.splprep, geometry_utils.py:78) for k=3,
// idim=3, unit weights, s=0.1, as pinned in oracle/fitpack_port.py.  One CTA per frame.  FITPACK rotates the
// collocation rows into a banded triangle one by one; the triangle (with a positive diagonal) is unique, so the
// same least-squares spline is obtained here from a two-level reduction: (1) every knot interval's rows only touch
// 4 columns and are reduced to a 4x4 triangle by one thread each, in parallel; (2) one thread merges the interval
// triangles, interleaved by last column with the smoothing-penalty rows when there are any, into the band - in
// that order no fill-in ever runs past a row's own last column, so the merge is O(n) instead of FITPACK's O(n^2)
// smoothing step.  Knot selection (fpknot) works on (residual, point-count, begin) nodes and a flag per data point.
constexpr int KN_THREADS = 256;  // most threads per CTA (array sizes); the kernel runs with blockDim.x = 256 or 128
#define KN_NT ((int)blockDim.x)   // threads of this launch: 128 lets two CTAs share an SM when there are more frames than SMs
constexpr int KN_MAXIT = 20;

// FITPACK's bracket search multiplies and divides p by 0.04 and then compares with the bracket end it started from:
// (p1 / 0.04) * 0.04 <= p1 holds in exact arithmetic and is decided by the last bit in floating point (true for about
// 93 % of the values of p1), where p1 = nk1 / sum(diag R) carries the rounding of every rotation before it.  The order
// of the rotations differs here, so that bit cannot be reproduced; the tie is resolved as exact arithmetic resolves
// it, which is also FITPACK's own outcome in the large majority of cases.
constexpr double KN_TIE_UP = 1.0 + 1e-12, KN_TIE_DOWN = 1.0 - 1e-12;
constexpr int KN_CH = 16;  // data points per triangle task (an interval with more points is cut into several)
// Two interval residuals closer than this (in units of the last place: 2^23 ulp is 1-2e-9 relative) cannot be told
// apart by solves that order their rotations differently: measured noise between the cyclic reduction, the one-thread
// sweep and FITPACK's row-by-row order is 1e-13 ... 1e-11 relative.
constexpr unsigned long long KN_NEAR_TIE = 1ull << 23;

struct KnotWs {
  double *t, *q, *term, *fpint, *tri, *R, *z, *c, *D;
  int *lidx, *ibegin, *nrd, *nbeg, *tbeg, *tint, *tfirst;
  unsigned char* flag;
  double *cr_ga, *cr_gb, *cr_e, *cr_c, *cr_scr;  // cyclic-reduction buffers, see cr_solve
};
__host__ __device__ inline int knot_task_cap(int cap) { return cap + cap / KN_CH + 2; }

__device__ __forceinline__ void givens(double piv, double& ww, double& cs, double& sn) {  // fpgivs
  const double store = fabs(piv);
  double dd;
  if (store >= ww) {
    const double r = ww / piv;
    dd = store * sqrt(1.0 + r * r);
  } else {
    const double r = piv / ww;
    dd = ww * sqrt(1.0 + r * r);
  }
  cs = ww / dd;
  sn = piv / dd;
  ww = dd;
}
__device__ __forceinline__ void rota(double cs, double sn, double& a, double& b) {  // fprota
  const double s1 = a, s2 = b;
  b = cs * s2 + sn * s1;
  a = cs * s1 - sn * s2;
}
// cubic B-spline basis at x in [t[l], t[l+1]) (fpbspl)
__device__ __forceinline__ void bspl3(const double* t, double x, int l, double (&h)[4]) {
  double hh[3];
  h[0] = 1.0;
  h[1] = h[2] = h[3] = 0.0;
#pragma unroll
  for (int j = 1; j <= 3; ++j) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
      if (i < j) hh[i] = h[i];
    h[0] = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      if (i < j) {
        const int li = l + 1 + i, lj = li - j;
        const double f = hh[i] / (t[li] - t[lj]);
        h[i] = h[i] + f * (t[li] - x);
        h[i + 1] = f * (x - t[lj]);
      }
    }
  }
}

// Rotate one row (first column f, `len` consecutive entries h, right-hand side x) into the banded triangle.
__device__ __forceinline__ void band_rotate_in(double* R, double* z, int f, int len, double* h, double (&x)[3]) {
  for (int s = 0; s < len; ++s) {
    const int j = f + s;
    const double piv = h[s];
    if (piv == 0.0) continue;
    double cs, sn;
    givens(piv, R[j * 5], cs, sn);
#pragma unroll
    for (int d = 0; d < 3; ++d) rota(cs, sn, x[d], z[j * 3 + d]);
    for (int e = s + 1; e < len; ++e) rota(cs, sn, h[e], R[j * 5 + (e - s)]);
  }
}

// The same plane rotation as fpgivs up to rounding, with one rsqrt on the dependency chain instead of two
// divisions and a square root: dd = sqrt(ww^2 + piv^2), cs = ww/dd, sn = piv/dd.  piv == 0 is the identity
// (FITPACK skips such pivots).  Magnitudes here (coordinates in metres, B-spline values, scaled penalty rows)
// are far from the fp64 range limits, so the unscaled sum of squares is safe.
// 1/sqrt(a) for a > 0 in the normal range without the library's special-case branch (a branch per rotation would
// cut the merge into basic blocks and stop ptxas from interleaving independent rotations): hardware seed (about 22
// bits) and two coupled Newton steps, g -> sqrt(a), h -> 1/(2 sqrt(a)); a few ulp.
__device__ __forceinline__ double rsqrt_pos(double a) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  double g = a * y, h = 0.5 * y;
  double r = fma(-g, h, 0.5);
  g = fma(g, r, g);
  h = fma(h, r, h);
  r = fma(-g, h, 0.5);
  h = fma(h, r, h);
  return h + h;
}
__device__ __forceinline__ void givens_fast(double piv, double& ww, double& cs, double& sn) {
  const double s2 = fma(piv, piv, ww * ww);
  const double inv = rsqrt_pos(s2);
  // Entries that are zero in exact arithmetic survive as rounding noise and, rotated level after level by the
  // cyclic reduction, shrink until their square underflows (1e-166 seen): below 1e-150 a pivot counts as zero.
  // Nothing legitimate is that small here (coordinates in metres, basis values, scaled knot differences).
  const bool nz = fabs(piv) > 1e-150;  // also covers s2 == 0 (inv = inf, discarded)
  cs = nz ? ww * inv : 1.0;
  sn = nz ? piv * inv : 0.0;
  ww = nz ? s2 * inv : ww;
}
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// Rotate a row of LEN entries h (right-hand side x) into rows S0.. of the register window of the band.
template <int S0, int LEN>
__device__ __forceinline__ void window_rotate_in(double (&Rw)[5][5], double (&zw)[5][3], double (&h)[5], double (&x)[3]) {
#pragma unroll
  for (int s = 0; s < LEN; ++s) {
    double cs, sn;
    givens_fast(h[s], Rw[S0 + s][0], cs, sn);
#pragma unroll
    for (int d = 0; d < 3; ++d) rota(cs, sn, x[d], zw[S0 + s][d]);
#pragma unroll
    for (int e = s + 1; e < LEN; ++e) rota(cs, sn, h[e], Rw[S0 + s][e - s]);
  }
}

// One thread merges the task triangles (in point order) and, when pinv > 0, the penalty rows into the banded
// triangle R (5 entries per row) with right-hand side z, then back-substitutes into c.  Interval i's rows touch band
// rows i..i+3 and penalty row i-1 touches i-1..i+3, so the five rows in flight live in registers and row i-1 is
// final once interval i is done.  Inputs are prefetched into L1 a few intervals ahead: they were written by other
// threads (or, for the back substitution, by this one: stores do not allocate in L1).
// (No __restrict__/const-cache loads: every input was written earlier in this kernel.)
// DIAG_ONLY: only the triangle is wanted (FITPACK's first smoothing parameter needs sum(diag R)): the right-hand sides
// are not rotated and nothing is solved.
// rem_out (may be null): sum of squares of what is left of the right-hand sides once their rows have been rotated in -
// FITPACK's own fp of the least-squares stage (fppara accumulates it the same way), see solve().
template <bool PEN, bool DIAG_ONLY = false>
__device__ __noinline__ void band_merge_solve(const double* tri, const int* tfirst, const double* D, double pinv, int nrint,
                                              double* R, double* z, double* c, double* rem_out = nullptr) {
  double rem = 0.0;
  double Rw[5][5], zw[5][3];
#pragma unroll
  for (int r = 0; r < 5; ++r) {
#pragma unroll
    for (int e = 0; e < 5; ++e) Rw[r][e] = 0.0;
#pragma unroll
    for (int d = 0; d < 3; ++d) zw[r][d] = 0.0;
  }
  const int nk1 = nrint + 3;
  // Rotates one task triangle (and, with WITH_PEN, the penalty row behind it) into the window as a wavefront.  Row q
  // (q = 0: the triangle's last row, one entry ... q = 3: its first row, four entries; q = 4: the penalty row, five)
  // starts at window row 4 - q, so its s-th rotation touches window row 4 - q + s: within one step s the rows touch
  // different window rows, and a window row sees the rows in the order q = 0, 1, .. - the same result as rotating the
  // rows in one after the other, but the (up to five) rotations of a step are independent.  Their Newton chains are
  // written side by side, array-wise, so that the in-order issue stream interleaves them instead of stalling on one.
  auto wave_in = [&](const double* o, const double* dr, double scale, auto with_pen) {
    constexpr int NQ = decltype(with_pen)::value ? 5 : 4;
    double v[22];
#pragma unroll
    for (int q = 0; q < 22; ++q) v[q] = o[q];
    double h[5][5], x[5][3];
    h[0][0] = v[9];
    h[1][0] = v[7]; h[1][1] = v[8];
    h[2][0] = v[4]; h[2][1] = v[5]; h[2][2] = v[6];
    h[3][0] = v[0]; h[3][1] = v[1]; h[3][2] = v[2]; h[3][3] = v[3];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int d = 0; d < 3; ++d) x[q][d] = v[10 + 3 * (3 - q) + d];
    if (NQ == 5) {
#pragma unroll
      for (int e = 0; e < 5; ++e) h[4][e] = dr[e] * scale;  // a zero row rotates as the identity
      x[4][0] = x[4][1] = x[4][2] = 0.0;
    }
#pragma unroll
    for (int st = 0; st < NQ; ++st) {
      double s2[5], y[5], g[5], hh[5], r[5], cs[5], sn[5];
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) s2[q] = fma(h[q][st], h[q][st], Rw[4 - q + st][0] * Rw[4 - q + st][0]);
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y[q]) : "d"(s2[q]));
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) {
          g[q] = s2[q] * y[q];
          hh[q] = 0.5 * y[q];
        }
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) r[q] = fma(-g[q], hh[q], 0.5);
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) {
          g[q] = fma(g[q], r[q], g[q]);
          hh[q] = fma(hh[q], r[q], hh[q]);
        }
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) r[q] = fma(-g[q], hh[q], 0.5);
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) {
          hh[q] = fma(hh[q], r[q], hh[q]);
          const double inv = hh[q] + hh[q];
          const bool nz = fabs(h[q][st]) > 1e-150;
          double& ww = Rw[4 - q + st][0];
          cs[q] = nz ? ww * inv : 1.0;
          sn[q] = nz ? h[q][st] * inv : 0.0;
          ww = nz ? s2[q] * inv : ww;
        }
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        if (st <= q) {
#pragma unroll
          for (int e = st + 1; e <= q; ++e) rota(cs[q], sn[q], h[q][e], Rw[4 - q + st][e - st]);
#pragma unroll
          for (int d = 0; d < 3; ++d)
            if (!DIAG_ONLY) rota(cs[q], sn[q], x[q][d], zw[4 - q + st][d]);
        }
    }
    if (!DIAG_ONLY) {
#pragma unroll
      for (int q = 0; q < 4; ++q) rem += x[q][0] * x[q][0] + x[q][1] * x[q][1] + x[q][2] * x[q][2];
    }
  };
  int t_next = tfirst[0], t_next2 = nrint > 1 ? tfirst[1] : 0;
  for (int i = 0; i < nrint; ++i) {  // window row r = band row i - 1 + r
    const double* o = tri + (size_t)t_next * 22;  // interval i's triangle (reduced onto its first task)
    t_next = t_next2;
    if (i + 2 < nrint) {
      t_next2 = tfirst[i + 2];
      const double* pf = tri + (size_t)t_next2 * 22;
      prefetch_l1(pf);
      prefetch_l1(pf + 16);
    }
    if (PEN) {
      const double* dr = D + (size_t)max(i - 1, 0) * 5;
      prefetch_l1(dr + 40);
      wave_in(o, dr, i >= 1 ? pinv : 0.0, std::true_type{});
    } else {
      wave_in(o, nullptr, 0.0, std::false_type{});
    }
    if (i >= 1) {
#pragma unroll
      for (int e = 0; e < 5; ++e) R[(size_t)(i - 1) * 5 + e] = Rw[0][e];
#pragma unroll
      for (int d = 0; d < 3; ++d) z[(size_t)(i - 1) * 3 + d] = zw[0][d];
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int e = 0; e < 5; ++e) Rw[r][e] = Rw[r + 1][e];
#pragma unroll
      for (int d = 0; d < 3; ++d) zw[r][d] = zw[r + 1][d];
    }
#pragma unroll
    for (int e = 0; e < 5; ++e) Rw[4][e] = 0.0;
#pragma unroll
    for (int d = 0; d < 3; ++d) zw[4][d] = 0.0;
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {  // rows nrint-1 .. nk1-1 are still in the window
#pragma unroll
    for (int e = 0; e < 5; ++e) R[(size_t)(nrint - 1 + r) * 5 + e] = Rw[r][e];
#pragma unroll
    for (int d = 0; d < 3; ++d) z[(size_t)(nrint - 1 + r) * 3 + d] = zw[r][d];
  }
  if (rem_out) *rem_out = rem;
  if (DIAG_ONLY) return;
  // fpback: c_i = (z_i - sum_e R[i][e] c_{i+e}) / R[i][0]; the newest coefficient enters last (shortest chain).
  // Entries of R past the last column were never touched and are zero.
  double c1[3] = {0.0, 0.0, 0.0}, c2[3] = {0.0, 0.0, 0.0}, c3[3] = {0.0, 0.0, 0.0}, c4[3] = {0.0, 0.0, 0.0};
  for (int i = nk1 - 1; i >= 0; --i) {
    if (i >= 24) {
      prefetch_l1(R + (size_t)(i - 24) * 5);
      prefetch_l1(z + (size_t)(i - 24) * 3);
    }
    const double r0 = R[(size_t)i * 5], r1 = R[(size_t)i * 5 + 1], r2 = R[(size_t)i * 5 + 2], r3 = R[(size_t)i * 5 + 3],
                 r4 = R[(size_t)i * 5 + 4];
    const double inv = 1.0 / r0;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      double a = z[(size_t)i * 3 + d];
      a = fma(-c4[d], r4, a);
      a = fma(-c3[d], r3, a);
      a = fma(-c2[d], r2, a);
      a = fma(-c1[d], r1, a) * inv;
      c4[d] = c3[d];
      c3[d] = c2[d];
      c2[d] = c1[d];
      c1[d] = a;
      c[(size_t)i * 3 + d] = a;
    }
  }
}

// ---- the same least-squares solve by block cyclic reduction (all threads of the CTA) -------------------------
// The sweep above is one dependent chain of about 15 rotations per knot interval.  Cut the columns into blocks of
// four: an interval's triangle (4 columns) and a penalty row (5 columns) touch at most two neighbouring blocks, so
// the problem is block bidiagonal, "group" g coupling blocks g and g+1.  A group is kept as an 8x8 triangle with
// its right-hand side.  One level eliminates every other block: the two groups around block b (columns ordered
// [c_b | c_left | c_right]) are rotated into four pivot rows, which are kept for the back substitution, and the
// remainders are compressed into the 8x8 triangle of the group that now couples b's neighbours.  log2(blocks)
// levels, then one 8x8 triangular solve and the back substitution down the levels.  Every transformation is a plane
// rotation, so the solution is the least-squares solution to rounding, like the sweep's; the triangular factor is a
// different one (other column order), which is why sum(diag R) for FITPACK's first p still comes from the sweep.
// Blocks are padded to 2^L + 1 with identity rows.  Prototype with the same indexing: tools/proto/cr_banded_lsq.py.
struct CrWs {
  double *Ga, *Gb, *E, *cpad, *scr;
};
constexpr int KN_CR_MIN = 48;  // intervals from which the cyclic reduction replaces the one-thread sweep
__host__ __device__ inline int knot_group_cap(int cap) { return cap / 2 + 8; }
__device__ __forceinline__ constexpr int tri8(int j, int e) { return j * 8 - j * (j - 1) / 2 + (e - j); }

__device__ __forceinline__ void rotate_in_tri8(double (&R)[36], double (&Z)[8][3], double (&h)[8], double (&x)[3]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (fabs(h[j]) > 1e-150) {
      double cs, sn;
      givens_fast(h[j], R[tri8(j, j)], cs, sn);
#pragma unroll
      for (int e = j + 1; e < 8; ++e) rota(cs, sn, h[e], R[tri8(j, e)]);
#pragma unroll
      for (int d = 0; d < 3; ++d) rota(cs, sn, x[d], Z[j][d]);
    }
  }
}

// Returns this thread's share of the sum of squares of the right-hand-side remainders (see band_merge_solve).
template <bool PEN>
__device__ __noinline__ double cr_solve(const double* tri, const int* tfirst, const double* D, double pinv, int nrint,
                                        CrWs cw, double* c) {
  const int tid = threadIdx.x;
  double rem = 0.0;
  const int nk1 = nrint + 3;
  int L = 1;
  while ((4 << L) < nrint) ++L;  // every interval index below 4 * 2^L, so each interval has its group; 2^L + 1 >= NB follows
  const int ng0 = 1 << L, nbp = ng0 + 1;
  // ---- level 0: groups from the interval triangles, the penalty rows and the padding ----
  for (int g = tid; g < ng0; g += KN_NT) {
    double R[36], Z[8][3];
#pragma unroll
    for (int q = 0; q < 36; ++q) R[q] = 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j) Z[j][0] = Z[j][1] = Z[j][2] = 0.0;
    for (int t = 0; t < 4; ++t) {
      const int i = 4 * g + t;  // local columns t.. of the group
      if (i >= nrint) break;
      const double* o = tri + (size_t)tfirst[i] * 22;
      const double* dr = D + (size_t)i * 5;
      const int nrow = (PEN && i < nrint - 1) ? 5 : 4;
      for (int r = 0; r < nrow; ++r) {
        double h[8], x[3];
        if (r < 4) {
          const int base = r * 4 - r * (r - 1) / 2 - r;  // packed index of entry (r, e) of the 4x4 triangle = base + e
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int e = k - t;
            h[k] = (e >= r && e <= 3) ? o[base + e] : 0.0;
          }
#pragma unroll
          for (int d = 0; d < 3; ++d) x[d] = o[10 + 3 * r + d];
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int e = k - t;
            h[k] = (e >= 0 && e <= 4) ? dr[e] * pinv : 0.0;
          }
          x[0] = x[1] = x[2] = 0.0;
        }
        rotate_in_tri8(R, Z, h, x);
        rem += x[0] * x[0] + x[1] * x[1] + x[2] * x[2];
      }
    }
#pragma unroll
    for (int lc = 4; lc < 8; ++lc)
      if (4 * g + lc >= nk1) R[tri8(lc, lc)] = 1.0;  // padding column: no row touches it, identity row
#pragma unroll
    for (int q = 0; q < 36; ++q) cw.Ga[(size_t)q * ng0 + g] = R[q];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int d = 0; d < 3; ++d) cw.Ga[(size_t)(36 + 3 * j + d) * ng0 + g] = Z[j][d];
  }
  __syncthreads();
  // ---- levels: eliminate block (2j+1) * 2^lvl between groups 2j and 2j+1 ----
  double* Gin = cw.Ga;
  double* Gout = cw.Gb;
  int ngin = ng0;
  for (int lvl = 0; lvl < L; ++lvl) {
    const int ngout = ngin >> 1;
    for (int j = tid; j < ngout; j += KN_NT) {
      double P[4][12], PZ[4][3];
#pragma unroll
      for (int a = 0; a < 4; ++a) {
#pragma unroll
        for (int e = 0; e < 12; ++e) P[a][e] = 0.0;
        PZ[a][0] = PZ[a][1] = PZ[a][2] = 0.0;
      }
      for (int k = 0; k < 16; ++k) {  // rows of the left group (columns c_left | c_b), then of the right one (c_b | c_right)
        const int gsel = 2 * j + (k >> 3), kk = k & 7;
        const int base = kk * 8 - kk * (kk - 1) / 2 - kk;  // tri8(kk, e) = base + e
        double r8[8], h[12], x[3];
#pragma unroll
        for (int e = 0; e < 8; ++e) r8[e] = e >= kk ? Gin[(size_t)(base + e) * ngin + gsel] : 0.0;
#pragma unroll
        for (int d = 0; d < 3; ++d) x[d] = Gin[(size_t)(36 + 3 * kk + d) * ngin + gsel];
        const bool left = k < 8;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          h[e] = left ? r8[4 + e] : r8[e];
          h[4 + e] = left ? r8[e] : 0.0;
          h[8 + e] = left ? 0.0 : r8[4 + e];
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          if (fabs(h[a]) > 1e-150) {
            double cs, sn;
            givens_fast(h[a], P[a][a], cs, sn);
#pragma unroll
            for (int e = a + 1; e < 12; ++e) rota(cs, sn, h[e], P[a][e]);
#pragma unroll
            for (int d = 0; d < 3; ++d) rota(cs, sn, x[d], PZ[a][d]);
          }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) cw.scr[(size_t)(k * 11 + e) * KN_NT + tid] = h[4 + e];
#pragma unroll
        for (int d = 0; d < 3; ++d) cw.scr[(size_t)(k * 11 + 8 + d) * KN_NT + tid] = x[d];
      }
      const int b = (2 * j + 1) << lvl;
      {
        int q = 0;
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int e = 0; e < 12; ++e)
            if (e >= a) cw.E[(size_t)(q++) * nbp + b] = P[a][e];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int d = 0; d < 3; ++d) cw.E[(size_t)(42 + 3 * a + d) * nbp + b] = PZ[a][d];
      }
      double R[36], Z[8][3];
#pragma unroll
      for (int q = 0; q < 36; ++q) R[q] = 0.0;
#pragma unroll
      for (int a = 0; a < 8; ++a) Z[a][0] = Z[a][1] = Z[a][2] = 0.0;
      for (int k = 0; k < 16; ++k) {
        double h[8], x[3];
#pragma unroll
        for (int e = 0; e < 8; ++e) h[e] = cw.scr[(size_t)(k * 11 + e) * KN_NT + tid];
#pragma unroll
        for (int d = 0; d < 3; ++d) x[d] = cw.scr[(size_t)(k * 11 + 8 + d) * KN_NT + tid];
        rotate_in_tri8(R, Z, h, x);
        rem += x[0] * x[0] + x[1] * x[1] + x[2] * x[2];
      }
#pragma unroll
      for (int q = 0; q < 36; ++q) Gout[(size_t)q * ngout + j] = R[q];
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int d = 0; d < 3; ++d) Gout[(size_t)(36 + 3 * a + d) * ngout + j] = Z[a][d];
    }
    __syncthreads();
    double* sw = Gin;
    Gin = Gout;
    Gout = sw;
    ngin = ngout;
  }
  // ---- the last group couples block 0 and block 2^L: 8x8 triangular solve ----
  double* cp = cw.cpad;  // [4 * nbp][3]
  if (tid == 0) {
    double top[8][3];
#pragma unroll
    for (int j = 7; j >= 0; --j) {
      const double dg = Gin[tri8(j, j)];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        double a = Gin[36 + 3 * j + d];
#pragma unroll
        for (int e = j + 1; e < 8; ++e) a -= Gin[tri8(j, e)] * top[e][d];
        top[j][d] = dg != 0.0 ? a / dg : 0.0;
        const int col = j < 4 ? j : 4 * ng0 + (j - 4);
        cp[(size_t)col * 3 + d] = top[j][d];
      }
    }
  }
  __syncthreads();
  // ---- back substitution down the levels ----
  for (int lvl = L - 1; lvl >= 0; --lvl) {
    const int cnt = 1 << (L - lvl - 1);
    for (int j = tid; j < cnt; j += KN_NT) {
      const int b = (2 * j + 1) << lvl, bl = b - (1 << lvl), br = b + (1 << lvl);
      double side[8][3], cb[4][3];
#pragma unroll
      for (int e = 0; e < 4; ++e)
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          side[e][d] = cp[(size_t)(4 * bl + e) * 3 + d];
          side[4 + e][d] = cp[(size_t)(4 * br + e) * 3 + d];
        }
      int q = 42;  // packed index walks backwards: row 3 (9 entries), row 2 (10), row 1 (11), row 0 (12)
#pragma unroll
      for (int a = 3; a >= 0; --a) {
        q -= 12 - a;
        const double dg = cw.E[(size_t)q * nbp + b];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          double acc = cw.E[(size_t)(42 + 3 * a + d) * nbp + b];
#pragma unroll
          for (int e = a + 1; e < 4; ++e) acc -= cw.E[(size_t)(q + e - a) * nbp + b] * cb[e][d];
#pragma unroll
          for (int e = 0; e < 8; ++e) acc -= cw.E[(size_t)(q + 4 - a + e) * nbp + b] * side[e][d];
          cb[a][d] = dg != 0.0 ? acc / dg : 0.0;
        }
      }
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int d = 0; d < 3; ++d) cp[(size_t)(4 * b + a) * 3 + d] = cb[a][d];
    }
    __syncthreads();
  }
  for (int i = tid; i < nk1 * 3; i += KN_NT) c[i] = cp[i];
  __syncthreads();
  return rem;
}

__device__ __forceinline__ double block_sum1(double v, double* s_red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  double a = 0.0;
  for (int w = 0; w < KN_NT / 32; ++w) a += s_red[w];
  return a;
}

__global__ void __launch_bounds__(KN_THREADS) k_fit_knots(const Params p, const KnotWs base, int cap, int debug) {
  __shared__ double s_red[KN_THREADS / 32];
  __shared__ int s_scan[KN_THREADS / 32], s_scan2[KN_THREADS / 32];
  __shared__ int s_i[8];
  __shared__ unsigned long long s_best[KN_THREADS / 32];
  __shared__ int s_besti[KN_THREADS / 32];
  __shared__ unsigned long long s_knot_key[2][KN_THREADS / 32], s_knot_key2[2][KN_THREADS / 32];
  __shared__ int4 s_knot_meta[2][KN_THREADS / 32];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (p.out.status[b] != RDP_FRAME_NEEDS_KNOTS) return;
  const int m = p.fs[b].n_edge;
  const double* X = p.edge_xyz + 3 * (size_t)b * p.ecap;
  double* U = p.edge_u + (size_t)b * p.ecap;
  KnotWs w;
  w.t = base.t + (size_t)b * cap;
  w.q = base.q + (size_t)b * p.ecap * 4;
  w.term = base.term + (size_t)b * p.ecap;
  w.fpint = base.fpint + (size_t)b * cap;
  const int tcap = knot_task_cap(cap);
  w.tri = base.tri + (size_t)b * tcap * 22;
  w.tbeg = base.tbeg + (size_t)b * tcap;
  w.tint = base.tint + (size_t)b * tcap;
  w.tfirst = base.tfirst + (size_t)b * cap;
  const size_t gcap = (size_t)knot_group_cap(cap);
  CrWs cw;
  cw.Ga = base.cr_ga + (size_t)b * gcap * 60;
  cw.Gb = base.cr_gb + (size_t)b * (gcap / 2 + 4) * 60;
  cw.E = base.cr_e + (size_t)b * (gcap + 1) * 54;
  cw.cpad = base.cr_c + (size_t)b * (gcap + 1) * 12;
  cw.scr = base.cr_scr + (size_t)b * 16 * 11 * KN_THREADS;
  w.R = base.R + (size_t)b * cap * 5;
  w.z = base.z + (size_t)b * cap * 3;
  w.c = base.c + (size_t)b * cap * 3;
  w.D = base.D + (size_t)b * cap * 5;
  w.lidx = base.lidx + (size_t)b * p.ecap;
  w.ibegin = base.ibegin + (size_t)b * cap;
  w.nrd = base.nrd + (size_t)b * cap;
  w.nbeg = base.nbeg + (size_t)b * cap;
  w.flag = base.flag + (size_t)b * p.ecap;

  const double s = SMOOTH_S, acc = FIT_TOL * SMOOTH_S;
  const int nmax = m + 4, nest = m + 6;
  {  // normalise the chord-length parameter in place (parcur): u[0] = 0, u[m-1] = 1
    const double total = U[m - 1];
    __syncthreads();
    for (int i = tid; i < m; i += KN_NT) {
      U[i] = i == 0 ? 0.0 : (i == m - 1 ? 1.0 : U[i] / total);
      w.flag[i] = 0;
    }
    __syncthreads();
  }

  int n = 8, nrint = 1, nk1 = 4, nplus = 0, ier = 0, ntasks = 0, max_tasks = 1;
  long long tm_tri = 0, tm_merge = 0, tm_res = 0, tm_setup = 0, tm_knot = 0, tm_mark = 0;  // RDP_DEBUG_KNOTS
  int n_solve = 0, n_iter = 0;
  bool tri_valid = false;        // w.tri holds the per-interval triangles of the current knots
  bool band_R_valid = false;     // w.R holds the banded triangle of the last solve (the sweep's, not the cyclic reduction's)
  const bool no_cr = (debug & 4) != 0;  // RDP_DEBUG_KNOTS=4: one-thread sweep for every solve (A/B timing)
  auto lap = [&](long long& acc) {
    if (debug) {
      const long long now = clock64();
      acc += now - tm_mark;
      tm_mark = now;
    }
  };
  double fp = 0.0, fp0 = 0.0, fpold = 0.0, fpms = 0.0;
  bool smoothing = false, finished = false;
  int fit_flags = 0;  // RDP_FIT_FLAG_*: decisions of this fit that rounding noise could have taken the other way

  // ---- one least-squares (pinv = 0) or smoothing (pinv > 0) solve on the current knots: fills w.c, returns fp ----
  auto solve = [&](double pinv) -> double {
    lap(tm_setup);
    ++n_solve;
    double rem_acc = 0.0;  // this thread's share of sum(remainder^2) over every right-hand side rotated in below
    // (1) task triangles, one thread per task (at most KN_CH consecutive points of one interval).  They depend on the
    // knots only, not on the smoothing parameter: the root find reuses those of the last least-squares fit.
    if (!tri_valid) {
    for (int k = tid; k < ntasks; k += KN_NT) {
      double T[4][4] = {}, Z[4][3] = {};
      const int p0 = w.tbeg[k], p1 = k + 1 < ntasks ? w.tbeg[k + 1] : m;
      for (int it = p0; it < p1; ++it) {
        double h[4] = {w.q[4 * it], w.q[4 * it + 1], w.q[4 * it + 2], w.q[4 * it + 3]};
        double x[3] = {X[3 * it], X[3 * it + 1], X[3 * it + 2]};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          double cs, sn;
          givens_fast(h[r], T[r][r], cs, sn);
#pragma unroll
          for (int d = 0; d < 3; ++d) rota(cs, sn, x[d], Z[r][d]);
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e > r) rota(cs, sn, h[e], T[r][e]);
        }
        rem_acc += x[0] * x[0] + x[1] * x[1] + x[2] * x[2];
      }
      double* o = w.tri + (size_t)k * 22;
      int q = 0;
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (e >= r) o[q++] = T[r][e];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int d = 0; d < 3; ++d) o[q++] = Z[r][d];
    }
    __syncthreads();
    // (1b) the triangles of one interval share their four columns: pairwise tree reduction down to one per interval
    for (int r = 1; r < max_tasks; r <<= 1) {
      for (int k = tid; k < ntasks; k += KN_NT) {
        const int i = w.tint[k], j = k - w.tfirst[i];
        if ((j & (2 * r - 1)) != 0 || k + r >= w.tfirst[i + 1]) continue;
        double* a = w.tri + (size_t)k * 22;
        const double* o = w.tri + (size_t)(k + r) * 22;
        double T[4][4], Z[4][3];
        {
          int q = 0;
#pragma unroll
          for (int rr = 0; rr < 4; ++rr)
#pragma unroll
            for (int e = 0; e < 4; ++e) T[rr][e] = e >= rr ? a[q++] : 0.0;
#pragma unroll
          for (int rr = 0; rr < 4; ++rr)
#pragma unroll
            for (int d = 0; d < 3; ++d) Z[rr][d] = a[q++];
        }
        int q = 0;
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {  // row rr of the other triangle: columns rr..3
          double h[4], x[3];
#pragma unroll
          for (int e = 0; e < 4; ++e) h[e] = e >= rr ? o[q++] : 0.0;
#pragma unroll
          for (int d = 0; d < 3; ++d) x[d] = o[10 + rr * 3 + d];
#pragma unroll
          for (int pc = 0; pc < 4; ++pc) {
            if (pc < rr) continue;
            double cs, sn;
            givens_fast(h[pc], T[pc][pc], cs, sn);
#pragma unroll
            for (int d = 0; d < 3; ++d) rota(cs, sn, x[d], Z[pc][d]);
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (e > pc) rota(cs, sn, h[e], T[pc][e]);
          }
          rem_acc += x[0] * x[0] + x[1] * x[1] + x[2] * x[2];
        }
        q = 0;
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e >= rr) a[q++] = T[rr][e];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
#pragma unroll
          for (int d = 0; d < 3; ++d) a[q++] = Z[rr][d];
      }
      __syncthreads();
    }
    tri_valid = true;
    }
    lap(tm_tri);
    // (2) merge into the band, penalty rows interleaved by last column; back substitution
    if (nrint >= KN_CR_MIN && !no_cr) {
      if (pinv > 0.0) cr_solve<true>(w.tri, w.tfirst, w.D, pinv, nrint, cw, w.c);
      else rem_acc += cr_solve<false>(w.tri, w.tfirst, w.D, 0.0, nrint, cw, w.c);
      band_R_valid = false;
    } else {
      if (tid == 0) {
        double rem_sweep = 0.0;
        if (pinv > 0.0) band_merge_solve<true>(w.tri, w.tfirst, w.D, pinv, nrint, w.R, w.z, w.c);
        else band_merge_solve<false>(w.tri, w.tfirst, w.D, 0.0, nrint, w.R, w.z, w.c, &rem_sweep);
        rem_acc += rem_sweep;
      }
      band_R_valid = true;
      __syncthreads();
    }
    lap(tm_merge);
    // (3) residuals
    double part = 0.0;
    for (int it = tid; it < m; it += KN_NT) {
      const int l0 = w.lidx[it] - 3;
      double tt = 0.0;
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        double fac = 0.0;
#pragma unroll
        for (int j = 0; j < 4; ++j) fac += w.c[(l0 + j) * 3 + d] * w.q[4 * it + j];
        const double e = fac - X[3 * it + d];
        tt += e * e;
      }
      w.term[it] = tt;
      part += tt;
    }
    const double total = block_sum1(part, s_red);
    lap(tm_res);
    // Least-squares stage: FITPACK's fp is what the rotations leave of the right-hand sides (fppara adds the squared
    // remainder of every data row it has rotated in), not the residual of the evaluated spline.  The two agree unless
    // the triangle is numerically singular (more B-splines than data near an end: pivots of 1e-23 seen) - then the
    // coefficients are meaningless, the evaluated residual explodes, and FITPACK carries on with the rotation-based
    // value.  Orthogonal transformations preserve it, so the sum over the tasks, the trees and the solver is the same
    // number.  The smoothing stage evaluates the spline, here as there.
    if (pinv == 0.0) return block_sum1(rem_acc, s_red);
    return total;
  };

  if (debug) tm_mark = clock64();
  for (int iter = 0; iter < m && !finished && !smoothing; ++iter) {
    // ---- knot vector, interval of every point and interval starts from the flags (block scan over the points) ----
    tri_valid = false;
    int carry = 0, carry2 = 0;
    for (int i0 = 0; i0 < m; i0 += KN_NT) {
      const int i = i0 + tid;
      const int f = (i < m) ? w.flag[i] : 0;
      const int cut = (i < m) && (f || (i % KN_CH) == 0);  // a triangle task starts here
      int incl = f, incl2 = cut;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o), v2 = __shfl_up_sync(0xffffffffu, incl2, o);
        if (lane >= o) {
          incl += v;
          incl2 += v2;
        }
      }
      if (lane == 31) {
        s_scan[warp] = incl;
        s_scan2[warp] = incl2;
      }
      __syncthreads();
      int woff = 0, woff2 = 0;
      for (int k = 0; k < warp; ++k) {
        woff += s_scan[k];
        woff2 += s_scan2[k];
      }
      int tot = 0, tot2 = 0;
      for (int k = 0; k < KN_NT / 32; ++k) {
        tot += s_scan[k];
        tot2 += s_scan2[k];
      }
      const int cnt = carry + woff + incl;  // knots at or before this point = its interval
      if (i < m) {
        w.lidx[i] = 3 + cnt;
        if (f) {
          w.ibegin[cnt] = i;
          w.t[3 + cnt] = U[i];
        }
        if (cut) {
          const int k = carry2 + woff2 + incl2 - 1;
          w.tbeg[k] = i;
          w.tint[k] = cnt;
          if (f || i == 0) w.tfirst[cnt] = k;  // the first task of interval cnt
        }
      }
      carry += tot;
      carry2 += tot2;
      __syncthreads();
    }
    ntasks = carry2;
    if (tid == 0) {
      w.tfirst[carry + 1] = ntasks;  // sentinel behind the last interval
      s_i[1] = 1;
    }
    __syncthreads();
    for (int i = tid; i < carry + 1; i += KN_NT) {  // most tasks any interval has (block-uniform loop bound below)
      const int c = w.tfirst[i + 1] - w.tfirst[i];
      if (c > 1) atomicMax(&s_i[1], c);
    }
    __syncthreads();
    max_tasks = s_i[1];
    n = 8 + carry;
    nrint = n - 7;
    nk1 = n - 4;
    if (tid < 4) {
      w.t[tid] = 0.0;
      w.t[n - 4 + tid] = 1.0;
    }
    if (tid == 0) w.ibegin[0] = 0;
    __syncthreads();
    // the last data point (u = 1) belongs to the last interval (FITPACK: l stops at nk1)
    if (tid == 0) w.lidx[m - 1] = nk1 - 1;
    for (int it = tid; it < m; it += KN_NT) {
      if (it == m - 1) continue;
      double h[4];
      bspl3(w.t, U[it], w.lidx[it], h);
      w.q[4 * it] = h[0]; w.q[4 * it + 1] = h[1]; w.q[4 * it + 2] = h[2]; w.q[4 * it + 3] = h[3];
    }
    __syncthreads();
    if (tid == 0) {
      double h[4];
      bspl3(w.t, U[m - 1], nk1 - 1, h);
      w.q[4 * (m - 1)] = h[0]; w.q[4 * (m - 1) + 1] = h[1]; w.q[4 * (m - 1) + 2] = h[2]; w.q[4 * (m - 1) + 3] = h[3];
    }
    __syncthreads();

    fp = solve(0.0);
    if ((debug & 2) && tid == 0) {
      double sd = 0.0, tk = 0.0, dmin = 1e300, ck = 0.0;
      for (int i = 0; i < n; ++i) tk += w.t[i] * (double)(i + 1);
      for (int i = 0; i < nk1 * 3; ++i) ck += w.c[i] * (double)(1 + i % 5);
      if (band_R_valid) {
        for (int i = 0; i < nk1; ++i) {
          sd += w.R[i * 5];
          dmin = fmin(dmin, w.R[i * 5]);
        }
        printf("  frame %d lsq n %d fp %.17g sumdiag %.17g knot checksum %.17g min diag %.3g coefficient checksum %.17g\n", b, n, fp, sd, tk, dmin, ck);
      } else {
        printf("  frame %d lsq n %d fp %.17g (cyclic reduction) knot checksum %.17g coefficient checksum %.17g\n", b, n, fp, tk, ck);
      }
    }
    if (n == 8) {
      ier = -2;
      fp0 = fp;
    }
    fpms = fp - s;
    if (fabs(fpms) < acc) { finished = true; break; }
    if (fpms < 0.0) {
      if (ier == -2) finished = true;
      else smoothing = true;
      break;
    }
    if (n == nmax) { ier = -1; finished = true; break; }
    if (n == nest) { ier = 1; finished = true; break; }
    if (ier == 0) {
      int npl1 = nplus * 2;
      if (fpold - fp > acc) {
        const double raw = (double)nplus * fpms / (fpold - fp);
        npl1 = (int)raw;
        if (fabs(raw - rint(raw)) <= 1e-9 * fabs(raw)) fit_flags |= RDP_FIT_FLAG_NPLUS_TIE;  // truncation on an integer's edge
      }
      nplus = min(nplus * 2, max(npl1, max(nplus / 2, 1)));
    } else {
      nplus = 1;
      ier = 0;
    }
    fpold = fp;
    ++n_iter;
    // ---- residual per interval; the point sitting on a knot is shared half/half by its two intervals ----
    for (int i = tid; i < nrint; i += KN_NT) {
      const int p0 = w.ibegin[i], p1 = i + 1 < nrint ? w.ibegin[i + 1] : m;
      double a = i == 0 ? w.term[p0] : 0.5 * w.term[p0];
      for (int it = p0 + 1; it < p1; ++it) a += w.term[it];
      if (i + 1 < nrint) a += 0.5 * w.term[p1];
      w.fpint[i] = a;
      w.nrd[i] = (i + 1 < nrint ? p1 : m - 1) - p0 - 1;
      w.nbeg[i] = p0 + 1;
    }
    __syncthreads();
    // ---- fpknot x nplus: split the interval with the largest residual that still has interior points ----
    // Node i (residual, interior point count, first interior point) belongs to thread i mod KN_THREADS: only its owner
    // reads or writes it, so the nodes need no barrier.  Every thread keeps the best of its own nodes in registers;
    // a split changes two nodes, and only their owners look through theirs again.  Per knot: one shuffle reduction,
    // one barrier (the per-warp results are double-buffered), and every thread derives the same winner.
    int nn = nrint;
    unsigned long long my_key = 0;  // residual as an integer (positive doubles order as their bit patterns); 0 = nothing
    int my_beg = 0x7fffffff, my_nrd = 0, my_idx = -1;
    // The runner-up (key and point count only) rides along: when it is within KN_NEAR_TIE of the winner without being
    // its bit-identical twin, the choice of interval hangs on rounding noise and the frame is flagged (fit_flags).
    unsigned long long my_key2 = 0;
    int my_nrd2 = 0;
    auto rescan = [&]() {
      my_key = my_key2 = 0;
      my_beg = 0x7fffffff;
      my_nrd = my_nrd2 = 0;
      my_idx = -1;
      for (int i = tid; i < nn; i += KN_NT) {
        const int nr = w.nrd[i];
        if (nr == 0) continue;
        const double v = w.fpint[i];
        if (!(v > 0.0)) continue;  // FITPACK starts from fpmax = 0 with a strict comparison
        const unsigned long long key = (unsigned long long)__double_as_longlong(v);
        const int nb = w.nbeg[i];
        if (key > my_key || (key == my_key && nb < my_beg)) {  // larger residual first, earlier interval on ties
          my_key2 = my_key;
          my_nrd2 = my_nrd;
          my_key = key;
          my_beg = nb;
          my_nrd = nr;
          my_idx = i;
        } else if (key > my_key2) {
          my_key2 = key;
          my_nrd2 = nr;
        }
      }
    };
    rescan();
    for (int l = 0; l < nplus; ++l) {
      unsigned long long bk = my_key, bk2 = my_key2;
      int bb = my_beg, br = my_nrd, bi = my_idx, br2 = my_nrd2;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long ok = __shfl_xor_sync(0xffffffffu, bk, o), ok2 = __shfl_xor_sync(0xffffffffu, bk2, o);
        const int ob = __shfl_xor_sync(0xffffffffu, bb, o), orr = __shfl_xor_sync(0xffffffffu, br, o),
                  oi = __shfl_xor_sync(0xffffffffu, bi, o), orr2 = __shfl_xor_sync(0xffffffffu, br2, o);
        if (ok > bk || (ok == bk && ob < bb)) {  // the other side wins: its runner-up competes with our winner
          if (bk > ok2) { bk2 = bk; br2 = br; } else { bk2 = ok2; br2 = orr2; }
          bk = ok;
          bb = ob;
          br = orr;
          bi = oi;
        } else if (ok > bk2) {
          bk2 = ok;
          br2 = orr;
        }
      }
      const int buf = l & 1;
      if (lane == 0) {
        s_knot_key[buf][warp] = bk;
        s_knot_meta[buf][warp] = make_int4(bb, br, bi, br2);
        s_knot_key2[buf][warp] = bk2;
      }
      __syncthreads();
      bk = bk2 = 0;
      bb = 0x7fffffff;
      br = br2 = 0;
      bi = -1;
#pragma unroll
      for (int k = 0; k < KN_NT / 32; ++k) {
        const unsigned long long ok = s_knot_key[buf][k], ok2 = s_knot_key2[buf][k];
        const int4 om = s_knot_meta[buf][k];
        if (ok > bk || (ok == bk && om.x < bb)) {
          if (bk > ok2) { bk2 = bk; br2 = br; } else { bk2 = ok2; br2 = om.w; }
          bk = ok;
          bb = om.x;
          br = om.y;
          bi = om.z;
        } else if (ok > bk2) {
          bk2 = ok;
          br2 = om.y;
        }
      }
      if (bi < 0) break;  // nothing left to split (block-uniform: every thread derived the same winner)
      // bit-identical residuals with equal point counts come from one expression (the halves of an odd split and their
      // descendants) and resolve by position in every implementation; anything else this close is decided by rounding
      if (bk2 != 0 && bk - bk2 <= KN_NEAR_TIE && !(bk == bk2 && br == br2)) fit_flags |= RDP_FIT_FLAG_KNOT_TIE;
      const double fpmax = __longlong_as_double((long long)bk);
      const int maxpt = br, maxbeg = bb;
      const int ihalf = maxpt / 2 + 1, nrx = maxbeg + ihalf;
      if ((debug & 16) && tid == 0) printf("  frame %d n %d split %d: node %d begin %d points %d residual %.17g -> knot at point %d\n", b, n, l, bi, maxbeg, maxpt, fpmax, nrx - 1);
      const bool own_old = (bi % KN_NT) == tid, own_new = (nn % KN_NT) == tid;
      if (own_old) {
        w.flag[nrx - 1] = 1;
        w.nrd[bi] = ihalf - 1;
        w.fpint[bi] = fpmax * (double)(ihalf - 1) / (double)maxpt;
      }
      if (own_new) {
        w.nrd[nn] = maxpt - ihalf;
        w.nbeg[nn] = nrx;
        w.fpint[nn] = fpmax * (double)(maxpt - ihalf) / (double)maxpt;
      }
      const int stop = (n + 1 == nmax) ? 2 : ((n + 1 == nest) ? 3 : 0);  // nmax: interpolation knots, nest: storage limit
      ++n;
      ++nn;
      if (stop == 2) {  // n == nmax: knots as for interpolation, t[4+i] = u[2+i]
        __syncthreads();  // the flag set above by another thread must not land after the rewrite below
        for (int i = tid; i < m; i += KN_NT) w.flag[i] = (i >= 2 && i <= m - 3) ? 1 : 0;
        break;
      }
      if (stop == 3) break;
      if (own_old || own_new) rescan();
    }
    __syncthreads();
    lap(tm_knot);
  }

  if (smoothing) {
    // ---- Part 2: find p with F(p) = s, F = residual of the spline minimising residual + (1/p) * derivative jumps ----
    for (int i = tid; i < nrint - 1; i += KN_NT) {  // fpdisc
      const double fac = (double)nrint / (w.t[nk1] - w.t[3]);
      const int l = i + 5;  // 1-based knot index k2 + i
      double h[8];
#pragma unroll
      for (int j = 1; j <= 4; ++j) {
        h[j - 1] = w.t[l - 1] - w.t[l + j - 5 - 1];
        h[j + 3] = w.t[l - 1] - w.t[l + j - 1];
      }
      int lp = l - 4;
#pragma unroll
      for (int j = 1; j <= 5; ++j) {
        double prod = h[j - 1];
#pragma unroll
        for (int e = 1; e <= 3; ++e) prod = prod * h[j - 1 + e] * fac;
        w.D[(size_t)i * 5 + (j - 1)] = (w.t[lp + 4 - 1] - w.t[lp - 1]) / prod;
        ++lp;
      }
    }
    if (!band_R_valid) {  // FITPACK's first p needs the diagonal of the banded triangle of the last least-squares fit
      if (tid == 0) band_merge_solve<false, true>(w.tri, w.tfirst, w.D, 0.0, nrint, w.R, w.z, w.c);
      __syncthreads();
    }
    lap(tm_merge);
    double sumdiag = 0.0;
    for (int i = tid; i < nk1; i += KN_NT) sumdiag += w.R[i * 5];
    sumdiag = block_sum1(sumdiag, s_red);
    double p1 = 0.0, f1 = fp0 - s, p3 = -1.0, f3 = fpms, pp = (double)nk1 / sumdiag;
    bool ich1 = false, ich3 = false;
    ier = 0;
    for (int it = 1; it <= KN_MAXIT; ++it) {
      fp = solve(1.0 / pp);
      fpms = fp - s;
      if ((debug & 2) && tid == 0) printf("  frame %d smoothing it %d p %.17g fp %.17g (p1 %.17g p3 %.17g)\n", b, it, pp, fp, p1, p3);
      if (fabs(fpms) < acc) break;
      if (it == KN_MAXIT) { ier = 3; break; }
      const double p2 = pp, f2 = fpms;
      if (!ich3) {
        if ((f2 - f3) <= acc) {
          p3 = p2; f3 = f2;
          pp = pp * 0.04;
          if (p1 > 0.0 && fabs(pp - p1) <= 1e-12 * p1) fit_flags |= RDP_FIT_FLAG_BRACKET_TIE;
          if (pp <= p1 * KN_TIE_UP) pp = p1 * 0.9 + p2 * 0.1;
          continue;
        }
        if (f2 < 0.0) ich3 = true;
      }
      if (!ich1) {
        if ((f1 - f2) <= acc) {
          p1 = p2; f1 = f2;
          pp = pp / 0.04;
          if (p3 < 0.0) continue;
          if (fabs(pp - p3) <= 1e-12 * p3) fit_flags |= RDP_FIT_FLAG_BRACKET_TIE;
          if (pp >= p3 * KN_TIE_DOWN) pp = p2 * 0.1 + p3 * 0.9;
          continue;
        }
        if (f2 > 0.0) ich1 = true;
      }
      if (f2 >= f1 || f2 <= f3) { ier = 2; break; }
      // fprati
      if (p3 > 0.0) {
        const double h1 = f1 * (f2 - f3), h2 = f2 * (f3 - f1), h3 = f3 * (f1 - f2);
        pp = -(p1 * p2 * h3 + p2 * p3 * h1 + p3 * p1 * h2) / (p1 * h1 + p2 * h2 + p3 * h3);
      } else {
        pp = (p1 * (f1 - f3) * f2 - p2 * (f2 - f3) * f1) / ((f1 - f2) * f3);
      }
      if (f2 < 0.0) { p3 = p2; f3 = f2; }
      else { p1 = p2; f1 = f2; }
    }
  }

  // ---- curvature and spline samples from (t, c): splev / splder at 100 parameters ----
  double samp0 = 0.0, samp1 = 0.0;
  unsigned long long kmax_key = 0;
  if (tid < NUM_SAMPLES) {
    const double u = tid == NUM_SAMPLES - 1 ? 1.0 : (double)tid * (1.0 / (double)(NUM_SAMPLES - 1));
    int l = 3;
    while (u >= w.t[l + 1] && l != nk1 - 1) ++l;
    double h3[4];
    bspl3(w.t, u, l, h3);
    // degree-2 and degree-1 bases on the same interval (derivative splines share the knots)
    const double tl = w.t[l], tr = w.t[l + 1];
    double b1[2] = {(tr - u) / (tr - tl), (u - tl) / (tr - tl)};
    double b2[3];
    {
      const double f0 = b1[0] / (w.t[l + 1] - w.t[l - 1]), f1 = b1[1] / (w.t[l + 2] - w.t[l]);
      b2[0] = f0 * (w.t[l + 1] - u);
      b2[1] = f0 * (u - w.t[l - 1]) + f1 * (w.t[l + 2] - u);
      b2[2] = f1 * (u - w.t[l]);
    }
    double pt[3], d1[3], d2[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const double c0 = w.c[(l - 3) * 3 + d], c1 = w.c[(l - 2) * 3 + d], c2 = w.c[(l - 1) * 3 + d], c3 = w.c[l * 3 + d];
      pt[d] = h3[0] * c0 + h3[1] * c1 + h3[2] * c2 + h3[3] * c3;
      // splder: c'_j = 3 (c_{j+1} - c_j) / (t_{j+4} - t_{j+1}), c''_j = 2 (c'_{j+1} - c'_j) / (t_{j+4} - t_{j+2})
      const double e0 = 3.0 * (c1 - c0) / (w.t[l + 1] - w.t[l - 2]);
      const double e1 = 3.0 * (c2 - c1) / (w.t[l + 2] - w.t[l - 1]);
      const double e2 = 3.0 * (c3 - c2) / (w.t[l + 3] - w.t[l]);
      d1[d] = b2[0] * e0 + b2[1] * e1 + b2[2] * e2;
      const double g0 = 2.0 * (e1 - e0) / (w.t[l + 1] - w.t[l - 1]);
      const double g1 = 2.0 * (e2 - e1) / (w.t[l + 2] - w.t[l]);
      d2[d] = b1[0] * g0 + b1[1] * g1;
    }
    const double cx = d1[1] * d2[2] - d1[2] * d2[1], cy = d1[2] * d2[0] - d1[0] * d2[2], cz = d1[0] * d2[1] - d1[1] * d2[0];
    const double n1 = sqrt(d1[0] * d1[0] + d1[1] * d1[1] + d1[2] * d1[2]);
    if (n1 > 1e-6) {
      const double kappa = sqrt(cx * cx + cy * cy + cz * cz) / (n1 * n1 * n1);
      samp0 = kappa;
      samp1 = 1.0;
      kmax_key = ordered_key(kappa);
    }
    double* dst = p.out.spline_points + ((size_t)b * NUM_SAMPLES + tid) * 3;
    dst[0] = pt[0]; dst[1] = pt[1]; dst[2] = pt[2];
  }
  const double ksum = block_sum1(samp0, s_red);
  const double kcnt = block_sum1(samp1, s_red);
  kmax_key = warp_max(kmax_key);
  __syncthreads();
  if (lane == 0) s_best[warp] = kmax_key;
  __syncthreads();
  if (tid == 0) {
    for (int k = 1; k < KN_NT / 32; ++k) kmax_key = s_best[k] > kmax_key ? s_best[k] : kmax_key;
    p.out.mean_curvature[b] = kcnt > 0.0 ? ksum / kcnt : 0.0;
    p.out.max_curvature[b] = kcnt > 0.0 ? key_to_double(kmax_key) : 0.0;
    p.out.status[b] = RDP_FRAME_OK;
    if (debug && ((debug & 2) || b == 0))
      printf("k_fit_knots frame %d: m %d knots %d solves %d knot-iterations %d | kcycles setup %lld triangles %lld merge %lld residual %lld fpknot %lld\n",
             b, m, n, n_solve, n_iter, tm_setup / 1000, tm_tri / 1000, tm_merge / 1000, tm_res / 1000, tm_knot / 1000);
    if (p.out.n_knots) p.out.n_knots[b] = n;
    if (p.out.fit_residual) p.out.fit_residual[b] = fp;
    if (p.out.fit_ier) p.out.fit_ier[b] = ier;
    if (p.out.fit_flags) p.out.fit_flags[b] = fit_flags;
  }
}


struct Plan {
  size_t header_bytes, total_bytes;
  int tiles, ecap;
};

static void carve_knots(Carver& c, KnotWs& k, int B, int ecap) {
  const size_t cap = (size_t)ecap + 8, e = (size_t)ecap;
  k.t = c.take<double>(B * cap);
  k.q = c.take<double>(B * e * 4);
  k.term = c.take<double>(B * e);
  k.fpint = c.take<double>(B * cap);
  const size_t tcap = (size_t)knot_task_cap((int)cap);
  k.tri = c.take<double>(B * tcap * 22);
  k.R = c.take<double>(B * cap * 5);
  k.z = c.take<double>(B * cap * 3);
  k.c = c.take<double>(B * cap * 3);
  k.D = c.take<double>(B * cap * 5);
  k.lidx = c.take<int>(B * e);
  k.ibegin = c.take<int>(B * cap);
  k.nrd = c.take<int>(B * cap);
  k.nbeg = c.take<int>(B * cap);
  k.tbeg = c.take<int>(B * tcap);
  k.tint = c.take<int>(B * tcap);
  k.tfirst = c.take<int>(B * cap);
  k.flag = c.take<unsigned char>(B * e);
  const size_t gcap = (size_t)knot_group_cap((int)cap);
  k.cr_ga = c.take<double>(B * gcap * 60);
  k.cr_gb = c.take<double>(B * (gcap / 2 + 4) * 60);
  k.cr_e = c.take<double>(B * (gcap + 1) * 54);
  k.cr_c = c.take<double>(B * (gcap + 1) * 12);
  k.cr_scr = c.take<double>((size_t)B * 16 * 11 * KN_THREADS);
}

static Plan carve(Params& p, KnotWs& kn, void* ws, int B, int H, int W) {
  const size_t P = (size_t)H * W;
  Plan pl;
  pl.tiles = (int)((P + TILE_PX - 1) / TILE_PX);
  pl.ecap = (int)(P / 20 + 64);
  Carver c(ws);
  p.fs = c.take<FrameState>(B);
  p.tile_state = c.take<unsigned long long>((size_t)B * pl.tiles);
  p.ticket = c.take<unsigned int>(64);
  pl.header_bytes = c.total();  // everything up to here is zeroed before each call
  p.pix = c.take<uint32_t>(B * P);
  p.bins = c.take<uint8_t>(B * P);
  p.binned = c.take<uint32_t>(B * P);
  p.xkey_c = c.take<unsigned long long>(B * P);
  p.ykey_c = c.take<unsigned long long>(B * P);
  p.ykey_b = c.take<unsigned long long>(B * P);
  p.edge_pix = c.take<uint32_t>((size_t)B * pl.ecap);
  p.edge_key = c.take<unsigned long long>((size_t)B * pl.ecap);
  p.edge_xyz = c.take<double>((size_t)B * pl.ecap * 3);
  p.edge_u = c.take<double>((size_t)B * pl.ecap);
  carve_knots(c, kn, B, pl.ecap);
  pl.total_bytes = c.total();
  p.tiles = pl.tiles;
  p.ecap = pl.ecap;
  return pl;
}

}  // namespace geom
}  // namespace rdp

using namespace rdp;
using namespace rdp::geom;

extern "C" int rdp_geom_workspace_bytes(int B, int H, int W, size_t* bytes) {
  RDP_REQUIRE(bytes && B > 0 && H > 0 && W > 0, RDP_ERR_BAD_ARGUMENT, "rdp_geom_workspace_bytes: bad shape %dx%dx%d", B, H, W);
  RDP_REQUIRE((size_t)H * W < (1ull << 31) / 4, RDP_ERR_UNSUPPORTED, "rdp_geom_workspace_bytes: frame too large");
  Params p{};
  KnotWs kn{};
  *bytes = carve(p, kn, nullptr, B, H, W).total_bytes;
  return RDP_OK;
}

extern "C" int rdp_geom_launch_count(void) { return 7; }

static int curvature_impl(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                          const double* intrinsics, double depth_scale, const RdpCurvatureOut* out,
                          void* workspace, size_t workspace_bytes, void* stream_, cudaEvent_t* marks) {
  RDP_REQUIRE(mask && depth && intrinsics && out && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_geom_curvature: null pointer");
  RDP_REQUIRE(B > 0 && H > 0 && W > 0, RDP_ERR_BAD_ARGUMENT, "rdp_geom_curvature: bad shape %dx%dx%d", B, H, W);
  RDP_REQUIRE((size_t)H * W < (1ull << 31) / 4, RDP_ERR_UNSUPPORTED, "rdp_geom_curvature: frame too large");
  RDP_REQUIRE(depth_fmt >= RDP_DEPTH_U16 && depth_fmt <= RDP_DEPTH_F64, RDP_ERR_BAD_ARGUMENT,
              "rdp_geom_curvature: unknown depth format %d", depth_fmt);
  RDP_REQUIRE(out->mean_curvature && out->max_curvature && out->spline_points && out->status, RDP_ERR_BAD_ARGUMENT,
              "rdp_geom_curvature: mean/max/spline_points/status outputs are required");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  Params p{};
  KnotWs kn{};
  const Plan pl = carve(p, kn, workspace, B, H, W);
  RDP_REQUIRE(workspace_bytes >= pl.total_bytes, RDP_ERR_WORKSPACE_TOO_SMALL,
              "rdp_geom_curvature: workspace %zu < required %zu bytes", workspace_bytes, pl.total_bytes);
  p.mask = mask;
  p.depth = depth;
  p.depth_fmt = depth_fmt;
  p.B = B; p.H = H; p.W = W; p.P = H * W;
  p.fx = intrinsics[0]; p.fy = intrinsics[1]; p.cx = intrinsics[2]; p.cy = intrinsics[3];
  p.scale = depth_scale;
  p.scale_f32 = (float)depth_scale;
  p.out = *out;

  int n_marks = 0;
  auto mark = [&]() {
    if (marks) cudaEventRecord(marks[n_marks++], stream);
  };
  mark();
  RDP_CUDA_TRY(cudaMemsetAsync(workspace, 0, pl.header_bytes, stream));
  mark();
  const bool vec = depth_fmt != RDP_DEPTH_F64 && (p.P % 16 == 0) && ((uintptr_t)mask % 16 == 0) && ((uintptr_t)depth % 16 == 0);
  const int tiles_total = B * pl.tiles;
  if (vec) k_deproject<true><<<tiles_total, DP_THREADS, 0, stream>>>(p);
  else k_deproject<false><<<tiles_total, DP_THREADS, 0, stream>>>(p);
  mark();
  const dim3 pgrid((p.P + PT_PER_BLOCK - 1) / PT_PER_BLOCK, B);
  k_bin_hist<<<pgrid, PT_THREADS, 0, stream>>>(p);
  mark();
  k_scatter<<<pgrid, PT_THREADS, 0, stream>>>(p);
  mark();
  k_select_sort_warp<<<(B * NUM_BINS + WS_WARPS - 1) / WS_WARPS, WS_WARPS * 32, 0, stream>>>(p);
  k_select_sort<<<dim3(NUM_BINS, B), SEL_THREADS, 0, stream>>>(p);  // the bins the warp kernel left (large, or heavily tied)
  mark();
  k_fit<<<B, FIT_THREADS, 0, stream>>>(p);
  mark();
  static const int knots_debug = dev_env("RDP_DEBUG_KNOTS") ? atoi(dev_env("RDP_DEBUG_KNOTS")) : 0;  // phase cycle counts
  // One CTA per frame and 255 registers per thread: 256 threads fill an SM's register file.  With more frames than SMs,
  // 128 threads per CTA put two frames on an SM; the parallel phases take longer, the serial chains (most of the time)
  // do not, so the batch finishes in about one CTA lifetime instead of two.
  int n_sm = 148, dev_id = 0;
  RDP_CUDA_TRY(cudaGetDevice(&dev_id));
  RDP_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev_id));
  static const int knots_threads_env = dev_env("RDP_KNOTS_THREADS") ? atoi(dev_env("RDP_KNOTS_THREADS")) : 0;  // 128 or 256 (A/B)
  const int kn_threads = knots_threads_env == 128 || knots_threads_env == 256 ? knots_threads_env : (B > n_sm ? 128 : KN_THREADS);
  k_fit_knots<<<B, kn_threads, 0, stream>>>(p, kn, pl.ecap + 8, knots_debug);  // exits at once for frames that need no extra knots
  mark();
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}

extern "C" int rdp_geom_curvature(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                                  const double* intrinsics, double depth_scale, const RdpCurvatureOut* out,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  return curvature_impl(mask, depth, depth_fmt, B, H, W, intrinsics, depth_scale, out, workspace, workspace_bytes, stream, nullptr);
}

extern "C" int rdp_geom_curvature_profiled(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                                           const double* intrinsics, double depth_scale, const RdpCurvatureOut* out,
                                           void* workspace, size_t workspace_bytes, void* stream, float* launch_ms) {
  RDP_REQUIRE(launch_ms, RDP_ERR_BAD_ARGUMENT, "rdp_geom_curvature_profiled: launch_ms is null");
  const int n = rdp_geom_launch_count();
  cudaEvent_t ev[16];
  for (int i = 0; i <= n; ++i) RDP_CUDA_TRY(cudaEventCreate(&ev[i]));
  int rc = curvature_impl(mask, depth, depth_fmt, B, H, W, intrinsics, depth_scale, out, workspace, workspace_bytes, stream, ev);
  if (rc == RDP_OK && cudaEventSynchronize(ev[n]) != cudaSuccess) {
    set_error("rdp_geom_curvature_profiled: %s", cudaGetErrorString(cudaGetLastError()));
    rc = RDP_ERR_CUDA;
  }
  for (int i = 0; i < n && rc == RDP_OK; ++i) cudaEventElapsedTime(&launch_ms[i], ev[i], ev[i + 1]);
  for (int i = 0; i <= n; ++i) cudaEventDestroy(ev[i]);
  return rc;
}

// ---------------------------------------------------------------------------------------------------------
// rdp_geom_deproject_static: the deprojection stage for callers that must not allocate device memory
// ---------------------------------------------------------------------------------------------------------
// Some hosts cannot hand this library a workspace: a serving process that gave every byte of device memory to
// its own pool at start-up and forbids cudaMalloc on the request path, or a control loop that cannot absorb the
// latency (or the failure) of an allocation.  For them the bookkeeping k_deproject needs - per-frame counters and
// min/max, the look-back descriptors of the ordered compaction, the tile ticket - lives in the fixed array below
// instead of a caller-provided workspace.  The caller passes only inputs and outputs.
//
// The array is part of the library image: every GPU that loads the library carries it until the process exits.
// All calls share it and it keeps what the last call left behind, so each call clears the part it uses first
// (stream-ordered), and two calls that overlap on the GPU would race: issue them on one stream, or otherwise make
// sure one has finished before the next starts.
constexpr int STATIC_MAX_FRAMES = 64;
constexpr int STATIC_MAX_TILES = 4096;  // sum over the batch of ceil(H*W / 16384)
constexpr size_t STATIC_SCRATCH_BYTES =
    ((sizeof(rdp::geom::FrameState) * STATIC_MAX_FRAMES + 255) / 256) * 256 + sizeof(unsigned long long) * STATIC_MAX_TILES + 256;
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

namespace {

__global__ void k_export_counts(const rdp::geom::FrameState* fs, int B, int32_t* n_points) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) n_points[b] = fs[b].n_points;
}

constexpr int STATIC_MAX_DEVICES = 64;

// In host code `g_static_scratch` is only a handle for the runtime, not an address a kernel can use: the device
// address has to be asked for, once per device, with that device current.
int static_scratch_address(void** out) {
  static void* cache[STATIC_MAX_DEVICES] = {};
  static bool known[STATIC_MAX_DEVICES] = {};
  static std::mutex lock;
  int dev = 0;
  RDP_CUDA_TRY(cudaGetDevice(&dev));
  RDP_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, RDP_ERR_UNSUPPORTED, "rdp_geom_deproject_static: device index %d out of range", dev);
  std::lock_guard<std::mutex> guard(lock);
  if (!known[dev]) {
    void* ptr = nullptr;
    RDP_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));
    cache[dev] = ptr;
    known[dev] = true;
  }
  *out = cache[dev];
  return RDP_OK;
}

}  // namespace

extern "C" int rdp_geom_deproject_static(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                                         const double* intrinsics, double depth_scale, float* points_xyz,
                                         uint32_t* pixel_ids, int32_t* n_points, void* stream_) {
  RDP_REQUIRE(mask && depth && intrinsics && points_xyz && pixel_ids && n_points, RDP_ERR_BAD_ARGUMENT,
              "rdp_geom_deproject_static: null pointer");
  RDP_REQUIRE(B > 0 && H > 0 && W > 0, RDP_ERR_BAD_ARGUMENT, "rdp_geom_deproject_static: bad shape %dx%dx%d", B, H, W);
  RDP_REQUIRE((size_t)H * W < (1ull << 31) / 4, RDP_ERR_UNSUPPORTED, "rdp_geom_deproject_static: frame too large");
  RDP_REQUIRE(depth_fmt >= RDP_DEPTH_U16 && depth_fmt <= RDP_DEPTH_F64, RDP_ERR_BAD_ARGUMENT,
              "rdp_geom_deproject_static: unknown depth format %d", depth_fmt);
  const int P = H * W;
  const int tiles = (P + TILE_PX - 1) / TILE_PX;
  RDP_REQUIRE(B <= STATIC_MAX_FRAMES && (long long)B * tiles <= STATIC_MAX_TILES, RDP_ERR_UNSUPPORTED,
              "rdp_geom_deproject_static: batch of %d frames x %d tiles exceeds the built-in scratch (at most %d frames and "
              "%d tiles of 16384 pixels per call); use rdp_geom_curvature with a workspace for larger batches",
              B, tiles, STATIC_MAX_FRAMES, STATIC_MAX_TILES);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  // The scratch address is looked up for the CURRENT device; a stream that belongs to another device would make the
  // memsets below and the kernels disagree about which GPU's array they touch.  Refuse before anything is issued.
  if (stream != nullptr && stream != cudaStreamLegacy && stream != cudaStreamPerThread) {
    int cur = -1, sdev = -1;
    RDP_CUDA_TRY(cudaGetDevice(&cur));
    if (cudaStreamGetDevice(stream, &sdev) == cudaSuccess) {
      RDP_REQUIRE(sdev == cur, RDP_ERR_BAD_ARGUMENT,
                  "rdp_geom_deproject_static: the stream belongs to device %d but device %d is current; make the stream's "
                  "device current before every call", sdev, cur);
    } else {
      (void)cudaGetLastError();  // older drivers: the check is advisory
    }
  }
  void* scratch = nullptr;
  const int rc = static_scratch_address(&scratch);
  if (rc != RDP_OK) return rc;

  Params p{};
  Carver c(scratch);
  p.fs = c.take<FrameState>(STATIC_MAX_FRAMES);
  p.tile_state = c.take<unsigned long long>(STATIC_MAX_TILES);
  p.ticket = c.take<unsigned int>(64);
  // counters, look-back descriptors and the ticket must start at zero; the rest of the array is not read
  RDP_CUDA_TRY(cudaMemsetAsync(p.fs, 0, sizeof(FrameState) * B, stream));
  RDP_CUDA_TRY(cudaMemsetAsync(p.tile_state, 0, sizeof(unsigned long long) * (size_t)B * tiles, stream));
  RDP_CUDA_TRY(cudaMemsetAsync(p.ticket, 0, sizeof(unsigned int), stream));
  p.mask = mask;
  p.depth = depth;
  p.depth_fmt = depth_fmt;
  p.B = B; p.H = H; p.W = W; p.P = P;
  p.tiles = tiles;
  p.fx = intrinsics[0]; p.fy = intrinsics[1]; p.cx = intrinsics[2]; p.cy = intrinsics[3];
  p.scale = depth_scale;
  p.scale_f32 = (float)depth_scale;
  p.pix = pixel_ids;
  p.out.points_xyz = points_xyz;
  const bool vec = depth_fmt != RDP_DEPTH_F64 && (P % 16 == 0) && ((uintptr_t)mask % 16 == 0) && ((uintptr_t)depth % 16 == 0);
  if (vec) k_deproject<true><<<B * tiles, DP_THREADS, 0, stream>>>(p);
  else k_deproject<false><<<B * tiles, DP_THREADS, 0, stream>>>(p);
  k_export_counts<<<(B + 63) / 64, 64, 0, stream>>>(p.fs, B, n_points);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}
