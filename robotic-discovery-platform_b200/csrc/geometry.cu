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
//   k_deproject    HBM-bound   reads 3P (u16) bytes/frame once, writes 4N (pixel ids) + 12N (fp32 xyz)
//   k_bin_hist     L2/HBM      reads 4N + gathers 2N, writes N (bin ids)
//   k_bin_offsets  tiny        50 bins per frame -> k per bin, offsets, status
//   k_scatter      L2/HBM      reads 5N, writes 4N (pixel ids grouped by bin)
//   k_select_sort  latency     one CTA per (frame, bin): radix select of the k largest y, bitonic sort by x
//   k_fit          latency     one CTA per frame: chord-length scan, 4x4 normal equations in fp64, curvature
#include <stdarg.h>

#include <mutex>

#include "rdp_common.cuh"

namespace rdp {
namespace geom {

constexpr int NUM_BINS = 50;      // geometry_utils.py:119
constexpr int MIN_POINTS = 100;   // :64
constexpr int MIN_EDGE = 20;      // :69
constexpr int NUM_SAMPLES = 100;  // :83,144
constexpr double SMOOTH_S = 0.1;  // :78
constexpr double FIT_TOL = 1e-3;  // FITPACK parcur tol

constexpr int DP_THREADS = 256;
constexpr int DP_PX_PER_THREAD = 16;
constexpr int TILE_PX = DP_THREADS * DP_PX_PER_THREAD;  // 4096 pixels per deprojection tile
constexpr int PT_THREADS = 256;
constexpr int PT_PER_BLOCK = 2048;  // points per block in the histogram / scatter passes
constexpr int SEL_THREADS = 256;
constexpr int STAGE_CAP = 6144;  // bin segment staged in shared memory up to this many points
constexpr int SORT_CAP = 2048;   // selected points sorted in shared memory up to this many

struct FrameState {
  unsigned long long xmax_key;      // atomicMax over ordered keys of x
  unsigned long long xmin_inv_key;  // atomicMax over ~key, so that a zeroed struct is the identity
  int n_points;
  int n_edge;
  int status;
  int pad_;
  int bin_count[NUM_BINS];
  int bin_k[NUM_BINS];
  int bin_off[NUM_BINS + 1];
  int edge_off[NUM_BINS + 1];
  int bin_cursor[NUM_BINS];
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
  __shared__ uint32_t s_pix[TILE_PX];
  __shared__ int s_warp[DP_THREADS / 32];
  __shared__ unsigned long long s_red[2][DP_THREADS / 32];
  __shared__ int s_ticket, s_base, s_total;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_ticket = (int)atomicAdd(p.ticket, 1u);  // tiles are claimed in launch order: no look-back deadlock
  __syncthreads();
  const int b = s_ticket / p.tiles, t = s_ticket % p.tiles;
  const size_t fbase = (size_t)b * p.P;
  const int px0 = t * TILE_PX + tid * DP_PX_PER_THREAD;

  unsigned valid = 0;
  if (VEC && px0 + DP_PX_PER_THREAD <= p.P) {
    // 128-bit loads: 16 mask bytes, then the 16 depths
    const uint4 m4 = __ldg(reinterpret_cast<const uint4*>(p.mask + fbase + px0));
    const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
    if (p.depth_fmt == RDP_DEPTH_U16) {
      const uint4* dp = reinterpret_cast<const uint4*>((const uint16_t*)p.depth + fbase + px0);
      const uint4 d0 = __ldg(dp), d1 = __ldg(dp + 1);
      const uint32_t dw[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t mb = (mw[i >> 2] >> (8 * (i & 3))) & 0xffu;
        const uint32_t dv = (dw[i >> 1] >> (16 * (i & 1))) & 0xffffu;
        if (mb && __dmul_rn((double)dv, p.scale) > 0.0) valid |= 1u << i;
      }
    } else {
      const float4* dp = reinterpret_cast<const float4*>((const float*)p.depth + fbase + px0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 d = __ldg(dp + q);
        const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const int i = q * 4 + r;
          const uint32_t mb = (mw[i >> 2] >> (8 * (i & 3))) & 0xffu;
          const double z = p.depth_fmt == RDP_DEPTH_F32 ? (double)__fmul_rn(dv[r], p.scale_f32)
                                                        : __dmul_rn((double)dv[r], p.scale);
          if (mb && z > 0.0) valid |= 1u << i;
        }
      }
    }
  } else {
    for (int i = 0; i < DP_PX_PER_THREAD; ++i) {
      const int px = px0 + i;
      if (px < p.P && p.mask[fbase + px] && depth_to_z(p, fbase + px) > 0.0) valid |= 1u << i;
    }
  }

  // block-wide exclusive scan of the per-thread counts
  const int cnt = __popc(valid);
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
  while (valid) {
    const int i = __ffs(valid) - 1;
    valid &= valid - 1;
    s_pix[local++] = (uint32_t)(px0 + i);
  }
  __syncthreads();

  // cooperative, coalesced emission of the tile's points
  const int total = s_total, base = s_base;
  unsigned long long kmax = 0, kmin_inv = 0;
  for (int j = tid; j < total; j += DP_THREADS) {
    const uint32_t px = s_pix[j];
    const int v = (int)(px / (uint32_t)p.W), u = (int)(px - (uint32_t)v * (uint32_t)p.W);
    const double z = depth_to_z(p, fbase + px);
    const double x = deproject_axis((double)u, p.cx, z, p.fx);
    const size_t o = fbase + base + j;
    p.pix[o] = px;
    if (p.out.points_xyz) {
      const double y = deproject_axis((double)v, p.cy, z, p.fy);
      float* dst = p.out.points_xyz + 3 * o;
      dst[0] = (float)x;
      dst[1] = (float)y;
      dst[2] = (float)z;
    }
    const unsigned long long k = ordered_key(x);
    kmax = k > kmax ? k : kmax;
    kmin_inv = ~k > kmin_inv ? ~k : kmin_inv;
  }
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

// ---------------------------------------------------------------------------------------------------------
// k_bin_hist / k_bin_offsets / k_scatter: counting sort of the points by x-bin
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PT_THREADS) k_bin_hist(const Params p) {
  __shared__ int s_hist[NUM_BINS];
  const int b = blockIdx.y;
  const FrameState& f = p.fs[b];
  const int n = f.n_points, i0 = blockIdx.x * PT_PER_BLOCK;
  double x_min, width;
  if (i0 >= n || !frame_bins(f, x_min, width)) return;
  if (threadIdx.x < NUM_BINS) s_hist[threadIdx.x] = 0;
  __syncthreads();
  const size_t fbase = (size_t)b * p.P;
  const int i1 = min(n, i0 + PT_PER_BLOCK);
  for (int i = i0 + threadIdx.x; i < i1; i += PT_THREADS) {
    const uint32_t px = p.pix[fbase + i];
    const int v = (int)(px / (uint32_t)p.W), u = (int)(px - (uint32_t)v * (uint32_t)p.W);
    const double x = deproject_axis((double)u, p.cx, depth_to_z(p, fbase + px), p.fx);
    const int bin = bin_of(x, x_min, width);
    p.bins[fbase + i] = (uint8_t)bin;
    atomicAdd(&s_hist[bin], 1);
  }
  __syncthreads();
  if (threadIdx.x < NUM_BINS && s_hist[threadIdx.x]) atomicAdd(&p.fs[b].bin_count[threadIdx.x], s_hist[threadIdx.x]);
}

__global__ void k_bin_offsets(const Params p) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= p.B) return;
  FrameState& f = p.fs[b];
  double x_min, width;
  int status = RDP_FRAME_OK, off = 0, eoff = 0;
  if (f.n_points < MIN_POINTS) {
    status = RDP_FRAME_TOO_FEW_POINTS;
  } else if (!frame_bins(f, x_min, width)) {
    status = RDP_FRAME_DEGENERATE_BINS;
  }
  for (int i = 0; i < NUM_BINS; ++i) {
    const int c = status == RDP_FRAME_OK ? f.bin_count[i] : 0;
    int k = 0;
    if (c > 0) {
      k = (int)__dmul_rn((double)c, 0.05);  // int(cnt * top_k_percent), geometry_utils.py:138
      k = k < 1 ? 1 : k;
    }
    f.bin_k[i] = k;
    f.bin_off[i] = off;
    f.edge_off[i] = eoff;
    off += c;
    eoff += k;
    if (p.out.bin_counts) p.out.bin_counts[b * NUM_BINS + i] = c;
  }
  f.bin_off[NUM_BINS] = off;
  f.edge_off[NUM_BINS] = eoff;
  f.n_edge = eoff;
  if (status == RDP_FRAME_OK && eoff < MIN_EDGE) status = RDP_FRAME_TOO_FEW_EDGE;
  f.status = status;
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
  for (int i = i0 + threadIdx.x; i < i1; i += PT_THREADS) atomicAdd(&s_cnt[p.bins[fbase + i]], 1);
  __syncthreads();
  if (threadIdx.x < NUM_BINS) {
    const int c = s_cnt[threadIdx.x];
    s_base[threadIdx.x] = f.bin_off[threadIdx.x] + (c ? atomicAdd(&f.bin_cursor[threadIdx.x], c) : 0);
    s_cnt[threadIdx.x] = 0;
  }
  __syncthreads();
  for (int i = i0 + threadIdx.x; i < i1; i += PT_THREADS) {
    const int bin = p.bins[fbase + i];
    p.binned[fbase + s_base[bin] + atomicAdd(&s_cnt[bin], 1)] = p.pix[fbase + i];
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

__global__ void __launch_bounds__(SEL_THREADS) k_select_sort(const Params p) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  unsigned long long* s_yk = reinterpret_cast<unsigned long long*>(s_raw);                  // [STAGE_CAP]
  uint32_t* s_px = reinterpret_cast<uint32_t*>(s_yk + STAGE_CAP);                            // [STAGE_CAP]
  unsigned long long* s_sk = reinterpret_cast<unsigned long long*>(s_px + STAGE_CAP);        // [SORT_CAP]
  uint32_t* s_sp = reinterpret_cast<uint32_t*>(s_sk + SORT_CAP);                             // [SORT_CAP]
  __shared__ int s_hist[256];
  __shared__ unsigned long long s_ph, s_mm[2][SEL_THREADS / 32];
  __shared__ uint32_t s_pl;
  __shared__ int s_krem, s_done, s_nsel, s_start;

  const int bin = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const FrameState& f = p.fs[b];
  if (f.status != RDP_FRAME_OK) return;
  const int cnt = f.bin_count[bin], k = f.bin_k[bin];
  if (cnt == 0) return;
  const size_t fbase = (size_t)b * p.P;
  const uint32_t* seg = p.binned + fbase + f.bin_off[bin];
  const bool staged = cnt <= STAGE_CAP;

  auto y_key_of = [&](uint32_t px) {
    const int v = (int)(px / (uint32_t)p.W);
    return ordered_key(deproject_axis((double)v, p.cy, depth_to_z(p, fbase + px), p.fy));
  };

  // stage the segment and find the common leading digits of the y keys
  unsigned long long kmax = 0, kmin_inv = 0;
  for (int i = tid; i < cnt; i += SEL_THREADS) {
    const uint32_t px = seg[i];
    const unsigned long long yk = y_key_of(px);
    if (staged) {
      s_yk[i] = yk;
      s_px[i] = px;
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
    s_done = (k >= cnt);  // everything is selected (cnt < 20 -> k = 1 = cnt, or tiny bins)
    s_nsel = 0;
  }
  __syncthreads();

  if (!s_done) {
    for (int d = s_start; d < 12; ++d) {
      s_hist[tid] = 0;
      __syncthreads();
      const unsigned long long ph = s_ph;
      const uint32_t pl = s_pl;
      for (int i = tid; i < cnt; i += SEL_THREADS) {
        unsigned long long yk;
        uint32_t px;
        if (staged) {
          yk = s_yk[i];
          px = s_px[i];
        } else {
          px = seg[i];
          yk = y_key_of(px);
        }
        if (key_has_prefix(yk, ~px, d, ph, pl)) atomicAdd(&s_hist[key_digit(yk, ~px, d)], 1);
      }
      __syncthreads();
      if (tid == 0) {
        int krem = s_krem, cum = 0, dig = 255;
        for (; dig > 0; --dig) {
          if (cum + s_hist[dig] >= krem) break;
          cum += s_hist[dig];
        }
        krem -= cum;
        if (d < 8) s_ph |= (unsigned long long)dig << (56 - 8 * d);
        else s_pl |= (uint32_t)dig << (24 - 8 * (d - 8));
        s_krem = krem;
        if (s_hist[dig] == krem) s_done = 1;  // the whole bucket is taken: threshold = prefix with zeros below
      }
      __syncthreads();
      if (s_done) break;
    }
  }

  // gather the selected points with their x sort key
  const unsigned long long th = s_done && k >= cnt ? 0ull : s_ph;
  const uint32_t tl = s_done && k >= cnt ? 0u : s_pl;
  const bool sort_in_smem = k <= SORT_CAP;
  const size_t ebase = (size_t)b * p.ecap + f.edge_off[bin];
  unsigned long long* sk = sort_in_smem ? s_sk : p.edge_key + ebase;
  uint32_t* sp = sort_in_smem ? s_sp : p.edge_pix + ebase;
  for (int i = tid; i < cnt; i += SEL_THREADS) {
    unsigned long long yk;
    uint32_t px;
    if (staged) {
      yk = s_yk[i];
      px = s_px[i];
    } else {
      px = seg[i];
      yk = y_key_of(px);
    }
    if (yk > th || (yk == th && ~px >= tl)) {
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
  }
}

constexpr size_t SELECT_SMEM = (size_t)STAGE_CAP * 12 + (size_t)SORT_CAP * 12;

struct Plan {
  size_t header_bytes, total_bytes;
  int tiles, ecap;
};

static Plan carve(Params& p, void* ws, int B, int H, int W) {
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
  p.edge_pix = c.take<uint32_t>((size_t)B * pl.ecap);
  p.edge_key = c.take<unsigned long long>((size_t)B * pl.ecap);
  p.edge_xyz = c.take<double>((size_t)B * pl.ecap * 3);
  p.edge_u = c.take<double>((size_t)B * pl.ecap);
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
  *bytes = carve(p, nullptr, B, H, W).total_bytes;
  return RDP_OK;
}

extern "C" int rdp_geom_launch_count(void) { return 7; }

static int curvature_impl(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                          const double* intrinsics, double depth_scale, const RdpCurvatureOut* out,
                          void* workspace, size_t workspace_bytes, void* stream_, cudaEvent_t* marks) {
  RDP_REQUIRE(mask && depth && intrinsics && out && workspace, RDP_ERR_BAD_ARGUMENT, "rdp_geom_curvature: null pointer");
  RDP_REQUIRE(B > 0 && H > 0 && W > 0, RDP_ERR_BAD_ARGUMENT, "rdp_geom_curvature: bad shape %dx%dx%d", B, H, W);
  RDP_REQUIRE((size_t)H * W < (1ull << 31) / 4, RDP_ERR_UNSUPPORTED, "rdp_geom_curvature: frame too large");
  RDP_REQUIRE(depth_fmt >= RDP_DEPTH_U16 && depth_fmt <= RDP_DEPTH_F32_WIDE, RDP_ERR_BAD_ARGUMENT,
              "rdp_geom_curvature: unknown depth format %d", depth_fmt);
  RDP_REQUIRE(out->mean_curvature && out->max_curvature && out->spline_points && out->status, RDP_ERR_BAD_ARGUMENT,
              "rdp_geom_curvature: mean/max/spline_points/status outputs are required");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  Params p{};
  const Plan pl = carve(p, workspace, B, H, W);
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

  static bool smem_opt_in = false;  // idempotent attribute; a benign race sets it twice at worst
  if (!smem_opt_in) {
    RDP_CUDA_TRY(cudaFuncSetAttribute(k_select_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SELECT_SMEM));
    smem_opt_in = true;
  }
  int n_marks = 0;
  auto mark = [&]() {
    if (marks) cudaEventRecord(marks[n_marks++], stream);
  };
  mark();
  RDP_CUDA_TRY(cudaMemsetAsync(workspace, 0, pl.header_bytes, stream));
  mark();
  const bool vec = (p.P % 16 == 0) && ((uintptr_t)mask % 16 == 0) && ((uintptr_t)depth % 16 == 0);
  const int tiles_total = B * pl.tiles;
  if (vec) k_deproject<true><<<tiles_total, DP_THREADS, 0, stream>>>(p);
  else k_deproject<false><<<tiles_total, DP_THREADS, 0, stream>>>(p);
  mark();
  const dim3 pgrid((p.P + PT_PER_BLOCK - 1) / PT_PER_BLOCK, B);
  k_bin_hist<<<pgrid, PT_THREADS, 0, stream>>>(p);
  mark();
  k_bin_offsets<<<(B + 63) / 64, 64, 0, stream>>>(p);
  mark();
  k_scatter<<<pgrid, PT_THREADS, 0, stream>>>(p);
  mark();
  k_select_sort<<<dim3(NUM_BINS, B), SEL_THREADS, SELECT_SMEM, stream>>>(p);
  mark();
  k_fit<<<B, FIT_THREADS, 0, stream>>>(p);
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
constexpr int STATIC_MAX_TILES = 16384;  // sum over the batch of ceil(H*W / 4096)
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
  RDP_REQUIRE(depth_fmt >= RDP_DEPTH_U16 && depth_fmt <= RDP_DEPTH_F32_WIDE, RDP_ERR_BAD_ARGUMENT,
              "rdp_geom_deproject_static: unknown depth format %d", depth_fmt);
  const int P = H * W;
  const int tiles = (P + TILE_PX - 1) / TILE_PX;
  RDP_REQUIRE(B <= STATIC_MAX_FRAMES && (long long)B * tiles <= STATIC_MAX_TILES, RDP_ERR_UNSUPPORTED,
              "rdp_geom_deproject_static: batch of %d frames x %d tiles exceeds the built-in scratch (at most %d frames and "
              "%d tiles of 4096 pixels per call); use rdp_geom_curvature with a workspace for larger batches",
              B, tiles, STATIC_MAX_FRAMES, STATIC_MAX_TILES);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
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
  const bool vec = (P % 16 == 0) && ((uintptr_t)mask % 16 == 0) && ((uintptr_t)depth % 16 == 0);
  if (vec) k_deproject<true><<<B * tiles, DP_THREADS, 0, stream>>>(p);
  else k_deproject<false><<<B * tiles, DP_THREADS, 0, stream>>>(p);
  k_export_counts<<<(B + 63) / 64, 64, 0, stream>>>(p.fs, B, n_points);
  RDP_CUDA_TRY(cudaGetLastError());
  return RDP_OK;
}
