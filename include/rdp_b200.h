/*
 * rdp_b200.h - C ABI of the B200-native vision hot path (U-Net forward + curvature profile).
 *
 * The reference (xuanjiangliu/robotic-discovery-platform) has no FFI of its own: the hot path sits behind two
 * Python symbols imported by its gRPC server (services/vision_analysis/server.py:47-48).  This header is the
 * boundary underneath our drop-in `pkg.segmentation_model.UNet` / `pkg.geometry_utils.compute_curvature_profile`:
 * plain pointers and sizes, no torch types, callable from ctypes or any C caller.
 *
 * Conventions
 *  - every data pointer is a DEVICE pointer owned by the caller unless a comment says "host";
 *  - calls are asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy default stream);
 *  - return value: 0 on success, a negative RDP_ERR_* otherwise; rdp_last_error() gives the text for the
 *    calling thread.  No exceptions cross this boundary;
 *  - per-frame *data* conditions (too few points, degenerate bins, ...) are not errors: they are reported in
 *    the per-frame `status` output (RDP_FRAME_*), mirroring the reference's "return an empty CurvatureResult";
 *  - handles are immutable after creation; workspaces are per call, so concurrent callers (the reference server
 *    runs 10 worker threads, server.py:172) only need distinct workspaces and streams.
 */
#ifndef RDP_B200_H
#define RDP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RDP_OK 0
#define RDP_ERR_BAD_ARGUMENT (-1)
#define RDP_ERR_UNSUPPORTED (-2)
#define RDP_ERR_CUDA (-3)
#define RDP_ERR_WORKSPACE_TOO_SMALL (-4)

/* per-frame status written by rdp_geom_curvature (same numbering as oracle/geometry_port.py) */
#define RDP_FRAME_OK 0
#define RDP_FRAME_TOO_FEW_POINTS 1  /* < 100 valid points             geometry_utils.py:64  */
#define RDP_FRAME_DEGENERATE_BINS 2 /* < 50 points or bin width <= 0  geometry_utils.py:121,127 */
#define RDP_FRAME_TOO_FEW_EDGE 3    /* < 20 edge points               geometry_utils.py:69  */
#define RDP_FRAME_FIT_INVALID 4     /* splprep would raise (parameter not strictly increasing) :95-97 */
#define RDP_FRAME_NEEDS_KNOTS 5     /* internal: set between the 8-knot fit and the knot-insertion kernel, never returned */

/* depth element formats */
#define RDP_DEPTH_U16 0 /* z = (double)d * scale                     (uint16 image * Python float) */
#define RDP_DEPTH_F32 1 /* z = (double)(float)(d * (float)scale)     (float32 image * Python float) */
#define RDP_DEPTH_F32_WIDE 2 /* z = (double)d * scale                (float32 image * numpy float64 scalar) */
#define RDP_DEPTH_F64 3 /* z = d * scale, d a double                  (float64 image; other integer dtypes widened by the host) */

/* U-Net input formats */
#define RDP_IN_RGB_U8_NHWC 0 /* (B,H,W,C) uint8, scaled by 1/255 on load (ToTensor, server.py:108) */
#define RDP_IN_F32_NCHW 1    /* (B,C,H,W) float32, the tensor the reference module receives (server.py:121) */
/* OR into in_fmt: keep every conv's activation in the workspace, each in storage of its own (parity hook:
 * rdp_unet_read_activation needs it; the workspace must then have the size rdp_unet_workspace_bytes_ex reports) */
#define RDP_IN_FLAG_KEEP_ACTIVATIONS 0x100

int rdp_abi_version(void);
const char* rdp_last_error(void);

/* ------------------------------------------------------------------------------------------------------ */
/* U-Net: replaces UNet.forward (pkg/segmentation_model.py:109-120) and, optionally, the server's          */
/* `sigmoid(out) > 0.5` mask rule (services/vision_analysis/server.py:124).                                 */
/* ------------------------------------------------------------------------------------------------------ */

typedef struct RdpUnet RdpUnet;

/* One 3x3 conv + eval-mode BatchNorm + ReLU (DoubleConv halves, segmentation_model.py:31-36), already folded:
 * weight = bf16 [cout][9][cin_pad] with gamma/sqrt(var+eps) multiplied in, tap index = 3*ky+kx, channels padded
 * with zeros to a multiple of 64 (the first layer is stored im2col-style as [cout][64], k = (3*ky+kx)*cin + c);
 * shift = fp32 [cout] = beta - mean*gamma/sqrt(var+eps). */
typedef struct {
  const void* weight; /* device, bf16 */
  const float* shift; /* device, fp32 */
  int32_t cin;        /* logical input channels (before padding) */
  int32_t cout;
} RdpConvLayer;

typedef struct {
  int32_t n_channels;      /* UNet(n_channels, ...) ; 9*n_channels must be <= 64 */
  int32_t n_classes;       /* UNet(..., n_classes)  ; 1..8 */
  RdpConvLayer conv[18];   /* inc.0, inc.3, down1.0, down1.3, ... down4.3, up1.0, ... up4.3 in forward order */
  const float* outc_weight; /* device fp32 [n_classes][64]   OutConv, segmentation_model.py:81 */
  const float* outc_bias;   /* device fp32 [n_classes] */
} RdpUnetWeights;

/* Builds an immutable handle (copies the small descriptor, keeps the caller's device pointers).  A handle belongs to
 * the device that was current when it was created (its weights live there); a process may hold handles on several
 * devices and call each with its own device current. */
int rdp_unet_create(const RdpUnetWeights* weights, RdpUnet** out);
void rdp_unet_destroy(RdpUnet* net);

/* Scratch for activations of one forward call of shape (B,H,W).  Buffers whose lifetimes do not overlap share
 * storage (8.2 GB for 64 frames of 640x480).  A forward called with RDP_IN_FLAG_KEEP_ACTIVATIONS keeps every layer's
 * output instead and needs the larger size rdp_unet_workspace_bytes_ex(net, RDP_IN_FLAG_KEEP_ACTIVATIONS, ...) gives. */
int rdp_unet_workspace_bytes(const RdpUnet* net, int B, int H, int W, size_t* bytes);
int rdp_unet_workspace_bytes_ex(const RdpUnet* net, int in_fmt_flags, int B, int H, int W, size_t* bytes);

/* logits: fp32 (B, n_classes, H, W) or NULL.  mask: uint8 (B, n_classes, H, W) in {0,1} or NULL; it is
 * [sigmoidf(logit) > 0.5f] evaluated in fp32 like the server does.  At least one of them must be non-NULL. */
int rdp_unet_forward(const RdpUnet* net, const void* input, int in_fmt, int B, int H, int W, float* logits,
                     uint8_t* mask, void* workspace, size_t workspace_bytes, void* stream);

/* Number of kernel launches one rdp_unet_forward call issues (for bench.py's gpu_launches). */
int rdp_unet_launch_count(const RdpUnet* net);

/* Measurement aid (no counterpart in the reference): the same forward with a CUDA event recorded on `stream` between
 * launches; blocks until the call has finished and fills launch_ms (HOST, rdp_unet_launch_count entries) in launch
 * order: im2col, inc.0, inc.3, then (maxpool, conv, conv) x4, then (upsample, conv, conv) x4. */
int rdp_unet_forward_profiled(const RdpUnet* net, const void* input, int in_fmt, int B, int H, int W, float* logits,
                              uint8_t* mask, void* workspace, size_t workspace_bytes, void* stream, float* launch_ms);

/* Debug / parity hook: copy of an intermediate activation as fp32 NCHW.  `index` = 0..17 (conv outputs in the
 * order of RdpUnetWeights.conv).  Must be called after an rdp_unet_forward with RDP_IN_FLAG_KEEP_ACTIVATIONS on the
 * same workspace and stream. */
int rdp_unet_read_activation(const RdpUnet* net, int index, int B, int H, int W, const void* workspace,
                             float* out_nchw, void* stream);

/* ------------------------------------------------------------------------------------------------------ */
/* Geometry: replaces compute_curvature_profile (pkg/geometry_utils.py:42-97) for a batch of frames.        */
/* ------------------------------------------------------------------------------------------------------ */

typedef struct {
  double* mean_curvature;  /* [B]          geometry_utils.py:160 */
  double* max_curvature;   /* [B] */
  double* spline_points;   /* [B][100][3]  geometry_utils.py:83-87 */
  int32_t* status;         /* [B] RDP_FRAME_* */
  int32_t* n_points;       /* [B] valid masked points N                      (NULL allowed) */
  int32_t* n_edge;         /* [B] edge points E                              (NULL allowed) */
  double* fit_residual;    /* [B] fp of the 8-knot least-squares cubic        (NULL allowed) */
  float* points_xyz;       /* [B][H*W][3] fp32 point cloud, row-major order   (NULL allowed) _get_pcd_from_mask */
  uint32_t* edge_pixels;   /* [B][H*W/20+64] pixel index v*W+u of the edge points, ordered by x (NULL allowed) */
  int32_t* bin_counts;     /* [B][50] points per x-bin                        (NULL allowed) */
  int32_t* n_knots;        /* [B] knots of the fitted spline (8 = single cubic)  (NULL allowed) */
  int32_t* fit_ier;        /* [B] FITPACK's ier: -2, 0, or 1/2/3 (warnings)     (NULL allowed) */
  int32_t* fit_flags;      /* [B] RDP_FIT_FLAG_* bits, 0 for the single-cubic regime (NULL allowed; ABI version 2) */
} RdpCurvatureOut;

/* fit_flags: decisions inside FITPACK's knot-insertion regime (splprep, geometry_utils.py:78) that were taken on a
 * margin smaller than the rounding noise of the least-squares solves.  FITPACK itself (scipy) takes the same
 * decisions on the same margins, so on a flagged frame two correct implementations - or scipy on two machines - may
 * place a knot differently and report curvatures that differ by far more than 1e-3; an
 * unflagged frame follows scipy's knot sequence exactly. */
#define RDP_FIT_FLAG_KNOT_TIE 1    /* fpknot: largest and second-largest interval residual within ~1e-9 relative
                                      (and not the bit-identical halves of one split) */
#define RDP_FIT_FLAG_BRACKET_TIE 2 /* smoothing root find: (p/0.04)*0.04 compared with p, an exact-arithmetic
                                      equality FITPACK decides by the last bit; resolved here as exact arithmetic does */
#define RDP_FIT_FLAG_NPLUS_TIE 4   /* the number of knots to add, int(nplus*(fp-s)/(fpold-fp)), sat on an integer */

int rdp_geom_workspace_bytes(int B, int H, int W, size_t* bytes);

/* mask: uint8 (B,H,W), pixel selected when > 0.  depth: (B,H,W) of `depth_fmt`.  intrinsics: HOST pointer to
 * 4 doubles fx, fy, cx, cy = K[0,0], K[1,1], K[0,2], K[1,2] (geometry_utils.py:111-112), shared by the batch. */
int rdp_geom_curvature(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                       const double* intrinsics, double depth_scale, const RdpCurvatureOut* out, void* workspace,
                       size_t workspace_bytes, void* stream);

/* Number of kernel launches (incl. the workspace memset) one rdp_geom_curvature call issues. */
int rdp_geom_launch_count(void);

/* Measurement aid (no counterpart in the reference): as rdp_geom_curvature, with a CUDA event between launches;
 * blocks, then fills launch_ms (HOST, rdp_geom_launch_count entries): memset, deproject, bin_hist, scatter,
 * select_sort, fit, fit_knots. */
int rdp_geom_curvature_profiled(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                                const double* intrinsics, double depth_scale, const RdpCurvatureOut* out,
                                void* workspace, size_t workspace_bytes, void* stream, float* launch_ms);

/* ------------------------------------------------------------------------------------------------------ */
/* Deprojection without a workspace.  (No counterpart in the reference: it exists for hosts that cannot      */
/* allocate device memory, not to mirror one of its functions; the arithmetic is _get_pcd_from_mask's.)      */
/* ------------------------------------------------------------------------------------------------------ */

/* For callers that may not call cudaMalloc - a server that reserved the whole GPU for its own pool and bans
 * allocation on the request path, a control loop that cannot wait for (or survive the failure of) an allocation.
 * The scratch the deprojection kernel needs lives in a fixed array inside the library instead of a workspace.
 *
 *   mask, depth, depth_fmt, intrinsics (host, 4 doubles), depth_scale: as for rdp_geom_curvature
 *   points_xyz  [B][H*W][3] fp32   the point cloud, row-major pixel order        (output, device)
 *   pixel_ids   [B][H*W]    uint32 flat pixel index v*W+u of every point         (output, device)
 *   n_points    [B]         int32  number of valid points per frame              (output, device)
 *
 * Limits: at most 64 frames and at most 4096 tiles of 16384 pixels per call (B * ceil(H*W/16384)); larger batches
 * are refused with RDP_ERR_UNSUPPORTED.  The array (about 100 KB) is held by every GPU that loads the library
 * until the process exits.  The call itself performs no allocation of any kind.
 *
 * The first call on each device looks up the array's device address; make that first call before any stream
 * capture begins.  On EVERY call the device that `stream` belongs to must be the current device (the array is per
 * device and is found through the current one); a stream of another device is refused with RDP_ERR_BAD_ARGUMENT
 * before any work is issued.  All calls share the array and each clears what it uses
 * before running, so calls whose GPU work could overlap (different streams, different threads) race: keep them
 * on one stream or wait for one to finish before issuing the next. */
int rdp_geom_deproject_static(const uint8_t* mask, const void* depth, int depth_fmt, int B, int H, int W,
                              const double* intrinsics, double depth_scale, float* points_xyz, uint32_t* pixel_ids,
                              int32_t* n_points, void* stream);

/* ------------------------------------------------------------------------------------------------------ */
/* Server-side pre/post-processing around the hot path (services/vision_analysis/server.py)                 */
/* ------------------------------------------------------------------------------------------------------ */

/* Replaces `self.preprocess(image_rgb).unsqueeze(0)` (server.py:107-110,121): torchvision ToTensor (uint8/255,
 * HWC -> CHW) followed by Resize((out_h,out_w), antialias=True) - ATen's separable antialiased bilinear filter,
 * horizontal pass first, weights normalised per output index - in one pass, fp32.
 *   rgb_nhwc [B][H][W][C] uint8 (device), C in 1..8      out_nchw [B][C][out_h][out_w] fp32 (device)
 *   flags: 0 or RDP_PRE_SWAP_RB
 * One launch, no workspace.  Agreement with torchvision: a few fp32 ulp (the tests hold it to 1e-6 absolute). */
#define RDP_PRE_SWAP_RB 1 /* input is BGR (cv2.imdecode): also does cv2.cvtColor(.., COLOR_BGR2RGB), server.py:120 */
int rdp_pre_resize(const uint8_t* rgb_nhwc, int B, int H, int W, int C, int out_h, int out_w, int flags,
                   float* out_nchw, void* stream);

/* Replaces `cv2.resize(mask_resized, (W, H), interpolation=cv2.INTER_NEAREST)` (server.py:125) and the count
 * behind `mask_coverage_percent` (server.py:133): source index = min(floor(dst * (1 / (dst_size / src_size))),
 * src_size - 1) in double, OpenCV's rule (not the half-pixel-centre one), exact.
 *   mask [B][h][w] uint8 (device)     out [B][H][W] uint8 (device)
 *   nonzero [B] int32 (device, optional): number of non-zero pixels of each up-sized mask; cleared by the call. */
int rdp_post_mask_nearest(const uint8_t* mask, int B, int h, int w, int H, int W, uint8_t* out, int32_t* nonzero,
                          void* stream);

/* Replaces `cv2.imencode('.png', final_mask * 255)` (server.py:142-143): the mask leaves the GPU as the PNG the
 * response carries, not as H*W bytes for a host encoder.  The client reads it with cv2.imdecode(.., IMREAD_GRAYSCALE)
 * (client.py:111-112); the bytes are not OpenCV's (1 bit per pixel, filter 0, one fixed-Huffman deflate block of
 * literals and distance-1 matches), the decoded picture is: 255 where mask != 0, else 0.
 *   mask [B][H][W] uint8 (device)
 *   png  [B][png_stride] bytes (device, 4-byte aligned, png_stride a multiple of 4 and >= max_png_bytes)
 *   png_bytes [B] int32 (device): length of each frame's PNG
 * rdp_post_mask_png_sizes gives max_png_bytes (per frame, worst case) and the workspace the call needs.  Two launches. */
int rdp_post_mask_png_sizes(int B, int H, int W, size_t* max_png_bytes, size_t* workspace_bytes);
int rdp_post_mask_png(const uint8_t* mask, int B, int H, int W, uint8_t* png, size_t png_stride, int32_t* png_bytes,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------ */
/* Wire-format decoders: replace the two cv2.imdecode calls at the top of the server loop */
/* (services/vision_analysis/server.py:117-118) for a batch of frames.                                       */
/* ------------------------------------------------------------------------------------------------------ */

/* per-image status of the decoders */
#define RDP_DECODE_OK 0
#define RDP_DECODE_NOT_AN_IMAGE 1     /* signature / SOI missing */
#define RDP_DECODE_TRUNCATED 2        /* marker or chunk structure broken */
#define RDP_DECODE_UNSUPPORTED 3      /* JPEG: not baseline 8-bit YCbCr 4:4:4/4:2:2/4:2:0 or grey; PNG: not 8/16-bit grey */
#define RDP_DECODE_WRONG_SIZE 4       /* the image is not H x W */
#define RDP_DECODE_CORRUPT 5          /* bad Huffman code / zlib stream */
#define RDP_DECODE_LENGTH_MISMATCH 6  /* PNG: inflated data shorter or longer than H rows */
#define RDP_DECODE_BAD_FILTER 7       /* PNG: filter type > 4 */

/* Workspace sizes of the two decoders for a batch of B images of H x W (either pointer may be NULL). */
int rdp_pre_decode_sizes(int B, int H, int W, size_t* jpeg_workspace_bytes, size_t* png_workspace_bytes);

/* Replaces `cv2.imdecode(np.frombuffer(request.color_image.data, np.uint8), cv2.IMREAD_COLOR)` (server.py:117) for the
 * JPEGs `cv2.imencode('.jpg', ...)` writes (client.py:65): baseline sequential Huffman JPEG, decoded with libjpeg's
 * accurate integer inverse DCT, "fancy" chroma up-sampling and fixed-point colour conversion, so the BGR bytes equal
 * OpenCV's (libjpeg-turbo) bit for bit.
 *   data     concatenated JPEG files (device)            offsets  [B+1] int64 byte offsets into data (device)
 *   bgr      [B][H][W][3] uint8 (device)                 status   [B] int32 RDP_DECODE_* (device); a failed image is all zero
 * Two launches.  A Huffman-coded scan is one serial bit stream: one warp per image, so throughput comes from the batch. */
int rdp_pre_decode_jpeg(const uint8_t* data, const int64_t* offsets, int B, int H, int W, uint8_t* bgr, int32_t* status,
                        void* workspace, size_t workspace_bytes, void* stream);

/* Replaces `cv2.imdecode(np.frombuffer(request.depth_image.data, np.uint8), cv2.IMREAD_UNCHANGED)` (server.py:118) for
 * the PNGs `cv2.imencode('.png', depth_uint16)` writes (client.py:66): non-interlaced 16-bit (or 8-bit) greyscale, any
 * zlib stream (stored / fixed / dynamic Huffman blocks), all five scan-line filters.
 *   depth    [B][H][W] uint16 (device; 8-bit images are widened)   status [B] int32 RDP_DECODE_* (device)
 * Two launches; one warp per image inflates, one CTA per image undoes the filters. */
int rdp_pre_decode_png16(const uint8_t* data, const int64_t* offsets, int B, int H, int W, uint16_t* depth, int32_t* status,
                         void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------ */
/* Training step: what `model(images)` in train mode and `loss.backward()` compute in the  */
/* reference's loop (scripts/train_segmenter.py:151-165).  The loss and the optimiser stay the caller's        */
/* (torch's BCEWithLogitsLoss and Adam there); BatchNorm running statistics are updated by the caller from the */
/* batch statistics returned here (momentum 0.1, unbiased variance, pkg/segmentation_model.py:32,35).          */
/* ------------------------------------------------------------------------------------------------------ */

typedef struct {
  const void* w;      /* bf16 [cout][9][cin_pad], tap = 3*ky+kx, NOT folded with BatchNorm (first layer: [cout][64], k = tap*cin + c) */
  const void* w_t;    /* bf16 [cin_pad][9][cout]: w_t[ci][3*ky+kx][co] = W[co][ci][2-ky][2-kx], for the input gradient (NULL for layer 0) */
  const float* gamma; /* fp32 [cout]  BatchNorm2d.weight */
  const float* beta;  /* fp32 [cout]  BatchNorm2d.bias   */
  int32_t cin, cout;
} RdpTrainLayer;

typedef struct {
  int32_t n_channels, n_classes;
  RdpTrainLayer conv[18];
  const float* outc_weight; /* fp32 [n_classes][64] */
  const float* outc_bias;   /* fp32 [n_classes] */
  float bn_eps;             /* 1e-5 when 0 */
} RdpTrainWeights;

typedef struct {  /* batch statistics of every BatchNorm layer (device, fp32 [cout]; entries may be NULL) */
  float* mean[18];
  float* var_biased[18];
} RdpTrainBatchStats;

typedef struct {  /* parameter gradients (device, fp32), each overwritten by the call */
  float* dw[18];      /* [cout][cin][3][3], the layout of Conv2d.weight */
  float* dgamma[18];  /* [cout] */
  float* dbeta[18];   /* [cout] */
  float* d_outc_weight; /* [n_classes][64] */
  float* d_outc_bias;   /* [n_classes] */
} RdpTrainGrads;

/* Conv2d.weight as the optimiser holds it (fp32 [cout][cin][3][3], pkg/segmentation_model.py:26-33) -> the two bf16
 * layouts above, once per step after the optimiser's update.  first_layer: w is [cout][64]; w_t may be NULL. */
int rdp_unet_train_pack(const float* w_oihw, int cout, int cin, int first_layer, void* w, void* w_t, void* stream);

/* H and W must be divisible by 16 (the reference trains at 256 x 256, train_segmenter.py:49). */
int rdp_unet_train_workspace_bytes(const RdpTrainWeights* w, int B, int H, int W, size_t* bytes);

/* x_nchw: fp32 (B, n_channels, H, W) as the DataLoader yields it; logits: fp32 (B, n_classes, H, W).  Keeps every
 * activation the backward pass needs in `workspace`. */
int rdp_unet_train_forward(const RdpTrainWeights* w, const float* x_nchw, int B, int H, int W, float* logits,
                           const RdpTrainBatchStats* stats, void* workspace, size_t workspace_bytes, void* stream);

/* dlogits: fp32 (B, n_classes, H, W), the gradient of the loss with respect to the logits.  Must follow an
 * rdp_unet_train_forward with the same weights, shape and workspace. */
int rdp_unet_train_backward(const RdpTrainWeights* w, const float* dlogits, int B, int H, int W, const RdpTrainGrads* grads,
                            void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RDP_B200_H */
