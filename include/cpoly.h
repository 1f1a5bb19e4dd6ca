/*
 * cpoly.h — C ABI of libcpoly.so, the B200 (sm_100a) implementation of CenterPoly v2's
 * data-parallel hot path.  Plain pointers, sizes and ints only: no torch / C++ types.
 *
 * What each entry point replaces in the reference (paths relative to the reference checkout):
 *
 *   cpoly_dcn_forward    <- `_ext.dcn_v2_forward(input, weight, bias, offset, mask, kh,kw, sh,sw, ph,pw,
 *                           dh,dw, deformable_group)` of the DCNv2 extension, reached through
 *                           `DCN.forward` from src/lib/models/networks/pose_dla_dcn.py:354-359 (DeformConv)
 *                           and src/lib/models/networks/resnet_dcn.py:221-223.  The extension itself is the
 *                           git submodule src/lib/models/networks/DCNv2 (pinned dee2e77e…, not vendored).
 *   cpoly_dcn_backward   <- `_ext.dcn_v2_backward(input, weight, bias, offset, mask, grad_output, …)` of the
 *                           same extension (autograd of the 16 DCN layers; driven by
 *                           src/lib/trains/base_trainer.py:97-98 `loss.backward()`).
 *   cpoly_dcn_pack_weights  (new) derived tap-major / MMA-swizzled weight cache; the stored parameter stays
 *                           (Cout,Cin,kh,kw) so checkpoints (src/lib/models/model.py:31-130) are unaffected.
 *   cpoly_offset_conv_forward <- `conv_offset_mask(input)` inside upstream DCNv2's `DCN.forward` (the module built at
 *                           src/lib/models/networks/pose_dla_dcn.py:354); cuDNN in the reference.
 *   cpoly_idaup_upsample_forward <- the depthwise `nn.ConvTranspose2d` `up_i` of IDAUp
 *                           (src/lib/models/networks/pose_dla_dcn.py:372-375,385) fused with the skip add of :387.
 *   cpoly_decode         <- `polydet_decode(heat, polys, depth, reg, cat_spec_poly, K, rep)`  [+ optionally the
 *                           coordinate back-mapping of `polydet_post_process`, src/lib/utils/post_process.py:105-112]
 *                           src/lib/models/decode.py:512-670, i.e. `_nms` :13-19, `_topk` :117-133,
 *                           `_transpose_and_gather_feat` src/lib/models/utils.py:22-26, and (optionally, with
 *                           heat_is_logit=1) the `hm.sigmoid_()` of src/lib/detectors/polydet.py:28.
 *   cpoly_rasterise      <- the mask construction inside `format_and_write_to_cityscapes`,
 *                           src/lib/datasets/dataset/cityscapes.py:226-279 (PIL ImageDraw + `bresenham` + numpy).
 *   cpoly_poly_iou_forward / _backward <- the polygon-IoU arm of `PolyLoss.forward`, src/lib/models/losses.py:870-888,
 *                           :909-910 (`WeilPolygonClipper` :373-628, `area` :25-41), including the
 *                           `_transpose_and_gather_feat` of :855 and, in backward, what autograd does with the
 *                           reference's per-object graph.
 *   cpoly_stem_conv7x7_forward (+ cpoly_stem_pack_weights) <- `DLA.base_layer` = Conv2d(3,16,7,1,3) + BatchNorm2d + ReLU,
 *                           src/lib/models/networks/pose_dla_dcn.py:231-236,289 (BN folded by the caller, eval mode).
 *   cpoly_maxpool2x2_nhwc_forward <- `Tree.downsample` / `project` = nn.MaxPool2d(2, stride=2), pose_dla_dcn.py:201,207,262.
 *
 * Ownership: the caller allocates every input, output and workspace in device memory and passes raw
 * device pointers; the library allocates nothing, keeps no global state and never synchronises the
 * host.  All work is enqueued on `stream`.  Re-entrant (one Python thread per GPU under
 * nn.DataParallel, src/lib/models/data_parallel.py:80-81, is fine).
 *
 * Errors: every int-returning function returns 0 on success or a CPOLY_E* code; a human-readable
 * message for the calling thread is then available from cpoly_last_error().  The Python shim turns
 * a non-zero code into RuntimeError (what the upstream extension's AT_ASSERTM did).
 *
 * Tensors are described by a base pointer plus four ELEMENT strides in logical (b, c, h, w) order,
 * so NCHW-contiguous (the reference's default) and channels_last (NHWC physical) are both accepted.
 */
#ifndef CPOLY_H_
#define CPOLY_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CPOLY_ABI_VERSION 10

/* element types */
#define CPOLY_F32 0
#define CPOLY_BF16 1

/* error codes */
#define CPOLY_OK 0
#define CPOLY_EINVAL 1      /* bad argument (shape, dtype, NULL pointer, unsupported geometry) */
#define CPOLY_EWORKSPACE 2  /* workspace too small */
#define CPOLY_ECUDA 3       /* a CUDA runtime call or kernel launch failed */
#define CPOLY_EUNSUPPORTED 4

/* polygon representation of the poly head (src/lib/opts.py:72-74, decode.py:582,603) */
#define CPOLY_REP_CARTESIAN 0   /* also what 'polar_confined' / any other string decodes as */
#define CPOLY_REP_POLAR 1
#define CPOLY_REP_POLAR_FIXED 2

/* which DCN kernel family cpoly_dcn_forward / backward should use */
#define CPOLY_PATH_AUTO 0    /* tensor-core path when the shape allows, else SIMT */
#define CPOLY_PATH_SIMT 1    /* generic CUDA-core implicit GEMM (any geometry, any strides) */
#define CPOLY_PATH_TCGEN05 2 /* tcgen05/TMEM path; CPOLY_EUNSUPPORTED if the shape does not qualify */

typedef struct CUstream_st* cpoly_stream_t; /* == cudaStream_t */

typedef struct {
  void* ptr;          /* device pointer to element (0,0,0,0) */
  int64_t sb, sc, sh, sw; /* element strides of the logical (batch, channel, row, column) axes */
} cpoly_tensor4;

/* ------------------------------------------------------------------------------------------ DCN */

typedef struct {
  int32_t B, Cin, Cout, H, W;              /* input is (B,Cin,H,W); output (B,Cout,Ho,Wo) */
  int32_t kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w;
  int32_t deformable_groups;
  int32_t dtype;         /* CPOLY_F32 | CPOLY_BF16: element type of x and y (and of gout in backward) */
  int32_t om_dtype;      /* element type of offset and mask */
  int32_t mask_is_logit; /* 1: `mask` holds pre-sigmoid logits; the sigmoid of DCN.forward is fused */
  int32_t path;          /* CPOLY_PATH_* */
  cpoly_tensor4 x;       /* (B,Cin,H,W)                       */
  cpoly_tensor4 offset;  /* (B,dg*2*kh*kw,Ho,Wo): channel 2k = dy, 2k+1 = dx of tap k = i*kw+j */
  cpoly_tensor4 mask;    /* (B,dg*kh*kw,Ho,Wo)                */
  const float* weight;   /* (Cout,Cin,kh,kw) dense fp32 master weights (generic path and backward) */
  const float* bias;     /* (Cout) fp32 or NULL                                                 */
  const void* weight_packed; /* output of cpoly_dcn_pack_weights for `dtype`, or NULL (=> SIMT)  */
  cpoly_tensor4 y;       /* (B,Cout,Ho,Wo), element type `dtype`                                */
  /* Fused epilogue (all optional): y = act(out_scale[o] * conv + bias[o]).  With out_scale = gamma/sqrt(var+eps)
   * and bias = out_scale*dcn_bias + beta - out_scale*mean this is the eval-mode DeformConv of
   * pose_dla_dcn.py:347-359 (DCN -> BatchNorm2d -> ReLU) in one pass.  NULL / 0 => plain DCN. */
  const float* out_scale; /* (Cout) fp32 or NULL */
  int32_t relu;          /* 1: clamp at zero */
  int32_t reserved0;
  /* Optional bias of the offset-producing convolution (DCN.conv_offset_mask.bias, a2 in SURVEY 8), added to the
   * raw offsets / mask logits inside the kernel so the producer conv can run bias-free: offset_bias has
   * dg*2*kh*kw floats, mask_bias dg*kh*kw floats (added BEFORE the sigmoid when mask_is_logit). */
  const float* offset_bias;
  const float* mask_bias;
  /* Scratch for the tap-split schedule of small feature maps (cpoly_dcn_forward_workspace_bytes); may be
   * NULL, in which case the unsplit schedule is used. */
  void* workspace;
  size_t workspace_bytes;
} cpoly_dcn_fwd_args;

typedef struct {
  cpoly_dcn_fwd_args f;  /* forward problem (f.y is ignored) */
  cpoly_tensor4 gout;    /* (B,Cout,Ho,Wo), element type f.dtype */
  /* outputs — fp32 (goffset / gmask: see gom_dtype), any may have ptr == NULL to skip it (ctx.needs_input_grad) */
  cpoly_tensor4 gx;      /* (B,Cin,H,W)      */
  cpoly_tensor4 goffset; /* like offset      */
  cpoly_tensor4 gmask;   /* like mask; d/d(logit) when f.mask_is_logit */
  float* gweight;        /* (Cout,Cin,kh,kw) dense */
  float* gbias;          /* (Cout) */
  void* workspace;       /* >= cpoly_dcn_backward_workspace_bytes(&args) bytes, 256-byte aligned */
  size_t workspace_bytes;
  int32_t gom_dtype;     /* ABI v10: element type of goffset / gmask, CPOLY_F32 (0, default) or CPOLY_BF16 - lets an autocast
                          * training step receive d/d(offset conv output) in the dtype that conv ran in, without a cast pass */
  int32_t reserved1;
} cpoly_dcn_bwd_args;

int cpoly_abi_version(void);
const char* cpoly_last_error(void);
/* 1 when the running device is sm_100 (tcgen05 available), 0 otherwise, <0 on CUDA error */
int cpoly_device_supports_tcgen05(void);

/* 1 if (args) qualifies for the tcgen05 path: 3x3, stride 1, pad 1, dil 1, dg 1, channels_last x and y
 * (sc == 1), Cin % 64 == 0 (bf16) or % 32 == 0 (fp32), Cout % 16 == 0 and <= 256. */
int cpoly_dcn_tcgen05_eligible(const cpoly_dcn_fwd_args* args);

size_t cpoly_dcn_packed_weight_bytes(int32_t Cin, int32_t Cout, int32_t kh, int32_t kw, int32_t dtype);
/* weight (Cout,Cin,kh,kw) fp32 -> tap-major, K-chunked, 128B-swizzled smem image in `dtype`
 * (bf16, or fp32 rounded to tf32) that the tcgen05 kernels bulk-copy straight into shared memory. */
int cpoly_dcn_pack_weights(const float* weight, void* packed, int32_t Cin, int32_t Cout, int32_t kh,
                           int32_t kw, int32_t dtype, cpoly_stream_t stream);

size_t cpoly_dcn_forward_workspace_bytes(const cpoly_dcn_fwd_args* args);
int cpoly_dcn_forward(const cpoly_dcn_fwd_args* args, cpoly_stream_t stream);

size_t cpoly_dcn_backward_workspace_bytes(const cpoly_dcn_bwd_args* args);
int cpoly_dcn_backward(const cpoly_dcn_bwd_args* args, cpoly_stream_t stream);

/* ------------------------------------------------------------------------------ IDAUp upsampling */

/* The `up_i` of IDAUp (src/lib/models/networks/pose_dla_dcn.py:372-375, applied at :385): depthwise
 * ConvTranspose2d(C, C, 2f, stride=f, padding=f/2, groups=C, bias=False), optionally fused with the
 * `layers[i] + layers[i-1]` of :387:   y = conv_transpose(x, weight) [+ skip].   y and skip are (B,C,H*f,W*f). */
typedef struct {
  int32_t B, C, H, W;   /* input (B,C,H,W) */
  int32_t f;            /* up-sampling factor: kernel 2f, stride f, padding f/2 */
  int32_t dtype;        /* element type of x, skip and y */
  int32_t weight_tap_major; /* 0: weight is the parameter layout (C,1,2f,2f); 1: derived (2f*2f, C) copy (vector loads) */
  int32_t reserved0;
  cpoly_tensor4 x;
  const float* weight;  /* dense fp32 */
  cpoly_tensor4 skip;   /* ptr == NULL => no addend */
  cpoly_tensor4 y;
} cpoly_upsample_args;

int cpoly_idaup_upsample_forward(const cpoly_upsample_args* args, cpoly_stream_t stream);

/* Backward of the same depthwise transposed convolution (what autograd does with nn.ConvTranspose2d in
 * IDAUp during training, src/lib/models/networks/pose_dla_dcn.py:372-375,385; driven by base_trainer.py:97-98).
 * channels_last tensors only; weight / gweight are the tap-major (4f^2, C) fp32 form. */
typedef struct {
  int32_t B, C, H, W, f, dtype;
  cpoly_tensor4 x;        /* (B,C,H,W) forward input                                   */
  cpoly_tensor4 gy;       /* (B,C,H*f,W*f) gradient of the output                      */
  const float* weight;    /* (4f^2, C) fp32                                            */
  cpoly_tensor4 gx;       /* (B,C,H,W) out, or ptr NULL                                */
  float* gweight;         /* (4f^2, C) fp32, or NULL.  Without a workspace it must be ZERO-INITIALISED by the caller
                           * (fp32 atomics, summation order varies at the 1e-7 level) */
  void* workspace;        /* ABI v10: >= cpoly_idaup_upsample_backward_workspace_bytes(&args), 16-byte aligned: the weight
                           * gradient is then summed in a fixed order (bit-reproducible, no atomics); NULL keeps the atomics */
  size_t workspace_bytes;
} cpoly_upsample_bwd_args;

size_t cpoly_idaup_upsample_backward_workspace_bytes(const cpoly_upsample_bwd_args* args);
int cpoly_idaup_upsample_backward(const cpoly_upsample_bwd_args* args, cpoly_stream_t stream);

/* --------------------------------------------------------------------------------------- decode */

typedef struct {
  int32_t B, C, H, W;   /* heat is (B,C,H,W) */
  int32_t n_poly;       /* 2N: channels of the poly head */
  int32_t K;            /* detections per image, 1..1024 and <= H*W */
  int32_t rep;          /* CPOLY_REP_* */
  int32_t heat_is_logit; /* 1: apply the logistic first (fuses detectors/polydet.py:28) */
  int32_t dtype;        /* element type of heat/polys/depth/reg: CPOLY_F32 | CPOLY_BF16 */
  cpoly_tensor4 heat;   /* (B,C,H,W); post-sigmoid unless heat_is_logit */
  cpoly_tensor4 polys;  /* (B,2N,H,W) */
  cpoly_tensor4 depth;  /* (B,1,H,W)  */
  cpoly_tensor4 reg;    /* (B,2,H,W) or ptr == NULL => centres get +0.5 (decode.py:529-530) */
  float* out;           /* (B,K,2N+7) dense fp32: xmin,ymin,xmax,ymax,score,class,x0,y0,…,depth */
  int64_t* out_inds;    /* optional (B,K) flat y*W+x of every detection, or NULL */
  void* workspace;      /* >= cpoly_decode_workspace_bytes(B,C,H,W,K), 256-byte aligned */
  size_t workspace_bytes;
  /* Optional fused `polydet_post_process` back-mapping (src/lib/utils/post_process.py:105-112 calling
   * `transform_preds` / `affine_transform`, src/lib/utils/image.py:19-24,63-66): (B,6) fp64 on the device, the
   * row-major 2x3 inverse affine `get_affine_transform(c, s, 0, (out_w, out_h), inv=1)` of every image.  When
   * given, the two bounding-box corners and all N vertices of every detection are written in ORIGINAL-image
   * coordinates, x' = fp32(t00*x + t01*y + t02) evaluated in fp64 like the reference's np.dot; score, class and
   * pseudo-depth are untouched.  NULL => feature-map coordinates, exactly `polydet_decode`. */
  const double* post_affine;
} cpoly_decode_args;

size_t cpoly_decode_workspace_bytes(int32_t B, int32_t C, int32_t H, int32_t W, int32_t K);
int cpoly_decode(const cpoly_decode_args* args, cpoly_stream_t stream);

/* ---- cpoly_decode without any allocation (ABI v9) ----------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Same inputs, outputs and results as cpoly_decode; `args->workspace` / `workspace_bytes` are ignored: the scratch is a
 * file-scope `static __device__` array of CPOLY_DECODE_STATIC_BYTES (4 MiB) inside the library, which every GPU that
 * loads libcpoly.so holds until the process exits.  Shapes whose cpoly_decode_workspace_bytes(B,C,H,W,K) exceeds that
 * (e.g. B > 15 at the 128x256 x 8-class map of the headline frame) are rejected with CPOLY_EWORKSPACE - use cpoly_decode.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  The first call on each device
 * looks the array's address up (cudaGetSymbolAddress): make that call outside a capture.  All calls on a device share
 * the array: calls that overlap on the GPU (different streams) race - serialise them on one stream or with events. */
#define CPOLY_DECODE_STATIC_BYTES (4u << 20)
size_t cpoly_decode_static_max_bytes(void);
int cpoly_decode_static(const cpoly_decode_args* args, cpoly_stream_t stream);

/* --------------------------------------------------------------------------- instance rasterisation */

/* Depth-ordered rasterisation of the decoded polygons into per-instance masks: the compute half of
 * `Cityscapes.format_and_write_to_cityscapes` (src/lib/datasets/dataset/cityscapes.py:226-279), i.e. what that
 * function does between reading `all_bboxes` and writing the PNG / txt files:
 *   - keep detections with score > thresh (:228), round the vertices to 2 decimals and truncate (:120-121,231-232), stable sort by
 *     pseudo-depth, ties in (class, detection index) order like the reference's per-class dict walk (:239);
 *   - mask = PIL polygon fill  U  radius-2 discs on every point of the Bresenham contour (:253-264) - evaluated
 *     per pixel in closed form (even-odd crossing count on pixel centres U distance test against the exact
 *     Bresenham steps), no scan conversion;
 *   - remove what nearer instances with score >= occluder_score already cover (:265-271);
 *   - keep = class is drawn && count_nonzero(mask) > min_area (:273), confidence = min(1, 1.2*score) (:277).
 * Bit-exact to the reference's PIL result when every truncated vertex lies in [-8192, 8191] (vertices outside
 * are clamped and counted in `status[b]`); W, H <= 8192; K <= 256; N <= 64. */
typedef struct {
  int32_t B, K, N;        /* dets is (B,K,2N+7) fp32 as written by cpoly_decode (with post_affine): box4, score, class, 2N, depth */
  int32_t W, H;           /* mask size = original image size (2048 x 1024 for Cityscapes) */
  double thresh;          /* opt.thresh, compared in fp64 like the reference's Python float */
  double occluder_score;  /* 0.5 */
  int32_t min_area;       /* 100 */
  uint32_t skip_classes;  /* bit c set => 0-based class c is never drawn (the reference's pole / traffic sign / traffic light) */
  const float* dets;
  int32_t* order;         /* (B,K)  slot -> detection index, nearest first; slots >= count[b] are -1 */
  int32_t* count;         /* (B)    instances with score > thresh */
  int16_t* verts;         /* (B,K,N,2) truncated vertices per slot */
  int32_t* area;          /* (B,K)  visible pixels per slot */
  uint8_t* keep;          /* (B,K)  1 => the reference would write this instance */
  double* conf;           /* (B,K)  min(1, 1.2*score) */
  int32_t* status;        /* (B)    vertices that had to be clamped (0 => exact) */
  uint8_t* masks;         /* (B,K,H,W) 0 / 255, slots >= count[b] untouched; NULL => areas / keep only */
  void* workspace;        /* >= cpoly_rasterise_workspace_bytes(B,K) */
  size_t workspace_bytes;
} cpoly_raster_args;

size_t cpoly_rasterise_workspace_bytes(int32_t B, int32_t K);
int cpoly_rasterise(const cpoly_raster_args* args, cpoly_stream_t stream);

/* --------------------------------------------------------------------------- polygon-IoU loss */

/* per-object status (`status` below); anything but OK contributes IoU 0 and no gradient */
#define CPOLY_IOU_OK 0
#define CPOLY_IOU_NO_OUTBOUND 1 /* in-bound but no out-bound crossings: the reference raises IndexError (losses.py:590) */
#define CPOLY_IOU_RUNAWAY 2     /* the reference's walk does not terminate (more than 8N vertices / in-bound list exhausted) */
#define CPOLY_IOU_TOO_MANY 3    /* more than 4N in-bound or out-bound crossings */
#define CPOLY_IOU_NONFINITE 4   /* NaN / Inf in the prediction or target row */

typedef struct {
  int32_t B, M, N;        /* images, object slots per image (max_objs), vertices per polygon (2 <= N <= 64) */
  int32_t H, W;           /* head-map size */
  int32_t dtype;          /* element type of `poly`: CPOLY_F32 | CPOLY_BF16 */
  cpoly_tensor4 poly;     /* (B,2N,H,W) polar poly head: channel 2v = r_v, 2v+1 = theta_v (any strides) */
  const int64_t* ind;     /* (B,M) flat cell index y*W+x of each object (batch['ind']) */
  const uint8_t* mask;    /* (B,M) batch['reg_mask'] != 0 */
  const float* target;    /* (B,M,2N) batch['poly'], polar, angles increasing */
  float* iou;             /* (B,M)   out: intersection / (union + 1e-6) per object, 0 where masked out */
  float* dpred;           /* (B,M,2N) out: d iou / d prediction row (in head-channel order) */
  int32_t* status;        /* (B,M)   out: CPOLY_IOU_* */
  float* stats;           /* (4)     out: {loss = 1 - sum iou/(n+1e-6), -1/(n+1e-6), n = sum mask, objects with status != 0} */
  /* backward only */
  const float* grad_loss; /* (1) upstream gradient of the scalar loss, on the device */
  cpoly_tensor4 grad_poly;/* (B,2N,H,W) fp32, zero-initialised by the caller; receives += grad_loss * d loss / d poly */
} cpoly_poly_iou_args;

int cpoly_poly_iou_forward(const cpoly_poly_iou_args* args, cpoly_stream_t stream);
int cpoly_poly_iou_backward(const cpoly_poly_iou_args* args, cpoly_stream_t stream);

/* ------------------------------------------------------------------- backbone stem and max-pool */

#define CPOLY_STEM_PACKED_BYTES (16 * 224 * 2)

typedef struct {
  int32_t B, H, W;
  int32_t in_dtype;   /* element type of x: CPOLY_F32 | CPOLY_BF16 (rounded to bf16 on load, like autocast) */
  int32_t out_dtype;  /* element type of y */
  int32_t relu;
  cpoly_tensor4 x;    /* (B,3,H,W), any strides */
  const void* weight_packed; /* CPOLY_STEM_PACKED_BYTES from cpoly_stem_pack_weights */
  const float* bias;  /* (16) fp32 (folded BatchNorm shift) or NULL */
  void* y;            /* (B,16,H,W) channels_last contiguous (NHWC physical) */
} cpoly_stem_args;

/* weight: (16,3,7,7) contiguous, fp32 or bf16 -> [16][224] bf16, k = ky*32 + kx*4 + c */
int cpoly_stem_pack_weights(const void* weight, int32_t weight_dtype, void* packed, cpoly_stream_t stream);
int cpoly_stem_conv7x7_forward(const cpoly_stem_args* args, cpoly_stream_t stream);

typedef struct {
  int32_t B, C, H, W; /* input (B,C,H,W) channels_last contiguous; output (B,C,H/2,W/2) channels_last contiguous */
  int32_t dtype;      /* C * sizeof(element) must be a multiple of 16 */
  int32_t reserved0;
  const void* x;
  void* y;
} cpoly_maxpool_args;

int cpoly_maxpool2x2_nhwc_forward(const cpoly_maxpool_args* args, cpoly_stream_t stream);

/* ---- the DCN module's offset / mask producer (ABI v8) ------------------------------------------------------------
 * `conv_offset_mask` of upstream DCNv2's `DCN` (a regular 3x3, stride 1, pad 1 Conv2d(Cin, 27) constructed with the
 * module at src/lib/models/networks/pose_dla_dcn.py:354 / resnet_dcn.py:221-223; `DCN.forward` chunks its output into
 * 18 offsets + 9 mask logits).  om = conv3x3(x, W) + bias, 32 output channels (27 + 5 zero), NHWC, fp32 or bf16 - the
 * layout cpoly_dcn_forward reads with offset = om[:, :18], mask = om[:, 18:27] (mask_is_logit = 1).
 * x: dense channels_last bf16 (B,Cin,H,W), Cin a multiple of 64. */
typedef struct {
  int32_t B, Cin, H, W;
  int32_t om_dtype;          /* CPOLY_F32 or CPOLY_BF16 */
  int32_t reserved0;
  const void* x;             /* bf16, dense NHWC, 16-byte aligned */
  const void* weight_packed; /* cpoly_offset_conv_pack_weights image */
  const float* bias;         /* 32 floats (27 + zeros) */
  void* om;                  /* dense NHWC (B,H,W,32), 16-byte aligned */
} cpoly_offset_conv_args;

size_t cpoly_offset_conv_packed_bytes(int32_t Cin);   /* 0 when the shape has no fast path */
/* weight: fp32 contiguous (n_out <= 32, Cin, 3, 3) */
int cpoly_offset_conv_pack_weights(const float* weight, void* packed, int32_t Cin, int32_t n_out, cpoly_stream_t stream);
int cpoly_offset_conv_forward(const cpoly_offset_conv_args* args, cpoly_stream_t stream);

/* ---- instance masks of the IDD / KITTI writers (ABI v10) ---------------------------------------------------------------
 * The mask construction inside `IDD.format_and_write_to_IDD` (src/lib/datasets/dataset/IDD.py:123-170) and
 * `KITTIPOLY.format_and_write_to_kitti` (src/lib/datasets/dataset/kitti_poly.py:95-136), i.e. what those functions do
 * between reading `all_bboxes` and writing the PNG / txt files:
 *   - keep detections with score >= thresh (IDD :141, inclusive = 1) or score > thresh (KITTI :113, inclusive = 0);
 *     file slot = the reference's counter: classes ascending, detection order inside a class;
 *   - vertices int(float("{:.2f}".format(v))) (:62-63,159-160); per class, stable depth order;
 *   - mask = PIL `polygon(points, outline=0, fill=255)` (scan-line fill of Pillow's polygon_generic minus the line8
 *     Bresenham outline) times the class's running `to_remove_mask`, which instances with score >= occluder_score clear
 *     with `polygon(points, outline=0, fill=0)` (the bare fill).
 * Pillow's algorithms are restated in oracle/raster_fill_oracle.py and pinned to the unmodified reference writers run with
 * Pillow 12.2 (tests/golden/fillraster_*.npz).  Vertices outside [-8192, 8191] are clamped and counted in status[b].
 * W, H <= 8192; K <= 256; N <= 64; B * K <= 65535. */
typedef struct {
  int32_t B, K, N;        /* dets is (B,K,2N+7) fp32 as written by cpoly_decode (with post_affine) */
  int32_t W, H;           /* mask size = image size */
  int32_t inclusive;      /* 1: score >= thresh (IDD), 0: score > thresh (KITTI) */
  double thresh;          /* opt.thresh, compared in fp64 like the reference's Python float */
  double occluder_score;  /* 0.5 */
  const float* dets;
  int32_t* count;         /* (B)   instances written */
  int32_t* slot_det;      /* (B,K) file slot -> detection index; -1 past count[b] */
  int32_t* area;          /* (B,K) visible pixels per slot */
  int32_t* status;        /* (B)   vertices that had to be clamped (0 => exact) */
  uint8_t* masks;         /* (B,K,H,W) 0 / 255 in file-slot order, slots >= count[b] untouched; NULL => count / slot_det only */
  void* workspace;        /* >= cpoly_rasterise_fill_workspace_bytes(B,K,N), 256-byte aligned */
  size_t workspace_bytes;
} cpoly_raster_fill_args;

size_t cpoly_rasterise_fill_workspace_bytes(int32_t B, int32_t K, int32_t N);
int cpoly_rasterise_fill(const cpoly_raster_fill_args* args, cpoly_stream_t stream);

/* ---- training-mode BatchNorm2d (+ fused ReLU) on dense NHWC activations (ABI v10) -----------------------------------
 * Replaces `nn.BatchNorm2d(planes, momentum=BN_MOMENTUM)` in training mode at the sites of the DLA-34 / DLAUp / IDAUp
 * network (src/lib/models/networks/pose_dla_dcn.py:31-47 BasicBlock, :143-160 Root, :231-236 base_layer, :247-252 conv
 * levels) and the `actf = Sequential(BatchNorm2d, ReLU)` of DeformConv (:345-359; relu = 1 fuses its ReLU).  PyTorch
 * semantics: batch statistics with the biased variance for the normalisation, running_mean / running_var updated in
 * place with `momentum` and the UNBIASED variance; backward returns the input gradient, gweight = sum(g * xhat) and
 * gbias = sum(g) with g = gy * [y > 0] when relu.  x, y, gy, gx: dense NHWC (N = B*H*W rows of C channels), fp32 or
 * bf16; weight / bias / running stats / save / gweight / gbias fp32.  C / (16 / sizeof(element)) must be a power of two
 * <= 256 (cpoly_bn_eligible); every reduction runs in a fixed order (bit-reproducible). */
typedef struct {
  int64_t N;                 /* rows (B*H*W) */
  int32_t C;
  int32_t dtype;             /* CPOLY_F32 or CPOLY_BF16 */
  int32_t relu;
  int32_t reserved0;
  const void* x;
  void* y;
  const float* weight;       /* gamma, NULL = 1 */
  const float* bias;         /* beta, NULL = 0 */
  float* running_mean;       /* updated in place, may be NULL */
  float* running_var;
  float momentum, eps;
  float* save;               /* out: [4][C] = mean, invstd, scale, shift (kept for the backward) */
  void* workspace;           /* >= cpoly_bn_workspace_bytes(C), 16-byte aligned */
  size_t workspace_bytes;
} cpoly_bn_fwd_args;

typedef struct {
  int64_t N;
  int32_t C;
  int32_t dtype;
  int32_t relu;
  int32_t reserved0;
  const void* x;             /* the forward's input */
  const void* gy;
  void* gx;                  /* may be NULL (no input gradient wanted) */
  const float* save;         /* the forward's save */
  float* gweight;            /* [C], may be NULL */
  float* gbias;              /* [C], may be NULL */
  void* workspace;
  size_t workspace_bytes;
} cpoly_bn_bwd_args;

int cpoly_bn_eligible(int32_t C, int32_t dtype);
size_t cpoly_bn_workspace_bytes(int32_t C);
int cpoly_bn_train_forward(const cpoly_bn_fwd_args* args, cpoly_stream_t stream);
int cpoly_bn_train_backward(const cpoly_bn_bwd_args* args, cpoly_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* CPOLY_H_ */
