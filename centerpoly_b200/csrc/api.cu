// extern "C" surface of libcpoly.so — see include/cpoly.h for the contract and the reference
// interfaces each entry point replaces.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace cpoly {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

static int validate_dcn(const cpoly_dcn_fwd_args& a, bool need_y) {
  CPOLY_REQUIRE(a.B >= 0 && a.Cin > 0 && a.Cout > 0 && a.H > 0 && a.W > 0, CPOLY_EINVAL, "bad tensor dims");
  CPOLY_REQUIRE(a.kh > 0 && a.kw > 0 && a.stride_h > 0 && a.stride_w > 0 && a.dil_h > 0 && a.dil_w > 0 &&
                    a.pad_h >= 0 && a.pad_w >= 0, CPOLY_EINVAL, "bad convolution geometry");
  CPOLY_REQUIRE(a.deformable_groups > 0 && a.Cin % a.deformable_groups == 0, CPOLY_EINVAL,
                "in_channels (%d) must be divisible by deformable_groups (%d)", a.Cin, a.deformable_groups);
  CPOLY_REQUIRE(a.dtype == CPOLY_F32 || a.dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a.dtype);
  CPOLY_REQUIRE(a.om_dtype == CPOLY_F32 || a.om_dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported offset/mask dtype %d", a.om_dtype);
  CPOLY_REQUIRE(a.x.ptr && a.offset.ptr && a.mask.ptr && (a.weight || a.weight_packed), CPOLY_EINVAL, "NULL input pointer");
  CPOLY_REQUIRE(!need_y || a.y.ptr, CPOLY_EINVAL, "NULL output pointer");
  const DcnGeom g = make_geom(a);
  CPOLY_REQUIRE(g.Ho > 0 && g.Wo > 0, CPOLY_EINVAL, "empty output (%d x %d)", g.Ho, g.Wo);
  return CPOLY_OK;
}

// Forward sampler of the tcgen05 path: "gather" (dcn_tc.cu, corners gathered from global memory through L1; default)
// or "window" (dcn_tc_win.cu, TMA-staged shared-memory window), selected with CPOLY_TC_SAMPLER=window.  The two are
// within +-5 % of each other per layer (both saturate the same LSU/shared-memory data pipe).
static bool want_window_sampler() {
  const char* e = getenv("CPOLY_TC_SAMPLER");
  return e && strcmp(e, "window") == 0;
}
static int dcn_forward_tc_any(const cpoly_dcn_fwd_args& a, cudaStream_t s) {
  if (want_window_sampler() && dcn_tc_win_eligible(a)) return dcn_forward_tc_win(a, s);
  return dcn_forward_tc(a, s);
}

}  // namespace cpoly

using namespace cpoly;

extern "C" {

int cpoly_abi_version(void) { return CPOLY_ABI_VERSION; }

const char* cpoly_last_error(void) { return g_error; }

int cpoly_device_supports_tcgen05(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return -1;
  return major == 10 ? 1 : 0;
}

int cpoly_dcn_tcgen05_eligible(const cpoly_dcn_fwd_args* args) {
  if (!args) return 0;
  return dcn_tc_eligible(*args) ? 1 : 0;
}

size_t cpoly_dcn_packed_weight_bytes(int32_t Cin, int32_t Cout, int32_t kh, int32_t kw, int32_t dtype) {
  return dcn_tc_packed_bytes(Cin, Cout, kh, kw, dtype);
}

int cpoly_dcn_pack_weights(const float* weight, void* packed, int32_t Cin, int32_t Cout, int32_t kh, int32_t kw,
                           int32_t dtype, cpoly_stream_t stream) {
  CPOLY_REQUIRE(weight && packed, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE(dcn_tc_packed_bytes(Cin, Cout, kh, kw, dtype) > 0, CPOLY_EUNSUPPORTED,
                "shape Cin=%d Cout=%d k=%dx%d dtype=%d has no packed layout", Cin, Cout, kh, kw, dtype);
  return dcn_tc_pack_weights(weight, packed, Cin, Cout, kh, kw, dtype, reinterpret_cast<cudaStream_t>(stream));
}

size_t cpoly_dcn_forward_workspace_bytes(const cpoly_dcn_fwd_args* args) {
  if (!args || !dcn_tc_eligible(*args)) return 256;
  const size_t g = dcn_tc_forward_workspace(*args), w = dcn_tc_win_eligible(*args) ? dcn_tc_win_forward_workspace(*args) : 0;
  return g > w ? g : w;
}

int cpoly_dcn_forward(const cpoly_dcn_fwd_args* args, cpoly_stream_t stream) {
  CPOLY_REQUIRE(args, CPOLY_EINVAL, "NULL args");
  if (int rc = validate_dcn(*args, true)) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool tc_ok = args->weight_packed && dcn_tc_eligible(*args);
  if (args->path == CPOLY_PATH_TCGEN05) {
    CPOLY_REQUIRE(tc_ok, CPOLY_EUNSUPPORTED, "tcgen05 path requested but the problem does not qualify (see cpoly_dcn_tcgen05_eligible)");
    return dcn_forward_tc_any(*args, s);
  }
  if (args->path == CPOLY_PATH_AUTO && tc_ok) return dcn_forward_tc_any(*args, s);
  CPOLY_REQUIRE(args->weight, CPOLY_EINVAL, "the generic path needs the dense fp32 weights");
  return dcn_forward_simt(*args, s);
}

static bool use_tc_wgrad(const cpoly_dcn_bwd_args& a) {
  return a.gweight && a.f.path != CPOLY_PATH_SIMT && dcn_tc_wgrad_eligible(a);
}
static bool use_tc_dgrad(const cpoly_dcn_bwd_args& a) {
  return (a.gx.ptr || a.goffset.ptr || a.gmask.ptr) && a.f.path != CPOLY_PATH_SIMT && dcn_tc_dgrad_eligible(a);
}
static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

size_t cpoly_dcn_backward_workspace_bytes(const cpoly_dcn_bwd_args* args) {
  if (!args) return 0;
  const size_t simt = dcn_backward_simt_workspace(*args);
  const size_t tc = (use_tc_wgrad(*args) ? align256(dcn_tc_wgrad_workspace(*args)) : 0) +
                    (use_tc_dgrad(*args) ? align256(dcn_tc_dgrad_workspace(*args)) : 0);
  return simt > tc ? simt : tc;
}

int cpoly_dcn_backward(const cpoly_dcn_bwd_args* args, cpoly_stream_t stream) {
  CPOLY_REQUIRE(args, CPOLY_EINVAL, "NULL args");
  if (int rc = validate_dcn(args->f, false)) return rc;
  CPOLY_REQUIRE(args->gout.ptr && args->f.weight, CPOLY_EINVAL, "NULL grad_output / weight");
  CPOLY_REQUIRE(args->gom_dtype == CPOLY_F32 || args->gom_dtype == CPOLY_BF16, CPOLY_EINVAL, "bad gom_dtype %d", args->gom_dtype);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool tcw = use_tc_wgrad(*args), tcd = use_tc_dgrad(*args);
  CPOLY_REQUIRE(args->f.path != CPOLY_PATH_TCGEN05 || ((tcw || !args->gweight) && (tcd || !(args->gx.ptr || args->goffset.ptr || args->gmask.ptr))),
                CPOLY_EUNSUPPORTED, "tcgen05 path requested but the backward problem does not qualify");
  cpoly_dcn_bwd_args rest = *args;   // what remains for the generic path
  char* ws = reinterpret_cast<char*>(args->workspace);
  size_t left = args->workspace_bytes;
  if (tcw) {   // weight gradient on the tensor cores
    const size_t need = align256(dcn_tc_wgrad_workspace(*args));
    CPOLY_REQUIRE(ws && left >= need, CPOLY_EWORKSPACE, "dcn backward workspace too small");
    if (int rc = dcn_wgrad_tc(*args, ws, need, s)) return rc;
    ws += need; left -= need;
    rest.gweight = nullptr;
  }
  if (tcd) {   // data / coordinate gradients on the tensor cores
    const size_t need = align256(dcn_tc_dgrad_workspace(*args));
    CPOLY_REQUIRE(ws && left >= need, CPOLY_EWORKSPACE, "dcn backward workspace too small");
    if (int rc = dcn_dgrad_tc(*args, ws, need, s)) return rc;
    rest.gx.ptr = nullptr; rest.goffset.ptr = nullptr; rest.gmask.ptr = nullptr;
  }
  if (!tcw && !tcd) return dcn_backward_simt(*args, s);
  if (rest.gweight || rest.gbias || rest.gx.ptr || rest.goffset.ptr || rest.gmask.ptr) return dcn_backward_simt(rest, s);
  return CPOLY_OK;
}

size_t cpoly_decode_workspace_bytes(int32_t B, int32_t C, int32_t H, int32_t W, int32_t K) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || K <= 0) return 256;
  return decode_workspace_bytes(B, C, H, W, K);
}

int cpoly_decode(const cpoly_decode_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->C > 0 && a->H > 0 && a->W > 0, CPOLY_EINVAL, "bad heat dims");
  CPOLY_REQUIRE(a->n_poly > 0 && a->n_poly % 2 == 0, CPOLY_EINVAL, "poly head must have an even channel count, got %d", a->n_poly);
  CPOLY_REQUIRE(a->K >= 1 && a->K <= 1024, CPOLY_EINVAL, "K must be in [1, 1024], got %d", a->K);
  CPOLY_REQUIRE((int64_t)a->K <= (int64_t)a->H * a->W, CPOLY_EINVAL,
                "K (%d) exceeds H*W (%d): the reference's per-class torch.topk would raise", a->K, a->H * a->W);
  CPOLY_REQUIRE((int64_t)a->C * a->H * a->W < 0xffffffffLL, CPOLY_EINVAL, "C*H*W must fit in 32 bits");
  CPOLY_REQUIRE(a->rep >= CPOLY_REP_CARTESIAN && a->rep <= CPOLY_REP_POLAR_FIXED, CPOLY_EINVAL, "bad rep %d", a->rep);
  CPOLY_REQUIRE(a->dtype == CPOLY_F32 || a->dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a->dtype);
  CPOLY_REQUIRE(a->heat.ptr && a->polys.ptr && a->depth.ptr && a->out, CPOLY_EINVAL, "NULL pointer");
  if (a->B == 0) return CPOLY_OK;
  return decode_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

size_t cpoly_rasterise_fill_workspace_bytes(int32_t B, int32_t K, int32_t N) { return raster_fill_workspace_bytes(B, K, N); }

int cpoly_rasterise_fill(const cpoly_raster_fill_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->K >= 1 && a->K <= 256 && a->N >= 3 && a->N <= 64, CPOLY_EINVAL, "need 1 <= K <= 256 and 3 <= N <= 64 (got K=%d N=%d)", a->K, a->N);
  CPOLY_REQUIRE(a->W >= 1 && a->W <= 8192 && a->H >= 1 && a->H <= 8192, CPOLY_EINVAL, "mask size must be within 8192 x 8192");
  CPOLY_REQUIRE((int64_t)a->B * a->K <= 65535, CPOLY_EINVAL, "B * K must not exceed 65535");
  CPOLY_REQUIRE(a->dets && a->count && a->slot_det && a->area && a->status, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE(!a->masks || (reinterpret_cast<uintptr_t>(a->masks) & 15) == 0, CPOLY_EINVAL, "masks must be 16-byte aligned");
  CPOLY_REQUIRE(a->workspace && a->workspace_bytes >= raster_fill_workspace_bytes(a->B, a->K, a->N) && (reinterpret_cast<uintptr_t>(a->workspace) & 255) == 0,
                CPOLY_EWORKSPACE, "rasterise_fill workspace too small or misaligned: need %zu bytes", raster_fill_workspace_bytes(a->B, a->K, a->N));
  if (a->B == 0) return CPOLY_OK;
  return raster_fill_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_bn_eligible(int32_t C, int32_t dtype) { return bn_eligible(C, dtype) ? 1 : 0; }
size_t cpoly_bn_workspace_bytes(int32_t C) { return bn_workspace_bytes(C); }

int cpoly_bn_train_forward(const cpoly_bn_fwd_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->N >= 0 && a->C > 0, CPOLY_EINVAL, "bad dims");
  CPOLY_REQUIRE(a->dtype == CPOLY_F32 || a->dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a->dtype);
  CPOLY_REQUIRE(a->x && a->y && a->save, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE(((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->y)) & 15) == 0, CPOLY_EINVAL, "x / y must be 16-byte aligned");
  CPOLY_REQUIRE(a->N > 1, CPOLY_EINVAL, "batch norm in training mode needs more than one value per channel");
  CPOLY_REQUIRE(a->workspace && a->workspace_bytes >= bn_workspace_bytes(a->C) && (reinterpret_cast<uintptr_t>(a->workspace) & 15) == 0, CPOLY_EWORKSPACE,
                "batch norm workspace too small or misaligned: need %zu bytes", bn_workspace_bytes(a->C));
  return bn_train_forward_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_bn_train_backward(const cpoly_bn_bwd_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->N > 0 && a->C > 0, CPOLY_EINVAL, "bad dims");
  CPOLY_REQUIRE(a->dtype == CPOLY_F32 || a->dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a->dtype);
  CPOLY_REQUIRE(a->x && a->gy && a->save, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE(((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->gy) | reinterpret_cast<uintptr_t>(a->gx)) & 15) == 0, CPOLY_EINVAL,
                "x / gy / gx must be 16-byte aligned");
  CPOLY_REQUIRE(a->workspace && a->workspace_bytes >= bn_workspace_bytes(a->C) && (reinterpret_cast<uintptr_t>(a->workspace) & 15) == 0, CPOLY_EWORKSPACE,
                "batch norm workspace too small or misaligned: need %zu bytes", bn_workspace_bytes(a->C));
  return bn_train_backward_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

size_t cpoly_decode_static_max_bytes(void) { return CPOLY_DECODE_STATIC_BYTES; }

int cpoly_decode_static(const cpoly_decode_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->C > 0 && a->H > 0 && a->W > 0 && a->K >= 1, CPOLY_EINVAL, "bad heat dims");
  const size_t need = cpoly_decode_workspace_bytes(a->B, a->C, a->H, a->W, a->K);
  CPOLY_REQUIRE(need <= (size_t)CPOLY_DECODE_STATIC_BYTES, CPOLY_EWORKSPACE,
                "cpoly_decode_static: B=%d C=%d H=%d W=%d needs %zu bytes of scratch, the built-in array holds %zu; use cpoly_decode "
                "with a caller-allocated workspace", a->B, a->C, a->H, a->W, need, (size_t)CPOLY_DECODE_STATIC_BYTES);
  cpoly_decode_args b = *a;
  void* scratch = nullptr;
  if (int rc = decode_static_scratch(&scratch)) return rc;
  b.workspace = scratch;
  b.workspace_bytes = CPOLY_DECODE_STATIC_BYTES;
  return cpoly_decode(&b, stream);
}

size_t cpoly_rasterise_workspace_bytes(int32_t B, int32_t K) {
  if (B <= 0 || K <= 0) return 256;
  return raster_workspace_bytes(B, K);
}

int cpoly_rasterise(const cpoly_raster_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->K >= 1 && a->K <= 256, CPOLY_EINVAL, "K must be in [1, 256], got %d", a->K);
  CPOLY_REQUIRE(a->N >= 3 && a->N <= 64, CPOLY_EINVAL, "N (vertices) must be in [3, 64], got %d", a->N);
  CPOLY_REQUIRE(a->W >= 1 && a->W <= 8192 && a->H >= 1 && a->H <= 8192, CPOLY_EINVAL, "mask size must be in [1, 8192]^2, got %dx%d", a->W, a->H);
  CPOLY_REQUIRE(a->dets && a->order && a->count && a->verts && a->area && a->keep && a->conf && a->status, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE((reinterpret_cast<uintptr_t>(a->verts) & 3) == 0 && (reinterpret_cast<uintptr_t>(a->masks) & 7) == 0 &&
                (reinterpret_cast<uintptr_t>(a->workspace) & 15) == 0, CPOLY_EINVAL, "verts / masks / workspace must be 4 / 8 / 16-byte aligned");
  if (a->B == 0) return CPOLY_OK;
  CPOLY_REQUIRE(a->workspace && a->workspace_bytes >= raster_workspace_bytes(a->B, a->K), CPOLY_EWORKSPACE, "raster workspace too small");
  return raster_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_idaup_upsample_forward(const cpoly_upsample_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->C > 0 && a->H > 0 && a->W > 0 && a->f >= 1, CPOLY_EINVAL, "bad upsample dims");
  CPOLY_REQUIRE(a->dtype == CPOLY_F32 || a->dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a->dtype);
  CPOLY_REQUIRE(a->x.ptr && a->weight && a->y.ptr, CPOLY_EINVAL, "NULL pointer");
  if (a->B == 0) return CPOLY_OK;
  return idaup_upsample_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

static int poly_iou_check(const cpoly_poly_iou_args* a) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->M >= 1 && a->H > 0 && a->W > 0, CPOLY_EINVAL, "bad poly-IoU dims");
  CPOLY_REQUIRE(a->N >= 2 && a->N <= 64, CPOLY_EINVAL, "N (vertices) must be in [2, 64], got %d", a->N);
  CPOLY_REQUIRE((int64_t)a->B * a->M < (1 << 28), CPOLY_EINVAL, "too many object slots");
  CPOLY_REQUIRE(a->ind && a->mask && a->dpred && a->stats, CPOLY_EINVAL, "NULL pointer");
  return CPOLY_OK;
}

int cpoly_poly_iou_forward(const cpoly_poly_iou_args* a, cpoly_stream_t stream) {
  if (int rc = poly_iou_check(a)) return rc;
  CPOLY_REQUIRE(a->dtype == CPOLY_F32 || a->dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a->dtype);
  CPOLY_REQUIRE(a->poly.ptr && a->target && a->iou && a->status, CPOLY_EINVAL, "NULL pointer");
  if (a->B == 0) return CPOLY_OK;
  return poly_iou_forward_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_poly_iou_backward(const cpoly_poly_iou_args* a, cpoly_stream_t stream) {
  if (int rc = poly_iou_check(a)) return rc;
  CPOLY_REQUIRE(a->grad_loss && a->grad_poly.ptr, CPOLY_EINVAL, "NULL pointer");
  if (a->B == 0) return CPOLY_OK;
  return poly_iou_backward_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_stem_pack_weights(const void* weight, int32_t weight_dtype, void* packed, cpoly_stream_t stream) {
  CPOLY_REQUIRE(weight && packed, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE(weight_dtype == CPOLY_F32 || weight_dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", weight_dtype);
  return stem_pack_weights_launch(weight, weight_dtype, packed, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_stem_conv7x7_forward(const cpoly_stem_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->B <= 65535 && a->H > 0 && a->W > 0, CPOLY_EINVAL, "bad stem dims");
  CPOLY_REQUIRE((a->in_dtype == CPOLY_F32 || a->in_dtype == CPOLY_BF16) && (a->out_dtype == CPOLY_F32 || a->out_dtype == CPOLY_BF16),
                CPOLY_EINVAL, "unsupported dtype");
  CPOLY_REQUIRE(a->x.ptr && a->weight_packed && a->y, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE((reinterpret_cast<uintptr_t>(a->y) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->weight_packed) & 3) == 0, CPOLY_EINVAL, "y must be 16-byte aligned");
  if (a->B == 0) return CPOLY_OK;
  return stem_conv_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_maxpool2x2_nhwc_forward(const cpoly_maxpool_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->C > 0 && a->H >= 2 && a->W >= 2, CPOLY_EINVAL, "bad max-pool dims");
  CPOLY_REQUIRE(a->dtype == CPOLY_F32 || a->dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a->dtype);
  CPOLY_REQUIRE((a->C * (a->dtype == CPOLY_BF16 ? 2 : 4)) % 16 == 0, CPOLY_EUNSUPPORTED, "C * element size must be a multiple of 16 bytes");
  CPOLY_REQUIRE(a->x && a->y && (reinterpret_cast<uintptr_t>(a->x) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->y) & 15) == 0, CPOLY_EINVAL, "x / y must be 16-byte aligned");
  return maxpool2x2_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

size_t cpoly_idaup_upsample_backward_workspace_bytes(const cpoly_upsample_bwd_args* a) {
  if (!a || a->C <= 0 || a->f < 2 || a->B <= 0 || a->H <= 0 || a->W <= 0 || !a->gweight) return 256;
  return idaup_upsample_backward_workspace(*a);
}

int cpoly_idaup_upsample_backward(const cpoly_upsample_bwd_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->C > 0 && a->H > 0 && a->W > 0 && a->f >= 2, CPOLY_EINVAL, "bad upsample dims");
  CPOLY_REQUIRE(a->dtype == CPOLY_F32 || a->dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported dtype %d", a->dtype);
  CPOLY_REQUIRE(a->x.ptr && a->gy.ptr && a->weight, CPOLY_EINVAL, "NULL pointer");
  if (a->B == 0) return CPOLY_OK;
  return idaup_upsample_backward_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

size_t cpoly_offset_conv_packed_bytes(int32_t Cin) { return Cin > 0 ? offset_conv_packed_bytes(Cin) : 0; }

int cpoly_offset_conv_pack_weights(const float* weight, void* packed, int32_t Cin, int32_t n_out, cpoly_stream_t stream) {
  CPOLY_REQUIRE(weight && packed, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE(offset_conv_packed_bytes(Cin) > 0 && n_out >= 1 && n_out <= 32, CPOLY_EUNSUPPORTED,
                "offset conv: Cin (%d) must be a multiple of 64 and n_out (%d) <= 32", Cin, n_out);
  return offset_conv_pack_launch(weight, packed, Cin, n_out, reinterpret_cast<cudaStream_t>(stream));
}

int cpoly_offset_conv_forward(const cpoly_offset_conv_args* a, cpoly_stream_t stream) {
  CPOLY_REQUIRE(a, CPOLY_EINVAL, "NULL args");
  CPOLY_REQUIRE(a->B >= 0 && a->H > 0 && a->W > 0 && offset_conv_packed_bytes(a->Cin) > 0, CPOLY_EUNSUPPORTED, "offset conv: bad dims / Cin %% 64 != 0");
  CPOLY_REQUIRE(a->om_dtype == CPOLY_F32 || a->om_dtype == CPOLY_BF16, CPOLY_EINVAL, "unsupported om dtype %d", a->om_dtype);
  CPOLY_REQUIRE(a->x && a->weight_packed && a->bias && a->om, CPOLY_EINVAL, "NULL pointer");
  CPOLY_REQUIRE(((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->om) | reinterpret_cast<uintptr_t>(a->weight_packed)) & 15) == 0,
                CPOLY_EINVAL, "x / om / weight_packed must be 16-byte aligned");
  CPOLY_REQUIRE((int64_t)a->B * a->H * a->W < (1LL << 31), CPOLY_EINVAL, "too many pixels");
  if (a->B == 0) return CPOLY_OK;
  return offset_conv_launch(*a, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
