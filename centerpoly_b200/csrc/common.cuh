// Shared helpers for libcpoly.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/cpoly.h"

namespace cpoly {

// ---- per-thread error string -------------------------------------------------------------------
void set_error(const char* fmt, ...);

#define CPOLY_REQUIRE(cond, code, ...)          \
  do {                                          \
    if (!(cond)) {                              \
      ::cpoly::set_error(__VA_ARGS__);          \
      return (code);                            \
    }                                           \
  } while (0)

#define CPOLY_CUDA_OK(expr)                                                               \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ::cpoly::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return CPOLY_ECUDA;                                                                 \
    }                                                                                     \
  } while (0)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("launch of %s failed: %s", what, cudaGetErrorString(e));
    return CPOLY_ECUDA;
  }
  return CPOLY_OK;
}

// ---- element access ----------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float ld_as_float(const T* p);
template <> __device__ __forceinline__ float ld_as_float<float>(const float* p) { return __ldg(p); }
template <> __device__ __forceinline__ float ld_as_float<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}
template <typename T> __device__ __forceinline__ void st_from_float(T* p, float v);
template <> __device__ __forceinline__ void st_from_float<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void st_from_float<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16_rn(v);
}

// read element `idx` of a buffer whose element type is only known at run time (offset / mask)
__device__ __forceinline__ float ld_dyn(const void* base, int64_t idx, int dtype) {
  return dtype == CPOLY_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx])
                             : __ldg(reinterpret_cast<const float*>(base) + idx);
}

// write element `idx` of a buffer whose element type is only known at run time (goffset / gmask)
__device__ __forceinline__ void st_dyn(void* base, int64_t idx, float v, int dtype) {
  if (dtype == CPOLY_BF16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16(v);
  else reinterpret_cast<float*>(base)[idx] = v;
}

__device__ __forceinline__ float sigmoidf_accurate(float v) { return 1.0f / (1.0f + expf(-v)); }

// ---- geometry shared by every DCN kernel ---------------------------------------------------------
struct DcnGeom {
  int B, Cin, Cout, H, W, Ho, Wo;
  int kh, kw, sh, sw, ph, pw, dh, dw, dg;
  int om_dtype, mask_is_logit;
  int gom_dtype;   // backward only: element type of goffset / gmask
  const float* offset_bias;
  const float* mask_bias;
};

inline DcnGeom make_geom(const cpoly_dcn_fwd_args& a) {
  DcnGeom g;
  g.B = a.B; g.Cin = a.Cin; g.Cout = a.Cout; g.H = a.H; g.W = a.W;
  g.kh = a.kh; g.kw = a.kw; g.sh = a.stride_h; g.sw = a.stride_w; g.ph = a.pad_h; g.pw = a.pad_w;
  g.dh = a.dil_h; g.dw = a.dil_w; g.dg = a.deformable_groups;
  g.Ho = (a.H + 2 * a.pad_h - (a.dil_h * (a.kh - 1) + 1)) / a.stride_h + 1;
  g.Wo = (a.W + 2 * a.pad_w - (a.dil_w * (a.kw - 1) + 1)) / a.stride_w + 1;
  g.om_dtype = a.om_dtype; g.mask_is_logit = a.mask_is_logit; g.gom_dtype = CPOLY_F32;
  g.offset_bias = a.offset_bias; g.mask_bias = a.mask_bias;
  return g;
}

// One deformed tap of one output pixel: four bilinear corners (element offset into x WITHOUT the
// channel term, or -1 when that corner is outside the image) and the bilinear weights.
// DCNv2 rule: the whole sample is dead unless -1 < p_y < H and -1 < p_x < W.
struct TapSample {
  int64_t off[4];  // (y0,x0) (y0,x1) (y1,x0) (y1,x1)
  float w[4];
  float ly, lx;
  bool alive;
};

__device__ __forceinline__ TapSample make_tap(float py, float px, int H, int W, int64_t base, int64_t sh,
                                              int64_t sw) {
  TapSample t;
  t.alive = (py > -1.0f) && (px > -1.0f) && (py < (float)H) && (px < (float)W);
  const float fy = floorf(py), fx = floorf(px);
  const int y0 = (int)fy, x0 = (int)fx, y1 = y0 + 1, x1 = x0 + 1;
  t.ly = py - fy;
  t.lx = px - fx;
  const float hy = 1.0f - t.ly, hx = 1.0f - t.lx;
  const bool vy0 = t.alive && y0 >= 0, vy1 = t.alive && y1 <= H - 1;
  const bool vx0 = x0 >= 0, vx1 = x1 <= W - 1;
  t.off[0] = (vy0 && vx0) ? base + y0 * sh + x0 * sw : -1;
  t.off[1] = (vy0 && vx1) ? base + y0 * sh + x1 * sw : -1;
  t.off[2] = (vy1 && vx0) ? base + y1 * sh + x0 * sw : -1;
  t.off[3] = (vy1 && vx1) ? base + y1 * sh + x1 * sw : -1;
  t.w[0] = hy * hx; t.w[1] = hy * t.lx; t.w[2] = t.ly * hx; t.w[3] = t.ly * t.lx;
  return t;
}

}  // namespace cpoly

// ---- internal entry points (one per translation unit) ---------------------------------------------
namespace cpoly {
int dcn_forward_simt(const cpoly_dcn_fwd_args& a, cudaStream_t stream);
int dcn_backward_simt(const cpoly_dcn_bwd_args& a, cudaStream_t stream);
size_t dcn_backward_simt_workspace(const cpoly_dcn_bwd_args& a);
bool dcn_tc_eligible(const cpoly_dcn_fwd_args& a);
size_t dcn_tc_packed_bytes(int Cin, int Cout, int kh, int kw, int dtype);
int dcn_tc_pack_weights(const float* weight, void* packed, int Cin, int Cout, int kh, int kw, int dtype, cudaStream_t stream);
int dcn_forward_tc(const cpoly_dcn_fwd_args& a, cudaStream_t stream);
size_t dcn_tc_forward_workspace(const cpoly_dcn_fwd_args& a);
bool dcn_tc_win_eligible(const cpoly_dcn_fwd_args& a);
size_t dcn_tc_win_forward_workspace(const cpoly_dcn_fwd_args& a);
int dcn_forward_tc_win(const cpoly_dcn_fwd_args& a, cudaStream_t stream);
bool dcn_tc_wgrad_eligible(const cpoly_dcn_bwd_args& a);
size_t dcn_tc_wgrad_workspace(const cpoly_dcn_bwd_args& a);
int dcn_wgrad_tc(const cpoly_dcn_bwd_args& a, void* workspace, size_t workspace_bytes, cudaStream_t stream);
bool dcn_tc_dgrad_eligible(const cpoly_dcn_bwd_args& a);
size_t dcn_tc_dgrad_workspace(const cpoly_dcn_bwd_args& a);
int dcn_dgrad_tc(const cpoly_dcn_bwd_args& a, void* workspace, size_t workspace_bytes, cudaStream_t stream);
bool dcn_col2im_eligible(const cpoly_dcn_bwd_args& a);
size_t dcn_col2im_gcol_bytes(const cpoly_dcn_fwd_args& f);
size_t dcn_col2im_workspace(const cpoly_dcn_bwd_args& a);
int dcn_col2im_launch(const cpoly_dcn_bwd_args& a, void* workspace, size_t workspace_bytes, cudaStream_t stream);
int idaup_upsample_launch(const cpoly_upsample_args& a, cudaStream_t stream);
size_t idaup_upsample_backward_workspace(const cpoly_upsample_bwd_args& a);
int idaup_upsample_backward_launch(const cpoly_upsample_bwd_args& a, cudaStream_t stream);
size_t decode_workspace_bytes(int B, int C, int H, int W, int K);
int decode_launch(const cpoly_decode_args& a, cudaStream_t stream);
int decode_static_scratch(void** ptr);
size_t raster_fill_workspace_bytes(int B, int K, int N);
int raster_fill_launch(const cpoly_raster_fill_args& a, cudaStream_t stream);
bool bn_eligible(int C, int dtype);
size_t bn_workspace_bytes(int C);
int bn_train_forward_launch(const cpoly_bn_fwd_args& a, cudaStream_t stream);
int bn_train_backward_launch(const cpoly_bn_bwd_args& a, cudaStream_t stream);   // device address of the file-scope scratch of cpoly_decode_static (this device)
size_t raster_workspace_bytes(int B, int K);
int raster_launch(const cpoly_raster_args& a, cudaStream_t stream);
int poly_iou_forward_launch(const cpoly_poly_iou_args& a, cudaStream_t stream);
int poly_iou_backward_launch(const cpoly_poly_iou_args& a, cudaStream_t stream);
int stem_pack_weights_launch(const void* w, int w_dtype, void* out, cudaStream_t stream);
int stem_conv_launch(const cpoly_stem_args& a, cudaStream_t stream);
int maxpool2x2_launch(const cpoly_maxpool_args& a, cudaStream_t stream);
size_t offset_conv_packed_bytes(int Cin);
int offset_conv_pack_launch(const float* weight, void* packed, int Cin, int n_out, cudaStream_t stream);
int offset_conv_launch(const cpoly_offset_conv_args& a, cudaStream_t stream);
}  // namespace cpoly
