// Deformable bilinear sampler pieces shared by the tcgen05 backward kernels (dcn_tc_bwd.cu): pixel-tile geometry,
// per-(pixel, tap) parameter staging and the packed blend.  Same arithmetic as the forward sampler of dcn_tc.cu
// (DCNv2 rule: a tap is dead unless -1 < p_y < H and -1 < p_x < W; out-of-image corners contribute 0).
#pragma once

#include "tc_common.cuh"

namespace cpoly {
namespace smp {

using namespace tc;

constexpr int TILE_H = 8, TILE_W = 16, TILE_PX = TILE_H * TILE_W;  // 128 pixels = one UMMA M (or K) extent
constexpr int NUM_TAPS = 9;

struct Geom {
  const void* x;
  cpoly_tensor4 offset, mask;
  const float* offset_bias;
  const float* mask_bias;
  int B, Cin, Cout, H, W;
  int om_dtype, mask_is_logit;
  int gom_dtype;   // backward only: element type of goffset / gmask
  int tiles_x, tiles_y, num_tiles;
};

__device__ __forceinline__ void tile_coords(const Geom& g, int tile, int& b, int& ty0, int& tx0) {
  const int tiles_per_img = g.tiles_x * g.tiles_y;
  b = tile / tiles_per_img;
  const int trem = tile - b * tiles_per_img;
  const int tyi = trem / g.tiles_x;
  ty0 = tyi * TILE_H;
  tx0 = (trem - tyi * g.tiles_x) * TILE_W;
}

// One deformed tap of one pixel, everything the backward needs.
struct TapGeom {
  uint32_t off[4];   // byte offsets of the corner channel vectors in x (dead corner: 0)
  float w[4];        // bilinear weights, NOT multiplied by the mask (dead corner: 0)
  float m;           // mask value (post-sigmoid)
  float ly, lx;
  int alive;         // the DCNv2 "whole sample" rule
  int valid[4];      // corner inside the image (and tap alive)
};

template <typename T>
__device__ __forceinline__ TapGeom make_tap_geom(const Geom& g, int b, int y, int x, int k) {
  TapGeom t;
#pragma unroll
  for (int q = 0; q < 4; ++q) { t.off[q] = 0u; t.w[q] = 0.f; t.valid[q] = 0; }
  t.m = 0.f; t.ly = t.lx = 0.f; t.alive = 0;
  if (y >= g.H || x >= g.W) return t;
  const int64_t opix = (int64_t)b * g.offset.sb + (int64_t)y * g.offset.sh + (int64_t)x * g.offset.sw;
  float dy = ld_dyn(g.offset.ptr, opix + (int64_t)(2 * k) * g.offset.sc, g.om_dtype);
  float dx = ld_dyn(g.offset.ptr, opix + (int64_t)(2 * k + 1) * g.offset.sc, g.om_dtype);
  float m = ld_dyn(g.mask.ptr, (int64_t)b * g.mask.sb + (int64_t)k * g.mask.sc + (int64_t)y * g.mask.sh + (int64_t)x * g.mask.sw, g.om_dtype);
  if (g.offset_bias) { dy += __ldg(g.offset_bias + 2 * k); dx += __ldg(g.offset_bias + 2 * k + 1); }
  if (g.mask_bias) m += __ldg(g.mask_bias + k);
  if (g.mask_is_logit) m = sigmoidf_accurate(m);
  t.m = m;
  const int ki = k / 3;
  const float py = (float)(y - 1 + ki) + dy;
  const float px = (float)(x - 1 + (k - 3 * ki)) + dx;
  if (!(py > -1.0f && px > -1.0f && py < (float)g.H && px < (float)g.W)) return t;
  t.alive = 1;
  const float fy = floorf(py), fx = floorf(px);
  const int y0 = (int)fy, x0 = (int)fx;
  t.ly = py - fy; t.lx = px - fx;
  const float hy = 1.0f - t.ly, hx = 1.0f - t.lx;
  const bool vy0 = y0 >= 0, vy1 = y0 + 1 <= g.H - 1, vx0 = x0 >= 0, vx1 = x0 + 1 <= g.W - 1;
  const uint32_t pixb = (uint32_t)g.Cin * (uint32_t)sizeof(T), rowb = (uint32_t)g.W * pixb;
  const uint32_t o00 = (uint32_t)(b * g.H * g.W) * pixb + (uint32_t)(y0 * g.W + x0) * pixb;  // may wrap for dead corners
  if (vy0 && vx0) { t.off[0] = o00;               t.w[0] = hy * hx;     t.valid[0] = 1; }
  if (vy0 && vx1) { t.off[1] = o00 + pixb;        t.w[1] = hy * t.lx;   t.valid[1] = 1; }
  if (vy1 && vx0) { t.off[2] = o00 + rowb;        t.w[2] = t.ly * hx;   t.valid[2] = 1; }
  if (vy1 && vx1) { t.off[3] = o00 + rowb + pixb; t.w[3] = t.ly * t.lx; t.valid[3] = 1; }
  return t;
}

// ---- blend of four 16-byte corner vectors into one 16-byte piece of a sampled row (mask folded into w)
template <typename T> struct Blend;
template <> struct Blend<__nv_bfloat16> {
  __device__ static __forceinline__ uint32_t weight(float w) { return pack_bf16x2(w, w); }
  __device__ static __forceinline__ uint4 run(const uint4 (&c)[4], const uint4 w) {
    uint4 out;
    out.x = bf16x2_fma(w.w, c[3].x, bf16x2_fma(w.z, c[2].x, bf16x2_fma(w.y, c[1].x, bf16x2_mul(w.x, c[0].x))));
    out.y = bf16x2_fma(w.w, c[3].y, bf16x2_fma(w.z, c[2].y, bf16x2_fma(w.y, c[1].y, bf16x2_mul(w.x, c[0].y))));
    out.z = bf16x2_fma(w.w, c[3].z, bf16x2_fma(w.z, c[2].z, bf16x2_fma(w.y, c[1].z, bf16x2_mul(w.x, c[0].z))));
    out.w = bf16x2_fma(w.w, c[3].w, bf16x2_fma(w.z, c[2].w, bf16x2_fma(w.y, c[1].w, bf16x2_mul(w.x, c[0].w))));
    return out;
  }
};
template <> struct Blend<float> {
  __device__ static __forceinline__ uint32_t weight(float w) { return __float_as_uint(w); }
  __device__ static __forceinline__ uint4 run(const uint4 (&c)[4], const uint4 w) {
    const float w0 = __uint_as_float(w.x), w1 = __uint_as_float(w.y), w2 = __uint_as_float(w.z), w3 = __uint_as_float(w.w);
    uint4 out;
#define CPOLY_BLEND1(f) \
  out.f = to_tf32(fmaf(w3, __uint_as_float(c[3].f), fmaf(w2, __uint_as_float(c[2].f), fmaf(w1, __uint_as_float(c[1].f), w0 * __uint_as_float(c[0].f)))))
    CPOLY_BLEND1(x); CPOLY_BLEND1(y); CPOLY_BLEND1(z); CPOLY_BLEND1(w);
#undef CPOLY_BLEND1
    return out;
  }
};

}  // namespace smp
}  // namespace cpoly
