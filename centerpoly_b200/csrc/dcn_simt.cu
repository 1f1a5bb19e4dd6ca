// Generic CUDA-core (SIMT) modulated deformable convolution: forward, data/coordinate backward and
// weight backward as implicit GEMMs whose A operand is produced on the fly by the bilinear sampler.
// Handles every geometry of the upstream DCNv2 op (any kernel size / stride / pad / dilation /
// deformable groups) and arbitrary tensor strides.  The 3x3/s1/p1 channels_last layers of the
// DLA-34 neck take the tcgen05 path (dcn_tc.cu) instead; this file is the general path and the
// first-correct implementation every other kernel was validated against.
//
// Semantics follow DCNv2 (missing submodule of the reference; see oracle/dcn_oracle.c header):
// reference call sites src/lib/models/networks/pose_dla_dcn.py:354-359, resnet_dcn.py:221-223.
#include "common.cuh"

namespace cpoly {

namespace {

constexpr int BM = 64;   // output pixels per CTA
constexpr int BN = 64;   // GEMM-N per CTA (Cout in forward, Cin in data-backward)
constexpr int BK = 16;   // reduction chunk
constexpr int NT = 256;  // threads per CTA: 16 x 16, each owns a 4 x 4 accumulator block

struct PixelTap {  // per (pixel-of-tile) parameters of the current (group, tap)
  int64_t off[4];
  float w[4];      // bilinear weights (NOT multiplied by the mask)
  float m;         // modulation mask (post-sigmoid)
  float ly, lx;
  int alive;
};

__device__ __forceinline__ void decode_pixel(int64_t p, const DcnGeom& g, int& b, int& ho, int& wo) {
  const int64_t hw = (int64_t)g.Ho * g.Wo;
  b = (int)(p / hw);
  const int r = (int)(p - (int64_t)b * hw);
  ho = r / g.Wo;
  wo = r - ho * g.Wo;
}

__device__ __forceinline__ void fill_pixel_tap(PixelTap& pt, int64_t p, int64_t npix, int grp, int tap,
                                               const DcnGeom& g, const cpoly_tensor4& x,
                                               const cpoly_tensor4& offset, const cpoly_tensor4& mask) {
  if (p >= npix) {
    pt.alive = 0; pt.m = 0.f; pt.ly = pt.lx = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) { pt.off[q] = -1; pt.w[q] = 0.f; }
    return;
  }
  int b, ho, wo;
  decode_pixel(p, g, b, ho, wo);
  const int T = g.kh * g.kw;
  const int i = tap / g.kw, j = tap - i * g.kw;
  const int64_t opix = (int64_t)b * offset.sb + (int64_t)ho * offset.sh + (int64_t)wo * offset.sw;
  float dy = ld_dyn(offset.ptr, opix + (int64_t)(grp * 2 * T + 2 * tap) * offset.sc, g.om_dtype);
  float dx = ld_dyn(offset.ptr, opix + (int64_t)(grp * 2 * T + 2 * tap + 1) * offset.sc, g.om_dtype);
  float m = ld_dyn(mask.ptr, (int64_t)b * mask.sb + (int64_t)(grp * T + tap) * mask.sc +
                                 (int64_t)ho * mask.sh + (int64_t)wo * mask.sw, g.om_dtype);
  if (g.offset_bias) { dy += __ldg(g.offset_bias + grp * 2 * T + 2 * tap); dx += __ldg(g.offset_bias + grp * 2 * T + 2 * tap + 1); }
  if (g.mask_bias) m += __ldg(g.mask_bias + grp * T + tap);
  if (g.mask_is_logit) m = sigmoidf_accurate(m);
  const float py = (float)(ho * g.sh - g.ph + i * g.dh) + dy;
  const float px = (float)(wo * g.sw - g.pw + j * g.dw) + dx;
  TapSample t = make_tap(py, px, g.H, g.W, (int64_t)b * x.sb, x.sh, x.sw);
#pragma unroll
  for (int q = 0; q < 4; ++q) { pt.off[q] = t.off[q]; pt.w[q] = t.w[q]; }
  pt.m = m; pt.ly = t.ly; pt.lx = t.lx; pt.alive = t.alive ? 1 : 0;
}

template <typename T>
__device__ __forceinline__ float sample_channel(const T* xp, const PixelTap& pt, int64_t coff) {
  float v = 0.f;
#pragma unroll
  for (int q = 0; q < 4; ++q)
    if (pt.off[q] >= 0) v = fmaf(pt.w[q], ld_as_float(xp + pt.off[q] + coff), v);
  return v;
}

// ------------------------------------------------------------------------------------- forward
template <typename T>
__global__ void __launch_bounds__(NT) dcn_fwd_simt_kernel(DcnGeom g, cpoly_tensor4 x, cpoly_tensor4 offset,
                                                          cpoly_tensor4 mask, const float* __restrict__ weight,
                                                          const float* __restrict__ bias, cpoly_tensor4 y,
                                                          const float* __restrict__ out_scale, int relu) {
  __shared__ PixelTap taps[BM];
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  const int64_t p0 = (int64_t)blockIdx.x * BM;
  const int o0 = blockIdx.y * BN;
  const int Tn = g.kh * g.kw, cpg = g.Cin / g.dg;
  const T* xp = reinterpret_cast<const T*>(x.ptr);
  float acc[4][4] = {};
  for (int grp = 0; grp < g.dg; ++grp)
    for (int tap = 0; tap < Tn; ++tap) {
      __syncthreads();
      if (tid < BM) fill_pixel_tap(taps[tid], p0 + tid, npix, grp, tap, g, x, offset, mask);
      __syncthreads();
      for (int c0 = 0; c0 < cpg; c0 += BK) {
        // A chunk: BK channels x BM pixels (value = mask * bilinear sample)
#pragma unroll
        for (int r = 0; r < (BK * BM) / NT; ++r) {
          const int e = tid + r * NT, pix = e % BM, kk = e / BM;
          const int c = c0 + kk;
          float v = 0.f;
          if (c < cpg && taps[pix].alive) v = taps[pix].m * sample_channel(xp, taps[pix], (int64_t)(grp * cpg + c) * x.sc);
          As[kk][pix] = v;
        }
        // B chunk: BK channels x BN output channels of this tap
#pragma unroll
        for (int r = 0; r < (BK * BN) / NT; ++r) {
          const int e = tid + r * NT, kk = e % BK, n = e / BK;
          const int c = c0 + kk, o = o0 + n;
          Bs[kk][n] = (c < cpg && o < g.Cout) ? __ldg(weight + ((int64_t)o * g.Cin + grp * cpg + c) * Tn + tap) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
          float a[4], bb[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; bb[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        __syncthreads();
      }
    }
  T* yp = reinterpret_cast<T*>(y.ptr);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t p = p0 + ty * 4 + i;
    if (p >= npix) continue;
    int b, ho, wo;
    decode_pixel(p, g, b, ho, wo);
    const int64_t ybase = (int64_t)b * y.sb + (int64_t)ho * y.sh + (int64_t)wo * y.sw;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = o0 + tx * 4 + j;
      if (o < g.Cout) {
        float v = fmaf(acc[i][j], out_scale ? __ldg(out_scale + o) : 1.f, bias ? __ldg(bias + o) : 0.f);
        if (relu) v = fmaxf(v, 0.f);
        st_from_float(yp + ybase + (int64_t)o * y.sc, v);
      }
    }
  }
}

// ---------------------------------------------------------------- backward: data + coordinates
// Per CTA: BM output pixels, every (group, tap), every input channel of the group in tiles of BN.
//   gcol[px, c] = sum_o gout[px, o] W[o, c, tap]          (implicit GEMM, K = Cout)
//   gmask      += gcol * v_c;   goffset += gcol * dv_c/dp;   gx[corner] += gcol * m * w_corner
// goffset / gmask are written once per (pixel, tap) without atomics; gx is scattered with fp32
// atomics in this generic path (the tcgen05 path has the atomics-free col2im).
template <typename T>
__global__ void __launch_bounds__(NT) dcn_bwd_data_simt_kernel(DcnGeom g, cpoly_tensor4 x, cpoly_tensor4 offset,
                                                               cpoly_tensor4 mask, const float* __restrict__ weight,
                                                               cpoly_tensor4 gout, cpoly_tensor4 gx,
                                                               cpoly_tensor4 goffset, cpoly_tensor4 gmask) {
  __shared__ PixelTap taps[BM];
  __shared__ float As[BK][BM + 4];  // gout chunk: [o][pixel]
  __shared__ float Bs[BK][BN + 4];  // weight chunk: [o][c]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  const int64_t p0 = (int64_t)blockIdx.x * BM;
  const int Tn = g.kh * g.kw, cpg = g.Cin / g.dg;
  const T* xp = reinterpret_cast<const T*>(x.ptr);
  const T* gop = reinterpret_cast<const T*>(gout.ptr);
  float* gxp = reinterpret_cast<float*>(gx.ptr);
  void* goffp = goffset.ptr;
  void* gmp = gmask.ptr;

  // gout addressing of this thread's A-load slots is tap independent
  for (int grp = 0; grp < g.dg; ++grp)
    for (int tap = 0; tap < Tn; ++tap) {
      __syncthreads();
      if (tid < BM) fill_pixel_tap(taps[tid], p0 + tid, npix, grp, tap, g, x, offset, mask);
      __syncthreads();
      float am[4] = {}, ay[4] = {}, ax[4] = {};
      for (int c0 = 0; c0 < cpg; c0 += BN) {
        float acc[4][4] = {};
        for (int o0 = 0; o0 < g.Cout; o0 += BK) {
#pragma unroll
          for (int r = 0; r < (BK * BM) / NT; ++r) {
            const int e = tid + r * NT, pix = e % BM, kk = e / BM;
            const int64_t p = p0 + pix;
            const int o = o0 + kk;
            float v = 0.f;
            if (p < npix && o < g.Cout) {
              int b, ho, wo;
              decode_pixel(p, g, b, ho, wo);
              v = ld_as_float(gop + (int64_t)b * gout.sb + (int64_t)o * gout.sc + (int64_t)ho * gout.sh + (int64_t)wo * gout.sw);
            }
            As[kk][pix] = v;
          }
#pragma unroll
          for (int r = 0; r < (BK * BN) / NT; ++r) {
            const int e = tid + r * NT, n = e % BN, kk = e / BN;
            const int c = c0 + n, o = o0 + kk;
            Bs[kk][n] = (c < cpg && o < g.Cout) ? __ldg(weight + ((int64_t)o * g.Cin + grp * cpg + c) * Tn + tap) : 0.f;
          }
          __syncthreads();
#pragma unroll
          for (int kk = 0; kk < BK; ++kk) {
            float a[4], bb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; bb[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
          }
          __syncthreads();
        }
        // epilogue of this channel tile: combine gcol with the four corners of x
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const PixelTap& pt = taps[ty * 4 + i];
          if (!pt.alive) continue;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = c0 + tx * 4 + j;
            if (c >= cpg) continue;
            const int64_t coff = (int64_t)(grp * cpg + c) * x.sc;
            float v[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) v[q] = pt.off[q] >= 0 ? ld_as_float(xp + pt.off[q] + coff) : 0.f;
            const float gc = acc[i][j];
            const float val = pt.w[0] * v[0] + pt.w[1] * v[1] + pt.w[2] * v[2] + pt.w[3] * v[3];
            am[i] = fmaf(gc, val, am[i]);
            ay[i] = fmaf(gc, (1.f - pt.lx) * (v[2] - v[0]) + pt.lx * (v[3] - v[1]), ay[i]);
            ax[i] = fmaf(gc, (1.f - pt.ly) * (v[1] - v[0]) + pt.ly * (v[3] - v[2]), ax[i]);
            if (gxp) {
              const float s = gc * pt.m;
              const int64_t gcoff = (int64_t)(grp * cpg + c) * gx.sc;
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (pt.off[q] >= 0) {
                  // gx shares x's strides (checked on the host), so the corner offsets carry over
                  atomicAdd(gxp + pt.off[q] + gcoff, s * pt.w[q]);
                }
            }
          }
        }
      }
      // reduce the per-pixel sums over the 16 threads (tx) that share a pixel row: they are the
      // 16 consecutive lanes of a half warp.
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int s = 8; s >= 1; s >>= 1) {
          am[i] += __shfl_xor_sync(0xffffffffu, am[i], s);
          ay[i] += __shfl_xor_sync(0xffffffffu, ay[i], s);
          ax[i] += __shfl_xor_sync(0xffffffffu, ax[i], s);
        }
      if (tx == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int64_t p = p0 + ty * 4 + i;
          if (p >= npix) continue;
          int b, ho, wo;
          decode_pixel(p, g, b, ho, wo);
          const float m = taps[ty * 4 + i].m;
          if (gmp) {
            float gm = am[i];
            if (g.mask_is_logit) gm *= m * (1.f - m);
            st_dyn(gmp, (int64_t)b * gmask.sb + (int64_t)(grp * Tn + tap) * gmask.sc + (int64_t)ho * gmask.sh + (int64_t)wo * gmask.sw, gm, g.gom_dtype);
          }
          if (goffp) {
            const int64_t ob = (int64_t)b * goffset.sb + (int64_t)ho * goffset.sh + (int64_t)wo * goffset.sw;
            st_dyn(goffp, ob + (int64_t)(grp * 2 * Tn + 2 * tap) * goffset.sc, m * ay[i], g.gom_dtype);
            st_dyn(goffp, ob + (int64_t)(grp * 2 * Tn + 2 * tap + 1) * goffset.sc, m * ax[i], g.gom_dtype);
          }
        }
      }
    }
}

// ---------------------------------------------------------------------- backward: weights
// gW[o, c, tap] = sum_px gout[px, o] * m * v_c(p_tap).  grid = (Cout tiles, (group, tap, Cin tiles), pixel
// splits); every CTA writes its partial tile to the workspace; a fixed-order reduction follows, so
// the result is deterministic (no atomics).
template <typename T>
__global__ void __launch_bounds__(NT) dcn_bwd_weight_simt_kernel(DcnGeom g, cpoly_tensor4 x, cpoly_tensor4 offset,
                                                                 cpoly_tensor4 mask, cpoly_tensor4 gout,
                                                                 float* __restrict__ partial, int nsplit,
                                                                 int64_t pix_per_split) {
  __shared__ PixelTap taps[BK];
  __shared__ float As[BK][BM + 4];  // gout chunk: [pixel][o]
  __shared__ float Bs[BK][BN + 4];  // sampled chunk: [pixel][c]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  const int Tn = g.kh * g.kw, cpg = g.Cin / g.dg;
  const int ctiles = (cpg + BN - 1) / BN;
  const int o0 = blockIdx.x * BM;
  int yy = blockIdx.y;
  const int ct = yy % ctiles; yy /= ctiles;
  const int tap = yy % Tn;
  const int grp = yy / Tn;
  const int c0 = ct * BN;
  const int split = blockIdx.z;
  const int64_t pbeg = (int64_t)split * pix_per_split;
  const int64_t pend = min(npix, pbeg + pix_per_split);
  const T* xp = reinterpret_cast<const T*>(x.ptr);
  const T* gop = reinterpret_cast<const T*>(gout.ptr);
  float acc[4][4] = {};
  for (int64_t pc = pbeg; pc < pend; pc += BK) {
    __syncthreads();
    if (tid < BK) fill_pixel_tap(taps[tid], (pc + tid < pend) ? pc + tid : npix, npix, grp, tap, g, x, offset, mask);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < (BK * BM) / NT; ++r) {
      const int e = tid + r * NT, n = e % BM, kk = e / BM;
      const int64_t p = pc + kk;
      const int o = o0 + n;
      float v = 0.f;
      if (p < pend && o < g.Cout) {
        int b, ho, wo;
        decode_pixel(p, g, b, ho, wo);
        v = ld_as_float(gop + (int64_t)b * gout.sb + (int64_t)o * gout.sc + (int64_t)ho * gout.sh + (int64_t)wo * gout.sw);
      }
      As[kk][n] = v;
    }
#pragma unroll
    for (int r = 0; r < (BK * BN) / NT; ++r) {
      const int e = tid + r * NT, n = e % BN, kk = e / BN;
      const int c = c0 + n;
      float v = 0.f;
      if (c < cpg && taps[kk].alive) v = taps[kk].m * sample_channel(xp, taps[kk], (int64_t)(grp * cpg + c) * x.sc);
      Bs[kk][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; bb[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
  }
  // partial layout: [split][Cout][Cin][T]
  float* dst = partial + (int64_t)split * g.Cout * g.Cin * Tn;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = o0 + ty * 4 + i;
    if (o >= g.Cout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c0 + tx * 4 + j;
      if (c < cpg) dst[((int64_t)o * g.Cin + grp * cpg + c) * Tn + tap] = acc[i][j];
    }
  }
}

__global__ void reduce_splits_kernel(const float* __restrict__ partial, float* __restrict__ out, int64_t n, int nsplit) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = 0.f;
  for (int k = 0; k < nsplit; ++k) s += partial[(int64_t)k * n + i];
  out[i] = s;
}

// gb[o] = sum over (b, ho, wo) of gout — one CTA per output channel, fixed-order tree reduction.
template <typename T>
__global__ void __launch_bounds__(256) dcn_bwd_bias_kernel(DcnGeom g, cpoly_tensor4 gout, float* __restrict__ gb) {
  __shared__ float red[256];
  const int o = blockIdx.x;
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  const T* gop = reinterpret_cast<const T*>(gout.ptr);
  float s = 0.f;
  for (int64_t p = threadIdx.x; p < npix; p += blockDim.x) {
    int b, ho, wo;
    decode_pixel(p, g, b, ho, wo);
    s += ld_as_float(gop + (int64_t)b * gout.sb + (int64_t)o * gout.sc + (int64_t)ho * gout.sh + (int64_t)wo * gout.sw);
  }
  red[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w >= 1; w >>= 1) {
    if ((int)threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) gb[o] = red[0];
}

// Channels-last fast path of the bias gradient: a thread owns one 16-byte piece of the channel vector (8 bf16 / 4 fp32
// channels) and walks the pixels of its CTA's fixed range with four independent 16-byte loads in flight; the pieces
// of a pixel are consecutive threads (coalesced).  partial[blk][o] is then summed over the blocks by one warp per
// channel in a fixed order (dcn_bwd_bias_final_kernel).  Fixed partition + fixed order = deterministic.
// (The one-CTA-per-channel kernel above reads NHWC with a stride of Cout elements: 341 us for 64 ch @256x512.)
constexpr int BIAS_MAX_BLOCKS = 592;
template <typename T>
__global__ void __launch_bounds__(256) dcn_bwd_bias_partial_kernel(const T* __restrict__ gout, float* __restrict__ partial, int64_t npix,
                                                                  int Cout, int64_t pix_per_block) {
  constexpr int kVec = 16 / sizeof(T);
  __shared__ float red[256][kVec];
  const int lpp = Cout / kVec;                       // 16-byte pieces (threads) per pixel; divides 256
  const int rows = 256 / lpp, piece = threadIdx.x % lpp, row = threadIdx.x / lpp;
  const int64_t p0 = (int64_t)blockIdx.x * pix_per_block;
  const int64_t p1 = p0 + pix_per_block < npix ? p0 + pix_per_block : npix;
  float acc[4][kVec];
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int e = 0; e < kVec; ++e) acc[u][e] = 0.f;
  auto add = [&](int u, int64_t p) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(gout + p * Cout) + piece);
    if (sizeof(T) == 4) {
      acc[u][0] += __uint_as_float(q.x); acc[u][1] += __uint_as_float(q.y);
      acc[u][2 % kVec] += __uint_as_float(q.z); acc[u][3 % kVec] += __uint_as_float(q.w);
    } else {
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        acc[u][(2 * e) % kVec] += __uint_as_float(w[e] << 16);
        acc[u][(2 * e + 1) % kVec] += __uint_as_float(w[e] & 0xffff0000u);
      }
    }
  };
  int64_t p = p0 + row;
  for (; p + 3 * rows < p1; p += 4 * rows) {
    add(0, p); add(1, p + rows); add(2, p + 2 * rows); add(3, p + 3 * rows);
  }
  for (; p < p1; p += rows) add(0, p);
#pragma unroll
  for (int e = 0; e < kVec; ++e) red[threadIdx.x][e] = (acc[0][e] + acc[1][e]) + (acc[2][e] + acc[3][e]);
  __syncthreads();
  if (row == 0) {
#pragma unroll
    for (int e = 0; e < kVec; ++e) {
      float s = red[piece][e];
      for (int r = 1; r < rows; ++r) s += red[r * lpp + piece][e];
      partial[(int64_t)blockIdx.x * Cout + piece * kVec + e] = s;
    }
  }
}
// one warp per channel: lane l sums blocks l, l+32, ... in order, then a fixed shuffle tree
__global__ void __launch_bounds__(32) dcn_bwd_bias_final_kernel(const float* __restrict__ partial, float* __restrict__ gb, int Cout, int nblk) {
  const int o = blockIdx.x, lane = threadIdx.x;
  float s = 0.f;
  for (int b = lane; b < nblk; b += 32) s += partial[(int64_t)b * Cout + o];
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if (lane == 0) gb[o] = s;
}

__global__ void zero_strided_kernel(cpoly_tensor4 t, int B, int C, int H, int W) {
  const int64_t n = (int64_t)B * C * H * W;
  float* p = reinterpret_cast<float*>(t.ptr);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H); r /= H;
    const int c = (int)(r % C);
    const int b = (int)(r / C);
    p[b * t.sb + c * t.sc + h * t.sh + w * t.sw] = 0.f;
  }
}

int weight_splits(const DcnGeom& g, int64_t* pix_per_split) {
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  const int Tn = g.kh * g.kw, cpg = g.Cin / g.dg;
  const int64_t base_ctas = (int64_t)((g.Cout + BM - 1) / BM) * g.dg * Tn * ((cpg + BN - 1) / BN);
  int64_t want = (148 * 4 + base_ctas - 1) / base_ctas;  // aim for ~4 CTAs per SM
  int64_t maxs = (npix + 255) / 256;                     // at least 256 pixels per split
  int64_t ns = want < 1 ? 1 : want;
  if (ns > maxs) ns = maxs;
  if (ns < 1) ns = 1;
  if (ns > 64) ns = 64;
  int64_t pps = (npix + ns - 1) / ns;
  pps = (pps + BK - 1) / BK * BK;
  ns = (npix + pps - 1) / pps;
  *pix_per_split = pps;
  return (int)ns;
}

}  // namespace

int dcn_forward_simt(const cpoly_dcn_fwd_args& a, cudaStream_t stream) {
  const DcnGeom g = make_geom(a);
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  if (npix == 0) return CPOLY_OK;
  dim3 grid((unsigned)((npix + BM - 1) / BM), (unsigned)((g.Cout + BN - 1) / BN));
  if (a.dtype == CPOLY_BF16)
    dcn_fwd_simt_kernel<__nv_bfloat16><<<grid, NT, 0, stream>>>(g, a.x, a.offset, a.mask, a.weight, a.bias, a.y, a.out_scale, a.relu);
  else
    dcn_fwd_simt_kernel<float><<<grid, NT, 0, stream>>>(g, a.x, a.offset, a.mask, a.weight, a.bias, a.y, a.out_scale, a.relu);
  return check_launch("dcn_fwd_simt_kernel");
}

size_t dcn_backward_simt_workspace(const cpoly_dcn_bwd_args& a) {
  const DcnGeom g = make_geom(a.f);
  int64_t pps;
  const int ns = weight_splits(g, &pps);
  const size_t wneed = (size_t)ns * g.Cout * g.Cin * g.kh * g.kw * sizeof(float);
  const size_t bneed = (size_t)BIAS_MAX_BLOCKS * g.Cout * sizeof(float);   // bias-gradient partials (reuse the same scratch)
  return (wneed > bneed ? wneed : bneed) + 256;
}

int dcn_backward_simt(const cpoly_dcn_bwd_args& a, cudaStream_t stream) {
  DcnGeom g = make_geom(a.f);
  g.gom_dtype = a.gom_dtype;
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  const bool bf = a.f.dtype == CPOLY_BF16;
  if (a.gx.ptr) {
    CPOLY_REQUIRE(a.gx.sb == a.f.x.sb && a.gx.sc == a.f.x.sc && a.gx.sh == a.f.x.sh && a.gx.sw == a.f.x.sw,
                  CPOLY_EINVAL, "grad_input must have the strides of input");
    zero_strided_kernel<<<148 * 4, 256, 0, stream>>>(a.gx, g.B, g.Cin, g.H, g.W);
    if (int rc = check_launch("zero_strided_kernel")) return rc;
  }
  if (npix == 0) return CPOLY_OK;
  if (a.gx.ptr || a.goffset.ptr || a.gmask.ptr) {
    dim3 grid((unsigned)((npix + BM - 1) / BM));
    if (bf)
      dcn_bwd_data_simt_kernel<__nv_bfloat16><<<grid, NT, 0, stream>>>(g, a.f.x, a.f.offset, a.f.mask, a.f.weight, a.gout, a.gx, a.goffset, a.gmask);
    else
      dcn_bwd_data_simt_kernel<float><<<grid, NT, 0, stream>>>(g, a.f.x, a.f.offset, a.f.mask, a.f.weight, a.gout, a.gx, a.goffset, a.gmask);
    if (int rc = check_launch("dcn_bwd_data_simt_kernel")) return rc;
  }
  if (a.gweight) {
    int64_t pps;
    const int ns = weight_splits(g, &pps);
    const size_t need = (size_t)ns * g.Cout * g.Cin * g.kh * g.kw * sizeof(float);
    CPOLY_REQUIRE(a.workspace && a.workspace_bytes >= need, CPOLY_EWORKSPACE,
                  "dcn backward workspace too small: need %zu bytes, have %zu", need, a.workspace_bytes);
    float* partial = reinterpret_cast<float*>(a.workspace);
    const int Tn = g.kh * g.kw, cpg = g.Cin / g.dg;
    dim3 grid((unsigned)((g.Cout + BM - 1) / BM), (unsigned)(g.dg * Tn * ((cpg + BN - 1) / BN)), (unsigned)ns);
    if (bf)
      dcn_bwd_weight_simt_kernel<__nv_bfloat16><<<grid, NT, 0, stream>>>(g, a.f.x, a.f.offset, a.f.mask, a.gout, partial, ns, pps);
    else
      dcn_bwd_weight_simt_kernel<float><<<grid, NT, 0, stream>>>(g, a.f.x, a.f.offset, a.f.mask, a.gout, partial, ns, pps);
    if (int rc = check_launch("dcn_bwd_weight_simt_kernel")) return rc;
    const int64_t n = (int64_t)g.Cout * g.Cin * Tn;
    reduce_splits_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(partial, a.gweight, n, ns);
    if (int rc = check_launch("reduce_splits_kernel")) return rc;
  }
  const bool gout_nhwc = a.gout.sc == 1 && a.gout.sw == g.Cout && a.gout.sh == (int64_t)g.Wo * g.Cout && a.gout.sb == (int64_t)g.Ho * g.Wo * g.Cout;
  const int bias_vec = bf ? 8 : 4;
  const bool bias_fast = gout_nhwc && g.Cout % bias_vec == 0 && 256 % (g.Cout / bias_vec) == 0 && g.Cout / bias_vec <= 256 &&
                         (reinterpret_cast<uintptr_t>(a.gout.ptr) & 15) == 0;
  if (a.gbias && bias_fast && a.workspace && a.workspace_bytes >= (size_t)BIAS_MAX_BLOCKS * g.Cout * sizeof(float)) {
    const int rows = 256 / (g.Cout / bias_vec);
    const int64_t min_ppb = (int64_t)rows * 8;   // >= 8 loads per thread, else as many CTAs as the cap allows
    int64_t nblk = (npix + min_ppb - 1) / min_ppb;
    if (nblk > BIAS_MAX_BLOCKS) nblk = BIAS_MAX_BLOCKS;
    const int64_t ppb = (npix + nblk - 1) / nblk;
    nblk = (npix + ppb - 1) / ppb;
    float* partial = reinterpret_cast<float*>(a.workspace);   // stream-ordered after the weight-gradient reduction that used it
    if (bf) dcn_bwd_bias_partial_kernel<__nv_bfloat16><<<(unsigned)nblk, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a.gout.ptr), partial, npix, g.Cout, ppb);
    else dcn_bwd_bias_partial_kernel<float><<<(unsigned)nblk, 256, 0, stream>>>(reinterpret_cast<const float*>(a.gout.ptr), partial, npix, g.Cout, ppb);
    if (int rc = check_launch("dcn_bwd_bias_partial_kernel")) return rc;
    dcn_bwd_bias_final_kernel<<<g.Cout, 32, 0, stream>>>(partial, a.gbias, g.Cout, (int)nblk);
    if (int rc = check_launch("dcn_bwd_bias_final_kernel")) return rc;
  } else if (a.gbias) {
    if (bf) dcn_bwd_bias_kernel<__nv_bfloat16><<<g.Cout, 256, 0, stream>>>(g, a.gout, a.gbias);
    else dcn_bwd_bias_kernel<float><<<g.Cout, 256, 0, stream>>>(g, a.gout, a.gbias);
    if (int rc = check_launch("dcn_bwd_bias_kernel")) return rc;
  }
  return CPOLY_OK;
}

}  // namespace cpoly
