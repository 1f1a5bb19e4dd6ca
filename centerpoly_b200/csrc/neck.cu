// IDAUp up-sampling: depthwise ConvTranspose2d(C, C, 2f, stride f, padding f/2, groups=C, bias=False)
// (reference src/lib/models/networks/pose_dla_dcn.py:372-375, applied at :385) fused with the skip add
// `layers[i] + layers[i-1]` of :387.  cuDNN runs this as a generic dgrad (170 us per call at 256x512x64);
// it is a pure streaming op: every output pixel reads a 2x2 input neighbourhood and 4 weights per channel.
//
//   y[b,c,oy,ox] = sum_{iy,ix} x[b,c,iy,ix] * w[c, oy + pad - iy*f, ox + pad - ix*f]  (+ skip[b,c,oy,ox])
//
// With kernel 2f and stride f exactly two input rows / columns contribute: iy1 = (oy + pad) / f with
// ky = (oy + pad) % f, and iy0 = iy1 - 1 with ky + f.  HBM bound: algorithmic bytes = s * B*C*(H*W + [2] * Ho*Wo).
#include <stdlib.h>

#include "common.cuh"

namespace cpoly {
namespace {

struct UpParams {
  int B, C, H, W, f, Ho, Wo;
  cpoly_tensor4 x, skip, y;
  const float* w;
  int w_tapmajor;
};

template <typename T, int VEC>
struct Vec;
template <> struct Vec<float, 4> {
  __device__ static __forceinline__ void load(const float* p, float (&v)[4]) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  }
  __device__ static __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Vec<__nv_bfloat16, 8> {
  __device__ static __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[2 * j] = __uint_as_float(w[j] << 16);
      v[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
    }
  }
  __device__ static __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 q;
    uint32_t* w = &q.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __nv_bfloat162 pk = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
      w[j] = *reinterpret_cast<const uint32_t*>(&pk);
    }
    *reinterpret_cast<uint4*>(p) = q;
  }
};

template <typename T> __device__ __forceinline__ void unpack16(const uint4 q, float* v);
template <> __device__ __forceinline__ void unpack16<float>(const uint4 q, float* v) {
  v[0] = __uint_as_float(q.x); v[1] = __uint_as_float(q.y); v[2] = __uint_as_float(q.z); v[3] = __uint_as_float(q.w);
}
template <> __device__ __forceinline__ void unpack16<__nv_bfloat16>(const uint4 q, float* v) {
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[2 * j] = __uint_as_float(w[j] << 16);
    v[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
  }
}

// channels_last fast path: one thread per (output pixel, VEC-channel piece); pieces of a pixel are adjacent
// threads, so loads and stores are 16-byte vectors over contiguous channels.
template <typename T, int VEC>
__global__ void __launch_bounds__(256) upsample_nhwc_kernel(UpParams p) {
  const int pieces = p.C / VEC;
  const int64_t total = (int64_t)p.B * p.Ho * p.Wo * pieces;
  const int K = 2 * p.f, pad = p.f / 2;
  const T* xg = reinterpret_cast<const T*>(p.x.ptr);
  const T* sg = reinterpret_cast<const T*>(p.skip.ptr);
  T* yg = reinterpret_cast<T*>(p.y.ptr);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int piece = (int)(i % pieces);
    int64_t t = i / pieces;
    const int ox = (int)(t % p.Wo); t /= p.Wo;
    const int oy = (int)(t % p.Ho);
    const int b = (int)(t / p.Ho);
    const int c0 = piece * VEC;
    const int iy1 = (oy + pad) / p.f, ky1 = (oy + pad) - iy1 * p.f;
    const int ix1 = (ox + pad) / p.f, kx1 = (ox + pad) - ix1 * p.f;
    float acc[VEC];
    if (sg) Vec<T, VEC>::load(sg + (int64_t)b * p.skip.sb + (int64_t)oy * p.skip.sh + (int64_t)ox * p.skip.sw + c0, acc);
    else {
#pragma unroll
      for (int j = 0; j < VEC; ++j) acc[j] = 0.f;
    }
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int iy = iy1 - dy, ky = ky1 + dy * p.f;
      if (iy < 0 || iy >= p.H) continue;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int ix = ix1 - dx, kx = kx1 + dx * p.f;
        if (ix < 0 || ix >= p.W) continue;
        float v[VEC];
        Vec<T, VEC>::load(xg + (int64_t)b * p.x.sb + (int64_t)iy * p.x.sh + (int64_t)ix * p.x.sw + c0, v);
        if (p.w_tapmajor) {   // (2f*2f, C): the piece's channels are contiguous
          const float4* wp = reinterpret_cast<const float4*>(p.w + (int64_t)(ky * K + kx) * p.C + c0);
#pragma unroll
          for (int j = 0; j < VEC / 4; ++j) {
            const float4 w4 = __ldg(wp + j);
            acc[4 * j] = fmaf(v[4 * j], w4.x, acc[4 * j]);
            acc[4 * j + 1] = fmaf(v[4 * j + 1], w4.y, acc[4 * j + 1]);
            acc[4 * j + 2] = fmaf(v[4 * j + 2], w4.z, acc[4 * j + 2]);
            acc[4 * j + 3] = fmaf(v[4 * j + 3], w4.w, acc[4 * j + 3]);
          }
        } else {
          const float* wp = p.w + ((int64_t)c0 * K + ky) * K + kx;
#pragma unroll
          for (int j = 0; j < VEC; ++j) acc[j] = fmaf(v[j], __ldg(wp + (int64_t)j * K * K), acc[j]);
        }
      }
    }
    Vec<T, VEC>::store(yg + (int64_t)b * p.y.sb + (int64_t)oy * p.y.sh + (int64_t)ox * p.y.sw + c0, acc);
  }
}

// channels_last, tap-major weights: one thread per (2x2 output quad, VEC-channel piece).  A quad whose origin has
// the parity of pad shares its 2x2 input neighbourhood (f is even, so (o + pad) and (o + pad + 1) never straddle a
// multiple of f): 4 input loads, one unpack and one set of index arithmetic serve 4 outputs - the one-pixel kernel
// above is bound by instruction issue (57 % issue slots busy, 10 % of DRAM), not by bytes.
template <typename T, int VEC>
__global__ void __launch_bounds__(256) upsample_nhwc_quad_kernel(UpParams p) {
  const int pieces = p.C / VEC;
  const int K = 2 * p.f, pad = p.f / 2, par = pad & 1;
  const int qh = p.Ho / 2 + par, qw = p.Wo / 2 + par;          // Ho, Wo are even (f is)
  const int64_t total = (int64_t)p.B * qh * qw * pieces;
  const T* xg = reinterpret_cast<const T*>(p.x.ptr);
  const T* sg = reinterpret_cast<const T*>(p.skip.ptr);
  T* yg = reinterpret_cast<T*>(p.y.ptr);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int piece = (int)(i % pieces);
    int t = (int)(i / pieces);
    const int qx = t % qw; t /= qw;
    const int qy = t % qh;
    const int b = t / qh;
    const int c0 = piece * VEC;
    const int oy0 = 2 * qy - par, ox0 = 2 * qx - par;
    const int iy1 = (oy0 + pad) / p.f, ky0 = (oy0 + pad) - iy1 * p.f;      // oy0 + pad >= 0
    const int ix1 = (ox0 + pad) / p.f, kx0 = (ox0 + pad) - ix1 * p.f;
    // every load of the quad is issued before anything is used: 8 x 16 bytes in flight per thread (with one output
    // pixel and a grid-stride loop the kernel sat at 1.4 TB/s, bound by memory latency, not by bytes or issue)
    uint4 xr[2][2], sr[2][2];
    bool ok[2][2], oko[2][2];
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int iy = iy1 - dy, ix = ix1 - dx;
        ok[dy][dx] = iy >= 0 && iy < p.H && ix >= 0 && ix < p.W;
        xr[dy][dx] = make_uint4(0u, 0u, 0u, 0u);
        if (ok[dy][dx]) xr[dy][dx] = __ldg(reinterpret_cast<const uint4*>(xg + (int64_t)b * p.x.sb + (int64_t)iy * p.x.sh + (int64_t)ix * p.x.sw + c0));
        const int oy = oy0 + dy, ox = ox0 + dx;
        oko[dy][dx] = oy >= 0 && oy < p.Ho && ox >= 0 && ox < p.Wo;
        sr[dy][dx] = make_uint4(0u, 0u, 0u, 0u);
        if (sg && oko[dy][dx]) sr[dy][dx] = __ldg(reinterpret_cast<const uint4*>(sg + (int64_t)b * p.skip.sb + (int64_t)oy * p.skip.sh + (int64_t)ox * p.skip.sw + c0));
      }
    float v[2][2][VEC];
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) unpack16<T>(xr[dy][dx], v[dy][dx]);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        if (!oko[r][q]) continue;
        const int oy = oy0 + r, ox = ox0 + q;
        float acc[VEC];
        unpack16<T>(sr[r][q], acc);            // zeros when there is no skip tensor
        // same tap order as the one-pixel kernel (dy, dx ascending): identical fp32 results
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
            if (!ok[dy][dx]) continue;
            const float4* wp = reinterpret_cast<const float4*>(p.w + (int64_t)((ky0 + r + dy * p.f) * K + kx0 + q + dx * p.f) * p.C + c0);
#pragma unroll
            for (int j = 0; j < VEC / 4; ++j) {
              const float4 w4 = __ldg(wp + j);
              acc[4 * j] = fmaf(v[dy][dx][4 * j], w4.x, acc[4 * j]);
              acc[4 * j + 1] = fmaf(v[dy][dx][4 * j + 1], w4.y, acc[4 * j + 1]);
              acc[4 * j + 2] = fmaf(v[dy][dx][4 * j + 2], w4.z, acc[4 * j + 2]);
              acc[4 * j + 3] = fmaf(v[dy][dx][4 * j + 3], w4.w, acc[4 * j + 3]);
            }
          }
        Vec<T, VEC>::store(yg + (int64_t)b * p.y.sb + (int64_t)oy * p.y.sh + (int64_t)ox * p.y.sw + c0, acc);
      }
    }
  }
}

// ---- backward of the up-sampler (training: the reference's nn.ConvTranspose2d autograd; cuDNN runs the depthwise
// weight gradient as `wgrad2d_grouped_direct_kernel` at ~1 ms per call and the data gradient as `dgrad2d_c1_k1_nhwc` at
// ~0.5 ms: 17 % of the configs[3] training step).  channels_last, tap-major fp32 weights.
// grad_x[b,c,iy,ix] = sum_{ky,kx} w[c,ky,kx] * gy[b,c, iy*f + ky - pad, ix*f + kx - pad]: one thread per (input pixel, piece)
template <typename T, int VEC>
__global__ void __launch_bounds__(256) upsample_bwd_input_kernel(UpParams p, cpoly_tensor4 gy, cpoly_tensor4 gx) {
  const int pieces = p.C / VEC;
  const int64_t total = (int64_t)p.B * p.H * p.W * pieces;
  const int K = 2 * p.f, pad = p.f / 2;
  const T* gyp = reinterpret_cast<const T*>(gy.ptr);
  T* gxp = reinterpret_cast<T*>(gx.ptr);
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int piece = (int)(i % pieces);
  int t = (int)(i / pieces);
  const int ix = t % p.W; t /= p.W;
  const int iy = t % p.H;
  const int b = t / p.H;
  const int c0 = piece * VEC;
  float acc[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) acc[j] = 0.f;
  for (int ky = 0; ky < K; ++ky) {
    const int oy = iy * p.f + ky - pad;
    if (oy < 0 || oy >= p.Ho) continue;
    for (int kx = 0; kx < K; ++kx) {
      const int ox = ix * p.f + kx - pad;
      if (ox < 0 || ox >= p.Wo) continue;
      float g[VEC];
      Vec<T, VEC>::load(gyp + (int64_t)b * gy.sb + (int64_t)oy * gy.sh + (int64_t)ox * gy.sw + c0, g);
      const float4* wp = reinterpret_cast<const float4*>(p.w + (int64_t)(ky * K + kx) * p.C + c0);
#pragma unroll
      for (int j = 0; j < VEC / 4; ++j) {
        const float4 w4 = __ldg(wp + j);
        acc[4 * j] = fmaf(g[4 * j], w4.x, acc[4 * j]);
        acc[4 * j + 1] = fmaf(g[4 * j + 1], w4.y, acc[4 * j + 1]);
        acc[4 * j + 2] = fmaf(g[4 * j + 2], w4.z, acc[4 * j + 2]);
        acc[4 * j + 3] = fmaf(g[4 * j + 3], w4.w, acc[4 * j + 3]);
      }
    }
  }
  Vec<T, VEC>::store(gxp + (int64_t)b * gx.sb + (int64_t)iy * gx.sh + (int64_t)ix * gx.sw + c0, acc);
}

// grad_w[(ky,kx), c] += sum over outputs of one phase (oy % f, ox % f): the 4 contributing taps are fixed per phase, so a
// thread keeps 4 x VEC partial sums in registers over its share of the phase's pixels, then warp-shuffle + atomicAdd.
// blockIdx.y = phase, blockIdx.x = pixel chunk; thread = (pixel lane, piece).  fp32 atomics: summation order varies
// between runs at the 1e-7 level (as in cuDNN's own grouped weight gradient).
template <typename T, int VEC>
__global__ void __launch_bounds__(256) upsample_bwd_weight_kernel(UpParams p, cpoly_tensor4 gy, float* __restrict__ gw, int px_per_block, float* __restrict__ partial) {
  const int pieces = p.C / VEC;                   // 8, 16 or 32 (power of two <= 32)
  const int K = 2 * p.f, pad = p.f / 2;
  const int py = blockIdx.y / p.f, pxp = blockIdx.y - py * p.f;
  const int piece = threadIdx.x % pieces, plane = threadIdx.x / pieces, nlanes = 256 / pieces;
  const int c0 = piece * VEC;
  const int uy = (py + pad) / p.f, ky1 = (py + pad) - uy * p.f;     // iy1 = m + uy
  const int ux = (pxp + pad) / p.f, kx1 = (pxp + pad) - ux * p.f;
  const T* xg = reinterpret_cast<const T*>(p.x.ptr);
  const T* gyp = reinterpret_cast<const T*>(gy.ptr);
  const int64_t npix = (int64_t)p.B * p.H * p.W;                    // phase pixels (m, n) per image = H x W
  const int64_t beg = (int64_t)blockIdx.x * px_per_block, end = beg + px_per_block < npix ? beg + px_per_block : npix;
  float acc[2][2][VEC];
#pragma unroll
  for (int dy = 0; dy < 2; ++dy)
#pragma unroll
    for (int dx = 0; dx < 2; ++dx)
#pragma unroll
      for (int j = 0; j < VEC; ++j) acc[dy][dx][j] = 0.f;
  for (int64_t q = beg + plane; q < end; q += nlanes) {
    int t = (int)q;
    const int n = t % p.W; t /= p.W;
    const int m = t % p.H;
    const int b = t / p.H;
    const int oy = m * p.f + py, ox = n * p.f + pxp;
    float g[VEC];
    Vec<T, VEC>::load(gyp + (int64_t)b * gy.sb + (int64_t)oy * gy.sh + (int64_t)ox * gy.sw + c0, g);
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int iy = m + uy - dy;
      if (iy < 0 || iy >= p.H) continue;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int ix = n + ux - dx;
        if (ix < 0 || ix >= p.W) continue;
        float v[VEC];
        Vec<T, VEC>::load(xg + (int64_t)b * p.x.sb + (int64_t)iy * p.x.sh + (int64_t)ix * p.x.sw + c0, v);
#pragma unroll
        for (int j = 0; j < VEC; ++j) acc[dy][dx][j] = fmaf(v[j], g[j], acc[dy][dx][j]);
      }
    }
  }
  // lanes with the same piece inside a warp: lane, lane + pieces, ... -> xor-reduce over the multiples of `pieces`
#pragma unroll
  for (int dy = 0; dy < 2; ++dy)
#pragma unroll
    for (int dx = 0; dx < 2; ++dx)
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        float a = acc[dy][dx][j];
        for (int o = pieces; o < 32; o <<= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        acc[dy][dx][j] = a;
      }
  if (partial) {
    // deterministic form: the 8 warps' sums are added in warp order, and the block's 4 taps x C values go to
    // partial[blockIdx.x][tap][C]; upsample_bwd_weight_finalize_kernel adds the blocks in order
    __shared__ float red[8][4][32 * VEC];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < pieces) {
#pragma unroll
      for (int dy = 0; dy < 2; ++dy)
#pragma unroll
        for (int dx = 0; dx < 2; ++dx)
#pragma unroll
          for (int j = 0; j < VEC; ++j) red[warp][dy * 2 + dx][lane * VEC + j] = acc[dy][dx][j];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 4 * p.C; i += 256) {
      const int t = i / p.C, c = i - t * p.C;
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w][t][c];
      const int dy = t >> 1, dx = t & 1;
      partial[((int64_t)blockIdx.x * K * K + (ky1 + dy * p.f) * K + kx1 + dx * p.f) * p.C + c] = s;
    }
    return;
  }
  if ((threadIdx.x & 31) < pieces) {
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float* dst = gw + (int64_t)((ky1 + dy * p.f) * K + kx1 + dx * p.f) * p.C + c0;
#pragma unroll
        for (int j = 0; j < VEC; ++j) atomicAdd(dst + j, acc[dy][dx][j]);
      }
  }
}

// gw[t][c] = sum over the pixel blocks, in block order
__global__ void upsample_bwd_weight_finalize_kernel(const float* __restrict__ partial, float* __restrict__ gw, int n_blocks, int n_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_out) return;
  float s = 0.f;
  for (int b = 0; b < n_blocks; ++b) s += partial[(int64_t)b * n_out + i];
  gw[i] = s;
}

// any strides / channel counts: one thread per output element
template <typename T>
__global__ void __launch_bounds__(256) upsample_generic_kernel(UpParams p) {
  const int64_t total = (int64_t)p.B * p.C * p.Ho * p.Wo;
  const int K = 2 * p.f, pad = p.f / 2;
  const T* xg = reinterpret_cast<const T*>(p.x.ptr);
  const T* sg = reinterpret_cast<const T*>(p.skip.ptr);
  T* yg = reinterpret_cast<T*>(p.y.ptr);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t t = i;
    const int ox = (int)(t % p.Wo); t /= p.Wo;
    const int oy = (int)(t % p.Ho); t /= p.Ho;
    const int c = (int)(t % p.C);
    const int b = (int)(t / p.C);
    const int iy1 = (oy + pad) / p.f, ky1 = (oy + pad) - iy1 * p.f;
    const int ix1 = (ox + pad) / p.f, kx1 = (ox + pad) - ix1 * p.f;
    float acc = sg ? ld_as_float(sg + (int64_t)b * p.skip.sb + (int64_t)c * p.skip.sc + (int64_t)oy * p.skip.sh + (int64_t)ox * p.skip.sw) : 0.f;
    for (int dy = 0; dy < 2; ++dy) {
      const int iy = iy1 - dy, ky = ky1 + dy * p.f;
      if (iy < 0 || iy >= p.H) continue;
      for (int dx = 0; dx < 2; ++dx) {
        const int ix = ix1 - dx, kx = kx1 + dx * p.f;
        if (ix < 0 || ix >= p.W) continue;
        acc = fmaf(ld_as_float(xg + (int64_t)b * p.x.sb + (int64_t)c * p.x.sc + (int64_t)iy * p.x.sh + (int64_t)ix * p.x.sw),
                   __ldg(p.w_tapmajor ? p.w + (int64_t)(ky * K + kx) * p.C + c : p.w + ((int64_t)c * K + ky) * K + kx), acc);
      }
    }
    st_from_float(yg + (int64_t)b * p.y.sb + (int64_t)c * p.y.sc + (int64_t)oy * p.y.sh + (int64_t)ox * p.y.sw, acc);
  }
}

bool vec_ok(const cpoly_tensor4& t, int vec_bytes_elems) {
  return t.ptr == nullptr || (t.sc == 1 && (reinterpret_cast<uintptr_t>(t.ptr) & 15) == 0 && t.sw % vec_bytes_elems == 0 &&
                              t.sh % vec_bytes_elems == 0 && t.sb % vec_bytes_elems == 0);
}

}  // namespace

int idaup_upsample_launch(const cpoly_upsample_args& a, cudaStream_t stream) {
  UpParams p;
  p.B = a.B; p.C = a.C; p.H = a.H; p.W = a.W; p.f = a.f; p.Ho = a.H * a.f; p.Wo = a.W * a.f;
  p.x = a.x; p.skip = a.skip; p.y = a.y; p.w = a.weight; p.w_tapmajor = a.weight_tap_major;
  const bool bf = a.dtype == CPOLY_BF16;
  const int vec = bf ? 8 : 4;
  const bool fast = a.C % vec == 0 && vec_ok(a.x, vec) && vec_ok(a.skip, vec) && vec_ok(a.y, vec);
  const int64_t total = (int64_t)a.B * p.Ho * p.Wo * (fast ? a.C / vec : a.C);
  const unsigned blocks = (unsigned)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  static const bool quad_on = [] { const char* e = getenv("CPOLY_UPSAMPLE_QUAD"); return !(e && e[0] == '0'); }();
  if (quad_on && fast && a.weight_tap_major && a.f % 2 == 0 && (int64_t)a.B * (p.Ho / 2 + 1) * (p.Wo / 2 + 1) * (a.C / vec) < 0x7fffffffLL) {
    const int par = (a.f / 2) & 1;
    const int64_t quads = (int64_t)a.B * (p.Ho / 2 + par) * (p.Wo / 2 + par) * (a.C / vec);
    const unsigned qblocks = (unsigned)((quads + 255) / 256);   // one quad per thread: every load of a wave in flight at once
    if (bf) upsample_nhwc_quad_kernel<__nv_bfloat16, 8><<<qblocks, 256, 0, stream>>>(p);
    else upsample_nhwc_quad_kernel<float, 4><<<qblocks, 256, 0, stream>>>(p);
  } else if (fast) {
    if (bf) upsample_nhwc_kernel<__nv_bfloat16, 8><<<blocks, 256, 0, stream>>>(p);
    else upsample_nhwc_kernel<float, 4><<<blocks, 256, 0, stream>>>(p);
  } else {
    if (bf) upsample_generic_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(p);
    else upsample_generic_kernel<float><<<blocks, 256, 0, stream>>>(p);
  }
  return check_launch("idaup_upsample_kernel");
}

// pixel blocks of the deterministic weight gradient: about 4 CTAs per SM over all f^2 phases, >= 8 pixels per thread
int idaup_bwd_weight_blocks(int B, int H, int W, int C, int f, int dtype) {
  const int vec = dtype == CPOLY_BF16 ? 8 : 4;
  const int pieces = C / vec > 0 ? C / vec : 1;
  const int nlanes = 256 / pieces > 0 ? 256 / pieces : 1;
  const int64_t npix = (int64_t)B * H * W;
  int64_t g = (148 * 4) / (f * f);
  const int64_t cap = (npix + (int64_t)nlanes * 8 - 1) / ((int64_t)nlanes * 8);
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

size_t idaup_upsample_backward_workspace(const cpoly_upsample_bwd_args& a) {
  return (size_t)idaup_bwd_weight_blocks(a.B, a.H, a.W, a.C, a.f, a.dtype) * 4 * a.f * a.f * a.C * sizeof(float) + 256;
}

int idaup_upsample_backward_launch(const cpoly_upsample_bwd_args& a, cudaStream_t stream) {
  UpParams p;
  p.B = a.B; p.C = a.C; p.H = a.H; p.W = a.W; p.f = a.f; p.Ho = a.H * a.f; p.Wo = a.W * a.f;
  p.x = a.x; p.skip = cpoly_tensor4{nullptr, 0, 0, 0, 0}; p.y = cpoly_tensor4{nullptr, 0, 0, 0, 0}; p.w = a.weight; p.w_tapmajor = 1;
  const bool bf = a.dtype == CPOLY_BF16;
  const int vec = bf ? 8 : 4;
  const int pieces = a.C / vec;
  const bool ok = a.C % vec == 0 && pieces >= 1 && pieces <= 32 && (pieces & (pieces - 1)) == 0 && vec_ok(a.x, vec) && vec_ok(a.gy, vec) &&
                  vec_ok(a.gx, vec) && a.f % 2 == 0 && (int64_t)a.B * p.Ho * p.Wo * pieces < 0x7fffffffLL;
  if (!ok) {
    set_error("idaup_upsample_backward needs channels_last tensors with C/%d a power of two <= 32 and an even f", vec);
    return CPOLY_EUNSUPPORTED;
  }
  if (a.gx.ptr) {
    const int64_t total = (int64_t)a.B * a.H * a.W * pieces;
    const unsigned blocks = (unsigned)((total + 255) / 256);
    if (bf) upsample_bwd_input_kernel<__nv_bfloat16, 8><<<blocks, 256, 0, stream>>>(p, a.gy, a.gx);
    else upsample_bwd_input_kernel<float, 4><<<blocks, 256, 0, stream>>>(p, a.gy, a.gx);
    if (int rc = check_launch("upsample_bwd_input_kernel")) return rc;
  }
  if (a.gweight) {
    const int64_t npix = (int64_t)a.B * a.H * a.W;
    const int nlanes = 256 / pieces;
    const int n_out = 4 * a.f * a.f * a.C;
    int gx_blocks = idaup_bwd_weight_blocks(a.B, a.H, a.W, a.C, a.f, a.dtype);
    const bool det = a.workspace && a.workspace_bytes >= (size_t)gx_blocks * n_out * sizeof(float) && (reinterpret_cast<uintptr_t>(a.workspace) & 15) == 0;
    if (det) {
      // a few hundred CTAs in total, each summing its pixel range in registers: no atomics, fixed order (bit-reproducible)
      int px_per_block = (int)((npix + gx_blocks - 1) / gx_blocks);
      px_per_block = (px_per_block + nlanes - 1) / nlanes * nlanes;
      float* partial = reinterpret_cast<float*>(a.workspace);
      const dim3 grid((unsigned)gx_blocks, (unsigned)(a.f * a.f));
      if (bf) upsample_bwd_weight_kernel<__nv_bfloat16, 8><<<grid, 256, 0, stream>>>(p, a.gy, a.gweight, px_per_block, partial);
      else upsample_bwd_weight_kernel<float, 4><<<grid, 256, 0, stream>>>(p, a.gy, a.gweight, px_per_block, partial);
      if (int rc = check_launch("upsample_bwd_weight_kernel")) return rc;
      upsample_bwd_weight_finalize_kernel<<<(n_out + 255) / 256, 256, 0, stream>>>(partial, a.gweight, gx_blocks, n_out);
      return check_launch("upsample_bwd_weight_finalize_kernel");
    }
    int px_per_block = nlanes * 16;                     // 16 pixels per thread; fp32 atomics into the zero-initialised gweight
    const dim3 grid((unsigned)((npix + px_per_block - 1) / px_per_block), (unsigned)(a.f * a.f));
    if (bf) upsample_bwd_weight_kernel<__nv_bfloat16, 8><<<grid, 256, 0, stream>>>(p, a.gy, a.gweight, px_per_block, nullptr);
    else upsample_bwd_weight_kernel<float, 4><<<grid, 256, 0, stream>>>(p, a.gy, a.gweight, px_per_block, nullptr);
    if (int rc = check_launch("upsample_bwd_weight_kernel")) return rc;
  }
  return CPOLY_OK;
}

}  // namespace cpoly
