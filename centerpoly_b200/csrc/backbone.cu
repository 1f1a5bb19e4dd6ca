// Two bandwidth-bound pieces of the DLA-34 backbone that the measured 2048x1024 frame spends 19 % of its time in
// when left to the framework kernels (profiles/r01j_frame_bf16_launches.txt: 330 us for the stem in a legacy
// tf32 kernel + layout conversions, 219 us for six NHWC max-pools):
//
//   stem_conv7x7_kernel   <- DLA.base_layer: Conv2d(3,16,7,stride 1,pad 3) + BatchNorm (folded) + ReLU
//                            src/lib/models/networks/pose_dla_dcn.py:231-236, :289
//   maxpool2x2_nhwc_kernel <- Tree.downsample = nn.MaxPool2d(2, stride=2)   pose_dla_dcn.py:201, :207
//
// Stem: implicit GEMM on the warp-level tensor-core path (mma.sync m16n8k16, bf16 in / fp32 accumulate).  N = 16
// output channels is far too narrow for tcgen05 to matter and the op is bound by the frame read + 67 MB write, not
// by the tensor pipe.  The fp32 (or bf16) frame is read ONCE with whatever strides it has (NCHW or channels_last),
// rounded to bf16 (what autocast does) into a shared window with 4-channel pixels, so that the K index
// k = ky*32 + kx*4 + c of the im2col row of an output pixel is a contiguous run of the window row: A fragments are
// plain 32-bit shared loads, no im2col buffer.  K = 7*32 = 224 (147 live); weights pre-packed [16][224] bf16.
#include <cuda_bf16.h>
#include <mma.h>

#include "common.cuh"

namespace cpoly {
namespace {

constexpr int ST = 32;              // output tile: ST x ST pixels
constexpr int WIN_H = ST + 6;       // rows y-3 .. y+ST+2
constexpr int WIN_W = ST + 12;      // cols x-4 .. x+ST+7: starts on a 16-byte boundary of an fp32 NHWC row (kx = 7 has zero weights but is read)
constexpr int WIN_X0 = 4;           // window column of the tile's first output column
constexpr int STEM_K = 224, STEM_N = 16;

__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

template <typename TOut>
__global__ void __launch_bounds__(256) stem_conv7x7_kernel(cpoly_tensor4 x, int in_dtype, const __nv_bfloat16* __restrict__ wpk,
                                                           const float* __restrict__ bias, TOut* __restrict__ y, int H, int W, int relu) {
  __shared__ __align__(16) __nv_bfloat16 win[WIN_H][WIN_W][4];
  __shared__ __align__(16) __nv_bfloat16 wsm[STEM_N][STEM_K + 8];   // +8: rows land on different banks
  const int b = blockIdx.z, ty0 = blockIdx.y * ST, tx0 = blockIdx.x * ST;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  {
    const uint4* src = reinterpret_cast<const uint4*>(wpk);                    // 16 rows x 28 uint4
    for (int e = tid; e < STEM_N * (STEM_K / 8); e += 256)
      *reinterpret_cast<uint4*>(&wsm[e / (STEM_K / 8)][(e % (STEM_K / 8)) * 8]) = __ldg(src + e);
  }
  // frame window -> bf16, zero outside the image (the conv's zero padding) and in the 4th channel
  for (int e = tid; e < WIN_H * WIN_W; e += 256) win[e / WIN_W][e % WIN_W][3] = __float2bfloat16_rn(0.0f);
  const int wx0 = tx0 - WIN_X0;
  const bool fast = in_dtype == CPOLY_F32 && x.sc == 1 && x.sw == 3 && (x.sh & 3) == 0 && (x.sb & 3) == 0 && wx0 >= 0 && wx0 + WIN_W <= W &&
                    (reinterpret_cast<uintptr_t>(x.ptr) & 15) == 0;
  if (fast) {
    // fp32 channels_last frame, interior columns: a window row is WIN_W*3 = 132 contiguous floats = 33 aligned float4
    constexpr int V = WIN_W * 3 / 4;
    const float* xb = reinterpret_cast<const float*>(x.ptr) + (int64_t)b * x.sb + (int64_t)wx0 * 3;
    for (int e = tid; e < WIN_H * V; e += 256) {
      const int r = e / V, j = e - r * V;
      const int yy = ty0 - 3 + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (yy >= 0 && yy < H) v = __ldg(reinterpret_cast<const float4*>(xb + (int64_t)yy * x.sh) + j);
      const float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int fl = 4 * j + q, px = fl / 3;
        win[r][px][fl - 3 * px] = __float2bfloat16_rn(f[q]);
      }
    }
  } else {
    for (int e = tid; e < WIN_H * WIN_W * 3; e += 256) {
      const int r = e / (WIN_W * 3), xc = e - r * (WIN_W * 3), cx = xc / 3, c = xc - cx * 3;
      const int yy = ty0 - 3 + r, xx = wx0 + cx;
      float v = 0.0f;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) v = ld_dyn(x.ptr, (int64_t)b * x.sb + (int64_t)c * x.sc + (int64_t)yy * x.sh + (int64_t)xx * x.sw, in_dtype);
      win[r][cx][c] = __float2bfloat16_rn(v);
    }
  }
  __syncthreads();

  const int g = lane >> 2, t = lane & 3;
  // B fragments for all 14 k-steps x 2 channel halves stay in registers
  uint32_t bf[14][2][2];
#pragma unroll
  for (int ks = 0; ks < 14; ++ks)
#pragma unroll
    for (int nh = 0; nh < 2; ++nh) {
      bf[ks][nh][0] = *reinterpret_cast<const uint32_t*>(&wsm[nh * 8 + g][ks * 16 + 2 * t]);
      bf[ks][nh][1] = *reinterpret_cast<const uint32_t*>(&wsm[nh * 8 + g][ks * 16 + 2 * t + 8]);
    }
  const float bz0 = bias ? bias[2 * t] : 0.f, bz1 = bias ? bias[2 * t + 1] : 0.f;
  const float bz8 = bias ? bias[8 + 2 * t] : 0.f, bz9 = bias ? bias[9 + 2 * t] : 0.f;

  // warp w owns rows 4w .. 4w+3 of the tile; an m16 tile = 16 consecutive pixels of one row.  Two rows are in
  // flight at once (4 independent accumulator chains per warp: the MMA chain latency, not issue, is the limit)
  for (int mp = 0; mp < 4; ++mp) {
    const int col0 = (mp & 1) * 16, rbase = warp * 4 + (mp >> 1) * 2;
    float accs[2][2][4];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int nh = 0; nh < 2; ++nh)
#pragma unroll
        for (int e = 0; e < 4; ++e) accs[u][nh][e] = 0.0f;
#pragma unroll
    for (int ks = 0; ks < 14; ++ks) {
      const int ky = ks >> 1, half = ks & 1;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        // k = half*16 + 2t (+8): offset in bf16 elements from the window pixel (row+ky, col0+g) channel 0
        const __nv_bfloat16* p0 = &win[rbase + u + ky][col0 + g + WIN_X0 - 3][0] + half * 16 + 2 * t;
        uint32_t a[4];
        a[0] = *reinterpret_cast<const uint32_t*>(p0);
        a[1] = *reinterpret_cast<const uint32_t*>(p0 + 8 * 4);
        a[2] = *reinterpret_cast<const uint32_t*>(p0 + 8);
        a[3] = *reinterpret_cast<const uint32_t*>(p0 + 8 * 4 + 8);
        mma_bf16(accs[u][0], a, bf[ks][0][0], bf[ks][0][1]);
        mma_bf16(accs[u][1], a, bf[ks][1][0], bf[ks][1][1]);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
    const int row = rbase + u;
    const float (&acc)[2][4] = accs[u];
    // epilogue: bias, ReLU; lane (g,t) holds channels {2t,2t+1} and {8+2t,9+2t} of pixels g and g+8
    float v[2][4];   // [pixel g | g+8][ch 2t, 2t+1, 8+2t, 9+2t]
    v[0][0] = acc[0][0] + bz0; v[0][1] = acc[0][1] + bz1; v[0][2] = acc[1][0] + bz8; v[0][3] = acc[1][1] + bz9;
    v[1][0] = acc[0][2] + bz0; v[1][1] = acc[0][3] + bz1; v[1][2] = acc[1][2] + bz8; v[1][3] = acc[1][3] + bz9;
    if (relu) {
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) v[i][e] = fmaxf(v[i][e], 0.0f);
    }
    const int yy = ty0 + row;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int xx = tx0 + col0 + g + 8 * i;
      if constexpr (sizeof(TOut) == 2) {
        // regroup so that lane t writes channels 4t .. 4t+3 (8 contiguous bytes; the 4 lanes of a pixel cover its 32 B)
        const uint32_t lo = pack2(v[i][0], v[i][1]), hi = pack2(v[i][2], v[i][3]);
        const int src0 = (lane & ~3) | ((2 * t) & 3), src1 = (lane & ~3) | ((2 * t + 1) & 3);
        const uint32_t lo0 = __shfl_sync(0xffffffffu, lo, src0), hi0 = __shfl_sync(0xffffffffu, hi, src0);
        const uint32_t lo1 = __shfl_sync(0xffffffffu, lo, src1), hi1 = __shfl_sync(0xffffffffu, hi, src1);
        const uint2 o = t < 2 ? make_uint2(lo0, lo1) : make_uint2(hi0, hi1);
        if (yy < H && xx < W)
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(y) + (((int64_t)b * H + yy) * W + xx) * STEM_N + 4 * t) = o;
      } else {
        if (yy < H && xx < W) {
          float* o = reinterpret_cast<float*>(y) + (((int64_t)b * H + yy) * W + xx) * STEM_N;
          *reinterpret_cast<float2*>(o + 2 * t) = make_float2(v[i][0], v[i][1]);
          *reinterpret_cast<float2*>(o + 8 + 2 * t) = make_float2(v[i][2], v[i][3]);
        }
      }
    }
    }
  }
}

// (16,3,7,7) fp32 or bf16 -> [16][224] bf16 with k = ky*32 + kx*4 + c
__global__ void stem_pack_weights_kernel(const void* w, int w_dtype, __nv_bfloat16* out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= STEM_N * STEM_K) return;
  const int n = e / STEM_K, k = e - n * STEM_K, ky = k >> 5, kx = (k & 31) >> 2, c = k & 3;
  float v = 0.0f;
  if (kx < 7 && c < 3) v = ld_dyn(w, ((n * 3 + c) * 7 + ky) * 7 + kx, w_dtype);
  out[e] = __float2bfloat16_rn(v);
}

// 2x2 / stride 2 max-pool over an NHWC tensor, 16 bytes (8 bf16 or 4 fp32 channels) per thread
template <typename T>
__global__ void __launch_bounds__(256) maxpool2x2_nhwc_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int H, int W, int vec_per_px, int64_t total) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const int Wo = W >> 1, Ho = H >> 1;
  const int v = (int)(i % vec_per_px);
  const int64_t px = i / vec_per_px;
  const int xo = (int)(px % Wo);
  const int64_t t2 = px / Wo;
  const int yo = (int)(t2 % Ho);
  const int64_t b = t2 / Ho;
  const int64_t row = ((b * H + 2 * yo) * W + 2 * xo) * vec_per_px + v;
  const uint4 a = __ldg(x + row), bq = __ldg(x + row + vec_per_px);
  const uint4 c = __ldg(x + row + (int64_t)W * vec_per_px), d = __ldg(x + row + (int64_t)W * vec_per_px + vec_per_px);
  uint4 o;
  if constexpr (sizeof(T) == 2) {
    auto mx = [](uint32_t p, uint32_t q) {
      const __nv_bfloat162 r = __hmax2_nan(*reinterpret_cast<const __nv_bfloat162*>(&p), *reinterpret_cast<const __nv_bfloat162*>(&q));
      return *reinterpret_cast<const uint32_t*>(&r);
    };
    o.x = mx(mx(a.x, bq.x), mx(c.x, d.x)); o.y = mx(mx(a.y, bq.y), mx(c.y, d.y));
    o.z = mx(mx(a.z, bq.z), mx(c.z, d.z)); o.w = mx(mx(a.w, bq.w), mx(c.w, d.w));
  } else {
    auto mx = [](uint32_t p, uint32_t q) {
      const float fp = __uint_as_float(p), fq = __uint_as_float(q);
      return __float_as_uint((fp != fp) ? fp : ((fq != fq) ? fq : fmaxf(fp, fq)));
    };
    o.x = mx(mx(a.x, bq.x), mx(c.x, d.x)); o.y = mx(mx(a.y, bq.y), mx(c.y, d.y));
    o.z = mx(mx(a.z, bq.z), mx(c.z, d.z)); o.w = mx(mx(a.w, bq.w), mx(c.w, d.w));
  }
  y[i] = o;
}

}  // namespace

int stem_pack_weights_launch(const void* w, int w_dtype, void* out, cudaStream_t stream) {
  stem_pack_weights_kernel<<<(STEM_N * STEM_K + 255) / 256, 256, 0, stream>>>(w, w_dtype, reinterpret_cast<__nv_bfloat16*>(out));
  return check_launch("stem_pack_weights_kernel");
}

int stem_conv_launch(const cpoly_stem_args& a, cudaStream_t stream) {
  const dim3 grid((a.W + ST - 1) / ST, (a.H + ST - 1) / ST, a.B);
  const __nv_bfloat16* wpk = reinterpret_cast<const __nv_bfloat16*>(a.weight_packed);
  if (a.out_dtype == CPOLY_BF16)
    stem_conv7x7_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(a.x, a.in_dtype, wpk, a.bias, reinterpret_cast<__nv_bfloat16*>(a.y), a.H, a.W, a.relu);
  else
    stem_conv7x7_kernel<float><<<grid, 256, 0, stream>>>(a.x, a.in_dtype, wpk, a.bias, reinterpret_cast<float*>(a.y), a.H, a.W, a.relu);
  return check_launch("stem_conv7x7_kernel");
}

int maxpool2x2_launch(const cpoly_maxpool_args& a, cudaStream_t stream) {
  const int esz = a.dtype == CPOLY_BF16 ? 2 : 4;
  const int vec_per_px = a.C * esz / 16;
  const int64_t total = (int64_t)a.B * (a.H / 2) * (a.W / 2) * vec_per_px;
  if (total == 0) return CPOLY_OK;
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (a.dtype == CPOLY_BF16)
    maxpool2x2_nhwc_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<uint4*>(a.y), a.H, a.W, vec_per_px, total);
  else
    maxpool2x2_nhwc_kernel<float><<<blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<uint4*>(a.y), a.H, a.W, vec_per_px, total);
  return check_launch("maxpool2x2_nhwc_kernel");
}

}  // namespace cpoly
