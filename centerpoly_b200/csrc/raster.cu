// Depth-ordered instance rasterisation (cpoly_rasterise): the mask construction of
// `Cityscapes.format_and_write_to_cityscapes`, /root/reference/src/lib/datasets/dataset/cityscapes.py:226-279.
//
// The reference scan-converts every polygon with PIL (fill + outline), stamps a radius-2 disc on every point of the
// Bresenham contour, and then runs ~6 full-image numpy passes per instance for the occlusion bookkeeping.  Here the
// mask is evaluated per pixel in closed form, so no instance ever needs a second pass over the image:
//   covered(p)  =  even-odd crossing count of the ray p -> +x over the integer polygon (pixel centres)
//               OR  p within the 21-pixel disc (5x5 minus corners = PIL's radius-2 ellipse) of a Bresenham step,
//                   the step's minor coordinate being floor((2*dmin*k + dmaj) / (2*dmaj)) for major step k.
// Both were checked against PIL on random polygons (tests/test_oracle_raster.py).  One thread owns PX=8 horizontally
// adjacent pixels and walks the instances nearest-first, carrying the 8 "already removed" bits in a register, so
// the occlusion rule costs nothing and the only HBM traffic is the (B,K,H,W) byte masks, written once with 8-byte
// stores (a warp writes 256 contiguous bytes per instance).  Algorithmic bytes: count*H*W per image (HBM writes).
// Measured (tools/raster_bench.py, profiles/r01l_*): 58-63 % of the HBM peak on a batch of 8 frames; a single frame
// (one wave of CTAs) is bound by the serial instruction chain of the warps that own the most crowded row segments.
#include <limits.h>
#include <stdlib.h>

#include "common.cuh"

namespace cpoly {
namespace {

constexpr int TILE_H = 8;         // 8 warps per CTA
constexpr int VCLAMP = 8192;      // vertices are clamped to [-VCLAMP, VCLAMP-1] so every product below fits in 32 bits
constexpr int FLAG_DRAWN = 1, FLAG_OCCLUDER = 2;

struct RasterParams {
  int B, K, N, W, H;
  double thresh, occluder_score;
  int min_area;
  uint32_t skip_classes;
  const float* dets;
  int32_t* order;
  int32_t* count;
  int16_t* verts;
  int32_t* area;
  uint8_t* keep;
  double* conf;
  int32_t* status;
  uint8_t* masks;
  int4* bbox;      // workspace: dilated integer bounding box per slot (empty when the class is not drawn)
  int32_t* flags;  // workspace
};

// int(float("{:.2f}".format(x))) of cityscapes.py:120-121,231-232: decimal rounding to two places, then truncation.
// The integer only differs from trunc(x) when the fraction rounds up to 1.00, i.e. frac(|x|) > 0.995 (no fp32 value
// sits between the double nearest to 0.995 and 0.995 itself, so the comparison in fp64 is exact; sign symmetric).
__device__ __forceinline__ int round2_trunc(float x) {
  const double a = fabs((double)x);
  if (!(a < 1.0e9)) return a != a ? 0 : (x < 0.f ? -1000000000 : 1000000000);
  const double fl = floor(a);
  const int n = (int)fl + ((a - fl) > 0.995 ? 1 : 0);
  return x < 0.f ? -n : n;
}

// ---- 1. filter, depth sort, vertex truncation ------------------------------------------------------------------
// One CTA per image, one thread per detection.  Rank sort on (depth, class, index): Python's sorted() is stable and
// the reference builds its list class by class (cityscapes.py:226-239), so equal depths keep (class, index) order.
__global__ void __launch_bounds__(256) raster_prep_kernel(const RasterParams p) {
  __shared__ float s_depth[256];
  __shared__ int s_cls[256];
  __shared__ uint8_t s_valid[256];
  __shared__ int s_clamped;
  const int b = blockIdx.x, i = threadIdx.x;
  const int row_len = 2 * p.N + 7;
  const float* row = p.dets + ((size_t)b * p.K + (i < p.K ? i : 0)) * row_len;
  const float score = row[4];
  const bool valid = i < p.K && (double)score > p.thresh;
  float depth = row[row_len - 1];
  if (depth != depth) depth = __int_as_float(0x7f800000);   // NaN sorts last
  const int cls = (int)row[5];
  s_depth[i] = depth; s_cls[i] = cls; s_valid[i] = valid;
  if (i == 0) s_clamped = 0;
  const int n = __syncthreads_count(valid);
  const size_t base = (size_t)b * p.K;
  if (valid) {
    int rank = 0;
    for (int j = 0; j < p.K; ++j) {
      if (!s_valid[j]) continue;
      const float dj = s_depth[j];
      const int cj = s_cls[j];
      rank += (dj < depth) || (dj == depth && (cj < cls || (cj == cls && j < i)));
    }
    const bool drawn = cls < 0 || cls > 31 || !((p.skip_classes >> cls) & 1u);
    int xmin = INT_MAX, xmax = INT_MIN, ymin = INT_MAX, ymax = INT_MIN, clamped = 0;
    int16_t* v = p.verts + (base + rank) * p.N * 2;
    for (int k = 0; k < p.N; ++k) {
      const int x = round2_trunc(row[6 + 2 * k]), y = round2_trunc(row[7 + 2 * k]);
      const int cx = min(max(x, -VCLAMP), VCLAMP - 1), cy = min(max(y, -VCLAMP), VCLAMP - 1);
      clamped += (cx != x) + (cy != y);
      v[2 * k] = (int16_t)cx; v[2 * k + 1] = (int16_t)cy;
      xmin = min(xmin, cx); xmax = max(xmax, cx); ymin = min(ymin, cy); ymax = max(ymax, cy);
    }
    if (clamped) atomicAdd(&s_clamped, clamped);
    p.order[base + rank] = i;
    p.bbox[base + rank] = drawn ? make_int4(xmin - 2, ymin - 2, xmax + 2, ymax + 2) : make_int4(1, 1, 0, 0);
    p.flags[base + rank] = (drawn ? FLAG_DRAWN : 0) | ((double)score >= p.occluder_score ? FLAG_OCCLUDER : 0);
    const double c = (double)score * 1.2;
    p.conf[base + rank] = c < 1.0 ? c : 1.0;
  }
  if (i < p.K) {
    p.area[base + i] = 0;
    if (i >= n) {
      p.order[base + i] = -1; p.conf[base + i] = 0.0; p.flags[base + i] = 0; p.bbox[base + i] = make_int4(1, 1, 0, 0);
      int16_t* v = p.verts + (base + i) * p.N * 2;
      for (int k = 0; k < 2 * p.N; ++k) v[k] = 0;
    }
  }
  __syncthreads();
  if (i == 0) { p.count[b] = n; p.status[b] = s_clamped; }
}

// ---- 2. per-pixel coverage ---------------------------------------------------------------------------------------
// A warp owns one row segment of 32*PX pixels; lane l owns pixels X0..X0+PX-1.  The per-edge work that only depends
// on the row (does the edge cross this scanline, where, is it within 2 rows) is done ONCE per warp with lane e
// handling edge e; only the few edges that matter for the row are then broadcast and applied to every lane's pixels.
template <int PX>
__device__ __forceinline__ uint32_t span_bits(int lo, int hi) {   // bits lo..hi of the thread's PX pixels, clipped
  if (hi < 0 || lo > PX - 1) return 0u;
  lo = max(lo, 0); hi = min(hi, PX - 1);
  return ((2u << hi) - 1u) & ~((1u << lo) - 1u);
}

// discs around the Bresenham steps of edge (x0,y0)->(x1,y1) (walked in that direction, as the reference does)
template <int PX>
__device__ __forceinline__ uint32_t near_bits(int x0, int y0, int x1, int y1, int X0, int Y) {
  const int exmin = min(x0, x1), exmax = max(x0, x1), eymin = min(y0, y1), eymax = max(y0, y1);
  if (X0 + PX - 1 < exmin - 2 || X0 > exmax + 2) return 0u;
  uint32_t near = 0u;
  const int adx = exmax - exmin, ady = eymax - eymin;
  const int sx = x1 > x0 ? 1 : -1, sy = y1 > y0 ? 1 : -1;
  if (adx > ady) {   // x-major: one step per column, row = y0 + sy*floor((2*ady*k + adx) / (2*adx))
    const int c_lo = max(X0 - 2, exmin), c_hi = min(X0 + PX + 1, exmax);
    const int k0 = sx > 0 ? c_lo - x0 : x0 - c_hi;
    const int num = 2 * ady * k0 + adx;
    int m = num / (2 * adx), r = num - m * 2 * adx;
    for (int k = k0; k <= k0 + (c_hi - c_lo); ++k) {
      const int cx = x0 + sx * k, d = abs(Y - (y0 + sy * m));
      if (d <= 2) { const int rad = d == 2 ? 1 : 2; near |= span_bits<PX>(cx - rad - X0, cx + rad - X0); }
      r += 2 * ady;
      if (r >= 2 * adx) { r -= 2 * adx; ++m; }
    }
  } else {           // y-major (and the single-point edge): one step per row
    const int r_lo = max(Y - 2, eymin), r_hi = min(Y + 2, eymax);
    const int k0 = sy > 0 ? r_lo - y0 : y0 - r_hi;
    int m = 0, r = 0;
    if (ady > 0) { const int num = 2 * adx * k0 + ady; m = num / (2 * ady); r = num - m * 2 * ady; }
    for (int k = k0; k <= k0 + (r_hi - r_lo); ++k) {
      const int cx = x0 + sx * m, d = abs(Y - (y0 + sy * k));
      const int rad = d == 2 ? 1 : 2;
      near |= span_bits<PX>(cx - rad - X0, cx + rad - X0);
      r += 2 * adx;
      if (ady > 0 && r >= 2 * ady) { r -= 2 * ady; ++m; }
    }
  }
  return near;
}

// coverage bits of pixels (X0..X0+PX-1, Y) for the polygon whose N vertices (x | y<<16) sit at sv[0..N); whole warp
template <int PX>
__device__ __forceinline__ uint32_t cover_bits(const uint32_t* sv, int N, int X0, int Y, int lane, int XW0) {
  constexpr uint32_t FULL = 0xffffffffu;
  uint32_t inside = 0u, near = 0u;
  for (int e0 = 0; e0 < N; e0 += 32) {
    const int e = e0 + lane;
    const bool have = e < N;
    const uint32_t cv = have ? sv[e] : 0u, pv = have ? sv[e == 0 ? N - 1 : e - 1] : 0u;
    const int x0 = (int16_t)(pv & 0xffffu), y0 = (int16_t)(pv >> 16);
    const int x1 = (int16_t)(cv & 0xffffu), y1 = (int16_t)(cv >> 16);
    // even-odd rule on pixel centres: the edge crosses this scanline (half-open in y) and pixel X lies left of it:
    //   (Y-y0)*dx - (X-x0)*dy > 0 (dy > 0)   <=>   X*dy < (Y-y0)*dx + x0*dy   <=>   X < ceil(num / dy) =: xc
    const bool cross = have && ((y0 <= Y) != (y1 <= Y));
    int xc = 0;
    if (cross) {
      int dy = y1 - y0, num = (Y - y0) * (x1 - x0) + x0 * dy;
      if (dy < 0) { dy = -dy; num = -num; }
      const int t = num + dy - 1;
      xc = t / dy;
      if (t % dy < 0) --xc;   // floor division
    }
    for (uint32_t cm = __ballot_sync(FULL, cross); cm; cm &= cm - 1) {
      const int t = __shfl_sync(FULL, xc, __ffs(cm) - 1);
      inside ^= span_bits<PX>(0, t - 1 - X0);
    }
    const bool close = have && Y >= min(y0, y1) - 2 && Y <= max(y0, y1) + 2 && max(x0, x1) + 2 >= XW0 && min(x0, x1) - 2 <= XW0 + 32 * PX - 1;
    for (uint32_t nm = __ballot_sync(FULL, close); nm; nm &= nm - 1) {
      const int src = __ffs(nm) - 1;
      const uint32_t a = __shfl_sync(FULL, pv, src), b = __shfl_sync(FULL, cv, src);
      near |= near_bits<PX>((int16_t)(a & 0xffffu), (int16_t)(a >> 16), (int16_t)(b & 0xffffu), (int16_t)(b >> 16), X0, Y);
    }
  }
  return inside | near;
}

__device__ __forceinline__ uint32_t bits_to_bytes(uint32_t b4) {   // 4 bits -> 4 bytes of 0x00 / 0xff
  return ((b4 * 0x00204081u) & 0x01010101u) * 0xffu;
}

template <int PX>
__global__ void __launch_bounds__(32 * TILE_H) raster_cover_kernel(const RasterParams p) {
  constexpr uint32_t FULL = 0xffffffffu;
  constexpr int TILE_W = 32 * PX;
  extern __shared__ uint4 smem_raw[];
  const int b = blockIdx.z;
  const int n = p.count[b];
  if (n == 0) return;
  int4* s_bbox = reinterpret_cast<int4*>(smem_raw);                      // K dilated boxes (.x > .z when not drawn)
  uint32_t* s_occ = reinterpret_cast<uint32_t*>(s_bbox + p.K);           // K occluder flags
  uint32_t* sv = s_occ + p.K;                                            // n * N packed vertices
  const int tx0 = blockIdx.x * TILE_W;
  const size_t base = (size_t)b * p.K;
  const uint32_t* gv = reinterpret_cast<const uint32_t*>(p.verts) + base * p.N;
  for (int t = threadIdx.x; t < n * p.N; t += blockDim.x) sv[t] = gv[t];
  for (int s = threadIdx.x; s < n; s += blockDim.x) {
    const int fl = p.flags[base + s];
    s_bbox[s] = (fl & FLAG_DRAWN) ? p.bbox[base + s] : make_int4(1, 1, 0, 0);
    s_occ[s] = (fl & FLAG_OCCLUDER) ? 1u : 0u;
  }
  __syncthreads();
  // the CTA's 8 rows are spread over the image (stride gridDim.y) so that every CTA gets the same mix of crowded
  // and empty regions: with one wave of CTAs the kernel ends when the slowest one does
  const int lane = threadIdx.x & 31, Y = blockIdx.y + (threadIdx.x >> 5) * gridDim.y, X0 = tx0 + lane * PX;
  if (Y >= p.H) return;
  constexpr uint32_t ALL = (1u << PX) - 1u;
  const uint32_t vm = X0 >= p.W ? 0u : (X0 + PX <= p.W ? ALL : (1u << (p.W - X0)) - 1u);
  // 0: no masks wanted, 1: one aligned PX-byte store per instance, 2: ragged width, byte stores
  const int mode = !p.masks ? 0 : ((p.W % PX == 0) ? (vm ? 1 : 0) : 2);
  uint8_t* out = p.masks ? p.masks + ((base * p.H + Y) * (size_t)p.W + X0) : nullptr;
  const size_t plane = (size_t)p.H * p.W;
  uint32_t removed = 0u;
  for (int g = 0; g < n; g += 32) {
    // which of the next 32 instances touch this row segment (lane j looks at instance g + j)
    bool h = false, o = false;
    if (g + lane < n) {
      const int4 bb = s_bbox[g + lane];
      h = bb.x <= bb.z && bb.x <= tx0 + TILE_W - 1 && bb.z >= tx0 && bb.y <= Y && bb.w >= Y;
      o = s_occ[g + lane] != 0u;
    }
    const uint32_t hit = __ballot_sync(FULL, h), occ = __ballot_sync(FULL, o);
    const int cnt = min(32, n - g);
    if (mode == 1 && hit == 0u) {   // nothing here: a run of zero stores
#pragma unroll 4
      for (int j = 0; j < cnt; ++j, out += plane) {
        if (PX == 8) __stcs(reinterpret_cast<uint2*>(out), make_uint2(0u, 0u));
        else __stcs(reinterpret_cast<uint4*>(out), make_uint4(0u, 0u, 0u, 0u));
      }
      continue;
    }
    for (int j = 0; j < cnt; ++j) {
      uint32_t m = 0u;
      if ((hit >> j) & 1u) {
        const int s = g + j;
        m = cover_bits<PX>(sv + s * p.N, p.N, X0, Y, lane, tx0) & ~removed & vm;
        if ((occ >> j) & 1u) removed |= m;
        const int c = __reduce_add_sync(FULL, __popc(m));
        if (lane == 0 && c) atomicAdd(p.area + base + s, c);
      }
      if (mode == 1) {
        if (PX == 8) {
          __stcs(reinterpret_cast<uint2*>(out), m ? make_uint2(bits_to_bytes(m & 15u), bits_to_bytes(m >> 4)) : make_uint2(0u, 0u));
        } else {
          __stcs(reinterpret_cast<uint4*>(out), m ? make_uint4(bits_to_bytes(m & 15u), bits_to_bytes((m >> 4) & 15u),
                                                              bits_to_bytes((m >> 8) & 15u), bits_to_bytes(m >> 12))
                                                  : make_uint4(0u, 0u, 0u, 0u));
        }
        out += plane;
      } else if (mode == 2) {
        for (int k = 0; k < PX; ++k)
          if ((vm >> k) & 1u) out[k] = ((m >> k) & 1u) ? 255 : 0;
        out += plane;
      }
    }
  }
}

// ---- 3. area filter (cityscapes.py:273) -----------------------------------------------------------------------------
__global__ void raster_keep_kernel(const RasterParams p) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p.B * p.K) return;
  p.keep[idx] = (p.flags[idx] & FLAG_DRAWN) && p.area[idx] > p.min_area;
}

}  // namespace

size_t raster_workspace_bytes(int B, int K) { return (size_t)B * K * (sizeof(int4) + sizeof(int32_t)) + 256; }

int raster_launch(const cpoly_raster_args& a, cudaStream_t stream) {
  RasterParams p;
  p.B = a.B; p.K = a.K; p.N = a.N; p.W = a.W; p.H = a.H;
  p.thresh = a.thresh; p.occluder_score = a.occluder_score; p.min_area = a.min_area; p.skip_classes = a.skip_classes;
  p.dets = a.dets; p.order = a.order; p.count = a.count; p.verts = a.verts; p.area = a.area; p.keep = a.keep;
  p.conf = a.conf; p.status = a.status; p.masks = a.masks;
  p.bbox = reinterpret_cast<int4*>(a.workspace);
  p.flags = reinterpret_cast<int32_t*>(p.bbox + (size_t)a.B * a.K);
  raster_prep_kernel<<<a.B, 256, 0, stream>>>(p);
  if (int rc = check_launch("raster_prep_kernel")) return rc;
  const size_t smem = (size_t)a.K * (a.N * 4 + sizeof(int4) + sizeof(uint32_t));
  // 8 pixels per thread (8-byte stores); CPOLY_RASTER_PX=16 selects 16 (16-byte stores, needs W % 16 == 0 and 16-byte masks)
  const char* env = getenv("CPOLY_RASTER_PX");
  const bool px16 = env && atoi(env) == 16 && a.W % 16 == 0 && (reinterpret_cast<uintptr_t>(a.masks) & 15) == 0;
  auto kernel = px16 ? raster_cover_kernel<16> : raster_cover_kernel<8>;
  if (smem > 48 * 1024) CPOLY_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int tile_w = 32 * (px16 ? 16 : 8);
  const dim3 grid((a.W + tile_w - 1) / tile_w, (a.H + TILE_H - 1) / TILE_H, a.B);
  CPOLY_REQUIRE(grid.y <= 65535 && grid.z <= 65535, CPOLY_EINVAL, "raster grid too large");
  kernel<<<grid, 32 * TILE_H, smem, stream>>>(p);
  if (int rc = check_launch("raster_cover_kernel")) return rc;
  raster_keep_kernel<<<(a.B * a.K + 255) / 256, 256, 0, stream>>>(p);
  return check_launch("raster_keep_kernel");
}

}  // namespace cpoly
