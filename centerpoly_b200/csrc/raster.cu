// Depth-ordered instance rasterisation (cpoly_rasterise): the mask construction of
// `Cityscapes.format_and_write_to_cityscapes`, /root/reference/src/lib/datasets/dataset/cityscapes.py:226-279.
//
// The reference scan-converts every polygon with PIL (fill + outline), stamps a radius-2 disc on every point of the
// Bresenham contour, and then runs ~6 full-image numpy passes per instance for the occlusion bookkeeping.  Here the
// mask is evaluated per pixel in closed form, so no instance ever needs a second pass over the image:
//   covered(p)  =  even-odd crossing count of the ray p -> +x over the integer polygon (pixel centres)
//               OR  p within the 21-pixel disc (5x5 minus corners = PIL's radius-2 ellipse) of a Bresenham step,
//                   the step's minor coordinate being floor((2*dmin*k + dmaj) / (2*dmaj)) for major step k.
// Both were checked against PIL on random polygons (tests/test_oracle_raster.py).  One thread owns PX = 16 (or 8)
// horizontally adjacent pixels and walks the instances nearest-first, carrying the "already removed" bits in a
// register, so the occlusion rule costs nothing and the only HBM traffic is the (B,K,H,W) byte masks, written once
// with 16-byte stores (a warp writes 512 contiguous bytes per instance).  Algorithmic bytes: count*H*W per image.
// Measured (tools/raster_bench.py, profiles/r01l_*): 88 % of the measured HBM peak on a batch of 8 Cityscapes frames,
// 45 % on a single frame (one wave; launch + prep + keep are ~10 of its 68 us).
#include <limits.h>
#include <stdlib.h>

#include "common.cuh"

namespace cpoly {
namespace {

constexpr int TILE_H = 4;         // warps (= rows in flight) per CTA
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
  int32_t* clamped;  // workspace: clamped vertices per slot
  int rows_per_warp;
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
// Grid (B, ceil(K/8)), one WARP per detection (a single thread per detection walked its 2N strided vertex loads one
// DRAM round trip at a time: 30 us).  Every CTA reads the K (score, class, depth) triples of its image; the warp of
// detection i rank-sorts i on (depth, class, index) - Python's sorted() is stable and the reference builds its list
// class by class (cityscapes.py:226-239), so equal depths keep (class, index) order - and writes slot `rank`.
constexpr int PREP_WARPS = 8;
__global__ void __launch_bounds__(32 * PREP_WARPS) raster_prep_kernel(const RasterParams p) {
  constexpr uint32_t FULL = 0xffffffffu;
  __shared__ float s_depth[256];
  __shared__ int s_cls[256];
  __shared__ uint8_t s_valid[256];
  const int b = blockIdx.x, t = threadIdx.x, lane = t & 31;
  const int row_len = 2 * p.N + 7;
  const float* __restrict__ img = p.dets + (size_t)b * p.K * row_len;
  bool tv = false;
  if (t < p.K) {
    const float* row = img + (size_t)t * row_len;
    float depth = row[row_len - 1];
    if (depth != depth) depth = __int_as_float(0x7f800000);   // NaN sorts last
    tv = (double)row[4] > p.thresh;
    s_depth[t] = depth; s_cls[t] = (int)row[5]; s_valid[t] = tv;
  }
  const int n = __syncthreads_count(tv);
  const size_t base = (size_t)b * p.K;
  if (blockIdx.y == 0) {   // per-image bookkeeping and the unused slots
    if (t == 0) p.count[b] = n;
    if (t < p.K) {
      p.area[base + t] = 0;
      if (t >= n) {
        p.order[base + t] = -1; p.conf[base + t] = 0.0; p.flags[base + t] = 0; p.bbox[base + t] = make_int4(1, 1, 0, 0);
        p.clamped[base + t] = 0;
        uint32_t* v = reinterpret_cast<uint32_t*>(p.verts) + (base + t) * p.N;
        for (int k = 0; k < p.N; ++k) v[k] = 0u;
      }
    }
  }
  const int i = blockIdx.y * PREP_WARPS + (t >> 5);
  if (i >= p.K || !s_valid[i]) return;
  const float depth = s_depth[i];
  const int cls = s_cls[i];
  int rank = 0;
  for (int j = lane; j < p.K; j += 32) {
    const float dj = s_depth[j];
    const int cj = s_cls[j];
    rank += s_valid[j] && ((dj < depth) || (dj == depth && (cj < cls || (cj == cls && j < i))));
  }
  rank = __reduce_add_sync(FULL, rank);
  const float* row = img + (size_t)i * row_len;
  int xmin = INT_MAX, xmax = INT_MIN, ymin = INT_MAX, ymax = INT_MIN, clamped = 0;
  uint32_t* v = reinterpret_cast<uint32_t*>(p.verts) + (base + rank) * p.N;
  for (int k = lane; k < p.N; k += 32) {
    const int x = round2_trunc(row[6 + 2 * k]), y = round2_trunc(row[7 + 2 * k]);
    const int cx = min(max(x, -VCLAMP), VCLAMP - 1), cy = min(max(y, -VCLAMP), VCLAMP - 1);
    clamped += (cx != x) + (cy != y);
    v[k] = ((uint32_t)cx & 0xffffu) | ((uint32_t)cy << 16);
    xmin = min(xmin, cx); xmax = max(xmax, cx); ymin = min(ymin, cy); ymax = max(ymax, cy);
  }
  xmin = __reduce_min_sync(FULL, xmin); xmax = __reduce_max_sync(FULL, xmax);
  ymin = __reduce_min_sync(FULL, ymin); ymax = __reduce_max_sync(FULL, ymax);
  clamped = __reduce_add_sync(FULL, clamped);
  if (lane == 0) {
    const float score = row[4];
    const bool drawn = cls < 0 || cls > 31 || !((p.skip_classes >> cls) & 1u);
    p.order[base + rank] = i;
    p.bbox[base + rank] = drawn ? make_int4(xmin - 2, ymin - 2, xmax + 2, ymax + 2) : make_int4(1, 1, 0, 0);
    p.flags[base + rank] = (drawn ? FLAG_DRAWN : 0) | ((double)score >= p.occluder_score ? FLAG_OCCLUDER : 0);
    p.clamped[base + rank] = clamped;
    const double c = (double)score * 1.2;
    p.conf[base + rank] = c < 1.0 ? c : 1.0;
  }
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

// floor(a / b) for b > 0 when the quotient is known to be small (|q| < 2^16): one approximate fp32 division and a
// +-1 fix-up instead of the ~25-instruction integer division sequence
__device__ __forceinline__ int floor_div(int a, int b) {
  int q = __float2int_rd(__fdividef((float)a, (float)b));
  const int r = a - q * b;
  if (r < 0) --q; else if (r >= b) ++q;
  return q;
}
__device__ __forceinline__ int ceil_div(int a, int b) { return -floor_div(-a, b); }

// Steps k in [0, adx] of an x-major edge whose row offset m(k) = floor((2*ady*k + adx) / (2*adx)) lies in [ta, tb]
// (m is non-decreasing and moves by at most 1 per step).  Empty => lo > hi.
__device__ __forceinline__ void step_range(int adx, int ady, int ta, int tb, int& lo, int& hi) {
  ta = max(ta, 0); tb = min(tb, ady);   // 0 <= m(k) <= ady: keeps the quotients below within [0, adx + 1]
  lo = 1; hi = 0;
  if (tb < ta) return;
  if (ady == 0) { if (ta == 0) { lo = 0; hi = adx; } return; }
  lo = max(0, ceil_div(2 * adx * ta - adx, 2 * ady));
  hi = min(adx, ceil_div(2 * adx * (tb + 1) - adx, 2 * ady) - 1);
}

// Pixels of row Y covered by the radius-2 discs (5x5 minus corners) on the Bresenham steps of edge (x0,y0)->(x1,y1),
// walked in that direction as the reference does: always ONE span [xl, xr] (steps within one row of Y reach 2 pixels
// sideways, steps two rows away reach 1, and consecutive steps move by at most one pixel).  Empty => xl > xr.
__device__ __forceinline__ void near_span(int x0, int y0, int x1, int y1, int Y, int& xl, int& xr) {
  const int adx = abs(x1 - x0), ady = abs(y1 - y0);
  const int sx = x1 > x0 ? 1 : -1, sy = y1 > y0 ? 1 : -1;
  xl = 1; xr = 0;
  if (adx > ady) {   // x-major: step k sits at (x0 + sx*k, y0 + sy*m(k))
    const int t = (Y - y0) * sy;
    int alo, ahi, clo, chi;
    step_range(adx, ady, t - 2, t + 2, alo, ahi);
    if (alo > ahi) return;
    step_range(adx, ady, t - 1, t + 1, clo, chi);
    int klo = alo - 1, khi = ahi + 1;
    if (clo <= chi) { klo = min(klo, clo - 2); khi = max(khi, chi + 2); }
    xl = sx > 0 ? x0 + klo : x0 - khi;
    xr = sx > 0 ? x0 + khi : x0 - klo;
  } else {           // y-major (and the single-point edge): step k sits at (x0 + sx*m(k), y0 + sy*k), one per row
    int lo = INT_MAX, hi = INT_MIN;
    const int k0 = max((Y - y0) * sy - 2, 0), k1 = min((Y - y0) * sy + 2, ady);   // the rows Y-2..Y+2 that hold a step
    if (k0 <= k1) {
      int m = 0, r = 0;
      if (ady) { const int num = 2 * adx * k0 + ady; m = floor_div(num, 2 * ady); r = num - m * 2 * ady; }
      for (int k = k0; k <= k1; ++k) {
        const int d = k - (Y - y0) * sy;
        const int cx = x0 + sx * m, rad = (d == 2 || d == -2) ? 1 : 2;
        lo = min(lo, cx - rad); hi = max(hi, cx + rad);
        r += 2 * adx;
        if (r >= 2 * ady && ady) { r -= 2 * ady; ++m; }
      }
    }
    if (lo <= hi) { xl = lo; xr = hi; }
  }
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
      xc = ceil_div(num, dy);
    }
    for (uint32_t cm = __ballot_sync(FULL, cross && xc > XW0); cm; cm &= cm - 1) {
      const int t = __shfl_sync(FULL, xc, __ffs(cm) - 1);
      inside ^= span_bits<PX>(0, t - 1 - X0);
    }
    int xl = 1, xr = 0;
    if (have && Y >= min(y0, y1) - 2 && Y <= max(y0, y1) + 2) near_span(x0, y0, x1, y1, Y, xl, xr);
    for (uint32_t nm = __ballot_sync(FULL, xl <= xr && xr >= XW0 && xl <= XW0 + 32 * PX - 1); nm; nm &= nm - 1) {
      const int src = __ffs(nm) - 1;
      const int l = __shfl_sync(FULL, xl, src), r = __shfl_sync(FULL, xr, src);
      near |= span_bits<PX>(l - X0, r - X0);
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
  uint32_t* sv = s_occ + ((p.K + 3) & ~3);                                          // n * N packed vertices
  const int tx0 = blockIdx.x * TILE_W;
  const size_t base = (size_t)b * p.K;
  const uint32_t* gv = reinterpret_cast<const uint32_t*>(p.verts) + base * p.N;
  // all copies in flight at once (a plain load/store loop walked one L2 round trip per iteration: ~15 % of a frame)
  if ((p.N & 3) == 0 && (reinterpret_cast<uintptr_t>(gv) & 15) == 0) {
    for (int t = threadIdx.x; t < n * p.N / 4; t += blockDim.x)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(sv + 4 * t)), "l"(gv + 4 * t));
  } else {
    for (int t = threadIdx.x; t < n * p.N; t += blockDim.x)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(sv + t)), "l"(gv + t));
  }
  for (int s = threadIdx.x; s < n; s += blockDim.x) {
    const int fl = p.flags[base + s];
    s_bbox[s] = (fl & FLAG_DRAWN) ? p.bbox[base + s] : make_int4(1, 1, 0, 0);
    s_occ[s] = (fl & FLAG_OCCLUDER) ? 1u : 0u;
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();
  // The CTA's rows are spread over the image (stride gridDim.y) so that every CTA sees the same mix of crowded and
  // empty regions; a warp takes rows_per_warp of them, chosen by the host so that a small batch is ONE balanced wave.
  const int lane = threadIdx.x & 31, X0 = tx0 + lane * PX;
  for (int rr = 0; rr < p.rows_per_warp; ++rr) {
  const int Y = blockIdx.y + ((threadIdx.x >> 5) + TILE_H * rr) * gridDim.y;
  if (Y >= p.H) break;
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
      // four independent address registers per round: with one pointer bumped in place every store waits for the
      // previous one to leave the LSU queue before its address register may be overwritten
      int j = 0;
      for (; j + 4 <= cnt; j += 4) {
        uint8_t* o0 = out + (size_t)j * plane;
        uint8_t* o1 = o0 + plane;
        uint8_t* o2 = o1 + plane;
        uint8_t* o3 = o2 + plane;
        if (PX == 8) {
          __stcs(reinterpret_cast<uint2*>(o0), make_uint2(0u, 0u)); __stcs(reinterpret_cast<uint2*>(o1), make_uint2(0u, 0u));
          __stcs(reinterpret_cast<uint2*>(o2), make_uint2(0u, 0u)); __stcs(reinterpret_cast<uint2*>(o3), make_uint2(0u, 0u));
        } else {
          __stcs(reinterpret_cast<uint4*>(o0), make_uint4(0u, 0u, 0u, 0u)); __stcs(reinterpret_cast<uint4*>(o1), make_uint4(0u, 0u, 0u, 0u));
          __stcs(reinterpret_cast<uint4*>(o2), make_uint4(0u, 0u, 0u, 0u)); __stcs(reinterpret_cast<uint4*>(o3), make_uint4(0u, 0u, 0u, 0u));
        }
      }
      for (; j < cnt; ++j) {
        uint8_t* o0 = out + (size_t)j * plane;
        if (PX == 8) __stcs(reinterpret_cast<uint2*>(o0), make_uint2(0u, 0u));
        else __stcs(reinterpret_cast<uint4*>(o0), make_uint4(0u, 0u, 0u, 0u));
      }
      out += (size_t)cnt * plane;
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
      uint8_t* o = out + (size_t)j * plane;
      if (mode == 1) {
        if (PX == 8) {
          __stcs(reinterpret_cast<uint2*>(o), m ? make_uint2(bits_to_bytes(m & 15u), bits_to_bytes(m >> 4)) : make_uint2(0u, 0u));
        } else {
          __stcs(reinterpret_cast<uint4*>(o), m ? make_uint4(bits_to_bytes(m & 15u), bits_to_bytes((m >> 4) & 15u),
                                                            bits_to_bytes((m >> 8) & 15u), bits_to_bytes(m >> 12))
                                                : make_uint4(0u, 0u, 0u, 0u));
        }
      } else if (mode == 2) {
        for (int k = 0; k < PX; ++k)
          if ((vm >> k) & 1u) o[k] = ((m >> k) & 1u) ? 255 : 0;
      }
    }
    if (mode) out += (size_t)cnt * plane;
  }
  }
}

// ---- 3. area filter (cityscapes.py:273) and the per-image count of clamped vertices -----------------------------------
__global__ void __launch_bounds__(256) raster_keep_kernel(const RasterParams p) {
  const int b = blockIdx.x, t = threadIdx.x;
  int c = 0;
  if (t < p.K) {
    const size_t idx = (size_t)b * p.K + t;
    p.keep[idx] = (p.flags[idx] & FLAG_DRAWN) && p.area[idx] > p.min_area;
    c = p.clamped[idx];
  }
  __shared__ int s_sum[8];
  c = __reduce_add_sync(0xffffffffu, c);
  if ((t & 31) == 0) s_sum[t >> 5] = c;
  __syncthreads();
  if (t == 0) {
    int tot = 0;
    for (int w = 0; w < 8; ++w) tot += s_sum[w];
    p.status[b] = tot;
  }
}

}  // namespace

size_t raster_workspace_bytes(int B, int K) { return (size_t)B * K * (sizeof(int4) + 2 * sizeof(int32_t)) + 256; }

int raster_launch(const cpoly_raster_args& a, cudaStream_t stream) {
  RasterParams p;
  p.B = a.B; p.K = a.K; p.N = a.N; p.W = a.W; p.H = a.H;
  p.thresh = a.thresh; p.occluder_score = a.occluder_score; p.min_area = a.min_area; p.skip_classes = a.skip_classes;
  p.dets = a.dets; p.order = a.order; p.count = a.count; p.verts = a.verts; p.area = a.area; p.keep = a.keep;
  p.conf = a.conf; p.status = a.status; p.masks = a.masks;
  p.bbox = reinterpret_cast<int4*>(a.workspace);
  p.flags = reinterpret_cast<int32_t*>(p.bbox + (size_t)a.B * a.K);
  p.clamped = p.flags + (size_t)a.B * a.K;
  raster_prep_kernel<<<dim3(a.B, (a.K + PREP_WARPS - 1) / PREP_WARPS), 32 * PREP_WARPS, 0, stream>>>(p);
  if (int rc = check_launch("raster_prep_kernel")) return rc;
  const size_t smem = (size_t)a.K * (a.N * 4 + sizeof(int4)) + (size_t)((a.K + 3) & ~3) * sizeof(uint32_t);
  // 16 pixels per thread (one 16-byte store per instance; a warp writes 512 contiguous bytes) when the row pitch and
  // the buffer allow it, else 8 (8-byte stores).  CPOLY_RASTER_PX=8 forces the narrow variant (tests, profiling).
  const char* env = getenv("CPOLY_RASTER_PX");
  const bool px16 = !(env && atoi(env) == 8) && a.W % 16 == 0 && (reinterpret_cast<uintptr_t>(a.masks) & 15) == 0;
  auto kernel = px16 ? raster_cover_kernel<16> : raster_cover_kernel<8>;
  if (smem > 48 * 1024) CPOLY_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int tile_w = 32 * (px16 ? 16 : 8);
  // rows per warp: 1 when the grid is many waves deep (or fits in one), else just enough for one balanced wave
  int dev = 0, sms = 0, per_sm = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CPOLY_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 32 * TILE_H, smem));
  const long long capacity = (long long)sms * (per_sm > 0 ? per_sm : 1);
  const int tiles_x = (a.W + tile_w - 1) / tile_w, row_groups = (a.H + TILE_H - 1) / TILE_H;
  const long long ctas = (long long)tiles_x * row_groups * a.B;
  p.rows_per_warp = (ctas > capacity && ctas <= 4 * capacity) ? (int)((ctas + capacity - 1) / capacity) : 1;
  const dim3 grid(tiles_x, (row_groups + p.rows_per_warp - 1) / p.rows_per_warp, a.B);
  CPOLY_REQUIRE(grid.y <= 65535 && grid.z <= 65535, CPOLY_EINVAL, "raster grid too large");
  kernel<<<grid, 32 * TILE_H, smem, stream>>>(p);
  if (int rc = check_launch("raster_cover_kernel")) return rc;
  raster_keep_kernel<<<a.B, 256, 0, stream>>>(p);
  return check_launch("raster_keep_kernel");
}

}  // namespace cpoly
