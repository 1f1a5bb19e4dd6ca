// Instance masks of the IDD and KITTI writers on the GPU, sm_100a.
//
// Replaces the mask construction inside `IDD.format_and_write_to_IDD` (src/lib/datasets/dataset/IDD.py:123-170) and
// `KITTIPOLY.format_and_write_to_kitti` (src/lib/datasets/dataset/kitti_poly.py:95-136): per image, the classes in
// ascending order; inside a class the detections with `score >= thresh` (IDD :141) / `score > thresh` (KITTI :113) get
// their file index from a counter in detection order, are drawn in stable depth order as
// `ImageDraw.polygon(points, outline=0, fill=255)` times the class's running `to_remove_mask`, and those with
// `score >= 0.5` clear that mask with `polygon(points, outline=0, fill=0)`.  Points are
// `int(float("{:.2f}".format(v)))` of the decoded vertices.  In the reference that is K full-frame Pillow images and numpy
// products per frame on one host core.
//
// What Pillow draws (third-party dependency, 12.2.0 in the build container; restated and pinned in
// oracle/raster_fill_oracle.py): `fill=255, outline=0` = polygon_generic's scan-line fill followed by one `line8`
// Bresenham per edge with ink 0; `fill=0, outline=0` = the bare fill (equal inks: ImageDraw skips the outline).  Here:
//   fill_prep_kernel     threshold (fp64), vertex truncation, file index = rank on (class, detection index), occluder set
//                        of every instance = same class, score >= 0.5, earlier in (depth, detection index)
//   fill_rows_kernel     one warp per (instance, image row): lane 0 restates polygon_generic for that scan line (float32
//                        edge intersections, end-of-edge duplicate, corner patch, sorted pairs with Pillow's rounding and
//                        x_pos rule, horizontal edges), all lanes write the row with 16-byte stores: byte = 1 inside
//   fill_outline_kernel  one thread per (instance, edge) walks line8 and ORs 2 into the bytes it visits
//   fill_compose_kernel  per pixel of the instance's box: visible = fill & !outline & !(fill of any occluder) -> bit 4
//   fill_final_kernel    byte -> 0 / 255, per-instance area
#include <limits.h>

#include "common.cuh"

namespace cpoly {
namespace {

constexpr int FK_MAX = 256;      // detections per image
constexpr int FN_MAX = 64;       // vertices per polygon
constexpr int VCLAMP = 8192;

struct FillParams {
  int B, K, N, W, H, inclusive;
  double thresh, occluder_score;
  const float* dets;
  int32_t* count;       // (B)
  int32_t* slot_det;    // (B,K) file slot -> detection index
  int32_t* area;        // (B,K)
  int32_t* status;      // (B)
  uint8_t* masks;       // (B,K,H,W)
  // workspace
  uint32_t* verts;      // (B,K,N) packed int16 x | y << 16, per slot
  int4* bbox;           // (B,K) inclusive integer box of the vertices, per slot
  unsigned long long* occ;   // (B,K,4) bitmask over slots
};

// int(float("{:.2f}".format(x))): decimal rounding to two places, then truncation (IDD.py:62-63,159-160).  The integer
// differs from trunc(x) only when the fraction rounds up to 1.00, i.e. frac(|x|) > 0.995 (no fp32 value lies between
// the double nearest to 0.995 and 0.995 itself, so the comparison in fp64 is exact; sign symmetric).
__device__ __forceinline__ int round2_trunc(float x) {
  const double a = fabs((double)x);
  if (!(a < 1.0e9)) return a != a ? 0 : (x < 0.f ? -1000000000 : 1000000000);
  const double fl = floor(a);
  const int n = (int)fl + ((a - fl) > 0.995 ? 1 : 0);
  return x < 0.f ? -n : n;
}

// ---- 1. filter, file order, occluder sets, vertex truncation ---------------------------------------------------------
__global__ void __launch_bounds__(FK_MAX) fill_prep_kernel(const FillParams p) {
  __shared__ float s_depth[FK_MAX], s_score[FK_MAX];
  __shared__ int s_cls[FK_MAX], s_slot[FK_MAX];
  __shared__ uint8_t s_valid[FK_MAX];
  __shared__ int s_clamped;
  const int b = blockIdx.x, t = threadIdx.x;
  const int row_len = 2 * p.N + 7;
  const float* img = p.dets + (size_t)b * p.K * row_len;
  if (t == 0) s_clamped = 0;
  bool valid = false;
  if (t < p.K) {
    const float* row = img + (size_t)t * row_len;
    const float score = row[4];
    float depth = row[row_len - 1];
    if (depth != depth) depth = __int_as_float(0x7f800000);
    s_depth[t] = depth;
    s_score[t] = score;
    s_cls[t] = (int)row[5];
    valid = p.inclusive ? ((double)score >= p.thresh) : ((double)score > p.thresh);
    s_valid[t] = valid;
  }
  __syncthreads();
  int slot = -1, n_valid = 0;
  if (t < p.K) {
    int rank = 0;
    for (int j = 0; j < p.K; ++j) {
      if (!s_valid[j]) continue;
      ++n_valid;
      // the reference's counter: classes ascending, detection order inside a class
      if (s_cls[j] < s_cls[t] || (s_cls[j] == s_cls[t] && j < t)) ++rank;
    }
    slot = valid ? rank : -1;
    s_slot[t] = slot;
    if (t == 0) p.count[b] = n_valid;
  }
  __syncthreads();
  if (t < p.K) {
    const size_t base = (size_t)b * p.K;
    if (t >= n_valid) { p.slot_det[base + t] = -1; p.area[base + t] = 0; }
    if (valid) {
      p.slot_det[base + slot] = t;
      // occluders: same class, score >= occluder_score, drawn earlier = stable sort on depth (ties: list order)
      unsigned long long m[4] = {0ull, 0ull, 0ull, 0ull};
      for (int j = 0; j < p.K; ++j) {
        if (!s_valid[j] || j == t || s_cls[j] != s_cls[t] || !((double)s_score[j] >= p.occluder_score)) continue;
        if (s_depth[j] < s_depth[t] || (s_depth[j] == s_depth[t] && j < t)) m[s_slot[j] >> 6] |= 1ull << (s_slot[j] & 63);
      }
      for (int q = 0; q < 4; ++q) p.occ[(base + slot) * 4 + q] = m[q];
      const float* row = img + (size_t)t * row_len;
      int xmin = INT_MAX, xmax = INT_MIN, ymin = INT_MAX, ymax = INT_MIN, clamped = 0;
      uint32_t* v = p.verts + (base + slot) * p.N;
      for (int k = 0; k < p.N; ++k) {
        const int x = round2_trunc(row[6 + 2 * k]), y = round2_trunc(row[7 + 2 * k]);
        const int cx = min(max(x, -VCLAMP), VCLAMP - 1), cy = min(max(y, -VCLAMP), VCLAMP - 1);
        clamped += (cx != x) + (cy != y);
        v[k] = ((uint32_t)cx & 0xffffu) | ((uint32_t)cy << 16);
        xmin = min(xmin, cx); xmax = max(xmax, cx); ymin = min(ymin, cy); ymax = max(ymax, cy);
      }
      p.bbox[base + slot] = make_int4(xmin, ymin, xmax, ymax);
      if (clamped) atomicAdd(&s_clamped, clamped);
    }
  }
  __syncthreads();
  if (t == 0) p.status[b] = s_clamped;
}

// ---- 2. polygon_generic, one warp per (instance, row) ---------------------------------------------------------------------
__device__ __forceinline__ int rnd_up(float f) { return f >= 0.f ? (int)floorf(__fadd_rn(f, 0.5f)) : -(int)floorf(__fadd_rn(fabsf(f), 0.5f)); }
__device__ __forceinline__ int rnd_down(float f) { return f >= 0.f ? (int)ceilf(__fadd_rn(f, -0.5f)) : -(int)ceilf(__fadd_rn(fabsf(f), -0.5f)); }

struct FEdge {
  int x0, y0, x1, y1, ymin, ymax;
  float dx;
};
__device__ __forceinline__ FEdge make_edge(const uint32_t* v, int i, int n) {
  const uint32_t a = v[i], b = v[i + 1 == n ? 0 : i + 1];
  FEdge e;
  e.x0 = (int)(short)(a & 0xffffu); e.y0 = (int)(short)(a >> 16);
  e.x1 = (int)(short)(b & 0xffffu); e.y1 = (int)(short)(b >> 16);
  e.ymin = min(e.y0, e.y1); e.ymax = max(e.y0, e.y1);
  e.dx = e.y0 == e.y1 ? 0.f : __fdiv_rn((float)(e.x1 - e.x0), (float)(e.y1 - e.y0));
  return e;
}
// float32 intersection with two roundings like the C expression (no fused multiply-add)
__device__ __forceinline__ float edge_x(const FEdge& e, int y) { return __fadd_rn(__fmul_rn((float)(y - e.y0), e.dx), (float)e.x0); }

constexpr int FR_WARPS = 8;
constexpr int XX_MAX = 2 * FN_MAX + 2;

__global__ void __launch_bounds__(32 * FR_WARPS) fill_rows_kernel(const FillParams p) {
  __shared__ uint32_t s_v[FR_WARPS][FN_MAX];
  __shared__ float s_xx[FR_WARPS][XX_MAX];
  __shared__ short s_owner[FR_WARPS][XX_MAX];
  __shared__ int s_span[FR_WARPS][2 * (XX_MAX / 2 + FN_MAX)];
  __shared__ int s_nspan[FR_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * FR_WARPS + warp;
  const int64_t total = (int64_t)p.B * p.K * p.H;
  if (wid >= total) return;
  const int y = (int)(wid % p.H);
  const int64_t bs = wid / p.H;                 // b * K + slot
  const int b = (int)(bs / p.K), slot = (int)(bs - (int64_t)b * p.K);
  if (slot >= p.count[b]) return;
  const int n = p.N;
  for (int k = lane; k < n; k += 32) s_v[warp][k] = p.verts[bs * n + k];
  __syncwarp();
  if (lane == 0) {
    const uint32_t* v = s_v[warp];
    float* xx = s_xx[warp];
    short* owner = s_owner[warp];
    int* span = s_span[warp];
    const int4 bb = p.bbox[bs];
    const int gymin = bb.y, gymax = bb.w;
    const bool closed = v[n - 1] == v[0];          // Pillow adds the closing edge only when last != first
    const int n_edges = closed ? n - 1 : n;
    int ns = 0, j = 0;
    if (y >= gymin && y <= gymax) {
      for (int i = 0; i < n_edges; ++i) {
        const FEdge cur = make_edge(v, i, n);
        if (cur.y0 == cur.y1) {                    // horizontal edge: drawn as a span of its own
          if (cur.y0 == y) { span[2 * ns] = min(cur.x0, cur.x1); span[2 * ns + 1] = max(cur.x0, cur.x1); ++ns; }
          continue;
        }
        if (y < cur.ymin || y > cur.ymax) continue;
        const float xv = edge_x(cur, y);
        xx[j] = xv; owner[j] = (short)i; ++j;
        if (y == cur.ymax && y < gymax) { xx[j] = xv; owner[j] = -1; ++j; continue; }   // "needed to draw consistent polygons"
        if (cur.dx != 0.f && truncf(xv) == xv) {   // connect discontiguous corners
          for (int jx = 0; jx < j - 1; ++jx) {
            const int k = owner[jx];
            if (k < 0 || k == i) continue;
            const FEdge oth = make_edge(v, k, n);
            if ((cur.dx > 0.f && oth.dx <= 0.f) || (cur.dx < 0.f && oth.dx >= 0.f)) continue;
            if (xv != edge_x(oth, y)) continue;
            const int xi = (int)xv;
            if (!(((cur.x0 == xi && cur.y0 == y) || (cur.x1 == xi && cur.y1 == y)) && ((oth.x0 == xi && oth.y0 == y) || (oth.x1 == xi && oth.y1 == y))))
              continue;                                // two edges that merely cross at an integer x are not a corner
            const int ya = y == gymax ? y - 1 : y + 1;
            if (ya < cur.ymin || ya > cur.ymax || ya < oth.ymin || ya > oth.ymax) continue;
            const float a = edge_x(cur, ya), bq = edge_x(oth, ya);
            const float lo = fminf(a, bq), hi = fmaxf(a, bq);
            if ((float)(rnd_up(lo) - 1) > xv) xx[jx] = (float)(rnd_up(lo) - 1);
            else if ((float)rnd_up(__fadd_rn(hi, 1.f)) < xv) xx[jx] = (float)rnd_up(__fadd_rn(hi, 1.f));
            break;
          }
        }
      }
      for (int a = 1; a < j; ++a) {                // insertion sort (a handful of entries per row)
        const float key = xx[a];
        int q = a - 1;
        while (q >= 0 && xx[q] > key) { xx[q + 1] = xx[q]; --q; }
        xx[q + 1] = key;
      }
      int x_pos = j ? 0 : -1;
      for (int i = 1; i < j; i += 2) {
        const int x_end = rnd_down(xx[i]);
        if (x_end < x_pos) continue;
        int x_start = rnd_up(xx[i - 1]);
        if (x_pos > x_start) {
          x_start = x_pos;
          if (x_end < x_start) continue;
        }
        span[2 * ns] = x_start; span[2 * ns + 1] = x_end; ++ns;
        x_pos = x_end + 1;
      }
    }
    s_nspan[warp] = ns;
  }
  __syncwarp();
  const int ns = s_nspan[warp];
  const int* span = s_span[warp];
  uint8_t* row = p.masks + (size_t)wid * p.W;
  if ((p.W & 15) == 0) {
    for (int c = lane; c < (p.W >> 4); c += 32) {
      const int x0 = c << 4;
      uint32_t w[4] = {0u, 0u, 0u, 0u};
      for (int s = 0; s < ns; ++s) {
        const int lo = max(span[2 * s], x0), hi = min(span[2 * s + 1], x0 + 15);
        for (int x = lo; x <= hi; ++x) w[(x - x0) >> 2] |= 1u << (8 * ((x - x0) & 3));
      }
      reinterpret_cast<uint4*>(row)[c] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  } else {
    for (int x = lane; x < p.W; x += 32) {
      uint8_t val = 0;
      for (int s = 0; s < ns; ++s) val |= (x >= span[2 * s] && x <= span[2 * s + 1]) ? 1 : 0;
      row[x] = val;
    }
  }
}

// ---- 3. outline: Pillow's line8 per edge, end point excluded; OR 2 into the visited bytes ----------------------------------
__global__ void __launch_bounds__(128) fill_outline_kernel(const FillParams p) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)p.B * p.K * p.N;
  if (tid >= total) return;
  const int e = (int)(tid % p.N);
  const int64_t bs = tid / p.N;
  const int b = (int)(bs / p.K), slot = (int)(bs - (int64_t)b * p.K);
  if (slot >= p.count[b]) return;
  const uint32_t* v = p.verts + bs * p.N;
  const uint32_t a = v[e], c = v[e + 1 == p.N ? 0 : e + 1];
  int x0 = (int)(short)(a & 0xffffu), y0 = (int)(short)(a >> 16);
  const int x1 = (int)(short)(c & 0xffffu), y1 = (int)(short)(c >> 16);
  uint8_t* plane = p.masks + (size_t)bs * p.H * p.W;
  auto point = [&](int x, int y) {
    if (x >= 0 && x < p.W && y >= 0 && y < p.H) plane[(size_t)y * p.W + x] |= 2;   // (same value from every writer of a byte)
  };
  int dx = x1 - x0, xs = 1, dy = y1 - y0, ys = 1;
  if (dx < 0) { dx = -dx; xs = -1; }
  if (dy < 0) { dy = -dy; ys = -1; }
  if (dx == 0) {
    for (int i = 0; i < dy; ++i) { point(x0, y0); y0 += ys; }
  } else if (dy == 0) {
    for (int i = 0; i < dx; ++i) { point(x0, y0); x0 += xs; }
  } else if (dx > dy) {
    const int n = dx;
    dy += dy;
    int err = dy - dx;
    dx += dx;
    for (int i = 0; i < n; ++i) {
      point(x0, y0);
      if (err >= 0) { y0 += ys; err -= dx; }
      err += dy; x0 += xs;
    }
  } else {
    const int n = dy;
    dx += dx;
    int err = dx - dy;
    dy += dy;
    for (int i = 0; i < n; ++i) {
      point(x0, y0);
      if (err >= 0) { x0 += xs; err -= dy; }
      err += dx; y0 += ys;
    }
  }
}

// ---- 4. visible = fill & !outline & !(fill of an occluder): bit 4 -------------------------------------------------------------
// Grid (row chunks, B * K).  Only the instance's own box can hold fill; occluders are tested through their boxes first.
__global__ void __launch_bounds__(256) fill_compose_kernel(const FillParams p) {
  __shared__ int4 s_box[FK_MAX];
  __shared__ int s_list[FK_MAX];
  __shared__ int s_n;
  const int64_t bs = blockIdx.y;
  const int b = (int)(bs / p.K), slot = (int)(bs - (int64_t)b * p.K);
  if (slot >= p.count[b]) return;
  const int4 bb = p.bbox[bs];
  const int x0 = max(bb.x, 0), x1 = min(bb.z, p.W - 1), y0 = max(bb.y, 0), y1 = min(bb.w, p.H - 1);
  if (x0 > x1 || y0 > y1) return;
  if (threadIdx.x == 0) {
    int n = 0;
    for (int q = 0; q < 4; ++q) {
      unsigned long long m = p.occ[bs * 4 + q];
      while (m) {
        const int j = q * 64 + __ffsll((long long)m) - 1;
        m &= m - 1;
        const int4 ob = p.bbox[(int64_t)b * p.K + j];
        if (ob.x <= x1 && ob.z >= x0 && ob.y <= y1 && ob.w >= y0) { s_box[n] = ob; s_list[n] = j; ++n; }
      }
    }
    s_n = n;
  }
  __syncthreads();
  const int n_occ = s_n;
  const int bw = x1 - x0 + 1;
  const int64_t npx = (int64_t)bw * (y1 - y0 + 1);
  const size_t plane_px = (size_t)p.H * p.W;
  uint8_t* mine = p.masks + (size_t)bs * plane_px;
  const uint8_t* img = p.masks + (size_t)b * p.K * plane_px;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npx; i += (int64_t)gridDim.x * blockDim.x) {
    const int y = y0 + (int)(i / bw), x = x0 + (int)(i % bw);
    const size_t off = (size_t)y * p.W + x;
    const uint8_t v = mine[off];
    if ((v & 3) != 1) continue;                 // not filled, or on the outline
    bool hidden = false;
    for (int q = 0; q < n_occ && !hidden; ++q) {
      const int4 ob = s_box[q];
      if (x >= ob.x && x <= ob.z && y >= ob.y && y <= ob.w) hidden = (img[(size_t)s_list[q] * plane_px + off] & 1) != 0;
    }
    if (!hidden) mine[off] = v | 4;
  }
}

// ---- 5. byte -> 0 / 255 and the area of every instance ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fill_final_kernel(const FillParams p) {
  const int64_t bs = blockIdx.y;
  const int b = (int)(bs / p.K), slot = (int)(bs - (int64_t)b * p.K);
  if (slot >= p.count[b]) return;
  const size_t plane_px = (size_t)p.H * p.W;
  uint8_t* mine = p.masks + (size_t)bs * plane_px;
  int local = 0;
  if ((plane_px & 15) == 0) {
    uint4* q = reinterpret_cast<uint4*>(mine);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (plane_px >> 4); i += (size_t)gridDim.x * blockDim.x) {
      uint4 v = q[i];
      uint32_t* w = &v.x;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t vis = (w[k] >> 2) & 0x01010101u;      // bit 4 of every byte
        local += __popc(vis);
        w[k] = vis * 255u;
      }
      q[i] = v;
    }
  } else {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < plane_px; i += (size_t)gridDim.x * blockDim.x) {
      const uint8_t vis = (mine[i] >> 2) & 1;
      local += vis;
      mine[i] = vis ? 255 : 0;
    }
  }
  local = __reduce_add_sync(0xffffffffu, local);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(p.area + bs, local);   // integer: order-independent
}

size_t align256f(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

size_t raster_fill_workspace_bytes(int B, int K, int N) {
  if (B <= 0 || K <= 0 || N <= 0) return 256;
  const size_t bk = (size_t)B * K;
  return align256f(bk * N * sizeof(uint32_t)) + align256f(bk * sizeof(int4)) + align256f(bk * 4 * sizeof(unsigned long long)) + 256;
}

int raster_fill_launch(const cpoly_raster_fill_args& a, cudaStream_t stream) {
  FillParams p;
  p.B = a.B; p.K = a.K; p.N = a.N; p.W = a.W; p.H = a.H; p.inclusive = a.inclusive;
  p.thresh = a.thresh; p.occluder_score = a.occluder_score;
  p.dets = a.dets; p.count = a.count; p.slot_det = a.slot_det; p.area = a.area; p.status = a.status; p.masks = a.masks;
  char* ws = reinterpret_cast<char*>(a.workspace);
  const size_t bk = (size_t)a.B * a.K;
  p.verts = reinterpret_cast<uint32_t*>(ws); ws += align256f(bk * a.N * sizeof(uint32_t));
  p.bbox = reinterpret_cast<int4*>(ws); ws += align256f(bk * sizeof(int4));
  p.occ = reinterpret_cast<unsigned long long*>(ws);
  CPOLY_CUDA_OK(cudaMemsetAsync(a.area, 0, bk * sizeof(int32_t), stream));
  fill_prep_kernel<<<a.B, FK_MAX, 0, stream>>>(p);
  if (int rc = check_launch("fill_prep_kernel")) return rc;
  if (!a.masks) return CPOLY_OK;
  const int64_t row_warps = (int64_t)a.B * a.K * a.H;
  fill_rows_kernel<<<(unsigned)((row_warps + FR_WARPS - 1) / FR_WARPS), 32 * FR_WARPS, 0, stream>>>(p);
  if (int rc = check_launch("fill_rows_kernel")) return rc;
  const int64_t edges = (int64_t)a.B * a.K * a.N;
  fill_outline_kernel<<<(unsigned)((edges + 127) / 128), 128, 0, stream>>>(p);
  if (int rc = check_launch("fill_outline_kernel")) return rc;
  const dim3 grid(16, (unsigned)bk);
  fill_compose_kernel<<<grid, 256, 0, stream>>>(p);
  if (int rc = check_launch("fill_compose_kernel")) return rc;
  const dim3 fgrid(32, (unsigned)bk);
  fill_final_kernel<<<fgrid, 256, 0, stream>>>(p);
  return check_launch("fill_final_kernel");
}

}  // namespace cpoly
