// Polygon-IoU loss of CenterPoly's PolyLoss (poly_loss = 'iou' / 'l1+iou', polar representation) on the GPU:
// forward value AND the analytic gradient w.r.t. the poly head in one launch, one warp per object slot.
//
// Reference: src/lib/models/losses.py
//   area                 :25-41     shoelace over the polar polygon with one extra wrap term
//   WeilPolygonClipper   :373-628   is_inside :378-400, compute_intersection :404-478, clip :480-617
//   PolyLoss.forward     :870-888   sort by angle, abs(r), clip, intersection / union; :909-910 loss = 1 - sum/(n + 1e-6)
//   _transpose_and_gather_feat (src/lib/models/utils.py:22-26) is fused: the prediction row is read straight from
//   the (B,2N,H,W) head map at `ind`, and the backward kernel scatters into the map's gradient.
// The reference spends ~25-400 ms of host Python per object (0-dim tensor arithmetic, np.where, torch.cat growth);
// here the N x N edge classification is lane-parallel, the in/out-bound walk runs on integer lists in shared
// memory, and the backward pass is closed-form (no tape): shoelace -> polar chain rule -> line-intersection
// chain rule -> the two prediction vertices of the crossing edge.
//
// fp32 arithmetic mirrors the reference operation by operation where a rounding could change a DECISION
// (side tests use separately rounded products: __fmul_rn / __fsub_rn, never FMA-contracted).
#include <math_constants.h>

#include "common.cuh"

namespace cpoly {
namespace {

constexpr int kWarps = 4;        // object slots per CTA
constexpr int kNMax = 64;        // vertices per polygon
constexpr int kLMax = 256;       // in-bound / out-bound crossings kept per object (4 * kNMax)
constexpr int kVMax = 8 * kNMax; // clipped-polygon vertices before the walk is declared a runaway

enum : int { ST_OK = 0, ST_NO_OUTBOUND = 1, ST_RUNAWAY = 2, ST_TOO_MANY = 3, ST_NONFINITE = 4 };
enum : unsigned char { V_SUBJECT = 0, V_CLIP = 1, V_CROSS = 2 };

struct Scratch {
  float sr[kNMax], st[kNMax], sx[kNMax], sy[kNMax];   // prediction, sorted by angle: |r|, theta, x, y
  float cr[kNMax], ct[kNMax], cx[kNMax], cy[kNMax];   // target
  float gr[kNMax], gt[kNMax];                         // d iou / d (sorted |r|, theta)
  float vx[kVMax + 4], vy[kVMax + 4];                 // clipped polygon, cartesian (as area() recomputes it)
  int first_out[kNMax];
  unsigned char perm[kNMax];
  signed char sgn[kNMax];
  unsigned char side[kNMax];
  unsigned char in_j[kLMax], in_i[kLMax], out_j[kLMax], out_i[kLMax], alive[kLMax];
  unsigned char in_cnt[kNMax];                        // alive in-bounds per clip index
  unsigned char vtype[kVMax + 4], va[kVMax + 4], vb[kVMax + 4];
  int n_in, n_out, n_v, status;
};

__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }

// is_inside (:378-400): 1 if R < 0, 2 if R == 0, else 0
__device__ __forceinline__ int side_of(float p1x, float p1y, float p2x, float p2y, float qx, float qy) {
  const float r = fsub(fmul(fsub(p2x, p1x), fsub(qy, p1y)), fmul(fsub(p2y, p1y), fsub(qx, p1x)));
  return r < 0.0f ? 1 : (r == 0.0f ? 2 : 0);
}

struct Cross {
  float x, y;        // intersection, cartesian
  float m1, b1, m2, b2;
  int kind;          // 0: first line vertical, 1: second line vertical, 2: general
  float rho, phi, u, xe;
};

// compute_intersection (:404-478)
__device__ __forceinline__ Cross intersect(float p1x, float p1y, float p2x, float p2y, float p3x, float p3y, float p4x, float p4y) {
  Cross c;
  c.m1 = c.b1 = c.m2 = c.b2 = 0.0f;
  if (fsub(p2x, p1x) == 0.0f) {
    c.kind = 0;
    c.x = p1x;
    c.m2 = fdiv(fsub(p4y, p3y), fsub(p4x, p3x));
    c.b2 = fsub(p3y, fmul(c.m2, p3x));
    c.y = fadd(fmul(c.m2, c.x), c.b2);
  } else if (fsub(p4x, p3x) == 0.0f) {
    c.kind = 1;
    c.x = p3x;
    c.m1 = fdiv(fsub(p2y, p1y), fsub(p2x, p1x));
    c.b1 = fsub(p1y, fmul(c.m1, p1x));
    c.y = fadd(fmul(c.m1, c.x), c.b1);
  } else {
    c.kind = 2;
    c.m1 = fdiv(fsub(p2y, p1y), fsub(p2x, p1x));
    c.b1 = fsub(p1y, fmul(c.m1, p1x));
    c.m2 = fdiv(fsub(p4y, p3y), fsub(p4x, p3x));
    c.b2 = fsub(p3y, fmul(c.m2, p3x));
    c.x = fdiv(fsub(c.b2, c.b1), fsub(c.m1, c.m2));
    c.y = fadd(fmul(c.m1, c.x), c.b1);
  }
  c.rho = sqrtf(fadd(fmul(c.x, c.x), fmul(c.y, c.y)));
  c.xe = fadd(c.x, 1e-8f);
  c.u = fdiv(fadd(c.y, 1e-8f), c.xe);
  c.phi = atanf(c.u);
  if (c.x < 0.0f) c.phi = fadd(c.phi, CUDART_PI_F);
  else if (c.y < 0.0f) c.phi = fadd(c.phi, 2.0f * CUDART_PI_F);
  return c;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// area (:25-41) of n cartesian vertices: terms 0..n of the doubled polygon. Returns right - left (signed, not halved).
__device__ __forceinline__ float signed_shoelace(const float* x, const float* y, int n, int lane) {
  float right = 0.0f, left = 0.0f;
  if (n > 0) {
    for (int t = lane; t <= n; t += 32) {
      const int a = t % n, b = (t + 1) % n;
      right += fmul(y[a], x[b]);
      left += fmul(x[a], y[b]);
    }
  }
  return fsub(warp_sum(right), warp_sum(left));
}

// d(right - left)/d(x_k, y_k) for the same n+1 terms (n >= 2)
__device__ __forceinline__ void shoelace_grad(const float* x, const float* y, int n, int k, float& gx, float& gy) {
  const int kp = (k + 1) % n, km = (k + n - 1) % n;
  gx = y[km] - y[kp];
  gy = x[kp] - x[km];
  if (k == 1 % n) { gx += y[0]; gy -= x[0]; }   // extra term t = n: a = 0, b = 1
  if (k == 0) { gx -= y[1 % n]; gy += x[1 % n]; }
}

struct Params {
  cpoly_tensor4 poly;
  int dtype, B, M, N, H, W;
  const int64_t* ind;
  const uint8_t* mask;
  const float* target;
  float* iou;
  float* dpred;
  int32_t* status;
};

__global__ void __launch_bounds__(kWarps * 32) poly_iou_kernel(Params p) {
  __shared__ Scratch scratch[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * kWarps + warp;
  if (slot >= p.B * p.M) return;
  Scratch& s = scratch[warp];
  const int N = p.N;
  float* dp = p.dpred + (size_t)slot * 2 * N;
  if (!p.mask[slot]) {
    for (int c = lane; c < 2 * N; c += 32) dp[c] = 0.0f;
    if (lane == 0) { p.iou[slot] = 0.0f; p.status[slot] = ST_OK; }
    return;
  }
  const int b = slot / p.M;
  int64_t cell = p.ind[slot];
  cell = cell < 0 ? 0 : (cell >= (int64_t)p.H * p.W ? (int64_t)p.H * p.W - 1 : cell);
  const int64_t base = b * p.poly.sb + (cell / p.W) * p.poly.sh + (cell % p.W) * p.poly.sw;
  const float* tg = p.target + (size_t)slot * 2 * N;

  // ---- load, sort the prediction by angle (stable), abs(r)                              (:876-877)
  float rr[2], rt[2];
  bool finite = true;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int v = lane + 32 * h;
    rr[h] = rt[h] = 0.0f;
    if (v < N) {
      rr[h] = ld_dyn(p.poly.ptr, base + (int64_t)(2 * v) * p.poly.sc, p.dtype);
      rt[h] = ld_dyn(p.poly.ptr, base + (int64_t)(2 * v + 1) * p.poly.sc, p.dtype);
      s.cr[v] = tg[2 * v];
      s.ct[v] = tg[2 * v + 1];
      s.sx[v] = rt[h];                       // staging: raw angles for the rank sort
      finite = finite && isfinite(rr[h]) && isfinite(rt[h]) && isfinite(s.cr[v]) && isfinite(s.ct[v]);
      s.gr[v] = 0.0f;
      s.gt[v] = 0.0f;
      s.first_out[v] = 0x7fffffff;
      s.in_cnt[v] = 0;
    }
  }
  if (lane == 0) { s.n_in = 0; s.n_out = 0; s.n_v = 0; s.status = ST_OK; }
  __syncwarp();
  if (!__all_sync(0xffffffffu, finite)) {
    for (int c = lane; c < 2 * N; c += 32) dp[c] = 0.0f;
    if (lane == 0) { p.iou[slot] = 0.0f; p.status[slot] = ST_NONFINITE; }
    return;
  }
  int rank[2] = {0, 0};
  for (int u = 0; u < N; ++u) {
    const float tu = s.sx[u];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int v = lane + 32 * h;
      rank[h] += (tu < rt[h] || (tu == rt[h] && u < v)) ? 1 : 0;
    }
  }
  __syncwarp();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int v = lane + 32 * h;
    if (v < N) {
      const int k = rank[h];
      s.sr[k] = fabsf(rr[h]);
      s.st[k] = rt[h];
      s.perm[k] = (unsigned char)v;
      s.sgn[k] = rr[h] > 0.0f ? 1 : (rr[h] < 0.0f ? -1 : 0);
    }
  }
  __syncwarp();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int v = lane + 32 * h;
    if (v < N) {
      float sn, cs;
      sincosf(s.st[v], &sn, &cs);
      s.sx[v] = fmul(s.sr[v], cs);
      s.sy[v] = fmul(s.sr[v], sn);
      sincosf(s.ct[v], &sn, &cs);
      s.cx[v] = fmul(s.cr[v], cs);
      s.cy[v] = fmul(s.cr[v], sn);
    }
  }
  __syncwarp();

  // ---- crossings: the double loop of clip() (:497-566), clip edge i outer, prediction edge j inner
  const unsigned lt = (1u << lane) - 1u;
  bool too_many = false;
  for (int i = 0; i < N; ++i) {
    const int ip = (i + N - 1) % N;
    const float c0x = s.cx[ip], c0y = s.cy[ip], c1x = s.cx[i], c1y = s.cy[i];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int j = lane + 32 * h;
      if (j < N) s.side[j] = (unsigned char)side_of(c0x, c0y, c1x, c1y, s.sx[j], s.sy[j]);
    }
    __syncwarp();
    for (int h = 0; h * 32 < N; ++h) {
      const int j = lane + 32 * h;
      int kind = 0;   // 1 = out-bound, 2 = in-bound
      if (j < N) {
        const int jp = (j + N - 1) % N;
        const int to_end = s.side[j], to_start = s.side[jp];
        if (to_start == 0 && to_end != 0) kind = 1;
        else if (to_end == 0 && to_start != 0) kind = 2;
        if (kind) {
          const int a = side_of(s.sx[jp], s.sy[jp], s.sx[j], s.sy[j], c1x, c1y);
          const int bb = side_of(s.sx[jp], s.sy[jp], s.sx[j], s.sy[j], c0x, c0y);
          if (a == bb) kind = 0;
        }
      }
      const unsigned mo = __ballot_sync(0xffffffffu, kind == 1), mi = __ballot_sync(0xffffffffu, kind == 2);
      const int n_out = s.n_out, n_in = s.n_in;
      if (kind == 1) {
        const int pos = n_out + __popc(mo & lt);
        if (pos < kLMax) { s.out_j[pos] = (unsigned char)j; s.out_i[pos] = (unsigned char)i; }
      } else if (kind == 2) {
        const int pos = n_in + __popc(mi & lt);
        if (pos < kLMax) { s.in_j[pos] = (unsigned char)j; s.in_i[pos] = (unsigned char)i; s.alive[pos] = 1; }
      }
      too_many = too_many || n_out + __popc(mo) > 4 * N || n_in + __popc(mi) > 4 * N;
      __syncwarp();
      if (lane == 0) { s.n_out = min(n_out + __popc(mo), kLMax); s.n_in = min(n_in + __popc(mi), kLMax); }
      __syncwarp();
    }
  }
  const int n_in0 = s.n_in, n_out0 = s.n_out;
  for (int k = lane; k < n_out0; k += 32) atomicMin(&s.first_out[s.out_j[k]], k);
  __syncwarp();

  // ---- the walk (:577-612), serial on integer lists
  if (lane == 0) {
    int status = ST_OK, nv = 0;
    for (int k = 0; k < n_in0; ++k) s.in_cnt[s.in_i[k]]++;
    if (too_many) status = ST_TOO_MANY;
    else if (n_in0 == 0) status = ST_OK;
    else if (n_out0 == 0) status = ST_NO_OUTBOUND;
    else {
      int alive_cnt = n_in0, used = 0, head = 0;
      const int max_len = 8 * N;
      auto push = [&](unsigned char t, int a, int bq) {
        if (nv < kVMax + 4) { s.vtype[nv] = t; s.va[nv] = (unsigned char)a; s.vb[nv] = (unsigned char)bq; }
        ++nv;
      };
      while (status == ST_OK && used < alive_cnt) {
        while (!s.alive[head]) ++head;             // inbounds[0]
        const int stop_j = s.in_j[head], stop_i = s.in_i[head];
        int j = stop_j, i = stop_i;
        bool start = true;
        while (status == ST_OK && (j != stop_j || i != stop_i || start)) {
          start = false;
          while (s.first_out[j] == 0x7fffffff) {
            push(V_SUBJECT, j, 0);
            j = (j + 1) % N;
            if (nv > max_len) { status = ST_RUNAWAY; break; }
          }
          if (status != ST_OK) break;
          const int ko = s.first_out[j];
          push(V_CROSS, s.out_j[ko], s.out_i[ko]);
          i = s.out_i[ko];
          while (s.in_cnt[i] == 0) {
            push(V_CLIP, i, 0);
            i = (i + 1) % N;
            if (nv > max_len || alive_cnt == 0) { status = ST_RUNAWAY; break; }
          }
          if (status != ST_OK) break;
          int k = head;
          while (!(s.alive[k] && s.in_i[k] == i)) ++k;   // np.where(inbounds[:,1] == i)[0][0]
          j = s.in_j[k];
          push(V_CROSS, s.in_j[k], s.in_i[k]);
          s.alive[k] = 0;
          s.in_cnt[i]--;
          --alive_cnt;
          ++used;
          if (nv > max_len) status = ST_RUNAWAY;
        }
      }
    }
    s.n_v = status == ST_OK ? nv : 0;
    s.status = status;
  }
  __syncwarp();
  const int status = s.status, nv = s.n_v;
  if (status != ST_OK) {
    for (int c = lane; c < 2 * N; c += 32) dp[c] = 0.0f;
    if (lane == 0) { p.iou[slot] = 0.0f; p.status[slot] = status; }
    return;
  }

  // ---- clipped polygon -> cartesian, as area() recomputes it from (r, theta)             (:31-32)
  for (int k = lane; k < nv; k += 32) {
    const int t = s.vtype[k], a = s.va[k];
    float x, y;
    if (t == V_SUBJECT) { x = s.sx[a]; y = s.sy[a]; }
    else if (t == V_CLIP) { x = s.cx[a]; y = s.cy[a]; }
    else {
      const int i = s.vb[k], jp = (a + N - 1) % N, ip = (i + N - 1) % N;
      const Cross c = intersect(s.sx[jp], s.sy[jp], s.sx[a], s.sy[a], s.cx[ip], s.cy[ip], s.cx[i], s.cy[i]);
      float sn, cs;
      sincosf(c.phi, &sn, &cs);
      x = fmul(c.rho, cs);
      y = fmul(c.rho, sn);
    }
    s.vx[k] = x;
    s.vy[k] = y;
  }
  __syncwarp();
  const float d_int = signed_shoelace(s.vx, s.vy, nv, lane);
  const float d_s = signed_shoelace(s.sx, s.sy, N, lane);
  const float d_c = signed_shoelace(s.cx, s.cy, N, lane);
  const float a_int = fabsf(0.5f * d_int), a_s = fabsf(0.5f * d_s), a_c = fabsf(0.5f * d_c);
  // :883-887
  const bool empty = a_int == 0.0f;
  const float inter = empty ? fadd(fminf(a_s, a_c), a_int) : a_int;
  const float uni = fsub(fadd(a_c, a_s), inter);
  const float den = fadd(uni, 1e-6f);
  const float iou = fdiv(inter, den);
  // d iou / d inter (through the union too) and d iou / d a_s through the union alone
  const float g_inter = 1.0f / den + inter / (den * den);
  float g_as = -inter / (den * den);
  float g_aint = 0.0f;
  if (empty) g_as += g_inter * (a_s < a_c ? 1.0f : (a_s == a_c ? 0.5f : 0.0f));
  else g_aint = g_inter;
  const float k_s = g_as * 0.5f * (d_s > 0.0f ? 1.0f : (d_s < 0.0f ? -1.0f : 0.0f));
  const float k_i = g_aint * 0.5f * (d_int > 0.0f ? 1.0f : (d_int < 0.0f ? -1.0f : 0.0f));

  // ---- backward: area(sorted_pred)
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int j = lane + 32 * h;
    if (j < N) {
      float gx, gy;
      shoelace_grad(s.sx, s.sy, N, j, gx, gy);
      gx *= k_s; gy *= k_s;
      float sn, cs;
      sincosf(s.st[j], &sn, &cs);
      s.gr[j] = gx * cs + gy * sn;
      s.gt[j] = -gx * s.sy[j] + gy * s.sx[j];
    }
  }
  __syncwarp();
  // ---- backward: area(clipped polygon); contributions added in vertex order (deterministic)
  if (k_i != 0.0f) {
    for (int k0 = 0; k0 < nv; k0 += 32) {
      const int k = k0 + lane;
      int ja = -1, jb = -1;
      float gar = 0.0f, gat = 0.0f, gbr = 0.0f, gbt = 0.0f;
      if (k < nv && s.vtype[k] != V_CLIP) {
        float gx, gy;
        shoelace_grad(s.vx, s.vy, nv, k, gx, gy);
        gx *= k_i; gy *= k_i;
        const int a = s.va[k];
        if (s.vtype[k] == V_SUBJECT) {
          float sn, cs;
          sincosf(s.st[a], &sn, &cs);
          ja = a;
          gar = gx * cs + gy * sn;
          gat = -gx * s.sy[a] + gy * s.sx[a];
        } else {
          const int i = s.vb[k], jp = (a + N - 1) % N, ip = (i + N - 1) % N;
          const float p1x = s.sx[jp], p1y = s.sy[jp], p2x = s.sx[a], p2y = s.sy[a];
          const Cross c = intersect(p1x, p1y, p2x, p2y, s.cx[ip], s.cy[ip], s.cx[i], s.cy[i]);
          float sn, cs;
          sincosf(c.phi, &sn, &cs);
          const float g_rho = gx * cs + gy * sn;
          const float g_phi = -gx * s.vy[k] + gy * s.vx[k];
          float g_x = c.rho > 0.0f ? g_rho * c.x / c.rho : 0.0f;
          float g_y = c.rho > 0.0f ? g_rho * c.y / c.rho : 0.0f;
          const float g_u = g_phi / (1.0f + c.u * c.u);
          g_y += g_u / c.xe;
          g_x -= g_u * c.u / c.xe;
          float g1x = 0.0f, g1y = 0.0f, g2x = 0.0f, g2y = 0.0f;
          if (c.kind == 0) {
            g1x = g_x + g_y * c.m2;                 // the reference's branch: x = p1x, y = m2 x + b2
          } else {
            // Closed form of what autograd yields for x = (b2-b1)/(m1-m2), y = m1 x + b1 (and for the second-line-
            // vertical branch), without slopes: the crossing moves along d2 by n1.((1-t) dP1 + t dP2) / (n1.d2).
            // The slope form cancels catastrophically in fp32 on steep prediction edges (the reference's own
            // fp32 gradient is off by up to ~10 % there; tests/test_gpu_polyiou.py measures both against fp64).
            const float d1x = p2x - p1x, d1y = p2y - p1y;
            const float d2x = s.cx[i] - s.cx[ip], d2y = s.cy[i] - s.cy[ip];
            const float t = ((c.x - p1x) * d1x + (c.y - p1y) * d1y) / (d1x * d1x + d1y * d1y);
            const float cc = (g_x * d2x + g_y * d2y) / (d1x * d2y - d1y * d2x);
            g1x = -(1.0f - t) * cc * d1y;
            g1y = (1.0f - t) * cc * d1x;
            g2x = -t * cc * d1y;
            g2y = t * cc * d1x;
          }
          sincosf(s.st[jp], &sn, &cs);
          ja = jp;
          gar = g1x * cs + g1y * sn;
          gat = -g1x * p1y + g1y * p1x;
          sincosf(s.st[a], &sn, &cs);
          jb = a;
          gbr = g2x * cs + g2y * sn;
          gbt = -g2x * p2y + g2y * p2x;
        }
      }
      for (int l = 0; l < 32 && k0 + l < nv; ++l) {
        if (lane == l) {
          if (ja >= 0) { s.gr[ja] += gar; s.gt[ja] += gat; }
          if (jb >= 0) { s.gr[jb] += gbr; s.gt[jb] += gbt; }
        }
        __syncwarp();
      }
    }
  }
  __syncwarp();
  // ---- un-sort, abs'(r) = sign(r)
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int j = lane + 32 * h;
    if (j < N) {
      const int v = s.perm[j];
      dp[2 * v] = s.gr[j] * (float)s.sgn[j];
      dp[2 * v + 1] = s.gt[j];
    }
  }
  if (lane == 0) { p.iou[slot] = iou; p.status[slot] = ST_OK; }
}

// loss = 1 - sum(iou) / (sum(mask) + 1e-6) (:909-910); stats = {loss, coef = -1/(sum(mask)+1e-6), n_mask, n_bad}
__global__ void __launch_bounds__(256) poly_iou_reduce_kernel(const float* iou, const uint8_t* mask, const int32_t* status, int n, float* stats) {
  __shared__ float s_iou[256], s_m[256], s_bad[256];
  float a = 0.0f, m = 0.0f, bad = 0.0f;
  for (int k = threadIdx.x; k < n; k += 256) {
    if (mask[k]) { a += iou[k]; m += 1.0f; bad += status[k] != 0 ? 1.0f : 0.0f; }
  }
  s_iou[threadIdx.x] = a; s_m[threadIdx.x] = m; s_bad[threadIdx.x] = bad;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      s_iou[threadIdx.x] += s_iou[threadIdx.x + o];
      s_m[threadIdx.x] += s_m[threadIdx.x + o];
      s_bad[threadIdx.x] += s_bad[threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float den = s_m[0] + 1e-6f;
    stats[0] = 1.0f - s_iou[0] / den;
    stats[1] = -1.0f / den;
    stats[2] = s_m[0];
    stats[3] = s_bad[0];
  }
}

// grad_map[b, c, ind] += grad_loss * coef * dpred[b, m, c]
__global__ void __launch_bounds__(256) poly_iou_scatter_kernel(const float* dpred, const uint8_t* mask, const int64_t* ind, const float* stats,
                                                               const float* grad_loss, cpoly_tensor4 gmap, int B, int M, int C, int H, int W) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= (int64_t)B * M * C) return;
  const int c = (int)(t % C);
  const int64_t slot = t / C;
  if (!mask[slot]) return;
  const float g = dpred[t];
  if (g == 0.0f) return;
  int64_t cell = ind[slot];
  cell = cell < 0 ? 0 : (cell >= (int64_t)H * W ? (int64_t)H * W - 1 : cell);
  const int b = (int)(slot / M);
  float* dst = reinterpret_cast<float*>(gmap.ptr) + b * gmap.sb + c * gmap.sc + (cell / W) * gmap.sh + (cell % W) * gmap.sw;
  atomicAdd(dst, grad_loss[0] * stats[1] * g);
}

}  // namespace

int poly_iou_forward_launch(const cpoly_poly_iou_args& a, cudaStream_t stream) {
  Params p;
  p.poly = a.poly; p.dtype = a.dtype; p.B = a.B; p.M = a.M; p.N = a.N; p.H = a.H; p.W = a.W;
  p.ind = a.ind; p.mask = a.mask; p.target = a.target; p.iou = a.iou; p.dpred = a.dpred; p.status = a.status;
  const int slots = a.B * a.M;
  poly_iou_kernel<<<(slots + kWarps - 1) / kWarps, kWarps * 32, 0, stream>>>(p);
  int rc = check_launch("poly_iou_kernel");
  if (rc) return rc;
  poly_iou_reduce_kernel<<<1, 256, 0, stream>>>(a.iou, a.mask, a.status, slots, a.stats);
  return check_launch("poly_iou_reduce_kernel");
}

int poly_iou_backward_launch(const cpoly_poly_iou_args& a, cudaStream_t stream) {
  const int64_t n = (int64_t)a.B * a.M * 2 * a.N;
  poly_iou_scatter_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(a.dpred, a.mask, a.ind, a.stats, a.grad_loss, a.grad_poly,
                                                                          a.B, a.M, 2 * a.N, a.H, a.W);
  return check_launch("poly_iou_scatter_kernel");
}

}  // namespace cpoly
