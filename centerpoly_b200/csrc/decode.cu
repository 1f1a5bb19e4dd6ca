// Fused polydet decode: [sigmoid] -> 3x3 max-pool NMS -> per-image top-K -> gather of the reg /
// poly / pseudo-depth heads -> [polar -> cartesian] -> centre add -> bbox -> (B,K,2N+7).
//
// Replaces (reference paths): _nms decode.py:13-19, _topk decode.py:117-133,
// _transpose_and_gather_feat models/utils.py:22-26, polydet_decode decode.py:512-670 and optionally
// hm.sigmoid_() detectors/polydet.py:28.  The reference runs ~25 ATen kernels and transposes the whole
// 64-channel poly map (33.5 MB at 2048x1024) to fetch K x 64 values; here the heat map is read from HBM once
// (a second time from L2) and only K x (2N+3) head values are gathered.
//
// Exact top-K in five small launches:
//   1 bounds             the only pass over every pixel: per item (128 columns x 8 rows of one class plane) the largest
//                        value, and whether it is certainly an NMS peak (>= the one-pixel ring around the item).
//   2 T0                 per image: the K-th largest certified peak is a lower bound T0 of the K-th best peak score.
//   3 collect            items whose maximum reaches T0's bin (a few percent) are re-scanned with the full NMS: 3x3
//                        maximum by warp shuffles (horizontal) and rolling registers (vertical); a pixel is a peak iff it
//                        equals its 3x3 max (plateaus keep every member, like `hmax == heat`); peaks at or above that bin
//                        go to the image's candidate list as 64-bit keys (score << 32 | ~flat index).
//   4 select             one CTA per image: radix-select the top K candidates, bitonic-sort them (score desc, flat index
//                        asc - the oracle's tie rule).
//   5 gather             one warp per detection: reg / poly / depth at the K centres, polar -> cartesian, centre add,
//                        bbox, optional back-mapping, write (K, 2N+7).
// Degenerate maps whose candidates overflow the list (e.g. a constant map) take a slow exact path in launch 4 (the CTA
// re-scans the image itself).
//
// The two-stage top-K of the reference (per class, then over C*K) equals the global top-K over
// (class, y, x) whenever scores are tie-free; heat must be non-negative (it is post-sigmoid).
#include <math_constants.h>

#include <mutex>

#include "common.cuh"

namespace cpoly {
namespace {

constexpr int DNT = 256;           // threads per CTA
constexpr int DWARPS = DNT / 32;
constexpr int KMAX = 1024;         // largest supported K
constexpr int BAND = 128;          // columns per warp (4 per lane)
constexpr int SCAN_RH = 8;         // rows per warp work item of the two scan passes
constexpr int CAP = 4096;          // shared key list capacity of the select kernel
constexpr int HBITS = 13;          // histogram over score bits [30:18]: exponent + 5 mantissa bits
constexpr int HBINS = 1 << HBITS;
constexpr int HSHIFT = 31 - HBITS;
constexpr int CAND_CAP = 32768;    // candidate keys per image (global)

struct DecodeParams {
  int B, C, H, W, n_poly, K, rep, heat_is_logit, dtype;
  cpoly_tensor4 heat, polys, depth, reg;
  float* out;
  int64_t* out_inds;
  const double* post_affine;   // (B,6) inverse affine of polydet_post_process, or NULL
  unsigned long long* cand;    // [B][CAND_CAP]
  unsigned int* cand_count;    // [B]
  int* tbin;                   // [B] score bin (bits [30:18]) of T0, the lower bound of the K-th best peak (decode_t0_kernel)
  unsigned int* item_max;      // [B][items_per_image] upper bound (float bits) of the peak scores of every scan item
  unsigned int* item_cert;     // [B][items_per_image] that maximum when it is certainly an NMS peak, else 0 (float bits)
  unsigned int* todo_count;    // [1] items on the re-scan list (decode_t0_kernel appends, decode_collect_kernel walks)
  unsigned int* todo;          // [B * items_per_image] b * items_per_image + item of every item whose bound reaches its image's threshold bin
  int rh;                      // rows per warp work item
  int nbands, nrowchunks, items_per_image;
  int b0, nb;                  // image group handled by this launch (kernels 1 and 2)
};

__device__ __forceinline__ int score_bin(float v) { return (int)(__float_as_uint(v) >> HSHIFT) & (HBINS - 1); }

__device__ __forceinline__ unsigned int float_to_ordered(float f) {
  const unsigned int u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(unsigned int o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
__device__ __forceinline__ unsigned long long make_key(float score, unsigned int flat) {
  return ((unsigned long long)float_to_ordered(score) << 32) | (unsigned long long)(0xffffffffu - flat);
}

struct SelectScratch {
  unsigned int hist[256];
  unsigned long long outbuf[KMAX];
  int out_count;
  int digit;
  int kth;
};

// Keeps the K largest keys of list[0..n) (keys are unique) in list[0..K), unsorted.  All DNT threads
// must call it.  Returns the smallest kept key.
__device__ unsigned long long block_select_topk(unsigned long long* list, int n, int K, SelectScratch& s) {
  const int tid = threadIdx.x;
  unsigned long long prefix = 0ull, mask = 0ull;
  int kth = K;
  for (int pass = 7; pass >= 0; --pass) {
    const int shift = pass * 8;
    s.hist[tid & 255] = 0;  // DNT == 256
    __syncthreads();
    for (int i = tid; i < n; i += DNT) {
      const unsigned long long key = list[i];
      if ((key & mask) == prefix) atomicAdd(&s.hist[(unsigned int)(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      // lane l owns digits 255-8l .. 248-8l (descending); find where the running count reaches kth
      unsigned int mine[8], lane_sum = 0;
#pragma unroll
      for (int q = 0; q < 8; ++q) { mine[q] = s.hist[255 - 8 * tid - q]; lane_sum += mine[q]; }
      unsigned int incl = lane_sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (tid >= d) incl += t;
      }
      const unsigned int excl = incl - lane_sum;
      if (excl < (unsigned)kth && incl >= (unsigned)kth) {
        unsigned int run = excl;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (run + mine[q] >= (unsigned)kth) { s.digit = 255 - 8 * tid - q; s.kth = kth - (int)run; break; }
          run += mine[q];
        }
      }
    }
    __syncthreads();
    const int digit = s.digit;
    const unsigned int in_bucket = s.hist[digit];
    kth = s.kth;
    prefix |= (unsigned long long)digit << shift;
    mask |= 0xffull << shift;
    __syncthreads();
    if ((int)in_bucket == kth) break;  // the whole bucket is taken: (key & mask) >= prefix selects exactly K
  }
  if (tid == 0) s.out_count = 0;
  __syncthreads();
  for (int i = tid; i < n; i += DNT) {
    const unsigned long long key = list[i];
    if ((key & mask) >= prefix) {
      const int pos = atomicAdd(&s.out_count, 1);
      if (pos < KMAX) s.outbuf[pos] = key;
    }
  }
  __syncthreads();
  unsigned long long lo = ~0ull;
  for (int i = tid; i < K; i += DNT) {
    const unsigned long long key = s.outbuf[i];
    list[i] = key;
    lo = key < lo ? key : lo;
  }
  // block-wide min of the kept keys
  __shared__ unsigned long long red[DWARPS];
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    const unsigned long long o = __shfl_xor_sync(0xffffffffu, lo, d);
    lo = o < lo ? o : lo;
  }
  if ((tid & 31) == 0) red[tid >> 5] = lo;
  __syncthreads();
  lo = red[0];
#pragma unroll
  for (int w = 1; w < DWARPS; ++w) lo = red[w] < lo ? red[w] : lo;
  __syncthreads();
  return lo;
}

template <typename T>
__device__ __forceinline__ float load_heat(const T* p, bool logit) {
  float v = ld_as_float(p);
  return logit ? sigmoidf_accurate(v) : v;
}

// Row `y` of a band as 4 values per lane plus the two halo columns; out-of-image -> -inf.
template <typename T>
__device__ __forceinline__ void load_row(const T* plane, int64_t sh, int64_t sw, int y, int H, int W, int x0, int lane,
                                         bool logit, float v[4], float& left, float& right) {
  const float NEG = -CUDART_INF_F;
  const int xb = x0 + lane * 4;
  if (y < 0 || y >= H) {
    v[0] = v[1] = v[2] = v[3] = NEG; left = right = NEG;
    return;
  }
  const T* row = plane + (int64_t)y * sh;
  if (sizeof(T) == 4 && sw == 1 && xb + 3 < W && ((reinterpret_cast<uintptr_t>(row + xb) & 15) == 0)) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(row + xb));
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
    if (logit) {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = sigmoidf_accurate(v[j]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = (xb + j < W) ? load_heat(row + (int64_t)(xb + j) * sw, logit) : NEG;
  }
  // neighbours inside the band come from the adjacent lanes; the band edges are loaded
  left = __shfl_up_sync(0xffffffffu, v[3], 1);
  right = __shfl_down_sync(0xffffffffu, v[0], 1);
  if (lane == 0) left = (x0 > 0) ? load_heat(row + (int64_t)(x0 - 1) * sw, logit) : NEG;
  if (lane == 31) right = (x0 + BAND < W) ? load_heat(row + (int64_t)(x0 + BAND) * sw, logit) : NEG;
}

__device__ __forceinline__ void hmax3(const float v[4], float left, float right, float h[4]) {
  h[0] = fmaxf(fmaxf(left, v[0]), v[1]);
  h[1] = fmaxf(fmaxf(v[0], v[1]), v[2]);
  h[2] = fmaxf(fmaxf(v[1], v[2]), v[3]);
  h[3] = fmaxf(fmaxf(v[2], v[3]), right);
}

template <typename T>
__device__ __forceinline__ float head_value(const cpoly_tensor4& t, int b, int ch, int y, int x) {
  return ld_as_float(reinterpret_cast<const T*>(t.ptr) + (int64_t)b * t.sb + (int64_t)ch * t.sc + (int64_t)y * t.sh +
                     (int64_t)x * t.sw);
}


// One warp scans its item (a 128-column band x RH rows of one class plane) and calls f(is_peak, score, max3x3, flat)
// for EVERY pixel slot (flat = c*H*W + y*W + x; is_peak = NMS peak with a positive score) - unconditionally, so the
// caller decides between predicated work (histogram: one predicated RED.shared, no branch) and a rare warp-uniform
// branch (candidate collection).  Out-of-image rows / columns hold -inf and can never be peaks: no bounds tests.
// All RH + 2 rows (one halo row above and below) are LOADED FIRST - RH + 2 independent 16-byte loads per lane in
// flight (5 KB per warp) - and only then shuffled / compared: with one row per loop iteration a warp paid one DRAM
// latency per row, and a 64-image batch ran at 5 % of the HBM roof (tools/decode_bench.py).
// Loads of one item (a 128-column band x RH rows of one class plane, plus one halo row above / below and one halo column
// left / right): v[i][j] = row r0 - 1 + i, columns xb .. xb + 3 of this lane; hv[i] = the halo column of that row (lane 0:
// left, lane 31: right).  Raw values (logits stay logits); out-of-image -> -inf.  ALL RH + 2 rows are requested before
// anything is consumed.
template <typename T, int RH>
__device__ __forceinline__ void load_item(const DecodeParams& p, int b, int item, int lane, float (&v)[RH + 2][4], float (&hv)[RH + 2],
                                          int& c, int& r0, int& xb) {
  constexpr int NR = RH + 2;
  const float NEG = -CUDART_INF_F;
  int t = item;
  const int band = t % p.nbands; t /= p.nbands;
  const int rc = t % p.nrowchunks;
  c = t / p.nrowchunks;
  r0 = rc * RH;
  const int x0 = band * BAND;
  const int64_t sh = p.heat.sh, sw = p.heat.sw;
  const T* plane = reinterpret_cast<const T*>(p.heat.ptr) + (int64_t)b * p.heat.sb + (int64_t)c * p.heat.sc;
  xb = x0 + lane * 4;
  const int xbl = xb;
  // vector path: unit column stride, every row start aligned like the plane
  const bool aligned = sw == 1 && (sh % 4 == 0) && ((reinterpret_cast<uintptr_t>(plane + x0) % (4 * sizeof(T))) == 0);
  const bool full = aligned && x0 + BAND <= p.W;   // warp-uniform: the whole band is inside the row
  // lane 0 fetches the column left of the band, lane 31 the one right of it: one halo register per row
  const bool want_h = (lane == 0 && x0 > 0) || (lane == 31 && x0 + BAND < p.W);
  const int hx = lane == 0 ? x0 - 1 : x0 + BAND;
  auto vec_load = [&](const T* row, float (&o)[4]) {
    if (sizeof(T) == 4) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(row + xbl));
      o[0] = q.x; o[1] = q.y; o[2] = q.z; o[3] = q.w;
    } else {
      const uint2 q = __ldg(reinterpret_cast<const uint2*>(row + xbl));
      o[0] = __uint_as_float(q.x << 16); o[1] = __uint_as_float(q.x & 0xffff0000u);
      o[2] = __uint_as_float(q.y << 16); o[3] = __uint_as_float(q.y & 0xffff0000u);
    }
  };
  if (full) {
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      const int y = r0 - 1 + i;
      const bool rv = (unsigned)y < (unsigned)p.H;
      const T* row = plane + (int64_t)(rv ? y : 0) * sh;
      v[i][0] = v[i][1] = v[i][2] = v[i][3] = NEG;
      hv[i] = NEG;
      if (rv) {
        vec_load(row, v[i]);
        if (want_h) hv[i] = ld_as_float(row + hx);
      }
    }
  } else if (aligned && p.W % 4 == 0) {   // ragged band, but every lane's 4 columns are all inside or all outside the row
    const bool in = xbl < p.W;
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      const int y = r0 - 1 + i;
      const bool rv = (unsigned)y < (unsigned)p.H;
      const T* row = plane + (int64_t)(rv ? y : 0) * sh;
      v[i][0] = v[i][1] = v[i][2] = v[i][3] = NEG;
      hv[i] = NEG;
      if (rv) {
        if (in) vec_load(row, v[i]);
        if (want_h) hv[i] = ld_as_float(row + hx);
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      const int y = r0 - 1 + i;
      const bool rv = (unsigned)y < (unsigned)p.H;
      const T* row = plane + (int64_t)(rv ? y : 0) * sh;
      v[i][0] = v[i][1] = v[i][2] = v[i][3] = NEG;
      hv[i] = NEG;
      if (rv) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (xbl + j < p.W) v[i][j] = ld_as_float(row + (int64_t)(xbl + j) * sw);
        if (want_h) hv[i] = ld_as_float(row + (int64_t)hx * sw);
      }
    }
  }
}

template <typename T, int RH, typename F>
__device__ __forceinline__ void scan_item(const DecodeParams& p, int b, int item, int lane, F&& f) {
  constexpr int NR = RH + 2;
  const float NEG = -CUDART_INF_F;
  const bool logit = p.heat_is_logit != 0;
  float v[NR][4], hv[NR];
  int c, r0, xb;
  load_item<T, RH>(p, b, item, lane, v, hv, c, r0, xb);
  // ---- phase 2 (rolling, three rows of horizontal maxima live): logistic (fused hm.sigmoid_), horizontal
  // neighbours + 3-max of row i, then vertical 3-max + peak test of row i - 1
  const unsigned int flat0 = (unsigned int)(c * (p.H * p.W) + r0 * p.W + xb);
  float h_prev[4], h_cur[4], h_next[4];
#pragma unroll
  for (int i = 0; i < NR; ++i) {
    if (logit) {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[i][j] = v[i][j] == NEG ? NEG : sigmoidf_accurate(v[i][j]);
      if (hv[i] != NEG) hv[i] = sigmoidf_accurate(hv[i]);
    }
    float left = __shfl_up_sync(0xffffffffu, v[i][3], 1);
    float right = __shfl_down_sync(0xffffffffu, v[i][0], 1);
    if (lane == 0) left = hv[i];
    if (lane == 31) right = hv[i];
    hmax3(v[i], left, right, h_next);
    if (i >= 2) {
      const int rr = i - 2;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float m = fmaxf(fmaxf(h_prev[j], h_cur[j]), h_next[j]);
        const float cur = v[rr + 1][j];
        f(cur == m && cur > 0.f, cur, m, flat0 + (unsigned int)(rr * p.W + j));
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { h_prev[j] = h_cur[j]; h_cur[j] = h_next[j]; }
  }
}

// Work items of the image group [b0, b0+nb) are dealt to warps in CTA-contiguous blocks.
__device__ __forceinline__ void cta_item_range(const DecodeParams& p, int64_t& beg, int64_t& end) {
  const int64_t total = (int64_t)p.nb * p.items_per_image;
  const int64_t per = (total + gridDim.x - 1) / gridDim.x;
  beg = (int64_t)blockIdx.x * per;
  end = beg + per < total ? beg + per : total;
}

// ------------------------------------------------------------------ launch 1: per-item bounds (the streaming pass)
// The one pass that touches every pixel does only load + max (3 instructions per pixel; the histogram scan it replaces
// needed 31 and was issue-bound at 39 % of the HBM roof).  Per item (128 columns x 8 rows of one class plane):
//   item_max  = largest value inside the item        -> upper bound of its peak scores (the collection pass skips items
//                                                       whose bound is below the image's threshold)
//   item_cert = item_max if it is >= every value of the one-pixel ring around the item (and > 0), else 0.  A certified
//               maximum IS an NMS peak (it equals its 3x3 maximum; plateaus keep every member, decode.py:13-19), so the
//               K-th largest certificate of an image is a LOWER bound T0 of its K-th best peak score: every top-K
//               detection has score >= T0, and for an untrained heat map only about K pixels in a million reach T0.
// Logits: maxima are taken on the logits (the logistic function is monotone) and converted once per item.
template <typename T>
__global__ void __launch_bounds__(DNT, 3) decode_bound_kernel(DecodeParams p) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int64_t beg, end;
  cta_item_range(p, beg, end);
  for (int64_t it = beg + warp; it < end; it += DWARPS) {
    const int lb = (int)(it / p.items_per_image), item = (int)(it - (int64_t)lb * p.items_per_image);
    constexpr int NR = SCAN_RH + 2;
    float v[NR][4], hv[NR];
    int c, r0, xb;
    load_item<T, SCAN_RH>(p, p.b0 + lb, item, lane, v, hv, c, r0, xb);
    float m_in = -CUDART_INF_F, m_ring = fmaxf(hv[0], hv[NR - 1]);
#pragma unroll
    for (int j = 0; j < 4; ++j) m_ring = fmaxf(m_ring, fmaxf(v[0][j], v[NR - 1][j]));
#pragma unroll
    for (int i = 1; i < NR - 1; ++i) {
      m_ring = fmaxf(m_ring, hv[i]);
#pragma unroll
      for (int j = 0; j < 4; ++j) m_in = fmaxf(m_in, v[i][j]);
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
      m_in = fmaxf(m_in, __shfl_xor_sync(0xffffffffu, m_in, d));
      m_ring = fmaxf(m_ring, __shfl_xor_sync(0xffffffffu, m_ring, d));
    }
    if (lane == 0) {
      const bool cert = m_in >= m_ring;
      float mx = m_in;
      if (p.heat_is_logit && mx != -CUDART_INF_F) mx = sigmoidf_accurate(mx);
      mx = mx > 0.f ? mx : 0.f;
      const int64_t slot = (int64_t)(p.b0 + lb) * p.items_per_image + item;
      p.item_max[slot] = __float_as_uint(mx);
      p.item_cert[slot] = __float_as_uint(cert ? mx : 0.f);
    }
  }
}

// ------------------------------------------------------------------ launch 2: T0 of every image (one CTA per image)
// tbin[b] = histogram bin (score bits [30:18]) of the K-th largest certificate, 0 when the image has fewer than K.
template <typename T>
__global__ void __launch_bounds__(DNT) decode_t0_kernel(DecodeParams p) {
  __shared__ unsigned long long list[CAP];
  __shared__ SelectScratch scratch;
  __shared__ int s_pos;
  const int tid = threadIdx.x;
  const int b = p.b0 + blockIdx.x;
  const unsigned int* cert = p.item_cert + (int64_t)b * p.items_per_image;
  const int K = p.K;
  if (tid == 0) s_pos = 0;
  __syncthreads();
  int n = 0, mine = 0;
  for (int off = 0; off < p.items_per_image;) {
    if (n == CAP) {   // uniform: keep the K largest so far
      block_select_topk(list, n, K, scratch);
      n = K;
    }
    const int take = min(CAP - n, p.items_per_image - off);
    for (int i = tid; i < take; i += DNT) {
      const float c = __uint_as_float(__ldcg(cert + off + i));
      list[n + i] = make_key(c, (unsigned int)(off + i));
      mine += c > 0.f ? 1 : 0;
    }
    n += take;
    off += take;
    __syncthreads();
  }
  atomicAdd(&s_pos, mine);
  __syncthreads();
  int tb = 0;
  if (s_pos >= K) {   // uniform; at least K certified peaks: the K-th largest of them bounds the K-th best peak from below
    const unsigned long long lo = block_select_topk(list, n, K, scratch);
    const float t0 = ordered_to_float((unsigned int)(lo >> 32));
    tb = t0 > 0.f ? score_bin(t0) : 0;
  }
  if (tid == 0) p.tbin[b] = tb;
  // the items of this image whose best peak can reach the threshold bin go on ONE batch-wide list: the collection pass
  // then deals them out one per warp, whatever their distribution over the images (a warp used to re-scan the flagged
  // ones among ITS 22 consecutive items serially: 34 us at B = 64).  The list order depends on the atomics; the
  // candidate list it feeds is unordered anyway (select sorts).
  if (!p.todo) return;          // small batch: decode_collect_direct_kernel needs no list
  const unsigned int* imax = p.item_max + (int64_t)b * p.items_per_image;
  for (int i0 = 0; i0 < p.items_per_image; i0 += DNT) {
    const int i = i0 + tid;
    bool need = false;
    if (i < p.items_per_image) {
      const float mx = __uint_as_float(__ldcg(imax + i));
      need = mx > 0.f && score_bin(mx) >= tb;
    }
    const unsigned int m = __ballot_sync(0xffffffffu, need);
    if (m) {
      unsigned int base = 0;
      if ((tid & 31) == 0) base = atomicAdd(p.todo_count, (unsigned int)__popc(m));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (need) p.todo[base + __popc(m & ((1u << (tid & 31)) - 1u))] = (unsigned int)b * (unsigned int)p.items_per_image + (unsigned int)i;
    }
  }
}

// ------------------------------------------------------------------ launch 3: candidate collection
// Only the items whose best peak reaches the image's threshold bin are re-scanned (about K + one bin's worth of peaks
// per image lie there), so this pass touches a few percent of the heat map instead of all of it again.
template <typename T>
__global__ void __launch_bounds__(DNT, 3) decode_collect_kernel(DecodeParams p) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned int n_todo = __ldcg(p.todo_count);
  for (unsigned int w = blockIdx.x * DWARPS + warp; w < n_todo; w += gridDim.x * DWARPS) {   // one listed item per warp and round
    const unsigned int e = __ldcg(p.todo + w);
    const int b = (int)(e / (unsigned int)p.items_per_image), item = (int)(e - (unsigned int)b * (unsigned int)p.items_per_image);
    const int tb = __ldcg(p.tbin + b);
    unsigned long long* cand = p.cand + (int64_t)b * CAND_CAP;
    unsigned int* cnt = p.cand_count + b;
    scan_item<T, SCAN_RH>(p, b, item, lane, [&](bool pk, float s, float, unsigned int flat) {
      const bool take = pk && score_bin(s) >= tb;
      if (__any_sync(0xffffffffu, take)) {
        if (take) {
          const unsigned int pos = atomicAdd(cnt, 1u);   // the compiler aggregates same-address atomics per warp
          if (pos < (unsigned)CAND_CAP) cand[pos] = make_key(s, flat);
        }
      }
    });
  }
}

// Small batches (every item can have a warp of its own): no list, the warp of an item tests its bound itself.
template <typename T>
__global__ void __launch_bounds__(DNT, 3) decode_collect_direct_kernel(DecodeParams p, int ipw) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t total = (int64_t)p.nb * p.items_per_image;
  // a warp looks at `ipw` (<= 32) consecutive items at once (one per lane: coalesced reads of item_max) and re-scans
  // the few whose best peak reaches their image's threshold bin.  ipw shrinks with the batch so that a single frame
  // (1024 items) still spreads over ~1000 warps: the re-scans of one warp are serial, and a flat heat map (bf16
  // logits of an untrained network) can qualify most items.
  for (int64_t base = ((int64_t)blockIdx.x * DWARPS + warp) * ipw; base < total; base += (int64_t)gridDim.x * DWARPS * ipw) {
    const int64_t mine = base + lane;
    bool need = false;
    if (lane < ipw && mine < total) {
      const int lb = (int)(mine / p.items_per_image);
      const float mx = __uint_as_float(__ldcg(p.item_max + (int64_t)p.b0 * p.items_per_image + mine));
      need = mx > 0.f && score_bin(mx) >= __ldcg(p.tbin + p.b0 + lb);
    }
    unsigned int todo = __ballot_sync(0xffffffffu, need);
    while (todo) {   // warp-uniform
      const int i = __ffs(todo) - 1;
      todo &= todo - 1;
      const int64_t it = base + i;
      const int lb = (int)(it / p.items_per_image), item = (int)(it - (int64_t)lb * p.items_per_image);
      const int b = p.b0 + lb;
      const int tb = __ldcg(p.tbin + b);
      unsigned long long* cand = p.cand + (int64_t)b * CAND_CAP;
      unsigned int* cnt = p.cand_count + b;
      scan_item<T, SCAN_RH>(p, b, item, lane, [&](bool pk, float s, float, unsigned int flat) {
        const bool take = pk && score_bin(s) >= tb;
        if (__any_sync(0xffffffffu, take)) {
          if (take) {
            const unsigned int pos = atomicAdd(cnt, 1u);   // the compiler aggregates same-address atomics per warp
            if (pos < (unsigned)CAND_CAP) cand[pos] = make_key(s, flat);
          }
        }
      });
    }
  }
}

// ------------------------------------------------------------------ launch 4: select + sort (one CTA per image)
template <typename T>
__global__ void __launch_bounds__(DNT) decode_select_kernel(DecodeParams p) {
  __shared__ unsigned long long list[CAP];
  __shared__ SelectScratch scratch;
  __shared__ int s_count;
  __shared__ unsigned char s_taken[KMAX];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int K = p.K;
  const int HW = p.H * p.W;
  const unsigned int total = __ldcg(p.cand_count + b);
  int n = 0;
  if (total <= (unsigned)CAND_CAP) {
    // ---- normal path: the candidate list holds every peak at or above the threshold bin
    const unsigned long long* src = p.cand + (int64_t)b * CAND_CAP;
    for (unsigned int off = 0; off < total;) {
      if (n == CAP) {  // uniform
        block_select_topk(list, n, K, scratch);
        n = K;
      }
      const unsigned int take = min((unsigned)(CAP - n), total - off);
      for (unsigned int i = tid; i < take; i += DNT) list[n + i] = __ldcg(src + off + i);
      n += (int)take;
      off += take;
      __syncthreads();
    }
  } else {
    // ---- slow exact path (degenerate map: the boundary bin alone overflows the candidate list): this CTA re-scans
    // the image two rows per warp and round, pruning the shared list with a block radix select whenever it could
    // overflow; after the first prune the K-th best key is a threshold that rejects most peaks at once.
    DecodeParams q = p;
    q.rh = 2;
    q.nrowchunks = (p.H + 1) / 2;
    q.items_per_image = p.C * q.nrowchunks * q.nbands;
    constexpr int ROUND_MAX = DWARPS * BAND * 2;
    unsigned long long threshold = 0ull;
    if (tid == 0) s_count = 0;
    __syncthreads();
    for (int base = 0; base < q.items_per_image; base += DWARPS) {
      const int it = base + warp;
      if (it < q.items_per_image) {
        scan_item<T, 2>(q, b, it, lane, [&](bool pk, float s, float, unsigned int flat) {
          const unsigned long long key = make_key(s, flat);
          if (pk && key > threshold) {
            const int pos = atomicAdd(&s_count, 1);
            if (pos < CAP) list[pos] = key;  // cannot overflow: a prune precedes any round that could
          }
        });
      }
      __syncthreads();
      if (s_count > CAP - ROUND_MAX) {  // uniform
        threshold = block_select_topk(list, s_count, K, scratch);
        if (tid == 0) s_count = K;
        __syncthreads();
      }
    }
    n = s_count;
    __syncthreads();
  }
  if (n > K) {
    block_select_topk(list, n, K, scratch);
    n = K;
  }
  // Fewer than K positive peaks: fill with zero-score pixels in ascending flat order (our tie rule;
  // the reference returns arbitrary zero-score indices here).
  if (n < K) {
    for (int i = tid; i < K; i += DNT) s_taken[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += DNT) {
      const unsigned int flat = 0xffffffffu - (unsigned int)(list[i] & 0xffffffffull);
      if (flat < (unsigned int)K) s_taken[flat] = 1;
    }
    __syncthreads();
    const int need = K - n;
    for (int f = tid; f < K; f += DNT) {
      if (s_taken[f]) continue;
      int rank = 0;
      for (int u = 0; u < f; ++u) rank += s_taken[u] ? 0 : 1;
      if (rank < need) list[n + rank] = make_key(0.f, (unsigned int)f);
    }
    __syncthreads();
    n = K;
  }
  // bitonic sort, descending, on the next power of two (padding keys are 0 = smallest)
  int npow = 1;
  while (npow < K) npow <<= 1;
  for (int i = K + tid; i < npow; i += DNT) list[i] = 0ull;
  __syncthreads();
  for (int size = 2; size <= npow; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < npow; i += DNT) {
        const int partner = i ^ stride;
        if (partner > i) {
          const bool desc = ((i & size) == 0);
          const unsigned long long a = list[i], bb = list[partner];
          if ((a < bb) == desc) { list[i] = bb; list[partner] = a; }
        }
      }
      __syncthreads();
    }

  // ---- the K sorted keys go back to the head of this image's candidate list (all of it is in shared memory by now);
  // decode_gather_kernel, one warp per detection over the whole batch, turns them into output rows.
  unsigned long long* dst = p.cand + (int64_t)b * CAND_CAP;
  for (int k = tid; k < K; k += DNT) dst[k] = list[k];
}

// ------------------------------------------------------------------ launch 5: gather + polygon assembly
// One warp per detection, B*K warps in flight: inside the select kernel (one CTA per image, 8 warps) the ~13 rounds
// of dependent head reads of a K = 100 image cost 25-30 us of pure DRAM latency.
template <typename T>
__global__ void __launch_bounds__(DNT) decode_gather_kernel(DecodeParams p) {
  const int lane = threadIdx.x & 31;
  const int K = p.K;
  const int HW = p.H * p.W;
  const int64_t det = (int64_t)blockIdx.x * DWARPS + (threadIdx.x >> 5);
  if (det >= (int64_t)p.B * K) return;
  const int b = (int)(det / K), k = (int)(det - (int64_t)b * K);
  const int N = p.n_poly / 2;
  const int row_len = p.n_poly + 7;
  {
    const unsigned long long key = __ldcg(p.cand + (int64_t)b * CAND_CAP + k);
    const unsigned int flat = 0xffffffffu - (unsigned int)(key & 0xffffffffull);
    const float score = ordered_to_float((unsigned int)(key >> 32));
    const int cls = (int)(flat / (unsigned int)HW);
    const int pix = (int)(flat - (unsigned int)cls * (unsigned int)HW);
    const int y = pix / p.W, x = pix - y * p.W;
    const float dep = head_value<T>(p.depth, b, 0, y, x);   // issued with the other head reads, used last
    float xs, ys;
    if (p.reg.ptr) {  // decode.py:523-527
      xs = __fadd_rn((float)x, head_value<T>(p.reg, b, 0, y, x));
      ys = __fadd_rn((float)y, head_value<T>(p.reg, b, 1, y, x));
    } else {          // decode.py:529-530
      xs = (float)x + 0.5f;
      ys = (float)y + 0.5f;
    }
    float* orow = p.out + ((int64_t)b * K + k) * row_len;
    // optional back-mapping to the original image (utils/post_process.py:105-112): fp64 like the reference's np.dot
    double t00 = 1.0, t01 = 0.0, t02 = 0.0, t10 = 0.0, t11 = 1.0, t12 = 0.0;
    const bool post = p.post_affine != nullptr;
    if (post) {
      const double* t = p.post_affine + (int64_t)b * 6;
      t00 = t[0]; t01 = t[1]; t02 = t[2]; t10 = t[3]; t11 = t[4]; t12 = t[5];
    }
    float xmin = CUDART_INF_F, ymin = CUDART_INF_F, xmax = -CUDART_INF_F, ymax = -CUDART_INF_F;
    for (int vtx = lane; vtx < N; vtx += 32) {
      float px = head_value<T>(p.polys, b, 2 * vtx, y, x);
      float py = head_value<T>(p.polys, b, 2 * vtx + 1, y, x);
      if (p.rep != CPOLY_REP_CARTESIAN) {  // decode.py:582-614: cos/sin in double, product in fp32
        const double ang = (p.rep == CPOLY_REP_POLAR_FIXED)
                               ? (2 * 3.14 - 2 * 3.14 / (double)p.n_poly * (double)(2 * vtx))
                               : (double)py;
        const float r = px;
        px = __fmul_rn(r, (float)cos(ang));
        py = __fmul_rn(r, (float)sin(ang));
      }
      px = __fadd_rn(px, xs);  // decode.py:624-625
      py = __fadd_rn(py, ys);
      if (post) {   // the box below is the min/max of the UNMAPPED vertices (decode.py:647-659), mapped afterwards
        orow[6 + 2 * vtx] = (float)(t00 * (double)px + t01 * (double)py + t02);
        orow[6 + 2 * vtx + 1] = (float)(t10 * (double)px + t11 * (double)py + t12);
      } else {
        orow[6 + 2 * vtx] = px;
        orow[6 + 2 * vtx + 1] = py;
      }
      xmin = fminf(xmin, px); xmax = fmaxf(xmax, px);
      ymin = fminf(ymin, py); ymax = fmaxf(ymax, py);
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
      xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, d));
      ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, d));
      xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, d));
      ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, d));
    }
    if (lane == 0) {
      if (post) {   // post_process.py:109-110: the two corners are mapped as points
        orow[0] = (float)(t00 * (double)xmin + t01 * (double)ymin + t02);
        orow[1] = (float)(t10 * (double)xmin + t11 * (double)ymin + t12);
        orow[2] = (float)(t00 * (double)xmax + t01 * (double)ymax + t02);
        orow[3] = (float)(t10 * (double)xmax + t11 * (double)ymax + t12);
      } else {
        orow[0] = xmin; orow[1] = ymin; orow[2] = xmax; orow[3] = ymax;  // decode.py:647-659
      }
      orow[4] = score;
      orow[5] = (float)cls;
      orow[6 + p.n_poly] = dep;
      if (p.out_inds) p.out_inds[(int64_t)b * K + k] = pix;
    }
    }
}

void plan(int B, int C, int H, int W, DecodeParams& p) {
  (void)B;
  p.nbands = (W + BAND - 1) / BAND;
  p.rh = SCAN_RH;
  p.nrowchunks = (H + p.rh - 1) / p.rh;
  p.items_per_image = C * p.nrowchunks * p.nbands;
}

size_t align256(size_t v) { return (v + 255) / 256 * 256; }

}  // namespace

static size_t decode_zero_bytes(int B) { return (2 * (size_t)B + 1) * sizeof(unsigned int); }   // cand_count + tbin + todo_count
static size_t decode_items_per_image(int C, int H, int W) {
  return (size_t)C * ((H + SCAN_RH - 1) / SCAN_RH) * ((W + BAND - 1) / BAND);
}
size_t decode_workspace_bytes(int B, int C, int H, int W, int K) {
  (void)K;
  return align256(decode_zero_bytes(B)) + 3 * align256((size_t)B * decode_items_per_image(C, H, W) * sizeof(unsigned int)) +
         align256((size_t)B * CAND_CAP * sizeof(unsigned long long)) + 256;
}

int decode_launch(const cpoly_decode_args& a, cudaStream_t stream) {
  DecodeParams p;
  p.B = a.B; p.C = a.C; p.H = a.H; p.W = a.W; p.n_poly = a.n_poly; p.K = a.K; p.rep = a.rep;
  p.heat_is_logit = a.heat_is_logit; p.dtype = a.dtype;
  p.heat = a.heat; p.polys = a.polys; p.depth = a.depth; p.reg = a.reg;
  p.out = a.out; p.out_inds = a.out_inds; p.post_affine = a.post_affine;
  plan(a.B, a.C, a.H, a.W, p);
  const size_t need = decode_workspace_bytes(a.B, a.C, a.H, a.W, a.K);
  CPOLY_REQUIRE(a.workspace && a.workspace_bytes >= need, CPOLY_EWORKSPACE,
                "decode workspace too small: need %zu bytes, have %zu", need, a.workspace_bytes);
  CPOLY_REQUIRE((reinterpret_cast<uintptr_t>(a.workspace) & 255) == 0, CPOLY_EINVAL, "decode workspace must be 256-byte aligned");
  CPOLY_REQUIRE((uint64_t)a.B * decode_items_per_image(a.C, a.H, a.W) < 0xffffffffull, CPOLY_EINVAL, "too many scan items for 32-bit work-list entries");
  char* ws = reinterpret_cast<char*>(a.workspace);
  const size_t zero_bytes = decode_zero_bytes(a.B);
  p.cand_count = reinterpret_cast<unsigned int*>(ws);
  p.tbin = reinterpret_cast<int*>(p.cand_count + a.B);
  p.todo_count = p.cand_count + 2 * a.B;
  ws += align256(zero_bytes);
  p.item_max = reinterpret_cast<unsigned int*>(ws);
  ws += align256((size_t)a.B * decode_items_per_image(a.C, a.H, a.W) * sizeof(unsigned int));
  p.item_cert = reinterpret_cast<unsigned int*>(ws);
  ws += align256((size_t)a.B * decode_items_per_image(a.C, a.H, a.W) * sizeof(unsigned int));
  p.todo = reinterpret_cast<unsigned int*>(ws);
  ws += align256((size_t)a.B * decode_items_per_image(a.C, a.H, a.W) * sizeof(unsigned int));
  p.cand = reinterpret_cast<unsigned long long*>(ws);
  CPOLY_CUDA_OK(cudaMemsetAsync(p.cand_count, 0, zero_bytes, stream));
  int dev = 0, sms = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const bool bf = a.dtype == CPOLY_BF16;
  {
    // one pass over the whole batch: the collection pass re-reads only the few items that hold top-K candidates,
    // so there is no L2-residency reason to split the batch into groups any more
    p.b0 = 0;
    p.nb = a.B;
    const int64_t items = (int64_t)p.nb * p.items_per_image;
    // every item gets a warp of its own in one resident wave? then the collection needs no work list
    const bool use_list = items > (int64_t)3 * sms * DWARPS;
    if (!use_list) p.todo = nullptr;
    int64_t ctas = (items + DWARPS - 1) / DWARPS;
    if (ctas > 3 * sms) ctas = 3 * sms;   // one resident wave: 24 warps per SM, each with (SCAN_RH + 2) x 512 B of loads in flight
    if (bf) decode_bound_kernel<__nv_bfloat16><<<(unsigned)ctas, DNT, 0, stream>>>(p);
    else decode_bound_kernel<float><<<(unsigned)ctas, DNT, 0, stream>>>(p);
    if (int rc = check_launch("decode_bound_kernel")) return rc;
    if (bf) decode_t0_kernel<__nv_bfloat16><<<(unsigned)a.B, DNT, 0, stream>>>(p);
    else decode_t0_kernel<float><<<(unsigned)a.B, DNT, 0, stream>>>(p);
    if (int rc = check_launch("decode_t0_kernel")) return rc;
    if (use_list) {
      int64_t cctas = (items + DWARPS - 1) / DWARPS;      // at most one listed item per warp in the first round
      if (cctas > 3 * sms) cctas = 3 * sms;
      if (bf) decode_collect_kernel<__nv_bfloat16><<<(unsigned)cctas, DNT, 0, stream>>>(p);
      else decode_collect_kernel<float><<<(unsigned)cctas, DNT, 0, stream>>>(p);
    } else {
      const int64_t cctas = (items + DWARPS - 1) / DWARPS;   // one item per warp
      if (bf) decode_collect_direct_kernel<__nv_bfloat16><<<(unsigned)cctas, DNT, 0, stream>>>(p, 1);
      else decode_collect_direct_kernel<float><<<(unsigned)cctas, DNT, 0, stream>>>(p, 1);
    }
    if (int rc = check_launch("decode_collect_kernel")) return rc;
  }
  if (bf) decode_select_kernel<__nv_bfloat16><<<(unsigned)a.B, DNT, 0, stream>>>(p);
  else decode_select_kernel<float><<<(unsigned)a.B, DNT, 0, stream>>>(p);
  if (int rc = check_launch("decode_select_kernel")) return rc;
  const unsigned gblocks = (unsigned)(((int64_t)a.B * a.K + DWARPS - 1) / DWARPS);
  if (bf) decode_gather_kernel<__nv_bfloat16><<<gblocks, DNT, 0, stream>>>(p);
  else decode_gather_kernel<float><<<gblocks, DNT, 0, stream>>>(p);
  return check_launch("decode_gather_kernel");
}

// ---- allocation-free variant: the scratch is a file-scope device array -------------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, serving loops, harnesses that forbid it).
// Every GPU that loads the library holds these 4 MiB until the process exits.  Every call on a device shares the array:
// two cpoly_decode_static calls that overlap on the GPU (different streams) race.  The array keeps the last call's
// contents; decode_launch clears the only part that must start at zero (candidate counters + threshold bins) itself.
static __device__ __align__(256) unsigned char g_decode_static_scratch[CPOLY_DECODE_STATIC_BYTES];

// In host code `g_decode_static_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
int decode_static_scratch(void** ptr) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static void* addr[MAX_DEV] = {};
  int dev = -1;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_REQUIRE(dev >= 0 && dev < MAX_DEV, CPOLY_EUNSUPPORTED, "device ordinal %d outside the static-scratch table", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* q = nullptr;
    CPOLY_CUDA_OK(cudaGetSymbolAddress(&q, g_decode_static_scratch));
    CPOLY_REQUIRE(q && (reinterpret_cast<uintptr_t>(q) & 255) == 0, CPOLY_ECUDA, "static decode scratch has no aligned device address");
    addr[dev] = q;
  }
  *ptr = addr[dev];
  return CPOLY_OK;
}

}  // namespace cpoly
