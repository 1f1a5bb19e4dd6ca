// Atomics-free, bit-reproducible col2im of the DCNv2 backward (grad_input), sm_100a.
//
// Replaces upstream DCNv2's `modulated_deformable_col2im_gpu_kernel` (one fp32 atomicAdd per (channel, tap, pixel,
// corner); the extension is the missing submodule src/lib/models/networks/DCNv2, reached from loss.backward() in
// src/lib/trains/base_trainer.py:97-98).  The data-gradient kernel (dcn_tc_bwd.cu, tcgen05) stores the column
// gradient gcol[pixel][tap][c] = sum_o gout[pixel,o] * W[o,c,tap] with plain coalesced stores; this file turns it into
//
//   grad_input[q, c] = sum over (pixel p, tap k, corner j) with corner_j(p,k) == q of  m(p,k) * w_j(p,k) * gcol[p,k,c]
//
// in GATHER form: a CTA owns one 8x16 destination tile D of grad_input, finds every (p,k,j) that lands in D, sorts them
// by destination pixel with a deterministic counting sort, and each destination pixel is then summed by ONE thread
// group in a fixed order and written exactly once.  No floating-point atomic, no pre-zeroing, same bits every run.
//
//   dcn_bbox_kernel     per source tile: one 16-byte record (y0, x0, ly, lx, m) per (pixel, tap) pair, laid out
//                       [tile][tap][pixel] so that a warp of the col2im reads 32 pairs with one coalesced 512-byte load,
//                       plus the bounding box of the corner positions of every (tap, 32-pixel row pair) ("unit", 36 per
//                       tile) and of the whole tile, as int16 x 4 (8 bytes).
//   dcn_col2im_kernel   per destination tile D:
//     1. candidate source tiles = tiles of the image whose box meets D; candidate units likewise (ordered block
//        compactions): with ordinary offsets that is D's own 36 units plus the rim of its neighbours, with far
//        offsets (|d| of tens of pixels) any tile of the image - no halo limit, no spill list.
//     2. pass 1 counts, per (warp, corner, destination pixel), the corners that land in D: the 32 lanes of a warp hold
//        32 source pixels of one unit, ONE __match_any_sync per unit on the base corner groups the lanes, and the
//        four corners of a lane go to four separate slot counters (corners of one base pixel are distinct pixels),
//        so counters are warp-private and need no atomics.  A scan gives every (warp, corner, pixel) its slot range;
//        pass 2 repeats the walk and writes (source row, m*w) into its slot: position = base + rank among the
//        matching lanes - a function of the DATA only.
//     3. eight lanes per destination pixel walk its slot range in order, 16 bytes of gcol per lane and entry,
//        fp32 accumulation, one store.
//     A destination tile that receives more than CAP contributions (offset fields that collapse onto one spot) is
//     processed in windows of CAP sorted slots; later windows add to what the first one stored (same thread, fixed order).
#include "dcn_sampler.cuh"

namespace cpoly {
namespace {

using namespace tc;
using namespace smp;

constexpr int C2I_THREADS = 512;
constexpr int C2I_WARPS = C2I_THREADS / 32;
constexpr int UNITS_PER_TILE = NUM_TAPS * 4;        // (tap, 32-pixel chunk = two tile rows)
constexpr int CAND_BATCH = 16;                      // candidate tiles whose units are tested together
constexpr int UNIT_CAP = CAND_BATCH * UNITS_PER_TILE;   // 576
constexpr int CAP = 6144;                           // contribution slots resident in shared memory (8 B each)

struct BBox { short ymin, ymax, xmin, xmax; };      // inclusive; empty when ymin > ymax

struct Col2imParams {
  Geom g;
  const void* gcol;          // [B*H*W][9][Cin] element type T, rows of dead (pixel, tap) pairs are never read
  float* gx;                 // dense NHWC fp32 (B,H,W,Cin)
  const uint4* pair;         // [num_tiles][9][128] (y0 | x0 << 16, ly, lx, m); dead pair: ly < 0
  const BBox* unit_box;      // [num_tiles][36]
  const BBox* tile_box;      // [num_tiles]
};

// Sample position of one (pixel, tap): exactly the arithmetic of make_tap_geom (dcn_sampler.cuh), without the loads
// the col2im does not need.
struct PairPos {
  int y0, x0;
  float ly, lx;
  bool alive;
};
__device__ __forceinline__ PairPos pair_pos(const Geom& g, int b, int y, int x, int k) {
  PairPos t;
  t.alive = false; t.y0 = t.x0 = 0; t.ly = t.lx = 0.f;
  if (y >= g.H || x >= g.W) return t;
  const int64_t opix = (int64_t)b * g.offset.sb + (int64_t)y * g.offset.sh + (int64_t)x * g.offset.sw;
  float dy = ld_dyn(g.offset.ptr, opix + (int64_t)(2 * k) * g.offset.sc, g.om_dtype);
  float dx = ld_dyn(g.offset.ptr, opix + (int64_t)(2 * k + 1) * g.offset.sc, g.om_dtype);
  if (g.offset_bias) { dy += __ldg(g.offset_bias + 2 * k); dx += __ldg(g.offset_bias + 2 * k + 1); }
  const int ki = k / 3;
  const float py = (float)(y - 1 + ki) + dy;
  const float px = (float)(x - 1 + (k - 3 * ki)) + dx;
  if (!(py > -1.0f && px > -1.0f && py < (float)g.H && px < (float)g.W)) return t;
  t.alive = true;
  const float fy = floorf(py), fx = floorf(px);
  t.y0 = (int)fy; t.x0 = (int)fx;
  t.ly = py - fy; t.lx = px - fx;
  return t;
}
__device__ __forceinline__ float pair_mask(const Geom& g, int b, int y, int x, int k) {
  float m = ld_dyn(g.mask.ptr, (int64_t)b * g.mask.sb + (int64_t)k * g.mask.sc + (int64_t)y * g.mask.sh + (int64_t)x * g.mask.sw, g.om_dtype);
  if (g.mask_bias) m += __ldg(g.mask_bias + k);
  return g.mask_is_logit ? sigmoidf_accurate(m) : m;
}

// ---------------------------------------------------------------------------------------------- boxes
__global__ void __launch_bounds__(TILE_PX) dcn_bbox_kernel(const Geom g, uint4* __restrict__ pair, BBox* __restrict__ unit_box,
                                                           BBox* __restrict__ tile_box) {
  __shared__ int tb[4];
  const int tile = blockIdx.x, r = threadIdx.x, warp = r >> 5, lane = r & 31;
  int b, ty0, tx0;
  tile_coords(g, tile, b, ty0, tx0);
  if (r == 0) { tb[0] = 32767; tb[1] = -32768; tb[2] = 32767; tb[3] = -32768; }
  __syncthreads();
  const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
  int tymin = 32767, tymax = -32768, txmin = 32767, txmax = -32768;
  for (int k = 0; k < NUM_TAPS; ++k) {
    const PairPos t = pair_pos(g, b, y, x, k);
    int ymin = 32767, ymax = -32768, xmin = 32767, xmax = -32768;
    uint4 rec = make_uint4(0u, __float_as_uint(-1.f), 0u, 0u);
    if (t.alive) {   // alive => at least one corner row and one corner column inside the image
      ymin = max(t.y0, 0); ymax = min(t.y0 + 1, g.H - 1);
      xmin = max(t.x0, 0); xmax = min(t.x0 + 1, g.W - 1);
      rec = make_uint4(((uint32_t)t.y0 & 0xffffu) | ((uint32_t)t.x0 << 16), __float_as_uint(t.ly), __float_as_uint(t.lx),
                       __float_as_uint(pair_mask(g, b, y, x, k)));
    }
    pair[((size_t)tile * NUM_TAPS + k) * TILE_PX + r] = rec;
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
      ymin = min(ymin, __shfl_xor_sync(0xffffffffu, ymin, s)); ymax = max(ymax, __shfl_xor_sync(0xffffffffu, ymax, s));
      xmin = min(xmin, __shfl_xor_sync(0xffffffffu, xmin, s)); xmax = max(xmax, __shfl_xor_sync(0xffffffffu, xmax, s));
    }
    if (lane == 0) {
      BBox bb; bb.ymin = (short)ymin; bb.ymax = (short)ymax; bb.xmin = (short)xmin; bb.xmax = (short)xmax;
      unit_box[(size_t)tile * UNITS_PER_TILE + k * 4 + warp] = bb;
    }
    tymin = min(tymin, ymin); tymax = max(tymax, ymax); txmin = min(txmin, xmin); txmax = max(txmax, xmax);
  }
  if (lane == 0) {   // integer min / max: order independent
    atomicMin(&tb[0], tymin); atomicMax(&tb[1], tymax); atomicMin(&tb[2], txmin); atomicMax(&tb[3], txmax);
  }
  __syncthreads();
  if (r == 0) {
    BBox bb; bb.ymin = (short)tb[0]; bb.ymax = (short)tb[1]; bb.xmin = (short)tb[2]; bb.xmax = (short)tb[3];
    tile_box[tile] = bb;
  }
}

__device__ __forceinline__ bool box_meets(const BBox bb, int ty0, int tx0) {
  return bb.ymin <= bb.ymax && (int)bb.ymax >= ty0 && (int)bb.ymin <= ty0 + TILE_H - 1 && (int)bb.xmax >= tx0 &&
         (int)bb.xmin <= tx0 + TILE_W - 1;
}

// Ordered compaction over the CTA: thread i with flag set writes `value` at out[#flagged threads j < i]; returns the
// total.  Two __syncthreads; every thread must call it.
__device__ __forceinline__ int block_compact(bool flag, uint32_t value, uint32_t* out, int* wsum) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bal = __ballot_sync(0xffffffffu, flag);
  if (lane == 0) wsum[warp] = __popc(bal);
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int w = 0; w < C2I_WARPS; ++w) {
    const int c = wsum[w];
    before += w < warp ? c : 0;
    total += c;
  }
  if (flag) out[before + __popc(bal & ((1u << lane) - 1u))] = value;
  __syncthreads();
  return total;
}

// --------------------------------------------------------------------------------------------- col2im
template <typename T>
__global__ void __launch_bounds__(C2I_THREADS, 2) dcn_col2im_kernel(const Col2imParams p) {
  constexpr int kVec = 16 / (int)sizeof(T);          // channels per 16-byte piece: 8 bf16 / 4 fp32
  constexpr int kChunk = ROW_BYTES / (int)sizeof(T); // channels per 128-byte chunk
  constexpr int NSEG = C2I_WARPS * 4;                // (warp, corner) segments inside a destination pixel's slot range
  // pass 0: corners per (warp, corner index, destination pixel); then, in place, the next sorted slot of that triple
  __shared__ uint32_t slot[C2I_WARPS][4][TILE_PX];
  __shared__ uint32_t start[TILE_PX + 1];
  __shared__ uint32_t cand[C2I_THREADS];
  __shared__ uint32_t units[UNIT_CAP];
  __shared__ int wsum[C2I_WARPS];
  extern __shared__ uint2 contrib[];                 // [CAP] (gcol row = pixel * 9 + tap, m * w_corner); dynamic: 48 KB

  const Geom& g = p.g;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x;
  int b, ty0, tx0;
  tile_coords(g, tile, b, ty0, tx0);
  const int tpi = g.tiles_x * g.tiles_y;
  const BBox* tboxes = p.tile_box + (size_t)b * tpi;
  const BBox* uboxes = p.unit_box + (size_t)b * tpi * UNITS_PER_TILE;
  const uint4* pairs = p.pair + (size_t)b * tpi * NUM_TAPS * TILE_PX;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int hmax = min(TILE_H, g.H - ty0), wmax = min(TILE_W, g.W - tx0);    // the part of D inside the image
  const uint32_t img_row0 = (uint32_t)(b * g.H) * (uint32_t)g.W;

  uint32_t total = 0, win_lo = 0;
  // pass 0 counts; pass 1 places the sorted slots [0, CAP) and accumulates them.  A destination tile with more than CAP
  // corners repeats (count, place window w, accumulate) for w = 1, 2, ...: rare (offset fields collapsing onto a spot).
  for (int pass = 0;; ++pass) {
    const bool counting = (pass & 1) == 0;   // even passes count, odd passes place one window and accumulate it
    if (counting) {
      for (int i = tid; i < NSEG * TILE_PX; i += C2I_THREADS) (&slot[0][0][0])[i] = 0u;
      __syncthreads();
    }
    for (int tbase = 0; tbase < tpi; tbase += C2I_THREADS) {
      const int tc_ = tbase + tid;
      const int ncand = block_compact(tc_ < tpi && box_meets(tboxes[tc_ < tpi ? tc_ : 0], ty0, tx0), (uint32_t)tc_, cand, wsum);
      for (int cb = 0; cb < ncand; cb += CAND_BATCH) {
        const int nu = min(CAND_BATCH, ncand - cb) * UNITS_PER_TILE;
        int nunits = 0;
#pragma unroll 1
        for (int rr = 0; rr < (UNIT_CAP + C2I_THREADS - 1) / C2I_THREADS; ++rr) {
          const int i = rr * C2I_THREADS + tid;
          bool hit = false;
          uint32_t code = 0;
          if (i < nu) {
            const uint32_t t = cand[cb + i / UNITS_PER_TILE], kc = (uint32_t)(i % UNITS_PER_TILE);
            hit = box_meets(uboxes[(size_t)t * UNITS_PER_TILE + kc], ty0, tx0);
            const uint32_t sty = t / (uint32_t)g.tiles_x, stx = t - sty * (uint32_t)g.tiles_x;   // one division per unit, here
            code = (sty << 20) | (stx << 6) | kc;      // 12 + 14 + 6 bits (H, W <= 32767)
          }
          nunits += block_compact(hit, code, units + nunits, wsum);
        }
        // ---- every warp walks its share of the units; lane = source pixel
        // (the record of the warp's NEXT unit is requested before the current one is walked: one L2 latency per unit
        //  would otherwise be exposed, there is nothing else for the warp to do)
        auto fetch = [&](int ui, uint32_t& code, uint4& rec) {
          code = units[ui];
          const uint32_t st = (code >> 20) * (uint32_t)g.tiles_x + ((code >> 6) & 0x3fffu);
          rec = __ldg(pairs + ((size_t)st * NUM_TAPS + ((code & 63u) >> 2)) * TILE_PX + (code & 3u) * 32 + lane);
        };
        uint32_t code_n = 0;
        uint4 rec_n = make_uint4(0u, 0u, 0u, 0u);
        if (warp < nunits) fetch(warp, code_n, rec_n);
        for (int ui = warp; ui < nunits; ui += C2I_WARPS) {
          const uint32_t code = code_n;
          const uint4 rec = rec_n;
          if (ui + C2I_WARPS < nunits) fetch(ui + C2I_WARPS, code_n, rec_n);
          const float ly = __uint_as_float(rec.y), lx = __uint_as_float(rec.z);
          // top-left corner relative to D; the four corners are (ry, rx) + {0,1}^2
          const int ry = (int)(short)(rec.x & 0xffffu) - ty0, rx = (int)(short)(rec.x >> 16) - tx0;
          const bool near = ly >= 0.f && ry >= -1 && ry < hmax && rx >= -1 && rx < wmax;   // alive and some corner may be in D
          const uint32_t act = __ballot_sync(0xffffffffu, near);
          if (act == 0u) continue;
          // lanes with the same top-left corner share all four destinations: ONE match per unit
          uint32_t same = 0u;
          if (near) same = __match_any_sync(act, (ry + 1) * 32 + (rx + 1));
          const float m = __uint_as_float(rec.w);
          const uint32_t k = (code & 63u) >> 2, r = (code & 3u) * 32 + (uint32_t)lane;
          const uint32_t sy = (code >> 20) * TILE_H + (r >> 4), sx = ((code >> 6) & 0x3fffu) * TILE_W + (r & (TILE_W - 1));
          const uint32_t src = (img_row0 + sy * (uint32_t)g.W + sx) * NUM_TAPS + k;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int cy = ry + (j >> 1), cx = rx + (j & 1);
            const float w = ((j >> 1) ? ly : 1.0f - ly) * ((j & 1) ? lx : 1.0f - lx);
            const bool ok = near && (unsigned)cy < (unsigned)hmax && (unsigned)cx < (unsigned)wmax && w != 0.f;
            const uint32_t okb = __ballot_sync(0xffffffffu, ok);
            if (okb == 0u) continue;
            if (ok) {
              // same destination pixel <=> same top-left corner (for a fixed corner index j)
              const uint32_t grp = same & okb;
              const uint32_t rank = (uint32_t)__popc(grp & lt_mask);
              uint32_t* sl = &slot[warp][j][cy * TILE_W + cx];
              if (counting) {
                if (rank == 0) *sl += (uint32_t)__popc(grp);
              } else {
                const uint32_t pos = *sl + rank;
                if (pos >= win_lo && pos < win_lo + CAP) contrib[pos - win_lo] = make_uint2(src, __float_as_uint(m * w));
                __syncwarp(grp);
                if (rank == 0) *sl += (uint32_t)__popc(grp);
              }
            }
          }
          __syncwarp();   // the slot cursors written by this unit's group leaders are read by the warp's next unit
        }
        __syncthreads();   // units[] / cand[] are rewritten by the next batch
      }
    }
    __syncthreads();
    if (counting) {
      // ---- slot ranges: start[q] = corners landing on pixels < q; inside a pixel's range (warp, corner index) order
      if (tid < TILE_PX) {
        uint32_t s = 0;
#pragma unroll
        for (int w = 0; w < NSEG; ++w) s += (&slot[0][0][0])[w * TILE_PX + tid];
        start[tid + 1] = s;
      }
      __syncthreads();
      if (warp == 0) {   // inclusive scan of 128 totals: 4 per lane + shuffle scan
        uint32_t v[4], s = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { s += start[lane * 4 + i + 1]; v[i] = s; }
        uint32_t incl = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
          if (lane >= d) incl += o;
        }
        const uint32_t excl = incl - s;
#pragma unroll
        for (int i = 0; i < 4; ++i) start[lane * 4 + i + 1] = excl + v[i];
        if (lane == 0) start[0] = 0u;
      }
      __syncthreads();
      total = start[TILE_PX];
      // counts -> first slot of every (warp, corner, pixel) segment, in place
      if (tid < TILE_PX) {
        uint32_t s = start[tid];
#pragma unroll
        for (int w = 0; w < NSEG; ++w) {
          uint32_t* c = &(&slot[0][0][0])[w * TILE_PX + tid];
          const uint32_t n = *c;
          *c = s;
          s += n;
        }
      }
      __syncthreads();
    } else {
      // ---- accumulate the resident window: 8 lanes per destination pixel, 16 bytes of a gcol row per lane
      const int grp = tid >> 3, l8 = tid & 7;
      const uint8_t* gcb = reinterpret_cast<const uint8_t*>(p.gcol);
      const size_t rowb = (size_t)g.Cin * sizeof(T);
      const int nchunk = g.Cin / kChunk;
      for (int q = grp; q < TILE_PX; q += C2I_THREADS / 8) {
        const int y = ty0 + (q >> 4), x = tx0 + (q & (TILE_W - 1));
        if (y >= g.H || x >= g.W) continue;
        const uint32_t s0 = start[q], s1 = start[q + 1];
        const bool empty = s0 == s1;
        const uint32_t e0 = s0 < win_lo ? win_lo : s0, e1 = s1 > win_lo + CAP ? win_lo + CAP : s1;
        if (empty ? pass != 1 : e0 >= e1) continue;   // nothing of q in this window (a pixel nobody samples is zero-filled in window 0)
        const bool first = empty || s0 >= win_lo;     // this window holds the head of q's range: store; later windows add
        float* dst = p.gx + ((size_t)((b * g.H + y) * g.W + x)) * g.Cin + l8 * kVec;
        for (int cc = 0; cc < nchunk; ++cc) {
          float acc[kVec];
#pragma unroll
          for (int i = 0; i < kVec; ++i) acc[i] = 0.f;
          const size_t coff = (size_t)cc * ROW_BYTES + (size_t)l8 * 16;
          uint32_t e = e0;
          for (; e + 4 <= e1; e += 4) {        // four independent 16-byte loads in flight per lane
            uint2 ce[4];
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) ce[u] = contrib[e + u - win_lo];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = ldg16(gcb + (size_t)ce[u].x * rowb + coff);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float w = __uint_as_float(ce[u].y);
              const uint32_t* vw = &v[u].x;
              if constexpr (sizeof(T) == 4) {
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[i] = fmaf(w, __uint_as_float(vw[i]), acc[i]);
              } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  acc[2 * i] = fmaf(w, __uint_as_float(vw[i] << 16), acc[2 * i]);
                  acc[2 * i + 1] = fmaf(w, __uint_as_float(vw[i] & 0xffff0000u), acc[2 * i + 1]);
                }
              }
            }
          }
          for (; e < e1; ++e) {
            const uint2 ce = contrib[e - win_lo];
            const uint4 v = ldg16(gcb + (size_t)ce.x * rowb + coff);
            const float w = __uint_as_float(ce.y);
            const uint32_t* vw = &v.x;
            if constexpr (sizeof(T) == 4) {
#pragma unroll
              for (int i = 0; i < 4; ++i) acc[i] = fmaf(w, __uint_as_float(vw[i]), acc[i]);
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                acc[2 * i] = fmaf(w, __uint_as_float(vw[i] << 16), acc[2 * i]);
                acc[2 * i + 1] = fmaf(w, __uint_as_float(vw[i] & 0xffff0000u), acc[2 * i + 1]);
              }
            }
          }
          float* d = dst + (size_t)cc * kChunk;
          if (!first) {     // a later window of a long range: add to what the earlier window stored (same thread)
#pragma unroll
            for (int i = 0; i < kVec; i += 4) {
              const float4 o = *reinterpret_cast<const float4*>(d + i);
              acc[i] += o.x; acc[i + 1] += o.y; acc[i + 2] += o.z; acc[i + 3] += o.w;
            }
          }
#pragma unroll
          for (int i = 0; i < kVec; i += 4) *reinterpret_cast<float4*>(d + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
        }
      }
      win_lo += CAP;
      if (win_lo >= total) break;
    }
    __syncthreads();
  }
}

bool dense_nhwc_t(const cpoly_tensor4& t, int C, int H, int W) {
  return t.sc == 1 && t.sw == C && t.sh == (int64_t)W * C && t.sb == (int64_t)H * W * C;
}

int tiles_of(const cpoly_dcn_fwd_args& f) { return f.B * ((f.W + TILE_W - 1) / TILE_W) * ((f.H + TILE_H - 1) / TILE_H); }

}  // namespace

// ---- host side -------------------------------------------------------------------------------------------------
bool dcn_col2im_eligible(const cpoly_dcn_bwd_args& a) {
  const cpoly_dcn_fwd_args& f = a.f;
  const int esize = f.dtype == CPOLY_BF16 ? 2 : 4;
  if (f.kh != 3 || f.kw != 3 || f.stride_h != 1 || f.stride_w != 1 || f.pad_h != 1 || f.pad_w != 1 || f.dil_h != 1 || f.dil_w != 1 ||
      f.deformable_groups != 1)
    return false;
  if (f.Cin % (ROW_BYTES / esize) != 0 || f.H > 32767 || f.W > 32767) return false;
  if ((int64_t)f.B * f.H * f.W * NUM_TAPS >= (1LL << 32)) return false;                  // 32-bit gcol row index
  if (!a.gx.ptr || !dense_nhwc_t(a.gx, f.Cin, f.H, f.W) || (reinterpret_cast<uintptr_t>(a.gx.ptr) & 15)) return false;
  return true;
}

size_t dcn_col2im_gcol_bytes(const cpoly_dcn_fwd_args& f) {
  return ((size_t)f.B * f.H * f.W * NUM_TAPS * f.Cin * (f.dtype == CPOLY_BF16 ? 2 : 4) + 255) / 256 * 256;
}

static size_t pair_bytes(const cpoly_dcn_fwd_args& f) { return (size_t)tiles_of(f) * NUM_TAPS * TILE_PX * sizeof(uint4); }

size_t dcn_col2im_workspace(const cpoly_dcn_bwd_args& a) {
  return dcn_col2im_gcol_bytes(a.f) + pair_bytes(a.f) + (size_t)tiles_of(a.f) * (UNITS_PER_TILE + 1) * sizeof(BBox) + 256;
}

// gcol lives at the START of `workspace` (dcn_col2im_gcol_bytes), the pair records and the boxes behind it.
int dcn_col2im_launch(const cpoly_dcn_bwd_args& a, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  const cpoly_dcn_fwd_args& f = a.f;
  CPOLY_REQUIRE(workspace && workspace_bytes >= dcn_col2im_workspace(a) - 256 && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0,
                CPOLY_EWORKSPACE, "dcn col2im workspace too small or misaligned");
  Col2imParams p;
  p.g.x = f.x.ptr; p.g.offset = f.offset; p.g.mask = f.mask; p.g.offset_bias = f.offset_bias; p.g.mask_bias = f.mask_bias;
  p.g.B = f.B; p.g.Cin = f.Cin; p.g.Cout = f.Cout; p.g.H = f.H; p.g.W = f.W;
  p.g.om_dtype = f.om_dtype; p.g.mask_is_logit = f.mask_is_logit;
  p.g.tiles_x = (f.W + TILE_W - 1) / TILE_W; p.g.tiles_y = (f.H + TILE_H - 1) / TILE_H;
  p.g.num_tiles = f.B * p.g.tiles_x * p.g.tiles_y;
  if (p.g.num_tiles == 0) return CPOLY_OK;
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  p.gcol = ws;
  uint4* pr = reinterpret_cast<uint4*>(ws + dcn_col2im_gcol_bytes(f));
  p.pair = pr;
  BBox* ub = reinterpret_cast<BBox*>(ws + dcn_col2im_gcol_bytes(f) + pair_bytes(f));
  BBox* tb = ub + (size_t)p.g.num_tiles * UNITS_PER_TILE;
  p.unit_box = ub; p.tile_box = tb;
  p.gx = reinterpret_cast<float*>(a.gx.ptr);
  dcn_bbox_kernel<<<p.g.num_tiles, TILE_PX, 0, stream>>>(p.g, pr, ub, tb);
  if (int rc = check_launch("dcn_bbox_kernel")) return rc;
  constexpr int dyn = CAP * (int)sizeof(uint2);
  if (f.dtype == CPOLY_BF16) {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_col2im_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    dcn_col2im_kernel<__nv_bfloat16><<<p.g.num_tiles, C2I_THREADS, dyn, stream>>>(p);
  } else {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_col2im_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    dcn_col2im_kernel<float><<<p.g.num_tiles, C2I_THREADS, dyn, stream>>>(p);
  }
  return check_launch("dcn_col2im_kernel");
}

}  // namespace cpoly
