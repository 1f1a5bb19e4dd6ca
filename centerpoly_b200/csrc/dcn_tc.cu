// DCNv2 forward on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// Replaces the DCNv2 extension's `dcn_v2_forward` for the 16 DeformConv layers of the DLA-34 neck
// (reference call site src/lib/models/networks/pose_dla_dcn.py:354-359; extension = missing submodule
// src/lib/models/networks/DCNv2).  Upstream materialises columns[9*Cin, H*W] in HBM (9x the input)
// and calls cuBLAS SGEMM; here the column tile never leaves the SM.
//
//   out[px, o] = act(scale[o] * sum_{tap, c} A[px, (tap, c)] * W[o, c, tap] + bias[o]),  A = mask * bilinear(x, p_tap)
//
// Persistent, warp-specialised CTA (one per SM).  A work item is (128-pixel tile, tap group): tiles are
// 8 rows x 16 columns of the image so that neighbouring pixels share bilinear corners in L1; the 9 taps are
// processed by one item (big maps) or split over 3 / 9 items whose fp32 partial sums are combined by a
// second, fixed-order kernel (small maps, which could not otherwise fill 148 SMs).
//
//   warps 0-3   epilogue   tcgen05.ld accumulator rows -> scale/bias/ReLU -> bf16/fp32 NHWC stores (or fp32 partials)
//   warp  4     B loader   one thread: cp.async.bulk (TMA engine, UBLKCP) of pre-swizzled weight chunks
//   warp  5     MMA        TMEM alloc; one thread issues tcgen05.mma (M=128, N=Cout, K=16 bf16 / 8 tf32)
//   warps 6..   samplers   per item: stage (corner offsets, mask*bilinear weights) of the item's taps in smem,
//                          then per (tap, 128-byte channel chunk) gather 4 corners with 16-byte loads, blend
//                          (packed bf16x2 FMAs / fp32 FMAs) and write the A tile straight into the 128B-swizzled
//                          K-major UMMA layout
//
// Pipelines: smem ring (fullA/fullB/empty mbarriers) between samplers+loader and the MMA thread; two TMEM
// accumulators (tmem_full/tmem_empty) between the MMA thread and the epilogue, so the epilogue of item t
// overlaps the MMAs of item t+1.  Every mbarrier wait is a hardware-suspended try_wait with a clock-based
// timeout that traps instead of hanging the GPU.
#include <stdlib.h>

#include "tc_common.cuh"

namespace cpoly {
namespace {

using namespace tc;

constexpr int TILE_H = 8, TILE_W = 16, TILE_PX = TILE_H * TILE_W;  // 128 = UMMA M
constexpr int NUM_TAPS = 9;
constexpr int A_STAGE_BYTES = TILE_PX * ROW_BYTES;  // 16 KB
constexpr int MAX_STAGES = 6;
constexpr int EPI_WARPS = 4;
constexpr int SAMPLER_WARP0 = 6;
constexpr int SAMPLER_WARPS = 16;
constexpr int NS = SAMPLER_WARPS * 32;         // sampler threads
constexpr int NTHREADS = (SAMPLER_WARP0 + SAMPLER_WARPS) * 32;
constexpr int ROWS_PER_PASS = NS / 8;          // 8 lanes x 16 B cover one 128-byte row
constexpr int PASSES = TILE_PX / ROWS_PER_PASS;
constexpr int MAX_ITEMS = (NUM_TAPS * TILE_PX + NS - 1) / NS;  // (pixel, tap) parameter items per sampler thread

// shared-memory map (dynamic, 1024-byte aligned)
constexpr int SMEM_BAR_OFF = 0;        // mbarriers + tmem pointer
constexpr int SMEM_VEC_OFF = 1024;     // scale[256] + bias[256] floats
constexpr int SMEM_POFF_OFF = 3072;    // corner offsets: NUM_TAPS * TILE_PX * 16 B
constexpr int SMEM_PW_OFF = SMEM_POFF_OFF + NUM_TAPS * TILE_PX * 16;  // corner weights: same size
constexpr int SMEM_STAGE_OFF = ((SMEM_PW_OFF + NUM_TAPS * TILE_PX * 16 + 1023) / 1024) * 1024;
constexpr int SMEM_LIMIT = 226 * 1024;  // 227 KB per CTA minus the alignment slack

struct TcParams {
  const void* x;
  cpoly_tensor4 offset, mask;
  const void* wpacked;
  const float* bias;
  const float* scale;
  const float* offset_bias;
  const float* mask_bias;
  void* y;
  float* partial;  // [split][B*H*W][Cout] fp32 when split > 1
  int B, Cin, Cout, H, W;
  int om_dtype, mask_is_logit, relu;
  int tiles_x, tiles_y, num_tiles, num_items;
  int split, taps_per_item;
  int nchunk;      // channel chunks per tap
  int stages, b_stage_bytes, stage_bytes;
  int acc_cols, tmem_cols;
  int om_esize;    // element size of the fused 32-channel offset/mask tensor (kOmFast kernels)
  int late_issue;  // sampler: corner loads of the next chunk issued after (1) or before (0) the proxy fence
};

// ---- blend of the four corner vectors (16 B each) into one 16-B piece of the A row ----------------
// bf16: weights arrive as bf16x2 pairs (w, w); products and sums are packed bf16x2 FMAs — half the issue slots of
// unpack + fp32 FMA + repack, and the A operand is rounded to bf16 anyway (tolerance 1e-2 in bf16 mode).
// fp32: fp32 FMAs, result rounded to tf32 (round-to-nearest) for kind::tf32.
template <typename T> struct Blend;
template <> struct Blend<__nv_bfloat16> {
  static constexpr int kVec = 8;  // channels per 16 bytes
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
  static constexpr int kVec = 4;
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

template <typename T> struct OutConv;
template <> struct OutConv<__nv_bfloat16> {
  // 16 fp32 values -> 32 bytes
  __device__ static __forceinline__ void store(__nv_bfloat16* dst, const float (&v)[16]) {
    uint4 a, b;
    a.x = pack_bf16x2(v[0], v[1]);  a.y = pack_bf16x2(v[2], v[3]);   a.z = pack_bf16x2(v[4], v[5]);   a.w = pack_bf16x2(v[6], v[7]);
    b.x = pack_bf16x2(v[8], v[9]);  b.y = pack_bf16x2(v[10], v[11]); b.z = pack_bf16x2(v[12], v[13]); b.w = pack_bf16x2(v[14], v[15]);
    reinterpret_cast<uint4*>(dst)[0] = a;
    reinterpret_cast<uint4*>(dst)[1] = b;
  }
};
template <> struct OutConv<float> {
  __device__ static __forceinline__ void store(float* dst, const float (&v)[16]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) reinterpret_cast<float4*>(dst)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  }
};

__device__ __forceinline__ void item_coords(const TcParams& p, int item, int& b, int& ty0, int& tx0, int& tap0) {
  const int tile = item / p.split;
  tap0 = (item - tile * p.split) * p.taps_per_item;
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  b = tile / tiles_per_img;
  const int trem = tile - b * tiles_per_img;
  const int tyi = trem / p.tiles_x;
  ty0 = tyi * TILE_H;
  tx0 = (trem - tyi * p.tiles_x) * TILE_W;
}

// ---------------------------------------------------------------------------------------- kernel
// kOmFast: offsets and mask logits are channels 0..17 / 18..26 of ONE dense 32-channel NHWC tensor (what
// cpoly_offset_conv_forward and the padded cuDNN producer write).  The item's 128 x 32 values are then fetched with
// coalesced 16-byte loads one item ahead (8 registers), transposed through shared memory (odd word pitch: conflict-free)
// and the parameters computed from there - instead of three strided 2/4-byte loads per (pixel, tap), which cost ~180 of
// the ~1500 LSU wavefronts of every (tile, tap) step (ncu, profiles/r01m_dcn_fwd_tc_64x64_raw.txt).
template <typename T, bool kOmFast>
__global__ void __maxnreg__(80) dcn_fwd_tc_kernel(const TcParams p) {
  constexpr bool kTF32 = sizeof(T) == 4;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operand tiles need 1024-byte alignment: align the window ourselves (+1 KB is requested)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_fullA = smem_base + SMEM_BAR_OFF;                 // [MAX_STAGES]
  const uint32_t bar_fullB = bar_fullA + 8 * MAX_STAGES;               // [MAX_STAGES]
  const uint32_t bar_empty = bar_fullB + 8 * MAX_STAGES;               // [MAX_STAGES]
  const uint32_t bar_tfull = bar_empty + 8 * MAX_STAGES;               // [2]
  const uint32_t bar_tempty = bar_tfull + 16;                          // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SMEM_BAR_OFF + 8 * (3 * MAX_STAGES + 4));
  float* scale_s = reinterpret_cast<float*>(smem + SMEM_VEC_OFF);
  float* bias_s = scale_s + 256;
  const uint32_t poff_s = smem_base + SMEM_POFF_OFF, pw_s = smem_base + SMEM_PW_OFF;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- one-time setup
  if (warp == 4 && lane == 0) {
    for (int s = 0; s < MAX_STAGES; ++s) {
      mbar_init(bar_fullA + 8 * s, SAMPLER_WARPS);
      mbar_init(bar_fullB + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar_tfull + 8 * a, 1);
      mbar_init(bar_tempty + 8 * a, EPI_WARPS * 32);
    }
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(smem_u32(tmem_ptr_smem), (uint32_t)p.tmem_cols);
  for (int i = threadIdx.x; i < p.Cout; i += NTHREADS) {
    scale_s[i] = p.scale ? __ldg(p.scale + i) : 1.f;
    bias_s[i] = p.bias ? __ldg(p.bias + i) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int chunks_per_item = p.taps_per_item * p.nchunk;
  const uint32_t nstages = (uint32_t)p.stages;
  // A CTA walks a CONTIGUOUS range of items: consecutive tiles of a tile row share 43 % of their bilinear footprint
  // (and the tap groups of one tile all of it), so the halo of the next item is already in this SM's L1 - with the
  // round-robin deal of round 1 every item started cold (L1 sector hit rate 83 %, long-scoreboard stalls dominant).
  const int item_beg = (int)(((long long)p.num_items * blockIdx.x) / gridDim.x);
  const int item_end = (int)(((long long)p.num_items * (blockIdx.x + 1)) / gridDim.x);

  if (warp >= SAMPLER_WARP0) {
    // =========================================================================== samplers
    const int st = threadIdx.x - SAMPLER_WARP0 * 32;
    const int row_in_pass = st >> 3, cg = st & 7;
    const int n_param = p.taps_per_item * TILE_PX;
    float raw_dy[MAX_ITEMS], raw_dx[MAX_ITEMS], raw_m[MAX_ITEMS];
    uint4 omv[2];   // kOmFast: this thread's 16-byte pieces of the NEXT item's offset/mask tile
    // raw offsets / mask of one item -> registers (issued one item ahead so the latency hides behind sampling)
    auto fetch_raw = [&](int item) {
      int b, ty0, tx0, tap0;
      item_coords(p, item, b, ty0, tx0, tap0);
      if constexpr (kOmFast) {
        const int ppp = p.om_esize * 2;                  // 16-byte pieces per pixel: 32 channels * esize / 16
        const uint8_t* base = reinterpret_cast<const uint8_t*>(p.offset.ptr);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int idx = st + i * NS;
          const int r = idx / ppp, j = idx - r * ppp;
          const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
          omv[i] = make_uint4(0u, 0u, 0u, 0u);
          if (r < TILE_PX && y < p.H && x < p.W)
            omv[i] = ldg16(base + ((size_t)((b * p.H + y) * p.W + x) * 32) * p.om_esize + j * 16);
        }
      } else {
#pragma unroll
        for (int i = 0; i < MAX_ITEMS; ++i) {
          const int e = st + i * NS;
          const int k = tap0 + (e >> 7), r = e & (TILE_PX - 1);
          const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
          raw_dy[i] = raw_dx[i] = raw_m[i] = 0.f;
          if (e < n_param && y < p.H && x < p.W) {
            const int64_t opix = (int64_t)b * p.offset.sb + (int64_t)y * p.offset.sh + (int64_t)x * p.offset.sw;
            raw_dy[i] = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k) * p.offset.sc, p.om_dtype);
            raw_dx[i] = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k + 1) * p.offset.sc, p.om_dtype);
            raw_m[i] = ld_dyn(p.mask.ptr, (int64_t)b * p.mask.sb + (int64_t)k * p.mask.sc + (int64_t)y * p.mask.sh + (int64_t)x * p.mask.sw, p.om_dtype);
            if (p.offset_bias) { raw_dy[i] += __ldg(p.offset_bias + 2 * k); raw_dx[i] += __ldg(p.offset_bias + 2 * k + 1); }
            if (p.mask_bias) raw_m[i] += __ldg(p.mask_bias + k);
          }
        }
      }
    };
    // kOmFast: registers -> shared tile [pixel][32 channels] with an odd word pitch (33 words fp32 / 17 words bf16), then
    // every thread picks the (dy, dx, mask) of its own (pixel, tap) entries.  The tile aliases the weight-parameter
    // array, which is dead between the first barrier of an item and the parameter stores.
    auto stage_raw = [&](int tap0) {
      const int ppp = p.om_esize * 2;
      const uint32_t pitch = p.om_esize == 4 ? 132u : 68u;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int idx = st + i * NS;
        const int r = idx / ppp, j = idx - r * ppp;
        if (r < TILE_PX) {
          const uint32_t a = pw_s + (uint32_t)r * pitch + (uint32_t)j * 16u;
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(omv[i].x) : "memory");
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(a + 4u), "r"(omv[i].y) : "memory");
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(a + 8u), "r"(omv[i].z) : "memory");
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(a + 12u), "r"(omv[i].w) : "memory");
        }
      }
      named_bar_sync(1, NS);
#pragma unroll
      for (int i = 0; i < MAX_ITEMS; ++i) {
        const int e = st + i * NS;
        raw_dy[i] = raw_dx[i] = raw_m[i] = 0.f;
        if (e < n_param) {
          const int k = tap0 + (e >> 7), r = e & (TILE_PX - 1);
          const uint32_t a = pw_s + (uint32_t)r * pitch;
          uint32_t u0, u1;
          if (p.om_esize == 4) {
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u0) : "r"(a + (uint32_t)(8 * k)));
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u1) : "r"(a + (uint32_t)(8 * k + 4)));
            raw_dy[i] = __uint_as_float(u0); raw_dx[i] = __uint_as_float(u1);
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u0) : "r"(a + (uint32_t)(72 + 4 * k)));
            raw_m[i] = __uint_as_float(u0);
          } else {
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u0) : "r"(a + (uint32_t)(4 * k)));          // (dy, dx) bf16 pair
            raw_dy[i] = __uint_as_float(u0 << 16); raw_dx[i] = __uint_as_float(u0 & 0xffff0000u);
            asm volatile("ld.shared.u16 %0, [%1];" : "=r"(u1) : "r"(a + (uint32_t)(36 + 2 * k)));
            raw_m[i] = __uint_as_float(u1 << 16);
          }
          if (p.offset_bias) { raw_dy[i] += __ldg(p.offset_bias + 2 * k); raw_dx[i] += __ldg(p.offset_bias + 2 * k + 1); }
          if (p.mask_bias) raw_m[i] += __ldg(p.mask_bias + k);
        }
      }
      named_bar_sync(1, NS);   // the tile is dead: its bytes become weight parameters again
    };
    RingCursor rc;
    if (item_beg < item_end) fetch_raw(item_beg);
    for (int item = item_beg; item < item_end; ++item) {
      int b, ty0, tx0, tap0;
      item_coords(p, item, b, ty0, tx0, tap0);
      const int img_base = b * p.H * p.W * p.Cin;   // element offset; bytes < 2^32 (checked on the host)
      named_bar_sync(1, NS);  // everyone finished reading the previous item's params
      if constexpr (kOmFast) stage_raw(tap0);
#pragma unroll
      for (int i = 0; i < MAX_ITEMS; ++i) {
        const int e = st + i * NS;
        if (e >= n_param) continue;
        const int kl = e >> 7, k = tap0 + kl, r = e & (TILE_PX - 1);
        const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
        uint4 po = make_uint4(0u, 0u, 0u, 0u), pw = make_uint4(0u, 0u, 0u, 0u);  // dead corner: offset 0, weight 0
        if (y < p.H && x < p.W) {
          const float m = p.mask_is_logit ? sigmoidf_accurate(raw_m[i]) : raw_m[i];
          const int ki = k / 3;
          const float py = (float)(y - 1 + ki) + raw_dy[i];
          const float px = (float)(x - 1 + (k - 3 * ki)) + raw_dx[i];
          if (py > -1.0f && px > -1.0f && py < (float)p.H && px < (float)p.W) {
            const float fy = floorf(py), fx = floorf(px);
            const int y0 = (int)fy, x0 = (int)fx;
            const float ly = py - fy, lx = px - fx, hy = 1.0f - ly, hx = 1.0f - lx;
            const bool vy0 = y0 >= 0, vy1 = y0 + 1 <= p.H - 1, vx0 = x0 >= 0, vx1 = x0 + 1 <= p.W - 1;
            const uint32_t pixb = (uint32_t)p.Cin * (uint32_t)sizeof(T), rowb = (uint32_t)p.W * pixb;
            const uint32_t o00 = (uint32_t)img_base * (uint32_t)sizeof(T) + (uint32_t)(y0 * p.W + x0) * pixb;  // may wrap for dead corners
            if (vy0 && vx0) { po.x = o00;               pw.x = Blend<T>::weight(hy * hx * m); }
            if (vy0 && vx1) { po.y = o00 + pixb;        pw.y = Blend<T>::weight(hy * lx * m); }
            if (vy1 && vx0) { po.z = o00 + rowb;        pw.z = Blend<T>::weight(ly * hx * m); }
            if (vy1 && vx1) { po.w = o00 + rowb + pixb; pw.w = Blend<T>::weight(ly * lx * m); }
          }
        }
        sts16(poff_s + (uint32_t)(kl * TILE_PX + r) * 16u, po);
        sts16(pw_s + (uint32_t)(kl * TILE_PX + r) * 16u, pw);
      }
      named_bar_sync(1, NS);
      if (item + 1 < item_end) fetch_raw(item + 1);
      // Rolling software pipeline over the item's chunks: the corner loads of chunk q+1 are issued while chunk q
      // is blended, so every sampler warp keeps one pass (4 x 16 B per thread) in flight at all times instead of
      // alternating between a load phase and a blend phase.
      uint4 c[PASSES][4], pw[PASSES];
      const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.x);
      auto issue = [&](uint32_t pidx, uint32_t cbyte, int pass) {   // params + corner loads of one (chunk, pass)
        const uint32_t idx = pidx + (uint32_t)(pass * ROWS_PER_PASS) * 16u;
        const uint4 po = lds16(poff_s + idx);   // BYTE offsets of the four corner vectors
        pw[pass] = lds16(pw_s + idx);
        c[pass][0] = ldg16(xb + (po.x + cbyte));
        c[pass][1] = ldg16(xb + (po.y + cbyte));
        c[pass][2] = ldg16(xb + (po.z + cbyte));
        c[pass][3] = ldg16(xb + (po.w + cbyte));
      };
      // (tap, channel chunk) cursor of the NEXT chunk to issue: parameter row index and channel byte offset
      const uint32_t cbyte0 = (uint32_t)(cg * 16);
      uint32_t n_pidx = (uint32_t)row_in_pass * 16u, n_cbyte = cbyte0;
      int n_cc = 0;
      auto next_chunk = [&]() {
        n_cbyte += ROW_BYTES;
        if (++n_cc == p.nchunk) { n_cc = 0; n_cbyte = cbyte0; n_pidx += TILE_PX * 16u; }
      };
#pragma unroll
      for (int pass = 0; pass < PASSES; ++pass) issue(n_pidx, n_cbyte, pass);
      next_chunk();
      for (int q = 0; q < chunks_per_item; ++q) {
        mbar_wait(bar_empty + 8 * rc.s, rc.ph ^ 1u);
        const uint32_t a_base = smem_base + SMEM_STAGE_OFF + rc.s * p.stage_bytes;
        const bool more = q + 1 < chunks_per_item;
        const bool late = p.late_issue != 0;
#pragma unroll
        for (int pass = 0; pass < PASSES; ++pass) {
          const int r = pass * ROWS_PER_PASS + row_in_pass;
          sts16(a_base + r * ROW_BYTES + ((cg ^ (r & 7)) << 4), Blend<T>::run(c[pass], pw[pass]));
          if (more && !late) issue(n_pidx, n_cbyte, pass);
        }
        fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core (async proxy)
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_fullA + 8 * rc.s);
        // (fence.proxy.async is MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC in SASS: it waits for every load this thread has in
        //  flight, so loads issued before it are drained at every step; issuing them after it keeps them in flight
        //  across the wait for the next free stage)
        if (more && late) {
#pragma unroll
          for (int pass = 0; pass < PASSES; ++pass) issue(n_pidx, n_cbyte, pass);
        }
        next_chunk();
        rc.advance(nstages);
      }
    }
  } else if (warp == 4) {
    // =========================================================================== weight loader
    // (whole warp walks the ring, one elected lane issues: see elect_one() in tc_common.cuh)
    RingCursor rc;
    const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.wpacked);
    for (int item = item_beg; item < item_end; ++item) {
      const int q0 = (item % p.split) * chunks_per_item;
      for (int q = 0; q < chunks_per_item; ++q) {
        mbar_wait(bar_empty + 8 * rc.s, rc.ph ^ 1u, 100);
        if (elect_one()) {
          mbar_arrive_expect_tx(bar_fullB + 8 * rc.s, (uint32_t)p.b_stage_bytes);
          bulk_copy_g2s(smem_base + SMEM_STAGE_OFF + rc.s * p.stage_bytes + A_STAGE_BYTES,
                        wsrc + (size_t)(q0 + q) * p.b_stage_bytes, (uint32_t)p.b_stage_bytes, bar_fullB + 8 * rc.s);
        }
        __syncwarp();
        rc.advance(nstages);
      }
    }
  } else if (warp == 5) {
    // =========================================================================== MMA issuer
    const uint32_t idesc = umma_idesc(kTF32, TILE_PX, p.Cout);
    RingCursor rc;
    uint32_t tcount = 0;
    for (int item = item_beg; item < item_end; ++item, ++tcount) {
      const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
      mbar_wait(bar_tempty + 8 * acc, accph ^ 1u, 50);  // epilogue drained this accumulator
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * (uint32_t)p.acc_cols;
      for (int q = 0; q < chunks_per_item; ++q) {
        mbar_wait(bar_fullB + 8 * rc.s, rc.ph, 32);
        mbar_wait(bar_fullA + 8 * rc.s, rc.ph, 32);
        tc_fence_after();
        const uint32_t a_addr = smem_base + SMEM_STAGE_OFF + rc.s * p.stage_bytes;
        const uint64_t adesc = umma_desc_sw128(a_addr);
        const uint64_t bdesc = umma_desc_sw128(a_addr + A_STAGE_BYTES);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // 4 x 32 bytes of K per 128-byte row
            umma<kTF32>(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (q | kk) != 0 ? 1u : 0u);
          umma_commit(bar_empty + 8 * rc.s);       // frees the smem stage when these MMAs retire
          if (q == chunks_per_item - 1) umma_commit(bar_tfull + 8 * acc);   // accumulator complete -> epilogue
        }
        __syncwarp();
        rc.advance(nstages);
      }
    }
  } else {
    // =========================================================================== epilogue (warps 0-3)
    T* yg = reinterpret_cast<T*>(p.y);
    const int64_t npix = (int64_t)p.B * p.H * p.W;
    uint32_t tcount = 0;
    for (int item = item_beg; item < item_end; ++item, ++tcount) {
      const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
      int b, ty0, tx0, tap0;
      item_coords(p, item, b, ty0, tx0, tap0);
      const int r = warp * 32 + lane;
      const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
      const bool valid = (y < p.H) && (x < p.W);
      const int64_t pix = ((int64_t)b * p.H + y) * p.W + x;
      mbar_wait(bar_tfull + 8 * acc, accph, 1000);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + acc * (uint32_t)p.acc_cols;
      if (p.split == 1) {
        T* yrow = yg + pix * p.Cout;
        for (int c0 = 0; c0 < p.Cout; c0 += 16) {
          uint32_t raw[16];
          tmem_ld16(t_row + (uint32_t)c0, raw);
          tmem_ld_wait();
          if (valid) {
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              v[j] = fmaf(__uint_as_float(raw[j]), scale_s[c0 + j], bias_s[c0 + j]);
              if (p.relu) v[j] = fmaxf(v[j], 0.f);
            }
            OutConv<T>::store(yrow + c0, v);
          }
        }
      } else {
        float* prow = p.partial + ((int64_t)(item % p.split) * npix + pix) * p.Cout;
        for (int c0 = 0; c0 < p.Cout; c0 += 16) {
          uint32_t raw[16];
          tmem_ld16(t_row + (uint32_t)c0, raw);
          tmem_ld_wait();
          if (valid) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              reinterpret_cast<uint4*>(prow + c0)[j] = make_uint4(raw[4 * j], raw[4 * j + 1], raw[4 * j + 2], raw[4 * j + 3]);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(bar_tempty + 8 * acc);
    }
  }

  // ---- teardown
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}

// Fixed-order combination of the tap-split partial sums: y = act(scale * sum_g partial[g] + bias).
template <typename T>
__global__ void __launch_bounds__(256) dcn_fwd_tc_combine_kernel(const float* __restrict__ partial, T* __restrict__ y,
                                                                 const float* __restrict__ scale, const float* __restrict__ bias,
                                                                 int64_t npix, int Cout, int split, int relu) {
  const int64_t n4 = npix * Cout / 4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int o = (int)((i * 4) % Cout);
    float4 s = __ldcs(reinterpret_cast<const float4*>(partial) + i);
    for (int g = 1; g < split; ++g) {
      const float4 t = __ldcs(reinterpret_cast<const float4*>(partial + (int64_t)g * npix * Cout) + i);
      s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
    }
    float v[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[j] = fmaf(v[j], scale ? __ldg(scale + o + j) : 1.f, bias ? __ldg(bias + o + j) : 0.f);
      if (relu) v[j] = fmaxf(v[j], 0.f);
    }
    if (sizeof(T) == 4) {
      reinterpret_cast<float4*>(y)[i] = make_float4(v[0], v[1], v[2], v[3]);
    } else {
      reinterpret_cast<uint2*>(y)[i] = make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
    }
  }
}

// weight (Cout,Cin,3,3) fp32 -> [tap][Cin/chunk] blobs of Cout rows x 128 B, 16-byte pieces XOR-swizzled by
// (row & 7): exactly the smem image the UMMA descriptor above expects, so the loader is one bulk copy.
template <typename T>
__global__ void pack_weights_kernel(const float* __restrict__ w, uint4* __restrict__ out, int Cin, int Cout) {
  constexpr int kVec = 16 / sizeof(T);
  constexpr int kChunk = ROW_BYTES / sizeof(T);
  const int nchunk = Cin / kChunk;
  const int64_t total = (int64_t)NUM_TAPS * nchunk * Cout * 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i & 7);
    int64_t t = i >> 3;
    const int n = (int)(t % Cout); t /= Cout;
    const int cc = (int)(t % nchunk);
    const int tap = (int)(t / nchunk);
    uint32_t words[4];
    if (sizeof(T) == 4) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int c = cc * kChunk + j * kVec + e;
        words[e] = to_tf32(w[((int64_t)n * Cin + c) * NUM_TAPS + tap]);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int c = cc * kChunk + j * kVec + 2 * e;
        words[e] = pack_bf16x2(w[((int64_t)n * Cin + c) * NUM_TAPS + tap], w[((int64_t)n * Cin + c + 1) * NUM_TAPS + tap]);
      }
    }
    const int64_t blob = (int64_t)tap * nchunk + cc;
    out[blob * Cout * 8 + (int64_t)n * 8 + (j ^ (n & 7))] = make_uint4(words[0], words[1], words[2], words[3]);
  }
}

bool dense_nhwc(const cpoly_tensor4& t, int C, int H, int W) {
  return t.sc == 1 && t.sw == C && t.sh == (int64_t)W * C && t.sb == (int64_t)H * W * C;
}

int pow2_at_least(int v) {
  int p = 32;
  while (p < v) p <<= 1;
  return p;
}

// Tap split of the schedule: 1 when the pixel tiles alone fill the machine, else 3 or 9.
int choose_split(int num_tiles) {
  if (num_tiles >= 148) return 1;
  if (num_tiles * 3 >= 148) return 3;
  return 9;
}

size_t split_workspace(int split, int64_t npix, int Cout) {
  return split > 1 ? (size_t)split * (size_t)npix * Cout * sizeof(float) : 0;
}

}  // namespace

bool dcn_tc_eligible(const cpoly_dcn_fwd_args& a) {
  const int esize = a.dtype == CPOLY_BF16 ? 2 : 4;
  const int chunk = ROW_BYTES / esize;
  if (a.kh != 3 || a.kw != 3 || a.stride_h != 1 || a.stride_w != 1 || a.pad_h != 1 || a.pad_w != 1 || a.dil_h != 1 ||
      a.dil_w != 1 || a.deformable_groups != 1)
    return false;
  if (a.Cin % chunk != 0 || a.Cout % 16 != 0 || a.Cout < 16 || a.Cout > 256) return false;
  if (!dense_nhwc(a.x, a.Cin, a.H, a.W) || !dense_nhwc(a.y, a.Cout, a.H, a.W)) return false;
  if ((reinterpret_cast<uintptr_t>(a.x.ptr) & 15) || (reinterpret_cast<uintptr_t>(a.y.ptr) & 15) ||
      (reinterpret_cast<uintptr_t>(a.weight_packed) & 15))
    return false;
  if ((int64_t)a.B * a.H * a.W * a.Cin * esize >= (1LL << 32)) return false;  // 32-bit byte offsets in the sampler
  return true;
}

size_t dcn_tc_forward_workspace(const cpoly_dcn_fwd_args& a) {
  const int tiles = a.B * ((a.W + TILE_W - 1) / TILE_W) * ((a.H + TILE_H - 1) / TILE_H);
  return split_workspace(choose_split(tiles), (int64_t)a.B * a.H * a.W, a.Cout) + 256;
}

size_t dcn_tc_packed_bytes(int Cin, int Cout, int kh, int kw, int dtype) {
  if (kh != 3 || kw != 3) return 0;
  const int esize = dtype == CPOLY_BF16 ? 2 : 4;
  if (Cin % (ROW_BYTES / esize) != 0 || Cout % 16 != 0 || Cout > 256) return 0;
  return (size_t)NUM_TAPS * Cin * Cout * esize;
}

int dcn_tc_pack_weights(const float* weight, void* packed, int Cin, int Cout, int kh, int kw, int dtype, cudaStream_t stream) {
  (void)kh; (void)kw;
  const int64_t total = (int64_t)NUM_TAPS * Cin * Cout * (dtype == CPOLY_BF16 ? 2 : 4) / 16;
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (dtype == CPOLY_BF16) pack_weights_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(weight, reinterpret_cast<uint4*>(packed), Cin, Cout);
  else pack_weights_kernel<float><<<blocks, 256, 0, stream>>>(weight, reinterpret_cast<uint4*>(packed), Cin, Cout);
  return check_launch("pack_weights_kernel");
}

int dcn_tc_launch_combine(const float* partial, void* y, int dtype, const float* scale, const float* bias, int64_t npix,
                          int Cout, int split, int relu, cudaStream_t stream) {
  const int64_t n4 = npix * Cout / 4;
  const unsigned blocks = (unsigned)((n4 + 255) / 256 < 148 * 8 ? (n4 + 255) / 256 : 148 * 8);
  if (dtype == CPOLY_BF16)
    dcn_fwd_tc_combine_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(partial, reinterpret_cast<__nv_bfloat16*>(y), scale, bias, npix, Cout, split, relu);
  else
    dcn_fwd_tc_combine_kernel<float><<<blocks, 256, 0, stream>>>(partial, reinterpret_cast<float*>(y), scale, bias, npix, Cout, split, relu);
  return check_launch("dcn_fwd_tc_combine_kernel");
}

int dcn_forward_tc(const cpoly_dcn_fwd_args& a, cudaStream_t stream) {
  TcParams p;
  const int esize = a.dtype == CPOLY_BF16 ? 2 : 4;
  p.x = a.x.ptr; p.offset = a.offset; p.mask = a.mask; p.wpacked = a.weight_packed; p.bias = a.bias; p.scale = a.out_scale; p.offset_bias = a.offset_bias; p.mask_bias = a.mask_bias;
  p.y = a.y.ptr;
  p.B = a.B; p.Cin = a.Cin; p.Cout = a.Cout; p.H = a.H; p.W = a.W;
  p.om_dtype = a.om_dtype; p.mask_is_logit = a.mask_is_logit; p.relu = a.relu;
  p.tiles_x = (a.W + TILE_W - 1) / TILE_W;
  p.tiles_y = (a.H + TILE_H - 1) / TILE_H;
  p.num_tiles = a.B * p.tiles_x * p.tiles_y;
  if (p.num_tiles == 0) return CPOLY_OK;
  const int64_t npix = (int64_t)a.B * a.H * a.W;
  p.split = choose_split(p.num_tiles);
  const size_t need = split_workspace(p.split, npix, a.Cout);
  if (p.split > 1 && (!a.workspace || a.workspace_bytes < need || (reinterpret_cast<uintptr_t>(a.workspace) & 15)))
    p.split = 1;  // no scratch supplied: unsplit schedule (same result, fewer busy SMs)
  p.partial = p.split > 1 ? reinterpret_cast<float*>(a.workspace) : nullptr;
  p.taps_per_item = NUM_TAPS / p.split;
  p.num_items = p.num_tiles * p.split;
  p.nchunk = a.Cin / (ROW_BYTES / esize);
  p.b_stage_bytes = a.Cout * ROW_BYTES;
  p.stage_bytes = A_STAGE_BYTES + p.b_stage_bytes;
  int stages = (SMEM_LIMIT - SMEM_STAGE_OFF) / p.stage_bytes;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  // 4 stages are as fast as 6 (r02 sweep: 2 / 3 / 4 / 6 stages = 79.8 / 77.9 / 78.0 / 79.9 us on 64->64 @256x512) and
  // leave 48 KB more of the SM's 256 KB to L1, which is what the corner gathers live on
  int want = 4;
  if (const char* e = getenv("CPOLY_TC_STAGES")) want = atoi(e);  // profiling knob
  if (want >= 2 && want < stages) stages = want;
  CPOLY_REQUIRE(stages >= 2, CPOLY_EUNSUPPORTED, "not enough shared memory for a 2-stage pipeline");
  p.stages = stages;
  p.acc_cols = a.Cout;
  p.tmem_cols = pow2_at_least(2 * a.Cout);
  const int smem_bytes = SMEM_STAGE_OFF + stages * p.stage_bytes + 1024;  // +1 KB alignment slack

  int dev = 0, sms = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.num_items < sms ? p.num_items : sms;
  // offsets + mask as ONE dense 32-channel NHWC tensor (channels 0..17 / 18..26)?  -> coalesced staging
  const int oes = a.om_dtype == CPOLY_BF16 ? 2 : 4;
  p.om_esize = oes;
  p.late_issue = 1;
  if (const char* e = getenv("CPOLY_TC_LATE")) p.late_issue = atoi(e) != 0;
  const cpoly_tensor4 &o = a.offset, &m = a.mask;
  bool om_fast = o.sc == 1 && o.sw == 32 && o.sh == (int64_t)a.W * 32 && o.sb == (int64_t)a.H * a.W * 32 && m.sc == 1 && m.sw == 32 &&
                 m.sh == o.sh && m.sb == o.sb && reinterpret_cast<const uint8_t*>(m.ptr) == reinterpret_cast<const uint8_t*>(o.ptr) + 18 * oes &&
                 (reinterpret_cast<uintptr_t>(o.ptr) & 15) == 0 && (int64_t)a.B * a.H * a.W < (1LL << 26);
  if (const char* e = getenv("CPOLY_TC_OM")) om_fast = om_fast && atoi(e) != 0;   // profiling knob: CPOLY_TC_OM=0 keeps the strided loads
#define CPOLY_LAUNCH_FWD(TT, FAST)                                                                                            \
  do {                                                                                                                        \
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_fwd_tc_kernel<TT, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes)); \
    dcn_fwd_tc_kernel<TT, FAST><<<grid, NTHREADS, smem_bytes, stream>>>(p);                                                   \
  } while (0)
  if (a.dtype == CPOLY_BF16) {
    if (om_fast) CPOLY_LAUNCH_FWD(__nv_bfloat16, true); else CPOLY_LAUNCH_FWD(__nv_bfloat16, false);
  } else {
    if (om_fast) CPOLY_LAUNCH_FWD(float, true); else CPOLY_LAUNCH_FWD(float, false);
  }
#undef CPOLY_LAUNCH_FWD
  if (int rc = check_launch("dcn_fwd_tc_kernel")) return rc;
  if (p.split > 1) return dcn_tc_launch_combine(p.partial, a.y.ptr, a.dtype, a.out_scale, a.bias, npix, a.Cout, p.split, a.relu, stream);
  return CPOLY_OK;
}

}  // namespace cpoly
