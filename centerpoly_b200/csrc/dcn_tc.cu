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
constexpr int MAX_A_STAGES = 8, MAX_B_STAGES = 4;
constexpr int EPI_WARPS = 4;
constexpr int SAMPLER_WARP0 = 6;
constexpr int SAMPLER_WARPS = 16;
constexpr int NS = SAMPLER_WARPS * 32;         // sampler threads
constexpr int NTHREADS = (SAMPLER_WARP0 + SAMPLER_WARPS) * 32;
constexpr int ROWS_PER_PASS = NS / 8;          // 8 lanes x 16 B cover one 128-byte row
constexpr int PASSES = TILE_PX / ROWS_PER_PASS;
constexpr int GROUPS = 4;                      // sampler groups: group g owns the steps t = g (mod GROUPS)
constexpr int GROUP_THREADS = NS / GROUPS;     // 128
constexpr int GROUP_ROWS = GROUP_THREADS / 8;  // 16 rows of the tile per pass (8 lanes x 16 B per row)
constexpr int GROUP_PASSES = TILE_PX / GROUP_ROWS;   // 8 passes per step
constexpr int PASS_OFF = GROUP_ROWS * 4;       // bytes of corner-offset words between two passes
constexpr int PARAM_THREADS = EPI_WARPS * 32;  // warps 0-3: thread r owns pixel r of the tile (= TMEM lane r)

// shared-memory map (dynamic, 1024-byte aligned)
constexpr int SMEM_BAR_OFF = 0;        // mbarriers + tmem pointer
constexpr int SMEM_VEC_OFF = 1024;     // scale[256] + bias[256] floats
constexpr int SMEM_OM_OFF = 3072;      // staged offset/mask tile of the next item: 128 pixels x (128 + 16) bytes
constexpr int OM_PITCH_F32 = 144, OM_PITCH_BF16 = 80;   // 16-byte aligned rows; 8 consecutive rows cover all 32 banks
constexpr int SMEM_PAR_OFF = SMEM_OM_OFF + TILE_PX * OM_PITCH_F32;   // two parameter buffers of taps_per_item * 128 * (4 + 8 | 16) B
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
  int a_stages, b_stages, b_stage_bytes;   // A ring (16 KB tiles, >= GROUPS stages) and the independent weight ring
  int b_off;       // byte offset of the weight ring (1024-aligned, after the A ring)
  int st_off;      // byte offset of the 16 KB output staging tile (after the weight ring)
  int acc_cols, tmem_cols;
  int om_esize;    // element size of the fused 32-channel offset/mask tensor (kOmFast kernels)
  int pbuf_bytes;  // one parameter buffer: taps_per_item * 128 base-offset words, then as many weight entries (8 B bf16 / 16 B fp32)
  int stage_off;   // byte offset of the operand ring (1024-aligned, after the parameter buffers)
  int pf_halo, pf_dist;   // L1 prefetch of a coming item's window: halo in pixels (< 0: off), distance in items
  long long* trace;  // profiling only (CPOLY_TC_TRACE): clock64 event log of CTA 0, 5 roles x 4096 entries
};

// ---- blend of the four corner vectors (16 B each) into one 16-B piece of the A row ----------------
// bf16: weights arrive as bf16x2 pairs (w, w); products and sums are packed bf16x2 FMAs — half the issue slots of
// unpack + fp32 FMA + repack, and the A operand is rounded to bf16 anyway (tolerance 1e-2 in bf16 mode).
// fp32: fp32 FMAs, result rounded to tf32 (round-to-nearest) for kind::tf32.
template <typename T> struct Blend;
template <> struct Blend<__nv_bfloat16> {
  static constexpr int kVec = 8;  // channels per 16 bytes
  // parameter entry: the four mask * bilinear weights as 4 x bf16 (8 bytes); the (w, w) pairs of the packed FMAs are made
  // with one PRMT each at blend time
  using Packed = uint2;
  static constexpr int kWBytes = 8;
  __device__ static __forceinline__ Packed pack(float w0, float w1, float w2, float w3) { return make_uint2(pack_bf16x2(w0, w1), pack_bf16x2(w2, w3)); }
  __device__ static __forceinline__ Packed lds(uint32_t a) {
    uint2 v;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
  }
  __device__ static __forceinline__ void sts(uint32_t a, Packed v) { asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a), "r"(v.x), "r"(v.y) : "memory"); }
  __device__ static __forceinline__ uint4 run(const uint4 (&c)[4], const Packed w) {
    return run(c, make_uint4(__byte_perm(w.x, 0u, 0x1010), __byte_perm(w.x, 0u, 0x3232), __byte_perm(w.y, 0u, 0x1010), __byte_perm(w.y, 0u, 0x3232)));
  }
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
  using Packed = uint4;   // 4 x fp32
  static constexpr int kWBytes = 16;
  __device__ static __forceinline__ Packed pack(float w0, float w1, float w2, float w3) { return make_uint4(__float_as_uint(w0), __float_as_uint(w1), __float_as_uint(w2), __float_as_uint(w3)); }
  __device__ static __forceinline__ Packed lds(uint32_t a) { return lds16(a); }
  __device__ static __forceinline__ void sts(uint32_t a, Packed v) { sts16(a, v); }
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
  __device__ static __forceinline__ void pieces(const float (&v)[16], uint4 (&o)[2]) {
    o[0] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    o[1] = make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]), pack_bf16x2(v[12], v[13]), pack_bf16x2(v[14], v[15]));
  }
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
  __device__ static __forceinline__ void pieces(const float (&v)[16], uint4 (&o)[4]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = make_uint4(__float_as_uint(v[4 * j]), __float_as_uint(v[4 * j + 1]), __float_as_uint(v[4 * j + 2]), __float_as_uint(v[4 * j + 3]));
  }
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
  const uint32_t bar_fullA = smem_base + SMEM_BAR_OFF;                 // [MAX_A_STAGES]
  const uint32_t bar_empty = bar_fullA + 8 * MAX_A_STAGES;             // [MAX_A_STAGES] A stage consumed
  const uint32_t bar_fullB = bar_empty + 8 * MAX_A_STAGES;             // [MAX_B_STAGES]
  const uint32_t bar_emptyB = bar_fullB + 8 * MAX_B_STAGES;            // [MAX_B_STAGES] weight stage consumed
  const uint32_t bar_tfull = bar_emptyB + 8 * MAX_B_STAGES;            // [2]
  const uint32_t bar_tempty = bar_tfull + 16;                          // [2]
  const uint32_t bar_pfull = bar_tempty + 16;                          // [2] parameter buffer written (PARAM_THREADS arrivals)
  const uint32_t bar_pempty = bar_pfull + 16;                          // [2] parameter buffer released (one arrival per sampler warp)
  const uint32_t bar_first = bar_pempty + 16;                          // [NUM_TAPS] first item only: tap k written (PARAM_THREADS arrivals)
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SMEM_BAR_OFF + 8 * (2 * MAX_A_STAGES + 2 * MAX_B_STAGES + 8 + NUM_TAPS));
  float* scale_s = reinterpret_cast<float*>(smem + SMEM_VEC_OFF);
  float* bias_s = scale_s + 256;
  const uint32_t om_s = smem_base + SMEM_OM_OFF, par_s = smem_base + SMEM_PAR_OFF;
  const uint32_t pw_off = (uint32_t)(p.taps_per_item * TILE_PX) * 4u;   // the weight entries follow the base-offset words inside a buffer
  const uint32_t pixb = (uint32_t)p.Cin * (uint32_t)sizeof(T), rowb = (uint32_t)p.W * pixb;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- one-time setup
  if (warp == 4 && lane == 0) {
    for (int s = 0; s < MAX_A_STAGES; ++s) {
      mbar_init(bar_fullA + 8 * s, SAMPLER_WARPS / GROUPS);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < MAX_B_STAGES; ++s) {
      mbar_init(bar_fullB + 8 * s, 1);
      mbar_init(bar_emptyB + 8 * s, 1);
    }
    for (int k = 0; k < NUM_TAPS; ++k) mbar_init(bar_first + 8 * k, PARAM_THREADS);
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar_tfull + 8 * a, 1);
      mbar_init(bar_tempty + 8 * a, EPI_WARPS * 32);
      mbar_init(bar_pfull + 8 * a, PARAM_THREADS);
      mbar_init(bar_pempty + 8 * a, (uint32_t)(p.taps_per_item * p.nchunk) * (SAMPLER_WARPS / GROUPS));
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
  const uint32_t nstages = (uint32_t)p.a_stages, nbstages = (uint32_t)p.b_stages;
  // A CTA walks a CONTIGUOUS range of items: consecutive tiles of a tile row share 43 % of their bilinear footprint
  // (and the tap groups of one tile all of it), so the halo of the next item is already in this SM's L1 - with the
  // round-robin deal of round 1 every item started cold (L1 sector hit rate 83 %, long-scoreboard stalls dominant).
  const int item_beg = (int)(((long long)p.num_items * blockIdx.x) / gridDim.x);
  const int item_end = (int)(((long long)p.num_items * (blockIdx.x + 1)) / gridDim.x);

  const int n_items = item_end - item_beg;
  int tr_n = 0;
  auto tr = [&](int role, int tag) {
    if (p.trace && blockIdx.x == 0 && lane == 0 && tr_n < 2048) { p.trace[role * 4096 + 2 * tr_n] = clock64(); p.trace[role * 4096 + 2 * tr_n + 1] = tag; ++tr_n; }
  };

  if (warp >= SAMPLER_WARP0) {
    // =========================================================================== samplers
    // The samplers only gather and blend; the per-(pixel, tap) parameters of item j are written by warps 0-3 into
    // parameter buffer j & 1 while item j-1 is being sampled (bar_pfull / bar_pempty; the CTA's first item is handed
    // over tap by tap through bar_first so that sampling starts after one tap's worth of parameter math).
    // The 16 warps form GROUPS = 4 groups of 4; group g builds the WHOLE A tile of the steps t = g (mod 4) of the CTA's
    // step sequence (step = one 128-byte channel chunk of one tap of one item) in 8 passes of 16 rows.  A warp pays the
    // per-step costs (stage wait, cursor update, the proxy fence that drains its loads, arrives) once per 8 passes, the
    // pass loop is straight-line code with immediate offsets, and the groups run out of phase instead of hitting the
    // LSU in lockstep.  One rolling software pipeline runs across all passes, steps and items of a group: two passes
    // (8 x 16 B per thread) are in flight while the previous one is blended.
    const int st = threadIdx.x - SAMPLER_WARP0 * 32;
    const int grp = st / GROUP_THREADS, gt = st - grp * GROUP_THREADS;
    const int row0 = gt >> 3, cg = gt & 7;
    const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.x);
    const uint32_t cbyte0 = (uint32_t)(cg * 16);
    const int total = n_items * chunks_per_item;
    const uint32_t st_off = (uint32_t)(row0 * ROW_BYTES + ((cg ^ (row0 & 7)) << 4));   // + pass * 16 rows * 128 B
    uint4 c[2][4];
    typename Blend<T>::Packed pw[2];
    constexpr uint32_t kWB = Blend<T>::kWBytes, PASS_W = GROUP_ROWS * kWB;
    // cursor: (item, tap, chunk) of a step; `par` / `cbyte` are what the passes need
    struct Cur { int t, j, kl, cc; uint32_t par, parw, cbyte; };
    auto seek = [&](Cur& u, int adv) {            // move u by `adv` steps (adv <= GROUPS), waiting for the parameters it lands on
      u.t += adv;
      if (u.t >= total) return;
      u.cc += adv;
      while (u.cc >= p.nchunk) { u.cc -= p.nchunk; ++u.kl; }
      while (u.kl >= p.taps_per_item) {           // every item's hand-over is observed in order (phase parity stays unambiguous)
        if (u.j == 0) mbar_wait(bar_pfull, 0u);   // (the first item was consumed through bar_first)
        u.kl -= p.taps_per_item; ++u.j;
        mbar_wait(bar_pfull + 8 * (u.j & 1), (uint32_t)(u.j >> 1) & 1u);
      }
      if (u.j == 0) mbar_wait(bar_first + 8 * u.kl, 0u);
      const uint32_t buf = par_s + (uint32_t)(u.j & 1) * (uint32_t)p.pbuf_bytes, e = (uint32_t)(u.kl * TILE_PX + row0);
      u.par = buf + e * 4u;
      u.parw = buf + pw_off + e * kWB;
      u.cbyte = cbyte0 + (uint32_t)u.cc * ROW_BYTES;
    };
    // parameters + corner loads of one pass.  Entry = one word (byte offset of the clamped base corner | bit 0: the
    // x+1 corner is a different pixel | bit 1: the y+1 corner is) + the four weights; a corner that falls outside the
    // image has weight 0 and re-reads an in-range neighbour.
    auto issue = [&](uint4 (&cb)[4], typename Blend<T>::Packed& pwb, const Cur& u, int pass) {
      uint32_t w;
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(u.par + (uint32_t)(pass * PASS_OFF)));
      pwb = Blend<T>::lds(u.parw + (uint32_t)pass * PASS_W);
      const uint32_t o00 = (w & ~3u) + u.cbyte;
      const uint32_t dx = (w & 1u) ? pixb : 0u, dy = (w & 2u) ? rowb : 0u;
      cb[0] = ldg16(xb + o00);
      cb[1] = ldg16(xb + (o00 + dx));
      cb[2] = ldg16(xb + (o00 + dy));
      cb[3] = ldg16(xb + (o00 + dy + dx));
    };
    Cur cur{-1, 0, 0, -1, 0u, 0u, 0u};
    seek(cur, grp + 1);                           // the group's first step (grp < GROUPS <= total is not guaranteed: checked below)
    if (cur.t < total) {
      issue(c[0], pw[0], cur, 0);
      issue(c[1], pw[1], cur, 1);
    }
    uint32_t s_idx = (uint32_t)grp % nstages, s_ph = ((uint32_t)grp / nstages) & 1u;
    while (cur.t < total) {
      // Look ahead to the group's next step only when its parameters can exist before this step is finished: the
      // parameter buffers hold two items, so a step two or more items ahead (items of fewer than GROUPS steps: tap-split
      // schedules of narrow layers) is produced only after THIS step has released its item - seeking to it now would wait
      // for ourselves.  Such a step is looked up after the arrive below, at the price of an empty pipeline.
      Cur nxt = cur;
      const int nt = cur.t + GROUPS;
      const bool ahead = nt >= total || nt / chunks_per_item <= cur.j + 1;
      if (ahead) seek(nxt, GROUPS);
      const bool more = ahead && nxt.t < total;
      mbar_wait(bar_empty + 8 * s_idx, s_ph ^ 1u);
      if (warp == SAMPLER_WARP0) tr(0, 1); else if (warp == SAMPLER_WARP0 + SAMPLER_WARPS - 1) tr(4, 1);
      const uint32_t a_st = smem_base + p.stage_off + s_idx * A_STAGE_BYTES + st_off;
#pragma unroll
      for (int i = 0; i < GROUP_PASSES; ++i) {
        sts16(a_st + (uint32_t)(i * GROUP_ROWS * ROW_BYTES), Blend<T>::run(c[i & 1], pw[i & 1]));
        if (i + 2 < GROUP_PASSES) issue(c[i & 1], pw[i & 1], cur, i + 2);
        else if (more) issue(c[i & 1], pw[i & 1], nxt, i + 2 - GROUP_PASSES);
      }
      fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core (async proxy); drains this thread's loads
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(bar_fullA + 8 * s_idx);
        mbar_arrive(bar_pempty + 8 * (cur.j & 1));   // this warp has consumed the step's parameters
      }
      if (warp == SAMPLER_WARP0) tr(0, 2); else if (warp == SAMPLER_WARP0 + SAMPLER_WARPS - 1) tr(4, 2);
      if (!ahead) {
        seek(nxt, GROUPS);
        if (nxt.t < total) {
          issue(c[0], pw[0], nxt, 0);
          issue(c[1], pw[1], nxt, 1);
        }
      }
      cur = nxt;
      s_idx += (uint32_t)GROUPS;
      while (s_idx >= nstages) { s_idx -= nstages; s_ph ^= 1u; }
    }
  } else if (warp == 4) {
    // =========================================================================== weight loader
    // (whole warp walks the ring, one elected lane issues: see elect_one() in tc_common.cuh)
    RingCursor rc;
    const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.wpacked);
    for (int item = item_beg; item < item_end; ++item) {
      const int q0 = (item % p.split) * chunks_per_item;
      for (int q = 0; q < chunks_per_item; ++q) {
        mbar_wait(bar_emptyB + 8 * rc.s, rc.ph ^ 1u, 100);
        tr(1, 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(bar_fullB + 8 * rc.s, (uint32_t)p.b_stage_bytes);
          bulk_copy_g2s(smem_base + p.b_off + rc.s * p.b_stage_bytes,
                        wsrc + (size_t)(q0 + q) * p.b_stage_bytes, (uint32_t)p.b_stage_bytes, bar_fullB + 8 * rc.s);
        }
        __syncwarp();
        rc.advance(nbstages);
      }
    }
  } else if (warp == 5) {
    // =========================================================================== MMA issuer
    const uint32_t idesc = umma_idesc(kTF32, TILE_PX, p.Cout);
    RingCursor rc, rb;
    uint32_t tcount = 0;
    for (int item = item_beg; item < item_end; ++item, ++tcount) {
      const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
      mbar_wait(bar_tempty + 8 * acc, accph ^ 1u, 50);  // epilogue drained this accumulator
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * (uint32_t)p.acc_cols;
      for (int q = 0; q < chunks_per_item; ++q) {
        mbar_wait(bar_fullB + 8 * rb.s, rb.ph, 32);
        tr(2, 1);
        mbar_wait(bar_fullA + 8 * rc.s, rc.ph, 32);
        tr(2, 2);
        tc_fence_after();
        const uint32_t a_addr = smem_base + p.stage_off + rc.s * A_STAGE_BYTES;
        const uint64_t adesc = umma_desc_sw128(a_addr);
        const uint64_t bdesc = umma_desc_sw128(smem_base + p.b_off + rb.s * p.b_stage_bytes);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // 4 x 32 bytes of K per 128-byte row
            umma<kTF32>(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (q | kk) != 0 ? 1u : 0u);
          umma_commit(bar_empty + 8 * rc.s);       // frees the A stage and the weight stage when these MMAs retire
          umma_commit(bar_emptyB + 8 * rb.s);
          if (q == chunks_per_item - 1) umma_commit(bar_tfull + 8 * acc);   // accumulator complete -> epilogue
        }
        __syncwarp();
        tr(2, 3);
        rc.advance(nstages);
        rb.advance(nbstages);
      }
    }
  } else {
    // =========================================================================== parameters + epilogue (warps 0-3)
    // Thread r owns pixel r of the tile: it computes the 9 (or 3, or 1) parameter entries of that pixel for the NEXT
    // item, then drains accumulator row r (TMEM lane r) of the CURRENT one.
    T* yg = reinterpret_cast<T*>(p.y);
    const int64_t npix = (int64_t)p.B * p.H * p.W;
    const int r = threadIdx.x;
    const int ppp = p.om_esize * 2;                       // kOmFast: 16-byte pieces per pixel (32 channels): 8 or 4
    const int ppp_shift = p.om_esize == 4 ? 3 : 2;
    const uint32_t om_pitch = p.om_esize == 4 ? (uint32_t)OM_PITCH_F32 : (uint32_t)OM_PITCH_BF16;
    // kOmFast: the item's 128 x 32 offset/mask values -> shared memory with coalesced 16-byte async copies (no
    // registers held while they fly; issued one item ahead)
    auto om_issue = [&](int item) {
      if constexpr (kOmFast) {
        int b, ty0, tx0, tap0;
        item_coords(p, item, b, ty0, tx0, tap0);
        const uint8_t* base = reinterpret_cast<const uint8_t*>(p.offset.ptr);
        for (int idx = r; idx < TILE_PX * ppp; idx += PARAM_THREADS) {
          const int rr = idx >> ppp_shift, jj = idx & (ppp - 1);
          const int y = ty0 + (rr >> 4), x = tx0 + (rr & (TILE_W - 1));
          if (y < p.H && x < p.W) {
            const uint8_t* src = base + ((size_t)((b * p.H + y) * p.W + x) * 32) * p.om_esize + jj * 16;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(om_s + (uint32_t)rr * om_pitch + (uint32_t)jj * 16u), "l"(src) : "memory");
          }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      }
    };
    auto produce = [&](int jl) {   // parameters of local item jl -> buffer jl & 1
      const int item = item_beg + jl;
      int b, ty0, tx0, tap0;
      item_coords(p, item, b, ty0, tx0, tap0);
      const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
      const bool inside = (y < p.H) && (x < p.W);
      if (warp == 0) tr(3, 10);
      float dyv[NUM_TAPS], dxv[NUM_TAPS], mv[NUM_TAPS];
      if constexpr (kOmFast) {
        asm volatile("cp.async.wait_all;" ::: "memory");
        named_bar_sync(1, PARAM_THREADS);                 // every thread's pieces have landed
        const uint32_t row = om_s + (uint32_t)r * om_pitch;
        if (p.om_esize == 4) {
          uint32_t w[28];
#pragma unroll
          for (int i = 0; i < 7; ++i) { const uint4 v = lds16(row + 16u * i); w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w; }
#pragma unroll
          for (int k = 0; k < NUM_TAPS; ++k) { dyv[k] = __uint_as_float(w[2 * k]); dxv[k] = __uint_as_float(w[2 * k + 1]); mv[k] = __uint_as_float(w[18 + k]); }
        } else {
          uint32_t w[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) { const uint4 v = lds16(row + 16u * i); w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w; }
#pragma unroll
          for (int k = 0; k < NUM_TAPS; ++k) {
            dyv[k] = __uint_as_float(w[k] << 16); dxv[k] = __uint_as_float(w[k] & 0xffff0000u);
            const uint32_t h = w[9 + (k >> 1)];
            mv[k] = __uint_as_float((k & 1) ? (h & 0xffff0000u) : (h << 16));
          }
        }
        named_bar_sync(1, PARAM_THREADS);                 // the staging tile is free again
        if (jl + 1 < n_items) om_issue(item + 1);
      } else {
#pragma unroll
        for (int k = 0; k < NUM_TAPS; ++k) {
          dyv[k] = dxv[k] = mv[k] = 0.f;
          if (inside && k >= tap0 && k < tap0 + p.taps_per_item) {
            const int64_t opix = (int64_t)b * p.offset.sb + (int64_t)y * p.offset.sh + (int64_t)x * p.offset.sw;
            dyv[k] = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k) * p.offset.sc, p.om_dtype);
            dxv[k] = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k + 1) * p.offset.sc, p.om_dtype);
            mv[k] = ld_dyn(p.mask.ptr, (int64_t)b * p.mask.sb + (int64_t)k * p.mask.sc + (int64_t)y * p.mask.sh + (int64_t)x * p.mask.sw, p.om_dtype);
          }
        }
      }
      // the samplers have released this buffer (it held item jl - 2)
      if (warp == 0) tr(3, 11);
      mbar_wait(bar_pempty + 8 * (jl & 1), ((uint32_t)(jl >> 1) & 1u) ^ 1u, 100);
      if (warp == 0) tr(3, 12);
      const uint32_t buf = par_s + (uint32_t)(jl & 1) * (uint32_t)p.pbuf_bytes;
      const uint32_t img_byte = (uint32_t)(b * p.H * p.W * p.Cin) * (uint32_t)sizeof(T);   // < 2^32 (checked on the host)
      // branch-free entry math, three taps per batch so that their dependent chains interleave (one warp per scheduler
      // has nobody else to hide latency behind)
      const float Hf = (float)p.H, Wf = (float)p.W;
      const int tap_end = tap0 + p.taps_per_item;
#pragma unroll
      for (int kb = 0; kb < NUM_TAPS; kb += 3) {
        if (kb + 3 <= tap0 || kb >= tap_end) continue;     // (uniform)
        uint32_t po[3];
        typename Blend<T>::Packed pw[3];
#pragma unroll
        for (int kk = 0; kk < 3; ++kk) {
          const int k = kb + kk;
          float dy = dyv[k], dx = dxv[k], mr = mv[k];
          if (p.offset_bias) { dy += __ldg(p.offset_bias + 2 * k); dx += __ldg(p.offset_bias + 2 * k + 1); }
          if (p.mask_bias) mr += __ldg(p.mask_bias + k);
          float m = p.mask_is_logit ? __fdividef(1.0f, 1.0f + __expf(-mr)) : mr;   // <= 2e-7 from the exact sigmoid
          const float py = (float)(y - 1 + k / 3) + dy;
          const float px = (float)(x - 1 + k % 3) + dx;
          const bool ok = inside && py > -1.0f && px > -1.0f && py < Hf && px < Wf;
          m = ok ? m : 0.0f;
          const float fy = floorf(py), fx = floorf(px);
          const int y0 = ok ? (int)fy : 0, x0 = ok ? (int)fx : 0;          // in [-1, H-1] x [-1, W-1]
          const float ly = py - fy, lx = px - fx, hy = 1.0f - ly, hx = 1.0f - lx;
          const bool vy0 = y0 >= 0, vy1 = y0 + 1 <= p.H - 1, vx0 = x0 >= 0, vx1 = x0 + 1 <= p.W - 1;
          // base corner clamped into the image; the +1 corners are distinct pixels only when both neighbours exist, else
          // they re-read the base row / column (their weight, or the base's, is zero then)
          const int yb = vy0 ? y0 : 0, xq = vx0 ? x0 : 0;
          po[kk] = (img_byte + (uint32_t)(yb * p.W + xq) * pixb) | ((vx0 && vx1) ? 1u : 0u) | ((vy0 && vy1) ? 2u : 0u);
          const float lym = vy1 ? ly * m : 0.0f, hym = vy0 ? hy * m : 0.0f;
          pw[kk] = Blend<T>::pack(vx0 ? hym * hx : 0.0f, vx1 ? hym * lx : 0.0f, vx0 ? lym * hx : 0.0f, vx1 ? lym * lx : 0.0f);
        }
#pragma unroll
        for (int kk = 0; kk < 3; ++kk) {
          const int k = kb + kk;
          if (k < tap0 || k >= tap_end) continue;
          const uint32_t e = (uint32_t)((k - tap0) * TILE_PX + r);
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(buf + e * 4u), "r"(po[kk]) : "memory");
          Blend<T>::sts(buf + pw_off + e * (uint32_t)Blend<T>::kWBytes, pw[kk]);
          if (jl == 0) mbar_arrive(bar_first + 8 * (k - tap0));   // the CTA's first item is handed over tap by tap
        }
      }
      mbar_arrive(bar_pfull + 8 * (jl & 1));   // release: the stores above are visible to the samplers that acquire it
      if (warp == 0) tr(3, 13);
    };
    // order: the parameters of item j+2 go first (the samplers look one step ahead and reach them before item j+1 is
    // finished), the epilogue of item j second (its accumulator is not needed again before item j+2's first MMA)
    // pull the bilinear footprint (tile + halo) of a coming item towards this SM's L1: a sampler wait covers 32 cache
    // lines and is as long as the slowest of them, so even the compulsory misses put an L2 round trip into most waits
    auto prefetch = [&](int jl) {
      if (p.pf_halo < 0 || jl >= n_items) return;
      int b, ty0, tx0, tap0;
      item_coords(p, item_beg + jl, b, ty0, tx0, tap0);
      const int hw = p.pf_halo, ww = TILE_W + 2 * hw, wh = TILE_H + 2 * hw;
      const int lpp = (p.Cin * (int)sizeof(T)) >> 7;                     // 128-byte lines per pixel
      const uint8_t* img = reinterpret_cast<const uint8_t*>(p.x) + (size_t)b * p.H * p.W * p.Cin * sizeof(T);
      for (int i = r; i < ww * wh * lpp; i += PARAM_THREADS) {
        const int px = i / lpp, l = i - px * lpp;
        const int wy = px / ww, wx = px - wy * ww;
        const int yy = ty0 - hw + wy, xx = tx0 - hw + wx;
        if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W)
          asm volatile("prefetch.global.L1 [%0];" ::"l"(img + ((size_t)(yy * p.W + xx) * p.Cin) * sizeof(T) + l * 128));
      }
    };
    if (n_items > 0) { om_issue(item_beg); produce(0); }
    if (n_items > 1) produce(1);
    uint32_t tcount = 0;
    for (int item = item_beg; item < item_end; ++item, ++tcount) {
      prefetch((int)tcount + p.pf_dist);
      if (item + 2 < item_end) produce((int)tcount + 2);
      const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
      int b, ty0, tx0, tap0;
      item_coords(p, item, b, ty0, tx0, tap0);
      const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
      const bool valid = (y < p.H) && (x < p.W);
      const int64_t pix = ((int64_t)b * p.H + y) * p.W + x;
      mbar_wait(bar_tfull + 8 * acc, accph, 1000);
      if (warp == 0) tr(3, 20);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + acc * (uint32_t)p.acc_cols;
      if (p.split == 1) {
        // Rows go through a swizzled 16 KB staging tile (128 rows x one 128-byte channel chunk) and leave as coalesced
        // 16-byte stores, 4 whole rows per warp instruction: a row-per-lane STG.128 touches 32 lines per instruction
        // (114 L1 tag requests per step on the 64-channel layer, a tenth of the LSU traffic of the whole CTA).
        constexpr int CE = ROW_BYTES / (int)sizeof(T);          // channels per staging chunk: 64 bf16 / 32 fp32
        constexpr int PG = 16 * (int)sizeof(T) / 16;            // 16-byte pieces per 16 channels: 2 / 4
        const uint32_t stg = smem_base + (uint32_t)p.st_off;
        uint8_t* ybase = reinterpret_cast<uint8_t*>(yg);
        for (int cc = 0; cc < p.Cout; cc += CE) {
          const int width = (p.Cout - cc < CE) ? p.Cout - cc : CE;
          for (int c0 = cc; c0 < cc + width; c0 += 32) {   // two 16-column TMEM loads in flight per wait
            uint32_t raw[2][16];
            const bool two = c0 + 16 < cc + width;
            tmem_ld16(t_row + (uint32_t)c0, raw[0]);
            if (two) tmem_ld16(t_row + (uint32_t)(c0 + 16), raw[1]);
            tmem_ld_wait();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              if (h == 1 && !two) continue;
              const int cb = c0 + 16 * h;
              float v[16];
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4) {
                const float4 sc = *reinterpret_cast<const float4*>(scale_s + cb + 4 * j4);
                const float4 bi = *reinterpret_cast<const float4*>(bias_s + cb + 4 * j4);
                v[4 * j4 + 0] = fmaf(__uint_as_float(raw[h][4 * j4 + 0]), sc.x, bi.x);
                v[4 * j4 + 1] = fmaf(__uint_as_float(raw[h][4 * j4 + 1]), sc.y, bi.y);
                v[4 * j4 + 2] = fmaf(__uint_as_float(raw[h][4 * j4 + 2]), sc.z, bi.z);
                v[4 * j4 + 3] = fmaf(__uint_as_float(raw[h][4 * j4 + 3]), sc.w, bi.w);
              }
              if (p.relu) {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
              }
              uint4 pc[PG];
              OutConv<T>::pieces(v, pc);
              const int p0 = ((cb - cc) / 16) * PG;
#pragma unroll
              for (int i = 0; i < PG; ++i) sts16(stg + (uint32_t)(r * ROW_BYTES) + (uint32_t)(((p0 + i) ^ (r & 7)) << 4), pc[i]);
            }
          }
          named_bar_sync(2, PARAM_THREADS);
          const int np = width * (int)sizeof(T) / 16;          // pieces per row in this chunk (8 for a full chunk)
          for (int q = r; q < TILE_PX * np; q += PARAM_THREADS) {
            const int row = (np == 8) ? (q >> 3) : q / np, j = q - row * np;
            const int yy = ty0 + (row >> 4), xx = tx0 + (row & (TILE_W - 1));
            if (yy < p.H && xx < p.W) {
              const uint4 val = lds16(stg + (uint32_t)(row * ROW_BYTES) + (uint32_t)((j ^ (row & 7)) << 4));
              *reinterpret_cast<uint4*>(ybase + ((((int64_t)b * p.H + yy) * p.W + xx) * p.Cout + cc) * (int64_t)sizeof(T) + j * 16) = val;
            }
          }
          named_bar_sync(2, PARAM_THREADS);                     // the staging tile is free again
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
      if (warp == 0) tr(3, 21);
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
int choose_split(int num_tiles, int Cout) {
  if (const char* e = getenv("CPOLY_TC_SPLIT")) {   // profiling knob
    const int v = atoi(e);
    if (v == 1 || v == 3 || v == 9) return v;
  }
  if (num_tiles >= 148) return 1;
  // 64..147 tiles with wide outputs: the unsplit schedule on fewer SMs beats 3 tap groups, whose fp32 partial tiles
  // (Cout x 512 B per item, written and re-read by the combine kernel) cost more than the idle SMs (r02, 64x128 maps:
  // 256->256 36.1 vs 42.2 us, 256->128 31.6 vs 32.1, 256->64 27.8 vs 26.6)
  if (num_tiles >= 64 && Cout >= 128) return 1;
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
  return split_workspace(choose_split(tiles, a.Cout), (int64_t)a.B * a.H * a.W, a.Cout) + 256;
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
  p.split = choose_split(p.num_tiles, a.Cout);
  const size_t need = split_workspace(p.split, npix, a.Cout);
  if (p.split > 1 && (!a.workspace || a.workspace_bytes < need || (reinterpret_cast<uintptr_t>(a.workspace) & 15)))
    p.split = 1;  // no scratch supplied: unsplit schedule (same result, fewer busy SMs)
  p.partial = p.split > 1 ? reinterpret_cast<float*>(a.workspace) : nullptr;
  p.taps_per_item = NUM_TAPS / p.split;
  p.num_items = p.num_tiles * p.split;
  p.nchunk = a.Cin / (ROW_BYTES / esize);
  p.b_stage_bytes = a.Cout * ROW_BYTES;
  p.pbuf_bytes = p.taps_per_item * TILE_PX * (4 + (a.dtype == CPOLY_BF16 ? 8 : 16));
  p.stage_off = ((SMEM_PAR_OFF + 2 * p.pbuf_bytes + 1023) / 1024) * 1024;
  // A ring: at least one stage per sampler group, two when they fit; weight ring: 2..4 stages of Cout x 128 B
  const int room = SMEM_LIMIT - p.stage_off - A_STAGE_BYTES;   // (the output staging tile)
  // One A stage per sampler group and two weight stages: with the 12-byte parameter entries more A stages measured
  // SLOWER (4 / 5 / 6 / 8 stages: 63.5 / 63.5 / 65.6 / 69.7 us) - the sampler is bound by L1 / LSU throughput, not by
  // waiting for a free stage, and every 32 KB of shared memory is 32 KB less L1 (carveout steps 132 / 164 / 196 / 228 KB).
  int b_stages = 2, a_stages = GROUPS;
  if (const char* e = getenv("CPOLY_TC_STAGES")) a_stages = atoi(e);    // profiling knobs
  if (const char* e = getenv("CPOLY_TC_BSTAGES")) b_stages = atoi(e);
  if (b_stages < 2) b_stages = 2;
  if (b_stages > MAX_B_STAGES) b_stages = MAX_B_STAGES;
  if (a_stages > MAX_A_STAGES) a_stages = MAX_A_STAGES;
  if (a_stages < GROUPS) a_stages = GROUPS;
  while (b_stages > 2 && GROUPS * A_STAGE_BYTES + b_stages * p.b_stage_bytes > room) --b_stages;
  while (a_stages > GROUPS && a_stages * A_STAGE_BYTES + b_stages * p.b_stage_bytes > room) --a_stages;
  CPOLY_REQUIRE(a_stages * A_STAGE_BYTES + b_stages * p.b_stage_bytes <= room, CPOLY_EUNSUPPORTED, "not enough shared memory for the operand rings");
  p.a_stages = a_stages; p.b_stages = b_stages;
  p.b_off = p.stage_off + a_stages * A_STAGE_BYTES;
  p.st_off = ((p.b_off + b_stages * p.b_stage_bytes + 1023) / 1024) * 1024;
  p.acc_cols = a.Cout;
  p.tmem_cols = pow2_at_least(2 * a.Cout);
  const int smem_bytes = p.st_off + A_STAGE_BYTES + 1024;  // +1 KB alignment slack

  int dev = 0, sms = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.num_items < sms ? p.num_items : sms;
  // offsets + mask as ONE dense 32-channel NHWC tensor (channels 0..17 / 18..26)?  -> coalesced staging
  const int oes = a.om_dtype == CPOLY_BF16 ? 2 : 4;
  p.om_esize = oes;
  p.pf_halo = -1; p.pf_dist = 1;
  if (const char* e = getenv("CPOLY_TC_PF")) p.pf_halo = atoi(e);     // profiling knobs
  if (const char* e = getenv("CPOLY_TC_PFD")) p.pf_dist = atoi(e);
  p.trace = nullptr;
  static long long* trace_buf = nullptr;
  const bool tracing = getenv("CPOLY_TC_TRACE") != nullptr;
  if (tracing) {
    if (!trace_buf) cudaMalloc(&trace_buf, 5 * 4096 * sizeof(long long));
    cudaMemsetAsync(trace_buf, 0, 5 * 4096 * sizeof(long long), stream);
    p.trace = trace_buf;
  }
  const cpoly_tensor4 &o = a.offset, &m = a.mask;
  bool om_fast = o.sc == 1 && o.sw == 32 && o.sh == (int64_t)a.W * 32 && o.sb == (int64_t)a.H * a.W * 32 && m.sc == 1 && m.sw == 32 &&
                 m.sh == o.sh && m.sb == o.sb && reinterpret_cast<const uint8_t*>(m.ptr) == reinterpret_cast<const uint8_t*>(o.ptr) + 18 * oes &&
                 (reinterpret_cast<uintptr_t>(o.ptr) & 15) == 0 && (int64_t)a.B * a.H * a.W < (1LL << 26);
  if (const char* e = getenv("CPOLY_TC_OM")) om_fast = om_fast && atoi(e) != 0;   // profiling knob: CPOLY_TC_OM=0 keeps the strided loads
  // shared-memory carveout: just what the CTA needs, so that the rest of the SM's 256 KB is L1 for the corner gathers
  int carve = (int)(((long long)(smem_bytes + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024));
  if (carve > 100) carve = 100;
  if (const char* e = getenv("CPOLY_TC_CARVE")) carve = atoi(e);   // profiling knob (-1: driver default)
#define CPOLY_LAUNCH_FWD(TT, FAST)                                                                                            \
  do {                                                                                                                        \
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_fwd_tc_kernel<TT, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes)); \
    if (carve >= 0) CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_fwd_tc_kernel<TT, FAST>, cudaFuncAttributePreferredSharedMemoryCarveout, carve)); \
    dcn_fwd_tc_kernel<TT, FAST><<<grid, NTHREADS, smem_bytes, stream>>>(p);                                                   \
  } while (0)
  if (a.dtype == CPOLY_BF16) {
    if (om_fast) CPOLY_LAUNCH_FWD(__nv_bfloat16, true); else CPOLY_LAUNCH_FWD(__nv_bfloat16, false);
  } else {
    if (om_fast) CPOLY_LAUNCH_FWD(float, true); else CPOLY_LAUNCH_FWD(float, false);
  }
#undef CPOLY_LAUNCH_FWD
  if (int rc = check_launch("dcn_fwd_tc_kernel")) return rc;
  if (tracing) {   // debugging aid only: dumps CTA 0's event log (role, tag, cycles since the first event) to stderr
    static int dumps = 0;
    cudaStreamSynchronize(stream);
    if (dumps++ == atoi(getenv("CPOLY_TC_TRACE"))) {
      static long long h[5 * 4096];
      cudaMemcpy(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost);
      long long t0 = 0;
      for (int r = 0; r < 5; ++r) if (h[r * 4096] && (!t0 || h[r * 4096] < t0)) t0 = h[r * 4096];
      for (int r = 0; r < 5; ++r)
        for (int i = 0; i < 2048 && h[r * 4096 + 2 * i]; ++i) fprintf(stderr, "TRACE %d %lld %lld\n", r, h[r * 4096 + 2 * i + 1], h[r * 4096 + 2 * i] - t0);
    }
  }
  if (p.split > 1) return dcn_tc_launch_combine(p.partial, a.y.ptr, a.dtype, a.out_scale, a.bias, npix, a.Cout, p.split, a.relu, stream);
  return CPOLY_OK;
}

}  // namespace cpoly
