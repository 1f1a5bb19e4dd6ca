// DCNv2 forward on tcgen05 with a TMA-staged input WINDOW in shared memory (sm_100a).  Opt-in: CPOLY_TC_SAMPLER=window.
//
// Same contract as dcn_tc.cu (replaces the DCNv2 extension's `dcn_v2_forward` for the DeformConv layers of
// src/lib/models/networks/pose_dla_dcn.py:347-359; the extension is the un-vendored submodule
// src/lib/models/networks/DCNv2), different sampler and pipeline.  dcn_tc.cu gathers the four bilinear corners of every
// (pixel, tap) from global memory through L1.  Here the TMA engine (cp.async.bulk.tensor.4d, SASS UTMALDG) stages the
// tile's input neighbourhood ONCE per (tile, 128-byte channel chunk) in shared memory:
//
//     window = rows [ty0-1-h, ty0+8+1+h) x cols [tx0-1-h, tx0+16+1+h) x 128 B,   h = halo (3; less if smem is short)
//
// out-of-image elements are zero-filled by the TMA unit (borders need no predicates), and the samplers gather the corners
// with LDS.128 - conflict-free because the 8 lanes of a pixel read one whole 128-byte row.  A (pixel, tap) whose corners
// leave the window (|offset| >~ h) takes a predicated global-memory path, so any offset is still exact.
//
// Pipeline (what the event traces of tools/trace_summary.py led to):
//   warps 0-3   parameter + epilogue warps: thread t owns pixel t; turns the raw offsets/mask of its 9 taps into sampler
//               parameters one item AHEAD (double buffer, pfull/pempty), later drains row t of the accumulator
//               (tcgen05.ld -> scale/bias/ReLU -> NHWC stores).  The first item's parameters are computed by all 640
//               sampler + parameter threads together.
//   warp  4     TMA window loads (one window ahead, wfull/wempty) + bulk copies of the weight chunks (own B ring)
//   warp  5     MMA issuer.  Warps 4/5 walk their loops as whole warps and issue under elect.sync, so the single-thread
//               instructions keep uniform operands (a `lane == 0` branch makes ptxas wrap every tcgen05.mma / TMA copy
//               in an ELECT/R2UR/BRA.U.ANY loop - ~12 dependent instructions each).
//   warps 6-21  samplers: ONE rolling pipeline over all steps of the CTA (item-major, chunk-major, tap-minor) with no
//               per-item drain.  fence.proxy.async compiles to MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC, i.e. it waits for
//               every outstanding load of the thread: the loads of step n+1 are therefore issued AFTER the fence/arrive
//               of step n (LDS latency is short; the other sampler warps cover it).
// Small maps split the channel chunks (not the taps) of a tile over CTAs; fp32 partials are combined in fixed order.
#include <cuda.h>
#include <stdlib.h>

#include "dcn_sampler.cuh"

namespace cpoly {

// dcn_tc.cu
int dcn_tc_launch_combine(const float* partial, void* y, int dtype, const float* scale, const float* bias, int64_t npix,
                          int Cout, int split, int relu, cudaStream_t stream);

namespace {

using namespace tc;
using namespace smp;

constexpr int MAX_A_STAGES = 4;   // sampled-column (A operand) ring
constexpr int MAX_B_STAGES = 8;   // weight (B operand) ring: deeper, a bulk copy from L2 takes ~1 us to land
constexpr int EPI_WARPS = 4;
constexpr int SAMPLER_WARP0 = 6;
constexpr int SAMPLER_WARPS = 16;
constexpr int NS = SAMPLER_WARPS * 32;
constexpr int NTHREADS = (SAMPLER_WARP0 + SAMPLER_WARPS) * 32;
constexpr int ROWS_PER_PASS = NS / 8;
constexpr int PASSES = TILE_PX / ROWS_PER_PASS;
constexpr int A_STAGE_BYTES = TILE_PX * ROW_BYTES;

constexpr int SMEM_BAR_OFF = 0;
constexpr int SMEM_VEC_OFF = 1024;                                   // scale[256] + bias[256]
constexpr int SMEM_PO_OFF = 3072;                                        // 2 x (window offset | flag, y0:x0) per (tap, pixel)
constexpr int SMEM_PW_OFF = SMEM_PO_OFF + 2 * NUM_TAPS * TILE_PX * 8;    // 2 x 4 corner weights x mask per (tap, pixel)
constexpr int SMEM_WIN_OFF = SMEM_PW_OFF + 2 * NUM_TAPS * TILE_PX * 16;  // 58368: two windows, then the A and B rings
constexpr int SMEM_LIMIT = 226 * 1024;

struct WinParams {
  const void* x;
  cpoly_tensor4 offset, mask;
  const void* wpacked;
  const float* bias;
  const float* scale;
  const float* offset_bias;
  const float* mask_bias;
  void* y;
  float* partial;
  int B, Cin, Cout, H, W;
  int om_dtype, mask_is_logit, relu;
  int tiles_x, tiles_y, num_tiles, num_items;
  int split, chunks_per_item, nchunk;
  int halo, win_h, win_w, win_bytes;
  int a_off, b_off, a_stages, b_stages, b_stage_bytes;
  int acc_cols, tmem_cols;
  long long* trace;   // profiling only (CPOLY_TC_TRACE): clock64 event log of CTA 0, 4 roles x 4096 entries
  int wait_mode;   // profiling only (CPOLY_TC_WAIT): 1 = plain polling waits
};

__device__ __forceinline__ uint2 lds8(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts8(uint32_t addr, uint2 v) {
  asm volatile("st.shared.v2.b32 [%0], {%1,%2};" ::"r"(addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tmap, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
      : "memory");
}

template <typename T> struct OutConv;
template <> struct OutConv<__nv_bfloat16> {
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

__device__ __forceinline__ void item_coords(const WinParams& p, int item, int& b, int& ty0, int& tx0, int& chunk0) {
  const int tile = item / p.split;
  chunk0 = (item - tile * p.split) * p.chunks_per_item;
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  b = tile / tiles_per_img;
  const int trem = tile - b * tiles_per_img;
  const int tyi = trem / p.tiles_x;
  ty0 = tyi * TILE_H;
  tx0 = (trem - tyi * p.tiles_x) * TILE_W;
}

// Raw offsets / mask logit of tap k at pixel (y, x) of image b (biases of the producer conv added here).
__device__ __forceinline__ void load_raw(const WinParams& p, int b, int y, int x, int k, float& dy, float& dx, float& m) {
  const int64_t opix = (int64_t)b * p.offset.sb + (int64_t)y * p.offset.sh + (int64_t)x * p.offset.sw;
  dy = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k) * p.offset.sc, p.om_dtype);
  dx = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k + 1) * p.offset.sc, p.om_dtype);
  m = ld_dyn(p.mask.ptr, (int64_t)b * p.mask.sb + (int64_t)k * p.mask.sc + (int64_t)y * p.mask.sh + (int64_t)x * p.mask.sw, p.om_dtype);
}

// Sampler parameters of one (pixel, tap): po = (window byte offset of corner 00 | bit 0 = a corner leaves the window,
// y0:x0 as int16 pair), pw = the four bilinear weights x mask in the blend's format.  Dead sample: offset 0, weights 0.
template <typename T>
__device__ __forceinline__ void make_param(const WinParams& p, int y, int x, int wy0, int wx0, int k, float dy, float dx, float m,
                                           uint2& po, uint4& pw) {
  po = make_uint2(0u, 0u);
  pw = make_uint4(0u, 0u, 0u, 0u);
  if (p.offset_bias) { dy += __ldg(p.offset_bias + 2 * k); dx += __ldg(p.offset_bias + 2 * k + 1); }
  if (p.mask_bias) m += __ldg(p.mask_bias + k);
  if (p.mask_is_logit) m = __frcp_rn(1.0f + __expf(-m));
  const int ki = k / 3;
  const float py = (float)(y - 1 + ki) + dy;
  const float px = (float)(x - 1 + (k - 3 * ki)) + dx;
  if (py > -1.0f && px > -1.0f && py < (float)p.H && px < (float)p.W) {   // DCNv2's whole-sample rule
    const float fy = floorf(py), fx = floorf(px);
    const int y0 = (int)fy, x0 = (int)fx;   // in [-1, H-1] x [-1, W-1]
    const float ly = py - fy, lx = px - fx, hy = 1.0f - ly, hx = 1.0f - lx;
    const int ry = y0 - wy0, rx = x0 - wx0;
    // corners outside the IMAGE but inside the window are zero-filled by the TMA unit: no validity needed
    po.x = ((unsigned)ry <= (unsigned)(p.win_h - 2) && (unsigned)rx <= (unsigned)(p.win_w - 2))
               ? (uint32_t)(ry * p.win_w + rx) * ROW_BYTES
               : 1u;   // bit 0: a corner leaves the window -> global-memory path
    po.y = ((uint32_t)(y0 & 0xffff) << 16) | (uint32_t)(x0 & 0xffff);
    pw.x = Blend<T>::weight(hy * hx * m);
    pw.y = Blend<T>::weight(hy * lx * m);
    pw.z = Blend<T>::weight(ly * hx * m);
    pw.w = Blend<T>::weight(ly * lx * m);
  }
}

// The FIRST item's parameters are computed by samplers and parameter warps together (640 threads, <= 2 entries each):
// four warps alone need ~10k cycles for 1152 entries, which nothing hides at kernel start.
constexpr int FIRST_PARAM_THREADS = (EPI_WARPS + SAMPLER_WARPS) * 32;
template <typename T>
__device__ __forceinline__ void first_item_params(const WinParams& p, int tid, uint32_t po_s, uint32_t pw_s) {
  if ((int)blockIdx.x < p.num_items) {
    int b, ty0, tx0, chunk0;
    item_coords(p, blockIdx.x, b, ty0, tx0, chunk0);
    const int wy0 = ty0 - 1 - p.halo, wx0 = tx0 - 1 - p.halo;
    float dy[2], dx[2], m[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int e = tid + i * FIRST_PARAM_THREADS, k = e >> 7, r = e & (TILE_PX - 1);
      const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
      dy[i] = dx[i] = m[i] = 0.f;
      if (e < NUM_TAPS * TILE_PX && y < p.H && x < p.W) load_raw(p, b, y, x, k, dy[i], dx[i], m[i]);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int e = tid + i * FIRST_PARAM_THREADS, k = e >> 7, r = e & (TILE_PX - 1);
      const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
      if (e >= NUM_TAPS * TILE_PX) continue;
      uint2 po = make_uint2(0u, 0u);
      uint4 pw = make_uint4(0u, 0u, 0u, 0u);
      if (y < p.H && x < p.W) make_param<T>(p, y, x, wy0, wx0, k, dy[i], dx[i], m[i], po, pw);
      sts8(po_s + (uint32_t)e * 8u, po);
      sts16(pw_s + (uint32_t)e * 16u, pw);
    }
  }
  named_bar_sync(2, FIRST_PARAM_THREADS);
}

template <typename T>
__global__ void __maxnreg__(80) dcn_fwd_tc_win_kernel(const WinParams p, const __grid_constant__ CUtensorMap tmap) {
  constexpr bool kTF32 = sizeof(T) == 4;
  constexpr int kChunk = ROW_BYTES / (int)sizeof(T);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_fullA = smem_base + SMEM_BAR_OFF;         // [MAX_A_STAGES] samplers wrote the A tile
  const uint32_t bar_emptyA = bar_fullA + 8 * MAX_A_STAGES;    // [MAX_A_STAGES] the MMAs reading it retired
  const uint32_t bar_fullB = bar_emptyA + 8 * MAX_A_STAGES;    // [MAX_B_STAGES] weight chunk landed
  const uint32_t bar_emptyB = bar_fullB + 8 * MAX_B_STAGES;    // [MAX_B_STAGES]
  const uint32_t bar_tfull = bar_emptyB + 8 * MAX_B_STAGES;    // [2]
  const uint32_t bar_tempty = bar_tfull + 16;                  // [2]
  const uint32_t bar_wfull = bar_tempty + 16;                  // [2] window landed (TMA complete_tx)
  const uint32_t bar_wempty = bar_wfull + 16;                  // [2] every sampler warp is done reading the window
  const uint32_t bar_pfull = bar_wempty + 16;                  // [2] parameters of an item written (4 parameter warps)
  const uint32_t bar_pempty = bar_pfull + 16;                  // [2] every sampler warp is done with them
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SMEM_BAR_OFF + 8 * (2 * MAX_A_STAGES + 2 * MAX_B_STAGES + 12));
  float* scale_s = reinterpret_cast<float*>(smem + SMEM_VEC_OFF);
  float* bias_s = scale_s + 256;
  const uint32_t po_s = smem_base + SMEM_PO_OFF, pw_s = smem_base + SMEM_PW_OFF, win_s = smem_base + SMEM_WIN_OFF;
  const uint32_t a_ring = smem_base + (uint32_t)p.a_off, b_ring = smem_base + (uint32_t)p.b_off;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto wait = [&](uint32_t bar, uint32_t parity, uint32_t sleep_ns) {
    if (p.wait_mode) mbar_wait_poll(bar, parity); else mbar_wait(bar, parity, sleep_ns);
  };

  int tr_n = 0;
  auto tr = [&](int role, int tag) {
    if (p.trace && blockIdx.x == 0 && tr_n < 2048) { p.trace[role * 4096 + 2 * tr_n] = clock64(); p.trace[role * 4096 + 2 * tr_n + 1] = tag; ++tr_n; }
  };

  if (warp == 4 && lane == 0) {
    for (int s = 0; s < MAX_A_STAGES; ++s) { mbar_init(bar_fullA + 8 * s, SAMPLER_WARPS); mbar_init(bar_emptyA + 8 * s, 1); }
    for (int s = 0; s < MAX_B_STAGES; ++s) { mbar_init(bar_fullB + 8 * s, 1); mbar_init(bar_emptyB + 8 * s, 1); }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar_tfull + 8 * a, 1);
      mbar_init(bar_tempty + 8 * a, EPI_WARPS * 32);
      mbar_init(bar_wfull + 8 * a, 1);
      mbar_init(bar_wempty + 8 * a, SAMPLER_WARPS);
      mbar_init(bar_pfull + 8 * a, EPI_WARPS);
      mbar_init(bar_pempty + 8 * a, SAMPLER_WARPS);
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
  const int steps_per_item = p.chunks_per_item * NUM_TAPS;
  const uint32_t na = (uint32_t)p.a_stages, nb = (uint32_t)p.b_stages;

  if (warp >= SAMPLER_WARP0) {
    // =========================================================================== samplers
    // One rolling pipeline over ALL steps of this CTA (item-major, chunk-major, tap-minor): the corner loads of step
    // n+1 are issued while step n is blended, across chunk and item boundaries.  The (pixel, tap) parameters come from
    // the parameter/epilogue warps through a double buffer, the input windows from the TMA thread.
    const int st = threadIdx.x - SAMPLER_WARP0 * 32;
    const int row_in_pass = st >> 3, cg = st & 7;
    const uint32_t win_rowb = (uint32_t)p.win_w * ROW_BYTES;
    const uint32_t pixb = (uint32_t)p.Cin * (uint32_t)sizeof(T), img_rowb = (uint32_t)p.W * pixb;
    const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.x);
    RingCursor rc;                 // A ring
    RingCursor w_issue, w_done;    // window buffers: issue side (one step ahead) / blend side
    RingCursor p_issue, p_done;    // parameter buffers, likewise
    uint4 c[PASSES][4] = {}, pw[PASSES] = {};
    uint32_t img_base = 0, po_cur = po_s, pw_cur = pw_s;
    auto issue = [&](int tap, int chunk, uint32_t wbuf, int pass) {   // parameters + four corner loads of one (step, pass)
      const uint32_t e = (uint32_t)(tap * TILE_PX + pass * ROWS_PER_PASS + row_in_pass);
      const uint2 po = lds8(po_cur + e * 8u);
      pw[pass] = lds16(pw_cur + e * 16u);
      if (!(po.x & 1u)) {
        const uint32_t a = win_s + wbuf * (uint32_t)p.win_bytes + po.x + (uint32_t)(cg * 16);
        c[pass][0] = lds16(a);
        c[pass][1] = lds16(a + ROW_BYTES);
        c[pass][2] = lds16(a + win_rowb);
        c[pass][3] = lds16(a + win_rowb + ROW_BYTES);
      } else {
        const int y0 = (int)(short)(po.y >> 16), x0 = (int)(short)(po.y & 0xffffu);
        const bool vy0 = (unsigned)y0 < (unsigned)p.H, vy1 = (unsigned)(y0 + 1) < (unsigned)p.H;
        const bool vx0 = (unsigned)x0 < (unsigned)p.W, vx1 = (unsigned)(x0 + 1) < (unsigned)p.W;
        // 32-bit byte offsets (they wrap for dead corners, exactly like the additions below; dead corners are not loaded)
        const uint32_t o00 = img_base + (uint32_t)(y0 * p.W + x0) * pixb + (uint32_t)(chunk * ROW_BYTES + cg * 16);
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        c[pass][0] = (vy0 && vx0) ? ldg16(xb + o00) : z;
        c[pass][1] = (vy0 && vx1) ? ldg16(xb + (o00 + pixb)) : z;
        c[pass][2] = (vy1 && vx0) ? ldg16(xb + (o00 + img_rowb)) : z;
        c[pass][3] = (vy1 && vx1) ? ldg16(xb + (o00 + img_rowb + pixb)) : z;
      }
    };
    int n_item = blockIdx.x, n_tap = 0, n_chunk = 0, n_chunk_end = 0;   // the step whose loads were issued last
    auto enter_item = [&](int item) {   // issue side switches to a new item: its parameters and first window must have landed
      int b, ty0, tx0, chunk0;
      item_coords(p, item, b, ty0, tx0, chunk0);
      img_base = (uint32_t)(b * p.H * p.W) * pixb;   // bytes < 2^32 (checked on the host)
      n_tap = 0; n_chunk = chunk0; n_chunk_end = chunk0 + p.chunks_per_item;
      po_cur = po_s + p_issue.s * (uint32_t)(NUM_TAPS * TILE_PX * 8);
      pw_cur = pw_s + p_issue.s * (uint32_t)(NUM_TAPS * TILE_PX * 16);
      wait(bar_pfull + 8 * p_issue.s, p_issue.ph, 32);
      wait(bar_wfull + 8 * w_issue.s, w_issue.ph, 32);
    };
    first_item_params<T>(p, EPI_WARPS * 32 + st, po_s, pw_s);   // with the parameter warps; buffer 0
    bool more = n_item < p.num_items;
    if (more) {
      enter_item(n_item);
#pragma unroll
      for (int pass = 0; pass < PASSES; ++pass) issue(n_tap, n_chunk, w_issue.s, pass);
    }
    // fence.proxy.async compiles to MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC: it waits for EVERY outstanding access of the
    // thread, loads included.  The loads of the next step are therefore issued after the fence/arrive of this one (not
    // between its blends, where the barrier would expose their full latency every step); LDS latency (~30 cycles) is
    // covered by the other sampler warps of the scheduler.
    while (more) {
      // flags of the step being blended now (= the one issued last)
      const bool last_tap = n_tap == NUM_TAPS - 1;
      const bool last_of_item = last_tap && (n_chunk + 1 == n_chunk_end);
      wait(bar_emptyA + 8 * rc.s, rc.ph ^ 1u, 32);
      if (st == 0) tr(0, 1);
      const uint32_t a_base = a_ring + rc.s * (uint32_t)A_STAGE_BYTES;
#pragma unroll
      for (int pass = 0; pass < PASSES; ++pass) {
        const int r = pass * ROWS_PER_PASS + row_in_pass;
        sts16(a_base + r * ROW_BYTES + ((cg ^ (r & 7)) << 4), Blend<T>::run(c[pass], pw[pass]));
      }
      fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(bar_fullA + 8 * rc.s);
        if (last_tap) mbar_arrive(bar_wempty + 8 * w_done.s);     // this warp's reads of the window have all landed
        if (last_of_item) mbar_arrive(bar_pempty + 8 * p_done.s); // ... and of the item's parameters
      }
      if (st == 0) tr(0, 2);
      if (last_tap) w_done.advance(2u);
      if (last_of_item) p_done.advance(2u);
      rc.advance(na);
      // advance the issue cursor by one step and issue its loads
      if (!last_tap) {
        ++n_tap;
      } else if (!last_of_item) {
        n_tap = 0; ++n_chunk;
        w_issue.advance(2u);
        wait(bar_wfull + 8 * w_issue.s, w_issue.ph, 32);
      } else {
        n_item += gridDim.x;
        more = n_item < p.num_items;
        w_issue.advance(2u);
        p_issue.advance(2u);
        if (more) enter_item(n_item);
      }
      if (more) {
#pragma unroll
        for (int pass = 0; pass < PASSES; ++pass) issue(n_tap, n_chunk, w_issue.s, pass);
      }
    }
  } else if (warp == 4) {
    // =========================================================================== window + weight loader
    RingCursor rc, wc;
    const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.wpacked);
    auto load_window = [&](int item, int cc) {
      int b, ty0, tx0, chunk0;
      item_coords(p, item, b, ty0, tx0, chunk0);
      wait(bar_wempty + 8 * wc.s, wc.ph ^ 1u, 20);
      if (elect_one()) {
        mbar_arrive_expect_tx(bar_wfull + 8 * wc.s, (uint32_t)p.win_bytes);
        tma_load_4d(win_s + wc.s * (uint32_t)p.win_bytes, &tmap, (chunk0 + cc) * kChunk, tx0 - 1 - p.halo, ty0 - 1 - p.halo, b,
                    bar_wfull + 8 * wc.s);
      }
      __syncwarp();
      wc.advance(2u);
    };
    if ((int)blockIdx.x < p.num_items) load_window(blockIdx.x, 0);
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      const int chunk0 = (item % p.split) * p.chunks_per_item;
      for (int cc = 0; cc < p.chunks_per_item; ++cc) {
        // the window one ahead, before this chunk's weights (its buffer was released two windows ago)
        if (cc + 1 < p.chunks_per_item) load_window(item, cc + 1);
        else if (item + (int)gridDim.x < p.num_items) load_window(item + gridDim.x, 0);
        for (int tap = 0; tap < NUM_TAPS; ++tap) {
          wait(bar_emptyB + 8 * rc.s, rc.ph ^ 1u, 20);
          if (elect_one()) {
            mbar_arrive_expect_tx(bar_fullB + 8 * rc.s, (uint32_t)p.b_stage_bytes);
            bulk_copy_g2s(b_ring + rc.s * (uint32_t)p.b_stage_bytes, wsrc + (size_t)(tap * p.nchunk + chunk0 + cc) * p.b_stage_bytes,
                          (uint32_t)p.b_stage_bytes, bar_fullB + 8 * rc.s);
          }
          __syncwarp();
          rc.advance(nb);
        }
      }
    }
  } else if (warp == 5) {
    // =========================================================================== MMA issuer
    // The whole warp walks the pipeline (uniform control flow and operands); one elected lane issues.
    const uint32_t idesc = umma_idesc(kTF32, TILE_PX, p.Cout);
    RingCursor ra, rb;
    uint32_t tcount = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x, ++tcount) {
      const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
      wait(bar_tempty + 8 * acc, accph ^ 1u, 20);   // epilogue drained this accumulator
      if (lane == 0) tr(2, 10);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * (uint32_t)p.acc_cols;
      for (int q = 0; q < steps_per_item; ++q) {
        wait(bar_fullB + 8 * rb.s, rb.ph, 20);
        wait(bar_fullA + 8 * ra.s, ra.ph, 20);
        if (lane == 0) tr(2, 2);
        tc_fence_after();
        const uint64_t adesc = umma_desc_sw128(a_ring + ra.s * (uint32_t)A_STAGE_BYTES);
        const uint64_t bdesc = umma_desc_sw128(b_ring + rb.s * (uint32_t)p.b_stage_bytes);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // 4 x 32 bytes of K per 128-byte row
            umma<kTF32>(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (q | kk) != 0 ? 1u : 0u);
          umma_commit(bar_emptyA + 8 * ra.s);   // both rings are released when these MMAs retire
          umma_commit(bar_emptyB + 8 * rb.s);
          if (q == steps_per_item - 1) umma_commit(bar_tfull + 8 * acc);   // accumulator complete -> epilogue
        }
        __syncwarp();
        if (lane == 0) tr(2, 3);
        ra.advance(na);
        rb.advance(nb);
      }
    }
  } else {
    // =========================================================================== parameter + epilogue warps (0-3)
    // Thread t owns pixel t of the tile: it turns the raw offsets / mask of its 9 taps into sampler parameters one
    // item AHEAD of the samplers (double buffer), and later drains row t of the item's accumulator.
    T* yg = reinterpret_cast<T*>(p.y);
    const int64_t npix = (int64_t)p.B * p.H * p.W;
    const int t = warp * 32 + lane;
    RingCursor pc;
    auto make_params = [&](int item) {
      int b, ty0, tx0, chunk0;
      item_coords(p, item, b, ty0, tx0, chunk0);
      const int wy0 = ty0 - 1 - p.halo, wx0 = tx0 - 1 - p.halo;
      const int y = ty0 + (t >> 4), x = tx0 + (t & (TILE_W - 1));
      const bool inside = y < p.H && x < p.W;
      float rdy[NUM_TAPS], rdx[NUM_TAPS], rm[NUM_TAPS];
#pragma unroll
      for (int k = 0; k < NUM_TAPS; ++k) {   // all 27 loads in flight before the first use
        rdy[k] = rdx[k] = rm[k] = 0.f;
        if (inside) load_raw(p, b, y, x, k, rdy[k], rdx[k], rm[k]);
      }
      wait(bar_pempty + 8 * pc.s, pc.ph ^ 1u, 64);   // the samplers are done with this buffer's previous item
      const uint32_t po_b = po_s + pc.s * (uint32_t)(NUM_TAPS * TILE_PX * 8), pw_b = pw_s + pc.s * (uint32_t)(NUM_TAPS * TILE_PX * 16);
#pragma unroll
      for (int k = 0; k < NUM_TAPS; ++k) {
        uint2 po = make_uint2(0u, 0u);
        uint4 pw = make_uint4(0u, 0u, 0u, 0u);
        if (inside) make_param<T>(p, y, x, wy0, wx0, k, rdy[k], rdx[k], rm[k], po, pw);
        sts8(po_b + (uint32_t)(k * TILE_PX + t) * 8u, po);
        sts16(pw_b + (uint32_t)(k * TILE_PX + t) * 16u, pw);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_pfull + 8 * pc.s);
      pc.advance(2u);
    };
    uint32_t tcount = 0;
    first_item_params<T>(p, t, po_s, pw_s);   // with the samplers; buffer 0
    if ((int)blockIdx.x < p.num_items) {
      if (lane == 0) mbar_arrive(bar_pfull);
      pc.advance(2u);
    }
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x, ++tcount) {
      if (item + (int)gridDim.x < p.num_items) make_params(item + gridDim.x);
      const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
      int b, ty0, tx0, chunk0;
      item_coords(p, item, b, ty0, tx0, chunk0);
      const int y = ty0 + (t >> 4), x = tx0 + (t & (TILE_W - 1));
      const bool valid = (y < p.H) && (x < p.W);
      const int64_t pix = ((int64_t)b * p.H + y) * p.W + x;
      wait(bar_tfull + 8 * acc, accph, 256);
      if (threadIdx.x == 0) tr(3, 1);
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
      if (threadIdx.x == 0) tr(3, 2);
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}

// ---- host side ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<EncodeTiledFn>(sym);
  }();
  return fn;
}

struct WinPlan {
  bool ok;
  int halo, win_h, win_w, win_bytes, a_off, b_off, a_stages, b_stages, smem_bytes;
  int nchunk, split, chunks_per_item, num_tiles;
  size_t partial_bytes;
};

int pow2_at_least(int v) {
  int p = 32;
  while (p < v) p <<= 1;
  return p;
}

WinPlan plan(const cpoly_dcn_fwd_args& a) {
  WinPlan w{};
  const int esize = a.dtype == CPOLY_BF16 ? 2 : 4;
  w.nchunk = a.Cin / (ROW_BYTES / esize);
  const int b_stage = a.Cout * ROW_BYTES;
  int max_halo = 4;
  if (const char* e = getenv("CPOLY_TC_HALO")) {   // profiling knob
    const int v = atoi(e);
    if (v >= 1 && v <= 6) max_halo = v;
  }
  // Shared-memory budget: two windows, an A ring of >= 2 tiles and a B ring as deep as possible (a weight chunk takes
  // ~1 us to arrive from L2, a step is meant to take a fraction of that).  Prefer the widest halo that still leaves
  // >= 4 weight stages; never go below halo 3 unless nothing else fits.
  auto fit = [&](int h, int min_b) {
    const int wh = TILE_H + 2 + 2 * h, ww = TILE_W + 2 + 2 * h;
    const int wb = wh * ww * ROW_BYTES;
    const int off = ((SMEM_WIN_OFF + 2 * wb + 1023) / 1024) * 1024;
    int na = 2;
    if (const char* e = getenv("CPOLY_TC_NA")) {   // profiling knob: depth of the A ring
      const int v = atoi(e);
      if (v >= 2 && v <= MAX_A_STAGES) na = v;
    }
    int nb = (SMEM_LIMIT - off - na * A_STAGE_BYTES) / b_stage;
    if (nb < min_b) return false;
    if (nb > MAX_B_STAGES) nb = MAX_B_STAGES;
    na += (SMEM_LIMIT - off - na * A_STAGE_BYTES - nb * b_stage) / A_STAGE_BYTES;
    if (na > MAX_A_STAGES) na = MAX_A_STAGES;
    w.ok = true; w.halo = h; w.win_h = wh; w.win_w = ww; w.win_bytes = wb;
    w.a_off = off; w.a_stages = na; w.b_off = off + na * A_STAGE_BYTES; w.b_stages = nb;
    w.smem_bytes = w.b_off + nb * b_stage + 1024;
    return true;
  };
  for (int h = max_halo; h >= 3 && !w.ok; --h) fit(h, 4);
  for (int h = max_halo < 3 ? max_halo : 3; h >= 1 && !w.ok; --h) fit(h, 2);
  w.num_tiles = a.B * ((a.W + TILE_W - 1) / TILE_W) * ((a.H + TILE_H - 1) / TILE_H);
  // small maps: split the channel chunks of a tile over CTAs (fp32 partial sums, fixed-order combine)
  w.split = 1;
  if (w.num_tiles < 148) {
    w.split = w.nchunk;
    for (int d = 2; d < w.nchunk; ++d)
      if (w.nchunk % d == 0 && w.num_tiles * d >= 148) { w.split = d; break; }
  }
  w.chunks_per_item = w.nchunk / w.split;
  w.partial_bytes = w.split > 1 ? (size_t)w.split * (size_t)a.B * a.H * a.W * a.Cout * sizeof(float) : 0;
  return w;
}

}  // namespace

bool dcn_tc_win_eligible(const cpoly_dcn_fwd_args& a) {
  if (!dcn_tc_eligible(a)) return false;
  if (a.H > 32000 || a.W > 32000) return false;   // (y0, x0) travel as int16
  if (!encode_tiled_fn()) return false;
  return plan(a).ok;
}

size_t dcn_tc_win_forward_workspace(const cpoly_dcn_fwd_args& a) { return plan(a).partial_bytes + 256; }

int dcn_forward_tc_win(const cpoly_dcn_fwd_args& a, cudaStream_t stream) {
  WinPlan w = plan(a);
  CPOLY_REQUIRE(w.ok, CPOLY_EUNSUPPORTED, "window plan does not fit in shared memory");
  if (w.num_tiles == 0) return CPOLY_OK;
  const int esize = a.dtype == CPOLY_BF16 ? 2 : 4;
  const int64_t npix = (int64_t)a.B * a.H * a.W;
  if (w.split > 1 && (!a.workspace || a.workspace_bytes < w.partial_bytes || (reinterpret_cast<uintptr_t>(a.workspace) & 15))) {
    w.split = 1; w.chunks_per_item = w.nchunk;   // no scratch supplied: unsplit schedule (same result, fewer busy SMs)
  }
  WinParams p;
  p.x = a.x.ptr; p.offset = a.offset; p.mask = a.mask; p.wpacked = a.weight_packed; p.bias = a.bias; p.scale = a.out_scale;
  p.offset_bias = a.offset_bias; p.mask_bias = a.mask_bias;
  p.y = a.y.ptr;
  p.partial = w.split > 1 ? reinterpret_cast<float*>(a.workspace) : nullptr;
  p.B = a.B; p.Cin = a.Cin; p.Cout = a.Cout; p.H = a.H; p.W = a.W;
  p.om_dtype = a.om_dtype; p.mask_is_logit = a.mask_is_logit; p.relu = a.relu;
  p.tiles_x = (a.W + TILE_W - 1) / TILE_W; p.tiles_y = (a.H + TILE_H - 1) / TILE_H;
  p.num_tiles = w.num_tiles; p.split = w.split; p.chunks_per_item = w.chunks_per_item; p.nchunk = w.nchunk;
  p.num_items = w.num_tiles * w.split;
  p.halo = w.halo; p.win_h = w.win_h; p.win_w = w.win_w; p.win_bytes = w.win_bytes;
  p.a_off = w.a_off; p.b_off = w.b_off; p.a_stages = w.a_stages; p.b_stages = w.b_stages;
  p.b_stage_bytes = a.Cout * ROW_BYTES;
  p.acc_cols = a.Cout; p.tmem_cols = pow2_at_least(2 * a.Cout);
  p.wait_mode = 0; p.trace = nullptr;
  static long long* trace_buf = nullptr;
  const bool tracing = getenv("CPOLY_TC_TRACE") != nullptr;
  if (tracing) {
    if (!trace_buf) cudaMalloc(&trace_buf, 4 * 4096 * sizeof(long long));
    cudaMemsetAsync(trace_buf, 0, 4 * 4096 * sizeof(long long), stream);
    p.trace = trace_buf;
  }
  if (const char* e = getenv("CPOLY_TC_WAIT")) p.wait_mode = atoi(e);

  // tensor map over x viewed as (B, H, W, Cin), innermost first; box = one window of one 128-byte channel chunk
  CUtensorMap tmap;
  const cuuint64_t dims[4] = {(cuuint64_t)a.Cin, (cuuint64_t)a.W, (cuuint64_t)a.H, (cuuint64_t)a.B};
  const cuuint64_t strides[3] = {(cuuint64_t)a.Cin * esize, (cuuint64_t)a.W * a.Cin * esize, (cuuint64_t)a.H * a.W * a.Cin * esize};
  const cuuint32_t box[4] = {(cuuint32_t)(ROW_BYTES / esize), (cuuint32_t)w.win_w, (cuuint32_t)w.win_h, 1u};
  const cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
  const CUresult cr = encode_tiled_fn()(&tmap, a.dtype == CPOLY_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                                        a.x.ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CPOLY_REQUIRE(cr == CUDA_SUCCESS, CPOLY_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)cr);

  int dev = 0, sms = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.num_items < sms ? p.num_items : sms;
  if (a.dtype == CPOLY_BF16) {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_fwd_tc_win_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, w.smem_bytes));
    dcn_fwd_tc_win_kernel<__nv_bfloat16><<<grid, NTHREADS, w.smem_bytes, stream>>>(p, tmap);
  } else {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_fwd_tc_win_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, w.smem_bytes));
    dcn_fwd_tc_win_kernel<float><<<grid, NTHREADS, w.smem_bytes, stream>>>(p, tmap);
  }
  if (int rc = check_launch("dcn_fwd_tc_win_kernel")) return rc;
  if (tracing) {   // debugging aid only: dumps CTA 0's event log (role, tag, cycles since the first event) to stderr
    static int dumps = 0;
    cudaStreamSynchronize(stream);
    if (dumps++ == atoi(getenv("CPOLY_TC_TRACE"))) {
      static long long h[4 * 4096];
      cudaMemcpy(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost);
      long long t0 = 0;
      for (int r = 0; r < 4; ++r) if (h[r * 4096] && (!t0 || h[r * 4096] < t0)) t0 = h[r * 4096];
      for (int r = 0; r < 4; ++r)
        for (int i = 0; i < 2048 && h[r * 4096 + 2 * i]; ++i) fprintf(stderr, "TRACE %d %lld %lld\n", r, h[r * 4096 + 2 * i + 1], h[r * 4096 + 2 * i] - t0);
    }
  }
  if (w.split > 1) return dcn_tc_launch_combine(p.partial, a.y.ptr, a.dtype, a.out_scale, a.bias, npix, a.Cout, w.split, a.relu, stream);
  return CPOLY_OK;
}

}  // namespace cpoly
