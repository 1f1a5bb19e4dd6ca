// DCNv2 forward on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// Replaces the DCNv2 extension's `dcn_v2_forward` for the 16 DeformConv layers of the DLA-34 neck
// (reference call site src/lib/models/networks/pose_dla_dcn.py:354-359; extension = missing submodule
// src/lib/models/networks/DCNv2).  Upstream materialises columns[9*Cin, H*W] in HBM (9x the input)
// and calls cuBLAS SGEMM; here the column tile never leaves the SM.
//
//   out[px, o] = bias[o] + sum_{tap, c} A[px, (tap, c)] * W[o, c, tap],   A = mask * bilinear(x, p_tap)
//
// Persistent, warp-specialised CTA (one per SM), 128-pixel tiles (8 rows x 16 columns of the image so
// that neighbouring pixels share bilinear corners in L1):
//
//   warps 0-3   epilogue   tcgen05.ld accumulator rows -> +bias -> bf16/fp32 -> NHWC global stores
//   warp  4     B loader   one thread: cp.async.bulk (TMA engine, UBLKCP) of pre-swizzled weight chunks
//   warp  5     MMA        TMEM alloc; one thread issues tcgen05.mma (M=128, N=Cout, K=16 bf16 / 8 tf32)
//   warps 6..   samplers   per tile: stage (corner offsets, mask*bilinear weights) of all 9 taps in smem,
//                          then per (tap, 128-byte channel chunk) gather 4 corners with 16-byte loads, blend in
//                          fp32 and write the A tile straight into the 128B-swizzled K-major UMMA layout
//
// Pipelines: smem ring (fullA/fullB/empty mbarriers) between samplers+loader and the MMA thread; two TMEM
// accumulators (tmem_full/tmem_empty) between the MMA thread and the epilogue, so the epilogue of tile t
// overlaps the MMAs of tile t+1.  Every mbarrier wait has a clock-based timeout that traps instead of
// hanging the GPU.
#include "common.cuh"

namespace cpoly {
namespace {

constexpr int TILE_H = 8, TILE_W = 16, TILE_PX = TILE_H * TILE_W;  // 128 = UMMA M
constexpr int NUM_TAPS = 9;
constexpr int ROW_BYTES = 128;                 // one swizzle row: 64 bf16 or 32 fp32 channels
constexpr int A_STAGE_BYTES = TILE_PX * ROW_BYTES;  // 16 KB
constexpr int MAX_STAGES = 6;
constexpr int EPI_WARPS = 4;
constexpr int SAMPLER_WARP0 = 6;
constexpr int SAMPLER_WARPS = 16;
constexpr int NS = SAMPLER_WARPS * 32;         // sampler threads
constexpr int NTHREADS = (SAMPLER_WARP0 + SAMPLER_WARPS) * 32;
constexpr int ROWS_PER_PASS = NS / 8;          // 8 lanes x 16 B cover one 128-byte row
constexpr int PASSES = TILE_PX / ROWS_PER_PASS;
constexpr long long WAIT_TIMEOUT_CYCLES = 4000000000LL;  // ~2 s: trap instead of hanging the box

// shared-memory map (dynamic, 1024-byte aligned)
constexpr int SMEM_BAR_OFF = 0;        // mbarriers + tmem pointer
constexpr int SMEM_BIAS_OFF = 1024;    // Cout floats (<= 256)
constexpr int SMEM_PARAM_OFF = 2048;   // NUM_TAPS * TILE_PX * 32 B
constexpr int SMEM_PARAM_BYTES = NUM_TAPS * TILE_PX * 32;
constexpr int SMEM_STAGE_OFF = ((SMEM_PARAM_OFF + SMEM_PARAM_BYTES + 1023) / 1024) * 1024;
constexpr int SMEM_LIMIT = 226 * 1024;  // 227 KB per CTA minus the alignment slack

struct TapParam {   // 32 bytes: what a sampler needs for one (pixel, tap)
  int off[4];       // element offset of the corner's channel vector in x (dead corner -> 0, weight 0)
  float w[4];       // mask * bilinear weight
};

struct TcParams {
  const void* x;
  cpoly_tensor4 offset, mask;
  const void* wpacked;
  const float* bias;
  void* y;
  int B, Cin, Cout, H, W;
  int om_dtype, mask_is_logit;
  int tiles_x, tiles_y, num_tiles;
  int nchunk;      // channel chunks per tap
  int stages, b_stage_bytes, stage_bytes;
  int acc_cols, tmem_cols;
};

// ------------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok;
}
__device__ __noinline__ void mbar_timeout(uint32_t bar, uint32_t parity) {
  printf("cpoly dcn_tc: mbarrier wait timed out (block %d thread %d bar 0x%x parity %u)\n", blockIdx.x, threadIdx.x, bar, parity);
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > WAIT_TIMEOUT_CYCLES) mbar_timeout(bar, parity);
  }
}
__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// K-major, 128-byte-swizzled operand tile: rows of 128 B, 8-row groups 1024 B apart (SBO), version 1.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffffu) >> 4);
  d |= (uint64_t)1 << 16;            // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;  // stride byte offset
  d |= (uint64_t)1 << 46;            // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;            // SWIZZLE_128B
  return d;
}

template <bool kTF32>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if (kTF32)
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void sts16(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t to_tf32(float f) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(f));
  return r;
}

// ---- blend of the four corner vectors (16 B each) into one 16-B piece of the A row ----------------
template <typename T> struct Blend;
template <> struct Blend<__nv_bfloat16> {
  static constexpr int kVec = 8;  // channels per 16 bytes
  __device__ static __forceinline__ uint4 run(const uint4 (&c)[4], const float (&w)[4]) {
    uint4 out;
    const uint32_t* cw[4] = {&c[0].x, &c[1].x, &c[2].x, &c[3].x};
    uint32_t* o = &out.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float lo = 0.f, hi = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t word = cw[q][j];
        lo = fmaf(w[q], __uint_as_float(word << 16), lo);
        hi = fmaf(w[q], __uint_as_float(word & 0xffff0000u), hi);
      }
      const __nv_bfloat162 pk = __floats2bfloat162_rn(lo, hi);
      o[j] = *reinterpret_cast<const uint32_t*>(&pk);
    }
    return out;
  }
};
template <> struct Blend<float> {
  static constexpr int kVec = 4;
  __device__ static __forceinline__ uint4 run(const uint4 (&c)[4], const float (&w)[4]) {
    uint4 out;
    const uint32_t* cw[4] = {&c[0].x, &c[1].x, &c[2].x, &c[3].x};
    uint32_t* o = &out.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) v = fmaf(w[q], __uint_as_float(cw[q][j]), v);
      o[j] = to_tf32(v);
    }
    return out;
  }
};

template <typename T> struct OutConv;
template <> struct OutConv<__nv_bfloat16> {
  // 16 fp32 accumulators -> 32 bytes
  __device__ static __forceinline__ void store(__nv_bfloat16* dst, const float (&v)[16]) {
    uint4 a, b;
    uint32_t* pa = &a.x;
    uint32_t* pb = &b.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __nv_bfloat162 p0 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
      const __nv_bfloat162 p1 = __floats2bfloat162_rn(v[8 + 2 * j], v[8 + 2 * j + 1]);
      pa[j] = *reinterpret_cast<const uint32_t*>(&p0);
      pb[j] = *reinterpret_cast<const uint32_t*>(&p1);
    }
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

// ---------------------------------------------------------------------------------------- kernel
template <typename T>
__global__ void __maxnreg__(80) dcn_fwd_tc_kernel(const TcParams p) {
  constexpr bool kTF32 = sizeof(T) == 4;
  constexpr int kVec = Blend<T>::kVec;       // channels per 16 B
  constexpr int kChunk = ROW_BYTES / sizeof(T);  // channels per K chunk (64 bf16 / 32 fp32)
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operand tiles need 1024-byte alignment: align the window ourselves (+1 KB is requested)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t smem_base = smem_u32(smem);
  // barrier indices
  const uint32_t bar_fullA = smem_base + SMEM_BAR_OFF;                 // [MAX_STAGES]
  const uint32_t bar_fullB = bar_fullA + 8 * MAX_STAGES;               // [MAX_STAGES]
  const uint32_t bar_empty = bar_fullB + 8 * MAX_STAGES;               // [MAX_STAGES]
  const uint32_t bar_tfull = bar_empty + 8 * MAX_STAGES;               // [2]
  const uint32_t bar_tempty = bar_tfull + 16;                          // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SMEM_BAR_OFF + 8 * (3 * MAX_STAGES + 4));
  float* bias_s = reinterpret_cast<float*>(smem + SMEM_BIAS_OFF);
  TapParam* params = reinterpret_cast<TapParam*>(smem + SMEM_PARAM_OFF);

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
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"((uint32_t)p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < p.Cout; i += NTHREADS) bias_s[i] = p.bias ? __ldg(p.bias + i) : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int chunks_per_tile = NUM_TAPS * p.nchunk;
  const int tiles_per_img = p.tiles_x * p.tiles_y;

  if (warp >= SAMPLER_WARP0) {
    // =========================================================================== samplers
    const int st = threadIdx.x - SAMPLER_WARP0 * 32;
    const int row_in_pass = st >> 3, cg = st & 7;
    const T* xg = reinterpret_cast<const T*>(p.x);
    constexpr int ITEMS = (NUM_TAPS * TILE_PX + NS - 1) / NS;  // (pixel, tap) items per thread and tile
    float raw_dy[ITEMS], raw_dx[ITEMS], raw_m[ITEMS];
    // raw offsets / mask of one tile -> registers (issued one tile ahead so the latency hides behind sampling)
    auto fetch_raw = [&](int tile) {
      const int b = tile / tiles_per_img;
      const int trem = tile - b * tiles_per_img;
      const int ty0 = (trem / p.tiles_x) * TILE_H, tx0 = (trem % p.tiles_x) * TILE_W;
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        const int item = st + i * NS;
        const int k = item / TILE_PX, r = item - k * TILE_PX;
        const int y = ty0 + r / TILE_W, x = tx0 + r % TILE_W;
        raw_dy[i] = raw_dx[i] = raw_m[i] = 0.f;
        if (item < NUM_TAPS * TILE_PX && y < p.H && x < p.W) {
          const int64_t opix = (int64_t)b * p.offset.sb + (int64_t)y * p.offset.sh + (int64_t)x * p.offset.sw;
          raw_dy[i] = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k) * p.offset.sc, p.om_dtype);
          raw_dx[i] = ld_dyn(p.offset.ptr, opix + (int64_t)(2 * k + 1) * p.offset.sc, p.om_dtype);
          raw_m[i] = ld_dyn(p.mask.ptr, (int64_t)b * p.mask.sb + (int64_t)k * p.mask.sc + (int64_t)y * p.mask.sh + (int64_t)x * p.mask.sw, p.om_dtype);
        }
      }
    };
    uint32_t it = 0;
    if ((int)blockIdx.x < p.num_tiles) fetch_raw(blockIdx.x);
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      const int b = tile / tiles_per_img;
      const int trem = tile - b * tiles_per_img;
      const int ty0 = (trem / p.tiles_x) * TILE_H, tx0 = (trem % p.tiles_x) * TILE_W;
      named_bar_sync(1, NS);  // everyone finished reading the previous tile's params
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        const int item = st + i * NS;
        if (item >= NUM_TAPS * TILE_PX) break;
        const int k = item / TILE_PX, r = item - k * TILE_PX;
        const int y = ty0 + r / TILE_W, x = tx0 + r % TILE_W;
        TapParam tp;
#pragma unroll
        for (int q = 0; q < 4; ++q) { tp.off[q] = 0; tp.w[q] = 0.f; }
        if (y < p.H && x < p.W) {
          const float m = p.mask_is_logit ? sigmoidf_accurate(raw_m[i]) : raw_m[i];
          const float py = (float)(y - 1 + k / 3) + raw_dy[i];
          const float px = (float)(x - 1 + k % 3) + raw_dx[i];
          const TapSample t = make_tap(py, px, p.H, p.W, (int64_t)b * p.H * p.W * p.Cin, (int64_t)p.W * p.Cin, (int64_t)p.Cin);
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (t.off[q] >= 0) { tp.off[q] = (int)t.off[q]; tp.w[q] = t.w[q] * m; }
        }
        uint4* dst = reinterpret_cast<uint4*>(&params[k * TILE_PX + r]);
        dst[0] = make_uint4((uint32_t)tp.off[0], (uint32_t)tp.off[1], (uint32_t)tp.off[2], (uint32_t)tp.off[3]);
        dst[1] = make_uint4(__float_as_uint(tp.w[0]), __float_as_uint(tp.w[1]), __float_as_uint(tp.w[2]), __float_as_uint(tp.w[3]));
      }
      named_bar_sync(1, NS);
      if (tile + (int)gridDim.x < p.num_tiles) fetch_raw(tile + gridDim.x);
      for (int k = 0; k < NUM_TAPS; ++k) {
        for (int cc = 0; cc < p.nchunk; ++cc, ++it) {
          const uint32_t s = it % (uint32_t)p.stages, ph = (it / (uint32_t)p.stages) & 1u;
          const int chan = cc * kChunk + cg * kVec;
          // phase 1: every corner load of this thread's rows goes out before anything is consumed
          uint4 c[PASSES][4];
          uint4 pw[PASSES];
#pragma unroll
          for (int pass = 0; pass < PASSES; ++pass) {
            const int r = pass * ROWS_PER_PASS + row_in_pass;
            const uint4* pp = reinterpret_cast<const uint4*>(&params[k * TILE_PX + r]);
            const uint4 po = pp[0];
            pw[pass] = pp[1];
            c[pass][0] = ldg16(xg + (int)po.x + chan);
            c[pass][1] = ldg16(xg + (int)po.y + chan);
            c[pass][2] = ldg16(xg + (int)po.z + chan);
            c[pass][3] = ldg16(xg + (int)po.w + chan);
          }
          mbar_wait(bar_empty + 8 * s, ph ^ 1u);
          const uint32_t a_base = smem_base + SMEM_STAGE_OFF + s * p.stage_bytes;
          // phase 2: blend and write the swizzled A rows
#pragma unroll
          for (int pass = 0; pass < PASSES; ++pass) {
            const int r = pass * ROWS_PER_PASS + row_in_pass;
            const float w[4] = {__uint_as_float(pw[pass].x), __uint_as_float(pw[pass].y), __uint_as_float(pw[pass].z),
                                __uint_as_float(pw[pass].w)};
            const uint4 v = Blend<T>::run(c[pass], w);
            sts16(a_base + r * ROW_BYTES + ((cg ^ (r & 7)) << 4), v);
          }
          fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core (async proxy)
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_fullA + 8 * s);
        }
      }
    }
  } else if (warp == 4) {
    // =========================================================================== weight loader
    if (lane == 0) {
      uint32_t it = 0;
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.wpacked);
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        for (int q = 0; q < chunks_per_tile; ++q, ++it) {
          const uint32_t s = it % (uint32_t)p.stages, ph = (it / (uint32_t)p.stages) & 1u;
          mbar_wait(bar_empty + 8 * s, ph ^ 1u);
          mbar_arrive_expect_tx(bar_fullB + 8 * s, (uint32_t)p.b_stage_bytes);
          bulk_copy_g2s(smem_base + SMEM_STAGE_OFF + s * p.stage_bytes + A_STAGE_BYTES, wsrc + (size_t)q * p.b_stage_bytes,
                        (uint32_t)p.b_stage_bytes, bar_fullB + 8 * s);
        }
      }
    }
  } else if (warp == 5) {
    // =========================================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | ((kTF32 ? 2u : 1u) << 7) | ((kTF32 ? 2u : 1u) << 10) |
                             ((uint32_t)(p.Cout >> 3) << 17) | ((uint32_t)(TILE_PX >> 4) << 24);
      uint32_t it = 0, tcount = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++tcount) {
        const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
        mbar_wait(bar_tempty + 8 * acc, accph ^ 1u);  // epilogue drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)p.acc_cols;
        for (int q = 0; q < chunks_per_tile; ++q, ++it) {
          const uint32_t s = it % (uint32_t)p.stages, ph = (it / (uint32_t)p.stages) & 1u;
          mbar_wait(bar_fullA + 8 * s, ph);
          mbar_wait(bar_fullB + 8 * s, ph);
          tc_fence_after();
          const uint32_t a_addr = smem_base + SMEM_STAGE_OFF + s * p.stage_bytes;
          const uint64_t adesc = umma_desc_sw128(a_addr);
          const uint64_t bdesc = umma_desc_sw128(a_addr + A_STAGE_BYTES);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // 4 x 32 bytes of K per 128-byte row
            umma<kTF32>(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (q | kk) != 0 ? 1u : 0u);
          umma_commit(bar_empty + 8 * s);          // frees the smem stage when these MMAs retire
        }
        umma_commit(bar_tfull + 8 * acc);          // accumulator complete -> epilogue
      }
    }
  } else {
    // =========================================================================== epilogue (warps 0-3)
    T* yg = reinterpret_cast<T*>(p.y);
    uint32_t tcount = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++tcount) {
      const uint32_t acc = tcount & 1u, accph = (tcount >> 1) & 1u;
      const int b = tile / tiles_per_img;
      const int trem = tile - b * tiles_per_img;
      const int r = warp * 32 + lane;
      const int y = (trem / p.tiles_x) * TILE_H + r / TILE_W, x = (trem % p.tiles_x) * TILE_W + r % TILE_W;
      const bool valid = (y < p.H) && (x < p.W);
      T* yrow = yg + (((int64_t)b * p.H + y) * p.W + x) * p.Cout;
      mbar_wait(bar_tfull + 8 * acc, accph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + acc * (uint32_t)p.acc_cols;
      for (int c0 = 0; c0 < p.Cout; c0 += 16) {
        uint32_t raw[16];
        tmem_ld16(t_row + (uint32_t)c0, raw);
        tmem_ld_wait();
        if (valid) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(raw[j]) + bias_s[c0 + j];
          OutConv<T>::store(yrow + c0, v);
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
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols) : "memory");
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
        const __nv_bfloat162 pk = __floats2bfloat162_rn(w[((int64_t)n * Cin + c) * NUM_TAPS + tap],
                                                        w[((int64_t)n * Cin + c + 1) * NUM_TAPS + tap]);
        words[e] = *reinterpret_cast<const uint32_t*>(&pk);
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
  if ((int64_t)a.B * a.H * a.W * a.Cin >= (1LL << 31)) return false;
  return true;
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

int dcn_forward_tc(const cpoly_dcn_fwd_args& a, cudaStream_t stream) {
  TcParams p;
  const int esize = a.dtype == CPOLY_BF16 ? 2 : 4;
  p.x = a.x.ptr; p.offset = a.offset; p.mask = a.mask; p.wpacked = a.weight_packed; p.bias = a.bias; p.y = a.y.ptr;
  p.B = a.B; p.Cin = a.Cin; p.Cout = a.Cout; p.H = a.H; p.W = a.W;
  p.om_dtype = a.om_dtype; p.mask_is_logit = a.mask_is_logit;
  p.tiles_x = (a.W + TILE_W - 1) / TILE_W;
  p.tiles_y = (a.H + TILE_H - 1) / TILE_H;
  p.num_tiles = a.B * p.tiles_x * p.tiles_y;
  if (p.num_tiles == 0) return CPOLY_OK;
  p.nchunk = a.Cin / (ROW_BYTES / esize);
  p.b_stage_bytes = a.Cout * ROW_BYTES;
  p.stage_bytes = A_STAGE_BYTES + p.b_stage_bytes;
  int stages = (SMEM_LIMIT - SMEM_STAGE_OFF) / p.stage_bytes;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  CPOLY_REQUIRE(stages >= 2, CPOLY_EUNSUPPORTED, "not enough shared memory for a 2-stage pipeline");
  p.stages = stages;
  p.acc_cols = a.Cout;
  p.tmem_cols = pow2_at_least(2 * a.Cout);
  const int smem_bytes = SMEM_STAGE_OFF + stages * p.stage_bytes + 1024;  // +1 KB alignment slack

  int dev = 0, sms = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.num_tiles < sms ? p.num_tiles : sms;
  if (a.dtype == CPOLY_BF16) {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_fwd_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    dcn_fwd_tc_kernel<__nv_bfloat16><<<grid, NTHREADS, smem_bytes, stream>>>(p);
  } else {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_fwd_tc_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    dcn_fwd_tc_kernel<float><<<grid, NTHREADS, smem_bytes, stream>>>(p);
  }
  return check_launch("dcn_fwd_tc_kernel");
}

}  // namespace cpoly
