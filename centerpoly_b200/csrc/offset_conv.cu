// The offset / mask producer of the DCN module: `conv_offset_mask`, a regular 3x3 / stride 1 / pad 1
// convolution Cin -> 27 whose output is chunked into 18 offsets and 9 mask logits by upstream DCNv2's `DCN.forward`
// (module constructed at src/lib/models/networks/pose_dla_dcn.py:354 and resnet_dcn.py:221-223; the extension is the
// missing submodule src/lib/models/networks/DCNv2).  The reference runs it as cuDNN; here it is one implicit-GEMM kernel:
//
//   om[px, n] = bias[n] + sum_{tap, c} x[px + tap, c] * W[n, c, tap]        n < 32 (27 real channels + zero padding)
//
// NHWC bf16 activations, fp32 accumulation, fp32 (or bf16) NHWC output with the bias already added - the DCN kernels
// then read exact fp32 offsets instead of bf16-rounded ones (an offset of 5.3 px rounded to bf16 moves the sampling
// point by up to 0.02 px).  The GEMM is M = pixels, N = 32, K = 9 * Cin.  Two kernels:
//
//   offset_conv3x3_tc_kernel (default)  tcgen05 / TMEM, further down in this file.
//   offset_conv3x3_kernel (CPOLY_OFFSET_CONV=mma)  the first version: warp-level mma.sync (HMMA) fed by ldmatrix from a
//   shared-memory halo window.  CTA = 8 warps = one 8 x 16 pixel tile per work item; warp w owns tile row w (16 pixels =
//   one m16 block) x 32 outputs.  Per (tile, 64-channel chunk): the 10 x 18 pixel halo window (23 KB, out-of-image
//   pixels zero-filled by cp.async) and the chunk's 9 x 32 x 64 weights (36 KB, pre-swizzled image) are staged with
//   16-byte cp.async, double-buffered against the MMAs of the previous stage; 16-byte pieces are XOR-swizzled by
//   (row & 7) so every ldmatrix phase is conflict-free.  Measured: HMMA-bound at ~180 TFLOP/s (27 us for the 64-channel
//   256x512 layer, 338 us per frame) - kept for comparison.
#include <stdlib.h>
#include <string.h>

#include <cuda.h>

#include "tc_common.cuh"

namespace cpoly {
namespace {

constexpr int OC_THREADS = 256;
constexpr int OC_TH = 8, OC_TW = 16;                   // pixel tile
constexpr int OC_WH = OC_TH + 2, OC_WW = OC_TW + 2;    // halo window
constexpr int OC_N = 32;                               // padded output channels
constexpr int OC_KC = 64;                              // input channels per stage (128 bytes of bf16)
constexpr int OC_WIN_BYTES = OC_WH * OC_WW * 128;      // 23040
constexpr int OC_WGT_BYTES = 9 * OC_N * 128;           // 36864

struct OffsetConvParams {
  const __nv_bfloat16* x;        // dense NHWC (B,H,W,Cin)
  const uint4* wpacked;          // [Cin/64][9][32][8] 16-byte pieces, piece index XOR (n & 7)
  const float* bias;             // [32] (padded with zeros)
  void* om;                      // dense NHWC (B,H,W,32), fp32 or bf16
  int B, Cin, H, W, nchunk, out_bf16;
  int tiles_x, tiles_y, num_tiles, wgt_stages;
};

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(OC_THREADS) offset_conv3x3_kernel(const OffsetConvParams p) {
  extern __shared__ __align__(128) uint8_t oc_smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint8_t* win_s[2] = {oc_smem, oc_smem + OC_WIN_BYTES};
  uint8_t* wgt_s[2] = {oc_smem + 2 * OC_WIN_BYTES, oc_smem + 2 * OC_WIN_BYTES + (p.wgt_stages == 2 ? OC_WGT_BYTES : 0)};
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const int my_tiles = ((int)blockIdx.x < p.num_tiles) ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int n_items = my_tiles * p.nchunk;
  const size_t pixb = (size_t)p.Cin * sizeof(__nv_bfloat16);

  // stage `it` = (tile, chunk): window + (when the weights are not resident) the chunk's weights
  auto issue = [&](int it) {
    const int tile = (int)blockIdx.x + (it / p.nchunk) * (int)gridDim.x, cc = it % p.nchunk;
    const int b = tile / tiles_per_img, trem = tile - b * tiles_per_img;
    const int ty0 = (trem / p.tiles_x) * OC_TH, tx0 = (trem % p.tiles_x) * OC_TW;
    const uint32_t wbase = s_u32(win_s[it & 1]);
    for (int i = tid; i < OC_WH * OC_WW * 8; i += OC_THREADS) {
      const int pidx = i >> 3, j = i & 7;
      const int y = ty0 - 1 + pidx / OC_WW, x = tx0 - 1 + pidx % OC_WW;
      const bool in = (unsigned)y < (unsigned)p.H && (unsigned)x < (unsigned)p.W;
      const uint8_t* src = reinterpret_cast<const uint8_t*>(p.x) + ((size_t)(b * p.H + (in ? y : 0)) * p.W + (in ? x : 0)) * pixb + (size_t)cc * 128 + j * 16;
      cp_async16(wbase + pidx * 128 + ((j ^ (pidx & 7)) << 4), src, in ? 16 : 0);     // src-size 0: zero fill (the conv's padding)
    }
    if (p.wgt_stages == 2 || it == 0) {
      const uint32_t gbase = s_u32(wgt_s[p.wgt_stages == 2 ? (it & 1) : 0]);
      const uint4* src = p.wpacked + (size_t)cc * (OC_WGT_BYTES / 16);
      for (int i = tid; i < OC_WGT_BYTES / 16; i += OC_THREADS) cp_async16(gbase + i * 16, src + i, 16);
    }
    cp_async_commit();
  };

  if (n_items > 0) issue(0);
  float acc[4][4];
  // ldmatrix lane roles (constant): A x4 = (rows 0-7, k 0-7) (rows 8-15, k 0-7) (rows 0-7, k 8-15) (rows 8-15, k 8-15)
  const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_kh = lane >> 4;
  // B x4 = (n 0-7, k 0-7) (n 0-7, k 8-15) (n 8-15, k 0-7) (n 8-15, k 8-15) of a 16-output pair
  const int b_n = (lane & 7) + (lane >> 4) * 8, b_kh = (lane >> 3) & 1;
  for (int it = 0; it < n_items; ++it) {
    if (it + 1 < n_items) { issue(it + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const int tile = (int)blockIdx.x + (it / p.nchunk) * (int)gridDim.x, cc = it % p.nchunk;
    if (cc == 0) {
#pragma unroll
      for (int t = 0; t < 4; ++t) acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
    }
    const uint32_t wbase = s_u32(win_s[it & 1]);
    const uint32_t gbase = s_u32(wgt_s[p.wgt_stages == 2 ? (it & 1) : 0]);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
      const int ti = tap / 3, tj = tap - 3 * ti;
      const int pidx = (warp + ti) * OC_WW + tj + a_row;          // window pixel this lane addresses for the A rows
      const uint32_t a_addr = wbase + pidx * 128;
      const uint32_t b_addr = gbase + tap * (OC_N * 128) + b_n * 128;
#pragma unroll
      for (int ks = 0; ks < OC_KC / 16; ++ks) {
        uint32_t a[4], b0[4], b1[4];
        ldmatrix_x4(a_addr + ((((ks << 1) | a_kh) ^ (pidx & 7)) << 4), a[0], a[1], a[2], a[3]);
        ldmatrix_x4(b_addr + ((((ks << 1) | b_kh) ^ (b_n & 7)) << 4), b0[0], b0[1], b0[2], b0[3]);
        ldmatrix_x4(b_addr + 16 * 128 + ((((ks << 1) | b_kh) ^ (b_n & 7)) << 4), b1[0], b1[1], b1[2], b1[3]);
        mma_bf16_16816(acc[0], a, b0[0], b0[1]);
        mma_bf16_16816(acc[1], a, b0[2], b0[3]);
        mma_bf16_16816(acc[2], a, b1[0], b1[1]);
        mma_bf16_16816(acc[3], a, b1[2], b1[3]);
      }
    }
    if (cc == p.nchunk - 1) {
      // C fragment: c0,c1 -> row lane/4, columns 2*(lane%4), +1 of each 8-wide output block; c2,c3 -> row lane/4 + 8
      const int b = tile / tiles_per_img, trem = tile - b * tiles_per_img;
      const int y = (trem / p.tiles_x) * OC_TH + warp, tx0 = (trem % p.tiles_x) * OC_TW;
      if (y < p.H) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int x = tx0 + (lane >> 2) + half * 8;
          if (x >= p.W) continue;
          const size_t pix = (size_t)(b * p.H + y) * p.W + x;
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int n = t * 8 + (lane & 3) * 2;
            const float v0 = acc[t][half * 2] + __ldg(p.bias + n), v1 = acc[t][half * 2 + 1] + __ldg(p.bias + n + 1);
            if (p.out_bf16) {
              const __nv_bfloat162 pk = __floats2bfloat162_rn(v0, v1);
              *reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(p.om) + pix * OC_N + n) = pk;
            } else {
              *reinterpret_cast<float2*>(reinterpret_cast<float*>(p.om) + pix * OC_N + n) = make_float2(v0, v1);
            }
          }
        }
      }
    }
    __syncthreads();   // the buffers of this stage are refilled by the next issue()
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// tcgen05 version (default).  mma.sync tops out near 180 TFLOP/s on this part (the kernel above: 27 us for the
// 64-channel 256x512 layer, 338 us per frame against cuDNN's ~190), so the GEMM moves to the 5th-generation tensor
// cores: per (tile, 64-channel chunk) 9 taps x 4 tcgen05.mma (M = 128 pixels, N = 32, K = 16) accumulate in 32 TMEM
// columns.  The A operand of a tap must be a K-major SWIZZLE_128B tile whose 8-row groups are 1024 B apart, which a
// halo window with an 18-pixel pitch cannot offer; so the window is staged THREE times, shifted by -1 / 0 / +1 pixel
// (10 rows x 16 pixels x 128 B each, cp.async with zero fill = the convolution's padding): tap (i, j) is then the
// contiguous 16 KB starting at row i of window j - a plain descriptor, 1024-byte aligned.  One stage (3 windows +
// the chunk's weights, 96 KB) per CTA, two CTAs per SM overlap each other's load and MMA phases.
constexpr int OT_THREADS = 128;
constexpr int OT_WIN1 = OC_WH * OC_TW * 128;           // one shifted window: 10 x 16 pixels x 128 B = 20480
constexpr int OT_WIN_BYTES = 3 * OT_WIN1;              // 61440
constexpr int OT_SMEM = OT_WIN_BYTES + OC_WGT_BYTES + 1024;   // + alignment slack

__global__ void __launch_bounds__(OT_THREADS) offset_conv3x3_tc_kernel(const OffsetConvParams p) {
  using namespace tc;
  extern __shared__ uint8_t ot_smem_raw[];
  __shared__ __align__(8) uint64_t bar_storage, wbar_storage;
  __shared__ uint32_t tmem_ptr_s;
  uint8_t* smem = ot_smem_raw + ((1024u - (smem_u32(ot_smem_raw) & 1023u)) & 1023u);
  const uint32_t win0 = smem_u32(smem), wgt0 = win0 + OT_WIN_BYTES;
  const uint32_t bar = smem_u32(&bar_storage), wbar = smem_u32(&wbar_storage);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const int my_tiles = ((int)blockIdx.x < p.num_tiles) ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int n_items = my_tiles * p.nchunk;
  const size_t pixb = (size_t)p.Cin * sizeof(__nv_bfloat16);
  uint32_t wphase = 0;

  if (tid == 0) { mbar_init(bar, 1); mbar_init(wbar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_ptr_s), 32u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = tmem_ptr_s;
  const uint32_t idesc = umma_idesc(false, 128, OC_N);
  uint32_t phase = 0;

  for (int it = 0; it < n_items; ++it) {
    const int tile = (int)blockIdx.x + (it / p.nchunk) * (int)gridDim.x, cc = it % p.nchunk;
    const int b = tile / tiles_per_img, trem = tile - b * tiles_per_img;
    const int ty0 = (trem / p.tiles_x) * OC_TH, tx0 = (trem % p.tiles_x) * OC_TW;
    // ---- stage: three shifted windows (+ the chunk's weights unless they are resident).  The 128 threads cover exactly
    // one window row (16 pixels x 8 pieces) per step, so the address of a step is the previous one plus a row pitch: no
    // per-piece index arithmetic (the first version spent ~3000 cycles per stage on it).  The weights are ONE bulk copy.
    if (tid == 0 && (p.nchunk > 1 || it == 0)) {
      mbar_arrive_expect_tx(wbar, (uint32_t)OC_WGT_BYTES);
      bulk_copy_g2s(wgt0, p.wpacked + (size_t)cc * (OC_WGT_BYTES / 16), (uint32_t)OC_WGT_BYTES, wbar);
    }
    {
      const int px = tid >> 3, j = tid & 7;
      const uint32_t dst0 = win0 + (uint32_t)px * 128u + (uint32_t)((j ^ (px & 7)) << 4);
      const uint8_t* img = reinterpret_cast<const uint8_t*>(p.x) + (size_t)b * p.H * p.W * pixb + (size_t)cc * 128 + j * 16;
#pragma unroll
      for (int sh = 0; sh < 3; ++sh) {
        const int x = tx0 - 1 + sh + px;
        const bool xin = (unsigned)x < (unsigned)p.W;
        const uint8_t* col = img + (size_t)(xin ? x : 0) * pixb;
#pragma unroll
        for (int wy = 0; wy < OC_WH; ++wy) {
          const int y = ty0 - 1 + wy;
          const bool in = xin && (unsigned)y < (unsigned)p.H;
          cp_async16(dst0 + (uint32_t)(sh * OT_WIN1 + wy * (OC_TW * 128)), col + (size_t)(in ? y : 0) * p.W * pixb, in ? 16 : 0);
        }
      }
    }
    cp_async_commit();
    cp_async_wait<0>();
    fence_proxy_async_smem();     // generic-proxy (cp.async) writes -> visible to the tensor core
    if (p.nchunk > 1 || it == 0) { mbar_wait(wbar, wphase, 20); wphase ^= 1u; }
    __syncthreads();
    // ---- 9 taps x 4 K-steps
    if (warp == 0) {
      tc_fence_after();
      if (elect_one()) {
#pragma unroll 1
        for (int tap = 0; tap < 9; ++tap) {
          const int ti = tap / 3, tj = tap - 3 * ti;
          const uint64_t adesc = umma_desc_sw128(win0 + tj * OT_WIN1 + ti * (OC_TW * 128));
          const uint64_t bdesc = umma_desc_sw128(wgt0 + tap * (OC_N * 128));
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma<false>(tmem_acc, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (cc | tap | kk) != 0 ? 1u : 0u);
        }
        umma_commit(bar);
      }
      __syncwarp();
    }
    mbar_wait(bar, phase, 20);    // the MMAs have read the stage (and, after the last chunk, the accumulator is complete)
    phase ^= 1u;
    if (cc == p.nchunk - 1) {
      tc_fence_after();
      const int r = warp * 32 + lane;                          // TMEM lane = M row = pixel (r / 16, r % 16) of the tile
      const int y = ty0 + (r >> 4), x = tx0 + (r & 15);
      const uint32_t t_row = tmem_acc + ((uint32_t)(warp * 32) << 16);
      uint32_t v0[16], v1[16];
      tmem_ld16(t_row, v0);
      tmem_ld16(t_row + 16u, v1);
      tmem_ld_wait();
      if (y < p.H && x < p.W) {
        const size_t pix = (size_t)(b * p.H + y) * p.W + x;
        float o[32];
#pragma unroll
        for (int n = 0; n < 16; ++n) { o[n] = __uint_as_float(v0[n]) + __ldg(p.bias + n); o[16 + n] = __uint_as_float(v1[n]) + __ldg(p.bias + 16 + n); }
        if (p.out_bf16) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.om) + pix * OC_N);
#pragma unroll
          for (int q = 0; q < 4; ++q)
            dst[q] = make_uint4(pack_bf16x2(o[8 * q], o[8 * q + 1]), pack_bf16x2(o[8 * q + 2], o[8 * q + 3]), pack_bf16x2(o[8 * q + 4], o[8 * q + 5]),
                                pack_bf16x2(o[8 * q + 6], o[8 * q + 7]));
        } else {
          float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.om) + pix * OC_N);
#pragma unroll
          for (int q = 0; q < 8; ++q) dst[q] = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
        }
      }
      tc_fence_before();
    }
    __syncthreads();   // every warp has read the accumulator / the stage may be refilled
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_acc, 32u); }
}

// ---------------------------------------------------------------------------------------------------------------------
// TMA-fed, pipelined version (default when the driver offers cuTensorMapEncodeTiled).  Same operand layout and MMAs as
// offset_conv3x3_tc_kernel, but the three shifted windows arrive as three `cp.async.bulk.tensor.4d` box loads
// (64 channels x 16 pixels x 10 rows, SWIZZLE_128B, out-of-image pixels zero-filled by the TMA unit = the convolution's
// padding) issued by one thread, into a two-stage ring; one persistent CTA per SM:
//   warp 4     producer + MMA issuer (one elected lane): load of item t+1 is in flight while the 36 MMAs of item t run
//   warps 0-3  epilogue: accumulator (two of them, 2 x 32 TMEM columns) -> + bias -> NHWC stores, overlapping the next tile
// The cp.async version spent its time issuing 30 16-byte copies per thread and then waiting for them with nothing else
// to do (22 us for the 64-channel 256x512 layer, launch list of r02).
constexpr int OP_THREADS = 160;
constexpr int OP_MAX_STAGES = 3;
// one 64-channel chunk: the weights stay resident and three window stages fit (216 KB); more chunks: a weight chunk
// travels with every item, two stages of windows + weights (192 KB)
constexpr int OP_SMEM_1 = 3 * OT_WIN_BYTES + OC_WGT_BYTES + 1024, OP_SMEM_N = 2 * (OT_WIN_BYTES + OC_WGT_BYTES) + 1024;

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tmap, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
      : "memory");
}

__global__ void __launch_bounds__(OP_THREADS) offset_conv3x3_tma_kernel(const OffsetConvParams p, const __grid_constant__ CUtensorMap tmap) {
  using namespace tc;
  extern __shared__ uint8_t op_smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * OP_MAX_STAGES + 4];   // full[3], empty[3], accfull[2], accempty[2]
  __shared__ uint32_t tmem_ptr_s;
  uint8_t* smem = op_smem_raw + ((1024u - (smem_u32(op_smem_raw) & 1023u)) & 1023u);
  const uint32_t stage0 = smem_u32(smem);
  const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[OP_MAX_STAGES]), bar_afull = smem_u32(&bars[2 * OP_MAX_STAGES]),
                 bar_aempty = smem_u32(&bars[2 * OP_MAX_STAGES + 2]);
  const int NS = p.wgt_stages;                                        // window stages (3 when one chunk, else 2)
  const uint32_t wgt0 = stage0 + (uint32_t)(NS * OT_WIN_BYTES);       // weight slots follow the windows
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const int my_tiles = ((int)blockIdx.x < p.num_tiles) ? (p.num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int n_items = my_tiles * p.nchunk;

  if (tid == 0) {
    for (int i = 0; i < OP_MAX_STAGES; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_afull + 8 * i, 1);
      mbar_init(bar_aempty + 8 * i, 128);
    }
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(smem_u32(&tmem_ptr_s), 64u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_ptr_s;

  if (warp == 4) {
    // ============================================================ producer + MMA issuer
    const uint32_t idesc = umma_idesc(false, 128, OC_N);
    // item t -> window stage t % NS; with several chunks a weight chunk travels with the item into the stage's slot
    auto issue_load = [&](int t) {
      const int s = t % NS;
      const int tile = (int)blockIdx.x + (t / p.nchunk) * (int)gridDim.x, cc = t % p.nchunk;
      const int b = tile / tiles_per_img, trem = tile - b * tiles_per_img;
      const int ty0 = (trem / p.tiles_x) * OC_TH, tx0 = (trem % p.tiles_x) * OC_TW;
      const bool wload = p.nchunk > 1 || t == 0;
      const uint32_t dst = stage0 + (uint32_t)(s * OT_WIN_BYTES);
      if (elect_one()) {
        mbar_arrive_expect_tx(bar_full + 8 * s, (uint32_t)(OT_WIN_BYTES + (wload ? OC_WGT_BYTES : 0)));
#pragma unroll
        for (int sh = 0; sh < 3; ++sh) tma_load_4d(dst + (uint32_t)(sh * OT_WIN1), &tmap, cc * OC_KC, tx0 - 1 + sh, ty0 - 1, b, bar_full + 8 * s);
        if (wload)
          bulk_copy_g2s(wgt0 + (uint32_t)((p.nchunk > 1 ? s : 0) * OC_WGT_BYTES), p.wpacked + (size_t)cc * (OC_WGT_BYTES / 16), (uint32_t)OC_WGT_BYTES,
                        bar_full + 8 * s);
      }
      __syncwarp();
    };
    for (int t = 0; t < NS - 1 && t < n_items; ++t) issue_load(t);
    for (int t = 0; t < n_items; ++t) {
      const int s = t % NS, cc = t % p.nchunk, tl = t / p.nchunk, acc = tl & 1;
      if (t + NS - 1 < n_items) {
        // the stage of item t + NS - 1 is the one item t - 1 used: its MMAs (committed one iteration ago) must have read it
        if (t >= 1) mbar_wait(bar_empty + 8 * ((t - 1) % NS), (uint32_t)((t - 1) / NS) & 1u, 20);
        issue_load(t + NS - 1);
      }
      mbar_wait(bar_full + 8 * s, (uint32_t)(t / NS) & 1u, 20);
      if (cc == 0 && tl >= 2) mbar_wait(bar_aempty + 8 * acc, (uint32_t)((tl >> 1) - 1) & 1u, 20);   // epilogue drained this accumulator
      tc_fence_after();
      const uint32_t win = stage0 + (uint32_t)(s * OT_WIN_BYTES), wgt = wgt0 + (uint32_t)((p.nchunk > 1 ? s : 0) * OC_WGT_BYTES),
                     d_tmem = tmem_base + (uint32_t)(acc * OC_N);
      if (elect_one()) {
#pragma unroll 1
        for (int tap = 0; tap < 9; ++tap) {
          const int ti = tap / 3, tj = tap - 3 * ti;
          const uint64_t adesc = umma_desc_sw128(win + tj * OT_WIN1 + ti * (OC_TW * 128));
          const uint64_t bdesc = umma_desc_sw128(wgt + tap * (OC_N * 128));
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma<false>(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (cc | tap | kk) != 0 ? 1u : 0u);
        }
        umma_commit(bar_empty + 8 * s);
        if (cc == p.nchunk - 1) umma_commit(bar_afull + 8 * acc);
      }
      __syncwarp();
    }
  } else {
    // ============================================================ epilogue (warps 0-3: TMEM lanes 32 * warp ..)
    float bias_r[OC_N];
#pragma unroll
    for (int n = 0; n < OC_N; ++n) bias_r[n] = __ldg(p.bias + n);
    for (int tl = 0; tl < my_tiles; ++tl) {
      const int acc = tl & 1;
      const int tile = (int)blockIdx.x + tl * (int)gridDim.x;
      const int b = tile / tiles_per_img, trem = tile - b * tiles_per_img;
      const int ty0 = (trem / p.tiles_x) * OC_TH, tx0 = (trem % p.tiles_x) * OC_TW;
      const int r = warp * 32 + lane;                          // TMEM lane = M row = pixel (r / 16, r % 16) of the tile
      const int y = ty0 + (r >> 4), x = tx0 + (r & 15);
      mbar_wait(bar_afull + 8 * acc, (uint32_t)(tl >> 1) & 1u, 50);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(acc * OC_N);
      uint32_t v0[16], v1[16];
      tmem_ld16(t_row, v0);
      tmem_ld16(t_row + 16u, v1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar_aempty + 8 * acc);                       // the values are in registers: the accumulator is free
      if (y < p.H && x < p.W) {
        const size_t pix = (size_t)(b * p.H + y) * p.W + x;
        float o[32];
#pragma unroll
        for (int n = 0; n < 16; ++n) { o[n] = __uint_as_float(v0[n]) + bias_r[n]; o[16 + n] = __uint_as_float(v1[n]) + bias_r[16 + n]; }
        if (p.out_bf16) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.om) + pix * OC_N);
#pragma unroll
          for (int q = 0; q < 4; ++q)
            dst[q] = make_uint4(pack_bf16x2(o[8 * q], o[8 * q + 1]), pack_bf16x2(o[8 * q + 2], o[8 * q + 3]), pack_bf16x2(o[8 * q + 4], o[8 * q + 5]),
                                pack_bf16x2(o[8 * q + 6], o[8 * q + 7]));
        } else {
          float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.om) + pix * OC_N);
#pragma unroll
          for (int q = 0; q < 8; ++q) dst[q] = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem_base, 64u); }
}

typedef CUresult (*OcEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
OcEncodeTiledFn oc_encode_tiled_fn() {
  static OcEncodeTiledFn fn = []() -> OcEncodeTiledFn {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<OcEncodeTiledFn>(sym);
  }();
  return fn;
}
// conv_offset_mask.weight (n_out <= 32, Cin, 3, 3), any float dtype as fp32 -> [Cin/64][9][32][8 pieces] bf16, piece j of
// row n stored at j ^ (n & 7); rows >= n_out are zero.
__global__ void offset_conv_pack_kernel(const float* __restrict__ w, uint4* __restrict__ out, int Cin, int n_out) {
  const int total = (Cin / OC_KC) * 9 * OC_N * 8;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int j = i & 7;
    int t = i >> 3;
    const int n = t % OC_N; t /= OC_N;
    const int tap = t % 9;
    const int cc = t / 9;
    uint32_t words[4] = {0u, 0u, 0u, 0u};
    if (n < n_out) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int c = cc * OC_KC + j * 8 + 2 * e;
        const __nv_bfloat162 pk = __floats2bfloat162_rn(w[((size_t)n * Cin + c) * 9 + tap], w[((size_t)n * Cin + c + 1) * 9 + tap]);
        words[e] = *reinterpret_cast<const uint32_t*>(&pk);
      }
    }
    out[((size_t)(cc * 9 + tap) * OC_N + n) * 8 + (j ^ (n & 7))] = make_uint4(words[0], words[1], words[2], words[3]);
  }
}

}  // namespace

size_t offset_conv_packed_bytes(int Cin) { return (Cin % OC_KC == 0) ? (size_t)(Cin / OC_KC) * OC_WGT_BYTES : 0; }

int offset_conv_pack_launch(const float* weight, void* packed, int Cin, int n_out, cudaStream_t stream) {
  const int total = (Cin / OC_KC) * 9 * OC_N * 8;
  offset_conv_pack_kernel<<<(total + 255) / 256, 256, 0, stream>>>(weight, reinterpret_cast<uint4*>(packed), Cin, n_out);
  return check_launch("offset_conv_pack_kernel");
}

int offset_conv_launch(const cpoly_offset_conv_args& a, cudaStream_t stream) {
  OffsetConvParams p;
  p.x = reinterpret_cast<const __nv_bfloat16*>(a.x);
  p.wpacked = reinterpret_cast<const uint4*>(a.weight_packed);
  p.bias = a.bias;
  p.om = a.om;
  p.B = a.B; p.Cin = a.Cin; p.H = a.H; p.W = a.W; p.nchunk = a.Cin / OC_KC; p.out_bf16 = a.om_dtype == CPOLY_BF16;
  p.tiles_x = (a.W + OC_TW - 1) / OC_TW; p.tiles_y = (a.H + OC_TH - 1) / OC_TH;
  p.num_tiles = a.B * p.tiles_x * p.tiles_y;
  if (p.num_tiles == 0) return CPOLY_OK;
  int dev = 0, sms = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const char* e = getenv("CPOLY_OFFSET_CONV");
  const bool aligned = (reinterpret_cast<uintptr_t>(a.x) & 15) == 0 && (int64_t)a.H * a.W * a.Cin * 2 < (1LL << 40);
  if (!e && aligned && oc_encode_tiled_fn()) {     // default: TMA-fed pipeline; CPOLY_OFFSET_CONV=cpasync | mma keep the older kernels
    CUtensorMap tmap;
    const cuuint64_t dims[4] = {(cuuint64_t)a.Cin, (cuuint64_t)a.W, (cuuint64_t)a.H, (cuuint64_t)a.B};
    const cuuint64_t strides[3] = {(cuuint64_t)a.Cin * 2, (cuuint64_t)a.W * a.Cin * 2, (cuuint64_t)a.H * a.W * a.Cin * 2};
    const cuuint32_t box[4] = {(cuuint32_t)OC_KC, (cuuint32_t)OC_TW, (cuuint32_t)OC_WH, 1u};
    const cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
    const CUresult cr = oc_encode_tiled_fn()(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(a.x), dims, strides, box, estr,
                                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CPOLY_REQUIRE(cr == CUDA_SUCCESS, CPOLY_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)cr);
    p.wgt_stages = p.nchunk == 1 ? 3 : 2;           // (window stages of this kernel)
    if (const char* st = getenv("CPOLY_OC_STAGES")) { if (atoi(st) == 2) p.wgt_stages = 2; }   // profiling knob
    const int smem = p.nchunk == 1 ? p.wgt_stages * OT_WIN_BYTES + OC_WGT_BYTES + 1024 : OP_SMEM_N;
    CPOLY_CUDA_OK(cudaFuncSetAttribute(offset_conv3x3_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OP_SMEM_1 > OP_SMEM_N ? OP_SMEM_1 : OP_SMEM_N));
    const int grid = p.num_tiles < sms ? p.num_tiles : sms;
    offset_conv3x3_tma_kernel<<<grid, OP_THREADS, smem, stream>>>(p, tmap);
    return check_launch("offset_conv3x3_tma_kernel");
  }
  if (!(e && strcmp(e, "mma") == 0)) {            // default: tcgen05; CPOLY_OFFSET_CONV=mma keeps the warp-level kernel
    p.wgt_stages = 1;
    CPOLY_CUDA_OK(cudaFuncSetAttribute(offset_conv3x3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OT_SMEM));
    CPOLY_CUDA_OK(cudaFuncSetAttribute(offset_conv3x3_tc_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));   // two CTAs per SM
    const int grid = p.num_tiles < 2 * sms ? p.num_tiles : 2 * sms;
    offset_conv3x3_tc_kernel<<<grid, OT_THREADS, OT_SMEM, stream>>>(p);
    return check_launch("offset_conv3x3_tc_kernel");
  }
  p.wgt_stages = p.nchunk == 1 ? 1 : 2;          // one chunk: the weights stay resident for the whole kernel
  const int smem = 2 * OC_WIN_BYTES + p.wgt_stages * OC_WGT_BYTES;
  CPOLY_CUDA_OK(cudaFuncSetAttribute(offset_conv3x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int per_sm = p.wgt_stages == 1 ? 2 : 1;
  const int grid = p.num_tiles < sms * per_sm ? p.num_tiles : sms * per_sm;
  offset_conv3x3_kernel<<<grid, OC_THREADS, smem, stream>>>(p);
  return check_launch("offset_conv3x3_kernel");
}

}  // namespace cpoly
