// DCNv2 backward on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// Replaces the DCNv2 extension's `dcn_v2_backward` (missing submodule src/lib/models/networks/DCNv2; driven by
// loss.backward() in src/lib/trains/base_trainer.py:97-98 through the 16 DeformConv layers of
// src/lib/models/networks/pose_dla_dcn.py:347-413).  Upstream re-materialises columns[9*Cin, H*W] in HBM and runs
// cuBLAS SGEMMs plus two atomic scatter kernels; here both GEMMs run on tcgen05 with sampler-produced operands.
//
//   wgrad   gW[o,(tap,c)] = sum_px gout[px,o] * A[px,(tap,c)],   A = mask * bilinear(x, p_tap)      (this file, part 1)
//
// wgrad kernel.  K = pixels: both operands are "MN-major" (rows = K index, 128 B of M / N per row), which is exactly
// how the data lies anyway: the gout tile of a 128-pixel image tile (NHWC rows) is the A operand, the sampled column
// tile written by the forward-style sampler is the B operand.  A job = (pixel split s, block group g); a block is
// one (tap, 128-byte channel chunk).  The CTA loops over the tiles of its split, accumulating all its blocks in TMEM
// (<= 512 columns), and finally writes ONE fp32 partial [tap][o][c]; a fixed-order reduction over the splits follows
// (deterministic, no atomics).
#include "dcn_sampler.cuh"

namespace cpoly {
namespace {

using namespace tc;
using namespace smp;

constexpr int EPI_WARPS = 4;
constexpr int SAMPLER_WARP0 = 6;
constexpr int SAMPLER_WARPS = 16;
constexpr int NS = SAMPLER_WARPS * 32;
constexpr int NTHREADS = (SAMPLER_WARP0 + SAMPLER_WARPS) * 32;
constexpr int ROWS_PER_PASS = NS / 8;
constexpr int PASSES = TILE_PX / ROWS_PER_PASS;
constexpr int ATOM_BYTES = TILE_PX * ROW_BYTES;   // 128 rows x 128 B = 16 KB: one swizzle atom column
constexpr int MAX_B_STAGES = 6;
constexpr int MAX_A_BUFS = 2;

constexpr int SMEM_BAR_OFF = 0;
constexpr int SMEM_POFF_OFF = 1024;
constexpr int SMEM_PW_OFF = SMEM_POFF_OFF + NUM_TAPS * TILE_PX * 16;
constexpr int SMEM_BUF_OFF = ((SMEM_PW_OFF + NUM_TAPS * TILE_PX * 16 + 1023) / 1024) * 1024;
constexpr int SMEM_LIMIT = 226 * 1024;

struct WgradParams {
  Geom g;
  const void* gout;      // dense NHWC (B,H,W,Cout), element type T
  float* partial;        // [nsplit][9][Cout][Cin]
  int nchunk, nblocks;   // channel chunks per tap, blocks = 9 * nchunk
  int gb, ng;            // blocks per group, groups
  int nsplit, tiles_per_split;
  int mh;                // M halves (Cout > 128 -> 2)
  int a_atoms;           // 128-byte atoms (columns of 16 KB) per gout buffer
  int a_bufs, b_stages;
  int tmem_cols;
};

template <typename T>
__global__ void __maxnreg__(80) dcn_wgrad_tc_kernel(const WgradParams p) {
  constexpr bool kTF32 = sizeof(T) == 4;
  constexpr int kChunk = ROW_BYTES / sizeof(T);   // channels per block (= UMMA N): 64 bf16 / 32 tf32
  constexpr int kInstrK = kTF32 ? 8 : 16;         // pixels per tcgen05.mma
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_fullB = smem_base + SMEM_BAR_OFF;              // [MAX_B_STAGES]
  const uint32_t bar_emptyB = bar_fullB + 8 * MAX_B_STAGES;         // [MAX_B_STAGES]
  const uint32_t bar_fullG = bar_emptyB + 8 * MAX_B_STAGES;         // [MAX_A_BUFS]
  const uint32_t bar_emptyG = bar_fullG + 8 * MAX_A_BUFS;           // [MAX_A_BUFS]
  const uint32_t bar_done = bar_emptyG + 8 * MAX_A_BUFS;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SMEM_BAR_OFF + 8 * (2 * MAX_B_STAGES + 2 * MAX_A_BUFS + 1));
  const uint32_t poff_s = smem_base + SMEM_POFF_OFF, pw_s = smem_base + SMEM_PW_OFF;
  const uint32_t a_bytes = (uint32_t)p.a_atoms * ATOM_BYTES;
  const uint32_t abuf0 = smem_base + SMEM_BUF_OFF;
  const uint32_t bstage0 = abuf0 + (uint32_t)p.a_bufs * a_bytes;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = blockIdx.x / p.ng, grp = blockIdx.x - split * p.ng;
  const int blk0 = grp * p.gb, blk1 = min(p.nblocks, blk0 + p.gb);
  const int tile0 = split * p.tiles_per_split, tile1 = min(p.g.num_tiles, tile0 + p.tiles_per_split);
  const int tap_first = blk0 / p.nchunk, tap_last = (blk1 - 1) / p.nchunk;

  if (warp == 4 && lane == 0) {
    for (int s = 0; s < MAX_B_STAGES; ++s) { mbar_init(bar_fullB + 8 * s, SAMPLER_WARPS); mbar_init(bar_emptyB + 8 * s, 1); }
    for (int a = 0; a < MAX_A_BUFS; ++a) { mbar_init(bar_fullG + 8 * a, SAMPLER_WARPS); mbar_init(bar_emptyG + 8 * a, 1); }
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(smem_u32(tmem_ptr_smem), (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp >= SAMPLER_WARP0) {
    // =========================================================================== samplers (+ gout tile loader)
    const int st = threadIdx.x - SAMPLER_WARP0 * 32;
    const int row_in_pass = st >> 3, cg = st & 7;
    const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.g.x);
    const uint8_t* gb = reinterpret_cast<const uint8_t*>(p.gout);
    const int ntaps = tap_last - tap_first + 1;
    const int row_pieces = p.g.Cout * (int)sizeof(T) / 16;         // 16-byte pieces per gout pixel row
    RingCursor rb, ra;
    for (int tile = tile0; tile < tile1; ++tile) {
      int b, ty0, tx0;
      tile_coords(p.g, tile, b, ty0, tx0);
      named_bar_sync(1, NS);   // previous tile's parameters are no longer read
      // ---- (pixel, tap) parameters of the job's taps
      for (int e = st; e < ntaps * TILE_PX; e += NS) {
        const int kl = e >> 7, r = e & (TILE_PX - 1);
        const TapGeom t = make_tap_geom<T>(p.g, b, ty0 + (r >> 4), tx0 + (r & (TILE_W - 1)), tap_first + kl);
        sts16(poff_s + (uint32_t)e * 16u, make_uint4(t.off[0], t.off[1], t.off[2], t.off[3]));
        sts16(pw_s + (uint32_t)e * 16u, make_uint4(Blend<T>::weight(t.w[0] * t.m), Blend<T>::weight(t.w[1] * t.m),
                                                   Blend<T>::weight(t.w[2] * t.m), Blend<T>::weight(t.w[3] * t.m)));
      }
      // ---- gout tile -> A buffer (MN-major: row = pixel, 128-byte atoms along Cout, pieces XOR-swizzled by row & 7)
      mbar_wait(bar_emptyG + 8 * ra.s, ra.ph ^ 1u);
      const uint32_t abuf = abuf0 + ra.s * a_bytes;
      for (int pc = st; pc < TILE_PX * row_pieces; pc += NS) {
        const int r = pc / row_pieces, j = pc - r * row_pieces;
        const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (y < p.g.H && x < p.g.W)
          v = ldg16(gb + ((size_t)((b * p.g.H + y) * p.g.W + x) * row_pieces + j) * 16);
        sts16(abuf + (uint32_t)(j >> 3) * ATOM_BYTES + (uint32_t)r * ROW_BYTES + (uint32_t)(((j & 7) ^ (r & 7)) << 4), v);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_fullG + 8 * ra.s);
      ra.advance((uint32_t)p.a_bufs);
      named_bar_sync(1, NS);   // parameters visible to every sampler
      // ---- one sampled column tile per block, rolling pipeline (loads of block j+1 in flight while j is blended)
      uint4 c[PASSES][4], pw[PASSES];
      auto issue = [&](int blk, int pass) {
        const int tap = blk / p.nchunk, cc = blk - tap * p.nchunk;
        const uint32_t idx = (uint32_t)((tap - tap_first) * TILE_PX + pass * ROWS_PER_PASS + row_in_pass) * 16u;
        const uint4 po = lds16(poff_s + idx);
        pw[pass] = lds16(pw_s + idx);
        const uint32_t cbyte = (uint32_t)(cc * ROW_BYTES + cg * 16);
        c[pass][0] = ldg16(xb + (po.x + cbyte));
        c[pass][1] = ldg16(xb + (po.y + cbyte));
        c[pass][2] = ldg16(xb + (po.z + cbyte));
        c[pass][3] = ldg16(xb + (po.w + cbyte));
      };
#pragma unroll
      for (int pass = 0; pass < PASSES; ++pass) issue(blk0, pass);
      for (int blk = blk0; blk < blk1; ++blk) {
        mbar_wait(bar_emptyB + 8 * rb.s, rb.ph ^ 1u);
        const uint32_t bst = bstage0 + rb.s * ATOM_BYTES;
#pragma unroll
        for (int pass = 0; pass < PASSES; ++pass) {
          const int r = pass * ROWS_PER_PASS + row_in_pass;
          sts16(bst + r * ROW_BYTES + ((cg ^ (r & 7)) << 4), Blend<T>::run(c[pass], pw[pass]));
          if (blk + 1 < blk1) issue(blk + 1, pass);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_fullB + 8 * rb.s);
        rb.advance((uint32_t)p.b_stages);
      }
    }
  } else if (warp == 5) {
    // =========================================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(kTF32, 128, kChunk, /*a_mn_major=*/true, /*b_mn_major=*/true);
      const int atoms_per_half = 128 * (int)sizeof(T) / ROW_BYTES;   // 128 output channels = 2 bf16 / 4 tf32 atoms
      RingCursor rb, ra;
      for (int tile = tile0; tile < tile1; ++tile) {
        mbar_wait(bar_fullG + 8 * ra.s, ra.ph, 100);
        const uint32_t abuf = abuf0 + ra.s * a_bytes;
        for (int blk = blk0; blk < blk1; ++blk) {
          mbar_wait(bar_fullB + 8 * rb.s, rb.ph, 100);
          tc_fence_after();
          const uint32_t bst = bstage0 + rb.s * ATOM_BYTES;
          for (int mhalf = 0; mhalf < p.mh; ++mhalf) {
            const uint32_t d_tmem = tmem_base + (uint32_t)(((blk - blk0) * p.mh + mhalf) * kChunk);
            const uint32_t a_addr = abuf + (uint32_t)(mhalf * atoms_per_half) * ATOM_BYTES;
#pragma unroll 4
            for (int kk = 0; kk < TILE_PX / kInstrK; ++kk) {
              const uint64_t adesc = umma_desc_sw128(a_addr + kk * kInstrK * ROW_BYTES, ATOM_BYTES, 1024);
              const uint64_t bdesc = umma_desc_sw128(bst + kk * kInstrK * ROW_BYTES, ATOM_BYTES, 1024);
              umma<kTF32>(d_tmem, adesc, bdesc, idesc, (tile > tile0 || kk > 0) ? 1u : 0u);
            }
          }
          umma_commit(bar_emptyB + 8 * rb.s);
          rb.advance((uint32_t)p.b_stages);
        }
        umma_commit(bar_emptyG + 8 * ra.s);
        ra.advance((uint32_t)p.a_bufs);
      }
      umma_commit(bar_done);
    }
  } else if (warp < EPI_WARPS) {
    // =========================================================================== epilogue: TMEM -> fp32 partial
    mbar_wait(bar_done, 0u, 2000);
    tc_fence_after();
    const int row = warp * 32 + lane;
    float* part = p.partial + (size_t)split * NUM_TAPS * p.g.Cout * p.g.Cin;
    for (int blk = blk0; blk < blk1; ++blk) {
      const int tap = blk / p.nchunk, cc = blk - tap * p.nchunk;
      for (int mhalf = 0; mhalf < p.mh; ++mhalf) {
        const int o = mhalf * 128 + row;
        const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(((blk - blk0) * p.mh + mhalf) * kChunk);
        float* dst = part + ((size_t)tap * p.g.Cout + o) * p.g.Cin + cc * kChunk;
        for (int c0 = 0; c0 < kChunk; c0 += 16) {
          uint32_t raw[16];
          tmem_ld16(t_row + (uint32_t)c0, raw);
          tmem_ld_wait();
          if (o < p.g.Cout) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              reinterpret_cast<uint4*>(dst + c0)[j] = make_uint4(raw[4 * j], raw[4 * j + 1], raw[4 * j + 2], raw[4 * j + 3]);
          }
        }
      }
    }
    tc_fence_before();
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}

// gW[(o*Cin + c)*9 + tap] = sum_s partial[s][tap][o][c]  (fixed order)
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ gw,
                                                           int Cout, int Cin, int nsplit) {
  const int64_t n = (int64_t)NUM_TAPS * Cout * Cin;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < nsplit; ++k) s += __ldcs(partial + (int64_t)k * n + i);
    const int c = (int)(i % Cin);
    const int64_t t = i / Cin;
    const int o = (int)(t % Cout), tap = (int)(t / Cout);
    gw[((int64_t)o * Cin + c) * NUM_TAPS + tap] = s;
  }
}

int pow2_cols(int v) {
  int p = 32;
  while (p < v) p <<= 1;
  return p;
}

struct WgradPlan {
  int nchunk, nblocks, gb, ng, nsplit, tiles_per_split, mh, a_atoms, a_bufs, b_stages, tmem_cols, smem_bytes;
  size_t partial_bytes;
  bool ok;
};

WgradPlan plan_wgrad(const cpoly_dcn_fwd_args& a) {
  WgradPlan w{};
  const int esize = a.dtype == CPOLY_BF16 ? 2 : 4;
  const int kchunk = ROW_BYTES / esize;
  w.nchunk = a.Cin / kchunk;
  w.nblocks = NUM_TAPS * w.nchunk;
  w.mh = a.Cout > 128 ? 2 : 1;
  const int gbmax = 512 / (kchunk * w.mh);
  w.ng = (w.nblocks + gbmax - 1) / gbmax;
  w.gb = (w.nblocks + w.ng - 1) / w.ng;
  w.ng = (w.nblocks + w.gb - 1) / w.gb;
  const int tiles = a.B * ((a.W + TILE_W - 1) / TILE_W) * ((a.H + TILE_H - 1) / TILE_H);
  int nsplit = 148 / w.ng;
  if (nsplit < 1) nsplit = 1;
  if (nsplit > tiles) nsplit = tiles;
  w.tiles_per_split = (tiles + nsplit - 1) / nsplit;
  w.nsplit = (tiles + w.tiles_per_split - 1) / w.tiles_per_split;
  const int atoms_filled = a.Cout * esize / ROW_BYTES;            // atoms holding real gout channels
  const int atoms_read = w.mh * 128 * esize / ROW_BYTES;          // atoms the M=128 MMAs touch
  w.a_atoms = atoms_filled > atoms_read ? atoms_filled : atoms_read;
  const int a_bytes = w.a_atoms * ATOM_BYTES;
  const int avail = SMEM_LIMIT - SMEM_BUF_OFF;
  w.a_bufs = (avail - 2 * ATOM_BYTES) / a_bytes >= 2 ? 2 : 1;
  w.b_stages = (avail - w.a_bufs * a_bytes) / ATOM_BYTES;
  if (w.b_stages > MAX_B_STAGES) w.b_stages = MAX_B_STAGES;
  w.tmem_cols = pow2_cols(w.gb * w.mh * kchunk);
  w.smem_bytes = SMEM_BUF_OFF + w.a_bufs * a_bytes + w.b_stages * ATOM_BYTES + 1024;
  w.partial_bytes = (size_t)w.nsplit * NUM_TAPS * a.Cout * a.Cin * sizeof(float);
  w.ok = w.b_stages >= 2 && w.tmem_cols <= 512 && a.Cout * esize % ROW_BYTES == 0;
  return w;
}

bool dense_nhwc4(const cpoly_tensor4& t, int C, int H, int W) {
  return t.sc == 1 && t.sw == C && t.sh == (int64_t)W * C && t.sb == (int64_t)H * W * C;
}

}  // namespace

bool dcn_tc_wgrad_eligible(const cpoly_dcn_bwd_args& a) {
  cpoly_dcn_fwd_args f = a.f;
  f.y = a.gout;            // same shape/layout requirements as the forward output
  f.weight_packed = nullptr;
  if (!dcn_tc_eligible(f)) return false;
  if (!dense_nhwc4(a.gout, a.f.Cout, a.f.H, a.f.W)) return false;
  return plan_wgrad(a.f).ok;
}

size_t dcn_tc_wgrad_workspace(const cpoly_dcn_bwd_args& a) { return plan_wgrad(a.f).partial_bytes + 256; }

int dcn_wgrad_tc(const cpoly_dcn_bwd_args& a, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  const cpoly_dcn_fwd_args& f = a.f;
  const WgradPlan w = plan_wgrad(f);
  CPOLY_REQUIRE(w.ok, CPOLY_EUNSUPPORTED, "wgrad plan does not fit");
  CPOLY_REQUIRE(workspace && workspace_bytes >= w.partial_bytes, CPOLY_EWORKSPACE,
                "dcn wgrad workspace too small: need %zu bytes, have %zu", w.partial_bytes, workspace_bytes);
  WgradParams p;
  p.g.x = f.x.ptr; p.g.offset = f.offset; p.g.mask = f.mask; p.g.offset_bias = f.offset_bias; p.g.mask_bias = f.mask_bias;
  p.g.B = f.B; p.g.Cin = f.Cin; p.g.Cout = f.Cout; p.g.H = f.H; p.g.W = f.W;
  p.g.om_dtype = f.om_dtype; p.g.mask_is_logit = f.mask_is_logit;
  p.g.tiles_x = (f.W + TILE_W - 1) / TILE_W; p.g.tiles_y = (f.H + TILE_H - 1) / TILE_H;
  p.g.num_tiles = f.B * p.g.tiles_x * p.g.tiles_y;
  p.gout = a.gout.ptr;
  p.partial = reinterpret_cast<float*>(workspace);
  p.nchunk = w.nchunk; p.nblocks = w.nblocks; p.gb = w.gb; p.ng = w.ng; p.nsplit = w.nsplit;
  p.tiles_per_split = w.tiles_per_split; p.mh = w.mh; p.a_atoms = w.a_atoms; p.a_bufs = w.a_bufs; p.b_stages = w.b_stages;
  p.tmem_cols = w.tmem_cols;
  const int grid = w.nsplit * w.ng;
  if (f.dtype == CPOLY_BF16) {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_wgrad_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, w.smem_bytes));
    dcn_wgrad_tc_kernel<__nv_bfloat16><<<grid, NTHREADS, w.smem_bytes, stream>>>(p);
  } else {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_wgrad_tc_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, w.smem_bytes));
    dcn_wgrad_tc_kernel<float><<<grid, NTHREADS, w.smem_bytes, stream>>>(p);
  }
  if (int rc = check_launch("dcn_wgrad_tc_kernel")) return rc;
  const int64_t n = (int64_t)NUM_TAPS * f.Cout * f.Cin;
  wgrad_reduce_kernel<<<(unsigned)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184), 256, 0, stream>>>(p.partial, a.gweight, f.Cout, f.Cin, w.nsplit);
  return check_launch("wgrad_reduce_kernel");
}

}  // namespace cpoly
