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
        sts16(abuf + (uint32_t)(j >> 3) * ATOM_BYTES + (uint32_t)r * ROW_BYTES + (mn_major_piece<kTF32>(j & 7, r) << 4), v);
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
          sts16(bst + r * ROW_BYTES + (mn_major_piece<kTF32>(cg, r) << 4), Blend<T>::run(c[pass], pw[pass]));
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
    // (whole warp walks the loops, one elected lane issues: see elect_one() in tc_common.cuh)
    {
      const uint32_t idesc = umma_idesc(kTF32, 128, kChunk, /*a_mn_major=*/true, /*b_mn_major=*/true);
      const int atoms_per_half = 128 * (int)sizeof(T) / ROW_BYTES;   // 128 output channels = 2 bf16 / 4 tf32 atoms
      RingCursor rb, ra;
      for (int tile = tile0; tile < tile1; ++tile) {
        mbar_wait(bar_fullG + 8 * ra.s, ra.ph, 32);
        const uint32_t abuf = abuf0 + ra.s * a_bytes;
        for (int blk = blk0; blk < blk1; ++blk) {
          mbar_wait(bar_fullB + 8 * rb.s, rb.ph, 32);
          tc_fence_after();
          const uint32_t bst = bstage0 + rb.s * ATOM_BYTES;
          if (elect_one()) {
            for (int mhalf = 0; mhalf < p.mh; ++mhalf) {
              const uint32_t d_tmem = tmem_base + (uint32_t)(((blk - blk0) * p.mh + mhalf) * kChunk);
              // MN-major operands: 16-bit elements use SWIZZLE_128B (8-row atoms); 32-bit elements (tf32) need
              // SWIZZLE_128B_BASE32B (4-row atoms, 512 B apart) - see umma_desc_sw128
              constexpr uint32_t kSbo = kTF32 ? 512u : 1024u, kLayout = kTF32 ? 1u : 2u;
              const uint64_t adesc0 = umma_desc_sw128(abuf + (uint32_t)(mhalf * atoms_per_half) * ATOM_BYTES, ATOM_BYTES, kSbo, kLayout);
              const uint64_t bdesc0 = umma_desc_sw128(bst, ATOM_BYTES, kSbo, kLayout);
#pragma unroll
              for (int kk = 0; kk < TILE_PX / kInstrK; ++kk) {   // the start address advances by whole 128-byte rows
                const uint64_t step = (uint64_t)((kk * kInstrK * ROW_BYTES) >> 4);
                umma<kTF32>(d_tmem, adesc0 + step, bdesc0 + step, idesc, (tile > tile0 || kk > 0) ? 1u : 0u);
              }
            }
            umma_commit(bar_emptyB + 8 * rb.s);
            if (blk == blk1 - 1) {
              umma_commit(bar_emptyG + 8 * ra.s);
              if (tile == tile1 - 1) umma_commit(bar_done);
            }
          }
          __syncwarp();
          rb.advance((uint32_t)p.b_stages);
        }
        ra.advance((uint32_t)p.a_bufs);
      }
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


// =====================================================================================================================
// Part 2: data / coordinate gradients.
//
//   gcol[px,(tap,c)] = sum_o gout[px,o] * W[o,c,tap]                       (tcgen05: M = 128 pixels, N = channel chunk, K = Cout)
//   grad_mask[px,tap]      = sum_c gcol * v_c(p)                           v = bilinear sample (unmasked)
//   grad_offset[px,2tap..] = m * sum_c gcol * dv_c/dp                      floor-based corner differences
//   gcol is ALSO stored, one coalesced 16-byte store per lane, for the atomics-free col2im that follows
//   (dcn_col2im.cu: grad_input[corner,c] += gcol * m * w_corner in gather form, bit-reproducible).
//
// Persistent warp-specialised CTA; a work item is (128-pixel tile, tap group) as in the forward kernel.
//   warps 0-3  TMEM -> shared: the gcol tile of one block, converted to T, in the sampler's row layout
//   warp  4    bulk-copies the block's pre-packed dgrad weights (K-major, rows = input channels) into a 2-stage ring
//   warp  5    issues the MMAs (A = the item's gout tile, loaded once per item by the samplers)
//   warps 6-21 gather the 4 corners of every (pixel, tap) exactly like the forward sampler, combine them with gcol,
//              reduce the coordinate gradients over the 8 lanes of a pixel row (written once, no atomics) and store
//              their piece of the gcol row.
namespace {

constexpr int D_BUFS = 2;          // TMEM accumulators and shared gcol tiles
constexpr int DG_POFF_OFF = 1024;  // three parameter arrays of taps_per_item * 128 * 16 bytes follow (runtime size)

struct DgradParams {
  Geom g;
  const void* gout;        // dense NHWC (B,H,W,Cout), element type T
  const void* wpacked;     // dgrad weight image, see pack_dgrad_weights_kernel
  void* gcol;              // [B*H*W][9][Cin] element type T for dcn_col2im.cu; or NULL (grad_input not requested)
  cpoly_tensor4 goffset, gmask;   // fp32, any strides; ptr may be NULL
  int nchunk;              // channel chunks per tap
  int split, taps_per_item, num_items;
  int k_atoms;             // 128-byte atoms along K = Cout
  int a_bufs;
  int w_stage_bytes;
  int w_stages;            // 2, or 1 when shared memory is short (fp32 with Cout = 256: the gout tile alone is 128 KB)
  int param_bytes;         // bytes of ONE parameter array = taps_per_item * 128 * 16
};

template <typename T>
__global__ void __maxnreg__(80) dcn_dgrad_tc_kernel(const DgradParams p) {
  constexpr bool kTF32 = sizeof(T) == 4;
  constexpr int kChunk = ROW_BYTES / sizeof(T);   // channels per block (= UMMA N): 64 bf16 / 32 tf32
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_fullG = smem_base, bar_emptyG = bar_fullG + 16, bar_Wfull = bar_emptyG + 16, bar_Wempty = bar_Wfull + 16;
  const uint32_t bar_Dfull = bar_Wempty + 16, bar_Dempty = bar_Dfull + 16, bar_gfull = bar_Dempty + 16, bar_gempty = bar_gfull + 16;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + 8 * 16);
  const uint32_t poff_s = smem_base + DG_POFF_OFF, pw_s = poff_s + (uint32_t)p.param_bytes, pm_s = pw_s + (uint32_t)p.param_bytes;
  const uint32_t a_bytes = (uint32_t)p.k_atoms * ATOM_BYTES;
  const uint32_t abuf0 = smem_base + ((DG_POFF_OFF + 3u * (uint32_t)p.param_bytes + 1023u) & ~1023u);
  const uint32_t wst0 = abuf0 + (uint32_t)p.a_bufs * a_bytes;
  const uint32_t gcol0 = wst0 + (uint32_t)p.w_stages * (uint32_t)p.w_stage_bytes;
  const bool w2 = p.w_stages == 2;   // weight ring cursor of block n: stage n & 1 / phase (n >> 1) & 1, or stage 0 / phase n & 1
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int blocks_per_item = p.taps_per_item * p.nchunk;

  if (warp == 4 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_fullG + 8 * i, SAMPLER_WARPS); mbar_init(bar_emptyG + 8 * i, 1);
      mbar_init(bar_Wfull + 8 * i, 1);             mbar_init(bar_Wempty + 8 * i, 1);
      mbar_init(bar_Dfull + 8 * i, 1);             mbar_init(bar_Dempty + 8 * i, EPI_WARPS * 32);
      mbar_init(bar_gfull + 8 * i, EPI_WARPS);     mbar_init(bar_gempty + 8 * i, SAMPLER_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(smem_u32(tmem_ptr_smem), (uint32_t)(kChunk * D_BUFS < 32 ? 32 : kChunk * D_BUFS));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp >= SAMPLER_WARP0) {
    // =========================================================================== samplers
    const int st = threadIdx.x - SAMPLER_WARP0 * 32;
    const int row_in_pass = st >> 3, cg = st & 7;
    const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.g.x);
    const uint8_t* gb = reinterpret_cast<const uint8_t*>(p.gout);
    const int row_pieces = p.g.Cout * (int)sizeof(T) / 16;
    void* gmp = p.gmask.ptr;
    void* gop = p.goffset.ptr;
    RingCursor ra;
    uint32_t nblk = 0;   // blocks processed by this CTA so far (D / gcol buffer cursor)
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      const int tile = item / p.split, tap0 = (item - tile * p.split) * p.taps_per_item;
      int b, ty0, tx0;
      tile_coords(p.g, tile, b, ty0, tx0);
      named_bar_sync(1, NS);
      for (int e = st; e < p.taps_per_item * TILE_PX; e += NS) {
        const int kl = e >> 7, r = e & (TILE_PX - 1);
        const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
        const TapGeom t = make_tap_geom<T>(p.g, b, y, x, tap0 + kl);
        const uint32_t flags = (uint32_t)(t.valid[0] | (t.valid[1] << 1) | (t.valid[2] << 2) | (t.valid[3] << 3)) |
                               ((y < p.g.H && x < p.g.W) ? 16u : 0u);
        sts16(poff_s + (uint32_t)e * 16u, make_uint4(t.off[0], t.off[1], t.off[2], t.off[3]));
        sts16(pw_s + (uint32_t)e * 16u, make_uint4(__float_as_uint(t.w[0]), __float_as_uint(t.w[1]), __float_as_uint(t.w[2]), __float_as_uint(t.w[3])));
        sts16(pm_s + (uint32_t)e * 16u, make_uint4(__float_as_uint(t.m), __float_as_uint(t.ly), __float_as_uint(t.lx), flags));
      }
      // gout tile -> A buffer (K-major: row = pixel, 128-byte atoms along Cout)
      mbar_wait(bar_emptyG + 8 * ra.s, ra.ph ^ 1u);
      const uint32_t abuf = abuf0 + ra.s * a_bytes;
      for (int pc = st; pc < TILE_PX * row_pieces; pc += NS) {
        const int r = pc / row_pieces, j = pc - r * row_pieces;
        const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1));
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (y < p.g.H && x < p.g.W) v = ldg16(gb + ((size_t)((b * p.g.H + y) * p.g.W + x) * row_pieces + j) * 16);
        sts16(abuf + (uint32_t)(j >> 3) * ATOM_BYTES + (uint32_t)r * ROW_BYTES + (uint32_t)(((j & 7) ^ (r & 7)) << 4), v);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_fullG + 8 * ra.s);
      ra.advance((uint32_t)p.a_bufs);
      named_bar_sync(1, NS);

      uint4 c[PASSES][4];
      auto issue = [&](int q, int pass) {   // corner loads of block q (tap-major, chunk-minor) of this item
        const int kl = q / p.nchunk, cc = q - kl * p.nchunk;
        const uint32_t idx = (uint32_t)(kl * TILE_PX + pass * ROWS_PER_PASS + row_in_pass) * 16u;
        const uint4 po = lds16(poff_s + idx);
        const uint32_t cbyte = (uint32_t)(cc * ROW_BYTES + cg * 16);
        c[pass][0] = ldg16(xb + (po.x + cbyte));
        c[pass][1] = ldg16(xb + (po.y + cbyte));
        c[pass][2] = ldg16(xb + (po.z + cbyte));
        c[pass][3] = ldg16(xb + (po.w + cbyte));
      };
#pragma unroll
      for (int pass = 0; pass < PASSES; ++pass) issue(0, pass);
      // Per (pixel, tap) the three gradients are linear in the four per-corner dot products s_j = sum_c gcol[c] * x_j[c]:
      //   gmask = sum_j w_j s_j,  goff_y = m (lx (s3 - s1) + hx (s2 - s0)),  goff_x = m (ly (s3 - s2) + hy (s1 - s0)),
      // so the channel loop is 4 FMAs per element and the bilinear algebra happens once, after the channel reduction
      // (the first version evaluated val / d val / dy / d val / dx per ELEMENT: 15 operations instead of 4).
      float sc0[PASSES], sc1[PASSES], sc2[PASSES], sc3[PASSES];
      for (int q = 0; q < blocks_per_item; ++q, ++nblk) {
        const int kl = q / p.nchunk, cc = q - kl * p.nchunk;
        const uint32_t d = nblk & 1u, dph = (nblk >> 1) & 1u;
        if (cc == 0) {
#pragma unroll
          for (int pass = 0; pass < PASSES; ++pass) sc0[pass] = sc1[pass] = sc2[pass] = sc3[pass] = 0.f;
        }
        mbar_wait(bar_gfull + 8 * d, dph);
        const uint32_t gt = gcol0 + d * ATOM_BYTES;
#pragma unroll
        for (int pass = 0; pass < PASSES; ++pass) {
          const int r = pass * ROWS_PER_PASS + row_in_pass;
          const uint32_t idx = (uint32_t)(kl * TILE_PX + r) * 16u;
          const uint4 pm = lds16(pm_s + idx);
          const uint32_t flags = pm.w;
          if (flags & 15u) {
            const uint4 gq = lds16(gt + (uint32_t)r * ROW_BYTES + (uint32_t)((cg ^ (r & 7)) << 4));
            const uint32_t* gw = &gq.x;
            const uint32_t* c0w = &c[pass][0].x;
            const uint32_t* c1w = &c[pass][1].x;
            const uint32_t* c2w = &c[pass][2].x;
            const uint32_t* c3w = &c[pass][3].x;
            float a0 = sc0[pass], a1 = sc1[pass], a2 = sc2[pass], a3 = sc3[pass];
            if (kTF32) {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float g = __uint_as_float(gw[e]);
                a0 = fmaf(g, __uint_as_float(c0w[e]), a0); a1 = fmaf(g, __uint_as_float(c1w[e]), a1);
                a2 = fmaf(g, __uint_as_float(c2w[e]), a2); a3 = fmaf(g, __uint_as_float(c3w[e]), a3);
              }
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e) {   // one 32-bit word = two bf16 channels
                const float glo = __uint_as_float(gw[e] << 16), ghi = __uint_as_float(gw[e] & 0xffff0000u);
                a0 = fmaf(glo, __uint_as_float(c0w[e] << 16), a0); a0 = fmaf(ghi, __uint_as_float(c0w[e] & 0xffff0000u), a0);
                a1 = fmaf(glo, __uint_as_float(c1w[e] << 16), a1); a1 = fmaf(ghi, __uint_as_float(c1w[e] & 0xffff0000u), a1);
                a2 = fmaf(glo, __uint_as_float(c2w[e] << 16), a2); a2 = fmaf(ghi, __uint_as_float(c2w[e] & 0xffff0000u), a2);
                a3 = fmaf(glo, __uint_as_float(c3w[e] << 16), a3); a3 = fmaf(ghi, __uint_as_float(c3w[e] & 0xffff0000u), a3);
              }
            }
            sc0[pass] = a0; sc1[pass] = a1; sc2[pass] = a2; sc3[pass] = a3;
            if (p.gcol) {   // this lane's 16 bytes of the (pixel, tap) row of gcol, as the tensor cores produced it
              const int yy = ty0 + (r >> 4), xx = tx0 + (r & (TILE_W - 1));
              const size_t row = (size_t)((b * p.g.H + yy) * p.g.W + xx) * NUM_TAPS + (size_t)(tap0 + kl);
              *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.gcol) + (row * p.g.Cin + (size_t)(cc * kChunk)) * sizeof(T) + (size_t)cg * 16) = gq;
            }
          }
          if (q + 1 < blocks_per_item) issue(q + 1, pass);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_gempty + 8 * d);
        if (cc == p.nchunk - 1) {   // tap complete: reduce over the 8 lanes of each pixel row and write once
#pragma unroll
          for (int pass = 0; pass < PASSES; ++pass) {
#pragma unroll
            for (int sft = 1; sft <= 4; sft <<= 1) {
              sc0[pass] += __shfl_xor_sync(0xffffffffu, sc0[pass], sft);
              sc1[pass] += __shfl_xor_sync(0xffffffffu, sc1[pass], sft);
              sc2[pass] += __shfl_xor_sync(0xffffffffu, sc2[pass], sft);
              sc3[pass] += __shfl_xor_sync(0xffffffffu, sc3[pass], sft);
            }
            float sy_p = 0.f, sx_p = 0.f;
            if (cg == 0) {
              const int r = pass * ROWS_PER_PASS + row_in_pass;
              const uint4 pm = lds16(pm_s + (uint32_t)(kl * TILE_PX + r) * 16u);
              if (pm.w & 16u) {
                const int y = ty0 + (r >> 4), x = tx0 + (r & (TILE_W - 1)), k = tap0 + kl;
                const float m = __uint_as_float(pm.x), ly = __uint_as_float(pm.y), lx = __uint_as_float(pm.z);
                const float hy = 1.0f - ly, hx = 1.0f - lx;
                const uint4 pwv = lds16(pw_s + (uint32_t)(kl * TILE_PX + r) * 16u);
                // a corner outside the image contributes nothing (its loads read a dummy in-range address)
                const float s0 = (pm.w & 1u) ? sc0[pass] : 0.f, s1 = (pm.w & 2u) ? sc1[pass] : 0.f;
                const float s2 = (pm.w & 4u) ? sc2[pass] : 0.f, s3 = (pm.w & 8u) ? sc3[pass] : 0.f;
                sy_p = fmaf(lx, s3 - s1, hx * (s2 - s0));
                sx_p = fmaf(ly, s3 - s2, hy * (s1 - s0));
                if (gmp) {
                  float gm = fmaf(__uint_as_float(pwv.w), s3, fmaf(__uint_as_float(pwv.z), s2, fmaf(__uint_as_float(pwv.y), s1, __uint_as_float(pwv.x) * s0)));
                  if (p.g.mask_is_logit) gm *= m * (1.f - m);
                  st_dyn(gmp, (int64_t)b * p.gmask.sb + (int64_t)k * p.gmask.sc + (int64_t)y * p.gmask.sh + (int64_t)x * p.gmask.sw, gm, p.g.gom_dtype);
                }
                if (gop) {
                  const int64_t ob = (int64_t)b * p.goffset.sb + (int64_t)y * p.goffset.sh + (int64_t)x * p.goffset.sw;
                  st_dyn(gop, ob + (int64_t)(2 * k) * p.goffset.sc, m * sy_p, p.g.gom_dtype);
                  st_dyn(gop, ob + (int64_t)(2 * k + 1) * p.goffset.sc, m * sx_p, p.g.gom_dtype);
                }
              }
            }
          }
        }
      }
    }
  } else if (warp == 4) {
    // =========================================================================== dgrad weight loader
    if (lane == 0) {
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.wpacked);
      uint32_t nblk = 0;
      for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
        const int tap0 = (item % p.split) * p.taps_per_item;
        for (int q = 0; q < blocks_per_item; ++q, ++nblk) {
          const uint32_t ws = w2 ? (nblk & 1u) : 0u, wph = w2 ? ((nblk >> 1) & 1u) : (nblk & 1u);
          mbar_wait(bar_Wempty + 8 * ws, wph ^ 1u, 200);
          mbar_arrive_expect_tx(bar_Wfull + 8 * ws, (uint32_t)p.w_stage_bytes);
          bulk_copy_g2s(wst0 + ws * (uint32_t)p.w_stage_bytes, wsrc + (size_t)(tap0 * p.nchunk + q) * p.w_stage_bytes,
                        (uint32_t)p.w_stage_bytes, bar_Wfull + 8 * ws);
        }
      }
    }
  } else if (warp == 5) {
    // =========================================================================== MMA issuer
    // (whole warp walks the loops, one elected lane issues)
    {
      const uint32_t idesc = umma_idesc(kTF32, TILE_PX, kChunk);
      RingCursor ra;
      uint32_t nblk = 0;
      for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
        mbar_wait(bar_fullG + 8 * ra.s, ra.ph, 32);
        const uint32_t abuf = abuf0 + ra.s * a_bytes;
        for (int q = 0; q < blocks_per_item; ++q, ++nblk) {
          const uint32_t d = nblk & 1u, dph = (nblk >> 1) & 1u;
          const uint32_t ws = w2 ? d : 0u, wph = w2 ? dph : (nblk & 1u);
          mbar_wait(bar_Wfull + 8 * ws, wph, 32);
          mbar_wait(bar_Dempty + 8 * d, dph ^ 1u, 32);
          tc_fence_after();
          const uint32_t wst = wst0 + ws * (uint32_t)p.w_stage_bytes;
          const uint32_t d_tmem = tmem_base + d * (uint32_t)kChunk;
          if (elect_one()) {
            for (int ka = 0; ka < p.k_atoms; ++ka) {
              const uint64_t adesc = umma_desc_sw128(abuf + (uint32_t)ka * ATOM_BYTES);
              const uint64_t bdesc = umma_desc_sw128(wst + (uint32_t)ka * (uint32_t)(kChunk * ROW_BYTES));
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                umma<kTF32>(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (ka | kk) != 0 ? 1u : 0u);
            }
            umma_commit(bar_Wempty + 8 * ws);
            umma_commit(bar_Dfull + 8 * d);
            if (q == blocks_per_item - 1) umma_commit(bar_emptyG + 8 * ra.s);
          }
          __syncwarp();
        }
        ra.advance((uint32_t)p.a_bufs);
      }
    }
  } else {
    // =========================================================================== TMEM -> shared gcol tile
    const int r = warp * 32 + lane;
    uint32_t nblk = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      for (int q = 0; q < blocks_per_item; ++q, ++nblk) {
        const uint32_t d = nblk & 1u, dph = (nblk >> 1) & 1u;
        mbar_wait(bar_Dfull + 8 * d, dph, 50);
        mbar_wait(bar_gempty + 8 * d, dph ^ 1u, 50);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + d * (uint32_t)kChunk;
        const uint32_t grow = gcol0 + d * ATOM_BYTES + (uint32_t)r * ROW_BYTES;
#pragma unroll
        for (int c0 = 0; c0 < kChunk; c0 += 16) {
          uint32_t raw[16];
          tmem_ld16(t_row + (uint32_t)c0, raw);
          tmem_ld_wait();
          if (kTF32) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              sts16(grow + (uint32_t)((((c0 >> 2) + j) ^ (r & 7)) << 4), make_uint4(raw[4 * j], raw[4 * j + 1], raw[4 * j + 2], raw[4 * j + 3]));
          } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              uint4 v;
              v.x = pack_bf16x2(__uint_as_float(raw[8 * j]), __uint_as_float(raw[8 * j + 1]));
              v.y = pack_bf16x2(__uint_as_float(raw[8 * j + 2]), __uint_as_float(raw[8 * j + 3]));
              v.z = pack_bf16x2(__uint_as_float(raw[8 * j + 4]), __uint_as_float(raw[8 * j + 5]));
              v.w = pack_bf16x2(__uint_as_float(raw[8 * j + 6]), __uint_as_float(raw[8 * j + 7]));
              sts16(grow + (uint32_t)((((c0 >> 3) + j) ^ (r & 7)) << 4), v);
            }
          }
        }
        tc_fence_before();
        mbar_arrive(bar_Dempty + 8 * d);
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_gfull + 8 * d);
      }
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)(kChunk * D_BUFS < 32 ? 32 : kChunk * D_BUFS));
  }
}

// weight (Cout,Cin,3,3) fp32 -> dgrad image: per (tap, channel chunk) a blob of k_atoms x [kChunk rows (input channel)
// x 128 B of K = output channels], 16-byte pieces XOR-swizzled by (row & 7): the K-major SWIZZLE_128B B operand.
template <typename T>
__global__ void pack_dgrad_weights_kernel(const float* __restrict__ w, uint4* __restrict__ out, int Cin, int Cout) {
  constexpr int kVec = 16 / sizeof(T);
  constexpr int kChunk = ROW_BYTES / sizeof(T);
  const int nchunk = Cin / kChunk, k_atoms = Cout / kChunk;
  const int64_t total = (int64_t)NUM_TAPS * nchunk * k_atoms * kChunk * 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i & 7);
    int64_t t = i >> 3;
    const int row = (int)(t % kChunk); t /= kChunk;
    const int ka = (int)(t % k_atoms); t /= k_atoms;
    const int cc = (int)(t % nchunk);
    const int tap = (int)(t / nchunk);
    const int c = cc * kChunk + row;
    uint32_t words[4];
    if (sizeof(T) == 4) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int o = ka * kChunk + j * kVec + e;
        words[e] = to_tf32(w[((int64_t)o * Cin + c) * NUM_TAPS + tap]);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int o = ka * kChunk + j * kVec + 2 * e;
        words[e] = pack_bf16x2(w[((int64_t)o * Cin + c) * NUM_TAPS + tap], w[((int64_t)(o + 1) * Cin + c) * NUM_TAPS + tap]);
      }
    }
    const int64_t blob = (int64_t)tap * nchunk + cc;
    out[((blob * k_atoms + ka) * kChunk + row) * 8 + (j ^ (row & 7))] = make_uint4(words[0], words[1], words[2], words[3]);
  }
}

struct DgradPlan {
  int nchunk, k_atoms, a_bufs, w_stage_bytes, w_stages, split, param_bytes, smem_bytes;
  size_t packed_bytes;
  bool ok;
};

DgradPlan plan_dgrad(const cpoly_dcn_fwd_args& a) {
  DgradPlan d{};
  const int esize = a.dtype == CPOLY_BF16 ? 2 : 4;
  const int kchunk = ROW_BYTES / esize;
  d.nchunk = a.Cin / kchunk;
  d.k_atoms = a.Cout / kchunk;
  d.w_stage_bytes = d.k_atoms * kchunk * ROW_BYTES;
  d.packed_bytes = (size_t)NUM_TAPS * a.Cin * a.Cout * esize;
  const int a_bytes = d.k_atoms * ATOM_BYTES;
  const int tiles = a.B * ((a.W + TILE_W - 1) / TILE_W) * ((a.H + TILE_H - 1) / TILE_H);
  d.split = tiles >= 148 ? 1 : (tiles * 3 >= 148 ? 3 : 9);
  d.ok = false;
  if (a.Cout % kchunk != 0 || a.Cin % kchunk != 0) return d;
  // first choice: the natural tap split and a 2-stage weight ring; when that does not fit (fp32 with Cout = 256: the gout
  // tile alone is 128 KB) stage at most 3 taps of parameters per item and keep ONE weight stage
  for (int attempt = 0; attempt < 2 && !d.ok; ++attempt) {
    if (attempt == 1) { d.split = d.split < 3 ? 3 : d.split; }
    d.w_stages = attempt == 0 ? 2 : 1;
    d.param_bytes = (NUM_TAPS / d.split) * TILE_PX * 16;
    const int buf_off = (DG_POFF_OFF + 3 * d.param_bytes + 1023) / 1024 * 1024;
    const int fixed = buf_off + d.w_stages * d.w_stage_bytes + D_BUFS * ATOM_BYTES;
    if (fixed + a_bytes > SMEM_LIMIT) continue;
    d.a_bufs = (SMEM_LIMIT - fixed) / a_bytes >= 2 ? 2 : 1;
    d.smem_bytes = fixed + d.a_bufs * a_bytes + 1024;
    d.ok = true;
  }
  return d;
}

}  // namespace

bool dcn_tc_dgrad_eligible(const cpoly_dcn_bwd_args& a) {
  cpoly_dcn_fwd_args f = a.f;
  f.y = a.gout;
  f.weight_packed = nullptr;
  if (!dcn_tc_eligible(f)) return false;
  if (!dense_nhwc4(a.gout, a.f.Cout, a.f.H, a.f.W)) return false;
  if (a.gx.ptr && !dcn_col2im_eligible(a)) return false;
  return plan_dgrad(a.f).ok;
}

static size_t dgrad_packed_aligned(const cpoly_dcn_fwd_args& f) { return (plan_dgrad(f).packed_bytes + 255) / 256 * 256; }

// [dgrad weight image | gcol + boxes of the col2im (only when grad_input is requested)]
size_t dcn_tc_dgrad_workspace(const cpoly_dcn_bwd_args& a) {
  return dgrad_packed_aligned(a.f) + (a.gx.ptr ? dcn_col2im_workspace(a) : 0) + 256;
}

int dcn_dgrad_tc(const cpoly_dcn_bwd_args& a, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  const cpoly_dcn_fwd_args& f = a.f;
  const DgradPlan d = plan_dgrad(f);
  CPOLY_REQUIRE(d.ok, CPOLY_EUNSUPPORTED, "dgrad plan does not fit");
  const size_t need = dcn_tc_dgrad_workspace(a) - 256;
  CPOLY_REQUIRE(workspace && workspace_bytes >= need && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0, CPOLY_EWORKSPACE,
                "dcn dgrad workspace too small or misaligned: need %zu bytes, have %zu", need, workspace_bytes);
  CPOLY_REQUIRE(!a.gx.ptr || dcn_col2im_eligible(a), CPOLY_EUNSUPPORTED, "grad_input layout not supported by the col2im");
  uint8_t* c2i_ws = reinterpret_cast<uint8_t*>(workspace) + dgrad_packed_aligned(f);   // gcol first, boxes behind it
  const bool bf = f.dtype == CPOLY_BF16;
  const int64_t pieces = (int64_t)d.packed_bytes / 16;
  const unsigned pblocks = (unsigned)((pieces + 255) / 256);
  if (bf) pack_dgrad_weights_kernel<__nv_bfloat16><<<pblocks, 256, 0, stream>>>(f.weight, reinterpret_cast<uint4*>(workspace), f.Cin, f.Cout);
  else pack_dgrad_weights_kernel<float><<<pblocks, 256, 0, stream>>>(f.weight, reinterpret_cast<uint4*>(workspace), f.Cin, f.Cout);
  if (int rc = check_launch("pack_dgrad_weights_kernel")) return rc;
  DgradParams p;
  p.g.x = f.x.ptr; p.g.offset = f.offset; p.g.mask = f.mask; p.g.offset_bias = f.offset_bias; p.g.mask_bias = f.mask_bias;
  p.g.B = f.B; p.g.Cin = f.Cin; p.g.Cout = f.Cout; p.g.H = f.H; p.g.W = f.W;
  p.g.om_dtype = f.om_dtype; p.g.mask_is_logit = f.mask_is_logit; p.g.gom_dtype = a.gom_dtype;
  p.g.tiles_x = (f.W + TILE_W - 1) / TILE_W; p.g.tiles_y = (f.H + TILE_H - 1) / TILE_H;
  p.g.num_tiles = f.B * p.g.tiles_x * p.g.tiles_y;
  p.gout = a.gout.ptr; p.wpacked = workspace; p.gcol = a.gx.ptr ? c2i_ws : nullptr;
  p.goffset = a.goffset; p.gmask = a.gmask;
  p.nchunk = d.nchunk; p.split = d.split; p.taps_per_item = NUM_TAPS / d.split; p.num_items = p.g.num_tiles * d.split;
  p.k_atoms = d.k_atoms; p.a_bufs = d.a_bufs; p.w_stage_bytes = d.w_stage_bytes; p.w_stages = d.w_stages; p.param_bytes = d.param_bytes;
  if (p.num_items == 0) return CPOLY_OK;
  int dev = 0, sms = 0;
  CPOLY_CUDA_OK(cudaGetDevice(&dev));
  CPOLY_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.num_items < sms ? p.num_items : sms;
  if (bf) {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_dgrad_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, d.smem_bytes));
    dcn_dgrad_tc_kernel<__nv_bfloat16><<<grid, NTHREADS, d.smem_bytes, stream>>>(p);
  } else {
    CPOLY_CUDA_OK(cudaFuncSetAttribute(dcn_dgrad_tc_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, d.smem_bytes));
    dcn_dgrad_tc_kernel<float><<<grid, NTHREADS, d.smem_bytes, stream>>>(p);
  }
  if (int rc = check_launch("dcn_dgrad_tc_kernel")) return rc;
  // grad_input: gather-form col2im over the stored gcol (no atomics, every element written exactly once)
  if (a.gx.ptr) return dcn_col2im_launch(a, c2i_ws, workspace_bytes - dgrad_packed_aligned(f), stream);
  return CPOLY_OK;
}

}  // namespace cpoly
