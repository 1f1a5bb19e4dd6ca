// Training-mode BatchNorm2d (+ optional fused ReLU) for dense NHWC activations, forward and backward, sm_100a.
//
// Replaces `nn.BatchNorm2d(planes, momentum=BN_MOMENTUM)` in training mode at every site of the DLA-34 / DLAUp / IDAUp
// network: src/lib/models/networks/pose_dla_dcn.py:31-47 (BasicBlock bn1 / bn2), :143-160 (Root), :231-236 (base_layer),
// :247-252 (conv levels) and the `actf = Sequential(BatchNorm2d, ReLU)` of DeformConv (:345-359), whose ReLU is fused
// here.  In the reference these run as PyTorch's native channels_last kernels: 55 launches each of
// batch_norm_collect_statistics / transform_input / backward_reduce / backward_elemt = 17.5 ms of a 68 ms training step
// on one B200 (tools/train_trace.py), i.e. ~6 % of the HBM roof - the 16..64-channel maps of the backbone leave most
// of their threads idle.  These are pure streaming kernels: every thread owns one 16-byte channel group of a pixel row,
// loads are 16-byte and coalesced for ANY channel count (rows are contiguous in NHWC), several rows are in flight per
// thread, and the grid is a multiple of the SM count.
//
//   forward   bn_stats_kernel     per-CTA partial sum / sum of squares per channel (fixed order: deterministic)
//             bn_finalize_kernel  partials -> mean, invstd (biased variance, fp64), scale = gamma * invstd,
//                                 shift = beta - mean * scale; running_mean / running_var (unbiased) updated in place
//             bn_apply_kernel     y = x * scale + shift (ReLU optional)
//   backward  bn_bwd_reduce_kernel   g = gy * [y > 0]; per-CTA partial sum(g), sum(g * xhat)
//             bn_bwd_finalize_kernel gweight = sum(g * xhat), gbias = sum(g), coefficients of the input gradient
//             bn_bwd_apply_kernel    gx = scale * (g - mean(g) - xhat * mean(g * xhat))
//
// Algorithmic bytes: forward 3 passes (read, read, write), backward 5 (2 reads, 2 reads + 1 write) of N * C * sizeof(T).
#include "common.cuh"

namespace cpoly {
namespace {

constexpr int BN_THREADS = 256;
constexpr int BN_MAX_CTAS = 148 * 4;

template <typename T> struct Vec16;
template <> struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float (&f)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { f[2 * i] = __uint_as_float(w[i] << 16); f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
  }
  __device__ static __forceinline__ uint4 pack(const float (&f)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 p = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&p);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <> struct Vec16<float> {
  static constexpr int N = 4;
  __device__ static __forceinline__ void unpack(const uint4& v, float (&f)[4]) {
    f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
  }
  __device__ static __forceinline__ uint4 pack(const float (&f)[4]) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
  }
};

// Rows [r0, r1) of this CTA, thread -> (row within the block step, channel group)
struct RowSplit {
  int64_t r0, r1;
  int tpr, rpb, row, cg;
};
__device__ __forceinline__ RowSplit row_split(int64_t n_rows, int tpr) {
  RowSplit s;
  s.tpr = tpr;
  s.rpb = BN_THREADS / tpr;
  s.row = threadIdx.x / tpr;
  s.cg = threadIdx.x - s.row * tpr;
  const int64_t steps = (n_rows + s.rpb - 1) / s.rpb;
  const int64_t per = (steps + gridDim.x - 1) / gridDim.x;
  s.r0 = (int64_t)blockIdx.x * per * s.rpb;
  s.r1 = s.r0 + per * s.rpb;
  if (s.r1 > n_rows) s.r1 = n_rows;
  return s;
}

// Sum over the rows of the block for every channel group (tree over the row index, fixed order), then the first `tpr`
// threads write 2 * V partial sums of their channels: partial[cta][which][C].
template <int V>
__device__ __forceinline__ void block_reduce_write(float (&a)[V], float (&b)[V], int tpr, int cg, float* partial, int C) {
  __shared__ float red[BN_THREADS][2 * V + 1];
#pragma unroll
  for (int i = 0; i < V; ++i) { red[threadIdx.x][i] = a[i]; red[threadIdx.x][V + i] = b[i]; }
  __syncthreads();
  for (int s = BN_THREADS / 2; s >= tpr; s >>= 1) {
    if ((int)threadIdx.x < s) {
#pragma unroll
      for (int i = 0; i < 2 * V; ++i) red[threadIdx.x][i] += red[threadIdx.x + s][i];
    }
    __syncthreads();
  }
  if ((int)threadIdx.x < tpr) {
    float* p0 = partial + (size_t)blockIdx.x * 2 * C + cg * V;
#pragma unroll
    for (int i = 0; i < V; ++i) { p0[i] = red[threadIdx.x][i]; p0[C + i] = red[threadIdx.x][V + i]; }
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_stats_kernel(const uint4* __restrict__ x, float* __restrict__ partial, int64_t n_rows, int C) {
  constexpr int V = Vec16<T>::N;
  const RowSplit s = row_split(n_rows, C / V);
  float sum[V], sq[V];
#pragma unroll
  for (int i = 0; i < V; ++i) sum[i] = sq[i] = 0.f;
  const uint4* px = x + s.cg;
  int64_t r = s.r0 + s.row;
  // four rows in flight per thread
  for (; r + 3 * s.rpb < s.r1; r += 4 * s.rpb) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = __ldcs(px + (r + (int64_t)u * s.rpb) * s.tpr);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float f[V];
      Vec16<T>::unpack(v[u], f);
#pragma unroll
      for (int i = 0; i < V; ++i) { sum[i] += f[i]; sq[i] = fmaf(f[i], f[i], sq[i]); }
    }
  }
  for (; r < s.r1; r += s.rpb) {
    float f[V];
    Vec16<T>::unpack(__ldcs(px + r * s.tpr), f);
#pragma unroll
    for (int i = 0; i < V; ++i) { sum[i] += f[i]; sq[i] = fmaf(f[i], f[i], sq[i]); }
  }
  block_reduce_write<V>(sum, sq, s.tpr, s.cg, partial, C);
}

// One warp per channel: lanes stride over the CTAs' partial sums, then a fixed shuffle tree (bit-reproducible).
__device__ __forceinline__ void warp_sum_partials(const float* __restrict__ partial, int n_ctas, int C, int c, double& s, double& q) {
  const int lane = threadIdx.x & 31;
  s = 0.0; q = 0.0;
  for (int i = lane; i < n_ctas; i += 32) { s += (double)partial[(size_t)i * 2 * C + c]; q += (double)partial[(size_t)i * 2 * C + C + c]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); q += __shfl_xor_sync(0xffffffffu, q, o); }
}

// save: [4][C] = mean, invstd, scale, shift
__global__ void bn_finalize_kernel(const float* __restrict__ partial, int n_ctas, int64_t n_rows, int C, const float* __restrict__ weight,
                                   const float* __restrict__ bias, float* running_mean, float* running_var, float momentum, float eps,
                                   float* __restrict__ save) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (c >= C) return;
  double s, q;
  warp_sum_partials(partial, n_ctas, C, c, s, q);
  if ((threadIdx.x & 31) != 0) return;
  const double n = (double)n_rows;
  const double mean = s / n;
  double var = q / n - mean * mean;
  if (var < 0.0) var = 0.0;
  const float invstd = (float)(1.0 / sqrt(var + (double)eps));
  const float g = weight ? weight[c] : 1.f, b = bias ? bias[c] : 0.f;
  const float scale = g * invstd;
  save[c] = (float)mean;
  save[C + c] = invstd;
  save[2 * C + c] = scale;
  save[3 * C + c] = b - (float)mean * scale;
  if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
  if (running_var) running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)(n > 1.0 ? var * n / (n - 1.0) : var);
}

// y = x * scale + shift (ReLU optional).  The grid stride is a multiple of the threads per row, so a thread keeps its
// channel group: scale / shift live in registers.
template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_apply_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, const float* __restrict__ save,
                                                              int64_t n_vec, int C, int relu) {
  constexpr int V = Vec16<T>::N;
  const int tpr = C / V;
  const int cg = threadIdx.x % tpr;
  float sc[V], sh[V];
#pragma unroll
  for (int i = 0; i < V; ++i) { sc[i] = save[2 * C + cg * V + i]; sh[i] = save[3 * C + cg * V + i]; }
  const int64_t stride = (int64_t)gridDim.x * BN_THREADS;
  int64_t i0 = (int64_t)blockIdx.x * BN_THREADS + threadIdx.x;
  for (; i0 + 3 * stride < n_vec; i0 += 4 * stride) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = __ldcs(x + i0 + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float f[V];
      Vec16<T>::unpack(v[u], f);
#pragma unroll
      for (int i = 0; i < V; ++i) { f[i] = fmaf(f[i], sc[i], sh[i]); if (relu) f[i] = fmaxf(f[i], 0.f); }
      y[i0 + u * stride] = Vec16<T>::pack(f);
    }
  }
  for (; i0 < n_vec; i0 += stride) {
    float f[V];
    Vec16<T>::unpack(__ldcs(x + i0), f);
#pragma unroll
    for (int i = 0; i < V; ++i) { f[i] = fmaf(f[i], sc[i], sh[i]); if (relu) f[i] = fmaxf(f[i], 0.f); }
    y[i0] = Vec16<T>::pack(f);
  }
}

// g = gy * [x * scale + shift > 0] (the forward's own arithmetic, so the mask is the forward's); partial sum(g), sum(g * xhat)
template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_reduce_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gy, const float* __restrict__ save,
                                                                   float* __restrict__ partial, int64_t n_rows, int C, int relu) {
  constexpr int V = Vec16<T>::N;
  const RowSplit s = row_split(n_rows, C / V);
  float mean[V], istd[V], sc[V], sh[V], sg[V], sgx[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = s.cg * V + i;
    mean[i] = save[c]; istd[i] = save[C + c]; sc[i] = save[2 * C + c]; sh[i] = save[3 * C + c];
    sg[i] = sgx[i] = 0.f;
  }
  const uint4* px = x + s.cg;
  const uint4* pg = gy + s.cg;
  int64_t r = s.r0 + s.row;
  for (; r + s.rpb < s.r1; r += 2 * s.rpb) {
    uint4 vx[2], vg[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) { vx[u] = __ldcs(px + (r + (int64_t)u * s.rpb) * s.tpr); vg[u] = __ldcs(pg + (r + (int64_t)u * s.rpb) * s.tpr); }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float fx[V], fg[V];
      Vec16<T>::unpack(vx[u], fx);
      Vec16<T>::unpack(vg[u], fg);
#pragma unroll
      for (int i = 0; i < V; ++i) {
        const float g = (relu && !(fmaf(fx[i], sc[i], sh[i]) > 0.f)) ? 0.f : fg[i];
        sg[i] += g;
        sgx[i] = fmaf(g, (fx[i] - mean[i]) * istd[i], sgx[i]);
      }
    }
  }
  for (; r < s.r1; r += s.rpb) {
    float fx[V], fg[V];
    Vec16<T>::unpack(__ldcs(px + r * s.tpr), fx);
    Vec16<T>::unpack(__ldcs(pg + r * s.tpr), fg);
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float g = (relu && !(fmaf(fx[i], sc[i], sh[i]) > 0.f)) ? 0.f : fg[i];
      sg[i] += g;
      sgx[i] = fmaf(g, (fx[i] - mean[i]) * istd[i], sgx[i]);
    }
  }
  block_reduce_write<V>(sg, sgx, s.tpr, s.cg, partial, C);
}

// coef: [2][C] = mean(g), mean(g * xhat); gweight = sum(g * xhat), gbias = sum(g)
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partial, int n_ctas, int64_t n_rows, int C, float* __restrict__ gweight,
                                       float* __restrict__ gbias, float* __restrict__ coef) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (c >= C) return;
  double s, q;
  warp_sum_partials(partial, n_ctas, C, c, s, q);
  if ((threadIdx.x & 31) != 0) return;
  if (gbias) gbias[c] = (float)s;
  if (gweight) gweight[c] = (float)q;
  coef[c] = (float)(s / (double)n_rows);
  coef[C + c] = (float)(q / (double)n_rows);
}

// gx = scale * (g - mean(g) - xhat * mean(g * xhat))
template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_apply_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gy, uint4* __restrict__ gx,
                                                                  const float* __restrict__ save, const float* __restrict__ coef, int64_t n_vec, int C,
                                                                  int relu) {
  constexpr int V = Vec16<T>::N;
  const int tpr = C / V;
  const int cg = threadIdx.x % tpr;
  float mean[V], istd[V], sc[V], sh[V], mg[V], mgx[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = cg * V + i;
    mean[i] = save[c]; istd[i] = save[C + c]; sc[i] = save[2 * C + c]; sh[i] = save[3 * C + c];
    mg[i] = coef[c]; mgx[i] = coef[C + c];
  }
  const int64_t stride = (int64_t)gridDim.x * BN_THREADS;
  int64_t i0 = (int64_t)blockIdx.x * BN_THREADS + threadIdx.x;
  for (; i0 + stride < n_vec; i0 += 2 * stride) {
    uint4 vx[2], vg[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) { vx[u] = __ldcs(x + i0 + u * stride); vg[u] = __ldcs(gy + i0 + u * stride); }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float fx[V], fg[V];
      Vec16<T>::unpack(vx[u], fx);
      Vec16<T>::unpack(vg[u], fg);
#pragma unroll
      for (int i = 0; i < V; ++i) {
        const float g = (relu && !(fmaf(fx[i], sc[i], sh[i]) > 0.f)) ? 0.f : fg[i];
        fg[i] = sc[i] * (g - mg[i] - (fx[i] - mean[i]) * istd[i] * mgx[i]);
      }
      gx[i0 + u * stride] = Vec16<T>::pack(fg);
    }
  }
  for (; i0 < n_vec; i0 += stride) {
    float fx[V], fg[V];
    Vec16<T>::unpack(__ldcs(x + i0), fx);
    Vec16<T>::unpack(__ldcs(gy + i0), fg);
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float g = (relu && !(fmaf(fx[i], sc[i], sh[i]) > 0.f)) ? 0.f : fg[i];
      fg[i] = sc[i] * (g - mg[i] - (fx[i] - mean[i]) * istd[i] * mgx[i]);
    }
    gx[i0] = Vec16<T>::pack(fg);
  }
}

int tpr_of(int C, int dtype) {   // threads per row, or 0 when the channel count has no fast path
  const int V = dtype == CPOLY_BF16 ? 8 : 4;
  if (C <= 0 || C % V != 0) return 0;
  const int t = C / V;
  if (t > BN_THREADS || (t & (t - 1)) != 0) return 0;
  return t;
}

int grid_rows(int64_t n_rows, int tpr) {   // CTAs of the two reduction kernels: >= 8 block steps each, at most 4 per SM
  const int rpb = BN_THREADS / tpr;
  const int64_t steps = (n_rows + rpb - 1) / rpb;
  int64_t g = (steps + 7) / 8;
  if (g > BN_MAX_CTAS) g = BN_MAX_CTAS;
  if (g < 1) g = 1;
  return (int)g;
}

int grid_vec(int64_t n_vec) {
  int64_t g = (n_vec + (int64_t)BN_THREADS * 4 - 1) / ((int64_t)BN_THREADS * 4);
  if (g > BN_MAX_CTAS * 2) g = BN_MAX_CTAS * 2;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

bool bn_eligible(int C, int dtype) { return (dtype == CPOLY_BF16 || dtype == CPOLY_F32) && tpr_of(C, dtype) > 0; }

// partial sums of the larger of the two reductions + the backward's coefficient rows
size_t bn_workspace_bytes(int C) { return (size_t)BN_MAX_CTAS * 2 * (size_t)(C > 0 ? C : 0) * sizeof(float) + 2 * (size_t)(C > 0 ? C : 0) * sizeof(float) + 256; }

int bn_train_forward_launch(const cpoly_bn_fwd_args& a, cudaStream_t stream) {
  const int tpr = tpr_of(a.C, a.dtype);
  CPOLY_REQUIRE(tpr > 0, CPOLY_EUNSUPPORTED, "batch norm: C=%d has no fast path (needs C / %d a power of two <= 256)", a.C, a.dtype == CPOLY_BF16 ? 8 : 4);
  const int64_t n_rows = a.N, n_vec = n_rows * tpr;
  float* partial = reinterpret_cast<float*>(a.workspace);
  const int g = grid_rows(n_rows, tpr);
  if (a.dtype == CPOLY_BF16) bn_stats_kernel<__nv_bfloat16><<<g, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), partial, n_rows, a.C);
  else bn_stats_kernel<float><<<g, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), partial, n_rows, a.C);
  if (int rc = check_launch("bn_stats_kernel")) return rc;
  bn_finalize_kernel<<<(a.C + 7) / 8, 256, 0, stream>>>(partial, g, n_rows, a.C, a.weight, a.bias, a.running_mean, a.running_var, a.momentum, a.eps, a.save);
  if (int rc = check_launch("bn_finalize_kernel")) return rc;
  const int gv = grid_vec(n_vec);
  if (a.dtype == CPOLY_BF16)
    bn_apply_kernel<__nv_bfloat16><<<gv, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<uint4*>(a.y), a.save, n_vec, a.C, a.relu);
  else
    bn_apply_kernel<float><<<gv, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<uint4*>(a.y), a.save, n_vec, a.C, a.relu);
  return check_launch("bn_apply_kernel");
}

int bn_train_backward_launch(const cpoly_bn_bwd_args& a, cudaStream_t stream) {
  const int tpr = tpr_of(a.C, a.dtype);
  CPOLY_REQUIRE(tpr > 0, CPOLY_EUNSUPPORTED, "batch norm: C=%d has no fast path", a.C);
  const int64_t n_rows = a.N, n_vec = n_rows * tpr;
  float* partial = reinterpret_cast<float*>(a.workspace);
  float* coef = partial + (size_t)BN_MAX_CTAS * 2 * a.C;
  const int g = grid_rows(n_rows, tpr);
  if (a.dtype == CPOLY_BF16)
    bn_bwd_reduce_kernel<__nv_bfloat16><<<g, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<const uint4*>(a.gy), a.save, partial, n_rows, a.C, a.relu);
  else
    bn_bwd_reduce_kernel<float><<<g, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<const uint4*>(a.gy), a.save, partial, n_rows, a.C, a.relu);
  if (int rc = check_launch("bn_bwd_reduce_kernel")) return rc;
  bn_bwd_finalize_kernel<<<(a.C + 7) / 8, 256, 0, stream>>>(partial, g, n_rows, a.C, a.gweight, a.gbias, coef);
  if (int rc = check_launch("bn_bwd_finalize_kernel")) return rc;
  if (!a.gx) return CPOLY_OK;
  const int gv = grid_vec(n_vec);
  if (a.dtype == CPOLY_BF16)
    bn_bwd_apply_kernel<__nv_bfloat16><<<gv, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<const uint4*>(a.gy), reinterpret_cast<uint4*>(a.gx), a.save, coef, n_vec, a.C, a.relu);
  else
    bn_bwd_apply_kernel<float><<<gv, BN_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(a.x), reinterpret_cast<const uint4*>(a.gy), reinterpret_cast<uint4*>(a.gx), a.save, coef, n_vec, a.C, a.relu);
  return check_launch("bn_bwd_apply_kernel");
}

}  // namespace cpoly
