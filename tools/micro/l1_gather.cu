// Microbenchmark: cycles per warp instruction of L1-resident gathers (LDG.128 touching 4 lines, LDG.64 / 2, LDG.32 / 1)
// against the same pattern from shared memory (LDS.128, 4 rows).  nvcc -arch=sm_100a -O3 -o l1_gather l1_gather.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ROWS = 256;   // 256 rows x 128 B = 32 KB working set (L1 resident)

template <int MODE>
__global__ void k(const uint8_t* __restrict__ g, long long* out, int iters, uint32_t* sink) {
  extern __shared__ uint8_t sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < ROWS * 8; i += blockDim.x) reinterpret_cast<uint4*>(sm)[i] = reinterpret_cast<const uint4*>(g)[i];
  __syncthreads();
  uint32_t acc = 0;
  uint32_t r = (warp * 37 + (lane >> 3) * 11) & (ROWS - 1);
  // warm L1
  for (int i = lane; i < ROWS * 8; i += 32) acc += __ldg(reinterpret_cast<const uint4*>(g) + i).x;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      r = (r * 5 + 17 + u) & (ROWS - 1);
      if (MODE == 0) {            // LDG.128, 4 rows per warp instruction (8 lanes per row)
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(g + r * 128 + (lane & 7) * 16));
        acc += v.x ^ v.w;
      } else if (MODE == 1) {     // LDS.128, same pattern
        const uint4 v = *reinterpret_cast<const uint4*>(sm + r * 128 + (lane & 7) * 16);
        acc += v.x ^ v.w;
      } else if (MODE == 2) {     // LDG.64, 2 rows per instruction (16 lanes per row)
        const uint32_t r2 = (r & ~1u) | (lane >> 4);
        const uint2 v = __ldg(reinterpret_cast<const uint2*>(g + r2 * 128 + (lane & 15) * 8));
        acc += v.x ^ v.y;
      } else if (MODE == 3) {     // LDG.32, 1 row per instruction
        const uint32_t r1 = __shfl_sync(0xffffffffu, r, 0);
        acc += __ldg(reinterpret_cast<const uint32_t*>(g + r1 * 128 + lane * 4));
      } else if (MODE == 5) {     // LDS.64, 2 rows per instruction
        const uint32_t r2 = (r & ~1u) | (lane >> 4);
        const uint2 v = *reinterpret_cast<const uint2*>(sm + r2 * 128 + (lane & 15) * 8);
        acc += v.x ^ v.y;
      } else if (MODE == 6) {     // LDS.32 broadcast, 2 distinct words per warp
        acc += *reinterpret_cast<const uint32_t*>(sm + r * 4 + (lane >> 4) * 4);
      } else if (MODE == 7) {     // LDS.64 broadcast, 2 distinct
        const uint2 v = *reinterpret_cast<const uint2*>(sm + (r & ~1u) * 8 + (lane >> 4) * 8);
        acc += v.x ^ v.y;
      } else if (MODE == 8) {     // LDS.128 broadcast, 2 distinct
        const uint4 v = *reinterpret_cast<const uint4*>(sm + (r & ~1u) * 16 + (lane >> 4) * 16);
        acc += v.x ^ v.w;
      } else if (MODE == 9) {     // LDS.128 broadcast, 4 distinct (8 lanes each)
        const uint4 v = *reinterpret_cast<const uint4*>(sm + (r & ~3u) * 16 + (lane >> 3) * 16);
        acc += v.x ^ v.w;
      } else if (MODE == 10) {    // STS.128, 4 rows, swizzled pieces
        const uint32_t rr = (r & ~3u) | (lane >> 3);
        *reinterpret_cast<uint4*>(sm + rr * 128 + (((lane & 7) ^ (rr & 7)) << 4)) = make_uint4(acc, r, u, lane);
      } else if (MODE == 11) {    // STS.64, 2 rows
        const uint32_t rr = (r & ~1u) | (lane >> 4);
        *reinterpret_cast<uint2*>(sm + rr * 128 + (lane & 15) * 8) = make_uint2(acc, r);
      } else if (MODE == 12) {    // LDG.64, 2 rows, 4 in a row then use (deeper MLP)
        const uint32_t r2 = (r & ~1u) | (lane >> 4);
        const uint2 v = __ldg(reinterpret_cast<const uint2*>(g + r2 * 128 + (lane & 15) * 8));
        const uint2 v2 = __ldg(reinterpret_cast<const uint2*>(g + ((r2 + 64) & (ROWS - 1)) * 128 + (lane & 15) * 8));
        acc += v.x ^ v.y ^ v2.x ^ v2.y;
      } else if (MODE == 4) {     // LDG.128, 512 contiguous bytes (4 consecutive lines)
        const uint32_t r4 = __shfl_sync(0xffffffffu, r, 0) & ~3u;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(g + r4 * 128 + lane * 16));
        acc += v.x ^ v.w;
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) *sink = acc;
}

int main() {
  uint8_t* g; long long* out; uint32_t* sink;
  cudaMalloc(&g, ROWS * 128); cudaMemset(g, 1, ROWS * 128);
  cudaMalloc(&out, 148 * 8); cudaMalloc(&sink, 4);
  const int iters = 2000;
  const char* names[] = {"LDG.128 4 rows", "LDS.128 4 rows", "LDG.64 2 rows", "LDG.32 1 row", "LDG.128 contiguous 512 B", "LDS.64 2 rows", "LDS.32 bcast 2", "LDS.64 bcast 2", "LDS.128 bcast 2", "LDS.128 bcast 4", "STS.128 4 rows swz", "STS.64 2 rows", "2x LDG.64 2 rows"};
  for (int warps : {16}) {
    for (int mode = 0; mode < 13; ++mode) {
      long long h[148];
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 1) k<1><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 2) k<2><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 3) k<3><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 4) k<4><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 5) k<5><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 6) k<6><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 7) k<7><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 8) k<8><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 9) k<9><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 10) k<10><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 11) k<11><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        if (mode == 12) k<12><<<148, warps * 32, ROWS * 128>>>(g, out, iters, sink);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      const double per_instr = (double)h[0] / ((double)iters * 8 * warps);
      printf("%2d warps  %-26s %6.2f cycles per warp instruction (SM-wide)  err=%s\n", warps, names[mode], per_instr, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
