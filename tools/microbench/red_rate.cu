// Microbenchmark: no-return global reductions to pseudo-random rows of an 8 MiB-per-level gradient slab.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
template <int MODE>  // 0: 8 x RED.64 (f32x2), 1: 4 x RED.128, 2: 8 x RED.32, 3: 16 x RED.32 (same bytes as mode 0)
__global__ void __launch_bounds__(256) k(const float* __restrict__ ld_src, float* g, int64_t n_pts, int levels, uint32_t slab_rows_mask) {
  int l = blockIdx.y;
  float2* g2 = reinterpret_cast<float2*>(g) + ((size_t)l << 20);
  for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < n_pts; m += (int64_t)gridDim.x * blockDim.x) {
    uint32_t h = mix((uint32_t)m * 16u + l);
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { uint32_t r = mix(h + i) & slab_rows_mask; atomicAdd(g2 + r, make_float2(1.f, 2.f)); }
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 4; ++i) { uint32_t r = mix(h + i) & slab_rows_mask; atomicAdd(reinterpret_cast<float4*>(g2) + (r >> 1), make_float4(1.f, 2.f, 3.f, 4.f)); }
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { uint32_t r = mix(h + i) & slab_rows_mask; atomicAdd(reinterpret_cast<float*>(g2 + r), 1.f); }
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { uint32_t r = mix(h + i) & slab_rows_mask; atomicAdd(reinterpret_cast<float*>(g2 + r), 1.f); atomicAdd(reinterpret_cast<float*>(g2 + r) + 1, 1.f); }
    } else {
      // real hash structure: random cell in a 2048^3 grid
      uint32_t ix = h & 2047u, iy = (h >> 11) & 2047u, iz = mix(h) & 2047u;
      float2 gv = make_float2(1.f, 2.f);
      if (MODE == 6) gv = make_float2(__ldg(ld_src + (size_t)(2 * l) * n_pts + m), __ldg(ld_src + (size_t)(2 * l + 1) * n_pts + m));
      uint32_t hy0 = iy * 2654435761u, hy1 = (iy + 1u) * 2654435761u, hz0 = iz * 805459861u, hz1 = (iz + 1u) * 805459861u;
      uint32_t rows[8] = {(ix ^ hy0 ^ hz0), (ix ^ hy0 ^ hz1), (ix ^ hy1 ^ hz0), (ix ^ hy1 ^ hz1), ((ix + 1) ^ hy0 ^ hz0), ((ix + 1) ^ hy0 ^ hz1), ((ix + 1) ^ hy1 ^ hz0), ((ix + 1) ^ hy1 ^ hz1)};
      if (MODE >= 5 && (ix & 1u) == 0u) {
#pragma unroll
        for (int i = 0; i < 4; ++i) atomicAdd(reinterpret_cast<float4*>(g2) + ((rows[i] & slab_rows_mask) >> 1), make_float4(gv.x, gv.y, 3.f, 4.f));
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) atomicAdd(g2 + (rows[i] & slab_rows_mask), gv);
      }
    }
  }
}
int main() {
  const int L = 16; const int64_t P = 786432;
  float* g; cudaMalloc(&g, (size_t)L << 23);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float* src; cudaMalloc(&src, (size_t)P * 32 * 4); cudaMemset(src, 0, (size_t)P * 32 * 4);
  for (int mode = 0; mode < 7; ++mode) for (int ctas_per_sm = 4; ctas_per_sm <= 8; ctas_per_sm *= 2) for (int lv = 16; lv >= 16; lv /= 4) {
    dim3 grid(148 * ctas_per_sm, lv);
    float best = 1e9;
    for (int it = 0; it < 5; ++it) {
      cudaMemsetAsync(g, 0, (size_t)L << 23);
      cudaEventRecord(a);
      if (mode == 0) k<0><<<grid, 256>>>(src, g, P * (16 / lv), lv, 0xFFFFF);
      if (mode == 1) k<1><<<grid, 256>>>(src, g, P * (16 / lv), lv, 0xFFFFF);
      if (mode == 2) k<2><<<grid, 256>>>(src, g, P * (16 / lv), lv, 0xFFFFF);
      if (mode == 3) k<3><<<grid, 256>>>(src, g, P * (16 / lv), lv, 0xFFFFF);
      if (mode == 4) k<4><<<grid, 256>>>(src, g, P, lv, 0xFFFFF);
      if (mode == 5) k<5><<<grid, 256>>>(src, g, P, lv, 0xFFFFF);
      if (mode == 6) k<6><<<grid, 256>>>(src, g, P, lv, 0xFFFFF);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    double ops = (double)P * 16 * (mode == 1 ? 4 : (mode == 3 ? 16 : (mode >= 5 ? 6 : 8)));
    printf("mode %d ctas/sm %d levels-concurrent %2d : %.1f us  %.1f Gops/s  %.1f GB/s payload\n", mode, ctas_per_sm, lv, best * 1e3, ops / best / 1e6,
           (double)P * 16 * 64 / best / 1e6 * (mode == 2 ? 0.5 : 1.0));
  }
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
