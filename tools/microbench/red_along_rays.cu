// Scatter along rays: 8192 random rays x 96 samples through the unit cube, one level of resolution `res`, 16 slabs (grid.y) per launch.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
__device__ __forceinline__ float u01(uint32_t h) { return (h >> 8) * (1.f / 16777216.f); }
template <int MODE>  // 0: split RED.128/RED.64 ; 1: all RED.64 ; 2: compute only rows, RED to (row) but rows randomised per lane (no sharing)
__global__ void __launch_bounds__(256) k(float* g, int64_t n_pts, int res, int S) {
  int l = blockIdx.y;
  float2* g2 = reinterpret_cast<float2*>(g) + ((size_t)l << 20);
  const uint32_t mask = 0xFFFFF;
  for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < n_pts; m += (int64_t)gridDim.x * blockDim.x) {
    uint32_t r = (uint32_t)(m / S), s = (uint32_t)(m % S);
    uint32_t h = mix(r * 7919u + l * 104729u + 1u);
    float ox = u01(h), oy = u01(mix(h + 1)), oz = u01(mix(h + 2));
    float dx = u01(mix(h + 3)) - 0.5f, dy = u01(mix(h + 4)) - 0.5f, dz = u01(mix(h + 5)) - 0.5f;
    float inv = rsqrtf(dx * dx + dy * dy + dz * dz + 1e-9f);
    float t = ((float)s / S - 0.5f) * 1.2f;
    float x = fminf(fmaxf(ox + dx * inv * t, 1e-6f), 1.f - 1e-6f), y = fminf(fmaxf(oy + dy * inv * t, 1e-6f), 1.f - 1e-6f), z = fminf(fmaxf(oz + dz * inv * t, 1e-6f), 1.f - 1e-6f);
    float sx = x * res, sy = y * res, sz = z * res;
    uint32_t ix = (uint32_t)sx, iy = (uint32_t)sy, iz = (uint32_t)sz;
    float wx = sx - ix, wy = sy - iy, wz = sz - iz;
    if (MODE == 2) { uint32_t q = mix((uint32_t)m + 12345u * l); ix = q & 2047u; iy = (q >> 11) & 2047u; iz = mix(q) & 2047u; }
    uint32_t hy0 = iy * 2654435761u, hy1 = (iy + 1u) * 2654435761u, hz0 = iz * 805459861u, hz1 = (iz + 1u) * 805459861u;
    uint32_t rows[8] = {(ix ^ hy0 ^ hz0), (ix ^ hy0 ^ hz1), (ix ^ hy1 ^ hz0), (ix ^ hy1 ^ hz1), ((ix + 1) ^ hy0 ^ hz0), ((ix + 1) ^ hy0 ^ hz1), ((ix + 1) ^ hy1 ^ hz0), ((ix + 1) ^ hy1 ^ hz1)};
    float2 gv = make_float2(wx, wy + wz);
    if (MODE != 1 && (ix & 1u) == 0u) {
#pragma unroll
      for (int i = 0; i < 4; ++i) atomicAdd(reinterpret_cast<float4*>(g2) + ((rows[i] & mask) >> 1), make_float4(gv.x, gv.y, 3.f, 4.f));
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) atomicAdd(g2 + (rows[i] & mask), gv);
    }
  }
}
int main() {
  const int L = 16; const int64_t P = 786432;
  float* g; cudaMalloc(&g, (size_t)L << 23);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  int ress[] = {16, 23, 33, 48, 70, 101, 147, 212, 307, 445, 645, 933, 1351, 1955, 2830, 4095};
  for (int mode = 0; mode < 3; ++mode) for (int ri = 0; ri < 16; ++ri) {
    dim3 grid(148 * 4, 16);
    float best = 1e9;
    for (int it = 0; it < 4; ++it) {
      cudaMemsetAsync(g, 0, (size_t)L << 23);
      cudaEventRecord(a);
      if (mode == 0) k<0><<<grid, 256>>>(g, P, ress[ri], 96);
      if (mode == 1) k<1><<<grid, 256>>>(g, P, ress[ri], 96);
      if (mode == 2) k<2><<<grid, 256>>>(g, P, ress[ri], 96);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    printf("mode %d res %4d : %.1f us per 16 slabs (%.1f us/level)\n", mode, ress[ri], best * 1e3, best * 1e3 / 16);
  }
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
