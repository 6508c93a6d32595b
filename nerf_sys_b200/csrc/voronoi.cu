// (f)-4: the Voronoi mask builder's per-ray work (scripts/create_clusters.py:386-556 compute_voronoi_opt, :559-634
// compute_voronoi_orig): S samples per pixel ray between near and far, every sample assigned to the centroids within the
// distance-ratio band (the a6 routing rule in the squared domain: d2 <= margin^2 * min d2; margin == 1: argmin), a ray is
// a member of expert c if any of its samples is; optionally the per-expert bounding boxes and counts of the assigned sample
// positions (the streams the script all-reduces with MIN / MAX / SUM, :928-932).  One thread per ray; the boxes go through a
// warp reduction and then ordered-integer atomics.  Compiled -fmad=false (sample positions as torch computes them).
#include <math.h>

#include "common.cuh"

namespace acn {

struct VorCfg {
  float c[ACN_MAX_EXPERTS][3];
  int C, dims;  // dims = 2: components (1,2); 3: (0,1,2)
  float m2;     // margin^2; strict Voronoi when margin == 1
  int strict;
};

// order-preserving map float -> uint (for atomicMin / atomicMax on floats)
__device__ __forceinline__ uint32_t f2o(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void __launch_bounds__(128) k_voronoi(const float* __restrict__ rays, int64_t N, int S, VorCfg cfg, uint8_t* __restrict__ mask,
                                                 uint32_t* __restrict__ mins_o, uint32_t* __restrict__ maxs_o, unsigned long long* __restrict__ counts) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const bool live = r < N;
  float o[3] = {0, 0, 0}, d[3] = {0, 0, 0}, near = 0.f, far = 0.f;
  if (live) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(rays + r * 8)), b = __ldg(reinterpret_cast<const float4*>(rays + r * 8) + 1);
    o[0] = a.x, o[1] = a.y, o[2] = a.z, d[0] = a.w, d[1] = b.x, d[2] = b.y, near = b.z, far = b.w;
  }
  const bool boxes = mins_o != nullptr;
  uint32_t ray_mask = 0;
  float mn[ACN_MAX_EXPERTS][3], mx[ACN_MAX_EXPERTS][3];
  uint32_t cnt[ACN_MAX_EXPERTS];
  if (boxes) {
#pragma unroll
    for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
      cnt[k] = 0;
#pragma unroll
      for (int a = 0; a < 3; ++a) mn[k][a] = INFINITY, mx[k][a] = -INFINITY;
    }
  }
  const float inv = S > 1 ? 1.0f / (float)(S - 1) : 0.f;
  for (int s = 0; live && s < S; ++s) {
    // torch.linspace(0, 1, S)[s]: start + s*step in the first half, end - (S-1-s)*step in the second
    const float step = inv;
    const float z = (s < S / 2) ? (float)s * step : 1.0f - (float)(S - 1 - s) * step;
    // torch.lerp(near, far, z)
    const float t = (z < 0.5f) ? near + z * (far - near) : far - (far - near) * (1.0f - z);
    const float x[3] = {o[0] + d[0] * t, o[1] + d[1] * t, o[2] + d[2] * t};
    float d2[ACN_MAX_EXPERTS];
    float dmin = INFINITY;
    int amin = 0;
#pragma unroll
    for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
      if (k < cfg.C) {
        const float dy = x[1] - cfg.c[k][1], dz = x[2] - cfg.c[k][2];
        float v = dy * dy + dz * dz;
        if (cfg.dims == 3) {
          const float dx = x[0] - cfg.c[k][0];
          v = dx * dx + v;
        }
        d2[k] = v;
        if (v < dmin) {
          dmin = v;
          amin = k;
        }
      } else {
        d2[k] = INFINITY;
      }
    }
    uint32_t m = 0;
    if (cfg.strict) {
      m = 1u << amin;
    } else {
      const float thr = cfg.m2 * dmin;
#pragma unroll
      for (int k = 0; k < ACN_MAX_EXPERTS; ++k)
        if (k < cfg.C && d2[k] <= thr) m |= 1u << k;
    }
    ray_mask |= m;
    if (boxes) {
#pragma unroll
      for (int k = 0; k < ACN_MAX_EXPERTS; ++k)
        if (m & (1u << k)) {
          ++cnt[k];
#pragma unroll
          for (int a = 0; a < 3; ++a) mn[k][a] = fminf(mn[k][a], x[a]), mx[k][a] = fmaxf(mx[k][a], x[a]);
        }
    }
  }
  if (live)
    for (int k = 0; k < cfg.C; ++k) mask[r * cfg.C + k] = (ray_mask >> k) & 1u;
  if (!boxes) return;
  // warp reduction per expert that any lane touched, then one set of atomics per warp
  const uint32_t lane = threadIdx.x & 31u;
#pragma unroll
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
    if (k >= cfg.C) break;
    if (!__any_sync(0xffffffffu, cnt[k] > 0)) continue;
    uint32_t c = cnt[k];
    float lo[3] = {mn[k][0], mn[k][1], mn[k][2]}, hi[3] = {mx[k][0], mx[k][1], mx[k][2]};
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      c += __shfl_xor_sync(0xffffffffu, c, off);
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], off));
        hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], off));
      }
    }
    if (lane == 0) {
      atomicAdd(counts + k, (unsigned long long)c);
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        atomicMin(mins_o + k * 3 + a, f2o(lo[a]));
        atomicMax(maxs_o + k * 3 + a, f2o(hi[a]));
      }
    }
  }
}

// ordered-integer boxes <-> floats (elementwise, tiny)
__global__ void k_vor_boxes(uint32_t* __restrict__ mins_o, uint32_t* __restrict__ maxs_o, float* __restrict__ mins, float* __restrict__ maxs, int n, int to_ordered) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (to_ordered) {
    mins_o[i] = f2o(mins[i]);
    maxs_o[i] = f2o(maxs[i]);
  } else {
    uint32_t a = mins_o[i], b = maxs_o[i];
    mins[i] = __uint_as_float((a & 0x80000000u) ? (a & 0x7fffffffu) : ~a);
    maxs[i] = __uint_as_float((b & 0x80000000u) ? (b & 0x7fffffffu) : ~b);
  }
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_voronoi_masks(const float* rays, int64_t N, int ray_samples, const float* centroids_host, int C, int cluster_2d, float margin, uint8_t* mask,
                      float* mins_io, float* maxs_io, int64_t* counts_io, void* ws, void* stream) {
  ACN_REQUIRE(rays && centroids_host && mask && N >= 0 && ray_samples >= 1, "acn_voronoi_masks: bad arguments");
  ACN_REQUIRE(C >= 1 && C <= ACN_MAX_EXPERTS && margin >= 1.0f, "acn_voronoi_masks: 1..%d centroids, margin >= 1", ACN_MAX_EXPERTS);
  const bool boxes = mins_io != nullptr;
  ACN_REQUIRE(!boxes || (maxs_io && counts_io && ws), "acn_voronoi_masks: boxes need mins, maxs, counts and the workspace");
  if (N == 0) return ACN_OK;
  VorCfg cfg;
  cfg.C = C;
  cfg.dims = cluster_2d ? 2 : 3;
  cfg.m2 = margin * margin;
  cfg.strict = margin == 1.0f;
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k)
    for (int a = 0; a < 3; ++a) cfg.c[k][a] = k < C ? centroids_host[k * 3 + a] : 0.f;
  cudaStream_t st = as_stream(stream);
  uint32_t* mo = boxes ? reinterpret_cast<uint32_t*>(ws) : nullptr;
  uint32_t* xo = boxes ? mo + 3 * ACN_MAX_EXPERTS : nullptr;
  if (boxes) {
    k_vor_boxes<<<1, 64, 0, st>>>(mo, xo, mins_io, maxs_io, 3 * C, 1);
    ACN_LAUNCH_CHECK();
  }
  static_assert(sizeof(unsigned long long) == sizeof(int64_t), "count width");
  k_voronoi<<<(unsigned)((N + 127) / 128), 128, 0, st>>>(rays, N, ray_samples, cfg, mask, mo, xo, reinterpret_cast<unsigned long long*>(counts_io));
  ACN_LAUNCH_CHECK();
  if (boxes) {
    k_vor_boxes<<<1, 64, 0, st>>>(mo, xo, mins_io, maxs_io, 3 * C, 0);
    ACN_LAUNCH_CHECK();
  }
  return ACN_OK;
}

int64_t acn_voronoi_ws_bytes(void) { return (int64_t)sizeof(uint32_t) * 6 * ACN_MAX_EXPERTS; }

}  // extern "C"
