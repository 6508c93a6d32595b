// a6-a7: sample -> spatial-expert routing (centroid assignment) and the stable
// pack-by-expert that replaces the reference's K x (nonzero, index_select) loop
// (models/inr/meta_container.py:97-134, 305-337).
//
// Distances are the direct fp32 form sqrt(sum((x-c)^2)) (compiled -fmad=false); the
// integer decisions (argmin / membership mask) are the bit-exact contract.
// HBM-bound: 12 B read per point (36 B/ray + 4 B/point in the implicit form),
// 8 B written per (point, expert) pair.
#include "common.cuh"

namespace acn {

constexpr int RB = 1024;      // points per routing block
constexpr int RT = 256;       // threads per CTA
constexpr int RPT = RB / RT;  // points per thread

struct RouteCfg {
  float c[ACN_MAX_EXPERTS][3];
  int K;
  float margin;
  int dims;  // 2 -> axes (1,2); 3 -> axes (0,1,2)
};

// Returns the membership mask; w[k] = routing weight (0 when not a member).
// Hard routing (margin <= 1): mask has the single argmin bit, w = 1.
__device__ __forceinline__ uint32_t route_point(const RouteCfg& cfg, float x, float y, float z, float* w, int* hard) {
  float dist[ACN_MAX_EXPERTS];
  float mind = INFINITY;
  int amin = 0;
#pragma unroll
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
    if (k < cfg.K) {
      float dy = y - cfg.c[k][1], dz = z - cfg.c[k][2];
      float s;
      if (cfg.dims == 3) {
        float dx = x - cfg.c[k][0];
        s = dx * dx + dy * dy;
        s = s + dz * dz;
      } else {
        s = dy * dy + dz * dz;
      }
      float d = sqrtf(s);
      if (cfg.margin > 1.0f) d = fmaxf(d, 1e-6f);
      dist[k] = d;
      if (d < mind) {  // strict: first index wins ties, as torch.argmin / min
        mind = d;
        amin = k;
      }
    } else {
      dist[k] = INFINITY;
    }
  }
  if (hard) *hard = amin;
  uint32_t mask = 0;
  if (!(cfg.margin > 1.0f)) {
#pragma unroll
    for (int k = 0; k < ACN_MAX_EXPERTS; ++k) w[k] = (k == amin) ? 1.0f : 0.0f;
    return 1u << amin;
  }
  float thr = cfg.margin * mind;
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
    float inv = 0.f;
    if (k < cfg.K) {
      inv = 1.0f / dist[k];
      bool in = dist[k] <= thr;
      inv = in ? inv : 0.f;
      sum = sum + inv;
      // membership is "weight > 0" (meta_container.py:308)
    }
    w[k] = inv;
  }
  float den = fmaxf(sum, 1e-6f);
#pragma unroll
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
    if (k < cfg.K) {
      w[k] = w[k] / den;
      if (w[k] > 0.f) mask |= 1u << k;
    }
  }
  return mask;
}

__global__ void __launch_bounds__(RT) k_routing(PointSrc src, int64_t P, RouteCfg cfg, float* __restrict__ weights,
                                                int64_t* __restrict__ hard) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    float x, y, z, w[ACN_MAX_EXPERTS];
    int h;
    load_xyz(src, p, x, y, z);
    route_point(cfg, x, y, z, w, &h);
    if (cfg.margin > 1.0f) {
#pragma unroll
      for (int k = 0; k < ACN_MAX_EXPERTS; ++k)
        if (k < cfg.K) weights[p * cfg.K + k] = w[k];
    } else {
      hard[p] = h;
    }
  }
}

// ---- pass 1: per-block, per-expert member counts: counts[k*nb + b] ------------
__global__ void __launch_bounds__(RT) k_route_count(PointSrc src, int64_t P, RouteCfg cfg, int nb, int32_t* __restrict__ counts) {
  __shared__ int cnt[ACN_MAX_EXPERTS];
  for (int b = blockIdx.x; b < nb; b += gridDim.x) {
    if (threadIdx.x < ACN_MAX_EXPERTS) cnt[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      int64_t p = (int64_t)b * RB + j * RT + threadIdx.x;
      uint32_t mask = 0;
      if (p < P) {
        float x, y, z, w[ACN_MAX_EXPERTS];
        load_xyz(src, p, x, y, z);
        mask = route_point(cfg, x, y, z, w, nullptr);
      }
      for (int k = 0; k < cfg.K; ++k) {
        unsigned bal = __ballot_sync(0xffffffffu, (mask >> k) & 1u);
        if ((threadIdx.x & 31) == 0 && bal) atomicAdd(&cnt[k], __popc(bal));
      }
    }
    __syncthreads();
    if (threadIdx.x < cfg.K) counts[(int64_t)threadIdx.x * nb + b] = cnt[threadIdx.x];
    __syncthreads();
  }
}

// ---- pass 2: exclusive scan over blocks per expert (in place) + tile-aligned segments
__global__ void __launch_bounds__(1024) k_route_scan(int32_t* __restrict__ counts, int nb, int K, int32_t* __restrict__ seg) {
  __shared__ int warp_sum[32];
  __shared__ int carry;
  __shared__ int totals[ACN_MAX_EXPERTS];
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k = 0; k < K; ++k) {
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += 1024) {
      int i = base + threadIdx.x;
      int v = (i < nb) ? counts[(int64_t)k * nb + i] : 0;
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
      }
      if (lane == 31) warp_sum[warp] = inc;
      __syncthreads();
      if (warp == 0) {
        int ws = warp_sum[lane];
        int winc = ws;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          int n = __shfl_up_sync(0xffffffffu, winc, o);
          if (lane >= o) winc += n;
        }
        warp_sum[lane] = winc - ws;  // exclusive
      }
      __syncthreads();
      int excl = carry + warp_sum[warp] + inc - v;
      if (i < nb) counts[(int64_t)k * nb + i] = excl;
      __syncthreads();
      if (threadIdx.x == 1023) carry = excl + v;
      __syncthreads();
    }
    if (threadIdx.x == 0) totals[k] = carry;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    int s = 0;
    seg[0] = 0;
    for (int k = 0; k < K; ++k) {
      s += (totals[k] + ACN_TILE - 1) / ACN_TILE * ACN_TILE;
      seg[k + 1] = s;
    }
  }
}

// ---- pass 3: stable write of (point, weight) pairs into each expert's segment ----
__global__ void __launch_bounds__(RT) k_route_write(PointSrc src, int64_t P, RouteCfg cfg, int nb, const int32_t* __restrict__ base,
                                                    const int32_t* __restrict__ seg, int32_t* __restrict__ pair_point,
                                                    float* __restrict__ pair_w, int64_t Mcap) {
  __shared__ int run[ACN_MAX_EXPERTS];
  __shared__ int wcnt[ACN_MAX_EXPERTS][RT / 32];
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int b = blockIdx.x; b < nb; b += gridDim.x) {
    if (threadIdx.x < cfg.K) run[threadIdx.x] = seg[threadIdx.x] + base[(int64_t)threadIdx.x * nb + b];
    __syncthreads();
#pragma unroll 1
    for (int j = 0; j < RPT; ++j) {
      int64_t p = (int64_t)b * RB + j * RT + threadIdx.x;
      uint32_t mask = 0;
      float w[ACN_MAX_EXPERTS];
      if (p < P) {
        float x, y, z;
        load_xyz(src, p, x, y, z);
        mask = route_point(cfg, x, y, z, w, nullptr);
      }
      unsigned bals[ACN_MAX_EXPERTS];
#pragma unroll
      for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
        bals[k] = 0;
        if (k < cfg.K) {
          bals[k] = __ballot_sync(0xffffffffu, (mask >> k) & 1u);
          if (lane == 0) wcnt[k][warp] = __popc(bals[k]);
        }
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
        if (k < cfg.K && ((mask >> k) & 1u)) {
          int off = run[k];
          for (int ww = 0; ww < warp; ++ww) off += wcnt[k][ww];
          off += __popc(bals[k] & ((1u << lane) - 1u));
          if (off < Mcap) {
            pair_point[off] = (int32_t)p;
            pair_w[off] = w[k];
          }
        }
      }
      __syncthreads();
      if (threadIdx.x < cfg.K) {
        int s = 0;
        for (int ww = 0; ww < RT / 32; ++ww) s += wcnt[threadIdx.x][ww];
        run[threadIdx.x] += s;
      }
      __syncthreads();
    }
  }
}

static int fill_cfg(RouteCfg& cfg, const float* centroids_host, int K, float margin, int cluster_2d) {
  cfg.K = K;
  cfg.margin = margin;
  cfg.dims = cluster_2d ? 2 : 3;
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k)
    for (int a = 0; a < 3; ++a) cfg.c[k][a] = (k < K) ? centroids_host[k * 3 + a] : 0.f;
  return 0;
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_routing(const AcnPoints* pts, const float* centroids_host, int K, float margin, int cluster_2d, float* weights,
                int64_t* hard, void* stream) {
  ACN_REQUIRE(pts && centroids_host && K >= 1 && K <= ACN_MAX_EXPERTS, "acn_routing: K must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(margin > 1.0f ? weights != nullptr : hard != nullptr, "acn_routing: missing output");
  if (pts->P == 0) return ACN_OK;
  RouteCfg cfg;
  fill_cfg(cfg, centroids_host, K, margin, cluster_2d);
  k_routing<<<persistent_grid((pts->P + RT - 1) / RT, 8), RT, 0, as_stream(stream)>>>(make_src(pts), pts->P, cfg, weights, hard);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int64_t acn_route_pack_capacity(int64_t P, int K) {
  // typical (margin 1.05): ~1.03 memberships per point; callers re-run with seg[K] if exceeded
  return P + P / 4 + (int64_t)K * ACN_TILE;
}

int64_t acn_route_pack_ws_bytes(int64_t P, int K) {
  int64_t nb = (P + RB - 1) / RB;
  return sizeof(int32_t) * (nb * K + 64);
}

int acn_route_pack(const AcnPoints* pts, const float* centroids_host, int K, float margin, int cluster_2d, int32_t* pair_point,
                   float* pair_weight, int32_t* seg, int64_t Mcap, void* ws, void* stream) {
  ACN_REQUIRE(pts && centroids_host && K >= 1 && K <= ACN_MAX_EXPERTS, "acn_route_pack: K must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(pair_point && pair_weight && seg && ws && Mcap > 0, "acn_route_pack: missing buffers");
  ACN_REQUIRE(pts->P < (int64_t)1 << 31, "acn_route_pack: chunk the points below 2^31");
  cudaStream_t st = as_stream(stream);
  int64_t P = pts->P;
  int nb = (int)((P + RB - 1) / RB);
  RouteCfg cfg;
  fill_cfg(cfg, centroids_host, K, margin, cluster_2d);
  int32_t* counts = reinterpret_cast<int32_t*>(ws);
  ACN_CUDA(cudaMemsetAsync(pair_point, 0xFF, sizeof(int32_t) * Mcap, st));
  ACN_CUDA(cudaMemsetAsync(pair_weight, 0, sizeof(float) * Mcap, st));
  if (nb > 0) {
    k_route_count<<<persistent_grid(nb, 8), RT, 0, st>>>(make_src(pts), P, cfg, nb, counts);
    ACN_LAUNCH_CHECK();
  }
  k_route_scan<<<1, 1024, 0, st>>>(counts, nb, K, seg);
  ACN_LAUNCH_CHECK();
  if (nb > 0) {
    k_route_write<<<persistent_grid(nb, 8), RT, 0, st>>>(make_src(pts), P, cfg, nb, counts, seg, pair_point, pair_weight, Mcap);
    ACN_LAUNCH_CHECK();
  }
  return ACN_OK;
}

}  // extern "C"
