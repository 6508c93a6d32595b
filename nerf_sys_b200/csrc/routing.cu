// a6-a7: sample -> spatial-expert routing (centroid assignment) and the stable
// pack-by-expert that replaces the reference's K x (nonzero, index_select) loop
// (models/inr/meta_container.py:97-134, 305-337).
//
// Distances are the direct fp32 form sqrt(sum((x-c)^2)) (compiled -fmad=false); the
// integer decisions (argmin / membership mask) are the bit-exact contract.
// Algorithmic traffic: 12 B read per point (36 B/ray + 4 B/point in the implicit form), 8 B written per (point, expert)
// pair; the pack additionally keeps one 4 B decision word per point between its count and write passes.  Measured
// binding resource: instruction issue in the count pass (profiles/r2_route_kernels.json), not HBM.
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
// KT = experts rounded up to a power of two (loops are unrolled to KT with no per-k "k < K" test: the centroids of the
// padding experts sit at +inf, so their distance is +inf and they are never the nearest, never members and add an exact 0
// to the weight sum); D3 = distances over (x,y,z) instead of (y,z).
template <int KT, bool D3>
__device__ __forceinline__ uint32_t route_point(const RouteCfg& cfg, float x, float y, float z, float* w, int* hard) {
  // squared distances, in the operation order the distance itself uses
  float sq[KT];
  float smin = INFINITY;
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    float dy = y - cfg.c[k][1], dz = z - cfg.c[k][2];
    float s;
    if (D3) {
      float dx = x - cfg.c[k][0];
      s = dx * dx + dy * dy;
      s = s + dz * dz;
    } else {
      s = dy * dy + dz * dz;
    }
    sq[k] = s;
    smin = fminf(smin, s);
  }
  // Fast path (19 of 20 samples): exactly one centroid inside the membership band.  The band is the membership test in the
  // squared domain widened by 2e-5 (1e-5 in distance), two orders of magnitude above the rounding of sqrt and of
  // margin*min, so a centroid outside it fails the exact test d_k <= margin*d_min too; the lone member's exact weight is
  // (1/d)/max(1/d, 1e-6) = 1 for 1e-6 < d < 1e6.  Everything else (two candidates, a tie, a sample on a centroid,
  // non-finite positions of invalid rays) takes the exact arithmetic below, so the integer decisions stay bit-exact
  // (tests/test_gpu_parity.py).
  {
    const float m2 = cfg.margin > 1.0f ? cfg.margin * cfg.margin : 1.0f;
    const float band = m2 * smin * 1.00002f;
    int ncand = 0, c = 0;
#pragma unroll
    for (int k = 0; k < KT; ++k) {
      const bool in = sq[k] <= band;
      ncand += in ? 1 : 0;
      c = in ? k : c;
    }
    if (ncand == 1 && smin > 1e-11f && smin < 1e11f) {
#pragma unroll
      for (int k = 0; k < KT; ++k) w[k] = (k == c) ? 1.0f : 0.0f;
      if (hard) *hard = c;
      return 1u << c;
    }
  }
  // exact path, branch-free over the KT centroids
  float dist[KT];
  float mind = INFINITY;
  int amin = 0;
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    float d = sqrtf(sq[k]);
    if (cfg.margin > 1.0f) d = (d < 1e-6f) ? 1e-6f : d;  // clamp_min that propagates NaN, as torch's does (fmaxf would return 1e-6)
    dist[k] = d;
    const bool lt = d < mind;  // strict: first index wins ties, as torch.argmin / min
    mind = lt ? d : mind;
    amin = lt ? k : amin;
  }
  if (hard) *hard = amin;
  if (!(cfg.margin > 1.0f)) {
#pragma unroll
    for (int k = 0; k < KT; ++k) w[k] = (k == amin) ? 1.0f : 0.0f;
    return 1u << amin;
  }
  const float thr = cfg.margin * mind;
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    float inv = 1.0f / dist[k];
    inv = (dist[k] <= thr) ? inv : 0.f;
    sum = sum + inv;
    w[k] = inv;  // membership is "weight > 0" (meta_container.py:308)
  }
  const float den = fmaxf(sum, 1e-6f);
  uint32_t mask = 0;
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    w[k] = w[k] / den;
    mask |= (w[k] > 0.f) ? (1u << k) : 0u;
  }
  return mask;
}

template <int KT, bool D3>
__global__ void __launch_bounds__(RT) k_routing(PointSrc src, int64_t P, RouteCfg cfg, float* __restrict__ weights,
                                                int64_t* __restrict__ hard) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    float x, y, z, w[KT];
    int h;
    load_xyz(src, p, x, y, z);
    route_point<KT, D3>(cfg, x, y, z, w, &h);
    if (cfg.margin > 1.0f) {
#pragma unroll
      for (int k = 0; k < KT; ++k)
        if (k < cfg.K) weights[p * cfg.K + k] = w[k];
    } else {
      hard[p] = h;
    }
  }
}

// ---- the same decisions for a whole warp (the pack kernels) -----------------------------------------------------
// Lanes are consecutive samples of a ray, so 4 of 5 warps hold a sample or two inside a membership band while the other
// lanes take the fast path; run per thread, the exact arithmetic then executes with ~2 of 32 lanes active (measured:
// 81 % of the warps, 2.3 active lanes, two thirds of the kernel's instructions).  Here every lane first classifies its own
// sample (squared distances, candidate count), then the warp works through its "exact" samples together: 32/KT samples per
// pass, ONE (sample, centroid) pair per lane - one sqrt, one reciprocal, one division per lane - with the weight sum
// accumulated in ascending centroid order by a shuffle chain, i.e. the same operations in the same order as route_point().
constexpr uint32_t UNIT_W = 0x80000000u;  // decision word: membership mask | UNIT_W when every member weight is exactly 1

// Returns the decision word of this lane's sample p; the weights of a sample with a member weight != 1 are written to
// pw[p*KT + k] by the lanes that computed them (one 4*KT-byte row per sample).
template <int KT, bool D3>
__device__ __forceinline__ uint32_t route_warp(const RouteCfg& cfg, bool valid, int p, float x, float y, float z, float* __restrict__ pw) {
  constexpr uint32_t FULL = 0xffffffffu;
  constexpr int PPP = 32 / KT;  // samples per pass
  const int lane = threadIdx.x & 31, g = lane / KT, k = lane % KT;
  const bool soft = cfg.margin > 1.0f;
  float sq[KT];
  float smin = INFINITY;
#pragma unroll
  for (int kk = 0; kk < KT; ++kk) {
    float dy = y - cfg.c[kk][1], dz = z - cfg.c[kk][2];
    float s;
    if (D3) {
      float dx = x - cfg.c[kk][0];
      s = dx * dx + dy * dy;
      s = s + dz * dz;
    } else {
      s = dy * dy + dz * dz;
    }
    sq[kk] = s;
    smin = fminf(smin, s);
  }
  const float band = (soft ? cfg.margin * cfg.margin : 1.0f) * smin * 1.00002f;
  int ncand = 0, c = 0;
#pragma unroll
  for (int kk = 0; kk < KT; ++kk) {
    const bool in = sq[kk] <= band;
    ncand += in ? 1 : 0;
    c = in ? kk : c;
  }
  const bool fast = ncand == 1 && smin > 1e-11f && smin < 1e11f;  // see route_point()
  uint32_t word = (valid && fast) ? ((1u << c) | UNIT_W) : UNIT_W;
  uint32_t slow = __ballot_sync(FULL, valid && !fast);
  while (slow) {  // warp-uniform
    // this pass takes the first PPP exact samples; group g of KT lanes works for the g-th of them
    uint32_t rem = slow;
    int src = -1;
#pragma unroll
    for (int i = 0; i < PPP; ++i) {
      const int l = rem ? (__ffs(rem) - 1) : -1;
      src = (i == g) ? l : src;
      rem = rem ? (rem & (rem - 1)) : 0u;
    }
    const uint32_t taken = slow & ~rem;
    slow = rem;
    const bool active = src >= 0;
    const int from = active ? src : 0;
    float v = INFINITY;
#pragma unroll
    for (int kk = 0; kk < KT; ++kk) {
      const float t = __shfl_sync(FULL, sq[kk], from);
      v = (kk == k) ? t : v;
    }
    const int p_from = __shfl_sync(FULL, p, from);
    float d = sqrtf(v);
    if (soft) d = (d < 1e-6f) ? 1e-6f : d;  // NaN stays NaN (torch.clamp_min): a non-finite sample is a member of no expert
    if (!active) d = INFINITY;
    float mind = d;  // min over the group; NaN never wins, as with the strict "d < min" chain
#pragma unroll
    for (int o = KT / 2; o > 0; o >>= 1) mind = fminf(mind, __shfl_xor_sync(FULL, mind, o));
    const uint32_t gmask = (KT == 32) ? FULL : ((1u << KT) - 1u);
    const int gshift = g * KT;
    float wv;
    bool member;
    if (!soft) {
      // first index that reaches the minimum (torch.argmin); index 0 when no distance is below +inf
      const uint32_t hit = (__ballot_sync(FULL, active && d == mind && mind < INFINITY) >> gshift) & gmask;
      const int amin = hit ? (__ffs(hit) - 1) : 0;
      member = active && k == amin;
      wv = member ? 1.0f : 0.0f;
    } else {
      const float thr = cfg.margin * mind;
      float inv = 1.0f / d;
      inv = (d <= thr) ? inv : 0.f;
      float sum = 0.f;
#pragma unroll
      for (int kk = 0; kk < KT; ++kk) sum = sum + __shfl_sync(FULL, inv, gshift + kk);  // ascending centroid order
      const float den = fmaxf(sum, 1e-6f);
      wv = inv / den;
      member = active && wv > 0.f;  // membership is "weight > 0" (meta_container.py:308)
    }
    const uint32_t bal = __ballot_sync(FULL, member);
    const uint32_t odd = __ballot_sync(FULL, member && wv != 1.0f);
    if (active && ((odd >> gshift) & gmask)) pw[(int64_t)p_from * KT + k] = wv;
    // hand the decision to the lane that owns the sample
    const bool mine = (taken >> lane) & 1u;
    const int my_shift = __popc(taken & ((1u << lane) - 1u)) * KT;
    if (mine) word = ((bal >> my_shift) & gmask) | (((odd >> my_shift) & gmask) ? 0u : UNIT_W);
  }
  return word;
}

// member counts of up to 4 experts as one byte each: bit i of a nibble -> bit 8*i
__device__ __forceinline__ uint32_t spread4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

// ---- pass 1: per-block, per-expert member counts: counts[k*nb + b].  The decisions are kept for pass 3: pmask[p] = decision
// word, pw[p*KT ..] = the weights of the samples whose word lacks UNIT_W (1 in 20).
template <int KT, bool D3>
__global__ void __launch_bounds__(RT) k_route_count(PointSrc src, int64_t P, RouteCfg cfg, int nb, int32_t* __restrict__ counts,
                                                    uint32_t* __restrict__ pmask, float* __restrict__ pw) {
  constexpr int NW = (KT + 3) / 4;
  __shared__ int cnt[ACN_MAX_EXPERTS];
  const uint32_t keep = ((1u << cfg.K) - 1u) | UNIT_W;  // never a padding expert
  for (int b = blockIdx.x; b < nb; b += gridDim.x) {
    if (threadIdx.x < ACN_MAX_EXPERTS) cnt[threadIdx.x] = 0;
    __syncthreads();
    uint32_t acc[NW];  // this thread's member count per expert, one byte each (<= RPT)
#pragma unroll
    for (int i = 0; i < NW; ++i) acc[i] = 0u;
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      const int64_t p = (int64_t)b * RB + j * RT + threadIdx.x;
      float x = 0.f, y = 0.f, z = 0.f;
      if (p < P) load_xyz(src, p, x, y, z);
      const uint32_t word = route_warp<KT, D3>(cfg, p < P, (int)p, x, y, z, pw) & keep;
      if (p < P) pmask[p] = word;
#pragma unroll
      for (int i = 0; i < NW; ++i) acc[i] += spread4((word >> (4 * i)) & 0xFu);
    }
#pragma unroll
    for (int i = 0; i < NW; ++i) {
      const uint32_t s = __reduce_add_sync(0xffffffffu, acc[i]);  // bytes stay <= 32*RPT = 128: no carry between experts
      if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int c = (s >> (8 * q)) & 0xFF;
          if (c) atomicAdd(&cnt[4 * i + q], c);
        }
      }
    }
    __syncthreads();
    if (threadIdx.x < cfg.K) counts[(int64_t)threadIdx.x * nb + b] = cnt[threadIdx.x];
    __syncthreads();
  }
}

// ---- pass 2: exclusive scan over blocks, one CTA per expert (in place); totals[k] -> tile-aligned segments
__global__ void __launch_bounds__(1024) k_route_scan(int32_t* __restrict__ counts, int nb, int32_t* __restrict__ totals) {
  __shared__ int warp_sum[32];
  __shared__ int carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, k = blockIdx.x;
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
}

__global__ void k_route_seg(const int32_t* __restrict__ totals, int K, int32_t* __restrict__ seg) {
  if (threadIdx.x == 0) {
    int s = 0;
    seg[0] = 0;
    for (int k = 0; k < K; ++k) {
      s += (totals[k] + ACN_TILE - 1) / ACN_TILE * ACN_TILE;
      seg[k + 1] = s;
    }
  }
}

// ---- pass 3: stable write of (point, weight) pairs into each expert's segment, from the decisions pass 1 kept (no distance
// arithmetic, no point fetch): per-warp member counts of the block's 4 x 8 warp rows -> one serial scan per expert -> offsets
template <int KT>
__global__ void __launch_bounds__(RT) k_route_write(int64_t P, int K, int nb, const int32_t* __restrict__ base, const int32_t* __restrict__ seg,
                                                    const uint32_t* __restrict__ pmask, const float* __restrict__ pw,
                                                    int32_t* __restrict__ pair_point, float* __restrict__ pair_w, int64_t Mcap) {
  constexpr int ROWS = RPT * (RT / 32);  // warp rows of a block, in point order
  __shared__ int wcnt[KT][ROWS + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int b = blockIdx.x; b < nb; b += gridDim.x) {
    uint32_t m[RPT];
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      const int64_t p = (int64_t)b * RB + j * RT + threadIdx.x;
      m[j] = (p < P) ? pmask[p] : 0u;
    }
#pragma unroll
    for (int j = 0; j < RPT; ++j)
#pragma unroll
      for (int k = 0; k < KT; ++k) {
        const unsigned bal = __ballot_sync(0xffffffffu, (m[j] >> k) & 1u);
        if (lane == 0) wcnt[k][j * (RT / 32) + warp] = __popc(bal);
      }
    __syncthreads();
    if (threadIdx.x < K) {
      int run = seg[threadIdx.x] + base[(int64_t)threadIdx.x * nb + b];
      for (int i = 0; i < ROWS; ++i) {
        const int c = wcnt[threadIdx.x][i];
        wcnt[threadIdx.x][i] = run;
        run += c;
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      const int64_t p = (int64_t)b * RB + j * RT + threadIdx.x;
      const bool unit = (m[j] & UNIT_W) != 0u;
#pragma unroll
      for (int k = 0; k < KT; ++k) {
        const unsigned bal = __ballot_sync(0xffffffffu, (m[j] >> k) & 1u);
        if ((m[j] >> k) & 1u) {
          const int off = wcnt[k][j * (RT / 32) + warp] + __popc(bal & ((1u << lane) - 1u));
          if (off < Mcap) {
            pair_point[off] = (int32_t)p;
            pair_w[off] = unit ? 1.0f : pw[p * KT + k];
          }
        }
      }
    }
    __syncthreads();
  }
}

// kernel<KT, D3> with KT = K rounded up to 2, 4, 8 or 16
#define ROUTE_LAUNCH_D(KT, c2d, kernel, grid, st, ...)                   \
  do {                                                                   \
    if (c2d)                                                             \
      kernel<KT, false><<<grid, RT, 0, st>>>(__VA_ARGS__);               \
    else                                                                 \
      kernel<KT, true><<<grid, RT, 0, st>>>(__VA_ARGS__);                \
  } while (0)
#define ROUTE_DISPATCH(K, c2d, kernel, grid, st, ...)                    \
  do {                                                                   \
    if ((K) <= 2)                                                        \
      ROUTE_LAUNCH_D(2, c2d, kernel, grid, st, __VA_ARGS__);             \
    else if ((K) <= 4)                                                   \
      ROUTE_LAUNCH_D(4, c2d, kernel, grid, st, __VA_ARGS__);             \
    else if ((K) <= 8)                                                   \
      ROUTE_LAUNCH_D(8, c2d, kernel, grid, st, __VA_ARGS__);             \
    else                                                                 \
      ROUTE_LAUNCH_D(16, c2d, kernel, grid, st, __VA_ARGS__);            \
  } while (0)

static int fill_cfg(RouteCfg& cfg, const float* centroids_host, int K, float margin, int cluster_2d) {
  cfg.K = K;
  cfg.margin = margin;
  cfg.dims = cluster_2d ? 2 : 3;
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k)
    for (int a = 0; a < 3; ++a) cfg.c[k][a] = (k < K) ? centroids_host[k * 3 + a] : INFINITY;  // padding experts: distance +inf
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
  ROUTE_DISPATCH(K, cluster_2d, k_routing, persistent_grid((pts->P + RT - 1) / RT, 8), as_stream(stream), make_src(pts), pts->P, cfg, weights, hard);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int64_t acn_route_pack_capacity(int64_t P, int K) {
  // typical (margin 1.05): ~1.03 memberships per point; callers re-run with seg[K] if exceeded
  return P + P / 4 + (int64_t)K * ACN_TILE;
}

static inline int route_kt(int K) { return K <= 2 ? 2 : (K <= 4 ? 4 : (K <= 8 ? 8 : 16)); }
static inline int64_t align256(int64_t v) { return (v + 255) / 256 * 256; }

// workspace: per-block counts (nb*K + 64 ints) | per-point decision word (P) | per-point weights of the non-unit samples (P*KT)
int64_t acn_route_pack_ws_bytes(int64_t P, int K) {
  int64_t nb = (P + RB - 1) / RB;
  return align256(sizeof(int32_t) * (nb * K + 64)) + align256(sizeof(uint32_t) * P) + sizeof(float) * P * route_kt(K) + 256;
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
  ACN_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, "acn_route_pack: ws must be 16-byte aligned");
  int32_t* counts = reinterpret_cast<int32_t*>(ws);
  uint32_t* pmask = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(ws) + align256(sizeof(int32_t) * ((int64_t)nb * K + 64)));
  float* pw = reinterpret_cast<float*>(reinterpret_cast<char*>(pmask) + align256(sizeof(uint32_t) * P));
  ACN_CUDA(cudaMemsetAsync(pair_point, 0xFF, sizeof(int32_t) * Mcap, st));
  ACN_CUDA(cudaMemsetAsync(pair_weight, 0, sizeof(float) * Mcap, st));
  if (nb > 0) {
    ROUTE_DISPATCH(K, cluster_2d, k_route_count, persistent_grid(nb, 8), st, make_src(pts), P, cfg, nb, counts, pmask, pw);
    ACN_LAUNCH_CHECK();
  }
  int32_t* totals = counts + (int64_t)nb * K;  // the 64 spare ints of the workspace
  k_route_scan<<<K, 1024, 0, st>>>(counts, nb, totals);
  ACN_LAUNCH_CHECK();
  k_route_seg<<<1, 32, 0, st>>>(totals, K, seg);
  ACN_LAUNCH_CHECK();
  if (nb > 0) {
    const int grid = persistent_grid(nb, 8);
    switch (route_kt(K)) {
      case 2: k_route_write<2><<<grid, RT, 0, st>>>(P, K, nb, counts, seg, pmask, pw, pair_point, pair_weight, Mcap); break;
      case 4: k_route_write<4><<<grid, RT, 0, st>>>(P, K, nb, counts, seg, pmask, pw, pair_point, pair_weight, Mcap); break;
      case 8: k_route_write<8><<<grid, RT, 0, st>>>(P, K, nb, counts, seg, pmask, pw, pair_point, pair_weight, Mcap); break;
      default: k_route_write<16><<<grid, RT, 0, st>>>(P, K, nb, counts, seg, pmask, pw, pair_point, pair_weight, Mcap); break;
    }
    ACN_LAUNCH_CHECK();
  }
  return ACN_OK;
}

}  // extern "C"
