// Expert-parallel dispatch / combine around the sample all-to-all (new design, the reference is single-GPU:
// its per-expert nonzero / index_select / index_add_ loop is models/inr/meta_container.py:305-337).
//
//   k_a2a_count   : valid pairs per expert of the routed pair list (groups are stable: valid slots first, -1 padding last),
//                   found by a binary search per expert; also the exclusive offsets of the groups in the send buffer.
//   k_a2a_pack    : compacts the pair list (drops the padding) and builds the send rows [xyz, dir] straight from the
//                   implicit point form (rays, t) - the (N*S, 6) tensor is never materialised.  Rows are grouped by
//                   expert, and experts are placed on ranks in contiguous blocks, so the buffer is already grouped by
//                   destination rank: it is the input of ONE all_to_all_single.
//   k_a2a_combine : out[point[j]] += weight[j] * y[j] for the returned [rgb, sigma] rows (REDG.128 per pair).
//   k_a2a_combine_bwd : d_y[j] = weight[j] * d_out[point[j]] (the rows that travel back to the owners).
// All HBM-bound streaming kernels: 28 B out per routed pair for the dispatch, 16 B + one RED.128 for the combine.
#include <string.h>

#include "common.cuh"

namespace acn {

// counts[e] (e < K), counts[K] = total; offs[e] = exclusive prefix.  One warp per expert, K <= ACN_MAX_EXPERTS.
__global__ void __launch_bounds__(32 * ACN_MAX_EXPERTS) k_a2a_count(const int32_t* __restrict__ point, const int32_t* __restrict__ seg, int K,
                                                                    int64_t Mcap, int32_t* __restrict__ counts, int32_t* __restrict__ offs) {
  __shared__ int s_cnt[ACN_MAX_EXPERTS];
  const int e = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (e < K && lane == 0) {
    int64_t lo = seg[e], hi = seg[e + 1];
    if (hi > Mcap) hi = Mcap;
    if (lo > hi) lo = hi;
    const int64_t base = lo;
    while (lo < hi) {  // first slot with point < 0
      int64_t mid = (lo + hi) >> 1;
      if (__ldg(point + mid) >= 0)
        lo = mid + 1;
      else
        hi = mid;
    }
    s_cnt[e] = (int)(lo - base);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int k = 0; k < K; ++k) {
      offs[k] = run;
      counts[k] = s_cnt[k];
      run += s_cnt[k];
    }
    counts[K] = ((int64_t)seg[K] > Mcap) ? -1 : run;  // -1: the pair list was truncated at its capacity
    offs[K] = run;
  }
}

__global__ void __launch_bounds__(256) k_a2a_pack(PointSrc src, const int32_t* __restrict__ point, const float* __restrict__ weight,
                                                  const int32_t* __restrict__ seg, int K, int64_t Mcap, const int32_t* __restrict__ offs,
                                                  float* __restrict__ rows6, int32_t* __restrict__ out_point, float* __restrict__ out_weight) {
  __shared__ int s_seg[ACN_MAX_EXPERTS + 1], s_off[ACN_MAX_EXPERTS + 1];
  if (threadIdx.x <= K) {
    s_seg[threadIdx.x] = seg[threadIdx.x];
    s_off[threadIdx.x] = offs[threadIdx.x];
  }
  __syncthreads();
  int64_t M = s_seg[K];
  if (M > Mcap) M = Mcap;
  for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x) {
    const int p = __ldg(point + m);
    if (p < 0) continue;
    int e = 0;
    for (int k = 1; k < K; ++k)
      if (m >= s_seg[k]) e = k;
    const int64_t j = (int64_t)s_off[e] + (m - s_seg[e]);
    float x, y, z, dx, dy, dz;
    load_xyz(src, p, x, y, z);
    load_dir(src, p, dx, dy, dz);
    float2* r = reinterpret_cast<float2*>(rows6 + j * 6);
    r[0] = make_float2(x, y);
    r[1] = make_float2(z, dx);
    r[2] = make_float2(dy, dz);
    out_point[j] = p;
    out_weight[j] = __ldg(weight + m);
  }
}

__global__ void __launch_bounds__(256) k_a2a_combine(const float4* __restrict__ y, const int32_t* __restrict__ point, const float* __restrict__ weight,
                                                     int64_t M, float4* __restrict__ out) {
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < M; j += (int64_t)gridDim.x * blockDim.x) {
    const float w = __ldg(weight + j);
    const float4 v = __ldg(y + j);
    red_add(out + __ldg(point + j), make_float4(v.x * w, v.y * w, v.z * w, v.w * w));
  }
}

__global__ void __launch_bounds__(256) k_a2a_combine_bwd(const float4* __restrict__ d_out, const int32_t* __restrict__ point,
                                                         const float* __restrict__ weight, int64_t M, float4* __restrict__ d_y) {
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < M; j += (int64_t)gridDim.x * blockDim.x) {
    const float w = __ldg(weight + j);
    const float4 g = __ldg(d_out + __ldg(point + j));
    d_y[j] = make_float4(g.x * w, g.y * w, g.z * w, g.w * w);
  }
}

// ---- the same three steps with the rows living in the OWNER's memory (NVLink peer mappings) ---------------------
// route.base[e] is the owner's buffer as mapped into this process (or a local buffer when this rank owns e) and
// route.row0[e] the first row of this rank's block of expert e in it.  Pair j of the compact list belongs to expert
// e = max{k : offs[k] <= j} and is row route.row0[e] + (j - offs[e]) there.  Stores and loads are 8- and 16-byte
// vectors of consecutive rows across a warp: full 32-byte sectors on the NVLink path.
struct RouteDev {
  void* base[ACN_MAX_EXPERTS];
  int64_t row0[ACN_MAX_EXPERTS];
};

__device__ __forceinline__ int expert_of(const int* s_off, int K, int64_t j) {
  int e = 0;
  for (int k = 1; k < K; ++k)
    if (j >= s_off[k]) e = k;
  return e;
}

// row0 of expert e: the host's value, or the one acn_p2p_plan left on the device
__device__ __forceinline__ void load_row0(const RouteDev& route, const int64_t* __restrict__ row0_dev, int K, int64_t* s_row0) {
  if (threadIdx.x < ACN_MAX_EXPERTS) s_row0[threadIdx.x] = (row0_dev && (int)threadIdx.x < K) ? row0_dev[threadIdx.x] : route.row0[threadIdx.x];
}

__global__ void __launch_bounds__(256) k_a2a_push_rows(PointSrc src, const int32_t* __restrict__ point, const float* __restrict__ weight,
                                                       const int32_t* __restrict__ seg, int K, int64_t Mcap, const int32_t* __restrict__ offs,
                                                       RouteDev route, const int64_t* __restrict__ row0_dev, int64_t cap_rows,
                                                       int32_t* __restrict__ out_point, float* __restrict__ out_weight) {
  __shared__ int s_seg[ACN_MAX_EXPERTS + 1], s_off[ACN_MAX_EXPERTS + 1];
  __shared__ int64_t s_row0[ACN_MAX_EXPERTS];
  if (threadIdx.x <= K) {
    s_seg[threadIdx.x] = seg[threadIdx.x];
    s_off[threadIdx.x] = offs[threadIdx.x];
  }
  load_row0(route, row0_dev, K, s_row0);
  __syncthreads();
  int64_t M = s_seg[K];
  if (M > Mcap) M = Mcap;
  for (int64_t m = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x) {
    const int p = __ldg(point + m);
    if (p < 0) continue;
    int e = 0;
    for (int k = 1; k < K; ++k)
      if (m >= s_seg[k]) e = k;
    const int64_t i = m - s_seg[e];
    const int64_t row = s_row0[e] + i;
    if (cap_rows <= 0 || row < cap_rows) {
      float x, y, z, dx, dy, dz;
      load_xyz(src, p, x, y, z);
      load_dir(src, p, dx, dy, dz);
      float2* r = reinterpret_cast<float2*>(static_cast<float*>(route.base[e]) + row * 6);
      r[0] = make_float2(x, y);
      r[1] = make_float2(z, dx);
      r[2] = make_float2(dy, dz);
    }
    const int64_t j = (int64_t)s_off[e] + i;
    out_point[j] = p;
    out_weight[j] = __ldg(weight + m);
  }
}

__global__ void __launch_bounds__(256) k_a2a_pull_combine(RouteDev route, const int64_t* __restrict__ row0_dev, const int32_t* __restrict__ point,
                                                          const float* __restrict__ weight, const int32_t* __restrict__ offs, int K, int64_t M,
                                                          float4* __restrict__ out) {
  __shared__ int s_off[ACN_MAX_EXPERTS + 1];
  __shared__ int64_t s_row0[ACN_MAX_EXPERTS];
  if (threadIdx.x <= K) s_off[threadIdx.x] = offs[threadIdx.x];
  load_row0(route, row0_dev, K, s_row0);
  __syncthreads();
  if (M > s_off[K]) M = s_off[K];  // the live pair count is the device's
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < M; j += (int64_t)gridDim.x * blockDim.x) {
    const int e = expert_of(s_off, K, j);
    const float4 v = __ldcg(static_cast<const float4*>(route.base[e]) + s_row0[e] + (j - s_off[e]));  // L2-coherent: a peer wrote it
    const float w = __ldg(weight + j);
    red_add(out + __ldg(point + j), make_float4(v.x * w, v.y * w, v.z * w, v.w * w));
  }
}

__global__ void __launch_bounds__(256) k_a2a_push_grad(const float4* __restrict__ d_out, const int32_t* __restrict__ point,
                                                       const float* __restrict__ weight, const int32_t* __restrict__ offs, int K, int64_t M,
                                                       RouteDev route, const int64_t* __restrict__ row0_dev, int64_t cap_rows) {
  __shared__ int s_off[ACN_MAX_EXPERTS + 1];
  __shared__ int64_t s_row0[ACN_MAX_EXPERTS];
  if (threadIdx.x <= K) s_off[threadIdx.x] = offs[threadIdx.x];
  load_row0(route, row0_dev, K, s_row0);
  __syncthreads();
  if (M > s_off[K]) M = s_off[K];
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < M; j += (int64_t)gridDim.x * blockDim.x) {
    const int e = expert_of(s_off, K, j);
    const int64_t row = s_row0[e] + (j - s_off[e]);
    if (cap_rows > 0 && row >= cap_rows) continue;
    const float w = __ldg(weight + j);
    const float4 g = __ldg(d_out + __ldg(point + j));
    *(static_cast<float4*>(route.base[e]) + row) = make_float4(g.x * w, g.y * w, g.z * w, g.w * w);
  }
}

// ---- device-side barrier (count exchange, small all-reduce) between the ranks of one box --------------------------
// Every slab starts with a sync area of ACN_P2P_SYNC_BYTES: int32 flag[16] | status | epoch | pad to 32 ints | table[16][17] |
// pad to 4096 B | float payload[2][16][ACN_P2P_PAYLOAD_FLOATS].
// The barrier number (1, 2, ... the same sequence on every rank) lives in the rank's OWN slab (int 17) and is advanced by
// the kernel: a launch carries no per-call state and replays from a CUDA graph.  Rank r stores the epoch into flag[r] of
// EVERY rank's slab (st.release.sys after a system fence, so the peer stores of this rank's earlier kernels are visible to
// whoever acquires the flag), then waits until all flags of its OWN slab have reached it (ld.acquire.sys).
//  counts : written into row `rank` of every slab's table first - one launch replaces an all-gather plus a barrier.
//  payload: written into slot [epoch & 1][rank] of every slab; after the wait every rank sums the slots in rank order
//           (bitwise identical on all ranks).  Two buffers: a rank can run at most one barrier ahead of the slowest
//           reader, so the slot it overwrites two barriers later has been consumed.
// The wait is bounded; on expiry the status word gets bit 0 and the kernel traps (nothing downstream may consume a
// half-written slab).
struct PeerSync {
  int32_t* base[ACN_MAX_PEERS];
};

__device__ __forceinline__ void st_release_sys(int32_t* p, int32_t v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_acquire_sys(const int32_t* p) {
  int32_t v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

constexpr int SYNC_STATUS = 16, SYNC_EPOCH = 17, SYNC_TABLE = 32, SYNC_PAYLOAD = 1024;  // int / float offsets in the sync area
constexpr int BARRIER_THREADS = 128;

__global__ void __launch_bounds__(BARRIER_THREADS) k_p2p_barrier(PeerSync peers, int world, int rank, const int32_t* __restrict__ counts,
                                                                 int n_counts, const float* __restrict__ payload, int n_payload,
                                                                 float* __restrict__ reduced, float scale, unsigned long long timeout_ns,
                                                                 int phase) {
  // phase 0: arrive + wait.  Split form: phase 1 = ARRIVE (advance the epoch, publish counts / payload, raise this rank's flag in
  // every slab, return at once); phase 2 = WAIT for the epoch of the last arrive, then reduce.  Work a rank enqueues between its
  // arrive and its wait absorbs the skew between the ranks (the reduction of gradients that are ready early in a backward pass).
  // Every arrive is followed by its wait before the next barrier of any kind on the same exchange.
  __shared__ int32_t s_epoch;
  const int t = threadIdx.x;
  if (t == 0) {
    int32_t* e = peers.base[rank] + SYNC_EPOCH;
    s_epoch = (phase == 2) ? *e : *e + 1;
    *e = s_epoch;
  }
  __syncthreads();
  const int32_t epoch = s_epoch;
  const int slot = (epoch & 1) * ACN_MAX_PEERS * ACN_P2P_PAYLOAD_FLOATS;
  if (phase == 2) goto wait_phase;
  if (counts != nullptr && t < world) {
    int32_t* row = peers.base[t] + SYNC_TABLE + rank * (ACN_MAX_EXPERTS + 1);
    for (int i = 0; i < n_counts; ++i) row[i] = counts[i];
  }
  if (payload != nullptr) {
    for (int s = 0; s < world; ++s) {
      float* dst = reinterpret_cast<float*>(peers.base[s]) + SYNC_PAYLOAD + slot + rank * ACN_P2P_PAYLOAD_FLOATS;
      for (int i = t; i < n_payload; i += BARRIER_THREADS) dst[i] = payload[i];
    }
  }
  __threadfence_system();
  __syncthreads();
  if (t < world) st_release_sys(peers.base[t] + rank, epoch);
  if (phase == 1) return;
wait_phase:
  if (t < world) {
    const int32_t* mine = peers.base[rank] + t;
    const unsigned long long t0 = global_ns();
    unsigned spins = 0;
    while (ld_acquire_sys(mine) - epoch < 0) {
      if ((++spins & 1023u) == 0u && global_ns() - t0 > timeout_ns) {  // a peer never arrived
        atomicOr(peers.base[rank] + SYNC_STATUS, 1);
        __threadfence_system();
        __trap();
      }
    }
  }
  __syncthreads();
  if (n_payload > 0 && reduced != nullptr) {
    const float* src = reinterpret_cast<const float*>(peers.base[rank]) + SYNC_PAYLOAD + slot;
    for (int i = t; i < n_payload; i += BARRIER_THREADS) {
      float acc = 0.f;
      for (int s = 0; s < world; ++s) acc += ld_relaxed_sys(src + s * ACN_P2P_PAYLOAD_FLOATS + i);  // rank order: same bits everywhere
      reduced[i] = acc * scale;
    }
  }
}

// ---- device-side exchange layout (parallel.peer_layout on the GPU) --------------------------------------------------
struct OwnerMap {
  int32_t owner[ACN_MAX_EXPERTS];
};

__global__ void __launch_bounds__(256) k_p2p_plan(const int32_t* __restrict__ sync_own, int world, int rank, int K, OwnerMap om, int64_t cap_rows,
                                                  int64_t* __restrict__ row0_dev, int32_t* __restrict__ seg_own, int32_t* __restrict__ point_own,
                                                  int32_t* __restrict__ live_own, float4* __restrict__ g_rows, int32_t* __restrict__ status) {
  __shared__ int s_tot[ACN_MAX_EXPERTS], s_seg[ACN_MAX_EXPERTS + 1], s_totown[ACN_MAX_EXPERTS];
  __shared__ int s_nown;
  if (threadIdx.x == 0) {
    const int32_t* table = sync_own + SYNC_TABLE;
    int used[ACN_MAX_PEERS];
    for (int r = 0; r < ACN_MAX_PEERS; ++r) used[r] = 0;
    int st = 0, nown = 0;
    for (int s = 0; s < world; ++s)
      if (table[s * (ACN_MAX_EXPERTS + 1) + K] < 0) st |= 4;
    for (int e = 0; e < K; ++e) {
      int tot = 0, before = 0;
      for (int s = 0; s < world; ++s) {
        const int c = table[s * (ACN_MAX_EXPERTS + 1) + e];
        tot += c;
        before += (s < rank) ? c : 0;
      }
      const int o = om.owner[e];
      const int start = used[o];
      used[o] += (tot + ACN_TILE - 1) / ACN_TILE * ACN_TILE;
      if (blockIdx.x == 0) row0_dev[e] = (int64_t)start + before;
      if (o == rank) {
        s_seg[nown] = start;
        s_totown[nown] = tot;
        ++nown;
      }
      s_tot[e] = tot;
    }
    for (int r = 0; r < world; ++r)
      if (used[r] > cap_rows) st |= 2;
    int end = used[rank];
    if (end > cap_rows) end = (int)(cap_rows / ACN_TILE * ACN_TILE);
    s_seg[nown] = end;
    for (int j = 0; j < nown; ++j)
      if (s_seg[j] > end) s_seg[j] = end;
    s_nown = nown;
    if (blockIdx.x == 0) {
      for (int j = 0; j <= nown; ++j) seg_own[j] = s_seg[j];
      for (int j = 0; j < nown; ++j) live_own[j] = s_totown[j] > 0 ? 1 : 0;
      if (st) atomicOr(status, st);
    }
  }
  __syncthreads();
  const int nown = s_nown;
  const int end = s_seg[nown];
  for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < end; m += gridDim.x * blockDim.x) {
    int j = 0;
    for (int k = 1; k < nown; ++k)
      if (m >= s_seg[k]) j = k;
    const bool live = (m - s_seg[j]) < s_totown[j];
    point_own[m] = live ? m : -1;
    if (!live && g_rows) g_rows[m] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void k_route_status(const int32_t* __restrict__ seg, int K, int64_t Mcap, int32_t* __restrict__ live, int32_t* __restrict__ status) {
  const int k = threadIdx.x;
  if (k < K && live) live[k] = seg[k + 1] > seg[k] ? 1 : 0;
  if (k == 0 && status && (int64_t)seg[K] > Mcap) atomicOr(status, 1);
}

static inline bool make_route(const AcnA2aRoute* r, int K, RouteDev& d) {
  for (int e = 0; e < ACN_MAX_EXPERTS; ++e) {
    d.base[e] = e < K ? r->base[e] : nullptr;
    d.row0[e] = e < K ? r->row0[e] : 0;
    if (e < K && (!d.base[e] || d.row0[e] < 0)) return false;
  }
  return true;
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_a2a_pack(const AcnPoints* pts, const AcnPairs* pairs, int K, float* rows6, int32_t* out_point, float* out_weight, int32_t* counts,
                 int32_t* offs, void* stream) {
  ACN_REQUIRE(pts && pairs && rows6 && out_point && out_weight && counts && offs, "acn_a2a_pack: NULL argument");
  ACN_REQUIRE(K >= 1 && K <= ACN_MAX_EXPERTS, "acn_a2a_pack: K must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(pairs->point && pairs->weight && pairs->seg, "acn_a2a_pack: needs a routed pair list (not the identity list)");
  ACN_REQUIRE(pts->x6 || (pts->rays && pts->t && pts->S >= 1), "acn_a2a_pack: points need x6 or (rays,t,S)");
  ACN_REQUIRE(pts->P < ((int64_t)1 << 31), "acn_a2a_pack: chunk the point set below 2^31");
  cudaStream_t st = as_stream(stream);
  k_a2a_count<<<1, 32 * ACN_MAX_EXPERTS, 0, st>>>(pairs->point, pairs->seg, K, pairs->Mcap, counts, offs);
  ACN_LAUNCH_CHECK();
  if (pairs->Mcap > 0) {
    k_a2a_pack<<<persistent_grid((pairs->Mcap + 255) / 256, 8), 256, 0, st>>>(make_src(pts), pairs->point, pairs->weight, pairs->seg, K, pairs->Mcap,
                                                                             offs, rows6, out_point, out_weight);
    ACN_LAUNCH_CHECK();
  }
  return ACN_OK;
}

int acn_a2a_combine(const float* y, const int32_t* point, const float* weight, int64_t M, float* out, void* stream) {
  ACN_REQUIRE(M >= 0 && out && (M == 0 || (y && point && weight)), "acn_a2a_combine: bad arguments");
  if (M == 0) return ACN_OK;
  k_a2a_combine<<<persistent_grid((M + 255) / 256, 8), 256, 0, as_stream(stream)>>>(reinterpret_cast<const float4*>(y), point, weight, M,
                                                                                   reinterpret_cast<float4*>(out));
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_a2a_combine_bwd(const float* d_out, const int32_t* point, const float* weight, int64_t M, float* d_y, void* stream) {
  ACN_REQUIRE(M >= 0 && (M == 0 || (d_out && point && weight && d_y)), "acn_a2a_combine_bwd: bad arguments");
  if (M == 0) return ACN_OK;
  k_a2a_combine_bwd<<<persistent_grid((M + 255) / 256, 8), 256, 0, as_stream(stream)>>>(reinterpret_cast<const float4*>(d_out), point, weight, M,
                                                                                       reinterpret_cast<float4*>(d_y));
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_a2a_count(const AcnPairs* pairs, int K, int32_t* counts, int32_t* offs, void* stream) {
  ACN_REQUIRE(pairs && counts && offs, "acn_a2a_count: NULL argument");
  ACN_REQUIRE(K >= 1 && K <= ACN_MAX_EXPERTS, "acn_a2a_count: K must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(pairs->point && pairs->seg, "acn_a2a_count: needs a routed pair list (not the identity list)");
  k_a2a_count<<<1, 32 * ACN_MAX_EXPERTS, 0, as_stream(stream)>>>(pairs->point, pairs->seg, K, pairs->Mcap, counts, offs);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_a2a_push_rows(const AcnPoints* pts, const AcnPairs* pairs, int K, const int32_t* offs, const AcnA2aRoute* route, const int64_t* row0_dev,
                      int64_t cap_rows, int32_t* out_point, float* out_weight, void* stream) {
  ACN_REQUIRE(pts && pairs && offs && route && out_point && out_weight, "acn_a2a_push_rows: NULL argument");
  ACN_REQUIRE(K >= 1 && K <= ACN_MAX_EXPERTS, "acn_a2a_push_rows: K must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(pairs->point && pairs->weight && pairs->seg, "acn_a2a_push_rows: needs a routed pair list (not the identity list)");
  ACN_REQUIRE(pts->x6 || (pts->rays && pts->t && pts->S >= 1), "acn_a2a_push_rows: points need x6 or (rays,t,S)");
  ACN_REQUIRE(pts->P < ((int64_t)1 << 31), "acn_a2a_push_rows: chunk the point set below 2^31");
  RouteDev rd;
  ACN_REQUIRE(make_route(route, K, rd), "acn_a2a_push_rows: route has a NULL base or a negative row");
  if (pairs->Mcap == 0) return ACN_OK;
  k_a2a_push_rows<<<persistent_grid((pairs->Mcap + 255) / 256, 8), 256, 0, as_stream(stream)>>>(make_src(pts), pairs->point, pairs->weight, pairs->seg,
                                                                                              K, pairs->Mcap, offs, rd, row0_dev, cap_rows, out_point,
                                                                                              out_weight);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_a2a_pull_combine(const AcnA2aRoute* route, const int64_t* row0_dev, const int32_t* point, const float* weight, const int32_t* offs, int K,
                         int64_t M, float* out, void* stream) {
  ACN_REQUIRE(route && offs && out && M >= 0 && (M == 0 || (point && weight)), "acn_a2a_pull_combine: bad arguments");
  ACN_REQUIRE(K >= 1 && K <= ACN_MAX_EXPERTS, "acn_a2a_pull_combine: K must be in [1,%d]", ACN_MAX_EXPERTS);
  RouteDev rd;
  ACN_REQUIRE(make_route(route, K, rd), "acn_a2a_pull_combine: route has a NULL base or a negative row");
  if (M == 0) return ACN_OK;
  k_a2a_pull_combine<<<persistent_grid((M + 255) / 256, 8), 256, 0, as_stream(stream)>>>(rd, row0_dev, point, weight, offs, K, M,
                                                                                        reinterpret_cast<float4*>(out));
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_a2a_push_grad(const float* d_out, const int32_t* point, const float* weight, const int32_t* offs, int K, int64_t M,
                      const AcnA2aRoute* route, const int64_t* row0_dev, int64_t cap_rows, void* stream) {
  ACN_REQUIRE(route && offs && M >= 0 && (M == 0 || (d_out && point && weight)), "acn_a2a_push_grad: bad arguments");
  ACN_REQUIRE(K >= 1 && K <= ACN_MAX_EXPERTS, "acn_a2a_push_grad: K must be in [1,%d]", ACN_MAX_EXPERTS);
  RouteDev rd;
  ACN_REQUIRE(make_route(route, K, rd), "acn_a2a_push_grad: route has a NULL base or a negative row");
  if (M == 0) return ACN_OK;
  k_a2a_push_grad<<<persistent_grid((M + 255) / 256, 8), 256, 0, as_stream(stream)>>>(reinterpret_cast<const float4*>(d_out), point, weight, offs, K, M,
                                                                                     rd, row0_dev, cap_rows);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_route_status(const int32_t* seg, int K, int64_t Mcap, int32_t* live, int32_t* status, void* stream) {
  ACN_REQUIRE(seg && K >= 1 && K <= ACN_MAX_EXPERTS && (live || status), "acn_route_status: bad arguments");
  k_route_status<<<1, 32, 0, as_stream(stream)>>>(seg, K, Mcap, live, status);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_p2p_plan(const void* sync_base_own, int world, int rank, int K, const int32_t* owner_host, int64_t cap_rows, int64_t* row0_dev,
                 int32_t* seg_own, int32_t* point_own, int32_t* live_own, float* g_rows, int32_t* status, void* stream) {
  ACN_REQUIRE(sync_base_own && owner_host && row0_dev && seg_own && point_own && live_own && status, "acn_p2p_plan: NULL argument");
  ACN_REQUIRE(world >= 1 && world <= ACN_MAX_PEERS && rank >= 0 && rank < world, "acn_p2p_plan: bad world / rank");
  ACN_REQUIRE(K >= 1 && K <= ACN_MAX_EXPERTS && cap_rows >= ACN_TILE && cap_rows < ((int64_t)1 << 31), "acn_p2p_plan: bad K / capacity");
  OwnerMap om;
  for (int e = 0; e < ACN_MAX_EXPERTS; ++e) {
    om.owner[e] = e < K ? owner_host[e] : 0;
    ACN_REQUIRE(om.owner[e] >= 0 && om.owner[e] < world, "acn_p2p_plan: expert %d has owner %d outside the world", e, om.owner[e]);
  }
  k_p2p_plan<<<persistent_grid((cap_rows + 255) / 256, 2), 256, 0, as_stream(stream)>>>(static_cast<const int32_t*>(sync_base_own), world, rank, K, om,
                                                                                       cap_rows, row0_dev, seg_own, point_own, live_own,
                                                                                       reinterpret_cast<float4*>(g_rows), status);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

/* ---- exchange slabs: memory one rank allocates and its peers map over NVLink (CUDA IPC, one process per GPU) ---- */
int acn_p2p_alloc(int64_t bytes, void** ptr, void* handle64) {
  ACN_REQUIRE(bytes > 0 && ptr && handle64, "acn_p2p_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  void* p = nullptr;
  ACN_CUDA(cudaMalloc(&p, (size_t)bytes));
  cudaError_t e = cudaMemset(p, 0, (size_t)bytes);  // rows a peer never wrote are read (and ignored) by whole-tile kernels: keep them finite
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("acn_p2p_alloc: cudaMemset: %s", cudaGetErrorString(e));
    return ACN_ECUDA;
  }
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("acn_p2p_alloc: cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
    return ACN_ECUDA;
  }
  memcpy(handle64, &h, 64);
  *ptr = p;
  return ACN_OK;
}

int acn_p2p_free(void* ptr) {
  if (ptr) ACN_CUDA(cudaFree(ptr));
  return ACN_OK;
}

int acn_p2p_open(const void* handle64, void** ptr) {
  ACN_REQUIRE(handle64 && ptr, "acn_p2p_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  ACN_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return ACN_OK;
}

int acn_p2p_close(void* ptr) {
  if (ptr) ACN_CUDA(cudaIpcCloseMemHandle(ptr));
  return ACN_OK;
}

static int launch_barrier(void* const* sync_bases_host, int world, int rank, const int32_t* counts, int n_counts, const float* payload, int n_payload,
                          float* reduced, float scale, float timeout_s, int phase, void* stream);

int acn_p2p_barrier(void* const* sync_bases_host, int world, int rank, const int32_t* counts, int n_counts, const float* payload, int n_payload,
                    float* reduced, float scale, float timeout_s, void* stream) {
  ACN_REQUIRE(n_payload == 0 || (payload && reduced), "acn_p2p_barrier: bad payload");
  return launch_barrier(sync_bases_host, world, rank, counts, n_counts, payload, n_payload, reduced, scale, timeout_s, 0, stream);
}

int acn_p2p_arrive(void* const* sync_bases_host, int world, int rank, const float* payload, int n_payload, void* stream) {
  ACN_REQUIRE(n_payload == 0 || payload, "acn_p2p_arrive: bad payload");
  return launch_barrier(sync_bases_host, world, rank, nullptr, 0, payload, n_payload, nullptr, 1.f, 0.f, 1, stream);
}

int acn_p2p_wait(void* const* sync_bases_host, int world, int rank, int n_payload, float* reduced, float scale, float timeout_s, void* stream) {
  ACN_REQUIRE(n_payload == 0 || reduced, "acn_p2p_wait: bad payload");
  return launch_barrier(sync_bases_host, world, rank, nullptr, 0, nullptr, n_payload, reduced, scale, timeout_s, 2, stream);
}

static int launch_barrier(void* const* sync_bases_host, int world, int rank, const int32_t* counts, int n_counts, const float* payload, int n_payload,
                          float* reduced, float scale, float timeout_s, int phase, void* stream) {
  ACN_REQUIRE(sync_bases_host && world >= 1 && world <= ACN_MAX_PEERS && rank >= 0 && rank < world, "acn_p2p_barrier: bad world / rank");
  ACN_REQUIRE(n_counts >= 0 && n_counts <= ACN_MAX_EXPERTS + 1 && (n_counts == 0 || counts), "acn_p2p_barrier: bad counts");
  ACN_REQUIRE(n_payload >= 0 && n_payload <= ACN_P2P_PAYLOAD_FLOATS, "acn_p2p_barrier: bad payload size");
  PeerSync ps;
  for (int r = 0; r < ACN_MAX_PEERS; ++r) {
    ps.base[r] = r < world ? static_cast<int32_t*>(sync_bases_host[r]) : nullptr;
    ACN_REQUIRE(r >= world || ps.base[r], "acn_p2p_barrier: NULL slab for rank %d", r);
  }
  if (!(timeout_s > 0.f)) timeout_s = 120.f;
  k_p2p_barrier<<<1, BARRIER_THREADS, 0, as_stream(stream)>>>(ps, world, rank, n_counts ? counts : nullptr, n_counts, n_payload ? payload : nullptr,
                                                             n_payload, reduced, scale, (unsigned long long)((double)timeout_s * 1e9), phase);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
