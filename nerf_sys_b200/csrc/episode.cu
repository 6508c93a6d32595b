// (f)-3: the episode sampler's ray -> micro-cell routing (data/task_dataset.py:130-171, 242-351, 353-418, 544-627).
// The reference runs it on the CPU over every training ray before the first step (a 64-iteration vectorised DDA with ~30
// elementwise torch ops per iteration); here it is one thread per ray.  Every decision is an integer (cell id, valid / keep
// flags) compared bit for bit with the reference, so the arithmetic follows the torch expressions operation by operation
// (compiled -fmad=false; torch.minimum / maximum / max propagate NaN, comparisons with NaN are false).
#include <math.h>

#include "common.cuh"

namespace acn {

struct EpGrid {
  float lo[3], hi[3];
  int n[3];
};

__device__ __forceinline__ float t_min(float a, float b) { return (a != a || b != b) ? NAN : fminf(a, b); }  // torch.minimum
__device__ __forceinline__ float t_max(float a, float b) { return (a != a || b != b) ? NAN : fmaxf(a, b); }  // torch.maximum

// TaskDataset._aabb_intersect (:130-153)
__device__ __forceinline__ bool aabb_intersect(const float* o, const float* d, const float* lo, const float* hi, float& t_entry, float& t_exit) {
  bool miss_parallel = false;
  float te = -INFINITY, tx = INFINITY;
  bool te_nan = false, tx_nan = false;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const bool parallel = fabsf(d[a]) < 1e-12f;
    const float inv = 1.0f / d[a];
    const float t0 = (lo[a] - o[a]) * inv, t1 = (hi[a] - o[a]) * inv;
    const float mn = t_min(t0, t1), mx = t_max(t0, t1);
    miss_parallel = miss_parallel || (parallel && !((o[a] >= lo[a]) && (o[a] <= hi[a])));
    te_nan = te_nan || (mn != mn);
    tx_nan = tx_nan || (mx != mx);
    te = fmaxf(te, mn);  // max / min over the axis dimension propagate NaN (handled by the flags)
    tx = fminf(tx, mx);
  }
  t_entry = te_nan ? NAN : te;
  t_exit = tx_nan ? NAN : tx;
  return (t_exit >= t_entry) && !miss_parallel;
}

// _overlap_len_with_cell (:216-232) for one ray and one cell box [lo(3), hi(3)]
__device__ __forceinline__ float overlap_len(const float* o, const float* d, bool have_bounds, float near, float far, const float* cb) {
  float te, tx;
  const bool hit = aabb_intersect(o, d, cb, cb + 3, te, tx);
  float t0 = t_max(te, 0.f), t1 = tx;
  if (have_bounds) {
    t0 = t_max(t0, near);
    t1 = t_min(t1, far);
  }
  float len = t1 - t0;
  len = (len != len) ? len : fmaxf(len, 0.f);  // clamp_min keeps NaN
  return hit ? len : 0.f;
}

__device__ __forceinline__ float big_if_nonfinite(float v) { return (v != v || isinf(v)) ? 1e30f : v; }  // nan_to_num_(nan=big, posinf=big, neginf=big)

__device__ __forceinline__ int clampi(long long v, int hi) { return v < 0 ? 0 : (v > hi ? hi : (int)v); }

__global__ void __launch_bounds__(128) k_episode_route(const float* __restrict__ rays, int ncol, int64_t N, EpGrid g, const float* __restrict__ cell_bounds,
                                                       const float* __restrict__ tol_cell, int policy, float alpha, float tol_abs, int max_steps,
                                                       uint8_t* __restrict__ valid_out, int32_t* __restrict__ cid_out, float* __restrict__ len_out,
                                                       uint8_t* __restrict__ keep_out) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= N) return;
  const float* q = rays + r * ncol;
  const float o[3] = {q[0], q[1], q[2]}, d[3] = {q[3], q[4], q[5]};
  const bool have_bounds = ncol >= 8;
  const float near = have_bounds ? q[6] : 0.f, far = have_bounds ? q[7] : 0.f;
  // _region_segment (:155-171)
  float te, tx;
  const bool hit = aabb_intersect(o, d, g.lo, g.hi, te, tx);
  float t0 = t_max(te, 0.f), t1 = tx;
  if (have_bounds) {
    t0 = t_max(t0, near);
    t1 = t_min(t1, far);
  }
  const float seg = t1 - t0;
  const bool valid = hit && (seg > 0.f);
  valid_out[r] = valid ? 1 : 0;
  if (!valid) {
    cid_out[r] = -1;
    len_out[r] = 0.f;
    keep_out[r] = 0;
    return;
  }
  const int nx = g.n[0], ny = g.n[1], nz = g.n[2], nyz = ny * nz;
  int cid_final;
  float len_best;
  if (policy == 0) {
    // ---- _dda_transform + _dda_init + _dda_maxoverlap (:242-351)
    float g_o[3], g_d[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      float cell = (g.hi[a] - g.lo[a]) / (float)g.n[a];
      cell = fmaxf(cell, 1e-12f);
      g_o[a] = (o[a] - g.lo[a]) / cell;
      g_d[a] = d[a] / cell;
    }
    const float ts = t0 + 1e-6f;
    float p[3], tMax[3], tDelta[3];
    long long idx[3];
    int stp[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      p[a] = g_o[a] + g_d[a] * ts;
      idx[a] = (long long)floorf(p[a]);
      stp[a] = g_d[a] > 0.f ? 1 : (g_d[a] < 0.f ? -1 : 0);
      const float nb = stp[a] > 0 ? floorf(p[a]) + 1.0f : ceilf(p[a]) - 1.0f;
      const float inv = 1.0f / g_d[a];
      tMax[a] = big_if_nonfinite((nb - p[a]) * inv);
      tDelta[a] = big_if_nonfinite((float)stp[a] * inv);
    }
    int ix = clampi(idx[0], nx - 1), iy = clampi(idx[1], ny - 1), iz = clampi(idx[2], nz - 1);
    float t = t0, best_len = 0.f;
    int best_cid = ix * nyz + iy * nz + iz;
    for (int it = 0; it < max_steps; ++it) {
      const float m = t_min(t_min(tMax[0], tMax[1]), tMax[2]);
      const float t_next = t_min(m, t1);
      float dt = t_next - t;
      dt = (dt != dt) ? dt : fmaxf(dt, 0.f);
      const int cid = ix * nyz + iy * nz + iz;
      if (dt > best_len) {
        best_len = dt;
        best_cid = cid;
      }
      if (t_next >= t1) break;  // the reference stops once every ray is done; further iterations add dt = 0
      const bool adv_x = (tMax[0] <= tMax[1]) && (tMax[0] <= tMax[2]);
      const bool adv_y = !(tMax[0] <= tMax[1]) && (tMax[1] <= tMax[2]);
      const bool adv_z = !(adv_x || adv_y);
      if (adv_x) {
        ix = clampi((long long)ix + stp[0], nx - 1);
        tMax[0] = tMax[0] + tDelta[0];
      }
      if (adv_y) {
        iy = clampi((long long)iy + stp[1], ny - 1);
        tMax[1] = tMax[1] + tDelta[1];
      }
      if (adv_z) {
        iz = clampi((long long)iz + stp[2], nz - 1);
        tMax[2] = tMax[2] + tDelta[2];
      }
      t = t_next;
    }
    cid_final = best_cid;
    len_best = best_len;
  } else {
    // ---- _alpha_local_max_overlap (:353-418): cell of the alpha point, then the best of it and its 6 neighbours by overlap
    float ta = t0 + alpha * (t1 - t0);
    ta = ta + 1e-6f * (t1 - t0);
    int ic[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float pa = o[a] + d[a] * ta;
      float rel = (pa - g.lo[a]) / fmaxf(g.hi[a] - g.lo[a], 1e-9f);
      rel = (rel != rel) ? rel : fminf(fmaxf(rel, 0.0f), 1.0f - 1e-7f);
      ic[a] = clampi((long long)floorf(rel * (float)g.n[a]), g.n[a] - 1);
    }
    const int cid_primary = ic[0] * nyz + ic[1] * nz + ic[2];
    const int dx[7] = {-1, 1, 0, 0, 0, 0, 0}, dy[7] = {0, 0, -1, 1, 0, 0, 0}, dz[7] = {0, 0, 0, 0, -1, 1, 0};
    float best = 0.f;
    int cid_best = 0;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const int cx = clampi((long long)ic[0] + dx[k], nx - 1), cy = clampi((long long)ic[1] + dy[k], ny - 1), cz = clampi((long long)ic[2] + dz[k], nz - 1);
      const int c = cx * nyz + cy * nz + cz;
      const float len = overlap_len(o, d, have_bounds, near, far, cell_bounds + (int64_t)c * 6);
      // argmax: first maximal value wins; a NaN is maximal for torch.argmax
      const bool better = (k == 0) || (len > best) || (len != len && !(best != best));
      if (better) {
        best = len;
        cid_best = c;
      }
    }
    const float thr = t_max(tol_abs, 1e-6f * seg);
    const bool ok = best >= thr;
    cid_final = ok ? cid_best : cid_primary;
    len_best = best;
  }
  cid_out[r] = cid_final;
  len_out[r] = len_best;
  // _route_and_bin (:585-597): overlap with the SELECTED cell against that cell's tolerance
  const float len_sel = overlap_len(o, d, have_bounds, near, far, cell_bounds + (int64_t)cid_final * 6);
  keep_out[r] = (len_sel >= tol_cell[cid_final]) ? 1 : 0;
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_episode_route(const float* rays, int ncol, int64_t N, const float* aabb_host, const int32_t* cells_host, const float* cell_bounds,
                      const float* tol_per_cell, int policy, float alpha, float tol_abs, int max_steps, uint8_t* valid, int32_t* cid, float* len_best,
                      uint8_t* keep, void* stream) {
  ACN_REQUIRE(rays && aabb_host && cells_host && cell_bounds && tol_per_cell && valid && cid && len_best && keep, "acn_episode_route: NULL argument");
  ACN_REQUIRE((ncol == 6 || ncol >= 8) && N >= 0 && (policy == 0 || policy == 1) && max_steps >= 1, "acn_episode_route: bad arguments");
  ACN_REQUIRE(cells_host[0] >= 1 && cells_host[1] >= 1 && cells_host[2] >= 1 &&
                  (int64_t)cells_host[0] * cells_host[1] * cells_host[2] < ((int64_t)1 << 31),
              "acn_episode_route: bad cell counts");
  if (N == 0) return ACN_OK;
  EpGrid g;
  for (int a = 0; a < 3; ++a) {
    g.lo[a] = aabb_host[a];
    g.hi[a] = aabb_host[3 + a];
    g.n[a] = cells_host[a];
  }
  k_episode_route<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(rays, ncol, N, g, cell_bounds, tol_per_cell, policy, alpha, tol_abs, max_steps,
                                                                             valid, cid, len_best, keep);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
