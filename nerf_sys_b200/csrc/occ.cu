// a21 / (f)-1: the occupancy-guided renderer's per-ray work (nerfs/ray_rendering.py:349-558, models/inr/meta_ngp.py:389-443).
// The reference delegates the grid traversal, the packed transmittance scan and the segmented sums to nerfacc 0.5.3 (not
// vendored, absent here: parity unpinned) and merges the experts' segments in a Python loop over rays (:196-256).  Here:
//   k_occ_march     multi-level occupancy-grid traversal (Amanatides-Woo DDA per level, finest containing level wins, samples of
//                   length clamp(t*cone_angle, step, inf) tiled through runs of occupied cells), count pass + write pass
//   k_occ_visible   nerfacc's render_visibility_from_density: transmittance >= early_stop_eps and alpha >= alpha_thre
//   k_seg_union     per-ray K-way merge of the experts' interval boundaries into one partition (torch.unique semantics)
//   k_packed_fwd/bwd  weights from densities over packed segments + the three segmented sums, and their gradients
// One thread per ray everywhere: rays are independent, lists are short (tens to hundreds of segments), the work is a serial scan.
// Compiled with -fmad=false: cell ids and segment counts are integer outputs and must match oracle/acnerf_oracle.py bit for bit.
#include <math.h>

#include "common.cuh"

namespace acn {

struct OccGrid {
  const uint8_t* bin;  // (levels, R, R, R) occupancy bits as bytes, x-major (cell = (ix*R + iy)*R + iz)
  int levels, R;
  float aabb[ACN_OCC_MAX_LEVELS][6];
};

__device__ __forceinline__ float occ_dt(float t, float cone, float step) { return fmaxf(t * cone, step); }

// slab test of one level's box, nerfacc's convention: a zero direction component gives +-inf and drops out of the min / max
__device__ __forceinline__ bool box_span(const float* o, const float* d, const float* bb, float& t0, float& t1) {
  t0 = -INFINITY;
  t1 = INFINITY;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float inv = 1.0f / d[a];
    float ta = (bb[a] - o[a]) * inv, tb = (bb[3 + a] - o[a]) * inv;
    if (d[a] == 0.f) {  // parallel to the slab: inside -> no constraint, outside -> miss
      if (o[a] < bb[a] || o[a] > bb[3 + a]) return false;
      continue;
    }
    t0 = fmaxf(t0, fminf(ta, tb));
    t1 = fminf(t1, fmaxf(ta, tb));
  }
  return t1 > t0;
}

// The traversal of one ray.  EMIT(t0, t1, level, cell) is called for every sample; returns the number of samples.
template <typename Emit>
__device__ __forceinline__ int march_ray(const OccGrid& g, const float* o, const float* d, float near, float far, float step, float cone,
                                         int max_samples, Emit emit) {
  // spans of every level (nested boxes), clipped to [near, far]
  float lo[ACN_OCC_MAX_LEVELS], hi[ACN_OCC_MAX_LEVELS];
  bool hit[ACN_OCC_MAX_LEVELS];
  for (int l = 0; l < g.levels; ++l) {
    hit[l] = box_span(o, d, g.aabb[l], lo[l], hi[l]);
    lo[l] = fmaxf(lo[l], near);
    hi[l] = fminf(hi[l], far);
    hit[l] = hit[l] && hi[l] > lo[l];
  }
  int n = 0;
  bool continuous = false;
  float t_last = near;
  float cur = near;  // everything before `cur` is done
  while (n < max_samples) {
    // finest level whose span contains `cur`, and the end of the stretch on which it stays the finest
    int lvl = -1;
    for (int l = 0; l < g.levels; ++l)
      if (hit[l] && lo[l] <= cur && cur < hi[l]) {
        lvl = l;
        break;
      }
    float end;
    if (lvl < 0) {  // outside every box: jump to the next box entry
      end = INFINITY;
      for (int l = 0; l < g.levels; ++l)
        if (hit[l] && lo[l] > cur) end = fminf(end, lo[l]);
      if (!(end < INFINITY)) break;
      cur = end;
      continuous = false;
      continue;
    }
    end = hi[lvl];
    for (int l = 0; l < lvl; ++l)
      if (hit[l] && lo[l] > cur) end = fminf(end, lo[l]);  // a finer box begins
    // DDA through level `lvl` over [cur, end)
    const float* bb = g.aabb[lvl];
    const float R = (float)g.R;
    int cell[3], stepi[3];
    float tmax[3], tdelta[3];
    {
      // enter slightly inside the stretch so that the start cell is the one being entered
      const float te = cur + (end - cur) * 1e-6f;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const float ext = bb[3 + a] - bb[a];
        const float gpos = ((o[a] + d[a] * te) - bb[a]) / ext * R;
        int c = (int)floorf(gpos);
        c = c < 0 ? 0 : (c > g.R - 1 ? g.R - 1 : c);
        cell[a] = c;
        const float cs = ext / R;
        if (d[a] > 0.f) {
          stepi[a] = 1;
          tmax[a] = ((bb[a] + (float)(c + 1) * cs) - o[a]) / d[a];
          tdelta[a] = cs / d[a];
        } else if (d[a] < 0.f) {
          stepi[a] = -1;
          tmax[a] = ((bb[a] + (float)c * cs) - o[a]) / d[a];
          tdelta[a] = -cs / d[a];
        } else {
          stepi[a] = 0;
          tmax[a] = INFINITY;
          tdelta[a] = INFINITY;
        }
      }
    }
    float ta = cur;
    while (ta < end && n < max_samples) {
      const float tcross = fminf(tmax[0], fminf(tmax[1], tmax[2]));
      const float tb = fminf(tcross, end);
      const int cid = (cell[0] * g.R + cell[1]) * g.R + cell[2];
      if (g.bin[(size_t)lvl * g.R * g.R * g.R + cid]) {
        while (n < max_samples) {
          float t0, t1;
          if (continuous) {
            const float dt = occ_dt(t_last, cone, step);
            if (t_last + dt * 0.5f >= tb) break;
            t0 = t_last;
            t1 = t_last + dt;
          } else {
            // a run starts at the cell's entry - or where the previous sample ended, if that one overshot a short empty
            // stretch: the samples of a ray never overlap, so its boundary sequence is non-decreasing (k_seg_union relies on it)
            const float ts = fmaxf(ta, t_last);
            const float dt = occ_dt(ts, cone, step);
            if (ts + dt * 0.5f >= tb) break;
            t0 = ts;
            t1 = ts + dt;
          }
          emit(n, t0, t1, lvl, cid);
          ++n;
          t_last = t1;
          continuous = true;
        }
      } else {
        continuous = false;
      }
      if (!(tcross < end)) break;
      // step into the next cell (ties advance one axis at a time, x before y before z)
      const int ax = (tmax[0] <= tmax[1] && tmax[0] <= tmax[2]) ? 0 : (tmax[1] <= tmax[2] ? 1 : 2);
      cell[ax] += stepi[ax];
      tmax[ax] += tdelta[ax];
      ta = tcross;
      if (cell[ax] < 0 || cell[ax] >= g.R) break;  // left the box through rounding
    }
    cur = end;
  }
  return n;
}

__global__ void __launch_bounds__(128) k_occ_march(OccGrid g, const float* __restrict__ rays, const float* __restrict__ near_jit, int64_t N,
                                                   float step, float cone, int max_samples, const int64_t* __restrict__ offsets,
                                                   int32_t* __restrict__ counts, int32_t* __restrict__ ray_idx, float* __restrict__ t0s,
                                                   float* __restrict__ t1s, int32_t* __restrict__ cells) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= N) return;
  const float4 a = __ldg(reinterpret_cast<const float4*>(rays + r * 8)), b = __ldg(reinterpret_cast<const float4*>(rays + r * 8) + 1);
  const float o[3] = {a.x, a.y, a.z}, d[3] = {a.w, b.x, b.y};
  float near = b.z, far = b.w;
  if (near_jit) near = near + __ldg(near_jit + r) * step;  // stratified: near += U[0,1) * step
  if (!(far > near)) {  // also NaN / inf rays
    if (counts) counts[r] = 0;
    return;
  }
  if (offsets == nullptr) {
    counts[r] = march_ray(g, o, d, near, far, step, cone, max_samples, [](int, float, float, int, int) {});
    return;
  }
  const int64_t base = offsets[r];
  march_ray(g, o, d, near, far, step, cone, max_samples, [&](int i, float t0, float t1, int lvl, int cid) {
    ray_idx[base + i] = (int32_t)r;
    t0s[base + i] = t0;
    t1s[base + i] = t1;
    if (cells) cells[base + i] = lvl * g.R * g.R * g.R + cid;
  });
}

// nerfacc.render_visibility_from_density over packed segments: alpha = 1 - exp(-sigma * (t1 - t0)), T = exclusive product of
// (1 - alpha) along the ray; visible = T >= early_stop_eps and alpha >= alpha_thre.
__global__ void __launch_bounds__(128) k_occ_visible(const int64_t* __restrict__ offsets, const int32_t* __restrict__ counts, int64_t N,
                                                     const float* __restrict__ t0s, const float* __restrict__ t1s, const float* __restrict__ sigma,
                                                     float early_stop_eps, float alpha_thre, uint8_t* __restrict__ vis) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= N) return;
  const int64_t base = offsets[r];
  const int n = counts[r];
  float T = 1.f;
  for (int i = 0; i < n; ++i) {
    const float alpha = 1.f - expf(-sigma[base + i] * (t1s[base + i] - t0s[base + i]));
    vis[base + i] = (T >= early_stop_eps && alpha >= alpha_thre) ? 1 : 0;
    T = T * (1.f - alpha);
  }
}

// Union of interval boundaries per ray (ray_rendering.py:196-256): the sorted, de-duplicated set of every expert's t0 / t1 on
// the ray; consecutive boundaries form the merged segments (gaps between disjoint segments become segments too, as in the
// reference).  Each expert's list is ordered along the ray, so its boundary sequence t0_0, t1_0, t0_1, ... is non-decreasing:
// a K-way merge of K cursors.
struct UnionLists {
  const int64_t* off[ACN_MAX_EXPERTS];  // (N,) first segment of ray r in list k
  const int32_t* cnt[ACN_MAX_EXPERTS];  // (N,) segments of ray r in list k (0 = the ray was not marched by expert k)
  const float* t0[ACN_MAX_EXPERTS];
  const float* t1[ACN_MAX_EXPERTS];
  int K;
};

__global__ void __launch_bounds__(128) k_seg_union(UnionLists u, int64_t N, const int64_t* __restrict__ out_off, int32_t* __restrict__ out_cnt,
                                                   int32_t* __restrict__ ray_idx, float* __restrict__ t0s, float* __restrict__ t1s) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= N) return;
  int pos[ACN_MAX_EXPERTS], len[ACN_MAX_EXPERTS];
  int64_t base[ACN_MAX_EXPERTS];
#pragma unroll
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
    pos[k] = 0;
    len[k] = k < u.K ? 2 * u.cnt[k][r] : 0;
    base[k] = k < u.K ? u.off[k][r] : 0;
  }
  const int64_t ob = out_off ? out_off[r] : 0;
  int nb = 0;  // distinct boundaries so far
  float prev = 0.f;
  while (true) {
    float best = INFINITY;
    int kb = -1;
#pragma unroll
    for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
      if (k < u.K && pos[k] < len[k]) {
        const int64_t s = base[k] + (pos[k] >> 1);
        const float v = (pos[k] & 1) ? u.t1[k][s] : u.t0[k][s];
        if (kb < 0 || v < best) {
          best = v;
          kb = k;
        }
      }
    }
    if (kb < 0) break;
    ++pos[kb];
    if (nb > 0 && best == prev) continue;  // torch.unique
    if (nb > 0 && out_off) {
      ray_idx[ob + nb - 1] = (int32_t)r;
      t0s[ob + nb - 1] = prev;
      t1s[ob + nb - 1] = best;
    }
    prev = best;
    ++nb;
  }
  if (!out_off) out_cnt[r] = nb >= 2 ? nb - 1 : 0;
}

// nerfacc.render_weight_from_density + accumulate_along_rays over packed segments (ray_rendering.py:477-487, 543-554):
// w_i = alpha_i * prod_{j<i} (1 - alpha_j); rgb = sum w c, depth = sum w t_mid, acc = sum w.  No background here.
__global__ void __launch_bounds__(128) k_packed_fwd(const int64_t* __restrict__ offsets, const int32_t* __restrict__ counts, int64_t N,
                                                    const float* __restrict__ t0s, const float* __restrict__ t1s, const float* __restrict__ sigma,
                                                    const float* __restrict__ rgb, float* __restrict__ w, float* __restrict__ rgb_map,
                                                    float* __restrict__ depth, float* __restrict__ acc) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= N) return;
  const int64_t base = offsets[r];
  const int n = counts[r];
  float T = 1.f, cr = 0.f, cg = 0.f, cb = 0.f, dp = 0.f, ac = 0.f;
  for (int i = 0; i < n; ++i) {
    const int64_t s = base + i;
    const float a = t0s[s], b = t1s[s];
    const float alpha = 1.f - expf(-sigma[s] * (b - a));
    const float wi = alpha * T;
    w[s] = wi;
    cr += wi * rgb[s * 3];
    cg += wi * rgb[s * 3 + 1];
    cb += wi * rgb[s * 3 + 2];
    dp += wi * (0.5f * (a + b));
    ac += wi;
    T = T * (1.f - alpha);
  }
  rgb_map[r * 3] = cr;
  rgb_map[r * 3 + 1] = cg;
  rgb_map[r * 3 + 2] = cb;
  depth[r] = dp;
  acc[r] = ac;
}

// d(loss)/d(sigma_i), d/d(rgb_i) from d_rgb_map (N,3), d_depth (N,), d_acc (N,), d_w (M,) (any may be NULL).
// With q_i = d_rgb.c_i + d_depth t_i + d_acc + d_w_i (the gradient reaching w_i):
//   d/d sigma_i = delta_i * ( (1 - alpha_i) T_i q_i  -  sum_{j>i} w_j q_j )       [d alpha_i / d sigma_i = delta_i (1 - alpha_i)]
__global__ void __launch_bounds__(128) k_packed_bwd(const int64_t* __restrict__ offsets, const int32_t* __restrict__ counts, int64_t N,
                                                    const float* __restrict__ t0s, const float* __restrict__ t1s, const float* __restrict__ sigma,
                                                    const float* __restrict__ rgb, const float* __restrict__ w, const float* __restrict__ d_rgb,
                                                    const float* __restrict__ d_depth, const float* __restrict__ d_acc, const float* __restrict__ d_w,
                                                    float* __restrict__ g_sigma, float* __restrict__ g_rgb) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= N) return;
  const int64_t base = offsets[r];
  const int n = counts[r];
  const float gr = d_rgb ? d_rgb[r * 3] : 0.f, gg = d_rgb ? d_rgb[r * 3 + 1] : 0.f, gb = d_rgb ? d_rgb[r * 3 + 2] : 0.f;
  const float gd = d_depth ? d_depth[r] : 0.f, ga = d_acc ? d_acc[r] : 0.f;
  // T_i = exp(-sum_{j<i} sigma_j delta_j): the optical depth is summed forward once and peeled off in the reverse loop
  // (w_i / alpha_i would be 0/0 exactly where the gradient matters most: samples whose density is still ~0)
  float S = 0.f;
  for (int i = 0; i < n; ++i) S += sigma[base + i] * (t1s[base + i] - t0s[base + i]);
  float suffix = 0.f;  // sum_{j>i} w_j q_j
  for (int i = n - 1; i >= 0; --i) {
    const int64_t s = base + i;
    const float a = t0s[s], b = t1s[s], delta = b - a;
    const float wi = w[s];
    const float q = gr * rgb[s * 3] + gg * rgb[s * 3 + 1] + gb * rgb[s * 3 + 2] + gd * (0.5f * (a + b)) + ga + (d_w ? d_w[s] : 0.f);
    const float od = sigma[s] * delta;
    S -= od;                             // optical depth in front of sample i
    const float T_next = expf(-(S + od));  // (1 - alpha_i) T_i
    g_sigma[s] = delta * (T_next * q - suffix);
    g_rgb[s * 3] = wi * gr;
    g_rgb[s * 3 + 1] = wi * gg;
    g_rgb[s * 3 + 2] = wi * gb;
    suffix += wi * q;
  }
}

static int fill_grid(OccGrid& g, const uint8_t* binaries, const float* aabbs_host, int levels, int R, const char* who) {
  ACN_REQUIRE(binaries && aabbs_host && levels >= 1 && levels <= ACN_OCC_MAX_LEVELS && R >= 1 && R <= 1024, "%s: bad occupancy grid (levels %d, R %d)",
              who, levels, R);
  g.bin = binaries;
  g.levels = levels;
  g.R = R;
  for (int l = 0; l < ACN_OCC_MAX_LEVELS; ++l)
    for (int a = 0; a < 6; ++a) g.aabb[l][a] = l < levels ? aabbs_host[l * 6 + a] : 0.f;
  return ACN_OK;
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_occ_march_count(const float* rays, const float* near_jitter, int64_t N, const uint8_t* binaries, const float* aabbs_host, int levels, int R,
                        float step, float cone_angle, int max_samples, int32_t* counts, void* stream) {
  ACN_REQUIRE(rays && counts && N >= 0 && step > 0.f && cone_angle >= 0.f && max_samples >= 1, "acn_occ_march_count: bad arguments");
  if (N == 0) return ACN_OK;
  OccGrid g;
  int rc = fill_grid(g, binaries, aabbs_host, levels, R, "acn_occ_march_count");
  if (rc) return rc;
  k_occ_march<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(g, rays, near_jitter, N, step, cone_angle, max_samples, nullptr, counts, nullptr,
                                                                         nullptr, nullptr, nullptr);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_occ_march_write(const float* rays, const float* near_jitter, int64_t N, const uint8_t* binaries, const float* aabbs_host, int levels, int R,
                        float step, float cone_angle, int max_samples, const int64_t* offsets, int32_t* ray_idx, float* t0, float* t1, int32_t* cells,
                        void* stream) {
  ACN_REQUIRE(rays && offsets && ray_idx && t0 && t1 && N >= 0 && step > 0.f && cone_angle >= 0.f && max_samples >= 1,
              "acn_occ_march_write: bad arguments");
  if (N == 0) return ACN_OK;
  OccGrid g;
  int rc = fill_grid(g, binaries, aabbs_host, levels, R, "acn_occ_march_write");
  if (rc) return rc;
  k_occ_march<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(g, rays, near_jitter, N, step, cone_angle, max_samples, offsets, nullptr, ray_idx,
                                                                         t0, t1, cells);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_occ_visibility(const int64_t* offsets, const int32_t* counts, int64_t N, const float* t0, const float* t1, const float* sigma,
                       float early_stop_eps, float alpha_thre, uint8_t* visible, void* stream) {
  ACN_REQUIRE(offsets && counts && t0 && t1 && sigma && visible && N >= 0, "acn_occ_visibility: bad arguments");
  if (N == 0) return ACN_OK;
  k_occ_visible<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(offsets, counts, N, t0, t1, sigma, early_stop_eps, alpha_thre, visible);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_seg_union(const AcnSegLists* lists, int64_t N, const int64_t* out_offsets, int32_t* out_counts, int32_t* ray_idx, float* t0, float* t1,
                  void* stream) {
  ACN_REQUIRE(lists && lists->K >= 1 && lists->K <= ACN_MAX_EXPERTS && N >= 0, "acn_seg_union: 1..%d lists", ACN_MAX_EXPERTS);
  ACN_REQUIRE(out_offsets ? (ray_idx && t0 && t1) : (out_counts != nullptr), "acn_seg_union: count pass needs out_counts, write pass needs the outputs");
  if (N == 0) return ACN_OK;
  UnionLists u;
  u.K = lists->K;
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k) {
    const bool on = k < lists->K;
    u.off[k] = on ? lists->offsets[k] : nullptr;
    u.cnt[k] = on ? lists->counts[k] : nullptr;
    u.t0[k] = on ? lists->t0[k] : nullptr;
    u.t1[k] = on ? lists->t1[k] : nullptr;
    ACN_REQUIRE(!on || (u.off[k] && u.cnt[k] && u.t0[k] && u.t1[k]), "acn_seg_union: list %d is incomplete", k);
  }
  k_seg_union<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(u, N, out_offsets, out_counts, ray_idx, t0, t1);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_packed_render_fwd(const int64_t* offsets, const int32_t* counts, int64_t N, const float* t0, const float* t1, const float* sigma,
                          const float* rgb, float* weights, float* rgb_map, float* depth, float* acc, void* stream) {
  ACN_REQUIRE(offsets && counts && t0 && t1 && sigma && rgb && weights && rgb_map && depth && acc && N >= 0, "acn_packed_render_fwd: bad arguments");
  if (N == 0) return ACN_OK;
  k_packed_fwd<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(offsets, counts, N, t0, t1, sigma, rgb, weights, rgb_map, depth, acc);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_packed_render_bwd(const int64_t* offsets, const int32_t* counts, int64_t N, const float* t0, const float* t1, const float* sigma,
                          const float* rgb, const float* weights, const float* d_rgb_map, const float* d_depth, const float* d_acc,
                          const float* d_weights, float* g_sigma, float* g_rgb, void* stream) {
  ACN_REQUIRE(offsets && counts && t0 && t1 && sigma && rgb && weights && g_sigma && g_rgb && N >= 0, "acn_packed_render_bwd: bad arguments");
  if (N == 0) return ACN_OK;
  k_packed_bwd<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(offsets, counts, N, t0, t1, sigma, rgb, weights, d_rgb_map, d_depth, d_acc,
                                                                          d_weights, g_sigma, g_rgb);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
