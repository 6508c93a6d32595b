// a1-a5: ray generation, AABB slab test, near/far clamp, stratified sampling, point build.
// Compiled with -fmad=false: every multiply/add rounds separately, as the reference's
// elementwise torch ops do (nerfs/ray_sampling.py, nerfs/scene_box.py, nerfs/ray_rendering.py:262-319).
// All kernels are HBM-streaming: grid-stride, one element per thread, 32-byte ray rows written as 2x float4.
#include "common.cuh"

namespace acn {

struct Cam {
  float R[9];  // row-major c2w[:3,:3]
  float o[3];
  float bmin[3], bmax[3];
  int has_box;
  float near_c, far_c, max_bound, invalid;
};

__device__ __forceinline__ float nan_min(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fminf(a, b); }
__device__ __forceinline__ float nan_max(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fmaxf(a, b); }

// nerfs/ray_sampling.py:111-136
__device__ __forceinline__ void pixel_dir(int i, int j, float fx, float fy, float cx, float cy, int center, float& x,
                                          float& y, float& z) {
  float fi = (float)i, fj = (float)j;
  if (center) {
    fi += 0.5f;
    fj += 0.5f;
  }
  x = (fi - cx) / fx;
  y = -((fj - cy) / fy);
  z = -1.0f;
  float n = sqrtf(x * x + y * y + z * z);
  n = fmaxf(n, 1e-12f);
  x = x / n;
  y = y / n;
  z = z / n;
}

// scene_box.py:45-107
__device__ __forceinline__ void slab(const Cam& c, float ox, float oy, float oz, float dx, float dy, float dz,
                                     float& tn, float& tf) {
  const float eps = 1e-8f;
  float d[3] = {dx, dy, dz}, o[3] = {ox, oy, oz};
  float lo = -INFINITY, hi = INFINITY;
  bool first = true;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float ds = (fabsf(d[a]) < eps) ? ((d[a] >= 0.f) ? eps : -eps) : d[a];
    float inv = 1.0f / ds;
    float t0 = (c.bmin[a] - o[a]) * inv;
    float t1 = (c.bmax[a] - o[a]) * inv;
    float mn = nan_min(t0, t1), mx = nan_max(t0, t1);
    lo = first ? mn : nan_max(lo, mn);
    hi = first ? mx : nan_min(hi, mx);
    first = false;
  }
  // clamp(min=0, max=max_bound) propagates NaN
  if (lo == lo) lo = fminf(fmaxf(lo, 0.f), c.max_bound);
  if (hi == hi) hi = fminf(fmaxf(hi, 0.f), c.max_bound);
  bool invalid = hi <= lo;
  tn = invalid ? c.invalid : lo;
  tf = invalid ? c.invalid : hi;
}

__device__ __forceinline__ void cam_to_world(const Cam& c, float vx, float vy, float vz, float& dx, float& dy, float& dz) {
  // dirs @ R.T : d_j = sum_k v_k R[j][k]   (ray_sampling.py:10-24; not re-normalised)
  dx = vx * c.R[0] + vy * c.R[1] + vz * c.R[2];
  dy = vx * c.R[3] + vy * c.R[4] + vz * c.R[5];
  dz = vx * c.R[6] + vy * c.R[7] + vz * c.R[8];
}

__device__ __forceinline__ void store_ray(float* rays, int64_t n, const Cam& c, float dx, float dy, float dz) {
  float tn, tf;
  if (c.has_box) {
    slab(c, c.o[0], c.o[1], c.o[2], dx, dy, dz, tn, tf);
  } else {
    tn = c.near_c;
    tf = c.far_c;
  }
  float4* q = reinterpret_cast<float4*>(rays + n * 8);
  q[0] = make_float4(c.o[0], c.o[1], c.o[2], dx);
  q[1] = make_float4(dy, dz, tn, tf);
}

__global__ void __launch_bounds__(256) k_ray_dirs(int H, int W, float fx, float fy, float cx, float cy, int center, float* dirs) {
  int64_t n = (int64_t)H * W;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    int j = (int)(p / W), i = (int)(p % W);
    float x, y, z;
    pixel_dir(i, j, fx, fy, cx, cy, center, x, y, z);
    dirs[p * 3 + 0] = x;
    dirs[p * 3 + 1] = y;
    dirs[p * 3 + 2] = z;
  }
}

__global__ void __launch_bounds__(256) k_get_rays(const float* __restrict__ dirs, int64_t n, Cam c, float* __restrict__ rays) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    float dx, dy, dz;
    cam_to_world(c, dirs[p * 3], dirs[p * 3 + 1], dirs[p * 3 + 2], dx, dy, dz);
    store_ray(rays, p, c, dx, dy, dz);
  }
}

__global__ void __launch_bounds__(256) k_raygen(int W, int row0, int64_t n, float fx, float fy, float cx, float cy, int center,
                                                Cam c, float* __restrict__ rays) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    int j = row0 + (int)(p / W), i = (int)(p % W);
    float vx, vy, vz, dx, dy, dz;
    pixel_dir(i, j, fx, fy, cx, cy, center, vx, vy, vz);
    cam_to_world(c, vx, vy, vz, dx, dy, dz);
    store_ray(rays, p, c, dx, dy, dz);
  }
}

__global__ void __launch_bounds__(256) k_ray_aabb(const float* __restrict__ o, const float* __restrict__ d, int64_t n, Cam c,
                                                  float* __restrict__ tn, float* __restrict__ tf) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    float a, b;
    slab(c, o[p * 3], o[p * 3 + 1], o[p * 3 + 2], d[p * 3], d[p * 3 + 1], d[p * 3 + 2], a, b);
    tn[p] = a;
    tf[p] = b;
  }
}

// nerfs/ray_sampling.py:139-176
__global__ void __launch_bounds__(256) k_clamp(const float* __restrict__ in, int64_t n, int none, int has_near, float nv,
                                               int has_far, float fv, float eps, float invalid, float* __restrict__ out,
                                               uint8_t* __restrict__ valid) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const float4* q = reinterpret_cast<const float4*>(in + p * 8);
    float4 a = q[0], b = q[1];
    float near = b.z, far = b.w;
    if (!none) {
      if (has_near) near = nan_max(near, nv);
      if (has_far) far = nan_min(far, fv);
    }
    bool ok = isfinite(near) && isfinite(far) && (far > near + eps);
    if (!none) {
      b.z = ok ? near : invalid;
      b.w = ok ? far : invalid;
    }
    if (out) {
      float4* w = reinterpret_cast<float4*>(out + p * 8);
      w[0] = a;
      w[1] = b;
    }
    valid[p] = ok ? 1 : 0;
  }
}

// nerfs/ray_rendering.py:262-287.  One thread per (ray, sample); near/far re-read through L1.
__global__ void __launch_bounds__(256) k_strat_t(const float* __restrict__ rays, int64_t N, int S, const float* __restrict__ tau,
                                                 const float* __restrict__ u, float* __restrict__ t) {
  int64_t total = N * S;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = ray_of(p, S);
    int s = (int)(p - r * S);
    float near = __ldg(rays + r * 8 + 6), far = __ldg(rays + r * 8 + 7);
    float ta = __ldg(tau + s);
    float tc = near * (1.0f - ta) + far * ta;
    if (u) {
      float lo = tc, hi = tc;
      if (s > 0) {
        float tb = __ldg(tau + s - 1);
        float tp = near * (1.0f - tb) + far * tb;
        lo = 0.5f * (tp + tc);
      }
      if (s < S - 1) {
        float tb = __ldg(tau + s + 1);
        float tn = near * (1.0f - tb) + far * tb;
        hi = 0.5f * (tc + tn);
      }
      tc = lo + (hi - lo) * u[p];
    }
    t[p] = tc;
  }
}

__global__ void __launch_bounds__(256) k_points(PointSrc src, int64_t P, float* __restrict__ x6) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    float x, y, z, dx, dy, dz;
    load_xyz(src, p, x, y, z);
    load_dir(src, p, dx, dy, dz);
    float2* q = reinterpret_cast<float2*>(x6 + p * 6);
    q[0] = make_float2(x, y);
    q[1] = make_float2(z, dx);
    q[2] = make_float2(dy, dz);
  }
}

static int fill_cam(Cam& c, const float* c2w, const float* aabb, float near_c, float far_c, float max_bound, float invalid) {
  if (c2w) {
    for (int r = 0; r < 3; ++r) {
      for (int k = 0; k < 3; ++k) c.R[r * 3 + k] = c2w[r * 4 + k];
      c.o[r] = c2w[r * 4 + 3];
    }
  } else {
    for (int k = 0; k < 9; ++k) c.R[k] = (k % 4 == 0) ? 1.f : 0.f;
    c.o[0] = c.o[1] = c.o[2] = 0.f;
  }
  c.has_box = aabb != nullptr;
  for (int k = 0; k < 3; ++k) {
    c.bmin[k] = aabb ? aabb[k] : 0.f;
    c.bmax[k] = aabb ? aabb[3 + k] : 0.f;
  }
  c.near_c = near_c;
  c.far_c = far_c;
  c.max_bound = max_bound;
  c.invalid = invalid;
  return 0;
}


// Proposal (hierarchical) resampling, the inverse-CDF step of NeRF's coarse-to-fine sampler.  ADDITIVE: the reference has no
// proposal sampler (its importance sampler is the occupancy grid); north_star names one.  Per ray: bins = midpoints of the S
// coarse distances (S-1 edges), weights = w[1..S-2] + 1e-5, cdf = [0, cumsum(pdf)]; each u in [0,1) lands in the bin where the
// cdf passes it (searchsorted, right side) and is placed linearly inside it; then the S2 new distances (u ascending -> ascending)
// are merged with the S coarse ones into one ascending list of S + S2.  One thread per ray; the cdf lives in local memory.
constexpr int PDF_MAX_S = 256;
__global__ void __launch_bounds__(128) k_sample_pdf(const float* __restrict__ t, const float* __restrict__ w, const float* __restrict__ u, int64_t N,
                                                    int S, int S2, float* __restrict__ t_new, float* __restrict__ t_merged) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= N) return;
  const float* tr = t + r * S;
  const float* wr = w + r * S;
  float cdf[PDF_MAX_S];
  const int nb = S - 1;  // bin edges; nb - 1 intervals
  // running sums in double, rounded to fp32 per entry (what torch.cumsum does for fp32 on the CPU): a sample in a bin of mass m
  // moves by (cdf error / m) of the bin width, and peaked weights leave bins with m ~ 1e-5
  double dsum = 0.0;
  for (int i = 0; i < nb - 1; ++i) dsum += (double)(wr[i + 1] + 1e-5f);
  const float sum = (float)dsum;
  cdf[0] = 0.f;
  double acc = 0.0;
  for (int i = 0; i < nb - 1; ++i) {
    acc += (double)((wr[i + 1] + 1e-5f) / sum);
    cdf[i + 1] = (float)acc;
  }
  int ci = 0;      // merge cursor into the coarse list
  int64_t mo = r * (int64_t)(S + S2);
  int mcount = 0;
  for (int j = 0; j < S2; ++j) {
    const float uj = u[r * (int64_t)S2 + j];
    // searchsorted(cdf, u, right=True): first index with cdf[idx] > u
    int lo = 0, hi = nb;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (cdf[mid] <= uj) lo = mid + 1; else hi = mid;
    }
    const int below = lo - 1 < 0 ? 0 : lo - 1, above = lo > nb - 1 ? nb - 1 : lo;
    float denom = cdf[above] - cdf[below];
    if (denom < 1e-5f) denom = 1.f;
    const float frac = (uj - cdf[below]) / denom;
    const float b0 = 0.5f * (tr[below] + tr[below + 1]), b1 = 0.5f * (tr[above] + tr[above + 1]);
    const float tn = b0 + frac * (b1 - b0);
    t_new[r * (int64_t)S2 + j] = tn;
    if (t_merged) {
      while (ci < S && tr[ci] <= tn) t_merged[mo + mcount++] = tr[ci++];
      t_merged[mo + mcount++] = tn;
    }
  }
  if (t_merged)
    while (ci < S) t_merged[mo + mcount++] = tr[ci++];
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_ray_directions(int H, int W, float fx, float fy, float cx, float cy, int center_pixels, float* dirs, void* stream) {
  ACN_REQUIRE(H > 0 && W > 0 && dirs, "acn_ray_directions: bad arguments");
  int64_t n = (int64_t)H * W;
  k_ray_dirs<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(H, W, fx, fy, cx, cy, center_pixels, dirs);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_get_rays(const float* dirs, int64_t n, const float* c2w_host, const float* aabb_host, float near_c, float far_c,
                 float aabb_max_bound, float aabb_invalid_value, float* rays, void* stream) {
  ACN_REQUIRE(dirs && rays && c2w_host && n >= 0, "acn_get_rays: bad arguments");
  if (n == 0) return ACN_OK;
  Cam c;
  fill_cam(c, c2w_host, aabb_host, near_c, far_c, aabb_max_bound, aabb_invalid_value);
  k_get_rays<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(dirs, n, c, rays);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_raygen(int H, int W, int row0, int rows, float fx, float fy, float cx, float cy, int center_pixels,
               const float* c2w_host, const float* aabb_host, float near_c, float far_c, float* rays, void* stream) {
  ACN_REQUIRE(H > 0 && W > 0 && row0 >= 0 && rows >= 0 && row0 + rows <= H && rays && c2w_host, "acn_raygen: bad arguments");
  int64_t n = (int64_t)rows * W;
  if (n == 0) return ACN_OK;
  Cam c;
  fill_cam(c, c2w_host, aabb_host, near_c, far_c, 1e10f, 1e10f);
  k_raygen<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(W, row0, n, fx, fy, cx, cy, center_pixels, c, rays);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_ray_aabb(const float* o, const float* d, int64_t n, const float* aabb_host, float max_bound, float invalid_value,
                 float* near_out, float* far_out, void* stream) {
  ACN_REQUIRE(o && d && aabb_host && near_out && far_out && n >= 0, "acn_ray_aabb: bad arguments");
  if (n == 0) return ACN_OK;
  Cam c;
  fill_cam(c, nullptr, aabb_host, 0.f, 0.f, max_bound, invalid_value);
  k_ray_aabb<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(o, d, n, c, near_out, far_out);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_clamp_near_far(const float* rays_in, int64_t n, int override_none, int has_near, float near_v, int has_far,
                       float far_v, float eps, float invalid_value, float* rays_out, uint8_t* valid, void* stream) {
  ACN_REQUIRE(rays_in && valid && n >= 0, "acn_clamp_near_far: bad arguments");
  ACN_REQUIRE(override_none || rays_out, "acn_clamp_near_far: rays_out required with an override");
  if (n == 0) return ACN_OK;
  k_clamp<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(rays_in, n, override_none, has_near, near_v, has_far,
                                                                              far_v, eps, invalid_value,
                                                                              override_none ? nullptr : rays_out, valid);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_stratified_t(const float* rays, int64_t N, int S, const float* tau, const float* u, float* t, void* stream) {
  ACN_REQUIRE(rays && tau && t && N >= 0 && S >= 1, "acn_stratified_t: bad arguments");
  if (N == 0) return ACN_OK;
  int64_t total = N * S;
  k_strat_t<<<persistent_grid((total + 255) / 256, 8), 256, 0, as_stream(stream)>>>(rays, N, S, tau, u, t);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_sample_pdf(const float* t, const float* weights, const float* u, int64_t N, int S, int S2, float* t_new, float* t_merged, void* stream) {
  ACN_REQUIRE(t && weights && u && t_new && N >= 0 && S2 >= 1, "acn_sample_pdf: bad arguments");
  ACN_REQUIRE(S >= 3 && S <= PDF_MAX_S, "acn_sample_pdf: 3 <= S <= %d coarse samples", PDF_MAX_S);
  if (N == 0) return ACN_OK;
  k_sample_pdf<<<(unsigned)((N + 127) / 128), 128, 0, as_stream(stream)>>>(t, weights, u, N, S, S2, t_new, t_merged);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_sample_points(const float* rays, const float* t, int64_t N, int S, float* x6, void* stream) {
  ACN_REQUIRE(rays && t && x6 && N >= 0 && S >= 1, "acn_sample_points: bad arguments");
  if (N == 0) return ACN_OK;
  PointSrc src{nullptr, rays, t, S};
  int64_t P = N * S;
  k_points<<<persistent_grid((P + 255) / 256, 8), 256, 0, as_stream(stream)>>>(src, P, x6);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
