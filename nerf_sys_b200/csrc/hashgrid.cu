// a9-a11 stand-alone encoder entry points (the HashGridEncoder / SHEncoder /
// FrequencyEncoder seams, models/encodings.py).  The fused render path uses the
// same device functions from hashgrid.cuh inside ngp.cu.
//
// Launch shape: grid.y = level, so one CTA touches one level's 8 MiB slab of the
// table (L2 locality); threads = consecutive points (consecutive samples along a
// ray in the render path => coarse-level corners coalesce / hit L1).
// Roofline: L2/HBM gather-bound, 1024 B of table rows per point (L*8*F*4) + 12 B in + 128 B out.
#include "hashgrid.cuh"

namespace acn {

__global__ void __launch_bounds__(256) k_hash_indices(const float* __restrict__ x01, int64_t P, GridParams g, int32_t* __restrict__ idx) {
  int l = blockIdx.y;
  uint32_t mask = (1u << g.log2_T) - 1u, base = (uint32_t)l << g.log2_T;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    Cell c = make_cell(x01[p * 3], x01[p * 3 + 1], x01[p * 3 + 2], g.res[l], 0);
    uint32_t rows[8];
    corner_rows(c, mask, base, rows);
    int4* o = reinterpret_cast<int4*>(idx + (p * g.L + l) * 8);
    o[0] = make_int4((int)rows[0], (int)rows[1], (int)rows[2], (int)rows[3]);
    o[1] = make_int4((int)rows[4], (int)rows[5], (int)rows[6], (int)rows[7]);
  }
}

__global__ void __launch_bounds__(256) k_hash_fwd(const float* __restrict__ x01, int64_t P, const float2* __restrict__ table,
                                                  GridParams g, float* __restrict__ y) {
  int l = blockIdx.y;
  uint32_t mask = (1u << g.log2_T) - 1u, base = (uint32_t)l << g.log2_T;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    Cell c = make_cell(__ldg(x01 + p * 3), __ldg(x01 + p * 3 + 1), __ldg(x01 + p * 3 + 2), g.res[l], g.interp);
    float2 v = level_encode(table, c, mask, base, g.interp);
    *reinterpret_cast<float2*>(y + p * (2 * g.L) + 2 * l) = v;
  }
}

__global__ void __launch_bounds__(256) k_hash_bwd(const float* __restrict__ x01, int64_t P, const float* __restrict__ dy, GridParams g,
                                                  float2* __restrict__ gtable) {
  int l = blockIdx.y;
  uint32_t mask = (1u << g.log2_T) - 1u, base = (uint32_t)l << g.log2_T;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    Cell c = make_cell(__ldg(x01 + p * 3), __ldg(x01 + p * 3 + 1), __ldg(x01 + p * 3 + 2), g.res[l], g.interp);
    float2 gv = __ldg(reinterpret_cast<const float2*>(dy + p * (2 * g.L) + 2 * l));
    level_scatter(gtable, c, mask, base, g.interp, gv);
  }
}

__global__ void __launch_bounds__(256) k_sh16(const float* __restrict__ d, int64_t P, float* __restrict__ out) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    float x = d[p * 3], y = d[p * 3 + 1], z = d[p * 3 + 2];
    normalize3(x, y, z, 1e-9f);
    float o[16];
    sh16_eval(x, y, z, o);
    float4* q = reinterpret_cast<float4*>(out + p * 16);
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
  }
}

// encodings.py:437-444: [x, per dim j: cos(x_j 2^k) k<n, sin(x_j 2^k) k<n]
__global__ void __launch_bounds__(256) k_freq(const float* __restrict__ x, int64_t P, int n, int incl, float* __restrict__ out) {
  int width = (incl ? 3 : 0) + 6 * n;
  int64_t total = P * 3 * n;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = i / (3 * n);
    int r = (int)(i - p * 3 * n);
    int j = r / n, k = r - j * n;
    float v = x[p * 3 + j];
    float a = v * exp2f((float)k);
    float* o = out + p * width + (incl ? 3 : 0) + j * 2 * n;
    o[k] = cosf(a);
    o[n + k] = sinf(a);
    if (incl && k == 0) out[p * width + j] = v;
  }
}

}  // namespace acn

using namespace acn;

static int check_grid(const AcnGridDesc* g, const char* who) {
  ACN_REQUIRE(g, "%s: grid descriptor is NULL", who);
  ACN_REQUIRE(g->levels >= 1 && g->levels <= ACN_MAX_LEVELS, "%s: levels must be in [1,%d]", who, ACN_MAX_LEVELS);
  ACN_REQUIRE(g->features == 2, "%s: features_per_level must be 2 (got %d)", who, g->features);
  ACN_REQUIRE(g->log2_T >= 1 && g->log2_T <= 24, "%s: log2_hashmap_size must be in [1,24]", who);
  ACN_REQUIRE(g->interp >= 0 && g->interp <= 2, "%s: unknown interpolation", who);
  ACN_REQUIRE(((int64_t)g->levels << g->log2_T) < ((int64_t)1 << 31), "%s: table too large for 32-bit rows", who);
  return ACN_OK;
}

extern "C" {

int acn_hashgrid_indices(const float* x01, int64_t P, const AcnGridDesc* g, int32_t* idx, void* stream) {
  int rc = check_grid(g, "acn_hashgrid_indices");
  if (rc) return rc;
  ACN_REQUIRE(x01 && idx && P >= 0, "acn_hashgrid_indices: bad arguments");
  if (P == 0) return ACN_OK;
  dim3 grid(persistent_grid((P + 255) / 256, 4), g->levels);
  k_hash_indices<<<grid, 256, 0, as_stream(stream)>>>(x01, P, make_grid(g), idx);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_hashgrid_fwd(const float* x01, int64_t P, const float* table, const AcnGridDesc* g, float* y, void* stream) {
  int rc = check_grid(g, "acn_hashgrid_fwd");
  if (rc) return rc;
  ACN_REQUIRE(x01 && table && y && P >= 0, "acn_hashgrid_fwd: bad arguments");
  if (P == 0) return ACN_OK;
  dim3 grid(persistent_grid((P + 255) / 256, 4), g->levels);
  k_hash_fwd<<<grid, 256, 0, as_stream(stream)>>>(x01, P, reinterpret_cast<const float2*>(table), make_grid(g), y);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_hashgrid_bwd(const float* x01, int64_t P, const float* dy, const AcnGridDesc* g, float* grad_table, void* stream) {
  int rc = check_grid(g, "acn_hashgrid_bwd");
  if (rc) return rc;
  ACN_REQUIRE(x01 && dy && grad_table && P >= 0, "acn_hashgrid_bwd: bad arguments");
  if (P == 0) return ACN_OK;
  dim3 grid(persistent_grid((P + 255) / 256, 4), g->levels);
  k_hash_bwd<<<grid, 256, 0, as_stream(stream)>>>(x01, P, dy, make_grid(g), reinterpret_cast<float2*>(grad_table));
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_sh16(const float* d, int64_t P, float* out, void* stream) {
  ACN_REQUIRE(d && out && P >= 0, "acn_sh16: bad arguments");
  if (P == 0) return ACN_OK;
  k_sh16<<<persistent_grid((P + 255) / 256, 8), 256, 0, as_stream(stream)>>>(d, P, out);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_freq_encode(const float* x, int64_t P, int n_freq, int include_input, float* out, void* stream) {
  ACN_REQUIRE(x && out && P >= 0 && n_freq >= 1 && n_freq <= 24, "acn_freq_encode: bad arguments");
  if (P == 0) return ACN_OK;
  int64_t total = P * 3 * n_freq;
  k_freq<<<persistent_grid((total + 255) / 256, 8), 256, 0, as_stream(stream)>>>(x, P, n_freq, include_input, out);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
