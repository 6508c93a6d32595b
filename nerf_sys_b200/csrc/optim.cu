// (f)2 fused optimiser pieces: torch.optim.Adam math on one flat tensor, the
// fast-weight SGD step of the meta inner loop, and the sum of squares that
// clip_grad_norm_ needs.  Pure HBM streaming: Adam reads 16 B and writes 12 B per
// element (the 128 MiB hash tables dominate), float4-vectorised, grid-stride.
#include <math.h>

#include "common.cuh"

namespace acn {

struct AdamK {
  float lr_step;   // lr / (1 - beta1^t)
  float inv_bc2s;  // 1 / sqrt(1 - beta2^t)
  float beta1, beta2, eps, wd;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamK& k, float gs) {
  g *= gs;
  if (k.wd != 0.f) g = fmaf(k.wd, p, g);
  m = m + (g - m) * (1.f - k.beta1);             // exp_avg.lerp_(grad, 1 - beta1)
  v = v * k.beta2 + (1.f - k.beta2) * g * g;     // mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  float denom = sqrtf(v) * k.inv_bc2s + k.eps;   // (sqrt(v) / sqrt(bc2)).add_(eps)
  p = p - k.lr_step * (m / denom);               // addcdiv_(exp_avg, denom, -step_size)
}

__global__ void __launch_bounds__(256) k_adam(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                              float* __restrict__ v, int64_t n, AdamK k, const float* __restrict__ gscale) {
  float gs = gscale ? __ldg(gscale) : 1.f;
  int64_t n4 = n >> 2;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 a = p4[i], b = g4[i], c = m4[i], d = v4[i];
    adam_one(a.x, b.x, c.x, d.x, k, gs);
    adam_one(a.y, b.y, c.y, d.y, k, gs);
    adam_one(a.z, b.z, c.z, d.z, k, gs);
    adam_one(a.w, b.w, c.w, d.w, k, gs);
    p4[i] = a;
    m4[i] = c;
    v4[i] = d;
  }
  for (int64_t i = (n4 << 2) + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    adam_one(p[i], g[i], m[i], v[i], k, gs);
}

__global__ void __launch_bounds__(256) k_axpy(const float* __restrict__ w, const float* __restrict__ g, float lr, int64_t n,
                                              float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = w[i] - lr * g[i];
}

__global__ void __launch_bounds__(256) k_sumsq(const float* __restrict__ x, int64_t n, float* __restrict__ acc) {
  __shared__ float red[8];
  float s = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s = fmaf(x[i], x[i], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    atomicAdd(acc, t);
  }
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1, float beta2,
                  float eps, float weight_decay, int64_t step, const float* grad_scale_dev, void* stream) {
  ACN_REQUIRE(param && grad && exp_avg && exp_avg_sq && n >= 0 && step >= 1, "acn_adam_step: bad arguments");
  if (n == 0) return ACN_OK;
  ACN_REQUIRE((((uintptr_t)param | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
              "acn_adam_step: tensors must be 16-byte aligned");
  AdamK k;
  double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  k.lr_step = (float)((double)lr / bc1);
  k.inv_bc2s = (float)(1.0 / sqrt(bc2));
  k.beta1 = beta1;
  k.beta2 = beta2;
  k.eps = eps;
  k.wd = weight_decay;
  k_adam<<<persistent_grid((n / 4 + 255) / 256 + 1, 8), 256, 0, as_stream(stream)>>>(param, grad, exp_avg, exp_avg_sq, n, k, grad_scale_dev);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_sgd_axpy(const float* w, const float* g, float lr, int64_t n, float* out, void* stream) {
  ACN_REQUIRE(w && g && out && n >= 0, "acn_sgd_axpy: bad arguments");
  if (n == 0) return ACN_OK;
  k_axpy<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(w, g, lr, n, out);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_sumsq_accum(const float* x, int64_t n, float* acc, void* stream) {
  ACN_REQUIRE(x && acc && n >= 0, "acn_sumsq_accum: bad arguments");
  if (n == 0) return ACN_OK;
  k_sumsq<<<persistent_grid((n + 255) / 256, 4), 256, 0, as_stream(stream)>>>(x, n, acc);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
