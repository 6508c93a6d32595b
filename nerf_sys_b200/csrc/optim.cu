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


// ---- multi-tensor variants: one launch for the whole parameter set -------------------------------
struct MultiDesc {
  AcnAdamTensor t[ACN_ADAM_MAX_TENSORS];
  int blk0[ACN_ADAM_MAX_TENSORS + 1];  // first block of each tensor
  int count;
  unsigned long long vec;  // bit i: every pointer of tensor i is 16-byte aligned (float4 path); else scalar
};

__device__ __forceinline__ int find_tensor(const MultiDesc& d, int b) {
  int t = 0;
  while (t + 1 < d.count && b >= d.blk0[t + 1]) ++t;
  return t;
}

// AcnAdamTensor.flags: low 8 bits = liveness slot + 1 (0 = unconditional), bit 8 = write zeros over the gradient once read
__device__ __forceinline__ bool tensor_live(const AcnAdamTensor& a, const int32_t* __restrict__ live) {
  const int slot = a.flags & 0xff;
  return slot == 0 || live == nullptr || __ldg(live + slot - 1) != 0;
}

__global__ void __launch_bounds__(256) k_sumsq_multi(const __grid_constant__ MultiDesc d, const int32_t* __restrict__ live, float* __restrict__ acc) {
  __shared__ float red[8];
  const int t = find_tensor(d, blockIdx.x);
  if (!tensor_live(d.t[t], live)) return;  // block-uniform: a tensor without a gradient this step (torch: p.grad is None)
  const int nb = d.blk0[t + 1] - d.blk0[t], b = blockIdx.x - d.blk0[t];
  const float* __restrict__ x = d.t[t].grad;
  const int64_t n = d.t[t].n, n4 = ((d.vec >> t) & 1ull) ? (n >> 2) : 0;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  float s = 0.f;
  for (int64_t i = b * 256 + threadIdx.x; i < n4; i += (int64_t)nb * 256) {
    float4 v = __ldg(x4 + i);
    s = fmaf(v.x, v.x, s);
    s = fmaf(v.y, v.y, s);
    s = fmaf(v.z, v.z, s);
    s = fmaf(v.w, v.w, s);
  }
  for (int64_t i = (n4 << 2) + b * 256 + threadIdx.x; i < n; i += (int64_t)nb * 256) s = fmaf(x[i], x[i], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int i = 0; i < 8; ++i) tot += red[i];
    atomicAdd(acc, tot);
  }
}

// coef[0] = gradient multiplier = (1/loss scale) * clip coefficient; coef[1] != 0 = skip the step (GradScaler found an
// inf / NaN, or the gradient norm itself is not finite).  The norm is taken of the UNSCALED gradients, as the reference
// unscales before it clips (runtime_adapt.py:261-268): norm = sqrt(sumsq) * |grad_scale|.
// steps[step_index] += 1 for every listed live tensor unless the step is skipped.
__global__ void k_adam_prep(const __grid_constant__ MultiDesc d, int32_t* __restrict__ steps, const float* __restrict__ sumsq, float max_norm,
                            const float* __restrict__ grad_scale, const float* __restrict__ found_inf, const int32_t* __restrict__ live,
                            float* __restrict__ coef, int write_coef) {
  __shared__ float s_skip;
  int i = threadIdx.x;
  if (i == 0) {
    if (write_coef) {
      const float gs = grad_scale ? *grad_scale : 1.f;
      float c = gs, skip = 0.f;
      if (sumsq && max_norm > 0.f) {
        const float norm = sqrtf(*sumsq) * fabsf(gs);
        if (!(norm <= 3.0e38f)) skip = 1.f;  // inf or NaN
        c = gs * fminf(1.f, max_norm / (norm + 1e-6f));  // torch.nn.utils.clip_grad_norm_
      }
      if (found_inf && *found_inf != 0.f) skip = 1.f;
      if (!(fabsf(gs) <= 3.0e38f)) skip = 1.f;
      coef[0] = c;
      coef[1] = skip;
    }
    s_skip = coef[1];
  }
  __syncthreads();
  if (i < d.count && s_skip == 0.f && tensor_live(d.t[i], live)) steps[d.t[i].step_index] += 1;
}

__global__ void __launch_bounds__(256) k_adam_multi(const __grid_constant__ MultiDesc d, float beta1, float beta2, float eps,
                                                    const int32_t* __restrict__ steps, const float* __restrict__ coef,
                                                    const int32_t* __restrict__ live) {
  __shared__ AdamK sk;
  const int t = find_tensor(d, blockIdx.x);
  const AcnAdamTensor& a = d.t[t];
  if (!tensor_live(a, live)) return;  // no gradient this step: parameter, moments and step count untouched (torch skips p.grad is None)
  const bool skip = coef && __ldg(coef + 1) != 0.f;
  const bool zero = (a.flags & 0x100) != 0;
  if (skip && !zero) return;
  if (threadIdx.x == 0) {
    double step = (double)steps[a.step_index];
    double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    sk.lr_step = (float)((double)a.lr / bc1);
    sk.inv_bc2s = (float)(1.0 / sqrt(bc2));
    sk.beta1 = beta1;
    sk.beta2 = beta2;
    sk.eps = eps;
    sk.wd = a.weight_decay;
  }
  __syncthreads();
  const AdamK k = sk;
  const float gs = coef ? __ldg(coef) : 1.f;
  const int nb = d.blk0[t + 1] - d.blk0[t], b = blockIdx.x - d.blk0[t];
  const int64_t n = a.n, n4 = ((d.vec >> t) & 1ull) ? (n >> 2) : 0;
  float4* p4 = reinterpret_cast<float4*>(a.param);
  float4* g4 = reinterpret_cast<float4*>(const_cast<float*>(a.grad));
  float4* m4 = reinterpret_cast<float4*>(a.exp_avg);
  float4* v4 = reinterpret_cast<float4*>(a.exp_avg_sq);
  if (skip) {  // a skipped step still hands the persistent gradient buffer back clean
    for (int64_t i = b * 256 + threadIdx.x; i < n4; i += (int64_t)nb * 256) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (n4 << 2) + b * 256 + threadIdx.x; i < n; i += (int64_t)nb * 256) const_cast<float*>(a.grad)[i] = 0.f;
    return;
  }
  for (int64_t i = b * 256 + threadIdx.x; i < n4; i += (int64_t)nb * 256) {
    float4 pa = p4[i], gb = g4[i], mc = m4[i], vd = v4[i];
    adam_one(pa.x, gb.x, mc.x, vd.x, k, gs);
    adam_one(pa.y, gb.y, mc.y, vd.y, k, gs);
    adam_one(pa.z, gb.z, mc.z, vd.z, k, gs);
    adam_one(pa.w, gb.w, mc.w, vd.w, k, gs);
    p4[i] = pa;
    m4[i] = mc;
    v4[i] = vd;
    // Zero the gradient it just read (replaces the next step's 128 MiB memset per table).  The store must not be issued while
    // the line's own fill is still in flight: a plain `g4[i] = 0` right behind the load runs the kernel at 2.1 TB/s instead of
    // 6.4 TB/s (tools/adam_bench.py, profiles/r2_adam_zero_store.txt).  Naming loaded values as operands holds it back until
    // the data has arrived.
    if (zero)
      asm volatile("{ .reg .f32 z; mov.f32 z, 0f00000000; st.global.v4.f32 [%0], {z, z, z, z}; }" ::"l"(g4 + i), "f"(pa.x), "f"(vd.w) : "memory");
  }
  for (int64_t i = (n4 << 2) + b * 256 + threadIdx.x; i < n; i += (int64_t)nb * 256) {
    float g = a.grad[i];
    if (zero) const_cast<float*>(a.grad)[i] = 0.f;
    adam_one(a.param[i], g, a.exp_avg[i], a.exp_avg_sq[i], k, gs);
  }
}

static int fill_multi(MultiDesc& d, const AcnAdamTensor* tensors, int count, const char* who) {
  ACN_REQUIRE(tensors && count >= 1 && count <= ACN_ADAM_MAX_TENSORS, "%s: 1..%d tensors per call", who, ACN_ADAM_MAX_TENSORS);
  int64_t per_wave = (int64_t)num_sms_cached() * 8;
  int blk = 0;
  d.vec = 0;
  for (int i = 0; i < count; ++i) {
    d.t[i] = tensors[i];
    ACN_REQUIRE(tensors[i].grad && tensors[i].n >= 0, "%s: tensor %d has no gradient pointer", who, i);
    // a gradient that is a view into a flat buffer (the PE-MLP's weight-gradient slab) may sit at any 4-byte offset
    ACN_REQUIRE((((uintptr_t)tensors[i].param | (uintptr_t)tensors[i].grad | (uintptr_t)tensors[i].exp_avg | (uintptr_t)tensors[i].exp_avg_sq) & 3) == 0,
                "%s: tensor %d is not 4-byte aligned", who, i);
    if ((((uintptr_t)tensors[i].param | (uintptr_t)tensors[i].grad | (uintptr_t)tensors[i].exp_avg | (uintptr_t)tensors[i].exp_avg_sq) & 15) == 0)
      d.vec |= 1ull << i;
    int64_t nb = (tensors[i].n / 4 + 1023) / 1024;  // >= 4 float4 per thread before adding blocks
    if (nb < 1) nb = 1;
    if (nb > per_wave) nb = per_wave;
    d.blk0[i] = blk;
    blk += (int)nb;
  }
  for (int i = count; i <= ACN_ADAM_MAX_TENSORS; ++i) d.blk0[i] = blk;
  d.count = count;
  return ACN_OK;
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

int acn_sumsq_multi(const AcnAdamTensor* tensors_host, int count, const int32_t* live_dev, float* acc, void* stream) {
  ACN_REQUIRE(acc, "acn_sumsq_multi: acc is NULL");
  MultiDesc d;
  int rc = fill_multi(d, tensors_host, count, "acn_sumsq_multi");
  if (rc) return rc;
  k_sumsq_multi<<<d.blk0[count], 256, 0, as_stream(stream)>>>(d, live_dev, acc);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_adam_multi(const AcnAdamTensor* tensors_host, int count, float beta1, float beta2, float eps, int32_t* steps_dev,
                   const float* sumsq_dev, float max_norm, const float* grad_scale_dev, const float* found_inf_dev,
                   const int32_t* live_dev, float* coef_dev, int write_coef, void* stream) {
  ACN_REQUIRE(steps_dev && coef_dev, "acn_adam_multi: steps_dev / coef_dev is NULL");
  MultiDesc d;
  int rc = fill_multi(d, tensors_host, count, "acn_adam_multi");
  if (rc) return rc;
  for (int i = 0; i < count; ++i)
    ACN_REQUIRE(tensors_host[i].param && tensors_host[i].exp_avg && tensors_host[i].exp_avg_sq && tensors_host[i].step_index >= 0,
                "acn_adam_multi: tensor %d is incomplete", i);
  cudaStream_t st = as_stream(stream);
  k_adam_prep<<<1, ACN_ADAM_MAX_TENSORS, 0, st>>>(d, steps_dev, sumsq_dev, max_norm, grad_scale_dev, found_inf_dev, live_dev, coef_dev, write_coef);
  ACN_LAUNCH_CHECK();
  k_adam_multi<<<d.blk0[count], 256, 0, st>>>(d, beta1, beta2, eps, steps_dev, coef_dev, live_dev);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
