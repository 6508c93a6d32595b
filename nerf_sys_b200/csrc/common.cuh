// Shared helpers for libacnerf_sm100a.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/acnerf.h"

namespace acn {

void set_error(const char* fmt, ...);
int num_sms_cached();
int launch_hint(int key);  // thread-local launch-sizing hints (acn_launch_hint)

#define ACN_REQUIRE(cond, ...)          \
  do {                                  \
    if (!(cond)) {                      \
      acn::set_error(__VA_ARGS__);      \
      return ACN_EINVAL;                \
    }                                   \
  } while (0)

#define ACN_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      acn::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return ACN_ECUDA;                                                                  \
    }                                                                                    \
  } while (0)

#define ACN_LAUNCH_CHECK()                                                               \
  do {                                                                                   \
    cudaError_t e__ = cudaGetLastError();                                                \
    if (e__ != cudaSuccess) {                                                            \
      acn::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
      return ACN_ECUDA;                                                                  \
    }                                                                                    \
  } while (0)

// Persistent grid: `per_sm` CTAs on each of the device's SMs (148 on B200),
// never more CTAs than there is work.
static inline int persistent_grid(int64_t work_items, int per_sm) {
  int64_t g = (int64_t)num_sms_cached() * per_sm;
  if (work_items < g) g = work_items;
  if (g < 1) g = 1;
  return (int)g;
}

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- device-side pieces shared by several kernels -------------------------
// A read-only load the compiler must issue HERE.  __ldg() is an invariant load that LLVM may sink to its first use, which
// silently turns a software prefetch (issue now, consume one tile later) into a load-then-stall.
__device__ __forceinline__ float ldg_now(const float* p) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int ldg_now(const int* p) {
  int v;
  asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ldg_now(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
// No-return reductions with an explicit .global state space.  atomicAdd() on a pointer the compiler cannot prove global
// (e.g. one picked from a kernel-parameter array by a runtime index) compiles to the generic ATOM.E form, which waits for
// the round trip through L2: measured 2x slower than REDG for the table-gradient scatter (profiles/r1_v5_red_microbench.txt).
__device__ __forceinline__ void red_add(float* p, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add(float2* p, float2 v) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(__cvta_generic_to_global(p)), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ void red_add(float4* p, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(__cvta_generic_to_global(p)), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

struct PointSrc {
  const float* x6;
  const float* rays;
  const float* t;
  int S;
  const float* tT;  // optional (S, n_rays) copy of t
  uint32_t n_rays;
};

static inline PointSrc make_src(const AcnPoints* p) {
  PointSrc s;
  s.x6 = p->x6;
  s.rays = p->rays;
  s.t = p->t;
  s.S = p->S;
  s.tT = p->x6 ? nullptr : p->t_T;
  s.n_rays = (p->S > 0) ? (uint32_t)(p->P / p->S) : 0u;
  return s;
}

// ray index of point p.  A 64-bit division is a ~500-cycle emulated subroutine on the GPU; point ids fit 32 bits in
// every launch the host layer makes (chunked below 2^31), so the unsigned 32-bit form is taken whenever it is exact.
__device__ __forceinline__ int64_t ray_of(int64_t p, int S) {
  if ((uint64_t)p <= 0xffffffffull) return (int64_t)((uint32_t)p / (uint32_t)S);
  return p / S;
}

// (ray, t) of point p of an implicit point set: ray-major ids (p = ray * S + sample, t[p]) or, when the (S,N) copy of t
// is given, sample-major ids (p = sample * N + ray, tT[p]).
__device__ __forceinline__ int64_t ray_and_t(const PointSrc& s, int64_t p, float& tv) {
  if (s.tT) {
    tv = __ldg(s.tT + p);
    return (int64_t)((uint32_t)p % s.n_rays);  // launches are chunked below 2^31 points
  }
  tv = __ldg(s.t + p);
  return ray_of(p, s.S);
}

// xyz of point p.  Implicit form: o + d*t with a separate multiply and add
// (ray_rendering.py:317) - no FMA contraction, so positions are bit-identical.
__device__ __forceinline__ void load_xyz(const PointSrc& s, int64_t p, float& x, float& y, float& z) {
  if (s.x6) {
    const float* q = s.x6 + p * 6;
    x = __ldg(q);
    y = __ldg(q + 1);
    z = __ldg(q + 2);
  } else {
    float tv;
    int64_t r = ray_and_t(s, p, tv);
    const float4* rr = reinterpret_cast<const float4*>(s.rays + r * 8);
    float4 a = __ldg(rr), b = __ldg(rr + 1);
    x = __fadd_rn(a.x, __fmul_rn(a.w, tv));
    y = __fadd_rn(a.y, __fmul_rn(b.x, tv));
    z = __fadd_rn(a.z, __fmul_rn(b.y, tv));
  }
}

__device__ __forceinline__ void load_dir(const PointSrc& s, int64_t p, float& x, float& y, float& z) {
  if (s.x6) {
    const float* q = s.x6 + p * 6 + 3;
    x = __ldg(q);
    y = __ldg(q + 1);
    z = __ldg(q + 2);
  } else {
    int64_t r = s.tT ? (int64_t)((uint32_t)p % s.n_rays) : ray_of(p, s.S);
    const float4* rr = reinterpret_cast<const float4*>(s.rays + r * 8);
    float4 a = __ldg(rr), b = __ldg(rr + 1);
    x = a.w;
    y = b.x;
    z = b.y;
  }
}

// meta_ngp.py:155-158: clamp((x - min) / extent, 1e-6, 1 - 1e-6), true division.
__device__ __forceinline__ float to_unit(float x, float mn, float ext) {
  float v = __fdiv_rn(__fsub_rn(x, mn), ext);
  const float lo = 1e-6f, hi = 1.0f - 1e-6f;  // same fp32 constants as torch (1 - 1e-6 rounded to fp32)
  return fminf(fmaxf(v, lo), hi);
}

// encodings.py:27-81 after d / max(|d|, 1e-9) (applied twice in the reference:
// meta_ngp.py:165-168 and encodings.py:141; the second is a no-op up to rounding).
__device__ __forceinline__ void sh16_eval(float x, float y, float z, float* o) {
  float xx = x * x, yy = y * y, zz = z * z;
  o[0] = 0.28209479177387814f;
  o[1] = 0.4886025119029199f * y;
  o[2] = 0.4886025119029199f * z;
  o[3] = 0.4886025119029199f * x;
  o[4] = 1.0925484305920792f * x * y;
  o[5] = 1.0925484305920792f * y * z;
  o[6] = 0.9461746957575601f * zz - 0.31539156525251999f;
  o[7] = 1.0925484305920792f * x * z;
  o[8] = 0.5462742152960396f * (xx - yy);
  o[9] = 0.5900435899266435f * y * (3.f * xx - yy);
  o[10] = 2.890611442640554f * x * y * z;
  o[11] = 0.4570457994644658f * y * (5.f * zz - 1.f);
  o[12] = 0.3731763325901154f * z * (5.f * zz - 3.f);
  o[13] = 0.4570457994644658f * x * (5.f * zz - 1.f);
  o[14] = 1.445305721320277f * z * (xx - yy);
  o[15] = 0.5900435899266435f * x * (xx - 3.f * yy);
}

__device__ __forceinline__ void normalize3(float& x, float& y, float& z, float eps) {
  float n = sqrtf(x * x + y * y + z * z);
  n = fmaxf(n, eps);
  x = __fdiv_rn(x, n);
  y = __fdiv_rn(y, n);
  z = __fdiv_rn(z, n);
}

// Generic "dW = dz^T x" reduction used by every small-MLP backward (wgrad.cu).
// dzT (OUT, ld) and xT (IN, ld) are feature-major with the sample index contiguous.
// Segment e covers samples [seg[e], seg[e+1]) (seg on the device; NULL = one segment [0,M)).
// partial: (E, gx, OUT*(IN+1)) scratch; results: dW (OUT,IN) row-major and db (OUT).
struct WgradJob {
  const float* dzT;
  const float* xT;  // may be NULL when IN == 0 (bias only)
  int OUT, IN;
  int64_t ld;
};
int wgrad_launch(const WgradJob& job, const int32_t* seg_dev, int E, int64_t M_single, float* const* dW,
                 float* const* db, float* partial, cudaStream_t st);
int64_t wgrad_partial_floats(int E);

// ngp.cu: hash-table gradient scatter over pair slots [m_lo, m_hi) (grid_x = 0: default persistent grid)
int launch_hash_bwd_range(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g,
                          const float* d_enc, const AcnNgpExpertGrads* grads, int64_t m_lo, int64_t m_hi, int grid_x, cudaStream_t st);
// ngp_tc.cu: max |x| over n4 float4s, atomically maxed into *out (caller zeroes)
int launch_absmax(const float* x, int64_t n4, float* out, cudaStream_t st);

}  // namespace acn
