// a16-a18: shared background head, alpha compositing (forward + backward), photometric loss.
//
// Compositing (nerfs/ray_rendering.py:114-165) is one warp per ray: lanes hold samples
// s = lane + 32*c, the transmittance T_s = prod_{j<s}(1 - alpha_j + 1e-10) is a
// warp-shuffle prefix product with a carry between 32-sample chunks, and the
// backward pass is the matching warp-shuffle suffix sum.  HBM-bound: 20 B read per
// sample (+4 B weights written when requested), 20 B written per ray.
#include <mutex>

#include "common.cuh"

// Scratch of acn_mse_loss_static (the entry point that allocates nothing): the per-CTA partial sums of k_mse_partial.
// One copy per GPU that loads the library, held until exit (8 KiB).  In HOST code this name is only a stand-in for the
// symbol, not a device address: the address a kernel may be given comes from cudaGetSymbolAddress (mse_static_scratch()).
static __device__ __align__(256) float g_mse_static_scratch[ACN_MSE_STATIC_PARTIALS];

namespace acn {

constexpr float ALPHA_MAX = (float)(1.0 - 1e-7);

__device__ __forceinline__ float warp_incl_prod(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v *= n;
  }
  return v;
}
__device__ __forceinline__ float warp_incl_suffix_sum(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float n = __shfl_down_sync(0xffffffffu, v, o);
    if (lane + o < 32) v += n;
  }
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// per-sample quantities shared by forward and backward
struct SampleEval {
  float alpha, fac, delta, e;  // fac = 1 - alpha + 1e-10 ; e = exp(-sigma*delta)
  bool alpha_pass, sigma_pass;
};

__device__ __forceinline__ SampleEval eval_sample(float sigma_in, float scale, const float* __restrict__ trow, int s, int S) {
  SampleEval r;
  float sg = fmaxf(sigma_in, 0.f);
  if (scale != 1.0f) sg = sg * scale;
  // dists: max(t[s+1]-t[s], 1e-4); the last sample repeats the previous distance
  int a = (s < S - 1) ? s : s - 1;
  float d = __fsub_rn(__ldg(trow + a + 1), __ldg(trow + a));
  r.delta = fmaxf(d, 1e-4f);
  r.e = expf(-(sg * r.delta));
  float araw = __fsub_rn(1.0f, r.e);
  r.alpha = fminf(fmaxf(araw, 0.f), ALPHA_MAX);
  r.alpha_pass = (araw >= 0.f) && (araw <= ALPHA_MAX);
  r.sigma_pass = sigma_in >= 0.f;
  r.fac = __fadd_rn(__fsub_rn(1.0f, r.alpha), 1e-10f);
  return r;
}

// rs element (ray r, sample s) sits at rs[r * ray_stride + s * sample_stride]: (S, 1) for the reference's (N,S,4) layout, (1, N) for
// the (S,N,4) output of a sample-major point set (AcnPoints.t_T)
__global__ void __launch_bounds__(256) k_composite_fwd(const float4* __restrict__ rs, int64_t ray_stride, int64_t sample_stride,
                                                       const float* __restrict__ t, const float* __restrict__ bg,
                                                       int64_t N, int S, float scale, float* __restrict__ rgb, float* __restrict__ depth,
                                                       float* __restrict__ weights, float* __restrict__ acc) {
  int lane = threadIdx.x & 31;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < N; r += nwarps) {
    const float* trow = t + r * S;
    float carry = 1.f, sr = 0.f, sg = 0.f, sb = 0.f, sd = 0.f, sa = 0.f;
    for (int c0 = 0; c0 < S; c0 += 32) {
      int s = c0 + lane;
      bool ok = s < S;
      float w = 0.f, fac = 1.f;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      float ts = 0.f;
      if (ok) {
        v = __ldg(rs + r * ray_stride + s * sample_stride);
        ts = __ldg(trow + s);
        SampleEval ev = eval_sample(v.w, scale, trow, s, S);
        fac = ev.fac;
        w = ev.alpha;
      }
      float incl = warp_incl_prod(fac, lane);
      float excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 1.f;
      float T = carry * excl;
      carry = carry * __shfl_sync(0xffffffffu, incl, 31);
      w = w * T;
      if (ok) {
        if (weights) weights[r * S + s] = w;
        sr += w * fminf(fmaxf(v.x, 0.f), 1.f);
        sg += w * fminf(fmaxf(v.y, 0.f), 1.f);
        sb += w * fminf(fmaxf(v.z, 0.f), 1.f);
        sd += w * ts;
        sa += w;
      }
    }
    sr = warp_sum(sr);
    sg = warp_sum(sg);
    sb = warp_sum(sb);
    sd = warp_sum(sd);
    sa = warp_sum(sa);
    if (lane == 0) {
      if (bg) {
        float k = 1.0f - sa;
        sr += k * __ldg(bg + r * 3);
        sg += k * __ldg(bg + r * 3 + 1);
        sb += k * __ldg(bg + r * 3 + 2);
      }
      rgb[r * 3] = sr;
      rgb[r * 3 + 1] = sg;
      rgb[r * 3 + 2] = sb;
      depth[r] = sd;
      acc[r] = sa;
    }
  }
}

// The same compositing for the (S,N) layouts of a sample-major point set (AcnPoints.t_T): one THREAD per ray walks its S samples
// front to back, so a warp reads 32 adjacent rays of one sample - full 512-byte rows of rs, 128-byte rows of t - and the
// transmittance is the sequential product the reference's cumprod forms.  No shuffles; 2 M rays of a frame fill the machine.
__global__ void __launch_bounds__(256) k_composite_fwd_sm(const float4* __restrict__ rs, const float* __restrict__ tT, const float* __restrict__ bg,
                                                          int64_t N, int S, float scale, float* __restrict__ rgb, float* __restrict__ depth,
                                                          float* __restrict__ wT, float* __restrict__ acc) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < N; r += (int64_t)gridDim.x * blockDim.x) {
    float T = 1.f, sr = 0.f, sg = 0.f, sb = 0.f, sd = 0.f, sa = 0.f;
    float t_cur = __ldg(tT + r), t_next = __ldg(tT + N + r), d_prev = 0.f;
    for (int s = 0; s < S; ++s) {
      const float4 v = __ldg(rs + (int64_t)s * N + r);
      const float d = (s < S - 1) ? __fsub_rn(t_next, t_cur) : d_prev;  // the last sample repeats the previous distance
      float sig = fmaxf(v.w, 0.f);
      if (scale != 1.0f) sig = sig * scale;
      const float delta = fmaxf(d, 1e-4f);
      const float e = expf(-(sig * delta));
      const float alpha = fminf(fmaxf(__fsub_rn(1.0f, e), 0.f), ALPHA_MAX);
      const float fac = __fadd_rn(__fsub_rn(1.0f, alpha), 1e-10f);
      const float w = alpha * T;
      T = T * fac;
      if (wT) wT[(int64_t)s * N + r] = w;
      sr += w * fminf(fmaxf(v.x, 0.f), 1.f);
      sg += w * fminf(fmaxf(v.y, 0.f), 1.f);
      sb += w * fminf(fmaxf(v.z, 0.f), 1.f);
      sd += w * t_cur;
      sa += w;
      d_prev = d;
      t_cur = t_next;
      if (s + 2 < S) t_next = __ldg(tT + (int64_t)(s + 2) * N + r);
    }
    if (bg) {
      const float k = 1.0f - sa;
      sr += k * __ldg(bg + r * 3);
      sg += k * __ldg(bg + r * 3 + 1);
      sb += k * __ldg(bg + r * 3 + 2);
    }
    rgb[r * 3] = sr;
    rgb[r * 3 + 1] = sg;
    rgb[r * 3 + 2] = sb;
    depth[r] = sd;
    acc[r] = sa;
  }
}

// CH = chunks of 32 samples held in registers (S <= 32*CH)
template <int CH>
__global__ void __launch_bounds__(256) k_composite_bwd(const float4* __restrict__ rs, const float* __restrict__ t, const float* __restrict__ bg,
                                                       int64_t N, int S, float scale, const float* __restrict__ g_rgb,
                                                       const float* __restrict__ g_depth, const float* __restrict__ g_acc,
                                                       const float* __restrict__ g_w, float4* __restrict__ d_rs, float* __restrict__ d_bg) {
  int lane = threadIdx.x & 31;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < N; r += nwarps) {
    const float* trow = t + r * S;
    float gr = __ldg(g_rgb + r * 3), gg = __ldg(g_rgb + r * 3 + 1), gb = __ldg(g_rgb + r * 3 + 2);
    float gd = g_depth ? __ldg(g_depth + r) : 0.f;
    float ga = g_acc ? __ldg(g_acc + r) : 0.f;
    float b0 = 0.f, b1 = 0.f, b2 = 0.f;
    if (bg) {
      b0 = __ldg(bg + r * 3);
      b1 = __ldg(bg + r * 3 + 1);
      b2 = __ldg(bg + r * 3 + 2);
      ga -= gr * b0 + gg * b1 + gb * b2;  // rgb += (1-acc)*bg
    }
    float T[CH], wv[CH], de[CH], fac[CH], alpha[CH], vterm[CH];
    float4 in[CH];
    float carry = 1.f, sa = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      int s = c * 32 + lane;
      bool ok = s < S;
      float f = 1.f;
      alpha[c] = 0.f;
      de[c] = 0.f;
      vterm[c] = 0.f;
      in[c] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c * 32 < S) {
        if (ok) {
          in[c] = __ldg(rs + r * S + s);
          SampleEval ev = eval_sample(in[c].w, scale, trow, s, S);
          f = ev.fac;
          alpha[c] = ev.alpha;
          de[c] = (ev.alpha_pass && ev.sigma_pass) ? ev.delta * ev.e * scale : 0.f;  // d alpha / d sigma_in
          float cr = fminf(fmaxf(in[c].x, 0.f), 1.f), cg = fminf(fmaxf(in[c].y, 0.f), 1.f), cb = fminf(fmaxf(in[c].z, 0.f), 1.f);
          vterm[c] = gr * cr + gg * cg + gb * cb + gd * __ldg(trow + s) + ga + (g_w ? __ldg(g_w + r * S + s) : 0.f);
        }
        float incl = warp_incl_prod(f, lane);
        float excl = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) excl = 1.f;
        T[c] = carry * excl;
        carry = carry * __shfl_sync(0xffffffffu, incl, 31);
      } else {
        T[c] = 0.f;
      }
      fac[c] = f;
      float w = alpha[c] * T[c];
      wv[c] = ok ? w * vterm[c] : 0.f;
      sa += ok ? w : 0.f;
    }
    float later = 0.f;  // sum of w_j v_j over later chunks
#pragma unroll
    for (int c = CH - 1; c >= 0; --c) {
      if (c * 32 < S) {
        int s = c * 32 + lane;
        bool ok = s < S;
        float suf = warp_incl_suffix_sum(wv[c], lane);
        float R = later + (suf - wv[c]);  // sum_{j>s} w_j v_j
        later += __shfl_sync(0xffffffffu, suf, 0);
        if (ok) {
          float dalpha = T[c] * vterm[c] - R / fac[c];
          float w = alpha[c] * T[c];
          float4 o;
          o.x = (in[c].x >= 0.f && in[c].x <= 1.f) ? w * gr : 0.f;  // clamp(0,1) passes the gradient inside the range
          o.y = (in[c].y >= 0.f && in[c].y <= 1.f) ? w * gg : 0.f;
          o.z = (in[c].z >= 0.f && in[c].z <= 1.f) ? w * gb : 0.f;
          o.w = dalpha * de[c];
          d_rs[r * S + s] = o;
        }
      }
    }
    if (d_bg && bg) {
      sa = warp_sum(sa);
      if (lane == 0) {
        float k = 1.0f - sa;
        d_bg[r * 3] = k * gr;
        d_bg[r * 3 + 1] = k * gg;
        d_bg[r * 3 + 2] = k * gb;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// background head: sigmoid(W1 relu(W0 SH16(normalize(d)) + b0) + b1), meta_container.py:347-382
// save rows (feature-major, ld = round128(N)): sh 16, h 32, rgb 3
// ---------------------------------------------------------------------------------------
constexpr int BG_R_SH = 0, BG_R_H = 16, BG_R_RGB = 48, BG_R_SAVE = 51;
constexpr int BG_R_DZ1 = 0, BG_R_DZ0 = 4, BG_R_WS = 36;

__global__ void __launch_bounds__(256) k_bg_fwd(const float* __restrict__ d, int64_t stride, int64_t N, const float* __restrict__ w0,
                                                const float* __restrict__ b0, const float* __restrict__ w1, const float* __restrict__ b1,
                                                float* __restrict__ rgb, float* __restrict__ sv, int64_t ld) {
  __shared__ float s_w0[32 * 16], s_b0[32], s_w1[3 * 32], s_b1[3];
  for (int i = threadIdx.x; i < 512; i += 256) s_w0[i] = __ldg(w0 + i);
  if (threadIdx.x < 32) s_b0[threadIdx.x] = __ldg(b0 + threadIdx.x);
  if (threadIdx.x < 96) s_w1[threadIdx.x] = __ldg(w1 + threadIdx.x);
  if (threadIdx.x < 3) s_b1[threadIdx.x] = __ldg(b1 + threadIdx.x);
  __syncthreads();
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < N; r += (int64_t)gridDim.x * blockDim.x) {
    float x = __ldg(d + r * stride), y = __ldg(d + r * stride + 1), z = __ldg(d + r * stride + 2);
    normalize3(x, y, z, 1e-12f);  // F.normalize
    normalize3(x, y, z, 1e-9f);   // SHEncoder.forward
    float sh[16];
    sh16_eval(x, y, z, sh);
    float o0 = s_b1[0], o1 = s_b1[1], o2 = s_b1[2];
#pragma unroll 4
    for (int j = 0; j < 32; ++j) {
      float a = s_b0[j];
#pragma unroll
      for (int i = 0; i < 16; ++i) a = fmaf(s_w0[j * 16 + i], sh[i], a);
      a = fmaxf(a, 0.f);
      if (sv) sv[(int64_t)(BG_R_H + j) * ld + r] = a;
      o0 = fmaf(s_w1[j], a, o0);
      o1 = fmaf(s_w1[32 + j], a, o1);
      o2 = fmaf(s_w1[64 + j], a, o2);
    }
    o0 = 1.f / (1.f + expf(-o0));
    o1 = 1.f / (1.f + expf(-o1));
    o2 = 1.f / (1.f + expf(-o2));
    rgb[r * 3] = o0;
    rgb[r * 3 + 1] = o1;
    rgb[r * 3 + 2] = o2;
    if (sv) {
#pragma unroll
      for (int i = 0; i < 16; ++i) sv[(int64_t)(BG_R_SH + i) * ld + r] = sh[i];
      sv[(int64_t)(BG_R_RGB + 0) * ld + r] = o0;
      sv[(int64_t)(BG_R_RGB + 1) * ld + r] = o1;
      sv[(int64_t)(BG_R_RGB + 2) * ld + r] = o2;
    }
  }
}

// Backward of the whole head in one kernel + a fixed-order reduce (was 5 launches, 54 us for 8192 rays).
// A CTA takes 128-ray tiles: thread = ray builds dz1 (3) and dz0 (32) and parks them with sh (16) and h (32) in shared
// memory (rows padded to 129 floats: conflict-free when a warp reads different rows at one ray); then thread = output
// accumulates its <= 6 of the 643 gradient entries over the tile's rays.  Per-CTA partials, summed in CTA order.
constexpr int BG_NOUT = 512 + 32 + 96 + 3;  // dW0 (32,16) | db0 (32) | dW1 (3,32) | db1 (3)
constexpr int BG_PSTRIDE = 648;
constexpr int BG_TILE = 128, BG_PAD = 129;
constexpr int BG_MAX_CTAS = 592;

__global__ void __launch_bounds__(BG_TILE) k_bg_bwd_fused(const float* __restrict__ d_rgb, int64_t N, const float* __restrict__ w1,
                                                          const float* __restrict__ sv, int64_t ld, float* __restrict__ partial) {
  __shared__ float s_w1[3 * 32];
  __shared__ float s_a[35][BG_PAD];  // rows 0..31 dz0, 32..34 dz1
  __shared__ float s_b[49][BG_PAD];  // rows 0..15 sh, 16..47 h, 48 ones
  const int t = threadIdx.x;
  if (t < 96) s_w1[t] = __ldg(w1 + t);
  s_b[48][t] = 1.f;
  // output o of this thread: A row ra[k], B row rb[k]
  int ra[6], rb[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    int o = t + k * BG_TILE;
    if (o < 512) { ra[k] = o >> 4; rb[k] = o & 15; }
    else if (o < 544) { ra[k] = o - 512; rb[k] = 48; }
    else if (o < 640) { ra[k] = 32 + (o - 544) / 32; rb[k] = 16 + (o - 544) % 32; }
    else if (o < BG_NOUT) { ra[k] = 32 + (o - 640); rb[k] = 48; }
    else { ra[k] = 0; rb[k] = 48; }
  }
  float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const int64_t tiles = (N + BG_TILE - 1) / BG_TILE;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t r = tile * BG_TILE + t;
    const bool ok = r < N;
    __syncthreads();  // previous tile's readers are done (and s_w1 / ones are in place on the first pass)
    float dz[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float o = ok ? sv[(int64_t)(BG_R_RGB + c) * ld + r] : 0.f;
      dz[c] = ok ? __ldg(d_rgb + r * 3 + c) * o * (1.f - o) : 0.f;
      s_a[32 + c][t] = dz[c];
    }
#pragma unroll 8
    for (int j = 0; j < 32; ++j) {
      float h = ok ? sv[(int64_t)(BG_R_H + j) * ld + r] : 0.f;
      float v = s_w1[j] * dz[0] + s_w1[32 + j] * dz[1] + s_w1[64 + j] * dz[2];
      s_a[j][t] = (h > 0.f) ? v : 0.f;
      s_b[16 + j][t] = h;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) s_b[i][t] = ok ? sv[(int64_t)(BG_R_SH + i) * ld + r] : 0.f;
    __syncthreads();
#pragma unroll 4
    for (int q = 0; q < BG_TILE; ++q) {
#pragma unroll
      for (int k = 0; k < 6; ++k) acc[k] = fmaf(s_a[ra[k]][q], s_b[rb[k]][q], acc[k]);
    }
  }
  float* out = partial + (int64_t)blockIdx.x * BG_PSTRIDE;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    int o = t + k * BG_TILE;
    if (o < BG_NOUT) out[o] = acc[k];
  }
}

__global__ void __launch_bounds__(128) k_bg_bwd_reduce(const float* __restrict__ partial, int G, float* __restrict__ g_w0, float* __restrict__ g_b0,
                                                       float* __restrict__ g_w1, float* __restrict__ g_b1) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= BG_NOUT) return;
  float acc = 0.f;
  int x = 0;
  for (; x + 8 <= G; x += 8) {  // eight loads in flight, added in CTA order (deterministic)
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = partial[(int64_t)(x + u) * BG_PSTRIDE + o];
#pragma unroll
    for (int u = 0; u < 8; ++u) acc += v[u];
  }
  for (; x < G; ++x) acc += partial[(int64_t)x * BG_PSTRIDE + o];
  if (o < 512) { if (g_w0) g_w0[o] = acc; }
  else if (o < 544) { if (g_b0) g_b0[o - 512] = acc; }
  else if (o < 640) { if (g_w1) g_w1[o - 544] = acc; }
  else { if (g_b1) g_b1[o - 640] = acc; }
}

// ---------------------------------------------------------------------------------------
// photometric loss: nerfs/losses.py:10-32 + nerfs/color_space.py:22-66
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float srgb_to_linear_f(float x) { return (x <= 0.04045f) ? x / 12.92f : powf((x + 0.055f) / 1.055f, 2.4f); }

__global__ void __launch_bounds__(256) k_mse_partial(const float* __restrict__ pred, const float* __restrict__ gt, int64_t n, int space,
                                                     float inv_n, float* __restrict__ d_pred, float* __restrict__ partial) {
  __shared__ float red[8];
  float acc = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float p = pred[i], g = fminf(fmaxf(gt[i], 0.f), 1.f);
    float pp, dpp;
    if (space == 0) {
      g = fminf(fmaxf(srgb_to_linear_f(g), 0.f), 1.f);
      pp = fminf(fmaxf(p, 0.f), 1.f);
      dpp = (p >= 0.f && p <= 1.f) ? 1.f : 0.f;
    } else if (space == 1) {
      float x = fminf(fmaxf(p, 0.f), 1.f);
      bool in = p >= 0.f && p <= 1.f;
      if (x <= 0.0031308f) {
        pp = 12.92f * x;
        dpp = in ? 12.92f : 0.f;
      } else {
        pp = 1.055f * powf(x, 1.f / 2.4f) - 0.055f;
        dpp = in ? 1.055f / 2.4f * powf(x, 1.f / 2.4f - 1.f) : 0.f;
      }
      if (!(pp >= 0.f && pp <= 1.f)) dpp = 0.f;
      pp = fminf(fmaxf(pp, 0.f), 1.f);
    } else {
      pp = p;
      dpp = 1.f;
    }
    float diff = pp - g;
    acc += diff * diff;
    if (d_pred) d_pred[i] = 2.f * diff * inv_n * dpp;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 8) {
    float v = red[threadIdx.x];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffu, v, o);
    if (threadIdx.x == 0) partial[blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(256) k_mse_final(const float* __restrict__ partial, int n, float inv_n, float* __restrict__ loss) {
  __shared__ float red[8];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) acc += partial[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = 0.f;
    for (int i = 0; i < 8; ++i) v += red[i];
    loss[0] = v * inv_n;
  }
}

}  // namespace acn

using namespace acn;

static inline int64_t r128(int64_t n) { return (n + 127) / 128 * 128; }

extern "C" {

int acn_composite_fwd_strided(const float* rgb_sigma, int64_t ray_stride, int64_t sample_stride, const float* t, const float* bg, int64_t N,
                              int S, float sigma_scale, float* rgb, float* depth, float* weights, float* acc, void* stream) {
  ACN_REQUIRE(rgb_sigma && t && rgb && depth && acc && N >= 0, "acn_composite_fwd: NULL argument");
  ACN_REQUIRE(S >= 2, "acn_composite_fwd: needs at least 2 samples per ray (got %d)", S);
  ACN_REQUIRE(ray_stride >= 1 && sample_stride >= 1, "acn_composite_fwd_strided: strides are in (rgb,sigma) rows and positive");
  if (N == 0) return ACN_OK;
  int grid = persistent_grid((N + 7) / 8, 8);
  k_composite_fwd<<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float4*>(rgb_sigma), ray_stride, sample_stride, t, bg, N, S,
                                                       sigma_scale, rgb, depth, weights, acc);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_composite_fwd_sm(const float* rgb_sigma_T, const float* t_T, const float* bg, int64_t N, int S, float sigma_scale, float* rgb,
                         float* depth, float* weights_T, float* acc, void* stream) {
  ACN_REQUIRE(rgb_sigma_T && t_T && rgb && depth && acc && N >= 0, "acn_composite_fwd_sm: NULL argument");
  ACN_REQUIRE(S >= 2, "acn_composite_fwd_sm: needs at least 2 samples per ray (got %d)", S);
  if (N == 0) return ACN_OK;
  k_composite_fwd_sm<<<persistent_grid((N + 255) / 256, 8), 256, 0, as_stream(stream)>>>(reinterpret_cast<const float4*>(rgb_sigma_T), t_T, bg, N, S,
                                                                                        sigma_scale, rgb, depth, weights_T, acc);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_composite_fwd(const float* rgb_sigma, const float* t, const float* bg, int64_t N, int S, float sigma_scale, float* rgb,
                      float* depth, float* weights, float* acc, void* stream) {
  return acn_composite_fwd_strided(rgb_sigma, S, 1, t, bg, N, S, sigma_scale, rgb, depth, weights, acc, stream);
}

int acn_composite_bwd(const float* rgb_sigma, const float* t, const float* bg, int64_t N, int S, float sigma_scale, const float* g_rgb,
                      const float* g_depth, const float* g_acc, const float* g_weights, float* d_rgb_sigma, float* d_bg, void* stream) {
  ACN_REQUIRE(rgb_sigma && t && g_rgb && d_rgb_sigma && N >= 0, "acn_composite_bwd: NULL argument");
  ACN_REQUIRE(S >= 2 && S <= 1024, "acn_composite_bwd: samples per ray must be in [2,1024] (got %d)", S);
  if (N == 0) return ACN_OK;
  int grid = persistent_grid((N + 7) / 8, 4);
  cudaStream_t st = as_stream(stream);
  const float4* rs = reinterpret_cast<const float4*>(rgb_sigma);
  float4* drs = reinterpret_cast<float4*>(d_rgb_sigma);
#define ACN_CB(CH) k_composite_bwd<CH><<<grid, 256, 0, st>>>(rs, t, bg, N, S, sigma_scale, g_rgb, g_depth, g_acc, g_weights, drs, d_bg)
  if (S <= 64) ACN_CB(2);
  else if (S <= 128) ACN_CB(4);
  else if (S <= 256) ACN_CB(8);
  else if (S <= 512) ACN_CB(16);
  else ACN_CB(32);
#undef ACN_CB
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int64_t acn_bg_save_bytes(int64_t N) { return sizeof(float) * BG_R_SAVE * r128(N); }
int64_t acn_bg_bwd_ws_bytes(int64_t N) {
  (void)N;
  return sizeof(float) * (int64_t)BG_MAX_CTAS * BG_PSTRIDE;
}

int acn_bg_forward(const float* d, int64_t stride, int64_t N, const float* w0, const float* b0, const float* w1, const float* b1,
                   float* rgb, void* save, void* stream) {
  ACN_REQUIRE(d && w0 && b0 && w1 && b1 && rgb && N >= 0 && stride >= 3, "acn_bg_forward: bad arguments");
  if (N == 0) return ACN_OK;
  k_bg_fwd<<<persistent_grid((N + 255) / 256, 4), 256, 0, as_stream(stream)>>>(d, stride, N, w0, b0, w1, b1, rgb,
                                                                               reinterpret_cast<float*>(save), r128(N));
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_bg_backward(const float* d_rgb, int64_t N, const float* w0, const float* w1, const void* save, float* g_w0, float* g_b0,
                    float* g_w1, float* g_b1, void* ws, void* stream) {
  (void)w0;
  ACN_REQUIRE(d_rgb && w1 && save && ws && N >= 0, "acn_bg_backward: bad arguments");
  cudaStream_t st = as_stream(stream);
  int64_t ld = r128(N);
  const float* sv = reinterpret_cast<const float*>(save);
  float* w = reinterpret_cast<float*>(ws);
  int G = (int)((N + BG_TILE - 1) / BG_TILE);
  if (G > BG_MAX_CTAS) G = BG_MAX_CTAS;
  if (G < 1) G = 1;  // N == 0: one CTA writes a zero partial, the gradients come out zero
  k_bg_bwd_fused<<<G, BG_TILE, 0, st>>>(d_rgb, N, w1, sv, ld, w);
  ACN_LAUNCH_CHECK();
  k_bg_bwd_reduce<<<(BG_NOUT + 127) / 128, 128, 0, st>>>(w, G, g_w0, g_b0, g_w1, g_b1);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_mse_loss(const float* pred, const float* gt, int64_t N, int space, float* loss, float* d_pred, float* ws, void* stream) {
  ACN_REQUIRE(pred && gt && loss && ws && N > 0 && space >= 0 && space <= 2, "acn_mse_loss: bad arguments");
  int64_t n = N * 3;
  int grid = persistent_grid((n + 255) / 256, 4);
  if (grid > 2048) grid = 2048;
  float inv_n = (float)(1.0 / (double)n);
  k_mse_partial<<<grid, 256, 0, as_stream(stream)>>>(pred, gt, n, space, inv_n, d_pred, ws);
  ACN_LAUNCH_CHECK();
  k_mse_final<<<1, 256, 0, as_stream(stream)>>>(ws, grid, inv_n, loss);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

// Device address of g_mse_static_scratch on the CURRENT device, looked up once per device (cudaGetSymbolAddress allocates
// nothing: the array is part of the loaded module).
static int mse_static_scratch(float** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static float* addr[MAX_DEV] = {};
  int dev = -1;
  ACN_CUDA(cudaGetDevice(&dev));
  ACN_REQUIRE(dev >= 0 && dev < MAX_DEV, "acn_mse_loss_static: device ordinal %d outside [0,%d)", dev, MAX_DEV);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    float* p = nullptr;
    ACN_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_mse_static_scratch));
    ACN_REQUIRE(p != nullptr, "acn_mse_loss_static: cudaGetSymbolAddress returned a null address");
    addr[dev] = p;
  }
  *out = addr[dev];
  return ACN_OK;
}

int acn_mse_loss_static(const float* pred, const float* gt, int64_t N, int space, float* loss, float* d_pred, void* stream) {
  ACN_REQUIRE(pred && gt && loss && N > 0 && space >= 0 && space <= 2, "acn_mse_loss_static: bad arguments");
  ACN_REQUIRE(N <= (int64_t)ACN_MSE_STATIC_MAX_RAYS, "acn_mse_loss_static: N = %lld rays exceeds the static scratch's limit of %lld (use acn_mse_loss with a caller workspace)",
              (long long)N, (long long)ACN_MSE_STATIC_MAX_RAYS);
  float* scratch = nullptr;
  int rc = mse_static_scratch(&scratch);
  if (rc != ACN_OK) return rc;
  const int64_t n = N * 3;
  int grid = persistent_grid((n + 255) / 256, 4);
  if (grid > ACN_MSE_STATIC_PARTIALS) grid = ACN_MSE_STATIC_PARTIALS;
  const float inv_n = (float)(1.0 / (double)n);
  // every partial the final kernel reads (0 .. grid-1) is overwritten by this call's first kernel: nothing has to start at zero
  k_mse_partial<<<grid, 256, 0, as_stream(stream)>>>(pred, gt, n, space, inv_n, d_pred, scratch);
  ACN_LAUNCH_CHECK();
  k_mse_final<<<1, 256, 0, as_stream(stream)>>>(scratch, grid, inv_n, loss);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
