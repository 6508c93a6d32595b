// Generic small-MLP weight gradient: dW (OUT,IN) = sum_m dz[m,:]^T x[m,:], db = sum_m dz[m,:],
// reduced over the samples of one expert segment.  Inputs are feature-major
// (sample index contiguous) so every global load is a coalesced 128-byte row.
// Deterministic: each CTA writes a partial, a second kernel sums the partials in a fixed order.
// fp32 CUDA-core version (exact mode); the tcgen05 path computes wgrad inside the fused backward.
#include "common.cuh"

namespace acn {

constexpr int WG_T = 32;         // samples per smem tile
constexpr int WG_MAXF = 64;      // max OUT / IN
constexpr int WG_PSTRIDE = WG_MAXF * (WG_MAXF + 1);
constexpr int WG_CTAS = 296;     // 2 per SM

static inline int wg_gx(int E) {
  int g = WG_CTAS / (E > 0 ? E : 1);
  return g < 1 ? 1 : g;
}
int64_t wgrad_partial_floats(int E) { return (int64_t)E * wg_gx(E) * WG_PSTRIDE; }

struct WgOut {
  float* dW[ACN_MAX_EXPERTS];
  float* db[ACN_MAX_EXPERTS];
};

__global__ void __launch_bounds__(256) k_wgrad_partial(const float* __restrict__ dzT, const float* __restrict__ xT, int OUT, int IN,
                                                       int64_t ld, const int32_t* __restrict__ seg, int64_t M_single,
                                                       float* __restrict__ partial) {
  __shared__ float s_dz[WG_MAXF][WG_T + 1];
  __shared__ float s_x[WG_MAXF][WG_T + 1];
  int e = blockIdx.y, gx = gridDim.x;
  int64_t m0 = seg ? seg[e] : 0, m1 = seg ? seg[e + 1] : M_single;
  int64_t ntiles = (m1 - m0 + WG_T - 1) / WG_T;
  int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4], accb[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    accb[r] = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;
  }
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gx) {
    int64_t mb = m0 + tile * WG_T;
    // cooperative load: 64 rows x 32 samples for each operand; lane = sample
    for (int i = threadIdx.x; i < WG_MAXF * WG_T; i += 256) {
      int f = i >> 5, s = i & 31;
      int64_t m = mb + s;
      bool ok = m < m1;
      s_dz[f][s] = (ok && f < OUT) ? __ldg(dzT + (int64_t)f * ld + m) : 0.f;
      s_x[f][s] = (ok && f < IN) ? __ldg(xT + (int64_t)f * ld + m) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int s = 0; s < WG_T; ++s) {
      float a[4], b[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) a[r] = s_dz[ty * 4 + r][s];
#pragma unroll
      for (int c = 0; c < 4; ++c) b[c] = s_x[tx * 4 + c][s];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        accb[r] += a[r];
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(a[r], b[c], acc[r][c]);
      }
    }
    __syncthreads();
  }
  float* out = partial + ((int64_t)e * gx + blockIdx.x) * WG_PSTRIDE;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    int o = ty * 4 + r;
#pragma unroll
    for (int c = 0; c < 4; ++c) out[o * (WG_MAXF + 1) + tx * 4 + c] = acc[r][c];
    if (tx == 0) out[o * (WG_MAXF + 1) + WG_MAXF] = accb[r];
  }
}

// one warp per output element: lanes stride over the gx partials, then a fixed xor tree (deterministic)
__global__ void __launch_bounds__(256) k_wgrad_reduce(const float* __restrict__ partial, int gx, int OUT, int IN, WgOut outp) {
  int e = blockIdx.y;
  int n = OUT * (IN + 1);
  int lane = threadIdx.x & 31;
  for (int i = blockIdx.x * 8 + (threadIdx.x >> 5); i < n; i += gridDim.x * 8) {
    int o = i / (IN + 1), c = i - o * (IN + 1);
    int col = (c == IN) ? WG_MAXF : c;
    const float* p = partial + (int64_t)e * gx * WG_PSTRIDE + o * (WG_MAXF + 1) + col;
    float s = 0.f;
    for (int x = lane; x < gx; x += 32) s += p[(int64_t)x * WG_PSTRIDE];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if (lane == 0) {
      if (c == IN) {
        if (outp.db[e]) outp.db[e][o] = s;
      } else {
        if (outp.dW[e]) outp.dW[e][o * IN + c] = s;
      }
    }
  }
}

int wgrad_launch(const WgradJob& job, const int32_t* seg_dev, int E, int64_t M_single, float* const* dW, float* const* db,
                 float* partial, cudaStream_t st) {
  ACN_REQUIRE(job.OUT >= 1 && job.OUT <= WG_MAXF && job.IN >= 0 && job.IN <= WG_MAXF, "wgrad: layer shape out of range");
  ACN_REQUIRE(E >= 1 && E <= ACN_MAX_EXPERTS, "wgrad: bad expert count");
  int gx = wg_gx(E);
  if (!seg_dev) {  // single segment of known length: no more CTAs than 8-tile work items (the reduce walks gx partials serially)
    int64_t want = (M_single + WG_T * 8 - 1) / (WG_T * 8);
    if (want < gx) gx = (int)(want < 1 ? 1 : want);
  }
  dim3 grid(gx, E);
  k_wgrad_partial<<<grid, 256, 0, st>>>(job.dzT, job.xT ? job.xT : job.dzT, job.OUT, job.xT ? job.IN : 0, job.ld, seg_dev, M_single, partial);
  ACN_LAUNCH_CHECK();
  WgOut o;
  for (int e = 0; e < ACN_MAX_EXPERTS; ++e) {
    o.dW[e] = (e < E && dW) ? dW[e] : nullptr;
    o.db[e] = (e < E && db) ? db[e] : nullptr;
  }
  dim3 g2((job.OUT * (job.IN + 1) + 7) / 8, E);
  k_wgrad_reduce<<<g2, 256, 0, st>>>(partial, gx, job.OUT, job.IN, o);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // namespace acn
